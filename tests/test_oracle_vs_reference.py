"""Oracle (restatement) vs the UNMODIFIED reference compiled by oracle/build_ref.sh (oracle/_ref/*.so).
Skipped where the reference library was not built (it needs /root/reference at build time)."""
import numpy as np
import pytest

import oracle_lib as ol

EBS = (0.1, 0.01, 1e-3, 1e-4, 1e-5, 1e-6)


def _ref(eb):
    if not ol.build_ref(eb):
        pytest.skip("oracle/_ref not built")
    return ol.RefLib(eb)


def _cases(rng):
    n = 1500
    yield "uniform10", (rng.random(n) * 10).astype(np.float32)
    yield "uniform1", rng.random(n).astype(np.float32)
    yield "ramp", (100 + 0.0049 * np.arange(n)).astype(np.float32)
    yield "sine", (1 + 0.5 * np.sin(np.arange(n) / 100.0)).astype(np.float32)
    yield "const", np.full(n, 0.123456789, np.float32)
    yield "zeros", np.zeros(n, np.float32)
    yield "tiny_n1", np.array([3.25], np.float32)
    yield "tiny_n2", np.array([3.25, 0.0], np.float32)
    yield "tiny_n4", np.array([1.0, 2.0, 3.0, 4.0], np.float32)
    yield "big", (rng.random(n) * 1e30).astype(np.float32)
    yield "small", (rng.random(n) * 1e-30).astype(np.float32)
    d = rng.random(n).astype(np.float32) * np.float32(1e-40)           # denormals
    yield "denormal", d
    w = (rng.random(n) * 4).astype(np.float32)
    w[rng.integers(0, n, 40)] = 0.0
    yield "with_zeros", w


@pytest.mark.parametrize("eb", EBS)
def test_streams_and_decode_match_reference(eb):
    R = _ref(eb)
    rng = np.random.default_rng(int(1 / eb))
    for name, x in _cases(rng):
        mn, y = R.min_subtract(x)
        mn2, y2 = ol.o_min_subtract(x)
        assert mn == mn2 and np.array_equal(y.view(np.uint32), y2.view(np.uint32))
        med, t = R.med(y, 0)
        med2, t2 = ol.o_med(y, 0)
        assert (med.view(np.uint32), t) == (med2.view(np.uint32), t2), name
        m17 = ol.mask17_of(med)
        for ct in ol.ALL_CTS:
            s, pos = R.compress(ct, y, t, m17)
            s2, pos2 = ol.o_compress(ct, y, eb, t, m17)
            assert pos == pos2 and np.array_equal(s, s2), (name, ct)
            assert R.crc32(s) == ol.o_crc32(s2)
            if ct == ol.CT_MASK and not ol.ref_mask_decode_defined(t, m17, eb):
                continue        # reference decoder undefined here; the oracle's round trip is tested below
            d = R.decompress(ct, s, y.size, t, m17)
            d2 = ol.o_decompress(ct, s2, y.size, eb, t, m17)
            assert np.array_equal(d.view(np.uint32), d2.view(np.uint32)), (name, ct)


@pytest.mark.parametrize("eb", (0.01, 1e-5))
def test_encoder_only_inputs(eb):
    """Inputs the reference cannot DECODE (negative values, -1.0 sentinel, NaN/Inf: SURVEY.md F3, F7) still
    have a deterministic ENCODING; the stream must match bit for bit."""
    R = _ref(eb)
    rng = np.random.default_rng(7)
    n = 800
    neg = (rng.random(n) * 8 - 4).astype(np.float32)
    sent = (rng.random(n) * 3).astype(np.float32)
    sent[[0, 5, 6, 7, 100, 101, 300, n - 1]] = -1.0
    sent2 = np.full(n, -1.0, np.float32)
    sent2[::7] = 2.5
    spec = (rng.random(n) * 3).astype(np.float32)
    spec[10], spec[20], spec[30] = np.inf, np.nan, -np.inf
    for name, y in (("neg", neg), ("sentinel", sent), ("sentinel2", sent2), ("special", spec)):
        for ct in (ol.CT_BITWISE, ol.CT_NP, ol.CT_OP, ol.CT_MASK):
            t, m17 = 2, ol.mask17_of(np.float32(1.7))
            s, pos = R.compress(ct, y, t, m17)
            s2, pos2 = ol.o_compress(ct, y, eb, t, m17)
            assert pos == pos2 and np.array_equal(s, s2), (name, ct)


def test_decoder_sentinel_state_machine():
    """Hand-built CT=5 streams that drive the decoder's -1 history sentinel (dataCompression.c:2930, :3110-3127):
    '101' before any value decodes to b1 = -1.0, which never enters the history."""
    eb = 0.01
    R = _ref(eb)
    def pack(bits):
        bits += "0" * (-len(bits) % 8)
        return np.frombuffer(int(bits, 2).to_bytes(len(bits) // 8, "big"), np.uint8)
    raw2 = "0" + format(128, "08b") + "10000000"      # 2.5-ish: E=128 -> m=8
    raw3 = "0" + format(128, "08b") + "11000000"
    for codes in (["101", raw2, "101", raw3, raw2, "110", "111", "101"],
                  [raw2, raw3, "110", "101", "111", raw2, "111"],
                  ["100", "100", "100", "111", "110", raw3, "110"],
                  ["110", "111", raw2, raw2, raw2, "111"]):
        s = pack("".join(codes))
        d = R.decompress(ol.CT_BITWISE, s, len(codes))
        d2 = ol.o_decompress(ol.CT_BITWISE, s, len(codes), eb)
        assert np.array_equal(d.view(np.uint32), d2.view(np.uint32)), codes


def _py_mask_decode(stream, n, n_eb, type_, mask17):
    """Independent code-by-code model of the CT=7 format as the ENCODER defines it."""
    bits = "".join(format(b, "08b") for b in stream) + "0" * 64
    f32 = np.float32
    u2f = lambda u: np.array([u], np.uint32).view(np.float32)[0]
    mid = lambda top, have: u2f(top if have >= 32 else top | (0x80000000 >> have))
    mm = lambda E: min(max(n_eb + E - 127, 0), 23)
    out, p, hist = [], 0, []
    b = [f32(-1), f32(-1), f32(-1)]                      # b3, b2, b1
    for _ in range(n):
        if bits[p] == "1":
            c = bits[p:p + 3]; p += 3
            b3, b2, b1 = b
            v = {"100": f32(0), "101": b1, "110": f32(f32(2) * b1 - b2),
                 "111": f32(f32(f32(3) * b1 - f32(3) * b2) + b3)}[c]
        elif bits[p + 1:p + 1 + type_] == "1" * type_:
            flag = bits[p + 1 + type_]; p += type_ + 2
            m = mm((mask17 >> 8) & 0xFF)
            k = m if flag == "1" else max(m - 8, 0)
            pay = int(bits[p:p + k], 2) if k else 0
            p += k
            if flag == "1":
                v = mid(((mask17 >> 8) << 23) | (pay << (23 - k)), 9 + k)
            else:
                v = mid((mask17 << 15) | (pay << (15 - k)), 17 + k)
        else:
            E = int(bits[p + 1:p + 9], 2); L = 9 + mm(E)
            v = mid(int(bits[p:p + L], 2) << (32 - L), L); p += L
        v = f32(v)
        out.append(v)
        if b[0] == -1: b[0] = v
        elif b[1] == -1: b[1] = v
        elif b[2] == -1: b[2] = v
        else: b = [b[1], b[2], v]
    return np.array(out, np.float32)


def test_mask_oracle_round_trip_outside_reference_domain():
    """Where the reference CT=7 decoder is undefined the oracle implements the evident intent;
    check it against an independent Python model of the format, including type==1 and 0 < m <= 8."""
    rng = np.random.default_rng(3)
    with np.errstate(all="ignore"):
        for eb in (0.1, 0.01, 1e-3):
            for scale in (1.0, 10.0, 1e3):
                y = (rng.random(1200) * scale).astype(np.float32)
                y[0] = 0
                y[100:400] = np.float32(scale / 2) + (rng.random(300) * scale / 512).astype(np.float32)
                med, t = ol.o_med(y, 0)
                m17 = ol.mask17_of(med)
                s, _ = ol.o_compress(ol.CT_MASK, y, eb, t, m17)
                d7 = ol.o_decompress(ol.CT_MASK, s, y.size, eb, t, m17)
                ref = _py_mask_decode(s, y.size, ol.o_n_eb(eb), t, m17)
                assert np.array_equal(d7.view(np.uint32), ref.view(np.uint32)), (eb, scale)
