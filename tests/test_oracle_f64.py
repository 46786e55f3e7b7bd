"""float64 oracle (oracle/dc_oracle_f64.c) vs the UNMODIFIED reference's *_double functions compiled into oracle/_ref
(SURVEY.md 8f row N1).  The reference holds no golden vectors for the double codecs, so this comparison -- and the committed
fixture tests/golden/f64_golden.npz generated from the reference by tests/golden/make_f64_golden.py -- is what pins the
oracle.  Skipped where oracle/_ref was not built (it needs /root/reference at build time)."""
import os

import numpy as np
import pytest

import oracle_lib as ol

EBS = (0.1, 0.01, 1e-3, 1e-4, 1e-5, 1e-6)
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "f64_golden.npz")


def _ref(eb):
    if not ol.build_ref(eb):
        pytest.skip("oracle/_ref not built")
    return ol.RefLibF64(eb)


def _cases(rng):
    n = 1200
    yield "uniform10", rng.random(n) * 10
    yield "uniform1", rng.random(n)
    yield "ramp", 100 + 0.0049 * np.arange(n)
    yield "sine", 1 + 0.5 * np.sin(np.arange(n) / 100.0)
    yield "const", np.full(n, 0.123456789)
    yield "zeros", np.zeros(n)
    yield "tiny_n1", np.array([3.25])
    yield "tiny_n2", np.array([3.25, 0.0])
    yield "tiny_n4", np.array([1.0, 2.0, 3.0, 4.0])
    yield "big", rng.random(n) * 1e200
    yield "small", rng.random(n) * 1e-200
    yield "denormal", rng.random(n) * 1e-310
    w = rng.random(n) * 4
    w[rng.integers(0, n, 40)] = 0.0
    yield "with_zeros", w
    yield "kmeans_like", np.sort(rng.random(n) * 1000.0)           # cluster centres, impl/k-means.c


@pytest.mark.parametrize("eb", EBS)
def test_streams_and_decode_match_reference(eb):
    R = _ref(eb)
    rng = np.random.default_rng(int(1 / eb) + 64)
    for name, x in _cases(rng):
        mn, y = R.min_subtract(x)
        mn2, y2 = ol.o_min_subtract_f64(x)
        assert mn == mn2 and np.array_equal(y.view(np.uint64), y2.view(np.uint64))
        med, t = R.med(y, 0)
        med2, t2 = ol.o_med_f64(y, 0)
        assert (med.view(np.uint64), t) == (med2.view(np.uint64), t2), name
        m20 = ol.mask20_of(med)
        for ct in ol.ALL_CTS:
            s, pos = R.compress(ct, y, t, m20)
            s2, pos2 = ol.o_compress_f64(ct, y, eb, t, m20)
            assert pos == pos2 and np.array_equal(s, s2), (name, ct)
            if ct == ol.CT_MASK and not ol.ref_mask_decode_defined_f64(t, m20, eb):
                continue
            d = R.decompress(ct, s, y.size, t, m20)
            d2 = ol.o_decompress_f64(ct, s2, y.size, eb, t, m20)
            assert np.array_equal(d.view(np.uint64), d2.view(np.uint64)), (name, ct)


@pytest.mark.parametrize("eb", (0.01, 1e-5))
def test_forced_masks_match_reference(eb):
    """CT=7 with types and masks the data would not choose itself: every escape shape (flag 0/1, payload 0..52)."""
    R = _ref(eb)
    rng = np.random.default_rng(11)
    y = rng.random(900) * 6
    for t in (2, 3, 5, 9):
        for medium in (1.7, 3.3, 0.004, 5.999, 700.0):
            m20 = ol.mask20_of(medium)
            s, pos = R.compress(ol.CT_MASK, y, t, m20)
            s2, pos2 = ol.o_compress_f64(ol.CT_MASK, y, eb, t, m20)
            assert pos == pos2 and np.array_equal(s, s2), (t, medium)
            if ol.ref_mask_decode_defined_f64(t, m20, eb):
                d = R.decompress(ol.CT_MASK, s, y.size, t, m20)
                d2 = ol.o_decompress_f64(ol.CT_MASK, s2, y.size, eb, t, m20)
                assert np.array_equal(d.view(np.uint64), d2.view(np.uint64)), (t, medium)


@pytest.mark.parametrize("eb", (0.01, 1e-5))
def test_encoder_only_inputs(eb):
    """Negative values, the -1.0 history sentinel, NaN/Inf: deterministic ENCODING (decoding is undefined in the reference)."""
    R = _ref(eb)
    rng = np.random.default_rng(7)
    n = 800
    neg = rng.random(n) * 8 - 4
    sent = rng.random(n) * 3
    sent[[0, 5, 6, 7, 100, 101, 300, n - 1]] = -1.0
    sent2 = np.full(n, -1.0)
    sent2[::7] = 2.5
    spec = rng.random(n) * 3
    spec[10], spec[20], spec[30] = np.inf, np.nan, -np.inf
    for name, y in (("neg", neg), ("sentinel", sent), ("sentinel2", sent2), ("special", spec)):
        for ct in ol.ALL_CTS:
            t, m20 = 2, ol.mask20_of(1.7)
            s, pos = R.compress(ct, y, t, m20)
            s2, pos2 = ol.o_compress_f64(ct, y, eb, t, m20)
            assert pos == pos2 and np.array_equal(s, s2), (name, ct)


def test_round_trip_and_chunking():
    rng = np.random.default_rng(3)
    x = np.cumsum(rng.standard_normal(10000)) * 0.01 + 50
    for eb in (0.01, 1e-6):
        for ct in ol.ALL_CTS:
            enc = ol.o_compress_chunked_f64(ct, x, 1024, eb)
            out = ol.o_decompress_chunked_f64(ct, enc, x.size, 1024, eb)
            assert np.isfinite(out).all()
            if ct == ol.CT_NP:
                assert np.abs(out - x).max() <= 2 * eb           # pure truncation codes respect the bound
            K = (x.size + 1023) // 1024
            assert enc["offsets"].size == K + 1 and enc["offsets"][-1] == enc["stream"].size


def test_golden_fixture():
    """tests/golden/f64_golden.npz: inputs and the REFERENCE's streams/decodes (made by tests/golden/make_f64_golden.py)."""
    if not os.path.exists(GOLD):
        pytest.skip("fixture not generated")
    g = np.load(GOLD)
    for key in [k for k in g.files if k.endswith("_x")]:
        tag = key[:-2]
        x, eb, ct = g[key], float(g[tag + "_eb"]), int(g[tag + "_ct"])
        t, m20 = int(g[tag + "_type"]), int(g[tag + "_mask20"])
        s2, _ = ol.o_compress_f64(ct, x, eb, t, m20)
        assert np.array_equal(s2, g[tag + "_stream"]), tag
        if (tag + "_decoded") in g.files:
            d2 = ol.o_decompress_f64(ct, s2, x.size, eb, t, m20)
            assert np.array_equal(d2.view(np.uint64), g[tag + "_decoded"].view(np.uint64)), tag
