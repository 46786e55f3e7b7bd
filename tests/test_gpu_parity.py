"""GPU parity tests proper: the CUDA path, called through the C ABI (data_compression_b200.api -> libdcb200.so),
against the CPU oracle on the same seeded inputs, bit for bit: stream bytes, chunk table, decoded floats.
Run on the B200 box with  python -m pytest tests -m gpu."""
import os

import numpy as np
import pytest

import oracle_lib as ol
from data_compression_b200 import api
from data_compression_b200.fields import make_field

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

EBS = (0.1, 0.01, 1e-4, 1e-6)
CTS = (5, 6, 7, 11)


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.fail("no CUDA device: the product has no CPU fallback and these tests need a B200")
    return torch.device("cuda:0")


def gpu_encode(x, ct, eb, chunk, dev, subtract_min=True, crc=True, fixed=None):
    xd = torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    kw = {}
    if fixed is not None:
        kw = dict(fixed_type=fixed[0], fixed_mask17=fixed[1])
    comp = api.compress(xd, ct=ct, abs_error_bound=eb, chunk_elems=chunk, subtract_min=subtract_min, crc=crc, **kw)
    torch.cuda.synchronize()
    return comp


def assert_same_as_oracle(x, ct, eb, chunk, dev, subtract_min=True, fixed=None, decode=True, tag=""):
    n = x.size
    comp = gpu_encode(x, ct, eb, chunk, dev, subtract_min=subtract_min, fixed=fixed)
    ochunk = chunk if (chunk and chunk < n) else max(n, 1)
    ft, fm = (fixed if fixed is not None else (-1, 0))
    enc = ol.o_compress_chunked(ct, x, ochunk, eb, subtract_min=subtract_min, fixed_type=ft, fixed_mask17=fm,
                                want_crc=True)
    K = len(enc["offsets"]) - 1
    offs = comp.chunk_offsets.cpu().numpy().astype(np.uint64)
    assert np.array_equal(offs, enc["offsets"]), (tag, "offsets")
    total = int(offs[-1])
    s = comp.stream[:total].cpu().numpy()
    if not np.array_equal(s, enc["stream"]):
        bad = np.nonzero(s != enc["stream"])[0]
        raise AssertionError("%s: stream differs at %d bytes, first at %d (chunk %d)" %
                             (tag, bad.size, bad[0], np.searchsorted(offs, bad[0], side="right") - 1))
    if subtract_min:
        assert np.array_equal(comp.chunk_min[:K].cpu().numpy().view(np.uint32), enc["chunk_min"].view(np.uint32)), tag
    if ct == 7 and fixed is None:
        assert np.array_equal(comp.chunk_medium[:K].cpu().numpy().view(np.uint32) >> 15,
                              enc["chunk_medium"].view(np.uint32) >> 15), (tag, "medium: only the 17 mask bits are defined")
        assert np.array_equal(comp.chunk_type[:K].cpu().numpy(), enc["chunk_type"]), (tag, "type")
    assert np.array_equal(comp.chunk_crc[:K].cpu().numpy().view(np.uint32), enc["chunk_crc"]), (tag, "crc")
    if decode:
        d = api.decompress(comp).cpu().numpy()
        od = ol.o_decompress_chunked(ct, enc, n, ochunk, eb, add_min=subtract_min, fixed_type=ft, fixed_mask17=fm)
        assert np.array_equal(d.view(np.uint32), od.view(np.uint32)), (tag, "decoded")
    return comp


@pytest.mark.parametrize("kind", ("smooth", "noisy", "constant"))
@pytest.mark.parametrize("ct", CTS)
def test_fields_chunked(kind, ct, dev):
    for n in (20000, 65536):
        x = make_field(kind, n)
        for eb in EBS:
            assert_same_as_oracle(x, ct, eb, 4096, dev, tag="%s n=%d ct=%d eb=%g" % (kind, n, ct, eb))


@pytest.mark.parametrize("ct", CTS)
def test_single_chunk_legacy_shapes(ct, dev):
    """chunk_elems=0: the whole array is ONE reference call (pingpong 131,072; a Himeno face 16,641; tiny n)."""
    rng = np.random.default_rng(5)
    for n in (1, 2, 3, 4, 5, 17, 100, 4095, 4096, 4097, 16641, 131072):
        x = (rng.random(n) * 10).astype(np.float32)
        if n > 1000:
            x[n // 3: n // 3 + 500] = make_field("smooth", 500)
        for eb in (0.01, 1e-6):
            assert_same_as_oracle(x, ct, eb, 0, dev, tag="single n=%d ct=%d eb=%g" % (n, ct, eb))


@pytest.mark.parametrize("ct", CTS)
def test_multi_tile_chunks(ct, dev):
    rng = np.random.default_rng(6)
    x = np.concatenate([make_field("smooth", 30000), (rng.random(21000) * 10).astype(np.float32)])
    for chunk in (8192, 16384):
        assert_same_as_oracle(x, ct, 1e-3, chunk, dev, tag="chunk=%d ct=%d" % (chunk, ct))


@pytest.mark.parametrize("ct", CTS)
def test_no_min_subtract_and_fixed_mask(ct, dev):
    rng = np.random.default_rng(8)
    x = (rng.random(30000) * 7).astype(np.float32)
    assert_same_as_oracle(x, ct, 0.01, 4096, dev, subtract_min=False, tag="nomin ct=%d" % ct)
    if ct == 7:
        fixed = (2, api.mask17_of(np.float32(3.3)))
        assert_same_as_oracle(x, ct, 0.01, 4096, dev, subtract_min=False, fixed=fixed, tag="fixedmask")
        assert_same_as_oracle(x, ct, 1e-5, 0, dev, subtract_min=True, fixed=fixed, tag="fixedmask single")


@pytest.mark.parametrize("ct", CTS)
def test_encoder_only_inputs(ct, dev):
    """Negative values, the -1.0f history sentinel, Inf/NaN, denormals, zeros of both signs:
    not decodable by the reference format, but the ENCODING is defined and must match bit for bit."""
    rng = np.random.default_rng(9)
    n = 12000
    neg = (rng.random(n) * 8 - 4).astype(np.float32)
    sent = (rng.random(n) * 3).astype(np.float32)
    sent[[0, 5, 6, 7, 100, 101, 300, 4095, 4096, 4097, 8191, n - 1]] = -1.0
    sent2 = np.full(n, -1.0, np.float32)
    sent2[::7] = 2.5
    spec = (rng.random(n) * 3).astype(np.float32)
    spec[10], spec[20], spec[30], spec[5000] = np.inf, np.nan, -np.inf, np.nan
    den = (rng.random(n).astype(np.float32) * np.float32(1e-40))
    zer = (rng.random(n) * 2).astype(np.float32)
    zer[::5] = 0.0
    zer[3::10] = -0.0
    fixed = (2, api.mask17_of(np.float32(1.7))) if ct == 7 else None
    for name, y in (("neg", neg), ("sentinel", sent), ("sentinel2", sent2), ("special", spec), ("denormal", den),
                    ("zeros", zer)):
        for chunk in (4096, 0):
            assert_same_as_oracle(y, ct, 0.01, chunk, dev, subtract_min=False, fixed=fixed, decode=False,
                                  tag="%s chunk=%d ct=%d" % (name, chunk, ct))
    # min-subtract corner case: the FIRST of several equal (+0 / -0) minima decides the sign of min (:3546-3553)
    mz = zer.copy(); mz[0] = 1.0
    if ct != 7:
        for chunk in (4096, 0):
            assert_same_as_oracle(mz, ct, 0.01, chunk, dev, subtract_min=True, decode=False,
                                  tag="minus-zero-min chunk=%d ct=%d" % (chunk, ct))
    # data[0] = NaN: every '<' fails and min stays NaN.  (The stream is not compared: x - NaN keeps the NaN
    # payload on x86 but is canonicalised to 0x7fffffff by the GPU; NaN is outside the format's domain.)
    nanfirst = spec.copy(); nanfirst[0] = np.nan
    comp = gpu_encode(nanfirst, 5 if ct == 7 else ct, 0.01, 4096, dev)
    assert np.isnan(comp.chunk_min[0].item()) and not np.isnan(comp.chunk_min[1].item())


def test_exact_medium_transducer_sum(dev):
    """DCB_F_EXACT_MEDIUM: med_dataset_float's sequential float32 mean, bit for bit, from the parallel
    binade-transducer sum (dcb_encode.cuh) -- on data that crosses many binades, has ties, zeros and tiny values."""
    rng = np.random.default_rng(31)
    n = 4096 * 24
    fields = {
        "noisy": make_field("noisy", n), "smooth": make_field("smooth", n), "constant": make_field("constant", n),
        "lognormal": np.exp(rng.normal(0, 4, n)).astype(np.float32),
        "small_ints": rng.integers(0, 5, n).astype(np.float32),                 # exact ties at every binade
        "halves": (rng.integers(0, 64, n) * 0.5).astype(np.float32),
        "sparse": np.where(rng.random(n) < 0.05, rng.random(n) * 1e6, 0).astype(np.float32),
        "growing": (np.arange(n) % 4096).astype(np.float32) ** 2,
        "tiny_then_big": np.where((np.arange(n) % 4096) < 2048, 1e-30, 3.0).astype(np.float32),
    }
    for name, x in fields.items():
        for sub in (True, False):
            xd = torch.from_numpy(x).to(dev)
            comp = api.compress(xd, ct=7, abs_error_bound=0.01, chunk_elems=4096, subtract_min=sub, exact_medium=True)
            torch.cuda.synchronize()
            enc = ol.o_compress_chunked(7, x, 4096, 0.01, subtract_min=sub)
            K = len(enc["offsets"]) - 1
            assert np.array_equal(comp.chunk_medium[:K].cpu().numpy().view(np.uint32), enc["chunk_medium"].view(np.uint32)), (name, sub)
            assert np.array_equal(comp.chunk_type[:K].cpu().numpy(), enc["chunk_type"]), (name, sub)
            assert np.array_equal(comp.stream[:comp.total_bytes()].cpu().numpy(), enc["stream"]), (name, sub)


def test_golden_fixtures(dev, golden_dir):
    """Committed outputs of the UNMODIFIED reference (tests/golden/make_golden.py)."""
    d = np.load(os.path.join(golden_dir, "datasets.npz"))
    for row in d["known_answers"]:
        name, eb, ct = str(row["name"]), float(row["eb"]), int(row["ct"])
        x = d[name]
        comp = gpu_encode(x, ct, eb, 0, dev)
        total = comp.total_bytes()
        assert total == row["bytes"], (name, eb, ct)
        assert np.array_equal(comp.stream[:total].cpu().numpy(), d["stream_%s_%g_%d" % (name, eb, ct)])
        assert np.uint32(comp.chunk_crc[0].item() & 0xFFFFFFFF) == row["crc32"]
        assert comp.chunk_min[0].item() == row["min"]
        if ct == 7:
            assert (comp.chunk_type[0].item(), api.mask17_of(comp.chunk_medium[0].item())) == (row["type"], row["mask17"])
        key = "decoded_%s_%g_%d" % (name, eb, ct)
        if key in d:
            # the reference tools add min back after decoding (tools/mycompress_bitwise_float.c:60)
            out = api.decompress(comp).cpu().numpy()
            want = (d[key] + np.float32(row["min"])).astype(np.float32)
            assert np.array_equal(out.view(np.uint32), want.view(np.uint32)), key
    g = np.load(os.path.join(golden_dir, "chunked.npz"))
    for kind in ("smooth", "noisy", "constant"):
        x = make_field(kind, 20000)
        for eb in (0.01, 1e-4):
            for ct in CTS:
                key = "%s_%g_%d_" % (kind, eb, ct)
                comp = gpu_encode(x, ct, eb, 4096, dev)
                assert np.array_equal(comp.chunk_offsets.cpu().numpy().astype(np.uint64), g[key + "offsets"]), key
                assert np.array_equal(comp.stream[:comp.total_bytes()].cpu().numpy(), g[key + "stream"]), key
                assert np.array_equal(comp.chunk_crc[:5].cpu().numpy().view(np.uint32), g[key + "chunk_crc"]), key
                if key + "decoded" in g:
                    out = api.decompress(comp).cpu().numpy()
                    assert np.array_equal(out.view(np.uint32), g[key + "decoded"].view(np.uint32)), key


def test_host_pointer_layer(dev):
    rng = np.random.default_rng(11)
    x = (rng.random(50000) * 10).astype(np.float32)
    for ct in (5, 7, 8, 9):
        enc = api.compress_host(x, ct=ct, abs_error_bound=0.01, chunk_elems=4096)
        base = 5 if ct in (5, 8) else 7
        o = ol.o_compress_chunked(base, x, 4096, 0.01, want_crc=True)
        assert np.array_equal(enc["stream"][:enc["total_bytes"]], o["stream"])
        if ct in (8, 9):
            assert np.array_equal(enc["chunk_crc"], o["chunk_crc"])
        out = api.decompress_host(enc)
        assert np.array_equal(out.view(np.uint32), ol.o_decompress_chunked(base, o, x.size, 4096, 0.01).view(np.uint32))


def test_crc_detects_bit_flip(dev):
    """CT=8 protocol of impl/pingpong.c:363-410: flip bits in the received stream, recompute, compare."""
    x = make_field("noisy", 40000)
    comp = gpu_encode(x, 8, 0.01, 4096, dev)
    before = comp.chunk_crc[:comp.num_chunks].clone()
    api.flip_bits(comp.stream, [3, 8 * int(comp.chunk_offsets[2].item()) + 5])
    after = api.crc32_chunks(comp.stream, comp.chunk_offsets)
    torch.cuda.synchronize()
    diff = (before != after).cpu().numpy()
    assert list(np.nonzero(diff)[0]) == [0, 2]


def test_malformed_ct6_stream_is_reported(dev):
    x = make_field("noisy", 8192)
    comp = gpu_encode(x, 6, 0.01, 4096, dev)
    api.flip_bits(comp.stream, [0])                      # leading 1: "Error leading bit 1", dataCompression.c:2501
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    api.decompress(comp, status=status)
    torch.cuda.synchronize()
    assert status.item() == -5


def test_error_bound_and_reference_drift(dev):
    """north_star (3): |x - x'| <= absErrorBound is a hard gate only for CT=6; CT=5/7/11 predict from ORIGINAL
    values in the encoder and RECONSTRUCTED values in the decoder, so the reference itself drifts.
    The drift must be the reference's own: max error equals the oracle's, bit for bit."""
    for kind in ("smooth", "noisy"):
        x = make_field(kind, 65536)
        for eb in (0.01, 1e-4):
            c6 = gpu_encode(x, 6, eb, 4096, dev)
            err6 = np.abs(api.decompress(c6).cpu().numpy().astype(np.float64) - x.astype(np.float64)).max()
            assert err6 <= eb * (1 + 1e-6) + 1e-6 * np.abs(x).max(), (kind, eb, err6)   # +min rounding slack
            c5 = gpu_encode(x, 5, eb, 4096, dev)
            d5 = api.decompress(c5).cpu().numpy()
            o5 = ol.o_decompress_chunked(5, ol.o_compress_chunked(5, x, 4096, eb), x.size, 4096, eb)
            assert np.abs(d5 - x).max() == np.abs(o5 - x).max()


@pytest.mark.parametrize("ct,kind", ((7, "smooth"), (5, "noisy")))
def test_full_size_properties(ct, kind, dev):
    """BASELINE.json configs[1] size (2^28 elements, 65,536 chunks): what can be checked without running the CPU oracle
    over 1 GiB.  (a) the offset table is monotone, byte aligned per chunk and ends at the stream length; (b) 96 chunks
    sampled over the whole range -- first, last, random -- are bit-exact against the oracle run on those chunks alone
    (chunks are independent: SURVEY 8e), stream bytes and decoded floats; (c) the CRC table computed on the GPU is a
    checksum of every chunk's bytes: recomputed after a decode -> re-encode of the ORIGINAL input it must not move
    (determinism: scanner/ticket order differs from launch to launch); (d) decoding never flags a malformed stream."""
    from data_compression_b200.fields import make_field_torch
    n, chunk, eb = 1 << 28, 4096, 0.01
    x = make_field_torch(kind, n, dev)
    comp = api.compress(x, ct=ct, abs_error_bound=eb, chunk_elems=chunk, crc=True)
    K = comp.num_chunks
    assert K == n // chunk
    offs = comp.chunk_offsets[:K + 1]
    assert int(offs[0].item()) == 0 and bool((offs[1:] > offs[:-1]).all().item())
    total = comp.total_bytes()
    assert int(offs[K].item()) == total and total <= api.lib().dcb_max_stream_bytes(n, chunk)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    out = api.decompress(comp, status=status)
    torch.cuda.synchronize()
    assert status.item() == 0
    rng = np.random.default_rng(5)
    picks = sorted(set([0, 1, K // 2, K - 2, K - 1] + [int(v) for v in rng.integers(0, K, 91)]))
    offs_h = offs.cpu().numpy().astype(np.int64)
    for c in picks:
        xc = x[c * chunk:(c + 1) * chunk].cpu().numpy()
        enc = ol.o_compress_chunked(ct, xc, chunk, eb, want_crc=True)
        got = comp.stream[offs_h[c]:offs_h[c + 1]].cpu().numpy()
        assert np.array_equal(got, enc["stream"]), (kind, ct, "chunk", c)
        assert int(comp.chunk_crc[c].item()) & 0xFFFFFFFF == int(enc["chunk_crc"][0]), (kind, ct, "crc", c)
        dec = ol.o_decompress_chunked(ct, enc, chunk, chunk, eb)
        assert np.array_equal(out[c * chunk:(c + 1) * chunk].cpu().numpy().view(np.uint32), dec.view(np.uint32)), (kind, ct, "decoded", c)
    crc1 = comp.chunk_crc[:K].clone()
    tot1 = total
    comp2 = api.compress(x, ct=ct, abs_error_bound=eb, chunk_elems=chunk, crc=True)
    torch.cuda.synchronize()
    assert comp2.total_bytes() == tot1 and torch.equal(comp2.chunk_crc[:K], crc1)
    assert torch.equal(comp2.chunk_offsets[:K + 1], offs)


@pytest.mark.parametrize("ct", CTS)
def test_one_tile_chunks_ragged_unaligned(ct, dev):
    """One chunk == one tile (the slot path: dcb_encode_warp / dcb_encode_slots + dcb_compact): a ragged last chunk, an input
    pointer that is only 4-byte aligned, sizes around the group size of the compaction kernel (64 chunks)."""
    rng = np.random.default_rng(11 + ct)
    for n in (4096 * 3 + 17, 4096 * 64, 4096 * 65 + 511, 4096 * 130 + 2049):
        x = np.concatenate([make_field("smooth", n // 2), make_field("noisy", n - n // 2)]).astype(np.float32)
        x[rng.integers(0, n, 50)] = 0.0
        assert_same_as_oracle(x, ct, 0.01, 4096, dev, tag="ragged n=%d ct=%d" % (n, ct))
    # unaligned device pointer: a view that starts one float into the allocation
    n = 4096 * 9 + 100
    base = torch.from_numpy(np.concatenate([[0.0], make_field("noisy", n)]).astype(np.float32)).to(dev)
    xd = base[1:]
    assert xd.data_ptr() % 16 == 4
    comp = api.compress(xd, ct=ct, abs_error_bound=1e-4, chunk_elems=4096, crc=True)
    torch.cuda.synchronize()
    x = xd.cpu().numpy()
    enc = ol.o_compress_chunked(ct, x, 4096, 1e-4, want_crc=True)
    total = comp.total_bytes()
    assert total == enc["stream"].size
    assert np.array_equal(comp.stream[:total].cpu().numpy(), enc["stream"])
    d = api.decompress(comp).cpu().numpy()
    od = ol.o_decompress_chunked(ct, enc, n, 4096, 1e-4)
    assert np.array_equal(d.view(np.uint32), od.view(np.uint32))


@pytest.mark.parametrize("encoder", ("warp", "slots", "single"))
def test_every_encoder_kernel(encoder):
    """The library picks one of three encoders for one-tile chunks (dcb_abi.cu encoder_choice).  DCB_ENCODER forces one for
    every CT; it is read once per process, so the parity tests of this module are rerun in a child process."""
    import subprocess
    import sys
    if os.environ.get("DCB_ENCODER"):
        pytest.skip("already inside a forced-encoder run")
    env = dict(os.environ, DCB_ENCODER=encoder)
    here = os.path.abspath(__file__)
    sel = "test_fields_chunked or test_no_min_subtract_and_fixed_mask or test_encoder_only_inputs or " \
          "test_exact_medium_transducer_sum or test_golden_fixtures or test_one_tile_chunks_ragged_unaligned"
    r = subprocess.run([sys.executable, "-m", "pytest", here, "-x", "-q", "-m", "gpu", "-k", sel, "-p", "no:cacheprovider"],
                       env=env, capture_output=True, text=True, cwd=os.path.dirname(os.path.dirname(here)))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


def test_ct7_mask_decision_at_scale(dev):
    """The CT=7 mask comes from a BOUND on the reference's sequential float32 sum whenever the bound pins its 17 bits (DESIGN.md
    4.3), from the exact transducer sum otherwise.  4 x 1024 chunks of differently distributed data: every chunk's mask and type
    must equal med_dataset_float's (computed by the oracle on the min-subtracted chunk), under the default encoder."""
    rng = np.random.default_rng(77)
    n, chunk = 1 << 22, 4096
    fields = {
        "lognormal": rng.lognormal(0.0, 2.0, n),
        "offset": 1000.0 + rng.uniform(0, 1e-2, n),                 # large minimum, tiny spread: cancellation in the raw sums
        "tiny": rng.uniform(0, 1e-30, n),
        "steps": np.repeat(rng.uniform(0, 100, n // 512), 512) + rng.normal(0, 1e-3, n),
    }
    for name, xf in fields.items():
        x = xf.astype(np.float32)
        comp = api.compress(torch.from_numpy(x).to(dev), ct=7, abs_error_bound=1e-3, chunk_elems=chunk)
        torch.cuda.synchronize()
        med = comp.chunk_medium[:n // chunk].cpu().numpy().view(np.uint32) >> 15
        typ = comp.chunk_type[:n // chunk].cpu().numpy()
        for c in range(n // chunk):
            seg = x[c * chunk:(c + 1) * chunk]
            _, y = ol.o_min_subtract(seg)
            m, t = ol.o_med(y, 0)
            assert int(med[c]) == ol.mask17_of(m) and int(typ[c]) == t, (name, c, hex(int(med[c])), hex(ol.mask17_of(m)))


@pytest.mark.parametrize("seed", (1, 2, 3))
def test_fuzz_mixed_fields(seed, dev):
    """Seeded mixtures of everything a chunk can hold -- smooth, noisy, constant, ramps, zeros, denormals, values near the
    error bound, lognormal magnitudes -- at ragged sizes, every CT, random error bounds: stream, tables, CRCs and decoded floats
    against the oracle (min-subtract on: values stay in the format's domain)."""
    rng = np.random.default_rng(1000 + seed)
    pieces = []
    for _ in range(40):
        m = int(rng.integers(100, 9000))
        kind = int(rng.integers(0, 8))
        if kind == 0:
            p = make_field("smooth", m, start=int(rng.integers(0, 60000)))
        elif kind == 1:
            p = make_field("noisy", m, start=int(rng.integers(0, 1 << 20)))
        elif kind == 2:
            p = np.full(m, rng.uniform(0, 50), np.float32)
        elif kind == 3:
            p = (np.arange(m) * rng.uniform(1e-4, 1e-1) + rng.uniform(0, 10)).astype(np.float32)
        elif kind == 4:
            p = np.zeros(m, np.float32)
        elif kind == 5:
            p = (rng.random(m) * 1e-39).astype(np.float32)
        elif kind == 6:
            p = (rng.uniform(0, 3, m) * rng.choice([1e-3, 1e-2, 1e-1])).astype(np.float32)
        else:
            p = (rng.lognormal(0, 3, m)).astype(np.float32)
        pieces.append(p.astype(np.float32))
    x = np.concatenate(pieces)
    for ct in CTS:
        eb = float(rng.choice(EBS))
        assert_same_as_oracle(x, ct, eb, 4096, dev, tag="fuzz seed=%d ct=%d eb=%g n=%d" % (seed, ct, eb, x.size))


def test_stream_offsets_beyond_4_gib(dev):
    """Just above BASELINE.json configs[2]'s size on ONE GPU (2^30 + 2^24 elements, 266,240 chunks) with the least compressible
    input of the sweep (CT=11 on the noisy field at 1e-6: 32 bits per element, the fallback is the float itself): the stream is
    longer than 2^32 bytes, so the offsets of the last chunks need their upper word.  Offsets monotone and ending at the stream length; chunks sampled on both sides
    of the 4 GiB line and at the very end bit-exact against the oracle (stream, CRC, decoded floats); nothing flagged."""
    from data_compression_b200.fields import make_field_torch
    n, chunk, eb, ct = (1 << 30) + (1 << 24), 4096, 1e-6, 11
    free, _ = torch.cuda.mem_get_info(dev)
    if free < 30 * (1 << 30):
        pytest.skip("needs 30 GiB of free device memory")
    x = make_field_torch("noisy", n, dev)
    comp = api.compress(x, ct=ct, abs_error_bound=eb, chunk_elems=chunk, crc=True)
    K = comp.num_chunks
    assert K == n // chunk
    offs = comp.chunk_offsets[:K + 1]
    total = comp.total_bytes()
    assert total > (1 << 32) and int(offs[K].item()) == total and int(offs[0].item()) == 0
    assert bool((offs[1:] > offs[:-1]).all().item())
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    out = api.decompress(comp, status=status)
    torch.cuda.synchronize()
    assert status.item() == 0
    offs_h = offs.cpu().numpy().astype(np.int64)
    line = int(np.searchsorted(offs_h, 1 << 32))                  # first chunk that ends beyond 4 GiB
    rng = np.random.default_rng(9)
    picks = sorted(set([0, line - 2, line - 1, line, line + 1, K - 2, K - 1] + [int(v) for v in rng.integers(line, K, 24)]))
    for c in picks:
        xc = x[c * chunk:(c + 1) * chunk].cpu().numpy()
        enc = ol.o_compress_chunked(ct, xc, chunk, eb, want_crc=True)
        got = comp.stream[offs_h[c]:offs_h[c + 1]].cpu().numpy()
        assert np.array_equal(got, enc["stream"]), ("chunk", c)
        assert int(comp.chunk_crc[c].item()) & 0xFFFFFFFF == int(enc["chunk_crc"][0]), ("crc", c)
        dec = ol.o_decompress_chunked(ct, enc, chunk, chunk, eb)
        assert np.array_equal(out[c * chunk:(c + 1) * chunk].cpu().numpy().view(np.uint32), dec.view(np.uint32)), ("decoded", c)
    # CT=11 stores what it cannot predict in full; (x - min) + min re-rounds, nothing more: one ulp of 10 plus the bound
    worst = float((out - x).abs_().max().item())
    assert worst <= 1e-6 + 2 * 9.5367431640625e-07, worst
    del out, comp, x
    torch.cuda.empty_cache()


def test_host_calls_with_large_pageable_buffers(dev):
    """dcb_compress_host / dcb_decompress_host on plain numpy (malloc'd, pageable) arrays of more than 8 MB: bulk copies take the
    staged path (four helpers, pinned staging, csrc/dcb_abi.cu copy_staged) -- sizes that are not a multiple of the 4 MB slice,
    results identical to the device-resident calls; twice in a row (staging reused), and from two threads at once."""
    import threading
    n = 5 * (1 << 20) + 12345                                 # 20 MB + a ragged tail
    x = np.concatenate([make_field("smooth", n // 2), make_field("noisy", n - n // 2)]).astype(np.float32)
    comp = api.compress(torch.from_numpy(x).to(dev), ct=5, abs_error_bound=0.01, chunk_elems=4096, crc=True)
    total = comp.total_bytes()
    want_stream = comp.stream[:total].cpu().numpy()
    want_out = api.decompress(comp).cpu().numpy()
    for rep in range(2):
        enc = api.compress_host(x, ct=5, abs_error_bound=0.01, chunk_elems=4096, crc=True)
        assert enc["total_bytes"] == total
        assert np.array_equal(enc["stream"][:total], want_stream)
        assert np.array_equal(enc["chunk_crc"], comp.chunk_crc[:comp.num_chunks].cpu().numpy().view(np.uint32))
        out = api.decompress_host(enc)
        assert np.array_equal(out.view(np.uint32), want_out.view(np.uint32))
    errs = []

    def worker():
        try:
            e = api.compress_host(x, ct=5, abs_error_bound=0.01, chunk_elems=4096)
            o = api.decompress_host(e)
            if not np.array_equal(o.view(np.uint32), want_out.view(np.uint32)) or e["total_bytes"] != total:
                errs.append("mismatch")
        except Exception as ex:                               # noqa: BLE001
            errs.append(repr(ex))
    ts = [threading.Thread(target=worker) for _ in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs


def test_chunked_decode_rejects_chunks_beyond_the_lane_readers_range(dev):
    """A decoding lane keeps its bit position in 32 bits: dcb_decompress refuses chunked streams with more than 2^26 elements per
    chunk (DCB_EINVAL) instead of wrapping; argument checks come before any memory is touched."""
    import ctypes as C
    d = torch.zeros(64, dtype=torch.uint8, device=dev)
    p = C.c_void_p(d.data_ptr())
    rc = api.lib().dcb_decompress(5, p, p, (1 << 28) + 1, 1 << 27, 0.01, 0, 0, 0, None, None, None, p, None,
                                  C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    assert rc == -1, rc
