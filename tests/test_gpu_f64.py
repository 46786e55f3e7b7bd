"""float64 codecs on the GPU against the float64 oracle (oracle/dc_oracle_f64.c, pinned against the
compiled reference by tests/test_oracle_f64.py): streams, chunk tables and decoded values bit for bit."""
import os

import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "f64_golden.npz")


def _fields(rng, n):
    yield "walk", np.cumsum(rng.standard_normal(n)) * 0.02 + 20.0
    yield "uniform", rng.random(n) * 10.0
    yield "sine", 1 + 0.5 * np.sin(np.arange(n) / 40.0)
    yield "centres", np.sort(rng.random(n) * 1000.0)
    yield "zeros", np.zeros(n)
    yield "big", rng.random(n) * 1e200


def _check(x, ct, eb, chunk, subtract_min=True, crc=False, fixed=None, decode=True):
    import torch
    from data_compression_b200 import api
    dev = torch.device("cuda:0")
    dx = torch.from_numpy(x).to(dev)
    kw = {}
    okw = {}
    if fixed is not None:
        kw = dict(fixed_type=fixed[0], fixed_mask20=fixed[1]); okw = dict(fixed_type=fixed[0], fixed_mask20=fixed[1])
    comp = api.compress_f64(dx, ct=ct, abs_error_bound=eb, chunk_elems=chunk, subtract_min=subtract_min, crc=crc, **kw)
    och = chunk if (chunk and chunk < x.size) else max(x.size, 1)
    enc = ol.o_compress_chunked_f64(ct, x, och, eb, subtract_min=subtract_min, want_crc=crc, **okw)
    K = enc["offsets"].size - 1
    offs = comp.chunk_offsets.cpu().numpy().astype(np.uint64)
    assert np.array_equal(offs, enc["offsets"]), "offsets"
    tot = int(offs[-1])
    assert np.array_equal(comp.stream[:tot].cpu().numpy(), enc["stream"]), "stream"
    if subtract_min:
        assert np.array_equal(comp.chunk_min.cpu().numpy()[:K].view(np.uint64), enc["chunk_min"].view(np.uint64))
    if ct == 7 and fixed is None:
        assert np.array_equal(comp.chunk_mask20.cpu().numpy()[:K].astype(np.uint32), enc["chunk_mask20"])
        assert np.array_equal(comp.chunk_type.cpu().numpy()[:K], enc["chunk_type"])
    if crc:
        assert np.array_equal(comp.chunk_crc.cpu().numpy()[:K].view(np.uint32), enc["chunk_crc"])
    if decode:
        if fixed is not None:
            enc["chunk_mask20"] = np.full(K, fixed[1], np.uint32); enc["chunk_type"] = np.full(K, fixed[0], np.int32)
        want = ol.o_decompress_chunked_f64(ct, enc, x.size, och, eb, add_min=subtract_min, **okw)
        got = api.decompress_f64(comp).cpu().numpy()
        assert np.array_equal(got.view(np.uint64), want.view(np.uint64)), "decode"
    return comp


@pytest.mark.parametrize("ct", [5, 6, 7, 11])
@pytest.mark.parametrize("eb", [0.01, 1e-6])
def test_chunked_parity(ct, eb):
    rng = np.random.default_rng(100 + ct)
    for name, x in _fields(rng, 3 * 4096 + 777):
        _check(x, ct, eb, 4096, crc=(ct in (5, 7)))


@pytest.mark.parametrize("ct", [5, 7])
def test_single_chunk_and_multi_tile_chunks(ct):
    rng = np.random.default_rng(5)
    for n in (1, 2, 3, 4, 100, 4095, 4096, 4097, 20000):
        x = np.cumsum(rng.standard_normal(n)) * 0.05 + 3.0
        _check(x, ct, 1e-3, 0)                       # one reference call over the whole array (k-means / mm / lu usage)
    x = np.cumsum(rng.standard_normal(5 * 8192 + 100)) * 0.05 + 3.0
    _check(x, ct, 1e-3, 8192)                        # two tiles per chunk: tile boundaries on arbitrary bits


@pytest.mark.parametrize("ct", [5, 6, 7, 11])
def test_encoder_only_inputs(ct):
    """Negative values, the -1.0 marker (also across thread / tile boundaries), NaN / Inf, signed zeros: the stream is defined."""
    rng = np.random.default_rng(9)
    n = 2 * 4096 + 50
    neg = rng.random(n) * 8 - 4
    sent = rng.random(n) * 3
    sent[[0, 5, 6, 7, 15, 16, 17, 31, 32, 100, 101, 300, 4095, 4096, 4097, 4111, 4112, n - 1]] = -1.0
    sent2 = np.full(n, -1.0); sent2[::7] = 2.5
    sent3 = np.full(n, -1.0); sent3[5000] = 1.0
    spec = rng.random(n) * 3
    spec[10], spec[20], spec[30], spec[4096] = np.inf, np.nan, -np.inf, np.nan
    zs = rng.random(n); zs[::5] = 0.0; zs[1::5] = -0.0
    for name, y in (("neg", neg), ("sentinel", sent), ("sentinel2", sent2), ("sentinel3", sent3), ("special", spec), ("zeros", zs)):
        for chunk in (4096, 0):
            _check(y, ct, 0.01, chunk, subtract_min=False, fixed=(2, ol.mask20_of(1.7)) if ct == 7 else None, decode=False)
    nanfirst = rng.random(n); nanfirst[0] = np.nan
    _check(nanfirst, 5, 0.01, 4096, subtract_min=True, decode=False)
    _check(zs - 0.0, ct, 0.01, 4096, subtract_min=True, decode=False)


def test_forced_masks():
    rng = np.random.default_rng(11)
    y = rng.random(9000) * 6
    for t in (0, 1, 2, 5, 9, 10):
        for medium in (1.7, 3.3, 0.004, 700.0):
            _check(y, 7, 1e-5, 4096, subtract_min=False, fixed=(t, ol.mask20_of(medium)))


def test_exact_medium_and_bracket_fallback():
    """The 20 mask bits must be those of the SEQUENTIAL sum, also where the tree sum's bracket straddles a mask boundary."""
    import torch
    from data_compression_b200 import api
    rng = np.random.default_rng(21)
    n = 64 * 4096
    x = rng.random(n)
    for c in range(0, 64, 2):                        # chunks whose mean sits right on a 2^-8 mantissa step of the mask
        seg = x[c * 4096:(c + 1) * 4096]
        target = 0.5 + (c + 1) / 256.0 * 0.5
        seg *= target / seg.mean()
        seg[0] -= seg.min()                          # keep min == 0 at index 0 so that the min-subtract leaves the values alone
        seg[0] = 0.0
    comp = api.compress_f64(torch.from_numpy(x).cuda(), ct=7, abs_error_bound=1e-4, chunk_elems=4096, exact_medium=True)
    enc = ol.o_compress_chunked_f64(7, x, 4096, 1e-4)
    assert np.array_equal(comp.chunk_medium.cpu().numpy().view(np.uint64), enc["chunk_medium"].view(np.uint64))
    _check(x, 7, 1e-4, 4096)


def test_golden_fixture():
    import torch
    from data_compression_b200 import api
    g = np.load(GOLD)
    dev = torch.device("cuda:0")
    for key in [k for k in g.files if k.endswith("_x")]:
        tag = key[:-2]
        x, eb, ct = g[key], float(g[tag + "_eb"]), int(g[tag + "_ct"])
        t, m20 = int(g[tag + "_type"]), int(g[tag + "_mask20"])
        comp = api.compress_f64(torch.from_numpy(x).to(dev), ct=ct, abs_error_bound=eb, chunk_elems=0, subtract_min=False,
                                fixed_type=t if ct == 7 else None, fixed_mask20=m20 if ct == 7 else None)
        tot = comp.total_bytes()
        assert np.array_equal(comp.stream[:tot].cpu().numpy(), g[tag + "_stream"]), tag
        if (tag + "_decoded") in g.files:
            got = api.decompress_f64(comp).cpu().numpy()
            assert np.array_equal(got.view(np.uint64), g[tag + "_decoded"].view(np.uint64)), tag


def test_np_rejects_leading_one():
    import torch
    from data_compression_b200 import api
    x = torch.rand(4096, dtype=torch.float64, device="cuda:0")
    comp = api.compress_f64(x, ct=6, abs_error_bound=0.01, chunk_elems=4096, subtract_min=False)
    comp.stream[0] |= 0x80
    with pytest.raises(api.DcbError):
        api.decompress_f64(comp)


@pytest.mark.parametrize("ct", [5, 7])
def test_long_single_chunk_statistics(ct):
    """ONE chunk of >= 65,536 doubles takes the multi-block statistics (csrc/dcb_f64_stats_long.cuh): minimum with the reference's
    first-occurrence rule, type / mask from max and mean, and the exact sequential mean when the bracket straddles a mask
    boundary -- streams, min, mask and type bit for bit as the oracle's, on the shapes that exercise each rule."""
    import torch
    from data_compression_b200 import api
    rng = np.random.default_rng(77)
    n = 300001
    sine = 5.0 + 4.5 * np.sin(2 * np.pi * (np.arange(1 << 18) % 65536) / 65536.0)      # mean - min = 4.5 exactly: a mask boundary
    cases = {
        "sine_boundary": sine,
        "uniform": rng.random(n) * 10.0,
        "walk": np.cumsum(rng.standard_normal(n)) * 0.02 + 100.0,
        "min_is_last": np.concatenate([rng.random(n - 1) + 1.0, [0.5]]),
        "min_is_first": np.concatenate([[0.25], rng.random(n - 1) + 1.0]),
        "two_equal_minima": np.where(np.arange(n) % 100000 == 7, 0.125, rng.random(n) + 1.0),
        "zeros_pos_then_neg": np.where(np.arange(n) == 70000, 0.0, np.where(np.arange(n) == 200000, -0.0, rng.random(n) + 1.0)),
        "zeros_neg_then_pos": np.where(np.arange(n) == 70000, -0.0, np.where(np.arange(n) == 200000, 0.0, rng.random(n) + 1.0)),
        "nan_inside": np.where(np.arange(n) == 123456, np.nan, rng.random(n) * 3.0),
        "big": rng.random(n) * 1e200,
        "constant": np.full(n, 0.123456789),
    }
    for name, x in cases.items():
        for sub in (True, False):
            try:
                _check(x, ct, 1e-3, 0, subtract_min=sub, decode=(name != "nan_inside"))
            except AssertionError as e:
                raise AssertionError("%s subtract_min=%s: %s" % (name, sub, e))
    # the exact mean itself (DCB_F_EXACT_MEDIUM), against the reference's sequential sum restated in numpy
    dev = torch.device("cuda:0")
    for name in ("sine_boundary", "uniform", "walk"):
        x = cases[name]
        comp = api.compress_f64(torch.from_numpy(x).to(dev), ct=7, abs_error_bound=1e-3, chunk_elems=0, exact_medium=True)
        y = x - x.min()
        total = 0.0
        for v in y.tolist():
            total += v
        assert comp.chunk_medium[0].item() == total / y.size, name


@pytest.mark.parametrize("ct", [5, 6, 7, 11])
def test_stream_that_is_only_4_byte_aligned(ct):
    """dcb_decompress_f64 asks for a 4-byte aligned stream; the ring decoder needs 16 (cp.async), so a stream at 16k + 4 takes the
    register-window kernel (csrc/dcb_f64.cuh dcb_f64_decode_kernel).  Same doubles bit for bit."""
    import ctypes as C
    import torch
    from data_compression_b200 import api
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(60 + ct)
    n = 9 * 4096 + 321
    x = np.concatenate([np.cumsum(rng.standard_normal(n // 2)) * 0.02 + 20.0, rng.random(n - n // 2) * 10.0])
    comp = api.compress_f64(torch.from_numpy(x).to(dev), ct=ct, abs_error_bound=1e-3, chunk_elems=4096)
    want = api.decompress_f64(comp)
    total = comp.total_bytes()
    buf = torch.zeros(total + 64, dtype=torch.uint8, device=dev)
    buf[4:4 + total] = comp.stream[:total]
    assert (buf.data_ptr() + 4) % 16 == 4
    out = torch.empty(n, dtype=torch.float64, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    per_chunk = ct == 7
    rc = api.lib().dcb_decompress_f64(ct, C.c_void_p(buf.data_ptr() + 4), C.c_void_p(comp.chunk_offsets.data_ptr()), n, 4096, 1e-3,
                                      api.F_SUBTRACT_MIN, 0, 0, C.c_void_p(comp.chunk_min.data_ptr()),
                                      C.c_void_p(comp.chunk_mask20.data_ptr()) if per_chunk else None,
                                      C.c_void_p(comp.chunk_type.data_ptr()) if per_chunk else None,
                                      C.c_void_p(out.data_ptr()), C.c_void_p(status.data_ptr()),
                                      C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    assert status.item() == 0
    assert torch.equal(out.view(torch.int64), want.view(torch.int64))
