"""CT=1 (second half): the byte-wise codec myCompress / myDecompress on the device against the CPU oracle
(oracle/dc_oracle_bytewise.c, pinned to the compiled reference by tests/test_oracle_bytewise.py): raw floats, hit types and
displacements bit for bit, per chunk and for the legacy single call, through the device API and the reference's own symbols."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
from data_compression_b200 import api
from data_compression_b200.fields import make_field

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.fail("no CUDA device: the product has no CPU fallback")
    return torch.device("cuda:0")


def inputs():
    rng = np.random.default_rng(9)
    yield "smooth", make_field("smooth", 4096 * 5 + 300)
    yield "noisy", make_field("noisy", 4096 * 3)
    yield "mixed", np.concatenate([make_field("smooth", 6000), make_field("noisy", 3000), make_field("constant", 4000)])
    x = make_field("smooth", 4096 * 4).copy()
    x[[0, 3, 4100, 4101, 9000, 16383]] = -1.0                   # the empty-slot marker: those chunks take the state machine
    yield "sentinel", x
    x = rng.normal(0, 1, 4096 * 2 + 17).astype(np.float32)
    x[[5, 4096 + 9]] = np.nan
    x[7000] = np.inf
    yield "nonfinite", x
    yield "tiny", np.array([3.0, 1.0, 4.0], np.float32)


def check(x, eb, chunk, dev):
    n = x.size
    enc = api.bytewise_compress(torch.from_numpy(x).to(dev), abs_error_bound=eb, chunk_elems=chunk)
    out = api.bytewise_decompress(enc).cpu().numpy()
    torch.cuda.synchronize()
    c_el = chunk if (chunk and chunk < n) else n
    af, ac, ad, nf = (enc[k].cpu().numpy() for k in ("array_float", "array_char", "displacement", "nfloat"))
    for c in range((n + c_el - 1) // c_el):
        seg = x[c * c_el:(c + 1) * c_el]
        of, oc, od = ol.o_bytewise_compress(seg, eb)
        b = c * c_el
        assert int(nf[c]) == of.size, (c, int(nf[c]), of.size)
        assert np.array_equal(af[b:b + of.size].view(np.uint32), of.view(np.uint32)), ("floats", c)
        assert np.array_equal(ac[b:b + oc.size], oc) and np.array_equal(ad[b:b + od.size], od), ("hits", c)
        dec = ol.o_bytewise_decompress(of, oc, od, seg.size)
        assert np.array_equal(out[b:b + seg.size].view(np.uint32), dec.view(np.uint32)), ("decoded", c)
    return enc


@pytest.mark.parametrize("eb", (0.01, 1e-4))
def test_chunked_and_single_call(eb, dev):
    for name, x in inputs():
        check(x, eb, 4096, dev)
        check(x, eb, 0, dev)                                     # one chunk = the whole array: the legacy call


def test_error_bound_on_smooth_data(dev):
    x = make_field("smooth", 1 << 20)
    enc = check(x[:4096 * 8], 0.01, 4096, dev)
    assert int(enc["nfloat"].sum()) < 0.2 * 4096 * 8             # mostly hits: ratio near 4


def test_reference_symbols(dev):
    """myCompress / myDecompress as impl/pingpong.c:148-161 and :520-530 call them."""
    L = api.lib()
    f32p = C.POINTER(C.c_float)
    L.myCompress.argtypes = [C.c_void_p, C.POINTER(f32p), C.POINTER(C.c_char_p), C.POINTER(C.POINTER(C.c_int)), C.c_int]
    L.myCompress.restype = C.c_int
    L.myDecompress.argtypes = [f32p, C.c_char_p, C.POINTER(C.c_int), C.c_int]
    L.myDecompress.restype = f32p
    L.dc_set_abs_error_bound.argtypes = [C.c_double]
    L.dc_set_abs_error_bound(0.01)
    for name, x in (("smooth", make_field("smooth", 10000)), ("noisy", make_field("noisy", 5000))):
        af, ac, ad = f32p(), C.c_char_p(), C.POINTER(C.c_int)()
        nf = L.myCompress(x.ctypes.data, C.byref(af), C.byref(ac), C.byref(ad), x.size)
        of, oc, od = ol.o_bytewise_compress(x, 0.01)
        nc = x.size - nf
        assert nf == of.size
        assert np.array_equal(np.ctypeslib.as_array(af, (nf,)).view(np.uint32), of.view(np.uint32))
        if nc:
            assert C.string_at(ac, nc) == oc.tobytes()
            assert np.array_equal(np.ctypeslib.as_array(ad, (nc,)), od)
        p = L.myDecompress(af, ac, ad, x.size)
        dec = np.ctypeslib.as_array(p, (x.size,))
        assert np.array_equal(dec.view(np.uint32), ol.o_bytewise_decompress(of, oc, od, x.size).view(np.uint32))
        assert np.max(np.abs(dec - x)) < 1.0                     # open-loop encoder: the bound drifts, not unbounded here
    L.dc_set_abs_error_bound(0.000001)


def test_float64_device_and_symbols(dev):
    """myCompress_double / myDecompress_double (impl/k-means.c:898-971, impl/mm.c:821, impl/lu.c:431) on the device."""
    for name, x in inputs():
        xd = x.astype(np.float64) * 1.000001
        if name == "sentinel":
            xd[[0, 3, 4100, 4101, 9000, 16383]] = -1.0
        for chunk in (4096, 0):
            n = xd.size
            enc = api.bytewise_compress(torch.from_numpy(xd).to(dev), abs_error_bound=1e-4, chunk_elems=chunk)
            out = api.bytewise_decompress(enc).cpu().numpy()
            c_el = chunk if (chunk and chunk < n) else n
            af, ac, ad, nf = (enc[k].cpu().numpy() for k in ("array_float", "array_char", "displacement", "nfloat"))
            for c in range((n + c_el - 1) // c_el):
                seg = xd[c * c_el:(c + 1) * c_el]
                of, oc, od = ol.o_bytewise_compress_f64(seg, 1e-4)
                b = c * c_el
                assert int(nf[c]) == of.size
                assert np.array_equal(af[b:b + of.size].view(np.uint64), of.view(np.uint64)), (name, chunk, c)
                assert np.array_equal(ac[b:b + oc.size], oc) and np.array_equal(ad[b:b + od.size], od), (name, chunk, c)
                dec = ol.o_bytewise_decompress_f64(of, oc, od, seg.size)
                assert np.array_equal(out[b:b + seg.size].view(np.uint64), dec.view(np.uint64)), (name, chunk, c)
    L = api.lib()
    f64p = C.POINTER(C.c_double)
    L.myCompress_double.argtypes = [C.c_void_p, C.POINTER(f64p), C.POINTER(C.c_char_p), C.POINTER(C.POINTER(C.c_int)), C.c_int]
    L.myCompress_double.restype = C.c_int
    L.myDecompress_double.argtypes = [f64p, C.c_char_p, C.POINTER(C.c_int), C.c_int]
    L.myDecompress_double.restype = f64p
    L.dc_set_abs_error_bound.argtypes = [C.c_double]
    L.dc_set_abs_error_bound(0.001)
    x = np.cumsum(np.random.default_rng(2).normal(0, 1e-3, 6000)) + 2.0
    af, ac, ad = f64p(), C.c_char_p(), C.POINTER(C.c_int)()
    nf = L.myCompress_double(x.ctypes.data, C.byref(af), C.byref(ac), C.byref(ad), x.size)
    of, oc, od = ol.o_bytewise_compress_f64(x, 0.001)
    assert nf == of.size and np.array_equal(np.ctypeslib.as_array(af, (nf,)).view(np.uint64), of.view(np.uint64))
    nc = x.size - nf
    assert C.string_at(ac, nc) == oc.tobytes() and np.array_equal(np.ctypeslib.as_array(ad, (nc,)), od)
    p = L.myDecompress_double(af, ac, ad, x.size)
    dec = np.ctypeslib.as_array(p, (x.size,))
    assert np.array_equal(dec.view(np.uint64), ol.o_bytewise_decompress_f64(of, oc, od, x.size).view(np.uint64))
    L.dc_set_abs_error_bound(0.000001)
