"""The reference's own symbols served by libdcb200.so (include/dataCompression_b200.h), called exactly as
impl/pingpong.c:148-210 and tools/mycompress_bitwise_float.c:38-60 call them, against the CPU oracle."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
from data_compression_b200 import api
from data_compression_b200.fields import make_field

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

_u8p = C.POINTER(C.c_uint8)
_f32p = C.POINTER(C.c_float)
_libc = C.CDLL(None)
_libc.free.argtypes = [C.c_void_p]


@pytest.fixture(scope="module")
def L():
    if not torch.cuda.is_available():
        pytest.fail("no CUDA device: the product has no CPU fallback")
    lib = api.lib()
    enc = [_f32p, C.c_int, C.POINTER(_u8p), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    dec = [_u8p, C.c_int, C.c_int]
    for name in ("myCompress_bitwise", "myCompress_bitwise_np", "myCompress_bitwise_op"):
        getattr(lib, name).argtypes, getattr(lib, name).restype = enc, None
    lib.myCompress_bitwise_mask.argtypes, lib.myCompress_bitwise_mask.restype = enc + [C.c_int, C.c_char_p], None
    for name in ("myDecompress_bitwise", "myDecompress_bitwise_np", "myDecompress_bitwise_op"):
        getattr(lib, name).argtypes, getattr(lib, name).restype = dec, _f32p
    lib.myDecompress_bitwise_mask.argtypes, lib.myDecompress_bitwise_mask.restype = dec + [C.c_int, C.c_char_p], _f32p
    lib.toSmallDataset_float.argtypes, lib.toSmallDataset_float.restype = [_f32p, C.POINTER(_f32p), C.c_int], C.c_float
    lib.med_dataset_float.argtypes, lib.med_dataset_float.restype = [_f32p, C.c_int, C.POINTER(C.c_int)], C.c_float
    lib.do_crc32.argtypes, lib.do_crc32.restype = [_u8p, C.c_int], C.c_uint32
    lib.to_absErrorBound_binary.argtypes, lib.to_absErrorBound_binary.restype = [C.c_double], C.c_int
    lib.floattostr.argtypes = [_f32p, C.c_char_p]
    lib.strtofloat.argtypes, lib.strtofloat.restype = [C.c_char_p], C.c_float
    lib.dc_set_abs_error_bound.argtypes = [C.c_double]
    lib.dc_get_abs_error_bound.restype = C.c_double
    return lib


ENC = {5: "myCompress_bitwise", 6: "myCompress_bitwise_np", 11: "myCompress_bitwise_op", 7: "myCompress_bitwise_mask"}
DEC = {5: "myDecompress_bitwise", 6: "myDecompress_bitwise_np", 11: "myDecompress_bitwise_op", 7: "myDecompress_bitwise_mask"}


def legacy_compress(L, ct, y, type_=0, mask=b"", state=None):
    bits, nbytes, pos = state or (_u8p(), C.c_int(0), C.c_int(8))          # (NULL, 0, 8): impl/pingpong.c:164-166
    args = [y.ctypes.data_as(_f32p), y.size, C.byref(bits), C.byref(nbytes), C.byref(pos)]
    if ct == 7:
        args += [type_, mask]
    getattr(L, ENC[ct])(*args)
    return bits, nbytes, pos


def test_pingpong_call_sequence(L):
    """min-subtract -> [medium/type/mask] -> compress -> crc -> decompress, all through the legacy symbols."""
    rng = np.random.default_rng(21)
    datasets = [make_field("smooth", 16641), (rng.random(8192) * 10).astype(np.float32), make_field("constant", 5000),
                make_field("noisy", 131072)]
    for eb in (0.01, 1e-6):
        L.dc_set_abs_error_bound(eb)
        assert L.dc_get_abs_error_bound() == eb
        assert L.to_absErrorBound_binary(eb) == ol.o_n_eb(eb)
        for x in datasets:
            p = _f32p()
            mn = L.toSmallDataset_float(x.ctypes.data_as(_f32p), C.byref(p), x.size)
            y = np.ctypeslib.as_array(p, shape=(x.size,)).copy()
            _libc.free(p)
            omn, oy = ol.o_min_subtract(x)
            assert np.float32(mn) == omn and np.array_equal(y.view(np.uint32), oy.view(np.uint32))
            t = C.c_int(0)
            med = np.float32(L.med_dataset_float(y.ctypes.data_as(_f32p), y.size, C.byref(t)))
            omed, ot = ol.o_med(y, 0)
            assert med.view(np.uint32) == omed.view(np.uint32) and t.value == ot      # exact sequential float32 mean
            s = C.create_string_buffer(33)
            L.floattostr(C.byref(C.c_float(float(med))), s)
            mask = s.value[:17]                                                        # impl/pingpong.c:202-206
            assert int(mask, 2) == ol.mask17_of(med)
            assert np.float32(L.strtofloat(s.value)).view(np.uint32) == med.view(np.uint32)
            for ct in (5, 6, 11, 7):
                bits, nbytes, pos = legacy_compress(L, ct, y, t.value, mask)
                got = np.ctypeslib.as_array(bits, shape=(nbytes.value,)).copy()
                want, wpos = ol.o_compress(ct, y, eb, t.value, ol.mask17_of(med))
                assert nbytes.value == want.size and pos.value == wpos, (ct, eb, x.size)
                assert np.array_equal(got, want), (ct, eb, x.size)
                assert L.do_crc32(bits, nbytes.value) == ol.o_crc32(want)
                dargs = [bits, nbytes.value, y.size] + ([t.value, mask] if ct == 7 else [])
                dp = getattr(L, DEC[ct])(*dargs)
                dec = np.ctypeslib.as_array(dp, shape=(y.size,)).copy()
                _libc.free(dp)
                _libc.free(bits)
                odec = ol.o_decompress(ct, want, y.size, eb, t.value, ol.mask17_of(med))
                assert np.array_equal(dec.view(np.uint32), odec.view(np.uint32)), (ct, eb, x.size)
    L.dc_set_abs_error_bound(1e-6)


def test_append_to_existing_stream(L):
    """The encoders APPEND to (*data_bits, *bytes, *pos) (dataCompression.c:5456-5489): two calls == one stream of
    both code sequences, bit-contiguous (the second call starts mid-byte)."""
    L.dc_set_abs_error_bound(0.01)
    rng = np.random.default_rng(22)
    a = (rng.random(1001) * 10).astype(np.float32)
    b = (rng.random(777) * 10).astype(np.float32)
    for ct in (5, 6, 11):
        st = legacy_compress(L, ct, a)
        pos_a = st[2].value
        st = legacy_compress(L, ct, b, state=st)
        got = np.ctypeslib.as_array(st[0], shape=(st[1].value,)).copy()
        _libc.free(st[0])
        sa, pa = ol.o_compress(ct, a, 0.01)
        sb, pb = ol.o_compress(ct, b, 0.01)
        assert pos_a == pa
        bits_a = np.unpackbits(sa)[: sa.size * 8 - (0 if pa == 8 else pa)]
        bits_b = np.unpackbits(sb)[: sb.size * 8 - (0 if pb == 8 else pb)]
        want = np.packbits(np.concatenate([bits_a, bits_b]))
        assert np.array_equal(got, want), ct
        tot = bits_a.size + bits_b.size
        assert st[2].value == (8 if tot % 8 == 0 else 8 - tot % 8)
    L.dc_set_abs_error_bound(1e-6)


# ---- float64 symbols: the call sequence of impl/k-means.c:520-590 -------------------------------------------------
_f64p = C.POINTER(C.c_double)
ENC64 = {5: "myCompress_bitwise_double", 6: "myCompress_bitwise_double_np", 11: "myCompress_bitwise_double_op",
         7: "myCompress_bitwise_double_mask"}
DEC64 = {5: "myDecompress_bitwise_double", 6: "myDecompress_bitwise_double_np", 11: "myDecompress_bitwise_double_op",
         7: "myDecompress_bitwise_double_mask"}


@pytest.fixture(scope="module")
def L64(L):
    enc = [_f64p, C.c_int, C.POINTER(_u8p), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    dec = [_u8p, C.c_int, C.c_int]
    for ct in (5, 6, 11):
        getattr(L, ENC64[ct]).argtypes, getattr(L, ENC64[ct]).restype = enc, None
        getattr(L, DEC64[ct]).argtypes, getattr(L, DEC64[ct]).restype = dec, _f64p
    L.myCompress_bitwise_double_mask.argtypes, L.myCompress_bitwise_double_mask.restype = enc + [C.c_int, C.c_char_p], None
    L.myDecompress_bitwise_double_mask.argtypes, L.myDecompress_bitwise_double_mask.restype = dec + [C.c_int, C.c_char_p], _f64p
    L.toSmallDataset_double.argtypes, L.toSmallDataset_double.restype = [_f64p, C.POINTER(_f64p), C.c_int], C.c_double
    L.med_dataset_double.argtypes, L.med_dataset_double.restype = [_f64p, C.c_int, C.POINTER(C.c_int)], C.c_double
    L.doubletostr.argtypes = [_f64p, C.c_char_p]
    L.strtodbl.argtypes, L.strtodbl.restype = [C.c_char_p], C.c_double
    return L


def test_kmeans_call_sequence_f64(L64):
    L = L64
    rng = np.random.default_rng(64)
    datasets = [np.sort(rng.random(100) * 1000.0), rng.random(8192) * 10, np.full(5000, 2.5),
                np.cumsum(rng.standard_normal(20001)) * 0.01 + 7.0, np.array([3.25])]
    for eb in (0.01, 1e-6):
        L.dc_set_abs_error_bound(eb)
        for x in datasets:
            p = _f64p()
            mn = L.toSmallDataset_double(x.ctypes.data_as(_f64p), C.byref(p), x.size)
            y = np.ctypeslib.as_array(p, shape=(x.size,)).copy()
            _libc.free(p)
            omn, oy = ol.o_min_subtract_f64(x)
            assert np.float64(mn) == omn and np.array_equal(y.view(np.uint64), oy.view(np.uint64))
            t = C.c_int(0)
            med = np.float64(L.med_dataset_double(y.ctypes.data_as(_f64p), y.size, C.byref(t)))
            omed, ot = ol.o_med_f64(y, 0)
            assert med.view(np.uint64) == omed.view(np.uint64) and t.value == ot       # exact sequential mean
            s = C.create_string_buffer(65)
            L.doubletostr(C.byref(C.c_double(float(med))), s)
            mask = s.value[:20]                                                         # impl/k-means.c:527-530
            assert int(mask, 2) == ol.mask20_of(med)
            assert np.float64(L.strtodbl(s.value)).view(np.uint64) == med.view(np.uint64)
            for ct in (5, 6, 11, 7):
                bits, nbytes, pos = _u8p(), C.c_int(0), C.c_int(8)
                args = [y.ctypes.data_as(_f64p), y.size, C.byref(bits), C.byref(nbytes), C.byref(pos)]
                if ct == 7:
                    args += [t.value, mask]
                getattr(L, ENC64[ct])(*args)
                got = np.ctypeslib.as_array(bits, shape=(nbytes.value,)).copy()
                want, wpos = ol.o_compress_f64(ct, y, eb, t.value, ol.mask20_of(med))
                assert nbytes.value == want.size and pos.value == wpos, (ct, eb, x.size)
                assert np.array_equal(got, want), (ct, eb, x.size)
                dargs = [bits, nbytes.value, y.size] + ([t.value, mask] if ct == 7 else [])
                dp = getattr(L, DEC64[ct])(*dargs)
                dec = np.ctypeslib.as_array(dp, shape=(y.size,)).copy()
                _libc.free(dp)
                _libc.free(bits)
                odec = ol.o_decompress_f64(ct, want, y.size, eb, t.value, ol.mask20_of(med))
                assert np.array_equal(dec.view(np.uint64), odec.view(np.uint64)), (ct, eb, x.size)
    L.dc_set_abs_error_bound(1e-6)
