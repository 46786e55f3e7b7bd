"""ctypes bindings for the CPU oracle (oracle/dc_oracle.c) and, when it was built, the unmodified
reference hot path (oracle/_ref/libdcref_<eb>.so, compiled from /root/reference by oracle/build_ref.sh).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "_build", "libdc_oracle.so")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")

CT_BITWISE, CT_NP, CT_MASK, CT_OP = 5, 6, 7, 11
ALL_CTS = (CT_BITWISE, CT_NP, CT_MASK, CT_OP)

_u8p = C.POINTER(C.c_uint8)
_f32p = C.POINTER(C.c_float)


def eb_str(eb):
    """Decimal spelling of an error bound as the reference's sed recipe writes it (no exponent)."""
    s = ("%.12f" % eb).rstrip("0")
    return s + "0" if s.endswith(".") else s


def build_oracle():
    srcs = [os.path.join(ORACLE_DIR, f) for f in ("dc_oracle.c", "dc_oracle_f64.c", "dc_oracle_hamming.c", "dc_oracle_bytewise.c", "dc_oracle.h")]
    if (not os.path.exists(ORACLE_SO)) or os.path.getmtime(ORACLE_SO) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "_build/libdc_oracle.so"], stdout=subprocess.DEVNULL)
    return ORACLE_SO


_oracle = None


def oracle():
    global _oracle
    if _oracle is None:
        lib = C.CDLL(build_oracle())
        lib.dco_n_eb.restype = C.c_int
        lib.dco_n_eb.argtypes = [C.c_double]
        lib.dco_min_subtract.restype = C.c_float
        lib.dco_min_subtract.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        lib.dco_med.restype = C.c_float
        lib.dco_med.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_int)]
        lib.dco_mask17.restype = C.c_uint32
        lib.dco_mask17.argtypes = [C.c_float]
        lib.dco_crc32.restype = C.c_uint32
        lib.dco_crc32.argtypes = [C.c_void_p, C.c_size_t]
        lib.dco_compress.restype = C.c_size_t
        lib.dco_compress.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_double, C.c_int, C.c_uint32,
                                     C.c_void_p, C.POINTER(C.c_int)]
        lib.dco_decompress.restype = C.c_int
        lib.dco_decompress.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_size_t, C.c_double, C.c_int,
                                       C.c_uint32, C.c_void_p]
        lib.dco_compress_chunked.restype = C.c_size_t
        lib.dco_compress_chunked.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_size_t, C.c_double, C.c_int,
                                             C.c_int, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_void_p, C.c_void_p]
        lib.dco_decompress_chunked.restype = C.c_int
        lib.dco_decompress_chunked.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t,
                                               C.c_double, C.c_int, C.c_void_p, C.c_int, C.c_uint32,
                                               C.c_void_p, C.c_void_p, C.c_void_p]
        lib.dco_min_subtract_f64.restype = C.c_double
        lib.dco_min_subtract_f64.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        lib.dco_med_f64.restype = C.c_double
        lib.dco_med_f64.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_int)]
        lib.dco_mask20.restype = C.c_uint32
        lib.dco_mask20.argtypes = [C.c_double]
        lib.dco_compress_f64.restype = C.c_size_t
        lib.dco_compress_f64.argtypes = lib.dco_compress.argtypes
        lib.dco_decompress_f64.restype = C.c_int
        lib.dco_decompress_f64.argtypes = lib.dco_decompress.argtypes
        lib.dco_compress_chunked_f64.restype = C.c_size_t
        lib.dco_compress_chunked_f64.argtypes = lib.dco_compress_chunked.argtypes
        lib.dco_decompress_chunked_f64.restype = C.c_int
        lib.dco_decompress_chunked_f64.argtypes = lib.dco_decompress_chunked.argtypes
        _oracle = lib
    return _oracle


def _f32(x):
    return np.ascontiguousarray(x, dtype=np.float32)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


# ------------------------------------------------------------------ oracle (restatement) wrappers
def o_n_eb(eb):
    return oracle().dco_n_eb(eb)


def o_min_subtract(x):
    x = _f32(x)
    y = np.empty_like(x)
    mn = oracle().dco_min_subtract(_ptr(x), x.size, _ptr(y))
    return np.float32(mn), y


def o_med(x, type0=0):
    x = _f32(x)
    t = C.c_int(type0)
    med = oracle().dco_med(_ptr(x), x.size, C.byref(t))
    return np.float32(med), t.value


def mask17_of(medium):
    return int(np.array([medium], dtype=np.float32).view(np.uint32)[0] >> 15)


def o_crc32(buf):
    b = np.ascontiguousarray(buf, dtype=np.uint8)
    return oracle().dco_crc32(_ptr(b), b.size)


def o_compress(ct, x, eb, type_=0, mask17=0):
    x = _f32(x)
    out = np.zeros(4 * x.size + 8, dtype=np.uint8)
    pos = C.c_int(0)
    nb = oracle().dco_compress(ct, _ptr(x), x.size, eb, type_, mask17, _ptr(out), C.byref(pos))
    return out[:nb].copy(), pos.value


def o_decompress(ct, stream, n, eb, type_=0, mask17=0):
    s = np.ascontiguousarray(stream, dtype=np.uint8)
    out = np.empty(n, dtype=np.float32)
    rc = oracle().dco_decompress(ct, _ptr(s), s.size, n, eb, type_, mask17, _ptr(out))
    if rc:
        raise ValueError("oracle: malformed stream (rc=%d)" % rc)
    return out


def o_compress_chunked(ct, x, chunk, eb, subtract_min=True, fixed_type=-1, fixed_mask17=0, want_crc=False):
    """-> dict(stream, offsets, chunk_min, chunk_medium, chunk_type, chunk_crc)"""
    x = _f32(x)
    n = x.size
    K = (n + chunk - 1) // chunk if n else 0
    out = np.zeros(4 * n + 8 * max(K, 1), dtype=np.uint8)
    offs = np.zeros(K + 1, dtype=np.uint64)
    cmin = np.zeros(K, dtype=np.float32)
    cmed = np.zeros(K, dtype=np.float32)
    ctype = np.zeros(K, dtype=np.int32)
    ccrc = np.zeros(K, dtype=np.uint32) if want_crc else None
    tot = oracle().dco_compress_chunked(ct, _ptr(x), n, chunk, eb, int(subtract_min), fixed_type, fixed_mask17,
                                        _ptr(out), _ptr(offs), _ptr(cmin), _ptr(cmed), _ptr(ctype), _ptr(ccrc))
    return dict(stream=out[:tot].copy(), offsets=offs, chunk_min=cmin, chunk_medium=cmed, chunk_type=ctype,
                chunk_crc=ccrc)


def o_decompress_chunked(ct, enc, n, chunk, eb, add_min=True, fixed_type=-1, fixed_mask17=0):
    out = np.empty(n, dtype=np.float32)
    s = np.ascontiguousarray(enc["stream"], dtype=np.uint8)
    if s.size == 0:
        s = np.zeros(1, dtype=np.uint8)
    rc = oracle().dco_decompress_chunked(ct, _ptr(s), _ptr(enc["offsets"]), n, chunk, eb, int(add_min),
                                         _ptr(enc["chunk_min"]), fixed_type, fixed_mask17,
                                         _ptr(enc["chunk_medium"]), _ptr(enc["chunk_type"]), _ptr(out))
    if rc:
        raise ValueError("oracle: malformed stream (rc=%d)" % rc)
    return out


# ------------------------------------------------------------------ unmodified reference (oracle/_ref)
_libc = C.CDLL(None)
_libc.free.argtypes = [C.c_void_p]
_libc.free.restype = None


def ref_available(eb):
    return os.path.exists(os.path.join(REF_DIR, "libdcref_%s.so" % eb_str(eb)))


def build_ref(eb):
    """Build oracle/_ref/libdcref_<eb>.so if /root/reference is here (no-op on the GPU box)."""
    if not ref_available(eb):
        subprocess.call([os.path.join(ORACLE_DIR, "build_ref.sh"), eb_str(eb)], stdout=subprocess.DEVNULL)
    return ref_available(eb)


class RefLib:
    """The reference's own entry points (impl/dataCompression.h:67-99,143), called exactly as
    impl/pingpong.c:148-210 and tools/mycompress_bitwise_float.c:38-56 call them."""
    _cache = {}

    def __new__(cls, eb):
        key = eb_str(eb)
        if key in cls._cache:
            return cls._cache[key]
        if not build_ref(eb):
            raise FileNotFoundError("oracle/_ref/libdcref_%s.so not built" % key)
        self = super().__new__(cls)
        lib = C.CDLL(os.path.join(REF_DIR, "libdcref_%s.so" % key))
        self.lib, self.eb = lib, eb
        enc = [_f32p, C.c_int, C.POINTER(_u8p), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        dec = [_u8p, C.c_int, C.c_int]
        for name in ("myCompress_bitwise", "myCompress_bitwise_np", "myCompress_bitwise_op"):
            getattr(lib, name).argtypes = enc
            getattr(lib, name).restype = None
        lib.myCompress_bitwise_mask.argtypes = enc + [C.c_int, C.c_char_p]
        lib.myCompress_bitwise_mask.restype = None
        for name in ("myDecompress_bitwise", "myDecompress_bitwise_np", "myDecompress_bitwise_op"):
            getattr(lib, name).argtypes = dec
            getattr(lib, name).restype = _f32p
        lib.myDecompress_bitwise_mask.argtypes = dec + [C.c_int, C.c_char_p]
        lib.myDecompress_bitwise_mask.restype = _f32p
        lib.toSmallDataset_float.argtypes = [_f32p, C.POINTER(_f32p), C.c_int]
        lib.toSmallDataset_float.restype = C.c_float
        lib.med_dataset_float.argtypes = [_f32p, C.c_int, C.POINTER(C.c_int)]
        lib.med_dataset_float.restype = C.c_float
        lib.do_crc32.argtypes = [_u8p, C.c_int]
        lib.do_crc32.restype = C.c_uint32
        lib.to_absErrorBound_binary.argtypes = [C.c_double]
        lib.to_absErrorBound_binary.restype = C.c_int
        lib.floattostr.argtypes = [_f32p, C.c_char_p]
        cls._cache[key] = self
        return self

    _ENC = {CT_BITWISE: "myCompress_bitwise", CT_NP: "myCompress_bitwise_np", CT_OP: "myCompress_bitwise_op",
            CT_MASK: "myCompress_bitwise_mask"}
    _DEC = {CT_BITWISE: "myDecompress_bitwise", CT_NP: "myDecompress_bitwise_np",
            CT_OP: "myDecompress_bitwise_op", CT_MASK: "myDecompress_bitwise_mask"}

    @staticmethod
    def mask_chars(mask17):
        return ("{:017b}".format(mask17)).encode()

    def n_eb(self):
        return self.lib.to_absErrorBound_binary(self.eb)

    def min_subtract(self, x):
        x = _f32(x)
        p = _f32p()
        mn = self.lib.toSmallDataset_float(x.ctypes.data_as(_f32p), C.byref(p), x.size)
        y = np.ctypeslib.as_array(p, shape=(x.size,)).copy()
        _libc.free(p)
        return np.float32(mn), y

    def med(self, x, type0=0):
        x = _f32(x)
        t = C.c_int(type0)
        med = self.lib.med_dataset_float(x.ctypes.data_as(_f32p), x.size, C.byref(t))
        return np.float32(med), t.value

    def crc32(self, buf):
        b = np.ascontiguousarray(buf, dtype=np.uint8)
        return self.lib.do_crc32(b.ctypes.data_as(_u8p), b.size)

    def compress(self, ct, x, type_=0, mask17=0):
        x = _f32(x)
        bits, nbytes, pos = _u8p(), C.c_int(0), C.c_int(8)     # (NULL, 0, 8): impl/pingpong.c:164-166
        args = [x.ctypes.data_as(_f32p), x.size, C.byref(bits), C.byref(nbytes), C.byref(pos)]
        if ct == CT_MASK:
            args += [type_, self.mask_chars(mask17)]
        getattr(self.lib, self._ENC[ct])(*args)
        out = np.ctypeslib.as_array(bits, shape=(nbytes.value,)).copy() if nbytes.value else np.zeros(0, np.uint8)
        if nbytes.value:
            _libc.free(bits)
        return out, pos.value

    def decompress(self, ct, stream, n, type_=0, mask17=0, pad=True):
        """pad=True appends one zero byte: the reference only emits a 9-bit / zero-payload code when the
        NEXT bit arrives (dataCompression.c:3022-3025), so when such a code is the last one of a byte-aligned
        stream it leaves out[n-1] uninitialised.  Eight zero bits can never complete another code, so the
        padded call equals the plain call everywhere else."""
        s = np.ascontiguousarray(stream, dtype=np.uint8)
        if pad:
            s = np.concatenate([s, np.zeros(1, np.uint8)])
        args = [s.ctypes.data_as(_u8p), s.size, n]
        if ct == CT_MASK:
            args += [type_, self.mask_chars(mask17)]
        p = getattr(self.lib, self._DEC[ct])(*args)
        out = np.ctypeslib.as_array(p, shape=(n,)).copy()
        _libc.free(p)
        return out

    def compress_chunked(self, ct, x, chunk, subtract_min=True, want_crc=False):
        """The reference run chunk-by-chunk. Same dict as o_compress_chunked."""
        x = _f32(x)
        n = x.size
        K = (n + chunk - 1) // chunk if n else 0
        parts, offs = [], np.zeros(K + 1, dtype=np.uint64)
        cmin, cmed = np.zeros(K, np.float32), np.zeros(K, np.float32)
        ctype, ccrc = np.zeros(K, np.int32), np.zeros(K, np.uint32)
        for c in range(K):
            seg = x[c * chunk:(c + 1) * chunk]
            if subtract_min:
                cmin[c], seg = self.min_subtract(seg)
            t, m17 = 0, 0
            if ct == CT_MASK:
                cmed[c], t = self.med(seg, 0)
                ctype[c], m17 = t, mask17_of(cmed[c])
            s, _ = self.compress(ct, seg, t, m17)
            if want_crc:
                ccrc[c] = self.crc32(s)
            parts.append(s)
            offs[c + 1] = offs[c] + np.uint64(s.size)
        stream = np.concatenate(parts) if parts else np.zeros(0, np.uint8)
        return dict(stream=stream, offsets=offs, chunk_min=cmin, chunk_medium=cmed, chunk_type=ctype,
                    chunk_crc=ccrc if want_crc else None)


def ref_mask_decode_defined(type_, mask17, eb):
    """True where the reference CT=7 DECODER is well defined (SURVEY.md F5 + the 9-bit escape corner,
    dataCompression.c:1866): 2<=type<=6; escape payload rule m==0 or m>8; no escape of total length 9."""
    n_eb = o_n_eb(eb)
    m = min(max(n_eb + ((mask17 >> 8) & 0xFF) - 127, 0), 23)
    if not (2 <= type_ <= 6):
        return False
    if 0 < m <= 8:
        return False
    k0 = max(m - 8, 0)
    return (type_ + 2 + k0 != 9) and (type_ + 2 + m != 9)


# ------------------------------------------------------------------ float64 variants
_f64p = C.POINTER(C.c_double)


def _f64(x):
    return np.ascontiguousarray(x, dtype=np.float64)


def mask20_of(medium):
    return int(np.array([medium], dtype=np.float64).view(np.uint64)[0] >> np.uint64(44))


def o_min_subtract_f64(x):
    x = _f64(x)
    y = np.empty_like(x)
    mn = oracle().dco_min_subtract_f64(_ptr(x), x.size, _ptr(y))
    return np.float64(mn), y


def o_med_f64(x, type0=0):
    x = _f64(x)
    t = C.c_int(type0)
    med = oracle().dco_med_f64(_ptr(x), x.size, C.byref(t))
    return np.float64(med), t.value


def o_compress_f64(ct, x, eb, type_=0, mask20=0):
    x = _f64(x)
    out = np.zeros(8 * x.size + 16, dtype=np.uint8)
    pos = C.c_int(0)
    nb = oracle().dco_compress_f64(ct, _ptr(x), x.size, eb, type_, mask20, _ptr(out), C.byref(pos))
    return out[:nb].copy(), pos.value


def o_decompress_f64(ct, stream, n, eb, type_=0, mask20=0):
    s = np.ascontiguousarray(stream, dtype=np.uint8)
    out = np.empty(n, dtype=np.float64)
    rc = oracle().dco_decompress_f64(ct, _ptr(s), s.size, n, eb, type_, mask20, _ptr(out))
    if rc:
        raise ValueError("oracle: malformed stream (rc=%d)" % rc)
    return out


def o_compress_chunked_f64(ct, x, chunk, eb, subtract_min=True, fixed_type=-1, fixed_mask20=0, want_crc=False):
    """-> dict(stream, offsets, chunk_min, chunk_medium, chunk_mask20, chunk_type, chunk_crc)"""
    x = _f64(x)
    n = x.size
    K = (n + chunk - 1) // chunk if n else 0
    out = np.zeros(8 * n + 16 * max(K, 1), dtype=np.uint8)
    offs = np.zeros(K + 1, dtype=np.uint64)
    cmin = np.zeros(K, dtype=np.float64)
    cmed = np.zeros(K, dtype=np.float64)
    ctype = np.zeros(K, dtype=np.int32)
    ccrc = np.zeros(K, dtype=np.uint32) if want_crc else None
    tot = oracle().dco_compress_chunked_f64(ct, _ptr(x), n, chunk, eb, int(subtract_min), fixed_type, fixed_mask20,
                                            _ptr(out), _ptr(offs), _ptr(cmin), _ptr(cmed), _ptr(ctype), _ptr(ccrc))
    return dict(stream=out[:tot].copy(), offsets=offs, chunk_min=cmin, chunk_medium=cmed,
                chunk_mask20=(cmed.view(np.uint64) >> np.uint64(44)).astype(np.uint32), chunk_type=ctype, chunk_crc=ccrc)


def o_decompress_chunked_f64(ct, enc, n, chunk, eb, add_min=True, fixed_type=-1, fixed_mask20=0):
    out = np.empty(n, dtype=np.float64)
    s = np.ascontiguousarray(enc["stream"], dtype=np.uint8)
    if s.size == 0:
        s = np.zeros(1, np.uint8)
    offs = np.ascontiguousarray(enc["offsets"], dtype=np.uint64)
    cmin = _f64(enc["chunk_min"])
    m20 = np.ascontiguousarray(enc["chunk_mask20"], dtype=np.uint32)
    ctype = np.ascontiguousarray(enc["chunk_type"], dtype=np.int32)
    rc = oracle().dco_decompress_chunked_f64(ct, _ptr(s), _ptr(offs), n, chunk, eb, int(add_min), _ptr(cmin), fixed_type,
                                             fixed_mask20, _ptr(m20), _ptr(ctype), _ptr(out))
    if rc:
        raise ValueError("oracle: malformed stream (rc=%d)" % rc)
    return out


class RefLibF64:
    """The reference's float64 entry points (impl/dataCompression.h), called as impl/k-means.c:520-590 calls them."""
    _ENC = {CT_BITWISE: "myCompress_bitwise_double", CT_NP: "myCompress_bitwise_double_np",
            CT_OP: "myCompress_bitwise_double_op", CT_MASK: "myCompress_bitwise_double_mask"}
    _DEC = {CT_BITWISE: "myDecompress_bitwise_double", CT_NP: "myDecompress_bitwise_double_np",
            CT_OP: "myDecompress_bitwise_double_op", CT_MASK: "myDecompress_bitwise_double_mask"}

    def __init__(self, eb):
        self.base = RefLib(eb)
        lib = self.lib = self.base.lib
        enc = [_f64p, C.c_int, C.POINTER(_u8p), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        dec = [_u8p, C.c_int, C.c_int]
        for name in (self._ENC[CT_BITWISE], self._ENC[CT_NP], self._ENC[CT_OP]):
            getattr(lib, name).argtypes = enc
            getattr(lib, name).restype = None
        lib.myCompress_bitwise_double_mask.argtypes = enc + [C.c_int, C.c_char_p]
        lib.myCompress_bitwise_double_mask.restype = None
        for name in (self._DEC[CT_BITWISE], self._DEC[CT_NP], self._DEC[CT_OP]):
            getattr(lib, name).argtypes = dec
            getattr(lib, name).restype = _f64p
        lib.myDecompress_bitwise_double_mask.argtypes = dec + [C.c_int, C.c_char_p]
        lib.myDecompress_bitwise_double_mask.restype = _f64p
        lib.toSmallDataset_double.argtypes = [_f64p, C.POINTER(_f64p), C.c_int]
        lib.toSmallDataset_double.restype = C.c_double
        lib.med_dataset_double.argtypes = [_f64p, C.c_int, C.POINTER(C.c_int)]
        lib.med_dataset_double.restype = C.c_double

    @staticmethod
    def mask_chars(mask20):
        return ("{:020b}".format(mask20)).encode()

    def min_subtract(self, x):
        x = _f64(x)
        p = _f64p()
        mn = self.lib.toSmallDataset_double(x.ctypes.data_as(_f64p), C.byref(p), x.size)
        y = np.ctypeslib.as_array(p, shape=(x.size,)).copy()
        _libc.free(p)
        return np.float64(mn), y

    def med(self, x, type0=0):
        x = _f64(x)
        t = C.c_int(type0)
        med = self.lib.med_dataset_double(x.ctypes.data_as(_f64p), x.size, C.byref(t))
        return np.float64(med), t.value

    def compress(self, ct, x, type_=0, mask20=0):
        x = _f64(x)
        bits, nbytes, pos = _u8p(), C.c_int(0), C.c_int(8)
        args = [x.ctypes.data_as(_f64p), x.size, C.byref(bits), C.byref(nbytes), C.byref(pos)]
        if ct == CT_MASK:
            args += [type_, self.mask_chars(mask20)]
        getattr(self.lib, self._ENC[ct])(*args)
        out = np.ctypeslib.as_array(bits, shape=(nbytes.value,)).copy() if nbytes.value else np.zeros(0, np.uint8)
        if nbytes.value:
            _libc.free(bits)
        return out, pos.value

    def decompress(self, ct, stream, n, type_=0, mask20=0, pad=True):
        """pad: one zero byte, for the same reason as RefLib.decompress (a 12-bit / zero-payload last code)."""
        s = np.ascontiguousarray(stream, dtype=np.uint8)
        if pad:
            s = np.concatenate([s, np.zeros(1, np.uint8)])
        args = [s.ctypes.data_as(_u8p), s.size, n]
        if ct == CT_MASK:
            args += [type_, self.mask_chars(mask20)]
        p = getattr(self.lib, self._DEC[ct])(*args)
        out = np.ctypeslib.as_array(p, shape=(n,)).copy()
        _libc.free(p)
        return out


def ref_mask_decode_defined_f64(type_, mask20, eb):
    """Where the reference's float64 CT=7 DECODER is well defined (dataCompression.c:1290-1296, :1355): the float32
    rule with 9 -> 12: 2 <= type <= 9; escape payload m == 0 or m > 8; no escape of total length 12."""
    n_eb = o_n_eb(eb)
    m = min(max(n_eb + ((mask20 >> 8) & 0x7FF) - 1023, 0), 52)
    if not (2 <= type_ <= 9):
        return False
    if 0 < m <= 8:
        return False
    k0 = max(m - 8, 0)
    return (type_ + 2 + k0 != 12) and (type_ + 2 + m != 12)


# ---- Hamming SECDED of CT=10 (oracle/dc_oracle_hamming.c; reference impl/dataCompression.c:5544-5878) ----
def _hm_oracle():
    lib = oracle()
    lib.dco_hm_length.restype = C.c_int
    lib.dco_hm_length.argtypes = [C.c_int]
    lib.dco_hamming_encode.restype = C.c_int
    lib.dco_hamming_encode.argtypes = [C.c_void_p, C.c_int, C.c_char_p]
    lib.dco_hamming_decode.restype = C.c_int
    lib.dco_hamming_decode.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int]
    return lib


def o_hm_length(k):
    return int(_hm_oracle().dco_hm_length(int(k)))


def o_hamming_encode(bits):
    """-> (r, check string of r + 1 chars '0'/'1')"""
    bits = np.ascontiguousarray(bits, dtype=np.uint8)
    buf = C.create_string_buffer(64)
    r = _hm_oracle().dco_hamming_encode(bits.ctypes.data, int(bits.size), buf)
    return int(r), buf.raw[:r + 1]


def o_hamming_decode(bits, c, r):
    """bits (uint8 array) and c (bytes) may hold errors -> (error_type, corrected bits, corrected c)"""
    bits = np.array(bits, dtype=np.uint8, copy=True)
    cb = C.create_string_buffer(bytes(c), len(c) + 1)
    t = _hm_oracle().dco_hamming_decode(bits.ctypes.data, cb, int(bits.size), int(r))
    return int(t), bits, cb.raw[:r + 1]


def ref_hamming_encode(bits, eb=0.01):
    """The compiled reference's hamming_encode (impl/dataCompression.c:5742)."""
    lib = RefLib(eb).lib
    bits = np.ascontiguousarray(bits, dtype=np.uint8)
    cp, r = C.c_char_p(), C.c_int()
    lib.hamming_encode.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.c_int, C.POINTER(C.c_int)]
    lib.hamming_encode.restype = None
    lib.hamming_encode(bits.ctypes.data, C.byref(cp), int(bits.size), C.byref(r))
    return r.value, C.string_at(cp, r.value + 1)


def ref_hamming_decode(bits, c, r, eb=0.01):
    """The compiled reference's hamming_decode (:5752); its messages go to the process's stdout."""
    lib = RefLib(eb).lib
    bits = np.array(bits, dtype=np.uint8, copy=True)
    cb = C.create_string_buffer(bytes(c), len(c) + 1)
    lib.hamming_decode.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int]
    lib.hamming_decode.restype = C.c_int
    t = lib.hamming_decode(bits.ctypes.data, cb, int(bits.size), int(r))
    return int(t), bits, cb.raw[:r + 1]


# ---- byte-wise codec CT=1 (oracle/dc_oracle_bytewise.c; reference myCompress :3977, myDecompress :3942) ----
def o_bytewise_compress(x, eb):
    """-> (array_float[nf], array_char[nc] as uint8, array_char_displacement[nc] int32)"""
    x = _f32(x)
    n = x.size
    af, ac, ad = np.zeros(max(n, 1), np.float32), np.zeros(max(n, 1), np.uint8), np.zeros(max(n, 1), np.int32)
    lib = oracle()
    lib.dco_bytewise_compress.restype = C.c_size_t
    lib.dco_bytewise_compress.argtypes = [C.c_void_p, C.c_size_t, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p]
    nf = lib.dco_bytewise_compress(_ptr(x), n, eb, _ptr(af), _ptr(ac), _ptr(ad))
    return af[:nf].copy(), ac[:n - nf].copy(), ad[:n - nf].copy()


def o_bytewise_decompress(af, ac, ad, n):
    out = np.zeros(max(n, 1), np.float32)
    af, ac, ad = _f32(af), np.ascontiguousarray(ac, np.uint8), np.ascontiguousarray(ad, np.int32)
    lib = oracle()
    lib.dco_bytewise_decompress.restype = None
    lib.dco_bytewise_decompress.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p]
    pad = lambda a: a if a.size else np.zeros(1, a.dtype)
    lib.dco_bytewise_decompress(_ptr(pad(af)), _ptr(pad(ac)), _ptr(pad(ad)), ac.size, n, _ptr(out))
    return out[:n]


def ref_bytewise_compress(x, eb):
    """The compiled reference's myCompress (impl/dataCompression.c:3977) with absErrorBound = eb."""
    lib = RefLib(eb).lib
    x = _f32(x)
    af, ac, ad = _f32p(), C.c_char_p(), C.POINTER(C.c_int)()
    lib.myCompress.argtypes = [C.c_void_p, C.POINTER(_f32p), C.POINTER(C.c_char_p), C.POINTER(C.POINTER(C.c_int)), C.c_int]
    lib.myCompress.restype = C.c_int
    nf = lib.myCompress(_ptr(x), C.byref(af), C.byref(ac), C.byref(ad), int(x.size))
    nc = x.size - nf
    f = np.ctypeslib.as_array(af, (nf,)).copy() if nf else np.zeros(0, np.float32)
    c = np.frombuffer(C.string_at(ac, nc), np.uint8).copy() if nc else np.zeros(0, np.uint8)
    d = np.ctypeslib.as_array(ad, (nc,)).copy() if nc else np.zeros(0, np.int32)
    return f, c, d


def ref_bytewise_decompress(af, ac, ad, n, eb):
    lib = RefLib(eb).lib
    af = _f32(np.concatenate([af, np.zeros(1, np.float32)]))
    ac = np.ascontiguousarray(np.concatenate([ac, np.zeros(1, np.uint8)]), np.uint8)
    ad = np.ascontiguousarray(np.concatenate([ad, np.zeros(1, np.int32)]), np.int32)       # the reference reads one past the end
    lib.myDecompress.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    lib.myDecompress.restype = _f32p
    p = lib.myDecompress(_ptr(af), _ptr(ac), _ptr(ad), int(n))
    return np.ctypeslib.as_array(p, (n,)).copy()


def o_bytewise_compress_f64(x, eb):
    x = _f64(x)
    n = x.size
    af, ac, ad = np.zeros(max(n, 1), np.float64), np.zeros(max(n, 1), np.uint8), np.zeros(max(n, 1), np.int32)
    lib = oracle()
    lib.dco_bytewise_compress_f64.restype = C.c_size_t
    lib.dco_bytewise_compress_f64.argtypes = [C.c_void_p, C.c_size_t, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p]
    nf = lib.dco_bytewise_compress_f64(_ptr(x), n, eb, _ptr(af), _ptr(ac), _ptr(ad))
    return af[:nf].copy(), ac[:n - nf].copy(), ad[:n - nf].copy()


def o_bytewise_decompress_f64(af, ac, ad, n):
    out = np.zeros(max(n, 1), np.float64)
    af, ac, ad = _f64(af), np.ascontiguousarray(ac, np.uint8), np.ascontiguousarray(ad, np.int32)
    lib = oracle()
    lib.dco_bytewise_decompress_f64.restype = None
    lib.dco_bytewise_decompress_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p]
    pad = lambda a: a if a.size else np.zeros(1, a.dtype)
    lib.dco_bytewise_decompress_f64(_ptr(pad(af)), _ptr(pad(ac)), _ptr(pad(ad)), ac.size, n, _ptr(out))
    return out[:n]


def ref_bytewise_compress_f64(x, eb):
    """The compiled reference's myCompress_double (impl/dataCompression.c:3815)."""
    lib = RefLib(eb).lib
    x = _f64(x)
    f64p = C.POINTER(C.c_double)
    af, ac, ad = f64p(), C.c_char_p(), C.POINTER(C.c_int)()
    lib.myCompress_double.argtypes = [C.c_void_p, C.POINTER(f64p), C.POINTER(C.c_char_p), C.POINTER(C.POINTER(C.c_int)), C.c_int]
    lib.myCompress_double.restype = C.c_int
    nf = lib.myCompress_double(_ptr(x), C.byref(af), C.byref(ac), C.byref(ad), int(x.size))
    nc = x.size - nf
    f = np.ctypeslib.as_array(af, (nf,)).copy() if nf else np.zeros(0, np.float64)
    c = np.frombuffer(C.string_at(ac, nc), np.uint8).copy() if nc else np.zeros(0, np.uint8)
    d = np.ctypeslib.as_array(ad, (nc,)).copy() if nc else np.zeros(0, np.int32)
    return f, c, d


def ref_bytewise_decompress_f64(af, ac, ad, n, eb):
    lib = RefLib(eb).lib
    af = _f64(np.concatenate([af, np.zeros(1, np.float64)]))
    ac = np.ascontiguousarray(np.concatenate([ac, np.zeros(1, np.uint8)]), np.uint8)
    ad = np.ascontiguousarray(np.concatenate([ad, np.zeros(1, np.int32)]), np.int32)
    lib.myDecompress_double.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    lib.myDecompress_double.restype = C.POINTER(C.c_double)
    p = lib.myDecompress_double(_ptr(af), _ptr(ac), _ptr(ad), int(n))
    return np.ctypeslib.as_array(p, (n,)).copy()
