"""Host-side binding of libdcb200.so (include/dc_b200.h) for Python callers.

PyTorch is plumbing only (device memory, streams, torch.distributed); every byte of codec work happens in
the hand-written sm_100a kernels behind the C ABI.  There is NO CPU path: loading fails loudly when the
library is missing, and every call raises when there is no CUDA device.

Naming follows the reference (impl/dataCompression.h): compress types CT=5/6/7/11 (+8/9 with CRC),
absErrorBound, streams / chunks / min / medium / type / mask.
"""
import ctypes as C
import os
from dataclasses import dataclass
from typing import Optional

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DCB200_LIB") or os.path.join(PKG, "libdcb200.so")   # override: A/B builds of the same ABI

CT_BITWISE, CT_BITWISE_NP, CT_BITWISE_MASK, CT_BITWISE_CRC, CT_MASK_CRC, CT_BITWISE_OP = 5, 6, 7, 8, 9, 11
F_SUBTRACT_MIN, F_FIXED_MASK, F_CRC, F_EXACT_MEDIUM = 1, 2, 4, 8
TILE_ELEMS = 4096
DEFAULT_CHUNK = 4096

ERRORS = {-1: "DCB_EINVAL (bad argument)", -2: "DCB_ECAPACITY (buffer too small)", -3: "DCB_ECUDA",
          -4: "DCB_ENODEVICE (no CUDA device; there is no CPU fallback)", -5: "DCB_EFORMAT (malformed stream)"}

EXPORTS = ["dcb_version", "dcb_last_cuda_error", "dcb_device_count", "dcb_n_eb", "dcb_num_chunks",
           "dcb_max_stream_bytes", "dcb_workspace_bytes", "dcb_compress", "dcb_decompress", "dcb_crc32_chunks",
           "dcb_chunk_stats", "dcb_chunk_medium", "dcb_flip_bits", "dcb_compress_host", "dcb_decompress_host",
           "dcb_last_stream_bits", "dcb_gather_face", "dcb_scatter_face", "dcb_himeno_jacobi",
           "dcb_himeno_copy_interior", "dcb_max_stream_bytes_f64", "dcb_workspace_bytes_f64", "dcb_compress_f64",
           "dcb_decompress_f64", "dcb_chunk_stats_f64", "dcb_profile_enable", "dcb_profile_last_ms",
           "dcb_hm_length", "dcb_hamming_encode", "dcb_hamming_decode",
           "dcb_bytewise_workspace_bytes", "dcb_bytewise_compress", "dcb_bytewise_decompress",
           "dcb_bytewise_compress_f64", "dcb_bytewise_decompress_f64", "dcb_static_max_elems", "dcb_compress_static", "dcb_host_release"]

_lib = None


class DcbError(RuntimeError):
    pass


def lib():
    """Loads libdcb200.so (no implicit build, no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise DcbError("libdcb200.so is not built: run `python -m data_compression_b200.build` "
                           "(or __graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        vp, u64, u32, i32, dbl = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int, C.c_double
        L.dcb_version.restype = C.c_char_p
        L.dcb_last_cuda_error.restype = C.c_char_p
        L.dcb_device_count.restype = i32
        L.dcb_n_eb.restype = i32
        L.dcb_n_eb.argtypes = [dbl]
        for f in ("dcb_num_chunks", "dcb_max_stream_bytes", "dcb_workspace_bytes"):
            getattr(L, f).restype = u64
            getattr(L, f).argtypes = [u64, u64]
        L.dcb_bytewise_workspace_bytes.restype = u64
        L.dcb_bytewise_workspace_bytes.argtypes = [u64, u64]
        L.dcb_bytewise_compress.restype = i32
        L.dcb_bytewise_compress.argtypes = [vp, u64, u64, dbl, vp, vp, vp, vp, vp, u64, vp]
        L.dcb_bytewise_decompress.restype = i32
        L.dcb_bytewise_decompress.argtypes = [vp, vp, vp, vp, u64, u64, vp, vp]
        L.dcb_bytewise_compress_f64.restype = i32
        L.dcb_bytewise_compress_f64.argtypes = [vp, u64, u64, dbl, vp, vp, vp, vp, vp, u64, vp]
        L.dcb_bytewise_decompress_f64.restype = i32
        L.dcb_bytewise_decompress_f64.argtypes = [vp, vp, vp, vp, u64, u64, vp, vp]
        L.dcb_hm_length.restype = i32
        L.dcb_hm_length.argtypes = [i32]
        L.dcb_hamming_encode.restype = i32
        L.dcb_hamming_encode.argtypes = [vp, u64, u32, vp, vp, vp]
        L.dcb_hamming_decode.restype = i32
        L.dcb_hamming_decode.argtypes = [vp, u64, u32, vp, vp, vp]
        L.dcb_profile_enable.restype = i32
        L.dcb_profile_enable.argtypes = [i32]
        L.dcb_profile_last_ms.restype = i32
        L.dcb_profile_last_ms.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.dcb_compress.restype = i32
        L.dcb_compress.argtypes = [i32, vp, u64, u64, dbl, u32, i32, u32, vp, u64, vp, vp, vp, vp, vp, vp, u64, vp]
        L.dcb_decompress.restype = i32
        L.dcb_decompress.argtypes = [i32, vp, vp, u64, u64, dbl, u32, i32, u32, vp, vp, vp, vp, vp, vp]
        L.dcb_static_max_elems.restype = u64
        L.dcb_compress_static.restype = i32
        L.dcb_compress_static.argtypes = [i32, vp, u64, u64, dbl, u32, i32, u32, vp, u64, vp, vp, vp, vp, vp, vp]
        L.dcb_crc32_chunks.restype = i32
        L.dcb_crc32_chunks.argtypes = [vp, vp, u64, vp, vp]
        L.dcb_chunk_stats.restype = i32
        L.dcb_chunk_stats.argtypes = [vp, u64, u64, i32, vp, vp, vp, vp]
        L.dcb_flip_bits.restype = i32
        L.dcb_flip_bits.argtypes = [vp, vp, u64, vp]
        L.dcb_gather_face.restype = i32
        L.dcb_gather_face.argtypes = [vp, i32, i32, i32, i32, i32, i32, i32, i32, vp, vp]
        L.dcb_himeno_jacobi.restype = i32
        L.dcb_himeno_jacobi.argtypes = [vp] * 7 + [i32] * 6 + [C.c_float, vp, vp]
        L.dcb_himeno_copy_interior.restype = i32
        L.dcb_himeno_copy_interior.argtypes = [vp, vp] + [i32] * 6 + [vp]
        L.dcb_scatter_face.restype = i32
        L.dcb_scatter_face.argtypes = [vp, i32, i32, i32, i32, i32, i32, i32, i32, vp, C.c_float, vp]
        for f in ("dcb_max_stream_bytes_f64", "dcb_workspace_bytes_f64"):
            getattr(L, f).restype = u64
            getattr(L, f).argtypes = [u64, u64]
        L.dcb_compress_f64.restype = i32
        L.dcb_compress_f64.argtypes = [i32, vp, u64, u64, dbl, u32, i32, u32, vp, u64, vp, vp, vp, vp, vp, vp, vp, u64, vp]
        L.dcb_chunk_stats_f64.restype = i32
        L.dcb_chunk_stats_f64.argtypes = [vp, u64, u64, i32, i32, vp, vp, vp, vp, vp]
        L.dcb_decompress_f64.restype = i32
        L.dcb_decompress_f64.argtypes = [i32, vp, vp, u64, u64, dbl, u32, i32, u32, vp, vp, vp, vp, vp, vp]
        L.dcb_compress_host.restype = i32
        L.dcb_compress_host.argtypes = [i32, vp, u64, u64, dbl, u32, i32, u32, vp, vp, vp, vp, vp, vp]
        L.dcb_decompress_host.restype = i32
        L.dcb_decompress_host.argtypes = [i32, vp, vp, u64, u64, dbl, u32, i32, u32, vp, vp, vp, vp]
        _lib = L
    return _lib


def check(rc, what):
    if rc != 0:
        extra = ""
        if rc == -3:
            extra = ": " + lib().dcb_last_cuda_error().decode()
        raise DcbError("%s failed: %s%s" % (what, ERRORS.get(rc, rc), extra))


def n_eb(abs_error_bound):
    return lib().dcb_n_eb(abs_error_bound)


def num_chunks(n, chunk_elems):
    return lib().dcb_num_chunks(n, chunk_elems)


def mask17_of(medium):
    """First 17 bits (sign, exponent, 8 mantissa bits) of `medium`: the char mask[17] of impl/pingpong.c:202-206."""
    return int(np.array([medium], dtype=np.float32).view(np.uint32)[0] >> 15)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _cur_stream(device):
    import torch
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


@dataclass
class Compressed:
    """Device-resident result of compress(): the reference's byte stream(s) plus the chunk table."""
    ct: int
    n: int
    chunk_elems: int
    abs_error_bound: float
    flags: int
    fixed_type: int
    fixed_mask17: int
    stream: "object"                  # uint8 [capacity]; valid bytes = total_bytes()
    chunk_offsets: "object"           # uint64-as-int64 [K+1]
    chunk_min: Optional["object"]     # float32 [K]
    chunk_medium: Optional["object"]  # float32 [K] (CT=7/9 per-chunk mask)
    chunk_type: Optional["object"]    # int32 [K]
    chunk_crc: Optional["object"]     # uint32-as-int32 [K] (CT=8/9)

    @property
    def num_chunks(self):
        return self.chunk_offsets.numel() - 1

    def total_bytes(self):
        return int(self.chunk_offsets[-1].item())

    def ratio(self):
        return 4.0 * self.n / max(self.total_bytes(), 1)

    def stream_bytes(self):
        return self.stream[: self.total_bytes()]


class Workspace:
    """Reusable device buffers for one (n, chunk_elems) shape, so that a timed loop does no allocation."""

    def __init__(self, n, chunk_elems, device):
        import torch
        L = lib()
        self.n, self.chunk_elems, self.device = n, chunk_elems, device
        K = L.dcb_num_chunks(n, chunk_elems)
        self.K = K
        self.capacity = L.dcb_max_stream_bytes(n, chunk_elems)
        self.stream = torch.empty(self.capacity, dtype=torch.uint8, device=device)
        self.offsets = torch.zeros(K + 1, dtype=torch.int64, device=device)
        self.cmin = torch.zeros(max(K, 1), dtype=torch.float32, device=device)
        self.cmed = torch.zeros(max(K, 1), dtype=torch.float32, device=device)
        self.ctype = torch.zeros(max(K, 1), dtype=torch.int32, device=device)
        self.ccrc = torch.zeros(max(K, 1), dtype=torch.int32, device=device)
        self.ws_bytes = L.dcb_workspace_bytes(n, chunk_elems)
        self.ws = torch.empty(self.ws_bytes, dtype=torch.uint8, device=device)
        self.status = torch.zeros(1, dtype=torch.int32, device=device)


def compress(x, ct=CT_BITWISE, abs_error_bound=1e-6, chunk_elems=DEFAULT_CHUNK, subtract_min=True,
             fixed_type=None, fixed_mask17=None, crc=False, workspace=None, exact_medium=False):
    """x: contiguous float32 CUDA tensor.  One chunk == one reference call myCompress_bitwise*(..., (NULL,0,8))."""
    import torch
    if not (x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()):
        raise DcbError("compress() needs a contiguous float32 CUDA tensor (there is no CPU path)")
    n = x.numel()
    ws = workspace or Workspace(n, chunk_elems, x.device)
    flags = (F_SUBTRACT_MIN if subtract_min else 0) | (F_CRC if crc else 0) | (F_EXACT_MEDIUM if exact_medium else 0)
    ft, fm = 0, 0
    if fixed_type is not None:
        flags |= F_FIXED_MASK
        ft, fm = int(fixed_type), int(fixed_mask17)
    with torch.cuda.device(x.device):
        rc = lib().dcb_compress(ct, _ptr(x), n, chunk_elems, abs_error_bound, flags, ft, fm, _ptr(ws.stream),
                                ws.capacity, _ptr(ws.offsets), _ptr(ws.cmin), _ptr(ws.cmed), _ptr(ws.ctype),
                                _ptr(ws.ccrc), _ptr(ws.ws), ws.ws_bytes, _cur_stream(x.device))
    check(rc, "dcb_compress")
    has_crc = crc or ct in (CT_BITWISE_CRC, CT_MASK_CRC)
    per_chunk_mask = ct in (CT_BITWISE_MASK, CT_MASK_CRC) and fixed_type is None
    return Compressed(ct, n, chunk_elems, abs_error_bound, flags, ft, fm, ws.stream, ws.offsets,
                      ws.cmin if subtract_min else None, ws.cmed if per_chunk_mask else None,
                      ws.ctype if per_chunk_mask else None, ws.ccrc if has_crc else None)


def decompress(comp, out=None, status=None, check_status=True):
    """status: device int32[1] receiving DCB_OK / DCB_EFORMAT (asynchronous; the caller reads it).  Without it the call owns
    the status word and, unless check_status=False, reads it back (one synchronisation) and raises on a malformed stream -- the
    condition on which the reference prints "Error leading bit 1" and exits (dataCompression.c:2501-2505)."""
    import torch
    dev = comp.stream.device
    if out is None:
        out = torch.empty(comp.n, dtype=torch.float32, device=dev)
    own_status = status is None
    if own_status:
        status = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = lib().dcb_decompress(comp.ct, _ptr(comp.stream), _ptr(comp.chunk_offsets), comp.n, comp.chunk_elems,
                                  comp.abs_error_bound, comp.flags & 3, comp.fixed_type, comp.fixed_mask17,
                                  _ptr(comp.chunk_min), _ptr(comp.chunk_medium), _ptr(comp.chunk_type), _ptr(out),
                                  _ptr(status), _cur_stream(dev))
    check(rc, "dcb_decompress")
    if own_status and check_status and int(status.item()) != 0:
        raise DcbError("dcb_decompress: %s" % ERRORS.get(int(status.item()), status.item()))
    return out


# ---- float64 codecs (the reference's *_double entry points, impl/k-means.c / mm.c / lu.c) ----------
def mask20_of(medium):
    """First 1+11+8 bits of `medium`: the char mask[20] of impl/k-means.c:527-530."""
    return int(np.array([medium], dtype=np.float64).view(np.uint64)[0] >> np.uint64(44))


@dataclass
class CompressedF64:
    ct: int
    n: int
    chunk_elems: int
    abs_error_bound: float
    flags: int
    fixed_type: int
    fixed_mask20: int
    stream: "object"                  # uint8 [capacity]
    chunk_offsets: "object"           # int64 [K+1]
    chunk_min: Optional["object"]     # float64 [K]
    chunk_mask20: Optional["object"]  # int32 [K] (CT=7/9 per-chunk mask)
    chunk_type: Optional["object"]    # int32 [K]
    chunk_medium: Optional["object"]  # float64 [K] (only with exact_medium)
    chunk_crc: Optional["object"]     # int32 [K]

    @property
    def num_chunks(self):
        return self.chunk_offsets.numel() - 1

    def total_bytes(self):
        return int(self.chunk_offsets[-1].item())

    def ratio(self):
        return 8.0 * self.n / max(self.total_bytes(), 1)


class WorkspaceF64:
    def __init__(self, n, chunk_elems, device):
        import torch
        L = lib()
        K = L.dcb_num_chunks(n, chunk_elems)
        self.K = K
        self.capacity = L.dcb_max_stream_bytes_f64(n, chunk_elems)
        self.stream = torch.empty(self.capacity, dtype=torch.uint8, device=device)
        self.offsets = torch.zeros(K + 1, dtype=torch.int64, device=device)
        self.cmin = torch.zeros(max(K, 1), dtype=torch.float64, device=device)
        self.cmed = torch.zeros(max(K, 1), dtype=torch.float64, device=device)
        self.cmask = torch.zeros(max(K, 1), dtype=torch.int32, device=device)
        self.ctype = torch.zeros(max(K, 1), dtype=torch.int32, device=device)
        self.ccrc = torch.zeros(max(K, 1), dtype=torch.int32, device=device)
        self.ws_bytes = L.dcb_workspace_bytes_f64(n, chunk_elems)
        self.ws = torch.empty(self.ws_bytes, dtype=torch.uint8, device=device)


def compress_f64(x, ct=CT_BITWISE, abs_error_bound=1e-6, chunk_elems=DEFAULT_CHUNK, subtract_min=True,
                 fixed_type=None, fixed_mask20=None, crc=False, workspace=None, exact_medium=False):
    """x: contiguous float64 CUDA tensor.  One chunk == one reference call myCompress_bitwise_double*(..., (NULL,0,8))."""
    import torch
    if not (x.is_cuda and x.dtype == torch.float64 and x.is_contiguous()):
        raise DcbError("compress_f64() needs a contiguous float64 CUDA tensor (there is no CPU path)")
    n = x.numel()
    ws = workspace or WorkspaceF64(n, chunk_elems, x.device)
    flags = (F_SUBTRACT_MIN if subtract_min else 0) | (F_CRC if crc else 0) | (F_EXACT_MEDIUM if exact_medium else 0)
    ft, fm = 0, 0
    if fixed_type is not None:
        flags |= F_FIXED_MASK
        ft, fm = int(fixed_type), int(fixed_mask20)
    with torch.cuda.device(x.device):
        rc = lib().dcb_compress_f64(ct, _ptr(x), n, chunk_elems, abs_error_bound, flags, ft, fm, _ptr(ws.stream),
                                    ws.capacity, _ptr(ws.offsets), _ptr(ws.cmin), _ptr(ws.cmask), _ptr(ws.ctype),
                                    _ptr(ws.cmed), _ptr(ws.ccrc), _ptr(ws.ws), ws.ws_bytes, _cur_stream(x.device))
    check(rc, "dcb_compress_f64")
    has_crc = crc or ct in (CT_BITWISE_CRC, CT_MASK_CRC)
    per_chunk_mask = ct in (CT_BITWISE_MASK, CT_MASK_CRC) and fixed_type is None
    return CompressedF64(ct, n, chunk_elems, abs_error_bound, flags, ft, fm, ws.stream, ws.offsets,
                         ws.cmin if subtract_min else None, ws.cmask if per_chunk_mask else None,
                         ws.ctype if per_chunk_mask else None, ws.cmed if (per_chunk_mask and exact_medium) else None,
                         ws.ccrc if has_crc else None)


def decompress_f64(comp, out=None, status=None):
    import torch
    dev = comp.stream.device
    if out is None:
        out = torch.empty(comp.n, dtype=torch.float64, device=dev)
    own_status = status is None
    if own_status:
        status = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = lib().dcb_decompress_f64(comp.ct, _ptr(comp.stream), _ptr(comp.chunk_offsets), comp.n, comp.chunk_elems,
                                      comp.abs_error_bound, comp.flags & 3, comp.fixed_type, comp.fixed_mask20,
                                      _ptr(comp.chunk_min), _ptr(comp.chunk_mask20), _ptr(comp.chunk_type), _ptr(out),
                                      _ptr(status), _cur_stream(dev))
    check(rc, "dcb_decompress_f64")
    if own_status and int(status.item()) != 0:
        raise DcbError("dcb_decompress_f64: malformed stream (status %d)" % int(status.item()))
    return out


def crc32_chunks(stream, chunk_offsets):
    import torch
    K = chunk_offsets.numel() - 1
    out = torch.zeros(max(K, 1), dtype=torch.int32, device=stream.device)
    with torch.cuda.device(stream.device):
        check(lib().dcb_crc32_chunks(_ptr(stream), _ptr(chunk_offsets), K, _ptr(out), _cur_stream(stream.device)),
              "dcb_crc32_chunks")
    return out[:K]


def flip_bits(stream, bit_indices):
    """bit_flip test hook (dataCompression.c:5858): flips the given stream bits in place."""
    import torch
    idx = torch.as_tensor(bit_indices, dtype=torch.int64, device=stream.device)
    with torch.cuda.device(stream.device):
        check(lib().dcb_flip_bits(_ptr(stream), _ptr(idx), idx.numel(), _cur_stream(stream.device)), "dcb_flip_bits")


def bytewise_compress(x, abs_error_bound=1e-6, chunk_elems=DEFAULT_CHUNK, buffers=None):
    """myCompress (dataCompression.c:3977) per chunk on the device.  -> dict(array_float, array_char, displacement: fixed-pitch
    tensors of n entries, chunk c at c*chunk_elems; nfloat[K]; n; chunk_elems; abs_error_bound).  `buffers`: the dict of an
    earlier call on an input of the same size and type, whose tensors (and workspace) are written again instead of new ones."""
    import torch
    n = x.numel()
    K = max(num_chunks(n, chunk_elems), 1)
    dev = x.device
    f64 = x.dtype == torch.float64
    wsb = lib().dcb_bytewise_workspace_bytes(n, chunk_elems)
    if buffers is not None:
        if buffers["n"] != n or buffers["chunk_elems"] != chunk_elems or buffers["array_float"].dtype != x.dtype:
            raise ValueError("bytewise_compress: buffers of another shape")
        af, ac, ad, nf, ws = (buffers[k] for k in ("array_float", "array_char", "displacement", "nfloat", "workspace"))
    else:
        af = torch.zeros(max(n, 1), dtype=x.dtype, device=dev)
        ac = torch.zeros(max(n, 1), dtype=torch.uint8, device=dev)
        ad = torch.zeros(max(n, 1), dtype=torch.int32, device=dev)
        nf = torch.zeros(K, dtype=torch.int32, device=dev)
        ws = torch.empty(max(wsb, 256), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        fn = lib().dcb_bytewise_compress_f64 if f64 else lib().dcb_bytewise_compress
        check(fn(_ptr(x), n, chunk_elems, abs_error_bound, _ptr(af), _ptr(ac), _ptr(ad), _ptr(nf), _ptr(ws), wsb,
                 _cur_stream(dev)), "dcb_bytewise_compress")
    return dict(array_float=af, array_char=ac, displacement=ad, nfloat=nf, n=n, chunk_elems=chunk_elems,
                abs_error_bound=abs_error_bound, workspace=ws)


def bytewise_decompress(enc, out=None):
    """myDecompress (dataCompression.c:3942) per chunk on the device."""
    import torch
    n = enc["n"]
    dev = enc["array_float"].device
    f64 = enc["array_float"].dtype == torch.float64
    if out is None:
        out = torch.empty(n, dtype=enc["array_float"].dtype, device=dev)
    with torch.cuda.device(dev):
        fn = lib().dcb_bytewise_decompress_f64 if f64 else lib().dcb_bytewise_decompress
        check(fn(_ptr(enc["array_float"]), _ptr(enc["array_char"]), _ptr(enc["displacement"]), _ptr(enc["nfloat"]), n,
                 enc["chunk_elems"], _ptr(out), _cur_stream(dev)), "dcb_bytewise_decompress")
    return out


def hm_length(k):
    """hmLength (dataCompression.c:5587): check bits of a Hamming block of k data bits."""
    return int(lib().dcb_hm_length(int(k)))


def hamming_encode(stream, total_bytes, block_bytes):
    """hamming_encode (dataCompression.c:5742) of every block of a finished device stream, as impl/pingpong.c:296-311 calls it.
    Returns (check, r): int32 tensors [blocks]; check[b] bit i = c[i] (i < r[b]), bit r[b] = the SECDED parity bit."""
    import torch
    nb = max((int(total_bytes) + int(block_bytes) - 1) // int(block_bytes), 1)
    chk = torch.zeros(nb, dtype=torch.int32, device=stream.device)
    r = torch.zeros(nb, dtype=torch.int32, device=stream.device)
    with torch.cuda.device(stream.device):
        check(lib().dcb_hamming_encode(_ptr(stream), int(total_bytes), int(block_bytes), _ptr(chk), _ptr(r),
                                       _cur_stream(stream.device)), "dcb_hamming_encode")
    return chk, r


def hamming_decode(stream, total_bytes, block_bytes, chk):
    """hamming_decode (dataCompression.c:5752) of every block: single-bit errors are rectified in `stream` / `chk` in place.
    Returns the error type of every block (0 none, 1 two-bit, 2 parity bit, 3 one bit)."""
    import torch
    et = torch.zeros(chk.numel(), dtype=torch.int32, device=stream.device)
    with torch.cuda.device(stream.device):
        check(lib().dcb_hamming_decode(_ptr(stream), int(total_bytes), int(block_bytes), _ptr(chk), _ptr(et),
                                       _cur_stream(stream.device)), "dcb_hamming_decode")
    return et


def gather_face(p, ijk, v, imax, jmax, kmax, out=None):
    """transform_3d_array_to_1d_array (dataCompression.c:3741-3775) on the device: p is a contiguous float32 CUDA tensor
    [MI, MJ, MK]; returns the face i/j/k == v (ijk = 1/2/3) as a 1-D tensor in the reference's loop order."""
    import torch
    MI, MJ, MK = p.shape
    n = {1: jmax * kmax, 2: imax * kmax, 3: imax * jmax}.get(ijk, 1)    # (an invalid ijk is rejected by the library)
    if out is None:
        out = torch.empty(n, dtype=torch.float32, device=p.device)
    with torch.cuda.device(p.device):
        check(lib().dcb_gather_face(_ptr(p), MI, MJ, MK, ijk, v, imax, jmax, kmax, _ptr(out), _cur_stream(p.device)),
              "dcb_gather_face")
    return out


def scatter_face(p, ijk, v, imax, jmax, kmax, values, add=0.0):
    """p[face] = values[pointer++] + add (himenoBMTxps.c:560-568): the write-back of a received halo face."""
    import torch
    MI, MJ, MK = p.shape
    with torch.cuda.device(p.device):
        check(lib().dcb_scatter_face(_ptr(p), MI, MJ, MK, ijk, v, imax, jmax, kmax, _ptr(values), float(add),
                                     _cur_stream(p.device)), "dcb_scatter_face")


def himeno_jacobi(p, a, b, c, bnd, wrk1, wrk2, imax, jmax, kmax, omega=0.8, gosa=None):
    """One Jacobi sweep of impl/himenoBMTxps.c:247-284 (without the halo exchange): wrk2 on the interior, then p = wrk2.
    Returns the device scalar (float64) sum ss^2 of this rank."""
    import torch
    MI, MJ, MK = p.shape
    if gosa is None:
        gosa = torch.zeros(1, dtype=torch.float64, device=p.device)
    with torch.cuda.device(p.device):
        st = _cur_stream(p.device)
        check(lib().dcb_himeno_jacobi(_ptr(p), _ptr(a), _ptr(b), _ptr(c), _ptr(bnd), _ptr(wrk1), _ptr(wrk2), MI, MJ, MK,
                                      imax, jmax, kmax, float(omega), _ptr(gosa), st), "dcb_himeno_jacobi")
        check(lib().dcb_himeno_copy_interior(_ptr(p), _ptr(wrk2), MI, MJ, MK, imax, jmax, kmax, st), "dcb_himeno_copy_interior")
    return gosa


# ---- host-pointer layer (numpy in, numpy out; H2D/D2H inside the C library) -------------------------------
def compress_host(x, ct=CT_BITWISE, abs_error_bound=1e-6, chunk_elems=DEFAULT_CHUNK, subtract_min=True,
                  fixed_type=None, fixed_mask17=None, crc=False, out=None):
    x = np.ascontiguousarray(x, dtype=np.float32)
    n = x.size
    L = lib()
    K = L.dcb_num_chunks(n, chunk_elems)
    o = out or {}
    stream = o.get("stream")
    if stream is None:
        stream = np.empty(L.dcb_max_stream_bytes(n, chunk_elems), dtype=np.uint8)
    offs = np.zeros(K + 1, dtype=np.uint64)
    cmin, cmed = np.zeros(max(K, 1), np.float32), np.zeros(max(K, 1), np.float32)
    ctype, ccrc = np.zeros(max(K, 1), np.int32), np.zeros(max(K, 1), np.uint32)
    flags = (F_SUBTRACT_MIN if subtract_min else 0) | (F_CRC if crc else 0)
    ft, fm = 0, 0
    if fixed_type is not None:
        flags |= F_FIXED_MASK
        ft, fm = int(fixed_type), int(fixed_mask17)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    check(L.dcb_compress_host(ct, vp(x), n, chunk_elems, abs_error_bound, flags, ft, fm, vp(stream), vp(offs),
                              vp(cmin), vp(cmed), vp(ctype), vp(ccrc)), "dcb_compress_host")
    return dict(ct=ct, n=n, chunk_elems=chunk_elems, abs_error_bound=abs_error_bound, flags=flags, fixed_type=ft,
                fixed_mask17=fm, stream=stream, total_bytes=int(offs[K]) if n else 0, offsets=offs,
                chunk_min=cmin[:K], chunk_medium=cmed[:K], chunk_type=ctype[:K], chunk_crc=ccrc[:K])


def decompress_host(enc, out=None):
    L = lib()
    n = enc["n"]
    if out is None:
        out = np.empty(n, dtype=np.float32)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    per_chunk_mask = enc["ct"] in (CT_BITWISE_MASK, CT_MASK_CRC) and not (enc["flags"] & F_FIXED_MASK)
    check(L.dcb_decompress_host(enc["ct"], vp(enc["stream"]), vp(enc["offsets"]), n, enc["chunk_elems"],
                                enc["abs_error_bound"], enc["flags"] & 3, enc["fixed_type"], enc["fixed_mask17"],
                                vp(enc["chunk_min"]) if (enc["flags"] & F_SUBTRACT_MIN) else None,
                                vp(enc["chunk_medium"]) if per_chunk_mask else None,
                                vp(enc["chunk_type"]) if per_chunk_mask else None, vp(out)), "dcb_decompress_host")
    return out
