"""dcb_compress_static (include/dc_b200.h): the allocation-free variant of dcb_compress whose scratch is a static device
array inside the library.  Called through the C ABI (ctypes) and compared with dcb_compress, bit for bit (the tolerance of
every codec test here), on two calls in a row -- the second call finds the first call's contents in the shared array.
Collected last on purpose (file name), and self-contained: it passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from data_compression_b200 import api
from data_compression_b200.fields import make_field

pytestmark = pytest.mark.gpu


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _compress_both(L, torch, dev, x_np, ct, eb, chunk, flags):
    """-> (outputs of dcb_compress, outputs of dcb_compress_static), each a dict of numpy arrays."""
    n = x_np.size
    K = L.dcb_num_chunks(n, chunk)
    cap = L.dcb_max_stream_bytes(n, chunk)
    x = torch.from_numpy(x_np).to(dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    res = []
    for static in (False, True):
        # outputs pre-filled with a pattern, so that anything the call does not write shows up as a difference
        s = torch.full((cap,), 0xA5, dtype=torch.uint8, device=dev)
        off = torch.full((K + 1,), -1, dtype=torch.int64, device=dev)
        mn = torch.full((K,), -7.0, dtype=torch.float32, device=dev)
        med = torch.full((K,), -7.0, dtype=torch.float32, device=dev)
        ty = torch.full((K,), -7, dtype=torch.int32, device=dev)
        crc = torch.full((K,), -7, dtype=torch.int32, device=dev)
        if static:
            rc = L.dcb_compress_static(ct, _ptr(x), n, chunk, eb, flags, 0, 0, _ptr(s), cap, _ptr(off), _ptr(mn), _ptr(med),
                                       _ptr(ty), _ptr(crc), stream)
        else:
            wsb = L.dcb_workspace_bytes(n, chunk)
            ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
            rc = L.dcb_compress(ct, _ptr(x), n, chunk, eb, flags, 0, 0, _ptr(s), cap, _ptr(off), _ptr(mn), _ptr(med),
                                _ptr(ty), _ptr(crc), _ptr(ws), wsb, stream)
        assert rc == 0, (static, rc, L.dcb_last_cuda_error())
        torch.cuda.synchronize()
        total = int(off[-1].item())
        res.append({"total": total, "stream": s[:total].cpu().numpy(), "off": off.cpu().numpy(), "min": mn.cpu().numpy(),
                    "med": med.cpu().numpy(), "type": ty.cpu().numpy(), "crc": crc.cpu().numpy()})
    return res


def test_static_scratch_matches_dcb_compress_twice_in_a_row():
    import torch
    assert torch.cuda.is_available()
    dev = torch.device("cuda:0")
    L = api.lib()
    assert L.dcb_static_max_elems() == 1 << 20
    # two calls in a row on different data and shapes: a large noisy field (long chunks fill the slots), then a shorter
    # smooth one with a ragged last chunk (short chunks; the slots still hold the first call's bytes), CT=9 = mask + CRC
    calls = [(np.concatenate([make_field("noisy", 300000), make_field("smooth", 100000)]), 9, 0.01, 4096),
             (make_field("smooth", 123457), 9, 1e-4, 4096),
             (make_field("noisy", 50000), 5, 0.01, 0)]                 # one chunk = the legacy call shape (single-pass encoder)
    for x_np, ct, eb, chunk in calls:
        ref, got = _compress_both(L, torch, dev, x_np, ct, eb, chunk, api.F_SUBTRACT_MIN)
        assert got["total"] == ref["total"] and got["total"] > 0
        for k in ("stream", "off", "min") + (("med", "type", "crc") if ct == 9 else ()):
            a, b = ref[k], got[k]
            assert np.array_equal(a.view(np.uint8), b.view(np.uint8)), (k, ct, eb, chunk)


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    import torch
    dev = torch.device("cuda:0")
    L = api.lib()
    n = (1 << 20) + 4096
    x = torch.zeros(n, dtype=torch.float32, device=dev)
    K = L.dcb_num_chunks(n, 4096)
    cap = L.dcb_max_stream_bytes(n, 4096)
    s = torch.empty(cap, dtype=torch.uint8, device=dev)
    off = torch.zeros(K + 1, dtype=torch.int64, device=dev)
    mn = torch.zeros(K, dtype=torch.float32, device=dev)
    rc = L.dcb_compress_static(5, _ptr(x), n, 4096, 0.01, api.F_SUBTRACT_MIN, 0, 0, _ptr(s), cap, _ptr(off), _ptr(mn), None, None,
                               None, C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    assert rc == -2                                                     # DCB_ECAPACITY
    msg = L.dcb_last_cuda_error().decode()
    assert "dcb_compress_static" in msg and "static array" in msg
    torch.cuda.synchronize()
