"""Stand-in for compute-sanitizer, which is closed on this GPU pool (profiles/sanitizer_r2.txt): guard bands around every buffer
the kernels write (out-of-bounds stores), and bit-identical results over repeated runs into pattern-filled scratch (races,
reads of stale or uninitialised scratch).  Correctness of the first result is what the parity tests establish."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

from data_compression_b200 import api
from data_compression_b200.fields import make_field

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
GUARD = 4096
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class Guarded:
    """A byte buffer of `nbytes` (256-byte aligned start) inside a larger one, with GUARD bytes of 0xC3 on both sides."""

    def __init__(self, nbytes, dev, fill=0x5A):
        self.nbytes = nbytes
        self.raw = torch.full((nbytes + 2 * GUARD + 512,), 0xC3, dtype=torch.uint8, device=dev)
        base = self.raw.data_ptr()
        self.off = ((base + GUARD + 255) // 256) * 256 - base
        self.view = self.raw[self.off:self.off + nbytes]
        self.view.fill_(fill)

    def ptr(self):
        return C.c_void_p(self.view.data_ptr())

    def intact(self):
        lo = self.raw[:self.off]
        hi = self.raw[self.off + self.nbytes:]
        return bool((lo == 0xC3).all().item()) and bool((hi == 0xC3).all().item())


def run_once(L, dev, x, ct, eb, chunk, fill):
    n = x.numel()
    K = L.dcb_num_chunks(n, chunk)
    cap = L.dcb_max_stream_bytes(n, chunk)
    wsb = L.dcb_workspace_bytes(n, chunk)
    g = {"stream": Guarded(cap, dev, fill), "off": Guarded(8 * (K + 1), dev, fill), "min": Guarded(4 * K, dev, fill),
         "med": Guarded(4 * K, dev, fill), "type": Guarded(4 * K, dev, fill), "crc": Guarded(4 * K, dev, fill),
         "ws": Guarded(wsb, dev, fill), "out": Guarded(4 * n, dev, fill), "status": Guarded(4, dev, fill)}
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    flags = api.F_SUBTRACT_MIN | api.F_CRC
    rc = L.dcb_compress(ct, C.c_void_p(x.data_ptr()), n, chunk, eb, flags, 0, 0, g["stream"].ptr(), cap, g["off"].ptr(),
                        g["min"].ptr(), g["med"].ptr(), g["type"].ptr(), g["crc"].ptr(), g["ws"].ptr(), wsb, st)
    assert rc == 0, (rc, L.dcb_last_cuda_error())
    rc = L.dcb_decompress(ct, g["stream"].ptr(), g["off"].ptr(), n, chunk, eb, api.F_SUBTRACT_MIN, 0, 0, g["min"].ptr(),
                          g["med"].ptr(), g["type"].ptr(), g["out"].ptr(), g["status"].ptr(), st)
    assert rc == 0, (rc, L.dcb_last_cuda_error())
    torch.cuda.synchronize()
    for name, b in g.items():
        assert b.intact(), "guard band of `%s` was written" % name
    total = int(g["off"].view.view(torch.int64)[K].item())
    per_chunk_mask = ct == 7
    res = [g["stream"].view[:total].cpu().numpy().tobytes(), g["off"].view.cpu().numpy().tobytes(),
           g["min"].view.cpu().numpy().tobytes(), g["crc"].view.cpu().numpy().tobytes(), g["out"].view.cpu().numpy().tobytes()]
    if per_chunk_mask:
        res.append(g["type"].view.cpu().numpy().tobytes())
        res.append((g["med"].view.view(torch.int32) >> 15).cpu().numpy().tobytes())     # the 17 defined bits
    return res


CASES = [("mixed", 20 * 4096 + 777, 4096), ("mixed", 50001, 0), ("mixed", 3 * 8192 + 5, 8192), ("smooth", 33929, 0),
         ("noisy", 4097, 4096), ("mixed", 1000, 0)]


@pytest.mark.parametrize("ct", (5, 6, 7, 11))
def test_guard_bands_and_repeatability(ct):
    dev = torch.device("cuda:0")
    L = api.lib()
    rng = np.random.default_rng(3)
    for kind, n, chunk in CASES:
        if kind == "mixed":
            xn = np.concatenate([make_field("smooth", n // 2), (rng.random(n - n // 2) * 10).astype(np.float32)])
        else:
            xn = make_field(kind, n)
        base = torch.zeros(n + 8, dtype=torch.float32, device=dev)
        for shift in (0, 1):                                      # 16-byte aligned and unaligned input
            x = base[shift:shift + n]
            x.copy_(torch.from_numpy(xn))
            first = None
            for rep in range(12):
                res = run_once(L, dev, x, ct, 0.01, chunk, fill=(0x5A, 0x00, 0xFF)[rep % 3])
                if first is None:
                    first = res
                else:
                    assert all(a == b for a, b in zip(first, res)), (ct, kind, n, chunk, shift, rep)


@pytest.mark.parametrize("encoder", ("slots", "single"))
def test_guard_bands_other_encoders(encoder):
    """DCB_ENCODER is read once per process: the same test in a subprocess for the two non-default encoder kernels."""
    env = dict(os.environ, DCB_ENCODER=encoder)
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-q", "-x", "-k",
                        "test_guard_bands_and_repeatability and (5 or 7)", "-p", "no:cacheprovider"],
                       capture_output=True, text=True, env=env, cwd=ROOT, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:]
