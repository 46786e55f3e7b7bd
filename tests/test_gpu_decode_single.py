"""The intra-chunk parallel decoder (csrc/dcb_decode_single.cuh): ONE chunk = ONE reference call on the whole array
(myDecompress_bitwise*, dataCompression.c:2922 / :2459 / :698 / :1703; impl/pingpong.c:553-566).  Decoded floats must equal the
oracle's bit for bit, and the lane-per-chunk kernel's (DCB_DECODER=lane is checked in a subprocess)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle_lib as ol
from data_compression_b200 import api
from data_compression_b200.fields import make_field

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.fail("no CUDA device")
    return torch.device("cuda:0")


def gpu_decode_stream(stream, n, ct, eb, dev, fixed=None, cmin=None):
    """Decode a raw byte stream as ONE chunk through dcb_decompress."""
    L = api.lib()
    cap = max(stream.size + 64, 64)
    s = torch.zeros(cap, dtype=torch.uint8, device=dev)
    s[:stream.size] = torch.from_numpy(stream).to(dev)
    off = torch.tensor([0, stream.size], dtype=torch.int64, device=dev)
    out = torch.full((n,), 12345.0, dtype=torch.float32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    flags, ft, fm = 0, 0, 0
    if fixed is not None:
        flags |= api.F_FIXED_MASK
        ft, fm = fixed
    mn = None
    if cmin is not None:
        flags |= api.F_SUBTRACT_MIN
        mn = torch.tensor([cmin], dtype=torch.float32, device=dev)
    rc = L.dcb_decompress(ct, C.c_void_p(s.data_ptr()), C.c_void_p(off.data_ptr()), n, 0, eb, flags, ft, fm,
                          C.c_void_p(mn.data_ptr()) if mn is not None else None, None, None, C.c_void_p(out.data_ptr()),
                          C.c_void_p(status.data_ptr()), C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    assert rc == 0, rc
    torch.cuda.synchronize()
    return out.cpu().numpy(), int(status.item())


def oracle_decode_stream(stream, n, ct, eb, fixed=None, cmin=None):
    enc = dict(stream=stream, offsets=np.array([0, stream.size], dtype=np.uint64),
               chunk_min=np.array([cmin if cmin is not None else 0.0], dtype=np.float32),
               chunk_medium=np.zeros(1, np.float32), chunk_type=np.zeros(1, np.int32))
    ft, fm = fixed if fixed is not None else (-1, 0)
    return ol.o_decompress_chunked(ct, enc, n, n, eb, add_min=cmin is not None, fixed_type=ft, fixed_mask17=fm)


@pytest.mark.parametrize("ct", (5, 6, 7, 11))
@pytest.mark.parametrize("kind", ("smooth", "noisy", "constant", "mixed"))
def test_fields_single_chunk(ct, kind, dev):
    rng = np.random.default_rng(17)
    for n in (4097, 33929, 131072, 300001):
        if kind == "mixed":
            x = (rng.random(n) * 10).astype(np.float32)
            for lo in range(0, n - 3000, 9000):
                x[lo:lo + 3000] = make_field("smooth", 3000, start=lo)
            x[n // 2: n // 2 + 1500] = 0.25
        else:
            x = make_field(kind, n)
        for eb in (0.01, 1e-4):
            xd = torch.from_numpy(x).to(dev)
            comp = api.compress(xd, ct=ct, abs_error_bound=eb, chunk_elems=0)
            d = api.decompress(comp).cpu().numpy()
            enc = ol.o_compress_chunked(ct, x, n, eb)
            od = ol.o_decompress_chunked(ct, enc, n, n, eb)
            assert np.array_equal(d.view(np.uint32), od.view(np.uint32)), (ct, kind, n, eb)


@pytest.mark.parametrize("ct", (5, 7, 11))
def test_random_bit_strings_decode_like_the_oracle(ct, dev):
    """Any bit string is a stream of CT=5/7/11 codes in some order: raw codes of every length, 3-bit codes on every history
    state, escapes.  (CT=6 rejects a leading 1; it is covered by the field tests.)"""
    rng = np.random.default_rng(23 + ct)
    for trial in range(6):
        nbytes = int(rng.integers(30000, 90000))
        stream = rng.integers(0, 256, nbytes, dtype=np.uint8)
        if trial % 2:                                   # lots of 3-bit codes: set the top bit of many bytes
            stream |= (rng.random(nbytes) < 0.6).astype(np.uint8) << 7
        n = nbytes * 8 // 32 - 4                         # every code has at most 32 bits: n codes certainly fit
        eb = (0.01, 1e-6, 0.1)[trial % 3]
        fixed = (2 + trial % 5, int(rng.integers(0, 1 << 16))) if ct == 7 else None
        got, st = gpu_decode_stream(stream, n, ct, eb, dev, fixed=fixed, cmin=0.5 if trial == 3 else None)
        ref = oracle_decode_stream(stream, n, ct, eb, fixed=fixed, cmin=0.5 if trial == 3 else None)
        # NaN payloads are outside the comparison: a raw code may carry exponent 255, and x86 and the GPU propagate payloads
        # differently through the predictors (NaN is outside the format's domain; the PARSE after it must still agree)
        differ = (got.view(np.uint32) != ref.view(np.uint32)) & ~(np.isnan(got) & np.isnan(ref))
        ne = np.nonzero(differ)[0]
        assert ne.size == 0, (ct, trial, ne[:5], got[ne[:5]], ref[ne[:5]])


def _bits_to_bytes(bits):
    bits = bits + "0" * (-len(bits) % 8)
    return np.frombuffer(int(bits, 2).to_bytes(len(bits) // 8, "big"), dtype=np.uint8).copy()


def test_marker_values_inside_a_long_chunk(dev):
    """-1.0f is the reference's 'empty history slot' marker (:3110-3127): a predicted value that comes out as exactly -1.0f is
    not shifted into the history like other values.  CT=11 stream built by hand: 1.0 (full word), 0 (100), 2*0 - 1 = -1 (110),
    then predicted codes on that state; repeated between stretches of other codes so that it also falls inside the batches of
    eight predicted codes."""
    rng = np.random.default_rng(31)
    w = lambda f: format(int(np.float32(f).view(np.uint32)), "032b")
    pieces, n = [], 0
    for rep in range(700):
        pad = int(rng.integers(0, 9))
        seq = [w(1.0), "100", "110"] + ["101", "110", "111", "110", "111", "101", "111", "110", "111"][: 4 + rep % 6]
        seq += ["111"] * pad + [w(rng.random() * 4), w(rng.random() * 4), w(rng.random() * 4)] + ["110", "111"] * (rep % 5)
        pieces += seq
        n += len(seq)
    stream = _bits_to_bytes("".join(pieces))
    assert n > 4096
    got, _ = gpu_decode_stream(stream, n, 11, 0.01, dev)
    ref = oracle_decode_stream(stream, n, 11, 0.01)
    assert (ref == -1.0).sum() >= 700
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))


def test_same_result_as_lane_per_chunk_kernel(dev):
    """DCB_DECODER=lane forces the serial kernel (the env var is read once per process, hence the subprocess)."""
    code = r'''
import sys, numpy as np, torch
sys.path.insert(0, %r); sys.path.insert(0, %r)
from data_compression_b200 import api
from data_compression_b200.fields import make_field
x = np.concatenate([make_field("noisy", 40000), make_field("smooth", 60000)])
xd = torch.from_numpy(x).cuda()
for ct in (5, 6, 7, 11):
    comp = api.compress(xd, ct=ct, abs_error_bound=0.001, chunk_elems=0)
    print(ct, int(api.decompress(comp).view(torch.int32).to(torch.int64).sum().item()))
''' % (ROOT, os.path.join(ROOT, "tests"))
    outs = []
    for mode in ("lane", "parallel"):
        env = dict(os.environ, DCB_DECODER=mode)
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(r.stdout)
    assert outs[0] == outs[1] and len(outs[0].split()) == 8


def test_long_chunk_medium_is_the_sequential_sum(dev):
    """CT=7 on ONE long chunk: chunk_medium must be med_dataset_float's sequential float32 mean bit for bit (:3593-3620) although
    a warp computes it (dcb_stats_kernel -> exact_medium), on data that crosses binades, ties, zeros, negatives."""
    rng = np.random.default_rng(53)
    n = 150001
    fields = {
        "noisy": make_field("noisy", n), "smooth": make_field("smooth", n), "constant": make_field("constant", n),
        "lognormal": np.exp(rng.normal(0, 4, n)).astype(np.float32),
        "small_ints": rng.integers(0, 5, n).astype(np.float32),
        "sparse": np.where(rng.random(n) < 0.05, rng.random(n) * 1e6, 0).astype(np.float32),
        "growing": (np.arange(n) % 4096).astype(np.float32) ** 2,
        "tiny_then_big": np.where(np.arange(n) < n // 2, 1e-30, 3.0).astype(np.float32),
        "negative": (rng.random(n) * 8 - 4).astype(np.float32),
    }
    # (no NaN field: a NaN sum becomes the mask, and x86 and the GPU propagate NaN payloads differently -- outside the format)
    for name, x in fields.items():
        for sub in (True, False):
            comp = api.compress(torch.from_numpy(x).to(dev), ct=7, abs_error_bound=0.01, chunk_elems=0, subtract_min=sub)
            torch.cuda.synchronize()
            enc = ol.o_compress_chunked(7, x, n, 0.01, subtract_min=sub)
            got, want = comp.chunk_medium[:1].cpu().numpy(), enc["chunk_medium"]
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)) or (np.isnan(got[0]) and np.isnan(want[0])), (name, sub, got, want)
            assert int(comp.chunk_type[0].item()) == int(enc["chunk_type"][0]), (name, sub)
            assert np.array_equal(comp.stream[:comp.total_bytes()].cpu().numpy(), enc["stream"]), (name, sub)
