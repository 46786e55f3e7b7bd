"""The float64 intra-chunk parallel decoder (csrc/dcb_decode_single_f64.cuh): ONE chunk = ONE reference call on the whole array
(myDecompress_bitwise_double{,_np,_op,_mask}, dataCompression.c:2656 / :2286 / :476 / :1199, as impl/k-means.c:520-590,
impl/mm.c:404-690 and impl/lu.c:199-345 call them).  Decoded doubles must equal the float64 oracle's bit for bit, and the
one-thread kernel's (DCB_DECODER=lane, checked in a subprocess)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle_lib as ol
from data_compression_b200 import api

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.fail("no CUDA device")
    return torch.device("cuda:0")


def fields(rng, n):
    i = np.arange(n)
    yield "sine", 5.0 + 4.5 * np.sin(2 * np.pi * (i % 65536) / 65536.0)
    yield "walk", np.cumsum(rng.standard_normal(n)) * 0.02 + 20.0
    yield "uniform", rng.random(n) * 10.0
    yield "constant", np.full(n, 0.123456789)
    mixed = rng.random(n) * 10.0
    for lo in range(0, n - 3000, 9000):
        mixed[lo:lo + 3000] = 3.0 + np.sin(np.arange(3000) / 50.0)
    mixed[n // 2: n // 2 + 1500] = 0.25
    yield "mixed", mixed
    yield "big", rng.random(n) * 1e200


@pytest.mark.parametrize("ct", (5, 6, 7, 11))
def test_fields_single_chunk(ct, dev):
    rng = np.random.default_rng(170 + ct)
    for n in (4097, 33929, 131072, 300001):
        for name, x in fields(rng, n):
            for eb in (0.01, 1e-6):
                for sub in ((True, False) if name == "uniform" else (True,)):   # (without the minimum only non-negative data decodes)
                    comp = api.compress_f64(torch.from_numpy(x).to(dev), ct=ct, abs_error_bound=eb, chunk_elems=0, subtract_min=sub)
                    got = api.decompress_f64(comp).cpu().numpy()
                    enc = ol.o_compress_chunked_f64(ct, x, n, eb, subtract_min=sub)
                    want = ol.o_decompress_chunked_f64(ct, enc, n, n, eb, add_min=sub)
                    assert np.array_equal(got.view(np.uint64), want.view(np.uint64)), (ct, name, n, eb, sub)


def gpu_decode_stream(stream, n, ct, eb, dev, fixed=None, cmin=None):
    """Decode a raw byte stream as ONE chunk through dcb_decompress_f64."""
    L = api.lib()
    s = torch.zeros(stream.size + 64, dtype=torch.uint8, device=dev)
    s[:stream.size] = torch.from_numpy(stream).to(dev)
    off = torch.tensor([0, stream.size], dtype=torch.int64, device=dev)
    out = torch.full((n,), 12345.0, dtype=torch.float64, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    flags, ft, fm = 0, 0, 0
    if fixed is not None:
        flags |= api.F_FIXED_MASK
        ft, fm = fixed
    mn = None
    if cmin is not None:
        flags |= api.F_SUBTRACT_MIN
        mn = torch.tensor([cmin], dtype=torch.float64, device=dev)
    rc = L.dcb_decompress_f64(ct, C.c_void_p(s.data_ptr()), C.c_void_p(off.data_ptr()), n, 0, eb, flags, ft, fm,
                              C.c_void_p(mn.data_ptr()) if mn is not None else None, None, None, C.c_void_p(out.data_ptr()),
                              C.c_void_p(status.data_ptr()), C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    assert rc == 0, rc
    torch.cuda.synchronize()
    return out.cpu().numpy(), int(status.item())


def oracle_decode_stream(stream, n, ct, eb, fixed=None, cmin=None):
    enc = dict(stream=stream, offsets=np.array([0, stream.size], dtype=np.uint64),
               chunk_min=np.array([cmin if cmin is not None else 0.0], dtype=np.float64),
               chunk_mask20=np.zeros(1, np.uint32), chunk_type=np.zeros(1, np.int32))
    ft, fm = fixed if fixed is not None else (-1, 0)
    return ol.o_decompress_chunked_f64(ct, enc, n, n, eb, add_min=cmin is not None, fixed_type=ft, fixed_mask20=fm)


@pytest.mark.parametrize("ct", (5, 7, 11))
def test_random_bit_strings_decode_like_the_oracle(ct, dev):
    """Any bit string is a stream of CT=5/7/11 codes in some order: raw codes of every length, 3-bit codes on every history
    state, escapes.  (CT=6 rejects a leading 1; it is covered by the field tests.)"""
    rng = np.random.default_rng(230 + ct)
    for trial in range(6):
        nbytes = int(rng.integers(60000, 160000))
        stream = rng.integers(0, 256, nbytes, dtype=np.uint8)
        if trial % 2:                                   # lots of 3-bit codes: set the top bit of many bytes
            stream |= (rng.random(nbytes) < 0.6).astype(np.uint8) << 7
        n = nbytes * 8 // 64 - 4                         # every code has at most 64 bits: n codes certainly fit
        eb = (0.01, 1e-6, 0.1)[trial % 3]
        fixed = (trial % 11, int(rng.integers(0, 1 << 20))) if ct == 7 else None
        got, st = gpu_decode_stream(stream, n, ct, eb, dev, fixed=fixed, cmin=0.5 if trial == 3 else None)
        ref = oracle_decode_stream(stream, n, ct, eb, fixed=fixed, cmin=0.5 if trial == 3 else None)
        # NaN payloads are outside the comparison (x86 and the GPU propagate them differently through the predictors; NaN is
        # outside the format's domain) -- the PARSE after them must still agree
        differ = (got.view(np.uint64) != ref.view(np.uint64)) & ~(np.isnan(got) & np.isnan(ref))
        ne = np.nonzero(differ)[0]
        assert ne.size == 0, (ct, trial, ne[:5], got[ne[:5]], ref[ne[:5]])


def _bits_to_bytes(bits):
    bits = bits + "0" * (-len(bits) % 8)
    return np.frombuffer(int(bits, 2).to_bytes(len(bits) // 8, "big"), dtype=np.uint8).copy()


def test_marker_values_inside_a_long_chunk(dev):
    """-1.0 is the reference's 'empty history slot' marker: a predicted value that comes out as exactly -1.0 takes the fill
    branch of the history update instead of the shift.  CT=11 stream built by hand: 1.0 (full word), 0 (100), 2*0 - 1 = -1 (110),
    then predicted codes on that state; repeated between stretches of other codes, also inside groups of 32 predicted codes."""
    rng = np.random.default_rng(31)
    w = lambda f: format(int(np.float64(f).view(np.uint64)), "064b")
    pieces, n = [], 0
    for rep in range(700):
        pad = int(rng.integers(0, 40))
        seq = [w(1.0), "100", "110"] + ["101", "110", "111", "110", "111", "101", "111", "110", "111"][: 4 + rep % 6]
        seq += ["111"] * pad + [w(rng.random() * 4), w(rng.random() * 4), w(rng.random() * 4)] + ["110", "111"] * (rep % 5)
        pieces += seq
        n += len(seq)
    stream = _bits_to_bytes("".join(pieces))
    assert n > 4096
    got, _ = gpu_decode_stream(stream, n, 11, 0.01, dev)
    ref = oracle_decode_stream(stream, n, 11, 0.01)
    assert (ref == -1.0).sum() >= 700
    differ = (got.view(np.uint64) != ref.view(np.uint64)) & ~(np.isnan(got) & np.isnan(ref))
    assert not differ.any(), np.nonzero(differ)[0][:5]


def test_same_result_as_one_thread_kernel(dev):
    """DCB_DECODER=lane forces the thread-per-chunk kernel (the env var is read once per process, hence the subprocess)."""
    code = r'''
import sys, numpy as np, torch
sys.path.insert(0, %r)
from data_compression_b200 import api
rng = np.random.default_rng(3)
x = np.concatenate([rng.random(40000) * 10, 5 + 4.5 * np.sin(np.arange(60000) / 700.0)])
xd = torch.from_numpy(x).cuda()
for ct in (5, 6, 7, 11):
    comp = api.compress_f64(xd, ct=ct, abs_error_bound=0.001, chunk_elems=0)
    print(ct, int(api.decompress_f64(comp).view(torch.int64).remainder(1000003).sum().item()))
''' % (ROOT,)
    outs = []
    for mode in ("lane", "parallel"):
        env = dict(os.environ, DCB_DECODER=mode)
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(r.stdout)
    assert outs[0] == outs[1] and len(outs[0].split()) == 8


def test_np_rejects_leading_one_in_a_long_chunk(dev):
    """CT=6 on a stream whose code starts with 1: status DCB_EFORMAT, as the one-thread kernel reports it (:2330-2334)."""
    x = np.random.default_rng(4).random(10000) * 10.0
    comp = api.compress_f64(torch.from_numpy(x).to(dev), ct=6, abs_error_bound=0.01, chunk_elems=0, subtract_min=False)
    total = comp.total_bytes()
    bad = comp.stream[:total].cpu().numpy().copy()
    bad[0] |= 0x80
    _, st = gpu_decode_stream(bad, x.size, 6, 0.01, dev)
    assert st == -5
