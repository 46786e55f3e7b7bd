"""CT=10: Hamming SECDED over blocks of the finished stream, on the device, against the CPU oracle
(oracle/dc_oracle_hamming.c, pinned to the compiled reference by tests/test_oracle_hamming.py): check bits of every block,
error classification, rectification in place -- through the device API and through the reference's own symbols called as
impl/pingpong.c:212-222, :296-311, :487-505 call them."""
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


def unpack(check, r):
    return bytes(ord("0") + ((int(check) >> i) & 1) for i in range(r + 1))


def pack(c):
    return sum((1 << i) for i, ch in enumerate(c) if ch == ord("1"))


@pytest.mark.parametrize("total,bs", ((1, 1), (5, 5), (1000, 64), (4099, 1000), (300001, 125000), (125000 * 3, 125000)))
def test_blocks_match_oracle(total, bs, dev):
    rng = np.random.default_rng(total + bs)
    host = rng.integers(0, 256, total + 3).astype(np.uint8)
    for shift in (0, 1):                                     # a stream that does not start on a 4-byte boundary, too
        if shift and total < 8:
            continue
        buf = torch.from_numpy(host).to(dev)
        s = buf[shift:shift + total]
        chk, r = api.hamming_encode(s, total, bs)
        torch.cuda.synchronize()
        nb = (total + bs - 1) // bs
        for b in range(nb):
            blk = host[shift + b * bs: shift + min(total, (b + 1) * bs)]
            orr, oc = ol.o_hamming_encode(blk)
            assert int(r[b]) == orr == api.hm_length(8 * blk.size)
            assert unpack(int(chk[b]) & 0xFFFFFFFF, orr) == oc, (total, bs, shift, b)


def test_decode_classifies_and_rectifies(dev):
    rng = np.random.default_rng(3)
    total, bs = 5000, 1000
    host = rng.integers(0, 256, total).astype(np.uint8)
    s = torch.from_numpy(host).to(dev)
    chk, r = api.hamming_encode(s, total, bs)
    r0 = int(r[0])
    clean = chk.clone()
    # block 0: clean; 1: one data bit; 2: two data bits; 3: the parity bit; 4: one check bit
    bad = host.copy()
    bad[1000 + 17] ^= 1 << 3
    bad[2000 + 5] ^= 1 << 7
    bad[2000 + 900] ^= 1
    sb = torch.from_numpy(bad).to(dev)
    cb = chk.clone()
    cb[3] ^= 1 << r0
    cb[4] ^= 1 << 2
    et = api.hamming_decode(sb, total, bs, cb)
    torch.cuda.synchronize()
    assert et.cpu().tolist() == [0, 3, 1, 2, 3]
    got = sb.cpu().numpy()
    want = bad.copy()
    for b in range(5):
        blk = bad[b * bs:(b + 1) * bs]
        c_in = unpack(int(chk[b]) & 0xFFFFFFFF, r0)
        if b == 3:
            c_in = unpack((int(chk[b]) ^ (1 << r0)) & 0xFFFFFFFF, r0)
        if b == 4:
            c_in = unpack((int(chk[b]) ^ 4) & 0xFFFFFFFF, r0)
        t, fixed, c_out = ol.o_hamming_decode(blk, c_in, r0)
        assert t == int(et[b])
        want[b * bs:(b + 1) * bs] = fixed
        assert unpack(int(cb[b]) & 0xFFFFFFFF, r0) == c_out, b
    assert np.array_equal(got, want)
    assert np.array_equal(got[:2000], host[:2000]) and np.array_equal(got[3000:], host[3000:])   # single errors are gone
    assert torch.equal(cb[[0, 1, 3, 4]], clean[[0, 1, 3, 4]])


def test_pingpong_ct10_call_sequence(dev, capfd):
    """impl/pingpong.c CT=10: compress (CT=5), block_size, hamming_encode per block; a bit flips on the wire; hamming_decode per
    block repairs it; the repaired stream decodes to the same floats."""
    L = api.lib()
    u8p = C.POINTER(C.c_uint8)
    L.hamming_encode.argtypes, L.hamming_encode.restype = [C.c_void_p, C.POINTER(C.c_char_p), C.c_int, C.POINTER(C.c_int)], None
    L.hamming_decode.argtypes, L.hamming_decode.restype = [C.c_void_p, C.c_char_p, C.c_int, C.c_int], C.c_int
    L.block_size.argtypes, L.block_size.restype = [C.c_int], C.c_int
    L.hmLength.argtypes, L.hmLength.restype = [C.c_int], C.c_int
    L.dc_set_ber.argtypes = [C.c_double]
    L.dc_set_ber(1e-4)                                       # blocks of 1250 bytes, so that the small stream has several
    x = make_field("noisy", 4000)
    enc = api.compress_host(x, ct=5, abs_error_bound=0.01, chunk_elems=0)
    stream = np.array(enc["stream"][:enc["total_bytes"]], dtype=np.uint8, copy=True)
    nbytes = stream.size
    bs = L.block_size(nbytes)
    assert bs == 1250
    nb = (nbytes + bs - 1) // bs
    cs, rs = [], []
    for i in range(nb):
        blk = stream[i * bs:min(nbytes, (i + 1) * bs)]
        cp, r = C.c_char_p(), C.c_int()
        L.hamming_encode(blk.ctypes.data, C.byref(cp), blk.size, C.byref(r))
        c = C.string_at(cp, r.value + 1)
        assert (r.value, c) == ol.o_hamming_encode(blk) and r.value == L.hmLength(8 * blk.size)
        cs.append(c)
        rs.append(r.value)
    wire = stream.copy()
    wire[bs + 77] ^= 1 << 2                                  # bit_flip on block 1
    types = []
    for i in range(nb):
        blk = wire[i * bs:min(nbytes, (i + 1) * bs)]        # a view: rectified in place
        cb = C.create_string_buffer(cs[i], len(cs[i]) + 1)
        types.append(L.hamming_decode(blk.ctypes.data, cb, blk.size, rs[i]))
    assert types == [0, 3] + [0] * (nb - 2)
    assert np.array_equal(wire, stream)
    out = capfd.readouterr().out
    assert "one bit error: pos = " in out
    want_pos = ol.o_hamming_decode(np.where(np.arange(bs) == 77, stream[bs:2 * bs] ^ 4, stream[bs:2 * bs]).astype(np.uint8), cs[1], rs[1])
    assert want_pos[0] == 3
    L.dc_set_ber(1e-6)
    assert L.block_size(10 ** 7) == 125000 and L.block_size(1000) == 1000
