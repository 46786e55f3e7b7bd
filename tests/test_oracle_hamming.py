"""CPU only: the Hamming SECDED restatement (oracle/dc_oracle_hamming.c) pinned against the UNMODIFIED reference
(oracle/_ref, impl/dataCompression.c:5544-5878) -- check bits, error classification and rectification, for the block
shapes impl/pingpong.c:296-311 produces (block = min(bytes, 1/(8 BER)), BER = 1e-6 -> 125000 bytes)."""
import numpy as np
import pytest

import oracle_lib as ol

needs_ref = pytest.mark.skipif(not ol.build_ref(0.01), reason="oracle/_ref not built (no /root/reference here)")


def test_hm_length_known_values():
    # 2^r - 1 - r - k >= 0 (:5587-5599)
    assert [ol.o_hm_length(k) for k in (1, 4, 8, 11, 12, 26, 57, 1000000)] == [2, 3, 4, 4, 5, 5, 6, 20]


@needs_ref
@pytest.mark.parametrize("nbytes", (1, 2, 3, 7, 64, 1000, 4097))
def test_encode_matches_reference(nbytes):
    rng = np.random.default_rng(nbytes)
    for fill in ("random", "zeros", "ones"):
        bits = {"random": rng.integers(0, 256, nbytes), "zeros": np.zeros(nbytes), "ones": np.full(nbytes, 255)}[fill].astype(np.uint8)
        assert ol.o_hamming_encode(bits) == ol.ref_hamming_encode(bits), (nbytes, fill)


@needs_ref
@pytest.mark.parametrize("nbytes", (1, 5, 300, 2048))
def test_decode_matches_reference_under_errors(nbytes, capfd):
    rng = np.random.default_rng(100 + nbytes)
    bits = rng.integers(0, 256, nbytes).astype(np.uint8)
    r, c = ol.o_hamming_encode(bits)
    k = 8 * nbytes
    cases = [("none", [], [])]
    cases += [("data1", [int(d)], []) for d in rng.integers(0, k, 6)] + [("data1", [0], []), ("data1", [k - 1], [])]
    cases += [("check1", [], [int(i)]) for i in range(r)] + [("parity", [], [r])]
    cases += [("data2", [int(a), int(b)], []) for a, b in zip(rng.integers(0, k, 4), rng.integers(0, k, 4)) if a != b]
    cases += [("data+check", [int(rng.integers(0, k))], [int(rng.integers(0, r))])]
    for name, dflip, cflip in cases:
        b2 = bits.copy()
        for d in dflip:
            b2[d // 8] ^= 1 << (7 - d % 8)
        c2 = bytearray(c)
        for i in cflip:
            c2[i] = ord("1") if c2[i] == ord("0") else ord("0")
        got = ol.o_hamming_decode(b2, bytes(c2), r)
        want = ol.ref_hamming_decode(b2, bytes(c2), r)
        assert got[0] == want[0] and np.array_equal(got[1], want[1]) and got[2] == want[2], (nbytes, name, dflip, cflip)
        if name in ("data1", "check1", "parity"):           # single errors are repaired
            assert np.array_equal(got[1], bits) and got[2] == c
    capfd.readouterr()                                      # the reference prints its diagnosis


@needs_ref
def test_full_block_125000_bytes():
    rng = np.random.default_rng(7)
    bits = rng.integers(0, 256, 125000).astype(np.uint8)
    r, c = ol.o_hamming_encode(bits)
    assert r == 20
    assert (r, c) == ol.ref_hamming_encode(bits)
