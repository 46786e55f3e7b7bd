"""CPU only: the byte-wise codec CT=1 restatement (oracle/dc_oracle_bytewise.c) pinned against the UNMODIFIED reference
(oracle/_ref: myCompress impl/dataCompression.c:3977, myDecompress :3942)."""
import numpy as np
import pytest

import oracle_lib as ol
from data_compression_b200.fields import make_field


def fields():
    rng = np.random.default_rng(4)
    yield "smooth", make_field("smooth", 5000)
    yield "noisy", make_field("noisy", 3000)
    yield "constant", make_field("constant", 777)
    yield "ramp", (np.arange(2000, dtype=np.float32) * 0.37 - 100).astype(np.float32)
    yield "cubic", ((np.arange(1500) * 1e-2) ** 3).astype(np.float32)
    x = make_field("smooth", 3000).copy()
    x[[0, 2, 50, 51, 52, 53, 900, 2999]] = -1.0              # the empty-slot marker inside the data
    yield "sentinel", x
    x = rng.normal(0, 1, 1000).astype(np.float32)
    x[[10, 500]] = np.nan
    x[700] = np.inf
    yield "nonfinite", x
    yield "tiny", np.array([1.5, 2.5], np.float32)
    yield "five", np.array([1, 2, 3, 4, 5], np.float32)


@pytest.mark.parametrize("eb", (0.1, 0.01, 1e-4))
def test_matches_reference(eb):
    if not ol.build_ref(eb):
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    for name, x in fields():
        of, oc, od = ol.o_bytewise_compress(x, eb)
        rf, rc, rd = ol.ref_bytewise_compress(x, eb)
        assert np.array_equal(of.view(np.uint32), rf.view(np.uint32)), (name, eb)
        assert np.array_equal(oc, rc) and np.array_equal(od, rd), (name, eb)
        dec = ol.o_bytewise_decompress(of, oc, od, x.size)
        ref = ol.ref_bytewise_decompress(rf, rc, rd, x.size, eb)
        assert np.array_equal(dec.view(np.uint32), ref.view(np.uint32)), (name, eb)
        if name in ("smooth", "ramp", "constant"):
            assert oc.size > x.size // 2                       # mostly hits


def test_layout_invariants():
    x = make_field("smooth", 4096)
    f, c, d = ol.o_bytewise_compress(x, 0.01)
    assert f.size + c.size == x.size and np.all(np.diff(d) > 0) and d.min() >= 5 and d.max() <= x.size
    assert set(np.unique(c)) <= set(b"abcd")


@pytest.mark.parametrize("eb", (0.01, 1e-4))
def test_float64_matches_reference(eb):
    if not ol.build_ref(eb):
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    for name, x in fields():
        xd = x.astype(np.float64) * 1.000001
        if name == "sentinel":
            xd[[0, 2, 50, 51, 52, 53, 900, 2999]] = -1.0
        of, oc, od = ol.o_bytewise_compress_f64(xd, eb)
        rf, rc, rd = ol.ref_bytewise_compress_f64(xd, eb)
        assert np.array_equal(of.view(np.uint64), rf.view(np.uint64)), (name, eb)
        assert np.array_equal(oc, rc) and np.array_equal(od, rd), (name, eb)
        dec = ol.o_bytewise_decompress_f64(of, oc, od, xd.size)
        ref = ol.ref_bytewise_decompress_f64(rf, rc, rd, xd.size, eb)
        assert np.array_equal(dec.view(np.uint64), ref.view(np.uint64)), (name, eb)
