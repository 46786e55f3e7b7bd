"""The GPU path against the COMPILED REFERENCE directly (oracle/_ref/libdcref_<eb>.so = the unmodified
impl/dataCompression.c, which travels to the GPU box), without the restatement in between: the three bench fields x CT 5/6/7/11
at C = 4096, streams + chunk tables + CRC bit for bit, and the decoded floats against the reference's own decoder (CT=7: where
the reference decoder is defined).  One driver record of `pytest -m gpu` thereby shows reference parity."""
import numpy as np
import pytest

import oracle_lib as ol
from data_compression_b200 import api
from data_compression_b200.fields import make_field

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
C = 4096


def reference(eb):
    if not ol.ref_available(eb) and not ol.build_ref(eb):
        pytest.skip("oracle/_ref/libdcref_%s.so was not built" % ol.eb_str(eb))
    return ol.RefLib(eb)


@pytest.mark.parametrize("kind", ("smooth", "noisy", "constant"))
@pytest.mark.parametrize("ct", (5, 6, 7, 11))
def test_chunked_streams_and_decodes_equal_the_compiled_reference(ct, kind):
    dev = torch.device("cuda:0")
    n = 11 * C                                             # whole chunks (the reference's decoder corrupts its heap on some
                                                           # short padded streams; ragged chunks are tested against the restatement)
    x = make_field(kind, n, start=7 * C)
    for eb in (0.01, 1e-4):
        R = reference(eb)
        ref = R.compress_chunked(ct, x, C, want_crc=True)
        comp = api.compress(torch.from_numpy(x).to(dev), ct=ct, abs_error_bound=eb, chunk_elems=C, crc=True)
        K = len(ref["offsets"]) - 1
        offs = comp.chunk_offsets.cpu().numpy().astype(np.uint64)
        assert np.array_equal(offs, ref["offsets"]), (ct, kind, eb)
        assert np.array_equal(comp.stream[:int(offs[-1])].cpu().numpy(), ref["stream"]), (ct, kind, eb, "stream")
        assert np.array_equal(comp.chunk_min[:K].cpu().numpy().view(np.uint32), ref["chunk_min"].view(np.uint32))
        assert np.array_equal(comp.chunk_crc[:K].cpu().numpy().view(np.uint32), ref["chunk_crc"]), (ct, kind, eb, "crc")
        if ct == 7:
            assert np.array_equal(comp.chunk_type[:K].cpu().numpy(), ref["chunk_type"])
            assert np.array_equal(comp.chunk_medium[:K].cpu().numpy().view(np.uint32) >> 15, ref["chunk_medium"].view(np.uint32) >> 15)
        got = api.decompress(comp).cpu().numpy()
        for c in range(K):
            lo, hi = c * C, min(n, (c + 1) * C)
            t, m17 = (int(ref["chunk_type"][c]), ol.mask17_of(ref["chunk_medium"][c])) if ct == 7 else (0, 0)
            if ct == 7 and not ol.ref_mask_decode_defined(t, m17, eb):
                continue                                   # the reference decoder is undefined there (documented)
            s = ref["stream"][int(ref["offsets"][c]):int(ref["offsets"][c + 1])]
            want = R.decompress(ct, s, hi - lo, t, m17) + ref["chunk_min"][c]
            assert np.array_equal(got[lo:hi].view(np.uint32), want.astype(np.float32).view(np.uint32)), (ct, kind, eb, c)


@pytest.mark.parametrize("ct", (5, 6, 11))
def test_single_call_equals_the_compiled_reference(ct):
    """One reference call on the whole array (impl/pingpong.c:148-169, :553-566) -- the intra-chunk parallel decoder."""
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(41)
    n = 40000
    x = np.concatenate([make_field("smooth", 25000), (rng.random(n - 25000) * 10).astype(np.float32)])
    R = reference(0.01)
    mn, y = R.min_subtract(x)
    s, _ = R.compress(ct, y)
    comp = api.compress(torch.from_numpy(x).to(dev), ct=ct, abs_error_bound=0.01, chunk_elems=0)
    assert comp.total_bytes() == s.size and np.array_equal(comp.stream[:s.size].cpu().numpy(), s)
    want = (R.decompress(ct, s, n) + mn).astype(np.float32)
    assert np.array_equal(api.decompress(comp).cpu().numpy().view(np.uint32), want.view(np.uint32))
