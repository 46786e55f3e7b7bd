"""Himeno halo faces on the device (BASELINE.json configs[3]): gather / scatter against numpy indexing in the
reference's loop order (dataCompression.c:3741-3775, himenoBMTxps.c:560-568), and a face compressed as ONE reference call
(toSmallDataset_float + myCompress_bitwise*) bit-exact against the oracle."""
import numpy as np
import pytest

import oracle_lib as ol
from data_compression_b200 import api, halo

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.fail("no CUDA device: the product has no CPU fallback and these tests need a B200")
    return torch.device("cuda:0")


def himeno_like(MI, MJ, MK, seed=3):
    rng = np.random.default_rng(seed)
    i, j, k = np.meshgrid(np.arange(MI), np.arange(MJ), np.arange(MK), indexing="ij")
    p = (i * i / float((MI - 1) ** 2)) + 0.05 * np.sin(0.1 * j) * np.cos(0.07 * k) + 1e-3 * rng.standard_normal((MI, MJ, MK))
    return p.astype(np.float32)


def np_face(p, ijk, v, imax, jmax, kmax):
    if ijk == 1:
        return p[v, :jmax, :kmax].reshape(-1).copy()
    if ijk == 2:
        return p[:imax, v, :kmax].reshape(-1).copy()
    return p[:imax, :jmax, v].reshape(-1).copy()


def test_gather_scatter_match_reference_loops(dev):
    MI, MJ, MK = 19, 35, 67
    imax, jmax, kmax = 18, 33, 65                      # the drivers allocate MIMAX > imax (himenoBMTxps.c param.h)
    p = himeno_like(MI, MJ, MK)
    pd = torch.from_numpy(p).to(dev)
    for ijk, ext in ((1, imax), (2, jmax), (3, kmax)):
        for v in (0, 1, ext - 2, ext - 1):
            got = api.gather_face(pd, ijk, v, imax, jmax, kmax).cpu().numpy()
            assert np.array_equal(got.view(np.uint32), np_face(p, ijk, v, imax, jmax, kmax).view(np.uint32)), (ijk, v)
    vals = np.arange(imax * jmax, dtype=np.float32) * 0.5
    q = pd.clone()
    api.scatter_face(q, 3, kmax - 1, imax, jmax, kmax, torch.from_numpy(vals).to(dev), add=0.25)
    want = p.copy()
    want[:imax, :jmax, kmax - 1] = (vals + np.float32(0.25)).reshape(imax, jmax)
    assert np.array_equal(q.cpu().numpy().view(np.uint32), want.view(np.uint32))
    with pytest.raises(api.DcbError):
        api.gather_face(pd, 4, 0, imax, jmax, kmax)
    with pytest.raises(api.DcbError):
        api.gather_face(pd, 3, MK, imax, jmax, kmax)


@pytest.mark.parametrize("ct", (5, 6, 7))
def test_face_compressed_as_one_reference_call(ct, dev):
    MI, MJ, MK = 66, 66, 130
    imax, jmax, kmax = 65, 65, 129
    p = himeno_like(MI, MJ, MK, seed=7)
    pd = torch.from_numpy(p).to(dev)
    for ijk, v in ((3, 1), (3, kmax - 2), (1, 1), (2, jmax - 2)):
        face = np_face(p, ijk, v, imax, jmax, kmax)
        comp = halo.compress_face(pd, ijk, v, imax, jmax, kmax, ct, 0.001)
        enc = ol.o_compress_chunked(ct, face, face.size, 0.001)
        total = comp.total_bytes()
        assert np.array_equal(comp.stream[:total].cpu().numpy(), enc["stream"]), (ct, ijk, v)
        q = pd.clone()
        halo.write_face(q, ijk, 0, imax, jmax, kmax, comp)
        dec = ol.o_decompress_chunked(ct, enc, face.size, face.size, 0.001)
        assert np.array_equal(np_face(q.cpu().numpy(), ijk, 0, imax, jmax, kmax).view(np.uint32), dec.view(np.uint32)), (ct, ijk, v)
