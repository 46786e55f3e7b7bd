"""Himeno Jacobi sweep on the device against a float32 numpy restatement of impl/himenoBMTxps.c:256-282
(every operation rounded on its own, in the C expression's order): wrk2 and the new p bit exact, gosa to 1e-6 relative
(the device sums ss^2 in double, block by block; the reference sums in float in loop order -- a diagnostic)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def jacobi_numpy(p, a, b, c, bnd, wrk1, imax, jmax, kmax, omega):
    f = np.float32
    I, J, K = slice(1, imax - 1), slice(1, jmax - 1), slice(1, kmax - 1)

    def sh(di, dj, dk):
        return p[1 + di:imax - 1 + di, 1 + dj:jmax - 1 + dj, 1 + dk:kmax - 1 + dk]

    s0 = a[0][I, J, K] * sh(1, 0, 0)
    s0 = s0 + a[1][I, J, K] * sh(0, 1, 0)
    s0 = s0 + a[2][I, J, K] * sh(0, 0, 1)
    s0 = s0 + b[0][I, J, K] * (((sh(1, 1, 0) - sh(1, -1, 0)) - sh(-1, 1, 0)) + sh(-1, -1, 0))
    s0 = s0 + b[1][I, J, K] * (((sh(0, 1, 1) - sh(0, -1, 1)) - sh(0, 1, -1)) + sh(0, -1, -1))
    s0 = s0 + b[2][I, J, K] * (((sh(1, 0, 1) - sh(-1, 0, 1)) - sh(1, 0, -1)) + sh(-1, 0, -1))
    s0 = s0 + c[0][I, J, K] * sh(-1, 0, 0)
    s0 = s0 + c[1][I, J, K] * sh(0, -1, 0)
    s0 = s0 + c[2][I, J, K] * sh(0, 0, -1)
    s0 = s0 + wrk1[I, J, K]
    ss = (s0 * a[3][I, J, K] - sh(0, 0, 0)) * bnd[I, J, K]
    assert ss.dtype == f
    w = sh(0, 0, 0) + f(omega) * ss
    return w, float((ss * ss).astype(np.float64).sum())


@pytest.mark.parametrize("dims", [(17, 17, 33), (33, 20, 19), (65, 65, 129)])
@pytest.mark.parametrize("coef", ["initmt", "random"])
def test_jacobi_bit_exact(dims, coef):
    import torch
    from data_compression_b200 import api
    imax, jmax, kmax = dims
    MI, MJ, MK = imax + 2, jmax + 1, kmax + 3            # MIMAX > imax as in the driver (padded allocation)
    rng = np.random.default_rng(7)
    f = np.float32
    p = np.zeros((MI, MJ, MK), f)
    if coef == "initmt":                                  # impl/himenoBMTxps.c:203-246
        a = np.zeros((4, MI, MJ, MK), f); b = np.zeros((3, MI, MJ, MK), f); c = np.zeros((3, MI, MJ, MK), f)
        bnd = np.zeros((MI, MJ, MK), f); wrk1 = np.zeros((MI, MJ, MK), f)
        a[:3, :imax, :jmax, :kmax] = 1.0; a[3, :imax, :jmax, :kmax] = f(1.0 / 6.0)
        c[:, :imax, :jmax, :kmax] = 1.0; bnd[:imax, :jmax, :kmax] = 1.0
        ii = np.arange(imax, dtype=f)
        p[:imax, :jmax, :kmax] = ((ii * ii) / f((imax - 1) * (imax - 1)))[:, None, None]
    else:
        a = rng.standard_normal((4, MI, MJ, MK)).astype(f); b = rng.standard_normal((3, MI, MJ, MK)).astype(f)
        c = rng.standard_normal((3, MI, MJ, MK)).astype(f); bnd = rng.integers(0, 2, (MI, MJ, MK)).astype(f)
        wrk1 = rng.standard_normal((MI, MJ, MK)).astype(f); p = rng.standard_normal((MI, MJ, MK)).astype(f)
    omega = 0.8
    dev = torch.device("cuda:0")
    t = lambda x: torch.from_numpy(x.copy()).to(dev)
    dp, da, db, dc, dbnd, dw1 = t(p), t(a), t(b), t(c), t(bnd), t(wrk1)
    dw2 = torch.full((MI, MJ, MK), -7.0, dtype=torch.float32, device=dev)
    want_p = p.copy()
    gosa_ref = None
    for sweep in range(3):
        w, gosa_ref = jacobi_numpy(want_p, a, b, c, bnd, wrk1, imax, jmax, kmax, omega)
        want_p[1:imax - 1, 1:jmax - 1, 1:kmax - 1] = w
        g = api.himeno_jacobi(dp, da, db, dc, dbnd, dw1, dw2, imax, jmax, kmax, omega)
        got_w = dw2.cpu().numpy()
        assert np.array_equal(got_w[1:imax - 1, 1:jmax - 1, 1:kmax - 1].view(np.uint32), w.view(np.uint32)), sweep
        assert np.array_equal(dp.cpu().numpy().view(np.uint32), want_p.view(np.uint32)), sweep
        assert abs(float(g.item()) - gosa_ref) <= 1e-6 * max(gosa_ref, 1e-30)
    # nothing outside the interior of wrk2 was written
    got_w = dw2.cpu().numpy()
    mask = np.ones((MI, MJ, MK), bool); mask[1:imax - 1, 1:jmax - 1, 1:kmax - 1] = False
    assert (got_w[mask] == -7.0).all()


def test_jacobi_rejects_bad_dims():
    import torch
    from data_compression_b200 import api
    dev = torch.device("cuda:0")
    z = torch.zeros((4, 4, 4), dtype=torch.float32, device=dev)
    z4 = torch.zeros((4, 4, 4, 4), dtype=torch.float32, device=dev)
    with pytest.raises(api.DcbError):
        api.himeno_jacobi(z, z4, z4[:3], z4[:3], z, z, z.clone(), 5, 4, 4)
    with pytest.raises(api.DcbError):
        api.himeno_jacobi(z, z4, z4[:3], z4[:3], z, z, z.clone(), 2, 4, 4)
