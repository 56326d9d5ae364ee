"""Self-consistency of the GP emulator oracle (oracle/gp.py), in extended precision.

``gp_emulator`` is not vendored by the reference and no golden vector exists for it
(oracle/gp.py header: parity unpinned), so the conventions the oracle restates -- sign of
``X - x`` in the Jacobian, ``theta`` as LOG hyper-parameters, where the nugget sits -- are
checked here against things that do not depend on the oracle's own derivative code:

* the defining sums of mean / variance / Jacobian / Hessian evaluated term by term in
  ``np.longdouble`` (80-bit on x86-64) from the stored fields ``inputs, theta, invQt, invQ``;
* central finite differences, in longdouble, of the MEAN (-> Jacobian) and of the
  JACOBIAN (-> Hessian): a wrong sign, a missing ``theta`` factor or a transposed index
  shows up at order 1, the tolerance is 1e-7 relative;
* the interpolation property of a GP with a small nugget: the mean at the training inputs
  reproduces the targets, which pins ``invQt = (K + s_n I)^-1 t`` with K built from
  ``exp(theta)``.

The call sites these conventions serve: kafka/inference/utils.py:86-90 (predict) and
kafka/inference/kf_tools.py:26-34 (hessian).
"""
import numpy as np
import pytest

from oracle.gp import GaussianProcess, make_synthetic_gp

LD = np.longdouble
pytestmark = pytest.mark.skipif(np.finfo(LD).eps > 1e-18, reason="no extended precision on this platform")

SAIL_MEAN = np.array([2.1, np.exp(-0.6), np.exp(-0.07), 0.1, np.exp(-0.88), np.exp(-0.2), np.exp(-2.0), 70. / 90., 0.5, 0.9])
SAIL_SIGMA = np.array([0.01, 0.2, 0.01, 0.05, 0.01, 0.01, 0.50, 0.1, 0.1, 0.1])


def _mean_ld(gp, x):
    """mu(x) = sum_j s_f exp(-1/2 sum_d e_d (X_jd - x_d)^2) alpha_j, all in longdouble; x: (D,) longdouble."""
    e = np.exp(gp.theta.astype(LD))
    diff = gp.inputs.astype(LD) - x[None, :]
    k = e[gp.D] * np.exp(-LD(0.5) * (e[:gp.D] * diff * diff).sum(axis=1))
    return (k * gp.invQt.astype(LD)).sum(), k, diff, e


def _jac_ld(gp, x):
    mu, k, diff, e = _mean_ld(gp, x)
    w = k * gp.invQt.astype(LD)
    return e[:gp.D] * (w[:, None] * diff).sum(axis=0)


def _hess_ld(gp, x):
    mu, k, diff, e = _mean_ld(gp, x)
    w = k * gp.invQt.astype(LD)
    sd = e[:gp.D] * diff
    return np.einsum("j,jd,je->de", w, sd, sd) - mu * np.diag(e[:gp.D])


def _var_ld(gp, x):
    mu, k, diff, e = _mean_ld(gp, x)
    return e[gp.D] - k @ (gp.invQ.astype(LD) @ k)


def _cases():
    rng = np.random.default_rng(2024)
    out = []
    gp = make_synthetic_gp(rng, SAIL_MEAN, SAIL_SIGMA, n_train=250)              # the bench's emulator shape
    out.append(("prosail", gp, SAIL_MEAN + SAIL_SIGMA * rng.standard_normal((6, 10)), SAIL_SIGMA))
    mean4, sig4 = SAIL_MEAN[[0, 1, 6, 2]], SAIL_SIGMA[[0, 1, 6, 2]]
    for outlier in (45.0, 8000.0):            # tests/test_gpu_gp.py: underflowing point / "wide table" case
        g0 = make_synthetic_gp(rng, mean4, sig4, n_train=40)
        X = g0.inputs.copy()
        X[7] = mean4 + outlier * 4.0 * sig4 / 2.0
        g = GaussianProcess(X, g0.targets, g0.theta)
        pts = np.vstack([mean4 + 0.5 * sig4 * rng.standard_normal((3, 4)),
                         X[7] + 0.5 * sig4 * rng.standard_normal((3, 4))])
        out.append(("outlier%g" % outlier, g, pts, sig4))
    return out


CASES = _cases()


@pytest.mark.parametrize("name,gp,pts,sigma", CASES, ids=[c[0] for c in CASES])
def test_predict_and_hessian_equal_the_defining_sums_in_extended_precision(name, gp, pts, sigma):
    mu, var, jac = gp.predict(pts)
    mu2, jac2 = gp.predict(pts, do_unc=False)
    assert np.array_equal(mu, mu2) and np.array_equal(jac, jac2)
    hess = gp.hessian(pts)
    s_f = np.exp(gp.theta[gp.D])
    for i, x in enumerate(pts):
        xl = x.astype(LD)
        m_ld = _mean_ld(gp, xl)[0]
        j_ld, h_ld, v_ld = _jac_ld(gp, xl), _hess_ld(gp, xl), _var_ld(gp, xl)
        # alpha alternates in sign with |alpha| up to 1e5 x the mean: the fp64 sums carry ~1e-16 x sum|terms|
        e = np.exp(gp.theta.astype(LD))
        k = _mean_ld(gp, xl)[1]
        cancel = float((np.abs(k * gp.invQt)).sum())
        assert abs(mu[i] - float(m_ld)) <= 1e-13 * cancel + 1e-300
        # the oracle forms the Jacobian as e_d (W^T Xc - (x - c) mu) on inputs centred at c: its rounding
        # level is set by the two terms it subtracts, not by their difference
        c = gp.inputs.mean(axis=0)
        wabs = np.abs(np.asarray(k, dtype=float) * gp.invQt)
        scale_j = float((np.asarray(e[:gp.D], dtype=float) * ((wabs[:, None] * np.abs(gp.inputs - c)).sum(axis=0)
                                                               + np.abs(x - c) * wabs.sum())).max())
        assert np.max(np.abs(jac[i] - j_ld.astype(float))) <= 1e-12 * scale_j + 1e-300
        assert np.max(np.abs(jac[i] - j_ld.astype(float)) / np.maximum(np.abs(jac[i]), np.abs(jac[i]).max() * 1e-3)) < 1e-9
        scale_h = max(np.max(np.abs(h_ld.astype(float))), 1e-300)
        assert np.max(np.abs(hess[i] - h_ld.astype(float))) <= 1e-9 * scale_h + 1e-12 * cancel * float(e[:gp.D].max())
        # s_f - k^T Q^-1 k cancels to ~1e-6 s_f near the training set: compare on the scale of s_f
        assert abs(var[i] - float(v_ld)) <= 1e-8 * s_f


@pytest.mark.parametrize("name,gp,pts,sigma", CASES, ids=[c[0] for c in CASES])
def test_jacobian_and_hessian_are_the_derivatives_of_the_mean(name, gp, pts, sigma):
    """Central differences in longdouble with step 1e-5 length scales: truncation ~1e-10,
    rounding ~1e-14 of the function scale."""
    _, jac = gp.predict(pts, do_unc=False)
    hess = gp.hessian(pts)
    ls = 1.0 / np.sqrt(np.exp(gp.theta[:gp.D]))
    for i, x in enumerate(pts):
        xl = x.astype(LD)
        fd_j = np.empty(gp.D)
        fd_h = np.empty((gp.D, gp.D))
        for d in range(gp.D):
            h = LD(1e-5) * LD(ls[d])
            step = np.zeros(gp.D, dtype=LD)
            step[d] = h
            fd_j[d] = float((_mean_ld(gp, xl + step)[0] - _mean_ld(gp, xl - step)[0]) / (2 * h))
            fd_h[:, d] = ((_jac_ld(gp, xl + step) - _jac_ld(gp, xl - step)) / (2 * h)).astype(float)
        # scale of a derivative: mean scale per length scale
        j_scale = np.maximum(np.abs(fd_j), np.max(np.abs(fd_j * ls)) / ls * 1e-3)
        assert np.max(np.abs(jac[i] - fd_j) / j_scale) < 1e-7, (name, i)
        h_scale = np.maximum(np.abs(fd_h), np.max(np.abs(fd_h * np.outer(ls, ls))) / np.outer(ls, ls) * 1e-3)
        assert np.max(np.abs(hess[i] - fd_h) / h_scale) < 1e-7, (name, i)
        assert np.max(np.abs(fd_h - fd_h.T) / h_scale) < 1e-7                  # and it is symmetric


def test_mean_interpolates_the_training_targets():
    """With nugget s_n << s_f the posterior mean at the training inputs is t - s_n alpha:
    pins invQt = (K + s_n I)^-1 t with K = s_f exp(-1/2 sum exp(theta_d) dx^2)."""
    rng = np.random.default_rng(8)
    gp = make_synthetic_gp(rng, SAIL_MEAN, SAIL_SIGMA, n_train=120)
    mu = gp.predict(gp.inputs, do_deriv=False, do_unc=False)
    s_n = np.exp(gp.theta[gp.D + 1])
    resid = gp.targets - mu
    assert np.max(np.abs(resid - s_n * gp.invQt)) < 1e-9 * np.abs(gp.targets).max()
    assert np.max(np.abs(resid)) < 1e-2 * np.abs(gp.targets).max()
    # the variance there is ~ s_n (<< s_f), and far away it returns to s_f
    var_train = gp.predict(gp.inputs[:5], do_deriv=False)[1]
    assert np.all(var_train < 100 * s_n) and np.all(var_train > -1e-9)
    far = gp.predict((SAIL_MEAN + 1e3 * SAIL_SIGMA)[None], do_deriv=False)[1]
    assert abs(far[0] - np.exp(gp.theta[gp.D])) < 1e-12
