"""Stand-in for ``gp_emulator.GaussianProcess`` (TEST INFRASTRUCTURE).

PARITY UNPINNED.  The reference never vendors its emulator: GP objects arrive by
unpickling files (``kafka/input_output/Sentinel2_Observations.py:158-159``,
``kafka/input_output/observations.py:286``) whose class lives in the third-party
package ``gp_emulator`` (github.com/jgomezdans/gp_emulator).  That package is
absent from ``/root/reference``, is listed in neither ``setup.py:5-6`` nor
``environment.yml`` (so there is NO pinned version), is not installed here and
the reference's tests hold no vector for it.  This module restates the
package's published algorithm (``GaussianProcess._set_params / predict /
hessian``) from the maths, keeps its attribute names (``inputs``, ``targets``,
``theta`` as LOG hyper-parameters, ``invQ``, ``invQt``, ``D``, ``n``) so that a
real un-pickled object exposes the same fields, and is anchored on the
reference's call sites:

* ``gp.predict(unique_vectors, do_unc=False)`` returning ``(mean, jacobian)`` or
  ``(mean, var, jacobian)``  -- ``kafka/inference/utils.py:86-90``
* ``gp.hessian(x)`` returning ``(n, D, D)``  -- ``kafka/inference/kf_tools.py:28``

Model (ARD squared exponential)::

    k(x, X_j) = s_f * exp(-0.5 * sum_d e_d (X_jd - x_d)^2)        e = exp(theta[:D])
    Q         = K(X, X) + s_n I                                   s_f = exp(theta[D])
    alpha     = Q^-1 t        ("invQt")                           s_n = exp(theta[D+1])
    mean(x)   = sum_j k_j alpha_j
    dmean/dx_d= e_d sum_j k_j alpha_j (X_jd - x_d)
    var(x)    = s_f - k^T Q^-1 k
    hess_de   = sum_j k_j alpha_j [ e_d e_e (X_jd-x_d)(X_je-x_e) - delta_de e_d ]
"""
import numpy as np


class GaussianProcess(object):
    def __init__(self, inputs, targets, theta=None):
        self.inputs = np.ascontiguousarray(inputs, dtype=np.float64)
        self.targets = np.ascontiguousarray(targets, dtype=np.float64)
        self.n, self.D = self.inputs.shape
        if theta is not None:
            self._set_params(np.asarray(theta, dtype=np.float64))

    # -- hyper-parameters -------------------------------------------------
    def _set_params(self, theta):
        assert theta.shape == (self.D + 2,)
        self.theta = theta.copy()
        e = np.exp(theta)
        Xs = self.inputs * np.sqrt(e[:self.D])
        d2 = ((Xs[:, None, :] - Xs[None, :, :]) ** 2).sum(axis=2)
        self.Z = e[self.D] * np.exp(-0.5 * d2)
        self.Q = self.Z + e[self.D + 1] * np.eye(self.n)
        self.invQ = np.linalg.inv(self.Q)
        # alpha through a solve (better conditioned than invQ @ t); both agree
        # to rounding and ``invQt`` is what every consumer reads.
        self.invQt = np.linalg.solve(self.Q, self.targets)

    # -- kernel row --------------------------------------------------------
    def _k(self, testing):
        e = np.exp(self.theta)
        diff = self.inputs[None, :, :] - testing[:, None, :]          # (nn, n, D)
        r2 = (e[:self.D] * diff * diff).sum(axis=2)                  # (nn, n)
        return e[self.D] * np.exp(-0.5 * r2), diff, e

    def predict(self, testing, do_deriv=True, do_unc=True):
        """Same structure as the package's ``predict``: scaled squared distances by
        ``cdist``, kernel matrix, then matrix products with ``invQt`` / ``invQ``.
        The derivative sum_j w_j (X_jd - x_d) is evaluated as W^T X - x mu on
        inputs centred on the training mean (algebraically identical, keeps the
        subtraction well conditioned)."""
        from scipy.spatial import distance as dist
        testing = np.atleast_2d(np.asarray(testing, dtype=np.float64))
        assert testing.shape[1] == self.D
        e = np.exp(self.theta)
        c = self.inputs.mean(axis=0)
        Xc = self.inputs - c
        root = np.sqrt(e[:self.D])
        out_mu = np.empty(testing.shape[0])
        out_var = np.empty(testing.shape[0]) if do_unc else None
        out_d = np.empty(testing.shape) if do_deriv else None
        step = 16384
        for s in range(0, testing.shape[0], step):
            sl = slice(s, s + step)
            tc = testing[sl] - c
            a = dist.cdist(root * Xc, root * tc, 'sqeuclidean')          # (n, nn)
            a = e[self.D] * np.exp(-0.5 * a)
            w = a * self.invQt[:, None]
            mu = w.sum(axis=0)
            out_mu[sl] = mu
            if do_unc:
                out_var[sl] = e[self.D] - np.sum(a * np.dot(self.invQ, a), axis=0)
            if do_deriv:
                out_d[sl] = e[:self.D] * (np.dot(w.T, Xc) - tc * mu[:, None])
        if do_deriv and do_unc:
            return out_mu, out_var, out_d
        if do_deriv:
            return out_mu, out_d
        if do_unc:
            return out_mu, out_var
        return out_mu

    def predict_direct(self, testing):
        """Term-by-term evaluation of the defining sums (slow; cross-check of
        ``predict`` in tests/test_oracle_golden.py)."""
        testing = np.atleast_2d(np.asarray(testing, dtype=np.float64))
        k, diff, e = self._k(testing)
        w = k * self.invQt[None, :]
        return w.sum(axis=1), e[:self.D] * np.einsum("ij,ijd->id", w, diff)

    def hessian(self, testing):
        testing = np.atleast_2d(np.asarray(testing, dtype=np.float64))
        k, diff, e = self._k(testing)
        w = k * self.invQt[None, :]
        sd = e[:self.D] * diff                                       # (nn, n, D)
        hess = np.einsum("ij,ijd,ije->ide", w, sd, sd)
        hess -= np.einsum("i,de->ide", w.sum(axis=1), np.diag(e[:self.D]))
        return hess


def make_synthetic_gp(rng, mean, sigma, n_train=250, w_scale=0.03,
                      sigma_f2=0.05, nugget=1e-6, offset=0.2, amp=0.1):
    """Smooth synthetic emulator around a prior (SURVEY.md section 8d recipe):
    training inputs ~ N(mean, (2 sigma)^2), targets offset + amp*tanh(z.w),
    inverse squared length scales 1/(4 sigma)^2."""
    mean = np.asarray(mean, dtype=np.float64)
    sigma = np.asarray(sigma, dtype=np.float64)
    D = mean.shape[0]
    X = mean + 2.0 * sigma * rng.standard_normal((n_train, D))
    w = rng.standard_normal(D) * w_scale
    t = offset + amp * np.tanh(((X - mean) / sigma) @ w)
    theta = np.concatenate([np.log(1.0 / (4.0 * sigma) ** 2),
                            [np.log(sigma_f2), np.log(nugget)]])
    return GaussianProcess(X, t, theta)
