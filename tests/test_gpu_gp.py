"""GPU parity: the GP emulator kernels against
the oracle's stand-in GaussianProcess (oracle/gp.py), through the C ABI."""
import numpy as np
import pytest
import torch

import golden_io as gio
import gpu_util as gu
import kafka_inferenceengine_b200 as kb
from kafka_inferenceengine_b200 import _lib
from kafka_inferenceengine_b200 import synthetic as syn
from kafka_inferenceengine_b200.device import DeviceVector, padded
from kafka_inferenceengine_b200.inference.utils import emulator_table, linearize_bands
from oracle import kafka_cpu as kc
from oracle.gp import GaussianProcess

pytestmark = pytest.mark.gpu

# fp64 tolerance on emulator outputs: the kernel evaluates an algebraically
# expanded form of the same sums, so agreement is at rounding level scaled by
# the cancellation in sum_j k_j alpha_j (worst case seen: 2e-9 for 500 training
# points in 4 dimensions; 1e-11 typical).  north_star asks 1e-6 on the posterior.
RTOL = 1e-8


def _gp(rng, D, T, w_scale=0.2):
    mean = rng.uniform(0.2, 2.0, D)
    sigma = rng.uniform(0.01, 0.5, D)
    return syn.make_emulator(rng, mean, sigma, n_train=T, w_scale=w_scale, cls=GaussianProcess), mean, sigma


def _scale(a):
    return np.maximum(np.maximum(np.abs(a), np.abs(a).max() * 1e-3), 1e-300)


@pytest.mark.parametrize("D,T", [(1, 5), (2, 33), (4, 120), (4, 500), (7, 64), (10, 250), (13, 40), (16, 100)])
def test_predict_mean_jacobian(D, T):
    rng = np.random.default_rng(100 * D + T)
    gp, mean, sigma = _gp(rng, D, T)
    n = 3001
    X = mean + 1.5 * sigma * rng.standard_normal((n, D))
    mu_o, jac_o = gp.predict(X, do_unc=False)
    mu, jac = kb.run_emulator(gp, X)                      # kafka_gp_predict through the C ABI
    assert np.max(np.abs(mu - mu_o) / _scale(mu_o)) < RTOL
    assert np.max(np.abs(jac - jac_o) / _scale(jac_o)) < RTOL


def test_predict_variance_and_hessian():
    rng = np.random.default_rng(11)
    for D, T in ((4, 60), (10, 250)):
        gp, mean, sigma = _gp(rng, D, T)
        n = 257
        X = mean + sigma * rng.standard_normal((n, D))
        _, var_o, _ = gp.predict(X, do_unc=True)
        hess_o = gp.hessian(X)
        tab = emulator_table(gp, torch.device(gu.DEV))
        ld = padded(n)
        xin = gu.planar(X, ld)
        mu = torch.empty(ld, dtype=torch.float64, device=gu.DEV)
        var = torch.empty(ld, dtype=torch.float64, device=gu.DEV)
        hess = torch.empty((D * D, ld), dtype=torch.float64, device=gu.DEV)
        lib = _lib.get_lib()
        _lib.check(lib.kafka_gp_predict(_lib.ptr(xin), ld, n, _lib.ptr(tab.table), T, D, _lib.ptr(tab.inv_q()),
                                        _lib.ptr(mu), None, _lib.ptr(var), gu.stream()), "predict")
        _lib.check(lib.kafka_gp_hessian(_lib.ptr(xin), ld, n, _lib.ptr(tab.table), T, D, _lib.ptr(hess),
                                        gu.stream()), "hessian")
        v = var[:n].cpu().numpy()
        # s_f - k^T Q^-1 k cancels to ~1e-6 of s_f: compare on the scale of s_f
        assert np.max(np.abs(v - var_o)) < 1e-8 * np.exp(gp.theta[D])
        h = hess[:, :n].t().reshape(n, D, D).cpu().numpy()
        assert np.max(np.abs(h - hess_o) / _scale(hess_o)) < RTOL


@pytest.mark.parametrize("D,T,n", [(3, 5, 7), (4, 33, 65), (10, 250, 1000), (7, 251, 129), (4, 500, 300), (2, 340, 64)])
def test_variance_tensor_core_tiles_vs_oracle(D, T, n):
    """kafka_gp_predict's variance through the fp64 tensor-core kernel (64-point tiles up to T ~ 340,
    32-point tiles beyond), odd and padded sizes, partial tiles: s_f - k^T Q^-1 k against the oracle
    on the scale of s_f, as the one-warp-per-point kernel is checked above."""
    rng = np.random.default_rng(1000 * D + T)
    gp, mean, sigma = _gp(rng, D, T)
    X = mean + sigma * rng.standard_normal((n, D))
    X[0] = mean + 50.0 * sigma                                   # a point far from the training set: variance = s_f
    _, var_o, _ = gp.predict(X, do_unc=True)
    tab = emulator_table(gp, torch.device(gu.DEV))
    ld = padded(n)
    xin = gu.planar(X, ld)
    mu = torch.empty(ld, dtype=torch.float64, device=gu.DEV)
    var = torch.full((ld,), float("nan"), dtype=torch.float64, device=gu.DEV)
    _lib.check(_lib.get_lib().kafka_gp_predict(_lib.ptr(xin), ld, n, _lib.ptr(tab.table), T, D, _lib.ptr(tab.inv_q()),
                                               _lib.ptr(mu), None, _lib.ptr(var), gu.stream()), "predict")
    v = var[:n].cpu().numpy()
    s_f = np.exp(gp.theta[D])
    assert np.isfinite(v).all()
    assert np.max(np.abs(v - var_o)) < 1e-8 * s_f
    assert abs(v[0] - s_f) < 1e-9 * s_f
    assert torch.isnan(var[n:]).all()                            # nothing written past the last point


def test_exp_underflow_and_far_points():
    """Inputs far from every training point: kernel values underflow to 0 (not NaN)."""
    rng = np.random.default_rng(12)
    gp, mean, sigma = _gp(rng, 4, 50)
    X = np.vstack([mean + 1e3 * sigma, mean - 40 * sigma, mean])
    mu_o, jac_o = gp.predict(X, do_unc=False)
    mu, jac = kb.run_emulator(gp, X)
    assert np.isfinite(mu).all() and np.isfinite(jac).all()
    assert mu[0] == 0.0 and np.all(jac[0] == 0.0)
    assert np.allclose(mu, mu_o, rtol=1e-9, atol=1e-300)
    assert np.allclose(jac, jac_o, rtol=1e-8, atol=1e-300)


@pytest.mark.parametrize("outlier", [30.0, 45.0, 8000.0])
def test_outlier_training_point(outlier):
    """One training point ``outlier`` length scales from the rest.  Near the centre its kernel
    value underflows (30: not quite; 45: yes, inside the unchecked exponent of the hot loop;
    8000: the packer flags the table as wide and the checked exponent runs); near the outlier
    it is the only point that counts (far-pixel path)."""
    rng = np.random.default_rng(13)
    D, T = 4, 40
    gp0, mean, sigma = _gp(rng, D, T)
    X = gp0.inputs.copy()
    ls = 4.0 * sigma                                        # length scales of make_emulator
    X[7] = mean + outlier * ls / np.sqrt(D)
    gp = GaussianProcess(X, gp0.targets, gp0.theta)
    tab = _lib.gp_pack(gp.inputs, gp.theta, gp.invQt)
    assert (tab[2 * D + 2] != 0.0) == (outlier > 1000.0)
    pts = np.vstack([mean + 0.5 * sigma * rng.standard_normal((64, D)),
                     X[7] + 0.5 * sigma * rng.standard_normal((64, D)),
                     mean + 0.5 * (X[7] - mean)[None, :] + sigma * rng.standard_normal((64, D))])
    mu_o, jac_o = gp.predict_direct(pts)
    mu, jac = kb.run_emulator(gp, pts)
    assert np.isfinite(mu).all() and np.isfinite(jac).all()
    # at a pixel next to an outlier 8000 length scales out, the expanded exponent e_j + z_j.x' - x'.x'/2
    # cancels terms of 1e11 (in units of ln2/1024): its rounding level is 1e-8 relative in the kernel value
    # (measured 1.2e-8); the direct form of the oracle does not cancel
    rtol = 3e-8 if outlier > 1000.0 else 1e-8
    assert np.allclose(mu, mu_o, rtol=rtol, atol=1e-12 * np.abs(mu_o).max())
    assert np.allclose(jac, jac_o, rtol=rtol, atol=1e-12 * np.abs(jac_o).max())


@pytest.mark.parametrize("n", [1, 127, 129, 4095, 40000])
def test_linearize_bands_mask_mapper_rounding(n):
    """kafka_gp_linearize on the TIP layout: two bands, per-band state mappers
    (utils.py:148-153), unobserved rows zero (utils.py:199-215), float32 rounding
    of the operator (utils.py:193-195).  n spans both pixels-per-thread variants."""
    rng = np.random.default_rng(n)
    P = 7
    mappers = [np.array([0, 1, 6, 2]), np.array([3, 4, 6, 5])]
    gps = [syn.make_emulator(rng, syn.TIP_MEAN[m], syn.TIP_SIGMA[m], n_train=90, w_scale=0.15, cls=GaussianProcess)
           for m in mappers]
    x = (syn.TIP_MEAN + 0.4 * syn.TIP_SIGMA * rng.standard_normal((n, P))).ravel()
    obs = rng.random((2, n)) < 0.8
    xv = DeviceVector.from_interleaved(x, P, gu.DEV)
    m = torch.zeros((2, xv.ld), dtype=torch.uint8, device=gu.DEV)
    m[:, :n] = torch.from_numpy(obs.astype(np.uint8)).to(gu.DEV)
    tabs = [emulator_table(g, torch.device(gu.DEV)) for g in gps]
    for r32 in (True, False):
        h0, jac = linearize_bands(xv, list(zip(tabs, mappers)), m, r32)
        for b in range(2):
            H0o, Jo = kc.emulator_operator(gps[b], x, P, obs[b], state_mapper=mappers[b], round_f32=r32)
            H0 = h0[b, :n].cpu().numpy()
            J = np.zeros((n, P))
            J[:, mappers[b]] = jac[b, :, :n].t().cpu().numpy()
            assert np.all(H0[~obs[b]] == 0) and np.all(J[~obs[b]] == 0)
            if r32:
                # both sides hold float32 values: equal up to one float32 ulp where
                # the fp64 results straddle a rounding boundary
                assert np.array_equal(H0.astype(np.float32).astype(np.float64), H0)
                assert np.allclose(H0, H0o, rtol=1.3e-7, atol=1e-30)
                assert np.allclose(J, Jo, rtol=1.3e-7, atol=1e-12)
                if n >= 1000:
                    assert np.mean(H0 == H0o) > 0.99
            else:
                assert np.max(np.abs(H0 - H0o) / _scale(H0o)) < RTOL
                assert np.max(np.abs(J - Jo) / _scale(Jo)) < RTOL


def test_operator_call_returns_reference_tuple():
    """Calling the descriptor like the reference's builder (utils.py:181-219)
    gives (H0 float32, H csr float32) equal to the unmodified reference's output
    stored in the solve_s2 fixture."""
    g = gio.load("solve_s2")
    mask = g["mask"]
    gps = gio.emulators(g, syn.EmulatorData)
    P = 10
    n = int(mask.sum())
    for b in (0, 4, 9):
        H0, H = kb.create_prosail_observation_operator(P, gps[b], None, g["obs_mask_0_%d" % b], mask, g["x_lin"], b)
        assert H0.dtype == np.float32 and H.dtype == np.float32 and H.shape == (n, n * P)
        assert np.allclose(H0, g["ref_H0"][b], rtol=2e-7, atol=1e-9)
        Hc = H.tocoo()
        J = np.zeros((n, P), dtype=np.float32)
        J[Hc.row, Hc.col - P * Hc.row] = Hc.data
        assert np.allclose(J, g["ref_J"][b], rtol=2e-7, atol=1e-9)
        obs = g["obs_mask_0_%d" % b][mask]
        assert H.getnnz(axis=1)[~obs].sum() == 0


def test_full_s2_tile_size_properties():
    """At a BASELINE-sized pixel count (one 1/64 S2 tile stripe) check
    size-independent properties: permutation equivariance and exact
    reproducibility of the emulator kernel."""
    rng = np.random.default_rng(21)
    P = 10
    gp = syn.make_emulator(rng, syn.SAIL_MEAN, syn.SAIL_SIGMA, n_train=250, cls=syn.EmulatorData)
    n = 10980 * 10980 // 64
    gen = torch.Generator(device=gu.DEV).manual_seed(5)
    x = DeviceVector.empty(n, P, gu.DEV)
    x.data.copy_(torch.from_numpy(syn.SAIL_MEAN).to(gu.DEV)[:, None] +
                 torch.from_numpy(syn.SAIL_SIGMA).to(gu.DEV)[:, None] *
                 torch.randn((P, x.ld), generator=gen, dtype=torch.float64, device=gu.DEV))
    tab = emulator_table(gp, torch.device(gu.DEV))
    h0, jac = linearize_bands(x, [(tab, np.arange(P))], None, False)
    h0b, jacb = linearize_bands(x, [(tab, np.arange(P))], None, False)
    assert torch.equal(h0[:, :n], h0b[:, :n]) and torch.equal(jac[:, :, :n], jacb[:, :, :n])
    perm = torch.randperm(n, generator=gen, device=gu.DEV)
    xp = DeviceVector.empty(n, P, gu.DEV)
    xp.data[:, :n] = x.data[:, :n][:, perm]
    h0p, jacp = linearize_bands(xp, [(tab, np.arange(P))], None, False)
    assert torch.equal(h0p[0, :n], h0[0, :n][perm]) and torch.equal(jacp[0, :, :n], jac[0, :, :n][:, perm])
    # spot check against the oracle
    sel = np.arange(0, n, n // 200)
    mu_o, jac_o = GaussianProcess(gp.inputs, gp.targets, gp.theta).predict(
        x.data[:, sel].t().cpu().numpy(), do_unc=False)
    assert np.max(np.abs(h0[0, sel].cpu().numpy() - mu_o) / _scale(mu_o)) < RTOL
    assert np.max(np.abs(jac[0][:, sel].t().cpu().numpy() - jac_o) / _scale(jac_o)) < RTOL
