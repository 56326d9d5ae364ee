"""GPU parity: state propagation / prior blending
against the reference's outputs (tests/golden/kat_propagators.npz, including the
numbers of the reference's own tests/test_kf.py) and the oracle."""
import numpy as np
import pytest
import scipy.sparse as sp

import golden_io as gio
import gpu_util as gu
import kafka_inferenceengine_b200 as kb
from kafka_inferenceengine_b200.device import BlockDiag, DeviceVector
from oracle import kafka_cpu as kc

pytestmark = pytest.mark.gpu


def _blocks(dense, n):
    N = dense.shape[0] // n
    return np.array([dense[i * n:(i + 1) * n, i * n:(i + 1) * n] for i in range(N)])


def test_reference_kat_standard_kalman():
    # /root/reference/tests/test_kf.py:19-27, same call
    x_analysis = np.ones(3)
    P_analysis = np.eye(3)
    M_matrix = 2. * np.eye(3)
    Q_matrix = np.eye(3) * 0.5
    x_forecast, P_forecast, P_forecast_inverse = kb.propagate_standard_kalman(
        x_analysis, P_analysis, None, M_matrix, Q_matrix)
    assert P_forecast_inverse is None
    assert np.all(np.asarray(x_forecast) == 2. * x_analysis)
    assert np.all(P_forecast.toarray() == (P_analysis + np.eye(3) * 0.5))


def test_reference_kat_information_filter():
    # /root/reference/tests/test_kf.py:30-54, same call (the name it imports does
    # not exist in the reference; the numbers are those of ..._approx_SLOW, the
    # comment matrix those of ..._SLOW)
    M_matrix = np.eye(7)
    sigma = np.array([0.12, 0.7, 0.0959, 0.15, 1.5, 0.2, 0.5])
    x_analysis = np.array([0.17, 1.0, 0.1, 0.7, 2.0, 0.18, np.exp(-0.5 * 1.5)])
    Pd = np.diag(sigma ** 2).astype(np.float32)
    Pd[5, 2] = 0.8862 * 0.0959 * 0.2
    Pd[2, 5] = 0.8862 * 0.0959 * 0.2
    Pi = np.linalg.inv(Pd)
    Q_matrix = np.eye(7) * 0.1
    x_forecast, P_forecast, P_forecast_inverse = kb.propagate_information_filter_approx_SLOW(
        x_analysis, None, Pi, M_matrix, Q_matrix)
    assert np.allclose(np.array(P_forecast_inverse.todense()).squeeze().diagonal(),
                       np.array([8.74, 1.69, 9.81, 8.16, 0.43, 9.21, 2.86]), atol=0.01)
    assert np.array_equal(np.asarray(x_forecast), x_analysis)
    _, _, exact = kb.propagate_information_filter_SLOW(x_analysis, None, Pi, M_matrix, Q_matrix)
    want = np.diag([8.74, 1.69, 9.33, 8.16, 0.43, 7.28, 2.86])
    want[2, 5] = want[5, 2] = -1.13
    assert np.allclose(exact.toarray(), want, atol=0.01)
    g = gio.load("kat_propagators")
    assert np.allclose(P_forecast_inverse.diagonal(), g["test_kf_approx_diag"], rtol=1e-13)
    assert np.allclose(exact.toarray(), g["test_kf_exact"], rtol=1e-9, atol=1e-11)


def test_propagators_vs_reference_outputs():
    g = gio.load("kat_propagators")
    x_a, q = g["x_a"], g["q"]
    Pi = sp.csr_matrix(g["Pinv"])
    M = sp.eye(21, format="csr")
    Q = sp.diags(q).tocsr()
    _, _, p = kb.propagate_information_filter_approx_SLOW(x_a, None, Pi, M, Q)
    assert np.allclose(p.toarray(), g["approx_Pinv"], rtol=1e-13, atol=0)
    xf, _, p = kb.propagate_information_filter_SLOW(x_a, None, Pi, M, Q)
    assert np.array_equal(np.asarray(xf), x_a)
    assert np.allclose(p.toarray(), g["exact_Pinv"], rtol=1e-9, atol=1e-10)
    xf, _, p = kb.propagate_information_filter_LAI(x_a, None, Pi, M, Q)
    assert np.array_equal(np.asarray(xf), g["lai_x"])
    assert np.array_equal(p.toarray().astype(np.float32), g["lai_Pinv"].astype(np.float32))
    xf, _, p = kb.no_propagation(x_a, None, Pi, M, Q)
    assert np.array_equal(np.asarray(xf), g["noprop_x"])
    assert np.array_equal(p.toarray().astype(np.float32), g["noprop_Pinv"].astype(np.float32))
    mean, cov, inv = kb.tip_prior()
    Pc = sp.block_diag([cov.astype(np.float64)] * 3).tocsr()
    xf, pf, none = kb.propagate_standard_kalman(x_a, Pc, None, M * 2.0, Q)
    assert none is None and np.array_equal(np.asarray(xf), g["std_x"])
    assert np.allclose(pf.toarray(), g["std_P"], rtol=1e-15)
    prior_mean = np.tile(mean, 3)
    prior_inv = sp.block_diag([inv.astype(np.float64)] * 3).tocsr()
    xc, C = kb.blend_prior(prior_mean, prior_inv, x_a, Pi)
    assert np.allclose(C.toarray(), g["blend_C"], rtol=1e-15)
    assert np.allclose(np.asarray(xc), g["blend_x"], rtol=1e-5)          # reference solves in float32
    xo, Co = kc.blend_prior(prior_mean, _blocks(prior_inv.toarray(), 7), x_a, _blocks(g["Pinv"], 7))
    assert np.allclose(np.asarray(xc), xo, rtol=1e-11)
    # broadcast prior path gives the same numbers
    xc2, C2 = kb.blend_prior(mean, BlockDiag.from_block(inv.astype(np.float64), 3, gu.DEV), x_a, Pi)
    assert np.array_equal(np.asarray(xc2), np.asarray(xc)) and np.array_equal(C2.blocks(), C.blocks())


@pytest.mark.parametrize("P", [1, 2, 5, 7, 10, 12, 16])
def test_propagators_vs_oracle_all_sizes(P):
    rng = np.random.default_rng(P)
    n = 1003
    Pinv = gu.random_spd(rng, n, P)
    x_a = rng.standard_normal(n * P)
    q = rng.uniform(0.0, 0.3, n * P)
    q[::5] = 0.0                                             # Q = 0 entries (kafka_test.py:207-208 style)
    A = BlockDiag.from_blocks(Pinv, gu.DEV)
    Q = kb.DiagonalMatrix(q, P, gu.DEV)
    idx = np.arange(P)
    diag = Pinv[:, idx, idx].ravel()
    _, _, p = kb.propagate_information_filter_approx_SLOW(x_a, None, A, None, Q)
    _, want = kc.propagate_information_filter_approx(x_a, diag, q)
    assert np.allclose(p.diagonal(), want, rtol=1e-14)
    off = p.blocks().copy()
    off[:, idx, idx] = 0
    assert not off.any()
    _, _, p = kb.propagate_information_filter_SLOW(x_a, None, A, None, Q)
    _, want = kc.propagate_information_filter_exact(x_a, Pinv, q)
    assert np.allclose(p.blocks(), want, rtol=1e-10, atol=1e-11)
    xf, pf, _ = kb.propagate_standard_kalman(x_a, A, None, None, Q)
    xo, po = kc.propagate_standard_kalman(x_a, Pinv, q)
    assert np.array_equal(np.asarray(xf), xo) and np.array_equal(pf.blocks(), po)
    slot = P - 1
    pm = rng.standard_normal(P)
    pinv = gu.random_spd(rng, 1, P)[0]
    xf, _, p = kb.propagate_information_filter_carry_slot(x_a, None, A, None, Q, pm, pinv, slot)
    xw = np.tile(pm, n)
    xw[slot::P] = x_a[slot::P]
    bw = np.repeat(pinv[None], n, axis=0)
    bw[:, slot, slot] = 1.0 / (1.0 / diag[slot::P] + q[slot::P])
    assert np.array_equal(np.asarray(xf), xw)
    assert np.array_equal(p.blocks(), bw.astype(np.float32).astype(np.float64))


def test_lai_advance_large_modis_tile_properties():
    """Config 3 size (2400 x 2400 = 5.76 M px, 7 params): idempotence of the
    prior reset and the closed form of the carried slot."""
    import torch
    n, P = 2400 * 2400, 7
    gen = torch.Generator(device=gu.DEV).manual_seed(1)
    x = DeviceVector.empty(n, P, gu.DEV)
    x.data.copy_(torch.rand(x.data.shape, generator=gen, dtype=torch.float64, device=gu.DEV))
    _, _, inv = kb.tip_prior()
    A = BlockDiag.from_block(inv.astype(np.float64), n, gu.DEV).materialise()
    A.data[27] = 1.0 + torch.rand(A.data.shape[1], generator=gen, dtype=torch.float64, device=gu.DEV)
    q = DeviceVector.empty(n, P, gu.DEV)
    q.data.zero_()
    q.data[6] = 0.04
    x1, _, p1 = kb.propagate_information_filter_LAI(x, None, A, None, q)
    x2, _, p2 = kb.propagate_information_filter_LAI(x1, None, A, None, q)
    assert torch.equal(x1.data[:, :n], x2.data[:, :n]) and torch.equal(p1.data[:, :n], p2.data[:, :n])
    want = (1.0 / (1.0 / A.data[27, :n] + 0.04)).float().double()
    assert torch.equal(p1.data[27, :n], want)
    assert torch.equal(x1.data[6, :n], x.data[6, :n])
    mean, _, _ = kb.tip_prior()
    assert torch.all(x1.data[0, :n] == mean[0])


def _make_indefinite(rng, blocks, which):
    """Flip the sign of one eigen-direction of the chosen SPD blocks."""
    out = blocks.copy()
    for i in which:
        w, V = np.linalg.eigh(out[i])
        k = int(rng.integers(0, len(w)))
        w[k] = -w[k]
        out[i] = (V * w) @ V.T
        out[i] = 0.5 * (out[i] + out[i].T)
    return out


@pytest.mark.parametrize("P", [1, 2, 3, 7, 10, 13, 16])
def test_indefinite_blocks_are_solved_like_the_reference_lu(P):
    """The reference's exact propagator and prior blend solve with a sparse LU (kf_tools.py:236-243,
    :93-94), so a block that is not positive definite still gets the solution of its equations.
    Here such pixels leave the Cholesky path and are solved by elimination with partial pivoting;
    a singular block is handed on unchanged / left at the forecast."""
    rng = np.random.default_rng(40 + P)
    n = 700
    Pinv = gu.random_spd(rng, n, P)
    which = np.array(sorted(set([0, 31, 32, 127, 128, n - 1]) | set(rng.integers(0, n, 30).tolist())))
    Pinv = _make_indefinite(rng, Pinv, which)
    x_a = rng.standard_normal(n * P)
    # inflation large enough for I + S Pinv S to lose positive definiteness on the flipped direction
    q = rng.uniform(2.0, 3.0, n * P) / np.repeat(np.abs(np.linalg.eigvalsh(Pinv)).min(axis=1), P)
    A = BlockDiag.from_blocks(Pinv, gu.DEV)
    Q = kb.DiagonalMatrix(q, P, gu.DEV)
    _, _, p = kb.propagate_information_filter_SLOW(x_a, None, A, None, Q)
    _, want = kc.propagate_information_filter_exact(x_a, Pinv, q)
    inner = np.eye(P)[None] + np.sqrt(q).reshape(n, P, 1) * Pinv * np.sqrt(q).reshape(n, 1, P)
    assert (np.linalg.eigvalsh(inner)[which].min(axis=1) < 0).all()          # these pixels did take the other path
    got = p.blocks()
    lower = np.tril(np.ones((P, P), dtype=bool))
    scale = np.abs(want).max(axis=(1, 2), keepdims=True)
    assert np.max(np.abs(got - want)[:, lower] / np.broadcast_to(scale, want.shape)[:, lower]) < 1e-8
    keep = np.ones(n, dtype=bool)
    keep[which] = False
    assert np.allclose(got[keep], want[keep], rtol=1e-10, atol=1e-11)

    # prior blend: C = Pinv + prior_inv indefinite on the same pixels
    prior_inv = 1e-3 * gu.random_spd(rng, n, P)
    prior_mean = rng.standard_normal(n * P)
    xc, C = kb.blend_prior(prior_mean, BlockDiag.from_blocks(prior_inv, gu.DEV), x_a, A)
    xo, Co = kc.blend_prior(prior_mean, prior_inv, x_a, Pinv)
    ev = np.linalg.eigvalsh(Co)
    assert (ev[which].min(axis=1) < 0).all()
    assert np.allclose(C.blocks(), Co, rtol=1e-14)
    err = np.abs(np.asarray(xc) - xo).reshape(n, P).max(axis=1) / np.abs(xo).reshape(n, P).max(axis=1)
    cond = np.abs(ev).max(axis=1) / np.abs(ev).min(axis=1)
    assert (err < 1e-13 * np.maximum(cond, 10.0)).all()
    assert (err[keep] < 1e-10).all()


def test_exact_propagator_on_indefinite_blocks_vs_unmodified_reference():
    """tests/golden/solve_indefinite.npz: the unmodified propagate_information_filter_SLOW
    (kf_tools.py:208-245, a float64 spsolve) on analysis blocks that are not positive definite for
    every third pixel; for 19 of the 55 pixels I + S P^-1 S is not positive definite either, so the
    kernel's Cholesky route fails and its pivoted elimination has to give the reference's answer."""
    g = gio.load("solve_indefinite")
    P = 7
    blocks, q = g["blocks"], g["q"]
    n = blocks.shape[0]
    inner = np.eye(P)[None] + np.sqrt(q).reshape(n, P, 1) * blocks * np.sqrt(q).reshape(n, 1, P)
    assert int((np.linalg.eigvalsh(inner).min(axis=1) < 0).sum()) == 19
    _, _, p = kb.propagate_information_filter_SLOW(g["x_f"], None, BlockDiag.from_blocks(blocks, gu.DEV), None,
                                                   kb.DiagonalMatrix(q, P, gu.DEV))
    got = p.blocks()
    lower = np.tril(np.ones((P, P), dtype=bool))
    scale = np.broadcast_to(np.abs(g["ref_exact_blocks"]).max(axis=(1, 2), keepdims=True), got.shape)
    assert np.max((np.abs(got - g["ref_exact_blocks"]) / scale)[:, lower]) < 1e-10
