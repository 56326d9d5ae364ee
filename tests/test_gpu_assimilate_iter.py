"""GPU parity of ``kafka_assimilate_iter`` (one do_all_bands iteration behind a
single C-ABI call, kafka/linear_kf.py:258-296): iterated from the forecast to
convergence by a plain ctypes loop, against oracle B's ``do_all_bands`` -- same
number of solves, posterior mean and diag(P^-1) within 1e-6 relative -- and
bit-equal to kafka_gp_linearize + kafka_solve_bands called one after the other."""
import numpy as np
import pytest
import torch

import gpu_util as gu
from kafka_inferenceengine_b200 import _lib
from kafka_inferenceengine_b200 import synthetic as syn
from kafka_inferenceengine_b200.device import padded, tri_count
from kafka_inferenceengine_b200.inference.utils import emulator_table
from oracle import kafka_cpu as kc
from oracle.gp import GaussianProcess

pytestmark = pytest.mark.gpu

TOL = 1e-6


def _setup(rng, n, P, mappers, mean, sigma, n_train, round_f32):
    B = len(mappers)
    gps = [syn.make_emulator(rng, mean[m], sigma[m], n_train=n_train, w_scale=0.1, cls=GaussianProcess)
           for m in mappers]
    x_true = mean + 0.7 * sigma * rng.standard_normal((n, P))
    x_f = np.tile(mean, n)
    Pinv = np.tile(np.diag(1.0 / sigma ** 2)[None], (n, 1, 1))
    obs = rng.random((B, n)) < 0.85
    ys, ws = [], []
    for b, (gp, m) in enumerate(zip(gps, mappers)):
        mu, _ = gp.predict(x_true[:, m], do_unc=False)
        sig = np.maximum(0.05 * np.abs(mu), 2.5e-3)
        ys.append(np.where(obs[b], mu + sig * rng.standard_normal(n), 0.0))
        ws.append(np.where(obs[b], 1.0 / sig ** 2, 0.0))
    ops = [(lambda xp, gp=gp, m=m, o=obs[b]: kc.emulator_operator(gp, xp, P, o, state_mapper=m, round_f32=round_f32))
           for b, (gp, m) in enumerate(zip(gps, mappers))]
    return gps, obs, ys, ws, x_f, Pinv, ops


def _gp_bands(gps, mappers):
    arr = (_lib.GpBand * len(gps))()
    tabs = []
    for b, (gp, m) in enumerate(zip(gps, mappers)):
        tab = emulator_table(gp, torch.device(gu.DEV))
        tabs.append(tab)
        arr[b].table = tab.table.data_ptr()
        arr[b].n_train = tab.n_train
        arr[b].n_dim = len(m)
        for d, s in enumerate(m):
            arr[b].mapper[d] = int(s)
    return arr, tabs


@pytest.mark.parametrize("case", ["tip_mappers", "prosail"])
@pytest.mark.parametrize("round_f32", [False, True])
def test_assimilate_iter_loop_matches_oracle(case, round_f32):
    rng = np.random.default_rng(77)
    if case == "tip_mappers":                       # utils.py:148-153
        P, mappers, mean, sigma, T, n = 7, [np.array([0, 1, 6, 2]), np.array([3, 4, 6, 5])], syn.TIP_MEAN, \
            syn.TIP_SIGMA, 120, 1500
    else:                                           # utils.py:181-219
        P, mappers, mean, sigma, T, n = 10, [np.arange(10)] * 3, syn.SAIL_MEAN, syn.SAIL_SIGMA, 100, 900
    B, D = len(mappers), len(mappers[0])
    gps, obs, ys, ws, x_f, Pinv, ops = _setup(rng, n, P, mappers, mean, sigma, T, round_f32)
    x_o, A_o, innov_o, n_o, norms_o = kc.do_all_bands(ops, ys, ws, obs, x_f, Pinv, P, round_f32=round_f32)

    lib = _lib.get_lib()
    ld = padded(n)
    L = tri_count(P)
    arr, tabs = _gp_bands(gps, mappers)
    y_t = gu.planar(np.stack(ys, axis=1), ld)
    w_t = gu.planar(np.stack(ws, axis=1), ld)
    m_t = torch.zeros((B, ld), dtype=torch.uint8, device=gu.DEV)
    m_t[:, :n] = torch.from_numpy(obs.astype(np.uint8)).to(gu.DEV)
    xf_t = gu.planar(x_f.reshape(n, P), ld)
    pinv_t = gu.pack_blocks(Pinv, ld)
    h0 = torch.empty((B, ld), dtype=torch.float64, device=gu.DEV)
    jac = torch.empty((B, D, ld), dtype=torch.float64, device=gu.DEV)
    xa = xf_t.clone()
    xb = torch.zeros_like(xa)
    a_out = torch.zeros((L, ld), dtype=torch.float64, device=gu.DEV)
    innov = torch.zeros((B, ld), dtype=torch.float64, device=gu.DEV)
    norm2 = torch.zeros(1, dtype=torch.float64, device=gu.DEV)
    n_bad = torch.zeros(1, dtype=torch.int32, device=gu.DEV)
    scratch = torch.empty(int(lib.kafka_solve_scratch_doubles(n)), dtype=torch.float64, device=gu.DEV)
    flags = _lib.ROUND_F32 if round_f32 else 0

    n_iter, norms = 1, []
    while True:                                     # the caller's loop, linear_kf.py:258-305
        _lib.check(lib.kafka_assimilate_iter(n, ld, P, arr, B, _lib.ptr(y_t), _lib.ptr(w_t), _lib.ptr(m_t),
                                             _lib.ptr(xa), _lib.ptr(xf_t), _lib.ptr(pinv_t), _lib.ptr(h0),
                                             _lib.ptr(jac), _lib.ptr(xb), _lib.ptr(a_out), _lib.ptr(innov), None,
                                             _lib.ptr(norm2), _lib.ptr(n_bad), _lib.ptr(scratch), flags,
                                             gu.stream()), "kafka_assimilate_iter")
        norm = float(norm2.item()) ** 0.5 / (n * P)
        norms.append(norm)
        if n_iter == 1:                             # same inputs through the two separate entry points
            h0_s = torch.empty_like(h0)
            jac_s = torch.empty_like(jac)
            _lib.check(lib.kafka_gp_linearize(_lib.ptr(xa), ld, n, arr, B, _lib.ptr(m_t), _lib.ptr(h0_s),
                                              _lib.ptr(jac_s), flags, gu.stream()), "kafka_gp_linearize")
            sb = (_lib.SolveBand * B)()
            for b in range(B):
                sb[b].kind = _lib.BAND_ARRAYS
                sb[b].n_dim = D
                for d in range(D):
                    sb[b].mapper[d] = int(mappers[b][d])
                sb[b].h0 = h0_s[b].data_ptr()
                sb[b].jac = jac_s[b].data_ptr()
                sb[b].y = y_t[b].data_ptr()
                sb[b].w = w_t[b].data_ptr()
            x_s = torch.zeros_like(xb)
            a_s = torch.zeros_like(a_out)
            n2_s = torch.zeros_like(norm2)
            _lib.check(lib.kafka_solve_bands(n, ld, P, sb, B, _lib.ptr(xa), _lib.ptr(xf_t), _lib.ptr(pinv_t),
                                             _lib.ptr(x_s), _lib.ptr(a_s), None, None, _lib.ptr(n2_s),
                                             _lib.ptr(n_bad), _lib.ptr(scratch), flags, gu.stream()),
                       "kafka_solve_bands")
            assert torch.equal(x_s[:, :n], xb[:, :n]) and torch.equal(a_s[:, :n], a_out[:, :n])
            assert float(n2_s.item()) == float(norm2.item())
        assert int(n_bad.item()) == 0
        stop = (norm < 1e-3 and n_iter >= 2) or n_iter > 25
        xa, xb = xb, xa
        n_iter += 1
        if stop:
            break

    assert n_iter - 1 == n_o
    np.testing.assert_allclose(norms, norms_o, rtol=1e-5)
    x = gu.unplanar(xa, n).ravel()
    A = gu.unpack_blocks(a_out, n, P)
    idx = np.arange(P)
    ex = np.abs(x - x_o) / np.maximum(np.abs(x_o), 1e-3)
    ed = gu.rel(A[:, idx, idx], A_o[:, idx, idx])
    assert ex.max() < TOL, ex.max()
    assert ed.max() < TOL, ed.max()
    got_innov = innov[:, :n].cpu().numpy().ravel()
    assert np.abs(got_innov - innov_o).max() < 1e-6


def test_assimilate_iter_rejects_bad_arguments():
    lib = _lib.get_lib()
    arr = (_lib.GpBand * 1)()
    assert lib.kafka_assimilate_iter(8, 32, 7, arr, 0, None, None, None, None, None, None, None, None, None, None,
                                     None, None, None, None, None, 0, None) == -1
    assert b"n_bands" in lib.kafka_last_error()
    assert lib.kafka_assimilate_iter(8, 32, 7, arr, 1, None, None, None, None, None, None, None, None, None, None,
                                     None, None, None, None, None, 0, None) == -1
    assert b"null" in lib.kafka_last_error()
