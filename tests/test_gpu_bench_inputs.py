"""Parity on the benchmark's OWN data: pixels drawn from the very tensors bench.py times
(``bench.synth_on_device(gps, n, 1234, device)`` with ``bench.make_emulators()``) go through
``LinearKalman.do_all_bands`` and are checked against oracle B (fp64 restatement of
kafka/linear_kf.py:245-323 + kafka/inference/solvers.py:100-145) to the north_star
tolerance, with equal iteration counts -- so the headline workload is covered by value,
not only by shape."""
import datetime as dt

import numpy as np
import pytest

import gpu_util as gu

pytestmark = pytest.mark.gpu

import bench                                                      # noqa: E402
import kafka_inferenceengine_b200 as kb                           # noqa: E402
from kafka_inferenceengine_b200 import synthetic as syn           # noqa: E402
from kafka_inferenceengine_b200.device import BlockDiag, DeviceVector   # noqa: E402
from oracle import kafka_cpu as kc                                # noqa: E402
from oracle.gp import GaussianProcess                             # noqa: E402

TOL = 1e-6


@pytest.mark.parametrize("host", [False, True])
def test_bench_workload_sample_matches_the_oracle(host):
    import torch
    device = torch.device(gu.DEV)
    n, P, B = 8192, bench.P, bench.B
    gps = bench.make_emulators()
    rho, inv, mask = bench.synth_on_device(gps, n, 1234, device)           # rank 0's seed in bench.py
    shape = (64, 128)
    date = dt.datetime(2017, 7, 3, 12)
    obs = (bench.HostObservations if host else bench.DeviceObservations)([date], gps, rho, inv, mask, shape)
    data = [obs.get_band_data(date, b) for b in range(B)]
    inv_block = np.diag(1.0 / syn.SAIL_SIGMA ** 2).astype(np.float32).astype(np.float64)
    kf = kb.LinearKalman(obs, None, np.ones(shape, dtype=bool), kb.create_prosail_observation_operator,
                         syn.S2_PARAMETERS, state_propagation=None, prior=None, linear=False, device=device)
    x, _, A, innov = kf.do_all_bands(date, data, DeviceVector.broadcast(syn.SAIL_MEAN, n, device), None,
                                     BlockDiag.from_block(inv_block, n, device))

    y = rho.cpu().numpy()
    m = mask.cpu().numpy().astype(bool)
    w = kc.gate_inverse_variance(inv.cpu().numpy(), m)
    ogps = [GaussianProcess(g.inputs, g.targets, g.theta) for g in gps]
    ops = [(lambda xp, gp=ogps[b], mm=m[b]: kc.emulator_operator(gp, xp, P, mm)) for b in range(B)]
    xo, Ao, innov_o, iters, norms = kc.do_all_bands(ops, list(y), list(w), list(m), np.tile(syn.SAIL_MEAN, n),
                                                    np.repeat(inv_block[None], n, axis=0), P)
    assert kf.last_iterations[-1] == iters
    assert np.allclose(kf.last_norms[-1], norms, rtol=1e-6)
    ex = np.abs(np.asarray(x) - xo) / np.maximum(np.abs(xo), 1e-3)
    idx = np.arange(P)
    ed = gu.rel(A.blocks()[:, idx, idx], Ao[:, idx, idx])
    assert ex.max() < TOL and ed.max() < TOL, (ex.max(), ed.max())
    assert np.max(gu.rel(A.blocks(), Ao, floor=1e-3 * np.abs(Ao).max())) < TOL
    assert np.allclose(np.asarray(innov), innov_o, rtol=1e-9, atol=1e-12)
    # the synthetic data are what bench.py says they are: ~90 % observed, 5 % relative sigma
    assert 0.85 < m.mean() < 0.95
    assert np.allclose(w[m], 1.0 / (0.05 * y[m]) ** 2, rtol=1e-12)


def test_full_shard_is_the_sum_of_its_chunks():
    """The whole bench shard (BASELINE configs[3] on one GPU: 1373 x 10980 = 15 075 540 px, 10 bands,
    10 parameters) through ``do_all_bands``, checked by a property that does not depend on the size:
    pixels are independent, so any window of the shard solved on its own -- with the same number of
    relinearisations -- must reproduce the shard's values for those pixels (the reference's drivers
    rely on exactly this when they cut a tile into 128 x 128 chunks, kafka_test_S2.py:202-205).
    Windows start off the 128-pixel tile grid and one ends at the ragged end of the shard.  A sample
    of the shard is also checked against oracle B."""
    import torch
    device = torch.device(gu.DEV)
    P, B = bench.P, bench.B
    shape = (bench.SHARD_ROWS, bench.TILE)
    n = shape[0] * shape[1]
    gps = bench.make_emulators()
    rho, inv, mask = bench.synth_on_device(gps, n, 1234, device)
    date = dt.datetime(2017, 7, 3, 12)
    inv_block = np.diag(1.0 / syn.SAIL_SIGMA ** 2).astype(np.float32).astype(np.float64)

    def solve(lo, hi, shp, min_iterations=2, tol=1e-3):
        r, i, m = (t[:, lo:hi].contiguous() for t in (rho, inv, mask))
        obs = bench.DeviceObservations([date], gps, r, i, m, shp)
        data = [obs.get_band_data(date, b) for b in range(B)]
        kf = kb.LinearKalman(obs, None, np.ones(shp, dtype=bool), kb.create_prosail_observation_operator,
                             syn.S2_PARAMETERS, state_propagation=None, prior=None, linear=False, device=device)
        x, _, A, innov = kf.do_all_bands(date, data, DeviceVector.broadcast(syn.SAIL_MEAN, hi - lo, device), None,
                                         BlockDiag.from_block(inv_block, hi - lo, device),
                                         convergence_tolerance=tol, min_iterations=min_iterations)
        return x, A, innov, kf.last_iterations[-1], kf.last_norms[-1]

    x, A, innov, k, norms = solve(0, n, shape)
    assert 2 <= k <= 5 and norms[-1] < 1e-3
    m_win = 300001
    for lo in (0, 7 * 1000 * 1000 + 77, n - m_win):
        xw, Aw, innov_w, kw, _ = solve(lo, lo + m_win, (1, m_win), min_iterations=k, tol=float("inf"))
        assert kw == k
        assert torch.equal(xw.data[:, :m_win], x.data[:, lo:lo + m_win])
        assert torch.equal(Aw.data[:, :m_win], A.data[:, lo:lo + m_win])
        assert torch.equal(innov_w.data[:, :m_win], innov.data[:, lo:lo + m_win])
    # value check of a sample spread over the shard
    sel = np.linspace(0, n - 1, 96).astype(np.int64)
    sel_t = torch.from_numpy(sel).to(device)
    y = rho[:, sel_t].cpu().numpy()
    m = mask[:, sel_t].cpu().numpy().astype(bool)
    w = kc.gate_inverse_variance(inv[:, sel_t].cpu().numpy(), m)
    ogps = [GaussianProcess(g.inputs, g.targets, g.theta) for g in gps]
    ops = [(lambda xp, gp=ogps[b], mm=m[b]: kc.emulator_operator(gp, xp, P, mm)) for b in range(B)]
    xo, Ao, _, _, _ = kc.do_all_bands(ops, list(y), list(w), list(m), np.tile(syn.SAIL_MEAN, len(sel)),
                                      np.repeat(inv_block[None], len(sel), axis=0), P, fixed_iterations=k)
    xs = x.data[:, sel_t].t().cpu().numpy().ravel()
    assert np.max(np.abs(xs - xo) / np.maximum(np.abs(xo), 1e-3)) < TOL
    idx = np.arange(P)
    diag_rows = [r * (r + 1) // 2 + r for r in range(P)]
    ds = A.data[diag_rows][:, sel_t].t().cpu().numpy()
    assert np.max(gu.rel(ds, Ao[:, idx, idx])) < TOL
