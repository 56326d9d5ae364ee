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
