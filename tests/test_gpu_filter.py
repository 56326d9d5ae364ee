"""GPU parity of the whole drop-in path: ``LinearKalman.run`` /
``do_all_bands`` (SURVEY section 8 rows a10/a11 and the b boundary) against

* oracle B (fp64 restatement)  -- posterior mean and diag(P^-1) within 1e-6
  relative (the north_star tolerance), iteration counts equal;
* the UNMODIFIED reference's outputs stored in tests/golden -- within the
  reference's own float32-LU reproducibility (see tests/test_oracle_golden.py).
"""
import datetime as dt

import numpy as np
import pytest

import golden_io as gio
import gpu_util as gu
import kafka_inferenceengine_b200 as kb
from kafka_inferenceengine_b200 import synthetic as syn
from kafka_inferenceengine_b200.device import BlockDiag, DeviceVector
from oracle import kafka_cpu as kc
from oracle.gp import GaussianProcess

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["one call per date", "one call per kernel"])
def host_path(request, monkeypatch):
    """Every test of this file runs twice: through the small-image path (kafka_assimilate_date, one
    C-ABI call per date, what images of up to 2^18 pixels take by default) and through the general
    path (one call per kernel, what the large shards of the benchmark take)."""
    import kafka_inferenceengine_b200 as kb_
    monkeypatch.setattr(kb_.LinearKalman, "SMALL_IMAGE_PATH_DEFAULT", request.param == "one call per date")
    return request.param

TOL = 1e-6           # north_star: posterior mean and posterior-variance diagonal, relative


def _iters_per_date(iter_log):
    out = []
    for n, _ in iter_log:
        if int(n) == 1:
            out.append(1)
        else:
            out[-1] = int(n)
    return out


def _oracle_ops(gps, P, bands, mappers=None):
    def make(date):
        return [(lambda xp, gp=gp, m=bands[date][b][2], mp=(None if mappers is None else mappers[b]):
                 kc.emulator_operator(gp, xp, P, m, state_mapper=mp))
                for b, gp in enumerate(gps)]
    return make


def _check_rows(out, rows, P, ref=None, tol=TOL):
    idx = np.arange(P)
    steps = sorted(out.output)
    assert [s.toordinal() for s in steps] == [r[0].toordinal() for r in rows]
    for k, (t, x, Pinv, _) in enumerate(rows):
        got = out.output[steps[k]]
        ex = np.abs(got["x"] - x) / np.maximum(np.abs(x), 1e-3)
        ed = gu.rel(got["diag"], Pinv[:, idx, idx].ravel())
        assert ex.max() < tol, (k, ex.max())
        assert ed.max() < tol, (k, ed.max())
        if ref is not None:                      # unmodified float32 reference
            rx = gu.rel(got["x"], ref["out_x"][k])
            rd = gu.rel(got["diag"], ref["out_diag"][k])
            assert np.median(rx) < 2e-5 and rx.max() < 5e-2
            assert np.median(rd) < 1e-4 and rd.max() < 5e-3
            # those bounds are the reference's own float32-LU error, measured by the fp64 oracle on the
            # same rows: the CUDA path must sit as close to the reference as the oracle does, so that a
            # regression of the kernels cannot hide inside the loose figures above
            bx = gu.rel(x, ref["out_x"][k])
            bd = gu.rel(Pinv[:, idx, idx].ravel(), ref["out_diag"][k])
            assert rx.max() <= 1.05 * bx.max() + 1e-6 and np.median(rx) <= 1.05 * np.median(bx) + 1e-8
            assert rd.max() <= 1.05 * bd.max() + 1e-6 and np.median(rd) <= 1.05 * np.median(bd) + 1e-8


def test_run_s2_prosail_vs_oracle_and_reference():
    """Config-2 shape (10 params, 10 bands, T=250, prior-only advance;
    kafka_test_S2.py:174-194) on the fixture the unmodified reference ran."""
    g = gio.load("run_s2")
    mask = g["mask"]
    P = 10
    N = int(mask.sum())
    gps = gio.emulators(g, syn.EmulatorData)
    obs = gio.FixtureObservations(g, gps)
    prior = syn.BroadcastPrior(g["x0"][:P], g["prior_inv_block"].astype(np.float64), N,
                               lambda blk, n: BlockDiag.from_block(blk, n, gu.DEV))
    out = syn.MemoryOutput(syn.S2_PARAMETERS)
    kf = kb.LinearKalman(obs, out, mask, kb.create_prosail_observation_operator, syn.S2_PARAMETERS,
                         state_propagation=None, prior=prior, linear=False)
    kf.set_trajectory_model()
    kf.set_trajectory_uncertainty(np.zeros(N * P))
    x0, Pinv0 = prior.process_prior(None)
    kf.run(gio.grid(g), x0, None, Pinv0, iter_obs_op=True)

    ogps = gio.emulators(g, GaussianProcess)
    bands = gio.compacted_bands(obs, mask)
    blocks = np.repeat(g["prior_inv_block"][None].astype(np.float64), N, axis=0)
    rows = list(kc.run_filter(gio.grid(g), obs.dates, bands, _oracle_ops(ogps, P, bands), g["x0"], blocks, P,
                              lambda x, A, d: (g["x0"].copy(), blocks.copy())))
    assert kf.last_iterations == sum((r[3] for r in rows), [])
    assert kf.last_iterations == _iters_per_date(g["iter_log"])          # same as the reference's log
    _check_rows(out, rows, P, ref=g)


def test_run_tip_state_mapper_lai_propagator():
    """Config-3 shape: 7-param TIP state, 2 bands through per-band state mappers
    (utils.py:130-177), propagate_information_filter_LAI with Q[6::7]=0.04
    (kafka_test.py:156-217)."""
    g = gio.load("run_tip")
    mask = g["mask"]
    P = 7
    N = int(mask.sum())
    gps = gio.emulators(g, syn.EmulatorData)
    obs = gio.FixtureObservations(g, gps)
    out = syn.MemoryOutput(syn.TIP_PARAMETERS)
    kf = kb.LinearKalman(obs, out, mask, kb.create_nonlinear_observation_operator, syn.TIP_PARAMETERS,
                         state_propagation=kb.propagate_information_filter_LAI, prior=None, linear=False)
    kf.set_trajectory_model()
    kf.set_trajectory_uncertainty(g["q"])
    mean, prior_inv = kb.tip_prior_full({"n_pixels": N})
    kf.run(gio.grid(g), mean, None, prior_inv, iter_obs_op=True)

    ogps = gio.emulators(g, GaussianProcess)
    bands = gio.compacted_bands(obs, mask)
    mappers = [g["mapper_0"], g["mapper_1"]]
    idx = np.arange(P)
    blocks = np.repeat(g["prior_inv_block"][None].astype(np.float64), N, axis=0)
    rows = list(kc.run_filter(gio.grid(g), obs.dates, bands, _oracle_ops(ogps, P, bands, mappers), g["x0"], blocks, P,
                              lambda x, A, d: kc.propagate_information_filter_lai(x, A[:, idx, idx].ravel(), g["q"])))
    assert kf.last_iterations == sum((r[3] for r in rows), []) == _iters_per_date(g["iter_log"])
    _check_rows(out, rows, P)
    steps = sorted(out.output)
    for k in range(len(steps)):                     # unmodified reference, float32
        assert np.allclose(out.output[steps[k]]["x"], g["out_x"][k], rtol=1e-4, atol=1e-6)
        assert np.allclose(out.output[steps[k]]["diag"], g["out_diag"][k], rtol=5e-6)


def test_run_linear_config1():
    """Config-1 shape: TIP state, ONE linear band in (H0,H) tuple form observing
    slot 6, 10 dates, LAI propagator."""
    g = gio.load("run_linear")
    mask = g["mask"]
    P = 7
    N = int(mask.sum())
    op = kb.LinearBandOperator(g["row"], 0.0)
    obs = gio.FixtureObservations(g, [op])
    out = syn.MemoryOutput(syn.TIP_PARAMETERS)
    kf = kb.LinearKalman(obs, out, mask, kb.create_linear_observation_operator, syn.TIP_PARAMETERS,
                         state_propagation=kb.propagate_information_filter_LAI, prior=None)
    kf.set_trajectory_model()
    kf.set_trajectory_uncertainty(g["q"])
    mean, prior_inv = kb.tip_prior_full({"n_pixels": N})
    kf.run(gio.grid(g), mean, None, prior_inv)
    assert kf.last_iterations == _iters_per_date(g["iter_log"])
    steps = sorted(out.output)
    assert [s.toordinal() for s in steps] == list(g["out_steps"])
    for k in range(len(steps)):
        assert np.allclose(out.output[steps[k]]["diag"], g["out_diag"][k], rtol=5e-6)
        assert np.allclose(out.output[steps[k]]["x"], g["out_x"][k], rtol=1e-4, atol=1e-6)
    bands = gio.compacted_bands(obs, mask)
    idx = np.arange(P)

    def make(date):
        m = bands[date][0][2]

        def lin(xp):
            J = np.where(m[:, None], op.row[None, :], 0.0)
            H0 = np.where(m, xp.reshape(N, P) @ op.row + op.offset, 0.0)
            return H0, J
        return [lin]
    blocks = np.repeat(g["prior_inv_block"][None].astype(np.float64), N, axis=0)
    rows = list(kc.run_filter(gio.grid(g), obs.dates, bands, make, g["x0"], blocks, P,
                              lambda x, A, d: kc.propagate_information_filter_lai(x, A[:, idx, idx].ravel(), g["q"])))
    _check_rows(out, rows, P)


def _barrax_s2_case(n_dates, seed=1234):
    rng = np.random.default_rng(seed)
    mask = gio.barrax_mask()
    P, B, T = 10, 10, 250
    N = int(mask.sum())
    gps = [syn.make_emulator(rng, syn.SAIL_MEAN, syn.SAIL_SIGMA, n_train=T, cls=GaussianProcess) for _ in range(B)]
    x_true = syn.SAIL_MEAN + syn.SAIL_SIGMA * rng.standard_normal((N, P))
    base = dt.datetime(2017, 7, 3)
    dates = [base + dt.timedelta(days=2 * d) for d in range(n_dates)]
    obs = syn.SyntheticObservations(mask, dates, gps, x_true, lambda gp, X: gp.predict(X, do_unc=False)[0], rng)
    grid = [base + dt.timedelta(days=2 * d) for d in range(n_dates + 1)]
    return mask, P, N, gps, obs, grid


def test_barrax_config2_two_dates_vs_oracle():
    """BASELINE config 2 (Barrax mask, 13 027 px, 10 bands, T=250) for two dates."""
    mask, P, N, gps, obs, grid = _barrax_s2_case(2)
    inv = np.diag(1.0 / syn.SAIL_SIGMA ** 2).astype(np.float32).astype(np.float64)
    prior = syn.BroadcastPrior(syn.SAIL_MEAN, inv, N, lambda blk, n: BlockDiag.from_block(blk, n, gu.DEV))
    out = syn.MemoryOutput(syn.S2_PARAMETERS)
    kf = kb.LinearKalman(obs, out, mask, kb.create_prosail_observation_operator, syn.S2_PARAMETERS,
                         state_propagation=None, prior=prior, linear=False)
    kf.set_trajectory_model()
    kf.set_trajectory_uncertainty(np.zeros(N * P))
    x0, Pinv0 = prior.process_prior(None)
    kf.run(grid, x0, None, Pinv0, iter_obs_op=True)
    bands = gio.compacted_bands(obs, mask)
    blocks = np.repeat(inv[None], N, axis=0)
    x0n = np.tile(syn.SAIL_MEAN, N)
    rows = list(kc.run_filter(grid, obs.dates, bands, _oracle_ops(gps, P, bands), x0n, blocks, P,
                              lambda x, A, d: (x0n.copy(), blocks.copy())))
    assert kf.last_iterations == sum((r[3] for r in rows), [])
    _check_rows(out, rows, P)
    # output stage: what KafkaOutput.dump_data would write (observations.py:374-375,392-393)
    last = sorted(out.output)[-1]
    xv = DeviceVector.from_interleaved(out.output[last]["x"], P, gu.DEV)
    A = BlockDiag.from_blocks(rows[-1][2], gu.DEV)
    mean_r, unc_r = kf.pixels.scatter_output(xv, A, P)
    assert np.array_equal(mean_r[6].cpu().numpy(), kc.scatter_param(out.output[last]["x"], 6, P, mask))


def test_do_all_bands_fixed_iteration_counts_and_blend():
    """Parity at FIXED iteration counts (1 and 3 solves) and with prior AND
    propagator both set (blend_prior path, kf_tools.py:161-164)."""
    g = gio.load("run_tip")
    mask = g["mask"]
    P = 7
    N = int(mask.sum())
    gps = gio.emulators(g, syn.EmulatorData)
    ogps = gio.emulators(g, GaussianProcess)
    obs = gio.FixtureObservations(g, gps)
    mappers = [g["mapper_0"], g["mapper_1"]]
    bands = gio.compacted_bands(obs, mask)
    date = obs.dates[0]
    blocks = np.repeat(g["prior_inv_block"][None].astype(np.float64), N, axis=0)
    kf = kb.LinearKalman(obs, syn.MemoryOutput(syn.TIP_PARAMETERS), mask, kb.create_nonlinear_observation_operator,
                         syn.TIP_PARAMETERS, prior=None)
    data = [obs.get_band_data(date, b) for b in range(2)]
    for k in (1, 3):
        # min_iterations=k with a huge tolerance stops after exactly k solves
        x, _, A, innov = kf.do_all_bands(date, data, g["x0"], None, BlockDiag.from_blocks(blocks, gu.DEV),
                                         convergence_tolerance=1e9, min_iterations=k)
        xo, Ao, innov_o, n, norms = kc.do_all_bands(_oracle_ops(ogps, P, bands, mappers)(date),
                                                    [b[0] for b in bands[date]], [b[1] for b in bands[date]],
                                                    [b[2] for b in bands[date]], g["x0"], blocks, P,
                                                    fixed_iterations=k)
        assert kf.last_iterations[-1] == k == n
        assert np.max(np.abs(np.asarray(x) - xo) / np.maximum(np.abs(xo), 1e-3)) < TOL
        assert np.max(gu.rel(A.blocks(), Ao, floor=1e-3)) < TOL
        assert np.allclose(np.asarray(innov), innov_o, rtol=1e-6, atol=1e-9)
        assert np.allclose(kf.last_norms[-1], norms, rtol=1e-6)
    # advance with both a propagator and a prior: blend
    mean, _, inv = kb.tip_prior()
    prior = syn.BroadcastPrior(mean, inv.astype(np.float64), N, lambda blk, n: BlockDiag.from_block(blk, n, gu.DEV))
    kf2 = kb.LinearKalman(obs, syn.MemoryOutput(syn.TIP_PARAMETERS), mask, kb.create_nonlinear_observation_operator,
                          syn.TIP_PARAMETERS, state_propagation=kb.propagate_information_filter_LAI, prior=prior)
    kf2.set_trajectory_model()
    kf2.set_trajectory_uncertainty(g["q"])
    kf2.current_timestep = date
    xa = np.asarray(x)
    xf, pf, pfi = kf2.advance(x, None, A, kf2.trajectory_model, kf2.trajectory_uncertainty)
    idx = np.arange(P)
    xl, bl = kc.propagate_information_filter_lai(xa, Ao[:, idx, idx].ravel(), g["q"])
    xb, Cb = kc.blend_prior(np.tile(mean, N), np.repeat(inv[None].astype(np.float64), N, axis=0), xl, bl)
    assert pf is None
    assert np.max(np.abs(np.asarray(xf) - xb) / np.maximum(np.abs(xb), 1e-3)) < TOL
    assert np.max(gu.rel(pfi.blocks(), Cb, floor=1e-3)) < TOL


def test_unsupported_operator_raises():
    mask = np.ones((4, 4), dtype=bool)
    with pytest.raises(NotImplementedError):
        kb.LinearKalman(None, None, mask, lambda *a: None, syn.S2_PARAMETERS)
    kf = kb.LinearKalman(None, None, mask, kb.create_prosail_observation_operator, syn.S2_PARAMETERS)
    with pytest.raises(ValueError):
        kf.do_all_bands(None, [], np.zeros(3 * 10), None, np.eye(10))


def test_uniform_first_iteration_shortcut_is_value_neutral():
    """When every pixel is linearised at the same point (first iteration after a
    prior reset) the emulators are evaluated once, like the reference's
    single-unique-vector branch (utils.py:72-74): same results and iteration
    counts as evaluating every pixel."""
    mask, P, N, gps, obs, grid = _barrax_s2_case(1, seed=77)
    inv = np.diag(1.0 / syn.SAIL_SIGMA ** 2).astype(np.float32).astype(np.float64)
    res = []
    for dedup in (True, False):
        kf = kb.LinearKalman(obs, syn.MemoryOutput(syn.S2_PARAMETERS), mask, kb.create_prosail_observation_operator,
                             syn.S2_PARAMETERS, state_propagation=None, prior=None, dedup_uniform_state=dedup)
        data = [obs.get_band_data(obs.dates[0], b) for b in range(10)]
        for x0 in (DeviceVector.broadcast(syn.SAIL_MEAN, N, gu.DEV), np.tile(syn.SAIL_MEAN, N)):
            x, _, A, innov = kf.do_all_bands(obs.dates[0], data, x0, None, BlockDiag.from_block(inv, N, gu.DEV))
            res.append((np.asarray(x), A.blocks(), np.asarray(innov), kf.last_iterations[-1], kf.last_norms[-1]))
    ref = res[-1]                                   # dedup off, host input
    for r in res[:-1]:
        assert r[3] == ref[3]
        assert np.max(np.abs(r[0] - ref[0]) / np.maximum(np.abs(ref[0]), 1e-3)) < 1e-12
        assert np.max(gu.rel(r[1], ref[1], floor=1e-3)) < 1e-12
        assert np.allclose(r[2], ref[2], rtol=1e-12, atol=1e-15)
        assert np.allclose(r[4], ref[4], rtol=1e-9)
    # and the two host paths against each other: one C-ABI call per date (kafka_assimilate_date, no
    # shared first linearisation) vs one call per kernel
    both = []
    for small in (True, False):
        kf = kb.LinearKalman(obs, syn.MemoryOutput(syn.S2_PARAMETERS), mask, kb.create_prosail_observation_operator,
                             syn.S2_PARAMETERS, state_propagation=None, prior=None, small_image_path=small)
        kf.want_forward_modelled = True
        data = [obs.get_band_data(obs.dates[0], b) for b in range(10)]
        x, _, A, innov = kf.do_all_bands(obs.dates[0], data, np.tile(syn.SAIL_MEAN, N), None,
                                         BlockDiag.from_block(inv, N, gu.DEV))
        both.append((np.asarray(x), A.blocks(), np.asarray(innov), kf.last_iterations[-1], kf.last_norms[-1],
                     np.asarray(kf.fwd_modelled)))
    assert both[0][3] == both[1][3] == ref[3]
    assert np.max(np.abs(both[0][0] - both[1][0]) / np.maximum(np.abs(both[1][0]), 1e-3)) < 1e-12
    assert np.max(gu.rel(both[0][1], both[1][1], floor=1e-3)) < 1e-12
    assert np.allclose(both[0][2], both[1][2], rtol=1e-12, atol=1e-15)
    assert np.allclose(both[0][4], both[1][4], rtol=1e-9)
    assert np.allclose(both[0][5], both[1][5], rtol=1e-10, atol=1e-13)
    # the detector itself
    v = DeviceVector.from_interleaved(np.tile(syn.SAIL_MEAN, 5000), P, gu.DEV)
    assert np.array_equal(v.uniform_value(), syn.SAIL_MEAN)
    x = np.tile(syn.SAIL_MEAN, 5000)
    x[P * 4999 + 3] = np.nextafter(x[P * 4999 + 3], 1.0)
    assert DeviceVector.from_interleaved(x, P, gu.DEV).uniform_value() is None


def test_raster_output_stage_matches_reference_arithmetic():
    """RasterOutput.dump_data = the arithmetic of KafkaOutput.dump_data
    (input_output/observations.py:374-375,392-393): float32 rasters of the mean
    and of 1/sqrt(diag(P^-1)), zero outside the state mask."""
    from kafka_inferenceengine_b200.output import RasterOutput
    rng = np.random.default_rng(3)
    mask = gio.barrax_mask()
    N, P = int(mask.sum()), 10
    x = rng.standard_normal(N * P)
    blocks = gu.random_spd(rng, N, P)
    got = {}
    out = RasterOutput(syn.S2_PARAMETERS, sink=lambda t, names, mean, unc: got.update(mean=mean.copy(), unc=unc.copy()))
    out.dump_data(dt.datetime(2017, 7, 3), x, None, kc.blocks_to_sparse(blocks), mask, P)
    assert not got                                   # overlapped: delivered at the next dump_data / flush
    out.flush()
    diag = blocks[:, np.arange(P), np.arange(P)].ravel()
    for ii in range(P):
        assert np.array_equal(got["mean"][ii], kc.scatter_param(x, ii, P, mask))
        assert np.allclose(got["unc"][ii], kc.scatter_unc(diag, ii, P, mask), rtol=2e-7, atol=0)
    assert out.d2h_bytes == 2 * P * mask.size * 4
    # and with the device objects LinearKalman hands over
    out.dump_data(dt.datetime(2017, 7, 4), DeviceVector.from_interleaved(x, P, gu.DEV), None,
                  BlockDiag.from_blocks(blocks, gu.DEV), mask, P)
    out.flush()
    assert np.array_equal(out.last[1], got["mean"])
    sync = RasterOutput(syn.S2_PARAMETERS, overlap=False)
    sync.dump_data(dt.datetime(2017, 7, 4), x, None, kc.blocks_to_sparse(blocks), mask, P)
    assert np.array_equal(sync.last[1], got["mean"]) and np.array_equal(sync.last[2], got["unc"])


def test_raster_output_stage_matches_the_unmodified_writer():
    """RasterOutput against the arrays the unmodified KafkaOutput.dump_data hands to
    GDAL's WriteArray (fixture kat_output_stage): float32 rasters, bit-exact, from host
    inputs and from the device objects LinearKalman passes."""
    from kafka_inferenceengine_b200.output import RasterOutput
    g = gio.load("kat_output_stage")
    mask, x, blocks = g["mask"], g["x"], g["blocks"]
    P = blocks.shape[1]
    for args in ((x, kc.blocks_to_sparse(blocks)),
                 (DeviceVector.from_interleaved(x, P, gu.DEV), BlockDiag.from_blocks(blocks, gu.DEV))):
        out = RasterOutput(["lai", "cab", "sm"], overlap=False)
        out.dump_data(dt.datetime(2017, 7, 3), args[0], None, args[1], mask, P)
        for ii in range(P):
            assert out.last[1][ii].dtype == np.float32
            assert np.array_equal(out.last[1][ii], g["mean_%d" % ii])
            assert np.array_equal(out.last[2][ii], g["unc_%d" % ii])


def test_run_with_prefetch_and_overlapped_output_matches_plain_run():
    """run() stages the next observation date while the current one is being
    solved and RasterOutput reads results back on a side stream: same rasters as
    the plain path, for a multi-date run with two dates in one grid interval."""
    from kafka_inferenceengine_b200.output import RasterOutput
    g = gio.load("run_s2")
    mask = g["mask"]
    P = 10
    N = int(mask.sum())
    gps = gio.emulators(g, syn.EmulatorData)
    obs = gio.FixtureObservations(g, gps)
    res = {}
    for mode in (True, False):
        prior = syn.BroadcastPrior(g["x0"][:P], g["prior_inv_block"].astype(np.float64), N,
                                   lambda blk, n: BlockDiag.from_block(blk, n, gu.DEV),
                                   (lambda m, n: DeviceVector.broadcast(m, n, gu.DEV)) if mode else None)
        got = []
        out = RasterOutput(syn.S2_PARAMETERS, overlap=mode,
                           sink=lambda t, names, mean, unc: got.append((t, mean.copy(), unc.copy())))
        kf = kb.LinearKalman(obs, out, mask, kb.create_prosail_observation_operator, syn.S2_PARAMETERS,
                             state_propagation=None, prior=prior, linear=False, prefetch_observations=mode)
        kf.set_trajectory_model()
        kf.set_trajectory_uncertainty(np.zeros(N * P))
        x0, Pinv0 = prior.process_prior(None)
        kf.run(gio.grid(g), x0, None, Pinv0, iter_obs_op=True)
        res[mode] = (got, list(kf.last_iterations))
    assert res[True][1] == res[False][1] == _iters_per_date(g["iter_log"])
    assert len(res[True][0]) == len(res[False][0]) == len(g["out_steps"])
    for (t1, m1, u1), (t2, m2, u2) in zip(res[True][0], res[False][0]):
        assert t1 == t2 and np.array_equal(m1, m2) and np.array_equal(u1, u2)
    # and equal to the reference-checked interleaved output
    last_x = res[True][0][-1][1]
    assert np.allclose(last_x[6][mask], g["out_x"][-1][6::P], rtol=2e-3, atol=1e-5)


def test_edge_cases_empty_mask_all_unobserved_and_maximum_sizes():
    """Empty state mask; a date where no pixel is observed; and the largest
    supported problem (16 parameters, 16 bands, 16-input emulators)."""
    # --- empty state mask: nothing to do, nothing breaks
    mask = np.zeros((5, 6), dtype=bool)
    kf = kb.LinearKalman(None, None, mask, kb.create_linear_observation_operator, syn.TIP_PARAMETERS, prior=None)
    assert kf.n_state_elems == 0
    band = syn.BandData(np.zeros(mask.shape), np.ones(mask.shape), np.ones(mask.shape, dtype=bool), None,
                        kb.LinearBandOperator(np.eye(7)[6]))
    x, _, A, innov = kf.do_all_bands(None, [band], np.zeros(0), None, BlockDiag.from_blocks(np.zeros((0, 7, 7)), gu.DEV))
    assert len(x) == 0 and np.asarray(x).size == 0 and kf.last_iterations[-1] == 2
    # --- nobody observed: analysis == forecast, A == P_f^-1, two iterations (min_iterations)
    rng = np.random.default_rng(2)
    mask = rng.random((9, 13)) < 0.7
    N = int(mask.sum())
    P = 7
    _, _, inv = kb.tip_prior()
    kf = kb.LinearKalman(None, None, mask, kb.create_linear_observation_operator, syn.TIP_PARAMETERS, prior=None)
    none_obs = np.zeros(mask.shape, dtype=bool)
    band = syn.BandData(np.zeros(mask.shape), syn.inverse_variance_matrix(np.full(mask.shape, np.inf)), none_obs, None,
                        kb.LinearBandOperator(np.eye(7)[6]))
    x0 = np.tile(syn.TIP_MEAN, N) + 0.01 * rng.standard_normal(N * P)
    x, _, A, innov = kf.do_all_bands(None, [band], x0, None, BlockDiag.from_block(inv.astype(np.float64), N, gu.DEV))
    assert np.allclose(np.asarray(x), x0, rtol=1e-12) and kf.last_iterations[-1] == 2
    assert np.array_equal(A.blocks()[3], inv.astype(np.float64)) and not np.asarray(innov).any()
    # --- maximum sizes
    P = B = D = 16
    mean = rng.uniform(0.5, 1.5, P)
    sigma = rng.uniform(0.05, 0.2, P)
    gps = [syn.make_emulator(rng, mean, sigma, n_train=40, cls=GaussianProcess) for _ in range(B)]
    N = int(mask.sum())
    x_true = mean + sigma * rng.standard_normal((N, P))
    date = dt.datetime(2017, 7, 3)
    obs = syn.SyntheticObservations(mask, [date], gps, x_true, lambda gp, X: gp.predict(X, do_unc=False)[0], rng)
    names = ["p%d" % k for k in range(P)]
    kf = kb.LinearKalman(obs, None, mask, kb.create_prosail_observation_operator, names, prior=None)
    blocks = np.repeat(np.diag(1.0 / sigma ** 2)[None], N, axis=0)
    data = [obs.get_band_data(date, b) for b in range(B)]
    x, _, A, _ = kf.do_all_bands(date, data, np.tile(mean, N), None, BlockDiag.from_blocks(blocks, gu.DEV))
    bands = gio.compacted_bands(obs, mask)[date]
    ops = [(lambda xp, gp=gp, m=bands[b][2]: kc.emulator_operator(gp, xp, P, m)) for b, gp in enumerate(gps)]
    xo, Ao, _, n, _ = kc.do_all_bands(ops, [b_[0] for b_ in bands], [b_[1] for b_ in bands], [b_[2] for b_ in bands],
                                      np.tile(mean, N), blocks, P)
    assert kf.last_iterations[-1] == n
    assert np.max(np.abs(np.asarray(x) - xo) / np.maximum(np.abs(xo), 1e-3)) < TOL
    idx = np.arange(P)
    assert np.max(gu.rel(A.blocks()[:, idx, idx], Ao[:, idx, idx])) < TOL            # posterior precision diagonal
    assert np.max(np.abs(A.blocks() - Ao)) < TOL * np.abs(Ao).max()                  # whole block, on its own scale


def test_legacy_band_by_band_assimilate():
    """assimilate (linear_kf.py:325-354): bands one after the other, each band's
    analysis the next band's forecast, Hessian correction applied per band."""
    g = gio.load("run_tip")
    mask = g["mask"]
    P = 7
    N = int(mask.sum())
    gps = gio.emulators(g, syn.EmulatorData)
    ogps = gio.emulators(g, GaussianProcess)
    obs = gio.FixtureObservations(g, gps)
    date = obs.dates[0]
    bands = gio.compacted_bands(obs, mask)[date]
    mappers = [g["mapper_0"], g["mapper_1"]]
    blocks = np.repeat(g["prior_inv_block"][None].astype(np.float64), N, axis=0)
    kf = kb.LinearKalman(obs, None, mask, kb.create_nonlinear_observation_operator, syn.TIP_PARAMETERS, prior=None)
    x, _, A = kf.assimilate([date], g["x0"], None, BlockDiag.from_blocks(blocks, gu.DEV))
    assert kf.previous_state.timestamp == date
    xo, Ao = g["x0"], blocks
    for b in range(2):
        op = [lambda xp, b=b: kc.emulator_operator(ogps[b], xp, P, bands[b][2], state_mapper=mappers[b])]
        xn, An, innov, n, _ = kc.do_all_bands(op, [bands[b][0]], [bands[b][1]], [bands[b][2]], xo, Ao, P, min_iterations=1)
        corr = kc.hessian_correction(ogps[b], xn, bands[b][1], innov, bands[b][2], P, mappers[b])
        xo, Ao = xn, An - corr
    assert np.max(np.abs(np.asarray(x) - xo) / np.maximum(np.abs(xo), 1e-3)) < TOL
    assert np.max(np.abs(A.blocks() - Ao)) < TOL * np.abs(Ao).max()


def test_indefinite_blocks_through_two_dates_with_hessian_correction(caplog):
    """What happens to a pixel whose information matrix is not positive definite, pinned
    over two dates of run().  With the second-order correction switched on
    (linear_kf.py:313-319, kf_tools.py:26-72) outlying observations make ``A - correction``
    indefinite for some pixels at date 1; the identity propagator hands those blocks to date
    2 as P_forecast_inverse.  The reference's LU (solvers.py:133-134) returns the solution of
    such a system; here the Cholesky factorisation fails and the pixel is solved by elimination
    with partial pivoting instead, in every iteration; its A is formed and stored as the
    equations say, the count is logged, and all other pixels are unaffected.  The expectation
    below is the fp64 oracle, whose dense LU (numpy.linalg.solve) follows the same rule."""
    import logging
    g = gio.load("run_tip")
    mask = g["mask"]
    P = 7
    N = int(mask.sum())
    gps = gio.emulators(g, syn.EmulatorData)
    ogps = gio.emulators(g, GaussianProcess)
    base_obs = gio.FixtureObservations(g, gps)
    dates = base_obs.dates[:2]
    sel = np.arange(N) % 5 == 0                       # every fifth state pixel sees outlying reflectances
    sel_raster = np.zeros(mask.shape, dtype=bool)
    sel_raster[mask] = sel

    class Outliers(object):
        def __init__(self):
            self.dates = dates
            self.bands_per_observation = {d: 2 for d in dates}

        def get_band_data(self, date, band):
            bd = base_obs.get_band_data(date, band)
            return bd._replace(observations=np.where(sel_raster, 1.8 * bd.observations, bd.observations))
    obs = Outliers()
    out = syn.MemoryOutput(syn.TIP_PARAMETERS)
    kf = kb.LinearKalman(obs, out, mask, kb.create_nonlinear_observation_operator, syn.TIP_PARAMETERS,
                         state_propagation=lambda x, Pm, Pinv, M, Q: (x, None, Pinv), prior=None,
                         hessian_correction=True, round_operator_f32=False)
    kf.set_trajectory_model()
    kf.set_trajectory_uncertainty(np.zeros(N * P))
    grid = [dates[0] - dt.timedelta(days=1), dates[0] + dt.timedelta(hours=12), dates[1] + dt.timedelta(hours=12)]
    blocks = np.repeat(g["prior_inv_block"][None].astype(np.float64), N, axis=0)
    with caplog.at_level(logging.WARNING):
        kf.run(grid, g["x0"], None, BlockDiag.from_blocks(blocks, gu.DEV), iter_obs_op=True)

    # ---- fp64 oracle with the documented rule
    mappers = [g["mapper_0"], g["mapper_1"]]
    bands = gio.compacted_bands(obs, mask)
    x_f, Pinv = g["x0"].astype(np.float64), blocks
    want = []
    for d in dates:
        ys, ws, ms = [b[0] for b in bands[d]], [b[1] for b in bands[d]], [b[2] for b in bands[d]]
        ops = [(lambda xp, gp=ogps[b], m=ms[b], mp=mappers[b]: kc.emulator_operator(gp, xp, P, m, state_mapper=mp,
                                                                                   round_f32=False)) for b in range(2)]
        x_prev, n_iter = x_f.copy(), 1
        bad_hist, norm_hist = [], []
        while True:
            H0s, Js = zip(*[op(x_prev) for op in ops])
            x, A, innov, _ = kc.solve_dense64(ys, ws, H0s, Js, x_prev, x_f, Pinv)
            ev = np.linalg.eigvalsh(A)
            assert np.min(np.abs(ev).min(axis=1) / np.abs(ev).max(axis=1)) > 1e-9      # no marginal block
            bad = ev.min(axis=1) <= 0
            bad_hist.append(int(bad.sum()))
            norm = np.linalg.norm(x - x_prev) / len(x)
            norm_hist.append(norm)
            stop = (norm < 1e-3 and n_iter >= 2) or n_iter > 25
            x_prev = x
            n_iter += 1
            if stop:
                break
        innov = innov.reshape(2, N)
        corr = sum(kc.hessian_correction(ogps[b], x, ws[b], innov[b], ms[b], P, mappers[b]) for b in range(2))
        want.append((x, A - corr, n_iter - 1, bad, bad_hist, norm_hist))
        x_f, Pinv = x, A - corr
    n_indef = int((np.linalg.eigvalsh(want[0][1]).min(axis=1) <= 0).sum())
    assert 0 < n_indef < N and want[1][4][0] == n_indef and not want[0][3].any()

    assert kf.last_iterations == [w[2] for w in want]
    steps = sorted(out.output)
    idx = np.arange(P)
    # Date 1: every block is positive definite, all pixels agree.  Date 2: a pixel whose forecast block is
    # indefinite is relinearised 26 times around the solutions of indefinite systems -- a chaotic map, in which
    # the last bit of one solve decides the state ten iterations later (the reference run twice with two BLAS
    # builds would not agree with itself either).  What can be pinned: the first solves of date 2, through the
    # convergence norm that sums over ALL pixels (same x_prev on both sides, the elimination's result inside);
    # and the pixels whose blocks stay positive definite, which no other pixel can touch.
    ever = np.zeros(N, dtype=bool)
    ever[np.arange(N) % 5 == 0] = True
    for k, (x, A2, _, bad, hist, norms) in enumerate(want):
        got = out.output[steps[k]]
        ex = (np.abs(got["x"] - x) / np.maximum(np.abs(x), 1e-3)).reshape(N, P)
        sel = np.ones(N, dtype=bool) if k == 0 else ~ever
        # 26 relinearisations of an oscillating problem amplify rounding ~50x: 1e-5 instead of 1e-6
        assert ex[sel].max() < 1e-5, (k, ex[sel].max())
        d_want = A2[:, idx, idx]
        d_got = got["diag"].reshape(N, P)
        assert np.max(np.abs(d_got - d_want)[sel]) < 1e-5 * np.abs(d_want[sel]).max()
        assert np.isfinite(got["x"]).all() and np.isfinite(got["diag"]).all()
        assert kf.last_norms[k][0] == pytest.approx(norms[0], rel=1e-9)
        assert kf.last_norms[k][1] == pytest.approx(norms[1], rel=1e-6)
    assert want[1][4][0] == n_indef and set(np.flatnonzero(want[1][3])) <= set(np.flatnonzero(ever))
    # no pixel of date 2 was left at its forecast (= the analysis of date 1)
    x1 = out.output[steps[0]]["x"].reshape(N, P)
    x2 = out.output[steps[1]]["x"].reshape(N, P)
    assert not np.all(x1 == x2, axis=1).any()
    msgs = [r.getMessage() for r in caplog.records if "not positive definite" in r.getMessage()]
    assert msgs and "partial pivoting" in msgs[-1]
    counts = [int(v) for v in __import__("re").findall(r"\d+", msgs[-1])]
    assert max(counts) >= n_indef                       # the first solve of date 2 met all of them
    assert not [r for r in caplog.records if "left at the forecast" in r.getMessage()]
