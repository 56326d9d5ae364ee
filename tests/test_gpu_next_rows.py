"""GPU parity for the rows SURVEY section 8f marks "next": the Water-Cloud SAR
operator, the Hessian (second-order) correction of the posterior precision, and
the observation ingest fused with the compaction gather."""
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


# ---------------------------------------------------------------- SAR / WCM
def test_sar_observation_operator_vs_reference_and_oracle():
    g = gio.load("kat_sar")
    n = g["x"].shape[0]
    for pol in ("VV", "VH"):
        s0, grad = kb.sar_observation_operator(g["x"], np.full(n, 23.0), pol)      # the reference's own call
        assert np.allclose(s0, g["sigma0_" + pol], rtol=1e-12)
        assert np.allclose(grad, g["grad_" + pol], rtol=1e-11, atol=1e-17)
    with pytest.raises(ValueError):
        kb.sar_observation_operator(np.array([[-1.0, 0.2]]), 23.0, "VV")          # "Negative LAI!"
    with pytest.raises(ValueError):
        kb.sar_observation_operator(np.array([[1.0, 0.2]]), 23.0, "HH")


def _sar_case(rng, n_dates=3):
    mask = rng.random((30, 41)) < 0.6
    N = int(mask.sum())
    x_true = np.stack([rng.uniform(1.5, 4.0, N), rng.uniform(0.15, 0.35, N)], axis=1)
    base = dt.datetime(2017, 5, 1)
    dates = [base + dt.timedelta(days=3 * d) for d in range(n_dates)]

    class Obs(object):
        pass
    obs = Obs()
    obs.dates = dates
    obs.bands_per_observation = {d: 2 for d in dates}
    store = {}
    for d in dates:
        bands = []
        for b, pol in enumerate(("VV", "VH")):
            s0, _ = kc.sar_forward(x_true, 23.0, pol)
            sig = 0.05 * s0
            y = np.zeros(mask.shape)
            y[mask] = s0 + sig * rng.standard_normal(N)
            om = rng.random(mask.shape) < 0.85
            iv = np.full(mask.shape, np.inf)
            iv[mask] = 1.0 / sig ** 2
            iv = np.where(om, iv, np.inf)
            bands.append(syn.BandData(np.where(om, y, 0.0), syn.inverse_variance_matrix(iv), om, {"band": b}, None))
        store[d] = bands
    obs.get_band_data = lambda date, band: store[date][band]
    return mask, N, obs, dates, base


def test_sar_filter_run_vs_oracle():
    """LinearKalman with create_sar_observation_operator (2-parameter state LAI, SM;
    VV + VH; sar_forward_model.py:116-173), prior-only advance, against oracle B."""
    rng = np.random.default_rng(17)
    mask, N, obs, dates, base = _sar_case(rng)
    P = 2
    mean = np.array([2.5, 0.25])
    inv = np.diag(1.0 / np.array([0.8, 0.08]) ** 2)       # keeps every iterate inside the model's domain
    prior = syn.BroadcastPrior(mean, inv, N, lambda blk, n: BlockDiag.from_block(blk, n, gu.DEV))
    out = syn.MemoryOutput(["lai", "sm"])
    kf = kb.LinearKalman(obs, out, mask, kb.create_sar_observation_operator, ["lai", "sm"],
                         state_propagation=None, prior=prior, linear=False)
    kf.set_trajectory_model()
    kf.set_trajectory_uncertainty(np.zeros(N * P))
    x0, Pinv0 = prior.process_prior(None)
    grid = [base + dt.timedelta(days=3 * d) for d in range(len(dates) + 1)]
    kf.run(grid, x0, None, Pinv0)
    bands = gio.compacted_bands(obs, mask)
    blocks = np.repeat(inv[None], N, axis=0)
    x0n = np.tile(mean, N)

    def ops(date):
        return [(lambda xp, b=b, m=bands[date][b][2]: kc.sar_operator(xp, P, m, b)) for b in range(2)]
    rows = list(kc.run_filter(grid, obs.dates, bands, ops, x0n, blocks, P, lambda x, A, d: (x0n.copy(), blocks.copy())))
    assert kf.last_iterations == sum((r[3] for r in rows), [])
    steps = sorted(out.output)
    for k, (t, x, Pinv, _) in enumerate(rows):
        got = out.output[steps[k]]
        assert np.max(np.abs(got["x"] - x) / np.maximum(np.abs(x), 1e-3)) < 1e-6
        assert np.max(gu.rel(got["diag"], Pinv[:, [0, 1], [0, 1]].ravel())) < 1e-6
    # the builder call of the reference's signature gives the same rows as the oracle
    d0 = dates[0]
    H0, H = kb.create_sar_observation_operator(P, None, None, obs.get_band_data(d0, 1).mask, mask, x0n, 1)
    H0o, Jo = kc.sar_operator(x0n, P, bands[d0][1][2], 1)
    assert np.array_equal(H0.astype(np.float64), H0o)
    Hc = H.tocoo()
    J = np.zeros((N, P))
    J[Hc.row, Hc.col - P * Hc.row] = Hc.data
    assert np.array_equal(J, Jo)


def test_sar_domain_violation_raises_like_reference():
    rng = np.random.default_rng(18)
    mask, N, obs, dates, base = _sar_case(rng, 1)
    kf = kb.LinearKalman(obs, syn.MemoryOutput(["lai", "sm"]), mask, kb.create_sar_observation_operator,
                         ["lai", "sm"], state_propagation=None, prior=None)
    x0 = np.tile([2.0, 0.25], N)
    data = [obs.get_band_data(dates[0], b) for b in range(2)]
    first_observed = int(np.flatnonzero(data[0].mask[mask])[0])
    x0[2 * first_observed] = -0.5                                  # negative LAI in an observed pixel
    with pytest.raises(ValueError):
        kf.do_all_bands(dates[0], data, x0, None, BlockDiag.from_block(np.eye(2) * 4.0, N, gu.DEV))


# ------------------------------------------------------ Hessian correction
def test_hessian_correction_vs_reference_fixture():
    g = gio.load("kat_hessian")
    mask = g["mask"]
    gps = gio.emulators(g, syn.EmulatorData)
    Rs, innovs, masks = [], [], []
    for b in range(2):
        R = syn.inverse_variance_matrix(g["inv_var_%d" % b])
        corr = kb.hessian_correction(gps[b], g["x0"], R, g["innov_%d" % b], g["obs_mask_%d" % b], mask, b, 7)
        ref = g["corr_%d" % b]
        assert np.max(np.abs(corr.blocks() - ref)) < 1e-8 * np.abs(ref).max()
        Rs.append(R)
        innovs.append(g["innov_%d" % b])
        masks.append(g["obs_mask_%d" % b])
    multi = kb.hessian_correction_multiband(gps[0], g["x0"], Rs, innovs, masks, mask, 2, 7)
    ref = g["corr_multi_gp0"]
    assert np.max(np.abs(multi.blocks() - ref)) < 1e-8 * np.abs(ref).max()
    one = kb.hessian_correction_pixel(gps[0], g["x0"][:7], 400.0, 0.03, 0, 7)
    ogp = gio.emulators(g, GaussianProcess)[0]
    want = kc.hessian_correction(ogp, g["x0"][:7], np.array([400.0]), np.array([0.03]), np.array([True]), 7,
                                 np.array([0, 1, 6, 2]))[0]
    assert np.allclose(one, want, rtol=1e-8, atol=1e-12)


def test_filter_with_hessian_correction_enabled():
    """do_all_bands with hessian_correction=True: A - sum_b corr_b at the analysis,
    the update the reference has commented out at linear_kf.py:313-319."""
    g = gio.load("run_tip")
    mask = g["mask"]
    P = 7
    N = int(mask.sum())
    gps = gio.emulators(g, syn.EmulatorData)
    ogps = gio.emulators(g, GaussianProcess)
    obs = gio.FixtureObservations(g, gps)
    date = obs.dates[0]
    bands = gio.compacted_bands(obs, mask)[date]
    blocks = np.repeat(g["prior_inv_block"][None].astype(np.float64), N, axis=0)
    data = [obs.get_band_data(date, b) for b in range(2)]
    res = {}
    for flag in (False, True):
        kf = kb.LinearKalman(obs, syn.MemoryOutput(syn.TIP_PARAMETERS), mask, kb.create_nonlinear_observation_operator,
                             syn.TIP_PARAMETERS, prior=None, hessian_correction=flag)
        x, _, A, innov = kf.do_all_bands(date, data, g["x0"], None, BlockDiag.from_blocks(blocks, gu.DEV))
        res[flag] = (np.asarray(x), A.blocks(), np.asarray(innov).reshape(2, N))
    assert np.array_equal(res[True][0], res[False][0])
    mappers = [g["mapper_0"], g["mapper_1"]]
    corr = sum(kc.hessian_correction(ogps[b], res[False][0], bands[b][1], res[False][2][b], bands[b][2], P, mappers[b])
               for b in range(2))
    want = res[False][1] - corr
    assert np.max(np.abs(res[True][1] - want)) < 1e-9 * np.abs(want).max()
    assert np.abs(corr).max() > 1e-6 * np.abs(want).max()          # the correction is not trivially zero


# ------------------------------------------------------ observation ingest
@pytest.mark.parametrize("dtype", [np.int16, np.uint16, np.float32])
def test_scaled_reflectance_ingest_matches_reader_arithmetic(dtype):
    rng = np.random.default_rng(4)
    mask = gio.barrax_mask()
    P, B = 10, 10
    N = int(mask.sum())
    gps = [syn.make_emulator(rng, syn.SAIL_MEAN, syn.SAIL_SIGMA, n_train=60) for _ in range(B)]
    raws, floats = [], []
    for b in range(B):
        lo = 0 if dtype == np.uint16 else -30
        raw = rng.integers(lo, 6000, size=mask.shape).astype(dtype)
        raw[rng.random(mask.shape) < 0.1] = 0
        rho, inv, m = kc.s2_ingest(raw)
        raws.append(syn.BandData(raw, kb.ScaledReflectance(), None, {"band": b}, gps[b]))
        floats.append(syn.BandData(rho.astype(np.float64), inv, m, {"band": b}, gps[b]))
    date = dt.datetime(2017, 7, 3)
    inv_block = np.diag(1.0 / syn.SAIL_SIGMA ** 2)
    out = []
    for data in (raws, floats):
        kf = kb.LinearKalman(None, None, mask, kb.create_prosail_observation_operator, syn.S2_PARAMETERS, prior=None)
        x, _, A, innov = kf.do_all_bands(date, data, DeviceVector.broadcast(syn.SAIL_MEAN, N, gu.DEV), None,
                                         BlockDiag.from_block(inv_block, N, gu.DEV))
        out.append((np.asarray(x), A.blocks(), kf._stager.h2d_bytes))
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])   # bit-identical
    assert out[0][2] == B * mask.size * np.dtype(dtype).itemsize                            # 2-4 B/px instead of 17
    assert out[1][2] == B * mask.size * 17
    # with an extra mask AND-ed in
    extra = rng.random(mask.shape) < 0.7
    kf = kb.LinearKalman(None, None, mask, kb.create_prosail_observation_operator, syn.S2_PARAMETERS, prior=None)
    bufs = kf._stager.upload([raws[0]._replace(mask=extra)])
    rho, inv, m = kc.s2_ingest(raws[0].observations, ext_mask=extra)
    y, r, mm = kc.gather_band(rho.astype(np.float64), inv.ravel(), m, mask)
    assert np.array_equal(bufs.y[0, :N].cpu().numpy(), y)
    assert np.array_equal(bufs.w[0, :N].cpu().numpy(), kc.gate_inverse_variance(r, mm))
    assert np.array_equal(bufs.m[0, :N].cpu().numpy().astype(bool), mm)


# ---------------------------------------------------------------- MODIS BHR ingest
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_bhr_kernel_weights_ingest_matches_reader_arithmetic(dtype):
    """BHRObservations.get_band_data (observations.py:288-310) on the device: two
    broadbands (TIP layout), fed once as kernel weights + quality level and once as
    the float rasters the reader would have produced: bit-identical posterior."""
    rng = np.random.default_rng(14)
    mask = gio.barrax_mask()
    P, B = 7, 2
    N = int(mask.sum())
    mappers = [np.array([0, 1, 6, 2]), np.array([3, 4, 6, 5])]
    gps = [syn.make_emulator(rng, syn.TIP_MEAN[m], syn.TIP_SIGMA[m], n_train=60, w_scale=0.1) for m in mappers]
    kernel_bands, float_bands = [], []
    for b in range(B):
        k = np.stack([rng.uniform(0.05, 0.6, mask.shape), rng.uniform(0.0, 0.3, mask.shape),
                      rng.uniform(0.0, 0.08, mask.shape)]).astype(dtype)
        k[:, 3, 5] = 1e-4                                           # sigma floor 2.5e-3 active
        qa = rng.integers(0, 2, size=mask.shape)
        obs_mask = rng.random(mask.shape) < 0.85
        bhr, inv, m = kc.bhr_ingest(k, obs_mask, qa)
        kernel_bands.append(syn.BandData(k, kb.BHRKernelWeights(qa), obs_mask, {"band": b}, gps[b]))
        float_bands.append(syn.BandData(np.where(m, bhr, 0.0), inv, m, {"band": b}, gps[b]))
    date = dt.datetime(2017, 7, 3)
    _, _, inv_block = kc.tip_prior()
    out = []
    for data in (kernel_bands, float_bands):
        kf = kb.LinearKalman(None, None, mask, kb.create_nonlinear_observation_operator, syn.TIP_PARAMETERS, prior=None)
        x, _, A, innov = kf.do_all_bands(date, data, DeviceVector.broadcast(syn.TIP_MEAN, N, gu.DEV), None,
                                         BlockDiag.from_block(inv_block.astype(np.float64), N, gu.DEV))
        out.append((np.asarray(x), A.blocks(), np.asarray(innov)))
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])
    assert np.array_equal(out[0][2], out[1][2])
    # the gathered rows themselves, incl. a pixel whose quality level is outside {0, 1}: w = inf like the reader
    k, marker, obs_mask = kernel_bands[0].observations, kernel_bands[0].uncertainty, kernel_bands[0].mask.copy()
    qa = marker.qa_level.copy()
    r0, c0 = np.argwhere(mask & obs_mask)[0]
    qa[r0, c0] = 3
    kf = kb.LinearKalman(None, None, mask, kb.create_nonlinear_observation_operator, syn.TIP_PARAMETERS, prior=None)
    bufs = kf._stager.upload([kernel_bands[0]._replace(uncertainty=kb.BHRKernelWeights(qa))])
    bhr, inv, m = kc.bhr_ingest(k, obs_mask, qa)
    y, r, mm = kc.gather_band(np.where(m, bhr, 0.0), inv.ravel(), m, mask)
    assert np.array_equal(bufs.y[0, :N].cpu().numpy(), y)
    assert np.array_equal(bufs.w[0, :N].cpu().numpy(), np.where(mm, r, 0.0))
    assert np.isinf(bufs.w[0, :N].cpu().numpy()).sum() == 1
    assert kf._stager.h2d_bytes == k.nbytes + 2 * mask.size
    with pytest.raises(TypeError):
        kf._stager.upload([kernel_bands[0]._replace(observations=k[:2])])


# ---------------------------------------------------------------- Sentinel-1 sigma0 ingest
@pytest.mark.parametrize("name", ["f32", "f64"])
def test_backscatter_ingest_matches_the_unmodified_reader(name):
    """S1Observations.get_band_data (Sentinel1_Observations.py:151-196) on the device:
    rows gathered from the raw sigma0 raster equal, bit for bit, the rows gathered from
    the unmodified reader's outputs (fixture kat_s1_ingest) and from the oracle."""
    g = gio.load("kat_s1_ingest")
    rng = np.random.default_rng(8)
    shape = g["theta"].shape
    mask = rng.random(shape) < 0.7
    mask[3, 5] = mask[4, 6] = True                       # the two infinite-weight pixels are in the state
    N = int(mask.sum())
    kf = kb.LinearKalman(None, None, mask, kb.create_sar_observation_operator, ["lai", "sm"], prior=None)
    for b, pol in enumerate(("VV", "VH")):
        raw = g["%s_%s_raw" % (name, pol)]
        marker = kb.BackscatterSigma0(missing_value=float(g["missing_value"]))
        before = kf._stager.h2d_bytes
        bufs = kf._stager.upload([syn.BandData(raw, marker, None, {"band": b}, None)])
        assert kf._stager.h2d_bytes - before == raw.nbytes               # 4 or 8 B/px instead of 17
        y, r, mm = kc.gather_band(g["%s_%s_obs" % (name, pol)].astype(np.float64), g["%s_%s_inv_var" % (name, pol)],
                                  g["%s_%s_mask" % (name, pol)], mask)
        assert np.array_equal(bufs.y[0, :N].cpu().numpy(), y)
        assert np.array_equal(bufs.w[0, :N].cpu().numpy(), kc.gate_inverse_variance(r, mm))
        assert np.array_equal(bufs.m[0, :N].cpu().numpy().astype(bool), mm)
        assert np.isinf(bufs.w[0, :N].cpu().numpy()).sum() == 2        # unmasked zero and underflowing square, as the reader
        obs_o, inv_o, m_o = kc.s1_ingest(raw.copy(), missing_value=float(g["missing_value"]))
        yo, ro, mo = kc.gather_band(obs_o.astype(np.float64), inv_o.ravel(), m_o, mask)
        assert np.array_equal(y, yo) and np.array_equal(r, ro) and np.array_equal(mm, mo)
    with pytest.raises(TypeError):
        kf._stager.upload([syn.BandData(np.zeros(shape, dtype=np.int16), kb.BackscatterSigma0(), None, {}, None)])


def test_sar_filter_run_from_raw_sigma0_equals_run_from_reader_outputs():
    """The Water-Cloud filter run fed with raw sigma0 rasters + BackscatterSigma0 markers
    gives the same bits as the run fed with what the reader would have produced."""
    rng = np.random.default_rng(23)
    mask, N, obs, dates, base = _sar_case(rng)
    P = 2
    raw_store, float_store = {}, {}
    for d in dates:
        raws, floats = [], []
        for b in range(2):
            src = obs.get_band_data(d, b)
            sigma0 = np.where(src.mask & mask, src.observations, -999.0).astype(np.float32)
            o, inv, m = kc.s1_ingest(sigma0.copy())
            raws.append(syn.BandData(sigma0, kb.BackscatterSigma0(), None, {"band": b}, None))
            floats.append(syn.BandData(o.astype(np.float64), syn.inverse_variance_matrix(inv), m, {"band": b}, None))
        raw_store[d], float_store[d] = raws, floats
    mean = np.array([2.5, 0.25])
    inv = np.diag(1.0 / np.array([0.8, 0.08]) ** 2)
    outs = []
    for store in (raw_store, float_store):
        obs.get_band_data = lambda date, band, s=store: s[date][band]
        prior = syn.BroadcastPrior(mean, inv, N, lambda blk, n: BlockDiag.from_block(blk, n, gu.DEV))
        out = syn.MemoryOutput(["lai", "sm"])
        kf = kb.LinearKalman(obs, out, mask, kb.create_sar_observation_operator, ["lai", "sm"],
                             state_propagation=None, prior=prior, linear=False)
        kf.set_trajectory_model()
        kf.set_trajectory_uncertainty(np.zeros(N * P))
        x0, Pinv0 = prior.process_prior(None)
        grid = [base + dt.timedelta(days=3 * d) for d in range(len(dates) + 1)]
        kf.run(grid, x0, None, Pinv0)
        outs.append((out, kf.last_iterations))
    assert outs[0][1] == outs[1][1]
    for t in outs[0][0].output:
        assert np.array_equal(outs[0][0].output[t]["x"], outs[1][0].output[t]["x"])
        assert np.array_equal(outs[0][0].output[t]["diag"], outs[1][0].output[t]["diag"])


def test_s2_and_bhr_ingest_match_the_unmodified_readers():
    """Device ingest of raw Sentinel-2 reflectance (int16 / uint16 / float32) and of MODIS
    kernel weights against rows gathered from the outputs of the unmodified readers
    (fixture kat_reader_ingest): bit-exact."""
    g = gio.load("kat_reader_ingest")
    rng = np.random.default_rng(12)
    shape = g["s2_i16_raw"].shape
    mask = rng.random(shape) < 0.75
    mask[2, 3] = True
    N = int(mask.sum())
    kf = kb.LinearKalman(None, None, mask, kb.create_prosail_observation_operator, syn.S2_PARAMETERS, prior=None)
    cases = [(g["s2_%s_raw" % n], kb.ScaledReflectance(), None, "s2_" + n) for n in ("i16", "u16", "f32")]
    cases += [(g["bhr_%s_kernels" % n], kb.BHRKernelWeights(g["bhr_%s_qa" % n]), g["bhr_%s_mask_in" % n], "bhr_" + n)
              for n in ("f32", "f64")]
    for obs, marker, in_mask, key in cases:
        bufs = kf._stager.upload([syn.BandData(obs, marker, in_mask, {}, None)])
        want_obs = np.nan_to_num(g[key + "_obs"].astype(np.float64), nan=0.0)
        y, r, mm = kc.gather_band(want_obs, g[key + "_inv_var"], g[key + "_mask"], mask)
        assert np.array_equal(bufs.y[0, :N].cpu().numpy(), y), key
        assert np.array_equal(bufs.w[0, :N].cpu().numpy(), kc.gate_inverse_variance(r, mm)), key
        assert np.array_equal(bufs.m[0, :N].cpu().numpy().astype(bool), mm), key
