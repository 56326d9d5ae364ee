"""Generate ``tests/golden/*.npz`` by running the UNMODIFIED reference
(TEST INFRASTRUCTURE; runs only in the build container, see ``ref_shim.py``).

    python -m oracle.make_golden

Every fixture stores the full inputs next to the reference's outputs so that the
tests never need ``/root/reference`` (the GPU box does not have it).

Fixtures
--------
kat_propagators.npz  reference propagators / priors on the TIP prior
                     (kf_tools.py:99-116,174-353) + the numbers quoted in the
                     reference's own tests/test_kf.py:19-54
kat_time_grid.npz    iterate_time_grid on tests/test_utils.py:18-38's case
barrax_mask.npz      Barrax_pivots.tif as packed bits + compaction index digest
solve_s2.npz         variational_kalman_multiband (solvers.py:100-145) on
                     operators built by create_prosail_observation_operator
run_s2.npz           LinearKalman.run, S2/PROSAIL style (kafka_test_S2.py:174-194)
run_tip.npz          LinearKalman.run, 2-band TIP with state_mapper operators and
                     propagate_information_filter_LAI (kafka_test.py:156-217)
run_linear.npz       LinearKalman.run, 1 linear band in (H0,H) tuple form
kat_sar.npz          sar_observation_operator, VV and VH
                     (observation_operators/sar_forward_model.py:13-113)
kat_hessian.npz      hessian_correction(_multiband) on the TIP layout (kf_tools.py:26-72)
solve_indefinite.npz variational_kalman_multiband and blend_prior on forecast blocks that are NOT
                     positive definite (what the Hessian correction can hand on): the sparse LU
                     returns the solution of the equations (solvers.py:133-134, kf_tools.py:93-94)
"""
import datetime as dt
import hashlib
import logging
import os
import re

import numpy as np
import scipy.sparse as sp

from oracle import ref_shim
from oracle.gp import GaussianProcess
from kafka_inferenceengine_b200 import synthetic as syn

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                   "tests", "golden")


class _IterLog(logging.Handler):
    """Collect 'Iteration # n, convergence norm: g' lines (linear_kf.py:294-296)."""
    rx = re.compile(r"Iteration # (\d+), convergence norm: (\S+)")

    def __init__(self):
        super().__init__()
        self.rows = []

    def emit(self, record):
        m = self.rx.search(record.getMessage())
        if m:
            self.rows.append((int(m.group(1)), float(m.group(2))))


def _predict(gp, X):
    return gp.predict(X, do_unc=False)[0]


def _gp_arrays(prefix, gps):
    out = {}
    for b, gp in enumerate(gps):
        out["%s_inputs_%d" % (prefix, b)] = gp.inputs
        out["%s_targets_%d" % (prefix, b)] = gp.targets
        out["%s_theta_%d" % (prefix, b)] = gp.theta
    return out


def _obs_arrays(obs):
    out = {}
    for di, d in enumerate(obs.dates):
        for b in range(obs.bands_per_observation[d]):
            bd = obs.get_band_data(d, b)
            out["obs_y_%d_%d" % (di, b)] = bd.observations
            out["obs_invvar_%d_%d" % (di, b)] = bd.uncertainty.diagonal().reshape(bd.mask.shape)
            out["obs_mask_%d_%d" % (di, b)] = bd.mask
    out["obs_dates"] = np.array([d.toordinal() for d in obs.dates])
    return out


def _small_mask(rng, ny, nx, frac=0.55):
    m = rng.random((ny, nx)) < frac
    m[0, :] = False           # ragged edges: an empty row and a full row
    m[1, :] = True
    return m


class _RefOutput(object):
    def __init__(self):
        self.rows = []

    def dump_data(self, timestep, x, P, P_inv, state_mask, n_params):
        self.rows.append((timestep.toordinal(), np.array(x, dtype=np.float64),
                          np.array(P_inv.diagonal(), dtype=np.float64)))


def _run_arrays(out, itlog):
    res = {"out_steps": np.array([r[0] for r in out.rows]),
           "out_x": np.array([r[1] for r in out.rows]),
           "out_diag": np.array([r[2] for r in out.rows]),
           "iter_log": np.array(itlog.rows, dtype=np.float64).reshape(-1, 2)}
    return res


def kat_propagators(kafka):
    kt = kafka.inference.kf_tools
    x0, cov, inv = kt.tip_prior()
    n_pix = 3
    x_a = np.tile(x0, n_pix) + 0.01 * np.arange(7 * n_pix)
    Pi = sp.block_diag([inv.astype(np.float64) * (1 + 0.1 * i) for i in range(n_pix)]).tocsr()
    M = sp.eye(7 * n_pix, format="csr")
    q = np.tile(np.linspace(0.05, 0.15, 7), n_pix)
    Q = sp.diags(q).tocsr()
    res = {"tip_mean": x0, "tip_cov": cov, "tip_inv": inv, "x_a": x_a,
           "Pinv": Pi.toarray(), "q": q}
    _, _, p = kt.propagate_information_filter_approx_SLOW(x_a, None, Pi, M, Q)
    res["approx_Pinv"] = p.toarray()
    _, _, p = kt.propagate_information_filter_SLOW(x_a, None, Pi, M, Q)
    res["exact_Pinv"] = p.toarray()
    xf, _, p = kt.propagate_information_filter_LAI(x_a, None, Pi, M, Q)
    res["lai_x"] = xf
    res["lai_Pinv"] = p.toarray()
    xf, _, p = kt.no_propagation(x_a, None, Pi, M, Q)
    res["noprop_x"] = xf
    res["noprop_Pinv"] = p.toarray()
    P = sp.block_diag([cov.astype(np.float64)] * n_pix).tocsr()
    xf, pf, _ = kt.propagate_standard_kalman(x_a, P, None, M * 2.0, Q)
    res["std_x"] = xf
    res["std_P"] = pf.toarray()
    # blend_prior (kf_tools.py:75-96)
    prior_mean = np.tile(x0, n_pix)
    prior_inv = sp.block_diag([inv.astype(np.float64)] * n_pix).tocsr()
    xc, C = kt.blend_prior(prior_mean, prior_inv, x_a, Pi)
    res["blend_x"] = xc
    res["blend_C"] = C.toarray()
    # numbers quoted in the reference's own test (tests/test_kf.py:30-54): single pixel, Q=0.1 I
    Pi1 = sp.csr_matrix(inv.astype(np.float64))
    _, _, p = kt.propagate_information_filter_approx_SLOW(x0, None, Pi1, np.eye(7),
                                                          sp.csr_matrix(np.eye(7) * 0.1))
    res["test_kf_approx_diag"] = p.toarray().diagonal()
    _, _, p = kt.propagate_information_filter_SLOW(x0, None, Pi1, sp.eye(7),
                                                   sp.csr_matrix(np.eye(7) * 0.1))
    res["test_kf_exact"] = np.asarray(p.todense())
    np.savez_compressed(os.path.join(OUT, "kat_propagators.npz"), **res)


def kat_time_grid(kafka):
    base = dt.datetime(2007, 7, 1)
    grid = [base + i * dt.timedelta(days=1) for i in range(0, 60, 16)]
    base = dt.datetime(2007, 1, 1)
    dates = [base + i * dt.timedelta(days=1) for i in range(1, 365 + 8, 8)]
    steps, locs, firsts = [], [], []
    for t, loc, first in kafka.inference.utils.iterate_time_grid(grid, dates):
        steps.append(t.toordinal())
        locs.append([d.toordinal() for d in loc])
        firsts.append(first)
    np.savez_compressed(os.path.join(OUT, "kat_time_grid.npz"),
                        steps=np.array(steps), locs=np.array(locs), firsts=np.array(firsts))


def barrax_mask():
    import cv2
    m = cv2.imread(os.path.join(ref_shim.REFERENCE_ROOT, "Barrax_pivots.tif"),
                   cv2.IMREAD_UNCHANGED)
    mask = m.astype(bool)
    idx = np.flatnonzero(mask.ravel())
    np.savez_compressed(os.path.join(OUT, "barrax_mask.npz"),
                        shape=np.array(mask.shape), bits=np.packbits(mask.ravel()),
                        n_true=np.array(idx.size),
                        idx_sha256=np.array(hashlib.sha256(idx.astype(np.int64).tobytes()).hexdigest()))
    return mask


def solve_and_run_s2(kafka):
    rng = np.random.default_rng(20240601)
    P, B, T = 10, 10, 250
    mask = _small_mask(rng, 18, 23)
    N = int(mask.sum())
    gps = [syn.make_emulator(rng, syn.SAIL_MEAN, syn.SAIL_SIGMA, n_train=T, cls=GaussianProcess)
           for _ in range(B)]
    x_true = syn.SAIL_MEAN + syn.SAIL_SIGMA * rng.standard_normal((N, P))
    base = dt.datetime(2017, 7, 3)
    dates = [base + dt.timedelta(days=d) for d in (0, 1, 4)]
    obs = syn.SyntheticObservations(mask, dates, gps, x_true, _predict, rng)
    inv_cov = np.diag(1.0 / syn.SAIL_SIGMA ** 2).astype(np.float32)
    ku = kafka.inference.utils

    class Prior(object):
        def process_prior(self, time, inv_cov=True):
            x0 = np.tile(syn.SAIL_MEAN, N)
            return x0, ku.block_diag([inv_cov_blk for _ in range(N)], dtype=np.float32)
    inv_cov_blk = inv_cov
    prior = Prior()
    x0, Pinv0 = prior.process_prior(None)

    # ---- single linearised solve around a perturbed state (solvers.py:100-145)
    x_lin = x0 + np.tile(0.3 * syn.SAIL_SIGMA, N) * rng.standard_normal(N * P)
    d0 = dates[0]
    Hs = [ku.create_prosail_observation_operator(P, gps[b], None, obs.get_band_data(d0, b).mask,
                                                 mask, x_lin, b) for b in range(B)]
    bd = [obs.get_band_data(d0, b) for b in range(B)]
    x, _, A, innov, fwd = kafka.inference.solvers.variational_kalman_multiband(
        [b_.observations for b_ in bd], [b_.mask for b_ in bd], mask,
        [b_.uncertainty for b_ in bd], Hs, P, x_lin, x0, None, Pinv0.tocsr(), None)
    res = {"mask": mask, "x_true": x_true, "x0": x0, "x_lin": x_lin,
           "prior_inv_block": inv_cov, "ref_x": x, "ref_A_diag": A.diagonal(),
           "ref_A_blocks_first": A[:3 * P, :3 * P].toarray(),
           "ref_innovations": innov, "ref_fwd": fwd,
           "ref_H0": np.array([h[0] for h in Hs]),
           "ref_J": np.array([np.asarray(h[1].tocsr().data) for h in Hs], dtype=object)}
    # J rows are ragged (unobserved rows are empty): store dense (B, N, P)
    Jd = np.zeros((B, N, P), dtype=np.float32)
    for b, h in enumerate(Hs):
        Hc = h[1].tocoo()
        Jd[b, Hc.row, Hc.col - P * Hc.row] = Hc.data
    res["ref_J"] = Jd
    res.update(_gp_arrays("gp", gps))
    res.update(_obs_arrays(obs))
    np.savez_compressed(os.path.join(OUT, "solve_s2.npz"), **res)

    # ---- full run (kafka_test_S2.py:174-194): prior-only advance, Q = 0
    itlog = _IterLog()
    logging.getLogger("kafka.linear_kf.linear_kf").addHandler(itlog)
    logging.getLogger("kafka.linear_kf.linear_kf").setLevel(logging.INFO)
    out = _RefOutput()
    kf = kafka.LinearKalman(obs, out, mask, ku.create_prosail_observation_operator,
                            syn.S2_PARAMETERS, state_propagation=None, prior=prior,
                            linear=False)
    kf.set_trajectory_model()
    kf.set_trajectory_uncertainty(np.zeros_like(x0))
    grid = [base + dt.timedelta(days=d) for d in (0, 2, 4, 6, 8)]
    kf.run(grid, x0, None, Pinv0, iter_obs_op=True)
    logging.getLogger("kafka.linear_kf.linear_kf").removeHandler(itlog)
    res2 = {"mask": mask, "x_true": x_true, "x0": x0, "prior_inv_block": inv_cov,
            "grid": np.array([g.toordinal() for g in grid])}
    res2.update(_gp_arrays("gp", gps))
    res2.update(_obs_arrays(obs))
    res2.update(_run_arrays(out, itlog))
    np.savez_compressed(os.path.join(OUT, "run_s2.npz"), **res2)


def run_tip(kafka):
    rng = np.random.default_rng(77)
    P, B, T = 7, 2, 120
    mask = _small_mask(rng, 14, 17)
    N = int(mask.sum())
    mappers = [np.array([0, 1, 6, 2]), np.array([3, 4, 6, 5])]
    gps = [syn.make_emulator(rng, syn.TIP_MEAN[m], syn.TIP_SIGMA[m], n_train=T, w_scale=0.15,
                             cls=GaussianProcess) for m in mappers]
    x_true = syn.TIP_MEAN + 0.5 * syn.TIP_SIGMA * rng.standard_normal((N, P))
    base = dt.datetime(2017, 1, 1)
    dates = [base + dt.timedelta(days=d) for d in (1, 9, 17, 18, 33)]
    obs = syn.SyntheticObservations(mask, dates, gps, x_true, _predict, rng,
                                    state_mappers=mappers, min_sigma=2.5e-3)
    kt = kafka.inference.kf_tools
    ku = kafka.inference.utils
    mean, prior_inv = kt.tip_prior_full({"n_pixels": N})
    itlog = _IterLog()
    lg = logging.getLogger("kafka.linear_kf.linear_kf")
    lg.addHandler(itlog)
    lg.setLevel(logging.INFO)
    out = _RefOutput()
    kf = kafka.LinearKalman(obs, out, mask, ku.create_nonlinear_observation_operator,
                            syn.TIP_PARAMETERS,
                            state_propagation=kt.propagate_information_filter_LAI,
                            prior=None, linear=False)
    q = np.zeros(N * P)
    q[6::7] = 0.04                                           # kafka_test.py:207-208
    kf.set_trajectory_model()
    kf.set_trajectory_uncertainty(q)
    grid = [base + dt.timedelta(days=d) for d in range(0, 49, 8)]
    kf.run(grid, mean, None, prior_inv.tocsr(), iter_obs_op=True)
    lg.removeHandler(itlog)
    res = {"mask": mask, "x_true": x_true, "x0": mean, "q": q,
           "prior_inv_block": kt.tip_prior()[2],
           "grid": np.array([g.toordinal() for g in grid]),
           "mapper_0": mappers[0], "mapper_1": mappers[1]}
    res.update(_gp_arrays("gp", gps))
    res.update(_obs_arrays(obs))
    res.update(_run_arrays(out, itlog))
    np.savez_compressed(os.path.join(OUT, "run_tip.npz"), **res)


def run_linear(kafka):
    """Config-1 shape: TIP state, one linear band observing slot 6, LAI propagator."""
    rng = np.random.default_rng(5)
    P = 7
    mask = _small_mask(rng, 12, 15)
    N = int(mask.sum())
    row = np.zeros(P)
    row[6] = 1.0
    op = syn.LinearBandOperator(row, 0.0)
    x_true = syn.TIP_MEAN + 0.5 * syn.TIP_SIGMA * rng.standard_normal((N, P))
    base = dt.datetime(2017, 1, 1)
    dates = [base + dt.timedelta(days=d) for d in range(1, 11)]
    obs = syn.LinearObservations(mask, dates, op, x_true, rng, sigma_obs=0.05, p_observed=0.85)
    kt = kafka.inference.kf_tools

    def create_linear_tuple_operator(n_params, emulator, metadata, mask_, state_mask, x_prev, band):
        # same contract as utils.py:181-219 with a linear forward model
        n = x_prev.shape[0] // n_params
        H = sp.lil_matrix((n, n_params * n), dtype=np.float32)
        H0 = np.zeros(n, dtype=np.float32)
        for i, m in enumerate(mask_[state_mask].flatten()):
            if m:
                H[i, n_params * i:n_params * (i + 1)] = emulator.row
                H0[i] = emulator.row @ x_prev[n_params * i:n_params * (i + 1)] + emulator.offset
        return (H0, H.tocsr())

    mean, prior_inv = kt.tip_prior_full({"n_pixels": N})
    itlog = _IterLog()
    lg = logging.getLogger("kafka.linear_kf.linear_kf")
    lg.addHandler(itlog)
    lg.setLevel(logging.INFO)
    out = _RefOutput()
    kf = kafka.LinearKalman(obs, out, mask, create_linear_tuple_operator, syn.TIP_PARAMETERS,
                            state_propagation=kt.propagate_information_filter_LAI, prior=None)
    q = np.zeros(N * P)
    q[6::7] = 0.04
    kf.set_trajectory_model()
    kf.set_trajectory_uncertainty(q)
    grid = [base + dt.timedelta(days=d) for d in range(0, 12)]
    kf.run(grid, mean, None, prior_inv.tocsr())
    lg.removeHandler(itlog)
    res = {"mask": mask, "x_true": x_true, "x0": mean, "q": q, "row": row,
           "prior_inv_block": kt.tip_prior()[2],
           "grid": np.array([g.toordinal() for g in grid])}
    res.update(_obs_arrays(obs))
    res.update(_run_arrays(out, itlog))
    np.savez_compressed(os.path.join(OUT, "run_linear.npz"), **res)


def kat_sar(kafka):
    """sar_observation_operator (observation_operators/sar_forward_model.py:13-113)."""
    from kafka.observation_operators.sar_forward_model import sar_observation_operator
    rng = np.random.default_rng(9)
    n = 400
    x = np.stack([rng.uniform(0.05, 7.0, n), rng.uniform(0.02, 0.6, n)], axis=1)
    res = {"x": x}
    for pol in ("VV", "VH"):
        s0, g = sar_observation_operator(x, np.full(n, 23.0), pol)
        res["sigma0_" + pol] = s0
        res["grad_" + pol] = g
    np.savez_compressed(os.path.join(OUT, "kat_sar.npz"), **res)


def kat_hessian(kafka):
    """hessian_correction / hessian_correction_multiband (kf_tools.py:26-72) on the
    TIP layout, with the oracle GP providing gp.hessian."""
    kt = kafka.inference.kf_tools
    rng = np.random.default_rng(31)
    P = 7
    mask = _small_mask(rng, 9, 11)
    N = int(mask.sum())
    mappers = [np.array([0, 1, 6, 2]), np.array([3, 4, 6, 5])]
    gps = [syn.make_emulator(rng, syn.TIP_MEAN[m], syn.TIP_SIGMA[m], n_train=80, w_scale=0.15,
                             cls=GaussianProcess) for m in mappers]
    x0 = (syn.TIP_MEAN + 0.4 * syn.TIP_SIGMA * rng.standard_normal((N, P))).ravel()
    res = {"mask": mask, "x0": x0}
    res.update(_gp_arrays("gp", gps))
    Rs, innovs, masks = [], [], []
    for b in range(2):
        omask = rng.random(mask.shape) < 0.8
        inv = np.where(omask, rng.uniform(100., 900., mask.shape), np.inf)
        R = syn.inverse_variance_matrix(inv)
        innov = 0.05 * rng.standard_normal(N)
        corr = kt.hessian_correction(gps[b], x0, R, innov, omask, mask, b, P)
        dense = corr.toarray()
        res["corr_%d" % b] = np.array([dense[i * P:(i + 1) * P, i * P:(i + 1) * P] for i in range(N)])
        res["obs_mask_%d" % b], res["inv_var_%d" % b], res["innov_%d" % b] = omask, inv, innov
        Rs.append(R)
        innovs.append(innov)
        masks.append(omask)
    # the reference passes ONE emulator for every band (linear_kf.py:314)
    multi = kt.hessian_correction_multiband(gps[0], x0, Rs, innovs, masks, mask, 2, P)
    dense = multi.toarray()
    res["corr_multi_gp0"] = np.array([dense[i * P:(i + 1) * P, i * P:(i + 1) * P] for i in range(N)])
    np.savez_compressed(os.path.join(OUT, "kat_hessian.npz"), **res)


def kat_s1_ingest(kafka):
    """S1Observations.get_band_data (input_output/Sentinel1_Observations.py:151-196),
    unmodified, on synthetic sigma0 rasters: the module's ``reproject_image`` (GDAL) is
    replaced by a stand-in that hands back the synthetic arrays."""
    import datetime as dt
    import kafka.input_output.Sentinel1_Observations as s1

    class _Ptr(object):
        def __init__(self, a):
            self.a = a

        def ReadAsArray(self):
            return self.a.copy()

    rng = np.random.default_rng(21)
    ny, nx = 37, 41
    theta = (30.0 + 10.0 * rng.random((ny, nx))).astype(np.float32)
    res = {"theta": theta, "missing_value": np.float64(s1.WRONG_VALUE)}
    date = dt.datetime(2017, 5, 4, 6, 2, 11)
    for name, dtype in (("f32", np.float32), ("f64", np.float64)):
        rasters = {}
        for pol in ("VV", "VH"):
            r = (-26.0 + 22.0 * rng.random((ny, nx))).astype(dtype)
            r[rng.random((ny, nx)) < 0.15] = s1.WRONG_VALUE
            r[3, 5] = 0.0                                  # unmasked zero: infinite weight in the reader
            r[4, 6] = 1e-30 if dtype == np.float32 else 1e-200   # square underflows -> inf as well
            rasters[pol] = r
        reader = object.__new__(s1.S1Observations)
        reader.date_data = {date: "S1A_IW_GRDH_1SDV_x_20170504T060211.nc"}
        reader.emulators = {"VV": "vv-model", "VH": "vh-model"}
        reader.state_mask = "state_mask.tif"

        def fake_reproject(fname, target, dstSRSs=None, _r=rasters):
            for pol in ("VV", "VH"):
                if fname.endswith("sigma0_" + pol):
                    return _Ptr(_r[pol])
            assert fname.endswith(":theta"), fname
            return _Ptr(theta)
        s1.reproject_image = fake_reproject
        for band, pol in enumerate(("VV", "VH")):
            with np.errstate(divide="ignore", over="ignore"):
                d = reader.get_band_data(date, band)
            assert d.emulator == reader.emulators[pol]
            assert np.array_equal(d.metadata["incidence_angle"], theta)
            res["%s_%s_raw" % (name, pol)] = rasters[pol]
            res["%s_%s_obs" % (name, pol)] = d.observations
            res["%s_%s_inv_var" % (name, pol)] = d.uncertainty.diagonal()
            res["%s_%s_mask" % (name, pol)] = d.mask
    np.savez_compressed(os.path.join(OUT, "kat_s1_ingest.npz"), **res)


def kat_reader_ingest(kafka):
    """Sentinel2Observations.get_band_data (input_output/Sentinel2_Observations.py:148-185)
    and BHRObservations.get_band_data (input_output/observations.py:288-310), unmodified,
    on synthetic rasters.  Stand-ins replace what needs GDAL or un-vendored packages:
    ``reproject_image`` / ``parse_xml`` / ``_find_emulator`` (S2, plus an emulator pickle
    written to a temporary file) and ``get_brdf_descriptors`` (MODIS)."""
    import datetime as dt
    import pickle
    import tempfile
    import kafka.input_output.Sentinel2_Observations as s2
    import kafka.input_output.observations as mod

    class _Ptr(object):
        def __init__(self, a):
            self.a = a

        def ReadAsArray(self):
            return self.a.copy()

    rng = np.random.default_rng(33)
    ny, nx = 29, 31
    res = {}
    date = dt.datetime(2017, 7, 3)
    band_names = [b"S2A_MSI_%02d" % k for k in (2, 3, 4, 5, 6, 7, 8, 9, 12, 13)]
    with tempfile.TemporaryDirectory() as tmp:
        emu_file = os.path.join(tmp, "emulators.pkl")
        with open(emu_file, "wb") as f:
            pickle.dump({k: "emulator-%s" % k.decode() for k in band_names}, f)
        for name, dtype in (("i16", np.int16), ("u16", np.uint16), ("f32", np.float32)):
            raw = rng.integers(0 if dtype == np.uint16 else -40, 7000, size=(ny, nx)).astype(dtype)
            raw[rng.random((ny, nx)) < 0.12] = 0
            reader = object.__new__(s2.Sentinel2Observations)
            reader.date_data = {date: os.path.join(tmp, "granule")}
            reader.band_map = ['02', '03', '04', '05', '06', '07', '08', '8A', '09', '12']
            reader.state_mask = "state_mask.tif"
            reader._find_emulator = lambda sza, saa, vza, vaa: emu_file
            s2.parse_xml = lambda fname: (30.0, 150.0, 5.0, 100.0)
            s2.reproject_image = lambda fname, target, dstSRSs=None, _r=raw: _Ptr(_r)
            with np.errstate(divide="ignore"):
                d = reader.get_band_data(date, 3)
            assert d.emulator == "emulator-S2A_MSI_05"
            res["s2_%s_raw" % name] = raw
            res["s2_%s_obs" % name] = d.observations
            res["s2_%s_inv_var" % name] = d.uncertainty.diagonal()
            res["s2_%s_mask" % name] = d.mask
    for name, dtype in (("f32", np.float32), ("f64", np.float64)):
        kernels = rng.uniform(0.0, 0.5, (3, ny, nx)).astype(dtype)
        kernels[:, 2, 3] = 0.0                               # bhr = 0 -> sigma clamps at 2.5e-3
        mask = rng.random((ny, nx)) < 0.8
        qa = rng.integers(0, 4, size=(ny, nx))
        reader = object.__new__(mod.BHRObservations)
        reader.emulator = "tip-emulator"
        reader.get_brdf_descriptors = lambda band_no, the_date, _r=(kernels, mask, qa): _r
        with np.errstate(divide="ignore", invalid="ignore"):
            d = reader.get_band_data(date, 0)
        assert d.emulator == "tip-emulator" and d.metadata is None
        res["bhr_%s_kernels" % name] = kernels
        res["bhr_%s_mask_in" % name] = mask
        res["bhr_%s_qa" % name] = qa
        res["bhr_%s_obs" % name] = d.observations
        res["bhr_%s_inv_var" % name] = d.uncertainty.diagonal()
        res["bhr_%s_mask" % name] = d.mask
    np.savez_compressed(os.path.join(OUT, "kat_reader_ingest.npz"), **res)


def kat_output_stage(kafka):
    """KafkaOutput.dump_data (input_output/observations.py:354-394), unmodified: the
    module's ``gdal`` is replaced by a stand-in whose driver keeps the arrays handed to
    WriteArray instead of writing GeoTIFFs."""
    import datetime as dt
    import scipy.sparse as sp
    import kafka.input_output.observations as mod

    written = {}

    class _Band(object):
        def __init__(self, name):
            self.name = name

        def WriteArray(self, a):
            written[self.name] = np.array(a)

    class _Dataset(object):
        def __init__(self, name):
            self.name = name

        def SetProjection(self, p):
            pass

        def SetGeoTransform(self, g):
            pass

        def GetRasterBand(self, k):
            return _Band(self.name)

    class _Driver(object):
        def Create(self, fname, nx, ny, nb, gdt, opts):
            return _Dataset(os.path.basename(fname))

    class _Gdal(object):
        GDT_Float32 = 6

        @staticmethod
        def GetDriverByName(fmt):
            return _Driver()

    mod.gdal = _Gdal
    rng = np.random.default_rng(41)
    mask = _small_mask(rng, 23, 19)
    N, P = int(mask.sum()), 3
    names = ["lai", "cab", "sm"]
    x = rng.standard_normal(N * P) * np.tile([3.0, 40.0, 0.2], N)
    blocks = []
    for _ in range(N):
        g = rng.standard_normal((P, P))
        blocks.append(g @ g.T + np.diag(rng.uniform(0.5, 50.0, P)))
    P_inv = sp.block_diag(blocks).tocsr()
    date = dt.datetime(2017, 7, 3)
    out = mod.KafkaOutput(names, [0, 10, 0, 0, 0, -10], "proj", "/nowhere")
    out.dump_data(date, x, None, P_inv, mask, P)
    res = {"mask": mask, "x": x, "blocks": np.array(blocks)}
    for ii, name in enumerate(names):
        res["mean_%d" % ii] = written["%s_A2017184.tif" % name]
        res["unc_%d" % ii] = written["%s_A2017184_unc.tif" % name]
        assert res["mean_%d" % ii].dtype == np.float32 and res["unc_%d" % ii].dtype == np.float32
    np.savez_compressed(os.path.join(OUT, "kat_output_stage.npz"), **res)


def solve_indefinite(kafka):
    """The unmodified solver and prior blend on a forecast whose inverse-covariance blocks are
    indefinite for every third pixel (one eigenvalue flipped); two bands with supplied
    linearisations in the (H0, dH) tuple form, some pixels unobserved."""
    rng = np.random.default_rng(4242)
    P, B = 7, 2
    mask = _small_mask(rng, 9, 11)
    N = int(mask.sum())
    ku = kafka.inference.utils
    blocks = np.zeros((N, P, P))
    for i in range(N):
        M = rng.standard_normal((P, P))
        S = M @ M.T / P + np.eye(P)
        if i % 3 == 0:                                   # flip the middle eigenvalue: indefinite, well conditioned
            w, V = np.linalg.eigh(S)
            w[P // 2] = -w[P // 2]
            S = (V * w) @ V.T
        blocks[i] = 0.5 * (S + S.T)
    blocks = blocks.astype(np.float32).astype(np.float64)
    Pinv = sp.block_diag([b for b in blocks], format="csr")
    x_f = rng.standard_normal(N * P)
    x_lin = x_f + 0.1 * rng.standard_normal(N * P)
    n_raster = mask.size
    Hs, obs_r, unc, masks, Jd, H0d = [], [], [], [], [], []
    rows = np.repeat(np.arange(N), P)
    cols = np.arange(N * P)
    for b in range(B):
        m = np.zeros(mask.shape, dtype=bool)
        m[mask] = rng.random(N) < 0.8
        J = rng.standard_normal((N, P)).astype(np.float32)
        J[~m[mask]] = 0.0                               # unobserved pixels carry an empty row (utils.py:199-215)
        H0 = (rng.standard_normal(N) * m[mask]).astype(np.float32)
        dH = sp.csr_matrix((J.ravel(), (rows, cols)), shape=(N, N * P), dtype=np.float32)
        y = np.zeros(mask.shape)
        y[mask] = rng.standard_normal(N)
        r = np.zeros(n_raster)
        r[m.ravel()] = 1.0 / (0.05 + 0.1 * rng.random(int(m.sum()))) ** 2
        Hs.append((H0, dH))
        obs_r.append(y)
        unc.append(sp.diags(r, format="csr"))
        masks.append(m)
        Jd.append(J)
        H0d.append(H0)
    x, _, A, innov, fwd = kafka.inference.solvers.variational_kalman_multiband(
        obs_r, masks, mask, unc, Hs, P, x_lin, x_f, None, Pinv, None)
    # prior blend with the same forecast blocks (kf_tools.py:75-96)
    prior_inv_blocks = (1e-2 * np.eye(P)[None] * (1.0 + rng.random((N, 1, 1)))).astype(np.float32).astype(np.float64)
    prior_mean = rng.standard_normal(N * P)
    xb, Cb = kafka.inference.kf_tools.blend_prior(prior_mean, sp.block_diag([b for b in prior_inv_blocks], format="csr"),
                                                  x_f, Pinv)
    # exact propagator on the same blocks (kf_tools.py:208-245: spsolve of (I + P^-1 Q) X = P^-1), with an
    # inflation large enough for the flipped direction to matter
    q = rng.uniform(0.5, 1.5, N * P) / np.repeat(np.abs(np.linalg.eigvalsh(blocks)).min(axis=1), P)
    _, _, Pf = kafka.inference.kf_tools.propagate_information_filter_SLOW(
        x_f, None, Pinv, sp.eye(N * P, format="csr"), sp.diags(q, format="csr"))
    Pf = sp.csr_matrix(Pf).toarray()
    res = {"q": q, "ref_exact_blocks": np.array([Pf[i * P:(i + 1) * P, i * P:(i + 1) * P] for i in range(N)]),
           "mask": mask, "blocks": blocks, "x_f": x_f, "x_lin": x_lin,
           "J": np.array(Jd), "H0": np.array(H0d), "obs": np.array(obs_r), "inv_var": np.array([u.diagonal() for u in unc]),
           "obs_mask": np.array(masks), "ref_x": np.asarray(x), "ref_A_diag": A.diagonal(),
           "ref_innovations": innov, "ref_fwd": fwd,
           "prior_inv_blocks": prior_inv_blocks, "prior_mean": prior_mean, "ref_blend_x": np.asarray(xb),
           "ref_blend_C_diag": Cb.diagonal()}
    np.savez_compressed(os.path.join(OUT, "solve_indefinite.npz"), **res)


def main():
    os.makedirs(OUT, exist_ok=True)
    kafka = ref_shim.load()
    kat_sar(kafka)
    kat_s1_ingest(kafka)
    kat_reader_ingest(kafka)
    kat_output_stage(kafka)
    kat_hessian(kafka)
    kat_propagators(kafka)
    kat_time_grid(kafka)
    barrax_mask()
    solve_and_run_s2(kafka)
    run_tip(kafka)
    run_linear(kafka)
    solve_indefinite(kafka)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
