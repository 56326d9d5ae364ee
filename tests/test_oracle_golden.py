"""Pin the CPU oracle (oracle/kafka_cpu.py) against the reference: its own
known-answer tests, the Barrax fixture and outputs of the unmodified reference
stored in tests/golden (CPU only)."""
import datetime as dt
import hashlib

import numpy as np
import pytest

import golden_io as gio
from oracle import kafka_cpu as kc
from oracle.gp import GaussianProcess


# ---- the reference's own three known-answer tests -------------------------
def test_kat_standard_kalman():
    # /root/reference/tests/test_kf.py:19-27
    x_a = np.ones(3)
    x_f, P_f = kc.propagate_standard_kalman(x_a, np.eye(3)[None], np.full(3, 0.5), m_diag=2.0 * np.ones(3))
    assert np.all(x_f == 2.0 * x_a)
    assert np.all(P_f[0] == np.eye(3) + 0.5 * np.eye(3))


def test_kat_information_filter_numbers():
    # /root/reference/tests/test_kf.py:30-54
    mean, cov, inv = kc.tip_prior()
    _, d = kc.propagate_information_filter_approx(mean, inv.diagonal().astype(np.float64), np.full(7, 0.1))
    assert np.allclose(d, [8.74, 1.69, 9.81, 8.16, 0.43, 9.21, 2.86], atol=0.01)
    _, ex = kc.propagate_information_filter_exact(mean, inv[None], np.full(7, 0.1))
    want = np.diag([8.74, 1.69, 9.33, 8.16, 0.43, 7.28, 2.86])
    want[2, 5] = want[5, 2] = -1.13
    assert np.allclose(ex[0], want, atol=0.01)
    g = gio.load("kat_propagators")
    assert np.allclose(d, g["test_kf_approx_diag"], rtol=1e-12)
    assert np.allclose(ex[0], g["test_kf_exact"], rtol=1e-9, atol=1e-12)


def test_kat_time_grid():
    # /root/reference/tests/test_utils.py:18-38
    base = dt.datetime(2007, 7, 1)
    grid = [base + i * dt.timedelta(days=1) for i in range(0, 60, 16)]
    base = dt.datetime(2007, 1, 1)
    dates = [base + i * dt.timedelta(days=1) for i in range(1, 365 + 8, 8)]
    good = [dt.datetime(2007, 7, 17), dt.datetime(2007, 8, 2), dt.datetime(2007, 8, 18)]
    obs = [[dt.datetime(2007, 7, 5), dt.datetime(2007, 7, 13)],
           [dt.datetime(2007, 7, 21), dt.datetime(2007, 7, 29)],
           [dt.datetime(2007, 8, 6), dt.datetime(2007, 8, 14)]]
    g = gio.load("kat_time_grid")
    for i, (t, loc, first) in enumerate(kc.iterate_time_grid(grid, dates)):
        assert t == good[i]
        assert list(loc) == obs[i]
        assert first == (i == 0) == bool(g["firsts"][i])
        assert t.toordinal() == g["steps"][i]
        assert [d.toordinal() for d in loc] == list(g["locs"][i])


def test_barrax_compaction():
    g = gio.load("barrax_mask")
    mask = gio.barrax_mask()
    assert mask.shape == (204, 235) and mask.sum() == 13027 == int(g["n_true"])
    idx = kc.compact_mask(mask)
    assert hashlib.sha256(idx.astype(np.int64).tobytes()).hexdigest() == str(g["idx_sha256"])
    raster = np.arange(mask.size, dtype=np.float64).reshape(mask.shape)
    assert np.array_equal(raster.ravel()[idx], raster[mask])


# ---- propagators against the reference's outputs --------------------------
def _blocks(dense, n):
    N = dense.shape[0] // n
    return np.array([dense[i * n:(i + 1) * n, i * n:(i + 1) * n] for i in range(N)])


def test_propagators_vs_reference():
    g = gio.load("kat_propagators")
    mean, cov, inv = kc.tip_prior()
    assert np.array_equal(mean, g["tip_mean"]) and np.array_equal(inv, g["tip_inv"])
    x_a, q = g["x_a"], g["q"]
    Pb = _blocks(g["Pinv"], 7)
    diag = g["Pinv"].diagonal()
    _, d = kc.propagate_information_filter_approx(x_a, diag, q)
    assert np.allclose(np.diag(d), g["approx_Pinv"], rtol=1e-13, atol=0)
    _, ex = kc.propagate_information_filter_exact(x_a, Pb, q)
    assert np.allclose(ex, _blocks(g["exact_Pinv"], 7), rtol=1e-9, atol=1e-10)
    xf, lb = kc.propagate_information_filter_lai(x_a, diag, q)
    assert np.array_equal(xf, g["lai_x"])
    assert np.array_equal(lb.astype(np.float32), _blocks(g["lai_Pinv"], 7).astype(np.float32))
    xf, nb = kc.no_propagation(3)
    assert np.array_equal(xf, g["noprop_x"])
    assert np.array_equal(nb.astype(np.float32), _blocks(g["noprop_Pinv"], 7).astype(np.float32))
    cb = np.repeat(cov[None].astype(np.float64), 3, axis=0)
    xf, pf = kc.propagate_standard_kalman(x_a, cb, q, m_diag=2.0 * np.ones(21))
    assert np.array_equal(xf, g["std_x"])
    assert np.allclose(pf, _blocks(g["std_P"], 7), rtol=1e-15)
    pm = np.tile(mean, 3)
    pinv = np.repeat(inv[None].astype(np.float64), 3, axis=0)
    xc, C = kc.blend_prior(pm, pinv, x_a, Pb, mode="sparse32")
    assert np.allclose(C, _blocks(g["blend_C"], 7), rtol=1e-15)
    assert np.allclose(xc, g["blend_x"], rtol=2e-6)
    xc64, _ = kc.blend_prior(pm, pinv, x_a, Pb)
    assert np.allclose(xc64, g["blend_x"], rtol=1e-5)


# ---- operators + one linearised solve --------------------------------------
def test_solve_s2_vs_reference():
    g = gio.load("solve_s2")
    mask = g["mask"]
    gps = gio.emulators(g, GaussianProcess)
    obs = gio.FixtureObservations(g, gps)
    bands = gio.compacted_bands(obs, mask)[obs.dates[0]]
    P = 10
    N = int(mask.sum())
    x_lin, x0 = g["x_lin"], g["x0"]
    H0s, Js = [], []
    for b, gp in enumerate(gps):
        H0, J = kc.emulator_operator(gp, x_lin, P, bands[b][2])
        # operator values are float32 in the reference (utils.py:193-195): bit-equal
        # up to fp64 summation order inside the GP, i.e. within 1 float32 ulp
        assert np.allclose(H0, g["ref_H0"][b], rtol=2e-7, atol=1e-9)
        assert np.allclose(J, g["ref_J"][b], rtol=2e-7, atol=1e-9)
        H0s.append(H0)
        Js.append(J)
    ys = [b_[0] for b_ in bands]
    ws = [b_[1] for b_ in bands]
    Pinv = np.repeat(g["prior_inv_block"][None].astype(np.float64), N, axis=0)
    xs, As, innov_s, fwd_s = kc.solve_sparse32(ys, ws, H0s, Js, x_lin, x0, kc.blocks_to_sparse(Pinv))
    xd, Ad, innov_d, fwd_d = kc.solve_dense64(ys, ws, H0s, Js, x_lin, x0, Pinv)
    # oracle A' reproduces the reference's float32 path
    assert np.allclose(As.diagonal(), g["ref_A_diag"], rtol=3e-7)
    assert np.allclose(xs, g["ref_x"], rtol=5e-5, atol=1e-6)
    assert np.allclose(innov_s, g["ref_innovations"], rtol=1e-6, atol=1e-7)
    assert np.allclose(fwd_s, g["ref_fwd"], rtol=1e-3, atol=1e-5)
    # oracle B agrees with the reference to the reference's own float32 accuracy
    idx = np.arange(P)
    assert np.allclose(Ad[:, idx, idx].ravel(), g["ref_A_diag"], rtol=3e-7)
    assert np.median(gio.rel_err(xd, g["ref_x"])) < 5e-6       # float32 LU error of the reference
    assert np.allclose(xd, g["ref_x"], rtol=1e-3, atol=1e-5)
    assert np.allclose(Ad[:3].reshape(-1), _blocks(g["ref_A_blocks_first"], P).reshape(-1), rtol=3e-7, atol=1e-4)


def _s2_operators(gps, P, bands, mapper=None):
    def make(date):
        return [(lambda xp, gp=gp, m=bands[date][b][2], mp=(None if mapper is None else mapper[b]):
                 kc.emulator_operator(gp, xp, P, m, state_mapper=mp))
                for b, gp in enumerate(gps)]
    return make


@pytest.mark.parametrize("mode", ["sparse32", "dense64"])
def test_run_s2_vs_reference(mode):
    g = gio.load("run_s2")
    mask = g["mask"]
    P = 10
    N = int(mask.sum())
    gps = gio.emulators(g, GaussianProcess)
    obs = gio.FixtureObservations(g, gps)
    bands = gio.compacted_bands(obs, mask)
    prior_blocks = np.repeat(g["prior_inv_block"][None].astype(np.float64), N, axis=0)
    mean = g["x0"]

    def advance(x_a, Pinv_a, date):             # prior only: kf_tools.py:165-166
        return mean.copy(), prior_blocks.copy()

    ref_iters = _iters_per_date(g["iter_log"])
    rows = list(kc.run_filter(gio.grid(g), obs.dates, bands, _s2_operators(gps, P, bands), mean,
                              prior_blocks, P, advance, mode=mode))
    assert [r[0].toordinal() for r in rows] == list(g["out_steps"])
    assert sum((r[3] for r in rows), []) == ref_iters
    idx = np.arange(P)
    for k, (t, x, Pinv, _) in enumerate(rows):
        # the reference solves in float32 (solvers.py:127-134); on these blocks
        # (cond ~1e5) its own result is reproducible to ~1e-2 worst case, ~1e-5 median
        ed = gio.rel_err(Pinv[:, idx, idx].ravel(), g["out_diag"][k])
        ex = gio.rel_err(x, g["out_x"][k])
        assert np.median(ed) < 1e-4 and ed.max() < 5e-3
        assert np.median(ex) < 2e-5 and ex.max() < 5e-2


def _iters_per_date(iter_log):
    out = []
    for n, _ in iter_log:
        if int(n) == 1:
            out.append(1)
        else:
            out[-1] = int(n)
    return out


@pytest.mark.parametrize("mode", ["sparse32", "dense64"])
def test_run_tip_vs_reference(mode):
    g = gio.load("run_tip")
    mask = g["mask"]
    P = 7
    N = int(mask.sum())
    gps = gio.emulators(g, GaussianProcess)
    obs = gio.FixtureObservations(g, gps)
    bands = gio.compacted_bands(obs, mask)
    mappers = [g["mapper_0"], g["mapper_1"]]
    q = g["q"]
    idx = np.arange(P)

    def advance(x_a, Pinv_a, date):             # kf_tools.py:292-314
        return kc.propagate_information_filter_lai(x_a, Pinv_a[:, idx, idx].ravel(), q)

    prior_blocks = np.repeat(g["prior_inv_block"][None].astype(np.float64), N, axis=0)
    rows = list(kc.run_filter(gio.grid(g), obs.dates, bands, _s2_operators(gps, P, bands, mappers),
                              g["x0"], prior_blocks, P, advance, mode=mode))
    assert sum((r[3] for r in rows), []) == _iters_per_date(g["iter_log"])
    for k, (t, x, Pinv, _) in enumerate(rows):
        assert np.allclose(Pinv[:, idx, idx].ravel(), g["out_diag"][k], rtol=5e-6)
        assert np.allclose(x, g["out_x"][k], rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("mode", ["sparse32", "dense64"])
def test_run_linear_vs_reference(mode):
    g = gio.load("run_linear")
    mask = g["mask"]
    P = 7
    N = int(mask.sum())
    from kafka_inferenceengine_b200 import synthetic as syn
    op = syn.LinearBandOperator(g["row"], 0.0)
    obs = gio.FixtureObservations(g, [op])
    bands = gio.compacted_bands(obs, mask)
    q = g["q"]
    idx = np.arange(P)

    def make(date):
        m = bands[date][0][2]

        def lin(xp):
            J = np.where(m[:, None], op.row[None, :], 0.0)
            H0 = np.where(m, xp.reshape(N, P) @ op.row + op.offset, 0.0)
            return (H0.astype(np.float32).astype(np.float64), J.astype(np.float32).astype(np.float64))
        return [lin]

    def advance(x_a, Pinv_a, date):
        return kc.propagate_information_filter_lai(x_a, Pinv_a[:, idx, idx].ravel(), q)

    prior_blocks = np.repeat(g["prior_inv_block"][None].astype(np.float64), N, axis=0)
    rows = list(kc.run_filter(gio.grid(g), obs.dates, bands, make, g["x0"], prior_blocks, P,
                              advance, mode=mode))
    assert sum((r[3] for r in rows), []) == _iters_per_date(g["iter_log"])
    for k, (t, x, Pinv, _) in enumerate(rows):
        assert np.allclose(Pinv[:, idx, idx].ravel(), g["out_diag"][k], rtol=5e-6)
        assert np.allclose(x, g["out_x"][k], rtol=1e-4, atol=1e-6)


# ---- "next" rows: SAR Water-Cloud operator, Hessian correction, S2 ingest ---
def test_sar_forward_vs_reference():
    g = gio.load("kat_sar")
    for pol in ("VV", "VH"):
        s0, grad = kc.sar_forward(g["x"], 23.0, pol)
        assert np.allclose(s0, g["sigma0_" + pol], rtol=1e-13)
        assert np.allclose(grad, g["grad_" + pol], rtol=1e-12, atol=1e-18)


def test_hessian_correction_vs_reference():
    g = gio.load("kat_hessian")
    mask = g["mask"]
    gps = gio.emulators(g, GaussianProcess)
    mappers = [np.array([0, 1, 6, 2]), np.array([3, 4, 6, 5])]
    total = 0
    for b in range(2):
        obs = g["obs_mask_%d" % b][mask]
        inv = g["inv_var_%d" % b][mask]
        c = kc.hessian_correction(gps[b], g["x0"], inv, g["innov_%d" % b], obs, 7, mappers[b])
        assert np.allclose(c, g["corr_%d" % b], rtol=1e-10, atol=1e-12)
        assert not c[~obs].any()
        total = total + kc.hessian_correction(gps[0], g["x0"], inv, g["innov_%d" % b], obs, 7, mappers[b])
    assert np.allclose(total, g["corr_multi_gp0"], rtol=1e-9, atol=1e-10)


def test_s2_ingest_matches_reader_arithmetic():
    # Sentinel2_Observations.py:166-179 restated; checked against the literal numpy expressions
    rng = np.random.default_rng(0)
    raw = rng.integers(-50, 9000, size=(40, 50)).astype(np.int16)
    rho, inv, m = kc.s2_ingest(raw)
    mask = raw > 0
    want = np.where(mask, raw / 10000., 0)
    R = want * 0.05
    R[~mask] = 0.
    with np.errstate(divide="ignore"):
        want_inv = 1. / R ** 2
    assert np.array_equal(m, mask) and np.array_equal(rho, want) and np.array_equal(inv, want_inv)
    assert np.isinf(inv[~mask]).all()


def test_bhr_ingest_restates_the_reader_line_by_line():
    """oracle bhr_ingest against an independent evaluation of observations.py:288-310
    (the reader itself needs GDAL and the un-vendored BRDF_descriptors package)."""
    rng = np.random.default_rng(3)
    k = rng.uniform(0.0, 0.5, (3, 9, 11))
    mask = rng.random((9, 11)) < 0.8
    qa = rng.integers(0, 3, size=(9, 11))
    bhr, inv, m = kc.bhr_ingest(k, mask, qa)
    for r in range(9):
        for c in range(11):
            if not mask[r, c]:
                assert np.isnan(bhr[r, c]) and np.isinf(inv[r, c])
                continue
            v = (k[0, r, c] * 1.0 + k[1, r, c] * 0.189184) + k[2, r, c] * -1.377622
            assert bhr[r, c] == v
            if qa[r, c] > 1:
                assert np.isinf(inv[r, c])
            else:
                sig = max(2.5e-3, v * (0.05 if qa[r, c] == 0 else 0.07))
                assert inv[r, c] == 1.0 / sig ** 2


def test_s1_ingest_matches_the_unmodified_reader():
    """oracle s1_ingest against outputs of the unmodified S1Observations.get_band_data
    (fixture made by oracle/make_golden.py:kat_s1_ingest with GDAL's reproject_image
    replaced by a stand-in): bit-exact for float32 and float64 rasters, including the
    infinite weights of masked pixels, of an unmasked zero and of an underflowing square."""
    g = gio.load("kat_s1_ingest")
    for name in ("f32", "f64"):
        for pol in ("VV", "VH"):
            raw = g["%s_%s_raw" % (name, pol)]
            obs, inv, m = kc.s1_ingest(raw.copy(), missing_value=float(g["missing_value"]))
            assert obs.dtype == raw.dtype
            assert np.array_equal(obs, g["%s_%s_obs" % (name, pol)])
            assert np.array_equal(m, g["%s_%s_mask" % (name, pol)])
            assert inv.dtype == np.float64
            assert np.array_equal(inv.ravel(), g["%s_%s_inv_var" % (name, pol)])
            assert np.isinf(inv[~m]).all() and np.isinf(inv[3, 5]) and np.isinf(inv[4, 6]) and m[3, 5] and m[4, 6]


def test_s2_and_bhr_ingest_match_the_unmodified_readers():
    """oracle s2_ingest / bhr_ingest against outputs of the unmodified
    Sentinel2Observations.get_band_data and BHRObservations.get_band_data (fixture made
    by oracle/make_golden.py:kat_reader_ingest with stand-ins for GDAL and the
    un-vendored packages): bit-exact, dtypes included."""
    g = gio.load("kat_reader_ingest")
    for name in ("i16", "u16", "f32"):
        rho, inv, m = kc.s2_ingest(g["s2_%s_raw" % name])
        want = g["s2_%s_obs" % name]
        assert rho.dtype == want.dtype and np.array_equal(rho, want)
        assert inv.dtype == np.float64 and np.array_equal(inv.ravel(), g["s2_%s_inv_var" % name])
        assert np.array_equal(m, g["s2_%s_mask" % name])
    for name in ("f32", "f64"):
        bhr, inv, m = kc.bhr_ingest(g["bhr_%s_kernels" % name], g["bhr_%s_mask_in" % name], g["bhr_%s_qa" % name])
        assert np.array_equal(bhr, g["bhr_%s_obs" % name], equal_nan=True)
        assert np.array_equal(inv.ravel(), g["bhr_%s_inv_var" % name])
        assert np.array_equal(m, g["bhr_%s_mask" % name])


def test_output_scatter_matches_the_unmodified_writer():
    """oracle scatter_param / scatter_unc against the arrays the unmodified
    KafkaOutput.dump_data hands to GDAL's WriteArray (fixture kat_output_stage)."""
    g = gio.load("kat_output_stage")
    P = g["blocks"].shape[1]
    diag = g["blocks"][:, np.arange(P), np.arange(P)].ravel()
    for ii in range(P):
        assert np.array_equal(kc.scatter_param(g["x"], ii, P, g["mask"]), g["mean_%d" % ii])
        assert np.array_equal(kc.scatter_unc(diag, ii, P, g["mask"]), g["unc_%d" % ii])


def _indefinite_inputs(g):
    """Compacted per-band rows of the solve_indefinite fixture as the oracle takes them."""
    mask = g["mask"]
    ys, ws, ms = [], [], []
    for b in range(g["obs"].shape[0]):
        y, r, m = kc.gather_band(g["obs"][b], g["inv_var"][b], g["obs_mask"][b], mask)
        ys.append(y)
        ws.append(kc.gate_inverse_variance(r, m))
        ms.append(m)
    return ys, ws, [g["H0"][b].astype(np.float64) for b in range(len(ys))], \
        [g["J"][b].astype(np.float64) for b in range(len(ys))]


def test_indefinite_blocks_vs_reference():
    """Forecast blocks that are not positive definite (every third pixel): the unmodified reference's
    sparse LU returns the solution of the equations for them like for any other pixel
    (solvers.py:133-134, kf_tools.py:93-94), and so do both oracles -- the rule the CUDA path is held
    to (tests/test_gpu_solve.py, test_gpu_advance.py)."""
    g = gio.load("solve_indefinite")
    P = 7
    blocks = g["blocks"]
    N = blocks.shape[0]
    ev = np.linalg.eigvalsh(blocks)
    indef = ev.min(axis=1) < 0
    assert indef.sum() == len(range(0, N, 3)) and (np.abs(ev).max(axis=1) / np.abs(ev).min(axis=1)).max() < 1e3
    ys, ws, H0s, Js = _indefinite_inputs(g)
    xd, Ad, innov_d, fwd_d = kc.solve_dense64(ys, ws, H0s, Js, g["x_lin"], g["x_f"], blocks)
    xs, As, _, _ = kc.solve_sparse32(ys, ws, H0s, Js, g["x_lin"], g["x_f"], kc.blocks_to_sparse(blocks))
    # most of those systems are still indefinite after the observations have been added
    still = np.linalg.eigvalsh(Ad).min(axis=1) < 0
    assert still.sum() >= 12 and not still[~indef].any()
    idx = np.arange(P)
    assert np.allclose(Ad[:, idx, idx].ravel(), g["ref_A_diag"], rtol=3e-6, atol=1e-5)
    # float32 LU of the reference: errors of the order of cond(A) * 6e-8
    cond = np.abs(np.linalg.eigvalsh(Ad)).max(axis=1) / np.abs(np.linalg.eigvalsh(Ad)).min(axis=1)
    err = np.abs(xd - g["ref_x"]).reshape(N, P).max(axis=1) / np.abs(g["ref_x"]).reshape(N, P).max(axis=1)
    assert (err < 2e-6 * np.maximum(cond, 10.0)).all(), (err / cond).max()
    assert np.allclose(xs, g["ref_x"], rtol=2e-3, atol=1e-4)
    assert np.allclose(innov_d, g["ref_innovations"], rtol=1e-6, atol=1e-6)
    assert np.allclose(fwd_d, g["ref_fwd"], rtol=2e-3, atol=2e-3)
    # prior blend on the same forecast blocks
    xb, Cb = kc.blend_prior(g["prior_mean"], g["prior_inv_blocks"], g["x_f"], blocks)
    assert np.allclose(Cb[:, idx, idx].ravel(), g["ref_blend_C_diag"], rtol=1e-6)
    evc = np.linalg.eigvalsh(Cb)
    assert (evc[indef].min(axis=1) < 0).all()
    condc = np.abs(evc).max(axis=1) / np.abs(evc).min(axis=1)
    errb = np.abs(xb - g["ref_blend_x"]).reshape(N, P).max(axis=1) / np.abs(g["ref_blend_x"]).reshape(N, P).max(axis=1)
    assert (errb < 2e-6 * np.maximum(condc, 10.0)).all(), (errb / condc).max()
    # exact propagator on the same blocks (kf_tools.py:208-245: a float64 spsolve)
    _, want = kc.propagate_information_filter_exact(g["x_f"], blocks, g["q"])
    scale = np.abs(g["ref_exact_blocks"]).max(axis=(1, 2), keepdims=True)
    assert np.max(np.abs(want - g["ref_exact_blocks"]) / scale) < 1e-12
