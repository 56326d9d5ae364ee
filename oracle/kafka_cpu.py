"""CPU restatement of the reference's assimilation path (TEST INFRASTRUCTURE).

Every function cites the reference lines it follows (paths relative to
``/root/reference``).  Two flavours of the solver are kept side by side:

* ``*_sparse32``  -- oracle A': the algorithm exactly as the reference runs it,
  i.e. global ``scipy.sparse`` matrices, ``A``/``b`` cast to float32 and one
  SuperLU factorisation (``kafka/inference/solvers.py:125-134``).  This is also
  what ``bench.py`` times as the CPU baseline ("the reference's scipy path").
* ``*_dense64``   -- oracle B: the same equations in fp64 with one dense
  ``P x P`` block per pixel.  This is the <=1e-6 parity target for the CUDA path
  (the reference's own float32 LU is less accurate than that).

Pinned against (tests/test_oracle_golden.py): the reference's three known-answer
tests (``tests/test_kf.py:19-54``, ``tests/test_utils.py:18-38``), the
``Barrax_pivots.tif`` fixture, and outputs of the unmodified reference run under
``oracle/ref_shim.py`` (fixtures in ``tests/golden/``, made by
``oracle/make_golden.py``).  The GP emulator itself is parity-unpinned, see
``oracle/gp.py``.
"""
import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spl

TIP_SIGMA = np.array([0.12, 0.7, 0.0959, 0.15, 1.5, 0.2, 0.5])
TIP_MEAN = np.array([0.17, 1.0, 0.1, 0.7, 2.0, 0.18, np.exp(-0.5 * 1.5)])
# kafka/inference/utils.py:148-153 and kf_tools.py:19-23
TIP_STATE_MAPPERS = (np.array([0, 1, 6, 2]), np.array([3, 4, 6, 5]))


# --------------------------------------------------------------------------
# a1  state-mask compaction and indexing
# --------------------------------------------------------------------------
def compact_mask(state_mask):
    """Row-major rank of every True pixel (linear_kf.py:79-80; the order is the
    one ``arr[state_mask]`` produces, solvers.py:90-92)."""
    return np.flatnonzero(np.asarray(state_mask, dtype=bool).ravel())


def gather_band(observations, inv_var_diag, mask, state_mask):
    """solvers.py:90-93: R^-1, y (zeroed where unobserved) and the compacted
    observation mask for one band."""
    sm = np.asarray(state_mask, dtype=bool)
    r = np.asarray(inv_var_diag)[sm.ravel()]
    m = np.asarray(mask)[sm]
    y = np.where(m, np.asarray(observations)[sm], 0.0)
    return y, r, m


def scatter_param(x, ii, n_params, state_mask, dtype=np.float32):
    """input_output/observations.py:374-375: one parameter back to a raster."""
    out = np.zeros(state_mask.shape, dtype=dtype)
    out[state_mask] = x[ii::n_params]
    return out


def scatter_unc(diag_pinv, ii, n_params, state_mask, dtype=np.float32):
    """input_output/observations.py:392-393: 1/sqrt(diag(P^-1))."""
    out = np.zeros(state_mask.shape, dtype=dtype)
    out[state_mask] = 1.0 / np.sqrt(diag_pinv[ii::n_params])
    return out


# --------------------------------------------------------------------------
# a2  time grid
# --------------------------------------------------------------------------
def iterate_time_grid(time_grid, the_dates):
    """kafka/inference/utils.py:44-65."""
    dates = np.array(the_dates)
    start = time_grid[0]
    first = True
    for timestep in time_grid[1:]:
        sel = np.logical_and(dates >= start, dates < timestep)
        yield timestep, dates[sel], first
        first = False
        start = timestep


# --------------------------------------------------------------------------
# a3/a4/a6  observation operators  (dense per-pixel Jacobian rows)
# --------------------------------------------------------------------------
def emulator_operator(gp, x_prev, n_params, obs_mask, state_mapper=None,
                      round_f32=True):
    """``create_prosail_observation_operator`` (utils.py:181-219) or, with a
    ``state_mapper``, ``create_nonlinear_observation_operator`` (utils.py:130-177).

    Returns ``H0[N]`` and the N non-zero row segments ``J[N, P]`` of the sparse
    operator (row i holds columns ``P*i .. P*i+P-1``, utils.py:213).  Unobserved
    pixels keep an empty row and ``H0 = 0`` (utils.py:199-215).  ``round_f32``
    mirrors the float32 storage of both (utils.py:193-195).  ``run_emulator``'s
    de-duplication (utils.py:72,97-101) only avoids repeated evaluation and is
    value-neutral, so every observed pixel is simply evaluated.
    """
    N = x_prev.shape[0] // n_params
    xs = x_prev.reshape(N, n_params)
    cols = np.arange(n_params) if state_mapper is None else np.asarray(state_mapper)
    obs = np.asarray(obs_mask, dtype=bool)
    H0 = np.zeros(N)
    J = np.zeros((N, n_params))
    if obs.any():
        mu, dmu = gp.predict(xs[obs][:, cols], do_unc=False)
        H0[obs] = mu
        Jo = np.zeros((int(obs.sum()), n_params))
        Jo[:, cols] = dmu
        J[obs] = Jo
    if round_f32:
        H0 = H0.astype(np.float32).astype(np.float64)
        J = J.astype(np.float32).astype(np.float64)
    return H0, J


def rows_to_csr(J):
    """The (N x P*N) sparse operator the reference builds (utils.py:193,213)."""
    N, P = J.shape
    indptr = np.arange(0, N * P + 1, P)
    indices = np.arange(N * P)
    return sp.csr_matrix((J.ravel(), indices, indptr), shape=(N, N * P))


def blocks_to_sparse(blocks, dtype=None):
    """``block_diag`` fast path for equal blocks (utils.py:329-337)."""
    N, P, _ = blocks.shape
    return sp.bsr_matrix((blocks if dtype is None else blocks.astype(dtype),
                          np.arange(N), np.arange(N + 1)),
                         shape=(N * P, N * P)).tocsr()


def sparse_to_blocks(A, n_params):
    A = sp.bsr_matrix(A, blocksize=(n_params, n_params))
    N = A.shape[0] // n_params
    out = np.zeros((N, n_params, n_params), dtype=A.dtype)
    A.sort_indices()
    rows = np.repeat(np.arange(N), np.diff(A.indptr))
    assert np.all(rows == A.indices), "matrix is not block diagonal"
    out[rows] = A.data
    return out


# --------------------------------------------------------------------------
# next rows: SAR Water-Cloud operator, Hessian correction, S2 ingest
# --------------------------------------------------------------------------
WCM_PARAMETERS = {'VV': [0.0846, 0.0615, -14.8465, 15.907, 1.],
                  'VH': [0.0795, 0.1464, -14.8332, 15.907, 0.]}


def sar_forward(x, theta_deg, polarisation):
    """kafka/observation_operators/sar_forward_model.py:56-98, vectorised: returns
    sigma_0 and the gradient with respect to (V, SM)."""
    x = np.atleast_2d(np.asarray(x, dtype=np.float64))
    A, B, C, D, E = WCM_PARAMETERS[polarisation]
    mu = np.cos(np.deg2rad(theta_deg))
    V, SM = x[:, 0], x[:, 1]
    tau = np.exp(-2. * B / mu * V)
    sigma_veg = A * np.power(V, E) * mu * (1. - tau)
    sigma_surf = 10. ** ((C + D * SM) / 10.)
    sigma_0 = sigma_veg + tau * sigma_surf
    z = np.power(V, E)
    z1 = np.power(V, E - 1.)
    g0 = A * E * mu * z1 * (1. - tau) + 2. * A * B * z * tau - (
        (2. ** (1 / 10. * (C + D * SM) + 1.)) * (5. ** (1 / 10. * (C + D * SM)) * B * tau)) / mu
    g1 = D * np.log(10.) * tau * 10. ** (1. / 10. * (C + D * SM) - 1.)
    return sigma_0, np.stack([g0, g1], axis=1)


def sar_operator(x_prev, n_params, obs_mask, band, theta_deg=23.0, slots=(0, 1), round_f32=True):
    """create_sar_observation_operator (sar_forward_model.py:116-173) as dense rows:
    band 0 = VV, band 1 = VH, incidence angle 23 degrees (:146-156)."""
    N = x_prev.shape[0] // n_params
    xs = x_prev.reshape(N, n_params)
    obs = np.asarray(obs_mask, dtype=bool)
    H0 = np.zeros(N)
    J = np.zeros((N, n_params))
    if obs.any():
        s0, g = sar_forward(xs[obs][:, list(slots)], theta_deg, ("VV", "VH")[band])
        H0[obs] = s0
        Jo = np.zeros((int(obs.sum()), n_params))
        Jo[:, slots[0]] = g[:, 0]
        Jo[:, slots[1]] = g[:, 1]
        J[obs] = Jo
    if round_f32:
        H0 = H0.astype(np.float32).astype(np.float64)
        J = J.astype(np.float32).astype(np.float64)
    return H0, J


def hessian_correction(gp, x0, inv_var, innovation, obs_mask, n_params, mapper):
    """kf_tools.py:26-60 for one band on compacted vectors: per observed pixel
    big_ddH * C_obs_inv * innovation with big_ddH the emulator Hessian embedded
    through ``mapper``; zero blocks for unobserved pixels."""
    N = x0.shape[0] // n_params
    xs = x0.reshape(N, n_params)
    obs = np.asarray(obs_mask, dtype=bool)
    out = np.zeros((N, n_params, n_params))
    if obs.any():
        mapper = np.asarray(mapper)
        ddH = gp.hessian(xs[obs][:, mapper])
        big = np.zeros((int(obs.sum()), n_params, n_params))
        big[:, mapper[:, None], mapper[None, :]] = ddH
        out[obs] = big * (np.asarray(inv_var)[obs] * np.asarray(innovation)[obs])[:, None, None]
    return out


def s2_ingest(raw, ext_mask=None, divisor=10000.0, rel_sigma=0.05):
    """Sentinel2_Observations.py:166-179: mask, reflectance and inverse variance
    (inf where masked) from the raw raster.  Scaled integers become float64 at the
    division; a float32 raster stays float32 through every expression and only the
    inverse variance is widened when it is stored on the sparse diagonal (:177-178)."""
    mask = raw > 0
    if ext_mask is not None:
        mask = mask & (np.asarray(ext_mask) != 0)
    rho = np.where(mask, raw / divisor, 0)
    R = rho * rel_sigma
    R[np.logical_not(mask)] = 0.
    with np.errstate(divide="ignore"):
        inv_var = 1. / (R) ** 2
    return rho, inv_var.astype(np.float64), mask


def s1_ingest(backscatter, missing_value=-999.0, rel_sigma=0.05):
    """kafka/input_output/Sentinel1_Observations.py:107-130,147-149,176-187
    (S1Observations.get_band_data): mask, standard deviation and inverse variance of
    one sigma0 raster.  The arithmetic runs in the raster's own type (float32 as read
    from the file, or float64) and the inverse variance is widened to float64 when it
    is stored on the sparse diagonal (:184-186).  inf where masked, and where an
    unmasked backscatter is zero or its square underflows."""
    backscatter = np.asarray(backscatter)
    unc = backscatter * rel_sigma                              # :124
    mask = np.ones_like(backscatter, dtype=bool)               # :147
    mask[backscatter == missing_value] = False                 # :148
    R_mat = unc                                                # :180-181
    R_mat[np.logical_not(mask)] = 0.                           # :182
    with np.errstate(divide="ignore", over="ignore"):
        inv_var = 1. / (R_mat.ravel()) ** 2                    # :186
    return backscatter, inv_var.astype(np.float64).reshape(backscatter.shape), mask


def bhr_ingest(kernels, mask, qa_level):
    """kafka/input_output/observations.py:288-310 (BHRObservations.get_band_data):
    bi-hemispherical reflectance from the three BRDF kernel weights, its
    standard deviation from the quality level and the inverse variance (inf where
    the standard deviation is 0: masked pixels and quality levels above 1)."""
    to_BHR = np.array([1.0, 0.189184, -1.377622])
    mask = np.asarray(mask, dtype=bool)
    with np.errstate(invalid="ignore"):
        bhr = np.where(mask, kernels * to_BHR[:, None, None], np.nan).sum(axis=0)
        R_mat = np.zeros_like(bhr)
        R_mat[qa_level == 0] = np.maximum(2.5e-3, bhr[qa_level == 0] * 0.05)
        R_mat[qa_level == 1] = np.maximum(2.5e-3, bhr[qa_level == 1] * 0.07)
    R_mat[np.logical_not(mask)] = 0.
    with np.errstate(divide="ignore"):
        inv_var = 1. / (R_mat.ravel()) ** 2
    return bhr, inv_var.reshape(bhr.shape), mask


# --------------------------------------------------------------------------
# a8/a9  one linearised solve
# --------------------------------------------------------------------------
def solve_sparse32(ys, ws, H0s, Js, x_prev, x_f, Pinv_f_sparse):
    """oracle A': ``variational_kalman_multiband`` as written
    (solvers.py:100-145) with per-band ``sort_band_data`` (solvers.py:81-96).
    ``ws`` is R^-1 with masked entries already gated to 0 (SURVEY "inf*0")."""
    Hs = [rows_to_csr(J) for J in Js]
    yp = [y + H.dot(x_prev) - H0 for y, H, H0 in zip(ys, Hs, H0s)]
    H = sp.vstack(Hs)
    R = sp.diags(np.hstack(ws))
    y = np.hstack(yp)
    A = H.T.dot(R).dot(H) + Pinv_f_sparse
    b = H.T.dot(R).dot(y) + Pinv_f_sparse.dot(x_f)
    b = b.astype(np.float32)
    A = A.astype(np.float32)
    x = spl.splu(sp.csc_matrix(A)).solve(b)
    fwd = H.dot(x - x_f) + np.hstack(H0s)
    innovations = np.hstack(ys) - np.hstack(H0s)
    return x, A, innovations, fwd


def solve_dense64(ys, ws, H0s, Js, x_prev, x_f, Pinv_f_blocks):
    """oracle B: same equations, fp64, one block per pixel.
    A_i = sum_b w_bi J_bi J_bi^T + Pinv_i ;  b_i = sum_b w_bi y'_bi J_bi + Pinv_i x_f,i
    with y'_bi = y_bi + J_bi . x_prev,i - H0_bi  (solvers.py:95,125-126)."""
    N, P, _ = Pinv_f_blocks.shape
    xp = x_prev.reshape(N, P)
    xf = x_f.reshape(N, P)
    A = Pinv_f_blocks.astype(np.float64).copy()
    b = np.einsum("npq,nq->np", A, xf)
    for y, w, H0, J in zip(ys, ws, H0s, Js):
        yp = y + np.einsum("np,np->n", J, xp) - H0
        A += w[:, None, None] * J[:, :, None] * J[:, None, :]
        b += (w * yp)[:, None] * J
    x = np.linalg.solve(A, b[:, :, None])[:, :, 0]
    fwd = np.hstack([np.einsum("np,np->n", J, x - xf) + H0 for J, H0 in zip(Js, H0s)])
    innovations = np.hstack(ys) - np.hstack(H0s)
    return x.ravel(), A, innovations, fwd


# --------------------------------------------------------------------------
# a10  iterated linearisation
# --------------------------------------------------------------------------
def do_all_bands(operators, ys, ws, obs_masks, x_f, Pinv_f, n_params,
                 mode="dense64", tol=1e-3, min_iterations=2, max_iter_cap=25,
                 round_f32=True, fixed_iterations=None):
    """``LinearKalman.do_all_bands`` (linear_kf.py:245-323).

    ``operators[b]`` is a callable ``x_prev -> (H0[N], J[N,P])``.
    Returns ``x, A, innovations, n_solves, norms``.  ``fixed_iterations`` forces
    exactly that many solves (parity at fixed iteration count)."""
    x_prev = x_f * 1.0
    n_iter = 1
    norms = []
    if mode == "sparse32" and not sp.issparse(Pinv_f):
        Pinv_f = blocks_to_sparse(Pinv_f)
    while True:
        H0s, Js = [], []
        for op in operators:
            H0, J = op(x_prev)
            H0s.append(H0)
            Js.append(J)
        if mode == "sparse32":
            x, A, innov, fwd = solve_sparse32(ys, ws, H0s, Js, x_prev, x_f, Pinv_f)
        else:
            x, A, innov, fwd = solve_dense64(ys, ws, H0s, Js, x_prev, x_f, Pinv_f)
        norm = np.linalg.norm(x - x_prev) / float(len(x))          # linear_kf.py:293
        norms.append(norm)
        if fixed_iterations is not None:
            stop = n_iter >= fixed_iterations
        else:
            stop = (norm < tol and n_iter >= min_iterations) or n_iter > max_iter_cap
        x_prev = x * 1.0
        n_iter += 1
        if stop:
            break
    return x, A, innov, n_iter - 1, norms


# --------------------------------------------------------------------------
# a12-a17  propagation / priors
# --------------------------------------------------------------------------
def tip_prior():
    """kf_tools.py:99-116 (covariance built in float32, inverse in float32)."""
    little_p = np.diag(TIP_SIGMA ** 2).astype(np.float32)
    little_p[5, 2] = 0.8862 * 0.0959 * 0.2
    little_p[2, 5] = 0.8862 * 0.0959 * 0.2
    inv_p = np.linalg.inv(little_p)
    return TIP_MEAN.copy(), little_p, inv_p


def sail_prior():
    """kafka_test_S2.py:84-94."""
    mean = np.array([2.1, np.exp(-60. / 100.), np.exp(-7.0 / 100.), 0.1,
                     np.exp(-50 * 0.0176), np.exp(-100. * 0.002),
                     np.exp(-4. / 2.), 70. / 90., 0.5, 0.9])
    sigma = np.array([0.01, 0.2, 0.01, 0.05, 0.01, 0.01, 0.50, 0.1, 0.1, 0.1])
    inv_covar = np.diag(1. / sigma ** 2).astype(np.float32)
    return mean, sigma, inv_covar


def propagate_standard_kalman(x_a, P_a_diag_or_blocks, q_diag, m_diag=None):
    """kf_tools.py:174-205 on block storage: x_f = M x_a, P_f = P_a + Q."""
    x_f = x_a if m_diag is None else m_diag * x_a
    P = np.array(P_a_diag_or_blocks, dtype=np.float64, copy=True)
    N, n, _ = P.shape
    idx = np.arange(n)
    P[:, idx, idx] += q_diag.reshape(N, n)
    return x_f, P


def propagate_information_filter_exact(x_a, Pinv_blocks, q_diag):
    """kf_tools.py:208-245: P_f^-1 = (I + P_a^-1 Q)^-1 P_a^-1, per block."""
    N, n, _ = Pinv_blocks.shape
    Pinv = Pinv_blocks.astype(np.float64)
    S = Pinv * q_diag.reshape(N, 1, n)                 # P^-1 . diag(q)
    A = np.eye(n)[None] + S
    return x_a.copy(), np.linalg.solve(A, Pinv)


def propagate_information_filter_approx(x_a, Pinv_diag, q_diag):
    """kf_tools.py:247-289: diagonal m/(1+m q); off-diagonals dropped."""
    return x_a.copy(), Pinv_diag / (1.0 + Pinv_diag * q_diag)


def propagate_information_filter_lai(x_a, Pinv_diag, q_diag):
    """kf_tools.py:292-314 (hard-codes 7 parameters, slot 6 = TLAI).  The
    forecast precision blocks are stored float32 (kf_tools.py:312)."""
    mean, _, inv_p = tip_prior()
    N = x_a.shape[0] // 7
    x_f = np.tile(mean, N)
    x_f[6::7] = x_a[6::7]
    blocks = np.repeat(inv_p[None].astype(np.float64), N, axis=0)
    lai = 1.0 / ((1.0 / Pinv_diag[6::7]) + q_diag[6::7])
    blocks[:, 6, 6] = lai
    blocks = blocks.astype(np.float32).astype(np.float64)
    return x_f, blocks


def no_propagation(n_pixels):
    """kf_tools.py:316-353."""
    mean, _, inv_p = tip_prior()
    blocks = np.repeat(inv_p[None].astype(np.float64), n_pixels, axis=0)
    return np.tile(mean, n_pixels), blocks


def blend_prior(prior_mean, prior_inv_blocks, x_f, Pinv_f_blocks, mode="dense64"):
    """kf_tools.py:75-96, INCLUDING its cross-weighted right-hand side
    ``b = P_f^-1 mu_prior + C_prior^-1 x_f`` (kf_tools.py:90)."""
    N, n, _ = Pinv_f_blocks.shape
    C = Pinv_f_blocks.astype(np.float64) + prior_inv_blocks.astype(np.float64)
    b = (np.einsum("npq,nq->np", Pinv_f_blocks.astype(np.float64), prior_mean.reshape(N, n)) +
         np.einsum("npq,nq->np", prior_inv_blocks.astype(np.float64), x_f.reshape(N, n)))
    if mode == "sparse32":
        Cs = blocks_to_sparse(C)
        x = spl.splu(sp.csc_matrix(Cs)).solve(b.ravel().astype(np.float32))
        return x, C
    x = np.linalg.solve(C, b[:, :, None])[:, :, 0].ravel()
    return x, C


# --------------------------------------------------------------------------
# run loop  (linear_kf.py:171-242)
# --------------------------------------------------------------------------
def gate_inverse_variance(inv_var, obs_mask):
    """Masked observations carry R^-1 = inf (Sentinel2_Observations.py:175-178)
    and drop out of the reference only because their operator rows are empty;
    on a per-pixel formulation the weight is gated to zero instead."""
    return np.where(obs_mask, inv_var, 0.0)


def run_filter(time_grid, dates, band_data, make_operators, x0, Pinv0_blocks,
               n_params, advance, mode="dense64", round_f32=True,
               fixed_iterations=None):
    """``LinearKalman.run`` + ``assimilate_multiple_bands``.

    ``band_data[date]`` -> list of (y[N], w[N], obs_mask[N]) compacted per band;
    ``make_operators(date)`` -> list of callables x_prev -> (H0, J);
    ``advance(x_a, Pinv_a_blocks, date)`` -> (x_f, Pinv_f_blocks).
    Yields (timestep, x_a, Pinv_a_blocks, [n_solves per obs date])."""
    x_f, Pinv_f = x0.astype(np.float64), np.asarray(Pinv0_blocks, dtype=np.float64)
    x_a = Pinv_a = None
    for timestep, locate_times, is_first in iterate_time_grid(time_grid, dates):
        if not is_first:
            x_f, Pinv_f = advance(x_a, Pinv_a, timestep)
        iters = []
        if len(locate_times) == 0:
            x_a, Pinv_a = x_f, Pinv_f
        else:
            for step in locate_times:
                ys = [b[0] for b in band_data[step]]
                ws = [b[1] for b in band_data[step]]
                ms = [b[2] for b in band_data[step]]
                x_a, A, _, n, _ = do_all_bands(make_operators(step), ys, ws, ms, x_f, Pinv_f,
                                               n_params, mode=mode, round_f32=round_f32,
                                               fixed_iterations=(None if fixed_iterations is None
                                                                 else fixed_iterations[len(iters)]))
                if sp.issparse(A):
                    A = sparse_to_blocks(A, n_params)
                Pinv_a = np.asarray(A, dtype=np.float64)
                x_a = np.asarray(x_a, dtype=np.float64)
                x_f, Pinv_f = x_a * 1.0, Pinv_a * 1.0
                iters.append(n)
        yield timestep, x_a, Pinv_a, iters
