"""Counterpart of kafka/inference/kf_tools.py: state propagation, prior blending
and the TIP prior, same function names and signatures, batched CUDA kernels
(csrc/advance.cu, csrc/solve.cu) behind them.

Inputs may be what the reference passes (numpy vectors, scipy block-diagonal
matrices) or the device containers this package returns (``DeviceVector``,
``BlockDiag``); outputs are always device containers, which quack like the
reference's return values (``np.asarray``, ``.diagonal()``, ``.dot``, ``+``, ``*1.``).
"""
import logging

import numpy as np
import scipy.sparse as sp
import torch

from .. import _lib
from ..device import BlockDiag, DeviceVector, default_device

LOG = logging.getLogger(__name__)

__all__ = ["NoHessianMethod", "band_selecta", "hessian_correction", "hessian_correction_multiband",
           "hessian_correction_pixel", "tip_prior", "tip_prior_full", "blend_prior",
           "propagate_and_blend_prior", "propagate_standard_kalman",
           "propagate_information_filter_SLOW", "propagate_information_filter_approx_SLOW",
           "propagate_information_filter_LAI", "propagate_information_filter_carry_slot",
           "no_propagation", "DiagonalMatrix", "IdentityModel"]


class NoHessianMethod(Exception):
    """kf_tools.py:13-17."""

    def __init__(self, message):
        self.message = message


def band_selecta(band):
    """kf_tools.py:19-23."""
    if band == 0:
        return np.array([0, 1, 6, 2])
    return np.array([3, 4, 6, 5])


def _hessian_mapper(tab, band, nparams, band_mapper=None):
    """State slots feeding the emulator of ``band``: the whole state when the
    emulator takes it all, else ``band_mapper[band]``, else the reference's
    hard-coded TIP selector (kf_tools.py:19-23)."""
    if band_mapper is not None:
        return np.asarray(band_mapper[band], dtype=np.int32)
    if tab.n_dim == nparams:
        return np.arange(nparams, dtype=np.int32)
    return band_selecta(band).astype(np.int32)


def _hessian_correction_device(gps, x0, ws, innovations, nparams, mappers, device, a_inout=None, want_corr=True):
    """kafka_gp_hessian_correction for a list of bands.  ``ws`` / ``innovations``
    are (n_bands, ld) device tensors (w gated to 0 where unobserved)."""
    from .utils import emulator_table
    lib = _lib.get_lib()
    x = DeviceVector.from_interleaved(x0, nparams, device)
    nb = len(gps)
    tabs = [emulator_table(g, device) for g in gps]
    arr = (_lib.GpBand * nb)()
    for b, (tab, mp) in enumerate(zip(tabs, mappers)):
        if len(mp) != tab.n_dim:
            raise ValueError("band %d: mapper has %d entries for a %d-input emulator" % (b, len(mp), tab.n_dim))
        arr[b].table, arr[b].n_train, arr[b].n_dim = tab.table.data_ptr(), tab.n_train, tab.n_dim
        for d, slot in enumerate(mp):
            arr[b].mapper[d] = int(slot)
    corr = BlockDiag.empty(x.n_pix, nparams, device) if want_corr else None
    with torch.cuda.device(device):
        _lib.check(lib.kafka_gp_hessian_correction(_lib.ptr(x.data), x.ld, x.n_pix, nparams, arr, nb,
                                                   _lib.ptr(innovations), _lib.ptr(ws),
                                                   _lib.ptr(corr.data) if corr is not None else None,
                                                   _lib.ptr(a_inout.data) if a_inout is not None else None,
                                                   _lib.stream_ptr()), "kafka_gp_hessian_correction")
    return corr


def _compact_rows(rasters_or_vectors, state_mask, n, ld, device, gate=None):
    out = torch.zeros((len(rasters_or_vectors), ld), dtype=torch.float64, device=device)
    sm = np.asarray(state_mask, dtype=bool)
    for b, r in enumerate(rasters_or_vectors):
        v = r.diagonal()[sm.flatten()] if sp.issparse(r) else np.asarray(r, dtype=np.float64)
        v = v[sm] if v.shape == sm.shape else v.reshape(-1)
        if gate is not None:
            v = np.where(gate[b], v, 0.0)
        out[b, :n] = torch.from_numpy(np.ascontiguousarray(v, dtype=np.float64)).to(device)
    return out


def hessian_correction_pixel(gp, x0, C_obs_inv, innovation, band, nparams):
    """kf_tools.py:26-34 for one pixel (through the batched kernel)."""
    x0 = np.asarray(x0, dtype=np.float64).reshape(-1)
    m = np.ones((1, 1), dtype=bool)
    R = sp.diags([float(C_obs_inv)])
    return hessian_correction(gp, x0, R, np.array([innovation]), m, m, band, nparams).blocks()[0]


def hessian_correction(gp, x0, R_mat, innovation, mask, state_mask, band, nparams, band_mapper=None):
    """kf_tools.py:37-60: block-diagonal second-order correction of one band,
    ``d2h/dx2 * R^-1 * innovation`` per observed pixel, zero elsewhere."""
    if not hasattr(gp, "inputs"):
        return 0.                                        # no Hessian available (kf_tools.py:41-44)
    from .utils import emulator_table
    device = default_device()
    sm = np.asarray(state_mask, dtype=bool)
    n = int(sm.sum())
    from ..device import padded
    ld = padded(n)
    obs = np.asarray(mask)[sm].astype(bool).reshape(-1)
    ws = _compact_rows([R_mat], sm, n, ld, device, gate=[obs])
    innov = _compact_rows([np.asarray(innovation, dtype=np.float64).reshape(-1)], sm, n, ld, device)
    tab = emulator_table(gp, device)
    return _hessian_correction_device([gp], x0, ws, innov, nparams, [_hessian_mapper(tab, band, nparams, band_mapper)],
                                      device)


def hessian_correction_multiband(gp, x0, R_mats, innovations, masks, state_mask, n_bands, nparams,
                                 band_mapper=None):
    """kf_tools.py:63-72: sum of the per-band corrections.  ``gp`` may be one
    emulator (what the reference passes, linear_kf.py:314) or one per band."""
    from .utils import emulator_table
    device = default_device()
    sm = np.asarray(state_mask, dtype=bool)
    n = int(sm.sum())
    from ..device import padded
    ld = padded(n)
    gps = list(gp) if isinstance(gp, (list, tuple)) else [gp] * n_bands
    obs = [np.asarray(m)[sm].astype(bool).reshape(-1) for m in masks]
    ws = _compact_rows(list(R_mats), sm, n, ld, device, gate=obs)
    innov = _compact_rows([np.asarray(v, dtype=np.float64).reshape(-1) for v in innovations], sm, n, ld, device)
    mappers = [_hessian_mapper(emulator_table(g, device), b, nparams, band_mapper) for b, g in enumerate(gps)]
    return _hessian_correction_device(gps, x0, ws, innov, nparams, mappers, device)


def tip_prior():
    """kf_tools.py:99-116 (covariance assembled in float32, as the reference)."""
    sigma = np.array([0.12, 0.7, 0.0959, 0.15, 1.5, 0.2, 0.5])
    x0 = np.array([0.17, 1.0, 0.1, 0.7, 2.0, 0.18, np.exp(-0.5 * 1.5)])
    little_p = np.diag(sigma ** 2).astype(np.float32)
    little_p[5, 2] = 0.8862 * 0.0959 * 0.2
    little_p[2, 5] = 0.8862 * 0.0959 * 0.2
    inv_p = np.linalg.inv(little_p)
    return x0, little_p, inv_p


def tip_prior_full(prior):
    """kf_tools.py:123-133: TIP prior for ``prior['n_pixels']`` pixels."""
    x_prior, _, c_inv_prior = tip_prior()
    n_pixels = int(prior['n_pixels'])
    mean = DeviceVector.broadcast(x_prior, n_pixels)
    return mean, BlockDiag.from_block(c_inv_prior.astype(np.float32), n_pixels)


# --------------------------------------------------------------------------
class IdentityModel(object):
    """``sp.eye(n)`` stand-in for the trajectory model (linear_kf.py:123-129)."""

    def __init__(self, n):
        self.shape = (n, n)

    def dot(self, x):
        return x

    def diagonal(self):
        return np.ones(self.shape[0])


class DiagonalMatrix(object):
    """Diagonal matrix kept as a device vector (trajectory uncertainty Q,
    linear_kf.py:131-146)."""

    def __init__(self, diag, n_params, device=None):
        self.vec = DeviceVector.from_interleaved(diag, n_params, device)
        n = len(self.vec)
        self.shape = (n, n)

    def diagonal(self):
        return np.asarray(self.vec)


def _infer_n_params(A):
    A = sp.coo_matrix(A)
    n = A.shape[0]
    bw = int(np.abs(A.row - A.col).max()) if A.nnz else 0
    for P in range(bw + 1, _lib.MAX_PARAMS + 1):
        if n % P == 0 and np.array_equal(A.row // P, A.col // P):
            return P
    raise ValueError("cannot infer the block size of the matrix (max %d)" % _lib.MAX_PARAMS)


def _blocks(A, n_params=None, n_pix=None, device=None):
    if isinstance(A, BlockDiag):
        return A
    if isinstance(A, np.ndarray) and A.ndim == 2 and A.shape[0] == A.shape[1] and A.shape[0] <= _lib.MAX_PARAMS \
            and n_params in (None, A.shape[0]):
        return BlockDiag.from_blocks(A[None], device)
    if n_params is None:
        n_params = _infer_n_params(A)
    return BlockDiag.from_scipy(A, n_params, device)


def _diag_vector(Q, n_params, device):
    """Main diagonal of Q as a planar DeviceVector."""
    if isinstance(Q, DiagonalMatrix):
        return Q.vec
    if isinstance(Q, DeviceVector):
        return Q
    if sp.issparse(Q):
        d = Q.diagonal()
    else:
        Q = np.asarray(Q, dtype=np.float64)
        if Q.ndim == 2:
            if np.count_nonzero(Q - np.diag(np.diagonal(Q))):
                raise NotImplementedError("only diagonal Q / M matrices are supported on the device")
            d = np.diagonal(Q)
        else:
            d = Q
    return DeviceVector.from_interleaved(d, n_params, device)


def _model_diag(M, n_params, n, device):
    """None for the identity, else the diagonal of M as a DeviceVector."""
    if M is None or isinstance(M, IdentityModel):
        return None
    d = M.diagonal() if sp.issparse(M) else np.asarray(M)
    if not sp.issparse(M) and d.ndim == 2:
        if np.count_nonzero(d - np.diag(np.diagonal(d))):
            raise NotImplementedError("only diagonal trajectory models are supported on the device")
        d = np.diagonal(d)
    elif sp.issparse(M) and M.nnz != np.count_nonzero(d):
        raise NotImplementedError("only diagonal trajectory models are supported on the device")
    if np.all(d == 1.0):
        return None
    return DeviceVector.from_interleaved(d, n_params, device)


def _apply_model(x, m):
    return x if m is None else DeviceVector(x.data * m.data, x.n_pix, x.n_params)


# --------------------------------------------------------------------------
def blend_prior(prior_mean, prior_cov_inverse, x_forecast, P_forecast_inverse):
    """kf_tools.py:75-96, INCLUDING its cross-weighted right-hand side
    ``P_f^-1 mu_prior + C_prior^-1 x_f`` (kf_tools.py:90); solved per pixel in fp64
    by Cholesky instead of one float32 sparse LU."""
    lib = _lib.get_lib()
    Pf = _blocks(P_forecast_inverse)
    P, n = Pf.n_params, Pf.n_pix
    dev = Pf.device
    Pf = Pf.materialise()
    xf = DeviceVector.from_interleaved(x_forecast, P, dev)
    Cp = _blocks(prior_cov_inverse, P, n, dev)
    flags = 0
    if Cp.is_broadcast and isinstance(prior_mean, np.ndarray) and prior_mean.size == P:
        mu_t = torch.from_numpy(np.ascontiguousarray(prior_mean, dtype=np.float64)).to(dev)
        cp_t = Cp.data
        flags = _lib.PRIOR_BROADCAST
    else:
        mu_t = DeviceVector.from_interleaved(prior_mean, P, dev).data
        cp_t = Cp.materialise().data
    x_out = DeviceVector.empty(n, P, dev)
    c_out = BlockDiag.empty(n, P, dev)
    n_bad = torch.zeros(2, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.kafka_blend_prior(_lib.ptr(mu_t), _lib.ptr(cp_t), _lib.ptr(xf.data), _lib.ptr(Pf.data),
                                         Pf.ld, n, P, _lib.ptr(x_out.data), _lib.ptr(c_out.data),
                                         _lib.ptr(n_bad), flags, _lib.stream_ptr()), "kafka_blend_prior")
        _warn_bad(n_bad, "blend_prior: %d pixels have a singular combined inverse covariance; "
                         "their mean is left at the forecast")
    return x_out, c_out


def _warn_bad(n_bad, message):
    """Read a kernel's failed-pixel counter back (kafka_read_small: no device->host memcpy) and
    log it the way do_all_bands does.  Blocks that are merely not positive definite are solved by
    elimination with partial pivoting, as the reference's LU solves them, and are not counted."""
    n = int(_lib.read_small(n_bad)[0])
    if n:
        LOG.warning(message, n)
    return n


def propagate_and_blend_prior(x_analysis, P_analysis, P_analysis_inverse, M_matrix, Q_matrix,
                              prior=None, state_propagator=None, date=None):
    """kf_tools.py:136-171, same dispatch."""
    if state_propagator is not None:
        x_forecast, P_forecast, P_forecast_inverse = state_propagator(
            x_analysis, P_analysis, P_analysis_inverse, M_matrix, Q_matrix)
    if prior is not None:
        prior_mean, prior_cov_inverse = prior.process_prior(date, inv_cov=True)
    if prior is not None and state_propagator is not None:
        x_combined, combined_cov_inv = blend_prior(prior_mean, prior_cov_inverse, x_forecast, P_forecast_inverse)
        return x_combined, None, combined_cov_inv
    elif prior is not None:
        return prior_mean, None, prior_cov_inverse
    elif state_propagator is not None:
        return x_forecast, P_forecast, P_forecast_inverse
    return None, None, None


def propagate_standard_kalman(x_analysis, P_analysis, P_analysis_inverse, M_matrix, Q_matrix,
                              prior=None, state_propagator=None, date=None):
    """kf_tools.py:174-205: x_f = M x_a, P_f = P_a + Q, no inverse."""
    lib = _lib.get_lib()
    Pa = _blocks(P_analysis).materialise()
    P, n, dev = Pa.n_params, Pa.n_pix, Pa.device
    xa = DeviceVector.from_interleaved(x_analysis, P, dev)
    m = _model_diag(M_matrix, P, n * P, dev)
    q = _diag_vector(Q_matrix, P, dev)
    x_f = DeviceVector.empty(n, P, dev)
    P_f = BlockDiag.empty(n, P, dev)
    with torch.cuda.device(dev):
        _lib.check(lib.kafka_advance_standard(_lib.ptr(xa.data), _lib.ptr(Pa.data), _lib.ptr(m.data) if m else None,
                                              _lib.ptr(q.data), Pa.ld, n, P, _lib.ptr(x_f.data), _lib.ptr(P_f.data),
                                              _lib.stream_ptr()), "kafka_advance_standard")
    return x_f, P_f, None


def propagate_information_filter_SLOW(x_analysis, P_analysis, P_analysis_inverse, M_matrix, Q_matrix,
                                      prior=None, state_propagator=None, date=None):
    """kf_tools.py:208-245: P_f^-1 = (I + P_a^-1 Q)^-1 P_a^-1, one Cholesky per pixel."""
    lib = _lib.get_lib()
    Pi = _blocks(P_analysis_inverse).materialise()
    P, n, dev = Pi.n_params, Pi.n_pix, Pi.device
    xa = DeviceVector.from_interleaved(x_analysis, P, dev)
    x_f = _apply_model(xa, _model_diag(M_matrix, P, n * P, dev))
    q = _diag_vector(Q_matrix, P, dev)
    out = BlockDiag.empty(n, P, dev)
    n_bad = torch.zeros(2, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.kafka_advance_exact(_lib.ptr(Pi.data), _lib.ptr(q.data), Pi.ld, n, P, _lib.ptr(out.data),
                                           _lib.ptr(n_bad), _lib.stream_ptr()), "kafka_advance_exact")
        _warn_bad(n_bad, "propagate_information_filter_SLOW: %d pixels have a singular inflated block; "
                         "their inverse covariance is carried over unchanged")
    return x_f, None, out


def propagate_information_filter_approx_SLOW(x_analysis, P_analysis, P_analysis_inverse, M_matrix, Q_matrix,
                                             prior=None, state_propagator=None, date=None):
    """kf_tools.py:247-289: diagonal inflation m/(1+m q), off-diagonals dropped."""
    lib = _lib.get_lib()
    Pi = _blocks(P_analysis_inverse).materialise()
    P, n, dev = Pi.n_params, Pi.n_pix, Pi.device
    xa = DeviceVector.from_interleaved(x_analysis, P, dev)
    x_f = _apply_model(xa, _model_diag(M_matrix, P, n * P, dev))
    q = _diag_vector(Q_matrix, P, dev)
    out = BlockDiag.empty(n, P, dev)
    with torch.cuda.device(dev):
        _lib.check(lib.kafka_advance_approx(_lib.ptr(Pi.data), _lib.ptr(q.data), Pi.ld, n, P, _lib.ptr(out.data),
                                            _lib.stream_ptr()), "kafka_advance_approx")
    return x_f, None, out


def propagate_information_filter_carry_slot(x_analysis, P_analysis, P_analysis_inverse, M_matrix, Q_matrix,
                                            prior_mean, prior_inv, carry_slot):
    """Generalisation of ``propagate_information_filter_LAI`` to any prior block
    and carried slot (see include/kafka_b200.h: kafka_advance_carry_slot)."""
    lib = _lib.get_lib()
    prior_mean = np.ascontiguousarray(prior_mean, dtype=np.float64)
    prior_inv = np.ascontiguousarray(prior_inv, dtype=np.float64)
    P = prior_mean.size
    Pi = _blocks(P_analysis_inverse, P).materialise()
    n, dev = Pi.n_pix, Pi.device
    xa = DeviceVector.from_interleaved(x_analysis, P, dev)
    xa = _apply_model(xa, _model_diag(M_matrix, P, n * P, dev))
    q = _diag_vector(Q_matrix, P, dev)
    x_f = DeviceVector.empty(n, P, dev)
    out = BlockDiag.empty(n, P, dev)
    with torch.cuda.device(dev):
        _lib.check(lib.kafka_advance_carry_slot(_lib.ptr(xa.data), _lib.ptr(Pi.data), _lib.ptr(q.data), Pi.ld, n, P,
                                                int(carry_slot), _lib.np_ptr(prior_mean), _lib.np_ptr(prior_inv),
                                                _lib.ptr(x_f.data), _lib.ptr(out.data), _lib.stream_ptr()),
                   "kafka_advance_carry_slot")
    return x_f, None, out


def propagate_information_filter_LAI(x_analysis, P_analysis, P_analysis_inverse, M_matrix, Q_matrix,
                                     prior=None, state_propagator=None, date=None):
    """kf_tools.py:292-314: reset to the TIP prior except slot 6 (TLAI), whose
    precision is inflated by Q; 7 parameters, as in the reference."""
    x_prior, _, c_inv_prior = tip_prior()
    return propagate_information_filter_carry_slot(x_analysis, P_analysis, P_analysis_inverse, M_matrix,
                                                   Q_matrix, x_prior, c_inv_prior, 6)


def no_propagation(x_analysis, P_analysis, P_analysis_inverse, M_matrix, Q_matrix,
                   prior=None, state_propagator=None, date=None):
    """kf_tools.py:316-353: return the TIP prior (inputs unused)."""
    x_prior, _, c_inv_prior = tip_prior()
    n_pixels = len(x_analysis) // 7
    x_f = DeviceVector.broadcast(x_prior, n_pixels)
    return x_f, None, BlockDiag.from_block(c_inv_prior.astype(np.float32), n_pixels)
