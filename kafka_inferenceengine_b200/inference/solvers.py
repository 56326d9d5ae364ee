"""Counterpart of kafka/inference/solvers.py: the linearised variational update
with SUPPLIED operators, same signatures and return tuples as the reference,
solved per pixel on the GPU (csrc/solve.cu) instead of one float32 sparse LU.

``H_matrix`` arguments are what the reference's operator builders return: a tuple
``(H0, H)`` with ``H`` an ``N x (n_params*N)`` sparse matrix whose row ``i`` only
touches the columns of pixel ``i`` (kafka/inference/utils.py:213).
"""
import logging

import numpy as np
import scipy.sparse as sp
import torch

from .. import engine
from ..device import BlockDiag, DeviceBands, DeviceVector, PixelIndex, default_device

LOG = logging.getLogger(__name__ + ".solvers")

__all__ = ["variational_kalman", "variational_kalman_multiband", "sort_band_data"]


def _rows_of(H, n_pix, n_params):
    """Dense per-pixel rows (n_params, n_pix) of a block-structured sparse operator."""
    Hc = sp.coo_matrix(H)
    if Hc.shape != (n_pix, n_pix * n_params):
        raise ValueError("operator has shape %s, expected %s" % (Hc.shape, (n_pix, n_pix * n_params)))
    slot = Hc.col - n_params * Hc.row
    if Hc.nnz and (slot.min() < 0 or slot.max() >= n_params):
        raise ValueError("operator row touches columns of another pixel")
    J = np.zeros((n_params, n_pix))
    np.add.at(J, (slot, Hc.row), Hc.data)
    return J


def sort_band_data(H_matrix, observations, uncertainty, mask, x0, x_forecast, state_mask):
    """solvers.py:81-96 on the host, for callers that use it directly: returns
    ``(H, H0, R, y', y_orig)``."""
    if len(H_matrix) == 2:
        H0, H_matrix_ = H_matrix
        non_linear = True
    else:
        H0, H_matrix_, non_linear = 0., H_matrix, False
    R = uncertainty.diagonal()[state_mask.flatten()]
    y = observations[state_mask]
    y = np.where(mask[state_mask], y, 0.)
    y_orig = y * 1.
    if non_linear:
        y = y + H_matrix_.dot(np.asarray(x0)) - H0
    return H_matrix_, H0, R, y, y_orig


def variational_kalman_multiband(observations_b, mask_b, state_mask, uncertainty_b, H_matrix_b, n_params,
                                 x0, x_forecast, P_forecast, P_forecast_inv, the_metadata_b,
                                 approx_diagonal=True):
    """solvers.py:100-145.  Returns ``(x_analysis, None, A, innovations,
    fwd_modelled)``: ``x_analysis`` a ``DeviceVector``, ``A`` a ``BlockDiag``,
    the two diagnostics as numpy arrays stacked over bands like the reference."""
    LOG.info("Solving")
    device = default_device()
    pix = PixelIndex(state_mask, device)
    n = pix.n_pix
    nb = len(observations_b)
    Band = engine.ObservationStager
    stager = Band(pix)

    class _D(object):
        __slots__ = ("observations", "uncertainty", "mask")

    data = []
    for b in range(nb):
        d = _D()
        d.observations, d.uncertainty, d.mask = observations_b[b], uncertainty_b[b], mask_b[b]
        data.append(d)
    bufs = stager.upload(data)
    ld = pix.ld
    descs = []
    keep = []
    for b in range(nb):
        Hm = H_matrix_b[b]
        if not isinstance(Hm, (tuple, list)) or len(Hm) != 2:
            # the reference cannot take a bare matrix either: len(sparse) raises
            # TypeError at solvers.py:83
            raise TypeError("H_matrix must be the (H0, H) tuple an operator builder returns")
        H0, H = Hm
        J = torch.zeros((n_params, ld), dtype=torch.float64, device=device)
        J[:, :n] = torch.from_numpy(_rows_of(H, n, n_params)).to(device)
        h0 = torch.zeros(ld, dtype=torch.float64, device=device)
        h0[:n] = torch.from_numpy(np.ascontiguousarray(H0, dtype=np.float64)).to(device)
        keep.append((J, h0))
        descs.append(("arrays", h0, J, np.arange(n_params)))
    ws = engine.SolveWorkspace(n, n_params, nb, device, want_fwd=True)
    x_prev = DeviceVector.from_interleaved(x0, n_params, device)
    x_f = DeviceVector.from_interleaved(x_forecast, n_params, device)
    Pinv = BlockDiag.coerce(P_forecast_inv, n_params, n, device)
    x_out = DeviceVector.empty(n, n_params, device)
    a_out = BlockDiag.empty(n, n_params, device)
    engine.solve_bands(ws, descs, bufs, x_prev, x_f, Pinv, x_out, a_out, want_fwd=True)
    innovations = np.asarray(DeviceBands(ws.innov, n))
    fwd = np.asarray(DeviceBands(ws.fwd, n))
    return x_out, None, a_out, innovations, fwd


def variational_kalman(observations, mask, state_mask, uncertainty, H_matrix, n_params,
                       x_forecast, P_forecast, P_forecast_inv, the_metadata, approx_diagonal=True):
    """solvers.py:41-78, the single-band legacy solver: relinearises around the
    FORECAST (solvers.py:56) and reports ``innovations = y - fwd`` (solvers.py:73)."""
    x, _, A, innov, fwd = variational_kalman_multiband([observations], [mask], state_mask, [uncertainty],
                                                       [H_matrix], n_params, x_forecast, x_forecast,
                                                       P_forecast, P_forecast_inv, [the_metadata])
    innov = (innov + np.asarray(H_matrix[0], dtype=np.float64)) - fwd
    return x, None, A, innov, fwd
