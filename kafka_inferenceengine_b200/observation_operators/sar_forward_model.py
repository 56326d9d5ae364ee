"""Counterpart of kafka/observation_operators/sar_forward_model.py: the Water
Cloud Model (WCM) SAR observation operator, evaluated on the GPU.

    tau        = exp(-2 B V / cos(theta))
    sigma_veg  = A V^E cos(theta) (1 - tau)
    sigma_soil = 10^((C + D SM) / 10)
    sigma_0    = sigma_veg + tau sigma_soil

``sar_observation_operator(x, theta, polarisation)`` keeps the reference's
signature and return values (sar_forward_model.py:13-113);
``create_sar_observation_operator`` is the operator descriptor ``LinearKalman``
takes (band 0 = VV, band 1 = VH, incidence angle 23 degrees as hard-coded at
sar_forward_model.py:146-156): inside the filter the model and its gradient are
evaluated per pixel inside the fused build+solve kernel (band kind
KAFKA_BAND_WCM), nothing is materialised.
"""
import logging

import numpy as np
import torch

from .. import _lib
from ..device import default_device, padded
from ..inference.utils import ObservationOperator

LOG = logging.getLogger(__name__)

__all__ = ["sar_observation_operator", "create_sar_observation_operator", "WCM_PARAMETERS", "wcm_parameters"]

# A, B, C, D, E per polarisation (sar_forward_model.py:60-61)
WCM_PARAMETERS = {'VV': [0.0846, 0.0615, -14.8465, 15.907, 1.],
                  'VH': [0.0795, 0.1464, -14.8332, 15.907, 0.]}


def wcm_parameters(polarisation, theta_deg):
    """{A, B, C, D, E, cos(theta)} for a polarisation and incidence angle (degrees)."""
    try:
        prm = list(WCM_PARAMETERS[polarisation])
    except KeyError:
        raise ValueError('Only VV and VH polarisations available!')      # sar_forward_model.py:66
    return np.array(prm + [np.cos(np.deg2rad(float(theta_deg)))], dtype=np.float64)


def sar_observation_operator(x, theta, polarisation):
    """sar_forward_model.py:13-113: ``(sigma_0, grad)`` for every row of ``x``
    (columns: vegetation descriptor, soil moisture).  ``theta`` (degrees) must be
    one angle for all rows, as everywhere in the reference."""
    x = np.atleast_2d(np.asarray(x, dtype=np.float64))
    th = np.unique(np.asarray(theta, dtype=np.float64))
    if th.size != 1:
        raise NotImplementedError("one incidence angle per call")
    if np.any(x[:, 0] <= 0.):
        raise ValueError("Negative LAI!")
    if np.any(x[:, 1] <= 0.):
        raise ValueError("Negative SM!")
    prm = wcm_parameters(polarisation, th[0])
    dev = default_device()
    n = x.shape[0]
    ld = padded(n)
    xin = torch.ones((2, ld), dtype=torch.float64, device=dev)
    xin[:, :n] = torch.from_numpy(np.ascontiguousarray(x[:, :2].T)).to(dev)
    s0 = torch.empty(ld, dtype=torch.float64, device=dev)
    grad = torch.empty((2, ld), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.get_lib().kafka_wcm_forward(_lib.ptr(xin), ld, n, _lib.np_ptr(prm), _lib.ptr(s0),
                                                    _lib.ptr(grad), _lib.stream_ptr()), "kafka_wcm_forward")
    sigma_0 = s0[:n].cpu().numpy()
    g = x * 0
    g[:, :2] = grad[:, :n].t().cpu().numpy()
    if np.any(np.isnan(sigma_0)):
        raise ValueError("Groan!")
    if np.any(np.isnan(g)):
        raise ValueError("More Groan!")
    return sigma_0, g


class _SarOperator(ObservationOperator):
    """Descriptor for the WCM operator: band -> polarisation, fixed incidence angle."""

    POLARISATIONS = ("VV", "VH")              # sar_forward_model.py:146-151
    THETA_DEG = 23.0                          # sar_forward_model.py:156

    def __init__(self):
        super().__init__("create_sar_observation_operator", "wcm")

    def band_parameters(self, band, metadata=None):
        if band >= len(self.POLARISATIONS):
            raise ValueError("the SAR operator knows bands 0 (VV) and 1 (VH), not %d" % band)
        theta = self.THETA_DEG
        if isinstance(metadata, dict) and "incidence_angle" in metadata:
            theta = float(metadata["incidence_angle"])
        return wcm_parameters(self.POLARISATIONS[band], theta)

    def __call__(self, n_params, forward_model, metadata, mask, state_mask, x_forecast, band):
        """The reference's builder call: returns (H0 float32, H csr float32)."""
        import scipy.sparse as sp
        LOG.info("Creating the ObsOp for band %d" % band)
        state_mask = np.asarray(state_mask, dtype=bool)
        obs = np.asarray(mask, dtype=bool)[state_mask]
        x = np.asarray(x_forecast, dtype=np.float64).reshape(-1, n_params)
        n = x.shape[0]
        LOG.info("Running SAR forward model")
        H0 = np.zeros(n, dtype=np.float32)
        J = np.zeros((n, n_params), dtype=np.float32)
        if obs.any():
            s0, g = sar_observation_operator(x[obs], self.band_parameters_theta(metadata),
                                             self.POLARISATIONS[band])
            H0[obs] = s0
            J[obs] = g
        LOG.info("Storing emulators in H matrix")
        indptr = np.concatenate([[0], np.cumsum(np.where(obs, n_params, 0))])
        cols = (n_params * np.flatnonzero(obs)[:, None] + np.arange(n_params)[None, :]).ravel()
        H = sp.csr_matrix((J[obs].ravel(), cols, indptr), shape=(n, n_params * n))
        return (H0, H)

    def band_parameters_theta(self, metadata):
        if isinstance(metadata, dict) and "incidence_angle" in metadata:
            return float(metadata["incidence_angle"])
        return self.THETA_DEG


create_sar_observation_operator = _SarOperator()
