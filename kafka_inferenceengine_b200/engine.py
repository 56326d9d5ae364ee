"""Device-side plumbing shared by ``LinearKalman`` and the solver functions:
observation upload + compaction gather, band descriptors and the call into the
fused build+solve kernel.  No arithmetic happens here."""
import ctypes as C

import numpy as np
import scipy.sparse as sp
import torch

from . import _lib
from .device import BlockDiag, DeviceBands, DeviceVector, PixelIndex, padded, tri_count


class BandBuffers(object):
    """Compacted per-band observation rows on the device."""

    def __init__(self, n_bands, pix, device):
        self.n_bands = n_bands
        self.pix = pix
        ld = pix.ld
        self.y = torch.zeros((n_bands, ld), dtype=torch.float64, device=device)
        self.w = torch.zeros((n_bands, ld), dtype=torch.float64, device=device)
        self.m = torch.zeros((n_bands, ld), dtype=torch.uint8, device=device)


class ObservationStager(object):
    """Pinned host staging + device raster buffers, reused across dates, for the
    host -> device copy of one date's rasters (observations, inverse variance,
    mask) followed by the compaction gather (solvers.py:90-93)."""

    def __init__(self, pix):
        self.pix = pix
        self.device = pix.device
        self._cap = 0
        self.h2d_bytes = 0

    def _ensure(self, n_bands):
        if n_bands <= self._cap:
            return
        n = self.pix.n_raster
        self.h_obs = torch.empty((n_bands, n), dtype=torch.float64).pin_memory()
        self.h_inv = torch.empty((n_bands, n), dtype=torch.float64).pin_memory()
        self.h_msk = torch.empty((n_bands, n), dtype=torch.uint8).pin_memory()
        self.d_obs = torch.empty((n_bands, n), dtype=torch.float64, device=self.device)
        self.d_inv = torch.empty((n_bands, n), dtype=torch.float64, device=self.device)
        self.d_msk = torch.empty((n_bands, n), dtype=torch.uint8, device=self.device)
        self._cap = n_bands

    @staticmethod
    def _inv_var(unc, n):
        """Main diagonal of the reference's (ny*nx)^2 sparse inverse-variance
        matrix (Sentinel2_Observations.py:176-179), or a plain raster of it."""
        if sp.issparse(unc):
            return unc.diagonal()
        a = np.asarray(unc)
        if a.ndim == 2 and a.shape == (n, n):
            return np.diagonal(a)
        return a.reshape(-1)

    def upload(self, current_data, out=None):
        pix = self.pix
        nb = len(current_data)
        out = out or BandBuffers(nb, pix, self.device)
        on_device = all(isinstance(d.observations, torch.Tensor) and d.observations.is_cuda for d in current_data)
        if on_device:
            # inputs already resident in HBM (device rasters): gather only
            for b, d in enumerate(current_data):
                pix.gather_band(d.observations.reshape(-1), d.uncertainty.reshape(-1),
                                d.mask.reshape(-1).view(torch.uint8) if d.mask.dtype == torch.bool else d.mask.reshape(-1),
                                out.y[b], out.w[b], out.m[b])
            return out
        self._ensure(nb)
        n = pix.n_raster
        for b, d in enumerate(current_data):
            self.h_obs[b].numpy()[:] = np.asarray(d.observations, dtype=np.float64).reshape(-1)
            self.h_inv[b].numpy()[:] = self._inv_var(d.uncertainty, n)
            self.h_msk[b].numpy()[:] = np.asarray(d.mask).reshape(-1) != 0
        self.d_obs[:nb].copy_(self.h_obs[:nb], non_blocking=True)
        self.d_inv[:nb].copy_(self.h_inv[:nb], non_blocking=True)
        self.d_msk[:nb].copy_(self.h_msk[:nb], non_blocking=True)
        self.h2d_bytes += nb * n * 17
        for b in range(nb):
            pix.gather_band(self.d_obs[b], self.d_inv[b], self.d_msk[b], out.y[b], out.w[b], out.m[b])
        return out


class SolveWorkspace(object):
    """Scratch and diagnostic buffers of kafka_solve_bands, allocated once per
    (n_pix, n_params, n_bands) and reused across iterations and dates.  The
    state and precision outputs are allocated per call by the caller (they are
    handed out to the user and must not be overwritten later)."""

    def __init__(self, n_pix, n_params, n_bands, device, want_fwd=False):
        lib = _lib.get_lib()
        self.n_pix, self.n_params, self.n_bands = n_pix, n_params, n_bands
        self.device = device
        ld = padded(n_pix)
        self.ld = ld
        self.innov = torch.zeros((max(n_bands, 1), ld), dtype=torch.float64, device=device)
        self.fwd = torch.zeros((max(n_bands, 1), ld), dtype=torch.float64, device=device) if want_fwd else None
        self.norm2 = torch.zeros(1, dtype=torch.float64, device=device)
        self.n_bad = torch.zeros(1, dtype=torch.int32, device=device)
        self.scratch = torch.empty(int(lib.kafka_solve_scratch_doubles(n_pix)), dtype=torch.float64, device=device)


def make_solve_bands(descs, bufs):
    """ctypes array of kafka_solve_band from python descriptors.

    ``descs[b]`` is ("arrays", h0_row, jac_rows(D, ld), mapper) or
    ("linear", row, offset)."""
    nb = len(descs)
    arr = (_lib.SolveBand * max(nb, 1))()
    for b, d in enumerate(descs):
        sb = arr[b]
        sb.y = bufs.y[b].data_ptr()
        sb.w = bufs.w[b].data_ptr()
        if d[0] == "linear":
            sb.kind = _lib.BAND_LINEAR
            row = np.asarray(d[1], dtype=np.float64)
            for p in range(row.size):
                sb.row[p] = float(row[p])
            sb.offset = float(d[2])
        else:
            sb.kind = _lib.BAND_ARRAYS
            _, h0, jac, mapper = d
            sb.h0 = h0.data_ptr()
            sb.jac = jac.data_ptr()
            sb.n_dim = len(mapper)
            for k, slot in enumerate(mapper):
                sb.mapper[k] = int(slot)
    return arr


def solve_bands(ws, descs, bufs, x_prev, x_f, pinv_f, x_out, a_out, want_fwd=False):
    """One linearised update (kafka_solve_bands).  Results land in ``x_out``,
    ``a_out``, ``ws.innov``, ``ws.fwd``; ``ws.norm2`` / ``ws.n_bad`` stay on the
    device until the caller reads them."""
    lib = _lib.get_lib()
    arr = make_solve_bands(descs, bufs)
    flags = _lib.PRIOR_BROADCAST if pinv_f.is_broadcast else 0
    with torch.cuda.device(ws.device):
        _lib.check(lib.kafka_solve_bands(ws.n_pix, ws.ld, ws.n_params, arr, len(descs), _lib.ptr(x_prev.data),
                                         _lib.ptr(x_f.data), _lib.ptr(pinv_f.data), _lib.ptr(x_out.data),
                                         _lib.ptr(a_out.data), _lib.ptr(ws.innov),
                                         _lib.ptr(ws.fwd) if want_fwd else None, _lib.ptr(ws.norm2),
                                         _lib.ptr(ws.n_bad), _lib.ptr(ws.scratch), flags, _lib.stream_ptr()),
                   "kafka_solve_bands")
