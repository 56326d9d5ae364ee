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
    """Host -> device path of one date's rasters (observations, inverse variance,
    mask), followed by the compaction gather (solvers.py:90-93).

    The collaborator hands over ordinary (pageable) numpy rasters.  They are
    copied band by band into pinned staging buffers by a small thread pool
    (numpy releases the GIL for these copies), each band's DMA is queued on a
    copy stream as soon as its staging is done, and the gather kernel of a band
    waits only for that band's event -- so host memcpy, PCIe transfer and the
    gathers overlap.  Buffers are allocated once and reused across dates."""

    N_THREADS = 8

    def __init__(self, pix):
        self.pix = pix
        self.device = pix.device
        self._cap = 0
        self.h2d_bytes = 0
        self._pool = None
        self._copy_stream = None

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
        self._events = [torch.cuda.Event() for _ in range(n_bands)]
        self._cap = n_bands
        if self._pool is None:
            import concurrent.futures
            self._pool = concurrent.futures.ThreadPoolExecutor(max_workers=self.N_THREADS)
            self._copy_stream = torch.cuda.Stream(device=self.device)

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

    def _stage(self, b, d, part, n_parts):
        n = self.pix.n_raster
        lo, hi = n * part // n_parts, n * (part + 1) // n_parts
        obs = np.asarray(d.observations).reshape(-1)
        np.copyto(self.h_obs[b].numpy()[lo:hi], obs[lo:hi], casting="unsafe")
        inv = d._kafka_inv_var
        np.copyto(self.h_inv[b].numpy()[lo:hi], inv[lo:hi], casting="unsafe")
        msk = np.asarray(d.mask).reshape(-1)
        np.not_equal(msk[lo:hi], 0, out=self.h_msk[b].numpy()[lo:hi].view(np.bool_))

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
        self._copy_stream.synchronize()              # DMA of the previous date has left the staging buffers
        n = pix.n_raster
        parts = max(1, min(self.N_THREADS, n // (1 << 20)))         # split big rasters across the pool
        holders = []
        for d in current_data:
            h = _Staged(d.observations, d.mask, self._inv_var(d.uncertainty, n))
            holders.append(h)
        futures = [[self._pool.submit(self._stage, b, holders[b], k, parts) for k in range(parts)] for b in range(nb)]
        main = torch.cuda.current_stream(self.device)
        self._copy_stream.wait_stream(main)          # staging buffers of the previous date are free
        for b in range(nb):
            for f in futures[b]:
                f.result()
            with torch.cuda.stream(self._copy_stream):
                self.d_obs[b].copy_(self.h_obs[b], non_blocking=True)
                self.d_inv[b].copy_(self.h_inv[b], non_blocking=True)
                self.d_msk[b].copy_(self.h_msk[b], non_blocking=True)
                self._events[b].record(self._copy_stream)
            main.wait_event(self._events[b])
            pix.gather_band(self.d_obs[b], self.d_inv[b], self.d_msk[b], out.y[b], out.w[b], out.m[b])
        self.h2d_bytes += nb * n * 17
        return out


class _Staged(object):
    __slots__ = ("observations", "mask", "_kafka_inv_var")

    def __init__(self, observations, mask, inv_var):
        self.observations, self.mask, self._kafka_inv_var = observations, mask, inv_var


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

    ``descs[b]`` is ("arrays", h0_row, jac_rows(D, ld), mapper),
    ("uniform", h0(1), jac(D), mapper) or ("linear", row, offset)."""
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
            sb.kind = _lib.BAND_UNIFORM if d[0] == "uniform" else _lib.BAND_ARRAYS
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
