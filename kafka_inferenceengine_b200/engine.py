"""Device-side plumbing shared by ``LinearKalman`` and the solver functions:
observation upload + compaction gather, band descriptors and the call into the
fused build+solve kernel.  No arithmetic happens here."""
import os

import numpy as np
import scipy.sparse as sp
import torch

from . import _lib
from .device import padded, rows_zero_padded


def pinned_raster(shape, dtype=np.float64):
    """Numpy array in page-locked host memory.  A reader that fills such arrays
    (e.g. GDAL's ``ReadAsArray(buf_obj=...)``) and hands them out through
    ``get_band_data`` lets the device read the rasters straight from them: the
    staging copy through this module's own pinned buffers -- one extra read and
    write of every byte by the host cores -- is skipped.  The array must stay
    untouched until the date it belongs to has been assimilated."""
    dtype = np.dtype(dtype)
    tdt = torch.bool if dtype == np.bool_ else getattr(torch, dtype.name)
    return torch.empty(tuple(np.atleast_1d(shape)), dtype=tdt).pin_memory().numpy()


def _pinned_view(a, dtype):
    """Flat torch view of a numpy raster when the device can read it where it lies
    (C-contiguous, of ``dtype``, page-locked), else None."""
    if not isinstance(a, np.ndarray) or a.dtype != dtype or not a.flags.c_contiguous or a.size == 0:
        return None
    if not a.flags.writeable:
        return None
    t = torch.from_numpy(a.reshape(-1))
    return t if t.is_pinned() else None


_F64, _BOOL = np.dtype(np.float64), np.dtype(np.bool_)


class BandBuffers(object):
    """Compacted per-band observation rows on the device."""

    def __init__(self, n_bands, pix, device):
        self.n_bands = n_bands
        self.pix = pix
        # every gather writes all pixels of its band row; only the padding has to start at zero
        self.y = rows_zero_padded(n_bands, pix.n_pix, device)
        self.w = rows_zero_padded(n_bands, pix.n_pix, device)
        self.m = rows_zero_padded(n_bands, pix.n_pix, device, torch.uint8)


class ScaledReflectance(object):
    """Marker passed in the ``uncertainty`` field of a band whose ``observations``
    are the RAW scaled-integer surface reflectance (int16 / uint16 / float32
    raster as read from the granule).  The arithmetic of
    ``Sentinel2Observations.get_band_data`` (Sentinel2_Observations.py:166-179:
    ``mask = raw > 0``, ``rho = raw / 10000``, ``sigma = 0.05 rho``, ``R^-1 =
    1/sigma^2``) then runs on the device, fused with the compaction gather, and
    only 2-4 bytes per pixel cross PCIe instead of 17.  ``mask`` of such a band
    may be None or an extra boolean raster that is AND-ed in."""

    RAW_TYPES = {np.dtype(np.int16): _lib.RAW_I16, np.dtype(np.uint16): _lib.RAW_U16,
                 np.dtype(np.float32): _lib.RAW_F32}

    def __init__(self, divisor=10000.0, rel_sigma=0.05):
        self.divisor = float(divisor)
        self.rel_sigma = float(rel_sigma)


class BHRKernelWeights(object):
    """Marker passed in the ``uncertainty`` field of a band whose ``observations``
    are the three BRDF kernel weights ``(3, ny, nx)`` (float32 or float64) of a
    MODIS broadband: the arithmetic of ``BHRObservations.get_band_data``
    (input_output/observations.py:288-310: bi-hemispherical reflectance, sigma =
    max(2.5e-3, 5 % or 7 % of it by quality level, 1/sigma^2) then runs on the
    device fused with the compaction gather.  ``qa_level`` is the MCD43A2 quality
    raster (integers); the band's ``mask`` is the reader's boolean mask."""

    KERNEL_TYPES = {np.dtype(np.float32): _lib.RAW_F32, np.dtype(np.float64): _lib.RAW_F64}

    def __init__(self, qa_level):
        self.qa_level = np.asarray(qa_level)


class BackscatterSigma0(object):
    """Marker passed in the ``uncertainty`` field of a band whose ``observations``
    are the Sentinel-1 sigma0 raster of one polarisation as read from the file
    (float32 or float64): the arithmetic of ``S1Observations.get_band_data``
    (Sentinel1_Observations.py:107-130,147-149,176-187: ``mask = sigma0 != -999``,
    ``sigma = 0.05 sigma0``, ``R^-1 = 1/sigma^2``, each step rounded in the raster's
    own type) then runs on the device, fused with the compaction gather; 4-8 bytes
    per pixel cross PCIe instead of 17.  ``mask`` of such a band is not read."""

    RAW_TYPES = {np.dtype(np.float32): _lib.RAW_F32, np.dtype(np.float64): _lib.RAW_F64}

    def __init__(self, missing_value=-999.0, rel_sigma=0.05):
        self.missing_value = float(missing_value)
        self.rel_sigma = float(rel_sigma)


class _Slot(object):
    """One set of pinned staging buffers + device rasters."""

    def __init__(self, n_bands, n, device):
        self.cap, self.n, self.device = n_bands, n, device
        self.d_obs = torch.empty((n_bands, n), dtype=torch.float64, device=device)
        self.d_inv = torch.empty((n_bands, n), dtype=torch.float64, device=device)
        self.d_msk = torch.empty((n_bands, n), dtype=torch.uint8, device=device)
        # page-locked staging buffers: allocated when the first raster arrives that has to be staged
        # (rasters handed over in page-locked memory never need them)
        self.h_obs = self.h_inv = self.h_msk = None
        self.np_obs = self.np_inv = self.np_msk = None
        self.h_raw = self.d_raw = None        # (n_bands, n) 4-byte cells for raw rasters, allocated on first use
        self.np_raw = None
        self.copied = [torch.cuda.Event() for _ in range(n_bands)]   # H2D of band b done
        self.gathered = None                                         # main stream finished reading d_*
        self.up = None                        # (n_bands, source / destination / size arrays) of kafka_upload_rasters

    def ensure_host(self, floats=True):
        if floats and self.h_obs is None:
            self.h_obs = torch.empty((self.cap, self.n), dtype=torch.float64).pin_memory()
            self.h_inv = torch.empty((self.cap, self.n), dtype=torch.float64).pin_memory()
            # numpy views made once (a fresh .numpy() per band and date is measurable on small rasters)
            self.np_obs, self.np_inv = self.h_obs.numpy(), self.h_inv.numpy()
        if self.h_msk is None:
            self.h_msk = torch.empty((self.cap, self.n), dtype=torch.uint8).pin_memory()
            self.np_msk = self.h_msk.numpy().view(np.bool_)

    def ensure_raw(self, host):
        if self.d_raw is None:
            self.d_raw = torch.empty((self.cap, self.n), dtype=torch.int32, device=self.device)
        if host and self.h_raw is None:
            self.h_raw = torch.empty((self.cap, self.n), dtype=torch.int32).pin_memory()
            self.np_raw = self.h_raw.numpy()


class StagedDate(object):
    """Hand-over between ``stage`` and ``finish``."""
    __slots__ = ("slot", "n_bands", "device_data", "raw", "one_event", "sources", "keep")

    def __init__(self, slot, n_bands, device_data=None, raw=None, one_event=False, sources=None, keep=None):
        self.keep = keep                     # host views the queued copies read from: alive until finish()
        # sources: band -> ("bhr", kernel_type, kernels, mask, qa) or ("s1", raw_type, sigma0, marker), device tensors
        self.slot, self.n_bands, self.device_data = slot, n_bands, device_data
        self.raw = raw or {}                 # band -> (raw_type, ScaledReflectance, has_mask)
        self.sources = sources or {}         # bands ingested from source rasters uploaded as they are
        self.one_event = one_event           # all bands went up in one copy: only copied[n_bands-1] was recorded


class ObservationStager(object):
    """Host -> device path of one date's rasters (observations, inverse variance,
    mask), followed by the compaction gather (solvers.py:90-93).

    The collaborator hands over ordinary (pageable) numpy rasters.  ``stage``
    copies them band by band into pinned staging buffers with a small thread
    pool (numpy releases the GIL for these copies) and queues each band's DMA on
    a copy stream as soon as its staging is done; ``finish`` makes the compute
    stream wait for each band's event and runs the gather kernels.  Two buffer
    sets alternate, so the NEXT date can be staged and copied while the kernels
    of the current date are still running (``LinearKalman`` does that from its
    iteration loop).  Buffers are allocated once and reused across dates."""

    # staging threads: up to 8, but never more than this rank's share of the host cores
    # (KAFKA_STAGE_THREADS overrides)
    N_THREADS = int(os.environ.get("KAFKA_STAGE_THREADS", 0)) or \
        max(2, min(8, (os.cpu_count() or 8) // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))))
    N_SLOTS = 2

    def __init__(self, pix):
        self.pix = pix
        self.device = pix.device
        self.h2d_bytes = 0
        self._slots = [None] * self.N_SLOTS
        self._next = 0
        self._pool = None
        self._copy_stream = None
        self._pinned_cache = {}              # address -> (byte size, page-locked?)
        self._band_cache = {}                # ids of a band's three arrays -> (their addresses, the arrays)

    def _slot(self, n_bands):
        k = self._next
        self._next = (k + 1) % self.N_SLOTS
        if self._slots[k] is None or self._slots[k].cap < n_bands:
            self._slots[k] = _Slot(n_bands, self.pix.n_raster, self.device)
        if self._pool is None:
            import concurrent.futures
            self._pool = concurrent.futures.ThreadPoolExecutor(max_workers=self.N_THREADS)
            self._copy_stream = torch.cuda.Stream(device=self.device)
        return self._slots[k]

    @staticmethod
    def _inv_var(unc, n):
        """Main diagonal of the reference's (ny*nx)^2 sparse inverse-variance
        matrix (Sentinel2_Observations.py:176-179), or a plain raster of it."""
        if sp.issparse(unc):
            if unc.format == "dia" and len(unc.offsets) == 1 and unc.offsets[0] == 0 and unc.shape[0] == unc.shape[1]:
                return unc.data[0]                # the stored diagonal itself, no copy
            return unc.diagonal()
        a = np.asarray(unc)
        if a.ndim == 2 and a.shape == (n, n):
            return np.diagonal(a)
        return a.reshape(-1)

    @staticmethod
    def _stage_part(slot, b, d, lo, hi):
        np.copyto(slot.np_obs[b, lo:hi], np.asarray(d.observations).reshape(-1)[lo:hi], casting="unsafe")
        np.copyto(slot.np_inv[b, lo:hi], d._kafka_inv_var[lo:hi], casting="unsafe")
        m = np.asarray(d.mask).reshape(-1)[lo:hi]
        if m.dtype == np.bool_:
            np.copyto(slot.np_msk[b, lo:hi], m)
        else:
            np.not_equal(m, 0, out=slot.np_msk[b, lo:hi])

    @staticmethod
    def _stage_raw(slot, b, d, info, lo, hi):
        dt_ = info[3]
        dst = slot.np_raw[b].view(np.uint8)[:slot.np_raw.shape[1] * dt_.itemsize].view(dt_)
        np.copyto(dst[lo:hi], np.asarray(d.observations).reshape(-1)[lo:hi])
        if info[2]:
            m = np.asarray(d.mask).reshape(-1)[lo:hi]
            if m.dtype == np.bool_:
                np.copyto(slot.np_msk[b, lo:hi], m)
            else:
                np.not_equal(m, 0, out=slot.np_msk[b, lo:hi])

    def stage(self, current_data):
        """Start moving one date's rasters to the device; returns a token for ``finish``."""
        nb = len(current_data)
        if all(isinstance(d.observations, torch.Tensor) and d.observations.is_cuda for d in current_data):
            return StagedDate(None, nb, list(current_data))       # already resident in HBM
        slot = self._slot(nb)
        n = self.pix.n_raster
        slot.copied[nb - 1].synchronize()            # host: the DMA that last read these pinned buffers is done
        if slot.gathered is not None:
            self._copy_stream.wait_event(slot.gathered)           # device rasters no longer being gathered
        parts = max(1, min(self.N_THREADS, n // (1 << 20)))       # split big rasters across the pool
        raw = {}
        sources = {}
        for b, d in enumerate(current_data):
            if isinstance(d.uncertainty, BHRKernelWeights):
                sources[b] = self._upload_bhr(d, n)
                continue
            if isinstance(d.uncertainty, BackscatterSigma0):
                sources[b] = self._upload_backscatter(d, n)
                continue
            if isinstance(d.uncertainty, ScaledReflectance):
                dt_ = np.asarray(d.observations).dtype
                if dt_ not in ScaledReflectance.RAW_TYPES:
                    raise TypeError("raw reflectance rasters must be int16, uint16 or float32, not %s" % dt_)
                raw[b] = (ScaledReflectance.RAW_TYPES[dt_], d.uncertainty, d.mask is not None, dt_)
        holders = [None if (b in raw or b in sources) else _Staged(d.observations, d.mask, self._inv_var(d.uncertainty, n))
                   for b, d in enumerate(current_data)]
        futures = []
        small = n < (1 << 18)                          # tiny rasters: the pool's hand-over costs more than the copy
        # rasters that already lie in page-locked memory (pinned_raster) are read by the DMA engine
        # where they are: no staging copy, the host cores do not touch the data at all
        if small:
            # tiny rasters: all in place (addresses, see below) or all through one staging copy
            direct = self._direct_addresses(holders, n, current_data) if not raw and not sources else {}
        else:
            direct = self._direct_sources(current_data, holders, raw, sources, n)
        staged_bands = [b for b in range(nb) if b not in sources and b not in direct]
        if raw:
            slot.ensure_raw(host=any(b in raw for b in staged_bands))
        if staged_bands:
            slot.ensure_host(floats=any(b not in raw for b in staged_bands))
        for b, d in enumerate(current_data):
            if small or b in sources or b in direct:
                futures.append([])
            elif b in raw:
                futures.append([self._pool.submit(self._stage_raw, slot, b, d, raw[b], n * k // parts,
                                                  n * (k + 1) // parts) for k in range(parts)])
            else:
                futures.append([self._pool.submit(self._stage_part, slot, b, holders[b], n * k // parts,
                                                  n * (k + 1) // parts) for k in range(parts)])
        if small and not raw and not sources and direct:
            # tiny page-locked rasters: three copies per band straight from the caller's arrays (the staging
            # memcpy of the whole date would cost more than the copies), queued by ONE library call
            # (kafka_upload_rasters), one event, one gather
            import ctypes as C
            if slot.up is None or slot.up[0] != nb:
                dst = (C.c_void_p * (3 * nb))(*([slot.d_obs[b].data_ptr() for b in range(nb)]
                                                + [slot.d_inv[b].data_ptr() for b in range(nb)]
                                                + [slot.d_msk[b].data_ptr() for b in range(nb)]))
                sizes = (C.c_int64 * (3 * nb))(*([8 * n] * (2 * nb) + [n] * nb))
                slot.up = (nb, (C.c_void_p * (3 * nb))(), dst, sizes)
            src = slot.up[1]
            for b in range(nb):
                src[b], src[nb + b], src[2 * nb + b] = direct[b]
            with torch.cuda.device(self.device):
                _lib.check(_lib.get_lib().kafka_upload_rasters(3 * nb, src, slot.up[2], slot.up[3],
                                                               C.c_void_p(self._copy_stream.cuda_stream)),
                           "kafka_upload_rasters")
            slot.copied[nb - 1].record(self._copy_stream)
            self.h2d_bytes += nb * n * 17
            return StagedDate(slot, nb, one_event=True, keep=holders)
        if small and not raw and not sources:
            # tiny rasters, the reference's float rasters: launch overhead is all there is, so
            # stage every band inline and queue ONE copy per array instead of three per band
            for b in range(nb):
                self._stage_part(slot, b, holders[b], 0, n)
            with torch.cuda.stream(self._copy_stream):
                slot.d_obs[:nb].copy_(slot.h_obs[:nb], non_blocking=True)
                slot.d_inv[:nb].copy_(slot.h_inv[:nb], non_blocking=True)
                slot.d_msk[:nb].copy_(slot.h_msk[:nb], non_blocking=True)
                slot.copied[nb - 1].record(self._copy_stream)
            self.h2d_bytes += nb * n * 17
            return StagedDate(slot, nb, raw=raw, one_event=True)
        for b in range(nb):
            if b in sources:
                slot.copied[b].record(self._copy_stream)           # uploaded above, on the copy stream
                continue
            if small and b not in direct:
                if b in raw:
                    self._stage_raw(slot, b, current_data[b], raw[b], 0, n)
                else:
                    self._stage_part(slot, b, holders[b], 0, n)
            for f in futures[b]:
                f.result()
            with torch.cuda.stream(self._copy_stream):
                if b in direct:
                    t_obs, t_inv, t_msk = direct[b]
                    if b in raw:
                        nbytes = n * raw[b][3].itemsize
                        slot.d_raw[b].view(torch.uint8)[:nbytes].copy_(t_obs.view(torch.uint8), non_blocking=True)
                        self.h2d_bytes += nbytes
                    else:
                        slot.d_obs[b].copy_(t_obs, non_blocking=True)
                        slot.d_inv[b].copy_(t_inv, non_blocking=True)
                        self.h2d_bytes += n * 16
                    if t_msk is not None:
                        slot.d_msk[b].copy_(t_msk.view(torch.uint8), non_blocking=True)
                        self.h2d_bytes += n
                elif b in raw:
                    item = raw[b][3].itemsize
                    nbytes = n * item
                    src = slot.h_raw[b].view(torch.uint8)[:nbytes]
                    slot.d_raw[b].view(torch.uint8)[:nbytes].copy_(src, non_blocking=True)
                    self.h2d_bytes += nbytes
                    if raw[b][2]:
                        slot.d_msk[b].copy_(slot.h_msk[b], non_blocking=True)
                        self.h2d_bytes += n
                else:
                    slot.d_obs[b].copy_(slot.h_obs[b], non_blocking=True)
                    slot.d_inv[b].copy_(slot.h_inv[b], non_blocking=True)
                    slot.d_msk[b].copy_(slot.h_msk[b], non_blocking=True)
                    self.h2d_bytes += n * 17
                slot.copied[b].record(self._copy_stream)
        return StagedDate(slot, nb, raw=raw, sources=sources, keep=direct)

    def _direct_addresses(self, holders, n, current_data=None):
        """band -> (address of obs, of the inverse variance, of the mask) when ALL bands' rasters are
        page-locked float64 / float64 / bool arrays of n elements, else {}.  Whether an address is page
        locked is remembered (asking the driver costs more than the rest of a small date's staging)."""
        out = {}
        cache = self._pinned_cache
        for b, h in enumerate(holders):
            # the same three array objects as on an earlier date (a source that keeps its buffers): same addresses
            d = current_data[b]
            key = (id(d.observations), id(d.uncertainty), id(d.mask))
            hit = self._band_cache.get(key)
            if hit is not None and hit[1] is d.observations and hit[2] is d.uncertainty and hit[3] is d.mask:
                out[b] = hit[0]
                continue
            trio = []
            for a, dt_ in ((h.observations, _F64), (h._kafka_inv_var, _F64), (h.mask, _BOOL)):
                if not isinstance(a, np.ndarray) or a.dtype != dt_ or a.size != n or not a.flags.c_contiguous:
                    return {}
                addr = a.ctypes.data
                hit = cache.get(addr)
                if hit is None or hit[0] != a.nbytes:
                    if len(cache) > 4096:
                        cache.clear()
                    hit = (a.nbytes, _pinned_view(a, dt_) is not None)
                    cache[addr] = hit
                if not hit[1]:
                    return {}
                trio.append(addr)
            out[b] = tuple(trio)
            if len(self._band_cache) > 256:
                self._band_cache.clear()
            self._band_cache[key] = (out[b], d.observations, d.uncertainty, d.mask)
        return out

    @staticmethod
    def _direct_sources(current_data, holders, raw, sources, n):
        """band -> (obs, inverse variance or None, mask or None) as flat torch views of the
        caller's own page-locked arrays, for the bands whose rasters can all be read in place."""
        direct = {}
        for b, d in enumerate(current_data):
            if b in sources:
                continue
            if b in raw:
                t_obs = _pinned_view(np.asarray(d.observations), raw[b][3])
                t_msk = None
                if raw[b][2]:
                    t_msk = _pinned_view(np.asarray(d.mask), np.dtype(np.bool_))
                    if t_msk is None:
                        continue
                if t_obs is not None and t_obs.numel() == n:
                    direct[b] = (t_obs, None, t_msk)
                continue
            h = holders[b]
            t_obs = _pinned_view(h.observations, np.dtype(np.float64))
            t_inv = _pinned_view(h._kafka_inv_var, np.dtype(np.float64))
            t_msk = _pinned_view(h.mask, np.dtype(np.bool_))
            if t_obs is not None and t_inv is not None and t_msk is not None and \
                    t_obs.numel() == t_inv.numel() == t_msk.numel() == n:
                direct[b] = (t_obs, t_inv, t_msk)
        return direct

    def _upload_bhr(self, d, n):
        """Kernel weights (3, ny, nx), quality level and mask of one MODIS band to the device
        (a 2400 x 2400 tile is 80 MB per band: plain copies on the copy stream)."""
        k = np.asarray(d.observations)
        if k.dtype not in BHRKernelWeights.KERNEL_TYPES or k.shape[0] != 3 or k.size != 3 * n:
            raise TypeError("BRDF kernel weights must be a float32/float64 array of shape (3, ny, nx)")
        qa = np.asarray(d.uncertainty.qa_level)
        if qa.size != n or np.asarray(d.mask).size != n:
            raise ValueError("quality level and mask must have the raster's shape")
        qa8 = np.where((qa >= 0) & (qa <= 254), qa, 255).astype(np.uint8).reshape(-1)   # only 0 and 1 carry meaning
        m8 = (np.asarray(d.mask).reshape(-1) != 0).astype(np.uint8)
        with torch.cuda.stream(self._copy_stream):
            kt = torch.from_numpy(np.ascontiguousarray(k.reshape(3, n))).to(self.device, non_blocking=True)
            qt = torch.from_numpy(qa8).to(self.device, non_blocking=True)
            mt = torch.from_numpy(m8).to(self.device, non_blocking=True)
        self.h2d_bytes += k.nbytes + 2 * n
        return ("bhr", BHRKernelWeights.KERNEL_TYPES[k.dtype], kt, mt, qt)

    def _upload_backscatter(self, d, n):
        """Sigma0 raster of one Sentinel-1 polarisation to the device, in its own type."""
        r = np.asarray(d.observations)
        if r.dtype not in BackscatterSigma0.RAW_TYPES or r.size != n:
            raise TypeError("a sigma0 raster must be a float32/float64 array of the state mask's shape")
        with torch.cuda.stream(self._copy_stream):
            rt = torch.from_numpy(np.ascontiguousarray(r.reshape(n))).to(self.device, non_blocking=True)
        self.h2d_bytes += r.nbytes
        return ("s1", BackscatterSigma0.RAW_TYPES[r.dtype], rt, d.uncertainty)

    def finish(self, token, out=None):
        """Gather the staged rasters into compacted rows on the compute stream."""
        pix = self.pix
        nb = token.n_bands
        out = out or BandBuffers(nb, pix, self.device)
        if token.slot is None:
            for b, d in enumerate(token.device_data):
                m = d.mask.reshape(-1)
                pix.gather_band(d.observations.reshape(-1), d.uncertainty.reshape(-1),
                                m.view(torch.uint8) if m.dtype == torch.bool else m, out.y[b], out.w[b], out.m[b])
            return out
        slot = token.slot
        main = torch.cuda.current_stream(self.device)
        lib = _lib.get_lib()
        if token.one_event:
            main.wait_event(slot.copied[nb - 1])
            pix.gather_bands(slot.d_obs, slot.d_inv, slot.d_msk, nb, out.y, out.w, out.m)
            slot.gathered = torch.cuda.Event()
            slot.gathered.record(main)
            return out
        for b in range(nb):
            main.wait_event(slot.copied[b])
            if b in token.sources and token.sources[b][0] == "s1":
                _, rtype, rt, marker = token.sources[b]
                with torch.cuda.device(self.device):
                    _lib.check(lib.kafka_gather_backscatter_band(_lib.ptr(rt), rtype, _lib.ptr(pix.idx), pix.n_pix,
                                                                 marker.missing_value, marker.rel_sigma,
                                                                 _lib.ptr(out.y[b]), _lib.ptr(out.w[b]),
                                                                 _lib.ptr(out.m[b]), _lib.stream_ptr()),
                               "kafka_gather_backscatter_band")
                rt.record_stream(main)           # allocated on the copy stream, read on the main stream
                continue
            if b in token.sources:
                _, ktype, kt, mt, qt = token.sources[b]
                with torch.cuda.device(self.device):
                    _lib.check(lib.kafka_gather_bhr_band(_lib.ptr(kt), ktype, kt.stride(0), _lib.ptr(mt), _lib.ptr(qt),
                                                         _lib.ptr(pix.idx), pix.n_pix, _lib.ptr(out.y[b]),
                                                         _lib.ptr(out.w[b]), _lib.ptr(out.m[b]), _lib.stream_ptr()),
                               "kafka_gather_bhr_band")
                for t in (kt, mt, qt):
                    t.record_stream(main)        # allocated on the copy stream, read on the main stream
                continue
            if b in token.raw:
                raw_type, sr, has_mask, _ = token.raw[b]
                with torch.cuda.device(self.device):
                    _lib.check(lib.kafka_gather_scaled_band(_lib.ptr(slot.d_raw[b]), raw_type,
                                                            _lib.ptr(slot.d_msk[b]) if has_mask else None,
                                                            _lib.ptr(pix.idx), pix.n_pix, sr.divisor, sr.rel_sigma,
                                                            _lib.ptr(out.y[b]), _lib.ptr(out.w[b]), _lib.ptr(out.m[b]),
                                                            _lib.stream_ptr()), "kafka_gather_scaled_band")
            else:
                pix.gather_band(slot.d_obs[b], slot.d_inv[b], slot.d_msk[b], out.y[b], out.w[b], out.m[b])
        slot.gathered = torch.cuda.Event()
        slot.gathered.record(main)
        return out

    def upload(self, current_data, out=None):
        return self.finish(self.stage(current_data), out)


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
        # the scalars every solve leaves behind, side by side so that one kafka_read_small brings them
        # to the host: [sum of squares (f64) | pixels left at the forecast (i32) | pixels solved by
        # pivoted elimination (i32)]
        self.scalars = torch.zeros(2, dtype=torch.float64, device=device)
        self.norm2 = self.scalars[0:1]
        self.n_bad = self.scalars[1:2].view(torch.int32)[0:2]
        self.scratch = torch.empty(int(lib.kafka_solve_scratch_doubles(n_pix)), dtype=torch.float64, device=device)

    def read_scalars(self):
        """(sum of squares, pixels left at the forecast, pixels solved by pivoted elimination) of
        the last solve, waiting for it."""
        host = _lib.read_small(self.scalars)
        counts = host[1:2].view(np.int32)
        return float(host[0]), int(counts[0]), int(counts[1])


def make_solve_bands(descs, bufs):
    """ctypes array of kafka_solve_band from python descriptors.

    ``descs[b]`` is ("arrays", h0_row, jac_rows(D, ld), mapper),
    ("uniform", h0(1), jac(D), mapper), ("linear", row, offset) or
    ("wcm", (A, B, C, D, E, cos_theta), veg_slot, sm_slot)."""
    nb = len(descs)
    arr = (_lib.SolveBand * max(nb, 1))()
    for b, d in enumerate(descs):
        sb = arr[b]
        sb.y = bufs.y[b].data_ptr()
        sb.w = bufs.w[b].data_ptr()
        if d[0] == "wcm":
            sb.kind = _lib.BAND_WCM
            for p_, v in enumerate(d[1]):
                sb.row[p_] = float(v)
            sb.mapper[0], sb.mapper[1] = int(d[2]), int(d[3])
        elif d[0] == "linear":
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


def solve_bands(ws, descs, bufs, x_prev, x_f, pinv_f, x_out, a_out, want_fwd=False, round_f32=False):
    """One linearised update (kafka_solve_bands).  Results land in ``x_out``,
    ``a_out``, ``ws.innov``, ``ws.fwd``; ``ws.norm2`` / ``ws.n_bad`` stay on the
    device until the caller reads them.  ``a_out=None`` marks an iteration whose
    A, innovations and forward-modelled values nobody will read: none is stored."""
    lib = _lib.get_lib()
    arr = make_solve_bands(descs, bufs)
    flags = (_lib.PRIOR_BROADCAST if pinv_f.is_broadcast else 0) | (_lib.ROUND_F32 if round_f32 else 0) | \
        _lib.SOLVE_COUNT_LU
    with torch.cuda.device(ws.device):
        _lib.check(lib.kafka_solve_bands(ws.n_pix, ws.ld, ws.n_params, arr, len(descs), _lib.ptr(x_prev.data),
                                         _lib.ptr(x_f.data), _lib.ptr(pinv_f.data), _lib.ptr(x_out.data),
                                         _lib.ptr(a_out.data) if a_out is not None else None,
                                         _lib.ptr(ws.innov) if a_out is not None else None,
                                         _lib.ptr(ws.fwd) if (want_fwd and a_out is not None) else None, _lib.ptr(ws.norm2),
                                         _lib.ptr(ws.n_bad), _lib.ptr(ws.scratch), flags, _lib.stream_ptr()),
                   "kafka_solve_bands")
