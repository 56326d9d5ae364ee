"""Output stage: device-side replacement for the arithmetic of
``KafkaOutput.dump_data`` (kafka/input_output/observations.py:354-394).

The reference scatters every parameter's mean ``x[ii::n_params]`` and its
uncertainty ``1/sqrt(diag(P^-1)[ii::n_params])`` into ``(ny, nx)`` float32
rasters on the host and writes GeoTIFFs with GDAL.  Here the scatter, the
``1/sqrt`` and the float32 cast run in one kernel (``kafka_scatter_output``) and
the rasters come back through pinned buffers; what is done with them (GeoTIFF,
NetCDF, nothing) is the ``sink``'s business -- GDAL itself is out of scope.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .device import BlockDiag, DeviceVector, PixelIndex


class RasterOutput(object):
    """``output`` collaborator for ``LinearKalman``: same ``dump_data`` signature
    as the reference's ``KafkaOutput``.  ``sink(timestep, parameter_list, mean,
    unc)`` receives two ``(n_params, ny, nx)`` float32 numpy arrays that are only
    valid during the call (they alias reusable pinned buffers).

    With ``overlap=True`` (default) the scatter kernel and the device->host copy
    run on a side stream into one of ``n_buffers`` pinned buffer pairs and the
    sink is called later, in date order, as soon as a ``dump_data`` call (or
    ``flush()``, which ``LinearKalman.run`` calls at the end) finds the copy
    finished -- so the read-back of one date overlaps the kernels of the next
    ones.  The calling thread only waits when every buffer pair is in flight.
    (With two pairs it waited for the previous date's copy at every call, and
    while it waited nothing new was queued on the GPU: on a host whose
    device->host rate is close to one date's output per date -- eight GPUs
    reading back 1.2 GB each -- that cost 20 ms per date.)"""

    def __init__(self, parameter_list, sink=None, keep_last=True, overlap=True, n_buffers=3):
        self.parameter_list = parameter_list
        self.sink = sink
        self.keep_last = keep_last
        self.overlap = overlap
        self.n_buffers = max(2, int(n_buffers))
        self.last = None
        self.d2h_bytes = 0
        self._pix = None
        self._free = []                  # pinned (mean, unc) pairs ready for reuse
        self._n_alloc = 0
        self._held = None                # the pair self.last aliases: recycled at the next delivery
        self._pending = []               # [(timestep, event, host pair, device tensors kept alive)], oldest first
        self._stream = None

    def _pixel_index(self, state_mask, device):
        if self._pix is None or self._pix.shape != np.shape(state_mask):
            self._pix = PixelIndex(state_mask, device)
        return self._pix

    def _deliver(self, block=True):
        """Hand finished dates to the sink, oldest first; ``block=False`` stops at the first
        date whose copy is still running."""
        while self._pending:
            timestep, event, host, keep = self._pending[0]
            if not block and not event.query():
                return
            event.synchronize()
            self._pending.pop(0)
            mean, unc = host[0].numpy(), host[1].numpy()
            if self.sink is not None:
                self.sink(timestep, self.parameter_list, mean, unc)
            if self._held is not None:
                self._free.append(self._held)
                self._held = None
            if self.keep_last:
                self.last = (timestep, mean, unc)
                self._held = host        # stays valid until the next date is delivered
            else:
                self._free.append(host)

    def flush(self):
        self._deliver()

    def _host_pair(self, shape):
        """A free buffer set of the ring: [host mean, host unc, host both (pinned), device both,
        event, the pixel index the device rasters were last written through]."""
        while True:
            while self._free:
                pair = self._free.pop()
                if pair[0].shape == shape:
                    return pair
                self._n_alloc -= 1       # wrong shape (another state mask): drop it
            if self._n_alloc < self.n_buffers or not self._pending:      # (no date in flight: the pair self.last holds)
                self._n_alloc += 1
                both = torch.empty((2,) + tuple(shape), dtype=torch.float32).pin_memory()
                return [both[0], both[1], both, None, torch.cuda.Event(), None]
            # every pair is in flight: wait for the oldest date
            timestep, event, host, keep = self._pending[0]
            event.synchronize()
            self._deliver(block=False)

    def dump_data(self, timestep, x_analysis, P_analysis, P_analysis_inv, state_mask, n_params):
        if not isinstance(x_analysis, DeviceVector):
            x_analysis = DeviceVector.from_interleaved(x_analysis, n_params)
        device = x_analysis.device
        P_analysis_inv = BlockDiag.coerce(P_analysis_inv, n_params, x_analysis.n_pix, device).materialise()
        pix = getattr(self, "pixels", None) or self._pixel_index(state_mask, device)
        self._deliver(block=False)                       # whatever has arrived in the meantime
        if self._stream is None:
            self._stream = torch.cuda.Stream(device=device)
            self._main_done = torch.cuda.Event()
        shape = (n_params,) + tuple(pix.shape)
        pair = self._host_pair(shape)
        # the device rasters belong to the buffer set: filled with zeros once, afterwards only the
        # state pixels are rewritten (through the same index), so no fill and no allocation per date
        zero_first = 0
        if pair[3] is None or pair[3].device != device:
            pair[3] = torch.empty((2, n_params, pix.n_raster), dtype=torch.float32, device=device)
            pair[5] = None
        if pair[5] is not pix:
            zero_first, pair[5] = 1, pix
        dev = pair[3]
        with torch.cuda.device(device):
            if self.overlap:
                self._main_done.record()                 # current stream: the analysis is complete there
                self._stream.wait_event(self._main_done)
                stream = C.c_void_p(self._stream.cuda_stream)
            else:
                stream = _lib.stream_ptr()
            _lib.check(_lib.get_lib().kafka_output_rasters(
                _lib.ptr(x_analysis.data), _lib.ptr(P_analysis_inv.data), x_analysis.ld, pix.n_pix, n_params,
                _lib.ptr(pix.idx), pix.n_raster, C.c_void_p(dev.data_ptr()), C.c_void_p(dev[1].data_ptr()), zero_first,
                C.c_void_p(pair[2].data_ptr()), stream), "kafka_output_rasters")
            if self.overlap:
                pair[4].record(self._stream)
            else:
                pair[4].record()
        # keep the inputs alive until the copy has been consumed
        self._pending.append((timestep, pair[4], pair, (x_analysis, P_analysis_inv)))
        self.d2h_bytes += 2 * n_params * pix.n_raster * 4
        if not self.overlap:
            self._deliver()
