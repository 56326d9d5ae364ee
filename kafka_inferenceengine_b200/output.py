"""Output stage: device-side replacement for the arithmetic of
``KafkaOutput.dump_data`` (kafka/input_output/observations.py:354-394).

The reference scatters every parameter's mean ``x[ii::n_params]`` and its
uncertainty ``1/sqrt(diag(P^-1)[ii::n_params])`` into ``(ny, nx)`` float32
rasters on the host and writes GeoTIFFs with GDAL.  Here the scatter, the
``1/sqrt`` and the float32 cast run in one kernel (``kafka_scatter_output``) and
the rasters come back through pinned buffers; what is done with them (GeoTIFF,
NetCDF, nothing) is the ``sink``'s business -- GDAL itself is out of scope.
"""
import numpy as np
import torch

from .device import BlockDiag, DeviceVector, PixelIndex


class RasterOutput(object):
    """``output`` collaborator for ``LinearKalman``: same ``dump_data`` signature
    as the reference's ``KafkaOutput``.  ``sink(timestep, parameter_list, mean,
    unc)`` receives two ``(n_params, ny, nx)`` float32 numpy arrays that are only
    valid during the call (they alias reusable pinned buffers).

    With ``overlap=True`` (default) the scatter kernel and the device->host copy
    run on a side stream and the sink is called one ``dump_data`` later (or at
    ``flush()``, which ``LinearKalman.run`` calls at the end), so the read-back
    of one date overlaps the kernels of the next."""

    def __init__(self, parameter_list, sink=None, keep_last=True, overlap=True):
        self.parameter_list = parameter_list
        self.sink = sink
        self.keep_last = keep_last
        self.overlap = overlap
        self.last = None
        self.d2h_bytes = 0
        self._pix = None
        self._host = [None, None]
        self._k = 0
        self._pending = None
        self._stream = None

    def _pixel_index(self, state_mask, device):
        if self._pix is None or self._pix.shape != np.shape(state_mask):
            self._pix = PixelIndex(state_mask, device)
        return self._pix

    def _deliver(self):
        if self._pending is None:
            return
        timestep, event, host, keep = self._pending
        self._pending = None
        event.synchronize()
        mean, unc = host[0].numpy(), host[1].numpy()
        if self.sink is not None:
            self.sink(timestep, self.parameter_list, mean, unc)
        if self.keep_last:
            self.last = (timestep, mean, unc)

    def flush(self):
        self._deliver()

    def dump_data(self, timestep, x_analysis, P_analysis, P_analysis_inv, state_mask, n_params):
        if not isinstance(x_analysis, DeviceVector):
            x_analysis = DeviceVector.from_interleaved(x_analysis, n_params)
        device = x_analysis.device
        P_analysis_inv = BlockDiag.coerce(P_analysis_inv, n_params, x_analysis.n_pix, device).materialise()
        pix = getattr(self, "pixels", None) or self._pixel_index(state_mask, device)
        self._deliver()                                  # the previous date's rasters are due now
        main = torch.cuda.current_stream(device)
        if self._stream is None:
            self._stream = torch.cuda.Stream(device=device)
        side = self._stream if self.overlap else main
        side.wait_stream(main)
        with torch.cuda.stream(side):
            mean_r, unc_r = pix.scatter_output(x_analysis, P_analysis_inv, n_params)
            k = self._k
            self._k ^= 1
            if self._host[k] is None or self._host[k][0].shape != mean_r.shape:
                self._host[k] = (torch.empty(mean_r.shape, dtype=torch.float32).pin_memory(),
                                 torch.empty(mean_r.shape, dtype=torch.float32).pin_memory())
            self._host[k][0].copy_(mean_r, non_blocking=True)
            self._host[k][1].copy_(unc_r, non_blocking=True)
            event = torch.cuda.Event()
            event.record(side)
        # keep the device tensors alive until the copy has been consumed
        self._pending = (timestep, event, self._host[k], (mean_r, unc_r, x_analysis, P_analysis_inv))
        self.d2h_bytes += 2 * mean_r.numel() * 4
        if not self.overlap:
            self._deliver()
