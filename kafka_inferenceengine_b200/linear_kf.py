"""Drop-in for ``kafka.linear_kf.LinearKalman`` (kafka/linear_kf.py:55-452) with
the per-pixel assimilation step on a B200.

Same constructor, ``run``, ``advance``, ``assimilate*``, ``do_all_bands``,
``solver*`` and ``set_trajectory_*`` as the reference; the state lives on the
GPU between calls as ``DeviceVector`` / ``BlockDiag`` objects, which behave like
the reference's numpy vector and scipy block-diagonal matrix towards the
caller-owned collaborators (``observations``, ``output``, ``prior``).

Two keyword-only additions: ``device`` (CUDA device) and ``comm`` (a
``torch.distributed`` process group): with ``comm`` every rank owns a shard of
the pixels and the convergence norm (linear_kf.py:293) is all-reduced so that
all ranks take the same stop decision as one process holding the whole image.
"""
import logging
from collections import namedtuple

import numpy as np
import torch

from . import _lib
from . import engine
from .device import BlockDiag, DeviceBands, DeviceVector, PixelIndex, default_device, rows_zero_padded
from .inference import kf_tools
from .inference.kf_tools import (DiagonalMatrix, IdentityModel, propagate_and_blend_prior,
                                 propagate_information_filter_LAI)
from .inference.utils import (ObservationOperator, emulator_table, iterate_time_grid, linearize_bands)

LOG = logging.getLogger(__name__ + ".linear_kf")

Metadata = namedtuple('Metadata', 'mask uncertainty')
Previous_State = namedtuple("Previous_State", "timestamp x_vect cov_m icov_mv")


class LinearKalman(object):
    """The raster Kalman / information filter of the reference, B200 inside."""

    def __init__(self, observations, output, state_mask,
                 create_observation_operator, parameters_list,
                 state_propagation=propagate_information_filter_LAI,
                 band_mapper=None,
                 linear=True, diagnostics=True, prior=None, *, device=None, comm=None,
                 round_operator_f32=True, dedup_uniform_state=True, prefetch_observations=True,
                 hessian_correction=False, small_image_path=None):
        _lib.get_lib()                                   # fail now if the CUDA library is missing
        if not isinstance(create_observation_operator, ObservationOperator):
            raise NotImplementedError(
                "create_observation_operator must be one of this package's operator descriptors "
                "(create_prosail_observation_operator, create_nonlinear_observation_operator, "
                "create_linear_observation_operator): arbitrary Python callables cannot run per "
                "pixel on the GPU and there is no CPU fallback")
        self.parameters_list = parameters_list
        self.n_params = len(self.parameters_list)
        if not 1 <= self.n_params <= _lib.MAX_PARAMS:
            raise ValueError("n_params=%d outside 1..%d" % (self.n_params, _lib.MAX_PARAMS))
        self.observations = observations
        self.output = output
        self.diagnostics = diagnostics
        self.state_mask = state_mask
        self.device = device or default_device()
        # state-mask compaction (linear_kf.py:79-80), done once on the device
        self.pixels = PixelIndex(state_mask, self.device)
        self.n_state_elems = self.pixels.n_pix
        self._state_propagator = state_propagation
        self._advance = propagate_and_blend_prior
        self.prior = prior
        self.band_mapper = band_mapper
        self._create_observation_operator = create_observation_operator
        self.comm = comm
        self.round_operator_f32 = bool(round_operator_f32)
        self.dedup_uniform_state = bool(dedup_uniform_state)
        # second-order correction of the posterior precision after convergence; the reference
        # has it switched off on the run() path (linear_kf.py:313-319) and on in assimilate_band
        self.hessian_correction = bool(hessian_correction)
        # stage the NEXT observation date (observations.get_band_data + host->device copy)
        # while the kernels of the current date run; calls get_band_data one date early
        self.prefetch_observations = bool(prefetch_observations)
        # one C-ABI call per date for small images (see _do_all_bands_small)
        self.small_image_path = self.SMALL_IMAGE_PATH_DEFAULT if small_image_path is None else bool(small_image_path)
        self._small = {}
        self._plan_cache = {}                # emulator objects of a date -> band plan
        self._schedule = []                  # observation dates still to come in run()
        self._staged = {}                    # date -> (token, current_data)
        self.want_forward_modelled = False
        self._stager = engine.ObservationStager(self.pixels)
        self._ws = {}
        self.trajectory_model = None
        self.trajectory_uncertainty = None
        self.last_iterations = []            # solver calls per assimilated date (diagnostic)
        self.last_norms = []
        self.profile = None                  # dict of CUDA-event pairs when bench.py asks for kernel times
        LOG.info("Starting KaFKA run!!!")

    # ------------------------------------------------------------------
    def advance(self, x_analysis, P_analysis, P_analysis_inverse,
                trajectory_model, trajectory_uncertainty):
        """linear_kf.py:99-108."""
        LOG.info("Calling state propagator...")
        x_forecast, P_forecast, P_forecast_inverse = \
            self._advance(x_analysis, P_analysis, P_analysis_inverse,
                          trajectory_model, trajectory_uncertainty,
                          prior=self.prior, date=self.current_timestep,
                          state_propagator=self._state_propagator)
        return x_forecast, P_forecast, P_forecast_inverse

    def _set_plot_view(self, diag_string, timestep, obs):
        pass

    def _plotter_iteration_start(self, plot_obj, x, obs, mask):
        pass

    def _plotter_iteration_end(self, plot_obj, x, P, innovation, mask):
        pass

    def set_trajectory_model(self):
        """linear_kf.py:123-129: identity."""
        n = self.n_state_elems
        self.trajectory_model = IdentityModel(self.n_params * n)

    def set_trajectory_uncertainty(self, Q):
        """linear_kf.py:131-146: ``Q`` is the main diagonal (length n_params*n
        or anything that broadcasts to it)."""
        n = self.n_state_elems * self.n_params
        q = np.broadcast_to(np.asarray(Q, dtype=np.float64), (n,)) if np.ndim(Q) <= 1 else np.asarray(Q).ravel()
        self.trajectory_uncertainty = DiagonalMatrix(np.ascontiguousarray(q), self.n_params, self.device)

    def _get_observations_timestep(self, timestep, band=None):
        """linear_kf.py:148-169."""
        data = self.observations.get_band_data(timestep, band)
        return (data.observations, data.uncertainty, data.mask, data.metadata, data.emulator)

    # ------------------------------------------------------------------
    def run(self, time_grid, x_forecast, P_forecast, P_forecast_inverse,
            diag_str="diagnostics",
            band=None, approx_diagonal=True, refine_diag=True,
            iter_obs_op=False, is_robust=False, dates=None):
        """linear_kf.py:171-212: the time loop.  State stays on the device from
        one time step to the next."""
        x_forecast = self._state(x_forecast)
        P_forecast_inverse = self._precision(P_forecast_inverse)
        plan = list(iterate_time_grid(time_grid, self.observations.dates))
        self._schedule = [d for _, locs, _ in plan for d in locs]
        self._staged = {}
        for timestep, locate_times, is_first in plan:
            self.current_timestep = timestep
            if not is_first:
                LOG.info("Advancing state, %s" % timestep.strftime("%Y-%m-%d"))
                x_forecast, P_forecast, P_forecast_inverse = self.advance(
                    x_analysis, P_analysis, P_analysis_inverse,
                    self.trajectory_model, self.trajectory_uncertainty)
            if len(locate_times) == 0:
                x_analysis = x_forecast
                P_analysis = P_forecast
                P_analysis_inverse = P_forecast_inverse
                LOG.info("No observations in this time")
            else:
                x_analysis, P_analysis, P_analysis_inverse = self.assimilate_multiple_bands(
                    locate_times, x_forecast, P_forecast, P_forecast_inverse,
                    approx_diagonal=approx_diagonal, refine_diag=refine_diag,
                    iter_obs_op=iter_obs_op, is_robust=is_robust, diag_str=diag_str)
            LOG.info("Dumping results to disk")
            self.output.dump_data(timestep, x_analysis, P_analysis,
                                  P_analysis_inverse, self.state_mask, self.n_params)
        if hasattr(self.output, "flush"):
            self.output.flush()

    def assimilate_multiple_bands(self, locate_times, x_forecast, P_forecast,
                                  P_forecast_inverse,
                                  approx_diagonal=True, refine_diag=False,
                                  iter_obs_op=False, is_robust=False, diag_str="diag"):
        """linear_kf.py:214-242: sequential update over the observation dates of
        one grid interval, all bands of a date at once."""
        for step in locate_times:
            LOG.info("Assimilating %s..." % step.strftime("%Y-%m-%d"))
            if self._schedule and self._schedule[0] == step:
                self._schedule.pop(0)
            staged = self._staged.pop(step, None)
            if staged is not None:
                self._token, current_data = staged
            else:
                current_data = self._read_bands(step)
            x_analysis, P_analysis, P_analysis_inverse, innovations = \
                self.do_all_bands(step, current_data, x_forecast, P_forecast, P_forecast_inverse)
            # the reference copies (x*1., linear_kf.py:232-240); device results are
            # never written again, so the next forecast may simply alias them
            x_forecast = x_analysis
            P_forecast = P_analysis
            P_forecast_inverse = P_analysis_inverse
        return x_analysis, P_analysis, P_analysis_inverse

    # ------------------------------------------------------------------
    def _read_bands(self, step):
        return [self.observations.get_band_data(step, band)
                for band in range(self.observations.bands_per_observation[step])]

    def _prefetch_next(self):
        """Read and start uploading the next observation date of run() while the
        GPU is busy with the current one."""
        if not (self.prefetch_observations and self._schedule):
            return
        nxt = self._schedule[0]
        if nxt in self._staged:
            return
        data = self._read_bands(nxt)
        self._staged[nxt] = (self._stager.stage(data), data)

    def _state(self, x):
        x = DeviceVector.from_interleaved(x, self.n_params, self.device)
        if x.n_pix != self.n_state_elems:
            raise ValueError("state vector holds %d pixels, the state mask %d" % (x.n_pix, self.n_state_elems))
        return x

    def _precision(self, A):
        if A is None:
            raise ValueError("the information filter needs P_forecast_inverse")
        A = BlockDiag.coerce(A, self.n_params, self.n_state_elems, self.device)
        if A.n_pix != self.n_state_elems or A.n_params != self.n_params:
            raise ValueError("inverse covariance does not match the state (%d px x %d params)"
                             % (self.n_state_elems, self.n_params))
        return A

    def _workspace(self, n_bands):
        key = (n_bands, self.want_forward_modelled)
        ws = self._ws.get(key)
        if ws is None:
            ws = engine.SolveWorkspace(self.n_state_elems, self.n_params, n_bands, self.device,
                                       want_fwd=self.want_forward_modelled)
            self._ws[key] = ws
        return ws

    def _band_plan(self, current_data):
        """Group the bands of one date by operator family.  Returns
        (gp_groups, descs_template): gp_groups = [(band indices, [(table, mapper)])]
        sharing an emulator input size, descs_template[b] = ("linear", row, offset)
        or None for emulator bands (filled per iteration)."""
        key = tuple(id(d.emulator) for d in current_data)
        hit = self._plan_cache.get(key)
        if hit is not None and all(e is d.emulator for e, d in zip(hit[2], current_data)):
            return hit[0], list(hit[1])          # same emulator objects as before: nothing to rebuild or re-check
        op = self._create_observation_operator
        descs = [None] * len(current_data)
        groups = {}
        for b, data in enumerate(current_data):
            if op.kind == "wcm":
                if self.n_params < 2:
                    raise ValueError("the SAR operator needs a state of (vegetation descriptor, soil moisture)")
                slots = (0, 1) if self.band_mapper is None else tuple(int(v) for v in self.band_mapper[b][:2])
                descs[b] = ("wcm", op.band_parameters(b, data.metadata), slots[0], slots[1])
            elif op.kind == "linear":
                em = data.emulator
                row = np.asarray(em.row, dtype=np.float64)
                if row.size != self.n_params:
                    raise ValueError("linear operator row has %d entries for %d parameters" % (row.size, self.n_params))
                descs[b] = ("linear", row, float(em.offset))
            else:
                tab = emulator_table(data.emulator, self.device)
                if op.state_mappers is None and self.band_mapper is not None:
                    mp = np.asarray(self.band_mapper[b], dtype=np.int32)
                else:
                    mp = op.mapper(self.n_params, b, tab.n_dim)
                if len(mp) != tab.n_dim or mp.min() < 0 or mp.max() >= self.n_params:
                    raise ValueError("band %d: state mapper %s does not fit emulator/state" % (b, mp))
                groups.setdefault(tab.n_dim, []).append((b, tab, mp))
        if op.kind == "gp" and self.band_mapper is None:
            if len(self._plan_cache) > 64:
                self._plan_cache.clear()
            self._plan_cache[key] = (list(groups.values()), list(descs), [d.emulator for d in current_data])
        return list(groups.values()), descs

    def _tick(self, name):
        if self.profile is None:
            return None
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        self.profile.setdefault(name, []).append((e0, e1))
        e0.record()
        return e1

    @staticmethod
    def _tock(ev):
        if ev is not None:
            ev.record()

    def _global_norm(self, ws):
        """Convergence norm over ALL pixels (all ranks): linear_kf.py:293."""
        n_elems = self.n_state_elems * self.n_params
        if self.comm is None:
            ssq, self._n_bad_last, self._n_lu_last = ws.read_scalars()
            return float(np.sqrt(ssq) / float(n_elems)), float(n_elems)
        from .dist import global_norm
        norm, count, (self._n_bad_last, self._n_lu_last) = global_norm(ws.norm2, n_elems, self.comm,
                                                                       n_bad_local=ws.n_bad)
        return norm, count

    @staticmethod
    def _warn_counts(n_bad, n_bad_max, n_lu, n_lu_max):
        """The counts of the last solve; an earlier linearisation point may have met more."""
        if n_lu_max:
            LOG.warning("%d pixels had an information matrix that is not positive definite; solved by elimination "
                        "with partial pivoting, as the reference's sparse LU solves them%s", n_lu,
                        "" if n_lu == n_lu_max else " (%d in an earlier iteration)" % n_lu_max)
        if n_bad_max:
            LOG.warning("%d pixels had a singular (or non-finite) information matrix; left at the forecast"
                        "%s", n_bad, "" if n_bad == n_bad_max else " (%d in an earlier iteration)" % n_bad_max)

    def do_all_bands(self, timestep, current_data, x_forecast, P_forecast,
                     P_forecast_inverse, convergence_tolerance=1e-3,
                     min_iterations=2):
        """linear_kf.py:245-323: iterated linearisation.  Per iteration: the GP
        emulator kernel linearises every band around ``x_prev``
        (utils.py:181-219), the fused kernel builds and solves every pixel's
        normal equations (solvers.py:100-145) and leaves sum((x-x_prev)^2) on the
        device; the stop rule is the reference's (linear_kf.py:293-307)."""
        x_forecast = self._state(x_forecast)
        P_forecast_inverse = self._precision(P_forecast_inverse)
        n_bands = len(current_data)
        token = getattr(self, "_token", None)
        self._token = None
        groups, descs = self._band_plan(current_data)
        if self._small_image_ok(groups, n_bands):
            return self._do_all_bands_small(current_data, groups[0], token, x_forecast, P_forecast_inverse,
                                            convergence_tolerance, min_iterations)
        ws = self._workspace(n_bands)
        # the per-band outputs are handed to the caller at the end: fresh buffers per call
        # instead of a copy out of the workspace
        # (the last solve of the loop writes every pixel of every band row)
        if n_bands:
            ws.innov = rows_zero_padded(n_bands, self.n_state_elems, self.device)
        else:
            ws.innov = torch.zeros_like(ws.innov)
        if ws.fwd is not None:
            ws.fwd = rows_zero_padded(n_bands, self.n_state_elems, self.device) if n_bands else torch.zeros_like(ws.fwd)
        bufs = self._stager.finish(token if token is not None else self._stager.stage(current_data))
        all_linear = not groups and all(d is not None and d[0] == "linear" for d in descs)
        n, P = self.n_state_elems, self.n_params

        bufsx = [DeviceVector.empty(n, P, self.device), DeviceVector.empty(n, P, self.device)]
        a_out = BlockDiag.empty(n, P, self.device)
        # operator outputs, one buffer per emulator group, reused by every iteration
        gbufs = []
        for group in groups:
            D = group[0][1].n_dim
            idxs = [g[0] for g in group]
            gbufs.append((torch.empty((len(group), ws.ld), dtype=torch.float64, device=self.device),
                          torch.empty((len(group), D, ws.ld), dtype=torch.float64, device=self.device),
                          bufs.m if len(idxs) == n_bands else bufs.m[idxs].contiguous()))

        x_prev = x_forecast
        n_iter = 1
        n_bad = n_bad_max = 0                 # pixels the solve left at the forecast: last iteration / worst
        n_lu = n_lu_max = 0                   # pixels solved by pivoted elimination (block not positive definite)
        is_wcm = self._create_observation_operator.kind == "wcm"
        norms = []
        not_converged = True
        x_analysis = None
        cached = None                         # (sum of squares, count) of a repeatable solve
        while not_converged:
            if cached is None:
                shared = x_prev.uniform_value() if (groups and self.dedup_uniform_state) else None
                for group, (hsub, jsub, msub) in zip(groups, gbufs):
                    if shared is not None:
                        # every pixel is linearised at the same point (e.g. first iteration
                        # after a prior reset): evaluate the emulators ONCE, as the reference's
                        # run_emulator does for a single unique vector (utils.py:72-74)
                        one = DeviceVector.broadcast(shared, 1, self.device)
                        h1, j1 = linearize_bands(one, [(g[1], g[2]) for g in group], None, self.round_operator_f32)
                        h1, j1 = h1[:, 0].contiguous(), j1[:, :, 0].contiguous()
                        for k, g in enumerate(group):
                            descs[g[0]] = ("uniform", h1[k], j1[k], g[2])
                        continue
                    ev = self._tick("gp_linearize")
                    linearize_bands(x_prev, [(g[1], g[2]) for g in group], msub, self.round_operator_f32,
                                    h0=hsub, jac=jsub)
                    self._tock(ev)
                    for k, g in enumerate(group):
                        descs[g[0]] = ("arrays", hsub[k], jsub[k], g[2])
                x_new = bufsx[(n_iter - 1) % 2]
                ev = self._tick("solve_bands" if shared is None else "solve_bands_shared_linearisation")
                # an iteration below min_iterations cannot be the last one (linear_kf.py:296): its A
                # would be overwritten unread, so it is not stored
                engine.solve_bands(ws, descs, bufs, x_prev, x_forecast, P_forecast_inverse, x_new,
                                   a_out if (n_iter >= min_iterations or n_iter > 25 or all_linear) else None,
                                   want_fwd=self.want_forward_modelled, round_f32=self.round_operator_f32)
                self._tock(ev)
                if shared is None or n_iter >= 2:
                    # a full emulator pass is now queued: use the wait to stage the next date
                    self._prefetch_next()
                convergence_norm, count = self._global_norm(ws)   # linear_kf.py:293
                n_bad, n_lu = self._n_bad_last, self._n_lu_last
                n_bad_max, n_lu_max = max(n_bad_max, n_bad), max(n_lu_max, n_lu)
                if n_bad and is_wcm:
                    # the reference's forward model refuses such a linearisation point in the iteration
                    # that meets it (sar_forward_model.py:66-71)
                    raise ValueError("%d pixels left the Water-Cloud model's domain (LAI <= 0 or SM <= 0) or have "
                                     "a singular information matrix (iteration %d)" % (n_bad, n_iter))
                x_analysis = x_new
                x_analysis._uniform = False      # a solved state: skip the shared-vector test next iteration
            else:
                # a linear operator relinearised at its own solution reproduces it:
                # same system, same result, zero step (what the reference computes again)
                convergence_norm = 0.0
            norms.append(convergence_norm)
            LOG.info("Band {:d}, Iteration # {:d}, convergence norm: {:g}".format(
                n_bands - 1, n_iter, convergence_norm))
            if (convergence_norm < convergence_tolerance) and (n_iter >= min_iterations):
                not_converged = False
            elif n_iter > 25:
                LOG.warning("Bailing out after 25 iterations!!!!!!")
                not_converged = False
            if all_linear:
                cached = count
            x_prev = x_analysis
            n_iter += 1

        self._warn_counts(n_bad, n_bad_max, n_lu, n_lu_max)
        if (self.hessian_correction or getattr(self, "_force_hessian", False)) and groups:
            # P_analysis_inverse - P_correction (linear_kf.py:319,416), evaluated at x_analysis
            for group in groups:
                idxs = [g[0] for g in group]
                kf_tools._hessian_correction_device(
                    [current_data[b].emulator for b in idxs], x_analysis,
                    bufs.w[idxs].contiguous() if len(idxs) != n_bands else bufs.w,
                    ws.innov[idxs].contiguous() if len(idxs) != n_bands else ws.innov,
                    P, [g[2] for g in group], self.device, a_inout=a_out, want_corr=False)
        self.last_iterations.append(n_iter - 1)
        self.last_norms.append(norms)
        innovations = DeviceBands(ws.innov, n)
        self.fwd_modelled = DeviceBands(ws.fwd, n) if self.want_forward_modelled else None
        return x_analysis, None, a_out, innovations

    # ------------------------------------------------------------------
    # Small images (the 128 x 128 .. 256 x 256 chunks the reference's drivers hand to LinearKalman,
    # kafka_test_S2.py:202): the kernels of a date take microseconds, so what counts is how many
    # calls the host makes.  The whole iterated linearisation of a date then goes through ONE C-ABI
    # call (kafka_assimilate_date: the same kernels, the reference's loop and stop rule in C) on
    # buffers kept across dates.
    SMALL_IMAGE_PIXELS = 1 << 18
    SMALL_IMAGE_PATH_DEFAULT = True

    def _small_image_ok(self, groups, n_bands):
        return (self.small_image_path and self.comm is None and self.profile is None
                and len(groups) == 1 and len(groups[0]) == n_bands
                and 0 < self.n_state_elems <= self.SMALL_IMAGE_PIXELS
                and not self.hessian_correction and not getattr(self, "_force_hessian", False))

    def _while_queued_hook(self):
        """kafka_host_hook of kafka_assimilate_date: the next date's rasters are read and sent once
        this date's first iterations are queued, so that host work hides behind the kernels.  An
        exception cannot cross the C frame: it is kept and raised when the call returns."""
        hook = getattr(self, "_hook", None)
        if hook is None:
            import weakref
            self._hook_error = None
            me = weakref.ref(self)                  # no reference cycle: the filter (and the page-locked rasters its
                                                    # caches hold) must go away when the caller drops it, not at some
                                                    # later collection in the middle of another run

            def run(_arg):
                kf = me()
                if kf is None:
                    return
                try:
                    kf._prefetch_next()
                except BaseException as e:          # noqa: B036 -- re-raised by the caller of the C function
                    kf._hook_error = e
            hook = self._hook = _lib.HOST_HOOK(run)
        return hook

    def _do_all_bands_small(self, current_data, group, token, x_forecast, P_forecast_inverse, tol, min_iterations):
        import ctypes as C
        lib = _lib.get_lib()
        n, P, nb = self.n_state_elems, self.n_params, len(group)
        D = group[0][1].n_dim
        dev = self.device
        st = self._small.get((nb, D))
        if st is None:
            ld = rows_zero_padded(1, n, dev).shape[1]
            st = {"ld": ld, "bufs": engine.BandBuffers(nb, self.pixels, dev),
                  "h0": torch.empty((nb, ld), dtype=torch.float64, device=dev),
                  "jac": torch.empty((nb, D, ld), dtype=torch.float64, device=dev),
                  "scalars": torch.zeros(2, dtype=torch.float64, device=dev),
                  "pinned": torch.zeros(2 * 26, dtype=torch.float64).pin_memory(),      # 16 bytes per solve
                  "scratch": torch.empty(int(lib.kafka_solve_scratch_doubles(n)), dtype=torch.float64, device=dev),
                  "norms": (C.c_double * 32)(), "ints": (C.c_int * 4)(), "arr_key": None, "arr": None}
            st["ints_at"] = C.addressof(st["ints"])
            self._small[(nb, D)] = st
        ld = st["ld"]
        key = tuple((g[1].table.data_ptr(), g[1].n_train, tuple(int(v) for v in g[2])) for g in group)
        if st["arr_key"] != key:
            arr = (_lib.GpBand * nb)()
            for b, (_, tab, mp) in enumerate(group):
                arr[b].table, arr[b].n_train, arr[b].n_dim = tab.table.data_ptr(), tab.n_train, D
                for d in range(D):
                    arr[b].mapper[d] = int(mp[d])
            st["arr"], st["arr_key"] = arr, key
        bufs = self._stager.finish(token if token is not None else self._stager.stage(current_data), out=st["bufs"])
        # handed to the caller: fresh per date (the output stage may still be reading the previous ones)
        x_work = torch.empty((2, P, ld), dtype=torch.float64, device=dev)
        a_out = BlockDiag.empty(n, P, dev)
        innov = torch.empty((nb, ld), dtype=torch.float64, device=dev)
        fwd = torch.empty((nb, ld), dtype=torch.float64, device=dev) if self.want_forward_modelled else None
        flags = (_lib.PRIOR_BROADCAST if P_forecast_inverse.is_broadcast else 0) | \
            (_lib.ROUND_F32 if self.round_operator_f32 else 0) | _lib.SOLVE_COUNT_LU
        ints = st["ints"]
        x_uni = None
        if self.dedup_uniform_state and isinstance(x_forecast._uniform, np.ndarray):
            # all pixels start from one state (a prior reset): the first iteration evaluates the emulators once
            x_uni = getattr(x_forecast, "_uniform_dev", None)
            if x_uni is None:
                x_uni = x_forecast._uniform_dev = x_forecast.data[:, 0].contiguous()
        with torch.cuda.device(dev):
            _lib.check(lib.kafka_assimilate_date(
                n, ld, P, st["arr"], nb, _lib.ptr(bufs.y), _lib.ptr(bufs.w), _lib.ptr(bufs.m), _lib.ptr(x_forecast.data),
                _lib.ptr(P_forecast_inverse.data), _lib.ptr(st["h0"]), _lib.ptr(st["jac"]), _lib.ptr(x_work),
                _lib.ptr(a_out.data), _lib.ptr(innov), _lib.ptr(fwd), _lib.ptr(st["scalars"]),
                C.c_void_p(st["pinned"].data_ptr()), _lib.ptr(st["scratch"]), float(tol), int(min_iterations), flags,
                _lib.stream_ptr(), C.addressof(st["norms"]), st["ints_at"], st["ints_at"] + 4, st["ints_at"] + 8,
                st["ints_at"] + 12, _lib.ptr(x_uni), self._while_queued_hook(), None),
                "kafka_assimilate_date")
        if self._hook_error is not None:
            err, self._hook_error = self._hook_error, None
            raise err
        n_solves, n_bad, n_bad_max, idx = ints[0], ints[1], ints[2], ints[3]
        _lib.launch_count += 4 * n_solves     # kernels of the call: (linearise, solve, second stage, read) per solve
        # slot k of the page-locked buffer: [sum of squares f64 | left at the forecast i32 | solved by elimination i32]
        lu = st["pinned"].numpy().view(np.int32)[3:4 * n_solves:4]
        n_lu, n_lu_max = int(lu[-1]), int(lu.max())
        norms = [st["norms"][k] for k in range(n_solves)]
        if LOG.isEnabledFor(logging.INFO):
            for k, v in enumerate(norms):
                LOG.info("Band {:d}, Iteration # {:d}, convergence norm: {:g}".format(nb - 1, k + 1, v))
        if n_solves > 25:
            LOG.warning("Bailing out after 25 iterations!!!!!!")
        self._warn_counts(n_bad, n_bad_max, n_lu, n_lu_max)
        x_analysis = DeviceVector(x_work[idx], n, P)
        x_analysis._uniform = False
        self.last_iterations.append(n_solves)
        self.last_norms.append(norms)
        self.fwd_modelled = DeviceBands(fwd, n) if self.want_forward_modelled else None
        return x_analysis, None, a_out, DeviceBands(innov, n)

    # ------------------------------------------------------------------
    def assimilate(self, locate_times, x_forecast, P_forecast,
                   P_forecast_inverse,
                   approx_diagonal=True, refine_diag=False,
                   iter_obs_op=False, is_robust=False, diag_str="diag"):
        """linear_kf.py:325-354 (legacy band-by-band path): every band is
        assimilated in turn, the analysis of one band being the forecast of the
        next.  The reference's version relinearises around the forecast, applies
        a 7-parameter Hessian correction and plots (linear_kf.py:412-423); here
        each band goes through the same fused kernels with a single band."""
        for step in locate_times:
            LOG.info("Assimilating %s..." % step.strftime("%Y-%m-%d"))
            for band in range(self.observations.bands_per_observation[step]):
                x_analysis, P_analysis, P_analysis_inverse, innovations = \
                    self.assimilate_band(band, step, x_forecast, P_forecast, P_forecast_inverse)
                x_forecast, P_forecast, P_forecast_inverse = x_analysis, P_analysis, P_analysis_inverse
        self.previous_state = Previous_State(step, x_analysis, P_analysis, P_analysis_inverse)
        return x_analysis, P_analysis, P_analysis_inverse

    def assimilate_band(self, band, timestep, x_forecast, P_forecast,
                        P_forecast_inverse, convergence_tolerance=1e-3,
                        min_iterations=1):
        """linear_kf.py:356-425 through the multiband kernels with one band,
        including the Hessian correction the reference applies here (:412-416)."""
        data = self.observations.get_band_data(timestep, band)
        self._force_hessian = True
        try:
            return self._assimilate_band(band, timestep, data, x_forecast, P_forecast, P_forecast_inverse,
                                         convergence_tolerance, min_iterations)
        finally:
            self._force_hessian = False

    def _assimilate_band(self, band, timestep, data, x_forecast, P_forecast, P_forecast_inverse,
                         convergence_tolerance, min_iterations):
        op = self._create_observation_operator
        if op.kind == "gp" and (op.state_mappers is not None or self.band_mapper is not None):
            # the band index selects the state mapper: keep it by padding the plan
            saved = self._band_plan

            def plan(current):
                tab = emulator_table(current[0].emulator, self.device)
                mp = (np.asarray(self.band_mapper[band], dtype=np.int32)
                      if (op.state_mappers is None and self.band_mapper is not None)
                      else op.mapper(self.n_params, band, tab.n_dim))
                return [[(0, tab, mp)]], [None]
            self._band_plan = plan
            try:
                return self.do_all_bands(timestep, [data], x_forecast, P_forecast, P_forecast_inverse,
                                         convergence_tolerance, min_iterations)
            finally:
                self._band_plan = saved
        return self.do_all_bands(timestep, [data], x_forecast, P_forecast, P_forecast_inverse,
                                 convergence_tolerance, min_iterations)

    def solver(self, observations, mask, H_matrix, x_forecast, P_forecast,
               P_forecast_inv, R_mat, the_metadata):
        """linear_kf.py:427-438."""
        from .inference.solvers import variational_kalman
        return variational_kalman(observations, mask, self.state_mask, R_mat, H_matrix, self.n_params,
                                  x_forecast, P_forecast, P_forecast_inv, the_metadata)

    def solver_multiband(self, observations, mask, H_matrix, x0, x_forecast, P_forecast,
                         P_forecast_inv, R_mat, the_metadata):
        """linear_kf.py:441-452."""
        from .inference.solvers import variational_kalman_multiband
        return variational_kalman_multiband(observations, mask, self.state_mask, R_mat, H_matrix,
                                            self.n_params, x0, x_forecast, P_forecast, P_forecast_inv,
                                            the_metadata)
