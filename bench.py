#!/usr/bin/env python
"""Benchmark of the KaFKA assimilation step on B200 (see DESIGN.md, section
"Measurement").

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Workload (config.workload): one GPU's shard of BASELINE.json configs[3] -- the
full Sentinel-2 tile (10980 x 10980 px) sharded by pixel over 8 GPUs, i.e.
1373 x 10980 = 15 075 540 px per GPU -- with the 10-parameter PROSAIL state, 10
bands, one 250-point GP emulator per band, iterated linearisation.  Weak
scaling: every rank holds one such shard, N=8 is the full tile.  A *step* is one
observation date assimilated for every pixel of the shard (all bands, all
iterations); metric = pixel-dates/s summed over ranks.

* value : inputs resident in HBM (device rasters), CUDA-event timed
* e2e   : LinearKalman.run() with HOST observation rasters; host->device copies
          and the device->host read of the posterior inside the timed region
* roofline / cpu_baseline : see DESIGN.md
"""
import argparse
import ctypes
import datetime as dt
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

P, B, T = 10, 10, 250
TILE = 10980
SHARD_ROWS = 1373                     # ceil(10980 / 8)
BYTES_PER_PIXEL_DATE = 16 * P + 8 * P * (P + 1) + 17 * B          # 1210
FLOPS_PER_PIXEL_ITER = B * T * (5 * P + 28) + B * (P * P + 3 * P) + P ** 3 / 3 + 2 * P * P
REF_CHUNK_SIDE = 64          # reference chunks: REF_CHUNK_SIDE^2 px, one LinearKalman each (drivers use <= 256^2)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=SHARD_ROWS, help="raster rows per GPU (x 10980 columns)")
    ap.add_argument("--cols", type=int, default=TILE)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ref-chunk", type=int, default=REF_CHUNK_SIDE,
                    help="side of the square chunks the CPU legs hand to the reference (<= 256, its drivers' practice)")
    ap.add_argument("--port-sample", type=int, default=65536, help="pixels the numpy port is timed on (cpu_baseline.port)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true",
                    help="leave configs2_modis_c3 / configs4_sweep_c5 out of the default line")
    ap.add_argument("--host-sources", default="pinned", choices=["pinned", "pageable"],
                    help="memory the e2e legs' host rasters live in (see e2e.api)")
    ap.add_argument("--comm", default="native", choices=["native", "torch"],
                    help="who carries the per-iteration convergence all-reduce at N > 1: the library's own NCCL "
                         "communicator behind the C ABI (kafka_ctx_create / kafka_allreduce_norm) or torch.distributed")
    ap.add_argument("--c3-days", type=int, default=365, help="daily steps of the c3 workload")
    ap.add_argument("--c3-side", type=int, default=2400, help="side of the square c3 tile per GPU")
    ap.add_argument("--workload", default="c4shard", choices=["c4shard", "c2", "c3", "c5"],
                    help="c4shard (default, the bench line): 1/8 of the full S2 tile per GPU; the others are the "
                         "remaining BASELINE.json configs, run on request for the record (single GPU)")
    return ap.parse_args()


# --------------------------------------------------------------------------
# synthetic workload
# --------------------------------------------------------------------------
def make_emulators(seed=1234):
    from kafka_inferenceengine_b200 import synthetic as syn
    rng = np.random.default_rng(seed)
    return [syn.make_emulator(rng, syn.SAIL_MEAN, syn.SAIL_SIGMA, n_train=T) for _ in range(B)]


def synth_on_device(gps, n_pix, seed, device):
    """x_true ~ N(prior), reflectance = emulator(x_true) + 5 % noise, observed with
    probability 0.9 and rho > 0, inverse variance 1/(0.05 rho)^2 (inf where
    masked) -- Sentinel2_Observations.py:168-178 shaped, generated on the device."""
    import torch
    from kafka_inferenceengine_b200 import synthetic as syn
    from kafka_inferenceengine_b200.device import DeviceVector
    from kafka_inferenceengine_b200.inference.utils import emulator_table, linearize_bands
    gen = torch.Generator(device=device).manual_seed(seed)
    xt = DeviceVector.empty(n_pix, P, device)
    mean = torch.from_numpy(syn.SAIL_MEAN).to(device)[:, None]
    sig = torch.from_numpy(syn.SAIL_SIGMA).to(device)[:, None]
    xt.data.copy_(mean + sig * torch.randn(xt.data.shape, generator=gen, dtype=torch.float64, device=device))
    tabs = [(emulator_table(g, device), np.arange(P)) for g in gps]
    rho = torch.empty((B, xt.ld), dtype=torch.float64, device=device)
    linearize_bands(xt, tabs, None, False, h0=rho, want_jac=False)       # forward model only
    rho = rho[:, :n_pix].contiguous()
    rho = rho * (1.0 + 0.05 * torch.randn(rho.shape, generator=gen, dtype=torch.float64, device=device))
    mask = (rho > 0) & (torch.rand(rho.shape, generator=gen, device=device) < 0.9)
    rho = torch.where(mask, rho, torch.zeros_like(rho))
    inv = torch.where(mask, 1.0 / (0.05 * rho) ** 2, torch.full_like(rho, float("inf")))
    return rho, inv, mask.to(torch.uint8)


class DeviceObservations(object):
    """Observation collaborator handing out DEVICE rasters (inputs resident in HBM)."""

    def __init__(self, dates, gps, rho, inv, mask, shape):
        from kafka_inferenceengine_b200 import synthetic as syn
        self.dates = list(dates)
        self.bands_per_observation = {d: B for d in self.dates}
        self._bands = [syn.BandData(rho[b].view(shape), inv[b].view(shape), mask[b].view(shape), {"band": b}, gps[b])
                       for b in range(B)]

    def get_band_data(self, date, band):
        return self._bands[band]


class HostObservations(object):
    """Same data as host numpy rasters (the reference's data contract:
    observations (ny,nx) float64, inverse-variance raster, boolean mask)."""

    def __init__(self, dates, gps, rho, inv, mask, shape, pinned=True):
        from kafka_inferenceengine_b200 import synthetic as syn
        self.dates = list(dates)
        self.bands_per_observation = {d: B for d in self.dates}
        self._bands = [syn.BandData(_host_raster(rho[b], shape, np.float64, pinned),
                                    _host_raster(inv[b], shape, np.float64, pinned),
                                    _host_raster(mask[b] != 0, shape, np.bool_, pinned), {"band": b}, gps[b])
                       for b in range(B)]

    def get_band_data(self, date, band):
        return self._bands[band]


def _host_raster(t, shape, dtype, pinned):
    """Host copy of a device row as a (ny, nx) numpy raster: in page-locked memory obtained from
    the package (kb.pinned_raster -- what a reader fills when it wants the DMA engine to fetch its
    rasters in place) or in ordinary pageable memory."""
    if not pinned:
        return t.cpu().numpy().reshape(shape).astype(dtype, copy=False)
    import torch
    from kafka_inferenceengine_b200 import pinned_raster
    a = pinned_raster(shape, dtype)
    torch.from_numpy(a.reshape(-1)).copy_(t.reshape(-1))
    return a


class RawObservations(object):
    """Raw scaled-integer reflectance as the S2 granule holds it (int16 = rho * 10000, <= 0 where
    there is no data): the ingest arithmetic of Sentinel2Observations.get_band_data runs on the device."""

    def __init__(self, dates, gps, rho, mask, shape, pinned=True):
        from kafka_inferenceengine_b200 import synthetic as syn
        from kafka_inferenceengine_b200 import ScaledReflectance
        self.dates = list(dates)
        self.bands_per_observation = {d: B for d in self.dates}
        raw = torch_round_i16(rho, mask)
        self._bands = [syn.BandData(_host_raster(raw[b], shape, np.int16, pinned), ScaledReflectance(10000.0, 0.05),
                                    None, {"band": b}, gps[b]) for b in range(B)]

    def get_band_data(self, date, band):
        return self._bands[band]


def torch_round_i16(rho, mask):
    import torch
    r = torch.clamp(torch.round(rho * 10000.0), 0, 32767)
    return torch.where(mask != 0, r, torch.zeros_like(r)).to(torch.int16)


class ReadbackOutput(object):
    """Output collaborator: reads the posterior mean and diag(P^-1) back to the
    host (what KafkaOutput.dump_data consumes, observations.py:374-375,392-393)."""

    def __init__(self):
        self.d2h_bytes = 0
        self.last = None

    def dump_data(self, timestep, x_analysis, P_analysis, P_analysis_inv, state_mask, n_params):
        x = np.asarray(x_analysis)
        d = P_analysis_inv.diagonal()
        self.d2h_bytes += x.nbytes + d.nbytes
        self.last = (x, d)


class NullOutput(object):
    def dump_data(self, *a):
        pass


# --------------------------------------------------------------------------
# clocks
# --------------------------------------------------------------------------
class ClockSampler(object):
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [c.strip() for c in r.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if f[2 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------
# CPU legs.  Nothing below imports kafka_inferenceengine_b200 (no CUDA library is mapped by
# the processes that time the CPU code): the workload comes from oracle/synthetic.py, the
# timed code is the UNMODIFIED reference (oracle/ref_runner.py -> baseline/_ref or
# /root/reference) and, beside it, the numpy port (oracle A': scipy sparse + float32 SuperLU).
# --------------------------------------------------------------------------
def _port_chunk(args):
    gp_arrays, y, w, m = args
    from oracle import kafka_cpu as kc
    from oracle import synthetic as osyn
    from oracle.gp import GaussianProcess
    gps = [GaussianProcess(a[0], a[1], a[2]) for a in gp_arrays]
    n = y.shape[1]
    inv = osyn.sail_prior_block().astype(np.float64)
    blocks = np.repeat(inv[None], n, axis=0)
    x0 = np.tile(osyn.SAIL_MEAN, n)
    ops = [(lambda xp, gp=gps[b], mm=m[b]: kc.emulator_operator(gp, xp, P, mm)) for b in range(B)]
    t0 = time.perf_counter()
    x, A, _, iters, _ = kc.do_all_bands(ops, [y[b] for b in range(B)], [w[b] for b in range(B)],
                                        [m[b] for b in range(B)], x0, blocks, P, mode="sparse32")
    return time.perf_counter() - t0, n, iters


def _ref_chunk(args):
    from oracle import ref_runner
    return ref_runner.time_do_all_bands(args)


def _port_job(args):
    """A reference-style job tuple through the numpy port (stand-in where the reference package
    is not installed: kind "port")."""
    gp_arrays, y, inv, m, side = args
    return _port_chunk((gp_arrays, y, np.where(m, inv, 0.0), m))


def _reference_installed():
    from oracle import ref_shim
    return ref_shim.available()


def _ref_solve_chunk(args):
    from oracle import ref_runner
    return ref_runner.time_multiband_solve(args)


def _cpu_workload(n_chunks, side, seed=99):
    """Chunks of the benchmark workload for the CPU legs, from the numpy generator."""
    from oracle import synthetic as osyn
    gps = osyn.make_emulators(B, T)
    gp_arrays = [(g.inputs, g.targets, g.theta) for g in gps]
    n = side * side
    y, inv, m = osyn.synth_observations(gps, n_chunks * n, seed)
    return [(gp_arrays, y[:, k * n:(k + 1) * n], inv[:, k * n:(k + 1) * n], m[:, k * n:(k + 1) * n], side)
            for k in range(n_chunks)]


def _single_thread_env():
    for var in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[var] = "1"              # one BLAS thread per worker process (spawned children inherit)


def cpu_baseline_leg(side=REF_CHUNK_SIDE, port_sample=65536):
    """cpu_baseline of the N=1 line: a bounded sample on ONE host core, run in a freshly spawned
    interpreter (no CUDA library in that process): (A) the unmodified reference's
    LinearKalman.do_all_bands on one chunk, (B) its variational_kalman_multiband alone on the same
    chunk, and the numpy port on a larger sample."""
    import multiprocessing as mp
    _single_thread_env()
    jobs = _cpu_workload(1, side)
    have_ref = _reference_installed()
    with mp.get_context("spawn").Pool(1) as pool:
        if have_ref:
            t_a, n_a, it_a = pool.apply(_ref_chunk, (jobs[0],))
            t_b, n_b = pool.apply(_ref_solve_chunk, (jobs[0],))
        gp_arrays = jobs[0][0]
        from oracle import synthetic as osyn
        gps = osyn.make_emulators(B, T)
        y, inv, m = osyn.synth_observations(gps, port_sample, 98)
        w = np.where(m, inv, 0.0)
        res = [pool.apply(_port_chunk, ((gp_arrays, y[:, k:k + 16384], w[:, k:k + 16384], m[:, k:k + 16384]),))
               for k in range(0, port_sample, 16384)]
    t_p = sum(r[0] for r in res)
    port = {"value": port_sample / t_p, "unit": "pixel-dates/s", "cores": 1, "kind": "port",
            "sample": "%d px in chunks of 16384 px, %.1f s; oracle A' = numpy GP + scipy.sparse build + "
                      "float32 SuperLU as solvers.py:100-145 (%s solves)"
                      % (port_sample, t_p, sorted(set(r[2] for r in res)))}
    if not have_ref:
        return port                                  # no reference package on this machine (baseline/_ref absent)
    return {"value": n_a / t_a, "unit": "pixel-dates/s", "cores": 1, "kind": "reference",
            "sample": "one %dx%d px chunk of the same workload (all bands, iterated to convergence: %d solves) "
                      "through the UNMODIFIED reference's LinearKalman.do_all_bands (kafka/linear_kf.py:245-323, "
                      "operators by create_prosail_observation_operator, emulator class = oracle.gp), %.1f s on one "
                      "core" % (side, side, it_a, t_a),
            "solve_only": {"value": n_b / t_b, "unit": "pixels/s", "seconds": t_b,
                           "what": "the unmodified variational_kalman_multiband (kafka/inference/solvers.py:100-145) "
                                   "alone on the same chunk, operators pre-built: the 'scipy spsolve path'"},
            "port": port}


def run_reference(args):
    """--impl reference: the UNMODIFIED reference (LinearKalman.do_all_bands from baseline/_ref) on all
    host cores, one chunk per core per step, one LinearKalman per chunk like the reference's dask fan-out
    (kafka_test_Py36.py:241-255).  Runs on rank 0 only; no GPU, no CUDA library, none of the product code."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    _single_thread_env()
    cores = os.cpu_count() or 1
    procs = max(1, min(cores, 64))
    side = args.ref_chunk
    jobs = _cpu_workload(procs, side)
    n_sample = procs * side * side
    times, iters = [], []
    have_ref = _reference_installed()
    step_fn = _ref_chunk if have_ref else _port_job
    with mp.get_context("spawn").Pool(procs) as pool:
        if have_ref:
            pool.map(_ref_solve_chunk, jobs)              # imports + first-call costs, outside every step
        for s in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            res = pool.map(step_fn, jobs, chunksize=1)
            wall = time.perf_counter() - t0
            if s >= args.warmup:
                times.append(wall)
                iters += [r[2] for r in res]
        # side figures: the slowest process's own timer (inputs and operators are built outside it)
        wall_b = max(r[0] for r in pool.map(_ref_solve_chunk, jobs, chunksize=1)) if have_ref else float("nan")
        w_jobs = [(j[0], j[1], np.where(j[3], j[2], 0.0), j[3]) for j in jobs]
        wall_p = max(r[0] for r in pool.map(_port_chunk, w_jobs, chunksize=1))
    ms = 1e3 * float(np.mean(times))
    value = n_sample / (ms * 1e-3)
    sample = ("per step %d chunks of %dx%d px of the same workload (one per process, %d processes = all host "
              "cores, BLAS pinned to one thread each), every chunk through the UNMODIFIED reference's "
              "LinearKalman.do_all_bands (kafka/linear_kf.py:245-323; %s solves per chunk); rate per pixel, "
              "the workload itself is %d px per GPU" % (procs, side, side, procs, sorted(set(iters)),
                                                         args.rows * args.cols))
    line = {"impl": "reference", "metric": "pixel-dates/sec", "value": value, "unit": "pixel-dates/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args),
            "cpu_baseline": {"value": value, "unit": "pixel-dates/s", "cores": procs,
                             "kind": "reference" if have_ref else "port",
                             "sample": sample if have_ref else sample.replace(
                                 "UNMODIFIED reference's LinearKalman.do_all_bands",
                                 "numpy port (oracle A'; the reference package is not installed here) of do_all_bands"),
                             "arithmetic": "the reference builds in float64 and solves in float32 (solvers.py:127-134)",
                             "solve_only": {"value": n_sample / wall_b, "unit": "pixels/s",
                                            "what": "unmodified variational_kalman_multiband alone, same chunks, "
                                                    "operators pre-built, all cores"},
                             "port": {"value": n_sample / wall_p, "unit": "pixel-dates/s", "kind": "port",
                                      "what": "numpy port (oracle A') on the same chunks, all cores"}},
            "e2e": {"value": value, "unit": "pixel-dates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(json.dumps(line))


def workload_config(args, note=None):
    cfg = {"workload": "BASELINE configs[3] shard: full Sentinel-2 tile 10980x10980 px sharded by pixel over 8 GPUs "
                       "= %d x %d px per GPU; 10-param PROSAIL state, 10 bands, 250-point GP emulator per band, "
                       "iterated linearisation (tol 1e-3, >=2 solves), prior-only advance" % (args.rows, args.cols),
           "pixels_per_gpu": args.rows * args.cols, "n_params": P, "n_bands": B, "n_train": T,
           "l2_policy": "inputs larger than L2 (>2 GB of rasters per step vs 126 MB L2)"}
    if note:
        cfg["note"] = note
    return cfg


# --------------------------------------------------------------------------
# --------------------------------------------------------------------------
# the other BASELINE.json configurations (on request; the default bench line is c4shard)
# --------------------------------------------------------------------------
def measure_c2(device, reps):
    """BASELINE configs[1]: Barrax mask (13 027 px), 10-band PROSAIL emulators, 20 dates through ONE
    LinearKalman.run() with host rasters, timed end to end on the host clock (the GPU is idle most of
    the time at this size: the number measures launch and host overhead, not the kernels)."""
    import torch
    import kafka_inferenceengine_b200 as kb
    from kafka_inferenceengine_b200 import synthetic as syn
    from kafka_inferenceengine_b200.device import BlockDiag, DeviceVector
    from kafka_inferenceengine_b200.output import RasterOutput
    base = dt.datetime(2017, 1, 1)
    g = np.load(os.path.join(ROOT, "tests", "golden", "barrax_mask.npz"))
    ny, nx = g["shape"]
    mask = np.unpackbits(g["bits"])[:ny * nx].astype(bool).reshape(ny, nx)
    N = int(mask.sum())
    gps = make_emulators()
    rho, inv, msk = synth_on_device(gps, ny * nx, 7, device)
    n_dates = 20
    dates = [base + dt.timedelta(days=2 * d, hours=12) for d in range(n_dates)]
    obs = HostObservations(dates, gps, rho, inv, msk, (ny, nx))
    inv_block = np.diag(1.0 / syn.SAIL_SIGMA ** 2).astype(np.float32).astype(np.float64)
    prior = syn.BroadcastPrior(syn.SAIL_MEAN, inv_block, N, lambda blk, n: BlockDiag.from_block(blk, n, device),
                               lambda m, n: DeviceVector.broadcast(m, n, device))
    out = RasterOutput(syn.S2_PARAMETERS)
    kf = kb.LinearKalman(obs, out, mask, kb.create_prosail_observation_operator, syn.S2_PARAMETERS,
                         state_propagation=None, prior=prior, linear=False, device=device)
    out.pixels = kf.pixels
    kf.set_trajectory_model()
    kf.set_trajectory_uncertainty(0.0)
    x0, P0 = prior.process_prior(None)
    grid = [base + dt.timedelta(days=2 * d) for d in range(n_dates + 1)]
    kf.run(grid, x0, None, P0, iter_obs_op=True)            # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        kf.run(grid, x0, None, P0, iter_obs_op=True)
    torch.cuda.synchronize()
    ms = 1e3 * (time.perf_counter() - t0) / reps
    return {"workload": "c2: Barrax mask 13027 px, P=10, B=10, T=250, 20 dates, LinearKalman.run with host "
                        "rasters, end to end", "metric": "pixel-dates/sec",
            "value": N * n_dates / (ms * 1e-3), "ms_per_run": ms, "ms_per_date": ms / n_dates,
            "iterations": kf.last_iterations[-n_dates:], "n_gpus": 1}


def _dist_setup(args):
    """(world, rank, device, comm, barrier, max_over_ranks) for the per-rank workloads."""
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1 and args.gpus > 1:
        raise SystemExit("launch with torch.distributed.run --nproc-per-node %d" % args.gpus)
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    comm = None
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=device)
        comm = _make_comm(args, device)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return world, rank, local, device, comm, barrier, max_over_ranks


def _make_comm(args, device):
    """torch.distributed is the plumbing (rendezvous, barrier, max over ranks); the hot path's all-reduce
    goes through the library's own communicator unless --comm torch asks otherwise."""
    import torch.distributed as dist
    if args.comm == "torch":
        return dist.group.WORLD
    from kafka_inferenceengine_b200.dist import NativeComm
    return NativeComm.from_process_group(dist.group.WORLD, device)


def _hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def run_c3(args, ctx=None):
    """BASELINE configs[2]: MODIS tile 2400 x 2400 (5.76 M px) per GPU, 7-parameter TIP state, 2 broadbands
    through per-band state mappers (D=4, T=500), --c3-days daily steps (365) with an observation every 8th
    day, propagate_information_filter_LAI with Q[6::7] = 0.04 (kafka_test.py:156-217,
    kafka/inference/kf_tools.py:292-314).  End to end: ONE LinearKalman.run() fed with HOST rasters on the
    observation days, RasterOutput reading the float32 mean / uncertainty rasters back EVERY day.  Weak
    scaling: one tile per rank, global convergence norm."""
    import torch
    import kafka_inferenceengine_b200 as kb
    from kafka_inferenceengine_b200 import synthetic as syn
    from kafka_inferenceengine_b200.device import DeviceVector
    from kafka_inferenceengine_b200.inference.utils import emulator_table, linearize_bands
    from kafka_inferenceengine_b200.output import RasterOutput
    world, rank, local, device, comm, barrier, max_over_ranks = ctx or _dist_setup(args)
    base = dt.datetime(2017, 1, 1)
    ny = nx = args.c3_side
    N = ny * nx
    rng = np.random.default_rng(3)
    mappers = [np.array([0, 1, 6, 2]), np.array([3, 4, 6, 5])]
    gps = [syn.make_emulator(rng, syn.TIP_MEAN[m], syn.TIP_SIGMA[m], n_train=500, w_scale=0.15) for m in mappers]
    gen = torch.Generator(device=device).manual_seed(11 + rank)
    xt = DeviceVector.empty(N, 7, device)
    xt.data.copy_(torch.from_numpy(syn.TIP_MEAN).to(device)[:, None] + 0.5 * torch.from_numpy(syn.TIP_SIGMA).to(device)[:, None]
                  * torch.randn(xt.data.shape, generator=gen, dtype=torch.float64, device=device))
    rho = torch.empty((2, xt.ld), dtype=torch.float64, device=device)
    linearize_bands(xt, [(emulator_table(g_, device), m) for g_, m in zip(gps, mappers)], None, False, h0=rho,
                    want_jac=False)
    rho = rho[:, :N].contiguous()
    sig = torch.clamp(0.05 * rho.abs(), min=2.5e-3)
    rho = rho + sig * torch.randn(rho.shape, generator=gen, dtype=torch.float64, device=device)
    msk = torch.rand(rho.shape, generator=gen, device=device) < 0.9
    inv = torch.where(msk, 1.0 / sig ** 2, torch.full_like(rho, float("inf")))
    n_steps = args.c3_days
    dates = [base + dt.timedelta(days=d, hours=12) for d in range(0, n_steps, 8)]
    pinned = args.host_sources == "pinned"
    obs = HostObservationsGeneric(dates, gps, rho, inv, msk, (ny, nx), pinned)
    del rho, inv, xt
    out = RasterOutput(syn.TIP_PARAMETERS)
    kf = kb.LinearKalman(obs, out, np.ones((ny, nx), dtype=bool), kb.create_nonlinear_observation_operator,
                         syn.TIP_PARAMETERS, state_propagation=kb.propagate_information_filter_LAI, prior=None,
                         device=device, comm=comm)
    out.pixels = kf.pixels
    kf.set_trajectory_model()
    q = DeviceVector.empty(N, 7, device)
    q.data.zero_()
    q.data[6] = 0.04
    kf.trajectory_uncertainty = kb.DiagonalMatrix(q, 7, device)
    mean, prior_inv = kb.tip_prior_full({"n_pixels": N})
    grid = [base + dt.timedelta(days=d) for d in range(n_steps + 1)]
    kf.run(grid[:10], mean, None, prior_inv, iter_obs_op=True)          # warm-up: two observation dates
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    kf.last_iterations = []
    h2d0, d2h0 = kf._stager.h2d_bytes, out.d2h_bytes
    t0 = time.perf_counter()
    kf.run(grid, mean, None, prior_inv, iter_obs_op=True)
    torch.cuda.synchronize()
    ms = max_over_ranks(1e3 * (time.perf_counter() - t0))
    clocks = sampler.stop() if sampler else None
    barrier()
    iters = list(kf.last_iterations)

    # parity inside the bench: the first 2048 pixels of this rank's tile, first observation date, against
    # oracle B with the iteration count the GPU run took on that date (the global norm decided it)
    parity = None
    if rank == 0:
        from oracle import kafka_cpu as kc
        from oracle.gp import GaussianProcess
        ns = 2048
        d0 = dates[0]
        bd = [obs.get_band_data(d0, b) for b in range(2)]
        ys = [np.asarray(b.observations).reshape(-1)[:ns] for b in bd]
        ms_ = [np.asarray(b.mask).reshape(-1)[:ns] for b in bd]
        ws = [kc.gate_inverse_variance(np.asarray(b.uncertainty).reshape(-1)[:ns], m_) for b, m_ in zip(bd, ms_)]
        ogps = [GaussianProcess(g_.inputs, g_.targets, g_.theta) for g_ in gps]
        ops = [(lambda xp, gp=ogps[b], m_=ms_[b], mp=mappers[b]: kc.emulator_operator(gp, xp, 7, m_, state_mapper=mp))
               for b in range(2)]
        x0n, _, inv_n = kc.tip_prior()
        xo, Ao, _, _, _ = kc.do_all_bands(ops, ys, ws, ms_, np.tile(x0n, ns), np.repeat(inv_n[None].astype(np.float64), ns, axis=0),
                                          7, fixed_iterations=iters[0])
        sub = [b._replace(observations=np.asarray(b.observations).reshape(-1)[:ns].reshape(1, ns),
                          uncertainty=np.asarray(b.uncertainty).reshape(-1)[:ns].reshape(1, ns),
                          mask=np.asarray(b.mask).reshape(-1)[:ns].reshape(1, ns)) for b in bd]
        kf_s = kb.LinearKalman(None, None, np.ones((1, ns), dtype=bool), kb.create_nonlinear_observation_operator,
                               syn.TIP_PARAMETERS, prior=None, device=device)
        m_s, p_s = kb.tip_prior_full({"n_pixels": ns})
        xg, _, Ag, _ = kf_s.do_all_bands(d0, sub, m_s, None, p_s, min_iterations=iters[0],
                                         convergence_tolerance=float("inf"))
        ex = float(np.max(np.abs(np.asarray(xg) - xo) / np.maximum(np.abs(xo), 1e-3)))
        idx = np.arange(7)
        ed = float(np.max(np.abs(Ag.blocks()[:, idx, idx] - Ao[:, idx, idx]) / Ao[:, idx, idx]))
        assert ex < 1e-6 and ed < 1e-6, (ex, ed)
        parity = {"pixels": ns, "date": 0, "solves": iters[0], "max_rel_err_x": ex, "max_rel_err_diag": ed,
                  "against": "oracle B (fp64), same number of solves"}
    if rank != 0:
        return None
    adv_bytes = N * (24 + 8 * 7 + 4 * 7 * 8)
    return {"workload": "c3: MODIS tile %dx%d (%.2f M px) per GPU, TIP P=7, B=2 (D=4, T=500), %d daily steps, "
                        "observations every 8th day (%d dates), LAI propagator with Q[6::7]=0.04; one "
                        "LinearKalman.run(): host rasters in on observation days (%s), float32 mean/uncertainty "
                        "rasters read back EVERY day" % (ny, nx, N / 1e6, n_steps, len(dates), args.host_sources),
            "metric": "pixel-steps/sec", "value": world * N * n_steps / (ms * 1e-3), "unit": "pixel-steps/s",
            "n_gpus": world, "scaling": "weak", "ms_per_run": ms, "ms_per_step": ms / n_steps,
            "pixel_dates_per_s": world * N * len(dates) / (ms * 1e-3), "iterations": iters,
            "h2d_bytes_per_run": kf._stager.h2d_bytes - h2d0, "d2h_bytes_per_run": out.d2h_bytes - d2h0,
            "d2h_gbs_per_rank": (out.d2h_bytes - d2h0) / (ms * 1e-3) / 1e9,
            "advance_algorithmic_bytes_per_step": adv_bytes, "clocks": clocks, "parity": parity}


C5_BANDS = (1, 4, 10, 13)


def _c5_cpu_arm(params, bands, procs):
    """Arm (B) beside the sweep: the UNMODIFIED variational_kalman_multiband (kafka/inference/solvers.py:100-145:
    sparse build, float32 cast, SuperLU) with supplied Jacobians, on chunks of 128 x 128 px (<= 256^2), three
    chunks on one core and one chunk per core on all cores; the rate per pixel is what the sweep's pixel
    counts are extrapolated with."""
    import multiprocessing as mp
    _single_thread_env()
    rows = []
    with mp.get_context("spawn").Pool(procs) as pool:
        pool.map(_c5_cpu_chunk, [(2, 1, 8, 0)] * procs)            # imports, outside the timing
        for Pn in params:
            for Bn in bands:
                one = [pool.apply(_c5_cpu_chunk, ((Pn, Bn, 128, k),)) for k in range(3)]
                t0 = time.perf_counter()
                allc = pool.map(_c5_cpu_chunk, [(Pn, Bn, 128, 10 + k) for k in range(procs)], chunksize=1)
                wall = time.perf_counter() - t0
                rows.append({"n_params": Pn, "n_bands": Bn, "chunk_px": 128 * 128,
                             "px_per_s_one_core": 3 * 128 * 128 / sum(r[0] for r in one),
                             "px_per_s_all_cores": procs * 128 * 128 / max(r[0] for r in allc), "cores": procs,
                             "wall_s_all_cores": wall})
    return rows


def _c5_cpu_chunk(args):
    Pn, Bn, side, seed = args
    from oracle import ref_runner
    return ref_runner.time_supplied_jacobian_solve(Pn, Bn, side, seed)


def run_c5(args, ctx=None, params=None, sizes=None, bands=None, cpu_arm=True):
    """BASELINE configs[4]: batched update with SUPPLIED Jacobians (the solve kernels alone), n_params 2-16 x
    bands {1,4,10,13} x pixels 1e4..1e8 (or what fits in memory) per GPU; with --gpus N every rank solves its
    own pixel range and the convergence norm is all-reduced after every launch, as do_all_bands does."""
    import torch
    import torch.distributed as dist
    from kafka_inferenceengine_b200 import _lib
    from kafka_inferenceengine_b200.device import padded, tri_count
    world, rank, local, device, comm, barrier, max_over_ranks = ctx or _dist_setup(args)
    lib = _lib.get_lib()
    hbm, peak_src = _hbm_peak()
    params = params or [int(v) for v in os.environ.get("KAFKA_C5_PARAMS", ",".join(str(k) for k in range(2, 17))).split(",")]
    sizes = sizes or [int(float(v)) for v in os.environ.get("KAFKA_C5_PIXELS", "1e4,1e5,1e6,1e7,1e8").split(",")]
    bands = bands or C5_BANDS
    # KAFKA_C5_FLAGS: kafka_solve_bands flags, e.g. 4 / 8 to force the two- / four-lanes-per-pixel kernel
    c5_flags = int(os.environ.get("KAFKA_C5_FLAGS", "0")) | (_lib.SOLVE_PAIR if os.environ.get("KAFKA_C5_PAIR") else 0)
    budget = float(os.environ.get("KAFKA_C5_MEM_GB", "120")) * 1e9
    gen = torch.Generator(device=device).manual_seed(3 + rank)
    sampler = ClockSampler(local) if rank == 0 else None
    rows, parity_worst = [], 0.0
    tot = torch.zeros(2, dtype=torch.float64, device=device)

    def fill(t):
        # one random plane per tensor, the others scaled and shifted copies of it: the kernel sees distinct
        # addresses and values, generating them costs one pass of writes instead of a normal variate per element
        base_ = torch.randn(t.shape[-1], generator=gen, dtype=torch.float64, device=device)
        for k in range(t.shape[0]):
            torch.mul(base_.roll(1 + 7 * k), 1.0 + 0.03 * k, out=t[k])
        return t

    for n_req in sizes:
        for Pn in params:
            Lp = tri_count(Pn)
            for Bn in bands:
                bytes_pp = 16 * Pn + 8 * Pn * (Pn + 1) + 8 * Pn + Bn * 8 * (3 + Pn)   # x_f,x_prev,x ; Pinv in + A out ; bands
                n = int(min(n_req, budget // bytes_pp // 1024 * 1024))
                if n < n_req and any(r["n_params"] == Pn and r["n_bands"] == Bn and r["pixels"] == n for r in rows):
                    continue                                   # the capped size was already measured
                ld = padded(n)
                J = [fill(torch.empty((Pn, ld), dtype=torch.float64, device=device)) for _ in range(Bn)]
                h0 = [fill(torch.empty((1, ld), dtype=torch.float64, device=device)) for _ in range(Bn)]
                y = [fill(torch.empty((1, ld), dtype=torch.float64, device=device)) for _ in range(Bn)]
                w = [torch.full((1, ld), 400.0, dtype=torch.float64, device=device) for _ in range(Bn)]
                xp = fill(torch.empty((Pn, ld), dtype=torch.float64, device=device))
                xf = fill(torch.empty((Pn, ld), dtype=torch.float64, device=device))
                pinv = torch.zeros((Lp, ld), dtype=torch.float64, device=device)
                r_, c_ = np.tril_indices(Pn)
                for k in range(Lp):
                    if r_[k] == c_[k]:
                        pinv[k] = 4.0
                arr = (_lib.SolveBand * Bn)()
                for b in range(Bn):
                    arr[b].kind, arr[b].n_dim = _lib.BAND_ARRAYS, Pn
                    for k in range(Pn):
                        arr[b].mapper[k] = k
                    arr[b].jac, arr[b].h0, arr[b].y, arr[b].w = J[b].data_ptr(), h0[b].data_ptr(), y[b].data_ptr(), w[b].data_ptr()
                xo, ao = torch.empty_like(xp), torch.empty_like(pinv)
                scal = torch.zeros(2, dtype=torch.float64, device=device)
                n2, nb = scal[0:1], scal[1:2].view(torch.int32)[0:1]
                scratch = torch.empty(int(lib.kafka_solve_scratch_doubles(n)), dtype=torch.float64, device=device)

                def call():
                    _lib.check(lib.kafka_solve_bands(n, ld, Pn, arr, Bn, _lib.ptr(xp), _lib.ptr(xf), _lib.ptr(pinv),
                                                     _lib.ptr(xo), _lib.ptr(ao), None, None, _lib.ptr(n2), _lib.ptr(nb),
                                                     _lib.ptr(scratch), c5_flags, _lib.stream_ptr()), "kafka_solve_bands")
                    if world > 1:                               # the global convergence norm of do_all_bands
                        tot[0] = n2[0]
                        tot[1] = float(n * Pn)
                        if args.comm == "native":
                            comm.all_reduce_sums(tot)
                        else:
                            dist.all_reduce(tot, op=dist.ReduceOp.SUM)
                reps = 5 if n >= 100000 else 20
                for _ in range(2):
                    call()
                barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(reps):
                    call()
                e1.record()
                barrier()
                ms = max_over_ranks(e0.elapsed_time(e1) / reps)
                if n_req == sizes[0] and rank == 0:
                    # parity inside the sweep: the first 512 pixels against oracle B (fp64, dense blocks)
                    from oracle import kafka_cpu as kc
                    ns = min(512, n)
                    Pm = np.zeros((ns, Pn, Pn))
                    Pm[:, np.arange(Pn), np.arange(Pn)] = 4.0
                    hx = lambda t: t[:, :ns].t().contiguous().cpu().numpy()
                    xo_, Ao_, _, _ = kc.solve_dense64([hx(v)[:, 0] for v in y], [hx(v)[:, 0] for v in w],
                                                      [hx(v)[:, 0] for v in h0], [hx(v) for v in J],
                                                      hx(xp).ravel(), hx(xf).ravel(), Pm)
                    got = hx(xo).ravel()
                    err = float(np.max(np.abs(got - xo_) / np.maximum(np.abs(xo_), 1e-6 * np.abs(xo_).max())))
                    assert err < 1e-6, (Pn, Bn, err)
                    parity_worst = max(parity_worst, err)
                rows.append({"pixels": n, "pixels_requested": n_req, "n_params": Pn, "n_bands": Bn, "ms": ms,
                             "Mpx_per_s": world * n / ms / 1e3, "GBs_per_gpu": n * bytes_pp / (ms * 1e-3) / 1e9,
                             "frac_of_hbm_peak": n * bytes_pp / (ms * 1e-3) / 1e9 / hbm})
                del J, h0, y, w, xp, xf, pinv, xo, ao, scratch
        torch.cuda.empty_cache()
    clocks = sampler.stop() if sampler else None
    if rank != 0:
        return None
    line = {"workload": "c5: batched update with supplied Jacobians, per-pixel prior blocks: n_params %s x bands %s x "
                        "pixels %s per GPU (capped at what fits %d GB)" % (params, list(bands), sizes, budget / 1e9),
            "metric": "pixels/sec", "n_gpus": world, "scaling": "weak", "hbm_peak": hbm, "peak_source": peak_src,
            "algorithmic_bytes_per_pixel": "16P + 8P(P+1) + 8P + B*8*(3+P)", "rows": rows, "clocks": clocks,
            "parity": {"checked": "first 512 pixels of every (n_params, bands) at the smallest size vs oracle B",
                       "max_rel_err_x": parity_worst}}
    if world == 1 and cpu_arm and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        cpu_params = [int(v) for v in os.environ.get("KAFKA_C5_CPU_PARAMS", "2,7,10,16").split(",")]
        line["cpu_arm"] = {"what": "unmodified reference variational_kalman_multiband (scipy sparse build + float32 "
                                   "SuperLU, kafka/inference/solvers.py:100-145), supplied Jacobians, chunks of 128x128 px; "
                                   "rates per pixel, EXTRAPOLATED linearly to the sweep's pixel counts (SuperLU runs "
                                   "out of memory at 1e6 px in one piece, SURVEY section 6)",
                           "kind": "reference" if _reference_installed() else "unavailable",
                           "rows": _c5_cpu_arm(cpu_params, bands, min(cores, 64)) if _reference_installed() else []}
    return line


def run_other_workload(args):
    import torch
    if args.workload == "c2":
        torch.cuda.set_device(0)
        emit(json.dumps(measure_c2(torch.device("cuda", 0), args.steps)))
        return
    line = run_c3(args) if args.workload == "c3" else run_c5(args)
    if line is not None:
        emit(json.dumps(line))
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


class HostObservationsGeneric(object):
    """Host numpy rasters (float64 observations, float64 inverse-variance raster, bool mask) of any number
    of bands, in page-locked or pageable memory."""

    def __init__(self, dates, gps, rho, inv, mask, shape, pinned=True):
        from kafka_inferenceengine_b200 import synthetic as syn
        self.dates = list(dates)
        nb = len(gps)
        self.bands_per_observation = {d: nb for d in self.dates}
        self._bands = [syn.BandData(_host_raster(rho[b], shape, np.float64, pinned),
                                    _host_raster(inv[b], shape, np.float64, pinned),
                                    _host_raster(mask[b] != 0, shape, np.bool_, pinned), {"band": b}, gps[b])
                       for b in range(nb)]

    def get_band_data(self, date, band):
        return self._bands[band]


class DeviceObservationsGeneric(object):
    def __init__(self, dates, gps, rho, inv, mask, shape):
        from kafka_inferenceengine_b200 import synthetic as syn
        self.dates = list(dates)
        nb = len(gps)
        self.bands_per_observation = {d: nb for d in self.dates}
        self._bands = [syn.BandData(rho[b].view(shape), inv[b].view(shape), mask[b].view(shape), {"band": b}, gps[b])
                       for b in range(nb)]

    def get_band_data(self, date, band):
        return self._bands[band]


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    if args.workload != "c4shard":
        return run_other_workload(args)

    import torch
    import torch.distributed as dist
    import kafka_inferenceengine_b200 as kb
    from kafka_inferenceengine_b200 import _lib
    from kafka_inferenceengine_b200 import synthetic as syn
    from kafka_inferenceengine_b200.device import BlockDiag, DeviceVector

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node %d" % args.gpus)
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    comm = None
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's own banner / debug lines go to stderr
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=device)
        comm = _make_comm(args, device)

    shape = (args.rows, args.cols)
    n_pix = args.rows * args.cols
    gps = make_emulators()
    rho, inv, mask = synth_on_device(gps, n_pix, 1234 + rank, device)
    state_mask = np.ones(shape, dtype=bool)
    base = dt.datetime(2017, 7, 3)
    date = base + dt.timedelta(hours=12)
    grid = [base, base + dt.timedelta(days=1)]
    inv_block = np.diag(1.0 / syn.SAIL_SIGMA ** 2).astype(np.float32).astype(np.float64)

    def make_filter(obs, out):
        prior = syn.BroadcastPrior(syn.SAIL_MEAN, inv_block, n_pix, lambda blk, n: BlockDiag.from_block(blk, n, device),
                                   lambda m, n: DeviceVector.broadcast(m, n, device))
        kf = kb.LinearKalman(obs, out, state_mask, kb.create_prosail_observation_operator, syn.S2_PARAMETERS,
                             state_propagation=None, prior=prior, linear=False, device=device, comm=comm)
        kf.set_trajectory_model()
        return kf, prior

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- value: inputs resident in HBM ------------------------
    obs_dev = DeviceObservations([date], gps, rho, inv, mask, shape)
    kf, prior = make_filter(obs_dev, NullOutput())
    x0 = DeviceVector.broadcast(syn.SAIL_MEAN, n_pix, device)
    Pinv0 = BlockDiag.from_block(inv_block, n_pix, device)
    data = [obs_dev.get_band_data(date, b) for b in range(B)]

    def step_resident():
        return kf.do_all_bands(date, data, x0, None, Pinv0)

    for _ in range(args.warmup):
        step_resident()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    kf.profile = {}
    kf.last_iterations = []
    launches0 = _lib.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step_resident()
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = _lib.launch_count - launches0
    clocks = sampler.stop() if sampler else None
    iters = list(kf.last_iterations)
    prof = {k: [a.elapsed_time(b) for a, b in v] for k, v in kf.profile.items()}
    kf.profile = None
    ms_step = ms_total / args.steps
    value = world * n_pix / (ms_step * 1e-3)

    # same step with the shared-first-linearisation shortcut switched off (every
    # pixel evaluated on every iteration), for transparency
    kf.dedup_uniform_state = False
    step_resident()
    barrier()
    e0.record()
    for _ in range(max(1, args.steps // 2)):
        step_resident()
    e1.record()
    barrier()
    ms_nodedup = max_over_ranks(e0.elapsed_time(e1)) / max(1, args.steps // 2)
    kf.dedup_uniform_state = True

    # ---------------- e2e: host rasters through LinearKalman.run -----------
    # ONE run() call over a time grid with one observation date per interval, as a user
    # processes a time series: per date the host rasters go host->device (pageable numpy,
    # staged through pinned memory), the state is reset to the prior (kafka_test_S2.py:177-178),
    # the date is assimilated and the float32 mean / uncertainty rasters are read back.
    e2e = e2e_raw = None
    if not args.no_e2e:
        from kafka_inferenceengine_b200.output import RasterOutput
        n_warm = max(2, args.warmup)
        all_dates = [base + dt.timedelta(days=d, hours=12) for d in range(n_warm + args.steps)]
        grid_warm = [base + dt.timedelta(days=d) for d in range(n_warm + 1)]
        grid_timed = [base + dt.timedelta(days=n_warm + d) for d in range(args.steps + 1)]

        class TimedRasterOutput(RasterOutput):
            """RasterOutput that notes when each date is handed over (host clock)."""
            stamps = None

            def dump_data(self, *a, **k):
                if self.stamps is not None:
                    self.stamps.append(time.perf_counter())
                return RasterOutput.dump_data(self, *a, **k)

        def run_e2e(obs):
            import gc
            gc.collect()                                   # what the previous leg left behind (page-locked rasters: freeing
            torch.cuda.synchronize()                       # them synchronises the device) goes now, not inside the timed run
            out = TimedRasterOutput(syn.S2_PARAMETERS)     # float32 mean + 1/sqrt(diag) rasters, read back to the host
            kf2, prior2 = make_filter(obs, out)
            out.pixels = kf2.pixels
            kf2.set_trajectory_uncertainty(0.0)
            # the prior (one mean vector, one block for all pixels) is a constant of the run and
            # stays on the device; the per-step inputs -- the date's rasters -- come from the host
            x0h, Pinv0h = prior2.process_prior(None)
            kf2.run(grid_warm, x0h, None, Pinv0h, iter_obs_op=True)
            barrier()
            h2d0, d2h0 = kf2._stager.h2d_bytes, out.d2h_bytes
            out.stamps = []
            t0 = time.perf_counter()
            kf2.run(grid_timed, x0h, None, Pinv0h, iter_obs_op=True)
            torch.cuda.synchronize()
            ms = max_over_ranks(1e3 * (time.perf_counter() - t0)) / args.steps
            gaps = np.diff(out.stamps)
            ms_steady = max_over_ranks(1e3 * float(np.median(gaps))) if len(gaps) else None
            out.stamps = None
            barrier()
            return {"value": world * n_pix / (ms * 1e-3), "unit": "pixel-dates/s", "ms_per_step": ms,
                    "h2d_bytes_per_step": (kf2._stager.h2d_bytes - h2d0) // args.steps,
                    "d2h_bytes_per_step": (out.d2h_bytes - d2h0) // args.steps,
                    "steady_state_ms_per_step": ms_steady,
                    "host_sources": args.host_sources}

        pinned = args.host_sources == "pinned"
        src_note = ("page-locked numpy rasters from kb.pinned_raster (what a reader fills when the DMA engine "
                    "is to fetch its rasters in place)" if pinned else
                    "ordinary pageable numpy rasters, staged through the package's pinned buffers by a thread pool")
        # ONE run() call over a time grid with one observation date per interval, as a user
        # processes a time series: per date the host rasters go host->device, the state is reset
        # to the prior (kafka_test_S2.py:177-178), the date is assimilated and the float32 mean /
        # uncertainty rasters are read back.
        obs_host = HostObservations(all_dates, gps, rho, inv, mask, shape, pinned=pinned)
        e2e = run_e2e(obs_host)
        del obs_host
        e2e["steady_state_note"] = ("median interval between successive dates inside the same run(); value/ms_per_step "
                                    "are the WHOLE call / K and also carry the first date's staging, which nothing can "
                                    "hide behind, and the last date's read-back")
        e2e["api"] = ("one LinearKalman.run(time_grid, x_forecast, None, P_forecast_inverse) over %d dates: host numpy "
                      "rasters from observations.get_band_data (float64 obs + float64 inverse variance + bool mask = "
                      "17 B per pixel and band; %s) copied host->device every date, prior-only advance, "
                      "output.dump_data = RasterOutput (float32 mean and 1/sqrt(diag P^-1) rasters read back to the "
                      "host every date, as KafkaOutput.dump_data writes them)" % (args.steps, src_note))
        # same run fed with the RAW int16 reflectance rasters (what the granule holds): the reader's
        # arithmetic (mask = raw > 0, /10000, sigma = 5 %, 1/sigma^2; Sentinel2_Observations.py:166-179)
        # runs on the device fused with the gather, 2 B/px/band cross PCIe instead of 17
        obs_raw = RawObservations(all_dates, gps, rho, mask, shape, pinned=pinned)
        e2e_raw = run_e2e(obs_raw)
        del obs_raw
        e2e_raw["api"] = ("same LinearKalman.run() fed with the raw int16 reflectance rasters the Sentinel-2 granule "
                          "holds + a ScaledReflectance marker: the reader's arithmetic "
                          "(Sentinel2_Observations.py:166-179) runs on the device fused with the compaction gather; "
                          "%s; same float32 output rasters read back every date" % src_note)

    # ---------------- the other BASELINE configs, inside the same line ------
    # configs[2] (MODIS year, every rank one tile) and a reduced configs[4] sweep, so that they are measured by
    # whoever runs the default command, at every N; the full versions are `--workload c3|c5`
    c3_line = c5_line = None
    if args.rows == SHARD_ROWS and not args.no_e2e and not args.no_other_configs:
        del obs_dev, kf, prior, x0, Pinv0, data, rho, inv, mask
        torch.cuda.empty_cache()
        ctx = (world, rank, local, device, comm, barrier, max_over_ranks)
        c3_line = run_c3(args, ctx)
        torch.cuda.empty_cache()
        c5_line = run_c5(args, ctx, params=[2, 7, 10, 13, 16], sizes=[10000, 1000000, 10000000], bands=(1, 10),
                         cpu_arm=False)
        torch.cuda.empty_cache()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---------------- roofline ---------------------------------------------
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    fp64_peak = fp64_mix_peak = None
    try:
        probe = ctypes.CDLL(os.path.join(ROOT, "benchmarks", "libkafka_probe.so"))
        probe.kafka_probe_fp64_tflops.restype = ctypes.c_double
        fp64_peak = float(probe.kafka_probe_fp64_tflops(5))
        probe.kafka_probe_fp64_loop_shape_tflops.restype = ctypes.c_double
        fp64_mix_peak = float(probe.kafka_probe_fp64_loop_shape_tflops(3))
    except (OSError, AttributeError):
        pass
    # DRAM traffic per pixel of each kernel from the committed ncu --set full capture
    # (profiles/r0N_ncu_kernels.json, written by profiles/summarize_ncu.py)
    ncu_bpp = {}
    ncu_file = next((f for f in ("r02_ncu_kernels.json", "r01_ncu_kernels.json")
                     if os.path.exists(os.path.join(ROOT, "profiles", f))), "r01_ncu_kernels.json")
    try:
        for k in json.load(open(os.path.join(ROOT, "profiles", ncu_file)))["kernels"]:
            if "dram_bytes_per_pixel" in k:
                ncu_bpp[k["kernel"].split("<")[0].replace("void ", "")] = k["dram_bytes_per_pixel"]
    except Exception:
        pass
    gp_ms = float(np.mean(prof.get("gp_linearize", [float("nan")])))
    sv_ms = float(np.mean(prof.get("solve_bands", [float("nan")])))
    n_iter_mean = float(np.mean(iters))
    gp_bytes = n_pix * (8 * P + B * (1 + 8 + 8 * P))            # x once, mask, h0, Jacobian out
    gp_flops = n_pix * B * T * (5 * P + 28)
    # solve launch with per-pixel operators (iteration >= 2): in x_prev, x_f (8P each), per band
    # w, y, H0, J[P] (8(3+P)); out x (8P), A packed (4P(P+1)), innovations (8B).  The prior block is
    # ONE block shared by all pixels in this workload (KAFKA_PRIOR_BROADCAST), so it is not counted.
    sv_bytes = n_pix * (3 * 8 * P + B * 8 * (3 + P) + 4 * P * (P + 1) + 8 * B)
    gp_tf = gp_flops / (gp_ms * 1e-3) / 1e12
    # the kernel EXECUTES 2D+8 fp64 instructions (2 flops each) per training point where the textbook
    # form counts 5D+28; "achieved" is in executed flops so that frac stays a fraction of the pipe
    gp_exec_tf = gp_tf * (2.0 * (2 * P + 8)) / (5 * P + 28)
    roofline = {
        "kernel": "gp_linearize_kernel<10,2> (GP emulator mean+Jacobian, all bands, one launch per iteration)",
        "bound": "fp64",
        "achieved": gp_exec_tf, "peak": fp64_peak, "unit": "TFLOP/s",
        "frac": (gp_exec_tf / fp64_peak) if fp64_peak else None,
        "peak_source": "fp64 DFMA probe measured in this run (benchmarks/probe.cu); MEASURED_PEAKS.json has no fp64 figure",
        "flops_counted": "executed: 2 x (2D+8) fp64 instructions per (pixel, band, training point)",
        # the same pipe fed with THIS loop's operand mix (most of its FMAs read three registers, and the register
        # file delivers a third 64-bit operand only every third cycle): benchmarks/probe.cu fp64_loop_shape_probe,
        # latency hidden by 16 warps per SM -- what the loop could reach if its dependent chain cost nothing
        "operand_mix": {"peak": fp64_mix_peak, "frac": (gp_exec_tf / fp64_mix_peak) if fp64_mix_peak else None,
                        "unit": "TFLOP/s", "peak_source": "loop-shaped probe measured in this run"},
        "ms_per_launch": gp_ms,
        "share_of_step": float(np.sum(prof.get("gp_linearize", [0.0]))) / args.steps / ms_step,
        "algorithmic": {"achieved": gp_tf, "frac": (gp_tf / fp64_peak) if fp64_peak else None,
                        "unit": "TFLOP/s by SURVEY 8d's count B*T*(5D+28) per pixel-iteration",
                        "flops_per_launch": gp_flops},
        "traffic": (ncu_bpp["gp_linearize_kernel"] * n_pix) if "gp_linearize_kernel" in ncu_bpp else None,
        "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum per pixel from profiles/%s "
                          "(ncu --set full) x pixels of this launch" % ncu_file,
        "hbm": {"achieved": gp_bytes / (gp_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                "frac": gp_bytes / (gp_ms * 1e-3) / 1e9 / hbm_peak, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": gp_bytes,
                "note": "for the record: this kernel is bound by the fp64 pipe, not by HBM "
                        "(SURVEY 8d: ~195 kflop vs ~1 kB per pixel-iteration)"},
        "solve_kernel": {"kernel": "solve_bands_kernel<10> (normal equations + Cholesky + partial norm; per-pixel "
                                   "operators, i.e. the launches of iteration >= 2) + finish_solve_kernel",
                         "bound": "hbm", "achieved": sv_bytes / (sv_ms * 1e-3) / 1e9, "peak": hbm_peak,
                         "unit": "GB/s", "frac": sv_bytes / (sv_ms * 1e-3) / 1e9 / hbm_peak,
                         "algorithmic_bytes_per_launch": sv_bytes, "ms_per_launch": sv_ms,
                         "traffic": (ncu_bpp["solve_bands_kernel"] * n_pix)
                         if "solve_bands_kernel" in ncu_bpp else None},
        "step_breakdown_ms": {k: float(np.sum(v)) / args.steps for k, v in prof.items()},
        "step": {"bytes_per_pixel_date": BYTES_PER_PIXEL_DATE,
                 "achieved": n_pix * BYTES_PER_PIXEL_DATE / (ms_step * 1e-3) / 1e9, "unit": "GB/s",
                 "frac": n_pix * BYTES_PER_PIXEL_DATE / (ms_step * 1e-3) / 1e9 / hbm_peak},
    }

    cpu = None
    if not args.no_cpu_baseline:
        cpu = cpu_baseline_leg(args.ref_chunk, args.port_sample)

    line = {"metric": "pixel-dates/sec", "value": value, "unit": "pixel-dates/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args), "solver_iterations_per_step": iters,
            "first_iteration": "all pixels start at the prior mean, so the first linearisation is evaluated once "
                               "and shared (reference: run_emulator single-unique-vector branch, utils.py:72-74); "
                               "later iterations evaluate every pixel",
            "value_without_shared_first_linearisation": world * n_pix / (ms_nodedup * 1e-3),
            "collective": None if world == 1 else ("4-double NCCL all-reduce (sum of squares, element count, two pixel counters) per solver iteration through %s"
                                                   % ("the library's communicator (kafka_allreduce_norm)"
                                                      if args.comm == "native" else "torch.distributed")),
            "clocks": clocks, "e2e": e2e, "e2e_raw_ingest": e2e_raw, "gpu_launches": launches, "roofline": roofline,
            "cpu_baseline": cpu}
    if world == 1 and not args.no_e2e and args.rows == SHARD_ROWS:
        # BASELINE configs[1] (the Barrax mask on one GPU) beside the headline workload: a parity-test size
        # at which a B200 is launch-bound, reported for completeness, measured after the timed region
        line["configs1_barrax"] = measure_c2(device, 3)
    if c3_line is not None:
        line["configs2_modis_c3"] = c3_line
        line["configs4_sweep_c5"] = c5_line
    emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


class _QuietStdout(object):
    """Everything written to file descriptor 1 while the bench runs -- by Python or by a C
    library such as NCCL's version banner -- goes to stderr; the JSON line is written to the
    real stdout at the end, so stdout carries exactly one line."""

    def __enter__(self):
        sys.stdout.flush()
        self.real = os.dup(1)
        os.dup2(2, 1)
        self.lines = []
        return self

    def emit(self, text):
        self.lines.append(text)

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self.real, 1)
        os.close(self.real)
        for text in self.lines:
            print(text, flush=True)
        return False


_OUT = None


def emit(text):
    if _OUT is None:
        print(text, flush=True)
    else:
        _OUT.emit(text)


if __name__ == "__main__":
    with _QuietStdout() as _OUT:
        main()
