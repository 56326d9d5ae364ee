"""Where does the host thread of the end-to-end run spend its time?

    python benchmarks/e2e_timeline.py [raw|host] [pinned|pageable] [rows]
    python -m torch.distributed.run --nproc-per-node N ... benchmarks/e2e_timeline.py raw pinned

Wraps the phases of LinearKalman.run (read bands, stage, finish, norm wait, output, advance)
with wall-clock timers -- no extra synchronisation, so the overlap is the one bench.py sees --
and prints milliseconds per date.  A phase that waits for the GPU (norm) shows the GPU time
the host could not hide behind.  Under torchrun every rank runs its own shard with the global
convergence norm (as bench.py does); rank 0 prints the per-phase mean and maximum over ranks
and, with KAFKA_TIMELINE_OUT=<file>, writes the table as JSON."""
import collections
import datetime as dt
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import kafka_inferenceengine_b200 as kb  # noqa: E402
from kafka_inferenceengine_b200 import engine, synthetic as syn  # noqa: E402
from kafka_inferenceengine_b200.device import BlockDiag, DeviceVector  # noqa: E402
from kafka_inferenceengine_b200.output import RasterOutput  # noqa: E402

ACC = collections.defaultdict(float)


LOG = []            # (tag, start, end) of every timed call, seconds


def timed(cls, name, tag):
    fn = getattr(cls, name)

    def wrap(*a, **k):
        t0 = time.perf_counter()
        try:
            return fn(*a, **k)
        finally:
            t1 = time.perf_counter()
            ACC[tag] += t1 - t0
            LOG.append((tag, t0, t1))
    setattr(cls, name, wrap)


def main():
    import torch.distributed as dist
    mode = sys.argv[1] if len(sys.argv) > 1 else "raw"
    pinned = (sys.argv[2] if len(sys.argv) > 2 else "pinned") == "pinned"
    rows = int(sys.argv[3]) if len(sys.argv) > 3 else bench.SHARD_ROWS
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    comm = None
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
        from kafka_inferenceengine_b200.dist import NativeComm
        comm = NativeComm.from_process_group(dist.group.WORLD, device)
    shape = (rows, bench.TILE)
    n_pix = rows * bench.TILE
    gps = bench.make_emulators()
    rho, inv, mask = bench.synth_on_device(gps, n_pix, 1234 + rank, device)
    base = dt.datetime(2017, 7, 3)
    n_warm, steps = 2, int(os.environ.get("KAFKA_TIMELINE_STEPS", "6"))
    dates = [base + dt.timedelta(days=d, hours=12) for d in range(n_warm + steps)]
    obs = (bench.RawObservations(dates, gps, rho, mask, shape, pinned=pinned) if mode == "raw"
           else bench.HostObservations(dates, gps, rho, inv, mask, shape, pinned=pinned))
    del rho, inv, mask
    out = RasterOutput(syn.S2_PARAMETERS)
    inv_block = np.diag(1.0 / syn.SAIL_SIGMA ** 2).astype(np.float32).astype(np.float64)
    prior = syn.BroadcastPrior(syn.SAIL_MEAN, inv_block, n_pix, lambda blk, n: BlockDiag.from_block(blk, n, device),
                               lambda m, n: DeviceVector.broadcast(m, n, device))
    kf = kb.LinearKalman(obs, out, np.ones(shape, dtype=bool), kb.create_prosail_observation_operator,
                         syn.S2_PARAMETERS, state_propagation=None, prior=prior, linear=False, device=device, comm=comm)
    kf.set_trajectory_model()
    kf.set_trajectory_uncertainty(0.0)
    out.pixels = kf.pixels
    x0, Pinv0 = prior.process_prior(None)
    kf.run([base + dt.timedelta(days=d) for d in range(n_warm + 1)], x0, None, Pinv0, iter_obs_op=True)
    torch.cuda.synchronize()

    # device-side timeline: when do the H2D copies of the NEXT date start/end relative to the
    # start of the emulator pass they are meant to hide behind?
    marks = []
    orig_stage = engine.ObservationStager.stage

    def stage_marked(self, data):
        e_gp = torch.cuda.Event(enable_timing=True)
        e_gp.record(torch.cuda.current_stream())          # main stream: emulator + solve of iteration 2 are queued
        tok = orig_stage(self, data)
        e_end = torch.cuda.Event(enable_timing=True)
        if self._copy_stream is not None:
            e_end.record(self._copy_stream)
        marks.append((e_gp, e_end))
        return tok
    engine.ObservationStager.stage = stage_marked
    timed(kb.LinearKalman, "_read_bands", "read_bands")
    timed(engine.ObservationStager, "stage", "stage (host copies + queue H2D)")
    timed(engine.ObservationStager, "finish", "finish (queue gathers)")
    timed(kb.LinearKalman, "_global_norm", "norm (waits for the GPU / all ranks)")
    timed(kb.LinearKalman, "advance", "advance")
    timed(RasterOutput, "dump_data", "dump_data (incl. wait for previous D2H)")
    timed(RasterOutput, "_deliver", "  of which _deliver")
    timed(kb.LinearKalman, "do_all_bands", "do_all_bands (total)")
    kf.profile = {}
    prof = None
    if os.environ.get("KAFKA_CPROFILE"):
        import cProfile
        prof = cProfile.Profile()
        prof.enable()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    kf.run([base + dt.timedelta(days=n_warm + d) for d in range(steps + 1)], x0, None, Pinv0, iter_obs_op=True)
    torch.cuda.synchronize()
    total = time.perf_counter() - t0
    if prof is not None:
        import pstats
        prof.disable()
        pstats.Stats(prof).sort_stats("tottime").print_stats(18)
    table = {"total ms per date": 1e3 * total / steps}
    for k, v in ACC.items():
        table[k] = 1e3 * v / steps
    h2d = []
    for e_gp, e_end in marks:
        try:
            h2d.append(e_gp.elapsed_time(e_end))
        except Exception:
            pass
    if h2d:
        table["H2D of the next date done .. ms after the current date's kernels (negative: before)"] = float(np.mean(h2d))
    for k, v in kf.profile.items():
        table["GPU " + k + " (CUDA events)"] = sum(a.elapsed_time(b) for a, b in v) / steps
    tables = [table]
    if world > 1:
        tables = [None] * world
        dist.all_gather_object(tables, table)
    if rank == 0:
        head = "mode %s, %s sources, %d ranks x %d px, %d staging threads per rank, %d dates" % (
            mode, "pinned" if pinned else "pageable", world, n_pix, engine.ObservationStager.N_THREADS, steps)
        print(head)
        rows_ = []
        for k in table:
            vals = [t.get(k, float("nan")) for t in tables]
            rows_.append({"phase": k, "mean_ms": float(np.mean(vals)), "max_ms": float(np.max(vals)),
                          "min_ms": float(np.min(vals))})
            print("  %-84s mean %8.2f  min %8.2f  max %8.2f ms/date" % (k, np.mean(vals), np.min(vals), np.max(vals)))
        print("  iterations", kf.last_iterations[-steps:])
        path = os.environ.get("KAFKA_TIMELINE_OUT")
        if path:
            with open(path, "a") as f:
                f.write(json.dumps({"what": head, "mode": mode, "pinned_sources": pinned, "n_ranks": world,
                                    "pixels_per_rank": n_pix, "staging_threads": engine.ObservationStager.N_THREADS,
                                    "cpu_count": os.cpu_count(), "dates": steps, "phases": rows_,
                                    "pixel_dates_per_s": world * n_pix / (max(t["total ms per date"] for t in tables) * 1e-3)})
                        + "\n")
    if os.environ.get("KAFKA_TIMELINE_LOG") and rank == 0:
        z = LOG[0][1]
        for tag, a, b in LOG:
            print("    %8.2f -> %8.2f ms  %s" % (1e3 * (a - z), 1e3 * (b - z), tag))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
