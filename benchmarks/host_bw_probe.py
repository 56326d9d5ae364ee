"""What the host side of one box can move: pinned H2D / D2H rates and pageable->pinned
memcpy rates, measured by ALL ranks at the same time (the end-to-end bench's ranks share
one host's memory system).

    python benchmarks/host_bw_probe.py
    python -m torch.distributed.run --nproc-per-node N ... benchmarks/host_bw_probe.py

Every figure is GB/s per rank; rank 0 prints the per-rank mean and the sum over ranks and,
with KAFKA_PROBE_OUT=<file>, appends one JSON line."""
import concurrent.futures
import json
import os
import time

import numpy as np
import torch


def main():
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    n = 1 << 30
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    d2 = torch.empty(n, dtype=torch.uint8, device=dev)
    res = {}

    def run(name, fn, reps=5, nbytes=n):
        fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        res[name] = reps * nbytes / (time.perf_counter() - t0) / 1e9
        barrier()

    run("h2d_pinned", lambda: d.copy_(h, non_blocking=True))
    run("d2h_pinned", lambda: h2.copy_(d2, non_blocking=True))
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def both():
        with torch.cuda.stream(s1):
            d.copy_(h, non_blocking=True)
        with torch.cuda.stream(s2):
            h2.copy_(d2, non_blocking=True)
    run("h2d_and_d2h_each", both)

    src = np.empty(n // 8)
    src[:] = 1.0                                          # touched, pageable
    dst = h.numpy().view(np.float64)
    for threads in (1, 2, 4, 8):
        pool = concurrent.futures.ThreadPoolExecutor(max_workers=threads)
        m = src.size

        def cp():
            fs = [pool.submit(np.copyto, dst[m * k // threads:m * (k + 1) // threads],
                              src[m * k // threads:m * (k + 1) // threads]) for k in range(threads)]
            for f in fs:
                f.result()
        run("memcpy_pageable_to_pinned_%dthr" % threads, cp, reps=3)
        pool.shutdown()

    # memcpy while the DMA engines run both ways (what the staged path does at steady state)
    pool = concurrent.futures.ThreadPoolExecutor(max_workers=4)
    dst2 = torch.empty(n, dtype=torch.uint8).pin_memory().numpy().view(np.float64)

    def cp_and_dma():
        both()
        m = src.size
        fs = [pool.submit(np.copyto, dst2[m * k // 4:m * (k + 1) // 4], src[m * k // 4:m * (k + 1) // 4]) for k in range(4)]
        for f in fs:
            f.result()
    run("memcpy_4thr_with_dma_both_ways", cp_and_dma, reps=3)

    allres = [res]
    if world > 1:
        allres = [None] * world
        dist.all_gather_object(allres, res)
    if rank == 0:
        line = {"n_ranks": world, "cpu_count": os.cpu_count(), "unit": "GB/s per rank (1 GiB transfers)",
                "per_rank_mean": {k: float(np.mean([r[k] for r in allres])) for k in res},
                "sum_over_ranks": {k: float(np.sum([r[k] for r in allres])) for k in res}}
        print(json.dumps(line, indent=1))
        path = os.environ.get("KAFKA_PROBE_OUT")
        if path:
            with open(path, "a") as f:
                f.write(json.dumps(line) + "\n")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
