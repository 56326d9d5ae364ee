"""CPU-only tests of the host side of the drop-in: time-grid iteration, pixel
sharding, the convergence-norm all-reduce (gloo, world size 2) and the loud
failure when no GPU is present."""
import datetime as dt
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import golden_io as gio
import kafka_inferenceengine_b200 as kb
from kafka_inferenceengine_b200 import dist as kdist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_iterate_time_grid_matches_reference_kat():
    # /root/reference/tests/test_utils.py:18-38 via the golden fixture
    g = gio.load("kat_time_grid")
    base = dt.datetime(2007, 7, 1)
    grid = [base + i * dt.timedelta(days=1) for i in range(0, 60, 16)]
    base = dt.datetime(2007, 1, 1)
    dates = [base + i * dt.timedelta(days=1) for i in range(1, 365 + 8, 8)]
    rows = list(kb.iterate_time_grid(grid, dates))
    assert [r[0].toordinal() for r in rows] == list(g["steps"])
    assert [[d.toordinal() for d in r[1]] for r in rows] == g["locs"].tolist()
    assert [r[2] for r in rows] == [bool(f) for f in g["firsts"]]


def test_operator_registry_and_names():
    for name in ("create_prosail_observation_operator", "create_nonlinear_observation_operator",
                 "create_linear_observation_operator", "variational_kalman_multiband", "variational_kalman",
                 "propagate_information_filter_LAI", "propagate_information_filter_SLOW",
                 "propagate_information_filter_approx_SLOW", "propagate_standard_kalman", "no_propagation",
                 "propagate_and_blend_prior", "blend_prior", "tip_prior", "block_diag", "run_emulator",
                 "iterate_time_grid", "LinearKalman"):
        assert hasattr(kb, name), name
    op = kb.create_nonlinear_observation_operator
    assert list(op.mapper(7, 0, 4)) == [0, 1, 6, 2] and list(op.mapper(7, 1, 4)) == [3, 4, 6, 5]
    assert list(kb.create_prosail_observation_operator.mapper(10, 3, 10)) == list(range(10))
    with pytest.raises(ValueError):
        kb.create_prosail_observation_operator.mapper(10, 0, 4)
    mean, cov, inv = kb.tip_prior()
    g = gio.load("kat_propagators")
    assert np.array_equal(mean, g["tip_mean"]) and np.array_equal(inv, g["tip_inv"])


@pytest.mark.skipif(torch.cuda.is_available(), reason="needs a machine without a GPU")
def test_product_path_fails_loudly_without_gpu():
    from kafka_inferenceengine_b200 import synthetic as syn
    mask = np.ones((4, 4), dtype=bool)
    with pytest.raises(kb._lib.KafkaCudaError):
        kb.LinearKalman(None, None, mask, kb.create_prosail_observation_operator, syn.S2_PARAMETERS)
    with pytest.raises(NotImplementedError):
        kb.LinearKalman(None, None, mask, lambda *a: None, syn.S2_PARAMETERS)


def test_shard_rows_partitions_and_balances():
    mask = gio.barrax_mask()
    total = int(mask.sum())
    for world in (1, 2, 3, 4, 8):
        got = 0
        prev = 0
        for r in range(world):
            r0, r1 = kdist.shard_rows(mask, r, world)
            assert r0 == prev and r1 >= r0
            prev = r1
            n = int(mask[r0:r1].sum())
            got += n
            assert abs(n - total / world) <= mask.shape[1] + 1
        assert prev == mask.shape[0] and got == total
    assert kdist.shard_pixels(10, 0, 4) == (0, 3) and kdist.shard_pixels(10, 3, 4) == (9, 10)
    assert kdist.shard_pixels(2, 3, 4) == (2, 2)


_WORKER = r"""
import os, sys
sys.path.insert(0, %(root)r)
import numpy as np, torch, torch.distributed as dist
from kafka_inferenceengine_b200 import dist as kdist
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%(port)d", rank=int(sys.argv[1]), world_size=2)
rank = dist.get_rank()
rng = np.random.default_rng(7)
d = rng.standard_normal(1000)               # the whole image's x - x_prev
i0, i1 = kdist.shard_pixels(1000, rank, 2)
local = torch.tensor(float((d[i0:i1] ** 2).sum()), dtype=torch.float64)
norm, count = kdist.global_norm(local, i1 - i0)
want = np.linalg.norm(d) / 1000.0           # linear_kf.py:293 on the unsharded vector
assert count == 1000.0, count
assert abs(norm - want) <= 1e-15 * want, (norm, want)
# the two pixel counters of a solve (left at the forecast, solved by elimination) ride in the same all-reduce
norm2, count2, counts = kdist.global_norm(local, i1 - i0, n_bad_local=torch.tensor([rank + 1, 10 * rank], dtype=torch.int32))
assert (norm2, count2) == (norm, count) and counts == (3, 10), counts
norm3, _, one = kdist.global_norm(float(local), i1 - i0, n_bad_local=2)
assert norm3 == norm and one == 4, one
g = kdist.gather_interleaved(torch.from_numpy(d[i0:i1].copy()))
assert np.array_equal(g.numpy(), d)
# ranks take the same stop decision
flag = torch.tensor([1.0 if norm < 1e-3 else 0.0])
both = [torch.zeros(1), torch.zeros(1)]
dist.all_gather(both, flag)
assert both[0].item() == both[1].item()
dist.destroy_process_group()
print("OK", rank)
"""


def test_global_norm_gloo_world2(tmp_path):
    port = 29500 + (os.getpid() % 500)
    script = tmp_path / "worker.py"
    script.write_text(_WORKER % {"root": ROOT, "port": port})
    procs = [subprocess.Popen([sys.executable, str(script), str(r)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                              env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
             for r in range(2)]
    outs = [p.communicate(timeout=180)[0].decode() for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
        assert "OK" in o


def test_bench_reference_arm_prints_contract_line():
    """bench.py --impl reference: the CPU arm (the unmodified reference on the host cores)
    prints one JSON line with the contract's keys; here on small chunks.  The processes of
    that arm never import the product package, so no CUDA library is mapped in them."""
    import json
    from oracle import ref_shim
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0", "--ref-chunk", "12"], capture_output=True, text=True, timeout=600,
                         env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "scaling", "config", "cpu_baseline", "e2e"):
        assert key in line, key
    assert line["impl"] == "reference" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == ("reference" if ref_shim.available() else "port")
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["value"] == line["value"]
    assert "note" not in line["config"]                 # same config dict as the GPU arm's line
    # ranks other than 0 print nothing and exit 0 (torchrun, N > 1)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                         capture_output=True, text=True, timeout=120, env=dict(os.environ, RANK="1", WORLD_SIZE="2"))
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_cpu_legs_do_not_import_the_product_package():
    """The functions bench.py runs in the CPU legs' worker processes leave
    kafka_inferenceengine_b200 (and with it libkafka_b200.so) unimported."""
    code = ("import sys; sys.path.insert(0, %r); import bench; "
            "jobs = bench._cpu_workload(1, 6); bench._port_job(jobs[0]); "
            "r = bench._ref_chunk(jobs[0]) if bench._reference_installed() else None; "
            "assert not any(m.startswith('kafka_inferenceengine_b200') for m in sys.modules), 'product imported'; "
            "assert not any('libkafka_b200' in l for l in open('/proc/self/maps')), 'library mapped'; print('OK')" % ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "OK" in out.stdout, out.stderr[-2000:]


def test_oracle_synthetic_restates_the_package_constants():
    """oracle/synthetic.py (used by the CPU legs instead of the product package) and
    kafka_inferenceengine_b200/synthetic.py describe the same workload."""
    from kafka_inferenceengine_b200 import synthetic as syn
    from oracle import synthetic as osyn
    assert np.array_equal(osyn.SAIL_MEAN, syn.SAIL_MEAN) and np.array_equal(osyn.SAIL_SIGMA, syn.SAIL_SIGMA)
    assert osyn.S2_PARAMETERS == syn.S2_PARAMETERS
    a = osyn.make_emulators(3, 40)
    rng = np.random.default_rng(1234)
    b = [syn.make_emulator(rng, syn.SAIL_MEAN, syn.SAIL_SIGMA, n_train=40) for _ in range(3)]
    for ga, gb in zip(a, b):
        assert np.array_equal(ga.inputs, gb.inputs) and np.array_equal(ga.targets, gb.targets)
        assert np.array_equal(ga.theta, gb.theta)


def test_get_chunks_matches_reference_tiling():
    """kafka/input_output/utils.py:12-40: same tuples in the same order as the
    reference's generator (transcribed expectations for 600 x 500 in 256 blocks,
    plus the unmodified function itself where /root/reference is present)."""
    from kafka_inferenceengine_b200.input_output import get_chunks
    got = list(get_chunks(600, 500, block_size=[256, 256]))
    assert got == [(0, 0, 256, 256, 1), (0, 256, 256, 244, 2), (256, 0, 256, 256, 3), (256, 256, 256, 244, 4),
                   (512, 0, 88, 256, 5), (512, 256, 88, 244, 6)]
    assert list(get_chunks(128, 128, block_size=[128, 128])) == [(0, 0, 128, 128, 1)]
    assert list(get_chunks(0, 10)) == []
    covered = np.zeros((500, 600), dtype=int)
    for x0, y0, w, h, _ in get_chunks(600, 500, block_size=[100, 37]):
        covered[y0:y0 + h, x0:x0 + w] += 1
    assert (covered == 1).all()
    from oracle import ref_shim
    if ref_shim.available():
        ref_shim.load()
        from kafka.input_output.utils import get_chunks as ref_chunks
        for nx, ny, bs in ((600, 500, [256, 256]), (235, 204, [128, 128]), (10980, 10980, [4096, 1000]), (5, 3, [7, 7])):
            assert list(get_chunks(nx, ny, bs)) == list(ref_chunks(nx, ny, block_size=bs))


def test_rows_zero_padded_clears_only_the_padding():
    """Band-row buffers are handed to kernels that write every pixel below n_pix; only
    the padding columns up to the aligned leading dimension have to start at zero."""
    import torch
    from kafka_inferenceengine_b200.device import padded, rows_zero_padded
    for n in (0, 1, 31, 32, 1000):
        t = rows_zero_padded(3, n, torch.device("cpu"))
        assert t.shape == (3, padded(n)) and t.dtype == torch.float64
        assert bool((t[:, n:] == 0).all())
    m = rows_zero_padded(2, 5, torch.device("cpu"), torch.uint8)
    assert m.dtype == torch.uint8 and bool((m[:, 5:] == 0).all())
