"""GPU test of the multi-process path: two ranks, each
owning a row band of the image, with the convergence norm all-reduced.  Both
processes share cuda:0 here (gloo carries the 2-element reduction; on the GPU box
bench.py uses NCCL, one GPU per rank); the ranks never wait on each other inside
a kernel.  The sharded result must equal the single-process run on the whole
image bit for bit, with the same iteration count."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_WORKER = r"""
import os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np, torch, torch.distributed as dist
import golden_io as gio
import kafka_inferenceengine_b200 as kb
from kafka_inferenceengine_b200 import synthetic as syn, dist as kdist
from kafka_inferenceengine_b200.device import BlockDiag
rank = int(sys.argv[1])
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%(port)d", rank=rank, world_size=2)
torch.cuda.set_device(0)
g = gio.load("run_s2")
mask = g["mask"]; P = 10
gps = gio.emulators(g, syn.EmulatorData)
obs = gio.FixtureObservations(g, gps)
date = obs.dates[0]
inv = g["prior_inv_block"].astype(np.float64)

def run(rows, comm):
    sub = mask[rows]
    n = int(sub.sum())
    data = [obs.get_band_data(date, b) for b in range(10)]
    data = [d._replace(observations=d.observations[rows], mask=d.mask[rows],
                       uncertainty=d.uncertainty.diagonal().reshape(mask.shape)[rows]) for d in data]
    kf = kb.LinearKalman(None, None, sub, kb.create_prosail_observation_operator, syn.S2_PARAMETERS, comm=comm)
    x, _, A, _ = kf.do_all_bands(date, data, np.tile(g["x0"][:P], n), None, BlockDiag.from_block(inv, n))
    return x.interleaved(), A.diagonal_device(), kf.last_iterations[-1], kf.last_norms[-1]

r0, r1 = kdist.shard_rows(mask, rank, 2)
xs, ds, it_s, norms_s = run(slice(r0, r1), dist.group.WORLD)
x_all = kdist.gather_interleaved(xs.cpu()); d_all = kdist.gather_interleaved(ds.cpu())
if rank == 0:
    xf, df, it_f, norms_f = run(slice(0, mask.shape[0]), None)
    assert it_s == it_f, (it_s, it_f)
    assert np.allclose(norms_s, norms_f, rtol=1e-12), (norms_s, norms_f)
    assert torch.equal(x_all, xf.cpu()) and torch.equal(d_all, df.cpu())
    assert 0 < r1 < mask.shape[0]
dist.barrier()
dist.destroy_process_group()
print("OK", rank, it_s)
"""


def test_two_ranks_sharded_by_rows_match_single_process(tmp_path):
    port = 29600 + (os.getpid() % 300)
    script = tmp_path / "worker.py"
    script.write_text(_WORKER % {"root": ROOT, "port": port})
    procs = [subprocess.Popen([sys.executable, str(script), str(r)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
             for r in range(2)]
    outs = [p.communicate(timeout=600)[0].decode() for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-3000:]
        assert "OK" in o
