"""The multi-GPU step behind the C ABI (kafka_ctx_create / kafka_allreduce_norm /
kafka_gather_outputs, include/kafka_b200.h): through ctypes, without torch.distributed.

* one rank: runs on any GPU box (all-reduce and gather over a communicator of size 1);
* two ranks, one GPU each (skipped with fewer than two GPUs): the unique id travels
  through a file; sums add up, ragged gathers concatenate in rank order, and the
  row-sharded filter run -- its convergence all-reduce taken through the library's own
  communicator -- equals the single-process run bit for bit."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_single_rank_context_through_the_c_abi():
    import ctypes as C
    from kafka_inferenceengine_b200 import _lib
    lib = _lib.get_lib()
    ident = C.create_string_buffer(_lib.COMM_ID_BYTES)
    assert lib.kafka_comm_unique_id(ident) == 0, lib.kafka_last_error()
    ctx = C.c_void_p()
    assert lib.kafka_ctx_create(0, 1, ident, C.byref(ctx)) == 0, lib.kafka_last_error()
    rank, size, ver = C.c_int(-1), C.c_int(-1), C.c_int(0)
    assert lib.kafka_ctx_info(ctx, C.byref(rank), C.byref(size), C.byref(ver)) == 0
    assert (rank.value, size.value) == (0, 1) and ver.value >= 20000
    sums = torch.tensor([3.5, 7.0, 2.0], dtype=torch.float64, device="cuda:0")
    assert lib.kafka_allreduce_norm(ctx, _lib.ptr(sums), 3, _lib.stream_ptr()) == 0, lib.kafka_last_error()
    assert sums.cpu().tolist() == [3.5, 7.0, 2.0]
    send = torch.arange(1000, dtype=torch.float64, device="cuda:0")
    recv = torch.zeros(1000, dtype=torch.float64, device="cuda:0")
    counts = (C.c_int64 * 1)(1000)
    assert lib.kafka_gather_outputs(ctx, _lib.ptr(send), counts, _lib.ptr(recv), _lib.stream_ptr()) == 0
    torch.cuda.synchronize()
    assert torch.equal(send, recv)
    # bad arguments are refused with a message, not passed to NCCL
    assert lib.kafka_allreduce_norm(ctx, _lib.ptr(sums), 0, None) == -1
    assert lib.kafka_ctx_create(2, 2, ident, C.byref(C.c_void_p())) == -1
    assert lib.kafka_ctx_destroy(ctx) == 0


_WORKER = r"""
import os, sys, time
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np, torch
import golden_io as gio
import kafka_inferenceengine_b200 as kb
from kafka_inferenceengine_b200 import synthetic as syn, dist as kdist
from kafka_inferenceengine_b200.device import BlockDiag
rank = int(sys.argv[1]); idfile = sys.argv[2]
torch.cuda.set_device(rank)
if rank == 0:
    ident = kdist.NativeComm.unique_id()
    with open(idfile + ".tmp", "wb") as f:
        f.write(ident)
    os.rename(idfile + ".tmp", idfile)
else:
    for _ in range(600):
        if os.path.exists(idfile):
            break
        time.sleep(0.1)
    ident = open(idfile, "rb").read()
comm = kdist.NativeComm(rank, 2, ident)
s = torch.tensor([1.0 + rank, 10.0 * (rank + 1), 0.25], dtype=torch.float64, device="cuda")
comm.all_reduce_sums(s)
assert s.cpu().tolist() == [3.0, 30.0, 0.5], s
mine = torch.arange(5 + 3 * rank, dtype=torch.float64, device="cuda") + 100 * rank
allv = comm.gather(mine, [5, 8])
want = torch.cat([torch.arange(5, dtype=torch.float64), torch.arange(8, dtype=torch.float64) + 100])
assert torch.equal(allv.cpu(), want)

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
xs, ds, it_s, norms_s = run(slice(r0, r1), comm)
x_all = kdist.gather_interleaved(xs, comm); d_all = kdist.gather_interleaved(ds, comm)
if rank == 0:
    xf, df, it_f, norms_f = run(slice(0, mask.shape[0]), None)
    assert it_s == it_f, (it_s, it_f)
    assert np.allclose(norms_s, norms_f, rtol=1e-12), (norms_s, norms_f)
    assert torch.equal(x_all.cpu(), xf.cpu()) and torch.equal(d_all.cpu(), df.cpu())
torch.cuda.synchronize()
comm.close()
print("OK", rank, it_s)
"""


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (one rank per GPU)")
def test_two_ranks_through_the_library_communicator(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER % {"root": ROOT})
    idfile = str(tmp_path / "nccl_id")
    procs = [subprocess.Popen([sys.executable, str(script), str(r), idfile], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT) for r in range(2)]
    outs = [p.communicate(timeout=600)[0].decode() for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-3000:]
        assert "OK" in o
