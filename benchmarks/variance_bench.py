"""Predictive variance (kafka_gp_predict with var): the fp64 tensor-core kernel against the
one-warp-per-point kernel (KAFKA_GP_VARIANCE_NAIVE=1), run on the GPU box:

    python benchmarks/variance_bench.py [out.json]
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import sys, json
sys.path.insert(0, %r)
import numpy as np, torch
from kafka_inferenceengine_b200 import _lib, synthetic as syn
from kafka_inferenceengine_b200.inference.utils import emulator_table
from kafka_inferenceengine_b200.device import padded
dev = torch.device("cuda", 0)
torch.manual_seed(0)
rng = np.random.default_rng(3)
rows = []
for D, T, n in ((10, 250, 400000), (4, 500, 200000)):
    mean, sigma = syn.SAIL_MEAN[:D], syn.SAIL_SIGMA[:D]
    gp = syn.make_emulator(rng, mean, sigma, n_train=T)
    tab = emulator_table(gp, dev)
    ld = padded(n)
    x = torch.from_numpy(mean).to(dev)[:, None] + torch.from_numpy(sigma).to(dev)[:, None] * torch.randn((D, ld), dtype=torch.float64, device=dev)
    mu = torch.empty(ld, dtype=torch.float64, device=dev)
    var = torch.empty(ld, dtype=torch.float64, device=dev)
    lib = _lib.get_lib()
    def call():
        _lib.check(lib.kafka_gp_predict(_lib.ptr(x), ld, n, _lib.ptr(tab.table), T, D, _lib.ptr(tab.inv_q()), _lib.ptr(mu), None,
                                        _lib.ptr(var), _lib.stream_ptr()), "predict")
    def mean_only():
        _lib.check(lib.kafka_gp_predict(_lib.ptr(x), ld, n, _lib.ptr(tab.table), T, D, None, _lib.ptr(mu), None, None,
                                        _lib.stream_ptr()), "predict")
    def timed(fn):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3): fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 3
    ms = timed(call) - timed(mean_only)
    rows.append({"n_dim": D, "n_train": T, "points": n, "variance_ms": ms, "tflops": 2.0 * n * T * T / (ms * 1e-3) / 1e12,
                 "checksum": float(var[:n].sum().item())})
print(json.dumps(rows))
""" % ROOT


def main():
    out = {}
    for name, env in (("tensor_core_tiles", {}), ("one_warp_per_point", {"KAFKA_GP_VARIANCE_NAIVE": "1"})):
        r = subprocess.run([sys.executable, "-c", CHILD], capture_output=True, text=True, env=dict(os.environ, **env))
        if r.returncode:
            raise SystemExit(r.stderr[-2000:])
        out[name] = json.loads(r.stdout.strip().splitlines()[-1])
    text = json.dumps({"what": "kafka_gp_predict variance s_f - k^T Q^-1 k: time of the variance launch alone (call with "
                               "variance minus call without), 2 T^2 flops per point", "kernels": out}, indent=1)
    print(text)
    if len(sys.argv) > 1:
        open(sys.argv[1], "w").write(text + "\n")


if __name__ == "__main__":
    main()
