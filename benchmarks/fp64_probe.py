"""fp64 figures of this GPU that the emulator kernel's design rests on (run on the GPU box):

    python benchmarks/fp64_probe.py [out.json]

* DFMA and DMMA (mma.sync m8n8k4) peak, TFLOP/s -- the roofline denominator of gp_linearize_kernel
  (MEASURED_PEAKS.json holds no fp64 figure) and the evidence that the fp64 tensor path offers nothing
  over the vector path on this part;
* cycles between successive dependent DFMA steps of a warp, for 1..8 independent chains per thread and
  1..4 warps per scheduler: with one chain and one warp this is the DFMA latency, and it tells how many
  independent chains a scheduler needs in flight before the pipe is the limit (2 cycles per warp-FMA).
"""
import ctypes
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    lib = ctypes.CDLL(os.path.join(ROOT, "benchmarks", "libkafka_probe.so"))
    for name in ("kafka_probe_fp64_tflops", "kafka_probe_dmma_tflops", "kafka_probe_fp64_fma3_tflops", "kafka_probe_fp64_loop_shape_tflops"):
        getattr(lib, name).restype = ctypes.c_double
    lib.kafka_probe_dfma_step_cycles.restype = ctypes.c_double
    lib.kafka_probe_dfma_step_cycles.argtypes = [ctypes.c_int, ctypes.c_int]
    q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_power_cap"
    smi = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "100"],
                           stdout=subprocess.PIPE, text=True)
    res = {"dfma_tflops": lib.kafka_probe_fp64_tflops(5), "dmma_m8n8k4_tflops": lib.kafka_probe_dmma_tflops(5)}
    res["dfma_three_register_operands_tflops"] = lib.kafka_probe_fp64_fma3_tflops(5)
    res["emulator_loop_operand_mix_tflops"] = lib.kafka_probe_fp64_loop_shape_tflops(5)
    res["dfma_tflops_again"] = lib.kafka_probe_fp64_tflops(5)
    steps = {}
    for warps in (1, 2, 3, 4):
        for ch in (1, 2, 3, 4, 6, 8):
            steps["chains%d_warps%d" % (ch, warps)] = lib.kafka_probe_dfma_step_cycles(ch, warps)
    res["dfma_cycles_per_dependent_step"] = steps
    smi.terminate()
    rows = [r.split(",") for r in smi.stdout.read().strip().splitlines() if r.count(",") >= 3]
    sm = sorted(float(r[0]) for r in rows) if rows else []
    res["clocks"] = {"sm_mhz_median": sm[len(sm) // 2] if sm else None, "sm_mhz_max_seen": sm[-1] if sm else None,
                     "sm_max_mhz": float(rows[0][1]) if rows else None, "samples": len(sm),
                     "hw_slowdown": any("Active" in r[2] and "Not" not in r[2] for r in rows),
                     "sw_power_cap": any("Active" in r[3] and "Not" not in r[3] for r in rows)}
    text = json.dumps(res, indent=1)
    print(text)
    if len(sys.argv) > 1:
        open(sys.argv[1], "w").write(text + "\n")


if __name__ == "__main__":
    main()
