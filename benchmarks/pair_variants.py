"""Sweep compile-time variants of the two-lanes-per-pixel solve kernel (csrc/solve_pair.cuh knobs).

    python benchmarks/pair_variants.py build     # here (nvcc, no GPU needed)
    python benchmarks/pair_variants.py run       # on the GPU box

Each variant is a full libkafka_b200.so whose solve.o was compiled with -D flags (the other objects are
the library's own, csrc/build/*.o); the run step puts each one behind `bench.py --workload c5` through
KAFKA_B200_LIB and prints the fraction of the HBM peak for n_params 14-16 (the bench checks parity
against the oracle on every row)."""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(ROOT, "kafka_inferenceengine_b200", "csrc")
OUT = os.path.join(HERE, "variants")
# Knobs that are still in the source.  Round 2 also measured -- and removed, see
# profiles/r02_pair_persistent_variants.txt -- a persistent form of the kernel (CTAs looping over tiles, the
# next tile's prior block prefetched into shared memory, per-warp sums, staggered starts): slower in every form.
VARIANTS = {
    "default": [],
    "split4": ["-DKAFKA_PAIR_SPLIT=4"],                 # 32-pixel CTAs, four per SM
    "stages2": ["-DKAFKA_ASYNC_STAGES=2"],
    "stages4": ["-DKAFKA_ASYNC_STAGES=4"],
    "pair_from_13": ["-DKAFKA_PAIR_FROM_PARAMS=13"],
}
if os.environ.get("KAFKA_PAIR_VARIANTS"):
    VARIANTS = {k: v for k, v in VARIANTS.items() if k in os.environ["KAFKA_PAIR_VARIANTS"].split(",")}
OTHERS = ["layout", "gp", "advance", "assimilate", "noalloc", "comm"]


def build():
    os.makedirs(OUT, exist_ok=True)
    procs = []
    for name, flags in VARIANTS.items():
        obj = os.path.join(OUT, "solve_%s.o" % name)
        cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler",
               "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v", "-c", os.path.join(CSRC, "solve.cu"), "-o", obj] + flags
        procs.append((name, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for name, obj, p in procs:
        out = p.communicate()[0].decode()
        lines = out.splitlines()
        info = [lines[k + 2].strip() + " | " + lines[k + 1].strip() for k, l in enumerate(lines)
                if "pair_kernelILi16ELb1" in l and "Compiling" in l and k + 2 < len(lines)]
        if p.returncode == 0:
            link = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o",
                    os.path.join(OUT, "pair_%s.so" % name), obj] + \
                   [os.path.join(CSRC, "build", o + ".o") for o in OTHERS] + ["-lcudart", "-ldl"]
            r = subprocess.run(link, capture_output=True, text=True)
            print(name, "ok" if r.returncode == 0 else "LINK FAILED\n" + r.stderr)
        else:
            print(name, "FAILED\n" + out[-3000:])
        for i in info:
            print("    ", i)


def run():
    params = os.environ.get("KAFKA_C5_PARAMS", "14,15,16")
    pixels = os.environ.get("KAFKA_C5_PIXELS", "4e6")
    libs = {"library in csrc/": None}
    libs.update({name: os.path.join(OUT, "pair_%s.so" % name) for name in VARIANTS})
    table = {}
    for name, lib in libs.items():
        if lib is not None and not os.path.exists(lib):
            continue
        env = dict(os.environ, KAFKA_C5_PARAMS=params, KAFKA_C5_PIXELS=pixels)
        if lib:
            env["KAFKA_B200_LIB"] = lib
        r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--workload", "c5", "--no-cpu-baseline"],
                           capture_output=True, text=True, env=env, cwd=ROOT)
        if r.returncode != 0:
            print(name, "FAILED", r.stderr[-1500:])
            continue
        rows = json.loads(r.stdout.strip().splitlines()[-1])["rows"]
        table[name] = {"%d params, %d bands" % (x["n_params"], x["n_bands"]): round(x["frac_of_hbm_peak"], 3) for x in rows}
        print("%-22s" % name, " ".join("%s=%.3f" % (k.replace(" params, ", "/").replace(" bands", ""), v)
                                       for k, v in table[name].items()), flush=True)
    print(json.dumps({"pixels": pixels, "fraction_of_hbm_peak": table}))


if __name__ == "__main__":
    {"build": build, "run": run}[sys.argv[1]]()
