"""Build libkafka_b200.so in-tree with nvcc for sm_100a (B200 only).

    python -m kafka_inferenceengine_b200.build [--force] [--verbose]

The library is git-ignored but travels to the GPU box with the working tree.
"""
import concurrent.futures
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(CSRC, "libkafka_b200.so")
SOURCES = ["layout.cu", "gp.cu", "solve.cu", "advance.cu", "assimilate.cu", "noalloc.cu", "comm.cu"]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def _sources():
    return [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def _headers():
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".inc"))]
    deps.append(os.path.join(os.path.dirname(HERE), "include", "kafka_b200.h"))
    return deps


def _obj(src):
    return os.path.join(CSRC, "build", src.replace(".cu", ".o"))


def _object_stale(src):
    """An object is rebuilt when its own source or any shared header is newer than it."""
    obj = _obj(src)
    if not os.path.exists(obj):
        return True
    t = os.path.getmtime(obj)
    return any(os.path.getmtime(d) > t for d in [os.path.join(CSRC, src)] + _headers())


def _compile(src, verbose):
    obj = _obj(src)
    cmd = [_nvcc()] + ARCH + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    return obj, r.stderr


def build_library(force=False, verbose=False):
    os.makedirs(os.path.join(CSRC, "build"), exist_ok=True)
    todo = [s for s in _sources() if force or _object_stale(s)]
    if not todo and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(_obj(s)) for s in _sources()):
        return LIB
    with concurrent.futures.ThreadPoolExecutor(max_workers=max(1, len(todo))) as ex:
        results = list(ex.map(lambda s: _compile(s, verbose), todo))
    if verbose:
        for _, log in results:
            sys.stderr.write(log)
    # NCCL is opened at run time by comm.cu (dlopen), not linked
    cmd = [_nvcc()] + ARCH + ["-shared", "-o", LIB] + [_obj(s) for s in _sources()] + ["-lcudart", "-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
