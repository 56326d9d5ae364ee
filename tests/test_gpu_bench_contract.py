"""bench.py prints ONE JSON line with every key of the measurement contract
(small shard, so that the check takes seconds)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*extra):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--gpus", "1", "--steps", "2", "--warmup", "3",
                          "--rows", "40", "--ref-chunk", "12", "--port-sample", "2048"] + list(extra),
                         capture_output=True, text=True, cwd=ROOT, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    return json.loads(lines[0])


def test_bench_line_has_the_contract_keys():
    d = _run()
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "e2e_raw_ingest", "gpu_launches", "roofline",
                "cpu_baseline"):
        assert key in d, key
    assert d["metric"] == "pixel-dates/sec" and d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 3
    assert d["dtype"] == "f64" and d["scaling"] == "weak" and d["higher_is_better"] is True
    assert "workload" in d["config"] and d["gpu_launches"] > 0 and d["value"] > 0
    e = d["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] == 40 * 10980 * 10 * 17 and e["d2h_bytes_per_step"] == 40 * 10980 * 10 * 8
    assert e["value"] < d["value"] * 1.05                     # never faster than the resident number (noise margin)
    raw = d["e2e_raw_ingest"]
    assert raw["value"] > 0 and raw["h2d_bytes_per_step"] == 40 * 10980 * 10 * 2 and raw["d2h_bytes_per_step"] == e["d2h_bytes_per_step"]
    r = d["roofline"]
    for key in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert key in r, key
    # the dominant kernel is bound by the fp64 pipe; its HBM figure and the HBM-bound solve kernel are nested
    assert r["bound"] == "fp64" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-12
    assert r["hbm"]["frac"] > 0 and r["solve_kernel"]["bound"] == "hbm" and 0.0 < r["solve_kernel"]["frac"] < 1.2
    c = d["cpu_baseline"]
    assert c["kind"] == "reference" and c["cores"] == 1 and c["value"] > 0 and "sample" in c
    assert c["port"]["kind"] == "port" and c["port"]["value"] > c["value"]
    assert set(("sm_mhz", "sm_max_mhz", "reasons")) <= set(d["clocks"])


def test_other_workloads_run():
    for w in ("c2", "c3"):
        out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--workload", w, "--steps", "1"],
                             capture_output=True, text=True, cwd=ROOT, timeout=600)
        assert out.returncode == 0, out.stderr[-2000:]
        d = json.loads(out.stdout.strip().splitlines()[-1])
        assert d["value"] > 0
