"""Turn an ncu report (ncu --set full ... -o X) into the small summaries committed here.

    python profiles/summarize_ncu.py gpurun_out/prof.ncu-rep profiles/r01_ncu_kernels 2196000
    python profiles/summarize_ncu.py gpurun_out/prof_raw.csv profiles/r02_ncu_kernels 2196000

Writes <out>.json (per kernel: duration, DRAM bytes, pipe/issue/occupancy figures, top stall
reasons, bytes per pixel) and <out>.txt (the same as a table).  The third argument is the pixel
count of the profiled launch, used for the per-pixel DRAM traffic that bench.py scales into
roofline.traffic.
"""
import csv
import json
import subprocess
import sys

METRICS = {
    "duration_ms": "gpu__time_duration.sum",
    "dram_read_gb": "dram__bytes_read.sum",
    "dram_write_gb": "dram__bytes_write.sum",
    "dram_pct_of_peak": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "fp64_pipe_pct": "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "registers_per_thread": "launch__registers_per_thread",
    "grid_size": "launch__grid_size",
    "block_size": "launch__block_size",
    "warp_instructions": "smsp__inst_executed.sum",
    "smem_bank_conflicts": "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "tensor_pipe_pct": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
}
SCALE = {"Gbyte": 1.0, "Mbyte": 1e-3, "Kbyte": 1e-6, "byte": 1e-9, "ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}


def main(rep, out, n_pix):
    if rep.endswith(".csv"):          # the raw page as exported on the GPU box (profiles/capture.sh)
        raw = open(rep).read()
    else:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    kernels = []
    for r in rows[2:]:
        k = {"kernel": r[idx["Kernel Name"]]}
        for name, metric in METRICS.items():
            if metric not in idx:
                continue
            try:
                v = float(r[idx[metric]].replace(",", ""))
            except ValueError:
                continue
            k[name] = v * SCALE.get(units[idx[metric]], 1.0) if name.endswith(("_gb", "_ms")) else v
        stalls = {}
        for h, i in idx.items():
            if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
                try:
                    stalls[h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]] = float(r[i])
                except ValueError:
                    pass
        k["top_stalls_per_issue"] = dict(sorted(stalls.items(), key=lambda kv: -kv[1])[:5])
        if n_pix and k.get("grid_size", 0) > 1000:
            k["dram_bytes_per_pixel"] = (k["dram_read_gb"] + k["dram_write_gb"]) * 1e9 / n_pix
        kernels.append(k)
    json.dump({"report": rep, "pixels_per_launch": n_pix, "kernels": kernels}, open(out + ".json", "w"), indent=1)
    with open(out + ".txt", "w") as f:
        for k in kernels:
            f.write(k["kernel"] + "\n")
            for key, v in k.items():
                if key != "kernel":
                    f.write("    %-24s %s\n" % (key, v))
    print(open(out + ".txt").read())


def launches(csv_path, out, command):
    """Launch list of `ncu --metrics gpu__time_duration.sum --csv --log-file X <command>` -> per-kernel
    count / total / mean / share table (python profiles/summarize_ncu.py launches X.csv out.txt "<command>")."""
    rows = [r for r in csv.reader(open(csv_path)) if len(r) > 5]
    hdr = rows[0]
    ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = {}
    for r in rows[1:]:
        try:
            ms = float(r[iv].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[iu].strip(), 1e-6)
        except ValueError:
            continue
        a = agg.setdefault(r[ik], [0, 0.0])
        a[0] += 1
        a[1] += ms
    ours = {k: v for k, v in agg.items() if "kafka::" in k and "probe" not in k}
    tot = sum(v[1] for v in ours.values())
    with open(out, "w") as f:
        f.write(command + "\n(per-launch times are cold-cache and serialised: compare SHARES)\n")
        for k, (n, ms) in sorted(ours.items(), key=lambda kv: -kv[1][1]):
            f.write("%-78s n=%3d total=%9.3f ms mean=%8.4f ms share_of_kafka_kernels=%5.1f%%\n"
                    % (k[:78], n, ms, ms / n, 100.0 * ms / tot))
        other = sum(v[1] for k, v in agg.items() if k not in ours)
        f.write("everything else (torch elementwise kernels for synthetic data, peak probes): %.3f ms\n" % other)
    print(open(out).read())


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "")
    else:
        main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 0)
