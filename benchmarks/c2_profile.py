"""Where do the 0.5 ms per date of the small-image run (BASELINE configs[1], Barrax mask) go on the host?

    python benchmarks/c2_profile.py [reps]

cProfile over `reps` runs of the 20-date LinearKalman.run that bench.py times as configs1_barrax;
prints the functions by own time and by cumulative time, and the un-profiled ms per date."""
import cProfile
import os
import pstats
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    device = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    first = bench.measure_c2(device, reps)
    print("un-profiled: %.4f ms per date" % first["ms_per_date"])
    prof = cProfile.Profile()
    prof.enable()
    bench.measure_c2(device, reps)
    prof.disable()
    st = pstats.Stats(prof)
    st.sort_stats("tottime").print_stats(40)
    st.sort_stats("cumulative").print_stats(45)


if __name__ == "__main__":
    main()
