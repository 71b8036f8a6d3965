"""cProfile of the host side of one per-op pass (tuning aid): python tools/prof_host.py [workload] [ntraj]"""
import cProfile
import os
import pstats
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "firstgenrepeater_clifford"
ntraj = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
g = bench.GpuRun(wl, ntraj, 0, 1, 0)
for _ in range(2):
    g.step(); g.finish_step(reduce=False)
pr = cProfile.Profile()
pr.enable()
for _ in range(3):
    g.step(); g.finish_step(reduce=False)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(28)
