"""Record the firstgenrepeater Clifford program and run it a few times (target for ncu captures)."""
import sys
sys.path.insert(0, ".")
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
g = bench.GpuRun("firstgenrepeater_clifford", n, 0, 1, 0)
g.record_program()
print(g.program.info())
for _ in range(int(sys.argv[2]) if len(sys.argv) > 2 else 2):
    g.run_program()
g.ctx.sync()
print("ok")
