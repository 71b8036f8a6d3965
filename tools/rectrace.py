import sys; sys.path.insert(0,'.')
import bench
g = bench.GpuRun("firstgenrepeater_clifford", 1024, 0, 1, 0)
g.record_program()
print(g.program.info())
