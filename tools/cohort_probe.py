"""Tuning aid: cohort tree of bench.py's firstgenrepeater Clifford workload -- programs, levels, time per batch."""
import sys, time, json
sys.path.insert(0, ".")
import torch
import bench
import qsb200  # noqa
from qsb200.backend import BatchContext, CudaBatchRepr
from qsb200.cohorts import CohortTree
from qsb200.registers import observable_options
from qsb200.sharding import StatsBuffer

ntraj, depth, batches = int(sys.argv[1]), (None if sys.argv[2] == "none" else int(sys.argv[2])), int(sys.argv[3])
stream = torch.cuda.Stream()
ctx = BatchContext(ntraj, seed=bench.SEED, stream=stream.cuda_stream)
stats = StatsBuffer(256, "cuda:0")
slot = [0]
def next_row():
    i = slot[0]; slot[0] += 1
    return stats.slot_ptr(i)
def schedule():
    slot[0] = 0
    with observable_options(stats_dev=next_row, sync=False):
        bench.wl_fgr(lambda: CudaBatchRepr(ctx, "clifford"), "clifford", lambda *a: None)
with torch.cuda.stream(stream):
    t0 = time.perf_counter()
    tree = CohortTree(ctx, schedule, max_depth=depth)
    print("root recorded in %.1f ms" % (1e3 * (time.perf_counter() - t0)), tree.root.program.info())
    for b in range(batches):
        stats.t.zero_()
        torch.cuda.synchronize()
        rec0 = tree.record_ms
        t0 = time.perf_counter()
        info = tree.run(traj_shift=b * ntraj)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        rows = stats.t.cpu().numpy()
        print(json.dumps({"batch": b, "ms": 1e3 * dt, "record_ms": tree.record_ms - rec0, **info, "alive_end": rows[251, 0], "zz_end": rows[251, 1] / max(rows[251, 0], 1),
                          "launch_shape": None}))
