#!/bin/bash
# the repeatergrid program (10^7 lanes): arena placement of the executor.  QSB_EXEC_GLOBAL=0 forces the shared-memory tile
# (42-thread blocks, 50 waves), QSB_EXEC_GLOBAL=1 the plane-major global arena (the default past 4 waves).
# Register-cap variants of the global-arena instantiation: rebuild qsb_exec_inst.cu with -DQSB_EXEC_T=0 -DQSB_EXEC_G_MINB=<4|6|8>
# and point QSB200_LIB at the variant library (measured: 113 / 93.5 / 100 ms).
G="timeout 300 python bench.py --workload repeatergrid_clifford --steps 2 --warmup 3 --extras= --no-cpu"
P='import sys,json; d=json.load(sys.stdin); print(sys.argv[1], "ms/step %.2f value %.4g" % (d["ms_per_step"], d["value"]), [(k["kernel"], round(k["ms"],2)) for k in d["kernels"][:2]], "global" if d["program"]["global_arena"] else "smem", "parity", (d.get("parity_check") or {}).get("equal"))'
for v in 0 1; do QSB_EXEC_GLOBAL=$v $G 2>/dev/null | python -c "$P" QSB_EXEC_GLOBAL=$v; done
