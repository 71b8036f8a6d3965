#!/bin/bash
# the repeatergrid program (10^7 lanes) under the global-arena executor, register-cap variants of the library
G="timeout 300 python bench.py --workload repeatergrid_clifford --steps 2 --warmup 3 --extras= --no-cpu"
P='import sys,json; d=json.load(sys.stdin); print(sys.argv[1], "ms/step %.2f value %.4g" % (d["ms_per_step"], d["value"]), [(k["kernel"], round(k["ms"],2)) for k in d["kernels"][:2]], "global" if d["program"]["global_arena"] else "smem", "parity", d.get("parity_check"))'
$G 2>/dev/null | python -c "$P" default-minb6
for v in 4 8; do QSB200_LIB=$PWD/quantumsavory.jl_b200/csrc/variants/libqsb200_g$v.so $G 2>/dev/null | python -c "$P" minb$v; done
