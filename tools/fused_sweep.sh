#!/bin/bash
B="timeout 200 python bench.py --workload firstgenrepeater_clifford_masks --steps 5 --warmup 3 --extras= --no-cpu"
for cfg in "1 8" "2 8" "2 1" "2 2" "2 4"; do
  set -- $cfg
  QSB_DRAW_WARPS_RATIO=$1 QSB_DRAW_SLICES=$2 $B 2>/dev/null | python -c "import sys,json; d=json.load(sys.stdin); print('ratio $1 slices $2: ms/step %.3f' % d['ms_per_step'], [(k['kernel'], round(k['ms'],3)) for k in d['kernels'][:2]])"
done
QSB_FUSED_DRAW=0 $B 2>/dev/null | python -c "import sys,json; d=json.load(sys.stdin); print('unfused: ms/step %.3f' % d['ms_per_step'], [(k['kernel'], round(k['ms'],3)) for k in d['kernels'][:2]])"
