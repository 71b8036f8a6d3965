#!/bin/bash
# tile / block-variant sweep of the dense fused executor (tuning aid): QSB_DEXEC_T x QSB_DEXEC_VARIANT (0: 256x1, 1: 256x2, 2: 128x4)
N=${1:-1000000}
IFS=";" read -ra LIST <<< "${CFGS:-auto auto;32 0;16 1;8 2;8 3;8 4;4 3;4 4}"
for cfg in "${LIST[@]}"; do
  set -- $cfg
  if [ $1 = auto ]; then unset QSB_DEXEC_T QSB_DEXEC_VARIANT; else export QSB_DEXEC_T=$1 QSB_DEXEC_VARIANT=$2; fi
  QSB_REC_TRACE=1 python bench.py --workload firstgenrepeater_dm --ntraj $N --steps 3 --extras= --no-cpu > gpurun_out/dx_$1_$2.json 2> gpurun_out/dx_$1_$2.err
  echo "== T=$1 variant=$2"; grep "dense program" gpurun_out/dx_$1_$2.err | tail -1; tail -1 gpurun_out/dx_$1_$2.err | cut -c1-200
  python tools/show.py gpurun_out/dx_$1_$2.json 2>/dev/null | head -1
done
