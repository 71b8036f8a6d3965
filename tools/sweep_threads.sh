#!/bin/bash
# block-size sweep of the fused executor (tuning aid): prints kernel times per forced block size
for t in 32 64 96 128 256 auto; do
  if [ $t = auto ]; then unset QSB_EXEC_THREADS; else export QSB_EXEC_THREADS=$t; fi
  python bench.py --steps 3 --warmup 3 --extras "" --no-cpu > gpurun_out/sweep_$t.json 2> gpurun_out/sweep_$t.err || tail -2 gpurun_out/sweep_$t.err
  echo "== T=$t"; python tools/show.py gpurun_out/sweep_$t.json | head -4
done
