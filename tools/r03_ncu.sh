#!/bin/bash
# ncu evidence of round 3 (run under gpurun, one GPU; every command first runs plain, then under ncu).  Reports are exported to
# CSV on the box and the .ncu-rep files dropped (gpurun_out/ comes back only when it stays under 64 MiB).
export QSB_PROFILING=1
exp() { ncu -i gpurun_out/$1.ncu-rep --page raw --csv > gpurun_out/$1_raw.csv 2>/dev/null; rm -f gpurun_out/$1.ncu-rep; }
# (1) launch list of the headline command
B="python bench.py --steps 2 --warmup 1 --extras= --no-cpu"
$B > gpurun_out/r03_plain_head.json 2> gpurun_out/r03_plain_head.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r03_launches_head.csv $B > /dev/null 2>&1
# (2) the cohort tree's OWN wave launches: two batches of 10^6 lanes, 4 levels x (draw pass + executor) each; the second batch is the
#     steady state (all 805 programs recorded), tools/ncu_kernels_json.py sums its launches (spec `last4`)
C="python tools/cohort_probe.py 1000000 3 2"
$C > gpurun_out/r03_plain_cohort.log 2>&1 &&
ncu --set full --clock-control none -k regex:'k_tab_exec|k_draw' -c 16 -o gpurun_out/r03_cohort $C > /dev/null 2>&1
exp r03_cohort
# (3) single-program launches (lockstep masks), the grid, the dense executor
M="python bench.py --workload firstgenrepeater_clifford_masks --steps 1 --warmup 1 --extras= --no-cpu"
$M > gpurun_out/r03_plain_masks.json 2>/dev/null &&
ncu --set full --clock-control none -k regex:'k_tab_exec|k_draw' -c 2 -o gpurun_out/r03_masks $M > /dev/null 2>&1
exp r03_masks
G="python bench.py --workload repeatergrid_clifford --steps 1 --warmup 1 --extras= --no-cpu"
$G > gpurun_out/r03_plain_grid.json 2>/dev/null &&
ncu --set full --clock-control none -k regex:'k_tab_exec|k_draw' -c 2 -o gpurun_out/r03_grid $G > /dev/null 2>&1
exp r03_grid
D="python bench.py --workload firstgenrepeater_dm --steps 1 --warmup 1 --extras= --no-cpu"
$D > gpurun_out/r03_plain_dm.json 2>/dev/null &&
ncu --set full --clock-control none -k regex:'k_dense_exec' -c 1 -o gpurun_out/r03_dm $D > /dev/null 2>&1
exp r03_dm
ls -la gpurun_out/; du -sh gpurun_out
