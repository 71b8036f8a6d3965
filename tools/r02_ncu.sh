#!/bin/bash
# ncu evidence of round 2 (run under gpurun; every command first runs plain, then under ncu).  Reports are exported to CSV on the
# box and the large .ncu-rep files dropped (gpurun_out/ comes back only when it stays under 64 MiB).
export QSB_PROFILING=1
exp() { ncu -i gpurun_out/$1.ncu-rep --page raw --csv > gpurun_out/$1_raw.csv 2>/dev/null; }
B="python bench.py --steps 2 --warmup 1 --extras= --no-cpu"
$B > gpurun_out/r02_plain_head.json 2> gpurun_out/r02_plain_head.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_head.csv $B > /dev/null 2>&1
ncu --set full --clock-control none -k regex:'k_tab_exec|k_draw|k_cohort' -c 16 -o gpurun_out/r02_head_cohort $B > /dev/null 2>&1
exp r02_head_cohort; rm -f gpurun_out/r02_head_cohort.ncu-rep
M="python bench.py --workload firstgenrepeater_clifford_masks --steps 1 --warmup 1 --extras= --no-cpu"
$M > gpurun_out/r02_plain_masks.json 2>/dev/null &&
ncu --set full --clock-control none --import-source on -k regex:'k_tab_exec|k_draw' -c 2 -o gpurun_out/r02_masks $M > /dev/null 2>&1
exp r02_masks; ncu -i gpurun_out/r02_masks.ncu-rep --page source --csv > gpurun_out/r02_masks_source.csv 2>/dev/null
G="python bench.py --workload repeatergrid_clifford --steps 1 --warmup 1 --extras= --no-cpu"
$G > gpurun_out/r02_plain_grid.json 2>/dev/null &&
ncu --set full --clock-control none -k regex:'k_tab_exec|k_draw' -c 2 -o gpurun_out/r02_grid $G > /dev/null 2>&1
exp r02_grid; rm -f gpurun_out/r02_grid.ncu-rep
D="python bench.py --workload firstgenrepeater_dm --steps 1 --warmup 1 --extras= --no-cpu"
$D > gpurun_out/r02_plain_dm.json 2>/dev/null &&
ncu --set full --clock-control none -k regex:'k_dense_exec' -c 1 -o gpurun_out/r02_dm $D > /dev/null 2>&1
exp r02_dm; rm -f gpurun_out/r02_dm.ncu-rep
gzip -f gpurun_out/r02_masks_source.csv
ls -la gpurun_out/; du -sh gpurun_out
