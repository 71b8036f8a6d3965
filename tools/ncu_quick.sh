#!/bin/bash
# one --set full capture of the level-0 launch pair of the masks workload (k_draw_bits, k_tab_exec), exported as raw CSV + per-line source CSV
tag=${1:-nq}
M="python bench.py --workload firstgenrepeater_clifford_masks --steps 1 --warmup 1 --extras= --no-cpu"
$M > gpurun_out/${tag}_plain.json 2>/dev/null || exit 1
ncu --set full --clock-control none --import-source on -k regex:'k_tab_exec|k_draw' -c 2 -o gpurun_out/${tag} $M > /dev/null 2>&1
ncu -i gpurun_out/${tag}.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>/dev/null
ncu -i gpurun_out/${tag}.ncu-rep --page source --csv > gpurun_out/${tag}_source.csv 2>/dev/null
ncu -i gpurun_out/${tag}.ncu-rep --page source --csv --print-source cuda -k regex:k_tab_exec > gpurun_out/${tag}_lines.csv 2>/dev/null
rm -f gpurun_out/${tag}.ncu-rep
gzip -f gpurun_out/${tag}_source.csv
python tools/ncu_summary.py gpurun_out/${tag}_raw.csv
python - <<PY
import csv
r = list(csv.reader(open("gpurun_out/${tag}_raw.csv")))
hdr = next(i for i, x in enumerate(r) if x and x[0] == "ID")
names = r[hdr]
for x in r[hdr + 2:]:
    if len(x) != len(names): continue
    print(x[names.index("Kernel Name")][:30])
    for n in names:
        if ("pipe" in n and "pct_of_peak_sustained_active" in n and "inst_executed" in n) or n.startswith("smsp__inst_executed_pipe") :
            v = x[names.index(n)]
            try:
                if float(v.replace(",", "")) > 3: print("   ", n, v)
            except ValueError: pass
PY
