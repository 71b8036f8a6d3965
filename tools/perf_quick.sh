#!/bin/bash
# quick perf + parity loop for the fused Clifford path (run under gpurun): parity tests of the fused kernels, then the
# masks workload (one launch pair per step) and the default cohort-tree headline, kernels broken out
tag=${1:-q}
timeout 400 python -m pytest tests/test_gpu_program.py tests/test_gpu_headline.py tests/test_gpu_cohorts.py tests/test_gpu_workloads.py -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest.log
timeout 200 python bench.py --workload firstgenrepeater_clifford_masks --steps 5 --warmup 3 --extras= --no-cpu > gpurun_out/${tag}_masks.json 2> gpurun_out/${tag}_masks.err; echo "masks rc=$?"
timeout 300 python bench.py --steps 5 --warmup 3 --extras= --no-cpu > gpurun_out/${tag}_head.json 2> gpurun_out/${tag}_head.err; echo "head rc=$?"
python - <<PY
import json
for f in ("gpurun_out/${tag}_masks.json", "gpurun_out/${tag}_head.json"):
    try:
        d = json.load(open(f))
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, "value %.4g  ms/step %.3f  e2e %.4g  parity %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], (d.get("parity_check") or {}).get("equal")))
    for k in d.get("kernels", [])[:6]:
        print("   ", k["kernel"], "launches", k["launches"], "ms %.3f" % k["ms"])
PY
