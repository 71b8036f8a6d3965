"""Print the headline fields of a bench.py JSON line (helper for reading gpurun output)."""
import json
import sys
d = json.load(open(sys.argv[1]))
print("value", d["value"], "ms/step", d["ms_per_step"], "e2e", d["e2e"]["value"], "launches", d["gpu_launches"])
print("program", d.get("program"))
for k in d.get("kernels") or []:
    print("  ", k["kernel"], k["launches"], round(k["ms"], 3), "ms", round(k["gbs"], 1), "GB/s")
print("cpu", d.get("cpu_baseline"))
for k, v in (d.get("extra") or {}).items():
    print(k, {a: b for a, b in v.items() if a not in ("kernels", "roofline")})
