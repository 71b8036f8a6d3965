"""Build profiles/ncu_kernels.json (what bench.py's roofline block reads) from `ncu --set full --page raw --csv` exports.

usage: python tools/ncu_kernels_json.py <workload>=<raw.csv>:<trajectories>[:<launch index per kernel>] ...
The first launch of each kernel in a capture is used unless an index is given (captures start after warm-up)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(__file__))
from ncu_summary import rows

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main(argv):
    path = os.path.join(ROOT, "profiles", "ncu_kernels.json")
    out = json.load(open(path)) if os.path.exists(path) else {}
    out["_about"] = ("Per-launch facts from `ncu --set full --clock-control none` captures on a B200 (tools/r02_ncu.sh); bench.py scales "
                     "warp_inst and dram_bytes by lanes processed / trajectories to fill roofline.achieved / frac / traffic. Key: workload|kernel.")
    for a in argv:
        wl, rest = a.split("=", 1)
        parts = rest.split(":")
        csv_path, traj = parts[0], int(float(parts[1]))
        idx = int(parts[2]) if len(parts) > 2 else 0
        seen = {}
        for d in rows(csv_path):
            k = d["kernel"].replace("void ", "").split("<")[0]
            k = {"k_draw_items": "k_draw_bits"}.get(k, k)   # the profile row of the draw pass keeps its round-1 name
            seen.setdefault(k, []).append(d)
        for k, lst in seen.items():
            d = lst[min(idx, len(lst) - 1)]
            out[f"{wl}|{k}"] = {"warp_inst": d["warp_inst"], "dram_bytes": d["dram_rd"] + d["dram_wr"], "dram_read": d["dram_rd"], "dram_write": d["dram_wr"],
                                "trajectories": traj, "grid": int(d["grid"]), "block": int(d["block"]), "regs": int(d["regs"]), "issue_active": d["issue_active"],
                                "warps_per_scheduler": d["warps_active_per_sched"], "ncu_time_us": d["ns"] / 1e3,
                                "stalls_per_issue": {s[6:]: round(d[s], 3) for s in d if s.startswith("stall_")},
                                "source": "profiles/" + os.path.basename(csv_path).replace("_raw.csv", "") + " (ncu --set full, see profiles/r02_ncu.md)"}
    json.dump(out, open(path, "w"), indent=1)
    print("wrote", path, len(out) - 1, "entries")


if __name__ == "__main__":
    main(sys.argv[1:])
