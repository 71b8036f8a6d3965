"""Build profiles/ncu_kernels.json (what bench.py's roofline block reads) from `ncu --set full --page raw --csv` exports.

usage: python tools/ncu_kernels_json.py <workload>=<raw.csv>:<trajectories>[:<launch index per kernel> | :last<N>] ...
The first launch of each kernel in a capture is used unless an index is given (captures start after warm-up).  `last<N>` SUMS
the last N launches of each kernel (one step of a cohort tree = N levels): warp instructions, DRAM bytes and time are totals of
the step, the shape / issue / stall columns are those of the largest launch, `launches` lists every launch.

The file also records where the facts come from: the commit checked out when this tool ran and a hash of every kernel source
file, so that bench.py can tell when a kernel was edited after its capture."""
import hashlib, json, os, subprocess, sys
sys.path.insert(0, os.path.dirname(__file__))
from ncu_summary import rows

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main(argv):
    path = os.path.join(ROOT, "profiles", "ncu_kernels.json")
    out = json.load(open(path)) if os.path.exists(path) else {}
    sources = {"k_tab_exec": ["qsb_exec_kernel.cuh", "qsb_exec_inst.cu", "qsb_tab_core.cuh", "qsb_exec.cu"],
               "k_draw_bits": ["qsb_exec.cu", "qsb_tab_core.cuh"], "k_dense_exec": ["qsb_dexec.cu", "qsb_dexec.cuh"]}
    csrc = os.path.join(ROOT, "quantumsavory.jl_b200", "csrc")
    commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], stdout=subprocess.PIPE, text=True).stdout.strip() or None
    out["_about"] = {"text": "Per-launch facts from `ncu --set full --clock-control none` captures on a B200 (tools/r03_ncu.sh); bench.py scales "
                             "warp_inst and dram_bytes by lanes processed / trajectories to fill roofline.achieved / frac / traffic. Key: workload|kernel.",
                     "captured_at_commit": commit, "kernel_sources": sources,
                     "sources_sha256_16": {f: hashlib.sha256(open(os.path.join(csrc, f), "rb").read()).hexdigest()[:16] for f in sorted({x for v in sources.values() for x in v})}}
    for a in argv:
        wl, rest = a.split("=", 1)
        parts = rest.split(":")
        csv_path, traj = parts[0], int(float(parts[1]))
        last = int(parts[2][4:]) if len(parts) > 2 and parts[2].startswith("last") else 0
        idx = int(parts[2]) if len(parts) > 2 and not last else 0
        seen = {}
        for d in rows(csv_path):
            k = d["kernel"].replace("void ", "").split("<")[0]
            k = {"k_draw_items": "k_draw_bits"}.get(k, k)   # the profile row of the draw pass keeps its round-1 name
            seen.setdefault(k, []).append(d)
        for k, lst in seen.items():
            d = lst[min(idx, len(lst) - 1)]
            tail = None
            if last:
                tail = lst[-last:]
                d = dict(max(tail, key=lambda r: r["grid"] * r["block"]))
                for f in ("warp_inst", "dram_rd", "dram_wr", "ns"):
                    d[f] = sum(r[f] for r in tail)
            out[f"{wl}|{k}"] = {"warp_inst": d["warp_inst"], "dram_bytes": d["dram_rd"] + d["dram_wr"], "dram_read": d["dram_rd"], "dram_write": d["dram_wr"],
                                "trajectories": traj, "grid": int(d["grid"]), "block": int(d["block"]), "regs": int(d["regs"]), "issue_active": d["issue_active"],
                                "warps_per_scheduler": d["warps_active_per_sched"], "ncu_time_us": d["ns"] / 1e3,
                                "stalls_per_issue": {s[6:]: round(d[s], 3) for s in d if s.startswith("stall_")},
                                "source": "profiles/" + os.path.basename(csv_path).replace("_raw.csv", "") + " (ncu --set full, see profiles/r03_ncu.md)"}
            if tail:
                out[f"{wl}|{k}"]["launches"] = [{"grid": int(r["grid"]), "block": int(r["block"]), "us": r["ns"] / 1e3, "warp_inst": r["warp_inst"], "issue_active": r["issue_active"],
                                                 "warps_per_scheduler": r["warps_active_per_sched"], "dram_bytes": r["dram_rd"] + r["dram_wr"]} for r in tail]
    json.dump(out, open(path, "w"), indent=1)
    print("wrote", path, len(out) - 1, "entries")


if __name__ == "__main__":
    main(sys.argv[1:])
