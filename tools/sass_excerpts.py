"""Regenerates profiles/r02_sass_excerpts.md: opcode histograms and the characteristic SASS lines of the round-2 kernels
(`cuobjdump -sass` of the objects under quantumsavory.jl_b200/csrc; run `make -C quantumsavory.jl_b200/csrc` first)."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "quantumsavory.jl_b200", "csrc")


def sass(obj, fn):
    out = subprocess.run(["cuobjdump", "-sass", os.path.join(CSRC, obj)], capture_output=True, text=True).stdout
    keep, on = [], False
    for line in out.splitlines():
        if "Function :" in line:
            on = fn in line
        if on:
            keep.append(line)
    return keep


def ops(lines):
    c = collections.Counter()
    for l in lines:
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(@!?U?P[0-9T]+\s+)?([A-Z0-9_.]+)", l)
        if m:
            c[m.group(2).split(".")[0]] += 1
    return c


KERNELS = (("qsb_exec.o", "k_draw_items", "k_draw_items (draw pass): one Philox block per item, round keys from the constant bank", r"IMAD\.WIDE|LOP3|VOTE|LDCU|STS|BAR"),
           ("qsb_exec_t1.o", "k_tab_exec", "k_tab_exec<1> (shared-memory executor)", r"LDGSTS|LDS|STS|LOP3|BAR|LDG|ATOM|RED"),
           ("qsb_dense.o", "k_apply_dmILi2", "k_apply_dm<2> (register path)", r"LDG|STG|DFMA|DMUL"),
           ("qsb_dense.o", "k_apply_dm2_tma", "k_apply_dm2_tma (bulk-copy staged variant, experiment)", r"UBLKCP|SYNCS|FENCE|LDS|STS"),
           ("qsb_evolve.o", "k_apply_superILi2", "k_apply_super<2> (qsb_evolve)", r"LDG|STG|DFMA|LDS"))

if __name__ == "__main__":
    doc = ["# SASS excerpts of the round-2 kernels (`cuobjdump -sass`, sm_100a, CUDA 12.9; regenerate with tools/sass_excerpts.py)", ""]
    for obj, fn, title, grep in KERNELS:
        lines = sass(obj, fn)
        c = ops(lines)
        doc += [f"## {title}", "", f"{sum(c.values())} instructions; opcode histogram: " + ", ".join(f"{k} {v}" for k, v in c.most_common(14)), "", "```"]
        shown = 0
        for l in lines:
            if re.search(grep, l) and shown < 28:
                doc.append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l).rstrip()[:140]); shown += 1
        doc += ["```", ""]
    open(os.path.join(ROOT, "profiles", "r02_sass_excerpts.md"), "w").write("\n".join(doc) + "\n")
    print("wrote profiles/r02_sass_excerpts.md")
