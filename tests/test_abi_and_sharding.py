"""CPU-only: the C-ABI library loads and exports every symbol include/qsb200.h declares, fails loudly
without a GPU (no fallback), and the multi-GPU host logic (trajectory sharding + one all-reduce of the
statistics) works at world_size 2 over gloo."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from qsb200 import _lib
from qsb200.sharding import StatsBuffer, shard_range

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "qsb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(qsb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = _lib.lib()
    names = declared_symbols()
    assert len(names) >= 36
    for n in names:
        assert hasattr(L, n), f"{n} declared in qsb200.h but not exported"
    assert set(_lib.EXPORTS) == set(names), set(_lib.EXPORTS) ^ set(names)
    assert L.qsb_version() == 100
    out = subprocess.run(["cuobjdump", "-lelf", _lib.SO_PATH], stdout=subprocess.PIPE, text=True).stdout
    assert "sm_100a" in out and "sm_90" not in out and "sm_80" not in out  # sm_100a only, no other arch


def test_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    L = _lib.lib()
    p = C.c_void_p()
    rc = L.qsb_ctx_create(0, 64, 0, 1, None, C.byref(p))
    assert rc == _lib.QSB_ERR_CUDA and b"no CPU fallback" in L.qsb_last_error()
    from qsb200.backend import BatchContext
    with pytest.raises(_lib.QsbError, match="no CUDA device"):
        BatchContext(8)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "quantumsavory.jl_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert "qsb_oracle" not in src and "libqsb_oracle" not in src, f


def test_shard_ranges_partition_all_trajectories():
    for total in (1, 31, 32, 100, 10**6, 10**6 + 7):
        for world in (1, 2, 3, 4, 8):
            covered = 0
            for r in range(world):
                off, cnt = shard_range(total, r, world)
                assert off == covered and cnt >= 0
                assert cnt == 0 or off % 32 == 0
                covered += cnt
            assert covered == total


_WORKER = r'''
import os, sys
sys.path.insert(0, sys.argv[1])
import numpy as np, torch, torch.distributed as dist
import qsb200
from qsb200.sharding import StatsBuffer, shard_range
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
total = 1000
off, cnt = shard_range(total, rank, world)
vals = np.cos(np.arange(off, off + cnt))          # a per-trajectory observable keyed by GLOBAL id
buf = StatsBuffer(2, "cpu")
buf.t[0] = torch.tensor([cnt, vals.sum(), (vals ** 2).sum(), 0.0], dtype=torch.float64)
buf.t[1] = torch.tensor([cnt, float(cnt), float(cnt), 0.0], dtype=torch.float64)
buf.allreduce()
h = buf.host()
ref = np.cos(np.arange(total))
assert h[0, 0] == total and abs(h[0, 1] - ref.sum()) < 1e-9 and abs(h[0, 2] - (ref ** 2).sum()) < 1e-9
assert abs(buf.means()[1] - 1.0) < 1e-15
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok")
'''


def test_stats_allreduce_world2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29531", WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script), ROOT], env=dict(env, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and f"rank {r} ok" in o, o
