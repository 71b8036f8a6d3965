"""TMA experiment: k_apply_dm<2> (register path, coalesced 128-bit loads) against k_apply_dm2_tma
(cp.async.bulk + mbarrier double-buffered shared-memory stages) on the shapes where a block row is short.
Usage (on a B200): python tools/tma_experiment.py > gpurun_out/tma_experiment.json"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qsb200  # noqa: F401,E402
from qsb200 import symbols as S  # noqa: E402
from qsb200._lib import DM  # noqa: E402
from qsb200.backend import BatchContext, CudaBatchState  # noqa: E402


def haar(rng, d):
    q, r = np.linalg.qr(rng.normal(size=(d, d)) + 1j * rng.normal(size=(d, d)))
    return q * (np.diag(r) / np.abs(np.diag(r)))


def run(n, ntraj, pairs, variants, reps=4):
    rng = np.random.default_rng(1)
    U = haar(rng, 4)
    out = []
    ctx = BatchContext(ntraj, seed=3)
    st = CudaBatchState.zeros(ctx, DM, n)
    for v in variants:
        if v:
            os.environ["QSB_APPLY_TMA"] = str(v)
        else:
            os.environ.pop("QSB_APPLY_TMA", None)
        for pair in pairs:
            st.apply(list(pair), U); ctx.sync()          # warm-up
            ctx.profile(True)
            for _ in range(reps):
                st.apply(list(pair), U)
            ctx.sync()
            rep = ctx.profile_report()
            ctx.profile(False)
            row = [r for r in rep if r["kernel"] == "k_apply_dm"][0]
            ms = row["ms"] / row["launches"]
            gbs = row["algo_bytes"] / row["launches"] / (ms * 1e-3) / 1e9
            out.append({"n": n, "ntraj": ntraj, "pair": pair, "variant": "tma x%d blocks/SM" % v if v else "registers", "ms": ms, "GBps": gbs})
    st.free(); ctx.close()
    return out


if __name__ == "__main__":
    res = []
    # parity of the two paths first (small state, bit-for-bit the same arithmetic order per element)
    ctx = BatchContext(200, seed=5)
    rng = np.random.default_rng(0)
    a = rng.normal(size=(16, 16)) + 1j * rng.normal(size=(16, 16)); rho = a @ a.conj().T; rho /= np.trace(rho)
    U = haar(rng, 4)
    got = {}
    for v in (0, 3):
        os.environ["QSB_APPLY_TMA"] = str(v)
        s = CudaBatchState.from_array(ctx, rho, DM)
        s.apply([4, 2], U); s.apply([1, 3], U)
        got[v] = np.stack([s.trajectory(t) for t in (0, 131, 199)])
        s.free()
    ctx.close()
    res.append({"parity_max_abs_diff": float(np.abs(got[0] - got[3]).max())})
    res += run(12, 256, [(1, 2), (6, 12), (11, 12)], (0, 3, 2, 1))
    res += run(8, 16384, [(1, 2), (4, 8), (7, 8)], (0, 3, 2))
    res += run(10, 1000, [(1, 2), (5, 10)], (0, 3, 2))
    res += run(4, 1000000, [(1, 2), (3, 4)], (0, 3, 2))
    print(json.dumps(res, indent=1))
