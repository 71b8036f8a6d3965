"""Workload driver: examples/repeatergrid of the reference (BASELINE.json configs[3]).

An n x n grid of repeater nodes; nearest neighbours generate Bell pairs, every node but the two corners
swaps "towards" the corners, the corners (vertex 1 and vertex n^2) consume the pairs they share, and a
cutoff process discards qubits that sat too long (examples/repeatergrid/setup.jl:40-95).

The protocols themselves (EntanglerProt, SwapperProt, EntanglementTracker, EntanglementConsumer, CutoffProt)
are the reference's control plane and stay on the host; this file restates just enough of
their classical logic to emit the op trace the backend sees:

  * entangler per edge  -- src/ProtocolZoo/ProtocolZoo.jl EntanglerProt: random free slot on both ends
    (`randomize=true`), Geometric(success_prob) attempts of `attempt_time`, then `initialize!` of the pair;
  * swapper per node    -- src/ProtocolZoo/swapping.jl:6-27,99-220 with the quadrant predicates and
    farthest-node choosers of setup.jl:8-37: `uptotime!`, `LocalEntanglementSwap`, then the two update
    messages; the remote EntanglementTracker applies Z (for the X outcome) / X (for the Z outcome) only
    where the reported outcome is 2 (ProtocolZoo.jl:505,569-573) -- a per-trajectory mask here.  Classical
    channels have zero delay, so the correction lands in the same instant and no forwarding is needed;
  * consumer            -- ProtocolZoo.jl:868-887: <ZZ>, <XX> of the corner pair, then `traceout!` of both;
  * cutoff per slot     -- src/ProtocolZoo/cutoff.jl:60-110 with `announce=true`: `traceout!` of a slot whose
    tag is older than `retention_time`, and of its counterpart on receipt of EntanglementDelete.

All classical randomness (attempt counts, slot picks) comes from one seeded numpy generator and is shared by
the lockstep ensemble; measurement outcomes are per trajectory.  BASELINE's config asks for the Clifford
backend, for which the reference has no `T1Decay` noise (src/backends/clifford/uptotime.jl:26-43), so the
background is `Depolarization(tau)`.
"""
from __future__ import annotations

import numpy as np

from .. import ops
from ..circuits import LocalEntanglementSwap
from ..registers import Register, apply, initialize, isassigned, observable, traceout, uptotime
from ..symbols import Depolarization
from .des import Resource, Simulation

ZZ = ops.tensor(ops.Z, ops.Z)
XX = ops.tensor(ops.X, ops.X)

DEFAULTS = dict(n=5, regsize=4, tau=10.0, success_prob=0.001, attempt_time=0.001, retry_lock_time=0.1,
                local_busy_time=0.0, retention_time=5.0, period_cons=0.1, period_dec=0.1, t_end=30.0, seed=1,
                pair=None)


def _xy(n, v):
    x = v // n if v % n == 0 else v // n + 1       # setup.jl:13-16
    return x, v - n * (x - 1)


def check_nodes(n, c_node, node, low=True):
    """setup.jl:11-18 -- is `node` in the low (high) quadrant of `c_node`?"""
    cx, cy = _xy(n, c_node)
    x, y = _xy(n, node)
    return (cx - x >= 0 and cy - y >= 0) if low else (cx - x <= 0 and cy - y <= 0)


def distance(n, a, b):
    """setup.jl:29-35"""
    x1, y1 = _xy(n, a)
    x2, y2 = _xy(n, b)
    return x1 - x2 + y1 - y2


def grid_edges(n):
    """Graphs.grid([n, n]): vertex v=(x-1)n+y is joined to its right and upper neighbours."""
    out = []
    for v in range(1, n * n + 1):
        if v % n != 0:
            out.append((v, v + 1))
        if v + n <= n * n:
            out.append((v, v + n))
    return out


class GridNet:
    def __init__(self, p, make_repr):
        n, rs = p["n"], p["regsize"]
        self.n, self.p = n, p
        self.sim = Simulation()
        bg = p.get("background") or Depolarization(p["tau"])
        self.registers = [Register(rs, make_repr(), bg) for _ in range(n * n)]
        self.locks = [[Resource(self.sim) for _ in range(rs)] for _ in range(n * n)]
        # EntanglementCounterpart tag of a slot: (remote node, remote slot, tag time)
        self.tag = [[None] * rs for _ in range(n * n)]
        self.rng = np.random.default_rng(p["seed"])
        self.log = {"pairs": 0, "swaps": 0, "consumed": 0, "cutoffs": 0}

    def reg(self, v):
        return self.registers[v - 1]

    def lock(self, v, i):
        return self.locks[v - 1][i - 1]

    def free_slots(self, v):
        reg = self.reg(v)
        return [i for i in range(1, len(reg) + 1)
                if not isassigned(reg, i) and not self.lock(v, i).islocked() and self.tag[v - 1][i - 1] is None]

    def settag(self, v, i, val):
        self.tag[v - 1][i - 1] = val


def entangler(net: GridNet, a, b):
    """EntanglerProt(sim, net, a, b; rounds=-1, randomize=true, success_prob) -- setup.jl:55-58."""
    sim, p = net.sim, net.p
    pair = p["pair"] or ops.StabilizerState("XX ZZ")
    while True:
        fa, fb = net.free_slots(a), net.free_slots(b)
        if not fa or not fb:
            yield sim.timeout(p["retry_lock_time"])
            continue
        ia, ib = int(net.rng.choice(fa)), int(net.rng.choice(fb))
        yield net.lock(a, ia).request() & net.lock(b, ib).request()
        attempts = int(net.rng.geometric(p["success_prob"]))
        yield sim.timeout(attempts * p["attempt_time"])
        initialize((net.reg(a)[ia], net.reg(b)[ib]), pair, time=sim.now)
        net.settag(a, ia, (b, ib, sim.now))
        net.settag(b, ib, (a, ia, sim.now))
        net.log["pairs"] += 1
        net.lock(a, ia).release()
        net.lock(b, ib).release()


def findswapablequbits(net: GridNet, node):
    """swapping.jl:6-27 with nodeL/nodeH = check_nodes and chooseL/chooseH = farthest node (setup.jl:66-71)."""
    n = net.n
    tags = net.tag[node - 1]
    cand = [(i, t) for i, t in enumerate(tags, 1)
            if t is not None and not net.lock(node, i).islocked() and isassigned(net.reg(node), i)]
    low = [(i, t) for i, t in cand if check_nodes(n, node, t[0], low=True)]
    high = [(i, t) for i, t in cand if check_nodes(n, node, t[0], low=False)]
    if not low or not high:
        return None
    il = int(np.argmax([distance(n, node, t[0]) for _, t in low]))
    ih = int(np.argmin([distance(n, node, t[0]) for _, t in high]))
    if low[il][0] == high[ih][0]:
        return None
    return low[il], high[ih]


def swapper(net: GridNet, node):
    """SwapperProt -- swapping.jl:99-220; the two update messages are delivered at once (zero-delay channels)."""
    sim, p = net.sim, net.p
    circuit = LocalEntanglementSwap()
    while True:
        found = findswapablequbits(net, node)
        if found is None:
            yield sim.timeout(p["retry_lock_time"])
            continue
        (i1, t1), (i2, t2) = found
        yield net.lock(node, i1).request() & net.lock(node, i2).request()
        if net.tag[node - 1][i1 - 1] != t1 or net.tag[node - 1][i2 - 1] != t2:   # re-query under lock, swapping.jl:155-163
            net.lock(node, i1).release()
            net.lock(node, i2).release()
            continue
        reg = net.reg(node)
        q1, q2 = reg[i1], reg[i2]
        net.settag(node, i1, None)
        net.settag(node, i2, None)
        uptotime((q1, q2), sim.now)
        xmeas, zmeas = circuit(q1, q2)
        (n1, s1, _), (n2, s2, _) = t1, t2
        apply(net.reg(n1)[s1], ops.Z, mask=(xmeas == 2))   # EntanglementUpdateX -> Z (ProtocolZoo.jl:505)
        apply(net.reg(n2)[s2], ops.X, mask=(zmeas == 2))   # EntanglementUpdateZ -> X
        net.settag(n1, s1, (n2, s2, sim.now))
        net.settag(n2, s2, (n1, s1, sim.now))
        net.log["swaps"] += 1
        yield sim.timeout(p["local_busy_time"])
        net.lock(node, i1).release()
        net.lock(node, i2).release()


def consumer(net: GridNet, a, b, record):
    """EntanglementConsumer(sim, net, 1, n^2; period) -- ProtocolZoo.jl:868-887."""
    sim, p = net.sim, net.p
    while True:
        found = [(i, t) for i, t in enumerate(net.tag[a - 1], 1)
                 if t is not None and t[0] == b and not net.lock(a, i).islocked() and not net.lock(b, t[1]).islocked()]
        if not found:
            yield sim.timeout(p["period_cons"])
            continue
        i, (_, j, _) = found[0]
        yield net.lock(a, i).request() & net.lock(b, j).request()
        if net.tag[a - 1][i - 1] is None or net.tag[a - 1][i - 1][:2] != (b, j):      # ProtocolZoo.jl:830-853
            net.lock(a, i).release()
            net.lock(b, j).release()
            continue
        qa, qb = net.reg(a)[i], net.reg(b)[j]
        net.settag(a, i, None)
        net.settag(b, j, None)
        ob1 = observable((qa, qb), ZZ)
        ob2 = observable((qa, qb), XX)
        traceout(qa, qb)
        if record is not None:
            record("ZZ", sim.now, ob1); record("XX", sim.now, ob2)
        net.log["consumed"] += 1
        net.lock(a, i).release()
        net.lock(b, j).release()
        yield sim.timeout(p["period_cons"])


def cutoff(net: GridNet, node, slot):
    """per_slot_cutoff with announce=true -- cutoff.jl:67-110; the counterpart's tracker answers
    EntanglementDelete with `traceout!(localslot)` (ProtocolZoo.jl:586-588)."""
    sim, p = net.sim, net.p
    while True:
        yield sim.timeout(p["period_dec"])
        yield net.lock(node, slot).request()
        t = net.tag[node - 1][slot - 1]
        if t is None or sim.now - t[2] < p["retention_time"]:
            net.lock(node, slot).release()
            continue
        net.settag(node, slot, None)
        traceout(net.reg(node)[slot])
        rn, rs, _ = t
        if net.tag[rn - 1][rs - 1] is not None and net.tag[rn - 1][rs - 1][:2] == (node, slot):
            net.settag(rn, rs, None)
            traceout(net.reg(rn)[rs])
        net.log["cutoffs"] += 1
        net.lock(node, slot).release()


def run(make_repr, record=None, **kw):
    """prepare_simulation + run(sim, t_end) -- setup.jl:40-95.  record(kind, time, value) gets the consumer's
    <ZZ>, <XX> (BatchValue for an ensemble).  Returns the GridNet (log = classical event counts)."""
    p = dict(DEFAULTS); p.update(kw)
    net = GridNet(p, make_repr)
    sim, n = net.sim, p["n"]
    for a, b in grid_edges(n):
        sim.process(entangler(net, a, b))
    for v in range(2, n * n):
        sim.process(swapper(net, v))
    sim.process(consumer(net, 1, n * n, record))
    for v in range(1, n * n + 1):
        for s in range(1, p["regsize"] + 1):
            sim.process(cutoff(net, v, s))
    sim.run(p["t_end"])
    return net
