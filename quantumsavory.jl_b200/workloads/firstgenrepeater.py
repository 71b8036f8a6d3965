"""Workload driver: examples/firstgenrepeater_lowlevel of the reference (BASELINE.json configs 0 and 1).

Restates the protocol of /root/reference/examples/firstgenrepeater_lowlevel/setup.jl (entangler :75-111,
swapper :124-164, purifier :185-240, helper queries :114-121,168-181,242-255) and the trajectory
driver 6.1_compare_formalisms_noplot.jl:10-57 on top of registers.py, for any backend representation.

Ensemble semantics (DESIGN.md "Lockstep ensemble"): with ntraj == 1 the host follows the measured
purification outcome exactly as the reference does.  With a lockstep ensemble the classical schedule
is shared: the host always continues on the success branch and the trajectories whose purification
failed lose that pair (`traceout(..., mask=failure)`: the slot is empty on those lanes), so every
later observable is averaged over the trajectories where the pair exists -- `something` on the others.
"""
from __future__ import annotations

from .. import ops
from ..circuits import EntanglementSwap, Purify2to1
from ..registers import Register, initialize, isassigned, observable, uptotime
from ..symbols import Depolarization, T1T2Noise, T2Dephasing  # noqa: F401
from .des import Resource, Simulation

XX = ops.tensor(ops.X, ops.X)
ZZ = ops.tensor(ops.Z, ops.Z)


def stab_noisy_pair_func(F):
    """F*stab_perfect_pair_dm + (1-F)*mixed_dm -- clifford_setup.jl:8-10."""
    return ops.Mixture([(F, ops.StabilizerState("XX ZZ")), (1 - F, ops.MixedState(2))])


class Network:
    def __init__(self, sizes, make_repr, background):
        self.sim = Simulation()
        self.registers = [Register(s, make_repr(), background) for s in sizes]
        self.locks = [[Resource(self.sim) for _ in range(s)] for s in sizes]
        self.enttrackers = [[None] * s for s in sizes]
        self.R = len(sizes)
        self.log = {"pairs": 0, "swaps": 0, "purifications": 0}

    def edges(self):
        return [(v, v + 1) for v in range(1, self.R)]

    def reg(self, v):
        return self.registers[v - 1]

    def lock(self, v, i):
        return self.locks[v - 1][i - 1]

    def islocked(self, v, i):
        return self.lock(v, i).islocked()


def findfreequbit(net, node):
    reg = net.reg(node)
    for i in range(1, len(reg) + 1):
        if not isassigned(reg, i) and not net.islocked(node, i):
            return i
    return None


def entangler(net, nodea, nodeb, noisy_pair, wait_time, busy_time):
    sim = net.sim
    while True:
        ia, ib = findfreequbit(net, nodea), findfreequbit(net, nodeb)
        if ia is None or ib is None:
            yield sim.timeout(wait_time)
            continue
        yield net.lock(nodea, ia).request() & net.lock(nodeb, ib).request()
        yield sim.timeout(busy_time)
        initialize((net.reg(nodea)[ia], net.reg(nodeb)[ib]), noisy_pair, time=sim.now)
        net.enttrackers[nodea - 1][ia - 1] = (nodeb, ib)
        net.enttrackers[nodeb - 1][ib - 1] = (nodea, ia)
        net.log["pairs"] += 1
        net.lock(nodea, ia).release()
        net.lock(nodeb, ib).release()


def findswapablequbits(net, node):
    trackers = net.enttrackers[node - 1]
    left = [(i, n) for i, n in enumerate(trackers, 1) if n is not None and n[0] < node and not net.islocked(node, i)]
    if not left:
        return None
    right = [(i, n) for i, n in enumerate(trackers, 1) if n is not None and n[0] > node and not net.islocked(node, i)]
    if not right:
        return None
    fl = min(left, key=lambda x: x[1][0])   # findmin: first minimal
    fr = max(right, key=lambda x: x[1][0])  # findmax: first maximal
    return fl[0], fr[0]


def swapper(net, node, wait_time, busy_time):
    sim = net.sim
    swapcircuit = EntanglementSwap()
    while True:
        pair = findswapablequbits(net, node)
        if pair is None:
            yield sim.timeout(wait_time)
            continue
        q1, q2 = pair
        yield net.lock(node, q1).request() & net.lock(node, q2).request()
        reg = net.reg(node)
        yield sim.timeout(busy_time)
        node1 = net.enttrackers[node - 1][q1 - 1]
        node2 = net.enttrackers[node - 1][q2 - 1]
        reg1, reg2 = net.reg(node1[0]), net.reg(node2[0])
        uptotime((reg[q1], reg1[node1[1]], reg[q2], reg2[node2[1]]), sim.now)
        swapcircuit(reg[q1], reg1[node1[1]], reg[q2], reg2[node2[1]])
        net.enttrackers[node1[0] - 1][node1[1] - 1] = node2
        net.enttrackers[node2[0] - 1][node2[1] - 1] = node1
        net.enttrackers[node - 1][q1 - 1] = None
        net.enttrackers[node - 1][q2 - 1] = None
        net.log["swaps"] += 1
        net.lock(node, q1).release()
        net.lock(node, q2).release()


def findqubitstopurify(net, nodea, nodeb):
    trackers = [(i, n) for i, n in enumerate(net.enttrackers[nodea - 1], 1)
                if n is not None and n[0] == nodeb and not net.islocked(nodea, i) and not net.islocked(nodeb, n[1])]
    if len(trackers) >= 2:
        (a1, n1), (a2, n2) = trackers[-2], trackers[-1]
        return a2, n2[1], a1, n1[1]
    return None


def _release_pair_tracking(net, node, slot):
    net.enttrackers[node - 1][slot - 1] = None


def purifier(net, nodea, nodeb, wait_time, busy_time):
    sim = net.sim
    rnd = 0
    while True:
        found = findqubitstopurify(net, nodea, nodeb)
        if found is None:
            yield sim.timeout(wait_time)
            continue
        p1a, p1b, p2a, p2b = found
        locks = [net.lock(nodea, p1a), net.lock(nodea, p2a), net.lock(nodeb, p1b), net.lock(nodeb, p2b)]
        ev = locks[0].request()
        for l in locks[1:]:
            ev = ev & l.request()
        yield ev
        yield sim.timeout(busy_time)
        rega, regb = net.reg(nodea), net.reg(nodeb)
        circuit = Purify2to1(("X", "Z")[rnd % 2])
        success = circuit(rega[p1a], regb[p1b], rega[p2a], regb[p2b])
        from ..circuits import _branches
        follow_success = bool(success) if _branches(success) else True     # masks mode: the ensemble shares the success branch
        if not follow_success:
            net.log["failures"] = net.log.get("failures", 0) + 1          # single-trajectory runs only (the twin of a lane)
            net.log.setdefault("first_failure_time", sim.now)
            _release_pair_tracking(net, nodea, p1a)
            _release_pair_tracking(net, nodeb, p1b)
        else:
            rnd += 1
        net.log["purifications"] += 1
        _release_pair_tracking(net, nodea, p2a)
        _release_pair_tracking(net, nodeb, p2b)
        for l in locks:
            l.release()


DEFAULTS = dict(sizes=(2, 3, 4, 3, 2), T2=100.0, F=0.97, t_end=25.0, dt_sample=0.2,
                entangler_wait_time=0.1, entangler_busy_time=1.0, swapper_wait_time=0.1, swapper_busy_time=0.15,
                purifier_wait_time=0.15, purifier_busy_time=0.2)


def monte_carlo_trajectory(make_repr, background=None, record=None, **kw):
    """monte_carlo_trajectory(...) -- 6.1_compare_formalisms_noplot.jl:10-57.

    make_repr() -> representation object for a slot.  record(kind, time, value) receives every sampled
    observable (BatchValue for an ensemble, number or `something` for one trajectory)."""
    p = dict(DEFAULTS); p.update(kw)
    bg = background if background is not None else T2Dephasing(p["T2"])
    net = Network(p["sizes"], make_repr, bg)
    sim = net.sim
    noisy_pair = p.get("noisy_pair") or stab_noisy_pair_func(p["F"])
    for a, b in net.edges():
        sim.process(entangler(net, a, b, noisy_pair, p["entangler_wait_time"], p["entangler_busy_time"]))
    for node in range(1, net.R + 1):
        sim.process(swapper(net, node, p["swapper_wait_time"], p["swapper_busy_time"]))
    for a in range(1, net.R + 1):
        for b in range(a + 1, net.R + 1):
            sim.process(purifier(net, a, b, p["purifier_wait_time"], p["purifier_busy_time"]))
    fidXX, fidZZ = [], []
    nsteps = int(round(p["t_end"] / p["dt_sample"]))
    ends = (net.reg(1)[2], net.reg(net.R)[2])
    for k in range(nsteps + 1):
        t = k * p["dt_sample"]
        sim.run(t)
        fxx = observable(ends, XX, something=0.0, time=t)
        fzz = observable(ends, ZZ, something=0.0, time=t)
        if record is not None:
            record("XX", t, fxx); record("ZZ", t, fzz)
        fidXX.append(fxx); fidZZ.append(fzz)
    return fidXX, fidZZ, net
