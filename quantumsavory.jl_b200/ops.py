"""Symbolic-level objects the register interface is driven with (the QuantumSymbolics surface that
QuantumSavory re-exports, /root/reference/src/QuantumSavory.jl:53-61), reduced to what the hot path
consumes: gates, Pauli-product observables, projectors, stabilizer / mixed / mixture states.

Every object knows how to lower itself for the dense representation (a numpy matrix / ket, what
`express(x, QuantumOpticsRepr())` returns) and, where defined, for the Clifford representation
(conjugation table / generators, what `express(x, CliffordRepr(), ...)` returns).
"""
from __future__ import annotations

import numpy as np

from . import symbols as S


class Gate:
    """A k-qubit operator.  `pauli` is set for tensor products of Paulis (fast observable path)."""

    def __init__(self, name, mat, pauli=None):
        self.name = name
        self.mat = np.asarray(mat, dtype=np.complex128)
        self.k = int(round(np.log2(self.mat.shape[0])))
        self.pauli = pauli
        self._table = None

    @property
    def table(self):
        """express(op, CliffordRepr(), UseAsOperation()); ValueError when not Clifford / k>2."""
        if self._table is None:
            if self.k > 2:
                raise ValueError("Clifford lowering is implemented for 1- and 2-qubit gates")
            self._table = S.clifford_table(self.mat)
        return self._table

    def __repr__(self):
        return f"Gate({self.name})"


def tensor(*ops):
    """a (x) b (x) ...: first factor acts on the first index (least significant)."""
    if all(isinstance(o, Gate) for o in ops):
        pauli = "".join(o.pauli for o in ops) if all(o.pauli is not None for o in ops) else None
        return Gate("⊗".join(o.name for o in ops), S.tensor(*[o.mat for o in ops]), pauli)
    if all(isinstance(o, StabilizerState) for o in ops):
        return StabilizerState.product(ops)
    return S.tensor(*[np.asarray(o) for o in ops])


I = Gate("I", S.Id, "I")
X = Gate("X", S.X, "X")
Y = Gate("Y", S.Y, "Y")
Z = Gate("Z", S.Z, "Z")
H = Gate("H", S.H)
CNOT = Gate("CNOT", S.CNOT)
CPHASE = Gate("CPHASE", S.CPHASE)
ZCX, ZCY, ZCZ = CNOT, Gate("ZCY", S.ZCY), CPHASE
XCX, XCY, XCZ = Gate("XCX", S.XCX), Gate("XCY", S.XCY), Gate("XCZ", S.XCZ)
YCX, YCY, YCZ = Gate("YCX", S.YCX), Gate("YCY", S.YCY), Gate("YCZ", S.YCZ)
σˣ, σʸ, σᶻ = X, Y, Z


class StabilizerState:
    """StabilizerState("XX ZZ"): a pure stabilizer state; lowers to a ket or to generators."""

    def __init__(self, spec: str):
        self.spec = spec
        self.xs, self.zs, self.signs = S.stabilizer_generators(spec)
        self.n = self.xs.shape[1]
        if self.xs.shape[0] != self.n:
            raise ValueError("a pure stabilizer state needs n generators on n qubits")
        self._ket = None

    @classmethod
    def product(cls, states):
        rows = []
        total = sum(s.n for s in states)
        off = 0
        for s in states:
            for g in range(s.n):
                row = ["I"] * total
                for q in range(s.n):
                    row[off + q] = {(0, 0): "I", (1, 0): "X", (0, 1): "Z", (1, 1): "Y"}[(int(s.xs[g, q]), int(s.zs[g, q]))]
                rows.append(("-" if s.signs[g] else "") + "".join(row))
            off += s.n
        return cls(" ".join(rows))

    @property
    def ket(self):
        if self._ket is None:
            self._ket = S.stabilizer_ket(self.spec)
        return self._ket

    @property
    def nsubsystems(self):
        return self.n


class MixedStabilizer:
    """MixedDestabilizer(S"ZZ"): fewer generators than qubits (test/general/observable_tests.jl:79-81)."""

    def __init__(self, spec: str, n: int | None = None):
        self.xs, self.zs, self.signs = S.stabilizer_generators(spec) if spec else (np.zeros((0, n), np.uint8),) * 2 + (np.zeros(0, np.uint8),)
        self.n = self.xs.shape[1] if n is None else n

    @property
    def nsubsystems(self):
        return self.n

    @property
    def dm(self):
        D = 2**self.n
        rho = np.eye(D, dtype=np.complex128) / D
        for g in range(self.xs.shape[0]):
            P = S._pauli_from_bits([(int(self.xs[g, q]), int(self.zs[g, q])) for q in range(self.n)]) * (-1) ** int(self.signs[g])
            rho = rho @ (np.eye(D) + P)
        return rho


class MixedState(MixedStabilizer):
    """MixedState(x): the maximally mixed state on x's Hilbert space, I/2^n."""

    def __init__(self, n: int):
        super().__init__("", n)


class Mixture:
    """sum_i w_i state_i (weights sum to 1), e.g. F*SProjector(bell) + (1-F)*MixedState(bell)
    (firstgenrepeater_lowlevel/clifford_setup.jl:10).  Dense lowering: the weighted density matrix.
    Clifford lowering: one term sampled per trajectory (one draw)."""

    def __init__(self, terms):
        self.terms = [(float(w), s) for w, s in terms]
        self.n = self.terms[0][1].nsubsystems

    @property
    def nsubsystems(self):
        return self.n

    @property
    def dm(self):
        rho = 0
        for w, s in self.terms:
            rho = rho + w * (s.dm if isinstance(s, MixedStabilizer) else S.dm(s.ket if isinstance(s, StabilizerState) else np.asarray(s)))
        return rho


BELL = StabilizerState("XX ZZ")  # stab_perfect_pair; its ket is (|00>+|11>)/sqrt2


def DepolarizedBellPair(p=None, F=None):
    """p|Phi+><Phi+| + (1-p) I/4, F = (3p+1)/4 -- src/StatesZoo/depolarized.jl:22-43."""
    if p is None:
        p = (4 * F - 1) / 3
    return Mixture([(p, StabilizerState("ZZ XX")), (1 - p, MixedState(2))])


class SProjector:
    """|psi><psi| used as an observable (projector(...) / SProjector(...))."""

    def __init__(self, state):
        self.state = state

    @property
    def ket(self):
        return self.state.ket if isinstance(self.state, StabilizerState) else np.asarray(self.state, dtype=np.complex128)


projector = SProjector


def ghz(k: int) -> StabilizerState:
    """QuantumClifford.ghz(k): XX..X, ZZI.., IZZ.., ..."""
    rows = ["X" * k] + ["I" * i + "ZZ" + "I" * (k - 2 - i) for i in range(k - 1)]
    return StabilizerState(" ".join(rows))


def nsubsystems(state) -> int:
    if hasattr(state, "nsubsystems"):
        return state.nsubsystems
    a = np.asarray(state)
    return int(round(np.log2(a.shape[0])))
