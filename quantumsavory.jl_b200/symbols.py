"""Host-side constants for the register-state path: basis kets, gates, observables, backgrounds.

Mirrors the symbols the reference's hot path is driven with (QuantumSymbolics names re-exported
by QuantumSavory, /root/reference/src/QuantumSavory.jl:53-61) and the dense constants of the
QuantumOptics adapter (/root/reference/src/backends/quantumoptics/uptotime.jl:66-86).

Conventions ("Conventions"):
  * |0> = Z1 = index 0, |1> = Z2 = index 1.
  * Multi-qubit matrices are column-major friendly numpy arrays M[row, col] whose own index has
    factor 0 (the *first* operand of the tensor product / first entry of `indices`) as the
    least-significant bit, i.e. tensor(a, b) = kron(b, a).
  * Two-qubit "ACB" gates: P+^A (x) I + P-^A (x) B with the FIRST index the control
    (CNOT = ZCX, CPHASE = ZCZ), uptotime.jl:83-84.
"""
from __future__ import annotations

import dataclasses
import math

import numpy as np

SQ2 = math.sqrt(2.0)
C = np.complex128

# ---- kets ---------------------------------------------------------------------------------
Z1 = np.array([1, 0], dtype=C)
Z2 = np.array([0, 1], dtype=C)
X1 = np.array([1, 1], dtype=C) / SQ2
X2 = np.array([1, -1], dtype=C) / SQ2
Y1 = np.array([1, 1j], dtype=C) / SQ2
Y2 = np.array([1, -1j], dtype=C) / SQ2
L0, L1 = Z1, Z2

# ---- single-qubit operators -----------------------------------------------------------------
Id = np.eye(2, dtype=C)
X = np.array([[0, 1], [1, 0]], dtype=C)
Y = np.array([[0, -1j], [1j, 0]], dtype=C)
Z = np.array([[1, 0], [0, -1]], dtype=C)
H = (Z + X) / SQ2  # _hadamard = (sigmaz+sigmax)/sqrt2, uptotime.jl:82
S = np.array([[1, 0], [0, 1j]], dtype=C)  # _phase
Pp = {"X": np.outer(X1, X1.conj()), "Y": np.outer(Y1, Y1.conj()), "Z": np.outer(Z1, Z1.conj())}
Pm = {"X": np.outer(X2, X2.conj()), "Y": np.outer(Y2, Y2.conj()), "Z": np.outer(Z2, Z2.conj())}
PAULI = {"I": Id, "X": X, "Y": Y, "Z": Z}


def tensor(*ops: np.ndarray) -> np.ndarray:
    """tensor(a, b, ...) with the first factor least significant: kron(..., b, a)."""
    out = np.array([[1]], dtype=C) if ops[0].ndim == 2 else np.array([1], dtype=C)
    for op in ops:
        out = np.kron(op, out)
    return out


def controlled(a: str, b: str) -> np.ndarray:
    """`aCb` gate: P+^a (x) I + P-^a (x) b, first factor (control) least significant."""
    return tensor(Pp[a], Id) + tensor(Pm[a], PAULI[b])


CNOT = controlled("Z", "X")
CPHASE = controlled("Z", "Z")
ZCX, ZCY, ZCZ = CNOT, controlled("Z", "Y"), CPHASE
XCX, XCY, XCZ = controlled("X", "X"), controlled("X", "Y"), controlled("X", "Z")
YCX, YCY, YCZ = controlled("Y", "X"), controlled("Y", "Y"), controlled("Y", "Z")

NAMED_GATES = {
    "X": X, "Y": Y, "Z": Z, "H": H, "S": S, "I": Id,
    "CNOT": CNOT, "CPHASE": CPHASE, "ZCX": ZCX, "ZCY": ZCY, "ZCZ": ZCZ,
    "XCX": XCX, "XCY": XCY, "XCZ": XCZ, "YCX": YCX, "YCY": YCY, "YCZ": YCZ,
}

# eigvecs(X|Y|Z) ordered (+1, -1): outcome 1 <-> +1 eigenstate (quantumoptics/express.jl:7-12)
EIGVECS = {"X": (X1, X2), "Y": (Y1, Y2), "Z": (Z1, Z2)}

BELL = (tensor(Z1, Z1) + tensor(Z2, Z2)) / SQ2  # perfect_pair, firstgenrepeater_lowlevel/setup.jl:64


def ghz_ket(k: int) -> np.ndarray:
    v = np.zeros(2**k, dtype=C)
    v[0] = v[-1] = 1 / SQ2
    return v


def dm(psi: np.ndarray) -> np.ndarray:
    return np.outer(psi, psi.conj())


def depolarized_bell_dm(p: float) -> np.ndarray:
    """DepolarizedBellPair(p) in QuantumOpticsRepr: p|Phi+><Phi+| + (1-p) I/4
    (/root/reference/src/StatesZoo/depolarized.jl:33-37; F = (3p+1)/4, :26)."""
    return p * dm(BELL) + (1 - p) * np.eye(4, dtype=C) / 4


# ---- backgrounds (src/backgrounds.jl) -------------------------------------------------------
BG_T1, BG_T2, BG_DEPOL, BG_T1T2 = 0, 1, 2, 3


@dataclasses.dataclass(frozen=True)
class T1Decay:
    t1: float = 1e9
    kind = BG_T1

    def __post_init__(self):
        if not self.t1 > 0:
            raise ValueError("t1 > 0")

    @property
    def params(self):
        return (float(self.t1), 0.0)


@dataclasses.dataclass(frozen=True)
class T2Dephasing:
    t2: float = 1e9
    kind = BG_T2

    def __post_init__(self):
        if not self.t2 > 0:
            raise ValueError("t2 > 0")

    @property
    def params(self):
        return (float(self.t2), 0.0)


@dataclasses.dataclass(frozen=True)
class Depolarization:
    tau: float = 1e9
    kind = BG_DEPOL

    def __post_init__(self):
        if not self.tau > 0:
            raise ValueError("tau > 0")

    @property
    def params(self):
        return (float(self.tau), 0.0)


@dataclasses.dataclass(frozen=True)
class T1T2Noise:
    t1: float = 1e9
    t2: float = 1e9
    kind = BG_T1T2

    def __post_init__(self):
        if not (self.t1 > 0 and self.t2 > 0):
            raise ValueError("t1, t2 > 0")

    @property
    def params(self):
        return (float(self.t1), float(self.t2))


_LH = np.array([[0, 1], [0, 0]], dtype=C)  # sigmap = |0><1| (uptotime.jl:72)


def lindbladop(bg):
    """Collapse operators of a background -- src/backends/quantumoptics/noninstant.jl:44-97."""
    if isinstance(bg, LindbladOnly):
        return lindbladop(bg.noise)
    if isinstance(bg, T1Decay):
        return [_LH / math.sqrt(bg.t1)]                      # 1/sqrt(T1) |0><1|
    if isinstance(bg, T2Dephasing):
        return [Z / math.sqrt(2 * bg.t2)]                    # 1/sqrt(2 T2) Z
    if isinstance(bg, T1T2Noise):
        tphi_inv = 1 / bg.t2 - 1 / (2 * bg.t1)
        if tphi_inv <= 0:                                    # no pure dephasing
            return [_LH / math.sqrt(bg.t1)]
        return [_LH / math.sqrt(bg.t1), Z / math.sqrt(2 / tphi_inv)]
    if isinstance(bg, Depolarization):
        raise RuntimeError("we do not have lindblad operators implemented for Depolarization")
    raise TypeError(f"no lindbladop method for {bg!r}")


@dataclasses.dataclass(frozen=True)
class LindbladOnly:
    """A background whose `krausops` is `nothing`: uptotime! falls back to the master equation / mcwf with
    `lindbladop(noise)` (src/backends/quantumoptics/uptotime.jl:24-29,51-54 -- in the reference the qumode
    AmplitudeDamping and PauliNoise take that route; this wrapper sends a qubit background down it)."""
    noise: object
    kind = -1

    @property
    def params(self):
        return self.noise.params


# ---- Clifford lowering ------------------------------------------------------------------------
_PAULI_BITS = {(0, 0): Id, (1, 0): X, (0, 1): Z, (1, 1): Y}


def _pauli_from_bits(bits):
    """Hermitian Pauli string from [(x0,z0),(x1,z1),...], factor 0 least significant."""
    return tensor(*[_PAULI_BITS[b] for b in bits])


def clifford_table(U: np.ndarray) -> np.ndarray:
    """Conjugation table of a 1- or 2-qubit Clifford unitary.

    entry t = x0 | z0<<1 | x1<<2 | z1<<3  ->  x0' | z0'<<1 | x1'<<2 | z1'<<3 | signflip<<(2k)
    such that U P(t) U^dagger = (-1)^signflip P(t').  Raises ValueError when U is not Clifford.
    This is the lowering express(op, CliffordRepr(), UseAsOperation()) performs symbolically in the
    reference (src/representations.jl:4-7); here it is derived numerically from the same unitary the
    dense path uses, so the two representations cannot disagree about a gate's meaning.
    """
    U = np.asarray(U, dtype=C)
    k = {2: 1, 4: 2}[U.shape[0]]
    table = np.zeros(4**k, dtype=np.uint8)
    for t in range(4**k):
        bits = [((t >> (2 * m)) & 1, (t >> (2 * m + 1)) & 1) for m in range(k)]
        img = U @ _pauli_from_bits(bits) @ U.conj().T
        found = False
        for t2 in range(4**k):
            bits2 = [((t2 >> (2 * m)) & 1, (t2 >> (2 * m + 1)) & 1) for m in range(k)]
            P2 = _pauli_from_bits(bits2)
            for sgn in (0, 1):
                if np.allclose(img, (-1) ** sgn * P2, atol=1e-9):
                    table[t] = t2 | (sgn << (2 * k))
                    found = True
        if not found:
            raise ValueError("operator is not a Clifford gate")
    return table


def clifford_linear_form(table: np.ndarray):
    """Split a conjugation table into its GF(2)-linear part and sign lookup.

    Returns (cols, lut): cols[j] is the image bit-vector of input basis bit j (bit j of t), so that
    t' = XOR_{j in t} cols[j]; lut bit t is the sign flip for input t.  The bit-sliced GPU kernels
    consume this form (one XOR network + one 4/16-entry truth table).
    """
    k2 = int(round(math.log2(len(table))))
    k = k2 // 2
    mask = (1 << k2) - 1
    cols = [int(table[1 << j]) & mask for j in range(k2)]
    lut = 0
    for t in range(len(table)):
        img = 0
        for j in range(k2):
            if (t >> j) & 1:
                img ^= cols[j]
        assert img == (int(table[t]) & mask), "table is not linear"
        if (int(table[t]) >> k2) & 1:
            lut |= 1 << t
    return cols, lut


def stabilizer_generators(spec: str):
    """Parse 'XX ZZ' / '-XI IZ' / 'X_Z _XZ ZZX' into (xs, zs, signs) byte arrays, qubit 0 first.

    StabilizerState("...") strings as used throughout the reference tests
    (e.g. test/general/observable_tests.jl:45,110)."""
    rows = spec.split()
    n = len(rows[0].lstrip("+-"))
    xs = np.zeros((len(rows), n), dtype=np.uint8)
    zs = np.zeros((len(rows), n), dtype=np.uint8)
    signs = np.zeros(len(rows), dtype=np.uint8)
    for g, row in enumerate(rows):
        if row[0] in "+-":
            signs[g] = row[0] == "-"
            row = row[1:]
        assert len(row) == n
        for q, ch in enumerate(row):
            xs[g, q] = ch in "XY"
            zs[g, q] = ch in "ZY"
    return xs, zs, signs


def stabilizer_ket(spec: str) -> np.ndarray:
    """Dense ket of a pure stabilizer state (the +1 joint eigenvector), global phase fixed so the
    first non-zero amplitude is real positive.  Used to compare the Clifford and dense paths the way
    test/general/observable_tests.jl:154-155 does (`Ket(clifford_state) ~ qo_state`)."""
    xs, zs, signs = stabilizer_generators(spec)
    n = xs.shape[1]
    proj = np.eye(2**n, dtype=C)
    for g in range(xs.shape[0]):
        P = _pauli_from_bits([(int(xs[g, q]), int(zs[g, q])) for q in range(n)]) * (-1) ** int(signs[g])
        proj = proj @ (np.eye(2**n, dtype=C) + P) / 2
    w, v = np.linalg.eigh((proj + proj.conj().T) / 2)
    psi = v[:, -1]
    assert abs(w[-1] - 1) < 1e-9 and (xs.shape[0] < n or abs(w[-2]) < 1e-9)
    k = int(np.argmax(np.abs(psi) > 1e-9))
    return psi * (abs(psi[k]) / psi[k])
