"""CPU: the oracle's restatement of timeevolution.master / schroedinger / mcwf (oracle/qsb_oracle.c, adaptive DP5)
against the exact propagator (scipy.linalg.expm of the embedded Liouvillian) and against itself (trajectory mean ==
master equation).  Pins the checker that tests/test_gpu_evolve.py compares the CUDA path with."""
import numpy as np
import pytest
from scipy.linalg import expm

from oracle import oracle as orc
from qsb200 import symbols as S


def embed(n, idx, op):
    """operator on subsystems idx (0-based, factor 0 least significant) of an n-qubit space"""
    k = len(idx)
    D = 2 ** n
    out = np.zeros((D, D), complex)
    rest = [q for q in range(n) if q not in idx]
    for c in range(D):
        cj = sum(((c >> q) & 1) << m for m, q in enumerate(idx))
        for i in range(2 ** k):
            r = c
            for m, q in enumerate(idx):
                r = (r & ~(1 << q)) | (((i >> m) & 1) << q)
            out[r, c] = op[i, cj]
    return out


def exact_master(rho, n, idx, H, lind, t):
    D = 2 ** n
    Hf = embed(n, idx, H) if H is not None else np.zeros((D, D), complex)
    Ls = [embed(n, [idx[p]], L) for p, L in lind]
    I = np.eye(D)
    # column-stacking: vec(A rho B) = kron(B.T, A) vec(rho)
    Lv = -1j * (np.kron(I, Hf) - np.kron(Hf.T, I))
    for L in Ls:
        G = L.conj().T @ L
        Lv += np.kron(L.conj(), L) - 0.5 * np.kron(I, G) - 0.5 * np.kron(G.T, I)
    v = expm(Lv * t) @ rho.reshape(-1, order="F")
    return v.reshape(D, D, order="F")


def rand_herm(rng, d):
    a = rng.normal(size=(d, d)) + 1j * rng.normal(size=(d, d))
    return (a + a.conj().T) / 2


def rand_dm(rng, n):
    a = rng.normal(size=(2 ** n, 2 ** n)) + 1j * rng.normal(size=(2 ** n, 2 ** n))
    rho = a @ a.conj().T
    return rho / np.trace(rho)


@pytest.mark.parametrize("k", [1, 2])
def test_master_matches_exact_propagator(k):
    rng = np.random.default_rng(7 + k)
    n = 3
    for trial in range(4):
        idx = list(rng.permutation(n)[:k])
        H = rand_herm(rng, 2 ** k)
        bgs = [S.T1T2Noise(1.3, 0.9), S.T2Dephasing(0.7)][:k]
        lind = [(p, L) for p, bg in enumerate(bgs) for L in S.lindbladop(bg)]
        rho = rand_dm(rng, n)
        t = float(rng.uniform(0.05, 1.5))
        got = orc.Dense.from_array(rho).evolve_master(idx, H, lind, t).array()
        want = exact_master(rho, n, idx, H, lind, t)
        assert np.abs(got - want).max() < 2e-6          # the solver's own tolerance (reltol 1e-6, abstol 1e-8)
        assert abs(np.trace(got) - 1) < 1e-6


def test_ket_is_promoted_and_no_collapse_is_unitary():
    rng = np.random.default_rng(3)
    psi = rng.normal(size=4) + 1j * rng.normal(size=4)
    psi /= np.linalg.norm(psi)
    H = rand_herm(rng, 2)
    d = orc.Dense.from_array(psi).evolve_master([1], H, [], 0.8)
    assert d.kind == 1                                   # apply_noninstant!(dm(state), ...), noninstant.jl:21-23
    U = embed(2, [1], expm(-1j * H * 0.8))
    assert np.abs(d.array() - U @ np.outer(psi, psi.conj()) @ U.conj().T).max() < 2e-6
    k = orc.Dense.from_array(psi)
    assert k.evolve_mcwf([1], H, [], 0.8, 1, 0, 0) == 0 and k.kind == 0   # schroedinger
    assert np.abs(k.array() - U @ psi).max() < 2e-6


def test_mcwf_trajectory_mean_is_the_master_equation():
    """the reference's own statistical check for MCKet noise (quantummc_repr_tests.jl:99-118 does it for Kraus jumps)"""
    H = 0.7 * S.X + 0.2 * S.Z
    lind = [(0, L) for L in S.lindbladop(S.T1T2Noise(1.0, 0.8))]
    t, N = 0.9, 1500
    want = exact_master(S.dm(S.X1), 1, [0], H, lind, t)
    acc = np.zeros((2, 2), complex)
    jumps = 0
    for traj in range(N):
        k = orc.Dense.from_array(S.X1)
        jumps += k.evolve_mcwf([0], H, lind, t, 0x77, traj, 5)
        v = k.array()
        assert abs(np.linalg.norm(v) - 1) < 1e-9
        acc += np.outer(v, v.conj())
    assert jumps > 0.2 * N
    assert np.abs(acc / N - want).max() < 4 / np.sqrt(N)
