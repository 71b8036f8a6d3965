"""Independent numpy reference for the dense path: full-operator kron algebra, no index tricks.
Used to pin the oracle's index conventions (subsystem 0 least significant) from a second angle."""
import numpy as np

C = np.complex128


def embed(n, idx, op):
    """Full 2^n operator for `op` (factor m on subsystem idx[m], factor 0 least significant)."""
    k = len(idx)
    D = 2**n
    out = np.zeros((D, D), dtype=C)
    rest = [q for q in range(n) if q not in idx]
    for c in range(D):
        cj = sum(((c >> idx[m]) & 1) << m for m in range(k))
        for i in range(2**k):
            r = sum(((c >> q) & 1) << q for q in rest) + sum(((i >> m) & 1) << idx[m] for m in range(k))
            out[r, c] = op[i, cj]
    return out


def kron_all(factors):
    """tensor(f0, f1, ...) with f0 least significant."""
    out = np.array([[1]], dtype=C) if factors[0].ndim == 2 else np.array([1], dtype=C)
    for f in factors:
        out = np.kron(f, out)
    return out


def ptrace_dm(rho, n, q):
    t = rho.reshape([2] * (2 * n))  # axes: row bits n-1..0, col bits n-1..0
    ax_r = n - 1 - q
    ax_c = 2 * n - 1 - q
    t = np.trace(t, axis1=ax_r, axis2=ax_c)
    D = 2 ** (n - 1)
    return t.reshape(D, D)


def project_dm(rho, n, q, psi):
    """(<psi|_q (x) I) rho (|psi>_q (x) I), unnormalised, subsystem q removed."""
    t = rho.reshape([2] * (2 * n))
    ax_r = n - 1 - q
    ax_c = 2 * n - 1 - q
    t = np.tensordot(psi.conj(), t, axes=([0], [ax_r]))  # contracted axis removed, others shift
    t = np.tensordot(t, psi, axes=([ax_c - 1], [0]))
    D = 2 ** (n - 1)
    return t.reshape(D, D)
