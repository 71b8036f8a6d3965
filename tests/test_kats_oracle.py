"""The reference's hot-path known-answer tests (tests/kat_suite.py) run on the CPU oracle through the
same register layer + circuits the CUDA backend is driven with.  This pins the oracle AND the host
logic (registers.py, circuits.py, ops.py) on every commit without a GPU."""
import numpy as np
import pytest

from tests import kat_suite
from tests.oracle_backend import OracleContext, OracleRepr


class OracleEnv:
    ntraj = 1
    ode_atol = 1e-6   # the oracle restates the reference's adaptive DP5 (reltol 1e-6, abstol 1e-8)

    def __init__(self, seed, traj):
        self.ctx = OracleContext(seed, traj)

    def repr(self, formalism):
        return OracleRepr(self.ctx, formalism)

    def values(self, x):
        return np.atleast_1d(np.real(complex(x)))

    def outcomes(self, j):
        return np.atleast_1d(int(j))

    def truth(self, b):
        return np.atleast_1d(bool(b))

    def sync(self):
        pass


@pytest.mark.parametrize("fn", kat_suite.ALL, ids=lambda f: f.__name__)
@pytest.mark.parametrize("traj", [0, 1, 2])
def test_kat(fn, traj):
    fn(OracleEnv(0x5153, traj))
