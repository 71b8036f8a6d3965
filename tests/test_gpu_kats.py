"""GPU: the reference's hot-path known-answer tests (tests/kat_suite.py) on the CUDA batched backend,
called through the C-ABI, 100 trajectories in lockstep (not a multiple of 32: exercises pad lanes).
Every assertion must hold on every trajectory although their measurement outcomes differ."""
import numpy as np
import pytest

from tests import kat_suite

pytestmark = pytest.mark.gpu


class CudaEnv:
    def __init__(self, ntraj, seed=0x5153):
        from qsb200.backend import BatchContext
        self.ctx = BatchContext(ntraj, seed=seed)
        self.ntraj = ntraj

    def repr(self, formalism):
        from qsb200.backend import CudaBatchRepr
        return CudaBatchRepr(self.ctx, formalism)

    def values(self, x):
        from qsb200.backend import BatchValue
        if isinstance(x, BatchValue):
            assert x.n_alive == self.ntraj
            # mean and variance pin every trajectory: all equal <=> var == 0
            m, v = x.mean, x.var
            return np.array([m - np.sqrt(max(v, 0.0)), m, m + np.sqrt(max(v, 0.0))])
        return np.atleast_1d(np.real(complex(x)))

    def outcomes(self, j):
        return j.numpy()

    def truth(self, b):
        return b.numpy()

    def sync(self):
        self.ctx.sync()


@pytest.fixture(scope="module")
def env():
    e = CudaEnv(100)
    yield e
    e.ctx.close()


@pytest.mark.parametrize("fn", kat_suite.ALL, ids=lambda f: f.__name__)
def test_kat(fn, env):
    fn(env)
    env.ctx.sync()


def test_single_trajectory_returns_plain_ints():
    """ntraj == 1 keeps the reference's scalar API: outcome converts to Int, masks to Bool."""
    from qsb200 import ops
    from qsb200.registers import Register, initialize, project_traceout
    e = CudaEnv(1)
    a = Register(2, e.repr("clifford"))
    initialize(a[1:2], ops.StabilizerState("XX ZZ"))
    m1 = project_traceout(a[1], ops.Y)
    m2 = project_traceout(a[2], ops.Y)
    assert int(m1) in (1, 2) and int(m1) != int(m2)
    assert bool(m1 == 2) == (int(m1) == 2)
    e.ctx.close()
