"""GPU: the optimised CPU baseline (oracle/qsb_bitsliced.cpp: the exported op records of a recorded tableau program replayed on
the host with the bit-sliced tableau core, OpenMP over 32-trajectory words) against the GPU executor on the same global
trajectory ids -- integer tallies, so every statistics row must be identical.  Ragged lane count, both trajectory offsets."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_bitsliced_cpu_replay_equals_the_gpu_program():
    import torch
    from oracle import bitsliced
    from qsb200 import symbols as S
    from qsb200.backend import BatchContext, CudaBatchRepr
    from qsb200.registers import dispose, observable
    from qsb200.sharding import StatsBuffer
    from tests.test_gpu_program import schedule
    ntraj, seed = 5000 - 7, 4242
    ctx = BatchContext(ntraj, seed=seed)
    stats = StatsBuffer(4, "cuda:0")

    def sink(i, refs, obs, t):
        observable(refs, obs, something=0.0, time=t, stats_dev=stats.slot_ptr(i), sync=False)
    with ctx.record() as rec:
        dispose(schedule(CudaBatchRepr(ctx, "clifford"), S.Depolarization(3.0), sink))
    ex = rec.program.export()
    assert ex["meta"][2] == rec.program.info()["ops"] and ex["meta"][6] == 4
    for shift in (0, 777 * 32):
        rec.program.run(traj_shift=shift); ctx.sync()
        gpu = stats.t[:4].cpu().numpy()
        cpu, threads = bitsliced.run(ex, seed, shift, ntraj)
        assert threads >= 1
        assert np.array_equal(gpu[:, :3], cpu[:, :3]), (shift, gpu, cpu)
        assert 0 < cpu[1, 0] <= ntraj
    rec.program.free(); ctx.close()
