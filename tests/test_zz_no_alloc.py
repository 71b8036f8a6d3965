"""GPU: qsb_expect_pauli_static, the Pauli observable for callers that are not allowed to allocate GPU memory.

Who needs it: a host whose HBM is already handed to another allocator (CUDA.jl's pool, torch's caching allocator) -- an
allocation this library makes behind its back can fail or trigger a pool trim that stalls the device -- and loops with a
latency bound per call, where qsb_expect's grow-on-demand scratch buffer would be freed and re-allocated mid-loop.  The
variant keeps its reduction partials in a 4 MiB device array compiled into libqsb200.so instead.

What is checked, all through the C-ABI (ctypes), against qsb_expect on the same state:
  - per-trajectory values and the {n_alive, sum re, sum re^2, sum im} row agree to 1e-12 (the split count, hence the
    summation order, may differ from qsb_expect's, so this is not a bit-for-bit comparison);
  - two calls IN A ROW with different shapes both agree: the array keeps the first call's partials, and the second call
    must not read them;
  - the call works on a context that never ran qsb_expect (it does not lean on the context's own scratch buffer) and does
    not lower the device's free memory;
  - the documented limit: QSB_STATIC_MAX_TRAJ trajectories work, one more is refused with a message naming the array.

This file sorts last on purpose (pytest -x stops at the first failure: a failure here must not hide any other test) and
passes both alone in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-12
STATIC_MAX_TRAJ = 260096     # QSB_STATIC_MAX_TRAJ in include/qsb200.h


def _pauli(ctx, st, idx0, codes, sign, static, check_mem=False):
    """One Pauli observable through either entry point -> (per-trajectory complex values, statistics row).
    check_mem: the device's free memory, read right around the library call, must not go down."""
    import torch
    from qsb200 import _lib
    npad = (ctx.ntraj + 31) // 32 * 32
    per = torch.full((2 * npad,), float("nan"), dtype=torch.float64, device=f"cuda:{ctx.device}")
    stats = torch.full((4,), float("nan"), dtype=torch.float64, device=f"cuda:{ctx.device}")
    torch.cuda.synchronize()                                   # the fills ran on torch's stream, the library uses its own
    idx = (C.c_int * len(idx0))(*idx0)
    cod = np.asarray(codes, np.uint8)
    free_before = torch.cuda.mem_get_info()[0]
    if static:
        _lib.check(ctx.L.qsb_expect_pauli_static(st.p, idx, len(idx0), cod.ctypes.data_as(C.POINTER(C.c_uint8)), sign,
                                                 C.c_void_p(per.data_ptr()), C.c_void_p(stats.data_ptr())))
    else:
        _lib.check(ctx.L.qsb_expect(st.p, idx, len(idx0), _lib.OBS_PAULI, cod.ctypes.data_as(C.c_void_p), sign,
                                    C.c_void_p(per.data_ptr()), C.c_void_p(stats.data_ptr()), None))
    ctx.sync()
    if check_mem:
        assert torch.cuda.mem_get_info()[0] >= free_before, "the call took memory from the device"
    return per.cpu().numpy().view(np.complex128)[: ctx.ntraj].copy(), stats.cpu().numpy().copy()


def _close(a, b):
    (va, sa), (vb, sb) = a, b
    assert np.all(np.isfinite(va.view(np.float64))) and np.all(np.isfinite(sa))
    assert np.max(np.abs(va - vb)) <= TOL
    assert np.max(np.abs(sa - sb) / np.maximum(1.0, np.abs(sb))) <= TOL


def _varied_state(ctx, kind, n, rng):
    """n-qubit state whose trajectories differ: a random (n+1)-qubit ket, one qubit measured away (one draw per lane)."""
    from qsb200 import ops
    from qsb200 import symbols as S
    from qsb200._lib import DM, KET
    from qsb200.backend import CudaBatchState
    psi = rng.normal(size=2 ** (n + 1)) + 1j * rng.normal(size=2 ** (n + 1))
    st = CudaBatchState.from_array(ctx, psi / np.linalg.norm(psi), KET)
    st.project_traceout(n + 1, ops.X)
    if kind == DM:
        st.uptotime(2, S.Depolarization(3.0), 0.7)             # promotes the ket to a density matrix and mixes it
        st.uptotime(1, S.T1Decay(2.0), 0.4)
    return st


def test_static_variant_matches_qsb_expect_on_two_calls_in_a_row():
    from qsb200._lib import DM, KET
    from qsb200.backend import BatchContext
    rng = np.random.default_rng(0x2A11)
    ctx = BatchContext(97, seed=0x57A7)                        # three full words + one lane
    try:
        ket = _varied_state(ctx, KET, 6, rng)
        rho = _varied_state(ctx, DM, 5, rng)
        dead = ctx.random_mask(0.2)
        rho.kill(dead)                                         # dead lanes must read 0 and stay out of the statistics
        # first the static variant, on a context whose own scratch buffer was never created ...
        got1 = _pauli(ctx, ket, [4, 0, 2], [2, 1, 3], 0, static=True)        # Y5 X1 Z3 on the ket: 64 splits x 128 lanes of partials
        got2 = _pauli(ctx, rho, [1, 3], [3, 2], 1, static=True)              # -Z2 Y4 on the dm, right after: 32 splits, same array
        got3 = _pauli(ctx, rho, [1, 3], [3, 2], 1, static=True, check_mem=True)   # again: same numbers bit for bit, no memory taken
        assert np.array_equal(got2[0], got3[0]) and np.array_equal(got2[1], got3[1])
        # ... then the allocating entry point on the same states
        _close(got1, _pauli(ctx, ket, [4, 0, 2], [2, 1, 3], 0, static=False))
        _close(got2, _pauli(ctx, rho, [1, 3], [3, 2], 1, static=False))
        assert got2[1][0] == 97 - dead.count()
        assert np.abs(got1[0]).max() > 1e-3 and np.ptp(got1[0].real) > 1e-3  # the comparison is not 0 == 0
        # and once more in the other order, now that qsb_expect has left its own partials behind
        _close(_pauli(ctx, ket, [5], [1], 0, static=True), _pauli(ctx, ket, [5], [1], 0, static=False))
        ket.free(); rho.free()
    finally:
        ctx.close()


def test_static_variant_size_limit_and_refusals():
    from qsb200._lib import ArgumentError, BoundsError, KET, MethodError, TAB
    from qsb200.backend import BatchContext, CudaBatchState
    # the largest ensemble the built-in array serves (one split, partials + block statistics fill it almost exactly)
    ctx = BatchContext(STATIC_MAX_TRAJ, seed=3)
    try:
        st = CudaBatchState.from_array(ctx, np.array([0.6, 0.8j, 0.0, 0.0]), KET)      # <Y1> = 0.96, <Z2> = 1
        a = _pauli(ctx, st, [0, 1], [2, 3], 0, static=True)
        _close(a, _pauli(ctx, st, [0, 1], [2, 3], 0, static=False))
        assert abs(a[1][1] / a[1][0] - 0.96) < 1e-12
        st.free()
    finally:
        ctx.close()
    ctx = BatchContext(STATIC_MAX_TRAJ + 1, seed=3)
    try:
        st = CudaBatchState.zeros(ctx, KET, 1)
        with pytest.raises(ArgumentError, match="built-in array"):
            _pauli(ctx, st, [0], [3], 0, static=True)
        _pauli(ctx, st, [0], [3], 0, static=False)              # the allocating entry point takes the same ensemble
        st.free()
    finally:
        ctx.close()
    ctx = BatchContext(40, seed=3)
    try:
        st = CudaBatchState.zeros(ctx, KET, 2)
        with pytest.raises(BoundsError):
            _pauli(ctx, st, [0, 2], [1, 1], 0, static=True)     # subsystem 3 of 2
        with pytest.raises(ArgumentError):
            _pauli(ctx, st, [0], [4], 0, static=True)           # not a Pauli code
        tab = CudaBatchState.zeros(ctx, TAB, 2)
        with pytest.raises(MethodError):
            _pauli(ctx, tab, [0], [3], 0, static=True)          # tableau expectations use no scratch: qsb_expect serves them
        st.free(); tab.free()
    finally:
        ctx.close()
