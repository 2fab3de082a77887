"""GPU: ufe_b200_calc_pc_truncation_error_noalloc (csrc/noalloc.cu) -- the truncation-error kernel with all of its device memory in a fixed
__device__ array inside the library, for callers that must not allocate -- against the existing entry point ufe_b200_calc_pc_truncation_error through
the C ABI, on calls in a row that share the array.  Bit-exact, like tests/test_pc_error.py: pointwise arithmetic, a maximum and two counts.

Collected last on purpose (the round's end runs pytest -x); the test builds its own handle, so it passes alone in a fresh process as well as after
the other tests.  No counterpart in the reference: the entry point was added at the request of the run's operators."""
import ctypes as C

import numpy as np
import pytest

import ufe_b200 as ub
import workloads as wl
from test_pc_error import _case

pytestmark = pytest.mark.gpu


def _existing(ctx, prm, case, optional=True):
    Hi_np1, Hi_star, grounded, gl_gr, frac, guilty = case
    tau = np.zeros(len(Hi_np1)); g = guilty.copy(); eta = C.c_double(); nt, ng = C.c_int32(), C.c_int32()
    ub.check(ub.lib().ufe_b200_calc_pc_truncation_error(ctx.h, *prm, ub.dp(Hi_np1), ub.dp(Hi_star), ub.ip(grounded), ub.ip(gl_gr), ub.dp(frac),
                                                        ub.dp(tau) if optional else None, ub.ip(g) if optional else None, C.byref(eta), C.byref(nt),
                                                        C.byref(ng)))
    return tau, g, eta.value, nt.value, ng.value


def _noalloc(ctx, prm, case, n=None, optional=True):
    Hi_np1, Hi_star, grounded, gl_gr, frac, guilty = case
    n = len(Hi_np1) if n is None else n
    tau = np.zeros(len(Hi_np1)); g = guilty.copy(); eta = C.c_double(); nt, ng = C.c_int32(), C.c_int32()
    ub.check(ub.lib().ufe_b200_calc_pc_truncation_error_noalloc(ctx.h, n, *prm, ub.dp(Hi_np1), ub.dp(Hi_star), ub.ip(grounded), ub.ip(gl_gr),
                                                                ub.dp(frac), ub.dp(tau) if optional else None, ub.ip(g) if optional else None,
                                                                C.byref(eta), C.byref(nt), C.byref(ng)))
    return tau, g, eta.value, nt.value, ng.value


def test_no_alloc_truncation_error_equals_the_existing_entry_point_on_calls_in_a_row():
    w = wl.mismip_mod(2.5e4)
    m = w.mesh
    assert m.nV <= ub.NOALLOC_MAX_VERTICES
    ctx = ub.Context(device=0)
    bare = ub.Context(device=0)            # a handle without a mesh: the new entry point needs none
    try:
        ctx.mesh_create(m); ctx.calc_matrix_operators()
        prm = (0.9, 0.2, 1e-8, 0.15)       # zeta_t, dt_n, C%pc_eta_min, C%pc_epsilon
        for call, seed in enumerate((8, 9)):          # two calls in a row on the shared array, different fields
            case = _case(m.nV, seed=seed)
            want = _existing(ctx, prm, case)
            got = _noalloc(bare if call else ctx, prm, case)
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), call
            assert got[2:] == want[2:] and want[4] > 0 and want[2] > prm[2], (call, got[2:], want[2:])
        # third call: nothing changed anywhere -> the maximum and the guilty count must not be the previous call's (cells re-initialised)
        Hi_np1, Hi_star, grounded, gl_gr, frac, guilty = case
        same = (Hi_np1, Hi_np1, grounded, gl_gr, frac, guilty)
        want = _existing(ctx, prm, same, optional=False)
        got = _noalloc(ctx, prm, same, optional=False)
        assert got[2:] == want[2:] == (prm[2], int((frac == 1.0).sum()), 0)
        # fourth call on a shorter range: the tail of the array still holds the longer call's fields and must not be counted
        k = m.nV // 3
        short = tuple(a[:k].copy() for a in _case(m.nV, seed=10))
        got = _noalloc(bare, prm, short)
        np.testing.assert_allclose(got[0], prm[0] * np.abs(short[0] - short[1]) / ((3.0 * prm[0] + 3.0) * prm[1]), rtol=1e-14)
        tau, full = got[0], short[4] == 1.0
        sel = (short[2] == 1) & (short[3] == 0) & full
        assert np.array_equal(got[1], short[5] + (full & (tau > prm[3])))
        assert got[2:] == (max(prm[2], tau[sel].max()), int(full.sum()), int((full & (tau > prm[3])).sum()))
        # n = 0 is an empty range, not an error
        assert _noalloc(bare, prm, short, n=0)[2:] == (prm[2], 0, 0)
    finally:
        bare.close()
        ctx.close()


def test_no_alloc_rejects_sizes_beyond_the_fixed_array():
    ctx = ub.Context(device=0)
    try:
        case = _case(16, seed=1)
        with pytest.raises(ub.UfeError, match="exceeds UFE_B200_NOALLOC_MAX_VERTICES = 262144"):
            _noalloc(ctx, (0.9, 0.2, 1e-8, 0.15), case, n=ub.NOALLOC_MAX_VERTICES + 1)
        with pytest.raises(ub.UfeError, match="negative"):
            _noalloc(ctx, (0.9, 0.2, 1e-8, 0.15), case, n=-1)
        # the handle and the array are still usable after a rejected call
        got = _noalloc(ctx, (0.9, 0.2, 1e-8, 0.15), case)
        assert got[3] == int((case[4] == 1.0).sum())
    finally:
        ctx.close()
