"""CPU: host logic of the multigrid pattern phase (ufemism2.0_b200/csrc/mg.cu: aggregation, coarse sliced-ELL pattern,
Galerkin gather lists), checked against a dense P^T A P in numpy.  No GPU, no oracle involved."""
import ctypes as C

import numpy as np
import pytest

import ufe_b200 as ub


def sliced_ell(n, off, nbrs):
    """nbrs[t] = list of extended column indices (own node = off + t first). Returns slice_ptr, col in the solver's layout."""
    nsl = (n + 31) // 32
    sp = np.zeros(nsl + 1, np.int32)
    for s in range(nsl):
        sp[s + 1] = sp[s] + max(len(nbrs[t]) for t in range(s * 32, min(n, s * 32 + 32)))
    col = np.zeros(int(sp[-1]) * 32, np.int32)
    for s in range(nsl):
        for lane in range(32):
            t = s * 32 + lane
            for k in range(sp[s], sp[s + 1]):
                j = k - sp[s]
                if t < n:
                    col[k * 32 + lane] = nbrs[t][j] if j < len(nbrs[t]) else off + t
    return sp, col


def host_transfer(n, off, sp, col):
    L = ub.lib()
    nc, cs, ge = C.c_int(), C.c_longlong(), C.c_longlong()
    ub.check(L.ufe_b200_mg_host_transfer(n, off, ub.ip(sp), ub.ip(col), C.byref(nc), C.byref(cs), C.byref(ge)))
    agg = np.zeros(n, np.int32); csp = np.zeros((nc.value + 31) // 32 + 1, np.int32); ccol = np.zeros(cs.value, np.int32)
    gp = np.zeros(cs.value + 1, np.int32); gs = np.zeros(ge.value, np.int32)
    ub.check(L.ufe_b200_mg_host_transfer_get(ub.ip(agg), ub.ip(csp), ub.ip(ccol), ub.ip(gp), ub.ip(gs)))
    return nc.value, agg, csp, ccol, gp, gs


def grid_pattern(nx, ny, rng, halo=0):
    """5-point + one random extra neighbour per node on an nx x ny grid; `halo` extra extended columns on each side."""
    n = nx * ny
    off = halo
    nbrs = []
    for t in range(n):
        i, j = divmod(t, ny)
        l = [off + t]
        for di, dj in ((1, 0), (-1, 0), (0, 1), (0, -1)):
            a, b = i + di, j + dj
            if 0 <= a < nx and 0 <= b < ny:
                l.append(off + a * ny + b)
        extra = int(rng.integers(0, n))
        if off + extra not in l:
            l.append(off + extra)
        if halo and (i == 0 or i == nx - 1):           # couplings to halo columns, dropped by the preconditioner
            l.append(int(rng.integers(0, halo)) if i == 0 else off + n + int(rng.integers(0, halo)))
        nbrs.append(l)
    return n, off, nbrs


def dense_from_ell(n, off, sp, col, val, n_ext):
    """val[(k*4 + q)*32 + lane]; returns the dense (2n x 2n_ext) matrix, duplicates added."""
    A = np.zeros((2 * n, 2 * n_ext))
    for t in range(n):
        s, lane = divmod(t, 32)
        for k in range(sp[s], sp[s + 1]):
            c = col[k * 32 + lane]
            v = [val[(k * 4 + q) * 32 + lane] for q in range(4)]
            A[2 * t, 2 * c] += v[0]; A[2 * t, 2 * c + 1] += v[1]; A[2 * t + 1, 2 * c] += v[2]; A[2 * t + 1, 2 * c + 1] += v[3]
    return A


@pytest.mark.parametrize("halo", [0, 7])
def test_transfer_is_galerkin_product(halo):
    rng = np.random.default_rng(5)
    n, off, nbrs = grid_pattern(23, 17, rng, halo)
    sp, col = sliced_ell(n, off, nbrs)
    nc, agg, csp, ccol, gp, gs = host_transfer(n, off, sp, col)
    # a partition into fewer, non-empty aggregates
    assert 0 < nc < n / 2
    assert agg.min() == 0 and agg.max() == nc - 1 and len(np.unique(agg)) == nc
    # values: random in the real slots, zero in the padding
    n_slots = int(sp[-1]) * 32
    val = np.zeros(n_slots * 4)
    for t in range(n):
        s, lane = divmod(t, 32)
        for j in range(len(nbrs[t])):
            for q in range(4):
                val[((sp[s] + j) * 4 + q) * 32 + lane] = rng.standard_normal()
    # the device kernel's sum, restated: coarse slot value = sum of its listed fine slots
    cval = np.zeros(len(ccol) * 4)
    for cs in range(len(ccol)):
        for fs in gs[gp[cs]:gp[cs + 1]]:
            for q in range(4):
                cval[((cs >> 5) * 4 + q) * 32 + (cs & 31)] += val[((fs >> 5) * 4 + q) * 32 + (fs & 31)]
    # every in-range fine slot is gathered exactly once
    in_range = sum(1 for t in range(n) for k in range(sp[t >> 5], sp[(t >> 5) + 1]) if 0 <= col[k * 32 + (t & 31)] - off < n)
    assert len(gs) == in_range and len(np.unique(gs)) == len(gs)
    A = dense_from_ell(n, off, sp, col, val, n + 2 * halo)[:, 2 * off:2 * (off + n)]      # owned columns only
    P = np.zeros((2 * n, 2 * nc))
    for t in range(n):
        P[2 * t, 2 * agg[t]] = 1.0; P[2 * t + 1, 2 * agg[t] + 1] = 1.0
    Ac = dense_from_ell(nc, 0, csp, ccol, cval, nc)
    np.testing.assert_allclose(Ac, P.T @ A @ P, rtol=1e-12, atol=1e-12)
    # coarse rows start with their own node (the diagonal block) and padding points at the own node
    for I in range(nc):
        assert ccol[csp[I >> 5] * 32 + (I & 31)] == I


def test_isolated_nodes_are_grouped():
    """identity rows nobody references (e.g. prescribed velocities in an ice-free corner) must not stall the coarsening"""
    n = 100
    nbrs = [[t] for t in range(n)]
    sp, col = sliced_ell(n, 0, nbrs)
    nc, agg, *_ = host_transfer(n, 0, sp, col)
    assert nc == 13 and np.bincount(agg).max() == 8


@pytest.mark.parametrize("nparts", [2, 3])
def test_split_rows_assemble_the_global_galerkin_product(nparts):
    """rows split over ranks (emulated): every part coarsens its own block, the coarse level is assembled from all parts and the sum of the
    parts' Galerkin gathers (= the all-reduce on the device) is P^T A P of the WHOLE matrix, couplings across the cuts included"""
    rng = np.random.default_rng(11)
    n, _off, nbrs = grid_pattern(29, 13, rng, halo=0)      # global ids
    cuts = [0] + sorted(int(x) for x in rng.choice(np.arange(40, n - 40), nparts - 1, replace=False)) + [n]
    sps, cols, sp_start, col_start, ns, offs = [], [], [], [], [], []
    for p in range(nparts):
        a, b = cuts[p], cuts[p + 1]
        sp, col = sliced_ell(b - a, a, nbrs[a:b])
        sp_start.append(sum(len(x) for x in sps)); col_start.append(sum(len(x) for x in cols))
        sps.append(sp); cols.append(col); ns.append(b - a); offs.append(a)
    i32 = lambda x: np.ascontiguousarray(x, np.int32)
    SP, COL = i32(np.concatenate(sps)), i32(np.concatenate(cols))
    L = ub.lib()
    nc, cs, ge = C.c_int(), C.c_longlong(), C.c_longlong()
    ub.check(L.ufe_b200_mg_host_transfer_parts(nparts, n, ub.ip(i32(ns)), ub.ip(i32(offs)), ub.ip(i32(sp_start)), ub.ip(i32(col_start)), ub.ip(SP),
                                               ub.ip(COL), C.byref(nc), C.byref(cs), C.byref(ge)))
    agg = np.zeros(n, np.int32); csp = np.zeros((nc.value + 31) // 32 + 1, np.int32); ccol = np.zeros(cs.value, np.int32)
    gp = np.zeros(nparts * (cs.value + 1), np.int32); gs = np.zeros(ge.value, np.int32); gstart = np.zeros(nparts + 1, np.int32)
    ub.check(L.ufe_b200_mg_host_transfer_parts_get(ub.ip(agg), ub.ip(csp), ub.ip(ccol), ub.ip(gp), ub.ip(gs), ub.ip(gstart)))
    assert agg.min() == 0 and agg.max() == nc.value - 1
    # aggregates never cross a cut and are numbered part after part
    for p in range(nparts - 1):
        assert agg[cuts[p]:cuts[p + 1]].max() < agg[cuts[p + 1]:cuts[p + 2]].min()
    # per-part values, per-part gathers, then the sum over parts
    A = np.zeros((2 * n, 2 * n)); cval = np.zeros(len(ccol) * 4)
    for p in range(nparts):
        a, b = cuts[p], cuts[p + 1]
        sp, col = sps[p], cols[p]
        val = np.zeros(int(sp[-1]) * 32 * 4)
        for t in range(b - a):
            s, lane = divmod(t, 32)
            for j in range(len(nbrs[a + t])):
                for q in range(4):
                    val[((sp[s] + j) * 4 + q) * 32 + lane] = rng.standard_normal()
        A[2 * a:2 * b, :] = dense_from_ell(b - a, a, sp, col, val, n)
        gpp = gp[p * (cs.value + 1):(p + 1) * (cs.value + 1)]; gsp = gs[gstart[p]:gstart[p + 1]]
        for c_ in range(len(ccol)):
            for fs in gsp[gpp[c_]:gpp[c_ + 1]]:
                for q in range(4):
                    cval[((c_ >> 5) * 4 + q) * 32 + (c_ & 31)] += val[((fs >> 5) * 4 + q) * 32 + (fs & 31)]
    P = np.zeros((2 * n, 2 * nc.value))
    for t in range(n):
        P[2 * t, 2 * agg[t]] = 1.0; P[2 * t + 1, 2 * agg[t] + 1] = 1.0
    Ac = dense_from_ell(nc.value, 0, csp, ccol, cval, nc.value)
    np.testing.assert_allclose(Ac, P.T @ A @ P, rtol=1e-12, atol=1e-12)


def test_chebyshev_steps_model_of_the_coarsest_level_kernel():
    """numpy model of k_mg_coarsest_cluster's Chebyshev mode (mg.cu): x_0 = 0, d_0 = Dw b / theta, x_{k+1} = x_k + d_k,
    d_k = rho_k rho_{k-1} d_{k-1} + (2 rho_k / delta) Dw (b - A x_k) on [lmax / ratio, lmax], lmax = 1.05 x the infinity-norm bound of Dw A.
    The map b -> x is linear (a fixed polynomial of Dw A: a valid stationary preconditioner), and 8 steps leave about the error of 32 damped
    Jacobi sweeps (and much less than 8 of them) on a diagonally dominant, slightly non-symmetric matrix."""
    rng = np.random.default_rng(21)
    n = 400
    # a 2-D Laplacian-like matrix with a weak shift and a non-symmetric perturbation (what the coarse SSA operators look like after Jacobi scaling)
    g = int(np.sqrt(n)); A = np.zeros((n, n))
    for i in range(g):
        for j in range(g):
            r = i * g + j
            A[r, r] = 4.0 + 0.02
            for di, dj in ((1, 0), (-1, 0), (0, 1), (0, -1)):
                ii, jj = i + di, j + dj
                if 0 <= ii < g and 0 <= jj < g:
                    A[r, ii * g + jj] = -1.0 + 0.05 * rng.standard_normal()
    Dw = 0.8 / np.diag(A)

    def cheb(b, steps, ratio=30.0):
        hi = 1.05 * np.abs(Dw[:, None] * A).sum(axis=1).max(); lo = hi / ratio
        theta, delta = 0.5 * (hi + lo), 0.5 * (hi - lo)
        sigma = theta / delta; rho_old = 1.0 / sigma
        d = Dw * b / theta; x = d.copy()
        for _ in range(steps):
            rho = 1.0 / (2.0 * sigma - rho_old)
            d = rho * rho_old * d + (2.0 * rho / delta) * (Dw * (b - A @ x))
            x = x + d
            rho_old = rho
        return x

    def jacobi(b, sweeps):
        x = Dw * b
        for _ in range(sweeps):
            x = x + Dw * (b - A @ x)
        return x

    b1, b2 = rng.standard_normal(n), rng.standard_normal(n)
    np.testing.assert_allclose(cheb(2.0 * b1 - 3.0 * b2, 8), 2.0 * cheb(b1, 8) - 3.0 * cheb(b2, 8), rtol=1e-12, atol=1e-12)   # linear in b
    x_ex = np.linalg.solve(A, b1)
    e_c = np.linalg.norm(cheb(b1, 8) - x_ex) / np.linalg.norm(x_ex)
    e_j = np.linalg.norm(jacobi(b1, 32) - x_ex) / np.linalg.norm(x_ex)
    e_j8 = np.linalg.norm(jacobi(b1, 8) - x_ex) / np.linalg.norm(x_ex)
    assert e_c < 1.1 * e_j < 1.0 and e_c < 0.85 * e_j8, (e_c, e_j, e_j8)      # 8 steps do what 32 sweeps do (the device probe says the same)
    assert np.linalg.norm(cheb(b1, 16) - x_ex) < np.linalg.norm(cheb(b1, 8) - x_ex)        # more steps, smaller error
