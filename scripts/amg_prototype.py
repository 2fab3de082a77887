"""Prototype (CPU, scipy): aggregation multigrid as the preconditioner of BiCGStab on the DIVA stiffness matrix.
Design study for the device preconditioner; not part of the product or of the tests."""
import os
import sys
import time

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ufemism2.0_b200")); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, ROOT)
import bench  # noqa
import oracle_lib as ol  # noqa
import ufe_b200 as ub  # noqa


def get_system(nv, nvisc=3):
    w = bench.make_workload(nv)
    m, cfg = w.mesh, w.cfg
    cfg.ksp_initial_guess_nonzero = 0
    o = ol.OracleMesh(m)
    o.calc_matrix_operators()
    st, A = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    ice = w.ice_inputs()
    t0 = time.time()
    rc, nvi, na = o.solve_velocity(cfg, 1, ice, st, max_visc_its=nvisc)
    print(f"oracle: nV={m.nV} {nvi} visc its {na} Krylov its {time.time()-t0:.1f}s", flush=True)
    ptr, ind, val, bb = o.assemble(cfg, A["N_b"], A["dN_dx_b"], A["dN_dy_b"], A["beta_eff_b"], A["tau_dx_b"], A["tau_dy_b"])
    n = len(ptr) - 1
    M = sp.csr_matrix((val, ind - 1, ptr - 1), shape=(n, n))
    M.sum_duplicates()
    return w, M, bb


def aggregate_graph(G, min_size=1):
    """Vanek-style greedy aggregation on the graph G (csr, symmetric pattern). Returns agg id per node, count."""
    n = G.shape[0]
    ptr, ind = G.indptr, G.indices
    agg = -np.ones(n, np.int64)
    na = 0
    # pass 1: roots whose whole neighbourhood is free
    for i in range(n):
        if agg[i] >= 0:
            continue
        nb = ind[ptr[i]:ptr[i + 1]]
        if np.all(agg[nb] < 0):
            agg[i] = na
            agg[nb] = na
            na += 1
    # pass 2: attach leftovers to a neighbouring aggregate (the one of the first aggregated neighbour)
    agg2 = agg.copy()
    for i in range(n):
        if agg[i] >= 0:
            continue
        nb = ind[ptr[i]:ptr[i + 1]]
        a = agg[nb]
        a = a[a >= 0]
        if len(a):
            # most common neighbour aggregate
            vals, cnt = np.unique(a, return_counts=True)
            agg2[i] = vals[np.argmax(cnt)]
    agg = agg2
    # pass 3: whatever is left forms new aggregates with free neighbours
    for i in range(n):
        if agg[i] >= 0:
            continue
        agg[i] = na
        nb = ind[ptr[i]:ptr[i + 1]]
        for j in nb:
            if agg[j] < 0:
                agg[j] = na
        na += 1
    return agg, na


class Level:
    pass


def build_hierarchy(A, tri_graph, max_levels=10, coarse_size=400, fine_graph="matrix", omega=0.7, smooth_P=False):
    """A: 2nT x 2nT with interleaved (u,v) unknowns per node.  Aggregation on nodes."""
    levels = []
    G = tri_graph
    while True:
        L = Level()
        L.A = A.tocsr()
        nT = A.shape[0] // 2
        # 2x2 block-diagonal inverse (block Jacobi)
        d = L.A.diagonal()
        L.Dinv = 1.0 / d
        Ab = sp.bsr_matrix(L.A, blocksize=(2, 2))
        Ab.sort_indices()
        # extract diagonal blocks
        D = np.zeros((nT, 2, 2))
        for i in range(nT):
            for k in range(Ab.indptr[i], Ab.indptr[i + 1]):
                if Ab.indices[k] == i:
                    D[i] = Ab.data[k]
        L.Dblk_inv = np.linalg.inv(D)
        levels.append(L)
        if nT <= coarse_size or len(levels) >= max_levels:
            L.lu = spla.splu(sp.csc_matrix(L.A))
            break
        if G is None:
            # node graph from the block matrix pattern
            G = sp.csr_matrix((np.ones(len(Ab.indices)), Ab.indices, Ab.indptr), shape=(nT, nT))
            G = (G + G.T).tocsr()
            G.setdiag(0); G.eliminate_zeros()
        agg, na = aggregate_graph(G)
        # P: piecewise constant per component
        rows = np.arange(2 * nT)
        cols = 2 * np.repeat(agg, 2) + np.tile([0, 1], nT)
        P = sp.csr_matrix((np.ones(2 * nT), (rows, cols)), shape=(2 * nT, 2 * na))
        if smooth_P:
            P = P - omega * sp.diags(L.Dinv) @ (L.A @ P)
        L.P = P
        L.R = P.T.tocsr()
        A = (L.R @ L.A @ L.P).tocsr()
        print(f"  level {len(levels)-1}: nodes {nT} -> {na} (ratio {nT/na:.1f}), nnz/row {L.A.nnz/L.A.shape[0]:.1f}", flush=True)
        G = None
    print(f"  coarsest: {levels[-1].A.shape[0]} rows, nnz/row {levels[-1].A.nnz/levels[-1].A.shape[0]:.1f}")
    print("  operator complexity", sum(l.A.nnz for l in levels) / levels[0].A.nnz)
    return levels


def smooth(L, x, b, nsweeps, omega, block=True):
    for _ in range(nsweeps):
        r = b - L.A @ x
        if block:
            rr = r.reshape(-1, 2)
            x = x + omega * np.einsum("nij,nj->ni", L.Dblk_inv, rr).reshape(-1)
        else:
            x = x + omega * L.Dinv * r
    return x


def cycle(levels, k, b, nu1, nu2, omega, gamma=1, overcorr=1.0):
    L = levels[k]
    if k == len(levels) - 1:
        return L.lu.solve(b)
    x = smooth(L, np.zeros_like(b), b, nu1, omega)
    r = b - L.A @ x
    rc = L.R @ r
    ec = cycle(levels, k + 1, rc, nu1, nu2, omega, gamma, overcorr)
    if gamma == 2 and k + 1 < len(levels) - 1:   # W-cycle: second visit with residual update
        rc2 = rc - levels[k + 1].A @ ec
        ec = ec + cycle(levels, k + 1, rc2, nu1, nu2, omega, gamma, overcorr)
    x = x + overcorr * (L.P @ ec)
    x = smooth(L, x, b, nu2, omega)
    return x


def bicgstab(A, b, prec, rtol, abstol, maxit, x0=None):
    """PETSc KSPBCGS-style left-preconditioned BiCGStab; returns x, iterations (each iteration = 2 matvec + 2 prec)."""
    n = len(b)
    x = np.zeros(n) if x0 is None else x0.copy()
    r = prec(b - A @ x)
    bn = np.linalg.norm(prec(b))
    tol = max(rtol * bn, abstol)
    rp = r.copy()
    rho = alpha = omega = 1.0
    p = np.zeros(n); v = np.zeros(n)
    if np.linalg.norm(r) <= tol:
        return x, 0
    for it in range(1, maxit + 1):
        rho_new = rp @ r
        beta = (rho_new / rho) * (alpha / omega)
        p = r + beta * (p - omega * v)
        v = prec(A @ p)
        alpha = rho_new / (rp @ v)
        s = r - alpha * v
        t = prec(A @ s)
        omega = (t @ s) / (t @ t)
        x = x + alpha * p + omega * s
        r = s - omega * t
        rho = rho_new
        if np.linalg.norm(r) <= tol:
            return x, it
    return x, maxit


if __name__ == "__main__":
    nv = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
    w, A, b = get_system(nv)
    m = w.mesh
    n = A.shape[0]
    print("system", n, "nnz", A.nnz)
    d = A.diagonal()
    t0 = time.time()
    x_j, it_j = bicgstab(A, b, lambda r: r / d, 1e-4, 1e-3, 4000)
    print(f"Jacobi-BiCGStab rtol 1e-4: {it_j} its ({time.time()-t0:.1f}s)")
    x_ref = spla.spsolve(sp.csc_matrix(A), b)
    print("  rel err vs direct", np.linalg.norm(x_j - x_ref) / np.linalg.norm(x_ref))
    # triangle face-adjacency graph
    TriC = m.TriC - 1
    rows = np.repeat(np.arange(m.nTri), 3)
    cols = TriC.reshape(-1)
    ok = cols >= 0
    Gt = sp.csr_matrix((np.ones(ok.sum()), (rows[ok], cols[ok])), shape=(m.nTri, m.nTri))
    Gt = ((Gt + Gt.T) > 0).astype(np.float64).tocsr()
    for name, g0, smoothP in (("matrix-graph", None, False), ("TriC-graph", Gt, False)):
        print(f"== hierarchy: {name} smoothP={smoothP}")
        t0 = time.time()
        lv = build_hierarchy(A, g0, smooth_P=smoothP)
        print(f"  setup {time.time()-t0:.1f}s")
        for (nu1, nu2, om, gamma, oc) in ((1, 1, 0.7, 1, 1.0), (2, 2, 0.7, 1, 1.0), (2, 2, 0.7, 1, 1.5), (1, 1, 0.7, 2, 1.0), (2, 2, 0.7, 2, 1.0)):
            prec = lambda r: cycle(lv, 0, r, nu1, nu2, om, gamma, oc)
            t0 = time.time()
            x, it = bicgstab(A, b, prec, 1e-4, 1e-30, 500)
            err = np.linalg.norm(x - x_ref) / np.linalg.norm(x_ref)
            x8, it8 = bicgstab(A, b, prec, 1e-8, 1e-30, 500)
            print(f"  nu=({nu1},{nu2}) omega={om} gamma={gamma} overcorr={oc}: its(1e-4)={it} err={err:.2e}  its(1e-8)={it8}  ({time.time()-t0:.1f}s)", flush=True)
