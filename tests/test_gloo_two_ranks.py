"""CPU, two processes over torch.distributed (gloo): the host-side model of the multi-GPU data path.
The device library talks NCCL directly; what is checked here, without a GPU, are the properties that path relies on and the collectives' use:
row blocks from partition_list, halos that are contiguous index ranges owned by the adjacent rank, y = A x on row blocks after a halo
exchange, all-reduced dot products, and the multigrid level-1 right-hand side completed by one all-reduce of per-rank restrictions."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    try:
        sys.path.insert(0, os.path.join(ROOT, "ufemism2.0_b200")); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, ROOT)
        import torch
        import torch.distributed as dist
        import oracle_lib as ol
        import ufe_b200 as ub
        import ufe_host as uh
        import workloads as wl
        from test_mg_host import sliced_ell
        import ctypes as C
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
        w = wl.mismip_mod(6.0e4)
        m = w.mesh
        o = ol.OracleMesh(m)
        rng = np.random.default_rng(5)          # same stream on both ranks
        other = 1 - rank
        checked = 0
        for name in ("map_a_b", "ddx_b_a", "ddy_a_a", "ddx_b_b"):
            ptr, ind, val, nrow, ncol = o.get_operator(ub.OP[name])
            ind = ind - 1 if ind.min() == 1 else ind
            r1, r2 = uh.partition_list(nrow, rank, world); c1, c2 = uh.partition_list(ncol, rank, world)      # 1-based inclusive
            o1, o2 = uh.partition_list(ncol, other, world)
            rows = slice(r1 - 1, r2)
            cols_needed = np.unique(ind[ptr[r1 - 1]:ptr[r2]])
            lo, hi = int(cols_needed.min()), int(cols_needed.max())
            halo = cols_needed[(cols_needed < c1 - 1) | (cols_needed > c2 - 1)]
            # the halo lies in the adjacent rank's block, next to the cut
            assert halo.size > 0 and ((halo >= o1 - 1) & (halo <= o2 - 1)).all()
            # halo exchange: tell the neighbour which contiguous range is wanted, receive exactly that
            x = rng.standard_normal(ncol)
            x_ext = np.full(hi - lo + 1, np.nan)
            own_lo, own_hi = max(lo, c1 - 1), min(hi, c2 - 1)
            x_ext[own_lo - lo:own_hi - lo + 1] = x[own_lo:own_hi + 1]
            want = (lo, c1 - 2) if rank == 1 else (c2, hi)                      # rank 1 needs a left halo, rank 0 a right halo
            wants = [None, None]
            dist.all_gather_object(wants, want)
            a, b = wants[other]
            send = torch.from_numpy(np.ascontiguousarray(x[a:b + 1]))
            assert a >= c1 - 1 and b <= c2 - 1                                   # the neighbour asks only for entries this rank owns
            recv = torch.empty(want[1] - want[0] + 1, dtype=torch.float64)
            reqs = [dist.isend(send, other), dist.irecv(recv, other)]
            for r_ in reqs:
                r_.wait()
            x_ext[want[0] - lo:want[1] - lo + 1] = recv.numpy()
            assert not np.isnan(x_ext[np.unique(ind[ptr[r1 - 1]:ptr[r2]]) - lo]).any()
            # y = A x on the row block with extended-local columns (the layout DevCsr uses) against the global product
            y = np.zeros(r2 - r1 + 1)
            for i in range(r1 - 1, r2):
                k = slice(ptr[i], ptr[i + 1])
                y[i - (r1 - 1)] = np.dot(val[k], x_ext[ind[k] - lo])
            y_glob = np.array([np.dot(val[ptr[i]:ptr[i + 1]], x[ind[ptr[i]:ptr[i + 1]]]) for i in range(nrow)])
            np.testing.assert_allclose(y, y_glob[rows], rtol=0, atol=0)
            # fused reductions: per-rank partial sums, one all-reduce
            part = torch.tensor([float(np.dot(y, y)), float(np.sum(y))], dtype=torch.float64)
            dist.all_reduce(part, op=dist.ReduceOp.SUM)
            assert abs(part[0].item() - float(np.dot(y_glob, y_glob))) <= 1e-12 * float(np.dot(y_glob, y_glob))
            checked += 1
        # multigrid: every rank coarsens its own rows (the emulated-parts entry of the library, same code as the device path), restricts its
        # residual into its slice of the level-1 right-hand side, one all-reduce completes it
        nb = m.nTri
        TriC = np.asarray(m.TriC) - 1
        nbrs = [[t] + [int(j) for j in TriC[t] if j >= 0] for t in range(nb)]
        cuts = [0, uh.partition_list(nb, 0, world)[1], nb]
        sps, cols, sp_start, col_start = [], [], [], []
        for p in range(world):
            sp, col = sliced_ell(cuts[p + 1] - cuts[p], cuts[p], nbrs[cuts[p]:cuts[p + 1]])
            sp_start.append(sum(len(x_) for x_ in sps)); col_start.append(sum(len(x_) for x_ in cols))
            sps.append(sp); cols.append(col)
        i32 = lambda a_: np.ascontiguousarray(a_, np.int32)
        L = ub.lib()
        nc, cs, ge = C.c_int(), C.c_longlong(), C.c_longlong()
        ub.check(L.ufe_b200_mg_host_transfer_parts(world, nb, ub.ip(i32([cuts[1] - cuts[0], cuts[2] - cuts[1]])), ub.ip(i32(cuts[:2])), ub.ip(i32(sp_start)),
                                                   ub.ip(i32(col_start)), ub.ip(i32(np.concatenate(sps))), ub.ip(i32(np.concatenate(cols))), C.byref(nc),
                                                   C.byref(cs), C.byref(ge)))
        agg = np.zeros(nb, np.int32); csp = np.zeros((nc.value + 31) // 32 + 1, np.int32); ccol = np.zeros(cs.value, np.int32)
        gp = np.zeros(world * (cs.value + 1), np.int32); gs = np.zeros(ge.value, np.int32); gstart = np.zeros(world + 1, np.int32)
        ub.check(L.ufe_b200_mg_host_transfer_parts_get(ub.ip(agg), ub.ip(csp), ub.ip(ccol), ub.ip(gp), ub.ip(gs), ub.ip(gstart)))
        res = rng.standard_normal((nb, 2))
        mine = slice(cuts[rank], cuts[rank + 1])
        assert agg[mine].min() >= (0 if rank == 0 else agg[:cuts[1]].max() + 1)            # aggregates do not cross the cut
        bc = np.zeros((nc.value, 2))
        np.add.at(bc, agg[mine], res[mine])
        t = torch.from_numpy(bc)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        full = np.zeros((nc.value, 2))
        np.add.at(full, agg, res)
        np.testing.assert_allclose(t.numpy(), full, rtol=1e-13, atol=1e-13)
        # copy boundary conditions ('infinite_SSA_icestream' / 'periodic_ISMIP-HOM'): a copy row reads the previous solution around a point that
        # usually lies in another rank's block, so every rank writes its own slice into a zero global vector and one all-reduce completes it
        # (velocity.cu: copy_bc_rhs); the load vector of the owned copy rows then equals the single-process assembly
        t1, t2 = uh.partition_list(m.nTri, rank, world)
        up = 100.0 * rng.standard_normal(m.nTri); vp = 100.0 * rng.standard_normal(m.nTri)
        glob = np.zeros(2 * m.nTri)
        glob[t1 - 1:t2] = up[t1 - 1:t2]; glob[m.nTri + t1 - 1:m.nTri + t2] = vp[t1 - 1:t2]
        tg = torch.from_numpy(glob)
        dist.all_reduce(tg, op=dist.ReduceOp.SUM)
        assert np.array_equal(tg.numpy(), np.concatenate([up, vp]))
        cfg = wl.scaling_config(); cfg.visc_it_relax = 0.3; cfg.refgeo_idealised_ISMIP_HOM_L = 1.0e6
        for k, (bu, bv) in enumerate([(2, 2), (3, 2), (0, 0), (3, 1)]):
            cfg.BC_u[k] = bu; cfg.BC_v[k] = bv
        coef = [1e10 * (1 + 0.1 * rng.random(m.nTri))] + [rng.standard_normal(m.nTri) for _ in range(5)]
        bb = o.assemble(cfg, *coef, u_b_prev=up, v_b_prev=vp)[3]
        side = {1: 0, 2: 0, 3: 1, 4: 1, 5: 2, 6: 2, 7: 3, 8: 3}
        n_copy = n_remote = 0
        for ti in range(t1, t2 + 1):
            if m.TriBI[ti - 1] == 0:
                continue
            for uv in range(2):
                choice = (cfg.BC_u if uv == 0 else cfg.BC_v)[side[int(m.TriBI[ti - 1])]]
                if choice < 2:
                    continue
                tj, wt = o.find_ti_copy(choice, 1.0e6, ti)
                src = tg.numpy()[uv * m.nTri:(uv + 1) * m.nTri]
                fixed = 0.0
                for n_ in range(len(tj)):
                    if tj[n_]:
                        fixed = fixed + wt[n_] * src[tj[n_] - 1]
                val_ = 0.3 * fixed + 0.7 * src[ti - 1]
                assert abs(val_ - bb[2 * (ti - 1) + uv]) <= 1e-12 * max(1.0, abs(val_))
                n_copy += 1
                n_remote += int(((tj[tj > 0] < t1) | (tj[tj > 0] > t2)).any())
        assert n_copy > 20 and n_remote > 0
        checked += 1
        # merged reductions of the multi-rank BiCGStab (mg.cu: k_ksmg_dot5 / k_ksmg_update_xr5): the second half-iteration all-reduces the five inner
        # products of s and t at once and forms the next rho and ||r||^2 from them -- 2 all-reduces per iteration instead of 3, same iterates
        import scipy.sparse as sp
        cfg2 = wl.scaling_config()
        coef2 = [1e10 * (1 + 0.1 * rng.random(m.nTri)), 1e3 * rng.standard_normal(m.nTri), 1e3 * rng.standard_normal(m.nTri), 1e3 * rng.random(m.nTri),
                 1e4 * rng.standard_normal(m.nTri), 1e4 * rng.standard_normal(m.nTri)]
        aptr, aind, aval, ab = o.assemble(cfg2, *coef2)[:4]
        A = sp.csr_matrix((aval, aind - 1, aptr - 1), shape=(len(aptr) - 1, len(aptr) - 1))
        n_rows = A.shape[0]
        q1, q2 = uh.partition_list(n_rows, rank, world)
        A_loc = A[q1 - 1:q2]; b_loc = ab[q1 - 1:q2]; dinv = 1.0 / A.diagonal()[q1 - 1:q2]
        calls = [0]

        def allsum(vals):
            tt_ = torch.tensor(vals, dtype=torch.float64)
            dist.all_reduce(tt_, op=dist.ReduceOp.SUM)
            calls[0] += 1
            return tt_.numpy()

        def gather(x_loc):                       # stands in for the halo exchange checked above
            parts = [None] * world
            dist.all_gather_object(parts, x_loc)
            return np.concatenate(parts)

        def bicgstab(merged, rtol=1e-10, maxit=3000):
            x = np.zeros(q2 - q1 + 1); r = dinv * b_loc; rp = r.copy(); p_ = np.zeros_like(r); v = np.zeros_like(r)
            rr0, = allsum([r @ r]); rho = rr = rr0
            alpha = omega = 1.0; rho_prev = 1.0; per_it = []
            for it in range(maxit):
                if np.sqrt(rr) <= rtol * np.sqrt(rr0):
                    return gather(x), it, per_it
                c0 = calls[0]
                p_ = r.copy() if it == 0 else r + (rho / rho_prev) * (alpha / omega) * (p_ - omega * v)
                v = dinv * (A_loc @ gather(p_))
                d1, = allsum([v @ rp])
                alpha = rho / d1
                s_ = r - alpha * v
                t_ = dinv * (A_loc @ gather(s_))
                if merged:
                    st, tt, ss, srp, trp = allsum([s_ @ t_, t_ @ t_, s_ @ s_, s_ @ rp, t_ @ rp])
                    omega = st / tt
                    rho_prev, rho, rr = rho, srp - omega * trp, max(0.0, ss - 2.0 * omega * st + omega * omega * tt)
                else:
                    st, tt = allsum([s_ @ t_, t_ @ t_])
                    omega = st / tt
                x = x + alpha * p_ + omega * s_
                r = s_ - omega * t_
                if not merged:
                    rho_prev = rho
                    rho, rr = allsum([r @ rp, r @ r])
                per_it.append(calls[0] - c0)
            return gather(x), maxit, per_it

        x_d, its_d, per_d = bicgstab(False)
        x_m, its_m, per_m = bicgstab(True)
        assert set(per_d) == {3} and set(per_m) == {2}, (set(per_d), set(per_m))
        assert 10 < its_d < 3000 and abs(its_m - its_d) <= max(3, its_d // 10), (its_d, its_m)
        x_ref = sp.linalg.spsolve(A.tocsc(), ab)
        for x_ in (x_d, x_m):
            assert np.linalg.norm(x_ - x_ref) <= 1e-7 * np.linalg.norm(x_ref), np.linalg.norm(x_ - x_ref) / np.linalg.norm(x_ref)
        checked += 1
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok", checked))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "error", traceback.format_exc() + str(e)))


def test_two_ranks_over_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, status, info in results:
        assert status == "ok", f"rank {rank}: {info}"
        assert info == 6


@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_halos_are_contiguous_and_adjacent(nranks):
    """the assumption behind halo_exchange (ncclSend/ncclRecv with the left and right neighbour only): on the x-sorted meshes every operator's
    columns outside a rank's block lie in the blocks of ranks r-1 and r+1"""
    sys.path.insert(0, os.path.join(ROOT, "ufemism2.0_b200"))
    import oracle_lib as ol
    import ufe_b200 as ub
    import ufe_host as uh
    import workloads as wl
    w = wl.mismip_mod(3.0e4)
    o = ol.OracleMesh(w.mesh)
    for name in ("map_a_b", "ddx_b_a", "ddy_a_a", "ddx_b_b", "ddx_a_b"):
        ptr, ind, val, nrow, ncol = o.get_operator(ub.OP[name])
        ind = ind - 1 if ind.min() == 1 else ind
        for r in range(nranks):
            r1, r2 = uh.partition_list(nrow, r, nranks)
            need = np.unique(ind[ptr[r1 - 1]:ptr[r2]])
            lo = uh.partition_list(ncol, max(r - 1, 0), nranks)[0] - 1
            hi = uh.partition_list(ncol, min(r + 1, nranks - 1), nranks)[1] - 1
            assert need.min() >= lo and need.max() <= hi, (name, r)
