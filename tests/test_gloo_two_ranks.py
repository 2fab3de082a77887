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
        assert info == 4


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
