"""CPU tests of the multi-GPU host logic (helfem_b200/shard.py): the family / unit partitions and the
world_size-2 all-reduce of partial operators over gloo.  The per-rank partial operators here are the
ORACLE's (the reference's family loop restricted to the rank's families, rebuilt from the oracle's
linearity in the stored integrals): rank r zeroes the integrals of the families it does not own, which
is what hfb200_set_shard does on the device.  The CUDA path of the same partition is test_shard_gpu.py."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def test_partitions_cover_exactly_once():
    from helfem_b200 import shard
    for nranks in (1, 2, 3, 4, 8):
        for nfam in (1, 5, 17, 49):
            seen = sum((shard.owned_families(nfam, r, nranks) for r in range(nranks)), [])
            assert sorted(seen) == list(range(nfam))
        for n in (0, 1, 7, 1024, 1025):
            sl = [shard.partition_units(n, r, nranks) for r in range(nranks)]
            assert sl[0][0] == 0 and sl[-1][1] == n
            assert all(sl[i][1] == sl[i + 1][0] for i in range(nranks - 1))
            sizes = [e - b for b, e in sl]
            assert max(sizes) - min(sizes) <= 1


def test_library_family_map_matches_the_python_rule():
    """hfb200_set_shard's map (host/famshard.h, shared by the host classes and the C ABI) is the rule the oracle-side tests use."""
    from helfem_b200 import shard
    for nranks in (1, 2, 3, 8):
        for nfam in (1, 5, 49, 585):
            own = shard.family_owners(nfam, nranks)
            assert own.tolist() == [shard.family_owner(r, nranks) for r in range(nfam)]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _atomic_oracle_partial(rank, nranks):
    """Oracle J/K of a small atomic case with the integrals of unowned families zeroed."""
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    import oracle
    from oracle import fem_oracle as fo
    from helfem_b200 import shard
    from util import splitmix64
    primbas, nnodes, nelem, lmax = 4, 6, 3, 1
    lval = np.array([0, 1, 1, 1], dtype=np.int32)
    mval = np.array([0, -1, 0, 1], dtype=np.int32)
    rs = fo.RadialSetup(primbas, nnodes, fo.get_grid(40.0, nelem), taylor_order=-1)
    nL = 2 * lmax + 1
    dL, dm1L, tei, ktei = [], [], [], []
    for L in range(nL):
        own = shard.family_owner(L, nranks) == rank
        for e in range(rs.nel):
            t = rs.twoe_integral(L, e)
            n = rs.basis(e).nbf
            k = fo.exchange_tei(t, n, n, n, n)
            a, b = rs.radial_integral(L, e), rs.radial_integral(-L - 1, e)
            if not own:
                t, k, a, b = 0 * t, 0 * k, 0 * a, 0 * b
            dL.append(a); dm1L.append(b); tei.append(t); ktei.append(k)
    o = oracle.AtomicOracle(lval, mval, rs.first, rs.last, dL, dm1L, tei, ktei)
    n = len(lval) * (int(rs.last[-1]) + 1)
    R = splitmix64(n * n).reshape(n, n)
    P = np.asfortranarray(0.5 * (R + R.T))
    return o, P


def _worker(rank, world, port, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        import torch.distributed as dist
        dist.init_process_group("gloo", rank=rank, world_size=world)
        sys.path.insert(0, ROOT)
        from helfem_b200 import shard
        o, P = _atomic_oracle_partial(rank, world)
        Jp, Kp = o.coulomb(P), o.exchange(P)
        J = shard.allreduce_operator(Jp)
        K = shard.allreduce_operator(Kp)
        assert J.flags.f_contiguous and J.shape == Jp.shape
        # block-sparse in-place all-reduce of a device-style tensor (here on the CPU): only the union of non-zero
        # angular blocks travels; the result equals the plain sum exactly
        import torch
        nang_t, nrad_t = 6, 4
        gen = torch.Generator().manual_seed(7)
        parts = [torch.zeros(nang_t * nrad_t, nang_t * nrad_t, dtype=torch.float64) for _ in range(world)]
        for r, blocks in enumerate(([(0, 0), (1, 1), (2, 3)], [(1, 1), (5, 5)])):
            v = parts[r].view(nang_t, nrad_t, nang_t, nrad_t)
            for (bi, bj) in blocks:
                v[bi, :, bj, :] = torch.rand(nrad_t, nrad_t, generator=gen, dtype=torch.float64)
        bs = shard.allreduce_blocksparse_(parts[rank].clone(), nang_t, nrad_t)
        assert torch.equal(bs, parts[0] + parts[1])
        # independent units: every rank takes its slice, results are gathered by rank 0 only for reporting
        b, e = shard.partition_units(7, rank, world)
        counts = [None] * world
        dist.all_gather_object(counts, e - b)
        # dynamic heavy-first queue of the sadatom sweep: two passes; every unit is claimed exactly once per pass
        costs = [1 + (u * 7) % 11 for u in range(23)]
        wq = shard.WorkQueue(shard.heavy_first(costs), chunk=3)
        claimed = []
        for _ in range(2):
            dist.barrier()
            wq.reset()
            dist.barrier()
            mine = []
            while True:
                c = wq.claim()
                if not c:
                    break
                mine += c
            allc = [None] * world
            dist.all_gather_object(allc, mine)
            claimed.append(allc)
        dist.destroy_process_group()
        q.put((rank, J, K, float(np.max(np.abs(Jp))), counts, claimed, costs))
    except Exception as ex:  # pragma: no cover
        import traceback
        q.put((rank, "ERR", traceback.format_exc(), 0, None, None, None))


@pytest.mark.timeout(300)
def test_allreduce_of_family_partials_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(60)
    for r in res:
        assert not isinstance(r[1], str), r[2]
    res.sort(key=lambda r: r[0])
    o, P = _atomic_oracle_partial(0, 1)  # all families
    Jfull, Kfull = o.coulomb(P), o.exchange(P)
    for rank, J, K, pmax, counts, claimed, costs in res:
        from helfem_b200 import shard
        for allc in claimed:
            flat = sum(allc, [])
            assert sorted(flat) == list(range(len(costs)))          # every unit exactly once
            for mine in allc:                                        # each rank sees its units heaviest first
                assert [costs[u] for u in mine] == sorted((costs[u] for u in mine), reverse=True)
        assert np.max(np.abs(J - Jfull)) <= 1e-12 * max(1.0, np.max(np.abs(Jfull)))
        assert np.max(np.abs(K - Kfull)) <= 1e-12 * max(1.0, np.max(np.abs(Kfull)))
        assert pmax > 0  # both ranks own something
        assert counts == [4, 3]
    assert np.array_equal(res[0][1], res[1][1])  # ranks hold bit-identical sums
