"""Family-sharded diatomic J / K with the IN-LIBRARY NCCL all-reduce (hfb200_comm_init), one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
        tools/shard_nccl_check.py [--lmax 4 --mmax 2 --nelem 4 --nnodes 8]

Every rank builds the same basis, owns the integral families r % N == rank (hfb200_set_shard), attaches the communicator
and calls the host-matrix operators TwoDBasis::coulomb / exchange, which return the COMPLETE J / K on every rank.  Every
rank compares with the CPU oracle (full family loop, unsharded second host object) for a symmetric AND a non-symmetric
density (the two item lists of the exchange engine); rank 0 prints one JSON line."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lmax", type=int, default=3)
    ap.add_argument("--mmax", type=int, default=2)
    ap.add_argument("--nelem", type=int, default=3)
    ap.add_argument("--nnodes", type=int, default=6)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import helfem_b200 as hb
    import oracle
    from util import splitmix64

    rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    mk = lambda: hb.DiatomicBasis(7, 7, 2.068, nelem=args.nelem, nnodes=args.nnodes, primbas=4, lmax=args.lmax, mmax=args.mmax, device=lr)
    b = mk()
    b.set_shard(rank, world, owned_only=True)
    b.compute_tei(True)
    uid = [hb.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    b.comm_init(uid[0], rank, world)
    n = b.nbf
    R = splitmix64(n * n).reshape(n, n)
    C = splitmix64(n * 4, seed=7).reshape(n, 4) * np.exp(-0.05 * np.arange(n))[:, None]
    Psym = np.asfortranarray(C @ C.T + 1e-2 * 0.5 * (R + R.T))
    Pns = np.asfortranarray(Psym + 0.3 * np.triu(splitmix64(n * n, seed=77).reshape(n, n), 1))
    bf = mk()  # unsharded: supplies ALL stored integrals to the oracle
    bf.compute_tei(True)
    dis = [[bf.tei(w, r, e) for r in range(bf.nlm) for e in range(bf.nel)] for w in range(4)]
    tei = [bf.tei(4 + ab, r, e) for r in range(bf.nlm) for e in range(bf.nel) for ab in range(4)]
    ktei = [bf.tei(8 + ab, r, e) for r in range(bf.nlm) for e in range(bf.nel) for ab in range(4)]
    o = oracle.DiatomicOracle(bf.lval, bf.mval, bf.first, bf.last, bf.Lmap, bf.Mmap, bf.Rhalf, dis, tei, ktei, bf.pure_indices())
    out = {"n_gpus": world, "nbf": n, "nlm": int(b.nlm), "nang": int(len(b.lval))}
    ok = True
    for name, P in (("sym", Psym), ("nonsym", Pns)):
        J, K = b.coulomb(P), b.exchange(P)
        Jo, Ko = o.coulomb(P), o.exchange(P)
        dJ, dK = float(np.max(np.abs(J - Jo))), float(np.max(np.abs(K - Ko)))
        t = torch.tensor([dJ, dK], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)  # worst rank: every rank must hold the complete operator
        out["max_dJ_" + name], out["max_dK_" + name] = float(t[0]), float(t[1])
        ok = ok and float(t[0]) <= max(1e-9, 5e-14 * np.max(np.abs(Jo))) and float(t[1]) <= max(1e-9, 5e-14 * np.max(np.abs(Ko)))
    # root-driven calls (hfb200_comm_set_root): only rank 0 passes host matrices, the others serve; rank 0 must get the same
    # complete operators (the workers' densities are deliberately garbage-free: they pass nothing)
    b.comm_set_root(0)
    for name, P in (("sym", Psym), ("nonsym", Pns)):
        if rank == 0:
            J, K = b.coulomb(P), b.exchange(P)
            Jo, Ko = o.coulomb(P), o.exchange(P)
            dJ, dK = float(np.max(np.abs(J - Jo))), float(np.max(np.abs(K - Ko)))
            good = dJ <= max(1e-9, 5e-14 * np.max(np.abs(Jo))) and dK <= max(1e-9, 5e-14 * np.max(np.abs(Ko)))
        else:
            b.serve(False)
            b.serve(True)
            dJ = dK = 0.0
            good = True
        t = torch.tensor([dJ, dK, 0.0 if good else 1.0], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["root_driven_max_dJ_" + name], out["root_driven_max_dK_" + name] = float(t[0]), float(t[1])
        ok = ok and float(t[2]) == 0.0
    b.comm_set_root(-1)
    out["parity_ok"] = bool(ok)
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    if not ok:
        sys.exit(1)


if __name__ == "__main__":
    main()
