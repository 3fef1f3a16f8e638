"""Family-sharded diatomic J / K over NCCL, one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
        tools/shard_nccl_check.py [--lmax 4 --mmax 2 --nelem 4 --nnodes 8]

Every rank builds the same basis, owns the integral families r % N == rank (hfb200_set_shard), forms its
partial J / K on its GPU and all-reduces them (helfem_b200/shard.py).  Rank 0 compares with the CPU
oracle (full family loop) and prints one JSON line with the device-timed partial build and the all-reduce."""
import argparse
import json
import os
import sys
import time

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
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--no-oracle", action="store_true")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import helfem_b200 as hb
    from helfem_b200 import shard
    from util import splitmix64

    rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    b = hb.DiatomicBasis(7, 7, 2.068, nelem=args.nelem, nnodes=args.nnodes, primbas=4, lmax=args.lmax, mmax=args.mmax, device=lr)
    op = shard.ShardedOperator(b, rank=rank, nranks=world)
    t0 = time.time()
    b.compute_tei(True)
    t_tei = time.time() - t0
    n = b.nbf
    R = splitmix64(n * n).reshape(n, n)
    C = splitmix64(n * 4, seed=7).reshape(n, 4) * np.exp(-0.05 * np.arange(n))[:, None]
    P = np.asfortranarray(C @ C.T + 1e-2 * 0.5 * (R + R.T))
    J, K = op.coulomb(P), op.exchange(P)  # warm-up + result
    tj = tk = 0.0
    for _ in range(args.reps):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.time(); op.coulomb(P); torch.cuda.synchronize(); tj += time.time() - t0
        t0 = time.time(); op.exchange(P); torch.cuda.synchronize(); tk += time.time() - t0
    t = torch.tensor([tj / args.reps, tk / args.reps], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out = {"n_gpus": world, "nbf": n, "nlm": int(b.nlm), "nang": int(len(b.lval)), "compute_tei_s": round(t_tei, 3),
           "coulomb_ms": round(1e3 * float(t[0]), 3), "exchange_ms": round(1e3 * float(t[1]), 3)}
    if rank == 0 and not args.no_oracle:
        import oracle
        # the oracle needs ALL families: a second, unsharded host object supplies the stored integrals
        bf = hb.DiatomicBasis(7, 7, 2.068, nelem=args.nelem, nnodes=args.nnodes, primbas=4, lmax=args.lmax, mmax=args.mmax, device=lr)
        bf.compute_tei(True)
        dis = [[bf.tei(w, r, e) for r in range(bf.nlm) for e in range(bf.nel)] for w in range(4)]
        tei = [bf.tei(4 + ab, r, e) for r in range(bf.nlm) for e in range(bf.nel) for ab in range(4)]
        ktei = [bf.tei(8 + ab, r, e) for r in range(bf.nlm) for e in range(bf.nel) for ab in range(4)]
        o = oracle.DiatomicOracle(bf.lval, bf.mval, bf.first, bf.last, bf.Lmap, bf.Mmap, bf.Rhalf, dis, tei, ktei, bf.pure_indices())
        Jo, Ko = o.coulomb(P), o.exchange(P)
        out["max_dJ"] = float(np.max(np.abs(J - Jo)))
        out["max_dK"] = float(np.max(np.abs(K - Ko)))
        out["parity_ok"] = bool(out["max_dJ"] <= max(1e-9, 5e-14 * np.max(np.abs(Jo))) and out["max_dK"] <= max(1e-9, 5e-14 * np.max(np.abs(Ko))))
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0 and not args.no_oracle and not out["parity_ok"]:
        sys.exit(1)


if __name__ == "__main__":
    main()
