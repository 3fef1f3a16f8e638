"""Debug: phase timeline of the fused exchange kernels on the first SMs (%globaltimer stamps written by thread 0 of every CTA).
Needs a library built with the trace points:  HFB200_NVCC_EXTRA="-DHFB_XF_TRACE" python -m helfem_b200.build --force
    python tools/xf_trace.py        # same-element kernel: phase 1 / phase 2 of every stack pair
    python tools/xf_trace.py pair   # element-pair kernel
Outputs of round 2: profiles/r02_xf_trace_*.txt (the pair trace there also had stamps inside the stage loop)."""
import ctypes, sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import helfem_b200 as hb
from helfem_b200 import lib
import torch
from util import diatomic_orbitals
b = hb.DiatomicBasis(7, 7, 2.068, nelem=5, nnodes=15, primbas=4, lmax=12, mmax=3)
b.compute_tei(True)
C = diatomic_orbitals(b)
pure = b.pure_indices()
P = np.asfortranarray(2.0 * C[pure] @ C[pure].T)
K = np.zeros_like(P, order="F")
b.exchange_into(P, K)
out = np.zeros(148 * 4096, dtype=np.uint64); cnt = np.zeros(148, dtype=np.int32)
lib().hfb200_debug_trace(out.ctypes.data_as(ctypes.c_void_p), cnt.ctypes.data_as(ctypes.c_void_p))
for sm in range(2):
    n = min(int(cnt[sm]), 4096)
    raw = [(int(v) >> 20, (int(v) >> 4) & 0xffff, int(v) & 15) for v in out[sm * 4096: sm * 4096 + n]]
    info = {blk: (t // 100000, t % 100000) for t, blk, e in raw if e == 5}
    ev = sorted(x for x in raw if x[2] in ((6, 7, 8) if "pair" in sys.argv else (1, 2, 3, 4)))
    t0 = ev[0][0]
    print("  cta: (tiles, slots)", {k: info[k] for k in list(info)[:12]})
    print("SM", sm, "events", n)
    for t, blk, e in [x for x in ev if x[1] == ev[0][1]][:70]:
        print("  %9.2f us  cta %5d  %s" % ((t - t0) / 1e3, blk, {1: "P1 begin", 2: "P1 end", 3: "P2 begin", 4: "P2 end", 6: "pair P1 begin (= previous P2 end)", 7: "pair P1 end", 8: "pair P2 begin"}[e]))
