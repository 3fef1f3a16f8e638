#!/usr/bin/env python
"""bench.py -- SCF Fock builds per second (J + K + Vxc, FP64) on B200.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun, one rank per GPU)
    python bench.py --impl reference ...                      (the CPU arm: oracle port on the host cores)

Default workload: BASELINE.json configs[3], the largest configuration that fits one GPU -- N2 Hartree-Fock in prolate
spheroidal coordinates, lmax 24, mmax 6, 5 LIP-15 elements (19 552 basis functions, 585 integral families).  A *step* is ONE
Fock build: coulomb(P) + exchange(P) (HF has no XC term) for an overlap-orthonormalised N2-like closed-shell density.
`value` = builds/s with P, J, K resident in HBM (hfb200_chan_fock_device = coulomb + exchange of one build);
`e2e`   = the same through the reference-facing host-matrix calls TwoDBasis::coulomb / TwoDBasis::exchange.
N > 1:   STRONG scaling of that one build: every rank holds the integral families r % N == rank and the library all-reduces
         J (per-channel radial results) and K (packed non-zero angular blocks) over NCCL inside the timed region.
`roofline` carries every stage of the build with its algorithmic flops / bytes, the dominant kernel on top.
`cpu_baseline` / `--impl reference`: the CPU port of the reference loops on a uniform SAMPLE of output blocks, scaled to
the full operator (kind "port-sampled"; the full port would need ~1e13 flop per exchange call).

Other workloads (--workload): the sadatom batches kr_pbe_cfg2 (configs[1]) and rn_r2scan_cfg3 (configs[2]): a step is `batch`
independent Fock builds (coulomb + eval_Fxc) on one basis, ranks process their own batches (weak scaling, no collective);
ne_hf_cfg1 (configs[0]); smaller diatomic cases; the product-grid XC cases.
"""
import os as _os
import sys as _sys
if int(_os.environ.get("WORLD_SIZE", "1")) > 1 and _os.environ.get("OMP_NUM_THREADS", "1") == "1":
    # torchrun exports OMP_NUM_THREADS=1, but the host side of the operator calls (block scan, pack, zero-fill) is OpenMP code.
    # Set before the native library and its OpenMP runtime are loaded: the sharded atomic / diatomic build is driven by rank 0
    # alone (the other ranks serve and do no host work), every other workload gives each rank its share of the cores.
    _wl = _sys.argv[_sys.argv.index("--workload") + 1] if "--workload" in _sys.argv[:-1] else "n2_hf_cfg4"
    _cores = len(_os.sched_getaffinity(0))
    _local = int(_os.environ.get("LOCAL_WORLD_SIZE", _os.environ["WORLD_SIZE"]))
    if "_hf_" in _wl and "--replicas" not in _sys.argv:
        _os.environ["OMP_NUM_THREADS"] = str(max(1, _cores - _local + 1) if int(_os.environ.get("RANK", "0")) == 0 else 1)
    else:
        _os.environ["OMP_NUM_THREADS"] = str(max(1, _cores // _local))
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if os.environ.get("NCCL_DEBUG", "").upper() not in ("INFO", "TRACE"):
    os.environ["NCCL_DEBUG"] = "NONE"  # keep NCCL's version banner off stdout: rank 0 prints exactly one JSON line
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # name: dict(Z, primbas, nnodes, nelem, taylor, lmax, method, nspin)
    "kr_pbe_cfg2": dict(Z=36, primbas=4, nnodes=15, nelem=20, taylor=-1, lmax=3, x=101, c=130, nspin=1,
                        desc="gensap --Z Kr --nelem 20 --nnodes 15 --primbas 4 --method gga_x_pbe-gga_c_pbe (configs[1])"),
    "rn_r2scan_cfg3": dict(Z=86, primbas=5, nnodes=15, nelem=40, taylor=20, lmax=3, x=497, c=498, nspin=2,
                           desc="gensap --Z Rn --nelem 40 --nnodes 15 --primbas 5 --method mgga_x_r2scan-mgga_c_r2scan (configs[2])"),
    # Hartree-Fock builds of the atomic / diatomic programs: coulomb(P) + exchange(P) through the channel engine
    "ne_hf_cfg1": dict(kind="atomic", Z=10, lmax=0, mmax=0, nelem=10, nnodes=15, primbas=4, nocc=5,
                       desc="atomic --Z Ne --lmax 0 --mmax 0 --nelem 10 --nnodes 15 --primbas 4 --method HF (configs[0])"),
    "n2_hf_cfg4": dict(kind="diatomic", Z1=7, Z2=7, Rbond=2.068, lmax=24, mmax=6, nelem=5, nnodes=15, primbas=4,
                       desc="diatomic --Z1 N --Z2 N --Rbond 2.068 --lmax 24 --mmax 6 --nelem 5 --nnodes 15 --primbas 4 --method HF (configs[3])"),
    "n2_hf_l12m3": dict(kind="diatomic", Z1=7, Z2=7, Rbond=2.068, lmax=12, mmax=3, nelem=5, nnodes=15, primbas=4,
                        desc="diatomic N2, lmax 12, mmax 3, 5 elements (half-size version of configs[3])"),
    "n2_hf_l4m1": dict(kind="diatomic", Z1=7, Z2=7, Rbond=2.068, lmax=4, mmax=1, nelem=3, nnodes=10, primbas=4,
                       desc="diatomic N2, lmax 4, mmax 1, 3 elements (small case the CPU oracle finishes in seconds)"),
    # product-grid DFT of the diatomic program (configs[3]b in SURVEY 8: the 2-D prolate grid), first correct version
    "n2_pbe_grid_l6m2": dict(kind="grid2d", Z1=7, Z2=7, Rbond=2.068, lmax=6, mmax=2, nelem=3, nnodes=10, primbas=4, x=101, c=130,
                             desc="diatomic N2 PBE eval_Fxc on the product grid, lmax 6, mmax 2, 3 elements, ldft 4 lmax+12, mdft 4 (mmax+1)+5"),
    "n2_pbe_cfg4b": dict(kind="grid2d", Z1=7, Z2=7, Rbond=2.068, lmax=24, mmax=6, nelem=5, nnodes=15, primbas=4, x=101, c=130,
                         desc="diatomic --Z1 N --Z2 N --Rbond 2.068 --lmax 24 --mmax 6 --nelem 5 --nnodes 15 --primbas 4 --method gga_x_pbe-gga_c_pbe: "
                              "eval_Fxc on the product grid, ldft 108 x mdft 33 = 3 564 angular x 375 mu points (configs[3]b of SURVEY 8)"),
    "n2_pbe_grid_l2m1": dict(kind="grid2d", Z1=7, Z2=7, Rbond=2.068, lmax=2, mmax=1, nelem=2, nnodes=5, primbas=4, x=101, c=130,
                             desc="diatomic N2 PBE eval_Fxc on the product grid, lmax 2, mmax 1 (oracle-sized)"),
    # the Z sweep (configs[4]): one shared Kr-type basis, every atom's trial configurations, dynamic queue over the ranks
    "sweep_z1_86": dict(kind="sweep", zmin=1, zmax=86, primbas=4, nnodes=15, nelem=20, taylor=-1, lmax=3, x=101, c=130,
                        desc="gensap --Z 1..86 --nelem 20 --nnodes 15 --primbas 4 --method gga_x_pbe-gga_c_pbe (configs[4]), one process group"),
    "ne_pbe_small": dict(Z=10, primbas=4, nnodes=15, nelem=5, taylor=-1, lmax=2, x=101, c=130, nspin=1,
                         desc="small smoke case"),
}


FP64_TFLOPS_MEASURED = 35.5  # cuBLAS DGEMM 8192^3 on this pool's B200, profiles/r01_fp64_peak.json (MEASURED_PEAKS.json has no FP64 entry)


def ncu_traffic(workload, kernel):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture of this same
    workload (profiles/ncu_traffic.json, written from the .ncu-rep by tools/ncu_summary.py); None if not captured."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        return json.load(open(p))[workload][kernel]
    except Exception:
        return None


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.samples, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_basis(w, device=0):
    import helfem_b200 as hb
    b = hb.SadatomBasis(Z=w["Z"], nelem=w["nelem"], nnodes=w["nnodes"], primbas=w["primbas"], taylor_order=w["taylor"],
                        lmax=w["lmax"], device=device)
    b.compute_tei(1, False)  # a pure functional only ever reads L = 0 (sadatom/basis.cpp:306)
    return b


def make_densities(b, w):
    """Aufbau density cube(s) of core-Hamiltonian orbitals, divided by 4 pi as FockBuild passes them
    (sadatom/solver.cpp:784,804)."""
    from util import sadatom_density
    S, T, Tl, V = b.overlap(), b.kinetic(), b.kinetic_l(), b.nuclear()
    if w["nspin"] == 1:
        return [sadatom_density(S, T, Tl, V, w["Z"], w["lmax"]) / (4 * np.pi)]
    return [sadatom_density(S, T, Tl, V, w["Z"], w["lmax"], spin=s) / (4 * np.pi) for s in ("a", "b")]


def algorithmic(b, w):
    """Algorithmic bytes / flops of ONE Fock build (restated in DESIGN.md)."""
    nsl, nspin = w["lmax"] + 1, w["nspin"]
    fam = {1: 1, 101: 2, 497: 3}[w["x"]]
    c = {1: 4, 2: 6, 3: 14}[fam]
    npe = (b.last - b.first + 1).astype(np.int64)
    np2 = b.np_max ** 2
    grid_flops = int(c * np.sum(npe ** 2) * b.nq * nspin)
    n_tab = 1 if fam == 1 else 2
    # batched: per build the P band comes in and the H band goes out; tables are shared (L2-resident)
    grid_bytes_batched = int(2 * nspin * nsl * b.nel * np2 * 8)
    grid_bytes_single = grid_bytes_batched + int(n_tab * np.sum(npe) * b.nq * 8)
    # Coulomb on triangle-packed integrals: (ij|kl) = (ji|kl) = (ij|lk), so ntri = np(np+1)/2 per side
    ntri = npe * (npe + 1) // 2
    j_flops = int(2 * np.sum(ntri ** 2))
    j_bytes_single = int(np.sum(ntri ** 2) * 8 + (nspin * nsl + 1) * b.nel * np2 * 8)
    j_bytes_batched = int((nspin * nsl + 1) * b.nel * np2 * 8)
    return dict(grid_flops=grid_flops, grid_bytes_batched=grid_bytes_batched, grid_bytes_single=grid_bytes_single,
                j_flops=j_flops, j_bytes_single=j_bytes_single, j_bytes_batched=j_bytes_batched)


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port (oracle/hotpath_oracle.cpp, OpenMP) fed with the same tables
# ------------------------------------------------------------------------------------------------
def oracle_from_basis(b):
    import oracle
    tabs = [b.tables(e) for e in range(b.nel)]
    o = oracle.SadatomOracle(b.first, b.last, [t[0] for t in tabs], [t[1] for t in tabs], [t[2] for t in tabs],
                             [t[3] for t in tabs], b.lmax)
    tei = [b.tei(0, e) for e in range(b.nel)]
    o.set_tei([t[0] for t in tei], [t[1] for t in tei], [t[2] for t in tei])
    return o


def cpu_builds(o, w, dens, nbuilds, nthreads=1):
    """nbuilds Fock builds on the host cores; returns seconds.  nthreads > 1: ONE BUILD PER THREAD (the reference runs the sweep /
    configuration search as independent processes, one per core), each build single-threaded inside (ctypes releases the GIL)."""
    def one(i):
        s = 1.0 + 0.01 * i
        if w["nspin"] == 1:
            P = dens[0] * s
            o.coulomb(np.asfortranarray(P.sum(axis=2)))
            o.eval_fxc(w["x"], w["c"], P)
        else:
            Pa, Pb = dens[0] * s, dens[1] * s
            o.coulomb(np.asfortranarray((Pa + Pb).sum(axis=2)))
            o.eval_fxc(w["x"], w["c"], Pa, Pb)
    t0 = time.perf_counter()
    if nthreads <= 1:
        for i in range(nbuilds):
            one(i)
    else:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=nthreads) as ex:
            list(ex.map(one, range(nbuilds)))
    return time.perf_counter() - t0


def sadatom_cpu_setup(w):
    """Oracle + densities of a sadatom workload from the numpy setup (oracle/fem_oracle.py): the CPU arm never maps the product library."""
    import oracle
    from oracle import fem_oracle as fo
    from util import sadatom_density
    rs = fo.RadialSetup(w["primbas"], w["nnodes"], fo.get_grid(40.0, w["nelem"]), taylor_order=w["taylor"])
    o = oracle.sadatom_from_setup(rs, w["lmax"], nL=1, exchange=False)
    S, T, Tl, V = rs.overlap(), rs.kinetic(), rs.kinetic_l(), rs.nuclear(w["Z"])
    if w["nspin"] == 1:
        dens = [sadatom_density(S, T, Tl, V, w["Z"], w["lmax"]) / (4 * np.pi)]
    else:
        dens = [sadatom_density(S, T, Tl, V, w["Z"], w["lmax"], spin=sp) / (4 * np.pi) for sp in ("a", "b")]
    return o, dens


def run_reference(args, w, rank):
    if rank != 0:
        return
    import oracle
    o, dens = sadatom_cpu_setup(w)
    nthreads = host_threads()
    oracle.lib().oracle_set_num_threads(1)  # one build per thread, every build single-threaded inside
    one = cpu_builds(o, w, dens, 1)
    per_step = max(nthreads, min(args.batch, int(2.0 * nthreads / max(one, 1e-6))))  # ~2 s of CPU work per step
    for _ in range(args.warmup):
        cpu_builds(o, w, dens, max(nthreads, per_step // 4), nthreads)
    t = 0.0
    for _ in range(args.steps):
        t += cpu_builds(o, w, dens, per_step, nthreads)
    value = per_step * args.steps / t
    line = {"impl": "reference", "metric": "fock_builds_per_sec", "value": value, "unit": "builds/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload, "desc": w["desc"], "builds_per_step": per_step,
                       "note": "CPU restatement of the reference loops (oracle/hotpath_oracle.cpp, -O3), tables from oracle/fem_oracle.py, one build "
                               "per host thread; the reference itself cannot be compiled here (no Armadillo/GSL/libxc/HDF5/gfortran)"},
            "cpu_baseline": {"value": value, "unit": "builds/s", "cores": nthreads, "kind": "port",
                             "sample": "%d builds per step x %d steps, one build per thread" % (per_step, args.steps)},
            "e2e": {"value": value, "unit": "builds/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def run_gpu(args, w, rank, world, local_rank):
    import torch
    import helfem_b200 as hb
    if not torch.cuda.is_available() or hb.device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: helfem_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    b = make_basis(w, device=local_rank)
    dens = make_densities(b, w)
    nsl, nspin, nb = w["lmax"] + 1, w["nspin"], args.batch
    blk = b.nel * b.np_max ** 2
    base = np.zeros((nspin, nsl, blk))
    for s in range(nspin):
        for l in range(nsl):
            base[s, l] = b.pack_band(dens[s][:, :, l])
    # batch of distinct trial densities: rank- and build-dependent rescaling (XC is nonlinear in P)
    rng = np.random.RandomState(20261104 + rank)
    scales = 1.0 + 0.05 * rng.uniform(-1, 1, nb)
    dbase = torch.from_numpy(base).cuda()
    dP = (torch.from_numpy(scales).cuda().view(nb, 1, 1, 1) * dbase.unsqueeze(0)).contiguous()
    dJ = torch.zeros((nb, blk), dtype=torch.float64, device="cuda")
    dH = torch.zeros((nb, nspin, nsl, blk), dtype=torch.float64, device="cuda")
    dS = torch.zeros((nb, 4), dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream()
    sh = stream.cuda_stream
    x, c, thr = w["x"], w["c"], 1e-12

    def step(J=True, H=True):
        b.fock_batch(nb, x, c, nspin, nsl, thr, dP.data_ptr(), dJ.data_ptr() if J else 0, dH.data_ptr() if H else 0,
                     dS.data_ptr() if H else 0, sh)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for e0, e1 in ev:
            e0.record(stream)
            fn()
            e1.record(stream)
        torch.cuda.synchronize()
        return sum(e0.elapsed_time(e1) for e0, e1 in ev)

    for _ in range(max(args.warmup, 3)):
        step()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    ms = timed(step, args.steps)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = nb * world * args.steps / (ms * 1e-3)

    if args.kernels_only:  # for ncu / tuning runs: just the timed device steps, nothing else
        out = {"kernels_only": True, "ms_per_step": ms / args.steps, "value": value}
        if args.split:
            out["ms_grid"] = timed(lambda: step(J=False), args.steps) / args.steps
            out["ms_coulomb"] = timed(lambda: step(H=False), args.steps) / args.steps
        if rank == 0:
            print(json.dumps(out), flush=True)
        return

    # ---- per-kernel timing (rank 0 reports): the grid/XC kernel and the Coulomb kernels on their own ----
    ms_grid = timed(lambda: step(J=False), args.steps) / args.steps
    ms_j = timed(lambda: step(H=False), args.steps) / args.steps

    # ---- end to end: the same batch through the C ABI with HOST buffers (pinned element bands in and out,
    # H2D + kernels + D2H inside the timed region) ----
    hP = torch.empty(dP.shape, dtype=torch.float64).pin_memory()
    hP.copy_(dP.cpu())
    hJ = torch.empty(dJ.shape, dtype=torch.float64).pin_memory()
    hH = torch.empty(dH.shape, dtype=torch.float64).pin_memory()
    hS = torch.empty(dS.shape, dtype=torch.float64).pin_memory()
    hPn, hJn, hHn, hSn = hP.numpy(), hJ.numpy(), hH.numpy(), hS.numpy()

    def e2e_pass():
        b.fock_host(nb, x, c, nspin, nsl, thr, hPn, hJn, hHn, hSn)

    e2e_pass()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_pass()
    t_e2e = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([t_e2e], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_e2e = float(t.item())
    e2e_value = nb * world * args.steps / t_e2e
    h2d = hP.numel() * 8
    d2h = (hJ.numel() + hH.numel() + hS.numel()) * 8
    # cross-check: host path reproduces the device-resident results bit for bit
    if not (torch.equal(hJ, dJ.cpu()) and torch.equal(hH, dH.cpu())):
        raise RuntimeError("host-buffer path and device-resident path disagree")

    # ---- latency of ONE build through the dense reference-facing calls (coulomb + eval_Fxc, dense host matrices) ----
    import helfem_b200 as hb_
    L = hb_.lib()
    ctx = b.context()
    nrad = b.nbf
    Pd = [np.asfortranarray(d) for d in dens]
    Ptot = np.asfortranarray(sum(d.sum(axis=2) for d in dens))
    Jd = np.zeros((nrad, nrad), order="F")
    Hd = [np.zeros((nrad, nrad, nsl), order="F") for _ in range(nspin)]
    exc_, nel_ = ctypes.c_double(), ctypes.c_double()
    dpp = ctypes.POINTER(ctypes.c_double)
    ptr = lambda a: a.ctypes.data_as(dpp)

    def dense_build():
        rc = L.hfb200_sadatom_coulomb(ctx, ptr(Ptot), nrad, ptr(Jd), nrad)
        rc |= L.hfb200_sadatom_eval_fxc(ctx, x, c, nspin, ptr(Pd[0]), ptr(Pd[1]) if nspin == 2 else None, nrad, nsl,
                                        ctypes.c_double(thr), ptr(Hd[0]), ptr(Hd[1]) if nspin == 2 else None, nrad, 1,
                                        ctypes.byref(exc_), ctypes.byref(nel_))
        if rc:
            raise RuntimeError("dense C-ABI call failed")

    for _ in range(3):
        dense_build()
    t0 = time.perf_counter()
    ndense = 50
    for _ in range(ndense):
        dense_build()
    dense_us = (time.perf_counter() - t0) / ndense * 1e6

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    alg = algorithmic(b, w)
    hbm_peak, peak_src = peaks()
    # dominant kernel of the batched step
    if ms_grid >= ms_j:
        kname, kms, kbytes = "sad_grid_kernel", ms_grid, alg["grid_bytes_batched"] * nb
        kflops = alg["grid_flops"] * nb
    else:
        kname, kms, kbytes = "sad_coulomb_kernel", ms_j, alg["j_bytes_batched"] * nb
        kflops = alg["j_flops"] * nb
    achieved = kbytes / (kms * 1e-3) / 1e9
    tflops = kflops / (kms * 1e-3) / 1e12
    hbm_frac, fp64_frac = achieved / hbm_peak, tflops / FP64_TFLOPS_MEASURED
    # the kernel's algorithmic intensity sits at or above the FP64 ridge, so both fractions are reported
    # and the larger one names the binding roof ("tensor" = the FP64 DMMA/DFMA pipe; tcgen05 has no FP64 kind)
    common = {"kernel": kname, "traffic": ncu_traffic(args.workload, kname), "algorithmic_bytes_per_launch": kbytes,
              "algorithmic_flops_per_launch": kflops, "ms_per_launch": kms, "ms_grid_kernel": ms_grid, "ms_coulomb_kernels": ms_j,
              "hbm": {"achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_frac, "peak_source": peak_src},
              "fp64": {"achieved": tflops, "peak": FP64_TFLOPS_MEASURED, "unit": "TFLOP/s", "frac": fp64_frac,
                       "peak_source": "measured cuBLAS DGEMM FP64 (profiles/r01_fp64_peak.json)"}}
    if fp64_frac >= hbm_frac:
        roofline = dict({"bound": "tensor", "achieved": tflops, "peak": FP64_TFLOPS_MEASURED, "unit": "TFLOP/s", "frac": fp64_frac}, **common)
    else:
        roofline = dict({"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_frac}, **common)
    # CPU baseline: the oracle port on this box's host cores, bounded sample
    cpu = None
    if world == 1:  # reported at N = 1 only
        import oracle
        nthr = host_threads()
        oracle.lib().oracle_set_num_threads(1)  # one build per thread, every build single-threaded inside
        o = oracle_from_basis(b)
        one = cpu_builds(o, w, dens, 1)
        nsample = max(nthr, min(200 * nthr, int(10.0 * nthr / max(one, 1e-6))))
        tc = cpu_builds(o, w, dens, nsample, nthr)
        cpu = {"value": nsample / tc, "unit": "builds/s", "cores": nthr, "kind": "port",
               "sample": "%d Fock builds (coulomb + eval_Fxc) of the same workload, one build per thread" % nsample}
    line = {"metric": "fock_builds_per_sec", "value": value, "unit": "builds/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload, "desc": w["desc"], "builds_per_step_per_gpu": nb,
                       "operators": "coulomb + eval_Fxc(%s) %s" % ("PBE" if x == 101 else "r2SCAN" if x == 497 else "LDA",
                                                                   "restricted" if nspin == 1 else "unrestricted"),
                       "l2": "inputs larger than L2: P+H+J bands = %.0f MB per step" % ((dP.numel() + dH.numel() + dJ.numel()) * 8 / 1e6),
                       "partition": "independent builds per rank, no data-path collective"},
            "roofline": roofline, "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": "builds/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "builds_per_step": nb, "api": "hfb200_sadatom_fock_host: pinned host element bands in, J/H bands + scalars out",
                    "dense_single_build_us": dense_us,
                    "dense_api": "hfb200_sadatom_coulomb + hfb200_sadatom_eval_fxc on dense host matrices, one build per call"},
            "gpu_launches": int(args.steps * 4), "clocks": clocks}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
# atomic / diatomic Hartree-Fock builds: J + K through the Gaunt-coupled channel engine
# ------------------------------------------------------------------------------------------------
def chan_orbitals(b, kind, rank):
    """Synthetic occupied orbitals, each pure in m (what an SCF of an atom / linear molecule produces).  Diatomic: seven
    N2-like orbitals (five sigma, pi+, pi-) Loewdin-ORTHONORMALISED in the overlap metric (tests/util.py), so J and K have the
    magnitude of a converged SCF density; atomic: five s-type orbitals.  Returns C (ndim x norb) in the operator's index space."""
    from util import splitmix64, diatomic_orbitals
    if kind == "diatomic" and int(np.max(np.abs(b.mval))) >= 1:
        return diatomic_orbitals(b, seed=100 + 17 * rank)
    nrad, nang = b.nrad, b.nang
    ndim = nang * nrad
    cols = []
    for q, m in enumerate([0] * 5):
        c = np.zeros((nang, nrad))
        sel = np.where(b.mval == m)[0]
        noise = splitmix64(len(sel) * nrad, seed=100 + q + 17 * rank).reshape(len(sel), nrad)
        c[sel] = noise * np.exp(-0.35 * b.lval[sel])[:, None] * np.exp(-0.06 * np.arange(nrad))[None, :]
        cols.append(c.reshape(ndim))
    return np.stack(cols, axis=1)


def make_chan_basis(w, device=0):
    import helfem_b200 as hb
    if w["kind"] == "atomic":
        return hb.AtomicBasis(w["Z"], nelem=w["nelem"], nnodes=w["nnodes"], primbas=w["primbas"], lmax=w["lmax"], mmax=w["mmax"], device=device)
    return hb.DiatomicBasis(w["Z1"], w["Z2"], w["Rbond"], nelem=w["nelem"], nnodes=w["nnodes"], primbas=w["primbas"],
                            lmax=w["lmax"], mmax=w["mmax"], device=device)


def chan_oracle(b, kind):
    """CPU port of the reference's coulomb / exchange (oracle/chan_oracle.cpp) on the SAME stored integrals (zero-copy views)."""
    import oracle
    if kind == "diatomic":
        dis = [[b.tei_view(ww, r, e) for r in range(b.nlm) for e in range(b.nel)] for ww in range(4)]
        tei = [b.tei_view(4 + ab, r, e) for r in range(b.nlm) for e in range(b.nel) for ab in range(4)]
        ktei = [b.tei_view(8 + ab, r, e) for r in range(b.nlm) for e in range(b.nel) for ab in range(4)]
        return oracle.DiatomicOracle(b.lval, b.mval, b.first, b.last, b.Lmap, b.Mmap, b.Rhalf, dis, tei, ktei, b.pure_indices())
    nL = 2 * int(np.max(b.lval)) + 1
    t = [b.tei(L, e) for L in range(nL) for e in range(b.nel)]
    return oracle.AtomicOracle(b.lval, b.mval, b.first, b.last, [x[0] for x in t], [x[1] for x in t], [x[2] for x in t], [x[3] for x in t])


def cpu_chan_build(b, kind, Ph, budget_s=12.0, seed=2026):
    """Seconds per Fock build (coulomb + exchange) of the CPU port on the host cores.  Small bases: one full build.  Large
    diatomic bases: a UNIFORM SAMPLE of output blocks through the block-restricted oracle (the same loops, evaluated for the
    sampled blocks only), scaled to the full operator: exchange by the number of output blocks with m_j = m_k (the others
    vanish for an m-diagonal density and cost the port nothing), coulomb by the number of channels.  Returns
    (seconds, kind, sample description, extra)."""
    import oracle
    from util import splitmix64
    oracle.lib().oracle_set_num_threads(host_threads())
    o = chan_oracle(b, kind)
    n = Ph.shape[0]
    if kind != "diatomic" or n <= 2500:
        t0 = time.perf_counter()
        Jo, Ko = o.coulomb(Ph), o.exchange(Ph)
        return time.perf_counter() - t0, "port", "1 Fock build (coulomb + exchange) of the same workload", {"J": Jo, "K": Ko}
    nang = b.nang
    mdiag = [(j, k) for j in range(nang) for k in range(nang) if b.mval[j] == b.mval[k]]
    u = splitmix64(4096, seed=seed)
    pick = lambda lst, cnt, off: [lst[int((u[off + q] + 1.0) * 0.5 * len(lst)) % len(lst)] for q in range(cnt)]
    t0 = time.perf_counter()
    o.exchange_blocks(Ph, [], [])  # the block-norm screen of the density alone (a fixed cost of every exchange call)
    t_mask = time.perf_counter() - t0
    nk = 16
    t0 = time.perf_counter()
    ks = pick(mdiag, nk, 0)
    o.exchange_blocks(Ph, [p[0] for p in ks], [p[1] for p in ks])
    t1 = time.perf_counter() - t0 - t_mask
    nk2 = int(max(16, min(len(mdiag), nk * 0.6 * budget_s / max(t1, 1e-3))))
    ks = pick(mdiag, nk2, 100)
    t0 = time.perf_counter()
    o.exchange_blocks(Ph, [p[0] for p in ks], [p[1] for p in ks])
    tK = max(time.perf_counter() - t0 - t_mask, 1e-6)
    est_K = t_mask + tK * len(mdiag) / nk2
    allp = [(i, j) for i in range(nang) for j in range(nang)]
    js = pick(allp, 6, 2000)
    chans = set()
    for i, j in js:
        li, mi, lj, mj = int(b.lval[i]), int(b.mval[i]), int(b.lval[j]), int(b.mval[j])
        for L in range(max(abs(lj - li) - 2, abs(mj - mi)), lj + li + 3):
            chans.add((L, mj - mi))
    t0 = time.perf_counter()
    o.coulomb_blocks(Ph, [p[0] for p in js], [p[1] for p in js])
    tJ = time.perf_counter() - t0
    nchan_all = len({(L, int(b.mval[k]) - int(b.mval[l])) for k in range(nang) for l in range(nang)
                     for L in range(max(abs(int(b.lval[k]) - int(b.lval[l])) - 2, abs(int(b.mval[k]) - int(b.mval[l]))), int(b.lval[k]) + int(b.lval[l]) + 3)})
    est_J = tJ * nchan_all / max(len(chans), 1)
    sample = ("exchange: %d of the %d output blocks with m_j = m_k (uniform, seeded) in %.1f s -> %.0f s per call; coulomb: %d of %d channels "
              "in %.1f s -> %.0f s per call" % (nk2, len(mdiag), tK + t_mask, est_K, len(chans), nchan_all, tJ, est_J))
    return est_K + est_J, "port-sampled", sample, {"fraction_exchange": nk2 / len(mdiag), "fraction_coulomb": len(chans) / nchan_all}


def run_chan_reference(args, w, rank):
    """CPU arm of the atomic / diatomic workloads: the oracle port on the host cores (stored integrals from the host library's
    compute_tei, which is setup, not the timed path; no device context is created).  A step = one (sampled) Fock build."""
    if rank != 0:
        return
    b = make_chan_basis(w)
    b.compute_tei(True)
    kind = w["kind"]
    C = chan_orbitals(b, kind, 0)
    P = 2.0 * C @ C.T
    if kind == "diatomic":
        idx = b.pure_indices()
        P = P[np.ix_(idx, idx)]
    Ph = np.asfortranarray(P)
    import oracle
    steps = max(1, min(args.steps, 3))
    ts, info = [], None
    for q in range(steps):
        t, k, sample, extra = cpu_chan_build(b, kind, Ph, budget_s=10.0, seed=2026 + q)
        ts.append(t)
        info = (k, sample, extra)
    sec = float(np.mean(ts))
    value = 1.0 / sec
    extra = {kk: vv for kk, vv in info[2].items() if not hasattr(vv, "shape")}
    line = {"impl": "reference", "metric": "fock_builds_per_sec", "value": value, "unit": "builds/s", "n_gpus": args.gpus, "steps": steps,
            "warmup": 0, "ms_per_step": 1e3 * sec, "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload, "desc": w["desc"], "operators": "coulomb + exchange (HF)", "nbf": int(Ph.shape[0]),
                       "note": "CPU restatement of the reference loops (oracle/chan_oracle.cpp, -O3 -fopenmp); the reference itself cannot be "
                               "compiled here (no Armadillo/GSL/libxc/HDF5/gfortran)"},
            "cpu_baseline": dict({"value": value, "unit": "builds/s", "cores": oracle.lib().oracle_num_threads(), "kind": info[0], "sample": info[1]}, **extra),
            "e2e": {"value": value, "unit": "builds/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_chan(args, w, rank, world, local_rank):
    import torch
    import helfem_b200 as hb
    if not torch.cuda.is_available() or hb.device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: helfem_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    kind = w["kind"]
    t0 = time.perf_counter()
    b = make_chan_basis(w, device=local_rank)
    # N > 1: STRONG scaling of one build.  The integral families r % world == rank live on this rank; the library all-reduces
    # J (per-channel radial results) and K (packed non-zero angular blocks) over NCCL inside the timed region.
    sharded = world > 1 and not args.replicas
    if sharded:
        if kind == "diatomic":
            b.set_shard(rank, world, owned_only=True)
        else:
            b.set_shard(rank, world)
    b.compute_tei(True)
    t_tei = time.perf_counter() - t0
    t0 = time.perf_counter()
    n = b.chan_dim()  # uploads the integral tables
    t_upload = time.perf_counter() - t0
    if sharded:
        uid = [hb.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        b.comm_init(uid[0], rank, world)
    C = torch.from_numpy(chan_orbitals(b, kind, 0 if sharded else rank)).cuda()
    dP = (2.0 * C @ C.T).contiguous()
    dP = (0.5 * (dP + dP.T)).contiguous()  # exactly symmetric (what C C^T through syrk gives a host driver)
    dJ = torch.zeros_like(dP)
    dK = torch.zeros_like(dP)
    torch.cuda.synchronize()
    # a real (non-default) stream: the C ABI reads a NULL stream as "the context's own stream"
    torch.cuda.set_stream(torch.cuda.Stream())
    stream = torch.cuda.current_stream()
    sh = stream.cuda_stream
    assert sh != 0

    def step(J=True, K=True):
        if J and K:  # one Fock build: hfb200_chan_fock_device (the Coulomb kernels on a second stream next to the exchange pipeline)
            b.fock_device(dP.data_ptr(), dJ.data_ptr(), dK.data_ptr(), n, sh)
        elif J:
            b.coulomb_device(dP.data_ptr(), dJ.data_ptr(), n, sh)
        elif K:
            b.exchange_device(dP.data_ptr(), dK.data_ptr(), n, sh)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    t0 = time.perf_counter()
    step()
    torch.cuda.synchronize()
    t_first = time.perf_counter() - t0  # includes the one-off output-list build for this sparsity pattern
    for _ in range(max(args.warmup, 3) - 1):
        step()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    ms = timed(step, args.steps)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    units = 1 if (sharded or world == 1) else world
    value = units * args.steps / (ms * 1e-3)
    barrier()
    ms_j = timed(lambda: step(K=False), args.steps) / args.steps
    barrier()
    ms_k = timed(lambda: step(J=False), args.steps) / args.steps
    barrier()
    b.set_profile(True)
    step(J=False)
    torch.cuda.synchronize()
    items, x3_flops, stage = b.exchange_stats()
    ms_ar = b.allreduce_ms()
    cnt = b.exchange_counts()
    b.set_profile(False)
    if args.kernels_only:
        if rank == 0:
            print(json.dumps({"kernels_only": True, "ms_per_step": ms / args.steps, "ms_coulomb": ms_j, "ms_exchange": ms_k,
                              "stage_ms": stage, "allreduce_ms": ms_ar, "items": items}), flush=True)
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- end to end: dense HOST matrices through the reference-facing calls TwoDBasis::coulomb + TwoDBasis::exchange (every rank
    #      passes the same host density and gets the complete host J and K, like an SPMD driver would) ----
    Ph = dP.cpu().numpy()
    if kind == "diatomic":
        idx = b.pure_indices()
        Ph = np.asfortranarray(Ph[np.ix_(idx, idx)])
    else:
        Ph = np.asfortranarray(Ph)
    ne2e = 2 if n > 8000 else 5
    driven = world > 1 and not args.replicas
    if driven:
        # N > 1: ONE process (rank 0) holds the host matrices and makes the reference-facing calls, like the reference's
        # single-process SCF driver would; the other ranks serve (hfb200_comm_set_root / hfb200_chan_serve): the non-zero blocks
        # of P reach them over NCCL, not through their own host memory
        b.comm_set_root(0)
    if rank == 0 or not driven:
        Jh, Kh = np.zeros_like(Ph, order="F"), np.zeros_like(Ph, order="F")  # caller-owned outputs, reused like an SCF loop's matrices

    def host_build():
        if rank == 0 or not driven:
            b.coulomb_into(Ph, Jh)
            b.exchange_into(Ph, Kh)
        else:
            b.serve(False)
            b.serve(True)

    host_build()
    barrier()
    t0 = time.perf_counter()
    for _ in range(ne2e):
        host_build()
    torch.cuda.synchronize()
    t_e2e = (time.perf_counter() - t0) / ne2e
    if driven:
        b.comm_set_root(-1)
    if dist is not None:
        t = torch.tensor([t_e2e], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_e2e = float(t.item())
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    nb = Ph.shape[0]
    # PCIe payload of the host-matrix calls: non-zero angular blocks of P down, m-compatible blocks of J / K back
    blkmax = dP.abs().reshape(b.nang, b.nrad, b.nang, b.nrad).amax(dim=(1, 3)).cpu().numpy() > 0
    mv = np.asarray(b.mval)
    dset = set((mv[i] - mv[l]) for i, l in zip(*np.nonzero(blkmax)))
    nz_in = int(np.count_nonzero(blkmax))
    nz_out = int(sum(1 for j in range(b.nang) for k in range(b.nang) if (mv[j] - mv[k]) in dset))
    cpu = None
    if world == 1:
        sec, ckind, sample, extra = cpu_chan_build(b, kind, Ph)
        import oracle
        cpu = {"value": 1.0 / sec, "unit": "builds/s", "cores": oracle.lib().oracle_num_threads(), "kind": ckind, "sample": sample}
        if "K" in extra:
            cpu["max_dJ"] = float(np.max(np.abs(np.asarray(Jh) - extra["J"])))
            cpu["max_dK"] = float(np.max(np.abs(np.asarray(Kh) - extra["K"])))
        else:
            cpu.update(extra)
    # ---- per-stage rooflines: algorithmic flops / bytes against the measured peaks ----
    hbm_peak, hbm_src = peaks()
    st2 = 4 if kind == "diatomic" else 1
    stack = 2 if kind == "diatomic" else 1
    nel, nrad, np2 = int(b.nel), int(b.nrad), int(np.max(b.last - b.first + 1)) ** 2
    npe = (b.last - b.first + 1).astype(np.float64)
    per_same = float(np.sum(2.0 * npe ** 4))
    per_pair = float(sum(2.0 * a * a * c + 2.0 * a * c * c for ia, a in enumerate(npe) for ic, c in enumerate(npe) if ia != ic))
    ang_pair = 2.0 * cnt[9] * np2 * st2  # right angular contraction per element pair
    f_same = (per_same * st2 * items + ang_pair * nel)
    f_pair = (per_pair * st2 * items + ang_pair * nel * (nel - 1))
    w_bytes = float(cnt[4]) * stack * cnt[5] * 8
    blocks_bytes = float(items) * nel * nel * np2 * 8
    stages = []

    def add(name, kernel, msv, bound, flops=None, nbytes=None):
        e = {"stage": name, "kernel": kernel, "ms": msv, "bound": bound}
        if flops is not None:
            e["algorithmic_flops"] = flops
            e["tflops"] = flops / (msv * 1e-3) / 1e12 if msv > 0 else None
            e["frac_fp64"] = e["tflops"] / FP64_TFLOPS_MEASURED if e["tflops"] else None
        if nbytes is not None:
            e["algorithmic_bytes"] = nbytes
            e["gbs"] = nbytes / (msv * 1e-3) / 1e9 if msv > 0 else None
            e["frac_hbm"] = e["gbs"] / hbm_peak if e["gbs"] else None
        tr = ncu_traffic(args.workload, kernel)
        if isinstance(tr, dict):  # bytes of ONE captured launch (a build has one launch per family batch) and that launch's algorithmic bytes
            e["traffic"] = tr["bytes"]
            e["traffic_launch"] = tr["launch"]
            e["traffic_over_algorithmic"] = tr["bytes"] / tr["algorithmic"]
        else:
            e["traffic"] = tr
        stages.append(e)

    add("mask", "chan_x0_mask_kernel", stage[0], "hbm", nbytes=float(n) * n * 8)
    add("x1_left_angular", "chan_x1_mma_kernel", stage[1], "hbm", nbytes=w_bytes + float(np.count_nonzero(np.asarray(b.mval)[:, None] == np.asarray(b.mval)[None, :])) * nrad * nrad * 8)
    add("xf_same_element", "chan_xf_same_kernel", stage[2], "tensor", flops=f_same, nbytes=w_bytes / nel + blocks_bytes / nel)
    add("xf_element_pairs", "chan_xf_pair_kernel", stage[3], "tensor", flops=f_pair, nbytes=w_bytes * (nel - 1) / nel + blocks_bytes * (nel - 1) / nel)
    add("x4_assemble", "chan_x4_smem_kernel" if nrad * (nrad + 1) * 8 <= 100 * 1024 else "chan_x4_kernel", stage[4], "hbm", nbytes=blocks_bytes + float(cnt[6]) * nrad * nrad * 8 * (2 if cnt[8] else 1))
    add("coulomb (pack, pre, DMMA, post, unpack)", "chan_pre_kernel", ms_j, "hbm", nbytes=2.0 * n * n * 8)
    if sharded:
        stages.append({"stage": "nccl_allreduce_K (gather + all-reduce + scatter)", "ms": ms_ar, "payload_bytes": float(cnt[10]) * nrad * nrad * 8})
    dom = max(stages[:5], key=lambda e: e["ms"])
    t_dmma = stage[2] + stage[3]
    tf = (f_same + f_pair) / (t_dmma * 1e-3) / 1e12 if t_dmma > 0 else None
    roofline = {"kernel": dom["kernel"], "bound": dom["bound"],
                "achieved": dom.get("tflops") if dom["bound"] == "tensor" else dom.get("gbs"),
                "peak": FP64_TFLOPS_MEASURED if dom["bound"] == "tensor" else hbm_peak, "unit": "TFLOP/s" if dom["bound"] == "tensor" else "GB/s",
                "frac": dom.get("frac_fp64") if dom["bound"] == "tensor" else dom.get("frac_hbm"), "traffic": dom["traffic"],
                "traffic_note": ("ncu dram bytes of ONE launch of the kernel (%s); its algorithmic bytes are %.3g x smaller; achieved / ms_per_launch_sum "
                                 "cover all launches of the build" % (dom.get("traffic_launch"), dom.get("traffic_over_algorithmic", float("nan")))) if dom.get("traffic_launch") else None,
                "peak_source": "FP64: measured cuBLAS DGEMM (profiles/r01_fp64_peak.json; MEASURED_PEAKS.json has no FP64 entry; tcgen05 has no FP64 kind, "
                               "DMMA.8x8x4 is the FP64 tensor path); HBM: " + hbm_src,
                "ms_per_launch_sum": dom["ms"], "stages": stages,
                "fused_dmma_tflops": tf, "fused_dmma_frac": tf / FP64_TFLOPS_MEASURED if tf else None,
                "whole_build_tflops": (f_same + f_pair) / (ms / args.steps * 1e-3) / 1e12,
                "whole_build_frac_fp64": (f_same + f_pair) / (ms / args.steps * 1e-3) / 1e12 / FP64_TFLOPS_MEASURED,
                "radial_flops_per_build": x3_flops, "ms_coulomb": ms_j, "ms_exchange": ms_k, "output_blocks": items,
                "symmetric_item_list": bool(cnt[8]), "family_batches": b.exchange_batches(), "w_bytes": w_bytes}
    # our kernels per build (ncu launch list, profiles/r02w_launches_n2_hf_cfg4.csv): Coulomb pack, pre, jmma, flag, post, unpack; the mask; x1, fused
    # same-element, fused pairs, x4 per family batch; sharded: + mask-flag conversion, block gather, block scatter (NCCL's own kernels not counted)
    nlaunch = 6 + 1 + 4 * b.exchange_batches() + (3 if sharded else 0)
    line = {"metric": "fock_builds_per_sec", "value": value, "unit": "builds/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong" if sharded else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload, "desc": w["desc"], "operators": "coulomb + exchange (HF)", "ndim": n, "nbf": nb,
                       "nang": int(b.nang), "nrad": int(b.nrad),
                       "l2": "dense P/J/K = %.0f MB per build, W = %.1f GB%s" % (3 * n * n * 8 / 1e6, w_bytes / 1e9, "" if 3 * n * n * 8 > 126e6 else " (fits L2: launch-bound regime)"),
                       "partition": ("one build by all ranks: integral families r %% %d == rank, NCCL all-reduce of J (channel results) and K (packed blocks) "
                                     "inside the library and inside the timed region" % world if sharded
                                     else "independent builds per rank, no data-path collective" if world > 1 else "one GPU"),
                       "setup_s": {"compute_tei_host": t_tei, "upload": t_upload, "first_build": t_first}},
            "roofline": roofline, "cpu_baseline": cpu,
            "e2e": {"value": units / t_e2e, "unit": "builds/s", "h2d_bytes_per_step": 2 * nz_in * nrad * nrad * 8, "d2h_bytes_per_step": 2 * nz_out * nrad * nrad * 8,
                    "host_matrix_bytes_per_step": 4 * nb * nb * 8,
                    "api": "TwoDBasis::coulomb + TwoDBasis::exchange on dense host matrices (pure basis, caller-owned outputs)" + (", called by rank 0 alone, the other ranks serve (root-driven communicator: the blocks of P are broadcast over NCCL)" if driven else "") + "; only the "
                           "non-zero angular blocks of P and the blocks of J / K that can be non-zero cross PCIe (packed), the dense host "
                           "matrices are scanned / zero-filled on the host inside the timed region"},
            "gpu_launches": int(args.steps * nlaunch), "clocks": clocks}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def sweep_units(b, w):
    """Units of the Z = 1..86 sweep (BASELINE configs[4]) on ONE shared radial basis (the basis does not depend on Z): for every
    atom the trial configurations of a gensap-style occupation search (tests/util.trial_configurations), each a mock SCF of
    niter(Z) Fock builds.  Returns (units, bands): units[u] = dict(Z, occ, niter, first build, ground), bands [nbuilds][2][nsl][blk]
    of spin densities / 4 pi as FockBuild passes them (sadatom/solver.cpp:784,804), heavy units first."""
    import scipy.linalg
    from util import trial_configurations
    from helfem_b200 import shard
    lmax, nsl = w["lmax"], w["lmax"] + 1
    S, T, Tl, V1 = b.overlap(), b.kinetic(), b.kinetic_l(), b.nuclear()  # the basis was built with Z = 1: V(Z) = Z V(1)
    blk = b.nel * b.np_max ** 2
    units = []
    for Z in range(w["zmin"], w["zmax"] + 1):
        orb = [scipy.linalg.eigh(T + l * (l + 1) * Tl + Z * V1, S)[1] for l in range(nsl)]
        for ic, occ in enumerate(trial_configurations(Z, lmax)):
            units.append(dict(Z=Z, occ=occ, niter=8 + Z // 8, ground=(ic == 0), orb=orb))
    order = shard.heavy_first([u["niter"] for u in units])
    units = [units[q] for q in order]
    nbuilds = sum(u["niter"] for u in units)
    bands = np.zeros((nbuilds, 2, nsl, blk))
    first = 0
    for u in units:
        u["first"] = first
        P = np.zeros((2, nsl, blk))
        for si, spin in enumerate("ab"):
            for l in range(nsl):
                cap = (2 * l + 1)
                numl = (u["occ"][l] + (1 if spin == "a" else 0)) // 2
                D = np.zeros_like(S)
                io = 0
                while numl > 0:  # sadatom's fractional shell filling (sadatom/solver.cpp:386-398)
                    k = min(cap, numl)
                    D += k * np.outer(u["orb"][l][:, io], u["orb"][l][:, io])
                    numl -= k
                    io += 1
                P[si, l] = b.pack_band(np.asfortranarray(D / (4 * np.pi)))
        for it in range(u["niter"]):  # the iterations of the mock SCF see slightly different densities
            bands[first + it] = P * (1.0 + 1e-3 * it)
        first += u["niter"]
        del u["orb"]
    return units, bands


def run_sweep(args, w, rank, world, local_rank):
    """sadatom sweep Z = 1..86 with PBE (BASELINE configs[4]): independent units, no data-path collective; the ranks draw chunks of
    units from one shared heavy-first queue (helfem_b200/shard.WorkQueue); rank 0 gathers Exc and Nel of every build."""
    import torch
    import helfem_b200 as hb
    from helfem_b200 import shard
    if not torch.cuda.is_available() or hb.device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: helfem_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    b = hb.SadatomBasis(Z=1, nelem=w["nelem"], nnodes=w["nnodes"], primbas=w["primbas"], taylor_order=w["taylor"], lmax=w["lmax"], device=local_rank)
    b.compute_tei(1, False)
    units, bands = sweep_units(b, w)
    nsl, blk = w["lmax"] + 1, b.nel * b.np_max ** 2
    nbuilds = bands.shape[0]
    hP = torch.from_numpy(bands).pin_memory()
    dP = hP.cuda()
    dJ = torch.zeros((nbuilds, blk), dtype=torch.float64, device="cuda")
    dH = torch.zeros((nbuilds, 2, nsl, blk), dtype=torch.float64, device="cuda")
    dS = torch.zeros((nbuilds, 4), dtype=torch.float64, device="cuda")
    hJ = torch.zeros((nbuilds, blk), dtype=torch.float64).pin_memory()
    hH = torch.zeros((nbuilds, 2, nsl, blk), dtype=torch.float64).pin_memory()
    hS = torch.zeros((nbuilds, 4), dtype=torch.float64).pin_memory()
    stream = torch.cuda.current_stream()
    sh = stream.cuda_stream
    x, c, thr = w["x"], w["c"], 1e-12
    chunk = max(1, len(units) // (4 * world))
    queue = shard.WorkQueue(list(range(len(units))), chunk=chunk)
    perP, perH = 2 * nsl * blk, 2 * nsl * blk

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def one_pass(host):
        """Every rank drains the shared queue once; returns the builds this rank did."""
        barrier()
        queue.reset()
        barrier()
        done = 0
        while True:
            ids = queue.claim()
            if not ids:
                break
            f0 = units[ids[0]]["first"]
            nb = sum(units[u]["niter"] for u in ids)  # consecutive queue entries are consecutive in memory
            if host:
                P_, J_, H_, S_ = (t.numpy().reshape(nbuilds, -1)[f0:f0 + nb] for t in (hP, hJ, hH, hS))
                tq = time.perf_counter()
                b.fock_host(nb, x, c, 2, nsl, thr, P_, J_, H_, S_)
                if os.environ.get("HFB200_SWEEP_DIAG"):
                    dt = time.perf_counter() - tq
                    sys.stderr.write("[sweep diag] rank %d fock_host nb %d: %.2f ms, %.1f GB/s (H2D + D2H)\n" % (rank, nb, 1e3 * dt, (P_.nbytes + J_.nbytes + H_.nbytes) / dt / 1e9))
            else:
                b.fock_batch(nb, x, c, 2, nsl, thr, dP.data_ptr() + f0 * perP * 8, dJ.data_ptr() + f0 * blk * 8, dH.data_ptr() + f0 * perH * 8,
                             dS.data_ptr() + f0 * 32, sh)
            done += nb
        torch.cuda.synchronize()
        return done

    for _ in range(max(args.warmup, 3)):
        one_pass(False)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    t0 = time.perf_counter()
    mine = 0
    for _ in range(args.steps):
        mine += one_pass(False)
    barrier()
    t_dev = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    one_pass(True)
    barrier()
    t0 = time.perf_counter()
    ne2e = max(1, args.steps // 2)
    for _ in range(ne2e):
        one_pass(True)
    barrier()
    t_e2e = (time.perf_counter() - t0) / ne2e
    share = [mine]
    if dist is not None:
        t = torch.tensor([t_dev, t_e2e], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_dev, t_e2e = float(t[0]), float(t[1])
        share = [None] * world
        dist.all_gather_object(share, mine)
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    # rank 0: one clean single-rank pass for the gathered table (Exc, Nel of every atom's ground configuration)
    dS.zero_()
    b.fock_batch(nbuilds, x, c, 2, nsl, thr, dP.data_ptr(), dJ.data_ptr(), dH.data_ptr(), dS.data_ptr(), sh)
    torch.cuda.synchronize()
    sc = dS.cpu().numpy()
    table = {u["Z"]: (float(sc[u["first"], 0]), float(sc[u["first"], 1])) for u in units if u["ground"]}
    max_dnel = max(abs(table[Z][1] - Z) for Z in table)
    hbm_peak, hbm_src = peaks()
    value = nbuilds * args.steps / t_dev
    line = {"metric": "fock_builds_per_sec", "value": value, "unit": "builds/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": 1e3 * t_dev / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload, "desc": w["desc"], "operators": "coulomb + eval_Fxc (unrestricted PBE)", "atoms": len(table),
                       "units": len(units), "builds_per_step": nbuilds, "queue_chunk_units": chunk,
                       "partition": "one shared heavy-first queue of (atom, trial configuration) units over the ranks (TCP store counter), no data-path "
                                    "collective; rank 0 gathers Exc / Nel", "builds_per_rank": share,
                       "timing": "wall clock around the drained queue (host claims + launches + device), max over ranks",
                       "l2": "P + J + H bands = %.0f MB per step" % ((bands.nbytes * 2 + nbuilds * blk * 8) / 1e6)},
            "roofline": None, "cpu_baseline": None,
            "e2e": {"value": nbuilds / t_e2e, "unit": "builds/s", "h2d_bytes_per_step": int(bands.nbytes), "d2h_bytes_per_step": int(bands.nbytes + nbuilds * (blk + 4) * 8),
                    "api": "hfb200_sadatom_fock_host per claimed chunk (pinned host bands)"},
            "checks": {"max_abs_Nel_minus_Z": max_dnel, "Exc_H": table[min(table)][0], "Exc_last": table[max(table)][0]},
            "gpu_launches": int(args.steps * 3 * ((len(units) + chunk - 1) // chunk)), "clocks": clocks}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def run_grid2d(args, w, rank, world, local_rank):
    """DFTGrid::eval_Fxc of the diatomic program through the host mirror class (dense host matrices in and out)."""
    import torch
    import helfem_b200 as hb
    if not torch.cuda.is_available() or hb.device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: helfem_b200 has no CPU fallback")
    from grid2d_cases import diatomic_density
    b = hb.DiatomicBasis(w["Z1"], w["Z2"], w["Rbond"], nelem=w["nelem"], nnodes=w["nnodes"], primbas=w["primbas"],
                         lmax=w["lmax"], mmax=w["mmax"], device=local_rank)
    lang, mang = 4 * w["lmax"] + 12, 4 * (w["mmax"] + 1) + 5  # diatomic/main.cpp:417,425 (lmmax.n_elem = mmax + 1)
    g = hb.DiatomicDFTGrid(b, lang, mang)
    n = b.nbf
    C = chan_orbitals(b, "diatomic", rank)
    pure = b.pure_indices()
    P = np.asfortranarray(2.0 * C[pure] @ C[pure].T)
    H = np.zeros((n, n), order="F")   # caller-owned result, overwritten by every call
    for _ in range(max(args.warmup, 3)):
        g.eval_Fxc(w["x"], w["c"], P, out=H)
    sampler = ClockSampler(local_rank)
    sampler.start()
    dev_ms = 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        _, exc, nel, ekin = g.eval_Fxc(w["x"], w["c"], P, out=H)
        dev_ms += g.last_call()["device_ms"]
    t = (time.perf_counter() - t0) / args.steps
    dev_ms /= args.steps
    acct = g.last_call()
    clocks = sampler.stop()
    if args.kernels_only:
        print(json.dumps({"kernels_only": True, "ms_per_step": 1e3 * t, "device_ms": dev_ms}), flush=True)
        return
    cpu = None
    if n <= 400:
        import oracle
        oracle.lib().oracle_set_num_threads(host_threads())
        tabs = [b.radial(e) for e in range(b.nel)]
        o = oracle.Grid2DOracle(1, b.lval, b.mval, b.first, b.last, [q[0] for q in tabs], [q[1] for q in tabs],
                                [q[2] for q in tabs], [q[3] for q in tabs], lang, mang, Rhalf=b.Rhalf, pure=pure)
        t0 = time.perf_counter()
        Ho, exco, nelo, ekino = o.eval_fxc(w["x"], w["c"], P)
        tc = time.perf_counter() - t0
        cpu = {"value": 1.0 / tc, "unit": "builds/s", "cores": oracle.lib().oracle_num_threads(), "kind": "port",
               "sample": "1 eval_Fxc of the same workload", "max_dH": float(np.max(np.abs(H - Ho))), "dExc": float(abs(exc - exco))}
    line = {"metric": "fock_builds_per_sec", "value": 1e3 / dev_ms, "unit": "builds/s", "n_gpus": 1, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload, "desc": w["desc"], "operators": "eval_Fxc (PBE) restricted, dense host matrices",
                       "nbf": n, "nang": int(b.nang), "nrad": int(b.nrad), "angular_points": lang * mang,
                       "timing": "value: CUDA events around the kernels of each call (packed blocks resident); e2e: the host call"},
            "roofline": None, "roofline_note": "seven short kernels (%d launches, %.2f ms together); per-kernel times in the committed launch list "
                                               "(profiles/), no single dominant kernel to quote a roofline for" % (acct["launches"], dev_ms),
            "cpu_baseline": cpu,
            "e2e": {"value": 1.0 / t, "unit": "builds/s", "ms_per_call": 1e3 * t, "h2d_bytes_per_step": acct["h2d_bytes"],
                    "d2h_bytes_per_step": acct["d2h_bytes"], "host_matrix_bytes_per_step": 2 * n * n * 8,
                    "api": "diatomic::DFTGrid::eval_Fxc mirror on dense host matrices (caller-owned H): non-zero angular blocks of P down, "
                           "the blocks of H that can be non-zero back; host scan of P and zero-fill of H inside the timed region"},
            "gpu_launches": acct["launches"] * args.steps, "clocks": clocks, "Exc": exc, "Nel": nel}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="n2_hf_cfg4", choices=list(WORKLOADS))
    ap.add_argument("--batch", type=int, default=1024, help="independent Fock builds per step per GPU")
    ap.add_argument("--kernels-only", action="store_true", help="run only the timed device steps (ncu captures)")
    ap.add_argument("--split", action="store_true", help="with --kernels-only: also time the grid and the Coulomb kernels alone")
    ap.add_argument("--replicas", action="store_true", help="atomic/diatomic workloads, N>1: independent builds per rank instead of the "
                    "default (one build sharded over the ranks by integral family, NCCL all-reduce inside the library: strong scaling)")
    ap.add_argument("--shard", action="store_true", help="(default for N>1; kept for older scripts)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    w = WORKLOADS[args.workload]
    if w.get("kind") == "sweep":
        if args.impl == "reference":
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": "the CPU arm of the sweep is the kr_pbe_cfg2 reference arm (same basis, same operators)"}))
            return
        run_sweep(args, w, rank, world, local_rank)
        return
    if w.get("kind") == "grid2d":
        if rank == 0 and args.impl != "reference":
            run_grid2d(args, w, rank, world, local_rank)
        return
    if w.get("kind") in ("atomic", "diatomic"):
        if args.impl == "reference":
            run_chan_reference(args, w, rank)
            return
        run_chan(args, w, rank, world, local_rank)
    elif args.impl == "reference":
        run_reference(args, w, rank)
    else:
        run_gpu(args, w, rank, world, local_rank)


if __name__ == "__main__":
    main()
