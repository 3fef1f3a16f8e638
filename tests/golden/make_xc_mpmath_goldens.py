"""Generates tests/golden/xc_mpmath.npz: seeded spin-resolved test points (above every threshold, sigma > 0) and the libxc-style
outputs of the six functionals (+ lda_c_pw_mod) from the 50-digit mpmath restatement tests/xc_mpmath.py, polarised and
unpolarised.  Usage (authoring container, CPU, all cores; the committed fixture has 1024 points, 8192 take ~3 minutes on 8 cores):
    python tests/golden/make_xc_mpmath_goldens.py 1024"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import xc_mpmath as xm  # noqa: E402

FIDS = (1, 12, 13, 101, 130, 497, 498)


def points(n, seed=20261104):
    """rho log-uniform over 9 decades, reduced gradient s in [1e-3, 30], alpha in [1e-3, 100] plus exact alpha = 0 and alpha = 1
    points, zeta from 0 to 1 - 2e-4; every fourth point exactly unpolarised."""
    rng = np.random.RandomState(seed)
    rho = 10.0 ** rng.uniform(-6, 3, (n, 2))
    ratio = 10.0 ** rng.uniform(-4, 0, n)  # the minority density stays two decades above the 1e-12 density threshold
    rho[:, 1] = rho[:, 0] * ratio
    rho[::4, 1] = rho[::4, 0]
    kf = (3 * np.pi**2 * 2 * rho) ** (1 / 3)
    s = 10.0 ** rng.uniform(-3, 1.5, (n, 2))
    s[::4, 1] = s[::4, 0]
    g = s * kf * 2 * rho
    cosab = 2 * rng.uniform(size=n) - 1
    cosab[::4] = 1.0
    sigma = np.stack([g[:, 0] ** 2, cosab * g[:, 0] * g[:, 1], g[:, 1] ** 2], axis=1)
    tw = g**2 / (8 * rho)
    tu = 0.3 * (6 * np.pi**2) ** (2 / 3) * rho ** (5 / 3)
    alpha = 10.0 ** rng.uniform(-3, 2, (n, 2))
    alpha[1::8] = 0.0
    alpha[2::8] = 1.0
    alpha[::4, 1] = alpha[::4, 0]
    tau = tw + alpha * tu
    return rho, sigma, tau


def _one(args):
    fid, r, sg, t = args
    e, vr, vs, vt = xm.evaluate_polarised(fid, r[0], r[1], sg[0], sg[1], sg[2], t[0], t[1])
    u = xm.evaluate_unpolarised(fid, 2 * r[0], 4 * sg[0], 2 * t[0])
    return (float(e), [float(v) for v in vr], [float(v) for v in vs], [float(v) for v in vt]) + tuple(float(v) for v in u)


def main():
    import multiprocessing as mp
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    rho, sigma, tau = points(n)
    out = {"rho": rho, "sigma": sigma, "tau": tau}
    with mp.Pool() as pool:
        for fid in FIDS:
            res = pool.map(_one, [(fid, rho[i], sigma[i], tau[i]) for i in range(n)], chunksize=16)
            out.update({"p%d_exc" % fid: np.array([r[0] for r in res]), "p%d_vrho" % fid: np.array([r[1] for r in res]),
                        "p%d_vsigma" % fid: np.array([r[2] for r in res]), "p%d_vtau" % fid: np.array([r[3] for r in res]),
                        "u%d_exc" % fid: np.array([r[4] for r in res]), "u%d_vrho" % fid: np.array([r[5] for r in res]),
                        "u%d_vsigma" % fid: np.array([r[6] for r in res]), "u%d_vtau" % fid: np.array([r[7] for r in res])})
            print("functional", fid, "done", flush=True)
    np.savez_compressed(os.path.join(HERE, "xc_mpmath.npz"), **out)


if __name__ == "__main__":
    main()
