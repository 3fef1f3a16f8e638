"""Generates tests/golden/xc_mpmath.npz: seeded spin-resolved test points (above every threshold, sigma > 0) and the libxc-style
outputs of the six functionals (+ lda_c_pw_mod) from the 50-digit mpmath restatement tests/xc_mpmath.py, polarised and
unpolarised.  Usage (authoring container, CPU, a few minutes):  python tests/golden/make_xc_mpmath_goldens.py [npoints]"""
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


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    rho, sigma, tau = points(n)
    out = {"rho": rho, "sigma": sigma, "tau": tau}
    for fid in FIDS:
        pe, pvr, pvs, pvt = np.zeros(n), np.zeros((n, 2)), np.zeros((n, 3)), np.zeros((n, 2))
        ue, uvr, uvs, uvt = np.zeros(n), np.zeros(n), np.zeros(n), np.zeros(n)
        for i in range(n):
            e, vr, vs, vt = xm.evaluate_polarised(fid, rho[i, 0], rho[i, 1], sigma[i, 0], sigma[i, 1], sigma[i, 2], tau[i, 0], tau[i, 1])
            pe[i], pvr[i], pvs[i], pvt[i] = e, vr, vs, vt
            ue[i], uvr[i], uvs[i], uvt[i] = xm.evaluate_unpolarised(fid, 2 * rho[i, 0], 4 * sigma[i, 0], 2 * tau[i, 0])
        out.update({"p%d_exc" % fid: pe, "p%d_vrho" % fid: pvr, "p%d_vsigma" % fid: pvs, "p%d_vtau" % fid: pvt,
                    "u%d_exc" % fid: ue, "u%d_vrho" % fid: uvr, "u%d_vsigma" % fid: uvs, "u%d_vtau" % fid: uvt})
        print("functional", fid, "done", flush=True)
    np.savez_compressed(os.path.join(HERE, "xc_mpmath.npz"), **out)


if __name__ == "__main__":
    main()
