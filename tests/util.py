"""Shared helpers of the test-suite: deterministic synthetic densities."""
import numpy as np
import scipy.linalg

SEED = 20261104


def splitmix64(n, seed=SEED):
    """n uniform doubles in (-1, 1) from splitmix64 (portable, no numpy RNG dependence)."""
    out = np.empty(n)
    x = seed & 0xFFFFFFFFFFFFFFFF
    mask = 0xFFFFFFFFFFFFFFFF
    for i in range(n):
        x = (x + 0x9E3779B97F4A7C15) & mask
        z = x
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & mask
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & mask
        z = z ^ (z >> 31)
        out[i] = (z >> 11) / float(1 << 53) * 2.0 - 1.0
    return out


# ground-state configurations by l-channel occupations (number of electrons in s, p, d, f)
def aufbau_occs(Z, lmax):
    order = [(1, 0), (2, 0), (2, 1), (3, 0), (3, 1), (4, 0), (3, 2), (4, 1), (5, 0), (4, 2), (5, 1), (6, 0), (4, 3), (5, 2),
             (6, 1), (7, 0), (5, 3), (6, 2), (7, 1)]
    occ = [0] * (lmax + 1)
    left = Z
    for n, l in order:
        if l > lmax or left <= 0:
            continue
        k = min(left, 2 * (2 * l + 1))
        occ[l] += k
        left -= k
    return occ


def sadatom_density(S, T, Tl, Vnuc, Z, lmax, perturb=1e-3, spin=None, seed=SEED, occs=None):
    """Density cube P_l (nbf x nbf x (lmax+1), Fortran order) from core-Hamiltonian orbitals filled by
    Aufbau with sadatom's fractional shell filling (sadatom/solver.cpp:386-398), plus a seeded symmetric
    perturbation so no block is accidentally zero.  spin: None (total), 'a' or 'b' (b has the last
    shell of the highest l emptied).  occs: electrons per l channel (default: Aufbau ground configuration)."""
    n = S.shape[0]
    P = np.zeros((n, n, lmax + 1), order="F")
    if occs is None:
        occs = aufbau_occs(Z, lmax)
    for l in range(lmax + 1):
        H = T + l * (l + 1) * Tl + Vnuc
        E, C = scipy.linalg.eigh(H, S)
        numl = occs[l]
        cap = 2 * (2 * l + 1)
        if spin is not None:
            cap //= 2
            numl = (numl + (1 if spin == "a" else 0)) // 2
        io = 0
        while numl > 0:
            k = min(cap, numl)
            P[:, :, l] += k * np.outer(C[:, io], C[:, io])
            numl -= k
            io += 1
    if perturb:
        # multiplicative, symmetric: every stored element moves by a relative O(perturb) amount, so no block
        # is accidentally special while the density stays positive in the exponential tails (an ABSOLUTE
        # perturbation swamps the tail density and drives meta-GGA potentials to 1e6 Eh there, where even the
        # CPU oracle only reproduces itself to 1e-6 under 1-ulp input noise)
        for l in range(lmax + 1):
            R = splitmix64(n * n, seed + l + (7 if spin == "b" else 0)).reshape(n, n)
            P[:, :, l] *= 1.0 + 0.5 * perturb * (R + R.T)
    return P


def xc_test_points(n=4096, seed=SEED):
    """Spin-resolved test points for the pointwise XC comparison: rho log-uniform over
    16 decades, reduced gradient s in [0, 50] incl. exact zeros, iso-orbital indicator alpha in [0, 100]
    incl. exact zeros, every third point spin-unpolarised.  Returns rho (n,2), sigma (n,3), tau (n,2)."""
    rng = np.random.RandomState(seed)
    rho = 10.0 ** rng.uniform(-13, 3, (n, 2))
    rho[::3, 1] = rho[::3, 0]
    kf = (3 * np.pi**2 * 2 * rho) ** (1 / 3)
    s = 10.0 ** rng.uniform(-3, 1.7, (n, 2))
    s[::7] = 0.0
    g = s * kf * 2 * rho
    sigma = np.stack([g[:, 0]**2, (2 * rng.uniform(size=n) - 1) * g[:, 0] * g[:, 1], g[:, 1]**2], axis=1)
    tw = g**2 / (8 * rho)
    tu = 0.3 * (6 * np.pi**2) ** (2 / 3) * rho ** (5 / 3)
    alpha = 10.0 ** rng.uniform(-3, 2, (n, 2))
    alpha[::5] = 0.0
    tau = tw + alpha * tu
    return rho, sigma, tau


def xc_pointwise_check(evaluate, oracle_xc, funcs=(1, 12, 101, 130, 497, 498), n=4096):
    """evaluate(func, nspin, n, rho, sigma, tau) -> (exc, vrho, vsigma, vtau) of ONE functional, compared with
    the oracle per functional (x and c separately: their sum cancels by ten orders of magnitude in v_sigma
    at low density, which would hide real errors behind an ill-conditioned total).

    Gate: relative 1e-9 (1e-7 where one spin channel holds < 1e-8 of the density: d/d zeta terms are
    ill-conditioned at |zeta| -> 1), absolute 1e-14 floor."""
    rho, sigma, tau = xc_test_points(n)
    worst = 0.0
    for f in funcs:
        for nspin in (1, 2):
            if nspin == 2:
                r_, s_, t_ = rho.ravel().copy(), sigma.ravel().copy(), tau.ravel().copy()
                extreme = np.minimum(rho[:, 0], rho[:, 1]) < 1e-8 * (rho[:, 0] + rho[:, 1])
            else:
                r_, s_, t_ = (2 * rho[:, 0]).copy(), (4 * sigma[:, 0]).copy(), (2 * tau[:, 0]).copy()
                extreme = np.zeros(n, dtype=bool)
            got = evaluate(f, nspin, n, r_, s_, t_)
            want = oracle_xc(f, r_, s_, t_, polarized=(nspin == 2))
            for name, g_, w_ in zip(("exc", "vrho", "vsigma", "vtau"), got, want):
                per = g_.size // n
                rel = np.where(np.repeat(extreme, per), 1e-7, 1e-9)
                bad = np.abs(g_ - w_) > rel * np.abs(w_) + 1e-14
                assert not bad.any(), (f, nspin, name, int(bad.sum()), g_[bad][:3], w_[bad][:3])
                m = np.abs(w_) > 1e-200
                if m.any():
                    worst = max(worst, float(np.max(np.abs(g_ - w_)[m] / np.abs(w_)[m])))
    return worst


# ---- diatomic program: overlap-normalised synthetic orbitals (full-size parity tests, bench) ----
def _cos2_matrix(lval, mval):
    """<Y_a| cos^2(theta) |Y_b> for the angular list: closed forms of the standard recursion
    cos(t) Y_lm = A_{l+1,m} Y_{l+1,m} + A_{l,m} Y_{l-1,m},  A_{l,m} = sqrt((l^2 - m^2) / (4 l^2 - 1))."""
    def A(l, m):
        return np.sqrt(max(l * l - m * m, 0) / (4.0 * l * l - 1.0)) if l > 0 else 0.0
    n = len(lval)
    c2 = np.zeros((n, n))
    for a in range(n):
        for b in range(n):
            if mval[a] != mval[b]:
                continue
            l, lp, m = int(lval[a]), int(lval[b]), int(mval[a])
            if l == lp:
                c2[a, b] = A(l + 1, m) ** 2 + A(l, m) ** 2
            elif l == lp + 2:
                c2[a, b] = A(l, m) * A(l - 1, m)
            elif lp == l + 2:
                c2[a, b] = A(lp, m) * A(lp - 1, m)
    return c2


def diatomic_overlap_apply(b, C):
    """S C for coefficient arrays C (nang, nrad, norb) in the dummy space of a DiatomicBasis harness object:
    S = Rh^3 [ delta_ab I2 - <a|cos^2|b> I0 ],  I_k = int B_i B_j sinh(mu) cosh^k(mu) dmu  (the prolate volume element
    Rh^3 sinh(mu) (cosh^2(mu) - cos^2(nu)); /root/reference/src/diatomic/basis.cpp overlap())."""
    nrad = b.nrad
    I0, I2 = np.zeros((nrad, nrad)), np.zeros((nrad, nrad))
    for e in range(b.nel):
        mu, w, bf, _ = b.radial(e)
        f, l = int(b.first[e]), int(b.last[e])
        I0[f:l + 1, f:l + 1] += bf.T @ (bf * (w * np.sinh(mu))[:, None])
        I2[f:l + 1, f:l + 1] += bf.T @ (bf * (w * np.sinh(mu) * np.cosh(mu) ** 2)[:, None])
    c2 = _cos2_matrix(b.lval, b.mval)
    return b.Rhalf ** 3 * (np.einsum("ij,ajo->aio", I2, C) - np.einsum("ab,ij,bjo->aio", c2, I0, C))


def diatomic_orbitals(b, occ_m=(0, 0, 0, 0, 0, 1, -1), seed=100, decay_l=0.35, decay_r=0.06):
    """Synthetic occupied orbitals of a linear molecule, each pure in m (N2-like by default: five sigma, pi+, pi-):
    smooth seeded coefficients decaying with l and the radial index, then Loewdin-ORTHONORMALISED in the overlap
    metric, so that P = 2 C C^T has tr(P S) = 2 norb like a converged closed-shell density and J, K have physical
    magnitudes.  Returns C (ndummy, norb); rows of the removed boundary functions are zero."""
    nrad, nang = b.nrad, b.nang
    cols = []
    for q, m in enumerate(occ_m):
        c = np.zeros((nang, nrad))
        sel = np.where(b.mval == m)[0]
        noise = splitmix64(len(sel) * nrad, seed=seed + q).reshape(len(sel), nrad)
        c[sel] = noise * np.exp(-decay_l * b.lval[sel])[:, None] * np.exp(-decay_r * np.arange(nrad))[None, :]
        if m != 0:
            c[sel, 0] = 0.0
        cols.append(c)
    C = np.stack(cols, axis=2)
    SC = diatomic_overlap_apply(b, C)
    G = np.einsum("aio,aip->op", C, SC)
    ev, U = np.linalg.eigh(G)
    C = C @ (U @ np.diag(ev ** -0.5) @ U.T)
    return C.reshape(nang * nrad, len(occ_m))


def trial_configurations(Z, lmax):
    """Occupations (electrons per l channel) a gensap-style configuration search visits for atom Z: the Aufbau ground
    configuration and its neighbours with one electron moved between adjacent l channels
    (/root/reference/src/sadatom/main.cpp:268-352 scans such occupation vectors).  Deterministic, ground state first."""
    base = aufbau_occs(Z, lmax)
    out = [tuple(base)]
    for l in range(lmax):
        for src, dst in ((l, l + 1), (l + 1, l)):
            if base[src] > 0:
                occ = list(base)
                occ[src] -= 1
                occ[dst] += 1
                if tuple(occ) not in out:
                    out.append(tuple(occ))
    return out
