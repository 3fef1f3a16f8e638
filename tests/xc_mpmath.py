"""TEST INFRASTRUCTURE: 110-digit mpmath restatement of the six functionals the device implements, written from the PUBLISHED
definitions in the papers' own variables (r_s, zeta, s, t, alpha) and differentiated NUMERICALLY (central differences at 110 digits), i.e.
independent of both the product's dual numbers and the C++ oracle's complex step:

  lda_x          Dirac / Slater exchange, spin-scaled
  lda_c_pw(_mod) Perdew, Wang, Phys. Rev. B 45, 13244 (1992), Table I ("_mod": more digits of the parameters and the exact f''(0))
  gga_x_pbe / gga_c_pbe   Perdew, Burke, Ernzerhof, Phys. Rev. Lett. 77, 3865 (1996)
  mgga_x_r2scan / mgga_c_r2scan   Furness, Kaplan, Ning, Perdew, Sun, J. Phys. Chem. Lett. 11, 8208 (2020) (+ SCAN: Sun, Ruzsinszky,
                 Perdew, Phys. Rev. Lett. 115, 036402 (2015) for the pieces r2SCAN inherits)

Constants with a closed form are DERIVED here (gamma = (1 - ln 2)/pi^2, mu = beta pi^2/3, f''(0) = 8/(9 (2^{4/3} - 2)), chi_inf from
its definition, C_2x from the polynomial, ...) instead of being typed; the fitted ones (PW92 table, b_1c..b_3c, the two degree-7
interpolation polynomials) are the papers' numbers and are additionally pinned by their smoothness conditions in
tests/test_xc_pins.py.  Conventions follow libxc (what the reference calls, /root/reference/src/sadatom/dftgrid.cpp:349-398):
inputs rho_s, sigma_ss', tau_s; outputs exc per particle and d(n exc)/d(input).  No thresholds here: only use above them.
libxc itself is not available in this image or on the GPU box (probed: no pylibxc, no libxc.so, no xc.h), so this is the
strongest pin the box allows; upstream parity stays "unpinned" (the reference does not pin libxc's version either)."""
import mpmath as mp

mp.mp.dps = 110  # the r2SCAN correlation nests one numerical derivative (d eps/d rs) inside the outer ones: both need head-room

PI = mp.pi
GAMMA = (1 - mp.log(2)) / PI**2
BETA_PBE = mp.mpf("0.06672455060314922")
KAPPA = mp.mpf("0.804")
MU_PBE = BETA_PBE * PI**2 / 3
CBRT_3PI2 = (3 * PI**2) ** (mp.mpf(1) / 3)
FZ20_EXACT = mp.mpf(8) / (9 * (2 ** (mp.mpf(4) / 3) - 2))

# PW92 Table I: (A, alpha1, beta1, beta2, beta3, beta4) for eps_c(rs,0), eps_c(rs,1), -alpha_c(rs)
PW_TAB = {
    False: [("0.031091", "0.21370", "7.5957", "3.5876", "1.6382", "0.49294"),
            ("0.015545", "0.20548", "14.1189", "6.1977", "3.3662", "0.62517"),
            ("0.016887", "0.11125", "10.357", "3.6231", "0.88026", "0.49671")],
    True: [("0.0310907", "0.21370", "7.5957", "3.5876", "1.6382", "0.49294"),
           ("0.01554535", "0.20548", "14.1189", "6.1977", "3.3662", "0.62517"),
           ("0.0168869", "0.11125", "10.357", "3.6231", "0.88026", "0.49671")],
}


def pw_G(rs, row):
    A, a1, b1, b2, b3, b4 = (mp.mpf(x) for x in row)
    q = 2 * A * (b1 * mp.sqrt(rs) + b2 * rs + b3 * rs ** mp.mpf("1.5") + b4 * rs**2)
    return -2 * A * (1 + a1 * rs) * mp.log(1 + 1 / q)


def f_zeta(z):
    return ((1 + z) ** (mp.mpf(4) / 3) + (1 - z) ** (mp.mpf(4) / 3) - 2) / (2 ** (mp.mpf(4) / 3) - 2)


def eps_pw92(rs, z, mod):
    e0, e1, mac = (pw_G(rs, r) for r in PW_TAB[mod])  # mac = -alpha_c
    fz20 = FZ20_EXACT if mod else mp.mpf("1.709921")
    fz = f_zeta(z)
    return e0 - mac * fz / fz20 * (1 - z**4) + (e1 - e0) * fz * z**4


def rs_of(n):
    return (3 / (4 * PI * n)) ** (mp.mpf(1) / 3)


# ---------------- exchange: E_x[na, nb] = (E_x[2 na] + E_x[2 nb]) / 2 ----------------
R2_ETA, R2_K1, R2_DP2, R2_A1, R2_H0X = mp.mpf("0.001"), mp.mpf("0.065"), mp.mpf("0.361"), mp.mpf("4.9479"), mp.mpf("1.174")
R2_C1X, R2_C2X, R2_DX = mp.mpf("0.667"), mp.mpf("0.8"), mp.mpf("1.24")
R2_CX = [mp.mpf(x) for x in ("1", "-0.667", "-0.4445555", "-0.663086601049", "1.451297044490", "-0.887998041597", "0.234528941479", "-0.023185843322")]
R2_C1C, R2_C2C, R2_DC = mp.mpf("0.64"), mp.mpf("1.5"), mp.mpf("0.7")
R2_CC = [mp.mpf(x) for x in ("1", "-0.64", "-0.4352", "-1.535685604549", "3.061560252175", "-1.915710236206", "0.516884468372", "-0.051848879792")]
R2_B1C, R2_B2C, R2_B3C = mp.mpf("0.0285764"), mp.mpf("0.0889"), mp.mpf("0.125541")


def f_alpha(a, c1, c2, d, poly):
    if a <= 0:
        return mp.exp(-c1 * a / (1 - a))
    if a <= mp.mpf("2.5"):
        return sum(c * a**i for i, c in enumerate(poly))
    return -d * mp.exp(c2 / (1 - a))


def ex_unpolarised(kind, n, sigma, tau):
    """Exchange energy per volume of a spin-UNPOLARISED density n with |grad n|^2 = sigma and kinetic density tau."""
    kf = CBRT_3PI2 * n ** (mp.mpf(1) / 3)
    ex_unif = -3 * kf * n / (4 * PI)
    if kind == "lda":
        return ex_unif
    p = sigma / (2 * kf * n) ** 2  # s^2
    if kind == "pbe":
        return ex_unif * (1 + KAPPA - KAPPA / (1 + MU_PBE * p / KAPPA))
    tau_w = sigma / (8 * n)
    tau_u = mp.mpf(3) / 10 * CBRT_3PI2**2 * n ** (mp.mpf(5) / 3)
    alpha = (tau - tau_w) / (tau_u + R2_ETA * tau_w)
    dfx1 = sum(i * c for i, c in enumerate(R2_CX))          # f_x'(1) of the interpolation polynomial
    c2x_big = -dfx1 * (1 - R2_H0X)                           # C_2x
    c_eta = mp.mpf(20) / 27 + R2_ETA * 5 / 3
    x = (c_eta * c2x_big * mp.exp(-p**2 / R2_DP2**4) + mp.mpf(10) / 81) * p
    h1 = 1 + R2_K1 - R2_K1 / (1 + x / R2_K1)
    gx = 1 - mp.exp(-R2_A1 / p ** (mp.mpf(1) / 4)) if p > 0 else mp.mpf(1)
    fa = f_alpha(alpha, R2_C1X, R2_C2X, R2_DX, R2_CX)
    return ex_unif * (h1 + fa * (R2_H0X - h1)) * gx


def energy_x(kind, na, nb, saa, sab, sbb, ta, tb):
    return (ex_unpolarised(kind, 2 * na, 4 * saa, 2 * ta) + ex_unpolarised(kind, 2 * nb, 4 * sbb, 2 * tb)) / 2


# ---------------- correlation ----------------
def energy_c(kind, na, nb, saa, sab, sbb, ta, tb):
    n = na + nb
    z = (na - nb) / n
    rs = rs_of(n)
    if kind in ("pw", "pw_mod"):
        return n * eps_pw92(rs, z, kind == "pw_mod")
    sigma, tau = saa + 2 * sab + sbb, ta + tb
    phi = ((1 + z) ** (mp.mpf(2) / 3) + (1 - z) ** (mp.mpf(2) / 3)) / 2
    kf = CBRT_3PI2 * n ** (mp.mpf(1) / 3)
    ks = mp.sqrt(4 * kf / PI)
    t2 = sigma / (2 * phi * ks * n) ** 2
    eps = eps_pw92(rs, z, True)
    if kind == "pbe":
        A = BETA_PBE / GAMMA / (mp.exp(-eps / (GAMMA * phi**3)) - 1)
        H = GAMMA * phi**3 * mp.log(1 + BETA_PBE / GAMMA * t2 * (1 + A * t2) / (1 + A * t2 + A**2 * t2**2))
        return n * (eps + H)
    # r2SCAN
    p = sigma / (2 * kf * n) ** 2
    ds = ((1 + z) ** (mp.mpf(5) / 3) + (1 - z) ** (mp.mpf(5) / 3)) / 2
    dx = ((1 + z) ** (mp.mpf(4) / 3) + (1 - z) ** (mp.mpf(4) / 3)) / 2
    gc = (1 - mp.mpf("2.3631") * (dx - 1)) * (1 - z**12)
    tau_w = sigma / (8 * n)
    tau_u = mp.mpf(3) / 10 * CBRT_3PI2**2 * n ** (mp.mpf(5) / 3) * ds
    alpha = (tau - tau_w) / (tau_u + R2_ETA * tau_w)
    # eps_c^0 (single-orbital limit)
    eps0_lda = -R2_B1C / (1 + R2_B2C * mp.sqrt(rs) + R2_B3C * rs)
    cx0 = 3 / (4 * PI) * (9 * PI / 4) ** (mp.mpf(1) / 3)
    beta_inf = mp.mpf("0.066725") * mp.mpf("0.1") / mp.mpf("0.1778")
    chi_inf = (3 * PI**2 / 16) ** (mp.mpf(2) / 3) * beta_inf / (mp.mpf("0.9") - cx0)
    w0 = mp.exp(-eps0_lda / R2_B1C) - 1
    ginf = (1 + 4 * chi_inf * p) ** (-mp.mpf(1) / 4)
    eps0 = (eps0_lda + R2_B1C * mp.log(1 + w0 * (1 - ginf))) * gc
    # eps_c^1 (PBE-like with beta(rs) and the r2SCAN correction Delta y)
    beta = mp.mpf("0.066725") * (1 + mp.mpf("0.1") * rs) / (1 + mp.mpf("0.1778") * rs)
    w1 = mp.exp(-eps / (GAMMA * phi**3)) - 1
    y = beta * t2 / (GAMMA * w1)
    deps1_drs = _ddx_inner(lambda r: eps_pw92(r, z, True), rs, rs)
    deps0_drs = _ddx_inner(lambda r: -R2_B1C / (1 + R2_B2C * mp.sqrt(r) + R2_B3C * r), rs, rs)
    dfc1 = sum(i * c for i, c in enumerate(R2_CC))  # f_c'(1)
    dy = dfc1 / (27 * GAMMA * ds * phi**3 * w1) * (20 * rs * (deps0_drs * gc - deps1_drs) - 45 * R2_ETA * (eps0_lda * gc - eps)) * p * mp.exp(-p**2 / R2_DP2**4)
    g1 = (1 + 4 * (y - dy)) ** (-mp.mpf(1) / 4)
    eps1 = eps + GAMMA * phi**3 * mp.log(1 + w1 * (1 - g1))
    fa = f_alpha(alpha, R2_C1C, R2_C2C, R2_DC, R2_CC)
    return n * (eps1 + fa * (eps0 - eps1))


FUNCS = {1: ("x", "lda"), 12: ("c", "pw"), 13: ("c", "pw_mod"), 101: ("x", "pbe"), 130: ("c", "pbe"), 497: ("x", "r2scan"), 498: ("c", "r2scan")}


def energy(fid, *v):
    side, kind = FUNCS[fid]
    return energy_x(kind, *v) if side == "x" else energy_c(kind, *v)


def _ddx(f, x, scale):
    """Central difference with a step RELATIVE to the variable's natural scale (mpmath.diff's default step is absolute, which
    is meaningless for sigma ~ 1e-40): error O(h^2) ~ 1e-48 relative, round-off far below at 110 digits."""
    h = abs(scale) * mp.mpf(2) ** -80
    return (f(x + h) - f(x - h)) / (2 * h)


def _ddx_inner(f, x, scale):
    """The nested derivative: a much smaller step, so that its round-off (2^-365 / 2^-150) survives the outer division by 2^-80."""
    h = abs(scale) * mp.mpf(2) ** -150
    return (f(x + h) - f(x - h)) / (2 * h)


def evaluate_polarised(fid, na, nb, saa, sab, sbb, ta, tb):
    """libxc-style outputs of the POLARISED call: exc, vrho[2], vsigma[3], vtau[2] (floats).  Needs sigma_aa, sigma_bb > 0."""
    v = [mp.mpf(x) for x in (na, nb, saa, sab, sbb, ta, tb)]
    scales = [v[0], v[1], v[2], mp.sqrt(v[2] * v[4]), v[4], v[5], v[6]]
    e = energy(fid, *v)
    grad = []
    for k in range(7):
        def f(x, k=k):
            w = list(v)
            w[k] = x
            return energy(fid, *w)
        grad.append(_ddx(f, v[k], scales[k]))
    return float(e / (v[0] + v[1])), [float(g) for g in grad[0:2]], [float(g) for g in grad[2:5]], [float(g) for g in grad[5:7]]


def evaluate_unpolarised(fid, n, sigma, tau):
    """libxc-style outputs of the UNPOLARISED call (total density, |grad n|^2 > 0, total tau)."""
    n, sigma, tau = mp.mpf(n), mp.mpf(sigma), mp.mpf(tau)
    f = lambda nn, ss, tt: energy(fid, nn / 2, nn / 2, ss / 4, ss / 4, ss / 4, tt / 2, tt / 2)
    e = f(n, sigma, tau)
    return (float(e / n), float(_ddx(lambda x: f(x, sigma, tau), n, n)), float(_ddx(lambda x: f(n, x, tau), sigma, sigma)),
            float(_ddx(lambda x: f(n, sigma, x), tau, tau)))
