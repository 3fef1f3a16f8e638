"""ORACLE (test infrastructure -- never imported by the product): numpy restatement of the
reference's finite-element SETUP for the Fock-build path, written independently of
helfem_b200/host/fem.cpp and evaluated in extended precision (np.longdouble) where rounding
matters, then rounded to float64.

Reference functions followed (read-only tree /root/reference):
  chebyshev()            libhelfem/src/chebyshev.cpp:22-53
  lobatto_nodes()        libhelfem/src/lobatto.cpp:588-743 (tabulated for n<20, Newton otherwise)
  get_grid()             libhelfem/src/grid.cpp:18-87
  PolyBasis              libhelfem/src/LIPBasis.cpp, HIPBasis.cpp, LIPBasis_eval.cpp, HIPBasis_eval.cpp,
                         PolynomialBasis.cpp:175-179
  RadialSetup.idx        libhelfem/src/FiniteElementBasis.cpp:39-51,269-277
  RadialSetup.get_bf/df/lf   libhelfem/src/RadialBasis.cpp:575-730 (Taylor patch :59-133)
  radial_integral etc.   libhelfem/src/RadialBasis.cpp:190-199,319-345; FiniteElementBasis.cpp:387-415
  twoe_integral          libhelfem/src/quadrature.cpp:23-126
  exchange_tei           libhelfem/src/utils.cpp:130-156
Pinned by the reference's own known-answer tests in tests/test_oracle_setup.py
(Maple 4x4 table of src/atomic/inttest.cpp:58-98, S/T identities of src/atomic/main.cpp:412-440).
"""
import math

import numpy as np

LD = np.longdouble


def chebyshev(n):
    i = np.arange(1, n + 1, dtype=np.float64)
    oonpp = 1.0 / (n + 1.0)
    sine = np.sin(i * math.pi * oonpp)
    sinesq = sine * sine
    cosine = np.cos(i * math.pi * oonpp)
    w = 16.0 / 3.0 / (n + 1.0) * sinesq * sinesq
    x = 1.0 - 2.0 * i * oonpp + (2.0 / math.pi) * (1.0 + 2.0 / 3.0 * sinesq) * cosine * sine
    return x[::-1].copy(), w[::-1].copy()


def lobatto_nodes(n):
    """Roots of (1-x^2) P'_{n-1}(x), ascending, via Newton on the Legendre recurrence (longdouble)."""
    x = np.cos(LD(math.pi) * np.arange(n, dtype=LD) / (n - 1))
    for _ in range(200):
        p0 = np.ones(n, dtype=LD)
        p1 = x.copy()
        for j in range(2, n):
            p0, p1 = p1, ((2 * j - 1) * x * p1 - (j - 1) * p0) / j
        xn = x - (x * p1 - p0) / (n * p1)
        err = np.max(np.abs(xn - x))
        x = xn
        if err < 4 * np.finfo(LD).eps:
            break
    x = np.sort(x)
    x = 0.5 * (x - x[::-1])  # symmetrise
    out = x.astype(np.float64)
    out[0], out[-1] = -1.0, 1.0
    return out


def get_grid(rmax, num_el, igrid=4, zexp=2.0):
    i = np.arange(num_el + 1, dtype=np.float64)
    if igrid == 1:
        b = rmax * i / num_el
    elif igrid == 2:
        b = i * i * rmax / (num_el * num_el)
    elif igrid == 3:
        b = rmax * (i / num_el) ** zexp
    elif igrid == 4:
        top = math.log(rmax + 1) ** (1.0 / zexp)
        b = np.exp((i * (top / num_el)) ** zexp) - 1.0
    else:
        raise ValueError("grid type not restated in the oracle")
    b[0], b[-1] = 0.0, rmax
    return b


class PolyBasis:
    """primbas 4 = LIP, 5 = HIP on Gauss-Lobatto nodes."""

    def __init__(self, primbas, nnodes):
        assert primbas in (4, 5)
        self.id = primbas
        self.x0 = lobatto_nodes(nnodes)
        self.noverlap = 1 if primbas == 4 else 2
        self.nprim = nnodes if primbas == 4 else 2 * nnodes
        self.enabled = list(range(self.nprim))
        if primbas == 5:
            d = self._lip(self.x0, 1)
            self.lipxi = np.array([d[1][i, i] for i in range(nnodes)], dtype=LD)

    def copy(self):
        p = PolyBasis.__new__(PolyBasis)
        p.__dict__.update(self.__dict__)
        p.enabled = list(self.enabled)
        return p

    def drop_first(self, func, deriv):
        if self.id == 4:
            if func:
                self.enabled = self.enabled[1:]
        else:
            if func and deriv:
                self.enabled = self.enabled[2:]
            elif func:
                self.enabled = self.enabled[1:]
            elif deriv:
                self.enabled = [self.enabled[0]] + self.enabled[2:]

    def drop_last(self, func, deriv):
        if self.id == 4:
            if func:
                self.enabled = self.enabled[:-1]
        else:
            if func and deriv:
                self.enabled = self.enabled[:-2]
            elif deriv:
                self.enabled = self.enabled[:-1]
            else:
                self.enabled = self.enabled[:-2] + [self.enabled[-1]]

    @property
    def nbf(self):
        return len(self.enabled)

    def _lip(self, x, nmax):
        """Derivatives 0..nmax of all Lagrange polynomials at x, by expanding
        prod_k ((x-x_k) + t)/(x_f-x_k) in powers of t (coefficient of t^n = l^(n)/n!)."""
        x = np.asarray(x, dtype=LD)
        x0 = self.x0.astype(LD)
        N = len(x0)
        out = [np.zeros((len(x), N), dtype=LD) for _ in range(nmax + 1)]
        for f in range(N):
            c = np.zeros((nmax + 1, len(x)), dtype=LD)
            c[0] = 1
            for k in range(N):
                if k == f:
                    continue
                a = (x - x0[k]) / (x0[f] - x0[k])
                s = 1 / (x0[f] - x0[k])
                new = a * c
                new[1:] += s * c[:-1]
                c = new
            fact = LD(1)
            for n in range(nmax + 1):
                if n > 0:
                    fact *= n
                out[n][:, f] = fact * c[n]
        return out

    def eval_dnf(self, x, n, element_length):
        """(len(x), nbf) float64: n-th derivative of the enabled functions / element_length^n."""
        if n > 20:
            raise ValueError("%dth derivatives not implemented!" % n)  # reference limit (generated code)
        x = np.asarray(x, dtype=LD)
        d = self._lip(x, n)
        if self.id == 4:
            prim = d[n]
        else:
            N = len(self.x0)
            x0 = self.x0.astype(LD)
            prim = np.zeros((len(x), 2 * N), dtype=LD)

            def dsq(m):  # d^m (l^2) by Leibniz
                return sum(math.comb(m, k) * d[k] * d[m - k] for k in range(m + 1))

            f2n = dsq(n)
            f2m = dsq(n - 1) if n > 0 else 0 * f2n
            f1 = 1 - 2 * (x[:, None] - x0[None, :]) * self.lipxi[None, :]
            df1 = -2 * self.lipxi[None, :]
            f3 = x[:, None] - x0[None, :]
            prim[:, 0::2] = n * df1 * f2m + f1 * f2n
            prim[:, 1::2] = (n * f2m + f3 * f2n) * LD(element_length)
        res = prim[:, self.enabled] / LD(element_length) ** n
        return res.astype(np.float64)


def exchange_tei(tei, Ni, Nj, Nk, Nl):
    t = tei.reshape((Ni, Nj, Nk, Nl), order="F")  # index (i,j,k,l), rows i fastest then j; cols k fastest then l
    # ktei[(j,k),(i,l)] = tei[(i,j),(k,l)]
    return np.transpose(t, (1, 2, 0, 3)).reshape((Nj * Nk, Ni * Nl), order="F")


def twoe_integral(rmin, rmax, xq, wq, poly, L):
    nq = len(xq)
    rmid0, rlen0 = 0.5 * (rmax + rmin), 0.5 * (rmax - rmin)
    r = rmid0 + rlen0 * xq
    npf = poly.nbf
    inner = np.zeros((nq, npf * npf))
    for ip in range(nq):
        a = rmin if ip == 0 else r[ip - 1]
        b = r[ip]
        rmid, rlen = 0.5 * (b + a), 0.5 * (b - a)
        rs = rmid + rlen * xq
        wp = wq * ((rs / b) ** L / b) * rlen
        xpoly = (rs - rmid0) / rlen0
        bf = poly.eval_dnf(xpoly, 0, rlen0)
        inner[ip] = (bf.T @ (bf * wp[:, None])).reshape(-1, order="F")
    for ip in range(1, nq):
        inner[ip] += inner[ip - 1] * (r[ip] ** (-L - 1) / r[ip - 1] ** (-L - 1))
    bf = poly.eval_dnf(xq, 0, rlen0)
    # bfprod column fi*np+fj = bf_fi * bf_fj
    bfprod = (bf[:, :, None] * bf[:, None, :]).reshape(nq, npf * npf) * (wq * rlen0)[:, None]
    ints = bfprod.T @ inner
    return ints + ints.T


class RadialSetup:
    """Radial FEM basis of the atomic / sadatom programs (point nucleus)."""

    def __init__(self, primbas, nnodes, bval, nquad=0, taylor_order=-1, zeroder=False):
        self.poly = PolyBasis(primbas, nnodes)
        self.bval = np.asarray(bval, dtype=np.float64)
        self.nel = len(bval) - 1
        self.nquad = nquad if nquad > 0 else 5 * self.poly.nbf
        self.taylor_order = taylor_order if taylor_order >= 0 else self.poly.nprim - 1
        self.flags = (True, False, True, bool(zeroder))
        self.xq, self.wq = chebyshev(self.nquad)
        self.first, self.last = [], []
        for iel in range(self.nel):
            f = 0 if iel == 0 else self.last[iel - 1] - self.poly.noverlap + 1
            self.first.append(f)
            self.last.append(f + self.basis(iel).nbf - 1)
        self.nbf = self.last[-1] + 1
        origin = np.array([-1.0])
        self.taylor_df = [self.eval_dnf(origin, i + 1, 0)[0] for i in range(self.taylor_order)]
        # RadialBasis.cpp:59-133: the cutoff ends up at the radius of the second node of element 0
        nodes = np.sort(self.poly.x0)
        self.cutoff = self.mid(0) + self.scal(0) * nodes[1]

    def basis(self, iel):
        p = self.poly.copy()
        if iel == 0:
            p.drop_first(self.flags[0], self.flags[1])
        if iel == self.nel - 1:
            p.drop_last(self.flags[2], self.flags[3])
        return p

    def mid(self, iel):
        return 0.5 * (self.bval[iel] + self.bval[iel + 1])

    def scal(self, iel):
        return (self.bval[iel + 1] - self.bval[iel]) / 2

    def r(self, iel, x=None):
        x = self.xq if x is None else x
        return self.mid(iel) + self.scal(iel) * x

    def wrad(self, iel):
        return self.scal(iel) * self.wq

    def eval_dnf(self, x, n, iel):
        return self.basis(iel).eval_dnf(x, n, self.scal(iel))

    def _taylor(self, r, ider):
        to = self.taylor_order
        coeff = np.zeros(to)
        rexp = np.arange(to)
        if to:
            coeff[0] = 1.0
        for i in range(1, to):
            coeff[i] = coeff[i - 1] / (i + 1)
        for i in range(to):
            for _ in range(ider):
                coeff[i] *= rexp[i]
                rexp[i] -= 1
        val = np.zeros((len(r), len(self.taylor_df[0]) if to else 0))
        for it in range(ider, to):
            val += coeff[it] * np.outer(r ** rexp[it], self.taylor_df[it])
        return val

    def _radial(self, iel, x, kind):
        x = self.xq if x is None else x
        r = self.r(iel, x)
        f = self.eval_dnf(x, 0, iel)
        d = self.eval_dnf(x, 1, iel) if kind >= 1 else None
        l2 = self.eval_dnf(x, 2, iel) if kind >= 2 else None
        invr = (1.0 / r)[:, None]
        if kind == 0:
            out = f * invr
        elif kind == 1:
            out = (-f * invr + d) * invr
        else:
            out = ((2.0 * f * invr - 2.0 * d) * invr + l2) * invr
        if iel == 0:
            nt = int(np.sum(r <= self.cutoff))
            if nt:
                out[:nt] = self._taylor(r[:nt], kind)
        return out

    def get_bf(self, iel, x=None):
        return self._radial(iel, x, 0)

    def get_df(self, iel, x=None):
        return self._radial(iel, x, 1)

    def get_lf(self, iel, x=None):
        return self._radial(iel, x, 2)

    def radial_integral(self, rexp, iel):
        r = self.r(iel)
        wp = self.wq * self.scal(iel) * r ** (rexp + 2)
        bf = self.get_bf(iel)
        return bf.T @ (bf * wp[:, None])

    def element_matrix(self, iel, which):
        wp = (self.wq * self.scal(iel))[:, None]
        if which == "overlap":
            f = self.eval_dnf(self.xq, 0, iel)
            return f.T @ (f * wp)
        if which == "kinetic":
            f = self.eval_dnf(self.xq, 1, iel)
            return 0.5 * (f.T @ (f * wp))
        if which == "kinetic_l":
            f = self.get_bf(iel)
            return 0.5 * (f.T @ (f * wp))
        raise ValueError(which)

    def assemble(self, fn):
        M = np.zeros((self.nbf, self.nbf))
        for iel in range(self.nel):
            f, l = self.first[iel], self.last[iel]
            M[f:l + 1, f:l + 1] += fn(iel)
        return M

    def overlap(self):
        return self.assemble(lambda e: self.radial_integral(0, e))

    def kinetic(self):
        return self.assemble(lambda e: self.element_matrix(e, "kinetic"))

    def kinetic_l(self):
        return self.assemble(lambda e: self.element_matrix(e, "kinetic_l"))

    def nuclear(self, Z):
        return -Z * self.assemble(lambda e: self.radial_integral(-1, e))

    def twoe_integral(self, L, iel):
        return twoe_integral(self.bval[iel], self.bval[iel + 1], self.xq, self.wq, self.basis(iel), L)
