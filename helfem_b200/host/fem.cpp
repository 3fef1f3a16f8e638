// Host-side finite-element setup (see fem.h for the reference map).
#include "fem.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <sstream>

namespace hfb {
namespace fem {

// ---------------------------------------------------------------------------
// Quadrature rules
// ---------------------------------------------------------------------------

// Reference: libhelfem/src/chebyshev.cpp:22-53.  The rule is generated for
// i = 1..n (descending nodes) and then reversed, so the result is ascending.
void chebyshev(int n, Vec& x, Vec& w) {
  x.assign(n, 0.0);
  w.assign(n, 0.0);
  const double oonpp = 1.0 / (n + 1.0);
  for (int i = 1; i <= n; i++) {
    const double sine = sin(i * M_PI * oonpp);
    const double sinesq = sine * sine;
    const double cosine = cos(i * M_PI * oonpp);
    const double wi = 16.0 / 3.0 / (n + 1.0) * sinesq * sinesq;
    const double xi = 1.0 - 2.0 * i * oonpp + M_2_PI * (1.0 + 2.0 / 3.0 * sinesq) * cosine * sine;
    x[n - i] = xi;
    w[n - i] = wi;
  }
}

// Reference: libhelfem/src/lobatto.cpp:588-743 (30-digit tables for n<20, a
// Newton iteration on the Legendre recurrence otherwise).  Here the Newton
// iteration is always used, in extended precision, and the result rounded to
// double; the nodes agree with the tabulated ones to the last bit or ulp.
Vec lobatto_nodes(int n) {
  if (n < 2) throw std::runtime_error("Lobatto rule needs n>=2");
  std::vector<long double> x(n), xold(n), pm1(n), pm2(n);
  const long double pi = acosl(-1.0L);
  for (int i = 0; i < n; i++) x[i] = cosl(pi * i / (n - 1));
  for (int iter = 0; iter < 200; iter++) {
    long double err = 0.0L;
    for (int i = 0; i < n; i++) {
      // Legendre recurrence up to P_{n-1}
      long double p0 = 1.0L, p1 = x[i];
      for (int j = 2; j <= n - 1; j++) {
        long double p2 = ((2 * j - 1) * x[i] * p1 - (j - 1) * p0) / j;
        p0 = p1;
        p1 = p2;
      }
      // p1 = P_{n-1}, p0 = P_{n-2}
      long double xn = x[i] - (x[i] * p1 - p0) / (n * p1);
      err = std::max(err, fabsl(xn - x[i]));
      x[i] = xn;
    }
    if (err < 4 * LDBL_EPSILON) break;
  }
  Vec out(n);
  for (int i = 0; i < n; i++) out[i] = (double)x[n - 1 - i];
  // enforce exact end points and symmetry
  out[0] = -1.0;
  out[n - 1] = 1.0;
  for (int i = 0; i < n / 2; i++) {
    double a = 0.5 * (out[n - 1 - i] - out[i]);
    out[i] = -a;
    out[n - 1 - i] = a;
  }
  if (n % 2) out[n / 2] = 0.0;
  return out;
}

// Reference: libhelfem/src/grid.cpp:18-87
Vec get_grid(double rmax, int num_el, int igrid, double zexp) {
  Vec b(num_el + 1, 0.0);
  switch (igrid) {
    case 1:
      for (int i = 0; i <= num_el; i++) b[i] = rmax * i / num_el;  // linspace
      break;
    case 2:
      for (int i = 0; i <= num_el; i++) b[i] = i * i * rmax / (num_el * num_el);
      break;
    case 3:
      for (int i = 0; i <= num_el; i++) b[i] = rmax * std::pow(i * 1.0 / num_el, zexp);
      break;
    case 4: {
      const double top = std::pow(log(rmax + 1), 1.0 / zexp);
      for (int i = 0; i <= num_el; i++) {
        // arma::linspace(0, top, n): start + i*delta
        const double t = (i == num_el) ? top : i * (top / num_el);
        b[i] = exp(std::pow(t, zexp)) - 1.0;
      }
    } break;
    case 5: {
      if (zexp <= 0.0 || zexp >= 1.0) throw std::logic_error("Invalid value for s parameter!\n");
      Vec hk(num_el);
      hk[num_el - 1] = (1.0 - zexp) / (1.0 - std::pow(zexp, num_el)) * rmax;
      for (int iel = num_el - 2; iel >= 0; iel--) hk[iel] = zexp * hk[iel + 1];
      for (int iel = 0; iel < num_el; iel++) b[iel + 1] = b[iel] + hk[iel];
    } break;
    default:
      throw std::logic_error("Invalid choice for grid\n");
  }
  b[0] = 0.0;
  b[num_el] = rmax;
  return b;
}

// ---------------------------------------------------------------------------
// Primitive polynomials
// ---------------------------------------------------------------------------

PolyBasis::PolyBasis(int primbas, int nnodes) {
  if (nnodes < 2) throw std::logic_error("Can't have finite element basis with less than two nodes per element.\n");
  if (primbas != 4 && primbas != 5)
    throw std::logic_error("Unsupported primitive basis (this build implements 4 = LIP and 5 = HIP).\n");
  id = primbas;
  x0 = lobatto_nodes(nnodes);
  if (primbas == 4) {
    noverlap = 1;
    nprim = nnodes;
  } else {
    noverlap = 2;
    nprim = 2 * nnodes;
    // l_i'(x_i)
    Mat dl = eval_lip(x0, 1);
    lipxi.resize(nnodes);
    for (int i = 0; i < nnodes; i++) lipxi[i] = dl(i, i);
  }
  enabled.resize(nprim);
  for (int i = 0; i < nprim; i++) enabled[i] = i;
}

// Reference: LIPBasis.cpp:42-52, HIPBasis.cpp:28-60
void PolyBasis::drop_first(bool func, bool deriv) {
  if (id == 4) {
    if (func) enabled.erase(enabled.begin());
  } else {
    if (func && deriv)
      enabled.erase(enabled.begin(), enabled.begin() + 2);
    else if (func)
      enabled.erase(enabled.begin());
    else if (deriv)
      enabled.erase(enabled.begin() + 1);
  }
}

void PolyBasis::drop_last(bool func, bool deriv) {
  if (id == 4) {
    if (func) enabled.pop_back();
  } else {
    if (func && deriv) {
      enabled.pop_back();
      enabled.pop_back();
    } else if (deriv) {
      enabled.pop_back();
    } else {
      // only the function: keep the derivative function (the last one).  Note that this
      // is also what the reference does when neither flag is set (HIPBasis.cpp:53-59).
      enabled.erase(enabled.end() - 2);
    }
  }
}

// n-th derivative of the Lagrange interpolating polynomials.
//   l_f(x) = prod_{k!=f} (x-x_k)/(x_f-x_k),
//   l_f^(n)(x) = n! * e_{m-n}({x-x_k}_{k!=f}) / prod_{k!=f}(x_f-x_k),  m = N-1,
// e_j being the elementary symmetric polynomial, i.e. the sum over all products that
// leave out n of the factors -- the same sum the reference's generated nested loops
// enumerate term by term (LIPBasis_eval.cpp:8-...).
Mat PolyBasis::eval_lip(const Vec& x, int n) const {
  const size_t N = x0.size();
  Mat out(x.size(), N);
  if (n < 0) throw std::logic_error("negative derivative order");
  if ((size_t)n > N - 1) return out;  // derivative of order > degree vanishes
  double nfac = 1.0;
  for (int i = 2; i <= n; i++) nfac *= i;
  const size_t m = N - 1;
  const size_t keep = m - n;  // factors kept in each product
  std::vector<double> e(keep + 1);
  for (size_t ix = 0; ix < x.size(); ix++)
    for (size_t fi = 0; fi < N; fi++) {
      std::fill(e.begin(), e.end(), 0.0);
      e[0] = 1.0;
      size_t cnt = 0;
      for (size_t ip = 0; ip < N; ip++) {
        if (ip == fi) continue;
        // ratio form keeps every partial product O(1)
        const double y = (x[ix] - x0[ip]) / (x0[fi] - x0[ip]);
        const double yd = 1.0 / (x0[fi] - x0[ip]);
        // Products either include the factor (ratio y) or have it removed by a
        // derivative (only its denominator 1/(x_f-x_k) survives).  Track by number kept.
        cnt++;
        // e[j]: sum over partial products with j factors kept and cnt-j removed (<= n)
        for (int j = (int)std::min(cnt, keep); j >= 0; j--) {
          const int removed = (int)cnt - j;
          double v = 0.0;
          if (j > 0) v += e[j - 1] * y;                        // keep this factor
          if (removed >= 1 && removed <= n) v += e[j] * yd;    // remove it
          e[j] = v;
        }
      }
      out(ix, fi) = nfac * e[keep];
    }
  return out;
}

// Reference: HIPBasis_eval.cpp (Leibniz expansion of [l_i^2]^(n), generated up to n=20)
Mat PolyBasis::eval_prim_dnf(const Vec& x, int n, double element_length) const {
  if (n > 20) {
    std::ostringstream oss;
    oss << n << "th derivatives not implemented!\n";  // same limit as the reference's generated code
    throw std::logic_error(oss.str());
  }
  if (id == 4) return eval_lip(x, n);

  const size_t N = x0.size();
  std::vector<Mat> dl(n + 1);
  for (int k = 0; k <= n; k++) dl[k] = eval_lip(x, k);
  // binomials
  std::vector<std::vector<double>> binom(n + 1, std::vector<double>(n + 1, 0.0));
  for (int a = 0; a <= n; a++) {
    binom[a][0] = 1.0;
    for (int b = 1; b <= a; b++) binom[a][b] = binom[a - 1][b - 1] + (b <= a - 1 ? binom[a - 1][b] : 0.0);
  }
  Mat out(x.size(), 2 * N);
  for (size_t ix = 0; ix < x.size(); ix++)
    for (size_t fi = 0; fi < N; fi++) {
      // d^m [l^2] for m = n and n-1
      auto dsq = [&](int m) {
        double s = 0.0;
        for (int k = 0; k <= m; k++) s += binom[m][k] * dl[k](ix, fi) * dl[m - k](ix, fi);
        return s;
      };
      const double f1 = 1.0 - 2.0 * (x[ix] - x0[fi]) * lipxi[fi];
      const double df1 = -2.0 * lipxi[fi];
      const double f3 = x[ix] - x0[fi];
      const double dnf2 = dsq(n);
      const double dnm1f2 = (n > 0) ? dsq(n - 1) : 0.0;
      out(ix, 2 * fi) = n * df1 * dnm1f2 + f1 * dnf2;
      out(ix, 2 * fi + 1) = (n * 1.0 * dnm1f2 + f3 * dnf2) * element_length;
    }
  return out;
}

// Reference: PolynomialBasis.cpp:175-179
Mat PolyBasis::eval_dnf(const Vec& x, int n, double element_length) const {
  Mat prim = eval_prim_dnf(x, n, element_length);
  Mat out(x.size(), enabled.size());
  const double scale = std::pow(element_length, n);
  for (size_t j = 0; j < enabled.size(); j++)
    for (size_t i = 0; i < x.size(); i++) out(i, j) = prim(i, enabled[j]) / scale;
  return out;
}

// ---------------------------------------------------------------------------
// Finite element bookkeeping
// ---------------------------------------------------------------------------

FEBasis::FEBasis(const PolyBasis& poly_, const Vec& bval_, bool zfl_, bool zdl_, bool zfr_, bool zdr_)
    : poly(poly_), bval(bval_), zfl(zfl_), zdl(zdl_), zfr(zfr_), zdr(zdr_) {
  if (bval.size() < 2) throw std::logic_error("Can't update basis function list since there are no elements!\n");
  const size_t nel = bval.size() - 1;
  first.assign(nel, 0);
  last.assign(nel, 0);
  // Reference: FiniteElementBasis.cpp:39-51
  for (size_t iel = 0; iel < nel; iel++) {
    first[iel] = (iel == 0) ? 0 : last[iel - 1] - poly.get_noverlap() + 1;
    last[iel] = first[iel] + get_basis(iel).get_nbf() - 1;
  }
  const double disc = check_bf_continuity();
  if (disc > sqrt(DBL_EPSILON)) throw std::logic_error("Finite element basis set is not continuous\n");
}

// Reference: FiniteElementBasis.cpp:269-277
PolyBasis FEBasis::get_basis(size_t iel) const {
  PolyBasis p(poly);
  if (iel == 0) p.drop_first(zfl, zdl);
  if (iel == bval.size() - 2) p.drop_last(zfr, zdr);
  return p;
}

Vec FEBasis::eval_coord(const Vec& x, size_t iel) const {
  Vec r(x.size());
  const double mid = element_midpoint(iel), sc = scaling_factor(iel);
  for (size_t i = 0; i < x.size(); i++) r[i] = mid * 1.0 + sc * x[i];
  return r;
}

Vec FEBasis::eval_prim(const Vec& r, size_t iel) const {
  Vec x(r.size());
  const double mid = element_midpoint(iel), sc = scaling_factor(iel);
  for (size_t i = 0; i < r.size(); i++) x[i] = (r[i] - mid * 1.0) / sc;
  return x;
}

Mat FEBasis::eval_dnf(const Vec& x, int n, size_t iel) const {
  return get_basis(iel).eval_dnf(x, n, scaling_factor(iel));
}

// Reference: FiniteElementBasis.cpp:52-115
double FEBasis::check_bf_continuity() const {
  const size_t nel = get_nelem();
  if (nel == 1) return 0.0;
  const int nov = poly.get_noverlap();
  double worst = 0.0;
  Vec xl(1, 1.0), xr(1, -1.0);
  for (size_t iel = 0; iel + 1 < nel; iel++) {
    double d2 = 0.0;
    for (int ider = 0; ider < nov; ider++) {
      Mat fl = eval_dnf(xl, ider, iel);
      Mat fr = eval_dnf(xr, ider, iel + 1);
      for (int k = 0; k < nov; k++) {
        const double diff = fl(0, fl.nc - nov + k) - fr(0, k);
        d2 += diff * diff;
      }
    }
    worst = std::max(worst, sqrt(d2));
  }
  return worst;
}

// ---------------------------------------------------------------------------
// Radial basis  chi_i(r) = B_i(r)/r
// ---------------------------------------------------------------------------

// Reference: RadialBasis.cpp:32-57
RadialBasis::RadialBasis(const FEBasis& fem_, int n_quad, int taylor_order_) : fem(fem_), taylor_order(taylor_order_) {
  chebyshev(n_quad, xq, wq);
  Vec origin(1, -1.0);
  taylor_df.resize(taylor_order);
  for (int i = 0; i < taylor_order; i++) {
    Mat d = fem.eval_dnf(origin, i + 1, 0);
    taylor_df[i].resize(d.nc);
    for (size_t f = 0; f < d.nc; f++) taylor_df[i][f] = d(0, f);
  }
  set_small_r_taylor_cutoff();
}

// Reference: RadialBasis.cpp:575-626.  B(r)/r = B'(0) + B''(0) r/2 + B'''(0) r^2/6 + ...
// evaluated for the first `ntaylor` points (the reference requires them to be consecutive
// starting at index 0).
void RadialBasis::get_taylor(const Vec& r, size_t ntaylor, Mat& val, int ider) const {
  Vec coeff(taylor_order);
  std::vector<int> rexp(taylor_order);
  if (taylor_order > 0) coeff[0] = 1.0;
  for (int i = 1; i < taylor_order; i++) coeff[i] = coeff[i - 1] / (i + 1);
  for (int i = 0; i < taylor_order; i++) rexp[i] = i;
  for (int i = 0; i < taylor_order; i++)
    for (int d = 0; d < ider; d++) {
      coeff[i] *= rexp[i];
      rexp[i]--;
    }
  for (size_t ifun = 0; ifun < val.nc; ifun++)
    for (size_t ir = 0; ir < ntaylor; ir++) {
      double v = 0.0;
      for (int it = ider; it < taylor_order; it++) v += coeff[it] * taylor_df[it][ifun] * std::pow(r[ir], rexp[it]);
      val(ir, ifun) = v;
    }
}

// Reference: RadialBasis.cpp:59-133.  The reference scans 1000 log-spaced trial radii for the
// first local minimum of the Taylor-vs-analytic mismatch, but then unconditionally selects
// the LAST trial radius (its `small_r_taylor_cutoff == DBL_MAX` test is always true at that
// point, RadialBasis.cpp:114-117), i.e. the radius of the second node of the first element.
// The behaviour (not the dead scan) is what is reproduced here; taylor_diff is the summed
// relative mismatch of value, first and second derivative at that radius.
void RadialBasis::set_small_r_taylor_cutoff() {
  PolyBasis p = fem.get_basis(0);
  Vec nodes = p.get_nodes();
  std::sort(nodes.begin(), nodes.end());
  Vec minx(1, nodes[1]);
  const double maxr = fem.eval_coord(minx, 0)[0];
  // last point of arma::logspace(-10,0,1000)*maxr
  Vec rcut(1, std::pow(10.0, 0.0) * maxr);
  Vec xprim = fem.eval_prim(rcut, 0);

  small_r_taylor_cutoff = -1.0;  // analytic values
  Mat bf0 = get_bf(xprim, 0), df0 = get_df(xprim, 0), lf0 = get_lf(xprim, 0);
  Mat bft(bf0), dft(df0), lft(lf0);
  get_taylor(rcut, 1, bft, 0);
  get_taylor(rcut, 1, dft, 1);
  get_taylor(rcut, 1, lft, 2);
  auto relnorm = [](const Mat& a, const Mat& b) {
    double n = 0.0, d = 0.0;
    for (size_t i = 0; i < a.size(); i++) {
      n += (a.d[i] - b.d[i]) * (a.d[i] - b.d[i]);
      d += b.d[i] * b.d[i];
    }
    return sqrt(n) / sqrt(d);
  };
  double diff = relnorm(bft, bf0);
  if (taylor_order >= 1) diff += relnorm(dft, df0);
  if (taylor_order > 1) diff += relnorm(lft, lf0);
  taylor_diff = diff;
  small_r_taylor_cutoff = rcut[0];
}

static size_t count_taylor(const Vec& r, double cutoff, size_t iel) {
  if (iel != 0) return 0;
  size_t n = 0;
  for (size_t i = 0; i < r.size(); i++)
    if (r[i] <= cutoff) n++;
  // the reference requires consecutive leading indices (RadialBasis.cpp:576-577)
  for (size_t i = 0; i < n; i++)
    if (!(r[i] <= cutoff)) throw std::logic_error("Taylor points not consecutive!\n");
  return n;
}

// Reference: RadialBasis.cpp:649-670
Mat RadialBasis::get_bf(const Vec& x, size_t iel) const {
  Mat val = fem.eval_dnf(x, 0, iel);
  Vec r = fem.eval_coord(x, iel);
  const size_t nt = count_taylor(r, small_r_taylor_cutoff, iel);
  if (nt) get_taylor(r, nt, val, 0);
  for (size_t f = 0; f < val.nc; f++)
    for (size_t ir = nt; ir < x.size(); ir++) val(ir, f) /= r[ir];
  return val;
}

// Reference: RadialBasis.cpp:676-699
Mat RadialBasis::get_df(const Vec& x, size_t iel) const {
  Mat fval = fem.eval_dnf(x, 0, iel);
  Mat dval = fem.eval_dnf(x, 1, iel);
  Vec r = fem.eval_coord(x, iel);
  Mat der(fval);
  const size_t nt = count_taylor(r, small_r_taylor_cutoff, iel);
  if (nt) get_taylor(r, nt, der, 1);
  for (size_t f = 0; f < der.nc; f++)
    for (size_t ir = nt; ir < x.size(); ir++) {
      const double invr = 1.0 / r[ir];
      der(ir, f) = (-fval(ir, f) * invr + dval(ir, f)) * invr;
    }
  return der;
}

// Reference: RadialBasis.cpp:705-730
Mat RadialBasis::get_lf(const Vec& x, size_t iel) const {
  Mat fval = fem.eval_dnf(x, 0, iel);
  Mat dval = fem.eval_dnf(x, 1, iel);
  Mat lval = fem.eval_dnf(x, 2, iel);
  Vec r = fem.eval_coord(x, iel);
  Mat lapl(fval);
  const size_t nt = count_taylor(r, small_r_taylor_cutoff, iel);
  if (nt) get_taylor(r, nt, lapl, 2);
  for (size_t f = 0; f < lapl.nc; f++)
    for (size_t ir = nt; ir < x.size(); ir++) {
      const double invr = 1.0 / r[ir];
      lapl(ir, f) = ((2.0 * fval(ir, f) * invr - 2.0 * dval(ir, f)) * invr + lval(ir, f)) * invr;
    }
  return lapl;
}

Vec RadialBasis::get_wrad(size_t iel) const {
  Vec w(wq.size());
  const double sc = fem.scaling_factor(iel);
  for (size_t i = 0; i < wq.size(); i++) w[i] = sc * wq[i];
  return w;
}

// Reference: RadialBasis.cpp:190-199 + FiniteElementBasis.cpp:387-415
Mat RadialBasis::radial_integral(int Rexp, size_t iel) const {
  Vec r = fem.eval_coord(xq, iel);
  Vec wp(wq.size());
  const double sc = fem.scaling_factor(iel);
  for (size_t i = 0; i < wq.size(); i++) wp[i] = wq[i] * sc * std::pow(r[i], Rexp + 2);
  Mat bf = get_bf(xq, iel);
  return atwb(bf, wp, bf);
}

Mat RadialBasis::overlap(size_t iel) const {
  Vec wp(wq.size());
  const double sc = fem.scaling_factor(iel);
  for (size_t i = 0; i < wq.size(); i++) wp[i] = wq[i] * sc;
  Mat f = fem.eval_dnf(xq, 0, iel);
  return atwb(f, wp, f);
}

// Reference: RadialBasis.cpp:319-327 (0.5 <B'|B'>)
Mat RadialBasis::kinetic(size_t iel) const {
  Vec wp(wq.size());
  const double sc = fem.scaling_factor(iel);
  for (size_t i = 0; i < wq.size(); i++) wp[i] = wq[i] * sc;
  Mat f = fem.eval_dnf(xq, 1, iel);
  Mat T = atwb(f, wp, f);
  for (auto& v : T.d) v *= 0.5;
  return T;
}

// Reference: RadialBasis.cpp:329-345 (0.5 <B/r|B/r>)
Mat RadialBasis::kinetic_l(size_t iel) const {
  Vec wp(wq.size());
  const double sc = fem.scaling_factor(iel);
  for (size_t i = 0; i < wq.size(); i++) wp[i] = wq[i] * sc;
  Mat f = get_bf(xq, iel);
  Mat T = atwb(f, wp, f);
  for (auto& v : T.d) v *= 0.5;
  return T;
}

// ---------------------------------------------------------------------------
// In-element two-electron integrals
// ---------------------------------------------------------------------------

// Reference: quadrature.cpp:23-126.
//   inner(p; k,l) = r_p^{-L-1} int_{rmin}^{r_p} r'^L B_k B_l dr'
// accumulated over the sub-intervals [r_{p-1}, r_p], each with a fresh Nq-point rule, and
// rescaled between points by (r_{p-1}/r_p)^{L+1}; outer integral over the element rule;
// the r<r' half is the transpose.
Mat twoe_integral(double rmin, double rmax, const Vec& x, const Vec& wx, const PolyBasis& poly, int L) {
  const size_t nq = x.size();
  const size_t np = poly.get_nbf();
  const double rmid0 = 0.5 * (rmax + rmin), rlen0 = 0.5 * (rmax - rmin);
  Vec r(nq);
  for (size_t i = 0; i < nq; i++) r[i] = rmid0 * 1.0 + rlen0 * x[i];

  // inner integrals, one row per outer point
  Mat inner(nq, np * np);
  Vec rs(nq), wp(nq), xpoly(nq);
  for (size_t ip = 0; ip < nq; ip++) {
    const double a = (ip == 0) ? rmin : r[ip - 1];
    const double b = r[ip];
    const double rmid = 0.5 * (b + a), rlen = 0.5 * (b - a);
    for (size_t i = 0; i < nq; i++) {
      rs[i] = rmid * 1.0 + rlen * x[i];
      const double fsb = std::pow(rs[i] / b, L) / b;
      wp[i] = wx[i] * fsb * rlen;
      xpoly[i] = (rs[i] - rmid0 * 1.0) / rlen0;
    }
    Mat bf = poly.eval_dnf(xpoly, 0, rlen0);
    for (size_t l = 0; l < np; l++)
      for (size_t k = 0; k <= l; k++) {
        double s = 0.0;
        for (size_t i = 0; i < nq; i++) s += bf(i, k) * wp[i] * bf(i, l);
        inner(ip, k + l * np) = s;
        inner(ip, l + k * np) = s;
      }
  }
  for (size_t ip = 1; ip < nq; ip++) {
    const double ratio = std::pow(r[ip], -L - 1) / std::pow(r[ip - 1], -L - 1);
    for (size_t c = 0; c < np * np; c++) inner(ip, c) += inner(ip - 1, c) * ratio;
  }

  // outer integral
  Mat bf = poly.eval_dnf(x, 0, rlen0);
  Mat ints(np * np, np * np);
  Vec prod(nq);
  for (size_t fi = 0; fi < np; fi++)
    for (size_t fj = 0; fj <= fi; fj++) {
      for (size_t i = 0; i < nq; i++) prod[i] = bf(i, fi) * bf(i, fj) * (wx[i] * rlen0);
      for (size_t l = 0; l < np; l++)
        for (size_t k = 0; k <= l; k++) {
          double s = 0.0;
          const double* col = &inner.d[(k + l * np) * nq];
          for (size_t i = 0; i < nq; i++) s += prod[i] * col[i];
          ints(fi * np + fj, k + l * np) = s;
          ints(fi * np + fj, l + k * np) = s;
          ints(fj * np + fi, k + l * np) = s;
          ints(fj * np + fi, l + k * np) = s;
        }
    }
  // ints += ints^T
  Mat out(np * np, np * np);
  for (size_t c = 0; c < np * np; c++)
    for (size_t rw = 0; rw < np * np; rw++) out(rw, c) = ints(rw, c) + ints(c, rw);
  return out;
}

// Reference: RadialBasis.cpp:478-490
Mat RadialBasis::twoe_integral(int L, size_t iel) const {
  return fem::twoe_integral(fem.element_begin(iel), fem.element_end(iel), xq, wq, fem.get_basis(iel), L);
}

Mat exchange_tei(const Mat& tei, size_t Ni, size_t Nj, size_t Nk, size_t Nl) {
  if (tei.nr != Ni * Nj || tei.nc != Nk * Nl) throw std::logic_error("Invalid input tei\n");
  Mat ktei(Nj * Nk, Ni * Nl);
  for (size_t ii = 0; ii < Ni; ii++)
    for (size_t jj = 0; jj < Nj; jj++)
      for (size_t kk = 0; kk < Nk; kk++)
        for (size_t ll = 0; ll < Nl; ll++) ktei(kk * Nj + jj, ll * Ni + ii) = tei(jj * Ni + ii, ll * Nk + kk);
  return ktei;
}

}  // namespace fem
}  // namespace hfb
