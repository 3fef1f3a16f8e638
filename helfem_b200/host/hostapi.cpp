// Flat C entry points over the host-side mirror classes, so the Python test / bench harness
// (ctypes) can drive exactly the objects a C++ driver would use.  Every function returns 0 on
// success; C++ exceptions are translated to a non-zero code (1 = std::logic_error,
// 2 = std::runtime_error, 3 = other) and the message kept for hfh_last_error().
#include <cmath>
#include <cstring>
#include <string>
#include "../../include/helfem_b200.h"
#include "angular.h"
#include "atomic.h"
#include "dftgrid.h"
#include "diatomic.h"
#include "legendre.h"
#include "sadatom.h"

using namespace hfb;

static thread_local std::string g_msg;

#define HFH_TRY try {
#define HFH_CATCH                                           \
  }                                                         \
  catch (const std::logic_error& e) { g_msg = e.what(); return 1; }   \
  catch (const std::runtime_error& e) { g_msg = e.what(); return 2; } \
  catch (const std::exception& e) { g_msg = e.what(); return 3; }     \
  return 0;

template <class Grid, class Basis>
static int grid_eval(void* g, void* b, int nrows, int x_func, int c_func, int nspin, const double* Pa, const double* Pb, double thr,
                     double* Ha, double* Hb, int beta, double* scal3) {
  HFH_TRY
  (void)b;
  double exc = 0, nel = 0, ekin = 0;
  // caller-owned storage, no copies; the mirror class checks the order against the basis like the reference does
  ((Grid*)g)->eval_Fxc(x_func, c_func, nspin, (size_t)nrows, Pa, Pb, Ha, Hb, exc, nel, ekin, beta != 0, thr);
  scal3[0] = exc; scal3[1] = nel; scal3[2] = ekin;
  HFH_CATCH
}
extern "C" {

const char* hfh_last_error(void) { return g_msg.c_str(); }

// ---- FEM primitives (exposed for the known-answer tests) ----
int hfh_chebyshev(int n, double* x, double* w) {
  HFH_TRY
  Vec xv, wv;
  fem::chebyshev(n, xv, wv);
  memcpy(x, xv.data(), n * sizeof(double));
  memcpy(w, wv.data(), n * sizeof(double));
  HFH_CATCH
}
int hfh_lobatto_nodes(int n, double* x) {
  HFH_TRY
  Vec xv = fem::lobatto_nodes(n);
  memcpy(x, xv.data(), n * sizeof(double));
  HFH_CATCH
}
int hfh_get_grid(double rmax, int nel, int igrid, double zexp, double* bval) {
  HFH_TRY
  Vec b = fem::get_grid(rmax, nel, igrid, zexp);
  memcpy(bval, b.data(), b.size() * sizeof(double));
  HFH_CATCH
}
/// values of the n-th derivative of the primitive basis at x: out is (nx x nbf) column-major
int hfh_poly_eval(int primbas, int nnodes, int nx, const double* x, int n, double element_length, double* out, int* nbf) {
  HFH_TRY
  fem::PolyBasis p(primbas, nnodes);
  Mat m = p.eval_dnf(Vec(x, x + nx), n, element_length);
  if (nbf) *nbf = (int)m.nc;
  if (out) memcpy(out, m.ptr(), m.size() * sizeof(double));
  HFH_CATCH
}
/// quadrature::twoe_integral on [rmin,rmax] for the full primitive basis (inttest.cpp known-answer test)
int hfh_twoe_integral(int primbas, int nnodes, int nquad, double rmin, double rmax, int L, double* out) {
  HFH_TRY
  fem::PolyBasis p(primbas, nnodes);
  Vec xq, wq;
  fem::chebyshev(nquad, xq, wq);
  Mat m = fem::twoe_integral(rmin, rmax, xq, wq, p, L);
  memcpy(out, m.ptr(), m.size() * sizeof(double));
  HFH_CATCH
}
double hfh_gaunt(int L, int M, int l, int m, int lp, int mp) { return angular::gaunt_coefficient(L, M, l, m, lp, mp); }
double hfh_modified_gaunt(int lj, int mj, int L, int M, int li, int mi) { return angular::modified_gaunt_coefficient(lj, mj, L, M, li, mi); }
void hfh_ylm(int l, int m, double cth, double phi, double* re, double* im) {
  std::complex<double> y = angular::spherical_harmonics(l, m, cth, phi);
  *re = y.real();
  *im = y.imag();
}

/// out: (Mmax+1) x (Lmax+1), out[M*(Lmax+1)+L]
void hfh_plm_array(int Lmax, int Mmax, double mu, double* out) { legendre::plm_array(Lmax, Mmax, mu, out); }
void hfh_qlm_array(int Lmax, int Mmax, double mu, double* out) { legendre::qlm_array(Lmax, Mmax, mu, out); }

// ---- sadatom ----
int hfh_sadatom_new(void** out, int Z, int primbas, int nnodes, int nquad, int nelem, double rmax, int igrid, double zexp,
                    int taylor_order, int lmax, int zeroder, int device) {
  HFH_TRY
  fem::PolyBasis poly(primbas, nnodes);
  if (nquad <= 0) nquad = 5 * poly.get_nbf();  // sadatom/main.cpp:211-213
  if (taylor_order < 0) taylor_order = poly.get_nprim() - 1;  // sadatom/main.cpp:219-220
  Vec bval = fem::get_grid(rmax, nelem, igrid, zexp);
  *out = new sadatom::TwoDBasis(Z, poly, zeroder != 0, nquad, bval, taylor_order, lmax, device);
  HFH_CATCH
}
void hfh_sadatom_free(void* h) { delete (sadatom::TwoDBasis*)h; }

int hfh_sadatom_dims(void* h, int* nbf, int* nel, int* nq, int* np_max, double* taylor_cutoff, double* taylor_diff) {
  HFH_TRY
  auto* b = (sadatom::TwoDBasis*)h;
  if (nbf) *nbf = (int)b->Nbf();
  if (nel) *nel = (int)b->get_rad_Nel();
  if (nq) *nq = b->get_radial().get_nquad();
  if (np_max) {
    *np_max = 0;
    for (size_t e = 0; e < b->get_rad_Nel(); e++) *np_max = std::max(*np_max, (int)b->get_radial().Nprim(e));
  }
  if (taylor_cutoff) *taylor_cutoff = b->get_radial().get_small_r_taylor_cutoff();
  if (taylor_diff) *taylor_diff = b->get_radial().get_taylor_diff();
  HFH_CATCH
}
int hfh_sadatom_idx(void* h, int* first, int* last) {
  HFH_TRY
  auto* b = (sadatom::TwoDBasis*)h;
  for (size_t e = 0; e < b->get_rad_Nel(); e++) {
    size_t f, l;
    b->get_radial().get_idx(e, f, l);
    first[e] = (int)f;
    last[e] = (int)l;
  }
  HFH_CATCH
}
/// which: 0 overlap, 1 kinetic, 2 kinetic_l, 3 nuclear
int hfh_sadatom_matrix(void* h, int which, double* out) {
  HFH_TRY
  auto* b = (sadatom::TwoDBasis*)h;
  Mat m = which == 0 ? b->overlap() : which == 1 ? b->kinetic() : which == 2 ? b->kinetic_l() : b->nuclear();
  memcpy(out, m.ptr(), m.size() * sizeof(double));
  HFH_CATCH
}
/// tables of element iel: r, wrad (nq), bf/df/lf (nq x np_iel column-major); NULLs are skipped
int hfh_sadatom_tables(void* h, int iel, double* r, double* wrad, double* bf, double* df, double* lf) {
  HFH_TRY
  auto* b = (sadatom::TwoDBasis*)h;
  if (r) { Vec v = b->get_r(iel); memcpy(r, v.data(), v.size() * sizeof(double)); }
  if (wrad) { Vec v = b->get_wrad(iel); memcpy(wrad, v.data(), v.size() * sizeof(double)); }
  if (bf) { Mat m = b->eval_bf(iel); memcpy(bf, m.ptr(), m.size() * sizeof(double)); }
  if (df) { Mat m = b->eval_df(iel); memcpy(df, m.ptr(), m.size() * sizeof(double)); }
  if (lf) { Mat m = b->eval_lf(iel); memcpy(lf, m.ptr(), m.size() * sizeof(double)); }
  HFH_CATCH
}
int hfh_sadatom_compute_tei(void* h, int nL, int exchange) {
  HFH_TRY
  ((sadatom::TwoDBasis*)h)->compute_tei(nL, exchange != 0);
  HFH_CATCH
}
int hfh_sadatom_get_tei(void* h, int L, int iel, double* disL, double* dism1L, double* tei, double* ktei) {
  HFH_TRY
  auto* b = (sadatom::TwoDBasis*)h;
  const size_t idx = (size_t)L * b->get_rad_Nel() + iel;
  if (L >= b->get_nL()) throw std::logic_error("Primitive teis have not been computed!\n");
  if (disL) memcpy(disL, b->get_disjoint_L()[idx].ptr(), b->get_disjoint_L()[idx].size() * sizeof(double));
  if (dism1L) memcpy(dism1L, b->get_disjoint_m1L()[idx].ptr(), b->get_disjoint_m1L()[idx].size() * sizeof(double));
  if (tei) memcpy(tei, b->get_prim_tei()[idx].ptr(), b->get_prim_tei()[idx].size() * sizeof(double));
  if (ktei) {
    if (b->get_prim_ktei().empty()) throw std::logic_error("Primitive teis have not been computed!\n");
    memcpy(ktei, b->get_prim_ktei()[idx].ptr(), b->get_prim_ktei()[idx].size() * sizeof(double));
  }
  HFH_CATCH
}
int hfh_atomic_context(void* h, hfb200_ctx** ctx) {
  HFH_TRY
  *ctx = ((atomic::TwoDBasis*)h)->context();
  HFH_CATCH
}
int hfh_diatomic_set_owned_only(void* h, int on) {
  HFH_TRY
  ((diatomic::TwoDBasis*)h)->set_owned_only(on != 0);
  HFH_CATCH
}
int hfh_diatomic_context(void* h, hfb200_ctx** ctx) {
  HFH_TRY
  *ctx = ((diatomic::TwoDBasis*)h)->context();
  HFH_CATCH
}
int hfh_sadatom_context(void* h, hfb200_ctx** ctx) {
  HFH_TRY
  *ctx = ((sadatom::TwoDBasis*)h)->context();
  HFH_CATCH
}
int hfh_sadatom_coulomb(void* h, const double* P, double* J) {
  HFH_TRY
  auto* b = (sadatom::TwoDBasis*)h;
  const size_t n = b->Nbf();
  Mat Pm(n, n);
  memcpy(Pm.ptr(), P, n * n * sizeof(double));
  Mat Jm = b->coulomb(Pm);
  memcpy(J, Jm.ptr(), n * n * sizeof(double));
  HFH_CATCH
}
int hfh_sadatom_exchange(void* h, const double* P, int nslices, double* K) {
  HFH_TRY
  auto* b = (sadatom::TwoDBasis*)h;
  const size_t n = b->Nbf();
  Cube Pc(n, n, nslices);
  memcpy(Pc.ptr(), P, n * n * nslices * sizeof(double));
  Cube Kc = b->exchange(Pc);
  memcpy(K, Kc.ptr(), n * n * nslices * sizeof(double));
  HFH_CATCH
}
int hfh_sadatom_eval_fxc(void* h, int x_func, int c_func, int nspin, const double* Pa, const double* Pb, int nslices,
                         double thr, double* Ha, double* Hb, int beta, double* Exc, double* Nel) {
  HFH_TRY
  auto* b = (sadatom::TwoDBasis*)h;
  const size_t n = b->Nbf(), tot = n * n * nslices;
  sadatom::DFTGrid grid(b);
  Cube P1(n, n, nslices), H1, H2;
  memcpy(P1.ptr(), Pa, tot * sizeof(double));
  if (nspin == 1) {
    grid.eval_Fxc(x_func, Vec(), c_func, Vec(), P1, H1, *Exc, *Nel, thr);
    memcpy(Ha, H1.ptr(), tot * sizeof(double));
  } else {
    Cube P2(n, n, nslices);
    memcpy(P2.ptr(), Pb, tot * sizeof(double));
    grid.eval_Fxc(x_func, Vec(), c_func, Vec(), P1, P2, H1, H2, *Exc, *Nel, beta != 0, thr);
    memcpy(Ha, H1.ptr(), tot * sizeof(double));
    if (Hb) memcpy(Hb, H2.ptr(), tot * sizeof(double));
  }
  HFH_CATCH
}

// ---- atomic ----
int hfh_atomic_new(void** out, int Z, int primbas, int nnodes, int nquad, int nelem, double rmax, int igrid, double zexp,
                   int taylor_order, int lmax, int mmax, int zeroder, int device) {
  HFH_TRY
  fem::PolyBasis poly(primbas, nnodes);
  if (nquad <= 0) nquad = 5 * poly.get_nbf();  // atomic/main.cpp:245-249
  if (taylor_order < 0) taylor_order = poly.get_nprim() - 1;
  Vec bval = fem::get_grid(rmax, nelem, igrid, zexp);
  std::vector<int> lval, mval;
  angular::angular_basis(lmax, mmax, lval, mval);  // atomic/main.cpp:268
  *out = new atomic::TwoDBasis(Z, poly, zeroder != 0, nquad, bval, taylor_order, lval, mval, device);
  HFH_CATCH
}
void hfh_atomic_free(void* h) { delete (atomic::TwoDBasis*)h; }
int hfh_atomic_dims(void* h, int* nbf, int* nrad, int* nang, int* nel, int* nL) {
  HFH_TRY
  auto* b = (atomic::TwoDBasis*)h;
  if (nbf) *nbf = (int)b->Nbf();
  if (nrad) *nrad = (int)b->Nrad();
  if (nang) *nang = (int)b->Nang();
  if (nel) *nel = (int)b->get_radial().Nel();
  if (nL) *nL = b->get_nL();
  HFH_CATCH
}
int hfh_atomic_lm(void* h, int* lval, int* mval) {
  HFH_TRY
  auto* b = (atomic::TwoDBasis*)h;
  for (size_t a = 0; a < b->Nang(); a++) { lval[a] = b->get_lval()[a]; mval[a] = b->get_mval()[a]; }
  HFH_CATCH
}
int hfh_atomic_idx(void* h, int* first, int* last) {
  HFH_TRY
  auto* b = (atomic::TwoDBasis*)h;
  for (size_t e = 0; e < b->get_radial().Nel(); e++) {
    size_t f, l;
    b->get_radial().get_idx(e, f, l);
    first[e] = (int)f;
    last[e] = (int)l;
  }
  HFH_CATCH
}
/// which: 0 overlap, 1 kinetic, 3 nuclear
int hfh_atomic_matrix(void* h, int which, double* out) {
  HFH_TRY
  auto* b = (atomic::TwoDBasis*)h;
  Mat m = which == 0 ? b->overlap() : which == 1 ? b->kinetic() : b->nuclear();
  memcpy(out, m.ptr(), m.size() * sizeof(double));
  HFH_CATCH
}
int hfh_atomic_compute_tei(void* h, int exchange) {
  HFH_TRY
  ((atomic::TwoDBasis*)h)->compute_tei(exchange != 0);
  HFH_CATCH
}
int hfh_atomic_get_tei(void* h, int L, int iel, double* disL, double* dism1L, double* tei, double* ktei) {
  HFH_TRY
  auto* b = (atomic::TwoDBasis*)h;
  const size_t idx = (size_t)L * b->get_radial().Nel() + iel;
  if (L >= b->get_nL()) throw std::logic_error("Primitive teis have not been computed!\n");
  if (disL) memcpy(disL, b->get_disjoint_L()[idx].ptr(), b->get_disjoint_L()[idx].size() * sizeof(double));
  if (dism1L) memcpy(dism1L, b->get_disjoint_m1L()[idx].ptr(), b->get_disjoint_m1L()[idx].size() * sizeof(double));
  if (tei) memcpy(tei, b->get_prim_tei()[idx].ptr(), b->get_prim_tei()[idx].size() * sizeof(double));
  if (ktei) {
    if (b->get_prim_ktei().empty()) throw std::logic_error("Primitive teis have not been computed!\n");
    memcpy(ktei, b->get_prim_ktei()[idx].ptr(), b->get_prim_ktei()[idx].size() * sizeof(double));
  }
  HFH_CATCH
}
int hfh_atomic_coulomb(void* h, const double* P, double* J) {
  HFH_TRY
  ((atomic::TwoDBasis*)h)->coulomb(P, J);
  HFH_CATCH
}
int hfh_atomic_exchange(void* h, const double* P, double* K) {
  HFH_TRY
  ((atomic::TwoDBasis*)h)->exchange(P, K);
  HFH_CATCH
}

// ---- diatomic ----
/// lmax_per_m: lmax for each |m| = 0..nm-1 (diatomic/main.cpp:255-273, lm_to_l_m basis.cpp:287-302)
int hfh_diatomic_new(void** out, int Z1, int Z2, double Rbond, int primbas, int nnodes, int nquad, int nelem, double rmax,
                     int igrid, double zexp, int nm, const int* lmax_per_m, int device) {
  HFH_TRY
  fem::PolyBasis poly(primbas, nnodes);
  if (nquad <= 0) nquad = 5 * poly.get_nbf();
  std::vector<int> lval, mval;
  for (int mabs = 0; mabs < nm; mabs++)
    for (int l = mabs; l <= lmax_per_m[mabs]; l++) {
      lval.push_back(l);
      mval.push_back(mabs);
      if (mabs > 0) { lval.push_back(l); mval.push_back(-mabs); }
    }
  const double Rhalf = 0.5 * Rbond;
  const double mumax = acosh(rmax / Rhalf);  // diatomic/main.cpp:275
  Vec bval = fem::get_grid(mumax, nelem, igrid, zexp);
  *out = new diatomic::TwoDBasis(Z1, Z2, Rhalf, poly, nquad, bval, lval, mval, device);
  HFH_CATCH
}
void hfh_diatomic_free(void* h) { delete (diatomic::TwoDBasis*)h; }
int hfh_diatomic_dims(void* h, int* nbf, int* ndummy, int* nrad, int* nang, int* nel, int* nlm) {
  HFH_TRY
  auto* b = (diatomic::TwoDBasis*)h;
  if (nbf) *nbf = (int)b->Nbf();
  if (ndummy) *ndummy = (int)b->Ndummy();
  if (nrad) *nrad = (int)b->Nrad();
  if (nang) *nang = (int)b->Nang();
  if (nel) *nel = (int)b->get_radial().Nel();
  if (nlm) *nlm = (int)b->get_lm_map().size();
  HFH_CATCH
}
int hfh_diatomic_tables(void* h, int* lval, int* mval, int* first, int* last, int* Lmap, int* Mmap) {
  HFH_TRY
  auto* b = (diatomic::TwoDBasis*)h;
  for (size_t a = 0; a < b->Nang(); a++) { lval[a] = b->get_lval()[a]; mval[a] = b->get_mval()[a]; }
  for (size_t e = 0; e < b->get_radial().Nel(); e++) {
    size_t f, l;
    b->get_radial().get_idx(e, f, l);
    first[e] = (int)f;
    last[e] = (int)l;
  }
  for (size_t r = 0; r < b->get_lm_map().size(); r++) { Lmap[r] = b->get_lm_map()[r].first; Mmap[r] = b->get_lm_map()[r].second; }
  HFH_CATCH
}
int hfh_diatomic_compute_tei(void* h, int exchange) {
  HFH_TRY
  ((diatomic::TwoDBasis*)h)->compute_tei(exchange != 0);
  HFH_CATCH
}
/// which: 0 P0, 1 P2, 2 Q0, 3 Q2 (np x np); 4+ab prim_tei block ab, 8+ab prim_ktei block ab (np^2 x np^2)
int hfh_diatomic_get_tei(void* h, int which, int ilm, int iel, double* out) {
  HFH_TRY
  auto* b = (diatomic::TwoDBasis*)h;
  const size_t idx = (size_t)ilm * b->get_radial().Nel() + iel;
  const Mat* m = nullptr;
  if (b->get_prim_tei().empty()) throw std::logic_error("Primitive teis have not been computed!\n");
  if (which < 4) m = &b->get_disjoint(which)[idx];
  else if (which < 8) m = &b->get_prim_tei()[4 * idx + which - 4];
  else {
    if (b->get_prim_ktei().empty()) throw std::logic_error("Primitive teis have not been computed!\n");
    m = &b->get_prim_ktei()[4 * idx + which - 8];
  }
  memcpy(out, m->ptr(), m->size() * sizeof(double));
  HFH_CATCH
}
// Zero-copy view of one stored block (same indexing as hfh_diatomic_get_tei): at BASELINE configs[3] the stored
// integrals are 9.5 GB, a second copy for the checker is not affordable.  The pointer lives as long as the basis.
int hfh_diatomic_tei_view(void* h, int which, int ilm, int iel, const double** ptr, long* n) {
  HFH_TRY
  auto* b = (diatomic::TwoDBasis*)h;
  const size_t idx = (size_t)ilm * b->get_radial().Nel() + iel;
  const Mat* m = nullptr;
  if (b->get_prim_tei().empty()) throw std::logic_error("Primitive teis have not been computed!\n");
  if (which < 4) m = &b->get_disjoint(which)[idx];
  else if (which < 8) m = &b->get_prim_tei()[4 * idx + which - 4];
  else {
    if (b->get_prim_ktei().empty()) throw std::logic_error("Primitive teis have not been computed!\n");
    m = &b->get_prim_ktei()[4 * idx + which - 8];
  }
  *ptr = m->ptr();
  *n = (long)m->size();
  HFH_CATCH
}
int hfh_diatomic_coulomb(void* h, const double* P, double* J) {
  HFH_TRY
  ((diatomic::TwoDBasis*)h)->coulomb(P, J);
  HFH_CATCH
}
int hfh_diatomic_exchange(void* h, const double* P, double* K) {
  HFH_TRY
  ((diatomic::TwoDBasis*)h)->exchange(P, K);
  HFH_CATCH
}

// ---- product-grid DFT (atomic / diatomic DFTGrid) ----
int hfh_atomic_grid_new(void** out, void* basis, int lang, int mang) {
  HFH_TRY
  *out = new atomic::DFTGrid((atomic::TwoDBasis*)basis, lang, mang);
  HFH_CATCH
}
void hfh_atomic_grid_free(void* g) { delete (atomic::DFTGrid*)g; }
int hfh_atomic_grid_eval(void* g, void* basis, int nrows, int x_func, int c_func, int nspin, const double* Pa, const double* Pb, double thr,
                         double* Ha, double* Hb, int beta, double* scal3) {
  return grid_eval<atomic::DFTGrid, atomic::TwoDBasis>(g, basis, nrows, x_func, c_func, nspin, Pa, Pb, thr, Ha, Hb, beta, scal3);
}
int hfh_diatomic_grid_new(void** out, void* basis, int lang, int mang) {
  HFH_TRY
  *out = new diatomic::DFTGrid((diatomic::TwoDBasis*)basis, lang, mang);
  HFH_CATCH
}
void hfh_diatomic_grid_free(void* g) { delete (diatomic::DFTGrid*)g; }
int hfh_diatomic_grid_eval(void* g, void* basis, int nrows, int x_func, int c_func, int nspin, const double* Pa, const double* Pb, double thr,
                           double* Ha, double* Hb, int beta, double* scal3) {
  return grid_eval<diatomic::DFTGrid, diatomic::TwoDBasis>(g, basis, nrows, x_func, c_func, nspin, Pa, Pb, thr, Ha, Hb, beta, scal3);
}
/// radial tables of one element: r, wrad (nq), bf, df (nq x np_el)
int hfh_diatomic_radial(void* h, int iel, double* r, double* wrad, double* bf, double* df) {
  HFH_TRY
  const auto& rad = ((diatomic::TwoDBasis*)h)->get_radial();
  const Vec rv = rad.get_r(iel), wv = rad.get_wrad(iel);
  const Mat b = rad.get_bf(iel), d = rad.get_df(iel);
  if (r) memcpy(r, rv.data(), rv.size() * sizeof(double));
  if (wrad) memcpy(wrad, wv.data(), wv.size() * sizeof(double));
  if (bf) memcpy(bf, b.ptr(), b.size() * sizeof(double));
  if (df) memcpy(df, d.ptr(), d.size() * sizeof(double));
  HFH_CATCH
}
int hfh_diatomic_nquad(void* h) { return ((diatomic::TwoDBasis*)h)->get_radial().get_nquad(); }

int hfh_atomic_set_shard(void* h, int rank, int n) {
  HFH_TRY
  ((atomic::TwoDBasis*)h)->set_shard(rank, n);
  HFH_CATCH
}
int hfh_diatomic_set_shard(void* h, int rank, int n) {
  HFH_TRY
  ((diatomic::TwoDBasis*)h)->set_shard(rank, n);
  HFH_CATCH
}

}  // extern "C"

// ---- test hooks of the host-side block scan / zero fill (host/blockscan.h; CPU-only, tests/test_blockscan_cpu.py) ----
#include "blockscan.h"
extern "C" {
void hfh_scan_nonzero_blocks(const double* M, int ld, int nang, const int* r0, const int* cnt, unsigned char* nz) {
  hfb::scan_nonzero_blocks(&M, 1, ld, nang, r0, cnt, nz);
}
void hfh_zero_fill(double* M, int ld, int n) { hfb::zero_fill(M, ld, n); }
}

// ---- the family -> rank map of the sharded operators (host/famshard.h), for tests and harnesses ----
#include "famshard.h"
extern "C" void hfh_family_owner(int nfam, int nranks, int* owner) {
  const std::vector<int> own = hfb::family_owner(nfam, nranks);
  for (int r = 0; r < nfam; r++) owner[r] = own[r];
}
