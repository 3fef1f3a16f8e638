// sm_100a kernels for the spherically-averaged-atom Fock build.
//
//   sad_grid_kernel      DFTGrid::eval_Fxc of the reference, fused per (build, element):
//                          compute_bf      sadatom/dftgrid.cpp:842-863   (tables are resident)
//                          update_density  sadatom/dftgrid.cpp:46-120 (R), :122-227 (U)
//                          compute_xc      sadatom/dftgrid.cpp:306-486   (device functionals)
//                          eval_Exc/Nel    sadatom/dftgrid.cpp:228-238,566-572
//                          eval_Fxc        sadatom/dftgrid.cpp:610-680 (R), :682-805 (U)
//   sad_trace_kernel     total radial density block + the two traces of coulomb()
//                          sadatom/basis.cpp:314-320
//   sad_coulomb_kernel   in-element GEMV + disjoint contributions   sadatom/basis.cpp:322-352
//
// Data layout ("element band"): every matrix of the path only ever touches the diagonal
// element blocks, so densities and Fock contributions live on the device as
// [build][spin][l][element][np_max x np_max] column-major zero-padded blocks.
#pragma once
#include <cuda_runtime.h>
#include "xc_point.cuh"

namespace hfb {

struct SadGridParams {
  // constant tables
  const double* Bt;    // [nel][np_max][ldq]  B_i(r_p)/r
  const double* dBt;   // [nel][np_max][ldq]  d/dr
  const double* r;     // [nel][nq]
  const double* wrad;  // [nel][nq]
  int nel, nq, ldq, np_max, nslices;
  int x_func, c_func;
  double dens_thr;
  // per batch
  const double* P;     // [nb][nspin][nslices][nel][np_max^2]
  double* H;           // same shape
  double* partial;     // [nb][nel][4]  Exc, Nel, tau integral, unused
};

__device__ __forceinline__ double quad_sum(double v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  return v;
}

// One CTA per (element, build).  FAM: highest functional family; NSPIN: 1 or 2;
// RPL: rows of the density block handled per lane of a 4-lane group (np_max <= 4*RPL).
template <int FAM, int NSPIN, int RPL>
__global__ void __launch_bounds__(256) sad_grid_kernel(SadGridParams prm) {
  extern __shared__ double smem[];
  const int el = blockIdx.x, b = blockIdx.y;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int np = prm.np_max, np2 = np * np, nq = prm.nq, ldq = prm.ldq;
  constexpr bool GGA = (FAM >= xc::FAM_GGA);
  constexpr bool MGGA = (FAM >= xc::FAM_MGGA);

  // ---- shared memory carve-up ----
  double* sB = smem;                                   // np*ldq
  double* sdB = sB + np * ldq;                         // np*ldq (GGA+)
  double* sP = sdB + (GGA ? np * ldq : 0);             // NSPIN*np2
  double* sPl = sP + NSPIN * np2;                      // NSPIN*np2 (MGGA)
  double* sRho = sPl + (MGGA ? NSPIN * np2 : 0);       // per-point arrays, each NSPIN*nq
  double* sGr = sRho + NSPIN * nq;
  double* sTau = sGr + (GGA ? NSPIN * nq : 0);
  double* sDl = sTau + (MGGA ? NSPIN * nq : 0);        // NSPIN*nq (MGGA)
  double* sRed = sDl + (MGGA ? NSPIN * nq : 0);        // 3*32 reduction scratch

  // ---- stage the element tables ----
  {
    const double* gB = prm.Bt + (size_t)el * np * ldq;
    for (int k = tid; k < np * ldq; k += nthr) sB[k] = gB[k];
    if (GGA) {
      const double* gdB = prm.dBt + (size_t)el * np * ldq;
      for (int k = tid; k < np * ldq; k += nthr) sdB[k] = gdB[k];
    }
  }
  // ---- sum the density blocks over l: P = sum_l P_l, Pl = sum_l l(l+1) P_l ----
  for (int s = 0; s < NSPIN; s++) {
    const double* gP = prm.P + ((size_t)(b * NSPIN + s) * prm.nslices * prm.nel + el) * np2;
    for (int k = tid; k < np2; k += nthr) {
      double p = 0.0, pl = 0.0;
      for (int l = 0; l < prm.nslices; l++) {
        const double v = gP[(size_t)l * prm.nel * np2 + k];
        p += v;
        pl += (double)(l * (l + 1)) * v;
      }
      sP[s * np2 + k] = p;
      if (MGGA) sPl[s * np2 + k] = pl;
    }
  }
  __syncthreads();

  // ---- density, gradient, tau on the quadrature points ----
  {
    const int quad = tid >> 2, ql = tid & 3, nquads = nthr >> 2;
    for (int base = 0; base < nq; base += nquads) {
      const int p = base + quad;
      const bool valid = p < nq;
      const int pp = valid ? p : 0;
      for (int s = 0; s < NSPIN; s++) {
        double v[RPL], vp[RPL], vl[RPL];
#pragma unroll
        for (int k = 0; k < RPL; k++) { v[k] = 0.0; vp[k] = 0.0; vl[k] = 0.0; }
        const double* Ps = sP + s * np2;
        const double* Pls = sPl + s * np2;
        for (int j = 0; j < np; j++) {
          const double bj = sB[j * ldq + pp];
          const double dbj = GGA ? sdB[j * ldq + pp] : 0.0;
#pragma unroll
          for (int k = 0; k < RPL; k++) {
            const int i = ql + 4 * k;
            if (i < np) {
              const double pij = Ps[i + j * np];
              v[k] = fma(pij, bj, v[k]);
              if (MGGA) {
                vp[k] = fma(pij, dbj, vp[k]);
                vl[k] = fma(Pls[i + j * np], bj, vl[k]);
              }
            }
          }
        }
        double rho = 0.0, g = 0.0, t1 = 0.0, t2 = 0.0;
#pragma unroll
        for (int k = 0; k < RPL; k++) {
          const int i = ql + 4 * k;
          if (i < np) {
            const double bi = sB[i * ldq + pp];
            rho = fma(v[k], bi, rho);
            if (GGA) {
              const double dbi = sdB[i * ldq + pp];
              g = fma(v[k], dbi, g);
              if (MGGA) {
                t1 = fma(vp[k], dbi, t1);
                t2 = fma(vl[k], bi, t2);
              }
            }
          }
        }
        rho = quad_sum(rho);
        if (GGA) g = quad_sum(g);
        if (MGGA) { t1 = quad_sum(t1); t2 = quad_sum(t2); }
        if (valid && ql == 0) {
          sRho[s * nq + p] = rho;
          if (GGA) sGr[s * nq + p] = 2.0 * g;
          if (MGGA) {
            const double rr = prm.r[(size_t)el * nq + p];
            sTau[s * nq + p] = 0.5 * (t1 + fmax(t2 / (rr * rr), 0.0));
          }
        }
      }
    }
  }
  __syncthreads();

  // ---- XC on the points; turn potentials into quadrature weights ----
  double my_exc = 0.0, my_nel = 0.0, my_tau = 0.0;
  {
    const xc::XcThresholds thr = xc::make_thresholds(prm.dens_thr);
    for (int p = tid; p < nq; p += nthr) {
      const double rr = prm.r[(size_t)el * nq + p];
      const double wr = prm.wrad[(size_t)el * nq + p];
      const double w4 = 4.0 * HFB_PI * wr;
      const double wtot = w4 * (rr * rr);
      if (NSPIN == 1) {
        const double rho = sRho[p];
        const double gr = GGA ? sGr[p] : 0.0;
        const double sigma = gr * gr;
        const double tau = MGGA ? sTau[p] : 0.0;
        double exc = 0.0, vrho = 0.0, vsigma = 0.0, vtau = 0.0;
        xc::xc_point_unpol<FAM>(prm.x_func, thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
        xc::xc_point_unpol<FAM>(prm.c_func, thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
        my_exc += wtot * exc * rho;
        my_nel += wtot * rho;
        my_tau += wtot * tau;
        sRho[p] = wtot * vrho;
        if (GGA) sGr[p] = 2.0 * wtot * vsigma * gr;
        if (MGGA) { sTau[p] = wtot * 0.5 * vtau; sDl[p] = vtau * (0.5 * w4); }
      } else {
        double rho[2] = {sRho[p], sRho[nq + p]};
        double gr[2] = {GGA ? sGr[p] : 0.0, GGA ? sGr[nq + p] : 0.0};
        double sigma[3] = {gr[0] * gr[0], gr[0] * gr[1], gr[1] * gr[1]};
        double tau[2] = {MGGA ? sTau[p] : 0.0, MGGA ? sTau[nq + p] : 0.0};
        double exc = 0.0, vrho[2] = {0.0, 0.0}, vsigma[3] = {0.0, 0.0, 0.0}, vtau[2] = {0.0, 0.0};
        xc::xc_point_pol<FAM>(prm.x_func, thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
        xc::xc_point_pol<FAM>(prm.c_func, thr, rho, sigma, tau, exc, vrho, vsigma, vtau);
        my_exc += wtot * exc * (rho[0] + rho[1]);
        my_nel += wtot * (rho[0] + rho[1]);
        my_tau += wtot * (tau[0] + tau[1]);
        sRho[p] = wtot * vrho[0];
        sRho[nq + p] = wtot * vrho[1];
        if (GGA) {
          sGr[p] = wtot * (2.0 * vsigma[0] * gr[0] + vsigma[1] * gr[1]);
          sGr[nq + p] = wtot * (2.0 * vsigma[2] * gr[1] + vsigma[1] * gr[0]);
        }
        if (MGGA) {
          sTau[p] = wtot * 0.5 * vtau[0];
          sTau[nq + p] = wtot * 0.5 * vtau[1];
          sDl[p] = vtau[0] * (0.5 * w4);
          sDl[nq + p] = vtau[1] * (0.5 * w4);
        }
      }
    }
  }
  // block reduction of the three scalars (fixed tree => run-to-run deterministic)
  {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      my_exc += __shfl_xor_sync(0xffffffffu, my_exc, o);
      my_nel += __shfl_xor_sync(0xffffffffu, my_nel, o);
      my_tau += __shfl_xor_sync(0xffffffffu, my_tau, o);
    }
    const int warp = tid >> 5, lane = tid & 31;
    if (lane == 0) { sRed[warp] = my_exc; sRed[32 + warp] = my_nel; sRed[64 + warp] = my_tau; }
  }
  __syncthreads();
  if (tid == 0) {
    double e = 0.0, n = 0.0, t = 0.0;
    for (int w = 0; w < (nthr >> 5); w++) { e += sRed[w]; n += sRed[32 + w]; t += sRed[64 + w]; }
    double* out = prm.partial + ((size_t)b * prm.nel + el) * 4;
    out[0] = e; out[1] = n; out[2] = t; out[3] = 0.0;
  }

  // ---- Fock block: H = B a B^T + (dB g B^T + B g dB^T) + dB c dB^T ; Hl = B dl B^T ----
  {
    const int npairs = np * (np + 1) / 2;
    for (int k = tid; k < npairs; k += nthr) {
      int j = (int)((sqrt(8.0 * k + 1.0) - 1.0) * 0.5);
      while ((j + 1) * (j + 2) / 2 <= k) j++;
      while (j * (j + 1) / 2 > k) j--;
      const int i = k - j * (j + 1) / 2;
      const double* bi = sB + i * ldq;
      const double* bj = sB + j * ldq;
      const double* dbi = sdB + i * ldq;
      const double* dbj = sdB + j * ldq;
      double h[NSPIN], hl[NSPIN];
#pragma unroll
      for (int s = 0; s < NSPIN; s++) { h[s] = 0.0; hl[s] = 0.0; }
      for (int p = 0; p < nq; p++) {
        const double bb = bi[p] * bj[p];
        double bd = 0.0, dd = 0.0;
        if (GGA) {
          bd = fma(dbi[p], bj[p], bi[p] * dbj[p]);
          if (MGGA) dd = dbi[p] * dbj[p];
        }
#pragma unroll
        for (int s = 0; s < NSPIN; s++) {
          h[s] = fma(sRho[s * nq + p], bb, h[s]);
          if (GGA) h[s] = fma(sGr[s * nq + p], bd, h[s]);
          if (MGGA) {
            h[s] = fma(sTau[s * nq + p], dd, h[s]);
            hl[s] = fma(sDl[s * nq + p], bb, hl[s]);
          }
        }
      }
#pragma unroll
      for (int s = 0; s < NSPIN; s++) {
        double* gH = prm.H + ((size_t)(b * NSPIN + s) * prm.nslices * prm.nel + el) * np2;
        for (int l = 0; l < prm.nslices; l++) {
          const double val = MGGA ? fma((double)(l * (l + 1)), hl[s], h[s]) : h[s];
          gH[(size_t)l * prm.nel * np2 + i + j * np] = val;
          gH[(size_t)l * prm.nel * np2 + j + i * np] = val;
        }
      }
    }
  }
}

inline size_t sad_grid_smem_bytes(int fam, int nspin, int np, int nq, int ldq) {
  const bool gga = fam >= xc::FAM_GGA, mgga = fam >= xc::FAM_MGGA;
  size_t n = (size_t)np * ldq * (gga ? 2 : 1);
  n += (size_t)nspin * np * np * (mgga ? 2 : 1);
  n += (size_t)nspin * nq * (1 + (gga ? 1 : 0) + (mgga ? 2 : 0));
  n += 96;
  return n * sizeof(double);
}

// ---------------------------------------------------------------------------
// Coulomb
// ---------------------------------------------------------------------------
struct SadCoulombParams {
  const double* dL;     // [nel][np2]   disjoint_L   (L=0)
  const double* dm1L;   // [nel][np2]   disjoint_m1L (L=0)
  const double* tei;    // [nel][np2*np2]
  int nel, np_max, nslices, nspin;
  const double* P;      // [nb][nspin][nslices][nel][np2]
  double* Ptot;         // [nb][nel][np2]   scratch: total density block
  double* traces;       // [nb][nel][2]     scratch: jsmall, jbig
  double* J;            // [nb][nel][np2]
};

// grid (nel, nb): total density block and the two traces.
__global__ void __launch_bounds__(256) sad_trace_kernel(SadCoulombParams prm) {
  __shared__ double red[64];
  const int el = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  const int np = prm.np_max, np2 = np * np;
  const double Lfac = 4.0 * HFB_PI;
  double ts = 0.0, tb = 0.0;
  for (int k = tid; k < np2; k += blockDim.x) {
    double p = 0.0;
    for (int s = 0; s < prm.nspin; s++)
      for (int l = 0; l < prm.nslices; l++)
        p += prm.P[(((size_t)(b * prm.nspin + s) * prm.nslices + l) * prm.nel + el) * np2 + k];
    prm.Ptot[((size_t)b * prm.nel + el) * np2 + k] = p;
    // trace(D * P) = sum_ij D(i,j) P(j,i)
    const int i = k % np, j = k / np;
    const double pt = 0.0;
    (void)pt;
    // P(j,i) is element i*np + j of the same block; re-read it summed below through symmetry-free path
    double pji = 0.0;
    for (int s = 0; s < prm.nspin; s++)
      for (int l = 0; l < prm.nslices; l++)
        pji += prm.P[(((size_t)(b * prm.nspin + s) * prm.nslices + l) * prm.nel + el) * np2 + j + i * np];
    ts = fma(prm.dL[(size_t)el * np2 + k], pji, ts);
    tb = fma(prm.dm1L[(size_t)el * np2 + k], pji, tb);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ts += __shfl_xor_sync(0xffffffffu, ts, o);
    tb += __shfl_xor_sync(0xffffffffu, tb, o);
  }
  if ((tid & 31) == 0) { red[tid >> 5] = ts; red[32 + (tid >> 5)] = tb; }
  __syncthreads();
  if (tid == 0) {
    double a = 0.0, c = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) { a += red[w]; c += red[32 + w]; }
    prm.traces[((size_t)b * prm.nel + el) * 2 + 0] = Lfac * a;
    prm.traces[((size_t)b * prm.nel + el) * 2 + 1] = Lfac * c;
  }
}

// grid (nel, nb): J_el = Lfac * tei_el * vec(P_el) + dL_el * sum_{j>el} jbig_j + dm1L_el * sum_{j<el} jsmall_j
__global__ void __launch_bounds__(256) sad_coulomb_kernel(SadCoulombParams prm) {
  extern __shared__ double sPv[];
  const int el = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  const int np = prm.np_max, np2 = np * np;
  const double Lfac = 4.0 * HFB_PI;
  const double* Pt = prm.Ptot + ((size_t)b * prm.nel + el) * np2;
  for (int k = tid; k < np2; k += blockDim.x) sPv[k] = Pt[k];
  __syncthreads();
  double below = 0.0, above = 0.0;  // sum of jsmall for j<el, of jbig for j>el
  const double* tr = prm.traces + (size_t)b * prm.nel * 2;
  for (int j = 0; j < el; j++) below += tr[2 * j];
  for (int j = el + 1; j < prm.nel; j++) above += tr[2 * j + 1];
  const double* T = prm.tei + (size_t)el * np2 * np2;
  for (int row = tid; row < np2; row += blockDim.x) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    int c = 0;
    for (; c + 3 < np2; c += 4) {
      a0 = fma(T[(size_t)c * np2 + row], sPv[c], a0);
      a1 = fma(T[(size_t)(c + 1) * np2 + row], sPv[c + 1], a1);
      a2 = fma(T[(size_t)(c + 2) * np2 + row], sPv[c + 2], a2);
      a3 = fma(T[(size_t)(c + 3) * np2 + row], sPv[c + 3], a3);
    }
    for (; c < np2; c++) a0 = fma(T[(size_t)c * np2 + row], sPv[c], a0);
    double acc = Lfac * ((a0 + a1) + (a2 + a3));
    acc = fma(above, prm.dL[(size_t)el * np2 + row], acc);
    acc = fma(below, prm.dm1L[(size_t)el * np2 + row], acc);
    prm.J[((size_t)b * prm.nel + el) * np2 + row] = acc;
  }
}

// [nb][nel][4] partial scalars -> [nb][4], fixed order
__global__ void sad_reduce_scalars_kernel(const double* partial, int nel, double* out) {
  const int b = blockIdx.x;
  if (threadIdx.x < 4) {
    double s = 0.0;
    for (int e = 0; e < nel; e++) s += partial[((size_t)b * nel + e) * 4 + threadIdx.x];
    out[(size_t)b * 4 + threadIdx.x] = s;
  }
}

}  // namespace hfb
