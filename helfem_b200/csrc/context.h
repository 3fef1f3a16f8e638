// Device context shared by the C-ABI entry points.
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>

namespace hfb {

void set_error(const char* fmt, ...);
const char* get_error();

#define HFB_CUDA_OK(call)                                                                   \
  do {                                                                                      \
    cudaError_t e_ = (call);                                                                \
    if (e_ != cudaSuccess) {                                                                \
      hfb::set_error("CUDA error %s at %s:%d: %s", cudaGetErrorName(e_), __FILE__, __LINE__, \
                     cudaGetErrorString(e_));                                               \
      return (e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver) ? -4 : -5;      \
    }                                                                                       \
  } while (0)

/// Growable device / pinned-host scratch buffer.
struct DevBuf {
  double* p = nullptr;
  size_t cap = 0;
  int reserve(size_t n) {
    if (n <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc(&p, n * sizeof(double));
    if (e != cudaSuccess) { set_error("cudaMalloc of %zu bytes failed: %s", n * sizeof(double), cudaGetErrorString(e)); return -5; }
    cap = n;
    return 0;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
struct PinBuf {
  double* p = nullptr;
  size_t cap = 0;
  int reserve(size_t n) {
    if (n <= cap) return 0;
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMallocHost(&p, n * sizeof(double));
    if (e != cudaSuccess) { set_error("cudaMallocHost of %zu bytes failed: %s", n * sizeof(double), cudaGetErrorString(e)); return -5; }
    cap = n;
    return 0;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

enum CtxKind { CTX_SADATOM = 1, CTX_ATOMIC = 2, CTX_DIATOMIC = 3 };
struct ChanEngine;
struct Grid2D;

}  // namespace hfb

// The opaque handle of include/helfem_b200.h
struct hfb200_ctx {
  int kind = 0;
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t xs[3] = {nullptr, nullptr, nullptr};  // chunk pipeline of the host-buffer batch calls
  // ---- radial tables ----
  int lmax = 0, nel = 0, nq = 0, nq8 = 0, ldq = 0, np_max = 0, np_pad = 0, nrad = 0;
  std::vector<int> first, last;
  double *dBt = nullptr, *ddBt = nullptr, *dR = nullptr, *dW = nullptr;
  int *dFirst = nullptr, *dNpe = nullptr, *dElemOf = nullptr;  // [nel], [nel], [nrad][2]
  double* dCpl = nullptr;  // sadatom exchange coupling [lout][lin][L]
  // ---- stored integrals (padded to np_max) ----
  int nL = 0;
  bool have_ktei = false;
  double *dDisL = nullptr, *dDisM1L = nullptr;  // [nL][nel][np2]            (exchange path)
  double* dKtei = nullptr;                      // [nL][nel][np2*np2]        (exchange path)
  // Coulomb path (L = 0): triangle-packed integrals in DMMA fragment order (sadatom_coulomb_mma.cuh)
  int ntri = 0, jMT = 0, jKS = 0;
  double *dTeiF = nullptr, *dLc = nullptr, *dm1Lc = nullptr;
  int2* dTri = nullptr;
  // ---- atomic / diatomic: angular basis + channel engine (chan_engine.h) ----
  std::vector<int> lval, mval;
  double Rhalf = 0.0;
  int shard_rank = 0, shard_n = 1;  // this context holds the integral families r with fam_owner[r] == shard_rank
  std::vector<int> fam_owner;       // [families] owning rank (host/famshard.h), filled by set_tei
  void* comm = nullptr;             // ncclComm_t of the sharded operators (comm.cu), or null: the caller sums the partial operators
  hfb::ChanEngine* chan = nullptr;
  hfb::Grid2D* g2d = nullptr;  // product-grid XC engine (grid2d.h)
  // ---- scratch ----
  hfb::DevBuf dP, dH, dJ, dPtot, dTraces, dPartial, dScalars, dK;
  hfb::PinBuf hIn, hOut;
};
