// Host-side driver of the channel-coupled Coulomb / exchange kernels (chan_kernels.cuh,
// chan_exchange.cuh) shared by the `atomic` and `diatomic` C-ABI entry points.
#pragma once
#include <functional>
#include <vector>
#include "chan_exchange.cuh"
#include "chan_exchange_fused.cuh"
#include "chan_kernels.cuh"
#include "context.h"

namespace hfb {

/// One coupling of an ordered angular pair (k,l) to a channel: Paux_c += coef[h] * P_{kl} (h = stack half).
struct ChanCoupling {
  int chan;
  double coef[2];
};

/// Host description of a coupled-channel problem (everything constant across SCF iterations).
struct ChanSpec {
  int nel = 0, nang = 0, stack = 1, nr = 0;
  std::vector<int> first, last;            // element index ranges (FiniteElementBasis::get_idx)
  std::vector<double> fac;                 // [nr]  4 pi/(2L+1)  resp.  LMfac
  std::vector<int> chan_r;                 // [nchan] family of each channel, sorted by family
  std::vector<std::vector<ChanCoupling>> pair_coupl;  // [nang*nang], index k*nang + l
  std::vector<int> ang_group;              // [nang] optional: angular functions with equal value (m) share their coupling
                                           // partners; only used to tile the exchange contractions efficiently
  std::vector<int> ang_m;                  // [nang] m of every angular function (which blocks of K can be non-zero for a given P)
  std::vector<char> fam_owned;             // [nr] multi-GPU: this rank holds the integrals of family r (empty: all).  pair_coupl lists the
                                           // couplings of ALL families: the bra contraction of a communicator-attached context needs them
  // in-element integrals of family r, element e, block (hout, hin) of the stacked matrix:
  // returns the (np_e^2 x np_e^2) column-major block and its sign, nullptr if the block is absent
  std::function<const double*(int r, int e, int hout, int hin, double* sign)> tei;
  // K-ordered in-element integrals (exchange), same indexing; may be empty
  std::function<const double*(int r, int e, int hout, int hin, double* sign)> ktei;
  // disjoint factors (np_e x np_e): "small" (r_<-type: dL, P0/P2) and "big" (r_>-type: dm1L, Q0/Q2)
  std::function<const double*(int r, int e, int h, double* sign)> small_f, big_f;
};

struct ChanEngine {
  struct DevList { void* p = nullptr; size_t cap = 0; };
  ChanParams prm{};                 // device pointers + sizes
  int nrad = 0, nchan = 0, npair = 0;
  std::vector<int> first, last;
  std::vector<void*> owned;         // device allocations to free
  DevBuf dP, dPc, dTr, dJc, dJ;
  PinBuf hIn, hOut;
  std::vector<int> fam_c0_host;
  void* comm = nullptr;             // ncclComm_t (comm.cu) or null
  int comm_size = 1;                // ranks of the communicator; comm_rank: this rank (set with comm)
  int comm_root = -1, comm_rank = 0; // root-driven host-matrix calls (hfb200_comm_set_root); -1: every rank passes the matrices
  DevBuf dHdr;                      // broadcast header of a root-driven call: [count, (i, l) ...]
  std::vector<double> hdr_host;
  std::vector<char> fam_owned;      // [nr]
  std::vector<int> ang_m;           // [nang]
  // bra-contraction CSR over the couplings of ALL families (prm.post_* covers the owned ones only)
  const int *dPostAllPtr = nullptr, *dPostAllChan = nullptr;
  const double* dPostAllCoef = nullptr;
  // exchange all-reduce: the angular blocks of K that can be non-zero for the cached mask of P, packed
  std::vector<int2> red_blocks;
  DevList dRedBlocks;
  DevBuf dRed;
  // exchange tables (chan_exchange.cuh)
  bool have_k = false;
  ChanXParams xprm{};
  int nch_max = 0;
  std::vector<double> fam_fac;                      // [nr]
  std::vector<int> xrowptr, xrow_i;                 // host copy of the coupling rows (item lists are derived from them)
  std::vector<double> xrow_coef;                    // [nnz][2]
  std::vector<int> fam_rows;                        // [nr] non-empty coupling rows of each family
  // per sparsity pattern of P (build_items): output blocks of every family, compact W layout, tiles, batches
  std::vector<unsigned char> mask_cached;
  struct XBatch { int t1_0, t1_1, t2_0, t2_1, gs0, gs1, gp0, gp1, item0, item1, out0, nout; long long wblocks; };
  std::vector<XBatch> batches;
  long long total_items = 0, max_wblocks = 0;
  int max_batch_items = 0;
  DevList dItems, dItemFam, dTiles2, dGroupsS, dGroupsP, dTileU, dSlotWblk, dSlotCoef, dWbase, dWbits, dOutPtr, dOutItem, dX1tab;
  int x1tab_ldp = -1;                               // leading dimension dX1tab was built for
  int ni_same = 32, ni_pair = 16;                   // items per CTA of the fused same-element / element-pair kernels
  bool last_sym = false;                            // the last exchange call used the symmetric (j <= k) item list
  const double* dFamFac = nullptr;
  std::vector<int> fam_t1;                          // [nr+1] x1 tiles of each family in dTiles1 ([c, nj, j x CHX_T] each)
  int* dTiles1 = nullptr;
  std::vector<int> ang_group;                       // [nang]
  unsigned char *dNzP = nullptr, *hNzP = nullptr;
  DevBuf dPdense, dKdense, dKblocks, dW;

  int build_items(const unsigned char* nz, bool sym);
  int build_exchange(const ChanSpec& spec);
  int build(const ChanSpec& spec, bool with_exchange);
  // dense HOST matrices; map (optional, length nang*nrad): operator index -> row/column of the caller's nout x nout
  // matrices, -1 = identically zero row (diatomic boundary functions, diatomic/basis.cpp:1735-1773)
  int fock_device(const double* dP, int ldp, double* dJ, int ldj, double* dK, int ldk, cudaStream_t st);  // J and K of one build, J on a second stream
  cudaStream_t st2 = nullptr;
  cudaEvent_t evFork = nullptr, evJoin = nullptr;
  int dense_apply(bool exchange, const double* P, int ldp, double* F, int ldf, cudaStream_t st, const int* map = nullptr, int nout = 0);
  int coulomb_dense(const double* P, int ldp, double* J, int ldj, cudaStream_t st, const int* map = nullptr, int nout = 0) { return dense_apply(false, P, ldp, J, ldj, st, map, nout); }
  int exchange_dense(const double* P, int ldp, double* K, int ldk, cudaStream_t st, const int* map = nullptr, int nout = 0) { return dense_apply(true, P, ldp, K, ldk, st, map, nout); }
  DevList dIoBlocks;                // block list of the host-matrix calls (input blocks, then output blocks)
  PinBuf hX;
  // device-resident variants: dense (nang*nrad)^2 matrices already in HBM; every kernel of the operator, no PCIe
  int coulomb_device(const double* dPm, int ldp, double* dJm, int ldj, cudaStream_t st, const unsigned char* nzx = nullptr);  // nzx: exact block mask of P
  std::function<int()> on_mask_ready;  // fock_device: called by exchange_device once the block mask of P is on the device
  unsigned char* dNzX = nullptr;       // [nang^2] the block of P has an entry that is not exactly zero (filled with the mask)
  int exchange_device(const double* dPm, int ldp, double* dKm, int ldk, cudaStream_t st);
  int coulomb_kernels(int nang, cudaStream_t st);
  const int *dFirstJ = nullptr, *dNpeJ = nullptr, *dElemOfJ = nullptr;
  unsigned char *dPnz = nullptr, *dCnz = nullptr, *dOnz = nullptr;  // zero-screening flags of the Coulomb kernels (chan_kernels.cuh)
  // statistics of the last exchange call (bench.py): output blocks summed over families, algorithmic flops of x3
  long long last_items = 0;
  long long counts[12] = {0};       // per-pattern list sizes, see hfb200_chan_exchange_counts
  int nbatches() const { return (int)batches.size(); }
  double last_x3_flops = 0.0, last_same_flops = 0.0;
  // optional per-stage device timing of exchange_device (x0 mask, x1, x2, x3, x4), summed over families
  bool profile_stages = false;
  float stage_ms[6] = {0, 0, 0, 0, 0, 0};  // mask, x1, fused same-element, fused pairs, x4, all-reduce
  std::vector<cudaEvent_t> ev_pool;
  void release();
};

}  // namespace hfb
