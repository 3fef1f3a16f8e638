/* helfem_b200 -- C ABI of the B200 Fock-build library (libhelfem_b200.so).
 *
 * The reference (kluonj/HelFEM) has no FFI: its hot path sits behind C++ member functions
 * that take and return Armadillo objects.  This header is the thin C layer those member
 * functions bind to in the drop-in build (see INTEGRATION.md for the 10-line shims):
 *
 *   helfem::sadatom::basis::TwoDBasis::coulomb(P)      sadatom/basis.cpp:296   -> hfb200_sadatom_coulomb
 *   helfem::sadatom::basis::TwoDBasis::exchange(P)     sadatom/basis.cpp:360   -> hfb200_sadatom_exchange
 *   helfem::sadatom::dftgrid::DFTGrid::eval_Fxc(...)   sadatom/dftgrid.cpp:874,941 -> hfb200_sadatom_eval_fxc
 *   helfem::atomic::basis::TwoDBasis::coulomb/exchange atomic/TwoDBasis.cpp:817,957 -> hfb200_atomic_coulomb/_exchange
 *   helfem::atomic::dftgrid::DFTGrid::eval_Fxc(...)    atomic/dftgrid.cpp:810,872  -> hfb200_atomic_eval_fxc
 *   helfem::diatomic::basis::TwoDBasis::coulomb/exchange diatomic/basis.cpp:1359,1532 -> hfb200_diatomic_*
 *   helfem::diatomic::dftgrid::DFTGrid::eval_Fxc(...)  diatomic/dftgrid.cpp:769,812 -> hfb200_diatomic_eval_fxc
 *
 * Conventions: all matrices are FP64 column-major with an explicit leading dimension (what
 * arma::mat::memptr() gives); every function returns 0 on success and a negative HFB200_E*
 * code otherwise and never throws; hfb200_last_error() returns the message the C++ shim
 * re-throws as the exception type the reference uses (std::logic_error for missing
 * integrals / size mismatch, std::runtime_error for an unsupported functional).
 * One opaque context per basis object holds every constant of the path on the device
 * (basis tables on the quadrature grid, weights, stored radial integrals); per call only the
 * density goes down and the Fock contribution comes back.  There is NO CPU fallback: without
 * a CUDA device every compute entry point fails with HFB200_ENODEVICE.
 */
#ifndef HELFEM_B200_H
#define HELFEM_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define HFB200_OK 0
#define HFB200_EINVAL (-1)    /* size / argument mismatch      (reference: std::logic_error)   */
#define HFB200_ENOTEI (-2)    /* integrals not uploaded        (reference: "Primitive teis have not been computed!") */
#define HFB200_EFUNC (-3)     /* unsupported functional        (reference: std::runtime_error)  */
#define HFB200_ENODEVICE (-4) /* no CUDA device / CUDA failure; the library never falls back to the CPU */
#define HFB200_ECUDA (-5)

/* XC functionals implemented on the device (named as on the reference's command line,
 * general/dftfuncs.cpp:94-117).  Anything else yields HFB200_EFUNC. */
#define HFB200_XC_NONE 0
#define HFB200_XC_LDA_X 1
#define HFB200_XC_LDA_C_PW 12
#define HFB200_XC_LDA_C_PW_MOD 13
#define HFB200_XC_GGA_X_PBE 101
#define HFB200_XC_GGA_C_PBE 130
#define HFB200_XC_MGGA_X_R2SCAN 497
#define HFB200_XC_MGGA_C_R2SCAN 498

typedef struct hfb200_ctx hfb200_ctx;

const char* hfb200_last_error(void);
const char* hfb200_version(void);
/* "lda_x" -> HFB200_XC_LDA_X ...; -1 if the functional has no device implementation. */
int hfb200_xc_id(const char* libxc_name);
/* Number of visible CUDA devices (0 if none; never an error). */
int hfb200_device_count(void);

/* ------------------------------------------------------------------------------------------
 * Radial context: the spherically averaged atom (sadatom) and the radial part of `atomic`.
 *
 * nel elements, nq quadrature points per element, element iel owns the radial functions
 * first[iel]..last[iel] (FiniteElementBasis::get_idx).  r, wrad: nel*nq values
 * (RadialBasis::get_r / get_wrad).  bf, df, lf: for each element the (nq x np_iel) column-major
 * matrix RadialBasis::get_bf/get_df/get_lf(iel) returns, concatenated over elements (lf may be
 * NULL: no supported functional needs the Laplacian).  lmax: largest l of the density cube.
 * ------------------------------------------------------------------------------------------ */
int hfb200_sadatom_create(hfb200_ctx** ctx, int device, int lmax, int nel, int nq, const int* first,
                          const int* last, const double* r, const double* wrad, const double* bf,
                          const double* df, const double* lf);
/* Stored integrals built by TwoDBasis::compute_tei (sadatom/basis.cpp:181-218).  Arrays of
 * nL*nel pointers, entry [L*nel+iel]: disjoint_L / disjoint_m1L (np x np), prim_tei / prim_ktei
 * (np^2 x np^2).  prim_ktei may be NULL if exchange() is never called. */
int hfb200_sadatom_set_tei(hfb200_ctx* ctx, int nL, const double* const* disjoint_L,
                           const double* const* disjoint_m1L, const double* const* prim_tei,
                           const double* const* prim_ktei);
void hfb200_destroy(hfb200_ctx* ctx);
/* Multi-GPU sharding of the atomic / diatomic operators (one process per GPU): the context only holds and
 * contracts the integral families r (atomic: r = L; diatomic: r = index of (L,|M|) in lm_map) with
 * r % nranks == rank (helfem_b200/host/famshard.h).  *_set_tei reads the stored integrals of the owned
 * families only (hfb200_shard_owner returns the map; the other pointers may be NULL).  The partial J / K
 * of the ranks add up to the full operator; they are all-reduced inside the library once a communicator is
 * attached (below), else by the caller.  Call before *_set_tei. */
int hfb200_set_shard(hfb200_ctx* ctx, int rank, int nranks);
/* owner[r] = rank holding family r, for r < n (valid after *_set_tei) */
int hfb200_shard_owner(hfb200_ctx* ctx, int* owner, int n);

/* NCCL communicator of a sharded context (one process per GPU, like `mpirun`-style SPMD drivers; the reference's own drivers
 * are single-process, diatomic/main.cpp:810-828).  Rank 0 obtains an id with hfb200_comm_unique_id (HFB200_COMM_ID_BYTES
 * bytes) and hands it to the other ranks by any means; every rank then calls hfb200_comm_init on its context, sharded
 * as the same (rank, nranks) with hfb200_set_shard.  From then on hfb200_chan_coulomb_device / hfb200_chan_exchange_device
 * and the host-matrix calls of that context return the COMPLETE operator on every rank: the partial results are
 * all-reduced inside the library on the call's stream over NVLink (coulomb: the per-channel radial results, a few MB;
 * exchange: the angular blocks of K that can be non-zero for the sparsity pattern of P, packed).  The density must be
 * the same on every rank.  NCCL is resolved with dlopen at the first hfb200_comm_* call. */
#define HFB200_COMM_ID_BYTES 128
int hfb200_comm_unique_id(void* id, int nbytes);
int hfb200_comm_init(hfb200_ctx* ctx, const void* id, int rank, int nranks);
int hfb200_comm_destroy(hfb200_ctx* ctx);
/* Root-driven host-matrix calls (no reference counterpart: the reference is a single-process program).  After
 * hfb200_comm_set_root(ctx, root) on EVERY rank, only rank `root` calls hfb200_{atomic,diatomic}_coulomb / _exchange with host
 * matrices -- one SCF process driving all GPUs; it broadcasts the non-zero angular blocks of P to the other ranks over
 * NCCL/NVLink, every rank computes its integral families, the all-reduce runs as before and the root returns the complete J / K.
 * The other ranks call hfb200_chan_serve(ctx, exchange) once per root call (0: coulomb, 1: exchange), in the same order.
 * root < 0 switches back to the default (every rank passes the same host matrix and gets the complete result). */
int hfb200_comm_set_root(hfb200_ctx* ctx, int root);
int hfb200_chan_serve(hfb200_ctx* ctx, int exchange);

/* J = coulomb(P): P, J are nrad x nrad (host). */
int hfb200_sadatom_coulomb(hfb200_ctx* ctx, const double* P, int ldp, double* J, int ldj);
/* K = exchange(P): P, K are nrad x nrad x nslices cubes (slice stride ld*nrad); returns -K
 * like the reference. */
int hfb200_sadatom_exchange(hfb200_ctx* ctx, const double* P, int ldp, int nslices, double* K, int ldk);
/* eval_Fxc: nspin=1 restricted (Pb, Hb ignored), nspin=2 unrestricted (beta!=0 also fills Hb).
 * P*, H* are nrad x nrad x nslices cubes; H is fully overwritten (zero outside the element
 * blocks, like H.zeros() in the reference). */
int hfb200_sadatom_eval_fxc(hfb200_ctx* ctx, int x_func, int c_func, int nspin, const double* Pa,
                            const double* Pb, int ldp, int nslices, double dens_thr, double* Ha,
                            double* Hb, int ldh, int beta, double* Exc, double* Nel);

/* ---- batched, device-resident form of the same three operators ---------------------------
 * Many independent Fock builds on the same basis (trial configurations of one atom, atoms of
 * a sweep on a common grid, spin channels) in one pass.  Densities and results stay on the
 * device in "element-band" layout: for build b, spin s, slice l, element e the np_max x np_max
 * block (column-major, zero padded) at offset (((b*nspin+s)*nslices+l)*nel+e)*np_max^2 holds
 * M(first[e]+i, first[e]+j).  J bands are [b][e][np_max^2]; scalars are [b][4] = Exc, Nel, Ekin-like
 * tau integral, reserved.  Pointers are DEVICE pointers; stream is a cudaStream_t (0 = default). */
int hfb200_sadatom_band_size(hfb200_ctx* ctx, int* nel, int* np_max);
int hfb200_sadatom_pack_band(hfb200_ctx* ctx, const double* M /*host nrad x nrad*/, int ld, double* band /*host*/);
int hfb200_sadatom_unpack_band_add(hfb200_ctx* ctx, const double* band /*host*/, double* M, int ld);
int hfb200_sadatom_fock_batch(hfb200_ctx* ctx, int nbatch, int x_func, int c_func, int nspin, int nslices,
                              double dens_thr, const double* dP_band, double* dJ_band, double* dH_band,
                              double* dScalars, void* stream);

/* Allocation-free variant of hfb200_sadatom_fock_batch.  It has no counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments, kernels and results as hfb200_sadatom_fock_batch, but
 * the per-build scratch (folded densities, disjoint traces, partial XC scalars) lives in one static device array of
 * HFB200_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit; the call makes no cudaMalloc /
 * cudaFree and calls no allocating library routine.  Batches whose scratch exceeds the array (about
 * (nel*(4*ceil8(np(np+1)/2)) + 6*nel)*8 bytes per build: 416 builds of configs[1], 54 of configs[2]) are refused with
 * HFB200_EINVAL.  All calls on a device share the array, which keeps the last call's contents (the range in use is
 * cleared on `stream` first): calls that overlap on the GPU race -- serialise them on one stream. */
#define HFB200_STATIC_SCRATCH_BYTES (8u << 20)
int hfb200_sadatom_fock_batch_static(hfb200_ctx* ctx, int nbatch, int x_func, int c_func, int nspin, int nslices,
                                     double dens_thr, const double* dP_band, double* dJ_band, double* dH_band,
                                     double* dScalars, void* stream);

/* The same batch with HOST element-band buffers (layout as above): one H2D copy of the densities, the
 * kernels, D2H copies of J / H bands and scalars; returns when the results are on the host.  hJ_band,
 * hH_band, hScalars may each be NULL to skip that operator. */
int hfb200_sadatom_fock_host(hfb200_ctx* ctx, int nbatch, int x_func, int c_func, int nspin, int nslices,
                             double dens_thr, const double* hP_band, double* hJ_band, double* hH_band,
                             double* hScalars);

/* ------------------------------------------------------------------------------------------
 * `atomic` program: product basis (l,m) x radial, global index iang*nrad + irad
 * (atomic/TwoDBasis.cpp:290-300).  lval/mval: the angular list of atomic::basis::angular_basis
 * (atomic/basis.cpp:174-198).  Stored integrals as built by TwoDBasis::compute_tei
 * (atomic/TwoDBasis.cpp:666-739) for L = 0..2 max(lval): arrays of nL*nel pointers, entry [L*nel+iel]
 * (the reference's in-element index Nel*Nel*L + iel*Nel + iel).  prim_ktei may be NULL.
 * P, J, K: dense (nang*nrad)^2 host matrices; exchange returns -K like the reference.
 * ------------------------------------------------------------------------------------------ */
int hfb200_atomic_create(hfb200_ctx** ctx, int device, int nang, const int* lval, const int* mval, int nel,
                         const int* first, const int* last);
int hfb200_atomic_set_tei(hfb200_ctx* ctx, int nL, const double* const* disjoint_L, const double* const* disjoint_m1L,
                          const double* const* prim_tei, const double* const* prim_ktei);
int hfb200_atomic_coulomb(hfb200_ctx* ctx, const double* P, int ldp, double* J, int ldj);
int hfb200_atomic_exchange(hfb200_ctx* ctx, const double* P, int ldp, double* K, int ldk);

/* ------------------------------------------------------------------------------------------
 * `diatomic` program (prolate spheroidal coordinates).  lval/mval: diatomic::basis::lm_to_l_m
 * (diatomic/basis.cpp:287-302); radial = FEM in mu.  Lmap/Mmap: the sorted (L,|M|) table lm_map of
 * TwoDBasis (diatomic/basis.cpp:336-376).  Stored integrals of TwoDBasis::compute_tei
 * (diatomic/basis.cpp:1166-1302): disjoint_{P0,P2,Q0,Q2} as arrays of nlm*nel pointers, entry
 * [ilm*nel+iel]; in-element blocks prim_tei{00,02,20,22} as one array of 4*nlm*nel pointers, entry
 * [4*(ilm*nel+iel) + 2a+b] for family "ab" (a, b in {0,1} standing for 0 and 2); prim_ktei likewise or NULL.
 * P, J, K: dense host matrices in the PURE basis (first radial function of every m != 0 shell
 * removed, diatomic/basis.cpp:482-499); exchange returns -K like the reference.
 * ------------------------------------------------------------------------------------------ */
int hfb200_diatomic_create(hfb200_ctx** ctx, int device, int nang, const int* lval, const int* mval, int nel,
                           const int* first, const int* last, double Rhalf);
int hfb200_diatomic_set_tei(hfb200_ctx* ctx, int nlm, const int* Lmap, const int* Mmap, const double* const* P0,
                            const double* const* P2, const double* const* Q0, const double* const* Q2,
                            const double* const* prim_tei, const double* const* prim_ktei);
int hfb200_diatomic_coulomb(hfb200_ctx* ctx, const double* P, int ldp, double* J, int ldj);
int hfb200_diatomic_exchange(hfb200_ctx* ctx, const double* P, int ldp, double* K, int ldk);

/* Device-resident variants of the atomic / diatomic operators: dP, dJ, dK are DEVICE pointers to dense
 * column-major matrices of dimension hfb200_chan_dim() = nang*nrad (atomic: the operator basis itself,
 * atomic/TwoDBasis.cpp:290-300; diatomic: the "dummy" basis of expand_boundaries, diatomic/basis.cpp:1755-1773),
 * `stream` a cudaStream_t (NULL: the context's stream).  Same kernels as the host-matrix calls above without the
 * PCIe legs; the work is enqueued on `stream` (exchange synchronises it once to read back the block mask of P).
 * hfb200_chan_exchange_stats: output blocks (summed over integral families) and algorithmic flops of the
 * radial contraction of the last exchange call; stage_ms (mask, x1 left angular, fused same-element, fused element pairs, x4 assembly) is filled when
 * hfb200_chan_set_profile(ctx, 1) is on. */
int hfb200_chan_dim(hfb200_ctx* ctx, int* ndim);
int hfb200_chan_coulomb_device(hfb200_ctx* ctx, const double* dP, int ldp, double* dJ, int ldj, void* stream);
int hfb200_chan_exchange_device(hfb200_ctx* ctx, const double* dP, int ldp, double* dK, int ldk, void* stream);
/* J and K of ONE Fock build from the same device-resident P (no reference counterpart: the reference calls coulomb and exchange one after
 * the other, diatomic/main.cpp SCF loop): the Coulomb kernels run on a second stream of the context next to the exchange pipeline; the
 * call is ordered on `stream` on both sides.  Results are identical to the two separate calls. */
int hfb200_chan_fock_device(hfb200_ctx* ctx, const double* dP, int ldp, double* dJ, int ldj, double* dK, int ldk, void* stream);
int hfb200_chan_exchange_stats(hfb200_ctx* ctx, long long* items, double* x3_flops, double* stage_ms /* [5] or NULL */);
/* sizes of the per-sparsity-pattern lists of the last exchange call (bench accounting): out[0] output blocks contracted (summed
 * over families), [1] tiles, [2] W sub-block slots read by the fused kernels, [3] the same padded to whole pipeline stages, [4] W blocks
 * written by x1, [5] doubles per stack half of a W block, [6] distinct output blocks summed over batches, [7] family batches,
 * [8] 1 if the symmetric (j <= k) list was used, [9] sum over tiles of slots x items (angular multiply-adds per sub-block entry and
 * stack pair), [10] angular blocks in the all-reduce payload. */
int hfb200_chan_exchange_counts(hfb200_ctx* ctx, long long* out, int n);
int hfb200_chan_allreduce_ms(hfb200_ctx* ctx, double* ms); /* with profiling on: gather + NCCL all-reduce + scatter of the last exchange call */
int hfb200_chan_exchange_batches(hfb200_ctx* ctx, int* nbatches); /* family batches of the last exchange call (5 launches each) */
int hfb200_chan_set_profile(hfb200_ctx* ctx, int on); /* per-stage CUDA-event timing of exchange (mask, x1..x4) */

/* ------------------------------------------------------------------------------------------
 * DFTGrid::eval_Fxc of the `atomic` and `diatomic` programs (angular x radial product grid).
 *   atomic::dftgrid::DFTGrid::eval_Fxc     atomic/dftgrid.cpp:810 (restricted), :872 (unrestricted)
 *   diatomic::dftgrid::DFTGrid::eval_Fxc   diatomic/dftgrid.cpp:769, :812
 * hfb200_set_grid uploads the constants DFTGridWorker::compute_bf rebuilds on every call in the
 * reference (atomic/dftgrid.cpp:710-796, diatomic/dftgrid.cpp:669-755), once per context (atomic or
 * diatomic):  r, wrad: nel*nq (radial.get_r / get_wrad);  bf, df: per element (nq x np_el) column-major,
 * concatenated (radial.get_bf / get_df);  the compound angular rule of angular::angular_chebyshev(lang, mang)
 * (general/angular.cpp:64-71): cth, wang [nangpts];  ylm, dth_ylm: (nang x nangpts) complex as (re,im)
 * pairs, entry [a + nang*ia]: Y_l^m and its theta derivative (atomic/TwoDBasis.cpp:1407-1419) at point ia.
 * eval_fxc: atomic matrices are dense (nang*nrad)^2; diatomic ones are in the PURE basis like
 * hfb200_diatomic_coulomb.  nspin = 1: Pa = P, Ha = H (Pb, Hb ignored).  nspin = 2: beta = 0 skips Hb
 * (the reference's `bool beta`).  Outputs are overwritten (H.zeros in the reference).  Ekin = integral of
 * tau (0 unless meta-GGA, atomic/dftgrid.cpp:262-275).
 * ------------------------------------------------------------------------------------------ */
int hfb200_set_grid(hfb200_ctx* ctx, int nq, const double* r, const double* wrad, const double* bf, const double* df,
                    int nangpts, const double* cth, const double* wang, const double* ylm, const double* dth_ylm);
int hfb200_atomic_eval_fxc(hfb200_ctx* ctx, int x_func, int c_func, int nspin, const double* Pa, const double* Pb, int ldp,
                           double dens_thr, double* Ha, double* Hb, int ldh, int beta, double* Exc, double* Nel, double* Ekin);
int hfb200_diatomic_eval_fxc(hfb200_ctx* ctx, int x_func, int c_func, int nspin, const double* Pa, const double* Pb, int ldp,
                             double dens_thr, double* Ha, double* Hb, int ldh, int beta, double* Exc, double* Nel, double* Ekin);
/* number of kernels the last product-grid eval_fxc launched (bench accounting) */
int hfb200_grid_launches(hfb200_ctx* ctx);
/* accounting of the last product-grid eval_fxc (no reference counterpart): out3 = { device milliseconds between the end of the
 * upload and the end of the last kernel (CUDA events on the context's stream), bytes uploaded, bytes downloaded } -- the packed
 * angular blocks, see DESIGN.md section 7 */
int hfb200_grid_last_call(hfb200_ctx* ctx, double* out3);

/* Pointwise XC on the device for validation (libxc array conventions, n points, host arrays;
 * outputs are overwritten with exchange + correlation). */
int hfb200_xc_eval(int device, int x_func, int c_func, int nspin, int n, double dens_thr, const double* rho,
                   const double* sigma, const double* tau, double* exc, double* vrho, double* vsigma,
                   double* vtau);

#ifdef __cplusplus
}
#endif
#endif
