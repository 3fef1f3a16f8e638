// C ABI: error handling, functional lookup, context life cycle (include/helfem_b200.h).
#include <cstring>
#include "../../include/helfem_b200.h"
#include "chan_engine.h"
#include "context.h"
#include "grid2d.h"
#include "xc_functionals.cuh"

namespace hfb {
static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* get_error() { return g_err; }
}  // namespace hfb

extern "C" {

const char* hfb200_last_error(void) { return hfb::get_error(); }
const char* hfb200_version(void) { return "helfem_b200 0.1 (sm_100a)"; }

// Names as the reference's --method option spells them (general/dftfuncs.cpp:94-117).
int hfb200_xc_id(const char* name) {
  if (!name) return -1;
  struct { const char* n; int id; } tab[] = {
      {"none", HFB200_XC_NONE},          {"lda_x", HFB200_XC_LDA_X},
      {"lda_c_pw", HFB200_XC_LDA_C_PW},  {"lda_c_pw_mod", HFB200_XC_LDA_C_PW_MOD},
      {"gga_x_pbe", HFB200_XC_GGA_X_PBE}, {"gga_c_pbe", HFB200_XC_GGA_C_PBE},
      {"mgga_x_r2scan", HFB200_XC_MGGA_X_R2SCAN}, {"mgga_c_r2scan", HFB200_XC_MGGA_C_R2SCAN}};
  for (auto& t : tab)
    if (strcasecmp(name, t.n) == 0) return t.id;
  hfb::set_error("functional '%s' has no device implementation", name);
  return -1;
}

int hfb200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

/* Multi-GPU sharding of the atomic / diatomic operators: this context only holds and contracts the
 * integral families r with r % nranks == rank (host/famshard.h; atomic: r = L; diatomic: r = index of
 * (L,|M|)); the results of the ranks add up to the full J / K (all-reduced inside the library
 * once a communicator is attached, comm.cu).  Must be called before *_set_tei. */
int hfb200_set_shard(hfb200_ctx* c, int rank, int nranks) {
  if (!c || nranks < 1 || rank < 0 || rank >= nranks) { hfb::set_error("hfb200_set_shard: invalid argument"); return HFB200_EINVAL; }
  if (c->chan) { hfb::set_error("hfb200_set_shard must precede set_tei"); return HFB200_EINVAL; }
  c->shard_rank = rank;
  c->shard_n = nranks;
  return HFB200_OK;
}

int hfb200_shard_owner(hfb200_ctx* c, int* owner, int n) {
  if (!c || !owner || n < 0 || (size_t)n > c->fam_owner.size()) { hfb::set_error("hfb200_shard_owner: invalid argument (call after *_set_tei; n <= number of families)"); return HFB200_EINVAL; }
  for (int r = 0; r < n; r++) owner[r] = c->fam_owner[r];
  return HFB200_OK;
}

void hfb200_destroy(hfb200_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->comm) hfb200_comm_destroy(c);
  double* ptrs[] = {c->dBt, c->ddBt, c->dR, c->dW, c->dDisL, c->dDisM1L, c->dTeiF, c->dLc, c->dm1Lc, c->dKtei, c->dCpl};
  if (c->dTri) cudaFree(c->dTri);
  int* iptrs[] = {c->dFirst, c->dNpe, c->dElemOf};
  for (int* p : iptrs)
    if (p) cudaFree(p);
  for (double* p : ptrs)
    if (p) cudaFree(p);
  hfb::DevBuf* bufs[] = {&c->dP, &c->dH, &c->dJ, &c->dPtot, &c->dTraces, &c->dPartial, &c->dScalars, &c->dK};
  for (auto* b : bufs) b->release();
  c->hIn.release();
  c->hOut.release();
  if (c->chan) { c->chan->release(); delete c->chan; }
  if (c->g2d) { c->g2d->release(); delete c->g2d; }
  if (c->stream) cudaStreamDestroy(c->stream);
  for (cudaStream_t q : c->xs)
    if (q) cudaStreamDestroy(q);
  delete c;
}

}  // extern "C"
