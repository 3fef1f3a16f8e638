// NCCL communicator of a sharded atomic / diatomic context (include/helfem_b200.h: hfb200_comm_*).
//
// The reference drivers are single-process programs (diatomic/main.cpp:810-828); the sharded operators run one process
// per GPU, every rank holding the integral families r % nranks == rank (hfb200_set_shard).  With a communicator attached
// the operators all-reduce inside the library, on the caller's stream, over NVLink:
//   coulomb:  the per-channel radial results Jaux_c (a few MB: nchan x nel x stacked triangle), BEFORE the bra
//             contraction, which every rank then carries out for the complete channel set
//   exchange: the (nrad x nrad) angular blocks of K that can be non-zero for the sparsity pattern of P, packed
// NCCL is resolved at run time (dlopen of libnccl.so.2: the copy the process already has, e.g. PyTorch's, or the system
// one), so the library has no link-time dependency on it and single-GPU users never load it.
#include <dlfcn.h>
#include <nccl.h>  // declarations only

#include <mutex>

#include "../../include/helfem_b200.h"
#include "chan_engine.h"
#include "context.h"

namespace hfb {

namespace {
struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

int load_nccl() {
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  if (g_nccl.handle) return 0;
  void* h = nullptr;
  for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
    h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  if (!h) { set_error("NCCL not found (dlopen libnccl.so.2): %s", dlerror()); return HFB200_ECUDA; }
  NcclApi a;
  a.handle = h;
  a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(h, "ncclGetUniqueId");
  a.CommInitRank = (decltype(a.CommInitRank))dlsym(h, "ncclCommInitRank");
  a.CommDestroy = (decltype(a.CommDestroy))dlsym(h, "ncclCommDestroy");
  a.AllReduce = (decltype(a.AllReduce))dlsym(h, "ncclAllReduce");
  a.Broadcast = (decltype(a.Broadcast))dlsym(h, "ncclBroadcast");
  a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
  if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllReduce || !a.Broadcast || !a.GetErrorString) {
    set_error("libnccl lacks a required entry point");
    return HFB200_ECUDA;
  }
  g_nccl = a;
  return 0;
}

#define HFB_NCCL_OK(call)                                                                                    \
  do {                                                                                                       \
    ncclResult_t r_ = (call);                                                                                \
    if (r_ != ncclSuccess) { set_error("NCCL error at %s:%d: %s", __FILE__, __LINE__, g_nccl.GetErrorString(r_)); return HFB200_ECUDA; } \
  } while (0)
}  // namespace

int comm_allreduce_sum(void* comm, double* buf, size_t n, cudaStream_t st) {
  if (n == 0) return 0;
  HFB_NCCL_OK(g_nccl.AllReduce(buf, buf, n, ncclDouble, ncclSum, (ncclComm_t)comm, st));
  return 0;
}
int comm_broadcast(void* comm, double* buf, size_t n, int root, cudaStream_t st) {
  if (n == 0) return 0;
  HFB_NCCL_OK(g_nccl.Broadcast(buf, buf, n, ncclDouble, root, (ncclComm_t)comm, st));
  return 0;
}

}  // namespace hfb

using namespace hfb;

extern "C" {

int hfb200_comm_unique_id(void* id, int nbytes) {
  if (!id || nbytes < (int)sizeof(ncclUniqueId)) { set_error("hfb200_comm_unique_id: need a buffer of %d bytes", (int)sizeof(ncclUniqueId)); return HFB200_EINVAL; }
  int rc;
  if ((rc = load_nccl())) return rc;
  ncclUniqueId u;
  HFB_NCCL_OK(g_nccl.GetUniqueId(&u));
  memcpy(id, &u, sizeof(u));
  return HFB200_OK;
}

int hfb200_comm_init(hfb200_ctx* c, const void* id, int rank, int nranks) {
  if (!c || !id || nranks < 1 || rank < 0 || rank >= nranks) { set_error("hfb200_comm_init: invalid argument"); return HFB200_EINVAL; }
  if (c->kind != CTX_ATOMIC && c->kind != CTX_DIATOMIC) { set_error("hfb200_comm_init: only the atomic / diatomic operators shard"); return HFB200_EINVAL; }
  if (c->shard_n != nranks || c->shard_rank != rank) {
    set_error("hfb200_comm_init: the context was sharded as rank %d of %d (hfb200_set_shard), not %d of %d", c->shard_rank, c->shard_n, rank, nranks);
    return HFB200_EINVAL;
  }
  int rc;
  if ((rc = load_nccl())) return rc;
  HFB_CUDA_OK(cudaSetDevice(c->device));
  if (c->comm) { g_nccl.CommDestroy((ncclComm_t)c->comm); c->comm = nullptr; }
  ncclUniqueId u;
  memcpy(&u, id, sizeof(u));
  ncclComm_t comm = nullptr;
  HFB_NCCL_OK(g_nccl.CommInitRank(&comm, nranks, u, rank));
  c->comm = comm;
  if (c->chan) { c->chan->comm = comm; c->chan->comm_rank = rank; c->chan->comm_size = nranks; }
  return HFB200_OK;
}

int hfb200_comm_destroy(hfb200_ctx* c) {
  if (!c) { set_error("hfb200_comm_destroy: null context"); return HFB200_EINVAL; }
  if (c->comm && g_nccl.CommDestroy) {
    cudaSetDevice(c->device);
    g_nccl.CommDestroy((ncclComm_t)c->comm);
  }
  c->comm = nullptr;
  if (c->chan) { c->chan->comm = nullptr; c->chan->comm_size = 1; c->chan->comm_root = -1; }
  return HFB200_OK;
}

}  // extern "C"
