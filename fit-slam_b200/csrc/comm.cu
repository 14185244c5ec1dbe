// comm.cu -- the multi-GPU plumbing of SURVEY.md 8e behind the C ABI: one process per GPU, one NCCL communicator per
// context.  The costmap (whole or dirty rows) and the map points are uploaded by rank 0 only and broadcast over NVLink
// (fitslam_set_costmap / fitslam_set_map_points call comm_broadcast), every rank scores the frontiers f with
// f % nranks == rank, and the 16-byte score records are merged by an all-reduce on the context stream
// (fitslam_cycle_merge; fitslam_cycle does it between its two halves).  The reference has no analogue (one CPU thread).
//
// NCCL is loaded at run time (dlopen of libnccl.so.2 in fitslam_comm_init): a single-GPU host needs no NCCL at all, and a
// process that already carries one (PyTorch's) shares it.  FITSLAM_NCCL_LIB names a particular file.
#include "ctx.h"

#include <dlfcn.h>
#include <nccl.h>

namespace fitslam {

struct CommState {
  void *lib = nullptr;
  ncclComm_t comm = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

static bool load_nccl(CommState *s, std::string *why) {
  if (s->lib) return true;
  // an NCCL the process already carries (a host linked against it, PyTorch's bundled one) is shared; else FITSLAM_NCCL_LIB
  // names the file, else the loader's search path decides.  RTLD_LOCAL: the library's symbols stay out of the global scope
  const char *named = getenv("FITSLAM_NCCL_LIB");
  s->lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL | RTLD_NOLOAD);
  if (!s->lib && named && *named) s->lib = dlopen(named, RTLD_NOW | RTLD_LOCAL);
  for (const char *name : {"libnccl.so.2", "libnccl.so"})
    if (!s->lib) s->lib = dlopen(name, RTLD_NOW | RTLD_LOCAL);
  if (!s->lib) { *why = std::string("cannot load libnccl: ") + dlerror(); return false; }
  s->GetUniqueId = reinterpret_cast<decltype(s->GetUniqueId)>(dlsym(s->lib, "ncclGetUniqueId"));
  s->CommInitRank = reinterpret_cast<decltype(s->CommInitRank)>(dlsym(s->lib, "ncclCommInitRank"));
  s->CommDestroy = reinterpret_cast<decltype(s->CommDestroy)>(dlsym(s->lib, "ncclCommDestroy"));
  s->Broadcast = reinterpret_cast<decltype(s->Broadcast)>(dlsym(s->lib, "ncclBroadcast"));
  s->AllReduce = reinterpret_cast<decltype(s->AllReduce)>(dlsym(s->lib, "ncclAllReduce"));
  s->GetErrorString = reinterpret_cast<decltype(s->GetErrorString)>(dlsym(s->lib, "ncclGetErrorString"));
  if (!s->GetUniqueId || !s->CommInitRank || !s->CommDestroy || !s->Broadcast || !s->AllReduce || !s->GetErrorString) {
    *why = "libnccl lacks an entry point";
    return false;
  }
  return true;
}

static CommState *g_loader = nullptr;  // fitslam_comm_unique_id has no context yet

#define FS_NCCL(ctx, s, expr)                                                                                     \
  do {                                                                                                            \
    ncclResult_t _r = (expr);                                                                                     \
    if (_r != ncclSuccess) return fail((ctx), FITSLAM_E_CUDA, std::string(#expr) + ": " + (s)->GetErrorString(_r)); \
  } while (0)

bool comm_active(const fitslam_ctx *c) { return c->comm && c->comm->comm; }

// rank 0's `bytes` at `dev` reach every rank's `dev`, on the context stream
fitslam_status comm_broadcast(fitslam_ctx *c, void *dev, size_t bytes) {
  if (!comm_active(c) || bytes == 0) return FITSLAM_OK;
  FS_NCCL(c, c->comm, c->comm->Broadcast(dev, dev, bytes, ncclUint8, 0, c->comm->comm, c->stream));
  return FITSLAM_OK;
}

void destroy_comm_state(fitslam_ctx *c) {
  if (!c->comm) return;
  if (c->comm->comm) c->comm->CommDestroy(c->comm->comm);
  delete c->comm;
  c->comm = nullptr;
}

}  // namespace fitslam

using namespace fitslam;

extern "C" {

fitslam_status fitslam_comm_unique_id(void *id, uint64_t bytes) {
  if (!id || bytes < sizeof(ncclUniqueId)) return FITSLAM_E_INVALID;
  if (!g_loader) g_loader = new CommState();
  std::string why;
  if (!load_nccl(g_loader, &why)) return FITSLAM_E_CUDA;
  ncclUniqueId u;
  if (g_loader->GetUniqueId(&u) != ncclSuccess) return FITSLAM_E_CUDA;
  memcpy(id, &u, sizeof(u));
  return FITSLAM_OK;
}

fitslam_status fitslam_comm_init(fitslam_ctx *c, const void *nccl_unique_id, int rank, int nranks) {
  if (!c || !nccl_unique_id || nranks < 1 || rank < 0 || rank >= nranks) return fail(c, FITSLAM_E_INVALID, "comm_init: bad argument");
  cudaSetDevice(c->device);
  destroy_comm_state(c);
  c->comm = new CommState();
  std::string why;
  if (!load_nccl(c->comm, &why)) {
    destroy_comm_state(c);
    return fail(c, FITSLAM_E_CUDA, "comm_init: " + why);
  }
  ncclUniqueId u;
  memcpy(&u, nccl_unique_id, sizeof(u));
  FS_NCCL(c, c->comm, c->comm->CommInitRank(&c->comm->comm, nranks, u, rank));
  c->rank = rank;
  c->nranks = nranks;
  return FITSLAM_OK;
}

fitslam_status fitslam_comm_destroy(fitslam_ctx *c) {
  if (!c) return FITSLAM_E_INVALID;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  destroy_comm_state(c);
  c->rank = 0;
  c->nranks = 1;
  return FITSLAM_OK;
}

// between fitslam_cycle_begin and fitslam_cycle_end: every rank's records (zero where another rank scored) summed as int32
fitslam_status fitslam_cycle_merge(fitslam_ctx *c) {
  if (!c) return FITSLAM_E_INVALID;
  if (!comm_active(c) || c->nranks == 1) return FITSLAM_OK;
  cudaSetDevice(c->device);
  const size_t words = (size_t)c->records_rows * (sizeof(ScoreRecord) / 4);
  if (words == 0) return FITSLAM_OK;
  FS_NCCL(c, c->comm, c->comm->AllReduce(c->records.p, c->records.p, words, ncclInt32, ncclSum, c->comm->comm, c->stream));
  return FITSLAM_OK;
}

}  // extern "C"
