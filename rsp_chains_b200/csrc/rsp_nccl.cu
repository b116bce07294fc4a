// rsp_nccl.cu -- the one inter-GPU exchange of the path: gathering the per-GPU detection lists.  Frames are sharded
// over the GPUs with no data-path collective; detections are tiny (16 bytes each), so the gather is two ncclAllGather calls on the
// caller's stream: the counts, then one fixed-size slot of entries per rank.  NCCL is bound at run time (dlopen of libnccl.so.2, the
// copy already loaded into the process when there is one, e.g. PyTorch's), so the library has no link-time dependency on it and a
// single-GPU user never needs it.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <cstring>
#include <string>

#include "../../include/rsp_chains.h"

namespace {

struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*CommCount)(const ncclComm_t, int *) = nullptr;
  ncclResult_t (*CommUserRank)(const ncclComm_t, int *) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  std::string error;
};

NcclApi &nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api;
  tried = true;
  void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);   // the copy the process already uses (PyTorch's), if any
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) { api.error = std::string("libnccl.so.2 cannot be loaded: ") + dlerror(); return api; }
  api.handle = h;
#define RSP_SYM(field, name) api.field = reinterpret_cast<decltype(api.field)>(dlsym(h, name)); if (!api.field) { api.error = std::string("libnccl lacks ") + name; api.handle = nullptr; return api; }
  RSP_SYM(GetUniqueId, "ncclGetUniqueId") RSP_SYM(CommInitRank, "ncclCommInitRank") RSP_SYM(CommDestroy, "ncclCommDestroy")
  RSP_SYM(CommCount, "ncclCommCount") RSP_SYM(CommUserRank, "ncclCommUserRank") RSP_SYM(AllGather, "ncclAllGather")
  RSP_SYM(GroupStart, "ncclGroupStart") RSP_SYM(GroupEnd, "ncclGroupEnd") RSP_SYM(GetErrorString, "ncclGetErrorString")
#undef RSP_SYM
  return api;
}

thread_local std::string g_nccl_error;

}  // namespace

// rsp_abi.cu keeps the thread's last error; this unit reports through the same channel
extern "C" void rsp_set_last_error_(const char *msg);
static rsp_status nfail(rsp_status st, const std::string &msg) { rsp_set_last_error_(msg.c_str()); return st; }
#define NCCL_TRY(api, expr)                                                                                          \
  do {                                                                                                               \
    ncclResult_t r__ = (expr);                                                                                       \
    if (r__ != ncclSuccess) return nfail(RSP_ECUDA, std::string(#expr) + ": " + (api).GetErrorString(r__));          \
  } while (0)

static_assert(sizeof(ncclUniqueId) == RSP_NCCL_UNIQUE_ID_BYTES, "ncclUniqueId is 128 bytes");

extern "C" rsp_status rsp_nccl_get_unique_id(uint8_t id[RSP_NCCL_UNIQUE_ID_BYTES]) {
  if (!id) return nfail(RSP_EINVAL, "null argument");
  NcclApi &api = nccl_api();
  if (!api.handle) return nfail(RSP_EUNSUPPORTED, api.error);
  ncclUniqueId u;
  NCCL_TRY(api, api.GetUniqueId(&u));
  std::memcpy(id, &u, sizeof(u));
  return RSP_OK;
}

extern "C" rsp_status rsp_nccl_comm_init_rank(void **comm, int n_ranks, const uint8_t id[RSP_NCCL_UNIQUE_ID_BYTES], int rank, int device) {
  if (!comm || !id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return nfail(RSP_EINVAL, "bad argument");
  NcclApi &api = nccl_api();
  if (!api.handle) return nfail(RSP_EUNSUPPORTED, api.error);
  int prev = -1;
  cudaGetDevice(&prev);
  if (cudaSetDevice(device) != cudaSuccess) return nfail(RSP_ECUDA, "cudaSetDevice failed");
  ncclUniqueId u;
  std::memcpy(&u, id, sizeof(u));
  ncclComm_t c = nullptr;
  const ncclResult_t r = api.CommInitRank(&c, n_ranks, u, rank);
  if (prev >= 0) cudaSetDevice(prev);
  if (r != ncclSuccess) return nfail(RSP_ECUDA, std::string("ncclCommInitRank: ") + api.GetErrorString(r));
  *comm = c;
  return RSP_OK;
}

extern "C" void rsp_nccl_comm_destroy(void *comm) {
  NcclApi &api = nccl_api();
  if (api.handle && comm) api.CommDestroy(static_cast<ncclComm_t>(comm));
}

extern "C" rsp_status rsp_gather_detections(rsp_chain *c, void *nccl_comm, const rsp_dets *local, uint32_t slot_capacity,
                                            rsp_detection *d_gathered, uint32_t *d_counts, void *stream) {
  (void)c;
  if (!nccl_comm || !local || !local->list || !local->count || !d_gathered || !d_counts) return nfail(RSP_EINVAL, "null argument");
  if (slot_capacity == 0 || slot_capacity > local->capacity) return nfail(RSP_EINVAL, "slot_capacity must be in 1..local->capacity (every rank sends one slot)");
  NcclApi &api = nccl_api();
  if (!api.handle) return nfail(RSP_EUNSUPPORTED, api.error);
  ncclComm_t comm = static_cast<ncclComm_t>(nccl_comm);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // counts first, then one slot per rank: rank r's list lands at d_gathered[r * slot_capacity ...], of which
  // min(d_counts[r], slot_capacity) entries are valid.  Both collectives are asynchronous on the caller's stream.
  NCCL_TRY(api, api.GroupStart());
  NCCL_TRY(api, api.AllGather(local->count, d_counts, 1, ncclUint32, comm, st));
  NCCL_TRY(api, api.AllGather(local->list, d_gathered, (size_t)slot_capacity * sizeof(rsp_detection), ncclUint8, comm, st));
  NCCL_TRY(api, api.GroupEnd());
  return RSP_OK;
}
