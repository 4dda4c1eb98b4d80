// ba_comm.cu — NCCL plumbing for the point-sharded multi-GPU solve: one
// process per GPU, cameras replicated, FP64 sum all-reduces of camera-keyed quantities over
// NVLink 5 / NVSwitch.  NCCL is resolved at run time (dlopen of libnccl.so.2, which in a torch
// process is the copy torch already loaded) so that the single-GPU library has no link-time
// dependency; a multi-rank request without NCCL fails loudly with HSFM_BA_ERR_COMM.
#include <dlfcn.h>

#include <cstdio>
#include <cstring>

#include "ba_internal.h"

namespace hsfm {
namespace {

typedef struct { char internal[128]; } ncclUniqueId_t;
typedef void* ncclComm_t_;
typedef int ncclResult_t_;
constexpr int kNcclInt32 = 2, kNcclFloat64 = 8;  // ncclDataType_t
constexpr int kNcclSum = 0, kNcclMax = 2, kNcclMin = 3;  // ncclRedOp_t

struct Api {
  void* lib = nullptr;
  ncclResult_t_ (*GetUniqueId)(ncclUniqueId_t*) = nullptr;
  ncclResult_t_ (*CommInitRank)(ncclComm_t_*, int, ncclUniqueId_t, int) = nullptr;
  ncclResult_t_ (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t_, cudaStream_t) = nullptr;
  ncclResult_t_ (*CommDestroy)(ncclComm_t_) = nullptr;
  bool ok = false;
};

Api& api() {
  static Api a;
  static bool tried = false;
  if (tried) return a;
  tried = true;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    a.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (a.lib) break;
  }
  if (!a.lib) return a;
  a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(a.lib, "ncclGetUniqueId");
  a.CommInitRank = (decltype(a.CommInitRank))dlsym(a.lib, "ncclCommInitRank");
  a.AllReduce = (decltype(a.AllReduce))dlsym(a.lib, "ncclAllReduce");
  a.CommDestroy = (decltype(a.CommDestroy))dlsym(a.lib, "ncclCommDestroy");
  a.ok = a.GetUniqueId && a.CommInitRank && a.AllReduce && a.CommDestroy;
  return a;
}

}  // namespace

int comm_unique_id(void* out128) {
  Api& a = api();
  if (!a.ok) return HSFM_BA_ERR_COMM;
  ncclUniqueId_t id;
  const int rc = a.GetUniqueId(&id);
  if (rc) return 2000 + rc;
  std::memcpy(out128, &id, 128);
  return 0;
}

int comm_init(Comm& c, const void* id128, int rank, int n_ranks) {
  if (n_ranks <= 1) { c.rank = 0; c.n_ranks = 1; c.handle = nullptr; return 0; }
  Api& a = api();
  if (!a.ok) return HSFM_BA_ERR_COMM;
  ncclUniqueId_t id;
  std::memcpy(&id, id128, 128);
  ncclComm_t_ h = nullptr;
  const int rc = a.CommInitRank(&h, n_ranks, id, rank);
  if (rc) return 2000 + rc;
  c.handle = h; c.rank = rank; c.n_ranks = n_ranks;
  return 0;
}

int Comm::allreduce_sum(double* buf, size_t count, cudaStream_t s) {
  if (n_ranks <= 1 || count == 0) return 0;
  const int rc = api().AllReduce(buf, buf, count, kNcclFloat64, kNcclSum, handle, s);
  return rc ? 2000 + rc : 0;
}
int Comm::allreduce_max(double* buf, size_t count, cudaStream_t s) {
  if (n_ranks <= 1 || count == 0) return 0;
  const int rc = api().AllReduce(buf, buf, count, kNcclFloat64, kNcclMax, handle, s);
  return rc ? 2000 + rc : 0;
}
int Comm::allreduce_sum_i32(int* buf, size_t count, cudaStream_t s) {
  if (n_ranks <= 1 || count == 0) return 0;
  const int rc = api().AllReduce(buf, buf, count, kNcclInt32, kNcclSum, handle, s);
  return rc ? 2000 + rc : 0;
}
int Comm::allreduce_min_i32(int* buf, size_t count, cudaStream_t s) {
  if (n_ranks <= 1 || count == 0) return 0;
  const int rc = api().AllReduce(buf, buf, count, kNcclInt32, kNcclMin, handle, s);
  return rc ? 2000 + rc : 0;
}
void Comm::destroy() {
  if (handle && api().ok) api().CommDestroy(handle);
  handle = nullptr;
  n_ranks = 1;
  rank = 0;
}

}  // namespace hsfm
