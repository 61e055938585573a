// NCCL collective hook in C++ (BASELINE north_star: "host code stays C++17"): the exchange steps of a sharded problem
// (include/viba_b200.h, vb_problem_create_sharded) call ncclAllReduce / ncclBroadcast directly on the problem's stream;
// no Python, no torch on the collective path.  NCCL is bound at run time (dlopen "libnccl.so.2": the copy already loaded
// in the process — e.g. the one torch bundles — or the system's), so single-GPU users of the library need no NCCL.
#include <dlfcn.h>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>
#include "../../include/viba_b200.h"

namespace {

// the few NCCL declarations the hook needs (nccl.h: ncclResult_t, ncclUniqueId, ncclDataType_t, ncclRedOp_t)
struct NcclUniqueId { char internal[128]; };
using NcclComm = void*;
enum { kNcclSuccess = 0, kNcclDouble = 8, kNcclSum = 0 };   // ncclFloat64 = 8, ncclSum = 0

struct NcclApi {
  int (*GetUniqueId)(NcclUniqueId*) = nullptr;
  int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
  int (*CommDestroy)(NcclComm) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, void*) = nullptr;
  int (*Broadcast)(const void*, void*, size_t, int, int, NcclComm, void*) = nullptr;
  int (*GetVersion)(int*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  std::string error;
  bool ok = false;
};

NcclApi& api() {
  static NcclApi a;
  static std::once_flag once;
  std::call_once(once, [] {
    void* h = nullptr;
    for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
      h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
      if (h) break;
    }
    if (!h) { a.error = std::string("NCCL not found: ") + dlerror(); return; }
    auto sym = [&](const char* n) { void* p = dlsym(h, n); if (!p && a.error.empty()) a.error = std::string("NCCL symbol missing: ") + n; return p; };
    a.GetUniqueId = (decltype(a.GetUniqueId))sym("ncclGetUniqueId");
    a.CommInitRank = (decltype(a.CommInitRank))sym("ncclCommInitRank");
    a.CommDestroy = (decltype(a.CommDestroy))sym("ncclCommDestroy");
    a.AllReduce = (decltype(a.AllReduce))sym("ncclAllReduce");
    a.Broadcast = (decltype(a.Broadcast))sym("ncclBroadcast");
    a.GetVersion = (decltype(a.GetVersion))sym("ncclGetVersion");
    a.GetErrorString = (decltype(a.GetErrorString))sym("ncclGetErrorString");
    a.ok = a.error.empty();
  });
  return a;
}

thread_local std::string g_ncclErr;

struct HookState { NcclComm comm; int64_t calls = 0; double bytes = 0; };

int ncclHook(void* user, int op, double* ptr, int64_t count, void* stream) {
  HookState* st = static_cast<HookState*>(user);
  NcclApi& a = api();
  int rc;
  if (op == VB_COLL_SUM) rc = a.AllReduce(ptr, ptr, (size_t)count, kNcclDouble, kNcclSum, st->comm, stream);
  else if (op == VB_COLL_BCAST0) rc = a.Broadcast(ptr, ptr, (size_t)count, kNcclDouble, 0, st->comm, stream);
  else return 1;
  st->calls++; st->bytes += 8.0 * (double)count;
  return rc == kNcclSuccess ? 0 : 1;
}

}  // namespace

extern "C" {

const char* vb_nccl_last_error(void) { return g_ncclErr.c_str(); }

int vb_nccl_version(int* version) {
  NcclApi& a = api();
  if (!a.ok) { g_ncclErr = a.error; return VB_ERR_STATE; }
  return a.GetVersion(version) == kNcclSuccess ? VB_OK : VB_ERR_CUDA;
}

int vb_nccl_unique_id(char* id128) {
  NcclApi& a = api();
  if (!a.ok) { g_ncclErr = a.error; return VB_ERR_STATE; }
  if (!id128) return VB_ERR_ARG;
  NcclUniqueId id;
  const int rc = a.GetUniqueId(&id);
  if (rc != kNcclSuccess) { g_ncclErr = a.GetErrorString(rc); return VB_ERR_CUDA; }
  std::memcpy(id128, id.internal, 128);
  return VB_OK;
}

// the caller has made `device` current (vb_ctx_create does); every rank calls this with the id rank 0 produced
int vb_nccl_comm_create(int rank, int nranks, const char* id128, void** comm_out) {
  NcclApi& a = api();
  if (!a.ok) { g_ncclErr = a.error; return VB_ERR_STATE; }
  if (!id128 || !comm_out || rank < 0 || rank >= nranks) return VB_ERR_ARG;
  NcclUniqueId id;
  std::memcpy(id.internal, id128, 128);
  NcclComm comm = nullptr;
  const int rc = a.CommInitRank(&comm, nranks, id, rank);
  if (rc != kNcclSuccess) { g_ncclErr = a.GetErrorString(rc); return VB_ERR_CUDA; }
  HookState* st = new HookState{comm};
  *comm_out = st;
  return VB_OK;
}

int vb_nccl_comm_destroy(void* comm) {
  if (!comm) return VB_OK;
  HookState* st = static_cast<HookState*>(comm);
  NcclApi& a = api();
  if (a.ok && st->comm) a.CommDestroy(st->comm);
  delete st;
  return VB_OK;
}

int vb_nccl_comm_stats(void* comm, int64_t* calls, double* bytes) {
  if (!comm) return VB_ERR_ARG;
  HookState* st = static_cast<HookState*>(comm);
  if (calls) *calls = st->calls;
  if (bytes) *bytes = st->bytes;
  return VB_OK;
}

int vb_problem_create_sharded_nccl(vb_ctx* ctx, const vb_graph_desc* g, int rank, int nranks, void* comm, vb_problem** out) {
  if (!comm) return VB_ERR_ARG;
  return vb_problem_create_sharded(ctx, g, rank, nranks, ncclHook, comm, out);
}

}  // extern "C"
