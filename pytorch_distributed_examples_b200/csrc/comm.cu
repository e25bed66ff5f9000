// Collectives over NCCL / NVLink 5: the one exchange step of the data-parallel path.
//   replaces ref: pytorch_elastic/mnist_ddp_elastic.py:26 (init_process_group("gloo")) + the
//   per-bucket allreduce DDP issues (:58), and Horovod's allreduce / broadcast
//   (ref: horovod/mnist_horovod.py:28,53,56 ; horovod_mnist_elastic.py:35,42,55).
// libnccl.so.2 is resolved at run time with dlopen (the copy torch already mapped is reused),
// so this library loads — and exports every symbol — on a box without NCCL or a GPU.
#include <dlfcn.h>
#include <stdlib.h>
#include <string.h>
#include <nccl.h>
#include "common.cuh"

namespace ddp {

struct NcclApi {
  void* handle = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclCommAbort) CommAbort = nullptr;
  decltype(&ncclCommGetAsyncError) CommGetAsyncError = nullptr;
  decltype(&ncclAllReduce) AllReduce = nullptr;
  decltype(&ncclBroadcast) Broadcast = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  bool tried = false;
};
static NcclApi g_nccl;

static int load_nccl() {
  if (g_nccl.handle) return 0;
  if (g_nccl.tried) { set_error("libnccl.so.2 could not be loaded earlier"); return DDP_ENOSYS; }
  g_nccl.tried = true;
  const char* override_path = getenv("DDP_NCCL_PATH");
  const char* names[] = {override_path, "libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* nm : names) {
    if (!nm) continue;
    h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  if (!h) { set_error("dlopen(libnccl.so.2) failed: %s", dlerror()); return DDP_ENOSYS; }
#define LOAD(sym)                                                                  \
  g_nccl.sym = reinterpret_cast<decltype(g_nccl.sym)>(dlsym(h, "nccl" #sym));      \
  if (!g_nccl.sym) { set_error("dlsym(nccl" #sym ") failed"); return DDP_ENOSYS; }
  LOAD(GetUniqueId) LOAD(CommInitRank) LOAD(CommDestroy) LOAD(CommAbort) LOAD(CommGetAsyncError)
  LOAD(AllReduce) LOAD(Broadcast) LOAD(GetErrorString)
#undef LOAD
  g_nccl.handle = h;
  return 0;
}

static int nccl_rc(ncclResult_t r, const char* what) {
  if (r == ncclSuccess) return 0;
  set_error("%s: NCCL error %d: %s", what, int(r), g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
  return -int(r);
}
#define DDP_NCCL(call)                                       \
  do {                                                       \
    ncclResult_t r__ = (call);                               \
    if (r__ != ncclSuccess) return nccl_rc(r__, #call);      \
  } while (0)

static bool map_dtype(int dtype, ncclDataType_t* out) {
  switch (dtype) {
    case 0: *out = ncclFloat32; return true;
    case 1: *out = ncclBfloat16; return true;
    case 2: *out = ncclInt64; return true;
    case 3: *out = ncclUint8; return true;
  }
  return false;
}

}  // namespace ddp

using namespace ddp;

extern "C" int ddp_comm_unique_id(void* uid_out) {
  DDP_REQUIRE(uid_out, "uid_out is NULL");
  static_assert(sizeof(ncclUniqueId) == DDP_UID_BYTES, "ncclUniqueId size");
  if (int rc = load_nccl()) return rc;
  ncclUniqueId id;
  DDP_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(uid_out, &id, sizeof(id));
  return 0;
}

extern "C" int ddp_comm_init(void** comm_out, const void* uid, int world, int rank) {
  DDP_REQUIRE(comm_out && uid, "null argument");
  DDP_REQUIRE(world >= 1 && rank >= 0 && rank < world, "bad world/rank");
  if (int rc = load_nccl()) return rc;
  ncclUniqueId id;
  memcpy(&id, uid, sizeof(id));
  ncclComm_t comm = nullptr;
  DDP_NCCL(g_nccl.CommInitRank(&comm, world, id, rank));
  *comm_out = comm;
  return 0;
}

extern "C" int ddp_comm_destroy(void* comm) {
  if (!comm) return 0;
  if (int rc = load_nccl()) return rc;
  DDP_NCCL(g_nccl.CommDestroy(static_cast<ncclComm_t>(comm)));
  return 0;
}

extern "C" int ddp_comm_abort(void* comm) {
  if (!comm) return 0;
  if (int rc = load_nccl()) return rc;
  DDP_NCCL(g_nccl.CommAbort(static_cast<ncclComm_t>(comm)));
  return 0;
}

extern "C" int ddp_comm_async_error(void* comm, int* nccl_result_out) {
  DDP_REQUIRE(comm && nccl_result_out, "null argument");
  if (int rc = load_nccl()) return rc;
  ncclResult_t async = ncclSuccess;
  DDP_NCCL(g_nccl.CommGetAsyncError(static_cast<ncclComm_t>(comm), &async));
  *nccl_result_out = int(async);
  if (async != ncclSuccess && async != ncclInProgress)
    set_error("asynchronous NCCL error %d: %s", int(async), g_nccl.GetErrorString(async));
  return 0;
}

static int allreduce_impl(void* comm, void* buf, int64_t count, int dtype, ncclRedOp_t op, void* stream) {
  DDP_REQUIRE(comm && (buf || count == 0), "null comm or buffer");
  DDP_REQUIRE(count >= 0, "negative count");
  ncclDataType_t dt;
  DDP_REQUIRE(map_dtype(dtype, &dt), "unknown dtype");
  if (count == 0) return 0;
  if (int rc = load_nccl()) return rc;
  DDP_NCCL(g_nccl.AllReduce(buf, buf, size_t(count), dt, op, static_cast<ncclComm_t>(comm),
                            static_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int ddp_allreduce_sum(void* comm, void* buf, int64_t count, int dtype, void* stream) {
  return allreduce_impl(comm, buf, count, dtype, ncclSum, stream);
}
extern "C" int ddp_allreduce_max(void* comm, void* buf, int64_t count, int dtype, void* stream) {
  return allreduce_impl(comm, buf, count, dtype, ncclMax, stream);
}

extern "C" int ddp_broadcast(void* comm, void* buf, int64_t nbytes, int root, void* stream) {
  DDP_REQUIRE(comm && (buf || nbytes == 0), "null comm or buffer");
  DDP_REQUIRE(nbytes >= 0 && root >= 0, "bad nbytes/root");
  if (nbytes == 0) return 0;
  if (int rc = load_nccl()) return rc;
  DDP_NCCL(g_nccl.Broadcast(buf, buf, size_t(nbytes), ncclUint8, root, static_cast<ncclComm_t>(comm),
                            static_cast<cudaStream_t>(stream)));
  return 0;
}
