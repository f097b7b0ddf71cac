// Data-parallel plumbing of the C ABI (include/nlbac.h): one NCCL communicator over the GPUs of the
// box and ONE in-place fp32 sum all-reduce per flat gradient buffer, enqueued on the caller's stream (no host sync).
// The reference's only analogue is the unused mpi4py scaffold (utils/mpi_pytorch.py:20-34: per-parameter host-buffer
// Allreduce); here the payloads (NODE 52 209 floats, actor / critic / Lyapunov / barrier 70-140 k floats each, the [m]
// constraint sums) are latency-bound messages over NVLink 5 / NVSwitch.
// NCCL is resolved at run time with dlopen("libnccl.so.2") -- the copy torch has already loaded when called from a
// torch process, the system one otherwise -- so the library has no link-time NCCL dependency.
#include <dlfcn.h>
#include <mutex>

#include "../../include/nlbac.h"
#include "common.cuh"

namespace nlbac {

struct NcclId { char internal[128]; };                 // ncclUniqueId
typedef void* NcclComm;                                // ncclComm_t
struct NcclApi {
  int (*GetUniqueId)(NcclId*);
  int (*CommInitRank)(NcclComm*, int, NcclId, int);
  int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
  int (*CommDestroy)(NcclComm);
  const char* (*GetErrorString)(int);
  bool ok = false;
};
static NcclApi g_nccl;
static std::mutex g_nccl_mu;

static int load_nccl() {
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  if (g_nccl.ok) return 0;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
  NLBAC_REQUIRE(h != nullptr, "nlbac_comm: cannot load libnccl.so.2 (%s)", dlerror());
  g_nccl.GetUniqueId = reinterpret_cast<decltype(g_nccl.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
  g_nccl.CommInitRank = reinterpret_cast<decltype(g_nccl.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
  g_nccl.AllReduce = reinterpret_cast<decltype(g_nccl.AllReduce)>(dlsym(h, "ncclAllReduce"));
  g_nccl.CommDestroy = reinterpret_cast<decltype(g_nccl.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
  g_nccl.GetErrorString = reinterpret_cast<decltype(g_nccl.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
  NLBAC_REQUIRE(g_nccl.GetUniqueId && g_nccl.CommInitRank && g_nccl.AllReduce && g_nccl.CommDestroy &&
                    g_nccl.GetErrorString, "nlbac_comm: libnccl.so.2 lacks an expected symbol");
  g_nccl.ok = true;
  return 0;
}

static int nccl_fail(int rc, const char* what) {
  set_error("NCCL error %d (%s) at %s", rc, g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?", what);
  return 1000 + rc;
}

}  // namespace nlbac

using namespace nlbac;

struct nlbac_comm {
  NcclComm comm;
  int world, rank;
};

extern "C" {

int nlbac_comm_unique_id(void* id128) {
  NLBAC_REQUIRE(id128 != nullptr, "null id buffer");
  if (int rc = load_nccl()) return rc;
  if (int rc = g_nccl.GetUniqueId(static_cast<NcclId*>(id128))) return nccl_fail(rc, "ncclGetUniqueId");
  return 0;
}

int nlbac_comm_init(int world, int rank, const void* id128, nlbac_comm** out) {
  NLBAC_REQUIRE(world >= 1 && rank >= 0 && rank < world && id128 != nullptr && out != nullptr,
                "bad communicator arguments (world=%d rank=%d)", world, rank);
  if (int rc = load_nccl()) return rc;
  NcclId id = *static_cast<const NcclId*>(id128);
  NcclComm c = nullptr;
  if (int rc = g_nccl.CommInitRank(&c, world, id, rank)) return nccl_fail(rc, "ncclCommInitRank");
  *out = new nlbac_comm{c, world, rank};
  return 0;
}

int nlbac_allreduce_grads(nlbac_comm* comm, float* flat, size_t n, void* stream) {
  NLBAC_REQUIRE(comm != nullptr && (flat != nullptr || n == 0), "null communicator or buffer");
  if (n == 0) return 0;
  // ncclFloat32 = 7, ncclSum = 0 (nccl.h); in place
  if (int rc = g_nccl.AllReduce(flat, flat, n, 7, 0, comm->comm, (cudaStream_t)stream))
    return nccl_fail(rc, "ncclAllReduce");
  return 0;
}

int nlbac_allreduce_sum_f64(nlbac_comm* comm, double* buf, size_t n, void* stream) {
  NLBAC_REQUIRE(comm != nullptr && (buf != nullptr || n == 0), "null communicator or buffer");
  if (n == 0) return 0;
  // ncclFloat64 = 8, ncclSum = 0 (nccl.h); in place
  if (int rc = g_nccl.AllReduce(buf, buf, n, 8, 0, comm->comm, (cudaStream_t)stream))
    return nccl_fail(rc, "ncclAllReduce(f64)");
  return 0;
}

int nlbac_comm_destroy(nlbac_comm* comm) {
  if (comm == nullptr) return 0;
  int rc = g_nccl.ok ? g_nccl.CommDestroy(comm->comm) : 0;
  delete comm;
  return rc ? nccl_fail(rc, "ncclCommDestroy") : 0;
}

}  // extern "C"
