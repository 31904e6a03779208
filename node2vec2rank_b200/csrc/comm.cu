// NCCL from C: the all-reduce of Y (n x l fp32) per operator application when the layers are sharded over
// processes (north_star: "only Y and the l x l Gram matrices are allreduced with NCCL over NVLink").
// The library does not link NCCL: it binds the copy torch has already loaded (libnccl.so.2) with dlopen/dlsym, creates
// its own communicator from a unique id the host side broadcasts, and the solver calls ncclAllReduce on its stream --
// no Python callback inside the solver loop.
#include "common.cuh"
#include <dlfcn.h>
#include <string.h>

namespace n2v2r {

struct NcclId { char internal[128]; };
typedef int (*nccl_get_id_fn)(NcclId*);
typedef int (*nccl_init_rank_fn)(void**, int, NcclId, int);
typedef int (*nccl_destroy_fn)(void*);
typedef int (*nccl_allreduce_fn)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef const char* (*nccl_errstr_fn)(int);

static struct {
  void* handle = nullptr;
  nccl_get_id_fn get_id = nullptr;
  nccl_init_rank_fn init_rank = nullptr;
  nccl_destroy_fn destroy = nullptr;
  nccl_allreduce_fn allreduce = nullptr;
  nccl_errstr_fn errstr = nullptr;
} g_nccl;

static int bind_nccl() {
  if (g_nccl.allreduce) return N2V2R_OK;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);        // the copy the process already uses (torch's)
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) { set_error("NCCL not found: %s", dlerror()); return N2V2R_ECOMM; }
  g_nccl.handle = h;
  g_nccl.get_id = (nccl_get_id_fn)dlsym(h, "ncclGetUniqueId");
  g_nccl.init_rank = (nccl_init_rank_fn)dlsym(h, "ncclCommInitRank");
  g_nccl.destroy = (nccl_destroy_fn)dlsym(h, "ncclCommDestroy");
  g_nccl.errstr = (nccl_errstr_fn)dlsym(h, "ncclGetErrorString");
  g_nccl.allreduce = (nccl_allreduce_fn)dlsym(h, "ncclAllReduce");
  if (!g_nccl.get_id || !g_nccl.init_rank || !g_nccl.destroy || !g_nccl.allreduce) {
    g_nccl.allreduce = nullptr;
    set_error("NCCL symbols missing in libnccl.so.2");
    return N2V2R_ECOMM;
  }
  return N2V2R_OK;
}

static int nccl_fail(int r, const char* what) {
  set_error("%s failed: %s", what, g_nccl.errstr ? g_nccl.errstr(r) : "NCCL error");
  return N2V2R_ECOMM;
}

// sum `count` floats over the communicator, in place, ordered on `s`
int comm_allreduce_f32(void* comm, float* d_buf, int64_t count, cudaStream_t s) {
  int rc = bind_nccl();
  if (rc) return rc;
  const int r = g_nccl.allreduce(d_buf, d_buf, (size_t)count, /*ncclFloat32*/ 7, /*ncclSum*/ 0, comm, s);
  return r == 0 ? N2V2R_OK : nccl_fail(r, "ncclAllReduce");
}

}  // namespace n2v2r

using namespace n2v2r;

extern "C" int n2v2r_comm_unique_id(void* h_id128) {
  N2V2R_REQUIRE(h_id128, "comm_unique_id: null pointer");
  int rc = bind_nccl();
  if (rc) return rc;
  NcclId id;
  const int r = g_nccl.get_id(&id);
  if (r) return nccl_fail(r, "ncclGetUniqueId");
  memcpy(h_id128, &id, sizeof(id));
  return N2V2R_OK;
}

extern "C" int n2v2r_comm_init(const void* h_id128, int n_ranks, int rank, void** comm) {
  N2V2R_REQUIRE(h_id128 && comm && n_ranks >= 1 && rank >= 0 && rank < n_ranks, "comm_init: bad arguments");
  int rc = bind_nccl();
  if (rc) return rc;
  NcclId id;
  memcpy(&id, h_id128, sizeof(id));
  const int r = g_nccl.init_rank(comm, n_ranks, id, rank);
  return r == 0 ? N2V2R_OK : nccl_fail(r, "ncclCommInitRank");
}

extern "C" int n2v2r_comm_destroy(void* comm) {
  if (!comm) return N2V2R_OK;
  int rc = bind_nccl();
  if (rc) return rc;
  const int r = g_nccl.destroy(comm);
  return r == 0 ? N2V2R_OK : nccl_fail(r, "ncclCommDestroy");
}

// test / bench hook: one all-reduce through the native path
extern "C" int n2v2r_comm_allreduce_f32(void* stream, void* comm, float* d_buf, int64_t count) {
  N2V2R_REQUIRE(comm && d_buf && count >= 0, "comm_allreduce: bad arguments");
  return comm_allreduce_f32(comm, d_buf, count, (cudaStream_t)stream);
}
