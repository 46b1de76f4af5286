// comm.cu — the one collective on the path: a NCCL sum-allreduce of the flattened gradient bucket
// for the batch-sharded MLP training step.  The reference has no distributed code at
// all; this is new work.  One process per GPU; NCCL is resolved at run time with dlopen so the
// library loads (and every single-GPU entry point works) on a box without NCCL.
#include <dlfcn.h>
#include <stdlib.h>
#include "common.cuh"

namespace wf2 {

struct NcclId { char internal[128]; };
typedef void* NcclComm;
typedef int (*fn_get_unique_id)(NcclId*);
typedef int (*fn_comm_init_rank)(NcclComm*, int, NcclId, int);
typedef int (*fn_all_reduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
typedef int (*fn_comm_destroy)(NcclComm);
typedef const char* (*fn_error_string)(int);

struct Nccl {
  void* handle = nullptr;
  fn_get_unique_id get_unique_id = nullptr;
  fn_comm_init_rank comm_init_rank = nullptr;
  fn_all_reduce all_reduce = nullptr;
  fn_comm_destroy comm_destroy = nullptr;
  fn_comm_destroy comm_abort = nullptr;
  fn_error_string error_string = nullptr;
  NcclComm comm = nullptr;
  int rank = -1, world = 0;
};
static Nccl g_nccl;

static int load_nccl() {
  if (g_nccl.handle) return 0;
  const char* cands[4] = {getenv("WF2CU_NCCL_PATH"), "libnccl.so.2", "libnccl.so", nullptr};
  for (int i = 0; i < 3 && !g_nccl.handle; ++i)
    if (cands[i] && cands[i][0]) g_nccl.handle = dlopen(cands[i], RTLD_NOW | RTLD_GLOBAL);
  if (!g_nccl.handle) return set_error(WF2CU_ERR_COMM, "NCCL not found (dlopen libnccl.so.2 failed: %s); set WF2CU_NCCL_PATH", dlerror());
  g_nccl.get_unique_id = (fn_get_unique_id)dlsym(g_nccl.handle, "ncclGetUniqueId");
  g_nccl.comm_init_rank = (fn_comm_init_rank)dlsym(g_nccl.handle, "ncclCommInitRank");
  g_nccl.all_reduce = (fn_all_reduce)dlsym(g_nccl.handle, "ncclAllReduce");
  g_nccl.comm_destroy = (fn_comm_destroy)dlsym(g_nccl.handle, "ncclCommDestroy");
  g_nccl.comm_abort = (fn_comm_destroy)dlsym(g_nccl.handle, "ncclCommAbort");
  g_nccl.error_string = (fn_error_string)dlsym(g_nccl.handle, "ncclGetErrorString");
  if (!g_nccl.get_unique_id || !g_nccl.comm_init_rank || !g_nccl.all_reduce || !g_nccl.comm_destroy)
    return set_error(WF2CU_ERR_COMM, "NCCL library lacks a required symbol");
  return 0;
}

static int nccl_fail(int code, const char* what) {
  return set_error(WF2CU_ERR_COMM, "NCCL error %d (%s) in %s", code,
                   g_nccl.error_string ? g_nccl.error_string(code) : "?", what);
}

}  // namespace wf2

using namespace wf2;

extern "C" {

int wf2cu_comm_unique_id(void* out128) {
  WF2_CHECK(out128, "wf2cu_comm_unique_id: null out");
  int rc = load_nccl();
  if (rc) return rc;
  NcclId id;
  int e = g_nccl.get_unique_id(&id);
  if (e) return nccl_fail(e, "ncclGetUniqueId");
  memcpy(out128, &id, sizeof(id));
  return 0;
}

int wf2cu_comm_init(int rank, int world, const void* id128) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(id128 && world >= 1 && rank >= 0 && rank < world, "wf2cu_comm_init: bad rank/world (%d/%d)", rank, world);
  int rc = load_nccl();
  if (rc) return rc;
  if (g_nccl.comm) return set_error(WF2CU_ERR_COMM, "wf2cu_comm_init: communicator already exists");
  NcclId id;
  memcpy(&id, id128, sizeof(id));
  int e = g_nccl.comm_init_rank(&g_nccl.comm, world, id, rank);
  if (e) { g_nccl.comm = nullptr; return nccl_fail(e, "ncclCommInitRank"); }
  g_nccl.rank = rank; g_nccl.world = world;
  return 0;
}

int wf2cu_allreduce_sum_f32(void* buf, int64_t offset, int64_t count) {
  WF2_REQUIRE_INIT();
  if (!g_nccl.comm) return set_error(WF2CU_ERR_COMM, "wf2cu_allreduce_sum_f32: no communicator (call wf2cu_comm_init)");
  WF2_CHECK(buf && count >= 0, "wf2cu_allreduce_sum_f32: bad arguments");
  if (count == 0) return 0;
  float* p = (float*)buf + offset;
  int e = g_nccl.all_reduce(p, p, (size_t)count, /*ncclFloat32*/ 7, /*ncclSum*/ 0, g_nccl.comm, S().stream);
  if (e) return nccl_fail(e, "ncclAllReduce");
  S().launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

int wf2cu_comm_destroy(void) {
  if (g_nccl.comm) {
    // All work on the library stream is drained first.  ncclCommDestroy blocks for ever while a CUDA
    // graph that captured a collective of this communicator is still instantiated (measured on
    // B200 / NCCL 2.28.9), so tear down with ncclCommAbort, which does not wait on graph references.
    cudaStreamSynchronize(S().stream);
    if (g_nccl.comm_abort) g_nccl.comm_abort(g_nccl.comm);
    else g_nccl.comm_destroy(g_nccl.comm);
    g_nccl.comm = nullptr;
  }
  return 0;
}

}  // extern "C"
