// comm.cu — C1: the tensor-parallel all-reduce the reference leaves as a TODO after o_proj / down_proj
// (src/layers/linear.rs:236), plus the all-gather used by the sharded argmax (embed_head.rs:321-336 TODO).
//
// One process per GPU.  NCCL is resolved at run time with dlopen (the copy torch already loaded wins, else the
// system libnccl.so.2), so libnvllm.so has no link-time NCCL dependency and loads on machines without it.
// Calls are enqueued on the caller's stream and are CUDA-graph capturable (NCCL >= 2.9).
#include <dlfcn.h>
#include <string.h>

#include "common.cuh"

namespace nvllm {

// Minimal NCCL ABI (stable since 2.x): opaque comm, 128-byte unique id, enums below.
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclSum = 0 };
enum { ncclInt8 = 0, ncclUint8 = 1, ncclFloat32 = 7, ncclBfloat16 = 9 };

struct NcclApi {
  void* handle = nullptr;
  int (*GetUniqueId)(ncclUniqueId*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool ok = false;
};

static NcclApi& nccl() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api;
  tried = true;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    api.handle = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);  // already mapped by torch?
    if (api.handle) break;
  }
  if (!api.handle)
    for (const char* n : names) {
      api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (api.handle) break;
    }
  if (!api.handle) return api;
#define NVLLM_SYM(field, sym) *reinterpret_cast<void**>(&api.field) = dlsym(api.handle, sym)
  NVLLM_SYM(GetUniqueId, "ncclGetUniqueId");
  NVLLM_SYM(CommInitRank, "ncclCommInitRank");
  NVLLM_SYM(CommDestroy, "ncclCommDestroy");
  NVLLM_SYM(AllReduce, "ncclAllReduce");
  NVLLM_SYM(AllGather, "ncclAllGather");
  NVLLM_SYM(GetErrorString, "ncclGetErrorString");
#undef NVLLM_SYM
  api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllReduce && api.AllGather;
  return api;
}

static int nccl_status(int rc, const char* what) {
  if (rc == ncclSuccess) return 0;
  set_error("%s: NCCL error %d (%s)", what, rc, nccl().GetErrorString ? nccl().GetErrorString(rc) : "?");
  return NVLLM_ERR_COMM;
}

}  // namespace nvllm

struct nvllm_comm {
  nvllm::ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
};

extern "C" {

int nvllm_comm_unique_id(void* h_id_out) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(h_id_out != nullptr, NVLLM_ERR_ARG, "comm_unique_id: NULL");
  NVLLM_CHECK_ARG(nccl().ok, NVLLM_ERR_COMM, "NCCL library not found (libnccl.so.2)");
  static_assert(sizeof(ncclUniqueId) == NVLLM_UNIQUE_ID_BYTES, "unique id size");
  ncclUniqueId id;
  if (int rc = nccl_status(nccl().GetUniqueId(&id), "ncclGetUniqueId")) return rc;
  memcpy(h_id_out, &id, sizeof(id));
  return 0;
}

int nvllm_comm_init(nvllm_comm** comm, const void* h_id, int32_t rank, int32_t world) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(comm && h_id && world >= 1 && rank >= 0 && rank < world, NVLLM_ERR_ARG, "comm_init: arguments");
  NVLLM_CHECK_ARG(nccl().ok, NVLLM_ERR_COMM, "NCCL library not found (libnccl.so.2)");
  ncclUniqueId id;
  memcpy(&id, h_id, sizeof(id));
  nvllm_comm* c = new nvllm_comm();
  c->rank = rank;
  c->world = world;
  if (int rc = nccl_status(nccl().CommInitRank(&c->comm, world, id, rank), "ncclCommInitRank")) {
    delete c;
    return rc;
  }
  *comm = c;
  return 0;
}

int nvllm_comm_destroy(nvllm_comm* comm) {
  if (!comm) return 0;
  if (comm->comm && nvllm::nccl().ok) nvllm::nccl().CommDestroy(comm->comm);
  delete comm;
  return 0;
}

int nvllm_allreduce(nvllm_comm* comm, void* buf, size_t count, int32_t dtype, void* stream) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(comm && comm->comm && buf, NVLLM_ERR_ARG, "allreduce: NULL");
  NVLLM_CHECK_ARG(dtype == NVLLM_BF16 || dtype == NVLLM_F32, NVLLM_ERR_DTYPE, "allreduce: dtype %d", dtype);
  if (count == 0 || comm->world == 1) return 0;
  return nccl_status(nccl().AllReduce(buf, buf, count, dtype == NVLLM_BF16 ? ncclBfloat16 : ncclFloat32, ncclSum,
                                      comm->comm, static_cast<cudaStream_t>(stream)),
                     "ncclAllReduce");
}

int nvllm_allgather(nvllm_comm* comm, const void* send, void* recv, size_t bytes_per_rank, void* stream) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(comm && comm->comm && send && recv, NVLLM_ERR_ARG, "allgather: NULL");
  if (bytes_per_rank == 0) return 0;
  return nccl_status(nccl().AllGather(send, recv, bytes_per_rank, ncclUint8, comm->comm,
                                      static_cast<cudaStream_t>(stream)),
                     "ncclAllGather");
}

}  // extern "C"

namespace nvllm {
int comm_world(const nvllm_comm* c) { return c ? c->world : 1; }
int comm_rank(const nvllm_comm* c) { return c ? c->rank : 0; }
}  // namespace nvllm
