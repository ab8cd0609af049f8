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

// Symmetric region: [control block 4 KiB | user data].  Control block (all uint32, written remotely by peers):
//   ready[8] at +0, done[8] at +256, then local-only: counter at +512, epoch at +516.
struct nvllm_comm {
  nvllm::ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
  uint8_t* sym_local = nullptr;       // this rank's region (cudaMalloc)
  uint8_t* sym_peer[8] = {nullptr};   // every rank's region as mapped into this process (own entry == sym_local)
  size_t sym_bytes = 0;
};

namespace nvllm {
constexpr size_t kSymCtrlBytes = 4096;

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void wait_flag(const uint32_t* p, uint32_t target) {
  const long long t0 = clock64();
  while (static_cast<int32_t>(ld_acquire_sys(p) - target) < 0) {
    if (clock64() - t0 > 8000000000ll) {  // a peer never arrived: fail loudly instead of hanging the GPU
      printf("nvllm: peer flag timeout (block %d thread %d)\n", blockIdx.x, threadIdx.x);
      __trap();
    }
  }
}

struct FusedArParams {
  uint8_t* peer[8];          // symmetric regions of all ranks as mapped here
  size_t x_off, res_off, out_off;  // byte offsets of the partial sums / residual / normed buffers inside a region
  const bf16* weight;
  int rank, world, num_tokens, hidden;
  float eps;
};

// Two-shot all-reduce fused with the residual add and the RMSNorm that always follow it (qwen3.rs:382-385,389):
//   shot 1  rank r owns rows [r*R, (r+1)*R): it reads those rows of every rank's partial sums over NVLink (16-byte
//           peer loads, fixed rank order -> bit-identical on every rank), rounds the sum to bf16 (the all-reduced tensor),
//           adds the residual (bf16 tensor add), normalises in f32;
//   shot 2  it writes the normed rows into EVERY rank's buffer (peer stores); the new residual rows stay on the owner
//           (the residual stream is row-sharded across ranks between the fused kernels of one forward).
// Flags in the control block: `ready` (my partial sums are complete / my buffers may be overwritten) before shot 1,
// `done` (all my remote writes are visible) before the kernel may complete.  Epoch counters live in device memory and
// advance inside the kernel, so a captured CUDA graph can replay it.
__global__ void __launch_bounds__(1024)
allreduce_add_rmsnorm_kernel(const FusedArParams p) {
  pdl_launch_dependents();
  pdl_wait();
  uint32_t* ctrl = reinterpret_cast<uint32_t*>(p.peer[p.rank]);
  const uint32_t epoch = ctrl[129] + 1;  // +516 bytes; written only by this rank's last CTA of the previous call
  if (blockIdx.x == 0 && threadIdx.x < p.world) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(p.peer[threadIdx.x]) + p.rank, epoch);  // ready[rank] on peer
  }
  if (threadIdx.x < p.world) wait_flag(ctrl + threadIdx.x, epoch);  // everybody's partial sums are complete
  __syncthreads();

  const int rows_per = (p.num_tokens + p.world - 1) / p.world;
  const int row0 = p.rank * rows_per;
  const int row1 = min(p.num_tokens, row0 + rows_per);
  const int nvec = p.hidden >> 3;
  const uint4* wv = reinterpret_cast<const uint4*>(p.weight);
  __shared__ float red[32];
  __shared__ float s_inv;
  for (int row = row0 + blockIdx.x; row < row1; row += gridDim.x) {
    const size_t roff = static_cast<size_t>(row) * p.hidden * 2;
    float v[2][8];
    uint32_t sw[2][4];
    float ss = 0.f;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int i = threadIdx.x + k * blockDim.x;
      if (i < nvec) {
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int q = 0; q < p.world; ++q) {  // fixed order: identical bits on every rank
          const uint4 a = ld_nc_16(p.peer[q] + p.x_off + roff + i * 16);
          const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) { acc[2 * j] += bf16_lo(aw[j]); acc[2 * j + 1] += bf16_hi(aw[j]); }
        }
        const uint4 r = *reinterpret_cast<const uint4*>(p.peer[p.rank] + p.res_off + roff + i * 16);
        const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t o = pack_bf16x2(acc[2 * j], acc[2 * j + 1]);                        // the all-reduced bf16 tensor
          sw[k][j] = pack_bf16x2(bf16_lo(o) + bf16_lo(rw[j]), bf16_hi(o) + bf16_hi(rw[j]));  // + residual (bf16 add)
          v[k][2 * j] = bf16_lo(sw[k][j]);
          v[k][2 * j + 1] = bf16_hi(sw[k][j]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) ss += v[k][j] * v[k][j];
      }
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x < 32) {
      float t = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.f;
      t = warp_sum(t);
      if (threadIdx.x == 0) s_inv = 1.0f / sqrtf(t / static_cast<float>(p.hidden) + p.eps);
    }
    __syncthreads();
    const float inv = s_inv;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int i = threadIdx.x + k * blockDim.x;
      if (i < nvec) {
        const uint4 w = __ldg(wv + i);
        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
        uint4 o;
        uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          ow[j] = pack_bf16x2(v[k][2 * j] * inv * bf16_lo(ww[j]), v[k][2 * j + 1] * inv * bf16_hi(ww[j]));
        // The residual stream stays ROW-SHARDED: its only readers are these fused kernels, and row `row` is always
        // reduced by this rank, so the new residual is written locally only (halves the NVLink write traffic).
        *reinterpret_cast<uint4*>(p.peer[p.rank] + p.res_off + roff + i * 16) = make_uint4(sw[k][0], sw[k][1], sw[k][2], sw[k][3]);
        for (int q = 0; q < p.world; ++q)  // shot 2: broadcast the normed row to every rank
          *reinterpret_cast<uint4*>(p.peer[q] + p.out_off + roff + i * 16) = o;
      }
    }
    __syncthreads();  // red / s_inv reuse
  }
  // ---- completion: the last CTA of this rank tells every peer, then waits for all of them -----------------------
  __threadfence_system();
  __syncthreads();
  __shared__ uint32_t s_last;
  if (threadIdx.x == 0) s_last = (atomicAdd(ctrl + 128, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (s_last) {
    if (threadIdx.x < p.world) {
      st_release_sys(reinterpret_cast<uint32_t*>(p.peer[threadIdx.x]) + 64 + p.rank, epoch);  // done[rank] on peer
      wait_flag(ctrl + 64 + threadIdx.x, epoch);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      ctrl[128] = 0;
      ctrl[129] = epoch;
      __threadfence();
    }
  }
}
}  // namespace nvllm

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

// Collective: every rank allocates `bytes` (+ control block), the CUDA IPC handles travel through an NCCL all-gather,
// peers are mapped with cudaIpcOpenMemHandle.  *local_base receives this rank's user-data pointer.
int nvllm_comm_symm_alloc(nvllm_comm* comm, size_t bytes, void** local_base) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(comm && comm->comm && local_base && bytes > 0, NVLLM_ERR_ARG, "comm_symm_alloc: arguments");
  NVLLM_CHECK_ARG(comm->sym_local == nullptr, NVLLM_ERR_STATE, "comm_symm_alloc: already allocated");
  NVLLM_CHECK_ARG(comm->world <= 8, NVLLM_ERR_ARG, "comm_symm_alloc: world %d > 8", comm->world);
  const size_t total = kSymCtrlBytes + ((bytes + 255) & ~static_cast<size_t>(255));
  void* base = nullptr;
  NVLLM_CUDA(cudaMalloc(&base, total));
  NVLLM_CUDA(cudaMemset(base, 0, total));
  cudaIpcMemHandle_t mine;
  NVLLM_CUDA(cudaIpcGetMemHandle(&mine, base));
  cudaIpcMemHandle_t* d_all = nullptr;
  NVLLM_CUDA(cudaMalloc(reinterpret_cast<void**>(&d_all), sizeof(mine) * (comm->world + 1)));
  NVLLM_CUDA(cudaMemcpy(d_all + comm->world, &mine, sizeof(mine), cudaMemcpyHostToDevice));
  cudaStream_t st;
  NVLLM_CUDA(cudaStreamCreate(&st));
  int rc = nccl_status(nccl().AllGather(d_all + comm->world, d_all, sizeof(mine), ncclUint8, comm->comm, st), "ncclAllGather(ipc)");
  if (rc == 0) rc = cuda_status(cudaStreamSynchronize(st), "ipc all-gather");
  cudaStreamDestroy(st);
  cudaIpcMemHandle_t all[8];
  if (rc == 0) rc = cuda_status(cudaMemcpy(all, d_all, sizeof(mine) * comm->world, cudaMemcpyDeviceToHost), "ipc handles D2H");
  cudaFree(d_all);
  if (rc != 0) { cudaFree(base); return rc; }
  for (int r = 0; r < comm->world; ++r) {
    if (r == comm->rank) { comm->sym_peer[r] = static_cast<uint8_t*>(base); continue; }
    void* pp = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&pp, all[r], cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      set_error("cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
      cudaFree(base);
      return NVLLM_ERR_COMM;
    }
    comm->sym_peer[r] = static_cast<uint8_t*>(pp);
  }
  comm->sym_local = static_cast<uint8_t*>(base);
  comm->sym_bytes = total;
  *local_base = comm->sym_local + kSymCtrlBytes;
  return 0;
}

int nvllm_allreduce_add_rmsnorm(nvllm_comm* comm, void* out, void* x, void* residual, const void* weight,
                                int32_t num_tokens, int32_t hidden, float eps, int32_t dtype, void* stream) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(comm && comm->sym_local, NVLLM_ERR_STATE, "allreduce_add_rmsnorm: no symmetric region (nvllm_comm_symm_alloc)");
  NVLLM_CHECK_ARG(dtype == NVLLM_BF16, NVLLM_ERR_DTYPE, "allreduce_add_rmsnorm: dtype %d", dtype);
  NVLLM_CHECK_ARG(num_tokens >= 0 && hidden > 0 && hidden % 8 == 0 && hidden <= 16384, NVLLM_ERR_ARG, "allreduce_add_rmsnorm: sizes");
  if (num_tokens == 0) return 0;
  auto off = [&](const void* ptr, size_t* o) {
    const uint8_t* b = static_cast<const uint8_t*>(ptr);
    if (b < comm->sym_local + kSymCtrlBytes || b + static_cast<size_t>(num_tokens) * hidden * 2 > comm->sym_local + comm->sym_bytes) return false;
    *o = static_cast<size_t>(b - comm->sym_local);
    return (*o & 15) == 0;
  };
  FusedArParams p;
  NVLLM_CHECK_ARG(off(out, &p.out_off) && off(x, &p.x_off) && off(residual, &p.res_off) && weight, NVLLM_ERR_ARG,
                  "allreduce_add_rmsnorm: out / x / residual must be 16-byte aligned buffers inside the symmetric region");
  for (int r = 0; r < 8; ++r) p.peer[r] = r < comm->world ? comm->sym_peer[r] : nullptr;
  p.weight = static_cast<const bf16*>(weight);
  p.rank = comm->rank; p.world = comm->world; p.num_tokens = num_tokens; p.hidden = hidden; p.eps = eps;
  const int nvec = hidden / 8;
  int threads = nvec <= 256 ? (nvec + 31) / 32 * 32 : ((nvec + 1) / 2 + 31) / 32 * 32;
  if (threads > 1024) threads = 1024;
  if (threads < 32 * ((comm->world + 31) / 32)) threads = 32;
  const int rows_per = (num_tokens + comm->world - 1) / comm->world;
  int grid = rows_per < 2 * sm_count() ? rows_per : 2 * sm_count();
  if (grid < 1) grid = 1;
  NVLLM_CUDA(launch_pdl(allreduce_add_rmsnorm_kernel, dim3(grid), dim3(threads), 0, static_cast<cudaStream_t>(stream), p));
  return 0;
}

int nvllm_comm_destroy(nvllm_comm* comm) {
  if (!comm) return 0;
  cudaDeviceSynchronize();  // no kernel of this rank may still be reading peer memory
  for (int r = 0; r < comm->world; ++r)
    if (r != comm->rank && comm->sym_peer[r]) cudaIpcCloseMemHandle(comm->sym_peer[r]);
  // Freeing exported memory while another process still has it mapped is undefined behaviour (observed as a hang at
  // process teardown): rendezvous once more — every rank has closed its peer mappings before anyone frees its region.
  if (comm->sym_local && comm->comm && nvllm::nccl().ok && comm->world > 1) {
    cudaStream_t st = nullptr;
    if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) == cudaSuccess) {
      nvllm::nccl().AllReduce(comm->sym_local, comm->sym_local, 1, nvllm::ncclFloat32, nvllm::ncclSum, comm->comm, st);
      cudaStreamSynchronize(st);
      cudaStreamDestroy(st);
    }
  }
  if (comm->sym_local) cudaFree(comm->sym_local);
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
