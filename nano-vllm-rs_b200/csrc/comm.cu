// comm.cu — C1: the tensor-parallel all-reduce the reference leaves as a TODO after o_proj / down_proj
// (src/layers/linear.rs:236), plus the all-gather used by the sharded argmax (embed_head.rs:321-336 TODO).
//
// One process per GPU.  NCCL is resolved at run time with dlopen (the copy torch already loaded wins, else the
// system libnccl.so.2), so libnvllm.so has no link-time NCCL dependency and loads on machines without it.
// Calls are enqueued on the caller's stream and are CUDA-graph capturable (NCCL >= 2.9).
#include <dlfcn.h>
#include <string.h>

#include <atomic>
#include <chrono>
#include <thread>

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
  int (*CommAbort)(ncclComm_t) = nullptr;  // optional
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
  NVLLM_SYM(CommAbort, "ncclCommAbort");
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


// Host-side deadline for work this library enqueued on a stream of its own (the NCCL collectives of set-up and
// tear-down): an NCCL kernel waiting for a rank that never joins has no timeout of its own and would keep the host in
// cudaStreamSynchronize for ever.  0 = finished, NVLLM_ERR_COMM = still running after 6 x NVLLM_COMM_TIMEOUT_MS (ranks reach
// set-up and tear-down less in step than they reach a decode step; the stream is left as it is: the caller must not free
// what the kernel may still touch), > 0 = cudaError_t.
static int stream_wait_bounded(cudaStream_t st, const char* what) {
  const auto t0 = std::chrono::steady_clock::now();
  const long long limit_ms = 6 * knobs().comm_timeout_ms;
  for (int spins = 0;; ++spins) {
    const cudaError_t e = cudaStreamQuery(st);
    if (e == cudaSuccess) return 0;
    if (e != cudaErrorNotReady) return cuda_status(e, what);
    const long long ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
    if (ms > limit_ms) {
      set_error("%s: still running after %lld ms (6 x NVLLM_COMM_TIMEOUT_MS) - a peer rank never joined the collective", what, ms);
      return NVLLM_ERR_COMM;
    }
    if (spins > 2000) std::this_thread::sleep_for(std::chrono::microseconds(200));
  }
}

}  // namespace nvllm

// Symmetric region: [control block 4 KiB | low-latency area: 3 x (stage, bcast) of ll_cap bytes | user data].
// Control block (all uint32; the first two arrays are written remotely by peers):  ready[8] at +0, done[8] at +256, then
// local-only: CTA counter at +512, epoch at +516, dead at +520, low-latency buffer index at +524, bytes in use of each
// low-latency buffer at +528..+539, low-latency CTA counter at +540.
struct nvllm_comm {
  nvllm::ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
  uint8_t* sym_local = nullptr;       // this rank's region (cudaMalloc)
  uint8_t* sym_peer[8] = {nullptr};   // every rank's region as mapped into this process (own entry == sym_local)
  size_t sym_bytes = 0;
  size_t ll_cap = 0;                  // bytes of ONE low-latency buffer (6 of them sit between the control block and user data)
  size_t user_off = 0;                // offset of the user data inside the region
  int* h_err = nullptr;               // pinned + mapped: a fused kernel that gave up waiting for a peer sets it
  int* d_err = nullptr;               // device alias of h_err
};

namespace nvllm {
int gemm_bf16_impl(void* y, const void* x, const void* w, int32_t m, int32_t n, int32_t k, int64_t ldx, int64_t ldw,
                   int64_t ldy, int32_t out_dtype, void* stream, const RowPush& ps);  // gemm_tcgen05.cu
constexpr size_t kSymCtrlBytes = 4096;

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}
// 16-byte COHERENT load of data another GPU wrote while this kernel may already be running (ordered after the flag
// acquire): never through the non-coherent path, never from L1 — ld.global.nc is outside the PTX memory model for
// memory that changes during the kernel's lifetime.
__device__ __forceinline__ uint4 ld_sys_16(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}
// What a kernel that gave up was waiting for, for nvllm_comm_error's message.  err[] is host-mapped: [0] sticky flag,
// [1] phase (1 / 2 = flag protocol waiting for ready / done, 3 = low-latency reduce waiting for staged partial rows,
// 4 = low-latency gather waiting for normed rows), [2] row (or the peer whose flag is missing), [3] CTA, [4] epoch or
// buffer pair, [5] bit mask of the source ranks still missing (phase 3).  Several threads may report: any one wins.
__device__ __forceinline__ void report_timeout(int* err, int phase, int what, int aux, int mask) {
  volatile int* e = err;
  if (e[0] == 0) { e[1] = phase; e[2] = what; e[3] = static_cast<int>(blockIdx.x); e[4] = aux; e[5] = mask; }
  e[0] = 1;
}
// Spins until *p reaches `target` (wrap-safe).  A peer that never arrives (crashed rank, a host-side stall longer than
// NVLLM_COMM_TIMEOUT_MS) must neither hang the GPU nor poison the CUDA context: the kernel raises the host-visible error
// flag, marks the communicator dead (later calls stop waiting) and runs to completion; the host reports NVLLM_ERR_COMM.
__device__ __forceinline__ void wait_flag(const uint32_t* p, uint32_t target, long long timeout_cycles, uint32_t* dead,
                                          int* err, int phase = 1) {
  if (static_cast<int32_t>(ld_acquire_sys(p) - target) >= 0) return;
  const long long t0 = clock64();
  while (static_cast<int32_t>(ld_acquire_sys(p) - target) < 0) {
    if (clock64() - t0 > timeout_cycles) {
      *reinterpret_cast<volatile uint32_t*>(dead) = 1u;
      report_timeout(err, phase, static_cast<int>(threadIdx.x), static_cast<int>(target), 0);
      __threadfence_system();
      return;
    }
  }
}

struct FusedArParams {
  uint8_t* peer[8];          // symmetric regions of all ranks as mapped here
  uint8_t* self;             // == peer[rank] (kept apart: a run-time index into peer[] would spill the parameter block)
  size_t x_off, res_off, out_off;  // byte offsets of the partial sums / residual / normed buffers inside a region
  const bf16* weight;
  int rank, num_tokens, hidden;
  float eps;
  long long timeout_cycles;
  int* err;
};

// Two-shot all-reduce fused with the residual add and the RMSNorm that always follow it (qwen3.rs:382-385,389):
//   shot 1  rank r owns rows [r*R, (r+1)*R): it reads those rows of every rank's partial sums over NVLink — WORLD is a
//           template parameter and a thread's WORLD x min(VPT, 2) 16-byte peer loads are all issued before the first add,
//           so one CTA keeps up to 64 KiB in flight per row (the round-1 kernel issued its peer loads one behind the other
//           through a run-time loop: ~8x less memory-level parallelism on a ~2 us NVLink round trip).  Fixed rank order of
//           the adds -> bit-identical on every rank.  The sum is rounded to bf16 (the all-reduced tensor), the residual
//           added (bf16 tensor add), the row normalised in f32;
//   shot 2  it writes the normed rows into EVERY rank's buffer (peer stores); the new residual rows stay on the owner
//           (the residual stream is row-sharded across ranks between the fused kernels of one forward).
// Flags in the control block: `ready` (my partial sums are complete / my buffers may be overwritten) before shot 1,
// `done` (all my remote writes are visible) before the kernel may complete.  Epoch counters live in device memory and
// advance inside the kernel, so a captured CUDA graph can replay it.
// p.peer[q] for a run-time q without indexing the parameter array dynamically (which would copy it to local memory).
__device__ __forceinline__ uint8_t* peer_of(const FusedArParams& p, unsigned q) {
  uint8_t* r = p.peer[0];
#pragma unroll
  for (unsigned i = 1; i < 8; ++i) r = (q == i) ? p.peer[i] : r;
  return r;
}

template <int WORLD, int VPT>
__global__ void __launch_bounds__(512)
allreduce_add_rmsnorm_kernel(const FusedArParams p) {
  pdl_launch_dependents();
  pdl_wait();
  uint32_t* ctrl = reinterpret_cast<uint32_t*>(p.self);
  const uint32_t epoch = ld_state(ctrl + 129) + 1;  // +516 bytes; written only by this rank's last CTA of the previous call
  const bool dead = ld_state(ctrl + 130) != 0;  // a previous call timed out: never wait again (the host reports the error)
  if (blockIdx.x == 0 && threadIdx.x < WORLD) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(peer_of(p, threadIdx.x)) + p.rank, epoch);  // ready[rank] on peer
  }
  if (threadIdx.x < WORLD && !dead) wait_flag(ctrl + threadIdx.x, epoch, p.timeout_cycles, ctrl + 130, p.err);
  __syncthreads();  // everybody's partial sums are complete (and ordered before the loads below: acquire + barrier)

  const int rows_per = (p.num_tokens + WORLD - 1) / WORLD;
  const int row0 = p.rank * rows_per;
  const int row1 = min(p.num_tokens, row0 + rows_per);
  const int nvec = p.hidden >> 3;
  const uint4* wv = reinterpret_cast<const uint4*>(p.weight);
  constexpr int KCH = VPT < 2 ? VPT : 2;  // vectors per thread whose peer loads are in flight together
  __shared__ float red[16];
  __shared__ float s_inv;
  for (int row = row0 + blockIdx.x; row < row1; row += gridDim.x) {
    const size_t roff = static_cast<size_t>(row) * p.hidden * 2;
    uint32_t sw[VPT][4];  // the new residual row (packed bf16), kept for the second pass
    float ss = 0.f;
#pragma unroll
    for (int k0 = 0; k0 < VPT; k0 += KCH) {
      uint4 a[KCH][WORLD], r[KCH];
#pragma unroll
      for (int k = 0; k < KCH; ++k) {
        const int i = threadIdx.x + (k0 + k) * blockDim.x;
        if (i < nvec) {
#pragma unroll
          for (int q = 0; q < WORLD; ++q) a[k][q] = ld_sys_16(p.peer[q] + p.x_off + roff + i * 16);
          r[k] = *reinterpret_cast<const uint4*>(p.self + p.res_off + roff + i * 16);
        }
      }
#pragma unroll
      for (int k = 0; k < KCH; ++k) {
        const int i = threadIdx.x + (k0 + k) * blockDim.x;
        if (i < nvec) {
          float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int q = 0; q < WORLD; ++q) {  // fixed order: identical bits whichever rank owns the row
            const uint32_t aw[4] = {a[k][q].x, a[k][q].y, a[k][q].z, a[k][q].w};
#pragma unroll
            for (int j = 0; j < 4; ++j) { acc[2 * j] += bf16_lo(aw[j]); acc[2 * j + 1] += bf16_hi(aw[j]); }
          }
          const uint32_t rw[4] = {r[k].x, r[k].y, r[k].z, r[k].w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t o = pack_bf16x2(acc[2 * j], acc[2 * j + 1]);                          // the all-reduced bf16 tensor
            const uint32_t h = pack_bf16x2(bf16_lo(o) + bf16_lo(rw[j]), bf16_hi(o) + bf16_hi(rw[j]));  // + residual (bf16 add)
            sw[k0 + k][j] = h;
            const float v0 = bf16_lo(h), v1 = bf16_hi(h);
            ss += v0 * v0 + v1 * v1;
          }
        }
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
    for (int k = 0; k < VPT; ++k) {
      const int i = threadIdx.x + k * blockDim.x;
      if (i < nvec) {
        const uint4 w = __ldg(wv + i);
        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
        uint4 o;
        uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          ow[j] = pack_bf16x2(bf16_lo(sw[k][j]) * inv * bf16_lo(ww[j]), bf16_hi(sw[k][j]) * inv * bf16_hi(ww[j]));
        // The residual stream stays ROW-SHARDED: its only readers are these fused kernels, and row `row` is always
        // reduced by this rank, so the new residual is written locally only (halves the NVLink write traffic).
        *reinterpret_cast<uint4*>(p.self + p.res_off + roff + i * 16) = make_uint4(sw[k][0], sw[k][1], sw[k][2], sw[k][3]);
#pragma unroll
        for (int q = 0; q < WORLD; ++q)  // shot 2: broadcast the normed row to every rank
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
    if (threadIdx.x < WORLD) {
      st_release_sys(reinterpret_cast<uint32_t*>(peer_of(p, threadIdx.x)) + 64 + p.rank, epoch);  // done[rank] on peer
      if (!dead) wait_flag(ctrl + 64 + threadIdx.x, epoch, p.timeout_cycles, ctrl + 130, p.err, 2);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      ctrl[128] = 0;
      ctrl[129] = epoch;
      __threadfence();
    }
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Low-latency variant for decode-sized messages: same arithmetic (bit for bit, up to the sign of zero), no flags.
//
// The flag protocol above costs two system-scope barriers per call (flag store -> peer poll, ~2-3 us each), a pull over
// NVLink (a full round trip) and a system fence that waits for every remote write to be acknowledged: ~17 us per call at
// 2 ranks however small the message.  Here the DATA is the signal: every 32-bit word of a buffer that has not been written
// yet holds kLLEmpty (two bf16 -0.0), senders replace that one bit pattern by +0.0 (numerically the same value), and a
// receiver spins on the 16-byte vector itself until none of its four words is kLLEmpty.  One NVLink one-way trip per shot:
//   A  push   every rank stores its partial rows straight into the OWNER's stage buffer   stage[src rank][row][hidden]
//   B  reduce the owner polls the world-1 staged copies of a vector (its own partial is local), adds in rank order, rounds,
//             adds the residual, normalises, stores the normed row locally and pushes it into every peer's bcast buffer
//   C  gather every rank polls the rows it does not own in its bcast buffer and copies them to `out`
// Three (stage, bcast) pairs rotate per call; a call clears the pair two calls ahead.  Why that is race-free: a peer can
// be at most ONE call ahead of this rank (to finish call i+1 it needs this rank's partial rows of call i+1), so during
// call i nothing can arrive for pair (i+2) % 3; everything that was ever pushed into that pair during call i-1 has been
// polled (every vector, phases B and C), so nothing can land late either; and the clear completes with this kernel,
// before this rank enters call i+1.  x / residual / out are ordinary local buffers for this kernel.
// Every CTA waits for remote CTAs that may not have started: the grid must be co-resident (launch_fused_ar_ll sizes it
// from the occupancy calculator), and phase A comes first in every CTA, unconditionally.

#ifdef NVLLM_TIMELINE
// Debug build only (make timeline, tools/ar_timeline.py): thread 0 of every CTA stamps %globaltimer at the phase boundaries.
__device__ unsigned long long* g_ar_tl_buf = nullptr;   // [cap][12]: {blockIdx, t0..t8, rows reduced, smid}
__device__ unsigned int g_ar_tl_cursor = 0;
__device__ unsigned int g_ar_tl_cap = 0;
__device__ __forceinline__ unsigned long long ar_tl_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
  return t;
}
#define ATL(i) do { if (threadIdx.x == 0) atl[i] = ar_tl_now(); } while (0)
#else
#define ATL(i) do { } while (0)
#endif

__device__ __forceinline__ bool ll_arrived(const uint4& v) {
  return v.x != kLLEmpty && v.y != kLLEmpty && v.z != kLLEmpty && v.w != kLLEmpty;
}
struct LLArParams {
  uint8_t* peer[8];          // symmetric regions of all ranks as mapped here
  uint8_t* self;
  const uint8_t* x;          // this rank's partial sums [num_tokens, hidden] (local)
  uint8_t* res;              // residual, row-sharded: only this rank's rows are read / written
  uint8_t* out;              // normed rows, all of them (local)
  const bf16* weight;
  size_t ll_off, ll_cap;     // low-latency area inside a region; bytes of one buffer
  int rank, num_tokens, hidden;
  int pushed;                // phase A already happened in the epilogue of the producing GEMM (RowPush)
  float eps;
  long long timeout_cycles;
  int* err;
};
__device__ __forceinline__ uint8_t* peer_of(const LLArParams& p, unsigned q) {
  uint8_t* r = p.peer[0];
#pragma unroll
  for (unsigned i = 1; i < 8; ++i) r = (q == i) ? p.peer[i] : r;
  return r;
}

template <int WORLD, int VPT>
__global__ void __launch_bounds__(512)
allreduce_add_rmsnorm_ll_kernel(const LLArParams p) {
#ifdef NVLLM_TIMELINE
  unsigned long long atl[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  unsigned int atl_rows = 0;
#endif
  ATL(0);
  pdl_launch_dependents();
  pdl_wait();
  ATL(1);
  uint32_t* ctrl = reinterpret_cast<uint32_t*>(p.self);
  const uint32_t b = ld_state(ctrl + 131);                  // pair of this call; advanced by the last CTA of the call
  const uint32_t bclr = b >= 1 ? b - 1 : 2;              // (b + 2) % 3
  const uint32_t clr_bytes = ld_state(ctrl + 132 + bclr);   // what the last call on that pair used
  uint32_t* dead_p = ctrl + 130;
  const bool dead = ld_state(dead_p) != 0;
  const int R = (p.num_tokens + WORLD - 1) / WORLD;      // rows per owner
  const size_t row_bytes = static_cast<size_t>(p.hidden) * 2;
  const size_t stage_off = p.ll_off + static_cast<size_t>(b) * 2 * p.ll_cap;
  const size_t bcast_off = stage_off + p.ll_cap;
  const int nvec = p.hidden >> 3;

  // ---- A: push my partial rows to their owners -------------------------------------------------------------------
  for (int row = blockIdx.x; row < p.num_tokens && !p.pushed; row += gridDim.x) {
    const int owner = row / R;
    if (owner == p.rank) continue;
    const uint8_t* src = p.x + row * row_bytes;
    uint8_t* dst = peer_of(p, owner) + stage_off + (static_cast<size_t>(p.rank) * R + (row - owner * R)) * row_bytes;
    uint4 v[VPT];
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int i = threadIdx.x + k * blockDim.x;
      if (i < nvec) v[k] = *reinterpret_cast<const uint4*>(src + i * 16);
    }
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int i = threadIdx.x + k * blockDim.x;
      if (i < nvec) st_sys_16(dst + i * 16, ll_scrub(v[k]));
    }
  }
  ATL(2);  // partial rows pushed (stores issued)
  // ---- housekeeping while the pushes fly: clear the pair two calls ahead ------------------------------------------
  {
    uint4* c0 = reinterpret_cast<uint4*>(p.self + p.ll_off + static_cast<size_t>(bclr) * 2 * p.ll_cap);
    uint4* c1 = reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(c0) + p.ll_cap);
    const uint4 e = make_uint4(kLLEmpty, kLLEmpty, kLLEmpty, kLLEmpty);
    const uint32_t n16 = clr_bytes >> 4;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += gridDim.x * blockDim.x) { c0[i] = e; c1[i] = e; }
  }
  ATL(3);  // clear issued
  // ---- B: reduce + add + norm the rows I own, broadcast them -----------------------------------------------------
  const int row0 = p.rank * R;
  const int row1 = min(p.num_tokens, row0 + R);
  const uint4* wv = reinterpret_cast<const uint4*>(p.weight);
  constexpr int KCH = VPT < 2 ? VPT : 2;
  __shared__ float red[16];
  __shared__ float s_inv;
  for (int row = row0 + blockIdx.x; row < row1; row += gridDim.x) {
    const size_t roff = row * row_bytes;
    const uint8_t* st_row = p.self + stage_off + static_cast<size_t>(row - row0) * row_bytes;  // + q * R * row_bytes
    uint32_t sw[VPT][4];
    float ss = 0.f;
    uint4 wreg[VPT];  // norm weights: requested before the polls, not after the row statistic (one L2 latency off the path)
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int i = threadIdx.x + k * blockDim.x;
      if (i < nvec) wreg[k] = __ldg(wv + i);
    }
#pragma unroll
    for (int k0 = 0; k0 < VPT; k0 += KCH) {
      uint4 a[KCH][WORLD], r[KCH];
#pragma unroll
      for (int k = 0; k < KCH; ++k) {
        const int i = threadIdx.x + (k0 + k) * blockDim.x;
        if (i < nvec) {
#pragma unroll
          for (int q = 0; q < WORLD; ++q)
            a[k][q] = (q == p.rank) ? *reinterpret_cast<const uint4*>(p.x + roff + i * 16)
                                    : ld_sys_16(st_row + static_cast<size_t>(q) * R * row_bytes + i * 16);
          r[k] = *reinterpret_cast<const uint4*>(p.res + roff + i * 16);
        }
      }
      if (!dead) {
        // Re-poll in ROUNDS: every vector that has not arrived is re-loaded in the same round (all in flight together), so
        // a round costs one L2 round trip however many of the KCH x (WORLD-1) vectors are still missing (spinning on them
        // one after the other could cost up to 14 round trips per thread at 8 ranks; measured it made no difference on an
        // idle box — 14.4 us per 640 KiB call either way, profiles/r02_scaling.md — but it bounds the worst case).
        long long t0 = 0;
        while (true) {
          bool miss = false;
#pragma unroll
          for (int k = 0; k < KCH; ++k)
#pragma unroll
            for (int q = 0; q < WORLD; ++q)
              miss |= (q != p.rank) && (threadIdx.x + (k0 + k) * blockDim.x < nvec) && !ll_arrived(a[k][q]);
          if (!miss) break;
          if (t0 == 0) t0 = clock64();
          if (clock64() - t0 > p.timeout_cycles || *reinterpret_cast<volatile uint32_t*>(dead_p) != 0) {
            *reinterpret_cast<volatile uint32_t*>(dead_p) = 1u;
            int missing = 0;
#pragma unroll
            for (int k = 0; k < KCH; ++k)
#pragma unroll
              for (int q = 0; q < WORLD; ++q)
                if (q != p.rank && threadIdx.x + (k0 + k) * blockDim.x < nvec && !ll_arrived(a[k][q])) missing |= 1 << q;
            report_timeout(p.err, 3, row, static_cast<int>(b), missing);
            __threadfence_system();
            break;
          }
#pragma unroll
          for (int k = 0; k < KCH; ++k) {
            const int i = threadIdx.x + (k0 + k) * blockDim.x;
#pragma unroll
            for (int q = 0; q < WORLD; ++q)
              if (q != p.rank && i < nvec && !ll_arrived(a[k][q]))
                a[k][q] = ld_sys_16(st_row + static_cast<size_t>(q) * R * row_bytes + i * 16);
          }
        }
      }
#pragma unroll
      for (int k = 0; k < KCH; ++k) {
        const int i = threadIdx.x + (k0 + k) * blockDim.x;
        if (i < nvec) {
          float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int q = 0; q < WORLD; ++q) {  // fixed order: identical bits whichever rank owns the row
            const uint32_t aw[4] = {a[k][q].x, a[k][q].y, a[k][q].z, a[k][q].w};
#pragma unroll
            for (int j = 0; j < 4; ++j) { acc[2 * j] += bf16_lo(aw[j]); acc[2 * j + 1] += bf16_hi(aw[j]); }
          }
          const uint32_t rw[4] = {r[k].x, r[k].y, r[k].z, r[k].w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t o = pack_bf16x2(acc[2 * j], acc[2 * j + 1]);
            const uint32_t h = pack_bf16x2(bf16_lo(o) + bf16_lo(rw[j]), bf16_hi(o) + bf16_hi(rw[j]));
            sw[k0 + k][j] = h;
            const float v0 = bf16_lo(h), v1 = bf16_hi(h);
            ss += v0 * v0 + v1 * v1;
          }
        }
      }
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
#ifdef NVLLM_TIMELINE
    if (atl_rows++ == 0) ATL(4);  // first owned row: every staged vector of the CTA has arrived
#endif
    if (threadIdx.x < 32) {
      float t = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.f;
      t = warp_sum(t);
      if (threadIdx.x == 0) s_inv = 1.0f / sqrtf(t / static_cast<float>(p.hidden) + p.eps);
    }
    __syncthreads();
    const float inv = s_inv;
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int i = threadIdx.x + k * blockDim.x;
      if (i < nvec) {
        const uint32_t ww[4] = {wreg[k].x, wreg[k].y, wreg[k].z, wreg[k].w};
        uint4 o;
        uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          ow[j] = pack_bf16x2(bf16_lo(sw[k][j]) * inv * bf16_lo(ww[j]), bf16_hi(sw[k][j]) * inv * bf16_hi(ww[j]));
        o = ll_scrub(o);  // the same bits on the owner and on the receivers
#pragma unroll
        for (int q = 0; q < WORLD; ++q)
          if (q != p.rank) st_sys_16(p.peer[q] + bcast_off + roff + i * 16, o);
        *reinterpret_cast<uint4*>(p.res + roff + i * 16) = make_uint4(sw[k][0], sw[k][1], sw[k][2], sw[k][3]);
        *reinterpret_cast<uint4*>(p.out + roff + i * 16) = o;
      }
    }
    __syncthreads();  // red / s_inv reuse
  }
  ATL(5);  // owned rows reduced and broadcast
  // ---- C: gather the rows the other ranks own ----------------------------------------------------------------------
  for (int row = blockIdx.x; row < p.num_tokens; row += gridDim.x) {
    if (row / R == p.rank) continue;
    const uint8_t* src = p.self + bcast_off + row * row_bytes;
    uint4 v[VPT];
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int i = threadIdx.x + k * blockDim.x;
      if (i < nvec) v[k] = ld_sys_16(src + i * 16);
    }
    if (!dead) {
      long long t0 = 0;
      while (true) {  // rounds, as above
        bool miss = false;
#pragma unroll
        for (int k = 0; k < VPT; ++k) miss |= (threadIdx.x + k * blockDim.x < nvec) && !ll_arrived(v[k]);
        if (!miss) break;
        if (t0 == 0) t0 = clock64();
        if (clock64() - t0 > p.timeout_cycles || *reinterpret_cast<volatile uint32_t*>(dead_p) != 0) {
          *reinterpret_cast<volatile uint32_t*>(dead_p) = 1u;
          report_timeout(p.err, 4, row, static_cast<int>(b), 0);
          __threadfence_system();
          break;
        }
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
          const int i = threadIdx.x + k * blockDim.x;
          if (i < nvec && !ll_arrived(v[k])) v[k] = ld_sys_16(src + i * 16);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int i = threadIdx.x + k * blockDim.x;
      if (i < nvec) *reinterpret_cast<uint4*>(p.out + row * row_bytes + i * 16) = v[k];
    }
  }
  ATL(6);  // foreign rows gathered
  // ---- the last CTA rotates the pair -------------------------------------------------------------------------------
  __syncthreads();
  ATL(7);
#ifdef NVLLM_TIMELINE
  if (threadIdx.x == 0 && g_ar_tl_buf != nullptr) {
    const unsigned int slot = atomicAdd(&g_ar_tl_cursor, 1u);
    if (slot < g_ar_tl_cap) {
      unsigned long long* e = g_ar_tl_buf + static_cast<size_t>(slot) * 12;
      unsigned int smid;
      asm volatile("mov.u32 %0, %%smid;\n" : "=r"(smid));
      e[0] = blockIdx.x;
      for (int i = 0; i < 9; ++i) e[1 + i] = atl[i];
      e[10] = atl_rows;
      e[11] = smid;
    }
  }
#endif
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(ctrl + 135, 1u) == gridDim.x - 1) {
      ctrl[135] = 0;
      ctrl[132 + b] = static_cast<uint32_t>(static_cast<size_t>(WORLD) * R * row_bytes);
      ctrl[131] = b == 2 ? 0u : b + 1;
      __threadfence();
    }
  }
}

template <int WORLD, int VPT>
static int launch_fused_ar_ll_v(const LLArParams& p, int threads, cudaStream_t st) {
  static std::atomic<int> per_sm_of[17];  // co-residency bound of this instantiation, by block size / 32
  int per_sm = per_sm_of[threads / 32].load(std::memory_order_relaxed);
  if (per_sm == 0) {
    int n = 0;
    NVLLM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, allreduce_add_rmsnorm_ll_kernel<WORLD, VPT>, threads, 0));
    per_sm = n < 1 ? 1 : n;
    per_sm_of[threads / 32].store(per_sm, std::memory_order_relaxed);
  }
  if (per_sm > 2) per_sm = 2;  // leaves room for the PDL successor's prologue; two rows per SM in flight are plenty
  int grid = p.num_tokens < per_sm * sm_count() ? p.num_tokens : per_sm * sm_count();
  NVLLM_CUDA(launch_pdl(allreduce_add_rmsnorm_ll_kernel<WORLD, VPT>, dim3(grid), dim3(threads), 0, st, p));
  return 0;
}
template <int WORLD>
static int launch_fused_ar_ll(const LLArParams& p, cudaStream_t st) {
  const int nvec = p.hidden / 8;
  const int vpt = nvec <= 128 ? 1 : (nvec <= 1024 ? 2 : 4);
  int threads = ((nvec + vpt - 1) / vpt + 31) / 32 * 32;
  if (threads < 32) threads = 32;
  if (vpt == 1) return launch_fused_ar_ll_v<WORLD, 1>(p, threads, st);
  if (vpt == 2) return launch_fused_ar_ll_v<WORLD, 2>(p, threads, st);
  return launch_fused_ar_ll_v<WORLD, 4>(p, threads, st);
}

__global__ void ll_fill_kernel(uint4* p, size_t n16) {
  const uint4 e = make_uint4(kLLEmpty, kLLEmpty, kLLEmpty, kLLEmpty);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n16; i += static_cast<size_t>(gridDim.x) * blockDim.x) p[i] = e;
}

template <int WORLD>
static int launch_fused_ar(const FusedArParams& p, int rows_per, cudaStream_t st) {
  const int nvec = p.hidden / 8;
  const int vpt = nvec <= 128 ? 1 : (nvec <= 1024 ? 2 : 4);
  int threads = ((nvec + vpt - 1) / vpt + 31) / 32 * 32;
  if (threads < 32) threads = 32;
  int grid = rows_per < 4 * sm_count() ? rows_per : 4 * sm_count();
  if (grid < 1) grid = 1;
  if (vpt == 1) NVLLM_CUDA(launch_pdl(allreduce_add_rmsnorm_kernel<WORLD, 1>, dim3(grid), dim3(threads), 0, st, p));
  else if (vpt == 2) NVLLM_CUDA(launch_pdl(allreduce_add_rmsnorm_kernel<WORLD, 2>, dim3(grid), dim3(threads), 0, st, p));
  else NVLLM_CUDA(launch_pdl(allreduce_add_rmsnorm_kernel<WORLD, 4>, dim3(grid), dim3(threads), 0, st, p));
  return 0;
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
  const size_t ll_cap = (static_cast<size_t>(knobs().ar_ll_max_bytes) + 255) & ~static_cast<size_t>(255);
  const size_t user_off = kSymCtrlBytes + 6 * ll_cap;
  const size_t total = user_off + ((bytes + 255) & ~static_cast<size_t>(255));
  void* base = nullptr;
  cudaIpcMemHandle_t* d_all = nullptr;
  cudaStream_t st = nullptr;
  bool stream_busy = false;  // a collective that never finished may still write d_all: leak it rather than free it
  // Every early exit below releases what was taken so far (base, d_all and the opened peer mappings leaked).
  auto fail = [&](int rc) {
    for (int r = 0; r < comm->world; ++r) {
      if (r != comm->rank && comm->sym_peer[r]) cudaIpcCloseMemHandle(comm->sym_peer[r]);
      comm->sym_peer[r] = nullptr;
    }
    if (!stream_busy) {
      if (st) cudaStreamDestroy(st);
      if (d_all) cudaFree(d_all);
      if (base) cudaFree(base);
    }
    return rc;
  };
#define NVLLM_CUDA_OR_FAIL(expr)                                       \
  do {                                                                 \
    cudaError_t _e = (expr);                                           \
    if (_e != cudaSuccess) return fail(::nvllm::cuda_status(_e, #expr)); \
  } while (0)
  NVLLM_CUDA_OR_FAIL(cudaMalloc(&base, total));
  NVLLM_CUDA_OR_FAIL(cudaMemset(base, 0, total));
  if (ll_cap > 0) {  // "nothing has arrived" everywhere, before any peer can learn this region's handle
    ll_fill_kernel<<<256, 256>>>(reinterpret_cast<uint4*>(static_cast<uint8_t*>(base) + kSymCtrlBytes), 6 * ll_cap / 16);
    NVLLM_CUDA_OR_FAIL(cudaGetLastError());
    NVLLM_CUDA_OR_FAIL(cudaDeviceSynchronize());
  }
  cudaIpcMemHandle_t mine;
  NVLLM_CUDA_OR_FAIL(cudaIpcGetMemHandle(&mine, base));
  NVLLM_CUDA_OR_FAIL(cudaMalloc(reinterpret_cast<void**>(&d_all), sizeof(mine) * (comm->world + 1)));
  NVLLM_CUDA_OR_FAIL(cudaMemcpy(d_all + comm->world, &mine, sizeof(mine), cudaMemcpyHostToDevice));
  NVLLM_CUDA_OR_FAIL(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  if (int rc = nccl_status(nccl().AllGather(d_all + comm->world, d_all, sizeof(mine), ncclUint8, comm->comm, st), "ncclAllGather(ipc)"))
    return fail(rc);
  if (int rc = stream_wait_bounded(st, "comm_symm_alloc: all-gather of the IPC handles")) {
    stream_busy = rc == NVLLM_ERR_COMM;
    return fail(rc);
  }
  cudaIpcMemHandle_t all[8];
  NVLLM_CUDA_OR_FAIL(cudaMemcpy(all, d_all, sizeof(mine) * comm->world, cudaMemcpyDeviceToHost));
  cudaStreamDestroy(st); st = nullptr;
  cudaFree(d_all); d_all = nullptr;
  for (int r = 0; r < comm->world; ++r) {
    if (r == comm->rank) continue;
    void* pp = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&pp, all[r], cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      set_error("cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
      return fail(NVLLM_ERR_COMM);
    }
    comm->sym_peer[r] = static_cast<uint8_t*>(pp);
  }
  if (comm->h_err == nullptr) {
    NVLLM_CUDA_OR_FAIL(cudaHostAlloc(reinterpret_cast<void**>(&comm->h_err), 16 * sizeof(int), cudaHostAllocMapped));
    memset(comm->h_err, 0, 16 * sizeof(int));
    NVLLM_CUDA_OR_FAIL(cudaHostGetDevicePointer(reinterpret_cast<void**>(&comm->d_err), comm->h_err, 0));
  }
#undef NVLLM_CUDA_OR_FAIL
  comm->sym_peer[comm->rank] = static_cast<uint8_t*>(base);
  comm->sym_local = static_cast<uint8_t*>(base);
  g_live_symm_regions.fetch_add(1, std::memory_order_relaxed);  // from here on kernels launch without PDL (common.cuh)
  comm->sym_bytes = total;
  comm->ll_cap = ll_cap;
  comm->user_off = user_off;
  *local_base = comm->sym_local + user_off;
  return 0;
}

// 1 when a call of this size takes the low-latency push protocol (and may therefore be fed by nvllm_gemm_bf16_push).
int nvllm_comm_ll_applies(const nvllm_comm* comm, int32_t num_tokens, int32_t hidden) {
  using namespace nvllm;
  if (!comm || !comm->sym_local || !knobs().ar_ll || num_tokens <= 0 || hidden <= 0) return 0;
  if (comm->world != 2 && comm->world != 4 && comm->world != 8) return 0;
  const int rows_per = (num_tokens + comm->world - 1) / comm->world;
  return static_cast<size_t>(comm->world) * rows_per * hidden * 2 <= comm->ll_cap ? 1 : 0;
}

// y = x . w^T for a row-parallel projection under TP (o_proj / down_proj, linear.rs:228-239) whose output feeds
// nvllm_allreduce_add_rmsnorm_pushed: rows this rank reduces itself land in y, every other row goes from the GEMM's
// epilogue straight into the owner's low-latency stage buffer (m == the all-reduce's num_tokens, n == hidden == ldy).
int nvllm_gemm_bf16_push(nvllm_comm* comm, void* y, const void* x, const void* w, int32_t m, int32_t n, int32_t k,
                         int64_t ldx, int64_t ldw, int64_t ldy, void* stream) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(comm && comm->sym_local, NVLLM_ERR_STATE, "gemm_bf16_push: no symmetric region");
  NVLLM_CHECK_ARG(m > 0 && m <= 512 && ldy == n && nvllm_comm_ll_applies(comm, m, n), NVLLM_ERR_ARG,
                  "gemm_bf16_push: m=%d n=%d ldy=%lld does not take the low-latency all-reduce", m, n, (long long)ldy);
  RowPush ps;
  for (int r = 0; r < 8; ++r) ps.peer[r] = r < comm->world ? comm->sym_peer[r] : nullptr;
  ps.ctrl = reinterpret_cast<const uint32_t*>(comm->sym_local);
  ps.ll_off = kSymCtrlBytes; ps.ll_cap = comm->ll_cap;
  ps.rank = comm->rank; ps.R = (m + comm->world - 1) / comm->world; ps.hidden = n;
  return gemm_bf16_impl(y, x, w, m, n, k, ldx, ldw, ldy, NVLLM_BF16, stream, ps);
}

static int allreduce_add_rmsnorm_any(nvllm_comm* comm, void* out, void* x, void* residual, const void* weight,
                                     int32_t num_tokens, int32_t hidden, float eps, int32_t dtype, void* stream, bool pushed);

int nvllm_allreduce_add_rmsnorm(nvllm_comm* comm, void* out, void* x, void* residual, const void* weight,
                                int32_t num_tokens, int32_t hidden, float eps, int32_t dtype, void* stream) {
  return allreduce_add_rmsnorm_any(comm, out, x, residual, weight, num_tokens, hidden, eps, dtype, stream, false);
}

// The same, after nvllm_gemm_bf16_push produced x: the foreign rows are already on their way to their owners.
int nvllm_allreduce_add_rmsnorm_pushed(nvllm_comm* comm, void* out, void* x, void* residual, const void* weight,
                                       int32_t num_tokens, int32_t hidden, float eps, int32_t dtype, void* stream) {
  NVLLM_CHECK_ARG(nvllm_comm_ll_applies(comm, num_tokens, hidden), NVLLM_ERR_ARG,
                  "allreduce_add_rmsnorm_pushed: %d x %d does not take the low-latency protocol", num_tokens, hidden);
  return allreduce_add_rmsnorm_any(comm, out, x, residual, weight, num_tokens, hidden, eps, dtype, stream, true);
}

static int allreduce_add_rmsnorm_any(nvllm_comm* comm, void* out, void* x, void* residual, const void* weight,
                                     int32_t num_tokens, int32_t hidden, float eps, int32_t dtype, void* stream, bool pushed) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(comm && comm->sym_local, NVLLM_ERR_STATE, "allreduce_add_rmsnorm: no symmetric region (nvllm_comm_symm_alloc)");
  NVLLM_CHECK_ARG(dtype == NVLLM_BF16, NVLLM_ERR_DTYPE, "allreduce_add_rmsnorm: dtype %d", dtype);
  NVLLM_CHECK_ARG(num_tokens >= 0 && hidden > 0 && hidden % 8 == 0 && hidden <= 16384, NVLLM_ERR_ARG, "allreduce_add_rmsnorm: sizes");
  if (num_tokens == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int rows_per = (num_tokens + comm->world - 1) / comm->world;
  const size_t staged = static_cast<size_t>(comm->world) * rows_per * hidden * 2;
  if (knobs().ar_ll && staged <= comm->ll_cap && (comm->world == 2 || comm->world == 4 || comm->world == 8)) {
    NVLLM_CHECK_ARG(out && x && residual && weight && ((reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(x) |
                                                         reinterpret_cast<uintptr_t>(residual)) & 15) == 0,
                    NVLLM_ERR_ARG, "allreduce_add_rmsnorm: out / x / residual must be 16-byte aligned");
    LLArParams q;
    for (int r = 0; r < 8; ++r) q.peer[r] = r < comm->world ? comm->sym_peer[r] : nullptr;
    q.self = comm->sym_local;
    q.x = static_cast<const uint8_t*>(x); q.res = static_cast<uint8_t*>(residual); q.out = static_cast<uint8_t*>(out);
    q.weight = static_cast<const bf16*>(weight);
    q.ll_off = kSymCtrlBytes; q.ll_cap = comm->ll_cap;
    q.rank = comm->rank; q.num_tokens = num_tokens; q.hidden = hidden; q.eps = eps; q.pushed = pushed ? 1 : 0;
    q.timeout_cycles = knobs().comm_timeout_ms * 2000000ll;
    q.err = comm->d_err;
    switch (comm->world) {
      case 2: return launch_fused_ar_ll<2>(q, st);
      case 4: return launch_fused_ar_ll<4>(q, st);
      default: return launch_fused_ar_ll<8>(q, st);
    }
  }
  auto off = [&](const void* ptr, size_t* o) {
    const uint8_t* b = static_cast<const uint8_t*>(ptr);
    if (b < comm->sym_local + comm->user_off || b + static_cast<size_t>(num_tokens) * hidden * 2 > comm->sym_local + comm->sym_bytes) return false;
    *o = static_cast<size_t>(b - comm->sym_local);
    return (*o & 15) == 0;
  };
  FusedArParams p;
  NVLLM_CHECK_ARG(off(out, &p.out_off) && off(x, &p.x_off) && off(residual, &p.res_off) && weight, NVLLM_ERR_ARG,
                  "allreduce_add_rmsnorm: out / x / residual must be 16-byte aligned buffers inside the symmetric region");
  for (int r = 0; r < 8; ++r) p.peer[r] = r < comm->world ? comm->sym_peer[r] : nullptr;
  p.weight = static_cast<const bf16*>(weight);
  p.self = comm->sym_local;
  p.rank = comm->rank; p.num_tokens = num_tokens; p.hidden = hidden; p.eps = eps;
  p.timeout_cycles = knobs().comm_timeout_ms * 2000000ll;  // clock64 ticks at <= 2 GHz
  p.err = comm->d_err;
  switch (comm->world) {
    case 2: return launch_fused_ar<2>(p, rows_per, st);
    case 4: return launch_fused_ar<4>(p, rows_per, st);
    case 8: return launch_fused_ar<8>(p, rows_per, st);
    default: break;
  }
  NVLLM_CHECK_ARG(false, NVLLM_ERR_ARG, "allreduce_add_rmsnorm: world %d (2, 4 or 8)", comm->world);
}

#ifdef NVLLM_TIMELINE
int nvllmdbg_ar_timeline(unsigned long long* d_buf, unsigned int cap) {
  unsigned int zero = 0;
  if (cudaMemcpyToSymbol(nvllm::g_ar_tl_buf, &d_buf, sizeof(d_buf)) != cudaSuccess) return 1;
  if (cudaMemcpyToSymbol(nvllm::g_ar_tl_cap, &cap, sizeof(cap)) != cudaSuccess) return 1;
  if (cudaMemcpyToSymbol(nvllm::g_ar_tl_cursor, &zero, sizeof(zero)) != cudaSuccess) return 1;
  return 0;
}
#endif

// 0, or NVLLM_ERR_COMM once a fused kernel of this communicator has given up waiting for a peer (sticky).
int nvllm_comm_error(const nvllm_comm* comm) {
  if (comm == nullptr || comm->h_err == nullptr) return 0;
  volatile int* e = comm->h_err;
  if (e[0] == 0) return 0;
  static const char* const kPhase[] = {"?", "flag protocol: a peer's ready flag", "flag protocol: a peer's done flag",
                                       "low-latency reduce: staged partial rows", "low-latency gather: normed rows"};
  const int ph = e[1] >= 1 && e[1] <= 4 ? e[1] : 0;
  nvllm::set_error("fused TP kernel on rank %d of %d gave up after NVLLM_COMM_TIMEOUT_MS waiting for %s (%s %d, CTA %d, %s %d, "
                   "missing source ranks 0x%x); the communicator is dead",
                   comm->rank, comm->world, kPhase[ph], ph <= 2 ? "peer" : "row", e[2], e[3], ph <= 2 ? "epoch" : "buffer pair",
                   e[4], e[5]);
  return NVLLM_ERR_COMM;
}

int nvllm_comm_destroy(nvllm_comm* comm) {
  if (!comm) return 0;
  cudaDeviceSynchronize();  // no kernel of this rank may still be reading peer memory
  for (int r = 0; r < comm->world; ++r)
    if (r != comm->rank && comm->sym_peer[r]) cudaIpcCloseMemHandle(comm->sym_peer[r]);
  // Freeing exported memory while another process still has it mapped is undefined behaviour (observed as a hang at
  // process teardown): rendezvous once more — every rank has closed its peer mappings before anyone frees its region.
  // The rendezvous is an NCCL all-reduce of one scratch word (never the live control block) with a host-side deadline.
  // A communicator that already reported NVLLM_ERR_COMM has a peer that is gone or stuck: no rendezvous can complete, so
  // none is tried, and — like after a rendezvous that timed out — the region is LEAKED (a peer may still have it mapped)
  // and the NCCL communicator is aborted instead of destroyed (ncclCommDestroy waits for the peers too).
  if (comm->sym_local) nvllm::g_live_symm_regions.fetch_sub(1, std::memory_order_relaxed);
  bool peers_ok = !(comm->h_err && *reinterpret_cast<volatile int*>(comm->h_err) != 0);
  if (peers_ok && comm->sym_local && comm->comm && nvllm::nccl().ok && comm->world > 1) {
    cudaStream_t st = nullptr;
    float* word = nullptr;
    if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) == cudaSuccess && cudaMalloc(reinterpret_cast<void**>(&word), 256) == cudaSuccess &&
        cudaMemset(word, 0, 256) == cudaSuccess &&
        nvllm::nccl().AllReduce(word, word, 1, nvllm::ncclFloat32, nvllm::ncclSum, comm->comm, st) == nvllm::ncclSuccess) {
      peers_ok = nvllm::stream_wait_bounded(st, "comm_destroy: rendezvous before freeing the symmetric region") == 0;
    } else {
      peers_ok = false;
    }
    if (peers_ok) {  // a kernel that never finished may still touch them
      if (word) cudaFree(word);
      if (st) cudaStreamDestroy(st);
    }
  }
  if (comm->sym_local && (peers_ok || comm->world == 1)) cudaFree(comm->sym_local);
  if (comm->h_err && peers_ok) cudaFreeHost(comm->h_err);
  if (comm->comm && nvllm::nccl().ok) {
    if (peers_ok) nvllm::nccl().CommDestroy(comm->comm);
    else if (nvllm::nccl().CommAbort) nvllm::nccl().CommAbort(comm->comm);
  }
  const int rc = peers_ok ? 0 : NVLLM_ERR_COMM;
  delete comm;
  return rc;
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
