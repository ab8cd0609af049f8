// common.cuh — shared host/device helpers for the sm_100a kernels behind include/nvllm.h.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>
#include <mutex>

#include "../../include/nvllm.h"

// ------------------------------------------------------------------------------------------------
// Error convention of the C ABI: 0 OK, <0 argument error, >0 cudaError_t; message is thread-local.
// ------------------------------------------------------------------------------------------------
namespace nvllm {

void set_error(const char* fmt, ...);  // api.cu

inline int cuda_status(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_error("%s: %s", what, cudaGetErrorString(e));
  return static_cast<int>(e);
}

#define NVLLM_CHECK_ARG(cond, code, ...) \
  do {                                   \
    if (!(cond)) {                       \
      ::nvllm::set_error(__VA_ARGS__);   \
      return (code);                     \
    }                                    \
  } while (0)

#define NVLLM_CUDA(expr)                                         \
  do {                                                           \
    cudaError_t _e = (expr);                                     \
    if (_e != cudaSuccess) return ::nvllm::cuda_status(_e, #expr); \
  } while (0)

#define NVLLM_LAUNCH_CHECK(name) \
  do {                           \
    cudaError_t _e = cudaGetLastError(); \
    if (_e != cudaSuccess) return ::nvllm::cuda_status(_e, name); \
  } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int sm_count();  // api.cu (cached)

// The NVLLM_* environment knobs (A/B measurements), parsed ONCE per process — never on a dispatch
// path.  nvllm_reload_env() (benchmark tools that sweep knobs inside one process) re-parses them.
struct Knobs {
  int pdl = 1;              // NVLLM_PDL: 0 disables programmatic dependent launch
  int pdl_tp = 0;           // NVLLM_PDL_TP: 1 keeps programmatic dependent launch while a TP symmetric region is live (round-2
                            // behaviour: ~1 in 10 tiny-model runs at 8 ranks ends with rank 0 one call behind)
  int gemm_bn = 0;          // NVLLM_GEMM_BN: forced tile width (0 = heuristic)
  int gemm_2cta = 1;        // NVLLM_GEMM_2CTA: 0 never, 1 default (M > 512), 2 every M > 128
  int gemm_mcast = 0;       // NVLLM_GEMM_MCAST: 1 enables the x-multicast clusters
  int gemm_stages = 0;      // NVLLM_GEMM_STAGES: ring depth cap of single-wave launches (0 = deepest)
  int sk_stages = 0;        // NVLLM_GEMM_SK_STAGES: ring depth of the split-K kernel (0 = deepest)
  int splitk_mode = 1;      // NVLLM_GEMM_SPLITK: 0 never, 1 cost model, 2 forced (splitk_bn, splitk_sk)
  int splitk_bn = 0, splitk_sk = 0;
  int decode_cfg = 0;       // NVLLM_DECODE_CFG: alternative (stages, warps) instantiations of the paged decode kernel
  int prefill_poly = 4;     // NVLLM_PREFILL_POLY: exp2 split of the prefill softmax
  int fuse_rope = 0;        // NVLLM_FUSE_ROPE: qkv GEMM epilogue with q/k-norm + RoPE + KV store
  int fuse_decode = 0;      // NVLLM_FUSE_DECODE: 1 = q/k-norm + RoPE + KV store inside the paged decode attention (first version measured 9 % slower)
  int fused_tp = 1;         // NVLLM_FUSED_TP: 0 = ncclAllReduce + separate add_rmsnorm
  int decode_persist = 0;   // NVLLM_DECODE_PERSIST: 1 = persistent paged decode kernel (pipelines across item boundaries; measured +-1 %: no gain)
  int gemm_push = 0;        // NVLLM_GEMM_PUSH: 1 = o_proj / down_proj push foreign rows from their epilogue (16-byte NVLink packets: measured 2.8 % slower at N=2)
  int ar_ll = 1;            // NVLLM_AR_LL: 0 = always the flag-protocol fused all-reduce (no low-latency push variant)
  long long ar_ll_max_bytes = 8ll << 20;  // NVLLM_AR_LL_MAX_BYTES: largest message of the low-latency variant (6x this is reserved)
  int profile = 0;          // NVLLM_PROFILE: 1 = the runner times every launch of eager steps with CUDA events
  long long comm_timeout_ms = 20000;   // NVLLM_COMM_TIMEOUT_MS: how long a fused TP kernel (or a set-up / tear-down collective) waits for a peer
};
const Knobs& knobs();  // api.cu
extern std::atomic<int> g_live_symm_regions;  // api.cu; nvllm_comm_symm_alloc ++, nvllm_comm_destroy --
// Programmatic dependent launch is OFF while this process holds a tensor-parallel symmetric region, unless NVLLM_PDL_TP=1:
// with it on, the fused all-reduce kernels' per-call device state (buffer pair / epoch, handed from one call to the next
// through the control block) was observed one call stale on the rank whose kernels never spin (8 ranks, 4-CTA grids, about
// one run in ten; tools/tp_parity_soak.py); with it off, 50 consecutive runs were clean.  Root cause open.
inline bool pdl_enabled() {
  return knobs().pdl != 0 && (knobs().pdl_tp != 0 || g_live_symm_regions.load(std::memory_order_relaxed) == 0);
}

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) exactly once per kernel instantiation, safe under concurrent callers
// (the header promises re-entrancy).  Use inside a function that returns the C-ABI status code.
#define NVLLM_SMEM_ATTR_ONCE(kern, bytes)                                                                     \
  do {                                                                                                        \
    static std::once_flag _once;                                                                              \
    static cudaError_t _once_err = cudaSuccess;                                                               \
    std::call_once(_once, [&] { _once_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes); }); \
    NVLLM_CUDA(_once_err);                                                                                    \
  } while (0)

}  // namespace nvllm

// ------------------------------------------------------------------------------------------------
// Device helpers
// ------------------------------------------------------------------------------------------------
#ifdef __CUDACC__

namespace nvllm {

// ---- Programmatic dependent launch (PDL) -----------------------------------------------------------
// Every kernel of the step is launched with cudaLaunchAttributeProgrammaticStreamSerialization: it may start while
// its predecessor is still running.  Contract inside each kernel: pdl_launch_dependents() first (lets the successor
// be scheduled), then only predecessor-independent work (barrier init, TMEM alloc, descriptor / weight prefetch),
// then pdl_wait() before the first access to anything a previous kernel wrote or still reads.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;\n" ::: "memory"); }
// A word of per-communicator device state that ONE CTA of an earlier kernel rewrites and EVERY CTA of a later kernel reads
// (call counters, buffer-pair index, dead flag in the symmetric region's control block).  Read it past L1: with
// programmatic dependent launch a CTA of the later kernel can be resident on an SM (launched early, parked in
// griddepcontrol.wait) BEFORE a CTA of the earlier kernel on the same SM pulls the old value of the line into that SM's
// L1 — the later kernel's ordinary ld.global then hits the stale line.  Observed as a rank that re-used the previous
// call's buffer pair / epoch (the round-2 8-GPU hang); a gpu-scope load always goes to L2.
__device__ __forceinline__ uint32_t ld_state(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

typedef __nv_bfloat16 bf16;
typedef __nv_bfloat162 bf162;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  bf162 v = __floats2bfloat162_rn(lo, hi);  // .x = lo (low 16 bits)
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

// ---- Low-latency tensor-parallel push (comm.cu) -------------------------------------------------------------------
// kLLEmpty (two bf16 -0.0) marks "not written yet" in the low-latency buffers of the fused all-reduce: senders map that
// one word to +0.0, receivers poll the data itself.  RowPush lets the PRODUCING GEMM's epilogue (o_proj / down_proj under
// TP) store the rows of its partial sums that another rank reduces straight into that rank's stage buffer — phase A of
// allreduce_add_rmsnorm_ll_kernel then happens tile by tile under the GEMM's tail instead of after it.
constexpr uint32_t kLLEmpty = 0x80008000u;
struct RowPush {
  uint8_t* peer[8];      // symmetric regions of all ranks as mapped here
  const uint32_t* ctrl;  // this rank's control block (pair index at [131]); nullptr = plain epilogue
  size_t ll_off, ll_cap; // low-latency area inside a region, bytes of one buffer
  int rank, R, hidden;   // rows per owner = ceil(tokens / world); row length in elements
};
__device__ __forceinline__ uint4 ll_scrub(uint4 v) {
  v.x = v.x == kLLEmpty ? 0u : v.x;
  v.y = v.y == kLLEmpty ? 0u : v.y;
  v.z = v.z == kLLEmpty ? 0u : v.z;
  v.w = v.w == kLLEmpty ? 0u : v.w;
  return v;
}
__device__ __forceinline__ void st_sys_16(void* p, const uint4& v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// Row `row` of the partial-sum tensor: nullptr when this rank reduces it itself (store to y as usual), else where it goes
// in the owner's stage buffer of the coming all-reduce call ([source rank][row within the owner's slice][hidden]).
__device__ __forceinline__ bf16* push_row(const RowPush& ps, int row) {
  const int owner = row / ps.R;
  if (owner == ps.rank) return nullptr;
  uint8_t* base = ps.peer[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) base = (owner == i) ? ps.peer[i] : base;  // no run-time index into the parameter block
  const uint32_t b = ld_state(ps.ctrl + 131);
  return reinterpret_cast<bf16*>(base + ps.ll_off + static_cast<size_t>(b) * 2 * ps.ll_cap +
                                 (static_cast<size_t>(ps.rank) * ps.R + (row - owner * ps.R)) * ps.hidden * 2);
}

// ---- cp.async (LDGSTS) ---------------------------------------------------------------------------
// 16-byte global->shared copy; src_bytes < 16 zero-fills the remainder (0 = pure zero fill).
__device__ __forceinline__ void cp_async_16(uint32_t smem_addr, const void* gptr, int src_bytes) {
  asm volatile("cp.async.cg.shared.global.L2::128B [%0], [%1], 16, %2;\n" ::"r"(smem_addr),
               "l"(gptr), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

// ---- ldmatrix / movmatrix / mma.sync (warp-level, used by the HBM-bound decode kernel) -----------
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t smem_addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t smem_addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_addr));
}
__device__ __forceinline__ void ldmatrix_x2(uint32_t (&r)[2], uint32_t smem_addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];\n"
               : "=r"(r[0]), "=r"(r[1])
               : "r"(smem_addr));
}
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;\n" : "=r"(d) : "r"(a));
  return d;
}
// D(16x8,f32) += A(16x16,bf16,row) * B(16x8,bf16,col)
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0,
                                               uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};\n"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}

// 16-byte streaming load / store (no L1 allocation) for once-touched data.
__device__ __forceinline__ uint4 ld_nc_16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_na_16(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

}  // namespace nvllm

#endif  // __CUDACC__
