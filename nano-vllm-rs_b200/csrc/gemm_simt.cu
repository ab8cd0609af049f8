// gemm_simt.cu — plain CUDA-core y = x . w^T with the contract of nvllm_gemm_bf16.
// TEST INFRASTRUCTURE for the tcgen05 kernel (checks large shapes on the device, where the numpy oracle
// would take minutes).  Nothing in the engine calls it.
#include "common.cuh"

namespace nvllm {

template <typename OutT>
__global__ void __launch_bounds__(256)
gemm_simt_kernel(OutT* __restrict__ y, const bf16* __restrict__ x, const bf16* __restrict__ w, int M, int N,
                 int K, int64_t ldx, int64_t ldw, int64_t ldy) {
  const int n = blockIdx.x * 16 + (threadIdx.x & 15);
  const int m = blockIdx.y * 16 + (threadIdx.x >> 4);
  if (m >= M || n >= N) return;
  const uint4* xr = reinterpret_cast<const uint4*>(x + m * ldx);
  const uint4* wr = reinterpret_cast<const uint4*>(w + n * ldw);
  float acc = 0.f;
  for (int k = 0; k < K / 8; ++k) {
    const uint4 a = xr[k], b = wr[k];
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) acc += bf16_lo(aw[j]) * bf16_lo(bw[j]) + bf16_hi(aw[j]) * bf16_hi(bw[j]);
  }
  if constexpr (sizeof(OutT) == 2) y[m * ldy + n] = __float2bfloat16_rn(acc); else y[m * ldy + n] = acc;
}

}  // namespace nvllm

extern "C" int nvllm_gemm_bf16_simt(void* y, const void* x, const void* w, int32_t m, int32_t n, int32_t k,
                                    int64_t ldx, int64_t ldw, int64_t ldy, int32_t out_dtype, void* stream) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(m >= 0 && n > 0 && k > 0 && k % 8 == 0 && ldx % 8 == 0 && ldw % 8 == 0, NVLLM_ERR_ARG, "gemm_simt: sizes");
  if (m == 0) return 0;
  dim3 grid((n + 15) / 16, (m + 15) / 16);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (out_dtype == NVLLM_BF16)
    gemm_simt_kernel<bf16><<<grid, 256, 0, st>>>(static_cast<bf16*>(y), static_cast<const bf16*>(x), static_cast<const bf16*>(w), m, n, k, ldx, ldw, ldy);
  else
    gemm_simt_kernel<float><<<grid, 256, 0, st>>>(static_cast<float*>(y), static_cast<const bf16*>(x), static_cast<const bf16*>(w), m, n, k, ldx, ldw, ldy);
  NVLLM_LAUNCH_CHECK("gemm_simt_kernel");
  return 0;
}
