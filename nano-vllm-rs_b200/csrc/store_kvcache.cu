// store_kvcache.cu — K1: scatter this step's K/V rows into the paged cache through slot_mapping.
// Replaces Attention::store_kv_cache (reference src/layers/attention.rs:150-174), which loops over
// tokens on the host with one blocking scalar read and two functional copies of the cache per token.
//
// HBM-bound, pure data movement (bit-exact).  One 16-byte vector per thread; a (token, kv-head) row
// of D bf16 is D/8 consecutive lanes, so loads are fully coalesced and stores are 256-byte segments
// (contiguous 2 KiB per token in the reference layout).  Grid is sized in waves of the SM count.
#include "common.cuh"

namespace nvllm {

__global__ void __launch_bounds__(256)
store_kvcache_kernel(const uint4* __restrict__ k, const uint4* __restrict__ v,
                     uint4* __restrict__ k_cache, uint4* __restrict__ v_cache,
                     const int32_t* __restrict__ slot_mapping, int num_tokens,
                     int64_t kv_token_stride_v,  // in 16-byte vectors
                     int block_size, int hkv, int vec_per_head, int64_t block_stride_v,
                     int64_t token_stride_v, int64_t head_stride_v) {
  pdl_launch_dependents();
  pdl_wait();
  const int vec_per_token = hkv * vec_per_head;
  const int64_t total = static_cast<int64_t>(num_tokens) * vec_per_token;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int tok = static_cast<int>(i / vec_per_token);
    const int rem = static_cast<int>(i - static_cast<int64_t>(tok) * vec_per_token);
    const int slot = __ldg(slot_mapping + tok);
    if (slot < 0) continue;
    const int head = rem / vec_per_head;
    const int c = rem - head * vec_per_head;
    const int blk = slot / block_size;
    const int off = slot - blk * block_size;
    const int64_t src = tok * kv_token_stride_v + rem;
    const int64_t dst = blk * block_stride_v + off * token_stride_v + head * head_stride_v + c;
    const uint4 kv = ld_nc_16(k + src);
    const uint4 vv = ld_nc_16(v + src);
    k_cache[dst] = kv;
    v_cache[dst] = vv;
  }
}

int check_layout(const nvllm_kv_layout* L) {
  NVLLM_CHECK_ARG(L != nullptr, NVLLM_ERR_ARG, "kv layout is NULL");
  NVLLM_CHECK_ARG(L->block_size > 0 && L->num_kv_heads > 0, NVLLM_ERR_ARG,
                  "kv layout: block_size=%d num_kv_heads=%d", L->block_size, L->num_kv_heads);
  NVLLM_CHECK_ARG(L->head_dim == 64 || L->head_dim == 128, NVLLM_ERR_ARG,
                  "kv layout: head_dim %d not in {64,128}", L->head_dim);
  NVLLM_CHECK_ARG(L->block_stride % 8 == 0 && L->token_stride % 8 == 0 && L->head_stride % 8 == 0,
                  NVLLM_ERR_ALIGN, "kv layout: strides must be multiples of 8 elements (16 bytes)");
  NVLLM_CHECK_ARG(L->token_stride >= L->head_dim && L->head_stride >= L->head_dim, NVLLM_ERR_ARG,
                  "kv layout: token/head stride smaller than head_dim");
  return 0;
}

}  // namespace nvllm

extern "C" int nvllm_store_kvcache(const void* k, const void* v, void* k_cache, void* v_cache,
                                   const int32_t* slot_mapping, int32_t num_tokens,
                                   int64_t kv_token_stride, const nvllm_kv_layout* layout,
                                   int32_t dtype, void* stream) {
  using namespace nvllm;
  NVLLM_CHECK_ARG(dtype == NVLLM_BF16, NVLLM_ERR_DTYPE, "store_kvcache: dtype %d (bf16 only)", dtype);
  if (int rc = check_layout(layout)) return rc;
  NVLLM_CHECK_ARG(num_tokens >= 0, NVLLM_ERR_ARG, "store_kvcache: num_tokens %d", num_tokens);
  if (num_tokens == 0) return 0;
  NVLLM_CHECK_ARG(k && v && k_cache && v_cache && slot_mapping, NVLLM_ERR_ARG,
                  "store_kvcache: NULL pointer");
  NVLLM_CHECK_ARG(aligned16(k) && aligned16(v) && aligned16(k_cache) && aligned16(v_cache),
                  NVLLM_ERR_ALIGN, "store_kvcache: pointers must be 16-byte aligned");
  NVLLM_CHECK_ARG(kv_token_stride % 8 == 0 &&
                      kv_token_stride >= static_cast<int64_t>(layout->num_kv_heads) * layout->head_dim,
                  NVLLM_ERR_ALIGN, "store_kvcache: kv_token_stride %lld", (long long)kv_token_stride);
  const int vec_per_head = layout->head_dim / 8;
  const int64_t total = static_cast<int64_t>(num_tokens) * layout->num_kv_heads * vec_per_head;
  const int threads = 256;
  int64_t blocks = (total + threads - 1) / threads;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 8;  // 8 resident CTAs of 256 threads / SM
  if (blocks > cap) blocks = cap;
  NVLLM_CUDA(launch_pdl(store_kvcache_kernel, dim3(static_cast<unsigned>(blocks)), dim3(threads), 0,
                        static_cast<cudaStream_t>(stream), static_cast<const uint4*>(k), static_cast<const uint4*>(v),
                        static_cast<uint4*>(k_cache), static_cast<uint4*>(v_cache), slot_mapping, num_tokens,
                        kv_token_stride / 8, layout->block_size, layout->num_kv_heads, vec_per_head,
                        layout->block_stride / 8, layout->token_stride / 8, layout->head_stride / 8));
  NVLLM_LAUNCH_CHECK("store_kvcache_kernel");
  return 0;
}
