// tc_host.cu — host side of the TMA plumbing: cuTensorMapEncodeTiled is fetched from the driver at run
// time (cudaGetDriverEntryPoint), so libnvllm.so has no link-time dependency on libcuda and loads on a
// machine without a GPU (the CPU-side ABI tests rely on that).
#include "tc.cuh"

namespace nvllm {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess || p == nullptr) {
    cudaGetLastError();
    return nullptr;
  }
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

int make_tmap_bf16_2d(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld,
                      int box_rows, int box_cols) {
  EncodeTiledFn fn = encode_fn();
  NVLLM_CHECK_ARG(fn != nullptr, NVLLM_ERR_STATE, "cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
  NVLLM_CHECK_ARG(aligned16(base) && (ld * 2) % 16 == 0, NVLLM_ERR_ALIGN,
                  "TMA: base must be 16-byte aligned and the row pitch a multiple of 16 bytes");
  NVLLM_CHECK_ARG(box_rows >= 1 && box_rows <= 256 && box_cols * 2 <= 128, NVLLM_ERR_ARG,
                  "TMA: box %d x %d", box_rows, box_cols);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_cols), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  NVLLM_CHECK_ARG(r == CUDA_SUCCESS, NVLLM_ERR_ARG,
                  "cuTensorMapEncodeTiled failed (%d) rows=%lld cols=%lld ld=%lld box=%dx%d", (int)r,
                  (long long)rows, (long long)cols, (long long)ld, box_rows, box_cols);
  return 0;
}

// General rank-N (<= 5) bf16 map: dims/box innermost first, strides in BYTES for dims 1..rank-1.
int make_tmap_bf16_nd(CUtensorMap* map, const void* base, int rank, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box) {
  EncodeTiledFn fn = encode_fn();
  NVLLM_CHECK_ARG(fn != nullptr, NVLLM_ERR_STATE, "cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
  NVLLM_CHECK_ARG(rank >= 2 && rank <= 5 && aligned16(base), NVLLM_ERR_ALIGN, "TMA nd: rank %d / base alignment", rank);
  cuuint64_t d[5], st[4];
  cuuint32_t b[5], es[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) {
    NVLLM_CHECK_ARG(strides_bytes[i] % 16 == 0, NVLLM_ERR_ALIGN, "TMA nd: stride %d = %llu bytes", i,
                    (unsigned long long)strides_bytes[i]);
    st[i] = strides_bytes[i];
  }
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), d, st, b, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  NVLLM_CHECK_ARG(r == CUDA_SUCCESS, NVLLM_ERR_ARG, "cuTensorMapEncodeTiled (rank %d) failed (%d)", rank, (int)r);
  return 0;
}

}  // namespace nvllm
