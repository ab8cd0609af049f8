// api.cu — version / error plumbing of the C ABI (include/nvllm.h).
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace nvllm {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int cached = 0;
  if (cached > 0) return cached;
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
    return 148;
  cached = n;
  return n;
}

static Knobs g_knobs;
static std::once_flag g_knobs_once;

static void parse_knobs() {
  Knobs k;
  auto num = [](const char* name, int dflt) {
    const char* e = getenv(name);
    return (e && e[0]) ? atoi(e) : dflt;
  };
  k.pdl = num("NVLLM_PDL", 1) != 0;
  k.pdl_tp = num("NVLLM_PDL_TP", 0) != 0;
  k.gemm_bn = num("NVLLM_GEMM_BN", 0);
  k.gemm_2cta = num("NVLLM_GEMM_2CTA", 1);
  k.gemm_mcast = num("NVLLM_GEMM_MCAST", 0) == 1;
  k.gemm_stages = num("NVLLM_GEMM_STAGES", 0);
  k.sk_stages = num("NVLLM_GEMM_SK_STAGES", 0);
  if (const char* e = getenv("NVLLM_GEMM_SPLITK")) {
    int bn = 0, sk = 0;
    if (e[0] == '0' && e[1] == 0) k.splitk_mode = 0;
    else if (sscanf(e, "%d,%d", &bn, &sk) == 2) { k.splitk_mode = 2; k.splitk_bn = bn; k.splitk_sk = sk; }
  }
  k.decode_cfg = num("NVLLM_DECODE_CFG", 0);
  k.prefill_poly = num("NVLLM_PREFILL_POLY", 4);
  k.fuse_rope = num("NVLLM_FUSE_ROPE", 0) == 1;
  k.fuse_decode = num("NVLLM_FUSE_DECODE", 0);
  k.fused_tp = num("NVLLM_FUSED_TP", 1);
  k.profile = num("NVLLM_PROFILE", 0);
  k.ar_ll = num("NVLLM_AR_LL", 1);
  k.gemm_push = num("NVLLM_GEMM_PUSH", 0);
  k.decode_persist = num("NVLLM_DECODE_PERSIST", 0);
  if (const char* e = getenv("NVLLM_AR_LL_MAX_BYTES")) { const long long v = atoll(e); if (v >= 0 && v <= (1ll << 30)) k.ar_ll_max_bytes = v; }
  if (const char* e = getenv("NVLLM_COMM_TIMEOUT_MS")) { const long long v = atoll(e); if (v > 0) k.comm_timeout_ms = v; }
  g_knobs = k;
}

std::atomic<int> g_live_symm_regions{0};

const Knobs& knobs() {
  std::call_once(g_knobs_once, parse_knobs);
  return g_knobs;
}

}  // namespace nvllm

extern "C" {

int nvllm_version(void) { return NVLLM_VERSION; }

int nvllm_reload_env(void) {
  nvllm::knobs();        // make sure the one-time parse has happened, then overwrite (single caller: benchmark tools)
  nvllm::parse_knobs();
  return 0;
}

const char* nvllm_last_error(void) { return nvllm::g_err; }

int nvllm_device_ok(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    return 0;
  }
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
  return major == 10 ? 1 : 0;
}

}  // extern "C"
