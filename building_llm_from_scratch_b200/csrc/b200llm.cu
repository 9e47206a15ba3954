// Library plumbing: error string, device cache, driver entry point for TMA descriptors,
// launch counter.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "common.cuh"

#include <cudaTypedefs.h>

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};
static std::mutex g_mu;
static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
static int g_sms[64] = {0};

int b200_set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

void b200_count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

static int ensure_driver() {
  if (g_encode) return 0;
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_encode) return 0;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e =
      cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || fn == nullptr)
    return b200_set_error(B200_ERR_INTERNAL, "cuTensorMapEncodeTiled entry point not found (%s)",
                          cudaGetErrorString(e));
  g_encode = (PFN_cuTensorMapEncodeTiled_v12000)fn;
  return 0;
}

// SMs the persistent kernels size their grids for. b200_set_sm_limit() lowers it below the device's count so that a
// few SMs stay free for the CTAs of a concurrent kernel — the NCCL all-reduce of a data-parallel step, which a grid
// of one persistent CTA per SM otherwise cannot overlap with (0 = no limit).
static std::atomic<int> g_sm_limit{0};

int b200_num_sms() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (g_sms[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    g_sms[dev] = n;
  }
  const int lim = g_sm_limit.load();
  return (lim > 0 && lim < g_sms[dev]) ? lim : g_sms[dev];
}

extern "C" int b200_set_sm_limit(int sms) {
  if (sms < 0 || (sms > 0 && sms < 8) || (sms & 1)) return b200_set_error(B200_ERR_ARG, "set_sm_limit: %d must be 0 or an even count >= 8", sms);
  g_sm_limit = sms;
  return 0;
}

int b200_make_tmap_2d_bf16(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols,
                           uint64_t ld, uint32_t box_rows, uint32_t box_cols) {
  int rc = ensure_driver();
  if (rc) return rc;
  // cuTensorMapEncodeTiled is a driver-API call: it needs the primary context bound to THIS thread.
  // Autograd worker threads may reach a GEMM before any runtime-API call has done that for them.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    B200_CHECK_CUDA(cudaFree(nullptr));
    ctx_bound = true;
  }
  if ((((uintptr_t)base) & 15u) != 0 || (ld * 2) % 16 != 0)
    return b200_set_error(B200_ERR_ARG, "TMA operand needs 16-B aligned base and row pitch");
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim,
                        gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return b200_set_error((int)r,
                          "cuTensorMapEncodeTiled failed (%d) rows=%llu cols=%llu ld=%llu box=%ux%u",
                          (int)r, (unsigned long long)rows, (unsigned long long)cols,
                          (unsigned long long)ld, box_rows, box_cols);
  return 0;
}

extern "C" {

int b200_version(void) { return B200LLM_VERSION; }

int b200_init(int device) {
  B200_CHECK_CUDA(cudaSetDevice(device));
  int major = 0, minor = 0;
  B200_CHECK_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  B200_CHECK_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
  if (major != 10)
    return b200_set_error(B200_ERR_UNSUPPORTED,
                          "b200llm needs an sm_100a device, found sm_%d%d (no fallback path)", major,
                          minor);
  int rc = ensure_driver();
  if (rc) return rc;
  (void)b200_num_sms();
  return 0;
}

const char* b200_last_error(void) { return g_err; }

uint64_t b200_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

}  // extern "C"
