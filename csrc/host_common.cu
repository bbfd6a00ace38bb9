// Error reporting, launch accounting and the TMA descriptor encoder for libomldm_kernels.so.
#include "host_common.h"

#include <cudaTypedefs.h>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>

namespace omldm {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

int set_error(const char* what, cudaError_t e) {
  snprintf(g_err, sizeof(g_err), "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
  return -1;
}
int set_error_msg(const char* msg) {
  snprintf(g_err, sizeof(g_err), "%s", msg);
  return -1;
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
int check_launch(const char* kernel_name) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(kernel_name, e);
  count_launch(1);
  return 0;
}

static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
static std::once_flag g_encode_once;

int tensormap_encode(CUtensorMap* tm, CUtensorMapDataType dtype, uint32_t rank, void* base,
                     const cuuint64_t* dims, const cuuint64_t* strides_bytes, const cuuint32_t* box,
                     const cuuint32_t* elem_strides, CUtensorMapSwizzle swizzle) {
  std::call_once(g_encode_once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess)
      g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
  });
  if (!g_encode) return set_error_msg("cuTensorMapEncodeTiled is not available from the driver");
  CUresult r = g_encode(tm, dtype, rank, base, dims, strides_bytes, box, elem_strides,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_err, sizeof(g_err), "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
    return -1;
  }
  return 0;
}

}  // namespace omldm

extern "C" const char* omldm_last_error() { return omldm::g_err; }
extern "C" long long omldm_launch_count() { return omldm::g_launches.load(); }
extern "C" void omldm_add_launch_count(long long n) { omldm::g_launches.fetch_add(n); }
extern "C" void omldm_reset_launch_count() { omldm::g_launches.store(0); }
extern "C" int omldm_abi_version() { return 1; }
