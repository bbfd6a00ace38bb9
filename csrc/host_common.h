// Host-side helpers shared by every .cu translation unit of libomldm_kernels.so:
// error reporting through the C ABI and a lazily resolved cuTensorMapEncodeTiled (looked up with
// cudaGetDriverEntryPoint so the library does not link libcuda and can be built on a CPU box).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace omldm {

int set_error(const char* what, cudaError_t e);
int set_error_msg(const char* msg);
int check_launch(const char* kernel_name);
void count_launch(int n = 1);

int tensormap_encode(CUtensorMap* tm, CUtensorMapDataType dtype, uint32_t rank, void* base,
                     const cuuint64_t* dims, const cuuint64_t* strides_bytes, const cuuint32_t* box,
                     const cuuint32_t* elem_strides, CUtensorMapSwizzle swizzle);

}  // namespace omldm

// Tuning builds only (-DOMLDM_TIMELINE, scripts/build_variant.sh): where a kernel sits in the step.  TL_BEGIN in block
// (0, 0) (and the latest block start in word 2), TL_END in every block (max); globaltimer ns, monotone, so the values left behind belong to the last launch.
#if defined(OMLDM_TIMELINE) && defined(__CUDACC__)
#define OMLDM_TL_DECLARE(name)                                                                          \
  __device__ unsigned long long g_tl_##name[4];                                                          \
  extern "C" int omldm_tl_##name(unsigned long long* out) {                                              \
    return cudaMemcpyFromSymbol(out, g_tl_##name, sizeof(unsigned long long) * 4) == cudaSuccess ? 0 : -1; \
  }
__device__ __forceinline__ unsigned long long omldm_tl_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define OMLDM_TL_BEGIN(name) do { if (threadIdx.x == 0) { const unsigned long long t_ = omldm_tl_now(); if (blockIdx.x == 0 && blockIdx.y == 0) g_tl_##name[0] = t_; atomicMax(&g_tl_##name[2], t_); } } while (0)
#define OMLDM_TL_END(name) do { if (threadIdx.x == 0) atomicMax(&g_tl_##name[1], omldm_tl_now()); } while (0)
#else
#define OMLDM_TL_DECLARE(name)
#define OMLDM_TL_BEGIN(name) do { } while (0)
#define OMLDM_TL_END(name) do { } while (0)
#endif
