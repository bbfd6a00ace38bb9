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
