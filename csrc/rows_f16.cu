// Kernels on fp16 staged record rows ("f16s" rows), the layout the Gram kernel (gram_f16.cu) consumes in place:
//     [ x_0 .. x_{d-1} | 1.0 bias/valid (0 = empty row) | y_hi | y_lo | zero pad ],   pitch = multiple of 64 halves
// The target travels as two halves (y = y_hi + y_lo, 22 significand bits).  A row in this layout is what the pinned
// host ring holds, what crosses PCIe and what the learner reads: no expansion pass on the ORR path.  The
// 2-in-10 holdout (operators/spoke/FlinkSpoke.scala:92-107) is a pure row move and uses holdout_swap_kernel on the
// same memory viewed as fp32 words.
//
//   expand_rows_f16s : f16s rows -> fp32 staged rows, for every learner that has no fp16 path (and for scoring the
//                      holdout rows); replaces DataPointParser.scala:16-55's boxed-list -> DenseVector conversion.
//   sq_residual_f16s : prequential squared error of ORR, sum over valid rows of (w.x + b - y)^2, read straight from
//                      the fp16 rows (predict-before-fit behind node.receiveTuple, FlinkSpoke.scala:101).
#include <cuda_fp16.h>
#include "ptx.cuh"
#include "host_common.h"
#include "omldm_kernels.h"

namespace omldm {

__global__ void __launch_bounds__(128)
expand_rows_f16s_kernel(const __half* __restrict__ Zh, int ldh, int B, int d, float* __restrict__ Z, int ld) {
  const int r = blockIdx.x;
  const __half* src = Zh + static_cast<size_t>(r) * ldh;
  const __half2* src2 = reinterpret_cast<const __half2*>(src);
  float2* dst2 = reinterpret_cast<float2*>(Z + static_cast<size_t>(r) * ld);
  const float y = __half2float(src[d + 1]) + __half2float(src[d + 2]);
  for (int c2 = threadIdx.x; c2 < (ld >> 1); c2 += blockDim.x) {
    const int c = 2 * c2;
    float2 v;
    if (c + 1 <= d) {                     // both columns are features or the bias flag: plain widening
      v = __half22float2(src2[c2]);
    } else {
      v.x = c <= d ? __half2float(src[c]) : (c == d + 1 ? y : 0.f);
      v.y = (c + 1 == d + 1) ? y : 0.f;
    }
    dst2[c2] = v;
  }
}

// One warp per row, two rows in flight, 16-byte loads (8 halves).  wt = [w_0..w_{d-1}, b, -1, -1] sits in shared memory
// as fp32; the row's dot product with it is the residual.
__global__ void __launch_bounds__(256)
sq_residual_f16s_kernel(const __half* __restrict__ Zh, int ldh, int B, int d, const float* __restrict__ wt,
                        double* __restrict__ stats) {
  extern __shared__ float s_w[];          // ldh floats, zero beyond d + 3
  const int n = d + 3;
  for (int c = threadIdx.x; c < ldh; c += blockDim.x) s_w[c] = c < n ? wt[c] : 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int n8 = (n + 7) >> 3;
  double acc = 0.0;
  int nv = 0;
  for (int r = warp; r < B; r += 2 * nwarps) {
    const int r2 = r + nwarps;
    const uint4* z0 = reinterpret_cast<const uint4*>(Zh + static_cast<size_t>(r) * ldh);
    const uint4* z1 = reinterpret_cast<const uint4*>(Zh + static_cast<size_t>(r2 < B ? r2 : r) * ldh);
    float d0 = 0.f, d1 = 0.f;
    for (int c = lane; c < n8; c += 32) {
      const uint4 a = __ldg(z0 + c), b = __ldg(z1 + c);
      const float4 w0 = *reinterpret_cast<const float4*>(s_w + 8 * c);
      const float4 w1 = *reinterpret_cast<const float4*>(s_w + 8 * c + 4);
      const __half2* ah = reinterpret_cast<const __half2*>(&a);
      const __half2* bh = reinterpret_cast<const __half2*>(&b);
      float2 t;
      t = __half22float2(ah[0]); d0 = fmaf(t.x, w0.x, d0); d0 = fmaf(t.y, w0.y, d0);
      t = __half22float2(ah[1]); d0 = fmaf(t.x, w0.z, d0); d0 = fmaf(t.y, w0.w, d0);
      t = __half22float2(ah[2]); d0 = fmaf(t.x, w1.x, d0); d0 = fmaf(t.y, w1.y, d0);
      t = __half22float2(ah[3]); d0 = fmaf(t.x, w1.z, d0); d0 = fmaf(t.y, w1.w, d0);
      t = __half22float2(bh[0]); d1 = fmaf(t.x, w0.x, d1); d1 = fmaf(t.y, w0.y, d1);
      t = __half22float2(bh[1]); d1 = fmaf(t.x, w0.z, d1); d1 = fmaf(t.y, w0.w, d1);
      t = __half22float2(bh[2]); d1 = fmaf(t.x, w1.x, d1); d1 = fmaf(t.y, w1.y, d1);
      t = __half22float2(bh[3]); d1 = fmaf(t.x, w1.z, d1); d1 = fmaf(t.y, w1.w, d1);
    }
    d0 = warp_sum(d0);
    d1 = warp_sum(d1);
    if (__half2float(Zh[static_cast<size_t>(r) * ldh + d]) != 0.f) { acc += static_cast<double>(d0) * d0; nv += 1; }
    if (r2 < B && __half2float(Zh[static_cast<size_t>(r2) * ldh + d]) != 0.f) { acc += static_cast<double>(d1) * d1; nv += 1; }
  }
  __shared__ double s_l[8];
  __shared__ int s_n[8];
  const int w = threadIdx.x >> 5;
  if (lane == 0) { s_l[w] = acc; s_n[w] = nv; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double l = 0; int c = 0;
    for (int i = 0; i < (blockDim.x >> 5); ++i) { l += s_l[i]; c += s_n[i]; }
    if (c) { atomicAdd(&stats[0], l); atomicAdd(&stats[2], static_cast<double>(c)); }
  }
}

}  // namespace omldm

using namespace omldm;

extern "C" int omldm_expand_rows_f16s(const void* Zh, int ldh, int B, int d, float* Z, int ld, void* stream) {
  if (B <= 0) return 0;
  if ((ldh & 7) || ldh < d + 3) return set_error_msg("expand_rows_f16s: the fp16 pitch must be a multiple of 8 and >= d + 3");
  if ((ld & 1) || ld < d + 2) return set_error_msg("expand_rows_f16s: ld must be even and >= d + 2");
  expand_rows_f16s_kernel<<<B, 128, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __half*>(Zh), ldh, B, d, Z, ld);
  return check_launch("expand_rows_f16s_kernel");
}

extern "C" int omldm_sq_residual_f16s(const void* Zh, int ldh, int B, int d, const float* wt, double* stats, void* stream) {
  if (B <= 0) return 0;
  if ((ldh & 7) || ldh < d + 3) return set_error_msg("sq_residual_f16s: the fp16 pitch must be a multiple of 8 and >= d + 3");
  if (reinterpret_cast<uintptr_t>(Zh) & 15) return set_error_msg("sq_residual_f16s: rows must be 16-byte aligned");
  const size_t smem = static_cast<size_t>(ldh) * sizeof(float);
  if (smem > 200 * 1024) return set_error_msg("sq_residual_f16s: row pitch too large for the smem weight copy");
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(sq_residual_f16s_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(sq_residual_f16s)", e);
  }
  int blocks = (B + 15) / 16;
  if (blocks > 148 * 4) blocks = 148 * 4;
  sq_residual_f16s_kernel<<<blocks, 256, smem, static_cast<cudaStream_t>(stream)>>>(static_cast<const __half*>(Zh), ldh, B, d, wt, stats);
  return check_launch("sq_residual_f16s_kernel");
}
