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

OMLDM_TL_DECLARE(loss)

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

__device__ __forceinline__ float dot8(const float4& a, const float4& w0, const float4& w1, float acc) {
  const __half2* h = reinterpret_cast<const __half2*>(&a);
  float2 t;
  t = __half22float2(h[0]); acc = fmaf(t.x, w0.x, acc); acc = fmaf(t.y, w0.y, acc);
  t = __half22float2(h[1]); acc = fmaf(t.x, w0.z, acc); acc = fmaf(t.y, w0.w, acc);
  t = __half22float2(h[2]); acc = fmaf(t.x, w1.x, acc); acc = fmaf(t.y, w1.y, acc);
  t = __half22float2(h[3]); acc = fmaf(t.x, w1.z, acc); acc = fmaf(t.y, w1.w, acc);
  return acc;
}

// One warp per row, RI rows in flight per warp, 16-byte loads (8 halves): RI x (row bytes / 512) independent loads per
// lane are issued before any of them is consumed.  wt = [w_0..w_{d-1}, b, -1, -1] sits in shared memory as fp32; the
// row's dot product with it is the residual.
constexpr int RI = 4;
// `rows_per_range` > 0: the batch is swept as ceil(B / rows_per_range) ranges side by side, blockIdx.y = range, each in
// row order - the order in which the split-K CTAs of the Gram kernel read the same rows, so that whichever kernel is
// ahead pulls a row from HBM and the other finds it in L2 when the two run concurrently.
// Optional tail of the pass: the last block to finish writes one learning-curve sample (cumulative loss sum, records
// fitted) into the device ring - the work of ingest.cu::curve_sample_kernel without a launch of its own.
struct CurveTail {
  double* ring;            // [cap][2], null = no sample
  long long* counter;      // samples written so far
  unsigned int* done;      // blocks finished (self-resetting)
  int cap;
};

__global__ void __launch_bounds__(256)
sq_residual_f16s_kernel(const __half* __restrict__ Zh_all, int ldh, int B_all, int d, const float* __restrict__ wt,
                        double* __restrict__ stats, int rows_per_range, const CurveTail tail) {
  OMLDM_TL_BEGIN(loss);
  const int row0 = rows_per_range > 0 ? blockIdx.y * rows_per_range : 0;
  const int B = rows_per_range > 0 ? min(rows_per_range, B_all - row0) : B_all;
  const __half* __restrict__ Zh = Zh_all + static_cast<size_t>(row0) * ldh;
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
  for (int r = warp; r < B; r += RI * nwarps) {
    const float4* z[RI];
    float dsum[RI];
#pragma unroll
    for (int u = 0; u < RI; ++u) {
      const int ru = r + u * nwarps;
      z[u] = reinterpret_cast<const float4*>(Zh + static_cast<size_t>(ru < B ? ru : r) * ldh);
      dsum[u] = 0.f;
    }
    for (int c0 = 0; c0 < n8; c0 += 64) {                   // two 16-byte loads per row and lane per trip
      const int ca = c0 + lane, cb = c0 + 32 + lane;
      float4 va[RI], vb[RI];
#pragma unroll
      for (int u = 0; u < RI; ++u) {
        va[u] = ca < n8 ? ld_nc_v4(z[u] + ca) : make_float4(0.f, 0.f, 0.f, 0.f);
        vb[u] = cb < n8 ? ld_nc_v4(z[u] + cb) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      if (ca < n8) {
        const float4 w0 = *reinterpret_cast<const float4*>(s_w + 8 * ca), w1 = *reinterpret_cast<const float4*>(s_w + 8 * ca + 4);
#pragma unroll
        for (int u = 0; u < RI; ++u) dsum[u] = dot8(va[u], w0, w1, dsum[u]);
      }
      if (cb < n8) {
        const float4 w0 = *reinterpret_cast<const float4*>(s_w + 8 * cb), w1 = *reinterpret_cast<const float4*>(s_w + 8 * cb + 4);
#pragma unroll
        for (int u = 0; u < RI; ++u) dsum[u] = dot8(vb[u], w0, w1, dsum[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < RI; ++u) {
      const float du = warp_sum(dsum[u]);
      const int ru = r + u * nwarps;
      if (ru < B && __half2float(Zh[static_cast<size_t>(ru) * ldh + d]) != 0.f) { acc += static_cast<double>(du) * du; nv += 1; }
    }
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
    if (tail.ring) {
      __threadfence();                                         // this block's sums before its arrival
      if (atomicAdd(tail.done, 1u) == gridDim.x * gridDim.y - 1) {
        __threadfence();
        *tail.done = 0;
        const long long i = *tail.counter;
        double* slot = tail.ring + 2 * (i % tail.cap);
        slot[0] = __ldcg(&stats[0]);                           // the other blocks' atomics live in L2
        slot[1] = __ldcg(&stats[2]);
        *tail.counter = i + 1;
      }
    }
  }
  OMLDM_TL_END(loss);
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

extern "C" int omldm_sq_residual_f16s(const void* Zh, int ldh, int B, int d, const float* wt, double* stats,
                                      int rows_per_range, void* stream) {
  return omldm_sq_residual_f16s_tail(Zh, ldh, B, d, wt, stats, rows_per_range, 0, nullptr, 0, nullptr, nullptr, stream);
}

extern "C" int omldm_sq_residual_f16s_tail(const void* Zh, int ldh, int B, int d, const float* wt, double* stats,
                                           int rows_per_range, int max_blocks, double* curve_ring, int curve_cap,
                                           long long* curve_counter, unsigned int* done_word, void* stream) {
  if (B <= 0) return 0;
  if ((ldh & 7) || ldh < d + 3) return set_error_msg("sq_residual_f16s: the fp16 pitch must be a multiple of 8 and >= d + 3");
  if (reinterpret_cast<uintptr_t>(Zh) & 15) return set_error_msg("sq_residual_f16s: rows must be 16-byte aligned");
  if (curve_ring && (curve_cap <= 0 || !curve_counter || !done_word))
    return set_error_msg("sq_residual_f16s: the curve sample needs a ring, its counter and a completion word");
  const size_t smem = static_cast<size_t>(ldh) * sizeof(float);
  if (smem > 200 * 1024) return set_error_msg("sq_residual_f16s: row pitch too large for the smem weight copy");
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(sq_residual_f16s_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(sq_residual_f16s)", e);
  }
  int ranges = 1;
  if (rows_per_range > 0) ranges = (B + rows_per_range - 1) / rows_per_range;
  if (ranges < 1 || ranges > 1024) return set_error_msg("sq_residual_f16s: bad rows_per_range");
  const int per = ranges > 1 ? rows_per_range : B;
  int blocks = (per + 31) / 32;
  // three blocks fit an SM alone; beside a kernel that holds registers on every SM the caller asks for fewer, so that
  // the pass stays one wave (max_blocks = blocks in total, 0 = 3 per SM)
  const int total = max_blocks > 0 ? max_blocks : 148 * 3;
  const int cap = (total + ranges - 1) / ranges;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  CurveTail tail{curve_ring, curve_counter, done_word, curve_cap};
  sq_residual_f16s_kernel<<<dim3(blocks, ranges), 256, smem, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __half*>(Zh), ldh, B, d, wt, stats, ranges > 1 ? rows_per_range : 0, tail);
  return check_launch("sq_residual_f16s_kernel");
}
