// Ingest kernels: holdout sampling, sparse->dense staging, raw->staged row conversion.
//
// Holdout rule being reproduced (operators/spoke/FlinkSpoke.scala:92-107): of every 10 training
// records, numbers 8 and 9 go to a bounded test set (default 256 rows); when the set is full the
// evicted (oldest) point is trained on instead.  Because the rule depends only on the running
// record count, it is a pure index map: holdout #h lives in ring slot h % cap and evicts holdout
// #(h - cap).  The kernel swaps rows in place, so after it runs the batch rows ARE the training
// sequence in the original order (an empty row, bias column 0, where nothing was evicted yet).
#include <cuda_fp16.h>
#include "ptx.cuh"
#include "host_common.h"
#include "omldm_kernels.h"

namespace omldm {

__host__ __device__ inline long long holdouts_before(long long c) {
  const long long r = c % 10;
  return 2 * (c / 10) + (r > 8 ? r - 8 : 0);
}

// One CTA per ring slot.  The holdouts of this batch that map to the slot form a chain
// h, h+cap, h+2cap, ... (batch rows i_0 < i_1 < ... < i_m).  Sequential semantics: every new
// holdout evicts the previous occupant, which is trained on in its place.  In place that is a
// shift by one along the chain: ring <- row i_m, row i_j <- row i_{j-1}, row i_0 <- old ring
// content (or an empty row while the ring is still filling).  Each thread owns a fixed set of
// columns, walks the chain from the tail and so never needs a barrier.
__global__ void __launch_bounds__(256)
holdout_swap_kernel(float* __restrict__ Z, int ld, int B, int d, float* __restrict__ ring, int cap,
                    long long count0_arg, const long long* __restrict__ count_dev) {
  const long long count0 = count_dev ? *count_dev : count0_arg;
  const int slot = blockIdx.x;
  if (slot >= cap) return;
  const long long h0 = holdouts_before(count0), h1 = holdouts_before(count0 + B);
  const long long hf = h0 + ((slot - (h0 % cap)) % cap + cap) % cap;   // first chain element
  if (hf >= h1) return;
  const int m = static_cast<int>((h1 - 1 - hf) / cap);                // chain has m + 1 elements
  auto row_of = [&](int j) -> float* {
    const long long h = hf + static_cast<long long>(j) * cap;
    const long long c = (h / 2) * 10 + 8 + (h & 1);
    return Z + static_cast<size_t>(c - count0) * ld;
  };
  const bool evict = hf >= cap;
  float* rs = ring + static_cast<size_t>(slot) * ld;
  const int ld4 = ld >> 2;   // ld is a multiple of 32
  constexpr int MAXC = 16;
  for (int c4 = threadIdx.x; c4 < ld4; c4 += blockDim.x) {
    float4* r4 = reinterpret_cast<float4*>(rs) + c4;
    const float4 old = evict ? *r4 : make_float4(0.f, 0.f, 0.f, 0.f);
    if (m < MAXC) {
      // short chain: issue every load first (independent, pipelined), then the shifted stores
      float4 v[MAXC];
#pragma unroll
      for (int j = 0; j < MAXC; ++j)
        if (j <= m) v[j] = reinterpret_cast<const float4*>(row_of(j))[c4];
#pragma unroll
      for (int j = 0; j < MAXC; ++j) {
        if (j == m) *r4 = v[j];
        else if (j < m) reinterpret_cast<float4*>(row_of(j + 1))[c4] = v[j];
      }
    } else {
      // long chain: walk it from the tail in segments of MAXC rows; inside a segment every load is
      // issued before the shifted stores, so only ceil(m / MAXC) phases are serialised
      *r4 = reinterpret_cast<float4*>(row_of(m))[c4];
      for (int hi = m; hi >= 1; hi -= MAXC) {
        const int lo = hi - MAXC + 1 > 1 ? hi - MAXC + 1 : 1;   // rows lo..hi receive rows lo-1..hi-1
        float4 v[MAXC];
#pragma unroll
        for (int t = 0; t < MAXC; ++t)
          if (lo + t <= hi) v[t] = reinterpret_cast<const float4*>(row_of(lo + t - 1))[c4];
#pragma unroll
        for (int t = 0; t < MAXC; ++t)
          if (lo + t <= hi) reinterpret_cast<float4*>(row_of(lo + t))[c4] = v[t];
      }
    }
    reinterpret_cast<float4*>(row_of(0))[c4] = old;
  }
}

__global__ void advance_count_kernel(long long* c, long long delta) { *c += delta; }

__global__ void densify_kernel(const int* __restrict__ indptr, const int* __restrict__ indices,
                               const float* __restrict__ values, int d, const float* __restrict__ targets,
                               float* __restrict__ Z, int ld, float* __restrict__ sqnorm) {
  const int r = blockIdx.x;
  float* row = Z + static_cast<size_t>(r) * ld;
  float4* row4 = reinterpret_cast<float4*>(row);
  for (int c = threadIdx.x; c < ld / 4; c += blockDim.x) row4[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  __syncthreads();
  float acc = 0.f;
  const int b = indptr[r], e = indptr[r + 1];
  for (int p = b + threadIdx.x; p < e; p += blockDim.x) {
    const int c = indices[p];
    const float v = values[p];
    if (c >= 0 && c < d) {
      atomicAdd(&row[c], v);  // duplicate indices accumulate
      acc = fmaf(v, v, acc);
    }
  }
  __shared__ float s[32];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 1.f;  // bias column contributes 1
    for (int i = 0; i < (blockDim.x >> 5); ++i) t += s[i];
    if (sqnorm) sqnorm[r] = t;
    row[d] = 1.f;
    row[d + 1] = targets ? targets[r] : 0.f;
  }
}

__global__ void stage_rows_kernel(const float* __restrict__ raw, int raw_ld, int B, int d,
                                  float* __restrict__ Z, int ld) {
  const int r = blockIdx.x;
  const float* src = raw + static_cast<size_t>(r) * raw_ld;
  float* dst = Z + static_cast<size_t>(r) * ld;
  for (int c = threadIdx.x; c < ld; c += blockDim.x) {
    float v = 0.f;
    if (c < d) v = src[c];
    else if (c == d) v = 1.f;
    else if (c == d + 1) v = src[d];
    dst[c] = v;
  }
}

// Compact wire format -> staged rows.  A wire row is [x as fp16 x d | (pad to an even count) | target as
// fp32 in two halves]; pitch `raw_ld` halves.  fp16 keeps 11 significand bits, one more than the tf32
// products the Gram/score GEMMs use, so for features inside the fp16 range the learners see the same
// inputs - at half the PCIe bytes per record (the e2e path is H2D bound).
__global__ void stage_rows_f16_kernel(const __half* __restrict__ raw, int raw_ld, int B, int d,
                                      float* __restrict__ Z, int ld) {
  const int r = blockIdx.x;
  const __half* src = raw + static_cast<size_t>(r) * raw_ld;
  const __half2* src2 = reinterpret_cast<const __half2*>(src);
  float2* dst2 = reinterpret_cast<float2*>(Z + static_cast<size_t>(r) * ld);
  const int d_even = (d + 1) & ~1;
  const float target = *reinterpret_cast<const float*>(src + d_even);
  for (int c2 = threadIdx.x; c2 < (ld >> 1); c2 += blockDim.x) {
    const int c = 2 * c2;
    float2 v;
    if (c + 1 < d) {
      v = __half22float2(src2[c2]);
    } else {
      v.x = c < d ? __half2float(src[c]) : (c == d ? 1.f : (c == d + 1 ? target : 0.f));
      v.y = (c + 1 == d) ? 1.f : (c + 1 == d + 1 ? target : 0.f);
    }
    dst2[c2] = v;
  }
}

}  // namespace omldm

using namespace omldm;

extern "C" int omldm_holdout_swap(float* Z, int ld, int B, int d, float* ring, int cap, long long count0,
                                  void* stream) {
  if (B <= 0 || cap <= 0) return 0;
  if (holdouts_before(count0 + B) == holdouts_before(count0)) return 0;
  if (ld & 3) return set_error_msg("holdout_swap: ld must be a multiple of 4");
  holdout_swap_kernel<<<cap, 256, 0, static_cast<cudaStream_t>(stream)>>>(Z, ld, B, d, ring, cap, count0, nullptr);
  return check_launch("holdout_swap_kernel");
}

// Graph-replayable variant: the running record count lives in device memory and is advanced by
// `omldm_advance_count` (a separate node, so every block of the swap kernel reads the old value).
extern "C" int omldm_holdout_swap_dev(float* Z, int ld, int B, int d, float* ring, int cap,
                                      const long long* count_dev, void* stream) {
  if (B <= 0 || cap <= 0) return 0;
  if (ld & 3) return set_error_msg("holdout_swap: ld must be a multiple of 4");
  holdout_swap_kernel<<<cap, 256, 0, static_cast<cudaStream_t>(stream)>>>(Z, ld, B, d, ring, cap, 0, count_dev);
  return check_launch("holdout_swap_kernel");
}

extern "C" int omldm_advance_count(long long* count_dev, long long delta, void* stream) {
  advance_count_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(count_dev, delta);
  return check_launch("advance_count_kernel");
}

extern "C" int omldm_densify(const int* indptr, const int* indices, const float* values, int B, int d,
                             const float* targets, float* Z, int ld, float* sqnorm, void* stream) {
  if (B <= 0) return 0;
  if (ld & 3) return set_error_msg("densify: ld must be a multiple of 4");
  densify_kernel<<<B, 256, 0, static_cast<cudaStream_t>(stream)>>>(indptr, indices, values, d, targets, Z,
                                                                 ld, sqnorm);
  return check_launch("densify_kernel");
}

extern "C" int omldm_stage_rows(const float* raw, int raw_ld, int B, int d, float* Z, int ld, void* stream) {
  if (B <= 0) return 0;
  stage_rows_kernel<<<B, 128, 0, static_cast<cudaStream_t>(stream)>>>(raw, raw_ld, B, d, Z, ld);
  return check_launch("stage_rows_kernel");
}

extern "C" int omldm_stage_rows_f16(const void* raw, int raw_ld, int B, int d, float* Z, int ld, void* stream) {
  if (B <= 0) return 0;
  if ((raw_ld & 1) || raw_ld < ((d + 1) & ~1) + 2) return set_error_msg("stage_rows_f16: raw_ld must be even and >= even(d) + 2");
  if ((ld & 1) || ld < d + 2) return set_error_msg("stage_rows_f16: ld must be even and >= d + 2");
  if (reinterpret_cast<uintptr_t>(raw) & 3) return set_error_msg("stage_rows_f16: raw must be 4-byte aligned");
  stage_rows_f16_kernel<<<B, 128, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __half*>(raw), raw_ld, B, d,
                                                                        Z, ld);
  return check_launch("stage_rows_f16_kernel");
}
