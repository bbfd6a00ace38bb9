// Ingest kernels: holdout sampling, sparse->dense staging, raw->staged row conversion.
//
// Holdout rule being reproduced (operators/spoke/FlinkSpoke.scala:92-107): of every 10 training
// records, numbers 8 and 9 go to a bounded test set (default 256 rows); when the set is full the
// evicted (oldest) point is trained on instead.  Because the rule depends only on the running
// record count, it is a pure index map: holdout #h lives in ring slot h % cap and evicts holdout
// #(h - cap).  The kernel swaps rows in place, so after it runs the batch rows ARE the training
// sequence in the original order (an empty row, bias column 0, where nothing was evicted yet).
#include "ptx.cuh"
#include "host_common.h"
#include "omldm_kernels.h"

namespace omldm {

__host__ __device__ inline long long holdouts_before(long long c) {
  const long long r = c % 10;
  return 2 * (c / 10) + (r > 8 ? r - 8 : 0);
}

__global__ void holdout_swap_kernel(float* __restrict__ Z, int ld, int B, int d, float* __restrict__ ring,
                                    int cap, long long count0) {
  const int warps_per_block = blockDim.x >> 5;
  const int slot = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (slot >= cap) return;
  const long long h0 = holdouts_before(count0), h1 = holdouts_before(count0 + B);
  long long h = h0 + ((slot - (h0 % cap)) % cap + cap) % cap;
  for (; h < h1; h += cap) {
    const long long c = (h / 2) * 10 + 8 + (h & 1);
    const int i = static_cast<int>(c - count0);
    float* row = Z + static_cast<size_t>(i) * ld;
    float* rs = ring + static_cast<size_t>(slot) * ld;
    const bool evict = h >= cap;
    for (int col = lane; col < ld; col += 32) {
      const float nv = row[col];
      const float ov = evict ? rs[col] : 0.f;
      rs[col] = nv;
      row[col] = ov;
    }
    __syncwarp();
  }
}

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

}  // namespace omldm

using namespace omldm;

extern "C" int omldm_holdout_swap(float* Z, int ld, int B, int d, float* ring, int cap, long long count0,
                                  void* stream) {
  if (B <= 0 || cap <= 0) return 0;
  if (holdouts_before(count0 + B) == holdouts_before(count0)) return 0;
  const int wpb = 8;
  holdout_swap_kernel<<<(cap + wpb - 1) / wpb, wpb * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      Z, ld, B, d, ring, cap, count0);
  return check_launch("holdout_swap_kernel");
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
