// Hoeffding-tree leaf statistics on the GPU (learner "HT", utils/parsers/requestStream/PipelineMap.scala:68;
// the reference never distributes it: FlinkSpoke.scala:204-206 forces SingleLearner).
//
// The tree itself is tiny and irregular, so split decisions stay on the host; what is hot is routing
// every record to its leaf and updating the per-leaf, per-class sufficient statistics of the Gaussian
// attribute observers (count, sum x, sum x^2 per feature).  One warp per record: the route is warp
// uniform, the feature updates are spread over the lanes.
#include "ptx.cuh"
#include "host_common.h"
#include "omldm_kernels.h"

namespace omldm {

__device__ __forceinline__ int ht_route(const float* __restrict__ zr, const int* __restrict__ feature,
                                        const float* __restrict__ threshold, const int* __restrict__ left,
                                        const int* __restrict__ right) {
  int node = 0;
  for (int depth = 0; depth < 64; ++depth) {
    const int f = feature[node];
    if (f < 0) break;
    node = zr[f] <= threshold[node] ? left[node] : right[node];
  }
  return node;
}

__device__ __forceinline__ int ht_class(float y, int C) {
  int c = static_cast<int>(rintf(y));
  if (C == 2 && c < 0) c = 0;          // {-1,+1} labels map to {0,1}
  return min(max(c, 0), C - 1);
}

__global__ void __launch_bounds__(256)
ht_accumulate_kernel(const float* __restrict__ Z, int ld, int d, int B, const int* __restrict__ feature,
                     const float* __restrict__ threshold, const int* __restrict__ left, const int* __restrict__ right,
                     const int* __restrict__ slot, int C, float* __restrict__ cnt, float* __restrict__ sum,
                     float* __restrict__ sq, int* __restrict__ leaf_out, double* __restrict__ stats) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  __shared__ int s_miss[8], s_n[8];
  int miss = 0, nfit = 0;
  if (warp < B) {
    const float* zr = Z + static_cast<size_t>(warp) * ld;
    if (zr[d] != 0.f) {
      const int node = ht_route(zr, feature, threshold, left, right);
      const int s = slot[node];
      const int c = ht_class(zr[d + 1], C);
      // prequential prediction = majority class of the leaf before this record is added
      float best = 0.f; int pred = -1;
      for (int k = 0; k < C; ++k) {
        const float v = __ldcg(&cnt[s * C + k]);   // other warps add to these counts concurrently (mini-batch semantics)
        if (v > best) { best = v; pred = k; }
      }
      miss = pred != c;
      nfit = 1;
      if (lane == 0) {
        atomicAdd(&cnt[s * C + c], 1.0f);
        if (leaf_out) leaf_out[warp] = node;
      }
      float* ps = sum + (static_cast<size_t>(s) * C + c) * d;
      float* pq = sq + (static_cast<size_t>(s) * C + c) * d;
      for (int f = lane; f < d; f += 32) {
        const float x = zr[f];
        atomicAdd(&ps[f], x);
        atomicAdd(&pq[f], x * x);
      }
    } else if (lane == 0 && leaf_out) {
      leaf_out[warp] = -1;
    }
  }
  const int w = threadIdx.x >> 5;
  if (lane == 0) { s_miss[w] = miss; s_n[w] = nfit; }
  __syncthreads();
  if (threadIdx.x == 0) {
    int m = 0, n = 0;
    for (int i = 0; i < (blockDim.x >> 5); ++i) { m += s_miss[i]; n += s_n[i]; }
    if (n) { atomicAdd(&stats[0], static_cast<double>(m)); atomicAdd(&stats[2], static_cast<double>(n)); }
  }
}

// forecast path: leaf index per record (the class is the arg-max of the leaf's counts)
__global__ void ht_route_kernel(const float* __restrict__ Z, int ld, int B, const int* __restrict__ feature,
                                const float* __restrict__ threshold, const int* __restrict__ left,
                                const int* __restrict__ right, int* __restrict__ leaf_out) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= B) return;
  leaf_out[r] = ht_route(Z + static_cast<size_t>(r) * ld, feature, threshold, left, right);
}

}  // namespace omldm

using namespace omldm;

extern "C" int omldm_ht_accumulate(const float* Z, int ld, int d, int B, const int* feature, const float* threshold,
                                   const int* left, const int* right, const int* slot, int C, float* cnt, float* sum,
                                   float* sq, int* leaf_out, double* stats, void* stream) {
  if (B <= 0) return 0;
  ht_accumulate_kernel<<<(B + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      Z, ld, d, B, feature, threshold, left, right, slot, C, cnt, sum, sq, leaf_out, stats);
  return check_launch("ht_accumulate_kernel");
}

extern "C" int omldm_ht_route(const float* Z, int ld, int B, const int* feature, const float* threshold,
                              const int* left, const int* right, int* leaf_out, void* stream) {
  if (B <= 0) return 0;
  ht_route_kernel<<<(B + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(Z, ld, B, feature, threshold, left,
                                                                                right, leaf_out);
  return check_launch("ht_route_kernel");
}
