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

// ------------------------------------------------------------------------------------------
// Split test on the device.  Phase 1, one CTA per node: a leaf that has seen `grace` records
// since its last check evaluates, for every feature, `n_points` thresholds between the class-conditional Gaussians
// (mean +- 2 sd), keeps the best information gain per feature, reduces the two best gains over the block and
// applies the Hoeffding bound  eps = sqrt(R^2 ln(1/delta) / (2 n)),  R = log2(classes): split if the best gain
// beats the runner-up by eps, or eps fell below the tie threshold.  Phase 2, one thread: the accepted splits are
// applied in ascending leaf order, so node numbering is deterministic and equals the host oracle's
// (oracle/learners.py::HoeffdingTree._try_split); children start from the parent's class prior cut by the Gaussian
// estimate of the chosen feature.
__device__ __forceinline__ float ht_entropy(const float* c, int C, float tot) {
  if (!(tot > 0.f)) return 0.f;
  float h = 0.f;
  for (int k = 0; k < C; ++k)
    if (c[k] > 0.f) { const float p = c[k] / tot; h -= p * log2f(p); }
  return h;
}

__device__ __forceinline__ float ht_cdf(float t, float mu, float sd) {
  return sd > 0.f ? 0.5f * (1.f + erff((t - mu) / (sd * 1.41421356f))) : (mu <= t ? 1.f : 0.f);
}

constexpr int HT_MAXC = 16;   // classes handled by the on-device split test

// decision[node] = (split?, feature, threshold bits, -)
__global__ void __launch_bounds__(128)
ht_split_eval_kernel(const int* __restrict__ feature, const int* __restrict__ n_nodes_dev, int cap, int C, int d,
                     const float* __restrict__ cnt, const float* __restrict__ sum, const float* __restrict__ sq,
                     float* __restrict__ checked, float grace, float delta, float tie, int n_points,
                     int4* __restrict__ decision) {
  const int node = blockIdx.x;
  if (threadIdx.x == 0 && node < cap) decision[node] = make_int4(0, -1, 0, 0);
  if (node >= *n_nodes_dev || node >= cap || feature[node] >= 0) return;
  __shared__ float s_cnt[HT_MAXC];
  __shared__ float s_g[128], s_g2[128], s_t[128];
  __shared__ int s_f[128], s_n[128];
  if (threadIdx.x < C) s_cnt[threadIdx.x] = cnt[node * C + threadIdx.x];
  __syncthreads();
  float tot = 0.f;
  int nz = 0;
  for (int k = 0; k < C; ++k) { tot += s_cnt[k]; nz += s_cnt[k] > 0.f; }
  if (tot - checked[node] < grace || nz < 2) return;            // block-uniform
  const float h0 = ht_entropy(s_cnt, C, tot);
  // thread-local best and runner-up over the features this thread visits
  float g1 = -INFINITY, t1 = 0.f, g2 = -INFINITY;
  int f1 = -1, n_eval = 0;
  for (int f = threadIdx.x; f < d; f += blockDim.x) {
    float mu[HT_MAXC], sd[HT_MAXC];
    float lo = INFINITY, hi = -INFINITY;
    for (int k = 0; k < C; ++k) {
      const float n = s_cnt[k];
      const size_t o = (static_cast<size_t>(node) * C + k) * d + f;
      mu[k] = n > 0.f ? sum[o] / n : 0.f;
      const float m2 = fmaxf(sq[o] - n * mu[k] * mu[k], 0.f);
      sd[k] = sqrtf(m2 / fmaxf(n - 1.f, 1.f));
      if (n > 0.f) { lo = fminf(lo, mu[k] - 2.f * sd[k]); hi = fmaxf(hi, mu[k] + 2.f * sd[k]); }
    }
    if (!(hi > lo)) continue;
    float g_f = -1.f, t_f = lo;
    for (int j = 1; j <= n_points; ++j) {
      const float t = lo + (hi - lo) * static_cast<float>(j) / static_cast<float>(n_points + 1);
      float l[HT_MAXC], r[HT_MAXC], lt = 0.f, rt = 0.f;
      for (int k = 0; k < C; ++k) {
        l[k] = s_cnt[k] * ht_cdf(t, mu[k], sd[k]);
        r[k] = s_cnt[k] - l[k];
        lt += l[k];
        rt += r[k];
      }
      const float g = h0 - (lt * ht_entropy(l, C, lt) + rt * ht_entropy(r, C, rt)) / tot;
      if (g > g_f) { g_f = g; t_f = t; }
    }
    ++n_eval;
    if (g_f > g1 || (g_f == g1 && f > f1)) { g2 = g1; g1 = g_f; t1 = t_f; f1 = f; }   // ties: larger feature index first
    else if (g_f > g2) g2 = g_f;
  }
  s_g[threadIdx.x] = g1;
  s_g2[threadIdx.x] = g2;
  s_t[threadIdx.x] = t1;
  s_f[threadIdx.x] = f1;
  s_n[threadIdx.x] = n_eval;
  __syncthreads();
  if (threadIdx.x == 0) {
    float b1 = -INFINITY, bt = 0.f, b2 = -INFINITY;
    int bf = -1, n_tot = 0;
    for (int i = 0; i < static_cast<int>(blockDim.x); ++i) {
      n_tot += s_n[i];
      if (s_f[i] < 0) continue;
      if (s_g[i] > b1 || (s_g[i] == b1 && s_f[i] > bf)) { b2 = fmaxf(b1, b2); b1 = s_g[i]; bt = s_t[i]; bf = s_f[i]; }
      else if (s_g[i] > b2) b2 = s_g[i];
      if (s_g2[i] > b2) b2 = s_g2[i];
    }
    checked[node] = tot;
    if (bf < 0) return;
    if (n_tot < 2) b2 = 0.f;
    const float R = log2f(static_cast<float>(C > 2 ? C : 2));
    const float eps = sqrtf(R * R * logf(1.f / delta) / (2.f * tot));
    const int split = (b1 > 0.f && (b1 - b2 > eps || eps < tie)) ? 1 : 0;
    decision[node] = make_int4(split, bf, __float_as_int(bt), __float_as_int(b1));
  }
}

// Phase 2: apply the accepted splits in ascending leaf order (single thread: the tree is tiny, the order matters).
__global__ void ht_split_apply_kernel(int* __restrict__ feature, float* __restrict__ threshold, int* __restrict__ left,
                                      int* __restrict__ right, int* __restrict__ n_nodes_dev, int cap, int C, int d,
                                      float* __restrict__ cnt, const float* __restrict__ sum,
                                      const float* __restrict__ sq, float* __restrict__ checked,
                                      const int4* __restrict__ decision, int* __restrict__ n_splits_out) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  int n = *n_nodes_dev, done = 0;
  const int n0 = n;
  for (int node = 0; node < n0; ++node) {
    const int4 dec = decision[node];
    if (!dec.x || feature[node] >= 0) continue;
    if (n + 2 > cap) break;
    const int f = dec.y;
    const float t = __int_as_float(dec.z);
    const int li = n, ri = n + 1;
    n += 2;
    float lsum = 0.f, rsum = 0.f;
    for (int k = 0; k < C; ++k) {
      const float c = cnt[node * C + k];
      const size_t o = (static_cast<size_t>(node) * C + k) * d + f;
      const float mu = c > 0.f ? sum[o] / c : 0.f;
      const float sd = c > 1.f ? sqrtf(fmaxf(sq[o] - c * mu * mu, 0.f) / fmaxf(c - 1.f, 1.f)) : 0.f;
      const float p = ht_cdf(t, mu, sd);
      cnt[li * C + k] = c * p;
      cnt[ri * C + k] = c * (1.f - p);
      lsum += c * p;
      rsum += c * (1.f - p);
    }
    feature[li] = -1; feature[ri] = -1;
    left[li] = right[li] = left[ri] = right[ri] = -1;
    checked[li] = lsum;
    checked[ri] = rsum;
    threshold[node] = t;
    left[node] = li;
    right[node] = ri;
    __threadfence();
    feature[node] = f;            // last: a concurrent router sees either the leaf or the complete split
    ++done;
  }
  *n_nodes_dev = n;
  if (n_splits_out) *n_splits_out = done;
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

// On-device Hoeffding split test + tree growth (two launches, no host round trip).  decision: cap x int4 scratch;
// n_nodes_dev: device int holding the number of nodes in use; n_splits_out (optional): splits applied by this call.
extern "C" int omldm_ht_split(int* feature, float* threshold, int* left, int* right, int* n_nodes_dev, int cap, int C,
                              int d, float* cnt, const float* sum, const float* sq, float* checked, float grace,
                              float delta, float tie, int n_points, void* decision, int* n_splits_out, void* stream) {
  if (cap <= 0) return 0;
  if (C < 2 || C > HT_MAXC) return set_error_msg("ht_split: 2 <= classes <= 16");
  if (n_points < 1) return set_error_msg("ht_split: at least one split point");
  if (reinterpret_cast<uintptr_t>(decision) & 15) return set_error_msg("ht_split: decision must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ht_split_eval_kernel<<<cap, 128, 0, st>>>(feature, n_nodes_dev, cap, C, d, cnt, sum, sq, checked, grace, delta, tie,
                                           n_points, static_cast<int4*>(decision));
  if (int rc = check_launch("ht_split_eval_kernel")) return rc;
  ht_split_apply_kernel<<<1, 32, 0, st>>>(feature, threshold, left, right, n_nodes_dev, cap, C, d, cnt, sum, sq, checked,
                                         static_cast<const int4*>(decision), n_splits_out);
  return check_launch("ht_split_apply_kernel");
}
