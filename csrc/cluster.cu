// Epilogue / bookkeeping kernels around the tcgen05 GEMMs for k-means and multiclass PA, plus a
// few small vector helpers.  Learner names: utils/parsers/requestStream/PipelineMap.scala:68;
// formulas are the literature ones (MacQueen 1967; Crammer et al. 2006).
#include "ptx.cuh"
#include "host_common.h"
#include "omldm_kernels.h"

namespace omldm {

// One warp per record: j = argmin_j (||c_j||^2 - 2 x.c_j); inertia += ||x||^2 + that minimum.
// Writes the one-hot row T[r, :] used by the outer-product update sums += T^T Z (the bias column
// of Z turns into per-centroid counts).
__global__ void kmeans_assign_kernel(const float* __restrict__ Z, int ld, int d, int B,
                                     const float* __restrict__ S, int lds, const float* __restrict__ cnorm,
                                     int K, float* __restrict__ T, int ldt, int* __restrict__ assign,
                                     double* __restrict__ stats) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  __shared__ double s_in[8];
  __shared__ int s_n[8];
  double inertia = 0.0;
  int nfit = 0;
  if (warp < B) {
    const float* zr = Z + static_cast<size_t>(warp) * ld;
    const bool valid = zr[d] != 0.f;
    float best = INFINITY;
    int bj = 0;
    if (valid) {
      const float* sr = S + static_cast<size_t>(warp) * lds;
      for (int j = lane; j < K; j += 32) {
        const float v = cnorm[j] - 2.f * sr[j];
        if (v < best) { best = v; bj = j; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oj = __shfl_xor_sync(0xffffffffu, bj, o);
        if (ob < best || (ob == best && oj < bj)) { best = ob; bj = oj; }
      }
      float nrm = 0.f;
      for (int c = lane; c < d; c += 32) nrm = fmaf(zr[c], zr[c], nrm);
      nrm = warp_sum(nrm);
      inertia = fmax(0.0, static_cast<double>(nrm) + static_cast<double>(best));
      nfit = 1;
    }
    float* tr = T + static_cast<size_t>(warp) * ldt;
    for (int j = lane; j < ldt; j += 32) tr[j] = (valid && j == bj) ? 1.f : 0.f;
    if (lane == 0 && assign) assign[warp] = valid ? bj : -1;
  }
  const int w = threadIdx.x >> 5;
  if (lane == 0) { s_in[w] = inertia; s_n[w] = nfit; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0; int n = 0;
    for (int i = 0; i < (blockDim.x >> 5); ++i) { a += s_in[i]; n += s_n[i]; }
    if (n) { atomicAdd(&stats[0], a); atomicAdd(&stats[2], static_cast<double>(n)); }
  }
}

__global__ void kmeans_finalize_kernel(const float* __restrict__ sums, const float* __restrict__ sums2, int lds,
                                       int K, int d, float* __restrict__ centroids, int ldc,
                                       float* __restrict__ cnorm) {
  const int j = blockIdx.x;
  if (j >= K) return;
  const float cnt = sums[static_cast<size_t>(j) * lds + d] + (sums2 ? sums2[static_cast<size_t>(j) * lds + d] : 0.f);
  const float inv = cnt > 0.f ? 1.f / cnt : 0.f;
  float acc = 0.f;
  for (int c = threadIdx.x; c < ldc; c += blockDim.x) {
    float v = 0.f;
    if (c < d) v = (sums[static_cast<size_t>(j) * lds + c] + (sums2 ? sums2[static_cast<size_t>(j) * lds + c] : 0.f)) * inv;
    centroids[static_cast<size_t>(j) * ldc + c] = v;
    acc = fmaf(v, v, acc);
  }
  __shared__ float s[32];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < (blockDim.x >> 5); ++i) t += s[i];
    // a centroid with no mass must never win the argmin
    cnorm[j] = cnt > 0.f ? t : INFINITY;
  }
}

// Exact MacQueen pass: one CTA, record by record.  sums rows hold S_j = n_j * c_j and n_j, so the
// update c_j += (x - c_j) / n_j is S_j += x, n_j += 1 - additive, which is what lets replicas
// merge by summation under the sync protocols.
__global__ void __launch_bounds__(256)
kmeans_exact_kernel(const float* __restrict__ Z, int ld, int d, int B, float* __restrict__ sums, int lds,
                    int K, double* __restrict__ stats) {
  extern __shared__ float sm[];
  float* s_x = sm;           // d
  float* s_best = sm + d;    // 8 warps
  int* s_bj = reinterpret_cast<int*>(s_best + 8);
  __shared__ int s_choice;
  double inertia = 0.0;
  int nfit = 0;
  for (int r = 0; r < B; ++r) {
    const float* zr = Z + static_cast<size_t>(r) * ld;
    if (zr[d] == 0.f) continue;  // uniform branch
    for (int c = threadIdx.x; c < d; c += blockDim.x) s_x[c] = zr[c];
    __syncthreads();
    float best = INFINITY;
    int bj = 0x7fffffff;
    for (int j = threadIdx.x; j < K; j += blockDim.x) {
      const float* sr = sums + static_cast<size_t>(j) * lds;
      const float cnt = sr[d];
      float dist = INFINITY;
      if (cnt > 0.f) {
        const float inv = 1.f / cnt;
        dist = 0.f;
        for (int c = 0; c < d; ++c) {
          const float df = s_x[c] - sr[c] * inv;
          dist = fmaf(df, df, dist);
        }
      } else {
        dist = -1.f - static_cast<float>(K - j);  // empty centroid: claim it (lowest index first)
      }
      if (dist < best) { best = dist; bj = j; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oj = __shfl_xor_sync(0xffffffffu, bj, o);
      if (ob < best || (ob == best && oj < bj)) { best = ob; bj = oj; }
    }
    if ((threadIdx.x & 31) == 0) { s_best[threadIdx.x >> 5] = best; s_bj[threadIdx.x >> 5] = bj; }
    __syncthreads();
    if (threadIdx.x == 0) {
      float b = s_best[0]; int j = s_bj[0];
      for (int i = 1; i < (blockDim.x >> 5); ++i)
        if (s_best[i] < b || (s_best[i] == b && s_bj[i] < j)) { b = s_best[i]; j = s_bj[i]; }
      s_choice = j;
      inertia += b > 0.f ? b : 0.f;
      nfit += 1;
    }
    __syncthreads();
    float* sr = sums + static_cast<size_t>(s_choice) * lds;
    for (int c = threadIdx.x; c <= d; c += blockDim.x) sr[c] += (c < d) ? s_x[c] : 1.f;
    __syncthreads();
  }
  if (threadIdx.x == 0) { stats[0] += inertia; stats[2] += static_cast<double>(nfit); }
}

// Multiclass PA coefficients.  One warp per record: r = argmax_{s != y} S_s,
// loss = max(0, 1 - (S_y - S_r)), tau from loss / (2 ||x||^2) (PA / PA-I / PA-II).
__global__ void mcpa_coef_kernel(const float* __restrict__ Z, int ld, int d, int B, const float* __restrict__ S,
                                 int lds, int C, const float* __restrict__ sqnorm, int variant, float Creg,
                                 float* __restrict__ T, int ldt, double* __restrict__ stats) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  __shared__ double s_l[8], s_m[8];
  __shared__ int s_n[8], s_a[8];
  double loss_d = 0.0, mist = 0.0;
  int nfit = 0, nact = 0;
  if (warp < B) {
    const float* zr = Z + static_cast<size_t>(warp) * ld;
    const bool valid = zr[d] != 0.f;
    int y = -1, rj = -1;
    float tau = 0.f;
    if (valid) {
      y = static_cast<int>(zr[d + 1] + 0.5f);
      const float* sr = S + static_cast<size_t>(warp) * lds;
      float best = -INFINITY;
      for (int j = lane; j < C; j += 32) {
        const float v = sr[j];
        if (j != y && v > best) { best = v; rj = j; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oj = __shfl_xor_sync(0xffffffffu, rj, o);
        if (ob > best || (ob == best && oj >= 0 && (rj < 0 || oj < rj))) { best = ob; rj = oj; }
      }
      const float sy = (y >= 0 && y < C) ? sr[y] : 0.f;
      const float loss = (y >= 0 && y < C && rj >= 0) ? fmaxf(0.f, 1.f - (sy - best)) : 0.f;
      const float n2 = 2.f * sqnorm[warp];
      if (loss > 0.f && n2 > 0.f) {
        tau = variant == 0 ? loss / n2 : variant == 1 ? fminf(Creg, loss / n2) : loss / (n2 + 1.f / (2.f * Creg));
      }
      loss_d = loss;
      mist = (best >= sy) ? 1.0 : 0.0;
      nfit = 1;
      nact = tau > 0.f ? 1 : 0;
    }
    float* tr = T + static_cast<size_t>(warp) * ldt;
    for (int j = lane; j < ldt; j += 32) tr[j] = (j == y) ? tau : (j == rj ? -tau : 0.f);
  }
  const int w = threadIdx.x >> 5;
  if (lane == 0) { s_l[w] = loss_d; s_m[w] = mist; s_n[w] = nfit; s_a[w] = nact; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double l = 0, m = 0; int n = 0, a = 0;
    for (int i = 0; i < (blockDim.x >> 5); ++i) { l += s_l[i]; m += s_m[i]; n += s_n[i]; a += s_a[i]; }
    if (n) {
      atomicAdd(&stats[0], l); atomicAdd(&stats[1], m); atomicAdd(&stats[2], static_cast<double>(n));
      atomicAdd(&stats[3], static_cast<double>(a));
    }
  }
}

__global__ void row_sqnorm_kernel(const float* __restrict__ Z, int ld, int n, int B, float* __restrict__ out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B) return;
  const float* zr = Z + static_cast<size_t>(warp) * ld;
  float acc = 0.f;
  if ((n & 3) == 0 && (ld & 3) == 0) {
    const float4* z4 = reinterpret_cast<const float4*>(zr);
    for (int c = lane; c < n / 4; c += 32) {
      const float4 v = z4[c];
      acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
  } else {
    for (int c = lane; c < n; c += 32) acc = fmaf(zr[c], zr[c], acc);
  }
  acc = warp_sum(acc);
  if (lane == 0) out[warp] = acc;
}

__global__ void axpy_kernel(float* __restrict__ y, const float* __restrict__ x, float alpha, size_t n) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x)
    y[i] = fmaf(alpha, x[i], y[i]);
}

// T[r, :] *= scale[0] (device scalar), e.g. 1 / n_active for averaged mini-batch updates.
__global__ void scale_rows_kernel(float* __restrict__ T, int ldt, int B, int C, const float* __restrict__ scale,
                                  int reciprocal) {
  const float s = reciprocal ? 1.f / fmaxf(*scale, 1.f) : *scale;
  const size_t total = static_cast<size_t>(B) * C;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t r = i / C, c = i % C;
    T[r * ldt + c] *= s;
  }
}

// sum over valid rows of (Z_r . wt)^2, wt has n entries (n <= ld).  One warp per row, float4
// loads, two rows in flight per warp; wt is staged in shared memory once per CTA.
__global__ void __launch_bounds__(256)
sq_residual_kernel(const float* __restrict__ Z, int ld, int B, int d, const float* __restrict__ wt, int n,
                   double* __restrict__ stats) {
  extern __shared__ float s_w[];   // ld floats, zero beyond n
  for (int c = threadIdx.x; c < ld; c += blockDim.x) s_w[c] = c < n ? wt[c] : 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int n4 = (n + 3) >> 2;
  const float4* w4 = reinterpret_cast<const float4*>(s_w);
  double acc = 0.0;
  int nv = 0;
  for (int r = warp; r < B; r += 2 * nwarps) {
    const int r2 = r + nwarps;
    const float4* z0 = reinterpret_cast<const float4*>(Z + static_cast<size_t>(r) * ld);
    const float4* z1 = reinterpret_cast<const float4*>(Z + static_cast<size_t>(r2 < B ? r2 : r) * ld);
    float d0 = 0.f, d1 = 0.f;
    for (int c = lane; c < n4; c += 32) {
      const float4 a = ld_nc_v4(z0 + c), b = ld_nc_v4(z1 + c), w = w4[c];
      d0 += a.x * w.x + a.y * w.y + a.z * w.z + a.w * w.w;
      d1 += b.x * w.x + b.y * w.y + b.z * w.z + b.w * w.w;
    }
    d0 = warp_sum(d0);
    d1 = warp_sum(d1);
    if (Z[static_cast<size_t>(r) * ld + d] != 0.f) { acc += static_cast<double>(d0) * d0; nv += 1; }
    if (r2 < B && Z[static_cast<size_t>(r2) * ld + d] != 0.f) { acc += static_cast<double>(d1) * d1; nv += 1; }
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

__global__ void mirror_upper_kernel(float* __restrict__ A, int ld, int n) {
  __shared__ float tile[32][33];
  const int bi = blockIdx.y, bj = blockIdx.x;
  if (bj < bi) return;  // read upper tiles only
  const int i = bi * 32 + threadIdx.y, j = bj * 32 + threadIdx.x;
  tile[threadIdx.y][threadIdx.x] = (i < n && j < n) ? A[static_cast<size_t>(i) * ld + j] : 0.f;
  __syncthreads();
  const int oi = bj * 32 + threadIdx.y, oj = bi * 32 + threadIdx.x;  // transposed position
  if (oi < n && oj < n && oi > oj) A[static_cast<size_t>(oi) * ld + oj] = tile[threadIdx.x][threadIdx.y];
}

}  // namespace omldm

using namespace omldm;

extern "C" int omldm_sq_residual(const float* Z, int ld, int B, int d, const float* wt, int n, double* stats,
                                 void* stream) {
  if (B <= 0) return 0;
  if ((ld & 3) || n > ld) return set_error_msg("sq_residual: ld must be a multiple of 4 and >= n");
  if (ld * sizeof(float) > 200 * 1024) return set_error_msg("sq_residual: row pitch too large for the smem weight copy");
  if (ld * sizeof(float) > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(sq_residual_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(ld * sizeof(float)));
    if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(sq_residual)", e);
  }
  int blocks = (B + 15) / 16;
  if (blocks > 148 * 4) blocks = 148 * 4;
  sq_residual_kernel<<<blocks, 256, ld * sizeof(float), static_cast<cudaStream_t>(stream)>>>(Z, ld, B, d, wt, n, stats);
  return check_launch("sq_residual_kernel");
}

extern "C" int omldm_mirror_upper(float* A, int ld, int n, void* stream) {
  dim3 grid((n + 31) / 32, (n + 31) / 32), block(32, 32);
  mirror_upper_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(A, ld, n);
  return check_launch("mirror_upper_kernel");
}

extern "C" int omldm_kmeans_assign(const float* Z, int ld, int d, int B, const float* S, int lds,
                                   const float* cnorm, int K, float* T, int ldt, int* assign, double* stats,
                                   void* stream) {
  if (B <= 0) return 0;
  kmeans_assign_kernel<<<(B + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(Z, ld, d, B, S, lds, cnorm,
                                                                                   K, T, ldt, assign, stats);
  return check_launch("kmeans_assign_kernel");
}

extern "C" int omldm_kmeans_finalize(const float* sums, const float* sums2, int lds, int K, int d, float* centroids,
                                     int ldc, float* cnorm, void* stream) {
  kmeans_finalize_kernel<<<K, 128, 0, static_cast<cudaStream_t>(stream)>>>(sums, sums2, lds, K, d, centroids, ldc, cnorm);
  return check_launch("kmeans_finalize_kernel");
}

extern "C" int omldm_kmeans_exact(const float* Z, int ld, int d, int B, float* sums, int lds, int K,
                                  double* stats, void* stream) {
  if (B <= 0) return 0;
  const size_t smem = (d + 8) * sizeof(float) + 8 * sizeof(int);
  kmeans_exact_kernel<<<1, 256, smem, static_cast<cudaStream_t>(stream)>>>(Z, ld, d, B, sums, lds, K, stats);
  return check_launch("kmeans_exact_kernel");
}

extern "C" int omldm_mcpa_coef(const float* Z, int ld, int d, int B, const float* S, int lds, int C,
                               const float* sqnorm, int variant, float Creg, float* T, int ldt, double* stats,
                               void* stream) {
  if (B <= 0) return 0;
  mcpa_coef_kernel<<<(B + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(Z, ld, d, B, S, lds, C, sqnorm,
                                                                               variant, Creg, T, ldt, stats);
  return check_launch("mcpa_coef_kernel");
}

extern "C" int omldm_row_sqnorm(const float* Z, int ld, int n, int B, float* out, void* stream) {
  if (B <= 0) return 0;
  row_sqnorm_kernel<<<(B + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(Z, ld, n, B, out);
  return check_launch("row_sqnorm_kernel");
}

extern "C" int omldm_axpy(float* y, const float* x, float alpha, size_t n, void* stream) {
  if (n == 0) return 0;
  int blocks = static_cast<int>((n + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  axpy_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(y, x, alpha, n);
  return check_launch("axpy_kernel");
}

extern "C" int omldm_scale_rows(float* T, int ldt, int B, int C, const float* scale_dev, int reciprocal, void* stream) {
  if (B <= 0) return 0;
  size_t total = static_cast<size_t>(B) * C;
  int blocks = static_cast<int>((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  scale_rows_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(T, ldt, B, C, scale_dev, reciprocal);
  return check_launch("scale_rows_kernel");
}
