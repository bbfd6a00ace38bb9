// Pipeline preprocessors on staged record rows: PolynomialFeatures, StandardScaler, MinMaxScaler
// (whitelist: utils/parsers/requestStream/PipelineMap.scala:67; the maths is out of tree).
//
// All three keep the bias/valid and target columns intact and skip empty rows (bias == 0).
#include "ptx.cuh"
#include "host_common.h"
#include "omldm_kernels.h"

namespace omldm {

// factors: [d_out][degree] input-column indices, -1 = unused factor.  Monomials are ordered
// highest degree first (FlinkML-style), no constant term.
__global__ void poly_features_kernel(const float* __restrict__ Zin, int ldin, int d_in,
                                     const short* __restrict__ factors, int degree, int d_out,
                                     float* __restrict__ Zout, int ldout, int B) {
  extern __shared__ float s_x[];  // d_in + 2
  const int r = blockIdx.x;
  if (r >= B) return;
  const float* src = Zin + static_cast<size_t>(r) * ldin;
  for (int c = threadIdx.x; c < d_in + 2; c += blockDim.x) s_x[c] = src[c];
  __syncthreads();
  float* dst = Zout + static_cast<size_t>(r) * ldout;
  const float valid = s_x[d_in];
  for (int c = threadIdx.x; c < ldout; c += blockDim.x) {
    float v = 0.f;
    if (c < d_out) {
      if (valid != 0.f) {
        v = 1.f;
        const short* f = factors + c * degree;
        for (int k = 0; k < degree; ++k) {
          const int idx = f[k];
          if (idx >= 0) v *= s_x[idx];
        }
      }
    } else if (c == d_out) {
      v = valid;
    } else if (c == d_out + 1) {
      v = s_x[d_in + 1];
    }
    dst[c] = v;
  }
}

// ---------------------------------------------------------------- StandardScaler
// pass 1: per-column sum and sum of squares of the batch (shifted by the running mean for
// stability), accumulated in double through `scratch` = [n_valid, done, sum[d], sumsq[d]].
__global__ void scaler_stats_kernel(const float* __restrict__ Z, int ld, int d, int B,
                                    const double* __restrict__ state, double* __restrict__ scratch) {
  // block = 32 x 8: threadIdx.x walks columns, threadIdx.y walks rows
  const int col = blockIdx.x * 32 + threadIdx.x;
  double s = 0.0, ss = 0.0;
  int nvalid = 0;
  const double shift = col < d ? state[1 + col] : 0.0;
  for (int r = blockIdx.y * blockDim.y + threadIdx.y; r < B; r += gridDim.y * blockDim.y) {
    const float* zr = Z + static_cast<size_t>(r) * ld;
    if (zr[d] == 0.f) continue;
    nvalid += 1;
    if (col < d) {
      const double v = static_cast<double>(zr[col]) - shift;
      s += v;
      ss += v * v;
    }
  }
  __shared__ double sh_s[8][33], sh_ss[8][33];
  __shared__ int sh_n[8];
  sh_s[threadIdx.y][threadIdx.x] = s;
  sh_ss[threadIdx.y][threadIdx.x] = ss;
  if (threadIdx.x == 0) sh_n[threadIdx.y] = nvalid;
  __syncthreads();
  if (threadIdx.y == 0) {
    for (int i = 1; i < 8; ++i) { s += sh_s[i][threadIdx.x]; ss += sh_ss[i][threadIdx.x]; }
    if (col < d) {
      atomicAdd(&scratch[2 + col], s);
      atomicAdd(&scratch[2 + d + col], ss);
    }
    if (threadIdx.x == 0 && blockIdx.x == 0) {
      int n = 0;
      for (int i = 0; i < 8; ++i) n += sh_n[i];
      atomicAdd(&scratch[0], static_cast<double>(n));
    }
  }
}

// pass 2: merge batch moments into (count, mean, M2) (Chan et al. parallel Welford) - done
// redundantly per block from the same inputs into registers - then transform the rows.
// The block with blockIdx.y == 0 for a column group commits the new state.
__global__ void scaler_apply_kernel(float* __restrict__ Z, int ld, int d, int B, double* __restrict__ state,
                                    const double* __restrict__ scratch, double* __restrict__ new_state,
                                    int fit) {
  const int col = blockIdx.x * 32 + threadIdx.x;
  double mean = 0.0, m2 = 0.0, n = state[0];
  if (col < d) { mean = state[1 + col]; m2 = state[1 + d + col]; }
  if (fit) {
    const double nb = scratch[0];
    if (nb > 0.0 && col < d) {
      const double s = scratch[2 + col], ss = scratch[2 + d + col];
      const double mb_shift = s / nb;               // batch mean relative to the old mean
      const double m2b = ss - s * mb_shift;         // batch M2 around its own mean
      const double tot = n + nb;
      const double delta = mb_shift;                // (batch mean) - (old mean)
      m2 = m2 + m2b + delta * delta * n * nb / tot;
      mean = mean + delta * nb / tot;
    }
    n += nb;
    if (blockIdx.y == 0 && threadIdx.y == 0) {
      if (col < d) { new_state[1 + col] = mean; new_state[1 + d + col] = m2; }
      if (col == 0) new_state[0] = n;
    }
  }
  const double var = n > 1.0 ? m2 / (n - 1.0) : 0.0;
  const float inv_std = var > 1e-24 ? static_cast<float>(1.0 / sqrt(var)) : 1.f;
  const float fmean = static_cast<float>(mean);
  if (col < d) {
    for (int r = blockIdx.y * blockDim.y + threadIdx.y; r < B; r += gridDim.y * blockDim.y) {
      float* zr = Z + static_cast<size_t>(r) * ld;
      if (zr[d] == 0.f) continue;
      zr[col] = (zr[col] - fmean) * inv_std;
    }
  }
}

__global__ void scaler_commit_kernel(double* __restrict__ state, const double* __restrict__ new_state,
                                     double* __restrict__ scratch, int d) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 1 + 2 * d) state[i] = new_state[i];
  if (i < 2 + 2 * d) scratch[i] = 0.0;
}

// ---------------------------------------------------------------- MinMaxScaler
__device__ __forceinline__ void atomic_min_f(float* a, float v) {
  int* ai = reinterpret_cast<int*>(a);
  int old = __ldcg(ai);     // other blocks update this word atomically: read it from L2, not as a plain load
  while (__int_as_float(old) > v) {
    const int assumed = old;
    old = atomicCAS(ai, assumed, __float_as_int(v));
    if (old == assumed) break;
  }
}
__device__ __forceinline__ void atomic_max_f(float* a, float v) {
  int* ai = reinterpret_cast<int*>(a);
  int old = __ldcg(ai);     // other blocks update this word atomically: read it from L2, not as a plain load
  while (__int_as_float(old) < v) {
    const int assumed = old;
    old = atomicCAS(ai, assumed, __float_as_int(v));
    if (old == assumed) break;
  }
}

__global__ void minmax_stats_kernel(const float* __restrict__ Z, int ld, int d, int B, float* __restrict__ state) {
  const int col = blockIdx.x * 32 + threadIdx.x;
  float mn = INFINITY, mx = -INFINITY;
  for (int r = blockIdx.y * blockDim.y + threadIdx.y; r < B; r += gridDim.y * blockDim.y) {
    const float* zr = Z + static_cast<size_t>(r) * ld;
    if (zr[d] == 0.f || col >= d) continue;
    const float v = zr[col];
    mn = fminf(mn, v);
    mx = fmaxf(mx, v);
  }
  __shared__ float s_mn[8][33], s_mx[8][33];
  s_mn[threadIdx.y][threadIdx.x] = mn;
  s_mx[threadIdx.y][threadIdx.x] = mx;
  __syncthreads();
  if (threadIdx.y == 0 && col < d) {
    for (int i = 1; i < 8; ++i) { mn = fminf(mn, s_mn[i][threadIdx.x]); mx = fmaxf(mx, s_mx[i][threadIdx.x]); }
    if (mn <= mx) { atomic_min_f(&state[col], mn); atomic_max_f(&state[d + col], mx); }
  }
}

__global__ void minmax_apply_kernel(float* __restrict__ Z, int ld, int d, int B, const float* __restrict__ state) {
  const int col = blockIdx.x * 32 + threadIdx.x;
  if (col >= d) return;
  const float mn = state[col], mx = state[d + col];
  const float range = mx - mn;
  const float inv = range > 1e-30f ? 1.f / range : 0.f;
  for (int r = blockIdx.y * blockDim.y + threadIdx.y; r < B; r += gridDim.y * blockDim.y) {
    float* zr = Z + static_cast<size_t>(r) * ld;
    if (zr[d] == 0.f) continue;
    zr[col] = (zr[col] - mn) * inv;
  }
}

}  // namespace omldm

using namespace omldm;

extern "C" int omldm_poly_features(const float* Zin, int ldin, int d_in, const short* factors, int degree,
                                   int d_out, float* Zout, int ldout, int B, void* stream) {
  if (B <= 0) return 0;
  if (d_out + 2 > ldout) return set_error_msg("poly_features: output pitch too small");
  poly_features_kernel<<<B, 256, (d_in + 2) * sizeof(float), static_cast<cudaStream_t>(stream)>>>(
      Zin, ldin, d_in, factors, degree, d_out, Zout, ldout, B);
  return check_launch("poly_features_kernel");
}

extern "C" int omldm_standard_scaler(float* Z, int ld, int d, int B, double* state, double* scratch, int fit,
                                     void* stream) {
  if (B <= 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 block(32, 8);
  int gy = (B + 63) / 64;
  if (gy > 64) gy = 64;
  if (gy < 1) gy = 1;
  dim3 grid((d + 31) / 32, gy);
  double* new_state = scratch + 2 + 2 * d;  // scratch tail: [1 + 2d] staged state
  if (fit) {
    scaler_stats_kernel<<<grid, block, 0, st>>>(Z, ld, d, B, state, scratch);
    if (check_launch("scaler_stats_kernel")) return -1;
  }
  scaler_apply_kernel<<<grid, block, 0, st>>>(Z, ld, d, B, state, scratch, new_state, fit);
  if (check_launch("scaler_apply_kernel")) return -1;
  if (fit) {
    const int n = 2 + 2 * d;
    scaler_commit_kernel<<<(n + 255) / 256, 256, 0, st>>>(state, new_state, scratch, d);
    if (check_launch("scaler_commit_kernel")) return -1;
  }
  return 0;
}

extern "C" int omldm_minmax_scaler(float* Z, int ld, int d, int B, float* state, int fit, void* stream) {
  if (B <= 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 block(32, 8);
  int gy = (B + 63) / 64;
  if (gy > 64) gy = 64;
  dim3 grid((d + 31) / 32, gy);
  if (fit) {
    minmax_stats_kernel<<<grid, block, 0, st>>>(Z, ld, d, B, state);
    if (check_launch("minmax_stats_kernel")) return -1;
  }
  minmax_apply_kernel<<<grid, block, 0, st>>>(Z, ld, d, B, state);
  return check_launch("minmax_apply_kernel");
}
