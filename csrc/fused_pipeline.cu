// Fused pipeline for BASELINE config 3: PolynomialFeatures -> StandardScaler -> linear learner
// (mini-batch PA / PA regressor / SVM) without ever materialising the expanded rows.
//
//   pass 1  poly_moments_kernel : per-column sum / sum-of-squares of the monomials of the batch,
//                                 computed on the fly from the raw d_in-wide rows (fp64 accumulation)
//   pass 2  poly_scale_linear_kernel : merge the moments into the running (count, mean, M2)
//                                 (Chan et al.), then per record recompute the monomials in registers,
//                                 standardise, take the margin against the frozen weights, and
//                                 accumulate the averaged update; the last CTA applies it
//   pass 3  scaler_commit (preprocess.cu's commit kernel is reused through omldm_scaler_commit)
//
// HBM traffic per step drops from ~5 passes over B x d_out floats to 2 passes over B x d_in floats
// (d_out = 434 vs d_in = 28 for degree 2).  Semantics are identical to running
// omldm_poly_features + omldm_standard_scaler(fit) + omldm_linear_minibatch one after the other.
#include "ptx.cuh"
#include "host_common.h"
#include "omldm_kernels.h"
#include "linear_math.cuh"

namespace omldm {

constexpr int FP_MAX_DEG = 4;

// Factor indices are decoded once per column; an unused factor points at the row's valid-flag column,
// which is exactly 1.0 for every row that is processed, so the product is branch free.
template <int DEG>
__device__ __forceinline__ void load_factors(const short* __restrict__ factors, int c, int d_in, int (&idx)[DEG]) {
#pragma unroll
  for (int k = 0; k < DEG; ++k) {
    const int f = factors[c * DEG + k];
    idx[k] = f >= 0 ? f : d_in;
  }
}
template <int DEG>
__device__ __forceinline__ float monomial(const float* __restrict__ x, const int (&idx)[DEG]) {
  float v = x[idx[0]];
#pragma unroll
  for (int k = 1; k < DEG; ++k) v *= x[idx[k]];
  return v;
}

constexpr int PM_ROWS = 64;

template <int DEG>
__global__ void __launch_bounds__(256)
poly_moments_kernel(const float* __restrict__ Z, int ld, int d_in, int B, const short* __restrict__ factors,
                    int d_out, const double* __restrict__ state, double* __restrict__ scratch) {
  extern __shared__ float s_x[];   // PM_ROWS x (d_in + 1): features + valid flag
  const int pitch = d_in + 1;
  double acc_s[2] = {0.0, 0.0}, acc_q[2] = {0.0, 0.0};   // up to 2 columns per thread (d_out <= 512)
  int idx[2][DEG];
  double shift[2] = {0.0, 0.0};
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int c = threadIdx.x + j * blockDim.x;
    if (c < d_out) {
      load_factors<DEG>(factors, c, d_in, idx[j]);
      shift[j] = state[1 + c];
    } else {
#pragma unroll
      for (int k = 0; k < DEG; ++k) idx[j][k] = d_in;
    }
  }
  int nvalid = 0;
  for (int r0 = blockIdx.x * PM_ROWS; r0 < B; r0 += gridDim.x * PM_ROWS) {
    const int rows = min(PM_ROWS, B - r0);
    for (int r = threadIdx.x >> 5; r < rows; r += blockDim.x >> 5)      // one warp per row: coalesced
      for (int c = threadIdx.x & 31; c < pitch; c += 32)
        s_x[r * pitch + c] = Z[static_cast<size_t>(r0 + r) * ld + c];
    __syncthreads();
    for (int r = 0; r < rows; ++r) {
      const float* x = s_x + r * pitch;
      if (x[d_in] == 0.f) continue;            // block-uniform
      if (threadIdx.x == 0) nvalid += 1;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const double v = static_cast<double>(monomial<DEG>(x, idx[j])) - shift[j];   // shifted: stable sums
        acc_s[j] += v;
        acc_q[j] += v * v;
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int c = threadIdx.x + j * blockDim.x;
    if (c < d_out) {
      atomicAdd(&scratch[2 + c], acc_s[j]);
      atomicAdd(&scratch[2 + d_out + c], acc_q[j]);
    }
  }
  if (threadIdx.x == 0 && nvalid) atomicAdd(&scratch[0], static_cast<double>(nvalid));
}

constexpr int FL_WARPS = 8;
constexpr int FL_MAXC = 16;   // columns per lane: d_out <= 512

// lin_scratch layout (floats) = omldm_linear_minibatch's: [0, dw) delta | dw active count | dw+1 blocks done
template <int DEG>
__global__ void __launch_bounds__(FL_WARPS * 32)
poly_scale_linear_kernel(const float* __restrict__ Z, int ld, int d_in, int B, const short* __restrict__ factors,
                         int d_out, const double* __restrict__ state, const double* __restrict__ scratch,
                         double* __restrict__ new_state, omldm_linear_desc D, float* __restrict__ lin_scratch,
                         int mean_update) {
  extern __shared__ float sm[];
  float* s_mean = sm;                  // d_out
  float* s_inv = sm + d_out;           // d_out
  float* s_w = s_inv + d_out;          // d_out + 1
  float* s_delta = s_w + d_out + 1;    // d_out + 1
  float* s_rows = s_delta + d_out + 1; // FL_WARPS x (d_in + 2): one staged raw row per warp
  __shared__ double s_loss[FL_WARPS], s_aux[FL_WARPS];
  __shared__ int s_fit[FL_WARPS], s_act[FL_WARPS];
  __shared__ bool s_last;
  const int dw = d_out + 1;
  // ---- merged scaler statistics (every CTA derives them from the same inputs)
  {
    const double n0 = state[0], nb = scratch[0], tot = n0 + nb;
    for (int c = threadIdx.x; c < d_out; c += blockDim.x) {
      double mean = state[1 + c], m2 = state[1 + d_out + c];
      if (nb > 0.0) {
        const double s = scratch[2 + c], ss = scratch[2 + d_out + c];
        const double delta = s / nb;
        m2 = m2 + (ss - s * delta) + delta * delta * n0 * nb / tot;
        mean = mean + delta * nb / tot;
      }
      if (blockIdx.x == 0) { new_state[1 + c] = mean; new_state[1 + d_out + c] = m2; }
      const double var = tot > 1.0 ? m2 / (tot - 1.0) : 0.0;
      s_mean[c] = static_cast<float>(mean);
      s_inv[c] = var > 1e-24 ? static_cast<float>(1.0 / sqrt(var)) : 1.f;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) new_state[0] = tot;
  }
  for (int c = threadIdx.x; c < dw; c += blockDim.x) { s_w[c] = D.w[c]; s_delta[c] = 0.f; }
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double loss_sum = 0.0, aux_sum = 0.0;
  int fitted = 0, active = 0;
  const float svm_eta = (D.kind == 2) ? 1.f / (D.lambda * static_cast<float>(D.stats[3] + 1.0)) : 0.f;
  // this lane's columns: factor indices and scaler constants live in registers for the whole kernel
  int idx[FL_MAXC][DEG];
  float cm[FL_MAXC], ci[FL_MAXC], cw[FL_MAXC];
#pragma unroll
  for (int j = 0; j < FL_MAXC; ++j) {
    const int c = lane + 32 * j;
    if (c < d_out) {
      load_factors<DEG>(factors, c, d_in, idx[j]);
      cm[j] = s_mean[c]; ci[j] = s_inv[c]; cw[j] = s_w[c];
    } else {
#pragma unroll
      for (int k = 0; k < DEG; ++k) idx[j][k] = d_in;
      cm[j] = 1.f; ci[j] = 0.f; cw[j] = 0.f;   // padding column: z = (1 - 1) * 0 = 0
    }
  }
  float* x = s_rows + warp * (d_in + 2);
  for (int r = blockIdx.x * FL_WARPS + warp; r < B; r += gridDim.x * FL_WARPS) {
    const float* zr = Z + static_cast<size_t>(r) * ld;
    __syncwarp();
    for (int c = lane; c < d_in + 2; c += 32) x[c] = zr[c];   // features | valid (= 1) | target
    __syncwarp();
    if (x[d_in] == 0.f) continue;
    float z[FL_MAXC];
    float dot = 0.f, nrm = 0.f;
#pragma unroll
    for (int j = 0; j < FL_MAXC; ++j) {
      z[j] = (monomial<DEG>(x, idx[j]) - cm[j]) * ci[j];
      dot = fmaf(cw[j], z[j], dot);
      nrm = fmaf(z[j], z[j], nrm);
    }
    dot = warp_sum(dot) + s_w[d_out];   // bias feature = 1
    nrm = warp_sum(nrm) + 1.f;
    const float y = x[d_in + 1];
    float loss, aux;
    float coef = linear_coef(D.kind, D.variant, D.C, D.eps, dot, y, nrm, loss, aux);
    if (D.kind == 2) {
      const float yy = y > 0.f ? 1.f : -1.f;
      coef = (yy * dot < 1.f) ? svm_eta * yy : 0.f;
    }
    if (coef != 0.f) {
#pragma unroll
      for (int j = 0; j < FL_MAXC; ++j) {
        const int c = lane + 32 * j;
        if (c < d_out) atomicAdd(&s_delta[c], coef * z[j]);
      }
      if (lane == 0) atomicAdd(&s_delta[d_out], coef);
      active += 1;
    }
    loss_sum += loss;
    aux_sum += aux;
    fitted += 1;
  }
  if (lane == 0) { s_loss[warp] = loss_sum; s_aux[warp] = aux_sum; s_fit[warp] = fitted; s_act[warp] = active; }
  __syncthreads();
  for (int c = threadIdx.x; c < dw; c += blockDim.x) {
    const float v = s_delta[c];
    if (v != 0.f) atomicAdd(&lin_scratch[c], v);
  }
  // every thread's additions must be out before thread 0 announces this block as done: without the barrier the last
  // block could read (and then clear) the accumulator while a slower warp of this block is still adding to it
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    double l = 0, a = 0; int f = 0, ac = 0;
    for (int i = 0; i < FL_WARPS; ++i) { l += s_loss[i]; a += s_aux[i]; f += s_fit[i]; ac += s_act[i]; }
    atomicAdd(&D.stats[0], l);
    atomicAdd(&D.stats[1], a);
    atomicAdd(&D.stats[2], static_cast<double>(f));
    if (ac) atomicAdd(&lin_scratch[dw], static_cast<float>(ac));
    __threadfence();
    const unsigned done = atomicAdd(reinterpret_cast<unsigned*>(&lin_scratch[dw + 1]), 1u);
    s_last = (done == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    __threadfence();
    const float n_active = __ldcg(&lin_scratch[dw]);
    float scale = 1.f;
    if (mean_update && D.kind != 2) scale = n_active > 0.f ? 1.f / n_active : 0.f;
    if (D.kind == 2) {
      const float shrink = 1.f - svm_eta * D.lambda;
      const float inv_b = 1.f / fmaxf(1.f, static_cast<float>(B));
      for (int c = threadIdx.x; c < dw; c += blockDim.x) {
        D.w[c] = shrink * D.w[c] + inv_b * __ldcg(&lin_scratch[c]);
        lin_scratch[c] = 0.f;
      }
      if (threadIdx.x == 0) D.stats[3] += 1.0;
    } else {
      for (int c = threadIdx.x; c < dw; c += blockDim.x) {
        D.w[c] += scale * __ldcg(&lin_scratch[c]);
        lin_scratch[c] = 0.f;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) { lin_scratch[dw] = 0.f; lin_scratch[dw + 1] = 0.f; }
  }
}

__global__ void fused_commit_kernel(double* __restrict__ state, const double* __restrict__ new_state,
                                    double* __restrict__ scratch, int d) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 1 + 2 * d) state[i] = new_state[i];
  if (i < 2 + 2 * d) scratch[i] = 0.0;
}

}  // namespace omldm

using namespace omldm;

extern "C" int omldm_fused_poly_scaler_linear(const float* Z, int ld, int d_in, int B, const short* factors, int degree,
                                              int d_out, double* scaler_state, double* scaler_scratch,
                                              const omldm_linear_desc* desc_host, float* lin_scratch, int mean_update,
                                              void* stream) {
  if (B <= 0) return 0;
  if (degree < 1 || degree > FP_MAX_DEG) return set_error_msg("fused pipeline: polynomial degree must be 1..4");
  if (d_out > 32 * FL_MAXC) return set_error_msg("fused pipeline: more than 512 expanded features; use the unfused kernels");
  if (desc_host->d != d_out) return set_error_msg("fused pipeline: learner width must equal the expanded width");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  double* new_state = scaler_scratch + 2 + 2 * d_out;
  int g1 = (B + PM_ROWS - 1) / PM_ROWS;
  if (g1 > 148 * 2) g1 = 148 * 2;
  int g2 = (B + FL_WARPS - 1) / FL_WARPS;
  if (g2 > 148 * 4) g2 = 148 * 4;
  const size_t smem1 = PM_ROWS * (d_in + 1) * sizeof(float);
  const size_t smem2 = (4 * static_cast<size_t>(d_out) + 2 + FL_WARPS * (d_in + 2)) * sizeof(float);
#define OMLDM_FUSED(DEG)                                                                                          \
  do {                                                                                                            \
    poly_moments_kernel<DEG><<<g1, 256, smem1, st>>>(Z, ld, d_in, B, factors, d_out, scaler_state, scaler_scratch); \
    if (check_launch("poly_moments_kernel")) return -1;                                                           \
    poly_scale_linear_kernel<DEG><<<g2, FL_WARPS * 32, smem2, st>>>(Z, ld, d_in, B, factors, d_out, scaler_state, \
                                                                    scaler_scratch, new_state, *desc_host,        \
                                                                    lin_scratch, mean_update);                    \
    if (check_launch("poly_scale_linear_kernel")) return -1;                                                      \
  } while (0)
  switch (degree) {
    case 1: OMLDM_FUSED(1); break;
    case 2: OMLDM_FUSED(2); break;
    case 3: OMLDM_FUSED(3); break;
    default: OMLDM_FUSED(4); break;
  }
#undef OMLDM_FUSED
  const int n = 2 + 2 * d_out;
  fused_commit_kernel<<<(n + 255) / 256, 256, 0, st>>>(scaler_state, new_state, scaler_scratch, d_out);
  return check_launch("fused_commit_kernel");
}
