// Multiclass PA straight from sparse records (BASELINE.json config 5: 100k-dim sparse stream, dense model).
//
// The dense route rebuilds a B x 100k fp32 batch every step (100 MB for 256 records) only to multiply it
// with a 10 x 100k weight matrix; the records have 128 non-zeros.  Here a record stays a fixed-width
// sparse row and the two GEMMs become gathers and scatters against the dense model:
//
//   sparse row (pitch ld floats, ld % 32 == 0, "pseudo feature count" 2*nnz so that the generic staged-row
//   conventions hold):  [ index bits x nnz | value x nnz | valid (1.0, 0 = empty row) | target | 0-pad ]
//
//   coef   : one warp per record - scores s_c = sum_j W[c, idx_j] v_j + W[c, d] (bias), rival class, hinge
//            loss, PA step size tau (same rule and statistics as cluster.cu::mcpa_coef_kernel).  With up to 16
//            classes the scores are gathered from a class-minor copy Wt[col][16] made by a small transpose kernel
//            in front: one non-zero then costs two 32-byte sectors of L2 traffic instead of one per class
//   update : one warp per record - W[y, idx_j] += s tau v_j, W[rival, idx_j] -= s tau v_j (+ bias column),
//            s = 1 / (#active records) for the averaged mini-batch update; red.global.add
//
// Because the rows keep the staged-row shape, the in-place holdout kernel (ingest.cu) works on them as is.
// The reference has no counterpart: its MultiClassPA lives out of tree on dense Breeze vectors.
#include <math_constants.h>

#include "host_common.h"
#include "omldm_kernels.h"
#include "ptx.cuh"

namespace omldm {

constexpr int SP_MAXC = 32;   // classes held in registers per lane

// ELL slot (SoA: idx[B*nnz] | val[B*nnz] | target[B]) -> sparse rows
// (one warp per record, 8 records per block: one block of 128 threads per 288-float row was 16 384 tiny blocks, 14.8 us)
__global__ void __launch_bounds__(256)
sparse_rows_kernel(const int* __restrict__ idx, const float* __restrict__ val,
                   const float* __restrict__ tgt, int B, int nnz, float* __restrict__ rows, int ld) {
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= B) return;
  float* row = rows + static_cast<size_t>(r) * ld;
  const int* ri = idx + static_cast<size_t>(r) * nnz;
  const float* rv = val + static_cast<size_t>(r) * nnz;
  for (int c = threadIdx.x & 31; c < ld; c += 32) {
    float v = 0.f;
    if (c < nnz) v = __int_as_float(ri[c]);
    else if (c < 2 * nnz) v = rv[c - nnz];
    else if (c == 2 * nnz) v = 1.f;
    else if (c == 2 * nnz + 1) v = tgt[r];
    row[c] = v;
  }
}

// Wt[col][16] <- W[c][col], col in [0, d] (bias column included), classes beyond C zero
__global__ void __launch_bounds__(256)
mcpa_transpose_kernel(const float* __restrict__ W, int ldw, int d, int C, float* __restrict__ Wt) {
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  if (col > d) return;
  float v[16];
#pragma unroll
  for (int c = 0; c < 16; ++c) v[c] = c < C ? W[static_cast<size_t>(c) * ldw + col] : 0.f;
  float4* out = reinterpret_cast<float4*>(Wt + static_cast<size_t>(col) * 16);
#pragma unroll
  for (int g = 0; g < 4; ++g) out[g] = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
}

// coef[r] = (tau, y, rival, 0).  WT: scores from the class-minor copy (C <= 16), else from W[c][col]
template <bool WT>
__global__ void __launch_bounds__(256)
mcpa_sparse_coef_kernel(const float* __restrict__ rows, int ld, int B, int nnz, const float* __restrict__ W, int ldw,
                        int d, int C, int variant, float Creg, float4* __restrict__ coef, double* __restrict__ stats) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  __shared__ double s_l[8], s_m[8];
  __shared__ int s_n[8], s_a[8];
  double loss_d = 0.0, mist = 0.0;
  int nfit = 0, nact = 0;
  if (warp < B) {
    const float* row = rows + static_cast<size_t>(warp) * ld;
    const bool valid = row[2 * nnz] != 0.f;
    float tau = 0.f;
    int y = -1, rj = -1;
    if (valid) {
      float acc[SP_MAXC];
#pragma unroll
      for (int c = 0; c < SP_MAXC; ++c) acc[c] = 0.f;
      float n2 = 0.f;
      for (int j = lane; j < nnz; j += 32) {
        const int col = __float_as_int(row[j]);
        const float v = row[nnz + j];
        if (col >= 0 && col < d) {
          n2 = fmaf(v, v, n2);
          if (WT) {
            const float4* wt = reinterpret_cast<const float4*>(W + static_cast<size_t>(col) * 16);
#pragma unroll
            for (int g = 0; g < 4; ++g)
              if (4 * g < C) {
                const float4 w4 = wt[g];
                acc[4 * g] = fmaf(w4.x, v, acc[4 * g]);
                acc[4 * g + 1] = fmaf(w4.y, v, acc[4 * g + 1]);
                acc[4 * g + 2] = fmaf(w4.z, v, acc[4 * g + 2]);
                acc[4 * g + 3] = fmaf(w4.w, v, acc[4 * g + 3]);
              }
          } else {
#pragma unroll
            for (int c = 0; c < SP_MAXC; ++c)
              if (c < C) acc[c] = fmaf(W[static_cast<size_t>(c) * ldw + col], v, acc[c]);
          }
        }
      }
      n2 = warp_sum(n2) + 1.f;              // the bias feature contributes 1
#pragma unroll
      for (int c = 0; c < SP_MAXC; ++c)
        if (c < C) acc[c] = warp_sum(acc[c]) + (WT ? W[static_cast<size_t>(d) * 16 + c] : W[static_cast<size_t>(c) * ldw + d]);
      y = static_cast<int>(row[2 * nnz + 1] + 0.5f);
      float best = -CUDART_INF_F;
#pragma unroll
      for (int c = 0; c < SP_MAXC; ++c)
        if (c < C && c != y && acc[c] > best) { best = acc[c]; rj = c; }
      float sy = 0.f;
#pragma unroll
      for (int c = 0; c < SP_MAXC; ++c)
        if (c == y) sy = acc[c];
      const bool ok = y >= 0 && y < C && rj >= 0;
      const float loss = ok ? fmaxf(0.f, 1.f - (sy - best)) : 0.f;
      const float den = 2.f * n2;
      if (loss > 0.f && den > 0.f)
        tau = variant == 0 ? loss / den : variant == 1 ? fminf(Creg, loss / den) : loss / (den + 1.f / (2.f * Creg));
      loss_d = loss;
      mist = (best >= sy) ? 1.0 : 0.0;
      nfit = 1;
      nact = tau > 0.f ? 1 : 0;
    }
    if (lane == 0) coef[warp] = make_float4(tau, static_cast<float>(y), static_cast<float>(rj), 0.f);
  }
  const int w = threadIdx.x >> 5;
  if (lane == 0) { s_l[w] = loss_d; s_m[w] = mist; s_n[w] = nfit; s_a[w] = nact; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double l = 0, m = 0;
    int n = 0, a = 0;
    for (int i = 0; i < (blockDim.x >> 5); ++i) { l += s_l[i]; m += s_m[i]; n += s_n[i]; a += s_a[i]; }
    if (n) {
      atomicAdd(&stats[0], l);
      atomicAdd(&stats[1], m);
      atomicAdd(&stats[2], static_cast<double>(n));
      atomicAdd(&stats[3], static_cast<double>(a));
    }
  }
}

__global__ void __launch_bounds__(256)
mcpa_sparse_update_kernel(const float* __restrict__ rows, int ld, int B, int nnz, const float4* __restrict__ coef,
                          float* __restrict__ W, int ldw, int d, int C, const double* __restrict__ stats,
                          const double* __restrict__ active_before, int mean) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B) return;
  const float4 cf = coef[warp];
  if (!(cf.x > 0.f)) return;
  const int y = static_cast<int>(cf.y), rj = static_cast<int>(cf.z);
  if (y < 0 || y >= C || rj < 0 || rj >= C) return;
  float s = cf.x;
  if (mean) {
    const double act = stats[3] - *active_before;
    if (act > 0.0) s = static_cast<float>(static_cast<double>(s) / act);
  }
  const float* row = rows + static_cast<size_t>(warp) * ld;
  float* wy = W + static_cast<size_t>(y) * ldw;
  float* wr = W + static_cast<size_t>(rj) * ldw;
  for (int j = lane; j < nnz; j += 32) {
    const int col = __float_as_int(row[j]);
    if (col >= 0 && col < d) {
      const float g = s * row[nnz + j];
      atomicAdd(wy + col, g);
      atomicAdd(wr + col, -g);
    }
  }
  if (lane == 0) {
    atomicAdd(wy + d, s);
    atomicAdd(wr + d, -s);
  }
}

}  // namespace omldm

using namespace omldm;

extern "C" int omldm_sparse_rows(const int* idx, const float* val, const float* targets, int B, int nnz, float* rows,
                                 int ld, void* stream) {
  if (B <= 0) return 0;
  if (nnz <= 0 || ld < 2 * nnz + 2) return set_error_msg("sparse_rows: ld must be >= 2 * nnz + 2");
  sparse_rows_kernel<<<(B + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(idx, val, targets, B, nnz, rows, ld);
  return check_launch("sparse_rows_kernel");
}

extern "C" size_t omldm_mcpa_sparse_wt_floats(int d) { return static_cast<size_t>(d + 1) * 16; }

extern "C" int omldm_mcpa_sparse_step(const float* rows, int ld, int B, int nnz, float* W, int ldw, int d, int C,
                                      int variant, float Creg, float* coef, double* stats,
                                      const double* active_before, int mean_update, void* stream) {
  return omldm_mcpa_sparse_step_wt(rows, ld, B, nnz, W, ldw, d, C, variant, Creg, coef, stats, active_before,
                                   mean_update, nullptr, stream);
}

extern "C" int omldm_mcpa_sparse_step_wt(const float* rows, int ld, int B, int nnz, float* W, int ldw, int d, int C,
                                         int variant, float Creg, float* coef, double* stats,
                                         const double* active_before, int mean_update, float* wt, void* stream) {
  if (B <= 0) return 0;
  if (C < 2 || C > SP_MAXC) return set_error_msg("mcpa_sparse_step: 2 <= classes <= 32");
  if (nnz <= 0 || ld < 2 * nnz + 2 || ldw < d + 1) return set_error_msg("mcpa_sparse_step: bad row pitch");
  if (reinterpret_cast<uintptr_t>(coef) & 15) return set_error_msg("mcpa_sparse_step: coef must be 16-byte aligned");
  if (mean_update && !active_before) return set_error_msg("mcpa_sparse_step: averaged update needs active_before");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int blocks = (B + 7) / 8;
  if (wt && C <= 16) {
    if (reinterpret_cast<uintptr_t>(wt) & 15) return set_error_msg("mcpa_sparse_step: wt must be 16-byte aligned");
    mcpa_transpose_kernel<<<(d + 1 + 255) / 256, 256, 0, st>>>(W, ldw, d, C, wt);
    if (int rc = check_launch("mcpa_transpose_kernel")) return rc;
    mcpa_sparse_coef_kernel<true><<<blocks, 256, 0, st>>>(rows, ld, B, nnz, wt, 16, d, C, variant, Creg,
                                                          reinterpret_cast<float4*>(coef), stats);
  } else {
    mcpa_sparse_coef_kernel<false><<<blocks, 256, 0, st>>>(rows, ld, B, nnz, W, ldw, d, C, variant, Creg,
                                                           reinterpret_cast<float4*>(coef), stats);
  }
  if (int rc = check_launch("mcpa_sparse_coef_kernel")) return rc;
  mcpa_sparse_update_kernel<<<blocks, 256, 0, st>>>(rows, ld, B, nnz, reinterpret_cast<const float4*>(coef), W, ldw, d,
                                                    C, stats, active_before, mean_update);
  return check_launch("mcpa_sparse_update_kernel");
}
