// Direct solve of the ridge system of online ridge regression, (A + lambda I) w = b, by a blocked right-looking
// Cholesky factorisation in ONE cooperative kernel - the fall-back of the conjugate-gradient refresh
// (ridge_cg.cu) and the query-time solve (the model read behind node.receiveQuery,
// src/main/scala/omldm/operators/spoke/FlinkSpoke.scala:163).  Replaces the library potrf / potrs pair on this
// path; the reference has no native code.
//
// Input: the (n + 1) x (n + 1) leading block of the mirrored Gram work matrix W (pitch ld): rows / columns 0..n-1 are
// A, row n is b^T.  The right-hand side rides along as that extra ROW: after the factorisation A = L L^T the row holds
// z^T = b^T L^-T, i.e. the forward substitution is done by the same triangular solves as every other row, and only
// the back substitution L^T w = z is left (one CTA, L streams once from L2).  W is destroyed (lower triangle <- L).
//
// Per 32-column panel: every CTA factors the 32 x 32 diagonal block itself in shared memory (no barrier needed to
// share it), solves its rows of the panel against it, grid barrier, rank-32 update of the trailing lower triangle
// in 32 x 32 tiles over all CTAs, grid barrier.
#include <cooperative_groups.h>

#include "host_common.h"
#include "omldm_kernels.h"
#include "ptx.cuh"

namespace cg = cooperative_groups;

namespace omldm {

constexpr int CH_NB = 32;
constexpr int CH_THREADS = 256;

// In-place Cholesky of the nb x nb block in s (lower triangle), all 256 threads; returns false on a non-positive pivot.
__device__ __forceinline__ bool chol_block(float (*s)[CH_NB + 1], int nb, int* s_bad) {
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int k = 0; k < nb; ++k) {
    __syncthreads();
    const float dkk = s[k][k];
    if (!(dkk > 0.f)) { if (threadIdx.x == 0) *s_bad = 1; }
    const float inv = dkk > 0.f ? 1.f / sqrtf(dkk) : 0.f;
    __syncthreads();
    if (threadIdx.x < nb && threadIdx.x >= k) s[threadIdx.x][k] *= inv;      // column k (and the pivot = sqrt)
    __syncthreads();
    for (int i = k + 1 + ty; i < nb; i += CH_THREADS / 32) {                 // trailing rank-1 update, lower part
      const int c = k + 1 + tx;
      if (c <= i) s[i][c] -= s[i][k] * s[c][k];
    }
  }
  __syncthreads();
  return *s_bad == 0;
}

__global__ void __launch_bounds__(CH_THREADS)
ridge_chol_kernel(float* __restrict__ W, int ld, int n, float lam, float* __restrict__ x, float* __restrict__ info) {
  cg::grid_group grid = cg::this_grid();
  __shared__ float s_d[CH_NB][CH_NB + 1];
  __shared__ float s_p[CH_NB][CH_NB + 1];
  __shared__ float s_q[CH_NB][CH_NB + 1];
  __shared__ int s_bad;
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int gthreads = gridDim.x * blockDim.x;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_bad = 0;
  for (int i = gtid; i < n; i += gthreads) W[static_cast<size_t>(i) * ld + i] += lam;
  grid.sync();

  const int rows = n + 1;                                    // A plus the right-hand-side row
  for (int j0 = 0; j0 < n; j0 += CH_NB) {
    const int nb = min(CH_NB, n - j0);
    // ---- (1) diagonal block, factored redundantly by every CTA
    for (int t = threadIdx.x; t < CH_NB * CH_NB; t += CH_THREADS) {
      const int i = t >> 5, c = t & 31;
      s_d[i][c] = (i < nb && c <= i) ? __ldcg(W + static_cast<size_t>(j0 + i) * ld + j0 + c) : (i == c ? 1.f : 0.f);
    }
    chol_block(s_d, nb, &s_bad);
    if (blockIdx.x == 0)
      for (int t = threadIdx.x; t < CH_NB * CH_NB; t += CH_THREADS) {
        const int i = t >> 5, c = t & 31;
        if (i < nb && c <= i) W[static_cast<size_t>(j0 + i) * ld + j0 + c] = s_d[i][c];
      }
    // ---- (2) panel rows below the block: X L11^T = A21, one thread per row
    for (int r = j0 + nb + gtid; r < rows; r += gthreads) {
      float* a = W + static_cast<size_t>(r) * ld + j0;
      float v[CH_NB];
#pragma unroll
      for (int k = 0; k < CH_NB; ++k) v[k] = k < nb ? __ldcg(a + k) : 0.f;
#pragma unroll
      for (int k = 0; k < CH_NB; ++k) {
        float acc = v[k];
#pragma unroll
        for (int m = 0; m < k; ++m) acc = fmaf(-v[m], s_d[k][m], acc);
        v[k] = acc / s_d[k][k];                              // rows >= nb of s_d are the identity
      }
#pragma unroll
      for (int k = 0; k < CH_NB; ++k)
        if (k < nb) a[k] = v[k];
    }
    grid.sync();
    // ---- (3) trailing update: C[bi, bj] -= L21[bi] L21[bj]^T for the lower tiles of what is left
    const int r0 = j0 + nb;
    if (r0 < rows) {
      const int RB = (rows - r0 + CH_NB - 1) / CH_NB;
      const int ntiles = RB * (RB + 1) / 2;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        int bi = 0, rem = t;
        while (rem > bi) { rem -= bi + 1; ++bi; }            // t -> (bi >= bj) in row-major lower order
        const int bj = rem;
        __syncthreads();
        for (int u = threadIdx.x; u < CH_NB * CH_NB; u += CH_THREADS) {
          const int i = u >> 5, c = u & 31;
          const int ri = r0 + bi * CH_NB + i, rj = r0 + bj * CH_NB + i;
          s_p[i][c] = (ri < rows && c < nb) ? __ldcg(W + static_cast<size_t>(ri) * ld + j0 + c) : 0.f;
          s_q[i][c] = (rj < rows && c < nb) ? __ldcg(W + static_cast<size_t>(rj) * ld + j0 + c) : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int u = 0; u < CH_NB / 8; ++u) {
          const int i = ty + 8 * u;
          const int ri = r0 + bi * CH_NB + i, cj = r0 + bj * CH_NB + tx;
          if (ri < rows && cj < n && cj <= ri) {
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k < CH_NB; ++k) acc = fmaf(s_p[i][k], s_q[tx][k], acc);
            float* cp = W + static_cast<size_t>(ri) * ld + cj;
            *cp = __ldcg(cp) - acc;
          }
        }
      }
    }
    grid.sync();
  }

  // ---- back substitution L^T w = z (z = row n), one CTA; w lives in x
  if (blockIdx.x == 0) {
    for (int i = threadIdx.x; i < n; i += CH_THREADS) x[i] = __ldcg(W + static_cast<size_t>(n) * ld + i);
    __syncthreads();
    const int last = (n - 1) / CH_NB * CH_NB;
    for (int k0 = last; k0 >= 0; k0 -= CH_NB) {
      const int nb = min(CH_NB, n - k0);
      for (int t = threadIdx.x; t < CH_NB * CH_NB; t += CH_THREADS) {
        const int i = t >> 5, c = t & 31;
        s_d[i][c] = (i < nb && c <= i) ? __ldcg(W + static_cast<size_t>(k0 + i) * ld + k0 + c) : 0.f;
      }
      __syncthreads();
      if (threadIdx.x < 32) {                               // 32 x 32 triangle by one warp, one unknown per lane
        float t = tx < nb ? x[k0 + tx] : 0.f;
        for (int m = nb - 1; m >= 0; --m) {
          float wm = 0.f;
          if (tx == m) { wm = t / s_d[m][m]; t = wm; }
          wm = __shfl_sync(0xffffffffu, wm, m);
          if (tx < m) t = fmaf(-s_d[m][tx], wm, t);
        }
        if (tx < nb) x[k0 + tx] = t;
      }
      __syncthreads();
      for (int c = threadIdx.x; c < k0; c += CH_THREADS) {   // fold the solved block into the rows above
        float acc = x[c];
#pragma unroll 8
        for (int m = 0; m < nb; ++m) acc = fmaf(-__ldcg(W + static_cast<size_t>(k0 + m) * ld + c), x[k0 + m], acc);
        x[c] = acc;
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) info[0] = s_bad ? 1.f : 0.f;
  }
}

}  // namespace omldm

using namespace omldm;

extern "C" int omldm_ridge_chol(float* W, int ld, int n, float lam, float* x, float* info, int blocks, void* stream) {
  if (n <= 0) return 0;
  if (n + 1 > ld) return set_error_msg("ridge_chol: the work matrix needs n + 1 columns");
  int dev = 0, sms = 0, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  static bool carveout_set[16] = {};     // same reason as in ridge_cg.cu: never make a Gram CTA wait for an SM to drain
  if (!carveout_set[dev & 15]) {
    cudaFuncSetAttribute(ridge_chol_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    carveout_set[dev & 15] = true;
  }
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ridge_chol_kernel, CH_THREADS, 0);
  int grid = blocks > 0 ? blocks : 64;
  if (grid > sms * per_sm) grid = sms * per_sm;
  if (grid < 1) return set_error_msg("ridge_chol: kernel cannot be resident");
#ifdef OMLDM_EMU
  emu::launcher_concurrent(ridge_chol_kernel, dim3(blocks > 0 && blocks <= 8 ? blocks : 2), dim3(CH_THREADS), 0, nullptr)(
      W, ld, n, lam, x, info);
  (void)grid;
  return check_launch("ridge_chol_kernel");
#else
  void* args[] = {&W, &ld, &n, &lam, &x, &info};
  cudaError_t e = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(ridge_chol_kernel), dim3(grid), dim3(CH_THREADS), args,
                                              0, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return set_error("ridge_chol_kernel", e);
  count_launch(1);
  return 0;
#endif
}
