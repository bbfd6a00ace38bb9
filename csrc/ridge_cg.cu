// Warm-started, Jacobi-preconditioned conjugate gradients for the ridge system of online ridge
// regression:  (A + lambda I) x = b,  A = the feature block of the Gram matrix (symmetric, full storage).
//
// Why: the weights of ORR are refreshed every few mini-batches.  A dense Cholesky of a 1000 x 1000
// system is a latency-bound chain of ~0.6 ms on any GPU, while between two refreshes A changes by a
// few mini-batches out of millions of records - the previous solution is an excellent starting point
// and a handful of CG iterations (one L2-resident 4 MB mat-vec each) restore the residual.
//
// One cooperative launch, a small grid (the refresh runs on a side stream beside the Gram GEMM and
// must not take its SMs), two grid-wide barriers per iteration:
//   phase M: every CTA rebuilds the search direction p = z + beta p_old in shared memory (z = M^-1 r),
//            its warps compute q_i = (A p)_i + lambda p_i for their rows, partial p.q -> one atomic per CTA
//   phase U: alpha = r.z / p.q;  x += alpha p;  r -= alpha q;  partial r.z and r.r -> atomics
// The reference has no counterpart (its ORR lives out of tree and solves on the JVM in fp64).
#include <cooperative_groups.h>

#include "host_common.h"
#include "omldm_kernels.h"
#include "ptx.cuh"

namespace cg = cooperative_groups;

namespace omldm {

constexpr int CG_THREADS = 256;
constexpr int CG_MAX_N = 4096;      // p lives in shared memory: 16 KB
constexpr int CG_MAX_ITER = 128;

struct __align__(32) CgScalars {
  double pq, rz, rr, bb;
};

__device__ __forceinline__ double block_sum(double v, double* sh) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();                      // sh may still be read from the previous reduction
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0.0;
  for (int i = 0; i < (CG_THREADS >> 5); ++i) t += sh[i];
  return t;
}

// vec layout in `work` (floats): r[n] | q[n] | minv[n] | p0[n] | p1[n]
__global__ void __launch_bounds__(CG_THREADS)
ridge_pcg_kernel(const float* __restrict__ A, int ld, int n, float lam, const float* __restrict__ b, int b_stride,
                 float* x, float* work, CgScalars* scal, int max_iter, float tol, float* info) {
  cg::grid_group grid = cg::this_grid();
  __shared__ float sp[CG_MAX_N];
  __shared__ double sred[CG_THREADS / 32];
  float* r = work;
  float* q = work + n;
  float* minv = work + 2 * n;
  float* pbuf[2] = {work + 3 * n, work + 4 * n};
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int gthreads = gridDim.x * blockDim.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gwarp = blockIdx.x * (CG_THREADS >> 5) + warp, gwarps = gridDim.x * (CG_THREADS >> 5);

  // ---- phase 0: clear the scalars, Jacobi preconditioner
  for (int i = gtid; i < (max_iter + 2) * 4; i += gthreads) reinterpret_cast<double*>(scal)[i] = 0.0;
  for (int i = gtid; i < n; i += gthreads) {
    const float dgl = A[static_cast<size_t>(i) * ld + i] + lam;
    minv[i] = dgl > 0.f ? 1.f / dgl : 1.f;
  }
  grid.sync();

  // ---- phase 1: r = b - (A + lam I) x,  z = M^-1 r,  p0 = z
  for (int i = threadIdx.x; i < n; i += blockDim.x) sp[i] = __ldcg(x + i);
  __syncthreads();
  {
    double l_rz = 0.0, l_rr = 0.0, l_bb = 0.0;
    for (int row = gwarp; row < n; row += gwarps) {
      const float* a = A + static_cast<size_t>(row) * ld;
      float acc = 0.f;
      for (int j = lane; j < n; j += 32) acc = fmaf(a[j], sp[j], acc);
      acc = warp_sum(acc);
      if (lane == 0) {
        const float bi = b[static_cast<size_t>(row) * b_stride];
        const float ri = bi - (acc + lam * sp[row]);
        const float zi = __ldcg(minv + row) * ri;
        r[row] = ri;
        pbuf[0][row] = zi;
        l_rz += static_cast<double>(ri) * zi;
        l_rr += static_cast<double>(ri) * ri;
        l_bb += static_cast<double>(bi) * bi;
      }
    }
    const double t_rz = block_sum(l_rz, sred), t_rr = block_sum(l_rr, sred), t_bb = block_sum(l_bb, sred);
    if (threadIdx.x == 0) {
      atomicAdd(&scal[0].rz, t_rz);
      atomicAdd(&scal[0].rr, t_rr);
      atomicAdd(&scal[0].bb, t_bb);
    }
  }
  grid.sync();
  const double bb = __ldcg(&scal[0].bb);
  const double stop = static_cast<double>(tol) * tol * (bb > 0.0 ? bb : 1.0);
  double rz = __ldcg(&scal[0].rz), rr = __ldcg(&scal[0].rr);
  int it = 0;
  // p of iteration k lives in pbuf[k & 1]; iteration 0 uses p0 = z as stored above (beta = 0)
  double beta = 0.0;
  while (it < max_iter && rr > stop) {
    float* pcur = pbuf[it & 1];
    const float* pold = pbuf[(it & 1) ^ 1];
    // ---- phase M
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      float pi;
      if (it == 0) {
        pi = __ldcg(pcur + i);
      } else {
        pi = __ldcg(minv + i) * __ldcg(r + i) + static_cast<float>(beta) * __ldcg(pold + i);
        if (blockIdx.x == 0) pcur[i] = pi;        // one CTA publishes the new direction for phase U
      }
      sp[i] = pi;
    }
    __syncthreads();
    double l_pq = 0.0;
    for (int row = gwarp; row < n; row += gwarps) {
      const float* a = A + static_cast<size_t>(row) * ld;
      float acc = 0.f;
      for (int j = lane; j < n; j += 32) acc = fmaf(a[j], sp[j], acc);
      acc = warp_sum(acc);
      if (lane == 0) {
        const float qi = acc + lam * sp[row];
        q[row] = qi;
        l_pq += static_cast<double>(qi) * sp[row];
      }
    }
    const double t_pq = block_sum(l_pq, sred);
    if (threadIdx.x == 0) atomicAdd(&scal[it + 1].pq, t_pq);
    grid.sync();
    // ---- phase U
    const double pq = __ldcg(&scal[it + 1].pq);
    if (!(pq > 0.0)) break;                        // not positive definite in fp32: give up, the caller falls back
    const float alpha = static_cast<float>(rz / pq);
    double l_rz = 0.0, l_rr = 0.0;
    for (int i = gtid; i < n; i += gthreads) {
      const float pi = __ldcg(pcur + i);
      x[i] = __ldcg(x + i) + alpha * pi;
      const float ri = __ldcg(r + i) - alpha * __ldcg(q + i);
      r[i] = ri;
      l_rz += static_cast<double>(ri) * (__ldcg(minv + i) * ri);
      l_rr += static_cast<double>(ri) * ri;
    }
    const double t_rz = block_sum(l_rz, sred), t_rr = block_sum(l_rr, sred);
    if (threadIdx.x == 0) {
      atomicAdd(&scal[it + 1].rz, t_rz);
      atomicAdd(&scal[it + 1].rr, t_rr);
    }
    grid.sync();
    const double rz_new = __ldcg(&scal[it + 1].rz);
    rr = __ldcg(&scal[it + 1].rr);
    beta = rz > 0.0 ? rz_new / rz : 0.0;
    rz = rz_new;
    ++it;
  }
  if (gtid == 0) {
    info[0] = static_cast<float>(it);
    info[1] = static_cast<float>(sqrt(rr / (bb > 0.0 ? bb : 1.0)));
  }
}

}  // namespace omldm

using namespace omldm;

extern "C" size_t omldm_ridge_cg_work_floats(int n) { return static_cast<size_t>(5) * n; }
extern "C" size_t omldm_ridge_cg_scalar_bytes(int max_iter) { return sizeof(CgScalars) * (static_cast<size_t>(max_iter) + 2); }

extern "C" int omldm_ridge_cg(const float* A, int ld, int n, float lam, const float* b, int b_stride, float* x,
                              float* work, void* scalars, int max_iter, float tol, float* info, int blocks,
                              void* stream) {
  if (n <= 0) return 0;
  if (n > CG_MAX_N) return set_error_msg("ridge_cg: n exceeds 4096");
  if (max_iter < 1 || max_iter > CG_MAX_ITER) return set_error_msg("ridge_cg: max_iter must be in [1, 128]");
  if (reinterpret_cast<uintptr_t>(scalars) & 31) return set_error_msg("ridge_cg: scalars must be 32-byte aligned");
  int dev = 0, sms = 0, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  // the refresh runs on a side stream while the Gram kernel (198 KB of shared memory per CTA) owns the SMs: keep every SM
  // it touches at the largest shared-memory carve-out, or a Gram CTA waits until the refresh blocks have left that SM
  static bool carveout_set[16] = {};
  if (!carveout_set[dev & 15]) {
    cudaFuncSetAttribute(ridge_pcg_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    carveout_set[dev & 15] = true;
  }
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ridge_pcg_kernel, CG_THREADS, 0);
  int grid = blocks > 0 ? blocks : 32;
  if (grid > sms * per_sm) grid = sms * per_sm;
  if (grid < 1) return set_error_msg("ridge_cg: kernel cannot be resident");
  CgScalars* sc = static_cast<CgScalars*>(scalars);
#ifdef OMLDM_EMU   // tools/cuda_emu: a few blocks run concurrently as host threads, the grid barrier spans all of them
  emu::launcher_concurrent(ridge_pcg_kernel, dim3(blocks > 0 && blocks <= 8 ? blocks : 2), dim3(CG_THREADS), 0, nullptr)(
      A, ld, n, lam, b, b_stride, x, work, sc, max_iter, tol, info);
  (void)grid;
  return check_launch("ridge_pcg_kernel");
#else
  void* args[] = {&A, &ld, &n, &lam, &b, &b_stride, &x, &work, &sc, &max_iter, &tol, &info};
  cudaError_t e = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(ridge_pcg_kernel), dim3(grid), dim3(CG_THREADS),
                                              args, 0, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return set_error("ridge_pcg_kernel", e);
  count_launch(1);
  return 0;
#endif
}
