// Element-wise glue of the small MLP learner ("NN", whitelist PipelineMap.scala:68; the reference's implementation
// lives out of tree).  The matrix products run on the tcgen05 GEMMs (umma_gemm.cu); what is left between them is fused
// here so that a training step is a handful of launches instead of a few dozen framework kernels:
//   nn_hidden_act   : H <- relu(H) on the first `od` columns, constant-1 column `od` (the next layer's bias input)
//   nn_out_grad     : output layer -> loss statistics and the gradient rows dL/d(out), invalid rows get zero;
//                     soft-max cross entropy for classifiers, squared error for regressors; counts the valid rows
//   nn_relu_mask    : g <- g * (activation > 0) for the backward pass through a hidden layer
//   nn_sgd_step     : W <- W - (lr / max(n_valid, 1)) * gW with n_valid read from device memory (no host round trip)
#include <math_constants.h>

#include "host_common.h"
#include "omldm_kernels.h"
#include "ptx.cuh"

namespace omldm {

__global__ void nn_hidden_act_kernel(float* __restrict__ H, int ldh, int B, int od) {
  const size_t total = static_cast<size_t>(B) * (od + 1);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t r = i / (od + 1);
    const int c = static_cast<int>(i % (od + 1));
    float* p = H + r * ldh + c;
    *p = c == od ? 1.f : fmaxf(*p, 0.f);
  }
}

// One warp per record.  out: B x ldo (first `od` columns are the network outputs); Z: staged rows (valid flag in
// column d, target in d + 1); G: B x ldg gradient rows (columns >= od are cleared); stats: (loss, -, fitted, -);
// n_valid: device float, this call's number of valid rows is ADDED to it (the caller clears it per step).
__global__ void __launch_bounds__(256)
nn_out_grad_kernel(const float* __restrict__ out, int ldo, const float* __restrict__ Z, int ld, int d, int B, int od,
                   int classes, float* __restrict__ G, int ldg, double* __restrict__ stats, float* __restrict__ n_valid) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  __shared__ double s_loss[8];
  __shared__ int s_n[8];
  double loss = 0.0;
  int nfit = 0;
  if (warp < B) {
    const float* o = out + static_cast<size_t>(warp) * ldo;
    const float* z = Z + static_cast<size_t>(warp) * ld;
    float* g = G + static_cast<size_t>(warp) * ldg;
    const bool valid = z[d] != 0.f;
    const float y = z[d + 1];
    if (classes > 1) {
      float m = -CUDART_INF_F;
      for (int c = lane; c < od; c += 32) m = fmaxf(m, o[c]);
      m = warp_max(m);
      float s = 0.f;
      for (int c = lane; c < od; c += 32) s += expf(o[c] - m);
      s = warp_sum(s);
      int yi = static_cast<int>(rintf(y));
      yi = yi < 0 ? 0 : (yi >= od ? od - 1 : yi);
      for (int c = lane; c < ldg; c += 32) {
        float v = 0.f;
        if (c < od && valid) v = expf(o[c] - m) / s - (c == yi ? 1.f : 0.f);
        g[c] = v;
      }
      if (valid) {
        const float py = fmaxf(expf(o[yi] - m) / s, 1e-30f);
        loss = -logf(py);
        nfit = 1;
      }
    } else {
      const float err = o[0] - y;
      for (int c = lane; c < ldg; c += 32) g[c] = (c == 0 && valid) ? 2.f * err : 0.f;
      if (valid) { loss = static_cast<double>(err) * err; nfit = 1; }
    }
  }
  const int w = threadIdx.x >> 5;
  if (lane == 0) { s_loss[w] = loss; s_n[w] = nfit; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double l = 0.0;
    int n = 0;
    for (int i = 0; i < (blockDim.x >> 5); ++i) { l += s_loss[i]; n += s_n[i]; }
    if (n) {
      atomicAdd(&stats[0], l);
      atomicAdd(&stats[2], static_cast<double>(n));
      atomicAdd(n_valid, static_cast<float>(n));
    }
  }
}

__global__ void nn_relu_mask_kernel(float* __restrict__ g, int ldg, const float* __restrict__ act, int lda, int B, int n) {
  const size_t total = static_cast<size_t>(B) * n;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t r = i / n, c = i % n;
    if (!(act[r * lda + c] > 0.f)) g[r * ldg + c] = 0.f;
  }
}

__global__ void nn_sgd_step_kernel(float* __restrict__ W, const float* __restrict__ gW, size_t n, float lr,
                                   const float* __restrict__ n_valid) {
  const float s = lr / fmaxf(*n_valid, 1.f);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x)
    W[i] = fmaf(-s, gW[i], W[i]);
}

static int grid_for(size_t total) {
  size_t b = (total + 255) / 256;
  return static_cast<int>(b < 1 ? 1 : (b > 148 * 8 ? 148 * 8 : b));
}

}  // namespace omldm

using namespace omldm;

extern "C" int omldm_nn_hidden_act(float* H, int ldh, int B, int od, void* stream) {
  if (B <= 0) return 0;
  if (ldh < od + 1) return set_error_msg("nn_hidden_act: the activation rows need room for the constant-1 column");
  nn_hidden_act_kernel<<<grid_for(static_cast<size_t>(B) * (od + 1)), 256, 0, static_cast<cudaStream_t>(stream)>>>(H, ldh, B, od);
  return check_launch("nn_hidden_act_kernel");
}

extern "C" int omldm_nn_out_grad(const float* out, int ldo, const float* Z, int ld, int d, int B, int od, int classes,
                                 float* G, int ldg, double* stats, float* n_valid, void* stream) {
  if (B <= 0) return 0;
  if (od < 1 || ldo < od || ldg < od) return set_error_msg("nn_out_grad: bad output width / pitch");
  nn_out_grad_kernel<<<(B + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(out, ldo, Z, ld, d, B, od, classes, G, ldg,
                                                                                stats, n_valid);
  return check_launch("nn_out_grad_kernel");
}

extern "C" int omldm_nn_relu_mask(float* g, int ldg, const float* act, int lda, int B, int n, void* stream) {
  if (B <= 0 || n <= 0) return 0;
  nn_relu_mask_kernel<<<grid_for(static_cast<size_t>(B) * n), 256, 0, static_cast<cudaStream_t>(stream)>>>(g, ldg, act, lda,
                                                                                                           B, n);
  return check_launch("nn_relu_mask_kernel");
}

extern "C" int omldm_nn_sgd_step(float* W, const float* gW, size_t n, float lr, const float* n_valid, void* stream) {
  if (n == 0) return 0;
  nn_sgd_step_kernel<<<grid_for(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(W, gW, n, lr, n_valid);
  return check_launch("nn_sgd_step_kernel");
}
