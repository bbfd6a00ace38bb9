// Linear online learners on staged record rows: PA classifier, PA regressor, Pegasos SVM.
//
// Staged record row ("Z row", pitch `ld` floats, multiple of 32):
//     [ x_0 .. x_{d-1} | 1.0 (bias / valid flag; 0.0 marks an empty row) | target | zero pad ]
// The weight vector has d+1 entries: w[d] is the intercept (bias handled as a constant feature).
//
// Two execution modes per learner:
//   exact     : one warp per pipeline walks the batch record by record, weights in registers -
//               the reference's per-record `receiveTuple` semantics
//               (src/main/scala/omldm/operators/spoke/FlinkSpoke.scala:101), many pipelines
//               (multi-tenant, FlinkSpoke.scala:97-105) advance in one launch.
//   minibatch : all margins against the frozen weights, one averaged update per batch.
//
// Formulas are the literature ones (Crammer et al. 2006; Shalev-Shwartz et al. 2007); the
// reference tree contains only the learner names (PipelineMap.scala:68).
#include <atomic>
#include "ptx.cuh"
#include "host_common.h"
#include "omldm_kernels.h"
#include "linear_math.cuh"

namespace omldm {

template <int R>
__global__ void __launch_bounds__(32)
linear_exact_kernel(const float* __restrict__ Z, int ld, int B, const omldm_linear_desc* descs) {
  const omldm_linear_desc D = descs[blockIdx.x];
  const int lane = threadIdx.x;
  const int dw = D.d + 1;  // features + bias column
  float w[R], xn[R], x[R];
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const int c = lane + 32 * j;
    w[j] = c < dw ? D.w[c] : 0.f;
    xn[j] = (B > 0 && c < dw) ? Z[c] : 0.f;
  }
  float yn = B > 0 ? Z[D.d + 1] : 0.f;
  float vn = B > 0 ? Z[D.d] : 0.f;
  double loss_sum = 0.0, aux_sum = 0.0;
  long long fitted = 0;
  long long t_svm = static_cast<long long>(D.stats[3]);
  for (int t = 0; t < B; ++t) {
    const float y = yn, valid = vn;
#pragma unroll
    for (int j = 0; j < R; ++j) x[j] = xn[j];
    if (t + 1 < B) {
      const float* zr = Z + static_cast<size_t>(t + 1) * ld;
#pragma unroll
      for (int j = 0; j < R; ++j) {
        const int c = lane + 32 * j;
        xn[j] = c < dw ? zr[c] : 0.f;
      }
      yn = zr[D.d + 1];
      vn = zr[D.d];
    }
    if (valid == 0.f) continue;  // empty row (holdout slot not yet evicting)
    float dot = 0.f, nrm = 0.f;
#pragma unroll
    for (int j = 0; j < R; ++j) {
      dot = fmaf(w[j], x[j], dot);
      nrm = fmaf(x[j], x[j], nrm);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      dot += __shfl_xor_sync(0xffffffffu, dot, o);
      nrm += __shfl_xor_sync(0xffffffffu, nrm, o);
    }
    float loss, aux;
    float c = linear_coef(D.kind, D.variant, D.C, D.eps, dot, y, nrm, loss, aux);
    if (D.kind == 2) {
      t_svm += 1;
      const float eta = 1.f / (D.lambda * static_cast<float>(t_svm));
      const float shrink = 1.f - eta * D.lambda;
      const float yy = y > 0.f ? 1.f : -1.f;
      const float cc = (yy * dot < 1.f) ? eta * yy : 0.f;
#pragma unroll
      for (int j = 0; j < R; ++j) w[j] = fmaf(cc, x[j], shrink * w[j]);
    } else if (c != 0.f) {
#pragma unroll
      for (int j = 0; j < R; ++j) w[j] = fmaf(c, x[j], w[j]);
    }
    loss_sum += loss;
    aux_sum += aux;
    fitted += 1;
  }
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const int c = lane + 32 * j;
    if (c < dw) D.w[c] = w[j];
  }
  if (lane == 0) {
    D.stats[0] += loss_sum;
    D.stats[1] += aux_sum;
    D.stats[2] += static_cast<double>(fitted);
    D.stats[3] = static_cast<double>(t_svm);
  }
}

// ---------------------------------------------------------------- minibatch
// scratch layout (floats): [0, dw) delta accumulators | dw: active count | dw+1: blocks done (as u32)
constexpr int MB_WARPS = 8;

// Block epilogue shared by the mini-batch kernels: per-warp statistics and the block's delta go to the global
// accumulator; the last block applies the averaged (or summed) update and clears the scratch for the next call.
__device__ __forceinline__ void minibatch_finish(const omldm_linear_desc& D, float* __restrict__ scratch, int B,
                                                 int mean_update, float svm_eta, float* s_delta, double loss_sum,
                                                 double aux_sum, int fitted, int active) {
  __shared__ double s_loss[MB_WARPS], s_aux[MB_WARPS];
  __shared__ int s_fit[MB_WARPS], s_act[MB_WARPS];
  __shared__ bool s_last;
  const int dw = D.d + 1;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { s_loss[warp] = loss_sum; s_aux[warp] = aux_sum; s_fit[warp] = fitted; s_act[warp] = active; }
  __syncthreads();
  for (int c = threadIdx.x; c < dw; c += blockDim.x) {
    const float v = s_delta[c];
    if (v != 0.f) atomicAdd(&scratch[c], v);
  }
  // every thread's additions must be out before thread 0 announces this block as done: without the barrier the last
  // block could read (and then clear) the accumulator while a slower warp of this block is still adding to it
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    double l = 0, a = 0; int f = 0, ac = 0;
    for (int i = 0; i < MB_WARPS; ++i) { l += s_loss[i]; a += s_aux[i]; f += s_fit[i]; ac += s_act[i]; }
    atomicAdd(&D.stats[0], l);
    atomicAdd(&D.stats[1], a);
    atomicAdd(&D.stats[2], static_cast<double>(f));
    if (ac) atomicAdd(&scratch[dw], static_cast<float>(ac));
    __threadfence();
    const unsigned done = atomicAdd(reinterpret_cast<unsigned*>(&scratch[dw + 1]), 1u);
    s_last = (done == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    // last block applies the averaged (or summed) update and resets the scratch for the next call
    __threadfence();
    const float n_active = __ldcg(&scratch[dw]);
    float scale = 1.f;
    if (mean_update && D.kind != 2) scale = n_active > 0.f ? 1.f / n_active : 0.f;
    if (D.kind == 2) {
      // Pegasos mini-batch step: w <- (1 - eta*lambda) w + (eta / B) sum_{violators} y x
      const float shrink = 1.f - svm_eta * D.lambda;
      const float inv_b = 1.f / fmaxf(1.f, static_cast<float>(B));
      for (int c = threadIdx.x; c < dw; c += blockDim.x) {
        D.w[c] = shrink * D.w[c] + inv_b * __ldcg(&scratch[c]);
        scratch[c] = 0.f;
      }
      if (threadIdx.x == 0) D.stats[3] += 1.0;
    } else {
      for (int c = threadIdx.x; c < dw; c += blockDim.x) {
        D.w[c] += scale * __ldcg(&scratch[c]);
        scratch[c] = 0.f;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) { scratch[dw] = 0.f; scratch[dw + 1] = 0.f; }
  }
}

__global__ void __launch_bounds__(MB_WARPS * 32)
linear_minibatch_kernel(const float* __restrict__ Z, int ld, int B, omldm_linear_desc D,
                        float* __restrict__ scratch, int mean_update) {
  extern __shared__ float s_delta[];  // dw floats
  const int dw = D.d + 1;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int c = threadIdx.x; c < dw; c += blockDim.x) s_delta[c] = 0.f;
  __syncthreads();
  double loss_sum = 0.0, aux_sum = 0.0;
  int fitted = 0, active = 0;
  const float svm_eta = (D.kind == 2) ? 1.f / (D.lambda * static_cast<float>(D.stats[3] + 1.0)) : 0.f;
  for (int r = blockIdx.x * MB_WARPS + warp; r < B; r += gridDim.x * MB_WARPS) {
    const float* zr = Z + static_cast<size_t>(r) * ld;
    if (zr[D.d] == 0.f) continue;
    float dot = 0.f, nrm = 0.f;
    for (int c = lane; c < dw; c += 32) {
      const float xv = zr[c];
      dot = fmaf(D.w[c], xv, dot);
      nrm = fmaf(xv, xv, nrm);
    }
    dot = warp_sum(dot);
    nrm = warp_sum(nrm);
    const float y = zr[D.d + 1];
    float loss, aux;
    float c = linear_coef(D.kind, D.variant, D.C, D.eps, dot, y, nrm, loss, aux);
    if (D.kind == 2) {
      const float yy = y > 0.f ? 1.f : -1.f;
      c = (yy * dot < 1.f) ? svm_eta * yy : 0.f;
    }
    if (c != 0.f) {
      for (int col = lane; col < dw; col += 32) atomicAdd(&s_delta[col], c * zr[col]);
      active += 1;
    }
    loss_sum += loss;
    aux_sum += aux;
    fitted += 1;
  }
  minibatch_finish(D, scratch, B, mean_update, svm_eta, s_delta, loss_sum, aux_sum, fitted, active);
}

// Thin rows (pitch 32 floats, d <= 30; BASELINE config 1 is d = 28): one LANE per record instead of one warp.  A lane
// holds its whole 128-byte row in registers (consecutive lanes = consecutive rows, so a warp reads 4 KB contiguous),
// needs no shuffle for the margin, and accumulates its share of the update in registers; the 32 partial sums are
// warp-reduced once at the end.  Same arithmetic and the same block epilogue as the general kernel.
__global__ void __launch_bounds__(MB_WARPS * 32)
linear_minibatch_thin_kernel(const float* __restrict__ Z, int B, omldm_linear_desc D, float* __restrict__ scratch,
                             int mean_update) {
  __shared__ float s_delta[32];
  __shared__ float s_w[32];
  const int dw = D.d + 1;
  if (threadIdx.x < 32) {
    s_delta[threadIdx.x] = 0.f;
    s_w[threadIdx.x] = threadIdx.x < dw ? D.w[threadIdx.x] : 0.f;
  }
  __syncthreads();
  double loss_sum = 0.0, aux_sum = 0.0;
  int fitted = 0, active = 0;
  const float svm_eta = (D.kind == 2) ? 1.f / (D.lambda * static_cast<float>(D.stats[3] + 1.0)) : 0.f;
  float acc[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) acc[j] = 0.f;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < B; r += gridDim.x * blockDim.x) {
    const float4* z4 = reinterpret_cast<const float4*>(Z + static_cast<size_t>(r) * 32);
    float x[32];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 t = ld_nc_v4(z4 + j);
      x[4 * j] = t.x; x[4 * j + 1] = t.y; x[4 * j + 2] = t.z; x[4 * j + 3] = t.w;
    }
    float valid = 0.f, y = 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {          // row[d] = valid flag, row[d + 1] = target; columns beyond the bias do not
      if (j == D.d) valid = x[j];           // take part in the margin
      if (j == D.d + 1) y = x[j];
      if (j >= dw) x[j] = 0.f;
    }
    if (valid == 0.f) continue;
    float dot = 0.f, nrm = 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      dot = fmaf(s_w[j], x[j], dot);
      nrm = fmaf(x[j], x[j], nrm);
    }
    float loss, aux;
    float c = linear_coef(D.kind, D.variant, D.C, D.eps, dot, y, nrm, loss, aux);
    if (D.kind == 2) {
      const float yy = y > 0.f ? 1.f : -1.f;
      c = (yy * dot < 1.f) ? svm_eta * yy : 0.f;
    }
    if (c != 0.f) {
#pragma unroll
      for (int j = 0; j < 32; ++j) acc[j] = fmaf(c, x[j], acc[j]);
      active += 1;
    }
    loss_sum += loss;
    aux_sum += aux;
    fitted += 1;
  }
  // per-warp totals (the epilogue expects them in lane 0) and the warp's share of the update
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    loss_sum += __shfl_xor_sync(0xffffffffu, loss_sum, o);
    aux_sum += __shfl_xor_sync(0xffffffffu, aux_sum, o);
    fitted += __shfl_xor_sync(0xffffffffu, fitted, o);
    active += __shfl_xor_sync(0xffffffffu, active, o);
  }
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const float t = warp_sum(acc[j]);
    if (lane == 0 && j < dw && t != 0.f) atomicAdd(&s_delta[j], t);
  }
  minibatch_finish(D, scratch, B, mean_update, svm_eta, s_delta, loss_sum, aux_sum, fitted, active);
}

// Predictions f = w . [x, 1] for a batch (forecast path, FlinkSpoke.scala:105).
__global__ void linear_predict_kernel(const float* __restrict__ Z, int ld, int B, const float* __restrict__ w,
                                      int d, float* __restrict__ out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B) return;
  const float* zr = Z + static_cast<size_t>(warp) * ld;
  float dot = 0.f;
  for (int c = lane; c < d; c += 32) dot = fmaf(w[c], zr[c], dot);
  dot = warp_sum(dot);
  if (lane == 0) out[warp] = dot + w[d];
}

}  // namespace omldm

using namespace omldm;

extern "C" int omldm_linear_exact(const float* Z, int ld, int B, const omldm_linear_desc* descs_dev,
                                  int n_pipelines, int max_d, void* stream) {
  if (n_pipelines <= 0 || B <= 0) return 0;
  const int dw = max_d + 1;
  const int R = (dw + 31) / 32;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define LAUNCH_R(RR) linear_exact_kernel<RR><<<n_pipelines, 32, 0, st>>>(Z, ld, B, descs_dev)
  if (R <= 1) LAUNCH_R(1);
  else if (R <= 2) LAUNCH_R(2);
  else if (R <= 4) LAUNCH_R(4);
  else if (R <= 8) LAUNCH_R(8);
  else if (R <= 16) LAUNCH_R(16);
  else if (R <= 32) LAUNCH_R(32);
  else return set_error_msg("linear_exact: d+1 > 1024 is not supported by the register-resident kernel");
#undef LAUNCH_R
  return check_launch("linear_exact_kernel");
}

extern "C" size_t omldm_linear_minibatch_scratch_floats(int d) { return static_cast<size_t>(d) + 1 + 2; }

extern "C" int omldm_linear_minibatch(const float* Z, int ld, int B, const omldm_linear_desc* desc_host,
                                      float* scratch, int mean_update, void* stream) {
  if (B <= 0) return 0;
  const int dw = desc_host->d + 1;
  if (dw * 4 > 200 * 1024) return set_error_msg("linear_minibatch: d too large for the smem accumulator");
  int blocks = (B + MB_WARPS - 1) / MB_WARPS;
  const int cap = 148 * 4;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dw * 4 > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(linear_minibatch_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, dw * 4);
    if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(linear_minibatch)", e);
  }
  if (ld == 32 && dw + 1 <= 32 && (reinterpret_cast<uintptr_t>(Z) & 15) == 0) {      // thin rows: one lane per record
    int tb = (B + MB_WARPS * 32 - 1) / (MB_WARPS * 32);
    if (tb > 148 * 2) tb = 148 * 2;
    linear_minibatch_thin_kernel<<<tb, MB_WARPS * 32, 0, st>>>(Z, B, *desc_host, scratch, mean_update);
    return check_launch("linear_minibatch_thin_kernel");
  }
  linear_minibatch_kernel<<<blocks, MB_WARPS * 32, dw * 4, st>>>(Z, ld, B, *desc_host, scratch, mean_update);
  return check_launch("linear_minibatch_kernel");
}

extern "C" int omldm_linear_predict(const float* Z, int ld, int B, const float* w, int d, float* out,
                                    void* stream) {
  if (B <= 0) return 0;
  const int threads = 256;
  const int blocks = (B * 32 + threads - 1) / threads;
  linear_predict_kernel<<<blocks, threads, 0, static_cast<cudaStream_t>(stream)>>>(Z, ld, B, w, d, out);
  return check_launch("linear_predict_kernel");
}

// ------------------------------------------------------------------------------------------
// Mini-batch update without any allocation: the accumulator the kernel needs lives in a
// file-scope device array instead of caller-provided scratch.  Meant for callers that cannot
// obtain device memory at call time - e.g. an ingest loop whose latency budget cannot absorb the
// device-wide synchronisation cudaMalloc/cudaFree imply, or an embedding process that handed all
// of its HBM to other components at start-up and forbids driver allocations afterwards.
//
// The array below is part of the loaded module: every GPU that loads this library holds its
// OMLDM_STATIC_SCRATCH_FLOATS * 4 bytes until the process exits.
static __device__ __align__(256) float g_static_scratch[OMLDM_STATIC_SCRATCH_FLOATS];

namespace {
constexpr int kMaxDevices = 64;
// Device address of g_static_scratch per device.  In host code the symbol name is only a
// stand-in, not an address a kernel can use, so it is resolved once per device with
// cudaGetSymbolAddress (with that device current) and cached.  Host threads may make their first call on a
// device at the same time: the slots are atomics (release store / acquire load), and two threads that both find
// a slot empty resolve the same address.
std::atomic<float*> g_static_scratch_ptr[kMaxDevices];
}  // namespace

extern "C" int omldm_linear_minibatch_static(const float* Z, int ld, int B, const omldm_linear_desc* desc_host,
                                             int mean_update, void* stream) {
  if (B <= 0) return 0;
  const int dw = desc_host->d + 1;
  if (desc_host->d < 1 || dw + 2 > OMLDM_STATIC_SCRATCH_FLOATS)
    return set_error_msg("linear_minibatch_static: d exceeds OMLDM_STATIC_SCRATCH_MAX_D (the built-in scratch "
                         "array is fixed-size); use omldm_linear_minibatch with caller-provided scratch");
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return set_error("cudaGetDevice", e);
  if (dev < 0 || dev >= kMaxDevices) return set_error_msg("linear_minibatch_static: device ordinal out of range");
  float* scratch = g_static_scratch_ptr[dev].load(std::memory_order_acquire);
  if (scratch == nullptr) {
    // one-time set-up for this device (the current one): resolve the symbol's device address
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&scratch), g_static_scratch);
    if (e != cudaSuccess) return set_error("cudaGetSymbolAddress(g_static_scratch)", e);
    g_static_scratch_ptr[dev].store(scratch, std::memory_order_release);
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // The array is shared by every call and keeps whatever the last call left behind; the
  // accumulators, the active count and the block counter must start at zero.
  e = cudaMemsetAsync(scratch, 0, sizeof(float) * (dw + 2), st);
  if (e != cudaSuccess) return set_error("cudaMemsetAsync(g_static_scratch)", e);
  int blocks = (B + MB_WARPS - 1) / MB_WARPS;
  if (blocks > 148 * 4) blocks = 148 * 4;
  linear_minibatch_kernel<<<blocks, MB_WARPS * 32, dw * 4, st>>>(Z, ld, B, *desc_host, scratch, mean_update);
  return check_launch("linear_minibatch_kernel(static scratch)");
}
