// Ingest kernels: holdout sampling, sparse->dense staging, raw->staged row conversion.
//
// Holdout rule being reproduced (operators/spoke/FlinkSpoke.scala:92-107): of every 10 training
// records, numbers 8 and 9 go to a bounded test set (default 256 rows); when the set is full the
// evicted (oldest) point is trained on instead.  Because the rule depends only on the running
// record count, it is a pure index map: holdout #h lives in ring slot h % cap and evicts holdout
// #(h - cap).  The kernel swaps rows in place, so after it runs the batch rows ARE the training
// sequence in the original order (an empty row, bias column 0, where nothing was evicted yet).
#include <cuda_fp16.h>
#include <cstring>
#include "ptx.cuh"
#include "host_common.h"
#include "omldm_kernels.h"

OMLDM_TL_DECLARE(holdout)

namespace omldm {

__host__ __device__ inline long long holdouts_before(long long c) {
  const long long r = c % 10;
  return 2 * (c / 10) + (r > 8 ? r - 8 : 0);
}

// One CTA per ring slot.  The holdouts of this batch that map to the slot form a chain
// h, h+cap, h+2cap, ... (batch rows i_0 < i_1 < ... < i_m).  Sequential semantics: every new
// holdout evicts the previous occupant, which is trained on in its place.  In place that is a
// shift by one along the chain: ring <- row i_m, row i_j <- row i_{j-1}, row i_0 <- old ring
// content (or an empty row while the ring is still filling).  Each thread owns a fixed set of
// columns, walks the chain from the tail and so never needs a barrier.
// V = the vector a thread moves (float4, or float2 for short rows so that all 256 threads of the CTA have a column:
// a 2 KB fp16 staged row is 256 float2); MAXC rows are in flight per thread (all loads issued before the shifted stores).
// `count_dev` (graph-replayable launches): the running record count is read from device memory and advanced by the
// LAST CTA to finish - every CTA has read the old value by then (`done` is a self-resetting arrival counter).
// `ordered` = 0: only the SET of records that is trained on matters (every mini-batch learner sums over the batch), so
// the chain collapses to one exchange per ring slot - the last holdout of the chain goes to the ring, the evicted
// occupant takes its place in the batch, and the holdouts in between stay where they are (they are all evicted, hence
// trained, within this batch anyway).  Same training set and same ring as the ordered shift, ~25x fewer rows moved
// at 32 768 records per batch.  `ordered` = 1 keeps the sequential positions (per-record learners).
template <typename V, int MAXC, bool EVEN_CAP>
__global__ void __launch_bounds__(256, 2)
holdout_swap_kernel(float* __restrict__ Z, int ld, int B, int d, float* __restrict__ ring, int cap,
                    long long count0_arg, long long* __restrict__ count_dev, unsigned int* __restrict__ done,
                    int ordered) {
  OMLDM_TL_BEGIN(holdout);
  const long long count0 = count_dev ? *reinterpret_cast<volatile long long*>(count_dev) : count0_arg;
  // narrow rows: several ring slots share a CTA so that every thread has a column (a 128-byte row is 16 float2)
  const int ldv_all = ld / static_cast<int>(sizeof(V) / sizeof(float));
  const int spc = ldv_all < static_cast<int>(blockDim.x) ? static_cast<int>(blockDim.x) / ldv_all : 1;
  const int slot = blockIdx.x * spc + (spc > 1 ? static_cast<int>(threadIdx.x) / ldv_all : 0);
  const int cv_first = spc > 1 ? static_cast<int>(threadIdx.x) % ldv_all : static_cast<int>(threadIdx.x);
  const int cv_step = spc > 1 ? ldv_all : static_cast<int>(blockDim.x);
  const bool lane_on = spc == 1 || static_cast<int>(threadIdx.x) < spc * ldv_all;
  const long long h0 = holdouts_before(count0), h1 = holdouts_before(count0 + B);
  const long long hf = h0 + ((slot - (h0 % cap)) % cap + cap) % cap;   // first chain element
  if (lane_on && slot < cap && hf < h1) {
    const int m = static_cast<int>((h1 - 1 - hf) / cap);                // chain has m + 1 elements
    // with an even ring size every element of a chain has the parity of the first one, so its rows are equally
    // spaced (5 cap records apart): one base pointer and a stride instead of a pointer per row
    const long long c0 = (hf / 2) * 10 + 8 + (hf & 1);
    V* const r0 = reinterpret_cast<V*>(Z + static_cast<size_t>(c0 - count0) * ld);
    const size_t rstride = static_cast<size_t>(cap / 2) * 10 * ld / (sizeof(V) / sizeof(float));
    auto row_of = [&](int j) -> V* {
      if (EVEN_CAP) return r0 + static_cast<size_t>(j) * rstride;
      const long long h = hf + static_cast<long long>(j) * cap;
      const long long c = (h / 2) * 10 + 8 + (h & 1);
      return reinterpret_cast<V*>(Z + static_cast<size_t>(c - count0) * ld);
    };
    const bool evict = hf >= cap;
    V* rs = reinterpret_cast<V*>(ring + static_cast<size_t>(slot) * ld);
    const int ldv = ld / static_cast<int>(sizeof(V) / sizeof(float));
    V zero;
    memset(&zero, 0, sizeof(V));
    for (int cv = cv_first; cv < ldv; cv += cv_step) {
      V* rv = rs + cv;
      const V old = evict ? *rv : zero;
      if (!ordered) {                      // exchange: ring <- last holdout of the chain, its row <- evicted occupant
        V* last = row_of(m) + cv;
        *rv = *last;
        *last = old;
        continue;
      }
      // walk the chain from the tail in segments of MAXC rows: row i_j <- row i_{j-1}.  Inside a segment every load is
      // issued before the shifted stores (independent, pipelined), so only ceil(m / MAXC) phases are serialised.
      *rv = row_of(m)[cv];
      for (int hi = m; hi >= 1; hi -= MAXC) {
        const int lo = hi - MAXC + 1 > 1 ? hi - MAXC + 1 : 1;   // rows lo..hi receive rows lo-1..hi-1
        V v[MAXC];
#pragma unroll
        for (int t = 0; t < MAXC; ++t)
          if (lo + t <= hi) v[t] = row_of(lo + t - 1)[cv];
#pragma unroll
        for (int t = 0; t < MAXC; ++t)
          if (lo + t <= hi) row_of(lo + t)[cv] = v[t];
      }
      row_of(0)[cv] = old;
    }
  }
  if (count_dev != nullptr) {
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned int prev = atomicAdd(done, 1u);
      if (prev == gridDim.x - 1) {              // every CTA has read the old count
        *done = 0;
        *count_dev = count0 + B;
      }
    }
  }
  OMLDM_TL_END(holdout);
}

// The exchange rule (`ordered` = 0) on its own: no chain walk, hence few registers - small enough to be placed on an SM
// that is already full of a GEMM CTA and two loss-pass blocks, which is where it runs when the engine pipelines the
// holdout of the next batch behind the current training step (Engine `pipelinedHoldout`).
template <typename V>
__global__ void __launch_bounds__(256)
holdout_exchange_kernel(float* __restrict__ Z, int ld, int B, float* __restrict__ ring, int cap,
                        long long* __restrict__ count_dev, unsigned int* __restrict__ done) {
  OMLDM_TL_BEGIN(holdout);
  const long long count0 = *reinterpret_cast<volatile long long*>(count_dev);
  const int ldv = ld / static_cast<int>(sizeof(V) / sizeof(float));
  const int spc = ldv < static_cast<int>(blockDim.x) ? static_cast<int>(blockDim.x) / ldv : 1;
  const int slot = blockIdx.x * spc + (spc > 1 ? static_cast<int>(threadIdx.x) / ldv : 0);
  const int cv_first = spc > 1 ? static_cast<int>(threadIdx.x) % ldv : static_cast<int>(threadIdx.x);
  const int cv_step = spc > 1 ? ldv : static_cast<int>(blockDim.x);
  const bool lane_on = spc == 1 || static_cast<int>(threadIdx.x) < spc * ldv;
  const long long h0 = holdouts_before(count0), h1 = holdouts_before(count0 + B);
  const long long hf = h0 + ((slot - (h0 % cap)) % cap + cap) % cap;   // first holdout of this batch that maps to the slot
  if (lane_on && slot < cap && hf < h1) {
    const long long hl = hf + (h1 - 1 - hf) / cap * cap;               // last one: it stays in the ring
    const long long c = (hl / 2) * 10 + 8 + (hl & 1);                  // its record index
    V* last = reinterpret_cast<V*>(Z + static_cast<size_t>(c - count0) * ld);
    V* rs = reinterpret_cast<V*>(ring + static_cast<size_t>(slot) * ld);
    const bool evict = hf >= cap;
    V zero;
    memset(&zero, 0, sizeof(V));
    for (int cv = cv_first; cv < ldv; cv += cv_step) {
      const V old = evict ? rs[cv] : zero;
      rs[cv] = last[cv];
      last[cv] = old;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(done, 1u);
    if (prev == gridDim.x - 1) {              // every CTA has read the old count
      *done = 0;
      *count_dev = count0 + B;
    }
  }
  OMLDM_TL_END(holdout);
}

__global__ void advance_count_kernel(long long* c, long long delta) { *c += delta; }

// Learning curve of the test mode (operators/hub/FlinkHub.scala:101-113: the hub records (loss, #fitted) per
// mini-batch): one sample of the learner's cumulative (loss sum, records fitted) per step into a device ring, so a
// default run keeps its curve without any host synchronisation.  counter = samples written so far.
__global__ void curve_sample_kernel(const double* __restrict__ stats, double* __restrict__ ring, int cap,
                                    long long* __restrict__ counter) {
  const long long i = *counter;
  double* slot = ring + 2 * (i % cap);
  slot[0] = stats[0];
  slot[1] = stats[2];
  *counter = i + 1;
}

__global__ void densify_kernel(const int* __restrict__ indptr, const int* __restrict__ indices,
                               const float* __restrict__ values, int d, const float* __restrict__ targets,
                               float* __restrict__ Z, int ld, float* __restrict__ sqnorm) {
  const int r = blockIdx.x;
  float* row = Z + static_cast<size_t>(r) * ld;
  float4* row4 = reinterpret_cast<float4*>(row);
  for (int c = threadIdx.x; c < ld / 4; c += blockDim.x) row4[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  __syncthreads();
  float acc = 0.f;
  const int b = indptr[r], e = indptr[r + 1];
  for (int p = b + threadIdx.x; p < e; p += blockDim.x) {
    const int c = indices[p];
    const float v = values[p];
    if (c >= 0 && c < d) {
      atomicAdd(&row[c], v);  // duplicate indices accumulate
      acc = fmaf(v, v, acc);
    }
  }
  __shared__ float s[32];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 1.f;  // bias column contributes 1
    for (int i = 0; i < (blockDim.x >> 5); ++i) t += s[i];
    if (sqnorm) sqnorm[r] = t;
    row[d] = 1.f;
    row[d + 1] = targets ? targets[r] : 0.f;
  }
}

__global__ void stage_rows_kernel(const float* __restrict__ raw, int raw_ld, int B, int d,
                                  float* __restrict__ Z, int ld) {
  const int r = blockIdx.x;
  const float* src = raw + static_cast<size_t>(r) * raw_ld;
  float* dst = Z + static_cast<size_t>(r) * ld;
  for (int c = threadIdx.x; c < ld; c += blockDim.x) {
    float v = 0.f;
    if (c < d) v = src[c];
    else if (c == d) v = 1.f;
    else if (c == d + 1) v = src[d];
    dst[c] = v;
  }
}

// Compact wire format -> staged rows.  A wire row is [x as fp16 x d | (pad to an even count) | target as
// fp32 in two halves]; pitch `raw_ld` halves.  fp16 keeps 11 significand bits, one more than the tf32
// products the Gram/score GEMMs use, so for features inside the fp16 range the learners see the same
// inputs - at half the PCIe bytes per record (the e2e path is H2D bound).
__global__ void stage_rows_f16_kernel(const __half* __restrict__ raw, int raw_ld, int B, int d,
                                      float* __restrict__ Z, int ld) {
  const int r = blockIdx.x;
  const __half* src = raw + static_cast<size_t>(r) * raw_ld;
  const __half2* src2 = reinterpret_cast<const __half2*>(src);
  float2* dst2 = reinterpret_cast<float2*>(Z + static_cast<size_t>(r) * ld);
  const int d_even = (d + 1) & ~1;
  const float target = *reinterpret_cast<const float*>(src + d_even);
  for (int c2 = threadIdx.x; c2 < (ld >> 1); c2 += blockDim.x) {
    const int c = 2 * c2;
    float2 v;
    if (c + 1 < d) {
      v = __half22float2(src2[c2]);
    } else {
      v.x = c < d ? __half2float(src[c]) : (c == d ? 1.f : (c == d + 1 ? target : 0.f));
      v.y = (c + 1 == d) ? 1.f : (c + 1 == d + 1 ? target : 0.f);
    }
    dst2[c2] = v;
  }
}

}  // namespace omldm

using namespace omldm;

static int launch_holdout(float* Z, int ld, int B, int d, float* ring, int cap, long long count0, long long* count_dev,
                          unsigned int* done, int ordered, cudaStream_t st) {
  if (ld & 3) return set_error_msg("holdout_swap: ld must be a multiple of 4");
  const bool even = (cap & 1) == 0;
  const int ldv = ld <= 512 ? ld / 2 : ld / 4;
  const int spc = ldv < 256 ? 256 / ldv : 1;           // ring slots per CTA (narrow rows)
  const int grid = (cap + spc - 1) / spc;
  if (!ordered && count_dev != nullptr) {
    if (ld <= 512) holdout_exchange_kernel<float2><<<grid, 256, 0, st>>>(Z, ld, B, ring, cap, count_dev, done);
    else holdout_exchange_kernel<float4><<<grid, 256, 0, st>>>(Z, ld, B, ring, cap, count_dev, done);
    return check_launch("holdout_exchange_kernel");
  }
  if (ld <= 512) {  // short rows (e.g. 2 KB fp16 staged rows viewed as 512 words): 8 bytes per thread keeps the CTA busy
    if (even) holdout_swap_kernel<float2, 32, true><<<grid, 256, 0, st>>>(Z, ld, B, d, ring, cap, count0, count_dev, done, ordered);
    else holdout_swap_kernel<float2, 16, false><<<grid, 256, 0, st>>>(Z, ld, B, d, ring, cap, count0, count_dev, done, ordered);
  } else {
    if (even) holdout_swap_kernel<float4, 16, true><<<grid, 256, 0, st>>>(Z, ld, B, d, ring, cap, count0, count_dev, done, ordered);
    else holdout_swap_kernel<float4, 16, false><<<grid, 256, 0, st>>>(Z, ld, B, d, ring, cap, count0, count_dev, done, ordered);
  }
  return check_launch("holdout_swap_kernel");
}

extern "C" int omldm_holdout_swap(float* Z, int ld, int B, int d, float* ring, int cap, long long count0,
                                  void* stream) {
  if (B <= 0 || cap <= 0) return 0;
  if (holdouts_before(count0 + B) == holdouts_before(count0)) return 0;
  return launch_holdout(Z, ld, B, d, ring, cap, count0, nullptr, nullptr, 1, static_cast<cudaStream_t>(stream));
}

// Graph-replayable variant: the running record count lives in device memory (count_dev[0]) and is advanced by the
// kernel itself once every CTA has read it; count_dev[1] is the kernel's arrival counter (zero before the first call).
extern "C" int omldm_holdout_swap_dev(float* Z, int ld, int B, int d, float* ring, int cap,
                                      long long* count_dev, void* stream) {
  if (B <= 0 || cap <= 0) return 0;
  return launch_holdout(Z, ld, B, d, ring, cap, 0, count_dev, reinterpret_cast<unsigned int*>(count_dev + 1), 1,
                        static_cast<cudaStream_t>(stream));
}

// Same, for consumers that only sum over the batch (every mini-batch learner): one exchange per ring slot instead of
// the shift along the whole chain.  Same set of trained records, same ring; the evicted records sit at other
// positions of the batch than the sequential rule would put them.
extern "C" int omldm_holdout_exchange_dev(float* Z, int ld, int B, int d, float* ring, int cap,
                                          long long* count_dev, void* stream) {
  if (B <= 0 || cap <= 0) return 0;
  return launch_holdout(Z, ld, B, d, ring, cap, 0, count_dev, reinterpret_cast<unsigned int*>(count_dev + 1), 0,
                        static_cast<cudaStream_t>(stream));
}

extern "C" int omldm_advance_count(long long* count_dev, long long delta, void* stream) {
  advance_count_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(count_dev, delta);
  return check_launch("advance_count_kernel");
}

extern "C" int omldm_curve_sample(const double* stats, double* ring, int cap, long long* counter, void* stream) {
  if (cap <= 0) return set_error_msg("curve_sample: the ring needs at least one slot");
  curve_sample_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(stats, ring, cap, counter);
  return check_launch("curve_sample_kernel");
}

extern "C" int omldm_densify(const int* indptr, const int* indices, const float* values, int B, int d,
                             const float* targets, float* Z, int ld, float* sqnorm, void* stream) {
  if (B <= 0) return 0;
  if (ld & 3) return set_error_msg("densify: ld must be a multiple of 4");
  densify_kernel<<<B, 256, 0, static_cast<cudaStream_t>(stream)>>>(indptr, indices, values, d, targets, Z,
                                                                 ld, sqnorm);
  return check_launch("densify_kernel");
}

extern "C" int omldm_stage_rows(const float* raw, int raw_ld, int B, int d, float* Z, int ld, void* stream) {
  if (B <= 0) return 0;
  stage_rows_kernel<<<B, 128, 0, static_cast<cudaStream_t>(stream)>>>(raw, raw_ld, B, d, Z, ld);
  return check_launch("stage_rows_kernel");
}

extern "C" int omldm_stage_rows_f16(const void* raw, int raw_ld, int B, int d, float* Z, int ld, void* stream) {
  if (B <= 0) return 0;
  if ((raw_ld & 1) || raw_ld < ((d + 1) & ~1) + 2) return set_error_msg("stage_rows_f16: raw_ld must be even and >= even(d) + 2");
  if ((ld & 1) || ld < d + 2) return set_error_msg("stage_rows_f16: ld must be even and >= d + 2");
  if (reinterpret_cast<uintptr_t>(raw) & 3) return set_error_msg("stage_rows_f16: raw must be 4-byte aligned");
  stage_rows_f16_kernel<<<B, 128, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __half*>(raw), raw_ld, B, d,
                                                                        Z, ld);
  return check_launch("stage_rows_f16_kernel");
}
