// Gram update of online ridge regression on fp16 record rows: G (+)= Z^T Z, straight from the rows as they arrive.
//
// Replaces the per-record `A += x x^T, b += y x` behind `node.receiveTuple`
// (reference call site: src/main/scala/omldm/operators/spoke/FlinkSpoke.scala:101) and, on more than one GPU, the
// worker -> hub push of the model delta (network/FlinkNetwork.scala:260-261): the kernel that computes the
// update also delivers it to the hub shards that own its rows.  The reference has no native code.
//
// Z is the mini-batch as it sits in the staged ring slot: [K records, ld halves], a record = [x | 1 | y_hi | y_lo | 0..]
// (fp16, 11-bit significand = the tf32 product precision; the target is split into two halves so that b = X^T y keeps
// 22 bits).  Both operands of Z^T Z are MN-major views of the SAME matrix, so no transpose pass exists anywhere.
//
// Blackwell structure
//   * CTA pairs (cluster 2x1x1, tcgen05 cta_group::2): one pair owns one 256 x 256 output tile of the upper triangle
//     and one slice of the batch axis (split-K).  Each CTA TMA-loads its 128 columns of the A side and its 128
//     columns of the B side per 64-record stage (plain loads from this GPU's memory, no multicast); the pair's MMAs
//     read both CTAs' shared memory, which halves the L2 -> smem traffic per flop against 128 x 256 single-CTA tiles.
//     On diagonal tiles the A side and the B side are the same columns: loaded once, used for both descriptors.
//   * one elected thread of the leader CTA issues tcgen05.mma.cta_group::2.kind::f16 (M 256, N 256, K 16);
//     accumulators live in TMEM (128 lanes x 256 columns per CTA); tcgen05.commit (multicast to both CTAs) frees the
//     smem stages and publishes the accumulator.
//   * epilogue: tcgen05.ld -> registers -> swizzled smem -> TMA reduce-add (fp32 adds at L2), upper 32 x 32 blocks only.
//   * hub push, lead mode (world > 1, default): while the MMAs of this launch run, the otherwise idle epilogue warps
//     copy the running sum AS IT WAS WHEN THE KERNEL STARTED (the result of the previous mini-batch), 32 x 32 block by
//     block, with TMA (load from L2 -> plain store over NVLink) into THIS worker's slot at the GPU that owns those rows;
//     the blocks of a half tile are dealt out over its batch-axis splits.  The transfer hides completely behind the main
//     loop; the slots trail the running sum by one mini-batch and are brought up to date by
//     csrc/sync.cu::sync_gram_push_kernel right before anybody reads them (the gather at a weight refresh or a query).
//   * hub push, tail mode: the split-K partials meet in this GPU's own L2-resident running sum; the LAST CTA to
//     finish a 128-row half tile then copies the updated rows, 32 x 32 block by block, with TMA (load from L2 ->
//     plain store over NVLink) into THIS worker's slot at the GPU that owns those rows.  Every slot has a single
//     writer, so the transfer needs no remote atomics, no flags and no zeroing, every element crosses NVLink once,
//     and the CTA only waits until the stores have read its shared memory.  The owners sum their slots when the
//     workers pull the global statistics (csrc/sync.cu::sync_gram_gather_kernel).  No CTA ever waits for another
//     CTA or GPU.
#include <cuda_fp16.h>
#include <cstring>
#include "ptx.cuh"
#include "host_common.h"
#include "omldm_kernels.h"

namespace omldm {

#ifndef OMLDM_GRAM_BK
#define OMLDM_GRAM_BK 64
#endif
#ifndef OMLDM_GRAM_STAGES
#define OMLDM_GRAM_STAGES 6
#endif
constexpr int G_BK = OMLDM_GRAM_BK;               // records per stage
constexpr int G_STAGES = OMLDM_GRAM_STAGES;
#ifndef OMLDM_GRAM_THREADS
#define OMLDM_GRAM_THREADS 384
#endif
constexpr int G_THREADS = OMLDM_GRAM_THREADS;                    // warp 0 TMA, warp 1 TMEM + MMA, warps 2, 3, 6..11 loss (four groups, only with the loss inside), warps 4..7 epilogue, warp 4 relay, warp 5 lead push
constexpr int G_HALF_BYTES = 2 * G_BK * 128;      // 128 columns = 2 chunks of 64 halves x 64 records = 16 KB
constexpr int G_STAGE_BYTES = 2 * G_HALF_BYTES;   // A side + B side
constexpr int G_SMEM = G_STAGES * G_STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
constexpr int G_LEAD_BYTES = 4 * 4096 + 768;      // lead-mode push: one 32 x 32 fp32 staging block per epilogue warp, 1024-aligned behind the barriers
constexpr int G_TMEM_COLS = 256;
enum { GS_TILE0 = 8 };
#ifdef OMLDM_GRAM_TIMING
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#endif
#ifdef OMLDM_GRAM_TIMING        // tuning builds only (scripts/build_variant.sh): per-CTA phase stamps in SM cycles, read back by
                                // tests/test_gram_f16_gpu.py::test_gram_f16_speed_report
__device__ long long g_gram_stamp[160 * 16];   // 0..6 SM cycles per phase, 8 / 9 globaltimer ns at start / end
__device__ long long g_gram_trace[2][5][24];   // tile 0, slice 0, both ranks: per stage use 30..53 the cycle of five events
#define GRAM_TRACE(ev, kb) do { if (blockIdx.x < 2 && blockIdx.y == 0 && (kb) >= 30 && (kb) < 54) g_gram_trace[blockIdx.x][ev][(kb) - 30] = clock64(); } while (0)
#define GRAM_STAMP(i) do { if ((threadIdx.x & 31) == 0) g_gram_stamp[(blockIdx.y * gridDim.x + blockIdx.x) % 160 * 16 + (i)] = (i) >= 8 ? static_cast<long long>(globaltimer_ns()) : clock64(); } while (0)
#else
#define GRAM_STAMP(i) do { } while (0)
#define GRAM_TRACE(ev, kb) do { } while (0)
#endif                            // scratch words: per half-tile arrival counters from word 8 on

struct GramPush {
  CUtensorMap owner[16];     // this worker's slot matrix at every owner (peer mapped), 32 x 32 fp32 boxes
  int world, rank, rows_per_owner, enabled;   // enabled: 0 off, 1 tail (exact at kernel end), 2 lead (during the main loop)
};

// Prequential loss of the batch from inside the Gram kernel (optional).  A 256-column block of the records sits in the
// shared memory of every pair whose tile touches it - as the A side of tiles (c, j >= c), as the B side of tiles
// (i < c, c): T pairs per block.  Each of them takes one T-th of the records of a stage: while the tensor cores work on the
// stage, two loss warps per CTA (one per operand side) read those records' 128 columns, form the partial dot products
// z_r . w and add them into dots[r] (fp32 red at L2; 2 T partials per record).  The last CTA to finish a batch slice
// squares and sums that slice's dots into stats[0], counts its valid rows into stats[2] and clears them; the last slice
// writes the learning-curve sample.  No second pass over the batch, no second kernel.
// Measured (B200, 32 768 x 1003): correct to 1e-9 relative; 39.3 us per launch against 34.9 us without the loss, and
// 48.0 against 46.8 us per engine step - a tie with the separate loss pass beside the kernel, which therefore stays the
// default; this path is `lossInGram: true`.  What is left of the difference is mostly the slice finish at the very end of
// the kernel (one warp reads and clears 4 681 dots: ~37 dependent-ish L2 round trips); the six loss warps that have no
// epilogue role could share it.
// Things that were tried on the way (each measured): the loss on the diagonal tiles' CTAs only (48.8 us per launch, those
// CTAs then pace the kernel); two groups of loss warps instead of four (47.6 us: a loss warp needs ~700 cycles of work
// plus ~650 cycles of barrier hand-over per stage use, so with 1 024 cycles available stages were released late and the
// main loop took 51 k instead of 38 k cycles); the remote "stage landed" arrive issued by the MMA thread (the mere presence
// of that branch in its loop: main loop 58 k cycles) or by one relay thread with a release at cluster scope (~1 000 cycles
// each); every T-th record per holder (16-way bank conflicts); generic instead of shared-space loads.
struct GramLoss {
  const float* wt;           // [w_0 .. w_{d-1}, b, -1, -1]: residual of a record = z . wt
  float* dots;               // K floats, zero between launches
  unsigned int* words;       // 3 * 128 words, zero between launches: valid rows per slice | arrivals per slice | slices done
  double* stats;             // stats[0] += sum of squared residuals, stats[2] += valid rows
  double* curve_ring;        // optional learning-curve ring [cap][2]
  long long* curve_counter;
  int curve_cap, d, enabled;
};

template <bool LOSS>     // LOSS = false: the loss code is compiled out (72 registers instead of 142, which is what lets two
                         // blocks of the separate loss pass fit beside a CTA of this kernel)
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(G_THREADS, 1)
gram_f16_kernel(const __grid_constant__ CUtensorMap tmZ, const __grid_constant__ CUtensorMap tmC,
                const __grid_constant__ GramPush push, const GramLoss loss, int n, int K, int k_per_split, int T,
                uint32_t* __restrict__ scratch) {
  const uint32_t rank = cluster_ctarank();        // 0 = leader of the pair
  const int tile = blockIdx.x >> 1;
  int ti = 0, tj = 0;
  {
    int t = tile;
    for (ti = 0; ti < T; ++ti) {
      const int len = T - ti;
      if (t < len) { tj = ti + t; break; }
      t -= len;
    }
  }
#ifdef OMLDM_GRAM_FORCE_DIAG     // traffic experiment only (wrong results): every tile loads one operand, like a diagonal tile
  const bool diag = true;
#else
  const bool diag = ti == tj;
#endif
  const int m0 = ti * 256 + static_cast<int>(rank) * 128;    // A-side columns of this CTA = its accumulator rows
  const int b0 = tj * 256 + static_cast<int>(rank) * 128;    // B-side columns this CTA feeds
  const int n0 = tj * 256;                                   // accumulator columns (all 256 in each CTA)
  const int k_begin = blockIdx.y * k_per_split;
  const int k_end = min(K, k_begin + k_per_split);
  const int num_kb = (k_end - k_begin + G_BK - 1) / G_BK;
  if (num_kb <= 0) return;                                   // same decision in both CTAs of the pair

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + G_STAGES * G_STAGE_BYTES);
  uint64_t* empty_bar = full_bar + G_STAGES;
  uint64_t* accum_bar = empty_bar + G_STAGES;
  uint64_t* push_bar = accum_bar + 1;                        // 4, one per epilogue warp
  uint64_t* ready_bar = push_bar + 4;                        // G_STAGES: "stage landed" for the loss warps of the non-leader
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(ready_bar + G_STAGES);
  uint32_t* s_flag = tmem_slot + 1;
  static_assert((3 * G_STAGES + 5) * 8 + 8 <= 256, "barrier area");
  const bool loss_on = LOSS && loss.enabled;                 // this CTA also forms residual dots (its share of the records)

  const int warp = threadIdx.x >> 5;
  if (warp == 0) { GRAM_STAMP(0); GRAM_STAMP(8); }
  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmZ);
    tma_prefetch_desc(&tmC);
    for (int s = 0; s < G_STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], loss_on ? (diag ? 2 : 3) : 1);  // the MMA commit (+ one arrival per loss warp of the stage's group)
      mbar_init(&ready_bar[s], 1);
    }
    mbar_init(accum_bar, 1);
    for (int q = 0; q < 4; ++q) mbar_init(&push_bar[q], 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_pair<G_TMEM_COLS>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();                 // barrier inits of both CTAs are visible before any remote arrival
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (warp == 0) GRAM_STAMP(1);

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (both CTAs)
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
#ifdef OMLDM_GRAM_TIMING
      long long t_empty = 0;
#endif
      for (int kb = 0; kb < num_kb; ++kb) {
#ifdef OMLDM_GRAM_TIMING
        const long long te0 = clock64();
#endif
        mbar_wait(&empty_bar[stage], phase ^ 1);
#ifdef OMLDM_GRAM_TIMING
        t_empty += clock64() - te0;
        if (kb == num_kb - 1) g_gram_stamp[(blockIdx.y * gridDim.x + blockIdx.x) % 160 * 16 + 12] = t_empty;
#endif
        GRAM_TRACE(0, kb);
        uint8_t* sa = smem + stage * G_STAGE_BYTES;
        uint8_t* sb = sa + G_HALF_BYTES;
        // the leader's barrier counts the bytes of both CTAs
        if (rank == 0) mbar_expect_tx(&full_bar[stage], diag ? 2 * G_HALF_BYTES : 2 * G_STAGE_BYTES);
        const int k0 = k_begin + kb * G_BK;
        tma_load_3d_pair(sa, &tmZ, &full_bar[stage], 0, k0, m0 / 64);
        if (!diag) tma_load_3d_pair(sb, &tmZ, &full_bar[stage], 0, k0, b0 / 64);
        if (++stage == G_STAGES) { stage = 0; phase ^= 1; }
      }
    }
    __syncwarp();
    if (loss_on) {
      // ---------------------------------------------------------- loss: close this batch slice if we are its last CTA half
      named_bar_sync(2, 288);                      // the eight loss warps of this CTA have added their dots
      const uint32_t lane = lane_id();
      unsigned int last = 0;
      if (lane == 0) {
        __threadfence();
        unsigned int* cnt = loss.words + 128 + blockIdx.y;
        last = atomicAdd(cnt, 1u) == gridDim.x - 1u;             // every CTA of the slice contributes
        if (last) *cnt = 0;
      }
      last = __shfl_sync(0xffffffffu, last, 0);
      if (last) {
        __threadfence();
        double sq = 0.0;
        const int k4 = (k_end - k_begin) >> 2;
        float4* d4 = reinterpret_cast<float4*>(loss.dots + k_begin);            // k_begin is a multiple of 64
        for (int i = static_cast<int>(lane); i < k4; i += 32) {
          const float4 v = __ldcg(d4 + i);
          d4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
          sq += static_cast<double>(v.x) * v.x + static_cast<double>(v.y) * v.y + static_cast<double>(v.z) * v.z +
                static_cast<double>(v.w) * v.w;
        }
        for (int i = k_begin + 4 * k4 + static_cast<int>(lane); i < k_end; i += 32) {
          const float v = __ldcg(loss.dots + i);
          loss.dots[i] = 0.f;
          sq += static_cast<double>(v) * v;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
        if (lane == 0) {
          const unsigned int nv = __ldcg(loss.words + blockIdx.y);
          loss.words[blockIdx.y] = 0;
          if (nv) {
            atomicAdd(loss.stats + 0, sq);
            atomicAdd(loss.stats + 2, static_cast<double>(nv));
          }
          __threadfence();
          if (atomicAdd(loss.words + 256, 1u) == gridDim.y - 1) {               // the last slice of this launch
            loss.words[256] = 0;
            __threadfence();
            if (loss.curve_ring) {
              const long long i = *loss.curve_counter;
              double* slot = loss.curve_ring + 2 * (i % loss.curve_cap);
              slot[0] = __ldcg(loss.stats + 0);
              slot[1] = __ldcg(loss.stats + 2);
              *loss.curve_counter = i + 1;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (leader CTA only)
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc_f16(256, 256, true, true);
      int stage = 0;
      uint32_t phase = 0;
#ifdef OMLDM_GRAM_TIMING
      long long t_full = 0;
#endif
      for (int kb = 0; kb < num_kb; ++kb) {
#ifdef OMLDM_GRAM_TIMING
        const long long tf0 = clock64();
#endif
        mbar_wait(&full_bar[stage], phase);
#ifdef OMLDM_GRAM_TIMING
        t_full += clock64() - tf0;
        if (kb == num_kb - 1 && lane_id() == 0) g_gram_stamp[(blockIdx.y * gridDim.x + blockIdx.x) % 160 * 16 + 13] = t_full;
        if (lane_id() == 0) GRAM_TRACE(2, kb);
#endif
        if (kb == 0) GRAM_STAMP(2);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + stage * G_STAGE_BYTES);
          const uint32_t sb = diag ? sa : sa + G_HALF_BYTES;
#pragma unroll
          for (int kk = 0; kk < G_BK / 16; ++kk) {
            // MN-major, 128-byte swizzle: 64-column chunks are G_BK * 128 B apart (LBO), 8-record groups 1024 B
            // apart (SBO); one MMA consumes 16 records = 2048 B
            const uint64_t da = make_smem_desc_sw128(sa + kk * 2048, G_BK * 128, 1024, 2);
            const uint64_t db = make_smem_desc_sw128(sb + kk * 2048, G_BK * 128, 1024, 2);
            umma_f16_pair(tmem_base, da, db, idesc, (kb | kk) != 0);
          }
          umma_commit_pair(&empty_bar[stage], 3);               // frees this stage in both CTAs
          if (kb == num_kb - 1) umma_commit_pair(accum_bar, 3);  // accumulators complete, both CTAs
        }
        __syncwarp();
        if (++stage == G_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
   // Warps 2..7.  Loss work (ALU heavy) runs on warps 2, 3, 6, 7 = schedulers 2 and 3 only: on the schedulers of the
   // TMA producer (warp 0) and the MMA issuer (warp 1) it delayed their few but urgent instructions and the main loop
   // of the diagonal tiles took 88 k instead of 38 k cycles.  Warps 4 and 5 (schedulers 0 and 1) only relay / push,
   // i.e. mostly sleep.  The epilogue needs one warp per TMEM lane quadrant (warp & 3): warps 4..7.
   if (warp == 4) {
    // ------------------------------------------------------------ relay (leader CTA, loss inside): tells the peer's loss
    // warps that a stage has landed.  The peer's TMA loads signal the LEADER's barrier, and putting this arrive into
    // the MMA-issuing thread's loop slowed the main loop from 38 k to 58 k cycles even when it was never executed.
    if (loss_on && rank == 0) {
      // one lane per stage: the release at cluster scope in front of the remote arrive takes ~1 000 cycles, so a single
      // relaying thread would pace the pipeline at one stage per 1 000 cycles (measured: main loop 88 k instead of 38 k)
      const int l = static_cast<int>(lane_id());
      if (l < G_STAGES) {
        uint32_t ph = 0;
        for (int kb = l; kb < num_kb; kb += G_STAGES) {
          mbar_wait(&full_bar[l], ph);
          mbar_arrive_remote_relaxed(&ready_bar[l], 1);
          ph ^= 1;
        }
      }
      __syncwarp();
    }
   } else if (warp == 5) {
    // ------------------------------------------------------------ hub push, lead mode: hidden behind the main loop
    if (push.enabled == 2 && elect_one()) {
      uint8_t* stage_blk = smem + G_STAGES * G_STAGE_BYTES + 1024;
      uint32_t ph = 0;
      for (int b = static_cast<int>(blockIdx.y); b < 32; b += static_cast<int>(gridDim.y)) {   // this split's share of the half tile
        const int row0 = m0 + (b >> 3) * 32, col0 = n0 + (b & 7) * 32;
        if (row0 >= n || col0 >= n || col0 + 31 < row0) continue;
        mbar_expect_tx(&push_bar[0], 4096);
        tma_load_2d(stage_blk, &tmC, &push_bar[0], col0, row0);
        mbar_wait(&push_bar[0], ph);
        ph ^= 1;
        tma_store_2d(&push.owner[row0 / push.rows_per_owner], stage_blk, col0, row0);   // this worker's slot at the owner (NVLink)
        tma_commit_group();
        tma_wait_group_read0();                    // the staging block is free again
      }
    }
    __syncwarp();
   }
   {
    // ------------------------------------------------------------ loss warps (2, 3, 6, 7) / epilogue (4..7), both CTAs
    const int q = warp & 3;                        // TMEM lane quadrant this warp may read
    const uint32_t r = lane_id();
    const int row0 = m0 + q * 32;
    // eight loss warps in four groups of two (A side / B side): warps 2 / 3, 6 / 7, 8 / 9, 10 / 11; group g takes the stage
    // uses g, g + 4, ... - a warp needs ~1 350 cycles per stage use (700 of work, 650 of barrier hand-over), the MMAs
    // 512 cycles, so two groups (1 024 cycles each) paced the pipeline and four (2 048) do not
    const bool loss_warp = warp == 2 || warp == 3 || warp >= 6;
    const int lgroup = warp < 4 ? 0 : warp < 8 ? 1 : warp < 10 ? 2 : 3, side = warp & 1;
    constexpr int LGROUPS = 4;
    if (loss_on && loss_warp) {
      // Side A: the CTA's accumulator-row columns [m0, m0 + 128), this pair is holder tj of that block; side B (off the
      // diagonal only): columns [b0, b0 + 128), holder ti.  Holder h of T takes the h-th T-th of a stage's records.
      // A lane owns one 64-column chunk (lanes 0-15 chunk 0, 16-31 chunk 1) and every 16th of those records: eight
      // 16-byte loads and 64 FMAs per record against weights held in registers.
      const bool active = side == 0 || !diag;
      const int cbase = side == 0 ? m0 : b0, holder = side == 0 ? tj : ti;
      const int ch = static_cast<int>(r) >> 4, slot = static_cast<int>(r) & 15;
      float w64[64];
#pragma unroll
      for (int j = 0; j < 64; ++j) {
        const int col = cbase + ch * 64 + j;
        w64[j] = (active && col < n) ? __ldg(loss.wt + col) : 0.f;
      }
      const int cd = loss.d - cbase;               // the bias / valid column, if this side holds it
      const bool has_valid = active && cd >= 0 && cd < 128 && (cd >> 6) == ch;
      const int uv = (cd & 63) >> 3, jv = cd & 7;
      unsigned int nvalid = 0;
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t side_off = static_cast<uint32_t>(side * G_HALF_BYTES + ch * (G_BK * 128));
#ifdef OMLDM_GRAM_TIMING
      long long t_wait = 0, t_work = 0;
#endif
      for (int kb = lgroup; kb < num_kb; kb += LGROUPS) {
        stage = kb % G_STAGES;
        phase = static_cast<uint32_t>(kb / G_STAGES) & 1u;
#ifdef OMLDM_GRAM_TIMING
        const long long t0 = clock64();
#endif
        if (active) {
          if (r == 0) {                             // one lane polls
            if (rank == 0) mbar_wait(&full_bar[stage], phase);
            else mbar_wait(&ready_bar[stage], phase);      // (an acquire at cluster scope here cost ~700 cycles per stage)
          }
          __syncwarp();
#ifdef OMLDM_GRAM_TIMING
          const long long t1 = clock64();
          t_wait += t1 - t0;
          if (warp == 2 && r == 0) GRAM_TRACE(3, kb);
#endif
          const uint32_t base = smem_u32(smem) + static_cast<uint32_t>(stage * G_STAGE_BYTES) + side_off;
          // holder h takes the records [h * share, (h + 1) * share) of the stage: consecutive records keep the 16-byte
          // loads of a half warp on different banks (every T-th record put them all on two: 917 cycles per stage)
          const int share = (G_BK + T - 1) / T, iters = (share + 15) / 16;   // the same count in every lane (shuffle below)
          for (int it = 0; it < iters; ++it) {
            const int in_share = slot + 16 * it, kr_raw = holder * share + in_share;
            const bool live = in_share < share && kr_raw < G_BK;
            const int kr = live ? kr_raw : 0;
            const uint32_t rowp = base + static_cast<uint32_t>(kr) * 128u;
            const uint32_t sw = static_cast<uint32_t>(kr & 7);
            uint4 raw[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) raw[u] = lds_v4(rowp + ((static_cast<uint32_t>(u) ^ sw) << 4));
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              const __half2* hp = reinterpret_cast<const __half2*>(&raw[u]);
              const float2 f0 = __half22float2(hp[0]), f1 = __half22float2(hp[1]), f2 = __half22float2(hp[2]),
                           f3 = __half22float2(hp[3]);
              a0 = fmaf(f0.x, w64[8 * u], a0);
              a1 = fmaf(f0.y, w64[8 * u + 1], a1);
              a2 = fmaf(f1.x, w64[8 * u + 2], a2);
              a3 = fmaf(f1.y, w64[8 * u + 3], a3);
              a0 = fmaf(f2.x, w64[8 * u + 4], a0);
              a1 = fmaf(f2.y, w64[8 * u + 5], a1);
              a2 = fmaf(f3.x, w64[8 * u + 6], a2);
              a3 = fmaf(f3.y, w64[8 * u + 7], a3);
            }
            if (has_valid && live) {                // bias half != +-0: the record is not an empty row
              uint32_t wv = 0;
#pragma unroll
              for (int u = 0; u < 8; ++u)
                if (u == uv) wv = jv < 2 ? raw[u].x : jv < 4 ? raw[u].y : jv < 6 ? raw[u].z : raw[u].w;
              nvalid += (((jv & 1) ? wv >> 16 : wv) & 0x7fffu) ? 1u : 0u;
            }
            // the two chunks of a record sit in lanes l and l ^ 16, which run this loop in lock step (same kr)
            float acc = (a0 + a1) + (a2 + a3);
            acc += __shfl_xor_sync(0xffffffffu, acc, 16);
            if (it + 1 == iters) {
              // Release the stage BEFORE the last adds go out: the arrive is a release, and a release waits for the
              // thread's outstanding global reds (an L2 round trip: 690 instead of ~200 cycles of work per stage).
              // The shared-memory reads of the stage are complete - their values have been used.
              __syncwarp();
              if (r == 0) mbar_arrive(&empty_bar[stage]);
            }
            const int kg = k_begin + kb * G_BK + kr;
            if (live && ch == 0 && kg < k_end) atomicAdd(loss.dots + kg, acc);
          }
#ifdef OMLDM_GRAM_TIMING
          t_work += clock64() - t1;
          if (warp == 2 && r == 0) GRAM_TRACE(4, kb);
#endif
        }
      }
#ifdef OMLDM_GRAM_TIMING
      if (warp == 2 && r == 0) {
        g_gram_stamp[(blockIdx.y * gridDim.x + blockIdx.x) % 160 * 16 + 10] = t_wait;
        g_gram_stamp[(blockIdx.y * gridDim.x + blockIdx.x) % 160 * 16 + 11] = t_work;
      }
#endif
      nvalid = __reduce_add_sync(0xffffffffu, nvalid);
      if (r == 0 && nvalid) atomicAdd(loss.words + blockIdx.y, nvalid);
      __threadfence();                              // this warp's adds before the CTA's arrival below
      named_bar_arrive(2, 288);                     // warp 0 collects the eight loss warps
    }
    if (loss_on) named_bar_sync(3, 320);            // warps 2..11: the epilogue below reuses the stage memory - not before
                                                    // the slowest loss warp has read its last stage
    if (warp >= 4 && warp < 8) {
    __syncwarp();
    mbar_wait(accum_bar, 0);
    if (warp == 4) GRAM_STAMP(3);
    tc_fence_after();
#pragma unroll 1
    for (int c = 0; c < 8; ++c) {
      const int col0 = n0 + c * 32;
      if (col0 >= n || row0 >= n || col0 + 31 < row0) continue;   // outside the matrix / strictly below the diagonal
      float v[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, v);
      tmem_ld_wait();
      uint8_t* slot = smem + (q * 8 + c) * 4096;
      const uint32_t srow = smem_u32(slot) + r * 128;          // shared-space stores (the pointer form compiled to generic ST.E.128)
#pragma unroll
      for (int j = 0; j < 8; ++j)
        sts_v4(srow + ((j ^ (r & 7)) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      fence_proxy_async();
      __syncwarp();
      if (r == 0) {
        tma_reduce_add_2d(&tmC, slot, col0, row0);
        tma_commit_group();
      }
    }
    if (warp == 4) GRAM_STAMP(4);
    if (r == 0) tma_wait_group0();
    __syncwarp();
    if (warp == 4) GRAM_STAMP(5);

    if (push.enabled == 1) {
      // -------------------------------------------------------- hub push, tail mode: by the last CTA of this half tile
      named_bar_sync(1, 128);                      // the reductions of all four epilogue warps are complete
      if (threadIdx.x == 128) {
        __threadfence();
        uint32_t* cnt = scratch + GS_TILE0 + 2 * tile + rank;
        const uint32_t prev = atomicAdd(cnt, 1u);
        const bool last = prev == gridDim.y - 1;
        if (last) *cnt = 0;                        // self-reset for the next launch
        *s_flag = last ? 1u : 0u;
      }
      named_bar_sync(1, 128);
      if (*s_flag) {
        __threadfence();
        if (r == 0 && row0 < n) {
          int nblk = 0;
          for (int c = 0; c < 8; ++c) {
            const int col0 = n0 + c * 32;
            if (col0 >= n || col0 + 31 < row0) continue;
            ++nblk;
          }
          if (nblk) {
            mbar_expect_tx(&push_bar[q], nblk * 4096);
            for (int c = 0; c < 8; ++c) {
              const int col0 = n0 + c * 32;
              if (col0 >= n || col0 + 31 < row0) continue;
              tma_load_2d(smem + (q * 8 + c) * 4096, &tmC, &push_bar[q], col0, row0);
            }
            mbar_wait(&push_bar[q], 0);
            const CUtensorMap* om = &push.owner[row0 / push.rows_per_owner];
            for (int c = 0; c < 8; ++c) {
              const int col0 = n0 + c * 32;
              if (col0 >= n || col0 + 31 < row0) continue;
              tma_store_2d(om, smem + (q * 8 + c) * 4096, col0, row0);    // into this worker's slot at the owner (NVLink)
            }
            tma_commit_group();
            tma_wait_group_read0();                // the stores have read smem; their arrival is ordered by the kernel end
          }
        }
      }
    }
    }   // epilogue warps
   }
  }

  tc_fence_before();
  cluster_sync_all();                 // the peer's MMAs no longer read this CTA's smem; both are done with TMEM
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair<G_TMEM_COLS>(tmem_base);
    GRAM_STAMP(6);
    GRAM_STAMP(9);
  }
}

// ------------------------------------------------------------------ host side
static int encode_z(CUtensorMap* tm, const void* Z, int K, int ld) {
  // [K, ld] halves row-major viewed as {64 halves, K records, ld / 64 chunks}; one box = 2 chunks x 64 records
  cuuint64_t dims[3] = {64, static_cast<cuuint64_t>(K), static_cast<cuuint64_t>(ld / 64)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(ld) * 2, 128};
  cuuint32_t box[3] = {64, static_cast<cuuint32_t>(G_BK), 2};
  cuuint32_t estr[3] = {1, 1, 1};
  return tensormap_encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(Z), dims, strides, box, estr,
                          CU_TENSOR_MAP_SWIZZLE_128B);
}
static int encode_g(CUtensorMap* tm, float* C, int rows, int cols, int ldc) {
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ldc) * 4};
  cuuint32_t box[2] = {32, 32};
  cuuint32_t estr[2] = {1, 1};
  return tensormap_encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, C, dims, strides, box, estr, CU_TENSOR_MAP_SWIZZLE_128B);
}

}  // namespace omldm

using namespace omldm;

extern "C" size_t omldm_gram_f16_scratch_words(void) { return 512; }

#ifdef OMLDM_GRAM_TIMING
extern "C" int omldm_gram_f16_trace(long long* out_host) {
  return cudaMemcpyFromSymbol(out_host, g_gram_trace, sizeof(long long) * 2 * 5 * 24) == cudaSuccess ? 0 : -1;
}
extern "C" int omldm_gram_f16_stamps(long long* out_host) {
  return cudaMemcpyFromSymbol(out_host, g_gram_stamp, sizeof(long long) * 160 * 16) == cudaSuccess ? 0 : -1;
}
#endif

// The batch-axis split omldm_gram_f16 uses for (K, n, ksplit): records per split (0 if K <= 0).
extern "C" int omldm_gram_f16_rows_per_split(int K, int n, int ksplit) {
  if (K <= 0 || n < 1) return 0;
  const int T = (n + 255) / 256;
  const int tiles = T * (T + 1) / 2;
  if (ksplit <= 0) {
    ksplit = 74 / tiles;
    if (ksplit < 1) ksplit = 1;
  }
  const int max_split = (K + 4 * G_BK - 1) / (4 * G_BK);
  if (ksplit > max_split) ksplit = max_split;
  return ((K + ksplit - 1) / ksplit + G_BK - 1) / G_BK * G_BK;
}

extern "C" size_t omldm_gram_f16_loss_words(void) { return 3 * 128; }

extern "C" int omldm_gram_f16(const void* Z, int ld, int K, int n, float* C, int ldc, const omldm_gram_peers* peers,
                              uint32_t* scratch, int ksplit, void* stream) {
  return omldm_gram_f16_loss(Z, ld, K, n, C, ldc, peers, scratch, ksplit, nullptr, stream);
}

extern "C" int omldm_gram_f16_loss(const void* Z, int ld, int K, int n, float* C, int ldc, const omldm_gram_peers* peers,
                                   uint32_t* scratch, int ksplit, const omldm_gram_loss* lossd, void* stream) {
  if (K <= 0) return 0;
  if (ld < 64 || (ld & 63)) return set_error_msg("gram_f16: the row pitch must be a multiple of 64 halves");
  if (reinterpret_cast<uintptr_t>(Z) & 127) return set_error_msg("gram_f16: rows must be 128-byte aligned");
  if (n < 1 || n > ld || n > ldc) return set_error_msg("gram_f16: n exceeds the row pitch or the pitch of C");
  if ((ldc & 3) || (reinterpret_cast<uintptr_t>(C) & 15)) return set_error_msg("gram_f16: C needs a 16-byte aligned base and pitch");
  const int T = (n + 255) / 256;
  const int tiles = T * (T + 1) / 2;
  if (GS_TILE0 + 2 * tiles > 512) return set_error_msg("gram_f16: n too large for the scratch words (n <= 5632)");
  if (peers && !scratch) return set_error_msg("gram_f16: the hub push needs the scratch words");
  CUtensorMap tmZ, tmC;
  if (encode_z(&tmZ, Z, K, ld)) return -1;
  if (encode_g(&tmC, C, n, n, ldc)) return -1;
  GramPush push;
  memset(&push, 0, sizeof(push));
  if (peers) {
    if (peers->world < 1 || peers->world > 16) return set_error_msg("gram_f16: world must be in [1,16]");
    if (peers->rows_per_owner < 32 || (peers->rows_per_owner & 31) ||
        static_cast<long long>(peers->rows_per_owner) * peers->world < n)
      return set_error_msg("gram_f16: rows_per_owner must be a multiple of 32 and cover n");
    push.world = peers->world; push.rank = peers->rank; push.rows_per_owner = peers->rows_per_owner;
    push.enabled = peers->lead ? 2 : 1;
    for (int p = 0; p < peers->world; ++p)
      if (encode_g(&push.owner[p], peers->owner_base[p], n, n, ldc)) return -1;
    for (int p = peers->world; p < 16; ++p) push.owner[p] = push.owner[0];
  }
  GramLoss loss;
  memset(&loss, 0, sizeof(loss));
  if (lossd) {
    if (!lossd->wt || !lossd->dots || !lossd->words || !lossd->stats) return set_error_msg("gram_f16: incomplete loss descriptor");
    if (lossd->d < 0 || lossd->d + 3 > n) return set_error_msg("gram_f16: loss needs n >= d + 3");
    if (reinterpret_cast<uintptr_t>(lossd->dots) & 15) return set_error_msg("gram_f16: dots must be 16-byte aligned");
    if (lossd->curve_ring && (lossd->curve_cap <= 0 || !lossd->curve_counter)) return set_error_msg("gram_f16: curve ring without counter");
    loss.wt = lossd->wt; loss.dots = lossd->dots; loss.words = lossd->words; loss.stats = lossd->stats;
    loss.curve_ring = lossd->curve_ring; loss.curve_counter = lossd->curve_counter; loss.curve_cap = lossd->curve_cap;
    loss.d = lossd->d; loss.enabled = 1;
  }
  // split the batch axis so that the pairs fill the 74 SM pairs; at least 4 stages of work per CTA
  const int k_per_split = omldm_gram_f16_rows_per_split(K, n, ksplit);
  const int gy = (K + k_per_split - 1) / k_per_split;
  static bool attr_set[16] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 15]) {                       // once per device, never inside a graph capture
    cudaError_t e = cudaFuncSetAttribute(gram_f16_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM + G_LEAD_BYTES);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(gram_f16_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM + G_LEAD_BYTES);
    if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(gram_f16)", e);
    // leave the SM at its largest shared-memory carve-out so that blocks of the loss pass fit beside a Gram CTA
    cudaFuncSetAttribute(gram_f16_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(gram_f16_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    attr_set[dev & 15] = true;
  }
  dim3 grid(2 * tiles, gy);
  const int smem_bytes = G_SMEM + (push.enabled == 2 ? G_LEAD_BYTES : 0);
  if (loss.enabled && gy > 128) return set_error_msg("gram_f16: more than 128 batch slices with the loss inside");
  if (loss.enabled)
    gram_f16_kernel<true><<<grid, G_THREADS, smem_bytes, static_cast<cudaStream_t>(stream)>>>(tmZ, tmC, push, loss, n, K,
                                                                                            k_per_split, T, scratch);
  else
    // (warps 8..11 exist only with the loss inside: 256 threads keep two blocks of the separate loss pass beside a CTA)
    gram_f16_kernel<false><<<grid, 256, smem_bytes, static_cast<cudaStream_t>(stream)>>>(tmZ, tmC, push, loss, n, K,
                                                                                       k_per_split, T, scratch);
  return check_launch("gram_f16_kernel");
}
