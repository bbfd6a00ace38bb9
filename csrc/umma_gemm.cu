// tcgen05 / TMEM / TMA GEMM core for the online learners (sm_100a only).
//
// Two operand arrangements, both fp32 in global memory and consumed by the tensor core as tf32
// with fp32 accumulation in TMEM:
//
//   gemm_tn   : C[M,N] (+)= A[M,K] * B[N,K]^T      (both operands K-major; "scores" shape:
//               X W^T for MultiClassPA / NN forward, X C^T for k-means distances)
//   gemm_outer: C[M,N] (+)= A[K,M]^T * B[K,N]      (both operands MN-major; "outer-product" shape:
//               Z^T Z Gram for ridge regression, T^T X for MultiClassPA / k-means / NN updates;
//               K is the mini-batch axis so records are consumed exactly as they sit in the
//               staged ring slot, no transpose pass)
//   gemm_nn   : C[M,N] (+)= A[M,K] * B[K,N]        (A K-major, B MN-major, both row-major in memory:
//               the back-propagated gradient dL/dA = G W of the NN learner)
//
// Replaces the per-record Breeze `Double` arithmetic behind `node.receiveTuple`
// (reference call site: src/main/scala/omldm/operators/spoke/FlinkSpoke.scala:101); the
// reference has no GEMM of its own.
//
// Structure: one CTA = one 128 x BN output tile x one K split.  Warp 0 = TMA producer,
// warp 1 = TMEM allocator + single-thread tcgen05.mma issuer, warps 2..5 = epilogue
// (tcgen05.ld -> registers -> st / red.add.v4 to global).  4-stage smem ring, mbarrier
// full/empty pairs, tcgen05.commit releases stages and publishes the accumulator.
#include "ptx.cuh"
#include "host_common.h"
#include <cstring>

namespace omldm {

constexpr int BM = 128;
constexpr int BK = 32;  // tf32 elements per stage along K (= one 128-byte swizzle row)
constexpr int STAGES = 4;
constexpr int GEMM_THREADS = 192;

enum : int { GEMM_ACCUMULATE = 1, GEMM_UPPER_ONLY = 2, GEMM_MN_SLAB2D = 4, GEMM_TMA_EPILOGUE = 8, GEMM_NO_TMA_EPILOGUE = 16 };

template <int BN>
struct GemmSmem {
  static constexpr int A_BYTES = BM * BK * 4;
  static constexpr int B_BYTES = BN * BK * 4;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int TOTAL = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

// Per-peer output maps of the scatter variant: the 128-row tile block-row [m0, m0 + 128) is reduced into the
// accumulation buffer of the GPU that owns those rows (rows_per_owner rows each, a multiple of 32).
struct PeerMaps {
  CUtensorMap tm[16];
  int rows_per_owner;
};

// K-means assignment as the epilogue of the distance GEMM S = X C^T (SURVEY K-L5): the scores never leave the SM.
// Each epilogue thread owns one record (one TMEM lane): pass 1 walks the accumulator columns and keeps
// argmin_j (||c_j||^2 - 2 S_ij); pass 2 writes the one-hot row of T in 32 x 32 blocks through shared memory and TMA
// (tmC maps T), which the update GEMM sums += T^T Z consumes.  stats += {inertia, -, #records}.
constexpr int KM_RI = 4;   // records whose exact distance is in flight per warp (k-means step: 46.7 us with 1, 34.4 with 4, 35.8 with 8)

struct KMeansEpi {
  const float* cnorm;   // ||c_j||^2, +inf for empty centroids
  const float* Z;       // the staged rows (valid flags; exact distance to the chosen centroid), pitch ldz
  const float* Cm;      // centroids, pitch ldcm
  int ldz, ldcm, d;
  int* assign;          // optional
  double* stats;
};

template <int BN, bool A_MN, bool B_MN, bool PEER>
__device__ __forceinline__ void umma_gemm_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                                               const PeerMaps* pm, float* __restrict__ C, int ldc, int M, int N, int K,
                                               int k_per_split, int flags, const KMeansEpi* km = nullptr) {
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  if ((flags & GEMM_UPPER_ONLY) && m0 > n0 + BN - 1) return;  // tile strictly below the diagonal
  const int k_begin = blockIdx.z * k_per_split;
  const int k_end = min(K, k_begin + k_per_split);
  const int num_kb = (k_end - k_begin + BK - 1) / BK;
  if (num_kb <= 0) return;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  using S = GemmSmem<BN>;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * S::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* accum_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

  const int warp = threadIdx.x >> 5;
  constexpr int TMEM_COLS = BN < 32 ? 32 : BN;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (!PEER && (flags & GEMM_TMA_EPILOGUE)) tma_prefetch_desc(&tmC);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(accum_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  constexpr bool MN_MAJOR = A_MN && B_MN;
  if (warp == 0 && MN_MAJOR && (flags & GEMM_MN_SLAB2D)) {
    // ------------------------------------------------------------ TMA producer, one 2-D box per
    // 32-float slab (lane l loads slab l): A has BM/32 slabs, B has BN/32.
    constexpr int A_SLABS = BM / 32, B_SLABS = BN / 32;
    const int l = static_cast<int>(lane_id());
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < num_kb; ++kb) {
      if (l == 0) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_expect_tx(&full_bar[stage], S::STAGE_BYTES);
      }
      __syncwarp();
      uint8_t* sa = smem + stage * S::STAGE_BYTES;
      uint8_t* sb = sa + S::A_BYTES;
      const int k0 = k_begin + kb * BK;
      if (l < A_SLABS) tma_load_2d(sa + l * (BK * 128), &tmA, &full_bar[stage], m0 + l * 32, k0);
      else if (l < A_SLABS + B_SLABS)
        tma_load_2d(sb + (l - A_SLABS) * (BK * 128), &tmB, &full_bar[stage], n0 + (l - A_SLABS) * 32, k0);
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + stage * S::STAGE_BYTES;
        uint8_t* sb = sa + S::A_BYTES;
        mbar_expect_tx(&full_bar[stage], S::STAGE_BYTES);
        const int k0 = k_begin + kb * BK;
        if constexpr (A_MN) tma_load_3d(sa, &tmA, &full_bar[stage], 0, k0, m0 / 32);
        else tma_load_2d(sa, &tmA, &full_bar[stage], k0, m0);
        if constexpr (B_MN) tma_load_3d(sb, &tmB, &full_bar[stage], 0, k0, n0 / 32);
        else tma_load_2d(sb, &tmB, &full_bar[stage], k0, n0);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = make_idesc_tf32(BM, BN, A_MN, B_MN);
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < num_kb; ++kb) {
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t sa = smem_u32(smem + stage * S::STAGE_BYTES);
        const uint32_t sb = sa + S::A_BYTES;
#pragma unroll
        for (int kk = 0; kk < BK / 8; ++kk) {
          // MN-major: K advance = 8 rows of 128 B; MN atoms (32 floats) are BK*128 B apart; 32-byte-atom
          //           swizzle: K groups are 4 rows (512 B) apart.
          // K-major : K advance = 8 tf32 = 32 B inside the 128-B swizzle row; 8-row groups 1024 B apart.
          const uint64_t da = A_MN ? make_smem_desc_sw128(sa + kk * 1024, BK * 128, 512, 1)
                                   : make_smem_desc_sw128(sa + kk * 32, 16, 1024);
          const uint64_t db = B_MN ? make_smem_desc_sw128(sb + kk * 1024, BK * 128, 512, 1)
                                   : make_smem_desc_sw128(sb + kk * 32, 16, 1024);
          umma_tf32(tmem_base, da, db, idesc, (kb | kk) != 0);
        }
        umma_commit(&empty_bar[stage]);                 // frees this smem stage when MMAs retire
        if (kb == num_kb - 1) umma_commit(accum_bar);   // accumulator complete
      }
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  } else {
    // ------------------------------------------------------------ epilogue (warps 2..5)
    const int q = warp & 3;  // TMEM lane quadrant this warp may read
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    const int row = m0 + q * 32 + static_cast<int>(lane_id());
    const bool accumulate = (flags & GEMM_ACCUMULATE) != 0;
    if (km != nullptr) {
      // ---- fused k-means assignment (N <= BN: one tile holds every centroid of a record)
      const uint32_t r = lane_id();
      const bool in_range = row < M;
      const float* zr = km->Z + static_cast<size_t>(in_range ? row : 0) * km->ldz;
      const bool valid = in_range && zr[km->d] != 0.f;
      float best = INFINITY;
      int bj = 0;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        if (c * 32 >= N) break;                      // warp-uniform
        float v[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, v);
        // ||c_j||^2 of this chunk: one coalesced load per chunk, handed round by shuffles (it was one load per score)
        const int mycol = c * 32 + static_cast<int>(r);
        const float cn_mine = mycol < N ? __ldg(km->cnorm + mycol) : INFINITY;
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const float cn = __shfl_sync(0xffffffffu, cn_mine, j);
          const float dist = cn - 2.f * v[j];        // +inf beyond N and for empty centroids: never the minimum
          if (dist < best) { best = dist; bj = c * 32 + j; }
        }
      }
      // The tf32 scores decide the argmin; the reported distance to the chosen centroid is recomputed in fp32
      // (||x||^2 - 2 x.c + ||c||^2 cancels, and tf32 truncation would bias the sum).  The warp walks its 32 records
      // together: every row and its centroid are read with coalesced 16-byte loads (a lane reading its own 512-byte row
      // touched 32 sectors per load instruction), each lane keeps the sum of its own record.
      double inertia = 0.0;
      {
        const int d4 = km->d >> 2;
        float mine = 0.f;
        const int vmask = static_cast<int>(__ballot_sync(0xffffffffu, valid));
#pragma unroll 1
        for (int i0 = 0; i0 < 32; i0 += KM_RI) {     // KM_RI records at a time: 2 KM_RI independent loads in flight
          float acc[KM_RI];
          const float* zi[KM_RI];
          const float* ci[KM_RI];
#pragma unroll
          for (int k = 0; k < KM_RI; ++k) {
            const int i = i0 + k;
            const bool ok = (vmask >> i) & 1;
            const int bj_i = __shfl_sync(0xffffffffu, bj, i);
            zi[k] = km->Z + static_cast<size_t>(ok ? m0 + q * 32 + i : 0) * km->ldz;
            ci[k] = km->Cm + static_cast<size_t>(ok ? bj_i : 0) * km->ldcm;
            acc[k] = 0.f;
          }
          for (int e = static_cast<int>(r); e < d4; e += 32) {
            float4 t[KM_RI], u[KM_RI];
#pragma unroll
            for (int k = 0; k < KM_RI; ++k) {
              t[k] = reinterpret_cast<const float4*>(zi[k])[e];
              u[k] = __ldg(reinterpret_cast<const float4*>(ci[k]) + e);
            }
#pragma unroll
            for (int k = 0; k < KM_RI; ++k) {
              const float a0 = t[k].x - u[k].x, a1 = t[k].y - u[k].y, a2 = t[k].z - u[k].z, a3 = t[k].w - u[k].w;
              acc[k] = fmaf(a0, a0, fmaf(a1, a1, fmaf(a2, a2, fmaf(a3, a3, acc[k]))));
            }
          }
          for (int e = (d4 << 2) + static_cast<int>(r); e < km->d; e += 32) {
#pragma unroll
            for (int k = 0; k < KM_RI; ++k) {
              const float a0 = zi[k][e] - ci[k][e];
              acc[k] = fmaf(a0, a0, acc[k]);
            }
          }
#pragma unroll
          for (int k = 0; k < KM_RI; ++k) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
            if (static_cast<int>(r) == i0 + k && ((vmask >> (i0 + k)) & 1)) mine = acc[k];
          }
        }
        inertia = static_cast<double>(mine);
      }
      if (in_range && km->assign) km->assign[row] = valid ? bj : -1;
      // one-hot rows of T, 32 x 32 blocks through swizzled shared memory and TMA (full 128-byte lines)
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        if (c * 32 >= ldc || m0 + q * 32 >= M) break;   // ldc = pitch of T: pad columns are written as zeros too
        uint8_t* slot = smem + (q * (BN / 32) + c) * 4096;
        const int hot = (valid && (bj >> 5) == c) ? (bj & 31) : -1;
        const uint32_t srow = smem_u32(slot) + r * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          sts_v4(srow + ((j ^ (r & 7)) << 4), hot == 4 * j ? 1.f : 0.f, hot == 4 * j + 1 ? 1.f : 0.f,
                 hot == 4 * j + 2 ? 1.f : 0.f, hot == 4 * j + 3 ? 1.f : 0.f);
        fence_proxy_async();
        __syncwarp();
        if (r == 0) {
          tma_store_2d(&tmC, slot, c * 32, m0 + q * 32);
          tma_commit_group();
        }
      }
      if (r == 0) tma_wait_group0();
      __syncwarp();
      // statistics: one pair of atomics per warp
      double w_in = inertia;
      int w_n = valid ? 1 : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        w_in += __shfl_xor_sync(0xffffffffu, w_in, o);
        w_n += __shfl_xor_sync(0xffffffffu, w_n, o);
      }
      if (r == 0 && w_n) { atomicAdd(&km->stats[0], w_in); atomicAdd(&km->stats[2], static_cast<double>(w_n)); }
    } else if (flags & GEMM_TMA_EPILOGUE) {
      // TMEM -> registers -> swizzled smem tile (32 rows x 128 B per warp and 32-column chunk) ->
      // TMA store / TMA reduce-add: the TMA unit writes full 128-B lines and performs the fp32
      // adds at L2 (the vector red.global is lowered to scalar REDGs on sm_100a).  The operand
      // stages are free by now (accum_bar fired), each (warp, chunk) owns a private 4 KB slot.
      const uint32_t r = lane_id();
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int col0 = n0 + c * 32;
        if (col0 >= N || m0 + q * 32 >= M) break;   // warp-uniform
        float v[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, v);
        tmem_ld_wait();
        uint8_t* slot = smem + (q * (BN / 32) + c) * 4096;
        const uint32_t srow = smem_u32(slot) + r * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          sts_v4(srow + ((j ^ (r & 7)) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        fence_proxy_async();
        __syncwarp();
        if (r == 0) {
          if constexpr (PEER) {
            // reduce-scatter inside the GEMM: this 32-row strip goes to its owner's buffer over NVLink
            const int row0 = m0 + q * 32;
            tma_reduce_add_2d(&pm->tm[row0 / pm->rows_per_owner], slot, col0, row0);
          } else {
            if (accumulate) tma_reduce_add_2d(&tmC, slot, col0, m0 + q * 32);
            else tma_store_2d(&tmC, slot, col0, m0 + q * 32);
          }
          tma_commit_group();
        }
      }
      if (r == 0) tma_wait_group0();
      __syncwarp();
    } else {
    const bool vec_ok = ((ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(C) & 15) == 0);
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      float v[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, v);
      tmem_ld_wait();
      if (row < M) {
        const int col0 = n0 + c * 32;
        float* dst = C + static_cast<size_t>(row) * ldc + col0;
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          if (vec_ok && col0 + j + 3 < N) {
            if (accumulate) red_add_v4(dst + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
            else *reinterpret_cast<float4*>(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
          } else {
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              if (col0 + j + t < N) {
                if (accumulate) atomicAdd(dst + j + t, v[j + t]);
                else dst[j + t] = v[j + t];
              }
            }
          }
        }
      }
    }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<TMEM_COLS>(tmem_base);
  }
}

template <int BN, bool MN_MAJOR>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
umma_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, float* __restrict__ C, int ldc, int M, int N, int K,
                 int k_per_split, int flags) {
  umma_gemm_body<BN, MN_MAJOR, MN_MAJOR, false>(tmA, tmB, tmC, nullptr, C, ldc, M, N, K, k_per_split, flags);
}

// distance GEMM with the k-means assignment as its epilogue (tmC maps the one-hot matrix T)
template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
umma_kmeans_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                   const __grid_constant__ CUtensorMap tmT, const __grid_constant__ KMeansEpi km, int ldt, int M, int N,
                   int K) {
  umma_gemm_body<BN, false, false, false>(tmA, tmB, tmT, nullptr, nullptr, ldt, M, N, K, (K + BK - 1) / BK * BK, 0, &km);
}

// A K-major, B MN-major (both row-major in memory): C = A B
template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
umma_gemm_nn_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, float* __restrict__ C, int ldc, int M, int N, int K,
                    int k_per_split, int flags) {
  umma_gemm_body<BN, false, true, false>(tmA, tmB, tmC, nullptr, C, ldc, M, N, K, k_per_split, flags);
}

// Gram / outer-product GEMM whose epilogue IS the reduce-scatter: every output strip is TMA-reduce-added into the
// buffer of the GPU that owns its rows (peer memory), so the transfer rides under the MMAs tile by tile.
template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
umma_gemm_scatter_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const __grid_constant__ PeerMaps pm, int M, int N, int K, int k_per_split, int flags) {
  umma_gemm_body<BN, true, true, true>(tmA, tmB, pm.tm[0], &pm, nullptr, 0, M, N, K, k_per_split,
                                 flags | GEMM_ACCUMULATE | GEMM_TMA_EPILOGUE);
}

// ------------------------------------------------------------------ host side
static int encode_kmajor(CUtensorMap* tm, const float* base, int rows, int k, int ld, int box_rows) {
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(k), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 4};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(BK), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  return tensormap_encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims,
                          strides, box, estr, CU_TENSOR_MAP_SWIZZLE_128B);
}
static int encode_mnmajor_2d(CUtensorMap* tm, const float* base, int k_rows, int ld) {
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(ld), static_cast<cuuint64_t>(k_rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 4};
  cuuint32_t box[2] = {32, static_cast<cuuint32_t>(BK)};
  cuuint32_t estr[2] = {1, 1};
  return tensormap_encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims,
                          strides, box, estr, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
}
static int encode_mnmajor(CUtensorMap* tm, const float* base, int k_rows, int ld, int box_chunks) {
  // view [k_rows, ld] row-major as {32 floats, k_rows, ld/32 chunks}; one box = box_chunks slabs
  // of (BK rows x 128 B), which is exactly the MN-major SWIZZLE_128B_BASE32B canonical layout.
  cuuint64_t dims[3] = {32, static_cast<cuuint64_t>(k_rows), static_cast<cuuint64_t>(ld / 32)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(ld) * 4, 128};
  cuuint32_t box[3] = {32, static_cast<cuuint32_t>(BK), static_cast<cuuint32_t>(box_chunks)};
  cuuint32_t estr[3] = {1, 1, 1};
  return tensormap_encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims,
                          strides, box, estr, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
}

static int encode_c(CUtensorMap* tm, float* C, int M, int N, int ldc) {
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(N), static_cast<cuuint64_t>(M)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ldc) * 4};
  cuuint32_t box[2] = {32, 32};
  cuuint32_t estr[2] = {1, 1};
  return tensormap_encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, C, dims, strides, box, estr,
                          CU_TENSOR_MAP_SWIZZLE_128B);
}

template <int BN, bool MN>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, float* C, int ldc, int M,
                       int N, int K, int ksplit, int flags, cudaStream_t stream) {
  CUtensorMap tmC;
  memset(&tmC, 0, sizeof(tmC));
  // the TMA epilogue needs a 16-byte aligned C with a pitch that is a multiple of 4 floats
  if (!(flags & GEMM_NO_TMA_EPILOGUE) && (ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0) {
    if (encode_c(&tmC, C, M, N, ldc)) return -1;
    flags |= GEMM_TMA_EPILOGUE;
  }
  static bool attr_set[16] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  auto kern = umma_gemm_kernel<BN, MN>;
  constexpr int smem = GemmSmem<BN>::TOTAL;
  if (!attr_set[dev & 15]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(gemm)", e);
    attr_set[dev & 15] = true;
  }
  if (ksplit < 1) ksplit = 1;
  int k_per_split = ((K + ksplit - 1) / ksplit + BK - 1) / BK * BK;
  dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN, (K + k_per_split - 1) / k_per_split);
  kern<<<grid, GEMM_THREADS, smem, stream>>>(tmA, tmB, tmC, C, ldc, M, N, K, k_per_split, flags);
  return check_launch("umma_gemm_kernel");
}

template <bool MN>
static int dispatch_bn(const CUtensorMap& tmA, const CUtensorMap& tmB, int bn, float* C, int ldc,
                       int M, int N, int K, int ksplit, int flags, cudaStream_t stream) {
  switch (bn) {
    case 32: return launch_gemm<32, MN>(tmA, tmB, C, ldc, M, N, K, ksplit, flags, stream);
    case 64: return launch_gemm<64, MN>(tmA, tmB, C, ldc, M, N, K, ksplit, flags, stream);
    case 128: return launch_gemm<128, MN>(tmA, tmB, C, ldc, M, N, K, ksplit, flags, stream);
    default: return launch_gemm<256, MN>(tmA, tmB, C, ldc, M, N, K, ksplit, flags, stream);
  }
}
static int pick_bn(int N) { return N <= 32 ? 32 : N <= 64 ? 64 : N <= 128 ? 128 : 256; }

}  // namespace omldm

using namespace omldm;

extern "C" int omldm_gemm_tn(const float* A, int lda, const float* B, int ldb, float* C, int ldc,
                             int M, int N, int K, int ksplit, int flags, void* stream) {
  if ((lda & 3) || (ldb & 3)) return set_error_msg("gemm_tn: lda/ldb must be multiples of 4 floats");
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15))
    return set_error_msg("gemm_tn: operands must be 16-byte aligned");
  if (ksplit > 1 && !(flags & GEMM_ACCUMULATE))
    return set_error_msg("gemm_tn: split-K requires the accumulate flag");
  const int bn = pick_bn(N);
  CUtensorMap tmA, tmB;
  if (encode_kmajor(&tmA, A, M, K, lda, BM)) return -1;
  if (encode_kmajor(&tmB, B, N, K, ldb, bn)) return -1;
  return dispatch_bn<false>(tmA, tmB, bn, C, ldc, M, N, K, ksplit, flags,
                            static_cast<cudaStream_t>(stream));
}

extern "C" int omldm_gemm_outer(const float* A, int lda, const float* B, int ldb, float* C, int ldc,
                                int M, int N, int K, int ksplit, int flags, void* stream) {
  if ((lda & 31) || (ldb & 31))
    return set_error_msg("gemm_outer: lda/ldb must be multiples of 32 floats (padded record rows)");
  if ((reinterpret_cast<uintptr_t>(A) & 127) || (reinterpret_cast<uintptr_t>(B) & 127))
    return set_error_msg("gemm_outer: operands must be 128-byte aligned");
  if (M > lda || N > ldb) return set_error_msg("gemm_outer: M/N exceed the row pitch");
  if (ksplit > 1 && !(flags & GEMM_ACCUMULATE))
    return set_error_msg("gemm_outer: split-K requires the accumulate flag");
  const int bn = pick_bn(N);
  CUtensorMap tmA, tmB;
  if (flags & GEMM_MN_SLAB2D) {
    if (encode_mnmajor_2d(&tmA, A, K, lda)) return -1;
    if (encode_mnmajor_2d(&tmB, B, K, ldb)) return -1;
  } else {
    if (encode_mnmajor(&tmA, A, K, lda, BM / 32)) return -1;
    if (encode_mnmajor(&tmB, B, K, ldb, bn / 32)) return -1;
  }
  return dispatch_bn<true>(tmA, tmB, bn, C, ldc, M, N, K, ksplit, flags,
                           static_cast<cudaStream_t>(stream));
}

// Fused k-means assignment: T[B, ldt] <- one-hot(argmin_j ||x_i - c_j||^2), stats += {inertia, -, #valid}; centroids
// Cm[Kc, ldc] (K-major rows of d floats), cnorm[Kc].  Kc <= 256, ldt % 32 == 0 and >= Kc.
extern "C" int omldm_kmeans_assign_fused(const float* Z, int ld, int d, int B, const float* Cm, int ldcm, const float* cnorm,
                                         int Kc, float* T, int ldt, int* assign, double* stats, void* stream) {
  if (B <= 0) return 0;
  if (Kc < 1 || Kc > 256) return set_error_msg("kmeans_assign_fused: 1 <= K <= 256 (one tile holds all centroids)");
  if ((ld & 3) || (ldcm & 3) || (ldt & 31) || ldt < Kc) return set_error_msg("kmeans_assign_fused: bad pitches");
  if ((reinterpret_cast<uintptr_t>(Z) & 15) || (reinterpret_cast<uintptr_t>(Cm) & 15) || (reinterpret_cast<uintptr_t>(T) & 15))
    return set_error_msg("kmeans_assign_fused: operands must be 16-byte aligned");
  const int bn = pick_bn(ldt < 256 ? ldt : 256);       // covers the pad columns of T as well (written as zeros)
  CUtensorMap tmA, tmB, tmT;
  if (encode_kmajor(&tmA, Z, B, d, ld, BM)) return -1;
  if (encode_kmajor(&tmB, Cm, Kc, d, ldcm, bn)) return -1;
  if (encode_c(&tmT, T, B, ldt, ldt)) return -1;
  KMeansEpi km;
  km.cnorm = cnorm; km.Z = Z; km.Cm = Cm; km.ldz = ld; km.ldcm = ldcm; km.d = d; km.assign = assign; km.stats = stats;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  static bool attr_set[16][4] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  auto go = [&](auto kern, int smem, int BNv) -> int {
    bool& done = attr_set[dev & 15][BNv == 32 ? 0 : BNv == 64 ? 1 : BNv == 128 ? 2 : 3];
    if (!done) {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(kmeans)", e);
      done = true;
    }
    dim3 grid((B + BM - 1) / BM, 1, 1);
    kern<<<grid, GEMM_THREADS, smem, st>>>(tmA, tmB, tmT, km, ldt, B, Kc, d);
    return check_launch("umma_kmeans_kernel");
  };
  switch (bn) {
    case 32: return go(umma_kmeans_kernel<32>, GemmSmem<32>::TOTAL, 32);
    case 64: return go(umma_kmeans_kernel<64>, GemmSmem<64>::TOTAL, 64);
    case 128: return go(umma_kmeans_kernel<128>, GemmSmem<128>::TOTAL, 128);
    default: return go(umma_kmeans_kernel<256>, GemmSmem<256>::TOTAL, 256);
  }
}

// C[M,N] (+)= A[M,K] B[K,N]: A row-major (pitch lda, multiple of 4 floats), B row-major (pitch ldb, multiple of 32
// floats, 128-byte aligned, N <= ldb).
extern "C" int omldm_gemm_nn(const float* A, int lda, const float* B, int ldb, float* C, int ldc, int M, int N, int K,
                             int ksplit, int flags, void* stream) {
  if (lda & 3) return set_error_msg("gemm_nn: lda must be a multiple of 4 floats");
  if (ldb & 31) return set_error_msg("gemm_nn: ldb must be a multiple of 32 floats (padded rows)");
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 127))
    return set_error_msg("gemm_nn: A must be 16-byte and B 128-byte aligned");
  if (N > ldb) return set_error_msg("gemm_nn: N exceeds the row pitch of B");
  if (ksplit > 1 && !(flags & GEMM_ACCUMULATE)) return set_error_msg("gemm_nn: split-K requires the accumulate flag");
  const int bn = pick_bn(N);
  CUtensorMap tmA, tmB, tmC;
  if (encode_kmajor(&tmA, A, M, K, lda, BM)) return -1;
  if (encode_mnmajor(&tmB, B, K, ldb, bn / 32)) return -1;
  memset(&tmC, 0, sizeof(tmC));
  flags &= ~GEMM_MN_SLAB2D;
  if (!(flags & GEMM_NO_TMA_EPILOGUE) && (ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0) {
    if (encode_c(&tmC, C, M, N, ldc)) return -1;
    flags |= GEMM_TMA_EPILOGUE;
  }
  if (ksplit < 1) ksplit = 1;
  const int k_per_split = ((K + ksplit - 1) / ksplit + BK - 1) / BK * BK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  static bool attr_set[16][4] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  auto go = [&](auto kern, int smem, int BNv) -> int {
    bool& done = attr_set[dev & 15][BNv == 32 ? 0 : BNv == 64 ? 1 : BNv == 128 ? 2 : 3];
    if (!done) {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(gemm_nn)", e);
      done = true;
    }
    dim3 grid((M + BM - 1) / BM, (N + BNv - 1) / BNv, (K + k_per_split - 1) / k_per_split);
    kern<<<grid, GEMM_THREADS, smem, st>>>(tmA, tmB, tmC, C, ldc, M, N, K, k_per_split, flags);
    return check_launch("umma_gemm_nn_kernel");
  };
  switch (bn) {
    case 32: return go(umma_gemm_nn_kernel<32>, GemmSmem<32>::TOTAL, 32);
    case 64: return go(umma_gemm_nn_kernel<64>, GemmSmem<64>::TOTAL, 64);
    case 128: return go(umma_gemm_nn_kernel<128>, GemmSmem<128>::TOTAL, 128);
    default: return go(umma_gemm_nn_kernel<256>, GemmSmem<256>::TOTAL, 256);
  }
}

// C_peers[p] = base of GPU p's accumulation matrix (M x N, pitch ldc, peer-mapped; the caller picks the rotating
// copy).  Rows [p * rows_per_owner, (p + 1) * rows_per_owner) of the product are reduced into C_peers[p] only.
extern "C" int omldm_gemm_outer_scatter(const float* A, int lda, const float* B, int ldb, float* const* C_peers,
                                        int n_peers, int rows_per_owner, int ldc, int M, int N, int K, int ksplit,
                                        int flags, void* stream) {
  if ((lda & 31) || (ldb & 31))
    return set_error_msg("gemm_outer_scatter: lda/ldb must be multiples of 32 floats (padded record rows)");
  if ((reinterpret_cast<uintptr_t>(A) & 127) || (reinterpret_cast<uintptr_t>(B) & 127))
    return set_error_msg("gemm_outer_scatter: operands must be 128-byte aligned");
  if (M > lda || N > ldb) return set_error_msg("gemm_outer_scatter: M/N exceed the row pitch");
  if (n_peers < 1 || n_peers > 16) return set_error_msg("gemm_outer_scatter: 1..16 peers");
  if (rows_per_owner < 32 || (rows_per_owner & 31) || static_cast<long long>(rows_per_owner) * n_peers < M)
    return set_error_msg("gemm_outer_scatter: rows_per_owner must be a multiple of 32 and cover M");
  if (ldc & 3) return set_error_msg("gemm_outer_scatter: ldc must be a multiple of 4 floats");
  if (flags & GEMM_MN_SLAB2D) return set_error_msg("gemm_outer_scatter: the 2-D slab producer is not supported");
  PeerMaps pm;
  memset(&pm, 0, sizeof(pm));
  pm.rows_per_owner = rows_per_owner;
  for (int p = 0; p < n_peers; ++p) {
    if (reinterpret_cast<uintptr_t>(C_peers[p]) & 15) return set_error_msg("gemm_outer_scatter: C must be 16-byte aligned");
    if (encode_c(&pm.tm[p], C_peers[p], M, N, ldc)) return -1;
  }
  for (int p = n_peers; p < 16; ++p) pm.tm[p] = pm.tm[0];
  const int bn = pick_bn(N);
  CUtensorMap tmA, tmB;
  if (encode_mnmajor(&tmA, A, K, lda, BM / 32)) return -1;
  if (encode_mnmajor(&tmB, B, K, ldb, bn / 32)) return -1;
  if (ksplit < 1) ksplit = 1;
  const int k_per_split = ((K + ksplit - 1) / ksplit + BK - 1) / BK * BK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  static bool attr_set[16][4] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  auto go = [&](auto kern, int smem, int BNv) -> int {
    bool& done = attr_set[dev & 15][BNv == 32 ? 0 : BNv == 64 ? 1 : BNv == 128 ? 2 : 3];
    if (!done) {                                   // once per device and tile shape, never inside a graph capture
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(gemm scatter)", e);
      done = true;
    }
    dim3 grid((M + BM - 1) / BM, (N + BNv - 1) / BNv, (K + k_per_split - 1) / k_per_split);
    kern<<<grid, GEMM_THREADS, smem, st>>>(tmA, tmB, pm, M, N, K, k_per_split, flags);
    return check_launch("umma_gemm_scatter_kernel");
  };
  switch (bn) {
    case 32: return go(umma_gemm_scatter_kernel<32>, GemmSmem<32>::TOTAL, 32);
    case 64: return go(umma_gemm_scatter_kernel<64>, GemmSmem<64>::TOTAL, 64);
    case 128: return go(umma_gemm_scatter_kernel<128>, GemmSmem<128>::TOTAL, 128);
    default: return go(umma_gemm_scatter_kernel<256>, GemmSmem<256>::TOTAL, 256);
  }
}
