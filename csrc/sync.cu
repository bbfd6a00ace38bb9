// Fused worker<->hub synchronisation kernels over NVLink peer memory (sm_100a).
//
// The reference moves every model between workers ("spokes") and parameter-server shards
// ("hubs") as serialized messages: spoke->hub through a Flink keyBy shuffle
// (network/FlinkNetwork.scala:260-261,279-281) and hub->spoke through a Kafka topic used as a
// feedback edge (Job.scala:77-87,136-142; FlinkNetwork.scala:262-271,282-291).  Here every GPU
// is one worker and one hub shard; buffers live in symmetric (peer-mapped) memory and ONE kernel
// per protocol step does reduce-scatter + averaging + all-gather (+ safe-zone check) with
// in-kernel peer loads/stores or NVLS multimem instructions.  No NCCL on this path.
//
// Cross-GPU ordering: data is written with plain/relaxed.sys stores, then
// __threadfence_system(), then a flag is published with st.release.sys; consumers poll with
// ld.acquire.sys.  All spin loops are bounded and record a timeout in the status word instead
// of hanging the GPU.
#include "ptx.cuh"
#include "host_common.h"
#include "omldm_kernels.h"
#include <cstring>

namespace omldm {

struct Comm {
  float* const* bufs;
  float* const* hubs;
  uint32_t* const* pads;
  uint32_t* local;
  float* mc_buf;
  int rank, world, emulate;
  int nhub;            // hub shards (HubParallelism): ranks 0 .. nhub-1 own a shard; 0 = every rank does
};

// number of hub shards the model vector is cut into (FlinkSpoke.scala:181-190, key net_hub FlinkLearning.scala:91-94)
__device__ __forceinline__ int hub_count(const Comm& c) { return c.nhub > 0 && c.nhub < c.world ? c.nhub : c.world; }

constexpr int LOCAL_WORDS = 64;  // private scratch words per (emulated) rank
// local scratch word indices
enum { L_DONE = 0, L_DONE2 = 1, L_NORM = 2, L_DECISION = 3, L_DONE3 = 4, L_EPOCH = 5, L_CLOCK = 6, L_STEP = 7, L_ROT = 8, L_QEPOCH = 9, L_GSTEP = 10, L_GEPOCH = 11, L_DONE4 = 12 };

// A sequence number passed as 0 means "device managed": the kernel uses (stored value + 1) and the
// block that finishes last stores it back, so the identical launch can be replayed from a CUDA graph.
__device__ __forceinline__ uint32_t resolve_seq(uint32_t given, const uint32_t* slot) {
  return given ? given : *reinterpret_cast<const volatile uint32_t*>(slot) + 1u;
}

__device__ __forceinline__ int my_rank(const Comm& c) { return c.emulate > 1 ? static_cast<int>(blockIdx.y) : c.rank; }
__device__ __forceinline__ uint32_t* my_local(const Comm& c) { return c.local + my_rank(c) * LOCAL_WORDS; }

// ~1.3 us per poll once the back-off reaches 256 ns: 2^21 polls bound every wait to about 3 s
constexpr uint32_t SPIN_LIMIT = 1u << 21;

__device__ __forceinline__ void flag_timeout(const Comm& c, int me, uint32_t code) {
  atomicOr(reinterpret_cast<unsigned int*>(c.pads[me] + OMLDM_PAD_STATUS), code);
}

// thread t < world publishes `seq` into slot base+me of peer t's pad
__device__ __forceinline__ void signal_all(const Comm& c, int me, int base, uint32_t seq) {
  if (threadIdx.x < c.world) st_release_sys(c.pads[threadIdx.x] + base + me, seq);
}
// every thread t < world waits for slot base+t of the own pad to reach seq; block-wide result
__device__ __forceinline__ void wait_all(const Comm& c, int me, int base, uint32_t seq) {
  if (threadIdx.x < c.world) {
    const uint32_t* p = c.pads[me] + base + threadIdx.x;
    uint32_t spins = 0;
    while (static_cast<int32_t>(ld_acquire_sys(p) - seq) < 0) {
      if (++spins > 64) __nanosleep(spins > 4096 ? 256 : 32);
      if (spins > SPIN_LIMIT) { flag_timeout(c, me, 1u); break; }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ float4 f4_add(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 f4_scale(float4 a, float s) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }

// ------------------------------------------------------------------------------------------
// Phase: owner-side reduce of this rank's shard over all peers and write-back to all peers.
// n4 = number of float4 elements.  Shards are contiguous: shard r = [r*per, min(n4,(r+1)*per)).
__device__ __forceinline__ void reduce_and_broadcast_shard(const Comm& c, int me, size_t n4, float scale) {
  const size_t per = (n4 + hub_count(c) - 1) / hub_count(c);
  const size_t lo = per * me < n4 ? per * me : n4, hi = lo + per < n4 ? lo + per : n4;   // ranks >= hub_count: empty shard
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  if (c.mc_buf != nullptr && c.emulate <= 1) {
    for (size_t i = lo + blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < hi; i += stride) {
      float4 acc = mc_ld_reduce(c.mc_buf + 4 * i);  // switch-side sum over all ranks
      mc_st(c.mc_buf + 4 * i, f4_scale(acc, scale));  // switch multicast to all ranks
    }
    return;
  }
  size_t i = lo + blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  for (; i + stride < hi; i += 2 * stride) {   // two elements (2 x world peer loads) in flight per thread
    float4 v0[16], v1[16];
#pragma unroll
    for (int p = 0; p < 16; ++p)
      if (p < c.world) {
        v0[p] = ld_peer_v4(reinterpret_cast<const float4*>(c.bufs[p]) + i);
        v1[p] = ld_peer_v4(reinterpret_cast<const float4*>(c.bufs[p]) + i + stride);
      }
    float4 a0 = v0[0], a1 = v1[0];
#pragma unroll
    for (int p = 1; p < 16; ++p)
      if (p < c.world) { a0 = f4_add(a0, v0[p]); a1 = f4_add(a1, v1[p]); }
    a0 = f4_scale(a0, scale);
    a1 = f4_scale(a1, scale);
#pragma unroll
    for (int p = 0; p < 16; ++p)
      if (p < c.world) {
        st_peer_v4(reinterpret_cast<float4*>(c.bufs[p]) + i, a0);
        st_peer_v4(reinterpret_cast<float4*>(c.bufs[p]) + i + stride, a1);
      }
  }
  for (; i < hi; i += stride) {
    float4 v[16];
#pragma unroll
    for (int p = 0; p < 16; ++p)
      if (p < c.world) v[p] = ld_peer_v4(reinterpret_cast<const float4*>(c.bufs[p]) + i);
    float4 acc = v[0];
#pragma unroll
    for (int p = 1; p < 16; ++p)
      if (p < c.world) acc = f4_add(acc, v[p]);
    acc = f4_scale(acc, scale);
#pragma unroll
    for (int p = 0; p < 16; ++p)
      if (p < c.world) st_peer_v4(reinterpret_cast<float4*>(c.bufs[p]) + i, acc);
  }
}

// grid-wide "all blocks of this rank finished": returns true in exactly one (the last) block
__device__ __forceinline__ bool grid_done(uint32_t* counter) {
  __shared__ bool s_last;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t prev = atomicAdd(counter, 1u);
    s_last = (prev == gridDim.x - 1);
    if (s_last) *counter = 0;  // self-reset for the next launch
  }
  __syncthreads();
  return s_last;
}

// grid-wide sum of squares of (buf - e); every block adds its partial to *acc (float bits)
__device__ __forceinline__ void partial_sqdist(const float* buf, const float* e, size_t n, float* acc) {
  float s = 0.f;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += stride) {
    const float d = buf[i] - (e ? e[i] : 0.f);
    s = fmaf(d, d, s);
  }
  __shared__ float sh[32];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < (blockDim.x >> 5); ++i) t += sh[i];
    atomicAdd(acc, t);
  }
  __syncthreads();
}

__device__ __forceinline__ void local_tail(float* buf, float* e, size_t n, int mode) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t n4 = n / 4;
  float4* b4 = reinterpret_cast<float4*>(buf);
  float4* e4 = reinterpret_cast<float4*>(e);
  constexpr int U = 4;   // independent loads in flight per thread
  size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  for (; i + (U - 1) * stride < n4; i += U * stride) {
    float4 b[U], ev[U];
#pragma unroll
    for (int u = 0; u < U; ++u) b[u] = ld_peer_v4(b4 + i + u * stride);   // written by peers: coherent path
    if (mode != 0) {
#pragma unroll
      for (int u = 0; u < U; ++u) ev[u] = e4[i + u * stride];
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (mode == 0) {
        if (e) e4[i + u * stride] = b[u];
      } else {
        e4[i + u * stride] = f4_add(ev[u], b[u]);
        b4[i + u * stride] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  for (; i < n4; i += stride) {
    const float4 b = ld_peer_v4(b4 + i);
    if (mode == 0) {
      if (e) e4[i] = b;
    } else {
      e4[i] = f4_add(e4[i], b);
      b4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

// ------------------------------------------------------------------------------------------
// Synchronous / GM kernel.  gm_threshold < 0: plain synchronous allreduce.
// gm_threshold >= 0: each rank first votes ||buf - e||^2 > threshold; the vote travels with the
// entry barrier and the allreduce only happens if any rank voted "violated".
__global__ void __launch_bounds__(256)
sync_allreduce_kernel(Comm c, float* e_base, size_t e_stride, size_t n, float scale, int mode,
                      uint32_t epoch_arg, float gm_threshold) {
  const int me = my_rank(c);
  uint32_t* loc = my_local(c);
  const uint32_t epoch = resolve_seq(epoch_arg, loc + L_EPOCH);
  float* buf = c.bufs[me];
  float* e = e_base ? e_base + e_stride * (c.emulate > 1 ? me : 0) : nullptr;
  const uint32_t seq_entry = 2 * epoch + 1, seq_exit = 2 * epoch + 2;

  if (gm_threshold >= 0.f) {
    // delta-mode buffers are monitored by their own norm (the delta IS buf)
    partial_sqdist(buf, mode == 0 ? e : nullptr, n, reinterpret_cast<float*>(loc + L_NORM));
    if (grid_done(loc + L_DONE2)) {
      const float nrm2 = __ldcg(reinterpret_cast<float*>(loc + L_NORM));
      __syncthreads();                    // everyone has read the norm before it is reset
      if (threadIdx.x == 0) { loc[L_NORM] = 0; loc[L_EPOCH] = epoch; }
      const uint32_t vote = 2 * epoch + (nrm2 > gm_threshold ? 1u : 0u);
      if (threadIdx.x < c.world) {
        st_relaxed_sys(c.pads[threadIdx.x] + OMLDM_PAD_VOTE + me, vote);
      }
      __threadfence_system();
      __syncthreads();
      signal_all(c, me, OMLDM_PAD_ENTRY, seq_entry);
    }
  } else if (blockIdx.x == 0) {
    signal_all(c, me, OMLDM_PAD_ENTRY, seq_entry);
  }
  wait_all(c, me, OMLDM_PAD_ENTRY, seq_entry);

  if (gm_threshold >= 0.f) {
    __shared__ int s_any;
    if (threadIdx.x == 0) s_any = 0;
    __syncthreads();
    if (threadIdx.x < c.world) {
      const uint32_t v = ld_acquire_sys(c.pads[me] + OMLDM_PAD_VOTE + threadIdx.x);
      if (v == 2 * epoch + 1) atomicOr(&s_any, 1);
    }
    __syncthreads();
    if (!s_any) return;  // every rank sees the same votes, so every rank returns here together
  }

  reduce_and_broadcast_shard(c, me, n / 4, scale);
  if (grid_done(loc + L_DONE)) {
    if (threadIdx.x == 0) loc[L_EPOCH] = epoch;
    signal_all(c, me, OMLDM_PAD_EXIT, seq_exit);
  }
  wait_all(c, me, OMLDM_PAD_EXIT, seq_exit);
  local_tail(buf, e, n, mode);
}

// ------------------------------------------------------------------------------------------
// Synchronous for additive statistics with the reduce-scatter moved INTO the producer
// (ROADMAP item 0; protocol modelled and race-tested in parallel/fake_backend.py::fused_scatter_step).
// The symmetric buffer of every rank holds THREE rotating copies of the model-sized delta, bufs[p] + r*n.
//   push   (step s, rotation cur = s mod 3): the producer adds its delta, shard by shard, straight into the shard
//          OWNER's copy `cur` with system-scope reductions - no barrier before it, so it overlaps the skew
//          between ranks.  Today a stand-alone kernel over the local delta; the Gram GEMM's TMA reduce-add
//          epilogue takes this role once it addresses the owners' buffers directly.
//   gather: ONE flag barrier, then every rank adds each owner's reduced shard to its private global vector e
//          and the owner clears its shard of copy (s+2) mod 3 - the copy read in step s-1.  The barrier of step s
//          proves that every peer has finished step s-1, and nobody accumulates into that copy before the barrier
//          of step s+1, which needs this rank's arrival: no exit barrier.
__global__ void __launch_bounds__(256)
sync_scatter_push_kernel(Comm c, float* delta_base, size_t delta_stride, size_t n, uint32_t step_arg) {
  const int me = my_rank(c);
  const uint32_t* loc = my_local(c);
  const uint32_t step = resolve_seq(step_arg, loc + L_ROT);
  const size_t cur = step % 3u;
  float4* d4 = reinterpret_cast<float4*>(delta_base + delta_stride * (c.emulate > 1 ? me : 0));
  const size_t n4 = n / 4;
  const size_t per = (n4 + hub_count(c) - 1) / hub_count(c);
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
    const float4 v = d4[i];
    const int owner = static_cast<int>(i / per);
    if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f)       // e.g. the unused lower triangle of a Gram delta
      red_add_sys_v4(c.bufs[owner] + cur * n + 4 * i, v);
    d4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

__global__ void __launch_bounds__(256)
sync_scatter_gather_kernel(Comm c, float* e_base, size_t e_stride, size_t n, float scale, uint32_t epoch_arg,
                           uint32_t step_arg) {
  const int me = my_rank(c);
  uint32_t* loc = my_local(c);
  const uint32_t epoch = resolve_seq(epoch_arg, loc + L_EPOCH);
  const uint32_t step = resolve_seq(step_arg, loc + L_ROT);
  const size_t cur = step % 3u, clr = (step + 2u) % 3u;
  float4* e4 = reinterpret_cast<float4*>(e_base + e_stride * (c.emulate > 1 ? me : 0));
  const size_t n4 = n / 4;
  const size_t per = (n4 + hub_count(c) - 1) / hub_count(c);
  const size_t lo = per * me, hi = lo + per < n4 ? lo + per : n4;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const uint32_t seq_entry = 2 * epoch + 1;
  if (blockIdx.x == 0) {
    __threadfence_system();                     // this rank's pushes (previous kernel on the stream) before its flag
    signal_all(c, me, OMLDM_PAD_ENTRY, seq_entry);
  }
  wait_all(c, me, OMLDM_PAD_ENTRY, seq_entry);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
    const int owner = static_cast<int>(i / per);
    const float4 v = ld_peer_v4(reinterpret_cast<const float4*>(c.bufs[owner] + cur * n) + i);
    e4[i] = f4_add(e4[i], f4_scale(v, scale));
    if (i >= lo && i < hi) reinterpret_cast<float4*>(c.bufs[me] + clr * n)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (grid_done(loc + L_DONE)) {
    if (threadIdx.x == 0) { loc[L_EPOCH] = epoch; loc[L_ROT] = step; }
  }
}

// ------------------------------------------------------------------------------------------
// Sharded-hub Synchronous protocol for the ORR Gram statistics (additive; C1 of network/FlinkNetwork.scala:260-261 with
// the hub->worker broadcast, :282-291, turned into a pull).  GPU p is the hub shard of rows [p * rpo, (p + 1) * rpo) of
// the ld x ld Gram matrix.  hubs[p] holds k SLOT matrices (ld x ld each): slot r is written by worker r only and
// carries worker r's running sum of those rows, so a push is a plain peer store - no remote atomics, no flags, and a
// newer value simply replaces the older one.  Workers need the global statistics only when they refresh their
// weights, so nothing is broadcast per step:
//   push   : every step, each worker copies the rows its update touched (upper 32 x 32 blocks of its running sum)
//            into its slot at the owners.  The fp16 Gram kernel does this in its own epilogue (gram_f16.cu);
//            sync_gram_push_kernel is the same step for running sums maintained by another kernel (the tf32 path).
//   gather : at a weight refresh / query.  Entry rendezvous (everybody's stores of the step have landed), every owner
//            sums its k slots and stores the result into every worker's global matrix (bufs[p]), exit rendezvous.
__global__ void __launch_bounds__(256)
sync_gram_push_kernel(Comm c, const float* acc_base, size_t acc_stride, int ld, int n, int rpo) {
  const int me = my_rank(c);
  const float* acc = acc_base + acc_stride * (c.emulate > 1 ? me : 0);
  const size_t slot = static_cast<size_t>(me) * ld * ld;
  const int nb = (n + 31) / 32;                        // 32 x 32 blocks per dimension; block (bi, bj) kept if bj >= bi
  const int items = nb * nb * 32 * 8;                  // (block, row in block, float4 in row)
  for (int it = blockIdx.x * blockDim.x + threadIdx.x; it < items; it += gridDim.x * blockDim.x) {
    const int q = it & 7, r = (it >> 3) & 31, b = it >> 8;
    const int bi = b / nb, bj = b % nb;
    if (bj < bi) continue;
    const int row = bi * 32 + r, col = bj * 32 + 4 * q;
    if (row >= n || col >= ld) continue;
    const size_t off = static_cast<size_t>(row) * ld + col;
    st_peer_v4(reinterpret_cast<float4*>(c.hubs[row / rpo] + slot + off), *reinterpret_cast<const float4*>(acc + off));
  }
}

__global__ void __launch_bounds__(256)
sync_gram_gather_kernel(Comm c, int ld, int n, int rpo, uint32_t epoch_arg) {
  const int me = my_rank(c);
  uint32_t* loc = my_local(c);
  const uint32_t epoch = resolve_seq(epoch_arg, loc + L_GEPOCH);
  if (blockIdx.x == 0) {
    __threadfence_system();                            // this worker's pushes (earlier kernels on the stream) before its flag
    signal_all(c, me, OMLDM_PAD_GENTRY, 2 * epoch + 1);
  }
  wait_all(c, me, OMLDM_PAD_GENTRY, 2 * epoch + 1);
  const int lo = me * rpo, hi = min(n, lo + rpo);
  const int row4 = ld / 4;
  const size_t mat = static_cast<size_t>(ld) * ld;
  const long long total = hi > lo ? static_cast<long long>(hi - lo) * row4 : 0;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int row = lo + static_cast<int>(i / row4), q = static_cast<int>(i % row4);
    if (4 * q + 3 < (row & ~31)) continue;             // left of the diagonal block: never written
    const size_t off = static_cast<size_t>(row) * ld + 4 * q;
    float4 acc = ld_peer_v4(reinterpret_cast<const float4*>(c.hubs[me] + off));
    for (int r = 1; r < c.world; ++r) acc = f4_add(acc, ld_peer_v4(reinterpret_cast<const float4*>(c.hubs[me] + r * mat + off)));
    for (int p = 0; p < c.world; ++p) st_peer_v4(reinterpret_cast<float4*>(c.bufs[p] + off), acc);
  }
  if (grid_done(loc + L_DONE4)) {                      // includes __threadfence_system(): the rows before the flag
    if (threadIdx.x == 0) loc[L_GEPOCH] = epoch;
    signal_all(c, me, OMLDM_PAD_GEXIT, 2 * epoch + 2);
  }
  wait_all(c, me, OMLDM_PAD_GEXIT, 2 * epoch + 2);
}

// ------------------------------------------------------------------------------------------
// SingleLearner forwarding (ForwardingWorker -> CentralizedMLServer, MLNodeGenerator.scala:27,56; forced for HT and
// centralised K-means, FlinkSpoke.scala:204-209): every worker stores its mini-batch into rank 0's symmetric
// buffer, slot (seq & 1), position `me`; rank 0 trains on the k concatenated batches.  Flags: arrival of batch
// `seq` in rank 0's ENTRY words; rank 0 announces "batches <= seq - 1 are consumed" in every pad's EXIT word 0 when
// it enters step seq, and a sender waits for "seq - 2 consumed" before it overwrites that slot - senders may run two
// steps ahead, nobody else waits.
__global__ void __launch_bounds__(256)
sync_forward_kernel(Comm c, const float* z_base, size_t z_stride, size_t n, uint32_t seq) {
  const int me = my_rank(c);
  uint32_t* loc = my_local(c);
  if (me == 0 && blockIdx.x == 0 && threadIdx.x < c.world)     // everything up to seq - 1 was trained on (stream order)
    st_release_sys(c.pads[threadIdx.x] + OMLDM_PAD_EXIT, seq - 1u);
  if (threadIdx.x == 0 && seq > 2u) {
    uint32_t spins = 0;
    while (static_cast<int32_t>(ld_acquire_sys(c.pads[me] + OMLDM_PAD_EXIT) - (seq - 2u)) < 0) {
      if (++spins > 64) __nanosleep(spins > 4096 ? 256 : 32);
      if (spins > SPIN_LIMIT) { flag_timeout(c, me, 1u); break; }
    }
  }
  __syncthreads();
  const float4* src = reinterpret_cast<const float4*>(z_base + z_stride * (c.emulate > 1 ? me : 0));
  float4* dst = reinterpret_cast<float4*>(c.bufs[0] + (static_cast<size_t>(seq & 1u) * c.world + me) * n);
  const size_t n4 = n / 4;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n4; i += stride)
    st_peer_v4(dst + i, src[i]);
  if (grid_done(loc + L_DONE)) {                 // includes __threadfence_system(): the rows before the flag
    if (threadIdx.x == 0) st_release_sys(c.pads[0] + OMLDM_PAD_ENTRY + me, seq);
  }
  if (me == 0) wait_all(c, me, OMLDM_PAD_ENTRY, seq);
}

// ------------------------------------------------------------------------------------------
// Asynchronous / SSP: one worker pushes scale*(buf - e) into the owning hub shards with
// returning vector atomics and installs the fresh hub value locally.  No rendezvous.
__global__ void __launch_bounds__(256)
sync_async_kernel(Comm c, float* e_base, size_t e_stride, size_t n, float scale, int mode, int ssp,
                  uint32_t clock_arg, int staleness) {
  const int me = my_rank(c);
  uint32_t* loc = my_local(c);
  const uint32_t clock = ssp ? resolve_seq(clock_arg, loc + L_CLOCK) : 0u;
  float4* b4 = reinterpret_cast<float4*>(c.bufs[me]);
  float4* e4 = reinterpret_cast<float4*>(e_base + e_stride * (c.emulate > 1 ? me : 0));
  const size_t n4 = n / 4;
  const size_t per = (n4 + hub_count(c) - 1) / hub_count(c);
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
    const int owner = static_cast<int>(i / per);
    const float4 w = b4[i], e = e4[i];
    // mode 0: buf is the model, push scale*(buf - e).  mode 1: buf is an additive delta, push it.
    const float4 d = mode == 0
        ? make_float4(scale * (w.x - e.x), scale * (w.y - e.y), scale * (w.z - e.z), scale * (w.w - e.w))
        : f4_scale(w, scale);
    const float4 old = atom_add_sys_v4(c.hubs[owner] + 4 * i, d);
    const float4 nv = f4_add(old, d);
    b4[i] = mode == 0 ? nv : make_float4(0.f, 0.f, 0.f, 0.f);
    e4[i] = nv;
  }
  if (!ssp) return;
  if (grid_done(loc + L_DONE3)) {
    if (threadIdx.x == 0) loc[L_CLOCK] = clock;
    signal_all(c, me, OMLDM_PAD_CLOCK, clock);
    // stale-synchronous wait: nobody may be more than `staleness` clocks behind me
    if (threadIdx.x < c.world) {
      const uint32_t* p = c.pads[me] + OMLDM_PAD_CLOCK + threadIdx.x;
      uint32_t spins = 0;
      while (static_cast<int32_t>(ld_acquire_sys(p) + static_cast<uint32_t>(staleness) - clock) < 0) {
        if (++spins > 64) __nanosleep(spins > 4096 ? 256 : 32);
        if (spins > SPIN_LIMIT) { flag_timeout(c, me, 2u); break; }
      }
    }
  }
}

// EASGD: x_i -= alpha (x_i - x~);  x~ += alpha (x_i - x~), x~ sharded over the hubs.
__global__ void __launch_bounds__(256) sync_easgd_kernel(Comm c, size_t n, float alpha) {
  const int me = my_rank(c);
  float4* b4 = reinterpret_cast<float4*>(c.bufs[me]);
  const size_t n4 = n / 4;
  const size_t per = (n4 + hub_count(c) - 1) / hub_count(c);
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
    const int owner = static_cast<int>(i / per);
    float* h = c.hubs[owner] + 4 * i;
    const float4 x = b4[i];
    const float4 t = ld_peer_v4(reinterpret_cast<const float4*>(h));
    const float4 d = make_float4(alpha * (x.x - t.x), alpha * (x.y - t.y), alpha * (x.z - t.z), alpha * (x.w - t.w));
    b4[i] = make_float4(x.x - d.x, x.y - d.y, x.z - d.z, x.w - d.w);
    red_add_sys_v4(h, d);
  }
}

// ------------------------------------------------------------------------------------------
// FGM (functional geometric monitoring), one launch per mini-batch and rank.
//   safe function  phi(D) = ||D|| - sqrt(T),  D = buf - e  (convex; phi(mean D) <= mean phi(D_i))
//   round          psi0 = k*phi(0); quantum theta = -psi/(2k); local counter c_i = floor((phi_i - zeta_i)/theta)
//   sub-round end  when the hub's global counter exceeds k it sums the published phi_i into psi;
//                  psi >= eps_psi * k * phi(0) ends the round (allreduce + reset), otherwise a new
//                  sub-round starts with the smaller quantum.
// Hub = rank 0's pad.  FGM words: [0] counter(s32) [1] subround id [2] theta bits [3] round id
// (mirrored into every pad) [4] the hub's step number when it announced the round (every pad)
// [5] terminate word of the current flush generation (every pad).  Hub pad, VOTE words: per-rank "draining" marks.
// state_local (float[8], private): [0] round id [1] subround seen [2] zeta [3] counts sent [4] rounds [5] subrounds
__global__ void __launch_bounds__(256)
sync_fgm_kernel(Comm c, float* e_base, size_t e_stride, size_t n, float scale, int mode, float threshold,
                float eps_psi, uint32_t step_arg, float* state_base, uint32_t drain) {
  const int me = my_rank(c);
  const int k = c.world;
  uint32_t* loc = my_local(c);
  const uint32_t step = resolve_seq(step_arg, loc + L_STEP);
  float* buf = c.bufs[me];
  float* e = e_base + e_stride * (c.emulate > 1 ? me : 0);
  float* st = state_base + 8 * (c.emulate > 1 ? me : 0);
  uint32_t* hub = c.pads[0] + OMLDM_PAD_FGM;
  const float phi0 = -sqrtf(threshold);

  partial_sqdist(buf, mode == 0 ? e : nullptr, n, reinterpret_cast<float*>(loc + L_NORM));
  if (grid_done(loc + L_DONE2)) {
    if (threadIdx.x == 0) {
      const float nrm2 = __ldcg(reinterpret_cast<float*>(loc + L_NORM));
      loc[L_NORM] = 0;
      loc[L_STEP] = step;
      const float phi = sqrtf(nrm2) + phi0;
      uint32_t round = static_cast<uint32_t>(st[0]);
      // Round-end flag is sampled BEFORE this step's hub logic, and a round announced by the hub in ITS
      // step s (word 4) is not executed by a rank whose own step number is s: ranks that advance in
      // lock-step (one launch for all emulated ranks) then all execute it in step s+1, whether they sampled
      // the flag before or after the hub wrote it; everybody else acts at first sight.
      const uint32_t seen_round = ld_acquire_sys(c.pads[me] + OMLDM_PAD_FGM + 3);
      bool act = false;
      if (seen_round != round) act = step != ld_relaxed_sys(c.pads[me] + OMLDM_PAD_FGM + 4);
      // end of stream (flush generation `drain`): tell the hub this rank has no more data
      if (drain) st_release_sys(c.pads[0] + OMLDM_PAD_VOTE + me, drain);
      // publish phi to the hub (push-based polling)
      st_relaxed_sys(c.pads[0] + OMLDM_PAD_PHI + me, __float_as_uint(phi));
      const uint32_t sub = ld_acquire_sys(hub + 1);
      float theta = __uint_as_float(ld_relaxed_sys(hub + 2));
      if (theta <= 0.f) theta = -phi0 / 2.f;  // first sub-round of the very first round
      if (static_cast<int>(st[1]) != static_cast<int>(sub)) {  // new sub-round (or first step): re-anchor
        st[1] = static_cast<float>(sub);
        st[2] = phi;
        st[3] = 0.f;
      }
      const int ci = max(static_cast<int>(st[3]), static_cast<int>(floorf((phi - st[2]) / theta)));
      const int inc = ci - static_cast<int>(st[3]);
      if (inc > 0) {
        atom_add_relaxed_sys_s32(reinterpret_cast<int32_t*>(hub), inc);
        st[3] = static_cast<float>(ci);
      }
      if (me == 0) {
        // hub logic
        const int counter = static_cast<int>(ld_relaxed_sys(hub));
        const uint32_t hub_round = ld_relaxed_sys(hub + 3);
        bool all_done = drain != 0u && hub_round == round && seen_round == round;
        for (int i = 0; all_done && i < k; ++i) all_done = ld_acquire_sys(c.pads[0] + OMLDM_PAD_VOTE + i) == drain;
        if (all_done) {
          // everybody is draining and no round is open: terminate this flush generation (the hub is the only
          // one that announces rounds, so nothing can be announced after this)
          for (int p = 0; p < k; ++p) st_release_sys(c.pads[p] + OMLDM_PAD_FGM + 5, drain);
        } else if (counter > k && hub_round == round && seen_round == round) {
          float psi = 0.f;
          for (int i = 0; i < k; ++i) psi += __uint_as_float(ld_relaxed_sys(c.pads[0] + OMLDM_PAD_PHI + i));
          if (psi >= eps_psi * k * phi0) {
            for (int p = 0; p < k; ++p)
              st_relaxed_sys(c.pads[p] + OMLDM_PAD_FGM + 4, step);
            for (int p = 0; p < k; ++p) st_release_sys(c.pads[p] + OMLDM_PAD_FGM + 3, round + 1);  // round end
          } else {
            st_relaxed_sys(hub + 2, __float_as_uint(-psi / (2.f * k)));
            st_relaxed_sys(hub, 0u);
            st_release_sys(hub + 1, sub + 1);
            st[5] += 1.f;
          }
        }
      }
      const uint32_t decision = 2 * step + (act ? 1u : 0u);
      __threadfence();
      atomicExch(loc + L_DECISION, decision);
    }
  }
  // every block learns the decision of this step
  __shared__ uint32_t s_dec;
  if (threadIdx.x == 0) {
    uint32_t v, spins = 0;
    while (((v = atomicAdd(loc + L_DECISION, 0u)) >> 1) != step) {
      if (++spins > 64) __nanosleep(64);
      if (spins > SPIN_LIMIT) { flag_timeout(c, me, 4u); v = 2 * step; break; }
    }
    s_dec = v & 1u;
  }
  __syncthreads();
  if (!s_dec) return;

  // ---------------- round end: synchronous allreduce keyed by the round number
  const uint32_t round = static_cast<uint32_t>(st[0]);
  const uint32_t seq_entry = 2 * round + 1, seq_exit = 2 * round + 2;
  if (blockIdx.x == 0) signal_all(c, me, OMLDM_PAD_ENTRY, seq_entry);
  wait_all(c, me, OMLDM_PAD_ENTRY, seq_entry);
  reduce_and_broadcast_shard(c, me, n / 4, scale);
  if (grid_done(loc + L_DONE)) {
    if (threadIdx.x == 0) {
      if (me == 0) {  // hub resets the monitoring state for the new round
        st_relaxed_sys(hub, 0u);
        st_relaxed_sys(hub + 2, __float_as_uint(-phi0 / 2.f));
        for (int i = 0; i < k; ++i)
          st_relaxed_sys(c.pads[0] + OMLDM_PAD_PHI + i, __float_as_uint(phi0));
        const uint32_t sub = ld_relaxed_sys(hub + 1);
        st_release_sys(hub + 1, sub + 1);
      }
    }
    __syncthreads();
    __threadfence_system();
    signal_all(c, me, OMLDM_PAD_EXIT, seq_exit);
  }
  wait_all(c, me, OMLDM_PAD_EXIT, seq_exit);
  local_tail(buf, e, n, mode);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st[0] = static_cast<float>(round + 1);
    st[1] = -1.f;  // force re-anchoring at the next step
    st[3] = 0.f;
    st[4] += 1.f;
  }
}

// ------------------------------------------------------------------------------------------
// Liveness beacon: one relaxed store per peer, no waiting (see omldm_heartbeat in the header).
__global__ void heartbeat_kernel(Comm c, uint32_t beat) {
  const int me = my_rank(c);
  if (threadIdx.x < c.world) st_relaxed_sys(c.pads[threadIdx.x] + OMLDM_PAD_HEART + me, beat);
}

// ------------------------------------------------------------------------------------------
// The termination probe and the query reduce rendezvous on their OWN pad words (QENTRY / QEXIT) with their OWN
// device-managed counter (L_QEPOCH): the all-reduce epoch, the FGM round number and the query number advance
// independently, so none of them may share a flag word with another.
__global__ void device_barrier_kernel(Comm c, uint32_t epoch_arg) {
  const int me = my_rank(c);
  uint32_t* loc = my_local(c);
  const uint32_t epoch = resolve_seq(epoch_arg, loc + L_QEPOCH);
  __syncthreads();
  if (threadIdx.x == 0) loc[L_QEPOCH] = epoch;
  signal_all(c, me, OMLDM_PAD_QENTRY, 2 * epoch + 1);
  wait_all(c, me, OMLDM_PAD_QENTRY, 2 * epoch + 1);
  signal_all(c, me, OMLDM_PAD_QEXIT, 2 * epoch + 2);
  wait_all(c, me, OMLDM_PAD_QEXIT, 2 * epoch + 2);
}

// Each rank scatters its 8 query scalars into every peer's pad, barrier, then sums locally.
__global__ void query_reduce_kernel(Comm c, const float* vals_base, float* out_base, uint32_t epoch_arg) {
  const int me = my_rank(c);
  uint32_t* loc = my_local(c);
  const uint32_t epoch = resolve_seq(epoch_arg, loc + L_QEPOCH);
  __syncthreads();
  if (threadIdx.x == 0) loc[L_QEPOCH] = epoch;
  const float* vals = vals_base + 8 * (c.emulate > 1 ? me : 0);
  float* out = out_base + 8 * (c.emulate > 1 ? me : 0);
  if (threadIdx.x < c.world * 8) {
    const int p = threadIdx.x / 8, j = threadIdx.x % 8;
    st_relaxed_sys(c.pads[p] + OMLDM_PAD_QUERY + 8 * me + j, __float_as_uint(vals[j]));
  }
  __threadfence_system();
  __syncthreads();
  signal_all(c, me, OMLDM_PAD_QENTRY, 2 * epoch + 1);
  wait_all(c, me, OMLDM_PAD_QENTRY, 2 * epoch + 1);
  if (threadIdx.x < 8) {
    float s = 0.f;
    for (int p = 0; p < c.world; ++p) s += __uint_as_float(ld_acquire_sys(c.pads[me] + OMLDM_PAD_QUERY + 8 * p + threadIdx.x));
    out[threadIdx.x] = s;
  }
  __syncthreads();
  signal_all(c, me, OMLDM_PAD_QEXIT, 2 * epoch + 2);
  wait_all(c, me, OMLDM_PAD_QEXIT, 2 * epoch + 2);
}

static Comm to_comm(const omldm_comm* h) {
  Comm c;
  c.bufs = h->bufs; c.hubs = h->hubs; c.pads = h->pads; c.local = h->local; c.mc_buf = h->mc_buf;
  c.rank = h->rank; c.world = h->world; c.emulate = h->emulate; c.nhub = h->nhub;
  return c;
}
static int clamp_blocks(int blocks, size_t n, int emulate) {
  int maxb = emulate > 1 ? (148 / emulate) : 148;
  if (maxb < 1) maxb = 1;
  int need = static_cast<int>((n / 4 + 1023) / 1024);   // ~4 float4 per thread
  if (need < 1) need = 1;
  if (blocks <= 0) blocks = need < 128 ? need : 128;
  if (blocks > maxb) blocks = maxb;
  return blocks;
}
// kernels whose blocks wait on one another must be co-resident: cooperative launch
template <typename... Args>
static int launch_coop(void (*kern)(Args...), dim3 grid, dim3 block, cudaStream_t st, const char* name, Args... args) {
#ifdef OMLDM_EMU   // tools/cuda_emu: every block of the launch runs concurrently (blocks and ranks wait on one another)
  emu::launcher_concurrent(kern, grid, block, 0, st)(args...);
  return check_launch(name);
#endif
  if (grid.y == 1) {
    // one rank per launch: the grid is far below one wave, every block is resident, and a plain
    // launch can be captured into a CUDA graph
    kern<<<grid, block, 0, st>>>(args...);
    return check_launch(name);
  }
  void* params[] = {(void*)&args...};
  cudaError_t e = cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(kern), grid, block, params, 0, st);
  if (e != cudaSuccess) return set_error(name, e);
  count_launch(1);
  return 0;
}

}  // namespace omldm

using namespace omldm;

static int check_world(const omldm_comm* h, size_t n) {
  if (h->world < 1 || h->world > 16) return set_error_msg("comm: world must be in [1,16]");
  if (n % 4) return set_error_msg("comm: element count must be a multiple of 4 (pad the model vector)");
  return 0;
}

extern "C" int omldm_sync_allreduce(const omldm_comm* h, float* e_local, size_t n, float scale, int mode,
                                    uint32_t epoch, int blocks, void* stream) {
  if (check_world(h, n)) return -1;
  Comm c = to_comm(h);
  blocks = clamp_blocks(blocks, n, h->emulate);
  dim3 grid(blocks, h->emulate > 1 ? h->emulate : 1);
  return launch_coop(sync_allreduce_kernel, grid, dim3(256), static_cast<cudaStream_t>(stream), "sync_allreduce_kernel",
                     c, e_local, n, n, scale, mode, epoch, -1.0f);
}

extern "C" int omldm_sync_gm(const omldm_comm* h, float* e_local, size_t n, float scale, int mode, float threshold,
                             uint32_t epoch, int blocks, void* stream) {
  if (check_world(h, n)) return -1;
  if (!e_local) return set_error_msg("sync_gm: the estimate vector e is required");
  Comm c = to_comm(h);
  blocks = clamp_blocks(blocks, n, h->emulate);
  dim3 grid(blocks, h->emulate > 1 ? h->emulate : 1);
  return launch_coop(sync_allreduce_kernel, grid, dim3(256), static_cast<cudaStream_t>(stream), "sync_gm_kernel",
                     c, e_local, n, n, scale, mode, epoch, threshold);
}

// `n` = model size; the symmetric buffers hold 3 * n floats.  step / epoch 0 = device managed.
extern "C" int omldm_sync_scatter_push(const omldm_comm* h, float* delta_local, size_t n, uint32_t step, int blocks,
                                       void* stream) {
  if (check_world(h, n)) return -1;
  if (!delta_local) return set_error_msg("sync_scatter_push: the local delta is required");
  Comm c = to_comm(h);
  blocks = clamp_blocks(blocks, n, h->emulate);
  dim3 grid(blocks, h->emulate > 1 ? h->emulate : 1);
  return launch_coop(sync_scatter_push_kernel, grid, dim3(256), static_cast<cudaStream_t>(stream),
                     "sync_scatter_push_kernel", c, delta_local, n, n, step);
}
extern "C" int omldm_sync_scatter_gather(const omldm_comm* h, float* e_local, size_t n, float scale, uint32_t epoch,
                                         uint32_t step, int blocks, void* stream) {
  if (check_world(h, n)) return -1;
  if (!e_local) return set_error_msg("sync_scatter_gather: the global vector e is required");
  Comm c = to_comm(h);
  blocks = clamp_blocks(blocks, n, h->emulate);
  dim3 grid(blocks, h->emulate > 1 ? h->emulate : 1);
  return launch_coop(sync_scatter_gather_kernel, grid, dim3(256), static_cast<cudaStream_t>(stream),
                     "sync_scatter_gather_kernel", c, e_local, n, n, scale, epoch, step);
}

// Sharded-hub Gram protocol.  hubs[p] = k slot matrices (ld x ld each) at owner p; rpo = rows per owner (multiple of 32).
extern "C" int omldm_sync_gram_push(const omldm_comm* h, const float* acc_local, int ld, int n, int rows_per_owner,
                                    int blocks, void* stream) {
  if (h->world < 1 || h->world > 16) return set_error_msg("comm: world must be in [1,16]");
  if (!acc_local) return set_error_msg("sync_gram_push: the running sum is required");
  if ((ld & 3) || n > ld || rows_per_owner < 32 || (rows_per_owner & 31) ||
      static_cast<long long>(rows_per_owner) * h->world < n)
    return set_error_msg("sync_gram_push: ld % 4 == 0, n <= ld, rows_per_owner a multiple of 32 that covers n");
  Comm c = to_comm(h);
  blocks = clamp_blocks(blocks, static_cast<size_t>(n) * ld / 2, h->emulate);
  dim3 grid(blocks, h->emulate > 1 ? h->emulate : 1);
  return launch_coop(sync_gram_push_kernel, grid, dim3(256), static_cast<cudaStream_t>(stream), "sync_gram_push_kernel",
                     c, acc_local, static_cast<size_t>(ld) * ld, ld, n, rows_per_owner);
}
extern "C" int omldm_sync_gram_gather(const omldm_comm* h, int ld, int n, int rows_per_owner, uint32_t epoch, int blocks,
                                      void* stream) {
  if (h->world < 1 || h->world > 16) return set_error_msg("comm: world must be in [1,16]");
  if ((ld & 3) || n > ld || rows_per_owner < 32 || (rows_per_owner & 31))
    return set_error_msg("sync_gram_gather: ld % 4 == 0, n <= ld, rows_per_owner a multiple of 32");
  Comm c = to_comm(h);
  blocks = clamp_blocks(blocks, static_cast<size_t>(rows_per_owner) * ld, h->emulate);
  dim3 grid(blocks, h->emulate > 1 ? h->emulate : 1);
  return launch_coop(sync_gram_gather_kernel, grid, dim3(256), static_cast<cudaStream_t>(stream),
                     "sync_gram_gather_kernel", c, ld, n, rows_per_owner, epoch);
}

// n floats per rank and batch (multiple of 4); rank 0's symmetric buffer holds 2 * world * n floats; seq = 1, 2, ...
extern "C" int omldm_sync_forward(const omldm_comm* h, const float* z_local, size_t n, uint32_t seq, int blocks,
                                  void* stream) {
  if (check_world(h, n)) return -1;
  if (!z_local || seq == 0u) return set_error_msg("sync_forward: a batch and a sequence number >= 1 are required");
  Comm c = to_comm(h);
  blocks = clamp_blocks(blocks, n, h->emulate);
  dim3 grid(blocks, h->emulate > 1 ? h->emulate : 1);
  return launch_coop(sync_forward_kernel, grid, dim3(256), static_cast<cudaStream_t>(stream), "sync_forward_kernel",
                     c, z_local, n, n, seq);
}

extern "C" int omldm_sync_async(const omldm_comm* h, float* e_local, size_t n, float scale, int mode, int blocks,
                                void* stream) {
  if (check_world(h, n)) return -1;
  Comm c = to_comm(h);
  blocks = clamp_blocks(blocks, n, h->emulate);
  dim3 grid(blocks, h->emulate > 1 ? h->emulate : 1);
  sync_async_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(c, e_local, n, n, scale, mode, 0, 0u, 0);
  return check_launch("sync_async_kernel");
}

extern "C" int omldm_sync_ssp(const omldm_comm* h, float* e_local, size_t n, float scale, int mode, uint32_t clock,
                              int staleness, int blocks, void* stream) {
  if (check_world(h, n)) return -1;
  Comm c = to_comm(h);
  blocks = clamp_blocks(blocks, n, h->emulate);
  dim3 grid(blocks, h->emulate > 1 ? h->emulate : 1);
  return launch_coop(sync_async_kernel, grid, dim3(256), static_cast<cudaStream_t>(stream), "sync_ssp_kernel",
                     c, e_local, n, n, scale, mode, 1, clock, staleness);
}

extern "C" int omldm_sync_easgd(const omldm_comm* h, size_t n, float alpha, int blocks, void* stream) {
  if (check_world(h, n)) return -1;
  Comm c = to_comm(h);
  blocks = clamp_blocks(blocks, n, h->emulate);
  dim3 grid(blocks, h->emulate > 1 ? h->emulate : 1);
  sync_easgd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(c, n, alpha);
  return check_launch("sync_easgd_kernel");
}

extern "C" int omldm_sync_fgm(const omldm_comm* h, float* e_local, size_t n, float scale, int mode, float threshold,
                              float eps_psi, uint32_t step, float* state_local, int blocks, void* stream) {
  if (check_world(h, n)) return -1;
  if (!e_local || !state_local) return set_error_msg("sync_fgm: e and the FGM state are required");
  Comm c = to_comm(h);
  blocks = clamp_blocks(blocks, n, h->emulate);
  dim3 grid(blocks, h->emulate > 1 ? h->emulate : 1);
  return launch_coop(sync_fgm_kernel, grid, dim3(256), static_cast<cudaStream_t>(stream), "sync_fgm_kernel",
                     c, e_local, n, n, scale, mode, threshold, eps_psi, step, state_local, 0u);
}

// End-of-stream step of flush generation `drain` (> 0): like omldm_sync_fgm, plus the drain handshake.
extern "C" int omldm_sync_fgm_drain(const omldm_comm* h, float* e_local, size_t n, float scale, int mode,
                                    float threshold, float eps_psi, uint32_t step, float* state_local,
                                    uint32_t drain, int blocks, void* stream) {
  if (check_world(h, n)) return -1;
  if (!e_local || !state_local) return set_error_msg("sync_fgm_drain: e and the FGM state are required");
  if (drain == 0u) return set_error_msg("sync_fgm_drain: the flush generation must be > 0");
  Comm c = to_comm(h);
  blocks = clamp_blocks(blocks, n, h->emulate);
  dim3 grid(blocks, h->emulate > 1 ? h->emulate : 1);
  return launch_coop(sync_fgm_kernel, grid, dim3(256), static_cast<cudaStream_t>(stream), "sync_fgm_kernel",
                     c, e_local, n, n, scale, mode, threshold, eps_psi, step, state_local, drain);
}

extern "C" int omldm_heartbeat(const omldm_comm* h, uint32_t beat, void* stream) {
  if (h->world < 1 || h->world > 16) return set_error_msg("comm: world must be in [1,16]");
  Comm c = to_comm(h);
  dim3 grid(1, h->emulate > 1 ? h->emulate : 1);
  return launch_coop(heartbeat_kernel, grid, dim3(32), static_cast<cudaStream_t>(stream), "heartbeat_kernel", c, beat);
}

extern "C" int omldm_device_barrier(const omldm_comm* h, uint32_t epoch, void* stream) {
  Comm c = to_comm(h);
  dim3 grid(1, h->emulate > 1 ? h->emulate : 1);
  return launch_coop(device_barrier_kernel, grid, dim3(32), static_cast<cudaStream_t>(stream), "device_barrier_kernel", c, epoch);
}

extern "C" int omldm_query_reduce(const omldm_comm* h, const float* vals8, float* out8, uint32_t epoch, void* stream) {
  Comm c = to_comm(h);
  dim3 grid(1, h->emulate > 1 ? h->emulate : 1);
  return launch_coop(query_reduce_kernel, grid, dim3(128), static_cast<cudaStream_t>(stream), "query_reduce_kernel",
                     c, vals8, out8, epoch);
}

#ifndef OMLDM_EMU   // CUDA IPC has no meaning in the CPU emulation
extern "C" int omldm_ipc_export(void* dev_ptr, void* handle64_host) {
  cudaIpcMemHandle_t hnd;
  cudaError_t e = cudaIpcGetMemHandle(&hnd, dev_ptr);
  if (e != cudaSuccess) return set_error("cudaIpcGetMemHandle", e);
  static_assert(sizeof(hnd) == 64, "ipc handle size");
  memcpy(handle64_host, &hnd, 64);
  return 0;
}
extern "C" int omldm_ipc_import(const void* handle64_host, void** dev_ptr_out) {
  cudaIpcMemHandle_t hnd;
  memcpy(&hnd, handle64_host, 64);
  cudaError_t e = cudaIpcOpenMemHandle(dev_ptr_out, hnd, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return set_error("cudaIpcOpenMemHandle", e);
  return 0;
}
extern "C" int omldm_ipc_close(void* dev_ptr) {
  cudaError_t e = cudaIpcCloseMemHandle(dev_ptr);
  if (e != cudaSuccess) return set_error("cudaIpcCloseMemHandle", e);
  return 0;
}
extern "C" int omldm_enable_peer_access(int peer_device) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return 0; }
  if (e != cudaSuccess) return set_error("cudaDeviceEnablePeerAccess", e);
  return 0;
}

#endif  // !OMLDM_EMU

// Arena memory for the symmetric buffers: a dedicated cudaMalloc so that the IPC handle maps the
// whole region at offset 0 in every peer.
extern "C" int omldm_dev_malloc(size_t bytes, void** out) {
  cudaError_t e = cudaMalloc(out, bytes);
  if (e != cudaSuccess) return set_error("cudaMalloc(arena)", e);
  e = cudaMemset(*out, 0, bytes);
  if (e != cudaSuccess) return set_error("cudaMemset(arena)", e);
  return 0;
}
extern "C" int omldm_dev_free(void* p) {
  cudaError_t e = cudaFree(p);
  if (e != cudaSuccess) return set_error("cudaFree(arena)", e);
  return 0;
}
