// CUDA-on-CPU emulation shim: lets the REAL kernel sources of csrc/*.cu (the ones without inline PTX) be compiled
// by g++ and run on host memory, one std::thread per CUDA thread, so that
//   * kernels can be compared against their PyTorch references on a box without a GPU, and
//   * AddressSanitizer / UBSan see every global and shared memory access of the kernel code
//     (compute-sanitizer is not available on the GPU pool this project is developed on).
// Semantics: blocks of a launch run one after the other; the threads of a block run concurrently and are
// NOT warp-synchronous (like independent thread scheduling); __syncthreads / __syncwarp / shuffles are real
// barriers that tolerate threads which already returned; every `__shared__` declaration gets per-block storage
// (filled with 0xCD, like uninitialised memory), dynamic shared memory is an exactly-sized heap block; atomics are
// host atomics.  `launcher_concurrent` runs all blocks of a grid at once, for kernels whose blocks (or emulated
// ranks) wait on one another; there a cooperative-groups grid barrier spans the whole grid.  Not emulated: tensor cores, TMA, NVLS multimem, streams (everything is synchronous).
#pragma once
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <type_traits>
#include <vector>

#define OMLDM_EMU 1

// ---------------------------------------------------------------- vector types
struct uint3 { unsigned x, y, z; };
struct dim3 {
  unsigned x = 1, y = 1, z = 1;
  dim3() = default;
  dim3(unsigned a, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {}
};
struct alignas(8) float2 { float x, y; };
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(16) double2 { double x, y; };
struct alignas(16) int4 { int x, y, z, w; };
struct alignas(8) int2 { int x, y; };
inline float2 make_float2(float x, float y) { return {x, y}; }
inline float4 make_float4(float x, float y, float z, float w) { return {x, y, z, w}; }
inline double2 make_double2(double x, double y) { return {x, y}; }
inline int4 make_int4(int x, int y, int z, int w) { return {x, y, z, w}; }
inline int2 make_int2(int x, int y) { return {x, y}; }

// ---------------------------------------------------------------- qualifiers
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __align__(n) __attribute__((aligned(n)))
#define __grid_constant__
#ifndef __restrict__
#define __restrict__ __restrict
#endif

namespace emu {

// barrier that tolerates participants leaving for good (threads that returned from the kernel)
struct Barrier {
  std::mutex m;
  std::condition_variable cv;
  int active = 0, waiting = 0;
  unsigned gen = 0;
  void reset(int n) { active = n; waiting = 0; }
  void sync() {
    std::unique_lock<std::mutex> l(m);
    const unsigned g = gen;
    if (++waiting >= active) { waiting = 0; ++gen; cv.notify_all(); }
    else cv.wait(l, [&] { return gen != g; });
  }
  void drop() {
    std::unique_lock<std::mutex> l(m);
    --active;
    if (active > 0 && waiting >= active) { waiting = 0; ++gen; cv.notify_all(); }
  }
};
struct Warp {
  Barrier bar;
  uint64_t slot[32];
};
struct BlockCtx {
  Barrier bar;
  std::vector<Warp> warps;
  char* dyn = nullptr;
  Barrier* grid_bar = nullptr;                      // set by run_grid_concurrent: barrier over every thread of the grid
  std::mutex shared_m;
  std::vector<std::pair<int, void*>> shared;        // (declaration id, storage) of the `__shared__` variables
  ~BlockCtx() { for (auto& kv : shared) std::free(kv.second); }
};

inline thread_local uint3 t_threadIdx, t_blockIdx;
inline thread_local dim3 t_blockDim, t_gridDim;
inline thread_local BlockCtx* t_ctx = nullptr;
inline thread_local int t_lane = 0, t_warp = 0;

inline void* dyn_smem() { return t_ctx->dyn; }

// storage of one `__shared__` declaration (build_emu.py rewrites `__shared__ T x[N];` into a reference bound to this):
// one instance per block, created by the first thread that reaches the declaration, filled with 0xCD like
// uninitialised memory so that a read-before-write shows up as garbage rather than as a convenient zero
template <class T>
inline T* block_shared(int id) {
  BlockCtx* c = t_ctx;
  std::lock_guard<std::mutex> l(c->shared_m);
  for (auto& kv : c->shared)
    if (kv.first == id) return static_cast<T*>(kv.second);
  void* p = std::malloc(sizeof(T));
  std::memset(p, 0xCD, sizeof(T));
  c->shared.emplace_back(id, p);
  return static_cast<T*>(p);
}

// OMLDM_EMU_WAVE=w (> 1): plain launches run their blocks in waves of w concurrent blocks instead of one after the
// other, which exposes the inter-block protocols ("the last block to finish applies the update") to real interleavings
// and to ThreadSanitizer.
inline int wave_size() {
  static const int w = [] { const char* e = std::getenv("OMLDM_EMU_WAVE"); const int v = e ? std::atoi(e) : 1; return v < 1 ? 1 : v; }();
  return w;
}

// OMLDM_EMU_JITTER=p (per mille, 0 = off): every atomic, every sys-scope access of ptx.cuh and every block-barrier exit
// pauses the calling thread with probability p / 1000 for 1..100 us.  Host threads that run freely interleave far less
// adversarially than warps on 148 SMs do; the pauses widen the windows in which an unordered pair of accesses can be
// observed in the wrong order (same idea as the `jitter` of parallel/fake_backend.py, applied to the real kernels).
inline int jitter_permille() {
  static const int p = [] { const char* e = std::getenv("OMLDM_EMU_JITTER"); const int v = e ? std::atoi(e) : 0; return v < 0 ? 0 : v; }();
  return p;
}
inline void jitter(int scale = 1) {                                // scale: 16 for per-element data accesses
  const int p = jitter_permille() / scale;
  if (p == 0) return;
  static std::atomic<uint64_t> streams{0x9E3779B97F4A7C15ull};
  thread_local uint64_t s = streams.fetch_add(0xBF58476D1CE4E5B9ull, std::memory_order_relaxed) | 1u;
  s ^= s << 13; s ^= s >> 7; s ^= s << 17;                       // xorshift64
  if (static_cast<int>(s % 1000u) < p)
    std::this_thread::sleep_for(std::chrono::microseconds(1 + static_cast<int>((s >> 32) % 100u)));
}

template <class Body>
void run_grid_waves(dim3 g, dim3 b, size_t smem, int wave, Body&& body) {
  const int T = static_cast<int>(b.x * b.y * b.z);
  const int nb = static_cast<int>(g.x * g.y * g.z);
  for (int b0 = 0; b0 < nb; b0 += wave) {
    const int w = std::min(wave, nb - b0);
    std::vector<BlockCtx> ctx(static_cast<size_t>(w));
    for (auto& c : ctx) {
      c.warps = std::vector<Warp>(static_cast<size_t>((T + 31) / 32));
      c.dyn = smem ? static_cast<char*>(std::malloc(smem)) : nullptr;
      c.bar.reset(T);
      for (size_t i = 0; i < c.warps.size(); ++i) c.warps[i].bar.reset(std::min(32, T - static_cast<int>(i) * 32));
    }
    // start gate: thread creation takes longer than a small kernel runs.  The waiters sleep (a thousand spinning
    // threads starve the creator, badly so under ThreadSanitizer).
    std::mutex gate_m;
    std::condition_variable gate_cv;
    bool go = false;
    auto worker = [&](int k, int t) {
      {
        std::unique_lock<std::mutex> l(gate_m);
        gate_cv.wait(l, [&] { return go; });
      }
      const int blk = b0 + k;
      t_blockDim = b;
      t_gridDim = g;
      t_ctx = &ctx[static_cast<size_t>(k)];
      t_threadIdx = {static_cast<unsigned>(t) % b.x, (static_cast<unsigned>(t) / b.x) % b.y, static_cast<unsigned>(t) / (b.x * b.y)};
      t_blockIdx = {static_cast<unsigned>(blk) % g.x, (static_cast<unsigned>(blk) / g.x) % g.y, static_cast<unsigned>(blk) / (g.x * g.y)};
      t_lane = t & 31;
      t_warp = t >> 5;
      body();
      t_ctx->warps[static_cast<size_t>(t_warp)].bar.drop();
      t_ctx->bar.drop();
    };
    std::vector<std::thread> th;
    th.reserve(static_cast<size_t>(w) * T);
    for (int k = 0; k < w; ++k)
      for (int t = 0; t < T; ++t) th.emplace_back(worker, k, t);
    {
      std::lock_guard<std::mutex> l(gate_m);
      go = true;
    }
    gate_cv.notify_all();
    for (auto& x : th) x.join();
    for (auto& c : ctx) std::free(c.dyn);
  }
}

template <class Body>
void run_grid(dim3 g, dim3 b, size_t smem, Body&& body) {
  const int T = static_cast<int>(b.x * b.y * b.z);
  if (T <= 0 || g.x * g.y * g.z == 0) return;
  if (wave_size() > 1 && g.x * g.y * g.z > 1) return run_grid_waves(g, b, smem, wave_size(), body);
  BlockCtx ctx;
  ctx.warps = std::vector<Warp>(static_cast<size_t>((T + 31) / 32));
  ctx.dyn = smem ? static_cast<char*>(std::malloc(smem)) : nullptr;   // exact size: overruns are visible to ASan
  Barrier fence;                                                      // never dropped: block boundaries
  fence.reset(T);
  auto arm = [&] {
    ctx.bar.reset(T);
    for (size_t w = 0; w < ctx.warps.size(); ++w)
      ctx.warps[w].bar.reset(std::min(32, T - static_cast<int>(w) * 32));
  };
  arm();
  auto worker = [&](int t) {
    t_blockDim = b;
    t_gridDim = g;
    t_ctx = &ctx;
    t_threadIdx = {static_cast<unsigned>(t) % b.x, (static_cast<unsigned>(t) / b.x) % b.y, static_cast<unsigned>(t) / (b.x * b.y)};
    t_lane = t & 31;
    t_warp = t >> 5;
    for (unsigned z = 0; z < g.z; ++z)
      for (unsigned y = 0; y < g.y; ++y)
        for (unsigned x = 0; x < g.x; ++x) {
          t_blockIdx = {x, y, z};
          body();
          ctx.warps[static_cast<size_t>(t_warp)].bar.drop();
          ctx.bar.drop();
          fence.sync();            // every thread has left block (x, y, z)
          if (t == 0) arm();
          fence.sync();
        }
  };
  std::vector<std::thread> th;
  th.reserve(static_cast<size_t>(T) - 1);
  for (int t = 1; t < T; ++t) th.emplace_back(worker, t);
  worker(0);
  for (auto& x : th) x.join();
  std::free(ctx.dyn);
}

template <class... KArgs>
struct Launcher {
  void (*k)(KArgs...);
  dim3 g, b;
  size_t smem;
  template <class... A>
  void operator()(A&&... a) const {
    auto kk = k;
    run_grid(g, b, smem, [&] { kk(static_cast<KArgs>(a)...); });
  }
};
template <class... KArgs>
Launcher<KArgs...> launcher(void (*k)(KArgs...), dim3 g, dim3 b, size_t smem = 0, void* = nullptr) {
  return Launcher<KArgs...>{k, g, b, smem};
}

// Every block of the grid at once (kernels whose blocks - or emulated ranks - wait on one another).
template <class Body>
void run_grid_concurrent(dim3 g, dim3 b, size_t smem, Body&& body) {
  const int T = static_cast<int>(b.x * b.y * b.z);
  const int nb = static_cast<int>(g.x * g.y * g.z);
  if (T <= 0 || nb <= 0) return;
  std::vector<BlockCtx> ctx(static_cast<size_t>(nb));
  Barrier grid_bar;
  grid_bar.reset(nb * T);
  for (auto& c : ctx) {
    c.grid_bar = &grid_bar;
    c.warps = std::vector<Warp>(static_cast<size_t>((T + 31) / 32));
    c.dyn = smem ? static_cast<char*>(std::malloc(smem)) : nullptr;
    c.bar.reset(T);
    for (size_t w = 0; w < c.warps.size(); ++w) c.warps[w].bar.reset(std::min(32, T - static_cast<int>(w) * 32));
  }
  auto worker = [&](int blk, int t) {
    t_blockDim = b;
    t_gridDim = g;
    t_ctx = &ctx[static_cast<size_t>(blk)];
    t_threadIdx = {static_cast<unsigned>(t) % b.x, (static_cast<unsigned>(t) / b.x) % b.y, static_cast<unsigned>(t) / (b.x * b.y)};
    t_blockIdx = {static_cast<unsigned>(blk) % g.x, (static_cast<unsigned>(blk) / g.x) % g.y, static_cast<unsigned>(blk) / (g.x * g.y)};
    t_lane = t & 31;
    t_warp = t >> 5;
    body();
    t_ctx->warps[static_cast<size_t>(t_warp)].bar.drop();
    t_ctx->bar.drop();
    grid_bar.drop();
  };
  std::vector<std::thread> th;
  th.reserve(static_cast<size_t>(nb) * T);
  for (int blk = 0; blk < nb; ++blk)
    for (int t = 0; t < T; ++t) th.emplace_back(worker, blk, t);
  for (auto& x : th) x.join();
  for (auto& c : ctx) std::free(c.dyn);
}
template <class... KArgs>
struct ConcurrentLauncher {
  void (*k)(KArgs...);
  dim3 g, b;
  size_t smem;
  template <class... A>
  void operator()(A&&... a) const {
    auto kk = k;
    run_grid_concurrent(g, b, smem, [&] { kk(static_cast<KArgs>(a)...); });
  }
};
template <class... KArgs>
ConcurrentLauncher<KArgs...> launcher_concurrent(void (*k)(KArgs...), dim3 g, dim3 b, size_t smem = 0, void* = nullptr) {
  return ConcurrentLauncher<KArgs...>{k, g, b, smem};
}

template <class T>
inline uint64_t to_bits(T v) { uint64_t u = 0; static_assert(sizeof(T) <= 8, "shuffle payload"); std::memcpy(&u, &v, sizeof(T)); return u; }
template <class T>
inline T from_bits(uint64_t u) { T v; std::memcpy(&v, &u, sizeof(T)); return v; }

template <class T>
inline T shfl(T v, int src_lane) {
  Warp& w = t_ctx->warps[static_cast<size_t>(t_warp)];
  w.slot[t_lane] = to_bits(v);
  w.bar.sync();
  const T r = from_bits<T>(w.slot[src_lane & 31]);
  w.bar.sync();
  return r;
}

}  // namespace emu

#define threadIdx (emu::t_threadIdx)
#define blockIdx (emu::t_blockIdx)
#define blockDim (emu::t_blockDim)
#define gridDim (emu::t_gridDim)
#define warpSize 32

// ---------------------------------------------------------------- synchronisation / warp primitives
inline void __syncthreads() { emu::t_ctx->bar.sync(); emu::jitter(); }
inline void __syncwarp(unsigned = 0xffffffffu) { emu::t_ctx->warps[static_cast<size_t>(emu::t_warp)].bar.sync(); }
// a real sleep: spin loops are bounded by iteration counts calibrated for GPU time, and thousands of host threads
// share a few cores - a waiting thread must get out of the way of the one it waits for
inline void __nanosleep(unsigned) { std::this_thread::sleep_for(std::chrono::microseconds(20)); }
inline void __threadfence() { std::atomic_thread_fence(std::memory_order_seq_cst); }
inline void __threadfence_block() { std::atomic_thread_fence(std::memory_order_seq_cst); }
inline void __threadfence_system() { std::atomic_thread_fence(std::memory_order_seq_cst); }
template <class T> inline T __shfl_xor_sync(unsigned, T v, int mask) { return emu::shfl(v, emu::t_lane ^ mask); }
template <class T> inline T __shfl_sync(unsigned, T v, int lane) { return emu::shfl(v, lane); }
template <class T> inline T __shfl_down_sync(unsigned, T v, int d) { return emu::shfl(v, emu::t_lane + d < 32 ? emu::t_lane + d : emu::t_lane); }

// ---------------------------------------------------------------- atomics (host atomics on plain memory)
template <class T>
inline T emu_atomic_add(T* p, T v) {
  emu::jitter();
  if constexpr (std::is_integral<T>::value) {
    return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST);
  } else {
    using U = typename std::conditional<sizeof(T) == 4, uint32_t, uint64_t>::type;
    U* up = reinterpret_cast<U*>(p);
    // floating-point accumulators: RELAXED like CUDA's atomicAdd, so that ThreadSanitizer does not take a chain of
    // additions to one address for synchronisation (it would hide a reader that is not ordered after every adder);
    // the integer atomics above stay seq_cst - they are the counters of the fence + atomic idioms
    U old = __atomic_load_n(up, __ATOMIC_RELAXED);
    for (;;) {
      T cur;
      std::memcpy(&cur, &old, sizeof(T));
      const T nv = cur + v;
      U nb;
      std::memcpy(&nb, &nv, sizeof(T));
      if (__atomic_compare_exchange_n(up, &old, nb, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) return cur;
    }
  }
}
inline float atomicAdd(float* p, float v) { return emu_atomic_add(p, v); }
inline double atomicAdd(double* p, double v) { return emu_atomic_add(p, v); }
inline int atomicAdd(int* p, int v) { return emu_atomic_add(p, v); }
inline unsigned atomicAdd(unsigned* p, unsigned v) { return emu_atomic_add(p, v); }
inline unsigned long long atomicAdd(unsigned long long* p, unsigned long long v) { return emu_atomic_add(p, v); }
template <class T>
inline T atomicCAS(T* p, T cmp, T val) {
  emu::jitter();
  __atomic_compare_exchange_n(p, &cmp, val, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
  return cmp;
}
template <class T> inline T atomicExch(T* p, T v) { emu::jitter(); return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
template <class T> inline T atomicOr(T* p, T v) { emu::jitter(); return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }
template <class T> inline T atomicMax(T* p, T v) {
  T old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
  while (old < v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
  return old;
}
template <class T> inline T atomicMin(T* p, T v) {
  T old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
  while (old > v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
  return old;
}

// ---------------------------------------------------------------- loads, bit casts, math
// ld.global.cg is a single-copy-atomic access at L2: kernels use it for words that other threads update with atomics
template <class T> inline T __ldcg(const T* p) {
  if constexpr (std::is_arithmetic<T>::value && (sizeof(T) == 4 || sizeof(T) == 8)) {
    using U = typename std::conditional<sizeof(T) == 4, uint32_t, uint64_t>::type;
    const U u = __atomic_load_n(reinterpret_cast<const U*>(p), __ATOMIC_RELAXED);
    T v;
    std::memcpy(&v, &u, sizeof(T));
    return v;
  } else {
    return *p;
  }
}
template <class T> inline T __ldg(const T* p) { return *p; }
template <class T> inline T __ldcs(const T* p) { return *p; }
template <class T> inline void __stcg(T* p, T v) { *p = v; }
inline int __float_as_int(float f) { int i; std::memcpy(&i, &f, 4); return i; }
inline float __int_as_float(int i) { float f; std::memcpy(&f, &i, 4); return f; }
inline unsigned __float_as_uint(float f) { unsigned i; std::memcpy(&i, &f, 4); return i; }
inline float __uint_as_float(unsigned i) { float f; std::memcpy(&f, &i, 4); return f; }
inline long long __double_as_longlong(double d) { long long i; std::memcpy(&i, &d, 8); return i; }
inline double __longlong_as_double(long long i) { double d; std::memcpy(&d, &i, 8); return d; }
template <class A, class B> inline typename std::common_type<A, B>::type min(A a, B b) { return b < a ? b : a; }
template <class A, class B> inline typename std::common_type<A, B>::type max(A a, B b) { return a < b ? b : a; }
inline float rsqrtf(float x) { return 1.0f / std::sqrt(x); }
inline double rsqrt(double x) { return 1.0 / std::sqrt(x); }
inline float __fdividef(float a, float b) { return a / b; }
inline float __expf(float x) { return std::exp(x); }
inline float __frcp_rn(float x) { return 1.0f / x; }
inline int __popc(unsigned x) { return __builtin_popcount(x); }
inline int __clz(int x) { return x ? __builtin_clz(static_cast<unsigned>(x)) : 32; }
#define CUDART_INF_F (__builtin_inff())
#define CUDART_INF (__builtin_inf())
using std::isfinite;
using std::isnan;

// ---------------------------------------------------------------- fp16 (storage type + conversions only)
struct __half { uint16_t x; };
struct alignas(4) __half2 { __half x, y; };
inline float __half2float(__half h) {
  const uint32_t s = (h.x & 0x8000u) << 16, e = (h.x >> 10) & 31u, m = h.x & 1023u;
  uint32_t u;
  if (e == 0) {
    if (m == 0) u = s;
    else { int k = 0; uint32_t mm = m; while (!(mm & 1024u)) { mm <<= 1; ++k; } u = s | ((113u - k) << 23) | ((mm & 1023u) << 13); }
  } else if (e == 31) u = s | 0x7F800000u | (m << 13);
  else u = s | ((e + 112u) << 23) | (m << 13);
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}
inline float2 __half22float2(__half2 h) { return {__half2float(h.x), __half2float(h.y)}; }

// ---------------------------------------------------------------- the slice of the runtime API the host wrappers use
typedef void* cudaStream_t;
typedef int cudaError_t;
enum { cudaSuccess = 0, cudaErrorInvalidValue = 1 };
enum { cudaDevAttrMultiProcessorCount = 16, cudaFuncAttributeMaxDynamicSharedMemorySize = 8,
       cudaFuncAttributePreferredSharedMemoryCarveout = 9, cudaSharedmemCarveoutMaxShared = 100 };
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
inline cudaError_t cudaPeekAtLastError() { return cudaSuccess; }
inline const char* cudaGetErrorString(cudaError_t) { return "emulated"; }
inline cudaError_t cudaGetDevice(int* d) { *d = 0; return cudaSuccess; }
inline cudaError_t cudaDeviceGetAttribute(int* v, int, int) { *v = 1; return cudaSuccess; }
inline cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t = nullptr) { std::memset(p, v, n); return cudaSuccess; }
inline cudaError_t cudaMemset(void* p, int v, size_t n) { std::memset(p, v, n); return cudaSuccess; }
inline cudaError_t cudaMalloc(void** p, size_t n) { *p = std::malloc(n); return *p ? cudaSuccess : cudaErrorInvalidValue; }
inline cudaError_t cudaFree(void* p) { std::free(p); return cudaSuccess; }
template <class F> inline cudaError_t cudaFuncSetAttribute(F, int, int) { return cudaSuccess; }
template <class F> inline cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int* n, F, int, size_t) { *n = 1; return cudaSuccess; }
inline cudaError_t cudaLaunchCooperativeKernel(const void*, dim3, dim3, void**, size_t, cudaStream_t) { return cudaErrorInvalidValue; }
template <class T> inline cudaError_t cudaGetSymbolAddress(void** p, T& sym) { *p = const_cast<void*>(static_cast<const void*>(&sym)); return cudaSuccess; }

// cooperative groups: under launcher_concurrent a grid barrier spans every thread of the grid; under the sequential
// launcher a grid has to be ONE block and the grid barrier is the block barrier
namespace cooperative_groups {
struct grid_group {
  void sync() const {
    std::atomic_thread_fence(std::memory_order_seq_cst);
    if (emu::t_ctx->grid_bar) emu::t_ctx->grid_bar->sync();
    else __syncthreads();
  }
};
inline grid_group this_grid() { return {}; }
}  // namespace cooperative_groups
