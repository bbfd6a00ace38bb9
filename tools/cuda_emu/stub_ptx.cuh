// Emulation stand-in for csrc/ptx.cuh: only the helpers that the PTX-free kernel files use.
#pragma once
#include "emu_cuda.h"
namespace omldm {
inline float warp_sum(float v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
inline float warp_max(float v) {
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
inline float warp_min(float v) {
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
inline float4 ld_nc_v4(const float4* p) { return *p; }
// system-scope loads / stores / atomics of the peer-memory protocol kernels: host atomics with the same ordering
inline void st_release_sys(uint32_t* a, uint32_t v) { emu::jitter(); __atomic_store_n(a, v, __ATOMIC_RELEASE); }
inline void st_relaxed_sys(uint32_t* a, uint32_t v) { emu::jitter(); __atomic_store_n(a, v, __ATOMIC_RELAXED); }
inline uint32_t ld_acquire_sys(const uint32_t* a) { emu::jitter(); return __atomic_load_n(a, __ATOMIC_ACQUIRE); }
inline uint32_t ld_relaxed_sys(const uint32_t* a) { return __atomic_load_n(a, __ATOMIC_RELAXED); }
inline uint32_t atom_add_release_sys(uint32_t* a, uint32_t v) { emu::jitter(); return __atomic_fetch_add(a, v, __ATOMIC_RELEASE); }
inline int32_t atom_add_relaxed_sys_s32(int32_t* a, int32_t v) { emu::jitter(); return __atomic_fetch_add(a, v, __ATOMIC_RELAXED); }
inline void red_add_sys_f32(float* a, float v) { atomicAdd(a, v); }
inline void red_add_v4(float* a, float x, float y, float z, float w) { atomicAdd(a, x); atomicAdd(a + 1, y); atomicAdd(a + 2, z); atomicAdd(a + 3, w); }
inline void red_add_sys_v4(float* a, float4 v) { red_add_v4(a, v.x, v.y, v.z, v.w); }
inline float4 atom_add_sys_v4(float* a, float4 v) {
  return make_float4(atomicAdd(a, v.x), atomicAdd(a + 1, v.y), atomicAdd(a + 2, v.z), atomicAdd(a + 3, v.w));
}
// ld.relaxed.sys / st.relaxed.sys: strong (atomic-like) accesses per 32-bit element, not plain loads / stores - the
// asynchronous protocols read a centre variable that peers are adding to, by design
inline float4 ld_peer_v4(const float4* p) {
  emu::jitter(16);
  const uint32_t* u = reinterpret_cast<const uint32_t*>(p);
  uint32_t w[4];
  for (int i = 0; i < 4; ++i) w[i] = __atomic_load_n(u + i, __ATOMIC_RELAXED);
  float4 r;
  std::memcpy(&r, w, sizeof(r));
  return r;
}
inline void st_peer_v4(float4* p, float4 v) {
  emu::jitter(16);
  uint32_t w[4];
  std::memcpy(w, &v, sizeof(v));
  uint32_t* u = reinterpret_cast<uint32_t*>(p);
  for (int i = 0; i < 4; ++i) __atomic_store_n(u + i, w[i], __ATOMIC_RELAXED);
}
// NVLS multimem needs a switch: never reached in the emulation (the multicast pointer is null)
inline float4 mc_ld_reduce(const float*) { std::abort(); }
inline void mc_st(float*, float4) { std::abort(); }
inline unsigned lane_id() { return static_cast<unsigned>(emu::t_lane); }
}  // namespace omldm
