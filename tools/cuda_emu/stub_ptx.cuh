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
inline unsigned lane_id() { return static_cast<unsigned>(emu::t_lane); }
}  // namespace omldm
