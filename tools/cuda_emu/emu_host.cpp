// Error reporting / launch accounting of the emulated kernel library (mirrors csrc/host_common.cu).
#include "host_common.h"

#include <atomic>
#include <cstdio>

namespace omldm {
static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};
int set_error(const char* what, cudaError_t e) { snprintf(g_err, sizeof(g_err), "%s: error %d", what, e); return -1; }
int set_error_msg(const char* msg) { snprintf(g_err, sizeof(g_err), "%s", msg); return -1; }
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
int check_launch(const char*) { count_launch(1); return 0; }
}  // namespace omldm

extern "C" const char* omldm_last_error() { return omldm::g_err; }
extern "C" long long omldm_launch_count() { return omldm::g_launches.load(); }
extern "C" void omldm_add_launch_count(long long n) { omldm::g_launches.fetch_add(n); }
extern "C" void omldm_reset_launch_count() { omldm::g_launches.store(0); }
extern "C" int omldm_abi_version() { return 1; }
extern "C" int omldm_is_emulator() { return 1; }

// ---- sanitizer self-tests: a kernel with a missing __syncthreads (data race on shared memory) and its fixed twin
namespace {
void race_kernel(float* out, int with_barrier) {
  float (&s)[64] = *emu::block_shared<float[64]>(900001);
  s[threadIdx.x] = static_cast<float>(threadIdx.x);
  if (with_barrier) __syncthreads();
  out[threadIdx.x] = s[(threadIdx.x + 1) & 63];
}
}  // namespace
extern "C" int omldm_emu_selftest_shared(float* out64, int with_barrier) {
  emu::launcher(race_kernel, dim3(1), dim3(64))(out64, with_barrier);
  return 0;
}

// ---- wave / jitter self-test: the "last block applies the sum" idiom of linear_minibatch_kernel, with and without the
// block barrier that must precede the completion counter.  scratch: 64 accumulators + 1 counter (zeroed by the caller).
// Every block adds 1 to each accumulator; the last block copies them to out and clears them.  Correct result:
// out[i] == gridDim.x.  Without the barrier the result is wrong in some launches (timing) and ThreadSanitizer reports
// the clearing store against a late addition in every run (no happens-before path).
namespace {
void lastblock_kernel(float* scratch, float* out, int with_barrier) {
  bool& s_last = *emu::block_shared<bool>(900002);
  atomicAdd(&scratch[threadIdx.x], 1.0f);
  if (with_barrier) { __threadfence(); __syncthreads(); }
  if (threadIdx.x == 0) {
    const unsigned done = atomicAdd(reinterpret_cast<unsigned*>(&scratch[64]), 1u);
    s_last = done == gridDim.x - 1;
  }
  __syncthreads();
  if (s_last) {
    out[threadIdx.x] = __ldcg(&scratch[threadIdx.x]);
    scratch[threadIdx.x] = 0.f;          // reset for the next call, as the real kernels do
  }
}
}  // namespace
extern "C" int omldm_emu_selftest_lastblock(float* scratch65, float* out64, int blocks, int with_barrier) {
  emu::launcher(lastblock_kernel, dim3(blocks), dim3(64))(scratch65, out64, with_barrier);
  return 0;
}
