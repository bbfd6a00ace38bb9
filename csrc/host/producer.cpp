// Synthetic stream generators and producer threads: the stand-in for the Kafka source subtasks
// that feed `trainingData` (Job.scala:42-50).  Records are written directly in the staged row
// layout so a ring slot can be copied to the GPU as is.
#include "omldm_host.h"

#include <atomic>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

namespace {

struct Rng {  // xoshiro256** - fast and good enough for synthetic features
  uint64_t s[4];
  explicit Rng(uint64_t seed) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull;
    for (int i = 0; i < 4; ++i) {
      z += 0x9E3779B97F4A7C15ull;
      uint64_t x = z;
      x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
      x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
      s[i] = x ^ (x >> 31);
    }
  }
  static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
  inline uint64_t next() {
    const uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
    return r;
  }
  inline float uniform() { return (next() >> 40) * (1.0f / 16777216.0f); }
  // approx N(0,1): sum of 4 uniforms, variance-corrected (cheap; tails do not matter here)
  inline float normal() {
    const uint64_t a = next();
    const float u = ((a & 0xFFFF) + ((a >> 16) & 0xFFFF) + ((a >> 32) & 0xFFFF) + (a >> 48)) * (1.0f / 65536.0f);
    return (u - 2.0f) * 1.7320508f;
  }
};

// hidden model shared by all batches of a stream: derived from the seed only
struct Truth {
  std::vector<float> w;  // kind 0/1: d weights; kind 2: n_aux centres x d; kind 3: n_aux planes x d
  float bias;
  Truth(int kind, int d, int n_aux, uint64_t seed) {
    Rng g(seed ^ 0xABCDEF1234ull);
    const int rows = (kind >= 2) ? (n_aux > 0 ? n_aux : 1) : 1;   // kind 3 / 4: one plane per class
    w.resize(static_cast<size_t>(rows) * d);
    const float scale = kind == 2 ? 4.0f : 1.0f / std::sqrt(static_cast<float>(d));
    for (auto& v : w) v = g.normal() * scale;
    bias = kind == 1 ? 0.5f : 0.0f;
  }
};

// fp32 -> fp16 bits, round to nearest even, overflow to infinity, subnormals kept
inline uint16_t f32_to_f16(float f) {
  uint32_t x;
  std::memcpy(&x, &f, 4);
  const uint32_t sign = (x >> 16) & 0x8000u;
  x &= 0x7FFFFFFFu;
  if (x >= 0x7F800000u) return static_cast<uint16_t>(sign | 0x7C00u | ((x > 0x7F800000u) ? 0x200u : 0u));
  if (x >= 0x477FF000u) return static_cast<uint16_t>(sign | 0x7C00u);          // rounds to >= 65520 -> inf
  if (x < 0x33000001u) return static_cast<uint16_t>(sign);                     // below half the smallest subnormal
  if (x < 0x38800000u) {                                                       // subnormal half
    const int shift = 126 - static_cast<int>(x >> 23);                         // 14..24
    const uint32_t m = (x & 0x7FFFFFu) | 0x800000u;
    const uint32_t q = m >> shift, rem = m & ((1u << shift) - 1u), half = 1u << (shift - 1);
    return static_cast<uint16_t>(sign | (q + ((rem > half || (rem == half && (q & 1u))) ? 1u : 0u)));
  }
  const uint32_t q = (x - 0x38000000u) >> 13, rem = x & 0x1FFFu;
  return static_cast<uint16_t>(sign | (q + ((rem > 0x1000u || (rem == 0x1000u && (q & 1u))) ? 1u : 0u)));
}

// fp16 bits -> fp32 (exact)
inline float f16_to_f32(uint16_t h) {
  const uint32_t sign = (static_cast<uint32_t>(h) & 0x8000u) << 16;
  uint32_t e = (h >> 10) & 0x1Fu, m = h & 0x3FFu, x;
  if (e == 0) {
    if (m == 0) x = sign;
    else {
      int sh = 0;
      while (!(m & 0x400u)) { m <<= 1; ++sh; }
      x = sign | ((113u - sh) << 23) | ((m & 0x3FFu) << 13);
    }
  } else if (e == 31) x = sign | 0x7F800000u | (m << 13);
  else x = sign | ((e + 112u) << 23) | (m << 13);
  float f;
  std::memcpy(&f, &x, 4);
  return f;
}

// one fp16 staged row ("f16s", include/omldm_kernels.h): [x | 1 | y_hi | y_lo | 0-pad], pitch `pitch` halves
inline void write_f16s_row(uint16_t* w, const float* x, float y, int d, int pitch) {
  for (int j = 0; j < d; ++j) w[j] = f32_to_f16(x[j]);
  w[d] = 0x3C00u;                                     // 1.0
  const uint16_t yh = f32_to_f16(y);
  w[d + 1] = yh;
  w[d + 2] = f32_to_f16(y - f16_to_f32(yh));
  for (int j = d + 3; j < pitch; ++j) w[j] = 0;
}
inline int f16s_pitch(int d) { return (d + 3 + 63) / 64 * 64; }

// wire = 0: staged fp32 rows (pitch ld floats).  wire = 1: compact rows [fp16 x d | pad | fp32 target],
// pitch even(d) + 2 halves.  wire = 2: fp16 staged rows [x | 1 | y_hi | y_lo | 0-pad], pitch pad64(d + 3) halves
// (include/omldm_host.h).
void fill_batch(void* dst_any, const Truth& tr, int kind, int d, int ld, int batch, int n_aux, uint64_t seed,
                uint64_t batch_index, int wire = 0) {
  Rng g(seed * 0x100000001B3ull + batch_index * 0x9E3779B97F4A7C15ull + 1);
  float* dst = static_cast<float*>(dst_any);
  std::vector<float> tmp(wire ? static_cast<size_t>(ld) : 0);
  const int wld = ((d + 1) & ~1) + 2;
  for (int r = 0; r < batch; ++r) {
    float* row = wire ? tmp.data() : dst + static_cast<size_t>(r) * ld;
    float y = 0.f;
    if (kind == 2) {
      const int c = static_cast<int>(g.next() % static_cast<uint64_t>(n_aux > 0 ? n_aux : 1));
      const float* ctr = tr.w.data() + static_cast<size_t>(c) * d;
      for (int j = 0; j < d; ++j) row[j] = ctr[j] + g.normal();
      y = static_cast<float>(c);
    } else {
      for (int j = 0; j < d; ++j) row[j] = g.normal();
      if (kind == 3) {
        int best = 0;
        float bs = -1e30f;
        for (int c = 0; c < n_aux; ++c) {
          const float* p = tr.w.data() + static_cast<size_t>(c) * d;
          float s = 0.f;
          for (int j = 0; j < d; ++j) s += p[j] * row[j];
          if (s > bs) { bs = s; best = c; }
        }
        y = static_cast<float>(best);
      } else {
        float s = tr.bias;
        for (int j = 0; j < d; ++j) s += tr.w[j] * row[j];
        y = kind == 0 ? (s >= 0.f ? 1.f : -1.f) : s + 0.01f * g.normal();
      }
    }
    row[d] = 1.0f;
    row[d + 1] = y;
    for (int j = d + 2; j < ld; ++j) row[j] = 0.f;
    if (wire == 2) {
      const int pitch = f16s_pitch(d);
      write_f16s_row(static_cast<uint16_t*>(dst_any) + static_cast<size_t>(r) * pitch, row, y, d, pitch);
    } else if (wire) {
      uint16_t* w = static_cast<uint16_t*>(dst_any) + static_cast<size_t>(r) * wld;
      for (int j = 0; j < d; ++j) w[j] = f32_to_f16(row[j]);
      if (d & 1) w[d] = 0;
      std::memcpy(w + wld - 2, &y, 4);
    }
  }
}

// Sparse multiclass batch in fixed-width (ELL) form: slot = [int32 idx B*nnz | float val B*nnz | float y B].
// The label is the argmax over n_aux hidden planes evaluated on the non-zeros only.
void fill_sparse_batch(uint8_t* slot, const Truth& tr, int d, int nnz, int batch, int n_aux, uint64_t seed,
                       uint64_t batch_index) {
  Rng g(seed * 0x100000001B3ull + batch_index * 0x9E3779B97F4A7C15ull + 7);
  int32_t* idx = reinterpret_cast<int32_t*>(slot);
  float* val = reinterpret_cast<float*>(slot + sizeof(int32_t) * static_cast<size_t>(batch) * nnz);
  float* tgt = val + static_cast<size_t>(batch) * nnz;
  for (int r = 0; r < batch; ++r) {
    float best = -1e30f;
    int bc = 0;
    int32_t* ri = idx + static_cast<size_t>(r) * nnz;
    float* rv = val + static_cast<size_t>(r) * nnz;
    for (int j = 0; j < nnz; ++j) {
      ri[j] = static_cast<int32_t>(g.next() % static_cast<uint64_t>(d));
      rv[j] = g.normal();
    }
    for (int c = 0; c < n_aux; ++c) {
      const float* p = tr.w.data() + static_cast<size_t>(c) * d;
      float s = 0.f;
      for (int j = 0; j < nnz; ++j) s += p[ri[j]] * rv[j];
      if (s > best) { best = s; bc = c; }
    }
    tgt[r] = static_cast<float>(bc);
  }
}

}  // namespace

struct omldm_producer {
  omldm_ring* ring;
  int kind, d, ld, batch, n_aux, mode, n_threads, wire = 0;
  uint64_t seed;
  int64_t max_batches;
  std::atomic<int64_t> next_batch{0};
  std::atomic<bool> stop{false};
  std::atomic<int> live{0};
  std::vector<std::thread> threads;
  std::vector<char> filled;  // replay mode: slot already holds data
  Truth truth;
  omldm_producer(int kind_, int d_, int n_aux_, uint64_t seed_) : truth(kind_, d_, n_aux_, seed_) {}
};

static void producer_loop(omldm_producer* p) {
  while (!p->stop.load()) {
    if (p->max_batches >= 0 && p->next_batch.load() >= p->max_batches) break;
    const int slot = omldm_ring_acquire_write(p->ring, 50);
    if (slot < 0) continue;
    const int64_t b = p->next_batch.fetch_add(1);
    if (p->max_batches >= 0 && b >= p->max_batches) {
      omldm_ring_commit_write(p->ring, slot, 0);   // empty marker batch, the reader skips it
      break;
    }
    float* dst = static_cast<float*>(omldm_ring_slot_ptr(p->ring, slot));
    if (p->mode == 0 || !p->filled[slot]) {
      if (p->kind == 4)   // sparse multiclass: `ld` carries the non-zeros per record
        fill_sparse_batch(reinterpret_cast<uint8_t*>(dst), p->truth, p->d, p->ld, p->batch, p->n_aux, p->seed,
                          static_cast<uint64_t>(b));
      else
        fill_batch(dst, p->truth, p->kind, p->d, p->ld, p->batch, p->n_aux, p->seed, static_cast<uint64_t>(b), p->wire);
      p->filled[slot] = 1;
    }
    omldm_ring_commit_write(p->ring, slot, p->batch);
  }
  if (p->live.fetch_sub(1) == 1) omldm_ring_close(p->ring);
}

extern "C" omldm_producer* omldm_producer_start_wire(omldm_ring* r, int kind, int d, int ld, int batch, int n_aux,
                                                     uint64_t seed, int mode, int64_t max_batches, int n_threads,
                                                     int wire) {
  if (!r || d <= 0 || batch <= 0 || (kind != 4 && ld < d + 2) || (kind == 4 && ld <= 0)) return nullptr;
  if (wire != 0 && ((wire != 1 && wire != 2) || kind == 4)) return nullptr;
  auto* p = new omldm_producer(kind, d, n_aux, seed);
  p->ring = r; p->kind = kind; p->d = d; p->ld = ld; p->batch = batch; p->n_aux = n_aux; p->mode = mode;
  p->wire = wire;
  p->seed = seed; p->max_batches = max_batches;
  p->n_threads = n_threads < 1 ? 1 : n_threads;
  p->filled.assign(1 << 16, 0);
  p->live.store(p->n_threads);
  for (int i = 0; i < p->n_threads; ++i) p->threads.emplace_back(producer_loop, p);
  return p;
}

// Same, resuming a stream: the first batch produced has index `first_batch` (the stream position stored in a
// checkpoint), max_batches still counts from 0.
extern "C" omldm_producer* omldm_producer_start_from(omldm_ring* r, int kind, int d, int ld, int batch, int n_aux,
                                                     uint64_t seed, int mode, int64_t max_batches, int n_threads,
                                                     int wire, int64_t first_batch) {
  if (!r || d <= 0 || batch <= 0 || (kind != 4 && ld < d + 2) || (kind == 4 && ld <= 0)) return nullptr;
  if (wire != 0 && ((wire != 1 && wire != 2) || kind == 4)) return nullptr;
  auto* p = new omldm_producer(kind, d, n_aux, seed);
  p->ring = r; p->kind = kind; p->d = d; p->ld = ld; p->batch = batch; p->n_aux = n_aux; p->mode = mode;
  p->wire = wire;
  p->seed = seed; p->max_batches = max_batches;
  p->next_batch.store(first_batch < 0 ? 0 : first_batch);
  p->n_threads = n_threads < 1 ? 1 : n_threads;
  p->filled.assign(1 << 16, 0);
  p->live.store(p->n_threads);
  for (int i = 0; i < p->n_threads; ++i) p->threads.emplace_back(producer_loop, p);
  return p;
}

extern "C" omldm_producer* omldm_producer_start(omldm_ring* r, int kind, int d, int ld, int batch, int n_aux,
                                                uint64_t seed, int mode, int64_t max_batches, int n_threads) {
  return omldm_producer_start_wire(r, kind, d, ld, batch, n_aux, seed, mode, max_batches, n_threads, 0);
}

extern "C" void omldm_producer_stop(omldm_producer* p) {
  if (!p) return;
  p->stop.store(true);
  for (auto& t : p->threads) t.join();
  delete p;
}

extern "C" void omldm_generate_sparse_batch(void* slot, int d, int nnz, int batch, int n_aux, uint64_t seed,
                                            uint64_t batch_index) {
  Truth tr(4, d, n_aux, seed);
  fill_sparse_batch(static_cast<uint8_t*>(slot), tr, d, nnz, batch, n_aux, seed, batch_index);
}

extern "C" void omldm_generate_batch(float* dst, int kind, int d, int ld, int batch, int n_aux, uint64_t seed,
                                     uint64_t batch_index) {
  Truth tr(kind, d, n_aux, seed);
  fill_batch(dst, tr, kind, d, ld, batch, n_aux, seed, batch_index);
}

extern "C" void omldm_generate_batch_f16(void* dst, int kind, int d, int batch, int n_aux, uint64_t seed,
                                         uint64_t batch_index) {
  Truth tr(kind, d, n_aux, seed);
  fill_batch(dst, tr, kind, d, d + 2, batch, n_aux, seed, batch_index, 1);
}

extern "C" void omldm_generate_batch_f16s(void* dst, int kind, int d, int batch, int n_aux, uint64_t seed,
                                          uint64_t batch_index) {
  Truth tr(kind, d, n_aux, seed);
  fill_batch(dst, tr, kind, d, d + 2, batch, n_aux, seed, batch_index, 2);
}

extern "C" void omldm_pack_f16s(const float* rows, int ld, int target_col, int batch, int d, void* dst) {
  const int pitch = f16s_pitch(d);
  for (int r = 0; r < batch; ++r) {
    const float* row = rows + static_cast<size_t>(r) * ld;
    uint16_t* w = static_cast<uint16_t*>(dst) + static_cast<size_t>(r) * pitch;
    if (ld > d + 1 && target_col == d + 1 && row[d] == 0.f) {       // empty staged row stays empty
      for (int j = 0; j < pitch; ++j) w[j] = 0;
      continue;
    }
    write_f16s_row(w, row, row[target_col], d, pitch);
  }
}

extern "C" void omldm_pack_wire_f16(const float* rows, int ld, int target_col, int batch, int d, void* dst) {
  const int wld = ((d + 1) & ~1) + 2;
  for (int r = 0; r < batch; ++r) {
    const float* row = rows + static_cast<size_t>(r) * ld;
    uint16_t* w = static_cast<uint16_t*>(dst) + static_cast<size_t>(r) * wld;
    for (int j = 0; j < d; ++j) w[j] = f32_to_f16(row[j]);
    if (d & 1) w[d] = 0;
    std::memcpy(w + wld - 2, row + target_col, 4);
  }
}
