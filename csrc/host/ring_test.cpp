// Stand-alone stress test of the ring buffer, the generators and the JSON scanner; built with
// -fsanitize=address,undefined by tests/test_host_sanitizers.py ("ASan/UBSan build of the
// host C++ ring buffer").
#include "omldm_host.h"

#include <atomic>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

int main() {
  const int d = 12, ld = 32, batch = 64, slots = 4;
  std::vector<float> mem(static_cast<size_t>(slots) * batch * ld);
  omldm_ring* r = omldm_ring_create(mem.data(), slots, sizeof(float) * batch * ld);
  if (!r) return 1;
  omldm_producer* p = omldm_producer_start(r, /*kind=*/1, d, ld, batch, 0, 42, /*mode=*/0, /*max_batches=*/50, /*threads=*/3);
  if (!p) return 2;
  long long got = 0;
  double checksum = 0;
  for (;;) {
    int64_t n = 0;
    int s = omldm_ring_acquire_read(r, 5000, &n);
    if (s < 0) break;
    const float* rows = static_cast<const float*>(omldm_ring_slot_ptr(r, s));
    for (int i = 0; i < n; ++i) {
      if (rows[static_cast<size_t>(i) * ld + d] != 1.0f) return 3;   // bias / valid column
      checksum += rows[static_cast<size_t>(i) * ld + d + 1];
    }
    got += n;
    omldm_ring_release_read(r, s);
  }
  omldm_producer_stop(p);
  if (got != 50LL * batch) { std::fprintf(stderr, "got %lld\n", got); return 4; }
  if (omldm_ring_consumed(r) != got || omldm_ring_produced(r) != got) return 5;
  omldm_ring_destroy(r);

  // sparse generator stays inside its slot
  const int nnz = 8;
  std::vector<unsigned char> slot(static_cast<size_t>(batch) * (2 * nnz + 1) * 4);
  omldm_generate_sparse_batch(slot.data(), 1000, nnz, batch, 5, 7, 3);
  const int* idx = reinterpret_cast<const int*>(slot.data());
  for (int i = 0; i < batch * nnz; ++i) if (idx[i] < 0 || idx[i] >= 1000) return 6;

  // JSON scanner: malformed input must be skipped, never read out of bounds
  const char* txt = "{\"numericalFeatures\":[1,2,3],\"target\":1}\nEOS\n{\"numericalFeatures\":[1,2\n{]]]]\n"
                    "{\"numericalFeatures\":[4,5,6],\"operation\":\"forecasting\"}";
  std::vector<float> out(8 * 32);
  int nf = 0;
  size_t used = 0;
  int rows = omldm_parse_json_records(txt, std::strlen(txt), 3, out.data(), 32, 8, &nf, &used);
  if (rows != 2 || nf != 1 || used != std::strlen(txt)) { std::fprintf(stderr, "rows %d nf %d\n", rows, nf); return 7; }
  // hostile fragments in an exactly-sized heap buffer (so ASan sees any read past the end): unterminated strings,
  // a backslash as the very last byte, unbalanced brackets, huge exponents / mantissas, missing values
  const char* bad[] = {"{\"numericalFeatures\":[1,2,3],\"s\":\"abc\\", "{\"numericalFeatures\":[1,2,3],\"t\":{\"a\":[1,2",
                       "{\"numericalFeatures\":[1e99999,123456789012345678901234567890.5e-400,3]}", "{\"numericalFeatures\":",
                       "{\"numericalFeatures\":[1,2,3],\"target\":", "{\"", "{\"a", "{\"numericalFeatures\":[-", "{\"operation\":\"tra"};
  for (const char* frag : bad) {
    const size_t n = std::strlen(frag);
    std::vector<char> exact(frag, frag + n);
    omldm_parse_json_records(exact.data(), n, 3, out.data(), 32, 8, &nf, &used);
    if (used != n) return 8;
  }
  // threaded scanner on a large text with garbage lines; row order and counts must match the serial one
  std::string big;
  for (int i = 0; i < 5000; ++i) {
    if (i % 11 == 0) big += "not json\n";
    big += "{\"numericalFeatures\":[" + std::to_string(i) + ",0.5,-1e-3],\"target\":" + std::to_string(i % 7) + "}\n";
  }
  std::vector<float> o1(6000 * 32), o4(6000 * 32);
  int nf1 = 0, nf4 = 0;
  size_t u1 = 0, u4 = 0;
  const int r1 = omldm_parse_json_records(big.data(), big.size(), 3, o1.data(), 32, 6000, &nf1, &u1);
  const int r4 = omldm_parse_json_records_mt(big.data(), big.size(), 3, o4.data(), 32, 6000, &nf4, &u4, 4);
  if (r1 != 5000 || r4 != 5000 || u1 != u4 || std::memcmp(o1.data(), o4.data(), sizeof(float) * 5000 * 32) != 0) return 9;

  // compact fp16 wire rows: generator and packer stay inside their buffers (odd d exercises the pad half)
  const int dw = 13, wld = ((dw + 1) & ~1) + 2;
  std::vector<uint16_t> wire(static_cast<size_t>(batch) * wld);
  omldm_generate_batch_f16(wire.data(), 1, dw, batch, 0, 9, 2);
  std::vector<float> staged(static_cast<size_t>(batch) * 32, 1.0f);
  omldm_pack_wire_f16(staged.data(), 32, dw + 1, batch, dw, wire.data());
  if (wire[0] != 0x3C00) return 10;                                   // 1.0 in fp16
  std::printf("ring_test ok %lld %.3f\n", got, checksum);
  return 0;
}
