// Bounded multi-producer / single-consumer ring of fixed-size slots over caller-owned memory.
// Slots are handed out in order; a slot is READY for the reader only after commit.
#include "omldm_host.h"

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <vector>

struct omldm_ring {
  uint8_t* base;
  int n_slots;
  size_t slot_bytes;
  std::mutex mu;
  std::condition_variable cv_space, cv_data;
  // state per slot: 0 free, 1 being written, 2 ready, 3 being read
  std::vector<int> state;
  std::vector<int64_t> nrec;
  int64_t w_next = 0;   // next slot sequence number to hand to a writer
  int64_t r_next = 0;   // number of slots the reader has released (space accounting)
  int64_t r_acq = 0;    // next slot sequence number to hand to the reader
  int64_t produced = 0, consumed = 0;
  bool closed = false;
};

extern "C" omldm_ring* omldm_ring_create(void* base, int n_slots, size_t slot_bytes) {
  if (!base || n_slots <= 0 || slot_bytes == 0) return nullptr;
  auto* r = new omldm_ring();
  r->base = static_cast<uint8_t*>(base);
  r->n_slots = n_slots;
  r->slot_bytes = slot_bytes;
  r->state.assign(n_slots, 0);
  r->nrec.assign(n_slots, 0);
  return r;
}
extern "C" void omldm_ring_destroy(omldm_ring* r) { delete r; }

extern "C" int omldm_ring_acquire_write(omldm_ring* r, int timeout_ms) {
  std::unique_lock<std::mutex> lk(r->mu);
  const int slot = static_cast<int>(r->w_next % r->n_slots);
  auto ok = [&] { return r->closed || (r->state[r->w_next % r->n_slots] == 0 && r->w_next - r->r_next < r->n_slots); };
  if (timeout_ms < 0) r->cv_space.wait(lk, ok);
  else if (!r->cv_space.wait_for(lk, std::chrono::milliseconds(timeout_ms), ok)) return -1;
  if (r->closed) return -1;
  (void)slot;
  const int s = static_cast<int>(r->w_next % r->n_slots);
  r->state[s] = 1;
  r->w_next += 1;
  return s;
}
extern "C" void omldm_ring_commit_write(omldm_ring* r, int slot, int64_t n_records) {
  {
    std::lock_guard<std::mutex> lk(r->mu);
    r->state[slot] = 2;
    r->nrec[slot] = n_records;
    r->produced += n_records;
  }
  r->cv_data.notify_all();
}
extern "C" int omldm_ring_acquire_read(omldm_ring* r, int timeout_ms, int64_t* n_records) {
  std::unique_lock<std::mutex> lk(r->mu);
  auto ok = [&] { return r->state[r->r_acq % r->n_slots] == 2 || (r->closed && r->r_acq == r->w_next); };
  if (timeout_ms < 0) r->cv_data.wait(lk, ok);
  else if (!r->cv_data.wait_for(lk, std::chrono::milliseconds(timeout_ms), ok)) return -1;
  const int s = static_cast<int>(r->r_acq % r->n_slots);
  if (r->state[s] != 2) return -1;  // closed and drained
  r->state[s] = 3;
  r->r_acq += 1;
  if (n_records) *n_records = r->nrec[s];
  return s;
}
extern "C" void omldm_ring_release_read(omldm_ring* r, int slot) {
  {
    std::lock_guard<std::mutex> lk(r->mu);
    r->state[slot] = 0;
    r->consumed += r->nrec[slot];
    r->r_next += 1;
  }
  r->cv_space.notify_all();
}
extern "C" void omldm_ring_close(omldm_ring* r) {
  {
    std::lock_guard<std::mutex> lk(r->mu);
    r->closed = true;
  }
  r->cv_space.notify_all();
  r->cv_data.notify_all();
}
extern "C" int64_t omldm_ring_produced(const omldm_ring* r) { return r->produced; }
extern "C" int64_t omldm_ring_consumed(const omldm_ring* r) { return r->consumed; }
extern "C" void* omldm_ring_slot_ptr(omldm_ring* r, int slot) { return r->base + static_cast<size_t>(slot) * r->slot_bytes; }
