/* C ABI of libomldm_host.so - host runtime of omldm_b200 (no CUDA dependency).
 *
 * Ring buffers over caller-provided (pinned) memory stand in for the Kafka topics of the
 * reference (Job.scala:42-67,127-142: trainingData, forecastingData, requests, psMessages ...).
 * A slot holds one mini-batch of staged record rows (see omldm_kernels.h) or one CSR batch.
 */
#ifndef OMLDM_HOST_H_
#define OMLDM_HOST_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct omldm_ring omldm_ring;

/* `base` = n_slots * slot_bytes bytes owned by the caller (pinned host memory). */
omldm_ring* omldm_ring_create(void* base, int n_slots, size_t slot_bytes);
void omldm_ring_destroy(omldm_ring* r);
/* producer side: returns slot index or -1 on timeout / closed */
int omldm_ring_acquire_write(omldm_ring* r, int timeout_ms);
void omldm_ring_commit_write(omldm_ring* r, int slot, int64_t n_records);
/* consumer side: returns slot index or -1 on timeout / end of stream; *n_records filled */
int omldm_ring_acquire_read(omldm_ring* r, int timeout_ms, int64_t* n_records);
void omldm_ring_release_read(omldm_ring* r, int slot);
void omldm_ring_close(omldm_ring* r);            /* producers done: readers drain then get -1 */
int64_t omldm_ring_produced(const omldm_ring* r);
int64_t omldm_ring_consumed(const omldm_ring* r);
void* omldm_ring_slot_ptr(omldm_ring* r, int slot);

/* Synthetic stream producer threads (stand-in for the Kafka source subtasks).
 * kind: 0 binary classification (y = sign(w.x)), 1 regression (y = w.x + b + noise),
 *       2 clusters (k centres + noise, y = cluster id), 3 multiclass (y = argmax of C planes).
 * mode: 0 generate every slot, 1 replay (fill each slot once, then re-publish it unchanged).
 * Rows are written in the staged layout [x(d) | 1 | y | 0-pad] with pitch ld floats. */
typedef struct omldm_producer omldm_producer;
omldm_producer* omldm_producer_start(omldm_ring* r, int kind, int d, int ld, int batch, int n_aux,
                                     uint64_t seed, int mode, int64_t max_batches, int n_threads);
/* Same, with a wire format: 0 = staged fp32 rows (above); 1 = compact rows for the host->device link,
 * [x as fp16 x d | pad to an even count | y as fp32 in two halves], pitch even(d) + 2 halves (the ring's slot
 * size must be batch * pitch * 2 bytes; `ld` >= d + 2 only sizes a scratch row).  Expanded on the GPU by
 * omldm_stage_rows_f16.  2 = fp16 staged rows [x | 1 | y_hi | y_lo | 0-pad], pitch pad64(d + 3) halves, consumed in
 * place by the fp16 Gram kernel.  Neither is available for kind 4. */
omldm_producer* omldm_producer_start_wire(omldm_ring* r, int kind, int d, int ld, int batch, int n_aux,
                                          uint64_t seed, int mode, int64_t max_batches, int n_threads, int wire);
/* Same, resuming at a stream position: the first batch produced has index first_batch (batches first_batch ..
 * max_batches - 1 are produced). */
omldm_producer* omldm_producer_start_from(omldm_ring* r, int kind, int d, int ld, int batch, int n_aux,
                                          uint64_t seed, int mode, int64_t max_batches, int n_threads, int wire,
                                          int64_t first_batch);
void omldm_producer_stop(omldm_producer* p);

/* kind 4 = sparse multiclass stream (BASELINE.json config 5): pass the non-zeros per record as `ld`;
 * a slot is [int32 idx batch*nnz | float val batch*nnz | float target batch] (fixed-width rows, so the
 * CSR row pointer is implicit) and is densified on the GPU (omldm_densify). */
void omldm_generate_sparse_batch(void* slot, int d, int nnz, int batch, int n_aux, uint64_t seed,
                                 uint64_t batch_index);
/* Fill one caller buffer with a synthetic staged batch (same generators, no ring). */
void omldm_generate_batch(float* dst, int kind, int d, int ld, int batch, int n_aux, uint64_t seed,
                          uint64_t batch_index);
/* ... in the compact fp16 wire format (dst: batch * (even(d) + 2) halves). */
void omldm_generate_batch_f16(void* dst, int kind, int d, int batch, int n_aux, uint64_t seed,
                              uint64_t batch_index);
/* ... as fp16 staged rows ("f16s": [x | 1 | y_hi | y_lo | 0-pad], pitch pad64(d + 3) halves; producer wire = 2).
 * This is the layout the fp16 Gram kernel reads in place, so a slot needs no expansion on the GPU. */
void omldm_generate_batch_f16s(void* dst, int kind, int d, int batch, int n_aux, uint64_t seed,
                               uint64_t batch_index);
/* fp32 rows (pitch ld, target in column target_col; staged rows with valid flag 0 stay empty) -> f16s rows */
void omldm_pack_f16s(const float* rows, int ld, int target_col, int batch, int d, void* dst);
/* fp32 rows (pitch ld floats, target in column `target_col`) -> compact fp16 wire rows (round to nearest even,
 * overflow to infinity, subnormals kept). */
void omldm_pack_wire_f16(const float* rows, int ld, int target_col, int batch, int d, void* dst);

/* Parse newline-delimited JSON DataInstances (utils/parsers/DataInstanceParser.scala:12-22,
 * dataStream/DataPointParser.scala:16-55) into staged rows.  Malformed records and "EOS" are
 * skipped silently, like the reference.  Returns the number of rows written;
 * *n_forecast counts rows whose operation was "forecasting" (their valid flag is 2.0). */
int omldm_parse_json_records(const char* text, size_t len, int d, float* dst, int ld, int max_rows,
                             int* n_forecast, size_t* consumed);
/* Same on n_threads threads (text split at line boundaries, row order = line order).  At most max_rows LINES are
 * taken, so dst needs max_rows rows and fewer rows may come back when lines are skipped; *consumed says where to
 * resume. */
int omldm_parse_json_records_mt(const char* text, size_t len, int d, float* dst, int ld, int max_rows,
                                int* n_forecast, size_t* consumed, int n_threads);

#ifdef __cplusplus
}
#endif
#endif
