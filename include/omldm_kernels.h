/* C ABI of libomldm_kernels.so - the sm_100a kernel library of omldm_b200.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - `stream` is a cudaStream_t passed as void*; all entry points are asynchronous on it;
 *   - return value 0 = launched, non-zero = rejected; `omldm_last_error()` has the reason;
 *   - "staged record row" (Z row, pitch `ld` floats, multiple of 32, 128-byte aligned base):
 *         [ x_0 .. x_{d-1} | 1.0 bias/valid (0.0 = empty row) | target | zero pad ]
 *
 * Each declaration cites the reference call site whose work it replaces (paths relative to
 * src/main/scala/omldm/ of ArisKonidaris/OMLDM); the reference itself has no native code.
 */
#ifndef OMLDM_KERNELS_H_
#define OMLDM_KERNELS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ library */
const char* omldm_last_error(void);
long long omldm_launch_count(void);       /* kernels launched by this library so far */
void omldm_reset_launch_count(void);
void omldm_add_launch_count(long long n); /* kernels run by replaying a captured graph */
int omldm_abi_version(void);

/* ------------------------------------------------------------------ tcgen05 GEMMs (umma_gemm.cu)
 * flags: 1 = accumulate into C (red.add), 2 = skip tiles strictly below the diagonal,
 *        4 = load MN-major slabs with per-slab 2-D TMA boxes, 16 = do not use the TMA store /
 *        reduce-add epilogue (it is used automatically when C is 16-byte aligned, ldc % 4 == 0).
 * Learner maths behind node.receiveTuple, operators/spoke/FlinkSpoke.scala:101. */
int omldm_gemm_tn(const float* A, int lda, const float* B, int ldb, float* C, int ldc,
                  int M, int N, int K, int ksplit, int flags, void* stream);
int omldm_gemm_outer(const float* A, int lda, const float* B, int ldb, float* C, int ldc,
                     int M, int N, int K, int ksplit, int flags, void* stream);
/* C[M,N] (+)= A[M,K] B[K,N], both row-major (A K-major, B MN-major: ldb % 32 == 0, 128-byte aligned): the
 * back-propagated gradient G W of the NN learner. */
int omldm_gemm_nn(const float* A, int lda, const float* B, int ldb, float* C, int ldc,
                  int M, int N, int K, int ksplit, int flags, void* stream);

/* ------------------------------------------------------------------ fp16 Gram update (gram_f16.cu)
 * G (+)= Z^T Z for a mini-batch of fp16 staged rows (ORR; per-record update behind node.receiveTuple,
 * operators/spoke/FlinkSpoke.scala:101; worker -> hub push network/FlinkNetwork.scala:260-261).
 * fp16 staged row, pitch `ld` halves (multiple of 64, 128-byte aligned base):
 *         [ x_0 .. x_{d-1} | 1.0 bias/valid (0 = empty row) | y_hi | y_lo | zero pad ],  target = y_hi + y_lo
 * n = d + 3 columns take part.  Only 32 x 32 blocks that touch the upper triangle are written.
 * The partial sums are reduce-added into C (n x n fp32, pitch ldc): this worker's running sum.
 *   peers != NULL : additionally, the last CTA to finish a 128-row half tile copies the updated rows of C into this
 *                   worker's slot at the GPU that owns those rows (owner_base[row / rows_per_owner] = base of the slot
 *                   matrix, same pitch ldc; plain TMA stores over NVLink, single writer per slot: no remote atomics,
 *                   no flags).  The copy is complete when the kernel is; owners sum their slots in
 *                   omldm_sync_gram_gather.  Nothing waits for another GPU.
 * scratch: omldm_gram_f16_scratch_words() uint32 words, zero before the first call, private to one (C, peers) pair.
 * ksplit 0 = automatic (fills the 74 SM pairs). */
typedef struct omldm_gram_peers {
  float* owner_base[16];     /* this worker's slot matrix at every owner (peer mapped), pitch ldc floats */
  int world;
  int rank;
  int rows_per_owner;        /* multiple of 32; owner p holds rows [p * rows_per_owner, (p + 1) * rows_per_owner) */
  int lead;                  /* 0: tail push - after the call the slots hold this call's result (last CTA of a half tile
                                copies it).  1: lead push - the slots receive C as it was BEFORE this call, copied while
                                the MMAs run (free); run omldm_sync_gram_push before the slots are read. */
} omldm_gram_peers;
/* Optional: the prequential loss of the batch from inside the same kernel (the CTAs of the diagonal tiles form the
 * residual dots z . wt from the shared-memory stages the MMAs are reading; no second pass over the batch).
 * stats[0] += sum over valid rows of (z . wt)^2, stats[2] += #valid rows; with curve_ring, one learning-curve sample
 * (stats[0], stats[2]) per call like omldm_curve_sample.  dots: K floats, words: omldm_gram_f16_loss_words() uint32 -
 * both zero before the first call (the kernel leaves them zero), private to one stream of calls. */
typedef struct omldm_gram_loss {
  const float* wt;           /* d + 3 floats: [w_0 .. w_{d-1}, b, -1, -1] */
  float* dots;
  unsigned int* words;
  double* stats;
  double* curve_ring;        /* [curve_cap][2] or NULL */
  long long* curve_counter;
  int curve_cap;
  int d;
} omldm_gram_loss;
size_t omldm_gram_f16_loss_words(void);
int omldm_gram_f16_loss(const void* Z, int ld, int K, int n, float* C, int ldc, const omldm_gram_peers* peers_host,
                        uint32_t* scratch, int ksplit, const omldm_gram_loss* loss_host, void* stream);
size_t omldm_gram_f16_scratch_words(void);
int omldm_gram_f16_rows_per_split(int K, int n, int ksplit);   /* records per batch-axis split the kernel will use */
int omldm_gram_f16(const void* Z, int ld, int K, int n, float* C, int ldc, const omldm_gram_peers* peers_host,
                   uint32_t* scratch, int ksplit, void* stream);

/* ------------------------------------------------------------------ fp16 staged rows (rows_f16.cu)
 * f16s rows -> fp32 staged rows (target = y_hi + y_lo); for learners without an fp16 path and for scoring.
 * Conversion behind utils/parsers/dataStream/DataPointParser.scala:16-55. */
int omldm_expand_rows_f16s(const void* Zh, int ldh, int B, int d, float* Z, int ld, void* stream);
/* stats[0] += sum over valid rows of (w.x + b - y)^2, stats[2] += #valid, straight from f16s rows;
 * wt = [w_0..w_{d-1}, b, -1, -1] (d + 3 floats).  Predict-before-fit of ORR, operators/spoke/FlinkSpoke.scala:101.
 * rows_per_range > 0: sweep the batch as ceil(B / rows_per_range) ranges side by side, each in row order (pass
 * omldm_gram_f16_rows_per_split so that the pass walks the rows in the order of the concurrent Gram kernel and the two
 * share them through L2); 0 = one range. */
int omldm_sq_residual_f16s(const void* Zh, int ldh, int B, int d, const float* wt, double* stats, int rows_per_range,
                           void* stream);
/* The same pass with (a) max_blocks > 0: at most that many blocks in total (the pass is a grid-stride loop; beside a
 * kernel that holds registers on every SM, e.g. omldm_gram_f16, 2 x 148 keeps it one wave), (b) curve_ring != NULL: the
 * last block to finish writes one learning-curve sample (stats[0], stats[2]) into ring[(counter % cap)] and advances
 * the counter - omldm_curve_sample without a launch of its own; done_word = one zero-initialised word of scratch. */
int omldm_sq_residual_f16s_tail(const void* Zh, int ldh, int B, int d, const float* wt, double* stats,
                                int rows_per_range, int max_blocks, double* curve_ring, int curve_cap,
                                long long* curve_counter, unsigned int* done_word, void* stream);

/* ------------------------------------------------------------------ linear learners (linear_learners.cu)
 * PA / RegressorPA / SVM named in utils/parsers/requestStream/PipelineMap.scala:68. */
typedef struct omldm_linear_desc {
  float* w;        /* d + 1 weights, w[d] = intercept */
  double* stats;   /* [loss_sum, aux_sum (mistakes | squared error), fitted, svm_step] */
  int d;
  int kind;        /* 0 PA classifier, 1 PA regressor, 2 SVM (Pegasos) */
  int variant;     /* PA family: 0 PA, 1 PA-I, 2 PA-II */
  float C;
  float eps;       /* regressor insensitivity */
  float lambda;    /* SVM regularisation */
} omldm_linear_desc;

/* exact per-record order; one warp per pipeline; descs_dev is a device array */
int omldm_linear_exact(const float* Z, int ld, int B, const omldm_linear_desc* descs_dev,
                       int n_pipelines, int max_d, void* stream);
/* mini-batch update; `scratch` must hold omldm_linear_minibatch_scratch_floats(d) floats and be
 * zero before the first call (the kernel re-zeroes it on exit) */
size_t omldm_linear_minibatch_scratch_floats(int d);
int omldm_linear_minibatch(const float* Z, int ld, int B, const omldm_linear_desc* desc_host,
                           float* scratch, int mean_update, void* stream);
/* Same update as omldm_linear_minibatch, but it allocates nothing and takes no scratch pointer:
 * the accumulator sits in a fixed-size array inside the library's device image.  This entry point
 * has no counterpart in the reference.  It exists for callers that cannot allocate device memory
 * when they call in - a latency-critical ingest loop that must not hit the device-wide
 * synchronisation of cudaMalloc/cudaFree, or a host process that gave away all of its HBM at
 * start-up and does not permit driver allocations afterwards.
 *   - limit: d <= OMLDM_STATIC_SCRATCH_MAX_D, larger models are rejected with an error (use the
 *     entry point above); every GPU that loads the library holds the
 *     OMLDM_STATIC_SCRATCH_FLOATS * 4 bytes until the process exits;
 *   - the first call on each device resolves the array's device address (cudaGetSymbolAddress);
 *     make that first call before starting any stream capture on that device;
 *   - all calls on a device share the one array: the call clears what it needs on `stream` first,
 *     and two calls that overlap on the GPU (different streams, no ordering) race on it;
 *   - calls no cudaMalloc / cudaFree / allocating library routine. */
#define OMLDM_STATIC_SCRATCH_FLOATS 4096
#define OMLDM_STATIC_SCRATCH_MAX_D (OMLDM_STATIC_SCRATCH_FLOATS - 3)
int omldm_linear_minibatch_static(const float* Z, int ld, int B, const omldm_linear_desc* desc_host,
                                  int mean_update, void* stream);
/* forecast path, operators/spoke/FlinkSpoke.scala:105 */
int omldm_linear_predict(const float* Z, int ld, int B, const float* w, int d, float* out,
                         void* stream);

/* ------------------------------------------------------------------ ingest (ingest.cu)
 * 2-in-10 holdout with evict-then-train, operators/spoke/FlinkSpoke.scala:92-107.
 * `count0` = number of training records seen before this batch; holdout ring = `cap` rows. */
int omldm_holdout_swap(float* Z, int ld, int B, int d, float* ring, int cap, long long count0,
                       void* stream);
/* CUDA-graph friendly variant: count_dev[0] = running record count in device memory, read by every CTA and advanced
 * by B by the last CTA to finish; count_dev[1] = arrival counter of the kernel (zero before the first call) */
int omldm_holdout_swap_dev(float* Z, int ld, int B, int d, float* ring, int cap,
                           long long* count_dev, void* stream);
/* same contract, for consumers that only SUM over the batch (every mini-batch learner): one exchange per ring slot
 * (ring <- last holdout of the slot's chain, that row <- evicted occupant) instead of the shift along the chain: the
 * same records are trained on and the same records stay in the ring, at a fraction of the rows moved */
int omldm_holdout_exchange_dev(float* Z, int ld, int B, int d, float* ring, int cap,
                               long long* count_dev, void* stream);
int omldm_advance_count(long long* count_dev, long long delta, void* stream);
/* learning curve sample (operators/hub/FlinkHub.scala:101-113): ring[2 * (*counter % cap)] = (stats[0], stats[2]), ++*counter */
int omldm_curve_sample(const double* stats, double* ring, int cap, long long* counter, void* stream);
/* sparse (CSR batch) -> dense staged rows + squared row norms (SparseVector, utils/CommonUtils.scala:23) */
int omldm_densify(const int* indptr, const int* indices, const float* values, int B, int d,
                  const float* targets, float* Z, int ld, float* sqnorm, void* stream);
/* raw ring slot [B, d+1] (features..., target) -> staged rows */
int omldm_stage_rows(const float* raw, int raw_ld, int B, int d, float* Z, int ld, void* stream);
/* Compact wire format -> staged rows.  A wire row is [x as fp16 x d | pad to an even count | target as fp32
 * stored in two halves], pitch `raw_ld` halves (even, >= even(d) + 2); `raw` 4-byte aligned.  Halves the
 * bytes per record on the host->device link; the reference's wire format is JSON text
 * (utils/parsers/DataInstanceParser.scala:12-22), so this has no counterpart there. */
int omldm_stage_rows_f16(const void* raw, int raw_ld, int B, int d, float* Z, int ld, void* stream);

/* ------------------------------------------------------------------ preprocessors (preprocess.cu)
 * PolynomialFeatures / StandardScaler / MinMaxScaler, PipelineMap.scala:67. */
int omldm_poly_features(const float* Zin, int ldin, int d_in, const short* factors, int degree,
                        int d_out, float* Zout, int ldout, int B, void* stream);
/* state: double[1 + 2*d] = [count, mean[d], M2[d]]; scratch: double[3 + 4*d], zero before first call */
int omldm_standard_scaler(float* Z, int ld, int d, int B, double* state, double* scratch, int fit,
                          void* stream);
/* state: float[2*d] = [min[d], max[d]] */
int omldm_minmax_scaler(float* Z, int ld, int d, int B, float* state, int fit, void* stream);

/* ------------------------------------------------------------------ fused pipeline (fused_pipeline.cu)
 * PolynomialFeatures -> StandardScaler(fit) -> mini-batch linear learner in two passes over the RAW rows,
 * never materialising the expanded features (BASELINE config 3).  State / scratch layouts are those of
 * omldm_standard_scaler (width d_out) and omldm_linear_minibatch; same result as running the three kernels. */
int omldm_fused_poly_scaler_linear(const float* Z, int ld, int d_in, int B, const short* factors, int degree,
                                   int d_out, double* scaler_state, double* scaler_scratch,
                                   const omldm_linear_desc* desc_host, float* lin_scratch, int mean_update,
                                   void* stream);

/* ------------------------------------------------------------------ k-means / multiclass PA (cluster.cu)
 * K-means and MultiClassPA, PipelineMap.scala:68. */
/* scores S[B,K] = X C^T (from omldm_gemm_tn); writes one-hot assignment T[B,ldt], stats += {inertia, n} */
int omldm_kmeans_assign(const float* Z, int ld, int d, int B, const float* S, int lds,
                        const float* cnorm, int K, float* T, int ldt, int* assign, double* stats,
                        void* stream);
/* the same assignment as the EPILOGUE of the tcgen05 distance GEMM X C^T (umma_gemm.cu): argmin straight from the TMEM
 * accumulator, the scores never reach global memory.  centroids Cm[K, ldcm] (rows of d floats), K <= 256,
 * ldt % 32 == 0 and >= K (pad columns of T are written as zeros). */
int omldm_kmeans_assign_fused(const float* Z, int ld, int d, int B, const float* Cm, int ldcm, const float* cnorm,
                              int K, float* T, int ldt, int* assign, double* stats, void* stream);
/* centroids = (sums + sums2) / counts (count column = d; sums2 may be NULL), refreshes ||c||^2 */
int omldm_kmeans_finalize(const float* sums, const float* sums2, int lds, int K, int d, float* centroids,
                          int ldc, float* cnorm, void* stream);
/* exact (MacQueen, per record) k-means on one CTA */
int omldm_kmeans_exact(const float* Z, int ld, int d, int B, float* sums, int lds, int K,
                       double* stats, void* stream);
/* scores S[B,C] -> update coefficients T[B,ldt] (+tau at y, -tau at runner-up), stats += {loss, mistakes, n} */
int omldm_mcpa_coef(const float* Z, int ld, int d, int B, const float* S, int lds, int C,
                    const float* sqnorm, int variant, float Creg, float* T, int ldt, double* stats,
                    void* stream);
int omldm_row_sqnorm(const float* Z, int ld, int d_plus_bias, int B, float* out, void* stream);
/* y[i] (+)= alpha * x[i]  /  generic helpers used by learners and protocols */
int omldm_axpy(float* y, const float* x, float alpha, size_t n, void* stream);
/* T[:, :C] *= (reciprocal ? 1/max(*scale_dev,1) : *scale_dev) */
int omldm_scale_rows(float* T, int ldt, int B, int C, const float* scale_dev, int reciprocal, void* stream);
/* stats[0] += sum_r (Z_r . wt)^2 over valid rows, stats[2] += #valid: squared residuals with wt = [w, b, -1] */
int omldm_sq_residual(const float* Z, int ld, int B, int d, const float* wt, int n, double* stats, void* stream);
/* symmetric completion: A[j][i] = A[i][j] for j > i (upper -> lower) */
int omldm_mirror_upper(float* A, int ld, int n, void* stream);

/* ------------------------------------------------------------------ Hoeffding tree statistics (tree.cu)
 * HT, PipelineMap.scala:68 (never distributed by the reference: FlinkSpoke.scala:204-206).
 * Tree arrays are indexed by node: feature (-1 = leaf), threshold, left, right, slot (leaf -> statistics
 * slot).  cnt[slot][C], sum/sq[slot][C][d] are updated with atomics; stats += {mistakes, -, fitted}. */
int omldm_ht_accumulate(const float* Z, int ld, int d, int B, const int* feature, const float* threshold,
                        const int* left, const int* right, const int* slot, int C, float* cnt, float* sum,
                        float* sq, int* leaf_out, double* stats, void* stream);
int omldm_ht_route(const float* Z, int ld, int B, const int* feature, const float* threshold, const int* left,
                   const int* right, int* leaf_out, void* stream);
/* Hoeffding split test + tree growth on the device (SURVEY K-L8): per leaf with >= grace new records, best / runner-up
 * information gain over features x n_points Gaussian thresholds, Hoeffding bound (delta, tie), then the accepted
 * splits are applied in ascending leaf order (node numbering as in the sequential algorithm).  2 <= C <= 16.
 * n_nodes_dev: device int, nodes in use (updated); checked: per-node record total at its last test;
 * decision: cap x int4 scratch (16-byte aligned); n_splits_out: optional device int. */
int omldm_ht_split(int* feature, float* threshold, int* left, int* right, int* n_nodes_dev, int cap, int C, int d,
                   float* cnt, const float* sum, const float* sq, float* checked, float grace, float delta, float tie,
                   int n_points, void* decision, int* n_splits_out, void* stream);

/* ------------------------------------------------------------------ fused protocol kernels (sync.cu)
 * worker<->hub synchronisation, network/FlinkNetwork.scala:242-293 (send / broadcast) and the
 * Kafka feedback loop Job.scala:77-87,136-142.  Peer pointer tables live in device memory. */
typedef struct omldm_comm {
  float* const* bufs;        /* k symmetric model/delta buffers (peer-mapped)            */
  float* const* hubs;        /* k symmetric hub-shard buffers (async / SSP / EASGD)      */
  uint32_t* const* pads;     /* k symmetric signal pads (uint32 words)                   */
  uint32_t* local;           /* private scratch words of this rank (>= 64 uint32)        */
  float* mc_buf;             /* NVLS multicast alias of bufs (NULL = use P2P)            */
  int rank;
  int world;
  int emulate;               /* >1: one launch plays `emulate` ranks (single-GPU tests)  */
  int nhub;                  /* hub shards (HubParallelism, FlinkSpoke.scala:181-190): ranks 0..nhub-1 own one; 0 = all */
} omldm_comm;

/* Sequence numbers (`epoch`, `clock`, `step`) passed as 0 are device managed: the kernel keeps the
 * counter in `local` and advances it itself, which makes the launch replayable from a CUDA graph.
 * Do not mix explicit and device-managed numbering on one omldm_comm. */
/* signal-pad word layout (per rank), uint32 indices */
enum {
  OMLDM_PAD_ENTRY = 0,     /* [0,16)   entry-barrier arrivals, one word per peer    */
  OMLDM_PAD_EXIT = 16,     /* [16,32)  exit-barrier arrivals                        */
  OMLDM_PAD_CLOCK = 32,    /* [32,48)  SSP clocks                                   */
  OMLDM_PAD_VOTE = 48,     /* [48,64)  GM / FGM round votes                         */
  OMLDM_PAD_PHI = 64,      /* [64,80)  FGM phi values (float bits), hub = rank 0    */
  OMLDM_PAD_FGM = 80,      /* [80,96)  FGM: counter, subround, theta, round id, announce step, terminate */
  OMLDM_PAD_STATUS = 96,   /* [96,100) status bits: 1 barrier, 2 SSP, 4 FGM timeouts  */
  OMLDM_PAD_HEART = 100,   /* [100,116) heartbeat counters, one word per peer (omldm_heartbeat) */
  OMLDM_PAD_QUERY = 128,   /* [128,256) query scalars, 8 per peer                   */
  OMLDM_PAD_QENTRY = 256,  /* [256,272) entry arrivals of the query / termination barrier (own words and own
                              sequence counter: an FGM round or an all-reduce never shares a flag with a query) */
  OMLDM_PAD_QEXIT = 272,   /* [272,288) exit arrivals of the query / termination barrier */
  OMLDM_PAD_GSTEP = 288,   /* [288,304) reserved */
  OMLDM_PAD_GENTRY = 304,  /* [304,320) sharded-Gram gather: entry arrivals */
  OMLDM_PAD_GEXIT = 320,   /* [320,336) sharded-Gram gather: exit arrivals  */
  OMLDM_PAD_WORDS = 512
};

/* Synchronous: every shard owner reduces all peers' buffers and writes the result back to all
 * peers. mode 0: buf <- scale * sum(buf_i); e <- buf (if e != NULL).
 * mode 1 (additive statistics): buf holds a delta; e += scale * sum(buf_i); buf <- 0. */
int omldm_sync_allreduce(const omldm_comm* comm_host, float* e_local, size_t n, float scale, int mode,
                         uint32_t epoch, int blocks, void* stream);
/* gemm_outer whose epilogue IS the reduce-scatter (ROADMAP item 0; no counterpart in the reference): the rows
 * [p * rows_per_owner, (p + 1) * rows_per_owner) of C (+)= A^T B are TMA-reduce-added into C_peers[p] (host array of
 * n_peers peer-mapped base pointers, all with pitch ldc), so the transfer to the shard owners overlaps the MMAs tile
 * by tile.  rows_per_owner: multiple of 32.  Pair with omldm_sync_scatter_gather (explicit step). */
int omldm_gemm_outer_scatter(const float* A, int lda, const float* B, int ldb, float* const* C_peers, int n_peers,
                             int rows_per_owner, int ldc, int M, int N, int K, int ksplit, int flags, void* stream);
/* Synchronous for additive statistics with the reduce-scatter moved into the producer (no counterpart in the
 * reference).  The symmetric buffers hold THREE rotating copies of the n-float delta (bufs[p] + r * n).
 * push  : adds the local delta (then cleared) shard by shard into the shard owners' copy `step mod 3` with
 *         system-scope reductions; needs no barrier.  Stand-in for a producer that reduces into the owners itself.
 * gather: one flag barrier, e += scale * (every owner's reduced shard), owner clears its shard of copy
 *         (step + 2) mod 3; no exit barrier.  Call push then gather on the same stream, once per step, on every rank.
 * step / epoch = 0: device managed (replayable from a CUDA graph). */
int omldm_sync_scatter_push(const omldm_comm* comm_host, float* delta_local, size_t n, uint32_t step, int blocks,
                            void* stream);
int omldm_sync_scatter_gather(const omldm_comm* comm_host, float* e_local, size_t n, float scale, uint32_t epoch,
                              uint32_t step, int blocks, void* stream);
/* Sharded-hub Synchronous protocol for the ORR Gram statistics (worker -> hub push network/FlinkNetwork.scala:260-261;
 * the hub -> worker broadcast :282-291 becomes a pull at weight refreshes).  GPU p is the hub shard of rows
 * [p * rows_per_owner, (p + 1) * rows_per_owner).  hubs[p] = k slot matrices (ld x ld each, slot r written by worker r
 * only) holding every worker's RUNNING SUM of those rows; bufs[p] = worker p's copy of the global matrix.
 * push  : copies the upper 32 x 32 blocks of the worker's running sum `acc_local` into its slot at every owner (plain
 *         peer stores).  Never waits.  The fp16 Gram kernel (omldm_gram_f16 with peers) does this itself.
 * gather: entry rendezvous (every worker's stores of the current step have landed), each owner sums its k slots and
 *         stores the result into every worker's global matrix, exit rendezvous.  Collective. */
int omldm_sync_gram_push(const omldm_comm* comm_host, const float* acc_local, int ld, int n, int rows_per_owner,
                         int blocks, void* stream);
int omldm_sync_gram_gather(const omldm_comm* comm_host, int ld, int n, int rows_per_owner, uint32_t epoch, int blocks,
                           void* stream);
/* SingleLearner forwarding (mlworkers.ForwardingWorker -> CentralizedMLServer, MLNodeGenerator.scala:27,56): every
 * rank stores its n-float mini-batch into rank 0's symmetric buffer (2 * world * n floats: slot seq & 1, position =
 * rank); rank 0 returns once all world batches of `seq` have arrived and trains on them.  seq = 1, 2, ... (same on
 * every rank); senders may run two steps ahead of rank 0. */
int omldm_sync_forward(const omldm_comm* comm_host, const float* z_local, size_t n, uint32_t seq, int blocks,
                       void* stream);
/* Asynchronous push/pull of one worker. mode 0: hub += scale*(buf - e); buf = e = hub.
 * mode 1 (additive delta): hub += scale*buf; e = hub; buf = 0.  No barrier. */
int omldm_sync_async(const omldm_comm* comm_host, float* e_local, size_t n, float scale, int mode,
                     int blocks, void* stream);
/* SSP: async push/pull, then clock++ and wait until min(clock) >= clock - staleness. */
int omldm_sync_ssp(const omldm_comm* comm_host, float* e_local, size_t n, float scale, int mode,
                   uint32_t clock, int staleness, int blocks, void* stream);
/* EASGD elastic step against the hub shards. */
int omldm_sync_easgd(const omldm_comm* comm_host, size_t n, float alpha, int blocks, void* stream);
/* GM: ||buf - e||^2 > threshold on any rank (vote carried by the entry barrier) => allreduce. */
int omldm_sync_gm(const omldm_comm* comm_host, float* e_local, size_t n, float scale, int mode,
                  float threshold, uint32_t epoch, int blocks, void* stream);
/* FGM monitor step: phi = ||buf - e|| - sqrt(threshold), counter increments to the hub,
 * sub-round / round bookkeeping, and the round-end allreduce, all inside one kernel.
 * `state_local` : float[8] private FGM state of this rank (zero-initialised); `step` counts from 1.
 * A round the hub announces in its step s is executed by a rank in its first step whose number is not s, so
 * ranks that advance in lock-step take the round-end branch in the same call and everybody else at first sight. */
int omldm_sync_fgm(const omldm_comm* comm_host, float* e_local, size_t n, float scale, int mode,
                   float threshold, float eps_psi, uint32_t step, float* state_local, int blocks,
                   void* stream);
/* End of stream.  The round-end all-reduce waits for every rank inside the kernel, so a rank must not simply stop
 * calling while peers may still announce a round.  Instead it keeps calling this entry point with the flush
 * generation `drain` (1, 2, ... per flush, same on all ranks): the step marks the rank as draining at the hub;
 * once the hub sees that mark from everybody and no round is open it writes `drain` into word OMLDM_PAD_FGM + 5
 * of every pad.  Call, wait for the stream, read that word of the own pad, repeat until it equals `drain`. */
int omldm_sync_fgm_drain(const omldm_comm* comm_host, float* e_local, size_t n, float scale, int mode,
                         float threshold, float eps_psi, uint32_t step, float* state_local, uint32_t drain,
                         int blocks, void* stream);
/* Liveness (no counterpart in the reference, which leaves failure detection to Flink): stores `beat` into word
 * OMLDM_PAD_HEART + rank of EVERY peer's pad.  Never waits.  A host watchdog calls it periodically on a side stream
 * and reads its own pad: a peer whose word stops advancing while the others move is stalled or gone. */
int omldm_heartbeat(const omldm_comm* comm_host, uint32_t beat, void* stream);
/* device barrier over the signal pads (termination probe, utils/statistics/StatisticsOperator.scala:135-142) */
int omldm_device_barrier(const omldm_comm* comm_host, uint32_t epoch, void* stream);
/* query reduce: out8 = sum over ranks of vals8 (utils/ResponseConstructor.scala:43-52).  Eight fp32 words per
 * rank; integer counts travel as 20-bit chunks so that their sum over up to 16 ranks stays exact in fp32. */
int omldm_query_reduce(const omldm_comm* comm_host, const float* vals8, float* out8, uint32_t epoch,
                       void* stream);

/* Sparse records against a dense model (BASELINE config 5; no counterpart in the reference, whose MultiClassPA
 * works on dense Breeze vectors out of tree).  A sparse row has pitch ld floats (multiple of 32, >= 2 nnz + 2):
 * [int32 column index bits x nnz | value x nnz | valid (1.0; 0 = empty row) | target | 0-pad], i.e. a staged row
 * with the pseudo feature count 2 nnz, so omldm_holdout_swap applies unchanged.
 * omldm_sparse_rows      : ELL slot (idx[B nnz], val[B nnz], target[B]) -> sparse rows.
 * omldm_mcpa_sparse_step : mini-batch MultiClassPA (2 <= C <= 32): per record scores by gathering W[c, idx],
 *   rival / hinge loss / step size and the statistics of omldm_mcpa_coef, then W[y, idx] += s tau v,
 *   W[rival, idx] -= s tau v (+ bias column d) with red.add; s = 1 / (#active records of this call) when
 *   mean_update (active_before = copy of stats[3] taken before the call).  coef: B x 4 floats scratch, 16-byte
 *   aligned.  Duplicate column indices accumulate in the scores / update; the step size uses sum v^2. */
int omldm_sparse_rows(const int* idx, const float* val, const float* targets, int B, int nnz, float* rows, int ld,
                      void* stream);
int omldm_mcpa_sparse_step(const float* rows, int ld, int B, int nnz, float* W, int ldw, int d, int C, int variant,
                           float Creg, float* coef, double* stats, const double* active_before, int mean_update,
                           void* stream);
/* Same step with a scratch of omldm_mcpa_sparse_wt_floats(d) floats (16-byte aligned): for C <= 16 the scores are
 * gathered from a class-minor copy Wt[col][16] rebuilt in front of the step (one 64-byte line per non-zero instead of
 * one 32-byte sector per class); wt == NULL or C > 16 = the plain step. */
size_t omldm_mcpa_sparse_wt_floats(int d);
int omldm_mcpa_sparse_step_wt(const float* rows, int ld, int B, int nnz, float* W, int ldw, int d, int C, int variant,
                              float Creg, float* coef, double* stats, const double* active_before, int mean_update,
                              float* wt, void* stream);

/* ------------------------------------------------------------------ neural network glue (nn.cu)
 * Element-wise steps between the GEMMs of the small MLP learner "NN" (whitelist PipelineMap.scala:68).
 * hidden_act : H[:, :od] <- relu, H[:, od] <- 1 (bias input of the next layer).
 * out_grad   : output layer (B x ldo, od outputs) + staged rows Z -> gradient rows G (B x ldg, columns >= od and
 *              invalid rows cleared), stats[0] += loss, stats[2] += #valid, *n_valid += #valid.
 *              classes > 1: soft-max cross entropy, target = class id; else squared error on output 0.
 * relu_mask  : g[r, c] <- 0 where act[r, c] <= 0 (first n columns).
 * sgd_step   : W <- W - lr / max(*n_valid, 1) * gW over n floats. */
int omldm_nn_hidden_act(float* H, int ldh, int B, int od, void* stream);
int omldm_nn_out_grad(const float* out, int ldo, const float* Z, int ld, int d, int B, int od, int classes, float* G,
                      int ldg, double* stats, float* n_valid, void* stream);
int omldm_nn_relu_mask(float* g, int ldg, const float* act, int lda, int B, int n, void* stream);
int omldm_nn_sgd_step(float* W, const float* gW, size_t n, float lr, const float* n_valid, void* stream);

/* Ridge refresh without a factorisation (no counterpart in the reference): solve (A + lam I) x = b by
 * Jacobi-preconditioned conjugate gradients started from the content of x.  A: n x n, symmetric, FULL storage,
 * pitch ld floats; b: n floats with stride b_stride (a column of the same matrix is fine); n <= 4096.
 * work: omldm_ridge_cg_work_floats(n) floats; scalars: omldm_ridge_cg_scalar_bytes(max_iter) bytes, 32-byte
 * aligned; info <- (iterations done, ||r|| / ||b||).  One cooperative launch of `blocks` CTAs (0 = 32),
 * stops at ||r|| <= tol ||b||, after max_iter (<= 128) iterations, or when p.Ap <= 0. */
size_t omldm_ridge_cg_work_floats(int n);
size_t omldm_ridge_cg_scalar_bytes(int max_iter);
int omldm_ridge_cg(const float* A, int ld, int n, float lam, const float* b, int b_stride, float* x, float* work,
                   void* scalars, int max_iter, float tol, float* info, int blocks, void* stream);

/* Direct ridge solve (no counterpart in the reference; replaces the library potrf / potrs pair): blocked Cholesky of
 * A + lam I with the right-hand side carried as an extra row, one cooperative launch.  W: (n + 1) x (n + 1) leading
 * block of a FULL symmetric matrix, pitch ld floats, rows / columns 0..n-1 = A, row n = b^T; destroyed.  x <- solution
 * (n floats, must not alias W); info[0] <- 1 when a pivot was not positive (the caller falls back to fp64), else 0.
 * blocks 0 = 64 CTAs.  Query-time model read, operators/spoke/FlinkSpoke.scala:163. */
int omldm_ridge_chol(float* W, int ld, int n, float lam, float* x, float* info, int blocks, void* stream);

/* CUDA IPC plumbing for the symmetric arena (one dedicated, zero-filled cudaMalloc per region) */
int omldm_dev_malloc(size_t bytes, void** out);
int omldm_dev_free(void* p);
int omldm_ipc_export(void* dev_ptr, void* handle64_host);
int omldm_ipc_import(const void* handle64_host, void** dev_ptr_out);
int omldm_ipc_close(void* dev_ptr);
int omldm_enable_peer_access(int peer_device);

#ifdef __cplusplus
}
#endif
#endif /* OMLDM_KERNELS_H_ */
