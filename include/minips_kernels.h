/*
 * minips_kernels.h — C ABI of libminips_kernels.so (hand-written sm_100a CUDA).
 *
 * Every entry point takes raw device pointers and a cudaStream_t (as void*), enqueues
 * work on that stream and returns a cudaError_t-compatible int (0 = success) unless
 * stated otherwise.  Python binds these through ctypes (minips_b200/ops/_lib.py); C++
 * callers include this header directly.
 *
 * "Reference" citations are file:line in the MiniPs reference tree and name the
 * behaviour an entry point reproduces (the reference itself has no device code).
 */
#ifndef MINIPS_KERNELS_H_
#define MINIPS_KERNELS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPS_MAX_SHARDS 64   /* server shards per table (8 GPUs x up to 8 servers per node) */
#define MPS_MAX_RANKS 16    /* GPUs (processes) in one NVLink domain */
#define MPS_MAX_WORKERS 256 /* user workers per task */

/* Signal-pad layout (int32 slots) inside each rank's symmetric pad, one pad per table. */
#define MPS_PAD_PROGRESS 0                                   /* [MPS_MAX_WORKERS] worker clocks   */
#define MPS_PAD_PULLED (MPS_PAD_PROGRESS + MPS_MAX_WORKERS)  /* [MPS_MAX_WORKERS] "Get of clock p done" = p+1 */
#define MPS_PAD_FOLDED (MPS_PAD_PULLED + MPS_MAX_WORKERS)    /* [MPS_MAX_SHARDS]  BSP fold clocks */
#define MPS_PAD_BARRIER (MPS_PAD_FOLDED + MPS_MAX_SHARDS)    /* [MPS_MAX_RANKS]   barrier epochs  */
#define MPS_PAD_ERROR (MPS_PAD_BARRIER + MPS_MAX_RANKS)      /* [1] sticky error (timeout) flag   */
#define MPS_PAD_MINCLOCK (MPS_PAD_ERROR + 1)                 /* [1] last min clock seen (debug)   */
#define MPS_PAD_LOCALCTR (MPS_PAD_MINCLOCK + 1)              /* [4] local CTA-completion counters */
/* LR exchange kernel: [2 sets][ready | arrived][MPS_MAX_RANKS] = CTAs of rank r that are done */
#define MPS_PAD_XCNT (MPS_PAD_LOCALCTR + 4)
/* bucketed sparse LR step: [2 sets][request | reply | delta | applied][MPS_MAX_RANKS] */
#define MPS_PAD_BCNT (MPS_PAD_XCNT + 4 * MPS_MAX_RANKS)
#define MPS_PAD_INTS 1024                                    /* pad size in int32 (4 KiB)         */

enum { MPS_MODEL_SSP = 0, MPS_MODEL_BSP = 1, MPS_MODEL_ASP = 2 }; /* driver/engine.hpp:25-27 */
enum { MPS_F32 = 0, MPS_F64 = 1 };

/*
 * Range-partitioned view of one table as seen from ONE device: the device analogue of
 * RangePartitionManager (base/range_partition_manager.hpp:16-66).  weights[s] is the
 * address, valid on the calling device, of shard s's first value (local memory or a
 * peer mapping obtained through the symmetric heap).  Shard s owns keys
 * [begin[s], begin[s+1]).  staging[s] is the BSP add buffer (server/consistency/
 * bsp_model.cpp:34-43 buffers adds until the clock edge); NULL under SSP/ASP where adds
 * land in place (server/consistency/ssp_model.cpp:54-56).
 */
typedef struct MpsTableView {
  void* weights[MPS_MAX_SHARDS];
  void* staging[MPS_MAX_SHARDS];
  uint32_t begin[MPS_MAX_SHARDS + 1];
  int32_t num_shards;
  uint32_t uniform_width; /* >0: every shard but the last has this many keys (owner = key / width) */
  int32_t vals_per_key;   /* values_ratio of base/range_partition_manager.hpp:47 (k values per key) */
  int32_t dtype;          /* MPS_F32 / MPS_F64 */
} MpsTableView;

/*
 * Device-side progress tracker + consistency gate: ProgressTracker
 * (server/util/progress_tracker.cpp:46-104) and the BSP/SSP/ASP get-gates
 * (bsp_model.cpp:45-56, ssp_model.cpp:58-85) as flags in peer memory.  pad[r] is rank
 * r's signal pad for this table, mapped on the calling device.
 */
typedef struct MpsClockView {
  int32_t* pad[MPS_MAX_RANKS];
  int32_t num_ranks;
  int32_t my_rank;
  int32_t num_workers;  /* workers registered in the tracker (ResetWorker)               */
  int32_t num_shards;   /* shards of the table                                           */
  int32_t own_first;    /* shards owned by this rank: [own_first, own_first + own_count) */
  int32_t own_count;
  int32_t model;        /* MPS_MODEL_*                                                   */
  int32_t staleness;    /* SSP staleness                                                 */
  int64_t timeout_cycles; /* spin-wait budget (SM clocks); on expiry the pad error flag is set */
} MpsClockView;

/* ---- runtime / symmetric heap (csrc/kernels/heap.cu) ------------------------------- */
int mps_device_count(void);
int mps_set_device(int dev);
int mps_sm_count(int dev);
int mps_heap_alloc(size_t bytes, void** out);           /* cudaMalloc + zero fill           */
int mps_heap_free(void* p);
int mps_ipc_get_handle(void* p, unsigned char out[64]); /* cudaIpcGetMemHandle              */
int mps_ipc_open_handle(const unsigned char h[64], void** out);
int mps_ipc_close_handle(void* p);
int mps_enable_peer(int peer_dev);                      /* same-process multi-device mapping */
int mps_memset32_async(void* p, int value, size_t n_ints, void* stream);
const char* mps_error_string(int err);

/* ---- generic KVClientTable ops (csrc/kernels/push_pull.cu) ------------------------- */
/* Get (worker/kv_client_table.hpp:175-195): vals[i*k + j] = shard(owner(keys[i]))[...].
 * keys sorted ascending is NOT required on device (routing is per key); out-of-range
 * keys read as 0 / are dropped on push (range_partition_manager.hpp:24-41 drops them).
 * If clk != NULL the kernel first gates on the consistency model for `worker` at clock
 * `progress` (spins on the local pad; no host round trip). */
int mps_pull(const MpsTableView* view, const MpsClockView* clk, int worker, int progress,
             const uint32_t* keys, void* vals, int64_t n, void* stream);
/* Add (kv_client_table.hpp:168-173): shard[...] += vals, red.add at the owner's L2. */
int mps_push(const MpsTableView* view, const uint32_t* keys, const void* vals, int64_t n,
             void* stream);
/* Dense full-table pull / push (k-means centroids, LR eval): contiguous slabs moved with
 * 16-byte peer loads / vector reds.  dst/src hold begin[num_shards]*vals_per_key values. */
int mps_pull_dense(const MpsTableView* view, const MpsClockView* clk, int worker, int progress,
                   void* dst, void* stream);
int mps_push_dense(const MpsTableView* view, void* src, int zero_src, int rot_shard, void* stream);

/* ---- clock / fold / barrier (csrc/kernels/clock.cu) -------------------------------- */
/* Clock (kv_client_table.hpp:123-134 + progress_tracker.cpp:46-72): publishes
 * progress[worker] = new_clock on every rank's pad with release.sys semantics (all prior
 * pushes of this device are visible first).  Under BSP, if this rank owns shards and
 * `fold` != 0, it then waits until every worker reached new_clock, folds staging into
 * the weights of its shards and publishes folded[s] = new_clock (bsp_model.cpp:14-32). */
int mps_clock(const MpsTableView* view, const MpsClockView* clk, int worker, int new_clock,
              int fold, void* stream);
/* Clock of two tables of the same worker in one launch (k-means: centres and counts). */
int mps_clock2(const MpsTableView* view0, const MpsClockView* clk0, const MpsTableView* view1,
               const MpsClockView* clk1, int worker, int new_clock, int fold, void* stream);
/* Resident fold server for a rank that owns BSP shards of the table but runs no worker of the task: one
 * persistent launch applies every clock edge after start_clock (wait for all workers, fold staging into
 * the weights, publish folded) until the host writes a non-zero value to *done_flag (pinned, device-
 * readable host memory) and no edge is pending.  bsp_model.cpp:14-32 without a host round trip. */
int mps_fold_server(const MpsTableView* view, const MpsClockView* clk, int start_clock, const int* done_flag,
                    void* stream);
/* Device-side barrier over the pads (comm/mailbox.cpp:362-396 semantics). */
int mps_barrier(const MpsClockView* clk, int epoch, void* stream);
/* Injected straggler (lr_example.cpp:347-353 sleeps the worker thread): stalls `stream` for ns
 * nanoseconds on the device, so the worker's next Clock is late for every peer. */
int mps_straggle(long long ns, void* stream);
/* ResetWorker (bsp_model.cpp:70-83): zero progress/folded on the LOCAL pad. */
int mps_reset_pad(const MpsClockView* clk, int start_clock, void* stream);

/* ---- sparse logistic regression (csrc/kernels/lr.cu) ------------------------------- */
typedef struct MpsCsr {
  const int64_t* row_ptr; /* [n_rows + 1] */
  const uint32_t* col;    /* [nnz] 0-based feature ids (lib/parser.hpp:33) */
  const float* val;       /* [nnz] */
  const float* label;     /* [n_rows] in {0,1} (lr_example.cpp:296-297 maps y<0 to 0) */
  int64_t n_rows;
  /* format = MPS_CSR_PACKED: the loader's compact host->device layout, read in place by
   * mps_lr_train_xchg (every other entry point wants MPS_CSR_PLAIN, the zero default):
   * row_len [n_rows] u8; coarse [n_rows/8 + 1] = offset (in non-zeros) of rows 0, 8, 16, ...;
   * packed_ids = little-endian bit stream, 4-byte aligned with a word of slack, id_bits bits
   * per feature id, the first one at bit bit_off; label_u8 [n_rows]; val as above. */
  int32_t format;
  int32_t id_bits;
  int32_t bit_off;
  int32_t reserved;
  const uint8_t* row_len;
  const int64_t* coarse;
  const void* packed_ids;
  const uint8_t* label_u8;
} MpsCsr;
enum { MPS_CSR_PLAIN = 0, MPS_CSR_PACKED = 1 };

/* One LR iteration, sparse regime: Get + forward SpMV + sigmoid + gradient + Add in ONE
 * kernel (lr_example.cpp:260,293-313): each warp walks a CSR row, loads the row's weights
 * straight from the owning shards (peer ld), reduces the logit, and issues the per-nnz
 * red.add of alpha*x*(y-pred) into the owner's staging/weights.  Rows are
 * [row_begin, row_begin + batch) modulo n_rows (BatchDataSampler wrap-around,
 * lib/batch_data_sampler.cpp:50-71).  loss_out[0] += sum logloss, loss_out[1] += #correct. */
int mps_lr_step_sparse(const MpsTableView* view, const MpsClockView* clk, int worker, int progress,
                       const MpsCsr* csr, int64_t row_begin, int64_t batch, double alpha,
                       float* loss_out, void* stream);

/* The same iteration as a request / reply / delta message flow between the GPUs (csrc/kernels/
 * lr_bucket.cu; the shape of the reference's kGet / kAdd traffic, worker/kv_client_table.hpp:168-195):
 * feature ids bucketed per owner, buckets shipped as contiguous peer stores, owners gather / apply in
 * their own HBM.  NVLink carries 3 x 4 bytes (fp32) per non-zero in long bursts instead of two
 * scattered small packets.  `bucket` describes the scratch: base[r] = rank r's symmetric scratch
 * (inbox_keys [R][cap] u32 | inbox_delta [R][cap] T | inbox_meta [R][2] i64 | reply [cap] T at the
 * given byte offsets), plus local buffers send_keys / slot [cap] u32, delta_send [cap] T, hist
 * [MPS_MAX_RANKS] u64; cap >= non-zeros of any batch; seq = number of bucketed launches on this table
 * so far (the same on every rank).  Collective: every rank runs exactly one worker through it, step by
 * step, with the same grid; batches must not wrap.  The Clock is part of the kernel: when it ends, every
 * rank's pad says that every worker reached progress + 1 and every shard is folded there (no mps_clock
 * afterwards).  gated != 0: wait for the table's Get gate of `progress` first (needed only when the
 * previous clock of the table was published by the generic path). */
typedef struct MpsBucket {
  void* base[MPS_MAX_RANKS];
  int64_t off_inbox_keys, off_inbox_delta, off_inbox_meta, off_reply;
  int64_t cap;
  uint32_t* send_keys;
  uint32_t* slot;
  void* delta_send;
  unsigned long long* hist;
  unsigned long long* timeline; /* NULL or 8 slots: %globaltimer at the phase boundaries (CTA 0) */
  int32_t seq;
  int32_t reserved;
} MpsBucket;
int mps_lr_step_bucketed(const MpsTableView* view, const MpsClockView* clk, int gated, int progress,
                         const MpsCsr* csr, int64_t row_begin, int64_t batch, double alpha,
                         const MpsBucket* bucket, float* loss_out, int avg_nnz, void* stream);

/* Dense regime, persistent cooperative kernel running `n_iters` whole iterations
 * (Get -> SpMV -> grad -> Add -> Clock [-> BSP fold]) without returning to the host:
 * slab-pull of all shards into `replica`, row pass accumulating into `gacc` (both scratch,
 * begin[num_shards] floats, gacc zero on entry and on exit), slab-push with vector reds,
 * then the clock / fold protocol of mps_clock.  batch_starts[i] is the first row of
 * iteration i; loss_out[2*i], loss_out[2*i+1] as above.  start_clock is the worker's clock
 * before the first iteration. */
int mps_lr_train_dense(const MpsTableView* view, const MpsClockView* clk, int worker,
                       int start_clock, const MpsCsr* csr, const int64_t* batch_starts,
                       int n_iters, int64_t batch, double alpha, void* replica, void* gacc,
                       float* loss_out, int max_ctas, void* stream);

/* mps_lr_train_dense plus observability / tuning knobs: `timeline` (NULL or 8 uint64 per
 * iteration) receives %globaltimer at the phase boundaries gate|pull|sync1|rows+sync2|
 * pulled-wait|push (the per-phase timers SURVEY 5.1 asks for); avg_nnz picks the lanes per
 * row (8/16/32; 0 = 40 nnz default).
 * Both take fp32 or fp64 tables (view->dtype; replica / gacc are of the table's type) — the
 * reference runs every table as Val = double (worker/kv_client_table.hpp:98-99).  Feature ids
 * outside the table's key range are skipped (range_partition_manager.hpp:24-41 drops them).
 * One dense-training worker per GPU and table: the kernel is a full-occupancy cooperative
 * launch that spins on its peers' clocks, so two of them cannot share a device. */
int mps_lr_train_dense_ex(const MpsTableView* view, const MpsClockView* clk, int worker,
                          int start_clock, const MpsCsr* csr, const int64_t* batch_starts,
                          int n_iters, int64_t batch, double alpha, void* replica, void* gacc,
                          float* loss_out, unsigned long long* timeline, int avg_nnz, int max_ctas,
                          void* stream);

/* BSP iterations as one exchange per step — reduce at the owner, owner broadcasts
 * (csrc/kernels/lr_xchg.cu; semantics of server/consistency/bsp_model.cpp:14-43: adds of a
 * clock become visible only after every worker clocked).  Every rank keeps a full weight
 * replica and two gradient accumulators in a symmetric buffer; replica[r] / gacc[r] address
 * rank r's copies from this device, replica_mc / gacc_mc are the NVLS multicast addresses of
 * the same buffers (NULL: plain peer loads / stores instead of multimem.ld_reduce /
 * multimem.st).  gacc_stride = elements between two of a rank's accumulators.
 * replica_valid != 0 promises that the replica already holds the weights of start_clock (the
 * previous launch on this table ended there and nothing else wrote the table since); otherwise
 * the kernel waits for the Get gate and copies every shard once.
 * flags: MPS_XCHG_REDUCE_P2P reduces with peer loads even when gacc_mc is set (fp64 has only a
 * scalar multimem.ld_reduce, which loses to 16-byte peer loads on few GPUs); MPS_XCHG_BCAST_P2P
 * broadcasts with peer stores even when replica_mc is set.
 * Collective: every rank of the table runs exactly one worker through this call with the same
 * n_iters and the same number of CTAs (same GPU model, same max_ctas); BSP tables with one value
 * per key; fp32 or fp64.  timeline slots per iteration: start | broadcast-wait | zero+rows |
 * clock+owner-wait | reduce+broadcast | signal. */
#define MPS_XCHG_REDUCE_P2P 1
#define MPS_XCHG_BCAST_P2P 2 /* broadcast with N-1 peer stores even when replica_mc is set */
typedef struct MpsXchg {
  void* replica[MPS_MAX_RANKS];
  void* gacc[MPS_MAX_RANKS];
  void* replica_mc;
  void* gacc_mc;
  int64_t gacc_stride;
  int32_t replica_valid;
  int32_t flags;
  int32_t seq;      /* number of exchange launches on this table so far (the same on every rank) */
  int32_t reserved;
} MpsXchg;
int mps_lr_train_xchg(const MpsTableView* view, const MpsClockView* clk, int worker, int start_clock,
                      const MpsCsr* csr, const int64_t* batch_starts, int n_iters, int64_t batch,
                      double alpha, const MpsXchg* xchg, float* loss_out, unsigned long long* timeline,
                      int avg_nnz, int max_ctas, void* stream);

/* The same persistent training kernel as mps_lr_train_dense, for callers that cannot
 * allocate device memory.  (No counterpart in the reference: its workers allocate freely
 * on the host heap.)  Who needs it: an embedder whose own pool already owns all free HBM,
 * so that a new cudaMalloc would fail; a latency-critical loop that must not take the
 * device-wide synchronisation an allocation implies; a plain C caller with no device
 * allocator to borrow scratch from.  The replica and the gradient accumulator live in a
 * fixed array inside the library image instead: 2 x MPS_NOALLOC_MAX_KEYS floats (8 MiB) that
 * every GPU loading the library holds until the process exits.
 *   - Inputs and outputs (table views, CSR arrays, batch_starts, loss_out) are still the
 *     caller's; nothing is allocated or freed by this call, here or in any library it uses.
 *   - Tables with more than MPS_NOALLOC_MAX_KEYS keys are refused: the call prints the two
 *     sizes to stderr and returns cudaErrorInvalidValue without launching anything.
 *   - The first call on each device looks up that device's address of the array (the
 *     device must be current).  Make that first call outside any stream capture.
 *   - All calls on one GPU share the one array.  It keeps the previous call's contents; the
 *     part that has to start at zero (the accumulator) is cleared on `stream` before the
 *     launch.  Two calls that overlap on the same GPU (different streams, different tables)
 *     race on it — serialise them. */
#define MPS_NOALLOC_MAX_KEYS (1u << 20)
int mps_lr_train_dense_noalloc(const MpsTableView* view, const MpsClockView* clk, int worker,
                               int start_clock, const MpsCsr* csr, const int64_t* batch_starts,
                               int n_iters, int64_t batch, float alpha, float* loss_out, int max_ctas,
                               void* stream);

/* Accuracy / logloss of local rows against a local full copy of the weights
 * (lr_example.cpp:61-87): weights[k - key0] for keys [key0, key0 + n_keys), dtype MPS_F32 /
 * MPS_F64; feature ids outside that range are skipped.  out[0] += logloss, out[1] += #correct. */
int mps_lr_eval(const void* weights, int dtype, uint32_t key0, uint32_t n_keys, const MpsCsr* csr,
                int64_t row_begin, int64_t n, float* out, void* stream);

/* ---- sparse helpers (csrc/kernels/sparse_utils.cu) ----------------------------------- */
/* Sorted unique feature ids of a batch (lib/batch_data_sampler.cpp:50-71) and, optionally,
 * the slot of every non-zero in that key list.  scratch_u32: 2*ceil(num_dims/32) +
 * ceil(ceil(num_dims/32)/4096) + 2 words.  keys_out ascending; count_out[0] = #keys.
 * slots_out[j - nnz_base] for the batch's non-zeros j (non-wrapping batches). */
int mps_batch_unique_keys(const MpsCsr* csr, int64_t row_begin, int64_t batch, uint32_t num_dims,
                          uint32_t* scratch_u32, uint32_t* keys_out, uint32_t* count_out,
                          int32_t* slots_out, int64_t nnz_base, void* stream);
/* libsvm text -> CSR on device (lib/parser.hpp:17-47, io/line_input_format.hpp:43-121), three
 * passes so the caller can size the outputs between them:
 *  1 scan_lines : line_end[k] = offset of the k-th line terminator, counts_out[0] = #lines
 *                 (scratch_u32: ceil((nbytes+1)/4096)+2 words)
 *  2 index_lines: valid/nnz per line, row_of_line, nnz_prefix, row_ptr[row];
 *                 counts_out[1] = #rows, counts_out[2] = #nnz
 *  3 parse      : col (0-based, file is 1-based), val, label (optionally y<0 -> 0) */
int mps_libsvm_scan_lines(const char* text, int64_t nbytes, int64_t* line_end, uint32_t* scratch_u32,
                          uint32_t* counts_out, void* stream);
int mps_libsvm_index_lines(const char* text, const int64_t* line_end, int64_t n_lines, uint32_t* valid,
                           uint32_t* nnz, int64_t* row_of_line, int64_t* nnz_prefix, int64_t* row_ptr,
                           uint32_t* scratch_u32, uint32_t* counts_out, void* stream);
int mps_libsvm_parse(const char* text, const int64_t* line_end, int64_t n_lines,
                     const int64_t* row_of_line, const int64_t* nnz_prefix, uint32_t* col, float* val,
                     float* label, int map_negative_label_to_zero, void* stream);

/* ---- MapStorage on device (csrc/kernels/hash_table.cu) -------------------------------- */
/* Range-partitioned open-addressing hash shards (server/map_storage.hpp:13-26): keys[s] /
 * vals[s] address shard s's [capacity[s]] u32 keys (0xFFFFFFFF = empty) and f32 values, local
 * or peer-mapped; capacity is a power of two. */
typedef struct MpsHashView {
  void* keys[MPS_MAX_SHARDS];
  void* vals[MPS_MAX_SHARDS];
  uint32_t capacity[MPS_MAX_SHARDS];
  uint32_t begin[MPS_MAX_SHARDS + 1];
  int32_t num_shards;
} MpsHashView;
int mps_hash_clear(void* keys, void* vals, int64_t capacity, void* stream);
int mps_hash_add(const MpsHashView* view, const uint32_t* keys, const float* vals, int64_t n,
                 int* overflow_flag, void* stream);
int mps_hash_get(const MpsHashView* view, const MpsClockView* clk, int progress, const uint32_t* keys,
                 float* vals, int64_t n, void* stream);
int mps_hash_export(const void* keys, const void* vals, int64_t capacity, uint32_t* out_keys,
                    float* out_vals, int* count, void* stream);

/* Rebuild CSR arrays from the compact host->device batch format: row_len [rows] u8; packed_ids = a
 * little-endian bit stream (4-byte aligned, one word of slack after the end) holding nnz feature ids of
 * id_bits bits each, the first one starting at bit bit_off; label_u8 [rows] in {0,1}  ->  row_ptr
 * [rows+1] int64 (starting at 0), col [nnz] u32 (16-byte aligned), label [rows] f32.
 * scratch_u32: rows/4096 + 2 words. */
int mps_csr_unpack(const uint8_t* row_len, const void* packed_ids, int id_bits, int bit_off,
                   const uint8_t* label_u8, int64_t rows, int64_t nnz, int64_t* row_ptr, uint32_t* col,
                   float* label, uint32_t* scratch_u32, void* stream);

/* ---- mini-batch k-means (csrc/kernels/kmeans_tc.cu, kmeans.cu) ----------------------- */
/* Nearest-centre assignment (apps/kmeans/kmeans_helper.hpp:45-66) as a tcgen05/TMEM/TMA GEMM
 * with the arg-min fused into the TMEM epilogue and the centroid pull (kmeans.cpp:223) fused
 * into the B-operand fill.  x_bf16: [n, D] row-major bf16 points (D = 64 or 128); centres come
 * either from a range-partitioned fp32 table view of K*D keys (row-major, peer loads, optional
 * consistency gate clk/progress) or, when centers_view is NULL, from the local dense array
 * centers_f32.  K must be a multiple of 128 and D 64 or 128 (the tcgen05 tile shapes); any other problem
 * is run PADDED: K_valid real centres (rows [K_valid, K) of the operand are zero with norm 2^127, so
 * they are never chosen) and D_valid real coordinates per centre (the table / centers_f32 rows are
 * D_valid long; x carries zeros in columns [D_valid, D)); 0 = no padding.  Scratch: centers_bf16_scratch [K*D] bf16,
 * cnorm_scratch [K] f32, sync_ctr [1] int.  Outputs: assign [n] int32, dist [n] f32 =
 * ||c||^2 - 2 x.c of the chosen centre (add ||x||^2 for the squared distance); optional
 * centers_f32_out [K*D] receives the pulled fp32 centres. */
int mps_kmeans_assign_tc(const MpsTableView* centers_view, const MpsClockView* clk, int progress,
                         const float* centers_f32, const void* x_bf16, int64_t n, int K, int D, int K_valid,
                         int D_valid, void* centers_bf16_scratch, float* cnorm_scratch, float* centers_f32_out,
                         int* sync_ctr, int32_t* assign, float* dist, int max_ctas, void* stream);
/* The tcgen05 kernels bound every mbarrier wait; one that expires (a protocol bug, never seen in the
 * tests) raises a flag instead of trapping the context.  These read and clear it (device sync). */
int mps_kmeans_tc_error(void);
int mps_dense_lr_tc_error(void);
/* CUDA-core kernel of the same operation on the same bf16 inputs (differential reference). */
int mps_kmeans_assign_ref(const void* x_bf16, const void* centers_bf16, const float* cnorm, int64_t n,
                          int K, int D, int32_t* assign, float* dist, void* stream);
/* Per-cluster partial sums / counts of a batch (kmeans.cpp:247-263 in batched form) and its SSE
 * (kmeans_helper.hpp:236-266): sums [K*D] += x, counts [K] += 1, sse[0] += ||x - c||^2. */
int mps_kmeans_accumulate(const void* x_bf16, const int32_t* assign, const float* dist, int64_t n,
                          int K, int D, float* sums, float* counts, float* sse, void* stream);
/* The same accumulation without one atomic per point and lane: the batch's row indices are counting-
 * sorted by centre (histogram, scan, scatter) and every warp sums a run of 64 sorted rows in registers,
 * issuing atomics only where the centre changes.  sort_scratch: 3*K + 1 + n ints, its first K zero on
 * entry (left zero).  D must be 64 or 128 and K <= 4096, otherwise (or with sort_scratch NULL) this is
 * mps_kmeans_accumulate. */
int mps_kmeans_accumulate_sorted(const void* x_bf16, const int32_t* assign, const float* dist, int64_t n,
                                 int K, int D, float* sums, float* counts, float* sse, int* sort_scratch,
                                 void* stream);
/* Centroid normalisation: delta_c = alpha (sum - m c) / (n_old + m), delta_n = m. */
int mps_kmeans_delta(const float* centers, const float* counts_old, const float* sums,
                     const float* counts_batch, int K, int D, float alpha, float* delta_c,
                     float* delta_n, void* stream);
/* The same normalisation fused with the Get of the counts table and the Add of BOTH tables
 * (kmeans.cpp:224, 247-274): old counts are read from the owning shards under counts_clk's gate
 * (NULL = ungated), delta_c / delta_n go out as red.add to the owners (staging under BSP), and
 * sums / counts_batch are cleared for the next batch.  centers: the fp32 centres this batch was
 * assigned against (centers_f32_out of mps_kmeans_assign_tc).  f32 tables, D % 4 == 0. */
int mps_kmeans_delta_push(const MpsTableView* centers_view, const MpsTableView* counts_view,
                          const MpsClockView* counts_clk, int progress, const float* centers,
                          float* sums, float* counts_batch, int K, int D, int K_valid, int D_valid, float alpha,
                          void* stream);
/* One whole mini-batch iteration of a worker (kmeans.cpp:216-275) enqueued by a single host call:
 * assign (centre pull fused) -> accumulate (sorted,
     see above) -> delta + push of both tables -> Clock of both tables at
 * progress + 1 (fold: this worker is the last local one to clock, see mps_clock). */
int mps_kmeans_step(const MpsTableView* centers_view, const MpsClockView* centers_clk,
                    const MpsTableView* counts_view, const MpsClockView* counts_clk, int worker,
                    int progress, int fold, const void* x_bf16, int64_t n, int K, int D, int K_valid, int D_valid,
                    float alpha,
                    void* centers_bf16_scratch, float* cnorm_scratch, float* centers_f32, int* sync_ctr,
                    int32_t* assign, float* dist, float* sums, float* counts_batch, float* sse,
                    int* sort_scratch, void* stream);

/* ---- dense logistic regression on tensor cores (csrc/kernels/dense_lr_tc.cu) ----------- */
/* Forward of a dense batch (lr_example.cpp:293-305 for dense features): logits = X.w on
 * tcgen05/TMEM fed by TMA, sigmoid / logloss / coef = alpha (y - p) fused in the TMEM epilogue,
 * weight pull fused into the prologue (weights_view: fp32 table of D keys; or w_f32 local).
 * x_bf16 [n, D] row-major, D % 8 == 0.  Scratch: w16_scratch 16 * roundup(D, 64) bf16,
 * sync_ctr [1].  Outputs: coef [n], loss_out[0..1] += (logloss sum, #correct), optional
 * w_f32_out [D]. */
int mps_dense_lr_forward_tc(const MpsTableView* weights_view, const MpsClockView* clk, int progress,
                            const float* w_f32, const void* x_bf16, const float* label, int64_t n, int D,
                            float alpha, void* w16_scratch, float* w_f32_out, int* sync_ctr, float* coef,
                            float* loss_out, int max_ctas, void* stream);
/* Forward AND gradient in one kernel that reads X once (csrc/kernels/dense_lr_fused.cu): 64-row tiles
 * stay in shared memory between the logits product and the gradient product, which takes the same
 * blocks as an MN-major A operand, so the whole kernel's gradient accumulates in TMEM and leaves it once:
 * grad[d] += sum_r alpha (y_r - sigmoid(x_r . w)) x[r, d].  Same arguments as mps_dense_lr_forward_tc with
 * `grad` [D] in place of `coef`; needs D % 8 == 0 and roundup(D, 128) <= 1024 (cudaErrorInvalidValue
 * otherwise: use the two-kernel path). */
int mps_dense_lr_step_tc(const MpsTableView* weights_view, const MpsClockView* clk, int progress,
                         const float* w_f32, const void* x_bf16, const float* label, int64_t n, int D,
                         float alpha, void* w16_scratch, float* w_f32_out, int* sync_ctr, float* grad,
                         float* loss_out, int max_ctas, void* stream);
/* The same one-pass iteration with the two matrix-vector products on the CUDA cores: TMA streams 64-row
 * tiles into shared memory, weights and gradient accumulators live in registers for the whole kernel
 * (csrc/kernels/dense_lr_stream.cu).  1 Mi x 1024 on one B200: 0.64 ms per iteration, the tcgen05
 * variant above 0.66 ms, two passes 0.75 ms, PyTorch 0.73 ms.  Same limits on D. */
int mps_dense_lr_step_stream(const MpsTableView* weights_view, const MpsClockView* clk, int progress,
                             const float* w_f32, const void* x_bf16, const float* label, int64_t n, int D,
                             float alpha, float* w_f32_out, float* grad, float* loss_out, int max_ctas,
                             void* stream);
/* Backward: grad[d] += sum_r coef[r] * x[r, d] (lr_example.cpp:305-312), feeds mps_push_dense. */
/* The same iteration with NO tile resident: one warp owns one row from the load to the gradient, the
 * shared-memory ring is filled by linear bulk copies and is only bytes in flight
 * (csrc/kernels/dense_lr_rows.cu).  1 Mi x 1024 on one B200: 0.349 ms per iteration = 0.94 of the HBM copy
 * rate.  D % 8 == 0, D <= 1024, x 16-byte aligned. */
int mps_dense_lr_step_rows(const MpsTableView* weights_view, const MpsClockView* clk, int progress,
                           const float* w_f32, const void* x_bf16, const float* label, int64_t n, int D,
                           float alpha, float* w_f32_out, float* grad, float* loss_out, int max_ctas,
                           void* stream);
int mps_dense_lr_grad(const void* x_bf16, const float* coef, int64_t n, int D, float* grad, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MINIPS_KERNELS_H_ */
