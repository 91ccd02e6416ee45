/* cudasp.h -- C ABI of the B200-native cudasp_scan hot path (libcudasp_b200.so).
 *
 * This is the drop-in boundary.  It replaces the reference's internal FFI between the DuckDB
 * table function and its GPU pipeline:
 *     void* LaunchBatchScan(...); int RunBatchScanKernels(...); void FreeBatchScanState(void*);
 *     (reference: src/cudasp_extension.cu:21-52, implemented in src/cudasp_gpu.cu:445-1023)
 * Plain C types only; every function returns CUDASP_OK (0) or a negative error code and
 * records a message retrievable with cudasp_last_error().  Unlike the reference, which printf()s
 * and silently yields no rows on a GPU failure (src/cudasp_extension.cu:325,378), errors are
 * always reported to the caller.  There is no CPU fallback: without a CUDA device every entry
 * point that needs one fails with CUDASP_ERR_CUDA.
 *
 * Byte layouts are the reference's:
 *   scan_private_key  32 B little-endian integer          (src/cudasp_extension.cu:77-85)
 *   points            64 B = x little-endian || y little-endian, no prefix byte   (:56-75)
 *   outputs           int64 = first 8 bytes of an output's x-only key, big-endian, as signed
 *   out_offsets       CSR: row r owns outputs[out_offsets[r] .. out_offsets[r+1])   (rows+1 entries;
 *                     the reference passes offset+length arrays, src/cudasp_gpu.cu:29-41)
 * Row semantics: a row matches iff  int64(x(P0) >> 192)  or, for any label i,
 * int64(x(P0 + L_i) >> 192)  is in the row's outputs, where  P0 = t*G + B_spend,
 * t = int_BE(SHA256(TH||TH||ser(S)||00000000)),  S = scan_sk * tweak_key.
 */
#ifndef CUDASP_H
#define CUDASP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CUDASP_OK 0
#define CUDASP_ERR_ARG (-1)    /* bad argument (sizes, NULLs, label count ...) */
#define CUDASP_ERR_CUDA (-2)   /* CUDA runtime error or no device */
#define CUDASP_ERR_STATE (-3)  /* call not valid in this state (e.g. slot not acquired) */
#define CUDASP_ERR_KEY (-4)    /* scan key for which no valid ladder schedule exists */

/* Tweak-key formats.  XY64 is the reference's (64 B = x_LE || y_LE).  SEC1_33 is this build's extension:
 * 33-byte compressed SEC1 keys (02/03 || x big-endian), decompressed on the GPU with the reference's CPU recipe
 * (scripts/decompress_tweak_key.py:26-46); a malformed key (bad prefix, x >= p, x not on the curve) never matches.
 * In a staging slot compressed keys are packed at 33 bytes per row from the start of slot.tweak_keys. */
#define CUDASP_KEYS_XY64 0
#define CUDASP_KEYS_SEC1_33 1

#define CUDASP_MAX_LABELS 64u
#define CUDASP_MAX_BATCH_ROWS 10000000u /* reference: batch_size <= 10,000,000 (src/cudasp_extension.cu:485) */

typedef struct cudasp_query cudasp_query;

/* Number of visible CUDA devices (reference: cudaGetDeviceCount in CudaspScanState,
 * src/cudasp_extension.cu:129-132).  Returns 0 when there is none. */
int cudasp_device_count(void);

/* Last error message of a query handle, or of the calling thread when q == NULL. */
const char *cudasp_last_error(const cudasp_query *q);

/* Once-per-query setup.  Replaces the per-batch key marshalling of ProcessBatch/LaunchBatchScan
 * (src/cudasp_extension.cu:277-297, src/cudasp_gpu.cu:445-669) and Solver::initialize
 * (gECC gsv.h:1094-1122): recodes the fixed scan key (GLV + wNAF) and brings the FIRST listed device up
 * (CUDA context, the process-wide fixed-base table -- built on the device on first use --, one staging slot).
 * The other devices are brought up when a staging slot on them is first needed, so a query of a few rows
 * touches one GPU only; a device that fails to come up is reported by the acquire/scan call that needed it.
 *   devices == NULL  -> use devices 0..n_devices-1;  n_devices == 0 -> all visible devices.
 *   max_outputs_per_batch == 0 -> 4 * batch_rows (grown on demand by the synchronous call; a streaming
 *   producer cuts a batch early when the next row's outputs would not fit).
 * Staging slots are materialised on demand (one at creation) and are parked in a process-wide
 * cache by cudasp_query_destroy, so later queries do not pay the pinned/device allocations again. */
int cudasp_query_create(const uint8_t scan_private_key[32], const uint8_t spend_public_key[64],
                        const uint8_t *label_keys /* n_labels * 64 B */, uint32_t n_labels,
                        const int *devices, int n_devices, uint32_t batch_rows,
                        uint64_t max_outputs_per_batch, cudasp_query **out);

/* Same, with the number of staging slots per device chosen by the caller (0 = default 6; max 64).
 * A host with T producer threads (DuckDB workers, src/cudasp_extension.cu:516-532) wants about
 * T / n_devices + 2 so that no producer waits for a slot while batches are in flight. */
int cudasp_query_create_ex(const uint8_t scan_private_key[32], const uint8_t spend_public_key[64],
                           const uint8_t *label_keys, uint32_t n_labels, const int *devices, int n_devices,
                           uint32_t batch_rows, uint64_t max_outputs_per_batch, uint32_t slots_per_device,
                           cudasp_query **out);

/* Streaming multiprocessor count of a device (no context is created): hosts size their batches in whole waves of the
 * ladder kernel (SMs x 4 blocks x 128 rows). 0 on error. */
int cudasp_device_sm_count(int device);

/* Frees the staging slots parked by destroyed queries (see above). */
void cudasp_release_cached_memory(void);

/* Replaces FreeBatchScanState (src/cudasp_gpu.cu:969-1023).  Waits for in-flight batches. */
void cudasp_query_destroy(cudasp_query *q);

/* ---- synchronous batch: host buffers in, matching row ids out ------------------------------
 * The call the reference's ProcessBatch makes per batch (LaunchBatchScan + RunBatchScanKernels +
 * the flag scan, src/cudasp_extension.cu:308-392), as one function.  Thread-safe: concurrent
 * callers are spread over the query's devices.  rows may exceed batch_rows (split internally).
 * match_ids (capacity >= rows) receives the ascending indices of matching rows. */
int cudasp_scan_batch(cudasp_query *q, const uint8_t *tweak_keys /* rows * 64 B */,
                      const int64_t *outputs, const uint32_t *out_offsets /* rows + 1 */,
                      uint32_t rows, uint32_t *match_ids, uint32_t *n_matches);

/* Same with an explicit key format (CUDASP_KEYS_*); tweak_keys holds rows * 33 bytes for SEC1_33. */
int cudasp_scan_batch_fmt(cudasp_query *q, const uint8_t *tweak_keys, int key_format, const int64_t *outputs,
                          const uint32_t *out_offsets, uint32_t rows, uint32_t *match_ids, uint32_t *n_matches);

/* Same with the number of host staging threads chosen by the caller (0 = automatic: up to 4 per device, as many as there are
 * pieces and cores; cudasp_scan_batch / _fmt use the automatic choice).  The call cuts the table into pieces of at most two
 * ladder waves, ramping up at the start and down at the end; each staging thread copies pieces into pinned slots (devices in
 * turn) and submits them, so one call over a handle that owns N devices feeds all N from one host table. */
int cudasp_scan_batch_mt(cudasp_query *q, const uint8_t *tweak_keys, int key_format, const int64_t *outputs,
                         const uint32_t *out_offsets, uint32_t rows, uint32_t *match_ids, uint32_t *n_matches, int host_threads);

/* ---- streaming: pinned double-buffered staging, async H2D on per-device streams -------------
 * The host fills a pinned slot directly from its column vectors (what AccumulateInput does into
 * std::vectors, src/cudasp_extension.cu:154-224), submits it, and collects results later. */
typedef struct cudasp_slot {
  uint32_t slot_id;
  int32_t device;
  uint8_t *tweak_keys;    /* pinned, cap_rows * 64 B */
  int64_t *outputs;       /* pinned, cap_outputs */
  uint32_t *out_offsets;  /* pinned, cap_rows + 1 */
  uint32_t cap_rows;
  uint64_t cap_outputs;
} cudasp_slot;

typedef struct cudasp_result {
  uint64_t tag;              /* caller's tag from cudasp_submit_slot */
  uint32_t slot_id;
  uint32_t rows;
  uint32_t n_matches;
  uint32_t flags;            /* CUDASP_RESULT_* */
  const uint32_t *match_ids; /* pinned; ascending row indices within the slot; valid until release */
} cudasp_result;
/* match details were enabled and the batch produced more records than max_details_per_batch: match_ids are complete,
 * cudasp_result_details returns the records that fit */
#define CUDASP_RESULT_DETAILS_TRUNCATED 1u

/* Preconditions the library checks on submit: out_offsets[0] == 0, out_offsets non-decreasing,
 * out_offsets[rows] == n_outputs (the kernels index outputs[] with them).
 * Threading: every function below may be called concurrently on one handle.
 * Blocks until a staging slot is free (a slot is recycled by cudasp_release_result). */
int cudasp_acquire_slot(cudasp_query *q, cudasp_slot *slot);
/* Non-blocking variant: returns 1 and fills *slot, or 0 when every slot is busy. */
int cudasp_try_acquire_slot(cudasp_query *q, cudasp_slot *slot);
/* Enqueue H2D + kernels + D2H for `rows` rows / `n_outputs` outputs of an acquired slot. */
int cudasp_submit_slot(cudasp_query *q, const cudasp_slot *slot, uint32_t rows, uint64_t n_outputs, uint64_t tag);
int cudasp_submit_slot_fmt(cudasp_query *q, const cudasp_slot *slot, uint32_t rows, uint64_t n_outputs, uint64_t tag,
                           int key_format);
/* Returns 1 and fills *r when a finished batch is available (ANY finished batch, whatever the submit order: a batch
 * finished on one device is not held back by an older one still running on another), 0 when none is finished yet.
 * On a negative return after a batch was taken, r->tag / slot_id / rows are filled (n_matches = 0) so that the caller
 * can hand the slot back with cudasp_release_result. */
int cudasp_poll(cudasp_query *q, cudasp_result *r);
/* Same, but blocks (for the oldest in-flight batch) when none has finished; returns 0 only when nothing is in flight. */
int cudasp_wait(cudasp_query *q, cudasp_result *r);
/* Hand the slot of a consumed result back to the pool. */
int cudasp_release_result(cudasp_query *q, const cudasp_result *r);

/* ---- device-resident entry (benchmarks; hosts that already hold columns in HBM) -------------
 * All pointers are device pointers on `device` (which must be one of the query's devices);
 * `stream` is a cudaStream_t (0 = the query's own stream for that device).  Asynchronous.  Thread-safe: the scratch it
 * borrows from the device's staging slots is handed from user to user with an event, recorded under a per-device lock.
 * d_out_offsets must be zero-based and non-decreasing (not checked: it lives on the device).
 * d_match_ids: capacity >= rows; *d_n_matches is reset and then holds the count; ids unordered. */
int cudasp_scan_device(cudasp_query *q, int device, const uint8_t *d_tweak_keys, const int64_t *d_outputs,
                       const uint32_t *d_out_offsets, uint32_t rows, uint32_t *d_match_ids,
                       uint32_t *d_n_matches, void *stream);

/* Synchronous variant of cudasp_scan_device that times each kernel with CUDA events on the query's
 * own stream: stage_ms[7] = ladder, inv(Z1), hash+fixed-base, inv(Z2), base, inv(labels), labels
 * (stages that do not run report 0).  Used by bench.py for the per-kernel roofline. */
int cudasp_profile_stages(cudasp_query *q, int device, const uint8_t *d_tweak_keys, const int64_t *d_outputs,
                          const uint32_t *d_out_offsets, uint32_t rows, uint32_t *d_match_ids, uint32_t *d_n_matches,
                          float *stage_ms);

/* ---- match details: the full BIP-352 receive loop on the rows that matched (extension, opt-in) ----------------
 * The reference stops at "this row has an output of mine": k = 0 only, no word on which output, which label, or the
 * tweak needed to spend (/root/reference/src/cudasp_gpu.cu:278-282, src/cudasp_extension.cu:383-391).  With details
 * enabled, every matching row is followed through BIP-352's scanning loop on the GPU (one extra kernel over the
 * handful of matched rows): k = 0, 1, 2, ... while an output equals P_k or P_k + label_i (64-bit x prefixes, the
 * reference's data model), t_k range-checked (0 < t_k < n).  match_ids are NOT affected: they stay bit-exact with the
 * reference; details only add information about those rows. */
typedef struct cudasp_match_detail {
  uint32_t row;          /* row id within the batch (cudasp_scan_batch_details: within the call) */
  uint32_t k;            /* BIP-352 output counter */
  uint32_t output_index; /* position in the row's outputs list (NULL children not counted) */
  int32_t label_index;   /* -1: P_k itself; i >= 0: P_k + label_keys[i] */
  uint8_t t_k[32];       /* big-endian; the output's private key is b_spend + t_k (+ the label's scalar) mod n */
} cudasp_match_detail;

/* Call after create, before the first batch.  max_details_per_batch = 0 picks min(batch_rows, 65536). */
int cudasp_query_enable_details(cudasp_query *q, uint32_t max_details_per_batch);
/* Details of a result handed out by cudasp_poll / cudasp_wait, sorted by (row, k); valid until the result is released. */
int cudasp_result_details(cudasp_query *q, const cudasp_result *r, const cudasp_match_detail **details, uint32_t *n_details);
/* cudasp_scan_batch_fmt plus details (rows numbered over the whole call). */
int cudasp_scan_batch_details(cudasp_query *q, const uint8_t *tweak_keys, int key_format, const int64_t *outputs,
                              const uint32_t *out_offsets, uint32_t rows, uint32_t *match_ids, uint32_t *n_matches,
                              cudasp_match_detail *details, uint32_t details_capacity, uint32_t *n_details);

/* ---- introspection used by the parity tests and bench.py -----------------------------------
 * Per-row candidates instead of match flags: cands[r*(1+L) + 0] = base, [.. + 1 + i] = label i. */
int cudasp_scan_candidates(cudasp_query *q, const uint8_t *tweak_keys, uint32_t rows, int64_t *cands);
/* Same for either key format; valid[r] (optional) = 1 unless row r's compressed key is malformed. */
int cudasp_scan_candidates_fmt(cudasp_query *q, const uint8_t *tweak_keys, int key_format, uint32_t rows, int64_t *cands,
                               uint8_t *valid);
/* Per-row intermediates: shared secret S (64 B LE x||y), t (32 B big-endian), P0 (64 B LE). Any may be NULL. */
int cudasp_scan_intermediates(cudasp_query *q, const uint8_t *tweak_keys, uint32_t rows,
                              uint8_t *S_le, uint8_t *t_be, uint8_t *P0_le);
/* Field KATs on the device: op 0 mul, 1 sqr(a), 2 add, 3 sub, 4 inv(a), 5 normalize(a); n elements of 32 B LE. */
int cudasp_test_field(int device, int op, const uint8_t *a, const uint8_t *b, uint8_t *out, uint32_t n);
/* ---- allocation-free entry (no counterpart in the reference) ----------
 * For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): a variant
 * of the pipeline's batch-inversion kernel (Montgomery's trick, one Fermat inversion per 16 elements) whose scratch is a
 * static __device__ array inside the library -- CUDASP_STATIC_INV_MAX elements = 2 MB per GPU that loads the library, held
 * until the process exits -- instead of a cudaMalloc'ed workspace.  d_out[i] = d_in[i]^-1 mod p (p = 2^256 - 2^32 - 977),
 * canonical, 0 -> 0; elements are 32 B little-endian, both pointers are DEVICE pointers on `device`, 16-byte aligned, and
 * may be equal (in place).  Asynchronous on `stream` (a cudaStream_t; 0 = the default stream).  n > CUDASP_STATIC_INV_MAX
 * is refused with CUDASP_ERR_ARG.  The call itself allocates nothing (no cudaMalloc / cudaFree, no allocating library
 * routine).  All calls on one device share the one scratch array: calls that overlap on the GPU (different streams, no
 * dependency between them) race -- serialise them.  Nothing in the scratch has to start at zero and nothing is cleared. */
#define CUDASP_STATIC_INV_MAX 65536u
int cudasp_batch_inverse_static(int device, const uint8_t *d_in, uint8_t *d_out, uint32_t n, void *stream);

/* n entries of the device's fixed-base table, window `window`, digits first_digit .. first_digit+n-1, 64 B LE x||y each:
 * entry (w, d) must be (d * 2^(bits*w) + r_w) * G (bits = cudasp_query_info's fixed_window_bits).  The table is built
 * on the device once per process (it replaces the reference's host-uploaded 2^i*G table, gECC gsv.h:1094-1122); the
 * first call builds it when no query has done so yet. */
int cudasp_debug_gtab(int device, uint32_t window, uint32_t first_digit, uint32_t n, uint8_t *out);
/* Counters: number of kernel launches issued by this query so far; ladder schedule shape. */
uint64_t cudasp_launch_count(const cudasp_query *q);
/* Bytes this query has handed to host->device and device->host copies so far (bench.py's e2e accounting). */
void cudasp_transfer_bytes(const cudasp_query *q, uint64_t *h2d, uint64_t *d2h);
int cudasp_query_info(const cudasp_query *q, int *n_devices, int *ladder_dbl, int *ladder_add, int *fixed_window_bits,
                      int *n_labels);
const char *cudasp_version(void);

#ifdef __cplusplus
}
#endif
#endif /* CUDASP_H */
