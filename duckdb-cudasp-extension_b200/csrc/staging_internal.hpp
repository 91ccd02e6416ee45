// Internal (NOT exported) interface between the C-ABI core (cudasp_capi.cu: slots, streams, kernels) and host-only ingest
// front-ends that fill pinned staging from some columnar format (arrow_ingest.cpp).  A front-end owns the slots it
// acquires here until it releases them; they are invisible to cudasp_poll / cudasp_wait, so several synchronous calls
// and the streaming API can share one query from different threads.
#pragma once
#include <cstdint>

struct cudasp_query;
struct ArrowSchema;
struct ArrowArray;

namespace csp {
namespace staging {

#define CSP_HIDDEN __attribute__((visibility("hidden")))

struct Staging {
  void *slot = nullptr;          // opaque
  uint8_t *tweaks = nullptr;     // pinned: rows * key bytes
  int64_t *outputs = nullptr;    // pinned: cap_outputs (may move in reserve_outputs)
  uint32_t *offsets = nullptr;   // pinned: cap_rows + 1
  const uint32_t *ids = nullptr; // after finish(): ascending matching row ids of the batch
  uint32_t cap_rows = 0;
  uint64_t cap_outputs = 0;
};

// 1 = *s is an acquired slot; 0 = every slot is busy (block == false only); < 0 = error
CSP_HIDDEN int acquire(cudasp_query *q, Staging *s, bool block);
// make room for n outputs in this slot (reallocates its output staging; pointers in *s are refreshed)
CSP_HIDDEN int reserve_outputs(cudasp_query *q, Staging *s, uint64_t n);
// async H2D + kernels + D2H on the slot's stream
CSP_HIDDEN int submit(cudasp_query *q, Staging *s, uint32_t rows, uint64_t n_outputs, int key_format);
// wait for the slot's batch; *n ids at s->ids
CSP_HIDDEN int finish(cudasp_query *q, Staging *s, uint32_t *n);
CSP_HIDDEN void release(cudasp_query *q, Staging *s);
// rows per piece a synchronous call should stream in (whole K1 waves, <= batch_rows)
CSP_HIDDEN uint32_t piece_rows(cudasp_query *q);
// devices the query spreads its slots over
CSP_HIDDEN int device_count(cudasp_query *q);
CSP_HIDDEN int fail(cudasp_query *q, int code, const char *msg);


// A scan that spans several (tweak_key, outputs) array pairs, for front-ends that produce them one after another on one thread
// (parquet_ingest.cpp: one pair per row group).  arrow_ingest.cpp: rows of consecutive pushes fill the same pieces, so the GPU
// sees whole pieces whatever the size of the pairs.  Column types and NULL rules are cudasp_scan_arrow_columns'.
struct ColumnScan;
// ramp: start with small pieces (the first rows reach the GPU sooner); a front-end with many threads ramps only a few
CSP_HIDDEN ColumnScan *column_scan_open(cudasp_query *q, bool ramp);
// stages the pair's rows (copied: the arrays may be reused on return); their row numbers count from row_base
CSP_HIDDEN int column_scan_push(ColumnScan *cs, const ArrowSchema *tweak_schema, const ArrowArray *tweak_key,
                                const ArrowSchema *outputs_schema, const ArrowArray *outputs, uint64_t row_base);
// the same with the keys read where they lie: n keys of `width` (64 or 33) bytes, `stride` bytes apart (PLAIN Parquet pages:
// fixed-length values stride = width, length-prefixed ones stride = width + 4), no NULLs; outputs as above (n rows: a slice)
CSP_HIDDEN int column_scan_push_keys(ColumnScan *cs, const uint8_t *keys, int64_t stride, int width, int64_t n,
                                     const ArrowSchema *outputs_schema, const ArrowArray *outputs, uint64_t row_base);
// The matching row numbers of the pieces that have come back so far (those whose slot the scan has recycled); the list only
// grows, in push order.  keys64 / key_widths (may be NULL): the key of match i = keys64[64 * i ..][0 .. key_widths[i]), copied
// from the staging slot -- the caller need not keep or re-read its key column.  Valid until the next call on the scan.
CSP_HIDDEN void column_scan_matches(ColumnScan *cs, const uint64_t **rows, uint64_t *n, const uint8_t **keys64, const uint8_t **key_widths);
// submits what is left and waits for every piece; column_scan_matches() then holds all matches
CSP_HIDDEN int column_scan_finish(ColumnScan *cs);
// nanoseconds the scan has spent waiting for the devices (a piece to come back, a slot to come free): tracing
CSP_HIDDEN uint64_t column_scan_wait_ns(ColumnScan *cs);
CSP_HIDDEN void column_scan_close(ColumnScan *cs);

}  // namespace staging
}  // namespace csp
