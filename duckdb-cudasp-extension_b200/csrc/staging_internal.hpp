// Internal (NOT exported) interface between the C-ABI core (cudasp_capi.cu: slots, streams, kernels) and host-only ingest
// front-ends that fill pinned staging from some columnar format (arrow_ingest.cpp).  A front-end owns the slots it
// acquires here until it releases them; they are invisible to cudasp_poll / cudasp_wait, so several synchronous calls
// and the streaming API can share one query from different threads.
#pragma once
#include <cstdint>

struct cudasp_query;

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
CSP_HIDDEN int fail(cudasp_query *q, int code, const char *msg);

}  // namespace staging
}  // namespace csp
