// CPU test double of the C-ABI core behind csrc/staging_internal.hpp, so that the Arrow ingest front-end
// (csrc/arrow_ingest.cpp) -- NULL rules, row maps, piece cutting, key-format switches, slot recycling -- is exercised
// without a GPU.  The "scan" here is a stand-in rule, not the EC path: a staged row matches iff one of its outputs
// equals the int64 read little-endian from key bytes [1, 9).  TEST INFRASTRUCTURE ONLY; never linked into the product.
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/cudasp.h"
#include "../../duckdb-cudasp-extension_b200/csrc/staging_internal.hpp"

struct MockSlot {
  std::vector<uint8_t> tweaks;
  std::vector<int64_t> outputs;
  std::vector<uint32_t> offsets, ids;
  bool busy = false, submitted = false;
  uint32_t rows = 0;
  int fmt = 0;
};

struct cudasp_query {
  uint32_t piece = 0, cap_rows = 0;
  uint64_t cap_outputs = 0;
  std::vector<MockSlot> slots;
  std::string err;
  std::condition_variable freed;
  std::mutex mu;               // the slot table and the counters: several front-end threads may share one query (Parquet decoder threads)
  uint64_t n_submits = 0, rows_staged = 0, n_submits_sec1 = 0, outputs_staged = 0, max_in_flight = 0, n_reserve = 0;
};

namespace csp {
namespace staging {
static void view(Staging *v, MockSlot *s, cudasp_query *q) {
  v->slot = s;
  v->tweaks = s->tweaks.data();
  v->outputs = s->outputs.data();
  v->offsets = s->offsets.data();
  v->ids = s->ids.data();
  v->cap_rows = q->cap_rows;
  v->cap_outputs = s->outputs.size();
}
int acquire(cudasp_query *q, Staging *v, bool block) {
  std::unique_lock<std::mutex> lk(q->mu);
  for (;;) {
    uint64_t busy = 0;
    for (MockSlot &s : q->slots) busy += s.busy;
    for (MockSlot &s : q->slots) {
      if (!s.busy) {
        s.busy = true;
        s.submitted = false;
        memset(s.tweaks.data(), 0xEE, s.tweaks.size());        // stale data must never be read
        if (busy + 1 > q->max_in_flight) q->max_in_flight = busy + 1;
        view(v, &s, q);
        return 1;
      }
    }
    if (!block) return 0;
    // another front-end thread may release a slot; a lone caller that blocks with every slot its own would wait for ever
    if (q->freed.wait_for(lk, std::chrono::seconds(2)) == std::cv_status::timeout) {
      lk.unlock();
      return fail(q, CUDASP_ERR_STATE, "mock: blocking acquire with every slot busy would deadlock");
    }
  }
}
int reserve_outputs(cudasp_query *q, Staging *v, uint64_t n) {
  MockSlot *s = static_cast<MockSlot *>(v->slot);
  { std::lock_guard<std::mutex> lk(q->mu); q->n_reserve++; }
  if (n > s->outputs.size()) s->outputs.resize(n + n / 4);
  view(v, s, q);
  return 0;
}
int submit(cudasp_query *q, Staging *v, uint32_t rows, uint64_t n_outputs, int key_format) {
  MockSlot *s = static_cast<MockSlot *>(v->slot);
  if (!s->busy || s->submitted) return fail(q, CUDASP_ERR_STATE, "mock: submit on a slot that is not acquired");
  if (rows > q->cap_rows || n_outputs > s->outputs.size()) return fail(q, CUDASP_ERR_ARG, "mock: piece exceeds the slot");
  if (s->offsets[0] != 0 || s->offsets[rows] != n_outputs) return fail(q, CUDASP_ERR_ARG, "mock: offsets do not frame the outputs");
  const size_t kb = key_format == CUDASP_KEYS_SEC1_33 ? 33 : 64;
  s->ids.clear();
  static const bool no_scan = getenv("MOCK_STAGING_NO_SCAN") != nullptr;       // timing aid: the front-end's own cost alone
  for (uint32_t r = 0; r < rows && !no_scan; r++) {
    if (s->offsets[r + 1] < s->offsets[r]) return fail(q, CUDASP_ERR_ARG, "mock: offsets not monotonic");
    int64_t want;
    memcpy(&want, s->tweaks.data() + r * kb + 1, 8);
    bool hit = false;
    for (uint32_t j = s->offsets[r]; j < s->offsets[r + 1]; j++) hit |= s->outputs[j] == want;
    if (hit) s->ids.push_back(r);
  }
  s->ids.push_back(0);          // keep data() non-null
  s->rows = rows;
  s->fmt = key_format;
  s->submitted = true;
  {
    std::lock_guard<std::mutex> lk(q->mu);
    q->n_submits++;
    q->rows_staged += rows;
    q->outputs_staged += n_outputs;
    q->n_submits_sec1 += key_format == CUDASP_KEYS_SEC1_33;
  }
  view(v, s, q);
  return 0;
}
int finish(cudasp_query *q, Staging *v, uint32_t *n) {
  MockSlot *s = static_cast<MockSlot *>(v->slot);
  if (!s->submitted) return fail(q, CUDASP_ERR_STATE, "mock: finish before submit");
  *n = (uint32_t)s->ids.size() - 1;
  v->ids = s->ids.data();
  return 0;
}
void release(cudasp_query *q, Staging *v) {
  if (v->slot) {
    std::lock_guard<std::mutex> lk(q->mu);
    static_cast<MockSlot *>(v->slot)->busy = false;
    q->freed.notify_all();
  }
  v->slot = nullptr;
}
uint32_t piece_rows(cudasp_query *q) { return q->piece; }
int device_count(cudasp_query *) { return 1; }
int fail(cudasp_query *q, int code, const char *msg) {
  if (q) {
    std::lock_guard<std::mutex> lk(q->mu);
    q->err = msg;
  }
  return code;
}
}  // namespace staging
}  // namespace csp

extern "C" {
cudasp_query *mock_query_create(uint32_t piece, uint32_t cap_rows, uint64_t cap_outputs, uint32_t n_slots) {
  cudasp_query *q = new cudasp_query;
  q->piece = piece;
  q->cap_rows = cap_rows;
  q->cap_outputs = cap_outputs;
  q->slots.resize(n_slots);
  for (MockSlot &s : q->slots) {
    s.tweaks.resize((size_t)cap_rows * 64);
    s.outputs.resize(cap_outputs);
    s.offsets.resize((size_t)cap_rows + 1);
  }
  return q;
}
void mock_query_destroy(cudasp_query *q) { delete q; }
int mock_busy_slots(cudasp_query *q) {
  int n = 0;
  for (MockSlot &s : q->slots) n += s.busy;
  return n;
}
void mock_stats(cudasp_query *q, uint64_t *out) {
  out[0] = q->n_submits;
  out[1] = q->rows_staged;
  out[2] = q->n_submits_sec1;
  out[3] = q->outputs_staged;
  out[4] = q->max_in_flight;
  out[5] = q->n_reserve;
}
const char *cudasp_last_error(const cudasp_query *q) { return q ? q->err.c_str() : ""; }
}
