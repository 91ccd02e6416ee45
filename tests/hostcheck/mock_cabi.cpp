// CPU test double of the PUBLIC C ABI (include/cudasp.h), the subset the DuckDB extension calls -- so that the extension's
// host logic (bind errors, NULL rules, batch cutting, slot recycling, result gathering, emit_early, many worker threads)
// runs through the real DuckDB engine on a box without a GPU, and so that its per-row host cost can be timed in isolation.
// The "scan" is a stand-in rule, not the EC path: a staged row matches iff one of its outputs equals the int64 read
// little-endian from key bytes [1, 9) (same rule as mock_staging.cpp).  TEST INFRASTRUCTURE ONLY: built as
// tests/hostcheck/mock_cabi/libcudasp_b200.so and selected with LD_LIBRARY_PATH; never linked into or shipped with the product.
//   MOCK_CUDASP_DEVICES  number of pretend devices (default 2)
//   MOCK_CUDASP_STATS    file that receives one line of counters per destroyed query
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/cudasp.h"

namespace {
struct MSlot {
  uint32_t id = 0;
  int device = 0;
  std::vector<uint8_t> tweaks;
  std::vector<int64_t> outputs;
  std::vector<uint32_t> offsets, ids;
  int state = 0;          // 0 free, 1 acquired, 2 in flight, 3 handed out
  uint64_t tag = 0;
  uint32_t rows = 0;
};
thread_local std::string tl_err;
// slots outlive queries (like the product's process-wide slot cache): a second query finds its buffers already touched
std::mutex g_pool_mu;
std::vector<MSlot *> g_pool;
}  // namespace

struct cudasp_query {
  std::mutex mu;
  std::condition_variable cv;
  std::vector<MSlot *> slots;
  std::deque<MSlot *> inflight;
  uint32_t cap_rows = 0, rr = 0, n_labels = 0;
  uint64_t cap_outputs = 0;
  std::string err;
  uint64_t n_submits = 0, rows_staged = 0, outputs_staged = 0, n_sec1 = 0, max_inflight = 0, max_rows = 0;
};

static int fail(cudasp_query *q, int code, const char *m) {
  tl_err = m;
  if (q) q->err = m;
  return code;
}
static int n_devices() {
  const char *e = getenv("MOCK_CUDASP_DEVICES");
  const int v = e ? atoi(e) : 2;
  return v > 0 ? v : 2;
}

extern "C" {
const char *cudasp_version(void) { return "cudasp-b200 MOCK (CPU test double)"; }
int cudasp_device_count(void) { return n_devices(); }
int cudasp_device_sm_count(int) { return 148; }
const char *cudasp_last_error(const cudasp_query *q) { return q ? q->err.c_str() : tl_err.c_str(); }

int cudasp_query_create_ex(const uint8_t *sk, const uint8_t *spend, const uint8_t *labels, uint32_t n_labels, const int *, int,
                           uint32_t batch_rows, uint64_t max_outputs, uint32_t slots_per_device, cudasp_query **out) {
  if (!sk || !spend || (n_labels && !labels) || !out || !batch_rows) return fail(nullptr, CUDASP_ERR_ARG, "mock: bad argument");
  bool zero = true;
  for (int i = 0; i < 32; i++) zero &= sk[i] == 0;
  if (zero) return fail(nullptr, CUDASP_ERR_KEY, "mock: all-zero scan key refused");      // exercises the host's error mapping
  cudasp_query *q = new cudasp_query();
  q->cap_rows = batch_rows;
  q->cap_outputs = max_outputs ? max_outputs : 4ull * batch_rows;
  q->n_labels = n_labels;
  if (slots_per_device == 0) slots_per_device = 6;
  for (int d = 0; d < n_devices(); d++)
    for (uint32_t i = 0; i < slots_per_device; i++) {
      MSlot *s = nullptr;
      {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        for (size_t k = 0; k < g_pool.size(); k++)
          if (g_pool[k]->tweaks.size() >= (size_t)batch_rows * 64 && g_pool[k]->outputs.size() >= q->cap_outputs) {
            s = g_pool[k];
            g_pool.erase(g_pool.begin() + k);
            break;
          }
      }
      if (!s) s = new MSlot();
      s->state = 0;
      s->id = (uint32_t)q->slots.size();
      s->device = d;
      q->slots.push_back(s);          // buffers are allocated on first acquisition, like the product's on-demand slots
    }
  *out = q;
  return CUDASP_OK;
}
void cudasp_query_destroy(cudasp_query *q) {
  if (!q) return;
  if (const char *p = getenv("MOCK_CUDASP_STATS")) {
    if (FILE *f = fopen(p, "a")) {
      fprintf(f, "{\"submits\": %llu, \"rows\": %llu, \"outputs\": %llu, \"sec1_submits\": %llu, \"max_inflight\": %llu, \"max_rows\": %llu}\n",
              (unsigned long long)q->n_submits, (unsigned long long)q->rows_staged, (unsigned long long)q->outputs_staged,
              (unsigned long long)q->n_sec1, (unsigned long long)q->max_inflight, (unsigned long long)q->max_rows);
      fclose(f);
    }
  }
  {
    std::lock_guard<std::mutex> lk(g_pool_mu);
    for (MSlot *s : q->slots) {
      if (s->tweaks.empty() || g_pool.size() >= 64) delete s;
      else g_pool.push_back(s);
    }
  }
  delete q;
}
int cudasp_query_enable_details(cudasp_query *q, uint32_t) { return fail(q, CUDASP_ERR_STATE, "mock: match details need the GPU build"); }
int cudasp_result_details(cudasp_query *q, const cudasp_result *, const cudasp_match_detail **, uint32_t *) {
  return fail(q, CUDASP_ERR_STATE, "mock: match details need the GPU build");
}

static void fill(cudasp_query *q, MSlot *s, cudasp_slot *out) {
  if (s->tweaks.size() < (size_t)q->cap_rows * 64 || s->outputs.size() < q->cap_outputs) {
    s->tweaks.resize((size_t)q->cap_rows * 64);
    s->outputs.resize(q->cap_outputs);
    s->offsets.resize((size_t)q->cap_rows + 1);
  }
  out->slot_id = s->id;
  out->device = s->device;
  out->tweak_keys = s->tweaks.data();
  out->outputs = s->outputs.data();
  out->out_offsets = s->offsets.data();
  out->cap_rows = q->cap_rows;
  out->cap_outputs = q->cap_outputs;
}
static MSlot *take_free(cudasp_query *q) {
  const size_t n = q->slots.size();
  for (size_t k = 0; k < n; k++) {
    MSlot *s = q->slots[(q->rr + k) % n];
    if (s->state == 0) {
      s->state = 1;
      q->rr = (uint32_t)((q->rr + k + 1) % n);
      return s;
    }
  }
  return nullptr;
}
int cudasp_try_acquire_slot(cudasp_query *q, cudasp_slot *slot) {
  if (!q || !slot) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  MSlot *s;
  {
    std::lock_guard<std::mutex> lk(q->mu);
    s = take_free(q);
  }
  if (!s) return 0;
  fill(q, s, slot);
  return 1;
}
int cudasp_acquire_slot(cudasp_query *q, cudasp_slot *slot) {
  if (!q || !slot) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  MSlot *s;
  {
    std::unique_lock<std::mutex> lk(q->mu);
    while (!(s = take_free(q))) q->cv.wait(lk);
  }
  fill(q, s, slot);
  return CUDASP_OK;
}
int cudasp_submit_slot_fmt(cudasp_query *q, const cudasp_slot *slot, uint32_t rows, uint64_t n_outputs, uint64_t tag, int fmt) {
  if (!q || !slot || slot->slot_id >= q->slots.size()) return fail(q, CUDASP_ERR_ARG, "bad slot");
  MSlot *s = q->slots[slot->slot_id];
  if (s->state != 1) return fail(q, CUDASP_ERR_STATE, "slot was not acquired");
  if (rows > q->cap_rows || n_outputs > q->cap_outputs) return fail(q, CUDASP_ERR_ARG, "rows/outputs exceed the slot capacity");
  if (rows && (s->offsets[0] != 0 || s->offsets[rows] != n_outputs)) return fail(q, CUDASP_ERR_ARG, "out_offsets[rows] != n_outputs");
  for (uint32_t i = 0; i < rows; i++)
    if (s->offsets[i + 1] < s->offsets[i]) return fail(q, CUDASP_ERR_ARG, "out_offsets must be monotonic");
  const size_t kb = fmt == CUDASP_KEYS_SEC1_33 ? 33 : 64;
  s->ids.clear();
  if (!getenv("MOCK_CUDASP_NO_SCAN"))
    for (uint32_t i = 0; i < rows; i++) {
      int64_t want;
      memcpy(&want, s->tweaks.data() + (size_t)i * kb + 1, 8);
      for (uint32_t o = s->offsets[i]; o < s->offsets[i + 1]; o++)
        if (s->outputs[o] == want) { s->ids.push_back(i); break; }
    }
  std::lock_guard<std::mutex> lk(q->mu);
  s->state = 2;
  s->tag = tag;
  s->rows = rows;
  q->inflight.push_back(s);
  q->n_submits++;
  q->rows_staged += rows;
  q->outputs_staged += n_outputs;
  q->n_sec1 += fmt == CUDASP_KEYS_SEC1_33;
  if (q->inflight.size() > q->max_inflight) q->max_inflight = q->inflight.size();
  if (rows > q->max_rows) q->max_rows = rows;
  return CUDASP_OK;
}
static int take(cudasp_query *q, cudasp_result *r) {
  if (!q || !r) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  std::lock_guard<std::mutex> lk(q->mu);
  if (q->inflight.empty()) return 0;
  MSlot *s = q->inflight.front();
  q->inflight.pop_front();
  s->state = 3;
  r->tag = s->tag;
  r->slot_id = s->id;
  r->rows = s->rows;
  r->n_matches = (uint32_t)s->ids.size();
  r->flags = 0;
  r->match_ids = s->ids.data();
  return 1;
}
int cudasp_poll(cudasp_query *q, cudasp_result *r) { return take(q, r); }
int cudasp_wait(cudasp_query *q, cudasp_result *r) { return take(q, r); }
int cudasp_release_result(cudasp_query *q, const cudasp_result *r) {
  if (!q || !r || r->slot_id >= q->slots.size()) return fail(q, CUDASP_ERR_ARG, "bad result");
  {
    std::lock_guard<std::mutex> lk(q->mu);
    MSlot *s = q->slots[r->slot_id];
    if (s->state != 3) return fail(q, CUDASP_ERR_STATE, "result was not handed out");
    s->state = 0;
  }
  q->cv.notify_all();
  return CUDASP_OK;
}
}  // extern "C"

// cudasp_scan_parquet (include/cudasp_parquet.h) needs the real library's page decoders and staging: not in this double
#include "../../include/cudasp_parquet.h"
extern "C" int cudasp_scan_parquet(cudasp_query *q, const char *const *paths, int n_paths, int, cudasp_parquet_result **) {
  std::string m = "mock: cudasp_scan_parquet needs the GPU build; paths:";       // the paths as the extension's bind passed them on
  for (int i = 0; i < n_paths; i++) {
    const char *slash = strrchr(paths[i], '/');
    m += std::string(" ") + (slash ? slash + 1 : paths[i]);
  }
  return fail(q, CUDASP_ERR_UNSUPPORTED, m.c_str());
}
extern "C" void cudasp_parquet_result_free(cudasp_parquet_result *) {}
