// cudasp_capi.cu -- implementation of include/cudasp.h (the drop-in C ABI).
//
// Replaces LaunchBatchScan / RunBatchScanKernels / FreeBatchScanState and their BatchScanState
// (src/cudasp_gpu.cu:425-1023): where the reference allocates ~20 buffers, creates a stream and
// synchronises >= 5 times per batch, this layer allocates everything once per query, keeps
// SLOTS_PER_DEVICE pinned+device staging slots per GPU, each with its own stream, and runs a batch
// as: 3 async H2D copies -> 5 (L = 0) or 7 (L > 0) kernels -> one small async D2H (count + first ids).
// No host synchronisation happens until the caller asks for a result.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/cudasp.h"
#include "launch.hpp"
#include "piece_plan.hpp"
#include "query_plan.hpp"
#include "staging_internal.hpp"

static_assert(sizeof(csp::MatchDetail) == sizeof(cudasp_match_detail) && sizeof(cudasp_match_detail) == 48, "detail record layout");

using namespace csp;

namespace {

thread_local std::string tl_err;

const int SLOTS_PER_DEVICE = 6, MAX_SLOTS_PER_DEVICE = 64;   // 6: measured, a 3-deep queue lets the GPU run dry (profiles/r01e)
const uint32_t EAGER_IDS = 1024;   // ids copied back together with the count
const uint32_t EAGER_DETAILS = 32; // detail records copied back together with their count
const uint32_t RAMP_DIV = 8;               // ... starting with 1/8 of a piece, doubling (see scan_batch_impl)
const uint32_t SYNC_PIECE_WAVES = 2;       // cudasp_scan_batch streams a large call in pieces of this many K1 waves
const uint32_t INV_ROWS_PER_THREAD = 16;

struct Workspace {          // device-only scratch for one batch in flight
  fe *X = nullptr, *Y = nullptr, *Z = nullptr, *S = nullptr;   // rows each
  fe *D = nullptr, *DS = nullptr;                              // L * rows each (labels only)
  uint8_t *flags = nullptr;
  uint8_t *invalid = nullptr;                                  // rows bytes: malformed compressed keys (KEYFMT 1)
  cudaEvent_t free_evt = nullptr;                              // recorded after the last kernel using it
  uint32_t cap_rows = 0;
  uint32_t cap_labels = 0;
};

struct Slot {
  uint32_t id = 0;
  int dev_index = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t done = nullptr;
  cudaEvent_t tr[3] = {nullptr, nullptr, nullptr};   // CUDASP_TRACE: before H2D, after H2D, after the last kernel
  // pinned host staging
  uint8_t *h_tweaks = nullptr;
  int64_t *h_outputs = nullptr;
  uint32_t *h_off = nullptr;
  uint32_t *h_result = nullptr;   // [0] = count, [1 ..] = ids (cap_rows)
  // device inputs / outputs
  uint8_t *d_tweaks = nullptr;
  int64_t *d_outputs = nullptr;
  uint32_t *d_off = nullptr;
  uint32_t *d_result = nullptr;   // [0] = count, [1 ..] = ids
  uint32_t cap_rows = 0;
  uint64_t cap_outputs = 0;
  Workspace ws;
  // match details (opt-in, cudasp_query_enable_details): entry 0 of each buffer is a header whose first word is the count
  MatchDetail *d_det = nullptr, *h_det = nullptr;
  fe *keep_sx = nullptr;          // cap_rows: affine x of the shared secret, kept by K3 for K8
  uint8_t *keep_sodd = nullptr;   // cap_rows: parity of its y
  uint32_t det_cap = 0;
  uint32_t n_det = 0;             // after finish()
  bool det_truncated = false;     // after finish(): K8 found more records than the query's details capacity
  // state
  int state = 0;                  // 0 free, 1 acquired, 2 in flight, 3 result handed out
  uint64_t tag = 0;
  uint32_t rows = 0;
  bool internal = false;          // owned by a synchronous call (not visible to poll/wait)
};

struct DevCtx {
  int device = 0;
  // A device is brought up (context, fixed-base table, labels, own stream) by the first slot materialised on it
  // (ensure_device): a query that never fills a second batch touches one GPU only.
  bool up = false;
  std::unique_ptr<std::mutex> init_mu{new std::mutex()};
  // Users of one workspace are ordered on the device by its free_evt: wait -> launches -> record.  Two host threads
  // interleaving those three steps on the same workspace (scan_device borrows the slots' workspaces) would both wait on
  // the OLD event and then run concurrently on the same scratch, so the sequence is made atomic per device.
  std::unique_ptr<std::mutex> launch_mu{new std::mutex()};
  ge *d_gtab = nullptr;
  ge *d_labels = nullptr;
  cudaStream_t own_stream = nullptr;
  std::vector<Slot *> slots;      // materialised so far (grown on demand up to max_slots)
  uint32_t max_slots = 0;
  uint32_t pending = 0;           // slots being allocated right now (outside the lock)
  uint32_t next_ws = 0;
};

}  // namespace

struct cudasp_query {
  host::QueryPlan plan;
  uint32_t n_labels = 0;
  uint32_t batch_rows = 0;
  uint64_t max_outputs = 0;
  uint32_t details_cap = 0;      // > 0: match details enabled, records per batch
  uint32_t wave_threads = 0;     // threads of one full wave of the 128-register kernels: SMs x 4 blocks x 128
  uint32_t piece_rows = 0;       // rows per piece of the synchronous call: whole K1 waves (SMs x 4 blocks x 128 rows)
  std::vector<DevCtx> devs;
  std::vector<Slot *> all_slots;
  std::mutex mu;
  std::condition_variable cv;
  std::deque<Slot *> inflight;   // public streaming API, submit order
  uint32_t rr = 0;               // round-robin cursor for slot acquisition
  std::string err;
  std::atomic<uint64_t> launches{0};
  std::atomic<uint64_t> h2d_bytes{0}, d2h_bytes{0};   // bytes handed to cudaMemcpyAsync by submit() / finish()
};

namespace {

int fail(cudasp_query *q, int code, const std::string &msg) {
  tl_err = msg;
  if (q) {
    std::lock_guard<std::mutex> lk(q->mu);
    q->err = msg;
  }
  return code;
}
// q->mu is already held by the caller (std::mutex is not recursive)
int fail_locked(cudasp_query *q, int code, const std::string &msg) {
  tl_err = msg;
  q->err = msg;
  return code;
}
#define CU(q, call)                                                                                      \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess)                                                                               \
      return fail(q, CUDASP_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));               \
  } while (0)

// ---- per-process, per-device fixed-base table, built ON the device (k_build_gtab) from the host's per-window seeds.
// start_device_gtab only enqueues (so that the devices of one query build concurrently); get_device_gtab waits.
struct DevGTab {
  ge *table = nullptr;
  ge *seeds = nullptr;       // [2][GTAB_NWIN]: firsts, bases
  uint32_t *bad = nullptr;   // exceptional additions met by the builder (must stay 0)
  bool pending = false, ready = false;
};
const int MAX_DEVICES = 64;
std::mutex g_gtab_mu[MAX_DEVICES];   // one per device ordinal: devices brought up by different threads build concurrently
DevGTab g_gtab_dev[MAX_DEVICES];

static void drop_device_gtab(DevGTab &t) {   // after a failed build: leave nothing half-made behind
  cudaFree(t.table); cudaFree(t.seeds); cudaFree(t.bad);
  t = DevGTab();
}
int start_device_gtab_locked(cudasp_query *q, int device) {
  DevGTab &t = g_gtab_dev[device];
  if (t.ready || t.pending) return 0;
  const host::GTabHost &g = host::gtab_host();
  auto body = [&]() -> int {
    CU(q, cudaSetDevice(device));
    CU(q, cudaMalloc(&t.table, GTAB_TOTAL * sizeof(ge)));
    CU(q, cudaMalloc(&t.seeds, 2 * (size_t)GTAB_NWIN * sizeof(ge)));
    CU(q, cudaMalloc(&t.bad, sizeof(uint32_t)));
    CU(q, cudaMemcpy(t.seeds, g.seeds.firsts.data(), GTAB_NWIN * sizeof(ge), cudaMemcpyHostToDevice));
    CU(q, cudaMemcpy(t.seeds + GTAB_NWIN, g.seeds.bases.data(), GTAB_NWIN * sizeof(ge), cudaMemcpyHostToDevice));
    CU(q, cudaMemset(t.bad, 0, sizeof(uint32_t)));
    CU(q, launch::build_gtab(t.table, t.seeds, t.seeds + GTAB_NWIN, t.bad));
    return 0;
  };
  const int rc = body();
  if (rc) { drop_device_gtab(t); return rc; }
  t.pending = true;
  return 0;
}
int get_device_gtab(cudasp_query *q, int device, ge **out) {
  if (device < 0 || device >= MAX_DEVICES) return fail(q, CUDASP_ERR_ARG, "device ordinal out of range");
  std::lock_guard<std::mutex> lk(g_gtab_mu[device]);
  int rc = start_device_gtab_locked(q, device);
  if (rc) return rc;
  DevGTab &t = g_gtab_dev[device];
  if (t.pending) {
    uint32_t bad = 0;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaMemcpy(&bad, t.bad, sizeof bad, cudaMemcpyDeviceToHost);   // default stream: waits for the builder
    if (e != cudaSuccess || bad) {
      drop_device_gtab(t);
      return e != cudaSuccess ? fail(q, CUDASP_ERR_CUDA, std::string("fixed-base table build: ") + cudaGetErrorString(e))
                              : fail(q, CUDASP_ERR_STATE, "fixed-base table builder met an exceptional addition");
    }
    t.pending = false;
    t.ready = true;
    cudaFree(t.seeds); cudaFree(t.bad);
    t.seeds = nullptr; t.bad = nullptr;
  }
  if (!t.ready) return fail(q, CUDASP_ERR_STATE, "fixed-base table is not available on this device");
  *out = t.table;
  return 0;
}

int alloc_workspace(cudasp_query *q, Workspace &w, uint32_t rows, uint32_t L) {
  w.cap_rows = rows;
  w.cap_labels = L;
  size_t n = (size_t)rows;
  CU(q, cudaMalloc(&w.X, n * sizeof(fe)));
  CU(q, cudaMalloc(&w.Y, n * sizeof(fe)));
  CU(q, cudaMalloc(&w.Z, n * sizeof(fe)));
  CU(q, cudaMalloc(&w.S, n * sizeof(fe)));
  CU(q, cudaMalloc(&w.invalid, n));
  if (L) {
    CU(q, cudaMalloc(&w.D, n * L * sizeof(fe)));
    CU(q, cudaMalloc(&w.DS, n * L * sizeof(fe)));
    CU(q, cudaMalloc(&w.flags, n));
  }
  CU(q, cudaEventCreateWithFlags(&w.free_evt, cudaEventDisableTiming));
  return 0;
}
void free_workspace(Workspace &w) {
  cudaFree(w.X); cudaFree(w.Y); cudaFree(w.Z); cudaFree(w.S); cudaFree(w.invalid); cudaFree(w.D); cudaFree(w.DS); cudaFree(w.flags);
  if (w.free_evt) cudaEventDestroy(w.free_evt);
  w = Workspace();
}

int alloc_slot_outputs(cudasp_query *q, Slot &s, uint64_t cap_outputs) {
  if (s.h_outputs) cudaFreeHost(s.h_outputs);
  if (s.d_outputs) cudaFree(s.d_outputs);
  s.h_outputs = nullptr; s.d_outputs = nullptr;
  s.cap_outputs = cap_outputs;
  CU(q, cudaHostAlloc((void **)&s.h_outputs, std::max<uint64_t>(cap_outputs, 1) * sizeof(int64_t), cudaHostAllocDefault));
  CU(q, cudaMalloc(&s.d_outputs, std::max<uint64_t>(cap_outputs, 1) * sizeof(int64_t)));
  return 0;
}

int alloc_slot(cudasp_query *q, Slot &s, int device, uint32_t rows, uint64_t cap_outputs, uint32_t L) {
  CU(q, cudaSetDevice(device));
  s.cap_rows = rows;
  CU(q, cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
  CU(q, cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
  for (int i = 0; i < 3; i++) CU(q, cudaEventCreate(&s.tr[i]));
  CU(q, cudaHostAlloc((void **)&s.h_tweaks, (size_t)rows * 64, cudaHostAllocDefault));
  CU(q, cudaHostAlloc((void **)&s.h_off, ((size_t)rows + 1) * 4, cudaHostAllocDefault));
  CU(q, cudaHostAlloc((void **)&s.h_result, ((size_t)rows + 1) * 4, cudaHostAllocDefault));
  CU(q, cudaMalloc(&s.d_tweaks, (size_t)rows * 64));
  CU(q, cudaMalloc(&s.d_off, ((size_t)rows + 1) * 4));
  CU(q, cudaMalloc(&s.d_result, ((size_t)rows + 1) * 4));
  int rc = alloc_slot_outputs(q, s, cap_outputs);
  if (rc) return rc;
  return alloc_workspace(q, s.ws, rows, L);
}
void free_slot(Slot &s) {
  cudaFreeHost(s.h_tweaks); cudaFreeHost(s.h_outputs); cudaFreeHost(s.h_off); cudaFreeHost(s.h_result);
  cudaFree(s.d_tweaks); cudaFree(s.d_outputs); cudaFree(s.d_off); cudaFree(s.d_result);
  cudaFreeHost(s.h_det); cudaFree(s.d_det); cudaFree(s.keep_sx); cudaFree(s.keep_sodd);
  free_workspace(s.ws);
  if (s.done) cudaEventDestroy(s.done);
  for (int i = 0; i < 3; i++) if (s.tr[i]) cudaEventDestroy(s.tr[i]);
  if (s.stream) cudaStreamDestroy(s.stream);
}

uint32_t inv_per_thread() {
  // elements per Fermat inversion (launch::batch_inv); CUDASP_INV_PER_THREAD is a tuning aid, read once
  static const uint32_t per = [] {
    const char *e = getenv("CUDASP_INV_PER_THREAD");
    const int v = e ? atoi(e) : 0;
    return v > 0 ? (uint32_t)v : INV_ROWS_PER_THREAD;
  }();
  return per;
}

// ---- process-wide cache of staging slots -------------------------------------------------------
// Pinned + device allocation of a slot costs tens of milliseconds; a host that runs many queries (DuckDB)
// would pay it per query.  Destroyed queries park their slots here; new queries take them back when the
// capacities fit.  cudasp_release_cached_memory() empties the cache.
std::mutex g_cache_mu;
std::vector<std::vector<Slot *>> g_cache;   // by device ordinal
const size_t CACHE_MAX_PER_DEVICE = 40;

Slot *cache_take(int device, uint32_t rows, uint64_t cap_outputs) {
  std::lock_guard<std::mutex> lk(g_cache_mu);
  if ((int)g_cache.size() <= device) return nullptr;
  auto &v = g_cache[device];
  for (size_t i = 0; i < v.size(); i++) {
    Slot *s = v[i];
    if (s->cap_rows >= rows && s->cap_rows <= std::max<uint64_t>(2ull * rows, 65536) && s->cap_outputs >= cap_outputs) {
      v.erase(v.begin() + i);
      return s;
    }
  }
  return nullptr;
}
void cache_put(int device, Slot *s) {
  {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    if ((int)g_cache.size() <= device) g_cache.resize(device + 1);
    if (g_cache[device].size() < CACHE_MAX_PER_DEVICE) {
      s->state = 0; s->internal = false; s->rows = 0; s->tag = 0;
      g_cache[device].push_back(s);
      return;
    }
  }
  free_slot(*s);
  delete s;
}

// Bring a device of the query up: CUDA context, the process-wide fixed-base table (built on first use), the query's
// labels, its own stream.  Called by the first slot materialised on the device, from whichever thread asks first;
// several devices come up concurrently when several threads ask (per-device locks only).
int ensure_device(cudasp_query *q, int di) {
  DevCtx &dc = q->devs[di];
  std::lock_guard<std::mutex> lk(*dc.init_mu);
  if (dc.up) return 0;
  int rc = get_device_gtab(q, dc.device, &dc.d_gtab);
  if (rc) return rc;
  CU(q, cudaSetDevice(dc.device));
  if (const char *e = getenv("CUDASP_L2_FETCH")) {       // tuning aid: L2 fetch granularity (32 / 64 / 128 B) for K3's random table reads
    const int v = atoi(e);
    if (v == 32 || v == 64 || v == 128) CU(q, cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)v));
  }
  if (!dc.own_stream) CU(q, cudaStreamCreateWithFlags(&dc.own_stream, cudaStreamNonBlocking));
  if (q->n_labels && !dc.d_labels) {
    CU(q, cudaMalloc(&dc.d_labels, q->n_labels * sizeof(ge)));
    CU(q, cudaMemcpy(dc.d_labels, q->plan.labels.data(), q->n_labels * sizeof(ge), cudaMemcpyHostToDevice));
  }
  {
    std::lock_guard<std::mutex> lk2(q->mu);       // readers hold q->mu (try_acquire) or init_mu (here)
    dc.up = true;
  }
  return 0;
}

// A ready-to-use slot for this query on device index `di`: from the cache or freshly allocated.
int materialise_slot(cudasp_query *q, int di, Slot **out) {
  DevCtx &dc = q->devs[di];
  const uint32_t L = q->n_labels;
  int rc0 = ensure_device(q, di);
  if (rc0) return rc0;
  CU(q, cudaSetDevice(dc.device));
  Slot *s = cache_take(dc.device, q->batch_rows, q->max_outputs);
  if (s) {
    if (L > s->ws.cap_labels) {      // label scratch too small for this query: regrow it
      cudaFree(s->ws.D); cudaFree(s->ws.DS); cudaFree(s->ws.flags);
      s->ws.D = s->ws.DS = nullptr; s->ws.flags = nullptr;
      const size_t n = (size_t)s->ws.cap_rows;
      cudaError_t e = cudaMalloc(&s->ws.D, n * L * sizeof(fe));
      if (e == cudaSuccess) e = cudaMalloc(&s->ws.DS, n * L * sizeof(fe));
      if (e == cudaSuccess) e = cudaMalloc(&s->ws.flags, n);
      if (e != cudaSuccess) { free_slot(*s); delete s; return fail(q, CUDASP_ERR_CUDA, cudaGetErrorString(e)); }
      s->ws.cap_labels = L;
    }
  } else {
    s = new Slot();
    int rc = alloc_slot(q, *s, dc.device, q->batch_rows, q->max_outputs, L);
    if (rc) { free_slot(*s); delete s; return rc; }
  }
  s->dev_index = di;
  s->state = 0;
  *out = s;
  return 0;
}

inline size_t key_bytes(int keyfmt) { return keyfmt == CUDASP_KEYS_SEC1_33 ? 33 : 64; }
int ladder_variant() {     // K1 launch shape (k_ladder.cu); 0 = production, others only in tuning builds
  static int v = [] { const char *e = getenv("CUDASP_LADDER_VARIANT"); return e ? atoi(e) : 0; }();
  return v;
}
cudaError_t launch_ladder(int keyfmt, const uint8_t *d_tweaks, fe *X, fe *Y, fe *Z, uint32_t rows, const LadderParams &lp,
                          uint8_t *invalid, cudaStream_t st) {
  return keyfmt == CUDASP_KEYS_SEC1_33 ? launch::ladder_sec1(ladder_variant(), d_tweaks, X, Y, Z, rows, lp, invalid, st)
                                       : launch::ladder_xy64(ladder_variant(), d_tweaks, X, Y, Z, rows, lp, invalid, st);
}

struct DetailSink {          // where K3 keeps the shared secrets and K8 appends its records
  fe *keep_sx;
  uint8_t *keep_sodd;
  MatchDetail *out;          // out[0] is the header (count in its first word), records start at out[1]
  uint32_t cap;
};
// (re)allocate the slot's detail buffers for the query's capacity; slots come from the cache or are materialised lazily,
// so this is checked on every submit (a compare when nothing is to do)
int ensure_details(cudasp_query *q, Slot &s) {
  if (s.det_cap >= q->details_cap && s.keep_sx) return 0;
  CU(q, cudaStreamSynchronize(s.stream));
  cudaFreeHost(s.h_det); cudaFree(s.d_det); cudaFree(s.keep_sx); cudaFree(s.keep_sodd);
  s.h_det = s.d_det = nullptr; s.keep_sx = nullptr; s.keep_sodd = nullptr; s.det_cap = 0;
  const size_t n = (size_t)q->details_cap + 1;
  CU(q, cudaHostAlloc((void **)&s.h_det, n * sizeof(MatchDetail), cudaHostAllocDefault));
  CU(q, cudaMalloc(&s.d_det, n * sizeof(MatchDetail)));
  CU(q, cudaMalloc(&s.keep_sx, (size_t)s.cap_rows * sizeof(fe)));
  CU(q, cudaMalloc(&s.keep_sodd, s.cap_rows));
  s.det_cap = q->details_cap;
  return 0;
}

// Enqueue the whole per-batch pipeline on `st`.  All pointers are device pointers.
// d_result: [0] = match count (reset here), [1..] = ids.  Either may be NULL (candidate mode).
int launch_pipeline(cudasp_query *q, DevCtx &dc, Workspace &w, const uint8_t *d_tweaks, const int64_t *d_outputs,
                    const uint32_t *d_off, uint32_t rows, uint32_t *d_ids, uint32_t *d_count, cudaStream_t st,
                    int64_t *d_cands, uint8_t *dbg_S, uint8_t *dbg_t, uint8_t *dbg_P0, cudaEvent_t *stage_evs = nullptr,
                    int keyfmt = CUDASP_KEYS_XY64, const DetailSink *det = nullptr) {
  // stage_evs (profiling entry only): 8 events recorded before each of the 7 possible kernels and
  // after the last one; a stage that does not run records its event anyway (zero duration).
  int ev_i = 0;
  auto mark = [&]() { if (stage_evs) cudaEventRecord(stage_evs[ev_i++], st); };
  if (det) CU(q, cudaMemsetAsync(det->out, 0, 4, st));
  if (rows == 0) {
    if (d_count) CU(q, cudaMemsetAsync(d_count, 0, 4, st));
    return 0;
  }
  if (rows > w.cap_rows) return fail(q, CUDASP_ERR_ARG, "rows exceed the workspace capacity");
  const uint32_t L = q->n_labels;
  const host::QueryPlan &pl = q->plan;
  const uint8_t *inv_flags = keyfmt == CUDASP_KEYS_SEC1_33 ? w.invalid : nullptr;
  uint64_t nl = 0;
  if (d_count) CU(q, cudaMemsetAsync(d_count, 0, 4, st));
  mark();
  const uint32_t per = inv_per_thread();
  if (!pl.fp.zero_scalar) {
    CU(q, launch_ladder(keyfmt, d_tweaks, w.X, w.Y, w.Z, rows, pl.lp, w.invalid, st));
    mark();
    CU(q, launch::batch_inv(w.Z, w.S, rows, per, st));
    nl += 2;
  } else {
    // scan key = 0 (mod n): S is the all-zero pair for every row; compressed keys are still validated (empty op list)
    if (keyfmt == CUDASP_KEYS_SEC1_33) {
      CU(q, launch_ladder(keyfmt, d_tweaks, w.X, w.Y, w.Z, rows, pl.lp, w.invalid, st));
      nl += 1;
    }
    mark();
  }
  mark();
  CU(q, launch::hash_fixedbase(w.X, w.Y, w.Z, rows, dc.d_gtab, pl.fp, dbg_S, dbg_t, det ? det->keep_sx : nullptr,
                               det ? det->keep_sodd : nullptr, st));
  mark();
  CU(q, launch::batch_inv(w.Z, w.S, rows, per, st));
  mark();
  CU(q, launch::final_base(w.X, w.Y, w.Z, rows, d_outputs, d_off, dc.d_labels, L, w.D, w.flags, d_ids, d_count, d_cands, dbg_P0,
                           inv_flags, st));
  nl += 3;
  mark();
  if (L) {
    const uint64_t n64 = (uint64_t)rows * L;
    if (n64 > 0xffffffffull) return fail(q, CUDASP_ERR_ARG, "rows * labels exceeds 2^32");
    // The label denominators are rows * L elements: with many labels there are threads to spare, so more elements share one
    // Fermat inversion (255 S + 15 M per thread whatever it inverts): 16 per thread costs (48 + 270) / 16 = 20 field operations
    // per element, 64 per thread 7.  Keep about four warps per scheduler (q->wave_threads) in flight, never fewer than `per`.
    static const bool fixed_per = getenv("CUDASP_INV_PER_THREAD") != nullptr;
    const uint32_t per_labels = fixed_per ? per : (uint32_t)std::min<uint64_t>(128, std::max<uint64_t>(per, n64 / std::max(q->wave_threads, 1u)));
    CU(q, launch::batch_inv(w.D, w.DS, (uint32_t)n64, per_labels, st));
    mark();
    CU(q, launch::final_labels(w.X, w.Y, w.D, rows, d_outputs, d_off, dc.d_labels, L, w.flags, d_ids, d_count, d_cands, inv_flags, st));
    nl += 2;
  } else {
    mark();
  }
  mark();
  if (det && d_ids && d_count && d_outputs) {      // K8: the BIP-352 k-loop over the rows that matched (a handful)
    CU(q, launch::details(det->keep_sx, det->keep_sodd, d_ids, d_count, d_outputs, d_off, dc.d_labels, L, dc.d_gtab, pl.fp,
                          det->out + 1, reinterpret_cast<uint32_t *>(det->out), det->cap, st));
    nl += 1;
  }
  q->launches += nl;
  return 0;
}

}  // namespace

// =====================================================================================
extern "C" {

const char *cudasp_version(void) { return "cudasp-b200 0.1 (sm_100a)"; }

int cudasp_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

const char *cudasp_last_error(const cudasp_query *q) {
  if (q) return q->err.c_str();
  return tl_err.c_str();
}

void cudasp_query_destroy(cudasp_query *q) {
  if (!q) return;
  for (DevCtx &dc : q->devs) {
    if (!dc.up && dc.slots.empty()) continue;      // never brought up: no context to touch
    cudaSetDevice(dc.device);
    cudaDeviceSynchronize();
    for (Slot *s : dc.slots) cache_put(dc.device, s);
    cudaFree(dc.d_labels);
    if (dc.own_stream) cudaStreamDestroy(dc.own_stream);
  }
  delete q;
}

int cudasp_query_create(const uint8_t scan_private_key[32], const uint8_t spend_public_key[64], const uint8_t *label_keys,
                        uint32_t n_labels, const int *devices, int n_devices, uint32_t batch_rows,
                        uint64_t max_outputs_per_batch, cudasp_query **out) {
  return cudasp_query_create_ex(scan_private_key, spend_public_key, label_keys, n_labels, devices, n_devices, batch_rows,
                                max_outputs_per_batch, 0, out);
}

int cudasp_query_create_ex(const uint8_t scan_private_key[32], const uint8_t spend_public_key[64], const uint8_t *label_keys,
                           uint32_t n_labels, const int *devices, int n_devices, uint32_t batch_rows,
                           uint64_t max_outputs_per_batch, uint32_t slots_per_device, cudasp_query **out) {
  if (slots_per_device == 0) {
    const char *e = getenv("CUDASP_SLOTS");                   // tuning aid
    slots_per_device = e && atoi(e) > 0 ? (uint32_t)atoi(e) : SLOTS_PER_DEVICE;
  }
  if (slots_per_device > (uint32_t)MAX_SLOTS_PER_DEVICE) return fail(nullptr, CUDASP_ERR_ARG, "slots_per_device out of range");
  if (!out) return fail(nullptr, CUDASP_ERR_ARG, "out is NULL");
  *out = nullptr;
  if (!scan_private_key || !spend_public_key) return fail(nullptr, CUDASP_ERR_ARG, "key pointer is NULL");
  if (n_labels > CUDASP_MAX_LABELS) return fail(nullptr, CUDASP_ERR_ARG, "too many label keys");
  if (n_labels && !label_keys) return fail(nullptr, CUDASP_ERR_ARG, "label_keys is NULL");
  if (batch_rows == 0 || batch_rows > CUDASP_MAX_BATCH_ROWS) return fail(nullptr, CUDASP_ERR_ARG, "batch_rows out of range");
  const int visible = cudasp_device_count();
  if (visible <= 0) return fail(nullptr, CUDASP_ERR_CUDA, "no CUDA device available (this library has no CPU path)");
  std::vector<int> devs;
  if (n_devices <= 0) n_devices = devices ? 0 : visible;
  for (int i = 0; i < n_devices; i++) {
    int d = devices ? devices[i] : i;
    if (d < 0 || d >= visible || d >= MAX_DEVICES) return fail(nullptr, CUDASP_ERR_ARG, "device ordinal out of range");
    devs.push_back(d);
  }
  if (devs.empty()) return fail(nullptr, CUDASP_ERR_ARG, "empty device list");

  cudasp_query *q = new cudasp_query();
  std::string err;
  int rc = host::make_query_plan(q->plan, scan_private_key, spend_public_key, label_keys, n_labels, err);
  if (rc) { delete q; return fail(nullptr, rc == -4 ? CUDASP_ERR_KEY : CUDASP_ERR_ARG, err); }
  q->n_labels = n_labels;
  q->batch_rows = batch_rows;
  if (max_outputs_per_batch == 0) max_outputs_per_batch = 4ull * batch_rows;
  q->devs.resize(devs.size());
  for (size_t i = 0; i < devs.size(); i++) {
    q->devs[i].device = devs[i];
    q->devs[i].max_slots = slots_per_device;
  }
  q->max_outputs = max_outputs_per_batch;
  {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, devs[0]);
    const char *pw = getenv("CUDASP_PIECE_WAVES");            // tuning aid
    const uint32_t waves = pw && atoi(pw) > 0 ? (uint32_t)atoi(pw) : SYNC_PIECE_WAVES;
    q->wave_threads = (uint32_t)sms * 4u * 128u;
    q->piece_rows = q->wave_threads * waves;
  }
  q->all_slots.reserve(devs.size() * slots_per_device);   // never reallocates: ids stay valid without the lock
  // Only the FIRST device is brought up here (context, fixed-base table, one slot), so that key and device errors surface
  // at creation; the others come up when a slot on them is first acquired (try_acquire): a query of a few rows -- every
  // statement of test/sql/cudasp.test -- costs one device's bring-up, not all of them (the reference's once-per-device
  // init is 16 KB, gsv.h:1094-1122; this build's is a 2.95 GB table, so it must not be paid where it is not used).
  {
    Slot *s = nullptr;
    rc = materialise_slot(q, 0, &s);
    if (rc) { std::string m = tl_err; cudasp_query_destroy(q); return fail(nullptr, rc, m); }
    s->id = (uint32_t)q->all_slots.size();
    q->devs[0].slots.push_back(s);
    q->all_slots.push_back(s);
  }
  *out = q;
  return CUDASP_OK;
}

int cudasp_device_sm_count(int device) {
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) { cudaGetLastError(); return 0; }
  return sms;
}

void cudasp_release_cached_memory(void) {
  std::vector<std::vector<Slot *>> old;
  {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    old.swap(g_cache);
  }
  for (size_t d = 0; d < old.size(); d++) {
    if (old[d].empty()) continue;
    cudaSetDevice((int)d);
    for (Slot *s : old[d]) { free_slot(*s); delete s; }
  }
}

uint64_t cudasp_launch_count(const cudasp_query *q) { return q ? q->launches.load() : 0; }
void cudasp_transfer_bytes(const cudasp_query *q, uint64_t *h2d, uint64_t *d2h) {
  if (h2d) *h2d = q ? q->h2d_bytes.load() : 0;
  if (d2h) *d2h = q ? q->d2h_bytes.load() : 0;
}

int cudasp_query_info(const cudasp_query *q, int *n_devices, int *ladder_dbl, int *ladder_add, int *fixed_window_bits,
                      int *n_labels) {
  if (!q) return CUDASP_ERR_ARG;
  if (n_devices) *n_devices = (int)q->devs.size();
  if (ladder_dbl) *ladder_dbl = q->plan.n_dbl;
  if (ladder_add) *ladder_add = q->plan.n_add;
  if (fixed_window_bits) *fixed_window_bits = GTAB_WBITS;
  if (n_labels) *n_labels = (int)q->n_labels;
  return CUDASP_OK;
}

}  // extern "C"

namespace {

// ---- slot pool ------------------------------------------------------------------------
// A free slot, devices taken in turn; when every materialised slot is busy and a device is below its
// limit, a new slot is materialised (cache or allocation) OUTSIDE the lock.  nullptr = all busy, at the limit.
Slot *try_acquire(cudasp_query *q, int *err) {
  *err = 0;
  int grow_dev = -1;
  {
    std::lock_guard<std::mutex> lk(q->mu);
    const size_t nd = q->devs.size();
    for (size_t k = 0; k < nd; k++) {
      DevCtx &dc = q->devs[(q->rr + k) % nd];
      for (Slot *s : dc.slots) {
        if (s->state == 0) {
          s->state = 1;
          q->rr = (uint32_t)((q->rr + k + 1) % nd);
          return s;
        }
      }
    }
    // grow where the fewest slots are; a device that is not up yet counts as if it had two, so that a caller
    // that never needs more than two batches in flight stays on the devices already paid for
    size_t best = SIZE_MAX;
    for (size_t k = 0; k < nd; k++) {
      DevCtx &dc = q->devs[(q->rr + k) % nd];
      const size_t have = dc.slots.size() + dc.pending;
      const size_t key = have + ((dc.up || have) ? 0 : 2);
      if (have < dc.max_slots && key < best) { best = key; grow_dev = (int)((q->rr + k) % nd); }
    }
    if (grow_dev < 0) return nullptr;
    q->devs[grow_dev].pending++;
  }
  Slot *s = nullptr;
  int rc = materialise_slot(q, grow_dev, &s);
  std::lock_guard<std::mutex> lk(q->mu);
  q->devs[grow_dev].pending--;
  if (rc) { *err = rc; return nullptr; }
  s->id = (uint32_t)q->all_slots.size();
  s->state = 1;
  q->devs[grow_dev].slots.push_back(s);
  q->all_slots.push_back(s);
  q->rr = (uint32_t)((grow_dev + 1) % q->devs.size());
  return s;
}
Slot *acquire_blocking(cudasp_query *q, int *err) {
  for (;;) {
    Slot *s = try_acquire(q, err);
    if (s || *err) return s;
    std::unique_lock<std::mutex> lk(q->mu);
    bool any_free = false;
    for (Slot *t : q->all_slots) any_free |= (t->state == 0);
    if (!any_free) q->cv.wait(lk);
  }
}
// at least min(n, max_slots) slots materialised on device index di (their workspaces are what scan_device borrows)
int ensure_slots(cudasp_query *q, int di, uint32_t n) {
  for (;;) {
    {
      std::lock_guard<std::mutex> lk(q->mu);
      DevCtx &dc = q->devs[di];
      if (dc.slots.size() + dc.pending >= std::min<uint32_t>(n, dc.max_slots)) return 0;
      dc.pending++;
    }
    Slot *s = nullptr;
    int rc = materialise_slot(q, di, &s);
    std::lock_guard<std::mutex> lk(q->mu);
    q->devs[di].pending--;
    if (rc) return rc;
    s->id = (uint32_t)q->all_slots.size();
    q->devs[di].slots.push_back(s);
    q->all_slots.push_back(s);
  }
}
Slot *slot_by_id(cudasp_query *q, uint32_t id) {
  std::lock_guard<std::mutex> lk(q->mu);
  return id < q->all_slots.size() ? q->all_slots[id] : nullptr;
}
void release_slot(cudasp_query *q, Slot *s) {
  {
    std::lock_guard<std::mutex> lk(q->mu);
    s->state = 0;
    s->internal = false;
  }
  q->cv.notify_all();
}

int submit(cudasp_query *q, Slot *s, uint32_t rows, uint64_t n_outputs, uint64_t tag, int keyfmt = CUDASP_KEYS_XY64) {
  DevCtx &dc = q->devs[s->dev_index];
  CU(q, cudaSetDevice(dc.device));
  cudaStream_t st = s->stream;
  s->rows = rows;
  s->tag = tag;
  static const bool trace = getenv("CUDASP_TRACE") != nullptr;
  // An event recorded on the slot's stream BEFORE its H2D copies makes the whole host-buffer call 7 % faster on B200
  // (cudasp_scan_batch over 9.4 M rows: 158.5 -> 147.7 ms, 59.3 -> 63.6 M rows/s; the events after the copies and after the
  // kernels change nothing -- profiles/r02e_submit_events.jsonl).  It used to be recorded only when tracing; now always
  // (CUDASP_SUBMIT_EVENTS=0 switches it off for the comparison).
  static const bool lead_event = [] { const char *e = getenv("CUDASP_SUBMIT_EVENTS"); return !e || (atoi(e) & 1); }();
  if (lead_event || trace) cudaEventRecord(s->tr[0], st);
  if (rows) {
    CU(q, cudaMemcpyAsync(s->d_tweaks, s->h_tweaks, (size_t)rows * key_bytes(keyfmt), cudaMemcpyHostToDevice, st));
    CU(q, cudaMemcpyAsync(s->d_off, s->h_off, ((size_t)rows + 1) * 4, cudaMemcpyHostToDevice, st));
    if (n_outputs) CU(q, cudaMemcpyAsync(s->d_outputs, s->h_outputs, n_outputs * 8, cudaMemcpyHostToDevice, st));
    q->h2d_bytes += (uint64_t)rows * key_bytes(keyfmt) + ((uint64_t)rows + 1) * 4 + n_outputs * 8;
  }
  if (trace) cudaEventRecord(s->tr[1], st);
  std::lock_guard<std::mutex> launch_lk(*dc.launch_mu);     // wait -> launches -> record is atomic per device (see DevCtx)
  CU(q, cudaStreamWaitEvent(st, s->ws.free_evt, 0));
  DetailSink sink{};
  const bool details = q->details_cap != 0;
  s->n_det = 0;
  s->det_truncated = false;
  if (details) {
    int r = ensure_details(q, *s);
    if (r) return r;
    sink = DetailSink{s->keep_sx, s->keep_sodd, s->d_det, q->details_cap};   // a cached slot may hold more
  }
  int rc = launch_pipeline(q, dc, s->ws, s->d_tweaks, s->d_outputs, s->d_off, rows, s->d_result + 1, s->d_result, st,
                           nullptr, nullptr, nullptr, nullptr, nullptr, keyfmt, details ? &sink : nullptr);
  if (rc) return rc;
  if (details)
    CU(q, cudaMemcpyAsync(s->h_det, s->d_det, ((size_t)std::min(q->details_cap, EAGER_DETAILS) + 1) * sizeof(MatchDetail),
                          cudaMemcpyDeviceToHost, st));
  if (trace) cudaEventRecord(s->tr[2], st);
  CU(q, cudaEventRecord(s->ws.free_evt, st));
  const uint32_t eager = std::min<uint32_t>(rows, EAGER_IDS);
  CU(q, cudaMemcpyAsync(s->h_result, s->d_result, ((size_t)eager + 1) * 4, cudaMemcpyDeviceToHost, st));
  q->d2h_bytes += ((uint64_t)eager + 1) * 4 + (details ? ((uint64_t)std::min(q->details_cap, EAGER_DETAILS) + 1) * sizeof(MatchDetail) : 0);
  CU(q, cudaEventRecord(s->done, st));
  return 0;
}

// wait for the slot's batch, bring back all ids, sort them ascending
int finish(cudasp_query *q, Slot *s, uint32_t *n_matches) {
  DevCtx &dc = q->devs[s->dev_index];
  CU(q, cudaSetDevice(dc.device));
  CU(q, cudaEventSynchronize(s->done));
  static const bool trace = getenv("CUDASP_TRACE") != nullptr && getenv("CUDASP_TRACE")[0] == '2';
  if (trace) {
    float a = 0, b = 0;
    cudaEventElapsedTime(&a, s->tr[0], s->tr[1]);
    cudaEventElapsedTime(&b, s->tr[1], s->tr[2]);
    fprintf(stderr, "[cudasp] slot %u: %u rows, H2D %.3f ms, kernels %.3f ms\n", s->id, s->rows, a, b);
  }
  uint32_t n = s->h_result[0];
  if (n > s->rows) return fail(q, CUDASP_ERR_STATE, "internal: match count exceeds rows");
  if (n > EAGER_IDS) {
    CU(q, cudaMemcpyAsync(s->h_result + 1 + EAGER_IDS, s->d_result + 1 + EAGER_IDS, ((size_t)n - EAGER_IDS) * 4,
                          cudaMemcpyDeviceToHost, s->stream));
    CU(q, cudaStreamSynchronize(s->stream));
    q->d2h_bytes += ((uint64_t)n - EAGER_IDS) * 4;
  }
  std::sort(s->h_result + 1, s->h_result + 1 + n);
  *n_matches = n;
  if (q->details_cap && s->h_det && s->rows) {
    uint32_t nd = *reinterpret_cast<const uint32_t *>(s->h_det);
    // K8 counts every record but stores only the first details_cap: the matches themselves are complete either way, so the
    // batch is delivered with the records that fit and the caller is told (CUDASP_RESULT_DETAILS_TRUNCATED)
    s->det_truncated = nd > q->details_cap;
    nd = std::min(nd, q->details_cap);
    if (nd > EAGER_DETAILS) {
      CU(q, cudaMemcpyAsync(s->h_det + 1 + EAGER_DETAILS, s->d_det + 1 + EAGER_DETAILS, ((size_t)nd - EAGER_DETAILS) * sizeof(MatchDetail),
                            cudaMemcpyDeviceToHost, s->stream));
      CU(q, cudaStreamSynchronize(s->stream));
    }
    std::sort(s->h_det + 1, s->h_det + 1 + nd, [](const MatchDetail &a, const MatchDetail &b) {
      return a.row != b.row ? a.row < b.row : a.k < b.k;
    });
    s->n_det = nd;
  }
  return 0;
}

}  // namespace

// ---- internal staging interface for host-only ingest front-ends (staging_internal.hpp) --------
namespace csp {
namespace staging {
static void view(Staging *v, Slot *s) {
  v->slot = s;
  v->tweaks = s->h_tweaks;
  v->outputs = s->h_outputs;
  v->offsets = s->h_off;
  v->ids = s->h_result + 1;
  v->cap_rows = s->cap_rows;
  v->cap_outputs = s->cap_outputs;
}
int acquire(cudasp_query *q, Staging *v, bool block) {
  int rc = 0;
  Slot *s = block ? acquire_blocking(q, &rc) : try_acquire(q, &rc);
  if (!s) return rc;
  s->internal = true;
  view(v, s);
  return 1;
}
int reserve_outputs(cudasp_query *q, Staging *v, uint64_t n) {
  Slot *s = static_cast<Slot *>(v->slot);
  if (n <= s->cap_outputs) return 0;
  cudaSetDevice(q->devs[s->dev_index].device);
  cudaStreamSynchronize(s->stream);
  int rc = alloc_slot_outputs(q, *s, n + n / 4);
  view(v, s);
  return rc;
}
int submit(cudasp_query *q, Staging *v, uint32_t rows, uint64_t n_outputs, int key_format) {
  return ::submit(q, static_cast<Slot *>(v->slot), rows, n_outputs, 0, key_format);
}
int finish(cudasp_query *q, Staging *v, uint32_t *n) { return ::finish(q, static_cast<Slot *>(v->slot), n); }
void release(cudasp_query *q, Staging *v) {
  if (v->slot) release_slot(q, static_cast<Slot *>(v->slot));
  v->slot = nullptr;
}
uint32_t piece_rows(cudasp_query *q) { return std::min<uint32_t>(q->batch_rows, q->piece_rows); }
int device_count(cudasp_query *q) { return (int)q->devs.size(); }
int fail(cudasp_query *q, int code, const char *msg) { return ::fail(q, code, msg); }
}  // namespace staging
}  // namespace csp

// =====================================================================================
extern "C" {

// ---- streaming API --------------------------------------------------------------------
int cudasp_acquire_slot(cudasp_query *q, cudasp_slot *slot) {
  if (!q || !slot) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  int err = 0;
  Slot *s = acquire_blocking(q, &err);
  if (!s) return err;
  slot->slot_id = s->id;
  slot->device = q->devs[s->dev_index].device;
  slot->tweak_keys = s->h_tweaks;
  slot->outputs = s->h_outputs;
  slot->out_offsets = s->h_off;
  slot->cap_rows = s->cap_rows;
  slot->cap_outputs = s->cap_outputs;
  return CUDASP_OK;
}

int cudasp_try_acquire_slot(cudasp_query *q, cudasp_slot *slot) {
  if (!q || !slot) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  int err = 0;
  Slot *s = try_acquire(q, &err);
  if (err) return err;
  if (!s) return 0;
  slot->slot_id = s->id;
  slot->device = q->devs[s->dev_index].device;
  slot->tweak_keys = s->h_tweaks;
  slot->outputs = s->h_outputs;
  slot->out_offsets = s->h_off;
  slot->cap_rows = s->cap_rows;
  slot->cap_outputs = s->cap_outputs;
  return 1;
}

int cudasp_submit_slot(cudasp_query *q, const cudasp_slot *slot, uint32_t rows, uint64_t n_outputs, uint64_t tag) {
  return cudasp_submit_slot_fmt(q, slot, rows, n_outputs, tag, CUDASP_KEYS_XY64);
}

int cudasp_submit_slot_fmt(cudasp_query *q, const cudasp_slot *slot, uint32_t rows, uint64_t n_outputs, uint64_t tag, int key_format) {
  if (key_format != CUDASP_KEYS_XY64 && key_format != CUDASP_KEYS_SEC1_33) return fail(q, CUDASP_ERR_ARG, "unknown key format");
  if (!q || !slot) return fail(q, CUDASP_ERR_ARG, "bad slot");
  Slot *s = slot_by_id(q, slot->slot_id);
  if (!s) return fail(q, CUDASP_ERR_ARG, "bad slot");
  if (s->state != 1) return fail(q, CUDASP_ERR_STATE, "slot was not acquired");
  if (rows > s->cap_rows || n_outputs > s->cap_outputs) return fail(q, CUDASP_ERR_ARG, "rows/outputs exceed the slot capacity");
  if (rows && s->h_off[rows] != n_outputs) return fail(q, CUDASP_ERR_ARG, "out_offsets[rows] != n_outputs");
  if (rows) {      // the kernels index outputs[] with these: zero-based and monotonic, or the batch is refused (1.2 MB pass)
    uint32_t prev = s->h_off[0], bad = prev;
    for (uint32_t i = 1; i <= rows; i++) { const uint32_t v = s->h_off[i]; bad |= (uint32_t)(v < prev); prev = v; }
    if (bad) return fail(q, CUDASP_ERR_ARG, "out_offsets must start at 0 and be monotonically non-decreasing");
  }
  int rc = submit(q, s, rows, n_outputs, tag, key_format);
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(q->mu);
  s->state = 2;
  q->inflight.push_back(s);
  return CUDASP_OK;
}

// Hands out a finished batch.  ANY in-flight batch whose `done` event has fired qualifies, not only the oldest: with
// several devices (or several streams of one device) a finished batch queued behind an unfinished one would otherwise
// keep its slot from being recycled (head-of-line blocking).  Blocking mode falls back to waiting for the oldest.
// On an error after the batch left the in-flight list, *r still names the slot (tag, slot_id, rows; no matches) so
// that the caller can release it -- otherwise the slot would be lost and a later acquire could block forever.
static int take_result(cudasp_query *q, cudasp_result *r, bool block) {
  if (!q || !r) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  Slot *s = nullptr;
  {
    std::lock_guard<std::mutex> lk(q->mu);
    if (q->inflight.empty()) return 0;
    // per device only the OLDEST batch is asked (batches of one device finish roughly in submit order): at most one
    // event query per device under the lock, however many batches are in flight
    uint64_t seen_devices = 0;
    size_t n_seen = 0;
    for (auto it = q->inflight.begin(); it != q->inflight.end() && n_seen < q->devs.size(); ++it) {
      const uint64_t bit = 1ull << ((*it)->dev_index & 63);
      if (seen_devices & bit) continue;
      seen_devices |= bit;
      n_seen++;
      const cudaError_t e = cudaEventQuery((*it)->done);
      if (e == cudaErrorNotReady) continue;
      if (e != cudaSuccess) return fail_locked(q, CUDASP_ERR_CUDA, cudaGetErrorString(e));
      s = *it;
      q->inflight.erase(it);
      break;
    }
    if (!s) {
      if (!block) return 0;
      s = q->inflight.front();
      q->inflight.pop_front();
    }
    s->state = 3;
  }
  uint32_t n = 0;
  const int rc = finish(q, s, &n);
  r->tag = s->tag;
  r->slot_id = s->id;
  r->rows = s->rows;
  r->flags = s->det_truncated ? CUDASP_RESULT_DETAILS_TRUNCATED : 0u;
  r->n_matches = rc ? 0 : n;
  r->match_ids = rc ? nullptr : s->h_result + 1;
  return rc ? rc : 1;
}
int cudasp_poll(cudasp_query *q, cudasp_result *r) { return take_result(q, r, false); }
int cudasp_wait(cudasp_query *q, cudasp_result *r) { return take_result(q, r, true); }

int cudasp_release_result(cudasp_query *q, const cudasp_result *r) {
  if (!q || !r) return fail(q, CUDASP_ERR_ARG, "bad result");
  Slot *s = slot_by_id(q, r->slot_id);
  if (!s) return fail(q, CUDASP_ERR_ARG, "bad result");
  if (s->state != 3) return fail(q, CUDASP_ERR_STATE, "result was not handed out");
  release_slot(q, s);
  return CUDASP_OK;
}

// ---- synchronous batch ----------------------------------------------------------------
int cudasp_scan_batch(cudasp_query *q, const uint8_t *tweak_keys, const int64_t *outputs, const uint32_t *out_offsets,
                      uint32_t rows, uint32_t *match_ids, uint32_t *n_matches) {
  return cudasp_scan_batch_fmt(q, tweak_keys, CUDASP_KEYS_XY64, outputs, out_offsets, rows, match_ids, n_matches);
}

static int scan_batch_impl(cudasp_query *q, const uint8_t *tweak_keys, int key_format, const int64_t *outputs,
                           const uint32_t *out_offsets, uint32_t rows, uint32_t *match_ids, uint32_t *n_matches,
                           cudasp_match_detail *details, uint32_t details_capacity, uint32_t *n_details, int host_threads);

int cudasp_scan_batch_fmt(cudasp_query *q, const uint8_t *tweak_keys, int key_format, const int64_t *outputs,
                          const uint32_t *out_offsets, uint32_t rows, uint32_t *match_ids, uint32_t *n_matches) {
  return scan_batch_impl(q, tweak_keys, key_format, outputs, out_offsets, rows, match_ids, n_matches, nullptr, 0, nullptr, 0);
}

int cudasp_scan_batch_details(cudasp_query *q, const uint8_t *tweak_keys, int key_format, const int64_t *outputs,
                              const uint32_t *out_offsets, uint32_t rows, uint32_t *match_ids, uint32_t *n_matches,
                              cudasp_match_detail *details, uint32_t details_capacity, uint32_t *n_details) {
  if (!q || !n_details || (details_capacity && !details)) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  if (!q->details_cap) return fail(q, CUDASP_ERR_STATE, "match details are not enabled: call cudasp_query_enable_details first");
  return scan_batch_impl(q, tweak_keys, key_format, outputs, out_offsets, rows, match_ids, n_matches, details, details_capacity,
                         n_details, 0);
}

int cudasp_query_enable_details(cudasp_query *q, uint32_t max_details_per_batch) {
  if (!q) return fail(q, CUDASP_ERR_ARG, "NULL query");
  std::lock_guard<std::mutex> lk(q->mu);
  for (Slot *s : q->all_slots)
    if (s->state != 0) return fail_locked(q, CUDASP_ERR_STATE, "enable details before the first batch is acquired");
  q->details_cap = max_details_per_batch ? max_details_per_batch : std::min<uint32_t>(q->batch_rows, 65536u);
  return CUDASP_OK;
}

int cudasp_result_details(cudasp_query *q, const cudasp_result *r, const cudasp_match_detail **details, uint32_t *n_details) {
  if (!q || !r || !details || !n_details) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  Slot *s = slot_by_id(q, r->slot_id);
  if (!s || s->state != 3) return fail(q, CUDASP_ERR_STATE, "result was not handed out");
  if (!q->details_cap) return fail(q, CUDASP_ERR_STATE, "match details are not enabled: call cudasp_query_enable_details first");
  *details = s->n_det ? reinterpret_cast<const cudasp_match_detail *>(s->h_det + 1) : nullptr;
  *n_details = s->n_det;
  return CUDASP_OK;
}

// The synchronous call: a host table in, matching row ids out.  The table is cut into pieces -- at most `piece_rows` rows
// (whole K1 waves), ramping up over the first pieces (1/8, 1/4, 1/2: the GPUs start after a short first copy) and down over
// the last ones (the tail the GPUs finish alone after the last copy is short) -- and `host_threads` staging threads take
// pieces in order: acquire a slot (devices in turn), memcpy into its pinned buffers, submit, and collect their own oldest
// piece whenever no slot is free.  One thread cannot keep one B200 busy from pageable memory (a 64 B/row memcpy at ~5 GB/s is
// as slow as the scan itself); several threads over one handle that owns several devices are the in-process form of the
// reference's "DuckDB workers round-robin onto GPUs" (src/cudasp_extension.cu:516-532).
static int scan_batch_impl(cudasp_query *q, const uint8_t *tweak_keys, int key_format, const int64_t *outputs,
                           const uint32_t *out_offsets, uint32_t rows, uint32_t *match_ids, uint32_t *n_matches,
                           cudasp_match_detail *details, uint32_t details_capacity, uint32_t *n_details, int host_threads) {
  if (n_details) *n_details = 0;
  if (key_format != CUDASP_KEYS_XY64 && key_format != CUDASP_KEYS_SEC1_33) return fail(q, CUDASP_ERR_ARG, "unknown key format");
  const size_t kb = key_bytes(key_format);
  if (!q || !n_matches) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  *n_matches = 0;
  if (rows == 0) return CUDASP_OK;
  if (!tweak_keys || !match_ids) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  if (out_offsets && !outputs && out_offsets[rows] != out_offsets[0]) return fail(q, CUDASP_ERR_ARG, "outputs is NULL");
  static const bool trace = getenv("CUDASP_TRACE") != nullptr;     // coarse host-side timings to stderr
  auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  const double t_begin = trace ? now() : 0;

  // ---- the piece list
  const uint32_t piece = std::min<uint32_t>(q->batch_rows, q->piece_rows);
  const uint32_t nd = (uint32_t)q->devs.size();
  const std::vector<uint32_t> starts = piece_plan(rows, piece, nd, RAMP_DIV);      // piece i = rows [starts[i], starts[i + 1])
  const uint32_t n_pieces = (uint32_t)starts.size() - 1;
  if (host_threads <= 0) {      // auto: enough copy threads to feed the devices, not more than there are pieces or cores
    const char *e = getenv("CUDASP_HOST_THREADS");
    host_threads = e && atoi(e) > 0 ? atoi(e) : (int)std::min<uint32_t>(4u * nd, std::max(1u, std::thread::hardware_concurrency() / 2));
  }
  host_threads = (int)std::max<uint32_t>(1u, std::min<uint32_t>((uint32_t)host_threads, std::min<uint32_t>(n_pieces / 2, 64u)));

  struct Pending { Slot *s; uint32_t base; };
  struct Worker {
    std::deque<Pending> pending;
    std::vector<uint32_t> ids;
    std::vector<cudasp_match_detail> det;
    uint64_t n_det = 0;          // found (may exceed what is kept when the caller's capacity is smaller)
    int rc = 0;
    double t_copy = 0, t_submit = 0, t_wait = 0;
  };
  std::vector<Worker> workers((size_t)host_threads);
  std::atomic<uint32_t> next_piece{0};
  std::atomic<bool> failed{false};

  auto collect = [&](Worker &w, Pending p) -> int {
    uint32_t n = 0;
    const double tw = trace ? now() : 0;
    int r = finish(q, p.s, &n);
    if (trace) w.t_wait += now() - tw;
    if (!r && p.s->det_truncated) r = fail(q, CUDASP_ERR_STATE, "more match details than max_details_per_batch: raise it");
    if (!r) {
      for (uint32_t i = 0; i < n; i++) w.ids.push_back(p.base + p.s->h_result[1 + i]);
      if (n_details) {
        for (uint32_t i = 0; i < p.s->n_det; i++, w.n_det++) {
          if (w.det.size() >= details_capacity) continue;       // counted, not stored
          cudasp_match_detail d;
          memcpy(&d, &p.s->h_det[1 + i], sizeof d);
          d.row += p.base;
          w.det.push_back(d);
        }
      }
    }
    release_slot(q, p.s);
    return r;
  };
  auto run = [&](Worker &w) {
    int rc = 0;
    while (!rc && !failed.load(std::memory_order_relaxed)) {
      const uint32_t pi = next_piece.fetch_add(1);
      if (pi >= n_pieces) break;
      const uint32_t base = starts[pi], cnt = starts[pi + 1] - base;
      Slot *s = nullptr;
      while (!s) {
        s = try_acquire(q, &rc);
        if (s || rc) break;
        if (!w.pending.empty()) {          // recycle our own oldest piece
          Pending p = w.pending.front();
          w.pending.pop_front();
          rc = collect(w, p);
          if (rc) break;
        } else {
          s = acquire_blocking(q, &rc);
          if (!s) break;
        }
      }
      if (rc || !s) break;
      s->internal = true;
      const uint32_t o0 = out_offsets ? out_offsets[base] : 0;
      if (out_offsets && out_offsets[base + cnt] < o0) {
        release_slot(q, s);
        rc = fail(q, CUDASP_ERR_ARG, "out_offsets is not monotonically non-decreasing");
        break;
      }
      const uint64_t no = out_offsets ? (uint64_t)(out_offsets[base + cnt] - o0) : 0;
      if (no > s->cap_outputs) {
        cudaSetDevice(q->devs[s->dev_index].device);
        cudaStreamSynchronize(s->stream);
        rc = alloc_slot_outputs(q, *s, no + no / 4);
        if (rc) { release_slot(q, s); break; }
      }
      const double tc = trace ? now() : 0;
      memcpy(s->h_tweaks, tweak_keys + (size_t)base * kb, (size_t)cnt * kb);
      if (out_offsets) {
        // rebase, and refuse a non-monotonic offset array: K5/K7/K8 index the outputs with these, unchecked
        uint32_t prev = 0, bad = 0;
        for (uint32_t i = 0; i <= cnt; i++) {
          const uint32_t v = out_offsets[base + i] - o0;
          bad |= (uint32_t)(v < prev);
          s->h_off[i] = prev = v;
        }
        if (bad) { release_slot(q, s); rc = fail(q, CUDASP_ERR_ARG, "out_offsets is not monotonically non-decreasing"); break; }
        if (no) memcpy(s->h_outputs, outputs + o0, no * 8);
      } else {
        memset(s->h_off, 0, ((size_t)cnt + 1) * 4);
      }
      const double ts = trace ? now() : 0;
      rc = submit(q, s, cnt, no, 0, key_format);
      if (trace) { w.t_copy += ts - tc; w.t_submit += now() - ts; }
      if (rc) { release_slot(q, s); break; }
      w.pending.push_back({s, base});
    }
    if (rc) failed.store(true);
    while (!w.pending.empty()) {
      Pending p = w.pending.front();
      w.pending.pop_front();
      int r = collect(w, p);
      if (!rc) rc = r;
    }
    w.rc = rc;
  };
  if (host_threads == 1) {
    run(workers[0]);
  } else {
    std::vector<std::thread> th;
    for (int t = 1; t < host_threads; t++) th.emplace_back([&, t] { run(workers[(size_t)t]); });
    run(workers[0]);
    for (auto &t : th) t.join();
  }
  int rc = 0;
  uint32_t total = 0;
  uint64_t total_det = 0;
  for (Worker &w : workers) {
    if (w.rc && !rc) rc = w.rc;
    if (total + w.ids.size() <= rows) memcpy(match_ids + total, w.ids.data(), w.ids.size() * 4);
    total += (uint32_t)w.ids.size();
    total_det += w.n_det;
  }
  if (!rc && total > rows) rc = fail(q, CUDASP_ERR_STATE, "internal: more matches than rows");
  if (!rc) std::sort(match_ids, match_ids + total);      // pieces are disjoint row ranges: ascending ids over the whole call
  *n_matches = rc ? 0 : total;
  if (n_details) {
    std::vector<cudasp_match_detail> all;
    for (Worker &w : workers) all.insert(all.end(), w.det.begin(), w.det.end());
    std::sort(all.begin(), all.end(), [](const cudasp_match_detail &a, const cudasp_match_detail &b) {
      return a.row != b.row ? a.row < b.row : a.k < b.k;
    });
    const size_t keep = std::min<size_t>(all.size(), details_capacity);
    if (keep) memcpy(details, all.data(), keep * sizeof(cudasp_match_detail));
    *n_details = (uint32_t)std::min<uint64_t>(total_det, 0xffffffffull);
    if (!rc && total_det > details_capacity) rc = fail(q, CUDASP_ERR_ARG, "details capacity too small (n_details holds the number found)");
  }
  if (trace) {
    double c = 0, sb = 0, wt = 0;
    for (Worker &w : workers) { c += w.t_copy; sb += w.t_submit; wt += w.t_wait; }
    fprintf(stderr, "[cudasp] scan_batch %u rows, %u pieces, %d host thread(s): total %.3f ms; summed over threads: host copy %.3f, "
                    "submit calls %.3f, waiting for the GPU %.3f\n", rows, n_pieces, host_threads, 1e3 * (now() - t_begin), 1e3 * c,
            1e3 * sb, 1e3 * wt);
  }
  return rc;
}

int cudasp_scan_batch_mt(cudasp_query *q, const uint8_t *tweak_keys, int key_format, const int64_t *outputs,
                         const uint32_t *out_offsets, uint32_t rows, uint32_t *match_ids, uint32_t *n_matches, int host_threads) {
  return scan_batch_impl(q, tweak_keys, key_format, outputs, out_offsets, rows, match_ids, n_matches, nullptr, 0, nullptr, host_threads);
}

// ---- device-resident entry ------------------------------------------------------------
int cudasp_scan_device(cudasp_query *q, int device, const uint8_t *d_tweak_keys, const int64_t *d_outputs,
                       const uint32_t *d_out_offsets, uint32_t rows, uint32_t *d_match_ids, uint32_t *d_n_matches,
                       void *stream) {
  if (!q) return fail(q, CUDASP_ERR_ARG, "NULL query");
  DevCtx *dc = nullptr;
  for (DevCtx &d : q->devs) if (d.device == device) dc = &d;
  if (!dc) return fail(q, CUDASP_ERR_ARG, "device is not part of this query");
  if (rows > q->batch_rows) return fail(q, CUDASP_ERR_ARG, "rows exceed batch_rows (split the call)");
  if (rows && (!d_tweak_keys || !d_match_ids || !d_n_matches)) return fail(q, CUDASP_ERR_ARG, "NULL device pointer");
  // borrow the scratch of one of the device's slots, round robin; ordering between users of the
  // same scratch is enforced with its free_evt (under the device's launch lock), so any stream may be used.
  int rc0 = ensure_slots(q, (int)(dc - q->devs.data()), 3);   // also brings the device up
  if (rc0) return rc0;
  CU(q, cudaSetDevice(device));
  cudaStream_t st = stream ? (cudaStream_t)stream : dc->own_stream;
  Workspace *w;
  {
    std::lock_guard<std::mutex> lk(q->mu);
    w = &dc->slots[dc->next_ws++ % dc->slots.size()]->ws;
  }
  std::lock_guard<std::mutex> launch_lk(*dc->launch_mu);
  CU(q, cudaStreamWaitEvent(st, w->free_evt, 0));
  int rc = launch_pipeline(q, *dc, *w, d_tweak_keys, d_outputs, d_out_offsets, rows, d_match_ids, d_n_matches, st, nullptr,
                           nullptr, nullptr, nullptr);
  if (rc) return rc;
  CU(q, cudaEventRecord(w->free_evt, st));
  return CUDASP_OK;
}

// ---- per-stage timing (bench.py roofline) -----------------------------------------------
int cudasp_profile_stages(cudasp_query *q, int device, const uint8_t *d_tweak_keys, const int64_t *d_outputs,
                          const uint32_t *d_out_offsets, uint32_t rows, uint32_t *d_match_ids, uint32_t *d_n_matches,
                          float *stage_ms) {
  if (!q || !stage_ms) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  DevCtx *dc = nullptr;
  for (DevCtx &d : q->devs) if (d.device == device) dc = &d;
  if (!dc) return fail(q, CUDASP_ERR_ARG, "device is not part of this query");
  if (rows == 0 || rows > q->batch_rows) return fail(q, CUDASP_ERR_ARG, "rows out of range");
  CU(q, cudaSetDevice(device));
  cudaEvent_t ev[8];
  for (int i = 0; i < 8; i++) CU(q, cudaEventCreate(&ev[i]));
  int rc0 = ensure_slots(q, (int)(dc - q->devs.data()), 1);
  if (rc0) return rc0;
  Workspace &w = dc->slots[0]->ws;
  cudaStream_t st = dc->own_stream;
  std::unique_lock<std::mutex> launch_lk(*dc->launch_mu);
  CU(q, cudaStreamWaitEvent(st, w.free_evt, 0));
  const char *kf = getenv("CUDASP_PROFILE_KEYFMT");      // tuning aid: "1" = d_tweak_keys holds 33-byte SEC1 keys
  int rc = launch_pipeline(q, *dc, w, d_tweak_keys, d_outputs, d_out_offsets, rows, d_match_ids, d_n_matches, st, nullptr, nullptr,
                           nullptr, nullptr, ev, kf && kf[0] == '1' ? CUDASP_KEYS_SEC1_33 : CUDASP_KEYS_XY64);
  if (!rc) {
    cudaEventRecord(w.free_evt, st);
    launch_lk.unlock();
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) rc = fail(q, CUDASP_ERR_CUDA, cudaGetErrorString(e));
    for (int i = 0; i < 7 && !rc; i++) cudaEventElapsedTime(&stage_ms[i], ev[i], ev[i + 1]);
  }
  for (int i = 0; i < 8; i++) cudaEventDestroy(ev[i]);
  return rc;
}

// ---- introspection --------------------------------------------------------------------
static int scan_debug(cudasp_query *q, const uint8_t *tweak_keys, uint32_t rows, int64_t *cands, uint8_t *S_le, uint8_t *t_be,
                      uint8_t *P0_le, int keyfmt = CUDASP_KEYS_XY64, uint8_t *valid = nullptr) {
  if (!q || (rows && !tweak_keys)) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  const uint32_t L = q->n_labels;
  const size_t kb = key_bytes(keyfmt);
  for (uint32_t base = 0; base < rows; base += q->batch_rows) {
    const uint32_t cnt = std::min<uint32_t>(q->batch_rows, rows - base);
    int rc = 0;
    Slot *s = acquire_blocking(q, &rc);
    if (!s) return rc;
    DevCtx &dc = q->devs[s->dev_index];
    int64_t *d_c = nullptr;
    uint8_t *d_S = nullptr, *d_t = nullptr, *d_P = nullptr;
    auto body = [&]() -> int {
      CU(q, cudaSetDevice(dc.device));
      if (cands) CU(q, cudaMalloc(&d_c, (size_t)cnt * (1 + L) * 8));
      if (S_le) CU(q, cudaMalloc(&d_S, (size_t)cnt * 64));
      if (t_be) CU(q, cudaMalloc(&d_t, (size_t)cnt * 32));
      if (P0_le) CU(q, cudaMalloc(&d_P, (size_t)cnt * 64));
      CU(q, cudaMemcpyAsync(s->d_tweaks, tweak_keys + (size_t)base * kb, (size_t)cnt * kb, cudaMemcpyHostToDevice, s->stream));
      {
        std::lock_guard<std::mutex> launch_lk(*dc.launch_mu);
        CU(q, cudaStreamWaitEvent(s->stream, s->ws.free_evt, 0));
        int r = launch_pipeline(q, dc, s->ws, s->d_tweaks, nullptr, nullptr, cnt, nullptr, nullptr, s->stream, d_c, d_S, d_t, d_P, nullptr,
                                keyfmt);
        if (r) return r;
        CU(q, cudaEventRecord(s->ws.free_evt, s->stream));
      }
      if (cands) CU(q, cudaMemcpyAsync(cands + (size_t)base * (1 + L), d_c, (size_t)cnt * (1 + L) * 8, cudaMemcpyDeviceToHost, s->stream));
      if (S_le) CU(q, cudaMemcpyAsync(S_le + (size_t)base * 64, d_S, (size_t)cnt * 64, cudaMemcpyDeviceToHost, s->stream));
      if (t_be) CU(q, cudaMemcpyAsync(t_be + (size_t)base * 32, d_t, (size_t)cnt * 32, cudaMemcpyDeviceToHost, s->stream));
      if (P0_le) CU(q, cudaMemcpyAsync(P0_le + (size_t)base * 64, d_P, (size_t)cnt * 64, cudaMemcpyDeviceToHost, s->stream));
      if (valid) {
        if (keyfmt == CUDASP_KEYS_SEC1_33) CU(q, cudaMemcpyAsync(valid + base, s->ws.invalid, cnt, cudaMemcpyDeviceToHost, s->stream));
        else memset(valid + base, 0, cnt);
      }
      CU(q, cudaStreamSynchronize(s->stream));
      if (valid) for (uint32_t i = 0; i < cnt; i++) valid[base + i] = valid[base + i] ? 0 : 1;   // invalid flag -> valid flag
      return 0;
    };
    rc = body();
    cudaFree(d_c); cudaFree(d_S); cudaFree(d_t); cudaFree(d_P);
    release_slot(q, s);
    if (rc) return rc;
  }
  return CUDASP_OK;
}
int cudasp_scan_candidates(cudasp_query *q, const uint8_t *tweak_keys, uint32_t rows, int64_t *cands) {
  if (!cands) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  return scan_debug(q, tweak_keys, rows, cands, nullptr, nullptr, nullptr);
}
int cudasp_scan_candidates_fmt(cudasp_query *q, const uint8_t *tweak_keys, int key_format, uint32_t rows, int64_t *cands,
                               uint8_t *valid) {
  if (!cands) return fail(q, CUDASP_ERR_ARG, "NULL argument");
  if (key_format != CUDASP_KEYS_XY64 && key_format != CUDASP_KEYS_SEC1_33) return fail(q, CUDASP_ERR_ARG, "unknown key format");
  return scan_debug(q, tweak_keys, rows, cands, nullptr, nullptr, nullptr, key_format, valid);
}
int cudasp_scan_intermediates(cudasp_query *q, const uint8_t *tweak_keys, uint32_t rows, uint8_t *S_le, uint8_t *t_be,
                              uint8_t *P0_le) {
  return scan_debug(q, tweak_keys, rows, nullptr, S_le, t_be, P0_le);
}

int cudasp_debug_gtab(int device, uint32_t window, uint32_t first_digit, uint32_t n, uint8_t *out) {
  if (!out) return fail(nullptr, CUDASP_ERR_ARG, "NULL argument");
  if (window >= (uint32_t)GTAB_NWIN ||
      (uint64_t)first_digit + n > (uint64_t)(window == (uint32_t)GTAB_NWIN - 1 ? GTAB_TOP_ENTRIES : GTAB_ENTRIES))
    return fail(nullptr, CUDASP_ERR_ARG, "window / digit out of range");
  ge *tab = nullptr;
  int rc = get_device_gtab(nullptr, device, &tab);
  if (rc) return rc;
  if (n == 0) return CUDASP_OK;
  CU(nullptr, cudaSetDevice(device));
  CU(nullptr, cudaMemcpy(out, tab + (size_t)window * GTAB_ENTRIES + first_digit, (size_t)n * sizeof(ge), cudaMemcpyDeviceToHost));
  return CUDASP_OK;
}

int cudasp_test_field(int device, int op, const uint8_t *a, const uint8_t *b, uint8_t *out, uint32_t n) {
  if (!a || !b || !out) return fail(nullptr, CUDASP_ERR_ARG, "NULL argument");
  if (n == 0) return CUDASP_OK;
  fe *da = nullptr, *db = nullptr, *dout = nullptr;
  auto body = [&]() -> int {
    CU(nullptr, cudaSetDevice(device));
    CU(nullptr, cudaMalloc(&da, (size_t)n * 32));
    CU(nullptr, cudaMalloc(&db, (size_t)n * 32));
    CU(nullptr, cudaMalloc(&dout, (size_t)n * 32));
    CU(nullptr, cudaMemcpy(da, a, (size_t)n * 32, cudaMemcpyHostToDevice));
    CU(nullptr, cudaMemcpy(db, b, (size_t)n * 32, cudaMemcpyHostToDevice));
    CU(nullptr, launch::test_field(op, da, db, dout, n));
    CU(nullptr, cudaMemcpy(out, dout, (size_t)n * 32, cudaMemcpyDeviceToHost));
    return 0;
  };
  int rc = body();
  cudaFree(da); cudaFree(db); cudaFree(dout);
  return rc;
}

}  // extern "C"
