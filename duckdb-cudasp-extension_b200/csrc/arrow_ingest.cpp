// Arrow C-data-interface front-end of the cudasp_scan hot path (include/cudasp_arrow.h): host-only code that fills the
// pinned staging slots of the C-ABI core straight from Arrow buffers.  Replaces, for columnar callers, the reference's
// per-row AccumulateInput + ProcessBatch marshalling (/root/reference/src/cudasp_extension.cu:154-330).
//
// Two staging paths per piece (a piece = up to piece_rows rows, whole K1 waves):
//   dense   : no NULL keys / lists / children in sight and one key width -> memcpy(keys), rebase(offsets), memcpy(values)
//   general : row by row with the reference's NULL rules, rows compacted, a row map kept to translate ids back
// A piece is filled across pushes (record batches, Parquet pages): what a push leaves over stays in an open piece.
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <string>
#include <vector>

#include "../../include/cudasp_arrow.h"
#include "staging_internal.hpp"
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

namespace {

using csp::staging::Staging;
namespace stg = csp::staging;

inline bool bit_set(const uint8_t *bits, int64_t i) { return (bits[i >> 3] >> (i & 7)) & 1; }

struct KeyColumn {
  int width = 0;                    // > 0: fixed_size_binary(width); 0: variable (offsets)
  int64_t stride = 0;               // width > 0 only: bytes from one key to the next; 0 = width (an Arrow array).  Larger for keys
                                    // read where they lie in a Parquet page (length-prefixed values: stride = width + 4)
  bool large = false;               // variable: 64-bit offsets
  const uint8_t *validity = nullptr;
  const void *offsets = nullptr;
  const uint8_t *data = nullptr;
  int64_t offset = 0, length = 0;
  bool may_have_nulls = false;

  bool valid(int64_t i) const { return !validity || bit_set(validity, offset + i); }
  int64_t begin(int64_t i) const {
    if (width) return (offset + i) * (stride ? stride : width);
    return large ? static_cast<const int64_t *>(offsets)[offset + i] : static_cast<const int32_t *>(offsets)[offset + i];
  }
  int64_t size(int64_t i) const { return width ? width : begin(i + 1) - begin(i); }
};

struct ListColumn {
  bool large = false;
  const uint8_t *validity = nullptr;
  const void *offsets = nullptr;
  int64_t offset = 0, length = 0;
  bool may_have_nulls = false;
  // child
  const uint8_t *child_validity = nullptr;
  const int64_t *values = nullptr;
  int64_t child_offset = 0;
  bool child_may_have_nulls = false;

  bool valid(int64_t i) const { return !validity || bit_set(validity, offset + i); }
  int64_t begin(int64_t i) const {
    return large ? static_cast<const int64_t *>(offsets)[offset + i] : static_cast<const int32_t *>(offsets)[offset + i];
  }
  bool child_valid(int64_t j) const { return !child_validity || bit_set(child_validity, child_offset + j); }
};

// null_count may be -1 (not computed): then the bitmap decides
bool may_have_nulls(const ArrowArray *a) { return a->null_count != 0 && a->n_buffers > 0 && a->buffers[0] != nullptr; }

int parse_key_column(cudasp_query *q, const ArrowSchema *sc, const ArrowArray *a, KeyColumn *k) {
  if (!sc || !a || !sc->format) return stg::fail(q, CUDASP_ERR_ARG, "tweak_key: NULL schema or array");
  const std::string f = sc->format;
  k->offset = a->offset;
  k->length = a->length;
  k->may_have_nulls = may_have_nulls(a);
  k->validity = k->may_have_nulls ? static_cast<const uint8_t *>(a->buffers[0]) : nullptr;
  if (f.rfind("w:", 0) == 0) {
    k->width = atoi(f.c_str() + 2);
    if (k->width <= 0 || a->n_buffers < 2) return stg::fail(q, CUDASP_ERR_ARG, "tweak_key: malformed fixed_size_binary column");
    k->data = static_cast<const uint8_t *>(a->buffers[1]);
  } else if (f == "z" || f == "Z") {
    if (a->n_buffers < 3) return stg::fail(q, CUDASP_ERR_ARG, "tweak_key: malformed binary column");
    k->large = f == "Z";
    k->offsets = a->buffers[1];
    k->data = static_cast<const uint8_t *>(a->buffers[2]);
  } else {
    return stg::fail(q, CUDASP_ERR_ARG, "tweak_key must be fixed_size_binary(64|33), binary or large_binary");
  }
  if (a->length && (!k->data || (!k->width && !k->offsets))) {
    // an all-empty binary column may legally carry a NULL data buffer; only the offsets are needed then
    if (k->width || !k->offsets) return stg::fail(q, CUDASP_ERR_ARG, "tweak_key: missing buffer");
  }
  return 0;
}

int parse_list_column(cudasp_query *q, const ArrowSchema *sc, const ArrowArray *a, ListColumn *l) {
  if (!sc || !a || !sc->format) return stg::fail(q, CUDASP_ERR_ARG, "outputs: NULL schema or array");
  const std::string f = sc->format;
  if (f != "+l" && f != "+L") return stg::fail(q, CUDASP_ERR_ARG, "outputs must be list<int64> or large_list<int64>");
  if (sc->n_children != 1 || a->n_children != 1 || !sc->children[0]->format || std::string(sc->children[0]->format) != "l")
    return stg::fail(q, CUDASP_ERR_ARG, "outputs must be a list of int64 (BIGINT)");
  if (a->n_buffers < 2 || (a->length && !a->buffers[1])) return stg::fail(q, CUDASP_ERR_ARG, "outputs: missing offsets buffer");
  l->large = f == "+L";
  l->offset = a->offset;
  l->length = a->length;
  l->may_have_nulls = may_have_nulls(a);
  l->validity = l->may_have_nulls ? static_cast<const uint8_t *>(a->buffers[0]) : nullptr;
  l->offsets = a->buffers[1];
  const ArrowArray *c = a->children[0];
  if (c->n_buffers < 2) return stg::fail(q, CUDASP_ERR_ARG, "outputs: malformed child array");
  l->child_offset = c->offset;
  l->child_may_have_nulls = may_have_nulls(c);
  l->child_validity = l->child_may_have_nulls ? static_cast<const uint8_t *>(c->buffers[0]) : nullptr;
  l->values = static_cast<const int64_t *>(c->buffers[1]);
  if (c->length && !l->values) return stg::fail(q, CUDASP_ERR_ARG, "outputs: missing values buffer");
  return 0;
}

inline uint64_t now_ns() {
  return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int format_of(int64_t key_size) {
  return key_size == 64 ? CUDASP_KEYS_XY64 : key_size == 33 ? CUDASP_KEYS_SEC1_33 : -1;
}

// Streams rows of any number of (key, list) column pairs through the query's staging slots; ids come back as global
// row numbers (row_base + index in the pair), ascending per pair, pairs in push order.
//
// Pieces are cut across pushes: the rows a push leaves over stay in an OPEN piece (an acquired slot that is still being
// filled) and the next push goes on filling it, so a stream of small record batches, or of Parquet row groups, reaches the
// GPU in whole pieces (whole waves of the ladder kernel) instead of one under-filled launch per batch.  drain() submits
// what is left.  A piece therefore holds one or more SEGMENTS, each a run of staged rows that came from one push.
class ArrowScanner {
 public:
  // ramp == false: full pieces from the start (the Parquet front-end runs one scanner per decoder thread and ramps only as
  // many as there are devices)
  explicit ArrowScanner(cudasp_query *q, bool ramp = true) : q_(q), piece_(stg::piece_rows(q)) {
    cur_ = ramp ? std::min(piece_, std::max<uint32_t>(piece_ / 8, 8192u)) : piece_;     // ramp-up: see piece_now()
  }
  ~ArrowScanner() {
    if (has_open_) stg::release(q_, &open_.s);       // never submitted
    for (Pending &p : pending_) {        // error exit: let the slot's stream finish before its staging is reused
      uint32_t n = 0;
      stg::finish(q_, &p.s, &n);
      stg::release(q_, &p.s);
    }
  }

  std::vector<uint64_t> matches;
  // keep_keys: the key of every match, copied from the slot it was staged in when its piece comes back: 64 bytes per match
  // (a 33-byte key is zero-padded) and its width
  bool keep_keys = false;
  std::vector<uint8_t> match_keys, match_key_widths;
  uint64_t wait_ns = 0;             // time spent waiting for the devices: for a piece to come back, or for a free slot

  int push(const KeyColumn &k, const ListColumn &l, const uint8_t *row_validity, int64_t row_validity_offset, uint64_t row_base) {
    int64_t a = 0;
    const int64_t n = k.length;
    while (a < n) {
      int rc = dense_ok(k, l, row_validity) ? stage_dense(k, l, row_base, &a) : 1;
      if (rc < 0) return rc;
      if (rc == 1) {
        rc = stage_general(k, l, row_validity, row_validity_offset, row_base, &a);
        if (rc < 0) return rc;
      }
    }
    return 0;
  }

  int drain() {
    int rc = flush_open();
    while (!pending_.empty()) {
      int r = collect_front();
      if (!rc) rc = r;
    }
    return rc;
  }

 private:
  static constexpr uint32_t NO_MAP = 0xFFFFFFFFu;
  struct Segment {
    uint32_t staged_first;            // first staged row of the segment
    uint32_t map_first;               // NO_MAP: staged row r is source row base + (r - staged_first); else rowmap[map_first + ...]
    uint64_t base;                    // global row number the segment counts from
  };
  struct Pending {
    Staging s;
    std::vector<Segment> segs;
    std::vector<uint32_t> rowmap;     // general path: source row - base of its segment, one per staged row of mapped segments
    int fmt = 0;                      // key format of the piece (set when it is submitted)
  };

  cudasp_query *q_;
  uint32_t piece_;
  uint32_t cur_ = 0;
  std::deque<Pending> pending_;
  // the piece being filled
  Pending open_;
  bool has_open_ = false;
  uint32_t open_rows_ = 0, open_target_ = 0;
  uint64_t open_nout_ = 0;
  int open_fmt_ = -1;

  // rows for the next piece: the first pieces of a scan ramp up (1/8, 1/4, 1/2 of a full piece) so that the GPU starts
  // after a short host copy instead of idling through a full one
  uint32_t piece_now() const { return cur_; }
  void piece_done() { cur_ = (uint32_t)std::min<uint64_t>(piece_, 2ull * cur_); }

  // cnt keys of w bytes from row `base` on, packed
  static void copy_keys(uint8_t *dst, const KeyColumn &k, int64_t base, int64_t cnt, int64_t w) {
    const uint8_t *src = k.data + k.begin(base);
    if (!k.width || !k.stride || k.stride == k.width) {
      memcpy(dst, src, (size_t)cnt * w);
    } else if (w == 64) {
#if defined(__SSE2__)
      if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
        // keys lifted out of a Parquet page: written once, read next by the DMA engine -- streaming stores keep them out of the
        // caches (the same finding as the DuckDB extension's CopyStreaming)
        for (int64_t i = 0; i < cnt; i++) {
          const __m128i *a = reinterpret_cast<const __m128i *>(src + k.stride * i);
          __m128i *d = reinterpret_cast<__m128i *>(dst + 64 * i);
          const __m128i v0 = _mm_loadu_si128(a), v1 = _mm_loadu_si128(a + 1), v2 = _mm_loadu_si128(a + 2), v3 = _mm_loadu_si128(a + 3);
          _mm_stream_si128(d, v0); _mm_stream_si128(d + 1, v1); _mm_stream_si128(d + 2, v2); _mm_stream_si128(d + 3, v3);
        }
        _mm_sfence();
        return;
      }
#endif
      for (int64_t i = 0; i < cnt; i++) memcpy(dst + 64 * i, src + k.stride * i, 64);
    } else {
      for (int64_t i = 0; i < cnt; i++) memcpy(dst + w * i, src + k.stride * i, (size_t)w);
    }
  }

  static bool dense_ok(const KeyColumn &k, const ListColumn &l, const uint8_t *row_validity) {
    return !row_validity && !k.may_have_nulls && !l.may_have_nulls && !l.child_may_have_nulls;
  }

  int collect_front() {
    Pending p = std::move(pending_.front());
    pending_.pop_front();
    uint32_t n = 0;
    const uint64_t t0 = now_ns();
    int rc = stg::finish(q_, &p.s, &n);
    wait_ns += now_ns() - t0;
    if (!rc) {
      size_t g = 0;                   // ids ascend: the segment cursor only moves forward
      for (uint32_t i = 0; i < n; i++) {
        const uint32_t id = p.s.ids[i];
        while (g + 1 < p.segs.size() && p.segs[g + 1].staged_first <= id) g++;
        const Segment &sg = p.segs[g];
        const uint32_t r = id - sg.staged_first;
        matches.push_back(sg.base + (sg.map_first == NO_MAP ? r : p.rowmap[sg.map_first + r]));
        if (keep_keys) {
          const size_t w = p.fmt == CUDASP_KEYS_SEC1_33 ? 33 : 64, at = match_keys.size();
          match_keys.resize(at + 64, 0);
          memcpy(match_keys.data() + at, p.s.tweaks + (size_t)id * w, w);
          match_key_widths.push_back((uint8_t)w);
        }
      }
    }
    stg::release(q_, &p.s);
    return rc;
  }

  int acquire(Staging *s) {
    for (;;) {
      int rc = stg::acquire(q_, s, false);
      if (rc != 0) return rc < 0 ? rc : 0;
      if (!pending_.empty()) {        // recycle our own oldest piece (keeps results in row order)
        rc = collect_front();
        if (rc) return rc;
      } else {
        const uint64_t t0 = now_ns();
        rc = stg::acquire(q_, s, true);
        wait_ns += now_ns() - t0;
        return rc < 0 ? rc : 0;
      }
    }
  }

  // makes sure there is an open piece of key format fmt (submits an open piece of the other format first)
  int open_piece(int fmt) {
    if (has_open_ && open_fmt_ != fmt) {
      int rc = flush_open();
      if (rc) return rc;
    }
    if (has_open_) return 0;
    open_ = Pending{};
    int rc = acquire(&open_.s);
    if (rc) return rc;
    has_open_ = true;
    open_rows_ = 0;
    open_nout_ = 0;
    open_fmt_ = fmt;
    open_target_ = std::min<uint32_t>(piece_now(), open_.s.cap_rows);
    open_.s.offsets[0] = 0;
    return 0;
  }

  // submits the open piece as it is (a full piece, a change of key format, no room for a row's outputs, or the end)
  int flush_open() {
    if (!has_open_) return 0;
    has_open_ = false;
    if (open_rows_ == 0) {
      stg::release(q_, &open_.s);
      return 0;
    }
    int rc = stg::submit(q_, &open_.s, open_rows_, open_nout_, open_fmt_);
    if (rc) {
      stg::release(q_, &open_.s);
      return rc;
    }
    open_.fmt = open_fmt_;
    pending_.push_back(std::move(open_));
    piece_done();
    return 0;
  }

  // room for n more outputs in the open piece: 0 = yes, 1 = the piece was submitted to make room (open a new one), < 0 = error
  int room_for_outputs(uint64_t n) {
    if (open_nout_ + n <= open_.s.cap_outputs && open_nout_ + n <= 0xFFFFFFFFull) return 0;
    if (open_rows_) {
      int rc = flush_open();
      return rc ? rc : 1;
    }
    if (n > 0xFFFFFFFFull) return stg::fail(q_, CUDASP_ERR_ARG, "outputs: more than 2^32 values in one piece");
    return stg::reserve_outputs(q_, &open_.s, n);     // an empty piece: its staging may be replaced
  }

  // returns 0 = staged rows (*a advanced), 1 = these rows need the general path, < 0 = error
  int stage_dense(const KeyColumn &k, const ListColumn &l, uint64_t row_base, int64_t *a) {
    const int64_t base = *a;
    const int64_t w = k.size(base);
    const int fmt = format_of(w);
    if (fmt < 0) return 1;
    for (;;) {
      int rc = open_piece(fmt);
      if (rc) return rc;
      const int64_t cnt = std::min<int64_t>(open_target_ - open_rows_, k.length - base);
      if (!k.width) {                   // binary column: the rows must be one uniform width
        bool uniform = true;
        if (k.large) {
          const int64_t *o = static_cast<const int64_t *>(k.offsets) + k.offset + base;
          for (int64_t i = 0; i < cnt; i++) uniform &= (o[i + 1] - o[i]) == w;
        } else {
          const int32_t *o = static_cast<const int32_t *>(k.offsets) + k.offset + base;
          for (int64_t i = 0; i < cnt; i++) uniform &= (o[i + 1] - o[i]) == w;
        }
        if (!uniform) return 1;
      }
      const int64_t o0 = l.begin(base), o1 = l.begin(base + cnt);
      const uint64_t no = (uint64_t)(o1 - o0);
      if (o1 < o0 || no > 0xFFFFFFFFull) return stg::fail(q_, CUDASP_ERR_ARG, "outputs: offsets out of range");
      rc = room_for_outputs(no);
      if (rc < 0) return rc;
      if (rc == 1) continue;            // a fresh piece has other room: cut again
      Staging &s = open_.s;
      copy_keys(s.tweaks + (size_t)open_rows_ * w, k, base, cnt, w);
      // rebase; a malformed batch (offsets going backwards) is refused here: the kernels index outputs[] with these
      const uint32_t first = (uint32_t)open_nout_;
      uint32_t prev = first, bad = 0;
      uint32_t *so = s.offsets + open_rows_;
      if (l.large) {
        const int64_t *o = static_cast<const int64_t *>(l.offsets) + l.offset + base;
        for (int64_t i = 0; i <= cnt; i++) { const uint32_t v = first + (uint32_t)(o[i] - o0); bad |= (uint32_t)(v < prev); so[i] = prev = v; }
      } else {
        const int32_t *o = static_cast<const int32_t *>(l.offsets) + l.offset + base;
        for (int64_t i = 0; i <= cnt; i++) { const uint32_t v = first + (uint32_t)(o[i] - (int32_t)o0); bad |= (uint32_t)(v < prev); so[i] = prev = v; }
      }
      if (bad || prev != first + (uint32_t)no) {
        so[0] = first;                  // the piece stays as it was before these rows
        return stg::fail(q_, CUDASP_ERR_ARG, "outputs: list offsets are not monotonically non-decreasing");
      }
      if (no) memcpy(s.outputs + open_nout_, l.values + l.child_offset + o0, no * 8);
      open_.segs.push_back(Segment{open_rows_, NO_MAP, row_base + (uint64_t)base});
      open_rows_ += (uint32_t)cnt;
      open_nout_ += no;
      *a = base + cnt;
      return open_rows_ == open_target_ ? flush_open() : 0;
    }
  }

  int stage_general(const KeyColumn &k, const ListColumn &l, const uint8_t *row_validity, int64_t rv_off, uint64_t row_base,
                    int64_t *a);
};

// Row by row with the reference's NULL rules (src/cudasp_extension.cu:195-221).  Consumes source rows from *a until the
// piece is full, the key format changes, or the column ends; rows that can never match are skipped.
int ArrowScanner::stage_general(const KeyColumn &k, const ListColumn &l, const uint8_t *row_validity, int64_t rv_off,
                                uint64_t row_base, int64_t *a) {
  const int64_t n = k.length;
  int64_t i = *a;
  // skip leading rows that are dropped anyway, so that an all-NULL stretch does not cost a slot
  auto dropped = [&](int64_t r) {
    return (row_validity && !bit_set(row_validity, rv_off + r)) || !k.valid(r) || format_of(k.size(r)) < 0;
  };
  while (i < n && dropped(i)) i++;
  if (i == n) {
    *a = n;
    return 0;
  }
  const int64_t first = i;
  const int64_t w = k.size(first);
  int rc = open_piece(format_of(w));
  if (rc) return rc;
  open_.segs.push_back(Segment{open_rows_, (uint32_t)open_.rowmap.size(), row_base + (uint64_t)first});
  const uint32_t seg_first = open_rows_;
  for (; i < n && open_rows_ < open_target_; i++) {
    if (dropped(i)) continue;
    if (k.size(i) != w) break;                       // the other key format: next piece
    int64_t b = 0, e = 0;
    if (l.valid(i)) {
      b = l.begin(i);
      e = l.begin(i + 1);
      if (e < b) return stg::fail(q_, CUDASP_ERR_ARG, "outputs: offsets not monotonic");
    }
    const uint64_t len = (uint64_t)(e - b);
    if (open_nout_ + len > open_.s.cap_outputs || open_nout_ + len > 0xFFFFFFFFull) {
      if (open_rows_) break;                         // cut the piece here
      if ((rc = room_for_outputs(len)) < 0) return rc;
    }
    Staging &s = open_.s;
    memcpy(s.tweaks + (size_t)open_rows_ * w, k.data + k.begin(i), (size_t)w);
    if (l.child_validity) {
      for (int64_t j = b; j < e; j++)
        if (l.child_valid(j)) s.outputs[open_nout_++] = l.values[l.child_offset + j];
    } else if (len) {
      memcpy(s.outputs + open_nout_, l.values + l.child_offset + b, len * 8);
      open_nout_ += len;
    }
    open_.rowmap.push_back((uint32_t)(i - first));
    s.offsets[++open_rows_] = (uint32_t)open_nout_;
  }
  *a = i;
  if (open_rows_ == seg_first) open_.segs.pop_back();          // nothing staged (no room): the piece is cut below
  // submit when full, or when it was cut short (no room for the next row's outputs, or the other key format follows)
  if (open_rows_ == open_target_ || i < n) return flush_open();
  return 0;
}

int copy_out(cudasp_query *q, const std::vector<uint64_t> &m, uint64_t *match_rows, uint64_t capacity, uint64_t *n_matches) {
  *n_matches = m.size();
  if (m.size() > capacity) return stg::fail(q, CUDASP_ERR_ARG, "match_rows capacity too small");
  if (!m.empty()) memcpy(match_rows, m.data(), m.size() * sizeof(uint64_t));
  return CUDASP_OK;
}

struct BatchColumns {
  KeyColumn key;
  ListColumn list;
  const uint8_t *row_validity = nullptr;
  int64_t row_validity_offset = 0;
};

// children of a struct array are NOT sliced by the parent's offset: logical row i = child row (parent.offset + i)
int parse_batch(cudasp_query *q, const ArrowSchema *schema, const ArrowArray *batch, int tweak_col, int outputs_col, BatchColumns *bc) {
  if (!schema || !batch || !schema->format) return stg::fail(q, CUDASP_ERR_ARG, "NULL schema or batch");
  if (std::string(schema->format) != "+s") return stg::fail(q, CUDASP_ERR_ARG, "a record batch must be a struct array (+s)");
  if (tweak_col < 0 || outputs_col < 0 || tweak_col >= schema->n_children || outputs_col >= schema->n_children ||
      schema->n_children != batch->n_children)
    return stg::fail(q, CUDASP_ERR_ARG, "column index out of range");
  int rc = parse_key_column(q, schema->children[tweak_col], batch->children[tweak_col], &bc->key);
  if (rc) return rc;
  rc = parse_list_column(q, schema->children[outputs_col], batch->children[outputs_col], &bc->list);
  if (rc) return rc;
  bc->key.offset += batch->offset;
  bc->list.offset += batch->offset;
  if (bc->key.length < batch->length || bc->list.length < batch->length)
    return stg::fail(q, CUDASP_ERR_ARG, "a column is shorter than the batch");
  bc->key.length = bc->list.length = batch->length;
  if (may_have_nulls(batch)) {
    bc->row_validity = static_cast<const uint8_t *>(batch->buffers[0]);
    bc->row_validity_offset = batch->offset;
  }
  return 0;
}


int parse_pair(cudasp_query *q, const ArrowSchema *tweak_schema, const ArrowArray *tweak_key, const ArrowSchema *outputs_schema,
               const ArrowArray *outputs, KeyColumn *k, ListColumn *l) {
  int rc = parse_key_column(q, tweak_schema, tweak_key, k);
  if (rc) return rc;
  if ((rc = parse_list_column(q, outputs_schema, outputs, l))) return rc;
  if (k->length != l->length) return stg::fail(q, CUDASP_ERR_ARG, "tweak_key and outputs differ in length");
  return 0;
}

}  // namespace

struct csp::staging::ColumnScan {
  cudasp_query *q;
  ArrowScanner sc;
  ColumnScan(cudasp_query *query, bool ramp) : q(query), sc(query, ramp) { sc.keep_keys = true; }
};
csp::staging::ColumnScan *csp::staging::column_scan_open(cudasp_query *q, bool ramp) { return new ColumnScan(q, ramp); }
int csp::staging::column_scan_push(ColumnScan *cs, const ArrowSchema *tweak_schema, const ArrowArray *tweak_key,
                                   const ArrowSchema *outputs_schema, const ArrowArray *outputs, uint64_t row_base) {
  KeyColumn k;
  ListColumn l;
  int rc = parse_pair(cs->q, tweak_schema, tweak_key, outputs_schema, outputs, &k, &l);
  return rc ? rc : cs->sc.push(k, l, nullptr, 0, row_base);
}
int csp::staging::column_scan_push_keys(ColumnScan *cs, const uint8_t *keys, int64_t stride, int width, int64_t n,
                                        const ArrowSchema *outputs_schema, const ArrowArray *outputs, uint64_t row_base) {
  if (!keys || (width != 64 && width != 33) || stride < width || n < 0) return stg::fail(cs->q, CUDASP_ERR_ARG, "column_scan_push_keys: bad key view");
  KeyColumn k;
  ListColumn l;
  k.width = width;
  k.stride = stride;
  k.data = keys;
  k.length = n;
  int rc = parse_list_column(cs->q, outputs_schema, outputs, &l);
  if (rc) return rc;
  if (k.length != l.length) return stg::fail(cs->q, CUDASP_ERR_ARG, "tweak_key and outputs differ in length");
  return cs->sc.push(k, l, nullptr, 0, row_base);
}
void csp::staging::column_scan_matches(ColumnScan *cs, const uint64_t **rows, uint64_t *n, const uint8_t **keys64, const uint8_t **key_widths) {
  *rows = cs->sc.matches.data();
  *n = cs->sc.matches.size();
  if (keys64) *keys64 = cs->sc.match_keys.data();
  if (key_widths) *key_widths = cs->sc.match_key_widths.data();
}
int csp::staging::column_scan_finish(ColumnScan *cs) { return cs->sc.drain(); }
uint64_t csp::staging::column_scan_wait_ns(ColumnScan *cs) { return cs->sc.wait_ns; }
void csp::staging::column_scan_close(ColumnScan *cs) { delete cs; }

extern "C" {

int cudasp_scan_arrow_columns(cudasp_query *q, const ArrowSchema *tweak_schema, const ArrowArray *tweak_key,
                              const ArrowSchema *outputs_schema, const ArrowArray *outputs, uint64_t *match_rows,
                              uint64_t capacity, uint64_t *n_matches) {
  if (!q || !n_matches || (capacity && !match_rows)) return stg::fail(q, CUDASP_ERR_ARG, "NULL argument");
  *n_matches = 0;
  KeyColumn k;
  ListColumn l;
  int rc = parse_pair(q, tweak_schema, tweak_key, outputs_schema, outputs, &k, &l);
  if (rc) return rc;
  ArrowScanner sc(q);
  if ((rc = sc.push(k, l, nullptr, 0, 0))) return rc;
  if ((rc = sc.drain())) return rc;
  return copy_out(q, sc.matches, match_rows, capacity, n_matches);
}

int cudasp_scan_arrow(cudasp_query *q, const ArrowSchema *schema, const ArrowArray *batch, int tweak_col, int outputs_col,
                      uint64_t *match_rows, uint64_t capacity, uint64_t *n_matches) {
  if (!q || !n_matches || (capacity && !match_rows)) return stg::fail(q, CUDASP_ERR_ARG, "NULL argument");
  *n_matches = 0;
  BatchColumns bc;
  int rc = parse_batch(q, schema, batch, tweak_col, outputs_col, &bc);
  if (rc) return rc;
  ArrowScanner sc(q);
  if ((rc = sc.push(bc.key, bc.list, bc.row_validity, bc.row_validity_offset, 0))) return rc;
  if ((rc = sc.drain())) return rc;
  return copy_out(q, sc.matches, match_rows, capacity, n_matches);
}

int cudasp_scan_arrow_stream(cudasp_query *q, ArrowArrayStream *stream, int tweak_col, int outputs_col, uint64_t **match_rows,
                             uint64_t *n_matches, uint64_t *rows_seen) {
  if (!q || !stream || !match_rows || !n_matches) return stg::fail(q, CUDASP_ERR_ARG, "NULL argument");
  *match_rows = nullptr;
  *n_matches = 0;
  if (rows_seen) *rows_seen = 0;
  ArrowSchema schema;
  memset(&schema, 0, sizeof schema);
  if (stream->get_schema(stream, &schema) != 0) {
    const char *e = stream->get_last_error ? stream->get_last_error(stream) : nullptr;
    return stg::fail(q, CUDASP_ERR_ARG, e ? e : "ArrowArrayStream.get_schema failed");
  }
  ArrowScanner sc(q);
  uint64_t base = 0;
  int rc = 0;
  for (;;) {
    ArrowArray batch;
    memset(&batch, 0, sizeof batch);
    if (stream->get_next(stream, &batch) != 0) {
      const char *e = stream->get_last_error ? stream->get_last_error(stream) : nullptr;
      rc = stg::fail(q, CUDASP_ERR_ARG, e ? e : "ArrowArrayStream.get_next failed");
      break;
    }
    if (!batch.release) break;          // end of stream
    BatchColumns bc;
    rc = parse_batch(q, &schema, &batch, tweak_col, outputs_col, &bc);
    if (!rc) rc = sc.push(bc.key, bc.list, bc.row_validity, bc.row_validity_offset, base);
    base += (uint64_t)batch.length;
    batch.release(&batch);              // staged into pinned memory: the batch is no longer needed
    if (rc) break;
  }
  if (schema.release) schema.release(&schema);
  int r2 = sc.drain();
  if (!rc) rc = r2;
  if (rc) return rc;
  if (rows_seen) *rows_seen = base;
  *n_matches = sc.matches.size();
  if (!sc.matches.empty()) {
    *match_rows = static_cast<uint64_t *>(malloc(sc.matches.size() * sizeof(uint64_t)));
    if (!*match_rows) return stg::fail(q, CUDASP_ERR_STATE, "out of memory");
    memcpy(*match_rows, sc.matches.data(), sc.matches.size() * sizeof(uint64_t));
  }
  return CUDASP_OK;
}

void cudasp_free(void *p) { free(p); }

}  // extern "C"
