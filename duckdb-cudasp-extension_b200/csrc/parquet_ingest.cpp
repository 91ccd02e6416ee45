// parquet_ingest.cpp -- include/cudasp_parquet.h: Parquet pages decoded on the host straight into the pinned staging slots.
//
// Host-only C++ (g++), linked into libcudasp_b200.so.  No reference counterpart (the reference reads tables through DuckDB,
// README.md:100-110); the Parquet format itself is restated from its public specification (parquet-format: thrift compact
// protocol footer, PageHeader, data pages V1 / V2, PLAIN / RLE_DICTIONARY encodings, RLE / bit-packed hybrid levels).
// Layering: this file only DECODES and walks pages; staging belongs to arrow_ingest.cpp (NULL rules, key formats, piece cutting,
// slots), reached through staging_internal.hpp's ColumnScan: one scan per decoder thread, fed page by page with either a view of
// the keys where they lie in the file mapping (column_scan_push_keys) or Arrow arrays of a decoded page (column_scan_push), and
// the row group's outputs as a list<int64> array sliced per page.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/cudasp_arrow.h"
#include "../../include/cudasp_parquet.h"
#include "staging_internal.hpp"

namespace {

namespace stg = csp::staging;

struct Unsupported { std::string what; };     // thrown by the decoders; becomes CUDASP_ERR_UNSUPPORTED
struct Corrupt { std::string what; };         // ... CUDASP_ERR_IO

[[noreturn]] void unsupported(const std::string &w) { throw Unsupported{w}; }
[[noreturn]] void corrupt(const std::string &w) { throw Corrupt{w}; }

// ---------------------------------------------------------------- the file, mapped read-only
struct MappedFile {
  const uint8_t *data = nullptr;
  size_t size = 0;
  int fd = -1;
  ~MappedFile() {
    if (data) munmap(const_cast<uint8_t *>(data), size);
    if (fd >= 0) close(fd);
  }
  bool open_path(const char *path, std::string &err) {
    fd = ::open(path, O_RDONLY);
    if (fd < 0) { err = std::string("cannot open ") + path; return false; }
    struct stat st;
    if (fstat(fd, &st) != 0 || st.st_size < 12) { err = std::string("not a Parquet file: ") + path; return false; }
    size = (size_t)st.st_size;
    void *p = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
    if (p == MAP_FAILED) { err = std::string("cannot map ") + path; return false; }
    data = static_cast<const uint8_t *>(p);
    madvise(p, size, MADV_SEQUENTIAL);
    if (memcmp(data, "PAR1", 4) != 0 || memcmp(data + size - 4, "PAR1", 4) != 0) {
      err = std::string(memcmp(data + size - 4, "PARE", 4) == 0 ? "encrypted Parquet file: " : "not a Parquet file: ") + path;
      return false;
    }
    return true;
  }
};

// ---------------------------------------------------------------- thrift compact protocol (only what the footer and the page headers use)
struct Thrift {
  const uint8_t *p, *end;
  Thrift(const uint8_t *b, const uint8_t *e) : p(b), end(e) {}
  uint8_t byte() {
    if (p >= end) corrupt("thrift: truncated");
    return *p++;
  }
  uint64_t varint() {
    uint64_t v = 0;
    for (int s = 0; s < 64; s += 7) {
      const uint8_t b = byte();
      v |= (uint64_t)(b & 0x7f) << s;
      if (!(b & 0x80)) return v;
    }
    corrupt("thrift: varint too long");
  }
  int64_t zigzag() {
    const uint64_t v = varint();
    return (int64_t)(v >> 1) ^ -(int64_t)(v & 1);
  }
  std::string binary() {
    const uint64_t n = varint();
    if (n > (uint64_t)(end - p)) corrupt("thrift: string past the end");
    std::string s(reinterpret_cast<const char *>(p), (size_t)n);
    p += n;
    return s;
  }
  // field header: returns false at STOP; id is updated, type is the compact type (1/2 = bool true/false)
  bool field(int &id, int &type) {
    const uint8_t h = byte();
    if (h == 0) return false;
    type = h & 0x0f;
    const int delta = h >> 4;
    id = delta ? id + delta : (int)zigzag();
    return true;
  }
  void list_header(int &elem_type, uint64_t &n) {
    const uint8_t h = byte();
    elem_type = h & 0x0f;
    n = h >> 4;
    if (n == 15) n = varint();
    if (n > (uint64_t)(end - p)) corrupt("thrift: list longer than the bytes left");       // every element takes at least one byte
  }
  int depth = 0;            // nesting of the value being skipped: a damaged footer must not recurse without bound
  void skip(int type) {
    struct Nest {
      int &d;
      explicit Nest(int &x) : d(x) { if (++d > 64) corrupt("thrift: nesting too deep"); }
      ~Nest() { --d; }
    } nest(depth);
    switch (type) {
      case 1: case 2: return;                       // bool inside a field header
      case 3: byte(); return;
      case 4: case 5: case 6: varint(); return;
      case 7: if (end - p < 8) corrupt("thrift: truncated double"); p += 8; return;
      case 8: binary(); return;
      case 9: case 10: {
        int et; uint64_t n;
        list_header(et, n);
        for (uint64_t i = 0; i < n; i++) { if (et == 1 || et == 2) byte(); else skip(et); }
        return;
      }
      case 11: {
        const uint64_t n = varint();
        if (n > (uint64_t)(end - p)) corrupt("thrift: map longer than the bytes left");
        if (n) {
          const uint8_t kv = byte();
          auto elem = [&](int t) { if (t == 1 || t == 2) byte(); else skip(t); };     // a bool in a container is one byte
          for (uint64_t i = 0; i < n; i++) { elem(kv >> 4); elem(kv & 0x0f); }
        }
        return;
      }
      case 12: {
        int id = 0, t;
        while (field(id, t)) skip(t);
        return;
      }
      default: corrupt("thrift: unknown type");
    }
  }
};

// ---------------------------------------------------------------- the metadata this reader needs
enum PhysType { T_BOOLEAN = 0, T_INT32 = 1, T_INT64 = 2, T_INT96 = 3, T_FLOAT = 4, T_DOUBLE = 5, T_BYTE_ARRAY = 6, T_FLBA = 7 };
enum Encoding { E_PLAIN = 0, E_PLAIN_DICT = 2, E_RLE = 3, E_BIT_PACKED = 4, E_RLE_DICT = 8 };
enum Codec { C_UNCOMPRESSED = 0, C_SNAPPY = 1 };
enum PageType { P_DATA = 0, P_INDEX = 1, P_DICT = 2, P_DATA_V2 = 3 };

struct SchemaElem {
  int type = -1;            // PhysType, -1 for groups
  int type_length = 0;
  int repetition = 0;       // 0 required, 1 optional, 2 repeated
  std::string name;
  int num_children = 0;
};
struct Leaf {                // one leaf column of the flattened schema
  std::vector<std::string> path;
  int type = -1, type_length = 0, max_def = 0, max_rep = 0;
  int top_field = -1;        // index of the top-level field it belongs to
  std::vector<int> reps;     // repetition of every node on the path
};
struct ChunkMeta {
  int type = -1, codec = 0;
  int64_t num_values = 0, total_compressed = 0, data_page_offset = 0, dict_page_offset = -1;
};
struct RowGroupMeta {
  int64_t num_rows = 0;
  std::vector<ChunkMeta> cols;   // one per leaf, schema order
};
struct FileMeta {
  int64_t num_rows = 0;
  std::vector<SchemaElem> schema;
  std::vector<Leaf> leaves;
  std::vector<RowGroupMeta> row_groups;
};

SchemaElem read_schema_elem(Thrift &t) {
  SchemaElem e;
  int id = 0, ty;
  while (t.field(id, ty)) {
    switch (id) {
      case 1: e.type = (int)t.zigzag(); break;
      case 2: e.type_length = (int)t.zigzag(); break;
      case 3: e.repetition = (int)t.zigzag(); break;
      case 4: e.name = t.binary(); break;
      case 5: e.num_children = (int)t.zigzag(); break;
      default: t.skip(ty);
    }
  }
  return e;
}
ChunkMeta read_column_meta(Thrift &t) {
  ChunkMeta c;
  int id = 0, ty;
  while (t.field(id, ty)) {
    switch (id) {
      case 1: c.type = (int)t.zigzag(); break;
      case 4: c.codec = (int)t.zigzag(); break;
      case 5: c.num_values = t.zigzag(); break;
      case 7: c.total_compressed = t.zigzag(); break;
      case 9: c.data_page_offset = t.zigzag(); break;
      case 11: c.dict_page_offset = t.zigzag(); break;
      default: t.skip(ty);
    }
  }
  return c;
}
ChunkMeta read_column_chunk(Thrift &t) {
  ChunkMeta c;
  bool have = false;
  int id = 0, ty;
  while (t.field(id, ty)) {
    if (id == 1 && ty == 8) { if (!t.binary().empty()) unsupported("column chunks in other files (file_path)"); }
    else if (id == 3 && ty == 12) { c = read_column_meta(t); have = true; }
    else if (id == 8 && ty == 12) unsupported("encrypted column chunks");
    else t.skip(ty);
  }
  if (!have) corrupt("column chunk without metadata");
  return c;
}
RowGroupMeta read_row_group(Thrift &t) {
  RowGroupMeta g;
  int id = 0, ty;
  while (t.field(id, ty)) {
    if (id == 1 && ty == 9) {
      int et; uint64_t n;
      t.list_header(et, n);
      for (uint64_t i = 0; i < n; i++) g.cols.push_back(read_column_chunk(t));
    } else if (id == 3) {
      g.num_rows = t.zigzag();
    } else {
      t.skip(ty);
    }
  }
  return g;
}

void flatten_schema(FileMeta &m) {
  if (m.schema.empty()) corrupt("empty schema");
  size_t pos = 1;
  struct Walker {
    FileMeta &m; size_t &pos;
    void walk(std::vector<std::string> &path, std::vector<int> &reps, int def, int rep, int top) {
      if (pos >= m.schema.size()) corrupt("schema: children past the end");
      const SchemaElem &e = m.schema[pos++];
      path.push_back(e.name);
      reps.push_back(e.repetition);
      const int d = def + (e.repetition != 0), r = rep + (e.repetition == 2);
      if (e.num_children == 0) {
        Leaf l;
        l.path = path; l.type = e.type; l.type_length = e.type_length; l.max_def = d; l.max_rep = r; l.top_field = top; l.reps = reps;
        m.leaves.push_back(l);
      } else {
        for (int c = 0; c < e.num_children; c++) walk(path, reps, d, r, top);
      }
      path.pop_back();
      reps.pop_back();
    }
  } w{m, pos};
  std::vector<std::string> path;
  std::vector<int> reps;
  for (int f = 0; f < m.schema[0].num_children; f++) w.walk(path, reps, 0, 0, f);
}

FileMeta read_footer(const MappedFile &f) {
  uint32_t len;
  memcpy(&len, f.data + f.size - 8, 4);
  if ((size_t)len + 12 > f.size) corrupt("footer length past the file");
  Thrift t(f.data + f.size - 8 - len, f.data + f.size - 8);
  FileMeta m;
  int id = 0, ty;
  while (t.field(id, ty)) {
    if (id == 2 && ty == 9) {
      int et; uint64_t n;
      t.list_header(et, n);
      for (uint64_t i = 0; i < n; i++) m.schema.push_back(read_schema_elem(t));
    } else if (id == 3) {
      m.num_rows = t.zigzag();
    } else if (id == 4 && ty == 9) {
      int et; uint64_t n;
      t.list_header(et, n);
      for (uint64_t i = 0; i < n; i++) m.row_groups.push_back(read_row_group(t));
    } else if (id == 8 && ty == 12) {
      unsupported("encrypted files");
    } else {
      t.skip(ty);
    }
  }
  flatten_schema(m);
  for (const RowGroupMeta &g : m.row_groups)
    if (g.cols.size() != m.leaves.size()) corrupt("row group with a different number of column chunks than the schema has leaves");
  return m;
}

}  // namespace

namespace {

// ---------------------------------------------------------------- Snappy (raw block format) -- the writer default of most tools
void snappy_uncompress(const uint8_t *src, size_t n, std::vector<uint8_t> &out, size_t expect) {
  const uint8_t *p = src, *end = src + n;
  uint64_t len = 0;
  for (int s = 0;; s += 7) {
    if (p >= end || s > 35) corrupt("snappy: bad length");
    const uint8_t b = *p++;
    len |= (uint64_t)(b & 0x7f) << s;
    if (!(b & 0x80)) break;
  }
  if (len != expect) corrupt("snappy: uncompressed size differs from the page header");
  if (len > 32 * (uint64_t)n + 64) corrupt("snappy: implausible uncompressed size");      // a 3-byte copy yields at most 64 bytes
  out.resize(len);
  uint8_t *o = out.data(), *oend = o + len;
  while (p < end) {
    const uint8_t tag = *p++;
    if ((tag & 3) == 0) {                       // literal
      size_t l = (tag >> 2) + 1;
      if (l > 60) {
        const int nb = (int)l - 60;
        if (end - p < nb) corrupt("snappy: truncated literal length");
        l = 0;
        for (int i = 0; i < nb; i++) l |= (size_t)p[i] << (8 * i);
        l += 1;
        p += nb;
      }
      if ((size_t)(end - p) < l || (size_t)(oend - o) < l) corrupt("snappy: literal past the end");
      memcpy(o, p, l);
      o += l; p += l;
    } else {
      size_t l, off;
      if ((tag & 3) == 1) {
        if (p >= end) corrupt("snappy: truncated copy");
        l = 4 + ((tag >> 2) & 7);
        off = ((size_t)(tag >> 5) << 8) | *p++;
      } else if ((tag & 3) == 2) {
        if (end - p < 2) corrupt("snappy: truncated copy");
        l = (tag >> 2) + 1;
        off = p[0] | ((size_t)p[1] << 8);
        p += 2;
      } else {
        if (end - p < 4) corrupt("snappy: truncated copy");
        l = (tag >> 2) + 1;
        off = p[0] | ((size_t)p[1] << 8) | ((size_t)p[2] << 16) | ((size_t)p[3] << 24);
        p += 4;
      }
      if (off == 0 || off > (size_t)(o - out.data()) || (size_t)(oend - o) < l) corrupt("snappy: bad copy");
      const uint8_t *from = o - off;
      for (size_t i = 0; i < l; i++) o[i] = from[i];      // byte-wise: source and destination may overlap
      o += l;
    }
  }
  if (o != oend) corrupt("snappy: short output");
}

// ---------------------------------------------------------------- RLE / bit-packed hybrid (levels, dictionary indices)
struct HybridDecoder {
  const uint8_t *p = nullptr, *end = nullptr;
  int bw = 0;
  // current run
  uint64_t rle_left = 0, packed_left = 0;
  uint32_t rle_value = 0;
  uint64_t bitpos = 0;           // bit offset into the packed run starting at `p`
  void init(const uint8_t *b, const uint8_t *e, int bit_width) {
    p = b; end = e; bw = bit_width; rle_left = packed_left = 0; bitpos = 0;
    if (bw < 0 || bw > 32) corrupt("hybrid: bit width out of range");
  }
  void next_run() {
    if (packed_left == 0 && bitpos) { p += (bitpos + 7) / 8; bitpos = 0; }
    uint64_t h = 0;
    for (int s = 0;; s += 7) {
      if (p >= end || s > 63) corrupt("hybrid: truncated run header");
      const uint8_t b = *p++;
      h |= (uint64_t)(b & 0x7f) << s;
      if (!(b & 0x80)) break;
    }
    if (h & 1) {
      packed_left = (h >> 1) * 8;
      if (packed_left == 0) corrupt("hybrid: empty bit-packed run");
      if ((uint64_t)(end - p) * 8 < packed_left * (uint64_t)bw) {
        // the last run of a page may be padded short: clamp to the bytes that exist
        packed_left = bw ? (uint64_t)(end - p) * 8 / bw : packed_left;
        if (packed_left == 0) corrupt("hybrid: bit-packed run past the end");
      }
      bitpos = 0;
    } else {
      rle_left = h >> 1;
      if (rle_left == 0) corrupt("hybrid: empty RLE run");
      const int nb = (bw + 7) / 8;
      if (end - p < nb) corrupt("hybrid: truncated RLE value");
      rle_value = 0;
      for (int i = 0; i < nb; i++) rle_value |= (uint32_t)p[i] << (8 * i);
      p += nb;
    }
  }
  // n values into out
  void get(uint32_t *out, uint64_t n) {
    while (n) {
      if (rle_left == 0 && packed_left == 0) {
        if (bw == 0 && p >= end) { for (uint64_t i = 0; i < n; i++) out[i] = 0; return; }   // zero-width levels: nothing stored
        next_run();
      }
      if (rle_left) {
        const uint64_t k = std::min(n, rle_left);
        for (uint64_t i = 0; i < k; i++) out[i] = rle_value;
        out += k; n -= k; rle_left -= k;
      } else {
        const uint64_t k = std::min(n, packed_left);
        const uint32_t mask = bw == 32 ? 0xffffffffu : ((1u << bw) - 1u);
        uint64_t i = 0;
        if (bw <= 8 && (bitpos & 7) == 0) {
          // whole groups of 8 values = bw bytes: one 64-bit load, eight shifts (levels are 1-3 bits wide, dictionary indices of small
          // dictionaries up to 8)
          const uint8_t *g = p + (bitpos >> 3);
          for (; i + 8 <= k && g + 8 <= end; i += 8, g += bw) {
            uint64_t w;
            memcpy(&w, g, 8);
            for (int j = 0; j < 8; j++) out[i + j] = (uint32_t)(w >> (j * bw)) & mask;
          }
        }
        for (; i < k; i++) {
          const uint64_t bit = bitpos + i * (uint64_t)bw;
          const uint8_t *q = p + (bit >> 3);
          uint64_t w = 0;
          const size_t avail = (size_t)(end - q);
          memcpy(&w, q, avail < 8 ? avail : 8);
          out[i] = (uint32_t)(w >> (bit & 7)) & mask;
        }
        bitpos += k * (uint64_t)bw;
        out += k; n -= k; packed_left -= k;
        if (packed_left == 0) { p += (bitpos + 7) / 8; bitpos = 0; }
      }
    }
  }
  // One-bit values (repetition levels of a one-level list): calls f(i) for every i in [0, n) whose value is 0 -- where a row
  // starts -- without expanding the levels: an RLE run is a loop or a skip, a bit-packed run is scanned 64 levels at a time.
  template <class F>
  void for_each_zero(uint64_t n, F f) {
    if (bw != 1) corrupt("hybrid: one-bit levels expected");
    uint64_t at = 0;
    while (at < n) {
      if (rle_left == 0 && packed_left == 0) next_run();
      if (rle_left) {
        const uint64_t k = std::min(n - at, rle_left);
        if (!(rle_value & 1)) for (uint64_t i = 0; i < k; i++) f(at + i);
        at += k; rle_left -= k;
      } else {
        const uint64_t k = std::min(n - at, packed_left);
        uint64_t i = 0;
        while (i < k) {
          const uint64_t bit = bitpos + i;
          const uint8_t *q = p + (bit >> 3);
          uint64_t w = 0;
          const size_t avail = (size_t)(end - q);
          memcpy(&w, q, avail < 8 ? avail : 8);
          const unsigned sh = (unsigned)(bit & 7);
          const uint64_t take = std::min<uint64_t>(k - i, 64 - sh);       // levels of this load that belong to the run
          uint64_t z = ~(w >> sh);
          if (take < 64) z &= (1ull << take) - 1;
          while (z) {
            f(at + i + (uint64_t)__builtin_ctzll(z));
            z &= z - 1;
          }
          i += take;
        }
        bitpos += k;
        at += k; packed_left -= k;
        if (packed_left == 0) { p += (bitpos + 7) / 8; bitpos = 0; }
      }
    }
  }
  // true (and consumes them) when the next n values are one RLE run of `value`: the all-valid fast path
  bool all_equal(uint64_t n, uint32_t value) {
    if (rle_left == 0 && packed_left == 0) {
      if (bw == 0) return value == 0;
      if (p >= end) return false;
      next_run();
    }
    if (rle_left >= n && rle_value == value) { rle_left -= n; return true; }
    return false;
  }
};
inline int bits_for(int max_level) {
  int b = 0;
  while ((1 << b) <= max_level) b++;
  return b;
}

// ---------------------------------------------------------------- pages of one column chunk
struct PageHeader {
  int type = -1;
  int32_t uncompressed = 0, compressed = 0;
  int32_t num_values = 0, encoding = 0;
  // V2
  int32_t def_bytes = 0, rep_bytes = 0;
  bool v2_compressed = true;
};
PageHeader read_page_header(Thrift &t) {
  PageHeader h;
  int id = 0, ty;
  while (t.field(id, ty)) {
    if (id == 1) h.type = (int)t.zigzag();
    else if (id == 2) h.uncompressed = (int32_t)t.zigzag();
    else if (id == 3) h.compressed = (int32_t)t.zigzag();
    else if ((id == 5 || id == 7) && ty == 12) {          // DataPageHeader / DictionaryPageHeader: num_values, encoding first
      int id2 = 0, ty2;
      while (t.field(id2, ty2)) {
        if (id2 == 1) h.num_values = (int32_t)t.zigzag();
        else if (id2 == 2) h.encoding = (int32_t)t.zigzag();
        else t.skip(ty2);
      }
    } else if (id == 8 && ty == 12) {                       // DataPageHeaderV2
      int id2 = 0, ty2;
      while (t.field(id2, ty2)) {
        if (id2 == 1) h.num_values = (int32_t)t.zigzag();
        else if (id2 == 4) h.encoding = (int32_t)t.zigzag();
        else if (id2 == 5) h.def_bytes = (int32_t)t.zigzag();
        else if (id2 == 6) h.rep_bytes = (int32_t)t.zigzag();
        else if (id2 == 7) h.v2_compressed = (ty2 == 1);
        else t.skip(ty2);
      }
    } else {
      t.skip(ty);
    }
  }
  return h;
}

// One data page, ready to decode: levels and values located, decompressed if need be.
struct Page {
  int32_t num_values = 0, encoding = 0;
  bool skipped = false;         // a page none of whose rows is wanted (ChunkReader::want): only num_values is set
  const uint8_t *rep = nullptr, *rep_end = nullptr, *def = nullptr, *def_end = nullptr, *values = nullptr, *values_end = nullptr;
  std::vector<uint8_t> buf;     // backing store when the page was compressed
};
struct Dictionary {
  std::vector<uint8_t> buf;     // backing store (decompressed page), or empty when it points into the file
  const uint8_t *data = nullptr, *end = nullptr;
  int32_t n = 0;
  std::vector<uint32_t> offs;   // BYTE_ARRAY: start of entry i's bytes, n + 1 entries (lengths implied)
};

struct ChunkReader {
  const MappedFile &f;
  const ChunkMeta &cm;
  const Leaf &leaf;
  const uint8_t *p, *end;
  int64_t values_left;
  Dictionary dict;
  // Row filter for FLAT columns (one value slot per row): when set, a data page that holds none of the wanted rows of the
  // row group (ascending) comes back as `skipped` -- not decompressed, not decoded.  collect_matches() reads the few pages
  // that hold a matching row this way instead of whole column chunks.
  const uint32_t *want = nullptr;
  size_t n_want = 0, want_pos = 0;
  int64_t row_pos = 0;
  ChunkReader(const MappedFile &file, const ChunkMeta &c, const Leaf &l, const uint32_t *wanted = nullptr, size_t n_wanted = 0)
      : f(file), cm(c), leaf(l), values_left(c.num_values), want(wanted), n_want(n_wanted) {
    if (c.codec != C_UNCOMPRESSED && c.codec != C_SNAPPY) unsupported("compression codec " + std::to_string(c.codec) + " (only uncompressed and Snappy pages are decoded)");
    int64_t start = c.data_page_offset;
    if (c.dict_page_offset > 0 && c.dict_page_offset < start) start = c.dict_page_offset;
    if (start < 4 || c.total_compressed < 0 || (uint64_t)start + (uint64_t)c.total_compressed > f.size) corrupt("column chunk outside the file");
    p = f.data + start;
    end = p + c.total_compressed;
    chunk_start = start;
    // map the chunk's pages in one call instead of one soft fault per 4 KB while decoding (Linux >= 5.14; harmless when refused)
#ifndef MADV_POPULATE_READ
#define MADV_POPULATE_READ 22
#endif
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)4095;
    if (!want) {
      madvise(reinterpret_cast<void *>(a0), (size_t)(reinterpret_cast<uintptr_t>(end) - a0), MADV_POPULATE_READ);
      populated = reinterpret_cast<void *>(a0);
      populated_len = (size_t)(reinterpret_cast<uintptr_t>(end) - a0);
    }
  }
  // ... and drop the page-table entries again when the chunk is done (the page cache keeps the data): otherwise the final munmap
  // of the file walks every entry of every chunk in one serial stretch at the end of the call -- measured 46 ms of a 0.40 s scan
  // of 100 M rows (profiles/r02p_parquet_native_8gpu_trace.txt)
  void *populated = nullptr;
  size_t populated_len = 0;
  int64_t chunk_start = 0;
  ~ChunkReader() {
    if (populated) madvise(populated, populated_len, MADV_DONTNEED);
  }
  ChunkReader(const ChunkReader &) = delete;
  ChunkReader &operator=(const ChunkReader &) = delete;
  void load_dictionary(const PageHeader &h, const uint8_t *body) {
    if (h.encoding != E_PLAIN && h.encoding != E_PLAIN_DICT) unsupported("dictionary page encoding " + std::to_string(h.encoding));
    if (cm.codec == C_SNAPPY) {
      snappy_uncompress(body, (size_t)h.compressed, dict.buf, (size_t)h.uncompressed);
      dict.data = dict.buf.data(); dict.end = dict.data + dict.buf.size();
    } else {
      dict.data = body; dict.end = body + h.compressed;
    }
    dict.n = h.num_values;
    if (dict.n < 0) corrupt("dictionary: negative entry count");
    if (leaf.type == T_FLBA && (leaf.type_length <= 0 || (int64_t)(dict.end - dict.data) < (int64_t)dict.n * leaf.type_length))
      corrupt("dictionary: fixed-length entries past the end");
    if (leaf.type == T_BYTE_ARRAY) {
      dict.offs.resize((size_t)dict.n + 1);
      const uint8_t *q = dict.data;
      for (int32_t i = 0; i < dict.n; i++) {
        if (dict.end - q < 4) corrupt("dictionary: truncated");
        uint32_t len;
        memcpy(&len, q, 4);
        q += 4;
        if ((uint64_t)(dict.end - q) < len) corrupt("dictionary: entry past the end");
        dict.offs[i] = (uint32_t)(q - dict.data);
        q += len;
      }
      dict.offs[dict.n] = (uint32_t)(q - dict.data) + 4;      // so that length i = offs[i + 1] - offs[i] - 4
    }
  }
  // the next DATA page, or false at the end of the chunk
  bool next(Page &pg) {
    for (;;) {
      if (values_left <= 0 || p >= end) return false;
      Thrift t(p, end);
      const PageHeader h = read_page_header(t);
      const uint8_t *body = t.p;
      if (h.compressed < 0 || (int64_t)(end - body) < h.compressed) corrupt("page past the end of its column chunk");
      p = body + h.compressed;
      if (h.type == P_DICT) { load_dictionary(h, body); continue; }
      if (h.type == P_INDEX) continue;
      if (h.type != P_DATA && h.type != P_DATA_V2) unsupported("page type " + std::to_string(h.type));
      if (h.num_values < 0 || (int64_t)h.num_values > values_left) corrupt("page: more values than its column chunk holds");
      pg.num_values = h.num_values;
      pg.encoding = h.encoding;
      pg.skipped = false;
      if (want) {
        while (want_pos < n_want && (int64_t)want[want_pos] < row_pos) want_pos++;
        pg.skipped = want_pos == n_want || (int64_t)want[want_pos] >= row_pos + h.num_values;
        row_pos += h.num_values;
        if (pg.skipped) {
          values_left -= h.num_values;
          return true;
        }
        // a page that will be read: map it in one call (one soft fault per 4 KB otherwise: measured 2.3 ms per row group with a
        // match, profiles/r02p_parquet_native_8gpu_trace.txt)
        const uintptr_t b0 = reinterpret_cast<uintptr_t>(body) & ~(uintptr_t)4095;
        madvise(reinterpret_cast<void *>(b0), (size_t)(reinterpret_cast<uintptr_t>(body) + (size_t)h.compressed - b0), MADV_POPULATE_READ);
        if (!populated) {
          const uintptr_t c0 = reinterpret_cast<uintptr_t>(f.data + chunk_start) & ~(uintptr_t)4095;
          populated = reinterpret_cast<void *>(c0);
          populated_len = (size_t)(reinterpret_cast<uintptr_t>(end) - c0);
        }
      }
      pg.rep = pg.rep_end = pg.def = pg.def_end = nullptr;
      const uint8_t *b = body, *be = body + h.compressed;
      if (h.type == P_DATA) {
        if (cm.codec == C_SNAPPY) {
          snappy_uncompress(body, (size_t)h.compressed, pg.buf, (size_t)h.uncompressed);
          b = pg.buf.data(); be = b + pg.buf.size();
        }
        for (int which = 0; which < 2; which++) {             // repetition levels first, then definition levels
          const int maxl = which == 0 ? leaf.max_rep : leaf.max_def;
          if (maxl == 0) continue;
          if (be - b < 4) corrupt("page: truncated level length");
          uint32_t len;
          memcpy(&len, b, 4);
          b += 4;
          if ((uint64_t)(be - b) < len) corrupt("page: levels past the end");
          if (which == 0) { pg.rep = b; pg.rep_end = b + len; } else { pg.def = b; pg.def_end = b + len; }
          b += len;
        }
        pg.values = b; pg.values_end = be;
      } else {
        if (h.rep_bytes < 0 || h.def_bytes < 0 || (int64_t)h.rep_bytes + h.def_bytes > h.compressed) corrupt("page v2: level lengths");
        pg.rep = body; pg.rep_end = body + h.rep_bytes;
        pg.def = pg.rep_end; pg.def_end = pg.def + h.def_bytes;
        const uint8_t *v = pg.def_end;
        if (cm.codec == C_SNAPPY && h.v2_compressed) {
          snappy_uncompress(v, (size_t)(be - v), pg.buf, (size_t)(h.uncompressed - h.rep_bytes - h.def_bytes));
          pg.values = pg.buf.data(); pg.values_end = pg.values + pg.buf.size();
        } else {
          pg.values = v; pg.values_end = be;
        }
      }
      values_left -= h.num_values;
      return true;
    }
  }
};

}  // namespace

namespace {

// ---------------------------------------------------------------- decoded columns of one row group (Arrow buffer shapes)
struct Bitmap {
  std::vector<uint8_t> bits;
  int64_t nulls = 0;
  void init(int64_t n) { bits.assign((size_t)(n + 7) / 8, 0xff); nulls = 0; }
  void clear(int64_t i) {
    if (bits[(size_t)i >> 3] & (1u << (i & 7))) { bits[(size_t)i >> 3] &= (uint8_t)~(1u << (i & 7)); nulls++; }
  }
  bool get(int64_t i) const { return bits[(size_t)i >> 3] & (1u << (i & 7)); }
};
struct RawBytes {             // growable, never value-initialised, keeps its capacity from row group to row group
  uint8_t *p = nullptr;
  size_t len = 0, cap = 0;
  RawBytes() = default;
  RawBytes(const RawBytes &) = delete;
  RawBytes &operator=(const RawBytes &) = delete;
  ~RawBytes() { free(p); }
  uint8_t *grow(size_t n) {   // room for n more bytes; returns where they go
    if (len + n > cap) {
      const size_t want = std::max(len + n, cap + cap / 2 + 4096);
      uint8_t *q = static_cast<uint8_t *>(realloc(p, want));
      if (!q) throw std::bad_alloc();
      p = q; cap = want;
    }
    uint8_t *at = p + len;
    len += n;
    return at;
  }
  size_t size() const { return len; }
  bool empty() const { return len == 0; }
  const uint8_t *data() const { return p; }
  void clear() { len = 0; }
};
struct BlobColumn {           // binary: int32 offsets + bytes + validity
  std::vector<int32_t> offsets;
  RawBytes data;
  Bitmap valid;
};
struct Int32Column {
  std::vector<int32_t> values;
  Bitmap valid;
};
struct ListColumn {           // list<int64>: int32 offsets + list validity; child values + child validity
  std::vector<int32_t> offsets;
  Bitmap valid;
  RawBytes value_bytes;       // int64 values, 8 bytes each
  std::vector<uint32_t> null_children;   // positions of NULL elements (rare): child_valid is built from them
  Bitmap child_valid;
  size_t n_values() const { return value_bytes.size() / 8; }
  const int64_t *values() const { return reinterpret_cast<const int64_t *>(value_bytes.data()); }
};

// values of one page of a BYTE_ARRAY / FIXED_LEN_BYTE_ARRAY column: calls put(ptr, len) once per non-null value, in order
template <class Put>
void page_blobs(const ChunkReader &cr, const Page &pg, int64_t n_present, Put put) {
  const Leaf &lf = cr.leaf;
  if (pg.encoding == E_PLAIN) {
    const uint8_t *q = pg.values;
    if (lf.type == T_FLBA) {
      const int64_t w = lf.type_length;
      if (w <= 0 || (int64_t)(pg.values_end - q) < w * n_present) corrupt("page: fixed-length values past the end");
      for (int64_t i = 0; i < n_present; i++, q += w) put(q, (uint32_t)w);
    } else {
      for (int64_t i = 0; i < n_present; i++) {
        if (pg.values_end - q < 4) corrupt("page: truncated value length");
        uint32_t len;
        memcpy(&len, q, 4);
        q += 4;
        if ((uint64_t)(pg.values_end - q) < len) corrupt("page: value past the end");
        put(q, len);
        q += len;
      }
    }
  } else if (pg.encoding == E_RLE_DICT || pg.encoding == E_PLAIN_DICT) {
    if (!cr.dict.data) corrupt("dictionary-encoded page without a dictionary page");
    if (pg.values >= pg.values_end) { if (n_present) corrupt("page: no dictionary indices"); return; }
    HybridDecoder ix;
    ix.init(pg.values + 1, pg.values_end, pg.values[0]);
    uint32_t idx[256];
    for (int64_t done = 0; done < n_present;) {
      const int64_t k = std::min<int64_t>(256, n_present - done);
      ix.get(idx, (uint64_t)k);
      for (int64_t i = 0; i < k; i++) {
        if (idx[i] >= (uint32_t)cr.dict.n) corrupt("dictionary index out of range");
        if (lf.type == T_FLBA) put(cr.dict.data + (size_t)idx[i] * lf.type_length, (uint32_t)lf.type_length);
        else put(cr.dict.data + cr.dict.offs[idx[i]], cr.dict.offs[idx[i] + 1] - cr.dict.offs[idx[i]] - 4);
      }
      done += k;
    }
  } else {
    unsupported("value encoding " + std::to_string(pg.encoding) + " of column " + lf.path[0]);
  }
}

// One data page of a flat BYTE_ARRAY / FIXED_LEN_BYTE_ARRAY column, appended to `out` as rows [row, row + pg.num_values): the
// caller has sized out.offsets and initialised out.valid for them.
void decode_blob_page(const ChunkReader &cr, const Page &pg, const Leaf &lf, BlobColumn &out, int64_t row, std::vector<uint32_t> &defs) {
  const int64_t n = pg.num_values;
  int64_t present = n;
  bool all_valid = true;
  if (lf.max_def) {
    HybridDecoder d;
    d.init(pg.def, pg.def_end, 1);
    if (!d.all_equal((uint64_t)n, 1)) {
      d.init(pg.def, pg.def_end, 1);
      defs.resize((size_t)n);
      d.get(defs.data(), (uint64_t)n);
      present = 0;
      for (int64_t i = 0; i < n; i++) present += defs[i] != 0;
      all_valid = present == n;
    }
  }
  int64_t i = 0;      // value slot within the page
  if (all_valid && pg.encoding == E_PLAIN && lf.type == T_FLBA) {
    // fixed-size values, no NULLs: the page's value area IS the column data
    const size_t w = (size_t)lf.type_length, bytes = w * (size_t)n;
    if (lf.type_length <= 0 || (size_t)(pg.values_end - pg.values) < bytes) corrupt("page: fixed-length values past the end");
    const size_t base = out.data.size();
    if (base + bytes > 0x7fffffffull) unsupported("more than 2 GiB of " + lf.path[0] + " in one row group");
    memcpy(out.data.grow(bytes), pg.values, bytes);
    int32_t *o = out.offsets.data() + row + 1;
    for (int64_t k = 0; k < n; k++) o[k] = (int32_t)(base + w * (size_t)(k + 1));
    i = n;
  } else {
    page_blobs(cr, pg, present, [&](const uint8_t *p, uint32_t len) {
      if (!all_valid) {
        while (defs[(size_t)i] == 0) { out.valid.clear(row + i); out.offsets[(size_t)(row + i) + 1] = (int32_t)out.data.size(); i++; }
      }
      uint8_t *dst = out.data.grow(len);
      if (len == 64) memcpy(dst, p, 64); else if (len == 32) memcpy(dst, p, 32); else memcpy(dst, p, len);
      out.offsets[(size_t)(row + i) + 1] = (int32_t)out.data.size();
      i++;
    });
    if (out.data.size() > 0x7fffffffull) unsupported("more than 2 GiB of " + lf.path[0] + " in one row group");
  }
  for (; i < n; i++) { out.valid.clear(row + i); out.offsets[(size_t)(row + i) + 1] = (int32_t)out.data.size(); }   // trailing NULLs
}

void check_blob_leaf(const ChunkMeta &cm, const Leaf &lf) {
  if (lf.type != T_BYTE_ARRAY && lf.type != T_FLBA) unsupported("column " + lf.path[0] + " is not BYTE_ARRAY / FIXED_LEN_BYTE_ARRAY");
  if (lf.max_rep != 0 || lf.max_def > 1) unsupported("column " + lf.path[0] + " is nested");
  if (lf.type == T_FLBA && cm.dict_page_offset > 0 && lf.type_length <= 0) corrupt("fixed-length column without a length");
}

// The wanted rows (ascending) of a flat BYTE_ARRAY / FIXED_LEN_BYTE_ARRAY column: out[k] = the bytes of row want[k], valid[k] = 0
// when it is NULL.  Pages without a wanted row are skipped unread; in the others the values are walked, not copied.
void fetch_blob_rows(const MappedFile &f, const ChunkMeta &cm, const Leaf &lf, int64_t rows, const uint32_t *want, size_t n_want,
                     std::vector<std::string> &out, std::vector<uint8_t> &valid, std::vector<uint32_t> &defs) {
  check_blob_leaf(cm, lf);
  out.assign(n_want, std::string());
  valid.assign(n_want, 0);
  ChunkReader cr(f, cm, lf, want, n_want);
  Page pg;
  int64_t row = 0;
  size_t k = 0;
  while (cr.next(pg)) {
    const int64_t n = pg.num_values;
    if (row + n > rows) corrupt("column " + lf.path[0] + ": more values than the row group has rows");
    if (pg.skipped) { row += n; continue; }
    while (k < n_want && (int64_t)want[k] < row) k++;
    int64_t present = n;
    bool all_valid = true;
    if (lf.max_def) {
      HybridDecoder d;
      d.init(pg.def, pg.def_end, 1);
      if (!d.all_equal((uint64_t)n, 1)) {
        d.init(pg.def, pg.def_end, 1);
        defs.resize((size_t)n);
        d.get(defs.data(), (uint64_t)n);
        present = 0;
        for (int64_t i = 0; i < n; i++) present += defs[i] != 0;
        all_valid = present == n;
      }
    }
    int64_t i = 0;      // value slot within the page
    page_blobs(cr, pg, present, [&](const uint8_t *p, uint32_t len) {
      if (!all_valid) {
        while (defs[(size_t)i] == 0) {
          if (k < n_want && (int64_t)want[k] == row + i) k++;          // NULL: valid[k] stays 0
          i++;
        }
      }
      if (k < n_want && (int64_t)want[k] == row + i) {
        out[k].assign(reinterpret_cast<const char *>(p), len);
        valid[k] = 1;
        k++;
      }
      i++;
    });
    row += n;
  }
  if (row != rows) corrupt("column " + lf.path[0] + ": fewer values than the row group has rows");
}

// A data page whose keys can be staged from where they lie: no NULLs, PLAIN, every value 64 or 33 bytes long.  Fixed-length
// values are packed (stride = width); BYTE_ARRAY values carry a 4-byte length each (stride = width + 4), all checked here.
bool page_key_view(const Page &pg, const Leaf &lf, const uint8_t **keys, int64_t *stride, int *width) {
  const int64_t n = pg.num_values;
  if (n <= 0 || pg.encoding != E_PLAIN) return false;
  if (lf.max_def) {
    HybridDecoder d;
    d.init(pg.def, pg.def_end, 1);
    if (!d.all_equal((uint64_t)n, 1)) return false;
  }
  const int64_t room = (int64_t)(pg.values_end - pg.values);
  if (lf.type == T_FLBA) {
    const int64_t w = lf.type_length;
    if ((w != 64 && w != 33) || room < w * n) return false;
    *keys = pg.values; *stride = w; *width = (int)w;
    return true;
  }
  if (room < 4) return false;
  uint32_t len0;
  memcpy(&len0, pg.values, 4);
  if ((len0 != 64 && len0 != 33) || room < (int64_t)(len0 + 4) * n) return false;
  const int64_t st = (int64_t)len0 + 4;
  uint32_t diff = 0;
  const uint8_t *q = pg.values;
  for (int64_t i = 0; i < n; i++, q += st) {
    uint32_t len;
    memcpy(&len, q, 4);
    diff |= len ^ len0;
  }
  if (diff) return false;
  *keys = pg.values + 4; *stride = st; *width = (int)len0;
  return true;
}

void decode_int32_column(const MappedFile &f, const ChunkMeta &cm, const Leaf &lf, int64_t rows, Int32Column &out,
                         const uint32_t *want = nullptr, size_t n_want = 0) {
  if (lf.type != T_INT32) unsupported("column " + lf.path[0] + " is not INT32");
  if (lf.max_rep != 0 || lf.max_def > 1) unsupported("column " + lf.path[0] + " is nested");
  ChunkReader cr(f, cm, lf, want, n_want);
  out.values.assign((size_t)rows, 0);
  out.valid.init(rows);
  int64_t row = 0;
  Page pg;
  std::vector<uint32_t> defs, idx;
  while (cr.next(pg)) {
    const int64_t n = pg.num_values;
    if (row + n > rows) corrupt("column " + lf.path[0] + ": more values than rows");
    if (pg.skipped) {
      row += n;
      continue;
    }
    defs.assign((size_t)n, 1);
    if (lf.max_def) {
      HybridDecoder d;
      d.init(pg.def, pg.def_end, 1);
      d.get(defs.data(), (uint64_t)n);
    }
    int64_t present = 0;
    for (int64_t i = 0; i < n; i++) present += defs[(size_t)i] != 0;
    const uint8_t *q = pg.values;
    if (pg.encoding == E_PLAIN) {
      if ((int64_t)(pg.values_end - q) < 4 * present) corrupt("page: INT32 values past the end");
      for (int64_t i = 0; i < n; i++) {
        if (defs[(size_t)i]) { memcpy(&out.values[(size_t)(row + i)], q, 4); q += 4; } else out.valid.clear(row + i);
      }
    } else if (pg.encoding == E_RLE_DICT || pg.encoding == E_PLAIN_DICT) {
      if (!cr.dict.data || (int64_t)(cr.dict.end - cr.dict.data) < 4 * (int64_t)cr.dict.n) corrupt("INT32 dictionary");
      idx.resize((size_t)present);
      if (present) {
        if (q >= pg.values_end) corrupt("page: no dictionary indices");
        HybridDecoder ix;
        ix.init(q + 1, pg.values_end, q[0]);
        ix.get(idx.data(), (uint64_t)present);
      }
      int64_t k = 0;
      for (int64_t i = 0; i < n; i++) {
        if (!defs[(size_t)i]) { out.valid.clear(row + i); continue; }
        if (idx[(size_t)k] >= (uint32_t)cr.dict.n) corrupt("dictionary index out of range");
        memcpy(&out.values[(size_t)(row + i)], cr.dict.data + 4 * (size_t)idx[(size_t)k++], 4);
      }
    } else {
      unsupported("value encoding " + std::to_string(pg.encoding) + " of column " + lf.path[0]);
    }
    row += n;
  }
  if (row != rows) corrupt("column " + lf.path[0] + ": fewer values than rows");
}

// LIST<INT64>: one repeated node on the path.  def < d0: an ancestor is NULL (NULL list); def == d0: empty list;
// d0 < def < max_def: NULL element; def == max_def: a value.  rep == 0 starts a row.
void decode_list_column(const MappedFile &f, const ChunkMeta &cm, const Leaf &lf, int64_t rows, ListColumn &out) {
  if (lf.type != T_INT64) unsupported("column " + lf.path[0] + " does not hold INT64 values");
  if (lf.max_rep != 1) unsupported("column " + lf.path[0] + " is not a one-level list");
  int d0 = 0;
  for (int r : lf.reps) {
    if (r == 2) break;
    d0 += r != 0;
  }
  ChunkReader cr(f, cm, lf);
  out.offsets.resize((size_t)rows + 1);
  out.offsets[0] = 0;
  out.valid.init(rows);
  out.value_bytes.clear();
  out.value_bytes.grow((size_t)std::max<int64_t>(cm.num_values, 1) * 8);
  out.value_bytes.clear();
  out.null_children.clear();
  int64_t row = -1;
  Page pg;
  std::vector<uint32_t> reps, defs, idx;
  const int def_bw = bits_for(lf.max_def);
  while (cr.next(pg)) {
    const int64_t n = pg.num_values;
    HybridDecoder r, d;
    r.init(pg.rep, pg.rep_end, 1);
    d.init(pg.def, pg.def_end, def_bw);
    const bool all_present = d.all_equal((uint64_t)n, (uint32_t)lf.max_def);
    int64_t present = n;
    if (!all_present) {
      reps.resize((size_t)n);
      defs.resize((size_t)n);
      r.get(reps.data(), (uint64_t)n);
      d.init(pg.def, pg.def_end, def_bw);
      d.get(defs.data(), (uint64_t)n);
      present = 0;
      for (int64_t i = 0; i < n; i++) present += defs[(size_t)i] == (uint32_t)lf.max_def;
    }
    const uint8_t *q = pg.values;
    const bool dict = pg.encoding == E_RLE_DICT || pg.encoding == E_PLAIN_DICT;
    if (dict) {
      if (!cr.dict.data || (int64_t)(cr.dict.end - cr.dict.data) < 8 * (int64_t)cr.dict.n) corrupt("INT64 dictionary");
      idx.resize((size_t)present);
      if (present) {
        if (q >= pg.values_end) corrupt("page: no dictionary indices");
        HybridDecoder ix;
        ix.init(q + 1, pg.values_end, q[0]);
        ix.get(idx.data(), (uint64_t)present);
        for (int64_t k = 0; k < present; k++) if (idx[(size_t)k] >= (uint32_t)cr.dict.n) corrupt("dictionary index out of range");
      }
    } else if (pg.encoding == E_PLAIN) {
      if ((int64_t)(pg.values_end - q) < 8 * present) corrupt("page: INT64 values past the end");
    } else {
      unsupported("value encoding " + std::to_string(pg.encoding) + " of column " + lf.path[0]);
    }
    if (out.n_values() + (size_t)n > 0x7fffffffull) unsupported("more than 2^31 outputs in one row group");
    if (present == n) {
      // every level entry is a value (no NULL list, empty list or NULL element on this page): values are copied in one go,
      // the repetition levels only say where rows start
      const size_t base = out.n_values();
      uint8_t *dst = out.value_bytes.grow((size_t)n * 8);
      if (dict) { for (int64_t k = 0; k < n; k++) memcpy(dst + 8 * k, cr.dict.data + 8 * (size_t)idx[(size_t)k], 8); }
      else memcpy(dst, q, (size_t)n * 8);
      r.for_each_zero((uint64_t)n, [&](uint64_t i) {
        if (row >= 0) out.offsets[(size_t)row + 1] = (int32_t)(base + (size_t)i);
        if (++row >= rows) corrupt("column " + lf.path[0] + ": more rows than the row group has");
      });
      if (row < 0) corrupt("column " + lf.path[0] + ": the row group starts inside a row");
      continue;
    }
    int64_t k = 0;
    for (int64_t i = 0; i < n; i++) {
      if (reps[(size_t)i] == 0) {
        if (row >= 0) out.offsets[(size_t)row + 1] = (int32_t)out.n_values();
        if (++row >= rows) corrupt("column " + lf.path[0] + ": more rows than the row group has");
      } else if (row < 0) {
        corrupt("column " + lf.path[0] + ": the row group starts inside a row");
      }
      const int dl = (int)defs[(size_t)i];
      if (dl < d0) { out.valid.clear(row); continue; }
      if (dl == d0) continue;                                  // empty list
      int64_t v = 0;
      if (dl == lf.max_def) {
        if (dict) memcpy(&v, cr.dict.data + 8 * (size_t)idx[(size_t)k], 8);
        else memcpy(&v, q + 8 * k, 8);
        k++;
      } else {
        out.null_children.push_back((uint32_t)out.n_values());  // NULL element
      }
      memcpy(out.value_bytes.grow(8), &v, 8);
    }
  }
  if (row >= 0) out.offsets[(size_t)row + 1] = (int32_t)out.n_values();
  if (row + 1 != rows) corrupt("column " + lf.path[0] + ": fewer rows than the row group has");
  out.child_valid.nulls = 0;
  if (!out.null_children.empty()) {
    out.child_valid.init((int64_t)out.n_values());
    for (uint32_t pos : out.null_children) out.child_valid.clear((int64_t)pos);
  }
}

}  // namespace

namespace {

// ---------------------------------------------------------------- which leaves are txid / height / tweak_key / outputs
struct ColumnMap { int txid = -1, height = -1, tweak = -1, outputs = -1; };
ColumnMap map_columns(const FileMeta &m) {
  ColumnMap c;
  const int n_top = m.schema[0].num_children;
  // by name first (top-level field names), else by position 0..3 like the reference (src/cudasp_extension.cu:159-163)
  int by_name[4] = {-1, -1, -1, -1};
  static const char *names[4] = {"txid", "height", "tweak_key", "outputs"};
  for (size_t l = 0; l < m.leaves.size(); l++)
    for (int k = 0; k < 4; k++)
      if (m.leaves[l].path[0] == names[k]) {
        if (by_name[k] >= 0) unsupported(std::string("column ") + names[k] + " has several leaves");
        by_name[k] = (int)l;
      }
  const bool named = by_name[0] >= 0 && by_name[1] >= 0 && by_name[2] >= 0 && by_name[3] >= 0;
  if (named) { c.txid = by_name[0]; c.height = by_name[1]; c.tweak = by_name[2]; c.outputs = by_name[3]; return c; }
  if (n_top < 4) unsupported("fewer than four columns (txid, height, tweak_key, outputs)");
  int by_pos[4] = {-1, -1, -1, -1};
  for (size_t l = 0; l < m.leaves.size(); l++) {
    const int f = m.leaves[l].top_field;
    if (f >= 0 && f < 4) {
      if (by_pos[f] >= 0) unsupported("one of the first four columns has several leaves");
      by_pos[f] = (int)l;
    }
  }
  for (int k = 0; k < 4; k++) if (by_pos[k] < 0) unsupported("the first four columns are not txid, height, tweak_key, outputs");
  c.txid = by_pos[0]; c.height = by_pos[1]; c.tweak = by_pos[2]; c.outputs = by_pos[3];
  return c;
}

// the four columns have the shapes the decoders handle: checked when the file is opened, before any row is scanned
void validate_columns(const FileMeta &m, const ColumnMap &c) {
  auto blob = [&](int l) {
    const Leaf &lf = m.leaves[(size_t)l];
    if (lf.type != T_BYTE_ARRAY && lf.type != T_FLBA) unsupported("column " + lf.path[0] + " is not BYTE_ARRAY / FIXED_LEN_BYTE_ARRAY");
    if (lf.max_rep != 0 || lf.max_def > 1) unsupported("column " + lf.path[0] + " is nested");
  };
  blob(c.txid);
  blob(c.tweak);
  const Leaf &h = m.leaves[(size_t)c.height];
  if (h.type != T_INT32) unsupported("column " + h.path[0] + " is not INT32");
  if (h.max_rep != 0 || h.max_def > 1) unsupported("column " + h.path[0] + " is nested");
  const Leaf &o = m.leaves[(size_t)c.outputs];
  if (o.type != T_INT64) unsupported("column " + o.path[0] + " does not hold INT64 values");
  if (o.max_rep != 1) unsupported("column " + o.path[0] + " is not a one-level list");
  for (const RowGroupMeta &g : m.row_groups)
    for (int l : {c.txid, c.height, c.tweak, c.outputs}) {
      const int codec = g.cols[(size_t)l].codec;
      if (codec != C_UNCOMPRESSED && codec != C_SNAPPY)
        unsupported("compression codec " + std::to_string(codec) + " (only uncompressed and Snappy pages are decoded)");
    }
  // value counts against row counts: a flat column has one value slot per row, the list column at least one
  for (const RowGroupMeta &g : m.row_groups) {
    if (g.num_rows < 0) corrupt("row group with a negative row count");
    for (int l : {c.txid, c.height, c.tweak})
      if (g.cols[(size_t)l].num_values != g.num_rows) corrupt("column " + m.leaves[(size_t)l].path[0] + ": value count differs from the row count");
    if (g.cols[(size_t)c.outputs].num_values < g.num_rows) corrupt("column " + o.path[0] + ": fewer values than rows");
  }
}

struct Match {
  uint32_t file;
  uint64_t row;
  int32_t height;
  std::string txid, key;
};

struct OpenFile {
  MappedFile map;
  FileMeta meta;
  ColumnMap cols;
  std::vector<int64_t> row_base;   // first row of every row group
};

// CUDASP_TRACE: seconds spent per stage, summed over threads, printed once per cudasp_scan_parquet call
struct StageClock {
  std::atomic<uint64_t> ns[6];
  StageClock() { for (auto &x : ns) x = 0; }
};
StageClock g_clock;
inline uint64_t now_ns() { return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// one row group: decode, hand to the Arrow ingest, keep what matched
struct Scratch {               // one per decoder thread: the decoded columns of the row group in hand, buffers reused from one to the next
  BlobColumn key;                          // a tweak_key page that cannot be staged from where it lies
  ListColumn outs;
  Int32Column height;                      // collect_matches()
  std::vector<std::string> txids;
  std::vector<uint8_t> txid_valid;
  std::vector<uint32_t> defs;
};
// Stages the rows of one row group into the thread's scan; their row numbers count from row_base.  outputs is decoded for the
// whole row group; tweak_key goes page by page: a page of plain 64- or 33-byte values without NULLs is staged from where it
// lies in the file mapping (no intermediate copy), any other page is decoded first.  txid and height are NOT read here: the
// scan does not need them; collect_matches() reads the pages that hold a matching row.
int push_row_group(stg::ColumnScan *scan, const OpenFile &of, size_t g, uint64_t row_base, Scratch &sc) {
  const RowGroupMeta &rg = of.meta.row_groups[g];
  const int64_t rows = rg.num_rows;
  if (rows == 0) return 0;
  if (rows > 0x7fffffff) unsupported("more than 2^31 rows in one row group");
  BlobColumn &key = sc.key;
  ListColumn &outs = sc.outs;
  uint64_t t0 = now_ns(), t1;
  decode_list_column(of.map, rg.cols[(size_t)of.cols.outputs], of.meta.leaves[(size_t)of.cols.outputs], rows, outs);
  t1 = now_ns(); g_clock.ns[1] += t1 - t0; t0 = t1;

  ArrowSchema ts{}, os{}, cs{};
  ArrowSchema *os_children[1] = {&cs};
  ts.format = "z"; ts.name = "tweak_key"; ts.flags = ARROW_FLAG_NULLABLE;
  cs.format = "l"; cs.name = "item"; cs.flags = ARROW_FLAG_NULLABLE;
  os.format = "+l"; os.name = "outputs"; os.flags = ARROW_FLAG_NULLABLE; os.n_children = 1; os.children = os_children;
  ArrowArray ta{}, oa{}, ca{};
  ArrowArray *oa_children[1] = {&ca};
  const void *cb[2] = {outs.child_valid.nulls ? outs.child_valid.bits.data() : nullptr, outs.n_values() ? (const void *)outs.values() : (const void *)""};
  ca.length = (int64_t)outs.n_values(); ca.null_count = outs.child_valid.nulls; ca.n_buffers = 2; ca.buffers = cb;
  const void *ob[2] = {outs.valid.nulls ? outs.valid.bits.data() : nullptr, outs.offsets.data()};
  oa.null_count = outs.valid.nulls; oa.n_buffers = 2; oa.buffers = ob; oa.n_children = 1; oa.children = oa_children;

  const ChunkMeta &kcm = rg.cols[(size_t)of.cols.tweak];
  const Leaf &klf = of.meta.leaves[(size_t)of.cols.tweak];
  check_blob_leaf(kcm, klf);
  ChunkReader cr(of.map, kcm, klf);
  Page pg;
  int64_t row = 0;
  while (cr.next(pg)) {
    const int64_t n = pg.num_values;
    if (row + n > rows) corrupt("column " + klf.path[0] + ": more values than the row group has rows");
    oa.offset = row;                      // this page's rows of the outputs column (an Arrow slice)
    oa.length = n;
    const uint8_t *keys = nullptr;
    int64_t stride = 0;
    int width = 0;
    int rc;
    if (page_key_view(pg, klf, &keys, &stride, &width)) {
      t1 = now_ns(); g_clock.ns[0] += t1 - t0; t0 = t1;
      rc = stg::column_scan_push_keys(scan, keys, stride, width, n, &os, &oa, row_base + (uint64_t)row);
    } else {
      key.offsets.resize((size_t)n + 1);
      key.offsets[0] = 0;
      key.data.clear();
      key.valid.init(n);
      decode_blob_page(cr, pg, klf, key, 0, sc.defs);
      const void *tb[3] = {key.valid.nulls ? key.valid.bits.data() : nullptr, key.offsets.data(), key.data.empty() ? (const void *)"" : (const void *)key.data.data()};
      ta.length = n; ta.null_count = key.valid.nulls; ta.n_buffers = 3; ta.buffers = tb;
      t1 = now_ns(); g_clock.ns[0] += t1 - t0; t0 = t1;
      rc = stg::column_scan_push(scan, &ts, &ta, &os, &oa, row_base + (uint64_t)row);
    }
    t1 = now_ns(); g_clock.ns[4] += t1 - t0; t0 = t1;
    if (rc) return rc;
    row += n;
  }
  if (row != rows) corrupt("column " + klf.path[0] + ": fewer values than the row group has rows");
  return 0;
}

// The matching rows ids[0..n) of one row group (ascending) -> Match records; their keys come from the scan (keys64, 64 bytes
// a match, and widths).  A row with a NULL txid or height, which the reference drops before the GPU
// (src/cudasp_extension.cu:195-197), is dropped here instead: the same result set.
void collect_matches(const OpenFile &of, uint32_t file_index, size_t g, const uint32_t *ids, size_t n, const uint8_t *keys64,
                     const uint8_t *widths, std::vector<Match> &matches, Scratch &sc) {
  const RowGroupMeta &rg = of.meta.row_groups[g];
  const int64_t rows = rg.num_rows;
  Int32Column &height = sc.height;
  uint64_t t0 = now_ns(), t1;
  for (size_t k = 0; k < n; k++)
    if ((int64_t)ids[k] >= rows || (k && ids[k] <= ids[k - 1])) corrupt("a match outside its row group");
  // only the pages that hold a matching row are read (ChunkReader::want)
  fetch_blob_rows(of.map, rg.cols[(size_t)of.cols.txid], of.meta.leaves[(size_t)of.cols.txid], rows, ids, n, sc.txids, sc.txid_valid, sc.defs);
  t1 = now_ns(); g_clock.ns[2] += t1 - t0; t0 = t1;
  decode_int32_column(of.map, rg.cols[(size_t)of.cols.height], of.meta.leaves[(size_t)of.cols.height], rows, height, ids, n);
  t1 = now_ns(); g_clock.ns[3] += t1 - t0;
  for (size_t k = 0; k < n; k++) {
    const size_t r = ids[k];
    if (!sc.txid_valid[k] || !height.valid.get((int64_t)r)) continue;
    Match m;
    m.file = file_index;
    m.row = (uint64_t)of.row_base[g] + r;
    m.height = height.values[r];
    m.txid = std::move(sc.txids[k]);
    m.key.assign(reinterpret_cast<const char *>(keys64) + 64 * k, (size_t)widths[k]);
    matches.push_back(std::move(m));
  }
}

struct ResultOwner : cudasp_parquet_result {           // the public struct and the buffers it points into
  std::vector<uint32_t> file_index_v;
  std::vector<uint64_t> row_in_file_v, txid_offsets_v, key_offsets_v;
  std::vector<int32_t> heights_v;
  std::vector<uint8_t> txids_v, keys_v;
};

}  // namespace

extern "C" int cudasp_scan_parquet(cudasp_query *q, const char *const *paths, int n_paths, int threads, cudasp_parquet_result **out) {
  if (!q || !out || n_paths < 0 || (n_paths && !paths)) return stg::fail(q, CUDASP_ERR_ARG, "cudasp_scan_parquet: NULL argument");
  const uint64_t call_t0 = now_ns();
  // CUDASP_TRACE: when (ns since the call began) the first / last thread had staged its first row group, finished its scan and
  // finished collecting its matches
  std::atomic<uint64_t> first_push_min{~0ull}, first_push_max{0}, scan_done_max{0}, collect_done_max{0};
  auto at_least = [](std::atomic<uint64_t> &a, uint64_t v) { uint64_t c = a.load(); while (c < v && !a.compare_exchange_weak(c, v)) {} };
  auto at_most = [](std::atomic<uint64_t> &a, uint64_t v) { uint64_t c = a.load(); while (c > v && !a.compare_exchange_weak(c, v)) {} };
  *out = nullptr;
  std::vector<OpenFile> files((size_t)n_paths);
  struct Item { uint32_t file; size_t group; };
  std::vector<Item> items;
  uint64_t total_rows = 0;
  try {
    for (int i = 0; i < n_paths; i++) {
      std::string err;
      if (!paths[i] || !files[(size_t)i].map.open_path(paths[i], err)) return stg::fail(q, CUDASP_ERR_IO, err.empty() ? "NULL path" : err.c_str());
      OpenFile &of = files[(size_t)i];
      of.meta = read_footer(of.map);
      of.cols = map_columns(of.meta);
      validate_columns(of.meta, of.cols);
      int64_t base = 0;
      for (size_t g = 0; g < of.meta.row_groups.size(); g++) {
        of.row_base.push_back(base);
        base += of.meta.row_groups[g].num_rows;
        items.push_back(Item{(uint32_t)i, g});
      }
      total_rows += (uint64_t)base;
    }
  } catch (const Unsupported &u) {
    return stg::fail(q, CUDASP_ERR_UNSUPPORTED, ("Parquet feature outside the fast path: " + u.what).c_str());
  } catch (const Corrupt &c) {
    return stg::fail(q, CUDASP_ERR_IO, ("malformed Parquet file: " + c.what).c_str());
  } catch (const std::exception &e) {       // a size read from a damaged footer (length_error, bad_alloc)
    return stg::fail(q, CUDASP_ERR_IO, (std::string("malformed Parquet file: ") + e.what()).c_str());
  }
  const uint64_t opened_ns = now_ns() - call_t0;
  if (threads <= 0) threads = (int)std::max(1u, std::thread::hardware_concurrency());
  threads = (int)std::max<size_t>(1, std::min<size_t>((size_t)threads, items.size()));
  std::vector<std::vector<Match>> found((size_t)threads);
  std::atomic<size_t> next{0};
  std::atomic<int> first_rc{0};
  std::mutex err_mu;
  std::string err_msg;
  // One scan per thread, spanning the row groups the thread takes: their rows fill the same pieces (csrc/arrow_ingest.cpp), so
  // the GPUs see whole pieces whatever the row-group size.  A row's number inside the scan is (item << 32) | row in the group.
  // Every thread starts with ramped-up (small) pieces: measured on one B200 against ramping one thread per device, 10 M rows
  // 0.172 against 0.182 s, 1 M rows 0.025 against 0.027 s (profiles/r02q_parquet_ramp.txt).
  int ramped = 1 << 30;
  if (const char *e = getenv("CUDASP_PARQUET_RAMP")) ramped = atoi(e) < 0 ? 1 << 30 : atoi(e);      // tuning aid: threads that ramp
  auto work = [&](int t) {
    Scratch scratch;
    stg::ColumnScan *scan = nullptr;
    int rc = 0;
    std::string msg;
    try {
      scan = stg::column_scan_open(q, t < ramped);
      bool pushed = false;
      uint64_t collected = 0;          // matches [0, collected) of the scan are in found[t]
      std::vector<uint32_t> in_group;
      // ids [collected, n): runs of one row group each (push order, ascending inside).  Done as the matches come back, between
      // two row groups, so that little of it is left for the end, when the GPUs have nothing else to do.  A row group whose
      // matches come back in two parts has its pages read twice.
      auto collect = [&]() {
        const uint64_t *ids = nullptr;
        const uint8_t *keys64 = nullptr, *widths = nullptr;
        uint64_t n = 0;
        stg::column_scan_matches(scan, &ids, &n, &keys64, &widths);
        for (uint64_t a = collected; a < n;) {
          const uint64_t a0 = a;
          const size_t i = (size_t)(ids[a] >> 32);
          if (i >= items.size()) corrupt("a match outside the files");
          in_group.clear();
          for (; a < n && (size_t)(ids[a] >> 32) == i; a++) in_group.push_back((uint32_t)ids[a]);
          collect_matches(files[items[i].file], items[i].file, items[i].group, in_group.data(), in_group.size(), keys64 + 64 * a0, widths + a0,
                          found[(size_t)t], scratch);
        }
        collected = n;
      };
      for (;;) {
        const size_t i = next.fetch_add(1);
        if (i >= items.size() || first_rc.load()) break;
        rc = push_row_group(scan, files[items[i].file], items[i].group, (uint64_t)i << 32, scratch);
        if (rc) break;
        collect();
        if (!pushed) {
          pushed = true;
          at_most(first_push_min, now_ns() - call_t0);
          at_least(first_push_max, now_ns() - call_t0);
        }
      }
      const uint64_t t0 = now_ns();
      const int rc2 = stg::column_scan_finish(scan);       // also after an error: lets the pieces in flight finish
      g_clock.ns[4] += now_ns() - t0;
      at_least(scan_done_max, now_ns() - call_t0);
      if (rc) {
        msg = cudasp_last_error(q);
      } else if (rc2) {
        rc = rc2;
        msg = cudasp_last_error(q);
      }
      if (!rc && !first_rc.load()) collect();
      at_least(collect_done_max, now_ns() - call_t0);
    } catch (const Unsupported &u) {
      rc = CUDASP_ERR_UNSUPPORTED; msg = "Parquet feature outside the fast path: " + u.what;
    } catch (const Corrupt &c) {
      rc = CUDASP_ERR_IO; msg = "malformed Parquet file: " + c.what;
    } catch (const std::bad_alloc &) {
      rc = CUDASP_ERR_IO; msg = "out of host memory while decoding a row group";
    } catch (const std::exception &e) {     // a size read from a damaged page (length_error)
      rc = CUDASP_ERR_IO; msg = std::string("malformed Parquet file: ") + e.what();
    }
    if (scan) {
      g_clock.ns[5] += stg::column_scan_wait_ns(scan);
      stg::column_scan_close(scan);         // waits for pieces still in flight
    }
    if (rc) {
      std::lock_guard<std::mutex> lk(err_mu);
      int expected = 0;
      if (first_rc.compare_exchange_strong(expected, rc)) err_msg = msg;
    }
  };
  if (threads == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int t = 1; t < threads; t++) th.emplace_back(work, t);
    work(0);
    for (auto &x : th) x.join();
  }
  if (getenv("CUDASP_TRACE")) {
    fprintf(stderr, "[cudasp] scan_parquet: %llu rows, %d threads; seconds summed over threads: tweak_key %.3f, outputs %.3f, txid %.3f, height %.3f, "
                    "staging + GPU %.3f (of which waiting for the GPUs %.3f); since the call began: files open %.4f s, first row group staged %.4f .. %.4f s, scans done %.4f s, "
                    "matches collected %.4f s, threads joined %.4f s\n", (unsigned long long)total_rows, threads, 1e-9 * g_clock.ns[0].exchange(0),
            1e-9 * g_clock.ns[1].exchange(0), 1e-9 * g_clock.ns[2].exchange(0), 1e-9 * g_clock.ns[3].exchange(0), 1e-9 * g_clock.ns[4].exchange(0),
            1e-9 * g_clock.ns[5].exchange(0), 1e-9 * opened_ns, 1e-9 * first_push_min.load(), 1e-9 * first_push_max.load(), 1e-9 * scan_done_max.load(), 1e-9 * collect_done_max.load(),
            1e-9 * (now_ns() - call_t0));
  }
  if (first_rc.load()) return stg::fail(q, first_rc.load(), err_msg.c_str());
  std::vector<Match> all;
  for (auto &v : found) { all.insert(all.end(), std::make_move_iterator(v.begin()), std::make_move_iterator(v.end())); }
  std::sort(all.begin(), all.end(), [](const Match &a, const Match &b) { return a.file != b.file ? a.file < b.file : a.row < b.row; });
  ResultOwner *r = new ResultOwner();
  memset(static_cast<cudasp_parquet_result *>(r), 0, sizeof(cudasp_parquet_result));
  r->txid_offsets_v.push_back(0);
  r->key_offsets_v.push_back(0);
  for (const Match &m : all) {
    r->file_index_v.push_back(m.file);
    r->row_in_file_v.push_back(m.row);
    r->heights_v.push_back(m.height);
    r->txids_v.insert(r->txids_v.end(), m.txid.begin(), m.txid.end());
    r->keys_v.insert(r->keys_v.end(), m.key.begin(), m.key.end());
    r->txid_offsets_v.push_back(r->txids_v.size());
    r->key_offsets_v.push_back(r->keys_v.size());
  }
  cudasp_parquet_result &pub = *r;
  pub.n_matches = all.size();
  pub.rows_scanned = total_rows;
  pub.file_index = r->file_index_v.data();
  pub.row_in_file = r->row_in_file_v.data();
  pub.heights = r->heights_v.data();
  pub.txids = r->txids_v.data();
  pub.txid_offsets = r->txid_offsets_v.data();
  pub.tweak_keys = r->keys_v.data();
  pub.key_offsets = r->key_offsets_v.data();
  *out = r;
  return CUDASP_OK;
}

extern "C" void cudasp_parquet_result_free(cudasp_parquet_result *r) {
  delete static_cast<ResultOwner *>(r);
}
