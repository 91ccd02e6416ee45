// piece_plan.hpp -- how the synchronous call cuts a host table into pieces (cudasp_capi.cu scan_batch_impl).
//
// Pieces are at most `piece` rows (whole waves of the ladder kernel).  The first pieces ramp up -- small, small*2, ... one
// round of devices per size -- so that every GPU starts after a short first copy instead of after a full piece; when the table
// is long enough to have a middle, the last pieces mirror that ramp, so that the tail the GPUs work off alone after the last
// copy has been enqueued is a small piece, not a full one.  Host-only, shared with the CPU tests (tests/hostcheck).
#pragma once
#include <algorithm>
#include <cstdint>
#include <vector>

namespace csp {

// -> starts[0] = 0 < starts[1] < ... < starts[n] = rows; piece i is rows [starts[i], starts[i + 1])
inline std::vector<uint32_t> piece_plan(uint32_t rows, uint32_t piece, uint32_t n_devices, uint32_t ramp_div = 8, uint32_t min_piece = 8192) {
  std::vector<uint32_t> starts(1, 0u);
  if (rows == 0) return starts;
  piece = std::max<uint32_t>(piece, 1);
  n_devices = std::max<uint32_t>(n_devices, 1);
  const uint32_t small = std::min(piece, std::max<uint32_t>(piece / std::max<uint32_t>(ramp_div, 1), min_piece));
  std::vector<uint32_t> head, tail;
  uint64_t used = 0;
  for (uint64_t sz = small; sz < piece; sz *= 2)
    for (uint32_t d = 0; d < n_devices; d++) { head.push_back((uint32_t)sz); used += sz; }
  if (used * 2 + (uint64_t)piece * n_devices <= rows) tail.assign(head.rbegin(), head.rend());
  uint64_t tail_rows = 0;
  for (uint32_t v : tail) tail_rows += v;
  uint64_t at = 0;
  size_t hi = 0;
  while (at < rows - tail_rows) {
    const uint32_t want = hi < head.size() ? head[hi++] : piece;
    at += std::min<uint64_t>(want, rows - tail_rows - at);
    starts.push_back((uint32_t)at);
  }
  for (uint32_t v : tail) { at += v; starts.push_back((uint32_t)at); }
  return starts;
}

}  // namespace csp
