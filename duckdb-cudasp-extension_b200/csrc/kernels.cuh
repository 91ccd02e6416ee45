// kernels.cuh -- the per-row scan pipeline as sm_100a kernels (thin wrappers over rowfn.cuh).
//
// Replaces the reference's stage list in RunBatchScanKernels (src/cudasp_gpu.cu:673-966):
//   processScalarPoint + scalarMulByCombinedDAA   -> k_ladder        (K1)
//   (one block-serial binary-GCD inversion per bit)-> k_batch_inv     (K2/K4/K6, Montgomery's trick)
//   SerializeToCompressedSEC1Kernel
//   + ComputeTaggedHashesKernel + host swizzle
//   + FixedPointMultiplyKernel + the "+ B_spend"  -> k_hash_fixedbase (K3)
//   CheckMatchesWithLabelsKernel                  -> k_final_base (K5), k_final_labels (K7)
// Intermediates are arrays of 32-byte field elements (one per row) in HBM; at ~300 B/row they are
// three orders of magnitude below the HBM roofline, the kernels are integer-multiply bound.
#pragma once
#include "rowfn.cuh"

namespace csp {

// ---------------------------------------------------------------- K1: S = scan_sk * tweak
// One row per thread.  Table {1,3,5,7}*P (+ beta*x for the lambda half) lives in shared memory,
// limb-major so that a warp's access to one limb of one entry is 32 consecutive words: the entry
// index is the same for every thread (same scan key), so there are no bank conflicts.
// smem words: [entry 0..3][component 0=x,1=beta*x,2=y][limb 0..7][thread]
template <int BLOCK>
struct SmemTab {
  uint32_t *base;   // already offset by threadIdx.x
  __device__ __forceinline__ uint32_t get(uint32_t e, uint32_t c, int l) const { return base[((e * 3 + c) * 8 + l) * BLOCK]; }
  __device__ __forceinline__ void set(uint32_t e, uint32_t c, int l, uint32_t v) { base[((e * 3 + c) * 8 + l) * BLOCK] = v; }
};
static const int LADDER_SMEM_WORDS_PER_THREAD = 4 * 3 * 8;

// KEYFMT 0: 64-byte keys (x_LE || y_LE, the reference's layout).  KEYFMT 1: 33-byte SEC1 compressed keys,
// decompressed here (decompress_tweak); `invalid[row]` then tells K5/K7 that the row can never match.
template <int BLOCK, int MINB, int SYNC, int KEYFMT>
__global__ void __launch_bounds__(BLOCK, MINB) k_ladder(const uint8_t *__restrict__ tweaks, fe *__restrict__ outX,
                                                        fe *__restrict__ outY, fe *__restrict__ outZ, uint32_t rows,
                                                        const __grid_constant__ LadderParams prm, uint8_t *__restrict__ invalid) {
  extern __shared__ uint32_t smem_tab[];
  uint32_t row = blockIdx.x * BLOCK + threadIdx.x;
  const bool live = row < rows;
  if (!SYNC && !live) return;
  if (!live) row = rows - 1;            // SYNC: every thread reaches every barrier; dead threads redo the last row
  SmemTab<BLOCK> tab;
  tab.base = smem_tab + threadIdx.x;
  ge P;
  if (KEYFMT == 0) {
    load_fe_ro(P.x, reinterpret_cast<const fe *>(tweaks + (size_t)row * 64));
    load_fe_ro(P.y, reinterpret_cast<const fe *>(tweaks + (size_t)row * 64 + 32));
  } else {
    uint8_t sec1[33];
    const uint8_t *src = tweaks + (size_t)row * 33;
#pragma unroll
    for (int i = 0; i < 33; i++) sec1[i] = __ldg(src + i);
    const bool ok = decompress_tweak(P, sec1);
    if (live) invalid[row] = ok ? 0 : 1;
  }
  gej acc;
  ladder_row<SmemTab<BLOCK>, SYNC>(acc, P, prm, tab);
  if (!live) return;
  store_fe(outX + row, acc.x);
  store_fe(outY + row, acc.y);
  store_fe(outZ + row, acc.z);
}

// The non-template kernels below are DEFINED only in the translation unit that launches them (k_stages.cu).
#if defined(CSP_DEFINE_STAGE_KERNELS)
// ---------------------------------------------------------------- K2/K4/K6: batch inversion
__global__ void __launch_bounds__(128) k_batch_inv(fe *__restrict__ vals, fe *__restrict__ scratch, uint32_t n, uint32_t T) {
  batch_inv_thread(vals, scratch, n, T, blockIdx.x * blockDim.x + threadIdx.x);
}

#endif  // CSP_DEFINE_STAGE_KERNELS

// ---------------------------------------------------------------- K3: hash, t*G, + B_spend (in place)
template <int BLOCK, int MINB>
__global__ void __launch_bounds__(BLOCK, MINB) k_hash_fixedbase(fe *__restrict__ X, fe *__restrict__ Y, fe *__restrict__ Z,
                                                                uint32_t rows, const ge *__restrict__ gtab,
                                                                const __grid_constant__ FinalParams fp,
                                                                uint8_t *__restrict__ dbg_S, uint8_t *__restrict__ dbg_t,
                                                                fe *__restrict__ keep_sx, uint8_t *__restrict__ keep_sodd) {
  const uint32_t row = blockIdx.x * BLOCK + threadIdx.x;
  if (row >= rows) return;
  fe x, y, zi, sx, sy, t;
  load_fe(x, X + row);
  load_fe(y, Y + row);
  load_fe(zi, Z + row);   // already inverted by k_batch_inv
  gej acc;
  const bool dbg = dbg_S || dbg_t || keep_sx;
  hash_fixedbase_row(acc, x, y, zi, gtab, fp, dbg ? &sx : nullptr, dbg ? &sy : nullptr, dbg ? &t : nullptr);
  if (keep_sx) {          // match details enabled: the compressed shared secret outlives this kernel (33 B/row)
    store_fe(keep_sx + row, sx);
    keep_sodd[row] = (uint8_t)(sy.v[0] & 1u);
  }
  if (dbg_S) {
    uint32_t *o = reinterpret_cast<uint32_t *>(dbg_S + (size_t)row * 64);
#pragma unroll
    for (int l = 0; l < 8; l++) { o[l] = sx.v[l]; o[8 + l] = sy.v[l]; }
  }
  if (dbg_t) {
    uint8_t *o = dbg_t + (size_t)row * 32;
#pragma unroll
    for (int l = 0; l < 8; l++) {
      uint32_t wv = t.v[7 - l];
      o[4 * l] = (uint8_t)(wv >> 24); o[4 * l + 1] = (uint8_t)(wv >> 16); o[4 * l + 2] = (uint8_t)(wv >> 8); o[4 * l + 3] = (uint8_t)wv;
    }
  }
  store_fe(X + row, acc.x);
  store_fe(Y + row, acc.y);
  store_fe(Z + row, acc.z);
}

// ---------------------------------------------------------------- match helpers
__device__ __forceinline__ bool outputs_contain(const int64_t *__restrict__ outputs, uint32_t lo, uint32_t hi, int64_t cand) {
  bool hit = false;
  for (uint32_t j = lo; j < hi; j++) hit |= (__ldg(outputs + j) == cand);   // src/cudasp_gpu.cu:29-41
  return hit;
}
// warp-aggregated append of matching row ids: ballot + prefix popcount, one atomic per warp
__device__ __forceinline__ void append_match(bool hit, uint32_t row, uint32_t *__restrict__ match_ids, uint32_t *__restrict__ n_matches) {
  const uint32_t m = __ballot_sync(0xffffffffu, hit);
  if (m == 0) return;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t leader = __ffs(m) - 1;
  uint32_t base = 0;
  if (lane == leader) base = atomicAdd(n_matches, __popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (hit) match_ids[base + __popc(m & ((1u << lane) - 1u))] = row;
}

// ---------------------------------------------------------------- K5: base candidate
// L == 0: matches are appended here.  L > 0: writes x0, y0 (into X, Y), the flag, and the L
// denominators d_i = xL_i - x0 (array D, label-major: D[i*rows + row]) for the shared inversion K6.
template <int BLOCK>
__global__ void __launch_bounds__(BLOCK) k_final_base(fe *__restrict__ X, fe *__restrict__ Y, const fe *__restrict__ Zinv,
                                                      uint32_t rows, const int64_t *__restrict__ outputs,
                                                      const uint32_t *__restrict__ out_off, const ge *__restrict__ labels,
                                                      uint32_t n_labels, fe *__restrict__ D, uint8_t *__restrict__ flags,
                                                      uint32_t *__restrict__ match_ids, uint32_t *__restrict__ n_matches,
                                                      int64_t *__restrict__ cands_out, uint8_t *__restrict__ dbg_P0,
                                                      const uint8_t *__restrict__ invalid) {
  const uint32_t row = blockIdx.x * BLOCK + threadIdx.x;   // grid is a whole number of warps
  const bool live = row < rows;
  bool hit = false;
  if (live) {
    fe zi, X1, Y1, x0, y0;
    load_fe(zi, Zinv + row);
    load_fe(X1, X + row);
    const bool need_y = n_labels || dbg_P0;
    if (need_y) load_fe(Y1, Y + row);
    const int64_t cand = base_candidate(x0, need_y ? &y0 : nullptr, X1, Y1, zi);
    if (cands_out) cands_out[(size_t)row * (1 + n_labels)] = cand;
    if (outputs) hit = outputs_contain(outputs, out_off[row], out_off[row + 1], cand);
    if (invalid && invalid[row]) hit = false;      // malformed compressed key: never matches
    if (need_y) {
      store_fe(X + row, x0);
      store_fe(Y + row, y0);
      if (dbg_P0) {
        uint32_t *o = reinterpret_cast<uint32_t *>(dbg_P0 + (size_t)row * 64);
#pragma unroll
        for (int l = 0; l < 8; l++) { o[l] = x0.v[l]; o[8 + l] = y0.v[l]; }
      }
      for (uint32_t i = 0; i < n_labels; i++) {
        fe lx, dd;
        load_fe_ro(lx, &labels[i].x);
        fe_sub(dd, lx, x0);
        store_fe(D + (size_t)i * rows + row, dd);
      }
      if (n_labels) flags[row] = hit ? 1 : 0;
    }
  }
  if (n_labels == 0 && match_ids) append_match(live && hit, row, match_ids, n_matches);
}

// ---------------------------------------------------------------- K7: label candidates
template <int BLOCK>
__global__ void __launch_bounds__(BLOCK) k_final_labels(const fe *__restrict__ X0, const fe *__restrict__ Y0,
                                                        const fe *__restrict__ Dinv, uint32_t rows,
                                                        const int64_t *__restrict__ outputs, const uint32_t *__restrict__ out_off,
                                                        const ge *__restrict__ labels, uint32_t n_labels,
                                                        const uint8_t *__restrict__ flags, uint32_t *__restrict__ match_ids,
                                                        uint32_t *__restrict__ n_matches, int64_t *__restrict__ cands_out,
                                                        const uint8_t *__restrict__ invalid) {
  const uint32_t row = blockIdx.x * BLOCK + threadIdx.x;
  const bool live = row < rows;
  bool hit = false;
  if (live) {
    hit = flags[row] != 0;
    fe x0, y0;
    load_fe(x0, X0 + row);
    load_fe(y0, Y0 + row);
    const uint32_t lo = outputs ? out_off[row] : 0, hi = outputs ? out_off[row + 1] : 0;
    for (uint32_t i = 0; i < n_labels; i++) {
      fe lx, ly, di;
      load_fe_ro(lx, &labels[i].x);
      load_fe_ro(ly, &labels[i].y);
      load_fe(di, Dinv + (size_t)i * rows + row);
      const int64_t cand = label_candidate(x0, y0, lx, ly, di);
      if (cands_out) cands_out[(size_t)row * (1 + n_labels) + 1 + i] = cand;
      if (outputs) hit |= outputs_contain(outputs, lo, hi, cand);
    }
    if (invalid && invalid[row]) hit = false;
  }
  if (match_ids) append_match(live && hit, row, match_ids, n_matches);
}

#if defined(CSP_DEFINE_STAGE_KERNELS)
// ---------------------------------------------------------------- field KAT kernel (tests)
__global__ void k_test_field(int op, const fe *__restrict__ a, const fe *__restrict__ b, fe *__restrict__ out, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  fe x, y, r;
  load_fe(x, a + i);
  load_fe(y, b + i);
  switch (op) {
    case 0: fe_mul(r, x, y); break;
    case 1: fe_sqr(r, x); break;
    case 2: fe_add(r, x, y); break;
    case 3: fe_sub(r, x, y); break;
    case 4: fe_inv(r, x); break;
    default: r = x; break;
  }
  fe_normalize(r);
  store_fe(out + i, r);
}

// ---------------------------------------------------------------- K8 (opt-in): match details
// One thread per MATCHED row (a handful per batch): the BIP-352 k-loop of details_row.  The grid is fixed and strides
// over the device-side match count, so no host round trip is needed to size it.  One warp per block spreads the few
// rows over the SMs; each record is appended with one atomic.
__global__ void __launch_bounds__(32) k_details(const fe *__restrict__ SX, const uint8_t *__restrict__ SODD,
                                                const uint32_t *__restrict__ match_ids, const uint32_t *__restrict__ n_matches,
                                                const int64_t *__restrict__ outputs, const uint32_t *__restrict__ out_off,
                                                const ge *__restrict__ labels, uint32_t n_labels, const ge *__restrict__ gtab,
                                                const __grid_constant__ FinalParams fp, MatchDetail *__restrict__ out,
                                                uint32_t *__restrict__ n_out, uint32_t cap) {
  const uint32_t n = *n_matches;
  for (uint32_t m = blockIdx.x * 32 + threadIdx.x; m < n; m += gridDim.x * 32) {
    const uint32_t row = match_ids[m];
    fe sx;
    load_fe(sx, SX + row);
    details_row(sx, SODD[row], outputs, out_off[row], out_off[row + 1], labels, n_labels, gtab, fp,
                [&](uint32_t k, uint32_t oi, int32_t li, const fe &t) {
                  const uint32_t at = atomicAdd(n_out, 1u);
                  if (at >= cap) return;          // counted, not stored: the host reports the overflow
                  MatchDetail d;
                  d.row = row; d.k = k; d.output_index = oi; d.label_index = li;
#pragma unroll
                  for (int l = 0; l < 8; l++) {
                    const uint32_t wv = t.v[7 - l];
                    d.t_k[4 * l] = (uint8_t)(wv >> 24); d.t_k[4 * l + 1] = (uint8_t)(wv >> 16);
                    d.t_k[4 * l + 2] = (uint8_t)(wv >> 8); d.t_k[4 * l + 3] = (uint8_t)wv;
                  }
                  out[at] = d;
                });
  }
}

// ---------------------------------------------------------------- fixed-base table, built where it is used
// Replaces the reference's 256 x 64 B __constant__ table of 2^i*G uploaded by Solver::initialize (gsv.h:1094-1122).
// One thread per GTAB_EPT consecutive entries of one window (rowfn.cuh gtab_chunk); runs once per process per device.
__global__ void __launch_bounds__(128) k_build_gtab(ge *__restrict__ table, const ge *__restrict__ firsts,
                                                    const ge *__restrict__ bases, uint32_t *__restrict__ bad) {
  const uint64_t gid = (uint64_t)blockIdx.x * 128 + threadIdx.x;
  const uint32_t chunks_per_window = GTAB_ENTRIES / GTAB_EPT;
  const uint32_t w = (uint32_t)(gid / chunks_per_window);
  if (w >= (uint32_t)GTAB_NWIN) return;
  const uint32_t d0 = (uint32_t)(gid % chunks_per_window) * GTAB_EPT;
  if (w == (uint32_t)GTAB_NWIN - 1 && d0 >= (uint32_t)GTAB_TOP_ENTRIES) return;   // the top window is short
  ge f, b;
  load_fe_ro(f.x, &firsts[w].x); load_fe_ro(f.y, &firsts[w].y);
  load_fe_ro(b.x, &bases[w].x); load_fe_ro(b.y, &bases[w].y);
  if (!gtab_chunk(table + (size_t)w * GTAB_ENTRIES + d0, f, b, d0)) atomicAdd(bad, 1u);
}

#endif  // CSP_DEFINE_STAGE_KERNELS

}  // namespace csp
