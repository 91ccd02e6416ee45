// rowfn.cuh -- the per-row stages of the scan as __host__ __device__ functions.
//
// The kernels in kernels.cuh are thin wrappers (thread <-> row mapping, memory layout, match
// compaction) around these functions; tests/hostcheck compiles the very same functions with g++
// so that the schedule logic (GLV/wNAF op list, isomorphic odd-multiples table, offset fixed-base
// windows, chord/tangent label rule, batch inversion) is verified on the CPU against the reference vectors
// before any GPU time is spent.  That host build is test infrastructure only: the C ABI never
// calls it and has no CPU path.
#pragma once
#include <stdint.h>
#include "fe.cuh"
#include "ge.cuh"
#include "sha256.cuh"

namespace csp {

static const int LADDER_MAX_OPS_DEV = 240;
// Fixed-base window width.  Every bit of width trades HBM for mixed additions in K3: 13 bits = 20 windows, 10 MB
// (L2-resident), 19 additions; 22 bits = 12 windows, 2.95 GB of the 180 GB HBM3e, 11 additions -- K3 0.647 -> 0.414 ms per
// 300 K rows on B200 (profiles/r01k_window_width.jsonl).  The table is built on the device (kernels.cuh k_build_gtab,
// 0.17 s once per process per device), entries are read once per addition with a one-ahead prefetch.
#ifndef CSP_GTAB_WBITS
#define CSP_GTAB_WBITS 22
#endif
static const int GTAB_WBITS = CSP_GTAB_WBITS;                          // fixed-base window width
static const int GTAB_NWIN = (256 + GTAB_WBITS - 1) / GTAB_WBITS;      // windows; NWIN - 1 mixed additions per row
static const int GTAB_ENTRIES = 1 << GTAB_WBITS;                       // entries per window (stride of the table)
static const int GTAB_TOP_BITS = 256 - GTAB_WBITS * (GTAB_NWIN - 1);   // the last window only sees the scalar's top bits
static const int GTAB_TOP_ENTRIES = 1 << GTAB_TOP_BITS;                // ... so only that many of its entries exist
static const size_t GTAB_TOTAL = (size_t)(GTAB_NWIN - 1) * GTAB_ENTRIES + GTAB_TOP_ENTRIES;

// digit w of the 256-bit scalar t (little-endian limbs): bits [w*WBITS, w*WBITS + WBITS)
CSP_HD uint32_t gtab_digit(const fe &t, int w) {
  const int bit = w * GTAB_WBITS, limb = bit >> 5, sh = bit & 31;
  uint32_t v = t.v[limb] >> sh;
  if (sh + GTAB_WBITS > 32 && limb < 7) v |= t.v[limb + 1] << (32 - sh);
  return v & (uint32_t)(GTAB_ENTRIES - 1);
}

struct LadderParams {          // kernel parameter (constant bank): the per-query op list
  uint8_t ops[LADDER_MAX_OPS_DEV];
  int32_t nops;
  uint32_t beta[8];
};

struct FinalParams {           // kernel parameter: per-query constants of K3
  uint32_t midstate[8];        // SHA-256 chaining value after TH||TH
  fe spend_adj_x, spend_adj_y; // B_spend minus the fixed-base table's offset point
  int32_t spend_adj_inf;       // that difference is the point at infinity: skip the add
  int32_t zero_scalar;         // scan key = 0 (mod n): S is the reference's all-zero pair
};

// ---------------------------------------------------------------- row <-> memory
CSP_HD void load_fe(fe &r, const fe *p) {
#if defined(__CUDA_ARCH__)
  const uint4 *q = reinterpret_cast<const uint4 *>(p);
  uint4 a = q[0], b = q[1];
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
#else
  r = *p;
#endif
}
CSP_HD void load_fe_ro(fe &r, const fe *p) {   // read-only path (tables, inputs)
#if defined(__CUDA_ARCH__)
  const uint4 *q = reinterpret_cast<const uint4 *>(p);
  uint4 a = __ldg(q), b = __ldg(q + 1);
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
#else
  r = *p;
#endif
}
CSP_HD void store_fe(fe *p, const fe &r) {
#if defined(__CUDA_ARCH__)
  uint4 *q = reinterpret_cast<uint4 *>(p);
  q[0] = make_uint4(r.v[0], r.v[1], r.v[2], r.v[3]);
  q[1] = make_uint4(r.v[4], r.v[5], r.v[6], r.v[7]);
#else
  *p = r;
#endif
}

// ---------------------------------------------------------------- compressed tweak keys
// 33-byte SEC1: prefix 02 / 03 (parity of y) || x big-endian.  Same result as the reference's CPU recipe
// (scripts/decompress_tweak_key.py:26-46): y = (x^3 + 7)^((p+1)/4), negated when its parity differs from the prefix.
// Returns false -- the row can never match -- for a bad prefix, x >= p, or x not on the curve.
CSP_HD bool decompress_tweak(ge &P, const uint8_t *sec1) {
  const uint32_t prefix = sec1[0];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int l = 0; l < 8; l++) {
    const uint8_t *b = sec1 + 1 + 4 * (7 - l);
    P.x.v[l] = ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | (uint32_t)b[3];
  }
  fe xc = P.x;
  fe_normalize(xc);
  bool ok = (prefix == 2u || prefix == 3u);
  bool canonical = true;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int l = 0; l < 8; l++) canonical &= (xc.v[l] == P.x.v[l]);      // x >= p is not a valid encoding
  ok &= canonical;
  fe y2, seven, y, chk;
  fe_sqr(y2, P.x);
  fe_mul(y2, y2, P.x);
  fe_set_u32(seven, 7);
  fe_add(y2, y2, seven);
  fe_sqrt_candidate(y, y2);
  fe_sqr(chk, y);
  ok &= fe_equal(chk, y2);
  fe_normalize(y);
  if ((y.v[0] & 1u) != (prefix & 1u)) {
    fe_neg(y, y);
    fe_normalize(y);
  }
  P.y = y;
  return ok;
}

// ---------------------------------------------------------------- K1 body: S = scan_sk * P
// Tab: storage for the 4-entry odd-multiples table, components 0 = x, 1 = beta*x, 2 = y;
//      needs  uint32_t get(e, c, l)  and  void set(e, c, l, v).
// Result: Jacobian S (not normalised to Z = 1; the caller batches the inversion).
// SYNC != 0 (device only): a block-wide barrier before every op keeps the block's warps on the same
// instructions, so that they share instruction-cache lines (the unrolled loop body is larger than the I-cache).
template <class Tab, int SYNC = 0>
CSP_HD void ladder_row(gej &acc, const ge &P, const LadderParams &prm, Tab &tab) {
  // d = 2P (Jacobian).  Work on the curve isomorphic by dz, where d is affine.
  gej d;
  {
    gej pj;
    pj.x = P.x; pj.y = P.y; fe_set_u32(pj.z, 1);
    gej_dbl(d, pj);
  }
  fe dz2, dz3;
  fe_sqr(dz2, d.z);
  fe_mul(dz3, dz2, d.z);
  gej e[4];
  fe h1, h2, h3;
  fe_mul(e[0].x, P.x, dz2);
  fe_mul(e[0].y, P.y, dz3);
  fe_set_u32(e[0].z, 1);
  gej_add_ge(e[1], e[0], d.x, d.y, &h1);   // 3P'
  gej_add_ge(e[2], e[1], d.x, d.y, &h2);   // 5P'
  gej_add_ge(e[3], e[2], d.x, d.y, &h3);   // 7P'
  // bring all four to the common Z = e[3].z:  scale entry i by s_i = Z3/Z_i
  //   s_2 = h3, s_1 = h2*h3, s_0 = h1*h2*h3 = Z3
  fe beta;
#pragma unroll
  for (int l = 0; l < 8; l++) beta.v[l] = prm.beta[l];
  {
    fe s, s2, s3, x, y, bx;
    // entry 3: as is
    fe_mul(bx, e[3].x, beta);
#pragma unroll
    for (int l = 0; l < 8; l++) { tab.set(3, 0, l, e[3].x.v[l]); tab.set(3, 1, l, bx.v[l]); tab.set(3, 2, l, e[3].y.v[l]); }
    // entry 2: s = h3
    s = h3;
    fe_sqr(s2, s); fe_mul(s3, s2, s);
    fe_mul(x, e[2].x, s2); fe_mul(y, e[2].y, s3); fe_mul(bx, x, beta);
#pragma unroll
    for (int l = 0; l < 8; l++) { tab.set(2, 0, l, x.v[l]); tab.set(2, 1, l, bx.v[l]); tab.set(2, 2, l, y.v[l]); }
    // entry 1: s = h2*h3
    fe_mul(s, h2, h3);
    fe_sqr(s2, s); fe_mul(s3, s2, s);
    fe_mul(x, e[1].x, s2); fe_mul(y, e[1].y, s3); fe_mul(bx, x, beta);
#pragma unroll
    for (int l = 0; l < 8; l++) { tab.set(1, 0, l, x.v[l]); tab.set(1, 1, l, bx.v[l]); tab.set(1, 2, l, y.v[l]); }
    // entry 0: s = Z3 (e[0].z = 1)
    s = e[3].z;
    fe_sqr(s2, s); fe_mul(s3, s2, s);
    fe_mul(x, e[0].x, s2); fe_mul(y, e[0].y, s3); fe_mul(bx, x, beta);
#pragma unroll
    for (int l = 0; l < 8; l++) { tab.set(0, 0, l, x.v[l]); tab.set(0, 1, l, bx.v[l]); tab.set(0, 2, l, y.v[l]); }
  }
  // Z of the final result must be multiplied by zscale = Z3 * dz
  fe zscale;
  fe_mul(zscale, e[3].z, d.z);

  fe_set_zero(acc.x); fe_set_zero(acc.y); fe_set_zero(acc.z);
  for (int i = 0; i < prm.nops; i++) {
    const uint32_t op = prm.ops[i];
#if defined(__CUDA_ARCH__)
    if (SYNC) __syncthreads();
#endif
    if (op & 0x80u) {
      gej_dbl(acc, acc);
    } else {
      const uint32_t en = op & 3u, comp = (op >> 2) & 1u;
      fe qx, qy;
#pragma unroll
      for (int l = 0; l < 8; l++) { qx.v[l] = tab.get(en, comp, l); qy.v[l] = tab.get(en, 2, l); }
      if (op & 0x08u) fe_neg(qy, qy);
      if (op & 0x40u) {
        acc.x = qx; acc.y = qy; fe_set_u32(acc.z, 1);
      } else {
        gej_add_ge(acc, acc, qx, qy);
      }
    }
  }
  fe_mul(acc.z, acc.z, zscale);
}

// ---------------------------------------------------------------- K2/K4/K6 body: batch inversion
// vals[i] <- vals[i]^-1 (0 stays 0) for i = tid, tid+T, tid+2T, ...  Prefix products go to
// scratch, one Fermat inversion, then the back pass (Montgomery's trick).
CSP_HD void batch_inv_thread(fe *vals, fe *scratch, uint32_t n, uint32_t T, uint32_t tid) {
  if (tid >= T || tid >= n) return;
  fe acc;
  fe_set_u32(acc, 1);
  uint32_t last = tid;
  for (uint64_t i = tid; i < n; i += T) {
    fe v;
    load_fe(v, vals + i);
    store_fe(scratch + i, acc);
    if (!fe_is_zero(v)) fe_mul(acc, acc, v);
    last = (uint32_t)i;
  }
  fe inv;
  fe_inv(inv, acc);
  for (uint32_t i = last;; i -= T) {
    fe v, pre, out;
    load_fe(v, vals + i);
    load_fe(pre, scratch + i);
    if (fe_is_zero(v)) {
      fe_set_zero(out);
    } else {
      fe_mul(out, inv, pre);
      fe_mul(inv, inv, v);
    }
    store_fe(vals + i, out);
    if (i == tid) break;
  }
}

// ---------------------------------------------------------------- fixed-base table builder (once per process per device)
// Entries d0 .. d0+GTAB_EPT-1 of one window:  out[i] = (d0 + i) * base + first, affine and canonical, where
// base = 2^(WBITS*w)*G and first = r_w*G come from the host (host_ec.hpp gtab_seeds).  d0*base by double-and-add,
// then a chain of mixed additions and ONE Fermat inversion for the chunk (Montgomery's trick).  Returns false when
// an incomplete addition met P = +-Q (cannot happen with the published offsets; the caller refuses to continue).
static const int GTAB_EPT = 16;
static_assert(GTAB_TOP_ENTRIES >= GTAB_EPT && GTAB_ENTRIES % GTAB_EPT == 0, "window shape vs builder chunk");
CSP_HD bool gtab_chunk(ge *out, const ge &first, const ge &base, uint32_t d0) {
  gej pts[GTAB_EPT];
  fe h;
  bool ok = true;
  if (d0 == 0) {
    pts[0].x = first.x; pts[0].y = first.y; fe_set_u32(pts[0].z, 1);
  } else {
    gej acc;
    acc.x = base.x; acc.y = base.y; fe_set_u32(acc.z, 1);
    int top = 31;
    while (!((d0 >> top) & 1u)) top--;
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
    for (int b = top - 1; b >= 0; b--) {
      gej_dbl(acc, acc);
      if ((d0 >> b) & 1u) {
        gej_add_ge(acc, acc, base.x, base.y, &h);
        ok &= !fe_is_zero(h);
      }
    }
    gej_add_ge(pts[0], acc, first.x, first.y, &h);
    ok &= !fe_is_zero(h);
  }
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
  for (int i = 1; i < GTAB_EPT; i++) {
    gej_add_ge(pts[i], pts[i - 1], base.x, base.y, &h);
    ok &= !fe_is_zero(h);
  }
  fe pre[GTAB_EPT];
  pre[0] = pts[0].z;
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
  for (int i = 1; i < GTAB_EPT; i++) fe_mul(pre[i], pre[i - 1], pts[i].z);
  fe inv;
  fe_inv(inv, pre[GTAB_EPT - 1]);
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
  for (int i = GTAB_EPT - 1; i >= 0; i--) {
    fe zi, zi2, zi3, x, y;
    if (i) {
      fe_mul(zi, inv, pre[i - 1]);
      fe_mul(inv, inv, pts[i].z);
    } else {
      zi = inv;
    }
    fe_sqr(zi2, zi);
    fe_mul(zi3, zi2, zi);
    fe_mul(x, pts[i].x, zi2);
    fe_mul(y, pts[i].y, zi3);
    fe_normalize(x);
    fe_normalize(y);
    store_fe(&out[i].x, x);
    store_fe(&out[i].y, y);
  }
  return ok;
}

// acc = t*G + B_spend from the offset window table (see the comment below); t is taken as a 256-bit integer, not reduced.
CSP_HD void fixedbase_plus_spend(gej &acc, const fe &t, const ge *gtab, const FinalParams &fp) {
  // all digits first (static limb indices: t stays in registers), then a one-entry-ahead prefetch pipeline
  uint32_t dg[GTAB_NWIN];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int w = 0; w < GTAB_NWIN; w++) dg[w] = gtab_digit(t, w);
  {
    const ge *q = gtab + dg[0];
    load_fe_ro(acc.x, &q->x);
    load_fe_ro(acc.y, &q->y);
    fe_set_u32(acc.z, 1);
  }
  fe qx, qy;
  {
    const ge *q = gtab + GTAB_ENTRIES + dg[1];
    load_fe_ro(qx, &q->x);
    load_fe_ro(qy, &q->y);
  }
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
  for (int w = 1; w < GTAB_NWIN; w++) {
    fe nx, ny;
    if (w + 1 < GTAB_NWIN) {   // prefetch the next window's entry while this addition runs
      const ge *q = gtab + (size_t)(w + 1) * GTAB_ENTRIES + dg[w + 1];
      load_fe_ro(nx, &q->x);
      load_fe_ro(ny, &q->y);
    } else {
      nx = qx; ny = qy;
    }
    gej_add_ge(acc, acc, qx, qy);
    qx = nx; qy = ny;
  }
  if (!fp.spend_adj_inf) gej_add_ge(acc, acc, fp.spend_adj_x, fp.spend_adj_y);
}

// ---------------------------------------------------------------- K3 body: hash, t*G, + B_spend
// in : shared secret S as Jacobian (X, Y) plus Zinv from the batch inversion
// out: P0 = t*G + B_spend as Jacobian.
// t*G uses the offset fixed-window table gtab[w][d] = (d * 2^(WBITS*w) + r_w) * G (affine, 64 B per entry; with
// WBITS = 22: 12 windows, 2.95 GB in HBM): GTAB_NWIN - 1 = 11 mixed additions, no zero digits, no doublings.  The
// constant (sum r_w)*G is removed by adding spend_adj = B_spend - (sum r_w)*G instead of B_spend.
// sx, sy, t_out (optional) receive the affine shared secret and the hash for the intermediates API.
CSP_HD void hash_fixedbase_row(gej &acc, const fe &X, const fe &Y, const fe &Zinv, const ge *gtab, const FinalParams &fp,
                               fe *sx, fe *sy, fe *t_out) {
  fe x, y;
  if (fp.zero_scalar) {
    fe_set_zero(x); fe_set_zero(y);
  } else {
    fe zi2, zi3;
    fe_sqr(zi2, Zinv);
    fe_mul(zi3, zi2, Zinv);
    fe_mul(x, X, zi2);
    fe_mul(y, Y, zi3);
    fe_normalize(x);
    fe_normalize(y);
  }
  if (sx) *sx = x;
  if (sy) *sy = y;
  fe t;
  bip352_shared_secret_hash(t, x, y.v[0] & 1u, fp.midstate);
  if (t_out) *t_out = t;
  fixedbase_plus_spend(acc, t, gtab, fp);
}

// ---------------------------------------------------------------- K5/K7 bodies
CSP_HD int64_t upper64(const fe &x_canonical) {   // src/cudasp_gpu.cu:67-71
  return (int64_t)(((uint64_t)x_canonical.v[7] << 32) | x_canonical.v[6]);
}
// base candidate: x0 = X * Zinv^2 (canonical), optionally y0
CSP_HD int64_t base_candidate(fe &x0, fe *y0, const fe &X, const fe &Y, const fe &Zinv) {
  fe zi2;
  fe_sqr(zi2, Zinv);
  fe_mul(x0, X, zi2);
  fe_normalize(x0);
  if (y0) {
    fe zi3;
    fe_mul(zi3, zi2, Zinv);
    fe_mul(*y0, Y, zi3);
    fe_normalize(*y0);
  }
  return upper64(x0);
}
// label candidate: int64(x(P0 + L) >> 192) with the affine chord rule and the inverted
// denominator dinv = (xL - x0)^-1 from the shared inversion   (src/cudasp_gpu.cu:153-172).
// dinv == 0 marks xL == x0: L = P0 -> tangent rule; L = -P0 -> infinity, whose affine form in the
// reference is (1, 0), i.e. candidate 0 (gECC ec.h:84-89,584-588).  Rare: private inversion.
CSP_HD int64_t label_candidate(const fe &x0, const fe &y0, const fe &lx, const fe &ly, const fe &dinv) {
  fe dy, lam, x;
  fe_sub(dy, ly, y0);
  if (fe_is_zero(dinv)) {
    if (fe_is_zero(dy) && !fe_is_zero(y0)) {
      fe num, den;
      fe_sqr(num, x0);
      fe_lin1<3>(num, num);
      fe_lin1<2>(den, y0);
      fe_inv(den, den);
      fe_mul(lam, num, den);
      fe_sqr(x, lam);
      fe_lin2<1, -2>(x, x, x0);
      fe_normalize(x);
      return upper64(x);
    }
    return 0;
  }
  fe_mul(lam, dy, dinv);
  fe_sqr(x, lam);
  fe_lin3<1, -1, -1>(x, x, x0, lx);
  fe_normalize(x);
  return upper64(x);
}

// ---------------------------------------------------------------- match details (opt-in, SURVEY 8f rank 3+4)
// Full BIP-352 receive loop for ONE row that already matched at k = 0 (BIP-352 "Scanning", with the reference's data
// model: outputs are 64-bit x prefixes, the label set is the query's label_keys):
//   k = 0;  loop:  t_k = hash(ser(S) || ser32(k));  stop if t_k == 0 or t_k >= n
//                  P_k = B_spend + t_k*G;  candidates c_0 = pre(x(P_k)), c_i = pre(x(P_k + L_i))
//                  first remaining output (list order) equal to c_0, else to the first matching c_i, is recorded
//                  as (row, k, output index, label or -1, t_k), removed from the set, k++;  none found -> stop
// The default path (k = 0 only, t not range-checked, src/cudasp_gpu.cu:278-282,869-876) is untouched by this pass:
// it only ADDS information about rows the default path reports.
struct MatchDetail {           // == cudasp_match_detail (include/cudasp.h)
  uint32_t row, k, output_index;
  int32_t label_index;
  uint8_t t_k[32];
};
static const uint32_t DETAILS_MAX_LABELS = 64;   // == CUDASP_MAX_LABELS
static const uint32_t DETAILS_MAX_K = 32;      // more outputs to one recipient in one transaction than this are not followed

CSP_HD bool tweak_in_range(const fe &t) {      // 0 < t < n  (n = group order)
  const uint32_t n[8] = {0xD0364141u, 0xBFD25E8Cu, 0xAF48A03Bu, 0xBAAEDCE6u, 0xFFFFFFFEu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
  uint32_t any = 0;
  bool lt = false;
  for (int i = 7; i >= 0; i--) {
    any |= t.v[i];
    if (t.v[i] != n[i]) { lt = t.v[i] < n[i]; break; }
  }
  for (int i = 0; i < 8; i++) any |= t.v[i];
  return any != 0 && lt;
}

// emit(k, output_index, label_index, t) is called once per recorded output; returns the number recorded.
template <class Emit>
CSP_HD uint32_t details_row(const fe &sx, uint32_t s_odd, const int64_t *outputs, uint32_t lo, uint32_t hi, const ge *labels,
                            uint32_t n_labels, const ge *gtab, const FinalParams &fp, Emit emit) {
  uint32_t used[DETAILS_MAX_K];
  uint32_t found = 0;
  for (uint32_t k = 0; k < DETAILS_MAX_K && lo + found < hi; k++) {
    fe t;
    bip352_shared_secret_hash_k(t, sx, s_odd, k, fp.midstate);
    if (!tweak_in_range(t)) break;
    gej acc;
    fixedbase_plus_spend(acc, t, gtab, fp);
    fe zi, x0, y0;
    fe_inv(zi, acc.z);
    const int64_t c0 = base_candidate(x0, &y0, acc.x, acc.y, zi);
    // the candidates of this k once (they do not depend on the output): c[0] = P_k, c[1 + i] = P_k + L_i, same rule
    // as K5/K7 (label_candidate) with a private inversion per label
    int64_t c[1 + DETAILS_MAX_LABELS];
    c[0] = c0;
    for (uint32_t i = 0; i < n_labels; i++) {
      fe lx, ly, d;
      load_fe_ro(lx, &labels[i].x);
      load_fe_ro(ly, &labels[i].y);
      fe_sub(d, lx, x0);
      fe_normalize(d);
      if (!fe_is_zero(d)) fe_inv(d, d);
      c[1 + i] = label_candidate(x0, y0, lx, ly, d);
    }
    int32_t hit_label = -2;
    uint32_t hit_out = 0;
    // outputs in list order; for each, P_k itself first, then the labels in order (BIP-352 "Scanning")
    for (uint32_t j = lo; j < hi && hit_label == -2; j++) {
      bool gone = false;
      for (uint32_t u = 0; u < found; u++) gone |= used[u] == j;
      if (gone) continue;
      const int64_t o = outputs[j];
      for (uint32_t i = 0; i <= n_labels; i++) {
        if (o == c[i]) { hit_label = (int32_t)i - 1; hit_out = j; break; }
      }
    }
    if (hit_label == -2) break;
    used[found++] = hit_out;
    emit(k, hit_out - lo, hit_label, t);
  }
  return found;
}

}  // namespace csp
