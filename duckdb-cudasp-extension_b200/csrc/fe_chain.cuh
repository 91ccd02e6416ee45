// fe_chain.cuh -- carry-chain implementations of the GF(p) primitives (included by fe.cuh).
//
// Design rule (measured on B200, profiles/r01a): ptxas interleaves every independent carry chain it can
// find; carries live in predicate registers (7 per thread), so more than ~6 chains in flight turn into
// P2R / ISETP / LOP3 predicate spill code -- 40 % of the dynamic instructions of the first ladder kernel.
// Every routine here is therefore ONE serial dependency chain with at most two carries alive:
//   * the 8x8 product runs row by row; row i cannot start before row i-1 has finished because its
//     multiplier word is or-ed with (last word of row i-1 & kzero), kzero being a __constant__ the
//     compiler cannot see through (always 0);
//   * additions, subtractions and small linear combinations are single add.cc / sub.cc chains with a
//     9th "top" word, folded by fold_top();
//   * the 512 -> 256 fold is two short chains.
// Latency is hidden by the other resident warps and by independent field operations of the same thread,
// not by interleaving the rows of one product.
//
// The same source runs on the host (tests/hostcheck, -DCSP_FE_EMUL) with the carry flag emulated in a
// variable, so the arithmetic is unit-tested on the CPU before it is run on the GPU.
#pragma once

namespace csp {
namespace cc {

#if defined(__CUDA_ARCH__)
// ---- device: PTX condition-code chains.  `CF` is a dummy; the carry lives in the PTX CC register.
struct CF {};
CSP_D uint32_t kzero() { return kzero_dev; }
CSP_D uint32_t add_cc(CF &, uint32_t a, uint32_t b) { uint32_t r; asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
CSP_D uint32_t addc_cc(CF &, uint32_t a, uint32_t b) { uint32_t r; asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
CSP_D uint32_t addc(CF &, uint32_t a, uint32_t b) { uint32_t r; asm volatile("addc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
CSP_D uint32_t sub_cc(CF &, uint32_t a, uint32_t b) { uint32_t r; asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
CSP_D uint32_t subc_cc(CF &, uint32_t a, uint32_t b) { uint32_t r; asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
CSP_D uint32_t subc(CF &, uint32_t a, uint32_t b) { uint32_t r; asm volatile("subc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
// (hi:lo) += a*b as a pair of mad.lo / madc.hi: ptxas fuses each pair into one IMAD.WIDE.U32
CSP_D void mad_wide_cc(CF &, uint32_t &lo, uint32_t &hi, uint32_t a, uint32_t b) {      // starts a chain, carry out
  asm volatile("mad.lo.cc.u32 %0, %2, %3, %0;\n\tmadc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(lo), "+r"(hi) : "r"(a), "r"(b));
}
CSP_D void madc_wide_cc(CF &, uint32_t &lo, uint32_t &hi, uint32_t a, uint32_t b) {     // carry in, carry out
  asm volatile("madc.lo.cc.u32 %0, %2, %3, %0;\n\tmadc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(lo), "+r"(hi) : "r"(a), "r"(b));
}
CSP_D void madc_wide(CF &, uint32_t &lo, uint32_t &hi, uint32_t a, uint32_t b) {        // carry in, ends the chain
  asm volatile("madc.lo.cc.u32 %0, %2, %3, %0;\n\tmadc.hi.u32 %1, %2, %3, %1;" : "+r"(lo), "+r"(hi) : "r"(a), "r"(b));
}
CSP_D void mad_wide(CF &, uint32_t &lo, uint32_t &hi, uint32_t a, uint32_t b) {         // no carry in, none out
  asm volatile("mad.lo.cc.u32 %0, %2, %3, %0;\n\tmadc.hi.u32 %1, %2, %3, %1;" : "+r"(lo), "+r"(hi) : "r"(a), "r"(b));
}
#else
// ---- host emulation of the same primitives (carry flag in CF.c)
struct CF { uint32_t c = 0; };
static volatile uint32_t kzero_host = 0;
inline uint32_t kzero() { return kzero_host; }
inline uint32_t add_cc(CF &f, uint32_t a, uint32_t b) { uint64_t s = (uint64_t)a + b; f.c = (uint32_t)(s >> 32); return (uint32_t)s; }
inline uint32_t addc_cc(CF &f, uint32_t a, uint32_t b) { uint64_t s = (uint64_t)a + b + f.c; f.c = (uint32_t)(s >> 32); return (uint32_t)s; }
inline uint32_t addc(CF &f, uint32_t a, uint32_t b) { return a + b + f.c; }
// PTX semantics: after sub.cc / subc.cc the flag holds the BORROW (1 = borrow); subc computes a - (b + CF)
inline uint32_t sub_cc(CF &f, uint32_t a, uint32_t b) { f.c = a < b ? 1u : 0u; return a - b; }
inline uint32_t subc_cc(CF &f, uint32_t a, uint32_t b) { uint64_t s = (uint64_t)a - b - f.c; f.c = (uint32_t)(s >> 63); return (uint32_t)s; }
inline uint32_t subc(CF &f, uint32_t a, uint32_t b) { return a - b - f.c; }
inline void mad_wide_cc(CF &f, uint32_t &lo, uint32_t &hi, uint32_t a, uint32_t b) {
  uint64_t p = (uint64_t)a * b, acc = ((uint64_t)hi << 32) | lo, s = acc + p;
  f.c = s < p ? 1u : 0u; lo = (uint32_t)s; hi = (uint32_t)(s >> 32);
}
inline void madc_wide_cc(CF &f, uint32_t &lo, uint32_t &hi, uint32_t a, uint32_t b) {
  uint64_t p = (uint64_t)a * b, acc = ((uint64_t)hi << 32) | lo;
  unsigned __int128 s = (unsigned __int128)acc + p + f.c;
  f.c = (uint32_t)(s >> 64); lo = (uint32_t)s; hi = (uint32_t)((uint64_t)s >> 32);
}
inline void madc_wide(CF &f, uint32_t &lo, uint32_t &hi, uint32_t a, uint32_t b) { madc_wide_cc(f, lo, hi, a, b); f.c = 0; }
inline void mad_wide(CF &f, uint32_t &lo, uint32_t &hi, uint32_t a, uint32_t b) { mad_wide_cc(f, lo, hi, a, b); f.c = 0; }
#endif

}  // namespace cc

// ------------------------------------------------------------------ small helpers
CSP_HD uint32_t funnel_l(uint32_t lo, uint32_t hi, int s) {   // high word of ((hi:lo) << s), 0 < s < 32
#if defined(__CUDA_ARCH__)
  return __funnelshift_l(lo, hi, s);
#else
  return (hi << s) | (lo >> (32 - s));
#endif
}
// b | (x & kzero): the value of b, made data-dependent on x (serialises two otherwise independent chains)
CSP_HD uint32_t after(uint32_t b, uint32_t x) { return b | (x & cc::kzero()); }

// r (8 words, any value) + top * 2^256  ->  congruent value in [0, 2^256).  top is a small signed number
// (|top| < 2^20).  2^256 = 2^32 + 977 (mod p).  One add chain + a 3-word correction for the rare second wrap.
CSP_HD void fold_top(uint32_t *r, int32_t top) {
  cc::CF f;
  const int32_t k = top * 977;
  const uint32_t sx = (uint32_t)(top >> 31);          // 0 or 0xffffffff: sign extension of top*(2^32+977)
  const uint32_t khi = (uint32_t)top + sx;            // second word of the 64-bit two's complement value
  r[0] = cc::add_cc(f, r[0], (uint32_t)k);
  r[1] = cc::addc_cc(f, r[1], khi);
#pragma unroll
  for (int i = 2; i < 8; i++) r[i] = cc::addc_cc(f, r[i], sx);
  const int32_t d = (int32_t)cc::addc(f, sx, 0);      // -1, 0 or +1: the result left [0, 2^256)
  // second wrap (|d| = 1 only when the value is within 2^40 of a multiple of 2^256): touches three words
#if defined(CSP_RARE_BRANCH) && defined(__CUDA_ARCH__)
  if (d == 0) return;                                 // all but ~2^-215 of the cases: two instructions instead of seven
#endif
  const int32_t k2 = d * 977;
  const uint32_t sx2 = (uint32_t)(d >> 31);
  r[0] = cc::add_cc(f, r[0], (uint32_t)k2);
  r[1] = cc::addc_cc(f, r[1], (uint32_t)d + sx2);
  r[2] = cc::addc(f, r[2], sx2);
}

// word i of (a << S) and the bits shifted out at the top
template <int S>
CSP_HD uint32_t shl_word(const fe &a, int i) {
  if (S == 0) return a.v[i];
  return i ? funnel_l(a.v[i - 1], a.v[i], S ? S : 1) : (a.v[0] << S);
}
template <int S>
CSP_HD uint32_t shl_top(const fe &a) { return S ? (a.v[7] >> ((32 - S) & 31)) : 0u; }

// acc (8 words + signed top) += / -= (a << S)
template <int S, bool NEG>
CSP_HD void lin_acc(uint32_t *r, int32_t &top, const fe &a) {
  cc::CF f;
  if (!NEG) {
    r[0] = cc::add_cc(f, r[0], shl_word<S>(a, 0));
#pragma unroll
    for (int i = 1; i < 8; i++) r[i] = cc::addc_cc(f, r[i], shl_word<S>(a, i));
    top = (int32_t)cc::addc(f, (uint32_t)top, shl_top<S>(a));
  } else {
    r[0] = cc::sub_cc(f, r[0], shl_word<S>(a, 0));
#pragma unroll
    for (int i = 1; i < 8; i++) r[i] = cc::subc_cc(f, r[i], shl_word<S>(a, i));
    top = (int32_t)cc::subc(f, (uint32_t)top, shl_top<S>(a));
  }
}
template <int S>
CSP_HD void lin_set(uint32_t *r, int32_t &top, const fe &a) {
#pragma unroll
  for (int i = 0; i < 8; i++) r[i] = shl_word<S>(a, i);
  top = (int32_t)shl_top<S>(a);
}
CSP_CE int lin_log2(int k) { return k == 1 ? 0 : k == 2 ? 1 : k == 4 ? 2 : k == 8 ? 3 : -1; }
CSP_CE int lin_abs(int k) { return k < 0 ? -k : k; }
CSP_CE bool lin_ok_first(int k) { return k == 1 || k == 2 || k == 3 || k == 4 || k == 8; }
CSP_CE bool lin_ok(int k) { return lin_log2(lin_abs(k)) >= 0; }
template <int K>
CSP_HD void lin_first(uint32_t *r, int32_t &top, const fe &a) {   // K in {1,2,3,4,8}
  if (K == 3) {
    lin_set<1>(r, top, a);
    lin_acc<0, false>(r, top, a);
  } else {
    lin_set<(lin_log2(K) < 0 ? 0 : lin_log2(K))>(r, top, a);
  }
}
template <int K>
CSP_HD void lin_next(uint32_t *r, int32_t &top, const fe &a) {    // K in {+-1, +-2, +-4, +-8}
  lin_acc<(lin_log2(lin_abs(K)) < 0 ? 0 : lin_log2(lin_abs(K))), (K < 0)>(r, top, a);
}

// ------------------------------------------------------------------ add / sub
CSP_HD void chain_add(fe &r, const fe &a, const fe &b) {
  uint32_t w[8];
  int32_t top;
  lin_set<0>(w, top, a);
  lin_acc<0, false>(w, top, b);
  fold_top(w, top);
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = w[i];
}
CSP_HD void chain_sub(fe &r, const fe &a, const fe &b) {
  uint32_t w[8];
  int32_t top;
  lin_set<0>(w, top, a);
  lin_acc<0, true>(w, top, b);
  fold_top(w, top);
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = w[i];
}

// ------------------------------------------------------------------ 256 x 256 -> 512
// Even/odd split: products whose word position i+j is even accumulate in e[], odd ones in o[] (o[k] is
// word k+1), so that every 64-bit product lands on an aligned register pair and each row is two carry
// chains of four IMAD.WIDE.  Rows are serialised with after(): two carries alive at any time.
CSP_HD void chain_row4(uint32_t *acc, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b, bool carry_out) {
  cc::CF f;
  cc::mad_wide_cc(f, acc[0], acc[1], a0, b);
  cc::madc_wide_cc(f, acc[2], acc[3], a1, b);
  cc::madc_wide_cc(f, acc[4], acc[5], a2, b);
  if (carry_out) {
    cc::madc_wide_cc(f, acc[6], acc[7], a3, b);
    acc[8] = cc::addc(f, 0, 0);
  } else {
    cc::madc_wide(f, acc[6], acc[7], a3, b);     // top pair was zero: cannot carry out
  }
}
// t = e + (o << 32)
CSP_HD void chain_merge_eo(uint32_t *t, const uint32_t *e, const uint32_t *o) {
  cc::CF f;
  t[0] = e[0];
  t[1] = cc::add_cc(f, e[1], o[0]);
#pragma unroll
  for (int k = 2; k < 15; k++) t[k] = cc::addc_cc(f, e[k], o[k - 1]);
  t[15] = cc::addc(f, e[15], o[14]);
}
CSP_HD void chain_mul_wide(uint32_t *t, const fe &a, const fe &b) {
  uint32_t e[16], o[16];
#pragma unroll
  for (int k = 0; k < 16; k++) e[k] = o[k] = 0;
  {
    cc::CF f;   // row 0: fresh accumulators, no carries
    cc::mad_wide(f, e[0], e[1], a.v[0], b.v[0]);
    cc::mad_wide(f, e[2], e[3], a.v[2], b.v[0]);
    cc::mad_wide(f, e[4], e[5], a.v[4], b.v[0]);
    cc::mad_wide(f, e[6], e[7], a.v[6], b.v[0]);
    cc::mad_wide(f, o[0], o[1], a.v[1], b.v[0]);
    cc::mad_wide(f, o[2], o[3], a.v[3], b.v[0]);
    cc::mad_wide(f, o[4], o[5], a.v[5], b.v[0]);
    cc::mad_wide(f, o[6], o[7], a.v[7], b.v[0]);
  }
#pragma unroll
  for (int i = 1; i < 8; i++) {
    // e[i + 6] and o[i + 5] hold the words the previous row wrote last
    const uint32_t bi = after(b.v[i], e[i + 6]);
    if (i & 1) {
      // odd row: a_odd * b_i -> even positions i+1 .. i+8 (top pair fresh); a_even * b_i -> o[i-1 .. i+6], carry -> o[i+7]
      chain_row4(e + i + 1, a.v[1], a.v[3], a.v[5], a.v[7], bi, false);
      chain_row4(o + i - 1, a.v[0], a.v[2], a.v[4], a.v[6], bi, true);
    } else {
      // even row: a_even * b_i -> e[i .. i+7], carry -> e[i+8]; a_odd * b_i -> o[i .. i+7] (top pair fresh)
      chain_row4(e + i, a.v[0], a.v[2], a.v[4], a.v[6], bi, true);
      chain_row4(o + i, a.v[1], a.v[3], a.v[5], a.v[7], bi, false);
    }
  }
  chain_merge_eo(t, e, o);
}

// ------------------------------------------------------------------ 256 -> 512 squaring
// upper triangle a_i * a_j (i < j) in even/odd accumulators, then t = 2 * cross + diagonal
template <int N>
CSP_HD void chain_rowN(uint32_t *acc, const uint32_t *a /* stride 2 */, uint32_t b, bool carry_out) {
  cc::CF f;
  if (N == 1 && !carry_out) { cc::mad_wide(f, acc[0], acc[1], a[0], b); return; }
  cc::mad_wide_cc(f, acc[0], acc[1], a[0], b);
#pragma unroll
  for (int k = 1; k < N - 1; k++) cc::madc_wide_cc(f, acc[2 * k], acc[2 * k + 1], a[2 * k], b);
  if (carry_out) {
    if (N > 1) cc::madc_wide_cc(f, acc[2 * N - 2], acc[2 * N - 1], a[2 * N - 2], b);
    acc[2 * N] = cc::addc(f, 0, 0);
  } else {
    cc::madc_wide(f, acc[2 * N - 2], acc[2 * N - 1], a[2 * N - 2], b);
  }
}
CSP_HD void chain_sqr_wide(uint32_t *t, const fe &a) {
  uint32_t e[16], o[16];
#pragma unroll
  for (int k = 0; k < 16; k++) e[k] = o[k] = 0;
  uint32_t m;
  // row 0
  chain_rowN<3>(e + 2, a.v + 2, a.v[0], false);   // j=2,4,6 -> e[2..7]
  chain_rowN<4>(o + 0, a.v + 1, a.v[0], false);   // j=1,3,5,7 -> o[0..7]
  m = after(a.v[1], e[7]);
  chain_rowN<3>(e + 4, a.v + 3, m, false);        // row 1: j=3,5,7 -> e[4..9]   (top fresh)
  chain_rowN<3>(o + 2, a.v + 2, m, true);         //        j=2,4,6 -> o[2..7], carry -> o[8]
  m = after(a.v[2], e[9]);
  chain_rowN<2>(e + 6, a.v + 4, m, true);         // row 2: j=4,6 -> e[6..9], carry -> e[10]
  chain_rowN<3>(o + 4, a.v + 3, m, false);        //        j=3,5,7 -> o[4..9]   (top fresh)
  m = after(a.v[3], e[10]);
  chain_rowN<2>(e + 8, a.v + 5, m, false);        // row 3: j=5,7 -> e[8..11]    (top fresh)
  chain_rowN<2>(o + 6, a.v + 4, m, true);         //        j=4,6 -> o[6..9], carry -> o[10]
  m = after(a.v[4], e[11]);
  chain_rowN<1>(e + 10, a.v + 6, m, true);        // row 4: j=6 -> e[10..11], carry -> e[12]
  chain_rowN<2>(o + 8, a.v + 5, m, false);        //        j=5,7 -> o[8..11]    (top fresh)
  m = after(a.v[5], e[12]);
  chain_rowN<1>(e + 12, a.v + 7, m, false);       // row 5: j=7 -> e[12..13]     (top fresh)
  chain_rowN<1>(o + 10, a.v + 6, m, true);        //        j=6 -> o[10..11], carry -> o[12]
  m = after(a.v[6], o[12]);
  chain_rowN<1>(o + 12, a.v + 7, m, false);       // row 6: j=7 -> o[12..13]     (top fresh)
  uint32_t x[16];
  chain_merge_eo(x, e, o);
  // double
#pragma unroll
  for (int k = 15; k > 0; k--) t[k] = funnel_l(x[k - 1], x[k], 1);
  t[0] = x[0] << 1;
  // + diagonal: one chain of eight
  cc::CF f;
  cc::mad_wide_cc(f, t[0], t[1], a.v[0], a.v[0]);
#pragma unroll
  for (int k = 1; k < 7; k++) cc::madc_wide_cc(f, t[2 * k], t[2 * k + 1], a.v[k], a.v[k]);
  cc::madc_wide(f, t[14], t[15], a.v[7], a.v[7]);
}

// ------------------------------------------------------------------ 512 -> 256 fold
// t = lo + hi * 2^256 = lo + hi * 977 + (hi << 32)  (mod p).
//   E (even-aligned pairs) = lo + 977 * hi_even ; O (odd-aligned pairs, word k+1) = (hi << 32) + 977 * hi_odd
// -- (hi << 32) at odd alignment is hi's own register pairs, so each is 4 IMAD.WIDE + one carry word --
// then u = E + O (9 adds), and the 33-bit top of u is folded once more.
CSP_HD void mad3_cc(cc::CF &f, uint32_t &dlo, uint32_t &dhi, uint32_t a, uint32_t b, uint32_t clo, uint32_t chi, int mode) {
  dlo = clo; dhi = chi;
  if (mode == 0) cc::mad_wide_cc(f, dlo, dhi, a, b);
  else cc::madc_wide_cc(f, dlo, dhi, a, b);
}
CSP_HD void chain_fold512(fe &r, const uint32_t *t) {
  uint32_t E[9], O[10];
  {
    cc::CF f;
    mad3_cc(f, E[0], E[1], t[8], 977u, t[0], t[1], 0);
    mad3_cc(f, E[2], E[3], t[10], 977u, t[2], t[3], 1);
    mad3_cc(f, E[4], E[5], t[12], 977u, t[4], t[5], 1);
    mad3_cc(f, E[6], E[7], t[14], 977u, t[6], t[7], 1);
    E[8] = cc::addc(f, 0, 0);
  }
  {
    cc::CF f;
    mad3_cc(f, O[1], O[2], t[9], 977u, t[8], t[9], 0);
    mad3_cc(f, O[3], O[4], t[11], 977u, t[10], t[11], 1);
    mad3_cc(f, O[5], O[6], t[13], 977u, t[12], t[13], 1);
    mad3_cc(f, O[7], O[8], t[15], 977u, t[14], t[15], 1);
    O[9] = cc::addc(f, 0, 0);
  }
  uint32_t u[8], u8, u9;
  {
    cc::CF f;
    u[0] = E[0];
    u[1] = cc::add_cc(f, E[1], O[1]);
#pragma unroll
    for (int i = 2; i < 8; i++) u[i] = cc::addc_cc(f, E[i], O[i]);
    u8 = cc::addc_cc(f, E[8], O[8]);
    u9 = cc::addc(f, O[9], 0);                      // (u9:u8) < 2^32 + 978
  }
  // second fold: (u9:u8) * (2^32 + 977) added at word 0
  uint32_t p0 = 0, p1 = u9 * 977u;                  // p1 < 2^10; after the product below p1 < 2^11
  {
    cc::CF f;
    cc::mad_wide(f, p0, p1, u8, 977u);              // (p1:p0) = u8 * 977 + (u9 * 977 << 32)
    const uint32_t k1 = cc::add_cc(f, p1, u8);      // word 1 of the addend: p1 + u8, carry into word 2
    const uint32_t k2 = cc::addc(f, u9, 0);
    u[0] = cc::add_cc(f, u[0], p0);
    u[1] = cc::addc_cc(f, u[1], k1);
    u[2] = cc::addc_cc(f, u[2], k2);
#pragma unroll
    for (int i = 3; i < 8; i++) u[i] = cc::addc_cc(f, u[i], 0);
    const uint32_t d = cc::addc(f, 0, 0);           // rare third wrap: the value is then tiny
#if defined(CSP_RARE_BRANCH) && defined(__CUDA_ARCH__)
    if (d != 0)
#endif
    {
      u[0] = cc::add_cc(f, u[0], d * 977u);
      u[1] = cc::addc_cc(f, u[1], d);
      u[2] = cc::addc(f, u[2], 0);
    }
  }
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = u[i];
}

// ------------------------------------------------------------------ public wrappers
CSP_HD void chain_mul(fe &r, const fe &a, const fe &b) {
  uint32_t t[16];
  chain_mul_wide(t, a, b);
  chain_fold512(r, t);
}
CSP_HD void chain_sqr(fe &r, const fe &a) {
  uint32_t t[16];
  chain_sqr_wide(t, a);
  chain_fold512(r, t);
}
template <int KA>
CSP_HD void chain_lin1(fe &r, const fe &a) {
  uint32_t w[8];
  int32_t top;
  lin_first<KA>(w, top, a);
  fold_top(w, top);
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = w[i];
}
template <int KA, int KB>
CSP_HD void chain_lin2(fe &r, const fe &a, const fe &b) {
  uint32_t w[8];
  int32_t top;
  lin_first<KA>(w, top, a);
  lin_next<KB>(w, top, b);
  fold_top(w, top);
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = w[i];
}
template <int KA, int KB, int KC>
CSP_HD void chain_lin3(fe &r, const fe &a, const fe &b, const fe &c) {
  uint32_t w[8];
  int32_t top;
  lin_first<KA>(w, top, a);
  lin_next<KB>(w, top, b);
  lin_next<KC>(w, top, c);
  fold_top(w, top);
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = w[i];
}

}  // namespace csp
