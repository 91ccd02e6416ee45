// scalar_host.hpp -- once-per-query host precompute for the fixed scan key.
//
// The reference replicates the single scan scalar per row and walks its 256 bits on the GPU
// for every row (src/cudasp_gpu.cu:719-732; gECC batch_ec.h:112-275).  Here the scalar is
// analysed ONCE per query on the host:
//   k  = scan_sk mod n                           (same point for order-n inputs)
//   k  = k1 + k2*lambda (mod n), |k1|,|k2| < 2^128   (GLV endomorphism lambda*(x,y) = (beta*x, y))
//   k1, k2 -> width-4 NAF digits, interleaved into one op list (DBL / ADD table[idx] +-),
// and the op list is handed to the ladder kernel as a kernel parameter, so that every thread of
// every warp executes the same branch-free-in-practice schedule.
#pragma once
#include <stdint.h>
#include <string.h>
#include <vector>

namespace csp {
namespace host {

typedef unsigned __int128 u128;

// little-endian 64-bit limbs, fixed capacity, value semantics; tiny and slow on purpose
struct Big {
  static const int W = 10;  // 640 bits
  uint64_t w[W];
  Big() { memset(w, 0, sizeof w); }
  explicit Big(uint64_t x) { memset(w, 0, sizeof w); w[0] = x; }
  static Big from_le(const uint8_t *b, int nbytes) {
    Big r;
    for (int i = 0; i < nbytes; i++) r.w[i / 8] |= (uint64_t)b[i] << (8 * (i % 8));
    return r;
  }
  static Big from_words_be(const uint64_t *be, int n) {  // most significant first
    Big r;
    for (int i = 0; i < n; i++) r.w[n - 1 - i] = be[i];
    return r;
  }
  void to_le(uint8_t *b, int nbytes) const {
    for (int i = 0; i < nbytes; i++) b[i] = (uint8_t)(w[i / 8] >> (8 * (i % 8)));
  }
  bool is_zero() const {
    for (int i = 0; i < W; i++) if (w[i]) return false;
    return true;
  }
  int bit(int i) const { return (int)((w[i / 64] >> (i % 64)) & 1); }
  int bitlen() const {
    for (int i = W * 64 - 1; i >= 0; i--) if (bit(i)) return i + 1;
    return 0;
  }
};
inline int cmp(const Big &a, const Big &b) {
  for (int i = Big::W - 1; i >= 0; i--) {
    if (a.w[i] > b.w[i]) return 1;
    if (a.w[i] < b.w[i]) return -1;
  }
  return 0;
}
inline Big add(const Big &a, const Big &b) {
  Big r; u128 c = 0;
  for (int i = 0; i < Big::W; i++) { c += (u128)a.w[i] + b.w[i]; r.w[i] = (uint64_t)c; c >>= 64; }
  return r;
}
inline Big sub(const Big &a, const Big &b) {  // a >= b
  Big r; u128 br = 0;
  for (int i = 0; i < Big::W; i++) { u128 d = (u128)a.w[i] - b.w[i] - br; r.w[i] = (uint64_t)d; br = (d >> 64) & 1; }
  return r;
}
inline Big mul(const Big &a, const Big &b) {  // truncated to 640 bits (callers stay below)
  Big r;
  for (int i = 0; i < Big::W; i++) {
    u128 c = 0;
    for (int j = 0; i + j < Big::W; j++) {
      c += (u128)a.w[i] * b.w[j] + r.w[i + j];
      r.w[i + j] = (uint64_t)c;
      c >>= 64;
    }
  }
  return r;
}
inline Big shr(const Big &a, int n) {
  Big r;
  int ws = n / 64, bs = n % 64;
  for (int i = 0; i + ws < Big::W; i++) {
    r.w[i] = a.w[i + ws] >> bs;
    if (bs && i + ws + 1 < Big::W) r.w[i] |= a.w[i + ws + 1] << (64 - bs);
  }
  return r;
}
inline Big shl1(const Big &a) {
  Big r; uint64_t c = 0;
  for (int i = 0; i < Big::W; i++) { r.w[i] = (a.w[i] << 1) | c; c = a.w[i] >> 63; }
  return r;
}
inline Big mod(const Big &a, const Big &m) {  // binary long division
  Big r;
  for (int i = a.bitlen() - 1; i >= 0; i--) {
    r = shl1(r);
    r.w[0] |= (uint64_t)a.bit(i);
    if (cmp(r, m) >= 0) r = sub(r, m);
  }
  return r;
}

// secp256k1 group order and GLV constants (checked numerically in tests/test_host_precompute.py)
inline Big N_ORDER() { static const uint64_t v[4] = {0xFFFFFFFFFFFFFFFFULL, 0xFFFFFFFFFFFFFFFEULL, 0xBAAEDCE6AF48A03BULL, 0xBFD25E8CD0364141ULL}; return Big::from_words_be(v, 4); }
inline Big LAMBDA() { static const uint64_t v[4] = {0x5363AD4CC05C30E0ULL, 0xA5261C028812645AULL, 0x122E22EA20816678ULL, 0xDF02967C1B23BD72ULL}; return Big::from_words_be(v, 4); }
inline Big GLV_G1() { static const uint64_t v[4] = {0x3086D221A7D46BCDULL, 0xE86C90E49284EB15ULL, 0x3DAA8A1471E8CA7FULL, 0xE893209A45DBB031ULL}; return Big::from_words_be(v, 4); }
inline Big GLV_G2() { static const uint64_t v[4] = {0xE4437ED6010E8828ULL, 0x6F547FA90ABFE4C4ULL, 0x221208AC9DF506C6ULL, 0x1571B4AE8AC47F71ULL}; return Big::from_words_be(v, 4); }
inline Big GLV_MINUS_B1() { static const uint64_t v[2] = {0xE4437ED6010E8828ULL, 0x6F547FA90ABFE4C3ULL}; return Big::from_words_be(v, 2); }
inline Big GLV_B2() { static const uint64_t v[2] = {0x3086D221A7D46BCDULL, 0xE86C90E49284EB15ULL}; return Big::from_words_be(v, 2); }
// beta: lambda*(x,y) = (beta*x mod p, y)
static const uint32_t BETA_LIMBS[8] = {0x719501EEu, 0xC1396C28u, 0x12F58995u, 0x9CF04975u, 0xAC3434E9u, 0x6E64479Eu, 0x657C0710u, 0x7AE96A2Bu};

struct SignedScalar {
  Big mag;
  bool neg;
};

// k (already < n) -> k1 + k2*lambda = k (mod n) with both magnitudes < 2^128 (+ a bit)
inline void glv_split(const Big &k, SignedScalar &k1, SignedScalar &k2) {
  const Big n = N_ORDER();
  Big half; half.w[5] = (uint64_t)1 << 63;          // 2^383, for round-to-nearest of >> 384
  Big c1 = shr(add(mul(k, GLV_G1()), half), 384);
  Big c2 = shr(add(mul(k, GLV_G2()), half), 384);
  // k2 = c1*(-b1) + c2*(-b2) = c1*(-b1) - c2*b2   (mod n)
  Big t1 = mod(mul(c1, GLV_MINUS_B1()), n);
  Big t2 = mod(mul(c2, GLV_B2()), n);
  Big r2 = cmp(t1, t2) >= 0 ? sub(t1, t2) : sub(add(t1, n), t2);
  // k1 = k - r2*lambda (mod n)
  Big t3 = mod(mul(r2, LAMBDA()), n);
  Big r1 = cmp(k, t3) >= 0 ? sub(k, t3) : sub(add(k, n), t3);
  Big hn = shr(n, 1);
  k1.neg = cmp(r1, hn) > 0; k1.mag = k1.neg ? sub(n, r1) : r1;
  k2.neg = cmp(r2, hn) > 0; k2.mag = k2.neg ? sub(n, r2) : r2;
}

// width-w NAF, least significant digit first; digits odd in (-2^(w-1), 2^(w-1)) or 0
inline std::vector<int> wnaf(Big k, int w) {
  std::vector<int> d;
  while (!k.is_zero()) {
    int di = 0;
    if (k.w[0] & 1) {
      di = (int)(k.w[0] & ((1u << w) - 1));
      if (di >= (1 << (w - 1))) {
        di -= (1 << w);
        k = add(k, Big((uint64_t)(-di)));
      } else {
        k = sub(k, Big((uint64_t)di));
      }
    }
    d.push_back(di);
    k = shr(k, 1);
  }
  return d;
}

// ---- ladder op list -------------------------------------------------------------------
// one byte per op:  0x80 = DBL;  otherwise ADD:  bits 0-1 table index (|digit| = 2*idx+1),
// bit 2 = lambda table (beta*x), bit 3 = negate y, bit 6 = SET (first op: acc <- table point).
enum { OP_DBL = 0x80, OP_LAMBDA = 0x04, OP_NEG = 0x08, OP_SET = 0x40 };
static const int LADDER_W = 4;        // wNAF width: table = {1,3,5,7} * P
static const int LADDER_MAX_OPS = 240;

struct LadderPlan {
  uint8_t ops[LADDER_MAX_OPS];
  int nops;
  bool zero;       // k = 0 (mod n): shared secret is the reference's all-zero pair
  int n_dbl, n_add;
};

// signed helper for the decomposition search: a + sign*b on (magnitude, sign) pairs
inline SignedScalar signed_add(const SignedScalar &a, const Big &bmag, bool bneg) {
  SignedScalar r;
  if (a.neg == bneg) { r.mag = add(a.mag, bmag); r.neg = a.neg; return r; }
  if (cmp(a.mag, bmag) >= 0) { r.mag = sub(a.mag, bmag); r.neg = a.neg; }
  else { r.mag = sub(bmag, a.mag); r.neg = bneg; }
  if (r.mag.is_zero()) r.neg = false;
  return r;
}
// lattice of pairs (a, b) with a + b*lambda = 0 (mod n): v1 = (A1, -MINUS_B1), v2 = (A2, B2), B2 = A1
inline Big GLV_A2() { static const uint64_t v[3] = {0x1ULL, 0x14CA50F7A8E2F3F6ULL, 0x57C1108D9D44CFD8ULL}; return Big::from_words_be(v, 3); }

// op list for k1 + k2*lambda with the halves given as signed magnitudes (interleaved width-4 NAFs, one doubling chain)
inline LadderPlan ladder_plan_from_halves(const SignedScalar &k1, const SignedScalar &k2) {
  LadderPlan p;
  memset(&p, 0, sizeof p);
  std::vector<int> d1 = wnaf(k1.mag, LADDER_W), d2 = wnaf(k2.mag, LADDER_W);
  int len = (int)(d1.size() > d2.size() ? d1.size() : d2.size());
  bool started = false;
  for (int i = len - 1; i >= 0; i--) {
    if (started) {
      if (p.nops >= LADDER_MAX_OPS) { p.nops = LADDER_MAX_OPS + 1; return p; }
      p.ops[p.nops++] = OP_DBL; p.n_dbl++;
    }
    for (int which = 0; which < 2; which++) {
      const std::vector<int> &d = which ? d2 : d1;
      int di = i < (int)d.size() ? d[i] : 0;
      if (!di) continue;
      bool neg = (di < 0) != (which ? k2.neg : k1.neg);
      int idx = ((di < 0 ? -di : di) - 1) / 2;
      uint8_t op = (uint8_t)(idx | (which ? OP_LAMBDA : 0) | (neg ? OP_NEG : 0));
      if (!started) { op |= OP_SET; started = true; } else p.n_add++;
      if (p.nops >= LADDER_MAX_OPS) { p.nops = LADDER_MAX_OPS + 1; return p; }
      p.ops[p.nops++] = op;
    }
  }
  return p;
}
// field multiplications K1 spends on a plan (3M + 4S per doubling, 8M + 3S per mixed addition)
inline int ladder_plan_cost(const LadderPlan &p) { return 7 * p.n_dbl + 11 * p.n_add; }

// variant (i, j), small integers: the balanced split moved by i*v1 + j*v2 -- the same scalar mod n, other digits.
// (0, 0) is the balanced split itself.
inline LadderPlan make_ladder_plan_variant(const uint8_t scan_sk_le[32], int i, int j) {
  LadderPlan p;
  memset(&p, 0, sizeof p);
  const Big n = N_ORDER();
  Big k = mod(Big::from_le(scan_sk_le, 32), n);
  if (k.is_zero()) { p.zero = true; return p; }
  SignedScalar k1, k2;
  glv_split(k, k1, k2);
  for (int t = 0; t < (i < 0 ? -i : i); t++) { k1 = signed_add(k1, GLV_B2(), i < 0); k2 = signed_add(k2, GLV_MINUS_B1(), i > 0); }   // v1 = (A1, b1), A1 = B2, b1 < 0
  for (int t = 0; t < (j < 0 ? -j : j); t++) { k1 = signed_add(k1, GLV_A2(), j < 0); k2 = signed_add(k2, GLV_B2(), j < 0); }          // v2 = (A2, B2)
  return ladder_plan_from_halves(k1, k2);
}
inline LadderPlan make_ladder_plan(const uint8_t scan_sk_le[32]) { return make_ladder_plan_variant(scan_sk_le, 0, 0); }

}  // namespace host
}  // namespace csp
