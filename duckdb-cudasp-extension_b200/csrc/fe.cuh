// fe.cuh -- arithmetic in GF(p), p = 2^256 - 2^32 - 977 (secp256k1 base field).
//
// Replaces the reference's Montgomery-form FpT (gECC/include/gecc/arith/fp.h:111-262,
// details/fp_mont_multiply.h:194-247, details/fp_mont_add_sub.h:6-51,
// details/fp_mont_inverse.h:74-183).  Differences by design:
//   * plain (non-Montgomery) residues: the special prime makes 2^256 = 2^32 + 977 (mod p),
//     so a 512-bit product folds with 8 small multiplies instead of a second 8x8 Montgomery pass;
//   * values are kept WEAKLY reduced in [0, 2^256) and only canonicalised (fe_normalize)
//     where bits are observed (parity, serialisation, x-prefix compare, zero test);
//   * inversion is a fixed Fermat addition chain (255 S + 15 M), constant control flow,
//     instead of the reference's data-dependent binary GCD;
//   * a dedicated squaring.
// 8 x 32-bit little-endian limbs.  Every function is __host__ __device__: the portable
// 64-bit-accumulator code path compiles to IMAD.WIDE.U32 on sm_100a and is also what the
// CPU-side unit tests run; the PTX carry-chain path (mad.lo.cc / madc.hi.cc, which ptxas
// fuses to IMAD.WIDE.U32 with predicate carries) is selected on the device.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define CSP_HD __host__ __device__ __forceinline__
#define CSP_D __device__ __forceinline__
#define CSP_CE __host__ __device__ constexpr
#else
#define CSP_HD inline
#define CSP_CE constexpr
#endif

#if defined(__CUDA_ARCH__) && !defined(CSP_FE_PORTABLE)
#define CSP_FE_PTX 1
#else
#define CSP_FE_PTX 0
#endif

// Carry-chain implementations (fe_chain.cuh): PTX on the device, emulated carry flag on the host when
// CSP_FE_EMUL is defined (tests/hostcheck builds both the portable and the emulated variant).
#if (defined(__CUDA_ARCH__) && !defined(CSP_FE_PORTABLE)) || defined(CSP_FE_EMUL)
#define CSP_CHAIN 1
#else
#define CSP_CHAIN 0
#endif

namespace csp {

struct fe {
  uint32_t v[8];
};

#if defined(__CUDACC__)
namespace cc {
__constant__ uint32_t kzero_dev;   // never written: 0.  Opaque to the compiler (see fe_chain.cuh, after()).
}
#endif

}  // namespace csp
#if CSP_CHAIN
#include "fe_chain.cuh"
#endif
namespace csp {

// 2^256 mod p = 2^32 + 977
#define CSP_PC_LO 977u

CSP_HD void fe_set_zero(fe &r) {
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = 0;
}
CSP_HD void fe_set_u32(fe &r, uint32_t x) {
  fe_set_zero(r);
  r.v[0] = x;
}

// r += k * (2^32 + 977) for a small k (k < 2^31), given that the caller knows the result
// cannot exceed 2^256 more than once; returns the carry out of limb 7.
CSP_HD uint32_t fe_add_kc(fe &r, uint32_t k) {
  uint64_t c = (uint64_t)r.v[0] + (uint64_t)k * CSP_PC_LO;
  r.v[0] = (uint32_t)c;
  c >>= 32;
  c += (uint64_t)r.v[1] + k;
  r.v[1] = (uint32_t)c;
  c >>= 32;
#pragma unroll
  for (int i = 2; i < 8; i++) {
    c += r.v[i];
    r.v[i] = (uint32_t)c;
    c >>= 32;
  }
  return (uint32_t)c;
}

// r = a + b (mod p), weakly reduced
CSP_HD void fe_add(fe &r, const fe &a, const fe &b) {
#if CSP_CHAIN
  chain_add(r, a, b);
  return;
#endif
  uint64_t c = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += (uint64_t)a.v[i] + b.v[i];
    r.v[i] = (uint32_t)c;
    c >>= 32;
  }
  // a + b = r + c*2^256 = r + c*(2^32+977) (mod p)
  uint32_t c2 = fe_add_kc(r, (uint32_t)c);
  if (c2) {  // only when r was within 2^33 of 2^256: the wrapped value is tiny, no third wrap
    uint64_t d = (uint64_t)r.v[0] + CSP_PC_LO;
    r.v[0] = (uint32_t)d;
    d = (d >> 32) + r.v[1] + 1;
    r.v[1] = (uint32_t)d;
    r.v[2] += (uint32_t)(d >> 32);
  }
}

// r = a - b (mod p), weakly reduced
CSP_HD void fe_sub(fe &r, const fe &a, const fe &b) {
#if CSP_CHAIN
  chain_sub(r, a, b);
  return;
#endif
  int64_t c = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += (int64_t)a.v[i] - (int64_t)b.v[i];
    r.v[i] = (uint32_t)c;
    c >>= 32;  // arithmetic: 0 or -1
  }
  // a - b = r - borrow*2^256 = r - borrow*(2^32+977) (mod p)
  uint32_t borrow = (uint32_t)(-c);  // 0 or 1
  int64_t d = (int64_t)r.v[0] - (int64_t)(borrow * CSP_PC_LO);
  r.v[0] = (uint32_t)d;
  d >>= 32;
  d += (int64_t)r.v[1] - (int64_t)borrow;
  r.v[1] = (uint32_t)d;
  d >>= 32;
#pragma unroll
  for (int i = 2; i < 8; i++) {
    d += (int64_t)r.v[i];
    r.v[i] = (uint32_t)d;
    d >>= 32;
  }
  if (d) {  // second borrow: r was < 2^33 before the fold, now r = 2^256 - small: subtract once more
    int64_t e = (int64_t)r.v[0] - (int64_t)CSP_PC_LO;
    r.v[0] = (uint32_t)e;
    e >>= 32;
    e += (int64_t)r.v[1] - 1;
    r.v[1] = (uint32_t)e;
    e >>= 32;
#pragma unroll
    for (int i = 2; i < 8; i++) {
      e += (int64_t)r.v[i];
      r.v[i] = (uint32_t)e;
      e >>= 32;
    }
  }
}

CSP_HD void fe_dbl(fe &r, const fe &a) {
#if CSP_CHAIN
  chain_lin1<2>(r, a);
#else
  fe_add(r, a, a);
#endif
}

// r = -a (mod p): 2p - a would overflow 2^256; use (p - a) when a <= p, else (2p - a) - 2^256 ... keep simple:
CSP_HD void fe_neg(fe &r, const fe &a) {
  fe z;
  fe_set_zero(z);
  fe_sub(r, z, a);
}

// r = a * k for small k (k <= 2^20)
CSP_HD void fe_mul_small(fe &r, const fe &a, uint32_t k) {
  uint64_t c = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += (uint64_t)a.v[i] * k;
    r.v[i] = (uint32_t)c;
    c >>= 32;
  }
  uint32_t c2 = fe_add_kc(r, (uint32_t)c);
  if (c2) {
    uint64_t d = (uint64_t)r.v[0] + CSP_PC_LO;
    r.v[0] = (uint32_t)d;
    d = (d >> 32) + r.v[1] + 1;
    r.v[1] = (uint32_t)d;
    r.v[2] += (uint32_t)(d >> 32);
  }
}

// ------------------------------------------------------------------ 512 -> 256 fold
// t[0..15] (little-endian words of a 512-bit value)  ->  r = t mod p, weakly reduced.
// lo + hi * (2^32 + 977): one pass gives 8 words + a top of < 2^34, folded a second time.
CSP_HD void fe_fold512(fe &r, const uint32_t *t) {
  uint64_t c = 0;
  uint32_t prev = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    c += (uint64_t)t[8 + j] * CSP_PC_LO + t[j] + prev;
    prev = t[8 + j];
    r.v[j] = (uint32_t)c;
    c >>= 32;
  }
  uint64_t top = c + prev;  // < 2^34
  // second fold: top * (2^32 + 977)
  uint64_t k = (top & 0xffffffffu) * CSP_PC_LO + ((top >> 32) * CSP_PC_LO << 32);  // top*977 < 2^44
  uint64_t d = (uint64_t)r.v[0] + (uint32_t)k;
  r.v[0] = (uint32_t)d;
  d >>= 32;
  d += (uint64_t)r.v[1] + (uint32_t)(k >> 32) + (uint32_t)top;
  r.v[1] = (uint32_t)d;
  d >>= 32;
  d += (uint64_t)r.v[2] + (uint32_t)(top >> 32);
  r.v[2] = (uint32_t)d;
  d >>= 32;
#pragma unroll
  for (int i = 3; i < 8; i++) {
    d += r.v[i];
    r.v[i] = (uint32_t)d;
    d >>= 32;
  }
  if (d) {  // wrapped 2^256 once more (value now tiny): add 2^32 + 977
    uint64_t e = (uint64_t)r.v[0] + CSP_PC_LO;
    r.v[0] = (uint32_t)e;
    e = (e >> 32) + r.v[1] + 1;
    r.v[1] = (uint32_t)e;
    e = (e >> 32) + r.v[2];
    r.v[2] = (uint32_t)e;
    r.v[3] += (uint32_t)(e >> 32);
  }
}

#if CSP_FE_PTX
// ---- device carry chains.  Four 32x32->64 products at word offsets 0,2,4,6 of acc[0..7]
// added in one carry chain; ptxas fuses each mad.lo.cc/madc.hi.cc pair into one
// IMAD.WIDE.U32 with predicate carry-in/out.  *_c also captures the carry into acc[8].
CSP_D void chain4(uint32_t *acc, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b) {
  asm("mad.lo.cc.u32 %0, %8, %12, %0;\n\tmadc.hi.cc.u32 %1, %8, %12, %1;\n\t"
      "madc.lo.cc.u32 %2, %9, %12, %2;\n\tmadc.hi.cc.u32 %3, %9, %12, %3;\n\t"
      "madc.lo.cc.u32 %4, %10, %12, %4;\n\tmadc.hi.cc.u32 %5, %10, %12, %5;\n\t"
      "madc.lo.cc.u32 %6, %11, %12, %6;\n\tmadc.hi.u32 %7, %11, %12, %7;"
      : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "+r"(acc[4]), "+r"(acc[5]), "+r"(acc[6]), "+r"(acc[7])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
}
CSP_D void chain4_c(uint32_t *acc, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b) {
  asm("mad.lo.cc.u32 %0, %9, %13, %0;\n\tmadc.hi.cc.u32 %1, %9, %13, %1;\n\t"
      "madc.lo.cc.u32 %2, %10, %13, %2;\n\tmadc.hi.cc.u32 %3, %10, %13, %3;\n\t"
      "madc.lo.cc.u32 %4, %11, %13, %4;\n\tmadc.hi.cc.u32 %5, %11, %13, %5;\n\t"
      "madc.lo.cc.u32 %6, %12, %13, %6;\n\tmadc.hi.cc.u32 %7, %12, %13, %7;\n\t"
      "addc.u32 %8, 0, 0;"
      : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "+r"(acc[4]), "+r"(acc[5]), "+r"(acc[6]), "+r"(acc[7]),
        "=r"(acc[8])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
}
CSP_D void mulwide(uint32_t &lo, uint32_t &hi, uint32_t a, uint32_t b) {
  uint64_t p = (uint64_t)a * b;
  lo = (uint32_t)p;
  hi = (uint32_t)(p >> 32);
}
// t = e + (o << 32): word k of t is e[k] + o[k-1]
CSP_D void merge_eo(uint32_t *t, const uint32_t *e, const uint32_t *o) {
  t[0] = e[0];
  asm("add.cc.u32 %0, %15, %30;\n\taddc.cc.u32 %1, %16, %31;\n\taddc.cc.u32 %2, %17, %32;\n\t"
      "addc.cc.u32 %3, %18, %33;\n\taddc.cc.u32 %4, %19, %34;\n\taddc.cc.u32 %5, %20, %35;\n\t"
      "addc.cc.u32 %6, %21, %36;\n\taddc.cc.u32 %7, %22, %37;\n\taddc.cc.u32 %8, %23, %38;\n\t"
      "addc.cc.u32 %9, %24, %39;\n\taddc.cc.u32 %10, %25, %40;\n\taddc.cc.u32 %11, %26, %41;\n\t"
      "addc.cc.u32 %12, %27, %42;\n\taddc.cc.u32 %13, %28, %43;\n\taddc.u32 %14, %29, %44;"
      : "=r"(t[1]), "=r"(t[2]), "=r"(t[3]), "=r"(t[4]), "=r"(t[5]), "=r"(t[6]), "=r"(t[7]), "=r"(t[8]), "=r"(t[9]),
        "=r"(t[10]), "=r"(t[11]), "=r"(t[12]), "=r"(t[13]), "=r"(t[14]), "=r"(t[15])
      : "r"(e[1]), "r"(e[2]), "r"(e[3]), "r"(e[4]), "r"(e[5]), "r"(e[6]), "r"(e[7]), "r"(e[8]), "r"(e[9]), "r"(e[10]),
        "r"(e[11]), "r"(e[12]), "r"(e[13]), "r"(e[14]), "r"(e[15]),
        "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]), "r"(o[4]), "r"(o[5]), "r"(o[6]), "r"(o[7]), "r"(o[8]), "r"(o[9]),
        "r"(o[10]), "r"(o[11]), "r"(o[12]), "r"(o[13]), "r"(o[14]));
}
#endif

// ------------------------------------------------------------------ 256 x 256 -> 512
CSP_HD void fe_mul_wide(uint32_t *t, const fe &a, const fe &b) {
#if CSP_FE_PTX
  // Even/odd column split: products whose word position i+j is even accumulate in e[],
  // odd ones in o[] (o[k] sits at word k+1), so every product lands on an aligned register
  // pair and each row is two independent 4-product carry chains.
  uint32_t e[16], o[16];
  mulwide(e[0], e[1], a.v[0], b.v[0]);
  mulwide(e[2], e[3], a.v[2], b.v[0]);
  mulwide(e[4], e[5], a.v[4], b.v[0]);
  mulwide(e[6], e[7], a.v[6], b.v[0]);
  mulwide(o[0], o[1], a.v[1], b.v[0]);
  mulwide(o[2], o[3], a.v[3], b.v[0]);
  mulwide(o[4], o[5], a.v[5], b.v[0]);
  mulwide(o[6], o[7], a.v[7], b.v[0]);
#pragma unroll
  for (int k = 8; k < 16; k++) e[k] = o[k] = 0;
#pragma unroll
  for (int i = 1; i < 8; i++) {
    if (i & 1) {
      // odd row: a_odd * b_i -> even positions i+1..i+8 (top word fresh: no carry out);
      //          a_even * b_i -> odd positions, o[i-1..i+6], carry into o[i+7]
      chain4(e + i + 1, a.v[1], a.v[3], a.v[5], a.v[7], b.v[i]);
      chain4_c(o + i - 1, a.v[0], a.v[2], a.v[4], a.v[6], b.v[i]);
    } else {
      // even row: a_even * b_i -> e[i..i+7], carry into e[i+8]; a_odd * b_i -> o[i..i+7] (top fresh)
      chain4_c(e + i, a.v[0], a.v[2], a.v[4], a.v[6], b.v[i]);
      chain4(o + i, a.v[1], a.v[3], a.v[5], a.v[7], b.v[i]);
    }
  }
  merge_eo(t, e, o);
#else
  uint64_t c = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    c += (uint64_t)a.v[j] * b.v[0];
    t[j] = (uint32_t)c;
    c >>= 32;
  }
  t[8] = (uint32_t)c;
#pragma unroll
  for (int i = 1; i < 8; i++) {
    c = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
      c += (uint64_t)a.v[j] * b.v[i] + t[i + j];
      t[i + j] = (uint32_t)c;
      c >>= 32;
    }
    t[i + 8] = (uint32_t)c;
  }
#endif
}

CSP_HD void fe_mul(fe &r, const fe &a, const fe &b) {
#if CSP_CHAIN
  chain_mul(r, a, b);
  return;
#endif
  uint32_t t[16];
  fe_mul_wide(t, a, b);
  fe_fold512(r, t);
}

#if CSP_FE_PTX
// variable-length carry chains for the squaring's upper triangle (see fe_sqr_wide)
template <int N, bool CARRY>
CSP_D void chainN(uint32_t *acc, const uint32_t *a /* stride 2 */, uint32_t b) {
  asm volatile("mad.lo.cc.u32 %0, %2, %3, %0;\n\tmadc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(acc[0]), "+r"(acc[1]) : "r"(a[0]), "r"(b));
#pragma unroll
  for (int k = 1; k < N; k++)
    asm volatile("madc.lo.cc.u32 %0, %2, %3, %0;\n\tmadc.hi.cc.u32 %1, %2, %3, %1;"
                 : "+r"(acc[2 * k]), "+r"(acc[2 * k + 1])
                 : "r"(a[2 * k]), "r"(b));
  if (CARRY) asm volatile("addc.u32 %0, 0, 0;" : "=r"(acc[2 * N]));
}
#endif

// ------------------------------------------------------------------ 256 -> 512 squaring
CSP_HD void fe_sqr_wide(uint32_t *t, const fe &a) {
#if CSP_FE_PTX
  // upper triangle a_i*a_j (i<j) in even/odd accumulators, then t = 2*cross + diagonal
  uint32_t e[16], o[16];
#pragma unroll
  for (int k = 0; k < 16; k++) e[k] = o[k] = 0;
  // row 0
  chainN<3, false>(e + 2, a.v + 2, a.v[0]);   // j=2,4,6 -> e[2..7]
  chainN<4, false>(o + 0, a.v + 1, a.v[0]);   // j=1,3,5,7 -> o[0..7]
  // row 1
  chainN<3, false>(e + 4, a.v + 3, a.v[1]);   // j=3,5,7 -> e[4..9]   (top fresh)
  chainN<3, true>(o + 2, a.v + 2, a.v[1]);    // j=2,4,6 -> o[2..7], carry -> o[8]
  // row 2
  chainN<2, true>(e + 6, a.v + 4, a.v[2]);    // j=4,6 -> e[6..9], carry -> e[10]
  chainN<3, false>(o + 4, a.v + 3, a.v[2]);   // j=3,5,7 -> o[4..9]   (top fresh)
  // row 3
  chainN<2, false>(e + 8, a.v + 5, a.v[3]);   // j=5,7 -> e[8..11]    (top fresh)
  chainN<2, true>(o + 6, a.v + 4, a.v[3]);    // j=4,6 -> o[6..9], carry -> o[10]
  // row 4
  chainN<1, true>(e + 10, a.v + 6, a.v[4]);   // j=6 -> e[10..11], carry -> e[12]
  chainN<2, false>(o + 8, a.v + 5, a.v[4]);   // j=5,7 -> o[8..11]    (top fresh)
  // row 5
  chainN<1, false>(e + 12, a.v + 7, a.v[5]);  // j=7 -> e[12..13]     (top fresh)
  chainN<1, true>(o + 10, a.v + 6, a.v[5]);   // j=6 -> o[10..11], carry -> o[12]
  // row 6
  chainN<1, false>(o + 12, a.v + 7, a.v[6]);  // j=7 -> o[12..13]     (top fresh)
  uint32_t x[16];
  merge_eo(x, e, o);
  // double
#pragma unroll
  for (int k = 15; k > 0; k--) t[k] = __funnelshift_l(x[k - 1], x[k], 1);
  t[0] = x[0] << 1;
  // + diagonal
  asm volatile("mad.lo.cc.u32 %0, %2, %2, %0;\n\tmadc.hi.cc.u32 %1, %2, %2, %1;" : "+r"(t[0]), "+r"(t[1]) : "r"(a.v[0]));
#pragma unroll
  for (int k = 1; k < 7; k++)
    asm volatile("madc.lo.cc.u32 %0, %2, %2, %0;\n\tmadc.hi.cc.u32 %1, %2, %2, %1;" : "+r"(t[2 * k]), "+r"(t[2 * k + 1]) : "r"(a.v[k]));
  asm volatile("madc.lo.cc.u32 %0, %2, %2, %0;\n\tmadc.hi.u32 %1, %2, %2, %1;" : "+r"(t[14]), "+r"(t[15]) : "r"(a.v[7]));
#else
  fe_mul_wide(t, a, a);
#endif
}

CSP_HD void fe_sqr(fe &r, const fe &a) {
#if CSP_CHAIN
  chain_sqr(r, a);
  return;
#endif
  uint32_t t[16];
  fe_sqr_wide(t, a);
  fe_fold512(r, t);
}

// ------------------------------------------------------------------ canonical form & predicates
// p = FFFFFFFF FFFFFFFF FFFFFFFF FFFFFFFF FFFFFFFF FFFFFFFF FFFFFFFE FFFFFC2F
CSP_HD void fe_normalize(fe &r) {
  // r in [0, 2^256) -> [0, p): r >= p  <=>  r + (2^32 + 977) carries out of 2^256
  uint64_t c = (uint64_t)r.v[0] + CSP_PC_LO;
  uint32_t s[8];
  s[0] = (uint32_t)c;
  c >>= 32;
  c += (uint64_t)r.v[1] + 1;
  s[1] = (uint32_t)c;
  c >>= 32;
#pragma unroll
  for (int i = 2; i < 8; i++) {
    c += r.v[i];
    s[i] = (uint32_t)c;
    c >>= 32;
  }
  if (c) {
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = s[i];
  }
}
CSP_HD bool fe_is_zero_normalized(const fe &a) {
  uint32_t z = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) z |= a.v[i];
  return z == 0;
}
// a == 0 (mod p) for a weakly reduced a: a is 0 or p
CSP_HD bool fe_is_zero(const fe &a) {
  fe t = a;
  fe_normalize(t);
  return fe_is_zero_normalized(t);
}
CSP_HD bool fe_equal(const fe &a, const fe &b) {
  fe t;
  fe_sub(t, a, b);
  return fe_is_zero(t);
}

// 32 bytes little-endian <-> limbs (src/cudasp_extension.cu:59-85 does this on the host,
// one byte at a time, up to ten times per row; here it is a plain word load).
CSP_HD void fe_from_le_words(fe &r, const uint32_t *w) {
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = w[i];
}
// inputs >= p are legal blobs: bring into [0, 2^256) (already) -- nothing to do, weak form.

// ------------------------------------------------------------------ inversion
template <int N>
CSP_HD void fe_sqr_n(fe &r, const fe &a) {
  fe_sqr(r, a);
  for (int i = 1; i < N; i++) fe_sqr(r, r);
}
// a^(p-2).  p-2 in binary: 223 ones, 0, 22 ones, 0000, 1, 0, 11, 0, 1 (low bits 0x...FFFFFC2D).
// Addition chain on runs of ones: x2,x3,x6,x9,x11,x22,x44,x88,x176,x220,x223 (255 S + 15 M).
CSP_HD void fe_inv(fe &r, const fe &a) {
  fe x2, x3, x6, x9, x11, x22, x44, x88, x176, x220, x223, t;
  fe_sqr(x2, a);
  fe_mul(x2, x2, a);
  fe_sqr(x3, x2);
  fe_mul(x3, x3, a);
  fe_sqr_n<3>(x6, x3);
  fe_mul(x6, x6, x3);
  fe_sqr_n<3>(x9, x6);
  fe_mul(x9, x9, x3);
  fe_sqr_n<2>(x11, x9);
  fe_mul(x11, x11, x2);
  fe_sqr_n<11>(x22, x11);
  fe_mul(x22, x22, x11);
  fe_sqr_n<22>(x44, x22);
  fe_mul(x44, x44, x22);
  fe_sqr_n<44>(x88, x44);
  fe_mul(x88, x88, x44);
  fe_sqr_n<88>(x176, x88);
  fe_mul(x176, x176, x88);
  fe_sqr_n<44>(x220, x176);
  fe_mul(x220, x220, x44);
  fe_sqr_n<3>(x223, x220);
  fe_mul(x223, x223, x3);
  // tail: 0, 22 ones, 0000, 1, 0, 11, 0, 1
  fe_sqr_n<23>(t, x223);
  fe_mul(t, t, x22);
  fe_sqr_n<5>(t, t);
  fe_mul(t, t, a);
  fe_sqr_n<3>(t, t);
  fe_mul(t, t, x2);
  fe_sqr_n<2>(t, t);
  fe_mul(r, t, a);
}


// a^((p+1)/4): the square root of a when a is a quadratic residue (p = 3 mod 4); otherwise the square of the
// result is -a.  (p+1)/4 = 2^254 - 2^30 - 244: blocks of ones of length 223, 22 and 2 -> 253 S + 13 M.
// Used only to decompress 33-byte SEC1 tweak keys (reference CPU semantics: scripts/decompress_tweak_key.py:34-46).
CSP_HD void fe_sqrt_candidate(fe &r, const fe &a) {
  fe x2, x3, x6, x9, x11, x22, x44, x88, x176, x220, x223, t;
  fe_sqr(x2, a);
  fe_mul(x2, x2, a);
  fe_sqr(x3, x2);
  fe_mul(x3, x3, a);
  fe_sqr_n<3>(x6, x3);
  fe_mul(x6, x6, x3);
  fe_sqr_n<3>(x9, x6);
  fe_mul(x9, x9, x3);
  fe_sqr_n<2>(x11, x9);
  fe_mul(x11, x11, x2);
  fe_sqr_n<11>(x22, x11);
  fe_mul(x22, x22, x11);
  fe_sqr_n<22>(x44, x22);
  fe_mul(x44, x44, x22);
  fe_sqr_n<44>(x88, x44);
  fe_mul(x88, x88, x44);
  fe_sqr_n<88>(x176, x88);
  fe_mul(x176, x176, x88);
  fe_sqr_n<44>(x220, x176);
  fe_mul(x220, x220, x44);
  fe_sqr_n<3>(x223, x220);
  fe_mul(x223, x223, x3);
  fe_sqr_n<23>(t, x223);
  fe_mul(t, t, x22);
  fe_sqr_n<6>(t, t);
  fe_mul(t, t, x2);
  fe_sqr_n<2>(r, t);
}

}  // namespace csp
