// ge.cuh -- secp256k1 group law (y^2 = x^3 + 7) on top of fe.cuh.
//
// Replaces the reference's generated point code (gECC/include/gecc/arith/ec.h:409-415
// mixed_add = madd-2007-bl, :676-699 dbl = dbl-2009-l, :584-602 to_affine, formulas in
// gECC/scripts/ec_ops.py:292-317,351-370).  Differences by design:
//   * Jacobian doubling written as 3M + 4S + 5 fused linear combinations (fe_lin*), because on
//     this field representation an add/sub costs ~1/10 of a multiply and is worth fusing;
//   * mixed addition as 8M + 3S (fewer temporaries than 7M + 4S; same cost when S ~ 0.6 M);
//   * no per-point inversion anywhere: callers batch Z (Montgomery's trick, batch_inv kernel).
// The exceptional cases (P = +-Q, infinity) are NOT handled inside gej_dbl / gej_add_ge:
// the callers guarantee they cannot occur (fixed-scalar ladder, offset comb) or detect them
// separately (label adds).  See DESIGN.md "Exceptional cases".
#pragma once
#include "fe.cuh"

namespace csp {

struct gej {  // Jacobian: x = X/Z^2, y = Y/Z^3
  fe x, y, z;
};
struct ge {  // affine
  fe x, y;
};

// ---- fused small linear combinations  r = KA*a + KB*b (+ KC*c)  (mod p), |K| <= 8, weakly reduced
CSP_HD void fe_fold_signed_top(fe &r, int64_t top) {
  // r (8 limbs, taken mod 2^256) + top * 2^256  ==  r + top * (2^32 + 977)   (mod p),  |top| small
  int64_t d = (int64_t)r.v[0] + top * (int64_t)CSP_PC_LO;
  r.v[0] = (uint32_t)d;
  d >>= 32;
  d += (int64_t)r.v[1] + top;
  r.v[1] = (uint32_t)d;
  d >>= 32;
#pragma unroll
  for (int i = 2; i < 8; i++) {
    d += (int64_t)r.v[i];
    r.v[i] = (uint32_t)d;
    d >>= 32;
  }
  if (d) {  // rare: wrapped once more by +-2^256; the remaining correction cannot wrap again
    int64_t e = (int64_t)r.v[0] + d * (int64_t)CSP_PC_LO;
    r.v[0] = (uint32_t)e;
    e >>= 32;
    e += (int64_t)r.v[1] + d;
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
template <int KA, int KB>
CSP_HD void fe_lin2(fe &r, const fe &a, const fe &b) {
#if CSP_CHAIN
  if (lin_ok_first(KA) && lin_ok(KB)) {
    chain_lin2<(lin_ok_first(KA) ? KA : 1), (lin_ok(KB) ? KB : 1)>(r, a, b);
    return;
  }
#endif
  int64_t c = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += (int64_t)KA * (int64_t)a.v[i] + (int64_t)KB * (int64_t)b.v[i];
    r.v[i] = (uint32_t)c;
    c >>= 32;
  }
  fe_fold_signed_top(r, c);
}
template <int KA, int KB, int KC>
CSP_HD void fe_lin3(fe &r, const fe &a, const fe &b, const fe &c3) {
#if CSP_CHAIN
  if (lin_ok_first(KA) && lin_ok(KB) && lin_ok(KC)) {
    chain_lin3<(lin_ok_first(KA) ? KA : 1), (lin_ok(KB) ? KB : 1), (lin_ok(KC) ? KC : 1)>(r, a, b, c3);
    return;
  }
#endif
  int64_t c = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += (int64_t)KA * (int64_t)a.v[i] + (int64_t)KB * (int64_t)b.v[i] + (int64_t)KC * (int64_t)c3.v[i];
    r.v[i] = (uint32_t)c;
    c >>= 32;
  }
  fe_fold_signed_top(r, c);
}
template <int KA>
CSP_HD void fe_lin1(fe &r, const fe &a) {
#if CSP_CHAIN
  if (lin_ok_first(KA)) {
    chain_lin1<(lin_ok_first(KA) ? KA : 1)>(r, a);
    return;
  }
#endif
  int64_t c = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += (int64_t)KA * (int64_t)a.v[i];
    r.v[i] = (uint32_t)c;
    c >>= 32;
  }
  fe_fold_signed_top(r, c);
}

// ---- doubling, a = 0.  Two formulas, selected at compile time (CSP_DBL_2M5S):
//   3M + 4S:  XX = X^2, YY = Y^2, YYYY = YY^2, T = X*YY, M = 3*XX,
//             X3 = M^2 - 8T,  Y3 = M*(4T - X3) - 8*YYYY,  Z3 = 2*Y*Z
//   2M + 5S:  T is obtained as ((X + YY)^2 - XX - YYYY) / 2, trading one multiplication for a squaring and
//             two more linear combinations (the shape of the reference's dbl-2009-l, gECC/scripts/ec_ops.py:351-370)
CSP_HD void gej_dbl(gej &r, const gej &p) {
#if defined(CSP_DBL_2M5S)
  fe xx, yy, yyyy, t2, m, u;
  fe_sqr(xx, p.x);
  fe_sqr(yy, p.y);
  fe_add(t2, p.x, yy);
  fe_mul(u, p.y, p.z);
  fe_lin1<2>(r.z, u);
  fe_sqr(yyyy, yy);
  fe_sqr(t2, t2);
  fe_lin3<1, -1, -1>(t2, t2, xx, yyyy);   // 2*X*YY
  fe_lin1<3>(m, xx);
  fe_sqr(u, m);
  fe_lin2<1, -4>(r.x, u, t2);
  fe_lin2<2, -1>(u, t2, r.x);
  fe_mul(u, m, u);
  fe_lin2<1, -8>(r.y, u, yyyy);
#else
  fe xx, yy, yyyy, t, m, u;
  fe_sqr(xx, p.x);
  fe_sqr(yy, p.y);
  fe_mul(t, p.x, yy);
  fe_mul(u, p.y, p.z);
  fe_lin1<2>(r.z, u);
  fe_sqr(yyyy, yy);
  fe_lin1<3>(m, xx);
  fe_sqr(u, m);
  fe_lin2<1, -8>(r.x, u, t);
  fe_lin2<4, -1>(u, t, r.x);
  fe_mul(u, m, u);
  fe_lin2<1, -8>(r.y, u, yyyy);
#endif
}

// ---- mixed addition r = p + q (q affine), p != +-q, neither infinity:  8M + 3S
//   Z1Z1 = Z1^2, U2 = x2*Z1Z1, S2 = y2*Z1*Z1Z1, H = U2 - X1, R = S2 - Y1,
//   HH = H^2, HHH = H*HH, V = X1*HH, X3 = R^2 - HHH - 2V, Y3 = R*(V - X3) - Y1*HHH, Z3 = Z1*H
// If h_out is given it receives Z3 / Z1 (= H here), so callers can test H == 0 or reuse Z ratios.
// With CSP_MADD_7M4S the reference's madd-2007-bl shape (7M + 4S, gECC/scripts/ec_ops.py:292-317) is compiled instead:
// Z3 = (Z1 + H)^2 - Z1Z1 - HH replaces the multiplication Z1*H by a squaring and three more linear combinations (measured:
// DESIGN.md section 6 -- on this field layer a linear combination costs a quarter of what a squaring saves over a product).
CSP_HD void gej_add_ge(gej &r, const gej &p, const fe &qx, const fe &qy, fe *h_out = nullptr) {
#if defined(CSP_MADD_7M4S)
  fe z1z1, u2, s2, h, rr, hh, i4, j, v, t, x3, zh;
  fe_sqr(z1z1, p.z);
  fe_mul(u2, qx, z1z1);
  fe_mul(s2, qy, p.z);
  fe_mul(s2, s2, z1z1);
  fe_sub(h, u2, p.x);
  fe_add(zh, p.z, h);
  fe_lin2<2, -2>(rr, s2, p.y);
  fe_sqr(hh, h);
  fe_sqr(zh, zh);
  fe_lin3<1, -1, -1>(r.z, zh, z1z1, hh);   // 2*Z1*H
  fe_lin1<4>(i4, hh);
  fe_mul(j, h, i4);
  fe_mul(v, p.x, i4);
  fe_sqr(t, rr);
  fe_lin3<1, -1, -2>(x3, t, j, v);
  fe_sub(t, v, x3);
  fe_mul(t, rr, t);
  fe_mul(u2, p.y, j);
  fe_lin2<1, -2>(r.y, t, u2);
  r.x = x3;
  if (h_out) fe_lin1<2>(*h_out, h);        // callers use it as the ratio Z3 / Z1, which is 2H here
#else
  fe z1z1, u2, s2, h, rr, hh, hhh, v, t, x3;
  fe_sqr(z1z1, p.z);
  fe_mul(u2, qx, z1z1);
  fe_mul(s2, qy, p.z);
  fe_mul(s2, s2, z1z1);
  fe_sub(h, u2, p.x);
  fe_sub(rr, s2, p.y);
  fe_sqr(hh, h);
  fe_mul(hhh, h, hh);
  fe_mul(v, p.x, hh);
  fe_mul(r.z, p.z, h);  // r may alias p: p.z is dead from here on
  fe_sqr(t, rr);
  fe_lin3<1, -1, -2>(x3, t, hhh, v);
  fe_sub(t, v, x3);
  fe_mul(t, rr, t);
  fe_mul(u2, p.y, hhh);
  fe_sub(r.y, t, u2);
  r.x = x3;
  if (h_out) *h_out = h;
#endif
}

}  // namespace csp
