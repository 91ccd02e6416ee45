// host_ec.hpp -- host-side (CPU) curve helpers used ONLY for once-per-process / once-per-query
// precomputation: the fixed-base table gtab[w][d] = (d*2^(13w) + r_w)*G, the spend-key adjustment
// B_spend - (sum r_w)*G, and the SHA-256 tag midstate.  They run the same fe.cuh / ge.cuh code the
// kernels run (portable path), with complete (checked) addition.  No row data ever goes through
// these functions; they are not a CPU fallback for the scan.
#pragma once
#include <algorithm>
#include <thread>
#include <atomic>
#include <stdint.h>
#include <string.h>
#include <vector>
#include "fe.cuh"
#include "ge.cuh"
#include "scalar_host.hpp"
#include "sha256.cuh"

namespace csp {
namespace host {

struct HPoint {  // affine with infinity flag
  fe x, y;
  bool inf;
};

inline void fe_from_le_bytes(fe &r, const uint8_t *b) {
  for (int i = 0; i < 8; i++)
    r.v[i] = (uint32_t)b[4 * i] | ((uint32_t)b[4 * i + 1] << 8) | ((uint32_t)b[4 * i + 2] << 16) | ((uint32_t)b[4 * i + 3] << 24);
}
inline void fe_to_le_bytes(uint8_t *b, const fe &a) {
  fe t = a;
  fe_normalize(t);
  for (int i = 0; i < 8; i++) { b[4 * i] = (uint8_t)t.v[i]; b[4 * i + 1] = (uint8_t)(t.v[i] >> 8); b[4 * i + 2] = (uint8_t)(t.v[i] >> 16); b[4 * i + 3] = (uint8_t)(t.v[i] >> 24); }
}

inline HPoint generator() {
  static const uint32_t gx[8] = {0x16F81798u, 0x59F2815Bu, 0x2DCE28D9u, 0x029BFCDBu, 0xCE870B07u, 0x55A06295u, 0xF9DCBBACu, 0x79BE667Eu};
  static const uint32_t gy[8] = {0xFB10D4B8u, 0x9C47D08Fu, 0xA6855419u, 0xFD17B448u, 0x0E1108A8u, 0x5DA4FBFCu, 0x26A3C465u, 0x483ADA77u};
  HPoint g;
  for (int i = 0; i < 8; i++) { g.x.v[i] = gx[i]; g.y.v[i] = gy[i]; }
  g.inf = false;
  return g;
}

inline HPoint h_neg(const HPoint &a) {
  HPoint r = a;
  if (!a.inf) { fe_neg(r.y, a.y); fe_normalize(r.y); }
  return r;
}
inline HPoint h_dbl(const HPoint &a) {
  HPoint r;
  r.inf = true;
  fe_set_zero(r.x); fe_set_zero(r.y);
  if (a.inf || fe_is_zero(a.y)) return r;
  fe num, den, lam, t;
  fe_sqr(num, a.x);
  fe_lin1<3>(num, num);
  fe_lin1<2>(den, a.y);
  fe_inv(den, den);
  fe_mul(lam, num, den);
  fe_sqr(t, lam);
  fe_lin2<1, -2>(r.x, t, a.x);
  fe_sub(t, a.x, r.x);
  fe_mul(t, lam, t);
  fe_sub(r.y, t, a.y);
  fe_normalize(r.x); fe_normalize(r.y);
  r.inf = false;
  return r;
}
inline HPoint h_add(const HPoint &a, const HPoint &b) {   // complete
  if (a.inf) return b;
  if (b.inf) return a;
  fe dx, dy;
  fe_sub(dx, b.x, a.x);
  fe_sub(dy, b.y, a.y);
  if (fe_is_zero(dx)) {
    if (fe_is_zero(dy)) return h_dbl(a);
    HPoint r; r.inf = true; fe_set_zero(r.x); fe_set_zero(r.y);
    return r;
  }
  HPoint r;
  fe lam, t;
  fe_inv(dx, dx);
  fe_mul(lam, dy, dx);
  fe_sqr(t, lam);
  fe_lin3<1, -1, -1>(r.x, t, a.x, b.x);
  fe_sub(t, a.x, r.x);
  fe_mul(t, lam, t);
  fe_sub(r.y, t, a.y);
  fe_normalize(r.x); fe_normalize(r.y);
  r.inf = false;
  return r;
}
// k*P, k a Big (any size); plain double-and-add with the complete affine law (slow, exact)
inline HPoint h_mul(const Big &k, const HPoint &p) {
  HPoint acc; acc.inf = true; fe_set_zero(acc.x); fe_set_zero(acc.y);
  for (int i = k.bitlen() - 1; i >= 0; i--) {
    acc = h_dbl(acc);
    if (k.bit(i)) acc = h_add(acc, p);
  }
  return acc;
}

// ---- SHA-256 on the host (tag midstate, table offsets)
inline void h_sha256(uint8_t out[32], const uint8_t *msg, size_t len) {
  uint32_t st[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
  std::vector<uint8_t> buf(msg, msg + len);
  buf.push_back(0x80);
  while (buf.size() % 64 != 56) buf.push_back(0);
  uint64_t bits = (uint64_t)len * 8;
  for (int j = 7; j >= 0; j--) buf.push_back((uint8_t)(bits >> (8 * j)));
  for (size_t o = 0; o < buf.size(); o += 64) {
    uint32_t w[16];
    for (int i = 0; i < 16; i++)
      w[i] = ((uint32_t)buf[o + 4 * i] << 24) | ((uint32_t)buf[o + 4 * i + 1] << 16) | ((uint32_t)buf[o + 4 * i + 2] << 8) | buf[o + 4 * i + 3];
    sha256_compress(st, w);
  }
  for (int j = 0; j < 8; j++) { out[4 * j] = (uint8_t)(st[j] >> 24); out[4 * j + 1] = (uint8_t)(st[j] >> 16); out[4 * j + 2] = (uint8_t)(st[j] >> 8); out[4 * j + 3] = (uint8_t)st[j]; }
}
// chaining value after the block SHA256(tag)||SHA256(tag), tag = "BIP0352/SharedSecret"
inline void bip352_midstate(uint32_t mid[8]) {
  static const char tag[] = "BIP0352/SharedSecret";
  uint8_t th[32];
  h_sha256(th, (const uint8_t *)tag, 20);
  uint32_t st[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
  uint32_t w[16];
  for (int i = 0; i < 16; i++) {
    const uint8_t *b = th + 4 * (i & 7);
    w[i] = ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3];
  }
  sha256_compress(st, w);
  memcpy(mid, st, sizeof st);
}

// ---- fixed-base table
// r_w = SHA256("cudasp-b200/gtab-offset" || byte(w)) mod n : fixed, public, nothing-up-my-sleeve
// offsets whose only role is to make every partial sum a "random" multiple of G, so that the
// incomplete mixed addition in k_hash_fixedbase never meets P = +-Q or infinity.
inline Big gtab_offset(int w) {
  uint8_t msg[32] = "cudasp-b200/gtab-offset";
  size_t len = strlen((const char *)msg);
  msg[len] = (uint8_t)w;
  uint8_t h[32], le[32];
  h_sha256(h, msg, len + 1);
  for (int i = 0; i < 32; i++) le[i] = h[31 - i];
  return mod(Big::from_le(le, 32), N_ORDER());
}
// The seeds of the table: per window w its first entry r_w*G and its step 2^(wbits*w)*G (both affine, canonical),
// plus sum_w r_w mod n.  The product builds the table from these ON THE DEVICE (k_build_gtab, kernels.cuh); the host
// builder below is kept for the CPU logic tests and as the independent check of the device-built table.
struct GTabSeeds {
  std::vector<ge> firsts, bases;
  Big offset_sum;
};
inline void gtab_seeds(GTabSeeds &sd, int wbits, int nwin) {
  const Big n = N_ORDER();
  sd.firsts.resize((size_t)nwin);
  sd.bases.resize((size_t)nwin);
  sd.offset_sum = Big();
  HPoint base = generator();   // 2^(wbits*w) * G
  for (int w = 0; w < nwin; w++) {
    Big rw = gtab_offset(w);
    sd.offset_sum = mod(add(sd.offset_sum, rw), n);
    HPoint f = h_mul(rw, generator());
    sd.firsts[w].x = f.x; sd.firsts[w].y = f.y;
    sd.bases[w].x = base.x; sd.bases[w].y = base.y;
    fe_normalize(sd.firsts[w].x); fe_normalize(sd.firsts[w].y);
    fe_normalize(sd.bases[w].x); fe_normalize(sd.bases[w].y);
    for (int b = 0; b < wbits; b++) base = h_dbl(base);
  }
}
// table[w * entries + d] = (d * 2^(wbits*w) + r_w) * G ; returns also sum_w r_w mod n.
// The windows are independent once their base points 2^(wbits*w)*G are known, so they are filled by a few host threads
// (0.25 s single-threaded for 20 windows of 13 bits, the shape the CPU tests use).  HOST builder: tests only, see gtab_seeds.
inline void build_gtab(std::vector<ge> &table, Big &offset_sum, int wbits, int nwin) {
  const int entries = 1 << wbits;
  table.resize((size_t)nwin * entries);
  GTabSeeds sd;
  gtab_seeds(sd, wbits, nwin);
  offset_sum = sd.offset_sum;
  const std::vector<ge> &firsts = sd.firsts, &bases = sd.bases;
  auto fill = [&](int w) {
    std::vector<gej> jac((size_t)entries);
    std::vector<fe> pre((size_t)entries);
    // Jacobian running sum cur += base, then ONE inversion for the whole window (Montgomery's trick).
    // The offsets make every partial sum a pseudo-random multiple of G: +-base or infinity cannot occur
    // (checked by the tests through Z != 0, which the incomplete addition would produce on a collision).
    jac[0].x = firsts[w].x; jac[0].y = firsts[w].y; fe_set_u32(jac[0].z, 1);
    for (int d = 1; d < entries; d++) gej_add_ge(jac[d], jac[d - 1], bases[w].x, bases[w].y);
    fe acc;
    fe_set_u32(acc, 1);
    for (int d = 0; d < entries; d++) { pre[d] = acc; fe_mul(acc, acc, jac[d].z); }
    fe inv;
    fe_inv(inv, acc);
    for (int d = entries - 1; d >= 0; d--) {
      fe zi, zi2, zi3;
      fe_mul(zi, inv, pre[d]);
      fe_mul(inv, inv, jac[d].z);
      fe_sqr(zi2, zi);
      fe_mul(zi3, zi2, zi);
      ge &e = table[(size_t)w * entries + d];
      fe_mul(e.x, jac[d].x, zi2);
      fe_mul(e.y, jac[d].y, zi3);
      fe_normalize(e.x);
      fe_normalize(e.y);
    }
  };
  unsigned nt = std::thread::hardware_concurrency();
  nt = std::max(1u, std::min({nt ? nt : 1u, 8u, (unsigned)nwin}));
  if (nt == 1) {
    for (int w = 0; w < nwin; w++) fill(w);
    return;
  }
  std::atomic<int> next{0};
  std::vector<std::thread> pool;
  for (unsigned t = 0; t < nt; t++)
    pool.emplace_back([&] { for (int w = next++; w < nwin; w = next++) fill(w); });
  for (std::thread &t : pool) t.join();
}

}  // namespace host
}  // namespace csp
