// cpu_scan_openssl.cpp -- the BIP-352 scan of one table on the host cores with OpenSSL 3 doing the curve work.
//
// BENCH / TEST INFRASTRUCTURE, NOT PRODUCT CODE: this is baseline B-cpu of BASELINE.md section 2 ("a libsecp256k1-class CPU
// scan if it installs offline" -- libsecp256k1 is not in the image, OpenSSL 3 is; the reference links OpenSSL but never calls
// it, CMakeLists.txt:15,92-93).  Only bench.py's cpu_baseline leg and tests/test_cpu_baseline.py load it; the product has no
// CPU path.  Row semantics are the reference's (SURVEY.md 8c rules 1-9, scripts/verify_bip352_test_vector.py:149-210):
//   S = scan_sk * tweak;  t = SHA256(TH || TH || ser(S) || 00000000);  P0 = t*G + B_spend;
//   cand_0 = int64(x(P0) >> 192), cand_i = int64(x(P0 + L_i) >> 192);  row matches iff a cand is among its outputs.
// EC_POINT_mul / EC_POINT_add / EC_POINT_get_affine_coordinates on NID_secp256k1, SHA256() for the hash; rows are split
// over std::threads in contiguous ranges.  Cross-checked against oracle/ on 10^4 rows (tests/test_cpu_baseline.py).
#include <openssl/bn.h>
#include <openssl/ec.h>
#include <openssl/obj_mac.h>
#include <openssl/sha.h>

#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace {

BIGNUM *bn_from_le(const uint8_t *le, int n) {
  uint8_t be[64];
  for (int i = 0; i < n; i++) be[i] = le[n - 1 - i];
  return BN_bin2bn(be, n, nullptr);
}
EC_POINT *point_from_blob(const EC_GROUP *g, const uint8_t b[64], BN_CTX *ctx) {      // x_LE || y_LE
  BIGNUM *x = bn_from_le(b, 32), *y = bn_from_le(b + 32, 32);
  EC_POINT *p = EC_POINT_new(g);
  const int ok = EC_POINT_set_affine_coordinates(g, p, x, y, ctx);
  BN_free(x); BN_free(y);
  if (!ok) { EC_POINT_free(p); return nullptr; }
  return p;
}
int64_t upper64(const BIGNUM *x) {               // src/cudasp_gpu.cu:67-71
  uint8_t be[32];
  BN_bn2binpad(x, be, 32);
  uint64_t v = 0;
  for (int i = 0; i < 8; i++) v = (v << 8) | be[i];
  return (int64_t)v;
}

struct Job {
  const uint8_t *sk, *spend, *labels, *tweaks;
  uint32_t L;
  const int64_t *outputs;
  const uint32_t *off;
  uint8_t *flags;
  int64_t *cands;
};

void scan_range(const Job &j, uint32_t lo, uint32_t hi) {
  EC_GROUP *g = EC_GROUP_new_by_curve_name(NID_secp256k1);
  BN_CTX *ctx = BN_CTX_new();
  BIGNUM *sk = bn_from_le(j.sk, 32), *x = BN_new(), *y = BN_new(), *t = BN_new();
  EC_POINT *B = point_from_blob(g, j.spend, ctx), *S = EC_POINT_new(g), *T = EC_POINT_new(g), *P0 = EC_POINT_new(g),
           *Q = EC_POINT_new(g);
  std::vector<EC_POINT *> lab(j.L);
  for (uint32_t i = 0; i < j.L; i++) lab[i] = point_from_blob(g, j.labels + 64 * (size_t)i, ctx);
  uint8_t th[32], buf[64 + 37], hash[32];
  SHA256(reinterpret_cast<const uint8_t *>("BIP0352/SharedSecret"), 20, th);
  memcpy(buf, th, 32);
  memcpy(buf + 32, th, 32);
  std::vector<int64_t> c(1 + j.L);
  for (uint32_t r = lo; r < hi; r++) {
    EC_POINT *P = point_from_blob(g, j.tweaks + 64 * (size_t)r, ctx);
    bool ok = P && B;
    if (ok) ok = EC_POINT_mul(g, S, nullptr, P, sk, ctx) && EC_POINT_get_affine_coordinates(g, S, x, y, ctx);
    if (ok) {
      buf[64] = (uint8_t)(0x02 + (BN_is_odd(y) ? 1 : 0));
      BN_bn2binpad(x, buf + 65, 32);
      memset(buf + 97, 0, 4);                                  // k = 0 (src/cudasp_gpu.cu:278-282)
      SHA256(buf, sizeof buf, hash);
      BN_bin2bn(hash, 32, t);                                  // not reduced by hand: EC_POINT_mul reduces mod n (same point)
      ok = EC_POINT_mul(g, T, t, nullptr, nullptr, ctx) && EC_POINT_add(g, P0, T, B, ctx) &&
           EC_POINT_get_affine_coordinates(g, P0, x, y, ctx);
    }
    if (ok) {
      c[0] = upper64(x);
      for (uint32_t i = 0; i < j.L; i++) {
        c[1 + i] = 0;                                          // infinity -> affine (1, 0) in the reference -> candidate 0
        if (lab[i] && EC_POINT_add(g, Q, P0, lab[i], ctx) && !EC_POINT_is_at_infinity(g, Q) &&
            EC_POINT_get_affine_coordinates(g, Q, x, y, ctx))
          c[1 + i] = upper64(x);
      }
    } else {
      for (auto &v : c) v = 0;
    }
    bool hit = false;
    if (ok && j.off)
      for (uint32_t o = j.off[r]; o < j.off[r + 1]; o++)
        for (uint32_t i = 0; i <= j.L; i++) hit |= j.outputs[o] == c[i];
    if (j.flags) j.flags[r] = hit ? 1 : 0;
    if (j.cands) memcpy(j.cands + (size_t)r * (1 + j.L), c.data(), (1 + j.L) * 8);
    EC_POINT_free(P);
  }
  for (EC_POINT *p : lab) EC_POINT_free(p);
  EC_POINT_free(B); EC_POINT_free(S); EC_POINT_free(T); EC_POINT_free(P0); EC_POINT_free(Q);
  BN_free(sk); BN_free(x); BN_free(y); BN_free(t);
  BN_CTX_free(ctx);
  EC_GROUP_free(g);
}

}  // namespace

extern "C" int cpu_scan_openssl(const uint8_t sk[32], const uint8_t spend[64], const uint8_t *labels, uint32_t L,
                                const uint8_t *tweaks, const int64_t *outputs, const uint32_t *off, uint32_t rows,
                                uint8_t *flags, int64_t *cands, int threads) {
  Job j{sk, spend, labels, tweaks, L, outputs, off, flags, cands};
  if (threads < 1) threads = 1;
  std::vector<std::thread> th;
  const uint32_t share = (rows + threads - 1) / threads;
  for (int t = 0; t < threads; t++) {
    const uint32_t lo = (uint32_t)t * share, hi = lo + share < rows ? lo + share : rows;
    if (lo >= hi) break;
    th.emplace_back(scan_range, std::cref(j), lo, hi);
  }
  for (auto &x : th) x.join();
  return 0;
}
