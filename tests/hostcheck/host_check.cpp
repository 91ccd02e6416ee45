// host_check.cpp -- TEST INFRASTRUCTURE.  Compiles the product's __host__ __device__ row functions
// (duckdb-cudasp-extension_b200/csrc/{fe,ge,sha256,rowfn}.cuh, portable arithmetic path) and the
// host precompute (scalar_host.hpp, host_ec.hpp, query_plan.hpp) with g++, and drives them row by
// row on the CPU, so that `pytest -m "not gpu"` can compare the exact schedule logic the kernels
// execute (GLV/wNAF op list, isomorphic table, offset windows, batch inversion, label rule)
// against the oracle.  It is never linked into libcudasp_b200.so and is not a CPU fallback.
#include <stdint.h>
#include <string.h>
#include <cstring>
#include <string>
#include <vector>

#include "query_plan.hpp"
#include "rowfn.cuh"

using namespace csp;
using namespace csp::host;

namespace {
struct ArrTab {
  uint32_t w[4][3][8];
  uint32_t get(uint32_t e, uint32_t c, int l) const { return w[e][c][l]; }
  void set(uint32_t e, uint32_t c, int l, uint32_t v) { w[e][c][l] = v; }
};
void batch_inv_all(std::vector<fe> &vals, uint32_t T) {
  std::vector<fe> scratch(vals.size());
  uint32_t n = (uint32_t)vals.size();
  if (T == 0) T = 1;
  for (uint32_t tid = 0; tid < T; tid++) batch_inv_thread(vals.data(), scratch.data(), n, T, tid);
}
}  // namespace

#include "piece_plan.hpp"
extern "C" {
// starts of the pieces the synchronous call cuts `rows` into (csrc/piece_plan.hpp); returns the number of pieces
int hc_piece_plan(uint32_t rows, uint32_t piece, uint32_t n_devices, uint32_t *starts, uint32_t cap) {
  const std::vector<uint32_t> s = csp::piece_plan(rows, piece, n_devices);
  for (size_t i = 0; i < s.size() && i < cap; i++) starts[i] = s[i];
  return (int)s.size() - 1;
}


// op 0 mul, 1 sqr, 2 add, 3 sub, 4 inv, 5 normalize, 6 neg, 7 lin2<1,-8>, 8 lin3<1,-1,-2>(a,b,b), 9 lin2<4,-1>
int hc_field(int op, const uint8_t *a, const uint8_t *b, uint8_t *out, uint32_t n) {
  for (uint32_t i = 0; i < n; i++) {
    fe x, y, r;
    fe_from_le_bytes(x, a + 32 * (size_t)i);
    fe_from_le_bytes(y, b + 32 * (size_t)i);
    switch (op) {
      case 0: fe_mul(r, x, y); break;
      case 1: fe_sqr(r, x); break;
      case 2: fe_add(r, x, y); break;
      case 3: fe_sub(r, x, y); break;
      case 4: fe_inv(r, x); break;
      case 6: fe_neg(r, x); break;
      case 7: fe_lin2<1, -8>(r, x, y); break;
      case 8: fe_lin3<1, -1, -2>(r, x, y, y); break;
      case 9: fe_lin2<4, -1>(r, x, y); break;
      case 10: fe_lin1<2>(r, x); break;
      case 11: fe_lin1<3>(r, x); break;
      case 12: fe_lin2<1, -2>(r, x, y); break;
      case 13: fe_lin3<1, -1, -1>(r, x, y, y); break;
      case 14: fe_dbl(r, x); break;
      case 15: fe_lin2<3, 8>(r, x, y); break;
      default: r = x; break;
    }
    fe_to_le_bytes(out + 32 * (size_t)i, r);
  }
  return 0;
}

int hc_glv(const uint8_t k_le[32], uint8_t k1_le[32], int *neg1, uint8_t k2_le[32], int *neg2) {
  Big k = mod(Big::from_le(k_le, 32), N_ORDER());
  SignedScalar a, b;
  glv_split(k, a, b);
  a.mag.to_le(k1_le, 32);
  b.mag.to_le(k2_le, 32);
  *neg1 = a.neg;
  *neg2 = b.neg;
  return 0;
}

int hc_plan(const uint8_t sk[32], uint8_t *ops, int *nops, int *n_dbl, int *n_add, int *zero, int *regular) {
  LadderPlan p = make_ladder_plan(sk);
  memcpy(ops, p.ops, (size_t)p.nops);
  *nops = p.nops; *n_dbl = p.n_dbl; *n_add = p.n_add; *zero = p.zero;
  *regular = p.zero ? 1 : (ladder_plan_is_regular(p) ? 1 : 0);
  return 0;
}

// the nine equivalent decompositions (scalar_host.hpp make_ladder_plan_variant) and the one make_query_plan picks
int hc_plan_variant(const uint8_t sk[32], int i, int j, uint8_t *ops, int *nops, int *n_dbl, int *n_add, int *regular) {
  LadderPlan p = make_ladder_plan_variant(sk, i, j);
  if (p.zero || p.nops > LADDER_MAX_OPS) { *nops = -1; return 0; }
  memcpy(ops, p.ops, (size_t)p.nops);
  *nops = p.nops; *n_dbl = p.n_dbl; *n_add = p.n_add;
  *regular = ladder_plan_is_regular(p) ? 1 : 0;
  return 0;
}
int hc_plan_chosen(const uint8_t sk[32], uint8_t *ops, int *nops, int *n_dbl, int *n_add) {
  QueryPlan q;
  std::string err;
  uint8_t spend[64];
  HPoint g = generator();
  fe_to_le_bytes(spend, g.x); fe_to_le_bytes(spend + 32, g.y);
  int rc = make_query_plan(q, sk, spend, nullptr, 0, err);
  if (rc) return rc;
  memcpy(ops, q.lp.ops, (size_t)q.lp.nops);
  *nops = q.lp.nops; *n_dbl = q.n_dbl; *n_add = q.n_add;
  return 0;
}

void hc_midstate(uint32_t out[8]) { bip352_midstate(out); }

// k * P for n points (64 B LE each) through ladder_row; out = affine 64 B LE
int hc_ladder(const uint8_t sk[32], const uint8_t *points, uint32_t n, uint8_t *out) {
  QueryPlan q;
  std::string err;
  uint8_t spend[64];
  HPoint g = generator();
  fe_to_le_bytes(spend, g.x); fe_to_le_bytes(spend + 32, g.y);
  int rc = make_query_plan(q, sk, spend, nullptr, 0, err);
  if (rc) return rc;
  if (q.fp.zero_scalar) return -100;
  for (uint32_t i = 0; i < n; i++) {
    ge P;
    fe_from_le_bytes(P.x, points + 64 * (size_t)i);
    fe_from_le_bytes(P.y, points + 64 * (size_t)i + 32);
    ArrTab tab;
    gej acc;
    ladder_row(acc, P, q.lp, tab);
    fe zi, zi2, zi3, x, y;
    fe_inv(zi, acc.z);
    fe_sqr(zi2, zi); fe_mul(zi3, zi2, zi);
    fe_mul(x, acc.x, zi2); fe_mul(y, acc.y, zi3);
    fe_to_le_bytes(out + 64 * (size_t)i, x);
    fe_to_le_bytes(out + 64 * (size_t)i + 32, y);
  }
  return 0;
}

// the whole pipeline, stage by stage like the kernels (including the three batch inversions with
// stride T), on the CPU.  cands: rows * (1 + L) int64.  Optional intermediates as in the C ABI.
int hc_scan_candidates(const uint8_t sk[32], const uint8_t spend[64], const uint8_t *labels, uint32_t L,
                       const uint8_t *tweaks, uint32_t rows, uint32_t T, int64_t *cands,
                       uint8_t *S_le, uint8_t *t_be, uint8_t *P0_le) {
  QueryPlan q;
  std::string err;
  int rc = make_query_plan(q, sk, spend, labels, L, err);
  if (rc) return rc;
  const GTabHost &g = gtab_host();
  std::vector<fe> X(rows), Y(rows), Z(rows);
  if (!q.fp.zero_scalar) {
    for (uint32_t r = 0; r < rows; r++) {
      ge P;
      fe_from_le_bytes(P.x, tweaks + 64 * (size_t)r);
      fe_from_le_bytes(P.y, tweaks + 64 * (size_t)r + 32);
      ArrTab tab;
      gej acc;
      ladder_row(acc, P, q.lp, tab);
      X[r] = acc.x; Y[r] = acc.y; Z[r] = acc.z;
    }
    batch_inv_all(Z, T);
  } else {
    for (uint32_t r = 0; r < rows; r++) { fe_set_zero(X[r]); fe_set_zero(Y[r]); fe_set_zero(Z[r]); }
  }
  for (uint32_t r = 0; r < rows; r++) {
    gej acc;
    fe sx, sy, t;
    hash_fixedbase_row(acc, X[r], Y[r], Z[r], gtab_host_table().data(), q.fp, &sx, &sy, &t);
    if (S_le) { fe_to_le_bytes(S_le + 64 * (size_t)r, sx); fe_to_le_bytes(S_le + 64 * (size_t)r + 32, sy); }
    if (t_be) for (int l = 0; l < 8; l++) {
      uint32_t wv = t.v[7 - l];
      uint8_t *o = t_be + 32 * (size_t)r + 4 * l;
      o[0] = (uint8_t)(wv >> 24); o[1] = (uint8_t)(wv >> 16); o[2] = (uint8_t)(wv >> 8); o[3] = (uint8_t)wv;
    }
    X[r] = acc.x; Y[r] = acc.y; Z[r] = acc.z;
  }
  batch_inv_all(Z, T);
  std::vector<fe> D((size_t)L * rows);
  for (uint32_t r = 0; r < rows; r++) {
    fe x0, y0;
    cands[(size_t)r * (1 + L)] = base_candidate(x0, &y0, X[r], Y[r], Z[r]);
    X[r] = x0; Y[r] = y0;
    if (P0_le) { fe_to_le_bytes(P0_le + 64 * (size_t)r, x0); fe_to_le_bytes(P0_le + 64 * (size_t)r + 32, y0); }
    for (uint32_t i = 0; i < L; i++) fe_sub(D[(size_t)i * rows + r], q.labels[i].x, x0);
  }
  if (L) {
    batch_inv_all(D, T);
    for (uint32_t r = 0; r < rows; r++)
      for (uint32_t i = 0; i < L; i++)
        cands[(size_t)r * (1 + L) + 1 + i] = label_candidate(X[r], Y[r], q.labels[i].x, q.labels[i].y, D[(size_t)i * rows + r]);
  }
  return 0;
}

// details_row (the K8 body) for one row given its affine shared secret; out: n records of
// {k, output_index, label_index, t_k[32] BE} = 3 x int32 + 32 bytes = 44 bytes each.  Returns the count.
int hc_details_row(const uint8_t sk[32], const uint8_t spend[64], const uint8_t *labels, uint32_t L, const uint8_t S_le[64],
                   const int64_t *outputs, uint32_t n_out, uint8_t *out, uint32_t cap) {
  QueryPlan q;
  std::string err;
  int rc = make_query_plan(q, sk, spend, labels, L, err);
  if (rc) return rc;
  const GTabHost &g = gtab_host();
  fe sx, sy;
  fe_from_le_bytes(sx, S_le);
  fe_from_le_bytes(sy, S_le + 32);
  uint32_t n = 0;
  details_row(sx, sy.v[0] & 1u, outputs, 0u, n_out, q.labels.data(), L, gtab_host_table().data(), q.fp,
              [&](uint32_t k, uint32_t oi, int32_t li, const fe &t) {
                if (n >= cap) return;
                uint8_t *o = out + 44 * (size_t)n++;
                int32_t h[3] = {(int32_t)k, (int32_t)oi, li};
                memcpy(o, h, 12);
                for (int l = 0; l < 8; l++) {
                  uint32_t wv = t.v[7 - l];
                  o[12 + 4 * l] = (uint8_t)(wv >> 24); o[13 + 4 * l] = (uint8_t)(wv >> 16);
                  o[14 + 4 * l] = (uint8_t)(wv >> 8); o[15 + 4 * l] = (uint8_t)wv;
                }
              });
  return (int)n;
}
int hc_tweak_in_range(const uint8_t t_le[32]) { fe t; fe_from_le_bytes(t, t_le); return tweak_in_range(t) ? 1 : 0; }

// gtab[w][d] as 64 B LE
void hc_gtab_entry(int w, int d, uint8_t out[64]) {
  const GTabHost &g = gtab_host();
  const ge &e = gtab_host_table()[(size_t)w * GTAB_ENTRIES + d];
  fe_to_le_bytes(out, e.x); fe_to_le_bytes(out + 32, e.y);
}
// the DEVICE builder's row function (rowfn.cuh gtab_chunk) on the CPU: GTAB_EPT entries from digit d0 of window w
int hc_gtab_chunk(int w, uint32_t d0, uint8_t *out /* GTAB_EPT x 64 B */) {
  const GTabHost &g = gtab_host();
  ge tmp[GTAB_EPT];
  bool ok = gtab_chunk(tmp, g.seeds.firsts[w], g.seeds.bases[w], d0);
  for (int i = 0; i < GTAB_EPT; i++) { fe_to_le_bytes(out + 64 * i, tmp[i].x); fe_to_le_bytes(out + 64 * i + 32, tmp[i].y); }
  return ok ? GTAB_EPT : -1;
}
void hc_gtab_offset(int w, uint8_t out[32]) { gtab_offset(w).to_le(out, 32); }
// 33-byte SEC1 -> 64 B LE x||y; returns 1 when valid
int hc_decompress(const uint8_t sec1[33], uint8_t out[64]) {
  ge P;
  bool ok = decompress_tweak(P, sec1);
  fe_to_le_bytes(out, P.x); fe_to_le_bytes(out + 32, P.y);
  return ok ? 1 : 0;
}
void hc_gtab_shape(int *wbits, int *nwin) { *wbits = GTAB_WBITS; *nwin = GTAB_NWIN; }
uint32_t hc_gtab_digit(const uint8_t t_le[32], int w) { fe t; fe_from_le_bytes(t, t_le); return gtab_digit(t, w); }
}
