// query_plan.hpp -- everything that is computed once per query (host side), shared by the C ABI
// (cudasp_capi.cu) and by the CPU-side logic tests (tests/hostcheck).
//
// Reference counterpart: the key conversions repeated per batch in ProcessBatch
// (src/cudasp_extension.cu:277-297,343-361) and per thread in CheckMatchesWithLabelsKernel
// (spend key / labels to Montgomery form, src/cudasp_gpu.cu:109-116,154-161).
#pragma once
#include <algorithm>
#include <mutex>
#include <string>
#include <vector>
#include "host_ec.hpp"
#include "rowfn.cuh"

namespace csp {
namespace host {

struct GTabHost {
  GTabSeeds seeds;         // per window: r_w*G and 2^(wbits*w)*G -- what the device builder starts from
  Big offset_sum;          // sum of r_w mod n
  HPoint offset_point;     // offset_sum * G
};
inline const GTabHost &gtab_host() {
  static GTabHost g;
  static std::once_flag once;
  std::call_once(once, [] {
    gtab_seeds(g.seeds, GTAB_WBITS, GTAB_NWIN);
    g.offset_sum = g.seeds.offset_sum;
    g.offset_point = h_mul(g.offset_sum, generator());
  });
  return g;
}
// the whole table built on the HOST: CPU logic tests only (the product builds it on the device)
inline const std::vector<ge> &gtab_host_table() {
  static std::vector<ge> table;
  static std::once_flag once;
  std::call_once(once, [] { Big s; build_gtab(table, s, GTAB_WBITS, GTAB_NWIN); });
  return table;
}

static const int LADDER_SEARCH = 2;   // decomposition search radius (make_query_plan)

struct QueryPlan {
  LadderParams lp;
  FinalParams fp;
  std::vector<ge> labels;
  int n_dbl, n_add;
};

inline bool point_from_blob(HPoint &p, const uint8_t b[64]) {
  fe_from_le_bytes(p.x, b);
  fe_from_le_bytes(p.y, b + 32);
  p.inf = false;
  return true;
}

// Checks that the op list never adds a point to itself or its negative: simulate the scalar
// (a + b*lambda mod n) the accumulator holds before every ADD.  With `expect` given, the scalar the whole list
// evaluates to must also equal it (k = scan_sk mod n): a wrong lattice constant or sign in a decomposition variant
// would otherwise give a wrong shared secret for every row without any error.
inline bool ladder_plan_is_regular(const LadderPlan &p, const Big *expect = nullptr) {
  const Big n = N_ORDER(), lam = LAMBDA();
  Big v;  // accumulator scalar mod n
  for (int i = 0; i < p.nops; i++) {
    uint8_t op = p.ops[i];
    if (op & OP_DBL) { v = mod(shl1(v), n); continue; }
    Big d((uint64_t)(2 * (op & 3) + 1));
    if (op & OP_LAMBDA) d = mod(mul(d, lam), n);
    if (op & OP_NEG) d = sub(n, d);
    if (!(op & OP_SET)) {
      if (cmp(v, d) == 0) return false;
      if (cmp(mod(add(v, d), n), Big()) == 0) return false;
      if (v.is_zero()) return false;
    }
    v = mod(add(v, d), n);
  }
  if (expect && cmp(v, *expect) != 0) return false;
  return !v.is_zero();
}

// returns 0, or a negative CUDASP_ERR_* with err set
inline int make_query_plan(QueryPlan &q, const uint8_t sk[32], const uint8_t spend[64], const uint8_t *labels,
                           uint32_t n_labels, std::string &err) {
  memset(&q.lp, 0, sizeof q.lp);
  memset(&q.fp, 0, sizeof q.fp);
  // The scan key is fixed for the whole query, so its recoding may be chosen at leisure: among the 25 GLV
  // decompositions k1 + i*a1 + j*a2, k2 + i*b1 + j*b2 (|i|, |j| <= 2; all equal k mod n, all about 128 bits) take the
  // one whose interleaved NAFs cost K1 the fewest field multiplications -- 2 % fewer on average than the balanced split.
  // Candidates are tried cheapest first; the regularity check (big-number arithmetic) runs until one passes.
  LadderPlan lp = make_ladder_plan(sk);
  const Big k_mod_n = mod(Big::from_le(sk, 32), N_ORDER());
  if (!lp.zero) {
    struct Cand { int cost, i, j; };
    std::vector<Cand> cands;
    for (int i = -LADDER_SEARCH; i <= LADDER_SEARCH; i++)
      for (int j = -LADDER_SEARCH; j <= LADDER_SEARCH; j++) {
        LadderPlan c = (i || j) ? make_ladder_plan_variant(sk, i, j) : lp;
        if (c.nops > LADDER_MAX_OPS_DEV) continue;
        cands.push_back({ladder_plan_cost(c), i, j});
      }
    std::stable_sort(cands.begin(), cands.end(), [](const Cand &a, const Cand &b) {
      if (a.cost != b.cost) return a.cost < b.cost;
      return (a.i * a.i + a.j * a.j) < (b.i * b.i + b.j * b.j);     // ties: the decomposition closest to the balanced one
    });
    for (const Cand &c : cands) {
      LadderPlan p = (c.i || c.j) ? make_ladder_plan_variant(sk, c.i, c.j) : lp;
      if (ladder_plan_is_regular(p, &k_mod_n)) { lp = p; break; }   // a variant that does not evaluate to k is skipped
    }
  }
  q.n_dbl = lp.n_dbl;
  q.n_add = lp.n_add;
  if (!lp.zero) {
    if (lp.nops > LADDER_MAX_OPS_DEV) { err = "internal: ladder op list too long"; return -4; }
    if (!ladder_plan_is_regular(lp, &k_mod_n)) { err = "scan_private_key yields a degenerate ladder schedule"; return -4; }
    memcpy(q.lp.ops, lp.ops, (size_t)lp.nops);
    q.lp.nops = lp.nops;
  }
  for (int i = 0; i < 8; i++) q.lp.beta[i] = BETA_LIMBS[i];
  q.fp.zero_scalar = lp.zero ? 1 : 0;
  bip352_midstate(q.fp.midstate);
  // spend_adj = B_spend - offset_point
  HPoint B;
  point_from_blob(B, spend);
  const GTabHost &g = gtab_host();
  HPoint adj = h_add(B, h_neg(g.offset_point));
  q.fp.spend_adj_inf = adj.inf ? 1 : 0;
  q.fp.spend_adj_x = adj.x;
  q.fp.spend_adj_y = adj.y;
  q.labels.resize(n_labels);
  for (uint32_t i = 0; i < n_labels; i++) {
    fe_from_le_bytes(q.labels[i].x, labels + 64 * (size_t)i);
    fe_from_le_bytes(q.labels[i].y, labels + 64 * (size_t)i + 32);
    // labels >= p are legal blobs; canonicalise once so that xL - x0 == 0 is an exact test
    fe_normalize(q.labels[i].x);
    fe_normalize(q.labels[i].y);
  }
  return 0;
}

}  // namespace host
}  // namespace csp
