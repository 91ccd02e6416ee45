// k_stages.cu -- every kernel but the ladder (kernels.cuh: K0, K2-K8, field KATs) and their launchers.
#define CSP_DEFINE_STAGE_KERNELS
#include "kernels.cuh"
#include "launch.hpp"

namespace csp {
namespace launch {
namespace {
const int HASH_BLOCK = 128, HASH_MINB = 3;
const int FINAL_BLOCK = 128;
const uint32_t DETAILS_GRID = 64;   // K8: 64 one-warp blocks stride over the matched rows
}  // namespace

cudaError_t batch_inv(fe *vals, fe *scratch, uint32_t n, uint32_t per_thread, cudaStream_t st) {
  // elements per Fermat inversion: more of them means less work (270 multiplications per inversion) but a longer
  // single-thread chain
  uint32_t T = (n + per_thread - 1) / per_thread;
  T = (T + 127u) / 128u * 128u;
  T = T < n ? T : n;
  if (T == 0) return cudaSuccess;
  k_batch_inv<<<(T + 127) / 128, 128, 0, st>>>(vals, scratch, n, T);
  return cudaGetLastError();
}

cudaError_t hash_fixedbase(fe *X, fe *Y, fe *Z, uint32_t rows, const ge *gtab, const FinalParams &fp, uint8_t *dbg_S,
                           uint8_t *dbg_t, fe *keep_sx, uint8_t *keep_sodd, cudaStream_t st) {
  const uint32_t grid = (rows + HASH_BLOCK - 1) / HASH_BLOCK;
  k_hash_fixedbase<HASH_BLOCK, HASH_MINB><<<grid, HASH_BLOCK, 0, st>>>(X, Y, Z, rows, gtab, fp, dbg_S, dbg_t, keep_sx, keep_sodd);
  return cudaGetLastError();
}

cudaError_t final_base(fe *X, fe *Y, const fe *Zinv, uint32_t rows, const int64_t *outputs, const uint32_t *out_off,
                       const ge *labels, uint32_t n_labels, fe *D, uint8_t *flags, uint32_t *match_ids, uint32_t *n_matches,
                       int64_t *cands_out, uint8_t *dbg_P0, const uint8_t *invalid, cudaStream_t st) {
  const uint32_t grid = (rows + FINAL_BLOCK - 1) / FINAL_BLOCK;
  k_final_base<FINAL_BLOCK><<<grid, FINAL_BLOCK, 0, st>>>(X, Y, Zinv, rows, outputs, out_off, labels, n_labels, D, flags, match_ids,
                                                         n_matches, cands_out, dbg_P0, invalid);
  return cudaGetLastError();
}

cudaError_t final_labels(const fe *X0, const fe *Y0, const fe *Dinv, uint32_t rows, const int64_t *outputs,
                         const uint32_t *out_off, const ge *labels, uint32_t n_labels, const uint8_t *flags, uint32_t *match_ids,
                         uint32_t *n_matches, int64_t *cands_out, const uint8_t *invalid, cudaStream_t st) {
  const uint32_t grid = (rows + FINAL_BLOCK - 1) / FINAL_BLOCK;
  k_final_labels<FINAL_BLOCK><<<grid, FINAL_BLOCK, 0, st>>>(X0, Y0, Dinv, rows, outputs, out_off, labels, n_labels, flags, match_ids,
                                                           n_matches, cands_out, invalid);
  return cudaGetLastError();
}

cudaError_t details(const fe *SX, const uint8_t *SODD, const uint32_t *match_ids, const uint32_t *n_matches,
                    const int64_t *outputs, const uint32_t *out_off, const ge *labels, uint32_t n_labels, const ge *gtab,
                    const FinalParams &fp, MatchDetail *out, uint32_t *n_out, uint32_t cap, cudaStream_t st) {
  k_details<<<DETAILS_GRID, 32, 0, st>>>(SX, SODD, match_ids, n_matches, outputs, out_off, labels, n_labels, gtab, fp, out, n_out, cap);
  return cudaGetLastError();
}

cudaError_t build_gtab(ge *table, const ge *firsts, const ge *bases, uint32_t *bad) {
  const uint64_t threads = (uint64_t)(GTAB_NWIN - 1) * (GTAB_ENTRIES / GTAB_EPT) + GTAB_TOP_ENTRIES / GTAB_EPT;
  k_build_gtab<<<(unsigned)((threads + 127) / 128), 128>>>(table, firsts, bases, bad);
  return cudaGetLastError();
}

cudaError_t test_field(int op, const fe *a, const fe *b, fe *out, uint32_t n) {
  k_test_field<<<(n + 127) / 128, 128>>>(op, a, b, out, n);
  return cudaGetLastError();
}

}  // namespace launch
}  // namespace csp
