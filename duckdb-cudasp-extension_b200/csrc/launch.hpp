// launch.hpp -- host-callable launchers of the scan kernels.
//
// Each kernel lives in its own translation unit (k_ladder.cu: K1, twice -- one object per key format; k_stages.cu: K0, K2-K8
// and the field KAT kernel) so that the objects build in parallel and a change to one kernel does not recompile the
// ladder (25 s of ptxas per instantiation).  cudasp_capi.cu (slots, streams, the C ABI) only sees these declarations.
// Every launcher enqueues on `st`, returns cudaGetLastError() of its own launch, and never synchronises.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "rowfn.cuh"

namespace csp {
namespace launch {

// K1: S = scan_sk * tweak, Jacobian, one row per thread.  `invalid` (rows bytes) is written for keyfmt 1 only.
cudaError_t ladder_xy64(int variant, const uint8_t *d_tweaks, fe *X, fe *Y, fe *Z, uint32_t rows, const LadderParams &lp,
                        uint8_t *invalid, cudaStream_t st);
cudaError_t ladder_sec1(int variant, const uint8_t *d_tweaks, fe *X, fe *Y, fe *Z, uint32_t rows, const LadderParams &lp,
                        uint8_t *invalid, cudaStream_t st);
// K2/K4/K6: vals[i] <- vals[i]^-1 (0 stays 0), n elements, `per_thread` of them share one Fermat inversion
cudaError_t batch_inv(fe *vals, fe *scratch, uint32_t n, uint32_t per_thread, cudaStream_t st);
// K3
cudaError_t hash_fixedbase(fe *X, fe *Y, fe *Z, uint32_t rows, const ge *gtab, const FinalParams &fp, uint8_t *dbg_S,
                           uint8_t *dbg_t, fe *keep_sx, uint8_t *keep_sodd, cudaStream_t st);
// K5 / K7
cudaError_t final_base(fe *X, fe *Y, const fe *Zinv, uint32_t rows, const int64_t *outputs, const uint32_t *out_off,
                       const ge *labels, uint32_t n_labels, fe *D, uint8_t *flags, uint32_t *match_ids, uint32_t *n_matches,
                       int64_t *cands_out, uint8_t *dbg_P0, const uint8_t *invalid, cudaStream_t st);
cudaError_t final_labels(const fe *X0, const fe *Y0, const fe *Dinv, uint32_t rows, const int64_t *outputs,
                         const uint32_t *out_off, const ge *labels, uint32_t n_labels, const uint8_t *flags, uint32_t *match_ids,
                         uint32_t *n_matches, int64_t *cands_out, const uint8_t *invalid, cudaStream_t st);
// K8 (opt-in match details)
cudaError_t details(const fe *SX, const uint8_t *SODD, const uint32_t *match_ids, const uint32_t *n_matches,
                    const int64_t *outputs, const uint32_t *out_off, const ge *labels, uint32_t n_labels, const ge *gtab,
                    const FinalParams &fp, MatchDetail *out, uint32_t *n_out, uint32_t cap, cudaStream_t st);
// K0: the fixed-base table (default stream of the current device)
cudaError_t build_gtab(ge *table, const ge *firsts, const ge *bases, uint32_t *bad);
// field KATs (tests)
cudaError_t test_field(int op, const fe *a, const fe *b, fe *out, uint32_t n);

}  // namespace launch
}  // namespace csp
