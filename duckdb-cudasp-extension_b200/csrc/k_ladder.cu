// k_ladder.cu -- K1 (kernels.cuh k_ladder) and its launcher, compiled once per key format:
//   -DCSP_LADDER_KEYFMT=0  64-byte keys x_LE || y_LE (the reference's layout)   -> launch::ladder_xy64
//   -DCSP_LADDER_KEYFMT=1  33-byte SEC1 compressed keys, decompressed in-kernel -> launch::ladder_sec1
#include <mutex>
#include "kernels.cuh"
#include "launch.hpp"

#ifndef CSP_LADDER_KEYFMT
#error "compile with -DCSP_LADDER_KEYFMT=0 or 1"
#endif

namespace csp {
namespace launch {
namespace {

const int MAX_DEVICES = 64;

template <int BLOCK, int MINB, int SYNC, int KEYFMT>
cudaError_t ladder_t(const uint8_t *d_tweaks, fe *X, fe *Y, fe *Z, uint32_t rows, const LadderParams &lp, uint8_t *invalid,
                     cudaStream_t st) {
  // the dynamic shared memory attribute is per DEVICE (and per instantiation): set once on each device this shape runs on
  static std::once_flag once[MAX_DEVICES];
  const int smem = LADDER_SMEM_WORDS_PER_THREAD * 4 * BLOCK;
  int dev = 0;
  cudaGetDevice(&dev);
  cudaError_t attr = cudaSuccess;
  std::call_once(once[dev >= 0 && dev < MAX_DEVICES ? dev : 0], [&] {
    attr = cudaFuncSetAttribute(k_ladder<BLOCK, MINB, SYNC, KEYFMT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  });
  if (attr != cudaSuccess) return attr;
  const uint32_t grid = (rows + BLOCK - 1) / BLOCK;
  k_ladder<BLOCK, MINB, SYNC, KEYFMT><<<grid, BLOCK, smem, st>>>(d_tweaks, X, Y, Z, rows, lp, invalid);
  return cudaGetLastError();
}

}  // namespace

// Launch shapes.  Variant 0 is the production default; the others exist for tuning (CUDASP_LADDER_VARIANT, tuning builds).
#if CSP_LADDER_KEYFMT == 0
cudaError_t ladder_xy64(int variant, const uint8_t *d_tweaks, fe *X, fe *Y, fe *Z, uint32_t rows, const LadderParams &lp,
                        uint8_t *invalid, cudaStream_t st) {
  switch (variant) {
    default:
    case 0: return ladder_t<128, 4, 1, 0>(d_tweaks, X, Y, Z, rows, lp, invalid, st);   // best measured (profiles/r01c)
#if defined(CSP_LADDER_TUNING)   // the other shapes cost 25 s of ptxas each: only in tuning builds (VARIANT_DEFS=-DCSP_LADDER_TUNING)
    case 1: return ladder_t<128, 3, 1, 0>(d_tweaks, X, Y, Z, rows, lp, invalid, st);
    case 2: return ladder_t<128, 4, 0, 0>(d_tweaks, X, Y, Z, rows, lp, invalid, st);
    case 3: return ladder_t<256, 2, 1, 0>(d_tweaks, X, Y, Z, rows, lp, invalid, st);
#endif
  }
}
#else
cudaError_t ladder_sec1(int, const uint8_t *d_tweaks, fe *X, fe *Y, fe *Z, uint32_t rows, const LadderParams &lp,
                        uint8_t *invalid, cudaStream_t st) {
  return ladder_t<128, 4, 1, 1>(d_tweaks, X, Y, Z, rows, lp, invalid, st);   // compressed keys: production shape only
}
#endif

}  // namespace launch
}  // namespace csp
