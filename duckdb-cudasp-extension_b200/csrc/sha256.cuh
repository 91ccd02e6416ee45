// sha256.cuh -- the one SHA-256 compression the scan needs per row.
//
// Replaces ComputeTaggedHashesKernel -> hash::tagged_hash -> hash::sha256
// (src/cudasp_gpu.cu:288-314; gECC/include/gecc/hash/sha256.h:271-303,187-225,68-120), which
// hashes the tag twice per row and runs three byte-oriented compressions through a 960-byte
// local buffer.  Here SHA256(tag)||SHA256(tag) is exactly one 64-byte block, so its chaining
// value is a per-process constant (computed once on the host with this same function), and the
// 37-byte message 02/03 || x_be || 00000000 plus padding is a single block built in registers
// from the limbs of x: ONE compression per row, no memory traffic.
#pragma once
#include <stdint.h>
#include "fe.cuh"

namespace csp {

#define CSP_K256_INIT { \
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, \
    0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, \
    0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da, \
    0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967, \
    0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85, \
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070, \
    0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3, \
    0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2 }
static const uint32_t h_k256[64] = CSP_K256_INIT;
#if defined(__CUDACC__)
static __constant__ uint32_t c_k256[64] = CSP_K256_INIT;
#endif
#if defined(__CUDA_ARCH__)
#define CSP_K256 c_k256
#else
#define CSP_K256 h_k256
#endif

CSP_HD uint32_t sha_ror(uint32_t x, int n) {
#if defined(__CUDA_ARCH__)
  return __funnelshift_r(x, x, n);
#else
  return (x >> n) | (x << (32 - n));
#endif
}

// st <- compress(st, w[0..15]); w is clobbered (rolling 16-word schedule)
CSP_HD void sha256_compress(uint32_t st[8], uint32_t w[16]) {
  uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
  for (int i = 0; i < 64; i++) {
    uint32_t wi;
    if (i < 16) {
      wi = w[i];
    } else {
      uint32_t w15 = w[(i - 15) & 15], w2 = w[(i - 2) & 15];
      uint32_t s0 = sha_ror(w15, 7) ^ sha_ror(w15, 18) ^ (w15 >> 3);
      uint32_t s1 = sha_ror(w2, 17) ^ sha_ror(w2, 19) ^ (w2 >> 10);
      wi = w[i & 15] + s0 + w[(i - 7) & 15] + s1;
      w[i & 15] = wi;
    }
    uint32_t t1 = h + (sha_ror(e, 6) ^ sha_ror(e, 11) ^ sha_ror(e, 25)) + ((e & f) ^ (~e & g)) + CSP_K256[i] + wi;
    uint32_t t2 = (sha_ror(a, 2) ^ sha_ror(a, 13) ^ sha_ror(a, 22)) + ((a & b) ^ (a & c) ^ (b & c));
    h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
  }
  st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

// t = int_BE( SHA256( TH || TH || (02|03) || x_be || 00000000 ) ) as field-style limbs
// (t.v[7] most significant).  x must be canonical; y_odd is the parity of the canonical y.
// midstate = chaining value after the TH||TH block.  k = 0 is hard-wired like the reference
// (src/cudasp_gpu.cu:278-282); the value is NOT reduced mod n (src/cudasp_gpu.cu:869-876).
// bip352_shared_secret_hash_k: the same with the output counter k of BIP-352 (ser_32(k), big-endian) -- only the
// opt-in match-details pass iterates k > 0.
CSP_HD void bip352_shared_secret_hash_k(fe &t, const fe &x, uint32_t y_odd, uint32_t k, const uint32_t midstate[8]) {
  uint32_t w[16];
  w[0] = ((0x02u + (y_odd & 1u)) << 24) | (x.v[7] >> 8);
#pragma unroll
  for (int j = 1; j < 8; j++) w[j] = (x.v[8 - j] << 24) | (x.v[7 - j] >> 8);
  w[8] = (x.v[0] << 24) | (k >> 8);        // last x byte, then the three leading bytes of k
  w[9] = (k << 24) | 0x00800000u;          // fourth byte of k, then the 0x80 pad
  w[10] = w[11] = w[12] = w[13] = w[14] = 0;
  w[15] = (64 + 37) * 8;    // 808 bits
  uint32_t st[8];
#pragma unroll
  for (int i = 0; i < 8; i++) st[i] = midstate[i];
  sha256_compress(st, w);
#pragma unroll
  for (int i = 0; i < 8; i++) t.v[i] = st[7 - i];
}
CSP_HD void bip352_shared_secret_hash(fe &t, const fe &x, uint32_t y_odd, const uint32_t midstate[8]) {
  bip352_shared_secret_hash_k(t, x, y_odd, 0u, midstate);
}

}  // namespace csp
