// fe29_bench.cu -- EXPERIMENT (not part of the product): a reduced-radix GF(p) multiplier / squarer, 9 limbs of 29 bits,
// measured against the product's 8 x 32-bit carry-chain field layer (csrc/fe.cuh) in the launch shape of K1
// (128 threads x 4 blocks per SM, one dependent chain per thread).
//
// Why: tools/imad_peak shows a plain IMAD.WIDE.U32 costs 2.2 dispatch cycles but one with a carry in/out 4.4 (DESIGN.md
// section 6).  With 29-bit limbs the 81 partial products accumulate in 64-bit columns without any carry; the price is a
// longer 522 -> 261 -> 256 bit fold.  This tool answers "how much faster is that, really" before anyone rewrites K1.
//
//   2^261 = 2^5 * (2^32 + 977) = 2^37 + 31264 (mod p)     (column k >= 9 folds into columns k-9, k-8)
//   a 64-bit column d = hi * 2^32 + lo is folded WITHOUT normalising it first:
//     lo * 2^(29k)      -> lo * 31264 at column k-9,  lo * 2^8  at column k-8
//     hi * 2^(29k + 32) -> hi * 8 * 31264 at column k-8,  hi * 2^11 at column k-7       (2^32 = 8 * 2^29)
//
// Output: one JSON object.  Every function is checked against fe_mul / fe_sqr on random and edge inputs first.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "../duckdb-cudasp-extension_b200/csrc/fe.cuh"

using namespace csp;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

struct f29 { uint32_t l[9]; };
// fold constants kept in constant memory: as immediates ptxas rewrites the power-of-two ones into shift + 64-bit add
// sequences (4 ALU instructions with carries instead of one plain IMAD.WIDE with a constant-bank operand)
__constant__ uint32_t K29[6] = {31264u, 256u, 250112u, 2048u, 977u, 8u};
static const uint32_t M29 = (1u << 29) - 1, M24 = (1u << 24) - 1;

// A 64-bit column is kept as two 32-bit registers.  (lo, hi) += x * y is written as a mad.lo.cc / madc.hi pair: ptxas fuses
// the pair into ONE plain IMAD.WIDE.U32 with accumulate, and -- unlike mad.wide.u32 on a 64-bit register -- it cannot
// re-associate the accumulation chain into fresh products + 3-input IADD3/IADD3.X trees (which it does otherwise:
// shorter dependency chains, but twice the instructions, and this code is dispatch-bound).
struct col { uint32_t lo, hi; };
__device__ __forceinline__ void madw(col &acc, uint32_t x, uint32_t y) {
  asm volatile("mad.lo.cc.u32 %0, %2, %3, %0;\n\tmadc.hi.u32 %1, %2, %3, %1;" : "+r"(acc.lo), "+r"(acc.hi) : "r"(x), "r"(y));
}
__device__ __forceinline__ col mulw(uint32_t x, uint32_t y) {
  col r;
  asm volatile("mul.lo.u32 %0, %2, %3;\n\tmul.hi.u32 %1, %2, %3;" : "=r"(r.lo), "=r"(r.hi) : "r"(x), "r"(y));
  return r;
}
// acc += (c >> 29)   (64-bit)
__device__ __forceinline__ void add_carry29(col &acc, const col &c) {
  const uint32_t cl = __funnelshift_r(c.lo, c.hi, 29), ch = c.hi >> 29;
  asm volatile("add.cc.u32 %0, %0, %2;\n\taddc.u32 %1, %1, %3;" : "+r"(acc.lo), "+r"(acc.hi) : "r"(cl), "r"(ch));
}

// 8 x 32 (any value < 2^256) -> 9 x 29
__device__ __forceinline__ void to29(f29 &r, const fe &a) {
#pragma unroll
  for (int k = 0; k < 9; k++) {
    const int bit = 29 * k, w = bit >> 5, s = bit & 31;
    uint32_t v = a.v[w] >> s;
    if (s + 29 > 32 && w < 7) v |= a.v[w + 1] << (32 - s);
    r.l[k] = v & M29;
  }
}
// 9 x 29 (limbs < 2^30, value < 2^257) -> 8 x 32 weakly reduced
__device__ __forceinline__ void from29(fe &r, const f29 &a) {
  uint64_t sum[10];
#pragma unroll
  for (int i = 0; i < 10; i++) sum[i] = 0;
#pragma unroll
  for (int k = 0; k < 9; k++) {
    const int bit = 29 * k, w = bit >> 5, s = bit & 31;
    const uint64_t v = (uint64_t)a.l[k] << s;
    sum[w] += (uint32_t)v;
    sum[w + 1] += v >> 32;
  }
#pragma unroll
  for (int i = 0; i < 9; i++) sum[i + 1] += sum[i] >> 32;
  const uint32_t w8 = (uint32_t)sum[8];                // bits >= 2^256 (a handful)
  uint64_t c = (uint64_t)(uint32_t)sum[0] + (uint64_t)w8 * 977u;
  r.v[0] = (uint32_t)c; c >>= 32;
  c += (uint64_t)(uint32_t)sum[1] + w8;
  r.v[1] = (uint32_t)c; c >>= 32;
#pragma unroll
  for (int i = 2; i < 8; i++) { c += (uint32_t)sum[i]; r.v[i] = (uint32_t)c; c >>= 32; }
  // a second wrap cannot happen for the values this tool produces; the comparison with fe_mul would show it
}

// fold columns 9..16 into 0..8, wrap at 2^256, carry-propagate: d[] -> 9 limbs (l0 < 2^29 + 2^22, l1 < 2^29 + 2^15,
// l2..l7 < 2^29, l8 < 2^24)
__device__ __forceinline__ void reduce29(f29 &r, col *d) {
#pragma unroll
  for (int k = 16; k >= 9; k--) {
    const uint32_t lo = d[k].lo, hi = d[k].hi;
    madw(d[k - 9], lo, K29[0]);
    madw(d[k - 8], lo, K29[1]);
    madw(d[k - 8], hi, K29[2]);
    madw(d[k - 7], hi, K29[3]);
  }
  // column 8 has weight 2^232: everything from its bit 24 up is >= 2^256
  {
    const uint32_t lo = d[8].lo, hi = d[8].hi;
    madw(d[0], hi, K29[2]);      // hi * 2^264 = hi * 2^8 * (2^32 + 977)
    madw(d[1], hi, K29[3]);
    const uint32_t q = lo >> 24;
    madw(d[0], q, K29[4]);
    madw(d[1], q, K29[5]);
    d[8].lo = lo & M24;
    d[8].hi = 0;
  }
#pragma unroll
  for (int k = 0; k < 8; k++) {
    r.l[k] = d[k].lo & M29;
    add_carry29(d[k + 1], d[k]);
  }
  const uint32_t q2 = __funnelshift_r(d[8].lo, d[8].hi, 24);          // < 2^12
  r.l[8] = d[8].lo & M24;
  r.l[0] += q2 * 977u;
  r.l[1] += q2 * 8u;
}

__device__ __forceinline__ void mul29(f29 &r, const f29 &a, const f29 &b) {
  col d[17];
#pragma unroll
  for (int i = 0; i < 9; i++) {
#pragma unroll
    for (int j = 0; j < 9; j++) {
      if (i == 0 || j == 8) d[i + j] = mulw(a.l[i], b.l[j]);
      else madw(d[i + j], a.l[i], b.l[j]);
    }
  }
  reduce29(r, d);
}

__device__ __forceinline__ void sqr29(f29 &r, const f29 &a) {
  col d[17];
  uint32_t a2[9];
#pragma unroll
  for (int i = 0; i < 9; i++) a2[i] = a.l[i] << 1;
#pragma unroll
  for (int i = 0; i < 9; i++) {
#pragma unroll
    for (int j = i; j < 9; j++) {
      const uint32_t y = (j == i) ? a.l[j] : a2[j];
      if (i == 0 || j == 8) d[i + j] = mulw(a.l[i], y);
      else madw(d[i + j], a.l[i], y);
    }
  }
  reduce29(r, d);
}

// ---------------------------------------------------------------- correctness: against the product's field layer
__global__ void k_check(const fe *a, const fe *b, fe *ref_mul, fe *got_mul, fe *ref_sqr, fe *got_sqr, fe *ref_chain,
                        fe *got_chain, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  fe x = a[i], y = b[i], r;
  f29 x9, y9, r9;
  to29(x9, x); to29(y9, y);
  fe_mul(r, x, y); fe_normalize(r); ref_mul[i] = r;
  mul29(r9, x9, y9); from29(r, r9); fe_normalize(r); got_mul[i] = r;
  fe_sqr(r, x); fe_normalize(r); ref_sqr[i] = r;
  sqr29(r9, x9); from29(r, r9); fe_normalize(r); got_sqr[i] = r;
  // a dependent chain without leaving the 29-bit form: x <- x*y, x <- x^2, 40 times
  fe cx = x;
  f29 c9 = x9;
  for (int t = 0; t < 40; t++) {
    fe_mul(cx, cx, y); fe_sqr(cx, cx);
    mul29(c9, c9, y9); sqr29(c9, c9);
  }
  fe_normalize(cx); ref_chain[i] = cx;
  from29(r, c9); fe_normalize(r); got_chain[i] = r;
}

// ---------------------------------------------------------------- timing: K1's launch shape, one dependent chain per thread
template <int MODE>
__global__ void __launch_bounds__(128, 4) k_time(fe *io, int iters) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  fe x = io[2 * i], y = io[2 * i + 1];
  if (MODE < 2) {
#pragma unroll 1
    for (int t = 0; t < iters; t++) {
      if (MODE == 0) fe_mul(x, x, y); else fe_sqr(x, x);
    }
  } else {
    f29 x9, y9;
    to29(x9, x); to29(y9, y);
#pragma unroll 1
    for (int t = 0; t < iters; t++) {
      if (MODE == 2) mul29(x9, x9, y9); else sqr29(x9, x9);
    }
    from29(x, x9);
  }
  io[2 * i] = x;
}

template <int MODE>
static int time_mode(const char *name, fe *d_io, int sms, int iters, double clock_hz, bool last) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  printf("  \"%s\": {", name);
  // blocks of 128 threads per SM: 4 is K1's occupancy (128 registers); more shows how much of the cost is exposed latency
  const int bps[4] = {2, 4, 8, 12};
  for (int b = 0; b < 4; b++) {
    const int blocks = sms * bps[b];
    k_time<MODE><<<blocks, 128>>>(d_io, iters);
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
      CK(cudaEventRecord(e0));
      k_time<MODE><<<blocks, 128>>>(d_io, iters);
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      if (ms < best) best = ms;
    }
    // bps warps per SM sub-partition run their chains side by side: scheduler cycles per operation = elapsed / (iters * bps)
    const double cyc = best * 1e-3 * clock_hz / ((double)iters * bps[b]);
    printf("\"cycles_per_op_at_%d_warps_per_scheduler\": %.1f%s", bps[b], cyc, b == 3 ? "" : ", ");
  }
  printf("}%s\n", last ? "" : ",");
  return 0;
}

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint32_t rnd32() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return (uint32_t)(rng_state >> 16); }

int main() {
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  const int n = 4096;
  std::vector<fe> a(n), b(n);
  for (int i = 0; i < n; i++)
    for (int l = 0; l < 8; l++) { a[i].v[l] = rnd32(); b[i].v[l] = rnd32(); }
  // edges: 0, 1, p-1, p, 2^256-1, 2^255, all-ones limbs patterns
  const uint32_t P[8] = {0xFFFFFC2Fu, 0xFFFFFFFEu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
  fe edge[6];
  for (int l = 0; l < 8; l++) {
    edge[0].v[l] = 0; edge[1].v[l] = l == 0; edge[2].v[l] = P[l] - (l == 0); edge[3].v[l] = P[l];
    edge[4].v[l] = 0xFFFFFFFFu; edge[5].v[l] = l == 7 ? 0x80000000u : 0;
  }
  for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) { a[i * 6 + j] = edge[i]; b[i * 6 + j] = edge[j]; }
  fe *d_a, *d_b, *d_out[6];
  CK(cudaMalloc(&d_a, n * sizeof(fe))); CK(cudaMalloc(&d_b, n * sizeof(fe)));
  for (int k = 0; k < 6; k++) CK(cudaMalloc(&d_out[k], n * sizeof(fe)));
  CK(cudaMemcpy(d_a, a.data(), n * sizeof(fe), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_b, b.data(), n * sizeof(fe), cudaMemcpyHostToDevice));
  k_check<<<(n + 127) / 128, 128>>>(d_a, d_b, d_out[0], d_out[1], d_out[2], d_out[3], d_out[4], d_out[5], n);
  CK(cudaDeviceSynchronize());
  std::vector<fe> o[6];
  for (int k = 0; k < 6; k++) { o[k].resize(n); CK(cudaMemcpy(o[k].data(), d_out[k], n * sizeof(fe), cudaMemcpyDeviceToHost)); }
  int bad[3] = {0, 0, 0};
  for (int i = 0; i < n; i++)
    for (int k = 0; k < 3; k++)
      for (int l = 0; l < 8; l++)
        if (o[2 * k][i].v[l] != o[2 * k + 1][i].v[l]) { bad[k]++; break; }
  int clk_khz = 0;
  CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
  const double clock_hz = 1e3 * clk_khz;
  printf("{\n  \"gpu\": \"%s\", \"sm_clock_mhz\": %.0f, \"checked\": %d, \"mismatch_mul\": %d, \"mismatch_sqr\": %d, \"mismatch_chain\": %d,\n",
         prop.name, clock_hz / 1e6, n, bad[0], bad[1], bad[2]);
  const int blocks = prop.multiProcessorCount * 12, iters = 4000;
  fe *d_io;
  CK(cudaMalloc(&d_io, (size_t)blocks * 128 * 2 * sizeof(fe)));
  for (size_t off = 0; off < (size_t)blocks * 128 * 2; off += n)
    CK(cudaMemcpy(d_io + off, a.data(), std::min<size_t>(n, (size_t)blocks * 128 * 2 - off) * sizeof(fe), cudaMemcpyHostToDevice));
  if (time_mode<0>("fe_mul_8x32_carry_chain", d_io, prop.multiProcessorCount, iters, clock_hz, false)) return 1;
  if (time_mode<1>("fe_sqr_8x32_carry_chain", d_io, prop.multiProcessorCount, iters, clock_hz, false)) return 1;
  if (time_mode<2>("mul29_9x29_plain_wide", d_io, prop.multiProcessorCount, iters, clock_hz, false)) return 1;
  if (time_mode<3>("sqr29_9x29_plain_wide", d_io, prop.multiProcessorCount, iters, clock_hz, true)) return 1;
  printf("}\n");
  return (bad[0] | bad[1] | bad[2]) ? 2 : 0;
}
