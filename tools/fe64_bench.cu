// fe64_bench.cu -- EXPERIMENT (not part of the product): GF(p) multiplication / squaring whose 25 (15) limb products run on the
// FP64 pipe (DFMA) instead of the integer multiply pipe, measured against the product's 8 x 32-bit carry-chain field layer
// (csrc/fe.cuh) in K1's launch shape, alone AND interleaved with it.
//
// Why: ncu shows K1 bound by ONE pipe -- fmaheavy, which executes every IMAD/IMAD.WIDE, 79 % busy (profiles/r02a_ncu_summary.md)
// -- while the FP64 pipe is idle and tools/imad_peak shows a DFMA issued beside an IMAD.WIDE costs little.  On B200 (not on
// B300) FP64 runs at 64 DFMA / clk / SM.  The question this tool answers before anyone rewrites K1: how many scheduler cycles
// does a field multiplication cost when its partial products are DFMAs, and does a mix of both kinds beat either alone?
//
// Method (Emmart / Zheng / Weems, "Faster modular exponentiation using double precision floating point arithmetic on the
// GPU", ARITH 2018): limbs of 52 bits held as exact doubles; for a, b < 2^52
//     hi = fma_rz(a, b, 2^104)                    bits(hi) = 0x467<<52 | floor(ab / 2^52)
//     lo = fma_rz(a, b, (2^104 + 2^52) - hi)      bits(lo) = 0x433<<52 | (ab mod 2^52)         (both exact)
// so the 64-bit bit patterns are added as INTEGERS into column accumulators (the exponent fields sum to a constant that is
// subtracted once), 3 FP64-pipe instructions per partial product, no float <-> int conversion instruction anywhere.
// The 520-bit column sum is carried into 16 x 32-bit words and folded by the product's own chain_fold512.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "../duckdb-cudasp-extension_b200/csrc/fe.cuh"

using namespace csp;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

struct f52 { double l[5]; };           // value = sum l[k] * 2^(52 k), every l[k] an integer in [0, 2^52)

static const uint64_t M52 = (1ull << 52) - 1;
#define BIAS_LO 0x4330000000000000ull   /* bit pattern of 2^52  */
#define BIAS_HI 0x4670000000000000ull   /* bit pattern of 2^104 */

// 8 x 32 (any value < 2^256) -> 5 x 52 as doubles: OR the 52 bits under the exponent of 2^52, subtract 2^52 (exact)
__device__ __forceinline__ void to52(f52 &r, const fe &a) {
#pragma unroll
  for (int k = 0; k < 5; k++) {
    const int bit = 52 * k, w = bit >> 5, s = bit & 31;      // s is 0, 20, 8, 28, 16
    uint32_t lo, hi;
    if (s == 0) { lo = a.v[w]; hi = w + 1 < 8 ? a.v[w + 1] : 0; }
    else {
      lo = __funnelshift_r(a.v[w], w + 1 < 8 ? a.v[w + 1] : 0, s);
      hi = __funnelshift_r(w + 1 < 8 ? a.v[w + 1] : 0, w + 2 < 8 ? a.v[w + 2] : 0, s);
    }
    hi = (hi & 0xFFFFFu) | 0x43300000u;
    r.l[k] = __dsub_rn(__hiloint2double((int)hi, (int)lo), 4503599627370496.0);
  }
}

// one partial product into two 64-bit column accumulators (bit patterns, see the header)
__device__ __forceinline__ void term(uint64_t &c_lo, uint64_t &c_hi, double a, double b) {
  const double hi = __fma_rz(a, b, 20282409603651670423947251286016.0);                       // 2^104
  const double lo = __fma_rz(a, b, __dsub_rn(20282409603651674927546878656512.0, hi));        // 2^104 + 2^52
  c_lo += (uint64_t)__double_as_longlong(lo);
  c_hi += (uint64_t)__double_as_longlong(hi);
}

// columns (52-bit spacing, exponent biases still inside) -> 16 x 32-bit words
template <class NLO, class NHI>
__device__ __forceinline__ void columns_to_words(uint32_t *t, uint64_t *col, NLO nlo, NHI nhi) {
  uint64_t L[10], carry = 0;
#pragma unroll
  for (int k = 0; k < 10; k++) {
    const uint64_t v = col[k] - (uint64_t)nlo(k) * BIAS_LO - (uint64_t)nhi(k) * BIAS_HI + carry;
    L[k] = v & M52;
    carry = v >> 52;
  }
#pragma unroll
  for (int w = 0; w < 16; w++) {
    const int bit = 32 * w, k = bit / 52, s = bit % 52;
    uint64_t x = L[k] >> s;
    if (s + 32 > 52 && k + 1 < 10) x |= L[k + 1] << (52 - s);
    t[w] = (uint32_t)x;
  }
}

__device__ __forceinline__ void mul52_wide(uint32_t *t, const f52 &a, const f52 &b) {
  uint64_t col[10];
#pragma unroll
  for (int k = 0; k < 10; k++) col[k] = 0;
#pragma unroll
  for (int i = 0; i < 5; i++)
#pragma unroll
    for (int j = 0; j < 5; j++) term(col[i + j], col[i + j + 1], a.l[i], b.l[j]);
  columns_to_words(t, col, [](int k) { return k < 5 ? k + 1 : 9 - k; }, [](int k) { return k == 0 ? 0 : (k <= 5 ? k : 10 - k); });
}
__device__ __forceinline__ void sqr52_wide(uint32_t *t, const f52 &a) {
  uint64_t col[10], cx[10];
#pragma unroll
  for (int k = 0; k < 10; k++) col[k] = cx[k] = 0;
#pragma unroll
  for (int i = 0; i < 5; i++) {
    term(col[2 * i], col[2 * i + 1], a.l[i], a.l[i]);
#pragma unroll
    for (int j = i + 1; j < 5; j++) term(cx[i + j], cx[i + j + 1], a.l[i], a.l[j]);
  }
#pragma unroll
  for (int k = 0; k < 10; k++) col[k] += cx[k] << 1;
  // lo terms in column k: diagonal (k even, k <= 8) + 2 x cross pairs i < j, i + j = k; hi terms: the same of column k - 1
  auto nlo = [](int k) { int n = 0; for (int i = 0; i < 5; i++) for (int j = i; j < 5; j++) if (i + j == k) n += (i == j) ? 1 : 2; return n; };
  auto nhi = [](int k) { int n = 0; for (int i = 0; i < 5; i++) for (int j = i; j < 5; j++) if (i + j + 1 == k) n += (i == j) ? 1 : 2; return n; };
  columns_to_words(t, col, nlo, nhi);
}

#if defined(__CUDA_ARCH__)
#define FOLD512 chain_fold512
#else
#define FOLD512 fe_fold512          /* host pass of nvcc only parses these device functions */
#endif
// drop-in shapes: 8 x 32 in, 8 x 32 out (weakly reduced, like fe_mul / fe_sqr)
__device__ __forceinline__ void fe_mul_f64(fe &r, const fe &a, const fe &b) {
  f52 x, y;
  to52(x, a); to52(y, b);
  uint32_t t[16];
  mul52_wide(t, x, y);
  FOLD512(r, t);
}
__device__ __forceinline__ void fe_sqr_f64(fe &r, const fe &a) {
  f52 x;
  to52(x, a);
  uint32_t t[16];
  sqr52_wide(t, x);
  FOLD512(r, t);
}

// ---------------------------------------------------------------- correctness: against the product's field layer
__global__ void k_check(const fe *a, const fe *b, fe *ref_mul, fe *got_mul, fe *ref_sqr, fe *got_sqr, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  fe x = a[i], y = b[i], r;
  fe_mul(r, x, y); fe_normalize(r); ref_mul[i] = r;
  fe_mul_f64(r, x, y); fe_normalize(r); got_mul[i] = r;
  fe_sqr(r, x); fe_normalize(r); ref_sqr[i] = r;
  fe_sqr_f64(r, x); fe_normalize(r); got_sqr[i] = r;
}

// ---------------------------------------------------------------- timing: K1's launch shape
// MODE: NI integer-pipe chains and NF FP64-pipe chains per thread, all independent of one another (x_c <- x_c * y or x_c^2)
template <int NI, int NF, bool SQR>
__global__ void __launch_bounds__(128, 4) k_time(fe *io, int iters) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  fe y = io[2 * i + 1];
  fe x[NI + NF];
#pragma unroll
  for (int c = 0; c < NI + NF; c++) { x[c] = io[2 * i]; x[c].v[0] += c; }
#pragma unroll 1
  for (int t = 0; t < iters; t++) {
#pragma unroll
    for (int c = 0; c < NI + NF; c++) {
      if (c < NI) { if (SQR) fe_sqr(x[c], x[c]); else fe_mul(x[c], x[c], y); }
      else { if (SQR) fe_sqr_f64(x[c], x[c]); else fe_mul_f64(x[c], x[c], y); }
    }
  }
  fe acc = x[0];
#pragma unroll
  for (int c = 1; c < NI + NF; c++) fe_add(acc, acc, x[c]);
  io[2 * i] = acc;
}

template <int NI, int NF, bool SQR>
static int time_mode(const char *name, fe *d_io, int sms, int iters, double clock_hz, bool last) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  cudaFuncAttributes at;
  CK(cudaFuncGetAttributes(&at, k_time<NI, NF, SQR>));
  printf("  \"%s\": {\"registers\": %d, \"local_bytes\": %d, ", name, at.numRegs, (int)at.localSizeBytes);
  const int bps[3] = {2, 4, 8};
  for (int b = 0; b < 3; b++) {
    const int blocks = sms * bps[b];
    k_time<NI, NF, SQR><<<blocks, 128>>>(d_io, iters);
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
      CK(cudaEventRecord(e0));
      k_time<NI, NF, SQR><<<blocks, 128>>>(d_io, iters);
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      if (ms < best) best = ms;
    }
    // scheduler cycles per field operation = elapsed / (iters * chains * warps per scheduler)
    const double cyc = best * 1e-3 * clock_hz / ((double)iters * (NI + NF) * bps[b]);
    printf("\"cycles_per_op_at_%d_warps_per_scheduler\": %.1f%s", bps[b], cyc, b == 2 ? "" : ", ");
  }
  printf("}%s\n", last ? "" : ",");
  return 0;
}

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint32_t rnd32() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return (uint32_t)(rng_state >> 16); }

int main() {
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  const int n = 8192;
  std::vector<fe> a(n), b(n);
  for (int i = 0; i < n; i++)
    for (int l = 0; l < 8; l++) { a[i].v[l] = rnd32(); b[i].v[l] = rnd32(); }
  // edges: 0, 1, p-1, p, 2^256-1, 2^255, limbs of all ones at the 52-bit boundaries
  const uint32_t P[8] = {0xFFFFFC2Fu, 0xFFFFFFFEu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
  fe edge[8];
  for (int l = 0; l < 8; l++) {
    edge[0].v[l] = 0; edge[1].v[l] = l == 0; edge[2].v[l] = P[l] - (l == 0); edge[3].v[l] = P[l];
    edge[4].v[l] = 0xFFFFFFFFu; edge[5].v[l] = l == 7 ? 0x80000000u : 0;
    edge[6].v[l] = (l & 1) ? 0xFFFFFFFFu : 0; edge[7].v[l] = l == 1 ? 0x000FFFFFu : (l == 0 ? 0xFFFFFFFFu : 0);
  }
  for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) { a[i * 8 + j] = edge[i]; b[i * 8 + j] = edge[j]; }
  fe *d_a, *d_b, *d_out[4];
  CK(cudaMalloc(&d_a, n * sizeof(fe))); CK(cudaMalloc(&d_b, n * sizeof(fe)));
  for (int k = 0; k < 4; k++) CK(cudaMalloc(&d_out[k], n * sizeof(fe)));
  CK(cudaMemcpy(d_a, a.data(), n * sizeof(fe), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_b, b.data(), n * sizeof(fe), cudaMemcpyHostToDevice));
  k_check<<<(n + 127) / 128, 128>>>(d_a, d_b, d_out[0], d_out[1], d_out[2], d_out[3], n);
  CK(cudaDeviceSynchronize());
  std::vector<fe> o[4];
  for (int k = 0; k < 4; k++) { o[k].resize(n); CK(cudaMemcpy(o[k].data(), d_out[k], n * sizeof(fe), cudaMemcpyDeviceToHost)); }
  int bad[2] = {0, 0};
  for (int i = 0; i < n; i++)
    for (int k = 0; k < 2; k++)
      for (int l = 0; l < 8; l++)
        if (o[2 * k][i].v[l] != o[2 * k + 1][i].v[l]) { bad[k]++; break; }
  int clk_khz = 0;
  CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
  const double clock_hz = 1e3 * clk_khz;
  printf("{\n  \"gpu\": \"%s\", \"sm_clock_mhz\": %.0f, \"checked\": %d, \"mismatch_mul\": %d, \"mismatch_sqr\": %d,\n", prop.name,
         clock_hz / 1e6, n, bad[0], bad[1]);
  const int sms = prop.multiProcessorCount, blocks = sms * 8, iters = 2000;
  fe *d_io;
  CK(cudaMalloc(&d_io, (size_t)blocks * 128 * 2 * sizeof(fe)));
  for (size_t off = 0; off < (size_t)blocks * 128 * 2; off += n)
    CK(cudaMemcpy(d_io + off, a.data(), std::min<size_t>(n, (size_t)blocks * 128 * 2 - off) * sizeof(fe), cudaMemcpyHostToDevice));
  if (time_mode<1, 0, false>("mul_int_x1", d_io, sms, iters, clock_hz, false)) return 1;
  if (time_mode<0, 1, false>("mul_f64_x1", d_io, sms, iters, clock_hz, false)) return 1;
  if (time_mode<2, 0, false>("mul_int_x2", d_io, sms, iters, clock_hz, false)) return 1;
  if (time_mode<1, 1, false>("mul_int_x1_f64_x1", d_io, sms, iters, clock_hz, false)) return 1;
  if (time_mode<0, 2, false>("mul_f64_x2", d_io, sms, iters, clock_hz, false)) return 1;
  if (time_mode<2, 1, false>("mul_int_x2_f64_x1", d_io, sms, iters, clock_hz, false)) return 1;
  if (time_mode<1, 0, true>("sqr_int_x1", d_io, sms, iters, clock_hz, false)) return 1;
  if (time_mode<0, 1, true>("sqr_f64_x1", d_io, sms, iters, clock_hz, false)) return 1;
  if (time_mode<1, 1, true>("sqr_int_x1_f64_x1", d_io, sms, iters, clock_hz, true)) return 1;
  printf("}\n");
  return (bad[0] || bad[1]) ? 2 : 0;
}
