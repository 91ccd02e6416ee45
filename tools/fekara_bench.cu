// fekara_bench.cu -- EXPERIMENT (not part of the product): one level of subtractive Karatsuba for the 8 x 32-bit field
// multiplication (48 + 9 IMAD.WIDE instead of 64 + 9, at the price of ~70 more carry-chain ALU instructions), written in the
// same serial-carry-chain discipline as csrc/fe_chain.cuh, measured against fe_mul in K1's launch shape.
//   a = a0 + a1 2^128, b = b0 + b1 2^128:   a b = z0 + (z0 + z2 + (a0 - a1)(b1 - b0)) 2^128 + z2 2^256,  z0 = a0 b0, z2 = a1 b1
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "../duckdb-cudasp-extension_b200/csrc/fe.cuh"

using namespace csp;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

#if defined(__CUDA_ARCH__)
// 4 x 4 words -> 8 words, even/odd accumulators, rows serialised (see fe_chain.cuh chain_mul_wide); `dep` makes the first row
// wait for an earlier value (keeps the three sub-products from being interleaved: at most two carries alive)
__device__ __forceinline__ void mul4_wide(uint32_t *t, const uint32_t *x, const uint32_t *y, uint32_t dep) {
  uint32_t e[8], o[8];
#pragma unroll
  for (int k = 0; k < 8; k++) e[k] = o[k] = 0;
  {
    cc::CF f;
    const uint32_t y0 = after(y[0], dep);
    cc::mad_wide(f, e[0], e[1], x[0], y0);
    cc::mad_wide(f, e[2], e[3], x[2], y0);
    cc::mad_wide(f, o[0], o[1], x[1], y0);
    cc::mad_wide(f, o[2], o[3], x[3], y0);
  }
#pragma unroll
  for (int i = 1; i < 4; i++) {
    const uint32_t yi = after(y[i], e[i + 2]);
    cc::CF f;
    if (i & 1) {
      cc::mad_wide_cc(f, e[i + 1], e[i + 2], x[1], yi);
      cc::madc_wide(f, e[i + 3], e[i + 4], x[3], yi);              // top pair fresh
      cc::mad_wide_cc(f, o[i - 1], o[i], x[0], yi);
      cc::madc_wide_cc(f, o[i + 1], o[i + 2], x[2], yi);
      o[i + 3] = cc::addc(f, 0, 0);
    } else {
      cc::mad_wide_cc(f, e[i], e[i + 1], x[0], yi);
      cc::madc_wide_cc(f, e[i + 2], e[i + 3], x[2], yi);
      e[i + 4] = cc::addc(f, 0, 0);
      cc::mad_wide_cc(f, o[i], o[i + 1], x[1], yi);
      cc::madc_wide(f, o[i + 2], o[i + 3], x[3], yi);              // top pair fresh
    }
  }
  cc::CF f;
  t[0] = e[0];
  t[1] = cc::add_cc(f, e[1], o[0]);
#pragma unroll
  for (int k = 2; k < 7; k++) t[k] = cc::addc_cc(f, e[k], o[k - 1]);
  t[7] = cc::addc(f, e[7], o[6]);
}
// d = |x - y| (4 words), returns 1 when x < y
__device__ __forceinline__ uint32_t absdiff4(uint32_t *d, const uint32_t *x, const uint32_t *y) {
  cc::CF f;
  d[0] = cc::sub_cc(f, x[0], y[0]);
  d[1] = cc::subc_cc(f, x[1], y[1]);
  d[2] = cc::subc_cc(f, x[2], y[2]);
  d[3] = cc::subc_cc(f, x[3], y[3]);
  const uint32_t m = cc::subc(f, 0, 0);          // 0 or 0xffffffff
  d[0] = cc::sub_cc(f, d[0] ^ m, m);
  d[1] = cc::subc_cc(f, d[1] ^ m, m);
  d[2] = cc::subc_cc(f, d[2] ^ m, m);
  d[3] = cc::subc(f, d[3] ^ m, m);
  return m & 1u;
}
__device__ __forceinline__ void fe_mul_kara(fe &r, const fe &a, const fe &b) {
  uint32_t t[16], mid[8], da[4], db[4];
  mul4_wide(t, a.v, b.v, 0);                           // z0
  mul4_wide(t + 8, a.v + 4, b.v + 4, t[7]);            // z2
  const uint32_t sa = absdiff4(da, a.v, a.v + 4);      // a0 - a1
  const uint32_t sb = absdiff4(db, b.v + 4, b.v);      // b1 - b0
  mul4_wide(mid, da, db, t[15]);
  const uint32_t neg = 0u - (sa ^ sb);                 // all ones: the middle product counts negative
  // z1 = z0 + z2 +- mid  (9 words, non-negative)
  uint32_t z1[9];
  {
    cc::CF f;
    z1[0] = cc::add_cc(f, t[0], t[8]);
#pragma unroll
    for (int k = 1; k < 8; k++) z1[k] = cc::addc_cc(f, t[k], t[8 + k]);
    z1[8] = cc::addc(f, 0, 0);
  }
  {
    cc::CF f;
    (void)cc::add_cc(f, neg, neg);                     // carry = 1 when the middle product is subtracted (~mid + 1)
    z1[0] = cc::addc_cc(f, z1[0], mid[0] ^ neg);
#pragma unroll
    for (int k = 1; k < 8; k++) z1[k] = cc::addc_cc(f, z1[k], mid[k] ^ neg);
    z1[8] = cc::addc(f, z1[8], neg);
  }
  // t += z1 << 128
  {
    cc::CF f;
    t[4] = cc::add_cc(f, t[4], z1[0]);
#pragma unroll
    for (int k = 1; k < 9; k++) t[4 + k] = cc::addc_cc(f, t[4 + k], z1[k]);
    t[13] = cc::addc_cc(f, t[13], 0);
    t[14] = cc::addc_cc(f, t[14], 0);
    t[15] = cc::addc(f, t[15], 0);
  }
  chain_fold512(r, t);
}
#else
__device__ __forceinline__ void fe_mul_kara(fe &r, const fe &a, const fe &b) { fe_mul(r, a, b); }   // host pass of nvcc only parses this
#endif

__global__ void k_check(const fe *a, const fe *b, fe *ref, fe *got, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  fe x = a[i], y = b[i], r;
  fe_mul(r, x, y); fe_normalize(r); ref[i] = r;
  fe_mul_kara(r, x, y); fe_normalize(r); got[i] = r;
}

// MODE 0: x <- x * y with fe_mul; 1: with fe_mul_kara; 2: alternate mul / sqr like the ladder (fe_mul); 3: the same with fe_mul_kara
template <int MODE>
__global__ void __launch_bounds__(128, 4) k_time(fe *io, int iters) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  fe x = io[2 * i], y = io[2 * i + 1];
#pragma unroll 1
  for (int t = 0; t < iters; t++) {
    if (MODE == 0) fe_mul(x, x, y);
    if (MODE == 1) fe_mul_kara(x, x, y);
    if (MODE == 2) { fe_mul(x, x, y); fe_sqr(y, x); fe_sub(x, x, y); }
    if (MODE == 3) { fe_mul_kara(x, x, y); fe_sqr(y, x); fe_sub(x, x, y); }
  }
  io[2 * i] = x;
}

template <int MODE>
static int time_mode(const char *name, fe *d_io, int sms, int iters, double clock_hz, bool last) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  cudaFuncAttributes at;
  CK(cudaFuncGetAttributes(&at, k_time<MODE>));
  printf("  \"%s\": {\"registers\": %d, \"local_bytes\": %d, ", name, at.numRegs, (int)at.localSizeBytes);
  const int bps[3] = {2, 4, 8};
  for (int b = 0; b < 3; b++) {
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
    const double cyc = best * 1e-3 * clock_hz / ((double)iters * bps[b]);
    printf("\"cycles_per_iteration_at_%d_warps_per_scheduler\": %.1f%s", bps[b], cyc, b == 2 ? "" : ", ");
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
  const uint32_t P[8] = {0xFFFFFC2Fu, 0xFFFFFFFEu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
  fe edge[8];
  for (int l = 0; l < 8; l++) {
    edge[0].v[l] = 0; edge[1].v[l] = l == 0; edge[2].v[l] = P[l] - (l == 0); edge[3].v[l] = P[l];
    edge[4].v[l] = 0xFFFFFFFFu; edge[5].v[l] = l == 7 ? 0x80000000u : 0;
    edge[6].v[l] = l < 4 ? 0xFFFFFFFFu : 0;          // a0 > a1
    edge[7].v[l] = l < 4 ? 0 : 0xFFFFFFFFu;          // a0 < a1
  }
  for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) { a[i * 8 + j] = edge[i]; b[i * 8 + j] = edge[j]; }
  for (int i = 64; i < 128; i++) for (int l = 0; l < 4; l++) a[i].v[l + 4] = a[i].v[l];     // a0 == a1: zero difference
  fe *d_a, *d_b, *d_out[2];
  CK(cudaMalloc(&d_a, n * sizeof(fe))); CK(cudaMalloc(&d_b, n * sizeof(fe)));
  for (int k = 0; k < 2; k++) CK(cudaMalloc(&d_out[k], n * sizeof(fe)));
  CK(cudaMemcpy(d_a, a.data(), n * sizeof(fe), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_b, b.data(), n * sizeof(fe), cudaMemcpyHostToDevice));
  k_check<<<(n + 127) / 128, 128>>>(d_a, d_b, d_out[0], d_out[1], n);
  CK(cudaDeviceSynchronize());
  std::vector<fe> o[2];
  for (int k = 0; k < 2; k++) { o[k].resize(n); CK(cudaMemcpy(o[k].data(), d_out[k], n * sizeof(fe), cudaMemcpyDeviceToHost)); }
  int bad = 0;
  for (int i = 0; i < n; i++)
    for (int l = 0; l < 8; l++)
      if (o[0][i].v[l] != o[1][i].v[l]) { bad++; break; }
  int clk_khz = 0;
  CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
  const double clock_hz = 1e3 * clk_khz;
  printf("{\n  \"gpu\": \"%s\", \"sm_clock_mhz\": %.0f, \"checked\": %d, \"mismatch_mul\": %d,\n", prop.name, clock_hz / 1e6, n, bad);
  const int sms = prop.multiProcessorCount, blocks = sms * 8, iters = 2000;
  fe *d_io;
  CK(cudaMalloc(&d_io, (size_t)blocks * 128 * 2 * sizeof(fe)));
  for (size_t off = 0; off < (size_t)blocks * 128 * 2; off += n)
    CK(cudaMemcpy(d_io + off, a.data(), std::min<size_t>(n, (size_t)blocks * 128 * 2 - off) * sizeof(fe), cudaMemcpyHostToDevice));
  if (time_mode<0>("mul_schoolbook", d_io, sms, iters, clock_hz, false)) return 1;
  if (time_mode<1>("mul_karatsuba", d_io, sms, iters, clock_hz, false)) return 1;
  if (time_mode<2>("mul_sqr_sub_schoolbook", d_io, sms, iters, clock_hz, false)) return 1;
  if (time_mode<3>("mul_sqr_sub_karatsuba", d_io, sms, iters, clock_hz, true)) return 1;
  printf("}\n");
  return bad ? 2 : 0;
}
