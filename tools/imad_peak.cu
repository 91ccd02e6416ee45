// Integer-multiply issue-rate microbenchmark for B200 (sm_100a).
//
// Measures, per SM per clock, the sustained lane throughput of the instruction
// classes the secp256k1 field arithmetic is built from, so that bench.py's
// roofline denominator ("peak IMAD32 lane-ops/s") is a
// MEASURED number and not the programming-guide figure:
//   imad_lo    mad.lo.u32           (IMAD, 32-bit result)
//   imad_hi    mad.hi.u32           (IMAD.HI)
//   imad_wide  mad.wide.u32         (IMAD.WIDE.U32, 64-bit result)
//   carry      mad.lo.cc/madc.hi.cc (the carry-chain pattern a 256-bit product uses)
//   dfma       fma.rn.f64
//   iadd3      add.u32 chains       (ALU pipe, for dual-issue headroom)
//   wide+iadd  IMAD.WIDE and IADD3 interleaved
//   wide+dfma  IMAD.WIDE and DFMA interleaved
// Output: one JSON object on stdout.  Build: see tools/Makefile target imad_peak.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

constexpr int ITERS = 4096;   // loop trips
constexpr int CHAINS = 8;     // independent dependency chains per thread

template <int MODE>
__global__ void __launch_bounds__(256) k_bench(uint32_t *out, uint32_t seed, long long *cycles) {
  uint32_t a = seed + threadIdx.x, b = seed * 3 + blockIdx.x;
  uint32_t r[CHAINS * 2];
  double d[CHAINS];
#pragma unroll
  for (int i = 0; i < CHAINS * 2; i++) r[i] = a + i;
#pragma unroll
  for (int i = 0; i < CHAINS; i++) d[i] = (double)(a + i);
  double da = (double)a * 1e-9, db = (double)b * 1e-9;
  uint32_t mm[8];
#pragma unroll
  for (int i = 0; i < 8; i++) mm[i] = a * (2 * i + 3) + b;
  long long t0 = clock64();
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < CHAINS; i++) {
      if (MODE == 0) {
        asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(r[i]) : "r"(a), "r"(b));
        asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(r[i + CHAINS]) : "r"(b), "r"(a));
      } else if (MODE == 1) {
        asm volatile("mad.hi.u32 %0, %1, %2, %0;" : "+r"(r[i]) : "r"(a), "r"(r[i]));
        asm volatile("mad.hi.u32 %0, %1, %2, %0;" : "+r"(r[i + CHAINS]) : "r"(b), "r"(r[i + CHAINS]));
      } else if (MODE == 2) {
        // the multiplier is the low word of the NEIGHBOURING chain's accumulator: with loop-invariant multipliers ptxas
        // computes a*b once and turns the loop into 64-bit additions (that is what the r01 figures of this mode measured)
        unsigned long long acc = ((unsigned long long)r[2 * i + 1] << 32) | r[2 * i];
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc) : "r"(r[(2 * i + 2) & 15]), "r"(b));
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc) : "r"(r[(2 * i + 4) & 15]), "r"(a));
        r[2 * i] = (uint32_t)acc; r[2 * i + 1] = (uint32_t)(acc >> 32);
      } else if (MODE == 3) {
        // a 4-word carry chain of two products, as in one row of a 256-bit operand-scanning product
        if ((i & 1) == 0)
          asm volatile("mad.lo.cc.u32 %0, %4, %5, %0;\n\tmadc.hi.cc.u32 %1, %4, %5, %1;\n\t"
                       "madc.lo.cc.u32 %2, %5, %4, %2;\n\tmadc.hi.u32 %3, %5, %4, %3;"
                       : "+r"(r[2 * i]), "+r"(r[2 * i + 1]), "+r"(r[2 * i + 2]), "+r"(r[2 * i + 3]) : "r"(a), "r"(b));
      } else if (MODE == 4) {
        asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(d[i]) : "d"(da), "d"(db));
        asm volatile("fma.rn.f64 %0, %2, %1, %0;" : "+d"(d[i]) : "d"(da), "d"(db));
      } else if (MODE == 5) {
        asm volatile("add.u32 %0, %0, %1;" : "+r"(r[i]) : "r"(a));
        asm volatile("add.u32 %0, %0, %1;" : "+r"(r[i + CHAINS]) : "r"(b));
      } else if (MODE == 6) {
        unsigned long long acc = ((unsigned long long)r[2 * i + 1] << 32) | r[2 * i];
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc) : "r"(r[(2 * i + 2) & 15]), "r"(b));
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc) : "r"(r[(2 * i + 4) & 15]), "r"(a));
        r[2 * i] = (uint32_t)acc; r[2 * i + 1] = (uint32_t)(acc >> 32);
        uint32_t x = (uint32_t)__double_as_longlong(d[i]);
        asm volatile("add.u32 %0, %0, %1;" : "+r"(x) : "r"(a));
        asm volatile("xor.b32 %0, %0, %1;" : "+r"(x) : "r"(b));
        d[i] = __longlong_as_double((long long)x);
      } else if (MODE == 8) {
        // 64-bit accumulate with carry-OUT captured (IMAD.WIDE Rd, P, ...) + IADD3.X into a third word
        if ((i & 1) == 0)
          asm volatile("mad.lo.cc.u32 %0, %3, %4, %0;\n\tmadc.hi.cc.u32 %1, %3, %4, %1;\n\taddc.u32 %2, %2, 0;"
                       : "+r"(r[2 * i]), "+r"(r[2 * i + 1]), "+r"(r[2 * i + 2]) : "r"(a), "r"(b));
      } else if (MODE == 9) {
        // full 8-limb row:  acc[0..8] += {a-limbs} * b, even columns only (4 products in one carry chain)
        if (i == 0) {
          asm volatile("mad.lo.cc.u32 %0, %9, %10, %0;\n\tmadc.hi.cc.u32 %1, %9, %10, %1;\n\t"
                       "madc.lo.cc.u32 %2, %10, %9, %2;\n\tmadc.hi.cc.u32 %3, %10, %9, %3;\n\t"
                       "madc.lo.cc.u32 %4, %9, %9, %4;\n\tmadc.hi.cc.u32 %5, %9, %9, %5;\n\t"
                       "madc.lo.cc.u32 %6, %10, %10, %6;\n\tmadc.hi.cc.u32 %7, %10, %10, %7;\n\t"
                       "addc.u32 %8, %8, 0;"
                       : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8])
                       : "r"(a), "r"(b));
        } else if (i == 1) {
          uint32_t t0 = (uint32_t)__double_as_longlong(d[0]), t1 = (uint32_t)__double_as_longlong(d[1]),
                   t2 = (uint32_t)__double_as_longlong(d[2]), t3 = (uint32_t)__double_as_longlong(d[3]),
                   t4 = (uint32_t)__double_as_longlong(d[4]), t5 = (uint32_t)__double_as_longlong(d[5]),
                   t6 = (uint32_t)__double_as_longlong(d[6]), t7 = (uint32_t)__double_as_longlong(d[7]);
          asm volatile("mad.lo.cc.u32 %0, %9, %10, %0;\n\tmadc.hi.cc.u32 %1, %9, %10, %1;\n\t"
                       "madc.lo.cc.u32 %2, %10, %9, %2;\n\tmadc.hi.cc.u32 %3, %10, %9, %3;\n\t"
                       "madc.lo.cc.u32 %4, %9, %9, %4;\n\tmadc.hi.cc.u32 %5, %9, %9, %5;\n\t"
                       "madc.lo.cc.u32 %6, %10, %10, %6;\n\tmadc.hi.cc.u32 %7, %10, %10, %7;\n\t"
                       "addc.u32 %8, %8, 0;"
                       : "+r"(t0), "+r"(t1), "+r"(t2), "+r"(t3), "+r"(t4), "+r"(t5), "+r"(t6), "+r"(t7), "+r"(r[9])
                       : "r"(a), "r"(b));
          d[0] = __longlong_as_double(t0); d[1] = __longlong_as_double(t1); d[2] = __longlong_as_double(t2); d[3] = __longlong_as_double(t3);
          d[4] = __longlong_as_double(t4); d[5] = __longlong_as_double(t5); d[6] = __longlong_as_double(t6); d[7] = __longlong_as_double(t7);
        }
      } else if (MODE == 10) {
        // unfused: separate IMAD (lo) and IMAD.HI with carries (different operands so ptxas cannot pair them)
        if ((i & 1) == 0)
          asm volatile("mad.lo.cc.u32 %0, %3, %4, %0;\n\tmadc.hi.cc.u32 %1, %4, %4, %1;\n\taddc.u32 %2, %2, 0;"
                       : "+r"(r[2 * i]), "+r"(r[2 * i + 1]), "+r"(r[2 * i + 2]) : "r"(a), "r"(b));
      } else if (MODE == 11) {
        // carry-IN only: one IADD3 produces the carry, the IMAD.WIDE.X consumes it and produces none
        if ((i & 1) == 0)
          asm volatile("add.cc.u32 %2, %2, %4;\n\tmadc.lo.cc.u32 %0, %3, %4, %0;\n\tmadc.hi.u32 %1, %3, %4, %1;"
                       : "+r"(r[2 * i]), "+r"(r[2 * i + 1]), "+r"(r[2 * i + 2]) : "r"(a), "r"(b));
      } else if (MODE == 12) {
        // carry IN and OUT: IADD3 (carry out) -> IMAD.WIDE.X (in, out) -> IADD3.X (in)
        if ((i & 1) == 0)
          asm volatile("add.cc.u32 %2, %2, %5;\n\tmadc.lo.cc.u32 %0, %4, %5, %0;\n\tmadc.hi.cc.u32 %1, %4, %5, %1;\n\taddc.u32 %3, %3, 0;"
                       : "+r"(r[2 * i]), "+r"(r[2 * i + 1]), "+r"(r[2 * i + 2]), "+r"(r[2 * i + 3]) : "r"(a), "r"(b));
      } else if (MODE == 13) {
        // reference for 11/12: the two IADD3 alone (add.cc + addc)
        if ((i & 1) == 0)
          asm volatile("add.cc.u32 %0, %0, %3;\n\taddc.u32 %1, %1, %2;"
                       : "+r"(r[2 * i]), "+r"(r[2 * i + 1]) : "r"(a), "r"(b));
      } else if (MODE == 14) {
        // plain IMAD.WIDE again, but every instruction reads DIFFERENT multiplier registers (no operand-reuse cache
        // hits, as in a real 8x8 or 9x9 limb product; mode 2 multiplies the same two registers over and over)
        unsigned long long acc = ((unsigned long long)r[2 * i + 1] << 32) | r[2 * i];
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc) : "r"(r[(2 * i + 2) & 15]), "r"(mm[(i + 3) & 7]));
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc) : "r"(r[(2 * i + 6) & 15]), "r"(mm[(i + 6) & 7]));
        r[2 * i] = (uint32_t)acc; r[2 * i + 1] = (uint32_t)(acc >> 32);
      } else if (MODE == 7) {
        unsigned long long acc = ((unsigned long long)r[2 * i + 1] << 32) | r[2 * i];
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc) : "r"(r[(2 * i + 2) & 15]), "r"(b));
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc) : "r"(r[(2 * i + 4) & 15]), "r"(a));
        r[2 * i] = (uint32_t)acc; r[2 * i + 1] = (uint32_t)(acc >> 32);
        asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(d[i]) : "d"(da), "d"(db));
        asm volatile("fma.rn.f64 %0, %2, %1, %0;" : "+d"(d[i]) : "d"(da), "d"(db));
      }
    }
  }
  long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < CHAINS * 2; i++) s ^= r[i];
#pragma unroll
  for (int i = 0; i < CHAINS; i++) s ^= (uint32_t)__double_as_longlong(d[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

struct Res { double per_clk_sm; double per_s; double ms; double mhz; };

template <int MODE>
static int run(const char *name, int ops_per_inner, int sms, int blocks_per_sm, Res *res, bool last) {
  int blocks = sms * blocks_per_sm;
  uint32_t *out; long long *cyc;
  CK(cudaMalloc(&out, (size_t)blocks * 256 * 4));
  CK(cudaMalloc(&cyc, (size_t)blocks * 8));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for (int w = 0; w < 3; w++) k_bench<MODE><<<blocks, 256>>>(out, 12345u + w, cyc);
  CK(cudaDeviceSynchronize());
  const int reps = 20;
  CK(cudaEventRecord(e0));
  for (int w = 0; w < reps; w++) k_bench<MODE><<<blocks, 256>>>(out, 777u + w, cyc);
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
  ms /= reps;
  long long *h = new long long[blocks];
  CK(cudaMemcpy(h, cyc, (size_t)blocks * 8, cudaMemcpyDeviceToHost));
  double mean_cyc = 0; for (int i = 0; i < blocks; i++) mean_cyc += (double)h[i]; mean_cyc /= blocks;
  delete[] h;
  double lane_ops = (double)blocks * 256 * ITERS * CHAINS * ops_per_inner;
  // per clock per SM from in-kernel cycle counters (blocks_per_sm CTAs resident together)
  double per_clk_sm = (double)blocks_per_sm * 256 * ITERS * CHAINS * ops_per_inner / mean_cyc;
  res->per_clk_sm = per_clk_sm; res->per_s = lane_ops / (ms * 1e-3); res->ms = ms;
  res->mhz = res->per_s / per_clk_sm / sms / 1e6;
  printf("  \"%s\": {\"lane_ops_per_clk_per_sm\": %.2f, \"lane_ops_per_s\": %.4e, \"ms\": %.4f, \"implied_sm_mhz\": %.0f}%s\n",
         name, per_clk_sm, res->per_s, ms, res->mhz, last ? "" : ",");
  CK(cudaFree(out)); CK(cudaFree(cyc));
  return 0;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int sms = p.multiProcessorCount;
  printf("{\n  \"gpu\": \"%s\", \"sms\": %d, \"blocks_per_sm\": 4, \"threads_per_block\": 256,\n", p.name, sms);
  Res r;
  // ops_per_inner = number of counted instructions per chain per trip
  if (run<0>("imad_lo", 2, sms, 4, &r, false)) return 1;
  if (run<1>("imad_hi", 2, sms, 4, &r, false)) return 1;
  if (run<2>("imad_wide", 2, sms, 4, &r, false)) return 1;
  if (run<3>("carry_chain_products", 1, sms, 4, &r, false)) return 1;  // 2 full 32x32->64 products
  if (run<4>("dfma", 2, sms, 4, &r, false)) return 1;
  if (run<5>("iadd", 2, sms, 4, &r, false)) return 1;
  if (run<6>("imad_wide_with_2alu_each", 2, sms, 4, &r, false)) return 1;       // counts the IMAD.WIDEs only
  if (run<7>("imad_wide_with_dfma_each", 2, sms, 4, &r, false)) return 1;       // counts the IMAD.WIDEs only
  if (run<8>("wide_carry_out_plus_addc", 1, sms, 4, &r, false)) return 1;       // CHAINS/2 per trip, 1 product each... counted x2 below
  if (run<9>("row4_carry_chain_x2", 1, sms, 4, &r, false)) return 1;            // 8 products per trip
  if (run<10>("unfused_lo_hi", 1, sms, 4, &r, false)) return 1;
  // the next three execute CHAINS/2 groups per trip but count CHAINS: divide the figure by 2 for groups per clock
  if (run<11>("addcc_then_wide_carry_in", 1, sms, 4, &r, false)) return 1;      // group = IADD3 + IMAD.WIDE.X(in)
  if (run<12>("addcc_wide_in_out_addc", 1, sms, 4, &r, false)) return 1;        // group = IADD3 + IMAD.WIDE.X(in,out) + IADD3.X
  if (run<13>("addcc_addc_pair", 1, sms, 4, &r, false)) return 1;               // group = IADD3 + IADD3.X
  if (run<14>("imad_wide_distinct_operands", 2, sms, 4, &r, true)) return 1;
  printf("}\n");
  return 0;
}
