// Tuning aid: one kernel per field operation, so that `python tools/sass_hist.py tools/fe_op_sass.cubin` shows the exact
// instruction mix of a single fe_mul / fe_sqr / fe_add / fe_sub / linear combination (cross-compiled, no GPU needed):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -cubin -o tools/fe_op_sass.cubin tools/fe_op_sass.cu
#include "../duckdb-cudasp-extension_b200/csrc/rowfn.cuh"
using namespace csp;
#define K(name, body) \
  __global__ void name(const fe *a, const fe *b, fe *out) { fe x, y, r; load_fe(x, a + threadIdx.x); load_fe(y, b + threadIdx.x); body; store_fe(out + threadIdx.x, r); }
K(op_mul, fe_mul(r, x, y))
K(op_sqr, fe_sqr(r, x))
K(op_add, fe_add(r, x, y))
K(op_sub, fe_sub(r, x, y))
K(op_lin3, (fe_lin3<1, -1, -2>(r, x, y, x)))
K(op_lin2, (fe_lin2<1, -8>(r, x, y)))
K(op_nop, r = x)
