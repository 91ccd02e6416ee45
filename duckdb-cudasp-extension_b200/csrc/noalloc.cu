// noalloc.cu -- cudasp_batch_inverse_static: a batch inversion in GF(p) that allocates NOTHING.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is a variant of K2/K4/K6
// (kernels.cuh k_batch_inv, Montgomery's trick: prefix products -> ONE Fermat inversion per thread -> back pass), whose
// scratch -- one prefix product per element -- normally lives in a cudaMalloc'ed workspace of the staging slot.  Here the
// scratch is a file-scope static __device__ array of CUDASP_STATIC_INV_MAX elements (2 MB), which every GPU that loads the
// library holds until the process exits.  The caller passes input and output device pointers.
//
// In host code the array's NAME is a host-side stand-in, not a device address (a kernel given it would fault or, where the
// GPU can reach host memory, silently use the wrong memory): the address is taken once per device, with that device
// current, by cudaGetSymbolAddress in set_up() below, and that pointer is what the kernel receives.
//
// Every call on a device shares the one array; a call leaves its prefix products behind.  Nothing has to start at zero
// (the forward pass writes every scratch word before the back pass reads it), so nothing is cleared -- but two calls that
// overlap on the GPU (different streams of one device, no dependency between them) race on the scratch: serialise them.
#include <cuda_runtime.h>
#include <functional>
#include <mutex>

#include "../../include/cudasp.h"
#include "rowfn.cuh"
#include "staging_internal.hpp"

namespace csp {
namespace {

const int NOALLOC_MAX_DEVICES = 64;
const uint32_t PER_THREAD = 16;      // elements per Fermat inversion, as in the pipeline's k_batch_inv launches

static_assert(sizeof(fe) == 32, "field element layout");
static __device__ __align__(256) fe g_inv_scratch[CUDASP_STATIC_INV_MAX];

// out[i] = in[i]^-1 mod p, canonical (0 -> 0), for i = tid, tid + T, ...; scratch[i] holds the prefix products of the thread's chain
// (in and out may be the same array: no __restrict__, plain loads)
__global__ void __launch_bounds__(128) k_batch_inv_static(const fe *in, fe *out, fe *__restrict__ scratch, uint32_t n, uint32_t T) {
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= T || tid >= n) return;
  fe acc;
  fe_set_u32(acc, 1);
  uint32_t last = tid;
  for (uint64_t i = tid; i < n; i += T) {
    fe v;
    load_fe(v, in + i);
    store_fe(scratch + i, acc);
    if (!fe_is_zero(v)) fe_mul(acc, acc, v);
    last = (uint32_t)i;
  }
  fe inv;
  fe_inv(inv, acc);
  for (uint32_t i = last;; i -= T) {
    fe v, pre, r;
    load_fe(v, in + i);
    load_fe(pre, scratch + i);
    if (fe_is_zero(v)) {
      fe_set_zero(r);
    } else {
      fe_mul(r, inv, pre);
      fe_mul(inv, inv, v);
    }
    fe_normalize(r);
    store_fe(out + i, r);
    if (i == tid) break;
  }
}

struct PerDevice {
  std::once_flag once;
  fe *scratch = nullptr;          // DEVICE address of g_inv_scratch on this device
  cudaError_t err = cudaSuccess;
};
PerDevice g_dev[NOALLOC_MAX_DEVICES];

// once per device, with that device current
void set_up(PerDevice &d) {
  void *p = nullptr;
  d.err = cudaGetSymbolAddress(&p, g_inv_scratch);
  if (d.err == cudaSuccess) d.scratch = static_cast<fe *>(p);
}

}  // namespace
}  // namespace csp

extern "C" int cudasp_batch_inverse_static(int device, const uint8_t *d_in, uint8_t *d_out, uint32_t n, void *stream) {
  using namespace csp;
  if (n == 0) return CUDASP_OK;
  if (!d_in || !d_out) return staging::fail(nullptr, CUDASP_ERR_ARG, "cudasp_batch_inverse_static: NULL device pointer");
  if (n > CUDASP_STATIC_INV_MAX)
    return staging::fail(nullptr, CUDASP_ERR_ARG, "cudasp_batch_inverse_static: n exceeds CUDASP_STATIC_INV_MAX (65536 elements): split the call");
  if (device < 0 || device >= NOALLOC_MAX_DEVICES) return staging::fail(nullptr, CUDASP_ERR_ARG, "cudasp_batch_inverse_static: device ordinal out of range");
  if ((reinterpret_cast<uintptr_t>(d_in) | reinterpret_cast<uintptr_t>(d_out)) & 15u)
    return staging::fail(nullptr, CUDASP_ERR_ARG, "cudasp_batch_inverse_static: device pointers must be 16-byte aligned");
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return staging::fail(nullptr, CUDASP_ERR_CUDA, cudaGetErrorString(e));
  PerDevice &d = g_dev[device];
  std::call_once(d.once, set_up, std::ref(d));
  if (d.err != cudaSuccess || !d.scratch) return staging::fail(nullptr, CUDASP_ERR_CUDA, cudaGetErrorString(d.err));
  uint32_t T = (n + PER_THREAD - 1) / PER_THREAD;
  T = (T + 127u) / 128u * 128u;
  T = T < n ? T : n;
  k_batch_inv_static<<<(T + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(reinterpret_cast<const fe *>(d_in),
                                                                                    reinterpret_cast<fe *>(d_out), d.scratch, n, T);
  e = cudaGetLastError();
  if (e != cudaSuccess) return staging::fail(nullptr, CUDASP_ERR_CUDA, cudaGetErrorString(e));
  return CUDASP_OK;
}
