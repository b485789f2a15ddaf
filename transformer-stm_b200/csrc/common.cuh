// Shared helpers for libcvt_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

namespace cvt {

// ---- error plumbing: every launcher returns a cudaError_t; the API layer turns it into cvt_status ----
void set_error(const std::string& msg);
extern thread_local long long g_launch_count;      // kernels launched by this thread (own kernels only)

#define CVT_CUDA(expr)                                                                      \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) {                                                                \
      ::cvt::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " @" + __FILE__ + ":" + std::to_string(__LINE__)); \
      return _e;                                                                            \
    }                                                                                       \
  } while (0)

#define CVT_LAUNCH_CHECK()                                                                  \
  do {                                                                                      \
    ::cvt::g_launch_count++;                                                                \
    cudaError_t _e = cudaGetLastError();                                                    \
    if (_e != cudaSuccess) {                                                                \
      ::cvt::set_error(std::string("kernel launch: ") + cudaGetErrorString(_e) + " @" + __FILE__ + ":" + std::to_string(__LINE__)); \
      return _e;                                                                            \
    }                                                                                       \
  } while (0)

// ---- programmatic dependent launch (PDL) ----
// The step is ~100 kernels of 5-100 us in one stream / one CUDA graph: without PDL every kernel boundary is a full drain (last CTA exits,
// memory flush) followed by the next grid's launch and prologue (shared-memory carve-up, mbarrier init, TMEM allocation, tensor-map
// fetch).  Every step kernel therefore
//   * is launched through CVT_LAUNCH with cudaLaunchAttributeProgrammaticStreamSerialization (in a captured graph: a programmatic edge),
//   * does everything that touches no global memory first, then pdl_wait() - returns once EVERY grid it depends on has completed and its
//     memory is visible, so nothing after it can race with a predecessor - and pdl_trigger(), which lets the NEXT grid's CTAs become
//     resident (as SMs free up) and run their own prologue while this grid is still computing.
// A kernel launched without the attribute sees both instructions as no-ops.  CVT_NO_PDL=1 launches everything without the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool pdl_enabled();

template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<Args&&>(args)...);
}
// launch + count + error check; every kernel launched this way starts with pdl_wait() before its first global-memory access
#define CVT_LAUNCH(kernel, grid, block, smem, stream, ...)                                                                      \
  do {                                                                                                                          \
    ::cvt::g_launch_count++;                                                                                                    \
    cudaError_t _e = ::cvt::launch_kernel(kernel, dim3(grid), dim3(block), (size_t)(smem), stream, __VA_ARGS__);                \
    if (_e == cudaSuccess) _e = cudaGetLastError();                                                                             \
    if (_e != cudaSuccess) {                                                                                                    \
      ::cvt::set_error(std::string("kernel launch: ") + cudaGetErrorString(_e) + " @" + __FILE__ + ":" + std::to_string(__LINE__)); \
      return _e;                                                                                                                \
    }                                                                                                                           \
  } while (0)

constexpr int kNumSMs = 148;

__host__ __device__ inline int cdiv(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ inline long long cdivl(long long a, long long b) { return (a + b - 1) / b; }

// token-row remap used to splice the cls row in and out of [B, Tc, D] tensors:
// r -> (r / period) * stride + r % period + off      (period == 0: identity)
struct RowMap {
  int period, stride, off;
  // 32-bit index arithmetic: token counts are far below 2^31 and a 64-bit division costs ~100 instructions on the device
  __host__ __device__ inline long long operator()(long long r) const {
    if (!period) return r;
    const int ri = (int)r, q = ri / period;
    return (long long)q * stride + (ri - q * period) + off;
  }
};
inline RowMap rowmap_none() { return RowMap{0, 0, 0}; }

// descriptor-table lookup of "multi-problem" launches: the last descriptor whose first tile (tile0, ascending) is <= bid.  One warp scans the
// table with independent loads (a serial scan is a chain of dependent global loads: ~0.3 us each, 10 us for the last entries of a 40-entry table)
template <typename Desc>
__device__ __forceinline__ int find_desc(const Desc* __restrict__ descs, int ndesc, int bid) {
  __shared__ int s_desc;
  if (threadIdx.x < 32) {
    int best = 0;
    for (int i = threadIdx.x; i < ndesc; i += 32)
      if (bid >= descs[i].tile0) best = i;
    best = __reduce_max_sync(0xffffffffu, best);
    if (threadIdx.x == 0) s_desc = best;
  }
  __syncthreads();
  return s_desc;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// round-to-nearest TF32 (10-bit mantissa): producers whose output only feeds tensor-core GEMMs store pre-rounded values, so that the
// tensor core's own truncation of the low 13 bits is exact and the rounding error stays unbiased (what cuBLAS' cvt.rna path gives TF)
__device__ __forceinline__ float round_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

// exact-erf GELU (tf.nn.gelu default, approximate=False) and its derivative
__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float gelu_erf_grad(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
  const float pdf = 0.39894228040143267794f * __expf(-0.5f * x * x);
  return cdf + x * pdf;
}

// GELU and its derivative from ONE exponential (TF32 training epilogue, gemm_epi.cuh): erfc(z) = t (a1 + t (a2 + ... a5 t)) e^{-z^2},
// t = 1 / (1 + p z), Abramowitz & Stegun 7.1.26, |error| <= 1.5e-7 (fp32 resolution); with z = |x| / sqrt 2 the same e^{-z^2} = e^{-x^2/2}
// is the Gaussian of the derivative.  ~15 instructions against ~35 for erff + expf.
__device__ __forceinline__ void gelu_and_grad(float x, float& g, float& dg) {
  const float z = fabsf(x) * 0.70710678118654752440f;
  const float t = __fdividef(1.0f, fmaf(0.3275911f, z, 1.0f));
  const float e = __expf(-z * z);
  float q = fmaf(t, 1.061405429f, -1.453152027f);
  q = fmaf(q, t, 1.421413741f); q = fmaf(q, t, -0.284496736f); q = fmaf(q, t, 0.254829592f);
  const float half_erfc = 0.5f * q * t * e;              // 0.5 erfc(|x| / sqrt 2)
  const float cdf = x >= 0.f ? 1.0f - half_erfc : half_erfc;
  g = x * cdf;
  dg = fmaf(x, 0.39894228040143267794f * e, cdf);
}

// ---- Philox4x32-10, bit-identical to oracle/cvt_oracle.py:philox4x32 ----
__host__ __device__ inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const unsigned long long p0 = (unsigned long long)c[0] * 0xD2511F53ull;
    const unsigned long long p1 = (unsigned long long)c[2] * 0xCD9E8D57ull;
    const uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
    const uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
    const uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}

}  // namespace cvt
