// Microbenchmark: cycles per tcgen05.mma kind::tf32 (and kind::f16) for a few shapes, SS and TS operand modes, one CTA per SM.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I transformer-stm_b200/csrc tools/micro/mma_rate.cu -o tools/micro/mma_rate
#include <cstdio>
#include <cuda_runtime.h>
#include "tc_common.cuh"
using namespace cvt::tc;

template <int N, int MODE>   // MODE 0: SS tf32, 1: TS tf32, 2: SS bf16, 3: TS tf32 with MN-major B (32B-atom swizzle), 4: SS tf32 MN-major B, 5: SS tf32 walking K like the kernels
__global__ void __launch_bounds__(128, 1) rate_kernel(long long* out, int iters) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  __shared__ uint64_t bar; __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += 128) reinterpret_cast<float*>(sm)[i] = 1.0f;
  if (threadIdx.x < 32) tmem_alloc(&slot, 512);
  if (threadIdx.x == 32) { mbar_init(&bar, 1); mbar_fence_init(); }
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc(MODE == 2 ? kFmtBF16 : kFmtTF32, 128, N, 0, (MODE == 3 || MODE == 4) ? 1 : 0);
    const uint64_t bmn = make_smem_desc(smem_u32(sm) + 16384, N * 128, 512, kLayoutSw128Base32);
    const uint64_t ad = make_smem_desc(smem_u32(sm), 16, 1024), bd = make_smem_desc(smem_u32(sm) + 16384, 16, 1024);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (MODE == 3) mma_tf32_ts(tmem, tmem + 256 + j * 8, bmn + (uint64_t)(j * 64), idesc, 1u);
        else if (MODE == 4) mma_tf32_ss(tmem, ad + (uint64_t)((j & 3) * 2), bmn + (uint64_t)(j * 64), idesc, 1u);
        else if (MODE == 5) mma_tf32_ss(tmem, ad + (uint64_t)((j >> 2) * 1024 + (j & 3) * 2), bd + (uint64_t)((j >> 2) * (N * 8) + (j & 3) * 2), idesc, 1u);
        else if (MODE == 1) mma_tf32_ts(tmem, tmem + 256 + j * 8, bd + (uint64_t)((j & 3) * 2), idesc, 1u);
        else if (MODE == 2) mma_bf16_ss(tmem, ad + (uint64_t)((j & 3) * 2), bd + (uint64_t)((j & 3) * 2), idesc, 1u);
        else mma_tf32_ss(tmem, ad + (uint64_t)((j & 3) * 2), bd + (uint64_t)((j & 3) * 2), idesc, 1u);
      }
    }
    mma_commit(&bar);
    mbar_wait(&bar, 0);
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tmem, 512);
}

template <int N, int MODE> void run(const char* name, int grid) {
  long long* d; cudaMalloc(&d, 8);
  cudaFuncSetAttribute(rate_kernel<N, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  const int iters = 2000;
  rate_kernel<N, MODE><<<grid, 128, 64 * 1024>>>(d, iters);
  rate_kernel<N, MODE><<<grid, 128, 64 * 1024>>>(d, iters);
  long long h = 0; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  cudaError_t e = cudaDeviceSynchronize();
  printf("%-28s grid %3d: %.1f cycles per MMA (128 x %d x %s)  %s\n", name, grid, (double)h / (iters * 8.0), N, MODE == 2 ? "16 bf16" : "8 tf32", cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  for (int grid : {148}) {
    run<64, 0>("tf32 SS N=64", grid); run<128, 0>("tf32 SS N=128", grid); run<256, 0>("tf32 SS N=256", grid);
    run<64, 1>("tf32 TS N=64", grid); run<128, 1>("tf32 TS N=128", grid);
    run<64, 2>("bf16 SS N=64", grid); run<256, 2>("bf16 SS N=256", grid);
    run<64, 3>("tf32 TS N=64 B mn-major", grid);
  }
  return 0;
}
