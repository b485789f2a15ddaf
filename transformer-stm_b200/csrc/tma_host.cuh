// Host-side TMA descriptor cache shared by the TMA-fed kernels (defined in gemm_tma.cu).
#pragma once
#include <cuda.h>

namespace cvt {
// fp32 tensor [d2][d1][d0] (d0 contiguous), strides in floats, box {32, box_rows, 1}, zero fill out of bounds.
// swizzle_mode: CU_TENSOR_MAP_SWIZZLE_128B (K-major operands) or CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B (MN-major tf32 operands).
bool make_tmap_3d(CUtensorMap* out, const float* base, long long d0, long long d1, long long d2, long long stride1, long long stride2,
                  int box_rows, int swizzle_mode);
}  // namespace cvt
