// Host-side TMA descriptor cache shared by the TMA-fed kernels (defined in gemm_tma.cu).
#pragma once
#include <cuda.h>

namespace cvt {
// fp32 tensor [d2][d1][d0] (d0 contiguous), strides in floats, box {32, box_rows, 1}, zero fill out of bounds.
// swizzle_mode: CU_TENSOR_MAP_SWIZZLE_128B (K-major operands) or CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B (MN-major tf32 operands).
bool make_tmap_3d(CUtensorMap* out, const float* base, long long d0, long long d1, long long d2, long long stride1, long long stride2,
                  int box_rows, int swizzle_mode);
// Implicit-GEMM view of a Conv2D(padding='same') input for TMA (ConvEmbed, models/CvT(Par).py:203-208): the NHWC map of image b starts at
// base + b * bstride (floats) and is described as the 4-D tensor [B][H][W][C] with element strides {1, s, s, 1}; one box = 32 channels of
// `tokens` consecutive output pixels (row-major over (b, i, j)) for ONE filter tap, addressed by the input coordinate of the tap at the
// first pixel.  Out-of-bounds coordinates (the SAME padding, the ragged last tile) read zeros.  tokens must be a whole number of output
// rows or of images: conv_tma_box() decides.
struct ConvBox { int bw, bh, bb; };      // output pixels per box along W, H, B (bw * bh * bb = tokens)
bool conv_tma_box(int ho, int wo, int tokens, ConvBox* box);
bool make_tmap_conv(CUtensorMap* out, const float* base, int hin, int win, int cin, long long bstride, int batch, int stride, ConvBox box,
                    int swizzle_mode);
}  // namespace cvt
