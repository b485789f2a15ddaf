// Stage-1 ConvEmbed of the reference spec as a direct convolution (inference): Conv2D(64, 7, strides 4, 'same') + bias on the 1-channel
// image (models/CvT(Par).py:203-208 with the spec of :66-72).  With C_in = 1 the contraction is only 49 long and each filter tap of a pixel is one
// scalar: as a GEMM it needs a patch matrix 4x the size of the image (im2col 12.6 us + GEMM 14 us at batch 64), as a direct kernel it is
// 3136 FMAs per output pixel out of shared memory.  CTA = 8 output rows x 32 output columns of one image: the 35 x 132 input window (zero
// = SAME padding) and the 49 x 64 filter live in shared memory; a thread owns 4 adjacent output columns x 16 channels, so one filter load
// (LDS.128) feeds 16 FMAs and the 4 threads of a pixel quad write 256 contiguous bytes per pixel.  fp32 FMA throughout (more exact than
// the TF32 GEMM it replaces).  Training keeps im2col + GEMM: the weight gradient reads the patch matrix.
#include "layers.cuh"
#include "gemm.cuh"

namespace cvt {
namespace {

constexpr int KS = 7, ST = 4, DD = 64, ROWS = 8, COLS_OUT = 32;
constexpr int IN_ROWS = (ROWS - 1) * ST + KS;           // 35
constexpr int IN_COLS = (COLS_OUT - 1) * ST + KS;       // 131
constexpr int IN_PITCH = 132;

__global__ void __launch_bounds__(256) conv1_direct_kernel(const float* __restrict__ x, ConvGeom g, const float* __restrict__ w /*[49][64]*/,
                                                           const float* __restrict__ bias, float* __restrict__ out, long long out_img_stride,
                                                           int out_row_off, int col_blocks, int row_blocks) {
  __shared__ __align__(16) float sIn[IN_ROWS * IN_PITCH];
  __shared__ __align__(16) float sW[KS * KS * DD];
  int bid = blockIdx.x;
  const int cb = bid % col_blocks; bid /= col_blocks;
  const int rb = bid % row_blocks; const int b = bid / row_blocks;
  const int i0 = rb * ROWS, j0 = cb * COLS_OUT;
  const float* xb = x + (long long)b * g.bstride + g.boff;
  const int y0 = i0 * ST - g.pt, x0 = j0 * ST - g.pl;
  for (int t = threadIdx.x; t < IN_ROWS * IN_PITCH; t += 256) {
    const int r = t / IN_PITCH, c = t - r * IN_PITCH;
    const int yy = y0 + r, xx = x0 + c;
    sIn[t] = (c < IN_COLS && yy >= 0 && yy < g.hin && xx >= 0 && xx < g.win) ? __ldg(xb + (long long)yy * g.win + xx) : 0.f;
  }
  for (int t = threadIdx.x; t < KS * KS * DD / 4; t += 256) reinterpret_cast<float4*>(sW)[t] = __ldg(reinterpret_cast<const float4*>(w) + t);
  __syncthreads();
  // thread -> (output row r, column quad q, channel group cg): lanes 0..3 = the 4 channel groups of one quad (256 contiguous output bytes per pixel)
  const int cg = threadIdx.x & 3, q = (threadIdx.x >> 2) & 7, r = threadIdx.x >> 5;
  float acc[4][16];
#pragma unroll
  for (int p = 0; p < 4; ++p)
#pragma unroll
    for (int c = 0; c < 16; ++c) acc[p][c] = 0.f;
#pragma unroll 1
  for (int kh = 0; kh < KS; ++kh) {
    // the 19 inputs of this filter row that the 4 pixels (input columns 16 q + 4 p + kw) touch, as 5 aligned float4s
    float in[20];
    const float4* src = reinterpret_cast<const float4*>(sIn + (r * ST + kh) * IN_PITCH + q * 16);
#pragma unroll
    for (int v = 0; v < 5; ++v) { const float4 t4 = src[v]; in[4 * v] = t4.x; in[4 * v + 1] = t4.y; in[4 * v + 2] = t4.z; in[4 * v + 3] = t4.w; }
#pragma unroll
    for (int kw = 0; kw < KS; ++kw) {
      const float4* wp = reinterpret_cast<const float4*>(sW + (kh * KS + kw) * DD + cg * 16);
      float wv[16];
#pragma unroll
      for (int v = 0; v < 4; ++v) { const float4 t4 = wp[v]; wv[4 * v] = t4.x; wv[4 * v + 1] = t4.y; wv[4 * v + 2] = t4.z; wv[4 * v + 3] = t4.w; }
#pragma unroll
      for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int c = 0; c < 16; ++c) acc[p][c] = fmaf(in[4 * p + kw], wv[c], acc[p][c]);
    }
  }
  const int oi = i0 + r;
  if (oi >= g.ho) return;
  float bv[16];
#pragma unroll
  for (int v = 0; v < 4; ++v) { const float4 t4 = __ldg(reinterpret_cast<const float4*>(bias + cg * 16) + v); bv[4 * v] = t4.x; bv[4 * v + 1] = t4.y; bv[4 * v + 2] = t4.z; bv[4 * v + 3] = t4.w; }
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const int oj = j0 + q * 4 + p;
    if (oj >= g.wo) continue;
    float* dst = out + (long long)b * out_img_stride + ((long long)(out_row_off + oi * g.wo + oj)) * DD + cg * 16;
#pragma unroll
    for (int v = 0; v < 4; ++v)
      *reinterpret_cast<float4*>(dst + 4 * v) = make_float4(acc[p][4 * v] + bv[4 * v], acc[p][4 * v + 1] + bv[4 * v + 1], acc[p][4 * v + 2] + bv[4 * v + 2],
                                                            acc[p][4 * v + 3] + bv[4 * v + 3]);
  }
}

}  // namespace

bool conv1_direct_supports(const ConvGeom& g, int D) {
  return g.cin == 1 && g.k == KS && g.s == ST && D == DD && (g.bstride & 3) == 0;
}

// out: token rows of image b start at out + b * out_img_stride + out_row_off * 64 (out_row_off = 1 when the stage carries a cls row)
cudaError_t launch_conv1_direct(const float* x, const ConvGeom& g, const float* w, const float* bias, float* out, long long out_img_stride,
                                int out_row_off, int B, cudaStream_t s) {
  if (!conv1_direct_supports(g, DD)) { set_error("conv1_direct: unsupported geometry"); return cudaErrorNotSupported; }
  const int col_blocks = cdiv(g.wo, COLS_OUT), row_blocks = cdiv(g.ho, ROWS);
  conv1_direct_kernel<<<B * row_blocks * col_blocks, 256, 0, s>>>(x, g, w, bias, out, out_img_stride, out_row_off, col_blocks, row_blocks);
  CVT_LAUNCH_CHECK();
  return cudaSuccess;
}

}  // namespace cvt
