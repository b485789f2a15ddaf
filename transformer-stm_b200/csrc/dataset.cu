// Input pipeline on the device: the part of preprocess_data that follows the JPEG decode
// (models/CvT(Par).py:420-423: cv2.resize(INTER_LINEAR) -> cv2.cvtColor(BGR2GRAY) -> / 255.0), the dataset kept resident in
// HBM, and the per-epoch shuffle of model.fit (:475) as a gather by index instead of a host-side copy of the whole set.
//
// The resize and the grey conversion are the 8-bit fixed-point algorithms of OpenCV 4.x (imgproc/resize.cpp HResizeLinear /
// VResizeLinear with 11-bit coefficients, color_rgb.simd.hpp RGB2Gray with 15-bit coefficients): integer work, reproduced bit for bit
// (oracle/preprocess_oracle.py is pinned against cv2 itself; tests/golden/preprocess_*.npz are cv2 outputs).
#include "layers.cuh"

namespace cvt {

namespace {

// rows of `row4` float4 each; idx arrays are nullable (identity).  grid = (chunks per row, rows)
__global__ void __launch_bounds__(256) move_rows_kernel(float4* __restrict__ dst, const int32_t* __restrict__ dst_idx, const float4* __restrict__ src,
                                                        const int32_t* __restrict__ src_idx, int row4,
                                                        float* __restrict__ dst_p, const float* __restrict__ src_p, int P,
                                                        float* __restrict__ dst_y, const float* __restrict__ src_y, int nc) {
  const int r = blockIdx.y;
  const long long so = src_idx ? (long long)src_idx[r] : r, dno = dst_idx ? (long long)dst_idx[r] : r;
  const float4* s = src + so * row4; float4* d = dst + dno * row4;
  // 4 x 16 bytes in flight per thread
  const int base = blockIdx.x * 1024 + threadIdx.x;
  float4 v[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) { const int i = base + j * 256; if (i < row4) v[j] = __ldg(s + i); }
#pragma unroll
  for (int j = 0; j < 4; ++j) { const int i = base + j * 256; if (i < row4) d[i] = v[j]; }
  if (blockIdx.x == 0) {
    if (src_p) for (int i = threadIdx.x; i < P; i += 256) dst_p[dno * P + i] = src_p[so * P + i];
    if (src_y) for (int i = threadIdx.x; i < nc; i += 256) dst_y[dno * nc + i] = src_y[so * nc + i];
  }
}

__global__ void __launch_bounds__(256) f64_to_f32_kernel(float* __restrict__ dst, const double* __restrict__ src, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) dst[i] = (float)src[i];
}

// one thread per output pixel.  tab = [sx | a0 | a1] (dw entries each) then [sy | b0 | b1] (dh entries each); lut[g] = (float)(g / 255.0)
__global__ void __launch_bounds__(256) preprocess_u8_kernel(const uint8_t* __restrict__ pix, int sh, int sw, int sc, const int32_t* __restrict__ tab,
                                                            const float* __restrict__ lut, int dh, int dw, long long count, float* __restrict__ out) {
  const long long total = count * dh * dw;
  const int32_t* sxp = tab; const int32_t* a0p = tab + dw; const int32_t* a1p = tab + 2 * dw;
  const int32_t* syp = tab + 3 * dw; const int32_t* b0p = syp + dh; const int32_t* b1p = syp + 2 * dh;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(i % dw); const long long t = i / dw; const int y = (int)(t % dh); const long long n = t / dh;
    const int x0 = sxp[x], x1 = min(x0 + 1, sw - 1), y0 = min(max(syp[y], 0), sh - 1), y1 = min(max(syp[y] + 1, 0), sh - 1);
    const int a0 = a0p[x], a1 = a1p[x], b0 = b0p[y], b1 = b1p[y];
    const uint8_t* im = pix + n * (long long)sh * sw * sc;
    const uint8_t* r0 = im + (long long)y0 * sw * sc; const uint8_t* r1 = im + (long long)y1 * sw * sc;
    int ch[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      if (c < sc) {
        const int h0 = (int)r0[x0 * sc + c] * a0 + (int)r0[x1 * sc + c] * a1;      // HResizeLinear: no shift
        const int h1 = (int)r1[x0 * sc + c] * a0 + (int)r1[x1 * sc + c] * a1;
        ch[c] = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;    // VResizeLinear, FixedPtCast<int, uchar, 22>
      } else ch[c] = ch[0];
    }
    const int g = sc >= 3 ? (ch[0] * 3735 + ch[1] * 19235 + ch[2] * 9798 + (1 << 14)) >> 15 : ch[0];   // BGR order, RGB2Gray<uchar>
    out[i] = lut[g];
  }
}

inline int blocks_for(long long n) { const long long b = (n + 255) / 256; return (int)(b < 1 ? 1 : (b > 148 * 16 ? 148 * 16 : b)); }

}  // namespace

cudaError_t launch_move_rows(float* dst_img, const int32_t* dst_idx, const float* src_img, const int32_t* src_idx, int rows, long long row_floats,
                             float* dst_p, const float* src_p, int P, float* dst_y, const float* src_y, int nc, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  if ((row_floats & 3) || (reinterpret_cast<uintptr_t>(dst_img) & 15) || (reinterpret_cast<uintptr_t>(src_img) & 15)) {
    set_error("move_rows: image rows must be multiples of 16 bytes"); return cudaErrorInvalidValue;
  }
  const int row4 = (int)(row_floats >> 2);
  dim3 grid((row4 + 1023) / 1024, rows);
  move_rows_kernel<<<grid, 256, 0, s>>>(reinterpret_cast<float4*>(dst_img), dst_idx, reinterpret_cast<const float4*>(src_img), src_idx, row4,
                                        dst_p, src_p, P, dst_y, src_y, nc);
  CVT_LAUNCH_CHECK();
  return cudaSuccess;
}

cudaError_t launch_f64_to_f32(float* dst, const double* src, long long n, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  f64_to_f32_kernel<<<blocks_for(n), 256, 0, s>>>(dst, src, n);
  CVT_LAUNCH_CHECK();
  return cudaSuccess;
}

// cv::resize(INTER_LINEAR) coordinate tables for one axis (resize.cpp): float source coordinate, floor, 11-bit weights rounded
// half-to-even (saturate_cast<short> == cvRound).  clamp_weights (the x axis): columns outside the image take weight (2048, 0) on the
// border column.  The y axis keeps its fractional weights; resizeGeneric_Invoker clips the two row indices instead (done in the kernel).
void resize_axis_table(int src, int dst, int clamp_weights, int32_t* s_out, int32_t* w0, int32_t* w1) {
  const double inv_scale = (double)dst / (double)src;
  const double scale = 1.0 / inv_scale;
  for (int d = 0; d < dst; ++d) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f -= (float)s;
    if (clamp_weights) {
      if (s < 0) { f = 0.f; s = 0; }
      if (s >= src - 1) { f = 0.f; s = src - 1; }
    }
    s_out[d] = s;
    w0[d] = (int32_t)lrintf((1.f - f) * 2048.f);
    w1[d] = (int32_t)lrintf(f * 2048.f);
  }
}

cudaError_t launch_preprocess_u8(const uint8_t* pix_dev, int sh, int sw, int sc, const int32_t* tab_dev, const float* lut_dev, int dh, int dw,
                                 long long count, float* out, cudaStream_t s) {
  if (count <= 0) return cudaSuccess;
  preprocess_u8_kernel<<<blocks_for(count * dh * dw), 256, 0, s>>>(pix_dev, sh, sw, sc, tab_dev, lut_dev, dh, dw, count, out);
  CVT_LAUNCH_CHECK();
  return cudaSuccess;
}

}  // namespace cvt
