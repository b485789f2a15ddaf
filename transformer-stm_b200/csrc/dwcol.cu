// Depthwise 3x3 q/k/v Projection (DepthwiseConv2D(3, 1, 'same', use_bias=False), models/CvT(Par).py:93,157-159), forward and
// backward, as "column walkers": a thread owns one (column, channel) and walks down a segment of image rows.  A warp = 32
// consecutive channels of two adjacent pixels (the filter taps in registers serve both), so every global access moves full 128-byte lines.  Rows stream through a shared-memory
// ring filled by 4-byte cp.async (each thread copies its own pixel; the first / last warp also copy the halo column), several rows
// ahead of the row being consumed, so ~2 KB per warp is in flight without holding prefetch registers; the left / right
// neighbours of a pixel are read from the ring slots of the neighbouring warps.  HBM-bound by design: forward reads xn once and
// writes the three maps; backward reads the three map gradients and xn once and writes dxn.  CTAs are persistent and keep their
// channel group, so BatchNorm statistics / depthwise-weight gradients stay in registers across work items and leave the CTA once,
// (statistics as per-CTA partial slices summed by bn_finalize_kernel, tap gradients by one atomic add per CTA and element).
#include "layers.cuh"

namespace cvt {
namespace {
constexpr int CT = 256, NW = 8, COLS = 2 * NW;          // block = 32 channels x 16 columns, two adjacent columns per thread
constexpr int FP = 6, FR = FP + 2;                      // forward: rows in flight, ring rows
constexpr int BP = 3, BR = BP + 2;                      // backward (ring = 5 x 18 x 512 B = 45 KB)

struct ColGeo { int CG, ncolg, nseg, RS, items_per_cg, pack; };   // channel groups, column groups, row segments (of RS rows); pack = 2: images
                                                                   // at most 8 columns wide go two to a CTA (warps 0-3 / 4-7)

// 4-byte asynchronous copy global -> shared address; sz = 0 writes a zero and reads nothing (the source address may then lie outside
// the tensor: it is never dereferenced)
__device__ __forceinline__ void cp4(uint32_t dst, const float* src, int sz) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ---- forward -------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(CT, 2) dwcol_fwd_kernel(const __grid_constant__ DwProjP p, const ColGeo g) {
  pdl_wait(); pdl_trigger();
  __shared__ float ring[FR][COLS + 2][32];               // column slot s <-> image column colg * COLS - 1 + s
  __shared__ float sred[NW][6][32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int D = p.D, H = p.H, W = p.W;
  const int cg = blockIdx.x % g.CG, c = cg * 32 + tx;
  const uint32_t ring_u32 = (uint32_t)__cvta_generic_to_shared(&ring[0][2 * ty + 1][tx]);
  float w[3][9];
#pragma unroll
  for (int z = 0; z < 3; ++z)
#pragma unroll
    for (int k = 0; k < 9; ++k) w[z][k] = p.w[(z * 9 + k) * D + c];
  float ssum[3] = {0.f, 0.f, 0.f}, ssq[3] = {0.f, 0.f, 0.f};
  for (int item = blockIdx.x / g.CG; item < g.items_per_cg; item += gridDim.x / g.CG) {
    int t = item;
    const int colg = t % g.ncolg; t /= g.ncolg;
    const int seg = t % g.nseg; const int b = (t / g.nseg) * g.pack + (g.pack == 2 ? ty >> 2 : 0);
    const int j = colg * COLS + 2 * (g.pack == 2 ? ty & 3 : ty);            // this thread's columns: j, j + 1
    const bool jok0 = b < p.B && j < W, jok1 = b < p.B && j + 1 < W;
    const bool lz = g.pack == 2 && (ty & 3) == 0, rz = g.pack == 2 && (ty & 3) == 3;   // packed: the neighbour slot belongs to the other image
    const int r0 = seg * g.RS, rend = min(H, r0 + g.RS);
    const long long base = ((long long)b * p.Tc + p.has_cls) * D + (long long)j * D + c;
    const float* x = p.xn + base;
    cp_wait<0>();                                        // (only zero rows past the segment end can still be pending)
    __syncthreads();                                     // the previous item's last rows have been read
    // running state of the copy front (row ir into ring slot islot) and of the row being consumed: no multiplications in the loop
    const int rs = W * D, rlo = max(r0 - 1, 0), rhi = min(rend, H - 1);
    int ir = r0 - 1, islot = 0, cslot = 0;
    const float* qx = x + (long long)ir * rs;
    float* po = p.out + base + (long long)(r0 - 2) * rs;
    auto issue = [&]() {                                 // rows r0-1 .. rend (halo rows included); anything else is a zero row
      const bool rok = ir >= rlo && ir <= rhi;
      const uint32_t dst = ring_u32 + islot * ((COLS + 2) * 128);
      cp4(dst, qx, (rok && jok0) ? 4 : 0);
      cp4(dst + 128, qx + D, (rok && jok1) ? 4 : 0);
      if (ty == 0) cp4(dst - 128, qx - D, (rok && j > 0) ? 4 : 0);
      if (ty == NW - 1) cp4(dst + 256, qx + 2 * D, (rok && j + 2 < W) ? 4 : 0);
      cp_commit();
      ++ir; qx += rs; islot = islot + 1 == FR ? 0 : islot + 1;
    };
#pragma unroll
    for (int k = 0; k < FP; ++k) issue();
    float a0[2][3], a1[2][3], a2[2][3];                  // outputs g-1, g, g+1 of the three branches, for both columns
#pragma unroll
    for (int q = 0; q < 2; ++q)
#pragma unroll
      for (int z = 0; z < 3; ++z) { a0[q][z] = 0.f; a1[q][z] = 0.f; a2[q][z] = 0.f; }
    for (int gi = r0 - 1; gi <= rend; ++gi) {
      issue();
      cp_wait<FP>();
      __syncthreads();
      const float* row = &ring[cslot][2 * ty][tx];
      cslot = cslot + 1 == FR ? 0 : cslot + 1;
      const float xv[4] = {lz ? 0.f : row[0], row[32], row[64], rz ? 0.f : row[96]};      // columns j-1, j, j+1, j+2
#pragma unroll
      for (int q = 0; q < 2; ++q)
#pragma unroll
        for (int z = 0; z < 3; ++z) {
          a0[q][z] = fmaf(xv[q], w[z][6], fmaf(xv[q + 1], w[z][7], fmaf(xv[q + 2], w[z][8], a0[q][z])));       // bottom tap row of output g-1
          a1[q][z] = fmaf(xv[q], w[z][3], fmaf(xv[q + 1], w[z][4], fmaf(xv[q + 2], w[z][5], a1[q][z])));       // middle tap row of output g
          a2[q][z] = fmaf(xv[q], w[z][0], fmaf(xv[q + 1], w[z][1], fmaf(xv[q + 2], w[z][2], a2[q][z])));       // top tap row of output g+1
        }
      const int io = gi - 1;
      if (io >= r0 && io < rend) {
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          if (q == 0 ? jok0 : jok1) {
#pragma unroll
            for (int z = 0; z < 3; ++z) {
              const float y = a0[q][z];
              po[z * p.out_bs + q * D] = p.round_out ? round_tf32(y) : y;
              ssum[z] += y; ssq[z] = fmaf(y, y, ssq[z]);
            }
          }
        }
      }
#pragma unroll
      for (int q = 0; q < 2; ++q)
#pragma unroll
        for (int z = 0; z < 3; ++z) { a0[q][z] = a1[q][z]; a1[q][z] = a2[q][z]; a2[q][z] = 0.f; }
      po += rs;
    }
  }
  cp_wait<0>();
  // cls rows pass through untouched (models/CvT(Par).py:150,168-170), or are zeroed when the cls row is handled outside (fold.cu)
  if (p.has_cls) {
    for (int i = blockIdx.x * CT + threadIdx.x; i < p.B * 3 * (D >> 2); i += gridDim.x * CT) {
      const int q = i % (D >> 2), z = (i / (D >> 2)) % 3, b = i / (3 * (D >> 2));
      const long long row = (long long)b * p.Tc * D + q * 4;
      *reinterpret_cast<float4*>(p.out + z * p.out_bs + row) = p.cls_zero ? make_float4(0.f, 0.f, 0.f, 0.f) : *reinterpret_cast<const float4*>(p.xn + row);
    }
  }
  if (p.stats) {   // partial slice (blockIdx / CG): [sum q,k,v | sum of squares q,k,v][D]; this CTA fills its 32 channels
#pragma unroll
    for (int z = 0; z < 3; ++z) { sred[ty][z][tx] = ssum[z]; sred[ty][3 + z][tx] = ssq[z]; }
    __syncthreads();
    for (int o = threadIdx.x; o < 6 * 32; o += CT) {
      const int k = o >> 5, l = o & 31;
      float a = 0.f;
#pragma unroll
      for (int y = 0; y < NW; ++y) a += sred[y][k][l];
      p.stats[(long long)(blockIdx.x / g.CG) * 6 * D + k * D + cg * 32 + l] = a;
    }
  }
}

// ---- backward ------------------------------------------------------------------------------------------------------------------
// dx[i,j] = sum_z sum_{kh,kw} d_z[i-kh+1, j-kw+1] w_z[kh,kw];   dw_z[kh,kw] += x[i+kh-1, j+kw-1] d_z[i,j]
__global__ void __launch_bounds__(CT, 2) dwcol_bwd_kernel(const __grid_constant__ DwProjBwdP p, const ColGeo g) {
  pdl_wait(); pdl_trigger();
  constexpr int RING = BR * (COLS + 2) * 4 * 32, RED = NW * 27 * 32;
  __shared__ float smem[RING > RED ? RING : RED];
  float (*ring)[COLS + 2][4][32] = reinterpret_cast<float (*)[COLS + 2][4][32]>(smem);      // [row slot][column slot][x, d_q, d_k, d_v][channel]
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int D = p.D, H = p.H, W = p.W;
  const int cg = blockIdx.x % g.CG, c = cg * 32 + tx;
  const uint32_t ring_u32 = (uint32_t)__cvta_generic_to_shared(&ring[0][2 * ty + 1][0][tx]);
  float w[3][9], dw[3][9];
#pragma unroll
  for (int z = 0; z < 3; ++z)
#pragma unroll
    for (int k = 0; k < 9; ++k) { w[z][k] = p.w[(z * 9 + k) * D + c]; dw[z][k] = 0.f; }
  for (int item = blockIdx.x / g.CG; item < g.items_per_cg; item += gridDim.x / g.CG) {
    int t = item;
    const int colg = t % g.ncolg; t /= g.ncolg;
    const int seg = t % g.nseg; const int b = (t / g.nseg) * g.pack + (g.pack == 2 ? ty >> 2 : 0);
    const int j = colg * COLS + 2 * (g.pack == 2 ? ty & 3 : ty);            // this thread's columns: j, j + 1
    const bool jok0 = b < p.B && j < W, jok1 = b < p.B && j + 1 < W;
    const bool lz = g.pack == 2 && (ty & 3) == 0, rz = g.pack == 2 && (ty & 3) == 3;   // packed: the neighbour slot belongs to the other image
    const int r0 = seg * g.RS, rend = min(H, r0 + g.RS);
    const long long base = ((long long)b * p.Tc + p.has_cls) * D + (long long)j * D + c;
    const float* x = p.xn + base;
    const float* d0 = p.dconv + base;
    cp_wait<0>();                                        // (only zero rows past the segment end can still be pending)
    __syncthreads();                                     // the previous item's last rows have been read
    const int rs = W * D, rlo = max(r0 - 1, 0), rhi = min(rend, H - 1);
    int ir = r0 - 1, islot = 0, cslot = 0;
    const float* qx = x + (long long)ir * rs;
    const float* q0 = d0 + (long long)ir * rs;
    const float* q1 = q0 + p.bs;
    const float* q2 = q1 + p.bs;
    float* po = p.dxn + base + (long long)(r0 - 2) * rs;
    auto issue = [&]() {                                 // rows r0-1 .. rend (halo rows included); anything else is a zero row
      const bool rok = ir >= rlo && ir <= rhi;
      const uint32_t dst = ring_u32 + islot * ((COLS + 2) * 512);
      const int s0 = (rok && jok0) ? 4 : 0, s1 = (rok && jok1) ? 4 : 0;
      cp4(dst, qx, s0); cp4(dst + 128, q0, s0); cp4(dst + 256, q1, s0); cp4(dst + 384, q2, s0);
      cp4(dst + 512, qx + D, s1); cp4(dst + 640, q0 + D, s1); cp4(dst + 768, q1 + D, s1); cp4(dst + 896, q2 + D, s1);
      if (ty == 0) {
        const int sz = (rok && j > 0) ? 4 : 0;
        cp4(dst - 512, qx - D, sz); cp4(dst - 384, q0 - D, sz); cp4(dst - 256, q1 - D, sz); cp4(dst - 128, q2 - D, sz);
      }
      if (ty == NW - 1) {
        const int sz = (rok && j + 2 < W) ? 4 : 0;
        cp4(dst + 1024, qx + 2 * D, sz); cp4(dst + 1152, q0 + 2 * D, sz); cp4(dst + 1280, q1 + 2 * D, sz); cp4(dst + 1408, q2 + 2 * D, sz);
      }
      cp_commit();
      ++ir; qx += rs; q0 += rs; q1 += rs; q2 += rs; islot = islot + 1 == BR ? 0 : islot + 1;
    };
#pragma unroll
    for (int k = 0; k < BP; ++k) issue();
    float a0[2] = {0.f, 0.f}, a1[2] = {0.f, 0.f}, a2[2] = {0.f, 0.f};      // dx of rows g-1, g, g+1 for both columns
    float dprev[2][3];                                   // centre d of row g-1, counted only inside [r0, rend)
#pragma unroll
    for (int q = 0; q < 2; ++q)
#pragma unroll
      for (int z = 0; z < 3; ++z) dprev[q][z] = 0.f;
    float px[4] = {0.f, 0.f, 0.f, 0.f}, qx2[4] = {0.f, 0.f, 0.f, 0.f};     // x of rows g-1 (px) and g-2 (qx2), columns j-1 .. j+2
    for (int gi = r0 - 1; gi <= rend; ++gi) {
      issue();
      cp_wait<BP>();
      __syncthreads();
      const float* row = &ring[cslot][2 * ty][0][tx];
      cslot = cslot + 1 == BR ? 0 : cslot + 1;
      const float xv[4] = {lz ? 0.f : row[0], row[128], row[256], rz ? 0.f : row[384]};           // columns j-1, j, j+1, j+2
      const bool cur_in = gi >= r0 && gi < rend;
#pragma unroll
      for (int z = 0; z < 3; ++z) {
        const float dv[4] = {lz ? 0.f : row[32 * (z + 1)], row[32 * (z + 1) + 128], row[32 * (z + 1) + 256], rz ? 0.f : row[32 * (z + 1) + 384]};
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          // input gradient: d row g feeds dx rows g-1 (tap row 0), g (tap row 1), g+1 (tap row 2); d[., j-1] pairs with tap column 2
          a0[q] = fmaf(dv[q], w[z][2], fmaf(dv[q + 1], w[z][1], fmaf(dv[q + 2], w[z][0], a0[q])));
          a1[q] = fmaf(dv[q], w[z][5], fmaf(dv[q + 1], w[z][4], fmaf(dv[q + 2], w[z][3], a1[q])));
          a2[q] = fmaf(dv[q], w[z][8], fmaf(dv[q + 1], w[z][7], fmaf(dv[q + 2], w[z][6], a2[q])));
          // weight gradient, one row late: the centre d of row g-1 pairs with x rows g-2 (tap row 0), g-1 (tap row 1), g (tap row 2)
          const float dp = dprev[q][z];
          dw[z][0] = fmaf(qx2[q], dp, dw[z][0]); dw[z][1] = fmaf(qx2[q + 1], dp, dw[z][1]); dw[z][2] = fmaf(qx2[q + 2], dp, dw[z][2]);
          dw[z][3] = fmaf(px[q], dp, dw[z][3]); dw[z][4] = fmaf(px[q + 1], dp, dw[z][4]); dw[z][5] = fmaf(px[q + 2], dp, dw[z][5]);
          dw[z][6] = fmaf(xv[q], dp, dw[z][6]); dw[z][7] = fmaf(xv[q + 1], dp, dw[z][7]); dw[z][8] = fmaf(xv[q + 2], dp, dw[z][8]);
          dprev[q][z] = cur_in ? dv[q + 1] : 0.f;
        }
      }
      const int io = gi - 1;
      if (io >= r0 && io < rend) {
        if (jok0) po[0] = a0[0];
        if (jok1) po[D] = a0[1];
      }
      po += rs;
#pragma unroll
      for (int q = 0; q < 2; ++q) { a0[q] = a1[q]; a1[q] = a2[q]; a2[q] = 0.f; }
#pragma unroll
      for (int q = 0; q < 4; ++q) { qx2[q] = px[q]; px[q] = xv[q]; }
    }
  }
  cp_wait<0>();
  if (p.has_cls) {   // the cls row bypassed conv and BN: its gradient is the plain sum of the three branches
    for (int i = blockIdx.x * CT + threadIdx.x; i < p.B * (D >> 2); i += gridDim.x * CT) {
      const int q = i % (D >> 2), b = i / (D >> 2);
      const long long row = (long long)b * p.Tc * D + q * 4;
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int z = 0; z < 3; ++z) { const float4 v = *reinterpret_cast<const float4*>(p.dconv + z * p.bs + row); s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
      *reinterpret_cast<float4*>(p.dxn + row) = s;
    }
  }
  if (p.ddw) {       // this CTA's tap gradients of its 32 channels, summed over its 8 warps, go straight into ddw [27][D]
    float (*sred)[27][32] = reinterpret_cast<float (*)[27][32]>(smem);
    __syncthreads();                                     // the ring is dead
#pragma unroll
    for (int z = 0; z < 3; ++z)
#pragma unroll
      for (int k = 0; k < 9; ++k) sred[ty][z * 9 + k][tx] = dw[z][k];
    __syncthreads();
    for (int o = threadIdx.x; o < 27 * 32; o += CT) {
      const int k = o >> 5, l = o & 31;
      float a = 0.f;
#pragma unroll
      for (int y = 0; y < NW; ++y) a += sred[y][k][l];
      atomicAdd(p.ddw + k * D + cg * 32 + l, a);          // <= 296 / CG adds per address, spread over the kernel's lifetime
    }
  }
}

bool make_colgeo(int B, int H, int W, int D, int cap_blocks, ColGeo* g, int* grid) {
  if (D % 32 != 0) return false;
  g->CG = D / 32; g->ncolg = cdiv(W, COLS);
  g->RS = H >= 32 ? 16 : 8;
  g->nseg = cdiv(H, g->RS);
  g->pack = W <= COLS / 2 ? 2 : 1;
  g->items_per_cg = cdiv(B, g->pack) * g->nseg * g->ncolg;
  int per_cg = cap_blocks / g->CG;
  if (per_cg < 1) return false;
  if (per_cg > g->items_per_cg) per_cg = g->items_per_cg;
  *grid = per_cg * g->CG;
  return true;
}
}  // namespace

// the three branches must all be depthwise convolutions (projection_method 'dw_bn'); otherwise cudaErrorNotSupported
cudaError_t launch_dwcol_fwd(const DwProjP& p, cudaStream_t s, int* stat_slices) {
  ColGeo g; int grid;
  if (p.method[0] != 0 || p.method[1] != 0 || p.method[2] != 0 || !make_colgeo(p.B, p.H, p.W, p.D, p.nparts, &g, &grid)) return cudaErrorNotSupported;
  *stat_slices = grid / g.CG;                            // slices of p.stats written by this launch (<= p.nparts): bn_finalize sums exactly these
  CVT_LAUNCH((dwcol_fwd_kernel), grid, CT, 0, s, p, g);
  return cudaSuccess;
}

cudaError_t launch_dwcol_bwd(const DwProjBwdP& p, cudaStream_t s) {
  ColGeo g; int grid;
  if (p.method[0] != 0 || p.method[1] != 0 || p.method[2] != 0 || p.coef || !make_colgeo(p.B, p.H, p.W, p.D, 2 * kNumSMs, &g, &grid)) return cudaErrorNotSupported;
  CVT_LAUNCH((dwcol_bwd_kernel), grid, CT, 0, s, p, g);
  return cudaSuccess;
}

}  // namespace cvt
