// Depthwise 3x3 q/k/v Projection (DepthwiseConv2D(3, 1, 'same', use_bias=False), models/CvT(Par).py:93,157-159), forward and
// backward, as shared-memory band kernels: a CTA stages a band of TR (+2 halo) image rows x W columns x CH channels with cp.async
// (every byte of the maps is fetched from HBM/L2 once, 64-byte or wider segments), then threads walk columns with a 3x3 register
// window fed by shared memory.  HBM-bound by design: forward reads xn once and writes the three maps; backward reads the three
// map gradients and xn once and writes dxn.  BatchNorm statistics / depthwise-weight gradients leave the CTA as per-CTA partial
// slices (plain stores), summed by bn_finalize_kernel / reduce_partials_kernel.
#include "layers.cuh"

namespace cvt {
namespace {

constexpr int NT = 384;

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

struct Band { int CH, TR, nchunk, nband, ntiles; };

// stage rows r0-1 .. r0+TR of channels [c0, c0+CH) of one map into plane[(row)(col)(ch)]; rows outside the image become zeros
__device__ __forceinline__ void load_band(float* plane, const float* src /* image base + c0 */, int D, int H, int W, int r0, int TR, int CH) {
  const int q4 = CH >> 2, per_row = W * q4, total = (TR + 2) * per_row;
  for (int i = threadIdx.x; i < total; i += NT) {
    const int rr = i / per_row, rem = i - rr * per_row, col = rem / q4, q = rem - col * q4;
    const int gi = r0 - 1 + rr;
    float* dst = plane + ((rr * W + col) * CH + q * 4);
    if (gi >= 0 && gi < H) cp_async16(dst, src + ((long long)gi * W + col) * D + q * 4);
    else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

constexpr int TRC = 8;      // band height (compile time: the row loops are fully unrolled so that the accumulators rotate by name)

__device__ __forceinline__ float4 f4z() { return make_float4(0.f, 0.f, 0.f, 0.f); }
__device__ __forceinline__ void fma4(float4& a, const float4& x, const float4& w) {
  a.x = fmaf(x.x, w.x, a.x); a.y = fmaf(x.y, w.y, a.y); a.z = fmaf(x.z, w.z, a.z); a.w = fmaf(x.w, w.w, a.w);
}

// Each work item = (branch z, column j, float4 channel group).  Plane row rr (image row r0-1+rr) is read once (left, mid, right
// neighbours) and immediately spread over the three output rows it touches, so only three float4 accumulators are live.
__global__ void __launch_bounds__(NT, 2) dwplane_fwd_kernel(const __grid_constant__ DwProjP p, Band bd) {
  extern __shared__ __align__(16) float smem[];
  const int D = p.D, H = p.H, W = p.W, CH = bd.CH;
  float* plane = smem;                                   // [(TRC+2)][W][CH]
  float* sstat = smem + (TRC + 2) * W * CH;              // [6][D] statistics of everything this CTA produced
  float* sitem = sstat + 6 * D;                          // [3][2][W][CH] per-item sums of this tile (reduced over columns without atomics)
  const int tid = threadIdx.x;
  if (p.stats) for (int i = tid; i < 6 * D; i += NT) sstat[i] = 0.f;
  const int q4n = CH >> 2, items = 3 * W * q4n, rowf = W * CH;
  for (int tile = blockIdx.x; tile < bd.ntiles; tile += gridDim.x) {
    int t = tile;
    const int chunk = t % bd.nchunk; t /= bd.nchunk;
    const int band = t % bd.nband; const int b = t / bd.nband;
    const int c0 = chunk * CH, r0 = band * TRC;
    const long long base = ((long long)b * p.Tc + p.has_cls) * D + c0;
    __syncthreads();                                     // previous tile's readers are done with the plane
    load_band(plane, p.xn + base, D, H, W, r0, TRC, CH);
    cp_async_wait_all();
    __syncthreads();
    const int rend = min(TRC, H - r0);
    for (int it = tid; it < items; it += NT) {
      const int z = it / (W * q4n), rem = it - z * (W * q4n), j = rem / q4n, c = (rem - j * q4n) << 2;
      float4 w[9];
#pragma unroll
      for (int k = 0; k < 9; ++k) w[k] = *reinterpret_cast<const float4*>(p.w + (z * 9 + k) * D + c0 + c);
      const bool lok = j > 0, rok = j + 1 < W;
      const float* pm = plane + j * CH + c;
      float* out = p.out + z * p.out_bs + base + (long long)r0 * W * D + (long long)j * D + c;
      float4 ssum = f4z(), ssq = f4z();
      float4 acc0 = f4z(), acc1 = f4z(), acc2 = f4z();   // outputs rr-2, rr-1, rr
#pragma unroll
      for (int rr = 0; rr < TRC + 2; ++rr) {
        const float4 xm = *reinterpret_cast<const float4*>(pm + rr * rowf);
        const float4 xl = lok ? *reinterpret_cast<const float4*>(pm + rr * rowf - CH) : f4z();
        const float4 xr = rok ? *reinterpret_cast<const float4*>(pm + rr * rowf + CH) : f4z();
        fma4(acc0, xl, w[6]); fma4(acc0, xm, w[7]); fma4(acc0, xr, w[8]);       // bottom row of output rr-2
        fma4(acc1, xl, w[3]); fma4(acc1, xm, w[4]); fma4(acc1, xr, w[5]);       // middle row of output rr-1
        fma4(acc2, xl, w[0]); fma4(acc2, xm, w[1]); fma4(acc2, xr, w[2]);       // top row of output rr
        if (rr >= 2 && rr - 2 < rend) {
          const float4 y = acc0;
          *reinterpret_cast<float4*>(out + (long long)(rr - 2) * W * D) =
              p.round_out ? make_float4(round_tf32(y.x), round_tf32(y.y), round_tf32(y.z), round_tf32(y.w)) : y;
          ssum.x += y.x; ssum.y += y.y; ssum.z += y.z; ssum.w += y.w; fma4(ssq, y, y);
        }
        acc0 = acc1; acc1 = acc2; acc2 = f4z();
      }
      if (p.stats) {
        *reinterpret_cast<float4*>(sitem + ((z * 2 + 0) * W + j) * CH + c) = ssum;
        *reinterpret_cast<float4*>(sitem + ((z * 2 + 1) * W + j) * CH + c) = ssq;
      }
    }
    if (p.stats) {
      __syncthreads();
      for (int o = tid; o < 6 * CH; o += NT) {           // o = (z, which, channel)
        const int zw = o / CH, c = o - zw * CH;
        float a = 0.f;
        for (int j = 0; j < W; ++j) a += sitem[(zw * W + j) * CH + c];
        sstat[((zw & 1) * 3 + (zw >> 1)) * D + c0 + c] += a;
      }
    }
  }
  // cls rows pass through untouched (models/CvT(Par).py:150,168-170), or are zeroed when the cls row is handled outside (fold.cu)
  if (p.has_cls) {
    for (int i = blockIdx.x * NT + tid; i < p.B * 3 * (D >> 2); i += gridDim.x * NT) {
      const int q = i % (D >> 2), z = (i / (D >> 2)) % 3, b = i / (3 * (D >> 2));
      const long long row = (long long)b * p.Tc * D + q * 4;
      *reinterpret_cast<float4*>(p.out + z * p.out_bs + row) = p.cls_zero ? f4z() : *reinterpret_cast<const float4*>(p.xn + row);
    }
  }
  if (p.stats) {
    __syncthreads();
    for (int i = tid; i < 6 * D; i += NT) p.stats[(long long)blockIdx.x * 6 * D + i] = sstat[i];
  }
}

// backward: per tile, for z = q, k, v in turn: stage the band of d(conv_z); half of the items add its transposed-conv contribution
// to the dx band held in shared memory, the other half accumulate the depthwise-weight gradient against the xn band
__global__ void __launch_bounds__(NT, 2) dwplane_bwd_kernel(const __grid_constant__ DwProjBwdP p, Band bd) {
  extern __shared__ __align__(16) float smem[];
  const int D = p.D, H = p.H, W = p.W, CH = bd.CH;
  const int psz = (TRC + 2) * W * CH;
  float* xpl = smem;                    // xn band
  float* dpl0 = smem + psz;             // d(conv_z) bands, double-buffered: branch z+1 streams in while branch z is consumed
  float* dxs = smem + 3 * psz;          // dx band [TRC][W][CH]
  float* sdw = dxs + TRC * W * CH;      // [27][D] weight-gradient partial of this CTA
  float* sitem = sdw + 27 * D;          // [9][W][CH] per-item weight-gradient sums of this (tile, z)
  const int tid = threadIdx.x;
  for (int i = tid; i < 27 * D; i += NT) sdw[i] = 0.f;
  const int q4n = CH >> 2, half = W * q4n, rowf = W * CH;      // items per role
  for (int tile = blockIdx.x; tile < bd.ntiles; tile += gridDim.x) {
    int t = tile;
    const int chunk = t % bd.nchunk; t /= bd.nchunk;
    const int band = t % bd.nband; const int b = t / bd.nband;
    const int c0 = chunk * CH, r0 = band * TRC;
    const long long base = ((long long)b * p.Tc + p.has_cls) * D + c0;
    const int rend = min(TRC, H - r0);
    __syncthreads();
    load_band(xpl, p.xn + base, D, H, W, r0, TRC, CH);
    load_band(dpl0, p.dconv + base, D, H, W, r0, TRC, CH);
    cp_async_commit();
    load_band(dpl0 + psz, p.dconv + p.bs + base, D, H, W, r0, TRC, CH);
    cp_async_commit();
    for (int z = 0; z < 3; ++z) {
      const float* dpl = dpl0 + (z & 1) * psz;
      if (z < 2) cp_async_wait<1>(); else cp_async_wait<0>();
      __syncthreads();
      for (int it = tid; it < 2 * half; it += NT) {
        const int role = it / half, rem = it - role * half, j = rem / q4n, c = (rem - j * q4n) << 2;
        const bool lok = j > 0, rok = j + 1 < W;
        if (role == 0) {
          // input gradient: dx[i,j] = sum_t d[i-di, j-dj] w[t]: plane row rr of d feeds dx rows rr-2 (tap row 0), rr-1 (1), rr (2)
          float4 w[9];
#pragma unroll
          for (int k = 0; k < 9; ++k) w[k] = *reinterpret_cast<const float4*>(p.w + (z * 9 + k) * D + c0 + c);
          const float* pm = dpl + j * CH + c;
          float* dst = dxs + j * CH + c;
          float4 acc0 = f4z(), acc1 = f4z(), acc2 = f4z();
#pragma unroll
          for (int rr = 0; rr < TRC + 2; ++rr) {
            const float4 dm = *reinterpret_cast<const float4*>(pm + rr * rowf);
            const float4 dl = lok ? *reinterpret_cast<const float4*>(pm + rr * rowf - CH) : f4z();     // d[., j-1] pairs with tap column 2
            const float4 dr = rok ? *reinterpret_cast<const float4*>(pm + rr * rowf + CH) : f4z();     // d[., j+1] pairs with tap column 0
            fma4(acc0, dl, w[2]); fma4(acc0, dm, w[1]); fma4(acc0, dr, w[0]);       // d row = dx row + 1  -> tap row 0
            fma4(acc1, dl, w[5]); fma4(acc1, dm, w[4]); fma4(acc1, dr, w[3]);       // d row = dx row      -> tap row 1
            fma4(acc2, dl, w[8]); fma4(acc2, dm, w[7]); fma4(acc2, dr, w[6]);       // d row = dx row - 1  -> tap row 2
            if (rr >= 2) {
              float4* q = reinterpret_cast<float4*>(dst + (rr - 2) * rowf);
              if (z == 0) *q = acc0; else { float4 o = *q; o.x += acc0.x; o.y += acc0.y; o.z += acc0.z; o.w += acc0.w; *q = o; }
            }
            acc0 = acc1; acc1 = acc2; acc2 = f4z();
          }
        } else if (p.ddw) {
          // weight gradient: dw[t] += x[i+di, j+dj] d[i,j]: plane row rr of x pairs with the centre d of plane rows rr+1 (tap row 0), rr (1), rr-1 (2)
          float4 aw[9];
#pragma unroll
          for (int k = 0; k < 9; ++k) aw[k] = f4z();
          const float* px = xpl + j * CH + c;
          const float* pd = dpl + j * CH + c;
          float4 dprev = f4z(), dcur = f4z();
          float4 dnext = (0 < rend) ? *reinterpret_cast<const float4*>(pd + 1 * rowf) : f4z();      // d of band row 0 = plane row 1
#pragma unroll
          for (int rr = 0; rr < TRC + 2; ++rr) {
            // dprev = d[plane row rr-1], dcur = d[plane row rr], dnext = d[plane row rr+1]; only band rows (plane rows 1..rend) count
            const float4 xm = *reinterpret_cast<const float4*>(px + rr * rowf);
            const float4 xl = lok ? *reinterpret_cast<const float4*>(px + rr * rowf - CH) : f4z();
            const float4 xr = rok ? *reinterpret_cast<const float4*>(px + rr * rowf + CH) : f4z();
            fma4(aw[0], xl, dnext); fma4(aw[1], xm, dnext); fma4(aw[2], xr, dnext);
            fma4(aw[3], xl, dcur);  fma4(aw[4], xm, dcur);  fma4(aw[5], xr, dcur);
            fma4(aw[6], xl, dprev); fma4(aw[7], xm, dprev); fma4(aw[8], xr, dprev);
            dprev = dcur; dcur = dnext;
            dnext = (rr + 2 >= 1 && rr + 2 <= rend) ? *reinterpret_cast<const float4*>(pd + (rr + 2) * rowf) : f4z();
          }
#pragma unroll
          for (int k = 0; k < 9; ++k) *reinterpret_cast<float4*>(sitem + (k * W + j) * CH + c) = aw[k];
        }
      }
      __syncthreads();                   // this d band may be overwritten; sitem is complete
      if (z == 0) { load_band(dpl0, p.dconv + 2 * p.bs + base, D, H, W, r0, TRC, CH); cp_async_commit(); }
      if (p.ddw) {
        for (int o = tid; o < 9 * CH; o += NT) {
          const int k = o / CH, c = o - k * CH;
          float a = 0.f;
          for (int j = 0; j < W; ++j) a += sitem[(k * W + j) * CH + c];
          sdw[(z * 9 + k) * D + c0 + c] += a;
        }
      }
    }
    // dx band -> global, coalesced float4s
    const int total = rend * W * q4n;
    for (int i = tid; i < total; i += NT) {
      const int rr = i / (W * q4n), rem = i - rr * (W * q4n), col = rem / q4n, q = rem - col * q4n;
      *reinterpret_cast<float4*>(p.dxn + base + ((long long)(r0 + rr) * W + col) * D + q * 4) =
          *reinterpret_cast<const float4*>(dxs + (rr * W + col) * CH + q * 4);
    }
  }
  if (p.has_cls) {   // the cls row bypassed conv and BN: its gradient is the plain sum of the three branches
    for (int i = blockIdx.x * NT + tid; i < p.B * (D >> 2); i += gridDim.x * NT) {
      const int q = i % (D >> 2), b = i / (D >> 2);
      const long long row = (long long)b * p.Tc * D + q * 4;
      float4 s = f4z();
#pragma unroll
      for (int z = 0; z < 3; ++z) { const float4 v = *reinterpret_cast<const float4*>(p.dconv + z * p.bs + row); s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
      *reinterpret_cast<float4*>(p.dxn + row) = s;
    }
  }
  if (p.ddw) {
    __syncthreads();
    for (int i = tid; i < 27 * D; i += NT) p.ddw_part[(long long)blockIdx.x * 27 * D + i] = sdw[i];
  }
}

bool make_band(int B, int H, int W, int D, Band* bd) {
  if (D % 16 != 0) return false;
  int CH = 512 / W; CH = CH / 16 * 16; if (CH < 16) CH = 16; if (CH > D) CH = D;
  while (D % CH != 0) CH -= 16;
  const int TR = TRC;
  bd->CH = CH; bd->TR = TR; bd->nchunk = D / CH; bd->nband = cdiv(H, TR); bd->ntiles = B * bd->nband * bd->nchunk;
  return true;
}

}  // namespace

// reduce_partials_kernel lives in dwproj.cu
cudaError_t launch_reduce_partials(const float* part, int nparts, int n, float* dst, cudaStream_t s);

cudaError_t launch_dwplane_fwd(const DwProjP& p, cudaStream_t s) {
  Band bd;
  if (p.method[0] != 0 || p.method[1] != 0 || p.method[2] != 0 || !make_band(p.B, p.H, p.W, p.D, &bd)) return cudaErrorNotSupported;
  const size_t smem = ((size_t)(bd.TR + 2) * p.W * bd.CH + 6 * p.D + (size_t)6 * p.W * bd.CH) * sizeof(float);
  if (smem > 200 * 1024) return cudaErrorNotSupported;
  static bool configured = false;
  if (!configured) { CVT_CUDA(cudaFuncSetAttribute(dwplane_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); configured = true; }
  const int grid = bd.ntiles < p.nparts ? bd.ntiles : p.nparts;
  if (p.stats && grid < p.nparts) CVT_CUDA(cudaMemsetAsync(p.stats + (size_t)grid * 6 * p.D, 0, (size_t)(p.nparts - grid) * 6 * p.D * sizeof(float), s));
  dwplane_fwd_kernel<<<grid, NT, smem, s>>>(p, bd);
  CVT_LAUNCH_CHECK();
  return cudaSuccess;
}

cudaError_t launch_dwplane_bwd(const DwProjBwdP& p, cudaStream_t s) {
  Band bd;
  if (p.method[0] != 0 || p.method[1] != 0 || p.method[2] != 0 || p.coef || !make_band(p.B, p.H, p.W, p.D, &bd)) return cudaErrorNotSupported;
  if (p.ddw && !p.ddw_part) return cudaErrorInvalidValue;
  const size_t smem = ((size_t)3 * (bd.TR + 2) * p.W * bd.CH + (size_t)bd.TR * p.W * bd.CH + 27 * p.D + (size_t)9 * p.W * bd.CH) * sizeof(float);
  if (smem > 200 * 1024) return cudaErrorNotSupported;
  static bool configured = false;
  if (!configured) { CVT_CUDA(cudaFuncSetAttribute(dwplane_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); configured = true; }
  const int cap = 2 * kNumSMs;
  const int grid = bd.ntiles < cap ? bd.ntiles : cap;
  dwplane_bwd_kernel<<<grid, NT, smem, s>>>(p, bd);
  CVT_LAUNCH_CHECK();
  if (p.ddw) CVT_CUDA(launch_reduce_partials(p.ddw_part, grid, 27 * p.D, p.ddw, s));
  return cudaSuccess;
}

}  // namespace cvt
