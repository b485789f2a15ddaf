// tcgen05 / TMEM contraction kernel (CVT_PREC_TF32): the tensor-core twin of gemm_simt.cu, same GemmP description.
//
//   tile       128 (M) x BN (64|128|256) x 32 (K, one 128-byte swizzle line of fp32), UMMA 128xBNx8 kind::tf32
//   operands   staged by all 128 threads global -> registers -> 128B-swizzled shared memory, so the A-side
//              prologues (im2col gather, BatchNorm affine, cls-row remap) cost nothing extra.  No transposes:
//              an operand that is contiguous along K is written K-major, one contiguous along M/N is written
//              MN-major, and the UMMA descriptors carry the major-ness.
//   pipeline   S smem stages; thread 0 issues the MMAs and commits them to the stage's mbarrier (frees the slot);
//              register prefetch of the next K-tile overlaps the tensor core.
//   epilogue   accumulator TMEM -> registers (tcgen05.ld 32x32b) -> per-warp smem transpose -> coalesced
//              bias/GELU/mask/residual/atomic/scatter epilogue (identical arithmetic to the FFMA kernel).
#include "gemm_fetch.cuh"
#include <cstdio>
#include <cstdlib>

#include "gemm_epi.cuh"
#include "model.cuh"
#include "tc_common.cuh"

namespace cvt {
using namespace fetch;
using namespace tc;

namespace {
constexpr int TBM = 128, TBK = 32, A_TILE_BYTES = TBM * TBK * 4, NT = 256;

template <int BN> struct Cfg {
  static constexpr int NS = BN == 64 ? 4 : 3;                       // pipeline slots: 96 KB either way -> 2 CTAs per SM
  static constexpr int B_TILE_BYTES = BN * TBK * 4;
  static constexpr int SMEM = 1024 + NS * (A_TILE_BYTES + B_TILE_BYTES) + 17 * 8 + 16;
};

// ---- operand sources: where the 4 floats of one 16-byte shared-memory chunk come from ------------------------------------------
// n = number of valid floats (0..4) at p; n == -1: the four floats are not contiguous in memory (im2col with cin % 4 != 0)
struct Src { const float* p; int n; };

template <int AM>
__device__ __forceinline__ Src srcA_row(const GemmP& p, const float* A, int k, int kend, bool mvalid, long long tok, int cb, int ci_, int cj) {
  if (!mvalid || k >= kend) return Src{A, 0};
  if (AM == A_PLAIN) return Src{A + tok * p.lda + k, min(4, kend - k)};
  const ConvGeom& g = p.cv;
  if ((g.cin & 3) != 0) return Src{A, -1};
  const int tap = k / g.cin, c = k - tap * g.cin;
  const int kh = tap / g.k, kw = tap - kh * g.k;
  const int yy = g.s * ci_ + kh - g.pt, xx = g.s * cj + kw - g.pl;
  if (yy < 0 || yy >= g.hin || xx < 0 || xx >= g.win) return Src{A, 0};
  return Src{A + (long long)cb * g.bstride + g.boff + ((long long)yy * g.win + xx) * g.cin + c, 4};
}
template <int AM>
__device__ __forceinline__ Src srcA_col(const GemmP& p, const float* A, int k, int kend, int m0) {
  if (k >= kend || m0 >= p.M) return Src{A, 0};
  if (AM == A_TRANS) return Src{A + p.a_map(k) * p.lda + m0, min(4, p.M - m0)};
  const ConvGeom& g = p.cv;
  if ((g.cin & 3) != 0) return Src{A, -1};
  const int hw = g.ho * g.wo;
  const int b = k / hw, r = k - b * hw;
  const int oi = r / g.wo, oj = r - oi * g.wo;
  const int tap = m0 / g.cin, c = m0 - tap * g.cin;
  const int kh = tap / g.k, kw = tap - kh * g.k;
  const int yy = g.s * oi + kh - g.pt, xx = g.s * oj + kw - g.pl;
  if (yy < 0 || yy >= g.hin || xx < 0 || xx >= g.win) return Src{A, 0};
  return Src{A + (long long)b * g.bstride + g.boff + ((long long)yy * g.win + xx) * g.cin + c, 4};
}
__device__ __forceinline__ Src srcB_kn(const GemmP& p, const float* B, int k, int kend, int n0) {
  if (k >= kend || n0 >= p.N) return Src{B, 0};
  return Src{B + p.b_map(k) * p.ldb + n0, min(4, p.N - n0)};
}
__device__ __forceinline__ Src srcB_nk(const GemmP& p, const float* B, int n, int k, int kend) {
  if (n >= p.N || k >= kend) return Src{B, 0};
  return Src{B + (long long)n * p.ldb + k, min(4, kend - k)};
}

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// 16-byte chunk: asynchronous copy when the source is a full aligned float4, else assembled through registers
__device__ __forceinline__ void copy_chunk(uint8_t* dst, const Src& s) {
  if (s.n == 4 && aligned16(s.p)) { cp_async16(smem_u32(dst), s.p); return; }
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (s.n > 0) { v.x = s.p[0]; if (s.n > 1) v.y = s.p[1]; if (s.n > 2) v.z = s.p[2]; if (s.n > 3) v.w = s.p[3]; }
  *reinterpret_cast<float4*>(dst) = v;
}

template <int AM, int BMODE, int BN, uint32_t E>
__global__ void __launch_bounds__(NT, 2) gemm_tc_kernel(const __grid_constant__ GemmP p) {
  pdl_wait(); pdl_trigger();
  using C = Cfg<BN>;
  constexpr int NS = C::NS;
  constexpr bool A_ROWFIXED = (AM == A_PLAIN || AM == A_CONV);     // K-contiguous A  -> K-major
  constexpr bool B_KMAJOR = (BMODE == B_NK);
  constexpr int NA_FRAGS = TBM * 8 / NT;       // 4
  constexpr int NB_FRAGS = BN * 8 / NT;        // 2 | 4
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sA = smem;
  uint8_t* sB = sA + NS * A_TILE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + NS * C::B_TILE_BYTES);      // [NS] stage-free, [1] accumulator-ready
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int zz = blockIdx.z;
  const int z = zz / p.splitk, ks = zz - z * p.splitk;
  const int zb = z / p.bdiv, zh = z - zb * p.bdiv;
  const float* A = p.A + zb * p.a_bs0 + zh * p.a_bs1;
  const float* B = p.B + zb * p.b_bs0 + zh * p.b_bs1;
  const float* sc = p.a_scale ? p.a_scale + z * p.a_aff_bs : nullptr;
  const float* sh = p.a_shift ? p.a_shift + z * p.a_aff_bs : nullptr;
  const int m_blk = blockIdx.y * TBM, n_blk = blockIdx.x * BN;
  const int ktiles = cdiv(p.K, TBK);
  const int tiles_per = cdiv(ktiles, p.splitk);
  const int kbeg = ks * tiles_per * TBK;
  const int kend = min(p.K, kbeg + tiles_per * TBK);
  if (kbeg >= kend) return;                                           // empty split (whole CTA exits together)
  const int nk = cdiv(kend - kbeg, TBK);

  if (warp == 0) tmem_alloc(tmem_slot, BN);
  if (tid == 32) {
    for (int i = 0; i <= NS; ++i) mbar_init(&bars[i], 1);
    mbar_fence_init();
  }

  // ---- per-thread operand coordinates ----
  // A, K-major:  chunk = tid & 7 (16 B along K), rows (tid >> 3) + 32 i
  // A, MN-major: m4 = tid & 31 (float4 along M),  k rows (tid >> 5) + 8 i
  long long a_tok[NA_FRAGS]; int a_cb[NA_FRAGS], a_cij[NA_FRAGS]; uint32_t a_valid = 0, a_skip = 0;
  if (A_ROWFIXED) {
#pragma unroll
    for (int i = 0; i < NA_FRAGS; ++i) {
      const int m = m_blk + (tid >> 3) + 32 * i;
      a_tok[i] = 0; a_cb[i] = 0; a_cij[i] = 0;
      if (m < p.M) {
        a_valid |= 1u << i;
        if (AM == A_PLAIN) {
          a_tok[i] = p.a_map(m);
          if (p.a_skip_period && ((int)a_tok[i] % p.a_skip_period) == 0) a_skip |= 1u << i;
        } else {
          const int hw = p.cv.ho * p.cv.wo; const int b = m / hw; const int r = m - b * hw; const int oi = r / p.cv.wo;
          a_cb[i] = b; a_cij[i] = (oi << 16) | (r - oi * p.cv.wo);
        }
      }
    }
  }
  auto a_off = [&](int i) -> uint32_t {
    return A_ROWFIXED ? sw128_kmajor_off((tid >> 3) + 32 * i, tid & 7) : sw128_mnmajor_off((tid >> 5) + 8 * i, tid & 31, TBM / 32);
  };
  auto b_off = [&](int i) -> uint32_t {
    const int idx = tid + NT * i;
    return B_KMAJOR ? sw128_kmajor_off(idx >> 3, idx & 7) : sw128_mnmajor_off(idx / (BN / 4), idx % (BN / 4), BN / 32);
  };
  // im2col element gather for Conv2D inputs whose channel count is not a multiple of 4 (stage 1: cin = 1)
  auto gather_conv = [&](int m, int k, bool k_is_pixel) -> float {
    // forward (A_CONV): m = output pixel, k = (tap, c);  wgrad (A_CONVT): k = output pixel, m = (tap, c)
    const ConvGeom& g = p.cv;
    const int pix = k_is_pixel ? k : m, tc = k_is_pixel ? m : k;
    const int hw = g.ho * g.wo; const int b = pix / hw, r = pix - b * hw; const int oi = r / g.wo, oj = r - oi * g.wo;
    const int tap = tc / g.cin, c = tc - tap * g.cin; const int kh = tap / g.k, kw = tap - kh * g.k;
    const int yy = g.s * oi + kh - g.pt, xx = g.s * oj + kw - g.pl;
    if (yy < 0 || yy >= g.hin || xx < 0 || xx >= g.win) return 0.f;
    return A[(long long)b * g.bstride + g.boff + ((long long)yy * g.win + xx) * g.cin + c];
  };
  auto issue_tile = [&](int slot, int k0) {
    uint8_t* a = sA + slot * A_TILE_BYTES; uint8_t* b = sB + slot * C::B_TILE_BYTES;
#pragma unroll
    for (int i = 0; i < NA_FRAGS; ++i) {
      Src s;
      if (A_ROWFIXED) {
        const int m = m_blk + (tid >> 3) + 32 * i, k = k0 + ((tid & 7) << 2);
        s = srcA_row<AM>(p, A, k, kend, (a_valid >> i) & 1, a_tok[i], a_cb[i], a_cij[i] >> 16, a_cij[i] & 0xFFFF);
        if (AM == A_CONV && s.n < 0) {
          float4 v;
          v.x = (k < kend) ? gather_conv(m, k, false) : 0.f; v.y = (k + 1 < kend) ? gather_conv(m, k + 1, false) : 0.f;
          v.z = (k + 2 < kend) ? gather_conv(m, k + 2, false) : 0.f; v.w = (k + 3 < kend) ? gather_conv(m, k + 3, false) : 0.f;
          *reinterpret_cast<float4*>(a + a_off(i)) = v;
          continue;
        }
      } else {
        const int k = k0 + (tid >> 5) + 8 * i, m0 = m_blk + ((tid & 31) << 2);
        s = srcA_col<AM>(p, A, k, kend, m0);
        if (AM == A_CONVT && s.n < 0) {
          float4 v;
          v.x = (m0 < p.M) ? gather_conv(m0, k, true) : 0.f; v.y = (m0 + 1 < p.M) ? gather_conv(m0 + 1, k, true) : 0.f;
          v.z = (m0 + 2 < p.M) ? gather_conv(m0 + 2, k, true) : 0.f; v.w = (m0 + 3 < p.M) ? gather_conv(m0 + 3, k, true) : 0.f;
          *reinterpret_cast<float4*>(a + a_off(i)) = v;
          continue;
        }
      }
      copy_chunk(a + a_off(i), s);
    }
#pragma unroll
    for (int i = 0; i < NB_FRAGS; ++i) {
      const int idx = tid + NT * i;
      const Src s = B_KMAJOR ? srcB_nk(p, B, n_blk + (idx >> 3), k0 + ((idx & 7) << 2), kend)
                             : srcB_kn(p, B, k0 + idx / (BN / 4), kend, n_blk + ((idx % (BN / 4)) << 2));
      copy_chunk(b + b_off(i), s);
    }
  };
  // second pass over the chunks this thread copied: optional per-feature affine on A (BatchNorm applied on load), then round to TF32
  auto convert_tile = [&](int slot, int k0) {
    uint8_t* a = sA + slot * A_TILE_BYTES; uint8_t* b = sB + slot * C::B_TILE_BYTES;
#pragma unroll
    for (int i = 0; i < NA_FRAGS; ++i) {
      float4* q = reinterpret_cast<float4*>(a + a_off(i));
      float4 v = *q;
      if ((AM == A_PLAIN || AM == A_TRANS) && sc != nullptr) {
        if (A_ROWFIXED) {
          const int k = k0 + ((tid & 7) << 2);
          if (((a_valid >> i) & 1) && !((a_skip >> i) & 1) && k < kend) {
            // K is a multiple of 4 whenever the affine is used (feature counts are multiples of 64)
            const float4 s4 = *reinterpret_cast<const float4*>(sc + k), h4 = *reinterpret_cast<const float4*>(sh + k);
            v.x = fmaf(v.x, s4.x, h4.x); v.y = fmaf(v.y, s4.y, h4.y); v.z = fmaf(v.z, s4.z, h4.z); v.w = fmaf(v.w, s4.w, h4.w);
          }
        } else {
          const int k = k0 + (tid >> 5) + 8 * i, m0 = m_blk + ((tid & 31) << 2);
          if (k < kend && m0 < p.M) {
            const int tok = (int)p.a_map(k);
            if (!(p.a_skip_period && (tok % p.a_skip_period) == 0)) {
              const float4 s4 = *reinterpret_cast<const float4*>(sc + m0), h4 = *reinterpret_cast<const float4*>(sh + m0);
              v.x = fmaf(v.x, s4.x, h4.x); v.y = fmaf(v.y, s4.y, h4.y); v.z = fmaf(v.z, s4.z, h4.z); v.w = fmaf(v.w, s4.w, h4.w);
            }
          }
        }
      }
      *q = make_float4(to_tf32(v.x), to_tf32(v.y), to_tf32(v.z), to_tf32(v.w));
    }
#pragma unroll
    for (int i = 0; i < NB_FRAGS; ++i) {
      float4* q = reinterpret_cast<float4*>(b + b_off(i));
      const float4 v = *q;
      *q = make_float4(to_tf32(v.x), to_tf32(v.y), to_tf32(v.z), to_tf32(v.w));
    }
  };

  // ---- pipeline: NS-1 k-tiles of cp.async in flight ahead of the tensor core ----
#pragma unroll 1
  for (int t = 0; t < NS - 1; ++t) {
    if (t < nk) issue_tile(t, kbeg + t * TBK);
    cp_async_commit();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  constexpr uint32_t idesc = make_idesc(kFmtTF32, TBM, BN, A_ROWFIXED ? 0u : 1u, B_KMAJOR ? 0u : 1u);
  for (int kt = 0; kt < nk; ++kt) {
    const int s = kt % NS;
    {  // refill the slot tile kt-1 used with tile kt+NS-1 once its MMAs have completed
      const int nt = kt + NS - 1;
      if (nt < nk) {
        if (kt > 0) mbar_wait(&bars[(kt - 1) % NS], ((kt - 1) / NS) & 1);
        issue_tile(nt % NS, kbeg + nt * TBK);
      }
      cp_async_commit();
    }
    cp_async_wait<NS - 1>();                                          // this thread's copies of tile kt have landed
    convert_tile(s, kbeg + kt * TBK);
    fence_proxy_async_smem();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const uint32_t a0 = smem_u32(sA + s * A_TILE_BYTES), b0 = smem_u32(sB + s * C::B_TILE_BYTES);
#pragma unroll
      for (int j = 0; j < TBK / 8; ++j) {
        const uint64_t ad = A_ROWFIXED ? make_smem_desc(a0 + j * 32, 16, 1024)
                                       : make_smem_desc(a0 + j * (TBM / 32) * 1024, 512, (TBM / 32) * 512, kLayoutSw128Base32);
        const uint64_t bd = B_KMAJOR ? make_smem_desc(b0 + j * 32, 16, 1024)
                                     : make_smem_desc(b0 + j * (BN / 32) * 1024, 512, (BN / 32) * 512, kLayoutSw128Base32);
        mma_tf32_ss(tmem_base, ad, bd, idesc, (kt > 0 || j > 0) ? 1u : 0u);
      }
      mma_commit(&bars[s]);
      if (kt == nk - 1) mma_commit(&bars[NS]);
    }
  }
  mbar_wait(&bars[NS], 0);
  tc_fence_after();

  // ---- epilogue: warp w reads TMEM lane quadrant w & 3, column half w >> 2 (one accumulator row per thread), transposes each
  //      32x32 block through its own staging tile (the pipeline slots are free now) and finishes in the coalesced domain ----
  float* Cp = p.C + zb * p.c_bs0 + zh * p.c_bs1;
  float* aux = p.aux ? p.aux + zb * p.c_bs0 + zh * p.c_bs1 : nullptr;
  const float* bias = (p.epi & EPI_BIAS) ? p.bias + z * p.bias_bs : nullptr;
  const float alpha = ((p.epi & EPI_ALPHA0) && z == 0) ? p.alpha0 : p.alpha;
  const int quad = warp & 3, half = warp >> 2;
  const uint32_t trow = tmem_base + ((uint32_t)(quad * 32) << 16);
  float* stg = reinterpret_cast<float*>(smem) + warp * (32 * 36);    // 32 rows x 36 floats (16-byte aligned rows, conflict-free both ways)
  const int er = lane >> 3, ec = (lane & 7) << 2;
#pragma unroll 1
  for (int c0 = half * (BN / 2); c0 < (half + 1) * (BN / 2); c0 += 32) {
    float v[32];
    tmem_ld32(trow + c0, v);
#pragma unroll
    for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(stg + lane * 36 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    __syncwarp();
#pragma unroll 2
    for (int i = 0; i < 8; ++i) {
      const int rr = er + 4 * i;
      const int m = m_blk + quad * 32 + rr;
      const float4 acc = *reinterpret_cast<const float4*>(stg + rr * 36 + ec);
      if (m < p.M) epilogue_f4<E>(p, acc, Cp, aux, bias, alpha, m, n_blk + c0 + ec);
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, BN);
}

template <int AM, int BMODE, int BN, uint32_t E>
cudaError_t launch_cfg(const GemmP& p, cudaStream_t s) {
  static bool configured = false;
  if (!configured) {
    CVT_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<AM, BMODE, BN, E>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM));
    configured = true;
  }
  dim3 grid(cdiv(p.N, BN), cdiv(p.M, TBM), p.batch * p.splitk);
  CVT_LAUNCH((gemm_tc_kernel<AM, BMODE, BN, E>), grid, NT, Cfg<BN>::SMEM, s, p);
  return cudaSuccess;
}

template <int AM, int BMODE, uint32_t E>
cudaError_t launch_bn(const GemmP& p, int bn, cudaStream_t s) {
  if (bn == 128) return launch_cfg<AM, BMODE, 128, E>(p, s);
  return launch_cfg<AM, BMODE, 64, E>(p, s);
}

constexpr uint32_t kEpiKinds = EPI_GELU | EPI_RESID | EPI_GELU_BWD | EPI_SCATTER | EPI_ATOMIC;

int pick_bn(const GemmP& p) {
  if (p.N % 128 != 0) return 64;
  const long long ctas128 = (long long)cdiv(p.M, TBM) * p.batch * p.splitk * (p.N / 128);
  return ctas128 >= 2 * kNumSMs ? 128 : 64;
}
}  // namespace

bool gemm_tc_supports(const GemmP& p) {
  if (p.epi & EPI_PHILOX) return false;          // in-epilogue dropout masks are instantiated for the TMA kernel only
  if (p.N % 64 != 0 || p.M < 1 || p.K < 1) return false;
  if (p.epi & (EPI_RELU | EPI_RELU_BWD)) return false;              // the tiny head layers stay on the FFMA kernel
  if (p.splitk > 1 && !(p.epi & (EPI_ATOMIC | EPI_SCATTER))) return false;
  return true;
}

cudaError_t launch_gemm_tc(const GemmP& p, cudaStream_t s) {
  if (!gemm_tc_supports(p)) { set_error("gemm_tc: unsupported shape / epilogue"); return cudaErrorNotSupported; }
  const int bn = pick_bn(p);
  const uint32_t kind = p.epi & kEpiKinds;
  const int modes = p.a_mode * 2 + p.b_mode;
  // the instantiations the CvT block uses (cvt_api.cu); anything else is refused loudly rather than silently slow
  if (modes == A_PLAIN * 2 + B_KN) {            // Dense / EinsumDense forward
    if (kind == 0) return launch_bn<A_PLAIN, B_KN, 0>(p, bn, s);
    if (kind == EPI_RESID) return launch_bn<A_PLAIN, B_KN, EPI_RESID>(p, bn, s);
    if (kind == EPI_GELU) return launch_bn<A_PLAIN, B_KN, EPI_GELU>(p, bn, s);
    if (kind == EPI_ATOMIC) return launch_bn<A_PLAIN, B_KN, EPI_ATOMIC>(p, bn, s);
  } else if (modes == A_PLAIN * 2 + B_NK) {     // dgrad
    if (kind == 0) return launch_bn<A_PLAIN, B_NK, 0>(p, bn, s);
    if (kind == EPI_GELU_BWD) return launch_bn<A_PLAIN, B_NK, EPI_GELU_BWD>(p, bn, s);
    if (kind == EPI_SCATTER) return launch_bn<A_PLAIN, B_NK, EPI_SCATTER>(p, bn, s);
    if (kind == EPI_ATOMIC) return launch_bn<A_PLAIN, B_NK, EPI_ATOMIC>(p, bn, s);
  } else if (modes == A_TRANS * 2 + B_KN) {     // wgrad
    if (kind == EPI_ATOMIC) return launch_bn<A_TRANS, B_KN, EPI_ATOMIC>(p, bn, s);
    if (kind == 0) return launch_bn<A_TRANS, B_KN, 0>(p, bn, s);
  } else if (modes == A_TRANS * 2 + B_NK) {
    if (kind == 0) return launch_bn<A_TRANS, B_NK, 0>(p, bn, s);
    if (kind == EPI_ATOMIC) return launch_bn<A_TRANS, B_NK, EPI_ATOMIC>(p, bn, s);
  } else if (modes == A_CONV * 2 + B_KN) {      // Conv2D forward (implicit GEMM)
    if (kind == 0) return launch_bn<A_CONV, B_KN, 0>(p, bn, s);
  } else if (modes == A_CONVT * 2 + B_KN) {     // Conv2D wgrad
    if (kind == EPI_ATOMIC) return launch_bn<A_CONVT, B_KN, EPI_ATOMIC>(p, bn, s);
  }
  set_error("gemm_tc: no instantiation for a_mode " + std::to_string(p.a_mode) + " b_mode " + std::to_string(p.b_mode) + " epilogue " + std::to_string(p.epi));
  return cudaErrorNotSupported;
}

}  // namespace cvt

// diagnostics: one raw contraction through either kernel (tests/test_gpu_gemm.py)
extern "C" CVT_API int cvt_debug_gemm(int a_mode, int b_mode, int M, int N, int K, int splitk, int use_tc, const float* A, const float* B,
                                      float* C) {
  using namespace cvt;
  if (a_mode > A_TRANS || a_mode < 0 || b_mode < 0 || b_mode > B_NK || M < 1 || N < 1 || K < 1 || splitk < 1) { set_error("cvt_debug_gemm: bad argument"); return CVT_ERR_INVALID; }
  float *dA = nullptr, *dB = nullptr, *dC = nullptr;
  const size_t na = (size_t)M * K, nb = (size_t)K * N, nc = (size_t)M * N;
  auto fail = [&](cudaError_t e) { set_error(std::string("cvt_debug_gemm: ") + cudaGetErrorString(e)); cudaFree(dA); cudaFree(dB); cudaFree(dC); return (int)CVT_ERR_CUDA; };
  cudaError_t e;
  if ((e = cudaMalloc(&dA, na * 4)) != cudaSuccess || (e = cudaMalloc(&dB, nb * 4)) != cudaSuccess || (e = cudaMalloc(&dC, nc * 4)) != cudaSuccess) return fail(e);
  if ((e = cudaMemcpy(dA, A, na * 4, cudaMemcpyHostToDevice)) != cudaSuccess || (e = cudaMemcpy(dB, B, nb * 4, cudaMemcpyHostToDevice)) != cudaSuccess ||
      (e = cudaMemset(dC, 0, nc * 4)) != cudaSuccess) return fail(e);
  GemmP p; p.M = M; p.N = N; p.K = K; p.splitk = splitk;
  p.A = dA; p.a_mode = a_mode; p.lda = a_mode == A_PLAIN ? K : M;
  p.B = dB; p.b_mode = b_mode; p.ldb = b_mode == B_KN ? N : K;
  p.C = dC; p.ldc = N; if (splitk > 1) p.epi = EPI_ATOMIC;
  p.dbg = (use_tc >> 1) & 1;
  const int x3 = (use_tc >> 4) & 1; use_tc &= 15; p.x3 = x3;      // bit 4: the error-compensated (3 x TF32) variant of the TMA-fed kernels
  if (use_tc & 8) {      // TMA-fed weight-gradient kernel: A is X[tokens=K, Kin=M], B is dY[tokens=K, N]
    WgradP w; w.X = dA; w.ldx = M; w.Kin = M; w.dY = dB; w.lddy = N; w.N = N; w.tokens = K; w.dW = dC; w.dbg = use_tc & 1; w.x3 = x3;
    if (a_mode != A_TRANS || b_mode != B_KN || !wgrad_tma_supports(w)) { cudaFree(dA); cudaFree(dB); cudaFree(dC); set_error("cvt_debug_gemm: shape not supported by the wgrad kernel"); return CVT_ERR_INVALID; }
    e = launch_wgrad_tma(w, 0);
    if (e != cudaSuccess) return fail(e);
    if ((e = cudaDeviceSynchronize()) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpy(C, dC, nc * 4, cudaMemcpyDeviceToHost)) != cudaSuccess) return fail(e);
    cudaFree(dA); cudaFree(dB); cudaFree(dC);
    return CVT_OK;
  }
  if (use_tc == 4) {
    if (!gemm_tma_supports(p)) { cudaFree(dA); cudaFree(dB); cudaFree(dC); set_error("cvt_debug_gemm: shape not supported by the TMA kernel"); return CVT_ERR_INVALID; }
    e = launch_gemm_tma(p, 0);
    if (e != cudaSuccess) return fail(e);
    if (getenv("CVT_GEMM_TRACE")) {                      // second (warm) launch with the clock64 timeline of CTA 0
      long long* tr = nullptr; static long long h[8 + 30 * 8];
      cudaMalloc(&tr, sizeof(h)); cudaMemset(tr, 0, sizeof(h));
      p.trace = tr; e = launch_gemm_tma(p, 0); p.trace = nullptr;
      cudaDeviceSynchronize(); cudaMemcpy(h, tr, sizeof(h), cudaMemcpyDeviceToHost); cudaFree(tr);
      printf("gemm_tma trace M=%d N=%d K=%d: start 0, setup done %lld, exit %lld (cycles)\n", M, N, K, h[1] - h[0], h[2] - h[0]);
      printf("tile | producer: slot issued | mma: acc_free full commit | epilogue: wait_begin acc_full done\n");
      for (int i = 0; i < 30 && h[8 + i * 8 + 1]; ++i) { printf("%3d |", i); for (int k = 0; k < 8; ++k) printf(" %7lld", h[8 + i * 8 + k] ? h[8 + i * 8 + k] - h[0] : -1); printf("\n"); }
    }
    if ((e = cudaDeviceSynchronize()) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpy(C, dC, nc * 4, cudaMemcpyDeviceToHost)) != cudaSuccess) return fail(e);
    cudaFree(dA); cudaFree(dB); cudaFree(dC);
    return CVT_OK;
  }
  if (use_tc && !gemm_tc_supports(p)) { cudaFree(dA); cudaFree(dB); cudaFree(dC); set_error("cvt_debug_gemm: shape not supported by the tcgen05 kernel"); return CVT_ERR_INVALID; }
  e = use_tc ? launch_gemm_tc(p, 0) : launch_gemm_simt(p, 0);
  if (e != cudaSuccess) return fail(e);
  if ((e = cudaDeviceSynchronize()) != cudaSuccess) return fail(e);
  if ((e = cudaMemcpy(C, dC, nc * 4, cudaMemcpyDeviceToHost)) != cudaSuccess) return fail(e);
  cudaFree(dA); cudaFree(dB); cudaFree(dC);
  return CVT_OK;
}

