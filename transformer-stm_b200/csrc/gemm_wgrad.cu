// Split-K weight-gradient kernel on tcgen05 fed by TMA (CVT_PREC_TF32):
//     dW[Kin, N] += X[tokens, Kin]^T . dY[tokens, N]        db[N] += sum_tokens dY[tokens, N]
// (Dense / EinsumDense backward of the CvT block; the reference gets these from TF autodiff).
// The contraction runs over tokens, so both operands are "MN-major" for the tensor core: a TMA box of 32 features x 32 tokens
// lands in shared memory as 128-byte token rows with the 32-byte-atom 128B swizzle (CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B), which
// is exactly the canonical MN-major tf32 layout, so no transpose ever happens.  HBM-bound: every activation is read once.
//
//   grid       (output tile, batch z, token split): each CTA accumulates its token range for one 128 x BN tile of dW in TMEM
//   warp 0     TMA producer (4 A boxes + BN/32 B boxes per 32-token step)
//   warp 1     MMA issuer (4 x tcgen05.mma 128 x BN x 8 per step)
//   warps 2-5  column sums of the dY stage straight from shared memory (bias gradient, m-tile 0 only), then the epilogue:
//              TMEM -> smem transpose -> vector red.global.add into dW
#include "gemm.cuh"
#include "tc_common.cuh"
#include "tma_host.cuh"

namespace cvt {
using namespace tc;

namespace {
constexpr int BM = 128, KT = 32, BOX = 32 * KT * 4, NTHREADS = 192;      // BOX: one 32-feature x 32-token box = 4 KB

// X3: error-compensated variant (CVT_PREC_TF32X3), as in gemm_tma.cu: four more warps write the low parts of both operand tiles of a
// stage (x - trunc(x)) into twin tiles, three MMAs per step (hi.hi + lo.hi + hi.lo)
constexpr int NSPLITW = 4;
template <int BN, bool X3> struct WCfg {
  static constexpr int A_STAGE = (BM / 32) * BOX, B_STAGE = (BN / 32) * BOX;
  static constexpr int NS = X3 ? (BN == 64 ? 4 : (BN == 128 ? 3 : 2)) : (BN == 64 ? 8 : (BN == 128 ? 6 : 4));
  static constexpr int STG = 4 * 32 * 36 * 4;
  static constexpr int SMEM = 1024 + NS * (A_STAGE + B_STAGE) * (X3 ? 2 : 1) + STG + (3 * NS + 1) * 8 + 16;
  static constexpr int THREADS = NTHREADS + (X3 ? 32 * NSPLITW : 0);
  static_assert(SMEM <= 232448, "wgrad_tma: shared memory budget");
};

template <int BN, bool X3>
__global__ void __launch_bounds__(WCfg<BN, X3>::THREADS, 1) wgrad_tma_kernel(const __grid_constant__ WgradP p, const __grid_constant__ CUtensorMap tmX,
                                                                              const __grid_constant__ CUtensorMap tmY, int m_tiles, int n_tiles,
                                                                              int splits, int steps_per_split) {
  using C = WCfg<BN, X3>;
  constexpr int NS = C::NS;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sA = smem;
  uint8_t* sB = sA + NS * C::A_STAGE;
  uint8_t* sAlo = sB + NS * C::B_STAGE;
  uint8_t* sBlo = sAlo + (X3 ? NS * C::A_STAGE : 0);
  float* stg_all = reinterpret_cast<float*>(sBlo + (X3 ? NS * C::B_STAGE : 0));
  uint64_t* full = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(stg_all) + C::STG);
  uint64_t* empty = full + NS;
  uint64_t* ready = empty + NS;
  uint64_t* acc_full = ready + NS;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int unit = blockIdx.x;
  const int split = unit % splits; unit /= splits;
  const int nt = unit % n_tiles; unit /= n_tiles;
  const int mt = unit % m_tiles; const int z = unit / m_tiles;
  const int total_steps = cdiv(p.tokens, KT);
  const int step0 = split * steps_per_split;
  const int nsteps = min(steps_per_split, total_steps - step0);          // >= 1 by construction
  const bool do_colsum = p.db != nullptr && mt == 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX); tma_prefetch_desc(&tmY);
    for (int i = 0; i < NS; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], do_colsum ? 5 : 1); mbar_init(&ready[i], 32 * NSPLITW); }
    mbar_init(acc_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait(); pdl_trigger();      // everything above touches no global memory
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int it = 0; it < nsteps; ++it) {
        const int tok = (step0 + it) * KT;
        mbar_wait(&empty[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full[stage], C::A_STAGE + C::B_STAGE);
#pragma unroll
        for (int b = 0; b < BM / 32; ++b) {
          if (p.conv) {
            const ConvGeom& g = p.cv;
            const int f0 = mt * BM + b * 32, tap = f0 / g.cin, c0 = f0 - tap * g.cin, kh = tap / g.k, kw = tap - kh * g.k;
            const int hw = g.ho * g.wo, b0 = tok / hw, i0 = (tok - b0 * hw) / g.wo;
            tma_load_4d(sA + stage * C::A_STAGE + b * BOX, &tmX, &full[stage], c0, kw - g.pl, g.s * i0 + kh - g.pt, b0);
          } else
          tma_load_3d(sA + stage * C::A_STAGE + b * BOX, &tmX, &full[stage], mt * BM + b * 32, tok, z);
        }
#pragma unroll
        for (int b = 0; b < BN / 32; ++b) tma_load_3d(sB + stage * C::B_STAGE + b * BOX, &tmY, &full[stage], nt * BN + b * 32, tok, z);
        if (++stage == NS) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(kFmtTF32, BM, BN, 1u, 1u);
      const uint32_t lbo = (p.dbg & 1) ? 512u : (uint32_t)BOX, sbo = (p.dbg & 1) ? (uint32_t)BOX : 512u;
      int stage = 0; uint32_t phase = 0;
      for (int it = 0; it < nsteps; ++it) {
        mbar_wait(X3 ? &ready[stage] : &full[stage], phase);
        tc_fence_after();
        const uint32_t a0 = smem_u32(sA + stage * C::A_STAGE), b0 = smem_u32(sB + stage * C::B_STAGE);
#pragma unroll
        for (int j = 0; j < KT / 8; ++j) {    // 8 tokens = two 4-token groups of 512 bytes inside every box
          const uint64_t da = make_smem_desc(a0 + j * 1024, lbo, sbo, kLayoutSw128Base32), db = make_smem_desc(b0 + j * 1024, lbo, sbo, kLayoutSw128Base32);
          mma_tf32_ss(tmem_base, da, db, idesc, (it > 0 || j > 0) ? 1u : 0u);
          if (X3) {
            mma_tf32_ss(tmem_base, make_smem_desc(smem_u32(sAlo + stage * C::A_STAGE) + j * 1024, lbo, sbo, kLayoutSw128Base32), db, idesc, 1u);
            mma_tf32_ss(tmem_base, da, make_smem_desc(smem_u32(sBlo + stage * C::B_STAGE) + j * 1024, lbo, sbo, kLayoutSw128Base32), idesc, 1u);
          }
        }
        mma_commit(&empty[stage]);
        if (++stage == NS) { stage = 0; phase ^= 1; }
      }
      mma_commit(acc_full);
    }
  } else if (X3 && warp >= 6) {
    // ===== X3 splitter warps =====
    const int t = threadIdx.x - 32 * 6;
    int stage = 0; uint32_t phase = 0;
    for (int it = 0; it < nsteps; ++it) {
      mbar_wait(&full[stage], phase);
      auto split = [&](const uint8_t* src, uint8_t* dst, int n4) {
        for (int i = t; i < n4; i += 32 * NSPLITW) {
          const float4 v = reinterpret_cast<const float4*>(src)[i];
          float4 lo;
          // the low part is itself rounded to nearest TF32 here: left to the tensor core's truncation its error would be one-sided and
          // grow linearly with K
          lo.x = to_tf32(v.x - __uint_as_float(__float_as_uint(v.x) & 0xFFFFE000u)); lo.y = to_tf32(v.y - __uint_as_float(__float_as_uint(v.y) & 0xFFFFE000u));
          lo.z = to_tf32(v.z - __uint_as_float(__float_as_uint(v.z) & 0xFFFFE000u)); lo.w = to_tf32(v.w - __uint_as_float(__float_as_uint(v.w) & 0xFFFFE000u));
          reinterpret_cast<float4*>(dst)[i] = lo;
        }
      };
      split(sA + stage * C::A_STAGE, sAlo + stage * C::A_STAGE, C::A_STAGE / 16);
      split(sB + stage * C::B_STAGE, sBlo + stage * C::B_STAGE, C::B_STAGE / 16);
      fence_proxy_async_smem();
      mbar_arrive(&ready[stage]);
      if (++stage == NS) { stage = 0; phase ^= 1; }
    }
  } else {
    const int ew = warp - 2, et = ew * 32 + lane;          // 0..127
    if (do_colsum) {
      // thread et sums dY columns et, et + 128 (when BN = 256) of every stage: box = column / 32, 32-byte chunk swizzled by (token & 3)
      float cs[(BN + 127) / 128];
#pragma unroll
      for (int q = 0; q < (BN + 127) / 128; ++q) cs[q] = 0.f;
      int stage = 0; uint32_t phase = 0;
      for (int it = 0; it < nsteps; ++it) {
        mbar_wait(&full[stage], phase);
#pragma unroll
        for (int q = 0; q < (BN + 127) / 128; ++q) {
          const int col = et + q * 128;
          if (col < BN) {
            const uint8_t* box = sB + stage * C::B_STAGE + (col >> 5) * BOX;
            const int c = col & 31;
            float a = 0.f;
#pragma unroll 8
            for (int k = 0; k < KT; ++k) a += *reinterpret_cast<const float*>(box + k * 128 + ((((c >> 3) ^ (k & 3))) << 5) + ((c & 7) << 2));
            cs[q] += a;
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[stage]);
        if (++stage == NS) { stage = 0; phase ^= 1; }
      }
#pragma unroll
      for (int q = 0; q < (BN + 127) / 128; ++q) {
        const int col = et + q * 128;
        if (col < BN && nt * BN + col < p.N) atomicAdd(p.db + z * p.db_bs + nt * BN + col, cs[q]);
      }
    }
    // ---- epilogue: accumulator rows = Kin features of this m-tile, columns = N features ----
    const int quad = warp & 3;
    float* stg = stg_all + ew * (32 * 36);
    const int er = lane >> 3, ec = (lane & 7) << 2;
    float* Cp = p.dW + z * p.dw_bs;
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const uint32_t trow = tmem_base + ((uint32_t)(quad * 32) << 16);
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      float v[32];
      tmem_ld32(trow + c0, v);
#pragma unroll
      for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(stg + lane * 36 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      __syncwarp();
#pragma unroll 2
      for (int i = 0; i < 8; ++i) {
        const int rr = er + 4 * i;
        const int m = mt * BM + quad * 32 + rr, n = nt * BN + c0 + ec;
        if (m < p.Kin && n < p.N) {
          const float4 acc = *reinterpret_cast<const float4*>(stg + rr * 36 + ec);
          atomicAdd(reinterpret_cast<float4*>(Cp + (long long)m * p.N + n), acc);
        }
      }
      __syncwarp();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, BN);
}

template <int BN, bool X3>
cudaError_t launch_w(const WgradP& p, cudaStream_t s) {
  static bool configured = false;
  if (!configured) {
    CVT_CUDA(cudaFuncSetAttribute(wgrad_tma_kernel<BN, X3>, cudaFuncAttributeMaxDynamicSharedMemorySize, WCfg<BN, X3>::SMEM));
    configured = true;
  }
  CUtensorMap tmX, tmY;
  if (p.conv) {
    ConvBox box;
    if (!conv_tma_box(p.cv.ho, p.cv.wo, KT, &box)) { set_error("wgrad_tma: conv geometry does not tile"); return cudaErrorInvalidValue; }
    if (!make_tmap_conv(&tmX, p.X + p.cv.boff, p.cv.hin, p.cv.win, p.cv.cin, p.cv.bstride, p.tokens / (p.cv.ho * p.cv.wo), p.cv.s, box,
                        CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)) return cudaErrorInvalidValue;
  } else
  if (!make_tmap_3d(&tmX, p.X, p.Kin, p.tokens, p.batch, p.ldx, p.x_bs, KT, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)) return cudaErrorInvalidValue;
  if (!make_tmap_3d(&tmY, p.dY, p.N, p.tokens, p.batch, p.lddy, p.dy_bs, KT, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)) return cudaErrorInvalidValue;
  const int m_tiles = cdiv(p.Kin, BM), n_tiles = cdiv(p.N, BN);
  const int tiles = m_tiles * n_tiles * p.batch;
  const int total_steps = cdiv(p.tokens, KT);
  int splits = kNumSMs / tiles;                                 // one CTA per SM (a single wave: prologue and the atomic epilogue are paid once) ...
  const int max_splits = total_steps / 8 > 0 ? total_steps / 8 : 1;   // ... but at least 8 steps (256 tokens) each
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  const int steps_per_split = cdiv(total_steps, splits);
  splits = cdiv(total_steps, steps_per_split);
  CVT_LAUNCH((wgrad_tma_kernel<BN, X3>), tiles * splits, (WCfg<BN, X3>::THREADS), (WCfg<BN, X3>::SMEM), s, p, tmX, tmY, m_tiles, n_tiles, splits, steps_per_split);
  return cudaSuccess;
}
}  // namespace

bool wgrad_tma_supports(const WgradP& p) {
  if (p.conv) {
    const ConvGeom& g = p.cv; ConvBox box;
    if (p.batch != 1 || g.cin < 32 || (g.cin & 31) || g.s < 1 || g.s > 8 || p.Kin != g.k * g.k * g.cin || p.tokens % (g.ho * g.wo) != 0) return false;
    if (!conv_tma_box(g.ho, g.wo, 32, &box) || box.bw * g.s > 256 || box.bh * g.s > 256) return false;
    if ((g.bstride & 3) || (g.boff & 3) || (p.N & 3) || (p.lddy & 3)) return false;
    return !((reinterpret_cast<uintptr_t>(p.X) & 15) || (reinterpret_cast<uintptr_t>(p.dY) & 15) || (reinterpret_cast<uintptr_t>(p.dW) & 15));
  }
  if (p.Kin < 1 || p.N < 4 || (p.N & 3) || p.tokens < 1 || p.batch < 1) return false;      // Kin need not be a multiple of 4 (ldx must)
  if ((p.ldx & 3) || (p.lddy & 3) || (p.x_bs & 3) || (p.dy_bs & 3) || (p.dw_bs & 3)) return false;
  if ((reinterpret_cast<uintptr_t>(p.X) & 15) || (reinterpret_cast<uintptr_t>(p.dY) & 15) || (reinterpret_cast<uintptr_t>(p.dW) & 15)) return false;
  return true;
}

cudaError_t launch_wgrad_tma(const WgradP& p, cudaStream_t s) {
  if (!wgrad_tma_supports(p)) { set_error("wgrad_tma: unsupported shape"); return cudaErrorNotSupported; }
  if (p.x3) {
    if (p.N > 128) return launch_w<256, true>(p, s);
    if (p.N > 64) return launch_w<128, true>(p, s);
    return launch_w<64, true>(p, s);
  }
  if (p.N > 128) return launch_w<256, false>(p, s);
  if (p.N > 64) return launch_w<128, false>(p, s);
  return launch_w<64, false>(p, s);
}

}  // namespace cvt
