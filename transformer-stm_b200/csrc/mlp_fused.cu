// Fused Mlp block of ConvTransformerBlock for inference (models/CvT(Par).py:278-281 with the Sequential of :253-258):
//     x2 = x1 + fc2(gelu(fc1(LayerNorm(x1))))                       (SURVEY kernel group K6)
// One CTA owns 128 tokens.  The [128, 4D] hidden activation never reaches HBM and never touches shared memory:
//   TMA      x1 tile -> shared memory (K-major, 128B swizzle); fc1 / fc2 weight chunks of 64 hidden units through a 2-stage ring
//   row warps LayerNorm the tile in place (it becomes the A operand of fc1), rounded to TF32
//   tcgen05  H_c [128 x 64] = y . W1_c^T into TMEM (double-buffered), row threads add the bias, apply the erf-GELU and write the
//            result back over H_c in TMEM (tcgen05.st), which is then the A operand (TMEM) of O [128 x D] += gelu(H_c) . W2_c
//   epilogue O + b2 + x1 -> x2 through a per-warp shared-memory transpose (coalesced 128-byte rows)
// Same pipeline shape as attn_fwd_kernel (attention_tc.cu): S -> softmax -> P.V there, H -> GELU -> H.W2 here.
// HBM traffic per token: D floats read (+ D re-read from L2 for the residual) and D floats written, against 2 * (D + 4D) + 4 * D for
// LayerNorm + two GEMM kernels.  D = 64: two CTAs per SM; D = 128: one (8 row warps split the 64 chunk columns).
#include <cuda.h>

#include "model.cuh"
#include "tc_common.cuh"
#include "tma_host.cuh"

namespace cvt {
using namespace tc;

namespace {

struct MlpP {
  const float* x1; float* x2;                          // [R, D]
  const float* gamma; const float* beta; float eps;    // norm1 (second use, :278)
  const float* b1; const float* b2;                    // fc1 bias [H4], fc2 bias [D]
  int R, H4;
  // training only: what the backward pass reads (y through a TMA store of the normalised tile)
  float* mean; float* rstd;                            // LayerNorm row statistics [R]
  float* hact; float* dsave;                           // [R, H4]: dropout(gelu(pre)) rounded to TF32; d hact / d pre = gelu'(pre) * keep / (1 - rate)
  uint8_t* mask1; uint8_t* mask2;                      // keep-masks [R, H4] (read in MASKS mode only), [R, D] (written in PHILOX mode, read in MASKS mode)
  int drop_mode;                                       // CVT_DROPOUT_OFF / PHILOX / MASKS
  float keep_scale; uint32_t ph_k0, ph_k1, ph_thr, ph_site1, ph_site2, ph_step; const float* ph_dyn;
};

template <int D> struct MfCfg {
  static constexpr int RW = 8;                         // row warps: warps w and w + 4 share a TMEM lane quadrant and split the 64 chunk columns
  static constexpr int NT = 64 + 32 * RW;
  static constexpr int KB = D / 32;                    // 32-float boxes along the LayerNorm / fc1 contraction axis
  static constexpr int Y_BYTES = 128 * D * 4;
  static constexpr int W1_STAGE = 64 * D * 4;          // KB boxes of 64 hidden rows x 32 floats
  static constexpr int W2_STAGE = D * 64 * 4;          // 2 boxes of D output rows x 32 floats
  static constexpr int STAGE = W1_STAGE + W2_STAGE;
  static constexpr int PAR_FLOATS = 3 * D + 1024;      // gamma, beta, b2, b1 (H4 <= 1024)
  static constexpr int SMEM = 1024 + Y_BYTES + 2 * STAGE + PAR_FLOATS * 4 + 16 * 8 + 16;
  static constexpr int TMEM_COLS = 256;                // H0 [0,64), H1 [64,128), O [128, 128 + D)
  static constexpr int COLS = 64 / (RW / 4);           // chunk columns per row thread
};
enum { MB_XF = 0, MB_YF = 1, MB_W1F = 2, MB_W1E = 4, MB_W2F = 6, MB_W2E = 8, MB_HF = 10, MB_GF = 12, MB_DONE = 14, MB_COUNT = 15 };

__device__ __forceinline__ bool elect_one_mf() {
  uint32_t pred;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(pred));
  return pred != 0;
}

// erf-GELU from one exponential (Abramowitz & Stegun 7.1.26, |error| <= 1.5e-7, as gelu_and_grad in common.cuh), without a select:
// x Phi(x) = max(x, 0) - |x| * (erfc(|x| / sqrt 2) / 2); the 1/2 is folded into the polynomial coefficients
__device__ __forceinline__ float gelu_fast(float x) {
  const float ax = fabsf(x);
  const float z = ax * 0.70710678118654752440f;
  const float t = __fdividef(1.0f, fmaf(0.3275911f, z, 1.0f));
  const float e = fast_ex2(-1.4426950408889634f * z * z);
  float q = fmaf(t, 0.5f * 1.061405429f, 0.5f * -1.453152027f);
  q = fmaf(q, t, 0.5f * 1.421413741f); q = fmaf(q, t, 0.5f * -0.284496736f); q = fmaf(q, t, 0.5f * 0.254829592f);
  return fmaf(-ax, q * t * e, fmaxf(x, 0.f));
}

__device__ __forceinline__ void st_v8(float* ptr, const float* a) {       // one 32-byte store per thread (st.global.v8.f32, sm_100)
  asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(ptr), "f"(a[0]), "f"(a[1]), "f"(a[2]), "f"(a[3]), "f"(a[4]),
               "f"(a[5]), "f"(a[6]), "f"(a[7])
               : "memory");
}
__device__ __forceinline__ void tma_store_3d(const void* tmap, const void* smem_src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
// keep bits (one byte per element, 0 / 1) of the 4 elements of quad `q` (= flat element index / 4) of a dropout site.  MODE is the kernel's
// compile-time variant: 2 = Philox4x32-10 (counter {q, site, step, 0}, key = seed: the stream of philox_mask_kernel, the GEMM epilogues and
// oracle.dropout_mask), 3 = caller-supplied masks, else no dropout
template <int MODE>
__device__ __forceinline__ uint32_t keep4(const MlpP& p, uint32_t site, const uint8_t* mask, uint32_t q, uint32_t step) {
  if (MODE == 2) {
    uint32_t c[4] = {q, site, step, 0u};
    philox4x32_10(c, p.ph_k0, p.ph_k1);
    return (c[0] >= p.ph_thr ? 1u : 0u) | (c[1] >= p.ph_thr ? 0x100u : 0u) | (c[2] >= p.ph_thr ? 0x10000u : 0u) | (c[3] >= p.ph_thr ? 0x1000000u : 0u);
  }
  if (MODE == 3) return __ldg(reinterpret_cast<const uint32_t*>(mask) + q);
  return 0x01010101u;
}

// TRAIN: additionally leaves what the backward pass reads - y (TMA store of the normalised tile), the LayerNorm row statistics, the
// dropped-out hidden activation and d hact / d pre (one exponential serves both, gelu_and_grad), the fc2 dropout mask - and applies
// the two dropouts (Philox counters identical to the GEMM epilogues' and to oracle.dropout_mask, or caller-supplied masks)
// MODE: 0 inference, 1 training without dropout, 2 training with Philox dropout, 3 training with caller-supplied masks
template <int D, int MODE>
__global__ void __launch_bounds__(MfCfg<D>::NT, D == 64 ? 2 : 1) mlp_fused_kernel(const __grid_constant__ MlpP p, const __grid_constant__ CUtensorMap tmX,
                                                                                  const __grid_constant__ CUtensorMap tmW1,
                                                                                  const __grid_constant__ CUtensorMap tmW2,
                                                                                  const __grid_constant__ CUtensorMap tmY) {
  using C = MfCfg<D>;
  constexpr int RW = C::RW, NT = C::NT, KB = C::KB, COLS = C::COLS;
  constexpr bool TRAIN = MODE > 0;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sY = sm;                                     // x1 tile, normalised in place
  uint8_t* ring = sY + C::Y_BYTES;                      // 2 x (W1 chunk, W2 chunk); the epilogue's transpose tiles alias it
  float* sG = reinterpret_cast<float*>(ring + 2 * C::STAGE); float* sBe = sG + D; float* sB2 = sBe + D; float* sB1 = sB2 + D;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB1 + 1024);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + MB_COUNT);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row0 = blockIdx.x * 128;
  const int nchunks = p.H4 >> 6;

  if (tid == 0) {
    tma_prefetch_desc(&tmX); tma_prefetch_desc(&tmW1); tma_prefetch_desc(&tmW2);
    for (int i = 0; i < MB_COUNT; ++i) mbar_init(&bars[i], (i == MB_YF || i == MB_GF || i == MB_GF + 1) ? RW * 32 : 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, C::TMEM_COLS);
  pdl_wait(); pdl_trigger();      // everything above touches no global memory
  if (tid == 0) {
    // the token tile is requested before the CTA has finished setting up (parameter staging overlaps its flight)
    mbar_arrive_expect_tx(&bars[MB_XF], C::Y_BYTES);
#pragma unroll
    for (int kb = 0; kb < KB; ++kb) tma_load_3d(sY + kb * 16384, &tmX, &bars[MB_XF], kb * 32, row0, 0);
  }
  for (int i = tid; i < D; i += NT) { sG[i] = __ldg(p.gamma + i); sBe[i] = __ldg(p.beta + i); sB2[i] = __ldg(p.b2 + i); }
  for (int i = tid; i < p.H4; i += NT) sB1[i] = __ldg(p.b1 + i);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      for (int c = 0; c < nchunks; ++c) {
        const int st = c & 1; const uint32_t ph = (c >> 1) & 1;
        uint8_t* sg = ring + st * C::STAGE;
        mbar_wait(&bars[MB_W1E + st], ph ^ 1);
        mbar_arrive_expect_tx(&bars[MB_W1F + st], C::W1_STAGE);
#pragma unroll
        for (int kb = 0; kb < KB; ++kb) tma_load_3d(sg + kb * 8192, &tmW1, &bars[MB_W1F + st], kb * 32, c * 64, 0);
        mbar_wait(&bars[MB_W2E + st], ph ^ 1);
        mbar_arrive_expect_tx(&bars[MB_W2F + st], C::W2_STAGE);
        tma_load_3d(sg + C::W1_STAGE, &tmW2, &bars[MB_W2F + st], c * 64, 0, 0);
        tma_load_3d(sg + C::W1_STAGE + D * 128, &tmW2, &bars[MB_W2F + st], c * 64 + 32, 0, 0);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: H_0, H_1, O_0, H_2, O_1, ... (tcgen05.mma executes in issue order) =====
    constexpr uint32_t idesc_h = make_idesc(kFmtTF32, 128, 64, 0, 0);
    constexpr uint32_t idesc_o = make_idesc(kFmtTF32, 128, D, 0, 0);
    const uint32_t y0 = smem_u32(sY);
    mbar_wait(&bars[MB_YF], 0);
    if (TRAIN && elect_one_mf()) {                      // y = norm1(x1), the operand of the fc1 weight gradient: the tile as it lies, rows past R clipped
#pragma unroll
      for (int kb = 0; kb < KB; ++kb) tma_store_3d(&tmY, sY + kb * 16384, kb * 32, row0, 0);
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    __syncwarp();
    auto issue_h = [&](int c) {
      const int st = c & 1;
      mbar_wait(&bars[MB_W1F + st], (c >> 1) & 1);
      tc_fence_after();
      if (elect_one_mf()) {
        const uint32_t w0 = smem_u32(ring + st * C::STAGE);
#pragma unroll
        for (int kb = 0; kb < KB; ++kb)
#pragma unroll
          for (int j = 0; j < 4; ++j)
            mma_tf32_ss(tmem + st * 64, make_smem_desc(y0 + kb * 16384 + j * 32, 16, 1024), make_smem_desc(w0 + kb * 8192 + j * 32, 16, 1024), idesc_h,
                        (kb > 0 || j > 0) ? 1u : 0u);
        mma_commit(&bars[MB_W1E + st]);
        mma_commit(&bars[MB_HF + st]);
      }
      __syncwarp();
    };
    issue_h(0);
    for (int c = 0; c < nchunks; ++c) {
      if (c + 1 < nchunks) issue_h(c + 1);
      const int st = c & 1; const uint32_t ph = (c >> 1) & 1;
      mbar_wait(&bars[MB_GF + st], ph);
      mbar_wait(&bars[MB_W2F + st], ph);
      tc_fence_after();
      if (elect_one_mf()) {
        const uint32_t w2 = smem_u32(ring + st * C::STAGE + C::W1_STAGE);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          mma_tf32_ts(tmem + 128, tmem + st * 64 + j * 8, make_smem_desc(w2 + (j >> 2) * (D * 128) + (j & 3) * 32, 16, 1024), idesc_o,
                      (c > 0 || j > 0) ? 1u : 0u);
        mma_commit(&bars[MB_W2E + st]);
        if (c == nchunks - 1) mma_commit(&bars[MB_DONE]);
      }
      __syncwarp();
    }
    if (TRAIN) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");      // (the thread that committed the group is the one that waits)
  } else {
    // ===== row warps =====
    {  // LayerNorm in place: TPR threads per row (adjacent lanes), 16-byte chunks of the swizzled K-major tile
      constexpr int TPR = RW / 4, NCH = D / 4;
      const int t = tid - 64, row = t / TPR, sub = t % TPR;
      mbar_wait(&bars[MB_XF], 0);
      auto chunk = [&](int c4) -> float4* { return reinterpret_cast<float4*>(sY + (c4 >> 3) * 16384 + sw128_kmajor_off(row, c4 & 7)); };
      float s = 0.f;
      for (int c4 = sub; c4 < NCH; c4 += TPR) { const float4 v = *chunk(c4); s += (v.x + v.y) + (v.z + v.w); }
      if (TPR == 2) s += __shfl_xor_sync(0xffffffffu, s, 1);
      const float mu = s * (1.f / (float)D);
      float q = 0.f;
      for (int c4 = sub; c4 < NCH; c4 += TPR) {
        const float4 v = *chunk(c4);
        const float a = v.x - mu, b = v.y - mu, c = v.z - mu, d = v.w - mu;
        q += (a * a + b * b) + (c * c + d * d);
      }
      if (TPR == 2) q += __shfl_xor_sync(0xffffffffu, q, 1);
      const float r = 1.0f / sqrtf(q * (1.f / (float)D) + p.eps);
      if (TRAIN && sub == 0 && row0 + row < p.R) { p.mean[row0 + row] = mu; p.rstd[row0 + row] = r; }
      for (int c4 = sub; c4 < NCH; c4 += TPR) {
        float4* ptr = chunk(c4);
        const float4 v = *ptr;
        const float4 g = *reinterpret_cast<const float4*>(sG + 4 * c4), be = *reinterpret_cast<const float4*>(sBe + 4 * c4);
        float4 o;
        o.x = to_tf32((v.x - mu) * r * g.x + be.x); o.y = to_tf32((v.y - mu) * r * g.y + be.y);
        o.z = to_tf32((v.z - mu) * r * g.z + be.z); o.w = to_tf32((v.w - mu) * r * g.w + be.w);
        *ptr = o;
      }
      fence_proxy_async_smem();
      mbar_arrive(&bars[MB_YF]);
    }
    const int quad = warp & 3, half = (warp - 2) >> 2;
    const uint32_t t_row = tmem + ((uint32_t)(quad * 32) << 16);
    const uint32_t step = (MODE == 2 && p.ph_dyn) ? __ldg(reinterpret_cast<const uint32_t*>(p.ph_dyn) + 2) : p.ph_step;
    for (int c = 0; c < nchunks; ++c) {
      const int st = c & 1;
      mbar_wait(&bars[MB_HF + st], (c >> 1) & 1);
      tc_fence_after();
      uint32_t h[COLS];
#pragma unroll
      for (int i = 0; i < COLS / 32; ++i) tmem_ld32_nowait(t_row + st * 64 + half * COLS + i * 32, h + i * 32);
      tmem_wait_ld();
      const float* bias = sB1 + c * 64 + half * COLS;
      if (TRAIN) {
        const int grow = row0 + quad * 32 + lane;       // this thread's token
        const bool valid = grow < p.R;
        const long long base = (long long)grow * p.H4 + c * 64 + half * COLS;
        const uint32_t q0 = valid ? (uint32_t)grow * (uint32_t)(p.H4 >> 2) + (uint32_t)((c * 64 + half * COLS) >> 2) : 0u;
#pragma unroll
        for (int i = 0; i < COLS; i += 8) {
          float hv[8], dv[8];
#pragma unroll
          for (int u = 0; u < 8; u += 4) {
            const float4 b4 = *reinterpret_cast<const float4*>(bias + i + u);
            const float bb[4] = {b4.x, b4.y, b4.z, b4.w};
            const uint32_t mk = keep4<MODE>(p, p.ph_site1, p.mask1, q0 + ((i + u) >> 2), step);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float g, dg;
              gelu_and_grad(__uint_as_float(h[i + u + j]) + bb[j], g, dg);
              const bool keep = ((mk >> (8 * j)) & 0xFFu) != 0u;
              hv[u + j] = to_tf32(keep ? g * p.keep_scale : 0.f);
              dv[u + j] = keep ? dg * p.keep_scale : 0.f;
              h[i + u + j] = __float_as_uint(hv[u + j]);
            }
          }
          if (valid) { st_v8(p.hact + base + i, hv); st_v8(p.dsave + base + i, dv); }
        }
      } else {
#pragma unroll
      for (int i = 0; i < COLS; i += 4) {
        const float4 b4 = *reinterpret_cast<const float4*>(bias + i);
        h[i] = tf32_rna_bits(gelu_fast(__uint_as_float(h[i]) + b4.x)); h[i + 1] = tf32_rna_bits(gelu_fast(__uint_as_float(h[i + 1]) + b4.y));
        h[i + 2] = tf32_rna_bits(gelu_fast(__uint_as_float(h[i + 2]) + b4.z)); h[i + 3] = tf32_rna_bits(gelu_fast(__uint_as_float(h[i + 3]) + b4.w));
      }
      }
#pragma unroll
      for (int i = 0; i < COLS / 32; ++i) tmem_st32(t_row + st * 64 + half * COLS + i * 32, h + i * 32);
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(&bars[MB_GF + st]);
    }
    // ===== epilogue: x2 = O + b2 + x1, 32 x 32 blocks through a padded transpose tile (the weight ring is idle by now).  The residual
    // rows are fetched (from L2: the tile was just read) before the wait for the last fc2 product, lane = column =====
    constexpr int NBLK = D / 32 / (RW / 4);            // 32-column blocks of O per warp
    float res[NBLK][32];
#pragma unroll
    for (int blk = 0; blk < NBLK; ++blk) {
      const int col = (half * NBLK + blk) * 32 + lane;
#pragma unroll
      for (int rr = 0; rr < 32; ++rr) {
        const int row = min(row0 + quad * 32 + rr, p.R - 1);
        res[blk][rr] = __ldg(p.x1 + (long long)row * D + col) + (TRAIN ? 0.f : sB2[col]);
      }
    }
    mbar_wait(&bars[MB_DONE], 0);
    tc_fence_after();
    float* stg = reinterpret_cast<float*>(ring) + (warp - 2) * (32 * 33);
#pragma unroll
    for (int blk = 0; blk < NBLK; ++blk) {
      const int cb = half * NBLK + blk;
      float v[32];
      tmem_ld32(t_row + 128 + cb * 32, v);
      if (TRAIN) {                                      // fc2 bias and dropout (:257) in the row layout (4 consecutive columns per Philox draw)
        const int grow = row0 + quad * 32 + lane;
        const bool valid = grow < p.R;
        const long long base = (long long)grow * D + cb * 32;
        const uint32_t q0 = valid ? (uint32_t)grow * (uint32_t)(D >> 2) + (uint32_t)(cb * 8) : 0u;
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const uint32_t mk = keep4<MODE>(p, p.ph_site2, p.mask2, q0 + (i >> 2), step);
          if (valid && MODE == 2) *reinterpret_cast<uint32_t*>(p.mask2 + base + i) = mk;
#pragma unroll
          for (int j = 0; j < 4; ++j) v[i + j] = ((mk >> (8 * j)) & 0xFFu) ? (v[i + j] + sB2[cb * 32 + i + j]) * p.keep_scale : 0.f;
        }
      }
#pragma unroll
      for (int i = 0; i < 32; ++i) stg[lane * 33 + i] = v[i];
      __syncwarp();
      const int col = cb * 32 + lane;
#pragma unroll
      for (int rr = 0; rr < 32; ++rr) {
        const int row = row0 + quad * 32 + rr;
        if (row < p.R) p.x2[(long long)row * D + col] = stg[rr * 33 + lane] + res[blk][rr];
      }
      __syncwarp();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, C::TMEM_COLS);
}

template <int D, int MODE>
cudaError_t launch_cfg(const MlpP& p, const float* w1t, const float* w2t, float* y_out, cudaStream_t s) {
  using C = MfCfg<D>;
  constexpr bool TRAIN = MODE > 0;
  static bool configured = false;
  if (!configured) {
    CVT_CUDA(cudaFuncSetAttribute(mlp_fused_kernel<D, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
    configured = true;
  }
  CUtensorMap tmX, tmW1, tmW2, tmY;
  if (!make_tmap_3d(&tmX, p.x1, D, p.R, 1, D, 0, 128, CU_TENSOR_MAP_SWIZZLE_128B)) return cudaErrorInvalidValue;
  if (!make_tmap_3d(&tmW1, w1t, D, p.H4, 1, D, 0, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return cudaErrorInvalidValue;      // fc1 kernel, K-major copy [H4][D]
  if (!make_tmap_3d(&tmW2, w2t, p.H4, D, 1, p.H4, 0, D, CU_TENSOR_MAP_SWIZZLE_128B)) return cudaErrorInvalidValue;    // fc2 kernel, K-major copy [D][H4]
  tmY = tmX;
  if (TRAIN && !make_tmap_3d(&tmY, y_out, D, p.R, 1, D, 0, 128, CU_TENSOR_MAP_SWIZZLE_128B)) return cudaErrorInvalidValue;
  CVT_LAUNCH((mlp_fused_kernel<D, MODE>), cdiv(p.R, 128), C::NT, C::SMEM, s, p, tmX, tmW1, tmW2, tmY);
  return cudaSuccess;
}

}  // namespace

bool mlp_fused_supports(const Model* m, int si) {
  static const bool off = getenv("CVT_NO_FUSED_MLP") != nullptr || getenv("CVT_NO_TMA") != nullptr;
  const int D = m->geo[si].D, H4 = m->cfg.mlp_ratio * D;
  return !off && m->cfg.precision != CVT_PREC_FP32 && !m->x3 && m->pt_valid && (D == 64 || D == 128) && H4 % 64 == 0 && H4 <= 1024;
}

// x2 = x1 + Mlp(norm1(x1)) of stage si for R tokens.  training == 0: nothing is kept.  training != 0 (folded TF32 mode): y, the LayerNorm
// statistics, hact, d hact / d pre (in the hpre buffer) and mask2 are left for stage_backward exactly as the three-kernel path leaves them.
cudaError_t launch_mlp_fused(const Model* m, int si, int R, int training, cudaStream_t s) {
  const StageGeo& g = m->geo[si]; const StageW& w = m->w[si]; const StageBuf& a = m->buf[si];
  MlpP p{};
  p.x1 = a.x1; p.x2 = a.x2; p.gamma = m->P + w.ln_g; p.beta = m->P + w.ln_b; p.eps = m->cfg.ln_eps;
  p.b1 = m->P + w.f1b; p.b2 = m->P + w.f2b; p.R = R; p.H4 = m->cfg.mlp_ratio * g.D;
  if (training) {
    const bool drop = m->dropout_mode != CVT_DROPOUT_OFF && m->cfg.drop_rate > 0.f;
    p.mean = a.lnb_mean; p.rstd = a.lnb_rstd; p.hact = a.hact; p.dsave = a.hpre; p.mask1 = a.mask1; p.mask2 = a.mask2;
    p.drop_mode = drop ? m->dropout_mode : CVT_DROPOUT_OFF;
    p.keep_scale = drop ? 1.f / (1.f - m->cfg.drop_rate) : 1.f;
    const double t = (double)m->cfg.drop_rate * 4294967296.0;
    p.ph_thr = t >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)t;
    p.ph_k0 = (uint32_t)(m->dropout_seed & 0xFFFFFFFFu); p.ph_k1 = (uint32_t)(m->dropout_seed >> 32);
    p.ph_site1 = (uint32_t)(si * 4 + 1); p.ph_site2 = (uint32_t)(si * 4 + 2); p.ph_step = m->dropout_step; p.ph_dyn = m->dyn;
    const int mode = p.drop_mode == CVT_DROPOUT_PHILOX ? 2 : (p.drop_mode == CVT_DROPOUT_MASKS ? 3 : 1);
#define MF_LAUNCH(DD) (mode == 2 ? launch_cfg<DD, 2>(p, m->Pt + w.f1w, m->Pt + w.f2w, a.y, s) : mode == 3 ? launch_cfg<DD, 3>(p, m->Pt + w.f1w, m->Pt + w.f2w, a.y, s) \
                                 : launch_cfg<DD, 1>(p, m->Pt + w.f1w, m->Pt + w.f2w, a.y, s))
    if (g.D == 64) return MF_LAUNCH(64);
    if (g.D == 128) return MF_LAUNCH(128);
#undef MF_LAUNCH
  } else {
    if (g.D == 64) return launch_cfg<64, 0>(p, m->Pt + w.f1w, m->Pt + w.f2w, nullptr, s);
    if (g.D == 128) return launch_cfg<128, 0>(p, m->Pt + w.f1w, m->Pt + w.f2w, nullptr, s);
  }
  set_error("mlp_fused: unsupported embed_dim");
  return cudaErrorNotSupported;
}

}  // namespace cvt
