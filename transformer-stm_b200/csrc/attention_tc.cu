// Flash-style attention on tcgen05 / TMEM (kind::tf32, fp32 accumulate) for keras.layers.MultiHeadAttention as called at
// models/CvT(Par).py:185:  S = Q K^T (Q already carries 1/sqrt(key_dim)), P = softmax(S), O = P V; key_dim = 64 in every spec.
// The [B, heads, Tc, Tc] score tensor never reaches HBM: forward keeps the row log-sum-exp, backward recomputes P from it.
//
// All three kernels are warp-specialised pipelines: warp 0 = TMA producer (K-major tiles with the 128B swizzle, MN-major tiles
// with the 32-byte-atom 128B swizzle, both straight from the [B, Tc, D] tensors, zero fill past Tc), warp 1 = tcgen05.mma issuer,
// the remaining warps own one TMEM lane (= one row of the 128-row tile) each.  The probabilities never touch shared memory:
// they are written back over the scores in TMEM (tcgen05.st) and consumed from there as the A operand of the next product.
// The score tile is double-buffered so that the tensor core works on tile j+1 while the row threads exponentiate tile j.
//
//   forward   CTA = 128 query rows of one (image, head), 64-key tiles, 2 CTAs per SM.  O accumulates in TMEM; the running max
//             is only refreshed when it grows by more than 2^8 (the exponentials stay in fp32 range, the final O / l is exact),
//             so O is rescaled in TMEM a handful of times per row block instead of once per tile.
//   backward  dQ kernel (CTA = 128 query rows, loops over key tiles) and dK/dV kernel (CTA = 128 key rows, loops over query
//             tiles, works on S^T so that P^T / dS^T are produced row-per-thread).  8 row warps: warps w and w+4 share a TMEM
//             lane quadrant and split the tile columns.  No atomics.  Two families: 64-wide score tiles, one CTA per 128-row
//             block (Tc <= 64, and the trace hooks), and persistent kernels with 128-wide score tiles (everything longer).
//   operands  Q/K/V/dO must already be rounded to TF32 by their producers (EPI_ROUND); TMA moves raw fp32 words.
#include <cuda.h>

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <string>

#include "model.cuh"
#include "tc_common.cuh"
#include "tma_host.cuh"

namespace cvt {
using namespace tc;

namespace {
constexpr int HD = 64;
constexpr float kLog2e = 1.4426950408889634f;
constexpr int TILE16K = 16384;                          // one 64-row x 64-float operand tile (two 32-float boxes of 8 KB)
constexpr int TILE32K = 32768;                          // one 128-row x 64-float operand tile (two boxes of 16 KB)
constexpr uint32_t kFullMask = 0xffffffffu;

struct AttnP {
  const float* Q; const float* K; const float* V;      // [B, Tc, D]
  float* O; float* LSE;                                // O [B, Tc, D]; LSE [B*heads, Tc]
  const float* dO; const float* Dd;                    // backward: dO [B, Tc, D]; Dd [B*heads, Tc] = rowsum(dO * O)
  float* dQ; float* dK; float* dV;
  int Tc, D, heads;
  float dq_scale;
  int round_out;                                       // store O / dQ / dK / dV rounded to TF32 (they only feed tensor-core GEMMs)
  long long* trace;                                    // diagnostics: clock64 timeline of CTA (0,0) of the dQ kernel (CVT_ATTN_TRACE)
};
#define ATTN_TRACE(slot) do { if (p.trace && blockIdx.x == 0 && blockIdx.y == 0 && lane == 0) p.trace[(slot)] = clock64(); } while (0)

// Descriptors are built once per operand tile; each MMA only adds a compile-time constant to the 14-bit start-address field
// (the single issuing thread's instruction stream is what bounds the MMA rate of these small tiles).
__device__ __forceinline__ uint64_t desc_add(uint64_t d, uint32_t bytes) { return d + (uint64_t)(bytes >> 4); }
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t saddr) { return make_smem_desc(saddr, 16, 1024); }
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t saddr) { return make_smem_desc(saddr, 8192, 512, kLayoutSw128Base32); }
__device__ __forceinline__ bool elect_one() {
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
// D[tmem] = A * B with A K-major [128 x 64] (two boxes of A_ROWS rows) and B K-major [B_ROWS x 64]; contraction = 64 floats
template <int A_ROWS, int B_ROWS>
__device__ __forceinline__ void mma_kk(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc) {
#pragma unroll
  for (int j = 0; j < 8; ++j)
    mma_tf32_ss(tmem_d, desc_add(adesc, (j >> 2) * (A_ROWS * 128) + (j & 3) * 32), desc_add(bdesc, (j >> 2) * (B_ROWS * 128) + (j & 3) * 32), idesc,
                j > 0 ? 1u : 0u);
}
// D[tmem] (+)= A[tmem, 128 lanes x 64 columns] * B with B MN-major: 64 contraction rows x 64 features as two TMA boxes
// (32 features x 64 rows, 8 KB each, 32-byte-atom swizzle): LBO = next feature box, SBO = next group of 4 contraction rows
__device__ __forceinline__ void mma_tmn(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, bool accumulate) {
#pragma unroll
  for (int j = 0; j < 8; ++j) mma_tf32_ts(tmem_d, tmem_a + j * 8, desc_add(bdesc, j * 1024), idesc, (accumulate || j > 0) ? 1u : 0u);
}

__device__ __forceinline__ uint8_t* align1024(uint8_t* p) {
  const uint32_t a = smem_u32(p);
  return p + (((a + 1023u) & ~1023u) - a);
}
// two 32-float boxes of `rows` rows: columns [col0, col0 + 64) of rows [row0, row0 + rows) of image z
__device__ __forceinline__ void tma_tile(uint8_t* dst, const CUtensorMap* tm, uint64_t* bar, int col0, int row0, int z, int rows) {
  tma_load_3d(dst, tm, bar, col0, row0, z);
  tma_load_3d(dst + rows * 128, tm, bar, col0 + 32, row0, z);
}
__device__ __forceinline__ float4 out4(const uint32_t* v, float sc, int round) {
  float4 t = make_float4(__uint_as_float(v[0]) * sc, __uint_as_float(v[1]) * sc, __uint_as_float(v[2]) * sc, __uint_as_float(v[3]) * sc);
  if (round) t = make_float4(to_tf32(t.x), to_tf32(t.y), to_tf32(t.z), to_tf32(t.w));
  return t;
}

// =====================================================================================================================
// forward
// =====================================================================================================================
constexpr int FWD_THREADS = 192;
constexpr int FWD_SMEM = 1024 + TILE32K + 4 * TILE16K + 128;
enum { FB_Q = 0, FB_KF = 1, FB_KE = 3, FB_VF = 5, FB_VE = 7, FB_SF = 9, FB_PF = 11, FB_PV = 13, FB_DONE = 14, FB_COUNT = 15 };

__global__ void __launch_bounds__(FWD_THREADS, 2) attn_fwd_kernel(const __grid_constant__ AttnP p, const __grid_constant__ CUtensorMap tmQ,
                                                                   const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = align1024(smem_raw);
  uint8_t* sQ = sm; uint8_t* sK = sQ + TILE32K; uint8_t* sV = sK + 2 * TILE16K;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + 2 * TILE16K);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + FB_COUNT);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int bh = blockIdx.y, b = bh / p.heads, h = bh - b * p.heads;
  const int q0 = blockIdx.x * 128;
  const int ntiles = cdiv(p.Tc, 64);

  if (tid == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
    for (int i = 0; i < FB_COUNT; ++i) mbar_init(&bars[i], (i == FB_PF || i == FB_PF + 1) ? 128 : 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait(); pdl_trigger();      // everything above touches no global memory
  const uint32_t tmem = *tmem_slot;                      // columns: S0 [0,64), S1 [64,128), O [128,192)

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      mbar_arrive_expect_tx(&bars[FB_Q], TILE32K);
      tma_tile(sQ, &tmQ, &bars[FB_Q], h * HD, q0, b, 128);
      for (int j = 0; j < ntiles; ++j) {
        const int st = j & 1; const uint32_t ph = (j >> 1) & 1;
        mbar_wait(&bars[FB_KE + st], ph ^ 1);
        mbar_arrive_expect_tx(&bars[FB_KF + st], TILE16K);
        tma_tile(sK + st * TILE16K, &tmK, &bars[FB_KF + st], h * HD, j * 64, b, 64);
        mbar_wait(&bars[FB_VE + st], ph ^ 1);
        mbar_arrive_expect_tx(&bars[FB_VF + st], TILE16K);
        tma_tile(sV + st * TILE16K, &tmV, &bars[FB_VF + st], h * HD, j * 64, b, 64);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: S_0, S_1, PV_0, S_2, PV_1, ... (tcgen05.mma executes in issue order).  The whole warp walks the loop
    // (uniform control flow keeps the descriptor arithmetic on the uniform datapath); one elected lane issues.
    constexpr uint32_t idesc_s = make_idesc(kFmtTF32, 128, 64, 0, 0);
    constexpr uint32_t idesc_o = make_idesc(kFmtTF32, 128, 64, 0, 1);
    const uint64_t dQ = desc_kmajor(smem_u32(sQ));
    const uint64_t dK[2] = {desc_kmajor(smem_u32(sK)), desc_kmajor(smem_u32(sK + TILE16K))};
    const uint64_t dV[2] = {desc_mnmajor(smem_u32(sV)), desc_mnmajor(smem_u32(sV + TILE16K))};
    mbar_wait(&bars[FB_Q], 0);
    auto issue_s = [&](int j) {
      const int st = j & 1;
      mbar_wait(&bars[FB_KF + st], (j >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        mma_kk<128, 64>(tmem + st * 64, dQ, dK[st], idesc_s);
        mma_commit(&bars[FB_KE + st]);
        mma_commit(&bars[FB_SF + st]);
      }
      __syncwarp();
    };
    issue_s(0);
    for (int j = 0; j < ntiles; ++j) {
      if (j + 1 < ntiles) issue_s(j + 1);
      const int st = j & 1; const uint32_t ph = (j >> 1) & 1;
      mbar_wait(&bars[FB_PF + st], ph);
      mbar_wait(&bars[FB_VF + st], ph);
      tc_fence_after();
      if (elect_one()) {
        mma_tmn(tmem + 128, tmem + st * 64, dV[st], idesc_o, j > 0);
        mma_commit(&bars[FB_VE + st]);
        mma_commit(&bars[FB_PV]);
        if (j == ntiles - 1) mma_commit(&bars[FB_DONE]);
      }
      __syncwarp();
    }
  } else {
    // ===== row threads: thread owns query row q0 + 32 * quad + lane = TMEM lane 32 * quad + lane =====
    const int quad = warp & 3;
    const uint32_t t_row = tmem + ((uint32_t)(quad * 32) << 16);
    float m_used = -INFINITY, l_run = 0.f;
    for (int j = 0; j < ntiles; ++j) {
      const int st = j & 1;
      mbar_wait(&bars[FB_SF + st], (j >> 1) & 1);
      tc_fence_after();
      uint32_t s[64];
      tmem_ld32_nowait(t_row + st * 64, s);
      tmem_ld32_nowait(t_row + st * 64 + 32, s + 32);
      tmem_wait_ld();
      const int nvalid = p.Tc - j * 64;                  // keys beyond Tc are masked out (last tile only)
      if (nvalid < 64) {
#pragma unroll
        for (int c = 0; c < 64; ++c) if (c >= nvalid) s[c] = __float_as_uint(-INFINITY);
      }
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < 64; ++c) mx = fmaxf(mx, __uint_as_float(s[c]));
      const float m_new = fmaxf(m_used, mx);
      const bool grow = (m_new - m_used) * kLog2e > 8.f;
      if (__any_sync(kFullMask, grow)) {
        const float alpha = exp2f((m_used - m_new) * kLog2e);
        if (j > 0) {                                     // rescale this warp's 32 rows of O once PV_{j-1} has landed
          mbar_wait(&bars[FB_PV], (j - 1) & 1);
          tc_fence_after();
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            uint32_t o[32];
            tmem_ld32_nowait(t_row + 128 + half * 32, o);
            tmem_wait_ld();
#pragma unroll
            for (int c = 0; c < 32; ++c) o[c] = __float_as_uint(__uint_as_float(o[c]) * alpha);
            tmem_st32(t_row + 128 + half * 32, o);
          }
        }
        l_run *= alpha;
        m_used = m_new;
      }
      const float mb = m_used * kLog2e;
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < 64; ++c) {
        const float e = fast_ex2(fmaf(__uint_as_float(s[c]), kLog2e, -mb));
        sum += e;
        s[c] = tf32_rna_bits(e);
      }
      l_run += sum;
      tmem_st32(t_row + st * 64, s);
      tmem_st32(t_row + st * 64 + 32, s + 32);
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(&bars[FB_PF + st]);
    }
    // A barrier of its own for "the last PV product has landed": a row warp that never rescaled has not followed FB_PV, whose one-bit
    // phase parity cannot tell "PV_{n-1} done" from "PV_{n-3} done" - with two CTAs sharing the tensor pipe a fast warp got here while
    // PV_{n-2} was still in flight and read O two tiles short (found by the batch-64 parity test; 1 CTA per SM never showed it).
    mbar_wait(&bars[FB_DONE], 0);
    tc_fence_after();
    const int q = q0 + quad * 32 + lane;
    const float inv = 1.f / l_run;
    float* dst = p.O + ((long long)b * p.Tc + q) * p.D + h * HD;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      uint32_t o[32];
      tmem_ld32_nowait(t_row + 128 + half * 32, o);
      tmem_wait_ld();
      if (q < p.Tc) {
#pragma unroll
        for (int c = 0; c < 8; ++c) *reinterpret_cast<float4*>(dst + half * 32 + 4 * c) = out4(o + 4 * c, inv, p.round_out);
      }
    }
    if (q < p.Tc && p.LSE) p.LSE[(long long)bh * p.Tc + q] = m_used + logf(l_run);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 256);
}

// =====================================================================================================================
// backward
// =====================================================================================================================
// Dd[bh, t] = sum_d dO[b,t,h,d] * O[b,t,h,d]   (half a warp per (token, head))
__global__ void __launch_bounds__(256) attn_dd_kernel(const float* __restrict__ dO, const float* __restrict__ O, float* __restrict__ Dd,
                                                      int B, int Tc, int D, int heads) {
  pdl_wait(); pdl_trigger();
  const long long gid = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 4;     // (b, t, h)
  const int l16 = threadIdx.x & 15;
  const long long total = (long long)B * Tc * heads;
  float s = 0.f;
  long long b = 0; int t = 0, h = 0;
  if (gid < total) {
    h = (int)(gid % heads); const long long bt = gid / heads; t = (int)(bt % Tc); b = bt / Tc;
    const long long off = (b * Tc + t) * D + h * HD + l16 * 4;
    const float4 a = *reinterpret_cast<const float4*>(dO + off);
    const float4 c = *reinterpret_cast<const float4*>(O + off);
    s = a.x * c.x + a.y * c.y + a.z * c.z + a.w * c.w;
  }
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (gid < total && l16 == 0) Dd[(b * heads + h) * Tc + t] = s;
}

constexpr int BWD_THREADS = 320;                       // producer, issuer, 8 row warps

// ---- dQ: CTA = 128 query rows, loop over 64-key tiles -----------------------------------------------------------------
constexpr int BQ_STAGE = 3 * TILE16K;                  // K (K-major), V (K-major), K (MN-major)
constexpr int BQ_NS = 3;                              // shared-memory stages (the TMEM score buffers stay double)
constexpr int BQ_SMEM = 1024 + 2 * TILE32K + BQ_NS * BQ_STAGE + 128;
enum { QB_QDO = 0, QB_KVF = 1, QB_KVE = QB_KVF + BQ_NS, QB_KMF = QB_KVE + BQ_NS, QB_KME = QB_KMF + BQ_NS, QB_SF = QB_KME + BQ_NS, QB_DSF = QB_SF + 2,
       QB_DONE = QB_DSF + 2, QB_COUNT = QB_DONE + 1 };

__global__ void __launch_bounds__(BWD_THREADS, 1) attn_bwd_dq_kernel(const __grid_constant__ AttnP p, const __grid_constant__ CUtensorMap tmQ,
                                                                      const __grid_constant__ CUtensorMap tmdO, const __grid_constant__ CUtensorMap tmK,
                                                                      const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmKmn) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = align1024(smem_raw);
  uint8_t* sQ = sm; uint8_t* sdO = sQ + TILE32K; uint8_t* stage0 = sdO + TILE32K;
  uint64_t* bars = reinterpret_cast<uint64_t*>(stage0 + BQ_NS * BQ_STAGE);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + QB_COUNT);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int bh = blockIdx.y, b = bh / p.heads, h = bh - b * p.heads;
  const int q0 = blockIdx.x * 128;
  const int ntiles = cdiv(p.Tc, 64);

  if (tid == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmdO); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmKmn);
    for (int i = 0; i < QB_COUNT; ++i) mbar_init(&bars[i], (i == QB_DSF || i == QB_DSF + 1) ? 256 : 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait(); pdl_trigger();      // everything above touches no global memory
  const uint32_t tmem = *tmem_slot;                      // columns: S0 0, S1 64, dP0 128, dP1 192, dQ 256

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(&bars[QB_QDO], 2 * TILE32K);
      tma_tile(sQ, &tmQ, &bars[QB_QDO], h * HD, q0, b, 128);
      tma_tile(sdO, &tmdO, &bars[QB_QDO], h * HD, q0, b, 128);
      for (int j = 0; j < ntiles; ++j) {
        const int st = j % BQ_NS; const uint32_t ph = (j / BQ_NS) & 1;
        uint8_t* sg = stage0 + st * BQ_STAGE;
        mbar_wait(&bars[QB_KVE + st], ph ^ 1);
        mbar_arrive_expect_tx(&bars[QB_KVF + st], 2 * TILE16K);
        tma_tile(sg, &tmK, &bars[QB_KVF + st], h * HD, j * 64, b, 64);
        tma_tile(sg + TILE16K, &tmV, &bars[QB_KVF + st], h * HD, j * 64, b, 64);
        mbar_wait(&bars[QB_KME + st], ph ^ 1);
        mbar_arrive_expect_tx(&bars[QB_KMF + st], TILE16K);
        tma_tile(sg + 2 * TILE16K, &tmKmn, &bars[QB_KMF + st], h * HD, j * 64, b, 64);
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc_kk = make_idesc(kFmtTF32, 128, 64, 0, 0);
    constexpr uint32_t idesc_kmn = make_idesc(kFmtTF32, 128, 64, 0, 1);
    const uint64_t dQ = desc_kmajor(smem_u32(sQ)), ddO = desc_kmajor(smem_u32(sdO));
    uint64_t dK[BQ_NS], dV[BQ_NS], dKmn[BQ_NS];
#pragma unroll
    for (int i = 0; i < BQ_NS; ++i) {
      const uint32_t sg = smem_u32(stage0 + i * BQ_STAGE);
      dK[i] = desc_kmajor(sg); dV[i] = desc_kmajor(sg + TILE16K); dKmn[i] = desc_mnmajor(sg + 2 * TILE16K);
    }
    mbar_wait(&bars[QB_QDO], 0);
    int st_s = 0; uint32_t ph_s = 0;                    // shared-memory stage / phase of the next S, dP product
    auto issue_s = [&](int j) {
      const int tb = j & 1;
      mbar_wait(&bars[QB_KVF + st_s], ph_s);
      ATTN_TRACE(j * 8 + 0);
      tc_fence_after();
      if (elect_one()) {
        mma_kk<128, 64>(tmem + tb * 64, dQ, dK[st_s], idesc_kk);              // S  = Q K^T
        mma_kk<128, 64>(tmem + 128 + tb * 64, ddO, dV[st_s], idesc_kk);       // dP = dO V^T
        mma_commit(&bars[QB_KVE + st_s]);
        mma_commit(&bars[QB_SF + tb]);
      }
      __syncwarp();
      if (++st_s == BQ_NS) { st_s = 0; ph_s ^= 1; }
    };
    issue_s(0);
    int st = 0; uint32_t ph = 0;
    for (int j = 0; j < ntiles; ++j) {
      if (j + 1 < ntiles) issue_s(j + 1);
      const int tb = j & 1;
      ATTN_TRACE(j * 8 + 1);
      mbar_wait(&bars[QB_DSF + tb], (j >> 1) & 1);
      ATTN_TRACE(j * 8 + 2);
      mbar_wait(&bars[QB_KMF + st], ph);
      ATTN_TRACE(j * 8 + 3);
      tc_fence_after();
      if (elect_one()) {
        mma_tmn(tmem + 256, tmem + 128 + tb * 64, dKmn[st], idesc_kmn, j > 0);   // dQ += dS K
        mma_commit(&bars[QB_KME + st]);
        if (j == ntiles - 1) mma_commit(&bars[QB_DONE]);
      }
      __syncwarp();
      if (++st == BQ_NS) { st = 0; ph ^= 1; }
    }
  } else {
    const int quad = warp & 3, half = (warp - 2) >> 2;
    const int q = q0 + quad * 32 + lane;
    const uint32_t t_row = tmem + ((uint32_t)(quad * 32) << 16) + half * 32;
    const float lse2 = q < p.Tc ? p.LSE[(long long)bh * p.Tc + q] * kLog2e : 0.f;
    const float dd = q < p.Tc ? p.Dd[(long long)bh * p.Tc + q] : 0.f;
    for (int j = 0; j < ntiles; ++j) {
      const int st = j & 1;
      if (warp == 2) ATTN_TRACE(j * 8 + 4);
      mbar_wait(&bars[QB_SF + st], (j >> 1) & 1);
      if (warp == 2) ATTN_TRACE(j * 8 + 5);
      tc_fence_after();
      uint32_t s[32], dp[32];
      tmem_ld32_nowait(t_row + st * 64, s);
      tmem_ld32_nowait(t_row + 128 + st * 64, dp);
      tmem_wait_ld();
      if (warp == 2) ATTN_TRACE(j * 8 + 6);
#pragma unroll
      for (int c = 0; c < 32; ++c)                                                       // dS = P * (dP - D)
        dp[c] = tf32_rna_bits(fast_ex2(fmaf(__uint_as_float(s[c]), kLog2e, -lse2)) * (__uint_as_float(dp[c]) - dd));
      tmem_st32(t_row + 128 + st * 64, dp);
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(&bars[QB_DSF + st]);
      if (warp == 2) ATTN_TRACE(j * 8 + 7);
    }
    mbar_wait(&bars[QB_DONE], 0);
    tc_fence_after();
    uint32_t acc[32];
    tmem_ld32_nowait(t_row + 256, acc);
    tmem_wait_ld();
    if (q < p.Tc) {
      float* dst = p.dQ + ((long long)b * p.Tc + q) * p.D + h * HD + half * 32;
#pragma unroll
      for (int c = 0; c < 8; ++c) *reinterpret_cast<float4*>(dst + 4 * c) = out4(acc + 4 * c, p.dq_scale, p.round_out);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ---- dK, dV: CTA = 128 key rows, loop over 64-query tiles; works on S^T ------------------------------------------------
constexpr int BK_STAGE = 4 * TILE16K;                  // Q (K-major), dO (K-major), Q (MN-major), dO (MN-major)
constexpr int BK_ROWSTAT = 2048;                        // LSE / D of up to this many queries are staged in shared memory once per CTA
constexpr int BK_SMEM = 1024 + 2 * TILE32K + 2 * BK_STAGE + 128 + 2 * BK_ROWSTAT * 4;
enum { KB_KV = 0, KB_QF = 1, KB_QE = 3, KB_MF = 5, KB_ME = 7, KB_SF = 9, KB_DSF = 11, KB_DONE = 13, KB_COUNT = 14 };

__global__ void __launch_bounds__(BWD_THREADS, 1) attn_bwd_dkv_kernel(const __grid_constant__ AttnP p, const __grid_constant__ CUtensorMap tmK,
                                                                       const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmQ,
                                                                       const __grid_constant__ CUtensorMap tmdO, const __grid_constant__ CUtensorMap tmQmn,
                                                                       const __grid_constant__ CUtensorMap tmdOmn) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = align1024(smem_raw);
  uint8_t* sK = sm; uint8_t* sV = sK + TILE32K; uint8_t* stage0 = sV + TILE32K;
  uint64_t* bars = reinterpret_cast<uint64_t*>(stage0 + 2 * BK_STAGE);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + KB_COUNT);
  float* s_lse = reinterpret_cast<float*>(stage0 + 2 * BK_STAGE + 128); float* s_dd = s_lse + BK_ROWSTAT;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int bh = blockIdx.y, b = bh / p.heads, h = bh - b * p.heads;
  const int kv0 = blockIdx.x * 128;
  const int ntiles = cdiv(p.Tc, 64);

  if (tid == 0) {
    tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmdO); tma_prefetch_desc(&tmQmn); tma_prefetch_desc(&tmdOmn);
    for (int i = 0; i < KB_COUNT; ++i) mbar_init(&bars[i], (i == KB_DSF || i == KB_DSF + 1) ? 256 : 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait(); pdl_trigger();      // everything above touches no global memory
  const uint32_t tmem = *tmem_slot;                      // columns: S^T0 0, S^T1 64, dP^T0 128, dP^T1 192, dV 256, dK 320

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(&bars[KB_KV], 2 * TILE32K);
      tma_tile(sK, &tmK, &bars[KB_KV], h * HD, kv0, b, 128);
      tma_tile(sV, &tmV, &bars[KB_KV], h * HD, kv0, b, 128);
      for (int i = 0; i < ntiles; ++i) {
        const int st = i & 1; const uint32_t ph = (i >> 1) & 1;
        uint8_t* sg = stage0 + st * BK_STAGE;
        mbar_wait(&bars[KB_QE + st], ph ^ 1);
        mbar_arrive_expect_tx(&bars[KB_QF + st], 2 * TILE16K);
        tma_tile(sg, &tmQ, &bars[KB_QF + st], h * HD, i * 64, b, 64);
        tma_tile(sg + TILE16K, &tmdO, &bars[KB_QF + st], h * HD, i * 64, b, 64);
        mbar_wait(&bars[KB_ME + st], ph ^ 1);
        mbar_arrive_expect_tx(&bars[KB_MF + st], 2 * TILE16K);
        tma_tile(sg + 2 * TILE16K, &tmQmn, &bars[KB_MF + st], h * HD, i * 64, b, 64);
        tma_tile(sg + 3 * TILE16K, &tmdOmn, &bars[KB_MF + st], h * HD, i * 64, b, 64);
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc_kk = make_idesc(kFmtTF32, 128, 64, 0, 0);
    constexpr uint32_t idesc_kmn = make_idesc(kFmtTF32, 128, 64, 0, 1);
    const uint64_t dK = desc_kmajor(smem_u32(sK)), dV = desc_kmajor(smem_u32(sV));
    uint64_t dQ[2], ddO[2], dQmn[2], ddOmn[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const uint32_t sg = smem_u32(stage0 + i * BK_STAGE);
      dQ[i] = desc_kmajor(sg); ddO[i] = desc_kmajor(sg + TILE16K); dQmn[i] = desc_mnmajor(sg + 2 * TILE16K); ddOmn[i] = desc_mnmajor(sg + 3 * TILE16K);
    }
    mbar_wait(&bars[KB_KV], 0);
    auto issue_s = [&](int i) {
      const int st = i & 1;
      mbar_wait(&bars[KB_QF + st], (i >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        mma_kk<128, 64>(tmem + st * 64, dK, dQ[st], idesc_kk);                // S^T  = K Q^T
        mma_kk<128, 64>(tmem + 128 + st * 64, dV, ddO[st], idesc_kk);         // dP^T = V dO^T
        mma_commit(&bars[KB_QE + st]);
        mma_commit(&bars[KB_SF + st]);
      }
      __syncwarp();
    };
    issue_s(0);
    for (int i = 0; i < ntiles; ++i) {
      if (i + 1 < ntiles) issue_s(i + 1);
      const int st = i & 1; const uint32_t ph = (i >> 1) & 1;
      mbar_wait(&bars[KB_DSF + st], ph);
      mbar_wait(&bars[KB_MF + st], ph);
      tc_fence_after();
      if (elect_one()) {
        mma_tmn(tmem + 256, tmem + st * 64, ddOmn[st], idesc_kmn, i > 0);          // dV += P^T dO
        mma_tmn(tmem + 320, tmem + 128 + st * 64, dQmn[st], idesc_kmn, i > 0);     // dK += dS^T Q
        mma_commit(&bars[KB_ME + st]);
        if (i == ntiles - 1) mma_commit(&bars[KB_DONE]);
      }
      __syncwarp();
    }
  } else {
    const int quad = warp & 3, half = (warp - 2) >> 2;
    const int kv = kv0 + quad * 32 + lane;
    const uint32_t t_row = tmem + ((uint32_t)(quad * 32) << 16) + half * 32;
    const float* LSE = p.LSE + (long long)bh * p.Tc; const float* Dd = p.Dd + (long long)bh * p.Tc;
    const bool staged = ntiles * 64 <= BK_ROWSTAT;
    if (staged) {                                        // LSE * log2(e) and D of every query of this (image, head), zero past Tc
      for (int t = tid - 64; t < ntiles * 64; t += BWD_THREADS - 64) {
        const bool ok = t < p.Tc;
        s_lse[t] = ok ? __ldg(LSE + t) * kLog2e : 0.f; s_dd[t] = ok ? __ldg(Dd + t) : 0.f;
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");     // the 8 row warps only
    }
    for (int i = 0; i < ntiles; ++i) {
      const int st = i & 1;
      const int qc = i * 64 + half * 32;                 // this thread's 32 tile columns are queries qc .. qc + 31 (warp-uniform addresses)
      float lse2[32], dd[32];
      if (staged) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float4 a = *reinterpret_cast<const float4*>(s_lse + qc + 4 * c), d4 = *reinterpret_cast<const float4*>(s_dd + qc + 4 * c);
          lse2[4 * c] = a.x; lse2[4 * c + 1] = a.y; lse2[4 * c + 2] = a.z; lse2[4 * c + 3] = a.w;
          dd[4 * c] = d4.x; dd[4 * c + 1] = d4.y; dd[4 * c + 2] = d4.z; dd[4 * c + 3] = d4.w;
        }
      } else {
#pragma unroll
        for (int c = 0; c < 32; ++c) { const bool ok = qc + c < p.Tc; lse2[c] = ok ? __ldg(LSE + qc + c) * kLog2e : 0.f; dd[c] = ok ? __ldg(Dd + qc + c) : 0.f; }
      }
      mbar_wait(&bars[KB_SF + st], (i >> 1) & 1);
      tc_fence_after();
      uint32_t s[32], dp[32];
      tmem_ld32_nowait(t_row + st * 64, s);
      tmem_ld32_nowait(t_row + 128 + st * 64, dp);
      tmem_wait_ld();
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        const float pr = fast_ex2(fmaf(__uint_as_float(s[c]), kLog2e, -lse2[c]));
        dp[c] = tf32_rna_bits(pr * (__uint_as_float(dp[c]) - dd[c]));
        s[c] = tf32_rna_bits(pr);
      }
      tmem_st32(t_row + st * 64, s);
      tmem_st32(t_row + 128 + st * 64, dp);
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(&bars[KB_DSF + st]);
    }
    mbar_wait(&bars[KB_DONE], 0);
    tc_fence_after();
    uint32_t acc[32];
    tmem_ld32_nowait(t_row + 256, acc);
    tmem_wait_ld();
    if (kv < p.Tc) {
      float* dst = p.dV + ((long long)b * p.Tc + kv) * p.D + h * HD + half * 32;
#pragma unroll
      for (int c = 0; c < 8; ++c) *reinterpret_cast<float4*>(dst + 4 * c) = out4(acc + 4 * c, 1.f, p.round_out);
    }
    tmem_ld32_nowait(t_row + 320, acc);
    tmem_wait_ld();
    if (kv < p.Tc) {
      float* dst = p.dK + ((long long)b * p.Tc + kv) * p.D + h * HD + half * 32;
#pragma unroll
      for (int c = 0; c < 8; ++c) *reinterpret_cast<float4*>(dst + 4 * c) = out4(acc + 4 * c, 1.f, p.round_out);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// =====================================================================================================================
// backward, 128-wide score tiles (Tc > 64).  The S / dP products above run with N = 64, where a 128 x N x 8 tcgen05.mma costs 48
// cycles whatever N is (shared-memory operand bandwidth: 4 KB of A per instruction; tools/micro/mma_rate.cu: N = 64 -> 48, N = 128 -> 64,
// N = 256 -> 128 cycles); with N = 128 the same instruction count does twice the work at the tensor core's own rate.  TMEM holds
// 512 columns, so only the S tile is double-buffered (2 x 128) and dP is single (128): the issuer orders the products so that the
// row threads still overlap - S_{j+1} is issued BEFORE the products that consume tile j, the row threads exponentiate S_{j+1} while
// the tensor core works on those, and only the short dS = P * (dP - D) step sits between dP_j and its consumers.
// =====================================================================================================================
// Both kernels are PERSISTENT: one CTA per SM walks the work items (128-row block, image, head) w = blockIdx.x, + gridDim.x, ...  A CTA of
// this size spends ~7k cycles outside its steady state (launch, TMEM allocation, barrier set-up, the first 128 KB of operands, the
// accumulator read-out) against ~18-24k inside; the roles therefore run straight on into the next item - every barrier keeps its phase
// across items (indexed by the global tile count g = item * ntiles + tile), the producer refills each buffer the moment the item's last
// product has read it, and the tensor core starts the next item's S / dP while the row threads still store the previous gradients.
// ---- dQ: work item = 128 query rows, loop over 128-key tiles -----------------------------------------------------------
constexpr int Q2_SMEM = 1024 + 2 * TILE32K + 2 * (2 * TILE32K) + TILE32K + 256;      // Q, dO | 2 stages of (K, V) K-major | K MN-major (two 64-key sub-tiles)
enum { Q2_QDO = 0, Q2_QDOE = 1, Q2_KVF = 2, Q2_KVE = 4, Q2_KMF = 6, Q2_KME = 8, Q2_SF = 10, Q2_DPF = 12, Q2_DSF = 13, Q2_DONE = 14, Q2_ACCE = 15, Q2_COUNT = 16 };

__global__ void __launch_bounds__(BWD_THREADS, 1) attn_bwd_dq128_kernel(const __grid_constant__ AttnP p, const __grid_constant__ CUtensorMap tmQ,
                                                                         const __grid_constant__ CUtensorMap tmdO, const __grid_constant__ CUtensorMap tmK,
                                                                         const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmKmn,
                                                                         int nblk, int nitems) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = align1024(smem_raw);
  uint8_t* sQ = sm; uint8_t* sdO = sQ + TILE32K; uint8_t* sKV = sdO + TILE32K; uint8_t* sKmn = sKV + 4 * TILE32K;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKmn + TILE32K);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + Q2_COUNT);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles = cdiv(p.Tc, 128);

  if (tid == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmdO); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmKmn);
    for (int i = 0; i < Q2_COUNT; ++i) mbar_init(&bars[i], (i == Q2_DSF || i == Q2_ACCE) ? 256 : 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait(); pdl_trigger();      // everything above touches no global memory
  const uint32_t tmem = *tmem_slot;                      // columns: S0 0, S1 128, dP 256, dQ 384

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;
      for (int w = blockIdx.x; w < nitems; w += gridDim.x, ++it) {
        const int bh = w / nblk, q0 = (w - bh * nblk) * 128, b = bh / p.heads, h = bh - b * p.heads;
        const int g0 = it * ntiles;
        mbar_wait(&bars[Q2_QDOE], (it & 1) ^ 1);         // the previous item's last S / dP product has read Q and dO
        mbar_arrive_expect_tx(&bars[Q2_QDO], 2 * TILE32K);
        tma_tile(sQ, &tmQ, &bars[Q2_QDO], h * HD, q0, b, 128);
        tma_tile(sdO, &tmdO, &bars[Q2_QDO], h * HD, q0, b, 128);
        auto load_kv = [&](int j) {
          const int g = g0 + j, st = g & 1; const uint32_t ph = (g >> 1) & 1;
          mbar_wait(&bars[Q2_KVE + st], ph ^ 1);
          mbar_arrive_expect_tx(&bars[Q2_KVF + st], 2 * TILE32K);
          tma_tile(sKV + st * 2 * TILE32K, &tmK, &bars[Q2_KVF + st], h * HD, j * 128, b, 128);
          tma_tile(sKV + st * 2 * TILE32K + TILE32K, &tmV, &bars[Q2_KVF + st], h * HD, j * 128, b, 128);
        };
        load_kv(0);
        for (int j = 0; j < ntiles; ++j) {
          if (j + 1 < ntiles) load_kv(j + 1);            // gated by dP_{j-1}: long done when dQ_{j-1} (the next wait) completes
          for (int sub = 0; sub < 2; ++sub) {            // the MN-major K tile is two independent 64-key sub-tiles: each is refilled as soon as its 8 products are done
            mbar_wait(&bars[Q2_KME + sub], ((g0 + j) & 1) ^ 1);
            mbar_arrive_expect_tx(&bars[Q2_KMF + sub], TILE16K);
            tma_tile(sKmn + sub * TILE16K, &tmKmn, &bars[Q2_KMF + sub], h * HD, j * 128 + sub * 64, b, 64);
          }
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc_s = make_idesc(kFmtTF32, 128, 128, 0, 0);
    constexpr uint32_t idesc_q = make_idesc(kFmtTF32, 128, 64, 0, 1);
    const uint64_t dQ = desc_kmajor(smem_u32(sQ)), ddO = desc_kmajor(smem_u32(sdO));
    const uint64_t dK[2] = {desc_kmajor(smem_u32(sKV)), desc_kmajor(smem_u32(sKV + 2 * TILE32K))};
    const uint64_t dV[2] = {desc_kmajor(smem_u32(sKV + TILE32K)), desc_kmajor(smem_u32(sKV + 3 * TILE32K))};
    const uint64_t dKmn[2] = {desc_mnmajor(smem_u32(sKmn)), desc_mnmajor(smem_u32(sKmn + TILE16K))};
    int it = 0;
    for (int w = blockIdx.x; w < nitems; w += gridDim.x, ++it) {
      const int g0 = it * ntiles;
      mbar_wait(&bars[Q2_QDO], it & 1);
      mbar_wait(&bars[Q2_KVF + (g0 & 1)], (g0 >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        mma_kk<128, 128>(tmem + (g0 & 1) * 128, dQ, dK[g0 & 1], idesc_s);       // S_0 = Q K_0^T
        mma_commit(&bars[Q2_SF + (g0 & 1)]);
      }
      __syncwarp();
      for (int j = 0; j < ntiles; ++j) {
        const int g = g0 + j, st = g & 1;
        if (elect_one()) {
          mma_kk<128, 128>(tmem + 256, ddO, dV[st], idesc_s);                   // dP_j = dO V_j^T (after dQ_{j-1}, which read dS_{j-1} from these columns)
          mma_commit(&bars[Q2_KVE + st]);
          mma_commit(&bars[Q2_DPF]);
          if (j + 1 == ntiles) mma_commit(&bars[Q2_QDOE]);                      // the item's last product that reads Q / dO
        }
        __syncwarp();
        if (j + 1 < ntiles) {
          mbar_wait(&bars[Q2_KVF + (st ^ 1)], ((g + 1) >> 1) & 1);
          tc_fence_after();
          if (elect_one()) {
            mma_kk<128, 128>(tmem + (st ^ 1) * 128, dQ, dK[st ^ 1], idesc_s);   // S_{j+1}
            mma_commit(&bars[Q2_SF + (st ^ 1)]);
          }
          __syncwarp();
        }
        mbar_wait(&bars[Q2_DSF], g & 1);
        if (j == 0 && it > 0) mbar_wait(&bars[Q2_ACCE], (it - 1) & 1);           // the row threads have read the previous item's dQ
        mbar_wait(&bars[Q2_KMF], g & 1);
        tc_fence_after();
        if (elect_one()) {
          mma_tmn(tmem + 384, tmem + 256, dKmn[0], idesc_q, j > 0);             // dQ += dS_j K_j (keys 0..63, then 64..127 of the tile)
          mma_commit(&bars[Q2_KME]);
        }
        __syncwarp();
        mbar_wait(&bars[Q2_KMF + 1], g & 1);
        tc_fence_after();
        if (elect_one()) {
          mma_tmn(tmem + 384, tmem + 256 + 64, dKmn[1], idesc_q, true);
          mma_commit(&bars[Q2_KME + 1]);
          if (j == ntiles - 1) mma_commit(&bars[Q2_DONE]);
        }
        __syncwarp();
      }
    }
  } else {
    const int quad = warp & 3, half = (warp - 2) >> 2;
    const uint32_t t_lane = tmem + ((uint32_t)(quad * 32) << 16);
    const uint32_t t_row = t_lane + half * 64;
    // The finished rows of an item leave one float4 per tile of the NEXT item: written in one burst after the accumulator read-out they held the
    // row threads for ~3.8k cycles (32 half-filled sectors per instruction) exactly while the tensor core waited for the next item's first dS.
    float4 pend[8];
    float* pend_dst = nullptr;
    const int cpt = cdiv(8, ntiles);                     // pending float4 chunks stored per tile
    auto row_stats = [&](int w, float& lse2, float& dd) {
      const int bh = w / nblk, q = (w - bh * nblk) * 128 + quad * 32 + lane;
      const bool ok = w < nitems && q < p.Tc;
      lse2 = ok ? __ldg(p.LSE + (long long)bh * p.Tc + q) * kLog2e : 0.f;
      dd = ok ? __ldg(p.Dd + (long long)bh * p.Tc + q) : 0.f;
    };
    float lse2, dd;
    row_stats(blockIdx.x, lse2, dd);
    int it = 0;
    for (int w = blockIdx.x; w < nitems; w += gridDim.x, ++it) {
      const int bh = w / nblk, q0 = (w - bh * nblk) * 128, b = bh / p.heads, h = bh - b * p.heads;
      const int g0 = it * ntiles;
      const int q = q0 + quad * 32 + lane;
      for (int j = 0; j < ntiles; ++j) {
        const int g = g0 + j, sb = g & 1;
        mbar_wait(&bars[Q2_SF + sb], (g >> 1) & 1);
        tc_fence_after();
        uint32_t s[64];
        tmem_ld32_nowait(t_row + sb * 128, s);
        tmem_ld32_nowait(t_row + sb * 128 + 32, s + 32);
        tmem_wait_ld();
#pragma unroll
        for (int c = 0; c < 64; ++c) s[c] = __float_as_uint(fast_ex2(fmaf(__uint_as_float(s[c]), kLog2e, -lse2)));      // P
        mbar_wait(&bars[Q2_DPF], g & 1);
        tc_fence_after();
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          uint32_t dp[32];
          tmem_ld32_nowait(t_row + 256 + hh * 32, dp);
          tmem_wait_ld();
#pragma unroll
          for (int c = 0; c < 32; ++c) dp[c] = tf32_rna_bits(__uint_as_float(s[hh * 32 + c]) * (__uint_as_float(dp[c]) - dd));      // dS = P * (dP - D)
          tmem_st32(t_row + 256 + hh * 32, dp);
        }
        tmem_wait_st();
        tc_fence_before();
        mbar_arrive(&bars[Q2_DSF]);
        if (pend_dst) {
#pragma unroll
          for (int c = 0; c < 8; ++c)
            if (c / cpt == j) *reinterpret_cast<float4*>(pend_dst + 4 * c) = pend[c];
        }
      }
      float lse2n, ddn;
      row_stats(w + gridDim.x, lse2n, ddn);              // the next item's row statistics, fetched under the last dQ product
      mbar_wait(&bars[Q2_DONE], it & 1);
      tc_fence_after();
      uint32_t acc[32];
      tmem_ld32_nowait(t_lane + 384 + half * 32, acc);
      tmem_wait_ld();
      tc_fence_before();
      mbar_arrive(&bars[Q2_ACCE]);                       // the accumulator may be overwritten by the next item
#pragma unroll
      for (int c = 0; c < 8; ++c) pend[c] = out4(acc + 4 * c, p.dq_scale, p.round_out);
      pend_dst = q < p.Tc ? p.dQ + ((long long)b * p.Tc + q) * p.D + h * HD + half * 32 : nullptr;
      lse2 = lse2n; dd = ddn;
    }
    if (pend_dst) {
#pragma unroll
      for (int c = 0; c < 8; ++c) *reinterpret_cast<float4*>(pend_dst + 4 * c) = pend[c];
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ---- dK, dV: work item = 128 key rows, loop over 128-query tiles; works on S^T ---------------------------------------
constexpr int K2_SMEM = 1024 + 2 * TILE32K + 4 * TILE32K + 256 + 2 * BK_ROWSTAT * 4;      // K, V | Q, dO K-major | Q, dO MN-major (all single-buffered)
enum { K2_KV = 0, K2_KVE = 1, K2_QF = 2, K2_QE = 3, K2_OF = 4, K2_OE = 5, K2_OMF = 6, K2_OME = 8, K2_QMF = 10, K2_QME = 12, K2_SF = 14, K2_DPF = 16, K2_DSF = 17,
       K2_DONE = 18, K2_ACCE = 19, K2_COUNT = 20 };

__global__ void __launch_bounds__(BWD_THREADS, 1) attn_bwd_dkv128_kernel(const __grid_constant__ AttnP p, const __grid_constant__ CUtensorMap tmK,
                                                                          const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmQ,
                                                                          const __grid_constant__ CUtensorMap tmdO, const __grid_constant__ CUtensorMap tmQmn,
                                                                          const __grid_constant__ CUtensorMap tmdOmn, int nblk, int nitems) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = align1024(smem_raw);
  uint8_t* sK = sm; uint8_t* sV = sK + TILE32K; uint8_t* sQk = sV + TILE32K; uint8_t* sOk = sQk + TILE32K; uint8_t* sQm = sOk + TILE32K; uint8_t* sOm = sQm + TILE32K;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sOm + TILE32K);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + K2_COUNT);
  float* s_lse = reinterpret_cast<float*>(sOm + TILE32K + 256); float* s_dd = s_lse + BK_ROWSTAT;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles = cdiv(p.Tc, 128);

  if (tid == 0) {
    tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmdO); tma_prefetch_desc(&tmQmn); tma_prefetch_desc(&tmdOmn);
    for (int i = 0; i < K2_COUNT; ++i) mbar_init(&bars[i], (i == K2_DSF || i == K2_ACCE) ? 256 : 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait(); pdl_trigger();      // everything above touches no global memory
  const uint32_t tmem = *tmem_slot;                      // columns: S^T0 0, S^T1 128, dP^T 256, dV 384, dK 448

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;
      for (int w = blockIdx.x; w < nitems; w += gridDim.x, ++it) {
        const int bh = w / nblk, kv0 = (w - bh * nblk) * 128, b = bh / p.heads, h = bh - b * p.heads;
        const int g0 = it * ntiles;
        mbar_wait(&bars[K2_KVE], (it & 1) ^ 1);          // the previous item's last S / dP product has read K and V
        mbar_arrive_expect_tx(&bars[K2_KV], 2 * TILE32K);
        tma_tile(sK, &tmK, &bars[K2_KV], h * HD, kv0, b, 128);
        tma_tile(sV, &tmV, &bars[K2_KV], h * HD, kv0, b, 128);
        // every operand tile is single-buffered: the loads are issued in the order their buffers come free on the tensor pipe
        // (S_0, dP_0, S_1, dV_0 dK_0, dP_1, S_2, dV_1 dK_1, ...)
        auto load_q = [&](int i) {
          mbar_wait(&bars[K2_QE], ((g0 + i) & 1) ^ 1);
          mbar_arrive_expect_tx(&bars[K2_QF], TILE32K);
          tma_tile(sQk, &tmQ, &bars[K2_QF], h * HD, i * 128, b, 128);
        };
        auto load_o = [&](int i) {
          mbar_wait(&bars[K2_OE], ((g0 + i) & 1) ^ 1);
          mbar_arrive_expect_tx(&bars[K2_OF], TILE32K);
          tma_tile(sOk, &tmdO, &bars[K2_OF], h * HD, i * 128, b, 128);
        };
        auto load_m = [&](int i) {                       // MN-major tiles: four independent 64-query sub-tiles, each refilled as soon as its 8 products are done
          for (int sub = 0; sub < 2; ++sub) {
            mbar_wait(&bars[K2_OME + sub], ((g0 + i) & 1) ^ 1);
            mbar_arrive_expect_tx(&bars[K2_OMF + sub], TILE16K);
            tma_tile(sOm + sub * TILE16K, &tmdOmn, &bars[K2_OMF + sub], h * HD, i * 128 + sub * 64, b, 64);
          }
          for (int sub = 0; sub < 2; ++sub) {
            mbar_wait(&bars[K2_QME + sub], ((g0 + i) & 1) ^ 1);
            mbar_arrive_expect_tx(&bars[K2_QMF + sub], TILE16K);
            tma_tile(sQm + sub * TILE16K, &tmQmn, &bars[K2_QMF + sub], h * HD, i * 128 + sub * 64, b, 64);
          }
        };
        load_q(0); load_o(0); load_m(0);
        if (ntiles > 1) load_q(1);
        for (int i = 1; i < ntiles; ++i) {
          load_o(i);                                     // after dP_{i-1}
          if (i + 1 < ntiles) load_q(i + 1);             // after S_i
          load_m(i);                                     // after dV_{i-1}, then dK_{i-1}
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc_s = make_idesc(kFmtTF32, 128, 128, 0, 0);
    constexpr uint32_t idesc_t = make_idesc(kFmtTF32, 128, 64, 0, 1);
    const uint64_t dK = desc_kmajor(smem_u32(sK)), dV = desc_kmajor(smem_u32(sV));
    const uint64_t dQk = desc_kmajor(smem_u32(sQk)), dOk = desc_kmajor(smem_u32(sOk));
    const uint64_t dQm[2] = {desc_mnmajor(smem_u32(sQm)), desc_mnmajor(smem_u32(sQm + TILE16K))};
    const uint64_t dOm[2] = {desc_mnmajor(smem_u32(sOm)), desc_mnmajor(smem_u32(sOm + TILE16K))};
    int it = 0;
    for (int w = blockIdx.x; w < nitems; w += gridDim.x, ++it) {
      const int g0 = it * ntiles;
      mbar_wait(&bars[K2_KV], it & 1);
      mbar_wait(&bars[K2_QF], g0 & 1);
      tc_fence_after();
      if (elect_one()) {
        mma_kk<128, 128>(tmem + (g0 & 1) * 128, dK, dQk, idesc_s);              // S^T_0 = K Q_0^T
        mma_commit(&bars[K2_QE]);
        mma_commit(&bars[K2_SF + (g0 & 1)]);
      }
      __syncwarp();
      for (int i = 0; i < ntiles; ++i) {
        const int g = g0 + i, sb = g & 1;
        mbar_wait(&bars[K2_OF], g & 1);
        tc_fence_after();
        if (elect_one()) {
          mma_kk<128, 128>(tmem + 256, dV, dOk, idesc_s);                       // dP^T_i = V dO_i^T (after dK_{i-1}, which read dS^T_{i-1} from these columns)
          mma_commit(&bars[K2_OE]);
          mma_commit(&bars[K2_DPF]);
          if (i + 1 == ntiles) mma_commit(&bars[K2_KVE]);                       // the item's last product that reads K / V
        }
        __syncwarp();
        if (i + 1 < ntiles) {
          mbar_wait(&bars[K2_QF], (g + 1) & 1);
          tc_fence_after();
          if (elect_one()) {
            mma_kk<128, 128>(tmem + (sb ^ 1) * 128, dK, dQk, idesc_s);          // S^T_{i+1}
            mma_commit(&bars[K2_QE]);
            mma_commit(&bars[K2_SF + (sb ^ 1)]);
          }
          __syncwarp();
        }
        mbar_wait(&bars[K2_DSF], g & 1);
        if (i == 0 && it > 0) mbar_wait(&bars[K2_ACCE], (it - 1) & 1);           // the row threads have read the previous item's dV, dK
#pragma unroll
        for (int sub = 0; sub < 2; ++sub) {                                      // dV += P^T dO
          mbar_wait(&bars[K2_OMF + sub], g & 1);
          tc_fence_after();
          if (elect_one()) {
            mma_tmn(tmem + 384, tmem + sb * 128 + sub * 64, dOm[sub], idesc_t, i > 0 || sub > 0);
            mma_commit(&bars[K2_OME + sub]);
          }
          __syncwarp();
        }
#pragma unroll
        for (int sub = 0; sub < 2; ++sub) {                                      // dK += dS^T Q
          mbar_wait(&bars[K2_QMF + sub], g & 1);
          tc_fence_after();
          if (elect_one()) {
            mma_tmn(tmem + 448, tmem + 256 + sub * 64, dQm[sub], idesc_t, i > 0 || sub > 0);
            mma_commit(&bars[K2_QME + sub]);
            if (i == ntiles - 1 && sub == 1) mma_commit(&bars[K2_DONE]);
          }
          __syncwarp();
        }
      }
    }
  } else {
    const int quad = warp & 3, half = (warp - 2) >> 2;
    const uint32_t t_lane = tmem + ((uint32_t)(quad * 32) << 16);
    const uint32_t t_row = t_lane + half * 64;
    const bool staged = ntiles * 128 <= BK_ROWSTAT;
    int it = 0, staged_bh = -1;
    for (int w = blockIdx.x; w < nitems; w += gridDim.x, ++it) {
      const int bh = w / nblk, kv0 = (w - bh * nblk) * 128, b = bh / p.heads, h = bh - b * p.heads;
      const int g0 = it * ntiles;
      const int kv = kv0 + quad * 32 + lane;
      const float* LSE = p.LSE + (long long)bh * p.Tc; const float* Dd = p.Dd + (long long)bh * p.Tc;
      if (staged && bh != staged_bh) {                   // LSE * log2(e) and D of every query of this (image, head), zero past Tc
        asm volatile("bar.sync 1, 256;" ::: "memory");   // (every row warp is done with the previous head's statistics)
        for (int t = tid - 64; t < ntiles * 128; t += BWD_THREADS - 64) {
          const bool ok = t < p.Tc;
          s_lse[t] = ok ? __ldg(LSE + t) * kLog2e : 0.f; s_dd[t] = ok ? __ldg(Dd + t) : 0.f;
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");   // the 8 row warps only
        staged_bh = bh;
      }
      for (int i = 0; i < ntiles; ++i) {
        const int g = g0 + i, sb = g & 1;
        int qc = i * 128 + half * 64;                    // this thread's 64 tile columns are queries qc .. qc + 63 (warp-uniform shared-memory addresses)
        if (!staged) {                                   // long sequences: the tile's 128 statistics are staged tile by tile (two alternating slots; one code path below -
          const int t = tid - 64, slot = sb * 128;       // the kernel has to stay inside the instruction cache)
          const int qq = i * 128 + (t & 127); const bool ok = qq < p.Tc;
          if (t < 128) s_lse[slot + t] = ok ? __ldg(LSE + qq) * kLog2e : 0.f;
          else s_dd[slot + t - 128] = ok ? __ldg(Dd + qq) : 0.f;
          asm volatile("bar.sync 1, 256;" ::: "memory");
          qc = slot + half * 64;
        }
        mbar_wait(&bars[K2_SF + sb], (g >> 1) & 1);
        tc_fence_after();
        uint32_t s[64];
        tmem_ld32_nowait(t_row + sb * 128, s);
        tmem_ld32_nowait(t_row + sb * 128 + 32, s + 32);
        tmem_wait_ld();
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          uint32_t pr[32];
#pragma unroll
          for (int c4 = 0; c4 < 8; ++c4) {
            const int c = hh * 32 + 4 * c4;
            const float4 l4 = *reinterpret_cast<const float4*>(s_lse + qc + c);
            const float p0 = fast_ex2(fmaf(__uint_as_float(s[c]), kLog2e, -l4.x)), p1 = fast_ex2(fmaf(__uint_as_float(s[c + 1]), kLog2e, -l4.y));
            const float p2 = fast_ex2(fmaf(__uint_as_float(s[c + 2]), kLog2e, -l4.z)), p3 = fast_ex2(fmaf(__uint_as_float(s[c + 3]), kLog2e, -l4.w));
            s[c] = __float_as_uint(p0); s[c + 1] = __float_as_uint(p1); s[c + 2] = __float_as_uint(p2); s[c + 3] = __float_as_uint(p3);
            pr[4 * c4] = tf32_rna_bits(p0); pr[4 * c4 + 1] = tf32_rna_bits(p1); pr[4 * c4 + 2] = tf32_rna_bits(p2); pr[4 * c4 + 3] = tf32_rna_bits(p3);
          }
          tmem_st32(t_row + sb * 128 + hh * 32, pr);     // P^T (the A operand of the dV product)
        }
        mbar_wait(&bars[K2_DPF], g & 1);
        tc_fence_after();
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          uint32_t dp[32];
          tmem_ld32_nowait(t_row + 256 + hh * 32, dp);
          tmem_wait_ld();
#pragma unroll
          for (int c4 = 0; c4 < 8; ++c4) {
            const int c = hh * 32 + 4 * c4;
            const float4 d4 = *reinterpret_cast<const float4*>(s_dd + qc + c);
            dp[4 * c4] = tf32_rna_bits(__uint_as_float(s[c]) * (__uint_as_float(dp[4 * c4]) - d4.x));
            dp[4 * c4 + 1] = tf32_rna_bits(__uint_as_float(s[c + 1]) * (__uint_as_float(dp[4 * c4 + 1]) - d4.y));
            dp[4 * c4 + 2] = tf32_rna_bits(__uint_as_float(s[c + 2]) * (__uint_as_float(dp[4 * c4 + 2]) - d4.z));
            dp[4 * c4 + 3] = tf32_rna_bits(__uint_as_float(s[c + 3]) * (__uint_as_float(dp[4 * c4 + 3]) - d4.w));
          }
          tmem_st32(t_row + 256 + hh * 32, dp);          // dS^T (the A operand of the dK product)
        }
        tmem_wait_st();
        tc_fence_before();
        mbar_arrive(&bars[K2_DSF]);
      }
      mbar_wait(&bars[K2_DONE], it & 1);
      tc_fence_after();
      uint32_t accv[32], acck[32];
      tmem_ld32_nowait(t_lane + 384 + half * 32, accv);
      tmem_ld32_nowait(t_lane + 448 + half * 32, acck);
      tmem_wait_ld();
      tc_fence_before();
      mbar_arrive(&bars[K2_ACCE]);                       // the accumulators may be overwritten by the next item
      if (kv < p.Tc) {
        float* dv = p.dV + ((long long)b * p.Tc + kv) * p.D + h * HD + half * 32;
        float* dk = p.dK + ((long long)b * p.Tc + kv) * p.D + h * HD + half * 32;
#pragma unroll
        for (int c = 0; c < 8; ++c) *reinterpret_cast<float4*>(dv + 4 * c) = out4(accv + 4 * c, 1.f, p.round_out);
#pragma unroll
        for (int c = 0; c < 8; ++c) *reinterpret_cast<float4*>(dk + 4 * c) = out4(acck + 4 * c, 1.f, p.round_out);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

__global__ void __launch_bounds__(256) round_tf32_inplace_kernel(float* x, long long n) {
  pdl_wait(); pdl_trigger();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) x[i] = to_tf32(x[i]);
}

cudaError_t configure_once() {
  static bool done = false;
  if (done) return cudaSuccess;
  CVT_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM));
  CVT_CUDA(cudaFuncSetAttribute(attn_bwd_dq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BQ_SMEM));
  CVT_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BK_SMEM));
  CVT_CUDA(cudaFuncSetAttribute(attn_bwd_dq128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Q2_SMEM));
  CVT_CUDA(cudaFuncSetAttribute(attn_bwd_dkv128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, K2_SMEM));
  done = true;
  return cudaSuccess;
}

// tensor maps over one [B, Tc, D] tensor: K-major boxes (32 floats x rows, 128B swizzle) or MN-major boxes (32-byte-atom swizzle)
bool tmap_kmajor(CUtensorMap* tm, const float* x, const AttnP& p, int B, int rows) {
  return make_tmap_3d(tm, x, p.D, p.Tc, B, p.D, (long long)p.Tc * p.D, rows, CU_TENSOR_MAP_SWIZZLE_128B);
}
bool tmap_mnmajor(CUtensorMap* tm, const float* x, const AttnP& p, int B) {
  return make_tmap_3d(tm, x, p.D, p.Tc, B, p.D, (long long)p.Tc * p.D, 64, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
}

cudaError_t run_fwd(const AttnP& p, int B, cudaStream_t s) {
  CVT_CUDA(configure_once());
  if (p.D % 4 != 0 || p.D != p.heads * HD) { set_error("attention_tc: D must equal heads * 64"); return cudaErrorInvalidValue; }
  CUtensorMap tmQ, tmK, tmV;
  if (!tmap_kmajor(&tmQ, p.Q, p, B, 128) || !tmap_kmajor(&tmK, p.K, p, B, 64) || !tmap_mnmajor(&tmV, p.V, p, B)) return cudaErrorInvalidValue;
  CVT_LAUNCH((attn_fwd_kernel), dim3(cdiv(p.Tc, 128), B * p.heads), FWD_THREADS, FWD_SMEM, s, p, tmQ, tmK, tmV);
  return cudaSuccess;
}
// side / ev_fork / ev_join (nullable): the dQ kernel runs on `side` next to the dK/dV kernel.  Both are grids of one ~170 KB CTA per SM, 512 of them on
// 148 SMs at the benched size: 3.46 waves each, i.e. two partial last waves when serialised; side by side the second kernel's CTAs take the SMs the
// first one's last wave leaves idle (6.9 waves in 7 slots instead of 8).  The kernels share only read-only inputs.
cudaError_t run_bwd(const AttnP& p, float* Dd, int B, cudaStream_t s, cudaStream_t side = nullptr, cudaEvent_t ev_fork = nullptr, cudaEvent_t ev_join = nullptr) {
  CVT_CUDA(configure_once());
  const long long groups = (long long)B * p.Tc * p.heads;
  CVT_LAUNCH((attn_dd_kernel), (unsigned)cdivl(groups * 16, 256), 256, 0, s, p.dO, p.O, Dd, B, p.Tc, p.D, p.heads);
  CUtensorMap q128, do128, k128, v128, q64, do64, k64, v64, qmn, domn, kmn;
  if (!tmap_kmajor(&q128, p.Q, p, B, 128) || !tmap_kmajor(&do128, p.dO, p, B, 128) || !tmap_kmajor(&k128, p.K, p, B, 128) ||
      !tmap_kmajor(&v128, p.V, p, B, 128) || !tmap_kmajor(&q64, p.Q, p, B, 64) || !tmap_kmajor(&do64, p.dO, p, B, 64) ||
      !tmap_kmajor(&k64, p.K, p, B, 64) || !tmap_kmajor(&v64, p.V, p, B, 64) || !tmap_mnmajor(&qmn, p.Q, p, B) ||
      !tmap_mnmajor(&domn, p.dO, p, B) || !tmap_mnmajor(&kmn, p.K, p, B))
    return cudaErrorInvalidValue;
  const dim3 grid(cdiv(p.Tc, 128), B * p.heads);
  static const bool serial = getenv("CVT_ATTN_BWD_SERIAL") != nullptr;
  const bool fork = side && ev_fork && ev_join && !serial;
  if (fork) { CVT_CUDA(cudaEventRecord(ev_fork, s)); CVT_CUDA(cudaStreamWaitEvent(side, ev_fork, 0)); }
  // 128-wide score tiles as soon as the 64-wide kernels would need two tiles (CVT_ATTN_BWD_TILE=64 / 128 force one family; the trace hooks live in the 64-wide dQ kernel)
  static const int force = getenv("CVT_ATTN_BWD_TILE") ? atoi(getenv("CVT_ATTN_BWD_TILE")) : 0;
  const bool wide = force == 128 || (force != 64 && p.Tc > 64 && !p.trace);
  const bool wide_dkv = wide;
  // the wide kernels are persistent (one CTA per SM walking 128-row work items, w = blockIdx.x, + gridDim.x, ...).  Side by side still pays: the
  // CTAs of the first grid that drew one item less exit early, and the second grid's lowest-numbered CTAs (one item more) start there
  const int nblk = cdiv(p.Tc, 128), nitems = nblk * B * p.heads, pgrid = nitems < kNumSMs ? nitems : kNumSMs;
  if (wide_dkv) { CVT_LAUNCH((attn_bwd_dkv128_kernel), pgrid, BWD_THREADS, K2_SMEM, s, p, k128, v128, q128, do128, qmn, domn, nblk, nitems); }
  else { CVT_LAUNCH((attn_bwd_dkv_kernel), grid, BWD_THREADS, BK_SMEM, s, p, k128, v128, q64, do64, qmn, domn); }
  if (wide) { CVT_LAUNCH((attn_bwd_dq128_kernel), pgrid, BWD_THREADS, Q2_SMEM, fork ? side : s, p, q128, do128, k128, v128, kmn, nblk, nitems); }
  else { CVT_LAUNCH((attn_bwd_dq_kernel), grid, BWD_THREADS, BQ_SMEM, fork ? side : s, p, q128, do128, k64, v64, kmn); }
  if (fork) { CVT_CUDA(cudaEventRecord(ev_join, side)); CVT_CUDA(cudaStreamWaitEvent(s, ev_join, 0)); }
  return cudaSuccess;
}
}  // namespace

bool attention_tc_enabled(const Model* m, int) { return m->cfg.precision != CVT_PREC_FP32; }

cudaError_t attention_forward_tc(Model* m, int si, int B, cudaStream_t s) {
  const StageGeo& g = m->geo[si]; StageBuf& a = m->buf[si];
  const long long RD = (long long)B * g.Tc * g.D;
  AttnP p{};
  p.Q = a.qkv2; p.K = a.qkv2 + RD; p.V = a.qkv2 + 2 * RD; p.O = a.attn_o; p.LSE = a.lse;
  p.Tc = g.Tc; p.D = g.D; p.heads = g.heads; p.dq_scale = 0.125f; p.round_out = m->folded && !m->x3;      // (TF32X3: O / dQ / dK / dV feed error-compensated GEMMs, keep them in full fp32)
  return run_fwd(p, B, s);
}

cudaError_t attention_backward_tc(Model* m, int si, int B, const float* dO, float* dQKV, cudaStream_t s) {
  const StageGeo& g = m->geo[si]; StageBuf& a = m->buf[si];
  const long long RD = (long long)B * g.Tc * g.D;
  AttnP p{};
  p.Q = a.qkv2; p.K = a.qkv2 + RD; p.V = a.qkv2 + 2 * RD; p.O = a.attn_o; p.LSE = a.lse;
  p.dO = dO; p.Dd = m->attn_dd; p.dQ = dQKV; p.dK = dQKV + RD; p.dV = dQKV + 2 * RD;
  p.Tc = g.Tc; p.D = g.D; p.heads = g.heads; p.dq_scale = 0.125f; p.round_out = m->folded && !m->x3;      // (TF32X3: O / dQ / dK / dV feed error-compensated GEMMs, keep them in full fp32)
  return run_bwd(p, m->attn_dd, B, s, m->side_stream, m->ev_fork, m->ev_join);
}

}  // namespace cvt

// diagnostics (tests/test_gpu_attention.py): the tcgen05 attention kernels on host arrays, device 0.
// Q (pre-scaled), K, V, dO: [B, Tc, heads*64]; O/dQ/dK/dV same shape; LSE [B*heads, Tc].  dO == NULL: forward only.
// Inputs are rounded to TF32 on the device first (in the model their producers do that in the GEMM epilogue).
extern "C" CVT_API int cvt_debug_attention(int B, int Tc, int heads, const float* Q, const float* K, const float* V, const float* dO,
                                           float* O, float* LSE, float* dQ, float* dK, float* dV) {
  using namespace cvt;
  if (B < 1 || Tc < 1 || heads < 1 || !Q || !K || !V || !O || !LSE) { set_error("cvt_debug_attention: bad argument"); return CVT_ERR_INVALID; }
  const int D = heads * 64;
  const size_t n = (size_t)B * Tc * D, nl = (size_t)B * heads * Tc;
  float* buf = nullptr;
  cudaError_t e = cudaMalloc(&buf, (9 * n + 2 * nl) * sizeof(float));
  if (e != cudaSuccess) { set_error(std::string("cvt_debug_attention: ") + cudaGetErrorString(e)); return CVT_ERR_CUDA; }
  float *dQ_ = buf, *dK_ = buf + n, *dV_ = buf + 2 * n, *dO_ = buf + 3 * n, *dOut = buf + 4 * n, *gQ = buf + 5 * n, *gK = buf + 6 * n, *gV = buf + 7 * n;
  float *dL = buf + 9 * n, *dD = dL + nl;
  auto fail = [&](cudaError_t err) { if (err != cudaSuccess) set_error(std::string("cvt_debug_attention: ") + cudaGetErrorString(err)); cudaFree(buf); return (int)CVT_ERR_CUDA; };
  if ((e = cudaMemcpy(dQ_, Q, n * 4, cudaMemcpyHostToDevice)) != cudaSuccess || (e = cudaMemcpy(dK_, K, n * 4, cudaMemcpyHostToDevice)) != cudaSuccess ||
      (e = cudaMemcpy(dV_, V, n * 4, cudaMemcpyHostToDevice)) != cudaSuccess) return fail(e);
  if (dO && (e = cudaMemcpy(dO_, dO, n * 4, cudaMemcpyHostToDevice)) != cudaSuccess) return fail(e);
  round_tf32_inplace_kernel<<<592, 256>>>(buf, (long long)(dO ? 4 : 3) * n);
  if ((e = cudaGetLastError()) != cudaSuccess) return fail(e);
  AttnP p{};
  p.Q = dQ_; p.K = dK_; p.V = dV_; p.O = dOut; p.LSE = dL; p.Tc = Tc; p.D = D; p.heads = heads; p.dq_scale = 0.125f;
  if ((e = run_fwd(p, B, 0)) != cudaSuccess) return fail(e);
  if (dO) {
    if (!dQ || !dK || !dV) { cudaFree(buf); set_error("cvt_debug_attention: backward needs dQ, dK, dV"); return CVT_ERR_INVALID; }
    p.dO = dO_; p.Dd = dD; p.dQ = gQ; p.dK = gK; p.dV = gV;
    long long* trace = nullptr;
    if (getenv("CVT_ATTN_TRACE")) { cudaMalloc(&trace, 8 * 256 * sizeof(long long)); cudaMemset(trace, 0, 8 * 256 * sizeof(long long)); p.trace = trace; }
    if ((e = run_bwd(p, dD, B, 0)) != cudaSuccess) return fail(e);
    if (trace) {
      static long long h[8 * 256];
      cudaMemcpy(h, trace, sizeof(h), cudaMemcpyDeviceToHost); cudaFree(trace);
      const int nt = (Tc + 63) / 64; const long long t0 = h[0];
      printf("dq trace (cycles since first kv_full): tile | kv_full  ds_wait ds_full kmn_full | row: s_wait s_full ld_done arrive\n");
      for (int j = 0; j < nt && j < 256; ++j) { printf("%3d |", j); for (int k = 0; k < 8; ++k) printf(" %7lld", h[j * 8 + k] ? h[j * 8 + k] - t0 : -1); printf("\n"); }
    }
  }
  if ((e = cudaDeviceSynchronize()) != cudaSuccess) return fail(e);
  if ((e = cudaMemcpy(O, dOut, n * 4, cudaMemcpyDeviceToHost)) != cudaSuccess || (e = cudaMemcpy(LSE, dL, nl * 4, cudaMemcpyDeviceToHost)) != cudaSuccess) return fail(e);
  if (dO && ((e = cudaMemcpy(dQ, gQ, n * 4, cudaMemcpyDeviceToHost)) != cudaSuccess || (e = cudaMemcpy(dK, gK, n * 4, cudaMemcpyDeviceToHost)) != cudaSuccess ||
             (e = cudaMemcpy(dV, gV, n * 4, cudaMemcpyDeviceToHost)) != cudaSuccess)) return fail(e);
  cudaFree(buf);
  return CVT_OK;
}

// ---- allocation-free variant -----------------------------------------------------------------------------------------------------------
// cvt_attention_backward_static: the flash-attention backward (attn_dd_kernel + attn_bwd_dkv_kernel + attn_bwd_dq_kernel, the same
// kernels run_bwd launches for the model) on caller-owned DEVICE buffers, for callers that cannot call cudaMalloc (CUDA graph capture,
// tight serving loops, harnesses that forbid it).  The one scratch buffer these kernels need - Dd[B*heads, Tc] = rowsum(dO * O) - lives
// in a file-scope __device__ array of CVT_STATIC_ATTN_MAX_ROWS floats that every GPU loading the library holds until exit.
// No counterpart in the reference.
// Every call shares the array: calls that overlap on the GPU (other streams or host threads of this process on the same
// device) race on it.  Nothing in it has to start at zero: attn_dd_kernel rewrites every element the two gradient kernels read.
namespace cvt {
namespace {
static __device__ __align__(256) float g_static_attn_dd[CVT_STATIC_ATTN_MAX_ROWS];

// In host code the array's name is only a host-side stand-in for the symbol: its device address comes from cudaGetSymbolAddress, looked
// up once per device with that device current.
cudaError_t static_attn_scratch(float** out) {
  constexpr int kMaxDevices = 64;
  static float* ptrs[kMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  CVT_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) { set_error("cvt_attention_backward_static: device ordinal out of range"); return cudaErrorInvalidDevice; }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    float* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_static_attn_dd);
    if (e != cudaSuccess) { set_error(std::string("cvt_attention_backward_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e)); return e; }
    if (!ptr) { set_error("cvt_attention_backward_static: cudaGetSymbolAddress returned a null address"); return cudaErrorInvalidSymbol; }
    ptrs[dev] = ptr;
  }
  *out = ptrs[dev];
  return cudaSuccess;
}
}  // namespace
}  // namespace cvt

extern "C" CVT_API int cvt_attention_backward_static(int B, int Tc, int heads, const float* Q_dev, const float* K_dev, const float* V_dev,
                                                     const float* O_dev, const float* LSE_dev, const float* dO_dev, float* dQ_dev,
                                                     float* dK_dev, float* dV_dev, void* stream) {
  using namespace cvt;
  if (B < 1 || Tc < 1 || heads < 1 || !Q_dev || !K_dev || !V_dev || !O_dev || !LSE_dev || !dO_dev || !dQ_dev || !dK_dev || !dV_dev) {
    set_error("cvt_attention_backward_static: bad argument");
    return CVT_ERR_INVALID;
  }
  const long long rows = (long long)B * heads * Tc;
  if (rows > (long long)CVT_STATIC_ATTN_MAX_ROWS) {
    set_error("cvt_attention_backward_static: batch * heads * tokens = " + std::to_string(rows) + " exceeds the static scratch of " +
              std::to_string((long long)CVT_STATIC_ATTN_MAX_ROWS) + " rows (CVT_STATIC_ATTN_MAX_ROWS)");
    return CVT_ERR_INVALID;
  }
  float* dd = nullptr;
  if (static_attn_scratch(&dd) != cudaSuccess) return CVT_ERR_CUDA;
  AttnP p{};
  p.Q = Q_dev; p.K = K_dev; p.V = V_dev; p.O = const_cast<float*>(O_dev); p.LSE = const_cast<float*>(LSE_dev);      // read-only in the backward kernels
  p.dO = dO_dev; p.Dd = dd; p.dQ = dQ_dev; p.dK = dK_dev; p.dV = dV_dev;
  p.Tc = Tc; p.D = heads * 64; p.heads = heads; p.dq_scale = 0.125f; p.round_out = 0;
  const cudaError_t e = run_bwd(p, dd, B, static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? CVT_OK : CVT_ERR_CUDA;      // run_bwd has recorded the message
}
