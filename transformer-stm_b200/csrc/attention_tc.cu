// Flash-style attention on tcgen05 / TMEM (kind::tf32, fp32 accumulate) for keras.layers.MultiHeadAttention as called at
// models/CvT(Par).py:185:  S = Q K^T (Q already carries 1/sqrt(key_dim)), P = softmax(S), O = P V; key_dim = 64 in every spec.
// The [B, heads, Tc, Tc] score tensor never reaches HBM: forward keeps the row log-sum-exp, backward recomputes P from it.
//
//   forward   CTA = 128 query rows of one (image, head); thread t owns row t (TMEM lane t).  Per 64-key tile:
//             K/V tile -> swizzled smem, S = Q K^T (8 UMMA 128x64x8) -> TMEM, tcgen05.ld, online softmax in registers,
//             P -> smem (K-major), P V -> TMEM, accumulated with the running rescale in registers.  2 CTAs per SM.
//   backward  dQ kernel (CTA = 128 query rows, loops over key tiles) and dK/dV kernel (CTA = 128 key rows, loops over query
//             tiles, works on S^T so that P^T / dS^T are produced row-per-thread as K-major A operands).  256 threads:
//             warps w and w+4 share a TMEM lane quadrant and split the 64 tile columns.  No atomics.
//   layouts   Q/K/V/O/dO are [B, Tc, D] with head h in columns [64h, 64h+64); operands that are contiguous along the
//             contraction index are staged K-major (SWIZZLE_128B), the others MN-major (SWIZZLE_128B_BASE32B).
#include "model.cuh"
#include "tc_common.cuh"

namespace cvt {
using namespace tc;

namespace {
constexpr int HD = 64;
constexpr float kLog2e = 1.4426950408889634f;

struct AttnP {
  const float* Q; const float* K; const float* V;      // [B, Tc, D]
  float* O; float* LSE;                                // O [B, Tc, D]; LSE [B*heads, Tc]
  const float* dO; const float* Dd;                    // backward: dO [B, Tc, D]; Dd [B*heads, Tc] = rowsum(dO * O)
  float* dQ; float* dK; float* dV;
  int Tc, D, heads;
  float dq_scale;
  int round_out;                                       // store O / dQ / dK / dV rounded to TF32 (they only feed tensor-core GEMMs)
};

// ---- tile movers: global fp32 rows -> tf32 -> swizzled shared memory -------------------------------------------------
// K-major tile of `rows` x 64 floats: two 32-float atoms along the contraction axis, atom a at a*rows*128 bytes.
template <int ROWS, int NT>
__device__ __forceinline__ void load_kmajor(uint8_t* dst, const float* src, long long ld, int row0, int nrows_valid, int tid) {
  // ROWS x 16 float4
#pragma unroll
  for (int i = 0; i < ROWS * 16 / NT; ++i) {
    const int idx = tid + NT * i;
    const int r = idx >> 4, c4 = idx & 15;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < nrows_valid) v = *reinterpret_cast<const float4*>(src + (long long)(row0 + r) * ld + c4 * 4);
    v.x = to_tf32(v.x); v.y = to_tf32(v.y); v.z = to_tf32(v.z); v.w = to_tf32(v.w);
    *reinterpret_cast<float4*>(dst + (c4 >> 3) * (ROWS * 128) + sw128_kmajor_off(r, c4 & 7)) = v;
  }
}
// MN-major tile: contraction index = row (ROWS of them), MN = the 64 features (2 atoms of 32 floats)
template <int ROWS, int NT>
__device__ __forceinline__ void load_mnmajor(uint8_t* dst, const float* src, long long ld, int row0, int nrows_valid, int tid) {
#pragma unroll
  for (int i = 0; i < ROWS * 16 / NT; ++i) {
    const int idx = tid + NT * i;
    const int r = idx >> 4, c4 = idx & 15;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < nrows_valid) v = *reinterpret_cast<const float4*>(src + (long long)(row0 + r) * ld + c4 * 4);
    v.x = to_tf32(v.x); v.y = to_tf32(v.y); v.z = to_tf32(v.z); v.w = to_tf32(v.w);
    *reinterpret_cast<float4*>(dst + sw128_mnmajor_off(r, c4, 2)) = v;
  }
}
// both layouts from one global read
template <int ROWS, int NT>
__device__ __forceinline__ void load_both(uint8_t* dst_k, uint8_t* dst_mn, const float* src, long long ld, int row0, int nrows_valid, int tid) {
#pragma unroll
  for (int i = 0; i < ROWS * 16 / NT; ++i) {
    const int idx = tid + NT * i;
    const int r = idx >> 4, c4 = idx & 15;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < nrows_valid) v = *reinterpret_cast<const float4*>(src + (long long)(row0 + r) * ld + c4 * 4);
    v.x = to_tf32(v.x); v.y = to_tf32(v.y); v.z = to_tf32(v.z); v.w = to_tf32(v.w);
    *reinterpret_cast<float4*>(dst_k + (c4 >> 3) * (ROWS * 128) + sw128_kmajor_off(r, c4 & 7)) = v;
    *reinterpret_cast<float4*>(dst_mn + sw128_mnmajor_off(r, c4, 2)) = v;
  }
}

// D[tmem] (+)= A * B with A K-major [128 x 64] (atoms of A_ROWS rows) and B K-major [64 rows x 64] (contraction = 64 floats)
template <int A_ROWS, int B_ROWS>
__device__ __forceinline__ void mma_kk(uint32_t tmem_d, uint32_t a0, uint32_t b0, uint32_t idesc, bool accumulate) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint64_t ad = make_smem_desc(a0 + (j >> 2) * (A_ROWS * 128) + (j & 3) * 32, 16, 1024);
    const uint64_t bd = make_smem_desc(b0 + (j >> 2) * (B_ROWS * 128) + (j & 3) * 32, 16, 1024);
    mma_tf32_ss(tmem_d, ad, bd, idesc, (accumulate || j > 0) ? 1u : 0u);
  }
}
// A K-major [128 x 64] and B MN-major [64 contraction rows x 64 features]
template <int A_ROWS>
__device__ __forceinline__ void mma_kmn(uint32_t tmem_d, uint32_t a0, uint32_t b0, uint32_t idesc, bool accumulate) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint64_t ad = make_smem_desc(a0 + (j >> 2) * (A_ROWS * 128) + (j & 3) * 32, 16, 1024);
    const uint64_t bd = make_smem_desc(b0 + j * 2048, 512, 1024, kLayoutSw128Base32);
    mma_tf32_ss(tmem_d, ad, bd, idesc, (accumulate || j > 0) ? 1u : 0u);
  }
}

__device__ __forceinline__ uint8_t* align1024(uint8_t* p) {
  const uint32_t a = smem_u32(p);
  return p + (((a + 1023u) & ~1023u) - a);
}

// one row of 32 tf32 values -> K-major swizzled tile (row-per-thread writers), atom `atom` of a tile with ROWS rows
template <int ROWS>
__device__ __forceinline__ void store_row32(uint8_t* tile, int row, int atom, const float v[32]) {
  uint8_t* base = tile + atom * (ROWS * 128);
#pragma unroll
  for (int c = 0; c < 8; ++c)
    *reinterpret_cast<float4*>(base + sw128_kmajor_off(row, c)) =
        make_float4(to_tf32(v[4 * c]), to_tf32(v[4 * c + 1]), to_tf32(v[4 * c + 2]), to_tf32(v[4 * c + 3]));
}

// =====================================================================================================================
// forward
// =====================================================================================================================
constexpr int FWD_SMEM = 1024 + 32768 + 16384 + 16384 + 32768 + 64;

__global__ void __launch_bounds__(128, 2) attn_fwd_kernel(const __grid_constant__ AttnP p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = align1024(smem_raw);
  uint8_t* sQ = sm; uint8_t* sK = sQ + 32768; uint8_t* sV = sK + 16384; uint8_t* sP = sV + 16384;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 32768);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);

  const int tid = threadIdx.x, warp = tid >> 5;
  const int bh = blockIdx.y, b = bh / p.heads, h = bh - b * p.heads;
  const int q0 = blockIdx.x * 128;
  const long long base = (long long)b * p.Tc * p.D + h * HD;
  const float* Q = p.Q + base; const float* K = p.K + base; const float* V = p.V + base;

  if (warp == 0) tmem_alloc(tmem_slot, 128);
  if (tid == 32) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_fence_init(); }
  load_kmajor<128, 128>(sQ, Q, p.D, q0, p.Tc, tid);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t t_s = tmem + ((uint32_t)(warp * 32) << 16), t_o = t_s + 64;
  constexpr uint32_t idesc_s = make_idesc(kFmtTF32, 128, 64, 0, 0);
  constexpr uint32_t idesc_o = make_idesc(kFmtTF32, 128, 64, 0, 1);

  float o[64];
#pragma unroll
  for (int i = 0; i < 64; ++i) o[i] = 0.f;
  float m_run = -INFINITY, l_run = 0.f;
  const int ntiles = cdiv(p.Tc, 64);
  for (int j = 0; j < ntiles; ++j) {
    const int kv0 = j * 64;
    load_kmajor<64, 128>(sK, K, p.D, kv0, p.Tc, tid);
    load_mnmajor<64, 128>(sV, V, p.D, kv0, p.Tc, tid);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      mma_kk<128, 64>(tmem, smem_u32(sQ), smem_u32(sK), idesc_s, false);
      mma_commit(&bars[0]);
    }
    mbar_wait(&bars[0], j & 1);
    tc_fence_after();
    float s[64];
    tmem_ld32(t_s, s);
    tmem_ld32(t_s + 32, s + 32);
    float mx = m_run;
    const int nvalid = p.Tc - kv0;                    // keys beyond Tc are masked out
#pragma unroll
    for (int c = 0; c < 64; ++c) { if (c >= nvalid) s[c] = -INFINITY; mx = fmaxf(mx, s[c]); }
    const float alpha = exp2f((m_run - mx) * kLog2e);
    const float mb = mx * kLog2e;
    float sum = 0.f;
#pragma unroll
    for (int c = 0; c < 64; ++c) { s[c] = exp2f(fmaf(s[c], kLog2e, -mb)); sum += s[c]; }
    l_run = l_run * alpha + sum;
    m_run = mx;
    store_row32<128>(sP, tid, 0, s);
    store_row32<128>(sP, tid, 1, s + 32);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      mma_kmn<128>(tmem + 64, smem_u32(sP), smem_u32(sV), idesc_o, false);
      mma_commit(&bars[1]);
    }
    mbar_wait(&bars[1], j & 1);
    tc_fence_after();
    float pv[32];
    tmem_ld32(t_o, pv);
#pragma unroll
    for (int c = 0; c < 32; ++c) o[c] = fmaf(o[c], alpha, pv[c]);
    tmem_ld32(t_o + 32, pv);
#pragma unroll
    for (int c = 0; c < 32; ++c) o[32 + c] = fmaf(o[32 + c], alpha, pv[c]);
  }
  const int q = q0 + tid;
  if (q < p.Tc) {
    const float inv = 1.f / l_run;
    float* dst = p.O + base + (long long)q * p.D;
#pragma unroll
    for (int c = 0; c < 16; ++c)
    {
      float4 t = make_float4(o[4 * c] * inv, o[4 * c + 1] * inv, o[4 * c + 2] * inv, o[4 * c + 3] * inv);
      if (p.round_out) t = make_float4(to_tf32(t.x), to_tf32(t.y), to_tf32(t.z), to_tf32(t.w));
      *reinterpret_cast<float4*>(dst + 4 * c) = t;
    }
    if (p.LSE) p.LSE[(long long)bh * p.Tc + q] = m_run + logf(l_run);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

// =====================================================================================================================
// backward
// =====================================================================================================================
template <int ROWS, int NT> struct TileRegs { float4 v[ROWS * 16 / NT]; };

template <int ROWS, int NT>
__device__ __forceinline__ void ldg_tile(TileRegs<ROWS, NT>& t, const float* src, long long ld, int row0, int nrows_valid, int tid) {
#pragma unroll
  for (int i = 0; i < ROWS * 16 / NT; ++i) {
    const int idx = tid + NT * i;
    const int r = idx >> 4, c4 = idx & 15;
    t.v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < nrows_valid) t.v[i] = *reinterpret_cast<const float4*>(src + (long long)(row0 + r) * ld + c4 * 4);
  }
}
template <int ROWS, int NT, bool KMAJ, bool MNMAJ>
__device__ __forceinline__ void sts_tile(const TileRegs<ROWS, NT>& t, uint8_t* dst_k, uint8_t* dst_mn, int tid) {
#pragma unroll
  for (int i = 0; i < ROWS * 16 / NT; ++i) {
    const int idx = tid + NT * i;
    const int r = idx >> 4, c4 = idx & 15;
    const float4 v = make_float4(to_tf32(t.v[i].x), to_tf32(t.v[i].y), to_tf32(t.v[i].z), to_tf32(t.v[i].w));
    if (KMAJ) *reinterpret_cast<float4*>(dst_k + (c4 >> 3) * (ROWS * 128) + sw128_kmajor_off(r, c4 & 7)) = v;
    if (MNMAJ) *reinterpret_cast<float4*>(dst_mn + sw128_mnmajor_off(r, c4, 2)) = v;
  }
}

// Dd[bh, t] = sum_d dO[b,t,h,d] * O[b,t,h,d]   (half a warp per (token, head))
__global__ void __launch_bounds__(256) attn_dd_kernel(const float* __restrict__ dO, const float* __restrict__ O, float* __restrict__ Dd,
                                                      int B, int Tc, int D, int heads) {
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

// ---- dQ: CTA = 128 query rows, loop over 64-key tiles -----------------------------------------------------------------
constexpr int BQ_SMEM = 1024 + 32768 + 32768 + 16384 + 16384 + 16384 + 32768 + 64;

__global__ void __launch_bounds__(256, 1) attn_bwd_dq_kernel(const __grid_constant__ AttnP p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = align1024(smem_raw);
  uint8_t* sQ = sm; uint8_t* sdO = sQ + 32768; uint8_t* sK = sdO + 32768; uint8_t* sKmn = sK + 16384;
  uint8_t* sV = sKmn + 16384; uint8_t* sdS = sV + 16384;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sdS + 32768);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int quad = warp & 3, half = warp >> 2;
  const int r = quad * 32 + lane;
  const int bh = blockIdx.y, b = bh / p.heads, h = bh - b * p.heads;
  const int q0 = blockIdx.x * 128;
  const long long base = (long long)b * p.Tc * p.D + h * HD;
  const float* Q = p.Q + base; const float* K = p.K + base; const float* V = p.V + base; const float* dO = p.dO + base;

  if (warp == 0) tmem_alloc(tmem_slot, 256);
  if (tid == 32) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_fence_init(); }
  load_kmajor<128, 256>(sQ, Q, p.D, q0, p.Tc, tid);
  load_kmajor<128, 256>(sdO, dO, p.D, q0, p.Tc, tid);
  const int q = q0 + r;
  const float lse2 = q < p.Tc ? p.LSE[(long long)bh * p.Tc + q] * kLog2e : 0.f;
  const float dd = q < p.Tc ? p.Dd[(long long)bh * p.Tc + q] : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t t_row = tmem + ((uint32_t)(quad * 32) << 16) + half * 32;
  constexpr uint32_t idesc_kk = make_idesc(kFmtTF32, 128, 64, 0, 0);
  constexpr uint32_t idesc_kmn = make_idesc(kFmtTF32, 128, 64, 0, 1);

  const int ntiles = cdiv(p.Tc, 64);
  for (int j = 0; j < ntiles; ++j) {
    const int kv0 = j * 64;
    TileRegs<64, 256> rk, rv;
    ldg_tile<64, 256>(rk, K, p.D, kv0, p.Tc, tid);
    ldg_tile<64, 256>(rv, V, p.D, kv0, p.Tc, tid);
    if (j > 0) { mbar_wait(&bars[1], (j - 1) & 1); tc_fence_after(); }     // the dQ MMAs that read sKmn / sdS have completed
    sts_tile<64, 256, true, true>(rk, sK, sKmn, tid);
    sts_tile<64, 256, true, false>(rv, sV, nullptr, tid);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      mma_kk<128, 64>(tmem, smem_u32(sQ), smem_u32(sK), idesc_kk, false);           // S  = Q K^T
      mma_kk<128, 64>(tmem + 64, smem_u32(sdO), smem_u32(sV), idesc_kk, false);     // dP = dO V^T
      mma_commit(&bars[0]);
    }
    mbar_wait(&bars[0], j & 1);
    tc_fence_after();
    float s[32], dp[32];
    tmem_ld32(t_row, s);
    tmem_ld32(t_row + 64, dp);
#pragma unroll
    for (int c = 0; c < 32; ++c) s[c] = exp2f(fmaf(s[c], kLog2e, -lse2)) * (dp[c] - dd);     // dS = P * (dP - D)
    store_row32<128>(sdS, r, half, s);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      mma_kmn<128>(tmem + 128, smem_u32(sdS), smem_u32(sKmn), idesc_kmn, j > 0);      // dQ += dS K
      mma_commit(&bars[1]);
    }
  }
  mbar_wait(&bars[1], (ntiles - 1) & 1);
  tc_fence_after();
  float acc[32];
  tmem_ld32(t_row + 128, acc);
  if (q < p.Tc) {
    float* dst = p.dQ + base + (long long)q * p.D + half * 32;
#pragma unroll
    for (int c = 0; c < 8; ++c)
    {
      float4 t = make_float4(acc[4 * c] * p.dq_scale, acc[4 * c + 1] * p.dq_scale, acc[4 * c + 2] * p.dq_scale, acc[4 * c + 3] * p.dq_scale);
      if (p.round_out) t = make_float4(to_tf32(t.x), to_tf32(t.y), to_tf32(t.z), to_tf32(t.w));
      *reinterpret_cast<float4*>(dst + 4 * c) = t;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// ---- dK, dV: CTA = 128 key rows, loop over 64-query tiles; works on S^T ------------------------------------------------
constexpr int BK_SMEM = 1024 + 2 * 32768 + 4 * 16384 + 2 * 32768 + 512 + 64;

__global__ void __launch_bounds__(256, 1) attn_bwd_dkv_kernel(const __grid_constant__ AttnP p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = align1024(smem_raw);
  uint8_t* sK = sm; uint8_t* sV = sK + 32768; uint8_t* sQ = sV + 32768; uint8_t* sQmn = sQ + 16384;
  uint8_t* sdO = sQmn + 16384; uint8_t* sdOmn = sdO + 16384; uint8_t* sPt = sdOmn + 16384; uint8_t* sdSt = sPt + 32768;
  float* s_lse = reinterpret_cast<float*>(sdSt + 32768); float* s_dd = s_lse + 64;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_dd + 64);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int quad = warp & 3, half = warp >> 2;
  const int r = quad * 32 + lane;
  const int bh = blockIdx.y, b = bh / p.heads, h = bh - b * p.heads;
  const int kv0 = blockIdx.x * 128;
  const long long base = (long long)b * p.Tc * p.D + h * HD;
  const float* Q = p.Q + base; const float* K = p.K + base; const float* V = p.V + base; const float* dO = p.dO + base;
  const float* LSE = p.LSE + (long long)bh * p.Tc; const float* Dd = p.Dd + (long long)bh * p.Tc;

  if (warp == 0) tmem_alloc(tmem_slot, 256);
  if (tid == 32) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_fence_init(); }
  load_kmajor<128, 256>(sK, K, p.D, kv0, p.Tc, tid);
  load_kmajor<128, 256>(sV, V, p.D, kv0, p.Tc, tid);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t t_row = tmem + ((uint32_t)(quad * 32) << 16) + half * 32;
  constexpr uint32_t idesc_kk = make_idesc(kFmtTF32, 128, 64, 0, 0);
  constexpr uint32_t idesc_kmn = make_idesc(kFmtTF32, 128, 64, 0, 1);

  const int ntiles = cdiv(p.Tc, 64);
  for (int i = 0; i < ntiles; ++i) {
    const int q0 = i * 64;
    TileRegs<64, 256> rq, rdo;
    ldg_tile<64, 256>(rq, Q, p.D, q0, p.Tc, tid);
    ldg_tile<64, 256>(rdo, dO, p.D, q0, p.Tc, tid);
    float aux = 0.f;
    if (tid < 128) { const int qq = q0 + (tid & 63); if (qq < p.Tc) aux = tid < 64 ? LSE[qq] * kLog2e : Dd[qq]; }
    if (i > 0) { mbar_wait(&bars[1], (i - 1) & 1); tc_fence_after(); }     // dV / dK MMAs of the previous tile have completed
    sts_tile<64, 256, true, true>(rq, sQ, sQmn, tid);
    sts_tile<64, 256, true, true>(rdo, sdO, sdOmn, tid);
    if (tid < 128) s_lse[tid] = aux;                                       // s_lse[0..63], s_dd = s_lse + 64
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      mma_kk<128, 64>(tmem, smem_u32(sK), smem_u32(sQ), idesc_kk, false);           // S^T  = K Q^T
      mma_kk<128, 64>(tmem + 64, smem_u32(sV), smem_u32(sdO), idesc_kk, false);     // dP^T = V dO^T
      mma_commit(&bars[0]);
    }
    mbar_wait(&bars[0], i & 1);
    tc_fence_after();
    float s[32], dp[32];
    tmem_ld32(t_row, s);
    tmem_ld32(t_row + 64, dp);
#pragma unroll
    for (int c = 0; c < 32; ++c) {
      const float pr = exp2f(fmaf(s[c], kLog2e, -s_lse[half * 32 + c]));
      dp[c] = pr * (dp[c] - s_dd[half * 32 + c]);
      s[c] = pr;
    }
    store_row32<128>(sPt, r, half, s);
    store_row32<128>(sdSt, r, half, dp);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      mma_kmn<128>(tmem + 128, smem_u32(sPt), smem_u32(sdOmn), idesc_kmn, i > 0);     // dV += P^T dO
      mma_kmn<128>(tmem + 192, smem_u32(sdSt), smem_u32(sQmn), idesc_kmn, i > 0);     // dK += dS^T Q
      mma_commit(&bars[1]);
    }
  }
  mbar_wait(&bars[1], (ntiles - 1) & 1);
  tc_fence_after();
  float acc[32];
  const int kv = kv0 + r;
  tmem_ld32(t_row + 128, acc);
  if (kv < p.Tc) {
    float* dst = p.dV + base + (long long)kv * p.D + half * 32;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      float4 t = make_float4(acc[4 * c], acc[4 * c + 1], acc[4 * c + 2], acc[4 * c + 3]);
      if (p.round_out) t = make_float4(to_tf32(t.x), to_tf32(t.y), to_tf32(t.z), to_tf32(t.w));
      *reinterpret_cast<float4*>(dst + 4 * c) = t;
    }
  }
  tmem_ld32(t_row + 192, acc);
  if (kv < p.Tc) {
    float* dst = p.dK + base + (long long)kv * p.D + half * 32;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      float4 t = make_float4(acc[4 * c], acc[4 * c + 1], acc[4 * c + 2], acc[4 * c + 3]);
      if (p.round_out) t = make_float4(to_tf32(t.x), to_tf32(t.y), to_tf32(t.z), to_tf32(t.w));
      *reinterpret_cast<float4*>(dst + 4 * c) = t;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

cudaError_t configure_once() {
  static bool done = false;
  if (done) return cudaSuccess;
  CVT_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM));
  CVT_CUDA(cudaFuncSetAttribute(attn_bwd_dq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BQ_SMEM));
  CVT_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BK_SMEM));
  done = true;
  return cudaSuccess;
}

cudaError_t run_fwd(const AttnP& p, int B, cudaStream_t s) {
  CVT_CUDA(configure_once());
  if (p.D % 4 != 0 || p.D != p.heads * HD) { set_error("attention_tc: D must equal heads * 64"); return cudaErrorInvalidValue; }
  attn_fwd_kernel<<<dim3(cdiv(p.Tc, 128), B * p.heads), 128, FWD_SMEM, s>>>(p);
  CVT_LAUNCH_CHECK();
  return cudaSuccess;
}
cudaError_t run_bwd(const AttnP& p, float* Dd, int B, cudaStream_t s) {
  CVT_CUDA(configure_once());
  const long long groups = (long long)B * p.Tc * p.heads;
  attn_dd_kernel<<<(unsigned)cdivl(groups * 16, 256), 256, 0, s>>>(p.dO, p.O, Dd, B, p.Tc, p.D, p.heads);
  CVT_LAUNCH_CHECK();
  attn_bwd_dkv_kernel<<<dim3(cdiv(p.Tc, 128), B * p.heads), 256, BK_SMEM, s>>>(p);
  CVT_LAUNCH_CHECK();
  attn_bwd_dq_kernel<<<dim3(cdiv(p.Tc, 128), B * p.heads), 256, BQ_SMEM, s>>>(p);
  CVT_LAUNCH_CHECK();
  return cudaSuccess;
}
}  // namespace

bool attention_tc_enabled(const Model* m, int) { return m->cfg.precision != CVT_PREC_FP32; }

cudaError_t attention_forward_tc(Model* m, int si, int B, cudaStream_t s) {
  const StageGeo& g = m->geo[si]; StageBuf& a = m->buf[si];
  const long long RD = (long long)B * g.Tc * g.D;
  AttnP p{};
  p.Q = a.qkv2; p.K = a.qkv2 + RD; p.V = a.qkv2 + 2 * RD; p.O = a.attn_o; p.LSE = a.lse;
  p.Tc = g.Tc; p.D = g.D; p.heads = g.heads; p.dq_scale = 0.125f; p.round_out = m->folded;
  return run_fwd(p, B, s);
}

cudaError_t attention_backward_tc(Model* m, int si, int B, const float* dO, float* dQKV, cudaStream_t s) {
  const StageGeo& g = m->geo[si]; StageBuf& a = m->buf[si];
  const long long RD = (long long)B * g.Tc * g.D;
  AttnP p{};
  p.Q = a.qkv2; p.K = a.qkv2 + RD; p.V = a.qkv2 + 2 * RD; p.O = a.attn_o; p.LSE = a.lse;
  p.dO = dO; p.Dd = m->attn_dd; p.dQ = dQKV; p.dK = dQKV + RD; p.dV = dQKV + 2 * RD;
  p.Tc = g.Tc; p.D = g.D; p.heads = g.heads; p.dq_scale = 0.125f; p.round_out = m->folded;
  return run_bwd(p, m->attn_dd, B, s);
}

}  // namespace cvt

// diagnostics (tests/test_gpu_attention.py): the tcgen05 attention kernels on host arrays, device 0.
// Q (pre-scaled), K, V, dO: [B, Tc, heads*64]; O/dQ/dK/dV same shape; LSE [B*heads, Tc].  dO == NULL: forward only.
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
  AttnP p{};
  p.Q = dQ_; p.K = dK_; p.V = dV_; p.O = dOut; p.LSE = dL; p.Tc = Tc; p.D = D; p.heads = heads; p.dq_scale = 0.125f;
  if ((e = run_fwd(p, B, 0)) != cudaSuccess) return fail(e);
  if (dO) {
    if (!dQ || !dK || !dV) { cudaFree(buf); set_error("cvt_debug_attention: backward needs dQ, dK, dV"); return CVT_ERR_INVALID; }
    if ((e = cudaMemcpy(dO_, dO, n * 4, cudaMemcpyHostToDevice)) != cudaSuccess) return fail(e);
    p.dO = dO_; p.Dd = dD; p.dQ = gQ; p.dK = gK; p.dV = gV;
    if ((e = run_bwd(p, dD, B, 0)) != cudaSuccess) return fail(e);
  }
  if ((e = cudaDeviceSynchronize()) != cudaSuccess) return fail(e);
  if ((e = cudaMemcpy(O, dOut, n * 4, cudaMemcpyDeviceToHost)) != cudaSuccess || (e = cudaMemcpy(LSE, dL, nl * 4, cudaMemcpyDeviceToHost)) != cudaSuccess) return fail(e);
  if (dO && ((e = cudaMemcpy(dQ, gQ, n * 4, cudaMemcpyDeviceToHost)) != cudaSuccess || (e = cudaMemcpy(dK, gK, n * 4, cudaMemcpyDeviceToHost)) != cudaSuccess ||
             (e = cudaMemcpy(dV, gV, n * 4, cudaMemcpyDeviceToHost)) != cudaSuccess)) return fail(e);
  cudaFree(buf);
  return CVT_OK;
}
