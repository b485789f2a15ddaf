// Persistent, warp-specialised tcgen05 contraction kernel fed by TMA (CVT_PREC_TF32) for the Dense / EinsumDense forward and
// dgrad products of the CvT block (models/CvT(Par).py:180-188,253-258):  C[M,N] = epilogue(A[M,K] . B[N,K]^T), both operands
// contiguous along K ("K-major"), fp32 in HBM, read by the tensor core as TF32 with fp32 accumulation in TMEM.
//
//   grid       one CTA per SM, each walks tiles (z, m-tile, n-tile) with stride gridDim.x
//   warp 0     TMA producer: 128x32 A box + BNx32 B box per k-step into a 128B-swizzled ring (mbarrier complete_tx)
//   warp 1     MMA issuer: 4 x tcgen05.mma 128xBNx8 kind::tf32 per k-step; tcgen05.commit frees the ring slot; the accumulator
//              is double-buffered in TMEM (2 x BN columns) so the epilogue of tile i overlaps the MMAs of tile i+1
//   warps 2-17 epilogue: tcgen05.ld -> per-warp smem transpose -> coalesced bias/GELU/mask/residual stores (gemm_epi.cuh)
// These GEMMs are HBM/L2-bound at the reference sizes (K, N <= 1024): the design goal is bytes in flight, not MMA rate.
#include <cuda.h>

#include <map>
#include <tuple>

#include "gemm_epi.cuh"
#include "model.cuh"
#include "tc_common.cuh"
#include "tma_host.cuh"

namespace cvt {
using namespace tc;

namespace {
constexpr int BM = 128, BK = 32, NEPI = 16, NTHREADS = 64 + 32 * NEPI;   // 16 epilogue warps: 4 TMEM lane quadrants x 4 column groups
constexpr int A_STAGE = BM * BK * 4;                 // 16 KB
constexpr int STG_BYTES = NEPI * 32 * 20 * 4;        // epilogue transpose tiles (32 rows x 16 columns, padded), one per epilogue warp

// X3: error-compensated products (CVT_PREC_TF32X3).  The tensor core truncates an fp32 word to TF32, so the word a TMA box delivers IS
// the high part; four extra warps write the low parts (x - trunc(x), exact in fp32) of both operand tiles of a stage into twin tiles and
// the issuer runs three MMAs per k-step: hi.hi + lo.hi + hi.lo (the dropped lo.lo term is 2^-22 relative).  fp32-grade results on tcgen05.
constexpr int NSPLIT = 4;                              // splitter warps of the X3 variant
template <int BN, bool X3> struct TCfg {
  static constexpr int B_STAGE = BN * BK * 4;
  static constexpr int NS = X3 ? (BN == 64 ? 3 : 2) : (BN == 64 ? 6 : (BN == 128 ? 5 : 3));      // X3: twice the bytes per stage, N tiles of at most 128
  static constexpr int STAGE_ALL = (A_STAGE + B_STAGE) * (X3 ? 2 : 1);
  static constexpr int SMEM = 1024 + NS * STAGE_ALL + STG_BYTES + (3 * NS + 4) * 8 + 16;
  static_assert(SMEM <= 232448, "gemm_tma: shared memory budget");
  static constexpr int THREADS = NTHREADS + (X3 ? 32 * NSPLIT : 32);      // single-pass kernel: one more warp issues the B operand's TMA loads
};

template <int BN, uint32_t E, bool X3>
__global__ void __launch_bounds__(TCfg<BN, X3>::THREADS, 1) gemm_tma_kernel(const __grid_constant__ GemmP p, const __grid_constant__ CUtensorMap tmA,
                                                                             const __grid_constant__ CUtensorMap tmB, int m_tiles, int n_tiles,
                                                                             int num_tiles) {
  using C = TCfg<BN, X3>;
  constexpr int NS = C::NS;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sA = smem;
  uint8_t* sB = sA + NS * A_STAGE;
  uint8_t* sAlo = sB + NS * C::B_STAGE;                // X3 only: low parts of the A / B tiles of every stage
  uint8_t* sBlo = sAlo + (X3 ? NS * A_STAGE : 0);
  float* stg_all = reinterpret_cast<float*>(sBlo + (X3 ? NS * C::B_STAGE : 0));
  uint64_t* full = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(stg_all) + STG_BYTES);
  uint64_t* empty = full + NS;
  uint64_t* ready = empty + NS;          // X3: the splitter warps are done with the stage
  uint64_t* acc_full = ready + NS;       // [2]
  uint64_t* acc_empty = acc_full + 2;    // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nk = cdiv(p.K, BK);
  // diagnostics (CVT_GEMM_TRACE, tools/gemm_trace.py): clock64 timeline of CTA 0; p.trace is null in the model
#define GEMM_TRACE(slot) do { if (p.trace && blockIdx.x == 0 && lane == 0) p.trace[(slot)] = clock64(); } while (0)
  if (threadIdx.x == 0) GEMM_TRACE(0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB);
    for (int i = 0; i < NS; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); mbar_init(&ready[i], 32 * NSPLIT); }
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], NEPI); }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 2 * BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait(); pdl_trigger();      // everything above touches no global memory
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) GEMM_TRACE(1);

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0; int itp = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++itp) {
        const int z = tile / (m_tiles * n_tiles), r = tile - z * (m_tiles * n_tiles);
        const int mt = r / n_tiles, nt = r - mt * n_tiles;
        for (int kt = 0; kt < nk; ++kt) {
          mbar_wait(&empty[stage], phase ^ 1);
          if (kt == 0 && itp < 30) GEMM_TRACE(8 + itp * 8 + 0);
          mbar_arrive_expect_tx(&full[stage], A_STAGE + C::B_STAGE);
          if (p.a_mode == A_CONV) {
            // implicit GEMM: k-step kt = (filter tap, 32-channel chunk); the 128 rows are consecutive output pixels = one strided box of the
            // input map, addressed by where the tap falls for the tile's first pixel
            const ConvGeom& g = p.cv;
            const int cpt = g.cin >> 5, tap = kt / cpt, c0 = (kt - tap * cpt) << 5;
            const int kh = tap / g.k, kw = tap - kh * g.k;
            const int m0 = mt * BM, hw = g.ho * g.wo, b0 = m0 / hw, i0 = (m0 - b0 * hw) / g.wo;
            tma_load_4d(sA + stage * A_STAGE, &tmA, &full[stage], c0, kw - g.pl, g.s * i0 + kh - g.pt, b0);
          } else
          tma_load_3d(sA + stage * A_STAGE, &tmA, &full[stage], kt * BK, mt * BM, z);
          if (X3) tma_load_3d(sB + stage * C::B_STAGE, &tmB, &full[stage], kt * BK, nt * BN, z);
          if (++stage == NS) { stage = 0; phase ^= 1; }
        }
        if (itp < 30) GEMM_TRACE(8 + itp * 8 + 1);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(kFmtTF32, BM, BN, 0u, 0u);
      int stage = 0; uint32_t phase = 0; int it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        mbar_wait(&acc_empty[buf], ((it >> 1) & 1) ^ 1);
        if (it < 30) GEMM_TRACE(8 + it * 8 + 2);
        tc_fence_after();
        const uint32_t d = tmem_base + buf * BN;
        for (int kt = 0; kt < nk; ++kt) {
          mbar_wait(X3 ? &ready[stage] : &full[stage], phase);
          if (kt == 0 && it < 30) GEMM_TRACE(8 + it * 8 + 3);
          tc_fence_after();
          const uint32_t a0 = smem_u32(sA + stage * A_STAGE), b0 = smem_u32(sB + stage * C::B_STAGE);
#pragma unroll
          for (int j = 0; j < BK / 8; ++j) {
            const uint64_t da = make_smem_desc(a0 + j * 32, 16, 1024), db = make_smem_desc(b0 + j * 32, 16, 1024);
            mma_tf32_ss(d, da, db, idesc, (kt > 0 || j > 0) ? 1u : 0u);
            if (X3) {
              mma_tf32_ss(d, make_smem_desc(smem_u32(sAlo + stage * A_STAGE) + j * 32, 16, 1024), db, idesc, 1u);
              mma_tf32_ss(d, da, make_smem_desc(smem_u32(sBlo + stage * C::B_STAGE) + j * 32, 16, 1024), idesc, 1u);
            }
          }
          mma_commit(&empty[stage]);
          if (++stage == NS) { stage = 0; phase ^= 1; }
        }
        mma_commit(&acc_full[buf]);
        if (it < 30) GEMM_TRACE(8 + it * 8 + 4);
      }
    }
  } else if (!X3 && warp == 2 + NEPI) {
    // ===== second TMA producer: the B operand =====
    // One thread issuing both operands' loads is what bounds the small-K products (K = 64: two k-steps per tile; measured ~450 cycles per
    // k-step in the single producer against a 128 x 64 tile's ~900 MMA / ~1300 epilogue cycles).  This warp waits on the same empty barrier
    // and only adds bytes to the full barrier the A producer arms: a complete_tx that lands before that expect_tx leaves the count
    // transiently negative, which the mbarrier allows; the phase cannot complete before the A producer's arrival.
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int z = tile / (m_tiles * n_tiles), r = tile - z * (m_tiles * n_tiles);
        const int nt = r % n_tiles;
        for (int kt = 0; kt < nk; ++kt) {
          mbar_wait(&empty[stage], phase ^ 1);
          tma_load_3d(sB + stage * C::B_STAGE, &tmB, &full[stage], kt * BK, nt * BN, z);
          if (++stage == NS) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (X3 && warp >= 2 + NEPI) {
    // ===== X3 splitter warps: low parts of both operand tiles of every stage, element for element (the swizzle is position-preserving) =====
    const int t = threadIdx.x - 32 * (2 + NEPI);
    int stage = 0; uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      for (int kt = 0; kt < nk; ++kt) {
        mbar_wait(&full[stage], phase);
        auto split = [&](const uint8_t* src, uint8_t* dst, int n4) {
          for (int i = t; i < n4; i += 32 * NSPLIT) {
            const float4 v = reinterpret_cast<const float4*>(src)[i];
            float4 lo;
            // the low part is itself rounded to nearest TF32 here: left to the tensor core's truncation its error would be one-sided and
            // grow linearly with K
            lo.x = to_tf32(v.x - __uint_as_float(__float_as_uint(v.x) & 0xFFFFE000u)); lo.y = to_tf32(v.y - __uint_as_float(__float_as_uint(v.y) & 0xFFFFE000u));
            lo.z = to_tf32(v.z - __uint_as_float(__float_as_uint(v.z) & 0xFFFFE000u)); lo.w = to_tf32(v.w - __uint_as_float(__float_as_uint(v.w) & 0xFFFFE000u));
            reinterpret_cast<float4*>(dst)[i] = lo;
          }
        };
        split(sA + stage * A_STAGE, sAlo + stage * A_STAGE, A_STAGE / 16);
        split(sB + stage * C::B_STAGE, sBlo + stage * C::B_STAGE, C::B_STAGE / 16);
        fence_proxy_async_smem();
        mbar_arrive(&ready[stage]);
        if (++stage == NS) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===== epilogue warps: TMEM lane quadrant = warp % 4, 16-column chunks cg, cg + 4, ... =====
    const int quad = warp & 3, cg = (warp - 2) >> 2;
    float* stg = stg_all + (warp - 2) * (32 * 20);
    const int er = lane >> 2, ec = (lane & 3) << 2;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int z = tile / (m_tiles * n_tiles), r = tile - z * (m_tiles * n_tiles);
      const int mt = r / n_tiles, nt = r - mt * n_tiles;
      const int buf = it & 1;
      float* Cp = p.C + z * p.c_bs0;
      float* aux = p.aux ? p.aux + z * p.c_bs0 : nullptr;
      const float* bias = (p.epi & EPI_BIAS) ? p.bias + z * p.bias_bs : nullptr;
      const float alpha = ((p.epi & EPI_ALPHA0) && z == 0) ? p.alpha0 : p.alpha;
      const float* coef = (E & EPI_BN_BWD) ? p.epi_coef + z * p.epi_coef_bs : nullptr;      // compile-time kind in this kernel
      const uint32_t trow = tmem_base + buf * BN + ((uint32_t)(quad * 32) << 16);
      bool waited = false;
      const bool wide = (p.ldc & 7) == 0 && (p.ldr & 7) == 0;       // 32-byte aligned rows: 8 columns per lane, 16 rows per instruction
#pragma unroll 1
      for (int c0 = cg * 16; c0 < BN; c0 += 64) {
        if (wide) {
          const int er8 = lane >> 1, ec8 = (lane & 1) << 3;
          const int n0 = nt * BN + c0 + ec8;
          EpiOps ops[2][2];
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int m = mt * BM + quad * 32 + er8 + 16 * i;
            if (m < p.M) { ops[i][0] = epilogue_load<E>(p, aux, m, n0, coef != nullptr); ops[i][1] = epilogue_load<E>(p, aux, m, n0 + 4, coef != nullptr); }
          }
          float4 b4[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
          if (bias) { b4[0] = __ldg(reinterpret_cast<const float4*>(bias + n0)); b4[1] = __ldg(reinterpret_cast<const float4*>(bias + n0 + 4)); }
          EpiCoef cf0{}, cf1{};
          if (coef) {
            cf0.a = __ldg(reinterpret_cast<const float4*>(coef + n0)); cf0.e = __ldg(reinterpret_cast<const float4*>(coef + p.N + n0)); cf0.f = __ldg(reinterpret_cast<const float4*>(coef + 2 * p.N + n0));
            cf1.a = __ldg(reinterpret_cast<const float4*>(coef + n0 + 4)); cf1.e = __ldg(reinterpret_cast<const float4*>(coef + p.N + n0 + 4)); cf1.f = __ldg(reinterpret_cast<const float4*>(coef + 2 * p.N + n0 + 4));
          }
          if (!waited) { if (warp == 2 && it < 30) GEMM_TRACE(8 + it * 8 + 5); mbar_wait(&acc_full[buf], (it >> 1) & 1); tc_fence_after(); waited = true; if (warp == 2 && it < 30) GEMM_TRACE(8 + it * 8 + 6); }
          float v[16];
          tmem_ld16(trow + c0, v);
#pragma unroll
          for (int j = 0; j < 16; j += 4) *reinterpret_cast<float4*>(stg + lane * 20 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
          __syncwarp();
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int rr = er8 + 16 * i;
            const int m = mt * BM + quad * 32 + rr;
            const float4 a0 = *reinterpret_cast<const float4*>(stg + rr * 20 + ec8), a1 = *reinterpret_cast<const float4*>(stg + rr * 20 + ec8 + 4);
            if (m < p.M) epilogue_apply8<E>(p, a0, a1, Cp, aux, b4[0], b4[1], alpha, m, n0, coef != nullptr, cf0, cf1, ops[i][0], ops[i][1]);
          }
          __syncwarp();
          continue;
        }
        // global operands of this chunk's 4 row groups first: they do not depend on the accumulator
        const int n0 = nt * BN + c0 + ec;
        EpiOps ops[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int m = mt * BM + quad * 32 + er + 8 * i;
          if (m < p.M) ops[i] = epilogue_load<E>(p, aux, m, n0, coef != nullptr);
        }
        const float4 b4 = bias ? __ldg(reinterpret_cast<const float4*>(bias + n0)) : make_float4(0.f, 0.f, 0.f, 0.f);
        EpiCoef cf{};
        if (coef) { cf.a = __ldg(reinterpret_cast<const float4*>(coef + n0)); cf.e = __ldg(reinterpret_cast<const float4*>(coef + p.N + n0)); cf.f = __ldg(reinterpret_cast<const float4*>(coef + 2 * p.N + n0)); }
        if (!waited) { if (warp == 2 && it < 30) GEMM_TRACE(8 + it * 8 + 5); mbar_wait(&acc_full[buf], (it >> 1) & 1); tc_fence_after(); waited = true; if (warp == 2 && it < 30) GEMM_TRACE(8 + it * 8 + 6); }
        float v[16];
        tmem_ld16(trow + c0, v);
#pragma unroll
        for (int j = 0; j < 16; j += 4) *reinterpret_cast<float4*>(stg + lane * 20 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int rr = er + 8 * i;
          const int m = mt * BM + quad * 32 + rr;
          const float4 acc = *reinterpret_cast<const float4*>(stg + rr * 20 + ec);
          if (m < p.M) epilogue_apply<E>(p, acc, Cp, aux, b4, alpha, m, n0, coef != nullptr, cf, ops[i]);
        }
        __syncwarp();
      }
      if (!waited) { mbar_wait(&acc_full[buf], (it >> 1) & 1); tc_fence_after(); }
      if (warp == 2 && it < 30) GEMM_TRACE(8 + it * 8 + 7);
      tc_fence_before();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) GEMM_TRACE(2);
  if (warp == 1) tmem_dealloc(tmem_base, 2 * BN);
#undef GEMM_TRACE
}

__global__ void __launch_bounds__(256) transpose_weights_kernel(const float* __restrict__ src, float* __restrict__ dst, float* __restrict__ dst2,
                                                                const TrDesc* __restrict__ descs, int ndesc, int round) {
  pdl_wait(); pdl_trigger();
  __shared__ float t[32][33];
  const TrDesc e = descs[find_desc(descs, ndesc, (int)blockIdx.x)];
  const int tile = blockIdx.x - e.tile0, tn = cdiv(e.N, 32);
  const int k0 = (tile / tn) * 32, n0 = (tile % tn) * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int i = ty; i < 32; i += 8) {
    const int k = k0 + i, n = n0 + tx;
    float v = (k < e.K && n < e.N) ? src[e.off + (long long)k * e.N + n] : 0.f;       // both copies are tensor-core operands
    if (round) v = round_tf32(v);
    t[i][tx] = v;
    if (e.copy && k < e.K && n < e.N) dst2[e.off + (long long)k * e.N + n] = v;
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) { const int n = n0 + i, k = k0 + tx; if (n < e.N && k < e.K) dst[e.toff + (long long)n * e.ldt + k] = t[tx][i]; }
}

// ---- tensor maps --------------------------------------------------------------------------------------------------------------
typedef CUresult (*encode_fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
encode_fn_t encode_fn() {
  static encode_fn_t fn = nullptr;
  if (!fn) {
    void* f = nullptr; cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<encode_fn_t>(f);
  }
  return fn;
}
}  // namespace

// fp32 tensor [d2][d1][d0] (d0 contiguous), strides in floats, box {32, box_rows, 1}, 128-byte swizzle, zero fill out of bounds
bool make_tmap_3d(CUtensorMap* out, const float* base, long long d0, long long d1, long long d2, long long stride1, long long stride2,
                  int box_rows, int swizzle_mode) {
  typedef std::tuple<const float*, long long, long long, long long, long long, long long, int, int> Key;
  static thread_local std::map<Key, CUtensorMap> cache;
  const Key key{base, d0, d1, d2, stride1, stride2, box_rows, swizzle_mode};
  auto it = cache.find(key);
  if (it != cache.end()) { *out = it->second; return true; }
  encode_fn_t fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled entry point not found"); return false; }
  if (d2 < 1) d2 = 1;
  if (d2 == 1) stride2 = d1 * stride1;
  const cuuint64_t dims[3] = {(cuuint64_t)d0, (cuuint64_t)d1, (cuuint64_t)d2};
  const cuuint64_t strides[2] = {(cuuint64_t)stride1 * 4, (cuuint64_t)stride2 * 4};
  const cuuint32_t box[3] = {32, (cuuint32_t)box_rows, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)swizzle_mode, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with code " + std::to_string((int)r)); return false; }
  cache[key] = *out;
  return true;
}

bool conv_tma_box(int ho, int wo, int tokens, ConvBox* box) {
  if (wo < 1 || ho < 1 || tokens % wo != 0) return false;
  const int rows = tokens / wo;                        // output rows per box
  if (rows <= ho) { if (ho % rows != 0) return false; *box = ConvBox{wo, rows, 1}; }
  else { if (rows % ho != 0) return false; *box = ConvBox{wo, ho, rows / ho}; }
  return true;
}

bool make_tmap_conv(CUtensorMap* out, const float* base, int hin, int win, int cin, long long bstride, int batch, int stride, ConvBox box,
                    int swizzle_mode) {
  typedef std::tuple<const float*, int, int, int, long long, int, int, int, int, int, int> Key;
  static thread_local std::map<Key, CUtensorMap> cache;
  const Key key{base, hin, win, cin, bstride, batch, stride, box.bw, box.bh, box.bb, swizzle_mode};
  auto it = cache.find(key);
  if (it != cache.end()) { *out = it->second; return true; }
  encode_fn_t fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled entry point not found"); return false; }
  const cuuint64_t dims[4] = {(cuuint64_t)cin, (cuuint64_t)win, (cuuint64_t)hin, (cuuint64_t)batch};
  const cuuint64_t strides[3] = {(cuuint64_t)cin * 4, (cuuint64_t)win * cin * 4, (cuuint64_t)bstride * 4};
  // a strided dimension loads ceil(boxDim / elementStride) elements: boxDim = pixels * stride
  const cuuint32_t bx[4] = {32, (cuuint32_t)(box.bw * stride), (cuuint32_t)(box.bh * stride), (cuuint32_t)box.bb};
  const cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  if (bx[1] > 256 || bx[2] > 256 || bx[3] > 256 || stride > 8) { set_error("make_tmap_conv: box exceeds the TMA limits"); return false; }
  const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(base), dims, strides, bx, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)swizzle_mode, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (conv) failed with code " + std::to_string((int)r)); return false; }
  cache[key] = *out;
  return true;
}

namespace {
template <int BN, uint32_t E, bool X3>
cudaError_t launch_tma_cfg(const GemmP& p, cudaStream_t s) {
  static bool configured = false;
  if (!configured) {
    CVT_CUDA(cudaFuncSetAttribute(gemm_tma_kernel<BN, E, X3>, cudaFuncAttributeMaxDynamicSharedMemorySize, TCfg<BN, X3>::SMEM));
    configured = true;
  }
  CUtensorMap tmA, tmB;
  if (p.a_mode == A_CONV) {
    ConvBox box;
    if (!conv_tma_box(p.cv.ho, p.cv.wo, BM, &box)) { set_error("gemm_tma: conv geometry does not tile"); return cudaErrorInvalidValue; }
    const int nimg = p.M / (p.cv.ho * p.cv.wo);
    if (!make_tmap_conv(&tmA, p.A + p.cv.boff, p.cv.hin, p.cv.win, p.cv.cin, p.cv.bstride, nimg, p.cv.s, box, CU_TENSOR_MAP_SWIZZLE_128B)) return cudaErrorInvalidValue;
  } else
  if (!make_tmap_3d(&tmA, p.A, p.K, p.M, p.batch, p.lda, p.a_bs0, BM, CU_TENSOR_MAP_SWIZZLE_128B)) return cudaErrorInvalidValue;
  if (!make_tmap_3d(&tmB, p.B, p.K, p.N, p.batch, p.ldb, p.b_bs0, BN, CU_TENSOR_MAP_SWIZZLE_128B)) return cudaErrorInvalidValue;
  const int m_tiles = cdiv(p.M, BM), n_tiles = p.N / BN, num_tiles = m_tiles * n_tiles * p.batch;
  const int grid = num_tiles < kNumSMs ? num_tiles : kNumSMs;
  CVT_LAUNCH((gemm_tma_kernel<BN, E, X3>), grid, (TCfg<BN, X3>::THREADS), (TCfg<BN, X3>::SMEM), s, p, tmA, tmB, m_tiles, n_tiles, num_tiles);
  return cudaSuccess;
}
template <uint32_t E>
cudaError_t launch_tma_bn(const GemmP& p, int bn, cudaStream_t s) {
  if (p.x3) {
    if (bn >= 128) return launch_tma_cfg<128, E, true>(p, s);
    return launch_tma_cfg<64, E, true>(p, s);
  }
  if (bn == 256) return launch_tma_cfg<256, E, false>(p, s);
  if (bn == 128) return launch_tma_cfg<128, E, false>(p, s);
  return launch_tma_cfg<64, E, false>(p, s);
}
constexpr uint32_t kTmaEpiKinds = EPI_GELU | EPI_RESID | EPI_GELU_BWD | EPI_PHILOX | EPI_BN_BWD | EPI_GELU_DSAVE;
}  // namespace

cudaError_t launch_transpose_weights(const float* src, float* dst, float* dst2, const TrDesc* descs_dev, int ndesc, int total_tiles, cudaStream_t s, int round) {
  if (ndesc < 1 || total_tiles < 1) return cudaSuccess;
  CVT_LAUNCH((transpose_weights_kernel), total_tiles, 256, 0, s, src, dst, dst2, descs_dev, ndesc, round);
  return cudaSuccess;
}

// ConvEmbed as an implicit GEMM: 32-channel chunks per tap, 128 consecutive output pixels = whole rows or whole images
bool gemm_tma_conv_supports(const GemmP& p) {
  const ConvGeom& g = p.cv; ConvBox box;
  if (p.a_mode != A_CONV || p.b_mode != B_NK || p.batch != 1 || p.splitk != 1 || p.a_map.period || p.a_scale) return false;
  if (g.cin < 32 || (g.cin & 31) || g.s < 1 || g.s > 8 || p.K != g.k * g.k * g.cin || p.M % (g.ho * g.wo) != 0) return false;
  if (!conv_tma_box(g.ho, g.wo, BM, &box) || box.bw * g.s > 256 || box.bh * g.s > 256) return false;
  if (p.N % 64 != 0 || (p.ldb & 3) || (g.bstride & 3) || (g.boff & 3)) return false;
  if ((reinterpret_cast<uintptr_t>(p.A) & 15) || (reinterpret_cast<uintptr_t>(p.B) & 15)) return false;
  if (p.epi & (EPI_RELU | EPI_RELU_BWD | EPI_ATOMIC | EPI_SCATTER)) return false;
  return true;
}

bool gemm_tma_supports(const GemmP& p) {
  if (p.a_mode == A_CONV) return gemm_tma_conv_supports(p);
  if (p.a_mode != A_PLAIN || p.b_mode != B_NK || p.a_map.period || p.a_scale || p.splitk != 1 || p.bdiv != 1) return false;
  if (p.N % 64 != 0 || p.K % 4 != 0 || p.M < 1 || (p.lda & 3) || (p.ldb & 3) || (p.a_bs0 & 3) || (p.b_bs0 & 3)) return false;
  if ((reinterpret_cast<uintptr_t>(p.A) & 15) || (reinterpret_cast<uintptr_t>(p.B) & 15)) return false;
  if (p.epi & (EPI_RELU | EPI_RELU_BWD | EPI_ATOMIC | EPI_SCATTER)) return false;
  if (p.batch > 1 && (p.c_bs1 || p.a_bs1 || p.b_bs1)) return false;
  return true;
}

cudaError_t launch_gemm_tma(const GemmP& p, cudaStream_t s) {
  if (!gemm_tma_supports(p)) { set_error("gemm_tma: unsupported shape / epilogue"); return cudaErrorNotSupported; }
  // widest N tile that still gives every SM a tile
  const long long mt = cdiv(p.M, BM) * (long long)p.batch;
  int bn = 64;
  if (p.N % 256 == 0 && mt * (p.N / 256) >= kNumSMs) bn = 256;
  else if (p.N % 128 == 0 && mt * (p.N / 128) >= kNumSMs) bn = 128;
  const uint32_t kind = p.epi & kTmaEpiKinds;
  if (kind == 0) return launch_tma_bn<0>(p, bn, s);
  if (kind == EPI_RESID) return launch_tma_bn<EPI_RESID>(p, bn, s);
  if (kind == EPI_GELU) return launch_tma_bn<EPI_GELU>(p, bn, s);
  if (kind == EPI_GELU_BWD) return launch_tma_bn<EPI_GELU_BWD>(p, bn, s);
  if (kind == EPI_BN_BWD) return launch_tma_bn<EPI_BN_BWD>(p, bn, s);
  if (kind == (EPI_RESID | EPI_PHILOX)) return launch_tma_bn<EPI_RESID | EPI_PHILOX>(p, bn, s);     // dropout mask drawn in the epilogue
  if (kind == (EPI_GELU | EPI_PHILOX)) return launch_tma_bn<EPI_GELU | EPI_PHILOX>(p, bn, s);
  if (kind == (EPI_GELU | EPI_GELU_DSAVE)) return launch_tma_bn<EPI_GELU | EPI_GELU_DSAVE>(p, bn, s);
  if (kind == (EPI_GELU | EPI_PHILOX | EPI_GELU_DSAVE)) return launch_tma_bn<EPI_GELU | EPI_PHILOX | EPI_GELU_DSAVE>(p, bn, s);
  if (kind == (EPI_GELU_BWD | EPI_GELU_DSAVE)) return launch_tma_bn<EPI_GELU_BWD | EPI_GELU_DSAVE>(p, bn, s);
  set_error("gemm_tma: no instantiation for epilogue " + std::to_string(p.epi));
  return cudaErrorNotSupported;
}

}  // namespace cvt
