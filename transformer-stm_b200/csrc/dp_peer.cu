// Data-parallel step tail as ONE kernel over NVLink peer memory: gradient reduce-scatter + Keras Adam + weight all-gather.
//
// Every rank of the box maps every other rank's workspace (cudaIpc handles exchanged by the host layer; all ranks build the same model,
// so an arena sits at the same offset in every workspace).  After its backward pass a rank
//   1. tells every peer "my gradients of step e are complete" (a flag in the peer's memory) and waits for the same word from all peers;
//   2. owns slice r of the flat arena: reads that slice of every rank's gradient arena through NVLink (rank order 0..N-1, so the sum is the
//      same bit pattern whoever computes it), applies the Adam update to its slice of the weights / moments, and writes the new weights
//      into every rank's weight arena;
//   3. tells every peer "my writes into your weights are done" and waits for all peers before the kernel ends (the next forward then reads
//      weights that are complete, and nobody overwrites gradients a peer is still reading).
// Per rank (N - 1) / N of the arena crosses NVLink once in each direction - 6.4 MB at N = 8 for the 7.35 MB reference model - against
// a ring all-reduce of the whole arena followed by a separate Adam pass over all of it on every rank.  Nothing here is captured into a
// CUDA graph (the epoch is a launch argument); the forward + backward graph replays before it.
#include "model.cuh"

namespace cvt {

namespace {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_v4(const float4* p) {       // never served from a stale L1 line
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
// bounded spin: a lost peer becomes a trap (reported as a launch failure) instead of a hung GPU
__device__ __forceinline__ void wait_flag(const uint32_t* p, uint32_t epoch) {
  for (long long spins = 0; (int32_t)(ld_acquire_sys(p) - epoch) < 0; ++spins) {
    __nanosleep(64);
    if (spins > (1ll << 26)) __trap();
  }
}

struct PeerArgs {
  float* G[CVT_MAX_PEERS]; float* P[CVT_MAX_PEERS]; uint32_t* flags[CVT_MAX_PEERS];      // the same arenas in every rank's workspace
  float* Mo; float* Vo;
  const float* dyn;                                    // {Adam step size, gradient scale (1 / world)}
  long long n4;                                        // float4s in the arena
  int rank, world, sync_grads;
  uint32_t epoch;
};
// flags of one rank: [0, 8) "gradients ready" written by peer p at index p, [8, 16) "weights written", [16] CTA counter
constexpr int kFlagReady = 0, kFlagDone = 8, kFlagCount = 16;

__global__ void __launch_bounds__(256) dp_reduce_adam_kernel(const PeerArgs a) {
  __shared__ int s_last;
  if (blockIdx.x == 0 && threadIdx.x < a.world) {
    __threadfence_system();
    st_release_sys(a.flags[threadIdx.x] + kFlagReady + a.rank, a.epoch);
  }
  if (threadIdx.x < a.world) wait_flag(a.flags[a.rank] + kFlagReady + threadIdx.x, a.epoch);
  __syncthreads();
  const float alpha = a.dyn[0], gs = a.dyn[1];
  const float b1 = 0.9f, b2 = 0.999f, eps = 1e-7f;
  const long long per = (a.n4 + a.world - 1) / a.world, lo = per * a.rank, hi = min(a.n4, lo + per);
  for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (long long)gridDim.x * blockDim.x) {
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (int p = 0; p < a.world; ++p) {
      const float4 t = ld_peer_v4(reinterpret_cast<const float4*>(a.G[p]) + i);
      g.x += t.x; g.y += t.y; g.z += t.z; g.w += t.w;
    }
    if (a.sync_grads)                                  // tests read the reduced gradient back from every rank
      for (int p = 0; p < a.world; ++p) reinterpret_cast<float4*>(a.G[p])[i] = g;
    float4 pp = reinterpret_cast<float4*>(a.P[a.rank])[i];
    float4 mm = reinterpret_cast<float4*>(a.Mo)[i];
    float4 vv = reinterpret_cast<float4*>(a.Vo)[i];
    g.x *= gs; g.y *= gs; g.z *= gs; g.w *= gs;
    mm.x = b1 * mm.x + (1.f - b1) * g.x; mm.y = b1 * mm.y + (1.f - b1) * g.y;
    mm.z = b1 * mm.z + (1.f - b1) * g.z; mm.w = b1 * mm.w + (1.f - b1) * g.w;
    vv.x = b2 * vv.x + (1.f - b2) * g.x * g.x; vv.y = b2 * vv.y + (1.f - b2) * g.y * g.y;
    vv.z = b2 * vv.z + (1.f - b2) * g.z * g.z; vv.w = b2 * vv.w + (1.f - b2) * g.w * g.w;
    pp.x -= alpha * mm.x / (sqrtf(vv.x) + eps); pp.y -= alpha * mm.y / (sqrtf(vv.y) + eps);
    pp.z -= alpha * mm.z / (sqrtf(vv.z) + eps); pp.w -= alpha * mm.w / (sqrtf(vv.w) + eps);
    reinterpret_cast<float4*>(a.Mo)[i] = mm;
    reinterpret_cast<float4*>(a.Vo)[i] = vv;
#pragma unroll 1
    for (int p = 0; p < a.world; ++p) reinterpret_cast<float4*>(a.P[p])[i] = pp;
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(a.flags[a.rank] + kFlagCount, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  // last CTA of this rank: every store of the rank has been fenced
  if (threadIdx.x == 0) a.flags[a.rank][kFlagCount] = 0;
  if (threadIdx.x < a.world) {
    st_release_sys(a.flags[threadIdx.x] + kFlagDone + a.rank, a.epoch);
    wait_flag(a.flags[a.rank] + kFlagDone + threadIdx.x, a.epoch);
  }
}

}  // namespace

cudaError_t launch_dp_reduce_adam(Model* m, cudaStream_t s) {
  if (!m->peer_ready) { set_error("dp_reduce_adam: peer memory is not mapped"); return cudaErrorInvalidValue; }
  PeerArgs a{};
  for (int p = 0; p < m->world; ++p) {
    char* base = reinterpret_cast<char*>(m->peer_base[p]);
    a.G[p] = reinterpret_cast<float*>(base + (reinterpret_cast<char*>(m->G) - reinterpret_cast<char*>(m->workspace)));
    a.P[p] = reinterpret_cast<float*>(base + (reinterpret_cast<char*>(m->P) - reinterpret_cast<char*>(m->workspace)));
    a.flags[p] = reinterpret_cast<uint32_t*>(base + (reinterpret_cast<char*>(m->dp_flags) - reinterpret_cast<char*>(m->workspace)));
  }
  a.Mo = m->Mo; a.Vo = m->Vo; a.dyn = m->lr_dev; a.n4 = m->n_train / 4; a.rank = m->rank; a.world = m->world;
  a.sync_grads = m->dp_sync_grads; a.epoch = ++m->dp_epoch;
  const long long per = (a.n4 + a.world - 1) / a.world;
  int grid = (int)((per + 255) / 256); if (grid > 2 * kNumSMs) grid = 2 * kNumSMs; if (grid < 1) grid = 1;
  dp_reduce_adam_kernel<<<grid, 256, 0, s>>>(a);
  CVT_LAUNCH_CHECK();
  return cudaSuccess;
}

}  // namespace cvt
