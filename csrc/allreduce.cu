// Hand-written sm_100a collectives over symmetric memory (K16 / C1-C6).
//
// The reference's gradient exchange is torch DDP calling ncclAllReduce per bucket
// (mnist_distributed.py:67,96) and its toy is ncclAllReduce on a 0-dim int32
// (allreduce_toy.py:30-31).  Here the same collectives are kernels of ours that load and
// store PEER GPU memory directly over NVLink / NVSwitch:
//
//   one-shot       every rank reads all peers' buffers and reduces locally   (<= ~256 KB:
//                  the 53 KB conv/BN bucket, BN buffers, the toy scalar)
//   two-shot P2P   reduce-scatter by pulling the own shard from every peer, then
//                  all-gather by pulling the reduced shards                   (fallback)
//   two-shot NVLS  multimem.ld_reduce on the own shard (the switch adds the n copies),
//                  multimem.st of the result (the switch replicates it)       (720 MB bucket)
//
// fused into the reduction epilogue: the 1/world_size gradient scale and - optionally - the
// SGD update itself: instead of all-gathering the reduced GRADIENT shard, rank r applies
// W[shard r] -= lr * g and multicasts the updated WEIGHT shard, so the separate optimizer
// pass over the 720 MB fc.weight (K14) disappears and all replicas stay bit-identical.
//
// Cross-GPU synchronisation: per-CTA flag words in every rank's symmetric flag page,
// written with st.release.sys by the peer and polled with ld.acquire.sys locally.  Flags
// carry a monotonically increasing epoch and are never reset; every wait is bounded (a dead
// peer traps after ~20 s instead of wedging the GPU).  All ranks launch identical grids.
#include "common.h"
#include "ptx.cuh"
#include <cuda_bf16.h>
#include <cstdint>

namespace {

constexpr int kMaxRanks = 16;
constexpr int kMaxBlocks = 256;

struct Peers {
  void* ptr[kMaxRanks];       // this buffer on every rank (VAs valid in this process)
  uint32_t* flag[kMaxRanks];  // flag page of every rank for this channel
  void* mc;                   // multicast VA of the buffer (nullptr if none)
  int rank, world;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer16(const void* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 mc_ld_reduce_f32(const void* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_f32(void* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// 8 bf16 per 16 bytes, accumulated in fp32 inside the switch
__device__ __forceinline__ uint4 mc_ld_reduce_bf16(const void* p) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_b32x4(void* p, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// How long an in-kernel wait for a peer may last before it is declared dead.  Set from the process group's
// timeout (init_process_group(timeout=), default 600 s like ProcessGroupNCCL) by b200_ar_set_timeout_ms: rank skew
// of minutes is legitimate (rank 0 writing a 720 MB checkpoint, a first-step build, an eval phase before a barrier).
__device__ unsigned long long g_wait_timeout_ns = 600ull * 1000000000ull;

// Barrier across the same-index CTA of every rank.  Slot layout: flag[(block*kMaxRanks)+src].
__device__ __forceinline__ void sys_barrier(const Peers& P, uint32_t epoch) {
  __threadfence_system();            // every thread publishes its own prior writes at system scope
  __syncthreads();
  const int t = threadIdx.x;
  if (t < P.world) {
    st_release_sys(P.flag[t] + (size_t)blockIdx.x * kMaxRanks + P.rank, epoch);
    const uint32_t* mine = P.flag[P.rank] + (size_t)blockIdx.x * kMaxRanks + t;
    const uint64_t t0 = b2::globaltimer_ns();
    const uint64_t limit = g_wait_timeout_ns;
    uint32_t spins = 0;
    // epochs are compared with wrap-safe signed distance
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if ((++spins & 0x3FF) == 0) {
        if (b2::globaltimer_ns() - t0 > limit) {
          printf("b200dp: rank %d block %d timed out after %llu ms waiting for rank %d (epoch %u)\n", P.rank,
                 blockIdx.x, (unsigned long long)(limit / 1000000ull), t, epoch);
          __trap();
        }
        __nanosleep(spins > 0x10000 ? 2000 : 100);      // a long wait should not hammer the peer's memory
      }
    }
  }
  __syncthreads();
}

__global__ void barrier_kernel(Peers P, uint32_t epoch) { sys_barrier(P, epoch); }

// ------------------------------------------------------------------ one-shot ---
// out (local, any memory) = scale * sum_p in_p   with `in` the symmetric staging buffer.
template <typename T, typename Acc>
__global__ void __launch_bounds__(512)
allreduce_oneshot_kernel(Peers P, T* __restrict__ out, size_t n, float scale, int op, uint32_t epoch) {
  sys_barrier(P, epoch);             // all inputs are in place
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    Acc acc = (Acc) reinterpret_cast<const volatile T*>(P.ptr[0])[i];
    for (int p = 1; p < P.world; ++p) {
      const Acc v = (Acc) reinterpret_cast<const volatile T*>(P.ptr[p])[i];
      if (op == 0) acc += v;
      else if (op == 2) acc = v > acc ? v : acc;
      else if (op == 3) acc = v < acc ? v : acc;
      else acc *= v;
    }
    out[i] = (T)(acc * (Acc)scale);
  }
  sys_barrier(P, epoch + 1);         // inputs may be overwritten after this
}
template <>
__global__ void __launch_bounds__(512)
allreduce_oneshot_kernel<__nv_bfloat16, float>(Peers P, __nv_bfloat16* __restrict__ out, size_t n, float scale, int op,
                                               uint32_t epoch) {
  sys_barrier(P, epoch);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    float acc = 0.f;
    for (int p = 0; p < P.world; ++p) {
      const unsigned short raw = reinterpret_cast<const volatile unsigned short*>(P.ptr[p])[i];
      const float v = __uint_as_float((uint32_t)raw << 16);
      if (p == 0) acc = v;
      else if (op == 0) acc += v;
      else if (op == 2) acc = fmaxf(acc, v);
      else if (op == 3) acc = fminf(acc, v);
      else acc *= v;
    }
    out[i] = __float2bfloat16_rn(acc * scale);
  }
  sys_barrier(P, epoch + 1);
}

// fp32, 16 bytes per thread, every peer's load in flight before the first add; optional fused SGD
// (param -= lr * scale * sum) for the small DDP bucket.  `in` is symmetric, out/param are local.
template <int MAXW>
__global__ void __launch_bounds__(512)
allreduce_oneshot_f32x4_kernel(Peers P, float4* __restrict__ out, float4* __restrict__ param, size_t n4, float scale,
                               float lr, uint32_t epoch, Peers Bc, int bc_n4, int bc_root) {
  sys_barrier(P, epoch);
  // folded broadcast (DDP's per-forward buffer sync, C4): every rank takes the root's copy of a second, tiny
  // symmetric tensor (the 512-byte BatchNorm blob) under the same pair of barriers
  if (Bc.ptr[0] != nullptr && P.rank != bc_root && blockIdx.x == 0 && (int)threadIdx.x < bc_n4)
    reinterpret_cast<float4*>(Bc.ptr[P.rank])[threadIdx.x] =
        ld_peer16(reinterpret_cast<const float4*>(Bc.ptr[bc_root]) + threadIdx.x);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 v[MAXW];
#pragma unroll
    for (int p = 0; p < MAXW; ++p)
      if (p < P.world) v[p] = ld_peer16(reinterpret_cast<const float4*>(P.ptr[p]) + i);
    float4 acc = v[0];
#pragma unroll
    for (int p = 1; p < MAXW; ++p)
      if (p < P.world) { acc.x += v[p].x; acc.y += v[p].y; acc.z += v[p].z; acc.w += v[p].w; }
    acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
    if (out) out[i] = acc;
    if (param) {
      float4 w = param[i];
      w.x = fmaf(-lr, acc.x, w.x); w.y = fmaf(-lr, acc.y, w.y); w.z = fmaf(-lr, acc.z, w.z); w.w = fmaf(-lr, acc.w, w.w);
      param[i] = w;
    }
  }
  sys_barrier(P, epoch + 1);
}


// ------------------------------------------------------------- push one-shot ---
// Latency path for small tensors that live ANYWHERE (no staging copy, no symmetric input): every rank writes its
// data into slot [rank] of every peer's scratch ring (peer stores over NVLink), signals, waits for everybody's
// signal, and reduces its own - now local - copies.  One kernel, one cross-GPU synchronisation.  The scratch has
// two halves used alternately (`buf`): a rank can only be one call ahead of a peer (it needs the peer's signal of
// the current call to finish it), so when it overwrites a half the peer has finished reading it two calls ago and
// no trailing barrier is needed.  The same grid runs on every rank, so the per-CTA flag barrier covers exactly the
// elements this CTA's peers pushed.
template <typename T, typename Acc>
__global__ void __launch_bounds__(512)
allreduce_push_kernel(Peers P, const T* __restrict__ in, T* __restrict__ out, size_t n, size_t slot_bytes, int buf,
                      float scale, int op, uint32_t epoch) {
  const size_t half = (size_t)buf * kMaxRanks * slot_bytes;
  const size_t g0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x, gs = (size_t)gridDim.x * blockDim.x;
  for (int k = 0; k < P.world; ++k) {
    const int p = (P.rank + k) % P.world;
    T* dst = reinterpret_cast<T*>(reinterpret_cast<char*>(P.ptr[p]) + half + (size_t)P.rank * slot_bytes);
    for (size_t i = g0; i < n; i += gs) dst[i] = in[i];
  }
  sys_barrier(P, epoch);             // my stores are visible to the peers, theirs to me
  const char* mine = reinterpret_cast<const char*>(P.ptr[P.rank]) + half;
  for (size_t i = g0; i < n; i += gs) {
    Acc acc = (Acc) reinterpret_cast<const volatile T*>(mine)[i];
    for (int p = 1; p < P.world; ++p) {
      const Acc v = (Acc) reinterpret_cast<const volatile T*>(mine + (size_t)p * slot_bytes)[i];
      if (op == 0) acc += v;
      else if (op == 2) acc = v > acc ? v : acc;
      else if (op == 3) acc = v < acc ? v : acc;
      else acc *= v;
    }
    out[i] = (T)(acc * (Acc)scale);
  }
}
template <>
__global__ void __launch_bounds__(512)
allreduce_push_kernel<__nv_bfloat16, float>(Peers P, const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                            size_t n, size_t slot_bytes, int buf, float scale, int op, uint32_t epoch) {
  const size_t half = (size_t)buf * kMaxRanks * slot_bytes;
  const size_t g0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x, gs = (size_t)gridDim.x * blockDim.x;
  for (int k = 0; k < P.world; ++k) {
    const int p = (P.rank + k) % P.world;
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(P.ptr[p]) + half + (size_t)P.rank * slot_bytes);
    for (size_t i = g0; i < n; i += gs) dst[i] = in[i];
  }
  sys_barrier(P, epoch);
  const char* mine = reinterpret_cast<const char*>(P.ptr[P.rank]) + half;
  for (size_t i = g0; i < n; i += gs) {
    float acc = 0.f;
    for (int p = 0; p < P.world; ++p) {
      const unsigned short raw = reinterpret_cast<const volatile unsigned short*>(mine + (size_t)p * slot_bytes)[i];
      const float v = __uint_as_float((uint32_t)raw << 16);
      if (p == 0) acc = v;
      else if (op == 0) acc += v;
      else if (op == 2) acc = fmaxf(acc, v);
      else if (op == 3) acc = fminf(acc, v);
      else acc *= v;
    }
    out[i] = __float2bfloat16_rn(acc * scale);
  }
}

// -------------------------------------------------------------- two-shot P2P ---
// In place on the symmetric fp32 buffer.  n4 = number of float4; shard r = [lo_r, hi_r).
// If w != nullptr the SGD update is fused: w is this rank's weight buffer (same layout, also
// symmetric: wpeers), phase 1 writes W[shard r] -= lr * g, phase 2 gathers WEIGHT shards.
__global__ void __launch_bounds__(512)
allreduce_twoshot_f32_kernel(Peers P, Peers Wp, size_t n4, float scale, float lr, uint32_t epoch) {
  const int r = P.rank, n = P.world;
  const size_t per = (n4 + n - 1) / n;
  const size_t lo = per * r, hi = min(n4, lo + per);
  const bool fused = Wp.ptr[0] != nullptr;
  sys_barrier(P, epoch);
  for (size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (size_t)gridDim.x * blockDim.x) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int p = 0; p < n; ++p) {    // fixed rank order: every rank would compute identical bits
      const float4 v = ld_peer16(reinterpret_cast<const float4*>(P.ptr[p]) + i);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
    if (fused) {
      float4* wp = reinterpret_cast<float4*>(Wp.ptr[r]) + i;
      float4 w = *wp;
      w.x = fmaf(-lr, acc.x, w.x); w.y = fmaf(-lr, acc.y, w.y); w.z = fmaf(-lr, acc.z, w.z); w.w = fmaf(-lr, acc.w, w.w);
      *wp = w;
    } else {
      reinterpret_cast<float4*>(P.ptr[r])[i] = acc;
    }
  }
  sys_barrier(P, epoch + 1);         // every shard is reduced (and every read of my shards is done)
  const Peers& G = fused ? Wp : P;
  for (int k = 1; k < n; ++k) {
    const int p = (r + k) % n;       // stagger peers so the n ranks do not all hit one GPU at once
    const size_t plo = per * p, phi = min(n4, plo + per);
    for (size_t i = plo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < phi; i += (size_t)gridDim.x * blockDim.x)
      reinterpret_cast<float4*>(G.ptr[r])[i] = ld_peer16(reinterpret_cast<const float4*>(G.ptr[p]) + i);
  }
  sys_barrier(P, epoch + 2);         // nobody still reads my reduced shard
}

// ------------------------------------------------------------- two-shot NVLS ---
// Register budget matters here: this kernel runs UNDER the backward/forward compute kernels.  At <= 32
// registers a 512-thread CTA takes 16 K of the SM's 64 K registers and co-resides with every persistent
// compute CTA of the step; at 46 it could not share an SM with conv1_*_tc and delayed those kernels.
template <int UNROLL>
__global__ void __launch_bounds__(512, 4)
allreduce_nvls_f32_kernel(Peers P, Peers Wp, size_t n4, float scale, float lr, uint32_t epoch) {
  const int r = P.rank, n = P.world;
  const size_t per = (n4 + n - 1) / n;
  const size_t lo = per * r, hi = min(n4, lo + per);
  const bool fused = Wp.ptr[0] != nullptr;
  const float4* src = reinterpret_cast<const float4*>(P.mc);
  float4* dst = reinterpret_cast<float4*>(fused ? Wp.mc : P.mc);
  const float4* wl = reinterpret_cast<const float4*>(Wp.ptr[r]);
  sys_barrier(P, epoch);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + (UNROLL - 1) * stride < hi; i += UNROLL * stride) {
    float4 v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) v[u] = mc_ld_reduce_f32(src + i + u * stride);
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      float4 g = make_float4(v[u].x * scale, v[u].y * scale, v[u].z * scale, v[u].w * scale);
      if (fused) {
        const float4 w = wl[i + u * stride];
        g = make_float4(fmaf(-lr, g.x, w.x), fmaf(-lr, g.y, w.y), fmaf(-lr, g.z, w.z), fmaf(-lr, g.w, w.w));
      }
      mc_st_f32(dst + i + u * stride, g);
    }
  }
  for (; i < hi; i += stride) {
    const float4 v = mc_ld_reduce_f32(src + i);
    float4 g = make_float4(v.x * scale, v.y * scale, v.z * scale, v.w * scale);
    if (fused) {
      const float4 w = wl[i];
      g = make_float4(fmaf(-lr, g.x, w.x), fmaf(-lr, g.y, w.y), fmaf(-lr, g.z, w.z), fmaf(-lr, g.w, w.w));
    }
    mc_st_f32(dst + i, g);
  }
  sys_barrier(P, epoch + 1);         // all multicast stores have landed everywhere
}

// "Exclusive" variant for the overlapped DDP bucket: few, fat CTAs (1024 threads, 8 x 16 B in flight per
// thread) that request kExclusiveSmem bytes of dynamic shared memory they never touch, so no persistent
// tensor-core CTA (all need > 67 KB) can share their SM.  See b200_set_reserved_sms (common.cu).
constexpr int kExclusiveSmem = 160 * 1024;
__global__ void __launch_bounds__(1024, 1)
allreduce_nvls_f32_excl_kernel(Peers P, Peers Wp, size_t n4, float scale, float lr, uint32_t epoch) {
  const int r = P.rank, n = P.world;
  const size_t per = (n4 + n - 1) / n;
  const size_t lo = per * r, hi = min(n4, lo + per);
  const bool fused = Wp.ptr[0] != nullptr;
  const float4* src = reinterpret_cast<const float4*>(P.mc);
  float4* dst = reinterpret_cast<float4*>(fused ? Wp.mc : P.mc);
  const float4* wl = reinterpret_cast<const float4*>(Wp.ptr[r]);
  sys_barrier(P, epoch);
  constexpr int U = 8;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + (U - 1) * stride < hi; i += U * stride) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = mc_ld_reduce_f32(src + i + u * stride);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float4 g = make_float4(v[u].x * scale, v[u].y * scale, v[u].z * scale, v[u].w * scale);
      if (fused) {
        const float4 w = wl[i + u * stride];
        g = make_float4(fmaf(-lr, g.x, w.x), fmaf(-lr, g.y, w.y), fmaf(-lr, g.z, w.z), fmaf(-lr, g.w, w.w));
      }
      mc_st_f32(dst + i + u * stride, g);
    }
  }
  for (; i < hi; i += stride) {
    const float4 v = mc_ld_reduce_f32(src + i);
    float4 g = make_float4(v.x * scale, v.y * scale, v.z * scale, v.w * scale);
    if (fused) {
      const float4 w = wl[i];
      g = make_float4(fmaf(-lr, g.x, w.x), fmaf(-lr, g.y, w.y), fmaf(-lr, g.z, w.z), fmaf(-lr, g.w, w.w));
    }
    mc_st_f32(dst + i, g);
  }
  sys_barrier(P, epoch + 1);
}


// ------------------------------------------------ tile-driven fused bucket ---
// K10 -> K16 pipeline: the fc backward kernel produces dW[10][F] slab by slab along F (all ten
// rows of a column range at a time) and bumps a per-slab counter in symmetric memory when a block has written its
// part (fc_bwd_kernel, `ready`).  This kernel is launched TOGETHER with the fc backward, on the side stream, and
// consumes slabs as they become ready ON EVERY RANK: for slab t, rank r owns the r-th 1/world of its columns,
// reduces them across ranks (NVLS: multimem.ld_reduce; P2P: one 16-byte load per peer, all in flight), applies
// the 1/world scale and the SGD update to its fp32 master weights and writes the new WEIGHTS to every rank
// (multimem.st / peer stores) - one pass, no reduce-scatter / all-gather split, no optimizer pass.
// A slab's weights are only overwritten after every rank's fc backward has finished reading them (the counter is
// bumped after the block's loads of that slab), and the trailing barrier orders the remote stores before the
// next forward.  Every wait is bounded by g_wait_timeout_ns.
struct TileGeom {
  size_t F4;           // float4 per row (F / 4)
  size_t slab4;        // float4 columns per slab
  int nslabs, rows;    // rows <= kTileRows
  uint32_t target;     // counter value that means "slab complete on that rank" (monotonic across steps)
};
constexpr int kTileRows = 10;      // rows of the fc.weight bucket (classes)

template <bool NVLS>
__global__ void __launch_bounds__(1024, 1)
allreduce_tile_f32_kernel(Peers P, Peers Wp, Peers Rdy, TileGeom tg, float scale, float lr, uint32_t epoch) {
  const int r = P.rank, n = P.world;
  const size_t tstride = (size_t)gridDim.x * blockDim.x;
  const float a = -lr * scale;
  for (int t = 0; t < tg.nslabs; ++t) {
    // ---- wait until slab t is complete on every rank
    if (threadIdx.x < n) {
      const uint32_t* cnt = reinterpret_cast<const uint32_t*>(Rdy.ptr[threadIdx.x]) + t;
      const uint64_t t0 = b2::globaltimer_ns();
      const uint64_t limit = g_wait_timeout_ns;
      uint32_t spins = 0;
      while ((int32_t)(ld_acquire_sys(cnt) - tg.target) < 0) {
        if ((++spins & 0xFF) == 0) {
          if (b2::globaltimer_ns() - t0 > limit) {
            printf("b200dp: rank %d timed out after %llu ms waiting for gradient slab %d of rank %d\n", r,
                   (unsigned long long)(limit / 1000000ull), t, (int)threadIdx.x);
            __trap();
          }
          __nanosleep(200);
        }
      }
    }
    __syncthreads();
    const size_t c0 = (size_t)t * tg.slab4;
    const size_t c1 = min(tg.F4, c0 + tg.slab4);
    const size_t per = (c1 - c0 + n - 1) / n;
    const size_t lo = min(c1, c0 + per * r), hi = min(c1, lo + per);
    if (NVLS) {
      // Eight independent switch round trips per thread (the register budget of 1024 threads): the link is latency
      // bound, ~128 KB in flight per SM is what the whole-bucket kernel above runs with.  Rows are F apart, so a
      // row's accesses stay coalesced across the threads.  Rows 0..7 of one column, then the remaining rows eight
      // columns at a time.
      const float4* src = reinterpret_cast<const float4*>(P.mc);
      float4* dst = reinterpret_cast<float4*>(Wp.mc);
      const float4* wl = reinterpret_cast<const float4*>(Wp.ptr[r]);
      const int full = tg.rows & ~7, rem = tg.rows & 7;
      for (int k0 = 0; k0 < full; k0 += 8)
        for (size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += tstride) {
          float4 v[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) v[k] = mc_ld_reduce_f32(src + (size_t)(k0 + k) * tg.F4 + i);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const size_t e = (size_t)(k0 + k) * tg.F4 + i;
            const float4 w = wl[e];                    // local HBM: short next to the switch round trip above
            mc_st_f32(dst + e, make_float4(fmaf(a, v[k].x, w.x), fmaf(a, v[k].y, w.y), fmaf(a, v[k].z, w.z),
                                           fmaf(a, v[k].w, w.w)));
          }
        }
      for (int k = full; k < full + rem; ++k) {        // left-over rows: eight columns of one row at a time
        const float4* srow = src + (size_t)k * tg.F4;
        float4* drow = dst + (size_t)k * tg.F4;
        const float4* wrow = wl + (size_t)k * tg.F4;
        size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + 7 * tstride < hi; i += 8 * tstride) {
          float4 v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) v[u] = mc_ld_reduce_f32(srow + i + u * tstride);
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const float4 w = wrow[i + u * tstride];
            mc_st_f32(drow + i + u * tstride, make_float4(fmaf(a, v[u].x, w.x), fmaf(a, v[u].y, w.y),
                                                          fmaf(a, v[u].z, w.z), fmaf(a, v[u].w, w.w)));
          }
        }
        for (; i < hi; i += tstride) {
          const float4 v = mc_ld_reduce_f32(srow + i);
          const float4 w = wrow[i];
          mc_st_f32(drow + i, make_float4(fmaf(a, v.x, w.x), fmaf(a, v.y, w.y), fmaf(a, v.z, w.z), fmaf(a, v.w, w.w)));
        }
      }
    } else {
      // P2P: five rows of one column per step; every peer's load is in flight before the first add
      for (size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += tstride) {
#pragma unroll 1
        for (int k0 = 0; k0 < tg.rows; k0 += 5) {
          float4 acc[5];
#pragma unroll
          for (int k = 0; k < 5; ++k) {
            acc[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (k0 + k < tg.rows) {
              const size_t e = (size_t)(k0 + k) * tg.F4 + i;
#pragma unroll 8
              for (int p = 0; p < n; ++p) {        // fixed rank order: every owner would compute identical bits
                const float4 v = ld_peer16(reinterpret_cast<const float4*>(P.ptr[p]) + e);
                acc[k].x += v.x; acc[k].y += v.y; acc[k].z += v.z; acc[k].w += v.w;
              }
            }
          }
#pragma unroll
          for (int k = 0; k < 5; ++k)
            if (k0 + k < tg.rows) {
              const size_t e = (size_t)(k0 + k) * tg.F4 + i;
              const float4 w = reinterpret_cast<const float4*>(Wp.ptr[r])[e];
              const float4 o = make_float4(fmaf(a, acc[k].x, w.x), fmaf(a, acc[k].y, w.y),
                                           fmaf(a, acc[k].z, w.z), fmaf(a, acc[k].w, w.w));
              for (int p = 0; p < n; ++p)          // push the new weights to every replica (own copy included)
                reinterpret_cast<float4*>(Wp.ptr[(r + p) % n])[e] = o;
            }
        }
      }
    }
  }
  sys_barrier(P, epoch);             // all weight stores have landed on every rank
}

// bf16 buffer (8 elements per 16 B), fp32 accumulation in the switch; scale applied here.
__global__ void __launch_bounds__(512)
allreduce_nvls_bf16_kernel(Peers P, size_t n16, float scale, uint32_t epoch) {
  const int r = P.rank, n = P.world;
  const size_t per = (n16 + n - 1) / n;
  const size_t lo = per * r, hi = min(n16, lo + per);
  uint4* mc = reinterpret_cast<uint4*>(P.mc);
  sys_barrier(P, epoch);
  for (size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (size_t)gridDim.x * blockDim.x) {
    uint4 v = mc_ld_reduce_bf16(mc + i);
    if (scale != 1.0f) {
      v.x = b2::pack_bf16x2(b2::bf16_lo(v.x) * scale, b2::bf16_hi(v.x) * scale);
      v.y = b2::pack_bf16x2(b2::bf16_lo(v.y) * scale, b2::bf16_hi(v.y) * scale);
      v.z = b2::pack_bf16x2(b2::bf16_lo(v.z) * scale, b2::bf16_hi(v.z) * scale);
      v.w = b2::pack_bf16x2(b2::bf16_lo(v.w) * scale, b2::bf16_hi(v.w) * scale);
    }
    mc_st_b32x4(mc + i, v);
  }
  sys_barrier(P, epoch + 1);
}

// DDP bucket exchanged in bf16 (ref: torch DDP's bf16_compress_hook on top of mnist_distributed.py:67 - the
// reference itself moves fp32): the fc backward writes its gradient as bf16 into a symmetric bucket, the switch adds
// the ranks' values in fp32 (multimem.ld_reduce ... acc::f32 ... bf16x2), and the 1/world scale, the SGD update of the
// fp32 master weights and the multicast of the new WEIGHTS are the epilogue of the same pass.  Half the gradient
// bytes on the wire and in HBM; the weights stay fp32 everywhere.  Few fat CTAs that own their SMs, like
// allreduce_nvls_f32_excl_kernel.
__global__ void __launch_bounds__(1024, 1)
allreduce_nvls_bf16_sgd_excl_kernel(Peers P, Peers Wp, size_t n8, float scale, float lr, uint32_t epoch) {
  const int r = P.rank, n = P.world;
  const size_t per = (n8 + n - 1) / n;
  const size_t lo = per * r, hi = min(n8, lo + per);
  const uint4* src = reinterpret_cast<const uint4*>(P.mc);
  float4* wdst = reinterpret_cast<float4*>(Wp.mc);
  const float4* wl = reinterpret_cast<const float4*>(Wp.ptr[r]);
  const float a = -lr * scale;
  sys_barrier(P, epoch);
  constexpr int U = 4;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  auto apply = [&](size_t idx, const uint4& v) {
    const float4 w0 = wl[2 * idx], w1 = wl[2 * idx + 1];
    mc_st_f32(wdst + 2 * idx, make_float4(fmaf(a, b2::bf16_lo(v.x), w0.x), fmaf(a, b2::bf16_hi(v.x), w0.y),
                                          fmaf(a, b2::bf16_lo(v.y), w0.z), fmaf(a, b2::bf16_hi(v.y), w0.w)));
    mc_st_f32(wdst + 2 * idx + 1, make_float4(fmaf(a, b2::bf16_lo(v.z), w1.x), fmaf(a, b2::bf16_hi(v.z), w1.y),
                                              fmaf(a, b2::bf16_lo(v.w), w1.z), fmaf(a, b2::bf16_hi(v.w), w1.w)));
  };
  for (; i + (U - 1) * stride < hi; i += U * stride) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = mc_ld_reduce_bf16(src + i + u * stride);
#pragma unroll
    for (int u = 0; u < U; ++u) apply(i + u * stride, v[u]);
  }
  for (; i < hi; i += stride) apply(i, mc_ld_reduce_bf16(src + i));
  sys_barrier(P, epoch + 1);
}

// broadcast: every rank copies `bytes` (multiple of 16) from the root's symmetric buffer
__global__ void __launch_bounds__(512)
broadcast_kernel(Peers P, float4* __restrict__ out, size_t n16, int root, uint32_t epoch) {
  sys_barrier(P, epoch);
  if (P.rank != root || out != reinterpret_cast<float4*>(P.ptr[root]))
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
      out[i] = ld_peer16(reinterpret_cast<const float4*>(P.ptr[root]) + i);
  sys_barrier(P, epoch + 1);
}

// all-gather: out[p*n16 + i] = peer p's buffer[i]
__global__ void __launch_bounds__(512)
allgather_kernel(Peers P, float4* __restrict__ out, size_t n16, uint32_t epoch) {
  sys_barrier(P, epoch);
  for (int k = 0; k < P.world; ++k) {
    const int p = (P.rank + k) % P.world;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
      out[(size_t)p * n16 + i] = ld_peer16(reinterpret_cast<const float4*>(P.ptr[p]) + i);
  }
  sys_barrier(P, epoch + 1);
}

Peers make_peers(void* const* ptrs, void* const* flags, void* mc, int rank, int world) {
  Peers P;
  for (int i = 0; i < kMaxRanks; ++i) {
    P.ptr[i] = (ptrs && i < world) ? ptrs[i] : nullptr;
    P.flag[i] = (flags && i < world) ? (uint32_t*)flags[i] : nullptr;
  }
  P.mc = mc; P.rank = rank; P.world = world;
  return P;
}

}  // namespace

// Bytes one channel's flag page needs on every rank.
// Per device (the symbol lives in each device's copy of the module): call with the communicator's device current.
B200_API int b200_ar_set_timeout_ms(long long ms) {
  if (ms < 1) return b200_fail(-2, "collective timeout must be >= 1 ms");
  const unsigned long long ns = (unsigned long long)ms * 1000000ull;
  B200_CUDA(cudaMemcpyToSymbol(g_wait_timeout_ns, &ns, sizeof(ns)));
  return 0;
}

B200_API size_t b200_ar_flag_bytes() { return (size_t)kMaxBlocks * kMaxRanks * sizeof(uint32_t); }
B200_API int b200_ar_max_blocks() { return kMaxBlocks; }

// Epoch convention: every call consumes `b200_ar_epochs(kind)` consecutive epochs starting at
// `epoch` (the Python layer keeps one counter per flag channel, identical on all ranks).
B200_API int b200_ar_barrier(void* const* flags, int rank, int world, uint32_t epoch, void* stream) {
  Peers P = make_peers(nullptr, flags, nullptr, rank, world);
  barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(P, epoch);
  B200_LAUNCH_CHECK();
  return 0;
}

// dtype: 0 f32, 2 i32, 3 i64, 5 bf16, 1 f64.  op: 0 SUM, 2 MAX, 3 MIN, 4 PROD (AVG = SUM with scale).
B200_API int b200_ar_oneshot(void* const* ptrs, void* const* flags, int rank, int world, void* out, size_t n,
                             int dtype, int op, float scale, int blocks, uint32_t epoch, void* stream) {
  B200_CHECK(world <= kMaxRanks && blocks >= 1 && blocks <= kMaxBlocks, "oneshot: bad world/blocks");
  Peers P = make_peers(ptrs, flags, nullptr, rank, world);
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case 0: allreduce_oneshot_kernel<float, float><<<blocks, 512, 0, s>>>(P, (float*)out, n, scale, op, epoch); break;
    case 1: allreduce_oneshot_kernel<double, double><<<blocks, 512, 0, s>>>(P, (double*)out, n, scale, op, epoch); break;
    case 2: allreduce_oneshot_kernel<int, int><<<blocks, 512, 0, s>>>(P, (int*)out, n, scale, op, epoch); break;
    case 3: allreduce_oneshot_kernel<long long, long long><<<blocks, 512, 0, s>>>(P, (long long*)out, n, scale, op, epoch); break;
    case 5: allreduce_oneshot_kernel<__nv_bfloat16, float><<<blocks, 512, 0, s>>>(P, (__nv_bfloat16*)out, n, scale, op, epoch); break;
    default: return b200_fail(-2, "oneshot: unsupported dtype code %d", dtype);
  }
  B200_LAUNCH_CHECK();
  return 0;
}

// Small all-reduce of an ordinary (non-symmetric) tensor through the push ring.  ptrs = every rank's mapping of the
// scratch ring ([2 halves][kMaxRanks slots][slot_bytes]); in/out are local (may alias); n * element size <= slot_bytes.
// dtype: 0 f32, 1 f64, 2 i32, 3 i64, 5 bf16; op: 0 sum, 2 max, 3 min, 4 prod.
B200_API int b200_ar_push(void* const* ptrs, void* const* flags, int rank, int world, const void* in, void* out,
                          size_t n, int dtype, int op, float scale, size_t slot_bytes, int buf, int blocks,
                          uint32_t epoch, void* stream) {
  B200_CHECK(world <= kMaxRanks && blocks >= 1 && blocks <= kMaxBlocks && (buf == 0 || buf == 1), "ar_push: bad arguments");
  Peers P = make_peers(ptrs, flags, nullptr, rank, world);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t es = dtype == 0 || dtype == 2 ? 4 : dtype == 5 ? 2 : 8;
  B200_CHECK(n * es <= slot_bytes, "ar_push: %zu bytes do not fit a %zu-byte slot", n * es, slot_bytes);
  switch (dtype) {
    case 0: allreduce_push_kernel<float, float><<<blocks, 512, 0, s>>>(P, (const float*)in, (float*)out, n, slot_bytes, buf, scale, op, epoch); break;
    case 1: allreduce_push_kernel<double, double><<<blocks, 512, 0, s>>>(P, (const double*)in, (double*)out, n, slot_bytes, buf, scale, op, epoch); break;
    case 2: allreduce_push_kernel<int, int><<<blocks, 512, 0, s>>>(P, (const int*)in, (int*)out, n, slot_bytes, buf, scale, op, epoch); break;
    case 3: allreduce_push_kernel<long long, long long><<<blocks, 512, 0, s>>>(P, (const long long*)in, (long long*)out, n, slot_bytes, buf, scale, op, epoch); break;
    case 5: allreduce_push_kernel<__nv_bfloat16, float><<<blocks, 512, 0, s>>>(P, (const __nv_bfloat16*)in, (__nv_bfloat16*)out, n, slot_bytes, buf, scale, op, epoch); break;
    default: return b200_fail(-2, "ar_push: unsupported dtype %d", dtype);
  }
  B200_LAUNCH_CHECK();
  return 0;
}

// Small fp32 SUM straight from a symmetric input (no staging copy): out (local, may be null) = scale*sum,
// and/or param -= lr * scale * sum.  n % 4 == 0.
static int oneshot_f32x4(void* const* ptrs, void* const* flags, int rank, int world, void* out, void* param, size_t n,
                         float scale, float lr, int blocks, uint32_t epoch, void* stream, void* const* bc_ptrs,
                         size_t bc_floats, int bc_root) {
  B200_CHECK(world <= kMaxRanks && blocks >= 1 && blocks <= kMaxBlocks && n % 4 == 0, "oneshot_f32x4: bad arguments");
  B200_CHECK(bc_floats % 4 == 0 && bc_floats <= 2048 && bc_root >= 0 && bc_root < world, "oneshot_f32x4: bad broadcast");
  Peers P = make_peers(ptrs, flags, nullptr, rank, world);
  Peers Bc = make_peers(bc_ptrs, nullptr, nullptr, rank, world);
  cudaStream_t s = (cudaStream_t)stream;
  if (world <= 8)
    allreduce_oneshot_f32x4_kernel<8><<<blocks, 512, 0, s>>>(P, (float4*)out, (float4*)param, n / 4, scale, lr, epoch, Bc,
                                                             (int)(bc_floats / 4), bc_root);
  else
    allreduce_oneshot_f32x4_kernel<16><<<blocks, 512, 0, s>>>(P, (float4*)out, (float4*)param, n / 4, scale, lr, epoch, Bc,
                                                              (int)(bc_floats / 4), bc_root);
  B200_LAUNCH_CHECK();
  return 0;
}
B200_API int b200_ar_oneshot_f32x4(void* const* ptrs, void* const* flags, int rank, int world, void* out, void* param,
                                   size_t n, float scale, float lr, int blocks, uint32_t epoch, void* stream) {
  return oneshot_f32x4(ptrs, flags, rank, world, out, param, n, scale, lr, blocks, epoch, stream, nullptr, 0, 0);
}
// Same, plus "every rank copies rank `bc_root`'s `bc_floats` floats of a second symmetric tensor" in the same kernel.
B200_API int b200_ar_oneshot_f32x4_bcast(void* const* ptrs, void* const* flags, int rank, int world, void* out,
                                         void* param, size_t n, float scale, float lr, int blocks, uint32_t epoch,
                                         void* stream, void* const* bc_ptrs, size_t bc_floats, int bc_root) {
  return oneshot_f32x4(ptrs, flags, rank, world, out, param, n, scale, lr, blocks, epoch, stream, bc_ptrs, bc_floats,
                       bc_root);
}

// In-place fp32 SUM over the symmetric buffer (n floats, n % 4 == 0).  algo: 1 = two-shot P2P,
// 2 = two-shot NVLS (needs mc), 3 = NVLS with SM-exclusive fat CTAs (DDP overlap).  wptrs/wmc != null fuses SGD: W -= lr * (scale * sum g).
B200_API int b200_ar_large_f32(void* const* ptrs, void* mc, void* const* flags, void* const* wptrs, void* wmc,
                               int rank, int world, size_t n, float scale, float lr, int algo, int blocks,
                               uint32_t epoch, void* stream) {
  B200_CHECK(world <= kMaxRanks && blocks >= 1 && blocks <= kMaxBlocks, "allreduce: bad world/blocks");
  B200_CHECK(n % 4 == 0, "allreduce: element count must be a multiple of 4");
  Peers P = make_peers(ptrs, flags, mc, rank, world);
  Peers Wp = make_peers(wptrs, nullptr, wmc, rank, world);
  cudaStream_t s = (cudaStream_t)stream;
  if (algo == 3) {
    B200_CHECK(mc != nullptr && (!wptrs || wmc), "allreduce: NVLS requested but no multicast mapping");
    static cudaError_t attr_err = cudaFuncSetAttribute(allreduce_nvls_f32_excl_kernel,
                                                       cudaFuncAttributeMaxDynamicSharedMemorySize, kExclusiveSmem);
    B200_CUDA(attr_err);
    allreduce_nvls_f32_excl_kernel<<<blocks, 1024, kExclusiveSmem, s>>>(P, Wp, n / 4, scale, lr, epoch);
  } else if (algo == 2) {
    B200_CHECK(mc != nullptr && (!wptrs || wmc), "allreduce: NVLS requested but no multicast mapping");
    allreduce_nvls_f32_kernel<4><<<blocks, 512, 0, s>>>(P, Wp, n / 4, scale, lr, epoch);
  } else {
    allreduce_twoshot_f32_kernel<<<blocks, 512, 0, s>>>(P, Wp, n / 4, scale, lr, epoch);
  }
  B200_LAUNCH_CHECK();
  return 0;
}

// Tile-driven fused bucket all-reduce + SGD for the fc.weight gradient (see allreduce_tile_f32_kernel).
//   ptrs / mc    gradient bucket [rows][F] on every rank / its multicast address (mc null -> P2P loads)
//   wptrs / wmc  fp32 master weights, same layout, symmetric (required: the update is fused)
//   rdy          per-rank pointers to the slab counters the fc backward bumps (b200_fc_bwd_tiles)
//   target       value every counter of a complete slab has reached (blocks-per-slab x steps so far)
B200_API int b200_ar_tile_f32(void* const* ptrs, void* mc, void* const* flags, void* const* wptrs, void* wmc,
                              void* const* rdy, int rank, int world, int rows, size_t F, size_t slab_floats,
                              uint32_t target, float scale, float lr, int blocks, uint32_t epoch, void* stream) {
  B200_CHECK(world <= kMaxRanks && blocks >= 1 && blocks <= kMaxBlocks, "allreduce(tile): bad world/blocks");
  B200_CHECK(F % 4 == 0 && slab_floats % 4 == 0 && slab_floats > 0, "allreduce(tile): F and slab must be multiples of 4");
  B200_CHECK(wptrs != nullptr && rdy != nullptr, "allreduce(tile): weights and slab counters are required");
  B200_CHECK(rows >= 1 && rows <= kTileRows, "allreduce(tile): at most %d rows", kTileRows);
  Peers P = make_peers(ptrs, flags, mc, rank, world);
  Peers Wp = make_peers(wptrs, nullptr, wmc, rank, world);
  Peers Rdy = make_peers(rdy, nullptr, nullptr, rank, world);
  TileGeom tg;
  tg.F4 = F / 4; tg.slab4 = slab_floats / 4;
  tg.nslabs = (int)((tg.F4 + tg.slab4 - 1) / tg.slab4);
  tg.rows = rows; tg.target = target;
  cudaStream_t s = (cudaStream_t)stream;
  if (mc != nullptr && wmc != nullptr) {
    B200_CUDA(set_max_smem(allreduce_tile_f32_kernel<true>, kExclusiveSmem));
    allreduce_tile_f32_kernel<true><<<blocks, 1024, kExclusiveSmem, s>>>(P, Wp, Rdy, tg, scale, lr, epoch);
  } else {
    B200_CUDA(set_max_smem(allreduce_tile_f32_kernel<false>, kExclusiveSmem));
    allreduce_tile_f32_kernel<false><<<blocks, 1024, kExclusiveSmem, s>>>(P, Wp, Rdy, tg, scale, lr, epoch);
  }
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_ar_nvls_bf16(void* mc, void* const* flags, int rank, int world, size_t n, float scale, int blocks,
                               uint32_t epoch, void* stream) {
  B200_CHECK(mc != nullptr && n % 8 == 0, "nvls_bf16: needs a multicast mapping and n %% 8 == 0");
  Peers P = make_peers(nullptr, flags, mc, rank, world);
  allreduce_nvls_bf16_kernel<<<blocks, 512, 0, (cudaStream_t)stream>>>(P, n / 8, scale, epoch);
  B200_LAUNCH_CHECK();
  return 0;
}

// bf16 gradient bucket (symmetric, multicast `mc`) -> fp32 weights (symmetric `wptrs`, multicast `wmc`):
// weights -= lr * scale * sum_over_ranks(grad), written to every rank.  n = number of gradient elements (multiple of 8).
B200_API int b200_ar_nvls_bf16_sgd(void* mc, void* const* flags, void* const* wptrs, void* wmc, int rank, int world,
                                   size_t n, float scale, float lr, int blocks, uint32_t epoch, void* stream) {
  B200_CHECK(mc != nullptr && wptrs != nullptr && wmc != nullptr, "nvls_bf16_sgd: needs multicast mappings of both buckets");
  B200_CHECK(n % 8 == 0 && world <= kMaxRanks && blocks >= 1 && blocks <= kMaxBlocks, "nvls_bf16_sgd: bad n/world/blocks");
  Peers P = make_peers(nullptr, flags, mc, rank, world);
  Peers Wp = make_peers(wptrs, nullptr, wmc, rank, world);
  B200_CUDA(set_max_smem(allreduce_nvls_bf16_sgd_excl_kernel, kExclusiveSmem));
  allreduce_nvls_bf16_sgd_excl_kernel<<<blocks, 1024, kExclusiveSmem, (cudaStream_t)stream>>>(P, Wp, n / 8, scale, lr,
                                                                                            epoch);
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_ar_broadcast(void* const* ptrs, void* const* flags, int rank, int world, void* out, size_t bytes,
                               int root, int blocks, uint32_t epoch, void* stream) {
  B200_CHECK(bytes % 16 == 0, "broadcast: byte count must be a multiple of 16");
  Peers P = make_peers(ptrs, flags, nullptr, rank, world);
  broadcast_kernel<<<blocks, 512, 0, (cudaStream_t)stream>>>(P, (float4*)out, bytes / 16, root, epoch);
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_ar_allgather(void* const* ptrs, void* const* flags, int rank, int world, void* out, size_t bytes,
                               int blocks, uint32_t epoch, void* stream) {
  B200_CHECK(bytes % 16 == 0, "allgather: byte count must be a multiple of 16");
  Peers P = make_peers(ptrs, flags, nullptr, rank, world);
  allgather_kernel<<<blocks, 512, 0, (cudaStream_t)stream>>>(P, (float4*)out, bytes / 16, epoch);
  B200_LAUNCH_CHECK();
  return 0;
}
