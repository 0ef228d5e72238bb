// Layer-1 forward of the precision-matched (tf32) mode, second generation: the 5x5 one-channel
// convolution as a HORIZONTAL Toeplitz product whose A operand is the raw input image.
// (Reference op: nn.Conv2d(1, 16, 5, padding=2) + BatchNorm2d + ReLU + MaxPool2d(2),
// mnist_distributed.py:25-29 - the batch-statistics pass and the conv/BN/ReLU/pool pass of
// b200_conv1_stats_tc_f32 / b200_conv1_fwd_pool_mask_tc_f32, same contracts.)
//
// Four horizontally adjacent output pixels (x = 4r .. 4r+3) of two image rows are one M row of the MMA:
//     D[r][(o, p, ch)] = sum over input rows d = 0..5 and window positions j = 0..11 of
//                        in[y0 - 2 + d][4r - 4 + j] * w[ch][d - o][j - p - 2]          (zero outside the 5x5 taps)
// K = 6 rows x 12 window positions = 72 = nine kind::tf32 MMAs of K = 8, N = 2 x 4 x 16 = 128, M = 128.
// The point of this arrangement is the A operand.  In the no-swizzle K-major layout an M row of one
// K chunk is 16 contiguous bytes and consecutive rows are 16 bytes apart - and the window of pixel
// group r+1 starts exactly 16 bytes after the window of group r.  So an input row, lying in shared memory
// as TMA delivered it, IS the operand: chunk c of input row d starts at row_base + 16 c, the three chunks
// overlap, no thread ever touches the input (the vertical-Toeplitz bf16 hi/lo kernel in conv1_tc.cu spends
// its time transposing the tile and issuing 15 full-rate MMAs per 1024 pixels; this one issues 9 at N = 128).
// The tensor core truncates the fp32 input to tf32; weights are rounded to nearest.
//
// Rows: one item = 2 output rows x 504 pixels (126 pixel groups; M rows 126/127 read their windows from the
// next row's bytes and are ignored - a 64-bit-element tensor map gives a 2048-byte box row = 512 pixels,
// 4 halo pixels left and 4 right).  A CTA walks DOWN a 504-pixel column strip, so each item needs only two
// new input rows: the rows live in a ring of 2-row slots, and because every MMA addresses either one row or
// the even/odd pair (d, d+1), the six rows of an item may sit in any three slots.
// Roles (608 threads): warps 0 and 18 issue the MMAs of alternate items (the issuing thread's waits + descriptor
// arithmetic per item are longer than the item's 576 tensor-pipe cycles; every tcgen05 instruction of a thread queues
// behind its MMAs, so there is ONE commit per item and the epilogue hands the ring slots back); warp 1 is the TMA
// producer; 16 epilogue warps = two groups alternating items, two warps per TMEM lane quarter, each thread one 2x2
// pooling window x 16 channels (both rows of the window are its own accumulator columns: no shuffles).
// Pass 1 (statistics): sums of conv, bias added analytically at the end.  Pass 2: the weights are multiplied by
// sign(BN scale), so the accumulator is sign*conv exactly and the BN map |scale|*v + shift' is monotone: the window
// maximum and the arg-max byte are taken on raw accumulators, one FMA + ReLU on the winner.  The pooled tile is staged
// in shared memory (128-byte swizzle, bank-conflict-free) and leaves through one TMA tensor store per item.
// Measured at batch 5, 3000x3000: pass 1 0.149 ms, pass 2 0.29 ms (conv1_tc.cu kernels: 0.238 / 0.476 ms); pass 2 is
// bound by the ALU pipe (FMNMX / FSETP / SEL of the arg-max) in the epilogue, pass 1 by the MMA issue.
#include <cuda.h>
#include <cuda_runtime.h>
#include <mutex>

#include "common.h"
#include "ptx.cuh"
#include <cstdint>
#include <cstdlib>

using namespace b2;

namespace {
namespace hz {
constexpr int kGroups = 126;                  // valid M rows (pixel groups of 4) per item
constexpr int kTilePx = 4 * kGroups;          // 504
constexpr int kRowBytes = 2048;               // one input row of the box: pixels x0-4 .. x0+507
constexpr int kSlotBytes = 2 * kRowBytes;     // a row pair
constexpr int kSlots = 8;
constexpr int kMmas = 9;
constexpr int kWBytes = kMmas * 4096;         // [mma][k chunk 2][n 128][16 B]
constexpr int kTStages = 4;
constexpr int kEpiWarps = 16;
constexpr int kThreads = 32 * (3 + kEpiWarps);   // warp 0 MMA (even items), warp 1 TMA, 16 epilogue warps, warp 18 MMA (odd items)
constexpr int kOutBytes = 128 * 128;          // pooled tile of one item: 126 rows of 2 pooled pixels x 16 channels (+2 unused)
constexpr int kCtlBytes = 4096;               // barriers, BN constants, per-warp statistics
constexpr int kSmemStats = kSlots * kSlotBytes + kWBytes + kCtlBytes + 1024;
constexpr int kSmemPool = kSmemStats + 4 * kOutBytes;       // two epilogue groups x two staging tiles
}  // namespace hz

struct HzGeom {
  int B, H, W;
  int nyp, nxt, nitems;
};

__device__ __forceinline__ void tma_load_3d_hz(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void st_global_v4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

__device__ __forceinline__ void named_bar_sync_hz(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
// shared -> global tensor store (bulk async group): the TMA engine un-swizzles the staged tile and clips it at the image edge
__device__ __forceinline__ void tma_store_3d(const void* tmap, const void* smem_src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(tmap), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void st_shared_v4(uint32_t saddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(saddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

template <bool POOL>
__global__ void __launch_bounds__(hz::kThreads, 1)
conv1_hz_kernel(const __grid_constant__ CUtensorMap tm_in, const __grid_constant__ CUtensorMap tm_out,
                const float* __restrict__ w1, const float* __restrict__ b1,
                const float* __restrict__ bn, float* __restrict__ pooled, uint8_t* __restrict__ amask,
                float* __restrict__ partials, HzGeom g) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* s_ring = smem;                                   // [kSlots][2 rows][2048 B]
  uint8_t* s_w = smem + hz::kSlots * hz::kSlotBytes;        // Toeplitz weights
  uint8_t* s_out = s_w + hz::kWBytes + hz::kCtlBytes;                // POOL: [group 2][buffer 2][128 rows][128 B], 128-byte swizzle
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_w + hz::kWBytes);
  uint64_t* rfull = bars;                       // [kSlots]
  uint64_t* rempty = rfull + hz::kSlots;        // [kSlots]
  uint64_t* tfull = rempty + hz::kSlots;        // [kTStages]
  uint64_t* tempty = tfull + hz::kTStages;      // [kTStages]
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(tempty + hz::kTStages);
  float* s_bn = reinterpret_cast<float*>(s_tmem + 4);       // [16][|scale|, shift'] (pass 2)  /  [bias 16] (pass 1)
  float* s_stat = s_bn + 32;                                // [16 warps][32]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  for (int idx = threadIdx.x; idx < hz::kMmas * 2 * 128 * 4; idx += hz::kThreads) {
    const int kk = idx & 3, n = (idx >> 2) & 127, h = (idx >> 9) & 1, i = idx >> 10;
    const int d = i < 6 ? i : 2 * (i - 6) + h, c = i < 6 ? h : 2;
    const int o = n >> 6, p = (n >> 4) & 3, ch = n & 15;
    const int dy = d - o, dx = 4 * c + kk - p - 2;
    float v = (dy >= 0 && dy <= 4 && dx >= 0 && dx <= 4) ? w1[ch * 25 + dy * 5 + dx] : 0.f;
    // pass 2 convolves with sign(BN scale) * w: the accumulator is then sign * conv EXACTLY (sign-symmetric arithmetic),
    // the scale applied to it is |scale| >= 0, and max-pool commutes with the (monotone) BN affine map - the epilogue
    // compares raw accumulators and applies one FMA to the winner instead of four
    if (POOL && bn[ch] < 0.f) v = -v;
    reinterpret_cast<float*>(s_w)[idx] = to_tf32_rna(v);
  }
  if (threadIdx.x < 16) {
    if (POOL) {
      // BN of (conv + bias): the bias moves into the shift
      const float sc = bn[threadIdx.x];
      s_bn[2 * threadIdx.x] = fabsf(sc);
      s_bn[2 * threadIdx.x + 1] = fmaf(b1[threadIdx.x], sc, bn[16 + threadIdx.x]);
    } else {
      s_bn[threadIdx.x] = b1[threadIdx.x];
    }
  }
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    for (int i = 0; i < hz::kSlots; ++i) { mbar_init(&rfull[i], 1); mbar_init(&rempty[i], 1); }
    for (int i = 0; i < hz::kTStages; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 256); }
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc<512>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  const int begin = (int)(((long long)blockIdx.x * g.nitems) / gridDim.x);
  const int end = (int)(((long long)(blockIdx.x + 1) * g.nitems) / gridDim.x);

  if (warp == 0 || warp == 18) {
    // ================= MMA issue: two warps, alternate items (the issuing thread's scalar work per item - waits, slot
    // arithmetic, nine descriptors - is longer than the 576 tensor-pipe cycles of the item's MMAs) =================
    const uint32_t mine = warp == 0 ? 0u : 1u;
    constexpr uint32_t idesc = umma_idesc(2, 128, 128, 0, 0);
    const uint64_t a_row = smem_desc_noswz(0, 16, 128);                  // chunks c, c+1 of one input row
    const uint64_t a_pair = smem_desc_noswz(0, hz::kRowBytes, 128);      // chunk 2 of rows d, d+1
    const uint64_t b_tmpl = smem_desc_noswz(smem_u32(s_w), 2048, 128);
    const uint32_t ring16 = smem_u32(s_ring) >> 4;
    // The issuing thread is the kernel's critical path: every mbarrier wait costs it ~100 cycles even when the phase
    // has long completed, and the tensor pipe only covers what is already queued.  So: no divisions, three slot bases
    // per item, and the waits of the warp's NEXT item sit between the fifth and sixth MMA of the current one, where
    // the queue is full anyway.  A row pair this warp waited for two items ago is not waited for again.
    auto step = [&](uint32_t& a_, int& yp_) {
      const bool last_ = yp_ + 1 == g.nyp;
      a_ += last_ ? 3u : 1u;
      yp_ = last_ ? 0 : yp_ + 1;
    };
    auto wait_item = [&](uint32_t a_, uint32_t q_, int yp_, int item_) {
      if (yp_ < 2 || item_ - begin < 2) mbar_wait(&rfull[a_ % hz::kSlots], (a_ / hz::kSlots) & 1);
      mbar_wait(&rfull[(a_ + 1) % hz::kSlots], ((a_ + 1) / hz::kSlots) & 1);
      mbar_wait(&rfull[(a_ + 2) % hz::kSlots], ((a_ + 2) / hz::kSlots) & 1);
      mbar_wait(&tempty[q_ % hz::kTStages], ((q_ / hz::kTStages) & 1) ^ 1);
      tc_fence_after();
    };
    uint32_t a = 0, q = mine;
    int yp = begin % g.nyp, item = begin + (int)mine;
    if (mine) step(a, yp);
    if (item < end) wait_item(a, q, yp, item);
    for (; item < end; item += 2, q += 2) {
      const uint32_t ts = q % hz::kTStages;
      const uint32_t dcol = tmem_base + ts * 128;
      const uint64_t sb0 = ring16 + (a % hz::kSlots) * (hz::kSlotBytes >> 4);
      const uint64_t sb1 = ring16 + ((a + 1) % hz::kSlots) * (hz::kSlotBytes >> 4);
      const uint64_t sb2 = ring16 + ((a + 2) % hz::kSlots) * (hz::kSlotBytes >> 4);
      constexpr uint64_t kRow = hz::kRowBytes >> 4, kW = 4096 >> 4;
      if (elect_one()) {
        umma_tf32(dcol, a_row + sb0, b_tmpl, idesc, 0u);
        umma_tf32(dcol, a_row + sb0 + kRow, b_tmpl + kW, idesc, 1u);
        umma_tf32(dcol, a_row + sb1, b_tmpl + 2 * kW, idesc, 1u);
        umma_tf32(dcol, a_row + sb1 + kRow, b_tmpl + 3 * kW, idesc, 1u);
        umma_tf32(dcol, a_row + sb2, b_tmpl + 4 * kW, idesc, 1u);
      }
      __syncwarp();
      uint32_t a2 = a;
      int yp2 = yp;
      step(a2, yp2);
      step(a2, yp2);
      if (item + 2 < end) wait_item(a2, q + 2, yp2, item + 2);
      if (elect_one()) {
        umma_tf32(dcol, a_row + sb2 + kRow, b_tmpl + 5 * kW, idesc, 1u);
        umma_tf32(dcol, a_pair + sb0 + 2, b_tmpl + 6 * kW, idesc, 1u);
        umma_tf32(dcol, a_pair + sb1 + 2, b_tmpl + 7 * kW, idesc, 1u);
        umma_tf32(dcol, a_pair + sb2 + 2, b_tmpl + 8 * kW, idesc, 1u);
        umma_commit(&tfull[ts]);     // the only commit: the epilogue hands the ring slots back (every tcgen05 instruction
      }                              // of this thread queues behind the MMAs)
      __syncwarp();
      a = a2;
      yp = yp2;
    }
  } else if (warp == 1) {
    // ================= TMA producer: two new input rows per item (six at the top of a column strip) =================
    if (lane == 0) {
      tma_prefetch_desc(&tm_in);
      uint32_t L = 0;
      int yp = begin % g.nyp, xt = (begin / g.nyp) % g.nxt, b = begin / (g.nyp * g.nxt);
      bool start = true;
      for (int item = begin; item < end; ++item) {
        const int c0 = (xt * hz::kTilePx - 4) >> 1;           // 8-byte elements
        for (int pr = start ? 0 : 2; pr < 3; ++pr, ++L) {
          const uint32_t s = L % hz::kSlots;
          mbar_wait(&rempty[s], ((L / hz::kSlots) & 1) ^ 1);
          mbar_expect_tx(&rfull[s], hz::kSlotBytes);
          tma_load_3d_hz(s_ring + s * hz::kSlotBytes, &tm_in, &rfull[s], c0, 2 * yp - 2 + 2 * pr, b);
        }
        start = false;
        if (++yp == g.nyp) { yp = 0; start = true; if (++xt == g.nxt) { xt = 0; ++b; } }
      }
    }
  } else {
    // ================= epilogue: 16 warps = two groups alternating items; in a group two warps share a TMEM lane
    // quarter and take one 2x2 pooling window (pixels 2h, 2h+1 of both rows = 64 accumulator columns) each: the
    // per-item work is a long dependent chain, four warps per scheduler hide it =================
    const int ew = warp - 2;
    const int quarter = warp & 3, half = (ew >> 2) & 1, grp = ew >> 3;
    const int m = quarter * 32 + lane;
    const bool leader = (ew & 7) == 0 && lane == 0;
    float ssum[16], ssq[16], cnt = 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c) { ssum[c] = 0.f; ssq[c] = 0.f; }
    uint32_t q = 0, a = 0;
    int yp = begin % g.nyp - 1, xt = (begin / g.nyp) % g.nxt, b = begin / (g.nyp * g.nxt);
    bool last = false;
    for (int item = begin; item < end; ++item, ++q) {
      if (item != begin) a += last ? 3u : 1u;
      if (++yp == g.nyp) { yp = 0; if (++xt == g.nxt) { xt = 0; ++b; } }
      last = item + 1 == end || yp + 1 == g.nyp;
      if ((int)(q & 1) != grp) continue;
      if (POOL && leader) bulk_wait_read0();                  // the group's previous tile has left shared memory
      const int x = xt * hz::kTilePx + 4 * m;
      const bool ok = m < hz::kGroups && x < g.W;             // W % 4 == 0: all four pixels or none
      const uint32_t ts = q % hz::kTStages;
      mbar_wait(&tfull[ts], (q / hz::kTStages) & 1);
      tc_fence_after();
      if (leader) {
        // the item's MMAs have completed: its top row pair (all three at the bottom of a strip) is free for the producer
        mbar_arrive(&rempty[a % hz::kSlots]);
        if (last) { mbar_arrive(&rempty[(a + 1) % hz::kSlots]); mbar_arrive(&rempty[(a + 2) % hz::kSlots]); }
      }
      // columns: [row o][pixel p][channel]; this thread: pixels 2 half, 2 half + 1 of row 0 (v[0..31]) and row 1 (v[32..63])
      const uint32_t tbase = tmem_base + ((uint32_t)(quarter * 32) << 16) + ts * 128 + half * 32;
      if constexpr (!POOL) {
        uint32_t v[32];
#pragma unroll
        for (int o = 0; o < 2; ++o) {
          tmem_ld16(tbase + o * 64, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
          tmem_ld16(tbase + o * 64 + 16, *reinterpret_cast<uint32_t(*)[16]>(&v[16]));
          tmem_ld_wait();
          if (o == 1) { tc_fence_before(); mbar_arrive(&tempty[ts]); }
          if (ok) {
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const float f = __uint_as_float(v[e]);
              ssum[e & 15] += f;
              ssq[e & 15] = fmaf(f, f, ssq[e & 15]);
            }
          }
        }
        if (ok) cnt += 4.f;
      } else {
        // pooled values: staged in shared memory (row = the 128 bytes of the thread pair that owns pixel group m, 16-byte
        // chunks XOR-swizzled so the 32 lanes of a store hit 32 different banks) and written by ONE tensor store per item -
        // as direct stores every lane touched its own line (ncu: 74 % of the L1 wavefront peak).
        uint32_t v[64];
        tmem_ld16(tbase, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
        tmem_ld16(tbase + 16, *reinterpret_cast<uint32_t(*)[16]>(&v[16]));
        tmem_ld16(tbase + 64, *reinterpret_cast<uint32_t(*)[16]>(&v[32]));
        tmem_ld16(tbase + 80, *reinterpret_cast<uint32_t(*)[16]>(&v[48]));
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(&tempty[ts]);
        const uint32_t buf = (uint32_t)(grp * 2 + ((q >> 1) & 1)) * hz::kOutBytes;
        if (ok) {
          uint32_t pm[16], mb[4] = {0u, 0u, 0u, 0u};
#pragma unroll
          for (int c = 0; c < 16; ++c) {
            const float2 ss = reinterpret_cast<const float2*>(s_bn)[c];
            const float v00 = __uint_as_float(v[c]), v01 = __uint_as_float(v[16 + c]);
            const float v10 = __uint_as_float(v[32 + c]), v11 = __uint_as_float(v[48 + c]);
            // first maximum in row-major order (ATen's max_pool2d): a later candidate wins only if strictly greater
            const float r0 = fmaxf(v00, v01), r1 = fmaxf(v10, v11);
            const uint32_t am = r1 > r0 ? (v11 > v10 ? 3u : 2u) : (v01 > v00 ? 1u : 0u);
            const float best = fmaf(fmaxf(r0, r1), ss.x, ss.y);
            pm[c] = __float_as_uint(fmaxf(best, 0.f));
            mb[c >> 2] |= (am | (best > 0.f ? 4u : 0u)) << (8 * (c & 3));
          }
          const uint32_t row = smem_u32(s_out) + buf + (uint32_t)m * 128u;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            st_shared_v4(row + (uint32_t)(((half * 4 + k) ^ (m & 7)) << 4), pm[4 * k], pm[4 * k + 1], pm[4 * k + 2], pm[4 * k + 3]);
          const size_t pix = ((size_t)b * g.nyp + yp) * (g.W >> 1) + (x >> 1) + half;
          st_global_v4(amask + pix * 16, mb[0], mb[1], mb[2], mb[3]);
        }
        fence_proxy_async_smem();
        named_bar_sync_hz(1 + grp, 256);
        if (leader) {
          tma_store_3d(&tm_out, reinterpret_cast<const void*>(s_out + buf), 0, xt * hz::kGroups, b * g.nyp + yp);
          bulk_commit();
        }
      }
    }
    if (POOL && leader) bulk_wait0();
    if constexpr (!POOL) {
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        // statistics of conv + bias from those of conv
        const float bc = s_bn[c];
        const float sb = fmaf(cnt, bc, ssum[c]);
        const float qb = fmaf(bc, fmaf(cnt, bc, 2.f * ssum[c]), ssq[c]);
        const float a1 = warp_sum(sb), a2 = warp_sum(qb);
        if (lane == 0) { s_stat[ew * 32 + c] = a1; s_stat[ew * 32 + 16 + c] = a2; }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (!POOL && threadIdx.x < 32) {
    float acc = 0.f;
#pragma unroll
    for (int w8 = 0; w8 < hz::kEpiWarps; ++w8) acc += s_stat[w8 * 32 + threadIdx.x];
    partials[(size_t)blockIdx.x * 32 + threadIdx.x] = acc;
  }
  if (warp == 0) { __syncwarp(); tmem_dealloc<512>(tmem_base); }
}

typedef CUresult (*EncodeTiledFnHz)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFnHz encode_fn_hz() {
  static EncodeTiledFnHz fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFnHz)p;
  });
  return fn;
}

template <bool POOL>
int conv1_hz_launch(const void* in_f32, const void* w1, const void* b1, const void* bn, void* pooled, void* amask,
                    void* partials, int B, int H, int W, int* nparts_out, cudaStream_t s) {
  B200_CHECK(H % 2 == 0 && W % 4 == 0, "conv1 (tf32 row-operand kernel): H must be even and W a multiple of 4");
  B200_CHECK((reinterpret_cast<uintptr_t>(in_f32) & 15) == 0, "conv1 (tf32 row-operand kernel): input must be 16-byte aligned");
  EncodeTiledFnHz enc = encode_fn_hz();
  if (!enc) return b200_fail(-3, "cuTensorMapEncodeTiled entry point not available");
  // the image as rows of 8-byte elements: a 256-element box row is 512 pixels (the box limit is 256 ELEMENTS)
  CUtensorMap tm;
  const cuuint64_t dims[3] = {(cuuint64_t)W / 2, (cuuint64_t)H, (cuuint64_t)B};
  const cuuint64_t strides[2] = {(cuuint64_t)W * 4, (cuuint64_t)H * W * 4};
  const cuuint32_t box[3] = {256, 2, 1}, estr[3] = {1, 1, 1};
  CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<void*>(in_f32), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return b200_fail((int)r, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  HzGeom g;
  g.B = B; g.H = H; g.W = W;
  g.nyp = H / 2;
  g.nxt = (W + hz::kTilePx - 1) / hz::kTilePx;
  g.nitems = B * g.nyp * g.nxt;
  // pooled output as rows of 128 bytes (two pooled pixels x 16 channels): the store box is one item's 126 rows
  CUtensorMap tmo = tm;
  if (POOL) {
    B200_CHECK((reinterpret_cast<uintptr_t>(pooled) & 15) == 0, "conv1 (tf32 row-operand kernel): p1 must be 16-byte aligned");
    const cuuint64_t od[3] = {32, (cuuint64_t)W / 4, (cuuint64_t)B * g.nyp};
    const cuuint64_t os[2] = {128, (cuuint64_t)(W / 4) * 128};
    const cuuint32_t ob[3] = {32, (cuuint32_t)hz::kGroups, 1};
    r = enc(&tmo, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, pooled, od, os, ob, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return b200_fail((int)r, "cuTensorMapEncodeTiled (p1) failed with CUresult %d", (int)r);
  }
  const int nsm = b200_compute_sms();
  const int grid = g.nitems < nsm ? g.nitems : nsm;
  constexpr int smem_bytes = POOL ? hz::kSmemPool : hz::kSmemStats;
  B200_CUDA(set_max_smem(conv1_hz_kernel<POOL>, smem_bytes));
  conv1_hz_kernel<POOL><<<grid, hz::kThreads, smem_bytes, s>>>(tm, tmo, (const float*)w1, (const float*)b1,
                                                              (const float*)bn, (float*)pooled, (uint8_t*)amask,
                                                              (float*)partials, g);
  B200_LAUNCH_CHECK();
  if (nparts_out) *nparts_out = grid;
  return 0;
}

}  // namespace

// Pass 1: BatchNorm batch statistics of conv1(in) + bias; partials [b200_sm_count()][32] = per-CTA [sum 16 | sumsq 16].
B200_API int b200_conv1_stats_tf32(const void* in_f32, const void* w1, const void* b1, void* partials, int B, int H,
                                   int W, int* nparts_out, void* stream) {
  return conv1_hz_launch<false>(in_f32, w1, b1, nullptr, nullptr, nullptr, partials, B, H, W, nparts_out,
                                (cudaStream_t)stream);
}
// Pass 2: conv again (same MMAs, same order: bit-identical values) -> BN (bn = [scale 16 | shift 16 | ...]) + ReLU +
// 2x2 max-pool -> p1 [B][H/2][W/2][16] fp32 and the arg-max | ReLU << 2 bytes of b200_conv1_wgrad_nox1_f32.
B200_API int b200_conv1_fwd_pool_mask_tf32(const void* in_f32, const void* w1, const void* b1, const void* bn,
                                           void* p1_f32, void* amask_u8, int B, int H, int W, void* stream) {
  B200_CHECK(bn != nullptr && p1_f32 != nullptr && amask_u8 != nullptr, "conv1_fwd_pool_mask_tf32: bn, p1 and the mask are required");
  return conv1_hz_launch<true>(in_f32, w1, b1, bn, p1_f32, amask_u8, nullptr, B, H, W, nullptr, (cudaStream_t)stream);
}
