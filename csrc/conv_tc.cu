// tcgen05 / TMEM / TMA implicit-GEMM kernels for the 5x5, pad-2, stride-1 conv2
// of the reference ConvNet (mnist_distributed.py:33-37: Conv2d(16, 32, 5, 1, 2))
// at the 1500x1500 post-pool resolution: forward (K6), data gradient (K12a)
// and weight gradient (K12b) of SURVEY.md 2.5.  The reference runs these as
// cuDNN TF32 NCHW library kernels; here they are written for sm_100a.
//
// Design (B200-first, not an im2col GEMM):
//  * Activations are bf16 NHWC.  A pixel's channel vector is split into 16-byte
//    "planes" of 8 channels, and TMA (5-D tensor map {8ch, plane, W, H, B})
//    drops one image row segment per plane into shared memory at a 16-byte
//    pixel pitch.  With the NO-SWIZZLE K-major UMMA layout (row r of the A
//    tile at r*16 B, the two 8-channel K-chunks one plane apart) a 128-pixel A
//    operand for filter tap (dy,dx) is just "row slot dy, start + dx*16 B":
//    the 5x5 halo is re-used straight from shared memory, every input row is
//    fetched from L2 once per strip (a ROLLING ring of row slots), and the
//    conv zero padding is TMA out-of-bounds fill.  No im2col tile is ever
//    materialised, in HBM or in smem.
//  * One CTA per SM, persistent over (image, 128-px column strip, row segment)
//    work items.  Warp 0 = TMA producer, warp 1 = single-thread tcgen05.mma
//    issuer (25 taps x CIN/16 MMAs of M=128,N=COUT,K=16 per output row),
//    warps 2-5 = epilogue: tcgen05.ld the fp32 accumulator row out of TMEM
//    (4-deep TMEM ring so MMA of row y+1..y+3 overlaps the epilogue of row y),
//    add bias, round to bf16, store NHWC and - for the forward - accumulate the
//    BatchNorm batch statistics (sum, sum of squares per channel) in registers
//    (K4/K7 fused into the producing kernel).
//  * wgrad contracts over PIXELS, so both operands are MN-major straight out of
//    the same planar row slots: A[m=(dx,ci8), k=pixel] starts at "slot dy,
//    plane, pixel k0" with a 16-byte M-stride (the dx shift IS the next core
//    matrix), B[n=co, k=pixel] is the dx2 row.  The ten (dy, plane) 64x32 fp32
//    accumulators live in TMEM for the whole kernel and are written once.
#include "common.h"
#include "ptx.cuh"
#include <cuda.h>
#include <cstdint>
#include <mutex>
#include <type_traits>

namespace {

using namespace b2;

constexpr int kBM = 128;          // output pixels per row tile (UMMA M)
constexpr int kPW = 136;          // pixels per smem plane row: 128 + 4 halo, padded to 8
constexpr int kPlaneBytes = kPW * 16;
constexpr int kTaps = 25;
constexpr int kNumTmemStages = 4;
constexpr int kThreads = 192;     // warp0 TMA, warp1 MMA, warps2-5 epilogue

template <int CIN, int COUT>
struct ConvCfg {
  static constexpr int kPlanes = CIN / 8;
  static constexpr int kSlotBytes = kPlanes * kPlaneBytes;
  static constexpr int kSlots = 12;
  static constexpr int kKSteps = CIN / 16;
  static constexpr int kWTapBytes = kKSteps * 2 * COUT * 16;
  static constexpr int kWBytes = kTaps * kWTapBytes;
  static constexpr int kTmemCols = (kNumTmemStages * COUT) < 32 ? 32 : (kNumTmemStages * COUT);
  static constexpr int kSmemBytes = kWBytes + kSlots * kSlotBytes + 2048 + 128;
  static_assert(kWBytes % 128 == 0 && kSlotBytes % 128 == 0, "TMA destinations must stay 128-B aligned");
  static_assert((kTmemCols & (kTmemCols - 1)) == 0, "TMEM allocation must be a power of two");
};

struct ConvGeom {
  int B, H, W;
  int nstrips;    // ceil(W / 128)
  int rseg;       // output rows per work item
  int nseg;       // ceil(H / rseg)
  int nitems;
};

// ============================================================ fwd / dgrad ===
// out[b,y,x,n] = bias[n] + sum_{dy,dx,k} act[b,y+dy-2,x+dx-2,k] * wsrc(n,k,dy,dx)
//   TRANSPOSED=false: wsrc = w[n][k][dy][dx]            (forward,  w is [COUT][CIN][5][5])
//   TRANSPOSED=true : wsrc = w[k][n][4-dy][4-dx]        (dgrad,    w is [CIN ][COUT][5][5])
template <int CIN, int COUT, bool TRANSPOSED, bool STATS>
__global__ void __launch_bounds__(kThreads, 1)
conv5x5_tc_kernel(const __grid_constant__ CUtensorMap tm_act, const float* __restrict__ w,
                  const float* __restrict__ bias, __nv_bfloat16* __restrict__ out,
                  float* __restrict__ partials, ConvGeom g) {
  using Cfg = ConvCfg<CIN, COUT>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);   // TMA destinations: 128-B aligned
  uint8_t* s_w = smem;
  uint8_t* s_slots = smem + Cfg::kWBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_slots + Cfg::kSlots * Cfg::kSlotBytes);
  uint64_t* full = bars;                       // [kSlots]   TMA -> MMA
  uint64_t* empty = bars + Cfg::kSlots;        // [kSlots]   MMA -> TMA
  uint64_t* tfull = empty + Cfg::kSlots;       // [4]        MMA -> epilogue
  uint64_t* tempty = tfull + kNumTmemStages;   // [4]        epilogue -> MMA
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(tempty + kNumTmemStages);
  float* s_bias = reinterpret_cast<float*>(s_tmem + 4);           // [COUT]
  float* s_stat = s_bias + COUT;                                  // [4][2*COUT]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- prologue: pack weights (fp32 -> bf16, UMMA K-major no-swizzle B tiles)
  // byte(tap, kk, c, n, e) = ((tap*KSTEPS + kk)*2 + c)*COUT*16 + n*16 + e*2,  k = kk*16 + c*8 + e
  for (int i = threadIdx.x; i < kTaps * CIN * COUT; i += kThreads) {
    const int e = i & 7;
    int r = i >> 3;
    const int n = r % COUT; r /= COUT;
    const int c = r & 1; r >>= 1;
    const int kk = r % Cfg::kKSteps;
    const int tap = r / Cfg::kKSteps;
    const int k = kk * 16 + c * 8 + e;
    const float v = TRANSPOSED ? w[((size_t)k * COUT + n) * kTaps + (kTaps - 1 - tap)]
                               : w[((size_t)n * CIN + k) * kTaps + tap];
    reinterpret_cast<__nv_bfloat16*>(s_w)[i] = __float2bfloat16_rn(v);
  }
  if (threadIdx.x < COUT) s_bias[threadIdx.x] = bias ? bias[threadIdx.x] : 0.f;
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    for (int i = 0; i < Cfg::kSlots; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < kNumTmemStages; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 128); }
    mbar_fence_init();
    tma_prefetch_desc(&tm_act);
  }
  if (warp == 1) tmem_alloc<Cfg::kTmemCols>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ===================== TMA producer (one lane) =====================
    if (lane == 0) {
      uint32_t gcount = 0;
      for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
        const int seg = item % g.nseg;
        const int strip = (item / g.nseg) % g.nstrips;
        const int b = item / (g.nseg * g.nstrips);
        const int y0 = seg * g.rseg;
        const int rows = min(g.rseg, g.H - y0);
        const int x0 = strip * kBM;
        for (int i = 0; i < rows + 4; ++i, ++gcount) {
          const uint32_t slot = gcount % Cfg::kSlots, use = gcount / Cfg::kSlots;
          mbar_wait(&empty[slot], (use & 1) ^ 1);
          mbar_expect_tx(&full[slot], Cfg::kSlotBytes);
          uint8_t* dst = s_slots + slot * Cfg::kSlotBytes;
#pragma unroll
          for (int p = 0; p < Cfg::kPlanes; ++p)
            tma_load_5d(dst + p * kPlaneBytes, &tm_act, &full[slot], 0, p, x0 - 2, y0 - 2 + i, b);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one lane) =====================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(kBM, COUT, 0, 0);
      const uint32_t slots_addr = smem_u32(s_slots);
      // descriptor templates: A: LBO = plane stride, SBO = 128 B;  B: LBO = COUT*16, SBO = 128 B
      const uint64_t a_tmpl = smem_desc_noswz(0, kPlaneBytes, 128);
      const uint64_t b_tmpl = smem_desc_noswz(smem_u32(s_w), COUT * 16, 128);
      uint32_t gbase = 0, q = 0;
      for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
        const int seg = item % g.nseg;
        const int y0 = seg * g.rseg;
        const int rows = min(g.rseg, g.H - y0);
        for (int j = 0; j < rows; ++j, ++q) {
          // input rows j..j+4 (local) must have landed; rows arrive in order
          for (int i = (j == 0 ? 0 : 4); i < 5; ++i) {
            const uint32_t gc = gbase + j + i;
            mbar_wait(&full[gc % Cfg::kSlots], (gc / Cfg::kSlots) & 1);
          }
          const uint32_t stage = q % kNumTmemStages, use = q / kNumTmemStages;
          mbar_wait(&tempty[stage], (use & 1) ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + stage * COUT;
          uint32_t accum = 0;
#pragma unroll
          for (int dy = 0; dy < 5; ++dy) {
            const uint32_t slot = (gbase + j + dy) % Cfg::kSlots;
            const uint32_t row_addr = slots_addr + slot * Cfg::kSlotBytes;
#pragma unroll
            for (int dx = 0; dx < 5; ++dx) {
#pragma unroll
              for (int kk = 0; kk < Cfg::kKSteps; ++kk) {
                const uint32_t a_addr = row_addr + (2 * kk) * kPlaneBytes + dx * 16;
                const uint64_t a_desc = a_tmpl | (uint64_t)((a_addr >> 4) & 0x3FFF);
                const uint64_t b_desc = b_tmpl + (uint64_t)((((dy * 5 + dx) * Cfg::kKSteps + kk) * 2 * COUT * 16) >> 4);
                umma_bf16(d_tmem, a_desc, b_desc, idesc, accum);
                accum = 1;
              }
            }
          }
          umma_commit(&tfull[stage]);                               // accumulator row ready
          umma_commit(&empty[(gbase + j) % Cfg::kSlots]);           // oldest input row is dead
        }
        for (int i = 0; i < 4; ++i) umma_commit(&empty[(gbase + rows + i) % Cfg::kSlots]);
        gbase += rows + 4;
      }
    }
  } else {
    // ===================== epilogue warps (2..5) =====================
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may read
    const int m = quarter * 32 + lane;            // pixel within the row tile
    float ssum[STATS ? COUT : 1], ssq[STATS ? COUT : 1];
    if (STATS) {
#pragma unroll
      for (int c = 0; c < COUT; ++c) { ssum[c] = 0.f; ssq[c] = 0.f; }
    }
    uint32_t q = 0;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
      const int seg = item % g.nseg;
      const int strip = (item / g.nseg) % g.nstrips;
      const int b = item / (g.nseg * g.nstrips);
      const int y0 = seg * g.rseg;
      const int rows = min(g.rseg, g.H - y0);
      const int x = strip * kBM + m;
      const bool valid = x < g.W;
      for (int j = 0; j < rows; ++j, ++q) {
        const uint32_t stage = q % kNumTmemStages, use = q / kNumTmemStages;
        mbar_wait(&tfull[stage], use & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + stage * COUT;
        uint32_t v[COUT];
#pragma unroll
        for (int c0 = 0; c0 < COUT; c0 += 16) tmem_ld16(taddr + c0, *reinterpret_cast<uint32_t(*)[16]>(&v[c0]));
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(&tempty[stage]);              // TMEM stage may be overwritten
        uint32_t pk[COUT / 2];
#pragma unroll
        for (int c = 0; c < COUT; c += 2) {
          pk[c / 2] = pack_bf16x2(__uint_as_float(v[c]) + s_bias[c], __uint_as_float(v[c + 1]) + s_bias[c + 1]);
          if (STATS && valid) {
            const float r0 = bf16_lo(pk[c / 2]), r1 = bf16_hi(pk[c / 2]);
            ssum[c] += r0; ssq[c] = fmaf(r0, r0, ssq[c]);
            ssum[c + 1] += r1; ssq[c + 1] = fmaf(r1, r1, ssq[c + 1]);
          }
        }
        if (valid) {
          uint4* o = reinterpret_cast<uint4*>(out + (((size_t)b * g.H + (y0 + j)) * g.W + x) * COUT);
#pragma unroll
          for (int i = 0; i < COUT / 8; ++i) o[i] = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
        }
      }
    }
    if (STATS) {
#pragma unroll
      for (int c = 0; c < COUT; ++c) {
        const float a = warp_sum(ssum[c]), s2 = warp_sum(ssq[c]);
        if (lane == 0) { s_stat[quarter * 2 * COUT + c] = a; s_stat[quarter * 2 * COUT + COUT + c] = s2; }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (STATS && threadIdx.x < 2 * COUT) {
    float a = 0.f;
#pragma unroll
    for (int w4 = 0; w4 < 4; ++w4) a += s_stat[w4 * 2 * COUT + threadIdx.x];
    partials[(size_t)blockIdx.x * 2 * COUT + threadIdx.x] = a;
  }
  if (warp == 1) { __syncwarp(); tmem_dealloc<Cfg::kTmemCols>(tmem_base); }
}

// ================================================================== wgrad ===
// dW[co][ci][dy][dx] = sum_{b,y,x} dxo[b,y,x,co] * act[b,y+dy-2,x+dx-2,ci]
// act: bf16 NHWC, 16 channels (2 planes);  dxo: bf16 NHWC, 32 channels (4 planes).
constexpr int kWgSlotsA = 10, kWgSlotsB = 4;
constexpr int kWgSlotABytes = 2 * kPlaneBytes;        // 4352
constexpr int kWgPlaneBBytes = kBM * 16;              // 2048
constexpr int kWgSlotBBytes = 4 * kWgPlaneBBytes;     // 8192
constexpr int kWgSmemBytes = kWgSlotsA * kWgSlotABytes + kWgSlotsB * kWgSlotBBytes + 1024 + 128;
constexpr int kWgTmemCols = 512;                      // 10 accumulators x 32 columns -> next pow2

__global__ void __launch_bounds__(kThreads, 1)
conv5x5_wgrad_tc_kernel(const __grid_constant__ CUtensorMap tm_act, const __grid_constant__ CUtensorMap tm_dxo,
                        float* __restrict__ partials, ConvGeom g) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  uint8_t* s_a = smem;
  uint8_t* s_b = smem + kWgSlotsA * kWgSlotABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_b + kWgSlotsB * kWgSlotBBytes);
  uint64_t* fullA = bars;
  uint64_t* emptyA = fullA + kWgSlotsA;
  uint64_t* fullB = emptyA + kWgSlotsA;
  uint64_t* emptyB = fullB + kWgSlotsB;
  uint64_t* done = emptyB + kWgSlotsB;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kWgSlotsA; ++i) { mbar_init(&fullA[i], 1); mbar_init(&emptyA[i], 1); }
    for (int i = 0; i < kWgSlotsB; ++i) { mbar_init(&fullB[i], 1); mbar_init(&emptyB[i], 1); }
    mbar_init(done, 1);
    mbar_fence_init();
    tma_prefetch_desc(&tm_act);
    tma_prefetch_desc(&tm_dxo);
  }
  if (warp == 1) tmem_alloc<kWgTmemCols>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  const bool has_work = (int)blockIdx.x < g.nitems;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t ga = 0, gb = 0;
      for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
        const int seg = item % g.nseg;
        const int strip = (item / g.nseg) % g.nstrips;
        const int b = item / (g.nseg * g.nstrips);
        const int y0 = seg * g.rseg;
        const int rows = min(g.rseg, g.H - y0);
        const int x0 = strip * kBM;
        for (int i = 0; i < rows + 4; ++i) {
          {
            const uint32_t slot = ga % kWgSlotsA, use = ga / kWgSlotsA;
            mbar_wait(&emptyA[slot], (use & 1) ^ 1);
            mbar_expect_tx(&fullA[slot], kWgSlotABytes);
            uint8_t* dst = s_a + slot * kWgSlotABytes;
            tma_load_5d(dst, &tm_act, &fullA[slot], 0, 0, x0 - 2, y0 - 2 + i, b);
            tma_load_5d(dst + kPlaneBytes, &tm_act, &fullA[slot], 0, 1, x0 - 2, y0 - 2 + i, b);
            ++ga;
          }
          // dxo row j is consumed together with act rows j..j+4: issue it after act row j+2
          if (i >= 2 && i - 2 < rows) {
            const uint32_t slot = gb % kWgSlotsB, use = gb / kWgSlotsB;
            mbar_wait(&emptyB[slot], (use & 1) ^ 1);
            mbar_expect_tx(&fullB[slot], kWgSlotBBytes);
            uint8_t* dst = s_b + slot * kWgSlotBBytes;
#pragma unroll
            for (int p = 0; p < 4; ++p)
              tma_load_5d(dst + p * kWgPlaneBBytes, &tm_dxo, &fullB[slot], 0, p, x0, y0 + i - 2, b);
            ++gb;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // A: M=64 (8 dx-shifts x 8 channels of one plane), MN-major: SBO = 16 B (next dx), LBO = 128 B (next 8 pixels)
      // B: N=32 (4 planes of 8 co),                      MN-major: SBO = plane stride,  LBO = 128 B
      constexpr uint32_t idesc = umma_idesc_bf16(64, 32, 1, 1);
      const uint64_t a_tmpl = smem_desc_noswz(0, 128, 16);
      const uint64_t b_tmpl = smem_desc_noswz(0, 128, kWgPlaneBBytes);
      const uint32_t a_base = smem_u32(s_a), b_base = smem_u32(s_b);
      uint32_t gbase = 0, gb = 0;
      uint32_t first = 1;
      for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
        const int seg = item % g.nseg;
        const int y0 = seg * g.rseg;
        const int rows = min(g.rseg, g.H - y0);
        for (int j = 0; j < rows; ++j, ++gb) {
          for (int i = (j == 0 ? 0 : 4); i < 5; ++i) {
            const uint32_t gc = gbase + j + i;
            mbar_wait(&fullA[gc % kWgSlotsA], (gc / kWgSlotsA) & 1);
          }
          const uint32_t slotB = gb % kWgSlotsB;
          mbar_wait(&fullB[slotB], (gb / kWgSlotsB) & 1);
          tc_fence_after();
          const uint32_t brow = b_base + slotB * kWgSlotBBytes;
#pragma unroll 1
          for (int kb = 0; kb < kBM / 16; ++kb) {
            const uint64_t b_desc = b_tmpl | (uint64_t)(((brow + kb * 256) >> 4) & 0x3FFF);
#pragma unroll
            for (int dy = 0; dy < 5; ++dy) {
              const uint32_t arow = a_base + ((gbase + j + dy) % kWgSlotsA) * kWgSlotABytes + kb * 256;
#pragma unroll
              for (int pl = 0; pl < 2; ++pl) {
                const uint64_t a_desc = a_tmpl | (uint64_t)(((arow + pl * kPlaneBytes) >> 4) & 0x3FFF);
                umma_bf16(tmem_base + (dy * 2 + pl) * 32, a_desc, b_desc, idesc, first ? 0u : 1u);
              }
            }
            first = 0;
          }
          umma_commit(&emptyB[slotB]);
          umma_commit(&emptyA[(gbase + j) % kWgSlotsA]);
        }
        for (int i = 0; i < 4; ++i) umma_commit(&emptyA[(gbase + rows + i) % kWgSlotsA]);
        gbase += rows + 4;
      }
      umma_commit(done);
    }
  } else {
    // epilogue: single pass at the very end.  M=64 accumulators occupy lanes
    // {0-15, 32-47, 64-79, 96-111}: row r -> lane (r%16) + 32*(r/16).
    const int quarter = warp & 3;
    float* dst = partials + (size_t)blockIdx.x * (32 * 16 * kTaps);
    if (has_work) {
      mbar_wait(done, 0);
      tc_fence_after();
    }
    const int r = quarter * 16 + (lane & 15);       // accumulator row for lanes 0..15
    const int dxs = r >> 3, e = r & 7;
    const bool row_valid = (lane < 16) && (dxs < 5);
#pragma unroll 1
    for (int acc = 0; acc < 10; ++acc) {
      const int dy = acc >> 1, pl = acc & 1;
      uint32_t v[32];
      if (has_work) {
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * 32;
        tmem_ld16(taddr, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
        tmem_ld16(taddr + 16, *reinterpret_cast<uint32_t(*)[16]>(&v[16]));
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int c = 0; c < 32; ++c) v[c] = 0u;
      }
      if (row_valid) {
        const int ci = pl * 8 + e;
#pragma unroll
        for (int co = 0; co < 32; ++co)
          dst[((size_t)co * 16 + ci) * kTaps + dy * 5 + dxs] = __uint_as_float(v[co]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) { __syncwarp(); tmem_dealloc<kWgTmemCols>(tmem_base); }
}

// ======================================================= v2: row-stationary ===
// Same operands, different loop nest.  v1 issues one M=128,N=COUT MMA per filter tap and
// output row (25 per row): every one of them re-reads its 4 KB A tile from shared memory for
// only 16 (N=32) cycles of math, so the kernel is bound by smem operand bandwidth and by the
// per-instruction cost, not by the tensor pipe.  v2 keeps the INPUT row stationary instead:
// for input row r and horizontal tap dx, ONE MMA with N = 5*COUT multiplies the row (shifted
// by dx) with the weights of all five vertical taps at once; column block jj = 4-dy of the
// result belongs to output row r-2+jj.  The accumulators of consecutive output rows sit in
// consecutive TMEM column blocks of a 16-deep ring, so that single MMA accumulates straight
// into five different output rows.  5 (x CIN/16) fat MMAs per row instead of 25 thin ones,
// each input row is needed only while its own MMAs run, and the epilogue is unchanged.
constexpr int kRsEpiWarps = 8;                       // two groups of 4 (one per TMEM lane quarter), alternating rows
constexpr int kRsThreads = 32 * (2 + kRsEpiWarps);  // warp0 feed, warp1 MMA issue, warps 2-9 epilogue

// ET = activation storage type: __nv_bfloat16 (kind::f16 MMAs, K = 16) or float (kind::tf32, K = 8: the
// reference's precision - fp32 tensors, TF32 convolution math).  A 16-byte plane holds kCPP channels.
template <int CIN, int COUT, typename ET = __nv_bfloat16>
struct RsCfg {
  static constexpr int kEB = (int)sizeof(ET);
  static constexpr int kCPP = 16 / kEB;                      // channels per 16-byte plane piece: 8 | 4
  static constexpr int kPlanes = CIN / kCPP;
  static constexpr int kSlotBytes = kPlanes * kPlaneBytes;
  // input-row ring = cp.async pipeline depth; power of two (fp32 dgrad rows are 17 KB: 8 of them)
  static constexpr int kSlots = (kEB == 4 && CIN == 32) ? 8 : 16;
  static constexpr int kKSteps = CIN / (2 * kCPP);           // MMAs per (row, dx): each spans two planes
  static constexpr int kNB = 5 * COUT;                       // B rows per (dx, kk): n = jj*COUT + co
  static constexpr int kWChunkBytes = kNB * 16;              // one kCPP-channel k-chunk of all n rows
  static constexpr int kWStepBytes = 2 * kWChunkBytes;       // one (dx, kk) B tile
  static constexpr int kWBytes = 5 * kKSteps * kWStepBytes;  // == 25*CIN*COUT*sizeof(ET)
  // The accumulator ring takes all 512 TMEM columns (16 slots at COUT=32, 32 at COUT=16).  A five-slot window that
  // reaches the end of the ring has to be split into two MMAs (a D range past column 511 faults), and each half
  // re-reads the whole A operand, so the longer ring of dgrad halves how often that happens.
  static constexpr int kTStages = 512 / COUT;
  static constexpr int kTmemCols = kTStages * COUT;          // 512
  static constexpr int kSmemBytes = kWBytes + kSlots * kSlotBytes + 4096 + 1024;
  static_assert(kWBytes % 128 == 0 && kSlotBytes % 128 == 0, "alignment");
  static_assert(kTmemCols <= 512 && (kTmemCols & (kTmemCols - 1)) == 0, "TMEM columns");
};

// Operand feed: LDGSTS=false uses TMA (one 5-D tensor-map request per plane and row; measured
// ~4-7 cycles per 16-byte box row, i.e. 1-3 k cycles per image row: the bottleneck of v2);
// LDGSTS=true has the whole producer warp issue 16-byte cp.async copies (zero-fill = padding) and
// then `cp.async.mbarrier.arrive.noinc` on the row's full barrier (count 32): the arrival fires when
// the lane's copies have landed, so the producer never waits for data and the number of rows in
// flight is the ring size.  (The earlier wait_group + fence + arrive scheme exposed the load latency
// once per row: ncu showed the warp parked on the scoreboard of the oldest LDGSTS.)
//
// Per-lane copy plan of one image-row segment.  One warp instruction moves 32/kPP consecutive pixels
// (all kPP planes of each): 512 contiguous bytes of global memory, landing as kPP contiguous runs in
// shared memory (one per plane) - whole sectors on the read side, conflict-free 128-byte wavefronts
// on the write side (splitting a 64-byte run of lanes across two planes doubled the LSU bank-conflict
// count in ncu).  Source and destination of a lane's k-th piece are its base plus a compile-time
// constant: the address registers of an in-flight LDGSTS stay reserved until its data returns, so the
// row loop must not recompute per-piece addresses.
template <int C, int NPIX, int EB = 2>
struct RowPlan {
  static constexpr int kPP = C * EB / 16;                // 16-byte pieces per pixel (= smem planes)
  static constexpr int kLP = 32 / kPP;                   // pixels per warp instruction
  static constexpr int kPer = (NPIX + kLP - 1) / kLP;    // warp instructions per row
  static_assert(kPer <= 64, "edge mask is 64 bits");
  using Mask = typename std::conditional<(kPer > 32), uint64_t, uint32_t>::type;
  uint32_t soff;           // byte offset of the lane's first piece inside the slot
  uint32_t goff;           // byte offset of the lane's first piece from pixel x_first of the row
  Mask ok;                 // bit k: piece k is inside the image horizontally
  int x_first, px0;
  __device__ __forceinline__ void build(int x_first_, int plane_bytes, int W, int lane) {
    x_first = x_first_;
    const int pl = lane / kLP;
    px0 = lane % kLP;
    soff = (uint32_t)(pl * plane_bytes + px0 * 16);
    goff = (uint32_t)(px0 * (C * EB) + pl * 16);
    ok = 0;
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      const int p = px0 + k * kLP, x = x_first + p;
      if (p < NPIX && x >= 0 && x < W) ok |= (Mask)1 << k;
    }
  }
  // dst: [C/8 planes][NPIX][16 B];  src pixel (b, y, x_first + p), channels plane*8..+7; zero-fill outside
  // (a piece with source size 0 reads nothing, so its address may lie outside the row)
  __device__ __forceinline__ void copy(uint8_t* dst, const void* __restrict__ src, int b, int y, int H,
                                       int W) const {
    const bool row_ok = (y >= 0) && (y < H);
    const char* row = reinterpret_cast<const char*>(src) +
                      (((int64_t)b * H + (row_ok ? y : 0)) * (int64_t)W + x_first) * (int64_t)(C * EB);
    const Mask okr = row_ok ? ok : (Mask)0;
    const char* lsrc = row + goff;
    uint8_t* ldst = dst + soff;
#pragma unroll
    for (int k = 0; k < kPer; ++k)
      if ((k + 1) * kLP <= NPIX || px0 + k * kLP < NPIX)
        cp_async16(ldst + k * (kLP * 16), lsrc + k * 512, ((okr >> k) & 1u) ? 16u : 0u);
  }
};

// STATS: 0 = none, 1 = BatchNorm forward statistics of the output (sum, sum of squares; conv2 forward),
// 2 = BatchNorm BACKWARD reductions of the layer below (dgrad): the output IS dp1, the gradient of the
//     pooled layer-1 activation, so with `pooled` = p1 (read before it is overwritten in place) and
//     bnprev = [scale|shift|mean|rstd] of BN1:  dy = dp1 where p1 > 0,
//     partials = [sum(dy) | sum(dy * xhat)],  xhat = ((p1 - shift)/scale - mean) * rstd.
//     This replaces a separate pass over the 1.44 GB conv1 output.
// SWZ (with LDGSTS = false): the input row is one TMA request.  A pixel's channel vector (CIN * sizeof(ET) = 32, 64
// or 128 bytes) is a shared-memory ROW of the K-major operand in the hardware's 32 / 64 / 128-byte swizzle, which is
// also how the pixel lies in NHWC memory: a 4-D tensor map {C, W, H, B} with box {C, 136, 1, 1} drops the whole row
// segment, halo and zero padding included, in one request with an inner extent of a full pixel (the planar layout
// needed one box row per 16-byte piece, which is what made TMA slow here).  The dx tap is still a start-address
// shift - by whole rows now; the swizzle is a function of the address bits, so a shifted start reads each row with
// the phase TMA wrote it with - and a K-step is +32 bytes inside the row.
template <int CIN, int COUT, bool TRANSPOSED, int STATS, bool LDGSTS, typename ET = __nv_bfloat16, bool SWZ = false>
__global__ void __launch_bounds__(kRsThreads, 1)
conv5x5_rs_kernel(const __grid_constant__ CUtensorMap tm_act, const ET* __restrict__ act,
                  const float* __restrict__ w,
                  const float* __restrict__ bias, ET* __restrict__ out,
                  float* __restrict__ partials, const ET* pooled, const float* __restrict__ bnprev,
                  ConvGeom g) {
  using Cfg = RsCfg<CIN, COUT, ET>;
  constexpr bool kF32 = sizeof(ET) == 4;
  static_assert(LDGSTS || SWZ || !kF32, "the planar TMA feed exists for bf16 only");
  static_assert(!(SWZ && LDGSTS), "SWZ is a TMA feed");
  constexpr int kRowB = CIN * (int)sizeof(ET);                       // bytes of one pixel = one operand row (SWZ)
  constexpr uint32_t kSwzLayout = kRowB == 128 ? 2u : kRowB == 64 ? 4u : 6u;
  static_assert(!SWZ || kRowB == 32 || kRowB == 64 || kRowB == 128, "a pixel must be one swizzle span");
  static_assert(!SWZ || (kPW * kRowB == Cfg::kSlotBytes && Cfg::kSlotBytes % (8 * kRowB) == 0 && Cfg::kWBytes % 1024 == 0),
                "row-major slots must keep the swizzle phase");
  constexpr int RT = Cfg::kTStages;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr uint32_t kAlign = SWZ ? 1024u : 128u;
  uint8_t* smem = smem_raw + ((kAlign - (smem_u32(smem_raw) & (kAlign - 1))) & (kAlign - 1));
  uint8_t* s_w = smem;
  uint8_t* s_slots = smem + Cfg::kWBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_slots + Cfg::kSlots * Cfg::kSlotBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + Cfg::kSlots;
  uint64_t* tfull = empty + Cfg::kSlots;       // [RT]
  uint64_t* tempty = tfull + RT;               // [RT]
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(tempty + RT);
  float* s_bias = reinterpret_cast<float*>(s_tmem + 4);
  float* s_stat = s_bias + COUT;               // [8 warps][2*COUT]
  float* s_bnp = s_stat + 2 * kRsEpiWarps * COUT;            // [4][COUT]: (shift, 1/scale, mean, rstd) of the layer below
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (STATS == 2 && threadIdx.x < COUT) {
    const float sc = bnprev[threadIdx.x];
    s_bnp[threadIdx.x] = bnprev[COUT + threadIdx.x];
    s_bnp[COUT + threadIdx.x] = sc != 0.f ? 1.f / sc : 0.f;
    s_bnp[2 * COUT + threadIdx.x] = bnprev[2 * COUT + threadIdx.x];
    s_bnp[3 * COUT + threadIdx.x] = bnprev[3 * COUT + threadIdx.x];
  }

  // weights: element(dx, kk, c, n=(jj,co), e) = (((dx*KSTEPS + kk)*2 + c)*kNB + n)*kCPP + e, dy = 4 - jj
  for (int i = threadIdx.x; i < kTaps * CIN * COUT; i += kRsThreads) {
    const int e = i % Cfg::kCPP;
    int r = i / Cfg::kCPP;
    const int n = r % Cfg::kNB; r /= Cfg::kNB;
    const int c = r & 1; r >>= 1;
    const int kk = r % Cfg::kKSteps;
    const int dx = r / Cfg::kKSteps;
    const int jj = n / COUT, co = n % COUT;
    const int tap = (4 - jj) * 5 + dx;
    const int k = (kk * 2 + c) * Cfg::kCPP + e;
    const float v = TRANSPOSED ? w[((size_t)k * COUT + co) * kTaps + (kTaps - 1 - tap)]
                               : w[((size_t)co * CIN + k) * kTaps + tap];
    if constexpr (kF32) reinterpret_cast<float*>(s_w)[i] = to_tf32_rna(v);
    else reinterpret_cast<__nv_bfloat16*>(s_w)[i] = __float2bfloat16_rn(v);
  }
  if (threadIdx.x < COUT) s_bias[threadIdx.x] = bias ? bias[threadIdx.x] : 0.f;
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    for (int i = 0; i < Cfg::kSlots; ++i) { mbar_init(&full[i], LDGSTS ? 32 : 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < RT; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 128); }
    mbar_fence_init();
    if (!LDGSTS) tma_prefetch_desc(&tm_act);
  }
  if (warp == 1) tmem_alloc<Cfg::kTmemCols>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    if (LDGSTS) {
      // whole warp: cp.async feed; completion of each lane's copies arrives on the row's full barrier
      uint32_t gcount = 0;          // rows issued
      for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
        const int seg = item % g.nseg;
        const int strip = (item / g.nseg) % g.nstrips;
        const int b = item / (g.nseg * g.nstrips);
        const int y0 = seg * g.rseg;
        const int rows = min(g.rseg, g.H - y0);
        const int x0 = strip * kBM;
        RowPlan<CIN, kPW, Cfg::kEB> plan;
        plan.build(x0 - 2, kPlaneBytes, g.W, lane);
        for (int i = 0; i < rows + 4; ++i, ++gcount) {
          const uint32_t slot = gcount % Cfg::kSlots, use = gcount / Cfg::kSlots;
          mbar_wait(&empty[slot], (use & 1) ^ 1);
          plan.copy(s_slots + slot * Cfg::kSlotBytes, act, b, y0 - 2 + i, g.H, g.W);
          cp_async_mbar_arrive_noinc(&full[slot]);
        }
      }
    } else if (lane == 0) {
      uint32_t gcount = 0, qbase = 0;
      for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
        const int seg = item % g.nseg;
        const int strip = (item / g.nseg) % g.nstrips;
        const int b = item / (g.nseg * g.nstrips);
        const int y0 = seg * g.rseg;
        const int rows = min(g.rseg, g.H - y0);
        const int x0 = strip * kBM;
        for (int i = 0; i < rows + 4; ++i, ++gcount) {
          if constexpr (SWZ) {
            // The input row that opens output row i is not fetched before that row's accumulator has been drained:
            // the issuing thread then needs no wait of its own for it (each mbarrier wait costs that thread, the one
            // serial role of the kernel, ~100 cycles even when the phase completed long ago; here it is free).
            if (i < rows) {
              const uint32_t qn = qbase + (uint32_t)i;
              mbar_wait(&tempty[qn % RT], ((qn / RT) & 1) ^ 1);
            }
          }
          const uint32_t slot = gcount % Cfg::kSlots, use = gcount / Cfg::kSlots;
          mbar_wait(&empty[slot], (use & 1) ^ 1);
          mbar_expect_tx(&full[slot], Cfg::kSlotBytes);
          uint8_t* dst = s_slots + slot * Cfg::kSlotBytes;
          if constexpr (SWZ) {
            tma_load_4d(dst, &tm_act, &full[slot], 0, x0 - 2, y0 - 2 + i, b);
          } else {
#pragma unroll
            for (int p = 0; p < Cfg::kPlanes; ++p)
              tma_load_5d(dst + p * kPlaneBytes, &tm_act, &full[slot], 0, p, x0 - 2, y0 - 2 + i, b);
          }
        }
        qbase += rows;
      }
    }
  } else if (warp == 1) {
    // ===== MMA issue.  The WHOLE warp runs this loop (warp-uniform control flow, descriptors in
    // uniform registers); only the tcgen05 instructions are predicated on one elected lane.
    // (Running the loop under `if (lane == 0)` made ptxas serialise every UTCHMMA operand through
    // ELECT / R2UR / BRA.U.ANY: ~285 SASS instructions per row on one lane - the bottleneck ncu showed.)
    constexpr uint32_t idesc0 = umma_idesc(UmmaKind<ET>::kFormat, kBM, 0, 0, 0);   // N filled in per MMA
    const uint32_t slots_addr = smem_u32(s_slots);
    // A: planar no-swizzle (LBO = plane stride, SBO = 8 pixels x 16 B) or swizzled rows (SBO = 8 rows)
    const uint64_t a_tmpl = SWZ ? smem_desc(0, 16, 8 * kRowB, kSwzLayout) : smem_desc_noswz(0, kPlaneBytes, 128);
    const uint32_t a_hi = (uint32_t)(a_tmpl >> 32), a_lo0 = (uint32_t)a_tmpl;
    // 16-byte units from the slot start to the operand of tap dx, K-step kk
    auto a_off = [](int dx, int kk) -> uint32_t {
      return SWZ ? (uint32_t)(dx * (kRowB >> 4) + kk * 2) : (uint32_t)(((2 * kk) * kPlaneBytes >> 4) + dx);
    };
    const uint64_t b_tmpl = smem_desc_noswz(smem_u32(s_w), Cfg::kWChunkBytes, 128);
    const uint32_t b_hi = (uint32_t)(b_tmpl >> 32), b_lo0 = (uint32_t)b_tmpl;
    uint32_t gcount = 0, qbase = 0;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
      const int seg = item % g.nseg;
      const int y0 = seg * g.rseg;
      const int rows = min(g.rseg, g.H - y0);
      for (int i = 0; i < rows + 4; ++i, ++gcount) {
        const int jj_lo = max(0, 4 - i), jj_hi = min(4, rows + 3 - i);     // output row j = i-4+jj in [0, rows)
        const bool starts_row = i < rows;                                  // block jj=4 opens output row i
        if (!SWZ && starts_row) {               // SWZ: implied by the row's full barrier (the feed waited for it)
          const uint32_t qn = qbase + i;
          mbar_wait(&tempty[qn % RT], ((qn / RT) & 1) ^ 1);
        }
        const uint32_t slot = gcount % Cfg::kSlots;
        mbar_wait(&full[slot], (gcount / Cfg::kSlots) & 1);
        tc_fence_after();
        // accumulating blocks [jj_lo, acc_hi] as (at most) two column ranges of the 16-slot TMEM ring
        const int acc_hi_first = starts_row ? min(jj_hi, 3) : jj_hi;       // on dx=0/kk=0, block 4 overwrites
        const uint32_t s0 = (qbase + (uint32_t)(i - 4 + jj_lo)) % RT;
        const uint32_t a_lo_row = a_lo0 + ((slots_addr + slot * Cfg::kSlotBytes) >> 4);
        if (elect_one()) {
          if (i >= 4 && starts_row && s0 + 5 <= (uint32_t)RT) {
            // steady state (all five target rows live, no TMEM-ring wrap): straight-line issue.  The issuing
            // warp is the one serial role of this kernel, so instructions per row are what bound it.
            const uint32_t d0 = tmem_base + s0 * COUT;
#pragma unroll
            for (int dx = 0; dx < 5; ++dx) {
#pragma unroll
              for (int kk = 0; kk < Cfg::kKSteps; ++kk) {
                const uint64_t a_desc = ((uint64_t)a_hi << 32) | (uint64_t)(a_lo_row + a_off(dx, kk));
                const uint32_t b_lo = b_lo0 + (((dx * Cfg::kKSteps + kk) * Cfg::kWStepBytes) >> 4);
                if (dx == 0 && kk == 0) {
                  UmmaKind<ET>::mma(d0, a_desc, ((uint64_t)b_hi << 32) | (uint64_t)b_lo, idesc0 | ((uint32_t)((4 * COUT) >> 3) << 17), 1u);
                  UmmaKind<ET>::mma(d0 + 4 * COUT, a_desc, ((uint64_t)b_hi << 32) | (uint64_t)(b_lo + 4 * COUT),
                            idesc0 | ((uint32_t)(COUT >> 3) << 17), 0u);
                } else {
                  UmmaKind<ET>::mma(d0, a_desc, ((uint64_t)b_hi << 32) | (uint64_t)b_lo, idesc0 | ((uint32_t)((5 * COUT) >> 3) << 17), 1u);
                }
              }
            }
          } else {
#pragma unroll
          for (int dx = 0; dx < 5; ++dx) {
#pragma unroll
            for (int kk = 0; kk < Cfg::kKSteps; ++kk) {
              const uint64_t a_desc = ((uint64_t)a_hi << 32) | (uint64_t)(a_lo_row + a_off(dx, kk));
              const uint32_t b_lo = b_lo0 + (((dx * Cfg::kKSteps + kk) * Cfg::kWStepBytes) >> 4);
              const bool first = (dx == 0 && kk == 0);
              const int acc_hi = first ? acc_hi_first : jj_hi;
              const int nacc = acc_hi - jj_lo + 1;                          // may be <= 0
              const int n1 = min(nacc, (int)(RT - s0));
              if (n1 > 0)
                UmmaKind<ET>::mma(tmem_base + s0 * COUT, a_desc, ((uint64_t)b_hi << 32) | (uint64_t)(b_lo + jj_lo * COUT),
                          idesc0 | ((uint32_t)((n1 * COUT) >> 3) << 17), 1u);
              if (nacc > n1)
                UmmaKind<ET>::mma(tmem_base, a_desc, ((uint64_t)b_hi << 32) | (uint64_t)(b_lo + (jj_lo + n1) * COUT),
                          idesc0 | ((uint32_t)(((nacc - n1) * COUT) >> 3) << 17), 1u);
              if (first && starts_row)
                UmmaKind<ET>::mma(tmem_base + ((qbase + (uint32_t)i) % RT) * COUT, a_desc,
                          ((uint64_t)b_hi << 32) | (uint64_t)(b_lo + 4 * COUT), idesc0 | ((uint32_t)(COUT >> 3) << 17), 0u);
            }
          }
          }
          // SWZ: one commit per row - the epilogue, which learns from tfull(j) that input rows <= j+4 are consumed,
          // hands the slots back (every tcgen05 instruction of this thread queues behind its MMAs)
          if (!SWZ) umma_commit(&empty[slot]);                              // this input row is consumed
          if (i >= 4) umma_commit(&tfull[(qbase + (uint32_t)(i - 4)) % RT]);  // output row i-4 is complete
        }
        __syncwarp();
      }
      qbase += rows;
    }
  } else {
    const int quarter = warp & 3, ew = warp - 2, grp = ew >> 2;      // group g drains rows with (q & 1) == g
    const int m = quarter * 32 + lane;
    float ssum[STATS ? COUT : 1], ssq[STATS ? COUT : 1];
    if (STATS) {
#pragma unroll
      for (int c = 0; c < COUT; ++c) { ssum[c] = 0.f; ssq[c] = 0.f; }
    }
    // STATS == 2: xhat = ((y - shift)/scale - mean) * rstd = y * kA + kB, constants in registers
    float kA[STATS == 2 ? COUT : 1], kB[STATS == 2 ? COUT : 1];
    if (STATS == 2) {
#pragma unroll
      for (int c = 0; c < COUT; ++c) {
        const float inv = s_bnp[COUT + c], rs = s_bnp[3 * COUT + c];
        kA[c] = inv * rs;
        kB[c] = -(s_bnp[c] * inv + s_bnp[2 * COUT + c]) * rs;
      }
    }
    uint32_t q = 0, gbase = 0;                    // gbase: input rows fed before this item
    const bool leader = (ew & 3) == 0 && lane == 0;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
      const int seg = item % g.nseg;
      const int strip = (item / g.nseg) % g.nstrips;
      const int b = item / (g.nseg * g.nstrips);
      const int y0 = seg * g.rseg;
      const int rows = min(g.rseg, g.H - y0);
      const int x = strip * kBM + m;
      const bool valid = x < g.W;
      const uint32_t gitem = gbase;
      gbase += (uint32_t)rows + 4u;
      for (int j = 0; j < rows; ++j, ++q) {
        if ((int)(q & 1u) != grp) continue;
        const uint32_t stage = q % RT, use = q / RT;
        // STATS == 2: fetch this pixel's pooled activation now, so the load latency overlaps the wait for
        // the accumulator (it is read here, before the gradient written below overwrites it in place)
        uint4 pvv[STATS == 2 ? COUT / Cfg::kCPP : 1];
        if (STATS == 2 && valid) {
          const uint4* pp = reinterpret_cast<const uint4*>(pooled + (((size_t)b * g.H + (y0 + j)) * g.W + x) * COUT);
#pragma unroll
          for (int i = 0; i < COUT / Cfg::kCPP; ++i) pvv[i] = pp[i];
        }
        mbar_wait(&tfull[stage], use & 1);
        tc_fence_after();
        if constexpr (SWZ) {
          if (leader) {      // output row j complete => input rows 0 .. j+4 of the item have been read by their MMAs
            if (j == 0) {
#pragma unroll
              for (uint32_t k = 0; k < 4; ++k) mbar_arrive(&empty[(gitem + k) % Cfg::kSlots]);
            }
            mbar_arrive(&empty[(gitem + (uint32_t)j + 4u) % Cfg::kSlots]);
          }
        }
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + stage * COUT;
        uint32_t v[COUT];
#pragma unroll
        for (int c0 = 0; c0 < COUT; c0 += 16) tmem_ld16(taddr + c0, *reinterpret_cast<uint32_t(*)[16]>(&v[c0]));
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(&tempty[stage]);
        if constexpr (kF32) {
          // fp32 storage: the accumulator (+ bias) is stored as is; statistics are those of the stored values
          if (!TRANSPOSED) {
#pragma unroll
            for (int c = 0; c < COUT; ++c) v[c] = __float_as_uint(__uint_as_float(v[c]) + s_bias[c]);
          }
          if (valid) {
            ET* o = out + (((size_t)b * g.H + (y0 + j)) * g.W + x) * COUT;
#pragma unroll
            for (int c = 0; c < COUT; ++c) {
              const float r0 = __uint_as_float(v[c]);
              if (STATS == 1) { ssum[c] += r0; ssq[c] = fmaf(r0, r0, ssq[c]); }
              if (STATS == 2) {
                const float y = __uint_as_float((&pvv[c / 4].x)[c & 3]);
                const float gg = y > 0.f ? r0 : 0.f;
                ssum[c] += gg;
                ssq[c] = fmaf(gg, fmaf(y, kA[c], kB[c]), ssq[c]);
              }
            }
#pragma unroll
            for (int i = 0; i < COUT / 8; ++i) st_global_v8(o + 8 * i, &v[8 * i]);
          }
          continue;
        }
        uint32_t pk[COUT / 2];
#pragma unroll
        for (int c = 0; c < COUT; c += 2) {
          // the transposed (dgrad) convolution has no bias
          pk[c / 2] = TRANSPOSED ? pack_bf16x2(__uint_as_float(v[c]), __uint_as_float(v[c + 1]))
                                 : pack_bf16x2(__uint_as_float(v[c]) + s_bias[c], __uint_as_float(v[c + 1]) + s_bias[c + 1]);
          if (STATS == 1 && valid) {
            const float r0 = bf16_lo(pk[c / 2]), r1 = bf16_hi(pk[c / 2]);
            ssum[c] += r0; ssq[c] = fmaf(r0, r0, ssq[c]);
            ssum[c + 1] += r1; ssq[c + 1] = fmaf(r1, r1, ssq[c + 1]);
          }
        }
        if (valid) {
          __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out) + (((size_t)b * g.H + (y0 + j)) * g.W + x) * COUT;
          if (STATS == 2) {
#pragma unroll
            for (int c = 0; c < COUT; ++c) {
              const uint32_t pw = (&pvv[(c / 8) % (COUT / Cfg::kCPP)].x)[(c & 7) >> 1];
              const float y = (c & 1) ? bf16_hi(pw) : bf16_lo(pw);
              const float gq = (c & 1) ? bf16_hi(pk[c / 2]) : bf16_lo(pk[c / 2]);
              const float gg = y > 0.f ? gq : 0.f;
              ssum[c] += gg;
              ssq[c] = fmaf(gg, fmaf(y, kA[c], kB[c]), ssq[c]);
            }
          }
#pragma unroll
          for (int i = 0; i < COUT / 16; ++i) st_global_v8(o + 16 * i, &pk[8 * i]);   // 32-byte stores: whole sectors
        }
      }
    }
    if (STATS) {
#pragma unroll
      for (int c = 0; c < COUT; ++c) {
        const float a = warp_sum(ssum[c]), s2 = warp_sum(ssq[c]);
        if (lane == 0) { s_stat[ew * 2 * COUT + c] = a; s_stat[ew * 2 * COUT + COUT + c] = s2; }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (STATS && threadIdx.x < 2 * COUT) {
    float a = 0.f;
#pragma unroll
    for (int w4 = 0; w4 < kRsEpiWarps; ++w4) a += s_stat[w4 * 2 * COUT + threadIdx.x];
    partials[(size_t)blockIdx.x * 2 * COUT + threadIdx.x] = a;
  }
  if (warp == 1) { __syncwarp(); tmem_dealloc<Cfg::kTmemCols>(tmem_base); }
}

// wgrad, row-stationary: the act (p1) row is the stationary operand; the five dx2 rows it
// pairs with (y = r-2 .. r+2) are five CONSECUTIVE slots of the dx2 ring, i.e. 20 consecutive
// 8-channel planes, i.e. one MN-major B operand with N = 160.  D[pl][(dx,ci8)][(jj,co)], dy = 4-jj.
constexpr int kWg2SlotsA = 12, kWg2SlotsB = 16;      // rows in flight: the rings are the cp.async pipeline depth
// The five dx2 rows of a window must be CONTIGUOUS in shared memory (one B operand, N = 160).  Slots 16..19 mirror
// slots 0..3 (the feed writes those rows twice), so a window that starts in slots 12..15 runs on into the mirror
// instead of wrapping: one MMA per K-step for every row.  (Splitting the wrapped windows into two MMAs re-read A and
// paid the 48-cycle MMA floor twice on a quarter of the rows of a kernel that is bound by the MMAs themselves.)
constexpr int kWg2PhysB = kWg2SlotsB + 4;
constexpr int kWg2SmemBytes = kWg2SlotsA * kWgSlotABytes + kWg2PhysB * kWgSlotBBytes + 1024 + 128;

template <bool LDGSTS>
__global__ void __launch_bounds__(kThreads, 1)
conv5x5_wgrad_rs_kernel(const __grid_constant__ CUtensorMap tm_act, const __grid_constant__ CUtensorMap tm_dxo,
                        const __nv_bfloat16* __restrict__ act, const __nv_bfloat16* __restrict__ dxo,
                        float* __restrict__ partials, ConvGeom g) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  uint8_t* s_a = smem;
  uint8_t* s_b = smem + kWg2SlotsA * kWgSlotABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_b + kWg2PhysB * kWgSlotBBytes);
  uint64_t* fullA = bars;
  uint64_t* emptyA = fullA + kWg2SlotsA;
  uint64_t* fullB = emptyA + kWg2SlotsA;
  uint64_t* emptyB = fullB + kWg2SlotsB;
  uint64_t* done = emptyB + kWg2SlotsB;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kWg2SlotsA; ++i) { mbar_init(&fullA[i], LDGSTS ? 32 : 1); mbar_init(&emptyA[i], 1); }
    for (int i = 0; i < kWg2SlotsB; ++i) { mbar_init(&fullB[i], LDGSTS ? 32 : 1); mbar_init(&emptyB[i], 1); }
    mbar_init(done, 1);
    mbar_fence_init();
    tma_prefetch_desc(&tm_act);
    tma_prefetch_desc(&tm_dxo);
  }
  if (warp == 1) tmem_alloc<kWgTmemCols>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    if (LDGSTS) {
      uint32_t ga = 0, gb = 0;
      for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
        const int seg = item % g.nseg;
        const int strip = (item / g.nseg) % g.nstrips;
        const int b = item / (g.nseg * g.nstrips);
        const int r0 = seg * g.rseg;
        const int rows = min(g.rseg, g.H - r0);
        const int x0 = strip * kBM;
        RowPlan<32, kBM> planB;
        planB.build(x0, kWgPlaneBBytes, g.W, lane);
        RowPlan<16, kPW> planA;
        planA.build(x0 - 2, kPlaneBytes, g.W, lane);
        for (int i = 0; i < rows + 4; ++i) {
          {
            const uint32_t slot = gb % kWg2SlotsB, use = gb / kWg2SlotsB;
            mbar_wait(&emptyB[slot], (use & 1) ^ 1);
            planB.copy(s_b + slot * kWgSlotBBytes, dxo, b, r0 - 2 + i, g.H, g.W);
            if (slot < 4) planB.copy(s_b + (kWg2SlotsB + slot) * kWgSlotBBytes, dxo, b, r0 - 2 + i, g.H, g.W);   // mirror
            cp_async_mbar_arrive_noinc(&fullB[slot]);
            ++gb;
          }
          if (i >= 4) {
            const uint32_t slot = ga % kWg2SlotsA, use = ga / kWg2SlotsA;
            mbar_wait(&emptyA[slot], (use & 1) ^ 1);
            planA.copy(s_a + slot * kWgSlotABytes, act, b, r0 + i - 4, g.H, g.W);
            cp_async_mbar_arrive_noinc(&fullA[slot]);
            ++ga;
          }
        }
      }
    } else if (lane == 0) {
      uint32_t ga = 0, gb = 0;
      for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
        const int seg = item % g.nseg;
        const int strip = (item / g.nseg) % g.nstrips;
        const int b = item / (g.nseg * g.nstrips);
        const int r0 = seg * g.rseg;
        const int rows = min(g.rseg, g.H - r0);
        const int x0 = strip * kBM;
        for (int i = 0; i < rows + 4; ++i) {
          {   // dx2 row r0-2+i (zero-filled outside the image)
            const uint32_t slot = gb % kWg2SlotsB, use = gb / kWg2SlotsB;
            mbar_wait(&emptyB[slot], (use & 1) ^ 1);
            mbar_expect_tx(&fullB[slot], slot < 4 ? 2 * kWgSlotBBytes : kWgSlotBBytes);
            uint8_t* dst = s_b + slot * kWgSlotBBytes;
#pragma unroll
            for (int p = 0; p < 4; ++p)
              tma_load_5d(dst + p * kWgPlaneBBytes, &tm_dxo, &fullB[slot], 0, p, x0, r0 - 2 + i, b);
            if (slot < 4) {   // mirror slot (see kWg2PhysB)
              uint8_t* dst2 = s_b + (kWg2SlotsB + slot) * kWgSlotBBytes;
#pragma unroll
              for (int p = 0; p < 4; ++p)
                tma_load_5d(dst2 + p * kWgPlaneBBytes, &tm_dxo, &fullB[slot], 0, p, x0, r0 - 2 + i, b);
            }
            ++gb;
          }
          if (i >= 4) {   // act row r0+i-4: all five dx2 rows it needs are now in flight
            const uint32_t slot = ga % kWg2SlotsA, use = ga / kWg2SlotsA;
            mbar_wait(&emptyA[slot], (use & 1) ^ 1);
            mbar_expect_tx(&fullA[slot], kWgSlotABytes);
            uint8_t* dst = s_a + slot * kWgSlotABytes;
            tma_load_5d(dst, &tm_act, &fullA[slot], 0, 0, x0 - 2, r0 + i - 4, b);
            tma_load_5d(dst + kPlaneBytes, &tm_act, &fullA[slot], 0, 1, x0 - 2, r0 + i - 4, b);
            ++ga;
          }
        }
      }
    }
  } else if (warp == 1) {
    // warp-uniform loop, tcgen05 instructions on one elected lane (see conv5x5_rs_kernel)
    constexpr uint32_t idesc0 = umma_idesc_bf16(64, 0, 1, 1);
    const uint64_t a_tmpl = smem_desc_noswz(0, 128, 16), b_tmpl = smem_desc_noswz(0, 128, kWgPlaneBBytes);
    const uint32_t a_hi = (uint32_t)(a_tmpl >> 32), b_hi = (uint32_t)(b_tmpl >> 32);
    const uint32_t a_base = (uint32_t)a_tmpl + (smem_u32(s_a) >> 4), b_base = (uint32_t)b_tmpl + (smem_u32(s_b) >> 4);
    uint32_t ga = 0, gbase = 0;
    uint32_t first = 1;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
      const int seg = item % g.nseg;
      const int r0 = seg * g.rseg;
      const int rows = min(g.rseg, g.H - r0);
      for (int i = 0; i < rows; ++i, ++ga) {
        for (int k = (i == 0 ? 0 : 4); k < 5; ++k) {          // dx2 local rows i..i+4 have landed
          const uint32_t gc = gbase + i + k;
          mbar_wait(&fullB[gc % kWg2SlotsB], (gc / kWg2SlotsB) & 1);
        }
        const uint32_t slotA = ga % kWg2SlotsA;
        mbar_wait(&fullA[slotA], (ga / kWg2SlotsA) & 1);
        tc_fence_after();
        const uint32_t s0 = (gbase + i) % kWg2SlotsB;          // ring slot of the window's first row
        const uint32_t arow = a_base + slotA * (kWgSlotABytes >> 4);
        const uint32_t brow = b_base + s0 * (kWgSlotBBytes >> 4);   // slots s0..s0+4 (<= 19: mirror slots) are contiguous
        const uint32_t acc0 = first ? 0u : 1u;
        if (elect_one()) {
#pragma unroll
          for (int kb = 0; kb < kBM / 16; ++kb) {
            const uint32_t acc = kb == 0 ? acc0 : 1u;
            const uint64_t bd = ((uint64_t)b_hi << 32) | (uint64_t)(brow + kb * 16);
#pragma unroll
            for (int pl = 0; pl < 2; ++pl) {
              const uint64_t a_desc = ((uint64_t)a_hi << 32) | (uint64_t)(arow + pl * (kPlaneBytes >> 4) + kb * 16);
              umma_bf16(tmem_base + pl * 160, a_desc, bd, idesc0 | ((uint32_t)(160 >> 3) << 17), acc);
            }
          }
          umma_commit(&emptyA[slotA]);
          umma_commit(&emptyB[(gbase + i) % kWg2SlotsB]);
        }
        __syncwarp();
        first = 0;
      }
      if (elect_one())
        for (int k = 0; k < 4; ++k) umma_commit(&emptyB[(gbase + rows + k) % kWg2SlotsB]);
      __syncwarp();
      gbase += rows + 4;
    }
    if (elect_one()) umma_commit(done);
    __syncwarp();
  } else {
    const int quarter = warp & 3;
    float* dst = partials + (size_t)blockIdx.x * (32 * 16 * kTaps);
    mbar_wait(done, 0);
    tc_fence_after();
    const int r = quarter * 16 + (lane & 15);
    const int dxs = r >> 3, e = r & 7;
    const bool row_valid = (lane < 16) && (dxs < 5);
#pragma unroll 1
    for (int acc = 0; acc < 10; ++acc) {
      const int pl = acc / 5, jj = acc % 5, dy = 4 - jj;
      uint32_t v[32];
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + pl * 160 + jj * 32;
      tmem_ld16(taddr, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
      tmem_ld16(taddr + 16, *reinterpret_cast<uint32_t(*)[16]>(&v[16]));
      tmem_ld_wait();
      if (row_valid) {
        const int ci = pl * 8 + e;
#pragma unroll
        for (int co = 0; co < 32; ++co)
          dst[((size_t)co * 16 + ci) * kTaps + dy * 5 + dxs] = __uint_as_float(v[co]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) { __syncwarp(); tmem_dealloc<kWgTmemCols>(tmem_base); }
}


// ===================================================== wgrad, dx-replicated A ===
// The kernels above give the weight gradient only half-filled MMAs: with the dx shift expressed as the
// M-stride of the act operand, one MMA covers one 8-channel plane, M = 64 with 40 useful rows, and a
// plane pair needs two of them.  Here the feed writes the act row FIVE times, each copy shifted by one
// pixel, as chunk rows [(dx, plane)][pixel][16 B].  The chunks of all (dx, ci) now sit at ONE constant
// stride, so a single M = 128 MMA (80 useful rows: m = dx*16 + ci) contracts a K-step against the five
// dx2 rows (N = 160): half the MMA count of the version above.  The extra copies cost cp.async wavefronts
// in the L1 data pipe (one per 32-byte sector), which is what bounds this version: 0.52 vs 0.56 ms at
// batch 5.  (The template also instantiates for float, but the hardware has no no-swizzle MN-major layout
// for TF32 - such MMAs accumulate zeros - so fp32 storage uses conv5x5_wgrad_tf32_kernel below.)
template <typename ET>
struct WgRepCfg {
  static constexpr int kEB = (int)sizeof(ET);
  static constexpr int kCPP = 16 / kEB;                 // channels per 16-byte chunk
  static constexpr int kSP = kEB == 2 ? 128 : 64;       // strip width in pixels
  static constexpr int kChunkRow = kSP * 16;            // bytes of one chunk over the strip
  static constexpr int kPlanesA = 16 / kCPP;            // act:  16 channels
  static constexpr int kPlanesB = 32 / kCPP;            // dx2:  32 channels
  static constexpr int kSlotA = 5 * kPlanesA * kChunkRow;   // 20 KB (both types)
  static constexpr int kSlotB = kPlanesB * kChunkRow;       //  8 KB
  static constexpr int kSlotsA = 4;
  static constexpr int kSlotsB = 12, kPhysB = kSlotsB + 4;  // slots 12..15 mirror 0..3 (windows never wrap)
  static constexpr int kKSteps = kSP / UmmaKind<ET>::kK;    // 8
  static constexpr int kKAdv = UmmaKind<ET>::kK * 16;       // bytes per K-step along a chunk row
  static constexpr int kTmemCols = 256;                     // one 128 x 160 fp32 accumulator
  static constexpr int kSmemBytes = kSlotsA * kSlotA + kPhysB * kSlotB + 1024 + 128;
  // an M = 128 operand spans 128 / kCPP chunk rows from the slot start; the rows past the 5 * kPlanesA real
  // ones belong to accumulator lanes >= 80 that nobody reads, but the bytes must exist: they fall into the
  // next A slots / the B ring, which follows the A ring in shared memory.
  static_assert((128 / kCPP) * kChunkRow <= kSlotA + kPhysB * kSlotB, "A operand overrun must stay inside the B ring");
};

template <typename ET>
__global__ void __launch_bounds__(kThreads, 1)
conv5x5_wgrad_rep_kernel(const ET* __restrict__ act, const ET* __restrict__ dxo, float* __restrict__ partials,
                         ConvGeom g) {
  using Cfg = WgRepCfg<ET>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  uint8_t* s_a = smem;
  uint8_t* s_b = smem + Cfg::kSlotsA * Cfg::kSlotA;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_b + Cfg::kPhysB * Cfg::kSlotB);
  uint64_t* fullA = bars;
  uint64_t* emptyA = fullA + Cfg::kSlotsA;
  uint64_t* fullB = emptyA + Cfg::kSlotsA;
  uint64_t* emptyB = fullB + Cfg::kSlotsB;
  uint64_t* done = emptyB + Cfg::kSlotsB;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < Cfg::kSlotsA; ++i) { mbar_init(&fullA[i], 32); mbar_init(&emptyA[i], 1); }
    for (int i = 0; i < Cfg::kSlotsB; ++i) { mbar_init(&fullB[i], 32); mbar_init(&emptyB[i], 1); }
    mbar_init(done, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<Cfg::kTmemCols>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ===== feed: whole warp, 16-byte cp.async, arrival on the row's barrier when the lane's copies land
    uint32_t ga = 0, gb = 0;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
      const int seg = item % g.nseg;
      const int strip = (item / g.nseg) % g.nstrips;
      const int b = item / (g.nseg * g.nstrips);
      const int r0 = seg * g.rseg;
      const int rows = min(g.rseg, g.H - r0);
      const int x0 = strip * Cfg::kSP;
      RowPlan<32, Cfg::kSP, Cfg::kEB> planB;
      planB.build(x0, Cfg::kChunkRow, g.W, lane);
      RowPlan<16, Cfg::kSP, Cfg::kEB> planA[5];          // one per dx: same lane offsets, different edge masks
#pragma unroll
      for (int dx = 0; dx < 5; ++dx) planA[dx].build(x0 - 2 + dx, Cfg::kChunkRow, g.W, lane);
      for (int i = 0; i < rows + 4; ++i) {
        {
          const uint32_t slot = gb % Cfg::kSlotsB, use = gb / Cfg::kSlotsB;
          mbar_wait(&emptyB[slot], (use & 1) ^ 1);
          planB.copy(s_b + slot * Cfg::kSlotB, dxo, b, r0 - 2 + i, g.H, g.W);
          if (slot < 4) planB.copy(s_b + (Cfg::kSlotsB + slot) * Cfg::kSlotB, dxo, b, r0 - 2 + i, g.H, g.W);
          cp_async_mbar_arrive_noinc(&fullB[slot]);
          ++gb;
        }
        if (i >= 4) {
          const uint32_t slot = ga % Cfg::kSlotsA, use = ga / Cfg::kSlotsA;
          mbar_wait(&emptyA[slot], (use & 1) ^ 1);
          uint8_t* dst = s_a + slot * Cfg::kSlotA;
#pragma unroll
          for (int dx = 0; dx < 5; ++dx)
            planA[dx].copy(dst + dx * Cfg::kPlanesA * Cfg::kChunkRow, act, b, r0 + i - 4, g.H, g.W);
          cp_async_mbar_arrive_noinc(&fullA[slot]);
          ++ga;
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issue: warp-uniform loop, tcgen05 instructions on one elected lane
    constexpr uint32_t idesc = umma_idesc(UmmaKind<ET>::kFormat, 128, 160, 1, 1);
    const uint64_t tmpl = smem_desc_noswz(0, 128, Cfg::kChunkRow);     // MN-major: LBO = 8 k-rows, SBO = next chunk row
    const uint32_t hi = (uint32_t)(tmpl >> 32);
    const uint32_t a_base = (uint32_t)tmpl + (smem_u32(s_a) >> 4), b_base = (uint32_t)tmpl + (smem_u32(s_b) >> 4);
    uint32_t ga = 0, gbase = 0;
    uint32_t first = 1;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
      const int seg = item % g.nseg;
      const int r0 = seg * g.rseg;
      const int rows = min(g.rseg, g.H - r0);
      for (int i = 0; i < rows; ++i, ++ga) {
        for (int k = (i == 0 ? 0 : 4); k < 5; ++k) {          // dx2 local rows i..i+4 have landed
          const uint32_t gc = gbase + i + k;
          mbar_wait(&fullB[gc % Cfg::kSlotsB], (gc / Cfg::kSlotsB) & 1);
        }
        const uint32_t slotA = ga % Cfg::kSlotsA;
        mbar_wait(&fullA[slotA], (ga / Cfg::kSlotsA) & 1);
        tc_fence_after();
        const uint32_t s0 = (gbase + i) % Cfg::kSlotsB;
        const uint32_t arow = a_base + slotA * (Cfg::kSlotA >> 4);
        const uint32_t brow = b_base + s0 * (Cfg::kSlotB >> 4);     // slots s0..s0+4 (<= 15: mirrors) are contiguous
        const uint32_t acc0 = first ? 0u : 1u;
        if (elect_one()) {
#pragma unroll
          for (int kb = 0; kb < Cfg::kKSteps; ++kb)
            UmmaKind<ET>::mma(tmem_base, ((uint64_t)hi << 32) | (uint64_t)(arow + kb * (Cfg::kKAdv >> 4)),
                              ((uint64_t)hi << 32) | (uint64_t)(brow + kb * (Cfg::kKAdv >> 4)), idesc,
                              kb == 0 ? acc0 : 1u);
          umma_commit(&emptyA[slotA]);
          umma_commit(&emptyB[(gbase + i) % Cfg::kSlotsB]);
        }
        __syncwarp();
        first = 0;
      }
      if (elect_one())
        for (int k = 0; k < 4; ++k) umma_commit(&emptyB[(gbase + rows + k) % Cfg::kSlotsB]);
      __syncwarp();
      gbase += rows + 4;
    }
    if (elect_one()) umma_commit(done);
    __syncwarp();
  } else {
    // ===== extraction: lane m = dx*16 + ci (m < 80), column n = jj*32 + co, dy = 4 - jj
    const int quarter = warp & 3;
    float* dst = partials + (size_t)blockIdx.x * (32 * 16 * kTaps);
    mbar_wait(done, 0);
    tc_fence_after();
    const int m = quarter * 32 + lane;
    const int dxs = m >> 4, ci = m & 15;
#pragma unroll 1
    for (int jj = 0; jj < 5; ++jj) {
      const int dy = 4 - jj;
      uint32_t v[32];
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + jj * 32;
      tmem_ld16(taddr, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
      tmem_ld16(taddr + 16, *reinterpret_cast<uint32_t(*)[16]>(&v[16]));
      tmem_ld_wait();
      if (m < 80) {
#pragma unroll
        for (int co = 0; co < 32; ++co)
          dst[((size_t)co * 16 + ci) * kTaps + dy * 5 + dxs] = __uint_as_float(v[co]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) { __syncwarp(); tmem_dealloc<Cfg::kTmemCols>(tmem_base); }
}

// =================================================== wgrad, fp32 storage, TF32 ===
// MN-major TF32 operands exist in exactly one shared-memory layout: 128-byte rows (32 fp32 MN elements per
// K row = pixel), 32-byte units XOR-swizzled by the row index (sw32_off, ptx.cuh).  That happens to be the
// NHWC memory of a 32-channel fp32 tensor, so a dx2 row segment is copied pixel by pixel with only the unit
// permutation applied, and five consecutive row slots are again one N = 160 operand (LBO = slot stride).
// The act tensor has 16 channels (64 B per pixel): an A row pairs two neighbouring pixels, row k holds
// [p1[x+k] | p1[x+k+1]].  MN atom j of the operand must hold [p1[x+k+2j] | p1[x+k+2j+1]] - which is the SAME tile
// two rows further down: the descriptor's atom stride (LBO) is simply 256 bytes.  The swizzle is a function of
// the shared-memory address bits, so a tile read from a shifted start sees each row with the phase it was
// written with.  One copy of the row (each pixel stored twice) serves dx = 0..7: M = 128, m = 16*dx + ci.
namespace wt {
constexpr int kSP = 64;                         // strip width in pixels
constexpr int kAtom = kSP * 128;                // 8 KB: one 32-element MN atom over the strip
constexpr int kRowsA = kSP + 8;                 // rows k = 0..69 are read (4 atoms, 2 rows apart); padded to the swizzle period
constexpr int kSlotA = kRowsA * 128;            // 9 KB
constexpr int kSlotB = kAtom;                   //  8 KB
constexpr int kSlotsA = 8;
constexpr int kSlotsB = 12, kPhysB = kSlotsB + 4;
constexpr int kKSteps = kSP / 8;
constexpr int kTmemCols = 256;
constexpr int kSmemBytes = kSlotsA * kSlotA + kPhysB * kSlotB + 1024 + 1024;
}  // namespace wt

// TMAB: the dx2 row is one TMA request (tensor map {32 ch, W, H, B}, box {32, 64, 1, 1}, swizzle 128B with 32-byte
// atoms = exactly the MN-major TF32 operand layout) instead of 512 cp.async pieces: the cp.async wavefronts in the L1
// data pipe (one per 32-byte sector) were a third of this kernel's shared-memory traffic.  The act row keeps cp.async:
// its rows pair two neighbouring pixels, and a tensor map cannot zero-fill on the SUM of two coordinates.
template <bool TMAB>
__global__ void __launch_bounds__(kThreads, 1)
conv5x5_wgrad_tf32_kernel(const __grid_constant__ CUtensorMap tm_dxo, const float* __restrict__ act,
                          const float* __restrict__ dxo, float* __restrict__ partials, ConvGeom g) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);     // swizzle atoms: 1 KB aligned
  uint8_t* s_a = smem;
  uint8_t* s_b = smem + wt::kSlotsA * wt::kSlotA;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_b + wt::kPhysB * wt::kSlotB);
  uint64_t* fullA = bars;
  uint64_t* emptyA = fullA + wt::kSlotsA;
  uint64_t* fullB = emptyA + wt::kSlotsA;
  uint64_t* emptyB = fullB + wt::kSlotsB;
  uint64_t* done = emptyB + wt::kSlotsB;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < wt::kSlotsA; ++i) { mbar_init(&fullA[i], 32); mbar_init(&emptyA[i], 1); }
    // fullB[row]: the dx2 row (TMA: 1 arrival + bytes, cp.async: 32) AND the act row fed in the same iteration (32
    // arrivals) - the issuing thread waits for one barrier per output row, not two
    for (int i = 0; i < wt::kSlotsB; ++i) { mbar_init(&fullB[i], TMAB ? 33 : 64); mbar_init(&emptyB[i], 1); }
    mbar_init(done, 1);
    mbar_fence_init();
    if (TMAB) tma_prefetch_desc(&tm_dxo);
  }
  if (warp == 1) tmem_alloc<wt::kTmemCols>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ===== feed.  dx2: one instruction = 4 pixels x 8 pieces (512 contiguous bytes of the row).
    //              act: one instruction = 8 pixels x 4 pieces (512 contiguous bytes), written to half h of atom j.
    const int bp = lane >> 3, bq = lane & 7;           // dx2: pixel-in-group, 16-byte piece (unit = bq >> 1)
    const int ap = lane >> 2, aq = lane & 3;           // act: pixel-in-group, 16-byte piece of the 64-byte pixel
    uint32_t ga = 0, gb = 0, xb_mine = 0;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
      const int seg = item % g.nseg;
      const int strip = (item / g.nseg) % g.nstrips;
      const int b = item / (g.nseg * g.nstrips);
      const int r0 = seg * g.rseg;
      const int rows = min(g.rseg, g.H - r0);
      const int x0 = strip * wt::kSP;
      const uint32_t gb0 = gb;                           // dx2-row counter at the start of the item
      for (int i = 0; i < rows + 4; ++i) {
        const uint32_t slotb = gb % wt::kSlotsB;
        if constexpr (TMAB) {
          const uint32_t slot = gb % wt::kSlotsB, use = gb / wt::kSlotsB;
          if (lane == 0) {
            mbar_wait(&emptyB[slot], (use & 1) ^ 1);
            mbar_expect_tx(&fullB[slot], slot < 4 ? 2 * wt::kSlotB : wt::kSlotB);
            tma_load_4d(s_b + slot * wt::kSlotB, &tm_dxo, &fullB[slot], 0, x0, r0 - 2 + i, b);
            if (slot < 4)      // mirror of slots 0..3
              tma_load_4d(s_b + (wt::kSlotsB + slot) * wt::kSlotB, &tm_dxo, &fullB[slot], 0, x0, r0 - 2 + i, b);
          }
          __syncwarp();
          ++gb;
        } else {
          const uint32_t slot = gb % wt::kSlotsB, use = gb / wt::kSlotsB;
          mbar_wait(&emptyB[slot], (use & 1) ^ 1);
          const int y = r0 - 2 + i;
          const bool row_ok = y >= 0 && y < g.H;
          const char* row = reinterpret_cast<const char*>(dxo) + (((int64_t)b * g.H + (row_ok ? y : 0)) * g.W + x0) * 128;
          uint8_t* dst = s_b + slot * wt::kSlotB;
          uint8_t* dst2 = s_b + (wt::kSlotsB + slot) * wt::kSlotB;            // mirror of slots 0..3
#pragma unroll
          for (int k4 = 0; k4 < wt::kSP / 4; ++k4) {
            const int k = k4 * 4 + bp;
            const uint32_t nbytes = (row_ok && x0 + k < g.W) ? 16u : 0u;
            const uint32_t off = sw32_off(k, bq >> 1) + (bq & 1) * 16;
            cp_async16(dst + off, row + k * 128 + bq * 16, nbytes);
            if (slot < 4) cp_async16(dst2 + off, row + k * 128 + bq * 16, nbytes);
          }
          cp_async_mbar_arrive_noinc(&fullB[slot]);
          ++gb;
        }
        if (i >= 4) {
          // An act slot is free when the MMAs of the output row it served have completed - which is what the commit on
          // that row's OLDEST dx2 slot (index xb) already says, so the issuing thread makes one commit per row, not two.
          // Lane s keeps xb of act slot s.  If the feed has since refilled that dx2 slot it has seen the phase complete.
          const uint32_t slot = ga % wt::kSlotsA;
          const uint32_t xb_old = __shfl_sync(0xffffffffu, xb_mine, (int)slot);
          if (ga >= (uint32_t)wt::kSlotsA && xb_old + (uint32_t)wt::kSlotsB >= gb)
            mbar_wait(&emptyB[xb_old % wt::kSlotsB], (xb_old / wt::kSlotsB) & 1);
          if (lane == (int)slot) xb_mine = gb0 + (uint32_t)(i - 4);
          const int y = r0 + i - 4;                                            // always inside the image
          const char* row = reinterpret_cast<const char*>(act) + ((int64_t)b * g.H + y) * g.W * 64;
          uint8_t* dst = s_a + slot * wt::kSlotA;
#pragma unroll
          for (int h = 0; h < 2; ++h) {                                        // half h of row k = pixel x0-2+k+h
#pragma unroll
            for (int k8 = 0; k8 < wt::kRowsA / 8; ++k8) {
              const int k = k8 * 8 + ap;
              const int x = x0 - 2 + h + k;
              const uint32_t nbytes = (x >= 0 && x < g.W) ? 16u : 0u;
              cp_async16(dst + sw32_off(k, (uint32_t)(h * 2 + (aq >> 1))) + (aq & 1) * 16,
                         row + (int64_t)(nbytes ? x : 0) * 64 + aq * 16, nbytes);
            }
          }
          cp_async_mbar_arrive_noinc(&fullB[slotb]);      // lands with this iteration's dx2 row: output row i-4 is ready
          ++ga;
        } else {
          mbar_arrive(&fullB[slotb]);                     // no act row in the first four iterations of an item
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc(2, 128, 160, 1, 1);
    // MN-major SW128_32B: LBO = distance between 32-element MN atoms, SBO = distance between 4-row K groups
    const uint64_t a_tmpl = smem_desc(0, 256, 512, 1), b_tmpl = smem_desc(0, wt::kSlotB, 512, 1);   // A: atom = +2 rows
    const uint32_t a_hi = (uint32_t)(a_tmpl >> 32), b_hi = (uint32_t)(b_tmpl >> 32);
    const uint32_t a_base = (uint32_t)a_tmpl + (smem_u32(s_a) >> 4), b_base = (uint32_t)b_tmpl + (smem_u32(s_b) >> 4);
    uint32_t ga = 0, gbase = 0;
    uint32_t first = 1;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
      const int seg = item % g.nseg;
      const int r0 = seg * g.rseg;
      const int rows = min(g.rseg, g.H - r0);
      for (int i = 0; i < rows; ++i, ++ga) {
        for (int k = (i == 0 ? 0 : 4); k < 5; ++k) {
          const uint32_t gc = gbase + i + k;
          mbar_wait(&fullB[gc % wt::kSlotsB], (gc / wt::kSlotsB) & 1);
        }
        const uint32_t slotA = ga % wt::kSlotsA;          // its act row arrived on the newest dx2 row's barrier
        tc_fence_after();
        const uint32_t s0 = (gbase + i) % wt::kSlotsB;
        const uint32_t arow = a_base + slotA * (wt::kSlotA >> 4);
        const uint32_t brow = b_base + s0 * (wt::kSlotB >> 4);
        const uint32_t acc0 = first ? 0u : 1u;
        if (elect_one()) {
#pragma unroll
          for (int kb = 0; kb < wt::kKSteps; ++kb)       // 8 pixels = 1 KB further down every atom
            umma_tf32(tmem_base, ((uint64_t)a_hi << 32) | (uint64_t)(arow + kb * 64),
                      ((uint64_t)b_hi << 32) | (uint64_t)(brow + kb * 64), idesc, kb == 0 ? acc0 : 1u);
          umma_commit(&emptyB[(gbase + i) % wt::kSlotsB]);     // frees the oldest dx2 row AND this row's act slot (feed)
        }
        __syncwarp();
        first = 0;
      }
      if (elect_one())
        for (int k = 0; k < 4; ++k) umma_commit(&emptyB[(gbase + rows + k) % wt::kSlotsB]);
      __syncwarp();
      gbase += rows + 4;
    }
    if (elect_one()) umma_commit(done);
    __syncwarp();
  } else {
    const int quarter = warp & 3;
    float* dst = partials + (size_t)blockIdx.x * (32 * 16 * kTaps);
    mbar_wait(done, 0);
    tc_fence_after();
    const int m = quarter * 32 + lane;
    const int dxs = m >> 4, ci = m & 15;
#pragma unroll 1
    for (int jj = 0; jj < 5; ++jj) {
      const int dy = 4 - jj;
      uint32_t v[32];
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + jj * 32;
      tmem_ld16(taddr, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
      tmem_ld16(taddr + 16, *reinterpret_cast<uint32_t(*)[16]>(&v[16]));
      tmem_ld_wait();
      if (m < 80) {
#pragma unroll
        for (int co = 0; co < 32; ++co)
          dst[((size_t)co * 16 + ci) * kTaps + dy * 5 + dxs] = __uint_as_float(v[co]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) { __syncwarp(); tmem_dealloc<wt::kTmemCols>(tmem_base); }
}

// ------------------------------------------------------------- host side ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  });
  return fn;
}

// 5-D map over a bf16 NHWC activation: {8 ch, C/8 planes, W, H, B}; box {8,1,boxW,1,1}.
int make_act_tmap(CUtensorMap* tm, const void* ptr, int B, int H, int W, int C, int boxW) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return b200_fail(-3, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t gdim[5] = {8, (cuuint64_t)(C / 8), (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t gstr[4] = {16, (cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t box[5] = {8, 1, (cuuint32_t)boxW, 1, 1};
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(ptr), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return b200_fail((int)r, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

// 4-D map over an NHWC activation of `eb`-byte elements: {C, W, H, B}; box {C, boxW, 1, 1}; the inner extent is one
// whole pixel (32 / 64 / 128 bytes) and selects the matching shared-memory swizzle.
int make_row_tmap(CUtensorMap* tm, const void* ptr, int B, int H, int W, int C, int eb, int boxW, bool atom32 = false) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return b200_fail(-3, "cuTensorMapEncodeTiled entry point not available");
  const int rowb = C * eb;
  cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t gstr[3] = {(cuuint64_t)rowb, (cuuint64_t)W * rowb, (cuuint64_t)H * W * rowb};
  cuuint32_t box[4] = {(cuuint32_t)C, (cuuint32_t)boxW, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  // atom32: 128-byte rows whose 32-byte units are permuted (the MN-major TF32 operand layout, sw32_off in ptx.cuh)
  const CUtensorMapSwizzle swz = atom32        ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                                 : rowb == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                 : rowb == 64  ? CU_TENSOR_MAP_SWIZZLE_64B
                                               : CU_TENSOR_MAP_SWIZZLE_32B;
  CUresult r = enc(tm, eb == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4,
                   const_cast<void*>(ptr), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return b200_fail((int)r, "cuTensorMapEncodeTiled (row map) failed with CUresult %d", (int)r);
  return 0;
}

// Pick the row-segment length that best fills the persistent grid.
ConvGeom make_geom(int B, int H, int W, int nsm, int halo_rows) {
  ConvGeom g;
  g.B = B; g.H = H; g.W = W;
  g.nstrips = (W + kBM - 1) / kBM;
  double best = -1.0;
  int best_rseg = H;
  for (int nseg = 1; nseg <= H; ++nseg) {
    const int rseg = (H + nseg - 1) / nseg;
    if (rseg < 8 && nseg > 1) break;
    const long items = (long)B * g.nstrips * ((H + rseg - 1) / rseg);
    const long waves = (items + nsm - 1) / nsm;
    const double eff = (double)items / (double)(waves * nsm) * (double)rseg / (double)(rseg + halo_rows);
    if (eff > best + 1e-9) { best = eff; best_rseg = rseg; }
  }
  g.rseg = best_rseg;
  g.nseg = (H + g.rseg - 1) / g.rseg;
  g.nitems = B * g.nstrips * g.nseg;
  return g;
}

// 1 = tap-stationary + TMA, 2 = row-stationary + TMA, 3 = row-stationary + cp.async feed,
// 4 = 3 with the dx-replicated weight-gradient kernel (M = 128),
// 5 = 4 with forward / dgrad fed by ONE TMA request per input row into swizzled pixel rows (default)
int g_conv_tc_version = 5;

template <int CIN, int COUT, bool TRANSPOSED, int STATS, bool LDGSTS, typename ET = __nv_bfloat16, bool SWZ = false>
int launch_conv_rs(const void* act, const void* w, const void* bias, void* out, void* partials, int B, int H, int W,
                   cudaStream_t s, int* grid_out, const void* pooled = nullptr, const void* bnprev = nullptr) {
  using Cfg = RsCfg<CIN, COUT, ET>;
  auto kern = conv5x5_rs_kernel<CIN, COUT, TRANSPOSED, STATS, LDGSTS, ET, SWZ>;
  B200_CUDA(set_max_smem(kern, Cfg::kSmemBytes));
  CUtensorMap tm;
  if (SWZ) {
    int rc = make_row_tmap(&tm, act, B, H, W, CIN, (int)sizeof(ET), kPW);
    if (rc) return rc;
  } else if (!LDGSTS) {
    int rc = make_act_tmap(&tm, act, B, H, W, CIN, kPW);
    if (rc) return rc;
  } else {
    memset(&tm, 0, sizeof(tm));          // the cp.async feed never touches the tensor map
  }
  const int nsm = b200_compute_sms();
  ConvGeom g = make_geom(B, H, W, nsm, 4);
  const int grid = g.nitems < nsm ? g.nitems : nsm;
  kern<<<grid, kRsThreads, Cfg::kSmemBytes, s>>>(tm, (const ET*)act, (const float*)w, (const float*)bias,
                                                (ET*)out, (float*)partials, (const ET*)pooled,
                                                (const float*)bnprev, g);
  B200_LAUNCH_CHECK();
  if (grid_out) *grid_out = grid;
  return 0;
}

template <int CIN, int COUT, bool TRANSPOSED, bool STATS>
int launch_conv(const void* act, const void* w, const void* bias, void* out, void* partials, int B, int H, int W,
                cudaStream_t s, int* grid_out) {
  if (g_conv_tc_version >= 5)
    return launch_conv_rs<CIN, COUT, TRANSPOSED, STATS ? 1 : 0, false, __nv_bfloat16, true>(act, w, bias, out, partials, B,
                                                                                            H, W, s, grid_out);
  if (g_conv_tc_version >= 3)
    return launch_conv_rs<CIN, COUT, TRANSPOSED, STATS ? 1 : 0, true>(act, w, bias, out, partials, B, H, W, s, grid_out);
  if (g_conv_tc_version == 2)
    return launch_conv_rs<CIN, COUT, TRANSPOSED, STATS ? 1 : 0, false>(act, w, bias, out, partials, B, H, W, s, grid_out);
  using Cfg = ConvCfg<CIN, COUT>;
  auto kern = conv5x5_tc_kernel<CIN, COUT, TRANSPOSED, STATS>;
  B200_CUDA(set_max_smem(kern, Cfg::kSmemBytes));
  CUtensorMap tm;
  int rc = make_act_tmap(&tm, act, B, H, W, CIN, kPW);
  if (rc) return rc;
  const int nsm = b200_compute_sms();
  ConvGeom g = make_geom(B, H, W, nsm, 4);
  const int grid = g.nitems < nsm ? g.nitems : nsm;
  kern<<<grid, kThreads, Cfg::kSmemBytes, s>>>(tm, (const float*)w, (const float*)bias, (__nv_bfloat16*)out,
                                                (float*)partials, g);
  B200_LAUNCH_CHECK();
  if (grid_out) *grid_out = grid;
  return 0;
}

ConvGeom make_geom_w(int B, int H, int W, int nsm, int halo_rows, int strip) {
  ConvGeom g = make_geom(B, H, (W + strip - 1) / strip * kBM, nsm, halo_rows);   // make_geom counts 128-px strips
  g.W = W;
  return g;
}

template <typename ET>
int launch_wgrad_rep(const void* p1, const void* dx2, void* partials, void* dw2, int B, int H, int W, float gscale,
                     cudaStream_t s) {
  using Cfg = WgRepCfg<ET>;
  B200_CUDA(set_max_smem(conv5x5_wgrad_rep_kernel<ET>, Cfg::kSmemBytes));
  const int nsm = b200_compute_sms();
  ConvGeom g = make_geom_w(B, H, W, nsm, 4, Cfg::kSP);
  const int grid = g.nitems < nsm ? g.nitems : nsm;
  conv5x5_wgrad_rep_kernel<ET><<<grid, kThreads, Cfg::kSmemBytes, s>>>((const ET*)p1, (const ET*)dx2, (float*)partials, g);
  B200_LAUNCH_CHECK();
  return b200_reduce_partials(partials, grid, 32 * 16 * kTaps, dw2, gscale, (void*)s);
}

}  // namespace

// Number of partial-stat rows conv2_fwd writes (= its grid size); callers size
// the partials buffer with b200_conv_tc_max_grid() x 2*COUT floats.
B200_API int b200_conv_tc_max_grid() { return b200_num_sms(); }
B200_API int b200_conv_tc_set_version(int v) {
  if (v < 1 || v > 5) return b200_fail(-2, "conv_tc version must be 1..5");
  g_conv_tc_version = v;
  return 0;
}
B200_API int b200_conv_tc_get_version() { return g_conv_tc_version; }

// *nparts_out receives the number of valid partial-stat rows (= grid size).
B200_API int b200_conv2_fwd_tc(const void* p1_bf16, const void* w2, const void* b2, void* x2_bf16,
                               void* partials, int B, int H, int W, int* nparts_out, void* stream) {
  return launch_conv<16, 32, false, true>(p1_bf16, w2, b2, x2_bf16, partials, B, H, W, (cudaStream_t)stream,
                                          nparts_out);
}

B200_API int b200_conv2_dgrad_tc(const void* dx2_bf16, const void* w2, void* dp1_bf16, int B, int H, int W,
                                 void* stream) {
  return launch_conv<32, 16, true, false>(dx2_bf16, w2, nullptr, dp1_bf16, nullptr, B, H, W, (cudaStream_t)stream,
                                          nullptr);
}

// dgrad with the BatchNorm-1 backward reductions fused into its epilogue (row-stationary kernel only).
// pooled = p1 (may alias dp1_bf16: each pixel is read before it is overwritten), bn1 = [scale|shift|mean|rstd],
// partials: [b200_conv_tc_max_grid()][32] = per-CTA [sum(dy) | sum(dy*xhat)]; *nparts_out = rows written.
B200_API int b200_conv2_dgrad_stats_tc(const void* dx2_bf16, const void* w2, const void* pooled, const void* bn1,
                                       void* dp1_bf16, void* partials, int B, int H, int W, int* nparts_out,
                                       void* stream) {
  if (g_conv_tc_version >= 5)
    return launch_conv_rs<32, 16, true, 2, false, __nv_bfloat16, true>(dx2_bf16, w2, nullptr, dp1_bf16, partials, B, H, W,
                                                                       (cudaStream_t)stream, nparts_out, pooled, bn1);
  if (g_conv_tc_version == 2)
    return launch_conv_rs<32, 16, true, 2, false>(dx2_bf16, w2, nullptr, dp1_bf16, partials, B, H, W,
                                                  (cudaStream_t)stream, nparts_out, pooled, bn1);
  return launch_conv_rs<32, 16, true, 2, true>(dx2_bf16, w2, nullptr, dp1_bf16, partials, B, H, W,
                                               (cudaStream_t)stream, nparts_out, pooled, bn1);
}

// partials: [b200_conv_tc_max_grid()][32*16*25] floats of scratch; dw2: [32][16][5][5].
B200_API int b200_conv2_wgrad_tc(const void* p1_bf16, const void* dx2_bf16, void* partials, void* dw2, int B,
                                 int H, int W, float gscale, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (g_conv_tc_version >= 4)
    return launch_wgrad_rep<__nv_bfloat16>(p1_bf16, dx2_bf16, partials, dw2, B, H, W, gscale, s);
  if (g_conv_tc_version == 3) B200_CUDA(set_max_smem(conv5x5_wgrad_rs_kernel<true>, kWg2SmemBytes));
  else if (g_conv_tc_version == 2) B200_CUDA(set_max_smem(conv5x5_wgrad_rs_kernel<false>, kWg2SmemBytes));
  else B200_CUDA(set_max_smem(conv5x5_wgrad_tc_kernel, kWgSmemBytes));
  CUtensorMap tma, tmb;
  int rc = make_act_tmap(&tma, p1_bf16, B, H, W, 16, kPW);
  if (rc) return rc;
  rc = make_act_tmap(&tmb, dx2_bf16, B, H, W, 32, kBM);
  if (rc) return rc;
  const int nsm = b200_compute_sms();
  ConvGeom g = make_geom(B, H, W, nsm, 4);
  const int grid = g.nitems < nsm ? g.nitems : nsm;
  if (g_conv_tc_version == 3)
    conv5x5_wgrad_rs_kernel<true><<<grid, kThreads, kWg2SmemBytes, s>>>(tma, tmb, (const __nv_bfloat16*)p1_bf16,
                                                                        (const __nv_bfloat16*)dx2_bf16, (float*)partials, g);
  else if (g_conv_tc_version == 2)
    conv5x5_wgrad_rs_kernel<false><<<grid, kThreads, kWg2SmemBytes, s>>>(tma, tmb, (const __nv_bfloat16*)p1_bf16,
                                                                         (const __nv_bfloat16*)dx2_bf16, (float*)partials, g);
  else
    conv5x5_wgrad_tc_kernel<<<grid, kThreads, kWgSmemBytes, s>>>(tma, tmb, (float*)partials, g);
  B200_LAUNCH_CHECK();
  return b200_reduce_partials(partials, grid, 32 * 16 * kTaps, dw2, gscale, stream);
}

// ---- fp32 storage + kind::tf32 (the reference's precision: fp32 tensors, TF32 convolution math) ------------
// Same contracts as the bf16 entry points above with every activation / gradient tensor in fp32.
B200_API int b200_conv2_fwd_tc_f32(const void* p1, const void* w2, const void* b2, void* x2, void* partials, int B,
                                   int H, int W, int* nparts_out, void* stream) {
  if (g_conv_tc_version >= 5)
    return launch_conv_rs<16, 32, false, 1, false, float, true>(p1, w2, b2, x2, partials, B, H, W, (cudaStream_t)stream,
                                                                nparts_out);
  return launch_conv_rs<16, 32, false, 1, true, float>(p1, w2, b2, x2, partials, B, H, W, (cudaStream_t)stream,
                                                       nparts_out);
}
B200_API int b200_conv2_dgrad_tc_f32(const void* dx2, const void* w2, void* dp1, int B, int H, int W, void* stream) {
  if (g_conv_tc_version >= 5)
    return launch_conv_rs<32, 16, true, 0, false, float, true>(dx2, w2, nullptr, dp1, nullptr, B, H, W,
                                                               (cudaStream_t)stream, nullptr);
  return launch_conv_rs<32, 16, true, 0, true, float>(dx2, w2, nullptr, dp1, nullptr, B, H, W, (cudaStream_t)stream,
                                                      nullptr);
}
B200_API int b200_conv2_dgrad_stats_tc_f32(const void* dx2, const void* w2, const void* pooled, const void* bn1,
                                           void* dp1, void* partials, int B, int H, int W, int* nparts_out,
                                           void* stream) {
  if (g_conv_tc_version >= 5)
    return launch_conv_rs<32, 16, true, 2, false, float, true>(dx2, w2, nullptr, dp1, partials, B, H, W,
                                                               (cudaStream_t)stream, nparts_out, pooled, bn1);
  return launch_conv_rs<32, 16, true, 2, true, float>(dx2, w2, nullptr, dp1, partials, B, H, W, (cudaStream_t)stream,
                                                      nparts_out, pooled, bn1);
}
B200_API int b200_conv2_wgrad_tc_f32(const void* p1, const void* dx2, void* partials, void* dw2, int B, int H, int W,
                                     float gscale, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  const int nsm = b200_compute_sms();
  ConvGeom g = make_geom_w(B, H, W, nsm, 4, wt::kSP);
  const int grid = g.nitems < nsm ? g.nitems : nsm;
  CUtensorMap tm;
  if (g_conv_tc_version >= 5) {
    int rc = make_row_tmap(&tm, dx2, B, H, W, 32, 4, wt::kSP, /*atom32=*/true);
    if (rc) return rc;
    B200_CUDA(set_max_smem(conv5x5_wgrad_tf32_kernel<true>, wt::kSmemBytes));
    conv5x5_wgrad_tf32_kernel<true><<<grid, kThreads, wt::kSmemBytes, s>>>(tm, (const float*)p1, (const float*)dx2,
                                                                           (float*)partials, g);
  } else {
    memset(&tm, 0, sizeof(tm));
    B200_CUDA(set_max_smem(conv5x5_wgrad_tf32_kernel<false>, wt::kSmemBytes));
    conv5x5_wgrad_tf32_kernel<false><<<grid, kThreads, wt::kSmemBytes, s>>>(tm, (const float*)p1, (const float*)dx2,
                                                                            (float*)partials, g);
  }
  B200_LAUNCH_CHECK();
  return b200_reduce_partials(partials, grid, 32 * 16 * kTaps, dw2, gscale, stream);
}
