// tcgen05 kernels for conv1 (Conv2d(1, 16, 5, 1, 2), mnist_distributed.py:29) - the layer with
// Cin = 1 that "has no tensor-core shape".  It has one if the GEMM is built differently.
//
// K13, weight gradient  dW1[c][dy][dx] = sum_{b,y,x} g[b,y,x,c] * in[b, y+dy-2, x+dx-2]
// where g, the gradient of the conv1 output, is the pool1/ReLU1/BN1 backward of dp1 and is
// never written to HBM.  The reduction runs over PIXELS, so pixels are the K dimension:
//
//   A (M=128, MN-major)  m = (image row r in 0..7, channel c in 0..15), k = pixel x
//        = g for 8 consecutive image rows; producer warps compute it on CUDA cores from
//        (x1, dp1, BN stats) and store it as bf16 straight in the UMMA layout;
//   B (N=96,  MN-major)  n = (input row r' in 0..11, horizontal tap dxs in 0..7), k = pixel x
//        = the input rows y0-2 .. y0+9 "expanded" so that E[r'][x][dxs] = in[y0-2+r'][x+dxs-2]
//        (a 1-channel row becomes 8 shifted copies: 16 bytes per pixel, cheap);
//   D[(r,c)][(r',dxs)] accumulates over every pixel of every work item of the CTA in ONE
//   128x96 fp32 TMEM tile; dW1[c][dy][dxs] = sum_r D[(r,c)][(r+dy, dxs)] is extracted once.
//
// One tcgen05.mma (M=128, N=96, K=16) contracts 16 pixels x 8 rows; 26 % of its MACs are
// useful, and it still replaces 16*8*16*25 = 51,200 FFMAs with a 56-cycle instruction.  The
// kernel is then bounded by producing g (reading x1 once, 1.44 GB at batch 5) - which is
// what the FFMA version (2.5 ms, csrc/convnet_kernels.cu) could not get near.
#include "common.h"
#include "ptx.cuh"
#include <cuda.h>
#include <cstdint>
#include <cstdlib>
#include <mutex>

namespace {

using namespace b2;

constexpr int kC1 = 16;
constexpr int kXW = 128;                    // pixels per strip
constexpr int kRows = 8;                    // image rows per work item (M = 8 rows x 16 channels)
constexpr int kChunkBytes = kXW * 16;       // one MN-chunk (8 elements) over the strip: 2048 B
constexpr int kABytes = 2 * kRows * kChunkBytes;     // 32 KB
constexpr int kERows = kRows + 4;           // 12 input rows
constexpr int kEBytes = kERows * kChunkBytes;        // 24 KB
constexpr int kStageBytes = kABytes + kEBytes;       // 56 KB
constexpr int kStages = 2;
constexpr int kInW = kXW + 8;               // staged fp32 input row: x0-2 .. x0+133
constexpr int kProducerWarps = 8;
constexpr int kProducers = 32 * kProducerWarps;      // 256 = one 2x2 window each
constexpr int kWgThreads = 32 + kProducers;          // warp 0 = MMA issue
constexpr int kNCols = kERows * 8;          // 96 accumulator columns
constexpr int kTmemCols = 128;
constexpr int kSmemBytes = kStages * kStageBytes + kStages * kERows * kInW * 4 + 1024 + 128;

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
// packed bf16x2 helpers on raw 32-bit words
__device__ __forceinline__ uint32_t bf2_max(uint32_t a, uint32_t b) {
  __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}
__device__ __forceinline__ uint32_t bf2_eq_mask(uint32_t a, uint32_t b) {
  return __heq2_mask(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
}
// one 32-byte pixel (16 bf16 channels) per 256-bit load: half the LSU wavefronts of two 128-bit loads
__device__ __forceinline__ void ldg32(const void* p, uint4& a, uint4& b) {
  asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(p));
}

struct WgGeom {
  int B, H, W;
  int nstrips, ngroups, nitems;
};

__global__ void __launch_bounds__(kWgThreads, 1)
conv1_wgrad_tc_kernel(const float* __restrict__ in, const __nv_bfloat16* __restrict__ x1,
                      const __nv_bfloat16* __restrict__ dp1, const float* __restrict__ bn,
                      const float* __restrict__ red, float* __restrict__ partials, WgGeom g) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  uint8_t* s_stage = smem;                                                    // [kStages][A | E]
  float* s_in = reinterpret_cast<float*>(smem + kStages * kStageBytes);       // [kStages][12][136]
  float4* s_const = reinterpret_cast<float4*>(s_in + kStages * kERows * kInW);  // [16] (sc, sh, ca, cb), 16-B aligned
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_const + kC1);
  uint64_t* full = bars;                 // [kStages] producers -> MMA (count = producer warps)
  uint64_t* empty = bars + kStages;      // [kStages] MMA -> producers
  uint64_t* done = empty + kStages;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kStages; ++i) { mbar_init(&full[i], kProducerWarps); mbar_init(&empty[i], 1); }
    mbar_init(done, 1);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc<kTmemCols>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  const int Ho = g.H / 2, Wo = g.W / 2;

  if (warp == 0) {
    // ================= MMA issue (warp-uniform, one elected lane issues) =================
    constexpr uint32_t idesc = umma_idesc_bf16(128, kNCols, 1, 1);
    const uint64_t tmpl = smem_desc_noswz(0, 128, kChunkBytes);      // MN-major: LBO = 8 pixels, SBO = next chunk
    const uint32_t hi = (uint32_t)(tmpl >> 32), lo0 = (uint32_t)tmpl;
    uint32_t n = 0;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x, ++n) {
      const uint32_t s = n % kStages;
      mbar_wait(&full[s], (n / kStages) & 1);
      tc_fence_after();
      const uint32_t a_lo = lo0 + (smem_u32(s_stage + s * kStageBytes) >> 4);
      const uint32_t b_lo = a_lo + (kABytes >> 4);
      const uint32_t acc0 = n == 0 ? 0u : 1u;
      if (elect_one()) {
#pragma unroll
        for (int kb = 0; kb < kXW / 16; ++kb)
          umma_bf16(tmem_base, ((uint64_t)hi << 32) | (uint64_t)(a_lo + kb * 16),
                    ((uint64_t)hi << 32) | (uint64_t)(b_lo + kb * 16), idesc, kb == 0 ? acc0 : 1u);
        umma_commit(&empty[s]);
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit(done);
    __syncwarp();
  } else {
    // ================= producers: 256 threads, one 2x2 window (x 16 channels) each =================
    // Software-pipelined: the global loads of item n+1 (its window of x1 / dp1 and its share of the
    // input rows) are issued into registers before item n is processed, so the ~1-2 us HBM latency
    // overlaps the CUDA-core work instead of being exposed twice per item.
    const int pt = threadIdx.x - 32;                 // 0..255
    const int wr = pt >> 6, wc = pt & 63;            // window row 0..3, window col 0..63
    for (int c = pt; c < kC1; c += kProducers) {     // per-channel constants: g = ca + cb*x + [argmax&relu] sc*dp
      const float s_ = bn[c], mu = bn[2 * kC1 + c], rs = bn[3 * kC1 + c];
      const float mdy = red[c], mdyx = red[kC1 + c];
      s_const[c] = make_float4(s_, bn[kC1 + c], s_ * (rs * mdyx * mu - mdy), -s_ * rs * mdyx);
    }
    named_bar_sync(1, kProducers);
    constexpr int kInPer = (kERows * kInW + kProducers - 1) / kProducers;   // 7 staged floats per thread

    struct Regs { uint4 xr[4][2], gr[2]; float pin[kInPer]; bool wvalid; };
    Regs nx;                                          // next item's data, in flight
    nx.wvalid = false;
    auto issue_loads = [&](int item) {
      uint4 (&xr)[4][2] = nx.xr; uint4 (&gr)[2] = nx.gr; float (&pin)[kInPer] = nx.pin; bool& wvalid = nx.wvalid;
      const int strip = item % g.nstrips;
      const int grp = (item / g.nstrips) % g.ngroups;
      const int b = item / (g.nstrips * g.ngroups);
      const int y0 = grp * kRows, x0 = strip * kXW;
#pragma unroll
      for (int j = 0; j < kInPer; ++j) {
        const int i = pt + j * kProducers;
        const int rr = i / kInW, cc = i - rr * kInW;
        const int gy = y0 - 2 + rr, gx = x0 - 2 + cc;
        pin[j] = (i < kERows * kInW && gy >= 0 && gy < g.H && gx >= 0 && gx < g.W)
                     ? __ldg(in + ((size_t)b * g.H + gy) * g.W + gx) : 0.f;
      }
      const int gy = y0 + 2 * wr, gx = x0 + 2 * wc;
      wvalid = gy < g.H && gx < g.W;
      if (wvalid) {
        const __nv_bfloat16* xb = x1 + (((size_t)b * g.H + gy) * g.W + gx) * kC1;
        ldg32(xb, xr[0][0], xr[0][1]);
        ldg32(xb + kC1, xr[1][0], xr[1][1]);
        ldg32(xb + (size_t)g.W * kC1, xr[2][0], xr[2][1]);
        ldg32(xb + (size_t)g.W * kC1 + kC1, xr[3][0], xr[3][1]);
        const __nv_bfloat16* gb = dp1 + (((size_t)b * Ho + (gy >> 1)) * Wo + (gx >> 1)) * kC1;
        ldg32(gb, gr[0], gr[1]);
      }
    };
    if ((int)blockIdx.x < g.nitems) issue_loads(blockIdx.x);
    uint32_t n = 0;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x, ++n) {
      const uint32_t s = n % kStages;
      uint8_t* sA = s_stage + s * kStageBytes;
      uint8_t* sE = sA + kABytes;
      float* sI = s_in + s * kERows * kInW;
      const Regs cur = nx;                            // (0) take this item's data, put the next item's loads in flight
      if (item + (int)gridDim.x < g.nitems) issue_loads(item + gridDim.x);
      const uint4 (&xr)[4][2] = cur.xr; const uint4 (&gr)[2] = cur.gr; const float (&pin)[kInPer] = cur.pin;
      const bool wvalid = cur.wvalid;
      mbar_wait(&empty[s], ((n / kStages) & 1) ^ 1);
      // (1) staged input rows: registers -> smem
#pragma unroll
      for (int j = 0; j < kInPer; ++j) {
        const int i = pt + j * kProducers;
        if (i < kERows * kInW) sI[i] = pin[j];
      }
      // (2) g for this thread's window (from the prefetched registers) -> A[(2r+pl)][k][e]
      // Two channels per instruction where the math allows it.  A 32-bit word of x1 holds channels (2p, 2p+1):
      // the 2x2 arg-max runs on packed bf16 (sign-flipped where the BN scale is negative, so it is always a
      // max; first maximum in row-major order like ATen) instead of four fp32 FMAs + compare/select chains per
      // channel; only the ReLU test, the routed gradient and the dense BN term use fp32.
      {
        uint32_t o[4][kC1 / 2];
#pragma unroll
        for (int cp = 0; cp < kC1 / 2; ++cp) {
          const float4 k0 = s_const[2 * cp], k1 = s_const[2 * cp + 1];      // (sc, sh, ca, cb)
          const uint32_t sgn = ((__float_as_uint(k0.x) >> 16) & 0x8000u) | (__float_as_uint(k1.x) & 0x80000000u);
          uint32_t xw[4], f[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            xw[i] = (&xr[i][cp >> 2].x)[cp & 3];
            f[i] = xw[i] ^ sgn;
          }
          const uint32_t m = bf2_max(bf2_max(f[0], f[1]), bf2_max(f[2], f[3]));
          const uint32_t e0 = bf2_eq_mask(f[0], m), e1 = bf2_eq_mask(f[1], m), e2 = bf2_eq_mask(f[2], m);
          uint32_t sel[4];
          sel[0] = e0;
          sel[1] = e1 & ~e0;
          sel[2] = e2 & ~(e0 | e1);
          sel[3] = ~(e0 | e1 | e2);                      // m is one of the four: if none of the first three, the last
          const uint32_t mx = m ^ sgn;                   // the pre-BN value at the arg-max
          const uint32_t gw = (&gr[cp >> 2].x)[cp & 3];
          const float g0 = fmaf(bf16_lo(mx), k0.x, k0.y) > 0.f ? bf16_lo(gw) * k0.x : 0.f;   // ReLU mask, dy * scale
          const float g1 = fmaf(bf16_hi(mx), k1.x, k1.y) > 0.f ? bf16_hi(gw) * k1.x : 0.f;
          // value at the arg-max pixel: dense term + routed gradient, summed in fp32 and rounded ONCE.  (Adding
          // two already-rounded bf16 numbers loses the small dense term at every arg-max pixel the same way,
          // a bias that grows with the pixel count and BN backward's cancellation then amplifies: 10 % on dW1.)
          const uint32_t r2 = pack_bf16x2(fmaf(k0.w, bf16_lo(mx), k0.z) + g0, fmaf(k1.w, bf16_hi(mx), k1.z) + g1);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const uint32_t d2 = pack_bf16x2(fmaf(k0.w, bf16_lo(xw[i]), k0.z), fmaf(k1.w, bf16_hi(xw[i]), k1.z));
            o[i][cp] = wvalid ? ((d2 & ~sel[i]) | (r2 & sel[i])) : 0u;
          }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int r = 2 * wr + (i >> 1), k = 2 * wc + (i & 1);
          uint8_t* dst = sA + (2 * r) * kChunkBytes + k * 16;
          *reinterpret_cast<uint4*>(dst) = make_uint4(o[i][0], o[i][1], o[i][2], o[i][3]);
          *reinterpret_cast<uint4*>(dst + kChunkBytes) = make_uint4(o[i][4], o[i][5], o[i][6], o[i][7]);
        }
      }
      named_bar_sync(1, kProducers);                 // staged input complete
      // (4) expanded input E[r'][k][dxs]: 12 x 128 pieces of 16 B, 6 per thread
      {
        const int k = pt & 127;
#pragma unroll
        for (int rr = pt >> 7; rr < kERows; rr += 2) {
          const float* src = sI + rr * kInW + k;     // in[.., x0 + k - 2 + dxs]
          uint4 o;
          o.x = pack_bf16x2(src[0], src[1]); o.y = pack_bf16x2(src[2], src[3]);
          o.z = pack_bf16x2(src[4], src[5]); o.w = pack_bf16x2(src[6], src[7]);
          *reinterpret_cast<uint4*>(sE + rr * kChunkBytes + k * 16) = o;
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&full[s]);
    }
  }
  // ================= extraction: dW1[c][dy][dxs] = sum_r D[(r,c)][(r+dy)*8 + dxs] =================
  float* s_T = reinterpret_cast<float*>(s_stage);    // [128][97] floats, aliases the (now idle) stages
  constexpr int kTP = kNCols + 1;
  if (warp >= 1 && warp <= 4) {
    mbar_wait(done, 0);
    tc_fence_after();
    const int q = warp & 3, m = q * 32 + lane;
#pragma unroll 1
    for (int c0 = 0; c0 < kNCols; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + c0, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; ++j) s_T[m * kTP + c0 + j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  for (int t = threadIdx.x; t < kC1 * 25; t += kWgThreads) {
    const int c = t / 25, tap = t % 25, dy = tap / 5, dxs = tap % 5;
    const int pl = c >> 3, e = c & 7;
    float a = 0.f;
#pragma unroll
    for (int r = 0; r < kRows; ++r) a += s_T[((2 * r + pl) * 8 + e) * kTP + (r + dy) * 8 + dxs];
    partials[(size_t)blockIdx.x * 400 + t] = a;
  }
  if (warp == 0) { __syncwarp(); tmem_dealloc<kTmemCols>(tmem_base); }
}


// ------------------------------------------------------------------------------------------------
// The same GEMM for fp32 storage (precision "tf32"): x1 and dp1 are fp32, the conv1-output gradient g
// and the expanded input E are written to shared memory as fp32 rounded to TF32, and the contraction
// runs as kind::tf32 MMAs (K = 8 pixels).  MN-major TF32 operands have one legal shared-memory layout
// (128-byte pixel rows of 32 MN elements, 32-byte units swizzled by the row index: sw32_off in ptx.cuh),
// and a strip is 64 pixels wide to keep the tiles the size of the bf16 kernel's:
//   A (M=128)  m = 16*r + c:    atom r/2,  unit 2*(r%2) + c/8   [atom][pixel k][128 B]   4 x 8 KB
//   B (N=96)   n = 8*r' + dxs:  atom r'/4, unit r'%4            [atom][pixel k][128 B]   3 x 8 KB
// D has the same (m, n) meaning as above, so the extraction is shared.  Producers work per
// (2x2 window, 8 channels) in plain fp32: y = sc*x + sh, first maximum in row-major order, ReLU mask,
// g = ca + cb*x (+ sc*dp at the arg-max).
namespace tf {
constexpr int kXW = 64;
constexpr int kAtom = kXW * 128;                      // 8 KB: 32 MN elements over the strip
constexpr int kABytes = 4 * kAtom;                    // 32 KB
constexpr int kEBytes = 3 * kAtom;                    // 24 KB
constexpr int kStageBytes = kABytes + kEBytes;
constexpr int kInW = kXW + 8;                         // staged fp32 input row: x0-2 .. x0+69
// 16 producer warps: one thread = one 2x2 window x 4 channels.  Its five 16-byte pieces (4 pixels of x1, 1 of dp1)
// reach it through a private two-slot cp.async ring, refilled for item n+2 as soon as item n has been read into
// registers, so two items of HBM latency are covered without holding a second item in registers.  (The first
// version prefetched one item ahead into registers: ncu showed 28 % of the stall samples on the first use of
// those registers - 148 SMs bursting 45 KB each at the same moment see ~2 us of latency, one item time.)
constexpr int kProdWarps = 16;
constexpr int kProd = 32 * kProdWarps;                // 512
constexpr int kThreads = 32 + kProd;
constexpr int kRawBytes = 5 * kProd * 16;             // one slot of the x1 / dp1 ring: 40 KB
constexpr int kInRing = 4;                            // input-row ring (shared by all producers; see the kernel)
constexpr int kInBytes = kERows * kInW * 4;           // 3456 B
constexpr int kSmemBytes = kStages * kStageBytes + 2 * kRawBytes + kInRing * kInBytes + 1024 + 1024;
}  // namespace tf

__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(src_bytes) : "memory");
}

__device__ __forceinline__ void ldg32f(const float* p, float (&v)[8]) {
  uint32_t u[8];
  asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]) : "l"(p));
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(u[i]);
}

__global__ void __launch_bounds__(tf::kThreads, 1)
conv1_wgrad_tf32_kernel(const float* __restrict__ in, const float* __restrict__ x1,
                        const float* __restrict__ dp1, const float* __restrict__ bn,
                        const float* __restrict__ red, float* __restrict__ partials, WgGeom g) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);    // swizzle atoms: 1 KB aligned
  uint8_t* s_stage = smem;                                                          // [kStages][A | E]
  uint8_t* s_rawx = smem + kStages * tf::kStageBytes;                               // [2][5 pieces][512 threads][16 B]
  float* s_in = reinterpret_cast<float*>(s_rawx + 2 * tf::kRawBytes);               // [kInRing][12][72]
  float4* s_const = reinterpret_cast<float4*>(s_in + tf::kInRing * kERows * tf::kInW);  // [16] (sc, sh, ca, cb)
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_const + kC1);
  uint64_t* full = bars;
  uint64_t* empty = bars + kStages;
  uint64_t* done = empty + kStages;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kStages; ++i) { mbar_init(&full[i], tf::kProdWarps); mbar_init(&empty[i], 1); }
    mbar_init(done, 1);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc<kTmemCols>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  const int Ho = g.H / 2, Wo = g.W / 2;

  if (warp == 0) {
    constexpr uint32_t idesc = umma_idesc(2, 128, kNCols, 1, 1);
    const uint64_t tmpl = smem_desc(0, tf::kAtom, 512, 1);           // MN-major SW128_32B: LBO = next atom, SBO = next 4 rows
    const uint32_t hi = (uint32_t)(tmpl >> 32), lo0 = (uint32_t)tmpl;
    uint32_t n = 0;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x, ++n) {
      const uint32_t s = n % kStages;
      mbar_wait(&full[s], (n / kStages) & 1);
      tc_fence_after();
      const uint32_t a_lo = lo0 + (smem_u32(s_stage + s * tf::kStageBytes) >> 4);
      const uint32_t b_lo = a_lo + (tf::kABytes >> 4);
      const uint32_t acc0 = n == 0 ? 0u : 1u;
      if (elect_one()) {
#pragma unroll
        for (int kb = 0; kb < tf::kXW / 8; ++kb)
          umma_tf32(tmem_base, ((uint64_t)hi << 32) | (uint64_t)(a_lo + kb * 64),
                    ((uint64_t)hi << 32) | (uint64_t)(b_lo + kb * 64), idesc, kb == 0 ? acc0 : 1u);
        umma_commit(&empty[s]);
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit(done);
    __syncwarp();
  } else {
    const int pt = threadIdx.x - 32;                 // 0..511
    const int q = pt & 3, w = pt >> 2;               // channels 4q..4q+3 of window w
    const int wr = w >> 5, wc = w & 31;              // window row 0..3, window col 0..31
    for (int c = pt; c < kC1; c += tf::kProd) {
      const float s_ = bn[c], mu = bn[2 * kC1 + c], rs = bn[3 * kC1 + c];
      const float mdy = red[c], mdyx = red[kC1 + c];
      s_const[c] = make_float4(s_, bn[kC1 + c], s_ * (rs * mdyx * mu - mdy), -s_ * rs * mdyx);
    }
    constexpr int kInTot = kERows * tf::kInW;                               // 864
    constexpr int kInPer = (kInTot + tf::kProd - 1) / tf::kProd;            // 2 staged floats per thread
    const int stride = (int)gridDim.x;
    // cp.async fetch of the n-th item of this CTA: own window pieces -> private slot n & 1, own share of the
    // input rows -> ring slot n & 3.  Always commits a group (possibly empty) so that wait_group counts items.
    // (strip, group, image) of the next item to fetch, advanced by `stride` items without divisions
    int f_strip = (int)blockIdx.x % g.nstrips, f_grp = ((int)blockIdx.x / g.nstrips) % g.ngroups,
        f_b = (int)blockIdx.x / (g.nstrips * g.ngroups);
    const int d_strip = stride % g.nstrips, d_grp = (stride / g.nstrips) % g.ngroups,
              d_b = stride / (g.nstrips * g.ngroups);
    uint32_t vq = 0;                                 // bit j: this thread's window is inside the image for item n+j
    auto fetch = [&](int item, uint32_t n, int ahead) {
      if (item < g.nitems) {
        const int strip = f_strip, grp = f_grp, b = f_b;
        f_strip += d_strip; f_grp += d_grp; f_b += d_b;
        if (f_strip >= g.nstrips) { f_strip -= g.nstrips; ++f_grp; }
        if (f_grp >= g.ngroups) { f_grp -= g.ngroups; ++f_b; }
        const int y0 = grp * kRows, x0 = strip * tf::kXW;
        float* sI = s_in + (n & (tf::kInRing - 1)) * kInTot;
#pragma unroll
        for (int j = 0; j < kInPer; ++j) {
          const int i = pt + j * tf::kProd;
          if (i < kInTot) {
            const int rr = i / tf::kInW, cc = i - rr * tf::kInW;
            const int gy = y0 - 2 + rr, gx = x0 - 2 + cc;
            const bool ok = gy >= 0 && gy < g.H && gx >= 0 && gx < g.W;
            cp_async4(sI + i, in + (ok ? ((size_t)b * g.H + gy) * g.W + gx : 0), ok ? 4u : 0u);
          }
        }
        const int gy = y0 + 2 * wr, gx = x0 + 2 * wc;
        const uint32_t nb = (gy < g.H && gx < g.W) ? 16u : 0u;
        vq |= (nb ? 1u : 0u) << ahead;
        const float* xb = x1 + (nb ? (((size_t)b * g.H + gy) * g.W + gx) * kC1 : 0) + q * 4;
        uint8_t* slot = s_rawx + (n & 1) * tf::kRawBytes + pt * 16;
        cp_async16(slot, xb, nb);
        cp_async16(slot + tf::kProd * 16, xb + kC1, nb);
        cp_async16(slot + 2 * tf::kProd * 16, xb + (size_t)g.W * kC1, nb);
        cp_async16(slot + 3 * tf::kProd * 16, xb + (size_t)g.W * kC1 + kC1, nb);
        cp_async16(slot + 4 * tf::kProd * 16,
                   dp1 + (nb ? (((size_t)b * Ho + (gy >> 1)) * Wo + (gx >> 1)) * kC1 : 0) + q * 4, nb);
      }
      cp_async_commit();
    };
    fetch(blockIdx.x, 0, 0);
    fetch(blockIdx.x + stride, 1, 1);
    named_bar_sync(1, tf::kProd);                    // s_const visible
    float4 kc[4];                                    // this thread's channel constants stay in registers
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) kc[cc] = s_const[q * 4 + cc];
    uint32_t n = 0;
    for (int item = blockIdx.x; item < g.nitems; item += stride, ++n) {
      const uint32_t s = n % kStages;
      uint8_t* sA = s_stage + s * tf::kStageBytes;
      uint8_t* sE = sA + tf::kABytes;
      const float* sI = s_in + (n & (tf::kInRing - 1)) * kInTot;
      cp_async_wait<1>();                            // this thread's pieces of item n have landed
      float4 xr[4], gr;
      {
        const uint8_t* slot = s_rawx + (n & 1) * tf::kRawBytes + pt * 16;
#pragma unroll
        for (int i = 0; i < 4; ++i) xr[i] = *reinterpret_cast<const float4*>(slot + i * tf::kProd * 16);
        gr = *reinterpret_cast<const float4*>(slot + 4 * tf::kProd * 16);
      }
      // refill the slot just read with item n+2.  Its input rows go to ring slot (n+2) & 3, last read in the E build
      // of item n-2: every thread is past that, because this thread has passed item n-1's barrier below.
      fetch(item + 2 * stride, n + 2, 2);
      const bool wvalid = vq & 1u;
      vq >>= 1;
      mbar_wait(&empty[s], ((n / kStages) & 1) ^ 1);
      {
        float o[4][4];
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          const float4 k = kc[cc];                                // (sc, sh, ca, cb)
          const float xv[4] = {(&xr[0].x)[cc], (&xr[1].x)[cc], (&xr[2].x)[cc], (&xr[3].x)[cc]};
          float best = fmaf(xv[0], k.x, k.y);
          int am = 0;
#pragma unroll
          for (int i = 1; i < 4; ++i) {
            const float y = fmaf(xv[i], k.x, k.y);
            if (y > best) { best = y; am = i; }
          }
          const float gq = best > 0.f ? (&gr.x)[cc] * k.x : 0.f;  // ReLU mask, dy * scale, routed to the arg-max
          // stored as fp32; the tensor core reads the top 19 bits (cvt.rna.tf32 is a 5-instruction emulation here)
#pragma unroll
          for (int i = 0; i < 4; ++i) o[i][cc] = fmaf(k.w, xv[i], k.z);
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (am == i) o[i][cc] += gq;
        }
        if (!wvalid) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) o[i][cc] = 0.f;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int kx = 2 * wc + (i & 1);                        // image row r = 2*wr + (i >> 1): atom wr
          *reinterpret_cast<float4*>(sA + wr * tf::kAtom + sw32_off(kx, 2 * (i >> 1) + (q >> 1)) + 16 * (q & 1)) =
              make_float4(o[i][0], o[i][1], o[i][2], o[i][3]);
        }
      }
      named_bar_sync(1, tf::kProd);                  // everyone's input-row pieces of item n have landed
      // expanded input E[r'][k][dxs]: 12 rows x 64 pixels x 2 pieces of 16 B, 3 per thread
#pragma unroll
      for (int j = 0; j < (2 * kERows * tf::kXW) / tf::kProd; ++j) {
        const int idx = pt + j * tf::kProd;
        const int kx = idx & (tf::kXW - 1), c2 = idx >> 6, rr = c2 >> 1;
        const float* src = sI + rr * tf::kInW + kx + 4 * (c2 & 1);          // in[.., x0 + kx - 2 + dxs]
        *reinterpret_cast<float4*>(sE + (rr >> 2) * tf::kAtom + sw32_off(kx, rr & 3) + 16 * (c2 & 1)) =
            make_float4(src[0], src[1], src[2], src[3]);
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&full[s]);
    }
    cp_async_wait<0>();
  }
  // extraction: identical (m, n) meaning to the bf16 kernel
  float* s_T = reinterpret_cast<float*>(s_stage);
  constexpr int kTP = kNCols + 1;
  if (warp >= 1 && warp <= 4) {
    mbar_wait(done, 0);
    tc_fence_after();
    const int q = warp & 3, m = q * 32 + lane;
#pragma unroll 1
    for (int c0 = 0; c0 < kNCols; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + c0, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; ++j) s_T[m * kTP + c0 + j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  for (int t = threadIdx.x; t < kC1 * 25; t += tf::kThreads) {
    const int c = t / 25, tap = t % 25, dy = tap / 5, dxs = tap % 5;
    float a = 0.f;
#pragma unroll
    for (int r = 0; r < kRows; ++r) a += s_T[(16 * r + c) * kTP + (r + dy) * 8 + dxs];
    partials[(size_t)blockIdx.x * 400 + t] = a;
  }
  if (warp == 0) { __syncwarp(); tmem_dealloc<kTmemCols>(tmem_base); }
}


// ------------------------------------------------------------------------------------------------
// Same kernel with every global load done by TMA (one elected thread, three requests per item) instead of
// 7 cp.async per producer thread: in the cp.async version the L1 data pipe was ~80 % busy, 1400 of its ~3200
// wavefronts per item being LDGSTS sectors.  Per item the loader warp fetches into a two-stage raw ring
//   x1 tile   8 rows x 64 px x 16 ch fp32 = 32 KB  - map {32 floats = 2 pixels, W/2, H, B}, 128-byte swizzle: the eight
//             windows a warp reads at once are eight consecutive 128-byte lines, the swizzle spreads their 16-byte
//             pieces over all banks (4 wavefronts per LDS.128 = the minimum for 512 bytes);
//   dp1 tile  4 x 32 pooled px x 16 ch = 8 KB     - map {16, W/2, H/2, B};
//   input     12 rows x 72 px fp32                - map {W, H, B}; out-of-image rows / columns arrive as zeros.
// Producers wait on the stage's mbarrier, so the per-item named barrier is gone too.
namespace tt {
constexpr int kXBytes = 8 * tf::kXW * 64;             // 32768
constexpr int kGBytes = 4 * (tf::kXW / 2) * 64;       // 8192
constexpr int kInW = 80;                              // input row x0-4 .. x0+75: the tile starts 16-byte aligned in global memory
constexpr int kIBytes = kERows * kInW * 4;            // 3840
constexpr int kRawStage = 45056;                      // 44 KB (1 KB multiple: the swizzled tile needs 1 KB alignment)
constexpr int kRawStages = 2;
constexpr int kThreads = 64 + tf::kProd;              // warp 0 MMA issue, warp 1 TMA loader, 16 producer warps
constexpr int kSmemBytes = kStages * tf::kStageBytes + kRawStages * kRawStage + 1024 + 1024;
static_assert(kXBytes + kGBytes + kIBytes <= kRawStage, "raw stage layout");
}  // namespace tt

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
      : "memory");
}

__global__ void __launch_bounds__(tt::kThreads, 1)
conv1_wgrad_tf32_tma_kernel(const __grid_constant__ CUtensorMap tm_x1, const __grid_constant__ CUtensorMap tm_dp1,
                            const __grid_constant__ CUtensorMap tm_in, const float* __restrict__ bn,
                            const float* __restrict__ red, float* __restrict__ partials, WgGeom g) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* s_stage = smem;                                                          // [kStages][A | E]
  uint8_t* s_rawb = smem + kStages * tf::kStageBytes;                               // [2][x1 | dp1 | in]
  float4* s_const = reinterpret_cast<float4*>(s_rawb + tt::kRawStages * tt::kRawStage);   // [16] (sc, sh, ca, cb)
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_const + kC1);
  uint64_t* full = bars;                       // [kStages]  producers -> MMA
  uint64_t* empty = full + kStages;            // [kStages]  MMA -> producers
  uint64_t* rfull = empty + kStages;           // [2]        TMA -> producers
  uint64_t* rempty = rfull + tt::kRawStages;   // [2]        producers -> loader
  uint64_t* done = rempty + tt::kRawStages;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kStages; ++i) { mbar_init(&full[i], tf::kProdWarps); mbar_init(&empty[i], 1); }
    for (int i = 0; i < tt::kRawStages; ++i) { mbar_init(&rfull[i], 1); mbar_init(&rempty[i], tf::kProdWarps); }
    mbar_init(done, 1);
    mbar_fence_init();
    tma_prefetch_desc(&tm_x1); tma_prefetch_desc(&tm_dp1); tma_prefetch_desc(&tm_in);
  }
  for (int c = threadIdx.x; c < kC1; c += tt::kThreads) {
    const float s_ = bn[c], mu = bn[2 * kC1 + c], rs = bn[3 * kC1 + c];
    const float mdy = red[c], mdyx = red[kC1 + c];
    s_const[c] = make_float4(s_, bn[kC1 + c], s_ * (rs * mdyx * mu - mdy), -s_ * rs * mdyx);
  }
  if (warp == 0) tmem_alloc<kTmemCols>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ================= MMA issue =================
    constexpr uint32_t idesc = umma_idesc(2, 128, kNCols, 1, 1);
    const uint64_t tmpl = smem_desc(0, tf::kAtom, 512, 1);
    const uint32_t hi = (uint32_t)(tmpl >> 32), lo0 = (uint32_t)tmpl;
    uint32_t n = 0;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x, ++n) {
      const uint32_t s = n % kStages;
      mbar_wait(&full[s], (n / kStages) & 1);
      tc_fence_after();
      const uint32_t a_lo = lo0 + (smem_u32(s_stage + s * tf::kStageBytes) >> 4);
      const uint32_t b_lo = a_lo + (tf::kABytes >> 4);
      const uint32_t acc0 = n == 0 ? 0u : 1u;
      if (elect_one()) {
#pragma unroll
        for (int kb = 0; kb < tf::kXW / 8; ++kb)
          umma_tf32(tmem_base, ((uint64_t)hi << 32) | (uint64_t)(a_lo + kb * 64),
                    ((uint64_t)hi << 32) | (uint64_t)(b_lo + kb * 64), idesc, kb == 0 ? acc0 : 1u);
        umma_commit(&empty[s]);
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit(done);
    __syncwarp();
  } else if (warp == 1) {
    // ================= TMA loader =================
    if (lane == 0) {
      uint32_t n = 0;
      for (int item = blockIdx.x; item < g.nitems; item += gridDim.x, ++n) {
        const int strip = item % g.nstrips;
        const int grp = (item / g.nstrips) % g.ngroups;
        const int b = item / (g.nstrips * g.ngroups);
        const int y0 = grp * kRows, x0 = strip * tf::kXW;
        const uint32_t rs = n % tt::kRawStages;
        mbar_wait(&rempty[rs], ((n / tt::kRawStages) & 1) ^ 1);
        uint8_t* dst = s_rawb + rs * tt::kRawStage;
        mbar_expect_tx(&rfull[rs], tt::kXBytes + tt::kGBytes + tt::kIBytes);
        tma_load_4d(dst, &tm_x1, &rfull[rs], 0, x0 >> 1, y0, b);
        tma_load_4d(dst + tt::kXBytes, &tm_dp1, &rfull[rs], 0, x0 >> 1, y0 >> 1, b);
        tma_load_3d(dst + tt::kXBytes + tt::kGBytes, &tm_in, &rfull[rs], x0 - 4, y0 - 2, b);
      }
    }
  } else {
    // ================= producers: 512 threads, one 2x2 window x 4 channels each =================
    const int pt = threadIdx.x - 64;                 // 0..511
    const int q = pt & 3, w = pt >> 2;               // channels 4q..4q+3 of window w
    const int wr = w >> 5, wc = w & 31;              // window row 0..3, window col 0..31
    float4 kc[4];
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) kc[cc] = s_const[q * 4 + cc];
    constexpr int kInTot = kERows * tf::kInW;
    const int stride = (int)gridDim.x;
    // (strip, group) of the current item, advanced without divisions
    int c_strip = (int)blockIdx.x % g.nstrips, c_grp = ((int)blockIdx.x / g.nstrips) % g.ngroups;
    const int d_strip = stride % g.nstrips, d_grp = (stride / g.nstrips) % g.ngroups;
    // byte offsets of this thread's pieces inside a raw stage (x1 tile: 128-byte swizzle, chunk ^= line index)
    uint32_t xoff[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
      xoff[i] = (uint32_t)((2 * wr + (i >> 1)) * (tf::kXW * 64) + wc * 128 + ((((i & 1) * 4 + q) ^ (wc & 7)) << 4));
    const uint32_t goff = (uint32_t)(tt::kXBytes + wr * (tf::kXW / 2) * 64 + wc * 64 + q * 16);
    uint32_t n = 0;
    for (int item = blockIdx.x; item < g.nitems; item += stride, ++n) {
      const uint32_t s = n % kStages, rs = n % tt::kRawStages;
      uint8_t* sA = s_stage + s * tf::kStageBytes;
      uint8_t* sE = sA + tf::kABytes;
      const uint8_t* raw = s_rawb + rs * tt::kRawStage;
      const float* sI = reinterpret_cast<const float*>(raw + tt::kXBytes + tt::kGBytes);
      const bool wvalid = (c_grp * kRows + 2 * wr) < g.H && (c_strip * tf::kXW + 2 * wc) < g.W;
      c_strip += d_strip; c_grp += d_grp;
      if (c_strip >= g.nstrips) { c_strip -= g.nstrips; ++c_grp; }
      if (c_grp >= g.ngroups) c_grp -= g.ngroups;
      mbar_wait(&rfull[rs], (n / tt::kRawStages) & 1);
      float4 xr[4], gr;
#pragma unroll
      for (int i = 0; i < 4; ++i) xr[i] = *reinterpret_cast<const float4*>(raw + xoff[i]);
      gr = *reinterpret_cast<const float4*>(raw + goff);
      mbar_wait(&empty[s], ((n / kStages) & 1) ^ 1);
      {
        float o[4][4];
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          const float4 k = kc[cc];                                // (sc, sh, ca, cb)
          const float xv[4] = {(&xr[0].x)[cc], (&xr[1].x)[cc], (&xr[2].x)[cc], (&xr[3].x)[cc]};
          float best = fmaf(xv[0], k.x, k.y);
          int am = 0;
#pragma unroll
          for (int i = 1; i < 4; ++i) {
            const float y = fmaf(xv[i], k.x, k.y);
            if (y > best) { best = y; am = i; }
          }
          const float gq = best > 0.f ? (&gr.x)[cc] * k.x : 0.f;  // ReLU mask, dy * scale, routed to the arg-max
#pragma unroll
          for (int i = 0; i < 4; ++i) o[i][cc] = fmaf(k.w, xv[i], k.z);
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (am == i) o[i][cc] += gq;
        }
        if (!wvalid) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) o[i][cc] = 0.f;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int kx = 2 * wc + (i & 1);                        // image row r = 2*wr + (i >> 1): atom wr
          *reinterpret_cast<float4*>(sA + wr * tf::kAtom + sw32_off(kx, 2 * (i >> 1) + (q >> 1)) + 16 * (q & 1)) =
              make_float4(o[i][0], o[i][1], o[i][2], o[i][3]);
        }
      }
      // expanded input E[r'][k][dxs]: 12 rows x 64 pixels x 2 pieces of 16 B, 3 per thread
#pragma unroll
      for (int j = 0; j < (2 * kERows * tf::kXW) / tf::kProd; ++j) {
        const int idx = pt + j * tf::kProd;
        const int kx = idx & (tf::kXW - 1), c2 = idx >> 6, rr = c2 >> 1;
        const float* src = sI + rr * tt::kInW + kx + 2 + 4 * (c2 & 1);      // in[.., x0 + kx - 2 + dxs]
        *reinterpret_cast<float4*>(sE + (rr >> 2) * tf::kAtom + sw32_off(kx, rr & 3) + 16 * (c2 & 1)) =
            make_float4(src[0], src[1], src[2], src[3]);
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) { mbar_arrive(&rempty[rs]); mbar_arrive(&full[s]); }
      (void)kInTot;
    }
  }
  // extraction: identical (m, n) meaning to the bf16 kernel
  float* s_T = reinterpret_cast<float*>(s_stage);
  constexpr int kTP = kNCols + 1;
  if (warp >= 2 && warp <= 5) {
    mbar_wait(done, 0);
    tc_fence_after();
    const int qq = warp & 3, m = qq * 32 + lane;
#pragma unroll 1
    for (int c0 = 0; c0 < kNCols; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(tmem_base + ((uint32_t)(qq * 32) << 16) + c0, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; ++j) s_T[m * kTP + c0 + j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  for (int t = threadIdx.x; t < kC1 * 25; t += tt::kThreads) {
    const int c = t / 25, tap = t % 25, dy = tap / 5, dxs = tap % 5;
    float a = 0.f;
#pragma unroll
    for (int r = 0; r < kRows; ++r) a += s_T[(16 * r + c) * kTP + (r + dy) * 8 + dxs];
    partials[(size_t)blockIdx.x * 400 + t] = a;
  }
  if (warp == 0) { __syncwarp(); tmem_dealloc<kTmemCols>(tmem_base); }
}


// ------------------------------------------------------------------------------------------------
// Layer-1 weight gradient WITHOUT the conv1 output.  g = ca + cb*x1 + routed, so
//   dW1[c][tap] = sum_p routed[p,c] * in[p+tap]                      (tensor cores, as above: D1 = A x E)
//               + ca_c * S[tap] + cb_c * R[c][tap],   S[tap] = sum_p in[p+tap],   R[c][tap] = sum_p x1[p,c] * in[p+tap]
// and since x1 = b + sum_tap' w[tap'] * in[p+tap'],   R[c][tap] = b_c * S[tap] + sum_tap' w[c][tap'] * M[tap'][tap]
// with the 25 x 25 Gram matrix M[tap'][tap] = sum_p in[p+tap'] * in[p+tap] of the zero-padded input.  M and S come out
// of ONE more MMA per K-step on the operand that is already in shared memory: D2 = E4 x E, where E4 is E plus a row of
// ones.  So the kernel reads dp1 (0.72 GB), the forward's arg-max/ReLU bytes (0.18 GB) and the input (0.18 GB) instead
// of x1 (2.88 GB), the producers only route (mask byte -> one non-zero of four), and the forward never stores x1.
// Needs H % 8 == 0 (a partial row group would put products of real input rows with no output pixel into M; partial
// strips are handled by zeroing the out-of-image columns of E).  Per-CTA partial row: [400 routed | 625 M | 25 S].
namespace tn {
constexpr int kEBytes = 4 * tf::kAtom;                // E4: three atoms of E + the "ones" atom (32 KB)
constexpr int kStageBytes = tf::kABytes + kEBytes;    // 64 KB
constexpr int kGBytes = 4 * (tf::kXW / 2) * 64;       // dp1 tile   8192
constexpr int kMBytes = 4 * (tf::kXW / 2) * 16;       // mask tile  2048
constexpr int kIBytes = kERows * tt::kInW * 4;        // input rows 3840
constexpr int kRawStage = 14336;
constexpr int kRawStages = 3;
constexpr int kThreads = 64 + tf::kProd;
constexpr int kPartial = 400 + 625 + 25;
constexpr int kSmemBytes = kStages * kStageBytes + kRawStages * kRawStage + 1024 + 1024;
static_assert(kGBytes + kMBytes + kIBytes <= kRawStage, "raw stage layout");
}  // namespace tn

__global__ void __launch_bounds__(tn::kThreads, 1)
conv1_wgrad_nox1_tf32_kernel(const __grid_constant__ CUtensorMap tm_dp1, const __grid_constant__ CUtensorMap tm_mask,
                             const __grid_constant__ CUtensorMap tm_in, const float* __restrict__ bn,
                             float* __restrict__ partials, WgGeom g) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* s_stage = smem;                                                          // [kStages][A | E4]
  uint8_t* s_rawb = smem + kStages * tn::kStageBytes;                               // [3][dp1 | mask | in]
  float* s_sc = reinterpret_cast<float*>(s_rawb + tn::kRawStages * tn::kRawStage);  // [16] BN scale
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_sc + kC1);
  uint64_t* full = bars;
  uint64_t* empty = full + kStages;
  uint64_t* rfull = empty + kStages;
  uint64_t* rempty = rfull + tn::kRawStages;
  uint64_t* done = rempty + tn::kRawStages;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kStages; ++i) { mbar_init(&full[i], tf::kProdWarps); mbar_init(&empty[i], 1); }
    for (int i = 0; i < tn::kRawStages; ++i) { mbar_init(&rfull[i], 1); mbar_init(&rempty[i], tf::kProdWarps); }
    mbar_init(done, 1);
    mbar_fence_init();
    tma_prefetch_desc(&tm_dp1); tma_prefetch_desc(&tm_mask); tma_prefetch_desc(&tm_in);
  }
  if (threadIdx.x < kC1) s_sc[threadIdx.x] = bn[threadIdx.x];
  // the "ones" atom of both stages: element 0 of unit 0 of every pixel row is 1, the rest 0 (written once)
  for (int i = threadIdx.x; i < kStages * tf::kXW * 8; i += tn::kThreads) {
    const int st = i / (tf::kXW * 8), r = i % (tf::kXW * 8), k = r >> 3, piece = r & 7;
    uint8_t* atom = s_stage + st * tn::kStageBytes + tf::kABytes + 3 * tf::kAtom;
    const float one = (piece == 0) ? 1.f : 0.f;       // logical unit 0, first 16 bytes
    *reinterpret_cast<float4*>(atom + sw32_off(k, piece >> 1) + 16 * (piece & 1)) = make_float4(one, 0.f, 0.f, 0.f);
  }
  fence_proxy_async_smem();
  if (warp == 0) tmem_alloc<256>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ================= MMA issue: D1 (cols 0..95) = A x E, D2 (cols 96..191) = E4 x E =================
    constexpr uint32_t idesc = umma_idesc(2, 128, kNCols, 1, 1);
    const uint64_t tmpl = smem_desc(0, tf::kAtom, 512, 1);
    const uint32_t hi = (uint32_t)(tmpl >> 32), lo0 = (uint32_t)tmpl;
    uint32_t n = 0;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x, ++n) {
      const uint32_t s = n % kStages;
      mbar_wait(&full[s], (n / kStages) & 1);
      tc_fence_after();
      const uint32_t a_lo = lo0 + (smem_u32(s_stage + s * tn::kStageBytes) >> 4);
      const uint32_t e_lo = a_lo + (tf::kABytes >> 4);
      const uint32_t acc0 = n == 0 ? 0u : 1u;
      if (elect_one()) {
#pragma unroll
        for (int kb = 0; kb < tf::kXW / 8; ++kb) {
          const uint64_t ed = ((uint64_t)hi << 32) | (uint64_t)(e_lo + kb * 64);
          umma_tf32(tmem_base, ((uint64_t)hi << 32) | (uint64_t)(a_lo + kb * 64), ed, idesc, kb == 0 ? acc0 : 1u);
          umma_tf32(tmem_base + kNCols, ed, ed, idesc, kb == 0 ? acc0 : 1u);
        }
        umma_commit(&empty[s]);
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit(done);
    __syncwarp();
  } else if (warp == 1) {
    // ================= TMA loader =================
    if (lane == 0) {
      uint32_t n = 0;
      for (int item = blockIdx.x; item < g.nitems; item += gridDim.x, ++n) {
        const int strip = item % g.nstrips;
        const int grp = (item / g.nstrips) % g.ngroups;
        const int b = item / (g.nstrips * g.ngroups);
        const int y0 = grp * kRows, x0 = strip * tf::kXW;
        const uint32_t rs = n % tn::kRawStages;
        mbar_wait(&rempty[rs], ((n / tn::kRawStages) & 1) ^ 1);
        uint8_t* dst = s_rawb + rs * tn::kRawStage;
        mbar_expect_tx(&rfull[rs], tn::kGBytes + tn::kMBytes + tn::kIBytes);
        tma_load_4d(dst, &tm_dp1, &rfull[rs], 0, x0 >> 1, y0 >> 1, b);
        tma_load_4d(dst + tn::kGBytes, &tm_mask, &rfull[rs], 0, x0 >> 1, y0 >> 1, b);
        tma_load_3d(dst + tn::kGBytes + tn::kMBytes, &tm_in, &rfull[rs], x0 - 4, y0 - 2, b);
      }
    }
  } else {
    // ================= producers: route dp1 to the arg-max pixel of its window =================
    const int pt = threadIdx.x - 64;                 // 0..511
    const int q = pt & 3, w = pt >> 2;               // channels 4q..4q+3 of window w
    const int wr = w >> 5, wc = w & 31;
    const float4 sc4 = *reinterpret_cast<const float4*>(s_sc + 4 * q);
    const int stride = (int)gridDim.x;
    int c_strip = (int)blockIdx.x % g.nstrips;
    const int d_strip = stride % g.nstrips;
    const uint32_t goff = (uint32_t)(wr * (tf::kXW / 2) * 64 + wc * 64 + q * 16);
    const uint32_t moff = (uint32_t)(tn::kGBytes + (wr * (tf::kXW / 2) + wc) * 16 + q * 4);
    uint32_t n = 0;
    for (int item = blockIdx.x; item < g.nitems; item += stride, ++n) {
      const uint32_t s = n % kStages, rs = n % tn::kRawStages;
      uint8_t* sA = s_stage + s * tn::kStageBytes;
      uint8_t* sE = sA + tf::kABytes;
      const uint8_t* raw = s_rawb + rs * tn::kRawStage;
      const float* sI = reinterpret_cast<const float*>(raw + tn::kGBytes + tn::kMBytes);
      const int kvalid = g.W - c_strip * tf::kXW;       // pixels of this strip inside the image (>= 64: all)
      c_strip += d_strip;
      if (c_strip >= g.nstrips) c_strip -= g.nstrips;
      mbar_wait(&rfull[rs], (n / tn::kRawStages) & 1);
      const float4 gr = *reinterpret_cast<const float4*>(raw + goff);
      const uint32_t mk = *reinterpret_cast<const uint32_t*>(raw + moff);
      mbar_wait(&empty[s], ((n / kStages) & 1) ^ 1);
      {
        float o[4][4];
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          const uint32_t mb = (mk >> (8 * cc)) & 0xFFu;
          const float gq = (mb & 4u) ? (&gr.x)[cc] * (&sc4.x)[cc] : 0.f;     // ReLU-active: dy * scale
          const uint32_t am = mb & 3u;
#pragma unroll
          for (int i = 0; i < 4; ++i) o[i][cc] = (am == (uint32_t)i) ? gq : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int kx = 2 * wc + (i & 1);
          *reinterpret_cast<float4*>(sA + wr * tf::kAtom + sw32_off(kx, 2 * (i >> 1) + (q >> 1)) + 16 * (q & 1)) =
              make_float4(o[i][0], o[i][1], o[i][2], o[i][3]);
        }
      }
      // expanded input E[r'][k][dxs]; columns past the image edge are zeroed (no output pixel there)
#pragma unroll
      for (int j = 0; j < (2 * kERows * tf::kXW) / tf::kProd; ++j) {
        const int idx = pt + j * tf::kProd;
        const int kx = idx & (tf::kXW - 1), c2 = idx >> 6, rr = c2 >> 1;
        const float* src = sI + rr * tt::kInW + kx + 2 + 4 * (c2 & 1);
        const float4 e = kx < kvalid ? make_float4(src[0], src[1], src[2], src[3]) : make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4*>(sE + (rr >> 2) * tf::kAtom + sw32_off(kx, rr & 3) + 16 * (c2 & 1)) = e;
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) { mbar_arrive(&rempty[rs]); mbar_arrive(&full[s]); }
    }
  }
  // ================= extraction: D1 -> routed dW, D2 -> Gram matrix M and column sums S =================
  float* s_T1 = reinterpret_cast<float*>(s_stage);               // [128][97]
  constexpr int kTP = kNCols + 1;
  float* s_T2 = s_T1 + 128 * kTP;                                // [128][97]
  if (warp >= 2 && warp <= 5) {
    mbar_wait(done, 0);
    tc_fence_after();
    const int qq = warp & 3, m = qq * 32 + lane;
#pragma unroll 1
    for (int c0 = 0; c0 < 2 * kNCols; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(tmem_base + ((uint32_t)(qq * 32) << 16) + c0, v);
      tmem_ld_wait();
      float* dstT = c0 < kNCols ? s_T1 + m * kTP + c0 : s_T2 + m * kTP + (c0 - kNCols);
#pragma unroll
      for (int j = 0; j < 16; ++j) dstT[j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  float* prow = partials + (size_t)blockIdx.x * tn::kPartial;
  for (int t = threadIdx.x; t < tn::kPartial; t += tn::kThreads) {
    float a = 0.f;
    if (t < 400) {                                   // routed: sum_r D1[(r, c)][(r + dy, dx)]
      const int c = t / 25, tap = t % 25, dy = tap / 5, dxs = tap % 5;
#pragma unroll
      for (int r = 0; r < kRows; ++r) a += s_T1[(16 * r + c) * kTP + (r + dy) * 8 + dxs];
    } else if (t < 1025) {                           // M[tap'][tap] = sum_r D2[(r + dy', dx')][(r + dy, dx)]
      const int u = t - 400, t1 = u / 25, t2 = u % 25;
#pragma unroll
      for (int r = 0; r < kRows; ++r) a += s_T2[((r + t1 / 5) * 8 + t1 % 5) * kTP + (r + t2 / 5) * 8 + t2 % 5];
    } else {                                         // S[tap] = sum_r D2[ones row 96][(r + dy, dx)]
      const int tap = t - 1025;
#pragma unroll
      for (int r = 0; r < kRows; ++r) a += s_T2[96 * kTP + (r + tap / 5) * 8 + tap % 5];
    }
    prow[t] = a;
  }
  if (warp == 0) { __syncwarp(); tmem_dealloc<256>(tmem_base); }
}

// dW1 = gscale * (routed + ca * S + cb * (w1 . M + b1 * S)), from the fp64-reduced partial rows.
// One block per tap: the 16 outputs of a tap need 42 column sums of the partial matrix (16 routed, column `tap` of the
// 25x25 Gram matrix, S[tap]) and nothing else, so 25 blocks reduce disjoint columns - eight threads per column, every
// load of a thread independent of the others (one block walking the 148 rows of all 1050 columns took 32 us, all of it
// load latency).
__global__ void __launch_bounds__(384)
conv1_wgrad_nox1_combine_kernel(const float* __restrict__ partials, int nparts,
                                const float* __restrict__ w1, const float* __restrict__ b1,
                                const float* __restrict__ bn, const float* __restrict__ red,
                                float* __restrict__ dw1, float gscale) {
  __shared__ double s_col[48];
  const int tap = blockIdx.x;
  const int j = threadIdx.x >> 3, part = threadIdx.x & 7;
  const bool valid = j < 42;
  const int idx = j < 16 ? j * 25 + tap : (j < 41 ? 400 + (j - 16) * 25 + tap : 1025 + tap);
  double a = 0.0;
  if (valid)
    for (int i = part; i < nparts; i += 8) a += (double)partials[(size_t)i * tn::kPartial + idx];
  a += __shfl_xor_sync(0xffffffffu, a, 1);                   // (every lane takes part: j >= 42 carries zeros)
  a += __shfl_xor_sync(0xffffffffu, a, 2);
  a += __shfl_xor_sync(0xffffffffu, a, 4);
  if (valid && part == 0) s_col[j] = a;
  __syncthreads();
  if (threadIdx.x < kC1) {
    const int c = threadIdx.x;
    const double sc = bn[c], mu = bn[2 * kC1 + c], rs = bn[3 * kC1 + c], mdy = red[c], mdyx = red[kC1 + c];
    const double ca = sc * (rs * mdyx * mu - mdy), cb = -sc * rs * mdyx;
    const double S = s_col[41];
    double R = (double)b1[c] * S;
    for (int t1 = 0; t1 < 25; ++t1) R += (double)w1[c * 25 + t1] * s_col[16 + t1];
    dw1[c * 25 + tap] = (float)((double)gscale * (s_col[c] + ca * S + cb * R));
  }
}

}  // namespace

// Same contract as b200_conv1_wgrad (partials: [b200_conv1_wgrad_grid()][400] scratch).
B200_API int b200_conv1_wgrad_tc(const void* in_f32, const void* x1, const void* dp1, const void* bn, const void* red,
                                 void* partials, void* dw1, int B, int H, int W, float gscale, void* stream) {
  B200_CHECK(H % 2 == 0 && W % 2 == 0, "conv1_wgrad_tc: H and W must be even");
  cudaStream_t s = (cudaStream_t)stream;
  B200_CUDA(set_max_smem(conv1_wgrad_tc_kernel, kSmemBytes));
  WgGeom g;
  g.B = B; g.H = H; g.W = W;
  g.nstrips = (W + kXW - 1) / kXW;
  g.ngroups = (H + kRows - 1) / kRows;
  g.nitems = B * g.nstrips * g.ngroups;
  const int nsm = b200_compute_sms();
  const int grid = g.nitems < nsm ? g.nitems : nsm;
  conv1_wgrad_tc_kernel<<<grid, kWgThreads, kSmemBytes, s>>>((const float*)in_f32, (const __nv_bfloat16*)x1,
                                                             (const __nv_bfloat16*)dp1, (const float*)bn,
                                                             (const float*)red, (float*)partials, g);
  B200_LAUNCH_CHECK();
  return b200_reduce_partials(partials, grid, 400, dw1, gscale, stream);
}

// fp32 storage, kind::tf32 (same contract; x1 and dp1 are fp32).
namespace {
typedef CUresult (*EncodeTiledFn1)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn1 encode_fn1() {
  static EncodeTiledFn1 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn1)p;
  });
  return fn;
}
// fp32 tensor map: dims / strides (bytes, for dims 1..n-1) / box, innermost first
int f32_map(CUtensorMap* tm, const void* ptr, int nd, const cuuint64_t* dims, const cuuint64_t* strides,
            const cuuint32_t* box, CUtensorMapSwizzle swz) {
  EncodeTiledFn1 enc = encode_fn1();
  if (!enc) return b200_fail(-3, "cuTensorMapEncodeTiled entry point not available");
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)nd, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return b200_fail((int)r, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}
}  // namespace

B200_API int b200_conv1_wgrad_tc_f32(const void* in_f32, const void* x1, const void* dp1, const void* bn,
                                     const void* red, void* partials, void* dw1, int B, int H, int W, float gscale,
                                     void* stream) {
  B200_CHECK(H % 2 == 0 && W % 2 == 0, "conv1_wgrad_tc_f32: H and W must be even");
  cudaStream_t s = (cudaStream_t)stream;
  static const bool use_tma = [] { const char* e = getenv("B200DP_CONV1_WGRAD_FEED"); return !(e && e[0] == 'c'); }();
  if (use_tma && W % 4 == 0) {      // TMA needs 16-byte row strides (W * 4 bytes for the input image)
    WgGeom g;
    g.B = B; g.H = H; g.W = W;
    g.nstrips = (W + tf::kXW - 1) / tf::kXW;
    g.ngroups = (H + kRows - 1) / kRows;
    g.nitems = B * g.nstrips * g.ngroups;
    const int nsm = b200_compute_sms();
    const int grid = g.nitems < nsm ? g.nitems : nsm;
    CUtensorMap tx, tg, ti;
    {
      cuuint64_t d[4] = {32, (cuuint64_t)W / 2, (cuuint64_t)H, (cuuint64_t)B};
      cuuint64_t st[3] = {128, (cuuint64_t)W * 64, (cuuint64_t)H * W * 64};
      cuuint32_t bx[4] = {32, tf::kXW / 2, kRows, 1};
      int rc = f32_map(&tx, x1, 4, d, st, bx, CU_TENSOR_MAP_SWIZZLE_128B);
      if (rc) return rc;
    }
    {
      cuuint64_t d[4] = {16, (cuuint64_t)W / 2, (cuuint64_t)H / 2, (cuuint64_t)B};
      cuuint64_t st[3] = {64, (cuuint64_t)(W / 2) * 64, (cuuint64_t)(H / 2) * (W / 2) * 64};
      cuuint32_t bx[4] = {16, tf::kXW / 2, kRows / 2, 1};
      int rc = f32_map(&tg, dp1, 4, d, st, bx, CU_TENSOR_MAP_SWIZZLE_NONE);
      if (rc) return rc;
    }
    {
      cuuint64_t d[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
      cuuint64_t st[2] = {(cuuint64_t)W * 4, (cuuint64_t)H * W * 4};
      cuuint32_t bx[3] = {tt::kInW, kERows, 1};
      int rc = f32_map(&ti, in_f32, 3, d, st, bx, CU_TENSOR_MAP_SWIZZLE_NONE);
      if (rc) return rc;
    }
    B200_CUDA(set_max_smem(conv1_wgrad_tf32_tma_kernel, tt::kSmemBytes));
    conv1_wgrad_tf32_tma_kernel<<<grid, tt::kThreads, tt::kSmemBytes, s>>>(tx, tg, ti, (const float*)bn, (const float*)red,
                                                                           (float*)partials, g);
    B200_LAUNCH_CHECK();
    return b200_reduce_partials(partials, grid, 400, dw1, gscale, stream);
  }
  B200_CUDA(set_max_smem(conv1_wgrad_tf32_kernel, tf::kSmemBytes));
  WgGeom g;
  g.B = B; g.H = H; g.W = W;
  g.nstrips = (W + tf::kXW - 1) / tf::kXW;
  g.ngroups = (H + kRows - 1) / kRows;
  g.nitems = B * g.nstrips * g.ngroups;
  const int nsm = b200_compute_sms();
  const int grid = g.nitems < nsm ? g.nitems : nsm;
  conv1_wgrad_tf32_kernel<<<grid, tf::kThreads, tf::kSmemBytes, s>>>((const float*)in_f32, (const float*)x1,
                                                                   (const float*)dp1, (const float*)bn,
                                                                   (const float*)red, (float*)partials, g);
  B200_LAUNCH_CHECK();
  return b200_reduce_partials(partials, grid, 400, dw1, gscale, stream);
}

// Layer-1 weight gradient without the conv output (see conv1_wgrad_nox1_tf32_kernel).  amask_u8 = the bytes
// b200_conv1_fwd_pool_mask_tc_f32 wrote; partials: [b200_conv1_wgrad_grid()][b200_conv1_wgrad_nox1_partial()] floats.
// Requires H % 8 == 0 and W % 4 == 0 (returns -4 otherwise: use b200_conv1_wgrad_tc_f32 with the stored x1).
B200_API int b200_conv1_wgrad_nox1_partial() { return tn::kPartial; }
B200_API int b200_conv1_wgrad_nox1_f32(const void* in_f32, const void* dp1, const void* amask_u8, const void* w1,
                                       const void* b1, const void* bn, const void* red, void* partials, void* dw1, int B,
                                       int H, int W, float gscale, void* stream) {
  if (H % 8 != 0 || W % 4 != 0) return b200_fail(-4, "conv1_wgrad_nox1: needs H %% 8 == 0 and W %% 4 == 0 (got %d x %d)", H, W);
  cudaStream_t s = (cudaStream_t)stream;
  WgGeom g;
  g.B = B; g.H = H; g.W = W;
  g.nstrips = (W + tf::kXW - 1) / tf::kXW;
  g.ngroups = H / kRows;
  g.nitems = B * g.nstrips * g.ngroups;
  const int nsm = b200_compute_sms();
  const int grid = g.nitems < nsm ? g.nitems : nsm;
  CUtensorMap tg, tmk, ti;
  {
    cuuint64_t d[4] = {16, (cuuint64_t)W / 2, (cuuint64_t)H / 2, (cuuint64_t)B};
    cuuint64_t st[3] = {64, (cuuint64_t)(W / 2) * 64, (cuuint64_t)(H / 2) * (W / 2) * 64};
    cuuint32_t bx[4] = {16, tf::kXW / 2, kRows / 2, 1};
    int rc = f32_map(&tg, dp1, 4, d, st, bx, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
  }
  {
    EncodeTiledFn1 enc = encode_fn1();
    if (!enc) return b200_fail(-3, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t d[4] = {16, (cuuint64_t)W / 2, (cuuint64_t)H / 2, (cuuint64_t)B};
    cuuint64_t st[3] = {16, (cuuint64_t)(W / 2) * 16, (cuuint64_t)(H / 2) * (W / 2) * 16};
    cuuint32_t bx[4] = {16, tf::kXW / 2, kRows / 2, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(&tmk, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(amask_u8), d, st, bx, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return b200_fail((int)r, "cuTensorMapEncodeTiled (mask map) failed with CUresult %d", (int)r);
  }
  {
    cuuint64_t d[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t st[2] = {(cuuint64_t)W * 4, (cuuint64_t)H * W * 4};
    cuuint32_t bx[3] = {tt::kInW, kERows, 1};
    int rc = f32_map(&ti, in_f32, 3, d, st, bx, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
  }
  B200_CUDA(set_max_smem(conv1_wgrad_nox1_tf32_kernel, tn::kSmemBytes));
  conv1_wgrad_nox1_tf32_kernel<<<grid, tn::kThreads, tn::kSmemBytes, s>>>(tg, tmk, ti, (const float*)bn, (float*)partials, g);
  B200_LAUNCH_CHECK();
  conv1_wgrad_nox1_combine_kernel<<<25, 384, 0, s>>>((const float*)partials, grid, (const float*)w1, (const float*)b1,
                                                    (const float*)bn, (const float*)red, (float*)dw1, gscale);
  B200_LAUNCH_CHECK();
  return 0;
}

// =====================================================================================
// K3 (+K4 statistics), conv1 forward as a Toeplitz GEMM on tcgen05.
//
//   out[y0+yo, x, c] = sum_{dx} sum_{k} in[y0-2+k, x+dx-2] * T_dx[k][(yo, c)],
//   T_dx[k][(yo,c)] = w[c][k - yo][dx]  (0 <= k - yo <= 4, else 0),   k = 0..15, yo = 0..7
//
// i.e. for 8 output rows at a time, M = 128 image COLUMNS, N = 8 rows x 16 channels = 128,
// K = 16 input rows, five MMAs (one per dx; A = the same input planes shifted right by dx
// columns = start address + dx * 16 B).  31 % of the MACs are useful; a 64-cycle MMA still
// replaces 409,600 FFMAs.  The input tile is written by producer warps as bf16 "row planes"
// (8 vertically consecutive pixels = 16 B per column, columns 16 B apart) so that it is directly
// the no-swizzle K-major A operand.
// Orientation matters for the epilogue: a TMEM lane is an image column, so a warp's 256-bit store
// of one output row covers 32 consecutive pixels = 1 KB of contiguous NHWC memory.  (The first
// version put image ROWS on the lanes: every store instruction touched 32 different lines and the
// kernel was bound by LSU wavefronts in the L1 data pipe, ncu: 65 %.)
// Accuracy: inputs and weights are split into bf16 hi + lo parts and three MMAs
// (hi*hi + hi*lo + lo*hi) are accumulated in fp32, which restores ~16 mantissa bits - the
// result matches the fp32 FFMA kernel to output rounding.  The kernel is bounded by writing
// its 16-channel bf16 output (1.44 GB at batch 5).
namespace {

constexpr int kFwCols = 128;                 // M: image columns per item
constexpr int kFwSegRows = 64;               // output rows per item (8 row-blocks of 8)
constexpr int kFwBlocks = kFwSegRows / 8;
constexpr int kFwPlanes = kFwBlocks + 1;     // input planes of 8 rows
constexpr int kFwPlaneCols = 136;            // 132 needed, padded so a plane is 17 * 128 B
constexpr int kFwPlaneBytes = kFwPlaneCols * 16;
constexpr int kFwXBytes = kFwPlanes * kFwPlaneBytes;          // one of hi / lo
constexpr int kFwStageBytes = 2 * kFwXBytes;                  // hi + lo
constexpr int kFwStages = 2;
constexpr int kFwWBytes = 5 * 2 * 128 * 16;                   // Toeplitz weights, one of hi / lo
constexpr int kFwProdWarps = 4, kFwEpiWarps = 8;
constexpr int kFwThreads = 32 * (1 + kFwProdWarps + kFwEpiWarps);   // 416
constexpr int kFwTStages = 4;
constexpr int kFwTileH = kFwPlanes * 8;      // 72 input rows
constexpr int kFwTileW = kFwCols + 4;        // 132 input columns
constexpr int kFwRawBytes = kFwTileH * kFwTileW * 4;          // fp32 staging of one input tile (38 KB)
constexpr int kFwSmemBytes = kFwStages * kFwStageBytes + 2 * kFwWBytes + 2 * kFwRawBytes + 2048 + 128;

struct FwGeom {
  int B, H, W;
  int nys, nxs, nitems;
};

// OT = storage type of the output: __nv_bfloat16 or float; POOL (float only) = second pass of the two-pass forward
template <typename OT, bool POOL = false>
__global__ void __launch_bounds__(kFwThreads, 1)
conv1_fwd_tc_kernel(const float* __restrict__ in, const float* __restrict__ w1, const float* __restrict__ b1,
                    OT* __restrict__ x1, float* __restrict__ partials, FwGeom g,
                    const float* __restrict__ bn = nullptr, float* __restrict__ pooled = nullptr,
                    uint8_t* __restrict__ amask = nullptr) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  uint8_t* s_x = smem;                                   // [kFwStages][hi | lo][row planes][columns][16 B]
  uint8_t* s_w = smem + kFwStages * kFwStageBytes;       // [hi | lo][dx][chunk][n][16 B]
  uint8_t* s_raw = s_w + 2 * kFwWBytes;                  // [2][72][132] fp32, filled by cp.async
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_raw + 2 * kFwRawBytes);
  uint64_t* xfull = bars;                   // [2]
  uint64_t* xempty = xfull + kFwStages;     // [2]
  uint64_t* tfull = xempty + kFwStages;     // [4]
  uint64_t* tempty = tfull + kFwTStages;    // [4]
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(tempty + kFwTStages);
  float* s_bias = reinterpret_cast<float*>(s_tmem + 4);  // [16]
  float* s_stat = s_bias + 16;                            // [8 warps][32]
  float* s_bn = s_stat + 8 * 32;                          // [scale 16 | shift 16] (fused BN+ReLU+pool epilogue)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (POOL && threadIdx.x < 32) s_bn[threadIdx.x] = bn[threadIdx.x];

  // Toeplitz weight tiles (hi / lo)
  for (int i = threadIdx.x; i < 5 * 2 * 128 * 8; i += kFwThreads) {
    const int e = i & 7;
    int r = i >> 3;
    const int n = r & 127; r >>= 7;
    const int c = r & 1;
    const int dx = r >> 1;
    const int k = c * 8 + e, yo = n >> 4, ch = n & 15, dy = k - yo;
    const float v = (dy >= 0 && dy <= 4) ? w1[ch * 25 + dy * 5 + dx] : 0.f;
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    reinterpret_cast<__nv_bfloat16*>(s_w)[i] = hi;
    reinterpret_cast<__nv_bfloat16*>(s_w + kFwWBytes)[i] = __float2bfloat16_rn(v - __bfloat162float(hi));
  }
  if (threadIdx.x < 16) s_bias[threadIdx.x] = b1[threadIdx.x];
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    for (int i = 0; i < kFwStages; ++i) { mbar_init(&xfull[i], kFwProdWarps); mbar_init(&xempty[i], 1); }
    for (int i = 0; i < kFwTStages; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 128); }
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc<512>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ================= MMA issue =================
    constexpr uint32_t idesc = umma_idesc_bf16(128, 128, 0, 0);
    const uint64_t a_tmpl = smem_desc_noswz(0, kFwPlaneBytes, 128);      // K-major: next k-chunk = next plane
    const uint64_t b_tmpl = smem_desc_noswz(0, 128 * 16, 128);
    const uint32_t a_hi32 = (uint32_t)(a_tmpl >> 32), a_lo0 = (uint32_t)a_tmpl;
    const uint32_t b_hi32 = (uint32_t)(b_tmpl >> 32);
    const uint32_t w_hi_lo = (uint32_t)b_tmpl + (smem_u32(s_w) >> 4), w_lo_lo = w_hi_lo + (kFwWBytes >> 4);
    uint32_t n = 0, q = 0;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x, ++n) {
      const uint32_t s = n % kFwStages;
      mbar_wait(&xfull[s], (n / kFwStages) & 1);
      const uint32_t xh = a_lo0 + (smem_u32(s_x + s * kFwStageBytes) >> 4), xl = xh + (kFwXBytes >> 4);
      for (int j = 0; j < kFwBlocks; ++j, ++q) {
        const uint32_t ts = q % kFwTStages;
        mbar_wait(&tempty[ts], ((q / kFwTStages) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d = tmem_base + ts * 128;
        const uint32_t pj = (uint32_t)j * (kFwPlaneBytes >> 4);
        if (elect_one()) {
#pragma unroll
          for (int dx = 0; dx < 5; ++dx) {
            const uint64_t ah = ((uint64_t)a_hi32 << 32) | (uint64_t)(xh + pj + dx);
            const uint64_t al = ((uint64_t)a_hi32 << 32) | (uint64_t)(xl + pj + dx);
            const uint64_t bh = ((uint64_t)b_hi32 << 32) | (uint64_t)(w_hi_lo + dx * (4096 >> 4));
            const uint64_t bl = ((uint64_t)b_hi32 << 32) | (uint64_t)(w_lo_lo + dx * (4096 >> 4));
            umma_bf16(d, ah, bh, idesc, dx == 0 ? 0u : 1u);
            umma_bf16(d, ah, bl, idesc, 1u);
            umma_bf16(d, al, bh, idesc, 1u);
          }
          umma_commit(&tfull[ts]);
          if (j == kFwBlocks - 1) umma_commit(&xempty[s]);
        }
        __syncwarp();
      }
    }
  } else if (warp <= kFwProdWarps) {
    // ================= producers: fp32 input tile -> bf16 hi/lo row planes =================
    // The tile of item n+1 is fetched with 8-byte cp.async (zero-fill outside the image) into an fp32
    // staging buffer while item n is converted, so HBM latency is off the critical path and costs no
    // registers.  The conversion transposes (8 vertically consecutive pixels become one 16-byte piece),
    // so a thread converts pixels other threads fetched: one named barrier per item orders both the
    // landing of item n and the reuse of the staging buffer item n-1 was converted from.
    const int pt = threadIdx.x - 32;                    // 0..127
    constexpr int kPieces = kFwTileH * kFwTileW / 2;    // 8-byte pieces (2 pixels) per tile
    auto fetch = [&](int item, int buf) {
      const int xs = item % g.nxs;
      const int ys = (item / g.nxs) % g.nys;
      const int b = item / (g.nxs * g.nys);
      const int y0 = ys * kFwSegRows, x0 = xs * kFwCols;
      const float* src = in + (size_t)b * g.H * g.W;
      uint8_t* raw = s_raw + buf * kFwRawBytes;
      for (int i = pt; i < kPieces; i += 32 * kFwProdWarps) {
        const int rho = i / (kFwTileW / 2), sx = 2 * (i - rho * (kFwTileW / 2));
        const int gy = y0 - 2 + rho, gx = x0 - 2 + sx;        // gx even, W even: both pixels in or both out
        const bool ok = gy >= 0 && gy < g.H && gx >= 0 && gx < g.W;
        cp_async8(raw + (size_t)i * 8, src + (ok ? (size_t)gy * g.W + gx : 0), ok ? 8u : 0u);
      }
      cp_async_commit();
    };
    if ((int)blockIdx.x < g.nitems) fetch(blockIdx.x, 0);
    uint32_t n = 0;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x, ++n) {
      const uint32_t s = n % kFwStages;
      uint8_t* xh = s_x + s * kFwStageBytes;
      uint8_t* xl = xh + kFwXBytes;
      cp_async_wait<0>();                                     // this thread's pieces of item n have landed
      named_bar_sync(2, 32 * kFwProdWarps);                   // ... everyone's have, and item n-1 is fully converted
      if (item + (int)gridDim.x < g.nitems) fetch(item + gridDim.x, (n + 1) & 1);
      mbar_wait(&xempty[s], ((n / kFwStages) & 1) ^ 1);
      const float* raw = reinterpret_cast<const float*>(s_raw + (n & 1) * kFwRawBytes);
      for (int p = pt; p < kFwPlanes * kFwTileW; p += 32 * kFwProdWarps) {
        const int pr = p / kFwTileW, x = p - pr * kFwTileW;
        const float* col = raw + (size_t)(8 * pr) * kFwTileW + x;
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const float v0 = col[(2 * r) * kFwTileW], v1 = col[(2 * r + 1) * kFwTileW];
          hi[r] = pack_bf16x2(v0, v1);
          lo[r] = pack_bf16x2(v0 - bf16_lo(hi[r]), v1 - bf16_hi(hi[r]));
        }
        const int off = pr * kFwPlaneBytes + x * 16;
        *reinterpret_cast<uint4*>(xh + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(xl + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&xfull[s]);
    }
  } else {
    // ================= epilogue: 8 warps, two groups alternating tiles =================
    const int ew = warp - (1 + kFwProdWarps);           // 0..7
    const int quarter = warp & 3, grp = ew >> 2;
    const int m = quarter * 32 + lane;
    float ssum[16], ssq[16], bias[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) { ssum[c] = 0.f; ssq[c] = 0.f; bias[c] = s_bias[c]; }
    uint32_t q = 0;
    for (int item = blockIdx.x; item < g.nitems; item += gridDim.x) {
      const int xs = item % g.nxs;
      const int ys = (item / g.nxs) % g.nys;
      const int b = item / (g.nxs * g.nys);
      const int x = xs * kFwCols + m;
      const bool col_ok = x < g.W;
      for (int j = 0; j < kFwBlocks; ++j, ++q) {
        if ((int)(q & 1) != grp) continue;
        const uint32_t ts = q % kFwTStages;
        mbar_wait(&tfull[ts], (q / kFwTStages) & 1);
        tc_fence_after();
        const int yb = ys * kFwSegRows + j * 8;
        OT* ocol = x1 + (((size_t)b * g.H + yb) * g.W + (col_ok ? x : 0)) * 16;
        // 4 chunks of 2 rows x 16 channels; the tcgen05.ld of chunk k+1 is in flight while chunk k is
        // converted (tcgen05.wait::ld waits for ALL outstanding loads, so it comes after the math)
        const uint32_t tbase = tmem_base + ((uint32_t)(quarter * 32) << 16) + ts * 128;
        uint32_t va[32], vb[32];
        tmem_ld16(tbase, *reinterpret_cast<uint32_t(*)[16]>(&va[0]));
        tmem_ld16(tbase + 16, *reinterpret_cast<uint32_t(*)[16]>(&va[16]));
        tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          uint32_t* v = (k & 1) ? vb : va;
          uint32_t* vn = (k & 1) ? va : vb;
          if (k < 3) {
            tmem_ld16(tbase + (k + 1) * 32, *reinterpret_cast<uint32_t(*)[16]>(&vn[0]));
            tmem_ld16(tbase + (k + 1) * 32 + 16, *reinterpret_cast<uint32_t(*)[16]>(&vn[16]));
          }
          const int c0 = k * 32;
          const bool ok = col_ok && (yb + (c0 >> 4)) < g.H;     // H % 2 == 0: both rows or neither
          if constexpr (sizeof(OT) == 4) {
            // fp32 storage: accumulator + bias stored as is (hi/lo split products: fp32-level accuracy)
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const float f0 = __uint_as_float(v[e]) + bias[e & 15];
              v[e] = __float_as_uint(f0);
              if (!POOL && ok) { ssum[e & 15] += f0; ssq[e & 15] = fmaf(f0, f0, ssq[e & 15]); }
            }
            if (ok && x1 != nullptr) {        // x1 == nullptr: statistics-only pass
              float* o = reinterpret_cast<float*>(ocol) + (size_t)(c0 >> 4) * g.W * 16;
              st_global_v8(o, &v[0]);
              st_global_v8(o + 8, &v[8]);
              st_global_v8(o + (size_t)g.W * 16, &v[16]);
              st_global_v8(o + (size_t)g.W * 16 + 8, &v[24]);
            }
            if constexpr (POOL) {
              // second pass of the two-pass forward: the batch statistics are known (first pass), so BN + ReLU + the
              // 2x2 max-pool run here and the separate pass over the 2.88 GB conv output (K5) disappears.  The two
              // rows of a window are this chunk's two rows, the two columns are this lane and its neighbour - and the
              // two lanes SHARE the window's work: the even lane takes channels 0..7, the odd lane 8..15; each sends
              // the other the half it does not keep (one shuffle per value instead of both lanes computing all 16).
              // amask (optional): one byte per pooled element = arg-max position (first maximum in row-major order,
              // as ATen picks it) | ReLU-active << 2 - all the backward needs to route dp1 without the conv output.
              const bool odd = lane & 1;
              uint32_t pm[8], mb[2] = {0u, 0u};
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                // v already holds conv + bias.  mine = my column, theirs = the neighbour's, channel c = 8 * odd + i
                const float m0 = __uint_as_float(odd ? v[8 + i] : v[i]), m1 = __uint_as_float(odd ? v[24 + i] : v[16 + i]);
                const float s0 = __uint_as_float(odd ? v[i] : v[8 + i]), s1 = __uint_as_float(odd ? v[16 + i] : v[24 + i]);
                const float t0 = __shfl_xor_sync(0xffffffffu, s0, 1), t1 = __shfl_xor_sync(0xffffffffu, s1, 1);
                const float sc = s_bn[(odd ? 8 : 0) + i], sh = s_bn[16 + (odd ? 8 : 0) + i];
                const float ym0 = fmaf(m0, sc, sh), ym1 = fmaf(m1, sc, sh), yt0 = fmaf(t0, sc, sh), yt1 = fmaf(t1, sc, sh);
                // window in row-major order: (row 0, left), (row 0, right), (row 1, left), (row 1, right); left = even lane
                const float l0 = odd ? yt0 : ym0, r0 = odd ? ym0 : yt0, l1 = odd ? yt1 : ym1, r1 = odd ? ym1 : yt1;
                const float a0 = fmaxf(l0, r0), a1 = fmaxf(l1, r1), best = fmaxf(a0, a1);
                // a later candidate wins only if strictly greater
                const uint32_t am = a1 > a0 ? (r1 > l1 ? 3u : 2u) : (r0 > l0 ? 1u : 0u);
                pm[i] = __float_as_uint(fmaxf(best, 0.f));
                mb[i >> 2] |= (am | (best > 0.f ? 4u : 0u)) << (8 * (i & 3));
              }
              if (ok) {
                const size_t pix = ((size_t)b * (g.H >> 1) + ((yb + (c0 >> 4)) >> 1)) * (g.W >> 1) + (x >> 1);
                st_global_v8(pooled + pix * 16 + (odd ? 8 : 0), &pm[0]);      // the pair writes the pixel's 64 bytes
                if (amask != nullptr) *reinterpret_cast<uint2*>(amask + pix * 16 + (odd ? 8 : 0)) = make_uint2(mb[0], mb[1]);
              }
            }
            if (k < 3) tmem_ld_wait();
            else { tc_fence_before(); mbar_arrive(&tempty[ts]); }
            continue;
          }
          uint32_t pk[16];
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            pk[e >> 1] = pack_bf16x2(__uint_as_float(v[e]) + bias[e & 15], __uint_as_float(v[e + 1]) + bias[(e + 1) & 15]);
            // statistics of the ROUNDED values: BN backward relies on mean/var being exactly those of the
            // stored tensor (sum dx = 0, sum dx*xhat = 0); unrounded stats leave a coherent residual
            const float f0 = bf16_lo(pk[e >> 1]), f1 = bf16_hi(pk[e >> 1]);
            if (ok) {
              ssum[e & 15] += f0; ssq[e & 15] = fmaf(f0, f0, ssq[e & 15]);
              ssum[(e + 1) & 15] += f1; ssq[(e + 1) & 15] = fmaf(f1, f1, ssq[(e + 1) & 15]);
            }
          }
          if (ok) {
            // lanes are image columns: each instruction writes 32 consecutive pixels of one row (1 KB contiguous)
            __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(ocol) + (size_t)(c0 >> 4) * g.W * 16;
            st_global_v8(o, &pk[0]);
            st_global_v8(o + (size_t)g.W * 16, &pk[8]);
          }
          if (k < 3) tmem_ld_wait();
          else { tc_fence_before(); mbar_arrive(&tempty[ts]); }
        }
      }
    }
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      const float a = warp_sum(ssum[c]), s2 = warp_sum(ssq[c]);
      if (lane == 0) { s_stat[ew * 32 + c] = a; s_stat[ew * 32 + 16 + c] = s2; }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (threadIdx.x < 32) {
    float a = 0.f;
#pragma unroll
    for (int w8 = 0; w8 < kFwEpiWarps; ++w8) a += s_stat[w8 * 32 + threadIdx.x];
    partials[(size_t)blockIdx.x * 32 + threadIdx.x] = a;
  }
  if (warp == 0) { __syncwarp(); tmem_dealloc<512>(tmem_base); }
}

}  // namespace

// Same contract as b200_conv1_fwd; partials must hold [b200_sm_count()][32]; *nparts_out = rows written.
template <typename OT>
static int conv1_fwd_tc_t(const void* in_f32, const void* w1, const void* b1, void* x1, void* partials, int B,
                          int H, int W, int* nparts_out, void* stream, const void* bn = nullptr, void* pooled = nullptr,
                          void* amask = nullptr) {
  B200_CHECK(H % 2 == 0 && W % 2 == 0, "conv1_fwd_tc: H and W must be even");
  cudaStream_t s = (cudaStream_t)stream;
  if (pooled != nullptr) {
    if constexpr (sizeof(OT) == 4) {
      B200_CUDA(set_max_smem(conv1_fwd_tc_kernel<OT, true>, kFwSmemBytes));
      FwGeom gp;
      gp.B = B; gp.H = H; gp.W = W;
      gp.nys = (H + kFwSegRows - 1) / kFwSegRows;
      gp.nxs = (W + kFwCols - 1) / kFwCols;
      gp.nitems = B * gp.nys * gp.nxs;
      const int nsmp = b200_compute_sms();
      const int gridp = gp.nitems < nsmp ? gp.nitems : nsmp;
      conv1_fwd_tc_kernel<OT, true><<<gridp, kFwThreads, kFwSmemBytes, s>>>((const float*)in_f32, (const float*)w1,
                                                                            (const float*)b1, (OT*)x1, (float*)partials,
                                                                            gp, (const float*)bn, (float*)pooled,
                                                                            (uint8_t*)amask);
      B200_LAUNCH_CHECK();
      if (nparts_out) *nparts_out = gridp;
      return 0;
    } else {
      return b200_fail(-2, "conv1_fwd_pool: fp32 storage only");
    }
  }
  B200_CUDA(set_max_smem(conv1_fwd_tc_kernel<OT>, kFwSmemBytes));
  FwGeom g;
  g.B = B; g.H = H; g.W = W;
  g.nys = (H + kFwSegRows - 1) / kFwSegRows;
  g.nxs = (W + kFwCols - 1) / kFwCols;
  g.nitems = B * g.nys * g.nxs;
  const int nsm = b200_compute_sms();
  const int grid = g.nitems < nsm ? g.nitems : nsm;
  conv1_fwd_tc_kernel<OT><<<grid, kFwThreads, kFwSmemBytes, s>>>((const float*)in_f32, (const float*)w1,
                                                                 (const float*)b1, (OT*)x1, (float*)partials, g,
                                                                 (const float*)bn, (float*)pooled);
  B200_LAUNCH_CHECK();
  if (nparts_out) *nparts_out = grid;
  return 0;
}
B200_API int b200_conv1_fwd_tc(const void* in_f32, const void* w1, const void* b1, void* x1_bf16, void* partials, int B,
                               int H, int W, int* nparts_out, void* stream) {
  return conv1_fwd_tc_t<__nv_bfloat16>(in_f32, w1, b1, x1_bf16, partials, B, H, W, nparts_out, stream);
}
// fp32 output (precision-matched mode)
B200_API int b200_conv1_fwd_tc_f32(const void* in_f32, const void* w1, const void* b1, void* x1_f32, void* partials,
                                   int B, int H, int W, int* nparts_out, void* stream) {
  return conv1_fwd_tc_t<float>(in_f32, w1, b1, x1_f32, partials, B, H, W, nparts_out, stream);
}

// Two-pass fp32 forward of layer 1 (the conv is cheap on the tensor cores, its 2.88 GB output is not):
//   b200_conv1_stats_tc_f32      pass 1 - batch statistics only, nothing stored (0.24 ms at batch 5);
//   b200_conv1_fwd_pool_tc_f32   pass 2 - recomputes the conv, stores it (the backward needs it) and applies
//                                BN (bn = [scale | shift | ...]) + ReLU + 2x2 max-pool in the epilogue -> p1.
// Both passes produce bit-identical conv values (same MMAs in the same order), so the statistics of pass 1 are those
// of the tensor pass 2 stores.
B200_API int b200_conv1_stats_tc_f32(const void* in_f32, const void* w1, const void* b1, void* partials, int B, int H,
                                     int W, int* nparts_out, void* stream) {
  return conv1_fwd_tc_t<float>(in_f32, w1, b1, nullptr, partials, B, H, W, nparts_out, stream);
}
B200_API int b200_conv1_fwd_pool_tc_f32(const void* in_f32, const void* w1, const void* b1, const void* bn, void* x1_f32,
                                        void* p1_f32, void* partials, int B, int H, int W, void* stream) {
  B200_CHECK(bn != nullptr && p1_f32 != nullptr && x1_f32 != nullptr, "conv1_fwd_pool: bn, x1 and p1 are required");
  return conv1_fwd_tc_t<float>(in_f32, w1, b1, x1_f32, partials, B, H, W, nullptr, stream, bn, p1_f32);
}
// Same, without the conv output: x1 may be null (nothing of the 2.88 GB tensor is written) and `amask`
// ([B][H/2][W/2][16] bytes: arg-max position | ReLU-active << 2) is what b200_conv1_wgrad_nox1_f32 routes dp1 with.
B200_API int b200_conv1_fwd_pool_mask_tc_f32(const void* in_f32, const void* w1, const void* b1, const void* bn,
                                             void* x1_f32_or_null, void* p1_f32, void* amask_u8, void* partials, int B,
                                             int H, int W, void* stream) {
  B200_CHECK(bn != nullptr && p1_f32 != nullptr && amask_u8 != nullptr, "conv1_fwd_pool_mask: bn, p1 and the mask are required");
  return conv1_fwd_tc_t<float>(in_f32, w1, b1, x1_f32_or_null, partials, B, H, W, nullptr, stream, bn, p1_f32, amask_u8);
}
