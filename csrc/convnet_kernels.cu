// CUDA-core (non tensor-core) kernels of the b200dp ConvNet training step.
//
// The reference model (mnist_onegpu.py:11-31 / mnist_distributed.py:25-45) is
//   conv(1->16,5x5,p2) BN ReLU pool2 | conv(16->32,5x5,p2) BN ReLU pool2 | fc(18M->10)
// and the reference runs it as ~30 separate cuDNN/ATen/cuBLAS launches, each a
// full HBM round trip over multi-GB fp32 NCHW tensors.
// Here the step is re-cut for B200: activations are bf16 NHWC, BN statistics
// are produced in the epilogue of the conv that feeds them, BN-apply+ReLU+pool
// is one pass, max-pool indices are never stored (recomputed from the pre-BN
// tensor in backward), the loss is fused with the fc tail and the fc backward
// can apply SGD in place.  The two layers that are GEMM-shaped (conv2
// fwd/dgrad/wgrad) live in conv_tc.cu on tcgen05; the ones here are either
// bandwidth-bound or have Cin=1 (conv1: K=25, no tensor-core shape).
#include "common.h"
#include "ptx.cuh"
#include <cstdint>
#include <mutex>

using b2::bf16_hi;
using b2::bf16_lo;
using b2::pack_bf16x2;
using b2::warp_sum;

namespace {

constexpr int kC1 = 16;       // conv1 output channels
constexpr int kTaps = 25;     // 5x5
__constant__ float c_w1[kC1 * kTaps];
__constant__ float c_b1[kC1];

__device__ __forceinline__ uint4 ldg_stream16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ldg_stream8(const void* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}

// Activation storage: bf16 (the fast mode) or fp32 (the reference's precision).  Kernels below are templated
// on the storage type and move 16-byte pieces: 8 bf16 or 4 fp32 channels.
template <typename T> struct Act;
template <> struct Act<__nv_bfloat16> {
  static constexpr int kN = 8;                         // channels per 16-byte piece
  __device__ static __forceinline__ void unpack(const uint4& v, float (&f)[8]) {
    f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
    f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
  }
  __device__ static __forceinline__ uint4 pack(const float (&f)[8]) {
    return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
  }
  // four consecutive elements (8 bytes)
  __device__ static __forceinline__ void load4(const __nv_bfloat16* p, float (&f)[4]) {
    const uint2 v = ldg_stream8(p);
    f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  }
  // stores four elements and returns them as stored (rounded) in f
  __device__ static __forceinline__ void store4(__nv_bfloat16* p, float (&f)[4]) {
    const uint2 o = make_uint2(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]));
    *reinterpret_cast<uint2*>(p) = o;
    f[0] = bf16_lo(o.x); f[1] = bf16_hi(o.x); f[2] = bf16_lo(o.y); f[3] = bf16_hi(o.y);
  }
};
template <> struct Act<float> {
  static constexpr int kN = 4;
  __device__ static __forceinline__ void unpack(const uint4& v, float (&f)[4]) {
    f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y); f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
  }
  __device__ static __forceinline__ uint4 pack(const float (&f)[4]) {
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
  }
  __device__ static __forceinline__ void load4(const float* p, float (&f)[4]) { unpack(ldg_stream16(p), f); }
  __device__ static __forceinline__ void store4(float* p, float (&f)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
  }
};

// ----------------------------------------------------------------- resize ---
// K1/K2: the reference resizes 28x28 -> 3000x3000 on the CPU
// with PIL (0.176 s/image) and ships 36 MB/image over PCIe.  Here the raw
// uint8 digits go to the device (784 B/image) and this kernel reproduces PIL's
// 8-bit bilinear resample bit for bit: horizontal pass then vertical pass,
// 22-bit fixed-point coefficients, rounding to uint8 after EACH pass, then
// ToTensor's /255.  The coefficient tables come from the host (data/resize.py).
__device__ __forceinline__ int clip8(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }

// One thread owns 4 consecutive output columns and walks kRzRows output rows: the horizontal pass of a
// source-row pair is recomputed only when the source row changes (every ~H/sh output rows), so an output
// pixel costs two integer multiply-adds, a table look-up for the exact v / 255.0f, and a quarter of a
// 16-byte store.  (One thread per pixel with its own table loads and a true fp32 division ran at 0.25 ms
// for a 180 MB write: 8x its memory floor.)
constexpr int kRzThreads = 128, kRzPx = 4, kRzMaxRows = 64, kRzMaxSrc = 64 * 64;

__global__ void __launch_bounds__(kRzThreads)
resize_pil_bilinear_kernel(const uint8_t* __restrict__ src, int sh, int sw,
                           const int* __restrict__ xtab, const int* __restrict__ ytab,
                           float* __restrict__ dst, int H, int W, int rows_per_block) {
  __shared__ uint8_t s_src[kRzMaxSrc];
  __shared__ int s_ytab[3 * kRzMaxRows];
  __shared__ float s_lut[256];
  const int b = blockIdx.z;
  for (int i = threadIdx.x; i < sh * sw; i += kRzThreads) s_src[i] = src[(size_t)b * sh * sw + i];
  for (int i = threadIdx.x; i < 256; i += kRzThreads) s_lut[i] = (float)i / 255.0f;   // ToTensor's division, exact
  const int y_begin = blockIdx.y * rows_per_block, y_end = min(y_begin + rows_per_block, H);
  for (int i = threadIdx.x; i < 3 * (y_end - y_begin); i += kRzThreads) s_ytab[i] = ytab[3 * y_begin + i];
  __syncthreads();
  const int x0 = (blockIdx.x * kRzThreads + threadIdx.x) * kRzPx;
  if (x0 >= W) return;
  int xm[kRzPx], xm1[kRzPx], kx0[kRzPx], kx1[kRzPx];
#pragma unroll
  for (int j = 0; j < kRzPx; ++j) {
    const int x = min(x0 + j, W - 1);
    xm[j] = xtab[3 * x]; kx0[j] = xtab[3 * x + 1]; kx1[j] = xtab[3 * x + 2];
    xm1[j] = min(xm[j] + 1, sw - 1);
  }
  int cur_ym = -1, h0[kRzPx], h1[kRzPx];
  const bool vec = (x0 + kRzPx <= W) && ((W & 3) == 0);
  for (int y = y_begin; y < y_end; ++y) {
    const int* yt = s_ytab + 3 * (y - y_begin);
    const int ym = yt[0], ky0 = yt[1], ky1 = yt[2];
    if (ym != cur_ym) {
      cur_ym = ym;
      const uint8_t* r0 = s_src + ym * sw;
      const uint8_t* r1 = s_src + min(ym + 1, sh - 1) * sw;
#pragma unroll
      for (int j = 0; j < kRzPx; ++j) {
        h0[j] = clip8((r0[xm[j]] * kx0[j] + r0[xm1[j]] * kx1[j] + (1 << 21)) >> 22);
        h1[j] = clip8((r1[xm[j]] * kx0[j] + r1[xm1[j]] * kx1[j] + (1 << 21)) >> 22);
      }
    }
    float o[kRzPx];
#pragma unroll
    for (int j = 0; j < kRzPx; ++j) o[j] = s_lut[clip8((h0[j] * ky0 + h1[j] * ky1 + (1 << 21)) >> 22)];
    float* d = dst + ((size_t)b * H + y) * W + x0;
    if (vec) {
      *reinterpret_cast<float4*>(d) = make_float4(o[0], o[1], o[2], o[3]);
    } else {
#pragma unroll
      for (int j = 0; j < kRzPx; ++j)
        if (x0 + j < W) d[j] = o[j];
    }
  }
}

// -------------------------------------------------------------- conv1 fwd ---
// K3 + K4-stats: direct 5x5 conv, Cin=1 -> 16 channels, fp32 FFMA with the 400
// weights as constant-bank operands.  Each thread owns 4 consecutive pixels of
// one row (a 5x8 register window) and all 16 channels; output is bf16 NHWC and
// the per-channel sum / sum-of-squares of the ROUNDED outputs are accumulated
// per thread and reduced once per block (deterministic: partials -> finalize).
constexpr int kT1H = 16, kT1W = 64;

__global__ void __launch_bounds__(256, 2)
conv1_fwd_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ x1,
                 float* __restrict__ partials, int B, int H, int W, int tilesX, int tilesY) {
  __shared__ __align__(16) float s_in[kT1H + 4][kT1W + 4];
  __shared__ float s_red[8][2 * kC1];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float ssum[kC1], ssq[kC1];
#pragma unroll
  for (int c = 0; c < kC1; ++c) { ssum[c] = 0.f; ssq[c] = 0.f; }

  const long total = (long)B * tilesX * tilesY;
  for (long tile = blockIdx.x; tile < total; tile += gridDim.x) {
    const int b = (int)(tile / (tilesX * tilesY));
    const int r = (int)(tile % (tilesX * tilesY));
    const int y0 = (r / tilesX) * kT1H, x0 = (r % tilesX) * kT1W;
    __syncthreads();
    for (int i = threadIdx.x; i < (kT1H + 4) * (kT1W + 4); i += 256) {
      int ry = i / (kT1W + 4), rx = i % (kT1W + 4);
      int gy = y0 + ry - 2, gx = x0 + rx - 2;
      float v = 0.f;
      if (gy >= 0 && gy < H && gx >= 0 && gx < W) v = __ldg(in + ((size_t)b * H + gy) * W + gx);
      s_in[ry][rx] = v;
    }
    __syncthreads();
    const int gy = y0 + ty, gx = x0 + 4 * tx;
    const bool valid = (gy < H) && (gx < W);   // W % 4 == 0: a 4-pixel group is all-or-nothing
    float win[5][8];
#pragma unroll
    for (int dy = 0; dy < 5; ++dy) {
      float4 a = *reinterpret_cast<const float4*>(&s_in[ty + dy][4 * tx]);
      float4 c = *reinterpret_cast<const float4*>(&s_in[ty + dy][4 * tx + 4]);
      win[dy][0] = a.x; win[dy][1] = a.y; win[dy][2] = a.z; win[dy][3] = a.w;
      win[dy][4] = c.x; win[dy][5] = c.y; win[dy][6] = c.z; win[dy][7] = c.w;
    }
    uint32_t packed[4][kC1 / 2];
#pragma unroll
    for (int c = 0; c < kC1; c += 2) {
      float acc[2][4];
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
#pragma unroll
        for (int p = 0; p < 4; ++p) acc[cc][p] = c_b1[c + cc];
#pragma unroll
        for (int dy = 0; dy < 5; ++dy)
#pragma unroll
          for (int dx = 0; dx < 5; ++dx) {
            const float w = c_w1[(c + cc) * kTaps + dy * 5 + dx];
#pragma unroll
            for (int p = 0; p < 4; ++p) acc[cc][p] = fmaf(win[dy][dx + p], w, acc[cc][p]);
          }
      }
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        uint32_t pk = pack_bf16x2(acc[0][p], acc[1][p]);
        packed[p][c / 2] = pk;
        if (valid) {
          float r0 = bf16_lo(pk), r1 = bf16_hi(pk);
          ssum[c] += r0; ssq[c] = fmaf(r0, r0, ssq[c]);
          ssum[c + 1] += r1; ssq[c + 1] = fmaf(r1, r1, ssq[c + 1]);
        }
      }
    }
    if (valid) {
      uint4* o = reinterpret_cast<uint4*>(x1 + (((size_t)b * H + gy) * W + gx) * kC1);
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        o[2 * p] = make_uint4(packed[p][0], packed[p][1], packed[p][2], packed[p][3]);
        o[2 * p + 1] = make_uint4(packed[p][4], packed[p][5], packed[p][6], packed[p][7]);
      }
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int c = 0; c < kC1; ++c) {
    float a = warp_sum(ssum[c]), q = warp_sum(ssq[c]);
    if (lane == 0) { s_red[warp][c] = a; s_red[warp][kC1 + c] = q; }
  }
  __syncthreads();
  if (threadIdx.x < 2 * kC1) {
    float a = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) a += s_red[w][threadIdx.x];
    partials[(size_t)blockIdx.x * 2 * kC1 + threadIdx.x] = a;
  }
}

// ------------------------------------------------------------ BN finalize ---
// K4: per-channel batch statistics from the conv epilogue partials (summed in
// fp64), the fused affine (scale = gamma*rstd, shift = beta - mean*scale) and
// the running-stat update torch's BatchNorm2d does in train mode: momentum
// 0.1, UNBIASED variance, num_batches_tracked += 1.
// out = [scale | shift | mean | rstd], each C floats.
__global__ void bn_finalize_kernel(const float* __restrict__ partials, int nparts, int C, double count,
                                   const float* __restrict__ gamma, const float* __restrict__ beta,
                                   float* running_mean, float* running_var, long long* nbt,
                                   float momentum, float eps, float* __restrict__ out) {
  const int c = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (c >= C) return;
  double s = 0.0, q = 0.0;
  for (int i = lane; i < nparts; i += 32) {
    s += (double)partials[(size_t)i * 2 * C + c];
    q += (double)partials[(size_t)i * 2 * C + C + c];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (lane == 0) {
    double mean = s / count;
    double var = q / count - mean * mean;
    if (var < 0.0) var = 0.0;
    double rstd = 1.0 / sqrt(var + (double)eps);
    float sc = (float)((double)gamma[c] * rstd);
    out[c] = sc;
    out[C + c] = (float)((double)beta[c] - mean * (double)gamma[c] * rstd);
    out[2 * C + c] = (float)mean;
    out[3 * C + c] = (float)rstd;
    if (running_mean) {
      double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
      running_mean[c] = (float)((1.0 - momentum) * (double)running_mean[c] + momentum * mean);
      running_var[c] = (float)((1.0 - momentum) * (double)running_var[c] + momentum * unbiased);
    }
    if (c == 0 && nbt) *nbt += 1;
  }
}

// ------------------------------------------------- BN-apply + ReLU + pool ---
// K5/K7: one pass, bf16 NHWC in -> bf16 NHWC out at quarter size; no argmax
// tensor is written (the reference's ATen max-pool stores int64 indices, 8 B
// per output element).  One thread = one pooled pixel x 8 channels (16 B).
template <int C, typename T>
__global__ void __launch_bounds__(256)
bn_relu_pool_kernel(const T* __restrict__ x, const float* __restrict__ bn, T* __restrict__ p, int B, int H, int W) {
  __shared__ float s_sc[C], s_sh[C];
  if (threadIdx.x < C) { s_sc[threadIdx.x] = bn[threadIdx.x]; s_sh[threadIdx.x] = bn[C + threadIdx.x]; }
  __syncthreads();
  constexpr int V = Act<T>::kN, CH = C / V;
  const int Ho = H / 2, Wo = W / 2;
  const size_t total = (size_t)B * Ho * Wo * CH;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (size_t)gridDim.x * blockDim.x) {
    const int ch = (int)(idx % CH);
    size_t pix = idx / CH;
    const int xo = (int)(pix % Wo); pix /= Wo;
    const int yo = (int)(pix % Ho);
    const int b = (int)(pix / Ho);
    const T* base = x + (((size_t)b * H + 2 * yo) * W + 2 * xo) * C + ch * V;
    uint4 v[4];
    v[0] = ldg_stream16(base);
    v[1] = ldg_stream16(base + C);
    v[2] = ldg_stream16(base + (size_t)W * C);
    v[3] = ldg_stream16(base + (size_t)W * C + C);
    float f[4][V], o[V];
#pragma unroll
    for (int i = 0; i < 4; ++i) Act<T>::unpack(v[i], f[i]);
#pragma unroll
    for (int e = 0; e < V; ++e) {
      float m = 0.f;   // ReLU folded into the max: start from 0
#pragma unroll
      for (int i = 0; i < 4; ++i) m = fmaxf(m, fmaf(f[i][e], s_sc[ch * V + e], s_sh[ch * V + e]));
      o[e] = m;
    }
    *reinterpret_cast<uint4*>(p + (((size_t)b * Ho + yo) * Wo + xo) * C + ch * V) = Act<T>::pack(o);
  }
}

// ------------------------------------ pool + ReLU + BN backward (two modes) --
// K11/K13: the gradient of the pooled tensor is routed to the arg-max of each
// 2x2 window (first maximum in row-major order, like ATen) when that maximum
// is positive.  The arg-max is recomputed from the stored pre-BN tensor.
//   MODE 0 (stats): per-channel  sum(dy), sum(dy * xhat)  -> partials
//   MODE 1 (apply): dx = scale * (dy - mean(dy) - xhat * mean(dy*xhat)) -> bf16
// bn = [scale | shift | mean | rstd]; red = [mean(dy) | mean(dy*xhat)].
template <int C, int MODE, typename T>
__global__ void __launch_bounds__(256)
pool_bn_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dp,
                   const float* __restrict__ bn, const float* __restrict__ red,
                   float* __restrict__ partials, T* __restrict__ dx, int B, int H, int W) {
  constexpr int V = Act<T>::kN, CH = C / V;
  __shared__ float s_bn[4 * C];
  __shared__ float s_red[2 * C];
  __shared__ float s_part[8][2 * C];
  for (int i = threadIdx.x; i < 4 * C; i += blockDim.x) s_bn[i] = bn[i];
  if (MODE == 1) for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) s_red[i] = red[i];
  __syncthreads();
  const int Ho = H / 2, Wo = W / 2;
  const size_t total = (size_t)B * Ho * Wo * CH;
  const int ch = threadIdx.x % CH;   // blockDim (256) is a multiple of CH: fixed per thread
  float a_dy[V], a_dyx[V];
#pragma unroll
  for (int e = 0; e < V; ++e) { a_dy[e] = 0.f; a_dyx[e] = 0.f; }

  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (size_t)gridDim.x * blockDim.x) {
    size_t pix = idx / CH;
    const int xo = (int)(pix % Wo); pix /= Wo;
    const int yo = (int)(pix % Ho);
    const int b = (int)(pix / Ho);
    const size_t off0 = (((size_t)b * H + 2 * yo) * W + 2 * xo) * C + ch * V;
    const size_t offs[4] = {off0, off0 + C, off0 + (size_t)W * C, off0 + (size_t)W * C + C};
    uint4 v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = ldg_stream16(x + offs[i]);
    const uint4 g4 = ldg_stream16(dp + (((size_t)b * Ho + yo) * Wo + xo) * C + ch * V);
    float xf[4][V], gf[V], o[4][V];
#pragma unroll
    for (int i = 0; i < 4; ++i) Act<T>::unpack(v[i], xf[i]);
    Act<T>::unpack(g4, gf);
#pragma unroll
    for (int e = 0; e < V; ++e) {
      const int c = ch * V + e;
      const float sc = s_bn[c], sh = s_bn[C + c], mu = s_bn[2 * C + c], rs = s_bn[3 * C + c];
      float xv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) xv[i] = xf[i][e];
      int am = 0;
      float best = fmaf(xv[0], sc, sh);
#pragma unroll
      for (int i = 1; i < 4; ++i) {
        float y = fmaf(xv[i], sc, sh);
        if (y > best) { best = y; am = i; }
      }
      const float g = best > 0.f ? gf[e] : 0.f;
      if (MODE == 0) {
        float xa = xv[0];
#pragma unroll
        for (int i = 1; i < 4; ++i) xa = (am == i) ? xv[i] : xa;
        a_dy[e] += g;
        a_dyx[e] = fmaf(g, (xa - mu) * rs, a_dyx[e]);
      } else {
        const float mdy = s_red[c], mdyx = s_red[C + c];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float xh = (xv[i] - mu) * rs;
          o[i][e] = sc * (((am == i) ? g : 0.f) - mdy - xh * mdyx);
        }
      }
    }
    if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 4; ++i) *reinterpret_cast<uint4*>(dx + offs[i]) = Act<T>::pack(o[i]);
    }
  }
  if (MODE == 0) {
    // threads with the same channel chunk are CH apart; reduce across the warp
    // over lanes that share (lane % CH), then across warps through smem.
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int e = 0; e < V; ++e) {
      float a = a_dy[e], q = a_dyx[e];
#pragma unroll
      for (int o2 = 16; o2 >= CH; o2 >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o2);
        q += __shfl_xor_sync(0xffffffffu, q, o2);
      }
      if (lane < CH) { s_part[warp][lane * V + e] = a; s_part[warp][C + lane * V + e] = q; }
    }
    __syncthreads();
    if (threadIdx.x < 2 * C) {
      float a = 0.f;
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) a += s_part[w][threadIdx.x];
      partials[(size_t)blockIdx.x * 2 * C + threadIdx.x] = a;
    }
  }
}

// red = [mean(dy) | mean(dy*xhat)], dgamma = sum(dy*xhat), dbeta = sum(dy).
__global__ void bn_bwd_finalize_kernel(const float* __restrict__ partials, int nparts, int C, double count,
                                       float* __restrict__ red, float* __restrict__ dgamma,
                                       float* __restrict__ dbeta, float gscale) {
  // thread = (row group, column of the [nparts][2C] partial matrix): every load of a warp is one contiguous row
  // piece (a warp per channel striding over rows touched 32 lines per load: 24 us for 1184 rows)
  __shared__ double s_acc[1024];
  const int ncol = 2 * C, ngrp = (int)blockDim.x / ncol;
  const int col = (int)threadIdx.x % ncol, grp = (int)threadIdx.x / ncol;
  double a0 = 0.0, a1 = 0.0;
  int i = grp;
  for (; i + ngrp < nparts; i += 2 * ngrp) {
    a0 += (double)partials[(size_t)i * ncol + col];
    a1 += (double)partials[(size_t)(i + ngrp) * ncol + col];
  }
  if (i < nparts) a0 += (double)partials[(size_t)i * ncol + col];
  s_acc[threadIdx.x] = a0 + a1;
  __syncthreads();
  if ((int)threadIdx.x < ncol) {
    double t = 0.0;
    for (int g2 = 0; g2 < ngrp; ++g2) t += s_acc[g2 * ncol + threadIdx.x];
    const int c = (int)threadIdx.x % C;
    if ((int)threadIdx.x < C) {
      red[c] = (float)(t / count);
      dbeta[c] = (float)(t * (double)gscale);
    } else {
      red[C + c] = (float)(t / count);
      dgamma[c] = (float)(t * (double)gscale);
    }
  }
}

// ------------------------------------------------------------ conv1 wgrad ---
// K13: pool1/ReLU1/BN1 backward fused with the conv1 weight gradient.  The
// dense conv1-output gradient (16 x 9M per image) is never materialised: each
// thread rebuilds it for a 2x2 window x 4 channels from (x1, dp1, BN stats) and
// immediately contracts it with the 6x6 input patch: 400 FMA per window.
// partials[block][c*25 + tap].
constexpr int kWgT = 32;   // full-res tile edge

template <typename T>
__global__ void __launch_bounds__(256, 1)
conv1_wgrad_kernel(const float* __restrict__ in, const T* __restrict__ x1,
                   const T* __restrict__ dp1, const float* __restrict__ bn,
                   const float* __restrict__ red, float* __restrict__ partials,
                   int B, int H, int W, int tilesX, int tilesY) {
  __shared__ float s_in[kWgT + 4][kWgT + 5];
  __shared__ float s_acc[8][4][100];
  const int cg = threadIdx.x & 3, wl = threadIdx.x >> 2;
  float sc[4], sh[4], mu[4], rs[4], mdy[4], mdyx[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const int ch = cg * 4 + c;
    sc[c] = bn[ch]; sh[c] = bn[kC1 + ch]; mu[c] = bn[2 * kC1 + ch]; rs[c] = bn[3 * kC1 + ch];
    mdy[c] = red[ch]; mdyx[c] = red[kC1 + ch];
  }
  float acc[4][kTaps];
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int t = 0; t < kTaps; ++t) acc[c][t] = 0.f;

  const int Ho = H / 2, Wo = W / 2;
  const long total = (long)B * tilesX * tilesY;
  for (long tile = blockIdx.x; tile < total; tile += gridDim.x) {
    const int b = (int)(tile / (tilesX * tilesY));
    const int r = (int)(tile % (tilesX * tilesY));
    const int y0 = (r / tilesX) * kWgT, x0 = (r % tilesX) * kWgT;
    __syncthreads();
    for (int i = threadIdx.x; i < (kWgT + 4) * (kWgT + 4); i += 256) {
      int ry = i / (kWgT + 4), rx = i % (kWgT + 4);
      int gy = y0 + ry - 2, gx = x0 + rx - 2;
      float v = 0.f;
      if (gy >= 0 && gy < H && gx >= 0 && gx < W) v = __ldg(in + ((size_t)b * H + gy) * W + gx);
      s_in[ry][rx] = v;
    }
    __syncthreads();
#pragma unroll 1
    for (int k = 0; k < 4; ++k) {
      const int widx = wl + 64 * k;           // 16 x 16 windows per tile
      const int wy = widx >> 4, wx = widx & 15;
      const int gy = y0 + 2 * wy, gx = x0 + 2 * wx;
      if (gy >= H || gx >= W) continue;       // H, W even: windows are all-or-nothing
      float dxv[4][4];                        // [pixel][channel]
      {
        float xr[4][4], gr[4];
        const T* xb = x1 + (((size_t)b * H + gy) * W + gx) * kC1 + cg * 4;
        Act<T>::load4(xb, xr[0]);
        Act<T>::load4(xb + kC1, xr[1]);
        Act<T>::load4(xb + (size_t)W * kC1, xr[2]);
        Act<T>::load4(xb + (size_t)W * kC1 + kC1, xr[3]);
        Act<T>::load4(dp1 + (((size_t)b * Ho + (gy >> 1)) * Wo + (gx >> 1)) * kC1 + cg * 4, gr);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float xv[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) xv[i] = xr[i][c];
          int am = 0;
          float best = fmaf(xv[0], sc[c], sh[c]);
#pragma unroll
          for (int i = 1; i < 4; ++i) {
            float y = fmaf(xv[i], sc[c], sh[c]);
            if (y > best) { best = y; am = i; }
          }
          const float g = best > 0.f ? gr[c] : 0.f;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float xh = (xv[i] - mu[c]) * rs[c];
            dxv[i][c] = sc[c] * (((am == i) ? g : 0.f) - mdy[c] - xh * mdyx[c]);
          }
        }
      }
      float patch[6][6];
#pragma unroll
      for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j < 6; ++j) patch[i][j] = s_in[2 * wy + i][2 * wx + j];
#pragma unroll
      for (int py = 0; py < 2; ++py)
#pragma unroll
        for (int px = 0; px < 2; ++px)
#pragma unroll
          for (int dy = 0; dy < 5; ++dy)
#pragma unroll
            for (int dxx = 0; dxx < 5; ++dxx) {
              const float v = patch[py + dy][px + dxx];
#pragma unroll
              for (int c = 0; c < 4; ++c)
                acc[c][dy * 5 + dxx] = fmaf(dxv[py * 2 + px][c], v, acc[c][dy * 5 + dxx]);
            }
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int t = 0; t < kTaps; ++t) {
      float a = acc[c][t];
      a += __shfl_xor_sync(0xffffffffu, a, 4);
      a += __shfl_xor_sync(0xffffffffu, a, 8);
      a += __shfl_xor_sync(0xffffffffu, a, 16);
      if (lane < 4) s_acc[warp][lane][c * kTaps + t] = a;   // lane == cg here
    }
  __syncthreads();
  for (int i = threadIdx.x; i < 400; i += 256) {
    const int g = i / 100, j = i % 100;    // channel = g*4 + j/25, tap = j%25  -> index i
    float a = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) a += s_acc[w][g][j];
    partials[(size_t)blockIdx.x * 400 + i] = a;
  }
}

// out[j] = scale * sum_i partials[i][j]   (fp64 accumulation, deterministic)
__global__ void reduce_partials_kernel(const float* __restrict__ partials, int nparts, int len,
                                       float* __restrict__ out, float scale) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= len) return;
  double s = 0.0;
  for (int i = 0; i < nparts; ++i) s += (double)partials[(size_t)i * len + j];
  out[j] = (float)(s * (double)scale);
}

// ------------------------------------------------------------------ fc fwd ---
// K8: logits[b][k] = sum_j p2[b][j] * W[k][j].  M = batch (5), N = 10,
// K = 18,000,000: a pure weight-streaming problem (720 MB of fp32 W per pass),
// so the roofline is HBM bandwidth, not the tensor pipe.  Split-K over the
// grid; partial logits are reduced by fc_reduce_kernel.
constexpr int kNC = 10;   // classes

template <int BT, typename T>
__global__ void __launch_bounds__(256)
fc_fwd_kernel(const T* __restrict__ p2, const float* __restrict__ W,
              float* __restrict__ partials, size_t F) {
  float acc[BT][kNC];
#pragma unroll
  for (int b = 0; b < BT; ++b)
#pragma unroll
    for (int k = 0; k < kNC; ++k) acc[b][k] = 0.f;
  const size_t n4 = F / 4;
  for (size_t j = (size_t)blockIdx.x * 256 + threadIdx.x; j < n4; j += (size_t)gridDim.x * 256) {
    uint4 w[kNC];
    float a[BT][4];
#pragma unroll
    for (int k = 0; k < kNC; ++k) w[k] = ldg_stream16(W + (size_t)k * F + 4 * j);
#pragma unroll
    for (int b = 0; b < BT; ++b) Act<T>::load4(p2 + (size_t)b * F + 4 * j, a[b]);
#pragma unroll
    for (int b = 0; b < BT; ++b) {
      const float a0 = a[b][0], a1 = a[b][1], a2 = a[b][2], a3 = a[b][3];
#pragma unroll
      for (int k = 0; k < kNC; ++k) {
        float s = acc[b][k];
        s = fmaf(a0, __uint_as_float(w[k].x), s);
        s = fmaf(a1, __uint_as_float(w[k].y), s);
        s = fmaf(a2, __uint_as_float(w[k].z), s);
        s = fmaf(a3, __uint_as_float(w[k].w), s);
        acc[b][k] = s;
      }
    }
  }
  __shared__ float s_red[8][BT * kNC];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int b = 0; b < BT; ++b)
#pragma unroll
    for (int k = 0; k < kNC; ++k) {
      float v = warp_sum(acc[b][k]);
      if (lane == 0) s_red[warp][b * kNC + k] = v;
    }
  __syncthreads();
  if (threadIdx.x < BT * kNC) {
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) v += s_red[w][threadIdx.x];
    partials[(size_t)blockIdx.x * BT * kNC + threadIdx.x] = v;
  }
}

// one warp per output (a single thread walking 592 partials in fp64 took 40 us: a serial tail on the critical path)
__global__ void fc_reduce_kernel(const float* __restrict__ partials, int nparts, int n,
                                 const float* __restrict__ bias, float* __restrict__ logits) {
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (i >= n) return;
  double s = 0.0;
  for (int p = lane; p < nparts; p += 32) s += (double)partials[(size_t)p * n + i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) logits[i] = (float)s + bias[i % kNC];
}

// K9: CrossEntropyLoss(reduction='mean') forward + backward in one block:
// loss, dlogits = (softmax - onehot) * gscale / B, dbias = sum_b dlogits.
// gscale carries DDP's 1/world_size pre-division for the bias gradient only;
// dlogits itself stays unscaled (the bucket all-reduce applies 1/world).
__global__ void ce_kernel(const float* __restrict__ logits, const long long* __restrict__ labels, int B,
                          float* __restrict__ loss, float* __restrict__ dlogits, float* __restrict__ dbias,
                          float gscale) {
  __shared__ float s_loss;
  __shared__ float s_db[kNC];
  if (threadIdx.x == 0) s_loss = 0.f;
  if (threadIdx.x < kNC) s_db[threadIdx.x] = 0.f;
  __syncthreads();
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    float l[kNC], m = -INFINITY;
#pragma unroll
    for (int k = 0; k < kNC; ++k) { l[k] = logits[b * kNC + k]; m = fmaxf(m, l[k]); }
    float se = 0.f;
#pragma unroll
    for (int k = 0; k < kNC; ++k) { l[k] = expf(l[k] - m); se += l[k]; }
    const int y = (int)labels[b];
    const float lse = logf(se) + m;
    atomicAdd(&s_loss, (lse - logits[b * kNC + y]) / (float)B);
#pragma unroll
    for (int k = 0; k < kNC; ++k) {
      float d = (l[k] / se - (k == y ? 1.f : 0.f)) / (float)B;
      dlogits[b * kNC + k] = d;
      atomicAdd(&s_db[k], d);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) *loss = s_loss;
  if (threadIdx.x < kNC) dbias[threadIdx.x] = s_db[threadIdx.x] * gscale;
}

// ------------------------------------------------------------------ fc bwd ---
// K10 (+K14): one streaming pass over the 18M axis produces
//   dW[k][j] = gscale * sum_b dl[b][k] * p2[b][j]   (written straight into the
//              gradient bucket: in DDP that buffer is the symmetric-memory
//              bucket the all-reduce kernel reads)
//   dp2[b][j] = sum_k dl[b][k] * W[k][j]            (bf16)
// and, on the single-GPU path (lr != 0), applies SGD in place (W -= lr * dW)
// so the 720 MB gradient is never written or re-read.
// Optionally (bn2 != nullptr) it also produces the BatchNorm-2 backward reductions, which otherwise cost a
// separate pass over the 1.44 GB conv2 output: with dy = dp2 where the pooled activation is positive,
// sum(dy) and sum(dy * xhat) only involve the arg-max element of each pooling window, and its xhat follows
// from the pooled value itself: y = p2 = scale*x + shift  =>  xhat = ((p2 - shift)/scale - mean) * rstd.
// bn2 = [scale | shift | mean | rstd] (32 channels; channel = j % 32 in the HWC flatten order).
template <int BT, typename T>
__global__ void __launch_bounds__(256)
fc_bwd_kernel(const float* __restrict__ dl, const T* __restrict__ p2, float* W,
              float* dW, T* __restrict__ dp2, size_t F, float lr, int accumulate, float gscale,
              const float* __restrict__ bn2, float* __restrict__ bwd_partials,
              unsigned int* ready, int sweeps_per_slab) {
  __shared__ float s_dl[BT * kNC];
  __shared__ float s_bp[8][64];
  if (threadIdx.x < BT * kNC) s_dl[threadIdx.x] = dl[threadIdx.x];
  __syncthreads();
  // this thread always sees channels c0..c0+3 (the grid stride is a multiple of 8 quads = 32 channels)
  const int c0 = 4 * (threadIdx.x & 7);
  float k_shift[4], k_inv[4], k_mu[4], k_rs[4], a_dy[4], a_dyx[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    a_dy[i] = 0.f; a_dyx[i] = 0.f;
    const float sc = bn2 ? bn2[c0 + i] : 1.f;
    k_inv[i] = sc != 0.f ? 1.f / sc : 0.f;
    k_shift[i] = bn2 ? bn2[32 + c0 + i] : 0.f;
    k_mu[i] = bn2 ? bn2[64 + c0 + i] : 0.f;
    k_rs[i] = bn2 ? bn2[96 + c0 + i] : 0.f;
  }
  float d[BT][kNC];
#pragma unroll
  for (int b = 0; b < BT; ++b)
#pragma unroll
    for (int k = 0; k < kNC; ++k) d[b][k] = s_dl[b * kNC + k];
  const size_t n4 = F / 4;
  // `ready` (data-parallel runs): a "slab" is `sweeps_per_slab` consecutive grid sweeps = a contiguous column range
  // of all ten rows of dW.  When this block has written its share of slab t it bumps ready[t] (release, system
  // scope); the tile-driven all-reduce kernel (allreduce.cu) running beside this one - here and on every peer -
  // starts on a slab as soon as all blocks of all ranks have reported it.  Uniform trip count for the barrier.
  const size_t sweep = (size_t)gridDim.x * 256;
  int slab = 0, in_slab = 0;
  for (size_t jb = 0; jb < n4; jb += sweep) {
    const size_t j = jb + (size_t)blockIdx.x * 256 + threadIdx.x;
    if (j < n4) {
    float4 w[kNC];
#pragma unroll
    for (int k = 0; k < kNC; ++k) w[k] = *reinterpret_cast<const float4*>(W + (size_t)k * F + 4 * j);
    float av[BT][4];
#pragma unroll
    for (int b = 0; b < BT; ++b) Act<T>::load4(p2 + (size_t)b * F + 4 * j, av[b]);
    // dp2
#pragma unroll
    for (int b = 0; b < BT; ++b) {
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int k = 0; k < kNC; ++k) {
        s.x = fmaf(d[b][k], w[k].x, s.x); s.y = fmaf(d[b][k], w[k].y, s.y);
        s.z = fmaf(d[b][k], w[k].z, s.z); s.w = fmaf(d[b][k], w[k].w, s.w);
      }
      float gq[4] = {s.x, s.y, s.z, s.w};
      Act<T>::store4(dp2 + (size_t)b * F + 4 * j, gq);
      if (bn2) {     // BN2 backward reductions from the STORED (rounded) gradient: what the apply pass will read
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float y = av[b][i];
          const float gg = y > 0.f ? gq[i] : 0.f;
          a_dy[i] += gg;
          a_dyx[i] = fmaf(gg, ((y - k_shift[i]) * k_inv[i] - k_mu[i]) * k_rs[i], a_dyx[i]);
        }
      }
    }
    // dW (+ optional in-place SGD)
#pragma unroll
    for (int k = 0; k < kNC; ++k) {
      float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int b = 0; b < BT; ++b) {
        g.x = fmaf(d[b][k], av[b][0], g.x); g.y = fmaf(d[b][k], av[b][1], g.y);
        g.z = fmaf(d[b][k], av[b][2], g.z); g.w = fmaf(d[b][k], av[b][3], g.w);
      }
      if (dW) {
        float4 o = make_float4(g.x * gscale, g.y * gscale, g.z * gscale, g.w * gscale);
        if (accumulate == 2) {          // dW is a bf16 bucket (gradient exchange in bf16: half the bytes on the wire)
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(dW) + (size_t)k * F + 4 * j) =
              make_uint2(pack_bf16x2(o.x, o.y), pack_bf16x2(o.z, o.w));
        } else {
          float4* gp = reinterpret_cast<float4*>(dW + (size_t)k * F + 4 * j);
          if (accumulate) { float4 prev = *gp; o.x += prev.x; o.y += prev.y; o.z += prev.z; o.w += prev.w; }
          *gp = o;
        }
      }
      if (lr != 0.f) {
        float4 o = make_float4(fmaf(-lr, g.x, w[k].x), fmaf(-lr, g.y, w[k].y),
                               fmaf(-lr, g.z, w[k].z), fmaf(-lr, g.w, w[k].w));
        *reinterpret_cast<float4*>(W + (size_t)k * F + 4 * j) = o;
      }
    }
    }   // j < n4
    if (ready != nullptr && (++in_slab == sweeps_per_slab || jb + sweep >= n4)) {
      __threadfence_system();
      __syncthreads();
      if (threadIdx.x == 0)
        asm volatile("red.release.sys.global.add.u32 [%0], 1;" ::"l"(ready + slab) : "memory");
      ++slab;
      in_slab = 0;
    }
  }
  if (bn2) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float a = a_dy[i], q = a_dyx[i];
      a += __shfl_xor_sync(0xffffffffu, a, 8);  q += __shfl_xor_sync(0xffffffffu, q, 8);
      a += __shfl_xor_sync(0xffffffffu, a, 16); q += __shfl_xor_sync(0xffffffffu, q, 16);
      if (lane < 8) { s_bp[warp][4 * lane + i] = a; s_bp[warp][32 + 4 * lane + i] = q; }
    }
    __syncthreads();
    if (threadIdx.x < 64) {
      float a = 0.f;
#pragma unroll
      for (int w8 = 0; w8 < 8; ++w8) a += s_bp[w8][threadIdx.x];
      float* dst = bwd_partials + (size_t)blockIdx.x * 64 + threadIdx.x;
      *dst = accumulate ? *dst + a : a;
    }
  }
}

// Any-batch variant (batch > 8): the batch is a run-time loop INSIDE the pass over the 18M axis, so fc.weight
// (720 MB) is read once and dW written once per step whatever the batch - the templated kernel above keeps a batch
// chunk in registers and needs one full pass (plus a read-modify-write of dW) per 8 rows, which at batch 64 cost
// 11 ms of a 50 ms step.  dlogits sits in shared memory (broadcast reads); a thread owns 4 columns of all ten rows.
template <typename T>
__global__ void __launch_bounds__(256)
fc_bwd_anybatch_kernel(const float* __restrict__ dl, const T* __restrict__ p2, float* W, float* dW,
                       T* __restrict__ dp2, int B, size_t F, float lr, float gscale,
                       const float* __restrict__ bn2, float* __restrict__ bwd_partials, int dw_bf16) {
  extern __shared__ float s_dl_dyn[];                  // [B][12]: ten dlogits per row, padded to 48 bytes
  __shared__ float s_bp[8][64];
  for (int i = threadIdx.x; i < B * kNC; i += 256) s_dl_dyn[(i / kNC) * 12 + (i % kNC)] = dl[i];
  __syncthreads();
  const int c0 = 4 * (threadIdx.x & 7);
  float k_shift[4], k_inv[4], k_mu[4], k_rs[4], a_dy[4], a_dyx[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    a_dy[i] = 0.f; a_dyx[i] = 0.f;
    const float sc = bn2 ? bn2[c0 + i] : 1.f;
    k_inv[i] = sc != 0.f ? 1.f / sc : 0.f;
    k_shift[i] = bn2 ? bn2[32 + c0 + i] : 0.f;
    k_mu[i] = bn2 ? bn2[64 + c0 + i] : 0.f;
    k_rs[i] = bn2 ? bn2[96 + c0 + i] : 0.f;
  }
  const size_t n4 = F / 4;
  for (size_t j = (size_t)blockIdx.x * 256 + threadIdx.x; j < n4; j += (size_t)gridDim.x * 256) {
    float4 w[kNC], g[kNC];
#pragma unroll
    for (int k = 0; k < kNC; ++k) {
      w[k] = *reinterpret_cast<const float4*>(W + (size_t)k * F + 4 * j);
      g[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // rows in chunks of four, the next chunk's loads in flight while this one is used (rows are F apart in memory:
    // every load is its own DRAM burst, so memory-level parallelism has to come from the row loop); two named
    // buffers, because a run-time index into a register array would put it in local memory
    float avA[4][4], avB[4][4];
    auto load_chunk = [&](int b0, float (&dst)[4][4]) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
        if (b0 + r < B) Act<T>::load4(p2 + (size_t)(b0 + r) * F + 4 * j, dst[r]);
    };
    auto use_chunk = [&](int b0, float (&av)[4][4]) {
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int b = b0 + r;
        if (b < B) {
          const float4 d0 = *reinterpret_cast<const float4*>(s_dl_dyn + b * 12);
          const float4 d1 = *reinterpret_cast<const float4*>(s_dl_dyn + b * 12 + 4);
          const float2 d2 = *reinterpret_cast<const float2*>(s_dl_dyn + b * 12 + 8);
          const float d[kNC] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w, d2.x, d2.y};
          float4 sacc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int k = 0; k < kNC; ++k) {
            sacc.x = fmaf(d[k], w[k].x, sacc.x); sacc.y = fmaf(d[k], w[k].y, sacc.y);
            sacc.z = fmaf(d[k], w[k].z, sacc.z); sacc.w = fmaf(d[k], w[k].w, sacc.w);
            g[k].x = fmaf(d[k], av[r][0], g[k].x); g[k].y = fmaf(d[k], av[r][1], g[k].y);
            g[k].z = fmaf(d[k], av[r][2], g[k].z); g[k].w = fmaf(d[k], av[r][3], g[k].w);
          }
          float gq[4] = {sacc.x, sacc.y, sacc.z, sacc.w};
          Act<T>::store4(dp2 + (size_t)b * F + 4 * j, gq);
          if (bn2) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float y = av[r][i];
              const float gg = y > 0.f ? gq[i] : 0.f;
              a_dy[i] += gg;
              a_dyx[i] = fmaf(gg, ((y - k_shift[i]) * k_inv[i] - k_mu[i]) * k_rs[i], a_dyx[i]);
            }
          }
        }
      }
    };
    load_chunk(0, avA);
    for (int b0 = 0; b0 < B; b0 += 8) {
      load_chunk(b0 + 4, avB);
      use_chunk(b0, avA);
      load_chunk(b0 + 8, avA);
      use_chunk(b0 + 4, avB);
    }
#pragma unroll
    for (int k = 0; k < kNC; ++k) {
      if (dW) {
        if (dw_bf16)
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(dW) + (size_t)k * F + 4 * j) =
              make_uint2(pack_bf16x2(g[k].x * gscale, g[k].y * gscale), pack_bf16x2(g[k].z * gscale, g[k].w * gscale));
        else
          *reinterpret_cast<float4*>(dW + (size_t)k * F + 4 * j) =
              make_float4(g[k].x * gscale, g[k].y * gscale, g[k].z * gscale, g[k].w * gscale);
      }
      if (lr != 0.f)
        *reinterpret_cast<float4*>(W + (size_t)k * F + 4 * j) =
            make_float4(fmaf(-lr, g[k].x, w[k].x), fmaf(-lr, g[k].y, w[k].y), fmaf(-lr, g[k].z, w[k].z),
                        fmaf(-lr, g[k].w, w[k].w));
    }
  }
  if (bn2) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float a = a_dy[i], q = a_dyx[i];
      a += __shfl_xor_sync(0xffffffffu, a, 8);  q += __shfl_xor_sync(0xffffffffu, q, 8);
      a += __shfl_xor_sync(0xffffffffu, a, 16); q += __shfl_xor_sync(0xffffffffu, q, 16);
      if (lane < 8) { s_bp[warp][4 * lane + i] = a; s_bp[warp][32 + 4 * lane + i] = q; }
    }
    __syncthreads();
    if (threadIdx.x < 64) {
      float a = 0.f;
#pragma unroll
      for (int w8 = 0; w8 < 8; ++w8) a += s_bp[w8][threadIdx.x];
      bwd_partials[(size_t)blockIdx.x * 64 + threadIdx.x] = a;
    }
  }
}

// K14: p -= lr * g (vectorised; n need not be a multiple of 4).
__global__ void sgd_kernel(float* __restrict__ p, const float* __restrict__ g, size_t n, float lr) {
  const size_t n4 = n / 4;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = reinterpret_cast<const float4*>(g)[i];
    pv.x = fmaf(-lr, gv.x, pv.x); pv.y = fmaf(-lr, gv.y, pv.y);
    pv.z = fmaf(-lr, gv.z, pv.z); pv.w = fmaf(-lr, gv.w, pv.w);
    reinterpret_cast<float4*>(p)[i] = pv;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    size_t i = n4 * 4 + threadIdx.x;
    p[i] = fmaf(-lr, g[i], p[i]);
  }
}

// fp32 NCHW-flatten <-> internal HWC-flatten permutation of fc.weight, used by
// checkpoint import/export (state_dict parity) and by the
// torch-module view of the engine.  dst[k][ (y*Wd + x)*C + c ] = src[k][ (c*Hd + y)*Wd + x ]
__global__ void fcw_chw_to_hwc_kernel(const float* __restrict__ src, float* __restrict__ dst,
                                      int C, int Hd, int Wd, int to_hwc) {
  const size_t F = (size_t)C * Hd * Wd;
  const size_t total = (size_t)kNC * F;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t k = i / F, j = i % F;        // j indexes the HWC side
    const int c = (int)(j % C);
    const size_t pix = j / C;
    const size_t chw = (size_t)c * Hd * Wd + pix;
    if (to_hwc) dst[k * F + j] = src[k * F + chw];
    else dst[k * F + chw] = src[k * F + j];
  }
}

// bf16 NHWC <-> fp32 NCHW activations (test / interop helper; not on the hot path)
__global__ void nhwc_bf16_to_nchw_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst,
                                             int B, int C, int H, int W) {
  const size_t total = (size_t)B * C * H * W;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    size_t pix = i / C;
    const int x = (int)(pix % W); pix /= W;
    const int y = (int)(pix % H);
    const int b = (int)(pix / H);
    dst[(((size_t)b * C + c) * H + y) * W + x] = __bfloat162float(src[i]);
  }
}
__global__ void nchw_f32_to_nhwc_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                             int B, int C, int H, int W) {
  const size_t total = (size_t)B * C * H * W;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    size_t pix = i / C;
    const int x = (int)(pix % W); pix /= W;
    const int y = (int)(pix % H);
    const int b = (int)(pix / H);
    dst[i] = __float2bfloat16_rn(src[(((size_t)b * C + c) * H + y) * W + x]);
  }
}

inline int grid_for(size_t work, int block, int max_blocks) {
  size_t g = (work + block - 1) / block;
  if (g < 1) g = 1;
  if (g > (size_t)max_blocks) g = max_blocks;
  return (int)g;
}

}  // namespace

// ============================================================ C ABI wrappers =
// (declarations + reference citations: include/b200dp.h)

B200_API int b200_resize_pil_bilinear(const void* src_u8, int B, int sh, int sw, const void* xtab,
                                      const void* ytab, void* dst_f32, int H, int W, void* stream) {
  if (sh * sw > kRzMaxSrc) return b200_fail(-2, "resize: source image larger than 64x64");
  // one wave: as many row bands as fill the resident-block slots (16 blocks of 128 threads per SM)
  const int gx = (W + kRzThreads * kRzPx - 1) / (kRzThreads * kRzPx);
  const int bands = std::max(1, (b200_num_sms() * 16) / std::max(1, gx * B));
  const int rows = std::min(kRzMaxRows, std::max(8, (H + bands - 1) / bands));
  dim3 grid(gx, (H + rows - 1) / rows, B);
  resize_pil_bilinear_kernel<<<grid, kRzThreads, 0, (cudaStream_t)stream>>>(
      (const uint8_t*)src_u8, sh, sw, (const int*)xtab, (const int*)ytab, (float*)dst_f32, H, W, rows);
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_conv1_fwd_grid() { return b200_num_sms() * 4; }

B200_API int b200_conv1_fwd(const void* in_f32, const void* w1, const void* b1, void* x1_bf16,
                            void* partials, int B, int H, int W, void* stream) {
  B200_CHECK(H % 4 == 0 && W % 4 == 0, "conv1_fwd: H and W must be multiples of 4 (got %d x %d)", H, W);
  cudaStream_t s = (cudaStream_t)stream;
  B200_CUDA(cudaMemcpyToSymbolAsync(c_w1, w1, sizeof(float) * kC1 * kTaps, 0, cudaMemcpyDeviceToDevice, s));
  B200_CUDA(cudaMemcpyToSymbolAsync(c_b1, b1, sizeof(float) * kC1, 0, cudaMemcpyDeviceToDevice, s));
  const int tilesX = (W + kT1W - 1) / kT1W, tilesY = (H + kT1H - 1) / kT1H;
  conv1_fwd_kernel<<<b200_conv1_fwd_grid(), 256, 0, s>>>((const float*)in_f32, (__nv_bfloat16*)x1_bf16,
                                                         (float*)partials, B, H, W, tilesX, tilesY);
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_bn_finalize(const void* partials, int nparts, int C, double count, const void* gamma,
                              const void* beta, void* running_mean, void* running_var, void* nbt,
                              float momentum, float eps, void* out, void* stream) {
  B200_CHECK(C <= 32, "bn_finalize: C must be <= 32");
  bn_finalize_kernel<<<1, 32 * C, 0, (cudaStream_t)stream>>>(
      (const float*)partials, nparts, C, count, (const float*)gamma, (const float*)beta,
      (float*)running_mean, (float*)running_var, (long long*)nbt, momentum, eps, (float*)out);
  B200_LAUNCH_CHECK();
  return 0;
}

template <typename T>
static int bn_relu_pool_t(const void* x, const void* bn, void* p, int B, int H, int W, int C, void* stream) {
  B200_CHECK(H % 2 == 0 && W % 2 == 0, "bn_relu_pool: H and W must be even");
  const size_t work = (size_t)B * (H / 2) * (W / 2) * (C / Act<T>::kN);
  const int grid = grid_for(work, 256, b200_num_sms() * 16);
  cudaStream_t s = (cudaStream_t)stream;
  if (C == 16) bn_relu_pool_kernel<16, T><<<grid, 256, 0, s>>>((const T*)x, (const float*)bn, (T*)p, B, H, W);
  else if (C == 32) bn_relu_pool_kernel<32, T><<<grid, 256, 0, s>>>((const T*)x, (const float*)bn, (T*)p, B, H, W);
  else return b200_fail(-2, "bn_relu_pool: C must be 16 or 32");
  B200_LAUNCH_CHECK();
  return 0;
}
B200_API int b200_bn_relu_pool(const void* x, const void* bn, void* p, int B, int H, int W, int C, void* stream) {
  return bn_relu_pool_t<__nv_bfloat16>(x, bn, p, B, H, W, C, stream);
}
B200_API int b200_bn_relu_pool_f32(const void* x, const void* bn, void* p, int B, int H, int W, int C, void* stream) {
  return bn_relu_pool_t<float>(x, bn, p, B, H, W, C, stream);
}

B200_API int b200_pool_bn_bwd_grid() { return b200_num_sms() * 8; }

// mode 0: stats -> partials [grid][2C];  mode 1: apply -> dx
template <typename T>
static int pool_bn_bwd_t(const void* x, const void* dp, const void* bn, const void* red, void* partials,
                         void* dx, int B, int H, int W, int C, int mode, void* stream) {
  B200_CHECK(H % 2 == 0 && W % 2 == 0, "pool_bn_bwd: H and W must be even");
  const int grid = b200_pool_bn_bwd_grid();
  cudaStream_t s = (cudaStream_t)stream;
#define B200_PBB(CC, MM)                                                                              \
  pool_bn_bwd_kernel<CC, MM, T><<<grid, 256, 0, s>>>((const T*)x, (const T*)dp,  \
                                                  (const float*)bn, (const float*)red, (float*)partials, \
                                                  (T*)dx, B, H, W)
  if (C == 16 && mode == 0) B200_PBB(16, 0);
  else if (C == 16 && mode == 1) B200_PBB(16, 1);
  else if (C == 32 && mode == 0) B200_PBB(32, 0);
  else if (C == 32 && mode == 1) B200_PBB(32, 1);
  else return b200_fail(-2, "pool_bn_bwd: unsupported C/mode");
#undef B200_PBB
  B200_LAUNCH_CHECK();
  return 0;
}
B200_API int b200_pool_bn_bwd(const void* x, const void* dp, const void* bn, const void* red, void* partials,
                              void* dx, int B, int H, int W, int C, int mode, void* stream) {
  return pool_bn_bwd_t<__nv_bfloat16>(x, dp, bn, red, partials, dx, B, H, W, C, mode, stream);
}
B200_API int b200_pool_bn_bwd_f32(const void* x, const void* dp, const void* bn, const void* red, void* partials,
                                  void* dx, int B, int H, int W, int C, int mode, void* stream) {
  return pool_bn_bwd_t<float>(x, dp, bn, red, partials, dx, B, H, W, C, mode, stream);
}

B200_API int b200_bn_bwd_finalize(const void* partials, int nparts, int C, double count, void* red,
                                  void* dgamma, void* dbeta, float gscale, void* stream) {
  B200_CHECK(C >= 1 && C <= 32 && 1024 % (2 * C) == 0, "bn_bwd_finalize: C must divide 512 and be <= 32");
  bn_bwd_finalize_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>((const float*)partials, nparts, C, count,
                                                                  (float*)red, (float*)dgamma, (float*)dbeta, gscale);
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_conv1_wgrad_grid() { return b200_num_sms(); }

template <typename T>
static int conv1_wgrad_t(const void* in_f32, const void* x1, const void* dp1, const void* bn,
                         const void* red, void* partials, void* dw1, int B, int H, int W,
                         float gscale, void* stream) {
  B200_CHECK(H % 2 == 0 && W % 2 == 0, "conv1_wgrad: H and W must be even");
  cudaStream_t s = (cudaStream_t)stream;
  const int tilesX = (W + kWgT - 1) / kWgT, tilesY = (H + kWgT - 1) / kWgT;
  const int grid = b200_conv1_wgrad_grid();
  conv1_wgrad_kernel<T><<<grid, 256, 0, s>>>((const float*)in_f32, (const T*)x1, (const T*)dp1, (const float*)bn,
                                             (const float*)red, (float*)partials, B, H, W, tilesX, tilesY);
  B200_LAUNCH_CHECK();
  reduce_partials_kernel<<<(400 + 127) / 128, 128, 0, s>>>((const float*)partials, grid, 400, (float*)dw1, gscale);
  B200_LAUNCH_CHECK();
  return 0;
}
B200_API int b200_conv1_wgrad(const void* in_f32, const void* x1, const void* dp1, const void* bn,
                              const void* red, void* partials, void* dw1, int B, int H, int W,
                              float gscale, void* stream) {
  return conv1_wgrad_t<__nv_bfloat16>(in_f32, x1, dp1, bn, red, partials, dw1, B, H, W, gscale, stream);
}
B200_API int b200_conv1_wgrad_f32(const void* in_f32, const void* x1, const void* dp1, const void* bn,
                                  const void* red, void* partials, void* dw1, int B, int H, int W,
                                  float gscale, void* stream) {
  return conv1_wgrad_t<float>(in_f32, x1, dp1, bn, red, partials, dw1, B, H, W, gscale, stream);
}

// ---------------------------------------------------------------------------------------
// Allocation-free variant of the conv1 weight gradient.
//
// b200_conv1_wgrad needs [grid][400] floats of scratch for its per-block partial sums and
// leaves it to the caller to provide them.  Some callers cannot: code running inside a CUDA
// stream capture or under a memory-pool policy that forbids new allocations, a host
// application that hands this library a fixed memory budget up front, or a thin C/Fortran
// driver that only owns its input and output arrays.  For them the scratch lives here, in a
// file-scope device array that every GPU loading this library carries until the process exits
// (kNoAllocMaxBlocks * 400 floats = 256,000 bytes per GPU).
//
// In host code the array's name is only a host-side stand-in for the symbol; the device
// address is obtained once per device with cudaGetSymbolAddress (with that device current) and
// is what the kernel receives.
constexpr int kNoAllocMaxBlocks = 160;
static __device__ __align__(256) float g_conv1_wgrad_scratch[kNoAllocMaxBlocks * 400];
static float* g_conv1_wgrad_scratch_addr[64];   // per device ordinal; filled by the first call on that device
static std::mutex g_conv1_wgrad_scratch_mu;

B200_API int b200_conv1_wgrad_noalloc_max_blocks() { return kNoAllocMaxBlocks; }

// Same inputs and output as b200_conv1_wgrad, without the `partials` argument.
//   nblocks  number of thread blocks (= partial rows) to use; 0 picks the device's SM count.
//            Values above b200_conv1_wgrad_noalloc_max_blocks() are rejected: the built-in
//            scratch cannot hold them - use b200_conv1_wgrad with your own scratch instead.
// The first call on each device looks up the scratch address (no allocation, but a driver
// query): make that first call outside any stream capture.  All calls on a device share the one
// array, which keeps the previous call's partial sums: nothing in it has to be zero on entry
// (every one of the nblocks rows is overwritten by its block before the reduction reads it, rows
// beyond nblocks are never read), but two calls whose kernels overlap on the GPU - different
// streams, or different host threads - race on it and must be serialised by the caller.
// Makes no call to cudaMalloc, cudaFree or any other allocating routine.
B200_API int b200_conv1_wgrad_noalloc(const void* in_f32, const void* x1, const void* dp1, const void* bn,
                                      const void* red, void* dw1, int B, int H, int W, float gscale, int nblocks,
                                      void* stream) {
  B200_CHECK(H % 2 == 0 && W % 2 == 0, "conv1_wgrad_noalloc: H and W must be even");
  B200_CHECK(nblocks >= 0, "conv1_wgrad_noalloc: nblocks must be >= 0 (got %d)", nblocks);
  const int grid = nblocks == 0 ? b200_num_sms() : nblocks;
  B200_CHECK(grid <= kNoAllocMaxBlocks,
             "conv1_wgrad_noalloc: %d partial rows of 400 floats requested but the built-in scratch holds %d "
             "(%zu bytes); call b200_conv1_wgrad with caller-provided scratch for larger grids",
             grid, kNoAllocMaxBlocks, sizeof(float) * 400 * (size_t)kNoAllocMaxBlocks);
  int dev = -1;
  B200_CUDA(cudaGetDevice(&dev));
  B200_CHECK(dev >= 0 && dev < 64, "conv1_wgrad_noalloc: device ordinal %d out of range", dev);
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_conv1_wgrad_scratch_mu);
    if (g_conv1_wgrad_scratch_addr[dev] == nullptr) {          // set-up: once per device, that device current
      float* p = nullptr;
      B200_CUDA(cudaGetSymbolAddress((void**)&p, g_conv1_wgrad_scratch));
      B200_CHECK(p != nullptr, "conv1_wgrad_noalloc: cudaGetSymbolAddress returned a null address");
      g_conv1_wgrad_scratch_addr[dev] = p;
    }
    scratch = g_conv1_wgrad_scratch_addr[dev];
  }
  cudaStream_t s = (cudaStream_t)stream;
  const int tilesX = (W + kWgT - 1) / kWgT, tilesY = (H + kWgT - 1) / kWgT;
  conv1_wgrad_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((const float*)in_f32, (const __nv_bfloat16*)x1,
                                                         (const __nv_bfloat16*)dp1, (const float*)bn, (const float*)red,
                                                         scratch, B, H, W, tilesX, tilesY);
  B200_LAUNCH_CHECK();
  reduce_partials_kernel<<<(400 + 127) / 128, 128, 0, s>>>(scratch, grid, 400, (float*)dw1, gscale);
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_reduce_partials(const void* partials, int nparts, int len, void* out, float scale, void* stream) {
  reduce_partials_kernel<<<(len + 127) / 128, 128, 0, (cudaStream_t)stream>>>((const float*)partials, nparts, len,
                                                                              (float*)out, scale);
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_fc_grid() { return b200_num_sms() * 4; }

// logits[bt][10] for a batch chunk of bt <= 8 rows.
template <typename T>
static int fc_fwd_t(const void* p2, const void* W, const void* bias, void* partials, void* logits,
                    int bt, size_t F, void* stream) {
  B200_CHECK(bt >= 1 && bt <= 8, "fc_fwd: batch chunk must be 1..8 (got %d)", bt);
  B200_CHECK(F % 4 == 0, "fc_fwd: F must be a multiple of 4");
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = b200_fc_grid();
  switch (bt) {
#define B200_FCF(N) case N: fc_fwd_kernel<N, T><<<grid, 256, 0, s>>>((const T*)p2, (const float*)W, (float*)partials, F); break;
    B200_FCF(1) B200_FCF(2) B200_FCF(3) B200_FCF(4) B200_FCF(5) B200_FCF(6) B200_FCF(7) B200_FCF(8)
#undef B200_FCF
  }
  B200_LAUNCH_CHECK();
  fc_reduce_kernel<<<(bt * kNC * 32 + 255) / 256, 256, 0, s>>>((const float*)partials, grid, bt * kNC, (const float*)bias,
                                                              (float*)logits);
  B200_LAUNCH_CHECK();
  return 0;
}
B200_API int b200_fc_fwd(const void* p2, const void* W, const void* bias, void* partials, void* logits,
                         int bt, size_t F, void* stream) {
  return fc_fwd_t<__nv_bfloat16>(p2, W, bias, partials, logits, bt, F, stream);
}
B200_API int b200_fc_fwd_f32(const void* p2, const void* W, const void* bias, void* partials, void* logits,
                             int bt, size_t F, void* stream) {
  return fc_fwd_t<float>(p2, W, bias, partials, logits, bt, F, stream);
}

B200_API int b200_cross_entropy(const void* logits, const void* labels_i64, int B, void* loss, void* dlogits,
                                void* dbias, float gscale, void* stream) {
  ce_kernel<<<1, 256, 0, (cudaStream_t)stream>>>((const float*)logits, (const long long*)labels_i64, B,
                                                  (float*)loss, (float*)dlogits, (float*)dbias, gscale);
  B200_LAUNCH_CHECK();
  return 0;
}

// bn2 / bwd_partials (both nullable): fused BatchNorm-2 backward reductions, partials [b200_fc_grid()][64].
template <typename T>
static int fc_bwd_t(const void* dlogits, const void* p2, void* W, void* dW, void* dp2, int bt, size_t F,
                    float lr, int accumulate, float gscale, const void* bn2, void* bwd_partials, void* stream,
                    void* ready = nullptr, int sweeps_per_slab = 0) {
  B200_CHECK(bt >= 1 && bt <= 4096, "fc_bwd: batch must be 1..4096 (got %d)", bt);
  B200_CHECK(F % 4 == 0, "fc_bwd: F must be a multiple of 4");
  B200_CHECK(!bn2 || (F % 32 == 0 && bwd_partials), "fc_bwd: fused BN statistics need F %% 32 == 0 and a partials buffer");
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = b200_fc_grid();
  if (bt > 8) {          // any-batch kernel: one pass over W for the whole batch (no accumulate, no slab counters)
    B200_CHECK(accumulate != 1 && !ready, "fc_bwd: batches > 8 run in one pass (accumulate / slab counters do not apply)");
    const size_t smem = (size_t)bt * 12 * sizeof(float);
    B200_CUDA(set_max_smem(fc_bwd_anybatch_kernel<T>, (int)smem));
    fc_bwd_anybatch_kernel<T><<<grid, 256, smem, s>>>((const float*)dlogits, (const T*)p2, (float*)W, (float*)dW, (T*)dp2,
                                                     bt, F, lr, gscale, (const float*)bn2, (float*)bwd_partials,
                                                     accumulate == 2 ? 1 : 0);
    B200_LAUNCH_CHECK();
    return 0;
  }
  switch (bt) {
#define B200_FCB(N) case N: fc_bwd_kernel<N, T><<<grid, 256, 0, s>>>((const float*)dlogits, (const T*)p2, (float*)W, (float*)dW, (T*)dp2, F, lr, accumulate, gscale, (const float*)bn2, (float*)bwd_partials, (unsigned int*)ready, sweeps_per_slab); break;
    B200_FCB(1) B200_FCB(2) B200_FCB(3) B200_FCB(4) B200_FCB(5) B200_FCB(6) B200_FCB(7) B200_FCB(8)
#undef B200_FCB
  }
  B200_LAUNCH_CHECK();
  return 0;
}
B200_API int b200_fc_bwd(const void* dlogits, const void* p2, void* W, void* dW, void* dp2, int bt, size_t F,
                         float lr, int accumulate, float gscale, const void* bn2, void* bwd_partials, void* stream) {
  return fc_bwd_t<__nv_bfloat16>(dlogits, p2, W, dW, dp2, bt, F, lr, accumulate, gscale, bn2, bwd_partials, stream);
}
B200_API int b200_fc_bwd_f32(const void* dlogits, const void* p2, void* W, void* dW, void* dp2, int bt, size_t F,
                             float lr, int accumulate, float gscale, const void* bn2, void* bwd_partials, void* stream) {
  return fc_bwd_t<float>(dlogits, p2, W, dW, dp2, bt, F, lr, accumulate, gscale, bn2, bwd_partials, stream);
}

B200_API int b200_sgd(void* p, const void* g, size_t n, float lr, void* stream) {
  const int grid = grid_for(n / 4 + 1, 256, b200_num_sms() * 8);
  sgd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((float*)p, (const float*)g, n, lr);
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_fcw_permute(const void* src, void* dst, int C, int Hd, int Wd, int to_hwc, void* stream) {
  const size_t total = (size_t)kNC * C * Hd * Wd;
  const int grid = grid_for(total, 256, b200_num_sms() * 8);
  fcw_chw_to_hwc_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)src, (float*)dst, C, Hd, Wd, to_hwc);
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_nhwc_bf16_to_nchw_f32(const void* src, void* dst, int B, int C, int H, int W, void* stream) {
  const int grid = grid_for((size_t)B * C * H * W, 256, b200_num_sms() * 8);
  nhwc_bf16_to_nchw_f32_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)src, (float*)dst, B, C, H, W);
  B200_LAUNCH_CHECK();
  return 0;
}
B200_API int b200_nchw_f32_to_nhwc_bf16(const void* src, void* dst, int B, int C, int H, int W, void* stream) {
  const int grid = grid_for((size_t)B * C * H * W, 256, b200_num_sms() * 8);
  nchw_f32_to_nhwc_bf16_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)src, (__nv_bfloat16*)dst, B, C, H, W);
  B200_LAUNCH_CHECK();
  return 0;
}

// fc backward that reports its progress slab by slab for the tile-driven bucket all-reduce (b200_ar_tile_f32).
//   ready            [ceil(F / b200_fc_bwd_tile_slab_floats(sweeps_per_slab))] uint32 counters in symmetric memory;
//                    every block adds 1 to ready[t] once its part of slab t is written (monotonic: never reset)
//   dtype            0 = bf16 activations, 1 = fp32
// b200_fc_grid() blocks report per slab.
B200_API size_t b200_fc_bwd_tile_slab_floats(int sweeps_per_slab) {
  return (size_t)b200_fc_grid() * 256 * 4 * (size_t)sweeps_per_slab;
}
B200_API int b200_fc_bwd_tiles(const void* dlogits, const void* p2, void* W, void* dW, void* dp2, int bt, size_t F,
                               float gscale, const void* bn2, void* bwd_partials, void* ready, int sweeps_per_slab,
                               int dtype, void* stream) {
  B200_CHECK(ready != nullptr && sweeps_per_slab >= 1 && dW != nullptr, "fc_bwd_tiles: needs counters, a slab size and dW");
  if (dtype == 1)
    return fc_bwd_t<float>(dlogits, p2, W, dW, dp2, bt, F, 0.f, 0, gscale, bn2, bwd_partials, stream, ready, sweeps_per_slab);
  return fc_bwd_t<__nv_bfloat16>(dlogits, p2, W, dW, dp2, bt, F, 0.f, 0, gscale, bn2, bwd_partials, stream, ready,
                                 sweeps_per_slab);
}
