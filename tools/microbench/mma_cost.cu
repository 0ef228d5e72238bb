// tcgen05.mma issue/throughput microbenchmark: cycles per MMA for M in {64,128}, N in {16..256},
// smem layout in {no-swizzle K-major, no-swizzle MN-major, 128B-swizzle K-major}, same vs rotating
// accumulator.  Operand contents are irrelevant (smem is zeroed).  One CTA per SM to see contention-free numbers.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include "../../csrc/ptx.cuh"
using namespace b2;

__device__ __forceinline__ uint64_t desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  uint64_t d = smem_desc_noswz(saddr, lbo, sbo);
  d |= (uint64_t)layout << 61;
  return d;
}

__global__ void __launch_bounds__(128, 1) mma_cost(int M, int N, int mode, int rotate, int iters, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tptr;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  fence_proxy_async_smem();
  if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
  if (threadIdx.x < 32) tmem_alloc<512>(&tptr);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  if (threadIdx.x == 0) {
    const uint32_t base = (smem_u32(smem) + 1023) & ~1023u;
    uint64_t a, b; uint32_t idesc;
    if (mode == 0) {        // K-major, no swizzle: planes 16 KB apart, rows 16 B
      a = desc(base, 16384, 128, 0); b = desc(base + 65536, 16384, 128, 0); idesc = umma_idesc_bf16(M, N, 0, 0);
    } else if (mode == 1) { // MN-major, no swizzle
      a = desc(base, 128, 16, 0); b = desc(base + 65536, 128, 2048, 0); idesc = umma_idesc_bf16(M, N, 1, 1);
    } else {                // K-major, 128B swizzle (layout type 2): rows 128 B, 8-row atoms 1024 B
      a = desc(base, 16, 1024, 2); b = desc(base + 65536, 16, 1024, 2); idesc = umma_idesc_bf16(M, N, 0, 0);
    }
    // warm-up
    umma_bf16(tptr, a, b, idesc, 0); umma_commit(&bar); mbar_wait(&bar, 0);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      uint32_t d = tptr + (rotate ? (uint32_t)((i & 1) * 256) : 0u);
      umma_bf16(d, a, b, idesc, 1);
    }
    long long t1 = clock64();
    umma_commit(&bar); mbar_wait(&bar, 1);
    long long t2 = clock64();
    if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc<512>(tptr);
}

int main() {
  long long* out; cudaMallocManaged(&out, 16);
  cudaFuncSetAttribute(mma_cost, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const char* names[3] = {"K-major noswz", "MN-major noswz", "K-major sw128"};
  const int iters = 2000;
  for (int grid : {1, 148}) {
    printf("== grid %d\n", grid);
    for (int mode = 0; mode < 3; ++mode)
      for (int M : {64, 128})
        for (int N : {16, 32, 80, 96, 128, 160, 256})
          for (int rot = 0; rot < 2; ++rot) {
            if (rot == 1 && N > 256) continue;
            if (mode == 1 && M == 128 && N % 16) continue;
            mma_cost<<<grid, 128, 200 * 1024>>>(M, N, mode, rot, iters, out);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("%s M=%d N=%d: %s\n", names[mode], M, N, cudaGetErrorString(e)); return 1; }
            printf("%-15s M=%3d N=%3d rot=%d  issue %6.1f cyc/mma   complete %6.1f cyc/mma  (ideal math %5.1f)\n", names[mode], M,
                   N, rot, (double)out[0] / iters, (double)out[1] / iters, (M == 64 ? 128.0 : 128.0) * N / 256.0);
          }
  }
  return 0;
}
