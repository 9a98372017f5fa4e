// Microbenchmark (GPU box): sustained cost of one tcgen05.mma (M = 128, K = 16, bf16) issued by one thread, for the
// shapes the attention kernel uses: SS with N = 64 / 128 / 256 and TS (A from TMEM) with N = 64.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../vitvs_b200/csrc/ptx.cuh"
using namespace vvs;

__device__ __forceinline__ void umma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d),
               "r"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

// mode 0: SS, B K-major (S = Q K^T);  mode 1: TS, B MN-major (O += P V)
template <int N, int MODE>
__global__ void __launch_bounds__(128, 1) k(float* out, int iters, int ctas_active) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = __shfl_sync(0xffffffffu, slot, 0);
  if (warp == 1 && (int)blockIdx.x < ctas_active) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, N, MODE == 1);
    const uint32_t a = smem_u32(smem), b = a + 16384;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          if (MODE == 0) umma_bf16(tb, umma_smem_desc(a + kk * 32, 16, 1024), umma_smem_desc(b + kk * 32, 16, 1024), idesc, 1);
          else umma_ts(tb + 256, tb + kk * 8, umma_smem_desc(b + kk * 2048, 1024, 1024), idesc, 1);
        }
        umma_commit(&bar);
      }
      __syncwarp();
      mbar_wait(&bar, it & 1);   // 4 MMAs in flight at a time (like one attention block), then the next 4
    }
    long long t1 = clock64();
    if ((threadIdx.x & 31) == 0 && blockIdx.x == 0) out[0] = float(t1 - t0) / (4.f * iters);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tb);
}

// same, but never waiting: MMAs and commits back to back (queue-limited issue rate)
template <int N, int MODE>
__global__ void __launch_bounds__(128, 1) kq(float* out, int iters) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = __shfl_sync(0xffffffffu, slot, 0);
  if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, N, MODE == 1);
    const uint32_t a = smem_u32(smem), b = a + 16384;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          if (MODE == 0) umma_bf16(tb, umma_smem_desc(a + kk * 32, 16, 1024), umma_smem_desc(b + kk * 32, 16, 1024), idesc, 1);
          else umma_ts(tb + 256, tb + kk * 8, umma_smem_desc(b + kk * 2048, 1024, 1024), idesc, 1);
        }
      }
      __syncwarp();
    }
    if ((threadIdx.x & 31) == 0) umma_commit(&bar);
    __syncwarp();
    mbar_wait(&bar, 0);
    long long t1 = clock64();
    if ((threadIdx.x & 31) == 0 && blockIdx.x == 0) out[0] = float(t1 - t0) / (4.f * iters);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tb);
}

template <int N, int MODE>
void run(const char* name, float* d) {
  const int smem = 16384 + 32768 + 1024;
  cudaFuncSetAttribute(k<N, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(kq<N, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  float a, b;
  k<N, MODE><<<148, 128, smem>>>(d, 2000, 148);
  cudaDeviceSynchronize();
  cudaMemcpy(&a, d, 4, cudaMemcpyDeviceToHost);
  kq<N, MODE><<<148, 128, smem>>>(d, 2000);
  cudaDeviceSynchronize();
  cudaMemcpy(&b, d, 4, cudaMemcpyDeviceToHost);
  printf("%-28s %7.1f clk/MMA in groups of 4 + commit + wait;  %7.1f clk/MMA back to back   (%s)\n", name, a, b,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  float* d;
  cudaMalloc(&d, 64);
  run<64, 0>("SS M128 N64  K16", d);
  run<128, 0>("SS M128 N128 K16", d);
  run<256, 0>("SS M128 N256 K16", d);
  run<64, 1>("TS M128 N64  K16 (B MN-major)", d);
  return 0;
}
