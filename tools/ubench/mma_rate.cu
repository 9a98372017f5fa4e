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

// SS N = 256 MMAs back to back (the GEMM main loop's shape) while other warps write shared memory: warps 2/3 with
// st.shared.v4 (LSU path) or one lane with 1-D bulk copies from an L2-resident buffer (async proxy, the TMA path), at
// a throttled rate.  Separates "the MMA operand reads share the 128 B/clk port with the operand writes" from "the
// L2 -> SM fabric is the limit".  out[0] = clk / MMA, out[1] = bytes written per clk.
template <int WMODE>   // 0: none, 1: st.shared, 2: bulk copy
__global__ void __launch_bounds__(128, 1) kw(float* out, const uint8_t* src, int iters, int gap) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar, cbar[8];
  __shared__ uint32_t slot;
  __shared__ volatile int done;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    for (int i = 0; i < 8; ++i) mbar_init(&cbar[i], 1);
    done = 0;
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = __shfl_sync(0xffffffffu, slot, 0);
  uint8_t* ring = smem + 16384 + 32768;   // 8 x 16 KB landing zone, never read by the MMAs
  if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, 256, false);
    const uint32_t a = smem_u32(smem), b = a + 16384;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_bf16(tb, umma_smem_desc(a + kk * 32, 16, 1024), umma_smem_desc(b + kk * 32, 16, 1024), idesc, 1);
      }
      __syncwarp();
    }
    if ((threadIdx.x & 31) == 0) umma_commit(&bar);
    __syncwarp();
    mbar_wait(&bar, 0);
    long long t1 = clock64();
    done = 1;
    if ((threadIdx.x & 31) == 0 && blockIdx.x == 0) out[0] = float(t1 - t0) / (4.f * iters);
    if ((threadIdx.x & 31) == 0 && blockIdx.x == 0) out[2] = float(t1 - t0);
  } else if (WMODE == 1 && warp >= 2) {
    const uint32_t dst = smem_u32(ring) + (warp - 2) * 65536 + (threadIdx.x & 31) * 16;
    long long n = 0;
    while (!done) {
#pragma unroll
      for (int i = 0; i < 32; ++i) st_shared_v4(dst + i * 512, i, i, i, i);   // 32 x 512 B
      n += 16384;
      if (gap > 0) {
        const long long t = clock64();
        while (clock64() - t < gap) {
        }
      }
    }
    if ((threadIdx.x & 31) == 0 && blockIdx.x == 0) out[4 + warp] = float(n);
  } else if (WMODE == 2 && warp == 2) {
    long long n = 0;
    int i = 0;
    const uint8_t* my = src + (size_t)blockIdx.x * 131072;
    if ((threadIdx.x & 31) == 0) {
      while (!done) {
        const int st = i & 7;
        if (i >= 8) mbar_wait(&cbar[st], ((i >> 3) - 1) & 1);
        mbar_expect_tx(&cbar[st], 16384);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         smem_u32(ring + st * 16384)),
                     "l"(my + st * 16384), "r"(16384), "r"(smem_u32(&cbar[st]))
                     : "memory");
        n += 16384;
        ++i;
        const long long t = clock64();
        while (clock64() - t < gap) {
        }
      }
      for (int k = (i > 8 ? i - 8 : 0); k < i; ++k) mbar_wait(&cbar[k & 7], (k >> 3) & 1);   // drain before exit
      if (blockIdx.x == 0) out[6] = float(n);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tb);
}

template <int WMODE>
void run_w(const char* name, float* d, const uint8_t* src, int gap) {
  const int smem = 16384 + 32768 + 131072 + 1024;
  cudaFuncSetAttribute(kw<WMODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  float h[8] = {0};
  cudaMemset(d, 0, 32);
  kw<WMODE><<<148, 128, smem>>>(d, src, 4000, gap);
  cudaDeviceSynchronize();
  cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
  const float bytes = WMODE == 1 ? h[6] + h[7] : h[6];
  printf("%-40s gap %5d : %7.1f clk/MMA (N=256 floor 128), %6.1f B/clk written to smem   (%s)\n", name, gap, h[0],
         h[2] > 0 ? bytes / h[2] : 0.f, cudaGetErrorString(cudaGetLastError()));
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
  uint8_t* src;
  cudaMalloc(&src, 148 * 131072);
  cudaMemset(src, 1, 148 * 131072);
  run_w<0>("SS N256 b2b, no smem writes", d, src, 0);
  for (int gap : {2000, 400, 200, 100, 50, 0}) run_w<1>("SS N256 b2b + st.shared.v4 (2 warps)", d, src, gap);
  for (int gap : {2000, 400, 200, 100, 50, 0}) run_w<2>("SS N256 b2b + 16 KB bulk copies (L2)", d, src, gap);
  run<64, 0>("SS M128 N64  K16", d);
  run<128, 0>("SS M128 N128 K16", d);
  run<256, 0>("SS M128 N256 K16", d);
  run<64, 1>("TS M128 N64  K16 (B MN-major)", d);
  return 0;
}
