// Microbenchmark (GPU box): how many SM sub-partition cycles the softmax inner math of the attention kernel costs
// in isolation - MUFU.EX2 alone, the packed exp body (FFMA2 + 2 MUFU + FADD2 + F2FP per 2 elements), and with the
// row max - at 1, 2 and 4 warps per sub-partition.   nvcc -arch=sm_100a -O3 -o softmax_math softmax_math.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../vitvs_b200/csrc/ptx.cuh"
using namespace vvs;

template <int MODE>
__global__ void __launch_bounds__(512, 1) k(float* out, int iters, float seed) {
  __shared__ uint32_t tmem_slot;
  uint32_t tbase = 0;
  if (MODE >= 4) {
    if (threadIdx.x < 32) tmem_alloc<512>(&tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    // warp w may touch lanes 32 (w % 4) ..; give every warp its own 32-column slice
    tbase = tmem_slot + ((uint32_t)((threadIdx.x >> 5) & 3) * 32 << 16) + (threadIdx.x >> 7) * 128;
  }
  float s[64];
#pragma unroll
  for (int i = 0; i < 64; ++i) s[i] = seed * (i + 1) + threadIdx.x * 1e-3f;
  float acc = 0.f;
  uint32_t xacc = 0;
  const f32x2 l2 = f2_pack(1.4426950408889634f, 1.4426950408889634f);
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {   // MUFU only
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        float y;
        asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(s[i]));
        s[i] = y * 0.5f - 1.0f;   // keep a dependency per slot (1 FFMA)
      }
    } else {
      float m = -1e30f;
      if (MODE >= 2) {
        float m0 = m, m1 = m, m2 = m, m3 = m;
#pragma unroll
        for (int i = 0; i < 64; i += 4) {
          m0 = fmaxf(m0, s[i]); m1 = fmaxf(m1, s[i + 1]); m2 = fmaxf(m2, s[i + 2]); m3 = fmaxf(m3, s[i + 3]);
        }
        m = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)) * 1.4426950408889634f;
      } else m = seed;
      const f32x2 nm = f2_pack(-m, -m);
      f32x2 a01 = f2_pack(0.f, 0.f), a23 = a01;
      uint32_t pk[32];
#pragma unroll
      for (int i = 0; i < 64; i += 4) {
        const f32x2 e01 = f2_ex2(f2_fma(f2_pack(s[i], s[i + 1]), l2, nm));
        const f32x2 e23 = f2_ex2(f2_fma(f2_pack(s[i + 2], s[i + 3]), l2, nm));
        a01 = f2_add(a01, e01);
        a23 = f2_add(a23, e23);
        pk[i / 2] = f2_to_bf16x2(e01);
        pk[i / 2 + 1] = f2_to_bf16x2(e23);
      }
      if (MODE >= 4) {   // + the P hand-off of generation 7/8: 4 x tcgen05.st x8 (+ wait::st)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint32_t p8[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) p8[i] = pk[q * 8 + i];
          tmem_st8(tbase + q * 8, p8);
        }
        tmem_st_wait();
        tc_fence_before();
        if (MODE == 5) {   // + S from TMEM: 2 x tcgen05.ld x32
          uint32_t a[32], b[32];
          tmem_ld32(tbase + 32, a);
          tmem_ld32(tbase + 64, b);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) { s[i] = __uint_as_float(a[i] & 0x3fffffff); s[32 + i] = __uint_as_float(b[i] & 0x3fffffff); }
        }
      } else if (MODE == 3) {   // + the P hand-off of generation 6: 8 x STS.128 into this thread's row
        extern __shared__ uint4 sm[];
#pragma unroll
        for (int q = 0; q < 8; ++q) sm[threadIdx.x * 8 + ((q ^ (threadIdx.x & 7)))] = make_uint4(pk[q * 4], pk[q * 4 + 1], pk[q * 4 + 2], pk[q * 4 + 3]);
      } else {
#pragma unroll
        for (int i = 0; i < 32; i += 2) xacc ^= pk[i] ^ pk[i + 1];
      }
      float x0, x1;
      f2_unpack(f2_add(a01, a23), x0, x1);
      acc += x0 + x1;
#pragma unroll
      for (int i = 0; i < 64; ++i) s[i] += acc * 1e-9f;   // new inputs every iteration (64 FFMA, counted below)
    }
  }
  long long t1 = clock64();
  float r = acc + __uint_as_float(xacc & 0x3fffff);
#pragma unroll
  for (int i = 0; i < 64; ++i) r += s[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0 && blockIdx.x == 0) out[1 << 20] = float(t1 - t0);
  if (MODE >= 4) {
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc<512>(tmem_slot);
  }
}

template <int MODE>
void run(const char* name, float* d) {
  for (int warps : {4, 8, 16}) {
    const int iters = 2000;
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    k<MODE><<<148, warps * 32, 65536>>>(d, 10, 0.01f);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k<MODE><<<148, warps * 32, 65536>>>(d, iters, 0.01f);
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms, clk;
    cudaEventElapsedTime(&ms, a, b);
    cudaMemcpy(&clk, d + (1 << 20), 4, cudaMemcpyDeviceToHost);
    printf("%-28s warps/SMSP %d : %8.1f clk per 64-element row block per warp, %8.1f clk per SMSP round (%.3f ms, %s)\n", name,
           warps / 4, clk / iters, clk / iters, ms, cudaGetErrorString(cudaGetLastError()));
  }
}

int main() {
  float* d;
  cudaMalloc(&d, ((1 << 20) + 16) * 4);
  run<0>("64 MUFU.EX2 (+64 FFMA)", d);
  run<1>("exp body (+64 FFMA feed)", d);
  run<2>("max + exp body (+64 FFMA)", d);
  run<3>("max + exp body + 8 STS.128", d);
  run<4>("max + exp body + 4 STTM.x8", d);
  run<5>("  ... + 2 LDTM.x32", d);
  return 0;
}
