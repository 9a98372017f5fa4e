// Microbenchmark (GPU box): can two CTAs that each tcgen05.alloc part of TMEM be resident on one SM?
// Reports the occupancy the driver computes and the overlap actually observed (%smid + globaltimer).
#include <cstdio>
#include <cstdint>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>
#include "../../vitvs_b200/csrc/ptx.cuh"
using namespace vvs;

template <bool TMEM>
__global__ void __launch_bounds__(128, 2) k(unsigned long long* rec, long long spin) {
  __shared__ uint32_t slot;
  if (TMEM) {
    if (threadIdx.x < 32) tmem_alloc<128>(&slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
  }
  unsigned long long t0, t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  const long long c0 = clock64();
  while (clock64() - c0 < spin) {
  }
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
  uint32_t smid;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  if (threadIdx.x == 0) {
    rec[blockIdx.x * 3 + 0] = smid;
    rec[blockIdx.x * 3 + 1] = t0;
    rec[blockIdx.x * 3 + 2] = t1;
  }
  if (TMEM) {
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc<128>(slot);
  }
}

template <bool TMEM>
void run(const char* name) {
  const int grid = 296;
  unsigned long long* d;
  cudaMalloc(&d, grid * 3 * 8);
  int nb = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k<TMEM>, 128, 0);
  k<TMEM><<<grid, 128>>>(d, 200000);
  cudaError_t e = cudaDeviceSynchronize();
  std::vector<unsigned long long> h(grid * 3);
  cudaMemcpy(h.data(), d, grid * 3 * 8, cudaMemcpyDeviceToHost);
  int overlaps = 0;
  for (int a = 0; a < grid; ++a)
    for (int b = a + 1; b < grid; ++b)
      if (h[a * 3] == h[b * 3] && h[a * 3 + 1] < h[b * 3 + 2] && h[b * 3 + 1] < h[a * 3 + 2]) ++overlaps;
  printf("%-10s occupancy API %d CTAs/SM; %d overlapping CTA pairs on the same SM out of %d CTAs (%s)\n", name, nb, overlaps,
         grid, cudaGetErrorString(e));
}

int main() {
  run<false>("no TMEM");
  run<true>("TMEM 128");
  return 0;
}
