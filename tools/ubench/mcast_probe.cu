// Probe (GPU box): where does the mbarrier complete_tx of a MULTICAST tensor load land when the load is issued
// with .cta_group::2 and the barrier operand names the pair LEADER's barrier?  Needed for a GEMM whose clusters are
// two CTA pairs sharing one operand tile: the MMA issuer of a pair waits on its leader's
// barrier only, so the bytes multicast to the non-leader CTAs must be countable there.
//
// Cluster of 4 CTAs = pairs (0,1) and (2,3).  Every CTA arms bar[0] with expect_tx(box bytes).  One CTA issues ONE
// load of a 64 x 128 bf16 box, multicast to the CTAs in `mask`, with the barrier operand pointing at the barrier of
// CTA `bar_rank`.  After a fixed delay every CTA reports whether its barrier completed and whether the data landed.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mcast_probe mcast_probe.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <vector>

#include "../../vitvs_b200/csrc/ptx.cuh"
using namespace vvs;

constexpr int kBoxBytes = 64 * 128 * 2;

template <int CTA_GROUP>
__global__ void __cluster_dims__(4, 1, 1) __launch_bounds__(32, 1)
probe(const __grid_constant__ CUtensorMap tm, int issuer, int mask, int bar_rank, int* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  const uint32_t rank = cluster_ctarank();
  if (threadIdx.x == 0) {
    for (int i = 0; i < kBoxBytes / 4; ++i) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    mbar_init(&bar, 1);
    fence_mbar_init();
    mbar_expect_tx(&bar, kBoxBytes);
  }
  cluster_sync_all();
  if (threadIdx.x == 0 && static_cast<int>(rank) == issuer) {
    const uint32_t bar_addr = mapa_shared(smem_u32(&bar), static_cast<uint32_t>(bar_rank));
    if (CTA_GROUP == 2)
      asm volatile(
          "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
          " [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(smem_u32(smem)),
          "l"(reinterpret_cast<uint64_t>(&tm)), "r"(bar_addr), "r"(0), "r"(0), "r"(0), "h"(static_cast<uint16_t>(mask))
          : "memory");
    else
      asm volatile(
          "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
          " [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(smem_u32(smem)),
          "l"(reinterpret_cast<uint64_t>(&tm)), "r"(bar_addr), "r"(0), "r"(0), "r"(0), "h"(static_cast<uint16_t>(mask))
          : "memory");
  }
  const long long t0 = clock64();
  while (clock64() - t0 < 200000) {
  }
  if (threadIdx.x == 0) {
    out[rank * 2 + 0] = mbar_test_wait(&bar, 0) ? 1 : 0;
    out[rank * 2 + 1] = reinterpret_cast<volatile uint32_t*>(smem)[5] != 0 ? 1 : 0;
  }
  cluster_sync_all();
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  void* d;
  cudaMalloc(&d, 1 << 20);
  std::vector<uint16_t> h((1 << 20) / 2, 0x3f80);
  cudaMemcpy(d, h.data(), 1 << 20, cudaMemcpyHostToDevice);
  int* out;
  cudaMalloc(&out, 64);
  EncodeFn enc = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", reinterpret_cast<void**>(&enc), cudaEnableDefault, &q);
  CUtensorMap tm;
  cuuint64_t dims[3] = {64, 1024, 1};
  cuuint64_t strides[2] = {128, 128 * 1024};
  cuuint32_t box[3] = {64, 128, 1};
  cuuint32_t es[3] = {1, 1, 1};
  CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    printf("encode failed %d\n", (int)r);
    return 1;
  }
  const int smem = kBoxBytes + 1024;
  cudaFuncSetAttribute(probe<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(probe<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  struct Case { int grp, issuer, mask, bar_rank; const char* what; };
  const Case cases[] = {
      {1, 0, 0x5, 0, "plain multicast from CTA 0 to {0,2}, barrier operand = own"},
      {1, 1, 0xA, 1, "plain multicast from CTA 1 to {1,3}, barrier operand = own"},
      {2, 0, 0x5, 0, "cta_group::2 multicast from CTA 0 to {0,2}, barrier operand = CTA 0 (own, leader)"},
      {2, 1, 0xA, 0, "cta_group::2 multicast from CTA 1 to {1,3}, barrier operand = CTA 0 (its pair leader)"},
      {2, 1, 0xA, 1, "cta_group::2 multicast from CTA 1 to {1,3}, barrier operand = CTA 1 (own)"},
      {2, 3, 0xA, 2, "cta_group::2 multicast from CTA 3 to {1,3}, barrier operand = CTA 2 (its pair leader)"},
      {2, 1, 0x2, 0, "cta_group::2, no fan-out (mask {1}), barrier operand = CTA 0 (what the GEMM does today)"},
  };
  for (const Case& c : cases) {
    cudaMemset(out, 0xff, 64);
    if (c.grp == 2) probe<2><<<4, 32, smem>>>(tm, c.issuer, c.mask, c.bar_rank, out);
    else probe<1><<<4, 32, smem>>>(tm, c.issuer, c.mask, c.bar_rank, out);
    cudaError_t e = cudaDeviceSynchronize();
    int h8[8];
    cudaMemcpy(h8, out, 32, cudaMemcpyDeviceToHost);
    printf("%-95s | barrier done c0..c3: %d %d %d %d | data landed: %d %d %d %d  (%s)\n", c.what, h8[0], h8[2], h8[4],
           h8[6], h8[1], h8[3], h8[5], h8[7], cudaGetErrorString(e));
    if (e != cudaSuccess) break;
  }
  return 0;
}
