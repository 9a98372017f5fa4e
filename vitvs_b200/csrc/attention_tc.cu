// tcgen05 flash-attention forward for DINOv2 blocks (reference semantics:
// dino_patch/attention.py:70-80 - non-causal SDPA, scale hd^-0.5, head dim 64, no mask, eval).
//
// Input is the fused-QKV GEMM output [n_img * n_tok, 3*D] bf16 exactly as the reference's
// `qkv(x).reshape(B, N, 3, H, hd)` lays it out (q already multiplied by hd^-0.5 in that GEMM's
// epilogue), so no head-major scatter is needed: the Q / K / V tiles of (image, head) are 64-column
// TMA boxes of that matrix.  Output goes straight into the [n_img * n_tok, D] layout attn.proj reads.
//
// One CTA = 128 query rows of one (image, head); 128 threads, thread t owns query row t
// (TMEM lane t).  Per 128-key block:  S = Q K^T (tcgen05, TMEM cols 0..127) -> online softmax in
// registers (two TMEM passes: max, then exp2 / sum / bf16 P to swizzled smem) -> PV = P V
// (tcgen05, V consumed MN-major straight from its row-major tile, TMEM cols 128..191) ->
// O = O * alpha + PV in registers.  Keys past n_tok (next image's rows or TMA zero fill) are masked
// to p = 0.  Two CTAs are resident per SM (80 KB smem, 256 TMEM columns each) so one CTA's softmax
// overlaps the other's MMAs and TMA.
#include <cuda.h>

#include <string>
#include <type_traits>

#include "ptx.cuh"
#include "vvs_internal.h"

namespace vvs {
namespace {

constexpr int kBQ = 128;   // query rows per CTA
constexpr int kBKV = 128;  // keys per block
constexpr int kHD = 64;    // head dim (every DINOv2 size)
constexpr uint32_t kTileBytes = kBQ * kHD * 2;        // 16 KB: Q, K and V tiles
constexpr uint32_t kPBytes = kBQ * kBKV * 2;          // 32 KB: P as two 64-key K-major atoms
constexpr uint32_t kAttnSmem = 3 * kTileBytes + kPBytes + 128 + 1024;
constexpr uint32_t kAttnTmemCols = 256;
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ float ex2_approx(float x) {   // single MUFU.EX2; ex2(-inf) = 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(128, 2)
attention_kernel(const __grid_constant__ CUtensorMap tmQKV, bf16* __restrict__ out, int n_tok, int D) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + kTileBytes;
  uint8_t* sV = smem + 2 * kTileBytes;
  uint8_t* sP = smem + 3 * kTileBytes;
  uint64_t* bar_k = reinterpret_cast<uint64_t*>(smem + 3 * kTileBytes + kPBytes);
  uint64_t* bar_v = bar_k + 1;
  uint64_t* bar_s = bar_k + 2;
  uint64_t* bar_o = bar_k + 3;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_k + 4);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int q0 = blockIdx.x * kBQ;
  const int head = blockIdx.y;
  const int img = blockIdx.z;
  const int row_base = img * n_tok;               // first row of this image in the qkv matrix
  const int nblk = (n_tok + kBKV - 1) / kBKV;
  const int colQ = head * kHD, colK = D + head * kHD, colV = 2 * D + head * kHD;

  if (tid == 0) {
    tma_prefetch_desc(&tmQKV);
    mbar_init(bar_k, 1);
    mbar_init(bar_v, 1);
    mbar_init(bar_s, 1);
    mbar_init(bar_o, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc<kAttnTmemCols>(tmem_slot);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
  const uint32_t tPV = tS + 128;

  constexpr uint32_t idescS = umma_idesc_bf16(128, 128, false);
  constexpr uint32_t idescPV = umma_idesc_bf16(128, 64, true);

  if (tid == 0) {
    mbar_expect_tx(bar_k, 2 * kTileBytes);
    tma_load_3d(sQ, &tmQKV, bar_k, colQ, row_base + q0, 0, kEvictNormal);
    tma_load_3d(sK, &tmQKV, bar_k, colK, row_base, 0, kEvictNormal);
    mbar_expect_tx(bar_v, kTileBytes);
    tma_load_3d(sV, &tmQKV, bar_v, colV, row_base, 0, kEvictNormal);
    mbar_wait(bar_k, 0);
    tc_fence_after();
    const uint32_t qa = smem_u32(sQ), ka = smem_u32(sK);
#pragma unroll
    for (int k = 0; k < kHD / 16; ++k)
      umma_bf16(tmem_base, umma_smem_desc(qa + k * 32, 16, 1024), umma_smem_desc(ka + k * 32, 16, 1024), idescS,
                k != 0);
    umma_commit(bar_s);
  }
  __syncwarp();

  float o[kHD];
#pragma unroll
  for (int d = 0; d < kHD; ++d) o[d] = 0.f;
  float m_run = -INFINITY;   // running max, already multiplied by log2(e)
  float l_run = 0.f;

  const uint32_t p_row_off = static_cast<uint32_t>(tid >> 3) * 1024 + static_cast<uint32_t>(tid & 7) * 128;
  const uint32_t sw = static_cast<uint32_t>(tid & 7);

  for (int j = 0; j < nblk; ++j) {
    const uint32_t par = j & 1;
    mbar_wait(bar_s, par);
    tc_fence_after();
    if (tid == 0 && j + 1 < nblk) {   // S_j is complete => the K tile may be overwritten
      mbar_expect_tx(bar_k, kTileBytes);
      tma_load_3d(sK, &tmQKV, bar_k, colK, row_base + (j + 1) * kBKV, 0, kEvictNormal);
    }
    __syncwarp();
    const int valid = n_tok - j * kBKV;   // keys [0, valid) of this block exist
    float alpha, psum;
    // Softmax of one 128-key block.  kTail = the last, partially filled block: keys >= valid are masked
    // (max ignores them, p = 0).  Full blocks carry no per-element masking at all.  TMEM loads are double
    // buffered; max / sum use 4 independent accumulators to keep the FMNMX / FADD chains short.
    auto softmax_block = [&](auto tail_tag) {
      constexpr bool kTail = decltype(tail_tag)::value;
      uint32_t raw[2][32];
      // ---- pass 1: row max
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
      tmem_ld32(tS, raw[0]);
#pragma unroll
      for (int c = 0; c < kBKV / 32; ++c) {
        tmem_ld_wait();
        if (c + 1 < kBKV / 32) tmem_ld32(tS + (c + 1) * 32, raw[(c + 1) & 1]);
        else tmem_ld32(tS, raw[(c + 1) & 1]);                 // first chunk again for pass 2
        if (kTail && c * 32 >= valid) continue;
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          float s0 = __uint_as_float(raw[c & 1][i]), s1 = __uint_as_float(raw[c & 1][i + 1]);
          float s2 = __uint_as_float(raw[c & 1][i + 2]), s3 = __uint_as_float(raw[c & 1][i + 3]);
          if (kTail) {
            if (c * 32 + i + 0 >= valid) s0 = -INFINITY;
            if (c * 32 + i + 1 >= valid) s1 = -INFINITY;
            if (c * 32 + i + 2 >= valid) s2 = -INFINITY;
            if (c * 32 + i + 3 >= valid) s3 = -INFINITY;
          }
          mx0 = fmaxf(mx0, s0);
          mx1 = fmaxf(mx1, s1);
          mx2 = fmaxf(mx2, s2);
          mx3 = fmaxf(mx3, s3);
        }
      }
      const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
      const float m_new = fmaxf(m_run, mx * kLog2e);
      alpha = ex2_approx(m_run - m_new);
      m_run = m_new;
      const float neg_m = -m_new;
      // ---- pass 2: p = exp2(s*log2e - m), row sum, bf16 P -> smem (K-major, SWIZZLE_128B, two 64-key atoms)
      float ps0 = 0.f, ps1 = 0.f, ps2 = 0.f, ps3 = 0.f;
#pragma unroll
      for (int c = 0; c < kBKV / 32; ++c) {
        tmem_ld_wait();
        if (c + 1 < kBKV / 32) tmem_ld32(tS + (c + 1) * 32, raw[(c + 1) & 1]);   // chunk c sits in raw[c & 1]
        const uint32_t(&rw)[32] = raw[c & 1];
        float pv[32];
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          float e0 = ex2_approx(fmaf(__uint_as_float(rw[i + 0]), kLog2e, neg_m));
          float e1 = ex2_approx(fmaf(__uint_as_float(rw[i + 1]), kLog2e, neg_m));
          float e2 = ex2_approx(fmaf(__uint_as_float(rw[i + 2]), kLog2e, neg_m));
          float e3 = ex2_approx(fmaf(__uint_as_float(rw[i + 3]), kLog2e, neg_m));
          if (kTail) {
            if (c * 32 + i + 0 >= valid) e0 = 0.f;
            if (c * 32 + i + 1 >= valid) e1 = 0.f;
            if (c * 32 + i + 2 >= valid) e2 = 0.f;
            if (c * 32 + i + 3 >= valid) e3 = 0.f;
          }
          pv[i + 0] = e0; pv[i + 1] = e1; pv[i + 2] = e2; pv[i + 3] = e3;
          ps0 += e0; ps1 += e1; ps2 += e2; ps3 += e3;
        }
        uint8_t* atom = sP + (c >> 1) * (kBQ * 128) + p_row_off;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 w;
          w.x = pack_bf16x2(pv[q * 8 + 0], pv[q * 8 + 1]);
          w.y = pack_bf16x2(pv[q * 8 + 2], pv[q * 8 + 3]);
          w.z = pack_bf16x2(pv[q * 8 + 4], pv[q * 8 + 5]);
          w.w = pack_bf16x2(pv[q * 8 + 6], pv[q * 8 + 7]);
          const uint32_t chunk = static_cast<uint32_t>((c & 1) * 4 + q);
          *reinterpret_cast<uint4*>(atom + ((chunk ^ sw) << 4)) = w;
        }
      }
      psum = (ps0 + ps1) + (ps2 + ps3);
    };
    if (valid >= kBKV) softmax_block(std::false_type{});
    else softmax_block(std::true_type{});
    l_run = l_run * alpha + psum;

    fence_proxy_async_smem();   // P (generic proxy) -> visible to the tensor core (async proxy)
    tc_fence_before();
    __syncthreads();

    if (tid == 0) {
      tc_fence_after();
      mbar_wait(bar_v, par);
      const uint32_t pa = smem_u32(sP), va = smem_u32(sV);
#pragma unroll
      for (int kk = 0; kk < kBKV / 16; ++kk) {
        const uint64_t da = umma_smem_desc(pa + (kk >> 2) * (kBQ * 128) + (kk & 3) * 32, 16, 1024);
        const uint64_t db = umma_smem_desc(va + kk * 2048, 1024, 1024);   // MN-major: 16 keys = 2 x 1024 B groups
        umma_bf16(tmem_base + 128, da, db, idescPV, kk != 0);
      }
      umma_commit(bar_o);
      if (j + 1 < nblk) {
        mbar_wait(bar_k, (j + 1) & 1);
        tc_fence_after();
        const uint32_t qa = smem_u32(sQ), ka = smem_u32(sK);
#pragma unroll
        for (int k = 0; k < kHD / 16; ++k)
          umma_bf16(tmem_base, umma_smem_desc(qa + k * 32, 16, 1024), umma_smem_desc(ka + k * 32, 16, 1024), idescS,
                    k != 0);
        umma_commit(bar_s);
      }
    }
    __syncwarp();

    mbar_wait(bar_o, par);
    tc_fence_after();
    if (tid == 0 && j + 1 < nblk) {   // PV_j is complete => the V tile may be overwritten
      mbar_expect_tx(bar_v, kTileBytes);
      tma_load_3d(sV, &tmQKV, bar_v, colV, row_base + (j + 1) * kBKV, 0, kEvictNormal);
    }
    __syncwarp();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint32_t raw[32];
      tmem_ld32(tPV + h * 32, raw);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) o[h * 32 + i] = fmaf(o[h * 32 + i], alpha, __uint_as_float(raw[i]));
    }
  }

  const int q = q0 + tid;
  if (q < n_tok) {
    const float inv = 1.0f / l_run;
    uint4* dst = reinterpret_cast<uint4*>(out + static_cast<size_t>(row_base + q) * D + head * kHD);
#pragma unroll
    for (int g = 0; g < 8; ++g) {
      uint4 w;
      w.x = pack_bf16x2(o[g * 8 + 0] * inv, o[g * 8 + 1] * inv);
      w.y = pack_bf16x2(o[g * 8 + 2] * inv, o[g * 8 + 3] * inv);
      w.z = pack_bf16x2(o[g * 8 + 4] * inv, o[g * 8 + 5] * inv);
      w.w = pack_bf16x2(o[g * 8 + 6] * inv, o[g * 8 + 7] * inv);
      dst[g] = w;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<kAttnTmemCols>(tmem_base);
}

}  // namespace

int launch_attention(const bf16* qkv, bf16* out, int n_img, int n_tok, int heads, cudaStream_t stream,
                     std::string* err) {
  const int D = heads * kHD;
  CUtensorMap tm;
  int rc = make_tmap_bf16(&tm, qkv, 3ull * D, static_cast<uint64_t>(n_img) * n_tok, 1, 3ull * D, 0, kBQ, err);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    VVS_CUDA_OK(cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(kAttnSmem)),
                err);
    attr_set = true;
  }
  dim3 grid((n_tok + kBQ - 1) / kBQ, heads, n_img);
  attention_kernel<<<grid, 128, kAttnSmem, stream>>>(tm, out, n_tok, D);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

}  // namespace vvs
