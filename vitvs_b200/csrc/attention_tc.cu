// tcgen05 flash-attention forward for DINOv2 blocks (reference semantics:
// dino_patch/attention.py:70-80 - non-causal SDPA, scale hd^-0.5, head dim 64, no mask, eval).
//
// Input is the fused-QKV GEMM output [n_img * n_tok, 3*D] bf16 exactly as the reference's
// `qkv(x).reshape(B, N, 3, H, hd)` lays it out (q already multiplied by hd^-0.5 in that GEMM's
// epilogue), so no head-major scatter is needed: the Q / K / V tiles of (image, head) are 64-column
// TMA boxes of that matrix.  Output goes straight into the [n_img * n_tok, D] layout attn.proj reads.
//
// One CTA = one 128-row query tile of one (image, head); keys are consumed in blocks of 64.
// The kernel is latency-structured rather than FLOP-structured (measured: removing all softmax arithmetic
// from a lock-step version changed its time by 15 %): the MMA warp runs one key block AHEAD of the softmax
// warpgroup - S is double buffered in TMEM and P double buffered in smem - so in steady state the softmax
// warps never wait for a tensor-core result, and two CTAs per SM (96 KB smem, 256 TMEM columns each)
// interleave their MUFU / TMEM / MMA phases.
//   S = Q K^T (tcgen05, fp32 in TMEM) -> the thread owning a query row reads its 64 scores ONCE into
//   registers -> max / exp2 / sum / bf16 P -> swizzled smem -> O += P V (tcgen05, V consumed MN-major
//   straight from its row-major tile, O accumulated in TMEM across blocks, lazily rescaled only when the
//   running max grows by more than 2^8) -> O / l once at the end.
// Keys past n_tok (the next image's rows or TMA zero fill) are masked to p = 0.
#include <cuda.h>

#include <string>

#include "ptx.cuh"
#include "vvs_internal.h"

namespace vvs {
namespace {

constexpr int kBQ = 128;   // query rows per CTA (one TMEM lane per row)
constexpr int kBKV = 64;   // keys per block
constexpr int kHD = 64;    // head dim (every DINOv2 size)
constexpr int kAttnThreads = 256;
constexpr int kKVStages = 3;
constexpr uint32_t kQBytes = kBQ * kHD * 2;           // 16 KB
constexpr uint32_t kKVBytes = kBKV * kHD * 2;         // 8 KB: one K or V block
constexpr uint32_t kPBytes = kBQ * kBKV * 2;          // 16 KB: P block, one 64-key K-major atom column
constexpr uint32_t kOffQ = 0, kOffK = kQBytes, kOffV = kOffK + kKVStages * kKVBytes,
                   kOffP = kOffV + kKVStages * kKVBytes, kOffBar = kOffP + 2 * kPBytes;
constexpr uint32_t kAttnSmem = kOffBar + 256 + 1024;
constexpr uint32_t kAttnTmemCols = 256;               // S0 0..63 | S1 64..127 | O 128..191
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kRescaleLog2 = 8.0f;

__device__ __forceinline__ float ex2_approx(float x) {   // single MUFU.EX2; ex2(-inf) = 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Warp roles (256 threads = 2 warpgroups so that setmaxnreg can move registers between them):
//   warps 0-3  softmax warpgroup (thread = query row = TMEM lane)
//   warp  4    MMA issuer (one lane)      warp 5  TMA producer (one lane)      warps 6-7 idle padding
__global__ void __launch_bounds__(kAttnThreads, 2)
attention_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                 bf16* __restrict__ out, int n_tok, int D) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* q_full = bars;                       // 1
  uint64_t* k_full = bars + 1;                   // kKVStages
  uint64_t* v_full = k_full + kKVStages;
  uint64_t* k_empty = v_full + kKVStages;
  uint64_t* v_empty = k_empty + kKVStages;
  uint64_t* s_full = v_empty + kKVStages;        // 2 (S buffers)
  uint64_t* p_full = s_full + 2;                 // 2 (P buffers), 128 arrivals
  uint64_t* pv_done = p_full + 2;                // 2 (P buffers)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pv_done + 2);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int q0 = blockIdx.x * kBQ;
  const int head = blockIdx.y;
  const int img = blockIdx.z;
  const int row_base = img * n_tok;               // first row of this image in the qkv matrix
  const int nblk = (n_tok + kBKV - 1) / kBKV;
  const int colQ = head * kHD, colK = D + head * kHD, colV = 2 * D + head * kHD;

  if (tid == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
    mbar_init(q_full, 1);
    for (int i = 0; i < kKVStages; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], kBQ);
      mbar_init(&pv_done[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc<kAttnTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    if (warp == 5 && lane == 0) {
      // ---------------------------------------------------------------- TMA producer
      mbar_expect_tx(q_full, kQBytes);
      tma_load_3d(smem + kOffQ, &tmQ, q_full, colQ, row_base + q0, 0, kEvictNormal);
      for (int j = 0; j < nblk; ++j) {
        const int st = j % kKVStages;
        const uint32_t ph = (j / kKVStages) & 1;
        mbar_wait(&k_empty[st], ph ^ 1);
        mbar_expect_tx(&k_full[st], kKVBytes);
        tma_load_3d(smem + kOffK + st * kKVBytes, &tmKV, &k_full[st], colK, row_base + j * kBKV, 0, kEvictLast);
        mbar_wait(&v_empty[st], ph ^ 1);
        mbar_expect_tx(&v_full[st], kKVBytes);
        tma_load_3d(smem + kOffV + st * kKVBytes, &tmKV, &v_full[st], colV, row_base + j * kBKV, 0, kEvictLast);
      }
    } else if (warp == 4 && lane == 0) {
      // ---------------------------------------------------------------- MMA issuer (one block ahead on S)
      constexpr uint32_t idescS = umma_idesc_bf16(128, kBKV, false);
      constexpr uint32_t idescPV = umma_idesc_bf16(128, kHD, true);
      const uint32_t sbase = smem_u32(smem);
      auto issue_S = [&](int j) {        // S(j) -> TMEM buffer j & 1
        const int st = j % kKVStages;
        mbar_wait(&k_full[st], (j / kKVStages) & 1);
        tc_fence_after();
        const uint32_t qa = sbase + kOffQ, ka = sbase + kOffK + st * kKVBytes;
#pragma unroll
        for (int k = 0; k < kHD / 16; ++k)
          umma_bf16(tmem_base + (j & 1) * kBKV, umma_smem_desc(qa + k * 32, 16, 1024),
                    umma_smem_desc(ka + k * 32, 16, 1024), idescS, k != 0);
        umma_commit(&s_full[j & 1]);
        umma_commit(&k_empty[st]);
      };
      mbar_wait(q_full, 0);
      issue_S(0);
      for (int j = 0; j < nblk; ++j) {
        if (j + 1 < nblk) {
          // S buffer (j+1)&1 last held S(j-1): free once P(j-1) has been published
          if (j >= 1) {
            mbar_wait(&p_full[(j - 1) & 1], ((j - 1) >> 1) & 1);
            tc_fence_after();
          }
          issue_S(j + 1);
        }
        const int st = j % kKVStages;
        mbar_wait(&p_full[j & 1], (j >> 1) & 1);       // P(j) in smem, O rescaled if it had to be
        mbar_wait(&v_full[st], (j / kKVStages) & 1);
        tc_fence_after();
        const uint32_t pa = sbase + kOffP + (j & 1) * kPBytes, va = sbase + kOffV + st * kKVBytes;
#pragma unroll
        for (int kk = 0; kk < kBKV / 16; ++kk)
          umma_bf16(tmem_base + 128, umma_smem_desc(pa + kk * 32, 16, 1024),
                    umma_smem_desc(va + kk * 2048, 1024, 1024),   // MN-major V: 16 keys = 2 x 1024 B groups
                    idescPV, (j > 0 || kk != 0) ? 1u : 0u);        // O accumulates in TMEM across key blocks
        umma_commit(&pv_done[j & 1]);
        umma_commit(&v_empty[st]);
      }
    }
    __syncwarp();
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 200;");
    // ------------------------------------------------------------------ softmax warpgroup
    const int row = warp * 32 + lane;                 // row inside the tile = TMEM lane
    const uint32_t lane_field = static_cast<uint32_t>(warp * 32) << 16;
    const uint32_t tO = tmem_base + 128 + lane_field;
    const uint32_t p_row_off = static_cast<uint32_t>(row >> 3) * 1024 + static_cast<uint32_t>(row & 7) * 128;
    const uint32_t sw = static_cast<uint32_t>(row & 7);
    float m_run = -INFINITY;   // reference max of the exponentials, already multiplied by log2(e)
    float l_run = 0.f;

    for (int j = 0; j < nblk; ++j) {
      const int sb = j & 1;
      const uint32_t ph2 = (j >> 1) & 1;
      mbar_wait(&s_full[sb], ph2);
      tc_fence_after();
      const int valid = n_tok - j * kBKV;   // keys [0, valid) of this block exist
      uint32_t sc[2][32];                   // the row of S (fp32 bits)
      tmem_ld32(tmem_base + sb * kBKV + lane_field, sc[0]);
      tmem_ld32(tmem_base + sb * kBKV + 32 + lane_field, sc[1]);
      tmem_ld_wait();
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
      if (valid < kBKV) {
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c * 32 + i >= valid) sc[c][i] = 0xff800000u;      // -inf: ignored by the max, exp2 -> 0
      }
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          mx0 = fmaxf(mx0, __uint_as_float(sc[c][i]));
          mx1 = fmaxf(mx1, __uint_as_float(sc[c][i + 1]));
          mx2 = fmaxf(mx2, __uint_as_float(sc[c][i + 2]));
          mx3 = fmaxf(mx3, __uint_as_float(sc[c][i + 3]));
        }
      const float m_blk = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * kLog2e;
      // Lazy rescale: the reference max only moves when a block exceeds it by more than 2^8; until then
      // p = exp2(s - m_run) may exceed 1 (<= 256: exact in the fp32 sum, representable in bf16) and O / l
      // is unchanged mathematically.  Always true for j == 0 (m_run = -inf).
      const bool raise = m_blk > m_run + kRescaleLog2;
      const float alpha = raise ? ex2_approx(m_run - m_blk) : 1.0f;
      if (raise) m_run = m_blk;
      const float neg_m = -m_run;
      float ps0 = 0.f, ps1 = 0.f, ps2 = 0.f, ps3 = 0.f;
      uint32_t pk[2][16];                   // bf16x2-packed P row
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float e0 = ex2_approx(fmaf(__uint_as_float(sc[c][i + 0]), kLog2e, neg_m));
          const float e1 = ex2_approx(fmaf(__uint_as_float(sc[c][i + 1]), kLog2e, neg_m));
          const float e2 = ex2_approx(fmaf(__uint_as_float(sc[c][i + 2]), kLog2e, neg_m));
          const float e3 = ex2_approx(fmaf(__uint_as_float(sc[c][i + 3]), kLog2e, neg_m));
          ps0 += e0; ps1 += e1; ps2 += e2; ps3 += e3;
          pk[c][i / 2] = pack_bf16x2(e0, e1);
          pk[c][i / 2 + 1] = pack_bf16x2(e2, e3);
        }
      l_run = l_run * alpha + ((ps0 + ps1) + (ps2 + ps3));

      if (j >= 2) mbar_wait(&pv_done[sb], ((j - 2) >> 1) & 1);    // PV(j-2) has finished reading P buffer sb
      if (j >= 1 && __any_sync(0xffffffffu, raise)) {
        mbar_wait(&pv_done[(j - 1) & 1], ((j - 1) >> 1) & 1);     // O must be quiescent while it is rescaled
        tc_fence_after();
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t ov[32];
          tmem_ld32(tO + h * 32, ov);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) ov[i] = __float_as_uint(__uint_as_float(ov[i]) * alpha);
          tmem_st32(tO + h * 32, ov);
        }
        tmem_st_wait();
      }
      uint8_t* prow = smem + kOffP + sb * kPBytes + p_row_off;
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int q = 0; q < 4; ++q)
          *reinterpret_cast<uint4*>(prow + ((static_cast<uint32_t>(c * 4 + q) ^ sw) << 4)) =
              make_uint4(pk[c][q * 4 + 0], pk[c][q * 4 + 1], pk[c][q * 4 + 2], pk[c][q * 4 + 3]);
      fence_proxy_async_smem();   // P (generic proxy) -> visible to the tensor core (async proxy)
      tc_fence_before();
      mbar_arrive(&p_full[sb]);
    }

    mbar_wait(&pv_done[(nblk - 1) & 1], ((nblk - 1) >> 1) & 1);
    tc_fence_after();
    const int q = q0 + row;
    const float inv = 1.0f / l_run;
    uint4* dst = reinterpret_cast<uint4*>(out + static_cast<size_t>(row_base + q) * D + head * kHD);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint32_t ov[32];
      tmem_ld32(tO + h * 32, ov);
      tmem_ld_wait();
      if (q < n_tok) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 w;
          w.x = pack_bf16x2(__uint_as_float(ov[g * 8 + 0]) * inv, __uint_as_float(ov[g * 8 + 1]) * inv);
          w.y = pack_bf16x2(__uint_as_float(ov[g * 8 + 2]) * inv, __uint_as_float(ov[g * 8 + 3]) * inv);
          w.z = pack_bf16x2(__uint_as_float(ov[g * 8 + 4]) * inv, __uint_as_float(ov[g * 8 + 5]) * inv);
          w.w = pack_bf16x2(__uint_as_float(ov[g * 8 + 6]) * inv, __uint_as_float(ov[g * 8 + 7]) * inv);
          dst[h * 4 + g] = w;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc<kAttnTmemCols>(tmem_base);
}

}  // namespace

int launch_attention(const bf16* qkv, bf16* out, int n_img, int n_tok, int heads, cudaStream_t stream,
                     std::string* err) {
  const int D = heads * kHD;
  CUtensorMap tmq, tmkv;
  const uint64_t rows = static_cast<uint64_t>(n_img) * n_tok;
  int rc = make_tmap_bf16(&tmq, qkv, 3ull * D, rows, 1, 3ull * D, 0, kBQ, err);
  if (rc) return rc;
  rc = make_tmap_bf16(&tmkv, qkv, 3ull * D, rows, 1, 3ull * D, 0, kBKV, err);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    VVS_CUDA_OK(cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(kAttnSmem)),
                err);
    attr_set = true;
  }
  dim3 grid((n_tok + kBQ - 1) / kBQ, heads, n_img);
  attention_kernel<<<grid, kAttnThreads, kAttnSmem, stream>>>(tmq, tmkv, out, n_tok, D);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

}  // namespace vvs
