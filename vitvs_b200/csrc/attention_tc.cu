// tcgen05 flash-attention forward for DINOv2 blocks (reference semantics:
// dino_patch/attention.py:70-80 - non-causal SDPA, scale hd^-0.5, head dim 64, no mask, eval).
//
// Input is the fused-QKV GEMM output [n_img * n_tok, 3*D] bf16 exactly as the reference's
// `qkv(x).reshape(B, N, 3, H, hd)` lays it out (q already multiplied by hd^-0.5 in that GEMM's
// epilogue), so no head-major scatter is needed: the Q / K / V tiles of (image, head) are 64-column
// TMA boxes of that matrix.  Output goes straight into the [n_img * n_tok, D] layout attn.proj reads.
//
// One CTA = 256 query rows (two 128-row tiles) of one (image, head), warp specialised (see the kernel).
// Per 128-key block and tile:  S = Q K^T (tcgen05, fp32 in TMEM) -> online softmax in registers (thread =
// row = TMEM lane; ONE TMEM pass: the row of S is held in registers for max, exp2, sum and the bf16 P that
// goes to swizzled smem) -> O += P V (tcgen05, V consumed MN-major straight from its row-major tile, O
// accumulated in TMEM across blocks with lazy rescaling) -> O / l once at the end.
// Keys past n_tok (the next image's rows or TMA zero fill) are masked to p = 0.
#include <cuda.h>

#include <string>
#include <type_traits>

#include "ptx.cuh"
#include "vvs_internal.h"

namespace vvs {
namespace {

constexpr int kBQ = 128;   // query rows per softmax warpgroup (one TMEM lane per row)
constexpr int kBKV = 128;  // keys per block
constexpr int kHD = 64;    // head dim (every DINOv2 size)
constexpr int kAttnThreads = 384;
constexpr uint32_t kTileBytes = kBQ * kHD * 2;        // 16 KB: one Q, K or V tile
constexpr uint32_t kPBytes = kBQ * kBKV * 2;          // 32 KB: P as two 64-key K-major atoms
// smem: Q[2] | K[2 stages] | V[2 stages] | P[2] | barriers
constexpr uint32_t kOffQ = 0, kOffK = 2 * kTileBytes, kOffV = 4 * kTileBytes, kOffP = 6 * kTileBytes;
constexpr uint32_t kOffBar = kOffP + 2 * kPBytes;
constexpr uint32_t kAttnSmem = kOffBar + 256 + 1024;
constexpr uint32_t kAttnTmemCols = 512;               // S_A 0..127 | S_B 128..255 | PV_A 256..319 | PV_B 320..383
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ float ex2_approx(float x) {   // single MUFU.EX2; ex2(-inf) = 0
#ifdef VVS_EXPERIMENT_NO_MUFU    // timing experiment only: wrong numerics
  return fmaf(x, 0.001f, 0.5f);
#else
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
#endif
}

// Warp roles (384 threads = 3 warpgroups so that setmaxnreg can move registers between them):
//   warps 0-3  softmax warpgroup A : query tile q0      .. q0+127   (thread = row = TMEM lane)
//   warps 4-7  softmax warpgroup B : query tile q0+128  .. q0+255
//   warp  8    MMA issuer (one lane): S = Q K^T and PV = P V for both tiles (tcgen05, accumulators in TMEM)
//   warp  9    TMA producer (one lane): Q tiles once, K / V blocks through a 2-stage ring
//   warps 10-11 idle (pad the producer warpgroup)
// The two softmax warpgroups ping-pong: while A runs exp2 on S_A(j) the tensor core computes S_B(j) /
// PV_B(j-1) and vice versa, and both tiles share every K / V block that is loaded.
__global__ void __launch_bounds__(kAttnThreads, 1)
attention_kernel(const __grid_constant__ CUtensorMap tmQKV, bf16* __restrict__ out, int n_tok, int D) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* q_full = bars;            // 1
  uint64_t* k_full = bars + 1;        // 2 stages
  uint64_t* v_full = bars + 3;
  uint64_t* k_empty = bars + 5;
  uint64_t* v_empty = bars + 7;
  uint64_t* s_full = bars + 9;        // per tile
  uint64_t* p_full = bars + 11;
  uint64_t* o_full = bars + 13;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int q0 = blockIdx.x * 2 * kBQ;
  const int head = blockIdx.y;
  const int img = blockIdx.z;
  const int row_base = img * n_tok;               // first row of this image in the qkv matrix
  const int nblk = (n_tok + kBKV - 1) / kBKV;
  const bool has_b = q0 + kBQ < n_tok;            // second query tile holds at least one real row
  const int colQ = head * kHD, colK = D + head * kHD, colV = 2 * D + head * kHD;

  if (tid == 0) {
    tma_prefetch_desc(&tmQKV);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], kBQ);
      mbar_init(&o_full[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc<kAttnTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= 8) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
    if (warp == 9 && lane == 0) {
      // ---------------------------------------------------------------- TMA producer
      mbar_expect_tx(q_full, has_b ? 2 * kTileBytes : kTileBytes);
      tma_load_3d(smem + kOffQ, &tmQKV, q_full, colQ, row_base + q0, 0, kEvictNormal);
      if (has_b) tma_load_3d(smem + kOffQ + kTileBytes, &tmQKV, q_full, colQ, row_base + q0 + kBQ, 0, kEvictNormal);
      for (int j = 0; j < nblk; ++j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        mbar_wait(&k_empty[st], ph ^ 1);
        mbar_expect_tx(&k_full[st], kTileBytes);
        tma_load_3d(smem + kOffK + st * kTileBytes, &tmQKV, &k_full[st], colK, row_base + j * kBKV, 0, kEvictLast);
        mbar_wait(&v_empty[st], ph ^ 1);
        mbar_expect_tx(&v_full[st], kTileBytes);
        tma_load_3d(smem + kOffV + st * kTileBytes, &tmQKV, &v_full[st], colV, row_base + j * kBKV, 0, kEvictLast);
      }
    } else if (warp == 8 && lane == 0) {
      // ---------------------------------------------------------------- MMA issuer
      constexpr uint32_t idescS = umma_idesc_bf16(128, 128, false);
      constexpr uint32_t idescPV = umma_idesc_bf16(128, 64, true);
      const uint32_t sbase = smem_u32(smem);
      auto issue_S = [&](int tile, int st) {
        const uint32_t qa = sbase + kOffQ + tile * kTileBytes, ka = sbase + kOffK + st * kTileBytes;
#pragma unroll
        for (int k = 0; k < kHD / 16; ++k)
          umma_bf16(tmem_base + tile * 128, umma_smem_desc(qa + k * 32, 16, 1024), umma_smem_desc(ka + k * 32, 16, 1024),
                    idescS, k != 0);
      };
      auto issue_PV = [&](int tile, int st, bool first_block) {
        const uint32_t pa = sbase + kOffP + tile * kPBytes, va = sbase + kOffV + st * kTileBytes;
#pragma unroll
        for (int kk = 0; kk < kBKV / 16; ++kk)
          umma_bf16(tmem_base + 256 + tile * 64, umma_smem_desc(pa + (kk >> 2) * (kBQ * 128) + (kk & 3) * 32, 16, 1024),
                    umma_smem_desc(va + kk * 2048, 1024, 1024),   // MN-major V: 16 keys = 2 x 1024 B groups
                    idescPV, (!first_block || kk != 0) ? 1u : 0u);   // O accumulates in TMEM across key blocks
      };
      mbar_wait(q_full, 0);
      mbar_wait(&k_full[0], 0);
      tc_fence_after();
      issue_S(0, 0);
      umma_commit(&s_full[0]);
      if (has_b) {
        issue_S(1, 0);
        umma_commit(&s_full[1]);
      }
      umma_commit(&k_empty[0]);
      for (int j = 0; j < nblk; ++j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        const bool more = j + 1 < nblk;
        const int nst = (j + 1) & 1;
        mbar_wait(&p_full[0], j & 1);              // P_A(j) written, S_A(j) fully read, O_A rescaled if needed
        mbar_wait(&v_full[st], ph);
        tc_fence_after();
        issue_PV(0, st, j == 0);
        umma_commit(&o_full[0]);
        if (more) {
          mbar_wait(&k_full[nst], ((j + 1) >> 1) & 1);
          tc_fence_after();
          issue_S(0, nst);
          umma_commit(&s_full[0]);
        }
        if (has_b) {
          mbar_wait(&p_full[1], j & 1);
          tc_fence_after();
          issue_PV(1, st, j == 0);
          umma_commit(&o_full[1]);
          if (more) {
            issue_S(1, nst);
            umma_commit(&s_full[1]);
          }
        }
        umma_commit(&v_empty[st]);
        if (more) umma_commit(&k_empty[nst]);
      }
    }
    __syncwarp();
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    // ------------------------------------------------------------------ softmax warpgroups
    const int tile = warp >> 2;                       // 0 = A, 1 = B
    const int row = (warp & 3) * 32 + lane;           // row inside the tile = TMEM lane
    if (tile == 0 || has_b) {
      const uint32_t lane_field = static_cast<uint32_t>((warp & 3) * 32) << 16;
      const uint32_t tS = tmem_base + tile * 128 + lane_field;
      const uint32_t tPV = tmem_base + 256 + tile * 64 + lane_field;
      uint8_t* sP = smem + kOffP + tile * kPBytes;
      const uint32_t p_row_off = static_cast<uint32_t>(row >> 3) * 1024 + static_cast<uint32_t>(row & 7) * 128;
      const uint32_t sw = static_cast<uint32_t>(row & 7);

      float m_run = -INFINITY;   // reference max of the exponentials, already multiplied by log2(e)
      float l_run = 0.f;

      // TMEM reads are the scarce resource of this kernel (~64 B/clk/SM), so every score is read exactly
      // once: the thread pulls its whole 128-key row of S into registers, and O stays in TMEM, accumulated
      // by the PV MMAs across key blocks.  The running max is only raised (and O / l rescaled) when a block's
      // max exceeds it by more than 2^kRescaleLog2 - until then p = exp2(s - m_run) may exceed 1 (bounded by
      // 256: exact in fp32 sums, representable in bf16), which leaves O / l unchanged mathematically.
      constexpr float kRescaleLog2 = 8.0f;
      for (int j = 0; j < nblk; ++j) {
        const uint32_t par = j & 1;
        mbar_wait(&s_full[tile], par);
        tc_fence_after();
        const int valid = n_tok - j * kBKV;   // keys [0, valid) of this block exist
        uint32_t sc[4][32];                   // the row of S (fp32 bits), later reused for packed bf16 P
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld32(tS + c * 32, sc[c]);
        tmem_ld_wait();
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
        if (valid >= kBKV) {
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              mx0 = fmaxf(mx0, __uint_as_float(sc[c][i]));
              mx1 = fmaxf(mx1, __uint_as_float(sc[c][i + 1]));
              mx2 = fmaxf(mx2, __uint_as_float(sc[c][i + 2]));
              mx3 = fmaxf(mx3, __uint_as_float(sc[c][i + 3]));
            }
        } else {
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              if (c * 32 + i >= valid) sc[c][i] = 0xff800000u;      // -inf: ignored by the max, exp2 -> 0
              mx0 = fmaxf(mx0, __uint_as_float(sc[c][i]));
            }
        }
#ifdef VVS_EXPERIMENT_NO_MAX
        mx0 = __uint_as_float(sc[0][0]); mx1 = __uint_as_float(sc[1][1]);
#endif
        const float m_blk = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * kLog2e;
        const bool raise = m_blk > m_run + kRescaleLog2;           // always true for j == 0 (m_run = -inf)
        const float alpha = raise ? ex2_approx(m_run - m_blk) : 1.0f;
        if (raise) m_run = m_blk;
        const float neg_m = -m_run;
        float ps0 = 0.f, ps1 = 0.f, ps2 = 0.f, ps3 = 0.f;
        uint32_t pk[4][16];                   // bf16x2-packed P row
#pragma unroll
        for (int c = 0; c < 4; ++c)
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

        if (j > 0) {
          // PV(j-1) must have finished: it reads the P buffer we are about to overwrite and writes the O
          // accumulator a rescale would touch.
          mbar_wait(&o_full[tile], (j - 1) & 1);
          tc_fence_after();
          if (__any_sync(0xffffffffu, raise)) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              uint32_t ov[32];
              tmem_ld32(tPV + h * 32, ov);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) ov[i] = __float_as_uint(__uint_as_float(ov[i]) * alpha);
              tmem_st32(tPV + h * 32, ov);
            }
            tmem_st_wait();
          }
        }
#ifndef VVS_EXPERIMENT_NO_PSTORE
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint8_t* atom = sP + (c >> 1) * (kBQ * 128) + p_row_off;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t chunk = static_cast<uint32_t>((c & 1) * 4 + q);
            *reinterpret_cast<uint4*>(atom + ((chunk ^ sw) << 4)) =
                make_uint4(pk[c][q * 4 + 0], pk[c][q * 4 + 1], pk[c][q * 4 + 2], pk[c][q * 4 + 3]);
          }
        }
#else
        if (pk[0][0] == 0x12345678u && pk[1][3] == 7u && pk[2][5] == 9u && pk[3][15] == 1u) sP[p_row_off] = 1;
#endif
        fence_proxy_async_smem();   // P (generic proxy) -> visible to the tensor core (async proxy)
        tc_fence_before();
        mbar_arrive(&p_full[tile]);
      }

      mbar_wait(&o_full[tile], (nblk - 1) & 1);
      tc_fence_after();
      const int q = q0 + tile * kBQ + row;
      const float inv = 1.0f / l_run;
      uint4* dst = reinterpret_cast<uint4*>(out + static_cast<size_t>(row_base + q) * D + head * kHD);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t ov[32];
        tmem_ld32(tPV + h * 32, ov);
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
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc<kAttnTmemCols>(tmem_base);
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
  dim3 grid((n_tok + 2 * kBQ - 1) / (2 * kBQ), heads, n_img);
  attention_kernel<<<grid, kAttnThreads, kAttnSmem, stream>>>(tm, out, n_tok, D);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

}  // namespace vvs
