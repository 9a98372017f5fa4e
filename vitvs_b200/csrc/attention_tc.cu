// tcgen05 flash-attention forward for DINOv2 blocks (reference semantics:
// dino_patch/attention.py:70-80 - non-causal SDPA, scale hd^-0.5, head dim 64, no mask, eval).
//
// Input is the fused-QKV GEMM output [n_img * n_tok, 3*D] bf16 exactly as the reference's
// `qkv(x).reshape(B, N, 3, H, hd)` lays it out (q already multiplied by hd^-0.5 in that GEMM's
// epilogue), so no head-major scatter is needed: the Q / K / V tiles of (image, head) are 64-column
// TMA boxes of that matrix.  Output goes straight into the [n_img * n_tok, D] layout attn.proj reads.
//
// One CTA (one per SM, 20 warps) = up to FOUR 128-row query tiles of one (image, head) - the whole
// sequence at the default 485 tokens - sharing every K / V block (64 keys) of a 4-stage TMA ring:
//   warps 0-15   four softmax warpgroups, one per query tile (thread = query row = TMEM lane)
//   warps 16-19  one MMA issuer thread per tile (issuer 0 is also the TMA producer)
// The kernel is not tensor-bound: per 128 x 64 block the MMAs need ~64 clk, the 8192 exponentials 512 clk of
// MUFU, and the softmax instruction stream (~350 warp instructions per block and warp) several times that
// (measured: removing every MUFU.EX2 shortens the kernel by 9 %, removing the TMEM reads by nothing).  Four
// softmax warps per SM sub-partition (one from each warpgroup) give the scheduler something to issue while
// a warp waits on a dependency / MUFU / an mbarrier, and each tile's S runs one block ahead of its warpgroup:
//   S_t(j+1) = Q_t K(j+1)^T is issued as soon as warpgroup t has pulled S_t(j) into registers;
//   softmax t: max / exp2 / sum on the row in registers (packed fp32x2 FMA / ADD), bf16 P -> swizzled smem;
//   O_t += P_t V(j) (V consumed MN-major from its row-major tile), O accumulated in TMEM across blocks and
//   lazily rescaled only when the running max grows by more than 2^8;  O / l once at the end.
// TMEM: tile t owns columns [128 t, 128 t + 64) for S and [128 t + 64, 128 t + 128) for O (512 in all).
// Keys past n_tok (the next image's rows or TMA zero fill) are masked to p = 0.
#include <cuda.h>

#include <cstdlib>
#include <string>

#include "ptx.cuh"
#include "vvs_internal.h"

namespace vvs {
namespace {

constexpr int kBQ = 128;   // query rows per tile (one TMEM lane per row)
constexpr int kBKV = 64;   // keys per block
constexpr int kHD = 64;    // head dim (every DINOv2 size)
constexpr int kQT = 4;     // query tiles per CTA
constexpr int kAttnThreads = (kQT * 4 + kQT) * 32;   // 640: 16 softmax warps + one MMA issuer warp per tile, 96 regs
constexpr int kKVStages = 4;
constexpr uint32_t kQBytes = kBQ * kHD * 2;           // 16 KB
constexpr uint32_t kKVBytes = kBKV * kHD * 2;         // 8 KB: one K or V block
constexpr uint32_t kPBytes = kBQ * kBKV * 2;          // 16 KB: P block, one 64-key K-major atom column
constexpr uint32_t kOffQ = 0, kOffK = kOffQ + kQT * kQBytes, kOffV = kOffK + kKVStages * kKVBytes,
                   kOffP = kOffV + kKVStages * kKVBytes, kOffBar = kOffP + kQT * kPBytes;
constexpr uint32_t kAttnSmem = kOffBar + 512 + 1024;
constexpr uint32_t kAttnTmemCols = 512;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kRescaleLog2 = 8.0f;
#ifndef VVS_ATT_STAGGER
#define VVS_ATT_STAGGER 0
#endif

__device__ __forceinline__ float ex2_approx(float x) {   // single MUFU.EX2; ex2(-inf) = 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(kAttnThreads, 1)
attention_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                 bf16* __restrict__ out, int n_tok, int D) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* q_full = bars;                       // kQT
  uint64_t* k_full = q_full + kQT;               // kKVStages
  uint64_t* v_full = k_full + kKVStages;
  uint64_t* k_empty = v_full + kKVStages;
  uint64_t* v_empty = k_empty + kKVStages;
  uint64_t* s_full = v_empty + kKVStages;        // kQT: S_t(j) is in TMEM
  uint64_t* s_free = s_full + kQT;               // kQT: warpgroup t holds S_t(j) in registers (4 warp arrivals)
  uint64_t* p_full = s_free + kQT;               // kQT: P_t(j) is in smem, O_t rescaled if needed (4 warp arrivals)
  uint64_t* pv_done = p_full + kQT;              // kQT: O_t += P_t(j) V(j) has completed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pv_done + kQT);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int qt0 = blockIdx.x * kQT;               // first query tile of this CTA
  const int head = blockIdx.y;
  const int img = blockIdx.z;
  const int row_base = img * n_tok;               // first row of this image in the qkv matrix
  const int nblk = (n_tok + kBKV - 1) / kBKV;
  const int n_qt = min(kQT, (n_tok + kBQ - 1) / kBQ - qt0);   // active tiles (the last CTA of a long sequence)
  const int colQ = head * kHD, colK = D + head * kHD, colV = 2 * D + head * kHD;

  if (tid == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
    for (int i = 0; i < kKVStages; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&k_empty[i], n_qt);    // one commit per tile's issuer
      mbar_init(&v_empty[i], n_qt);
    }
    for (int i = 0; i < kQT; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&s_free[i], 4);
      mbar_init(&p_full[i], 4);
      mbar_init(&pv_done[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == kQT * 4) tmem_alloc<kAttnTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= kQT * 4) {
    // ------------------------------------------------------------------ MMA issuer of tile t (one lane)
    // Every tile has its own issuer thread, so no tile ever waits behind another tile's barrier.  The events
    // an issuer waits for arrive in a fixed order - S_t(j) consumed, then P_t(j) published - hence blocking
    // waits never stall work that is ready.  Issuer 0 doubles as the TMA producer (Q tiles, K / V ring).
    const int t = warp - kQT * 4;
    if (lane == 0 && t < n_qt) {
      constexpr uint32_t idescS = umma_idesc_bf16(128, kBKV, false);
      constexpr uint32_t idescPV = umma_idesc_bf16(128, kHD, true);
      const uint32_t sbase = smem_u32(smem);
      const uint32_t qa = sbase + kOffQ + t * kQBytes;
      const uint32_t pa = sbase + kOffP + t * kPBytes;
      const bool producer = (t == 0);
      auto load_K = [&](int b) {      // block b -> stage b % kKVStages once every tile's S(b - kKVStages) is done
        const int st = b % kKVStages;
        mbar_wait(&k_empty[st], ((b / kKVStages) & 1) ^ 1);
        mbar_expect_tx(&k_full[st], kKVBytes);
        tma_load_3d(smem + kOffK + st * kKVBytes, &tmKV, &k_full[st], colK, row_base + b * kBKV, 0, kEvictLast);
      };
      auto load_V = [&](int b) {
        const int st = b % kKVStages;
        mbar_wait(&v_empty[st], ((b / kKVStages) & 1) ^ 1);
        mbar_expect_tx(&v_full[st], kKVBytes);
        tma_load_3d(smem + kOffV + st * kKVBytes, &tmKV, &v_full[st], colV, row_base + b * kBKV, 0, kEvictLast);
      };
      auto issue_S = [&](int b) {     // S_t(b) = Q_t K(b)^T
        const int st = b % kKVStages;
        mbar_wait(&k_full[st], (b / kKVStages) & 1);
        if (b == 0) mbar_wait(&q_full[t], 0);
        else mbar_wait(&s_free[t], (b - 1) & 1);      // warpgroup t holds S_t(b-1) in registers
        tc_fence_after();
        const uint32_t ka = sbase + kOffK + st * kKVBytes;
#pragma unroll
        for (int k = 0; k < kHD / 16; ++k)
          umma_bf16(tmem_base + t * 128, umma_smem_desc(qa + k * 32, 16, 1024), umma_smem_desc(ka + k * 32, 16, 1024),
                    idescS, k != 0);
        umma_commit(&s_full[t]);
        umma_commit(&k_empty[st]);
      };
      if (producer) {
        for (int u = 0; u < n_qt; ++u) {
          mbar_expect_tx(&q_full[u], kQBytes);
          tma_load_3d(smem + kOffQ + u * kQBytes, &tmQ, &q_full[u], colQ, row_base + (qt0 + u) * kBQ, 0, kEvictNormal);
          if (u == 0) load_K(0);      // first K block right behind the first Q tile
        }
        for (int b = 0; b < kKVStages - 1 && b < nblk; ++b) {
          if (b > 0) load_K(b);
          load_V(b);
        }
      }
      // The four tiles would otherwise run in lock step - every warp of an SM sub-partition in its TMEM-read
      // phase at once, then every warp in its MUFU phase - serialising the two bottleneck resources.  A one-off
      // start offset per tile keeps them out of phase (nothing couples the tiles tighter than the K/V ring).
      if (t > 0) {
        const long long t0 = clock64();
        while (clock64() - t0 < static_cast<long long>(t) * VVS_ATT_STAGGER) {
        }
      }
      issue_S(0);
      for (int j = 0; j < nblk; ++j) {
        if (producer && j + kKVStages - 1 < nblk) load_K(j + kKVStages - 1);
        if (j + 1 < nblk) issue_S(j + 1);             // one block ahead of the softmax warpgroup
        const int st = j % kKVStages;
        mbar_wait(&v_full[st], (j / kKVStages) & 1);
        mbar_wait(&p_full[t], j & 1);                 // P_t(j) in smem, O_t rescaled if it had to be
        tc_fence_after();
        const uint32_t va = sbase + kOffV + st * kKVBytes;
#pragma unroll
        for (int kk = 0; kk < kBKV / 16; ++kk)
          umma_bf16(tmem_base + t * 128 + 64, umma_smem_desc(pa + kk * 32, 16, 1024),
                    umma_smem_desc(va + kk * 2048, 1024, 1024),   // MN-major V: 16 keys = 2 x 1024 B groups
                    idescPV, (j > 0 || kk != 0) ? 1u : 0u);        // O accumulates in TMEM across key blocks
        umma_commit(&pv_done[t]);
        umma_commit(&v_empty[st]);
        if (producer && j + kKVStages - 1 < nblk) load_V(j + kKVStages - 1);
      }
    }
    __syncwarp();
  } else if ((warp >> 2) < n_qt) {
    // ------------------------------------------------------------------ softmax warpgroup of tile t
    const int t = warp >> 2;
    const int wq = warp & 3;                          // TMEM lane quarter of this warp
    const int row = wq * 32 + lane;                   // row inside the tile = TMEM lane
    const uint32_t lane_field = static_cast<uint32_t>(wq * 32) << 16;
    const uint32_t tS = tmem_base + t * 128 + lane_field;
    const uint32_t tO = tS + 64;
    const uint32_t sw = static_cast<uint32_t>(row & 7);
    const uint32_t p_row_s = smem_u32(smem + kOffP + t * kPBytes) + static_cast<uint32_t>(row >> 3) * 1024 + sw * 128;
    float m_run = -INFINITY;   // reference max of the exponentials, already multiplied by log2(e)
    float l_run = 0.f;
    const f32x2 log2e2 = f2_pack(kLog2e, kLog2e);

    for (int j = 0; j < nblk; ++j) {
      const uint32_t ph = j & 1;
      mbar_wait(&s_full[t], ph);
      tc_fence_after();
      uint32_t sc[2][32];                   // the row of S (fp32 bits)
      tmem_ld32(tS, sc[0]);
      tmem_ld32(tS + 32, sc[1]);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s_free[t]);       // S_t(j+1) may overwrite the TMEM buffer
      const int valid = n_tok - j * kBKV;   // keys [0, valid) of this block exist
      if (valid < kBKV) {
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c * 32 + i >= valid) sc[c][i] = 0xff800000u;      // -inf: ignored by the max, exp2 -> 0
      }
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
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
      // exponentials: the first 32 columns are computed before waiting for PV(j-1) (whose completion frees the
      // P buffer), so that wait is normally over by the time it is reached; the rest are packed and stored
      // 8 at a time (few live registers).
      const f32x2 neg_m2 = f2_pack(-m_run, -m_run);
      f32x2 ps01 = f2_pack(0.f, 0.f), ps23 = ps01;
      uint32_t pk0[16];
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const f32x2 e01 = f2_ex2(f2_fma(f2_pack(__uint_as_float(sc[0][i + 0]), __uint_as_float(sc[0][i + 1])), log2e2, neg_m2));
        const f32x2 e23 = f2_ex2(f2_fma(f2_pack(__uint_as_float(sc[0][i + 2]), __uint_as_float(sc[0][i + 3])), log2e2, neg_m2));
        ps01 = f2_add(ps01, e01);
        ps23 = f2_add(ps23, e23);
        pk0[i / 2] = f2_to_bf16x2(e01);
        pk0[i / 2 + 1] = f2_to_bf16x2(e23);
      }
      if (j >= 1) {
        mbar_wait(&pv_done[t], (j - 1) & 1);    // PV(j-1) has finished: P buffer free, O quiescent
        if (__any_sync(0xffffffffu, raise)) {
          tc_fence_after();
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            uint32_t ov[16];
            tmem_ld16(tO + h * 16, ov);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) ov[i] = __float_as_uint(__uint_as_float(ov[i]) * alpha);
            tmem_st16(tO + h * 16, ov);
          }
          tmem_st_wait();
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q)
        st_shared_v4(p_row_s + ((static_cast<uint32_t>(q) ^ sw) << 4), pk0[q * 4 + 0], pk0[q * 4 + 1], pk0[q * 4 + 2],
                     pk0[q * 4 + 3]);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint32_t pk[4];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int i = q * 8 + h * 4;
          const f32x2 e01 = f2_ex2(f2_fma(f2_pack(__uint_as_float(sc[1][i + 0]), __uint_as_float(sc[1][i + 1])), log2e2, neg_m2));
          const f32x2 e23 = f2_ex2(f2_fma(f2_pack(__uint_as_float(sc[1][i + 2]), __uint_as_float(sc[1][i + 3])), log2e2, neg_m2));
          ps01 = f2_add(ps01, e01);
          ps23 = f2_add(ps23, e23);
          pk[h * 2 + 0] = f2_to_bf16x2(e01);
          pk[h * 2 + 1] = f2_to_bf16x2(e23);
        }
        st_shared_v4(p_row_s + ((static_cast<uint32_t>(4 + q) ^ sw) << 4), pk[0], pk[1], pk[2], pk[3]);
      }
      float s0, s1;
      f2_unpack(f2_add(ps01, ps23), s0, s1);
      l_run = l_run * alpha + (s0 + s1);
      fence_proxy_async_smem();   // P (generic proxy) -> visible to the tensor core (async proxy)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[t]);
    }

    mbar_wait(&pv_done[t], (nblk - 1) & 1);
    tc_fence_after();
    const int q = (qt0 + t) * kBQ + row;
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
  if (warp == kQT * 4) tmem_dealloc<kAttnTmemCols>(tmem_base);
}

}  // namespace

int launch_attention(const bf16* qkv, bf16* out, int n_img, int n_tok, int heads, cudaStream_t stream,
                     std::string* err, bool pdl) {
  static const bool use_v6 = []() {
    const char* e = getenv("VITVS_ATTN");
    return e && e[0] == '6';
  }();
  if (!use_v6) return launch_attention_ts(qkv, out, n_img, n_tok, heads, stream, err, pdl);
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
  const int q_tiles = (n_tok + kBQ - 1) / kBQ;
  dim3 grid((q_tiles + kQT - 1) / kQT, heads, n_img);
  attention_kernel<<<grid, kAttnThreads, kAttnSmem, stream>>>(tmq, tmkv, out, n_tok, D);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

}  // namespace vvs
