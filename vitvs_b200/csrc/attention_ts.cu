// tcgen05 flash-attention forward, generation 9 (reference semantics:
// dino_patch/attention.py:70-80 - non-causal SDPA, scale hd^-0.5, head dim 64, no mask, eval).
//
// Input is the fused-QKV GEMM output [n_img * n_tok, 3*D] bf16 exactly as the reference's
// `qkv(x).reshape(B, N, 3, H, hd)` lays it out (q already multiplied by hd^-0.5 in that GEMM's epilogue):
// the Q / K / V tiles of (image, head) are 64-column TMA boxes of that matrix.  Output goes straight into
// the [n_img * n_tok, D] layout attn.proj reads.
//
// The kernel is bound by the exponentials: 128 x 64 scores cost 512 clk of MUFU.EX2 per SM sub-partition
// against 64 clk of tensor pipe, so everything is arranged to keep the MUFU fed (tools/ubench/softmax_math.cu:
// the softmax arithmetic alone runs at 1.13x the MUFU floor with two warps per sub-partition) and to keep
// every hand-off latency off the softmax warps' critical path (measured: an mbarrier wait costs 100-250 clk
// even when it succeeds at once, P hand-off -> next S in TMEM ~900 clk, CTA launch + prologue + epilogue ~9000 clk):
//   * PERSISTENT: one CTA per SM walks a static list of work items (image, head, pair of 128-row query tiles);
//     all pipelines - K / V ring, S buffers, Q tiles - run straight across item boundaries, so the next item's
//     Q / K are in shared memory and its first scores in TMEM before the current item's last block is done, and
//     the current item's O / l epilogue is deferred until after the first block of the next item.
//       warps 0-7   two softmax warpgroups, one per tile (thread = query row = TMEM lane)
//       warps 8-9   one MMA issuer thread per tile     warp 10   TMA producer (Q tiles, K / V rings)
//   * S is triple buffered in TMEM and runs THREE blocks ahead: S(g+3) is issued right behind PV(g), so the
//     scores are in TMEM a whole block before the softmax warpgroup asks for them; the warpgroup pulls S(g+1)
//     into a second register set and takes its row max in the middle of block g's exponentials (software
//     pipelining inside the warp), so consecutive blocks' MUFU work is back to back.
//   * P never leaves the tensor-core datapath: bf16 P(g) is written with tcgen05.st over the first 32 columns
//     of S(g)'s buffer and O += P V is a TS MMA (A = P from TMEM, B = V consumed MN-major from its row-major
//     tile): no shared-memory store, no generic->async proxy fence, no smem operand bandwidth.  tcgen05.mma
//     executes in issue order, which is what makes overwriting P(g) with S(g+3) safe.
//   * O accumulates in TMEM across blocks and is lazily rescaled only when the running max grows by more
//     than 2^8 (then, and in the epilogue, the warpgroup looks at pv_done);  O / l once per item.
//   * Warp indices come from a shuffle broadcast so that the compiler keeps every address, descriptor and
//     barrier on the uniform datapath (with threadIdx.x >> 5 each tcgen05.mma sat in a ~100 clk waterfall loop).
// TMEM: tile t owns columns [256 t, 256 t + 192) for S / P (three 64-column buffers) and
// [256 t + 192, 256 t + 256) for O.  Keys past n_tok (the next image's rows or TMA zero fill) are masked to p = 0.
#include <cuda.h>

#include <cstdio>
#include <cstdlib>
#include <string>

#include "ptx.cuh"
#include "vvs_internal.h"

namespace vvs {
namespace {

constexpr int kBQ = 128;   // query rows per tile (one TMEM lane per row)
constexpr int kBKV = 64;   // keys per block
constexpr int kHD = 64;    // head dim (every DINOv2 size)
constexpr int kQT = 2;     // query tiles per work item
constexpr int kThreads = (kQT * 4 + kQT + 1) * 32;   // 352: 8 softmax warps, 2 MMA issuers, 1 TMA producer
constexpr int kKVStages = 6;
#ifndef VVS_ATT_SBUFS
#define VVS_ATT_SBUFS 3
#endif
constexpr int kSBufs = VVS_ATT_SBUFS;  // S buffers per tile = how many blocks S runs ahead of PV
constexpr uint32_t kQBytes = kBQ * kHD * 2;      // 16 KB
constexpr uint32_t kKVBytes = kBKV * kHD * 2;    // 8 KB: one K or V block
constexpr uint32_t kOffQ = 0, kOffK = kOffQ + kQT * kQBytes, kOffV = kOffK + kKVStages * kKVBytes,
                   kOffBar = kOffV + kKVStages * kKVBytes;
constexpr uint32_t kSmem = kOffBar + 512 + 1024;   // + barriers + alignment slack
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kTileCols = 256;              // TMEM columns per tile: S0 | S1 | S2 | O
constexpr uint32_t kOCol = kSBufs * 64;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kRescaleLog2 = 8.0f;
#ifndef VVS_ATT_STAGGER
#define VVS_ATT_STAGGER 0
#endif
#ifndef VVS_EXP_TILE_BAR
#define VVS_EXP_TILE_BAR 4
#endif

struct AttnDev {
  bf16* out;
  int n_tok, D, heads;
  int n_half;      // work items per (image, head): pairs of query tiles
  int q_tiles;     // 128-row tiles per sequence
  int n_items;     // n_img * heads * n_half
};

__device__ __forceinline__ float ex2_approx(float x) {   // single MUFU.EX2; ex2(-inf) = 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// D[tmem] (+)= A[tmem] * B[smem]: A = 128 lanes x (K bf16 packed two per 32-bit column)
__device__ __forceinline__ void umma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// 16 exponentials of a row segment: p = exp2(s * log2e - m), packed to bf16 pairs, row sum accumulated
__device__ __forceinline__ void exp_pack16(const uint32_t* s, f32x2 log2e2, f32x2 neg_m2, f32x2& sum01, f32x2& sum23,
                                           uint32_t* pk) {
#pragma unroll
  for (int i = 0; i < 16; i += 4) {
    const f32x2 e01 = f2_ex2(f2_fma(f2_pack(__uint_as_float(s[i + 0]), __uint_as_float(s[i + 1])), log2e2, neg_m2));
    const f32x2 e23 = f2_ex2(f2_fma(f2_pack(__uint_as_float(s[i + 2]), __uint_as_float(s[i + 3])), log2e2, neg_m2));
    sum01 = f2_add(sum01, e01);
    sum23 = f2_add(sum23, e23);
    pk[i / 2] = f2_to_bf16x2(e01);
    pk[i / 2 + 1] = f2_to_bf16x2(e23);
  }
}

#ifdef VVS_ATT_TRACE   // timing experiment: per-block clock64 stamps of one CTA (tools/att_trace.py)
__device__ long long g_att_trace[4 * 3 * 16 * 8];
#define VVS_TR(ev)                                                                                         \
  do {                                                                                                     \
    if (tr_warp >= 0 && lane == 0 && g < 64) g_att_trace[((g >> 4) * 3 + tr_warp) * 128 + (g & 15) * 8 + (ev)] = clock64(); \
  } while (0)
#else
#define VVS_TR(ev) do { } while (0)
#endif

// row max of a 64-score block held in registers, times log2(e); keys >= valid are first set to -inf
__device__ __forceinline__ float masked_row_max(uint32_t (&sc)[2][32], int valid) {
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
  return fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * kLog2e;
}

// work item -> (first row of the image in the qkv matrix, head, first query tile, active tiles)
struct Item {
  int row_base, head, qt0, n_qt;
};
__device__ __forceinline__ Item decode_item(const AttnDev& p, int item) {
  const int half = item % p.n_half;
  const int ih = item / p.n_half;
  Item it;
  it.head = ih % p.heads;
  it.row_base = (ih / p.heads) * p.n_tok;
  it.qt0 = half * kQT;
  it.n_qt = min(kQT, p.q_tiles - it.qt0);
  return it;
}

__global__ void __launch_bounds__(kThreads, 1)
attention_ts_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV, const AttnDev p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* q_full = bars;                       // kQT: Q tile of the current item has landed
  uint64_t* q_empty = q_full + kQT;              // kQT: every S MMA of the item that read the Q tile is done
  uint64_t* k_full = q_empty + kQT;              // kKVStages
  uint64_t* v_full = k_full + kKVStages;
  uint64_t* k_empty = v_full + kKVStages;        // 2 arrivals: one per tile's issuer
  uint64_t* v_empty = k_empty + kKVStages;
  uint64_t* s_full = v_empty + kKVStages;        // kQT x kSBufs: S_t(g) is in TMEM buffer g % kSBufs
  // kQT x 2 (block parity): P_t(g) is in TMEM, O_t rescaled if needed (4 warp arrivals).  Two barriers because a
  // fast warp may finish block g+1 (its S is there: S runs kSBufs ahead) before a slow one finishes block g; with
  // one barrier its second arrival would complete block g's phase.
  uint64_t* p_full = s_full + kSBufs * kQT;
  uint64_t* pv_done = p_full + 2 * kQT;              // kQT: O_t += P_t(g) V(g) has completed
  uint64_t* o_free = pv_done + kQT;              // kQT: the item's O_t is in registers (4 warp arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_free + kQT);

  const int tid = threadIdx.x;
  // warp-uniform by construction AND provably so for the compiler (shuffle broadcast): everything derived from
  // it - tile index, barrier / smem / TMEM addresses, MMA descriptors - then lives on the uniform datapath.
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int lane = tid & 31;
  const int n_tok = p.n_tok, D = p.D;
  const int nblk = (n_tok + kBKV - 1) / kBKV;

  if (tid == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
    for (int i = 0; i < kKVStages; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&k_empty[i], kQT);
      mbar_init(&v_empty[i], kQT);
    }
    for (int i = 0; i < kQT; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
      for (int b = 0; b < kSBufs; ++b) mbar_init(&s_full[kSBufs * i + b], 1);
      mbar_init(&p_full[2 * i], 4);
      mbar_init(&p_full[2 * i + 1], 4);
      mbar_init(&pv_done[i], 1);
      mbar_init(&o_free[i], 4);
    }
    fence_mbar_init();
  }
  if (warp == kQT * 4) tmem_alloc<kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  if (warp == kQT * 4 + kQT) {
    // ------------------------------------------------------------------ TMA producer (one lane)
    if (lane == 0) {
      int g0 = 0;              // K / V blocks loaded before this item (ring position)
      int n_act[kQT] = {};     // items each tile has taken part in (Q buffer phases)
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, g0 += nblk) {
        const Item it = decode_item(p, item);
        const int colQ = it.head * kHD, colK = D + it.head * kHD, colV = 2 * D + it.head * kHD;
        for (int u = 0; u < it.n_qt; ++u) {
          mbar_wait(&q_empty[u], (n_act[u]++ & 1) ^ 1);
          mbar_expect_tx(&q_full[u], kQBytes);
          tma_load_3d(smem + kOffQ + u * kQBytes, &tmQ, &q_full[u], colQ, it.row_base + (it.qt0 + u) * kBQ, 0, kEvictNormal);
        }
        // consumption order: K(0) .. K(kSBufs-1) V(0) K(kSBufs) V(1) ... (S runs kSBufs blocks ahead of PV)
        auto load = [&](int b, bool is_v) {
          const int g = g0 + b, st = g % kKVStages;
          mbar_wait(is_v ? &v_empty[st] : &k_empty[st], ((g / kKVStages) & 1) ^ 1);
          uint64_t* full = is_v ? &v_full[st] : &k_full[st];
          mbar_expect_tx(full, kKVBytes);
          tma_load_3d(smem + (is_v ? kOffV : kOffK) + st * kKVBytes, &tmKV, full, is_v ? colV : colK,
                      it.row_base + b * kBKV, 0, kEvictLast);
        };
        for (int b = 0; b < kSBufs && b < nblk; ++b) load(b, false);
        for (int b = 0; b < nblk; ++b) {
          load(b, true);
          if (b + kSBufs < nblk) load(b + kSBufs, false);
        }
      }
    }
    __syncwarp();
  } else if (warp >= kQT * 4) {
    // ------------------------------------------------------------------ MMA issuer of tile t (one lane)
    const int t = warp - kQT * 4;
    if (lane == 0) {
      constexpr uint32_t idescS = umma_idesc_bf16(128, kBKV, false);
      constexpr uint32_t idescPV = umma_idesc_bf16(128, kHD, true);
      const uint32_t sbase = smem_u32(smem);
      const uint32_t qa = sbase + kOffQ + t * kQBytes;
      const uint32_t tT = tmem_base + t * kTileCols;
      const uint32_t tO = tT + kOCol;
      // Two cursors over this tile's block sequence h = 0, 1, 2, ... (continuous across work items):
      //   S cursor  - S(h) may be issued as soon as PV(h - kSBufs) has been (its TMEM buffer is free in issue order)
      //   PV cursor - PV(h) follows the warpgroup's P(h)
      // Inactive items (a sequence's last, single-tile pair) only return the shared K / V ring's stages.
      auto active = [&](int item) { return t < decode_item(p, item).n_qt; };
      int s_item = blockIdx.x, s_b = 0, s_h = 0, s_g0 = 0, s_nact = 0;
      while (s_item < p.n_items && !active(s_item)) {
        s_item += gridDim.x;
        s_g0 += nblk;
      }
      // issue S at the cursor; when !blocking, give up (false) if its Q tile or K block has not landed yet
      auto issue_S = [&](bool blocking) -> bool {
        const int g = s_g0 + s_b, st = g % kKVStages;
        if (s_b == 0) {
          if (blocking) mbar_wait(&q_full[t], s_nact & 1);
          else if (!mbar_test_wait(&q_full[t], s_nact & 1)) return false;
        }
        if (blocking) mbar_wait(&k_full[st], (g / kKVStages) & 1);
        else if (!mbar_test_wait(&k_full[st], (g / kKVStages) & 1)) return false;
        tc_fence_after();
        const uint32_t ka = sbase + kOffK + st * kKVBytes;
        const uint32_t tS = tT + (s_h % kSBufs) * 64;
#pragma unroll
        for (int k = 0; k < kHD / 16; ++k)
          umma_bf16(tS, umma_smem_desc(qa + k * 32, 16, 1024), umma_smem_desc(ka + k * 32, 16, 1024), idescS, k != 0);
        umma_commit(&k_empty[st]);
        umma_commit(&s_full[kSBufs * t + (s_h % kSBufs)]);
        ++s_h;
        if (++s_b == nblk) {
          umma_commit(&q_empty[t]);       // the Q tile may be replaced by the next item's
          s_b = 0;
          ++s_nact;
          do {
            s_item += gridDim.x;
            s_g0 += nblk;
          } while (s_item < p.n_items && !active(s_item));
        }
        return true;
      };
      int h = 0, g0 = 0, n_act = 0;
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, g0 += nblk) {
        if (!active(item)) {
          for (int b = 0; b < nblk; ++b) {
            const int g = g0 + b, st = g % kKVStages;
            mbar_wait(&k_full[st], (g / kKVStages) & 1);
            mbar_arrive(&k_empty[st]);
            mbar_wait(&v_full[st], (g / kKVStages) & 1);
            mbar_arrive(&v_empty[st]);
          }
          continue;
        }
        for (int j = 0; j < nblk; ++j, ++h) {
          // S(h) and S(h+1) must be on their way before this thread sleeps on P(h): the warpgroup asks for
          // S(h+1) in the middle of block h.  Further ahead only if the operands are already there.
          while (s_h <= h + 1 && s_item < p.n_items) issue_S(true);
#ifndef VVS_EXP_NO_RUNAHEAD
          while (s_h < h + kSBufs && s_item < p.n_items && issue_S(false)) {
          }
#endif
          const int g = g0 + j, st = g % kKVStages;
          mbar_wait(&v_full[st], (g / kKVStages) & 1);   // operands first: nothing but the MMA issue follows P
          if (j == 0 && n_act > 0) mbar_wait(&o_free[t], (n_act - 1) & 1);   // previous item's O has been read
          mbar_wait(&p_full[2 * t + (h & 1)], (h >> 1) & 1);   // P_t(h) in TMEM, O_t rescaled if it had to be
          tc_fence_after();
          const uint32_t va = sbase + kOffV + st * kKVBytes;
          const uint32_t tP = tT + (h % kSBufs) * 64;
#pragma unroll
          for (int kk = 0; kk < kBKV / 16; ++kk)
            umma_bf16_ts(tO, tP + kk * 8, umma_smem_desc(va + kk * 2048, 1024, 1024),   // MN-major V: 16 keys = 2 x 1024 B
                         idescPV, (j > 0 || kk != 0) ? 1u : 0u);                        // O accumulates across key blocks
          umma_commit(&v_empty[st]);
          umma_commit(&pv_done[t]);
          // in order behind PV(h): S(h + kSBufs) overwrites S(h) / P(h)
#ifdef VVS_EXP_WAIT_PV
          mbar_wait(&pv_done[t], h & 1);
#endif
#ifndef VVS_EXP_NO_POSTPV
          if (s_h <= h + kSBufs && s_item < p.n_items) issue_S(false);
#endif
        }
        ++n_act;
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ softmax warpgroup of tile t
    const int t = warp >> 2;
    const int wq = warp & 3;                          // TMEM lane quarter of this warp
    const int row = wq * 32 + lane;                   // row inside the tile = TMEM lane
    const uint32_t lane_field = static_cast<uint32_t>(wq * 32) << 16;
    const uint32_t tT = tmem_base + t * kTileCols + lane_field;
    const uint32_t tO = tT + kOCol;
    float m_run = -INFINITY;   // reference max of the exponentials, already multiplied by log2(e)
    float l_run = 0.f;
    float m_blk = 0.f;         // row max (x log2 e) of the block held in `cur`
    const f32x2 log2e2 = f2_pack(kLog2e, kLog2e);
    int g = 0;                 // this tile's block counter (S buffer / p_full / pv_done phases)
#ifdef VVS_ATT_TRACE
    const int tr_warp = blockIdx.x != 17 ? -1 : warp == 0 ? 0 : warp == 1 ? 1 : warp == 4 ? 2 : -1;
#endif

    // one 64-key block: `cur` holds S(g) (masked, max known); the next block's scores - of this item or, on the
    // last block, of the next one - are pulled into `nxt` on the way
    auto block = [&](uint32_t (&cur)[2][32], uint32_t (&nxt)[2][32], int j, bool more, int next_valid) {
      VVS_TR(0);
      // probe for S(g+1) now and look at the answer after half a block of exponentials: an mbarrier wait costs
      // 100-250 clk even when the phase completed long ago (it did: S runs kSBufs blocks ahead)
      uint64_t* s_next = &s_full[kSBufs * t + ((g + 1) % kSBufs)];
      const uint32_t s_next_par = ((g + 1) / kSBufs) & 1;
#ifdef VVS_EXP_NO_PROBE
      const bool s_next_ready = false;
#else
      const bool s_next_ready = more && mbar_test_wait(s_next, s_next_par);
#endif
      // Lazy rescale: the reference max only moves when a block exceeds it by more than 2^8; until then
      // p = exp2(s - m_run) may exceed 1 (<= 256: exact in the fp32 sum, representable in bf16) and O / l
      // is unchanged mathematically.  Always true for j == 0 (m_run = -inf).
      const bool raise = m_blk > m_run + kRescaleLog2;
      const float alpha = raise ? ex2_approx(m_run - m_blk) : 1.0f;
      if (raise) m_run = m_blk;
      if (j >= 1 && __any_sync(0xffffffffu, raise)) {
        mbar_wait(&pv_done[t], (g - 1) & 1);          // PV(g-1) has completed: O is quiescent
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
      }
      const f32x2 neg_m2 = f2_pack(-m_run, -m_run);
      f32x2 ps01 = f2_pack(0.f, 0.f), ps23 = ps01;
      const uint32_t tP = tT + (g % kSBufs) * 64;     // P(g) overwrites columns [0, 32) of S(g)'s buffer
      uint32_t pk[8];
#if VVS_EXP_TILE_BAR == 2
      asm volatile("bar.sync %0, 128;" ::"r"(1 + t) : "memory");
#endif
      exp_pack16(&cur[0][0], log2e2, neg_m2, ps01, ps23, pk);
      tmem_st8(tP, pk);
      exp_pack16(&cur[0][16], log2e2, neg_m2, ps01, ps23, pk);
      tmem_st8(tP + 8, pk);
      VVS_TR(1);
#if VVS_EXP_TILE_BAR == 3
      asm volatile("bar.sync %0, 128;" ::"r"(1 + t) : "memory");
#endif
      if (more) {
        if (!s_next_ready) mbar_wait(s_next, s_next_par);
        tc_fence_after();
        const uint32_t tN = tT + ((g + 1) % kSBufs) * 64;
        tmem_ld32(tN, nxt[0]);
        tmem_ld32(tN + 32, nxt[1]);
      }
      VVS_TR(2);
      exp_pack16(&cur[1][0], log2e2, neg_m2, ps01, ps23, pk);
      tmem_st8(tP + 16, pk);
      float m_next = 0.f;
      if (more) {
        tmem_ld_wait();
        m_next = masked_row_max(nxt, next_valid);
      }
      exp_pack16(&cur[1][16], log2e2, neg_m2, ps01, ps23, pk);
      tmem_st8(tP + 24, pk);
      float s0, s1;
      f2_unpack(f2_add(ps01, ps23), s0, s1);
      l_run = l_run * alpha + (s0 + s1);
      m_blk = m_next;
      VVS_TR(3);
#if VVS_EXP_TILE_BAR == 4
      asm volatile("bar.sync %0, 128;" ::"r"(1 + t) : "memory");
#endif
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[2 * t + (g & 1)]);
#if VVS_EXP_TILE_BAR == 1
      asm volatile("bar.sync %0, 128;" ::"r"(1 + t) : "memory");
#endif
      VVS_TR(4);
      ++g;
    };
    // O / l of a finished item -> global (its PV(last) completed long ago when this runs one block later)
    auto epilogue = [&](const Item& it, float l_item, int g_last) {
      mbar_wait(&pv_done[t], g_last & 1);
      tc_fence_after();
      const int q = (it.qt0 + t) * kBQ + row;
      const float inv = 1.0f / l_item;
      uint4* dst = reinterpret_cast<uint4*>(p.out + static_cast<size_t>(it.row_base + q) * D + it.head * kHD);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t ov[32];
        tmem_ld32(tO + h * 32, ov);
        tmem_ld_wait();
        if (h == 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&o_free[t]);     // PV(0) of the next item may overwrite O
        }
        if (q < n_tok) {
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint4 w;
            w.x = pack_bf16x2(__uint_as_float(ov[c * 8 + 0]) * inv, __uint_as_float(ov[c * 8 + 1]) * inv);
            w.y = pack_bf16x2(__uint_as_float(ov[c * 8 + 2]) * inv, __uint_as_float(ov[c * 8 + 3]) * inv);
            w.z = pack_bf16x2(__uint_as_float(ov[c * 8 + 4]) * inv, __uint_as_float(ov[c * 8 + 5]) * inv);
            w.w = pack_bf16x2(__uint_as_float(ov[c * 8 + 6]) * inv, __uint_as_float(ov[c * 8 + 7]) * inv);
            dst[h * 4 + c] = w;
          }
        }
      }
    };

    uint32_t sa[2][32], sb[2][32];
    // first active item of this tile
    int item = blockIdx.x;
    while (item < p.n_items && t >= decode_item(p, item).n_qt) item += gridDim.x;
    bool have_prev = false;
    Item prev_it = {};
    float prev_l = 0.f;
    int prev_g = 0;
    if (item < p.n_items) {
      mbar_wait(&s_full[kSBufs * t], 0);
      tc_fence_after();
#if VVS_ATT_STAGGER > 0
      // The two tiles share each sub-partition's MUFU; started together they stay in lock step (both in their
      // exponentials, then both in their hand-offs).  Half a block of offset lets one fill the other's gaps.
      if (t == 1) {
        const long long t0 = clock64();
        while (clock64() - t0 < VVS_ATT_STAGGER) {
        }
      }
#endif
      tmem_ld32(tT, sa[0]);
      tmem_ld32(tT + 32, sa[1]);
      tmem_ld_wait();
      m_blk = masked_row_max(sa, n_tok);
    }
    // the register sets alternate per block; `flip` tracks which one holds the current block
    bool flip = false;
    while (item < p.n_items) {
      const Item it = decode_item(p, item);
      int nitem = item + gridDim.x;
      while (nitem < p.n_items && t >= decode_item(p, nitem).n_qt) nitem += gridDim.x;
      const bool has_next = nitem < p.n_items;
      m_run = -INFINITY;
      l_run = 0.f;
      for (int j = 0; j < nblk; ++j) {
        const bool last = j == nblk - 1;
        const bool more = !last || has_next;
        const int next_valid = last ? n_tok : n_tok - (j + 1) * kBKV;
        if (!flip) block(sa, sb, j, more, next_valid);
        else block(sb, sa, j, more, next_valid);
        flip = !flip;
        if (j == 0 && have_prev) {
          epilogue(prev_it, prev_l, prev_g);
          have_prev = false;
        }
      }
      if (have_prev) epilogue(prev_it, prev_l, prev_g);   // (nblk == 1 never defers twice)
      prev_it = it;
      prev_l = l_run;
      prev_g = g - 1;
      have_prev = true;
      item = nitem;
    }
    if (have_prev) epilogue(prev_it, prev_l, prev_g);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kQT * 4) tmem_dealloc<kTmemCols>(tmem_base);
}

}  // namespace

#ifdef VVS_ATT_TRACE
extern "C" __attribute__((visibility("default"))) int vvs_dbg_att_trace(long long* out) {
  return static_cast<int>(cudaMemcpyFromSymbol(out, g_att_trace, sizeof(g_att_trace)));
}
#endif

int launch_attention_ts(const bf16* qkv, bf16* out, int n_img, int n_tok, int heads, cudaStream_t stream,
                        std::string* err) {
  const int D = heads * kHD;
  CUtensorMap tmq, tmkv;
  const uint64_t rows = static_cast<uint64_t>(n_img) * n_tok;
  int rc = make_tmap_bf16(&tmq, qkv, 3ull * D, rows, 1, 3ull * D, 0, kBQ, err);
  if (rc) return rc;
  rc = make_tmap_bf16(&tmkv, qkv, 3ull * D, rows, 1, 3ull * D, 0, kBKV, err);
  if (rc) return rc;
  static bool attr_set = false;
  static int num_sms = 0;
  if (!attr_set) {
    VVS_CUDA_OK(cudaFuncSetAttribute(attention_ts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(kSmem)),
                err);
    int dev = 0;
    VVS_CUDA_OK(cudaGetDevice(&dev), err);
    VVS_CUDA_OK(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev), err);
    attr_set = true;
  }
  AttnDev p;
  p.out = out;
  p.n_tok = n_tok;
  p.D = D;
  p.heads = heads;
  p.q_tiles = (n_tok + kBQ - 1) / kBQ;
  p.n_half = (p.q_tiles + kQT - 1) / kQT;
  p.n_items = n_img * heads * p.n_half;
  const int grid = p.n_items < num_sms ? p.n_items : num_sms;
  attention_ts_kernel<<<grid, kThreads, kSmem, stream>>>(tmq, tmkv, p);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

}  // namespace vvs
