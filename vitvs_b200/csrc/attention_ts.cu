// tcgen05 flash-attention forward, generations 7-9: P never leaves the tensor-core datapath; dedicated issuer warp.
// (reference semantics: dino_patch/attention.py:70-80 - non-causal SDPA, scale hd^-0.5,
// head dim 64, no mask, eval.)
//
// Input is the fused-QKV GEMM output [n_img * n_tok, 3*D] bf16 exactly as the reference's
// `qkv(x).reshape(B, N, 3, H, hd)` lays it out (q already multiplied by hd^-0.5 in that GEMM's epilogue):
// the Q / K / V tiles of (image, head) are 64-column TMA boxes of that matrix.  Output goes straight into
// the [n_img * n_tok, D] layout attn.proj reads.
//
// Generation 9 (the default, VVS_ATT_QT = 1): one CTA = ONE 128-row query tile of one (image, head), FOUR CTAs
// resident per SM (128 TMEM columns, ~49 KB of shared memory and 160 threads each).  A sequence of 485 tokens is
// only 8 key blocks long, so a CTA spends a third of its life outside its main loop - ~4500 clk until the first
// S is there (barriers, TMEM allocation, first operands from HBM) and ~5000 clk after the last P (last PV, O
// read-out, stores, teardown); measured with the per-block clock trace (tools/att_trace.py).  With four
// independent CTAs per SM those phases fall under the main loops of three others instead of one (2 tiles per CTA,
// 2 CTAs per SM: 180 us; 1 tile per CTA: 161 us), and the output leaves as one TMA store per tile from the tile's
// own Q buffer instead of 8 row-strided 16-byte stores per thread (146 us).  VVS_ATT_QT = 2 restores generation 8:
// two tiles per CTA sharing one K / V ring.
//   warps 0-3 (per tile)  softmax warpgroup (thread = query row = TMEM lane); it never issues an MMA or a TMA:
//               wait S -> TMEM load -> max / exp -> TMEM store -> arrive
//   warp 4 (4 kQT + u)    MMA issuer of tile u (generation 8).  The whole warp walks the loop so that every address
//               and descriptor lives on the uniform datapath; lane 0 issues.  The first issuer also feeds the K / V
//               ring.  When a softmax lane issued, its serial chain (poll P, poll V, 4 + 4 MMAs at ~75 clk each
//               through ELECT / R2UR waterfalls, three commits, TMA pump) cost 1500-2500 clk per block on the
//               tile's critical path; now ~600.
// Per tile and 64-key block:
//   S = Q K^T              (SS MMA, fp32 S in TMEM columns [0, 64) of the tile)
//   softmax warpgroup: S -> registers (one TMEM pass), row max, p = exp2(s log2e - m) (packed fp32x2 FMA, one
//   MUFU.EX2 per element), bf16 P written back with tcgen05.st over the first 32 columns of S
//   O += P V               (TS MMA: A = P straight from TMEM, B = V consumed MN-major from its row-major tile)
// so P costs no shared-memory store, no generic->async proxy fence and no smem operand bandwidth (an SS MMA
// with N = 64 reads 6 KB of operands per 32-clk step - over the 128 B/clk shared-memory port; the TS form reads 2 KB).
// S(j+1) is issued right behind PV(j) by the same thread: tcgen05.mma executes in issue order, which is
// what makes overwriting P(j) with S(j+1) safe, and one commit (s_full) tells the warpgroup both "S(j+1) is
// there" and "PV(j) is done" (O may be rescaled / read).  O accumulates in TMEM across blocks and is lazily
// rescaled only when the running max grows by more than 2^8;  O / l once at the end.
// TMEM: tile t owns columns [128 t, 128 t + 64) for S / P and [128 t + 64, 128 t + 128) for O.
// Keys past n_tok (the next image's rows or TMA zero fill) are masked to p = 0.
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
#ifndef VVS_ATT_QT
#define VVS_ATT_QT 1
#endif
#ifndef VVS_ATT_STAGES
#define VVS_ATT_STAGES (2 * VVS_ATT_QT)
#endif
#ifndef VVS_ATT_TMA_STORE
#define VVS_ATT_TMA_STORE 1
#endif

constexpr int kQT = VVS_ATT_QT;     // query tiles per CTA
constexpr int kSoftmaxWarps = kQT * 4;   // one warpgroup per tile
constexpr int kThreads = (kSoftmaxWarps + kQT) * 32;   // + one MMA issuer warp per tile (the first also feeds the K / V ring)
constexpr int kKVStages = VVS_ATT_STAGES;
constexpr uint32_t kQBytes = kBQ * kHD * 2;      // 16 KB
constexpr uint32_t kKVBytes = kBKV * kHD * 2;    // 8 KB: one K or V block
constexpr uint32_t kOffQ = 0, kOffK = kOffQ + kQT * kQBytes, kOffV = kOffK + kKVStages * kKVBytes,
                   kOffBar = kOffV + kKVStages * kKVBytes;
constexpr uint32_t kSmem = kOffBar + 256 + 1024;   // + barriers + alignment slack
constexpr uint32_t kTmemCols = 128 * kQT;
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

// exp2 on the FMA pipe (the MUFU pipe is the kernel's tightest resource: 4 lanes / clk / sub-partition): round to
// nearest integer with the 1.5 * 2^23 trick, degree-3 polynomial of 2^r on [-0.5, 0.5] (relative error 7.7e-5, far
// below the bf16 rounding of P), integer part added to the exponent bits.  Inputs are clamped at -125.
__device__ __forceinline__ f32x2 f2_ex2_poly(f32x2 x) {
  float x0, x1;
  f2_unpack(x, x0, x1);
  x = f2_pack(fmaxf(x0, -125.0f), fmaxf(x1, -125.0f));
  const f32x2 magic = f2_pack(12582912.0f, 12582912.0f), neg_magic = f2_pack(-12582912.0f, -12582912.0f);
  const f32x2 t = f2_add(x, magic);
  const f32x2 r = f2_fma(f2_add(t, neg_magic), f2_pack(-1.0f, -1.0f), x);
  f32x2 p = f2_fma(f2_pack(0.05508868396282196f, 0.05508868396282196f), r, f2_pack(0.24260404706001282f, 0.24260404706001282f));
  p = f2_fma(p, r, f2_pack(0.6932762265205383f, 0.6932762265205383f));
  p = f2_fma(p, r, f2_pack(0.9999289512634277f, 0.9999289512634277f));
  float p0, p1, t0, t1;
  f2_unpack(p, p0, p1);
  f2_unpack(t, t0, t1);
  return f2_pack(__uint_as_float(__float_as_uint(p0) + (__float_as_uint(t0) << 23)),
                 __uint_as_float(__float_as_uint(p1) + (__float_as_uint(t1) << 23)));
}

#ifndef VVS_ATT_POLY_MASK
#define VVS_ATT_POLY_MASK 0x00   // bit i: pair i of every 8 pairs (16 elements) takes the polynomial instead of MUFU.EX2
#endif

// 16 exponentials of a row segment: p = exp2(s * log2e - m), packed to bf16 pairs, row sum accumulated
__device__ __forceinline__ void exp_pack16(const uint32_t* s, f32x2 log2e2, f32x2 neg_m2, f32x2& sum01, f32x2& sum23,
                                           uint32_t* pk) {
#pragma unroll
  for (int i = 0; i < 16; i += 4) {
    const f32x2 a01 = f2_fma(f2_pack(__uint_as_float(s[i + 0]), __uint_as_float(s[i + 1])), log2e2, neg_m2);
    const f32x2 a23 = f2_fma(f2_pack(__uint_as_float(s[i + 2]), __uint_as_float(s[i + 3])), log2e2, neg_m2);
    const f32x2 e01 = ((VVS_ATT_POLY_MASK >> (i / 2)) & 1) ? f2_ex2_poly(a01) : f2_ex2(a01);
    const f32x2 e23 = ((VVS_ATT_POLY_MASK >> (i / 2 + 1)) & 1) ? f2_ex2_poly(a23) : f2_ex2(a23);
    sum01 = f2_add(sum01, e01);
    sum23 = f2_add(sum23, e23);
    pk[i / 2] = f2_to_bf16x2(e01);
    pk[i / 2 + 1] = f2_to_bf16x2(e23);
  }
}

#ifdef VVS_ATT_TRACE   // timing experiment: per-block clock64 stamps of a few CTAs (tools/att_trace.py)
__device__ long long g_att_trace[4 * 3 * 16 * 8];
#define VVS_TR(ev)                                                                                         \
  do {                                                                                                     \
    if (tr_slot >= 0 && lane == 0 && j < 16) g_att_trace[((tr_slot * 3 + tr_warp) * 16 + j) * 8 + (ev)] = clock64(); \
  } while (0)
#else
#define VVS_TR(ev) do { } while (0)
#endif

__global__ void __launch_bounds__(kThreads, 4 / kQT)
attention_ts_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                    const __grid_constant__ CUtensorMap tmO, bf16* __restrict__ out, int n_tok, int D,
                    unsigned long long* trace) {
#ifdef VVS_ATT_TRACE
  const long long tr_entry = clock64();
#endif
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* q_full = bars;                       // kQT
  uint64_t* k_full = q_full + kQT;               // kKVStages
  uint64_t* v_full = k_full + kKVStages;
  uint64_t* k_empty = v_full + kKVStages;
  uint64_t* v_empty = k_empty + kKVStages;
  uint64_t* s_full = v_empty + kKVStages;        // kQT: S_t(j) is in TMEM and PV_t(j-1) has completed
  uint64_t* p_full = s_full + kQT;               // kQT: P_t(j) is in TMEM, O_t rescaled if needed (4 warp arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(p_full + kQT);

  const int tid = threadIdx.x;
  // warp-uniform by construction AND provably so for the compiler (shuffle broadcast): tile index, barrier /
  // smem / TMEM addresses and MMA descriptors then live on the uniform datapath.  With a plain threadIdx.x >> 5
  // every tcgen05.mma / commit sat in an ELECT + R2UR.BROADCAST waterfall loop costing ~100 clk per instruction.
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int lane = tid & 31;
  const int qt0 = blockIdx.x * kQT;               // first query tile of this CTA
  const int head = blockIdx.y;
  // the QKV GEMM wrote the matrix in ascending row order and it is larger than L2: start with the images that
  // are still resident
  const int img = static_cast<int>(gridDim.z) - 1 - static_cast<int>(blockIdx.z);
  const int row_base = img * n_tok;               // first row of this image in the qkv matrix
  const int nblk = (n_tok + kBKV - 1) / kBKV;
  const int n_qt = min(kQT, (n_tok + kBQ - 1) / kBQ - qt0);   // active tiles (last CTA of a sequence)
  const int colQ = head * kHD, colK = D + head * kHD, colV = 2 * D + head * kHD;

  // The producer lane sets the barriers up and starts the first loads (Q tiles, the first kKVStages K / V blocks:
  // their stages are trivially free) BEFORE the TMEM allocation and the CTA-wide sync, so that the ~2000 clk
  // HBM / L2 latency of the first operands overlaps the rest of the prologue.
  const int n_early = min(nblk, kKVStages);
  if (tid == kSoftmaxWarps * 32) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
    for (int i = 0; i < kKVStages; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&k_empty[i], n_qt);    // one arrival per tile (its softmax warpgroup)
      mbar_init(&v_empty[i], n_qt);
    }
    for (int i = 0; i < kQT; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);
    }
    fence_mbar_init();
    pdl_wait();                        // qkv is the previous kernel's output
    for (int q = 0; q < n_qt; ++q) {
      mbar_expect_tx(&q_full[q], kQBytes);
      tma_load_3d(smem + kOffQ + q * kQBytes, &tmQ, &q_full[q], colQ, row_base + (qt0 + q) * kBQ, 0, kEvictNormal);
      if (q == 0) {                    // K(0) right behind Q(0): S_0(0) needs nothing else
        mbar_expect_tx(&k_full[0], kKVBytes);
        tma_load_3d(smem + kOffK, &tmKV, &k_full[0], colK, row_base, 0, kEvictLast);
      }
    }
    for (int b = 0; b < n_early; ++b) {
      if (b > 0) {
        mbar_expect_tx(&k_full[b], kKVBytes);
        tma_load_3d(smem + kOffK + b * kKVBytes, &tmKV, &k_full[b], colK, row_base + b * kBKV, 0, kEvictLast);
      }
      mbar_expect_tx(&v_full[b], kKVBytes);
      tma_load_3d(smem + kOffV + b * kKVBytes, &tmKV, &v_full[b], colV, row_base + b * kBKV, 0, kEvictLast);
    }
  }
  if (warp == 0) tmem_alloc<kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  pdl_launch_dependents();   // the prologue above may overlap the previous kernel's tail
  pdl_wait();
  trace_begin(trace);        // behind the wait, like gemm_tc.cu: the span must not include the QKV GEMM's tail

  if (warp >= kSoftmaxWarps) {
    // ---- issuer warps: MMA issuer of tile u; the first one also feeds the K / V ring.  The whole warp walks the
    // loops (every value warp-uniform => uniform datapath, no ELECT / R2UR waterfalls); only the issue instructions
    // are predicated on lane 0.  A softmax warp never issues: its timeline is wait S -> TMEM load -> max / exp ->
    // TMEM store -> arrive.  The issuer's serial chain per block is: see P -> 4 TS MMAs -> 4 SS MMAs -> one commit;
    // operand arrival is waited for right AFTER a block was issued (a free ~1000 clk), and the K / V stages are
    // released by the softmax side (below), not by extra commits here. -------------------------------------------
    const int u = warp - kSoftmaxWarps;            // the tile this warp issues for
    if (u < n_qt) {
      constexpr uint32_t idescS = umma_idesc_bf16(128, kBKV, false);
      constexpr uint32_t idescPV = umma_idesc_bf16(128, kHD, true);
      const uint32_t sbase = smem_u32(smem);
      const bool producer = (u == 0);
      int next_load = 2 * n_early;     // the prologue issued K / V blocks [0, n_early)
      // loads are issued in the order K(b) V(b) K(b+1) V(b+1) ... whenever their stage is free (never blocking)
      auto pump = [&]() {
        while (next_load < 2 * nblk) {
          const int b = next_load >> 1, st = b % kKVStages;
          const bool is_v = next_load & 1;
          uint64_t* empty = is_v ? &v_empty[st] : &k_empty[st];
          if (!mbar_test_wait(empty, ((b / kKVStages) & 1) ^ 1)) return;   // never suspends
          if (lane == 0) {
            uint64_t* full = is_v ? &v_full[st] : &k_full[st];
            mbar_expect_tx(full, kKVBytes);
            tma_load_3d(smem + (is_v ? kOffV : kOffK) + st * kKVBytes, &tmKV, full, is_v ? colV : colK,
                        row_base + b * kBKV, 0, kEvictLast);
          }
          __syncwarp();
          ++next_load;
        }
      };
      // the producer keeps the ring fed while it waits (a lagging sibling tile must never starve it of loads)
      auto wait_feed = [&](uint64_t* bar, uint32_t parity) {
        if (!producer) {
          mbar_wait(bar, parity);
          return;
        }
        uint32_t polls = 0;
        while (!mbar_try_wait(bar, parity)) {
          pump();
          if (++polls == (1u << 26)) __trap();
        }
      };
      const uint32_t tSi = tmem_base + u * 128, tOi = tSi + 64;   // lane field 0: operands address the whole tile
      const uint64_t qd = umma_smem_desc(sbase + kOffQ + u * kQBytes, 16, 1024);
      const uint64_t kd = umma_smem_desc(sbase + kOffK, 16, 1024);
      const uint64_t vd = umma_smem_desc(sbase + kOffV, 1024, 1024);   // MN-major V: 16 keys = 2 x 1024 B
      auto issue_S = [&](int b) {     // S_u(b) = Q_u K(b)^T   (lane 0)
        const uint64_t kb = kd + (((b % kKVStages) * kKVBytes) >> 4);
#pragma unroll
        for (int k = 0; k < kHD / 16; ++k) umma_bf16(tSi, qd + 2 * k, kb + 2 * k, idescS, k != 0);   // +32 B per k step
      };
      wait_feed(&q_full[u], 0);
      wait_feed(&k_full[0], 0);
      tc_fence_after();
      if (lane == 0) {
        issue_S(0);
        umma_commit(&s_full[u]);
      }
      __syncwarp();
      wait_feed(&v_full[0], 0);
      if (nblk > 1) wait_feed(&k_full[1 % kKVStages], (1 / kKVStages) & 1);
#ifdef VVS_ATT_TRACE
      const int tr8 = (blockIdx.z == 5 && blockIdx.y == 3 && blockIdx.x < 2) ? static_cast<int>(blockIdx.x)
                      : (blockIdx.z == 100 && blockIdx.y == 7 && blockIdx.x < 2) ? 2 + static_cast<int>(blockIdx.x) : -1;
#define VVS_TR8(ev)                                                                                            \
  do {                                                                                                         \
    if (tr8 >= 0 && tr8 < 4 && u == 0 && lane == 0 && j < 16) g_att_trace[((tr8 * 3 + 2) * 16 + j) * 8 + (ev)] = clock64(); \
  } while (0)
#else
#define VVS_TR8(ev) do { } while (0)
#endif
      // PV_u(j), then S_u(j+1) right behind it (tcgen05.mma executes in issue order, which is what makes overwriting
      // P(j) with S(j+1) safe); one commit tells the warpgroup both "S(j+1) is there" and "PV(j) is done"
      for (int j = 0; j < nblk; ++j) {
        mbar_wait(&p_full[u], j & 1);       // V(j) and K(j+1) have landed (waited for below, one block ahead)
        VVS_TR8(0);
        tc_fence_after();
        if (lane == 0) {
          const uint64_t vb = vd + (((j % kKVStages) * kKVBytes) >> 4);
#pragma unroll
          for (int kk = 0; kk < kBKV / 16; ++kk)
            umma_bf16_ts(tOi, tSi + kk * 8, vb + kk * (2048 >> 4), idescPV, (j > 0 || kk != 0) ? 1u : 0u);   // O accumulates
          if (j + 1 < nblk) issue_S(j + 1);
          umma_commit(&s_full[u]);                    // S(j+1) ready and PV(j) done (last block: O final)
        }
        __syncwarp();
        VVS_TR8(1);
        if (producer) pump();
        if (j + 1 < nblk) wait_feed(&v_full[(j + 1) % kKVStages], ((j + 1) / kKVStages) & 1);
        if (j + 2 < nblk) wait_feed(&k_full[(j + 2) % kKVStages], ((j + 2) / kKVStages) & 1);
        VVS_TR8(2);
      }
    }
  }

  const int t = warp >> 2;                          // query tile of this warpgroup
  if (warp < kSoftmaxWarps && t < n_qt) {
    const int wq = warp & 3;                          // TMEM lane quarter of this warp
    const int row = wq * 32 + lane;                   // row inside the tile = TMEM lane
    const uint32_t lane_field = static_cast<uint32_t>(wq * 32) << 16;
    const uint32_t tS = tmem_base + t * 128 + lane_field;
    const uint32_t tO = tS + 64;

    float m_run = -INFINITY;   // reference max of the exponentials, already multiplied by log2(e)
    float l_run = 0.f;
    const f32x2 log2e2 = f2_pack(kLog2e, kLog2e);
#ifdef VVS_ATT_TRACE
    const int tr_slot = (blockIdx.z == 5 && blockIdx.y == 3 && blockIdx.x < 2) ? static_cast<int>(blockIdx.x)
                        : (blockIdx.z == 100 && blockIdx.y == 7 && blockIdx.x < 2) ? 2 + static_cast<int>(blockIdx.x) : -1;
    const int tr_warp = warp == 0 ? 0 : warp == 1 ? 1 : -1;
    const int tr_ok = tr_warp >= 0 ? tr_slot : -1;
#undef VVS_TR
#define VVS_TR(ev)                                                                                         \
  do {                                                                                                     \
    if (tr_ok >= 0 && tr_ok < 4 && lane == 0 && j < 16) g_att_trace[((tr_ok * 3 + tr_warp) * 16 + j) * 8 + (ev)] = clock64(); \
  } while (0)
#endif

    for (int j = 0; j < nblk; ++j) {
      VVS_TR(0);
      mbar_wait(&s_full[t], j & 1);
      tc_fence_after();
      // S(j) complete => its K block is free; PV(j-1) complete => its V block is free (one arrival per tile)
      if (wq == 0 && lane == 0) {
        mbar_arrive(&k_empty[j % kKVStages]);
        if (j >= 1) mbar_arrive(&v_empty[(j - 1) % kKVStages]);
      }
#if VVS_ATT_STAGGER > 0
      // The tiles sharing an SM sub-partition's MUFU tend to run in lock step (all in their exponentials, then all
      // in their hand-offs); a one-off offset of the second tile lets one fill the other's gaps.
      if (j == 0 && t == 1) {
        const long long t0 = clock64();
        while (clock64() - t0 < VVS_ATT_STAGGER) {
        }
      }
#endif
      VVS_TR(1);
      uint32_t sc[2][32];                   // the row of S (fp32 bits)
      tmem_ld32(tS, sc[0]);
      tmem_ld32(tS + 32, sc[1]);
      tmem_ld_wait();
      VVS_TR(2);
#ifdef VVS_EXP_NOMATH   // timing experiment: MMA + hand-off chain only (wrong numerics)
      {
        uint32_t pk[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) pk[i] = sc[0][i] ^ sc[1][i + 8];
#pragma unroll
        for (int q = 0; q < 4; ++q) tmem_st8(tS + q * 8, pk);
        l_run = 1.f;
      }
#else
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
      if (j >= 1 && __any_sync(0xffffffffu, raise)) {   // PV(j-1) has completed (s_full(j)): O is quiescent
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
      // P(j) overwrites columns [0, 32) of S(j) (this thread's own lane only: its S row is in registers)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint32_t pk[8];
        exp_pack16(&sc[q >> 1][(q & 1) * 16], log2e2, neg_m2, ps01, ps23, pk);
        tmem_st8(tS + q * 8, pk);
      }
      float s0, s1;
      f2_unpack(f2_add(ps01, ps23), s0, s1);
      l_run = l_run * alpha + (s0 + s1);
#endif
      VVS_TR(3);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[t]);
      VVS_TR(4);
    }

    mbar_wait(&s_full[t], nblk & 1);      // PV(nblk - 1) has completed
    tc_fence_after();
    [[maybe_unused]] const int q = (qt0 + t) * kBQ + row;
    const float inv = 1.0f / l_run;
#if VVS_ATT_TMA_STORE
    // The tile's Q buffer is free (every S MMA has completed): the normalised bf16 rows go there in the 128-byte
    // swizzled layout of the output tensor map and leave as ONE TMA store per tile (rows past n_tok are clipped by
    // the map's per-image extent).  Per-thread global stores of a row each (32 distinct lines per instruction) kept
    // the warps ~2000 clk longer.
    {
      const uint32_t stage = smem_u32(smem + kOffQ + t * kQBytes) + static_cast<uint32_t>(row) * 128u;
      const uint32_t sw = static_cast<uint32_t>(row & 7);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t ov[32];
        tmem_ld32(tO + h * 32, ov);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 4; ++g)
          st_shared_v4(stage + ((static_cast<uint32_t>(h * 4 + g) ^ sw) << 4),
                       pack_bf16x2(__uint_as_float(ov[g * 8 + 0]) * inv, __uint_as_float(ov[g * 8 + 1]) * inv),
                       pack_bf16x2(__uint_as_float(ov[g * 8 + 2]) * inv, __uint_as_float(ov[g * 8 + 3]) * inv),
                       pack_bf16x2(__uint_as_float(ov[g * 8 + 4]) * inv, __uint_as_float(ov[g * 8 + 5]) * inv),
                       pack_bf16x2(__uint_as_float(ov[g * 8 + 6]) * inv, __uint_as_float(ov[g * 8 + 7]) * inv));
      }
      fence_proxy_async_smem();
      // immediate barrier ids: a run-time id makes the compiler reserve all 16 hardware barriers per CTA
      if (kQT == 1 || t == 0) named_bar_sync<1>(128);
      else named_bar_sync<2>(128);
      if (wq == 0 && lane == 0) {
        tma_store_3d(&tmO, smem_u32(smem + kOffQ + t * kQBytes), head * kHD, (qt0 + t) * kBQ, img);
        bulk_commit();
        bulk_wait_read0();     // the buffer must outlive the store's read of it
      }
    }
#else
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
#endif
  }

#ifdef VVS_ATT_TRACE
  const long long tr_loop_end = clock64();
#endif
  trace_end(trace);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<kTmemCols>(tmem_base);
#ifdef VVS_ATT_TRACE
  {   // CTA lifetime: entry, end of this thread's role, exit (slot j = 15 of the issuer row is otherwise unused)
    const int sl = (blockIdx.z == 5 && blockIdx.y == 3 && blockIdx.x < 2) ? static_cast<int>(blockIdx.x)
                   : (blockIdx.z == 100 && blockIdx.y == 7 && blockIdx.x < 2) ? 2 + static_cast<int>(blockIdx.x) : -1;
    if (sl >= 0 && sl < 4 && tid == 0) {
      long long* o = &g_att_trace[((sl * 3 + 2) * 16 + 15) * 8];
      o[5] = tr_entry;
      o[6] = tr_loop_end;
      o[7] = clock64();
    }
  }
#endif
}

}  // namespace

#ifdef VVS_ATT_TRACE
extern "C" __attribute__((visibility("default"))) int vvs_dbg_att_trace(long long* out) {
  return static_cast<int>(cudaMemcpyFromSymbol(out, g_att_trace, sizeof(g_att_trace)));
}
#endif

int launch_attention(const bf16* qkv, bf16* out, int n_img, int n_tok, int heads, cudaStream_t stream,
                        std::string* err, bool pdl, int num_sms) {
  const int D = heads * kHD;
  CUtensorMap tmq, tmkv;
  const uint64_t rows = static_cast<uint64_t>(n_img) * n_tok;
  int rc = make_tmap_bf16(&tmq, qkv, 3ull * D, rows, 1, 3ull * D, 0, kBQ, err);
  if (rc) return rc;
  rc = make_tmap_bf16(&tmkv, qkv, 3ull * D, rows, 1, 3ull * D, 0, kBKV, err);
  if (rc) return rc;
  CUtensorMap tmo;   // [n_img][n_tok][D]: the per-image row extent clips the last tile's rows past n_tok
  rc = make_tmap_bf16(&tmo, out, static_cast<uint64_t>(D), static_cast<uint64_t>(n_tok), static_cast<uint64_t>(n_img),
                      static_cast<uint64_t>(D), static_cast<uint64_t>(n_tok) * D, kBQ, err);
  if (rc) return rc;
  // Under programmatic dependent launch the CTAs of this kernel become resident while the QKV GEMM still owns most SMs:
  // four of them pack onto every SM that frees up first, and a grid of a hundred CTAs (a single frame pair) then runs on
  // a few dozen SMs at a quarter of an SM each (measured: 8.8 -> 21 us per launch).  A small grid therefore asks for
  // enough shared memory to be alone on its SM - the CTAs spread as the GEMM's CTAs retire.
  constexpr uint32_t kSmemAlone = 116 * 1024;
  static PerDeviceOnce once;
  VVS_CUDA_OK(once.run([] {
                cudaError_t e = cudaFuncSetAttribute(attention_ts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                     static_cast<int>(kSmemAlone));
                if (e != cudaSuccess) return e;
                return cudaFuncSetAttribute(attention_ts_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                            cudaSharedmemCarveoutMaxShared);
              }),
              err);
  const int q_tiles = (n_tok + kBQ - 1) / kBQ;
  dim3 grid((q_tiles + kQT - 1) / kQT, heads, n_img);
  const bool alone = pdl && static_cast<long long>(grid.x) * grid.y * grid.z <= num_sms;
  VVS_CUDA_OK(launch_kernel(attention_ts_kernel, grid, kThreads, alone ? kSmemAlone : kSmem, stream, pdl, tmq, tmkv, tmo, out,
                            n_tok, D, tl_trace_slot), err);
  return VVS_OK;
}

}  // namespace vvs
