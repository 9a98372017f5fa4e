// tcgen05 / TMEM / TMA GEMM family for sm_100a (SURVEY §2b k1,k3,k5,k6,k7,k9).
//
//   C[b] = A[b] (M x K, K-major bf16) * B[b|0] (N x K, K-major bf16)^T, fp32 accumulation in TMEM,
//   with the consumer of C fused into the epilogue (bias / GELU / SwiGLU / LayerScale+residual /
//   patch-embed row remap + pos-embed / similarity arg-max).
//
// Default: CTA PAIRS (cluster of 2 = one TPC) compute 256 x BN tiles with tcgen05.mma.cta_group::2 - each CTA
// stages its own 128 rows of A and half of the B tile, the leader CTA issues the MMAs, both CTAs run the
// epilogue on their own 128 accumulator rows in their own TMEM.  VITVS_GEMM_CTAS=1 keeps the 1-CTA kernels.
//
// Structure (one persistent CTA per SM, 320 threads):
//   warp 0  lane 0 : TMA producer   - 128x64 A box + BNx64 B box per stage, SWIZZLE_128B, mbarrier tx
//   warp 1  lane 0 : MMA issuer     - 4 x tcgen05.mma (M128, N=BN, K16) per stage, commit -> empty[stage]
//   warps 2..9     : epilogue       - tcgen05.ld (32 lanes x 32 columns, double buffered), fused op, global store
//   TMEM: two accumulators of BN columns (double buffered: epilogue of tile i overlaps MMA of tile i+1)
// Tiles are ordered N-fastest so the CTAs running at one time share A row-panels and the whole weight
// matrix through the 126 MB L2.
//
// Epilogue protocol (every variant): the tile's column operands (bias, gamma / colsum) are staged in shared memory by
// the 256 epilogue threads between two named barriers, fetched one tile ahead; a warp hands the accumulator back to the
// MMA warp as soon as its last chunk sits in registers (fence-free arrive: the tcgen05 fences order the TMEM accesses);
// outputs leave through per-warp swizzled slabs and TMA stores / reduce-adds.
// LayerNorm folded into the GEMMs around it (EPI_*_LN_*, dino_patch/block.py:90-95): the producers of the residual
// stream (patch embedding, attn.proj, mlp.fc2) also write bf16(x) and per-64-column (sum, sum of squares) partials;
// the consumers (qkv, fc1 / w12) multiply the raw rows by W diag(gamma) and apply rstd (acc - mean colsum) + b'.
// Tile regimes: mixed clusters / pairs of 256 x BN for the batched step, single CTAs of 128 x 128 and 128 x 64 for the
// launch-bound problems of a single frame pair (same accumulation order over K: results do not depend on the regime).
#include <cuda.h>

#include <algorithm>
#include <cstdlib>
#include <string>

#include "ptx.cuh"
#include "vvs_internal.h"

namespace vvs {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kGemmThreads = 320;   // TMA warp + MMA warp + 8 epilogue warps

struct GemmDev {
  int M, N, K, batch, b_batched;
  int m_tiles, n_tiles, num_kb;
  const float* bias;
  const float* gamma;
  float* x_f32;
  bf16* out_bf16;
  int ldo;
  int q_cols;
  float q_scale;
  const float* pos;
  int T;
  unsigned long long* row_best;
  unsigned long long* col_best;
  int best_pitch;
  int b_static;   // B does not depend on the previous kernel (weights): its first boxes may be requested ahead of griddepcontrol.wait
  int pin_row0;   // EPI_RESID_F32 / EPI_PATCH_F32: rows >= pin_row0 of x carry the L2 evict-last hint
  unsigned long long* trace;   // vvs_trace_step slot (null otherwise)
  // folded LayerNorm (EPI_*_LN_*)
  float2* ln_part_out;         // producers: [M, N / 64]
  bf16* xb_out;                // producers: bf16 copy of the new residual rows (pitch ldo)
  const float2* ln_part_in;    // consumers: [M, K / 64]
  const float* colsum;         // consumers: [N]
  float ln_inv_d, ln_eps;
};

// LayerNorm-folded epilogues share the code of the epilogue they extend
__host__ __device__ constexpr int epi_base(int e) {
  return e == EPI_RESID_LN_F32 ? EPI_RESID_F32
       : e == EPI_PATCH_LN_F32 ? EPI_PATCH_F32
       : e == EPI_BIAS_LN_BF16 ? EPI_BIAS_BF16
       : e == EPI_GELU_LN_BF16 ? EPI_GELU_BF16
       : e == EPI_SWIGLU_LN_BF16 ? EPI_SWIGLU_BF16 : e;
}

// NCTA = 2: a CTA pair (cluster of 2 on one TPC) computes a 256 x BN tile with cta_group::2 MMAs; each CTA
// stages its own 128 rows of A and HALF of the B tile, so the L2 -> SM operand traffic per flop drops by a
// third against the 1-CTA 128 x 256 tile (measured: the 1-CTA kernel saturates the L2 -> SM fabric at
// ~13-15 TB/s while the tensor pipe idles).
template <int BN, int NCTA, int EPI = EPI_STORE_F32>
struct Cfg {
  static constexpr int kPair = NCTA >= 2 ? 2 : 1;   // CTAs per MMA (cta_group); NCTA = 4: two pairs sharing the B tile
  static constexpr uint32_t kABytes = BM * BK * 2;
  static constexpr uint32_t kBBytes = (BN / kPair) * BK * 2;
  static constexpr uint32_t kStageBytes = kABytes + kBBytes;
  static constexpr uint32_t kTmemCols = 2 * BN;
  // one 32-row x 128-byte slab per epilogue warp (transpose buffer of the TMA stores); the epilogues that ADD to rows they
  // load (EPI_RESID_LN_F32: the old residual values; EPI_PATCH_LN_F32: the positional rows) have two: the operand of
  // chunk i + 1 lands in one (TMA load) while chunk i is updated in place in the other
  static constexpr uint32_t kSlabsPerWarp = (EPI == EPI_RESID_LN_F32 || EPI == EPI_PATCH_LN_F32) ? 2 : 1;
  static constexpr uint32_t kSlabWarps = (BN == 64) ? 4 : 8;   // 64-wide tiles: only the warps of half 0 write slabs
  static constexpr uint32_t kStagingBytes = kSlabWarps * 4096 * kSlabsPerWarp;
  static constexpr uint32_t kBarrierBytes = 512;
  static constexpr uint32_t kRingBudget = 224 * 1024 - kStagingBytes;   // what the slabs leave to the operand ring
  // 64-wide tiles serve single-pair problems: one K loop per CTA whose pace is the round trip of a stage (TMA issue ->
  // operands landed -> MMA -> commit -> producer) divided by the stages in flight; the big tiles are tensor-bound with six
#ifndef VVS_STAGES64
#define VVS_STAGES64 8
#endif
  static constexpr int kMaxStages = (BN == 64) ? VVS_STAGES64 : 6;
  static constexpr int kStages =
      (kRingBudget / kStageBytes) < kMaxStages ? static_cast<int>(kRingBudget / kStageBytes) : kMaxStages;
  // per-tile column operands (bias / gamma / colsum of the tile's BN columns): read from shared memory by the chunk
  // loops, so that no global-load latency sits between an accumulator chunk and its arithmetic
  static constexpr uint32_t kColopBytes = 2 * BN * 4;
  // the dynamic shared-memory window is declared 1024-byte aligned (SWIZZLE_128B atoms) and the kernel traps if it is not
  static constexpr uint32_t kSmemBytes = kStages * kStageBytes + kStagingBytes + kBarrierBytes + kColopBytes;
};

// exact (erf) GELU = x * Phi(x) with Phi(x) = 1 - q (x >= 0) or q (x < 0), q = 0.5 erfc(|x| / sqrt 2).
// q = exp2(P(|x|)): P is a degree-6 fit of log2(0.5 erfc(z / sqrt 2)) on z in [0, 9] (q(9) ~ 1e-19, clamp beyond),
// evaluated in fp32: |gelu - exact| <= 6e-7 over [-12, 12] (checked against scipy erf; the bf16 output ulp at
// |x| ~ 1 is 4e-3).  1 MUFU (ex2) + 10 FMA/ALU instructions instead of erff()'s ~30.
[[maybe_unused]] __device__ __forceinline__ float gelu_erf(float x) {
  const float z = fminf(fabsf(x), 9.0f);
  float p = fmaf(3.314364715e-05f, z, -7.678688386e-04f);
  p = fmaf(p, z, 8.068593625e-03f);
  p = fmaf(p, z, -5.337729771e-02f);
  p = fmaf(p, z, -4.588138689e-01f);
  p = fmaf(p, z, -1.151179458e+00f);
  p = fmaf(p, z, -9.999967071e-01f);
  float q;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(q) : "f"(p));
  return x * (x >= 0.f ? 1.0f - q : q);
}
// the same function on a register pair: the polynomial runs on the packed fp32x2 pipe (3 issue slots per
// element instead of 6) and x * Phi(x) is written as max(x, 0) - |x| q  (x >= 0: x - x q;  x < 0: x q).
__device__ __forceinline__ void gelu_erf2(f32x2 x2, float& y0, float& y1) {
  float x0, x1;
  f2_unpack(x2, x0, x1);
  const f32x2 z = f2_pack(fminf(fabsf(x0), 9.0f), fminf(fabsf(x1), 9.0f));
  f32x2 p = f2_fma(f2_pack(3.314364715e-05f, 3.314364715e-05f), z, f2_pack(-7.678688386e-04f, -7.678688386e-04f));
  p = f2_fma(p, z, f2_pack(8.068593625e-03f, 8.068593625e-03f));
  p = f2_fma(p, z, f2_pack(-5.337729771e-02f, -5.337729771e-02f));
  p = f2_fma(p, z, f2_pack(-4.588138689e-01f, -4.588138689e-01f));
  p = f2_fma(p, z, f2_pack(-1.151179458e+00f, -1.151179458e+00f));
  p = f2_fma(p, z, f2_pack(-9.999967071e-01f, -9.999967071e-01f));
  float p0, p1, q0, q1;
  f2_unpack(p, p0, p1);
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(q0) : "f"(p0));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(q1) : "f"(p1));
  y0 = fmaf(-fabsf(x0), q0, fmaxf(x0, 0.f));
  y1 = fmaf(-fabsf(x1), q1, fmaxf(x1, 0.f));
}
__device__ __forceinline__ float silu(float x) { return __fdividef(x, 1.0f + __expf(-x)); }

// monotone float -> uint32 map (larger float => larger key)
__device__ __forceinline__ uint32_t ord_f32(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}


// NCTA = 4 ("mixed"): the grid is cut into chunks of four consecutive CTAs, launched with cluster size 2 and PREFERRED
// cluster size 4 (cudaLaunchAttributePreferredClusterDimension).  The hardware groups a chunk into one 4-CTA cluster
// where a GPC has two free TPCs side by side and into two 2-CTA clusters where it has not (GPCs with an odd number of
// TPCs: a static 4-CTA launch fits only 33 clusters = 132 of the 148 SMs).  Either way the chunk computes the same
// 512 x BN work item: as a quad its two pairs share the B tile through TMA multicast (each CTA fetches a quarter), as
// two pairs each fetches its own half.  The cluster size is read at run time (%cluster_nctaid).
template <int BN, int EPI, int NCTA>
__device__ __forceinline__ void gemm_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmB2,
                                          const CUtensorMap& tmOut, const GemmDev& p) {
  using C = Cfg<BN, NCTA, EPI>;
  constexpr int EB = epi_base(EPI);     // the epilogue whose arithmetic this one extends
  constexpr bool kLnOut = (EPI == EPI_RESID_LN_F32 || EPI == EPI_PATCH_LN_F32);   // emits xb + row statistics
  constexpr bool kLnIn = (EPI == EPI_BIAS_LN_BF16 || EPI == EPI_GELU_LN_BF16 || EPI == EPI_SWIGLU_LN_BF16);
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem_raw) & 1023u) != 0u) __trap();   // the layout below has no alignment slack
  uint8_t* staging = smem + C::kStages * C::kStageBytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(staging + C::kStagingBytes);
  uint64_t* empty = full + C::kStages;
  uint64_t* tfull = empty + C::kStages;
  uint64_t* tempty = tfull + 2;
  uint64_t* xfull = tempty + 2;         // EPI_RESID_LN_F32: [epilogue warp][slab] the old residual values have landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(xfull + 16);
  float* colop = reinterpret_cast<float*>(staging + C::kStagingBytes + C::kBarrierBytes);   // [2][BN]

  // shuffle broadcast: warp-uniform for the compiler too, so descriptors / barrier addresses stay on the uniform
  // datapath (a plain threadIdx.x >> 5 wraps every tcgen05.mma / TMA issue in an ELECT + R2UR waterfall loop)
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = (NCTA >= 2) ? cluster_ctarank() : 0u;
  const uint32_t pair_rank = cta_rank & 1u;          // 0 = leader of its pair (issues the MMAs)
  const uint32_t leader_rank = cta_rank & ~1u;       // cluster rank of this CTA's pair leader
  const bool quad = (NCTA == 4) && cluster_nctaid_x() == 4u;   // mixed mode: this chunk runs as one 4-CTA cluster
  // 128-row block of the work item this CTA owns: its position inside the chunk (quad: the cluster rank)
  const uint32_t slot = (NCTA == 4 && !quad) ? ((blockIdx.x & 2u) | cta_rank) : cta_rank;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if constexpr (NCTA == 4 || EPI == EPI_PATCH_LN_F32) tma_prefetch_desc(&tmB2);
    tma_prefetch_desc(&tmOut);
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(&full[s], 1);     // pair: only the leader's is used; its expect_tx covers both CTAs' boxes
      mbar_init(&empty[s], quad ? 2 : 1);   // two pairs sharing B: a stage is free when BOTH have consumed it
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 8 * C::kPair);   // pair: the peer's epilogue warps arrive on the leader's barrier
    }
    if constexpr (kLnOut)
      for (int a = 0; a < 16; ++a) mbar_init(&xfull[a], 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    if constexpr (NCTA >= 2) tmem_alloc_2sm<C::kTmemCols>(tmem_slot);
    else tmem_alloc<C::kTmemCols>(tmem_slot);
  }
  tc_fence_before();
  if constexpr (NCTA >= 2) cluster_sync_all();   // peer barriers initialised before any remote arrive / TMA signal
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  // everything above (barriers, TMEM, descriptor prefetch) may overlap the previous kernel's tail
  pdl_launch_dependents();

  // p.m_tiles counts (128 * NCTA)-row tiles; this CTA owns 128-row block m_tile * NCTA + cta_rank of each
  const int tiles_per_batch = p.m_tiles * p.n_tiles;
  const int total_tiles = p.batch * tiles_per_batch;
  const int first_tile = blockIdx.x / NCTA;
  const int tile_step = gridDim.x / NCTA;

  // One stage of operands (lane 0 of the producer warp).  `arm`: post the stage's byte count on its barrier (once per
  // use of the stage); want_a / want_b: which of the two boxes to request now.
  const uint32_t full0 = (NCTA >= 2) ? mapa_shared(smem_u32(&full[0]), leader_rank) : 0u;   // leader's full[] in cluster space
  auto issue_stage = [&](uint32_t stage, int b, int m_blk, int n_blk, int kb, bool arm, bool want_a, bool want_b) {
    uint8_t* sa = smem + stage * C::kStageBytes;
    if (NCTA == 4 && quad) {
      // this CTA fetches ONE quarter of the B tile and multicasts it to the CTA of the same parity in the
      // other pair: every CTA still receives its 128-row half, the L2 -> SM traffic per CTA drops by a quarter
      if (arm && pair_rank == 0) mbar_expect_tx(&full[stage], 2 * C::kStageBytes);
      const uint32_t fb = full0 + stage * 8;
      const int quarter = static_cast<int>(cta_rank >> 1);
      if (want_a) tma_load_3d_2sm(sa, &tmA, fb, kb * BK, m_blk * BM, b, kEvictNormal);
      if (want_b)
        tma_load_3d_2sm_mc(sa + C::kABytes + quarter * (BN / 4) * BK * 2, &tmB, fb, kb * BK,
                           n_blk * BN + static_cast<int>(pair_rank) * (BN / 2) + quarter * (BN / 4), p.b_batched ? b : 0,
                           static_cast<uint16_t>((1u << cta_rank) | (1u << (cta_rank ^ 2u))), kEvictLast);
    } else if constexpr (NCTA >= 2) {
      if (arm && cta_rank == 0) mbar_expect_tx(&full[stage], 2 * C::kStageBytes);
      const uint32_t fb = full0 + stage * 8;
      if (want_a) tma_load_3d_2sm(sa, &tmA, fb, kb * BK, m_blk * BM, b, kEvictNormal);
      if (want_b)
        tma_load_3d_2sm(sa + C::kABytes, NCTA == 4 ? &tmB2 : &tmB, fb, kb * BK,
                        n_blk * BN + static_cast<int>(cta_rank) * (BN / 2), p.b_batched ? b : 0, kEvictLast);
    } else {
      if (arm) mbar_expect_tx(&full[stage], C::kStageBytes);
      if (want_a) tma_load_3d(sa, &tmA, &full[stage], kb * BK, m_blk * BM, b, kEvictNormal);
      if (want_b) tma_load_3d(sa + C::kABytes, &tmB, &full[stage], kb * BK, n_blk * BN, p.b_batched ? b : 0, kEvictLast);
    }
  };
  // Programmatic dependent launch: B is a WEIGHT matrix (b_static), not the predecessor's output, so the B boxes of the
  // first stages are requested before the grid dependency resolves - their latency runs under the predecessor's tail
  // and only the A boxes wait.  (Every stage is free at this point: no empty-barrier wait.)
  int pre = 0;
  if (warp == 0 && p.b_static && first_tile < total_tiles) {
    pre = p.num_kb < C::kStages ? p.num_kb : C::kStages;
    const int b0 = first_tile / tiles_per_batch, r0 = first_tile - b0 * tiles_per_batch;
    if (lane == 0)
      for (int kb = 0; kb < pre; ++kb)
        issue_stage(static_cast<uint32_t>(kb), b0, (r0 / p.n_tiles) * NCTA + static_cast<int>(slot), r0 % p.n_tiles, kb, true,
                    false, true);
    __syncwarp();
  }
  pdl_wait();
  trace_begin(p.trace);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    // The whole warp walks the loop (stage / phase / tile stay warp-uniform values on the uniform datapath);
    // only the issue instructions themselves are predicated on lane 0.
    {
      uint32_t stage = 0, phase = 0;
      for (int tile = first_tile; tile < total_tiles; tile += tile_step) {
        const int b = tile / tiles_per_batch;
        const int r = tile - b * tiles_per_batch;
        const int m_blk = (r / p.n_tiles) * NCTA + static_cast<int>(slot);
        const int n_blk = r % p.n_tiles;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          const bool early_b = tile == first_tile && kb < pre;   // armed, B on its way
          if (lane == 0) issue_stage(stage, b, m_blk, n_blk, kb, !early_b, true, !early_b);
          __syncwarp();
          if (++stage == C::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (whole warp walks, lane 0 issues)
    if (pair_rank == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM * C::kPair, BN, false);
      uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
      // Operand descriptors: only the 14-bit start-address field of the low word changes from one MMA to the next
      // (stage and 32-byte k step; the ring ends below 256 KB, so the field never carries).  Built once and advanced by
      // one 32-bit add per operand - rebuilding them from the address cost eight dependent uniform-datapath instructions
      // per MMA, which is what paces the K loop of the small tiles (four 128 x 64 x 16 MMAs of 32 clk each per k-block).
      const uint64_t desc0 = umma_smem_desc(smem_u32(smem), 16, 1024);
      const uint32_t desc_hi = static_cast<uint32_t>(desc0 >> 32);
      const uint32_t a_lo0 = static_cast<uint32_t>(desc0), b_lo0 = a_lo0 + (C::kABytes >> 4);
      for (int tile = first_tile; tile < total_tiles; tile += tile_step) {
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_lo = a_lo0 + stage * (C::kStageBytes >> 4), b_lo = b_lo0 + stage * (C::kStageBytes >> 4);
          if (lane == 0) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t da = pack_u64(a_lo + 2 * k, desc_hi), db = pack_u64(b_lo + 2 * k, desc_hi);
              if constexpr (NCTA >= 2) umma_bf16_2sm(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
              else umma_bf16(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
            }
            if (NCTA == 4 && quad) umma_commit_2sm(&empty[stage], 0xF);   // one of the two arrivals, in all four CTAs
            else if constexpr (NCTA >= 2) umma_commit_2sm(&empty[stage], 3);   // frees the stage in both CTAs
            else umma_commit(&empty[stage]);
          }
          __syncwarp();
          if (++stage == C::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (lane == 0) {
          if constexpr (NCTA >= 2) umma_commit_2sm(&tfull[acc], static_cast<uint16_t>(3u << leader_rank));
          else umma_commit(&tfull[acc]);
        }
        __syncwarp();
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (8 warps)
    // Two warps per TMEM lane quarter; each owns half of the tile's 32-column chunks.  TMEM loads are
    // double buffered and the per-column operands (bias, gamma) are fetched ahead of the TMEM wait, so
    // neither latency sits on the critical path.  Output leaves through TMA: a thread owns one accumulator
    // ROW, so each warp writes its 32 x 128-byte slab into a private smem buffer in the SWIZZLE_128B
    // pattern and one lane issues a bulk tensor store (bf16 outputs) or a bulk tensor REDUCE-ADD (the fp32
    // residual stream: x += d is performed by the L2, the SM never loads x).  TMA clips the M / N tails.
    const int quarter = warp & 3;             // TMEM lane quarter this warp may read
    const int half = (warp - 2) >> 2;         // which half of the chunks
    const uint32_t lane_field = static_cast<uint32_t>(quarter * 32) << 16;
    constexpr int NC = BN / 32;               // chunks per tile
    // 64-wide tiles (launch-bound small problems): both chunks belong to the warps of half 0, so that a warp still owns
    // whole 64-column groups (bf16 slabs, LayerNorm partials); the warps of half 1 only keep the protocol going
    constexpr int NCH = (BN == 64) ? 2 : NC / 2;   // chunks per working warp
    const bool idle_warp = (BN == 64) && half == 1;
    const int c_begin = (BN == 64) ? 0 : half * NCH;
    const uint32_t stg_s = smem_u32(staging + (warp - 2) * 4096 * C::kSlabsPerWarp);   // 1024-byte aligned: swizzle phase = row & 7
    const uint32_t my_row_s = stg_s + lane * 128;
    uint64_t* const xbar = xfull + (warp - 2) * 2;   // EPI_RESID_LN_F32
    uint32_t xphase = 0;                             // bit s: parity the next wait on xbar[s] expects
    const uint32_t sw = static_cast<uint32_t>(lane & 7);
    uint32_t acc = 0, acc_phase = 0;
    const uint32_t tempty0 = (NCTA >= 2) ? mapa_shared(smem_u32(&tempty[0]), leader_rank) : 0u;   // leader's tempty[]
    // folded LayerNorm, consumer side: (rstd, -rstd * mean) of this thread's row of tile t from the producer's
    // per-64-column partials, summed in a fixed order (bit-identical whatever the tiling of producer and consumer)
    float ln_a_next = 1.0f, ln_c_next = 0.0f;
    auto ln_stats = [&](int t) {
      if constexpr (kLnIn) {
        ln_a_next = 1.0f;
        ln_c_next = 0.0f;
        if (t >= total_tiles) return;
        const int tb = t / tiles_per_batch, tr = t - tb * tiles_per_batch;
        const int trow = ((tr / p.n_tiles) * NCTA + static_cast<int>(slot)) * BM + quarter * 32 + lane;
        if (trow < p.M) {
          const float4* pp = reinterpret_cast<const float4*>(p.ln_part_in + (static_cast<size_t>(tb) * p.M + trow) * p.num_kb);
          float s1 = 0.f, s2 = 0.f;
#pragma unroll 4
          for (int g = 0; g < (p.num_kb >> 1); ++g) {   // K / 64 partials, two per 16-byte load
            const float4 t4 = __ldg(pp + g);
            s1 += t4.x; s2 += t4.y;
            s1 += t4.z; s2 += t4.w;
          }
          const float mean = s1 * p.ln_inv_d;
          const float var = fmaxf(fmaf(-mean, mean, s2 * p.ln_inv_d), 0.f);
          ln_a_next = rsqrtf(var + p.ln_eps);
          ln_c_next = -ln_a_next * mean;
        }
      }
    };
    ln_stats(first_tile);
    // Column operands of a tile: thread t of the 256 epilogue threads owns tile column t.  Array 0 = bias (b' for the
    // folded LayerNorm), array 1 = gamma (residual epilogues) or colsum (folded LayerNorm).  The values of tile i + 1 are
    // requested while tile i is processed (two registers), and written to shared memory between two named barriers
    // at the top of tile i + 1: nobody still reads tile i's, everybody sees tile i + 1's.
    constexpr bool kColop = (EB == EPI_BIAS_BF16 || EB == EPI_GELU_BF16 || EB == EPI_RESID_F32 || EB == EPI_PATCH_F32 ||
                             EB == EPI_SWIGLU_BF16);
    constexpr bool kColop1 = (EB == EPI_RESID_F32 || kLnIn);
    const uint32_t colop_s = smem_u32(colop);
    const int ct = (warp - 2) * 32 + lane;      // tile column this thread stages
    float co0 = 0.f, co1 = 0.f;
    auto colop_fetch = [&](int t) {
      if constexpr (kColop) {
        if (t < total_tiles && ct < BN) {
          const int col = ((t % tiles_per_batch) % p.n_tiles) * BN + ct;
          if (col < p.N) {
            co0 = __ldg(p.bias + col);
            if constexpr (kColop1) co1 = __ldg((kLnIn ? p.colsum : p.gamma) + col);
          }
        }
      }
    };
    colop_fetch(first_tile);
    for (int tile = first_tile; tile < total_tiles; tile += tile_step) {
      const int b = tile / tiles_per_batch;
      const int r = tile - b * tiles_per_batch;
      const int m_blk = (r / p.n_tiles) * NCTA + static_cast<int>(slot);
      const int n_blk = r % p.n_tiles;
      if constexpr (kColop) {
        named_bar_sync<1>(256);
        if (ct < BN) {
          colop[ct] = co0;
          if constexpr (kColop1) colop[BN + ct] = co1;
        }
        named_bar_sync<2>(256);
        colop_fetch(tile + tile_step);
      }
      const int tile_row0 = m_blk * BM + quarter * 32;      // first row of this warp's slab (inside batch b)
      const int row = tile_row0 + lane;                     // row inside batch b
      const bool row_ok = row < p.M;
      const size_t grow = static_cast<size_t>(b) * p.M + row;
      // folded LayerNorm, consumer side:  y = rstd * (acc - mean * colsum) + b' = ln_a * acc + (ln_c * colsum + b').
      // The statistics of THIS tile's rows were computed at the end of the previous tile (ln_stats below), off the
      // accumulator's critical path; here only the next tile's partials are pulled into L1.
      float ln_a = 1.0f, ln_c = 0.0f;
      if constexpr (kLnIn) {
        ln_a = ln_a_next;
        ln_c = ln_c_next;
        const int tn = tile + tile_step;
        if (tn < total_tiles) {
          const int bnx = tn / tiles_per_batch, rn = tn - bnx * tiles_per_batch;
          const int rown = ((rn / p.n_tiles) * NCTA + static_cast<int>(slot)) * BM + quarter * 32 + lane;
          if (rown < p.M && (rn / p.n_tiles) != (r / p.n_tiles)) {   // same rows as this tile: already there
            const char* q = reinterpret_cast<const char*>(p.ln_part_in + (static_cast<size_t>(bnx) * p.M + rown) * p.num_kb);
            prefetch_l1(q);
            prefetch_l1(q + p.num_kb * 8 - 1);
          }
        }
      }
      // EPI_RESID_LN_F32 adds to x in the SM: the old values of a chunk arrive by TMA in one of this warp's two slabs
      // (the very boxes the store writes back), requested one chunk ahead; chunk 0 here, ahead of the accumulator wait
      // (EPI_PATCH_LN_F32: the same machinery fetches the positional rows - tmB2 is the map of pos', extended
      // periodically by 31 rows so that a slab that runs over the end of an image is still one box)
      [[maybe_unused]] const CUtensorMap* const xmap = (EPI == EPI_PATCH_LN_F32) ? &tmB2 : &tmOut;
      [[maybe_unused]] const int xrow0 = (EPI == EPI_PATCH_LN_F32) ? tile_row0 % (p.T + 1) : tile_row0;
      if constexpr (kLnOut) {
        const int col_first = n_blk * BN + c_begin * 32;
        if (lane == 0 && col_first < p.N && !idle_warp) {
          bulk_wait_read1();                 // slab 0 was last read by the store of the previous tile's chunk NCH - 2
          mbar_expect_tx(&xbar[0], 4096);
          tma_load_2d_s(stg_s, xmap, &xbar[0], col_first, xrow0, kEvictNormal);
        }
      }
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + acc * BN + lane_field;
      // The accumulator goes back to the MMA warp as soon as this warp's LAST chunk of it sits in registers - a chunk's
      // arithmetic and stores earlier than the end of the tile.  tcgen05.wait::ld is warp-collective, so every lane's
      // reads are complete; the arrive carries no memory fence (the tcgen05 fences order the TMEM accesses, and a
      // release at cluster scope would wait for every global store of the epilogue).
      auto release_acc = [&]() {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if constexpr (NCTA >= 2) mbar_arrive_cluster_nofence(tempty0 + acc * 8);
          else mbar_arrive(&tempty[acc]);
        }
      };

      if constexpr (EB == EPI_SWIGLU_BF16 && BN != 256) {
        __trap();   // never launched: the launcher only pairs the SwiGLU epilogue with 256-wide tiles
      } else if (idle_warp) {
        release_acc();
      } else if constexpr (EB == EPI_SWIGLU_BF16) {
        // tile columns [0,BN/2) = x1 slice, [BN/2,BN) = matching x2 slice (rows of w12 interleaved on load);
        // N % BN == 0 is enforced by the launcher, so every tile is complete.
        constexpr int NP = NC / 4;              // (x1, x2) chunk pairs per warp: 2 -> one 64-column slab
        if (lane == 0) bulk_wait_read0();       // the previous tile's store has finished reading the buffer
        __syncwarp();
#pragma unroll
        for (int i = 0; i < NP; ++i) {
          const int c = half * NP + i;
          uint32_t r1[32], r2[32];
          tmem_ld32(t_acc + c * 32, r1);
          tmem_ld32(t_acc + BN / 2 + c * 32, r2);
          tmem_ld_wait();
          if (i == NP - 1) release_acc();
          float v[32];
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            float4 b1 = ld_shared_v4f(colop_s + (c * 32 + q * 4) * 4);
            float4 b2 = ld_shared_v4f(colop_s + (BN / 2 + c * 32 + q * 4) * 4);
            if constexpr (kLnIn) {
              const float4 c1 = ld_shared_v4f(colop_s + (BN + c * 32 + q * 4) * 4);
              const float4 c2 = ld_shared_v4f(colop_s + (BN + BN / 2 + c * 32 + q * 4) * 4);
              b1 = make_float4(fmaf(ln_c, c1.x, b1.x), fmaf(ln_c, c1.y, b1.y), fmaf(ln_c, c1.z, b1.z), fmaf(ln_c, c1.w, b1.w));
              b2 = make_float4(fmaf(ln_c, c2.x, b2.x), fmaf(ln_c, c2.y, b2.y), fmaf(ln_c, c2.z, b2.z), fmaf(ln_c, c2.w, b2.w));
            }
            v[q * 4 + 0] = silu(fmaf(ln_a, __uint_as_float(r1[q * 4 + 0]), b1.x)) * fmaf(ln_a, __uint_as_float(r2[q * 4 + 0]), b2.x);
            v[q * 4 + 1] = silu(fmaf(ln_a, __uint_as_float(r1[q * 4 + 1]), b1.y)) * fmaf(ln_a, __uint_as_float(r2[q * 4 + 1]), b2.y);
            v[q * 4 + 2] = silu(fmaf(ln_a, __uint_as_float(r1[q * 4 + 2]), b1.z)) * fmaf(ln_a, __uint_as_float(r2[q * 4 + 2]), b2.z);
            v[q * 4 + 3] = silu(fmaf(ln_a, __uint_as_float(r1[q * 4 + 3]), b1.w)) * fmaf(ln_a, __uint_as_float(r2[q * 4 + 3]), b2.w);
          }
#pragma unroll
          for (int q = 0; q < 4; ++q)
            st_shared_v4(my_row_s + ((static_cast<uint32_t>(i * 4 + q) ^ sw) << 4), pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]),
                         pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]), pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]),
                         pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]));
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0 && tile_row0 < p.M) {
          tma_store_2d(&tmOut, stg_s, n_blk * (BN / 2) + half * NP * 32, tile_row0);
          bulk_commit();
        }
      } else {
        // EPI_SIM state: running row best over this warp's columns; ties -> lowest column
        float best_v = -INFINITY;
        int best_j = -1;                        // index inside this warp's NCH * 32 columns
        // EPI_PATCH_F32: the positional row of this GEMM row (row r of image i is token r of the residual stream)
        const float* pos_row = nullptr;
        if constexpr (EB == EPI_PATCH_F32) pos_row = p.pos + static_cast<size_t>(row_ok ? row % (p.T + 1) : 0) * p.ldo;
        // kLnOut: running (sum, sum of squares) of this thread's row over the current 64-column group, two packed lanes
        f32x2 st1a = 0ull, st1b = 0ull, st2a = 0ull, st2b = 0ull;
        uint32_t raw[2][32];
        tmem_ld32(t_acc + c_begin * 32, raw[0]);
        // folded LayerNorm, consumer side, 256-wide tiles: the NEXT tile's row statistics are gathered while this tile is
        // processed - chunk i requests a third of the partials, chunk i + 1 adds them up (same order as ln_stats), so
        // that no load latency is left at either end of the tile
        constexpr bool kStreamStats = kLnIn && NCH >= 4;
        const float4* spp = nullptr;
        float4 sp[4];
        float sx1 = 0.f, sx2 = 0.f;
        const int sG = p.num_kb >> 1, sper = (sG + NCH - 2) / (NCH - 1);
        if constexpr (kStreamStats) {
          const int tn = tile + tile_step;
          if (tn < total_tiles) {
            const int tb = tn / tiles_per_batch, tr = tn - tb * tiles_per_batch;
            const int trow = ((tr / p.n_tiles) * NCTA + static_cast<int>(slot)) * BM + quarter * 32 + lane;
            if (trow < p.M)
              spp = reinterpret_cast<const float4*>(p.ln_part_in + (static_cast<size_t>(tb) * p.M + trow) * p.num_kb);
          }
        }
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
          const int c = c_begin + i;
          const int col0 = n_blk * BN + c * 32;
          const bool live = col0 < p.N;         // warp-uniform (ragged last N tile)
          if constexpr (kStreamStats) {
            if (spp != nullptr) {
              if (i >= 1) {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  if (k < sper && (i - 1) * sper + k < sG) {
                    sx1 += sp[k].x; sx2 += sp[k].y;
                    sx1 += sp[k].z; sx2 += sp[k].w;
                  }
              }
              if (i < NCH - 1) {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  if (k < sper && i * sper + k < sG) sp[k] = __ldg(spp + i * sper + k);
              }
            }
          }
          // per-column operands, requested before the TMEM wait so their latency overlaps it
          float4 bq[8], gq[8];
          if constexpr (EB == EPI_BIAS_BF16 || EB == EPI_GELU_BF16 || EB == EPI_RESID_F32 || EB == EPI_PATCH_F32) {
            if (live) {
#pragma unroll
              for (int q = 0; q < 8; ++q) bq[q] = ld_shared_v4f(colop_s + (c * 32 + q * 4) * 4);
            }
          }
          if constexpr (EB == EPI_RESID_F32) {
            if (live) {
#pragma unroll
              for (int q = 0; q < 8; ++q) gq[q] = ld_shared_v4f(colop_s + (BN + c * 32 + q * 4) * 4);
            }
          }
          if constexpr (kLnIn) {
            // the column term of the folded LayerNorm, ln_c * colsum(W') + b', replaces the bias before the accumulator
            // arrives: one live operand array and one FMA per element, as in the plain epilogue
            if (live) {
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                const float4 cs = ld_shared_v4f(colop_s + (BN + c * 32 + q * 4) * 4);
                bq[q] = make_float4(fmaf(ln_c, cs.x, bq[q].x), fmaf(ln_c, cs.y, bq[q].y), fmaf(ln_c, cs.z, bq[q].z),
                                    fmaf(ln_c, cs.w, bq[q].w));
              }
            }
          }
          if constexpr (kLnOut) {
            if (lane == 0 && i + 1 < NCH && col0 + 32 < p.N) {   // the next chunk's old values -> the other slab
              bulk_wait_read0();             // ... which the previous chunk's store has finished reading
              mbar_expect_tx(&xbar[(i + 1) & 1], 4096);
              tma_load_2d_s(stg_s + ((i + 1) & 1) * 4096, xmap, &xbar[(i + 1) & 1], col0 + 32, xrow0, kEvictNormal);
            }
          }
          tmem_ld_wait();
          if (i + 1 < NCH) tmem_ld32(t_acc + (c + 1) * 32, raw[(i + 1) & 1]);
          else release_acc();
          const uint32_t(&rw)[32] = raw[i & 1];
          if (!live) continue;

          if constexpr (EB == EPI_SIM) {
            // Row best: a thread owns one accumulator row, so it scans its 32 columns in increasing order (strict >:
            // ties keep the lowest column).  Column best: the warp's 32 x 32 chunk is transposed through its private
            // swizzled smem slab (8 conflict-free 16-byte stores per lane, 32 conflict-free word loads), then lane L
            // scans column col0 + L over the 32 rows in increasing order (ties keep the lowest row) and raises ONE
            // packed key.  All 32-bit work: ~8 instructions per accumulator element less than a 64-bit shuffle butterfly.
            const bool chunk_full = (tile_row0 + 32 <= p.M) && (col0 + 32 <= p.N);   // warp-uniform
            if (chunk_full) {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const float v = __uint_as_float(rw[j]);
                if (v > best_v) {
                  best_v = v;
                  best_j = i * 32 + j;           // chunk-local: a compile-time immediate
                }
              }
#pragma unroll
              for (int q = 0; q < 8; ++q)
                st_shared_v4(my_row_s + ((static_cast<uint32_t>(q) ^ sw) << 4), rw[q * 4 + 0], rw[q * 4 + 1], rw[q * 4 + 2],
                             rw[q * 4 + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const float v = __uint_as_float(rw[j]);
                if ((col0 + j) < p.N && v > best_v) {
                  best_v = v;
                  best_j = i * 32 + j;
                }
              }
              // rows past M hold zero accumulators (TMA zero fill): they must never win a column
#pragma unroll
              for (int q = 0; q < 8; ++q)
                st_shared_v4(my_row_s + ((static_cast<uint32_t>(q) ^ sw) << 4), row_ok ? rw[q * 4 + 0] : 0xff800000u,
                             row_ok ? rw[q * 4 + 1] : 0xff800000u, row_ok ? rw[q * 4 + 2] : 0xff800000u,
                             row_ok ? rw[q * 4 + 3] : 0xff800000u);
            }
            __syncwarp();
            float cb = -INFINITY;
            int cr = 0;
            const uint32_t col_s = stg_s + ((static_cast<uint32_t>(lane) & 3u) << 2);
            const uint32_t unit = static_cast<uint32_t>(lane) >> 2;
#pragma unroll
            for (int rr = 0; rr < 32; ++rr) {
              const float v = __uint_as_float(ld_shared_u32(col_s + rr * 128 + ((unit ^ static_cast<uint32_t>(rr & 7)) << 4)));
              if (v > cb) {
                cb = v;
                cr = rr;
              }
            }
            __syncwarp();                         // the slab is rewritten by the next chunk
            if (col0 + lane < p.N && cb > -INFINITY)
              atomicMax(&p.col_best[static_cast<size_t>(b) * p.best_pitch + col0 + lane],
                        (static_cast<unsigned long long>(ord_f32(cb)) << 32) |
                            (0xFFFFFFFFu - static_cast<uint32_t>(tile_row0 + cr)));
          } else if constexpr (EB == EPI_BIAS_BF16 || EB == EPI_GELU_BF16) {
            // two 32-column chunks (64 bf16 = 128 bytes per row) form one slab
            if ((i & 1) == 0) {
              if (lane == 0) bulk_wait_read0();   // the previous slab's store has finished reading the buffer
              __syncwarp();
            }
            uint32_t pk[16];
            // pre-activation of a column pair: acc + b, or with the folded LayerNorm ln_a * acc + (ln_c * colsum + b')
            const f32x2 la2 = f2_pack(ln_a, ln_a);
            auto pre = [&](uint32_t r0, uint32_t r1, float b0, float b1) -> f32x2 {
              const f32x2 a = f2_pack(__uint_as_float(r0), __uint_as_float(r1));
              if constexpr (kLnIn) return f2_fma(la2, a, f2_pack(b0, b1));
              else return f2_add(a, f2_pack(b0, b1));
            };
            if constexpr (EB == EPI_BIAS_BF16) {
              const float s = (col0 < p.q_cols) ? p.q_scale : 1.0f;
              const f32x2 s2 = f2_pack(s, s);
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                float a0, a1, a2, a3;
                f2_unpack(f2_mul(pre(rw[q * 4 + 0], rw[q * 4 + 1], bq[q].x, bq[q].y), s2), a0, a1);
                f2_unpack(f2_mul(pre(rw[q * 4 + 2], rw[q * 4 + 3], bq[q].z, bq[q].w), s2), a2, a3);
                pk[q * 2 + 0] = pack_bf16x2(a0, a1);
                pk[q * 2 + 1] = pack_bf16x2(a2, a3);
              }
            } else {
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                float a0, a1, a2, a3;
                gelu_erf2(pre(rw[q * 4 + 0], rw[q * 4 + 1], bq[q].x, bq[q].y), a0, a1);
                gelu_erf2(pre(rw[q * 4 + 2], rw[q * 4 + 3], bq[q].z, bq[q].w), a2, a3);
                pk[q * 2 + 0] = pack_bf16x2(a0, a1);
                pk[q * 2 + 1] = pack_bf16x2(a2, a3);
              }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q)
              st_shared_v4(my_row_s + ((static_cast<uint32_t>((i & 1) * 4 + q) ^ sw) << 4), pk[q * 4 + 0], pk[q * 4 + 1],
                           pk[q * 4 + 2], pk[q * 4 + 3]);
            const bool slab_done = (i & 1) || (i + 1 == NCH) || (col0 + 32 >= p.N);
            if (slab_done) {
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0 && tile_row0 < p.M) {
                tma_store_2d(&tmOut, stg_s, col0 - (i & 1) * 32, tile_row0);   // columns >= N are clipped
                bulk_commit();
              }
            }
          } else if constexpr (kLnOut) {
            // The new residual values in registers: x + gamma * (acc + bias) with the old x loaded above (the very
            // arithmetic of the reduce-add path: one rounding per operation), or acc + bias + pos' (patch embed).
            // They leave three ways: the fp32 slab by TMA store, the bf16 copy the next GEMM multiplies by two 256-bit
            // stores of this thread's row, and the (sum, sum of squares) of every 64-column group for the consumer.
            const uint32_t slab_s = stg_s + (i & 1) * 4096;
            const uint32_t slab_row_s = slab_s + lane * 128;
            mbar_wait(&xbar[i & 1], (xphase >> (i & 1)) & 1u);   // the old values / positional rows of this chunk have landed
            xphase ^= 1u << (i & 1);
            if ((col0 & 63) == 0) st1a = st1b = st2a = st2b = 0ull;   // a new 64-column group starts
            uint32_t pk[8];
            // four columns at a time: a thread reads and rewrites only its own row of the slab (no cross-lane hazard)
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              float v0, v1, v2, v3;
              const uint32_t cell = slab_row_s + ((static_cast<uint32_t>(q) ^ sw) << 4);
              if constexpr (EPI == EPI_RESID_LN_F32) {
                const float4 xo = ld_shared_v4f(cell);
                v0 = __fadd_rn(xo.x, __fmul_rn(gq[q].x, __fadd_rn(__uint_as_float(rw[q * 4 + 0]), bq[q].x)));
                v1 = __fadd_rn(xo.y, __fmul_rn(gq[q].y, __fadd_rn(__uint_as_float(rw[q * 4 + 1]), bq[q].y)));
                v2 = __fadd_rn(xo.z, __fmul_rn(gq[q].z, __fadd_rn(__uint_as_float(rw[q * 4 + 2]), bq[q].z)));
                v3 = __fadd_rn(xo.w, __fmul_rn(gq[q].w, __fadd_rn(__uint_as_float(rw[q * 4 + 3]), bq[q].w)));
              } else {
                const float4 ps = ld_shared_v4f(cell);
                v0 = __uint_as_float(rw[q * 4 + 0]) + bq[q].x + ps.x;
                v1 = __uint_as_float(rw[q * 4 + 1]) + bq[q].y + ps.y;
                v2 = __uint_as_float(rw[q * 4 + 2]) + bq[q].z + ps.z;
                v3 = __uint_as_float(rw[q * 4 + 3]) + bq[q].w + ps.w;
              }
              st_shared_v4f(cell, v0, v1, v2, v3);
              const f32x2 va = f2_pack(v0, v1), vb = f2_pack(v2, v3);
              st1a = f2_add(st1a, va);
              st1b = f2_add(st1b, vb);
              st2a = f2_fma(va, va, st2a);
              st2b = f2_fma(vb, vb, st2b);
              pk[(2 * q) & 7] = pack_bf16x2(v0, v1);
              pk[(2 * q + 1) & 7] = pack_bf16x2(v2, v3);
              if ((q & 3) == 3 && row_ok)   // 16 columns of the bf16 copy: one 32-byte sector of this thread's row
                st_global_v8(p.xb_out + grow * p.ldo + col0 + (q >> 2) * 16, pk[0], pk[1], pk[2], pk[3], pk[4], pk[5], pk[6], pk[7]);
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && tile_row0 < p.M) {
              if (tile_row0 >= p.pin_row0) tma_store_2d_hint(&tmOut, slab_s, col0, tile_row0, kEvictLast);
              else tma_store_2d(&tmOut, slab_s, col0, tile_row0);
              bulk_commit();
            }
            if (row_ok) {
              if ((col0 & 63) == 32) {
                float a0, a1, b0, b1, c0, c1, d0, d1;
                f2_unpack(st1a, a0, a1);
                f2_unpack(st1b, b0, b1);
                f2_unpack(st2a, c0, c1);
                f2_unpack(st2b, d0, d1);
                p.ln_part_out[grow * static_cast<size_t>(p.N >> 6) + (col0 >> 6)] =
                    make_float2((a0 + a1) + (b0 + b1), (c0 + c1) + (d0 + d1));
              }
            }
          } else if constexpr (EPI == EPI_RESID_F32) {
            // x += gamma * (acc + bias): one 32-column fp32 chunk = one 128-byte-row slab, reduce-added by TMA
            if (lane == 0) bulk_wait_read0();
            __syncwarp();
#pragma unroll
            for (int q = 0; q < 8; ++q)
              st_shared_v4f(my_row_s + ((static_cast<uint32_t>(q) ^ sw) << 4),
                            gq[q].x * (__uint_as_float(rw[q * 4 + 0]) + bq[q].x),
                            gq[q].y * (__uint_as_float(rw[q * 4 + 1]) + bq[q].y),
                            gq[q].z * (__uint_as_float(rw[q * 4 + 2]) + bq[q].z),
                            gq[q].w * (__uint_as_float(rw[q * 4 + 3]) + bq[q].w));
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && tile_row0 < p.M) {
              if (tile_row0 >= p.pin_row0) tma_reduce_add_2d_hint(&tmOut, stg_s, col0, tile_row0, kEvictLast);
              else tma_reduce_add_2d(&tmOut, stg_s, col0, tile_row0);
              bulk_commit();
            }
          } else if constexpr (EPI == EPI_PATCH_F32) {
            // x = acc + bias + pos': one 32-column fp32 chunk = one 128-byte-row slab, stored by TMA
            if (lane == 0) bulk_wait_read0();
            __syncwarp();
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const float4 ps = __ldg(reinterpret_cast<const float4*>(pos_row + col0) + q);
              st_shared_v4f(my_row_s + ((static_cast<uint32_t>(q) ^ sw) << 4),
                            __uint_as_float(rw[q * 4 + 0]) + bq[q].x + ps.x, __uint_as_float(rw[q * 4 + 1]) + bq[q].y + ps.y,
                            __uint_as_float(rw[q * 4 + 2]) + bq[q].z + ps.z, __uint_as_float(rw[q * 4 + 3]) + bq[q].w + ps.w);
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && tile_row0 < p.M) {
              if (tile_row0 >= p.pin_row0) tma_store_2d_hint(&tmOut, stg_s, col0, tile_row0, kEvictLast);
              else tma_store_2d(&tmOut, stg_s, col0, tile_row0);
              bulk_commit();
            }
          } else if constexpr (EPI == EPI_STORE_F32) {
            if (row_ok) {
              float* dst = p.x_f32 + grow * p.ldo + col0;
              if (col0 + 32 <= p.N && (p.ldo & 3) == 0) {
                float4* x4 = reinterpret_cast<float4*>(dst);
#pragma unroll
                for (int q = 0; q < 8; ++q)
                  x4[q] = make_float4(__uint_as_float(rw[q * 4]), __uint_as_float(rw[q * 4 + 1]),
                                      __uint_as_float(rw[q * 4 + 2]), __uint_as_float(rw[q * 4 + 3]));
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (col0 + j < p.N) dst[j] = __uint_as_float(rw[j]);
              }
            }
          }
        }
        if constexpr (kStreamStats) {
          ln_a_next = 1.0f;
          ln_c_next = 0.0f;
          if (spp != nullptr) {
            const float mean = sx1 * p.ln_inv_d;
            const float var = fmaxf(fmaf(-mean, mean, sx2 * p.ln_inv_d), 0.f);
            ln_a_next = rsqrtf(var + p.ln_eps);
            ln_c_next = -ln_a_next * mean;
          }
        }
        if constexpr (EB == EPI_SIM) {
          if (row_ok && best_j >= 0) {
            const unsigned long long k = (static_cast<unsigned long long>(ord_f32(best_v)) << 32) |
                                         (0xFFFFFFFFu - static_cast<uint32_t>(n_blk * BN + c_begin * 32 + best_j));
            atomicMax(&p.row_best[static_cast<size_t>(b) * p.best_pitch + row], k);
          }
        }
      }
      if constexpr (kLnIn) {
        if constexpr (EB != EPI_SWIGLU_BF16 && BN / 64 >= 4) {
          // gathered by the chunk loop above
        } else {
          ln_stats(tile + tile_step);   // the accumulator is long released: this latency is nobody's critical path
        }
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (lane == 0) bulk_wait0();   // every bulk store / reduce of this warp has landed before the CTA retires
  }

  trace_end(p.trace);
  tc_fence_before();
  if constexpr (NCTA >= 2) {
    cluster_sync_all();   // the peer may still be reading operands / signalling barriers of this CTA
    if (warp == 1) tmem_dealloc_2sm<C::kTmemCols>(tmem_base);
  } else {
    __syncthreads();
    if (warp == 1) tmem_dealloc<C::kTmemCols>(tmem_base);
  }
}

template <int BN, int EPI, int NCTA>
__global__ void __cluster_dims__(NCTA, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmAux, const __grid_constant__ CUtensorMap tmOut, const GemmDev p) {
  gemm_body<BN, EPI, NCTA>(tmA, tmB, tmAux, tmOut, p);   // tmAux: the positional rows of EPI_PATCH_LN_F32 (else unused)
}

// mixed mode: no compile-time cluster size - the launch supplies {2, preferred 4}
template <int BN, int EPI>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_mixed_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB4,
                     const __grid_constant__ CUtensorMap tmB2, const __grid_constant__ CUtensorMap tmOut, const GemmDev p) {
  gemm_body<BN, EPI, 4>(tmA, tmB4, tmB2, tmOut, p);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn(std::string* err) {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || sym == nullptr) {
    set_error(err, std::string("cudaGetDriverEntryPoint(cuTensorMapEncodeTiled) failed: ") + cudaGetErrorString(e));
    return nullptr;
  }
  fn = reinterpret_cast<EncodeTiledFn>(sym);
  return fn;
}

template <int BN, int EPI, int NCTA>
int launch_one(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmAux, const CUtensorMap& tmOut,
               const GemmDev& d, int grid, cudaStream_t stream, bool pdl, std::string* err) {
  using C = Cfg<BN, NCTA, EPI>;
  static PerDeviceOnce once;
  VVS_CUDA_OK(once.run([] {
                return cudaFuncSetAttribute(gemm_tc_kernel<BN, EPI, NCTA>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            static_cast<int>(C::kSmemBytes));
              }),
              err);
  VVS_CUDA_OK(launch_kernel(gemm_tc_kernel<BN, EPI, NCTA>, grid, kGemmThreads, C::kSmemBytes, stream, pdl, tmA, tmB, tmAux, tmOut,
                            d),
              err);
  return VVS_OK;
}

// mixed mode launch: cluster size 2, preferred cluster size 4, one chunk of four CTAs per 512-row work item
template <int EPI>
int launch_mixed(const CUtensorMap& tmA, const CUtensorMap& tmB4, const CUtensorMap& tmB2, const CUtensorMap& tmOut,
                 const GemmDev& d, int grid, cudaStream_t stream, bool pdl, std::string* err) {
  using C = Cfg<256, 4, EPI>;
  static PerDeviceOnce once;
  VVS_CUDA_OK(once.run([] {
                cudaError_t e = cudaFuncSetAttribute(gemm_tc_mixed_kernel<256, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                     static_cast<int>(C::kSmemBytes));
                if (e != cudaSuccess) return e;
                return cudaFuncSetAttribute(gemm_tc_mixed_kernel<256, EPI>, cudaFuncAttributeNonPortableClusterSizeAllowed, 0);
              }),
              err);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = C::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute at[3];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributePreferredClusterDimension;
  at[1].val.preferredClusterDim.x = 4; at[1].val.preferredClusterDim.y = 1; at[1].val.preferredClusterDim.z = 1;
  at[2].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[2].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 3 : 2;
  VVS_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_tc_mixed_kernel<256, EPI>, tmA, tmB4, tmB2, tmOut, d), err);
  return VVS_OK;
}

template <int BN, int NCTA>
int dispatch_epi(int epi, const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmAux, const CUtensorMap& tmOut,
                 const GemmDev& d, int grid, cudaStream_t stream, bool pdl, std::string* err) {
  switch (epi) {
    case EPI_BIAS_BF16: return launch_one<BN, EPI_BIAS_BF16, NCTA>(tmA, tmB, tmAux, tmOut, d, grid, stream, pdl, err);
    case EPI_GELU_BF16: return launch_one<BN, EPI_GELU_BF16, NCTA>(tmA, tmB, tmAux, tmOut, d, grid, stream, pdl, err);
    case EPI_RESID_F32: return launch_one<BN, EPI_RESID_F32, NCTA>(tmA, tmB, tmAux, tmOut, d, grid, stream, pdl, err);
    case EPI_PATCH_F32: return launch_one<BN, EPI_PATCH_F32, NCTA>(tmA, tmB, tmAux, tmOut, d, grid, stream, pdl, err);
    case EPI_SWIGLU_BF16: return launch_one<BN, EPI_SWIGLU_BF16, NCTA>(tmA, tmB, tmAux, tmOut, d, grid, stream, pdl, err);
    case EPI_SIM: return launch_one<BN, EPI_SIM, NCTA>(tmA, tmB, tmAux, tmOut, d, grid, stream, pdl, err);
    case EPI_STORE_F32: return launch_one<BN, EPI_STORE_F32, NCTA>(tmA, tmB, tmAux, tmOut, d, grid, stream, pdl, err);
    case EPI_RESID_LN_F32: return launch_one<BN, EPI_RESID_LN_F32, NCTA>(tmA, tmB, tmAux, tmOut, d, grid, stream, pdl, err);
    case EPI_PATCH_LN_F32: return launch_one<BN, EPI_PATCH_LN_F32, NCTA>(tmA, tmB, tmAux, tmOut, d, grid, stream, pdl, err);
    case EPI_BIAS_LN_BF16: return launch_one<BN, EPI_BIAS_LN_BF16, NCTA>(tmA, tmB, tmAux, tmOut, d, grid, stream, pdl, err);
    case EPI_GELU_LN_BF16: return launch_one<BN, EPI_GELU_LN_BF16, NCTA>(tmA, tmB, tmAux, tmOut, d, grid, stream, pdl, err);
    case EPI_SWIGLU_LN_BF16: return launch_one<BN, EPI_SWIGLU_LN_BF16, NCTA>(tmA, tmB, tmAux, tmOut, d, grid, stream, pdl, err);
    default: set_error(err, "launch_gemm: unknown epilogue"); return VVS_ERR_INVALID;
  }
}

}  // namespace

int make_tmap_bf16(CUtensorMap* out, const void* base, uint64_t cols, uint64_t rows, uint64_t batch, uint64_t ld,
                   uint64_t batch_pitch, uint32_t box_rows, std::string* err) {
  EncodeTiledFn fn = get_encode_fn(err);
  if (!fn) return VVS_ERR_CUDA;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld * 2) % 16 || (batch_pitch * 2) % 16) {
    set_error(err, "make_tmap_bf16: base / pitches must be 16-byte aligned");
    return VVS_ERR_INVALID;
  }
  if (batch <= 1) batch_pitch = rows * ld;   // unused by the hardware, must still be a legal stride
  cuuint64_t dims[3] = {cols, rows, batch < 1 ? 1 : batch};
  cuuint64_t strides[2] = {ld * 2, batch_pitch * 2};
  cuuint32_t box[3] = {64, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error(err, "cuTensorMapEncodeTiled failed with CUresult " + std::to_string(static_cast<int>(r)));
    return VVS_ERR_CUDA;
  }
  return VVS_OK;
}

// row-major [rows, cols] tensor of 2- or 4-byte elements (pitch ld elements), box {box_cols, box_rows},
// SWIZZLE_128B (box_cols * elem_bytes must be 128): the epilogue's store / reduce-add target.
int make_tmap_out(CUtensorMap* out, const void* base, int elem_bytes, uint64_t cols, uint64_t rows, uint64_t ld,
                  uint32_t box_cols, uint32_t box_rows, std::string* err) {
  EncodeTiledFn fn = get_encode_fn(err);
  if (!fn) return VVS_ERR_CUDA;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld * elem_bytes) % 16) {
    set_error(err, "launch_gemm: output base / pitch must be 16-byte aligned");
    return VVS_ERR_INVALID;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * elem_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                  const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error(err, "cuTensorMapEncodeTiled (output) failed with CUresult " + std::to_string(static_cast<int>(r)));
    return VVS_ERR_CUDA;
  }
  return VVS_OK;
}

// 256-wide N tiles keep the per-stage smem operand traffic at 96 B/clk (128-wide tiles sit at the 128 B/clk
// smem limit); fall back to 128 only when padding N up to a multiple of 256 would waste more than 15 %.
int gemm_block_n(int N) {
  if (N <= 128) return 128;
  const long long padded = (static_cast<long long>(N) + 255) / 256 * 256;
  return (padded * 100 <= static_cast<long long>(N) * 115) ? 256 : 128;
}

int launch_gemm(const GemmArgs& a, int num_sms, cudaStream_t stream, std::string* err) {
  if (a.M <= 0 || a.N <= 0 || a.K <= 0 || a.batch <= 0) {
    set_error(err, "launch_gemm: empty problem");
    return VVS_ERR_INVALID;
  }
  if (a.K % 8 || a.lda % 8 || a.ldb % 8) {
    set_error(err, "launch_gemm: K, lda, ldb must be multiples of 8 (16-byte TMA rows)");
    return VVS_ERR_INVALID;
  }
  const int eb = epi_base(a.epilogue);
  const bool ln_out = a.epilogue == EPI_RESID_LN_F32 || a.epilogue == EPI_PATCH_LN_F32;
  const bool ln_in = a.epilogue == EPI_BIAS_LN_BF16 || a.epilogue == EPI_GELU_LN_BF16 || a.epilogue == EPI_SWIGLU_LN_BF16;
  if (ln_out && (!a.ln_part_out || !a.xb_out || a.N % 64 || a.ldo % 16)) {
    set_error(err, "launch_gemm: folded-LayerNorm producer needs ln_part_out, xb_out, N % 64 == 0 and ldo % 16 == 0");
    return VVS_ERR_INVALID;
  }
  if (ln_in && (!a.ln_part_in || !a.colsum || a.K % 128)) {
    set_error(err, "launch_gemm: folded-LayerNorm consumer needs ln_part_in, colsum and K % 128 == 0");
    return VVS_ERR_INVALID;
  }
  const bool masked_n = (a.epilogue == EPI_SIM || a.epilogue == EPI_STORE_F32);
  if (!masked_n && (a.N % 32 || a.ldo % 8)) {
    set_error(err, "launch_gemm: N must be a multiple of 32 and ldo of 8 for this epilogue");
    return VVS_ERR_INVALID;
  }
  int bn = gemm_block_n(a.N);
  // Default (VITVS_GEMM_CTAS unset or 4): big single-problem GEMMs with 256-wide N tiles run in MIXED mode - chunks of four
  // CTAs, each a multicast quad where the hardware can place one, two pairs where it cannot (measured on B200, ViT-B 64
  // pairs, two interleaved A/B rounds: step 13.24 -> 13.01 ms, QKV -3 %, fc1 -2.5 %) - everything else as CTA pairs.
  // VITVS_GEMM_CTAS=2: pairs everywhere; =1: single-CTA kernels (A/B experiments).
  static const int ncta_env = []() {
    const char* e = getenv("VITVS_GEMM_CTAS");
    return (e && e[0] == '1') ? 1 : (e && e[0] == '2') ? 2 : 4;
  }();
  // two-pair clusters only where they pay: 256-wide tiles, one problem, at least a few 512-row cluster tiles
  const bool mixed_epi = eb == EPI_BIAS_BF16 || eb == EPI_GELU_BF16 || eb == EPI_RESID_F32 || eb == EPI_SWIGLU_BF16;
  int ncta = (ncta_env == 4 && !(mixed_epi && bn == 256 && a.batch == 1 && a.M >= 4096)) ? 2 : ncta_env;
  // Small problems (a single frame pair: M = 970 rows): 256 x 256 pair tiles would occupy a dozen of the 74 CTA-pair
  // slots and run a K loop at twice the per-tile cost.  When the pair tiling fills less than half of the machine, use
  // single-CTA 128 x 128 tiles instead: four times the tiles, half the K-loop time each (VITVS_SMALL_TILES=0: off).
  // The accumulation order over K is unchanged, so results stay bit-identical to the batched path.
  static const bool small_tiles = []() {
    const char* e = getenv("VITVS_SMALL_TILES");
    return !(e && e[0] == '0');
  }();
  if (small_tiles && ncta == 2 && eb != EPI_SWIGLU_BF16) {
    const long long pair_tiles = static_cast<long long>(a.batch) * ((a.M + 2 * BM - 1) / (2 * BM)) * ((a.N + bn - 1) / bn);
    if (pair_tiles * 4 < num_sms) {
      ncta = 1;
      bn = 128;
      // ... and when even those leave half of the SMs without a tile, 128 x 64 ones: such a kernel is one K loop per CTA,
      // bound by what one SM can pull from L2 (24 KB instead of 32 KB per k-block), and twice as many SMs pull
      // (VITVS_SMALL_TILES=1: 128-wide only)
      static const bool tiles64 = []() {
        const char* e = getenv("VITVS_SMALL_TILES");
        return !(e && e[0] == '1');
      }();
      const long long tiles128 = static_cast<long long>(a.batch) * ((a.M + BM - 1) / BM) * ((a.N + 127) / 128);
      // (measured on B200, single pair, p50: ViT-B 0.843 -> 0.808 ms; under programmatic dependent launch - once the
      // attention CTAs spread instead of packing, attention_ts.cu - ViT-B 0.769 -> 0.734, ViT-S 0.583 -> 0.538)
      if (tiles64 && tiles128 * 2 <= num_sms && eb != EPI_SWIGLU_BF16) bn = 64;   // still one tile per CTA
    }
  }
  if (eb == EPI_SWIGLU_BF16 && (bn != 256 || a.N % bn)) {
    set_error(err, "launch_gemm: SwiGLU needs N to be a multiple of the 256-wide N tile");
    return VVS_ERR_INVALID;
  }
  const bool tma_out = eb == EPI_BIAS_BF16 || eb == EPI_GELU_BF16 || eb == EPI_SWIGLU_BF16 || eb == EPI_RESID_F32 ||
                       eb == EPI_PATCH_F32;
  if (tma_out && a.batch != 1) {
    set_error(err, "launch_gemm: this epilogue stores through a 2-D tensor map (batch must be 1)");
    return VVS_ERR_INVALID;
  }
  CUtensorMap tmA, tmB;
  int rc = make_tmap_bf16(&tmA, a.A, a.K, a.M, a.batch, a.lda, a.a_batch_pitch, BM, err);
  if (rc) return rc;
  const bool b_batched = a.b_batch_pitch != 0;
  rc = make_tmap_bf16(&tmB, a.B, a.K, a.N, b_batched ? a.batch : 1, a.ldb, a.b_batch_pitch, bn / ncta, err);   // pair: half, two pairs: quarter
  if (rc) return rc;
  CUtensorMap tmOut = tmA;   // unused by the epilogues that write with plain stores
  if (eb == EPI_RESID_F32 || eb == EPI_PATCH_F32)
    rc = make_tmap_out(&tmOut, a.x_f32, 4, a.N, a.M, a.ldo, 32, 32, err);
  else if (eb == EPI_SWIGLU_BF16)
    rc = make_tmap_out(&tmOut, a.out_bf16, 2, a.N / 2, a.M, a.ldo, 64, 32, err);
  else if (tma_out)
    rc = make_tmap_out(&tmOut, a.out_bf16, 2, a.N, a.M, a.ldo, 64, 32, err);
  if (rc) return rc;
  CUtensorMap tmAux = tmB;   // unused unless ...
  if (a.epilogue == EPI_PATCH_LN_F32) {   // ... the epilogue adds positional rows it fetches by TMA
    if (!a.pos || a.pos_rows < a.T + 1 + 31) {
      set_error(err, "launch_gemm: EPI_PATCH_LN_F32 needs pos with T + 1 + 31 rows (periodic extension)");
      return VVS_ERR_INVALID;
    }
    rc = make_tmap_out(&tmAux, a.pos, 4, a.N, a.pos_rows, a.ldo, 32, 32, err);
    if (rc) return rc;
  }
  GemmDev d;
  d.M = a.M; d.N = a.N; d.K = a.K; d.batch = a.batch; d.b_batched = b_batched ? 1 : 0;
  d.b_static = (a.b_static && a.pdl) ? 1 : 0;
  d.m_tiles = (a.M + BM * ncta - 1) / (BM * ncta);
  d.n_tiles = (a.N + bn - 1) / bn;
  d.num_kb = (a.K + BK - 1) / BK;
  d.bias = a.bias; d.gamma = a.gamma; d.x_f32 = a.x_f32; d.out_bf16 = a.out_bf16; d.ldo = a.ldo;
  d.q_cols = a.q_cols; d.q_scale = a.q_scale; d.pos = a.pos; d.T = a.T;
  d.row_best = a.row_best; d.col_best = a.col_best; d.best_pitch = a.best_pitch;
  d.pin_row0 = a.pin_row0 > 0 ? a.pin_row0 : 0x7fffffff;
  d.trace = tl_trace_slot;
  d.ln_part_out = a.ln_part_out; d.xb_out = a.xb_out; d.ln_part_in = a.ln_part_in; d.colsum = a.colsum;
  d.ln_inv_d = 1.0f / static_cast<float>(a.K); d.ln_eps = a.ln_eps;
  const long long tiles = static_cast<long long>(d.batch) * d.m_tiles * d.n_tiles;
  if (ncta == 4) {
    CUtensorMap tmB2;   // pair-mode chunks fetch half of the B tile per CTA (tmB holds the quad mode's quarter boxes)
    rc = make_tmap_bf16(&tmB2, a.B, a.K, a.N, 1, a.ldb, a.b_batch_pitch, bn / 2, err);
    if (rc) return rc;
    const int grid = 4 * static_cast<int>(std::min<long long>(tiles, num_sms / 4));
    const bool pdl = a.pdl != 0;
    switch (a.epilogue) {
      case EPI_BIAS_BF16: return launch_mixed<EPI_BIAS_BF16>(tmA, tmB, tmB2, tmOut, d, grid, stream, pdl, err);
      case EPI_GELU_BF16: return launch_mixed<EPI_GELU_BF16>(tmA, tmB, tmB2, tmOut, d, grid, stream, pdl, err);
      case EPI_RESID_F32: return launch_mixed<EPI_RESID_F32>(tmA, tmB, tmB2, tmOut, d, grid, stream, pdl, err);
      case EPI_SWIGLU_BF16: return launch_mixed<EPI_SWIGLU_BF16>(tmA, tmB, tmB2, tmOut, d, grid, stream, pdl, err);
      case EPI_RESID_LN_F32: return launch_mixed<EPI_RESID_LN_F32>(tmA, tmB, tmB2, tmOut, d, grid, stream, pdl, err);
      case EPI_BIAS_LN_BF16: return launch_mixed<EPI_BIAS_LN_BF16>(tmA, tmB, tmB2, tmOut, d, grid, stream, pdl, err);
      case EPI_GELU_LN_BF16: return launch_mixed<EPI_GELU_LN_BF16>(tmA, tmB, tmB2, tmOut, d, grid, stream, pdl, err);
      case EPI_SWIGLU_LN_BF16: return launch_mixed<EPI_SWIGLU_LN_BF16>(tmA, tmB, tmB2, tmOut, d, grid, stream, pdl, err);
      default: set_error(err, "launch_gemm: epilogue not built for mixed clusters"); return VVS_ERR_INVALID;
    }
  }
  if (ncta == 2) {
    const int grid = 2 * static_cast<int>(std::min<long long>(tiles, num_sms / 2));
    if (bn == 256) return dispatch_epi<256, 2>(a.epilogue, tmA, tmB, tmAux, tmOut, d, grid, stream, a.pdl != 0, err);
    return dispatch_epi<128, 2>(a.epilogue, tmA, tmB, tmAux, tmOut, d, grid, stream, a.pdl != 0, err);
  }
  const int grid = static_cast<int>(std::min<long long>(tiles, num_sms));
  if (bn == 256) return dispatch_epi<256, 1>(a.epilogue, tmA, tmB, tmAux, tmOut, d, grid, stream, a.pdl != 0, err);
  if (bn == 64) return dispatch_epi<64, 1>(a.epilogue, tmA, tmB, tmAux, tmOut, d, grid, stream, a.pdl != 0, err);
  return dispatch_epi<128, 1>(a.epilogue, tmA, tmB, tmAux, tmOut, d, grid, stream, a.pdl != 0, err);
}

}  // namespace vvs
