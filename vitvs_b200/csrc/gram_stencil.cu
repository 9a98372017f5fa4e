// Matching on 3x3 log-binned descriptors WITHOUT materialising them (SURVEY §8d "Gram-stencil").
//
// replaces: _log_bin (dinov2_extractor.py:265-311, hierarchy 1) + chunk_cosine_sim (vitvs_v2.py:49-56) + the two
// arg-maxes of find_correspondences_batch (vitvs_v2.py:80-81) for descriptors that are the nine clamped 3x3
// neighbours of the tokens, concatenated.  For such descriptors
//     <F_i, G_j> = sum over (dy,dx) of <f[nb(i,dy,dx)], g[nb(j,dy,dx)]>      and      |F_i|^2 = sum |f[nb(i,.)]|^2 ,
// so the [T, 9D] x [9D, T] similarity product is a 9-point stencil over the plain token Gram matrix f g^T
// ([T, D] x [D, T]: nine times fewer tensor flops) and the 9D-wide descriptors (856 MB for 64 ViT-B pairs) are
// never written.  Kernels here: tokens -> bf16 rows + squared norms; binned inverse norms; stencil + cosine scaling
// + row / column arg-max in the same packed-key format the fused similarity epilogue produces (gemm_tc.cu EPI_SIM),
// so select_kernel consumes either.  The Gram matrices (0.94 MB per pair) stay L2-resident between the two kernels.
#include <string>

#include "vvs_internal.h"

namespace vvs {
namespace {

__device__ __forceinline__ uint32_t ord_f32(float f) {     // monotone float -> uint32 (as gemm_tc.cu / match_ibvs.cu)
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// x: [n_img, 1+T, D] fp32 residual stream -> tb [n_img * T, D] bf16 (CLS dropped), n2 [n_img * T] = |f|^2 (fp32 tokens).
// One warp per token, 128-bit loads.
__global__ void __launch_bounds__(256)
tokens_bf16_sqnorm_kernel(const float* __restrict__ x, bf16* __restrict__ tb, float* __restrict__ n2, int n_img, int T,
                          int D) {
  const int lane = threadIdx.x & 31;
  const long long tok = blockIdx.x * 8ll + (threadIdx.x >> 5);
  if (tok >= static_cast<long long>(n_img) * T) return;
  const int img = static_cast<int>(tok / T), t = static_cast<int>(tok % T);
  const float4* r = reinterpret_cast<const float4*>(x + (static_cast<size_t>(img) * (T + 1) + 1 + t) * D);
  uint2* o = reinterpret_cast<uint2*>(tb + static_cast<size_t>(tok) * D);
  float s = 0.f;
  for (int i = lane; i < D / 4; i += 32) {
    const float4 v = r[i];
    s += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    o[i] = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
  }
#pragma unroll
  for (int w = 16; w > 0; w >>= 1) s += __shfl_xor_sync(0xffffffffu, s, w);
  if (lane == 0) n2[tok] = s;
}

// inv[img*T + t] = 1 / max(sqrt(sum of |f|^2 over the nine clamped neighbours), 1e-8)   (CosineSimilarity eps, vitvs_v2.py:55)
__global__ void binned_inv_norm_kernel(const float* __restrict__ n2, float* __restrict__ inv, int n_img, int hp, int wp) {
  const int T = hp * wp;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_img * T) return;
  const int img = idx / T, t = idx - img * T;
  const int y0 = t / wp, x0 = t - y0 * wp;
  float tot = 0.f;
#pragma unroll
  for (int p = 0; p < 9; ++p)
    tot += n2[static_cast<size_t>(img) * T + clampi(y0 + p / 3 - 1, 0, hp - 1) * wp + clampi(x0 + p % 3 - 1, 0, wp - 1)];
  inv[idx] = 1.0f / fmaxf(sqrtf(tot), 1e-8f);
}

constexpr int kRowsPerCta = 16;
constexpr int kStencilThreads = 256;
constexpr int kMaxColsPerThread = 8;     // T <= 2048 per launch configuration

// G: [n_pairs, T, T] fp32 token Gram matrices (rows = goal tokens, columns = current tokens).
// One CTA = kRowsPerCta goal tokens x all current tokens of one pair.
__global__ void __launch_bounds__(kStencilThreads)
gram_stencil_argmax_kernel(const float* __restrict__ G, const float* __restrict__ inv, unsigned long long* __restrict__ row_best,
                           unsigned long long* __restrict__ col_best, int n_pairs, int hp, int wp, int best_pitch) {
  const int T = hp * wp;
  const int pair = blockIdx.y;
  const int i0 = blockIdx.x * kRowsPerCta;
  const float* Gp = G + static_cast<size_t>(pair) * T * T;
  const float* inv_goal = inv + static_cast<size_t>(pair) * T;
  const float* inv_cur = inv + static_cast<size_t>(n_pairs + pair) * T;
  __shared__ unsigned long long rbest[kRowsPerCta];
  if (threadIdx.x < kRowsPerCta) rbest[threadIdx.x] = 0ull;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  for (int jc = 0; jc < kMaxColsPerThread; ++jc) {
    const int j = jc * kStencilThreads + threadIdx.x;
    if (jc * kStencilThreads >= T) break;
    const bool j_ok = j < T;
    const int jj = j_ok ? j : T - 1;
    const int jy = jj / wp, jx = jj - jy * wp;
    int cj[9];
#pragma unroll
    for (int p = 0; p < 9; ++p) cj[p] = clampi(jy + p / 3 - 1, 0, hp - 1) * wp + clampi(jx + p % 3 - 1, 0, wp - 1);
    const float ij = inv_cur[jj];
    unsigned long long cbest = 0ull;
    for (int r = 0; r < kRowsPerCta; ++r) {
      const int i = i0 + r;
      if (i >= T) break;
      const int iy = i / wp, ix = i - iy * wp;
      float acc = 0.f;
#pragma unroll
      for (int p = 0; p < 9; ++p) {
        const int ci = clampi(iy + p / 3 - 1, 0, hp - 1) * wp + clampi(ix + p % 3 - 1, 0, wp - 1);
        acc += __ldg(Gp + static_cast<size_t>(ci) * T + cj[p]);
      }
      const float s = acc * inv_goal[i] * ij;
      const uint32_t o = ord_f32(s);
      // row i: best over the columns (ties -> lowest column); column j: best over the rows (ties -> lowest row)
      unsigned long long rk = j_ok ? ((static_cast<unsigned long long>(o) << 32) | (0xFFFFFFFFu - static_cast<uint32_t>(j))) : 0ull;
#pragma unroll
      for (int w = 16; w > 0; w >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, rk, w);
        rk = other > rk ? other : rk;
      }
      if (lane == 0) atomicMax(&rbest[r], rk);
      const unsigned long long ck = (static_cast<unsigned long long>(o) << 32) | (0xFFFFFFFFu - static_cast<uint32_t>(i));
      cbest = ck > cbest ? ck : cbest;
    }
    if (j_ok) atomicMax(&col_best[static_cast<size_t>(pair) * best_pitch + j], cbest);
  }
  __syncthreads();
  if (threadIdx.x < kRowsPerCta && i0 + threadIdx.x < T)
    row_best[static_cast<size_t>(pair) * best_pitch + i0 + threadIdx.x] = rbest[threadIdx.x];
}

}  // namespace

int launch_tokens_bf16_sqnorm(const float* x, bf16* tb, float* n2, int n_img, int T, int D, cudaStream_t s,
                              std::string* err) {
  const long long toks = static_cast<long long>(n_img) * T;
  tokens_bf16_sqnorm_kernel<<<static_cast<int>((toks + 7) / 8), 256, 0, s>>>(x, tb, n2, n_img, T, D);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

int launch_binned_inv_norm(const float* n2, float* inv, int n_img, int hp, int wp, cudaStream_t s, std::string* err) {
  const int n = n_img * hp * wp;
  binned_inv_norm_kernel<<<(n + 255) / 256, 256, 0, s>>>(n2, inv, n_img, hp, wp);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

bool gram_stencil_supports(int T) { return T <= kMaxColsPerThread * kStencilThreads; }

int launch_gram_stencil_argmax(const float* G, const float* inv, unsigned long long* row_best,
                               unsigned long long* col_best, int n_pairs, int hp, int wp, int best_pitch, cudaStream_t s,
                               std::string* err) {
  const int T = hp * wp;
  if (!gram_stencil_supports(T)) {
    set_error(err, "gram stencil: too many tokens per image");
    return VVS_ERR_INVALID;
  }
  dim3 grid((T + kRowsPerCta - 1) / kRowsPerCta, n_pairs);
  gram_stencil_argmax_kernel<<<grid, kStencilThreads, 0, s>>>(G, inv, row_best, col_best, n_pairs, hp, wp, best_pitch);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

}  // namespace vvs
