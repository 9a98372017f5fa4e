// Latency-bound tail of the dino path:
//   select_kernel : packed arg-max keys -> nn1 / nn2 / sim1, same-image test, cyclic-distance candidate
//                   rule (== mutual-NN set), deterministic top-K selection, patch-centre keypoints in
//                   camera pixels (vitvs_v2.py:80-151, :511-513, :525-553)
//   ibvs_kernel   : depth lookup, normalised coordinates, interaction rows, 6x6 normal equations in
//                   fp64, v = -lambda * L^+ e, optional EMA (vitvs_v2.py:566-586, :610-623, :634-659)
#include <algorithm>
#include <string>

#include "ptx.cuh"
#include "vvs_internal.h"

namespace vvs {
namespace {

__device__ __forceinline__ float unord_f32(uint32_t k) {
  const uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  return __uint_as_float(u);
}
__device__ __forceinline__ uint32_t ord_f32(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

template <typename T, typename Op>
__device__ __forceinline__ T block_reduce(T v, Op op, T* scratch /* >= 32 */) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = op(v, __shfl_xor_sync(0xffffffffu, v, o));
  __syncthreads();   // scratch reuse across calls
  if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
  __syncthreads();
  T r = scratch[0];
  for (int i = 1; i < (blockDim.x >> 5); ++i) r = op(r, scratch[i]);
  return r;
}

// max of a 64-bit key over the warp with two pairs of redux.sync: high words first, then the low words of the lanes
// that hold the winning high word
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
  const uint32_t hi = static_cast<uint32_t>(v >> 32), lo = static_cast<uint32_t>(v);
  const uint32_t mh = __reduce_max_sync(0xffffffffu, hi);
  const uint32_t ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
  return (static_cast<unsigned long long>(mh) << 32) | ml;
}

// vitvs_v2.py:117-120 as evaluated by torch on CPU: fp32, acc = a*a; acc = fma(b,b,acc); sqrt
// (bit-exactness against PairwiseDistance is pinned by tests/test_oracle_vs_reference.py).
__device__ __forceinline__ float cyclic_distance(int dr, int dc) {
  const float a = __fadd_rn(static_cast<float>(dr), 1e-6f);
  const float b = __fadd_rn(static_cast<float>(dc), 1e-6f);
  return __fsqrt_rn(__fmaf_rn(b, b, __fmul_rn(a, a)));
}

__global__ void __launch_bounds__(1024) select_kernel(const SelectArgs a, unsigned long long* trace) {
  trace_begin(trace);
  __shared__ float s_f32[32];
  __shared__ double s_f64[32];
  __shared__ int s_i32[32];
  const int pair = blockIdx.x;
  const int T = a.T, w = a.grid_w;
  const unsigned long long* rb = a.row_best + static_cast<size_t>(pair) * a.best_pitch;
  const unsigned long long* cb = a.col_best + static_cast<size_t>(pair) * a.best_pitch;
  int32_t* nn1 = a.nn1 + static_cast<size_t>(pair) * T;
  int32_t* nn2 = a.nn2 + static_cast<size_t>(pair) * T;
  float* sim1 = a.sim1 + static_cast<size_t>(pair) * T;
  uint8_t* cand = a.candidate + static_cast<size_t>(pair) * T;

  // 1. decode packed keys (value in the high word, ~index in the low word)
  // A key that no atomicMax ever raised (still the memset 0) means the whole row / column of S was NaN - only
  // possible with NaN / Inf in caller-supplied descriptors.  torch.max would return some index and carry on, so
  // the index decodes to 0 (never out of bounds), the similarity to NaN, and flags bit 1 reports it.
  double ssum = 0.0;
  int unset = 0;
  for (int i = threadIdx.x; i < T; i += blockDim.x) {
    const unsigned long long kr = rb[i], kc = cb[i];
    const uint32_t i1 = 0xFFFFFFFFu - static_cast<uint32_t>(kr), i2 = 0xFFFFFFFFu - static_cast<uint32_t>(kc);
    nn1[i] = (kr != 0ull && i1 < static_cast<uint32_t>(T)) ? static_cast<int32_t>(i1) : 0;
    const float s = kr != 0ull ? unord_f32(static_cast<uint32_t>(kr >> 32)) : __int_as_float(0x7fc00000);
    sim1[i] = s;
    ssum += s;
    nn2[i] = (kc != 0ull && i2 < static_cast<uint32_t>(T)) ? static_cast<int32_t>(i2) : 0;
    unset |= (kr == 0ull || kc == 0ull) ? 1 : 0;
  }
  ssum = block_reduce(ssum, [](double x, double y) { return x + y; }, s_f64);
  unset = block_reduce(unset, [](int x, int y) { return x | y; }, s_i32);
  const float mean_sim = static_cast<float>(ssum / T);
  const bool same = mean_sim > a.same_thresh;          // vitvs_v2.py:84
  __syncthreads();   // nn1/nn2 of every token visible

  // 2. cyclic distances (vitvs_v2.py:105-124): d = -||ij(nn2[nn1[i]]) - ij(i) + 1e-6||
  float dmin = INFINITY, dmax = -INFINITY;
  for (int i = threadIdx.x; i < T; i += blockDim.x) {
    const int c = nn2[nn1[i]];
    const float d = -cyclic_distance(c / w - i / w, c % w - i % w);
    dmin = fminf(dmin, d);
    dmax = fmaxf(dmax, d);
  }
  dmin = block_reduce(dmin, [](float x, float y) { return fminf(x, y); }, s_f32);
  dmax = block_reduce(dmax, [](float x, float y) { return fmaxf(x, y); }, s_f32);
  const float denom = __fadd_rn(__fsub_rn(dmax, dmin), 1e-8f);   // max(d - min) + 1e-8
  int ncand = 0;
  for (int i = threadIdx.x; i < T; i += blockDim.x) {
    const int c = nn2[nn1[i]];
    const float d = -cyclic_distance(c / w - i / w, c % w - i % w);
    const bool ok = __fdiv_rn(__fsub_rn(d, dmin), denom) >= 1.0f;   // distance_threshold = 1 (:72,:130)
    cand[i] = ok ? 1 : 0;
    ncand += ok ? 1 : 0;
  }
  ncand = block_reduce(ncand, [](int x, int y) { return x + y; }, s_i32);
  __syncthreads();

  // 3. deterministic selection: K best by sim1 (ties -> lower token) among the candidates
  //    (same-image branch: among all tokens, identical points, sim = 1; vitvs_v2.py:86-101)
  const int K = a.K;
  int32_t* sel_tok = a.sel_tok + static_cast<size_t>(pair) * K;
  float* sel_sim = a.sel_sim + static_cast<size_t>(pair) * K;
  int32_t* uvs = a.uv_star + static_cast<size_t>(pair) * K * 2;
  int32_t* uvc = a.uv + static_cast<size_t>(pair) * K * 2;
  const int avail = same ? T : ncand;
  int nsel = avail < K ? avail : K;
  if (a.sel_order) {
    // reference-faithful selection (vitvs_v2.py:139-141 / :90-91): the caller supplies a priority list of goal tokens
    // (a host torch.randperm(T)); the first nsel of them that are candidates are taken, in list order - a uniform
    // random subset of the mutual-NN set in uniform random order, exactly the distribution of perm[:num_to_select]
    // applied to filtered_points; in the same-image branch it IS perm[:num_points].
    const int32_t* order = a.sel_order + static_cast<size_t>(pair) * T;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int base = 0;
    for (int i0 = 0; i0 < T && base < nsel; i0 += blockDim.x) {
      const int i = i0 + threadIdx.x;
      int tok = -1;
      bool f = false;
      if (i < T) {
        tok = order[i];
        f = tok >= 0 && tok < T && (same || cand[tok]);
      }
      const unsigned b = __ballot_sync(0xffffffffu, f);
      __syncthreads();
      if (lane == 0) s_i32[wid] = __popc(b);
      __syncthreads();
      int off = 0, tot = 0;
      for (int w2 = 0; w2 < (blockDim.x >> 5); ++w2) {
        if (w2 < wid) off += s_i32[w2];
        tot += s_i32[w2];
      }
      const int rank = base + off + __popc(b & ((1u << lane) - 1u));
      if (f && rank < nsel) {
        sel_tok[rank] = tok;
        sel_sim[rank] = same ? 1.0f : sim1[tok];
      }
      base += tot;
    }
    __syncthreads();
    nsel = base < nsel ? base : nsel;          // a list that is not a permutation may run short
    for (int k = nsel + threadIdx.x; k < K; k += blockDim.x) {
      sel_tok[k] = -1;
      sel_sim[k] = 0.f;
    }
  }
  if (!a.sel_order) {
    // K rounds of "largest key below the previous winner".  Keys (ordered sim1 << 32 | ~token: unique, ties -> lower
    // token) sit in shared memory when T * 8 bytes fit; every thread keeps the best key of its own strided tokens in a
    // register and only the round's winner rescans its tokens, so a round costs two redux pairs and ONE barrier
    // whatever T is (the alternating scratch rows make the second barrier unnecessary).
    extern __shared__ unsigned long long s_keys[];
    const bool in_smem = a.keys_in_smem != 0;
    auto key_at = [&](int i) -> unsigned long long {
      if (in_smem) return s_keys[i];
      return (same || cand[i])
                 ? (static_cast<unsigned long long>(ord_f32(sim1[i])) << 32) | (0xFFFFFFFFu - static_cast<uint32_t>(i))
                 : 0ull;
    };
    auto my_best_below = [&](unsigned long long prev) {
      unsigned long long best = 0ull;
      for (int i = threadIdx.x; i < T; i += blockDim.x) {
        const unsigned long long key = key_at(i);
        if (key < prev && key > best) best = key;
      }
      return best;
    };
    if (in_smem) {
      for (int i = threadIdx.x; i < T; i += blockDim.x)
        s_keys[i] = (same || cand[i])
                        ? (static_cast<unsigned long long>(ord_f32(sim1[i])) << 32) | (0xFFFFFFFFu - static_cast<uint32_t>(i))
                        : 0ull;
    }
    __shared__ unsigned long long s_best[2][32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    // Few tokens (the servo grid: 484): every key counts the keys above it - its rank IS its slot (keys are unique) -
    // with broadcast reads of the shared array: one barrier instead of K rounds of reduce + barrier.
    const bool by_rank = in_smem && T <= 2048;
    if (by_rank) {
      __syncthreads();
      for (int i = threadIdx.x; i < T; i += blockDim.x) {
        const unsigned long long key = s_keys[i];
        if (key == 0ull) continue;
        int rank = 0;
        for (int j = 0; j < T; ++j) rank += s_keys[j] > key ? 1 : 0;
        if (rank < nsel) {
          const int tok = static_cast<int>(0xFFFFFFFFu - static_cast<uint32_t>(key));
          sel_tok[rank] = tok;
          sel_sim[rank] = same ? 1.0f : sim1[tok];
        }
      }
    }
    unsigned long long cur = by_rank ? 0ull : my_best_below(~0ull);   // a thread only reads the keys it wrote itself: no barrier needed
    for (int k = 0; k < (by_rank ? 0 : nsel); ++k) {
      const unsigned long long wb = warp_max_u64(cur);
      if (lane == 0) s_best[k & 1][wid] = wb;
      __syncthreads();
      const unsigned long long best = warp_max_u64(lane < nwarps ? s_best[k & 1][lane] : 0ull);
      if (best == 0ull) {                              // cannot happen while k < nsel <= number of keys
        if (threadIdx.x == 0) { sel_tok[k] = -1; sel_sim[k] = 0.f; }
      } else if (cur == best) {                        // keys are unique: exactly one thread owns the winner
        const int tok = static_cast<int>(0xFFFFFFFFu - static_cast<uint32_t>(best));
        sel_tok[k] = tok;
        sel_sim[k] = same ? 1.0f : sim1[tok];
        cur = my_best_below(best);
      }
    }
    for (int k = nsel + threadIdx.x; k < K; k += blockDim.x) {
      sel_tok[k] = -1;
      sel_sim[k] = 0.f;
    }
  }
  __syncthreads();

  // 4. keypoints (vitvs_v2.py:511-513 in fp32 as torch evaluates `long * float + float`, then
  //    calculate_uv :525-553 in fp64 with round-half-even); < 4 pairs => all-zero arrays (:539-541)
  if (threadIdx.x < K) {
    const int k = threadIdx.x;
    int u_s = 0, v_s = 0, u_c = 0, v_c = 0;
    if (k < nsel && (nsel == K || nsel >= 4)) {
      const int t1 = sel_tok[k];
      const int t2 = same ? t1 : nn1[t1];
      const double scale_d = static_cast<double>(a.image_size) / static_cast<double>(w);
      const float scale = static_cast<float>(scale_d), half = static_cast<float>(scale_d / 2);
      const double sx = static_cast<double>(a.cam.width) / a.image_size;
      const double sy = static_cast<double>(a.cam.height) / a.image_size;
      const float r1 = __fadd_rn(__fmul_rn(static_cast<float>(t1 / w), scale), half);
      const float c1 = __fadd_rn(__fmul_rn(static_cast<float>(t1 % w), scale), half);
      const float r2 = __fadd_rn(__fmul_rn(static_cast<float>(t2 / w), scale), half);
      const float c2 = __fadd_rn(__fmul_rn(static_cast<float>(t2 % w), scale), half);
      u_s = static_cast<int>(rint(static_cast<double>(c1) * sx));   // flip (row,col) -> (u,v)
      v_s = static_cast<int>(rint(static_cast<double>(r1) * sy));
      u_c = static_cast<int>(rint(static_cast<double>(c2) * sx));
      v_c = static_cast<int>(rint(static_cast<double>(r2) * sy));
    }
    uvs[2 * k] = u_s;
    uvs[2 * k + 1] = v_s;
    uvc[2 * k] = u_c;
    uvc[2 * k + 1] = v_c;
  }
  if (threadIdx.x == 0) {
    a.mean_sim1[pair] = mean_sim;
    a.n_candidates[pair] = ncand;
    a.n_selected[pair] = nsel;
    a.flags[pair] = (same ? 1 : 0) | (unset ? 2 : 0);
  }
  trace_end(trace);
}

// Per-point terms of the control law with the reference's own argument lists (no depth frame involved):
//   transform_to_real_world(s_uv, s_uv_star)  (vitvs_v2.py:634-648)   uv / uv_star int32 -> xy / xy_star fp64
//   calculate_interaction_matrix(s_xy, Z)     (vitvs_v2.py:650-659)   xy, z fp64 -> L rows
// One thread per point.
__global__ void points_kernel(const int32_t* __restrict__ uv, const int32_t* __restrict__ uv_star, int n, Camera cam,
                              double* __restrict__ xy, double* __restrict__ xy_star) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (uv && xy) {
    xy[2 * i] = (uv[2 * i] - cam.cx) / cam.fx;
    xy[2 * i + 1] = (uv[2 * i + 1] - cam.cy) / cam.fy;
  }
  if (uv_star && xy_star) {
    xy_star[2 * i] = (uv_star[2 * i] - cam.cx) / cam.fx;
    xy_star[2 * i + 1] = (uv_star[2 * i + 1] - cam.cy) / cam.fy;
  }
}

__global__ void interaction_kernel(const double* __restrict__ xy, const double* __restrict__ z, int n,
                                   double* __restrict__ L) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x = xy[2 * i], y = xy[2 * i + 1], zz = z[i];
  double* r = L + static_cast<size_t>(i) * 12;
  // explicit round-to-nearest products: no FMA contraction, so every entry is the IEEE value numpy computes
  r[0] = -1.0 / zz; r[1] = 0.0; r[2] = x / zz; r[3] = __dmul_rn(x, y); r[4] = -__dadd_rn(1.0, __dmul_rn(x, x)); r[5] = y;
  r[6] = 0.0; r[7] = -1.0 / zz; r[8] = y / zz; r[9] = __dadd_rn(1.0, __dmul_rn(y, y)); r[10] = -__dmul_rn(x, y); r[11] = -x;
}

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// symmetric 6x6 pseudo-inverse solve by cyclic Jacobi (rank-deficient fall-back for np.linalg.pinv)
__device__ void pinv_solve6(double A[6][6], const double g[6], double x[6]) {
  double V[6][6];
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) V[i][j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 30; ++sweep) {
    double off = 0.0;
    for (int p = 0; p < 6; ++p)
      for (int q = p + 1; q < 6; ++q) off += A[p][q] * A[p][q];
    if (off < 1e-300) break;
    for (int p = 0; p < 6; ++p) {
      for (int q = p + 1; q < 6; ++q) {
        if (fabs(A[p][q]) < 1e-300) continue;
        const double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
        const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < 6; ++k) {
          const double akp = A[k][p], akq = A[k][q];
          A[k][p] = c * akp - s * akq;
          A[k][q] = s * akp + c * akq;
        }
        for (int k = 0; k < 6; ++k) {
          const double apk = A[p][k], aqk = A[q][k];
          A[p][k] = c * apk - s * aqk;
          A[q][k] = s * apk + c * aqk;
        }
        for (int k = 0; k < 6; ++k) {
          const double vkp = V[k][p], vkq = V[k][q];
          V[k][p] = c * vkp - s * vkq;
          V[k][q] = s * vkp + c * vkq;
        }
      }
    }
  }
  double lmax = 0.0;
  for (int i = 0; i < 6; ++i) lmax = fmax(lmax, A[i][i]);
  for (int i = 0; i < 6; ++i) x[i] = 0.0;
  for (int e = 0; e < 6; ++e) {
    if (A[e][e] <= lmax * 1e-13 || A[e][e] <= 0.0) continue;   // null-space direction (pinv drops it)
    double proj = 0.0;
    for (int i = 0; i < 6; ++i) proj += V[i][e] * g[i];
    proj /= A[e][e];
    for (int i = 0; i < 6; ++i) x[i] += V[i][e] * proj;
  }
}

// one warp per frame pair
__global__ void __launch_bounds__(128) ibvs_kernel(const IbvsArgs a, unsigned long long* trace) {
  trace_begin(trace);
  const int lane = threadIdx.x & 31;
  const int pair = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (pair >= a.n_pairs) return;
  const int K = a.K;
  const int32_t* uv = a.uv + static_cast<size_t>(pair) * K * 2;
  const int32_t* uvs = a.uv_star + static_cast<size_t>(pair) * K * 2;
  const uint16_t* depth = a.depth + static_cast<size_t>(pair) * a.cam.width * a.cam.height;

  double acc[27];   // 21 upper-triangular entries of L^T L, then 6 of L^T e
#pragma unroll
  for (int i = 0; i < 27; ++i) acc[i] = 0.0;
  for (int k = lane; k < K; k += 32) {
    const int u = uv[2 * k], v = uv[2 * k + 1];
    double z = 100.0;                                         // vitvs_v2.py:582-584
    if (u >= 0 && u < a.cam.width && v >= 0 && v < a.cam.height) {
      const uint16_t dmm = depth[static_cast<size_t>(v) * a.cam.width + u];
      if (dmm != 0) z = static_cast<double>(dmm) / 1000.0;
    }
    const double x = (u - a.cam.cx) / a.cam.fx, y = (v - a.cam.cy) / a.cam.fy;   // :634-648
    const double xs = (uvs[2 * k] - a.cam.cx) / a.cam.fx, ys = (uvs[2 * k + 1] - a.cam.cy) / a.cam.fy;
    const double e1 = x - xs, e2 = y - ys;
    const double r1[6] = {-1.0 / z, 0.0, x / z, x * y, -(1.0 + x * x), y};        // :650-659
    const double r2[6] = {0.0, -1.0 / z, y / z, 1.0 + y * y, -x * y, -x};
    if (a.z_out) a.z_out[static_cast<size_t>(pair) * K + k] = z;
    if (a.xy_out) {
      a.xy_out[(static_cast<size_t>(pair) * K + k) * 2] = x;
      a.xy_out[(static_cast<size_t>(pair) * K + k) * 2 + 1] = y;
    }
    if (a.xy_star_out) {
      a.xy_star_out[(static_cast<size_t>(pair) * K + k) * 2] = xs;
      a.xy_star_out[(static_cast<size_t>(pair) * K + k) * 2 + 1] = ys;
    }
    if (a.L_out) {
      double* Lr = a.L_out + (static_cast<size_t>(pair) * K + k) * 12;
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        Lr[i] = r1[i];
        Lr[6 + i] = r2[i];
      }
    }
    int n = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = i; j < 6; ++j) acc[n++] += r1[i] * r1[j] + r2[i] * r2[j];
#pragma unroll
    for (int i = 0; i < 6; ++i) acc[21 + i] += r1[i] * e1 + r2[i] * e2;
  }
#pragma unroll
  for (int i = 0; i < 27; ++i) acc[i] = warp_sum_f64(acc[i]);
  if (lane != 0) return;

  double A[6][6], g[6], sol[6];
  {
    int n = 0;
    for (int i = 0; i < 6; ++i)
      for (int j = i; j < 6; ++j) {
        A[i][j] = acc[n];
        A[j][i] = acc[n];
        ++n;
      }
    for (int i = 0; i < 6; ++i) g[i] = acc[21 + i];
    if (a.mu > 0.0)   // damped least squares (vvs_config.damping_mu); untouched arithmetic when 0
      for (int i = 0; i < 6; ++i) A[i][i] += a.mu;
  }
  // Cholesky A = C C^T when safely positive definite (rank 6), else Jacobi pseudo-inverse
  double C[6][6];
  bool ok = true;
  double dmax = 0.0;
  for (int i = 0; i < 6; ++i) dmax = fmax(dmax, A[i][i]);
  for (int i = 0; i < 6 && ok; ++i) {
    for (int j = 0; j <= i; ++j) {
      double s = A[i][j];
      for (int k = 0; k < j; ++k) s -= C[i][k] * C[j][k];
      if (i == j) {
        if (!(s > dmax * 1e-10)) {
          ok = false;
          break;
        }
        C[i][i] = sqrt(s);
      } else {
        C[i][j] = s / C[j][j];
      }
    }
  }
  if (ok) {
    double y[6];
    for (int i = 0; i < 6; ++i) {
      double s = g[i];
      for (int k = 0; k < i; ++k) s -= C[i][k] * y[k];
      y[i] = s / C[i][i];
    }
    for (int i = 5; i >= 0; --i) {
      double s = y[i];
      for (int k = i + 1; k < 6; ++k) s -= C[k][i] * sol[k];
      sol[i] = s / C[i][i];
    }
  } else {
    pinv_solve6(A, g, sol);
  }
  double* tw = a.twist + static_cast<size_t>(pair) * 6;
  for (int i = 0; i < 6; ++i) {
    double v = -a.lambda * sol[i];                                               // :622
    if (a.twist_raw) a.twist_raw[static_cast<size_t>(pair) * 6 + i] = v;
    if (a.ema_state) {                                                           // :329-343
      double* st = a.ema_state + static_cast<size_t>(pair) * 6 + i;
      v = a.ema_init[pair] ? a.ema_alpha * v + (1.0 - a.ema_alpha) * (*st) : v;
      *st = v;
    }
    tw[i] = v;
  }
  if (a.ema_state) a.ema_init[pair] = 1;
  trace_end(trace);
}

}  // namespace

int launch_select(const SelectArgs& a, cudaStream_t s, std::string* err) {
  if (a.K > 256 || a.K < 1) {
    set_error(err, "select: num_pairs must be in [1, 256]");
    return VVS_ERR_INVALID;
  }
  // up to 1024 threads per pair; the selection keys live in dynamic shared memory when they fit (T <= 25 600)
  constexpr size_t kMaxKeySmem = 200 * 1024;
  static PerDeviceOnce once;
  VVS_CUDA_OK(once.run([&] {
    return cudaFuncSetAttribute(select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kMaxKeySmem));
  }), err);
  SelectArgs b = a;
  const size_t key_bytes = static_cast<size_t>(a.T) * sizeof(unsigned long long);
  b.keys_in_smem = (!a.sel_order && key_bytes <= kMaxKeySmem) ? 1 : 0;
  const int threads = std::min(1024, std::max(128, (a.T + 127) / 128 * 128));   // a function of T alone (reduction orders)
  select_kernel<<<a.n_pairs, threads, b.keys_in_smem ? key_bytes : 0, s>>>(b, tl_trace_slot);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

int launch_points(const int32_t* uv, const int32_t* uv_star, int n, const Camera& cam, double* xy, double* xy_star,
                  cudaStream_t s, std::string* err) {
  points_kernel<<<(n + 127) / 128, 128, 0, s>>>(uv, uv_star, n, cam, xy, xy_star);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

int launch_interaction(const double* xy, const double* z, int n, double* L, cudaStream_t s, std::string* err) {
  interaction_kernel<<<(n + 127) / 128, 128, 0, s>>>(xy, z, n, L);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

int launch_ibvs(const IbvsArgs& a, cudaStream_t s, std::string* err) {
  ibvs_kernel<<<(a.n_pairs + 3) / 4, 128, 0, s>>>(a, tl_trace_slot);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

}  // namespace vvs
