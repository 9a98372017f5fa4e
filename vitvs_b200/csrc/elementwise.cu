// HBM-bound kernels of the dino path (SURVEY §2b k0,k2,k8): pre-processing + im2col for the
// patch-embed GEMM, LayerNorm, descriptor gather (CLS drop, 3x3 log-binning) and L2 normalisation.
// Vectorised (128-bit) coalesced access, one warp per row with shuffle reductions; no tensor work.
#include <string>

#include "ptx.cuh"
#include "vvs_internal.h"

namespace vvs {
namespace {

constexpr int kPatch = 14;
constexpr int kPatchK = 3 * kPatch * kPatch;   // 588

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

// ToTensor + Normalize (dinov2_extractor.py:186-190, constants :49-50), evaluated in fp32 as torch does:
// (u8 / 255 - mean) / std with IEEE division.
__device__ __forceinline__ float normalize_px(uint8_t v, int c) {
  const float mean[3] = {0.485f, 0.456f, 0.406f};
  const float stdv[3] = {0.229f, 0.224f, 0.225f};
  return __fdiv_rn(__fsub_rn(__fdiv_rn(static_cast<float>(v), 255.0f), mean[c]), stdv[c]);
}

// A_pe[img*T + t][k], k = c*196 + ky*14 + kx (Conv2d weight [D,3,14,14] flattened), zero padded to kpad.
// One thread = 8 consecutive k of one patch row -> one 16-byte store.
template <bool kU8>
__global__ void im2col_kernel(const void* __restrict__ img_, bf16* __restrict__ out, int n_img, int S, int kpad, int gp,
                              int stride, unsigned long long* trace) {
  trace_begin(trace);
  // u8 input: the 3 x 256 possible outputs of ToTensor + Normalize, evaluated once per block with the exact
  // arithmetic above (two IEEE divisions each) - the per-element work is then a table lookup
  __shared__ float lut[3][256];
  if (kU8) {
    for (int i = threadIdx.x; i < 768; i += blockDim.x) lut[i >> 8][i & 255] = normalize_px(static_cast<uint8_t>(i & 255), i >> 8);
    __syncthreads();
  }
  const int T = gp * gp;                      // gp = 1 + (S - 14) / stride patches per side (dinov2_extractor.py:262)
  const int k8_per_row = kpad / 8;
  const long long total = static_cast<long long>(n_img) * T * k8_per_row;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int k8 = static_cast<int>(idx % k8_per_row);
    const long long row = idx / k8_per_row;
    const int t = static_cast<int>(row % T);
    const int img = static_cast<int>(row / T);
    const int py = t / gp, px = t - py * gp;
    float v[8];
    // (c, ky, kx) of the first element once, then carried: k = c * 196 + ky * 14 + kx
    const int kfirst = k8 * 8;
    int c = kfirst / (kPatch * kPatch);
    int ky = (kfirst - c * kPatch * kPatch) / kPatch;
    int kx = kfirst - c * kPatch * kPatch - ky * kPatch;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      if (kfirst + e < kPatchK) {
        const int y = py * stride + ky, x = px * stride + kx;   // stride < 14: overlapping patches (:123-144)
        if (kU8) {
          const uint8_t* img8 = static_cast<const uint8_t*>(img_);
          v[e] = lut[c][img8[((static_cast<size_t>(img) * S + y) * S + x) * 3 + c]];
        } else {
          const float* imgf = static_cast<const float*>(img_);
          v[e] = imgf[((static_cast<size_t>(img) * 3 + c) * S + y) * S + x];
        }
      } else {
        v[e] = 0.f;
      }
      if (++kx == kPatch) {
        kx = 0;
        if (++ky == kPatch) {
          ky = 0;
          ++c;
        }
      }
    }
    uint4 w = make_uint4(pack2(v[0], v[1]), pack2(v[2], v[3]), pack2(v[4], v[5]), pack2(v[6], v[7]));
    // row img * (T + 1) + 1 + t: the matrix has one (all-zero) row per CLS slot, so that the patch-embed GEMM's rows
    // are the residual stream's rows and its epilogue can leave through a plain 2-D TMA store
    *reinterpret_cast<uint4*>(out + (row + img + 1) * kpad + k8 * 8) = w;
  }
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
       idx < static_cast<long long>(n_img) * k8_per_row; idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long img = idx / k8_per_row;
    *reinterpret_cast<uint4*>(out + img * (T + 1) * kpad + (idx - img * k8_per_row) * 8) = make_uint4(0u, 0u, 0u, 0u);
  }
  trace_end(trace);
}

// pos'[0] = cls + pos[0] - patch_bias (the CLS row of the patch-embed GEMM has a zero accumulator: acc + bias + pos'[0]
// = cls + pos[0]); pos'[r] = pos[r] for the patch rows.
__global__ void pos_prime_kernel(const float* __restrict__ pos, const float* __restrict__ cls,
                                 const float* __restrict__ bias, float* __restrict__ out, int n_tok, int D) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_tok * D) return;
  out[idx] = idx < D ? (cls[idx] + pos[idx]) - bias[idx] : pos[idx];
}

__global__ void cls_rows_kernel(float* __restrict__ x, const float* __restrict__ cls, const float* __restrict__ pos,
                                int n_img, int n_tok, int D) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_img * D) return;
  const int img = idx / D, d = idx - img * D;
  x[static_cast<size_t>(img) * n_tok * D + d] = cls[d] + pos[d];
}

// LayerNorm over D (eps 1e-6, affine), fp32 in -> bf16 out. One warp per row, NV float4 per lane.
template <int NV>
__global__ void __launch_bounds__(256)
layernorm_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                 bf16* __restrict__ out, int rows, float eps, int pin_row0, unsigned long long* trace) {
  constexpr int D = NV * 128;
  pdl_launch_dependents();
  pdl_wait();   // x is the previous kernel's output
  trace_begin(trace);
  const int lane = threadIdx.x & 31;
  // The producer (a persistent GEMM) wrote x in ascending row order and x is larger than L2: walk the rows from
  // the end so that the reads start on the lines that are still resident; the consumer GEMM starts at row 0,
  // which is then the part of the output written last.
  const int row = rows - 1 - (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5));
  if (row < 0) return;
  const float4* xr = reinterpret_cast<const float4*>(x + static_cast<size_t>(row) * D);
  float4 v[NV];
  float s = 0.f;
  // rows >= pin_row0 are the part of the residual stream every kernel asks L2 to keep (evict-last); the rest is
  // dead until the next residual GEMM rewrites it: evict-first, so that it does not push the kept part out
  const uint64_t pol = (row >= pin_row0) ? kEvictLast : kEvictFirst;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    v[i] = (pin_row0 < rows) ? ld_global_v4_hint(xr + i * 32 + lane, pol) : xr[i * 32 + lane];
    s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  }
  const float mean = warp_sum(s) * (1.0f / D);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const float a = v[i].x - mean, c = v[i].y - mean, e = v[i].z - mean, f = v[i].w - mean;
    q += (a * a + c * c) + (e * e + f * f);
  }
  const float rstd = rsqrtf(warp_sum(q) * (1.0f / D) + eps);
  uint2* orow = reinterpret_cast<uint2*>(out + static_cast<size_t>(row) * D);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const float4 g = __ldg(reinterpret_cast<const float4*>(w) + i * 32 + lane);
    const float4 h = __ldg(reinterpret_cast<const float4*>(b) + i * 32 + lane);
    const float y0 = (v[i].x - mean) * rstd * g.x + h.x;
    const float y1 = (v[i].y - mean) * rstd * g.y + h.y;
    const float y2 = (v[i].z - mean) * rstd * g.z + h.z;
    const float y3 = (v[i].w - mean) * rstd * g.w + h.w;
    orow[i * 32 + lane] = make_uint2(pack2(y0, y1), pack2(y2, y3));
  }
  trace_end(trace);
}

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// x: [n_img, 1+T, D] residual stream.  out: [n_img, T, P*D] fp32, P = 9 (bin) or 1.
// dinov2_extractor.py:326-336 + :265-311 (hierarchy 1): parts in (dy,dx) row-major order, replicate clamp.
__global__ void descriptors_f32_kernel(const float* __restrict__ x, float* __restrict__ out, int n_img, int hp, int wp,
                                       int D, int parts) {
  const int T = hp * wp;
  const int d4 = D / 4;
  const long long total = static_cast<long long>(n_img) * T * parts * d4;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c4 = static_cast<int>(idx % d4);
    long long r = idx / d4;
    const int part = static_cast<int>(r % parts);
    r /= parts;
    const int t = static_cast<int>(r % T);
    const int img = static_cast<int>(r / T);
    int sy = t / wp, sx = t - sy * wp;
    if (parts == 9) {
      sy = clampi(sy + part / 3 - 1, 0, hp - 1);
      sx = clampi(sx + part % 3 - 1, 0, wp - 1);
    }
    const float4 v = *reinterpret_cast<const float4*>(x + (static_cast<size_t>(img) * (T + 1) + 1 + sy * wp + sx) * D +
                                                      c4 * 4);
    *reinterpret_cast<float4*>(out + ((static_cast<size_t>(img) * T + t) * parts + part) * D + c4 * 4) = v;
  }
}

// Log-binned descriptors with any hierarchy (dinov2_extractor.py:265-311): bin (k, dy, dx) of token
// (y, x) = the 3^k x 3^k average (in-grid elements only: count_include_pad=False) around the replicate-clamped
// neighbour (y + dy 3^k, x + dx 3^k); the centre bin exists at k = 0 only.  Bins level-major, then (dy, dx) row-major.
// x: [n_img, 1+T, D] fp32 -> out [n_img, T, (1 + 8 h) D] fp32.  Thread = 4 channels of one bin of one token.
__global__ void descriptors_hier_f32_kernel(const float* __restrict__ x, float* __restrict__ out, int n_img, int hp,
                                            int wp, int D, int hierarchy) {
  const int T = hp * wp;
  const int bins = 1 + 8 * hierarchy;
  const int d4 = D / 4;
  const long long total = static_cast<long long>(n_img) * T * bins * d4;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c4 = static_cast<int>(idx % d4);
    long long r = idx / d4;
    const int bin = static_cast<int>(r % bins);
    r /= bins;
    const int t = static_cast<int>(r % T);
    const int img = static_cast<int>(r / T);
    int k = 0, slot = bin;                     // level 0 has 9 slots, deeper levels 8 (centre skipped)
    if (bin >= 9) {
      k = 1 + (bin - 9) / 8;
      slot = (bin - 9) % 8;
      if (slot >= 4) ++slot;                   // re-insert the skipped centre
    }
    int win = 1;
    for (int i = 0; i < k; ++i) win *= 3;
    const int y0 = t / wp, x0 = t - y0 * wp;
    const int cy = clampi(y0 + (slot / 3 - 1) * win, 0, hp - 1);
    const int cx = clampi(x0 + (slot % 3 - 1) * win, 0, wp - 1);
    const int half = win / 2;
    const int ya = max(cy - half, 0), yb = min(cy + half, hp - 1);
    const int xa = max(cx - half, 0), xb = min(cx + half, wp - 1);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int yy = ya; yy <= yb; ++yy)
      for (int xx = xa; xx <= xb; ++xx) {
        const float4 v = *reinterpret_cast<const float4*>(x + (static_cast<size_t>(img) * (T + 1) + 1 + yy * wp + xx) * D +
                                                          c4 * 4);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
    const float n = static_cast<float>((yb - ya + 1) * (xb - xa + 1));
    if (win > 1) { acc.x /= n; acc.y /= n; acc.z /= n; acc.w /= n; }
    *reinterpret_cast<float4*>(out + ((static_cast<size_t>(img) * T + t) * bins + bin) * D + c4 * 4) = acc;
  }
}

// facet descriptors (dinov2_extractor.py:203-218, 326-336): q / k / v of the block's attention input as
// [B, h, t, d] -> permute(0, 2, 3, 1).flatten(-2): channel c = dd * heads + hh reads qkv column facet*D + hh*64 + dd.
// Thread = one output channel of one (image, token, part): writes coalesced, reads gather from an L2-resident row.
__global__ void facet_descriptors_f32_kernel(const bf16* __restrict__ qkv, float* __restrict__ out, int n_img, int hp,
                                             int wp, int D, int heads, int facet, int parts, int include_cls) {
  const int T = hp * wp;
  const int rows_out = T + (include_cls ? 1 : 0);
  const long long total = static_cast<long long>(n_img) * rows_out * parts * D;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(idx % D);
    long long r = idx / D;
    const int part = static_cast<int>(r % parts);
    r /= parts;
    const int t = static_cast<int>(r % rows_out);
    const int img = static_cast<int>(r / rows_out);
    int src_tok;                                         // row inside the image's 1+T tokens
    if (include_cls) {
      src_tok = t;
    } else {
      int sy = t / wp, sx = t - sy * wp;
      if (parts == 9) {
        sy = clampi(sy + part / 3 - 1, 0, hp - 1);
        sx = clampi(sx + part % 3 - 1, 0, wp - 1);
      }
      src_tok = 1 + sy * wp + sx;
    }
    const int dd = c / heads, hh = c - dd * heads;
    out[idx] = __bfloat162float(qkv[(static_cast<size_t>(img) * (T + 1) + src_tok) * (3 * D) + facet * D + hh * 64 + dd]);
  }
}

// squared L2 norm of every patch token (CLS skipped): one warp per token.
// zero_a / zero_b (optional): the packed arg-max keys of the matching stage, zero_n entries each, cleared here by the
// first zero_n tokens - the two memset nodes a frame pair's graph would otherwise carry (2-3 us each at one pair).
__global__ void token_sqnorm_kernel(const float* __restrict__ x, float* __restrict__ n2, int n_img, int T, int D,
                                    unsigned long long* __restrict__ zero_a, unsigned long long* __restrict__ zero_b,
                                    long long zero_n, unsigned long long* trace) {
  trace_begin(trace);
  const int lane = threadIdx.x & 31;
  const long long tok = blockIdx.x * static_cast<long long>(blockDim.x >> 5) + (threadIdx.x >> 5);
  if (tok >= static_cast<long long>(n_img) * T) return;
  const int img = static_cast<int>(tok / T), t = static_cast<int>(tok % T);
  const float4* r = reinterpret_cast<const float4*>(x + (static_cast<size_t>(img) * (T + 1) + 1 + t) * D);
  float s = 0.f;
  for (int i = lane; i < D / 4; i += 32) {
    const float4 v = r[i];
    s += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
  }
  s = warp_sum(s);
  if (lane == 0) {
    n2[tok] = s;
    if (zero_a != nullptr && tok < zero_n) {
      zero_a[tok] = 0ull;
      zero_b[tok] = 0ull;
    }
  }
  trace_end(trace);
}

// unit-norm bf16 descriptors for the similarity GEMM: ||F_p||^2 = sum over the 9 clamped neighbours of
// ||f||^2 (log-binning is a gather), so the binned row is written already scaled by 1/||F_p||.
// One block (128 threads) per (image, token).
__global__ void __launch_bounds__(128)
descriptors_unit_bf16_kernel(const float* __restrict__ x, const float* __restrict__ n2, bf16* __restrict__ out, int hp,
                             int wp, int D, int parts, unsigned long long* trace) {
  trace_begin(trace);
  const int T = hp * wp;
  const int t = blockIdx.x % T;
  const int img = blockIdx.x / T;
  const int y0 = t / wp, x0 = t - y0 * wp;
  int src[9];
  float tot = 0.f;
#pragma unroll
  for (int p = 0; p < 9; ++p) {
    if (p < parts) {
      int sy = y0, sx = x0;
      if (parts == 9) {
        sy = clampi(y0 + p / 3 - 1, 0, hp - 1);
        sx = clampi(x0 + p % 3 - 1, 0, wp - 1);
      }
      src[p] = sy * wp + sx;
      tot += n2[static_cast<size_t>(img) * T + src[p]];
    }
  }
  const float inv = 1.0f / fmaxf(sqrtf(tot), 1e-8f);   // CosineSimilarity eps (vitvs_v2.py:55)
  const int d8 = D / 8;
  bf16* orow = out + (static_cast<size_t>(img) * T + t) * parts * D;
  for (int i = threadIdx.x; i < parts * d8; i += blockDim.x) {
    const int p = i / d8, c8 = i - p * d8;
    const float4* s4 =
        reinterpret_cast<const float4*>(x + (static_cast<size_t>(img) * (T + 1) + 1 + src[p]) * D + c8 * 8);
    const float4 a = s4[0], b = s4[1];
    const uint4 w = make_uint4(pack2(a.x * inv, a.y * inv), pack2(a.z * inv, a.w * inv), pack2(b.x * inv, b.y * inv),
                               pack2(b.z * inv, b.w * inv));
    *reinterpret_cast<uint4*>(orow + p * D + c8 * 8) = w;
  }
  trace_end(trace);
}

// arbitrary fp32 rows -> unit-norm bf16 rows (zero padded to dpad).  One WARP per row, 128-bit loads, the row kept in
// registers between the norm pass and the write pass when it fits (d <= 2048; longer rows are read a second time, from
// L2), 64-bit packed bf16 stores.  Bandwidth-bound: 4 d bytes in, 2 dpad bytes out per row.
constexpr int kNormKeep = 16;   // float4 per lane kept in registers (16 * 32 lanes * 4 = 2048 elements)
__global__ void __launch_bounds__(256)
normalize_rows_kernel(const float* __restrict__ in, bf16* __restrict__ out, long long rows, int d, int dpad,
                      unsigned long long* trace) {
  trace_begin(trace);
  const int lane = threadIdx.x & 31;
  const long long row = blockIdx.x * static_cast<long long>(blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float* r = in + row * d;
  bf16* o = out + row * dpad;
  const bool vec = (d & 3) == 0 && (reinterpret_cast<uintptr_t>(in) & 15) == 0;
  float s = 0.f;
  float4 keep[kNormKeep];
  const int n4 = d >> 2;
  if (vec) {
    const float4* r4 = reinterpret_cast<const float4*>(r);
#pragma unroll
    for (int i = 0; i < kNormKeep; ++i) {
      const int idx = i * 32 + lane;
      keep[i] = idx < n4 ? __ldg(r4 + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
      s += (keep[i].x * keep[i].x + keep[i].y * keep[i].y) + (keep[i].z * keep[i].z + keep[i].w * keep[i].w);
    }
    for (int idx = kNormKeep * 32 + lane; idx < n4; idx += 32) {
      const float4 v = __ldg(r4 + idx);
      s += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
    }
  } else {
    for (int i = lane; i < d; i += 32) s += r[i] * r[i];
  }
  s = warp_sum(s);
  const float inv = 1.0f / fmaxf(sqrtf(s), 1e-8f);   // CosineSimilarity eps (vitvs_v2.py:55)
  if (vec) {
    uint2* o2 = reinterpret_cast<uint2*>(o);          // dpad is a multiple of 8 and the workspace 256-byte aligned
    const float4* r4 = reinterpret_cast<const float4*>(r);
#pragma unroll
    for (int i = 0; i < kNormKeep; ++i) {
      const int idx = i * 32 + lane;
      if (idx < n4) o2[idx] = make_uint2(pack2(keep[i].x * inv, keep[i].y * inv), pack2(keep[i].z * inv, keep[i].w * inv));
    }
    for (int idx = kNormKeep * 32 + lane; idx < n4; idx += 32) {
      const float4 v = __ldg(r4 + idx);
      o2[idx] = make_uint2(pack2(v.x * inv, v.y * inv), pack2(v.z * inv, v.w * inv));
    }
    for (int i = d + lane; i < dpad; i += 32) o[i] = __float2bfloat16_rn(0.f);
  } else {
    for (int i = lane; i < dpad; i += 32) o[i] = __float2bfloat16_rn(i < d ? r[i] * inv : 0.f);
  }
  trace_end(trace);
}

int grid_for(long long work_items, int block) {
  long long g = (work_items + block - 1) / block;
  return static_cast<int>(g < 1 ? 1 : (g > 148 * 32 ? 148 * 32 : g));
}

}  // namespace

int launch_im2col_u8(const uint8_t* img, bf16* out, int n_img, int size, int kpad, int grid, int stride, cudaStream_t s,
                     std::string* err) {
  const long long items = static_cast<long long>(n_img) * grid * grid * (kpad / 8);
  im2col_kernel<true><<<grid_for(items, 256), 256, 0, s>>>(img, out, n_img, size, kpad, grid, stride, tl_trace_slot);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

int launch_im2col_f32(const float* img, bf16* out, int n_img, int size, int kpad, int grid, int stride, cudaStream_t s,
                      std::string* err) {
  const long long items = static_cast<long long>(n_img) * grid * grid * (kpad / 8);
  im2col_kernel<false><<<grid_for(items, 256), 256, 0, s>>>(img, out, n_img, size, kpad, grid, stride, tl_trace_slot);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

int launch_pos_prime(const float* pos, const float* cls, const float* bias, float* out, int n_tok, int D, cudaStream_t s,
                     std::string* err) {
  pos_prime_kernel<<<(n_tok * D + 255) / 256, 256, 0, s>>>(pos, cls, bias, out, n_tok, D);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

int launch_cls_rows(float* x, const float* cls, const float* pos, int n_img, int n_tok, int D, cudaStream_t s,
                    std::string* err) {
  cls_rows_kernel<<<(n_img * D + 255) / 256, 256, 0, s>>>(x, cls, pos, n_img, n_tok, D);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

int launch_layernorm(const float* x, const float* w, const float* b, bf16* out, int rows, int D, float eps, int pin_row0,
                     cudaStream_t s, std::string* err, bool pdl) {
  const int grid = (rows + 7) / 8;
  cudaError_t e;
  switch (D) {
    case 384: e = launch_kernel(layernorm_kernel<3>, grid, 256, 0, s, pdl, x, w, b, out, rows, eps, pin_row0, tl_trace_slot); break;
    case 768: e = launch_kernel(layernorm_kernel<6>, grid, 256, 0, s, pdl, x, w, b, out, rows, eps, pin_row0, tl_trace_slot); break;
    case 1024: e = launch_kernel(layernorm_kernel<8>, grid, 256, 0, s, pdl, x, w, b, out, rows, eps, pin_row0, tl_trace_slot); break;
    case 1536: e = launch_kernel(layernorm_kernel<12>, grid, 256, 0, s, pdl, x, w, b, out, rows, eps, pin_row0, tl_trace_slot); break;
    default: set_error(err, "layernorm: unsupported width (384/768/1024/1536)"); return VVS_ERR_INVALID;
  }
  VVS_CUDA_OK(e, err);
  return VVS_OK;
}

int launch_descriptors_f32(const float* x, float* out, int n_img, int hp, int wp, int D, int bin, cudaStream_t s,
                           std::string* err) {
  const int parts = bin ? 9 : 1;
  const long long items = static_cast<long long>(n_img) * hp * wp * parts * (D / 4);
  descriptors_f32_kernel<<<grid_for(items, 256), 256, 0, s>>>(x, out, n_img, hp, wp, D, parts);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

int launch_descriptors_hier_f32(const float* x, float* out, int n_img, int hp, int wp, int D, int hierarchy,
                                cudaStream_t s, std::string* err) {
  const long long items = static_cast<long long>(n_img) * hp * wp * (1 + 8 * hierarchy) * (D / 4);
  descriptors_hier_f32_kernel<<<grid_for(items, 256), 256, 0, s>>>(x, out, n_img, hp, wp, D, hierarchy);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

int launch_facet_descriptors_f32(const bf16* qkv, float* out, int n_img, int hp, int wp, int D, int heads, int facet,
                                 int bin, int include_cls, cudaStream_t s, std::string* err) {
  const int parts = bin ? 9 : 1;
  const long long items = static_cast<long long>(n_img) * (hp * wp + (include_cls ? 1 : 0)) * parts * D;
  facet_descriptors_f32_kernel<<<grid_for(items, 256), 256, 0, s>>>(qkv, out, n_img, hp, wp, D, heads, facet, parts,
                                                                     include_cls);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

int launch_descriptors_unit_bf16(const float* x, bf16* out, float* sqnorm_scratch, int n_img, int hp, int wp, int D,
                                 int bin, cudaStream_t s, std::string* err, unsigned long long* zero_a,
                                 unsigned long long* zero_b, long long zero_n) {
  const int T = hp * wp;
  const long long toks = static_cast<long long>(n_img) * T;
  if (zero_a != nullptr && (zero_b == nullptr || zero_n > toks)) {
    set_error(err, "descriptors: key clearing needs both arrays and at most one entry per token");
    return VVS_ERR_INVALID;
  }
  token_sqnorm_kernel<<<static_cast<int>((toks + 7) / 8), 256, 0, s>>>(x, sqnorm_scratch, n_img, T, D, zero_a, zero_b, zero_n,
                                                                       tl_trace_slot);
  VVS_CUDA_OK(cudaGetLastError(), err);
  descriptors_unit_bf16_kernel<<<static_cast<int>(toks), 128, 0, s>>>(x, sqnorm_scratch, out, hp, wp, D, bin ? 9 : 1,
                                                                      tl_trace_slot);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

int launch_normalize_rows_bf16(const float* in, bf16* out, int rows, int d, int dpad, cudaStream_t s,
                               std::string* err) {
  normalize_rows_kernel<<<(rows + 7) / 8, 256, 0, s>>>(in, out, rows, d, dpad, tl_trace_slot);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

}  // namespace vvs
