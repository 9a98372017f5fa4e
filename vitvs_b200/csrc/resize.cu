// Camera frame -> ViT input: Pillow-exact 8-bit antialiased bicubic resize.
//
// replaces: PIL Image.resize((S, S)) in Controller.detect_features (vitvs_v2.py:474-475).  Pillow is a third-party
// dependency of the reference (not vendored); the algorithm restated here is the published one of Pillow's
// src/libImaging/Resample.c (precompute_coeffs, normalize_coeffs_8bpc, ImagingResampleHorizontal/Vertical_8bpc):
// two separable passes, horizontal first, filter support scaled by the down-scale factor (antialiasing), coefficients
// normalised in double and rounded to 22-bit fixed point, accumulator seeded with 2^21, arithmetic shift and clip to
// [0, 255] after each pass.  Integer arithmetic => the result is bit-identical to Pillow's (tests/test_gpu_resize.py).
//
// Both passes are HBM/L2-streaming byte kernels: the horizontal pass stages whole input rows in shared memory with
// 128-bit loads and writes the narrow intermediate [rows, out_w, 3]; the vertical pass walks that intermediate (L2
// resident: 57 MB for 128 frames) four bytes per thread with fully coalesced 32-bit accesses.
#include <algorithm>
#include <cmath>
#include <string>
#include <vector>

#include "vvs_internal.h"

namespace vvs {
namespace {

constexpr int kPrecisionBits = 32 - 8 - 2;   // Resample.c PRECISION_BITS

double bicubic_filter(double x) {            // Resample.c bicubic_filter, a = -0.5
  const double a = -0.5;
  if (x < 0.0) x = -x;
  if (x < 1.0) return ((a + 2.0) * x - (a + 3.0)) * x * x + 1;
  if (x < 2.0) return (((x - 5) * x + 8) * x - 4) * a;
  return 0.0;
}

__device__ __forceinline__ uint32_t clip8(int acc) {
  return static_cast<uint32_t>(min(max(acc >> kPrecisionBits, 0), 255));
}

constexpr int kRowsPerBlock = 4;
constexpr int kHThreads = 320;

// in [n_rows, in_w, 3] u8 -> out [n_rows, out_w, 3] u8.  One block = kRowsPerBlock input rows staged in smem.
__global__ void __launch_bounds__(kHThreads)
resize_h_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int n_rows, int in_w, int out_w,
                const int32_t* __restrict__ bounds, const int32_t* __restrict__ kk, int ksize, int row_pitch) {
  extern __shared__ __align__(16) uint8_t srow[];
  const int row0 = blockIdx.x * kRowsPerBlock;
  const int rows = min(kRowsPerBlock, n_rows - row0);
  const size_t in_bytes = static_cast<size_t>(in_w) * 3;
  const uint8_t* src = in + static_cast<size_t>(row0) * in_bytes;
  if ((in_bytes & 15) == 0 && (reinterpret_cast<uintptr_t>(in) & 15) == 0) {
    const int vec_per_row = static_cast<int>(in_bytes >> 4);
    for (int i = threadIdx.x; i < rows * vec_per_row; i += kHThreads) {
      const int r = i / vec_per_row, v = i - r * vec_per_row;
      *reinterpret_cast<uint4*>(srow + r * row_pitch + v * 16) =
          __ldg(reinterpret_cast<const uint4*>(src + r * in_bytes) + v);
    }
  } else {
    for (int i = threadIdx.x; i < rows * static_cast<int>(in_bytes); i += kHThreads) {
      const int r = i / static_cast<int>(in_bytes), b = i - r * static_cast<int>(in_bytes);
      srow[r * row_pitch + b] = src[r * in_bytes + b];
    }
  }
  __syncthreads();
  for (int x = threadIdx.x; x < out_w; x += kHThreads) {
    const int xmin = __ldg(bounds + 2 * x), cnt = __ldg(bounds + 2 * x + 1);
    int acc[kRowsPerBlock][3];
#pragma unroll
    for (int r = 0; r < kRowsPerBlock; ++r) acc[r][0] = acc[r][1] = acc[r][2] = 1 << (kPrecisionBits - 1);
    const uint8_t* p = srow + xmin * 3;
    const int32_t* kx = kk + x * ksize;
    for (int t = 0; t < cnt; ++t) {
      const int k = __ldg(kx + t);
#pragma unroll
      for (int r = 0; r < kRowsPerBlock; ++r) {
        const uint8_t* q = p + r * row_pitch + t * 3;
        acc[r][0] += q[0] * k;
        acc[r][1] += q[1] * k;
        acc[r][2] += q[2] * k;
      }
    }
#pragma unroll
    for (int r = 0; r < kRowsPerBlock; ++r) {
      if (r < rows) {
        uint8_t* o = out + (static_cast<size_t>(row0 + r) * out_w + x) * 3;
        o[0] = static_cast<uint8_t>(clip8(acc[r][0]));
        o[1] = static_cast<uint8_t>(clip8(acc[r][1]));
        o[2] = static_cast<uint8_t>(clip8(acc[r][2]));
      }
    }
  }
}

// in [n_img, in_h, row_bytes] u8 -> out [n_img, out_h, row_bytes] u8; thread = kVec consecutive bytes of one output row.
template <int kVec>
__global__ void __launch_bounds__(256)
resize_v_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int n_img, int in_h, int out_h,
                int row_bytes, const int32_t* __restrict__ bounds, const int32_t* __restrict__ kk, int ksize) {
  const int groups = row_bytes / kVec;
  const long long total = static_cast<long long>(n_img) * out_h * groups;
  for (long long idx = blockIdx.x * 256ll + threadIdx.x; idx < total; idx += gridDim.x * 256ll) {
    const int g = static_cast<int>(idx % groups);
    const long long row = idx / groups;
    const int y = static_cast<int>(row % out_h);
    const int img = static_cast<int>(row / out_h);
    const int ymin = __ldg(bounds + 2 * y), cnt = __ldg(bounds + 2 * y + 1);
    const int32_t* ky = kk + y * ksize;
    const uint8_t* p = in + (static_cast<size_t>(img) * in_h + ymin) * row_bytes + g * kVec;
    int acc[kVec];
#pragma unroll
    for (int e = 0; e < kVec; ++e) acc[e] = 1 << (kPrecisionBits - 1);
    for (int t = 0; t < cnt; ++t) {
      const int k = __ldg(ky + t);
      if (kVec == 4) {
        const uint32_t w = *reinterpret_cast<const uint32_t*>(p + static_cast<size_t>(t) * row_bytes);
        acc[0] += static_cast<int>(w & 0xffu) * k;
        acc[1 % kVec] += static_cast<int>((w >> 8) & 0xffu) * k;
        acc[2 % kVec] += static_cast<int>((w >> 16) & 0xffu) * k;
        acc[3 % kVec] += static_cast<int>(w >> 24) * k;
      } else {
        acc[0] += p[static_cast<size_t>(t) * row_bytes] * k;
      }
    }
    uint8_t* o = out + static_cast<size_t>(row) * row_bytes + g * kVec;
    if (kVec == 4)
      *reinterpret_cast<uint32_t*>(o) =
          clip8(acc[0]) | (clip8(acc[1 % kVec]) << 8) | (clip8(acc[2 % kVec]) << 16) | (clip8(acc[3 % kVec]) << 24);
    else
      o[0] = static_cast<uint8_t>(clip8(acc[0]));
  }
}

}  // namespace

// Resample.c precompute_coeffs + normalize_coeffs_8bpc over the full box [0, in_size); tables go to the device.
int make_resize_axis(ResizeAxis* ax, int in_size, int out_size, std::string* err) {
  ax->in_size = in_size;
  ax->out_size = out_size;
  const double scale = static_cast<double>(in_size) / out_size;
  const double filterscale = scale < 1.0 ? 1.0 : scale;
  const double support = 2.0 * filterscale;
  const int ksize = static_cast<int>(ceil(support)) * 2 + 1;
  ax->ksize = ksize;
  std::vector<int32_t> bounds(static_cast<size_t>(out_size) * 2), kk(static_cast<size_t>(out_size) * ksize, 0);
  std::vector<double> w(ksize);
  const double ss = 1.0 / filterscale;
  for (int xx = 0; xx < out_size; ++xx) {
    const double center = (xx + 0.5) * scale;
    int xmin = static_cast<int>(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = static_cast<int>(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    double ww = 0.0;
    for (int x = 0; x < xmax; ++x) {
      w[x] = bicubic_filter((x + xmin - center + 0.5) * ss);
      ww += w[x];
    }
    for (int x = 0; x < xmax; ++x) {
      const double k = ww != 0.0 ? w[x] / ww : w[x];
      kk[static_cast<size_t>(xx) * ksize + x] = k < 0 ? static_cast<int>(-0.5 + k * (1 << kPrecisionBits))
                                                      : static_cast<int>(0.5 + k * (1 << kPrecisionBits));
    }
    bounds[2 * xx] = xmin;
    bounds[2 * xx + 1] = xmax;
  }
  VVS_CUDA_OK(cudaMalloc(reinterpret_cast<void**>(&ax->bounds), bounds.size() * sizeof(int32_t)), err);
  VVS_CUDA_OK(cudaMalloc(reinterpret_cast<void**>(&ax->kk), kk.size() * sizeof(int32_t)), err);
  VVS_CUDA_OK(cudaMemcpy(ax->bounds, bounds.data(), bounds.size() * sizeof(int32_t), cudaMemcpyHostToDevice), err);
  VVS_CUDA_OK(cudaMemcpy(ax->kk, kk.data(), kk.size() * sizeof(int32_t), cudaMemcpyHostToDevice), err);
  return VVS_OK;
}

void free_resize_axis(ResizeAxis* ax) {
  if (ax->bounds) cudaFree(ax->bounds);
  if (ax->kk) cudaFree(ax->kk);
  ax->bounds = nullptr;
  ax->kk = nullptr;
}

int launch_resize_h(const uint8_t* in, uint8_t* out, int n_rows, const ResizeAxis& ax, cudaStream_t s, std::string* err) {
  const int row_pitch = (ax.in_size * 3 + 15) / 16 * 16 + 16;
  const size_t smem = static_cast<size_t>(kRowsPerBlock) * row_pitch;
  if (smem > 200 * 1024) {
    set_error(err, "resize: input rows wider than the shared-memory staging allows");
    return VVS_ERR_INVALID;
  }
  static PerDeviceOnce once;
  if (smem > 48 * 1024)
    VVS_CUDA_OK(once.run([] {
                  return cudaFuncSetAttribute(resize_h_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
                }),
                err);
  const int grid = (n_rows + kRowsPerBlock - 1) / kRowsPerBlock;
  resize_h_kernel<<<grid, kHThreads, smem, s>>>(in, out, n_rows, ax.in_size, ax.out_size, ax.bounds, ax.kk, ax.ksize,
                                               row_pitch);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

int launch_resize_v(const uint8_t* in, uint8_t* out, int n_img, int row_bytes, const ResizeAxis& ax, cudaStream_t s,
                    std::string* err) {
  const bool vec = (row_bytes & 3) == 0 && (reinterpret_cast<uintptr_t>(in) & 3) == 0 &&
                   (reinterpret_cast<uintptr_t>(out) & 3) == 0;
  const long long items = static_cast<long long>(n_img) * ax.out_size * (vec ? row_bytes / 4 : row_bytes);
  const int grid = static_cast<int>(std::min<long long>((items + 255) / 256, 148ll * 64));
  if (vec)
    resize_v_kernel<4><<<grid, 256, 0, s>>>(in, out, n_img, ax.in_size, ax.out_size, row_bytes, ax.bounds, ax.kk, ax.ksize);
  else
    resize_v_kernel<1><<<grid, 256, 0, s>>>(in, out, n_img, ax.in_size, ax.out_size, row_bytes, ax.bounds, ax.kk, ax.ksize);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

// ---- the same resize for callers that must not allocate device memory (vvs_resize_frames_static) ------------------
//
// vvs_resize_frames keeps three things in cudaMalloc'ed memory of its handle: the tap tables of the two passes and the
// intermediate image between them.  Here all three live in ONE fixed array that is part of the library image, so the
// call never asks the driver (or torch, or anybody) for memory.  Layout of the array, byte offsets:
//   [0, kTabBytes)                     horizontal pass: bounds [S, 2] then taps [S, ksize], int32
//   [kTabBytes, 2 * kTabBytes)         vertical pass, same layout
//   [kMidOffset, kMidOffset + kMidMax) intermediate image [n, frame_height, S, 3] u8 (output of the horizontal pass)
// Every GPU that loads the library carries its own copy of the array from the moment the module is loaded until the
// process exits, which is why the limits are modest.
//
// Nothing in the array has to start at zero: the tables are rewritten whenever the geometry differs from the one the
// device currently holds, and each call's horizontal pass writes every intermediate byte its vertical pass reads.  What
// the array does keep is the previous call's contents, and there is only one of it per GPU: two calls that are not
// ordered one after the other on the device (different streams, no event between them) overwrite each other's data.
namespace {

constexpr int kStaticMaxSide = 1024;                                   // largest output side S
constexpr int kStaticMaxTaps = 512;                                    // largest filter length of one output sample
constexpr size_t kTabInts = 32768;                                     // per pass: 2 * S + S * ksize must fit
constexpr size_t kTabBytes = kTabInts * sizeof(int32_t);               // 128 KiB
constexpr size_t kMidOffset = 2 * kTabBytes;                           // 256 KiB, a multiple of 256
constexpr size_t kMidMax = 8u << 20;                                   // 8 MiB: 18 frames of 640x480 -> 308

static __device__ __align__(256) uint8_t g_static_scratch[kMidOffset + kMidMax];

struct StaticDeviceState {
  uint8_t* base = nullptr;       // device address of g_static_scratch on this GPU (cudaGetSymbolAddress), null = not set up
  int in_w = 0, in_h = 0, side = 0;   // geometry whose tables the device holds
  int ksize_w = 0, ksize_h = 0;
};
std::mutex g_static_mu;
StaticDeviceState g_static_dev[64];
int32_t g_static_host_tab[kTabInts];   // host staging of one pass's tables (guarded by g_static_mu)

// precompute_coeffs + normalize_coeffs_8bpc of one axis into `tab` = bounds [out, 2] | taps [out, ksize]; the same
// arithmetic as make_resize_axis above, written into fixed storage instead of vectors.  Returns ksize, 0 = too large.
int fill_static_table(int in_size, int out_size, int32_t* tab) {
  const double scale = static_cast<double>(in_size) / out_size;
  const double filterscale = scale < 1.0 ? 1.0 : scale;
  const double support = 2.0 * filterscale;
  const double ksize_d = ceil(support) * 2 + 1;
  if (ksize_d > kStaticMaxTaps || 2.0 * out_size + ksize_d * out_size > static_cast<double>(kTabInts)) return 0;
  const int ksize = static_cast<int>(ksize_d);
  int32_t* bounds = tab;
  int32_t* kk = tab + 2 * out_size;
  double w[kStaticMaxTaps];
  const double ss = 1.0 / filterscale;
  for (int xx = 0; xx < out_size; ++xx) {
    const double center = (xx + 0.5) * scale;
    int xmin = static_cast<int>(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = static_cast<int>(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    double ww = 0.0;
    for (int x = 0; x < xmax; ++x) {
      w[x] = bicubic_filter((x + xmin - center + 0.5) * ss);
      ww += w[x];
    }
    int32_t* row = kk + static_cast<size_t>(xx) * ksize;
    for (int x = 0; x < ksize; ++x) {
      const double k = x >= xmax ? 0.0 : (ww != 0.0 ? w[x] / ww : w[x]);
      row[x] = k < 0 ? static_cast<int>(-0.5 + k * (1 << kPrecisionBits)) : static_cast<int>(0.5 + k * (1 << kPrecisionBits));
    }
    bounds[2 * xx] = xmin;
    bounds[2 * xx + 1] = xmax;
  }
  return ksize;
}

}  // namespace

// Body of vvs_resize_frames_static (include/vitvs_b200.h); runs on the CURRENT device.  Device set-up happens inside the
// first call on each GPU (symbol address) and inside any call whose geometry differs from the previous one on that GPU
// (tables are computed on the host and copied into the array on `s`).  No cudaMalloc / cudaFree anywhere below.
int resize_frames_static(const uint8_t* frames, int n, int in_h, int in_w, int side, uint8_t* out, cudaStream_t s,
                         std::string* err) {
  if (!frames || !out) {
    set_error(err, "vvs_resize_frames_static: null argument");
    return VVS_ERR_INVALID;
  }
  if (n < 1 || in_h < 1 || in_w < 1 || side < 1) {
    set_error(err, "vvs_resize_frames_static: n, frame_height, frame_width and image_size must be positive");
    return VVS_ERR_INVALID;
  }
  if (side > kStaticMaxSide) {
    set_error(err, "vvs_resize_frames_static: image_size " + std::to_string(side) + " exceeds the built-in limit of " +
                       std::to_string(kStaticMaxSide));
    return VVS_ERR_INVALID;
  }
  const bool need_h = in_w != side, need_v = in_h != side;
  const size_t mid_bytes = static_cast<size_t>(n) * in_h * side * 3;
  if (need_h && need_v && mid_bytes > kMidMax) {
    set_error(err, "vvs_resize_frames_static: n * frame_height * image_size * 3 = " + std::to_string(mid_bytes) +
                       " bytes of intermediate image exceed the built-in scratch of " + std::to_string(kMidMax) +
                       " bytes; send fewer frames per call or use vvs_resize_frames");
    return VVS_ERR_INVALID;
  }
  if (static_cast<size_t>(n) * std::max(in_h, side) * std::max(in_w, side) * 3 > 0x7fffffffull) {
    set_error(err, "vvs_resize_frames_static: batch too large");
    return VVS_ERR_INVALID;
  }
  if (!need_h && !need_v) {
    VVS_CUDA_OK(cudaMemcpyAsync(out, frames, static_cast<size_t>(n) * side * side * 3, cudaMemcpyDeviceToDevice, s), err);
    return VVS_OK;
  }
  int dev = 0;
  VVS_CUDA_OK(cudaGetDevice(&dev), err);
  if (dev < 0 || dev >= 64) {
    set_error(err, "vvs_resize_frames_static: device ordinal outside [0, 64)");
    return VVS_ERR_INVALID;
  }
  std::lock_guard<std::mutex> lock(g_static_mu);
  StaticDeviceState& st = g_static_dev[dev];
  if (!st.base) {
    // g_static_scratch names a host-side shadow here; the kernels need the address the array has on THIS device
    uint8_t* p = nullptr;
    VVS_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch), err);
    st.base = p;
  }
  if (st.in_w != in_w || st.in_h != in_h || st.side != side) {
    st.in_w = st.in_h = st.side = 0;   // tables are in flux until both uploads are enqueued
    int kw = 0, kh = 0;
    if (need_h) {
      if (!(kw = fill_static_table(in_w, side, g_static_host_tab))) {
        set_error(err, "vvs_resize_frames_static: horizontal tap table exceeds the built-in limit (scale too large)");
        return VVS_ERR_INVALID;
      }
      // pageable source: the runtime has taken its copy of the host table by the time this returns
      VVS_CUDA_OK(cudaMemcpyAsync(st.base, g_static_host_tab, (2 + kw) * static_cast<size_t>(side) * sizeof(int32_t),
                                  cudaMemcpyHostToDevice, s), err);
    }
    if (need_v) {
      if (!(kh = fill_static_table(in_h, side, g_static_host_tab))) {
        set_error(err, "vvs_resize_frames_static: vertical tap table exceeds the built-in limit (scale too large)");
        return VVS_ERR_INVALID;
      }
      VVS_CUDA_OK(cudaMemcpyAsync(st.base + kTabBytes, g_static_host_tab,
                                  (2 + kh) * static_cast<size_t>(side) * sizeof(int32_t), cudaMemcpyHostToDevice, s), err);
    }
    st.ksize_w = kw;
    st.ksize_h = kh;
    st.in_w = in_w;
    st.in_h = in_h;
    st.side = side;
  }
  ResizeAxis ax_w, ax_h;
  ax_w.in_size = in_w;
  ax_w.out_size = side;
  ax_w.ksize = st.ksize_w;
  ax_w.bounds = reinterpret_cast<int32_t*>(st.base);
  ax_w.kk = ax_w.bounds + 2 * side;
  ax_h.in_size = in_h;
  ax_h.out_size = side;
  ax_h.ksize = st.ksize_h;
  ax_h.bounds = reinterpret_cast<int32_t*>(st.base + kTabBytes);
  ax_h.kk = ax_h.bounds + 2 * side;
  int rc = VVS_OK;
  const uint8_t* v_in = frames;
  if (need_h) {
    uint8_t* h_out = need_v ? st.base + kMidOffset : out;
    rc = launch_resize_h(frames, h_out, n * in_h, ax_w, s, err);
    v_in = h_out;
  }
  if (!rc && need_v) rc = launch_resize_v(v_in, out, n, side * 3, ax_h, s, err);
  return rc;
}

}  // namespace vvs
