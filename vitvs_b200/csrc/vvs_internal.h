// Internal C++ declarations shared by the vitvs_b200 translation units (not part of the C ABI).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>
#include <string>

namespace vvs {

typedef __nv_bfloat16 bf16;

// ---- error plumbing: no exceptions cross the C ABI -------------------------------------------
enum Status : int {
  VVS_OK = 0,
  VVS_ERR_INVALID = -1,    // bad argument / unsupported shape
  VVS_ERR_CUDA = -2,       // a CUDA runtime / driver call failed
  VVS_ERR_STATE = -3,      // weights missing, workspace too small, ...
  VVS_ERR_UNSUPPORTED = -4 // device is not sm_100
};
void set_error(std::string* slot, const std::string& msg);

#define VVS_CUDA_OK(expr, errslot)                                                                      \
  do {                                                                                                  \
    cudaError_t _e = (expr);                                                                            \
    if (_e != cudaSuccess) {                                                                            \
      vvs::set_error(errslot, std::string(#expr) + ": " + cudaGetErrorString(_e));                      \
      return vvs::VVS_ERR_CUDA;                                                                         \
    }                                                                                                   \
  } while (0)

// ---- per-device one-time setup ---------------------------------------------------------------------
// cudaFuncSetAttribute / occupancy queries are per DEVICE, not per process: a call site keeps one of these and runs
// its setup the first time it executes on each device (handles on several GPUs in one process, include/vitvs_b200.h:13).
// The lock is held across the setup so that a second host thread cannot launch before the attribute is in place.
struct PerDeviceOnce {
  std::mutex mu;
  unsigned long long done = 0;   // bit d: setup has run on device d
  template <typename F>
  cudaError_t run(F&& setup) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const unsigned long long bit = 1ull << (dev & 63);
    std::lock_guard<std::mutex> lock(mu);
    if (done & bit) return cudaSuccess;
    e = setup();
    if (e == cudaSuccess) done |= bit;
    return e;
  }
};

// ---- step tracing ------------------------------------------------------------------------------------
// Slot ([begin, end] device words) the NEXT kernel launched by this host thread stamps its %globaltimer span into
// (ptx.cuh trace_begin / trace_end); null outside vvs_trace_step.  Set by api.cu around every launch.
extern thread_local unsigned long long* tl_trace_slot;

// ---- launches ------------------------------------------------------------------------------------
// <<< >>> with an optional programmatic-dependent-launch attribute (ptx.cuh: pdl_wait / pdl_launch_dependents)
// programmatic dependent launch for a forward pass over `rows` token rows of width `dim`: on for every shape unless
// VITVS_PDL=0 (measured to help small batches most and never to lose, see api.cu); `rows` / `dim` are kept in the
// signature for a shape-dependent policy but are not consulted today
bool pdl_for_rows(long long rows, int dim);
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, bool pdl,
                                 Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---- TMA tensor maps ---------------------------------------------------------------------------
// bf16 tensor [batch, rows, cols] (cols contiguous, row pitch ld elements, batch pitch in elements),
// box = {64 cols, box_rows, 1}, SWIZZLE_128B, zero fill out of bounds.
int make_tmap_bf16(CUtensorMap* out, const void* base, uint64_t cols, uint64_t rows, uint64_t batch, uint64_t ld,
                   uint64_t batch_pitch, uint32_t box_rows, std::string* err);

// ---- tcgen05 GEMM family -----------------------------------------------------------------------
enum Epilogue : int {
  EPI_BIAS_BF16 = 0,   // out_bf16 = (acc + bias) * (col < q_cols ? q_scale : 1)              (qkv)
  EPI_GELU_BF16 = 1,   // out_bf16 = gelu_erf(acc + bias)                                      (mlp.fc1)
  EPI_RESID_F32 = 2,   // x_f32   += gamma * (acc + bias)                                      (attn.proj, mlp.fc2 / w3)
  EPI_PATCH_F32 = 3,   // x_f32[r] = acc + bias + pos'[r % (T+1)], A has a zero row per CLS slot        (patch embed)
  EPI_SWIGLU_BF16 = 4, // out_bf16 = silu(acc1 + b1) * (acc2 + b2), weight rows tile-interleaved (mlp.w12)
  EPI_SIM = 5,         // packed row / column arg-max of the similarity tile, S never stored    (matching)
  EPI_STORE_F32 = 6,   // out_f32 = acc  (plain fp32 store; unit tests / similarity dump)
  // LayerNorm folded into the GEMMs on either side of it (no LayerNorm pass over the residual stream):
  // PRODUCERS of the residual stream also emit its bf16 copy and per-row (sum, sum of squares) partials, one per
  // 64-column group; CONSUMERS multiply the RAW bf16 rows by gamma-folded weights and normalise in the epilogue:
  //   LN(x) W^T + b = rstd * (x W'^T - mean * colsum(W')) + b',  W' = W diag(gamma_ln),  b' = b + W beta_ln
  EPI_RESID_LN_F32 = 7,    // EPI_RESID_F32 computed in registers (x loaded, stored) + xb_out + ln_part_out
  EPI_PATCH_LN_F32 = 8,    // EPI_PATCH_F32 + xb_out + ln_part_out
  EPI_BIAS_LN_BF16 = 9,    // EPI_BIAS_BF16 on rstd * (acc - mean * colsum) instead of acc
  EPI_GELU_LN_BF16 = 10,   // EPI_GELU_BF16 likewise
  EPI_SWIGLU_LN_BF16 = 11  // EPI_SWIGLU_BF16 likewise
};

struct GemmArgs {
  // C[b] = A[b] (M x K, K-major) * B[b or 0] (N x K, K-major)^T
  const bf16* A;
  const bf16* B;
  int M, N, K;
  int lda, ldb;
  int batch;             // independent problems
  long long a_batch_pitch, b_batch_pitch;   // elements; b_batch_pitch == 0 => B shared by all batches
  int epilogue;
  // epilogue operands (unused ones may be null)
  const float* bias;     // [N]
  const float* gamma;    // [N]
  float* x_f32;          // EPI_RESID/PATCH/STORE_F32 destination, pitch ldo
  bf16* out_bf16;        // bf16 destination, pitch ldo
  int ldo;
  int q_cols;
  float q_scale;
  const float* pos;      // [T+1, ldo] (EPI_PATCH_F32): pos' with the CLS row folded in (launch_pos_prime)
  int T;                 // patches per image (EPI_PATCH_F32)
  int pos_rows;          // rows of `pos` (EPI_PATCH_LN_F32 fetches 32-row boxes: T + 1 rows extended periodically by 31)
  unsigned long long* row_best;   // [batch, best_pitch] (EPI_SIM)
  unsigned long long* col_best;   // [batch, best_pitch]
  int best_pitch;
  int pin_row0;          // EPI_RESID_F32: rows >= pin_row0 of x are reduce-added with the L2 evict-last hint (<= 0: no hint)
  int pdl;               // launch with the programmatic-dependent-launch attribute (predecessor is a PDL-aware kernel)
  int b_static;          // B is a weight matrix (never written by the preceding kernel): under PDL its first boxes are
                         // requested before the grid dependency resolves
  // folded LayerNorm (EPI_*_LN_*): partials are float2 (sum, sum of squares) per row and 64-column group
  float2* ln_part_out;        // producers: [M, N / 64] statistics of the NEW residual rows
  bf16* xb_out;               // producers: bf16 copy of the new residual rows, pitch ldo
  const float2* ln_part_in;   // consumers: [M, K / 64] statistics of A's rows
  const float* colsum;        // consumers: [N] sum over k of B[n, k] (of the bf16 values the tensor core multiplies)
  float ln_eps;
};
int launch_gemm(const GemmArgs& a, int num_sms, cudaStream_t stream, std::string* err);
int gemm_block_n(int N);   // the N tile the launcher picks for this N (128 or 256)

// ---- attention ---------------------------------------------------------------------------------
// qkv: [n_img * n_tok, 3*D] bf16 (q pre-scaled by hd^-0.5), out: [n_img * n_tok, D] bf16
// attention_ts.cu: P stays in TMEM (TS MMA), one 128-row query tile per CTA and four CTAs per SM
int launch_attention(const bf16* qkv, bf16* out, int n_img, int n_tok, int heads, cudaStream_t stream,
                     std::string* err, bool pdl = false, int num_sms = 148);

// ---- camera frame -> ViT input: Pillow-exact 8-bit bicubic resize (resize.cu) ---------------------
struct ResizeAxis {          // one separable pass: per output index the first tap, the tap count and the 22-bit taps
  int in_size = 0, out_size = 0, ksize = 0;
  int32_t* bounds = nullptr; // [out_size, 2] device
  int32_t* kk = nullptr;     // [out_size, ksize] device
};
int make_resize_axis(ResizeAxis* ax, int in_size, int out_size, std::string* err);
void free_resize_axis(ResizeAxis* ax);
// in [n_rows, ax.in_size, 3] -> out [n_rows, ax.out_size, 3]
int launch_resize_h(const uint8_t* in, uint8_t* out, int n_rows, const ResizeAxis& ax, cudaStream_t s, std::string* err);
// in [n_img, ax.in_size, row_bytes] -> out [n_img, ax.out_size, row_bytes]
int launch_resize_v(const uint8_t* in, uint8_t* out, int n_img, int row_bytes, const ResizeAxis& ax, cudaStream_t s,
                    std::string* err);
// both passes on the current device with tables and intermediate image in the library's built-in device array: no
// device allocation, no handle (resize.cu; vvs_resize_frames_static)
int resize_frames_static(const uint8_t* frames, int n, int in_h, int in_w, int side, uint8_t* out, cudaStream_t s,
                         std::string* err);

// ---- bandwidth-bound kernels -------------------------------------------------------------------
int launch_im2col_u8(const uint8_t* img, bf16* out, int n_img, int size, int kpad, int grid, int stride, cudaStream_t s,
                     std::string* err);
int launch_im2col_f32(const float* img, bf16* out, int n_img, int size, int kpad, int grid, int stride, cudaStream_t s,
                      std::string* err);
int launch_pos_prime(const float* pos, const float* cls, const float* bias, float* out, int n_tok, int D, cudaStream_t s,
                     std::string* err);
int launch_cls_rows(float* x, const float* cls, const float* pos, int n_img, int n_tok, int D, cudaStream_t s,
                    std::string* err);
// rows >= pin_row0 are read with the L2 evict-last hint, the others evict-first (pin_row0 >= rows: no hints)
int launch_layernorm(const float* x, const float* w, const float* b, bf16* out, int rows, int D, float eps, int pin_row0,
                     cudaStream_t s, std::string* err, bool pdl = false);
// residual stream [n_img, 1+T, D] -> descriptors (CLS dropped, optional 3x3 log-binning)
int launch_descriptors_f32(const float* x, float* out, int n_img, int hp, int wp, int D, int bin, cudaStream_t s,
                           std::string* err);
// residual stream -> log-binned descriptors with `hierarchy` levels, [n_img, T, (1 + 8 h) D] fp32
int launch_descriptors_hier_f32(const float* x, float* out, int n_img, int hp, int wp, int D, int hierarchy,
                                cudaStream_t s, std::string* err);
// qkv [n_img * (1+T), 3*D] bf16 (columns [facet][head][64]) -> key / query / value descriptors fp32
// [n_img, T (+1 with CLS), parts * D] in the reference's channel order (d major, head minor)
int launch_facet_descriptors_f32(const bf16* qkv, float* out, int n_img, int hp, int wp, int D, int heads, int facet,
                                 int bin, int include_cls, cudaStream_t s, std::string* err);
// zero_a / zero_b: optional arrays of zero_n (<= n_img * hp * wp) 64-bit words cleared by the same launch (the matching
// stage's packed arg-max keys)
int launch_descriptors_unit_bf16(const float* x, bf16* out, float* sqnorm_scratch, int n_img, int hp, int wp, int D,
                                 int bin, cudaStream_t s, std::string* err, unsigned long long* zero_a = nullptr,
                                 unsigned long long* zero_b = nullptr, long long zero_n = 0);
// arbitrary fp32 rows [rows, d] -> unit-norm bf16 rows [rows, dpad] (zero padded)
int launch_normalize_rows_bf16(const float* in, bf16* out, int rows, int d, int dpad, cudaStream_t s,
                               std::string* err);

// ---- Gram-stencil matching on 3x3-binned descriptors without materialising them (gram_stencil.cu) ----
int launch_tokens_bf16_sqnorm(const float* x, bf16* tb, float* n2, int n_img, int T, int D, cudaStream_t s,
                              std::string* err);
int launch_binned_inv_norm(const float* n2, float* inv, int n_img, int hp, int wp, cudaStream_t s, std::string* err);
bool gram_stencil_supports(int T);
int launch_gram_stencil_argmax(const float* G, const float* inv, unsigned long long* row_best,
                               unsigned long long* col_best, int n_pairs, int hp, int wp, int best_pitch, cudaStream_t s,
                               std::string* err);

// ---- matching tail + control law ----------------------------------------------------------------
struct Camera {
  int width, height;      // u_max, v_max
  double fx, fy, cx, cy;
};
// headless planar-scene RGB-D camera (render.cu): poses [n, 12] = camera-to-world R (row-major) + t, device memory
int launch_render_plane(const uint8_t* tex, int th, int tw, double plane_w, double plane_h, const double* poses, int n,
                        const Camera& cam, uint8_t* rgb, uint16_t* depth, cudaStream_t s, std::string* err);

struct SelectArgs {
  const unsigned long long* row_best;
  const unsigned long long* col_best;
  int best_pitch;
  int n_pairs, T, grid_w;
  int K;                  // num_pairs
  float same_thresh;      // 0.99
  int image_size;         // S
  Camera cam;
  // outputs
  int32_t* nn1;           // [n_pairs, T]
  int32_t* nn2;           // [n_pairs, T]
  float* sim1;            // [n_pairs, T]
  uint8_t* candidate;     // [n_pairs, T]
  float* mean_sim1;       // [n_pairs]
  int32_t* n_candidates;  // [n_pairs]
  int32_t* n_selected;    // [n_pairs]
  int32_t* flags;         // [n_pairs] bit0 = same-image branch
  int32_t* sel_tok;       // [n_pairs, K] chosen goal tokens (-1 padded)
  float* sel_sim;         // [n_pairs, K]
  int32_t* uv_star;       // [n_pairs, K, 2]
  int32_t* uv;            // [n_pairs, K, 2]
  const int32_t* sel_order;   // [n_pairs, T] optional priority list of goal tokens (reference-faithful selection)
  int keys_in_smem;           // set by launch_select: the T selection keys fit the kernel's dynamic shared memory
};
int launch_select(const SelectArgs& a, cudaStream_t s, std::string* err);

struct IbvsArgs {
  const int32_t* uv;        // [n_pairs, K, 2] current (u,v)
  const int32_t* uv_star;   // [n_pairs, K, 2] goal
  const uint16_t* depth;    // [n_pairs, height, width] millimetres
  int n_pairs, K;
  Camera cam;
  double lambda;
  double mu;                // damping added to the diagonal of L^T L (0 = the reference's pinv)
  double ema_alpha;         // used when ema_state != null
  double* ema_state;        // [n_pairs, 6] or null
  int32_t* ema_init;        // [n_pairs] 0 = first call copies
  double* twist;            // [n_pairs, 6] (after EMA when enabled)
  double* twist_raw;        // [n_pairs, 6] or null (before EMA)
  // optional per-point terms (Controller.get_depth / transform_to_real_world / calculate_interaction_matrix)
  double* z_out;            // [n_pairs, K]
  double* xy_out;           // [n_pairs, K, 2]
  double* xy_star_out;      // [n_pairs, K, 2]
  double* L_out;            // [n_pairs, 2K, 6]
};
int launch_ibvs(const IbvsArgs& a, cudaStream_t s, std::string* err);
// transform_to_real_world / calculate_interaction_matrix with the reference's own argument lists (n points)
int launch_points(const int32_t* uv, const int32_t* uv_star, int n, const Camera& cam, double* xy, double* xy_star,
                  cudaStream_t s, std::string* err);
int launch_interaction(const double* xy, const double* z, int n, double* L, cudaStream_t s, std::string* err);

}  // namespace vvs
