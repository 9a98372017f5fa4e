/* vitvs_b200 - C ABI of the B200-native ViT-VS `dino` servoing hot path.
 *
 * What this replaces.  The reference (begbaj/ViT-VS) has no FFI on this path: `--method dino` selects
 * catkin_ws/ibvs/src/vitvs_v2.py (run_ibvs.sh:57-62) whose servo loop reaches the feature method only
 * through Controller.detect_features() (vitvs_v2.py:464-523) and the control law through
 * Controller.ibvs() (:588-632).  The entry points below are what a ctypes binding inside those two
 * methods would call; each cites the reference code it stands in for.  INTEGRATION.md shows the stub.
 *
 * Conventions: plain C types; the CALLER owns every buffer passed in (device pointers unless the name
 * says _host); the library owns only its workspace and its bf16 copies of the weights inside the
 * handle.  Every call returns 0 or a negative vvs_status (no exceptions cross the ABI) and
 * vvs_last_error() gives the message.  Calls are stream-ordered and asynchronous unless documented
 * otherwise.  Any number of handles per process and per device (every call switches to its handle's device for
 * its duration), re-entrant across handles, not thread-safe per handle.  A handle's workspace is shared by its
 * calls: a call on a different stream than the previous one first waits (on the device) for that call to finish.
 * There is no CPU fallback: on a machine without an sm_100 GPU vvs_create() fails.
 */
#ifndef VITVS_B200_H_
#define VITVS_B200_H_

#include <stdint.h>

#if defined(__GNUC__)
#define VVS_API __attribute__((visibility("default")))
#else
#define VVS_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vvs_ctx* vvs_handle;
typedef void* vvs_stream; /* cudaStream_t */

enum vvs_status {
  VVS_SUCCESS = 0,
  VVS_INVALID_ARGUMENT = -1,
  VVS_CUDA_ERROR = -2,
  VVS_BAD_STATE = -3,
  VVS_UNSUPPORTED_DEVICE = -4
};

enum vvs_dtype { VVS_F32 = 0, VVS_BF16 = 1 };
enum vvs_image_kind {
  VVS_IMAGE_U8_HWC = 0,  /* [n, S, S, 3] uint8 RGB, already resized (vitvs_v2.py:474-475; see vvs_resize_frames) */
  VVS_IMAGE_F32_NCHW = 1 /* [n, 3, S, S] fp32, output of ViTExtractor.preprocess_pil (dinov2_extractor.py:177-191) */
};

/* Mirrors the keys Controller.load_parameters reads (vitvs_v2.py:272-323, config/config.yaml) plus the
 * constants the reference hard-codes (vitvs_v2.py:72,84,250,484-485). */
typedef struct vvs_config {
  /* architecture: dinov2_vit{s,b,l,g}14 */
  int32_t dim;          /* 384 / 768 / 1024 / 1536 */
  int32_t heads;        /* dim / 64 */
  int32_t depth;        /* 12 / 12 / 24 / 40 */
  int32_t ffn_hidden;   /* 4*dim, or 4096 for the SwiGLU giant */
  int32_t swiglu;       /* 0 = Mlp(GELU), 1 = SwiGLUFFNFused */
  /* geometry */
  int32_t image_size;   /* dino_input_size (308) - square ViT input; image_size - 14 a multiple of patch_stride */
  int32_t max_images;   /* workspace sizing: images in flight (= 2 x frame pairs per call) */
  /* camera / control (config.yaml) */
  int32_t frame_width;  /* u_max 640 */
  int32_t frame_height; /* v_max 480 */
  double f_x, f_y;      /* 502.3016357421875 */
  double c_x, c_y;      /* u_max/2, v_max/2 (vitvs_v2.py:282-283) */
  double lambda_;       /* 0.03 */
  double ema_alpha;     /* 0.8 */
  int32_t num_pairs;    /* 24 (48 in the best-pose search, vitvs_v2.py:1158) */
  int32_t layer;        /* 11 (vitvs_v2.py:484) - the forward stops after this block */
  int32_t use_feature_binning; /* 1 */
  float same_image_threshold;  /* 0.99 (vitvs_v2.py:84) */
  int32_t max_match_tokens;    /* workspace for vvs_match on caller descriptors (0 => tokens of image_size) */
  int32_t max_match_dim;       /* (0 => 9*dim) */
  int32_t patch_stride;        /* ViTExtractor stride (dinov2_extractor.py:25,123-144): 14 = non-overlapping patches
                                  (vitvs_v2.py:250), a divisor of 14 = overlapping patches, 1 + (S-14)/stride tokens
                                  per side; 0 => 14 */
  double damping_mu;           /* Levenberg-Marquardt damping of the control law: v = -lambda (L^T L + mu I)^-1 L^T e.
                                  0 (default) = the reference's undamped pinv(L) (vitvs_v2.py:622); > 0 keeps the twist
                                  bounded when the matched points leave L badly conditioned.  No counterpart in the
                                  reference; appended last so that the layout of the fields above is unchanged. */
} vvs_config;

/* replaces: Controller.__init__ -> ViTExtractor(...) model construction (vitvs_v2.py:249-250,
 * dinov2_extractor.py:25-55).  Allocates the workspace; weights arrive through vvs_set_weight. */
VVS_API int vvs_create(const vvs_config* cfg, vvs_handle* out);
VVS_API int vvs_destroy(vvs_handle h);
/* message of the last failing call on this handle (h may be NULL: last failure of vvs_create or of a handle-free
 * call such as vvs_resize_frames_static). */
VVS_API const char* vvs_last_error(vvs_handle h);
VVS_API int vvs_num_sms(vvs_handle h);
/* hash of the CUDA sources this library was compiled from (vitvs_b200/build.py computes it; smoke() compares it with
 * the sources next to the .so, so a stale binary cannot pass for the tree it travels with). */
VVS_API const char* vvs_build_id(void);

/* replaces: the state_dict torch.hub.load() fills (dinov2_extractor.py:65-67).  `name` uses the upstream
 * DINOv2 parameter names: patch_embed.proj.{weight,bias}, cls_token, blocks.<i>.norm{1,2}.{weight,bias},
 * blocks.<i>.attn.{qkv,proj}.{weight,bias}, blocks.<i>.ls{1,2}.gamma, blocks.<i>.mlp.{fc1,fc2}.{weight,bias}
 * or blocks.<i>.mlp.{w12,w3}.{weight,bias}; plus `pos_embed_interp` [1+T, dim]: pos_embed already
 * interpolated to the patch grid (done once per geometry by the host mirror with the maths of
 * dinov2_extractor.py:94-120).  dev_ptr is a device pointer, copied/converted (bf16 for GEMM operands). */
VVS_API int vvs_set_weight(vvs_handle h, const char* name, const void* dev_ptr, int32_t dtype, const int64_t* shape,
                   int32_t rank, vvs_stream stream);
/* 0 when every parameter of blocks[0..layer] and the embeddings has been set.  On success the fp32 copies of
 * attn.qkv.weight and mlp.fc1 / w12.weight that were held for folding norm1 / norm2 into them are released.  Re-setting
 * a tensor that is folded into one of these matrices afterwards (norm1.*, attn.qkv.bias; norm2.*, mlp.fc1 / w12.bias)
 * therefore needs that block's matrix set again as well: until it is, this function and every forward call report the
 * block as incomplete.  Re-loading a whole state_dict in any order works as before. */
VVS_API int vvs_weights_ready(vvs_handle h);

/* replaces: ViTExtractor._extract_features with facet 'token' (dinov2_extractor.py:245-263): runs
 * patch-embed .. blocks[layer] and leaves the residual stream [n_img, 1+T, dim] fp32 in the workspace. */
VVS_API int vvs_forward_tokens(vvs_handle h, const void* images, int32_t image_kind, int32_t n_img, int32_t layer,
                       vvs_stream stream);
/* copy of that residual stream (hook output, dinov2_extractor.py:198-201) to tokens_out [n_img, 1+T, dim] */
VVS_API int vvs_get_tokens(vvs_handle h, float* tokens_out, int32_t n_img, vvs_stream stream);
/* replaces: ViTExtractor.extract_descriptors(..., facet='token', bin, include_cls=False)
 * (dinov2_extractor.py:313-337, _log_bin :265-311) on the tokens of the last vvs_forward_tokens call:
 * desc_out [n_img, T, (bin ? 9 : 1) * dim] fp32. */
VVS_API int vvs_descriptors(vvs_handle h, int32_t n_img, int32_t bin, float* desc_out, vvs_stream stream);

/* replaces: ViTExtractor._log_bin(x, hierarchy) (dinov2_extractor.py:265-311) for hierarchy >= 1 on the tokens of the
 * last vvs_forward_tokens call: level k = the 3x3 neighbourhood at distance 3^k of the 3^k x 3^k average-pooled token
 * map (in-grid elements only), centre kept at k = 0 only, borders replicate-clamped.  desc_out
 * [n_img, T, (1 + 8 * hierarchy) * dim] fp32; hierarchy 1 equals vvs_descriptors(bin = 1). */
VVS_API int vvs_descriptors_hierarchy(vvs_handle h, int32_t n_img, int32_t hierarchy, float* desc_out, vvs_stream stream);
/* replaces: ViTExtractor._extract_features with facet 'query' / 'key' / 'value' (dinov2_extractor.py:203-218): runs
 * blocks[0..layer-1], then norm1 and the fused qkv projection of blocks[layer] - the tensor the reference's hook on
 * `blocks[layer].attn` recomputes - and stops.  Invalidates vvs_get_tokens / vvs_descriptors until the next
 * vvs_forward_tokens. */
VVS_API int vvs_forward_facet(vvs_handle h, const void* images, int32_t image_kind, int32_t n_img, int32_t layer,
                      vvs_stream stream);
/* replaces: ViTExtractor.extract_descriptors(..., facet in {'query','key','value'}, bin, include_cls)
 * (dinov2_extractor.py:313-337) on the last vvs_forward_facet pass.  facet: 0 query, 1 key, 2 value.  desc_out
 * [n_img, T + include_cls, (bin ? 9 : 1) * dim] fp32, channels in the reference's order (head-dim major, head minor:
 * `permute(0, 2, 3, 1).flatten(-2)`); values carry the bf16 rounding of the projection's output. */
VVS_API int vvs_facet_descriptors(vvs_handle h, int32_t n_img, int32_t facet, int32_t bin, int32_t include_cls,
                          float* desc_out, vvs_stream stream);

typedef struct vvs_match_out { /* every pointer optional (NULL = not wanted), device memory */
  int32_t* nn1;          /* [n_pairs, T]   argmax_j S[i,j]         (vitvs_v2.py:80) */
  int32_t* nn2;          /* [n_pairs, T]   argmax_i S[i,j]         (vitvs_v2.py:81) */
  float* sim1;           /* [n_pairs, T]   max_j S[i,j] */
  uint8_t* candidate;    /* [n_pairs, T]   passes the cyclic-distance rule (:105-131) == mutual NN */
  float* mean_sim1;      /* [n_pairs] */
  int32_t* n_candidates; /* [n_pairs] */
  int32_t* n_selected;   /* [n_pairs]   min(num_pairs, n_candidates) */
  int32_t* flags;        /* [n_pairs]   bit 0: same-image branch taken (:84-101); bit 1: a row / column of S had
                            no comparable value (NaN descriptors) - its indices are reported as 0 */
  int32_t* sel_tok;      /* [n_pairs, num_pairs] chosen goal tokens, best similarity first, -1 padded */
  float* sel_sim;        /* [n_pairs, num_pairs] */
  int32_t* uv_star;      /* [n_pairs, num_pairs, 2] goal keypoints (u,v) in camera pixels (:511-553) */
  int32_t* uv;           /* [n_pairs, num_pairs, 2] current keypoints */
  /* INPUT (optional): reference-faithful selection (vitvs_v2.py:139-141, :90-91).  NULL = the num_pairs best
   * candidates by similarity (deterministic; the benchmark default).  Otherwise a priority list of goal tokens per
   * pair [n_pairs, T] int32 - e.g. a host torch.randperm(T) - and the first num_pairs of its entries that are
   * candidates are selected, in list order: a uniform random subset of the mutual-NN set like the reference's
   * perm[:num_to_select]; in the same-image branch exactly perm[:num_points]. */
  const int32_t* sel_order;
} vvs_match_out;

/* replaces: chunk_cosine_sim + find_correspondences_batch + calculate_uv (vitvs_v2.py:49-155, 511-553) on
 * caller-supplied descriptors desc1/desc2 [n_pairs, T, d] fp32 (rows = goal / current tokens), patch grid
 * grid_w x grid_w.  The T x T similarity matrix is never written.  Selection is deterministic (num_pairs
 * best by similarity); the full candidate set comes back in `candidate` for callers that want to draw
 * their own random subset like the reference does (:139-141). */
VVS_API int vvs_match(vvs_handle h, const float* desc1, const float* desc2, int32_t n_pairs, int32_t T, int32_t d,
              int32_t grid_w, int32_t num_pairs, const vvs_match_out* out, vvs_stream stream);
/* debugging / tests: dense S [n_pairs, T, T] fp32 of the same descriptors (vitvs_v2.py:49-56). */
VVS_API int vvs_similarity(vvs_handle h, const float* desc1, const float* desc2, int32_t n_pairs, int32_t T, int32_t d,
                   float* S_out, vvs_stream stream);

/* replaces: get_depth + transform_to_real_world + calculate_interaction_matrix + pinv + EMA
 * (vitvs_v2.py:566-586, 610-623, 634-659).  uv / uv_star [n_pairs, K, 2] int32 (u,v), depth
 * [n_pairs, frame_height, frame_width] uint16 mm.  twist [n_pairs, 6] fp64 (EMA-smoothed when ema_state
 * != NULL: ema_state [n_pairs, 6] fp64 and ema_init [n_pairs] int32 are caller-owned episode state);
 * twist_raw (optional) is the value before smoothing. */
VVS_API int vvs_ibvs(vvs_handle h, const int32_t* uv, const int32_t* uv_star, const uint16_t* depth, int32_t n_pairs,
             int32_t K, double* twist, double* twist_raw, double* ema_state, int32_t* ema_init, vvs_stream stream);

/* replaces: Controller.transform_to_real_world(s_uv, s_uv_star) (vitvs_v2.py:634-648) with the reference's own
 * arguments - no depth frame involved: uv / uv_star [n_points, 2] int32 -> xy / xy_star [n_points, 2] fp64
 * (either pair may be NULL). */
VVS_API int vvs_transform_to_real_world(vvs_handle h, const int32_t* uv, const int32_t* uv_star, int32_t n_points,
                                double* xy, double* xy_star, vvs_stream stream);
/* replaces: Controller.calculate_interaction_matrix(s_xy, Z) (vitvs_v2.py:650-659) with the reference's own
 * arguments: xy [n_points, 2], z [n_points] fp64 -> L [2 * n_points, 6] fp64, bit-identical to numpy's IEEE result. */
VVS_API int vvs_interaction_matrix(vvs_handle h, const double* xy, const double* z, int32_t n_points, double* L,
                           vvs_stream stream);

/* replaces, term by term: Controller.get_depth -> z [n_pairs,K]; transform_to_real_world -> xy / xy_star
 * [n_pairs,K,2]; calculate_interaction_matrix -> L [n_pairs,2K,6] (vitvs_v2.py:566-586, 634-659); any output
 * may be NULL.  twist [n_pairs,6] as in vvs_ibvs without smoothing. */
VVS_API int vvs_ibvs_terms(vvs_handle h, const int32_t* uv, const int32_t* uv_star, const uint16_t* depth, int32_t n_pairs,
                   int32_t K, double* twist, double* z, double* xy, double* xy_star, double* L, vvs_stream stream);

typedef struct vvs_pair_out { /* device memory; twist required, the rest optional */
  double* twist;          /* [n_pairs, 6]  v_c = -lambda * pinv(L) e   (vitvs_v2.py:622) */
  int32_t* n_selected;    /* [n_pairs] */
  float* mean_sim_sel;    /* [n_pairs]  mean similarity of the selected pairs (best-pose search, :1170) */
  vvs_match_out match;    /* per-stage outputs for parity checks */
} vvs_pair_out;

/* replaces: one Controller.ibvs() iteration per frame pair (vitvs_v2.py:588-623), batched:
 * goal / current [n_pairs, S, S, 3] uint8 (device), depth [n_pairs, H, W] uint16 (device). */
VVS_API int vvs_frame_pairs_to_twist(vvs_handle h, const uint8_t* goal, const uint8_t* current, const uint16_t* depth,
                             int32_t n_pairs, const vvs_pair_out* out, vvs_stream stream);
/* same with HOST buffers: H2D of the frames, the whole path, D2H of twist [n_pairs,6] / n_selected
 * [n_pairs] / mean_sim_sel [n_pairs] (last two optional), then a stream synchronise.  Pinned host memory
 * makes the copies asynchronous. */
VVS_API int vvs_frame_pairs_to_twist_host(vvs_handle h, const uint8_t* goal_host, const uint8_t* current_host,
                                  const uint16_t* depth_host, int32_t n_pairs, double* twist_host,
                                  int32_t* n_selected_host, float* mean_sim_sel_host, vvs_stream stream);

/* The two host-buffer calls without the final wait: they return once the copies and kernels are enqueued, and the next
 * call's H2D copies (second landing zone, own copy stream) then run under this call's compute.  Outputs are valid after
 * vvs_wait(h, 0) (the latest asynchronous call) / vvs_wait(h, 1) (the one before it); at most two may be in flight, and
 * the host buffers of a call must stay untouched until it has been waited for.  A servo loop over many cameras submits
 * batch i+1 while batch i computes (independent units). */
VVS_API int vvs_frame_pairs_to_twist_host_async(vvs_handle h, const uint8_t* goal_host, const uint8_t* current_host,
                                        const uint16_t* depth_host, int32_t n_pairs, double* twist_host,
                                        int32_t* n_selected_host, float* mean_sim_sel_host, vvs_stream stream);
VVS_API int vvs_frames_to_twist_host_async(vvs_handle h, const uint8_t* goal_frames_host, const uint8_t* current_frames_host,
                                   const uint16_t* depth_host, int32_t n_pairs, double* twist_host,
                                   int32_t* n_selected_host, float* mean_sim_sel_host, vvs_stream stream);
VVS_API int vvs_wait(vvs_handle h, int32_t calls_back);

/* Cached-goal mode - an EXTRA, never the benchmarked path.  The reference recomputes the goal view's descriptors on
 * every servo iteration (vitvs_v2.py:482-487) although the goal image is loaded once per episode (:255) and never
 * changes; vvs_set_goal runs n goal views through the backbone once and keeps their unit descriptors in the handle,
 * vvs_current_to_twist then forwards only the current views and matches view i against goal i (i < n_pairs <= n).
 * Results are bit-identical to vvs_frame_pairs_to_twist / vvs_frames_to_twist on the same (goal, current) pairs.
 * from_frames = 0: [n, S, S, 3] uint8 images as in vvs_frame_pairs_to_twist; 1: [n, frame_height, frame_width, 3]
 * camera frames as in vvs_frames_to_twist.  Device pointers, stream-ordered; `out` as in vvs_frame_pairs_to_twist.
 * Setting weights or calling any other forward entry leaves the cached goal descriptors untouched; re-set the goal
 * after changing weights. */
VVS_API int vvs_set_goal(vvs_handle h, const uint8_t* goal, int32_t n, int32_t from_frames, vvs_stream stream);
VVS_API int vvs_current_to_twist(vvs_handle h, const uint8_t* current, const uint16_t* depth_mm, int32_t n_pairs,
                         int32_t from_frames, const vvs_pair_out* out, vvs_stream stream);

/* replaces: PIL `Image.resize((S, S))` of the goal / current camera frame (vitvs_v2.py:474-475; Pillow default
 * filter = antialiased BICUBIC, 8-bit fixed point, horizontal pass then vertical pass).  frames
 * [n, frame_height, frame_width, 3] uint8 RGB (device) -> out [n, S, S, 3] uint8 (device), bit-identical to
 * Pillow 12.2 (integer arithmetic).  n <= max_images. */
VVS_API int vvs_resize_frames(vvs_handle h, const uint8_t* frames, int32_t n, uint8_t* out, vvs_stream stream);
/* No counterpart in the reference - an addition of this library for callers that cannot obtain device memory at call
 * time.  vvs_resize_frames needs a handle, and vvs_create / the first frame call of a handle cudaMalloc tap tables and
 * an intermediate image.  Two kinds of caller cannot afford that: (1) a servo node that shares its GPU with a framework
 * which has already claimed the device's memory for its own pool (torch's caching allocator after a large model, an
 * inference server with a fixed memory budget) - cudaMalloc then fails or succeeds depending on what else ran first;
 * (2) a camera-callback / control thread with a deadline (vitvs_v2.py:455-462 feed a 30 Hz loop) - cudaMalloc and
 * cudaFree take a driver lock and can wait for all work on the device, so one allocation stalls every other stream for
 * an unbounded time.  This entry gives them the same resize with NO allocation of any kind on the device: no handle,
 * no cudaMalloc / cudaFree, no library allocator.  The tap tables and the intermediate image live in a fixed array
 * inside the library image (8.25 MiB; every GPU that loads the library holds it until the process exits).
 *
 * frames [n, frame_height, frame_width, 3] uint8 RGB (device) -> out [n, image_size, image_size, 3] uint8 (device), on
 * the caller's CURRENT device, stream-ordered on `stream`; output bit-identical to vvs_resize_frames / Pillow.
 * Limits (VVS_INVALID_ARGUMENT with a message naming the limit, read it with vvs_last_error(NULL)):
 * image_size <= 1024; when both sides change, n * frame_height * image_size * 3 <= 8 MiB (18 frames of 640x480 -> 308);
 * image_size * (2 + taps per output sample) <= 32768 per pass, i.e. down-scaling up to about 25x at image_size 308.
 * Set-up: the first call on each GPU looks up the array's device address, and the first call with a new
 * (frame_height, frame_width, image_size) computes the tap tables on the host and copies them into the array on
 * `stream` (a pageable-memory copy, which may wait for `stream`).  Make that first call once during start-up, before
 * any stream capture; later calls with the same geometry only launch the two kernels.
 * The array is shared by every call on that GPU and keeps what the last call left in it; nothing needs to start at
 * zero (each call writes all it reads).  Calls must therefore not overlap on the GPU: issue them on one stream, or order
 * them with events - two calls running concurrently on different streams race on the intermediate image and on the
 * tables.  Host-side the call is thread-safe. */
VVS_API int vvs_resize_frames_static(const uint8_t* frames, int32_t n, int32_t frame_height, int32_t frame_width,
                             int32_t image_size, uint8_t* out, vvs_stream stream);
/* vvs_frame_pairs_to_twist on raw camera frames: goal / current [n_pairs, frame_height, frame_width, 3] uint8
 * (device); the resize of vitvs_v2.py:474-475 runs on the GPU ahead of the patch embedding. */
VVS_API int vvs_frames_to_twist(vvs_handle h, const uint8_t* goal_frames, const uint8_t* current_frames,
                        const uint16_t* depth, int32_t n_pairs, const vvs_pair_out* out, vvs_stream stream);
/* same with HOST buffers (the whole of one Controller.ibvs() iteration from the camera callbacks' buffers,
 * vitvs_v2.py:455-462 -> :622): H2D, resize, the whole path, D2H, stream synchronise.  Like
 * vvs_frame_pairs_to_twist_host the batch is processed in sub-batches so that the H2D copies of one run under
 * the compute of the previous one. */
VVS_API int vvs_frames_to_twist_host(vvs_handle h, const uint8_t* goal_frames_host, const uint8_t* current_frames_host,
                             const uint16_t* depth_host, int32_t n_pairs, double* twist_host,
                             int32_t* n_selected_host, float* mean_sim_sel_host, vvs_stream stream);
/* number of kernels the last vvs_frame_pairs_to_twist* / vvs_frames_to_twist* call launched (bench.py gpu_launches). */
VVS_API int vvs_last_launch_count(vvs_handle h);
/* replaces: `controller.num_pairs = 48` / restore around the best-pose search (vitvs_v2.py:1156-1158, 1179): changes
 * the K of the fused frame-pair entry points, 1 <= num_pairs <= max(num_pairs at creation, 64). */
VVS_API int vvs_set_num_pairs(vvs_handle h, int32_t num_pairs);
/* 1 when the fused frame-pair path matches 3x3-binned descriptors as a 9-point stencil over the token Gram matrix
 * (the [T, 9*dim] descriptors of dinov2_extractor.py:265-311 are then never written and the similarity product of
 * vitvs_v2.py:49-56 runs with K = dim): same cosine similarities up to fp32 summation order.  0 = the binned
 * descriptors are materialised: the default (the stencil form is opt-in through VITVS_GRAM=1 in the environment
 * at vvs_create time; it trades 0.3 points of arg-max agreement with the fp32 reference for 1.3 % of a step). */
VVS_API int vvs_uses_gram_stencil(vvs_handle h);

/* ---- the step AFTER the path: twist -> next camera frame, headless ------------------------------- */
/* replaces: the simulated RealSense D435 of the reference's Gazebo world looking at its textured poster - a
 * plane_w x plane_h metre plane (models/viso/meshes/resized.dae:113: 0.8 x 0.6) in z = 0 of the world frame,
 * centred on the origin, texture [tex_h, tex_w, 3] uint8 (x right, y down) - through the pinhole camera of the
 * handle's config (frame_width x frame_height, f_x, f_y, c_x, c_y).  cam_to_world [n, 12] fp64 (device): rotation
 * R row-major then translation t; a camera at (0, 0, -h) with R = I looks along +z at the poster from h metres.
 * rgb_out [n, frame_height, frame_width, 3] uint8 (bilinear texture fetch, grey 128 off the poster), depth_out
 * [n, frame_height, frame_width] uint16 millimetres, 0 outside [0.2, 10] m (RealSensePlugin.cpp:252-262).  Device
 * buffers only: no window, no graphics API.  fp64, bit-identical to the CPU restatement in oracle/render.py. */
VVS_API int vvs_render_plane(vvs_handle h, const uint8_t* texture, int32_t tex_h, int32_t tex_w, double plane_w,
                     double plane_h, const double* cam_to_world, int32_t n, uint8_t* rgb_out, uint16_t* depth_out,
                     vvs_stream stream);

/* ---- per-kernel-class device timing (CUDA events on the launching stream around every launch) -------- */
enum vvs_kernel_class {
  VVS_K_OTHER = 0, VVS_K_IM2COL = 1, VVS_K_PATCH_GEMM = 2, VVS_K_LAYERNORM = 3, VVS_K_QKV_GEMM = 4,
  VVS_K_ATTENTION = 5, VVS_K_PROJ_GEMM = 6, VVS_K_FFN1_GEMM = 7, VVS_K_FFN2_GEMM = 8, VVS_K_DESCRIPTORS = 9,
  VVS_K_SIM_GEMM = 10, VVS_K_SELECT = 11, VVS_K_IBVS = 12, VVS_K_RESIZE = 13, VVS_K_STENCIL = 14, VVS_K_COUNT = 15
};
VVS_API int vvs_profile_enable(vvs_handle h, int32_t on);
/* Step accounting.  Re-captures the two CUDA graphs of vvs_frame_pairs_to_twist for n_pairs with a trace slot in every
 * kernel (first / last warp stamp %globaltimer), replays them `replays` times back to back on the inputs the previous
 * call left in the workspace, and returns for the LAST replay, per kernel in launch order: its vvs_kernel_class and its
 * device begin / end time in ns.  end[i] - begin[i] is the kernel, begin[i+1] - end[i] the gap behind it (launch latency
 * of the next graph node, plus the two memset nodes ahead of the similarity GEMM).  Synchronises. */
VVS_API int vvs_trace_step(vvs_handle h, int32_t n_pairs, int32_t replays, int32_t* classes, uint64_t* begin_ns,
                   uint64_t* end_ns, int32_t max_nodes, int32_t* n_nodes, vvs_stream stream);
/* sums (ms) and launch counts per vvs_kernel_class since the last read; synchronises on the events. */
VVS_API int vvs_profile_read(vvs_handle h, double* total_ms, int32_t* launches, int32_t n_classes);

/* ---- kernel-level entry points (unit tests, per-kernel benchmarks; device pointers) ---------------- */
/* C = A[M,K] * B[N,K]^T with one of the fused epilogues (see csrc/vvs_internal.h Epilogue). */
VVS_API int vvs_dbg_gemm(vvs_handle h, const void* A_bf16, const void* B_bf16, int32_t M, int32_t N, int32_t K,
                 int32_t epilogue, const float* bias, const float* gamma, float* x_f32, void* out_bf16, int32_t ldo,
                 int32_t q_cols, float q_scale, const float* pos, int32_t T, vvs_stream stream);
/* The same with a folded-LayerNorm epilogue (csrc/vvs_internal.h EPI_*_LN_*; dino_patch/block.py:90-95 norm1 / norm2):
 * producers (EPI_RESID_LN_F32, EPI_PATCH_LN_F32) also write xb_out = bf16(new x) [M, ldo] and ln_part_out [M, N/64, 2]
 * = (sum, sum of squares) of every 64-column group of the new rows; consumers (EPI_BIAS_LN_BF16, EPI_GELU_LN_BF16,
 * EPI_SWIGLU_LN_BF16) read ln_part_in [M, K/64, 2] and colsum [N] and compute rstd * (acc - mean * colsum) + bias.
 * EPI_PATCH_LN_F32 fetches its positional rows in 32-row boxes: `pos` must hold T + 1 + 31 rows, the last 31 repeating
 * the first ones. */
VVS_API int vvs_dbg_gemm_ln(vvs_handle h, const void* A_bf16, const void* B_bf16, int32_t M, int32_t N, int32_t K,
                    int32_t epilogue, const float* bias, const float* gamma, float* x_f32, void* out_bf16, int32_t ldo,
                    int32_t q_cols, float q_scale, const float* pos, int32_t T, float* ln_part_out, void* xb_out_bf16,
                    const float* ln_part_in, const float* colsum, float ln_eps, vvs_stream stream);
VVS_API int vvs_dbg_attention(vvs_handle h, const void* qkv_bf16, void* out_bf16, int32_t n_img, int32_t n_tok,
                      int32_t heads, vvs_stream stream);
VVS_API int vvs_dbg_layernorm(vvs_handle h, const float* x, const float* w, const float* b, void* out_bf16, int32_t rows,
                      int32_t dim, vvs_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* VITVS_B200_H_ */
