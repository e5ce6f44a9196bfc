/* gsr.h — C ABI of the B200-native Gaussian-splat forward renderer (libgsr.so).
 *
 * The reference (M-Gjerde/GaussianSplattingViewer) has no FFI: its render path sits behind the Python
 * class interface GaussianRenderBase (/root/reference/renderer_ogl.py:75-101) and is implemented by
 * PyOpenGL shaders (renderer_ogl.py:104-179) or the diff_gaussian_rasterization extension
 * (renderer_cuda.py:101-258).  This header is the boundary a maintainer binds UNDER that interface
 * (ctypes).  Every entry point names the reference call it stands in for.
 *
 * Conventions
 *   - every function returns 0 on success, a negative gsr_status otherwise; gsr_last_error() gives a
 *     thread-local message (the Python layer raises RuntimeError, like renderer_cuda.py:121-125).
 *   - matrices are 4x4 float32, row-major MATHS layout (M[row][col], translation in M[.][3]): what
 *     util.Camera.get_view_matrix()/get_project_matrix() return as NumPy arrays.
 *   - images are float32 RGBA, HWC, row 0 = TOP of the displayed image (the layout
 *     renderer_cuda.py:226-228 produces; the GL read-back flip of main.py:912 is already applied);
 *     channel 3 = accumulated coverage 1 - prod(1 - alpha).
 *   - a handle is bound to one CUDA device and is not thread-safe; `stream` is a cudaStream_t
 *     (NULL = legacy default stream).  No call synchronises the device unless it says so.
 *   - no CPU fallback exists anywhere behind this header.
 */
#ifndef GSR_H
#define GSR_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gsr_renderer gsr_t;

typedef enum {
    GSR_OK = 0,
    GSR_EINVAL = -1,   /* bad argument / call order */
    GSR_ECUDA = -2,    /* a CUDA runtime call failed */
    GSR_ENOMEM = -3,   /* device allocation failed */
    GSR_EOVERFLOW = -4 /* tile-instance buffer overflowed and could not be grown */
} gsr_status;

/* gsr_set_params flags */
#define GSR_FLAG_CLAMP_UNORM 1u /* clamp source colour to [0,1] (RGB8 target semantics, main.py:852); off = float target (R32F, main.py:870) */
#define GSR_FLAG_KEYS64 2u      /* sort duplicated 64-bit (tile<<32|depth) keys in one radix sort instead of the two-level depth-then-tile sort; identical result */
#define GSR_FLAG_DEBUG 4u       /* keep per-stage intermediates for gsr_debug_fetch */
#define GSR_FLAG_OFFICIAL 8u    /* CUDARenderer semantics (renderer_cuda.py:181-228): the forward pass of the out-of-tree,
                                 * unpinned diff_gaussian_rasterization package restated from its published source —
                                 * cull view z <= 0.2, radius ceil(3 sqrt(lambda_max)) tile squares, integer pixel
                                 * coordinates, colour clamped at 0 from below, full SH degree (render_mod ignored),
                                 * stop before T (1 - alpha) < 1e-4, alpha channel 1.  gsr_set_camera takes the same GL
                                 * matrices; the z-forward view and P * view are derived like update_camera_pose does.
                                 * With GSR_FLAG_DEBUG: VSOUT = visible, p_proj[3], radius x2, conic[3], color[3], opacity,
                                 * depth; SETUP = pixel x, y, depth, ok, tile rect (inclusive, rows top-down). */

/* gsr_debug_fetch selectors (GSR_FLAG_DEBUG must be set before the frame is rendered) */
#define GSR_DBG_VSOUT 0      /* N x 16 f32: visible, ndc[3], half[2], conic[3], color[3], alpha, depth, pad[2] (layout of GsoVsOut) */
#define GSR_DBG_SETUP 1      /* N x 8  32-bit: cx, cy, depth (f32), ok, imin, imax, jmin, jmax (i32) (layout of GsoSetup) */
#define GSR_DBG_INST_KEYS 2  /* D x u64: (tile<<32 | depth bits) of the sorted tile-instance list */
#define GSR_DBG_INST_VALS 3  /* D x u32: Gaussian index of each sorted instance */
#define GSR_DBG_RANGES 4     /* tiles x 2 i32: [start, end) into the instance list per tile */

typedef struct {
    long long n_gaussians; /* N */
    long long n_visible;   /* Gaussians that survive cull + clip + empty-rect */
    long long n_instances; /* D = sum of tiles touched */
    long long capacity;    /* instance capacity of the work buffers */
    int overflow;          /* 1 if D exceeded capacity in the last frame (image is incomplete) */
    int tiles_x, tiles_y;
    float ms_preprocess, ms_sort_depth, ms_duplicate, ms_sort_tiles, ms_ranges, ms_blend, ms_total; /* last gsr_render_timed */
} gsr_stats;

/* ---- life cycle -------------------------------------------------------------------------------- */
/* replaces OpenGLRenderer.__init__ / CUDARenderer.__init__ (renderer_ogl.py:105, renderer_cuda.py:102) */
int gsr_create(int device, gsr_t** out);
int gsr_destroy(gsr_t* h);
const char* gsr_last_error(void);
int gsr_version(void);

/* ---- scene ------------------------------------------------------------------------------------- */
/* replaces update_gaussian_data (renderer_ogl.py:132-137: SSBO upload of GaussianData.flat();
 * renderer_cuda.py:134-136 + gaus_cuda_from_cpu :87-98).  Arrays are GaussianData's fields:
 * xyz (N,3) rot (N,4 wxyz) scale (N,3) opacity (N,1) sh (N,sh_dim), C-contiguous float32, host
 * pointers (src_is_device = 0) or device pointers on the handle's device (1).  Copies; the caller
 * keeps ownership.  Synchronises the device. */
int gsr_set_gaussians(gsr_t* h, const float* xyz, const float* rot, const float* scale, const float* opacity,
                      const float* sh, long long n, int sh_dim, int src_is_device);

/* load_ply on the device (util_gau.py:63-125).  The host parses the PLY header (properties are looked up
 * by NAME, :66-112) into byte offsets; the vertex rows are decoded by one kernel straight into the
 * resident scene buffers: de-interleave, rot / ||rot|| and exp(scale) in float64 rounded to float32
 * (:116-119), float32 sigmoid (:120-121, NumPy's float32 exp restated), f_rest channel-major ->
 * coefficient-major (:99-100).  All 59 properties must be `property float`.
 * offset[]: x y z | f_dc_0..2 | f_rest_0..44 | opacity | scale_0..2 | rot_0..3.
 * rows: n * row_bytes bytes, HOST (pinned memory overlaps the upload with the decode, 64 MB chunks) or
 * DEVICE (16-byte aligned, readable up to the next 16-byte boundary).
 * bbox: xyz.min(axis=0), xyz.max(axis=0) (:81-83); center: xyz.mean(axis=0) with NumPy's float32 running
 * sum (:85; a serial ~2 ns/vertex kernel — pass center = NULL to skip it).  Replaces the scene like
 * gsr_set_gaussians (sh_dim 48).  Synchronises. */
typedef struct {
    int32_t row_bytes;
    int32_t big_endian;
    int32_t offset[59];
} gsr_ply_layout;
int gsr_load_ply_rows(gsr_t* h, const void* rows, long long n_vertices, const gsr_ply_layout* layout,
                      int src_is_device, float bbox[6], float center[3], void* stream);
/* inverse of gsr_set_gaussians: copies the resident scene into GaussianData's five host arrays (any may be
 * NULL); what OpenGLRenderer keeps as self.gaussians (renderer_ogl.py:133).  Synchronises. */
int gsr_get_gaussians(gsr_t* h, float* xyz, float* rot, float* scale, float* opacity, float* sh);

/* replaces set_render_reso / glViewport (renderer_ogl.py:157-158, renderer_cuda.py:175-179) */
int gsr_set_resolution(gsr_t* h, int width, int height);

/* replaces update_camera_pose + update_camera_intrin (renderer_ogl.py:160-173): uniforms view_matrix,
 * cam_pos, projection_matrix, hfovxy_focal = (htanx, htany, focal) */
int gsr_set_camera(gsr_t* h, const float view[16], const float proj[16], const float campos[3], float htanx,
                   float htany, float focal);

/* replaces set_scale_modifier / set_render_mod (renderer_ogl.py:148-152).  render_mod: 0..3 SH degree,
 * -1 disparity ("depth", 1.2x scale), -2 billboard, -3 flat ball, -4 gaussian ball (gau_frag.glsl:7) */
int gsr_set_params(gsr_t* h, float scale_modifier, int render_mod, unsigned flags);

/* Instance capacity of the work buffers (0 = automatic: max(4M, 16 N), x pixels/1080p up to 4x, grown on overflow by the
 * *_host entry points). */
int gsr_set_capacity(gsr_t* h, long long max_instances);

/* ---- draw -------------------------------------------------------------------------------------- */
/* replaces draw() (renderer_ogl.py:175-179 glDrawElementsInstanced; renderer_cuda.py:205-228).
 * Asynchronous; out_rgba is a DEVICE buffer of H*W*4 floats. */
int gsr_render(gsr_t* h, float* out_rgba, void* stream);
/* same, with CUDA events around every stage; synchronises; fills the ms_* fields of gsr_get_stats */
int gsr_render_timed(gsr_t* h, float* out_rgba, void* stream);
/* draw + read-back into HOST memory (pinned or pageable); synchronises; re-renders with a larger
 * capacity if the instance buffer overflowed.
 * gsr_render_host: H*W*4 floats.  gsr_render_rgb8_host: H*W*3 bytes, round(clamp(c,0,1)*255) — what
 * glReadPixels(GL_RGB, GL_UNSIGNED_BYTE) + FLIP_TOP_BOTTOM returns (main.py:852,912).
 * gsr_render_r32f_host: H*W floats, channel 0 — glReadPixels(GL_RED, GL_FLOAT) (main.py:870). */
int gsr_render_host(gsr_t* h, float* out_rgba_host, void* stream);
int gsr_render_rgb8_host(gsr_t* h, uint8_t* out_rgb_host, void* stream);
int gsr_render_r32f_host(gsr_t* h, float* out_r_host, void* stream);

/* Pipelined read-back for capture loops (main.py:793-917 renders and reads back hundreds of poses): enqueues
 * draw + RGB8 conversion + copy into out_rgb_host (pinned host memory, H*W*3 bytes) and returns.  The copy of
 * frame k runs on a second stream out of a second device buffer, so it overlaps the rendering of frame k+1.
 * Two frames may be in flight: the call first waits for the frame submitted two calls earlier.  A host
 * buffer is complete after the next-but-one call or after gsr_host_sync.  If a frame overflowed the
 * instance buffers the sync point grows them and returns GSR_EOVERFLOW: submit that frame again. */
int gsr_render_rgb8_host_async(gsr_t* h, uint8_t* out_rgb_host, void* stream);
int gsr_host_sync(gsr_t* h);

/* One pose of the stereo capture loop (main.py:844-917; one of the ~100 poses per scene of render_all.py:14-35),
 * three renders and their packing on the device:
 *   left      : scale/render_mod of gsr_set_params at (view_left, campos_left) on an RGB8 target ->
 *               H*W*3 bytes, glReadPixels(GL_RGB, GL_UNSIGNED_BYTE) + FLIP_TOP_BOTTOM (:846-852, :912)
 *   disparity : render_mod -1 (1.2x scale, gau_vert.glsl:151-156,182-211) at the LEFT pose on the R32F target
 *               (:858-870) -> H*W uint16 = (d * 65535).astype(uint16) (:878-879); the rotate-180 + mirror of
 *               :880-881 is the top-down row order every image of this ABI already has
 *   right     : like left at (view_right, campos_right) (:888-897); the reference builds view_right as the left
 *               view pre-multiplied by an x translation of -0.5 in view space (:376-380)
 * Any of the three host pointers may be NULL to skip that render.  share_order != 0: the depth order is
 * computed once per call and reused by the other renders when their view-space depths are identical (same
 * third row of the view matrix) — the result is bit-identical to three independent frames.
 * Pipelined exactly like gsr_render_rgb8_host_async (same two slots, same gsr_host_sync / GSR_EOVERFLOW
 * rules); host buffers should be pinned. */
int gsr_capture_stereo_async(gsr_t* h, const float view_left[16], const float campos_left[3],
                             const float view_right[16], const float campos_right[3], uint8_t* left_rgb_host,
                             uint16_t* disparity_host, uint8_t* right_rgb_host, int share_order, void* stream);

/* The render_all.py workload (render_all.py:14-35, main.py:793-840): n_views poses over one scene.
 * views: n_views x 16 floats, campos: n_views x 3 floats (HOST).  out_rgba: DEVICE buffer of
 * n_views*H*W*4 floats, or NULL to render every view into an internal scratch frame (throughput runs).
 * Asynchronous. */
int gsr_render_batch(gsr_t* h, const float* views, const float* campos, int n_views, float* out_rgba, void* stream);

/* ---- sort -------------------------------------------------------------------------------------- */
/* replaces _sort_gaussian(gaus, view_mat) (renderer_ogl.py:10-53) as called by sort_and_update
 * (renderer_ogl.py:139-146): out_index[k] = k-th Gaussian in ASCENDING view-space z (far -> near),
 * stable, NaN last; N int32, host (dst_is_device = 0, synchronises) or device memory. */
int gsr_sort_depth(gsr_t* h, const float view[16], int32_t* out_index, int dst_is_device, void* stream);

/* sort_and_update WITHOUT auto-sort (renderer_ogl.py:139-146 as driven by main.py:989-997: the order buffer `gi`
 * is rebuilt only when "sort Gaussians" is pressed or "auto sort" is ticked, and draw() keeps using it while the
 * camera moves).  gsr_freeze_order sorts the N Gaussians by their depth in `view` once (stable, front to back,
 * Gaussians behind that camera first — the reverse of the far->near argsort the reference uploads); every later
 * frame skips its own depth sort and builds its tile lists in that order (culling, footprints and colours still
 * follow the frame's own camera, exactly like the vertex shader reading a stale `gi`).  gsr_thaw_order, a new
 * scene or gsr_sort_depth return to one depth sort per frame.  Not available with GSR_FLAG_KEYS64. */
int gsr_freeze_order(gsr_t* h, const float view[16], void* stream);
int gsr_thaw_order(gsr_t* h);

/* The onesweep LSB radix sort primitive on its own (device pointers, n pairs, stable, bits
 * [begin_bit, end_bit) of the key).  key_bits is 32 or 64; values are uint32.  Sorted data ends up in
 * keys/vals; keys_tmp/vals_tmp are scratch of the same size.  vals may be NULL on input to mean
 * 0,1,2,... (vals_tmp and an output are still required).  Asynchronous. */
int gsr_sort_pairs(gsr_t* h, int key_bits, void* keys, void* keys_tmp, uint32_t* vals, uint32_t* vals_tmp,
                   long long n, int begin_bit, int end_bit, void* stream);

/* gsr_sort_pairs without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same arguments and result as gsr_sort_pairs minus the handle (it works on the CURRENT device; vals must not be
 * NULL); the scratch of the radix passes (digit histograms, tile tickets, look-back rows) is a static device array
 * inside libgsr.so, about 5.8 MB held by every GPU that loads the library until the process exits.
 * Limit: n <= GSR_SORT_STATIC_MAX_PAIRS, larger calls return GSR_EINVAL with a message.  The call makes no cudaMalloc /
 * cudaFree (nor any allocating library call).  Every call on a device shares the one array, which keeps the last call's
 * contents; the sort itself clears what must start at zero, but two calls that OVERLAP on the same GPU (different
 * streams, or threads) race on it - serialise them.  Asynchronous on `stream`. */
#define GSR_SORT_STATIC_MAX_PAIRS (1ll << 22)
int gsr_sort_pairs_static(int key_bits, void* keys, void* keys_tmp, uint32_t* vals, uint32_t* vals_tmp, long long n,
                          int begin_bit, int end_bit, void* stream);

/* ---- introspection ----------------------------------------------------------------------------- */
int gsr_get_stats(gsr_t* h, gsr_stats* out); /* synchronises */
int gsr_debug_fetch(gsr_t* h, int which, void* dst_host, long long capacity_bytes, long long* out_bytes);
/* device pointers of the resident SoA scene buffers, for torch.distributed.broadcast at scene load
 * (one NCCL broadcast per scene, none per frame): which = 0 xyz, 1 rot, 2 scale+opacity, 3 sh */
int gsr_scene_buffer(gsr_t* h, int which, void** dev_ptr, long long* bytes);
/* allocate resident scene buffers for n Gaussians without filling them (receiving side of a broadcast) */
int gsr_alloc_gaussians(gsr_t* h, long long n, int sh_dim);
/* number of kernels launched by this handle since creation (bench.py's gpu_launches) */
long long gsr_launch_count(gsr_t* h);

/* Per-handle variant knobs (measurement and cross-checks; images are identical for every setting).  Keys of round 1
 * that selected kernel variants which no longer exist (0-5, 8) are rejected with GSR_EINVAL. */
#define GSR_TUNE_FRAME_FORK 6 /* 1 (default) = scratch clears and the tile scan run on an auxiliary stream beside preprocess / depth sort */
#define GSR_TUNE_TILE 7 /* tile edge of the shader path in pixels, 32 (default) or 16; the instance list, ranges and tiles_x/tiles_y of gsr_stats follow it (GSR_FLAG_OFFICIAL frames always use 16) */
#define GSR_TUNE_FUSED_DEPTH_HIST 9 /* 1 (default) = fast depth sort: the preprocess kernel counts the digits of the rebased depth keys, three compacting 9-bit radix passes over the visible Gaussians; 0 = histogram kernel + four 8-bit passes over all N */
#define GSR_TUNE_PDL 10 /* bit mask of the kernels launched with programmatic dependent launch (resident behind their predecessor, griddepcontrol.wait before the first memory access): 1 radix passes, 2 duplicate, 4 blend, 8 preprocess; default 11, 0 = plain launches */
#define GSR_TUNE_FUSED_RGB8 11 /* 1 (default) = gsr_render_rgb8_host_async lets the blend kernel write the RGB8 frame (no k_pack_rgb8 launch) */
#define GSR_TUNE_GRAPH 12 /* 1 (default) = the stream operations of a frame configuration are captured once as a CUDA graph and replayed with new uniforms; 0 = every frame is enqueued kernel by kernel */
int gsr_tune(gsr_t* h, int key, int value);

/* The measurement entry points bench.py uses for its roofline denominators (gsr_measure_peaks, gsr_count_pairs) are
 * declared in include/gsr_measure.h: they are not part of what a maintainer binds. */

#ifdef __cplusplus
}
#endif
#endif /* GSR_H */
