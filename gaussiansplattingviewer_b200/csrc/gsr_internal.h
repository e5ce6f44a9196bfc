// Internal declarations shared by the CUDA translation units of libgsr.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define GSR_TILE 16  // smallest tile edge (buffers are sized for it); a frame uses 16 or 32, see FrameParams.tile_shift
#define GSR_GRID_REPLICAS 16  // copies of the corner-impulse grid (CTA b adds into copy b % 16)  // pixels per tile edge (16x16 tiles, one CTA per tile in the blend kernel)

// Per-frame uniforms.  Lives in device memory (written by k_set_params) so that a captured CUDA graph
// can be replayed with a new camera by updating one kernel node.
struct FrameParams {
    float view[16];  // row-major maths layout
    float proj[16];
    float htanx, htany, focal;
    float campos[3];
    float scale_modifier;
    int render_mod;
    int sh_dim;
    int W, H;
    int tiles_x, tiles_y;
    unsigned flags;
    int tile_shift;  // log2 of the tile edge in pixels: 4 (16x16) or 5 (32x32, shader path only)
    // digit plan of the frame's depth sort (the preprocess kernel counts these digits, the compacting first radix
    // pass applies the same key' = min(max(key, k0), khi) - k0): depth_npass passes of depth_bits[] bits, LSB first
    unsigned depth_k0, depth_khi;
    int depth_npass;
    int depth_bits[4];
};

// Device-side counters of one frame (zeroed at frame start).
struct FrameCounters {
    unsigned n_visible;    // Gaussians with ok = 1
    unsigned n_instances;  // D (sum of tiles touched), unclamped
    unsigned overflow;     // D > capacity
    unsigned n_sort;       // instances the tile sort handles: D, or 0 when the buffers overflowed
    unsigned n_keys;       // entries of the depth order the frame duplicates from (N, or the Gaussians with a depth key)
    unsigned depth_clamped;  // depth keys outside [depth_k0, depth_khi] (their mutual order is lost; none in a sane frustum)
    // not cleared per frame (pipelined paths read them at a later sync point and clear them there)
    unsigned sticky_overflow;  // frames that overflowed since the last clear
    unsigned sticky_max_d;     // largest D among them
};

// internal FrameParams.flags bit: the depth key is written for every Gaussian in front of the camera, visible or
// not, so that one depth order serves several renders that share the view-space z (stereo capture)
#define GSR_FLAG_INTERNAL_DEPTH_ALL 0x100u
#define GSR_FLAG_OFFICIAL_BIT 8u  // == GSR_FLAG_OFFICIAL of include/gsr.h

// Splat record, three float4 planes indexed by Gaussian:
//   rec0 = (cx, cy, hx, hy)            splat centre (window coordinates) and 3-sigma half extents
//   rec1 = (A', B', C', opacity)       conic pre-scaled: log2(e) * power = A' dx^2 + B' dx dy + C' dy^2
//   rec2 = (r, g, b, flag)             flag = 1: conic not safely positive definite (keep the power > 0 test)
// plus the tile rectangle trect = (tx0 | tiles_wide << 16, ty0 | tiles_high << 16), 0 x 0 when culled.
struct SceneBuffers {
    const float* xyz;      // N x 3
    const float4* rot;     // N (w, x, y, z)
    const float4* scale_op;  // N (sx, sy, sz, opacity)
    const float* sh;       // N x sh_dim
};

struct SplatBuffers {
    float4* rec0;
    float4* rec1;
    float4* rec2;
    uint32_t* depth_key;  // bits(depth) or 0xFFFFFFFF when not ok
    uint2* trect;         // tile rectangle (0 x 0 when not ok)
    int* tile_grid;       // (tiles_y + 1) x grid_w corner impulses of the tile rectangles (see k_tile_scan)
    int grid_w;           // row pitch of tile_grid: tiles_x + 1 rounded up to an odd number
    int grid_stride;      // ints per replica; GSR_GRID_REPLICAS replicas spread the atomics of hot tiles
    float* dbg_vsout;     // N x 16 or nullptr
    int32_t* dbg_setup;   // N x 8 or nullptr
    uint32_t* depth_hist; // [4][512] digit counts of the rebased depth keys (FrameParams.depth_*), accumulated by k_preprocess (nullptr: not wanted)
};

// ---- programmatic dependent launch (sm_90+) ---------------------------------------------------------
// A kernel launched with launch_pdl may become resident while the kernel in front of it on the stream is still
// running; it must call pdl_wait() before it touches ANY global memory (the call returns when the kernel in
// front has completed and its writes are visible; it is a no-op for a plain launch).  What this hides is the
// launch latency and block scheduling of the next kernel behind the tail of the current one.
// gsr_tune state of a handle; render_frame makes the handle's PDL mask current for the launches it issues
struct GsrTune {
    int pdl = 11;              // GSR_TUNE_PDL: bit 0 radix passes, 1 duplicate, 2 blend (measured slower: off), 3 preprocess
    int fast_depth_sort = 1;   // GSR_TUNE_FUSED_DEPTH_HIST: digit counts from the preprocess kernel + compacting 9-bit passes
    int fused_rgb8 = 1;        // GSR_TUNE_FUSED_RGB8
    int frame_fork = 1;        // GSR_TUNE_FRAME_FORK
    int graph = 1;             // GSR_TUNE_GRAPH: frames of a configuration are replayed from a captured CUDA graph
};
extern thread_local int g_pdl;  // PDL mask of the handle whose frame is being enqueued on this thread
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
template <typename... KArgs, typename... Args>
inline void launch_pdl_m(int which, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = (g_pdl & which) ? 1 : 0;
    cfg.attrs = attr; cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif

// ---- radix sort (radix_sort.cu) ----------------------------------------------------------------
// Scratch for one sort: histograms for up to 8 passes, ticket counters, two look-back status buffers.
struct SortScratch {
    uint32_t* hist;       // [8][256] raw digit counts per pass
    uint32_t* tickets;    // [8] tile tickets per pass (+ 1 for the histogram kernel)
    uint32_t* status[2];  // [max_tiles][256] decoupled look-back words, alternating per pass
    long long max_tiles;
    size_t zero_bytes;    // bytes to clear at sort start (hist + tickets are contiguous)
};
size_t sort_scratch_bytes(long long max_items, int key_bytes);
void sort_scratch_carve(SortScratch* s, void* base, long long max_items, int key_bytes);

// Sorts n pairs (n read from *n_dev when n_dev != nullptr, else n_host) on bits [begin_bit, end_bit).
// keys/vals are ping-ponged with keys_tmp/vals_tmp: returns 0 if the result is in keys/vals, 1 if in
// keys_tmp/vals_tmp.  vals_identity: the values of the first pass are 0, 1, 2, ... (vals is only written).
// keys_first: input of the first pass (may equal keys; it is only read), so a key array can be sorted without
// being overwritten.  opt (may be nullptr, all fields zero = defaults):
struct SortOptions {
    int n_pass;                  // > 0: explicit digit plan, pass p sorts pass_bits[p] bits (<= 9) upward from begin_bit
    int pass_bits[8];
    bool hist_ready;             // scratch->hist already holds the per-pass digit counts (k_preprocess / k_tile_scan):
                                 // no histogram kernel
    int hist_stride;             // words between the passes' histograms in scratch->hist (0 = 256)
    bool vals_only_last;         // the last pass does not write keys (the caller only needs the permuted values)
    unsigned hist_direct_mask;   // passes whose digits are spread over their bins: k_rs_hist counts them without the warp match
    bool scratch_cleared;        // the caller already ran sort_scratch_clear (e.g. on another stream)
    bool drop_rebase;            // 32-bit keys, identity values: the first pass drops keys == 0xFFFFFFFF and rewrites the
    uint32_t k0, khi;            // others as min(max(key, k0), khi) - k0; the later passes sort n_dev_rest keys
    const uint32_t* n_dev_rest;  // device count of the keys that survive the first pass
};
// the clears radix_sort_pairs starts with (histograms + tickets, or tickets + first look-back rows when hist_ready)
template <typename KeyT>
void sort_scratch_clear(SortScratch* scratch, long long upper, bool hist_ready, cudaStream_t stream, const SortOptions* opt = nullptr);
template <typename KeyT>
int radix_sort_pairs(const KeyT* keys_first, KeyT* keys, KeyT* keys_tmp, uint32_t* vals, uint32_t* vals_tmp, bool vals_identity,
                     long long n_host, const uint32_t* n_dev, long long capacity, int begin_bit, int end_bit,
                     SortScratch* scratch, cudaStream_t stream, long long* launch_count, const SortOptions* opt = nullptr);
int radix_sort_prepare_device();  // per-device kernel opt-ins (gsr_create)
// the same sort with its scratch in a static device array (no allocation); see radix_sort.cu
long long radix_sort_static_max_items();
int radix_sort_pairs_static(int key_bits, void* keys, void* keys_tmp, uint32_t* vals, uint32_t* vals_tmp, long long n,
                            int begin_bit, int end_bit, cudaStream_t stream, const char** err);
void launch_tile_scan(int* grid, int grid_stride, int gw, int tiles_x, int tiles_y, int bits0, long long capacity, int2* ranges,
                      uint32_t* hist, FrameCounters* counters, cudaStream_t s);

// ---- kernels' host launchers ---------------------------------------------------------------------
// n_keys_init: value of counters->n_keys at frame start (-1 = keep: the frame reuses a stored depth order)
void launch_set_params(FrameParams* dst, const FrameParams& fp, FrameCounters* counters, cudaStream_t s,
                       int* zero_a = nullptr, long long na = 0, uint32_t* zero_b = nullptr, int nb = 0,
                       long long n_keys_init = 0);
void launch_pack_scene(const float* rot, const float* scale, const float* opacity, float4* rot4, float4* scale_op,
                       long long n, cudaStream_t s);
void launch_preprocess(const FrameParams* fp, SceneBuffers scene, SplatBuffers out, FrameCounters* counters,
                       long long n, int sh_dim, cudaStream_t s);
// official-rasterizer semantics: fp->view / fp->proj hold the z-forward view and the FULL projection (P * view)
void launch_preprocess_official(const FrameParams* fp, SceneBuffers scene, SplatBuffers out, FrameCounters* counters,
                                long long n, int sh_dim, cudaStream_t s);
void launch_depth_keys(const float* xyz, const float* view16_host, uint32_t* keys, long long n, cudaStream_t s, int negate = 0);
// big_list: N uint4 entries, big_count: one zeroed counter (large rectangles, see binning.cu)
// n: upper bound of the entries of sorted_gid; the kernel reads counters->n_keys (<= n) entries
void launch_duplicate_sorted(const FrameParams* fp, const uint32_t* sorted_gid, SplatBuffers sp,
                             FrameCounters* counters, uint32_t* chain_status, uint32_t* chain_ticket,
                             uint32_t* out_tile, uint32_t* out_gid, uint4* big_list, uint32_t* big_count, long long n,
                             long long capacity, cudaStream_t s);
void launch_duplicate_keys64(const FrameParams* fp, SplatBuffers sp, FrameCounters* counters, uint32_t* chain_status,
                             uint32_t* chain_ticket, uint64_t* out_keys, uint32_t* out_gid, uint4* big_list,
                             uint32_t* big_count, long long n, long long capacity, cudaStream_t s);
template <typename KeyT>
void launch_tile_ranges(const KeyT* sorted_keys, int key_shift, const FrameCounters* counters, long long capacity,
                        int2* ranges, int n_tiles, cudaStream_t s);
void launch_blend(const FrameParams* fp, const int2* ranges, const uint32_t* inst_gid, SplatBuffers sp,
                  float* out_rgba, int render_mod, unsigned flags, int tiles_x, int tiles_y, cudaStream_t s, int tile = 16,
                  uint8_t* out_rgb8 = nullptr);  // out_rgb8: also write the frame as RGB8 (k_pack_rgb8's rule)
void launch_pack_rgb8(const float* rgba, uint8_t* rgb, long long n_px, cudaStream_t s);
void launch_extract_r(const float* rgba, float* r, long long n_px, cudaStream_t s);
void launch_pack_disp_u16(const float* rgba, uint16_t* d, long long n_px, cudaStream_t s);
void launch_keys64_from_ranges(const int2* ranges, int n_tiles, const uint32_t* gid, const uint32_t* depth_key,
                               uint64_t* out, cudaStream_t s);
void launch_blend_count(const FrameParams* fp, const int2* ranges, const uint32_t* inst_gid, SplatBuffers sp,
                        float* out_rgba, int render_mod, unsigned flags, int tiles_x, int tiles_y,
                        unsigned long long* pair_counts, cudaStream_t s, int tile = 16);
// the k_set_params kernel (to find its node in a captured frame) and the update of that node's arguments
const void* set_params_kernel();
int set_params_node_update(cudaGraphExec_t exec, cudaGraphNode_t node, const cudaKernelNodeParams& base, FrameParams* dst,
                           const FrameParams& fp, FrameCounters* counters, int* zero_a, long long na, uint32_t* zero_b, int nb,
                           long long n_keys_init);
int preprocess_prepare_device();  // per-device kernel opt-ins (gsr_create)
int sm_count_current();           // SM count of the current device (cached per device)
int measure_peaks_impl(float* fp32_tflops, float* ex2_tops, float* copy_gbs);

// ---- PLY decode (ply_decode.cu) ----------------------------------------------------------------
// Byte offsets of the 59 float32 columns load_ply reads, in canonical order:
// 0-2 x y z | 3-5 f_dc_0..2 | 6-50 f_rest_0..44 | 51 opacity | 52-54 scale_0..2 | 55-58 rot_0..3
struct PlyLayout {
    int row_bytes;
    int big_endian;
    int off[59];
};
int ply_decode_configure(int row_bytes);  // 0 ok, -1 row too wide for the staging window
void launch_ply_bbox_init(int* bbox_ord, cudaStream_t s);
// rows: n vertex rows (device, 16-byte aligned, readable up to the next multiple of 16 bytes); outputs are
// written at Gaussian index first + i
void launch_ply_decode(const unsigned char* rows, const PlyLayout& lay, long long n, long long first, float* xyz,
                       float4* rot4, float4* scale_op, float* sh, int* bbox_ord, int sm_count, cudaStream_t s);
// bbox_ord: 10 words (6 ordered min/max, NaN mask, 3 running float32 sums); center chunks must be launched in row
// order on one stream, each starting at a multiple of 4 rows
void launch_ply_center(const float* xyz_rows, long long n_rows, int* bbox_ord, cudaStream_t s);
void launch_ply_final(const int* bbox_ord, long long n, float* bbox6, float* center3, cudaStream_t s);
