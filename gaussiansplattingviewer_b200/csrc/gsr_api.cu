// C ABI of libgsr.so (include/gsr.h): resource management and the per-frame launch sequence
//   K0 set params -> K1 preprocess -> K3 depth sort (N pairs) -> K2 duplicate (depth order) ->
//   K3 tile sort (D pairs, tile bits only) -> K4 ranges -> K5 blend
// The two-level sort is an LSB radix sort of the 64-bit key (tile << 32 | depth bits) whose four
// depth-digit passes run on the N Gaussians BEFORE duplication instead of on the D instances after it:
// duplication preserves order and each Gaussian appears at most once per tile, so the final order is
// exactly the stable order by (tile, depth, index).  GSR_FLAG_KEYS64 selects the one-sort variant.
// No host synchronisation inside a frame: D lives on the device, buffers are capacity-bounded.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/gsr.h"
#include "../../include/gsr_measure.h"
#include "gsr_internal.h"

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}

#define CK(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess)                                                                       \
            return fail(e_ == cudaErrorMemoryAllocation ? GSR_ENOMEM : GSR_ECUDA,                   \
                        std::string(#call) + ": " + cudaGetErrorString(e_));                        \
    } while (0)

constexpr long long MAX_CAPACITY = (1ll << 30) - 1;

void drop_graphs(gsr_t* h);  // captured frames refer to buffers: every (re)allocation drops them

template <typename T>
void dfree(T*& p) {
    if (p) cudaFree(p);
    p = nullptr;
}

}  // namespace

// a captured frame (see render_frame)
struct GraphKey {
    long long n, capacity; const void *out, *out_rgb8, *dbg;
    int sh_dim, W, H, tile_req, order_mode, order_buf_in, blend_class, pdl, fused_hist, fork;
    unsigned flags, epoch;
};
struct FrameGraph {
    GraphKey key;
    cudaGraphExec_t exec = nullptr;     // nullptr: the key was seen once, not captured yet
    cudaGraph_t graph = nullptr;        // the captured graph: owns params_node, lives as long as exec
    cudaGraphNode_t params_node = nullptr;
    cudaKernelNodeParams params_base = {};
    long long launches = 0;             // kernels of one frame (gsr_launch_count)
    int order_buf_after = -1, last_result_buf = 0;
    bool order_count_dev_after = false, last_keys64 = false;
    unsigned long long last_use = 0;
};
constexpr int MAX_FRAME_GRAPHS = 12;

struct gsr_renderer {
    int device = 0;
    long long N = 0;
    int sh_dim = 0;
    // resident scene (SoA)
    float* d_xyz = nullptr;
    float4* d_rot = nullptr;
    float4* d_scale_op = nullptr;
    float* d_sh = nullptr;
    // per-Gaussian frame buffers
    SplatBuffers sp = {};
    uint32_t* d_dkey[2] = {nullptr, nullptr};  // depth-sort key ping-pong
    uint32_t* d_gid[2] = {nullptr, nullptr};   // depth-sort value ping-pong
    void* d_sortn = nullptr;                   // sort scratch sized for N
    SortScratch ss_n = {};
    unsigned long long* d_chain = nullptr;     // duplicate: look-back words per chunk + ticket + big-list counter
    uint4* d_big = nullptr;                    // duplicate: large rectangles (N entries)
    size_t chain_bytes = 0;
    // per-instance buffers
    long long capacity = 0, capacity_req = 0;
    bool cap_keys64 = false;
    void* d_ikeys[2] = {nullptr, nullptr};
    uint32_t* d_ivals[2] = {nullptr, nullptr};
    void* d_sorti = nullptr;
    SortScratch ss_i = {};
    uint64_t* d_dbgkeys = nullptr;
    // image-side
    int W = 0, H = 0, tiles_x = 0, tiles_y = 0, tile = 16;  // tiling of the last frame
    int tile_req = 32;  // gsr_tune(GSR_TUNE_TILE): tile edge of the shader path, 16 or 32 (official-mode frames always use 16)
    int2* d_ranges = nullptr;
    int* d_tile_grid = nullptr;  // (tiles_y + 1) x grid_w corner impulses
    int grid_w = 0;
    float* d_frame = nullptr;
    uint8_t* d_rgb8 = nullptr;
    // pipelined read-back: second RGB8 buffer, copy stream, per-slot events
    uint8_t* d_rgb8b = nullptr;
    uint8_t* d_cap[2] = {nullptr, nullptr};  // stereo capture slots: left RGB8 | right RGB8 | disparity u16
    int order_buf = -1;                      // d_gid[] buffer holding a shared depth order (stereo capture, frozen order)
    bool order_frozen = false;               // gsr_freeze_order: frames draw in the stored order
    bool order_count_dev = false;            // the stored order is compacted: its length lives in d_counters->n_keys
    float frozen_view[16] = {};              // the view of that order (gsr_sort_depth borrows its buffers and re-freezes)
    cudaStream_t copy_stream = nullptr;
    // frame-internal fork: scratch clears and the single-CTA k_tile_scan run beside preprocess / the depth sort
    cudaStream_t aux_stream = nullptr;
    cudaEvent_t ev_start = nullptr, ev_pre = nullptr, ev_cleared = nullptr, ev_scan = nullptr;
    cudaEvent_t ev_packed[2] = {}, ev_copied[2] = {};
    bool slot_busy[2] = {false, false};
    unsigned async_idx = 0;
    float* d_r32 = nullptr;
    // uniforms
    FrameParams fp = {};
    FrameParams* d_fp = nullptr;
    FrameCounters* d_counters = nullptr;
    FrameCounters* h_counters = nullptr;  // pinned
    bool have_camera = false;
    // last frame
    int last_result_buf = 0;  // which instance ping-pong buffer holds the sorted list
    bool last_keys64 = false;
    bool frame_valid = false;
    gsr_stats stats = {};
    cudaEvent_t ev[8] = {};
    long long launches = 0;
    GsrTune tune;  // gsr_tune state (per handle)
    // captured frames (render_frame)
    std::vector<FrameGraph> graphs;
    unsigned graph_epoch = 0;
    unsigned long long graph_clock = 0;
    cudaStream_t cap_stream = nullptr;
    // scratch for gsr_sort_pairs
    void* d_sortp = nullptr;
    long long sortp_items = 0;
    int sortp_keybytes = 0;
};

thread_local int g_pdl = 11;  // PDL mask of the handle whose frame this thread is enqueuing (GsrTune.pdl)

namespace {


int free_gaussian_buffers(gsr_t* h) {
    drop_graphs(h);
    dfree(h->d_xyz); dfree(h->d_rot); dfree(h->d_scale_op); dfree(h->d_sh);
    dfree(h->sp.rec0); dfree(h->sp.rec1); dfree(h->sp.rec2); dfree(h->sp.depth_key); dfree(h->sp.trect);
    dfree(h->sp.dbg_vsout); dfree(h->sp.dbg_setup);
    dfree(h->d_dkey[0]); dfree(h->d_dkey[1]); dfree(h->d_gid[0]); dfree(h->d_gid[1]);
    dfree(h->d_sortn); dfree(h->d_chain); dfree(h->d_big);
    h->N = 0;
    return 0;
}

int free_instance_buffers(gsr_t* h) {
    drop_graphs(h);
    dfree(h->d_ikeys[0]); dfree(h->d_ikeys[1]); dfree(h->d_ivals[0]); dfree(h->d_ivals[1]);
    dfree(h->d_sorti); dfree(h->d_dbgkeys);
    h->capacity = 0;
    return 0;
}

int alloc_gaussian_buffers(gsr_t* h, long long n, int sh_dim) {
    free_gaussian_buffers(h);
    const size_t n1 = (size_t)std::max<long long>(n, 1);
    CK(cudaMalloc(&h->d_xyz, n1 * 3 * sizeof(float)));
    CK(cudaMalloc(&h->d_rot, n1 * sizeof(float4)));
    CK(cudaMalloc(&h->d_scale_op, n1 * sizeof(float4)));
    CK(cudaMalloc(&h->d_sh, n1 * (size_t)sh_dim * sizeof(float)));
    CK(cudaMalloc(&h->sp.rec0, n1 * sizeof(float4)));
    CK(cudaMalloc(&h->sp.rec1, n1 * sizeof(float4)));
    CK(cudaMalloc(&h->sp.rec2, n1 * sizeof(float4)));
    CK(cudaMalloc(&h->sp.depth_key, n1 * 4));
    CK(cudaMalloc(&h->sp.trect, n1 * sizeof(uint2)));
    for (int i = 0; i < 2; ++i) {
        CK(cudaMalloc(&h->d_dkey[i], n1 * 4));
        CK(cudaMalloc(&h->d_gid[i], n1 * 4));
    }
    CK(cudaMalloc(&h->d_sortn, sort_scratch_bytes((long long)n1, 4)));
    sort_scratch_carve(&h->ss_n, h->d_sortn, (long long)n1, 4);
    h->chain_bytes = ((n1 + 1023) / 1024 + 3) * 8;
    CK(cudaMalloc(&h->d_chain, h->chain_bytes));
    CK(cudaMalloc(&h->d_big, n1 * sizeof(uint4)));
    h->N = n;
    h->sh_dim = sh_dim;
    h->frame_valid = false;
    h->order_buf = -1;
    h->order_frozen = false;
    return 0;
}

int ensure_debug_buffers(gsr_t* h) {
    const size_t n1 = (size_t)std::max<long long>(h->N, 1);
    if (h->fp.flags & GSR_FLAG_DEBUG) {
        if (!h->sp.dbg_vsout) CK(cudaMalloc(&h->sp.dbg_vsout, n1 * 16 * sizeof(float)));
        if (!h->sp.dbg_setup) CK(cudaMalloc(&h->sp.dbg_setup, n1 * 8 * sizeof(int)));
    } else if (h->sp.dbg_vsout || h->sp.dbg_setup) {
        CK(cudaDeviceSynchronize());  // frames in flight may still write them
        drop_graphs(h);
        dfree(h->sp.dbg_vsout);
        dfree(h->sp.dbg_setup);
    }
    return 0;
}

int ensure_capacity(gsr_t* h, long long want, bool keys64) {
    want = std::min(std::max<long long>(want, 1024), MAX_CAPACITY);
    if (h->capacity >= want && (h->cap_keys64 || !keys64)) return 0;
    CK(cudaDeviceSynchronize());
    keys64 = keys64 || h->cap_keys64;
    free_instance_buffers(h);
    const size_t kb = keys64 ? 8 : 4;
    for (int i = 0; i < 2; ++i) {
        CK(cudaMalloc(&h->d_ikeys[i], (size_t)want * kb));
        CK(cudaMalloc(&h->d_ivals[i], (size_t)want * 4));
    }
    CK(cudaMalloc(&h->d_sorti, sort_scratch_bytes(want, (int)kb)));
    sort_scratch_carve(&h->ss_i, h->d_sorti, want, (int)kb);
    h->capacity = want;
    h->cap_keys64 = keys64;
    h->frame_valid = false;
    return 0;
}

long long auto_capacity(const gsr_t* h) {
    if (h->capacity_req > 0) return h->capacity_req;
    // instances per Gaussian grow with the pixel count: 16 per Gaussian up to 1080p, 4x that at 4K and above
    const double px = (double)h->W * (double)h->H / (1920.0 * 1080.0);
    const long long per = (long long)(16.0 * std::min(4.0, std::max(1.0, px)));
    return std::min<long long>(std::max<long long>(4ll << 20, per * h->N), MAX_CAPACITY);
}

int tile_bits(int n_tiles) {
    int b = 1;
    while ((1ll << b) < n_tiles) ++b;
    return b;
}

// Enqueues one frame.  timed: record events between stages (h->ev[0..6]).
// order_mode: 0 = sort this frame's depth keys; 1 = same, with keys for every Gaussian in front of the camera, and
// remember the order; 2 = reuse the remembered order (same view-space depths: same view up to an in-plane shift)
// What a frame's kernels are launched with that changes from frame to frame (the uniforms) plus what decides WHICH
// kernels and buffers a frame uses (the key of a captured graph).
struct FramePlan {
    FrameParams fp;         // final uniforms (official matrices, tiling, depth plan)
    long long n_keys_init;  // k_set_params: counters->n_keys at frame start
    int* zero_a; long long na; uint32_t* zero_b; int nb;  // k_set_params: the two clears
    int order_mode;         // after the adjustments of frame_body
    int blend_class;        // which k_blend instance: 0 gaussian, 2 billboard, 3 flat ball, 4 gaussian ball, 5 official
    bool fused_hist, fork;
};

int frame_body(gsr_t* h, const FrameParams& fp_in, float* out, cudaStream_t s, bool timed, int order_mode,
               uint8_t* out_rgb8, FramePlan* plan_only) {
    FrameParams fp = fp_in;
    g_pdl = h->tune.pdl;
    if (fp.flags & GSR_FLAG_KEYS64) {
        if (h->order_frozen) return fail(GSR_EINVAL, "a frozen order needs the two-level sort (clear GSR_FLAG_KEYS64)");
        order_mode = 0;
    }
    if (h->order_frozen) order_mode = 2;
    if (order_mode == 2 && h->order_buf < 0) order_mode = 1;
    fp.flags &= ~GSR_FLAG_INTERNAL_DEPTH_ALL;
    const bool official = (fp.flags & GSR_FLAG_OFFICIAL) != 0;
    if (official) {
        // what CUDARenderer.update_camera_pose builds (renderer_cuda.py:181-194): view rows 0 and 2 negated
        // (GL -> z-forward), projmatrix = P_gl * view; the kernels read them from fp.view / fp.proj
        if (h->order_frozen) return fail(GSR_EINVAL, "the official-rasterizer mode sorts every frame (thaw the order first)");
        order_mode = 0;
        float vo[16], po[16];
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) vo[r * 4 + c] = (r == 0 || r == 2) ? -fp_in.view[r * 4 + c] : fp_in.view[r * 4 + c];
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) {
                float acc = fp_in.proj[r * 4 + 0] * vo[0 * 4 + c];
                acc = acc + fp_in.proj[r * 4 + 1] * vo[1 * 4 + c];
                acc = acc + fp_in.proj[r * 4 + 2] * vo[2 * 4 + c];
                acc = acc + fp_in.proj[r * 4 + 3] * vo[3 * 4 + c];
                po[r * 4 + c] = acc;
            }
        memcpy(fp.view, vo, 64);
        memcpy(fp.proj, po, 64);
    }
    if (order_mode == 1) fp.flags |= GSR_FLAG_INTERNAL_DEPTH_ALL;
    if (h->N <= 0 && !h->d_xyz) return fail(GSR_EINVAL, "no gaussians set");
    if (h->W <= 0) return fail(GSR_EINVAL, "no resolution set");
    if (!h->have_camera) return fail(GSR_EINVAL, "no camera set");
    if (!out) return fail(GSR_EINVAL, "null output");
    const bool keys64 = (fp.flags & GSR_FLAG_KEYS64) != 0;
    int rc = ensure_debug_buffers(h);
    if (rc) return rc;
    rc = ensure_capacity(h, std::max(h->capacity, auto_capacity(h)), keys64);
    if (rc) return rc;
    const long long N = h->N, cap = h->capacity;
    // tile edge of this frame: buffers are sized for 16; the official rasterizer's semantics depend on 16x16 tiles
    const int tile = (official || h->tile_req != 32) ? 16 : 32;
    fp.tile_shift = (tile == 32) ? 5 : 4;
    h->tiles_x = fp.tiles_x = (h->W + tile - 1) / tile;
    h->tiles_y = fp.tiles_y = (h->H + tile - 1) / tile;
    h->grid_w = (h->tiles_x + 1) | 1;  // odd pitch: conflict-free column walks in k_tile_scan
    h->tile = tile;
    const int n_tiles = h->tiles_x * h->tiles_y;
    const int tbits = tile_bits(n_tiles);
    SceneBuffers scene = {h->d_xyz, h->d_rot, h->d_scale_op, h->d_sh};

    // split of the tile-id bits between the two radix passes of the tile sort (k_tile_scan builds both
    // digit histograms from the per-tile counts); more than 16 tile bits fall back to the histogram kernel
    const int bits0 = (tbits > 8) ? (tbits + 1) / 2 : tbits;
    const bool scan_hist = (tbits <= 16) && !keys64;
    SplatBuffers sp = h->sp;
    sp.tile_grid = h->d_tile_grid;
    sp.grid_w = h->grid_w;
    sp.grid_stride = h->grid_w * (h->tiles_y + 1);

    // Untimed frames fork: the scratch clears run beside the preprocess kernel and the single-CTA k_tile_scan
    // beside the depth sort, on the handle's auxiliary stream (joined by events; no host synchronisation).
    // Timed frames stay on one stream so that the per-stage events bracket exactly one stage each.
    const bool fork = !timed && !keys64 && h->tune.frame_fork;
    cudaStream_t sa = fork ? h->aux_stream : s;
    const long long sort_upper = cap;
    // Fast depth sort of the shader path: a visible depth lies in [0.1, 100] (depth clip), so its bits minus those of
    // 0.05 are a 27-bit integer - three 9-bit radix passes.  The preprocess kernel counts those digits while it writes
    // the keys (no histogram kernel), the first pass drops the culled Gaussians (compaction, ties keep ascending
    // index), so passes two and three and the duplicate kernel only see the Gaussians that have a key.
    const bool fused_hist = h->tune.fast_depth_sort && !keys64 && !official && order_mode != 2;
    fp.depth_k0 = 0x3D4CCCCDu;                        // bits(0.05f)
    fp.depth_khi = fp.depth_k0 + ((1u << 27) - 1u);   // bits of ~3277: keys outside are clamped (and counted)
    fp.depth_npass = 3;
    fp.depth_bits[0] = fp.depth_bits[1] = fp.depth_bits[2] = 9; fp.depth_bits[3] = 0;
    sp.depth_hist = fused_hist ? h->ss_n.hist : nullptr;
    // entries of the depth order this frame duplicates from: counted by the preprocess kernel (fast path), all N
    // (generic path), or what the stored order has (reuse)
    long long n_keys_init = fused_hist ? 0 : N;
    if (order_mode == 2) n_keys_init = h->order_count_dev ? -1 : N;
    if (plan_only) {  // everything above is host work; nothing has been enqueued
        plan_only->fp = fp;
        plan_only->n_keys_init = n_keys_init;
        plan_only->zero_a = h->d_tile_grid; plan_only->na = (long long)sp.grid_stride * GSR_GRID_REPLICAS;
        plan_only->zero_b = sp.depth_hist; plan_only->nb = sp.depth_hist ? 3 * 512 : 0;
        plan_only->order_mode = order_mode;
        plan_only->blend_class = official ? 5 : (fp.render_mod == -2 ? 2 : (fp.render_mod == -3 ? 3 : (fp.render_mod == -4 ? 4 : 0)));
        plan_only->fused_hist = fused_hist;
        plan_only->fork = fork;
        return 0;
    }
    if (timed) cudaEventRecord(h->ev[0], s);
    if (fork) {
        CK(cudaEventRecord(h->ev_start, s));
        CK(cudaStreamWaitEvent(sa, h->ev_start, 0));
    }
    // uniforms, counters, and the clears of the corner-impulse grid and of those histograms in one launch
    launch_set_params(h->d_fp, fp, h->d_counters, s, h->d_tile_grid, (long long)sp.grid_stride * GSR_GRID_REPLICAS,
                      sp.depth_hist, 3 * 512, n_keys_init);
    if (official) launch_preprocess_official(h->d_fp, scene, sp, h->d_counters, N, h->sh_dim, s);
    else launch_preprocess(h->d_fp, scene, sp, h->d_counters, N, h->sh_dim, s);
    h->launches += 2;
    if (timed) cudaEventRecord(h->ev[1], s);
    CK(cudaMemsetAsync(h->d_chain, 0, h->chain_bytes, sa));
    uint32_t* chain_ticket = reinterpret_cast<uint32_t*>(h->d_chain + (h->chain_bytes / 8 - 1));
    uint32_t* big_count = reinterpret_cast<uint32_t*>(h->d_chain + (h->chain_bytes / 8 - 2));
    const uint32_t* n_sort = &h->d_counters->n_sort;
    int res = 0;
    if (!keys64) {
        // depth sort of the N Gaussians (culled ones carry key 0xFFFFFFFF and sort to the end)
        // (the two mantissa bytes of a depth key are spread over their bins, the exponent bytes are not)
        SortOptions so_n = {};
        so_n.vals_only_last = true;   // only the order (the Gaussian indices) is used downstream
        so_n.scratch_cleared = fork;
        if (fused_hist) {
            so_n.n_pass = 3;
            so_n.pass_bits[0] = so_n.pass_bits[1] = so_n.pass_bits[2] = 9;
            so_n.hist_ready = true;
            so_n.hist_stride = 512;
            so_n.drop_rebase = true;
            so_n.k0 = fp.depth_k0; so_n.khi = fp.depth_khi;
            so_n.n_dev_rest = &h->d_counters->n_keys;
        } else {
            so_n.hist_direct_mask = 0x3u;  // the two mantissa bytes of a depth key are spread over their bins
        }
        SortOptions so = {};
        so.vals_only_last = true;
        so.scratch_cleared = fork;
        if (scan_hist) {  // both digit histograms of the tile sort come from k_tile_scan
            so.hist_ready = true;
            so.n_pass = (tbits > bits0) ? 2 : 1;
            so.pass_bits[0] = bits0;
            so.pass_bits[1] = tbits - bits0;
        }
        if (fork) {
            if (order_mode != 2) sort_scratch_clear<uint32_t>(&h->ss_n, N, fused_hist, sa, &so_n);
            sort_scratch_clear<uint32_t>(&h->ss_i, sort_upper, scan_hist, sa, &so);
            CK(cudaEventRecord(h->ev_cleared, sa));
            CK(cudaEventRecord(h->ev_pre, s));
            CK(cudaStreamWaitEvent(sa, h->ev_pre, 0));
            launch_tile_scan(h->d_tile_grid, sp.grid_stride, h->grid_w, h->tiles_x, h->tiles_y, bits0, cap, h->d_ranges,
                             scan_hist ? h->ss_i.hist : nullptr, h->d_counters, sa);
            h->launches += 1;
            // the frame's counters are final after the tile scan: their read-back runs here, beside the depth sort,
            // not behind the blend kernel in front of the next frame
            CK(cudaMemcpyAsync(h->h_counters, h->d_counters, sizeof(FrameCounters), cudaMemcpyDeviceToHost, sa));
            CK(cudaEventRecord(h->ev_scan, sa));
            CK(cudaStreamWaitEvent(s, h->ev_cleared, 0));
        }
        int r0 = h->order_buf;
        if (order_mode != 2) {
            r0 = radix_sort_pairs<uint32_t>(h->sp.depth_key, h->d_dkey[0], h->d_dkey[1], h->d_gid[0], h->d_gid[1],
                                            true, N, nullptr, N, 0, 32, &h->ss_n, s, &h->launches, &so_n);
            h->order_buf = (order_mode == 1) ? r0 : -1;
            h->order_count_dev = fused_hist;
        }
        const uint32_t* sorted_gid = h->d_gid[r0];
        if (timed) cudaEventRecord(h->ev[2], s);
        // counts -> ranges, D, overflow flag, digit histograms of the tile sort
        if (fork) {
            CK(cudaStreamWaitEvent(s, h->ev_scan, 0));
        } else {
            launch_tile_scan(h->d_tile_grid, sp.grid_stride, h->grid_w, h->tiles_x, h->tiles_y, bits0, cap, h->d_ranges,
                             scan_hist ? h->ss_i.hist : nullptr, h->d_counters, s);
            h->launches += 1;
        }
        if (timed) cudaEventRecord(h->ev[3], s);
        launch_duplicate_sorted(h->d_fp, sorted_gid, sp, h->d_counters, reinterpret_cast<uint32_t*>(h->d_chain),
                                chain_ticket, (uint32_t*)h->d_ikeys[0], h->d_ivals[0], h->d_big, big_count, N, cap, s);
        h->launches += 2;
        if (timed) cudaEventRecord(h->ev[4], s);
        res = radix_sort_pairs<uint32_t>((uint32_t*)h->d_ikeys[0], (uint32_t*)h->d_ikeys[0], (uint32_t*)h->d_ikeys[1],
                                         h->d_ivals[0], h->d_ivals[1], false, 0, n_sort, cap, 0, tbits, &h->ss_i, s,
                                         &h->launches, &so);
    } else {
        if (timed) cudaEventRecord(h->ev[2], s);
        launch_tile_scan(h->d_tile_grid, sp.grid_stride, h->grid_w, h->tiles_x, h->tiles_y, bits0, cap, h->d_ranges, nullptr,
                         h->d_counters, s);
        h->launches += 1;
        if (timed) cudaEventRecord(h->ev[3], s);
        launch_duplicate_keys64(h->d_fp, sp, h->d_counters, reinterpret_cast<uint32_t*>(h->d_chain), chain_ticket,
                                (uint64_t*)h->d_ikeys[0], h->d_ivals[0], h->d_big, big_count, N, cap, s);
        h->launches += 2;
        if (timed) cudaEventRecord(h->ev[4], s);
        res = radix_sort_pairs<uint64_t>((uint64_t*)h->d_ikeys[0], (uint64_t*)h->d_ikeys[0], (uint64_t*)h->d_ikeys[1],
                                         h->d_ivals[0], h->d_ivals[1], false, 0, n_sort, cap, 0, 32 + tbits, &h->ss_i, s,
                                         &h->launches);
        // the one-sort variant re-derives the ranges from the sorted keys (cross-check of k_tile_scan)
        launch_tile_ranges<uint64_t>((const uint64_t*)h->d_ikeys[res], 32, h->d_counters, cap, h->d_ranges, n_tiles, s);
        h->launches += 1;
    }
    if (timed) cudaEventRecord(h->ev[5], s);
    launch_blend(h->d_fp, h->d_ranges, h->d_ivals[res], h->sp, out, fp.render_mod, fp.flags, h->tiles_x, h->tiles_y, s, tile, out_rgb8);
    h->launches += 1;
    if (timed) cudaEventRecord(h->ev[6], s);
    if (!fork) CK(cudaMemcpyAsync(h->h_counters, h->d_counters, sizeof(FrameCounters), cudaMemcpyDeviceToHost, s));
    h->last_result_buf = res;
    h->last_keys64 = keys64;
    h->frame_valid = true;
    CK(cudaGetLastError());
    return 0;
}

// ---- the frame as a CUDA graph ---------------------------------------------------------------------------------
// A frame is 12 kernels, a few memsets and a small copy, forked over two streams: enqueued one by one that is ~20
// driver calls per frame.  Frames of one configuration differ only in the uniforms, which live in device memory and
// are written by the frame's first kernel (k_set_params) - so the stream operations of a configuration are captured
// once and replayed with that kernel node's arguments updated (FramePlan).  What selects kernels or buffers is the key.
void drop_graphs(gsr_t* h) {
    for (FrameGraph& g : h->graphs) {
        if (g.exec) cudaGraphExecDestroy(g.exec);
        if (g.graph) cudaGraphDestroy(g.graph);
    }
    h->graphs.clear();
    ++h->graph_epoch;
}

int render_frame(gsr_t* h, const FrameParams& fp_in, float* out, cudaStream_t s, bool timed, int order_mode = 0,
                 uint8_t* out_rgb8 = nullptr) {
    if (timed || !h->tune.graph) return frame_body(h, fp_in, out, s, timed, order_mode, out_rgb8, nullptr);
    FramePlan plan;
    int rc = frame_body(h, fp_in, out, s, false, order_mode, out_rgb8, &plan);  // validates, (re)allocates, plans
    if (rc) return rc;
    GraphKey key;
    memset(&key, 0, sizeof(key));
    key.n = h->N; key.capacity = h->capacity; key.out = out; key.out_rgb8 = out_rgb8; key.dbg = h->sp.dbg_vsout;
    key.sh_dim = h->sh_dim; key.W = h->W; key.H = h->H; key.tile_req = h->tile_req; key.order_mode = plan.order_mode;
    key.order_buf_in = (plan.order_mode == 2) ? h->order_buf : -1;
    key.blend_class = plan.blend_class; key.pdl = h->tune.pdl; key.fused_hist = plan.fused_hist; key.fork = plan.fork;
    key.flags = plan.fp.flags; key.epoch = h->graph_epoch;
    FrameGraph* g = nullptr;
    for (FrameGraph& e : h->graphs) if (memcmp(&e.key, &key, sizeof(key)) == 0) { g = &e; break; }
    if (!g) {  // first frame of this configuration: enqueue directly (lazy one-time set-up happens here), remember the key
        if ((int)h->graphs.size() >= MAX_FRAME_GRAPHS) {
            size_t victim = 0;
            for (size_t i = 1; i < h->graphs.size(); ++i) if (h->graphs[i].last_use < h->graphs[victim].last_use) victim = i;
            if (h->graphs[victim].exec) cudaGraphExecDestroy(h->graphs[victim].exec);
            if (h->graphs[victim].graph) cudaGraphDestroy(h->graphs[victim].graph);
            h->graphs.erase(h->graphs.begin() + victim);
        }
        FrameGraph e;
        e.key = key;
        e.last_use = ++h->graph_clock;
        h->graphs.push_back(e);
        return frame_body(h, fp_in, out, s, false, order_mode, out_rgb8, nullptr);
    }
    g->last_use = ++h->graph_clock;
    bool fresh = false;
    if (!g->exec) {  // second frame: capture
        const long long l0 = h->launches;
        cudaGraph_t graph = nullptr;
        cudaError_t e = cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal);
        if (e == cudaSuccess) {
            rc = frame_body(h, fp_in, out, h->cap_stream, false, order_mode, out_rgb8, nullptr);
            e = cudaStreamEndCapture(h->cap_stream, &graph);
            if (rc != 0 && e == cudaSuccess) e = cudaErrorUnknown;
        }
        if (e == cudaSuccess) e = cudaGraphInstantiate(&g->exec, graph, 0);
        if (e == cudaSuccess) {  // find the k_set_params node
            size_t nn = 0;
            cudaGraphGetNodes(graph, nullptr, &nn);
            std::vector<cudaGraphNode_t> nodes(nn);
            cudaGraphGetNodes(graph, nodes.data(), &nn);
            for (cudaGraphNode_t nd : nodes) {
                cudaGraphNodeType ty;
                if (cudaGraphNodeGetType(nd, &ty) != cudaSuccess || ty != cudaGraphNodeTypeKernel) continue;
                cudaKernelNodeParams kp = {};
                if (cudaGraphKernelNodeGetParams(nd, &kp) == cudaSuccess && kp.func == set_params_kernel()) {
                    g->params_node = nd; g->params_base = kp; break;
                }
            }
            if (!g->params_node) e = cudaErrorUnknown;
        }
        if (e != cudaSuccess) {  // this driver / configuration cannot be captured: frames are enqueued directly from now on
            cudaGetLastError();
            if (g->exec) { cudaGraphExecDestroy(g->exec); g->exec = nullptr; }
            if (graph) cudaGraphDestroy(graph);
            h->tune.graph = 0;
            h->launches = l0;
            return frame_body(h, fp_in, out, s, false, order_mode, out_rgb8, nullptr);
        }
        g->graph = graph;
        g->launches = h->launches - l0;
        g->order_buf_after = h->order_buf; g->order_count_dev_after = h->order_count_dev;
        g->last_result_buf = h->last_result_buf; g->last_keys64 = h->last_keys64;
        fresh = true;
    }
    // replay with this frame's uniforms
    if (set_params_node_update(g->exec, g->params_node, g->params_base, h->d_fp, plan.fp, h->d_counters, plan.zero_a, plan.na,
                               plan.zero_b, plan.nb, plan.n_keys_init) != 0)
        return fail(GSR_ECUDA, std::string("graph parameter update failed: ") + cudaGetErrorString(cudaGetLastError()));
    CK(cudaGraphLaunch(g->exec, s));
    if (!fresh) h->launches += g->launches;
    if (plan.order_mode != 2) { h->order_buf = g->order_buf_after; h->order_count_dev = g->order_count_dev_after; }
    h->last_result_buf = g->last_result_buf;
    h->last_keys64 = g->last_keys64;
    h->frame_valid = true;
    return 0;
}

// The asynchronous entry points (gsr_render, gsr_render_batch) cannot know whether the frame they enqueue fits the
// instance buffers; k_tile_scan counts such frames in the sticky words, whose pinned copy the NEXT call looks at:
// it grows the buffers and tells the caller (GSR_EOVERFLOW) that an earlier frame was incomplete.
int repair_sticky_overflow(gsr_t* h, bool* grew) {
    *grew = false;
    const volatile FrameCounters* hc = h->h_counters;
    if (!hc->sticky_overflow) return 0;
    CK(cudaDeviceSynchronize());
    const long long d = std::max<long long>(hc->sticky_max_d, hc->n_instances);
    const long long want = std::max(h->capacity * 2, d + d / 8);
    CK(cudaMemset(&h->d_counters->sticky_overflow, 0, 8));
    h->h_counters->sticky_overflow = 0;
    h->h_counters->sticky_max_d = 0;
    h->h_counters->overflow = 0;
    h->slot_busy[0] = h->slot_busy[1] = false;
    h->capacity_req = 0;
    if (h->capacity >= MAX_CAPACITY) return fail(GSR_EOVERFLOW, "tile-instance buffer overflow at the maximum capacity");
    int rc = ensure_capacity(h, want, h->cap_keys64);
    if (rc) return rc;
    *grew = true;
    return 0;
}

// render + (optional conversion) + D2H, growing the instance buffers on overflow
int render_to_host(gsr_t* h, void* host_dst, int kind /*0 rgba f32, 1 rgb8, 2 r32f*/, cudaStream_t s) {
    if (!host_dst) return fail(GSR_EINVAL, "null output");
    const long long npx = (long long)h->W * h->H;
    for (int attempt = 0; attempt < 4; ++attempt) {
        int rc = render_frame(h, h->fp, h->d_frame, s, false);
        if (rc) return rc;
        if (kind == 0) {
            CK(cudaMemcpyAsync(host_dst, h->d_frame, (size_t)npx * 16, cudaMemcpyDeviceToHost, s));
        } else if (kind == 1) {
            launch_pack_rgb8(h->d_frame, h->d_rgb8, npx, s);
            h->launches += 1;
            CK(cudaMemcpyAsync(host_dst, h->d_rgb8, (size_t)npx * 3, cudaMemcpyDeviceToHost, s));
        } else {
            launch_extract_r(h->d_frame, h->d_r32, npx, s);
            h->launches += 1;
            CK(cudaMemcpyAsync(host_dst, h->d_r32, (size_t)npx * 4, cudaMemcpyDeviceToHost, s));
        }
        CK(cudaStreamSynchronize(s));
        if (!h->h_counters->overflow) {
            // the overflowed attempts of this call (and earlier synchronous ones) are repaired: forget them, or the
            // next pipelined sync point would report them again (nothing asynchronous is in flight on this path)
            if (h->h_counters->sticky_overflow && !h->slot_busy[0] && !h->slot_busy[1]) {
                CK(cudaMemsetAsync(&h->d_counters->sticky_overflow, 0, 8, s));
                CK(cudaStreamSynchronize(s));
                h->h_counters->sticky_overflow = 0;
                h->h_counters->sticky_max_d = 0;
            }
            return 0;
        }
        const long long d = h->h_counters->n_instances;
        if (h->capacity >= MAX_CAPACITY) break;
        long long want = std::max(h->capacity * 2, d + d / 8);
        h->capacity_req = 0;
        rc = ensure_capacity(h, want, h->cap_keys64);
        if (rc) return rc;
    }
    return fail(GSR_EOVERFLOW, "tile-instance buffer overflow");
}

}  // namespace

extern "C" {

const char* gsr_last_error(void) { return g_err.c_str(); }
int gsr_version(void) { return 100; }

int gsr_create(int device, gsr_t** out) {
    if (!out) return fail(GSR_EINVAL, "null out");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0)
        return fail(GSR_ECUDA, std::string("no CUDA device: ") + cudaGetErrorString(e));
    if (device < 0 || device >= count) return fail(GSR_EINVAL, "bad device index");
    CK(cudaSetDevice(device));
    if (radix_sort_prepare_device() || preprocess_prepare_device())
        return fail(GSR_ECUDA, "kernel set-up failed on this device (shared-memory opt-in)");
    gsr_t* h = new gsr_renderer();
    h->device = device;
    CK(cudaMalloc(&h->d_fp, sizeof(FrameParams) * 2));
    CK(cudaMalloc(&h->d_counters, sizeof(FrameCounters)));
    CK(cudaMemset(h->d_counters, 0, sizeof(FrameCounters)));
    CK(cudaMallocHost(&h->h_counters, sizeof(FrameCounters)));
    memset(h->h_counters, 0, sizeof(FrameCounters));
    for (auto& ev : h->ev) CK(cudaEventCreate(&ev));
    CK(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&h->aux_stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));
    for (cudaEvent_t* e : {&h->ev_start, &h->ev_pre, &h->ev_cleared, &h->ev_scan})
        CK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    for (int i = 0; i < 2; ++i) {
        CK(cudaEventCreateWithFlags(&h->ev_packed[i], cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&h->ev_copied[i], cudaEventDisableTiming));
    }
    h->fp.scale_modifier = 1.f;
    h->fp.render_mod = 3;
    h->fp.flags = GSR_FLAG_CLAMP_UNORM;
    *out = h;
    return 0;
}

int gsr_destroy(gsr_t* h) {
    if (!h) return 0;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    free_gaussian_buffers(h);
    free_instance_buffers(h);
    dfree(h->d_ranges); dfree(h->d_tile_grid); dfree(h->d_frame); dfree(h->d_rgb8); dfree(h->d_rgb8b); dfree(h->d_r32);
    dfree(h->d_cap[0]); dfree(h->d_cap[1]);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->aux_stream) cudaStreamDestroy(h->aux_stream);
    drop_graphs(h);
    if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
    for (cudaEvent_t e : {h->ev_start, h->ev_pre, h->ev_cleared, h->ev_scan}) if (e) cudaEventDestroy(e);
    for (int i = 0; i < 2; ++i) { if (h->ev_packed[i]) cudaEventDestroy(h->ev_packed[i]); if (h->ev_copied[i]) cudaEventDestroy(h->ev_copied[i]); }
    dfree(h->d_fp); dfree(h->d_counters); dfree(h->d_sortp);
    if (h->h_counters) cudaFreeHost(h->h_counters);
    for (auto& ev : h->ev) if (ev) cudaEventDestroy(ev);
    delete h;
    return 0;
}

int gsr_alloc_gaussians(gsr_t* h, long long n, int sh_dim) {
    if (!h || n < 0 || sh_dim < 3 || sh_dim > 48 || sh_dim % 3) return fail(GSR_EINVAL, "bad n / sh_dim");
    if (n >= (1ll << 31) - 1024) return fail(GSR_EINVAL, "too many gaussians");
    CK(cudaSetDevice(h->device));
    CK(cudaDeviceSynchronize());
    int rc = alloc_gaussian_buffers(h, n, sh_dim);
    if (rc) return rc;
    h->fp.sh_dim = sh_dim;
    return 0;
}

int gsr_set_gaussians(gsr_t* h, const float* xyz, const float* rot, const float* scale, const float* opacity,
                      const float* sh, long long n, int sh_dim, int src_is_device) {
    if (!h) return fail(GSR_EINVAL, "null handle");
    if (n > 0 && (!xyz || !rot || !scale || !opacity || !sh)) return fail(GSR_EINVAL, "null array");
    int rc = gsr_alloc_gaussians(h, n, sh_dim);
    if (rc) return rc;
    if (n == 0) return 0;
    const cudaMemcpyKind kind = src_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    const size_t n1 = (size_t)n;
    CK(cudaMemcpy(h->d_xyz, xyz, n1 * 12, kind));
    CK(cudaMemcpy(h->d_sh, sh, n1 * (size_t)sh_dim * 4, kind));
    // rot / scale / opacity are packed into two float4 planes on the device
    float *t_rot = nullptr, *t_scale = nullptr, *t_op = nullptr;
    const float *p_rot = rot, *p_scale = scale, *p_op = opacity;
    if (!src_is_device) {
        CK(cudaMalloc(&t_rot, n1 * 16)); CK(cudaMalloc(&t_scale, n1 * 12)); CK(cudaMalloc(&t_op, n1 * 4));
        CK(cudaMemcpy(t_rot, rot, n1 * 16, kind));
        CK(cudaMemcpy(t_scale, scale, n1 * 12, kind));
        CK(cudaMemcpy(t_op, opacity, n1 * 4, kind));
        p_rot = t_rot; p_scale = t_scale; p_op = t_op;
    }
    launch_pack_scene(p_rot, p_scale, p_op, h->d_rot, h->d_scale_op, n, 0);
    h->launches += 1;
    CK(cudaDeviceSynchronize());
    if (t_rot) { cudaFree(t_rot); cudaFree(t_scale); cudaFree(t_op); }
    CK(cudaGetLastError());
    return 0;
}

// load_ply on the device (util_gau.py:63-125): raw vertex rows -> resident scene planes, bbox, center.
int gsr_load_ply_rows(gsr_t* h, const void* rows, long long n, const gsr_ply_layout* layout, int src_is_device,
                      float bbox[6], float center[3], void* stream) {
    if (!h || !layout || (n > 0 && !rows)) return fail(GSR_EINVAL, "null argument");
    if (n < 0) return fail(GSR_EINVAL, "bad vertex count");
    PlyLayout lay;
    lay.row_bytes = layout->row_bytes;
    lay.big_endian = layout->big_endian ? 1 : 0;
    if (lay.row_bytes < 59 * 4 || lay.row_bytes % 4) return fail(GSR_EINVAL, "row_bytes must be a multiple of 4 and hold 59 floats");
    for (int i = 0; i < 59; ++i) {
        const int o = layout->offset[i];
        if (o < 0 || o % 4 || o + 4 > lay.row_bytes) return fail(GSR_EINVAL, "bad property offset");
        lay.off[i] = o;
    }
    if (ply_decode_configure(lay.row_bytes)) return fail(GSR_EINVAL, "vertex row too wide");
    int rc = gsr_alloc_gaussians(h, n, 48);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    int sm_count = 148;
    cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, h->device);
    int* d_red = nullptr;   // 7 ordered ints + 3 running sums | 6 bbox floats + 3 center floats
    CK(cudaMalloc(&d_red, 20 * 4));
    float* d_out = reinterpret_cast<float*>(d_red) + 10;
    launch_ply_bbox_init(d_red, s);
    h->launches += 1;
    const size_t rb = (size_t)lay.row_bytes;
    unsigned char* d_stage[2] = {nullptr, nullptr};
    cudaEvent_t ev_up[2] = {}, ev_done[2] = {};
    int ret = 0;
    if (src_is_device) {
        if (((uintptr_t)rows & 15) != 0) { cudaFree(d_red); return fail(GSR_EINVAL, "device rows must be 16-byte aligned"); }
        launch_ply_decode((const unsigned char*)rows, lay, n, 0, h->d_xyz, h->d_rot, h->d_scale_op, h->d_sh, d_red, sm_count, s);
        if (center) launch_ply_center(h->d_xyz, n, d_red, s);
        h->launches += (n > 0) * (center ? 2 : 1);
    } else if (n > 0) {
        // double-buffered upload: chunk k+1 crosses PCIe on the copy stream while chunk k is decoded
        const long long chunk_rows = std::max<long long>(128, ((64ll << 20) / (long long)rb) / 128 * 128);
        const size_t stage_bytes = ((size_t)chunk_rows * rb + 15) & ~(size_t)15;
        for (int i = 0; i < 2 && !ret; ++i) {
            if (cudaMalloc(&d_stage[i], stage_bytes) != cudaSuccess) ret = GSR_ENOMEM;
            else if (cudaEventCreateWithFlags(&ev_up[i], cudaEventDisableTiming) != cudaSuccess ||
                     cudaEventCreateWithFlags(&ev_done[i], cudaEventDisableTiming) != cudaSuccess) ret = GSR_ECUDA;
        }
        int k = 0;
        for (long long r0 = 0; r0 < n && !ret; r0 += chunk_rows, ++k) {
            const int b = k & 1;
            const long long nr = std::min(chunk_rows, n - r0);
            if (k >= 2) cudaStreamWaitEvent(h->copy_stream, ev_done[b], 0);  // buffer b was decoded
            if (cudaMemcpyAsync(d_stage[b], (const unsigned char*)rows + (size_t)r0 * rb, (size_t)nr * rb,
                                cudaMemcpyHostToDevice, h->copy_stream) != cudaSuccess) { ret = GSR_ECUDA; break; }
            cudaEventRecord(ev_up[b], h->copy_stream);
            cudaStreamWaitEvent(s, ev_up[b], 0);
            launch_ply_decode(d_stage[b], lay, nr, r0, h->d_xyz, h->d_rot, h->d_scale_op, h->d_sh, d_red, sm_count, s);
            cudaEventRecord(ev_done[b], s);
            // the serial running sum of this chunk hides behind the upload of the next one
            if (center) launch_ply_center(h->d_xyz + r0 * 3, nr, d_red, s);
            h->launches += center ? 2 : 1;
        }
    }
    if (!ret) {
        launch_ply_final(d_red, n, d_out, d_out + 6, s);
        h->launches += 1;
        float host[9] = {0};
        if (cudaMemcpyAsync(host, d_out, sizeof(host), cudaMemcpyDeviceToHost, s) != cudaSuccess ||
            cudaStreamSynchronize(s) != cudaSuccess) ret = GSR_ECUDA;
        else {
            if (bbox) memcpy(bbox, host, 6 * sizeof(float));
            if (center) {
                memcpy(center, host + 6, 3 * sizeof(float));
                if (n == 0) center[0] = center[1] = center[2] = nanf("");  // np.mean of an empty axis
            }
        }
    }
    cudaStreamSynchronize(h->copy_stream);
    cudaError_t e = cudaGetLastError();
    for (int i = 0; i < 2; ++i) {
        if (d_stage[i]) cudaFree(d_stage[i]);
        if (ev_up[i]) cudaEventDestroy(ev_up[i]);
        if (ev_done[i]) cudaEventDestroy(ev_done[i]);
    }
    cudaFree(d_red);
    if (ret == GSR_ENOMEM) return fail(GSR_ENOMEM, "staging buffers for the PLY upload");
    if (ret || e != cudaSuccess) return fail(GSR_ECUDA, std::string("PLY decode: ") + cudaGetErrorString(e));
    return 0;
}

// inverse of gsr_set_gaussians: resident planes -> GaussianData's five host arrays
int gsr_get_gaussians(gsr_t* h, float* xyz, float* rot, float* scale, float* opacity, float* sh) {
    if (!h) return fail(GSR_EINVAL, "null handle");
    if (!h->d_xyz) return fail(GSR_EINVAL, "no gaussians set");
    const size_t n = (size_t)h->N;
    if (n == 0) return 0;
    CK(cudaSetDevice(h->device));
    CK(cudaDeviceSynchronize());
    if (xyz) CK(cudaMemcpy(xyz, h->d_xyz, n * 12, cudaMemcpyDeviceToHost));
    if (rot) CK(cudaMemcpy(rot, h->d_rot, n * 16, cudaMemcpyDeviceToHost));
    if (scale) CK(cudaMemcpy2D(scale, 12, h->d_scale_op, 16, 12, n, cudaMemcpyDeviceToHost));
    if (opacity) CK(cudaMemcpy2D(opacity, 4, reinterpret_cast<const char*>(h->d_scale_op) + 12, 16, 4, n, cudaMemcpyDeviceToHost));
    if (sh) CK(cudaMemcpy(sh, h->d_sh, n * (size_t)h->sh_dim * 4, cudaMemcpyDeviceToHost));
    return 0;
}

int gsr_scene_buffer(gsr_t* h, int which, void** dev_ptr, long long* bytes) {
    if (!h || !dev_ptr || !bytes) return fail(GSR_EINVAL, "null argument");
    const long long n = h->N;
    switch (which) {
        case 0: *dev_ptr = h->d_xyz; *bytes = n * 12; break;
        case 1: *dev_ptr = h->d_rot; *bytes = n * 16; break;
        case 2: *dev_ptr = h->d_scale_op; *bytes = n * 16; break;
        case 3: *dev_ptr = h->d_sh; *bytes = n * (long long)h->sh_dim * 4; break;
        default: return fail(GSR_EINVAL, "bad buffer selector");
    }
    return 0;
}

int gsr_set_resolution(gsr_t* h, int width, int height) {
    if (!h || width <= 0 || height <= 0 || width > 65535 || height > 65535) return fail(GSR_EINVAL, "bad resolution");
    CK(cudaSetDevice(h->device));
    if (width == h->W && height == h->H) return 0;
    CK(cudaDeviceSynchronize());
    drop_graphs(h);
    dfree(h->d_ranges); dfree(h->d_tile_grid); dfree(h->d_frame); dfree(h->d_rgb8); dfree(h->d_rgb8b); dfree(h->d_r32);
    dfree(h->d_cap[0]); dfree(h->d_cap[1]);
    h->slot_busy[0] = h->slot_busy[1] = false;
    h->W = width; h->H = height;
    h->tiles_x = (width + GSR_TILE - 1) / GSR_TILE;
    h->tiles_y = (height + GSR_TILE - 1) / GSR_TILE;
    const size_t npx = (size_t)width * height;
    CK(cudaMalloc(&h->d_ranges, sizeof(int2) * (size_t)h->tiles_x * h->tiles_y));
    h->grid_w = (h->tiles_x + 1) | 1;  // odd pitch: conflict-free column walks in k_tile_scan
    CK(cudaMalloc(&h->d_tile_grid, sizeof(int) * (size_t)h->grid_w * (h->tiles_y + 1) * GSR_GRID_REPLICAS));
    CK(cudaMalloc(&h->d_frame, npx * 16));
    CK(cudaMalloc(&h->d_rgb8, npx * 3));
    CK(cudaMalloc(&h->d_rgb8b, npx * 3));
    CK(cudaMalloc(&h->d_r32, npx * 4));
    h->fp.W = width; h->fp.H = height; h->fp.tiles_x = h->tiles_x; h->fp.tiles_y = h->tiles_y;
    h->frame_valid = false;
    return 0;
}

int gsr_set_camera(gsr_t* h, const float view[16], const float proj[16], const float campos[3], float htanx,
                   float htany, float focal) {
    if (!h || !view || !proj || !campos) return fail(GSR_EINVAL, "null argument");
    memcpy(h->fp.view, view, 64);
    memcpy(h->fp.proj, proj, 64);
    memcpy(h->fp.campos, campos, 12);
    h->fp.htanx = htanx; h->fp.htany = htany; h->fp.focal = focal;
    h->have_camera = true;
    return 0;
}

int gsr_set_params(gsr_t* h, float scale_modifier, int render_mod, unsigned flags) {
    if (!h) return fail(GSR_EINVAL, "null handle");
    if (render_mod < -4) return fail(GSR_EINVAL, "render_mod must be >= -4");
    h->fp.scale_modifier = scale_modifier;
    h->fp.render_mod = render_mod;
    h->fp.flags = flags;
    return 0;
}

int gsr_set_capacity(gsr_t* h, long long max_instances) {
    if (!h || max_instances < 0 || max_instances > MAX_CAPACITY) return fail(GSR_EINVAL, "bad capacity");
    CK(cudaSetDevice(h->device));
    h->capacity_req = max_instances;
    if (max_instances > 0 && max_instances != h->capacity) {
        CK(cudaDeviceSynchronize());
        const bool k64 = h->cap_keys64;
        free_instance_buffers(h);
        h->cap_keys64 = false;
        return ensure_capacity(h, max_instances, k64);
    }
    return 0;
}

int gsr_render(gsr_t* h, float* out_rgba, void* stream) {
    if (!h) return fail(GSR_EINVAL, "null handle");
    CK(cudaSetDevice(h->device));
    h->fp.sh_dim = h->sh_dim;
    bool grew = false;
    int rc = repair_sticky_overflow(h, &grew);
    if (rc) return rc;
    if (grew) return fail(GSR_EOVERFLOW, "an earlier frame overflowed the tile-instance buffers and is incomplete (grown now): draw again");
    return render_frame(h, h->fp, out_rgba, (cudaStream_t)stream, false);
}

int gsr_render_timed(gsr_t* h, float* out_rgba, void* stream) {
    if (!h) return fail(GSR_EINVAL, "null handle");
    CK(cudaSetDevice(h->device));
    h->fp.sh_dim = h->sh_dim;
    cudaStream_t s = (cudaStream_t)stream;
    int rc = render_frame(h, h->fp, out_rgba, s, true);
    if (rc) return rc;
    CK(cudaStreamSynchronize(s));
    float ms[6];
    for (int i = 0; i < 6; ++i) CK(cudaEventElapsedTime(&ms[i], h->ev[i], h->ev[i + 1]));
    h->stats.ms_preprocess = ms[0];
    h->stats.ms_sort_depth = ms[1];
    h->stats.ms_ranges = ms[2];  // k_tile_scan: counts -> ranges + sort histograms
    h->stats.ms_duplicate = ms[3];
    h->stats.ms_sort_tiles = ms[4];
    h->stats.ms_blend = ms[5];
    CK(cudaEventElapsedTime(&h->stats.ms_total, h->ev[0], h->ev[6]));
    return 0;
}

int gsr_render_host(gsr_t* h, float* out, void* stream) {
    if (!h) return fail(GSR_EINVAL, "null handle");
    CK(cudaSetDevice(h->device));
    h->fp.sh_dim = h->sh_dim;
    return render_to_host(h, out, 0, (cudaStream_t)stream);
}
int gsr_render_rgb8_host(gsr_t* h, uint8_t* out, void* stream) {
    if (!h) return fail(GSR_EINVAL, "null handle");
    CK(cudaSetDevice(h->device));
    h->fp.sh_dim = h->sh_dim;
    return render_to_host(h, out, 1, (cudaStream_t)stream);
}
int gsr_render_r32f_host(gsr_t* h, float* out, void* stream) {
    if (!h) return fail(GSR_EINVAL, "null handle");
    CK(cudaSetDevice(h->device));
    h->fp.sh_dim = h->sh_dim;
    return render_to_host(h, out, 2, (cudaStream_t)stream);
}

namespace {
// waits for an in-flight read-back slot; reports (and repairs) an instance-buffer overflow seen since
int finish_slot(gsr_t* h, int slot) {
    if (!h->slot_busy[slot]) return 0;
    CK(cudaEventSynchronize(h->ev_copied[slot]));
    h->slot_busy[slot] = false;
    // the sticky words count every overflowed frame since the last clear, whichever frame the pinned copy is from
    if (h->h_counters->sticky_overflow) {
        const long long d = std::max<long long>(h->h_counters->sticky_max_d, h->h_counters->n_instances);
        const long long want = std::max(h->capacity * 2, d + d / 8);
        h->capacity_req = 0;
        CK(cudaDeviceSynchronize());
        CK(cudaMemset(&h->d_counters->sticky_overflow, 0, 8));
        h->h_counters->sticky_overflow = 0;
        h->h_counters->sticky_max_d = 0;
        h->h_counters->overflow = 0;
        h->slot_busy[0] = h->slot_busy[1] = false;
        int rc = ensure_capacity(h, want, h->cap_keys64);
        if (rc) return rc;
        return fail(GSR_EOVERFLOW, "a pipelined frame overflowed the tile-instance buffers (grown now): submit the frames in flight again");
    }
    return 0;
}
}  // namespace

int gsr_render_rgb8_host_async(gsr_t* h, uint8_t* out, void* stream) {
    if (!h || !out) return fail(GSR_EINVAL, "null argument");
    CK(cudaSetDevice(h->device));
    h->fp.sh_dim = h->sh_dim;
    cudaStream_t s = (cudaStream_t)stream;
    const int slot = (int)(h->async_idx++ & 1u);
    int rc = finish_slot(h, slot);
    if (rc) return rc;
    const long long npx = (long long)h->W * h->H;
    uint8_t* dev = slot ? h->d_rgb8b : h->d_rgb8;
    rc = render_frame(h, h->fp, h->d_frame, s, false, 0, h->tune.fused_rgb8 ? dev : nullptr);  // the blend kernel packs RGB8 itself
    if (rc) return rc;
    if (!h->tune.fused_rgb8) {
        launch_pack_rgb8(h->d_frame, dev, npx, s);
        h->launches += 1;
    }
    CK(cudaEventRecord(h->ev_packed[slot], s));
    CK(cudaStreamWaitEvent(h->copy_stream, h->ev_packed[slot], 0));
    CK(cudaMemcpyAsync(out, dev, (size_t)npx * 3, cudaMemcpyDeviceToHost, h->copy_stream));
    CK(cudaEventRecord(h->ev_copied[slot], h->copy_stream));
    h->slot_busy[slot] = true;
    return 0;
}

// One pose of the stereo capture loop (main.py:844-917): left RGB8, disparity u16 at the left pose, right RGB8.
int gsr_capture_stereo_async(gsr_t* h, const float view_left[16], const float campos_left[3], const float view_right[16],
                             const float campos_right[3], uint8_t* left_rgb, uint16_t* disparity, uint8_t* right_rgb,
                             int share_order, void* stream) {
    if (!h || !view_left || !campos_left) return fail(GSR_EINVAL, "null argument");
    if (right_rgb && (!view_right || !campos_right)) return fail(GSR_EINVAL, "right image without a right pose");
    CK(cudaSetDevice(h->device));
    if (h->W <= 0) return fail(GSR_EINVAL, "no resolution set");
    h->fp.sh_dim = h->sh_dim;
    cudaStream_t s = (cudaStream_t)stream;
    const int slot = (int)(h->async_idx++ & 1u);
    int rc = finish_slot(h, slot);
    if (rc) return rc;
    const size_t npx = (size_t)h->W * h->H;
    if (!h->d_cap[slot]) CK(cudaMalloc(&h->d_cap[slot], npx * 8));
    uint8_t* d_left = h->d_cap[slot];
    uint8_t* d_right = d_left + npx * 3;
    uint16_t* d_disp = reinterpret_cast<uint16_t*>(d_left + npx * 6);
    // the three renders have the same view-space depth per Gaussian (same view; the right view differs by an
    // x translation in view space, main.py:376-380), so one depth order serves all of them
    int mode = share_order ? 1 : 0;
    FrameParams fp = h->fp;
    memcpy(fp.view, view_left, 64);
    memcpy(fp.campos, campos_left, 12);
    if (left_rgb) {
        fp.flags |= GSR_FLAG_CLAMP_UNORM;  // RGB8 FBO (main.py:846-852)
        rc = render_frame(h, fp, h->d_frame, s, false, mode, h->tune.fused_rgb8 ? d_left : nullptr);
        if (rc) return rc;
        if (mode) mode = 2;
        if (!h->tune.fused_rgb8) { launch_pack_rgb8(h->d_frame, d_left, (long long)npx, s); h->launches += 1; }
    }
    if (disparity) {
        FrameParams fd = fp;
        fd.render_mod = -1;                  // main.py:866
        fd.flags &= ~GSR_FLAG_CLAMP_UNORM;   // R32F FBO (main.py:858-870)
        rc = render_frame(h, fd, h->d_frame, s, false, mode);
        if (rc) return rc;
        if (mode) mode = 2;
        launch_pack_disp_u16(h->d_frame, d_disp, (long long)npx, s);
        h->launches += 1;
    }
    if (right_rgb) {
        FrameParams fr = fp;
        fr.flags |= GSR_FLAG_CLAMP_UNORM;
        memcpy(fr.view, view_right, 64);
        memcpy(fr.campos, campos_right, 12);
        if (mode == 2) {  // the shared order needs equal depth rows; otherwise sort again
            for (int k = 8; k < 12; ++k) if (fr.view[k] != fp.view[k]) mode = 0;
        }
        rc = render_frame(h, fr, h->d_frame, s, false, mode, h->tune.fused_rgb8 ? d_right : nullptr);
        if (rc) return rc;
        if (!h->tune.fused_rgb8) { launch_pack_rgb8(h->d_frame, d_right, (long long)npx, s); h->launches += 1; }
    }
    if (!h->order_frozen) h->order_buf = -1;
    CK(cudaEventRecord(h->ev_packed[slot], s));
    CK(cudaStreamWaitEvent(h->copy_stream, h->ev_packed[slot], 0));
    if (left_rgb) CK(cudaMemcpyAsync(left_rgb, d_left, npx * 3, cudaMemcpyDeviceToHost, h->copy_stream));
    if (disparity) CK(cudaMemcpyAsync(disparity, d_disp, npx * 2, cudaMemcpyDeviceToHost, h->copy_stream));
    if (right_rgb) CK(cudaMemcpyAsync(right_rgb, d_right, npx * 3, cudaMemcpyDeviceToHost, h->copy_stream));
    CK(cudaEventRecord(h->ev_copied[slot], h->copy_stream));
    h->slot_busy[slot] = true;
    return 0;
}

int gsr_host_sync(gsr_t* h) {
    if (!h) return fail(GSR_EINVAL, "null handle");
    CK(cudaSetDevice(h->device));
    int rc0 = finish_slot(h, 0);
    int rc1 = finish_slot(h, 1);
    return rc0 ? rc0 : rc1;
}

int gsr_render_batch(gsr_t* h, const float* views, const float* campos, int n_views, float* out_rgba, void* stream) {
    if (!h || !views || !campos || n_views < 0) return fail(GSR_EINVAL, "bad argument");
    CK(cudaSetDevice(h->device));
    if (!h->have_camera) return fail(GSR_EINVAL, "set intrinsics with gsr_set_camera first");
    h->fp.sh_dim = h->sh_dim;
    bool grew = false;
    int rc0 = repair_sticky_overflow(h, &grew);
    if (rc0) return rc0;
    if (grew) return fail(GSR_EOVERFLOW, "an earlier frame overflowed the tile-instance buffers and is incomplete (grown now): draw again");
    const size_t frame = (size_t)h->W * h->H * 4;
    FrameParams fp = h->fp;
    for (int v = 0; v < n_views; ++v) {
        memcpy(fp.view, views + 16 * (size_t)v, 64);
        memcpy(fp.campos, campos + 3 * (size_t)v, 12);
        float* dst = out_rgba ? out_rgba + frame * (size_t)v : h->d_frame;
        int rc = render_frame(h, fp, dst, (cudaStream_t)stream, false);
        if (rc) return rc;
    }
    return 0;
}

int gsr_sort_depth(gsr_t* h, const float view[16], int32_t* out_index, int dst_is_device, void* stream) {
    if (!h || !view || !out_index) return fail(GSR_EINVAL, "null argument");
    if (!h->d_xyz) return fail(GSR_EINVAL, "no gaussians set");
    CK(cudaSetDevice(h->device));
    cudaStream_t s = (cudaStream_t)stream;
    const long long N = h->N;
    if (N == 0) return 0;
    launch_depth_keys(h->d_xyz, view, h->d_dkey[0], N, s);
    h->launches += 1;
    SortOptions so_n = {};
    so_n.hist_direct_mask = 0x3u;  // mantissa bytes counted directly, see render_frame
    const int r = radix_sort_pairs<uint32_t>(h->d_dkey[0], h->d_dkey[0], h->d_dkey[1], h->d_gid[0], h->d_gid[1], true, N,
                                             nullptr, N, 0, 32, &h->ss_n, s, &h->launches, &so_n);
    CK(cudaMemcpyAsync(out_index, h->d_gid[r], (size_t)N * 4,
                       dst_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, s));
    if (!dst_is_device) CK(cudaStreamSynchronize(s));
    h->frame_valid = false;
    h->order_buf = -1;
    CK(cudaGetLastError());
    // the sort borrowed the buffers a frozen order lives in: rebuild that order (one more sort of a rare call)
    if (h->order_frozen) {
        float fv[16];
        memcpy(fv, h->frozen_view, 64);
        return gsr_freeze_order(h, fv, stream);
    }
    return 0;
}

// The OpenGL backend draws in the order of the LAST sort_and_update, whatever the camera has done since
// (renderer_ogl.py:139-146; main.py:989-997: a "sort" button, or every frame when "auto sort" is ticked).
int gsr_freeze_order(gsr_t* h, const float view[16], void* stream) {
    if (!h || !view) return fail(GSR_EINVAL, "null argument");
    if (!h->d_xyz) return fail(GSR_EINVAL, "no gaussians set");
    CK(cudaSetDevice(h->device));
    cudaStream_t s = (cudaStream_t)stream;
    const long long N = h->N;
    h->order_buf = 0;
    h->order_frozen = true;
    h->order_count_dev = false;
    h->frame_valid = false;
    memcpy(h->frozen_view, view, 64);
    if (N == 0) return 0;
    launch_depth_keys(h->d_xyz, view, h->d_dkey[0], N, s, 1);
    h->launches += 1;
    SortOptions so_n = {};
    so_n.hist_direct_mask = 0x3u;
    h->order_buf = radix_sort_pairs<uint32_t>(h->d_dkey[0], h->d_dkey[0], h->d_dkey[1], h->d_gid[0], h->d_gid[1], true, N,
                                              nullptr, N, 0, 32, &h->ss_n, s, &h->launches, &so_n);
    CK(cudaGetLastError());
    return 0;
}

int gsr_thaw_order(gsr_t* h) {
    if (!h) return fail(GSR_EINVAL, "null handle");
    h->order_frozen = false;
    h->order_buf = -1;
    return 0;
}

int gsr_sort_pairs(gsr_t* h, int key_bits, void* keys, void* keys_tmp, uint32_t* vals, uint32_t* vals_tmp, long long n,
                   int begin_bit, int end_bit, void* stream) {
    if (!h) return fail(GSR_EINVAL, "null handle");
    if (n == 0) return 0;
    if (!keys || !keys_tmp || !vals_tmp) return fail(GSR_EINVAL, "null argument");
    if (key_bits != 32 && key_bits != 64) return fail(GSR_EINVAL, "key_bits must be 32 or 64");
    if (begin_bit < 0 || end_bit > key_bits || begin_bit > end_bit) return fail(GSR_EINVAL, "bad bit range");
    if (n < 0 || n > MAX_CAPACITY) return fail(GSR_EINVAL, "bad n");
    CK(cudaSetDevice(h->device));
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0 || begin_bit == end_bit) return 0;
    const int kb = key_bits / 8;
    if (h->sortp_items < n || h->sortp_keybytes != kb) {
        CK(cudaDeviceSynchronize());
        dfree(h->d_sortp);
        CK(cudaMalloc(&h->d_sortp, sort_scratch_bytes(n, kb)));
        h->sortp_items = n;
        h->sortp_keybytes = kb;
    }
    SortScratch ss;
    sort_scratch_carve(&ss, h->d_sortp, h->sortp_items, kb);
    const bool identity = (vals == nullptr);
    uint32_t* vals_a = vals;
    int r;
    if (identity) {
        // results must land somewhere: with identity input the caller passes the OUTPUT buffer as vals_tmp
        // and we need a second value buffer for the ping-pong; borrow it from the key scratch size class
        return fail(GSR_EINVAL, "vals == NULL (identity) is only supported through gsr_sort_depth");
    }
    if (kb == 4)
        r = radix_sort_pairs<uint32_t>((uint32_t*)keys, (uint32_t*)keys, (uint32_t*)keys_tmp, vals_a, vals_tmp, false, n,
                                       nullptr, n, begin_bit, end_bit, &ss, s, &h->launches);
    else
        r = radix_sort_pairs<uint64_t>((uint64_t*)keys, (uint64_t*)keys, (uint64_t*)keys_tmp, vals_a, vals_tmp, false, n,
                                       nullptr, n, begin_bit, end_bit, &ss, s, &h->launches);
    if (r == 1) {  // odd number of passes: bring the result back to keys / vals
        CK(cudaMemcpyAsync(keys, keys_tmp, (size_t)n * kb, cudaMemcpyDeviceToDevice, s));
        CK(cudaMemcpyAsync(vals, vals_tmp, (size_t)n * 4, cudaMemcpyDeviceToDevice, s));
    }
    CK(cudaGetLastError());
    return 0;
}

// Allocation-free variant of gsr_sort_pairs (no counterpart in the reference):
// the scratch of the radix passes is a static device array inside libgsr.so.
int gsr_sort_pairs_static(int key_bits, void* keys, void* keys_tmp, uint32_t* vals, uint32_t* vals_tmp, long long n,
                          int begin_bit, int end_bit, void* stream) {
    if (key_bits != 32 && key_bits != 64) return fail(GSR_EINVAL, "key_bits must be 32 or 64");
    if (n < 0) return fail(GSR_EINVAL, "bad n");
    if (n > GSR_SORT_STATIC_MAX_PAIRS)
        return fail(GSR_EINVAL, "gsr_sort_pairs_static: n exceeds GSR_SORT_STATIC_MAX_PAIRS (4194304); use gsr_sort_pairs");
    if (begin_bit < 0 || end_bit > key_bits || begin_bit > end_bit) return fail(GSR_EINVAL, "bad bit range");
    if (n == 0 || begin_bit == end_bit) return 0;
    if (!keys || !keys_tmp || !vals || !vals_tmp) return fail(GSR_EINVAL, "null argument");
    const char* err = "";
    const int rc = radix_sort_pairs_static(key_bits, keys, keys_tmp, vals, vals_tmp, n, begin_bit, end_bit, (cudaStream_t)stream, &err);
    if (rc == -1) return fail(GSR_EINVAL, err);
    if (rc) return fail(GSR_ECUDA, std::string("gsr_sort_pairs_static: ") + err);
    return 0;
}

int gsr_get_stats(gsr_t* h, gsr_stats* out) {
    if (!h || !out) return fail(GSR_EINVAL, "null argument");
    CK(cudaSetDevice(h->device));
    CK(cudaDeviceSynchronize());
    h->stats.n_gaussians = h->N;
    h->stats.n_visible = h->h_counters->n_visible;
    h->stats.n_instances = h->h_counters->n_instances;
    h->stats.overflow = (int)h->h_counters->overflow;
    h->stats.capacity = h->capacity;
    h->stats.tiles_x = h->tiles_x;
    h->stats.tiles_y = h->tiles_y;
    *out = h->stats;
    return 0;
}

int gsr_debug_fetch(gsr_t* h, int which, void* dst, long long capacity_bytes, long long* out_bytes) {
    if (!h || !out_bytes) return fail(GSR_EINVAL, "null argument");
    CK(cudaSetDevice(h->device));
    CK(cudaDeviceSynchronize());
    if (!h->frame_valid) return fail(GSR_EINVAL, "no frame rendered since the last state change");
    const long long D = std::min<long long>(h->h_counters->n_instances, h->capacity);
    const void* src = nullptr;
    long long bytes = 0;
    switch (which) {
        case GSR_DBG_VSOUT:
            if (!h->sp.dbg_vsout) return fail(GSR_EINVAL, "GSR_FLAG_DEBUG was not set for the last frame");
            src = h->sp.dbg_vsout; bytes = h->N * 64; break;
        case GSR_DBG_SETUP:
            if (!h->sp.dbg_setup) return fail(GSR_EINVAL, "GSR_FLAG_DEBUG was not set for the last frame");
            src = h->sp.dbg_setup; bytes = h->N * 32; break;
        case GSR_DBG_INST_KEYS:
            bytes = D * 8;
            if (h->last_keys64) src = h->d_ikeys[h->last_result_buf];
            else if (D > 0 && dst) {
                dfree(h->d_dbgkeys);
                CK(cudaMalloc(&h->d_dbgkeys, (size_t)D * 8));
                launch_keys64_from_ranges(h->d_ranges, h->tiles_x * h->tiles_y, h->d_ivals[h->last_result_buf],
                                          h->sp.depth_key, h->d_dbgkeys, 0);
                CK(cudaDeviceSynchronize());
                src = h->d_dbgkeys;
            }
            break;
        case GSR_DBG_INST_VALS: src = h->d_ivals[h->last_result_buf]; bytes = D * 4; break;
        case GSR_DBG_RANGES: src = h->d_ranges; bytes = (long long)h->tiles_x * h->tiles_y * 8; break;
        default: return fail(GSR_EINVAL, "bad selector");
    }
    *out_bytes = bytes;
    if (!dst) return 0;
    if (capacity_bytes < bytes) return fail(GSR_EINVAL, "destination too small");
    if (bytes > 0 && src) CK(cudaMemcpy(dst, src, (size_t)bytes, cudaMemcpyDeviceToHost));
    return 0;
}

long long gsr_launch_count(gsr_t* h) { return h ? h->launches : 0; }

int gsr_measure_peaks(gsr_t* h, float* fp32_tflops, float* ex2_tops, float* copy_gbs) {
    if (!h || !fp32_tflops || !ex2_tops || !copy_gbs) return fail(GSR_EINVAL, "null argument");
    CK(cudaSetDevice(h->device));
    CK(cudaDeviceSynchronize());
    if (measure_peaks_impl(fp32_tflops, ex2_tops, copy_gbs) != 0) return fail(GSR_ECUDA, "peak micro-benchmark failed");
    return 0;
}

int gsr_tune(gsr_t* h, int key, int value) {
    if (!h) return fail(GSR_EINVAL, "null handle");
    switch (key) {
        case GSR_TUNE_FUSED_DEPTH_HIST: h->tune.fast_depth_sort = value ? 1 : 0; return 0;
        case GSR_TUNE_PDL: h->tune.pdl = value; return 0;
        case GSR_TUNE_FUSED_RGB8: h->tune.fused_rgb8 = value ? 1 : 0; return 0;
        case GSR_TUNE_FRAME_FORK: h->tune.frame_fork = value ? 1 : 0; return 0;
        case GSR_TUNE_GRAPH: h->tune.graph = value ? 1 : 0; return 0;
        case GSR_TUNE_TILE:
            if (value != 16 && value != 32) return fail(GSR_EINVAL, "tile edge must be 16 or 32");
            h->tile_req = value;
            h->frame_valid = false;
            return 0;
        default: return fail(GSR_EINVAL, "unknown (or retired) tuning key");
    }
}

int gsr_count_pairs(gsr_t* h, unsigned long long* visited, unsigned long long* contributed) {
    if (!h || !visited || !contributed) return fail(GSR_EINVAL, "null argument");
    CK(cudaSetDevice(h->device));
    CK(cudaDeviceSynchronize());
    if (!h->frame_valid) return fail(GSR_EINVAL, "no frame rendered since the last state change");
    unsigned long long* d = nullptr;
    CK(cudaMalloc(&d, 16));
    CK(cudaMemset(d, 0, 16));
    launch_blend_count(h->d_fp, h->d_ranges, h->d_ivals[h->last_result_buf], h->sp, h->d_frame, h->fp.render_mod,
                       h->fp.flags, h->tiles_x, h->tiles_y, d, 0, h->tile);
    unsigned long long out[2] = {0, 0};
    CK(cudaMemcpy(out, d, 16, cudaMemcpyDeviceToHost));
    cudaFree(d);
    *visited = out[0];
    *contributed = out[1];
    return 0;
}

}  // extern "C"
