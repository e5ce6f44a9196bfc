// K5 — per-16x16-tile front-to-back alpha blend for sm_100a.
//
// One CTA per tile.  A thread owns PPT vertically adjacent pixels of one column (PPT = 4: 64 threads,
// a warp covers a compact 16x8 block), so the three shared-memory reads of a splat record and the
// x-dependent half of the quadratic form are paid once per PPT pixels.  The tile's instance list
// (nearest first) is streamed through shared memory in double-buffered batches with cp.async
// (LDGSTS): while batch b is blended, batch b+1's records are already in flight and batch b+2's
// Gaussian indices are being read.  The per-pixel body is branch-free (predicated), which keeps all 32
// lanes busy: the fragment rule of /root/reference/shaders/gau_frag.glsl:13-39
//     power = -0.5 (A dx^2 + C dy^2) - B dx dy ; discard if power > 0
//     alpha = min(0.99, opacity * exp(power))  ; discard if alpha < 1/255
// is evaluated as ex2 of a conic pre-scaled by -0.5*log2(e) / -log2(e) in the preprocess kernel, and
// composited front to back, which equals the reference's back-to-front SRC_ALPHA / ONE_MINUS_SRC_ALPHA
// blend over a black clear (renderer_ogl.py:128-130, main.py:821):
//     C = sum_i c_i a_i T_i,  T_i = prod_{j<i} (1 - a_j).
// A pixel stops once T < 1e-4 (its remaining contribution is below 1e-4 of any colour); a warp skips a
// batch when all its pixels have stopped and the CTA leaves when every pixel has.
// The quad test |dx| <= hx && |dy| <= hy is the oracle's coverage rule verbatim (pixel centre minus
// splat centre against the 3-sigma half extents, gau_vert.glsl:174-178).
// Modes: >= 0 SH colour, -1 disparity, -2 billboard, -3 flat ball, -4 gaussian ball.
#include "gsr_internal.h"

namespace {

constexpr int BL_BATCH = 128;  // records per shared-memory batch (64 and 256 measured 2-4 % slower)
constexpr int BL_RUN = 8;      // hits between two checks whether the warp's pixels are all finished (4: slower, 16: same)
struct TagTrue { static constexpr bool value = true; };
struct TagFalse { static constexpr bool value = false; };

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ float4 lds128(unsigned a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ unsigned lds16(unsigned a) {
    unsigned v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}

// Packed binary32 pairs (sm_100 FFMA2 / FMUL2 / FADD2): one issue slot does the IEEE operation on two
// pixels.  The blend is issue bound, and its FMA-pipe half packs; results are those of the scalar
// instructions bit for bit (round-to-nearest, no contraction on either side).
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float lo2(f32x2 v) { return __uint_as_float((unsigned)(v & 0xffffffffull)); }
__device__ __forceinline__ float hi2(f32x2 v) { return __uint_as_float((unsigned)(v >> 32)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// MODE: 0 = gaussian falloff (render_mod >= 0 or -1), 2 = billboard (-2), 3 = flat ball (-3), 4 = gaussian ball (-4),
// 5 = the official rasterizer's renderCUDA rule (GSR_FLAG_OFFICIAL; diff_gaussian_rasterization forward.cu, the
// kernel renderer_cuda.py:212-224 calls): pixel coordinates are integers, rows run top-down, every pixel of a
// touched tile evaluates the splat (the record carries infinite half extents), `power > 0` is always tested, a
// pixel stops BEFORE blending the splat that would take T * (1 - alpha) below 1e-4, and alpha is written as 1.
template <int MODE, int PPT, bool COUNT, bool SHAPE8, int SKIP, int TILE = 16>  // SKIP: 0 none, 1 warp vote per splat, 2 warp filter per batch
__global__ void __launch_bounds__(TILE * TILE / PPT) k_blend(const FrameParams* __restrict__ fp, const int2* __restrict__ ranges,
                                                     const uint32_t* __restrict__ inst_gid, SplatBuffers sp,
                                                     float* __restrict__ out_rgba, unsigned flags,
                                                     unsigned long long* __restrict__ pair_counts,
                                                     uint8_t* __restrict__ out_rgb8) {
    constexpr int THREADS = TILE * TILE / PPT;
    constexpr int PER_THREAD = (BL_BATCH + THREADS - 1) / THREADS;  // records staged per thread per batch
    // one record = three consecutive float4 (48 B): the three shared-memory reads of a splat are constant
    // offsets from one running pointer
    __shared__ __align__(16) float4 s_rec[2][BL_BATCH][3];
    __shared__ uint16_t s_hit[(SKIP >= 2) ? THREADS / 32 : 1][BL_BATCH + 2];  // per-warp hit lists (SKIP >= 2); + 2: the walk reads one entry ahead
    pdl_wait();  // may be resident behind the last radix pass: no global memory access before this
    const int W = fp->W, H = fp->H;
    const int tile = blockIdx.y * gridDim.x + blockIdx.x;
    // pixel block of a warp: SHAPE8 = 8 columns x (4 * PPT) rows (compact: better quad / termination
    // coherence), else 16 columns x (2 * PPT) rows
    int lx, lyb;
    if (SHAPE8) {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        constexpr int WARPS_Y = TILE / (4 * PPT);  // warps stacked vertically per 8-column strip
        lx = (warp / WARPS_Y) * 8 + (lane & 7);
        lyb = (warp % WARPS_Y) * 4 + (lane >> 3);
    } else {
        lx = threadIdx.x & (TILE - 1);
        lyb = threadIdx.x / TILE;
    }
    // pixel-centre bounds of the warp's block, widened by 0.01 px (SKIP == 2)
    float bx0, bx1, by0, by1;
    {
        const int warp = threadIdx.x >> 5;
        int wx0, wy0, ww, wh;
        if (SHAPE8) {
            constexpr int WARPS_Y = TILE / (4 * PPT);
            wx0 = (warp / WARPS_Y) * 8; wy0 = (warp % WARPS_Y) * 4 * PPT; ww = 8; wh = 4 * PPT;
        } else {
            wx0 = 0; wy0 = warp * (32 / TILE) * PPT; ww = TILE; wh = (32 / TILE) * PPT;
        }
        constexpr float C0 = (MODE == 5) ? 0.f : 0.5f;
        bx0 = (float)(blockIdx.x * TILE + wx0) + C0 - 0.01f;
        bx1 = (float)(blockIdx.x * TILE + wx0 + ww - 1) + C0 + 0.01f;
        by0 = (float)(blockIdx.y * TILE + wy0) + C0 - 0.01f;
        by1 = (float)(blockIdx.y * TILE + wy0 + wh - 1) + C0 + 0.01f;
    }
    const int px = blockIdx.x * TILE + lx;
    const int py0 = blockIdx.y * TILE + lyb * PPT;  // GL row, 0 = bottom
    constexpr float CENTRE = (MODE == 5) ? 0.f : 0.5f;
    const float pxf = (float)px + CENTRE;
    float pyf[PPT], T[PPT], cr[PPT], cg[PPT], cb[PPT];
#pragma unroll
    for (int k = 0; k < PPT; ++k) {
        pyf[k] = (float)(py0 + k) + CENTRE;
        // pixels outside the image start terminated
        T[k] = (px < W && py0 + k < H) ? 1.f : 0.f;
        cr[k] = cg[k] = cb[k] = 0.f;
    }
    bool outside[PPT];
#pragma unroll
    for (int k = 0; k < PPT; ++k) outside[k] = !(px < W && py0 + k < H);
    const int2 range = ranges[tile];
    const bool clamp = (flags & 1u) != 0u;
    unsigned n_eval = 0, n_hit = 0;  // COUNT only
    const int n_batches = (range.y - range.x + BL_BATCH - 1) / BL_BATCH;

    // ---- staging helpers: thread t stages records t, t + THREADS, ... of a batch
    uint32_t gid_next[PER_THREAD];  // Gaussian indices of the batch after the one in flight
    auto load_gids = [&](int b) {
#pragma unroll
        for (int i = 0; i < PER_THREAD; ++i) {
            const int slot = threadIdx.x + i * THREADS;
            const int idx = range.x + b * BL_BATCH + slot;
            gid_next[i] = (slot < BL_BATCH && idx < range.y) ? __ldg(inst_gid + idx) : 0xFFFFFFFFu;
        }
    };
    auto issue_batch = [&](int buf) {
#pragma unroll
        for (int i = 0; i < PER_THREAD; ++i) {
            const int slot = threadIdx.x + i * THREADS;
            const uint32_t g = gid_next[i];
            if (g != 0xFFFFFFFFu) {
                cp_async16(&s_rec[buf][slot][0], sp.rec0 + g);
                cp_async16(&s_rec[buf][slot][1], sp.rec1 + g);
                cp_async16(&s_rec[buf][slot][2], sp.rec2 + g);
            }
        }
        cp_async_commit();
    };

    if (n_batches > 0) {
        load_gids(0);
        issue_batch(0);
        if (n_batches > 1) load_gids(1);
    }
    for (int b = 0; b < n_batches; ++b) {
        const int buf = b & 1;
        if (b + 1 < n_batches) {
            issue_batch(buf ^ 1);
            if (b + 2 < n_batches) load_gids(b + 2);
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        const int cnt = min(BL_BATCH, range.y - range.x - b * BL_BATCH);
        // does the batch hold a splat whose conic is not safely positive definite (rec2.w, set by the preprocess
        // kernel)?  Only then can the shader's `power > 0 -> discard` fire; every thread looks at the records it
        // staged itself and the barrier ORs the answers, so the usual batch runs the loop without that test.
        bool flagged = false;
#pragma unroll
        for (int i = 0; i < PER_THREAD; ++i) {
            const int slot = threadIdx.x + i * THREADS;
            if (slot < cnt) flagged = flagged || (s_rec[buf][slot][2].w != 0.f);
        }
        const bool check_power = __syncthreads_or(flagged ? 1 : 0) != 0;
        bool live = false;
#pragma unroll
        for (int k = 0; k < PPT; ++k) live = live || (T[k] >= 1e-4f);
        auto run_batch = [&](auto tag) {
            constexpr bool CHECK_POWER = decltype(tag)::value;
            auto splat = [&](unsigned rec) {  // shared-window address of the record
                const float4 a = lds128(rec);  // cx, cy, hx, hy
                const float dx = pxf - a.x;
                const bool inx = fabsf(dx) <= a.z;
                float dy[PPT];
                bool in[PPT];
                bool any_in = false;
                if constexpr (MODE == 0 && !COUNT && PPT % 2 == 0) {
                    const f32x2 ay2 = pk2(a.y, a.y);
#pragma unroll
                    for (int k = 0; k < PPT; k += 2) {
                        const f32x2 d2 = sub2(pk2(pyf[k], pyf[k + 1]), ay2);
                        dy[k] = lo2(d2); dy[k + 1] = hi2(d2);
                    }
                }
#pragma unroll
                for (int k = 0; k < PPT; ++k) {
                    if constexpr (!(MODE == 0 && !COUNT && PPT % 2 == 0)) dy[k] = pyf[k] - a.y;
                    if (COUNT) n_eval += (T[k] >= 1e-4f) ? 1u : 0u;
                    // The shader path has no per-pixel stop (the ROP blends every fragment); a finished pixel
                    // (T < 1e-4) only matters for the warp-level exits below, so the gaussian mode leaves the
                    // transmittance out of the coverage predicate: one compare per pixel less, and w = alpha * T of a
                    // finished pixel is below 1e-4 anyway
                    constexpr bool T_TEST = !(MODE == 0 && !COUNT);
                    in[k] = inx && (fabsf(dy[k]) <= a.w) && (!T_TEST || T[k] >= 1e-4f);
                    any_in = any_in || in[k];
                }
                // no live pixel of this warp lies inside the quad: skip the splat for 3 instructions
                if (SKIP == 1 && !__any_sync(0xffffffffu, any_in)) return;
                const float4 q = lds128(rec + 16);  // A', B', C' (pre-scaled conic), opacity
                const float4 c = lds128(rec + 32);  // r, g, b, flag: conic not positive definite
                const float bdx = q.y * dx;
                const float adx2 = q.x * dx * dx;
                if constexpr (MODE == 0 && !COUNT && PPT % 2 == 0) {
                    // pixel pairs through the packed FMA pipe: 9 packed + 2 ex2 + 2 min + 4 compare/select per pair
                    // instead of 18 + 2 + 2 + 4 scalar instructions
                    const f32x2 bdx2 = pk2(bdx, bdx), adx22 = pk2(adx2, adx2), qz2 = pk2(q.z, q.z), qw2 = pk2(q.w, q.w);
                    const f32x2 cx2 = pk2(c.x, c.x), cy2 = pk2(c.y, c.y), cz2 = pk2(c.z, c.z);
                    // min(0.99, alpha) can only bite when opacity * 2^p2 can exceed 0.99: with p2 <= 0 guaranteed
                    // (!CHECK_POWER) that needs an opacity above 0.99.  The record is the same for every lane, so the
                    // branch does not diverge; two min per pixel pair less for the other splats.  (Choosing the variant
                    // per run of eight hits instead - a 46-instruction body - measured no faster than this 50.)
                    auto pairs = [&](auto clamp_tag) {
                        constexpr bool CLAMP = decltype(clamp_tag)::value;
#pragma unroll
                        for (int k = 0; k < PPT; k += 2) {
                            const f32x2 dy2 = pk2(dy[k], dy[k + 1]);
                            const f32x2 p2 = fma2(fma2(qz2, dy2, bdx2), dy2, adx22);  // log2(e) * power
                            const f32x2 ae = mul2(qw2, pk2(ex2_approx(lo2(p2)), ex2_approx(hi2(p2))));
                            float al0 = lo2(ae), al1 = hi2(ae);
                            if (CLAMP) { al0 = fminf(0.99f, al0); al1 = fminf(0.99f, al1); }
                            const f32x2 T2 = pk2(T[k], T[k + 1]);
                            const f32x2 aT = mul2(pk2(al0, al1), T2);
                            float w0, w1;
                            if (CHECK_POWER) {
                                w0 = (in[k] && !(al0 < 1.f / 255.f) && !(lo2(p2) > 0.f)) ? lo2(aT) : 0.f;
                                w1 = (in[k + 1] && !(al1 < 1.f / 255.f) && !(hi2(p2) > 0.f)) ? hi2(aT) : 0.f;
                            } else {
                                // one compare that ANDs in the coverage predicate, one select (the compiler otherwise
                                // selects twice): w = (in && !(alpha < 1/255)) ? alpha * T : 0
                                asm("{\n\t"
                                    ".reg .pred p, q;\n\t"
                                    "setp.ne.u32 q, %2, 0;\n\t"
                                    "setp.geu.and.f32 p, %1, 0f3B808081, q;\n\t"
                                    "selp.f32 %0, %3, 0f00000000, p;\n\t"
                                    "}"
                                    : "=f"(w0)
                                    : "f"(al0), "r"((unsigned)in[k]), "f"(lo2(aT)));
                                asm("{\n\t"
                                    ".reg .pred p, q;\n\t"
                                    "setp.ne.u32 q, %2, 0;\n\t"
                                    "setp.geu.and.f32 p, %1, 0f3B808081, q;\n\t"
                                    "selp.f32 %0, %3, 0f00000000, p;\n\t"
                                    "}"
                                    : "=f"(w1)
                                    : "f"(al1), "r"((unsigned)in[k + 1]), "f"(hi2(aT)));
                            }
                            const f32x2 w2 = pk2(w0, w1);
                            const f32x2 r2 = fma2(w2, cx2, pk2(cr[k], cr[k + 1]));
                            const f32x2 g2 = fma2(w2, cy2, pk2(cg[k], cg[k + 1]));
                            const f32x2 b2 = fma2(w2, cz2, pk2(cb[k], cb[k + 1]));
                            const f32x2 Tn = sub2(T2, w2);
                            cr[k] = lo2(r2); cr[k + 1] = hi2(r2);
                            cg[k] = lo2(g2); cg[k + 1] = hi2(g2);
                            cb[k] = lo2(b2); cb[k + 1] = hi2(b2);
                            T[k] = lo2(Tn); T[k + 1] = hi2(Tn);
                        }
                    };
                    if (CHECK_POWER || q.w > 0.98f) pairs(TagTrue{});
                    else pairs(TagFalse{});
                    return;
                }
#pragma unroll
                for (int k = 0; k < PPT; ++k) {
                    bool valid = in[k];
                    if (MODE == 2) {  // billboard: opaque quad
                        if (COUNT) n_hit += valid ? 1u : 0u;
                        const float w = valid ? T[k] : 0.f;
                        cr[k] = fmaf(w, c.x, cr[k]); cg[k] = fmaf(w, c.y, cg[k]); cb[k] = fmaf(w, c.z, cb[k]);
                        T[k] -= w;
                    } else {
                        const float p2 = fmaf(fmaf(q.z, dy[k], bdx), dy[k], adx2);  // log2(e) * power
                        const float e = ex2_approx(p2);
                        const float alpha = fminf(0.99f, q.w * e);
                        valid = valid && !(alpha < 1.f / 255.f);
                        if (CHECK_POWER) valid = valid && !(p2 > 0.f);
                        if (MODE == 5) {
                            const float test_T = T[k] * (1.f - alpha);
                            const bool stop = valid && (test_T < 0.0001f);
                            const bool hit = valid && !stop;
                            if (COUNT) n_hit += hit ? 1u : 0u;
                            const float w = hit ? alpha * T[k] : 0.f;
                            cr[k] = fmaf(w, c.x, cr[k]); cg[k] = fmaf(w, c.y, cg[k]); cb[k] = fmaf(w, c.z, cb[k]);
                            T[k] = stop ? 0.f : (hit ? test_T : T[k]);
                        } else if (MODE == 0) {
                            if (COUNT) n_hit += valid ? 1u : 0u;
                            float w;
                            if (!COUNT && !CHECK_POWER) {
                                // one compare that ANDs in the coverage predicate, one select (the compiler otherwise
                                // selects twice): w = (in && !(alpha < 1/255)) ? alpha * T : 0
                                asm("{\n\t"
                                    ".reg .pred p, q;\n\t"
                                    "setp.ne.u32 q, %2, 0;\n\t"
                                    "setp.geu.and.f32 p, %1, 0f3B808081, q;\n\t"
                                    "selp.f32 %0, %3, 0f00000000, p;\n\t"
                                    "}"
                                    : "=f"(w)
                                    : "f"(alpha), "r"((unsigned)in[k]), "f"(alpha * T[k]));
                            } else {
                                w = valid ? alpha * T[k] : 0.f;
                            }
                            cr[k] = fmaf(w, c.x, cr[k]); cg[k] = fmaf(w, c.y, cg[k]); cb[k] = fmaf(w, c.z, cb[k]);
                            T[k] -= w;
                        } else {  // flat / gaussian ball: alpha becomes 1 above 0.22, 0 below
                            const bool opaque = valid && (alpha > 0.22f);
                            if (COUNT) n_hit += valid ? 1u : 0u;
                            float r = c.x, g2 = c.y, bl = c.z;
                            if (MODE == 4) {
                                r *= e; g2 *= e; bl *= e;
                                if (clamp) { r = fminf(r, 1.f); g2 = fminf(g2, 1.f); bl = fminf(bl, 1.f); }
                            }
                            const float w = opaque ? T[k] : 0.f;
                            cr[k] = fmaf(w, r, cr[k]); cg[k] = fmaf(w, g2, cg[k]); cb[k] = fmaf(w, bl, cb[k]);
                            T[k] -= w;
                        }
                    }
                }
            };
            if (SKIP >= 2) {
                // Warp-level filter: each lane tests four records of the batch against the warp's pixel block
                // (conservatively: splat box vs block box, 0.01 px of slack against rounding), the ballots give the
                // hit list, and only hits run the per-pixel body.  A 16x16 tile's splat touches on average fewer
                // than half of the four 8x8 warp blocks; testing them one by one cost 16 instructions per miss.
                const int lane = threadIdx.x & 31;
                constexpr int NW = BL_BATCH / 32;
                // The hits go to a per-warp list of record offsets in shared memory, in list order: walking it costs
                // one LDS + pointer add per splat instead of find-first-set (BREV + FLO on the XU pipe, which the
                // four ex2 of the body already load to 50 %), clear-lowest-bit and an index multiply.
                uint16_t* hl = s_hit[threadIdx.x >> 5];
                const unsigned rec_base = (unsigned)__cvta_generic_to_shared(&s_rec[buf][0][0]);  // < 64 K: fits 16 bits
                const unsigned hl_base = (unsigned)__cvta_generic_to_shared(hl);
                int n_hit_w = 0;
#pragma unroll 4
                for (int i = 0; i < NW; ++i) {
                    const int j = i * 32 + lane;
                    const float4 a = s_rec[buf][j][0];
                    bool keep = (j < cnt) && (a.x + a.z >= bx0) && (a.x - a.z <= bx1) && (a.y + a.w >= by0) &&
                                (a.y - a.w <= by1);
                    if (SKIP == 3 && MODE != 2) {
                        // second test: can any pixel of the block reach alpha >= 1/255, i.e. is the maximum of the
                        // (concave) exponent p2 = A x^2 + B x y + C y^2 over the block's box of centre offsets at
                        // least -log2(255 opacity)?  The maximum over a box that does not hold the centre lies on an
                        // edge; each edge is a 1-D parabola with a clamped vertex.  Conservative by 0.05 in log2
                        // units; conics that are not positive definite with a margin (flag, or det < 1e-3 of its
                        // terms) are never culled.
                        const float4 q = s_rec[buf][j][1];
                        const float flag = s_rec[buf][j][2].w;
                        const float x0 = bx0 - a.x, x1 = bx1 - a.x, y0 = by0 - a.y, y1 = by1 - a.y;
                        const bool centre_in = (x0 <= 0.f) && (x1 >= 0.f) && (y0 <= 0.f) && (y1 >= 0.f);
                        const float ac4 = 4.f * q.x * q.z;
                        const bool pd = (flag == 0.f) && (q.x < 0.f) && (q.z < 0.f) && (ac4 - q.y * q.y > 1e-3f * ac4);
                        if (keep && pd && !centre_in) {
                            const float ry = __frcp_rn(-2.f * q.z), rx = __frcp_rn(-2.f * q.x);
                            auto edge_x = [&](float x) {  // max over y in [y0, y1] at fixed x
                                const float y = fminf(fmaxf(q.y * x * ry, y0), y1);
                                return fmaf(fmaf(q.z, y, q.y * x), y, q.x * x * x);
                            };
                            auto edge_y = [&](float y) {  // max over x in [x0, x1] at fixed y
                                const float x = fminf(fmaxf(q.y * y * rx, x0), x1);
                                return fmaf(fmaf(q.x, x, q.y * y), x, q.z * y * y);
                            };
                            const float pmax = fmaxf(fmaxf(edge_x(x0), edge_x(x1)), fmaxf(edge_y(y0), edge_y(y1)));
                            const float theta = __log2f(255.f * q.w);  // alpha >= 1/255  <=>  p2 >= -theta
                            keep = !(pmax < -theta - 0.05f);          // NaN anywhere keeps the splat
                        }
                    }
                    const uint32_t hit = __ballot_sync(0xffffffffu, keep);
                    if (keep) hl[n_hit_w + __popc(hit & ((1u << lane) - 1u))] = (uint16_t)(rec_base + j * 48);
                    n_hit_w += __popc(hit);
                }
                __syncwarp();
                for (int h0 = 0; h0 < n_hit_w; h0 += BL_RUN) {
                    // pixels die as the batch goes on: re-check every 8 hits
                    if (h0) {
                        bool alive = false;
#pragma unroll
                        for (int k = 0; k < PPT; ++k) alive = alive || (T[k] >= 1e-4f);
                        if (!__any_sync(0xffffffffu, alive)) break;
                    }
                    const int h1 = min(h0 + BL_RUN, n_hit_w);
                    // the next entry's offset is fetched while this splat is blended: the LDS.U16 -> LDS.128 pair at
                    // the head of every splat's dependency chain becomes one shared-memory latency (-1.7 % kernel time)
                    unsigned rec_next = lds16(hl_base + 2 * h0);
#pragma unroll 1
                    for (unsigned p = hl_base + 2 * h0, pe = hl_base + 2 * h1; p != pe; p += 2) {
                        const unsigned rec = rec_next;
                        rec_next = lds16(p + 2);
                        splat(rec);
                    }
                }
            } else {
#pragma unroll 2
                for (int j = 0; j < cnt; ++j) splat((unsigned)__cvta_generic_to_shared(&s_rec[buf][j][0]));
            }
        };
        if (__any_sync(0xffffffffu, live)) {
            if (check_power) run_batch(TagTrue{});
            else run_batch(TagFalse{});
        }
        bool live2 = false;
#pragma unroll
        for (int k = 0; k < PPT; ++k) live2 = live2 || (T[k] >= 1e-4f);
        // also the barrier that protects buffer `buf` from the cp.async of the next iteration
        if (__syncthreads_or(live2 ? 1 : 0) == 0) break;
    }
    cp_async_wait<0>();
#pragma unroll
    for (int k = 0; k < PPT; ++k) {
        if (!outside[k]) {
            const int row = (MODE == 5) ? (py0 + k) : (H - 1 - (py0 + k));
            float4* o = reinterpret_cast<float4*>(out_rgba) + (size_t)row * (size_t)W + px;
            *o = make_float4(cr[k], cg[k], cb[k], (MODE == 5) ? 1.f : 1.f - T[k]);
        }
    }
    // The capture loop wants the frame as RGB8 (main.py:852 glReadPixels): converted here with k_pack_rgb8's unorm8
    // rule instead of by a second kernel that re-reads the float frame.  A full tile goes through shared memory
    // (the record staging is free now) and leaves as aligned 32-bit words, 3 * TILE bytes per row; tiles cut by
    // the image edge, or a width that is not a multiple of four, fall back to byte stores.
    if (out_rgb8) {
        const int x0 = blockIdx.x * TILE, y0 = blockIdx.y * TILE;
        const bool full = ((W & 3) == 0) && (x0 + TILE <= W) && (y0 + TILE <= H);  // block-uniform
        unsigned rgb[PPT];  // r | g << 8 | b << 16
#pragma unroll
        for (int k = 0; k < PPT; ++k) {
            rgb[k] = (unsigned)__float2int_rn(fminf(fmaxf(cr[k], 0.f), 1.f) * 255.f) |
                     ((unsigned)__float2int_rn(fminf(fmaxf(cg[k], 0.f), 1.f) * 255.f) << 8) |
                     ((unsigned)__float2int_rn(fminf(fmaxf(cb[k], 0.f), 1.f) * 255.f) << 16);
        }
        if (full) {
            __syncthreads();  // every thread is past its last read of the staged records
            uint8_t* sb = reinterpret_cast<uint8_t*>(&s_rec[0][0][0]);  // [TILE rows (GL order)][3 * TILE]
#pragma unroll
            for (int k = 0; k < PPT; ++k) {
                uint8_t* d = sb + ((lyb * PPT + k) * TILE + lx) * 3;
                d[0] = (uint8_t)rgb[k]; d[1] = (uint8_t)(rgb[k] >> 8); d[2] = (uint8_t)(rgb[k] >> 16);
            }
            __syncthreads();
            constexpr int WPR = 3 * TILE / 4;  // words per tile row
            for (int w = threadIdx.x; w < TILE * WPR; w += THREADS) {
                const int ly = w / WPR, j = w - ly * WPR;
                const int row = (MODE == 5) ? (y0 + ly) : (H - 1 - (y0 + ly));
                uint32_t* dst = reinterpret_cast<uint32_t*>(out_rgb8 + 3 * ((size_t)row * (size_t)W + x0)) + j;
                *dst = reinterpret_cast<const uint32_t*>(sb)[ly * WPR + j];
            }
        } else {
#pragma unroll
            for (int k = 0; k < PPT; ++k) {
                if (!outside[k]) {
                    const int row = (MODE == 5) ? (py0 + k) : (H - 1 - (py0 + k));
                    uint8_t* d = out_rgb8 + 3 * ((size_t)row * (size_t)W + px);
                    d[0] = (uint8_t)rgb[k]; d[1] = (uint8_t)(rgb[k] >> 8); d[2] = (uint8_t)(rgb[k] >> 16);
                }
            }
        }
    }
    if (COUNT) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            n_eval += __shfl_xor_sync(0xffffffffu, n_eval, o);
            n_hit += __shfl_xor_sync(0xffffffffu, n_hit, o);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(&pair_counts[0], (unsigned long long)n_eval);
            atomicAdd(&pair_counts[1], (unsigned long long)n_hit);
        }
    }
}

__global__ void k_pack_rgb8(const float4* __restrict__ rgba, uint8_t* __restrict__ rgb, long long n_px) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_px) return;
    const float4 v = rgba[i];
    // unorm8 conversion of the GL spec: round(clamp(c, 0, 1) * 255)
    rgb[3 * i + 0] = (uint8_t)__float2int_rn(fminf(fmaxf(v.x, 0.f), 1.f) * 255.f);
    rgb[3 * i + 1] = (uint8_t)__float2int_rn(fminf(fmaxf(v.y, 0.f), 1.f) * 255.f);
    rgb[3 * i + 2] = (uint8_t)__float2int_rn(fminf(fmaxf(v.z, 0.f), 1.f) * 255.f);
}

__global__ void k_extract_r(const float4* __restrict__ rgba, float* __restrict__ r, long long n_px) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_px) r[i] = rgba[i].x;
}

// (disparity * 65535).astype(np.uint16) of main.py:878-879: float32 product, truncation toward zero, wrap
// modulo 2^16 like the x86 conversion NumPy's cast compiles to (out of the int32 range, or NaN -> 0)
__global__ void k_pack_disp_u16(const float4* __restrict__ rgba, uint16_t* __restrict__ d, long long n_px) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_px) return;
    const float v = __fmul_rn(rgba[i].x, 65535.f);
    int q = 0;
    if (fabsf(v) < 2147483648.f) q = __float2int_rz(v);
    d[i] = (uint16_t)(unsigned)q;
}

}  // namespace

// default configuration (every mode, and the counting variant)
constexpr int BL_PPT = 2;
constexpr bool BL_SHAPE8 = true;
constexpr int BL_SKIP = 3;

template <bool COUNT>
static void launch_blend_t(const FrameParams* fp, const int2* ranges, const uint32_t* inst_gid, SplatBuffers sp,
                           float* out_rgba, int render_mod, unsigned flags, int tiles_x, int tiles_y,
                           unsigned long long* pc, cudaStream_t s, int tile, uint8_t* out_rgb8) {
    const dim3 grid((unsigned)tiles_x, (unsigned)tiles_y);
    constexpr int TH = 256 / BL_PPT;
#define BL_ARGS fp, ranges, inst_gid, sp, out_rgba, flags, pc, out_rgb8
    if (tile == 32) {
        // 32x32 tiles (shader path only): a quarter of the tiles, less than half the instances to duplicate and
        // sort; the per-warp batch filter keeps the blend's work per pixel block what it was.  4 pixels per thread
        // (a warp owns 8x16 pixels, 256 threads per tile) measured 6 % faster than 2 here (0.254 -> 0.238 ms).
        // 64x64 tiles were tried too: duplicate + tile sort 0.18 -> 0.15 ms, blend 0.238 -> 0.330 ms - not kept.
        constexpr int SK32 = COUNT ? 0 : BL_SKIP;
        constexpr int P32 = 4, TH32 = 1024 / P32;
        if (render_mod == -2) launch_pdl_m(4, k_blend<2, P32, COUNT, true, SK32, 32>, dim3(grid), dim3(TH32), 0, s, BL_ARGS);
        else if (render_mod == -3) launch_pdl_m(4, k_blend<3, P32, COUNT, true, SK32, 32>, dim3(grid), dim3(TH32), 0, s, BL_ARGS);
        else if (render_mod == -4) launch_pdl_m(4, k_blend<4, P32, COUNT, true, SK32, 32>, dim3(grid), dim3(TH32), 0, s, BL_ARGS);
        else launch_pdl_m(4, k_blend<0, P32, COUNT, true, SK32, 32>, dim3(grid), dim3(TH32), 0, s, BL_ARGS);
        return;
    }
    // the counting variant visits every splat of a list (no skipping), so that "pairs visited" is the algorithmic
    // figure of SURVEY 8(d): per pixel, the list position at which it terminates
    constexpr int SK = COUNT ? 0 : BL_SKIP;
    if (flags & GSR_FLAG_OFFICIAL_BIT) launch_pdl_m(4, k_blend<5, BL_PPT, COUNT, BL_SHAPE8, 0>, dim3(grid), dim3(TH), 0, s, BL_ARGS);
    else if (render_mod == -2) launch_pdl_m(4, k_blend<2, BL_PPT, COUNT, BL_SHAPE8, SK>, dim3(grid), dim3(TH), 0, s, BL_ARGS);
    else if (render_mod == -3) launch_pdl_m(4, k_blend<3, BL_PPT, COUNT, BL_SHAPE8, SK>, dim3(grid), dim3(TH), 0, s, BL_ARGS);
    else if (render_mod == -4) launch_pdl_m(4, k_blend<4, BL_PPT, COUNT, BL_SHAPE8, SK>, dim3(grid), dim3(TH), 0, s, BL_ARGS);
    else launch_pdl_m(4, k_blend<0, BL_PPT, COUNT, BL_SHAPE8, SK>, dim3(grid), dim3(TH), 0, s, BL_ARGS);
    // (the pixels-per-thread / block-shape / skip-level variants measured in round 1 are gone: DESIGN.md section 5 keeps
    // their numbers)
#undef BL_ARGS
}

void launch_blend(const FrameParams* fp, const int2* ranges, const uint32_t* inst_gid, SplatBuffers sp, float* out_rgba,
                  int render_mod, unsigned flags, int tiles_x, int tiles_y, cudaStream_t s, int tile, uint8_t* out_rgb8) {
    launch_blend_t<false>(fp, ranges, inst_gid, sp, out_rgba, render_mod, flags, tiles_x, tiles_y, nullptr, s, tile, out_rgb8);
}

// Instrumented variant: same image, plus pair_counts[0] += pairs visited, [1] += pairs that contributed.
void launch_blend_count(const FrameParams* fp, const int2* ranges, const uint32_t* inst_gid, SplatBuffers sp,
                        float* out_rgba, int render_mod, unsigned flags, int tiles_x, int tiles_y,
                        unsigned long long* pair_counts, cudaStream_t s, int tile) {
    launch_blend_t<true>(fp, ranges, inst_gid, sp, out_rgba, render_mod, flags, tiles_x, tiles_y, pair_counts, s, tile, nullptr);
}

void launch_pack_rgb8(const float* rgba, uint8_t* rgb, long long n_px, cudaStream_t s) {
    k_pack_rgb8<<<(unsigned)((n_px + 255) / 256), 256, 0, s>>>(reinterpret_cast<const float4*>(rgba), rgb, n_px);
}
void launch_pack_disp_u16(const float* rgba, uint16_t* d, long long n_px, cudaStream_t s) {
    k_pack_disp_u16<<<(unsigned)((n_px + 255) / 256), 256, 0, s>>>(reinterpret_cast<const float4*>(rgba), d, n_px);
}
void launch_extract_r(const float* rgba, float* r, long long n_px, cudaStream_t s) {
    k_extract_r<<<(unsigned)((n_px + 255) / 256), 256, 0, s>>>(reinterpret_cast<const float4*>(rgba), r, n_px);
}
