// K5 — per-16x16-tile front-to-back alpha blend for sm_100a.
//
// One CTA per tile, one thread per pixel.  The tile's instance list (nearest first) is staged through
// shared memory in batches of 256 splat records; every thread evaluates the fragment rule of
// /root/reference/shaders/gau_frag.glsl:13-39 for its pixel and composites front to back, which equals
// the reference's back-to-front SRC_ALPHA / ONE_MINUS_SRC_ALPHA blend over a black clear
// (renderer_ogl.py:128-130, main.py:821):   C = sum_i c_i a_i T_i,  T_i = prod_{j<i} (1 - a_j).
// A pixel stops once T < 1e-4 (its remaining contribution is below 1e-4 of any colour); a warp skips a
// batch when all its pixels have stopped (ballot) and the CTA leaves when every pixel has.
// Modes: >= 0 SH colour, -1 disparity, -2 billboard, -3 flat ball, -4 gaussian ball.
#include "gsr_internal.h"

namespace {

constexpr int BL_THREADS = GSR_TILE * GSR_TILE;

// MODE: 0 = gaussian falloff (render_mod >= 0 or -1), 2 = billboard (-2), 3 = flat ball (-3), 4 = gaussian ball (-4)
template <int MODE, bool COUNT>
__global__ void __launch_bounds__(BL_THREADS) k_blend(const FrameParams* __restrict__ fp, const int2* __restrict__ ranges,
                                                     const uint32_t* __restrict__ inst_gid, SplatBuffers sp,
                                                     float* __restrict__ out_rgba, unsigned flags,
                                                     unsigned long long* __restrict__ pair_counts) {
    __shared__ float4 s0[BL_THREADS];
    __shared__ float4 s1[BL_THREADS];
    __shared__ float4 s2[BL_THREADS];
    const int W = fp->W, H = fp->H;
    const int tile = blockIdx.y * gridDim.x + blockIdx.x;
    const int lx = threadIdx.x & (GSR_TILE - 1), ly = threadIdx.x / GSR_TILE;
    const int px = blockIdx.x * GSR_TILE + lx;
    const int py = blockIdx.y * GSR_TILE + ly;  // GL row, 0 = bottom
    const bool inside = (px < W) && (py < H);
    const float pxf = (float)px + 0.5f, pyf = (float)py + 0.5f;
    const int2 range = ranges[tile];
    bool done = !inside;
    float T = 1.f, cr = 0.f, cg = 0.f, cb = 0.f;
    const bool clamp = (flags & 1u) != 0u;
    unsigned n_eval = 0, n_hit = 0;  // COUNT only: (pixel, splat) pairs visited / pairs that contributed

    for (int start = range.x; start < range.y; start += BL_THREADS) {
        if (__syncthreads_count(done) == BL_THREADS) break;
        const int idx = start + threadIdx.x;
        if (idx < range.y) {
            const uint32_t g = __ldg(inst_gid + idx);
            s0[threadIdx.x] = __ldg(sp.rec0 + g);
            s1[threadIdx.x] = __ldg(sp.rec1 + g);
            s2[threadIdx.x] = __ldg(sp.rec2 + g);
        }
        __syncthreads();
        const int cnt = min(BL_THREADS, range.y - start);
        if (__all_sync(0xffffffffu, done)) continue;  // whole warp finished: skip the batch
        for (int j = 0; j < cnt && !done; ++j) {
            const float4 a = s0[j];
            if (COUNT) ++n_eval;
            const uint32_t rx = __float_as_uint(a.z), ry = __float_as_uint(a.w);
            // pixel-rect test == |dx| <= hx && |dy| <= hy of the quad (exact, see preprocess.cu)
            if ((uint32_t)(px - (int)(rx & 0xFFFFu)) > (rx >> 16)) continue;
            if ((uint32_t)(py - (int)(ry & 0xFFFFu)) > (ry >> 16)) continue;
            const float4 c = s2[j];
            if (MODE == 2) {  // billboard: opaque quad
                if (COUNT) ++n_hit;
                cr += T * c.x; cg += T * c.y; cb += T * c.z;
                T = 0.f; done = true;
                break;
            }
            const float4 b = s1[j];
            const float dx = pxf - a.x, dy = pyf - a.y;
            const float power = -0.5f * (b.x * dx * dx + b.z * dy * dy) - b.y * dx * dy;
            if (power > 0.f) continue;
            const float e = __expf(power);
            const float alpha = fminf(0.99f, b.w * e);
            if (alpha < 1.f / 255.f) continue;
            if (COUNT) ++n_hit;
            if (MODE == 0) {
                const float w = alpha * T;
                cr += w * c.x; cg += w * c.y; cb += w * c.z;
                T = T * (1.f - alpha);
                if (T < 1e-4f) done = true;
            } else {
                if (alpha > 0.22f) {  // flat / gaussian ball: alpha becomes 1
                    float r = c.x, g2 = c.y, bl = c.z;
                    if (MODE == 4) {
                        r *= e; g2 *= e; bl *= e;
                        if (clamp) { r = fminf(r, 1.f); g2 = fminf(g2, 1.f); bl = fminf(bl, 1.f); }
                    }
                    cr += T * r; cg += T * g2; cb += T * bl;
                    T = 0.f; done = true;
                }
            }
        }
    }
    if (inside) {
        float4* o = reinterpret_cast<float4*>(out_rgba) + (size_t)(H - 1 - py) * (size_t)W + px;
        *o = make_float4(cr, cg, cb, 1.f - T);
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

}  // namespace

template <bool COUNT>
static void launch_blend_t(const FrameParams* fp, const int2* ranges, const uint32_t* inst_gid, SplatBuffers sp,
                           float* out_rgba, int render_mod, unsigned flags, int tiles_x, int tiles_y,
                           unsigned long long* pc, cudaStream_t s) {
    const dim3 grid((unsigned)tiles_x, (unsigned)tiles_y);
    if (render_mod == -2) k_blend<2, COUNT><<<grid, BL_THREADS, 0, s>>>(fp, ranges, inst_gid, sp, out_rgba, flags, pc);
    else if (render_mod == -3) k_blend<3, COUNT><<<grid, BL_THREADS, 0, s>>>(fp, ranges, inst_gid, sp, out_rgba, flags, pc);
    else if (render_mod == -4) k_blend<4, COUNT><<<grid, BL_THREADS, 0, s>>>(fp, ranges, inst_gid, sp, out_rgba, flags, pc);
    else k_blend<0, COUNT><<<grid, BL_THREADS, 0, s>>>(fp, ranges, inst_gid, sp, out_rgba, flags, pc);
}

void launch_blend(const FrameParams* fp, const int2* ranges, const uint32_t* inst_gid, SplatBuffers sp, float* out_rgba,
                  int render_mod, unsigned flags, int tiles_x, int tiles_y, cudaStream_t s) {
    launch_blend_t<false>(fp, ranges, inst_gid, sp, out_rgba, render_mod, flags, tiles_x, tiles_y, nullptr, s);
}

// Instrumented variant: same image, plus pair_counts[0] += pairs visited, [1] += pairs that contributed.
void launch_blend_count(const FrameParams* fp, const int2* ranges, const uint32_t* inst_gid, SplatBuffers sp,
                        float* out_rgba, int render_mod, unsigned flags, int tiles_x, int tiles_y,
                        unsigned long long* pair_counts, cudaStream_t s) {
    launch_blend_t<true>(fp, ranges, inst_gid, sp, out_rgba, render_mod, flags, tiles_x, tiles_y, pair_counts, s);
}

void launch_pack_rgb8(const float* rgba, uint8_t* rgb, long long n_px, cudaStream_t s) {
    k_pack_rgb8<<<(unsigned)((n_px + 255) / 256), 256, 0, s>>>(reinterpret_cast<const float4*>(rgba), rgb, n_px);
}
void launch_extract_r(const float* rgba, float* r, long long n_px, cudaStream_t s) {
    k_extract_r<<<(unsigned)((n_px + 255) / 256), 256, 0, s>>>(reinterpret_cast<const float4*>(rgba), r, n_px);
}
