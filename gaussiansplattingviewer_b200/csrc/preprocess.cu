// K1 — per-Gaussian preprocess for sm_100a.
//
// One thread per Gaussian: frustum cull, 3D covariance from scale/rotation, EWA projection to a 2D
// conic, 3-sigma axis-aligned footprint, exact pixel/tile rectangle, SH degree 0..3 or disparity
// colour.  Restates /root/reference/shaders/gau_vert.glsl:73-251 (per instance, once instead of once
// per quad vertex) plus the raster set-up rules of oracle/ref_driver.inc.
//
// THIS FILE IS COMPILED WITH -fmad=false: every operation is a separately rounded IEEE binary32
// operation in the order the shader writes it (conventions in oracle/glsl_shim.hpp), so the splat
// records are bit-identical to the CPU oracle's.  The kernel is HBM-bound (12 B per culled Gaussian,
// ~236 B in / 56 B out per visible one), so the unfused arithmetic costs nothing.
#include "gsr_internal.h"

namespace {

#define SH_C0 0.28209479177387814f
#define SH_C1 0.4886025119029199f
#define SH_C2_0 1.0925484305920792f
#define SH_C2_1 -1.0925484305920792f
#define SH_C2_2 0.31539156525252005f
#define SH_C2_3 -1.0925484305920792f
#define SH_C2_4 0.5462742152960396f
#define SH_C3_0 -0.5900435899266435f
#define SH_C3_1 2.890611442640554f
#define SH_C3_2 -0.4570457994644658f
#define SH_C3_3 0.3731763325901154f
#define SH_C3_4 -0.4570457994644658f
#define SH_C3_5 1.445305721320277f
#define SH_C3_6 -0.5900435899266435f

__device__ __forceinline__ float gl_min(float x, float y) { return (y < x) ? y : x; }
__device__ __forceinline__ float gl_max(float x, float y) { return (x < y) ? y : x; }

struct m3 { float m[3][3]; };  // column-major like GLSL: m[col][row]

__device__ __forceinline__ m3 m3_mul(const m3& A, const m3& B) {
    m3 R;
#pragma unroll
    for (int col = 0; col < 3; ++col)
#pragma unroll
        for (int row = 0; row < 3; ++row)
            R.m[col][row] = (A.m[0][row] * B.m[col][0] + A.m[1][row] * B.m[col][1]) + A.m[2][row] * B.m[col][2];
    return R;
}
__device__ __forceinline__ m3 m3_transpose(const m3& A) {
    m3 R;
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int r = 0; r < 3; ++r) R.m[c][r] = A.m[r][c];
    return R;
}
__device__ __forceinline__ void mat4_mul_vec4(const float* __restrict__ M, const float v[4], float out[4]) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
        out[r] = ((M[r * 4 + 0] * v[0] + M[r * 4 + 1] * v[1]) + M[r * 4 + 2] * v[2]) + M[r * 4 + 3] * v[3];
}

// pixel coverage interval, identical to cover_lo / cover_hi in oracle/gs_oracle.c
__device__ __forceinline__ int cover_lo(float c, float h, int n) {
    const float f = (c - h) - 0.5f;
    int i;
    if (!(f > 0.f)) i = 0;
    else if (f >= (float)n) i = n;
    else i = (int)floorf(f);
    while (i > 0 && (((float)(i - 1) + 0.5f) - c) >= -h) --i;
    while (i < n && (((float)i + 0.5f) - c) < -h) ++i;
    return i;
}
__device__ __forceinline__ int cover_hi(float c, float h, int n) {
    const float g = (c + h) - 0.5f;
    int i;
    if (!(g >= 0.f)) i = -1;
    else if (g >= (float)(n - 1)) i = n - 1;
    else i = (int)floorf(g);
    while (i < n - 1 && (((float)(i + 1) + 0.5f) - c) <= h) ++i;
    while (i >= 0 && (((float)i + 0.5f) - c) > h) --i;
    return i;
}

__device__ __forceinline__ bool finitef(float x) { return fabsf(x) <= 3.402823466e+38f; }

// Block-uniform copy of the frame uniforms into shared memory.
struct SharedParams {
    float view[16];
    float proj[16];
    float htanx, htany, focal;
    float campos[3];
    float scale_modifier;
    int render_mod, W, H, tiles_x, tiles_y;
    unsigned flags;
    int tile_shift;
    unsigned depth_k0, depth_khi;
    int depth_npass, depth_shift[4], depth_mask[4];
};

template <int SH_DIM, bool DEBUG>
__global__ void __launch_bounds__(256, 3) k_preprocess(const FrameParams* __restrict__ fp, SceneBuffers scene,
                                                    SplatBuffers out, FrameCounters* __restrict__ counters,
                                                    unsigned n, int sh_dim_rt) {
    // Persistent CTAs, one chunk of 256 Gaussians per iteration (a warp owns 32 consecutive rows), every load a
    // coalesced cp.async issued well before its use, so that the fetches run under the covariance maths:
    //   s_sh  (per warp 32 rows x 13 float4, SH_DIM == 48)  the SH rows of the warp's visible Gaussians; a lane reads
    //         its own row with LDS.128 (rows padded from 192 to 208 bytes: conflict-free).  A
    //         lane-per-row LDG.128 at a 192-byte stride costs 32 L1 wavefronts per instruction and 48 registers
    //   s_in  (per warp 24 + 32 + 32 float4)  xyz / rot / scale+opacity rows of the warp's NEXT chunk
    extern __shared__ float4 s_dyn[];
    __shared__ SharedParams P;
    __shared__ unsigned s_visible;
    // digit counts of the depth keys this CTA writes (the four passes of the depth sort need them): the CTAs are
    // persistent, so the counts stay in shared memory for the whole kernel and go out once per CTA - the depth
    // sort's histogram kernel (32 us over 3 M keys, bound by shared-memory atomics) is not launched
    __shared__ uint32_t s_dh[3 * 512];  // up to three 9-bit digits of the rebased depth key
    for (int i = threadIdx.x; i < 3 * 512; i += 256) s_dh[i] = 0;
    pdl_wait();  // resident behind k_set_params: uniforms and cleared scratch are read after this
    unsigned n_keys = 0, n_clamped = 0;
    if (threadIdx.x < 16) { P.view[threadIdx.x] = fp->view[threadIdx.x]; P.proj[threadIdx.x] = fp->proj[threadIdx.x]; }
    if (threadIdx.x == 32) {
        P.htanx = fp->htanx; P.htany = fp->htany; P.focal = fp->focal;
        P.campos[0] = fp->campos[0]; P.campos[1] = fp->campos[1]; P.campos[2] = fp->campos[2];
        P.scale_modifier = fp->scale_modifier; P.render_mod = fp->render_mod;
        P.W = fp->W; P.H = fp->H; P.tiles_x = fp->tiles_x; P.tiles_y = fp->tiles_y; P.flags = fp->flags; P.tile_shift = fp->tile_shift;
        P.depth_k0 = fp->depth_k0; P.depth_khi = fp->depth_khi;
        P.depth_npass = fp->depth_npass < 3 ? fp->depth_npass : 3;
        int sh = 0;
        for (int p = 0; p < 4; ++p) {
            const int b = (p < fp->depth_npass) ? fp->depth_bits[p] : 0;
            P.depth_shift[p] = sh; P.depth_mask[p] = (1 << b) - 1;
            sh += b;
        }
        s_visible = 0;
    }
    __syncthreads();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    float4* const s_row = s_dyn + warp * (32 * 13 + 88);  // SH rows of this warp
    float4* const s_in = s_row + 32 * 13;                   // xyz (24 float4), rot (32), scale+opacity (32)
    const unsigned n_chunks = (n + 255u) >> 8;
    // rows [first, first + 32) of the warp -> s_in (lanes past the end copy nothing; a partial xyz chunk is zero-filled)
    auto prefetch = [&](unsigned chunk) {
        if (chunk >= n_chunks) return;
        const unsigned first = chunk * 256u + warp * 32u;
        if (lane < 24u) {
            const long long left = (long long)n * 12 - ((long long)first * 12 + lane * 16);
            if (left > 0) {
                const unsigned dst = (unsigned)__cvta_generic_to_shared(s_in + lane);
                const int bytes = left < 16 ? (int)left : 16;
                asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(dst),
                             "l"(reinterpret_cast<const char*>(scene.xyz) + (size_t)first * 12 + lane * 16), "r"(bytes) : "memory");
            }
        }
        if (first + lane < n) {
            const unsigned d0 = (unsigned)__cvta_generic_to_shared(s_in + 24 + lane);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d0), "l"(scene.rot + first + lane) : "memory");
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d0 + 32 * 16), "l"(scene.scale_op + first + lane) : "memory");
        }
    };
    prefetch(blockIdx.x);
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncwarp();
    unsigned n_ok = 0;
    for (unsigned chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const unsigned g = chunk * 256u + threadIdx.x;
        const bool in = g < n;
        const unsigned gl = in ? g : n - 1;
        bool ok = false;
        const int sh_dim = (SH_DIM > 0) ? SH_DIM : sh_dim_rt;
        const int mod = P.render_mod;
        int ncoef = 1;
        if (sh_dim > 3 && mod >= 1) ncoef = 4;
        if (sh_dim > 12 && mod >= 2) ncoef = 9;
        if (sh_dim > 27 && mod >= 3) ncoef = 16;
        // ---- this chunk's rows out of shared memory, then the next chunk's rows on their way into it
        const float* const s_xyz = reinterpret_cast<const float*>(s_in);
        const float px = s_xyz[3 * lane + 0], py = s_xyz[3 * lane + 1], pz = s_xyz[3 * lane + 2];
        const float4 q = s_in[24 + lane];
        const float4 so = s_in[56 + lane];
        __syncwarp();
        prefetch(chunk + gridDim.x);
        // ---- gau_vert.glsl:134-146: project + early cull
        const float g_pos[4] = {px, py, pz, 1.f};
        float pv[4], pc[4];
        mat4_mul_vec4(P.view, g_pos, pv);
        mat4_mul_vec4(P.proj, pv, pc);
        const float ndc0 = pc[0] / pc[3], ndc1 = pc[1] / pc[3], ndc2 = pc[2] / pc[3];
        const float lim = (float)1.3;
        const bool culled = !in || (fabsf(ndc0) > lim || fabsf(ndc1) > lim || fabsf(ndc2) > lim);

        // ---- SH staging, round 1: rows whose centre is on screen (visible for certain unless a value is not
        // finite) are requested now, so that the fetch runs under the covariance maths; the border rows
        // (1 < |ndc| <= 1.3) wait until `ok` is known (round 2) - most of them end up invisible
        // 4 lanes per row, 8 rows per step: lane l copies float4 (l & 3) + 4k (k = 0..2) of row 8t + (l >> 2), so
        // every address is one base plus an immediate and every 32-byte sector that is fetched is used in full
        const float4* const lane_src = reinterpret_cast<const float4*>(scene.sh) +
                                       ((size_t)(g - lane) + (lane >> 2)) * 12 + (lane & 3u);
        const unsigned lane_dst = (unsigned)__cvta_generic_to_shared(s_row + (lane >> 2) * 13 + (lane & 3u));
        const int nv = (ncoef * 3 + 3) >> 2;  // float4s needed: 1, 3, 7, 12
        auto stage = [&](unsigned mask) {
            const unsigned mrow = mask >> (lane >> 2);
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                if ((mrow >> (8 * t)) & 1u) {
#pragma unroll
                    for (int k = 0; k < 3; ++k)
                        if ((int)(lane & 3u) + 4 * k < nv)
                            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(lane_dst + (t * 8 * 13 + 4 * k) * 16),
                                         "l"(lane_src + t * 8 * 12 + 4 * k) : "memory");
                }
            }
        };
        const bool use_stage = (SH_DIM == 48) && (mod != -1);
        bool staged = false;
        if (use_stage) {
            staged = !culled && (DEBUG || (fabsf(ndc0) <= 1.f && fabsf(ndc1) <= 1.f && ndc2 >= -1.f && ndc2 <= 1.f));
            const unsigned m1 = __ballot_sync(0xffffffffu, staged);
            if (m1) stage(m1);
        }

        float vs[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) vs[k] = 0.f;
        int imin = 0, imax = -1, jmin = 0, jmax = -1;
        float cx = 0.f, cy = 0.f, depth = 0.f, hx = 0.f, hy = 0.f;
        float conic0 = 0.f, conic1 = 0.f, conic2 = 0.f, alpha = 0.f, c0 = 0.f, c1 = 0.f, c2 = 0.f;

        if (!culled) {
            depth = -pv[2];
            const float ndcx = ndc0 + 0.f, ndcy = ndc1 + 0.f;
            // ---- :151-156 scale, computeCov3D :73-93
            float s[3] = {so.x * P.scale_modifier, so.y * P.scale_modifier, so.z * P.scale_modifier};
            if (mod == -1) { s[0] = s[0] * (float)1.2; s[1] = s[1] * (float)1.2; s[2] = s[2] * (float)1.2; }
            m3 S;
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
                for (int r = 0; r < 3; ++r) S.m[c][r] = 0.f;
            S.m[0][0] = s[0]; S.m[1][1] = s[1]; S.m[2][2] = s[2];
            const float r = q.x, x = q.y, y = q.z, z = q.w;
            m3 R;
            R.m[0][0] = 1.f - 2.f * (y * y + z * z); R.m[0][1] = 2.f * (x * y - r * z); R.m[0][2] = 2.f * (x * z + r * y);
            R.m[1][0] = 2.f * (x * y + r * z); R.m[1][1] = 1.f - 2.f * (x * x + z * z); R.m[1][2] = 2.f * (y * z - r * x);
            R.m[2][0] = 2.f * (x * z - r * y); R.m[2][1] = 2.f * (y * z + r * x); R.m[2][2] = 1.f - 2.f * (x * x + y * y);
            const m3 M = m3_mul(S, R);
            const m3 cov3d = m3_mul(m3_transpose(M), M);
            // ---- computeCov2D :95-120
            float tx = pv[0], ty = pv[1];
            const float tz = pv[2];
            const float limx = 1.3f * P.htanx, limy = 1.3f * P.htany;
            const float txtz = tx / tz, tytz = ty / tz;
            tx = gl_min(limx, gl_max(-limx, txtz)) * tz;
            ty = gl_min(limy, gl_max(-limy, tytz)) * tz;
            const float focal = P.focal;
            m3 J;
            J.m[0][0] = focal / tz; J.m[0][1] = 0.0f; J.m[0][2] = -(focal * tx) / (tz * tz);
            J.m[1][0] = 0.0f; J.m[1][1] = focal / tz; J.m[1][2] = -(focal * ty) / (tz * tz);
            J.m[2][0] = 0.f; J.m[2][1] = 0.f; J.m[2][2] = 0.f;
            m3 Wm;
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
                for (int rr = 0; rr < 3; ++rr) Wm.m[c][rr] = P.view[c * 4 + rr];
            const m3 T = m3_mul(Wm, J);
            const m3 A = m3_mul(m3_transpose(T), m3_transpose(cov3d));
            m3 cov = m3_mul(A, T);
            cov.m[0][0] += 0.3f;
            cov.m[1][1] += 0.3f;
            const float ca = cov.m[0][0], cb = cov.m[0][1], cc = cov.m[1][1];
            // ---- :167-178 conic + footprint
            const float det = (ca * cc - cb * cb);
            const float det_inv = 1.f / det;
            conic0 = cc * det_inv; conic1 = -cb * det_inv; conic2 = ca * det_inv;
            hx = 3.f * sqrtf(ca); hy = 3.f * sqrtf(cc);
            alpha = so.w;
            // ---- colour
            if (mod == -1) {  // :182-211 disparity
                const float baseline = -0.5f;
                const float xl = (ndc0 + 1.f) / 2.f;
                const float pr[4] = {px + baseline, py + 0.f, pz + 0.f, 1.f};
                float pvr[4], pcr[4];
                mat4_mul_vec4(P.view, pr, pvr);
                mat4_mul_vec4(P.proj, pvr, pcr);
                const float xr = ((pcr[0] / pcr[3]) + 1.f) / 2.f;
                const float d = fabsf(xl - xr);
                c0 = d; c1 = d; c2 = d;
            }
            // ---- raster set-up (oracle/gs_oracle.c setup_one): depth clip, viewport, pixel rect
            if (ndc2 >= -1.f && ndc2 <= 1.f) {
                cx = ((ndcx + 1.f) * 0.5f) * (float)P.W;
                cy = ((ndcy + 1.f) * 0.5f) * (float)P.H;
                if (finitef(cx) && finitef(cy) && finitef(hx) && finitef(hy)) {
                    imin = cover_lo(cx, hx, P.W); imax = cover_hi(cx, hx, P.W);
                    jmin = cover_lo(cy, hy, P.H); jmax = cover_hi(cy, hy, P.H);
                    ok = (imin <= imax && jmin <= jmax);
                }
            }
        }
        // ---- SH staging, round 2 (border rows that turned out visible), then the colour of the visible rows only
        if (use_stage) {
            const bool late = ok && !staged;
            const unsigned m2 = __ballot_sync(0xffffffffu, late);
            if (m2) stage(m2);
            asm volatile("cp.async.wait_all;" ::: "memory");
            __syncwarp();
        }
        if (!culled && mod != -1 && (ok || DEBUG)) {  // :214-250 SH
            float dx = px - P.campos[0], dy = py - P.campos[1], dz = pz - P.campos[2];
            const float len = sqrtf((dx * dx + dy * dy) + dz * dz);
            dx = dx / len; dy = dy / len; dz = dz / len;
            const float* __restrict__ shp = scene.sh + (size_t)sh_dim * (size_t)gl;
            float sh[48];
            if (SH_DIM == 48) {
                const float4* my = s_row + lane * 13;
#pragma unroll
                for (int k = 0; k < 12; ++k) {
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (k < nv) v = my[k];
                    sh[4 * k + 0] = v.x; sh[4 * k + 1] = v.y; sh[4 * k + 2] = v.z; sh[4 * k + 3] = v.w;
                }
            } else {
#pragma unroll
                for (int k = 0; k < 48; ++k) sh[k] = (k < ncoef * 3 && k < sh_dim) ? __ldg(shp + k) : 0.f;
            }
#define SHV(k, ch) sh[(k) * 3 + (ch)]
            float c[3];
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) c[ch] = SH_C0 * SHV(0, ch);
            if (ncoef >= 4) {
                const float X = dx, Y = dy, Z = dz;
#pragma unroll
                for (int ch = 0; ch < 3; ++ch)
                    c[ch] = c[ch] - SH_C1 * Y * SHV(1, ch) + SH_C1 * Z * SHV(2, ch) - SH_C1 * X * SHV(3, ch);
                if (ncoef >= 9) {
                    const float xx = X * X, yy = Y * Y, zz = Z * Z;
                    const float xy = X * Y, yz = Y * Z, xz = X * Z;
#pragma unroll
                    for (int ch = 0; ch < 3; ++ch)
                        c[ch] = c[ch] + SH_C2_0 * xy * SHV(4, ch) + SH_C2_1 * yz * SHV(5, ch) +
                                SH_C2_2 * (2.0f * zz - xx - yy) * SHV(6, ch) + SH_C2_3 * xz * SHV(7, ch) +
                                SH_C2_4 * (xx - yy) * SHV(8, ch);
                    if (ncoef >= 16) {
#pragma unroll
                        for (int ch = 0; ch < 3; ++ch)
                            c[ch] = c[ch] + SH_C3_0 * Y * (3.0f * xx - yy) * SHV(9, ch) + SH_C3_1 * xy * Z * SHV(10, ch) +
                                    SH_C3_2 * Y * (4.0f * zz - xx - yy) * SHV(11, ch) +
                                    SH_C3_3 * Z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * SHV(12, ch) +
                                    SH_C3_4 * X * (4.0f * zz - xx - yy) * SHV(13, ch) +
                                    SH_C3_5 * Z * (xx - yy) * SHV(14, ch) + SH_C3_6 * X * (xx - 3.0f * yy) * SHV(15, ch);
                    }
                }
            }
#undef SHV
            c0 = c[0] + 0.5f; c1 = c[1] + 0.5f; c2 = c[2] + 0.5f;
        }
        if (DEBUG && !culled) {
            vs[0] = 1.f; vs[1] = ndc0 + 0.f; vs[2] = ndc1 + 0.f; vs[3] = ndc2; vs[4] = hx; vs[5] = hy;
            vs[6] = conic0; vs[7] = conic1; vs[8] = conic2; vs[9] = c0; vs[10] = c1; vs[11] = c2;
            vs[12] = alpha; vs[13] = depth;
        }
        // ---- outputs
        uint2 trect = make_uint2(0u, 0u);  // tx0 | tiles wide << 16, ty0 | tiles high << 16 (0 x 0 when not ok)
        if (ok) {
            const int ts = P.tile_shift;
            const int tx0 = imin >> ts, tx1 = imax >> ts, ty0 = jmin >> ts, ty1 = jmax >> ts;
            trect.x = (unsigned)tx0 | ((unsigned)(tx1 - tx0 + 1) << 16);
            trect.y = (unsigned)ty0 | ((unsigned)(ty1 - ty0 + 1) << 16);
            // per-tile instance counts without one atomic per tile: the rectangle is the 2D prefix sum of four
            // corner impulses (+1, -1, -1, +1); k_tile_scan integrates the grid into counts, list offsets
            // and the digit histograms of the tile sort
            if (out.tile_grid) {
                const int gw = out.grid_w;
                int* grid = out.tile_grid + (size_t)(blockIdx.x & (GSR_GRID_REPLICAS - 1)) * out.grid_stride;
                atomicAdd(grid + ty0 * gw + tx0, 1);
                atomicAdd(grid + ty0 * gw + (tx1 + 1), -1);
                atomicAdd(grid + (ty1 + 1) * gw + tx0, -1);
                atomicAdd(grid + (ty1 + 1) * gw + (tx1 + 1), 1);
            }
            // source-colour clamp of a normalised target (ref_driver.inc); mode -4 scales the colour per
            // fragment first, so only the lower clamp can be applied here
            if (P.flags & 1u) {
                if (mod == -4) { c0 = gl_max(c0, 0.f); c1 = gl_max(c1, 0.f); c2 = gl_max(c2, 0.f); }
                else {
                    c0 = gl_min(gl_max(c0, 0.f), 1.f); c1 = gl_min(gl_max(c1, 0.f), 1.f); c2 = gl_min(gl_max(c2, 0.f), 1.f);
                }
            }
            // blend-ready record: centre + half extents (the quad test |dx| <= hx, |dy| <= hy is evaluated per
            // pixel exactly as the oracle does), conic pre-scaled so that the blend kernel gets
            // log2(e) * power = A' dx^2 + B' dx dy + C' dy^2 with one ex2
            const float L2E = 1.4426950408889634f;
            out.rec0[g] = make_float4(cx, cy, hx, hy);
            out.rec1[g] = make_float4(conic0 * (-0.5f * L2E), conic1 * (-L2E), conic2 * (-0.5f * L2E), alpha);
            // rec2.w: 0 when the conic is positive definite with a margin, i.e. power <= 0 at every pixel even
            // after rounding (A, C > 0 and det > 1e-5 A C: the cancellation error of the quadratic form is a few
            // 1e-7 of its terms); 1 tells the blend kernel to keep the shader's `power > 0` discard for this splat
            const bool pd = (conic0 > 0.f) && (conic2 > 0.f) &&
                            ((conic0 * conic2 - conic1 * conic1) > 1e-5f * (conic0 * conic2));
            out.rec2[g] = make_float4(c0, c1, c2, pd ? 0.f : 1.f);
        }
        if (in) {
            uint32_t dk;
            if (P.flags & GSR_FLAG_INTERNAL_DEPTH_ALL) {
                const float d_all = -pv[2];
                dk = (d_all > 0.f) ? __float_as_uint(d_all) : 0xFFFFFFFFu;
            } else {
                dk = ok ? __float_as_uint(depth) : 0xFFFFFFFFu;
            }
            out.depth_key[g] = dk;
            out.trect[g] = trect;
            if (out.depth_hist && dk != 0xFFFFFFFFu) {
                // digits of the key the compacting radix pass will write (culled Gaussians are dropped there)
                const unsigned kc = min(max(dk, P.depth_k0), P.depth_khi);
                n_clamped += (kc != dk) ? 1u : 0u;
                const unsigned kp = kc - P.depth_k0;
                ++n_keys;
#pragma unroll
                for (int p = 0; p < 3; ++p)
                    if (p < P.depth_npass) atomicAdd(&s_dh[p * 512 + ((kp >> P.depth_shift[p]) & P.depth_mask[p])], 1u);
            }
        }
        if (DEBUG && in) {
            float4* d = reinterpret_cast<float4*>(out.dbg_vsout + 16 * (size_t)g);
            d[0] = make_float4(vs[0], vs[1], vs[2], vs[3]);
            d[1] = make_float4(vs[4], vs[5], vs[6], vs[7]);
            d[2] = make_float4(vs[8], vs[9], vs[10], vs[11]);
            d[3] = make_float4(vs[12], vs[13], 0.f, 0.f);
            int4* e = reinterpret_cast<int4*>(out.dbg_setup + 8 * (size_t)g);
            const bool clip_ok = !culled && (ndc2 >= -1.f && ndc2 <= 1.f);
            e[0] = make_int4(__float_as_int(clip_ok ? cx : 0.f), __float_as_int(clip_ok ? cy : 0.f),
                             __float_as_int(clip_ok ? depth : 0.f), ok ? 1 : 0);
            e[1] = make_int4(imin, imax, jmin, jmax);
        }
        n_ok += ok ? 1u : 0u;
        // the next chunk's rows have landed; every lane is done with this chunk's SH rows
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncwarp();
    }
    // visible count: one atomic per CTA
    n_ok = __reduce_add_sync(0xffffffffu, n_ok);
    if ((threadIdx.x & 31) == 0 && n_ok) atomicAdd(&s_visible, n_ok);
    __syncthreads();
    if (threadIdx.x == 0 && s_visible) atomicAdd(&counters->n_visible, s_visible);
    if (out.depth_hist) {
        n_keys = __reduce_add_sync(0xffffffffu, n_keys);
        n_clamped = __reduce_add_sync(0xffffffffu, n_clamped);
        if ((threadIdx.x & 31) == 0) {
            if (n_keys) atomicAdd(&counters->n_keys, n_keys);
            if (n_clamped) atomicAdd(&counters->depth_clamped, n_clamped);
        }
        __syncthreads();
        for (int i = threadIdx.x; i < P.depth_npass * 512; i += 256) {
            const uint32_t c = s_dh[i];
            if (c) atomicAdd(&out.depth_hist[i], c);
        }
    }
}

// ---- official-rasterizer semantics (GSR_FLAG_OFFICIAL) ---------------------------------------------------
// What CUDARenderer.draw hands to diff_gaussian_rasterization (/root/reference/renderer_cuda.py:181-228): the
// package is NOT in the reference tree and not pinned (README.md:29), so this restates the published forward
// pass (preprocessCUDA / computeCov2D / computeColorFromSH / getRect of graphdeco-inria/diff-gaussian-rasterization,
// the pre-antialiasing revision whose call returns (color, radii)) — parity unpinned, checked against the CPU
// restatement in oracle/gs_oracle.c only.  Differences from the shader path: cull is view z <= 0.2 only; the
// footprint is every tile overlapping the square of radius ceil(3 sqrt(lambda_max)) around the pixel position
// ((ndc + 1) S - 1) / 2 (tiles anchored at the TOP-left, rows top-down because the viewer negates view rows 0 and
// 2); colour is clamped below at 0 only; the full SH degree of the scene is used whatever render_mod says
// (renderer_cuda.py:137, set_render_mod is a no-op :145-146).  P.view is the z-forward view, P.proj = P_gl * view.
template <bool DEBUG>
__global__ void __launch_bounds__(256) k_preprocess_official(const FrameParams* __restrict__ fp, SceneBuffers scene,
                                                             SplatBuffers out, FrameCounters* __restrict__ counters,
                                                             unsigned n, int sh_dim) {
    __shared__ SharedParams P;
    __shared__ unsigned s_visible;
    if (threadIdx.x < 16) { P.view[threadIdx.x] = fp->view[threadIdx.x]; P.proj[threadIdx.x] = fp->proj[threadIdx.x]; }
    if (threadIdx.x == 32) {
        P.htanx = fp->htanx; P.htany = fp->htany; P.focal = fp->focal;
        P.campos[0] = fp->campos[0]; P.campos[1] = fp->campos[1]; P.campos[2] = fp->campos[2];
        P.scale_modifier = fp->scale_modifier; P.render_mod = fp->render_mod;
        P.W = fp->W; P.H = fp->H; P.tiles_x = fp->tiles_x; P.tiles_y = fp->tiles_y; P.flags = fp->flags; P.tile_shift = fp->tile_shift;
        s_visible = 0;
    }
    __syncthreads();
    const unsigned g = blockIdx.x * blockDim.x + threadIdx.x;
    bool ok = false;
    if (g < n) {
        float vs[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) vs[k] = 0.f;
        int rminx = 0, rmaxx = 0, rminy = 0, rmaxy = 0;
        float pix = 0.f, piy = 0.f, depth = 0.f;
        float conic0 = 0.f, conic1 = 0.f, conic2 = 0.f, c0 = 0.f, c1 = 0.f, c2 = 0.f, opac = 0.f;
        const float px = __ldg(scene.xyz + 3 * (size_t)g + 0);
        const float py = __ldg(scene.xyz + 3 * (size_t)g + 1);
        const float pz = __ldg(scene.xyz + 3 * (size_t)g + 2);
        const float g_pos[4] = {px, py, pz, 1.f};
        float pv[4];
        mat4_mul_vec4(P.view, g_pos, pv);          // transformPoint4x3
        if (pv[2] > 0.2f) {                        // in_frustum: p_view.z <= 0.2f -> culled
            float ph[4];
            mat4_mul_vec4(P.proj, g_pos, ph);      // transformPoint4x4 with the full projection
            const float p_w = 1.0f / (ph[3] + 0.0000001f);
            const float pp0 = ph[0] * p_w, pp1 = ph[1] * p_w, pp2 = ph[2] * p_w;
            const float4 q = __ldg(scene.rot + g);
            const float4 so = __ldg(scene.scale_op + g);
            // computeCov3D
            m3 S;
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
                for (int r = 0; r < 3; ++r) S.m[c][r] = 0.f;
            S.m[0][0] = P.scale_modifier * so.x; S.m[1][1] = P.scale_modifier * so.y; S.m[2][2] = P.scale_modifier * so.z;
            const float r = q.x, x = q.y, y = q.z, z = q.w;
            m3 R;
            R.m[0][0] = 1.f - 2.f * (y * y + z * z); R.m[0][1] = 2.f * (x * y - r * z); R.m[0][2] = 2.f * (x * z + r * y);
            R.m[1][0] = 2.f * (x * y + r * z); R.m[1][1] = 1.f - 2.f * (x * x + z * z); R.m[1][2] = 2.f * (y * z - r * x);
            R.m[2][0] = 2.f * (x * z - r * y); R.m[2][1] = 2.f * (y * z + r * x); R.m[2][2] = 1.f - 2.f * (x * x + y * y);
            const m3 M = m3_mul(S, R);
            const m3 cov3d = m3_mul(m3_transpose(M), M);
            // computeCov2D
            const float focal_x = (float)P.W / (2.0f * P.htanx), focal_y = (float)P.H / (2.0f * P.htany);
            float tx = pv[0], ty = pv[1];
            const float tz = pv[2];
            const float limx = 1.3f * P.htanx, limy = 1.3f * P.htany;
            const float txtz = tx / tz, tytz = ty / tz;
            tx = gl_min(limx, gl_max(-limx, txtz)) * tz;
            ty = gl_min(limy, gl_max(-limy, tytz)) * tz;
            m3 J;
            J.m[0][0] = focal_x / tz; J.m[0][1] = 0.0f; J.m[0][2] = -(focal_x * tx) / (tz * tz);
            J.m[1][0] = 0.0f; J.m[1][1] = focal_y / tz; J.m[1][2] = -(focal_y * ty) / (tz * tz);
            J.m[2][0] = 0.f; J.m[2][1] = 0.f; J.m[2][2] = 0.f;
            m3 Wm;
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
                for (int rr = 0; rr < 3; ++rr) Wm.m[c][rr] = P.view[c * 4 + rr];
            const m3 T = m3_mul(Wm, J);
            const m3 A = m3_mul(m3_transpose(T), m3_transpose(cov3d));
            m3 cov = m3_mul(A, T);
            cov.m[0][0] += 0.3f;
            cov.m[1][1] += 0.3f;
            const float ca = cov.m[0][0], cb = cov.m[0][1], cc = cov.m[1][1];
            const float det = (ca * cc - cb * cb);
            if (det != 0.0f) {
                const float det_inv = 1.f / det;
                conic0 = cc * det_inv; conic1 = -cb * det_inv; conic2 = ca * det_inv;
                const float mid = 0.5f * (ca + cc);
                const float disc = sqrtf(gl_max(0.1f, mid * mid - det));
                const float lambda1 = mid + disc, lambda2 = mid - disc;
                const float my_radius = ceilf(3.f * sqrtf(gl_max(lambda1, lambda2)));
                // ndc2Pix: ((v + 1.0) * S - 1.0) * 0.5 is evaluated in double (the literals are doubles)
                pix = (float)((((double)pp0 + 1.0) * (double)P.W - 1.0) * 0.5);
                piy = (float)((((double)pp1 + 1.0) * (double)P.H - 1.0) * 0.5);
                if (finitef(pix) && finitef(piy) && finitef(my_radius)) {
                    // getRect (BLOCK_X = BLOCK_Y = 16); the casts are clamped first so they stay defined
                    const float lim = 1.0e9f;
                    const float ax = gl_min(lim, gl_max(-lim, (pix - my_radius) / 16.f));
                    const float bx = gl_min(lim, gl_max(-lim, (pix + my_radius + 15.f) / 16.f));
                    const float ay = gl_min(lim, gl_max(-lim, (piy - my_radius) / 16.f));
                    const float by = gl_min(lim, gl_max(-lim, (piy + my_radius + 15.f) / 16.f));
                    rminx = min(P.tiles_x, max(0, (int)ax)); rmaxx = min(P.tiles_x, max(0, (int)bx));
                    rminy = min(P.tiles_y, max(0, (int)ay)); rmaxy = min(P.tiles_y, max(0, (int)by));
                    ok = ((rmaxx - rminx) * (rmaxy - rminy) != 0);
                }
                // computeColorFromSH at the scene's full degree
                float dx = px - P.campos[0], dy = py - P.campos[1], dz = pz - P.campos[2];
                const float len = sqrtf((dx * dx + dy * dy) + dz * dz);
                dx = dx / len; dy = dy / len; dz = dz / len;
                const float* __restrict__ shp = scene.sh + (size_t)sh_dim * (size_t)g;
                int ncoef = 1;
                if (sh_dim > 3) ncoef = 4;
                if (sh_dim > 12) ncoef = 9;
                if (sh_dim > 27) ncoef = 16;
                float c[3];
#define SHV(k, ch) __ldg(shp + (k) * 3 + (ch))
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) c[ch] = SH_C0 * SHV(0, ch);
                if (ok && ncoef >= 4) {
                    const float X = dx, Y = dy, Z = dz;
#pragma unroll
                    for (int ch = 0; ch < 3; ++ch)
                        c[ch] = c[ch] - SH_C1 * Y * SHV(1, ch) + SH_C1 * Z * SHV(2, ch) - SH_C1 * X * SHV(3, ch);
                    if (ncoef >= 9) {
                        const float xx = X * X, yy = Y * Y, zz = Z * Z;
                        const float xy = X * Y, yz = Y * Z, xz = X * Z;
#pragma unroll
                        for (int ch = 0; ch < 3; ++ch)
                            c[ch] = c[ch] + SH_C2_0 * xy * SHV(4, ch) + SH_C2_1 * yz * SHV(5, ch) +
                                    SH_C2_2 * (2.0f * zz - xx - yy) * SHV(6, ch) + SH_C2_3 * xz * SHV(7, ch) +
                                    SH_C2_4 * (xx - yy) * SHV(8, ch);
                        if (ncoef >= 16) {
#pragma unroll
                            for (int ch = 0; ch < 3; ++ch)
                                c[ch] = c[ch] + SH_C3_0 * Y * (3.0f * xx - yy) * SHV(9, ch) + SH_C3_1 * xy * Z * SHV(10, ch) +
                                        SH_C3_2 * Y * (4.0f * zz - xx - yy) * SHV(11, ch) +
                                        SH_C3_3 * Z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * SHV(12, ch) +
                                        SH_C3_4 * X * (4.0f * zz - xx - yy) * SHV(13, ch) +
                                        SH_C3_5 * Z * (xx - yy) * SHV(14, ch) + SH_C3_6 * X * (xx - 3.0f * yy) * SHV(15, ch);
                        }
                    }
                }
#undef SHV
                c0 = gl_max(c[0] + 0.5f, 0.f); c1 = gl_max(c[1] + 0.5f, 0.f); c2 = gl_max(c[2] + 0.5f, 0.f);
                opac = so.w;
                depth = pv[2];
                if (DEBUG && ok) {
                    vs[0] = 1.f; vs[1] = pp0; vs[2] = pp1; vs[3] = pp2; vs[4] = my_radius; vs[5] = my_radius;
                    vs[6] = conic0; vs[7] = conic1; vs[8] = conic2; vs[9] = c0; vs[10] = c1; vs[11] = c2;
                    vs[12] = opac; vs[13] = depth;
                }
            }
        }
        uint2 trect = make_uint2(0u, 0u);
        if (ok) {
            trect.x = (unsigned)rminx | ((unsigned)(rmaxx - rminx) << 16);
            trect.y = (unsigned)rminy | ((unsigned)(rmaxy - rminy) << 16);
            if (out.tile_grid) {
                const int gw = out.grid_w;
                int* grid = out.tile_grid + (size_t)(blockIdx.x & (GSR_GRID_REPLICAS - 1)) * out.grid_stride;
                atomicAdd(grid + rminy * gw + rminx, 1);
                atomicAdd(grid + rminy * gw + rmaxx, -1);
                atomicAdd(grid + rmaxy * gw + rminx, -1);
                atomicAdd(grid + rmaxy * gw + rmaxx, 1);
            }
            const float L2E = 1.4426950408889634f;
            const float BIG = 3.0e38f;  // every pixel of a touched tile evaluates the splat
            out.rec0[g] = make_float4(pix, piy, BIG, BIG);
            out.rec1[g] = make_float4(conic0 * (-0.5f * L2E), conic1 * (-L2E), conic2 * (-0.5f * L2E), opac);
            out.rec2[g] = make_float4(c0, c1, c2, 1.f);  // renderCUDA always tests power > 0
        }
        out.depth_key[g] = ok ? __float_as_uint(depth) : 0xFFFFFFFFu;
        out.trect[g] = trect;
        if (DEBUG) {
            float4* d = reinterpret_cast<float4*>(out.dbg_vsout + 16 * (size_t)g);
            d[0] = make_float4(vs[0], vs[1], vs[2], vs[3]);
            d[1] = make_float4(vs[4], vs[5], vs[6], vs[7]);
            d[2] = make_float4(vs[8], vs[9], vs[10], vs[11]);
            d[3] = make_float4(vs[12], vs[13], 0.f, 0.f);
            int4* e = reinterpret_cast<int4*>(out.dbg_setup + 8 * (size_t)g);
            e[0] = make_int4(__float_as_int(ok ? pix : 0.f), __float_as_int(ok ? piy : 0.f),
                             __float_as_int(ok ? depth : 0.f), ok ? 1 : 0);
            e[1] = ok ? make_int4(rminx, rmaxx - 1, rminy, rmaxy - 1) : make_int4(0, -1, 0, -1);
        }
    }
    const unsigned b = __ballot_sync(0xffffffffu, ok);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(&s_visible, __popc(b));
    __syncthreads();
    if (threadIdx.x == 0 && s_visible) atomicAdd(&counters->n_visible, s_visible);
}

// Frame start: uniforms + counters, and the two scratch clears the preprocess kernel needs before it runs (the
// corner-impulse grid and the depth-digit histograms) - one launch instead of a kernel and two memsets in front
// of the frame's first real kernel.
__global__ void __launch_bounds__(256) k_set_params(FrameParams* dst, FrameParams fp, FrameCounters* counters,
                                                    int* __restrict__ zero_a, long long na, uint32_t* __restrict__ zero_b, int nb,
                                                    long long n_keys_init) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *dst = fp;
        if (counters) {
            counters->n_visible = 0; counters->n_instances = 0; counters->overflow = 0; counters->n_sort = 0;
            counters->depth_clamped = 0;
            if (n_keys_init >= 0) counters->n_keys = (unsigned)n_keys_init;
        }
    }
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long first = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (long long i = first; i < na; i += stride) zero_a[i] = 0;
    for (long long i = first; i < nb; i += stride) zero_b[i] = 0u;
}

__global__ void k_pack_scene(const float* __restrict__ rot, const float* __restrict__ scale,
                             const float* __restrict__ opacity, float4* __restrict__ rot4,
                             float4* __restrict__ scale_op, long long n) {
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    rot4[g] = make_float4(rot[4 * g + 0], rot[4 * g + 1], rot[4 * g + 2], rot[4 * g + 3]);
    scale_op[g] = make_float4(scale[3 * g + 0], scale[3 * g + 1], scale[3 * g + 2], opacity[g]);
}

// keys for the global far->near order of sort_and_update: ascending view z, -0 == +0, NaN last
// (/root/reference/renderer_ogl.py:10-19; oracle/gs_oracle.c f32_sort_key)
struct View16 { float m[16]; };
__global__ void k_depth_keys(const float* __restrict__ xyz, View16 vm, uint32_t* __restrict__ keys, long long n, int negate) {
    const float* view = vm.m;
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    const float x = __ldg(xyz + 3 * g), y = __ldg(xyz + 3 * g + 1), z = __ldg(xyz + 3 * g + 2);
    float d = ((view[8] * x + view[9] * y) + view[10] * z) + view[11];
    uint32_t k;
    if (d != d) k = 0xFFFFFFFFu;
    else {
        if (negate) d = -d;  // ascending depth = -z: the front-to-back order the tile lists are built in
        d = d + 0.0f;
        const uint32_t u = __float_as_uint(d);
        k = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    }
    keys[g] = k;
}

}  // namespace

void launch_set_params(FrameParams* dst, const FrameParams& fp, FrameCounters* counters, cudaStream_t s, int* zero_a,
                       long long na, uint32_t* zero_b, int nb, long long n_keys_init) {
    const long long work = (na > nb ? na : nb);
    long long grid = (work + 255) / 256;
    if (grid < 1) grid = 1;
    if (grid > 296) grid = 296;
    k_set_params<<<(unsigned)grid, 256, 0, s>>>(dst, fp, counters, zero_a, zero_a ? na : 0, zero_b, zero_b ? nb : 0, n_keys_init);
}

const void* set_params_kernel() { return (const void*)k_set_params; }

int set_params_node_update(cudaGraphExec_t exec, cudaGraphNode_t node, const cudaKernelNodeParams& base, FrameParams* dst,
                           const FrameParams& fp_in, FrameCounters* counters, int* zero_a, long long na, uint32_t* zero_b, int nb,
                           long long n_keys_init) {
    FrameParams fp = fp_in;
    if (!zero_a) na = 0;
    if (!zero_b) nb = 0;
    void* args[8] = {&dst, &fp, &counters, &zero_a, &na, &zero_b, &nb, &n_keys_init};  // k_set_params' parameter list
    cudaKernelNodeParams p = base;
    p.kernelParams = args;
    p.extra = nullptr;
    return cudaGraphExecKernelNodeSetParams(exec, node, &p) == cudaSuccess ? 0 : -1;
}

int sm_count_current() {
    static int cache[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (cache[dev] == 0) {
        cudaDeviceGetAttribute(&cache[dev], cudaDevAttrMultiProcessorCount, dev);
        if (cache[dev] <= 0) cache[dev] = 148;
    }
    return cache[dev];
}

// persistent CTAs: 3 per SM (8 warps x (32 x 13 + 88) float4 of staging each)
static constexpr int PRE_SMEM = 8 * (32 * 13 + 88) * (int)sizeof(float4);
int preprocess_prepare_device() {  // opt-ins for the CURRENT device (gsr_create)
    bool ok = true;
    ok = cudaFuncSetAttribute(k_preprocess<48, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, PRE_SMEM) == cudaSuccess && ok;
    ok = cudaFuncSetAttribute(k_preprocess<48, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, PRE_SMEM) == cudaSuccess && ok;
    ok = cudaFuncSetAttribute(k_preprocess<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, PRE_SMEM) == cudaSuccess && ok;
    ok = cudaFuncSetAttribute(k_preprocess<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, PRE_SMEM) == cudaSuccess && ok;
    ok = cudaFuncSetAttribute(k_preprocess<48, false>, cudaFuncAttributePreferredSharedMemoryCarveout, 100) == cudaSuccess && ok;
    if (!ok) cudaGetLastError();
    return ok ? 0 : -1;
}

void launch_pack_scene(const float* rot, const float* scale, const float* opacity, float4* rot4, float4* scale_op,
                       long long n, cudaStream_t s) {
    if (n <= 0) return;
    k_pack_scene<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(rot, scale, opacity, rot4, scale_op, n);
}

void launch_preprocess(const FrameParams* fp, SceneBuffers scene, SplatBuffers out, FrameCounters* counters,
                       long long n, int sh_dim, cudaStream_t s) {
    if (n <= 0) return;
    const bool dbg = out.dbg_vsout != nullptr;
    // persistent CTAs: 3 per SM, chunks of 256 Gaussians round-robin (opt-ins: preprocess_prepare_device)
    const int smem = PRE_SMEM;
    const int sm_count = sm_count_current();
    const long long chunks = (n + 255) / 256;
    const unsigned grid = (unsigned)(chunks < 3LL * sm_count ? chunks : 3LL * sm_count);
    if (sh_dim == 48) {
        if (dbg) launch_pdl_m(8, k_preprocess<48, true>, dim3(grid), dim3(256), (size_t)smem, s, fp, scene, out, counters, (unsigned)n, sh_dim);
        else launch_pdl_m(8, k_preprocess<48, false>, dim3(grid), dim3(256), (size_t)smem, s, fp, scene, out, counters, (unsigned)n, sh_dim);
    } else {
        if (dbg) launch_pdl_m(8, k_preprocess<0, true>, dim3(grid), dim3(256), (size_t)smem, s, fp, scene, out, counters, (unsigned)n, sh_dim);
        else launch_pdl_m(8, k_preprocess<0, false>, dim3(grid), dim3(256), (size_t)smem, s, fp, scene, out, counters, (unsigned)n, sh_dim);
    }
}

void launch_preprocess_official(const FrameParams* fp, SceneBuffers scene, SplatBuffers out, FrameCounters* counters,
                                long long n, int sh_dim, cudaStream_t s) {
    if (n <= 0) return;
    const unsigned grid = (unsigned)((n + 255) / 256);
    if (out.dbg_vsout) k_preprocess_official<true><<<grid, 256, 0, s>>>(fp, scene, out, counters, (unsigned)n, sh_dim);
    else k_preprocess_official<false><<<grid, 256, 0, s>>>(fp, scene, out, counters, (unsigned)n, sh_dim);
}

void launch_depth_keys(const float* xyz, const float* view16_host, uint32_t* keys, long long n, cudaStream_t s, int negate) {
    if (n <= 0) return;
    View16 vm;
    for (int i = 0; i < 16; ++i) vm.m[i] = view16_host[i];
    k_depth_keys<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(xyz, vm, keys, n, negate);
}
