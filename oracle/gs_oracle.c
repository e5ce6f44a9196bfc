/* TEST INFRASTRUCTURE — not product code.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product path (gaussiansplattingviewer_b200/csrc) never links or calls it.
 *
 * CPU restatement, in plain C / IEEE binary32, of the reference's forward render path:
 *   vertex stage   /root/reference/shaders/gau_vert.glsl:73-251   (cull, cov3D, EWA cov2D, conic,
 *                                                                   quad extents, SH / disparity colour)
 *   fragment stage /root/reference/shaders/gau_frag.glsl:13-39    (falloff, thresholds, modes -2/-3/-4)
 *   depth sort     /root/reference/renderer_ogl.py:10-19          (ascending view z, made stable)
 *   draw + blend   /root/reference/renderer_ogl.py:128-130,175-179 (SRC_ALPHA, ONE_MINUS_SRC_ALPHA,
 *                                                                   painter's order, black clear main.py:821)
 * plus the tile-instance list the B200 pipeline is specified to produce
 * (keys (tile << 32 | depth bits), stable order (tile, depth, index)).
 *
 * PINNING: this restatement is validated against oracle/_ref/libgsref.so, which is the reference's
 * own shader source compiled as C++ (oracle/ref_build.py), and against golden vectors generated
 * from it (tests/golden/, tests/golden/make_golden.py).  Arithmetic conventions (operation order,
 * no FMA contraction, GLSL min/max) are those of oracle/glsl_shim.hpp; fixed-function conventions
 * (viewport transform, pixel coverage, clip, blend) those of oracle/ref_driver.inc.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -fopenmp -fPIC -shared gs_oracle.c -o libgsoracle.so -lm
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "gs_tilecull.h"
#ifdef _OPENMP
#include <omp.h>
#endif

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

typedef struct {
    float view[16]; /* maths layout, row-major: view[r*4+c] */
    float proj[16];
    float hfovxy_focal[3];
    float cam_pos[3];
    float scale_modifier;
    int render_mod;
    int sh_dim;
} GsoCamera;

/* vertex-stage outputs, same layout as VsOut in oracle/ref_driver.inc; pad[0] carries depth = -z_view */
typedef struct {
    float visible;
    float ndc[3];
    float half[2];
    float conic[3];
    float color[3];
    float alpha;
    float depth;
    float pad[2];
} GsoVsOut;

/* raster setup per Gaussian: what the tile pipeline consumes */
typedef struct {
    float cx, cy;   /* window coordinates of the centre, GL y-up */
    float depth;    /* -z_view, the sort key (bits are monotone for depth > 0) */
    int ok;         /* 1: passes shader cull, depth clip, finite, non-empty pixel rect */
    int imin, imax; /* inclusive pixel-column interval covered by the quad, clipped to [0, W-1] */
    int jmin, jmax; /* inclusive pixel-row interval (GL rows, 0 = bottom) */
} GsoSetup;

static inline float gl_min(float x, float y) { return (y < x) ? y : x; }
static inline float gl_max(float x, float y) { return (x < y) ? y : x; }

/* M * v with M in maths layout, terms summed left to right (glsl_shim.hpp mat4*vec4) */
static inline void mat4_mul_vec4(const float* M, const float v[4], float out[4]) {
    for (int r = 0; r < 4; ++r)
        out[r] = ((M[r * 4 + 0] * v[0] + M[r * 4 + 1] * v[1]) + M[r * 4 + 2] * v[2]) + M[r * 4 + 3] * v[3];
}

/* column-major 3x3 helpers mirroring glsl_shim.hpp: m[c][r] */
typedef struct { float m[3][3]; } m3;
static inline m3 m3_mul(const m3* A, const m3* B) {
    m3 R;
    for (int col = 0; col < 3; ++col)
        for (int row = 0; row < 3; ++row)
            R.m[col][row] = (A->m[0][row] * B->m[col][0] + A->m[1][row] * B->m[col][1]) + A->m[2][row] * B->m[col][2];
    return R;
}
static inline m3 m3_transpose(const m3* A) {
    m3 R;
    for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 3; ++r) R.m[c][r] = A->m[r][c];
    return R;
}

/* gau_vert.glsl:73-93 */
static m3 compute_cov3d(const float scale[3], const float q[4]) {
    m3 S;
    memset(&S, 0, sizeof(S));
    S.m[0][0] = scale[0]; S.m[1][1] = scale[1]; S.m[2][2] = scale[2];
    const float r = q[0], x = q[1], y = q[2], z = q[3];
    m3 R;
    R.m[0][0] = 1.f - 2.f * (y * y + z * z); R.m[0][1] = 2.f * (x * y - r * z); R.m[0][2] = 2.f * (x * z + r * y);
    R.m[1][0] = 2.f * (x * y + r * z); R.m[1][1] = 1.f - 2.f * (x * x + z * z); R.m[1][2] = 2.f * (y * z - r * x);
    R.m[2][0] = 2.f * (x * z - r * y); R.m[2][1] = 2.f * (y * z + r * x); R.m[2][2] = 1.f - 2.f * (x * x + y * y);
    m3 M = m3_mul(&S, &R);
    m3 Mt = m3_transpose(&M);
    return m3_mul(&Mt, &M);
}

/* gau_vert.glsl:95-120 */
static void compute_cov2d(const float mean_view[4], float focal_x, float focal_y, float tan_fovx, float tan_fovy,
                          const m3* cov3D, const float* V /* maths layout 4x4 */, float out[3]) {
    float tx = mean_view[0], ty = mean_view[1];
    const float tz = mean_view[2];
    const float limx = 1.3f * tan_fovx;
    const float limy = 1.3f * tan_fovy;
    const float txtz = tx / tz;
    const float tytz = ty / tz;
    tx = gl_min(limx, gl_max(-limx, txtz)) * tz;
    ty = gl_min(limy, gl_max(-limy, tytz)) * tz;
    m3 J;
    J.m[0][0] = focal_x / tz; J.m[0][1] = 0.0f; J.m[0][2] = -(focal_x * tx) / (tz * tz);
    J.m[1][0] = 0.0f; J.m[1][1] = focal_y / tz; J.m[1][2] = -(focal_y * ty) / (tz * tz);
    J.m[2][0] = 0.f; J.m[2][1] = 0.f; J.m[2][2] = 0.f;
    /* W = transpose(mat3(viewmatrix)); viewmatrix[c][r] = V[r][c]  =>  W[c][r] = V[c][r] */
    m3 W;
    for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 3; ++r) W.m[c][r] = V[c * 4 + r];
    m3 T = m3_mul(&W, &J);
    m3 Tt = m3_transpose(&T);
    m3 Ct = m3_transpose(cov3D);
    m3 A = m3_mul(&Tt, &Ct);
    m3 cov = m3_mul(&A, &T);
    cov.m[0][0] += 0.3f;
    cov.m[1][1] += 0.3f;
    out[0] = cov.m[0][0]; out[1] = cov.m[0][1]; out[2] = cov.m[1][1];
}

/* gau_vert.glsl:131-251 for one Gaussian, quad-centre and (+1,+1)-corner outputs */
static void vertex_one(const float* p3, const float* rot4, const float* scale3, float opacity, const float* sh,
                       const GsoCamera* cam, GsoVsOut* o) {
    memset(o, 0, sizeof(*o));
    const float g_pos[4] = {p3[0], p3[1], p3[2], 1.f};
    float pv[4], pc[4];
    mat4_mul_vec4(cam->view, g_pos, pv);
    mat4_mul_vec4(cam->proj, pv, pc);
    const float ndc[3] = {pc[0] / pc[3], pc[1] / pc[3], pc[2] / pc[3]};
    const float lim = (float)1.3;
    if (fabsf(ndc[0]) > lim || fabsf(ndc[1]) > lim || fabsf(ndc[2]) > lim) return; /* :142-146 */
    o->visible = 1.f;
    /* quad centre: gl_Position.xy = ndc.xy + position * quadwh_ndc with position = (0,0); the
     * "+ 0" turns -0 into +0 (a non-finite quad size makes the centre NaN in the shader; such
     * splats are dropped in setup_one by the finite-extent test instead) */
    o->ndc[0] = ndc[0] + 0.f; o->ndc[1] = ndc[1] + 0.f; o->ndc[2] = ndc[2];
    o->depth = -pv[2];

    float s[3];
    for (int k = 0; k < 3; ++k) {
        s[k] = scale3[k] * cam->scale_modifier;
        if (cam->render_mod == -1) s[k] = s[k] * (float)1.2; /* :151-156 */
    }
    m3 cov3d = compute_cov3d(s, rot4);
    float cov2d[3];
    compute_cov2d(pv, cam->hfovxy_focal[2], cam->hfovxy_focal[2], cam->hfovxy_focal[0], cam->hfovxy_focal[1], &cov3d,
                  cam->view, cov2d);
    const float det = (cov2d[0] * cov2d[2] - cov2d[1] * cov2d[1]);
    const float det_inv = 1.f / det;
    o->conic[0] = cov2d[2] * det_inv;
    o->conic[1] = -cov2d[1] * det_inv;
    o->conic[2] = cov2d[0] * det_inv;
    o->half[0] = 3.f * sqrtf(cov2d[0]); /* coordxy at position = (+1,+1) */
    o->half[1] = 3.f * sqrtf(cov2d[2]);
    o->alpha = opacity;

    if (cam->render_mod == -1) { /* :182-211 */
        const float baseline = -0.5f;
        const float xl = (ndc[0] + 1.f) / 2.f; /* p_ndc_l.x is pc.x / pc.w, the same quotient as ndc[0] */
        const float pr[4] = {p3[0] + baseline, p3[1] + 0.f, p3[2] + 0.f, 1.f};
        float pvr[4], pcr[4];
        mat4_mul_vec4(cam->view, pr, pvr);
        mat4_mul_vec4(cam->proj, pvr, pcr);
        const float xr = ((pcr[0] / pcr[3]) + 1.f) / 2.f;
        const float d = fabsf(xl - xr);
        o->color[0] = d; o->color[1] = d; o->color[2] = d;
        return;
    }

    float dir[3] = {p3[0] - cam->cam_pos[0], p3[1] - cam->cam_pos[1], p3[2] - cam->cam_pos[2]};
    const float len = sqrtf((dir[0] * dir[0] + dir[1] * dir[1]) + dir[2] * dir[2]);
    dir[0] = dir[0] / len; dir[1] = dir[1] / len; dir[2] = dir[2] / len;
    const int sh_dim = cam->sh_dim, mod = cam->render_mod;
    float c[3];
#define SHV(k, ch) sh[(k) * 3 + (ch)]
    for (int ch = 0; ch < 3; ++ch) {
        c[ch] = SH_C0 * SHV(0, ch);
    }
    if (sh_dim > 3 && mod >= 1) {
        const float x = dir[0], y = dir[1], z = dir[2];
        for (int ch = 0; ch < 3; ++ch)
            c[ch] = c[ch] - SH_C1 * y * SHV(1, ch) + SH_C1 * z * SHV(2, ch) - SH_C1 * x * SHV(3, ch);
        if (sh_dim > 12 && mod >= 2) {
            const float xx = x * x, yy = y * y, zz = z * z;
            const float xy = x * y, yz = y * z, xz = x * z;
            for (int ch = 0; ch < 3; ++ch)
                c[ch] = c[ch] + SH_C2_0 * xy * SHV(4, ch) + SH_C2_1 * yz * SHV(5, ch) +
                        SH_C2_2 * (2.0f * zz - xx - yy) * SHV(6, ch) + SH_C2_3 * xz * SHV(7, ch) +
                        SH_C2_4 * (xx - yy) * SHV(8, ch);
            if (sh_dim > 27 && mod >= 3) {
                for (int ch = 0; ch < 3; ++ch)
                    c[ch] = c[ch] + SH_C3_0 * y * (3.0f * xx - yy) * SHV(9, ch) + SH_C3_1 * xy * z * SHV(10, ch) +
                            SH_C3_2 * y * (4.0f * zz - xx - yy) * SHV(11, ch) +
                            SH_C3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * SHV(12, ch) +
                            SH_C3_4 * x * (4.0f * zz - xx - yy) * SHV(13, ch) + SH_C3_5 * z * (xx - yy) * SHV(14, ch) +
                            SH_C3_6 * x * (xx - 3.0f * yy) * SHV(15, ch);
            }
        }
    }
#undef SHV
    for (int ch = 0; ch < 3; ++ch) o->color[ch] = c[ch] + 0.5f;
}

/* --- pixel coverage intervals (fixed-function convention of ref_driver.inc) ------------------------
 * Pixel i (centre i + 0.5) is covered iff -h <= fl((i + 0.5) - c) <= h.  fl((i+0.5) - c) is
 * monotone in i, so the covered set is an interval; a float estimate is corrected by direct tests
 * so that the interval agrees EXACTLY with the per-pixel rule. */
static inline int cover_lo(float c, float h, int n) {
    const float f = (c - h) - 0.5f;
    int i;
    if (!(f > 0.f)) i = 0;
    else if (f >= (float)n) i = n;
    else i = (int)floorf(f);
    while (i > 0 && (((float)(i - 1) + 0.5f) - c) >= -h) --i;
    while (i < n && (((float)i + 0.5f) - c) < -h) ++i;
    return i; /* first covered candidate; n if none */
}
static inline int cover_hi(float c, float h, int n) {
    const float g = (c + h) - 0.5f;
    int i;
    if (!(g >= 0.f)) i = -1;
    else if (g >= (float)(n - 1)) i = n - 1;
    else i = (int)floorf(g);
    while (i < n - 1 && (((float)(i + 1) + 0.5f) - c) <= h) ++i;
    while (i >= 0 && (((float)i + 0.5f) - c) > h) --i;
    return i; /* last covered candidate; -1 if none */
}

static void setup_one(const GsoVsOut* v, int W, int H, GsoSetup* s) {
    memset(s, 0, sizeof(*s));
    s->imin = 0; s->imax = -1; s->jmin = 0; s->jmax = -1;
    if (v->visible == 0.f) return;
    if (!(v->ndc[2] >= -1.f && v->ndc[2] <= 1.f)) return; /* depth clip, w = 1 */
    const float cx = ((v->ndc[0] + 1.f) * 0.5f) * (float)W;
    const float cy = ((v->ndc[1] + 1.f) * 0.5f) * (float)H;
    const float hx = v->half[0], hy = v->half[1];
    s->cx = cx; s->cy = cy; s->depth = v->depth;
    if (!isfinite(cx) || !isfinite(cy) || !isfinite(hx) || !isfinite(hy)) return;
    s->imin = cover_lo(cx, hx, W); s->imax = cover_hi(cx, hx, W);
    s->jmin = cover_lo(cy, hy, H); s->jmax = cover_hi(cy, hy, H);
    if (s->imin > s->imax || s->jmin > s->jmax) return;
    s->ok = 1;
}

/* Vertex stage + raster setup for n Gaussians (SoA inputs as in GaussianData). */
int gso_preprocess(const float* xyz, const float* rot, const float* scale, const float* opacity, const float* sh, int n,
                   const GsoCamera* cam, int W, int H, GsoVsOut* vs, GsoSetup* setup) {
#pragma omp parallel for schedule(static)
    for (int g = 0; g < n; ++g) {
        GsoVsOut o;
        vertex_one(xyz + 3 * (size_t)g, rot + 4 * (size_t)g, scale + 3 * (size_t)g, opacity[g],
                   sh + (size_t)cam->sh_dim * (size_t)g, cam, &o);
        if (vs) vs[g] = o;
        if (setup) setup_one(&o, W, H, &setup[g]);
    }
    return 0;
}

/* --- sorting ---------------------------------------------------------------------------------- */
static inline uint32_t f32_bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
/* order-preserving map float -> uint32 for ascending sort; -0 == +0; every NaN sorts last (numpy) */
static inline uint32_t f32_sort_key(float f) {
    if (f != f) return 0xFFFFFFFFu;
    f = f + 0.0f; /* -0 -> +0 */
    uint32_t u = f32_bits(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

/* stable LSD radix sort of (key, val) pairs, 8-bit digits */
static void radix_sort_u32(uint32_t* keys, uint32_t* vals, size_t n) {
    uint32_t* k2 = (uint32_t*)malloc(n * 4);
    uint32_t* v2 = (uint32_t*)malloc(n * 4);
    for (int pass = 0; pass < 4; ++pass) {
        size_t hist[256] = {0};
        const int sh = pass * 8;
        for (size_t i = 0; i < n; ++i) hist[(keys[i] >> sh) & 255]++;
        size_t sum = 0;
        for (int d = 0; d < 256; ++d) { size_t c = hist[d]; hist[d] = sum; sum += c; }
        for (size_t i = 0; i < n; ++i) {
            size_t dst = hist[(keys[i] >> sh) & 255]++;
            k2[dst] = keys[i]; v2[dst] = vals[i];
        }
        uint32_t* t = keys; keys = k2; k2 = t;
        t = vals; vals = v2; v2 = t;
    }
    free(k2); free(v2); /* 4 passes: results are back in the caller's arrays */
}
static void radix_sort_u64(uint64_t* keys, uint32_t* vals, size_t n) {
    uint64_t* k2 = (uint64_t*)malloc(n * 8);
    uint32_t* v2 = (uint32_t*)malloc(n * 4);
    for (int pass = 0; pass < 8; ++pass) {
        size_t hist[256] = {0};
        const int sh = pass * 8;
        for (size_t i = 0; i < n; ++i) hist[(keys[i] >> sh) & 255]++;
        size_t sum = 0;
        for (int d = 0; d < 256; ++d) { size_t c = hist[d]; hist[d] = sum; sum += c; }
        for (size_t i = 0; i < n; ++i) {
            size_t dst = hist[(keys[i] >> sh) & 255]++;
            k2[dst] = keys[i]; v2[dst] = vals[i];
        }
        uint64_t* t = keys; keys = k2; k2 = t;
        uint32_t* tv = vals; vals = v2; v2 = tv;
    }
    free(k2); free(v2);
}

/* renderer_ogl.py:10-19: depth = (V[:3,:3] @ xyz + V[:3,3]).z ; index = argsort(depth) (made stable).
 * out_index: n int32, ascending view z = far -> near.  Also returns the depths when out_depth != NULL. */
int gso_sort_depth(const float* xyz, int n, const float* view, int* out_index, float* out_depth) {
    uint32_t* keys = (uint32_t*)malloc((size_t)n * 4);
    uint32_t* vals = (uint32_t*)malloc((size_t)n * 4);
#pragma omp parallel for schedule(static)
    for (int g = 0; g < n; ++g) {
        const float* p = xyz + 3 * (size_t)g;
        const float z = ((view[8] * p[0] + view[9] * p[1]) + view[10] * p[2]) + view[11];
        if (out_depth) out_depth[g] = z;
        keys[g] = f32_sort_key(z);
        vals[g] = (uint32_t)g;
    }
    radix_sort_u32(keys, vals, (size_t)n);
    for (int g = 0; g < n; ++g) out_index[g] = (int)vals[g];
    free(keys); free(vals);
    return 0;
}

/* Stable sorts exported for direct parity tests of the device radix sort. */
int gso_sort_pairs_u32(uint32_t* keys, uint32_t* vals, long long n) { radix_sort_u32(keys, vals, (size_t)n); return 0; }
int gso_sort_pairs_u64(uint64_t* keys, uint32_t* vals, long long n) { radix_sort_u64(keys, vals, (size_t)n); return 0; }

/* Front-to-back order of the Gaussians that survive setup: stable by (depth, index).
 * out_order: indices, nearest first; returns the count. */
int gso_visible_order(const GsoSetup* setup, int n, int* out_order) {
    uint32_t* keys = (uint32_t*)malloc((size_t)n * 4);
    uint32_t* vals = (uint32_t*)malloc((size_t)n * 4);
    size_t m = 0;
    for (int g = 0; g < n; ++g)
        if (setup[g].ok) { keys[m] = f32_bits(setup[g].depth); vals[m] = (uint32_t)g; ++m; }
    radix_sort_u32(keys, vals, m);
    for (size_t i = 0; i < m; ++i) out_order[i] = (int)vals[i];
    free(keys); free(vals);
    return (int)m;
}

/* Tile-instance list: for every surviving Gaussian, one instance per tile x tile block of pixels its
 * pixel rect overlaps; key = (tile_id << 32) | bits(depth), value = Gaussian index; stable sort by key
 * (ties keep ascending index).  tile_id = ty * tiles_x + tx with GL rows (ty = 0 at the bottom).
 * Pass keys = NULL to get the count only.  ranges: 2 * tiles_x * tiles_y ints (start, end). */
long long gso_tile_instances(const GsoSetup* setup, const GsoVsOut* vs, int cull, int n, int W, int H, int tile,
                             uint64_t* keys, uint32_t* vals, long long capacity, int* ranges) {
    const int tiles_x = (W + tile - 1) / tile, tiles_y = (H + tile - 1) / tile;
    long long total = 0;
    if (tile != GS_TILE) cull = 0;
    for (int g = 0; g < n; ++g) {
        if (!setup[g].ok) continue;
        const int tx0 = setup[g].imin / tile, tx1 = setup[g].imax / tile;
        const int ty0 = setup[g].jmin / tile, ty1 = setup[g].jmax / tile;
        /* oracle/gs_tilecull.h: drop tiles in which no pixel can reach alpha >= 1/255 */
        const int test = cull && vs && ((tx1 - tx0 + 1) * (ty1 - ty0 + 1) <= GS_CULL_MAX_TILES);
        const float thr = test ? gs_cull_threshold(vs[g].alpha) : 0.f;
        for (int ty = ty0; ty <= ty1; ++ty)
            for (int tx = tx0; tx <= tx1; ++tx) {
                if (test && !gs_tile_keep(setup[g].cx, setup[g].cy, setup[g].imin, setup[g].imax, setup[g].jmin,
                                          setup[g].jmax, tx, ty, vs[g].conic[0], vs[g].conic[1], vs[g].conic[2], thr))
                    continue;
                if (keys) {
                    if (total >= capacity) return -1;
                    keys[total] = ((uint64_t)(uint32_t)(ty * tiles_x + tx) << 32) | (uint64_t)f32_bits(setup[g].depth);
                    vals[total] = (uint32_t)g;
                }
                ++total;
            }
    }
    if (!keys) return total;
    radix_sort_u64(keys, vals, (size_t)total);
    if (ranges) {
        for (int t = 0; t < tiles_x * tiles_y; ++t) { ranges[2 * t] = 0; ranges[2 * t + 1] = 0; }
        for (long long i = 0; i < total; ++i) {
            const int t = (int)(keys[i] >> 32);
            if (i == 0 || (int)(keys[i - 1] >> 32) != t) ranges[2 * t] = (int)i;
            if (i == total - 1 || (int)(keys[i + 1] >> 32) != t) ranges[2 * t + 1] = (int)(i + 1);
        }
    }
    return total;
}

/* --- fragment stage + blend ---------------------------------------------------------------------- */
#define FRAG_EPS 2e-5f /* relative width of the "fragile" band around a threshold */

/* gau_frag.glsl:13-39.  Returns 0 when the fragment is discarded.  *fragile is OR-ed with 1 when a
 * threshold decision in this fragment is within FRAG_EPS of flipping (so a different but equally
 * valid exp()/rounding may legitimately decide the other way). */
static inline int fragment(int mod, const float color[3], float alpha, const float conic[3], float dx, float dy,
                           float out[4], int* fragile) {
    if (mod == -2) { out[0] = color[0]; out[1] = color[1]; out[2] = color[2]; out[3] = 1.f; return 1; }
    const float t1 = -0.5f * (conic[0] * dx * dx + conic[2] * dy * dy);
    const float t2 = conic[1] * dx * dy;
    const float power = t1 - t2;
    if (fragile) {
        const float mag = fabsf(t1) + fabsf(t2);
        if (fabsf(power) <= 4.f * FRAG_EPS * mag && mag > 0.f) *fragile |= 1;
    }
    if (power > 0.f) return 0;
    const float e = expf(power);
    const float ae = alpha * e;
    const float opacity = gl_min(0.99f, ae);
    const float thr = 1.f / 255.f;
    if (fragile && fabsf(opacity - thr) <= FRAG_EPS * thr) *fragile |= 1;
    if (opacity < thr) return 0;
    out[0] = color[0]; out[1] = color[1]; out[2] = color[2]; out[3] = opacity;
    if (mod == -3 || mod == -4) {
        if (fragile && fabsf(opacity - 0.22f) <= FRAG_EPS * 0.22f) *fragile |= 1;
        out[3] = (out[3] > 0.22f) ? 1.f : 0.f;
        if (mod == -4) { out[0] = out[0] * e; out[1] = out[1] * e; out[2] = out[2] * e; }
    }
    return 1;
}

/* Draws Gaussians order[0..m) (painter's order: order[0] first = farthest) into a black W x H target.
 * out_rgba: H*W*4 floats, row 0 = top of the displayed image (read-back flip, main.py:912);
 * channel 3 = accumulated coverage.  fragile: optional H*W bytes, set to 1 where any fragment's
 * threshold decision was within FRAG_EPS of flipping.  Parallel over bands of 16 GL rows. */
int gso_draw(const GsoVsOut* vs, const GsoSetup* setup, const int* order, int m, int render_mod, int W, int H,
             int clamp_unorm, float* out_rgba, unsigned char* fragile) {
    const int band = 16;
    const int nb = (H + band - 1) / band;
    memset(out_rgba, 0, sizeof(float) * 4 * (size_t)W * (size_t)H);
    if (fragile) memset(fragile, 0, (size_t)W * (size_t)H);
    /* ordered per-band lists */
    size_t* start = (size_t*)calloc((size_t)nb + 1, sizeof(size_t));
    for (int k = 0; k < m; ++k) {
        const GsoSetup* s = &setup[order[k]];
        if (!s->ok) continue;
        for (int b = s->jmin / band; b <= s->jmax / band; ++b) start[b + 1]++;
    }
    for (int b = 0; b < nb; ++b) start[b + 1] += start[b];
    int* list = (int*)malloc((start[nb] ? start[nb] : 1) * sizeof(int));
    size_t* fill = (size_t*)malloc((size_t)nb * sizeof(size_t));
    for (int b = 0; b < nb; ++b) fill[b] = start[b];
    for (int k = 0; k < m; ++k) {
        const GsoSetup* s = &setup[order[k]];
        if (!s->ok) continue;
        for (int b = s->jmin / band; b <= s->jmax / band; ++b) list[fill[b]++] = order[k];
    }
#pragma omp parallel for schedule(dynamic, 1)
    for (int b = 0; b < nb; ++b) {
        const int j0 = b * band, j1 = (j0 + band < H) ? j0 + band : H;
        for (size_t q = start[b]; q < start[b + 1]; ++q) {
            const int g = list[q];
            const GsoVsOut* v = &vs[g];
            const GsoSetup* s = &setup[g];
            const int ja = s->jmin > j0 ? s->jmin : j0;
            const int jb = s->jmax < j1 - 1 ? s->jmax : j1 - 1;
            for (int j = ja; j <= jb; ++j) {
                const float dy = ((float)j + 0.5f) - s->cy;
                float* row = out_rgba + (size_t)(H - 1 - j) * (size_t)W * 4;
                unsigned char* frow = fragile ? fragile + (size_t)(H - 1 - j) * (size_t)W : NULL;
                for (int i = s->imin; i <= s->imax; ++i) {
                    const float dx = ((float)i + 0.5f) - s->cx;
                    float f[4];
                    int fr = 0;
                    const int keep = fragment(render_mod, v->color, v->alpha, v->conic, dx, dy, f, frow ? &fr : NULL);
                    if (fr) frow[i] = 1;
                    if (!keep) continue;
                    float r = f[0], gg = f[1], bl = f[2], a = f[3];
                    if (clamp_unorm) {
                        r = gl_min(gl_max(r, 0.f), 1.f); gg = gl_min(gl_max(gg, 0.f), 1.f); bl = gl_min(gl_max(bl, 0.f), 1.f);
                        a = gl_min(gl_max(a, 0.f), 1.f);
                    }
                    float* p = row + (size_t)i * 4;
                    const float ia = 1.f - a;
                    p[0] = a * r + ia * p[0];
                    p[1] = a * gg + ia * p[1];
                    p[2] = a * bl + ia * p[2];
                    p[3] = a + ia * p[3];
                }
            }
        }
    }
    free(start); free(list); free(fill);
    return 0;
}

/* Whole frame the way the reference does it: preprocess, global depth order, draw far -> near.
 * The painter's order used is the REVERSE of the stable front-to-back order (depth, index), i.e.
 * on exact depth ties the lower index is in front. */
int gso_render(const float* xyz, const float* rot, const float* scale, const float* opacity, const float* sh, int n,
               const GsoCamera* cam, int W, int H, int clamp_unorm, float* out_rgba, unsigned char* fragile) {
    GsoVsOut* vs = (GsoVsOut*)malloc(sizeof(GsoVsOut) * (size_t)(n ? n : 1));
    GsoSetup* st = (GsoSetup*)malloc(sizeof(GsoSetup) * (size_t)(n ? n : 1));
    int* order = (int*)malloc(sizeof(int) * (size_t)(n ? n : 1));
    gso_preprocess(xyz, rot, scale, opacity, sh, n, cam, W, H, vs, st);
    const int m = gso_visible_order(st, n, order);
    for (int a = 0, b = m - 1; a < b; ++a, --b) { int t = order[a]; order[a] = order[b]; order[b] = t; }
    gso_draw(vs, st, order, m, cam->render_mod, W, H, clamp_unorm, out_rgba, fragile);
    free(vs); free(st); free(order);
    return m;
}

int gso_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* torchrun exports OMP_NUM_THREADS=1; the reference arm of bench.py runs on rank 0 alone and asks for all cores */
void gso_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* ---- official-rasterizer semantics (GSR_FLAG_OFFICIAL) -------------------------------------------------
 * PARITY UNPINNED.  The viewer's CUDA backend (/root/reference/renderer_cuda.py:205-228) calls
 * diff_gaussian_rasterization, which is not in the reference tree, has no pinned version (README.md:29) and
 * cannot be installed offline.  This restates its published forward pass (graphdeco-inria/
 * diff-gaussian-rasterization, cuda_rasterizer/forward.cu + auxiliary.h, the revision whose Python call
 * returns (color, radii) as renderer_cuda.py:215 expects): preprocessCUDA, computeCov3D, computeCov2D,
 * computeColorFromSH, getRect, the (tile << 32 | depth) key order and renderCUDA, anchored on the call site's
 * settings (renderer_cuda.py:104-117: bg 0, prefiltered False, sh_degree sqrt(K)-1) and matrices (:181-203).
 * cam->view = the z-forward view (GL view with rows 0 and 2 negated), cam->proj = P_gl * that view.
 * Conventions as elsewhere in this file: binary32, no FMA contraction, left-to-right sums. */
static int official_one(const float* p3, const float* rot4, const float* scale3, float opacity, const float* sh,
                        const GsoCamera* cam, int W, int H, GsoVsOut* o, GsoSetup* s) {
    memset(o, 0, sizeof(*o));
    memset(s, 0, sizeof(*s));
    s->imin = 0; s->imax = -1; s->jmin = 0; s->jmax = -1;
    const int tiles_x = (W + 15) / 16, tiles_y = (H + 15) / 16;
    const float g_pos[4] = {p3[0], p3[1], p3[2], 1.f};
    float pv[4], ph[4];
    mat4_mul_vec4(cam->view, g_pos, pv);
    if (!(pv[2] > 0.2f)) return 0; /* in_frustum */
    mat4_mul_vec4(cam->proj, g_pos, ph);
    const float p_w = 1.0f / (ph[3] + 0.0000001f);
    const float pp[3] = {ph[0] * p_w, ph[1] * p_w, ph[2] * p_w};
    float sc[3];
    for (int k = 0; k < 3; ++k) sc[k] = cam->scale_modifier * scale3[k];
    m3 cov3d = compute_cov3d(sc, rot4);
    const float focal_x = (float)W / (2.0f * cam->hfovxy_focal[0]);
    const float focal_y = (float)H / (2.0f * cam->hfovxy_focal[1]);
    float cov[3];
    compute_cov2d(pv, focal_x, focal_y, cam->hfovxy_focal[0], cam->hfovxy_focal[1], &cov3d, cam->view, cov);
    const float det = (cov[0] * cov[2] - cov[1] * cov[1]);
    if (det == 0.0f) return 0;
    const float det_inv = 1.f / det;
    const float conic[3] = {cov[2] * det_inv, -cov[1] * det_inv, cov[0] * det_inv};
    const float mid = 0.5f * (cov[0] + cov[2]);
    const float disc = sqrtf(gl_max(0.1f, mid * mid - det));
    const float lambda1 = mid + disc, lambda2 = mid - disc;
    const float radius = ceilf(3.f * sqrtf(gl_max(lambda1, lambda2)));
    const float pix = (float)((((double)pp[0] + 1.0) * (double)W - 1.0) * 0.5); /* ndc2Pix, double literals */
    const float piy = (float)((((double)pp[1] + 1.0) * (double)H - 1.0) * 0.5);
    if (!isfinite(pix) || !isfinite(piy) || !isfinite(radius)) return 0;
    const float lim = 1.0e9f;
    const float ax = gl_min(lim, gl_max(-lim, (pix - radius) / 16.f));
    const float bx = gl_min(lim, gl_max(-lim, (pix + radius + 15.f) / 16.f));
    const float ay = gl_min(lim, gl_max(-lim, (piy - radius) / 16.f));
    const float by = gl_min(lim, gl_max(-lim, (piy + radius + 15.f) / 16.f));
    int rminx = (int)ax, rmaxx = (int)bx, rminy = (int)ay, rmaxy = (int)by;
    rminx = rminx < 0 ? 0 : (rminx > tiles_x ? tiles_x : rminx); rmaxx = rmaxx < 0 ? 0 : (rmaxx > tiles_x ? tiles_x : rmaxx);
    rminy = rminy < 0 ? 0 : (rminy > tiles_y ? tiles_y : rminy); rmaxy = rmaxy < 0 ? 0 : (rmaxy > tiles_y ? tiles_y : rmaxy);
    if ((rmaxx - rminx) * (rmaxy - rminy) == 0) return 0;
    float dir[3] = {p3[0] - cam->cam_pos[0], p3[1] - cam->cam_pos[1], p3[2] - cam->cam_pos[2]};
    const float len = sqrtf((dir[0] * dir[0] + dir[1] * dir[1]) + dir[2] * dir[2]);
    dir[0] = dir[0] / len; dir[1] = dir[1] / len; dir[2] = dir[2] / len;
    const int sh_dim = cam->sh_dim;
    float c[3];
#define SHV(k, ch) sh[(k) * 3 + (ch)]
    for (int ch = 0; ch < 3; ++ch) c[ch] = SH_C0 * SHV(0, ch);
    if (sh_dim > 3) {
        const float x = dir[0], y = dir[1], z = dir[2];
        for (int ch = 0; ch < 3; ++ch)
            c[ch] = c[ch] - SH_C1 * y * SHV(1, ch) + SH_C1 * z * SHV(2, ch) - SH_C1 * x * SHV(3, ch);
        if (sh_dim > 12) {
            const float xx = x * x, yy = y * y, zz = z * z;
            const float xy = x * y, yz = y * z, xz = x * z;
            for (int ch = 0; ch < 3; ++ch)
                c[ch] = c[ch] + SH_C2_0 * xy * SHV(4, ch) + SH_C2_1 * yz * SHV(5, ch) +
                        SH_C2_2 * (2.0f * zz - xx - yy) * SHV(6, ch) + SH_C2_3 * xz * SHV(7, ch) +
                        SH_C2_4 * (xx - yy) * SHV(8, ch);
            if (sh_dim > 27) {
                for (int ch = 0; ch < 3; ++ch)
                    c[ch] = c[ch] + SH_C3_0 * y * (3.0f * xx - yy) * SHV(9, ch) + SH_C3_1 * xy * z * SHV(10, ch) +
                            SH_C3_2 * y * (4.0f * zz - xx - yy) * SHV(11, ch) +
                            SH_C3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * SHV(12, ch) +
                            SH_C3_4 * x * (4.0f * zz - xx - yy) * SHV(13, ch) + SH_C3_5 * z * (xx - yy) * SHV(14, ch) +
                            SH_C3_6 * x * (xx - 3.0f * yy) * SHV(15, ch);
            }
        }
    }
#undef SHV
    o->visible = 1.f;
    o->ndc[0] = pp[0]; o->ndc[1] = pp[1]; o->ndc[2] = pp[2];
    o->half[0] = radius; o->half[1] = radius;
    o->conic[0] = conic[0]; o->conic[1] = conic[1]; o->conic[2] = conic[2];
    for (int ch = 0; ch < 3; ++ch) o->color[ch] = gl_max(c[ch] + 0.5f, 0.f);
    o->alpha = opacity;
    o->depth = pv[2];
    s->cx = pix; s->cy = piy; s->depth = pv[2]; s->ok = 1;
    s->imin = rminx; s->imax = rmaxx - 1; s->jmin = rminy; s->jmax = rmaxy - 1;
    return 1;
}

int gso_official_preprocess(const float* xyz, const float* rot, const float* scale, const float* opacity, const float* sh,
                            int n, const GsoCamera* cam, int W, int H, GsoVsOut* vs, GsoSetup* setup) {
#pragma omp parallel for schedule(static)
    for (int g = 0; g < n; ++g)
        official_one(xyz + 3 * (size_t)g, rot + 4 * (size_t)g, scale + 3 * (size_t)g, opacity[g],
                     sh + (size_t)cam->sh_dim * (size_t)g, cam, W, H, &vs[g], &setup[g]);
    return 0;
}

/* renderCUDA for every tile; out_rgba: H*W*4 floats, row 0 = top, alpha = 1 (renderer_cuda.py:227).
 * fragile: optional H*W bytes, 1 where a threshold decision (alpha < 1/255, power > 0) was within FRAG_EPS of
 * flipping.  Returns the number of Gaussians that passed preprocess. */
int gso_official_render(const float* xyz, const float* rot, const float* scale, const float* opacity, const float* sh,
                        int n, const GsoCamera* cam, int W, int H, float* out_rgba, unsigned char* fragile) {
    GsoVsOut* vs = (GsoVsOut*)malloc(sizeof(GsoVsOut) * (size_t)(n ? n : 1));
    GsoSetup* st = (GsoSetup*)malloc(sizeof(GsoSetup) * (size_t)(n ? n : 1));
    int* order = (int*)malloc(sizeof(int) * (size_t)(n ? n : 1));
    gso_official_preprocess(xyz, rot, scale, opacity, sh, n, cam, W, H, vs, st);
    const int m = gso_visible_order(st, n, order); /* stable by (depth bits, index) */
    const int tiles_x = (W + 15) / 16, tiles_y = (H + 15) / 16, nt = tiles_x * tiles_y;
    size_t* start = (size_t*)calloc((size_t)nt + 1, sizeof(size_t));
    for (int k = 0; k < m; ++k) {
        const GsoSetup* s = &st[order[k]];
        for (int ty = s->jmin; ty <= s->jmax; ++ty)
            for (int tx = s->imin; tx <= s->imax; ++tx) start[ty * tiles_x + tx + 1]++;
    }
    for (int t = 0; t < nt; ++t) start[t + 1] += start[t];
    int* list = (int*)malloc(sizeof(int) * (start[nt] ? start[nt] : 1));
    size_t* fill = (size_t*)malloc(sizeof(size_t) * (size_t)nt);
    memcpy(fill, start, sizeof(size_t) * (size_t)nt);
    for (int k = 0; k < m; ++k) {
        const GsoSetup* s = &st[order[k]];
        for (int ty = s->jmin; ty <= s->jmax; ++ty)
            for (int tx = s->imin; tx <= s->imax; ++tx) list[fill[ty * tiles_x + tx]++] = order[k];
    }
    if (fragile) memset(fragile, 0, (size_t)W * (size_t)H);
    const float FE = 2e-5f;
#pragma omp parallel for schedule(dynamic, 4)
    for (int t = 0; t < nt; ++t) {
        const int tx = t % tiles_x, ty = t / tiles_x;
        for (int j = ty * 16; j < ty * 16 + 16 && j < H; ++j)
            for (int i = tx * 16; i < tx * 16 + 16 && i < W; ++i) {
                float T = 1.f, C[3] = {0.f, 0.f, 0.f};
                unsigned char fr = 0;
                for (size_t k = start[t]; k < start[t + 1]; ++k) {
                    const int g = list[k];
                    const float dx = st[g].cx - (float)i, dy = st[g].cy - (float)j;
                    const float* con = vs[g].conic;
                    const float power = -0.5f * (con[0] * dx * dx + con[2] * dy * dy) - con[1] * dx * dy;
                    if (fabsf(power) <= FE) fr = 1;
                    if (power > 0.f) continue;
                    const float alpha = gl_min(0.99f, vs[g].alpha * expf(power));
                    if (fabsf(alpha - 1.f / 255.f) <= FE * (1.f / 255.f) * 4.f) fr = 1;
                    if (alpha < 1.f / 255.f) continue;
                    const float test_T = T * (1.f - alpha);
                    /* a flipped stop decision moves the pixel by less than 1e-4: not fragile */
                    if (test_T < 0.0001f) break;
                    for (int ch = 0; ch < 3; ++ch) C[ch] += vs[g].color[ch] * alpha * T;
                    T = test_T;
                }
                float* o = out_rgba + 4 * ((size_t)j * (size_t)W + (size_t)i);
                o[0] = C[0]; o[1] = C[1]; o[2] = C[2]; o[3] = 1.f; /* bg = 0: C + T * bg */
                if (fragile) fragile[(size_t)j * (size_t)W + (size_t)i] = fr;
            }
    }
    free(vs); free(st); free(order); free(start); free(list); free(fill);
    return m;
}
