/* gs_tilecull.h — TEST INFRASTRUCTURE (oracle/): a tile-exact footprint rule that was costed for the product
 * (lever 3) and is NOT used by any CUDA kernel; only oracle/gs_oracle.c includes it, to count what
 * the rule would save.  Original header text follows.
 * The tile-instance rule, written so that CUDA kernels and the CPU oracle could share it VERBATIM.
 *
 * A Gaussian whose pixel rectangle overlaps a 16x16 tile gets an instance in that tile's list unless
 * no pixel of the tile can pass the fragment shader's `opacity * exp(power) >= 1/255` test
 * (/root/reference/shaders/gau_frag.glsl:24-26).  Such an instance would contribute nothing, so
 * dropping it changes no pixel; it only shortens the lists that are sorted and blended.
 *
 * The test bounds the quadratic form q(d) = 0.5*(A dx^2 + C dy^2) + B dx dy  (= -power) from below
 * over the box of pixel-centre offsets the tile can see, and keeps the instance iff
 *      q_min <= ln(255 * opacity) + margin.
 * Everything is plain float +,-,*,/ in a fixed order (compile without FMA contraction: gcc
 * -ffp-contract=off, nvcc -fmad=false for the including file), and ln() is replaced by a bounded
 * rational approximation, so the CPU and the GPU take bit-identical decisions.  The margin makes the
 * rule conservative against (a) the approximation error of gs_ln_upper, (b) rounding of q, (c) the
 * approximate exp of the blend kernel.
 *
 * Only the oracle includes this header today; it contains no reference code.
 */
#ifndef GS_TILECULL_H
#define GS_TILECULL_H

#ifdef __CUDACC__
#define GS_HD __host__ __device__ __forceinline__
#else
#define GS_HD static inline
#endif

#define GS_TILE 16
#define GS_CULL_MAX_TILES 256 /* rects touching more tiles than this are not tested (kept whole) */

/* Upper bound of ln(x) for x > 0 with error < 0.01: x = m * 2^e, m in [1,2);
 * ln(m) <= (m-1)*(6+(m-1)) / (6+4(m-1)) + 0.003   (Pade(2,1) of ln(1+t), max error 2.6e-3 on [0,1]) */
GS_HD float gs_ln_upper(float x) {
    union { float f; unsigned u; } v;
    v.f = x;
    const int e = (int)((v.u >> 23) & 0xFFu) - 127;
    v.u = (v.u & 0x007FFFFFu) | 0x3F800000u;
    const float t = v.f - 1.0f;
    const float num = t * (6.0f + t);
    const float den = 6.0f + 4.0f * t;
    return (float)e * 0.69314718f + num / den + 0.003f;
}

GS_HD float gs_minf(float a, float b) { return (b < a) ? b : a; }
GS_HD float gs_maxf(float a, float b) { return (a < b) ? b : a; }

/* min over y in [y0,y1] of 0.5*(A x^2 + C y^2) + B x y for fixed x  (C > 0) */
GS_HD float gs_edge_min(float A, float B, float C, float x, float y0, float y1) {
    float y = -(B * x) / C;
    y = gs_minf(gs_maxf(y, y0), y1);
    return 0.5f * (A * x * x + C * y * y) + B * x * y;
}

/* Per-Gaussian threshold: keep a tile iff q_min <= thr.  opacity <= 0 or NaN -> keep nothing that can be
 * culled (thr = -1: only tiles containing the centre survive ... they cannot contribute either, but the
 * rule must stay monotone), inf/huge -> keep everything. */
GS_HD float gs_cull_threshold(float opacity) {
    if (!(opacity > 0.0f)) return -1.0f;
    const float x = 255.0f * opacity;
    if (!(x < 3.0e38f)) return 3.0e38f;
    const float l = gs_ln_upper(x);
    return l + 0.02f + 0.001f * (l < 0.0f ? -l : l);
}

/* 1 = keep.  (cx, cy): splat centre in window coordinates; rect [imin,imax]x[jmin,jmax]: its pixel
 * rectangle; (tx, ty): tile coordinates; conic (A, B, C); thr from gs_cull_threshold. */
GS_HD int gs_tile_keep(float cx, float cy, int imin, int imax, int jmin, int jmax, int tx, int ty, float A, float B,
                       float C, float thr) {
    int x0 = tx * GS_TILE, x1 = x0 + GS_TILE - 1, y0 = ty * GS_TILE, y1 = y0 + GS_TILE - 1;
    if (x0 < imin) x0 = imin;
    if (x1 > imax) x1 = imax;
    if (y0 < jmin) y0 = jmin;
    if (y1 > jmax) y1 = jmax;
    const float det = A * C - B * B;
    if (!(A > 0.0f) || !(C > 0.0f) || !(det > 0.0f)) return 1; /* not positive definite: no bound, keep */
    const float dx0 = ((float)x0 + 0.5f) - cx, dx1 = ((float)x1 + 0.5f) - cx;
    const float dy0 = ((float)y0 + 0.5f) - cy, dy1 = ((float)y1 + 0.5f) - cy;
    if (dx0 <= 0.0f && dx1 >= 0.0f && dy0 <= 0.0f && dy1 >= 0.0f) return 1; /* centre inside the box */
    float q = gs_edge_min(A, B, C, dx0, dy0, dy1);
    q = gs_minf(q, gs_edge_min(A, B, C, dx1, dy0, dy1));
    q = gs_minf(q, gs_edge_min(C, B, A, dy0, dx0, dx1));
    q = gs_minf(q, gs_edge_min(C, B, A, dy1, dx0, dx1));
    return (q <= thr) ? 1 : 0;
}

#endif /* GS_TILECULL_H */
