// ply_decode.cu — 3DGS point-cloud rows -> resident scene buffers, on the device (sm_100a).
//
// Stands in for the body of load_ply (/root/reference/util_gau.py:63-125) after the header has been
// parsed on the host: the reference pulls 59 named float columns out of the vertex element with Python
// per-column loops (:96-97, :105-106, :111-112) into float64 temporaries, then applies
//     rot   = rot / ||rot||         in float64, rounded to float32            (:116-117)
//     scale = exp(scale)            in float64, rounded to float32            (:118-119)
//     op    = 1 / (1 + exp(-op))    in float32 (NumPy's float32 exp)          (:120-121)
//     sh    = [f_dc (3) | f_rest channel-major (3,15) -> coefficient-major (15,3)]   (:99-100, :122-123)
// and returns the bounding box and the float32 mean of xyz (:80-85).  Here one kernel reads the raw rows
// once (coalesced, staged through shared memory), writes the four scene planes the renderer keeps
// resident, and reduces the bounding box; a second, deliberately serial kernel reproduces NumPy's
// float32 running sum for the mean bit for bit.
//
// Compiled with -fmad=false; every rounding below is the one NumPy performs:
//   * float64 products/sums/sqrt/divide are IEEE; the norm is ((r0^2 + r1^2) + r2^2) + r3^2.
//   * float64 exp: CUDA's exp() (<= 1 ulp) vs NumPy's; after rounding to float32 the two can differ only
//     when the float64 value sits within 1 ulp(double) of a float32 rounding boundary (p ~ 2^-29 per value).
//   * float32 exp: np_expf() restates NumPy's AVX512F/AVX2 float32 kernel (Cody-Waite reduction, 5/2
//     rational polynomial, scalef); checked bit-identical against NumPy on 3.9e8 inputs over all float32.
#include "gsr_internal.h"
#include <algorithm>

namespace {

constexpr int PLY_ROWS = 128;     // vertex rows per CTA iteration
constexpr int PLY_THREADS = 256;

__device__ __forceinline__ float np_expf(float x) {
    if (x != x) return x;
    if (x > 88.72283935546875f) return __int_as_float(0x7f800000);
    if (x < -103.97207642f) return 0.0f;
    const float magic = 12582912.0f;  // 0x1.8p+23: add/subtract rounds to the nearest integer
    float q = __fmul_rn(x, 1.44269504088896341f);
    q = __fadd_rn(q, magic);
    q = __fsub_rn(q, magic);
    float r = __fmaf_rn(q, -6.93145752e-1f, x);
    r = __fmaf_rn(q, -1.42860677e-6f, r);
    float n = __fmaf_rn(5.082762527590693718096e-04f, r, 6.757896990527504603057e-03f);
    n = __fmaf_rn(n, r, 5.114512081637298353406e-02f);
    n = __fmaf_rn(n, r, 2.473615434895520810817e-01f);
    n = __fmaf_rn(n, r, 7.257664613233124478488e-01f);
    n = __fmaf_rn(n, r, 9.999999999980870924916e-01f);
    float d = __fmaf_rn(2.159509375685829852307e-02f, r, -2.742335390411667452936e-01f);
    d = __fmaf_rn(d, r, 1.0f);
    const float p = __fdiv_rn(n, d);
    return ldexpf(p, (int)q);
}

__device__ __forceinline__ float ld_f32(const unsigned char* row, int off, bool swap) {
    // rows are staged 4-byte aligned when row_bytes and the offsets are multiples of 4 (the host checks)
    unsigned u = *reinterpret_cast<const unsigned*>(row + off);
    if (swap) u = __byte_perm(u, 0, 0x0123);
    return __uint_as_float(u);
}

// float -> int with the same total order (for atomicMin/atomicMax on the bounding box)
__device__ __forceinline__ int f2ord(float f) {
    const int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}

// bbox_ord: [0..2] min, [3..5] max as ordered ints, [6] = bit c set when axis c holds a NaN
__global__ void __launch_bounds__(PLY_THREADS)
k_ply_decode(const unsigned char* __restrict__ rows, PlyLayout lay, long long n, long long first,
             float* __restrict__ xyz, float4* __restrict__ rot4, float4* __restrict__ scale_op,
             float* __restrict__ sh, int* __restrict__ bbox_ord) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    __shared__ int s_shoff[48];
    __shared__ int s_off[59];
    __shared__ int s_red[7];
    const int t = threadIdx.x;
    const int rb = lay.row_bytes;
    const bool swap = lay.big_endian != 0;
    if (t < 48) {
        // output column j of sh: j < 3 -> f_dc_j; else coefficient k = (j-3)/3, channel c = (j-3)%3 -> f_rest_{c*15+k}
        const int src = t < 3 ? 3 + t : 6 + ((t - 3) % 3) * 15 + (t - 3) / 3;
        s_shoff[t] = lay.off[src];
    }
    if (t >= 64 && t < 64 + 59) s_off[t - 64] = lay.off[t - 64];
    if (t < 3) { s_red[t] = 0x7fffffff; s_red[3 + t] = (int)0x80000000; }
    if (t == 6) s_red[6] = 0;
    const long long n_chunks = (n + PLY_ROWS - 1) / PLY_ROWS;
    for (long long chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const long long row0 = chunk * PLY_ROWS;
        const int nrow = (int)min((long long)PLY_ROWS, n - row0);
        __syncthreads();
        {   // stage nrow * rb bytes; chunk starts are 16-byte aligned because PLY_ROWS * rb is
            const uint4* src = reinterpret_cast<const uint4*>(rows + row0 * rb);
            uint4* dst = reinterpret_cast<uint4*>(s_raw);
            const int n16 = (nrow * rb + 15) >> 4;  // the device buffer is padded to 16 bytes
#pragma unroll 4
            for (int i = t; i < n16; i += PLY_THREADS) dst[i] = __ldg(src + i);
        }
        __syncthreads();
        // spherical harmonics: a pure permutation, coalesced on the output side
        {
            float* o = sh + (first + row0) * 48;
            const int total = nrow * 48;
#pragma unroll 4
            for (int e = t; e < total; e += PLY_THREADS) {
                const int v = e / 48, j = e - v * 48;
                o[e] = ld_f32(s_raw + v * rb, s_shoff[j], swap);
            }
        }
        // positions (+ bounding box)
        for (int e = t; e < nrow * 3; e += PLY_THREADS) {
            const int v = e / 3, c = e - v * 3;
            const float p = ld_f32(s_raw + v * rb, s_off[c], swap);
            xyz[(first + row0) * 3 + e] = p;
            if (p != p) atomicOr(&s_red[6], 1 << c);
            else { atomicMin(&s_red[c], f2ord(p)); atomicMax(&s_red[3 + c], f2ord(p)); }
        }
        if (t < PLY_ROWS) {
            if (t < nrow) {  // rotation: float64 normalise, round to float32
                const unsigned char* row = s_raw + t * rb;
                const double r0 = (double)ld_f32(row, s_off[55], swap), r1 = (double)ld_f32(row, s_off[56], swap);
                const double r2 = (double)ld_f32(row, s_off[57], swap), r3 = (double)ld_f32(row, s_off[58], swap);
                const double ss = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(r0, r0), __dmul_rn(r1, r1)), __dmul_rn(r2, r2)),
                                            __dmul_rn(r3, r3));
                const double nrm = __dsqrt_rn(ss);
                rot4[first + row0 + t] = make_float4(__double2float_rn(__ddiv_rn(r0, nrm)), __double2float_rn(__ddiv_rn(r1, nrm)),
                                                     __double2float_rn(__ddiv_rn(r2, nrm)), __double2float_rn(__ddiv_rn(r3, nrm)));
            }
        } else {
            const int v = t - PLY_ROWS;
            if (v < nrow) {  // scale: float64 exp, round to float32; opacity: float32 sigmoid
                const unsigned char* row = s_raw + v * rb;
                const float s0 = __double2float_rn(exp((double)ld_f32(row, s_off[52], swap)));
                const float s1 = __double2float_rn(exp((double)ld_f32(row, s_off[53], swap)));
                const float s2 = __double2float_rn(exp((double)ld_f32(row, s_off[54], swap)));
                const float o = ld_f32(row, s_off[51], swap);
                const float sg = __fdiv_rn(1.0f, __fadd_rn(1.0f, np_expf(-o)));
                scale_op[first + row0 + v] = make_float4(s0, s1, s2, sg);
            }
        }
    }
    __syncthreads();
    if (t < 3) { atomicMin(&bbox_ord[t], s_red[t]); atomicMax(&bbox_ord[3 + t], s_red[3 + t]); }
    if (t == 6 && s_red[6]) atomicOr(&bbox_ord[6], s_red[6]);
}

// NumPy's xyz.mean(axis=0) on a float32 (N,3) array is a float32 running sum over the rows in order,
// divided by N in float32 (checked here against np.cumsum).  A running sum has no parallel form with the
// same roundings, so three threads (one per axis, one per warp) walk the rows while the other threads keep
// a double-buffered shared-memory window filled: ~4 cycles per row, ~6 ms for 3M rows, on its own stream.
constexpr int CEN_ROWS = 1024;
constexpr int CEN_THREADS = 128;

__device__ __forceinline__ void cen_fill(float* dst, const float* __restrict__ xyz, long long row0, long long n, int t) {
    const long long f0 = row0 * 3;
    const long long f1 = min(n, row0 + CEN_ROWS) * 3;
    // 16-byte cp.async where aligned (row0 * 12 is a multiple of 16 because CEN_ROWS is), scalar tail
    const long long nvec = (f1 - f0) >> 2;
    for (long long i = t; i < nvec; i += CEN_THREADS) {
        const unsigned sa = (unsigned)__cvta_generic_to_shared(dst + i * 4);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(xyz + f0 + i * 4));
    }
    for (long long i = f0 + nvec * 4 + t; i < f1; i += CEN_THREADS) dst[i - f0] = xyz[i];
    asm volatile("cp.async.commit_group;\n" ::);
}

// acc3: the running sums, carried from one chunk of rows to the next (chunks are launched in row order on one stream)
__global__ void __launch_bounds__(CEN_THREADS)
k_ply_center(const float* __restrict__ xyz, long long n, float* __restrict__ acc3) {
    __shared__ __align__(16) float s_win[2][CEN_ROWS * 3];
    const int t = threadIdx.x;
    const bool walker = (t & 31) == 0 && t < 96;
    const long long n_chunks = (n + CEN_ROWS - 1) / CEN_ROWS;
    float acc = walker ? acc3[t >> 5] : 0.0f;
    if (n_chunks > 0) cen_fill(s_win[0], xyz, 0, n, t);
    for (long long c = 0; c < n_chunks; ++c) {
        if (c + 1 < n_chunks) {
            cen_fill(s_win[(c + 1) & 1], xyz, (c + 1) * CEN_ROWS, n, t);
            asm volatile("cp.async.wait_group 1;\n" ::);
        } else {
            asm volatile("cp.async.wait_group 0;\n" ::);
        }
        __syncthreads();
        if (walker) {
            const float* w = s_win[c & 1] + (t >> 5);
            const int rows = (int)min((long long)CEN_ROWS, n - c * CEN_ROWS);
            int i = 0;
            if (rows >= 16) {
                // the adds are a 4-cycle dependent chain; the next 16 values are loaded while they run
                float v[16];
#pragma unroll
                for (int k = 0; k < 16; ++k) v[k] = w[k * 3];
                for (; i + 32 <= rows; i += 16) {
                    float nx[16];
#pragma unroll
                    for (int k = 0; k < 16; ++k) nx[k] = w[(i + 16 + k) * 3];
#pragma unroll
                    for (int k = 0; k < 16; ++k) acc = __fadd_rn(acc, v[k]);
#pragma unroll
                    for (int k = 0; k < 16; ++k) v[k] = nx[k];
                }
#pragma unroll
                for (int k = 0; k < 16; ++k) acc = __fadd_rn(acc, v[k]);
                i += 16;
            }
            for (; i < rows; ++i) acc = __fadd_rn(acc, w[i * 3]);
        }
        __syncthreads();
    }
    if (walker) acc3[t >> 5] = acc;
}

// bbox_ord: 7 ints (see k_ply_decode) followed by the 3 running sums
__global__ void k_ply_bbox_init(int* bbox_ord) {
    const int t = threadIdx.x;
    if (t < 3) bbox_ord[t] = 0x7fffffff;
    else if (t < 6) bbox_ord[t] = (int)0x80000000;
    else if (t < 10) bbox_ord[t] = 0;
}

__global__ void k_ply_final(const int* bbox_ord, long long n, float* bbox, float* center) {
    const int t = threadIdx.x;
    if (t < 6) {
        const int o = bbox_ord[t];
        float f = __int_as_float(o >= 0 ? o : o ^ 0x7fffffff);
        if (bbox_ord[6] & (1 << (t % 3))) f = __int_as_float(0x7fc00000);  // np.min/np.max propagate NaN
        bbox[t] = f;
    } else if (t < 9) {
        center[t - 6] = __fdiv_rn(__int_as_float(bbox_ord[7 + t - 6]), (float)n);
    }
}

}  // namespace

size_t ply_decode_smem(int row_bytes) { return ((size_t)PLY_ROWS * row_bytes + 15) & ~(size_t)15; }

int ply_decode_configure(int row_bytes) {
    const size_t smem = ply_decode_smem(row_bytes);
    if (smem > 200 * 1024) return -1;
    if (smem > 48 * 1024)
        if (cudaFuncSetAttribute(k_ply_decode, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -1;
    return 0;
}

void launch_ply_bbox_init(int* bbox_ord, cudaStream_t s) { k_ply_bbox_init<<<1, 32, 0, s>>>(bbox_ord); }

void launch_ply_decode(const unsigned char* rows, const PlyLayout& lay, long long n, long long first, float* xyz,
                       float4* rot4, float4* scale_op, float* sh, int* bbox_ord, int sm_count, cudaStream_t s) {
    if (n <= 0) return;
    const long long n_chunks = (n + PLY_ROWS - 1) / PLY_ROWS;
    const size_t smem = ply_decode_smem(lay.row_bytes);
    const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (200 * 1024) / (smem + 1024)));
    const unsigned grid = (unsigned)std::min<long long>(n_chunks, (long long)sm_count * per_sm);
    k_ply_decode<<<grid, PLY_THREADS, smem, s>>>(rows, lay, n, first, xyz, rot4, scale_op, sh, bbox_ord);
}

void launch_ply_center(const float* xyz_rows, long long n_rows, int* bbox_ord, cudaStream_t s) {
    if (n_rows > 0) k_ply_center<<<1, CEN_THREADS, 0, s>>>(xyz_rows, n_rows, reinterpret_cast<float*>(bbox_ord + 7));
}

void launch_ply_final(const int* bbox_ord, long long n, float* bbox6, float* center3, cudaStream_t s) {
    k_ply_final<<<1, 32, 0, s>>>(bbox_ord, n, bbox6, center3);
}
