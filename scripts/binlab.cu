// binlab - feasibility lab (development tool, not product code): "bin the tile instances with atomics, then sort every
// tile's segment by depth inside one CTA" against the pipeline in the tree (global depth sort + ordered duplication +
// two-pass tile sort).  Input: build/lab/frame_*.bin written by build/lab/gen.py (tile rectangles and depth keys of a frame).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a scripts/binlab.cu -o build/lab/binlab
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

constexpr uint32_t K0 = 0x3D4CCCCDu, KHI = K0 + ((1u << 27) - 1u);

// ---- binning: warp-cooperative emission (the loop of k_duplicate), positions from per-tile atomic cursors
__global__ void __launch_bounds__(256) k_bin(const uint32_t* __restrict__ rx, const uint32_t* __restrict__ ry,
                                             const uint32_t* __restrict__ dk, unsigned n, unsigned tiles_x,
                                             const uint32_t* __restrict__ seg_start, uint32_t* __restrict__ cursor,
                                             uint32_t* __restrict__ out_key, uint32_t* __restrict__ out_gid) {
    const unsigned lane = threadIdx.x & 31u;
    const unsigned warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned n_warps = (gridDim.x * blockDim.x) >> 5;
    for (unsigned base = warp_global * 32u; base < n; base += n_warps * 32u) {
        const unsigned g = base + lane;
        uint32_t x = 0, y = 0, key = 0;
        if (g < n) { x = __ldg(rx + g); y = __ldg(ry + g); key = __ldg(dk + g); }
        const uint32_t cnt = (x >> 16) * (y >> 16);
        key = min(max(key, K0), KHI) - K0;
        uint32_t incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (unsigned)o) incl += v;
        }
        const uint32_t oloc = incl - cnt;
        const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
        for (uint32_t j0 = 0; j0 < total; j0 += 32) {
            const uint32_t j = j0 + lane;
            int pos = 0;
#pragma unroll
            for (int step = 16; step > 0; step >>= 1) {
                const int c = pos + step;
                const uint32_t v = __shfl_sync(0xffffffffu, oloc, c & 31);
                if (c < 32 && v <= j) pos = c;
            }
            const uint32_t ol = __shfl_sync(0xffffffffu, oloc, pos);
            const uint32_t xx = __shfl_sync(0xffffffffu, x, pos);
            const uint32_t yy = __shfl_sync(0xffffffffu, y, pos);
            const uint32_t kk = __shfl_sync(0xffffffffu, key, pos);
            if (j < total) {
                const uint32_t k = j - ol, w = xx >> 16;
                const uint32_t row = k / w;
                const uint32_t tile = (yy & 0xFFFFu) * tiles_x + (xx & 0xFFFFu) + row * tiles_x + (k - row * w);
                const uint32_t dst = __ldg(seg_start + tile) + atomicAdd(cursor + tile, 1u);
                out_key[dst] = kk;
                out_gid[dst] = base + (unsigned)pos;
            }
        }
    }
}

template <int BITS>
__device__ __forceinline__ uint32_t match_digit(uint32_t d) {
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < BITS; ++b) {
        asm("{\n\t.reg .pred p;\n\t.reg .b32 t, m;\n\tand.b32 t, %1, %2;\n\tsetp.ne.b32 p, t, 0;\n\t"
            "vote.sync.ballot.b32 m, p, 0xffffffff;\n\t@p and.b32 %0, %0, m;\n\t@!p lop3.b32 %0, %0, m, 0, 0x30;\n\t}"
            : "+r"(peers) : "r"(d), "r"(1u << b));
    }
    return peers;
}

// ---- one CTA sorts one tile's segment (<= THREADS * ITEMS entries) by its 27-bit key: three 9-bit passes in shared memory
template <int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS) k_local_sort(const uint32_t* __restrict__ seg_list, const uint32_t* __restrict__ seg_start,
                                                        const uint32_t* __restrict__ keys, const uint32_t* __restrict__ gids,
                                                        uint32_t* __restrict__ out_gid, uint32_t* __restrict__ out_key) {
    constexpr int WARPS = THREADS / 32, CAP = THREADS * ITEMS, BINS = 512;
    extern __shared__ __align__(16) unsigned char s_dyn[];
    uint32_t* s_key = reinterpret_cast<uint32_t*>(s_dyn);
    uint32_t* s_gid = s_key + CAP;
    uint16_t* s_cnt = reinterpret_cast<uint16_t*>(s_gid + CAP);  // [WARPS][BINS]
    uint32_t* s_start = reinterpret_cast<uint32_t*>(s_cnt + WARPS * BINS);  // [BINS]
    __shared__ uint32_t s_tmp[10];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t tile = seg_list[blockIdx.x];
    const uint32_t s = seg_start[tile], L = seg_start[tile + 1] - s;
    uint32_t key[ITEMS], gid[ITEMS], rank[ITEMS];
    const int wfirst = warp * (ITEMS * 32) + lane;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const uint32_t idx = wfirst + i * 32;
        key[i] = idx < L ? keys[s + idx] : 0xFFFFFFFFu;
        gid[i] = idx < L ? gids[s + idx] : 0u;
    }
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll 1
    for (int pass = 0; pass < 3; ++pass) {
        const int shift = 9 * pass;
        for (int i = t; i < WARPS * BINS / 8; i += THREADS) reinterpret_cast<uint4*>(s_cnt)[i] = make_uint4(0u, 0u, 0u, 0u);
        __syncthreads();
        uint16_t* wh = s_cnt + warp * BINS;
        const bool warp_live = (uint32_t)(warp * ITEMS * 32) < L;
        if (warp_live) {
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) rank[i] = match_digit<9>((key[i] >> shift) & 511u);
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) {
                const uint32_t d = (key[i] >> shift) & 511u;
                const uint32_t peers = rank[i], below = peers & lt;
                const uint32_t old = wh[d];
                __syncwarp();
                if (below == 0u) wh[d] = (uint16_t)(old + __popc(peers));
                __syncwarp();
                rank[i] = old + __popc(below);
            }
        }
        __syncthreads();
        // digits: the first 256 threads own two each
        uint32_t loc0 = 0, loc1 = 0, sum = 0;
        if (t < 256) {
            uint32_t c0 = 0, c1 = 0;
#pragma unroll 4
            for (int w = 0; w < WARPS; ++w) {
                const uint32_t a = s_cnt[w * BINS + 2 * t], b = s_cnt[w * BINS + 2 * t + 1];
                s_cnt[w * BINS + 2 * t] = (uint16_t)c0; s_cnt[w * BINS + 2 * t + 1] = (uint16_t)c1;
                c0 += a; c1 += b;
            }
            loc0 = c0; loc1 = c1; sum = c0 + c1;
        }
        // exclusive scan over the first 256 threads
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
        if (lane == 31 && warp < 8) s_tmp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = lane < 8 ? s_tmp[lane] : 0u, wi = w;
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += v; }
            if (lane < 8) s_tmp[lane] = wi - w;
        }
        __syncthreads();
        if (t < 256) {
            const uint32_t run = s_tmp[warp] + incl - sum;
            s_start[2 * t] = run; s_start[2 * t + 1] = run + loc0;
        }
        __syncthreads();
        if (warp_live) {
            if (pass < 2) {
#pragma unroll
                for (int i = 0; i < ITEMS; ++i) {
                    const uint32_t d = (key[i] >> shift) & 511u;
                    const uint32_t pos = s_start[d] + wh[d] + rank[i];
                    if (pos < (uint32_t)CAP) { s_key[pos] = key[i]; s_gid[pos] = gid[i]; }
                }
            } else {  // last pass: straight to global memory
#pragma unroll
                for (int i = 0; i < ITEMS; ++i) {
                    const uint32_t d = (key[i] >> shift) & 511u;
                    const uint32_t pos = s_start[d] + wh[d] + rank[i];
                    if (pos < L) { out_gid[s + pos] = gid[i]; if (out_key) out_key[s + pos] = key[i]; }
                }
            }
        }
        if (pass < 2) {
            __syncthreads();
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) {
                const uint32_t idx = wfirst + i * 32;
                key[i] = s_key[idx < (uint32_t)CAP ? idx : 0];
                gid[i] = s_gid[idx < (uint32_t)CAP ? idx : 0];
                if (!(idx < L)) { key[i] = 0xFFFFFFFFu; }
            }
            __syncthreads();
        }
    }
}

template <int THREADS, int ITEMS>
static void launch_sort(const std::vector<uint32_t>& list, uint32_t* d_list, const uint32_t* d_start, const uint32_t* keys,
                        const uint32_t* gids, uint32_t* out_gid, uint32_t* out_key, cudaStream_t st) {
    if (list.empty()) return;
    constexpr size_t smem = (size_t)THREADS * ITEMS * 8 + (size_t)(THREADS / 32) * 512 * 2 + 512 * 4;
    static bool once = false;
    if (!once) { CK(cudaFuncSetAttribute(k_local_sort<THREADS, ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); once = true; }
    k_local_sort<THREADS, ITEMS><<<(unsigned)list.size(), THREADS, smem, st>>>(d_list, d_start, keys, gids, out_gid, out_key);
}

int main(int argc, char** argv) {
    const char* path = argc > 1 ? argv[1] : "build/lab/frame_1080.bin";
    FILE* f = fopen(path, "rb");
    if (!f) { printf("cannot open %s\n", path); return 1; }
    int hdr[4];
    if (fread(hdr, 4, 4, f) != 4) return 1;
    const unsigned n = hdr[0], TX = hdr[1], TY = hdr[2], T = TX * TY;
    std::vector<uint32_t> rx(n), ry(n), dk(n);
    if (fread(rx.data(), 4, n, f) != n || fread(ry.data(), 4, n, f) != n || fread(dk.data(), 4, n, f) != n) return 1;
    fclose(f);
    std::vector<uint32_t> count(T + 1, 0);
    for (unsigned g = 0; g < n; ++g) {
        const unsigned w = rx[g] >> 16, h = ry[g] >> 16, x0 = rx[g] & 0xFFFF, y0 = ry[g] & 0xFFFF;
        for (unsigned y = 0; y < h; ++y) for (unsigned x = 0; x < w; ++x) ++count[(y0 + y) * TX + x0 + x];
    }
    std::vector<uint32_t> start(T + 1, 0);
    for (unsigned i = 0; i < T; ++i) start[i + 1] = start[i] + count[i];
    const unsigned D = start[T];
    std::vector<uint32_t> la, lb, lc; unsigned lmax = 0, too_long = 0;
    std::vector<uint32_t> order(T);
    for (unsigned i = 0; i < T; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return count[a] > count[b]; });  // longest first
    for (unsigned k = 0; k < T; ++k) {
        const unsigned i = order[k], c = count[i];
        lmax = std::max(lmax, c);
        if (c == 0) continue;
        if (c <= 256 * 18) la.push_back(i); else if (c <= 512 * 18) lb.push_back(i); else if (c <= 1024 * 12) lc.push_back(i); else ++too_long;
    }
    printf("%s: N %u tiles %u instances %u longest segment %u; classes A %zu B %zu C %zu, too long %u\n", path, n, T, D, lmax, la.size(), lb.size(), lc.size(), too_long);
    uint32_t *d_rx, *d_ry, *d_dk, *d_start, *d_cursor, *d_key, *d_gid, *d_out, *d_outk, *d_la, *d_lb, *d_lc;
    CK(cudaMalloc(&d_rx, 4ull * n)); CK(cudaMalloc(&d_ry, 4ull * n)); CK(cudaMalloc(&d_dk, 4ull * n));
    CK(cudaMalloc(&d_start, 4ull * (T + 1))); CK(cudaMalloc(&d_cursor, 4ull * T));
    CK(cudaMalloc(&d_key, 4ull * D)); CK(cudaMalloc(&d_gid, 4ull * D)); CK(cudaMalloc(&d_out, 4ull * D)); CK(cudaMalloc(&d_outk, 4ull * D));
    CK(cudaMalloc(&d_la, 4ull * (la.size() + 1))); CK(cudaMalloc(&d_lb, 4ull * (lb.size() + 1))); CK(cudaMalloc(&d_lc, 4ull * (lc.size() + 1)));
    CK(cudaMemcpy(d_rx, rx.data(), 4ull * n, cudaMemcpyHostToDevice)); CK(cudaMemcpy(d_ry, ry.data(), 4ull * n, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_dk, dk.data(), 4ull * n, cudaMemcpyHostToDevice)); CK(cudaMemcpy(d_start, start.data(), 4ull * (T + 1), cudaMemcpyHostToDevice));
    if (!la.empty()) CK(cudaMemcpy(d_la, la.data(), 4ull * la.size(), cudaMemcpyHostToDevice));
    if (!lb.empty()) CK(cudaMemcpy(d_lb, lb.data(), 4ull * lb.size(), cudaMemcpyHostToDevice));
    if (!lc.empty()) CK(cudaMemcpy(d_lc, lc.data(), 4ull * lc.size(), cudaMemcpyHostToDevice));
    cudaStream_t s0, s1, s2;
    CK(cudaStreamCreate(&s0)); CK(cudaStreamCreate(&s1)); CK(cudaStreamCreate(&s2));
    cudaEvent_t e0, e1, e2, ef1, ef2;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreate(&e2));
    CK(cudaEventCreateWithFlags(&ef1, cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&ef2, cudaEventDisableTiming));
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float best_bin = 1e9f, best_sort = 1e9f, best_both = 1e9f;
    for (int rep = 0; rep < 12; ++rep) {
        CK(cudaMemsetAsync(d_cursor, 0, 4ull * T, s0));
        CK(cudaEventRecord(e0, s0));
        k_bin<<<sms * 8, 256, 0, s0>>>(d_rx, d_ry, d_dk, n, TX, d_start, d_cursor, d_key, d_gid);
        CK(cudaEventRecord(e1, s0));
        // the three classes side by side, longest segments first
        CK(cudaStreamWaitEvent(s1, e1, 0)); CK(cudaStreamWaitEvent(s2, e1, 0));
        launch_sort<1024, 12>(lc, d_lc, d_start, d_key, d_gid, d_out, d_outk, s0);
        launch_sort<512, 18>(lb, d_lb, d_start, d_key, d_gid, d_out, d_outk, s1);
        launch_sort<256, 18>(la, d_la, d_start, d_key, d_gid, d_out, d_outk, s2);
        CK(cudaEventRecord(ef1, s1)); CK(cudaEventRecord(ef2, s2));
        CK(cudaStreamWaitEvent(s0, ef1, 0)); CK(cudaStreamWaitEvent(s0, ef2, 0));
        CK(cudaEventRecord(e2, s0));
        CK(cudaStreamSynchronize(s0));
        float a, b, c;
        CK(cudaEventElapsedTime(&a, e0, e1)); CK(cudaEventElapsedTime(&b, e1, e2)); CK(cudaEventElapsedTime(&c, e0, e2));
        if (rep >= 2) { best_bin = std::min(best_bin, a); best_sort = std::min(best_sort, b); best_both = std::min(best_both, c); }
    }
    CK(cudaGetLastError());
    printf("bin %.1f us, local sort %.1f us, both %.1f us\n", best_bin * 1e3f, best_sort * 1e3f, best_both * 1e3f);
    // check: every segment ascending by key, and a permutation of what was binned
    std::vector<uint32_t> ok(D), og(D), bk(D), bg(D);
    CK(cudaMemcpy(ok.data(), d_outk, 4ull * D, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(og.data(), d_out, 4ull * D, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(bk.data(), d_key, 4ull * D, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(bg.data(), d_gid, 4ull * D, cudaMemcpyDeviceToHost));
    unsigned long long bad_order = 0, bad_perm = 0;
    for (unsigned i = 0; i < T; ++i) {
        if (count[i] > 1024 * 12) continue;
        for (unsigned k = start[i] + 1; k < start[i + 1]; ++k) if (ok[k - 1] > ok[k]) ++bad_order;
        unsigned long long s_in = 0, s_out = 0;
        for (unsigned k = start[i]; k < start[i + 1]; ++k) { s_in += bg[k] * 2654435761ull + bk[k]; s_out += og[k] * 2654435761ull + ok[k]; }
        if (s_in != s_out) ++bad_perm;
    }
    printf("check: order violations %llu, segments whose content changed %llu\n", bad_order, bad_perm);
    return 0;
}
