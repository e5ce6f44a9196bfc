// K2 duplicate-with-keys and K4 tile ranges for sm_100a.
//
// K2 turns the per-Gaussian tile rectangles into the tile-instance list.  One pass:
//   * each CTA takes a chunk of 1024 Gaussians (in the order given: depth-sorted for the two-level
//     path, index order for the 64-bit-key path), block-scans their tile counts, chains the chunk total
//     to the previous chunks with a decoupled look-back (64-bit status word: 2 flag bits + count), so the
//     prefix sum never makes a round trip through global memory;
//   * instances are emitted warp-cooperatively and load balanced over instances: lane l of a warp writes
//     output slot s * 32 + l of the warp's current 32 Gaussians and finds the owning Gaussian with a
//     5-step binary search over the lanes' offsets (shuffles); the tile index comes from a multiply-high
//     with a per-Gaussian magic reciprocal (no division); stores are contiguous.  Rectangles above
//     DUP_BIG tiles go to a global list that k_duplicate_big spreads over the whole grid: the depth sort
//     puts the Gaussians nearest to the camera - the ones with screen-sized rectangles - next to each
//     other, so expanding them inside their own chunk left one CTA with most of the frame's instances.
// Counterpart of the scan + duplicateWithKeys + identifyTileRanges steps of the tile rasteriser the
// reference's CUDA backend calls (/root/reference/renderer_cuda.py:211-224); the OpenGL backend has no
// tiles (its draw is one instanced quad per Gaussian, renderer_ogl.py:175-179).
#include "gsr_internal.h"

namespace {

constexpr int DUP_THREADS = 256;
constexpr int DUP_WARPS = DUP_THREADS / 32;
constexpr int DUP_ITEMS = 4;
constexpr int DUP_CHUNK = DUP_THREADS * DUP_ITEMS;
constexpr uint32_t DUP_BIG = 256;    // instances; larger rectangles are expanded by k_duplicate_big
constexpr unsigned long long CH_FLAG_AGG = 1ull << 62;
constexpr unsigned long long CH_FLAG_INC = 2ull << 62;
constexpr unsigned long long CH_FLAG_MASK = 3ull << 62;
constexpr unsigned long long CH_VAL_MASK = (1ull << 62) - 1ull;

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// k / w for k < 65536 by multiply-high with magic = ceil(2^32 / w) (exact for k, w < 2^16); w == 1 and
// huge rectangles take the plain paths.
__device__ __forceinline__ uint32_t magic_of(uint32_t w) { return (w > 1u) ? (0xFFFFFFFFu / w + 1u) : 0u; }
__device__ __forceinline__ uint32_t div_magic(uint32_t k, uint32_t w, uint32_t magic, bool small) {
    if (w == 1u) return k;
    return small ? __umulhi(k, magic) : k / w;
}

#ifdef GSR_DUP_TRACE  // development only: globaltimer stamps of the first chunk of every CTA (scripts/probe_dup_trace.py)
__device__ unsigned long long g_dup_trace[2048 * 8];
__device__ __forceinline__ unsigned long long dup_now() { unsigned long long v; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(v)); return v; }
#define DUP_T(i) do { if (threadIdx.x == 0 && dup_first && blockIdx.x < 2048) g_dup_trace[blockIdx.x * 8 + (i)] = dup_now(); } while (0)
#else
#define DUP_T(i) do { } while (0)
#endif

// KEYS64 = false: gid comes from sorted_gid, output key is the 32-bit tile id.
// KEYS64 = true : gid is the running index, output key is (tile << 32) | depth bits.
template <bool KEYS64>
__global__ void __launch_bounds__(DUP_THREADS) k_duplicate(const FrameParams* __restrict__ fp,
                                                          const uint32_t* __restrict__ sorted_gid, SplatBuffers sp,
                                                          FrameCounters* __restrict__ counters,
                                                          unsigned long long* __restrict__ chain,
                                                          uint32_t* __restrict__ ticket, void* __restrict__ out_keys,
                                                          uint32_t* __restrict__ out_gid, unsigned n,
                                                          unsigned long long capacity, uint4* __restrict__ big_list,
                                                          uint32_t* __restrict__ big_count) {
#ifdef GSR_DUP_TRACE
    bool dup_first = true; unsigned long long dup_chunks = 0;
    DUP_T(0);
#endif
    pdl_wait();  // no global memory access before this (programmatic dependent launch)
    DUP_T(1);
    __shared__ uint32_t s_warp[DUP_WARPS + 1];
    __shared__ uint32_t s_chunk;
    __shared__ unsigned long long s_prefix;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t tiles_x = (uint32_t)fp->tiles_x;
    // the two-level path walks a depth order of counters->n_keys entries (all N, or only the Gaussians with a key)
    if (!KEYS64) n = min(n, counters->n_keys);
    const unsigned n_chunks = (n + DUP_CHUNK - 1) / DUP_CHUNK;

    auto emit = [&](unsigned long long dst, uint32_t tile, uint32_t g) {
        if (dst < capacity) {
            if (KEYS64)
                reinterpret_cast<unsigned long long*>(out_keys)[dst] =
                    ((unsigned long long)tile << 32) | (unsigned long long)__ldg(sp.depth_key + g);
            else
                reinterpret_cast<uint32_t*>(out_keys)[dst] = tile;
            out_gid[dst] = g;
        }
    };

    while (true) {
        if (t == 0) s_chunk = atomicAdd(ticket, 1u);
        __syncthreads();
        const unsigned chunk = s_chunk;
        if (chunk >= n_chunks) break;
        DUP_T(2);
#ifdef GSR_DUP_TRACE
        ++dup_chunks;
#endif
        // ---- load: lane-strided inside the warp's 128 slots, so the gid reads coalesce
        const unsigned wbase = chunk * DUP_CHUNK + warp * (32 * DUP_ITEMS);
        uint32_t gid[DUP_ITEMS], cnt[DUP_ITEMS], rx[DUP_ITEMS], ry[DUP_ITEMS];
        // (all index loads first, then all rectangle loads: a warp issues in order, and with each rectangle load
        // right behind its index load the phase was DUP_ITEMS x 2 dependent round trips - 5.5 us per chunk in the
        // globaltimer trace, scripts/probe_dup_trace.py)
#pragma unroll
        for (int i = 0; i < DUP_ITEMS; ++i) {
            const unsigned r = wbase + i * 32 + lane;
            gid[i] = KEYS64 ? r : ((r < n) ? __ldg(sorted_gid + r) : 0u);
        }
#pragma unroll
        for (int i = 0; i < DUP_ITEMS; ++i) {
            const unsigned r = wbase + i * 32 + lane;
            uint2 tr = make_uint2(0u, 0u);
            if (r < n) tr = __ldg(sp.trect + gid[i]);
            rx[i] = tr.x; ry[i] = tr.y;
        }
#pragma unroll
        for (int i = 0; i < DUP_ITEMS; ++i) {
            if (KEYS64 && !(wbase + i * 32 + lane < n)) gid[i] = 0;
            cnt[i] = (rx[i] >> 16) * (ry[i] >> 16);
        }
        // ---- scan in slot order (slot = i * 32 + lane): per-round warp scans, then rounds, then warps
        uint32_t off[DUP_ITEMS];
        uint32_t wsum = 0;
#pragma unroll
        for (int i = 0; i < DUP_ITEMS; ++i) {
            uint32_t incl = cnt[i];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            off[i] = wsum + incl - cnt[i];
            wsum += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (lane == 0) s_warp[warp] = wsum;
        __syncthreads();
        if (warp == 0) {
            const uint32_t w = (lane < DUP_WARPS) ? s_warp[lane] : 0u;
            uint32_t wi = w;
#pragma unroll
            for (int o = 1; o < DUP_WARPS; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += v;
            }
            if (lane < DUP_WARPS) s_warp[lane] = wi - w;
            if (lane == DUP_WARPS - 1) s_warp[DUP_WARPS] = wi;
        }
        __syncthreads();
        const uint32_t chunk_total = s_warp[DUP_WARPS];
        const uint32_t wexcl = s_warp[warp];
        DUP_T(3);
        // ---- chain the chunk total: warp 0 probes 32 predecessors per round trip (a one-thread walk left the
        // other 255 threads at the barrier for most of the kernel: ncu barrier stall 9 per issue)
        if (warp == 0) {
            unsigned long long prefix = 0;
            if (chunk == 0) {
                if (lane == 0) st_volatile_u64(&chain[0], CH_FLAG_INC | (unsigned long long)chunk_total);
            } else {
                if (lane == 0) st_volatile_u64(&chain[chunk], CH_FLAG_AGG | (unsigned long long)chunk_total);
                long long p = (long long)chunk - 1;
                while (true) {
                    const unsigned long long wd = (p - lane >= 0) ? ld_volatile_u64(&chain[p - lane]) : CH_FLAG_INC;
                    const unsigned rdy = __ballot_sync(0xffffffffu, (wd & CH_FLAG_MASK) != 0ull);
                    const unsigned inc = __ballot_sync(0xffffffffu, (wd & CH_FLAG_INC) != 0ull);
                    const int n_ready = (rdy == 0xffffffffu) ? 32 : (__ffs(~rdy) - 1);  // usable: a run from lane 0
                    const unsigned inc_in = inc & ((n_ready == 32) ? 0xffffffffu : ((1u << n_ready) - 1u));
                    const int upto = inc_in ? __ffs(inc_in) : n_ready;  // lanes to add (the INC lane included)
                    unsigned long long v = (lane < upto) ? (wd & CH_VAL_MASK) : 0ull;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                    prefix += v;
                    if (inc_in) break;
                    p -= upto;
                    if (upto == 0) __nanosleep(128);
                }
                if (lane == 0) st_volatile_u64(&chain[chunk], CH_FLAG_INC | (prefix + chunk_total));
            }
            if (lane == 0) s_prefix = prefix;  // D and the overflow flag come from k_tile_scan
        }
        // per-Gaussian reciprocal (owner lane)
        uint32_t magic[DUP_ITEMS];
#pragma unroll
        for (int i = 0; i < DUP_ITEMS; ++i) {
            magic[i] = magic_of(rx[i] >> 16);
            off[i] += wexcl;
        }
        __syncthreads();
        DUP_T(4);
        const unsigned long long prefix = s_prefix;
        // large rectangles: (first output slot, rectangle, Gaussian) to the global list
#pragma unroll
        for (int i = 0; i < DUP_ITEMS; ++i) {
            if (cnt[i] > DUP_BIG) {
                const unsigned long long dst = prefix + off[i];
                const uint32_t b = atomicAdd(big_count, 1u);
                big_list[b] = make_uint4(dst > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)dst, rx[i], ry[i], gid[i]);
                cnt[i] = 0;
            }
        }
        // ---- warp-cooperative emission, load balanced over INSTANCES: in each round the 32 Gaussians of the
        // warp (one per lane) own consecutive runs of the round's output; lane l of step s writes output
        // s * 32 + l and finds its owner by a 5-step binary search over the lanes' offsets (shuffles), so all
        // 32 lanes store no matter how small the rectangles are (one Gaussian per step kept 24 % of the
        // lanes busy at 7.6 tiles per Gaussian)
#pragma unroll
        for (int i = 0; i < DUP_ITEMS; ++i) {
            // offsets of the round without the deferred rectangles
            uint32_t incl = cnt[i];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            const uint32_t oloc = incl - cnt[i];
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
                const uint32_t ot = __shfl_sync(0xffffffffu, off[i], pos);
                const uint32_t x = __shfl_sync(0xffffffffu, rx[i], pos);
                const uint32_t y = __shfl_sync(0xffffffffu, ry[i], pos);
                const uint32_t g = __shfl_sync(0xffffffffu, gid[i], pos);
                const uint32_t m = __shfl_sync(0xffffffffu, magic[i], pos);
                if (j < total) {
                    const uint32_t k = j - ol;
                    const uint32_t w = x >> 16;
                    const uint32_t row = div_magic(k, w, m, true);  // rectangle <= DUP_BIG < 65536 instances
                    emit(prefix + ot + k, (y & 0xFFFFu) * tiles_x + (x & 0xFFFFu) + row * tiles_x + (k - row * w), g);
                }
            }
        }
        __syncthreads();
        DUP_T(5);
#ifdef GSR_DUP_TRACE
        dup_first = false;
#endif
    }
#ifdef GSR_DUP_TRACE
    if (threadIdx.x == 0 && blockIdx.x < 2048) { g_dup_trace[blockIdx.x * 8 + 6] = dup_now(); g_dup_trace[blockIdx.x * 8 + 7] = dup_chunks; }
#endif
}

// Expands the large rectangles k_duplicate put on the global list: one warp per entry, entries dealt
// round-robin to all warps of the grid (an entry is at most one screen of tiles).
template <bool KEYS64>
__global__ void __launch_bounds__(DUP_THREADS) k_duplicate_big(const FrameParams* __restrict__ fp, SplatBuffers sp,
                                                              const uint4* __restrict__ big_list,
                                                              const uint32_t* __restrict__ big_count,
                                                              void* __restrict__ out_keys, uint32_t* __restrict__ out_gid,
                                                              unsigned long long capacity) {
    pdl_wait();  // resident behind k_duplicate, which fills the list
    const uint32_t nb = *big_count;
    const uint32_t tiles_x = (uint32_t)fp->tiles_x;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t n_warps = gridDim.x * (DUP_THREADS / 32);
    // consecutive entries to different CTAs: neighbours on the list are neighbours in depth and similar in size
    for (uint32_t b = blockIdx.x + gridDim.x * (threadIdx.x >> 5); b < nb; b += n_warps) {
        const uint4 e = __ldg(big_list + b);
        const uint32_t w = e.y >> 16, c = w * (e.z >> 16);
        const uint32_t tile0 = (e.z & 0xFFFFu) * tiles_x + (e.y & 0xFFFFu);
        const uint32_t m = magic_of(w);
        const bool small = c <= 65536u;
        uint32_t dk = 0;
        if (KEYS64) dk = __ldg(sp.depth_key + e.w);
        for (uint32_t k = lane; k < c; k += 32) {
            const uint32_t row = div_magic(k, w, m, small);
            const uint32_t tile = tile0 + row * tiles_x + (k - row * w);
            const unsigned long long dst = (unsigned long long)e.x + k;
            if (dst < capacity) {
                if (KEYS64)
                    reinterpret_cast<unsigned long long*>(out_keys)[dst] = ((unsigned long long)tile << 32) | dk;
                else
                    reinterpret_cast<uint32_t*>(out_keys)[dst] = tile;
                out_gid[dst] = e.w;
            }
        }
    }
}

// K4' — tile counts, list offsets and sort histograms from the corner-impulse grid (one CTA).
// The preprocess kernel added +1/-1/-1/+1 at the four corners of every visible Gaussian's tile rectangle;
// the 2D inclusive prefix sum of that grid is the number of instances per tile.  From the counts follow
//   ranges[tile] = [start, start + count)   (exclusive scan in tile-id order = order of the sorted list),
//   the digit histograms of the two radix passes of the tile sort (no pass over the D instances needed),
//   D itself and the overflow flag.
constexpr int TS_THREADS = 1024;
__global__ void __launch_bounds__(TS_THREADS) k_tile_scan(int* __restrict__ grid, int grid_stride, int gw, int TX, int TY, int bits0,
                                                         int use_smem, unsigned long long capacity,
                                                         int2* __restrict__ ranges, uint32_t* __restrict__ hist,
                                                         FrameCounters* __restrict__ counters) {
    extern __shared__ int s_dyn[];
    __shared__ uint32_t s_hist[512];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int T = TX * TY;
    int* w = use_smem ? s_dyn : grid;
    // sum the replicas; 4 elements x 16 replicas = 64 independent loads in flight per thread (one element at
    // a time made this loop a chain of L2 round trips: 8 us of the kernel's 21 at 1080p)
    for (int i0 = t; i0 < TY * gw; i0 += 4 * TS_THREADS) {
        int v[4] = {0, 0, 0, 0};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + u * TS_THREADS;
            if (i < TY * gw) {
#pragma unroll
                for (int r = 0; r < GSR_GRID_REPLICAS; ++r) v[u] += __ldg(grid + (size_t)r * grid_stride + i);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + u * TS_THREADS;
            if (i < TY * gw) w[i] = v[u];
        }
    }
    if (t < 512) s_hist[t] = 0;
    __syncthreads();
    // inclusive prefix along x: one warp per row, 32 columns per step
    for (int r = warp; r < TY; r += TS_THREADS / 32) {
        int carry = 0;
        for (int c0 = 0; c0 < TX; c0 += 32) {
            const int x = c0 + lane;
            int v = (x < TX) ? w[r * gw + x] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o) v += u;
            }
            v += carry;
            if (x < TX) w[r * gw + x] = v;
            carry = __shfl_sync(0xffffffffu, v, 31);
        }
    }
    __syncthreads();
    // inclusive prefix along y: one thread per column
    for (int x = t; x < TX; x += TS_THREADS) {
        int run = 0;
        for (int y = 0; y < TY; ++y) { run += w[y * gw + x]; w[y * gw + x] = run; }
    }
    __syncthreads();
    // exclusive scan of the counts in tile-id order; thread t owns ids [t * per, (t + 1) * per)
    const int per = (T + TS_THREADS - 1) / TS_THREADS;
    const int id0 = min(t * per, T), id1 = min(id0 + per, T);
    unsigned long long sum = 0;
    {
        int y = id0 / TX, x = id0 - y * TX;
        for (int id = id0; id < id1; ++id) {
            sum += (unsigned)w[y * gw + x];
            if (++x == TX) { x = 0; ++y; }
        }
    }
    // block exclusive scan (64-bit: D may exceed 2^32 before the overflow check)
    unsigned long long incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    __shared__ unsigned long long s_w64[TS_THREADS / 32 + 1];
    if (lane == 31) s_w64[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const unsigned long long v = s_w64[lane];
        unsigned long long vi = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long u = __shfl_up_sync(0xffffffffu, vi, o);
            if (lane >= o) vi += u;
        }
        s_w64[lane] = vi - v;
        if (lane == 31) s_w64[32] = vi;
    }
    __syncthreads();
    const unsigned long long total = s_w64[32];
    const bool overflow = total > capacity;
    unsigned long long start = s_w64[warp] + incl - sum;
    {
        // ranges + the high-digit histogram: the ids of a thread are consecutive, so they share their high
        // digit except across a 2^bits0 boundary -> accumulate in a register, one shared atomic per run
        int y = id0 / TX, x = id0 - y * TX;
        uint32_t run_digit = 0xFFFFFFFFu, run_sum = 0;
        for (int id = id0; id < id1; ++id) {
            const unsigned c = (unsigned)w[y * gw + x];
            ranges[id] = (overflow || c == 0) ? make_int2(0, 0) : make_int2((int)start, (int)(start + c));  // empty: (0, 0)
            const uint32_t hd = (uint32_t)id >> bits0;
            if (hist && hd != run_digit) {
                if (run_sum) atomicAdd(&s_hist[256 + run_digit], run_sum);
                run_digit = hd; run_sum = 0;
            }
            run_sum += c;
            start += c;
            if (++x == TX) { x = 0; ++y; }
        }
        if (hist && run_sum) atomicAdd(&s_hist[256 + run_digit], run_sum);
    }
    if (hist) {  // nullptr: more than 16 tile bits, the tile sort counts its digits itself (k_rs_hist)
        // low-digit histogram: thread t sums the tiles id = d + 2^bits0 * (slice + n_slices * j) of its digit d
        const int nd = 1 << bits0;
        const int n_slices = TS_THREADS >> bits0;  // >= 4 (bits0 <= 8)
        const int d = t & (nd - 1), slice = t >> bits0;
        uint32_t acc = 0;
        for (int id = d + nd * slice; id < T; id += nd * n_slices) {
            const int y = id / TX, x = id - y * TX;
            acc += (unsigned)w[y * gw + x];
        }
        if (acc) atomicAdd(&s_hist[d], acc);
    }
    __syncthreads();
    if (hist && t < 512) hist[t] = s_hist[t];
    if (t == 0) {
        counters->n_instances = (total > 0xFFFFFFFFull) ? 0xFFFFFFFFu : (unsigned)total;
        counters->overflow = overflow ? 1u : 0u;
        if (overflow) {
            counters->sticky_overflow += 1u;
            counters->sticky_max_d = max(counters->sticky_max_d, (unsigned)min(total, (unsigned long long)0xFFFFFFFFu));
        }
        counters->n_sort = overflow ? 0u : (unsigned)total;
    }
}

template <typename KeyT>
__global__ void __launch_bounds__(256) k_tile_ranges(const KeyT* __restrict__ keys, int key_shift,
                                                    const FrameCounters* __restrict__ counters, long long capacity,
                                                    int2* __restrict__ ranges) {
    long long n = counters->n_instances;
    if (n > capacity) n = capacity;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint32_t cur = (uint32_t)(keys[i] >> key_shift);
        if (i == 0) ranges[cur].x = 0;
        else {
            const uint32_t prev = (uint32_t)(keys[i - 1] >> key_shift);
            if (prev != cur) { ranges[prev].y = (int)i; ranges[cur].x = (int)i; }
        }
        if (i == n - 1) ranges[cur].y = (int)n;
    }
}

// debug view of the sorted list as 64-bit keys: the two-level path never materialises them
__global__ void k_keys64_from_ranges(const int2* __restrict__ ranges, int n_tiles, const uint32_t* __restrict__ gid,
                                     const uint32_t* __restrict__ depth_key, unsigned long long* __restrict__ out) {
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int2 r = ranges[tile];
        for (int i = r.x + threadIdx.x; i < r.y; i += blockDim.x)
            out[i] = ((unsigned long long)(unsigned)tile << 32) | (unsigned long long)depth_key[gid[i]];
    }
}

int sm_count() { return sm_count_current(); }

}  // namespace

#ifdef GSR_DUP_TRACE
extern "C" __attribute__((visibility("default"))) int gsr_debug_dup_trace(unsigned long long* out, long long n_words) {
    if (n_words > 2048 * 8) n_words = 2048 * 8;
    return (int)cudaMemcpyFromSymbol(out, g_dup_trace, (size_t)n_words * 8);
}
#endif


void launch_duplicate_sorted(const FrameParams* fp, const uint32_t* sorted_gid, SplatBuffers sp, FrameCounters* counters,
                             uint32_t* chain_status, uint32_t* chain_ticket, uint32_t* out_tile, uint32_t* out_gid,
                             uint4* big_list, uint32_t* big_count, long long n, long long capacity, cudaStream_t s) {
    if (n <= 0) return;
    const long long chunks = (n + DUP_CHUNK - 1) / DUP_CHUNK;
    long long grid = chunks < (long long)sm_count() * 6 ? chunks : (long long)sm_count() * 6;
    launch_pdl_m(2, k_duplicate<false>, dim3((unsigned)grid), dim3(DUP_THREADS), 0, s, fp, sorted_gid, sp, counters,
               reinterpret_cast<unsigned long long*>(chain_status), chain_ticket, (void*)out_tile, out_gid, (unsigned)n,
               (unsigned long long)capacity, big_list, (uint32_t*)big_count);
    launch_pdl_m(2, k_duplicate_big<false>, dim3((unsigned)(sm_count() * 4)), dim3(DUP_THREADS), 0, s, fp, sp,
               (const uint4*)big_list, (const uint32_t*)big_count, (void*)out_tile, out_gid, (unsigned long long)capacity);
}

void launch_duplicate_keys64(const FrameParams* fp, SplatBuffers sp, FrameCounters* counters, uint32_t* chain_status,
                             uint32_t* chain_ticket, uint64_t* out_keys, uint32_t* out_gid, uint4* big_list,
                             uint32_t* big_count, long long n, long long capacity, cudaStream_t s) {
    if (n <= 0) return;
    const long long chunks = (n + DUP_CHUNK - 1) / DUP_CHUNK;
    long long grid = chunks < (long long)sm_count() * 6 ? chunks : (long long)sm_count() * 6;
    k_duplicate<true><<<(unsigned)grid, DUP_THREADS, 0, s>>>(fp, nullptr, sp, counters,
                                                            reinterpret_cast<unsigned long long*>(chain_status),
                                                            chain_ticket, out_keys, out_gid, (unsigned)n,
                                                            (unsigned long long)capacity, big_list, big_count);
    k_duplicate_big<true><<<sm_count() * 4, DUP_THREADS, 0, s>>>(fp, sp, big_list, big_count, out_keys, out_gid,
                                                                (unsigned long long)capacity);
}

template <typename KeyT>
void launch_tile_ranges(const KeyT* sorted_keys, int key_shift, const FrameCounters* counters, long long capacity,
                        int2* ranges, int n_tiles, cudaStream_t s) {
    cudaMemsetAsync(ranges, 0, sizeof(int2) * (size_t)n_tiles, s);
    k_tile_ranges<KeyT><<<sm_count() * 8, 256, 0, s>>>(sorted_keys, key_shift, counters, capacity, ranges);
}
template void launch_tile_ranges<uint32_t>(const uint32_t*, int, const FrameCounters*, long long, int2*, int, cudaStream_t);
template void launch_tile_ranges<uint64_t>(const uint64_t*, int, const FrameCounters*, long long, int2*, int, cudaStream_t);

void launch_keys64_from_ranges(const int2* ranges, int n_tiles, const uint32_t* gid, const uint32_t* depth_key,
                               uint64_t* out, cudaStream_t s) {
    k_keys64_from_ranges<<<sm_count() * 8, 256, 0, s>>>(ranges, n_tiles, gid, depth_key,
                                                        reinterpret_cast<unsigned long long*>(out));
}

// Returns the dynamic shared memory the scan wants (0 = works in place in global memory).
static size_t tile_scan_smem(int gw, int TY) {
    const size_t need = (size_t)gw * TY * sizeof(int);
    return need <= 200 * 1024 ? need : 0;
}

void launch_tile_scan(int* grid, int grid_stride, int gw, int tiles_x, int tiles_y, int bits0, long long capacity, int2* ranges,
                      uint32_t* hist, FrameCounters* counters, cudaStream_t s) {
    const size_t smem = tile_scan_smem(gw, tiles_y);
    static bool configured[64] = {};  // per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (smem > 48 * 1024 && dev >= 0 && dev < 64 && !configured[dev]) {
        cudaFuncSetAttribute(k_tile_scan, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(200 * 1024));
        configured[dev] = true;
    }
    k_tile_scan<<<1, TS_THREADS, smem, s>>>(grid, grid_stride, gw, tiles_x, tiles_y, bits0, smem ? 1 : 0,
                                             (unsigned long long)capacity, ranges, hist, counters);
}
