// K2 duplicate-with-keys and K4 tile ranges for sm_100a.
//
// K2 turns the per-Gaussian tile rectangles into the tile-instance list.  One pass, load balanced:
//   * each CTA takes a chunk of 1024 Gaussians (in the order given: depth-sorted for the two-level
//     path, index order for the 64-bit-key path), block-scans their tile counts, chains the chunk total
//     to the previous chunks with a decoupled look-back (64-bit status word: 2 flag bits + count), so the
//     prefix sum never makes a round trip through global memory;
//   * instances are then emitted by INSTANCE index, not by Gaussian: thread j of the CTA finds the owner
//     of instance j by binary search in the chunk's shared-memory offsets, so a Gaussian covering
//     thousands of tiles is written by the whole CTA with fully coalesced stores.
// Counterpart of the scan + duplicateWithKeys + identifyTileRanges steps of the tile rasteriser the
// reference's CUDA backend calls (/root/reference/renderer_cuda.py:211-224); the OpenGL backend has no
// tiles (its draw is one instanced quad per Gaussian, renderer_ogl.py:175-179).
#include "gsr_internal.h"

namespace {

constexpr int DUP_THREADS = 256;
constexpr int DUP_ITEMS = 4;
constexpr int DUP_CHUNK = DUP_THREADS * DUP_ITEMS;
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

// KEYS64 = false: gid comes from sorted_gid, output key is the 32-bit tile id.
// KEYS64 = true : gid is the running index, output key is (tile << 32) | depth bits.
template <bool KEYS64>
__global__ void __launch_bounds__(DUP_THREADS) k_duplicate(const FrameParams* __restrict__ fp,
                                                          const uint32_t* __restrict__ sorted_gid, SplatBuffers sp,
                                                          FrameCounters* __restrict__ counters,
                                                          unsigned long long* __restrict__ chain,
                                                          uint32_t* __restrict__ ticket, void* __restrict__ out_keys,
                                                          uint32_t* __restrict__ out_gid, unsigned n,
                                                          unsigned long long capacity) {
    __shared__ uint32_t s_off[DUP_CHUNK];  // chunk-local exclusive instance offsets
    __shared__ uint32_t s_gid[DUP_CHUNK];
    __shared__ uint32_t s_rx[DUP_CHUNK];   // tx0 | (tiles wide) << 16
    __shared__ uint32_t s_ty0[DUP_CHUNK];
    __shared__ uint32_t s_warp[DUP_THREADS / 32 + 1];
    __shared__ uint32_t s_chunk;
    __shared__ unsigned long long s_prefix;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int tiles_x = fp->tiles_x;
    const unsigned n_chunks = (n + DUP_CHUNK - 1) / DUP_CHUNK;

    while (true) {
        if (t == 0) s_chunk = atomicAdd(ticket, 1u);
        __syncthreads();
        const unsigned chunk = s_chunk;
        if (chunk >= n_chunks) break;
        const unsigned base = chunk * DUP_CHUNK + t * DUP_ITEMS;
        uint32_t gid[DUP_ITEMS], cnt[DUP_ITEMS];
#pragma unroll
        for (int i = 0; i < DUP_ITEMS; ++i) {
            const unsigned r = base + i;
            gid[i] = 0; cnt[i] = 0;
            if (r < n) {
                gid[i] = KEYS64 ? r : __ldg(sorted_gid + r);
                cnt[i] = __ldg(sp.tiles + gid[i]);
            }
        }
        uint32_t tsum = 0;
#pragma unroll
        for (int i = 0; i < DUP_ITEMS; ++i) tsum += cnt[i];
        // block exclusive scan of tsum
        uint32_t incl = tsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = (lane < DUP_THREADS / 32) ? s_warp[lane] : 0u;
            uint32_t wi = w;
#pragma unroll
            for (int o = 1; o < DUP_THREADS / 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += v;
            }
            if (lane < DUP_THREADS / 32) s_warp[lane] = wi - w;
            if (lane == DUP_THREADS / 32 - 1) s_warp[DUP_THREADS / 32] = wi;
        }
        __syncthreads();
        uint32_t off = s_warp[warp] + incl - tsum;
        const uint32_t chunk_total = s_warp[DUP_THREADS / 32];
        // chain the chunk total (thread 0), while the others stage their entries
        if (t == 0) {
            unsigned long long prefix = 0;
            if (chunk == 0) {
                st_volatile_u64(&chain[0], CH_FLAG_INC | (unsigned long long)chunk_total);
            } else {
                st_volatile_u64(&chain[chunk], CH_FLAG_AGG | (unsigned long long)chunk_total);
                for (long long p = (long long)chunk - 1;; --p) {
                    unsigned long long w;
                    do { w = ld_volatile_u64(&chain[p]); } while ((w & CH_FLAG_MASK) == 0ull);
                    prefix += w & CH_VAL_MASK;
                    if (w & CH_FLAG_INC) break;
                }
                st_volatile_u64(&chain[chunk], CH_FLAG_INC | (prefix + chunk_total));
            }
            s_prefix = prefix;
            if (chunk == n_chunks - 1) {
                const unsigned long long total = prefix + chunk_total;
                counters->n_instances = (total > 0xFFFFFFFFull) ? 0xFFFFFFFFu : (unsigned)total;
                counters->overflow = (total > capacity) ? 1u : 0u;
            }
        }
#pragma unroll
        for (int i = 0; i < DUP_ITEMS; ++i) {
            const int slot = t * DUP_ITEMS + i;
            s_off[slot] = off;
            off += cnt[i];
            s_gid[slot] = gid[i];
            uint32_t rx = 0, ty0 = 0;
            if (cnt[i]) {
                const float4 r0 = __ldg(sp.rec0 + gid[i]);
                const uint32_t px = __float_as_uint(r0.z), py = __float_as_uint(r0.w);
                const uint32_t imin = px & 0xFFFFu, imax = imin + (px >> 16);
                const uint32_t jmin = py & 0xFFFFu;
                const uint32_t tx0 = imin / GSR_TILE, tx1 = imax / GSR_TILE;
                rx = tx0 | ((tx1 - tx0 + 1) << 16);
                ty0 = jmin / GSR_TILE;
            }
            s_rx[slot] = rx;
            s_ty0[slot] = ty0;
        }
        __syncthreads();
        const unsigned long long prefix = s_prefix;
        // ---- emit by instance index
        for (uint32_t j = t; j < chunk_total; j += DUP_THREADS) {
            // largest slot with s_off[slot] <= j and a non-zero count (zero-count slots share the offset
            // of their successor, so take the LAST slot whose offset <= j)
            int lo = 0, hi = DUP_CHUNK - 1;
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (s_off[mid] <= j) lo = mid; else hi = mid - 1;
            }
            const uint32_t k = j - s_off[lo];
            const uint32_t rx = s_rx[lo];
            const uint32_t wdt = rx >> 16;
            const uint32_t ty = s_ty0[lo] + k / wdt, tx = (rx & 0xFFFFu) + k % wdt;
            const uint32_t tile = ty * (uint32_t)tiles_x + tx;
            const unsigned long long dst = prefix + j;
            if (dst < capacity) {
                const uint32_t g = s_gid[lo];
                if (KEYS64)
                    reinterpret_cast<unsigned long long*>(out_keys)[dst] =
                        ((unsigned long long)tile << 32) | (unsigned long long)__ldg(sp.depth_key + g);
                else
                    reinterpret_cast<uint32_t*>(out_keys)[dst] = tile;
                out_gid[dst] = g;
            }
        }
        __syncthreads();
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

__global__ void k_keys64_from_sorted(const uint32_t* __restrict__ tile_keys, const uint32_t* __restrict__ gid,
                                     const uint32_t* __restrict__ depth_key, const FrameCounters* __restrict__ counters,
                                     long long capacity, unsigned long long* __restrict__ out) {
    long long n = counters->n_instances;
    if (n > capacity) n = capacity;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = ((unsigned long long)tile_keys[i] << 32) | (unsigned long long)depth_key[gid[i]];
}

int sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

}  // namespace

void launch_duplicate_sorted(const FrameParams* fp, const uint32_t* sorted_gid, SplatBuffers sp, FrameCounters* counters,
                             uint32_t* chain_status, uint32_t* chain_ticket, uint32_t* out_tile, uint32_t* out_gid,
                             uint32_t* /*tile_hist*/, long long n, long long capacity, cudaStream_t s) {
    if (n <= 0) return;
    const long long chunks = (n + DUP_CHUNK - 1) / DUP_CHUNK;
    long long grid = chunks < (long long)sm_count() * 6 ? chunks : (long long)sm_count() * 6;
    k_duplicate<false><<<(unsigned)grid, DUP_THREADS, 0, s>>>(fp, sorted_gid, sp, counters,
                                                             reinterpret_cast<unsigned long long*>(chain_status),
                                                             chain_ticket, out_tile, out_gid, (unsigned)n,
                                                             (unsigned long long)capacity);
}

void launch_duplicate_keys64(const FrameParams* fp, SplatBuffers sp, FrameCounters* counters, uint32_t* chain_status,
                             uint32_t* chain_ticket, uint64_t* out_keys, uint32_t* out_gid, long long n,
                             long long capacity, cudaStream_t s) {
    if (n <= 0) return;
    const long long chunks = (n + DUP_CHUNK - 1) / DUP_CHUNK;
    long long grid = chunks < (long long)sm_count() * 6 ? chunks : (long long)sm_count() * 6;
    k_duplicate<true><<<(unsigned)grid, DUP_THREADS, 0, s>>>(fp, nullptr, sp, counters,
                                                            reinterpret_cast<unsigned long long*>(chain_status),
                                                            chain_ticket, out_keys, out_gid, (unsigned)n,
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

void launch_keys64_from_sorted(const uint32_t* tile_keys, const uint32_t* gid, const uint32_t* depth_key,
                               const FrameCounters* counters, long long capacity, uint64_t* out, cudaStream_t s) {
    k_keys64_from_sorted<<<sm_count() * 8, 256, 0, s>>>(tile_keys, gid, depth_key, counters, capacity,
                                                        reinterpret_cast<unsigned long long*>(out));
}
