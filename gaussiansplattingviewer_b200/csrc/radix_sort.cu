// K3 — hand-written onesweep LSB radix sort (key/value pairs) for sm_100a.
//
// Replaces the torch / cupy / numpy argsort of /root/reference/renderer_ogl.py:10-53 and the CUB
// DeviceRadixSort inside diff_gaussian_rasterization.  Stable, 8-bit digits, 32- or 64-bit keys,
// 32-bit values.
//
//   k_rs_hist : ONE read of the keys builds the digit histograms of ALL passes (shared-memory
//               histograms, one global atomic per bin per CTA) and clears the first look-back buffer.
//   k_rs_pass : one kernel per digit.  Persistent CTAs take tiles in ticket order; each tile is ranked
//               with warp match_any + per-warp digit counters, its per-digit counts are chained to
//               the previous tiles with decoupled look-back (single 32-bit status word per
//               (tile, digit): 2 flag bits + 30 count bits), and keys/values are scattered through
//               shared memory so global stores are contiguous per digit run.  No global memory
//               round trip between "count" and "scatter": keys are read once and written once per
//               pass (the onesweep property).
//
// Item counts may live on the device (n_dev): grids are sized for the SM count, not for n, so the
// same launch (and the same captured CUDA graph) serves any n up to the buffer capacity.
#include "gsr_internal.h"

namespace {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr uint32_t RS_FLAG_AGG = 1u << 30;
constexpr uint32_t RS_FLAG_INC = 2u << 30;
constexpr uint32_t RS_FLAG_MASK = 3u << 30;
constexpr uint32_t RS_VAL_MASK = (1u << 30) - 1u;

// items per thread of the scatter pass: LARGE is the default, SMALL trades tile size for occupancy
// (48 registers, 5 CTAs per SM); gsr_tune(GSR_TUNE_SORT_ITEMS) switches between them
template <typename KeyT> struct RsCfg;
template <> struct RsCfg<uint32_t> { static constexpr int LARGE = 16, SMALL = 8; };
template <> struct RsCfg<uint64_t> { static constexpr int LARGE = 12, SMALL = 6; };
int g_rs_small = 0;
int g_rs_prefetch = 1;
int g_rs_balance = 1;  // k_rs_pass2: 1 = equal spans per CTA (gsr_tune GSR_TUNE_SORT_BALANCE)
int g_rs_mix = 0;  // k_rs_pass2: 1 = odd items are matched by MATCH.ANY (ADU pipe) instead of ballots (measured slower)

struct RsPasses {
    int n;
    unsigned direct_mask;  // passes whose digit is counted without the warp-level match (k_rs_hist)
    int shift[8];
    int bits[8];
};

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Lanes of the warp whose digit equals this lane's.  Built from one ballot per digit bit: the hardware
// MATCH.ANY instruction runs on the address-divergence unit at a fraction of ALU rate (ncu: pipe_adu 85 %
// in the histogram kernel, 72 % in the scatter pass), VOTE + LOP3 run at full rate.
__device__ __forceinline__ uint32_t match_digit(uint32_t d, int /*bits*/) {
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        // all 8 bits always: digits are masked to `bits` bits, so the upper ballots agree on every lane and
        // cost less than predicating each step on a run-time bit count.  Four instructions per bit:
        // LOP3 -> predicate, VOTE, and one of two predicated LOP3s.
        asm("{\n\t"
            ".reg .pred p;\n\t"
            ".reg .b32 t, m;\n\t"
            "and.b32 t, %1, %2;\n\t"
            "setp.ne.b32 p, t, 0;\n\t"
            "vote.sync.ballot.b32 m, p, 0xffffffff;\n\t"
            "@p and.b32 %0, %0, m;\n\t"
            "@!p lop3.b32 %0, %0, m, 0, 0x30;\n\t"
            "}"
            : "+r"(peers)
            : "r"(d), "r"(1u << b));
    }
    return peers;
}

// Decoupled look-back for digit t of tile `tile`: sum of the counts of all earlier tiles.  Walks back over
// the per-tile status words until one carries an inclusive prefix.  The walk must outrun the rate at which
// new tiles are started (hundreds of CTAs are in flight and none of the tiles right behind this one has an
// inclusive prefix yet): a walk of one dependent L2 read per step left 40 % of the pass waiting here
// (clock64 phase timing), so WIN status words are fetched per round trip - 4 first (the usual case ends
// there), then 16 at a time.
template <int WIN>
__device__ __forceinline__ bool look_back_round(const uint32_t* status, long long& p, int t, uint32_t& prefix) {
    uint32_t w[WIN];
#pragma unroll
    for (int k = 0; k < WIN; ++k) w[k] = (p - k >= 0) ? ld_volatile_u32(&status[(p - k) * 256 + t]) : RS_FLAG_INC;
    bool done = false;
    int used = 0;
#pragma unroll
    for (int k = 0; k < WIN; ++k) {
        if (!done && used == k && (w[k] & RS_FLAG_MASK) != 0u) {
            prefix += w[k] & RS_VAL_MASK;
            used = k + 1;
            if (w[k] & RS_FLAG_INC) done = true;
        }
    }
    p -= used;
    // nothing new since the last probe: back off instead of spinning - the CTAs sharing this SM are
    // ranking their tiles and need the issue slots (ALU pipe) this loop would burn
    if (!done && used == 0) __nanosleep(256);
    return done;
}
__device__ __forceinline__ uint32_t look_back(const uint32_t* status, long long tile, int t) {
    uint32_t prefix = 0;
    long long p = tile - 1;
    if (look_back_round<4>(status, p, t, prefix)) return prefix;
    while (!look_back_round<16>(status, p, t, prefix)) {}
    return prefix;
}

// exclusive scan of one value per thread over a 256-thread block; returns the exclusive prefix and the
// block total via *total.  s_tmp: RS_WARPS + 1 words.
__device__ __forceinline__ uint32_t block_excl_scan_256(uint32_t v, uint32_t* s_tmp, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_tmp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = (lane < RS_WARPS) ? s_tmp[lane] : 0u;
        uint32_t wi = w;
#pragma unroll
        for (int o = 1; o < RS_WARPS; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        if (lane < RS_WARPS) s_tmp[lane] = wi - w;
        if (lane == RS_WARPS - 1) s_tmp[RS_WARPS] = wi;
    }
    __syncthreads();
    const uint32_t res = s_tmp[warp] + incl - v;
    if (total) *total = s_tmp[RS_WARPS];
    __syncthreads();  // s_tmp may be reused by the caller
    return res;
}

template <typename KeyT>
__global__ void __launch_bounds__(RS_THREADS) k_rs_hist(const KeyT* __restrict__ keys, long long n_host,
                                                       const uint32_t* __restrict__ n_dev, long long capacity,
                                                       RsPasses passes, uint32_t* __restrict__ hist,
                                                       uint32_t* __restrict__ status0, int tile_items) {
    __shared__ uint32_t s_hist[8 * 256];
    long long n = n_dev ? (long long)(*n_dev) : n_host;
    if (n > capacity) n = capacity;
    for (int i = threadIdx.x; i < passes.n * 256; i += RS_THREADS) s_hist[i] = 0;
    __syncthreads();
    const long long stride = (long long)gridDim.x * RS_THREADS;
    // 4 keys per thread per step (one 16- or 32-byte load), 2 steps in flight: the loop is otherwise
    // latency bound (one dependent 4-byte load per iteration kept < 20 % of HBM bandwidth busy)
    constexpr int VEC = 4;
    const long long n_vec = (reinterpret_cast<uintptr_t>(keys) % (sizeof(KeyT) * VEC)) ? 0 : n / VEC;
    const int lane = threadIdx.x & 31;
    // Two ways to count a digit.  Shared-memory atomics retire about one lane per cycle per SM, so a digit
    // whose values repeat inside a warp (exponent bytes of depth keys, high tile bits) is first aggregated
    // with a ballot match (32 ALU instructions, then one atomic per distinct value); a digit that is spread
    // over its bins (mantissa bytes) gains nothing from the match and goes straight to the atomic.  The
    // caller says which passes are which (passes.direct_mask); padding keys (all ones: culled Gaussians)
    // are always aggregated.
    auto count_key = [&](KeyT k, bool valid) {
        const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
        const bool ones = valid && (k == (KeyT) ~(KeyT)0);
        const uint32_t omask = __ballot_sync(0xffffffffu, ones);
#pragma unroll
        for (int p = 0; p < 8; ++p) {
            if (p < passes.n) {
                const uint32_t d = (uint32_t)(k >> passes.shift[p]) & ((1u << passes.bits[p]) - 1u);
                if ((passes.direct_mask >> p) & 1) {
                    if (valid && !ones) atomicAdd(&s_hist[p * 256 + d], 1u);
                    if (omask && lane == (__ffs(omask) - 1)) atomicAdd(&s_hist[p * 256 + d], __popc(omask));
                } else {
                    const uint32_t peers = match_digit(d, passes.bits[p]) & vmask;
                    if (valid && lane == (__ffs(peers) - 1)) atomicAdd(&s_hist[p * 256 + d], __popc(peers));
                }
            }
        }
    };
    struct alignas(sizeof(KeyT) * VEC) KeyVec { KeyT k[VEC]; };
    const KeyVec* keys_v = reinterpret_cast<const KeyVec*>(keys);
    // warps stay converged: the trip count is rounded up to a multiple of the warp width
    const long long first = (long long)blockIdx.x * RS_THREADS + threadIdx.x;
    for (long long i = first - lane; i < n_vec; i += stride) {
        const long long iv = i + lane;
        const bool valid = iv < n_vec;
        KeyVec v;
        if (valid) v = keys_v[iv];
        else { for (int j = 0; j < VEC; ++j) v.k[j] = 0; }
#pragma unroll
        for (int j = 0; j < VEC; ++j) count_key(v.k[j], valid);
    }
    // scalar remainder (n % VEC keys, or everything when the key pointer is not vector aligned)
    for (long long i = n_vec * VEC + first - lane; i < n; i += stride) {
        const long long it = i + lane;
        const bool valid = it < n;
        count_key(valid ? keys[it] : (KeyT)0, valid);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes.n * 256; i += RS_THREADS) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&hist[i], c);
    }
    // clear the look-back rows the first pass will use
    const long long rows = (n + tile_items - 1) / tile_items;
    for (long long i = (long long)blockIdx.x * RS_THREADS + threadIdx.x; i < rows * 256; i += stride) status0[i] = 0;
}

template <typename KeyT, int ITEMS, bool IDENTITY_VALS>
__global__ void __launch_bounds__(RS_THREADS, (ITEMS <= 8 ? 5 : 3)) k_rs_pass(const KeyT* __restrict__ keys_in, KeyT* __restrict__ keys_out,
                                                       const uint32_t* __restrict__ vals_in,
                                                       uint32_t* __restrict__ vals_out, long long n_host,
                                                       const uint32_t* __restrict__ n_dev, long long capacity,
                                                       int shift, int bits, const uint32_t* __restrict__ hist,
                                                       uint32_t* __restrict__ ticket, uint32_t* __restrict__ status_cur,
                                                       uint32_t* __restrict__ status_next) {
    constexpr int TILE = RS_THREADS * ITEMS;
    __shared__ KeyT s_keys[TILE];
    __shared__ uint32_t s_vals[TILE];
    __shared__ uint32_t s_whist[RS_WARPS][256];
    __shared__ uint32_t s_digit_start[256];
    __shared__ uint32_t s_bin_base[256];
    __shared__ uint32_t s_global_excl[256];
    __shared__ uint32_t s_tmp[RS_WARPS + 1];
    __shared__ uint32_t s_tile;

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t mask = (1u << bits) - 1u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    long long n = n_dev ? (long long)(*n_dev) : n_host;
    if (n > capacity) n = capacity;

    s_global_excl[t] = block_excl_scan_256(hist[t], s_tmp, nullptr);

    while (true) {
        if (t == 0) s_tile = atomicAdd(ticket, 1u);
        __syncthreads();
        const long long tile = s_tile;
        const long long tile_base = tile * TILE;
        if (tile_base >= n) break;
        const int n_valid = (int)((n - tile_base < TILE) ? (n - tile_base) : TILE);
        status_next[tile * 256 + t] = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) s_whist[w][t] = 0;
        __syncthreads();

        // ---- load (warp-striped: item i of lane l is element warp_base + i*32 + l)
        KeyT key[ITEMS];
        uint32_t val[ITEMS];
        const long long warp_base = tile_base + (long long)warp * (32 * ITEMS);
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const long long idx = warp_base + i * 32 + lane;
            const bool valid = idx < n;
            key[i] = valid ? keys_in[idx] : (KeyT)~(KeyT)0;
            if (IDENTITY_VALS) val[i] = (uint32_t)idx;
            else val[i] = valid ? vals_in[idx] : 0u;
        }
        // ---- rank inside the warp, digit counters per warp
        uint32_t rank[ITEMS];
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & mask;
            const uint32_t peers = match_digit(d, bits);
            const int leader = __ffs(peers) - 1;
            uint32_t old = 0;
            if (lane == leader) {
                old = s_whist[warp][d];
                s_whist[warp][d] = old + __popc(peers);
            }
            old = __shfl_sync(0xffffffffu, old, leader);
            rank[i] = old + __popc(peers & lt_mask);
            __syncwarp();
        }
        __syncthreads();
        // ---- per digit (thread t owns digit t): exclusive scan over warps, tile count
        uint32_t count = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) {
            const uint32_t c = s_whist[w][t];
            s_whist[w][t] = count;
            count += c;
        }
        // padding keys (all ones) sit in digit `mask`, after every valid key of that digit
        const uint32_t count_pub = count - (((uint32_t)t == mask) ? (uint32_t)(TILE - n_valid) : 0u);
        if ((uint32_t)t <= mask)  // digits above the mask do not exist in this pass
            st_volatile_u32(&status_cur[tile * 256 + t], ((tile == 0) ? RS_FLAG_INC : RS_FLAG_AGG) | count_pub);
        const uint32_t dstart = block_excl_scan_256(count, s_tmp, nullptr);
        s_digit_start[t] = dstart;
        // ---- decoupled look-back over previous tiles for digit t
        uint32_t prefix = 0;
        if (tile > 0 && (uint32_t)t <= mask) {
            prefix = look_back(status_cur, tile, t);
            st_volatile_u32(&status_cur[tile * 256 + t], RS_FLAG_INC | (prefix + count_pub));
        }
        s_bin_base[t] = s_global_excl[t] + prefix - dstart;
        __syncthreads();
        // ---- scatter into shared memory in tile-sorted order
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & mask;
            const uint32_t pos = s_digit_start[d] + s_whist[warp][d] + rank[i];
            s_keys[pos] = key[i];
            s_vals[pos] = val[i];
        }
        __syncthreads();
        // ---- contiguous runs per digit go out to global memory
        for (int s = t; s < n_valid; s += RS_THREADS) {
            const KeyT k = s_keys[s];
            const uint32_t d = (uint32_t)(k >> shift) & mask;
            const uint32_t dst = s_bin_base[d] + (uint32_t)s;
            if (keys_out) keys_out[dst] = k;  // nullptr: the caller only needs the permuted values (last pass)
            vals_out[dst] = s_vals[s];
        }
        __syncthreads();
    }
}

// ---- k_rs_pass2: the same pass over SUPER-TILES of S sub-tiles ------------------------------------
// clock64 phase timing of k_rs_pass showed 40 % of a pass waiting in the look-back: the persistent CTAs
// run in lock step, so in every wave the inclusive prefix has to cross all ~444 tiles in flight before the
// last of them can write, and that chain repeats every wave.  Its total length is (tiles) / sqrt(tiles
// in flight), so the cure is fewer, larger tiles without fewer CTAs: a CTA ranks S sub-tiles one after
// the other, parks each one tile-sorted in shared memory (the registers are free again for the next),
// publishes ONE count row for all of them and does ONE look-back.  S = 2 uses the shared memory three
// CTAs per SM can have (2 x 32 KB staging each) and halves the chain.
template <int ITEMS, int S, bool IDENTITY_VALS>
__global__ void __launch_bounds__(RS_THREADS, 3) k_rs_pass2(const uint32_t* __restrict__ keys_in, uint32_t* __restrict__ keys_out,
                                                          const uint32_t* __restrict__ vals_in,
                                                          uint32_t* __restrict__ vals_out, long long n_host,
                                                          const uint32_t* __restrict__ n_dev, long long capacity,
                                                          int shift, int bits, const uint32_t* __restrict__ hist,
                                                          uint32_t* __restrict__ ticket, uint32_t* __restrict__ status_cur,
                                                          uint32_t* __restrict__ status_next, int mix, int balance) {
    constexpr int TILE = RS_THREADS * ITEMS;  // one sub-tile
    extern __shared__ __align__(16) unsigned char s_dyn[];
    uint32_t* s_keys = reinterpret_cast<uint32_t*>(s_dyn);  // [S][TILE] tile-sorted keys
    uint32_t* s_vals = s_keys + S * TILE;                   // [S][TILE]
    __shared__ uint16_t s_whist[RS_WARPS][256];  // per-warp digit counts (<= 32 * ITEMS each)
    __shared__ uint32_t s_digit_start[256];
    __shared__ uint32_t s_bin_base[S][256];
    __shared__ uint32_t s_global_excl[256];
    __shared__ uint32_t s_tmp[RS_WARPS + 1];
    __shared__ uint32_t s_tile;

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t mask = (1u << bits) - 1u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    pdl_wait();  // resident behind the previous pass; nothing of global memory is touched before this
    long long n = n_dev ? (long long)(*n_dev) : n_host;
    if (n > capacity) n = capacity;

    s_global_excl[t] = block_excl_scan_256(hist[t], s_tmp, nullptr);

    // Balanced spans: a super-tile covers `span` <= S * TILE items, chosen so that every CTA of the grid gets the
    // same number of equally long super-tiles.  With the full 8192-item span 3 M keys are 367 super-tiles on 444
    // CTA slots - SMs holding three of them rank 1.5x longer than SMs holding two, and in a single-wave pass
    // everybody's look-back waits for the slowest count row.
    long long span = (long long)S * TILE;
    if (balance) {
        const long long G = gridDim.x;
        const long long waves = (n + G * span - 1) / (G * span);
        long long sp = (n + G * waves - 1) / (G * (waves > 0 ? waves : 1));
        sp = (sp + 31) & ~31ll;  // 128-byte aligned warp loads
        span = sp < TILE ? TILE : (sp > span ? span : sp);  // >= TILE: the cleared look-back rows are n / TILE
    }

    while (true) {
        if (t == 0) s_tile = atomicAdd(ticket, 1u);
        __syncthreads();
        const long long tile = s_tile;  // super-tile index
        const long long super_base = tile * span;
        if (super_base >= n) break;
        const long long super_end = (super_base + span < n) ? super_base + span : n;
        status_next[tile * 256 + t] = 0;
        uint32_t cnt[S], dst[S];  // digit t of sub-tile s: valid count, start inside the sorted sub-tile
        int n_valid[S];
#pragma unroll
        for (int sub = 0; sub < S; ++sub) {
            const long long tile_base = super_base + (long long)sub * TILE;
            n_valid[sub] = (int)((super_end - tile_base < TILE) ? ((super_end - tile_base > 0) ? (super_end - tile_base) : 0) : TILE);
            cnt[sub] = 0; dst[sub] = 0;
            if (n_valid[sub] == 0) continue;  // block-uniform
#pragma unroll
            for (int w = 0; w < RS_WARPS; ++w) s_whist[w][t] = 0;
            __syncthreads();
            // ---- load (warp-striped: item i of lane l is element warp_base + i*32 + l)
            uint32_t key[ITEMS], val[ITEMS];
            const long long warp_base = tile_base + (long long)warp * (32 * ITEMS);
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) {
                const long long idx = warp_base + i * 32 + lane;
                const bool valid = idx < super_end;
                key[i] = valid ? keys_in[idx] : 0xFFFFFFFFu;
                if (IDENTITY_VALS) val[i] = (uint32_t)idx;
                else val[i] = valid ? vals_in[idx] : 0u;
            }
            // ---- rank inside the warp, digit counters per warp.  All matches first: they are independent of the
            // counters, so their ballots overlap the serial counter updates below (0.232 -> 0.226 ms per tile
            // sort).  `mix` sends odd items through the hardware MATCH.ANY instead, to use the otherwise idle
            // address-divergence unit next to the ALU pipe: measured SLOWER (0.240 ms), the unit is shared by the SM.
            uint32_t rank[ITEMS];  // first the peer mask, then the rank
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) {
                const uint32_t d = (key[i] >> shift) & mask;
                rank[i] = (mix && (i & 1)) ? __match_any_sync(0xffffffffu, d) : match_digit(d, bits);
            }
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) {
                const uint32_t d = (key[i] >> shift) & mask;
                const uint32_t peers = rank[i];
                const int leader = __ffs(peers) - 1;
                uint32_t old = 0;
                if (lane == leader) {
                    old = s_whist[warp][d];
                    s_whist[warp][d] = (uint16_t)(old + __popc(peers));
                }
                old = __shfl_sync(0xffffffffu, old, leader);
                rank[i] = old + __popc(peers & lt_mask);
                __syncwarp();
            }
            __syncthreads();
            // ---- per digit (thread t owns digit t): exclusive scan over warps, sub-tile count
            uint32_t count = 0;
#pragma unroll
            for (int w = 0; w < RS_WARPS; ++w) {
                const uint32_t c = s_whist[w][t];
                s_whist[w][t] = (uint16_t)count;
                count += c;
            }
            // padding keys (all ones) sit in digit `mask`, after every valid key of that digit
            cnt[sub] = count - (((uint32_t)t == mask) ? (uint32_t)(TILE - n_valid[sub]) : 0u);
            const uint32_t dstart = block_excl_scan_256(count, s_tmp, nullptr);
            dst[sub] = dstart;
            s_digit_start[t] = dstart;
            __syncthreads();
            // ---- park the sub-tile in shared memory in tile-sorted order
            uint32_t* sk = s_keys + sub * TILE;
            uint32_t* sv = s_vals + sub * TILE;
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) {
                const uint32_t d = (key[i] >> shift) & mask;
                const uint32_t pos = s_digit_start[d] + s_whist[warp][d] + rank[i];
                sk[pos] = key[i];
                sv[pos] = val[i];
            }
            __syncthreads();  // s_whist / s_digit_start are rewritten by the next sub-tile
        }
        // ---- one count row and one look-back for the whole super-tile
        uint32_t total = 0;
#pragma unroll
        for (int sub = 0; sub < S; ++sub) total += cnt[sub];
        uint32_t prefix = 0;
        if ((uint32_t)t <= mask) {  // digits above the mask do not exist in this pass
            if (tile == 0) st_volatile_u32(&status_cur[t], RS_FLAG_INC | total);
            else {
                st_volatile_u32(&status_cur[tile * 256 + t], RS_FLAG_AGG | total);
                prefix = look_back(status_cur, tile, t);
                st_volatile_u32(&status_cur[tile * 256 + t], RS_FLAG_INC | (prefix + total));
            }
        }
        uint32_t run = s_global_excl[t] + prefix;
#pragma unroll
        for (int sub = 0; sub < S; ++sub) {
            s_bin_base[sub][t] = run - dst[sub];
            run += cnt[sub];
        }
        __syncthreads();
        // ---- contiguous runs per digit go out to global memory
#pragma unroll
        for (int sub = 0; sub < S; ++sub) {
            const uint32_t* sk = s_keys + sub * TILE;
            const uint32_t* sv = s_vals + sub * TILE;
            for (int s = t; s < n_valid[sub]; s += RS_THREADS) {
                const uint32_t k = sk[s];
                const uint32_t d = (k >> shift) & mask;
                const uint32_t o = s_bin_base[sub][d] + (uint32_t)s;
                if (keys_out) keys_out[o] = k;
                vals_out[o] = sv[s];
            }
        }
        __syncthreads();  // staging and s_bin_base are rewritten by the next super-tile
    }
}

int g_num_sms = 0;
int num_sms() {
    if (g_num_sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
        if (g_num_sms <= 0) g_num_sms = 148;
    }
    return g_num_sms;
}

}  // namespace

void set_sort_small_tiles(int v) { g_rs_small = v ? 1 : 0; }
void set_sort_prefetch(int v) { g_rs_prefetch = v ? 1 : 0; }
void set_sort_mix(int v) { g_rs_mix = v ? 1 : 0; }
void set_sort_balance(int v) { g_rs_balance = v ? 1 : 0; }

size_t sort_scratch_bytes(long long max_items, int key_bytes) {
    const int items = (key_bytes == 8) ? RsCfg<uint64_t>::SMALL : RsCfg<uint32_t>::SMALL;
    const long long tiles = (max_items + RS_THREADS * items - 1) / (RS_THREADS * items) + 1;
    return (size_t)(8 * 256 + 64) * 4 + 2 * (size_t)tiles * 256 * 4;
}

void sort_scratch_carve(SortScratch* s, void* base, long long max_items, int key_bytes) {
    const int items = (key_bytes == 8) ? RsCfg<uint64_t>::SMALL : RsCfg<uint32_t>::SMALL;
    const long long tiles = (max_items + RS_THREADS * items - 1) / (RS_THREADS * items) + 1;
    uint32_t* p = reinterpret_cast<uint32_t*>(base);
    s->hist = p;
    s->tickets = p + 8 * 256;
    s->zero_bytes = (size_t)(8 * 256 + 64) * 4;
    s->status[0] = p + 8 * 256 + 64;
    s->status[1] = s->status[0] + (size_t)tiles * 256;
    s->max_tiles = tiles;
}

template <typename KeyT, int ITEMS>
void launch_pass(bool identity, unsigned grid, cudaStream_t stream, const KeyT* kin, KeyT* kout, const uint32_t* vin,
                 uint32_t* vout, long long n_host, const uint32_t* n_dev, long long capacity, int shift, int bits,
                 const uint32_t* hist, uint32_t* ticket, uint32_t* st_cur, uint32_t* st_next) {
    if (identity)
        k_rs_pass<KeyT, ITEMS, true><<<grid, RS_THREADS, 0, stream>>>(kin, kout, nullptr, vout, n_host, n_dev, capacity, shift,
                                                                     bits, hist, ticket, st_cur, st_next);
    else
        k_rs_pass<KeyT, ITEMS, false><<<grid, RS_THREADS, 0, stream>>>(kin, kout, vin, vout, n_host, n_dev, capacity, shift,
                                                                      bits, hist, ticket, st_cur, st_next);
}

// super-tile variant (32-bit keys): dynamic shared memory = S staged sub-tiles of keys + values
constexpr int RS2_ITEMS = 16, RS2_S = 2;
constexpr size_t RS2_SMEM = (size_t)RS2_S * RS_THREADS * RS2_ITEMS * (sizeof(uint32_t) + sizeof(uint32_t));
int rs2_ctas_per_sm() {  // 0 = the variant cannot run here
    static int n = -1;
    if (n < 0) {
        n = 0;
        bool ok = true;
        auto prep = [&](const void* f) {
            ok = ok && cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RS2_SMEM) == cudaSuccess;
            ok = ok && cudaFuncSetAttribute(f, cudaFuncAttributePreferredSharedMemoryCarveout, 100) == cudaSuccess;
        };
        prep((const void*)k_rs_pass2<RS2_ITEMS, RS2_S, true>);
        prep((const void*)k_rs_pass2<RS2_ITEMS, RS2_S, false>);
        int nb = 0;
        if (ok && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_rs_pass2<RS2_ITEMS, RS2_S, false>, RS_THREADS,
                                                                RS2_SMEM) == cudaSuccess)
            n = nb;
        cudaGetLastError();
    }
    return n;
}

template <typename KeyT>
void sort_scratch_clear(SortScratch* scratch, long long upper, bool hist_ready, cudaStream_t stream) {
    if (upper <= 0) return;
    if (hist_ready) {
        // the digit histograms come from the caller (k_tile_scan): clear the tickets and the look-back rows of
        // the first pass only
        const int TILE = RS_THREADS * (g_rs_small ? RsCfg<KeyT>::SMALL : RsCfg<KeyT>::LARGE);
        const long long tiles_upper = (upper + TILE - 1) / TILE;
        cudaMemsetAsync(scratch->tickets, 0, 64 * sizeof(uint32_t), stream);
        cudaMemsetAsync(scratch->status[0], 0, (size_t)tiles_upper * 256 * sizeof(uint32_t), stream);
    } else {
        cudaMemsetAsync(scratch->hist, 0, scratch->zero_bytes, stream);  // histograms + tickets; k_rs_hist clears the rows
    }
}
template void sort_scratch_clear<uint32_t>(SortScratch*, long long, bool, cudaStream_t);
template void sort_scratch_clear<uint64_t>(SortScratch*, long long, bool, cudaStream_t);

template <typename KeyT>
int radix_sort_pairs(const KeyT* keys_first, KeyT* keys, KeyT* keys_tmp, uint32_t* vals, uint32_t* vals_tmp, bool vals_identity,
                     long long n_host, const uint32_t* n_dev, long long capacity, int begin_bit, int end_bit,
                     SortScratch* scratch, cudaStream_t stream, long long* launch_count, const SortOptions* opt) {
    const int first_bits = (opt && opt->first_bits > 0) ? opt->first_bits : 8;
    const bool hist_ready = opt && opt->hist_ready;
    const bool vals_only_last = opt && opt->vals_only_last;
    RsPasses passes;
    passes.n = 0;
    passes.direct_mask = opt ? opt->hist_direct_mask : 0u;
    for (int b = begin_bit; b < end_bit && passes.n < 8;) {
        const int want = (passes.n == 0) ? first_bits : 8;
        passes.shift[passes.n] = b;
        passes.bits[passes.n] = (end_bit - b < want) ? (end_bit - b) : want;
        b += passes.bits[passes.n];
        ++passes.n;
    }
    for (int p = passes.n; p < 8; ++p) { passes.shift[p] = 0; passes.bits[p] = 1; }
    if (passes.n == 0) return 0;
    const long long upper = n_dev ? capacity : n_host;
    if (upper <= 0) return 0;
    const bool small = g_rs_small != 0;
    const int TILE = RS_THREADS * (small ? RsCfg<KeyT>::SMALL : RsCfg<KeyT>::LARGE);
    const long long tiles_upper = (upper + TILE - 1) / TILE;
    const int sms = num_sms();
    if (!(opt && opt->scratch_cleared)) sort_scratch_clear<KeyT>(scratch, upper, hist_ready, stream);
    if (!hist_ready) {
        long long g = (upper + RS_THREADS * 8 - 1) / (RS_THREADS * 8);
        if (g > sms * 8) g = sms * 8;
        if (g < 1) g = 1;
        k_rs_hist<KeyT><<<(unsigned)g, RS_THREADS, 0, stream>>>(keys_first, n_host, n_dev, capacity, passes, scratch->hist,
                                                               scratch->status[0], TILE);
        if (launch_count) ++*launch_count;
    }
    // super-tile variant: 32-bit keys, default tile size
    int pf_ctas = 0;
    if (sizeof(KeyT) == 4 && !small && g_rs_prefetch) pf_ctas = rs2_ctas_per_sm();
    const int per_sm = pf_ctas > 0 ? pf_ctas : (small ? 5 : 3);  // resident CTAs per SM: persistent CTAs, one wave
    // balanced spans are never shorter than one sub-tile, so up to tiles_upper CTAs can have work
    const long long work_tiles = pf_ctas > 0 ? (g_rs_balance ? tiles_upper : (tiles_upper + RS2_S - 1) / RS2_S) : tiles_upper;
    long long grid = work_tiles < (long long)sms * per_sm ? work_tiles : (long long)sms * per_sm;
    if (grid < 1) grid = 1;
    const KeyT* kin = keys_first; KeyT* kout = keys_tmp; KeyT* kother = keys;
    uint32_t* vin = vals; uint32_t* vout = vals_tmp;
    for (int p = 0; p < passes.n; ++p) {
        if (vals_only_last && p == passes.n - 1) kout = nullptr;
        uint32_t* st_cur = scratch->status[p & 1];
        uint32_t* st_next = scratch->status[(p + 1) & 1];
        const bool identity = (p == 0 && vals_identity);
        if (pf_ctas > 0) {
            if constexpr (sizeof(KeyT) == 4) {
                if (identity)
                    launch_pdl_m(1, k_rs_pass2<RS2_ITEMS, RS2_S, true>, dim3((unsigned)grid), dim3(RS_THREADS), RS2_SMEM, stream,
                               kin, kout, (const uint32_t*)nullptr, vout, n_host, n_dev, capacity, passes.shift[p], passes.bits[p],
                               (const uint32_t*)(scratch->hist + p * 256), scratch->tickets + p, st_cur, st_next, g_rs_mix, g_rs_balance);
                else
                    launch_pdl_m(1, k_rs_pass2<RS2_ITEMS, RS2_S, false>, dim3((unsigned)grid), dim3(RS_THREADS), RS2_SMEM, stream,
                               kin, kout, (const uint32_t*)vin, vout, n_host, n_dev, capacity, passes.shift[p], passes.bits[p],
                               (const uint32_t*)(scratch->hist + p * 256), scratch->tickets + p, st_cur, st_next, g_rs_mix, g_rs_balance);
            }
        } else if (small)
            launch_pass<KeyT, RsCfg<KeyT>::SMALL>(identity, (unsigned)grid, stream, kin, kout, vin, vout, n_host, n_dev, capacity,
                                                  passes.shift[p], passes.bits[p], scratch->hist + p * 256,
                                                  scratch->tickets + p, st_cur, st_next);
        else
            launch_pass<KeyT, RsCfg<KeyT>::LARGE>(identity, (unsigned)grid, stream, kin, kout, vin, vout, n_host, n_dev, capacity,
                                                  passes.shift[p], passes.bits[p], scratch->hist + p * 256,
                                                  scratch->tickets + p, st_cur, st_next);
        if (launch_count) ++*launch_count;
        kin = kout; { KeyT* tk = kout; kout = kother; kother = tk; }
        uint32_t* tv = vin; vin = vout; vout = tv;
    }
    return passes.n & 1;
}

template int radix_sort_pairs<uint32_t>(const uint32_t*, uint32_t*, uint32_t*, uint32_t*, uint32_t*, bool, long long, const uint32_t*,
                                        long long, int, int, SortScratch*, cudaStream_t, long long*, const SortOptions*);
template int radix_sort_pairs<uint64_t>(const uint64_t*, uint64_t*, uint64_t*, uint32_t*, uint32_t*, bool, long long, const uint32_t*,
                                        long long, int, int, SortScratch*, cudaStream_t, long long*, const SortOptions*);

// ---- allocation-free sort (gsr_sort_pairs_static) ---------------------------------------------------------------
// The radix passes need scratch (digit histograms, tile tickets, two buffers of look-back rows).  Callers that must
// not call cudaMalloc get the same sort with that scratch in a static device array of fixed size: every GPU that
// loads libgsr.so holds RS_STATIC_WORDS * 4 bytes (about 5.6 MB) until the process exits.  All calls on a device share
// the array; the sort clears what must start at zero (histograms + tickets by memset, the first look-back rows by the
// histogram kernel, the next pass's rows by each pass), so stale contents of an earlier call are harmless - but two
// calls that OVERLAP on one GPU (different streams) race on it.
constexpr long long RS_STATIC_MAX_ITEMS = 1ll << 22;  // pairs per call
constexpr size_t RS_STATIC_TILES = (size_t)(RS_STATIC_MAX_ITEMS / (RS_THREADS * 6) + 2);  // 64-bit keys: smallest tiles
constexpr size_t RS_STATIC_WORDS = (8 * 256 + 64) + 2 * RS_STATIC_TILES * 256;
static __device__ __align__(256) uint32_t g_rs_static_scratch[RS_STATIC_WORDS];

long long radix_sort_static_max_items() { return RS_STATIC_MAX_ITEMS; }

// Device address of the static scratch on the CURRENT device.  g_rs_static_scratch is, in host code, only a stand-in
// for the symbol: the address a kernel can use comes from cudaGetSymbolAddress, once per device with that device current.
static uint32_t* rs_static_scratch_ptr(const char** err) {
    static uint32_t* ptr[64] = {};
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { *err = "no current CUDA device"; return nullptr; }
    if (!ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_rs_static_scratch);
        if (e != cudaSuccess || !p) { *err = cudaGetErrorString(e); cudaGetLastError(); return nullptr; }
        ptr[dev] = static_cast<uint32_t*>(p);
    }
    return ptr[dev];
}

// 0 ok (result in keys / vals), -1 bad argument, -2 CUDA error; *err gets a static message.  Allocates nothing.
int radix_sort_pairs_static(int key_bits, void* keys, void* keys_tmp, uint32_t* vals, uint32_t* vals_tmp, long long n,
                            int begin_bit, int end_bit, cudaStream_t stream, const char** err) {
    *err = "";
    if (n > RS_STATIC_MAX_ITEMS) { *err = "n exceeds the static scratch (GSR_SORT_STATIC_MAX_PAIRS = 4194304 pairs)"; return -1; }
    uint32_t* base = rs_static_scratch_ptr(err);
    if (!base) return -2;
    const int kb = key_bits / 8;
    SortScratch ss;
    sort_scratch_carve(&ss, base, RS_STATIC_MAX_ITEMS, kb);
    if ((size_t)((ss.status[1] + (size_t)ss.max_tiles * 256) - base) > RS_STATIC_WORDS) { *err = "static scratch too small (internal)"; return -2; }
    int r;
    if (kb == 4)
        r = radix_sort_pairs<uint32_t>((uint32_t*)keys, (uint32_t*)keys, (uint32_t*)keys_tmp, vals, vals_tmp, false, n, nullptr, n,
                                       begin_bit, end_bit, &ss, stream, nullptr);
    else
        r = radix_sort_pairs<uint64_t>((uint64_t*)keys, (uint64_t*)keys, (uint64_t*)keys_tmp, vals, vals_tmp, false, n, nullptr, n,
                                       begin_bit, end_bit, &ss, stream, nullptr);
    if (r == 1) {  // odd number of passes: bring the result back to keys / vals
        if (cudaMemcpyAsync(keys, keys_tmp, (size_t)n * kb, cudaMemcpyDeviceToDevice, stream) != cudaSuccess ||
            cudaMemcpyAsync(vals, vals_tmp, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream) != cudaSuccess) {
            *err = "copy back failed"; cudaGetLastError(); return -2;
        }
    }
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { *err = cudaGetErrorString(e); return -2; }
    return 0;
}
