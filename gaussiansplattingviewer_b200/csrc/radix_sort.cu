// K3 — hand-written onesweep LSB radix sort (key/value pairs) for sm_100a.
//
// Replaces the torch / cupy / numpy argsort of /root/reference/renderer_ogl.py:10-53 and the CUB
// DeviceRadixSort inside diff_gaussian_rasterization.  Stable, digits of 5 to 9 bits, 32- or 64-bit keys,
// 32-bit values.
//
//   k_rs_hist  : ONE read of the keys builds the digit histograms of ALL passes (shared-memory
//                histograms, one global atomic per bin per CTA) and clears the first look-back buffer.  The
//                frame's two sorts do not run it: their histograms come from k_preprocess / k_tile_scan.
//   k_rs_pass3 : one kernel per digit of a 32-bit-key sort.  Persistent CTAs take equal spans of keys in ticket
//                order; a span is ranked in one or two sub-tiles (ballot matching + per-warp digit counters), parked
//                tile-sorted in shared memory, its per-digit counts are chained to the spans in front through
//                self-validating status words (AGG per tile, GROUP per 16 tiles, CKPT per group boundary:
//                tile_prefix / tile_prefix_thread), and keys/values leave in contiguous runs per digit.  Keys are read
//                once and written once per pass (the onesweep property).  The first pass of the frame's depth sort
//                also compacts (DROP: culled Gaussians have no key) and rebases the keys to 27 bits.
//   k_rs_pass  : the same for 64-bit keys (GSR_FLAG_KEYS64 cross-check path): tiles of 12 keys per thread, decoupled
//                look-back over one status word per (tile, digit).
//
// Item counts may live on the device (n_dev): grids are sized for the SM count, not for n, so the
// same launch (and the same captured CUDA graph) serves any n up to the buffer capacity.
#include <cstdlib>

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

template <int BITS>
__device__ __forceinline__ uint32_t match_digit_t(uint32_t d) {
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < BITS; ++b) {
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

// ---- inter-tile prefixes of k_rs_pass3: two-level look-back over self-validating words ------------------------------
// The decoupled look-back of k_rs_pass (every thread walks back over the count rows of its digit until it meets an
// inclusive prefix) is a poor fit for persistent CTAs that run in lock step: all tiles of a wave finish ranking at about
// the same time, nothing in front of them has an inclusive prefix yet, and every thread walks back over hundreds of rows,
// 16 volatile loads per round trip, while it takes issue slots from the CTAs that still rank (ncu: half of the executed
// instructions and 45 % of the stall samples of a pass were that loop).  Here the walk is over GROUPS of RS_GROUP tiles:
//   AGG[tile][d]   = count of digit d in the tile                      (published by the tile, first thing)
//   GROUP[g][d]    = sum of AGG over the tiles of group g              (published by the LAST tile of the group, as soon
//                                                                       as the 15 rows in front of it are there)
//   CKPT[g][d]     = exclusive prefix at the start of group g          (published by the last tile of group g - 1 once
//                                                                       it knows its own prefix; CKPT[0] = 0 implicitly)
//   prefix(tile)   = CKPT[h] + sum GROUP[h .. g-1] + sum AGG of the tiles in front of it in its own group g,
//                    h = the nearest group at or below g whose checkpoint is there.
// Every word carries its own ready bit (bit 31) next to a 31-bit value, so a word is complete the moment it is visible:
// no fences, no barriers, no atomics, every thread follows its own digits.  All loads of a step are issued before the
// first is consumed; what is not ready is read again after a back-off.  Every wait is for a tile with a smaller ticket.
// One status buffer: [AGG rows][GROUP rows][CKPT rows], rows of `stride` words, all zero at the start of a pass.
constexpr int RS_GROUP = 16;
constexpr uint32_t RS_READY = 0x80000000u;
struct RsStatus { uint32_t* agg; uint32_t* group; uint32_t* ckpt; long long groups; };
__host__ __device__ constexpr long long rs_status_groups(long long rows) { return rows / RS_GROUP + 2; }
__host__ __device__ constexpr long long rs_status_rows_total(long long rows) { return rows + 2 * rs_status_groups(rows); }
// words of one status buffer with rows of up to 512 words
__host__ __device__ constexpr long long rs_status_words(long long rows) { return rs_status_rows_total(rows) * 512; }
__host__ __device__ inline RsStatus rs_status_carve(uint32_t* base, long long rows, int stride) {
    RsStatus st;
    st.groups = rs_status_groups(rows);
    st.agg = base; st.group = base + rows * stride; st.ckpt = st.group + st.groups * stride;
    return st;
}

template <int DPT, int BINS>
__device__ __forceinline__ void tile_prefix_thread(const RsStatus& st, long long tile, long long n_tiles, int stride, uint32_t mask,
                                            const uint32_t (&total)[DPT], uint32_t (&prefix)[DPT]) {
    const int t = threadIdx.x;
    const long long g = tile / RS_GROUP;
    const int r = (int)(tile - g * RS_GROUP);
    const long long g_first = g * RS_GROUP;
    const int gsize = (int)((n_tiles - g_first < RS_GROUP) ? (n_tiles - g_first) : RS_GROUP);
    const bool last_of_group = (r == gsize - 1);
    bool exists[DPT];
    uint32_t sum_agg[DPT];
#pragma unroll
    for (int j = 0; j < DPT; ++j) {
        const int d = t * DPT + j;
        exists[j] = d < BINS && (uint32_t)d <= mask;  // digits above the mask do not exist in this pass
        if (exists[j]) st_volatile_u32(st.agg + tile * stride + d, RS_READY | total[j]);
        prefix[j] = 0; sum_agg[j] = 0;
    }
    bool any = false;
#pragma unroll
    for (int j = 0; j < DPT; ++j) any = any || exists[j];
    if (!any) return;
    // ---- the rows of the tiles in front of this one in its group
    if (r > 0) {
        uint32_t got[DPT];
        const uint32_t want = (1u << r) - 1u;
#pragma unroll
        for (int j = 0; j < DPT; ++j) got[j] = exists[j] ? 0u : want;
        unsigned backoff = 32;
        while (true) {
            // (unconditional loads from clamped rows, masked afterwards: one predicate per load exhausted the seven
            // predicate registers of a thread)
            uint32_t w[DPT][RS_GROUP - 1];
#pragma unroll
            for (int j = 0; j < DPT; ++j) {
                const uint32_t* col = st.agg + g_first * stride + (exists[j] ? t * DPT + j : 0);
#pragma unroll
                for (int i = 0; i < RS_GROUP - 1; ++i) w[j][i] = ld_volatile_u32(col + min(i, r - 1) * stride);  // (min, not a select: r is CTA-uniform and 15 selects exhaust the uniform predicate registers)
            }
            bool all = true;
#pragma unroll
            for (int j = 0; j < DPT; ++j) {
                uint32_t fresh = 0, add = 0;
#pragma unroll
                for (int i = 0; i < RS_GROUP - 1; ++i) {
                    const uint32_t take = (w[j][i] >> 31) & (want >> i) & ~(got[j] >> i) & 1u;  // ready, wanted, not yet counted
                    add += take ? (w[j][i] & ~RS_READY) : 0u;
                    fresh |= take << i;
                }
                sum_agg[j] += add;
                got[j] |= fresh;
                all = all && (got[j] == want);
            }
            if (all) break;
            // wait on ONE word - the first row that is still missing - not on the whole window
#pragma unroll
            for (int j = 0; j < DPT; ++j) {
                if (got[j] != want) {
                    const uint32_t* word = st.agg + (g_first + (__ffs(~got[j] & want) - 1)) * stride + t * DPT + j;
                    while (!(ld_volatile_u32(word) & RS_READY)) {
                        __nanosleep(backoff);
                        if (backoff < 1024) backoff *= 2;
                    }
                }
            }
        }
    }
    if (last_of_group) {
#pragma unroll
        for (int j = 0; j < DPT; ++j)
            if (exists[j]) st_volatile_u32(st.group + g * stride + t * DPT + j, RS_READY | (sum_agg[j] + total[j]));
    }
    // ---- back over the groups to the nearest checkpoint
    if (g > 0) {
        constexpr int WG = (DPT > 1) ? 8 : 16;  // groups per step: 2 * WG * DPT loads in flight
        long long at[DPT];   // next group boundary to examine: CKPT[at], then GROUP[at - 1]
        bool found[DPT];
        uint32_t base[DPT];
#pragma unroll
        for (int j = 0; j < DPT; ++j) { at[j] = g; found[j] = !exists[j]; base[j] = 0; }
        unsigned backoff = 32;
        while (true) {
            uint32_t c[DPT][WG], q[DPT][WG];
#pragma unroll
            for (int j = 0; j < DPT; ++j) {
                const int col = exists[j] ? t * DPT + j : 0;
#pragma unroll
                for (int k = 0; k < WG; ++k) {
                    long long hgrp = at[j] - k;  // unconditional loads from clamped rows (see above); rows below 1 are
                    if (hgrp < 1) hgrp = 1;      // never consumed: the walk ends at group boundary 0
                    c[j][k] = ld_volatile_u32(st.ckpt + hgrp * stride + col);
                    q[j][k] = ld_volatile_u32(st.group + (hgrp - 1) * stride + col);
                }
            }
            bool all = true, progress = false;
#pragma unroll
            for (int j = 0; j < DPT; ++j) {
                bool stop = found[j];
#pragma unroll
                for (int k = 0; k < WG; ++k) {
                    if (!stop) {
                        if (at[j] == 0) { found[j] = true; stop = true; }               // start of the keys
                        else if (c[j][k] & RS_READY) { base[j] += c[j][k] & ~RS_READY; found[j] = true; stop = true; progress = true; }
                        else if (q[j][k] & RS_READY) { base[j] += q[j][k] & ~RS_READY; at[j] -= 1; progress = true; }
                        else stop = true;                                               // not there yet
                    }
                }
                if (!found[j] && at[j] == 0) found[j] = true;
                all = all && found[j];
            }
            if (all) break;
            if (!progress) {  // blocked: wait on the two words at the boundary, not on the whole window
#pragma unroll
                for (int j = 0; j < DPT; ++j) {
                    if (!found[j]) {
                        const uint32_t* cw = st.ckpt + at[j] * stride + t * DPT + j;
                        const uint32_t* qw = st.group + (at[j] - 1) * stride + t * DPT + j;
                        while (!((ld_volatile_u32(cw) | ld_volatile_u32(qw)) & RS_READY)) {
                            __nanosleep(backoff);
                            if (backoff < 1024) backoff *= 2;
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int j = 0; j < DPT; ++j) prefix[j] = base[j];
    }
#pragma unroll
    for (int j = 0; j < DPT; ++j) prefix[j] += sum_agg[j];
    if (last_of_group && g_first + gsize < n_tiles) {  // the checkpoint of the next group
#pragma unroll
        for (int j = 0; j < DPT; ++j)
            if (exists[j]) st_volatile_u32(st.ckpt + (g + 1) * stride + t * DPT + j, RS_READY | (prefix[j] + total[j]));
    }
}


#ifdef GSR_RS_TRACE  // development only: globaltimer stamps of the first super-tile of every CTA, per pass of a frame
__device__ unsigned long long g_rs_trace[5 * 1024 * 10];
__device__ unsigned long long g_rs_trace2[5 * 1024 * 10];
__device__ int g_rs_slot_dummy;
#define RS_P(i, val) do { if (threadIdx.x == 0 && tr_on && blockIdx.x < 1024) g_rs_trace2[(tr_slot * 1024 + blockIdx.x) * 10 + (i)] = (val); } while (0)
__device__ __forceinline__ unsigned long long rs_now() { unsigned long long v; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(v)); return v; }
__device__ unsigned long long g_rs_acc[5 * 1024 * 8];  // per CTA, summed over ALL its tiles: ticket, clear, sub-tile 0, sub-tile 1, prefix, scatter; tiles
#define RS_T(i) do { if (threadIdx.x == 0) { const unsigned long long now_ = rs_now(); if ((i) >= 2) { rs_acc[(i) - 2] += now_ - rs_prev; } rs_prev = now_; if (rs_first && blockIdx.x < 1024) g_rs_trace[(rs_slot * 1024 + blockIdx.x) * 10 + (i)] = now_; } } while (0)
#else
#define RS_T(i) do { } while (0)
#define RS_P(i, val) do { } while (0)
#endif

__device__ __forceinline__ uint4 ld_volatile_v4(const uint32_t* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

// The walk is done by the CTA as a whole (globaltimer trace of the per-thread version, scripts/probe_rs_trace.py: in a
// single-wave pass the prefix of a tile was known 10-12 us after the last tile in front of it had counted - five
// dependent round trips of up to 32 scalar loads per thread, with all 113 k threads of the grid polling).  One STEP reads
// up to 16 rows with 128-bit loads spread over the 256 threads (a thread keeps one column of four digits and rows
// row0, row0 + rpt, ...), an OR over the CTA tells which rows are complete, the leading complete rows are summed in
// registers, and the owners of the digits add up the partial columns out of shared memory.  A blocked CTA waits with ONE
// thread polling one word.
//   s_part: RS_THREADS * 4 words, s_ctl: 2 words of shared memory.
template <int DPT, int BINS>
__device__ __forceinline__ void tile_prefix(const RsStatus& st, long long tile, long long n_tiles, int stride, uint32_t mask,
                                            const uint32_t (&total)[DPT], uint32_t (&prefix)[DPT], uint32_t* s_part,
                                            uint32_t* s_ctl, int tr_slot = 0, bool tr_on = false) {
    (void)tr_slot; (void)tr_on;
    unsigned long long n_steps = 0, n_waits = 0; (void)n_steps; (void)n_waits;
    RS_P(0, rs_now());
    constexpr int NCOL_MAX = (BINS >= 4) ? BINS / 4 : 1;
    constexpr int L = (16 * NCOL_MAX > RS_THREADS) ? 16 * NCOL_MAX / RS_THREADS : 1;  // 512 bins: 8, 256: 4, 128: 2, else 1
    const int t = threadIdx.x;
    const long long g = tile / RS_GROUP;
    const int r = (int)(tile - g * RS_GROUP);
    const long long g_first = g * RS_GROUP;
    const int gsize = (int)((n_tiles - g_first < RS_GROUP) ? (n_tiles - g_first) : RS_GROUP);
    const bool last_of_group = (r == gsize - 1);
    bool exists[DPT];
#pragma unroll
    for (int j = 0; j < DPT; ++j) {
        const int d = t * DPT + j;
        exists[j] = d < BINS && (uint32_t)d <= mask;  // digits above the mask do not exist in this pass
        if (exists[j]) st_volatile_u32(st.agg + tile * stride + d, RS_READY | total[j]);
        prefix[j] = 0;
    }
    if (tile == 0) return;  // CTA-uniform: nothing in front (a group of one tile has no GROUP / CKPT reader either)
    const int nd = (int)mask + 1;
    const int cshift = (nd >= 4) ? (29 - __clz(nd)) : 0;  // log2 of the number of four-word columns of a row
    const int ncol = 1 << cshift;
    const int col = t & (ncol - 1), row0 = t >> cshift, rpt = RS_THREADS >> cshift;
    const int nparts = rpt < 16 ? rpt : 16;  // partial columns that can hold a row
    uint32_t ex = 0;                          // which of the four words of this thread's column are digits of the pass
#pragma unroll
    for (int w = 0; w < 4; ++w) ex |= ((uint32_t)(4 * col + w) <= mask) ? (1u << w) : 0u;
    if (t == 0) { s_ctl[0] = 0u; s_ctl[1] = 0u; }
    __syncthreads();
    int par = 0;
    uint4 v[L];
    // loads rows i = 0 .. n_rows - 1 (row i at ptr(i)); returns a CTA-uniform mask of the rows that are NOT complete yet
    auto load_rows = [&](auto ptr, int n_rows) -> uint32_t {
        uint32_t missing = 0;
        // every load of the step is issued before the first one is looked at (a warp issues in order: a consumer right
        // behind its load makes the step L dependent round trips - measured: 4.8 us instead of 1)
#pragma unroll
        for (int k = 0; k < L; ++k) {
            const int i = row0 + k * rpt;
            if (i < n_rows) v[k] = ld_volatile_v4(ptr(i) + 4 * col);
            else v[k] = make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll
        for (int k = 0; k < L; ++k) {
            const int i = row0 + k * rpt;
            const uint32_t rdy = (v[k].x >> 31) | ((v[k].y >> 31) << 1) | ((v[k].z >> 31) << 2) | ((v[k].w >> 31) << 3);
            if (i < n_rows && (rdy & ex) != ex) missing |= 1u << i;
        }
        missing = __reduce_or_sync(0xffffffffu, missing);
        if ((t & 31) == 0) {
            if (missing) atomicOr(&s_ctl[par], missing);
        }
        if (t == 0) s_ctl[par ^ 1] = 0u;
        __syncthreads();
        const uint32_t m = s_ctl[par];
        par ^= 1;
        return m;
    };
    // this thread's column summed over the rows [lo, hi) of the last load
    auto sum_rows = [&](int lo, int hi) -> uint4 {
        uint4 a = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
        for (int k = 0; k < L; ++k) {
            const int i = row0 + k * rpt;
            if (i >= lo && i < hi) {
                a.x += v[k].x & ~RS_READY; a.y += v[k].y & ~RS_READY; a.z += v[k].z & ~RS_READY; a.w += v[k].w & ~RS_READY;
            }
        }
        return a;
    };
    // partial columns -> the owners of the digits
    auto to_owners = [&](const uint4& a, uint32_t (&dst)[DPT]) {
        if (row0 < nparts) reinterpret_cast<uint4*>(s_part)[row0 * ncol + col] = a;
        __syncthreads();
#pragma unroll
        for (int j = 0; j < DPT; ++j) {
            if (exists[j]) {
                uint32_t acc = 0;
                for (int q = 0; q < nparts; ++q) acc += s_part[q * (4 * ncol) + t * DPT + j];
                dst[j] += acc;
            }
        }
        // (s_part is written again only behind the barrier of the next load_rows)
    };
    auto wait_for = [&](const uint32_t* w0, const uint32_t* w1) {  // one thread polls; digit 0 exists in every pass
        if (t == 0) {
            unsigned backoff = 32;
            while (!((ld_volatile_u32(w0) | (w1 ? ld_volatile_u32(w1) : 0u)) & RS_READY)) {
                __nanosleep(backoff);
                if (backoff < 512) backoff *= 2;
            }
        }
        __syncthreads();
    };

    // ---- the rows of the tiles in front of this one in its group: tile - 1, tile - 2, ..., g_first
    uint32_t sum_agg[DPT];
#pragma unroll
    for (int j = 0; j < DPT; ++j) sum_agg[j] = 0;
    {
        long long top = tile - 1;
        int need = r;
        while (need > 0) {
            const uint32_t* row_top = st.agg + top * stride;
            const uint32_t missing = load_rows([&](int i) { return row_top - (long long)i * stride; }, need);
            ++n_steps;
            int run = missing ? __ffs(missing) - 1 : need;
            if (run > need) run = need;
            if (run > 0) {
                to_owners(sum_rows(0, run), sum_agg);
                top -= run; need -= run;
            } else {
                ++n_waits;
                wait_for(row_top, nullptr);
            }
        }
    }
    RS_P(1, rs_now()); RS_P(5, n_steps | (n_waits << 32)); n_steps = 0; n_waits = 0;
    if (last_of_group) {
#pragma unroll
        for (int j = 0; j < DPT; ++j)
            if (exists[j]) st_volatile_u32(st.group + g * stride + t * DPT + j, RS_READY | (sum_agg[j] + total[j]));
    }
    bool tr_firststep = true; (void)tr_firststep;
    // ---- back over the groups to the nearest checkpoint: row 0 of a step is CKPT[at], rows 1.. are GROUP[at - 1], ...
    // (letting only the group's publisher walk and the other tiles wait for their checkpoint was measured: a third less
    // status traffic, but the extra publish -> poll -> read hop costs more, depth sort 0.108 -> 0.112 ms)
    {
        long long at = g;
        while (at > 0) {
            const int n_rows = (int)((at + 1 < 16) ? at + 1 : 16);
            const uint32_t* ck = st.ckpt + at * stride;
            const uint32_t* gr = st.group + at * stride;
            const uint32_t missing = load_rows([&](int i) { return i == 0 ? ck : gr - (long long)i * stride; }, n_rows);
            ++n_steps;
            if (tr_firststep) { RS_P(2, rs_now()); tr_firststep = false; }
            if (!(missing & 1u)) {  // the checkpoint is there
                to_owners(sum_rows(0, 1), prefix);
                break;
            }
            const uint32_t mg = missing >> 1;
            int run = mg ? __ffs(mg) - 1 : n_rows - 1;
            if (run > n_rows - 1) run = n_rows - 1;
            if (run > 0) {
                to_owners(sum_rows(1, 1 + run), prefix);
                at -= run;
            } else {
                ++n_waits;
                wait_for(ck, gr - stride);
                RS_P(3, rs_now());
            }
        }
    }
    RS_P(4, rs_now()); RS_P(6, n_steps | (n_waits << 32)); RS_P(7, (unsigned long long)last_of_group);
#pragma unroll
    for (int j = 0; j < DPT; ++j) prefix[j] += sum_agg[j];
    if (last_of_group && g_first + gsize < n_tiles) {  // the checkpoint of the next group
#pragma unroll
        for (int j = 0; j < DPT; ++j)
            if (exists[j]) st_volatile_u32(st.ckpt + (g + 1) * stride + t * DPT + j, RS_READY | (prefix[j] + total[j]));
    }
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
                                                       uint32_t* __restrict__ status0, long long clear_words) {
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
    for (long long i = (long long)blockIdx.x * RS_THREADS + threadIdx.x; i < clear_words; i += stride) status0[i] = 0;
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

// ---- k_rs_pass3: the 32-bit-key pass over SUPER-TILES of S sub-tiles, digits of up to 9 bits -------------------------
// clock64 phase timing of k_rs_pass showed 40 % of a pass waiting in the look-back: the persistent CTAs run in lock
// step, so in every wave the inclusive prefix has to cross all ~444 tiles in flight before the last of them can
// write, and that chain repeats every wave.  Its total length is (tiles) / sqrt(tiles in flight), so the cure is
// fewer, larger tiles without fewer CTAs: a CTA ranks S sub-tiles one after the other, parks each one tile-sorted in
// shared memory (the registers are free again for the next), publishes ONE count row for all of them and does ONE
// look-back.
//   BITS  : digit width of the template instance (ballot rounds, bins); a pass with fewer bits uses the next instance
//           up (its upper ballots agree on every lane).  Thread t owns digits [t * DPT, (t + 1) * DPT).
//   partial sub-tiles: a warp owns ITEMS * 32 consecutive keys of a sub-tile; warps whose chunk lies behind the end of
//           the span skip the ranking altogether (a warp-uniform branch around the whole phase - per-item guards
//           were measured 1.5x slower: they cut the unrolled ballot chains into basic blocks).
//   DROP  : the compacting first pass of the frame's depth sort: keys equal to 0xFFFFFFFF (culled Gaussians) are
//           neither counted nor written, the others leave as key' = min(max(key, k0), khi) - k0, which makes the
//           visible depth range a 27-bit integer (three 9-bit passes over the visible Gaussians instead of four 8-bit
//           passes over all of them).
struct RsRebase { uint32_t k0, khi; };


template <int BITS, int ITEMS, int S, int MINB, bool IDENTITY_VALS, bool DROP>
__global__ void __launch_bounds__(RS_THREADS, MINB) k_rs_pass3(const uint32_t* __restrict__ keys_in, uint32_t* __restrict__ keys_out,
                                                          const uint32_t* __restrict__ vals_in,
                                                          uint32_t* __restrict__ vals_out, long long n_host,
                                                          const uint32_t* __restrict__ n_dev, long long capacity,
                                                          int shift, int bits, const uint32_t* __restrict__ hist,
                                                          uint32_t* __restrict__ ticket, uint32_t* __restrict__ status_cur,
                                                          uint32_t* __restrict__ status_next, int row_stride, int span_max,
                                                          long long status_rows, RsRebase rb) {
    constexpr int BINS = 1 << BITS;
    constexpr int DPT = (BINS + RS_THREADS - 1) / RS_THREADS;  // digits per thread in the per-digit phases
    extern __shared__ __align__(16) unsigned char s_dyn[];
    uint32_t* s_keys = reinterpret_cast<uint32_t*>(s_dyn);    // [span_max] tile-sorted keys, sub-tile after sub-tile
    uint32_t* s_vals = s_keys + span_max;                     // [span_max]
    __shared__ __align__(16) uint16_t s_whist[RS_WARPS * BINS];  // per-warp digit counts (<= 32 * ITEMS each)
    __shared__ uint16_t s_digit_start[BINS];
    __shared__ uint32_t s_bin_base[S][BINS];
    __shared__ uint32_t s_tmp[RS_WARPS + 1];
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_ctl[2];
    // partial columns of tile_prefix: the per-warp counters are free by then (they are 4 KB from 256 bins on)
    __shared__ __align__(16) uint32_t s_part_small[4];  // (narrower digits would need RS_THREADS * 4 words of their own)
    uint32_t* const s_part = (BINS >= 512) ? reinterpret_cast<uint32_t*>(s_whist) : s_part_small;
    const RsStatus st = rs_status_carve(status_cur, status_rows, row_stride);
    const RsStatus stn = rs_status_carve(status_next, status_rows, row_stride);

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t mask = (1u << bits) - 1u;
    const uint32_t lt_mask = (1u << lane) - 1u;
#ifdef GSR_RS_TRACE
    const int rs_slot = (bits == 9) ? shift / 9 : (shift > 0 ? 4 : 3);
    bool rs_first = true;
    unsigned long long rs_prev = 0, rs_acc[6] = {0, 0, 0, 0, 0, 0}, rs_tiles = 0;
    RS_T(0);
#endif
    pdl_wait();  // resident behind the previous pass; nothing of global memory is touched before this
    RS_T(1);
    long long n = n_dev ? (long long)(*n_dev) : n_host;
    if (n > capacity) n = capacity;

    // exclusive scan of the pass's global digit counts; thread t keeps the bases of its own digits
    uint32_t gexcl[DPT];
    {
        uint32_t loc[DPT], sum = 0;
#pragma unroll
        for (int j = 0; j < DPT; ++j) {
            const int d = t * DPT + j;
            loc[j] = (d < BINS) ? hist[d] : 0u;
            sum += loc[j];
        }
        uint32_t run = block_excl_scan_256(sum, s_tmp, nullptr);
#pragma unroll
        for (int j = 0; j < DPT; ++j) { gexcl[j] = run; run += loc[j]; }
    }

    // Balanced spans: every CTA of the grid gets the same number of equally long super-tiles (with fixed spans 3 M
    // keys are 367 super-tiles on 444 CTA slots, and in a single-wave pass everybody's look-back waits for the
    // slowest count row).
    const long long G = gridDim.x;
    const long long waves = (n + G * span_max - 1) / (G * span_max);
    long long span = (n + G * waves - 1) / (G * (waves > 0 ? waves : 1));
    span = (span + 31) & ~31ll;  // 128-byte aligned warp loads
    if (span > span_max) span = span_max;
    if (span < 256) span = 256;
    const long long n_tiles = (n + span - 1) / span;
    // the next pass has at most as many keys, hence at most this many tiles (and never more than status_rows)
    // (its instance may cut shorter spans than this one: count with the shortest span any instance uses)
    const long long span_floor = span_max < 27 * RS_THREADS ? span_max : 27 * RS_THREADS;
    const long long waves_next = (n + G * span_floor - 1) / (G * span_floor);
    long long rows_next = G * (waves_next > 0 ? waves_next : 1) + 1;
    if (rows_next > status_rows) rows_next = status_rows;
    const long long groups_next = rows_next / RS_GROUP + 2;

    while (true) {
        // (drawn behind the histogram scan on purpose: CTAs that get going first take the spans in front, so a tile
        // rarely waits for a slower predecessor; drawing the first ticket before the scan, beside the histogram loads, cost
        // 5 us per sort)
        if (t == 0) s_tile = atomicAdd(ticket, 1u);
        __syncthreads();
        const long long tile = s_tile;  // super-tile index
        const long long super_base = tile * span;
        if (super_base >= n) break;
        RS_T(2);
#ifdef GSR_RS_TRACE
        if (threadIdx.x == 0 && rs_first && blockIdx.x < 1024) {
            unsigned smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            g_rs_trace[(rs_slot * 1024 + blockIdx.x) * 10 + 8] = ((unsigned long long)smid << 32) | (unsigned long long)tile;
            g_rs_trace[(rs_slot * 1024 + blockIdx.x) * 10 + 9] = (unsigned long long)n_tiles;
        }
#endif
        const long long super_end = (super_base + span < n) ? super_base + span : n;
        // the status words of the next pass start at zero: its AGG rows [0, rows_next) and, behind the status_rows AGG rows
        // of the buffer, all GROUP rows and the CKPT rows [0, groups_next) - one loop over the two row ranges (two loops
        // made ptxas fail its register allocation, C7600)
        for (long long idx = tile; idx < rows_next + stn.groups + groups_next; idx += n_tiles) {
            const long long row = idx < rows_next ? idx : status_rows + (idx - rows_next);
            for (int j = t; j < row_stride; j += RS_THREADS) status_next[row * row_stride + j] = 0;
        }
        RS_T(3);
        constexpr int per_sub = ITEMS * RS_THREADS;
        uint32_t cnt[S][DPT];   // digit d of sub-tile s: keys
        int sub_out[S];         // keys parked by sub-tile s
        int out_base = 0;       // start of the sub-tile in the staging area
#pragma unroll
        for (int sub = 0; sub < S; ++sub) {
            const long long sub_base = super_base + (long long)sub * per_sub;
            const long long sub_end = (sub_base + per_sub < super_end) ? sub_base + per_sub : super_end;
#pragma unroll
            for (int j = 0; j < DPT; ++j) cnt[sub][j] = 0;
            sub_out[sub] = 0;
            if (sub_base >= super_end) continue;  // block-uniform
            for (int i = t; i < RS_WARPS * BINS / 8; i += RS_THREADS) reinterpret_cast<uint4*>(s_whist)[i] = make_uint4(0u, 0u, 0u, 0u);
            __syncthreads();
            // ---- load (warp-striped: item i of lane l is element warp_first + i*32 + l; 32-bit offsets from one pointer)
            // more than 16 keys per thread: the values are fetched after the ranking, into the registers of the keys
            // (key + peer mask/rank + value per item do not fit 80 registers)
            constexpr bool DEFER_VALS = !IDENTITY_VALS && ITEMS > 16;
            uint32_t key[ITEMS], val[DEFER_VALS ? 1 : ITEMS];
            const long long warp_first = sub_base + (long long)warp * (ITEMS * 32);
            const bool warp_has_keys = warp_first < sub_end;  // warp-uniform
            // items i with i * 32 < left hold a key of the sub-tile
            const int left = (int)((sub_end - warp_first < (long long)(ITEMS * 32)) ? (sub_end - warp_first) : (long long)(ITEMS * 32)) - lane;
            const uint32_t* kp = keys_in + warp_first + lane;
            const uint32_t* vp = IDENTITY_VALS ? nullptr : vals_in + warp_first + lane;
            const uint32_t idx0 = (uint32_t)warp_first + (uint32_t)lane;
            uint32_t vbits = 0;  // DROP only: bit i = item i takes part (a key that is not dropped)
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) {
                const bool in = i * 32 < left;
                key[i] = in ? kp[i * 32] : 0xFFFFFFFFu;  // padding keys: digit `mask` of every pass, behind the valid keys
                if constexpr (IDENTITY_VALS) val[i] = idx0 + i * 32;
                else if constexpr (!DEFER_VALS) val[i] = in ? vp[i * 32] : 0u;
                if (DROP) {
                    const bool keep = in && (key[i] != 0xFFFFFFFFu);
                    key[i] = min(max(key[i], rb.k0), rb.khi) - rb.k0;
                    vbits |= (keep ? 1u : 0u) << i;
                }
            }
            // padding keys that get counted: those of the warps that hold at least one key of the sub-tile
            const int n_sub = (int)(sub_end - sub_base);
            const int n_pad = ((n_sub + ITEMS * 32 - 1) / (ITEMS * 32)) * (ITEMS * 32) - n_sub;
            // ---- rank inside the warp, digit counters per warp.  All matches first: they are independent of the
            // counters, so their ballots overlap the serial counter updates below.  The update is branch free: every
            // lane reads its digit's counter (a broadcast among peers), the lowest peer writes it back.
            uint32_t rank[ITEMS];  // first the peer mask, then the rank
            if (warp_has_keys) {
                uint16_t* wh = s_whist + warp * BINS;
#pragma unroll
                for (int i = 0; i < ITEMS; ++i) {
                    const uint32_t d = (key[i] >> shift) & mask;
                    uint32_t m = match_digit_t<BITS>(d);
                    if (DROP) m &= __ballot_sync(0xffffffffu, (vbits >> i) & 1u);
                    rank[i] = m;
                }
#pragma unroll
                for (int i = 0; i < ITEMS; ++i) {
                    const uint32_t d = (key[i] >> shift) & mask;
                    const uint32_t peers = rank[i];
                    const uint32_t below = peers & lt_mask;
                    const uint32_t old = wh[d];
                    __syncwarp();  // every peer has read the counter
                    if (below == 0u && (!DROP || ((vbits >> i) & 1u))) wh[d] = (uint16_t)(old + __popc(peers));
                    __syncwarp();  // the update is visible to the next item's reads
                    rank[i] = old + __popc(below);
                }
            } else {
#pragma unroll
                for (int i = 0; i < ITEMS; ++i) rank[i] = 0;
            }
            __syncthreads();
            // ---- per digit (thread t owns digits t * DPT ...): exclusive scan over warps, sub-tile count
            uint32_t loc[DPT], sum = 0;
#pragma unroll
            for (int j = 0; j < DPT; ++j) {
                const int d = t * DPT + j;
                uint32_t count = 0;
                if (d < BINS) {
#pragma unroll
                    for (int w = 0; w < RS_WARPS; ++w) {
                        const uint32_t c = s_whist[w * BINS + d];
                        s_whist[w * BINS + d] = (uint16_t)count;
                        count += c;
                    }
                }
                loc[j] = count; sum += count;
                // padding keys (all ones) sit in digit `mask`, behind every valid key of that digit (not in the
                // compacting pass: it counts only the keys that take part)
                cnt[sub][j] = count - ((!DROP && (uint32_t)d == mask) ? (uint32_t)n_pad : 0u);
            }
            uint32_t total = 0;
            uint32_t run = block_excl_scan_256(sum, s_tmp, &total);
            sub_out[sub] = DROP ? (int)total : n_sub;
#pragma unroll
            for (int j = 0; j < DPT; ++j) {
                const int d = t * DPT + j;
                if (d < BINS) {
                    s_digit_start[d] = (uint16_t)run;
                    s_bin_base[sub][d] = 0u - run;  // completed below once the look-back is done
                }
                run += loc[j];
            }
            __syncthreads();
            // ---- park the sub-tile in shared memory in tile-sorted order
            uint32_t* sk = s_keys + out_base;
            uint32_t* sv = s_vals + out_base;
            if (warp_has_keys) {
                const uint16_t* wh = s_whist + warp * BINS;
                // padding keys (only the last warp with keys has any; they rank behind every valid key of the last
                // digit) are not parked: the staging area holds the span's keys and nothing else
                if constexpr (DEFER_VALS) {
#pragma unroll
                    for (int i = 0; i < ITEMS; ++i) {
                        const uint32_t d = (key[i] >> shift) & mask;
                        const uint32_t pos = (uint32_t)s_digit_start[d] + wh[d] + rank[i];
                        rank[i] = pos;
                        if (i * 32 < left) sk[pos] = key[i];
                    }
#pragma unroll
                    for (int i = 0; i < ITEMS; ++i) key[i] = (i * 32 < left) ? vp[i * 32] : 0u;
#pragma unroll
                    for (int i = 0; i < ITEMS; ++i)
                        if (i * 32 < left) sv[rank[i]] = key[i];
                } else {
#pragma unroll
                    for (int i = 0; i < ITEMS; ++i) {
                        // (these instances stage S * ITEMS * 256 keys, so padding keys may park behind the valid ones)
                        if (!DROP || ((vbits >> i) & 1u)) {
                            const uint32_t d = (key[i] >> shift) & mask;
                            const uint32_t pos = (uint32_t)s_digit_start[d] + wh[d] + rank[i];
                            sk[pos] = key[i];
                            sv[pos] = val[i];
                        }
                    }
                }
            }
            out_base += sub_out[sub];
            __syncthreads();  // s_whist / s_digit_start are rewritten by the next sub-tile
            if (sub == 0) RS_T(4);
        }
        RS_T(5);
        // ---- one count row for the whole super-tile, then its prefix over the tiles in front (tile_prefix)
        uint32_t total[DPT], prefix[DPT];
#pragma unroll
        for (int j = 0; j < DPT; ++j) {
            total[j] = 0;
#pragma unroll
            for (int sub = 0; sub < S; ++sub) total[j] += cnt[sub][j];
            prefix[j] = 0;
        }
        // 9-bit digits (the frame's depth sort: single-wave passes, 2 KB rows) walk as a CTA; narrower digits keep the
        // per-thread walk, which has no barriers and in a multi-wave pass finds a checkpoint at its first probe (the
        // CTA-wide walk cost the 4K tile sort 6 us)
        if constexpr (BINS >= 512) {
#ifdef GSR_RS_TRACE
            tile_prefix<DPT, BINS>(st, tile, n_tiles, row_stride, mask, total, prefix, s_part, s_ctl, rs_slot, rs_first);
#else
            tile_prefix<DPT, BINS>(st, tile, n_tiles, row_stride, mask, total, prefix, s_part, s_ctl);
#endif
        } else {
            tile_prefix_thread<DPT, BINS>(st, tile, n_tiles, row_stride, mask, total, prefix);
        }
        RS_T(6);
#pragma unroll
        for (int j = 0; j < DPT; ++j) {
            const int d = t * DPT + j;
            if (d < BINS) {
                uint32_t run = gexcl[j] + prefix[j];
#pragma unroll
                for (int sub = 0; sub < S; ++sub) {
                    s_bin_base[sub][d] += run;
                    run += cnt[sub][j];
                }
            }
        }
        __syncthreads();
        // ---- contiguous runs per digit go out to global memory
        int base = 0;
#pragma unroll
        for (int sub = 0; sub < S; ++sub) {
            const uint32_t* sk = s_keys + base;
            const uint32_t* sv = s_vals + base;
            for (int s = t; s < sub_out[sub]; s += RS_THREADS) {
                const uint32_t k = sk[s];
                const uint32_t d = (k >> shift) & mask;
                const uint32_t o = s_bin_base[sub][d] + (uint32_t)s;
                if (keys_out) keys_out[o] = k;
                vals_out[o] = sv[s];
            }
            base += sub_out[sub];
        }
        __syncthreads();  // staging and s_bin_base are rewritten by the next super-tile
        RS_T(7);
#ifdef GSR_RS_TRACE
        rs_first = false; ++rs_tiles;
#endif
    }
#ifdef GSR_RS_TRACE
    if (threadIdx.x == 0 && blockIdx.x < 1024) {
        for (int i = 0; i < 6; ++i) g_rs_acc[(rs_slot * 1024 + blockIdx.x) * 8 + i] = rs_acc[i];
        g_rs_acc[(rs_slot * 1024 + blockIdx.x) * 8 + 6] = rs_tiles;
        g_rs_acc[(rs_slot * 1024 + blockIdx.x) * 8 + 7] = rs_now();
    }
#endif
}

#ifdef GSR_RS_TRACE
}  // namespace
extern "C" __attribute__((visibility("default"))) int gsr_debug_rs_trace(unsigned long long* out, long long n_words) {
    if (n_words > 5 * 1024 * 10) n_words = 5 * 1024 * 10;
    return (int)cudaMemcpyFromSymbol(out, g_rs_trace, (size_t)n_words * 8);
}
extern "C" __attribute__((visibility("default"))) int gsr_debug_rs_acc(unsigned long long* out, long long n_words) {
    if (n_words > 5 * 1024 * 8) n_words = 5 * 1024 * 8;
    return (int)cudaMemcpyFromSymbol(out, g_rs_acc, (size_t)n_words * 8);
}
extern "C" __attribute__((visibility("default"))) int gsr_debug_rs_trace2(unsigned long long* out, long long n_words) {
    if (n_words > 5 * 1024 * 10) n_words = 5 * 1024 * 10;
    return (int)cudaMemcpyFromSymbol(out, g_rs_trace2, (size_t)n_words * 8);
}
namespace {
#endif

int num_sms() {  // of the CURRENT device (a process may hold handles on several)
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

// ---- k_rs_pass3 configuration: 15 keys per thread and sub-tile, two sub-tiles, three CTAs per SM -----------------
// 9-bit digits: 13 KB of counters per CTA leave 60 KB of staging (15 keys per thread); up to 8 bits: 16 keys
// The frame's depth sort runs single-wave passes (one super-tile per resident CTA), and the globaltimer trace of such
// a pass (scripts/probe_rs_trace.py) shows a sub-tile costing ~5 us whether it holds 832 or 3840 keys (global-load
// latency, four barriers, the counter scan): so those passes take their span as ONE sub-tile with more keys per thread
// - 27 in the compacting first pass (identity values: key + rank only), 19 (+ a second sub-tile for views where more
// Gaussians are visible) in the other two, which fetch the values after the ranking.
constexpr int RS3_MINB = 3;
template <int BITS, bool IDENT, bool DROP> struct Rs3Cfg {
    static constexpr int ITEMS = BITS > 8 ? 15 : 16, S = 2, SPAN = S * ITEMS * RS_THREADS;  // 7680 / 8192 keys
};
template <> struct Rs3Cfg<9, true, true> { static constexpr int ITEMS = 27, S = 1, SPAN = 27 * RS_THREADS; };    // 6912 (15 x 2 measured 1 % slower)
template <> struct Rs3Cfg<9, false, false> { static constexpr int ITEMS = 19, S = 2, SPAN = 30 * RS_THREADS; };  // 7680 (4864 + 2816)
constexpr int rs3_span(int bits) { return bits > 8 ? 27 * RS_THREADS : 2 * 16 * RS_THREADS; }  // smallest span of a plan: sizes grid and rows

template <int BITS, bool IDENT, bool DROP>
void rs3_launch(unsigned grid, cudaStream_t stream, const uint32_t* kin, uint32_t* kout, const uint32_t* vin, uint32_t* vout,
                long long n_host, const uint32_t* n_dev, long long capacity, int shift, int bits, const uint32_t* hist,
                uint32_t* ticket, uint32_t* st_cur, uint32_t* st_next, int row_stride, int span_max, long long status_rows,
                RsRebase rb) {
    using Cfg = Rs3Cfg<BITS, IDENT, DROP>;
    (void)span_max;  // every instance has its own span; the caller's figure only sizes the grid and the rows
    launch_pdl_m(1, k_rs_pass3<BITS, Cfg::ITEMS, Cfg::S, RS3_MINB, IDENT, DROP>, dim3(grid), dim3(RS_THREADS),
                 (size_t)Cfg::SPAN * 8, stream, kin, kout, vin, vout, n_host, n_dev, capacity, shift, bits, hist, ticket, st_cur,
                 st_next, row_stride, (int)Cfg::SPAN, status_rows, rb);
}
template <int BITS, bool IDENT, bool DROP>
bool rs3_prepare_one() {
    using Cfg = Rs3Cfg<BITS, IDENT, DROP>;
    const void* f = (const void*)k_rs_pass3<BITS, Cfg::ITEMS, Cfg::S, RS3_MINB, IDENT, DROP>;
    return cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SPAN * 8) == cudaSuccess &&
           cudaFuncSetAttribute(f, cudaFuncAttributePreferredSharedMemoryCarveout, 100) == cudaSuccess;
}
int g_rs3_ctas[64] = {};  // resident CTAs per SM of the pass kernel, per device (0 = not prepared yet)

}  // namespace

// Opt-ins of the pass kernels (dynamic shared memory above 48 KB) for the CURRENT device; gsr_create calls this
// for every handle, so a process with handles on several GPUs configures each of them.
int radix_sort_prepare_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return -1;
    bool ok = true;
    ok = rs3_prepare_one<5, false, false>() && ok;
    ok = rs3_prepare_one<6, false, false>() && ok;
    ok = rs3_prepare_one<7, false, false>() && ok;
    ok = rs3_prepare_one<8, false, false>() && ok;
    ok = rs3_prepare_one<9, false, false>() && ok;
    ok = rs3_prepare_one<8, true, false>() && ok;
    ok = rs3_prepare_one<8, true, true>() && ok;
    ok = rs3_prepare_one<9, true, true>() && ok;
    int nb = 0;
    if (!ok || cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_rs_pass3<9, Rs3Cfg<9, false, false>::ITEMS, Rs3Cfg<9, false, false>::S, RS3_MINB, false, false>,
                                                             RS_THREADS, (size_t)Rs3Cfg<9, false, false>::SPAN * 8) != cudaSuccess || nb <= 0) {
        cudaGetLastError();
        return -1;
    }
    g_rs3_ctas[dev] = nb;
    return 0;
}

namespace {
int rs3_ctas_per_sm() {
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return 1;
    if (g_rs3_ctas[dev] == 0) radix_sort_prepare_device();
    return g_rs3_ctas[dev] > 0 ? g_rs3_ctas[dev] : 1;
}
// grid and look-back rows of a pass3 sort over at most `upper` keys in spans of at most span_max keys
void rs3_shape(long long upper, int span_max, long long* grid, long long* rows) {
    const long long resident = (long long)num_sms() * rs3_ctas_per_sm();
    long long g = (upper + 2047) / 2048;  // never more CTAs than 2048-key spans
    if (g > resident) g = resident;
    if (g > 3 * 192) g = 3 * 192;  // what scratch_rows provides for
    if (g < 1) g = 1;
    const long long waves = (upper + g * span_max - 1) / (g * span_max);
    *grid = g;
    *rows = g * (waves > 0 ? waves : 1) + 1;
}
}  // namespace

// Look-back rows: the buffers are sized in rows of 512 words.  A 32-bit-key sort over max_items keys uses at
// most max_items / 7680 + resident CTAs + 1 rows of 256 or 512 words (rs3_shape), a 64-bit-key sort max_items / 3072
// rows of 256 words.
constexpr long long scratch_rows(long long max_items) { return max_items / 6144 + 3 * 192 + 4; }  // 3 resident CTAs per SM, <= 192 SMs

size_t sort_scratch_bytes(long long max_items, int /*key_bytes*/) {
    return (size_t)(8 * 256 + 64) * 4 + 2 * (size_t)rs_status_words(scratch_rows(max_items)) * 4;
}

void sort_scratch_carve(SortScratch* s, void* base, long long max_items, int /*key_bytes*/) {
    const long long rows = scratch_rows(max_items);
    uint32_t* p = reinterpret_cast<uint32_t*>(base);
    s->hist = p;
    s->tickets = p + 8 * 256;
    s->zero_bytes = (size_t)(8 * 256 + 64) * 4;
    s->status[0] = p + 8 * 256 + 64;
    s->status[1] = s->status[0] + (size_t)rs_status_words(rows);
    s->max_tiles = rows;
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

// words per look-back row of a sort: 512 when a pass has 9-bit digits, else 256
static int plan_row_stride(const SortOptions* opt) {
    if (opt && opt->n_pass > 0)
        for (int p = 0; p < opt->n_pass; ++p) if (opt->pass_bits[p] > 8) return 512;
    return 256;
}

template <typename KeyT>
void sort_scratch_clear(SortScratch* scratch, long long upper, bool hist_ready, cudaStream_t stream, const SortOptions* opt) {
    if (upper <= 0) return;
    if (hist_ready) {
        // the digit histograms come from the caller (k_preprocess / k_tile_scan): clear the tickets and the look-back
        // rows of the first pass only
        // 32-bit keys: the AGG / GROUP / CKPT rows of the first pass (k_rs_pass3); 64-bit keys: its look-back rows (k_rs_pass)
        size_t words;
        if (sizeof(KeyT) == 4) {
            long long g, rows;
            const int stride = plan_row_stride(opt);
            rs3_shape(upper, stride > 256 ? rs3_span(9) : rs3_span(8), &g, &rows);
            words = (size_t)rs_status_rows_total(rows) * stride;
        } else words = (size_t)((upper + RS_THREADS * RsCfg<KeyT>::LARGE - 1) / (RS_THREADS * RsCfg<KeyT>::LARGE)) * 256;
        cudaMemsetAsync(scratch->tickets, 0, 64 * sizeof(uint32_t), stream);
        cudaMemsetAsync(scratch->status[0], 0, words * sizeof(uint32_t), stream);
    } else {
        cudaMemsetAsync(scratch->hist, 0, scratch->zero_bytes, stream);  // histograms + tickets; k_rs_hist clears the rows
    }
}
template void sort_scratch_clear<uint32_t>(SortScratch*, long long, bool, cudaStream_t, const SortOptions*);
template void sort_scratch_clear<uint64_t>(SortScratch*, long long, bool, cudaStream_t, const SortOptions*);

template <typename KeyT>
int radix_sort_pairs(const KeyT* keys_first, KeyT* keys, KeyT* keys_tmp, uint32_t* vals, uint32_t* vals_tmp, bool vals_identity,
                     long long n_host, const uint32_t* n_dev, long long capacity, int begin_bit, int end_bit,
                     SortScratch* scratch, cudaStream_t stream, long long* launch_count, const SortOptions* opt) {
    const bool hist_ready = opt && opt->hist_ready;
    const bool vals_only_last = opt && opt->vals_only_last;
    const bool drop = opt && opt->drop_rebase;
    RsPasses passes;
    passes.n = 0;
    passes.direct_mask = opt ? opt->hist_direct_mask : 0u;
    if (opt && opt->n_pass > 0) {  // explicit digit plan
        int b = begin_bit;
        for (int p = 0; p < opt->n_pass && p < 8; ++p) {
            passes.shift[p] = b;
            passes.bits[p] = opt->pass_bits[p];
            b += opt->pass_bits[p];
            ++passes.n;
        }
    } else {
        for (int b = begin_bit; b < end_bit && passes.n < 8;) {
            passes.shift[passes.n] = b;
            passes.bits[passes.n] = (end_bit - b < 8) ? (end_bit - b) : 8;
            b += passes.bits[passes.n];
            ++passes.n;
        }
    }
    for (int p = passes.n; p < 8; ++p) { passes.shift[p] = 0; passes.bits[p] = 1; }
    if (passes.n == 0) return 0;
    const long long upper = n_dev ? capacity : n_host;
    if (upper <= 0) return 0;
    const int sms = num_sms();
    const int row_stride = plan_row_stride(opt);
    const int hist_stride = (opt && opt->hist_stride > 0) ? opt->hist_stride : 256;
    long long grid = 1, rows0 = 1;
    const int span_max = row_stride > 256 ? rs3_span(9) : rs3_span(8);  // one span size for all passes of a sort
    if (sizeof(KeyT) == 4) {
        rs3_shape(upper, span_max, &grid, &rows0);
    } else {
        const long long tiles_upper = (upper + RS_THREADS * RsCfg<KeyT>::LARGE - 1) / (RS_THREADS * RsCfg<KeyT>::LARGE);
        grid = tiles_upper < (long long)sms * 3 ? tiles_upper : (long long)sms * 3;
        if (grid < 1) grid = 1;
        rows0 = tiles_upper;
    }
    if (!(opt && opt->scratch_cleared)) sort_scratch_clear<KeyT>(scratch, upper, hist_ready, stream, opt);
    if (!hist_ready) {
        long long g = (upper + RS_THREADS * 8 - 1) / (RS_THREADS * 8);
        if (g > sms * 8) g = sms * 8;
        if (g < 1) g = 1;
        const long long clear_words = (sizeof(KeyT) == 4) ? rs_status_rows_total(rows0) * row_stride : rows0 * 256;
        k_rs_hist<KeyT><<<(unsigned)g, RS_THREADS, 0, stream>>>(keys_first, n_host, n_dev, capacity, passes, scratch->hist,
                                                               scratch->status[0], clear_words);
        if (launch_count) ++*launch_count;
    }
    const KeyT* kin = keys_first; KeyT* kout = keys_tmp; KeyT* kother = keys;
    uint32_t* vin = vals; uint32_t* vout = vals_tmp;
    for (int p = 0; p < passes.n; ++p) {
        if (vals_only_last && p == passes.n - 1) kout = nullptr;
        uint32_t* st_cur = scratch->status[p & 1];
        uint32_t* st_next = scratch->status[(p + 1) & 1];
        const bool identity = (p == 0 && vals_identity);
        const uint32_t* hist_p = scratch->hist + p * hist_stride;
        if constexpr (sizeof(KeyT) == 4) {
            const uint32_t* nd = (p > 0 && opt && opt->n_dev_rest) ? opt->n_dev_rest : n_dev;
            const int bits = passes.bits[p], shift = passes.shift[p];
            const RsRebase rb = {opt ? opt->k0 : 0u, opt ? opt->khi : 0xFFFFFFFFu};
#define RS3_GO(B, I, D) rs3_launch<B, I, D>((unsigned)grid, stream, kin, kout, vin, vout, n_host, nd, capacity, shift, bits, hist_p, \
                                           scratch->tickets + p, st_cur, st_next, row_stride, span_max, rows0, rb)
            const int nb = passes.bits[p];
            if (p == 0 && drop) {  // compacting first pass (identity values)
                if (nb > 8) RS3_GO(9, true, true); else RS3_GO(8, true, true);
            } else if (identity) {
                RS3_GO(8, true, false);  // (identity values with 9-bit digits only occur in the compacting pass above)
            } else if (nb > 8) RS3_GO(9, false, false);
            else if (nb == 8) RS3_GO(8, false, false);
            else if (nb == 7) RS3_GO(7, false, false);
            else if (nb == 6) RS3_GO(6, false, false);
            else RS3_GO(5, false, false);
#undef RS3_GO
        } else {
            launch_pass<KeyT, RsCfg<KeyT>::LARGE>(identity, (unsigned)grid, stream, kin, kout, vin, vout, n_host, n_dev, capacity,
                                                  passes.shift[p], passes.bits[p], hist_p, scratch->tickets + p, st_cur, st_next);
        }
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
// loads libgsr.so holds RS_STATIC_WORDS * 4 bytes (about 5.8 MB) until the process exits.  All calls on a device share
// the array; the sort clears what must start at zero (histograms + tickets by memset, the first look-back rows by the
// histogram kernel, the next pass's rows by each pass), so stale contents of an earlier call are harmless - but two
// calls that OVERLAP on one GPU (different streams) race on it.
constexpr long long RS_STATIC_MAX_ITEMS = 1ll << 22;  // pairs per call
constexpr size_t RS_STATIC_WORDS = (8 * 256 + 64) + 2 * (size_t)rs_status_words(scratch_rows(RS_STATIC_MAX_ITEMS));
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
    if ((size_t)((ss.status[1] + (size_t)rs_status_words(ss.max_tiles)) - base) > RS_STATIC_WORDS) { *err = "static scratch too small (internal)"; return -2; }
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
