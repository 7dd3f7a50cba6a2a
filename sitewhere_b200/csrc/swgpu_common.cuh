// swgpu_common.cuh — shared pieces of libswgpu.so: context, error plumbing, the single-pass
// ordered-compaction scan (decoupled look-back), byte readers over shared/global memory, hashes.
// sm_100a only; no CPU path.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <string>
#include <vector>

#include "../../include/swgpu.h"

// ---------------------------------------------------------------------------------------------
// context
struct ZoneTableHost;  // zone.cu

struct swgpu_ctx {
    int device = -1;
    int sm_count = 148;  // multiprocessors of the device (grid sizing of the ticketed kernels)
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    std::mutex mu;

    // scan scratch (tile states + ticket + totals), grown on demand
    void *scratch = nullptr;
    size_t scratch_bytes = 0;

    // staging buffers of the host-pointer entry points (device side), grown on demand
    struct Buf {
        void *p = nullptr;
        size_t cap = 0;
    };
    Buf d_bytes, d_off, d_status, d_first, d_cols, d_pairs, d_lon, d_lat, d_meta, d_ids;
    Buf d_zwork;  // zone test work area (parked counts / ids, tile totals / bases)
    Buf h_small;  // pinned, small results

    // zone table
    void *zone_blob = nullptr;  // device
    uint64_t zone_bytes = 0;
    void *zone_hdr = nullptr;  // host copy of the header (ZoneHeader)

    // device state (merge.cu)
    void *state = nullptr;
    void *dedup_state = nullptr;  // a second table of the same kind: the alternate ids seen (swgpu_dedup_probe_dev)
    uint64_t dedup_rows = 0;      // rows probed so far: earlier rows carry larger sequence numbers
    // chunked host-buffer pipeline (pipeline.cu)
    void *pipeline = nullptr;

    // swgpu_set_option
    uint32_t opt_zone_force_ray = 0;        // "zone.force_ray": 0 = choose per point, 1..4 = +x, -x, +y, -y (tests)
    uint32_t opt_zone_scan_sequential = 0;  // "zone.scan_sequential": 1 = single-CTA scan of the tile totals (tests)
    uint64_t opt_presence_now1 = 0;         // "state.presence_now_ms" + 1: folded events touch their device's presence row
    uint32_t opt_enriched_fast = 1;         // "enriched.fast_path": 0 = every enriched record takes the general parser (tests)
    uint32_t opt_json_fast = 1;             // "json.fast_path": 0 = no template recogniser, every record takes the general pass (tests)
    uint32_t opt_json_stats = 0;            // "json.stats": 1 = count the records the template recogniser takes
    uint32_t opt_zone_stats = 0;            // "zone.stats": 1 = count cell / edge records and straddle tests
    unsigned long long zone_stats[4] = {0, 0, 0, 0};  // of the last counted zone test (+ its location count)
};

#define SW_SET_ERR(ctx, ...)                         \
    do {                                             \
        char _b[512];                                \
        snprintf(_b, sizeof _b, __VA_ARGS__);        \
        if (ctx) (ctx)->err = _b;                    \
    } while (0)

#define SW_CUDA(ctx, call)                                                                         \
    do {                                                                                           \
        cudaError_t _e = (call);                                                                   \
        if (_e != cudaSuccess) {                                                                   \
            SW_SET_ERR(ctx, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return SWGPU_E_CUDA;                                                                   \
        }                                                                                          \
    } while (0)

int sw_ensure(swgpu_ctx *ctx, swgpu_ctx::Buf &b, size_t bytes);          // device buffer
int sw_ensure_scratch(swgpu_ctx *ctx, size_t bytes);
int sw_ensure_pinned(swgpu_ctx *ctx, swgpu_ctx::Buf &b, size_t bytes);   // pinned host buffer

// scratch layout: [0] ticket (u32) | [8] total (u64) | [16] aux (u64) | [64...] tile states (u64 each)
#define SW_SCRATCH_HDR 64

// ---------------------------------------------------------------------------------------------
// device helpers
#ifdef __CUDACC__

#define SW_FULL 0xffffffffu

__device__ __forceinline__ uint64_t sw_ld_volatile_u64(const uint64_t *p) {
    uint64_t v;
    // GPU-scope relaxed: served by L2, never by a stale L1 line; cheaper than .volatile (system scope)
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void sw_st_volatile_u64(uint64_t *p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// streaming 16-byte load that does not pollute L1 (payload bytes are read exactly once)
__device__ __forceinline__ uint4 sw_ld_stream_u4(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// ---- decoupled look-back: exclusive prefix of per-tile totals, tiles ordered by ticket ----------
// word = flag << 62 | value ; flag 0 = not ready, 1 = tile aggregate, 2 = inclusive prefix
#define SW_FLAG_AGG (1ull << 62)
#define SW_FLAG_PFX (2ull << 62)
#define SW_VAL_MASK ((1ull << 62) - 1)

#define SW_LB_K 8  // predecessors examined per lane per look-back round

// Called by all 32 lanes of ONE warp of the block.  Returns the exclusive prefix of `tile`.
__device__ __forceinline__ uint64_t sw_lookback(uint64_t *tile_state, uint32_t tile, uint64_t total) {
    const uint32_t lane = threadIdx.x & 31;
    if (tile == 0) {
        if (lane == 0) sw_st_volatile_u64(&tile_state[0], SW_FLAG_PFX | total);
        return 0;
    }
    if (lane == 0) sw_st_volatile_u64(&tile_state[tile], SW_FLAG_AGG | total);
    // Walk back 32 predecessors per round until one of them holds an inclusive prefix.  (Measured on
    // B200: wider windows and sleeping polls do not help — what a tile waits for is the slowest of the
    // few hundred tiles in flight, not the walk itself; see DESIGN.md "ordered placement".)
    uint64_t excl = 0;
    int64_t idx = (int64_t)tile - 1;
    for (;;) {
        const int64_t j = idx - (int64_t)lane;
        uint64_t s = (j >= 0) ? sw_ld_volatile_u64(&tile_state[j]) : SW_FLAG_PFX;
        while (__any_sync(SW_FULL, (s >> 62) == 0)) {
            if ((s >> 62) == 0) s = sw_ld_volatile_u64(&tile_state[j]);
        }
        const uint32_t pmask = __ballot_sync(SW_FULL, (s >> 62) == 2);
        const int first = pmask ? (__ffs(pmask) - 1) : 32;
        uint64_t c = ((int)lane <= first) ? (s & SW_VAL_MASK) : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(SW_FULL, c, o);
        excl += c;
        if (pmask) break;
        idx -= 32;
    }
    if (lane == 0) sw_st_volatile_u64(&tile_state[tile], SW_FLAG_PFX | (excl + total));
    return excl;
}

// Block-wide exclusive scan of one uint32 per thread + ordered global base through look-back.
// Returns (global exclusive prefix for this thread); *block_total_out gets the tile total.
// smem: needs THREADS/32 + 2 uint64 words.
template <int THREADS>
__device__ __forceinline__ uint64_t sw_block_ordered_offset(uint32_t count, uint64_t *tile_state,
                                                            uint32_t tile, uint64_t *s_warp /*[THREADS/32+2]*/,
                                                            uint64_t *tile_total_out) {
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(SW_FULL, inc, o);
        if (lane >= (uint32_t)o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    constexpr int NW = THREADS / 32;
    if (warp == 0) {
        uint64_t w = (lane < NW) ? s_warp[lane] : 0;
        uint64_t winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint64_t t = __shfl_up_sync(SW_FULL, winc, o);
            if (lane >= (uint32_t)o) winc += t;
        }
        uint64_t total = __shfl_sync(SW_FULL, winc, NW - 1);
        uint64_t base = sw_lookback(tile_state, tile, total);
        if (lane < NW) s_warp[lane] = winc - w;  // exclusive warp offsets
        if (lane == 0) {
            s_warp[NW] = base;
            s_warp[NW + 1] = total;
        }
    }
    __syncthreads();
    uint64_t r = s_warp[NW] + s_warp[warp] + (uint64_t)(inc - count);
    *tile_total_out = s_warp[NW + 1];
    return r;
}

// ---- hashes ------------------------------------------------------------------------------------------
// 64-bit row hash.  Low half: Kafka Utils.murmur2 (bit-identical: a row's shard is derivable from it).
// High half: an independent hash over the same little-endian 4-byte words murmur2 consumes (xor, multiply,
// rotate per word; the 1-3 tail bytes and the length folded in last), finished with fmix32.  One pass over
// the bytes feeds both; the high half is this repository's own definition (oracle: orc_hash64).
__host__ __device__ __forceinline__ uint32_t sw_hash_hi_step(uint32_t f, uint32_t word) {
    f = (f ^ word) * 0x9E3779B1u;
    return (f << 15) | (f >> 17);
}
__host__ __device__ __forceinline__ uint32_t sw_hash_hi_final(uint32_t f, uint32_t tail, uint32_t len) {
    f = sw_hash_hi_step(f, tail ^ (len << 24));
    f ^= f >> 16;
    f *= 0x85ebca6bu;
    f ^= f >> 13;
    f *= 0xc2b2ae35u;
    f ^= f >> 16;
    return f;
}
template <typename Rd>
__device__ __forceinline__ uint64_t sw_hash64(const Rd &rd, uint32_t off, uint32_t len) {
    const uint32_t m = 0x5bd1e995u;
    uint32_t h = 0x9747b28cu ^ len;
    uint32_t f = 0x811c9dc5u;
    uint32_t i = 0;
    // eight bytes per load while they last (one unaligned 64-bit read instead of eight byte reads)
    for (; i + 8 <= len; i += 8) {
        const uint64_t v = rd.le64(off + i);
#pragma unroll
        for (int half = 0; half < 2; half++) {
            uint32_t k = (uint32_t)(v >> (32 * half));
            f = sw_hash_hi_step(f, k);
            k *= m;
            k ^= k >> 24;
            k *= m;
            h *= m;
            h ^= k;
        }
    }
    for (; i + 4 <= len; i += 4) {
        uint32_t b0 = rd.byte(off + i), b1 = rd.byte(off + i + 1), b2 = rd.byte(off + i + 2),
                 b3 = rd.byte(off + i + 3);
        uint32_t k = b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
        f = sw_hash_hi_step(f, k);
        k *= m;
        k ^= k >> 24;
        k *= m;
        h *= m;
        h ^= k;
    }
    uint32_t rem = len - i, tail = 0;
    if (rem) {
        uint32_t b0 = rd.byte(off + i);
        uint32_t b1 = rem > 1 ? rd.byte(off + i + 1) : 0;
        uint32_t b2 = rem > 2 ? rd.byte(off + i + 2) : 0;
        tail = (b2 << 16) | (b1 << 8) | b0;
        h ^= tail;
        h *= m;
    }
    h ^= h >> 13;
    h *= m;
    h ^= h >> 15;
    return ((uint64_t)sw_hash_hi_final(f, tail, len) << 32) | h;
}

// ---- byte readers: the same parser runs over a shared-memory staged record or over global memory ----
struct SmemBytes {
    const uint8_t *s;  // record start inside the staged tile
    __device__ __forceinline__ uint32_t byte(uint32_t i) const { return s[i]; }
    __device__ __forceinline__ uint64_t le64(uint32_t i) const {
        // unaligned 8-byte little-endian read from three aligned words
        uint32_t a = (uint32_t)(uintptr_t)(s + i);
        const uint32_t *w = (const uint32_t *)(s + i - (a & 3));
        uint32_t sh = (a & 3) * 8;
        uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
        uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);
        return ((uint64_t)hi << 32) | lo;
    }
};
// either of the two behind one type (generic addressing), where one copy of a large parser matters more than
// LDS against LD: le64 reads the aligned words that hold the eight bytes, never one that holds none of them
struct AnyBytes {
    const uint8_t *s;
    __device__ __forceinline__ uint32_t byte(uint32_t i) const { return s[i]; }
    __device__ __forceinline__ uint64_t le64(uint32_t i) const {
        const uintptr_t a = (uintptr_t)(s + i);
        const uint32_t *w = (const uint32_t *)(a & ~(uintptr_t)3);
        const uint32_t sh = (uint32_t)(a & 3) * 8;
        const uint32_t w0 = w[0], w1 = w[1], w2 = sh ? w[2] : 0u;
        const uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);
        return ((uint64_t)hi << 32) | lo;
    }
};
struct GmemBytes {
    const uint8_t *s;
    __device__ __forceinline__ uint32_t byte(uint32_t i) const { return __ldg(s + i); }
    __device__ __forceinline__ uint64_t le64(uint32_t i) const {
        uint64_t v = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) v |= (uint64_t)__ldg(s + i + k) << (8 * k);
        return v;
    }
};

__device__ __forceinline__ uint64_t sw_mk_meta(uint32_t kind, uint32_t flags, uint32_t level,
                                               uint32_t source, uint32_t has_etype, uint32_t etype,
                                               uint32_t rec) {
    return (uint64_t)(kind & 0xff) | ((uint64_t)(flags & 0xffff) << 8) | ((uint64_t)(level & 3) << 24) |
           ((uint64_t)(source & 1) << 26) | ((uint64_t)(has_etype & 1) << 27) |
           ((uint64_t)(etype & 7) << 28) | ((uint64_t)rec << 32);
}
__device__ __forceinline__ uint64_t sw_mk_span(uint32_t off, uint32_t len) {
    return (uint64_t)off | ((uint64_t)len << 32);
}

#endif  // __CUDACC__
