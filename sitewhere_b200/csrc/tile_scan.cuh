// tile_scan.cuh — exclusive scan of per-tile totals (64-bit), the middle step of every count -> scan -> emit
// placement in this library: the zone test's alert pairs (zone.cu) and the encoder's record bytes (encode.cu).
#pragma once
#include "swgpu_common.cuh"

#define ZS_THREADS 1024
#define ZS_ITEMS 8
static __global__ void __launch_bounds__(ZS_THREADS) zone_scan_kernel(const uint64_t *tot, uint64_t *base, uint32_t ntiles,
                                                               uint64_t *total_out) {
    __shared__ uint64_t s_warp[ZS_THREADS / 32];
    __shared__ uint64_t s_carry;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (uint32_t start = 0; start < ntiles; start += ZS_THREADS * ZS_ITEMS) {
        const uint32_t i0 = start + tid * ZS_ITEMS;
        uint64_t v[ZS_ITEMS], sum = 0;
#pragma unroll
        for (int k = 0; k < ZS_ITEMS; k++) {
            v[k] = (i0 + k < ntiles) ? tot[i0 + k] : 0;
            sum += v[k];
        }
        uint64_t inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint64_t t = __shfl_up_sync(SW_FULL, inc, o);
            if (lane >= (uint32_t)o) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint64_t wv = s_warp[lane], winc = wv;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint64_t t = __shfl_up_sync(SW_FULL, winc, o);
                if (lane >= (uint32_t)o) winc += t;
            }
            s_warp[lane] = winc - wv;  // exclusive warp offsets
        }
        __syncthreads();
        uint64_t run = s_carry + s_warp[warp] + (inc - sum);
#pragma unroll
        for (int k = 0; k < ZS_ITEMS; k++) {
            if (i0 + k < ntiles) base[i0 + k] = run;
            run += v[k];
        }
        __syncthreads();
        if (tid == ZS_THREADS - 1) s_carry = run;
        __syncthreads();
    }
    if (tid == 0) *total_out = s_carry;
}

// Exclusive scan of the tile totals on several CTAs without any wait between them: CTA j scans its own
// ZS_THREADS totals and simply re-adds all the totals before its range (j coalesced loads per thread, all
// L2 hits).  Quadratic in the number of CTAs, so sw_zone_launch uses it up to ZS_MC_MAX_TILES tiles
// (67 M locations) and the sequential kernel above beyond that.
#define ZS_MC_MAX_TILES 262144u
static __global__ void __launch_bounds__(ZS_THREADS) zone_scan_mc_kernel(const uint64_t *tot, uint64_t *base, uint32_t ntiles,
                                                                  uint64_t *total_out) {
    __shared__ uint64_t s_warp[ZS_THREADS / 32], s_acc[ZS_THREADS / 32];
    __shared__ uint64_t s_pre;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t start = blockIdx.x * ZS_THREADS;
    uint64_t acc = 0;
#pragma unroll 8
    for (uint32_t i = tid; i < start; i += ZS_THREADS) acc += tot[i];
    const uint32_t i = start + tid;
    const uint64_t v = i < ntiles ? tot[i] : 0;
    uint64_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint64_t t = __shfl_up_sync(SW_FULL, inc, o);
        if (lane >= (uint32_t)o) inc += t;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(SW_FULL, acc, o);
    if (lane == 31) s_warp[warp] = inc;
    if (lane == 0) s_acc[warp] = acc;
    __syncthreads();
    if (warp == 0) {
        uint64_t wv = s_warp[lane], winc = wv, av = s_acc[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint64_t t = __shfl_up_sync(SW_FULL, winc, o);
            if (lane >= (uint32_t)o) winc += t;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) av += __shfl_xor_sync(SW_FULL, av, o);
        s_warp[lane] = winc - wv;  // exclusive warp offsets
        if (lane == 0) s_pre = av;
    }
    __syncthreads();
    const uint64_t excl = s_pre + s_warp[warp] + (inc - v);
    if (i < ntiles) base[i] = excl;
    if (i == ntiles - 1) *total_out = excl + v;
}

// Each thread places the pairs of ZE_PER = 8 consecutive locations (two 256-bit loads of their parked words,
// both in flight together); a warp therefore covers exactly one 256-location tile: its offsets come from one
// warp scan, with no shared memory and no CTA barrier.
