// zone_walk.cuh — the packed zone table's layout and the per-location evaluation (grid cell -> zone records ->
// ring walk) of zone.cu (count / emit kernels, table builder), kept apart so that another kernel can test a
// location it holds in registers.  See zone.cu for the reference citations and the exactness argument.
#pragma once
#include <math.h>

#include "swgpu_common.cuh"

#define ZN_THREADS 256
#define ZN_MAXIN 8  // inside-set kept in registers per location; overflow recomputes (exact)
#define ZN_MAGIC 0x335a4e4f5a575355ull  // "USWZONZ3": table layout version 3

struct ZoneHeader {  // start of the packed table
    uint64_t magic, total_bytes;
    uint32_t n_tests, n_outside, grid_nx, grid_ny;
    double gx0, gy0, gx1, gy1;  // union of envelopes
    double inv_cx, inv_cy;
    uint64_t off_tests, off_verts, off_cell_start, off_cell_items, off_bin_start, off_bin_items, off_outside;
};
static_assert(sizeof(ZoneHeader) == 136, "ZoneHeader layout");

struct ZoneTestRec {  // 64 bytes, one per test (exact envelope first: the rectangle path reads it as one 256-bit load)
    double minx, maxx, miny, maxy;
    uint32_t vtx_off, npts;  // closed ring in verts[]
    uint32_t flags, nbins;   // flags: bit0 cond == Outside, bit1 isRectangle
    double inv_bin;
    uint32_t bin_off, pad;
};
static_assert(sizeof(ZoneTestRec) == 64, "ZoneTestRec layout");

// What one location needs in order to test one zone: 32 bytes = one 256-bit load = one L1 wavefront.
// The grid holds one of these per cell, so the common case (a cell overlapped by a single zone
// envelope) costs a single dependent load before the edges.  A cell overlapped by several envelopes
// holds a list head instead (ZC_LIST: `bin_off` = first entry in the list array, `t` = entries), an
// empty cell a record with ZC_EMPTY.
//   minx, miny, inv16  the slab functions: x-slab = zn_cell(px, minx, scale_x, nbins) and
//               y-slab = zn_cell(py, miny, scale_y, nbins).  Any scale works as long as registration and
//               lookup use the same one, so 16 bits of it are plenty.  The same two products, in slab
//               units, also pick the ray direction.
//   bin_off     first edge record of slab 0 (slabs `stride` records apart, each list ended by a
//               sentinel edge at -inf), or with ZC_RANGES the first {start,end} slab range.
// There is no envelope in the record: Geometry.contains' envelope short-circuit only ever answers
// "outside", and for a point outside the envelope the ring walk says the same (its slab is clamped to the
// first or last one, where no edge straddles the ray and the point is on no edge) after a single record.
struct ZnCellRec {
    double minx, miny;
    uint32_t inv16;  // slab scales as the top 16 bits of a float: low half for y-slabs, high half for x-slabs
    uint32_t bin_off;
    uint32_t t;
    uint32_t info;  // bits 0..7 flags, 8..19 slab count, 20..31 slab stride
};
static_assert(sizeof(ZnCellRec) == 32, "ZnCellRec layout");
#define ZC_OUTSIDE 1u   // alert condition == Outside
#define ZC_RECT 2u      // JTS isRectangle: RectangleContains on the exact envelope
#define ZC_RANGES 4u    // uneven slabs: {start,end} ranges instead of a fixed stride
#define ZC_XONLY 8u     // rays along x only (see build_zone_blob: translations p_i - p are not always exact)
#define ZC_EMPTY 0x40u  // no zone envelope overlaps the cell
#define ZC_LIST 0x80u   // list head
#define ZC_MAX_STRIDE 4095u

// monotone cell function shared by host (registration) and device (lookup); separate IEEE
// sub and mul (no contraction) so both sides round identically.
__host__ __device__ __forceinline__ double zn_cell_pos(double v, double v0, double inv) {
#ifdef __CUDA_ARCH__
    return __dmul_rn(__dsub_rn(v, v0), inv);
#else
    volatile double d = v - v0;
    volatile double t = d * inv;
    return t;
#endif
}
__host__ __device__ __forceinline__ uint32_t zn_cell_clamp(double t, uint32_t n) {
    if (!(t >= 0.0)) return 0;
    if (t >= (double)n) return n - 1;
    return (uint32_t)t;
}
__host__ __device__ __forceinline__ uint32_t zn_cell(double v, double v0, double inv, uint32_t n) {
    return zn_cell_clamp(zn_cell_pos(v, v0, inv), n);
}


// ---------------------------------------------------------------------------------------------
// device
struct ZnArgs {
    const uint8_t *blob;
    ZoneHeader h;
    const double *lon, *lat;
    const uint64_t *meta;
    uint32_t n;
    const uint32_t *n_rows_dev;
    uint64_t max_pairs;
    uint64_t *pairs;
    uint64_t *n_pairs;
    uint32_t *ticket;
    uint64_t *tile_state;
    const uint32_t *row_base;  // chunked pipeline: global index of row 0 (NULL = 0)
    uint32_t force_ray;        // test knob (swgpu_set_option "zone.force_ray"): 0 = choose, 1..4 = +x, -x, +y, -y
};

// exact sign of x1*y2 - y1*x2: products with their rounding errors (fma), compared as the
// monotone rounding of two reals; ties broken by the error terms.
// Rounding is monotone, so rounded products that differ already order the exact ones — even when a
// product over- or underflowed.  Only a tie needs the rounding errors, and those are exact (fma) only
// while the products are well inside the normal range; a tie outside it (0 == -0 from an underflow,
// inf == inf) takes the scaled comparison below.
static __device__ __noinline__ int zn_sign_det_scaled(double x1, double y1, double x2, double y2) {
    const int sp = ((x1 > 0) - (x1 < 0)) * ((y2 > 0) - (y2 < 0));
    const int sq = ((y1 > 0) - (y1 < 0)) * ((x2 > 0) - (x2 < 0));
    if (sp != sq) return sp > sq ? 1 : -1;
    if (sp == 0) return 0;
    // same sign: compare |x1*y2| with |y1*x2| as (double-double mantissa product, exponent sum)
    int ea, eb, ec, ed;
    const double ma = frexp(fabs(x1), &ea), md = frexp(fabs(y2), &ed);
    const double mb = frexp(fabs(y1), &eb), mc = frexp(fabs(x2), &ec);
    const double hp = __dmul_rn(ma, md), lp = __fma_rn(ma, md, -hp);  // in [0.25, 1): no underflow, exact
    double hq = __dmul_rn(mb, mc), lq = __fma_rn(mb, mc, -hq);
    const int d = (ea + ed) - (eb + ec);
    int cmp;
    if (d >= 2)
        cmp = 1;
    else if (d <= -2)
        cmp = -1;
    else {
        hq = ldexp(hq, -d);  // by 1/2, 1 or 2: exact
        lq = ldexp(lq, -d);
        cmp = hp > hq ? 1 : hp < hq ? -1 : (lp > lq) - (lp < lq);
    }
    return sp * cmp;
}
static __device__ __noinline__ int zn_sign_det(double x1, double y1, double x2, double y2) {
    double p = __dmul_rn(x1, y2), q = __dmul_rn(y1, x2);
    if (p > q) return 1;
    if (p < q) return -1;
    const double ap = fabs(p);
    if (!(ap >= 0x1p-960 && ap <= 0x1p1000)) return zn_sign_det_scaled(x1, y1, x2, y2);
    double e = __fma_rn(x1, y2, -p), f = __fma_rn(y1, x2, -q);
    return (e > f) - (e < f);
}

// 256-bit read-only load (sm_100: LDG.E.256): a whole 32-byte edge record / half a test record per request
struct __align__(32) ZnD4 {
    double x, y, z, w;
};
__device__ __forceinline__ ZnD4 zn_ld256(const void *p) {
    ZnD4 r;
    asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
    return r;
}

// RayCrossingCounter.locatePointInRing(p, ring) == INTERIOR restricted to the slab's edges.
// countSegment's tests (RayCrossingCounter.java:121-161), regrouped so that an edge that does not
// straddle the ray costs five comparisons instead of ten:
//   both ends left of p                        -> ignored (and, by the x ordering, so is the rest)
//   the ends are on different sides of y = py  -> the crossing test; a zero determinant is "on the
//       segment", which is also what the reference's earlier `p == p2` test reports for such an edge
//   otherwise p can only touch the edge if p2.y == py (then p1.y <= py): at p2, or on a horizontal edge
//
// The walk runs in the frame of the chosen copy of the ring (see the table builder): `u` is the
// coordinate along the ray (x, -x, y or -y), `v` the one across it, `slab` the copy's slab number.
// All four copies give the same answer for every point, so the choice is free.
struct ZnCnt {  // work counters of one location; only zone_count_kernel<true> reads them (dead code elsewhere)
    uint32_t cells, records, straddles;
};
__device__ __forceinline__ bool zn_ring_interior(const ZnArgs &a, uint32_t info, uint32_t bin_off, uint32_t slab,
                                                 double px, double py, ZnCnt &cnt) {
    const ZnD4 *er = (const ZnD4 *)(a.blob + a.h.off_bin_items);  // edge records, 32 B each (one sector)
    uint32_t k, k1;
    if (info & ZC_RANGES) {
        const uint2 br = __ldg((const uint2 *)(a.blob + a.h.off_bin_start) + bin_off + slab);
        k = br.x;
        k1 = br.y;
    } else {  // fixed stride; the slab's sentinel ends the loop
        k = bin_off + slab * (info >> 20);
        k1 = 0xffffffffu;
    }
    uint32_t crossings = 0;
    for (; k < k1; k++) {
        // one 256-bit load per edge record: every lane reads a different record, so the cost is L1
        // wavefronts per load instruction, not bytes
        const ZnD4 e = zn_ld256(er + k);  // p1 = (x,y)  p2 = (z,w)
        cnt.records++;
        if (e.x < px && e.z < px) break;  // slab edges are ordered by max x, descending: the rest are left too
        const bool up1 = e.y > py, up2 = e.w > py;
        if (up1 != up2) {
            cnt.straddles++;
            // sign of x1*y2 - y1*x2 on the translated coordinates, flipped when y2 < y1.  With the ends on
            // different sides, y2 < y1 is exactly `up1` (a difference of unequal doubles never rounds to 0),
            // and rounded products that differ already order the exact ones (see zn_sign_det).
            const double x1 = __dsub_rn(e.x, px), y1 = __dsub_rn(e.y, py);
            const double x2 = __dsub_rn(e.z, px), y2 = __dsub_rn(e.w, py);
            const double p = __dmul_rn(x1, y2), q = __dmul_rn(y1, x2);
            const bool gt = p > q, lt = p < q;
            if (gt | lt) {
                crossings += (gt != up1) ? 1u : 0u;
            } else {  // tie (or not comparable): the exact path
                int sg = zn_sign_det(x1, y1, x2, y2);
                if (sg == 0) return false;  // on the boundary
                if (up1) sg = -sg;
                if (sg > 0) crossings++;
            }
        } else if (e.w == py) {
            if (px == e.z) return false;  // p is the vertex p2
            if (e.y == py && px >= fmin(e.x, e.z) && px <= fmax(e.x, e.z)) return false;  // on a horizontal edge
        }
    }
    return crossings & 1;
}

// Visit the tests whose polygon contains (px,py), in ascending test order.
template <typename F>
__device__ __forceinline__ void zn_for_each_inside(const ZnArgs &a, double px, double py, ZnCnt &cnt, F f) {
    const ZoneHeader &h = a.h;
    if (!(px >= h.gx0 && px <= h.gx1 && py >= h.gy0 && py <= h.gy1)) return;  // outside every envelope (or NaN)
    uint32_t cx = zn_cell(px, h.gx0, h.inv_cx, h.grid_nx), cy = zn_cell(py, h.gy0, h.inv_cy, h.grid_ny);
    const ZnD4 *ct = (const ZnD4 *)(a.blob + h.off_cell_start);
    ZnD4 r = zn_ld256(ct + (cy * h.grid_nx + cx));
    uint32_t n = 1;
    const ZnD4 *lp = nullptr;
    if ((uint32_t)__double2hiint(r.w) & (ZC_LIST | ZC_EMPTY)) {
        if ((uint32_t)__double2hiint(r.w) & ZC_EMPTY) return;
        n = (uint32_t)__double2loint(r.w);
        lp = (const ZnD4 *)(a.blob + h.off_cell_items) + (uint32_t)__double2hiint(r.z);
        r = zn_ld256(lp);
    }
    for (uint32_t k = 0;;) {
        const uint32_t t = (uint32_t)__double2loint(r.w), info = (uint32_t)__double2hiint(r.w);
        cnt.cells++;
        bool in;
        if (info & ZC_RECT) {  // RectangleContains: inside the exact envelope and not on its border
            const ZnD4 ev = zn_ld256((const ZoneTestRec *)(a.blob + h.off_tests) + t);
            in = px > ev.x && px < ev.y && py > ev.z && py < ev.w;
        } else {
            // position of the point in slab units from the envelope's low corner (both axes have nb slabs):
            // gives the slab on either axis and the ray that leaves the envelope soonest
            const uint32_t nb = (info >> 8) & 0xfffu, i16 = (uint32_t)__double2loint(r.z);
            const double sy = (double)__int_as_float(i16 << 16), sx = (double)__int_as_float(i16 & 0xffff0000u);
            const double tl = zn_cell_pos(px, r.x, sx), tb = zn_cell_pos(py, r.y, sy);
            const double tr = (double)nb - tl, tt = (double)nb - tb;
            bool alongy = !(info & ZC_XONLY) && fmin(tb, tt) < fmin(tl, tr);
            bool neg = alongy ? tb < tt : tl < tr;
            if (a.force_ray) {  // tests only: walk the given copy whatever the flags say
                alongy = ((a.force_ray - 1u) >> 1) != 0;
                neg = ((a.force_ray - 1u) & 1u) != 0;
            }
            // ray along x: slabs by y (copies 0 .. 2nb-1); along y: slabs by x (copies 2nb .. 4nb-1)
            const uint32_t sl = alongy ? nb + zn_cell_clamp(tl, nb) : zn_cell_clamp(tb, nb);
            const double u = alongy ? py : px, v = alongy ? px : py;
            in = zn_ring_interior(a, info, (uint32_t)__double2hiint(r.z), 2u * sl + (neg ? 1u : 0u), neg ? -u : u, v, cnt);
        }
        if (in) f(t, info & ZC_OUTSIDE);
        if (++k >= n) break;
        r = zn_ld256(lp + k);
    }
}


// work area of one zone test: what the count step leaves for the scan and the emit step
struct ZnWork {
    uint64_t *park;       // [n] bits 0..1: alerts (3 = "see bits 2.."), then either two 31-bit test ids or the count
    uint64_t *tile_tot;   // [ntiles]
    uint64_t *tile_base;  // [ntiles]
    unsigned long long *stats;  // [3] cell records, edge records, straddle tests (zone_count_kernel<true> only)
};

// One zone test over `n` row slots: sw_zone_plan validates, sizes the work area and fills the arguments;
// sw_zone_run launches count -> scan -> emit.
struct SwZonePlan {
    ZnArgs a;
    ZnWork w;
    uint32_t ntiles;
};
int sw_zone_plan(swgpu_ctx *ctx, const double *lon, const double *lat, const uint64_t *meta, uint32_t n,
                 const uint32_t *n_rows_dev, uint64_t max_pairs, uint64_t *pairs, uint64_t *n_pairs_dev,
                 const uint32_t *row_base, SwZonePlan *pl);
int sw_zone_run(swgpu_ctx *ctx, const SwZonePlan &pl);
