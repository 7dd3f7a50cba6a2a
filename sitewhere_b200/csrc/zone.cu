// zone.cu — batched ZoneTestRuleProcessor.onLocation
// (reference: service-rule-processing/src/main/java/com/sitewhere/rules/processors/geospatial/
//  ZoneTestRuleProcessor.java:47-63; polygons per sitewhere-core/.../geospatial/GeoUtils.java:55-67;
//  containment = JTS 1.13 Polygon.contains(Point): envelope short-circuit, isRectangle ->
//  RectangleContains, else RayCrossingCounter with RobustDeterminant.signOfDet2x2).
//
// Results are bit-identical to the index-free reference loop; three exact accelerations remove work
// that cannot change the answer:
//   * a uniform grid over zone envelopes: Geometry.contains starts with an envelope test, so a
//     test whose envelope misses the point's grid cell is Outside.  Cell ids come from the same
//     monotone expression on host (registration) and device (lookup), so a point inside an
//     envelope always finds that zone in its cell.
//   * per-zone slabs across the ray: an edge whose range misses the point's slab is strictly on one
//     side of the ray; RayCrossingCounter.countSegment neither counts it nor flags a boundary.
//   * a choice of ray: every ring is stored for the four axis-parallel ray directions and the walk
//     takes the one that leaves the envelope soonest; the crossing parity and the on-boundary
//     verdict do not depend on the direction (see build_zone_blob).
// Crossing side = exact sign of x1*y2 - y1*x2 on the translated coordinates via error-free products
// (fma), the value RobustDeterminant returns.  Bound by instruction issue (walk lengths differ per
// lane); the zone table lives in L2.
#include <math.h>

#include <algorithm>

#include "swgpu_common.cuh"

#include "zone_walk.cuh"

// ---------------------------------------------------------------------------------------------
// host: build the packed table
static bool host_is_rectangle(const double *x, const double *y, uint32_t npts, double minx, double maxx,
                              double miny, double maxy) {
    if (npts != 5) return false;
    for (int i = 0; i < 5; i++) {
        if (!(x[i] == minx || x[i] == maxx)) return false;
        if (!(y[i] == miny || y[i] == maxy)) return false;
    }
    double px = x[0], py = y[0];
    for (int i = 1; i <= 4; i++) {
        bool xc = x[i] != px, yc = y[i] != py;
        if (xc == yc) return false;
        px = x[i];
        py = y[i];
    }
    return true;
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// slab scale kept as the top 16 bits of its float value (sign, exponent, 7 mantissa bits)
static uint32_t inv_bits16(double invd) {
    if (!(invd > 0) || !std::isfinite(invd)) return 0;
    float f = (float)invd;
    if (!std::isfinite(f)) return 0;
    uint32_t b;
    memcpy(&b, &f, 4);
    return b >> 16;
}
static double inv_from16(uint32_t b16) {
    uint32_t b = b16 << 16;
    float f;
    memcpy(&f, &b, 4);
    return (double)f;
}

static int build_zone_blob(swgpu_ctx *ctx, uint32_t T, const uint32_t *vtx_off, const double *lon,
                           const double *lat, const uint8_t *cond, std::vector<uint8_t> &blob) {
    std::vector<ZoneTestRec> tests(T);
    std::vector<double> verts;  // x,y interleaved
    std::vector<uint32_t> outside;
    uint64_t total_pts = 0;
    for (uint32_t t = 0; t < T; t++) {
        if (vtx_off[t + 1] < vtx_off[t]) {
            SW_SET_ERR(ctx, "swgpu_set_zones: vtx_off not monotone at test %u", t);
            return SWGPU_E_INVALID;
        }
        uint32_t nv = vtx_off[t + 1] - vtx_off[t];
        if (nv == 0) {  // GeoUtils would throw on locations.get(0)
            SW_SET_ERR(ctx, "swgpu_set_zones: test %u has no bounds", t);
            return SWGPU_E_INVALID;
        }
        if (nv < 3) {
            // GeoUtils.createPolygonForLocations closes the ring with the first bound: 1 or 2 bounds give 2 or 3
            // points, and JTS 1.13 LinearRing.validateConstruction throws IllegalArgumentException ("Invalid
            // number of points in LinearRing (found N - must be 0 or >= 4)").  In the reference every
            // onLocation then fails at this test (the polygon is never cached); here the table is refused.
            SW_SET_ERR(ctx, "swgpu_set_zones: test %u has %u bounds; JTS needs at least 3 (a closed ring of >= 4 points)",
                       t, nv);
            return SWGPU_E_INVALID;
        }
        total_pts += nv + 1;
    }
    if (total_pts > 0x7fffffffull) {
        SW_SET_ERR(ctx, "swgpu_set_zones: too many vertices");
        return SWGPU_E_INVALID;
    }
    verts.reserve(total_pts * 2);
    double gx0 = INFINITY, gy0 = INFINITY, gx1 = -INFINITY, gy1 = -INFINITY;
    std::vector<uint32_t> bin_ranges;  // {start,end} per slab, for tests with ZC_RANGES
    std::vector<double> bin_items;     // per slab: the edges that can touch it, as (x1,y1,x2,y2) records
    std::vector<ZnCellRec> crecs(T);
    for (uint32_t t = 0; t < T; t++) {
        ZoneTestRec &z = tests[t];
        memset(&z, 0, sizeof z);
        uint32_t a = vtx_off[t], b = vtx_off[t + 1], nv = b - a;
        z.vtx_off = (uint32_t)(verts.size() / 2);
        z.npts = nv + 1;
        for (uint32_t k = a; k < b; k++) {
            verts.push_back(lon[k]);
            verts.push_back(lat[k]);
        }
        verts.push_back(lon[a]);
        verts.push_back(lat[a]);
        const double *xy = &verts[(size_t)z.vtx_off * 2];
        double minx = xy[0], maxx = xy[0], miny = xy[1], maxy = xy[1];
        for (uint32_t i = 1; i < z.npts; i++) {
            double x = xy[2 * i], y = xy[2 * i + 1];
            if (x < minx) minx = x;
            if (x > maxx) maxx = x;
            if (y < miny) miny = y;
            if (y > maxy) maxy = y;
        }
        z.minx = minx;
        z.maxx = maxx;
        z.miny = miny;
        z.maxy = maxy;
        if (!(std::isfinite(minx) && std::isfinite(maxx) && std::isfinite(miny) && std::isfinite(maxy))) {
            SW_SET_ERR(ctx, "swgpu_set_zones: test %u has non-finite bounds", t);
            return SWGPU_E_INVALID;
        }
        std::vector<double> xs(z.npts), ys(z.npts);
        for (uint32_t i = 0; i < z.npts; i++) {
            xs[i] = xy[2 * i];
            ys[i] = xy[2 * i + 1];
            if (!std::isfinite(xs[i]) || !std::isfinite(ys[i])) {
                SW_SET_ERR(ctx, "swgpu_set_zones: test %u has a non-finite vertex", t);
                return SWGPU_E_INVALID;
            }
        }
        bool rect = host_is_rectangle(xs.data(), ys.data(), z.npts, minx, maxx, miny, maxy);
        z.flags = (cond[t] ? ZC_OUTSIDE : 0u) | (rect ? ZC_RECT : 0u);
        if (cond[t]) outside.push_back(t);
        gx0 = std::min(gx0, minx);
        gx1 = std::max(gx1, maxx);
        gy0 = std::min(gy0, miny);
        gy1 = std::max(gy1, maxy);
        // the compact record the grid cells carry
        ZnCellRec &cr = crecs[t];
        memset(&cr, 0, sizeof cr);
        cr.minx = minx;
        cr.miny = miny;
        cr.t = t;
        // slabs over edges i = 1..npts-1 (edge i joins ring[i] and ring[i-1])
        if (rect) {
            z.nbins = 0;
            z.inv_bin = 0;
            cr.info = z.flags;
        } else {
            // The ring is stored four times, once per axis-parallel ray direction; the kernel walks the
            // copy whose ray leaves the envelope soonest, which is the shortest walk on average:
            //   +x  edges as they are, cut into y-slabs, rightmost first.  RayCrossingCounter ignores an
            //       edge with both ends left of the point, so the walk stops at the first such edge.
            //   -x  the mirror image (x negated), again "rightmost" first: the same walk on (-px, py)
            //       counts the crossings LEFT of the point.
            //   +y, -y  the same two with the axes exchanged (x-slabs; the walk sees (py, px), (-py, px)).
            // A closed ring crosses any line an even number of times (with the walk's half-open rule for
            // vertices on the line), so all four parities agree, and an edge the point lies on is in the
            // point's slab and is met before the walk stops, whichever way it goes.
            //
            // That argument is about EXACT orientation predicates.  JTS 1.13 decides a crossing from the
            // coordinates translated in double (x1 = p1.x - p.x, ...): the sign it takes is exact for the
            // ROUNDED differences.  The +x and -x walks test the same edges (those that straddle y = p.y) with
            // the same determinant (negating x commutes with IEEE subtraction), so each such edge is classed
            // left / right / on identically and the two parities agree whatever the rounding did.  The +-y
            // walks test a different set of edges (those that straddle x = p.x): they agree with JTS's +x
            // answer for sure only when every difference is exact, i.e. the predicate is the true orientation.
            // By Sterbenz's lemma a - b is exact when a/2 <= b <= 2a; a point that reaches a sign test lies
            // inside the envelope, so if on each axis the envelope keeps one sign and max <= 2 min in
            // magnitude, every vertex - point difference is exact.  Only such zones (any geofence that does
            // not touch an axis and is smaller than its distance from it) use the +-y copies; the others are
            // flagged ZC_XONLY and walk along x only.
            const uint32_t nb = std::max(1u, std::min(64u, (z.npts - 1 + 3) / 4));
            const double *uc[2] = {xs.data(), ys.data()};  // axis 0: ray along x, slabs along y
            const double v0[2] = {miny, minx};  // slab origins
            const double vmax[2] = {maxy, maxx};
            uint32_t inv16[2];
            double inv[2];
            std::vector<std::vector<uint32_t>> bins[2];
            size_t refs = 0, longest = 0;
            for (int ax = 0; ax < 2; ax++) {
                inv16[ax] = inv_bits16((double)nb / (vmax[ax] - v0[ax]));
                inv[ax] = inv_from16(inv16[ax]);
                const double *vc = uc[1 - ax];
                bins[ax].assign(nb, std::vector<uint32_t>());
                for (uint32_t i = 1; i < z.npts; i++) {
                    double a1 = vc[i], a2 = vc[i - 1];
                    uint32_t lo = zn_cell(std::min(a1, a2), v0[ax], inv[ax], nb), hi = zn_cell(std::max(a1, a2), v0[ax], inv[ax], nb);
                    for (uint32_t bb = lo; bb <= hi; bb++) bins[ax][bb].push_back(i);
                }
                for (uint32_t bb = 0; bb < nb; bb++) {
                    refs += bins[ax][bb].size();
                    longest = std::max(longest, bins[ax][bb].size());
                }
            }
            z.nbins = nb;
            z.inv_bin = inv[0];
            cr.inv16 = inv16[0] | (inv16[1] << 16);
            auto sterbenz = [](double lo, double hi) {  // every a, b in [lo, hi]:  a/2 <= b <= 2a
                return (lo > 0 && hi <= 2 * lo) || (hi < 0 && lo >= 2 * hi);
            };
            if (!(sterbenz(minx, maxx) && sterbenz(miny, maxy))) z.flags |= ZC_XONLY;
            auto put_slab = [&](std::vector<uint32_t> &v, int ax, int side) {
                const double sg = side ? -1.0 : 1.0;
                const double *u = uc[ax], *vc = uc[1 - ax];
                std::sort(v.begin(), v.end(), [&](uint32_t e1, uint32_t e2) {
                    return std::max(sg * u[e1], sg * u[e1 - 1]) > std::max(sg * u[e2], sg * u[e2 - 1]);
                });
                for (uint32_t e : v) {  // edge record: p1 = ring[e], p2 = ring[e-1], 32 bytes
                    bin_items.push_back(sg * u[e]);
                    bin_items.push_back(vc[e]);
                    bin_items.push_back(sg * u[e - 1]);
                    bin_items.push_back(vc[e - 1]);
                }
            };
            // Slabs of similar length are stored a fixed stride apart, each ended by a sentinel edge
            // whose ends are at -inf ("both ends behind the point": the loop's own exit test), so the
            // kernel needs no slab range.  Very uneven slabs would waste the padding: they keep ranges.
            // Slab (ax, bb, side) is number (ax * nb + bb) * 2 + side.
            const size_t stride = longest + 1;
            const bool fixed = stride <= ZC_MAX_STRIDE && 2 * (size_t)nb * stride <= 4 * refs + 64 &&
                               bin_items.size() / 4 + 4 * (size_t)nb * stride < 0xffffffffull;
            if (fixed) {
                cr.bin_off = z.bin_off = (uint32_t)(bin_items.size() / 4);
                for (int ax = 0; ax < 2; ax++)
                    for (uint32_t bb = 0; bb < nb; bb++)
                        for (int side = 0; side < 2; side++) {
                            put_slab(bins[ax][bb], ax, side);
                            for (size_t k = bins[ax][bb].size(); k < stride; k++) {
                                bin_items.push_back(-INFINITY);
                                bin_items.push_back(0.0);
                                bin_items.push_back(-INFINITY);
                                bin_items.push_back(0.0);
                            }
                        }
                cr.info = z.flags | (nb << 8) | ((uint32_t)stride << 20);
            } else {
                cr.bin_off = z.bin_off = (uint32_t)(bin_ranges.size() / 2);
                for (int ax = 0; ax < 2; ax++)
                    for (uint32_t bb = 0; bb < nb; bb++)
                        for (int side = 0; side < 2; side++) {
                            bin_ranges.push_back((uint32_t)(bin_items.size() / 4));
                            put_slab(bins[ax][bb], ax, side);
                            bin_ranges.push_back((uint32_t)(bin_items.size() / 4));
                        }
                cr.info = z.flags | ZC_RANGES | (nb << 8);
            }
            if (bin_items.size() / 4 >= 0xffffffffull) {
                SW_SET_ERR(ctx, "swgpu_set_zones: edge table too large");
                return SWGPU_E_INVALID;
            }
        }
    }
    // grid over envelopes
    uint32_t g = 1;
    while ((uint64_t)g * g < (uint64_t)T * 4 && g < 1024) g *= 2;
    uint32_t nx = g, ny = g;
    std::vector<uint32_t> cell_start, cell_items;
    double inv_cx = 0, inv_cy = 0;
    for (;;) {
        double w = gx1 - gx0, h = gy1 - gy0;
        inv_cx = (T && w > 0 && std::isfinite((double)nx / w)) ? (double)nx / w : 0.0;
        inv_cy = (T && h > 0 && std::isfinite((double)ny / h)) ? (double)ny / h : 0.0;
        if (inv_cx == 0.0) nx = 1;
        if (inv_cy == 0.0) ny = 1;
        std::vector<uint32_t> cnt((size_t)nx * ny + 1, 0);
        uint64_t items = 0;
        for (uint32_t t = 0; t < T; t++) {
            const ZoneTestRec &z = tests[t];
            uint32_t x0 = zn_cell(z.minx, gx0, inv_cx, nx), x1 = zn_cell(z.maxx, gx0, inv_cx, nx);
            uint32_t y0 = zn_cell(z.miny, gy0, inv_cy, ny), y1 = zn_cell(z.maxy, gy0, inv_cy, ny);
            items += (uint64_t)(x1 - x0 + 1) * (y1 - y0 + 1);
        }
        if (items > (64ull << 20) && (nx > 1 || ny > 1)) {  // large overlapping zones: coarsen
            nx = std::max(1u, nx / 2);
            ny = std::max(1u, ny / 2);
            continue;
        }
        for (uint32_t t = 0; t < T; t++) {
            const ZoneTestRec &z = tests[t];
            uint32_t x0 = zn_cell(z.minx, gx0, inv_cx, nx), x1 = zn_cell(z.maxx, gx0, inv_cx, nx);
            uint32_t y0 = zn_cell(z.miny, gy0, inv_cy, ny), y1 = zn_cell(z.maxy, gy0, inv_cy, ny);
            for (uint32_t cy = y0; cy <= y1; cy++)
                for (uint32_t cx = x0; cx <= x1; cx++) cnt[(size_t)cy * nx + cx + 1]++;
        }
        for (size_t i = 1; i < cnt.size(); i++) cnt[i] += cnt[i - 1];
        cell_start = cnt;
        cell_items.assign(items, 0);
        std::vector<uint32_t> fill(cnt.begin(), cnt.end() - 1);
        for (uint32_t t = 0; t < T; t++) {  // ascending t => each cell list is in test order
            const ZoneTestRec &z = tests[t];
            uint32_t x0 = zn_cell(z.minx, gx0, inv_cx, nx), x1 = zn_cell(z.maxx, gx0, inv_cx, nx);
            uint32_t y0 = zn_cell(z.miny, gy0, inv_cy, ny), y1 = zn_cell(z.maxy, gy0, inv_cy, ny);
            for (uint32_t cy = y0; cy <= y1; cy++)
                for (uint32_t cx = x0; cx <= x1; cx++) cell_items[fill[(size_t)cy * nx + cx]++] = t;
        }
        break;
    }
    ZoneHeader h;
    memset(&h, 0, sizeof h);
    h.magic = ZN_MAGIC;
    h.n_tests = T;
    h.n_outside = (uint32_t)outside.size();
    h.grid_nx = nx;
    h.grid_ny = ny;
    h.gx0 = gx0;
    h.gy0 = gy0;
    h.gx1 = gx1;
    h.gy1 = gy1;
    h.inv_cx = inv_cx;
    h.inv_cy = inv_cy;
    size_t off = align_up(sizeof(ZoneHeader), 256);
    h.off_tests = off;
    off = align_up(off + tests.size() * sizeof(ZoneTestRec), 256);
    h.off_verts = off;
    off = align_up(off + verts.size() * 8, 256);
    // flattened for the kernel: as few dependent loads per location as the data allows.
    //   cell table   one ZnCellRec per grid cell: the only zone whose envelope overlaps the cell, or a
    //                record no point passes (empty cell), or the head of a list
    //   cell lists   ZnCellRec copies, in test order, for cells overlapped by several envelopes
    //   slab ranges  {start,end} per slab, only for tests with ZC_RANGES
    std::vector<ZnCellRec> cell_tab(cell_start.size() - 1), cell_list;
    for (size_t c = 0; c + 1 < cell_start.size(); c++) {
        const uint32_t k0 = cell_start[c], k1 = cell_start[c + 1];
        ZnCellRec &e = cell_tab[c];
        memset(&e, 0, sizeof e);
        if (k1 == k0) {
            e.info = ZC_EMPTY;
        } else if (k1 - k0 == 1) {
            e = crecs[cell_items[k0]];
        } else {
            e.info = ZC_LIST;
            e.bin_off = (uint32_t)cell_list.size();
            e.t = k1 - k0;
            for (uint32_t k = k0; k < k1; k++) cell_list.push_back(crecs[cell_items[k]]);
        }
    }
    h.off_cell_start = off;
    off = align_up(off + cell_tab.size() * sizeof(ZnCellRec), 256);
    h.off_cell_items = off;
    off = align_up(off + cell_list.size() * sizeof(ZnCellRec), 256);
    h.off_bin_start = off;
    off = align_up(off + bin_ranges.size() * 4, 256);
    h.off_bin_items = off;
    off = align_up(off + bin_items.size() * 8, 256);
    h.off_outside = off;
    off = align_up(off + outside.size() * 4, 256);
    h.total_bytes = off;
    blob.assign(off, 0);
    memcpy(blob.data(), &h, sizeof h);
    if (!tests.empty()) memcpy(blob.data() + h.off_tests, tests.data(), tests.size() * sizeof(ZoneTestRec));
    if (!verts.empty()) memcpy(blob.data() + h.off_verts, verts.data(), verts.size() * 8);
    if (!cell_tab.empty()) memcpy(blob.data() + h.off_cell_start, cell_tab.data(), cell_tab.size() * sizeof(ZnCellRec));
    if (!cell_list.empty())
        memcpy(blob.data() + h.off_cell_items, cell_list.data(), cell_list.size() * sizeof(ZnCellRec));
    if (!bin_ranges.empty()) memcpy(blob.data() + h.off_bin_start, bin_ranges.data(), bin_ranges.size() * 4);
    if (!bin_items.empty()) memcpy(blob.data() + h.off_bin_items, bin_items.data(), bin_items.size() * 8);
    if (!outside.empty()) memcpy(blob.data() + h.off_outside, outside.data(), outside.size() * 4);
    return SWGPU_OK;
}

void sw_zone_free(swgpu_ctx *ctx) {
    if (ctx->zone_blob) cudaFree(ctx->zone_blob);
    ctx->zone_blob = nullptr;
    ctx->zone_bytes = 0;
    delete (ZoneHeader *)ctx->zone_hdr;
    ctx->zone_hdr = nullptr;
}

static int zone_install(swgpu_ctx *ctx, const void *src, uint64_t bytes, cudaMemcpyKind kind) {
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    void *nb = nullptr;
    cudaError_t e = cudaMalloc(&nb, bytes);
    if (e != cudaSuccess) {
        SW_SET_ERR(ctx, "cudaMalloc(zone table %llu) failed: %s", (unsigned long long)bytes, cudaGetErrorString(e));
        return SWGPU_E_NOMEM;
    }
    e = cudaMemcpyAsync(nb, src, bytes, kind, ctx->stream);
    ZoneHeader *h = new ZoneHeader();
    if (e == cudaSuccess) e = cudaMemcpyAsync(h, nb, sizeof(ZoneHeader), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess || h->magic != ZN_MAGIC || h->total_bytes != bytes) {
        if (e != cudaSuccess)
            SW_SET_ERR(ctx, "zone table upload failed: %s", cudaGetErrorString(e));
        else
            SW_SET_ERR(ctx, "zone table blob is not a swgpu zone table (bad magic/size)");
        cudaFree(nb);
        delete h;
        return e != cudaSuccess ? SWGPU_E_CUDA : SWGPU_E_INVALID;
    }
    sw_zone_free(ctx);
    ctx->zone_blob = nb;
    ctx->zone_bytes = bytes;
    ctx->zone_hdr = h;
    return SWGPU_OK;
}

extern "C" int32_t swgpu_set_zones(swgpu_ctx *ctx, uint32_t n_tests, const uint32_t *vtx_off, const double *lon,
                                   const double *lat, const uint8_t *cond) {
    if (!ctx) return SWGPU_E_INVALID;
    if (n_tests && (!vtx_off || !lon || !lat || !cond)) {
        SW_SET_ERR(ctx, "swgpu_set_zones: NULL argument");
        return SWGPU_E_INVALID;
    }
    std::vector<uint8_t> blob;
    uint32_t zero_off[1] = {0};
    if (int rc = build_zone_blob(ctx, n_tests, n_tests ? vtx_off : zero_off, lon, lat, cond, blob)) return rc;
    return zone_install(ctx, blob.data(), blob.size(), cudaMemcpyHostToDevice);
}

extern "C" int32_t swgpu_zone_table_bytes(swgpu_ctx *ctx, uint64_t *out_bytes) {
    if (!ctx || !out_bytes) return SWGPU_E_INVALID;
    if (!ctx->zone_blob) {
        SW_SET_ERR(ctx, "no zone table set");
        return SWGPU_E_STATE;
    }
    *out_bytes = ctx->zone_bytes;
    return SWGPU_OK;
}
extern "C" int32_t swgpu_zone_table_export_dev(swgpu_ctx *ctx, void *dst_dev, uint64_t capacity) {
    if (!ctx || !dst_dev) return SWGPU_E_INVALID;
    if (!ctx->zone_blob) {
        SW_SET_ERR(ctx, "no zone table set");
        return SWGPU_E_STATE;
    }
    if (capacity < ctx->zone_bytes) {
        SW_SET_ERR(ctx, "zone table export: capacity %llu < %llu", (unsigned long long)capacity,
                   (unsigned long long)ctx->zone_bytes);
        return SWGPU_E_CAPACITY;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    SW_CUDA(ctx, cudaMemcpyAsync(dst_dev, ctx->zone_blob, ctx->zone_bytes, cudaMemcpyDeviceToDevice, ctx->stream));
    SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SWGPU_OK;
}
extern "C" int32_t swgpu_zone_table_import_dev(swgpu_ctx *ctx, const void *src_dev, uint64_t bytes) {
    if (!ctx || !src_dev || bytes < sizeof(ZoneHeader)) return SWGPU_E_INVALID;
    return zone_install(ctx, src_dev, bytes, cudaMemcpyDeviceToDevice);
}

// what one location evaluates to: the tests that contain it (ascending, with their condition bit) and the
// number of alerts the reference would fire for it
struct ZnLoc {
    double px, py;
    uint32_t ins[ZN_MAXIN];
    uint32_t n_in, count;
};

// is row i a location the reference would route to onLocation?
__device__ __forceinline__ bool zn_row_valid(const ZnArgs &a, uint32_t i, uint32_t n_live) {
    bool valid = i < n_live;
    if (valid && a.meta) {
        uint64_t m = __ldg(a.meta + i);
        uint32_t fl = SWGPU_META_FLAGS(m);
        valid = SWGPU_META_KIND(m) == SWGPU_KIND_LOCATION && (fl & SWGPU_F_HAS_V0) && (fl & SWGPU_F_HAS_V1);
    }
    return valid;
}

__device__ __forceinline__ void zn_eval_point(const ZnArgs &a, bool valid, double px, double py, ZnLoc &L) {
    const uint32_t T_out = a.h.n_outside;
    L.px = px;
    L.py = py;
    L.n_in = 0;
    uint32_t n_in_icond = 0, n_in_ocond = 0;
    if (valid) {
        ZnCnt cnt{0, 0, 0};
        zn_for_each_inside(a, L.px, L.py, cnt, [&](uint32_t t, uint32_t outside_cond) {
            if (L.n_in < ZN_MAXIN) L.ins[L.n_in] = (t << 1) | outside_cond;
            L.n_in++;
            if (outside_cond)
                n_in_ocond++;
            else
                n_in_icond++;
        });
    }
    // alerts: Inside-condition tests that contain the point + Outside-condition tests that do not
    L.count = valid ? (n_in_icond + (T_out - n_in_ocond)) : 0;
}

__device__ __forceinline__ void zn_eval(const ZnArgs &a, uint32_t i, uint32_t n_live, ZnLoc &L) {
    const bool valid = zn_row_valid(a, i, n_live);
    zn_eval_point(a, valid, valid ? __ldg(a.lon + i) : 0.0, valid ? __ldg(a.lat + i) : 0.0, L);
}

// writes the location's alert pairs, in test order, starting at `pos`
__device__ __forceinline__ void zn_emit(const ZnArgs &a, uint32_t i, const ZnLoc &L, uint64_t pos) {
    const uint32_t T_out = a.h.n_outside;
    const uint32_t n_in = L.n_in;
    const uint32_t *ins = L.ins;
    const double px = L.px, py = L.py;
    {
        const uint64_t ev = (uint64_t)((a.row_base ? __ldg(a.row_base) : 0u) + i) << 32;
        if (T_out == 0 && n_in <= ZN_MAXIN) {
            for (uint32_t k = 0; k < n_in; k++, pos++)
                if (pos < a.max_pairs) a.pairs[pos] = ev | (ins[k] >> 1);
        } else if (n_in <= ZN_MAXIN) {
            // merge (ascending): inside hits with Inside condition, and the Outside-condition list
            // minus the tests that contain the point
            const uint32_t *ol = (const uint32_t *)(a.blob + a.h.off_outside);
            uint32_t k = 0, j = 0;
            while (k < n_in || j < T_out) {
                uint32_t ti = k < n_in ? (ins[k] >> 1) : 0xffffffffu;
                uint32_t to = j < T_out ? __ldg(ol + j) : 0xffffffffu;
                if (ti < to) {
                    if (!(ins[k] & 1u)) {
                        if (pos < a.max_pairs) a.pairs[pos] = ev | ti;
                        pos++;
                    }
                    k++;
                } else if (to < ti) {
                    if (pos < a.max_pairs) a.pairs[pos] = ev | to;
                    pos++;
                    j++;
                } else {  // same test: Outside condition but the point is inside -> no alert
                    k++;
                    j++;
                }
            }
        } else {
            // more containing zones than the register set holds (rare): walk every test in order and
            // re-derive containment through the same visitor — slow, exact
            const ZoneTestRec *tests = (const ZoneTestRec *)(a.blob + a.h.off_tests);
            for (uint32_t t = 0; t < a.h.n_tests; t++) {
                bool in = false;
                ZnCnt cnt{0, 0, 0};
                zn_for_each_inside(a, px, py, cnt, [&](uint32_t tt, uint32_t) {
                    if (tt == t) in = true;
                });
                bool outside_cond = (__ldg(&tests[t].flags) & 1u) != 0;
                if (in != outside_cond) {
                    if (pos < a.max_pairs) a.pairs[pos] = ev | t;
                    pos++;
                }
            }
        }
    }
}

// ---- single-launch form (ticketed tiles + decoupled look-back).  Needs only 8 bytes of scratch per
// 256 locations; kept for swgpu_zone_test_noalloc_dev.  Every tile waits for the slowest tile in flight,
// so the three-launch form below is what the other entry points use.
__global__ void __launch_bounds__(ZN_THREADS) zone_test_kernel(ZnArgs a) {
    __shared__ uint64_t s_warp[ZN_THREADS / 32 + 2];
    __shared__ uint32_t s_tile;
    const uint32_t tid = threadIdx.x;
    if (tid == 0) s_tile = atomicAdd(a.ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t n_live = a.n_rows_dev ? min(__ldg(a.n_rows_dev), a.n) : a.n;
    const uint32_t i = tile * ZN_THREADS + tid;
    ZnLoc L;
    zn_eval(a, i, n_live, L);
    uint64_t tile_total;
    uint64_t pos = sw_block_ordered_offset<ZN_THREADS>(L.count, a.tile_state, tile, s_warp, &tile_total);
    // tiles are ordered by ticket, so the tile with the highest id is last: it publishes the total
    if (tile == gridDim.x - 1 && tid == 0) *a.n_pairs = pos + tile_total;  // thread 0's offset = tile base
    if (L.count) zn_emit(a, i, L, pos);
}

// ---- three-launch form: count -> scan of tile totals -> emit.  No tile ever waits for another one.
//   zone_count_kernel  evaluates every location once and parks one 8-byte word for it: in the common
//                      case (no Outside-condition tests, at most two alerts) the count and the test ids,
//                      otherwise just the count; writes the tile's total.
//   zone_scan_kernel   one CTA: exclusive scan of the tile totals (a few tens of thousands of words).
//   zone_emit_kernel   places each location's pairs at tile base + its offset inside the tile; locations
//                      whose ids were not parked (rare) are evaluated again.
// STATS = true: the same kernel, also summing how many cell records, edge records and straddle tests the
// batch took (bench.py reports the zone stage in those units; the timed kernel is STATS = false).
template <bool STATS>
__global__ void __launch_bounds__(ZN_THREADS, 6) zone_count_kernel(ZnArgs a, ZnWork w) {
    __shared__ uint32_t s_part[ZN_THREADS / 32];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n_live = a.n_rows_dev ? min(__ldg(a.n_rows_dev), a.n) : a.n;
    const uint32_t i = blockIdx.x * ZN_THREADS + tid;
    // only the count and the first two containing tests are kept here; a location with more alerts (or
    // with Outside-condition tests) is evaluated again by the emit kernel
    uint32_t id0 = 0, id1 = 0, n_in = 0, n_oc = 0, count = 0;
    ZnCnt cnt{0, 0, 0};
    if (zn_row_valid(a, i, n_live)) {
        zn_for_each_inside(a, __ldg(a.lon + i), __ldg(a.lat + i), cnt, [&](uint32_t t, uint32_t outside_cond) {
            if (n_in == 0) id0 = t;
            if (n_in == 1) id1 = t;
            n_in++;
            n_oc += outside_cond;
        });
        // alerts: Inside-condition tests that contain the point + Outside-condition tests that do not
        count = (n_in - n_oc) + (a.h.n_outside - n_oc);
    }
    if (i < a.n) {
        if (a.h.n_outside == 0 && count <= 2)  // test ids are below 2^31 (one vertex each at least)
            w.park[i] = (uint64_t)count | ((uint64_t)id0 << 2) | ((uint64_t)id1 << 33);
        else
            w.park[i] = 3u | ((uint64_t)count << 2);
    }
    if (STATS) {
        uint32_t c0 = cnt.cells, c1 = cnt.records, c2 = cnt.straddles;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            c0 += __shfl_xor_sync(SW_FULL, c0, o);
            c1 += __shfl_xor_sync(SW_FULL, c1, o);
            c2 += __shfl_xor_sync(SW_FULL, c2, o);
        }
        if (lane == 0) {
            atomicAdd(w.stats + 0, (unsigned long long)c0);
            atomicAdd(w.stats + 1, (unsigned long long)c1);
            atomicAdd(w.stats + 2, (unsigned long long)c2);
        }
    }
    uint32_t s = count;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(SW_FULL, s, o);
    if (lane == 0) s_part[warp] = s;
    __syncthreads();
    if (tid == 0) {
        uint64_t t = 0;
#pragma unroll
        for (int k = 0; k < ZN_THREADS / 32; k++) t += s_part[k];
        w.tile_tot[blockIdx.x] = t;
    }
}

#include "tile_scan.cuh"

#define ZE_PER 8
static_assert(ZE_PER * 32 == ZN_THREADS, "one warp of the emit kernel = one tile");
__global__ void __launch_bounds__(ZN_THREADS) zone_emit_kernel(ZnArgs a, ZnWork w) {
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t tile = (uint64_t)blockIdx.x * (ZN_THREADS / 32) + warp;
    const uint64_t i0 = (tile * 32 + lane) * ZE_PER;
    if (tile * ZN_THREADS >= a.n) return;  // whole warp
    uint64_t parked[ZE_PER];
    if (i0 + ZE_PER <= a.n) {
        const ZnD4 v0 = zn_ld256(w.park + i0), v1 = zn_ld256(w.park + i0 + 4);
        parked[0] = (uint64_t)__double_as_longlong(v0.x);
        parked[1] = (uint64_t)__double_as_longlong(v0.y);
        parked[2] = (uint64_t)__double_as_longlong(v0.z);
        parked[3] = (uint64_t)__double_as_longlong(v0.w);
        parked[4] = (uint64_t)__double_as_longlong(v1.x);
        parked[5] = (uint64_t)__double_as_longlong(v1.y);
        parked[6] = (uint64_t)__double_as_longlong(v1.z);
        parked[7] = (uint64_t)__double_as_longlong(v1.w);
    } else {
#pragma unroll
        for (int j = 0; j < ZE_PER; j++) parked[j] = i0 + j < a.n ? w.park[i0 + j] : 0;
    }
    uint32_t tsum = 0;
#pragma unroll
    for (int j = 0; j < ZE_PER; j++) tsum += (parked[j] & 3u) == 3u ? (uint32_t)(parked[j] >> 2) : (uint32_t)(parked[j] & 3u);
    uint32_t inc = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(SW_FULL, inc, o);
        if (lane >= (uint32_t)o) inc += t;
    }
    if (!tsum) return;
    uint64_t pos = w.tile_base[tile] + (inc - tsum);
    const uint32_t row_base = a.row_base ? __ldg(a.row_base) : 0u;
#pragma unroll
    for (int j = 0; j < ZE_PER; j++) {
        const uint32_t count = (parked[j] & 3u) == 3u ? (uint32_t)(parked[j] >> 2) : (uint32_t)(parked[j] & 3u);
        if (!count) continue;
        const uint32_t i = (uint32_t)i0 + j;
        if (a.h.n_outside == 0 && count <= 2) {
            const uint64_t ev = (uint64_t)(row_base + i) << 32;
            if (pos < a.max_pairs) a.pairs[pos] = ev | ((parked[j] >> 2) & 0x7fffffffu);
            if (count == 2 && pos + 1 < a.max_pairs) a.pairs[pos + 1] = ev | (parked[j] >> 33);
        } else {
            const uint32_t n_live = a.n_rows_dev ? min(__ldg(a.n_rows_dev), a.n) : a.n;
            ZnLoc L;
            zn_eval(a, i, n_live, L);
            zn_emit(a, i, L, pos);
        }
        pos += count;
    }
}

int sw_zone_plan(swgpu_ctx *ctx, const double *lon, const double *lat, const uint64_t *meta, uint32_t n,
                 const uint32_t *n_rows_dev, uint64_t max_pairs, uint64_t *pairs, uint64_t *n_pairs_dev,
                 const uint32_t *row_base, SwZonePlan *pl) {
    if (!ctx) return SWGPU_E_INVALID;
    if (!ctx->zone_blob) {
        SW_SET_ERR(ctx, "swgpu_zone_test: no zone table (call swgpu_set_zones first)");
        return SWGPU_E_STATE;
    }
    if (!n_pairs_dev || (n && (!lon || !lat)) || (max_pairs && !pairs)) {
        SW_SET_ERR(ctx, "swgpu_zone_test_dev: NULL argument");
        return SWGPU_E_INVALID;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    pl->ntiles = (n + ZN_THREADS - 1) / ZN_THREADS;
    ZnArgs &a = pl->a;
    a.blob = (const uint8_t *)ctx->zone_blob;
    a.h = *(const ZoneHeader *)ctx->zone_hdr;
    a.lon = lon;
    a.lat = lat;
    a.meta = meta;
    a.n = n;
    a.n_rows_dev = n_rows_dev;
    a.max_pairs = max_pairs;
    a.pairs = pairs;
    a.n_pairs = n_pairs_dev;
    a.ticket = nullptr;
    a.tile_state = nullptr;
    a.row_base = row_base;
    a.force_ray = ctx->opt_zone_force_ray;
    ZnWork &w = pl->w;
    w.park = w.tile_tot = w.tile_base = nullptr;
    w.stats = nullptr;
    if (n == 0) return SWGPU_OK;
    const uint32_t ntiles = pl->ntiles;
    // work area: one parked word (8 B) per location, totals and bases (8 B each) per tile
    const size_t off_tot = ((size_t)n * 8 + 255) / 256 * 256;
    const size_t off_base = off_tot + (size_t)ntiles * 8;
    const size_t off_stats = off_base + (size_t)ntiles * 8;
    const size_t work_bytes = off_stats + 32;
    if (int rc = sw_ensure(ctx, ctx->d_zwork, work_bytes)) return rc;
    w.stats = (unsigned long long *)((uint8_t *)ctx->d_zwork.p + off_stats);
    w.park = (uint64_t *)ctx->d_zwork.p;
    w.tile_tot = (uint64_t *)((uint8_t *)ctx->d_zwork.p + off_tot);
    w.tile_base = (uint64_t *)((uint8_t *)ctx->d_zwork.p + off_base);
    return SWGPU_OK;
}

int sw_zone_run(swgpu_ctx *ctx, const SwZonePlan &pl) {
    const ZnArgs &a = pl.a;
    const ZnWork &w = pl.w;
    const uint32_t ntiles = pl.ntiles;
    if (a.n == 0) {
        SW_CUDA(ctx, cudaMemsetAsync(a.n_pairs, 0, sizeof(uint64_t), ctx->stream));
        return SWGPU_OK;
    }
    if (ctx->opt_zone_stats) {
        SW_CUDA(ctx, cudaMemsetAsync(w.stats, 0, 24, ctx->stream));
        zone_count_kernel<true><<<ntiles, ZN_THREADS, 0, ctx->stream>>>(a, w);
        SW_CUDA(ctx, cudaMemcpyAsync(ctx->zone_stats, w.stats, 24, cudaMemcpyDeviceToHost, ctx->stream));
        ctx->zone_stats[3] = a.n;
    } else
        zone_count_kernel<false><<<ntiles, ZN_THREADS, 0, ctx->stream>>>(a, w);
    SW_CUDA(ctx, cudaGetLastError());
    // (option "zone.scan_sequential" forces the sequential kernel: lets the tests cover it without a 67 M batch)
    if (ntiles <= ZS_MC_MAX_TILES && !ctx->opt_zone_scan_sequential)
        zone_scan_mc_kernel<<<(ntiles + ZS_THREADS - 1) / ZS_THREADS, ZS_THREADS, 0, ctx->stream>>>(w.tile_tot, w.tile_base,
                                                                                                  ntiles, a.n_pairs);
    else
        zone_scan_kernel<<<1, ZS_THREADS, 0, ctx->stream>>>(w.tile_tot, w.tile_base, ntiles, a.n_pairs);
    SW_CUDA(ctx, cudaGetLastError());
    zone_emit_kernel<<<(ntiles + ZN_THREADS / 32 - 1) / (ZN_THREADS / 32), ZN_THREADS, 0, ctx->stream>>>(a, w);
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
}

// Single-launch form for small batches (launches are what costs there): zone_test_kernel with in-kernel ordered
// placement; `scratch` = SW_SCRATCH_HDR + 8 bytes per 256 locations of device memory the caller has cleared.
int sw_zone_single_launch(swgpu_ctx *ctx, const double *lon, const double *lat, const uint64_t *meta, uint32_t n,
                          const uint32_t *n_rows_dev, uint64_t max_pairs, uint64_t *pairs, uint64_t *n_pairs_dev,
                          void *scratch) {
    SwZonePlan pl;
    if (int rc = sw_zone_plan(ctx, lon, lat, meta, 0, n_rows_dev, max_pairs, pairs, n_pairs_dev, nullptr, &pl)) return rc;
    if (n == 0) {
        SW_CUDA(ctx, cudaMemsetAsync(n_pairs_dev, 0, sizeof(uint64_t), ctx->stream));
        return SWGPU_OK;
    }
    pl.a.n = n;
    pl.a.ticket = (uint32_t *)scratch;
    pl.a.tile_state = (uint64_t *)((uint8_t *)scratch + SW_SCRATCH_HDR);
    zone_test_kernel<<<(n + ZN_THREADS - 1) / ZN_THREADS, ZN_THREADS, 0, ctx->stream>>>(pl.a);
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
}

int sw_zone_launch(swgpu_ctx *ctx, const double *lon, const double *lat, const uint64_t *meta, uint32_t n,
                   const uint32_t *n_rows_dev, uint64_t max_pairs, uint64_t *pairs, uint64_t *n_pairs_dev,
                   const uint32_t *row_base) {
    SwZonePlan pl;
    if (int rc = sw_zone_plan(ctx, lon, lat, meta, n, n_rows_dev, max_pairs, pairs, n_pairs_dev, row_base, &pl)) return rc;
    return sw_zone_run(ctx, pl);
}

// Work counters of the last zone test run with option "zone.stats" = 1 (synchronises the stream).
extern "C" int32_t swgpu_zone_stats(swgpu_ctx *ctx, uint64_t out[4]) {
    if (!ctx || !out) return SWGPU_E_INVALID;
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k < 4; k++) out[k] = ctx->zone_stats[k];
    return SWGPU_OK;
}

extern "C" int32_t swgpu_zone_test_dev(swgpu_ctx *ctx, const double *lon, const double *lat, const uint64_t *meta,
                                       uint32_t n, const uint32_t *n_rows_dev, uint64_t max_pairs,
                                       uint64_t *pairs, uint64_t *n_pairs_dev) {
    return sw_zone_launch(ctx, lon, lat, meta, n, n_rows_dev, max_pairs, pairs, n_pairs_dev, nullptr);
}

extern "C" int32_t swgpu_zone_test(swgpu_ctx *ctx, const double *lon, const double *lat, const uint64_t *meta,
                                   uint32_t n, uint64_t max_pairs, uint64_t *pairs, uint64_t *n_pairs) {
    if (!ctx) return SWGPU_E_INVALID;
    if (!n_pairs || (n && (!lon || !lat)) || (max_pairs && !pairs)) {
        SW_SET_ERR(ctx, "swgpu_zone_test: NULL argument");
        return SWGPU_E_INVALID;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (int rc = sw_ensure(ctx, ctx->d_lon, (size_t)n * 8 + 16)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_lat, (size_t)n * 8 + 16)) return rc;
    if (meta)
        if (int rc = sw_ensure(ctx, ctx->d_meta, (size_t)n * 8 + 16)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_pairs, (size_t)max_pairs * 8 + 16)) return rc;
    if (n) {
        SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_lon.p, lon, (size_t)n * 8, cudaMemcpyHostToDevice, st));
        SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_lat.p, lat, (size_t)n * 8, cudaMemcpyHostToDevice, st));
        if (meta) SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_meta.p, meta, (size_t)n * 8, cudaMemcpyHostToDevice, st));
    }
    uint64_t *d_np = (uint64_t *)((uint8_t *)ctx->d_pairs.p + (size_t)max_pairs * 8);
    int rc = swgpu_zone_test_dev(ctx, (const double *)ctx->d_lon.p, (const double *)ctx->d_lat.p,
                                 meta ? (const uint64_t *)ctx->d_meta.p : nullptr, n, nullptr, max_pairs,
                                 (uint64_t *)ctx->d_pairs.p, d_np);
    if (rc) return rc;
    SW_CUDA(ctx, cudaMemcpyAsync(n_pairs, d_np, 8, cudaMemcpyDeviceToHost, st));
    SW_CUDA(ctx, cudaStreamSynchronize(st));
    uint64_t live = *n_pairs < max_pairs ? *n_pairs : max_pairs;
    if (live) {
        SW_CUDA(ctx, cudaMemcpyAsync(pairs, ctx->d_pairs.p, live * 8, cudaMemcpyDeviceToHost, st));
        SW_CUDA(ctx, cudaStreamSynchronize(st));
    }
    if (*n_pairs > max_pairs) {
        SW_SET_ERR(ctx, "zone test produced %llu alerts but max_pairs is %llu", (unsigned long long)*n_pairs,
                   (unsigned long long)max_pairs);
        return SWGPU_E_CAPACITY;
    }
    return SWGPU_OK;
}

// ---------------------------------------------------------------------------------------------
// Allocation-free zone test.
//
// swgpu_zone_test_dev grows the context's scan scratch with cudaMalloc when a batch is larger than
// any before it.  Some callers of this library must not allocate at the moment they run a batch:
//   * the rule-processing consumer thread shares the GPU with the decode and state-merge contexts of
//     the same process; cudaMalloc / cudaFree synchronise the whole device, so a late scratch growth
//     on this thread would stall every other stream's batches for its duration;
//   * deployments that size the device-state table to fill the memory budget up front have no room
//     left: a scratch growth in the middle of consuming a partition could fail with out-of-memory
//     after offsets were already committed.
// For them the scan scratch lives in this file-scope device array: it is part of the library image,
// every GPU that loads libswgpu.so holds its 512 KiB + 64 B until the process exits, and a call
// never allocates.  The price is a fixed upper bound on the batch (SWGPU_NOALLOC_MAX_LOCATIONS).
#define ZN_NOALLOC_TILES (SWGPU_NOALLOC_MAX_LOCATIONS / ZN_THREADS)
#define ZN_NOALLOC_BYTES (SW_SCRATCH_HDR + (size_t)ZN_NOALLOC_TILES * 8)
static __device__ __align__(256) unsigned char g_zn_noalloc_scratch[ZN_NOALLOC_BYTES];

// The array's name is only a host-side handle; kernels need its device address, which differs per
// device.  It is looked up once per device (with that device current) and cached here.
#define ZN_NOALLOC_MAX_DEVICES 64
static std::mutex g_zn_noalloc_mu;
static void *g_zn_noalloc_addr[ZN_NOALLOC_MAX_DEVICES] = {nullptr};

extern "C" int32_t swgpu_zone_test_noalloc_dev(swgpu_ctx *ctx, const double *lon, const double *lat,
                                               const uint64_t *meta, uint32_t n, const uint32_t *n_rows_dev,
                                               uint64_t max_pairs, uint64_t *pairs, uint64_t *n_pairs_dev) {
    if (!ctx) return SWGPU_E_INVALID;
    if (!ctx->zone_blob) {
        SW_SET_ERR(ctx, "swgpu_zone_test_noalloc_dev: no zone table (call swgpu_set_zones first)");
        return SWGPU_E_STATE;
    }
    if (!n_pairs_dev || (n && (!lon || !lat)) || (max_pairs && !pairs)) {
        SW_SET_ERR(ctx, "swgpu_zone_test_noalloc_dev: NULL argument");
        return SWGPU_E_INVALID;
    }
    if (n > SWGPU_NOALLOC_MAX_LOCATIONS) {
        SW_SET_ERR(ctx,
                   "swgpu_zone_test_noalloc_dev: %u locations exceed the fixed scratch (at most %u per call); "
                   "split the batch or use swgpu_zone_test_dev",
                   n, (unsigned)SWGPU_NOALLOC_MAX_LOCATIONS);
        return SWGPU_E_CAPACITY;
    }
    if (ctx->device < 0 || ctx->device >= ZN_NOALLOC_MAX_DEVICES) {
        SW_SET_ERR(ctx, "swgpu_zone_test_noalloc_dev: device ordinal %d not supported (max %d)", ctx->device,
                   ZN_NOALLOC_MAX_DEVICES - 1);
        return SWGPU_E_INVALID;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    void *scratch = nullptr;
    {
        // first call on this device: resolve the device address of the static array (set-up step)
        std::lock_guard<std::mutex> lk(g_zn_noalloc_mu);
        if (!g_zn_noalloc_addr[ctx->device]) {
            void *p = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&p, g_zn_noalloc_scratch);
            if (e != cudaSuccess || !p) {
                SW_SET_ERR(ctx, "cudaGetSymbolAddress(noalloc scratch) failed on device %d: %s", ctx->device,
                           cudaGetErrorString(e));
                return SWGPU_E_CUDA;
            }
            g_zn_noalloc_addr[ctx->device] = p;
        }
        scratch = g_zn_noalloc_addr[ctx->device];
    }
    if (n == 0) {
        SW_CUDA(ctx, cudaMemsetAsync(n_pairs_dev, 0, sizeof(uint64_t), ctx->stream));
        return SWGPU_OK;
    }
    // the array still holds the previous call's ticket and tile states: both must start at zero
    const uint32_t ntiles = (n + ZN_THREADS - 1) / ZN_THREADS;
    SW_CUDA(ctx, cudaMemsetAsync(scratch, 0, SW_SCRATCH_HDR + (size_t)ntiles * 8, ctx->stream));
    ZnArgs a;
    a.blob = (const uint8_t *)ctx->zone_blob;
    a.h = *(const ZoneHeader *)ctx->zone_hdr;
    a.lon = lon;
    a.lat = lat;
    a.meta = meta;
    a.n = n;
    a.n_rows_dev = n_rows_dev;
    a.max_pairs = max_pairs;
    a.pairs = pairs;
    a.n_pairs = n_pairs_dev;
    a.ticket = (uint32_t *)scratch;
    a.tile_state = (uint64_t *)((uint8_t *)scratch + SW_SCRATCH_HDR);
    a.row_base = nullptr;
    a.force_ray = ctx->opt_zone_force_ray;
    zone_test_kernel<<<ntiles, ZN_THREADS, 0, ctx->stream>>>(a);
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
}
