// decode_pb.cu — batched ProtobufDeviceEventDecoder.decode
// (reference: service-event-sources/src/main/java/com/sitewhere/sources/decoder/protobuf/
//  ProtobufDeviceEventDecoder.java:75-304; schema sitewhere-communication/src/main/proto/sitewhere.proto).
//
// Two kernels, launched back to back.  pb_decode_dense_kernel is the optimistic pass: a warp takes four
// consecutive tiles of 32 records, each tile's contiguous bytes arrive in the warp's shared-memory slice by
// one TMA bulk copy (double-buffered, mbarrier completion), each lane walks one record (proto2 wire format,
// protobuf-java lite semantics) and writes its row at row index = record index; it raises a flag when some
// record yields 0 or several rows.  pb_decode_ordered_kernel returns at once unless that flag is up; then it
// redoes the batch with exact ordered placement (ticketed tiles + decoupled look-back).  HBM traffic =
// payload once + offsets + columns: the dense kernel is HBM-bound.
#include <algorithm>

#include "merge_table.cuh"
#include "pb_reader.cuh"
#include "swgpu_common.cuh"

#define PB_THREADS 256
#define PB_WARPS (PB_THREADS / 32)
#define PB_WSTAGE 4608  // bytes of payload staged per warp tile (32 records of up to 144 B on average)

struct PbArgs {
    const uint8_t *bytes;
    const uint64_t *rec_off;
    uint32_t n;
    uint32_t max_rows;
    swgpu_decoded_cols cols;
    uint8_t *status;
    uint32_t *first_row;
    uint32_t *out_rows;
    uint32_t *ticket;
    uint64_t *tile_state;
    // chunked pipeline: global index of record 0, running row total before / after this launch
    uint32_t rec_base;
    const uint32_t *row_base;
    uint32_t *row_next;
    uint32_t *not_dense;  // raised by the dense pass when some record does not yield exactly one row
    uint32_t force_ordered;  // small batches: no dense pass, the ordered pass decodes the batch (one launch)
    // fused decode + merge (pb_decode_dense_kernel<true>): the state table, per-RECORD event ids, sequence base
    MgTable mt;
    const uint64_t *id_hi, *id_lo;
    uint64_t seq_base;
};

struct PbOut {
    uint32_t status, count, cmd, flags;
    uint32_t hw_off, hw_len, orig_off, orig_len, name_off, name_len, msg_off, msg_len;
    uint32_t md_first, md_count;
    uint32_t body_pos, body_end;  // second message body (for re-walking extra measurements)
    int64_t date;
    uint64_t v0, v1, v2;  // raw bits
};

// Recogniser for the exact byte layouts ProtobufDeviceEventEncoder emits
// (ES/decoder/protobuf/ProtobufDeviceEventEncoder.java:60-176: Header{command}, then hardwareId, the event
// fields in field-number order and eventDate; every length below 128).  A record of that shape decodes
// to the same result through the generic wire walker below; this path only gets there in a few dozen
// instructions.  Anything else (originator, metadata, updateState, several measurements, long strings,
// unknown fields, damage) takes the generic path.
template <typename B>
__device__ __forceinline__ bool pb_fast_path(const B &bytes, uint32_t total, PbOut &o) {
    if (total < 15 || total > 131) return false;
    const uint64_t h = bytes.le64(0);  // 02 08 cmd LL 0A TL ..
    const uint32_t cmd = (uint32_t)(h >> 16) & 0xff, LL = (uint32_t)(h >> 24) & 0xff, TL = (uint32_t)(h >> 40) & 0xff;
    if ((h & 0xffff) != 0x0802u || ((h >> 32) & 0xff) != 0x0A || LL != total - 4 || (TL & 0x80)) return false;
    uint32_t p = 6 + TL;  // first byte after the token
    o.hw_off = 6;
    o.hw_len = TL;
    o.orig_off = o.orig_len = o.name_off = o.name_len = o.msg_off = o.msg_len = 0;
    o.md_first = o.md_count = 0;
    o.v0 = o.v1 = o.v2 = 0;
    o.body_pos = 4;
    o.body_end = total;
    if (cmd == 3) {
        if (LL != TL + 38) return false;
        if (bytes.byte(p) != 0x11 || bytes.byte(p + 9) != 0x19 || bytes.byte(p + 18) != 0x21 || bytes.byte(p + 27) != 0x29)
            return false;
        o.v0 = bytes.le64(p + 1);
        o.v1 = bytes.le64(p + 10);
        o.v2 = bytes.le64(p + 19);
        o.date = (int64_t)bytes.le64(p + 28);
        o.flags = SWGPU_F_HAS_DATE | SWGPU_F_HAS_V0 | SWGPU_F_HAS_V1 | SWGPU_F_HAS_V2;
    } else if (cmd == 5) {
        if (p + 4 > total || bytes.byte(p) != 0x12) return false;
        const uint32_t ML = bytes.byte(p + 1), NL = bytes.byte(p + 3);
        if (bytes.byte(p + 2) != 0x0A || ((NL | ML) & 0x80) || ML != NL + 11 || LL != TL + 2 + 2 + ML + 9) return false;
        const uint32_t q = p + 4 + NL;
        if (bytes.byte(q) != 0x11 || bytes.byte(q + 9) != 0x19) return false;
        o.name_off = p + 4;
        o.name_len = NL;
        o.v0 = bytes.le64(q + 1);
        o.date = (int64_t)bytes.le64(q + 10);
        o.flags = SWGPU_F_HAS_DATE | SWGPU_F_HAS_NAME | SWGPU_F_HAS_V0;
    } else if (cmd == 4) {
        if (p + 2 > total || bytes.byte(p) != 0x12) return false;
        const uint32_t AL = bytes.byte(p + 1);
        if (AL & 0x80) return false;
        const uint32_t q = p + 2 + AL;
        if (q + 2 > total || bytes.byte(q) != 0x1A) return false;
        const uint32_t ML = bytes.byte(q + 1);
        if ((ML & 0x80) || LL != TL + 2 + AL + 2 + ML + 2 + 9) return false;
        const uint32_t d = q + 2 + ML;
        if (bytes.byte(d) != 0x21) return false;
        o.name_off = p + 2;
        o.name_len = AL;
        o.msg_off = q + 2;
        o.msg_len = ML;
        o.date = (int64_t)bytes.le64(d + 1);
        o.flags = SWGPU_F_HAS_DATE | SWGPU_F_HAS_NAME | SWGPU_F_HAS_MESSAGE;
    } else
        return false;
    o.status = SWGPU_ST_OK;
    o.cmd = cmd;
    o.count = 1;
    return true;
}

template <typename B>
__device__ void pb_parse_record(B bytes, uint32_t total, PbOut &o) {
    if (pb_fast_path(bytes, total, o)) return;
    PbReader<B> r;
    r.b = bytes;
    o.count = 0;
    uint32_t spos = 0;
    if (!r.open_delimited(spos, total)) {
        o.status = SWGPU_ST_RUNTIME_EXCEPTION;  // header == null -> NPE
        return;
    }
    bool has_cmd = false;
    uint32_t cmd = 0;
    bool has_orig = false;
    o.orig_off = o.orig_len = 0;
    while (!r.err) {
        uint32_t tag = r.read_tag();
        if (r.err || tag == 0) break;
        if (tag == 8) {
            uint32_t raw = r.varint32();
            if (!r.err && raw >= 1 && raw <= 8) {
                has_cmd = true;
                cmd = raw;
            }
        } else if (tag == 18) {
            r.lenfield(o.orig_off, o.orig_len);
            has_orig = true;
        } else if (!r.unknown(tag))
            break;
    }
    if (r.err || r.last_tag != 0 || !has_cmd) {
        o.status = SWGPU_ST_DECODE_EXCEPTION;
        return;
    }
    if (cmd != 3 && cmd != 4 && cmd != 5) {
        o.status = SWGPU_ST_PASSTHROUGH;
        return;
    }
    if (!r.open_delimited(spos, total)) {
        o.status = SWGPU_ST_RUNTIME_EXCEPTION;  // message == null -> NPE
        return;
    }
    o.body_pos = r.pos;
    o.body_end = r.end;
    const uint32_t t_date = cmd == 3 ? 41u : (cmd == 4 ? 33u : 25u);
    const uint32_t t_md = cmd == 3 ? 50u : (cmd == 4 ? 42u : 34u);
    const uint32_t t_us = cmd == 3 ? 56u : (cmd == 4 ? 48u : 40u);
    bool has_hw = false, has_lat = false, has_lon = false, has_date = false, us = false;
    bool has_type = false, has_msg = false;
    uint32_t nmx = 0;
    o.hw_off = o.hw_len = o.name_off = o.name_len = o.msg_off = o.msg_len = 0;
    o.md_first = o.md_count = 0;
    o.date = 0;
    o.v0 = o.v1 = o.v2 = 0;
    while (!r.err) {
        uint32_t tag_pos = r.pos;
        uint32_t tag = r.read_tag();
        if (r.err || tag == 0) break;
        if (tag == 10) {
            r.lenfield(o.hw_off, o.hw_len);
            has_hw = true;
        } else if (tag == t_date) {
            o.date = (int64_t)r.fixed64();
            has_date = true;
        } else if (tag == t_md) {
            if (o.md_count == 0) o.md_first = tag_pos;
            o.md_count++;
            r.metadata_entry();
        } else if (tag == t_us) {
            us = r.varint64() != 0;
        } else if (cmd == 3 && tag == 17) {
            o.v0 = r.fixed64();
            has_lat = true;
        } else if (cmd == 3 && tag == 25) {
            o.v1 = r.fixed64();
            has_lon = true;
        } else if (cmd == 3 && tag == 33) {
            o.v2 = r.fixed64();
        } else if (cmd == 4 && tag == 18) {
            r.lenfield(o.name_off, o.name_len);
            has_type = true;
        } else if (cmd == 4 && tag == 26) {
            r.lenfield(o.msg_off, o.msg_len);
            has_msg = true;
        } else if (cmd == 5 && tag == 18) {
            uint32_t no = 0, nl = 0;
            uint64_t vb = 0;
            r.measurement(no, nl, vb);
            if (!r.err) {
                if (nmx == 0) {
                    o.name_off = no;
                    o.name_len = nl;
                    o.v0 = vb;
                }
                nmx++;
            }
        } else if (!r.unknown(tag))
            break;
    }
    bool ok = !r.err && r.last_tag == 0 && has_hw;
    if (cmd == 3) ok = ok && has_lat && has_lon;
    if (cmd == 4) ok = ok && has_type && has_msg;
    if (!ok) {
        o.status = SWGPU_ST_DECODE_EXCEPTION;
        return;
    }
    o.status = SWGPU_ST_OK;
    o.cmd = cmd;
    uint32_t f = 0;
    if (us) f |= SWGPU_F_UPDATE_STATE;
    if (has_date) f |= SWGPU_F_HAS_DATE;
    if (has_orig) f |= SWGPU_F_HAS_ORIGINATOR;
    if (o.md_count) f |= SWGPU_F_HAS_METADATA;
    if (cmd == 3) f |= SWGPU_F_HAS_V0 | SWGPU_F_HAS_V1 | SWGPU_F_HAS_V2;
    if (cmd == 4) f |= SWGPU_F_HAS_NAME | SWGPU_F_HAS_MESSAGE;
    if (cmd == 5) f |= SWGPU_F_HAS_NAME | SWGPU_F_HAS_V0;
    o.flags = f;
    o.count = cmd == 5 ? nmx : 1u;
}

// k-th (k >= 1) Measurement of an already validated DeviceMeasurements body
template <typename B>
__device__ void pb_nth_measurement(B bytes, const PbOut &o, uint32_t k, uint32_t &noff, uint32_t &nlen,
                                   uint64_t &vbits) {
    PbReader<B> r;
    r.b = bytes;
    r.pos = o.body_pos;
    r.end = o.body_end;
    r.limit = 0xffffffffu;
    r.err = false;
    r.last_tag = 0;
    uint32_t seen = 0;
    while (!r.err) {
        uint32_t tag = r.read_tag();
        if (r.err || tag == 0) break;
        if (tag == 18) {
            r.measurement(noff, nlen, vbits);
            if (seen == k) return;
            seen++;
        } else if (tag == 34) {
            r.metadata_entry();
        } else if (tag == 10) {
            uint32_t a, b2;
            r.lenfield(a, b2);
        } else if (tag == 25) {
            r.fixed64();
        } else if (tag == 40) {
            r.varint64();
        } else if (!r.unknown(tag))
            break;
    }
}

template <typename B>
__device__ __forceinline__ void pb_write_rows(const PbArgs &a, B bytes, const PbOut &o, uint32_t rec,
                                              uint64_t row0) {
    const uint32_t kind = o.cmd == 3 ? SWGPU_KIND_LOCATION : (o.cmd == 4 ? SWGPU_KIND_ALERT : SWGPU_KIND_MEASUREMENT);
    const uint64_t meta = sw_mk_meta(kind, o.flags, 0, 0, 1, kind, a.rec_base + rec);
    const uint64_t th = sw_hash64(bytes, o.hw_off, o.hw_len);
    const uint64_t tspan = sw_mk_span(o.hw_off, o.hw_len);
    const uint64_t ospan = (o.flags & SWGPU_F_HAS_ORIGINATOR) ? sw_mk_span(o.orig_off, o.orig_len) : 0;
    const uint64_t mdspan = o.md_count ? sw_mk_span(o.md_first, o.md_count) : 0;
    for (uint32_t k = 0; k < o.count; k++) {
        uint64_t row = row0 + k;
        if (row >= a.max_rows) return;
        uint32_t noff = o.name_off, nlen = o.name_len;
        uint64_t v0 = o.v0;
        if (k > 0) pb_nth_measurement(bytes, o, k, noff, nlen, v0);
        a.cols.meta[row] = meta;
        a.cols.token_hash[row] = th;
        a.cols.name_hash[row] = (kind == SWGPU_KIND_LOCATION) ? 0 : sw_hash64(bytes, noff, nlen);
        a.cols.event_date[row] = o.date;
        a.cols.v0[row] = __longlong_as_double((long long)v0);
        a.cols.v1[row] = __longlong_as_double((long long)o.v1);
        a.cols.v2[row] = __longlong_as_double((long long)o.v2);
        a.cols.token_span[row] = tspan;
        a.cols.name_span[row] = (kind == SWGPU_KIND_LOCATION) ? 0 : sw_mk_span(noff, nlen);
        if (a.cols.originator_span) a.cols.originator_span[row] = ospan;
        if (a.cols.message_span)
            a.cols.message_span[row] = kind == SWGPU_KIND_ALERT ? sw_mk_span(o.msg_off, o.msg_len) : 0;
        if (a.cols.metadata_span) a.cols.metadata_span[row] = mdspan;
        if (a.cols.alternate_id_span) a.cols.alternate_id_span[row] = 0;
    }
}

// Fused decode + merge: the record's rows go straight into the device-state table (merge_table.cuh) instead of
// the columns.  Row k of record `rec` has sequence number seq_base + (rec << SWGPU_SEQ_RECORD_SHIFT) + min(k,
// 4095): it does not depend on how many rows other records yield, so no ordered placement is needed.
template <typename B>
__device__ __forceinline__ void pb_merge_rows(const PbArgs &a, MgLocalCnt *local, B bytes, const PbOut &o, uint32_t rec) {
    const uint32_t kind = o.cmd == 3 ? SWGPU_KIND_LOCATION : (o.cmd == 4 ? SWGPU_KIND_ALERT : SWGPU_KIND_MEASUREMENT);
    const uint32_t kindbyte = mg_kindbyte(kind, o.flags);
    const uint64_t th = sw_hash64(bytes, o.hw_off, o.hw_len);
    const uint64_t ihi = __ldg(a.id_hi + rec), ilo = __ldg(a.id_lo + rec);
    const uint64_t seq0 = a.seq_base + ((uint64_t)(a.rec_base + rec) << SWGPU_SEQ_RECORD_SHIFT) + 1;
    for (uint32_t k = 0; k < o.count; k++) {
        uint32_t noff = o.name_off, nlen = o.name_len;
        uint64_t v0 = o.v0;
        if (k > 0) pb_nth_measurement(bytes, o, k, noff, nlen, v0);
        const uint32_t nl = kind == SWGPU_KIND_LOCATION ? 0u : nlen;
        if (o.hw_len > 0xffffu || nl > 0xffffu) {
            atomicAdd(&a.mt.ctr->key_too_long, 1ull);
            continue;
        }
        const uint64_t nh = kind == SWGPU_KIND_LOCATION ? 0ull : sw_hash64(bytes, noff, nlen);
        MgKeyBytes<B> key{bytes, o.hw_off, o.hw_len, noff, nl};
        mg_upsert(a.mt, local, mg_mix(th, nh, kindbyte), th, o.hw_len, nl, kindbyte, key,
                  seq0 + (k < (1u << SWGPU_SEQ_RECORD_SHIFT) - 1 ? k : (1u << SWGPU_SEQ_RECORD_SHIFT) - 1), ihi, ilo,
                  (uint64_t)o.date, v0);
    }
}

// ---- ordered pass (exact redo).  A CTA owns a tile of 256 consecutive records (its position in the ordered
// scan); inside it every WARP stages the contiguous bytes of its own 32 records into its private slice of
// shared memory (coalesced 16-byte streaming loads, __syncwarp only) and each lane walks one record.  The only
// block barriers are the ticket and the two of the ordered scan.  Rows are placed by ticketed tiles +
// decoupled look-back; every tile waits for the slowest of the few hundred tiles in flight, which is why this
// pass is ~2.5x slower than the dense one (measured; DESIGN.md "ordered placement") and only runs when the
// dense pass found a record that does not decode to exactly one row.
__device__ __forceinline__ void pb_decode_tile(const PbArgs &a, const uint32_t tile, uint8_t *smem, uint64_t *s_warp) {
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t rec0_64 = (uint64_t)tile * PB_THREADS + warp * 32;
    const uint32_t rec0 = rec0_64 < a.n ? (uint32_t)rec0_64 : a.n;
    const uint32_t nrec = min(32u, a.n - rec0);  // 0 for the padding warps of the last CTA
    uint8_t *ws = smem + warp * (PB_WSTAGE + 32);
    uint64_t r0 = 0, r1 = 0;
    if (lane < nrec) {
        r0 = a.rec_off[rec0 + lane];
        r1 = a.rec_off[rec0 + lane + 1];
    }
    const uint64_t b0 = __shfl_sync(SW_FULL, r0, 0);
    const uint64_t b1 = __shfl_sync(SW_FULL, r1, nrec ? nrec - 1 : 0);
    const uintptr_t addr = (uintptr_t)(a.bytes + b0);
    const uint32_t lead = (uint32_t)(addr & 15);
    const uint64_t span = (uint64_t)lead + (b1 - b0);
    const bool staged = span <= PB_WSTAGE;
    if (staged && nrec) {
        const uint4 *src = (const uint4 *)(addr - lead);
        uint4 *dst = (uint4 *)ws;
        const uint32_t nvec = (uint32_t)((span + 15) >> 4);
        for (uint32_t i = lane; i < nvec; i += 32) dst[i] = sw_ld_stream_u4(src + i);
    }
    __syncwarp();
    PbOut o;
    o.count = 0;
    o.status = 0;
    if (lane < nrec) {
        const uint64_t l64 = r1 - r0;
        const uint32_t len = l64 > 0x7fffffffull ? 0x7fffffffu : (uint32_t)l64;
        if (staged)
            pb_parse_record(SmemBytes{ws + lead + (uint32_t)(r0 - b0)}, len, o);
        else
            pb_parse_record(GmemBytes{a.bytes + r0}, len, o);
    }
    uint64_t tile_total = 0;
    const uint64_t row0 = sw_block_ordered_offset<PB_THREADS>(o.count, a.tile_state, tile, s_warp, &tile_total);
    const uint64_t row_base = a.row_base ? (uint64_t)__ldg(a.row_base) : 0ull;  // rows of earlier chunks
    if (lane < nrec) {
        const uint32_t rec = rec0 + lane;
        a.status[rec] = (uint8_t)o.status;
        a.first_row[rec] = (uint32_t)(row_base + row0);
        if (o.count != 0) {
            if (staged)
                pb_write_rows(a, SmemBytes{ws + lead + (uint32_t)(r0 - b0)}, o, rec, row0);
            else
                pb_write_rows(a, GmemBytes{a.bytes + r0}, o, rec, row0);
        }
    }
    if ((uint64_t)tile * PB_THREADS + PB_THREADS >= a.n && tid == 0) {  // the tile holding the last record
        const uint64_t total = row0 + tile_total;  // thread 0's offset is the tile base
        a.first_row[a.n] = (uint32_t)(row_base + total);
        *a.out_rows = (uint32_t)total;
        if (a.row_next) *a.row_next = (uint32_t)(row_base + total);
    }
}

// A few CTAs per SM take tiles by ticket (tickets are handed out in start order, which is what the look-back
// needs), so when the dense pass was exact — the usual case — the launch costs a few hundred CTAs that read
// one flag, not one CTA per tile.
__global__ void __launch_bounds__(PB_THREADS) pb_decode_ordered_kernel(PbArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];  // PB_WARPS * (PB_WSTAGE + 32)
    __shared__ uint64_t s_warp[PB_THREADS / 32 + 2];
    __shared__ uint32_t s_tile;
    if (!a.force_ordered && __ldg(a.not_dense) == 0) return;  // the dense pass was exact: nothing to redo
    const uint32_t ntiles = (uint32_t)(((uint64_t)a.n + PB_THREADS - 1) / PB_THREADS);
    for (;;) {
        if (threadIdx.x == 0) s_tile = atomicAdd(a.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= ntiles) return;
        pb_decode_tile(a, tile, smem, s_warp);
        __syncthreads();  // s_tile, s_warp and the staging buffers are reused
    }
}

// ---- dense pass, pipelined.  Same work per warp-tile (32 records) as pb_decode_tile, but a warp goes
// through PD_TILES consecutive tiles with the NEXT tile's payload already on its way into the other half
// of its shared-memory slice (one TMA bulk copy per tile, completion on an mbarrier) and the offsets of the tile after that
// already requested, so the two serial round trips at the head of a tile (offsets, then payload: half of
// the stall samples of the one-tile-per-warp kernel) are paid once per warp instead of once per tile.
#ifndef PB_MIN_CTAS
#define PB_MIN_CTAS 3  // resident CTAs per SM asked of the compiler (at most 80 registers): 3 CTAs for the plain pass as before, and the fused merge's random table accesses want warps in flight (128 registers and 2 CTAs without it: 1.63 against 1.54 ms per 10 M events)
#endif
#define PD_TILES 4
#define PD_STAGE 3072  // bytes staged per warp tile (32 records of up to 96 B on average); larger tiles read global memory

// One bulk copy (TMA unit, `cp.async.bulk`: SASS UBLKCP) per tile, issued by one lane, completion counted
// in bytes on an mbarrier in shared memory: no per-lane copy loop, no registers or LSU slots spent on staging.
__device__ __forceinline__ uint32_t pd_saddr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void pd_bar_init(uint64_t *bar) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(pd_saddr(bar)) : "memory");
}
__device__ __forceinline__ void pd_bulk_load(uint8_t *dst_smem, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(pd_saddr(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     pd_saddr(dst_smem)),
                 "l"(src), "r"(bytes), "r"(pd_saddr(bar))
                 : "memory");
}
__device__ __forceinline__ void pd_bar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t spin = 0; !done; spin++) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(pd_saddr(bar)), "r"(parity)
            : "memory");
        if (spin > (1u << 26)) __trap();  // a copy that never lands is a bug: fail loudly instead of hanging
    }
}

// A tile in flight is just this lane's record offsets; everything else (record count, first byte, alignment
// lead, whether it fits the stage) is recomputed from them when needed — a few shuffles and integer ops
// instead of a dozen live registers per tile.
struct PdTile {
    uint64_t r0, r1;  // this lane's record: [r0, r1) in the payload
};
struct PdGeom {
    uint64_t b0;  // first byte of the warp tile
    uint32_t rec0, nrec, lead;
    bool staged;
};
__device__ __forceinline__ uint32_t pd_nrec(const PbArgs &a, uint32_t wt, uint32_t nwt) {
    return wt < nwt ? min(32u, a.n - wt * 32u) : 0u;
}
__device__ __forceinline__ void pd_load_offsets(const PbArgs &a, uint32_t wt, uint32_t nwt, uint32_t lane, PdTile &t) {
    t.r0 = t.r1 = 0;
    if (lane < pd_nrec(a, wt, nwt)) {
        t.r0 = __ldg(a.rec_off + wt * 32u + lane);
        t.r1 = __ldg(a.rec_off + wt * 32u + lane + 1);
    }
}
__device__ __forceinline__ PdGeom pd_geom(const PbArgs &a, uint32_t wt, uint32_t nwt, const PdTile &t, uint64_t &span) {
    PdGeom g;
    g.nrec = pd_nrec(a, wt, nwt);
    g.rec0 = wt < nwt ? wt * 32u : a.n;
    g.b0 = __shfl_sync(SW_FULL, t.r0, 0);
    const uint64_t b1 = __shfl_sync(SW_FULL, t.r1, g.nrec ? g.nrec - 1 : 0);
    g.lead = (uint32_t)((uintptr_t)(a.bytes + g.b0) & 15);
    span = (uint64_t)g.lead + (b1 - g.b0);
    g.staged = span <= PD_STAGE;
    return g;
}
// request the tile's bytes (needs the offsets); true if a copy is now in flight on `bar`
__device__ __forceinline__ bool pd_issue(const PbArgs &a, uint32_t wt, uint32_t nwt, uint32_t lane, const PdTile &t,
                                         uint8_t *buf, uint64_t *bar) {
    uint64_t span;
    const PdGeom g = pd_geom(a, wt, nwt, t, span);
    const bool go = g.staged && g.nrec;
    if (go && lane == 0) {
        const uint8_t *src = a.bytes + g.b0 - g.lead;
        pd_bulk_load(buf, src, (uint32_t)((span + 15) >> 4) << 4, bar);
    }
    return go;
}

// MERGE = false: rows -> columns (swgpu_decode_pb[_dev]).  MERGE = true: rows -> device-state table
// (swgpu_decode_merge[_dev]): no columns, no first_row, no ordered redo; *out_rows counts the decoded rows.
template <bool MERGE>
__global__ void __launch_bounds__(PB_THREADS, PB_MIN_CTAS) pb_decode_dense_kernel(PbArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];  // PB_WARPS * 2 * (PD_STAGE + 32)
    __shared__ __align__(8) uint64_t s_bar[PB_WARPS][2];
    __shared__ MgLocalCnt s_cnt[PB_WARPS];
    uint32_t my_rows = 0;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t nwt = (a.n + 31u) / 32u;  // warp tiles in the batch
    // MERGE: newest records first (CTAs start roughly in index order, so CTA 0 takes the END of the batch and
    // each warp walks its tiles backwards): for a key with several rows in the batch the winner tends to arrive
    // first and the older rows lose their atomicMax at once, instead of each raising the maximum in turn
    const uint32_t blk = MERGE ? gridDim.x - 1 - blockIdx.x : blockIdx.x;
    const uint32_t first = (blk * PB_WARPS + (MERGE ? PB_WARPS - 1 - warp : warp)) * PD_TILES;
    if (first >= nwt) return;
    const uint32_t cnt = min((uint32_t)PD_TILES, nwt - first);  // tiles of this warp
    auto tile_of = [&](uint32_t k) -> uint32_t { return k < cnt ? (MERGE ? first + cnt - 1 - k : first + k) : nwt; };
    uint8_t *const slice = smem + (size_t)(warp * 2) * (PD_STAGE + 32);
    uint64_t *const bar = s_bar[warp];
    if (lane == 0) {
        pd_bar_init(&bar[0]);
        pd_bar_init(&bar[1]);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        s_cnt[warp].keys = s_cnt[warp].bytes = 0;
    }
    __syncwarp();
    const uint64_t row_base = a.row_base ? (uint64_t)__ldg(a.row_base) : 0ull;  // rows of earlier chunks
    PdTile cur, nxt, nn;
    pd_load_offsets(a, tile_of(0), nwt, lane, cur);
    pd_load_offsets(a, tile_of(1), nwt, lane, nxt);
    uint32_t parity = 0;                                           // bit b: phase of buffer b's barrier
    bool cur_copy = pd_issue(a, tile_of(0), nwt, lane, cur, slice, &bar[0]);
#pragma unroll 1
    for (uint32_t k = 0; k < cnt; k++) {
        const uint32_t wt = tile_of(k);
        // tile k+1: bytes on their way (the warp's last tile has no successor of its own); tile k+2: offsets requested
        const uint32_t b = k & 1;
        const bool nxt_copy = pd_issue(a, tile_of(k + 1), nwt, lane, nxt, slice + (b ^ 1) * (PD_STAGE + 32), &bar[b ^ 1]);
        pd_load_offsets(a, tile_of(k + 2), nwt, lane, nn);
        if (cur_copy) {
            pd_bar_wait(&bar[b], (parity >> b) & 1u);
            parity ^= 1u << b;
        }
        uint64_t span;
        const PdGeom g = pd_geom(a, wt, nwt, cur, span);
        const uint8_t *ws = slice + b * (PD_STAGE + 32);
        if (lane < g.nrec) {
            PbOut o;
            o.count = 0;
            o.status = 0;
            const uint64_t l64 = cur.r1 - cur.r0;
            const uint32_t len = l64 > 0x7fffffffull ? 0x7fffffffu : (uint32_t)l64;
            if (g.staged)
                pb_parse_record(SmemBytes{ws + g.lead + (uint32_t)(cur.r0 - g.b0)}, len, o);
            else
                pb_parse_record(GmemBytes{a.bytes + cur.r0}, len, o);
            const uint32_t rec = g.rec0 + lane;
            a.status[rec] = (uint8_t)o.status;
            if (MERGE) {
                my_rows += o.count;
                if (o.count) {
                    if (g.staged)
                        pb_merge_rows(a, &s_cnt[warp], SmemBytes{ws + g.lead + (uint32_t)(cur.r0 - g.b0)}, o, rec);
                    else
                        pb_merge_rows(a, &s_cnt[warp], GmemBytes{a.bytes + cur.r0}, o, rec);
                }
            } else {
                const uint64_t row0 = rec;  // row index = record index
                if (o.count != 1) atomicOr(a.not_dense, 1u);
                a.first_row[rec] = (uint32_t)(row_base + row0);
                if (o.count == 1) {
                    if (g.staged)
                        pb_write_rows(a, SmemBytes{ws + g.lead + (uint32_t)(cur.r0 - g.b0)}, o, rec, row0);
                    else
                        pb_write_rows(a, GmemBytes{a.bytes + cur.r0}, o, rec, row0);
                }
                if (rec == a.n - 1) {  // the last record of the batch
                    a.first_row[a.n] = (uint32_t)(row_base + a.n);
                    *a.out_rows = a.n;
                    if (a.row_next) *a.row_next = (uint32_t)(row_base + a.n);
                }
            }
        }
        __syncwarp();  // every lane is done reading this buffer: it is refilled two tiles from now
        cur = nxt;
        nxt = nn;
        cur_copy = nxt_copy;
    }
    if (MERGE) {
        my_rows = __reduce_add_sync(SW_FULL, my_rows);
        if (lane == 0) {
            if (my_rows) atomicAdd(a.out_rows, my_rows);
            mg_flush_counts(a.mt, &s_cnt[warp]);
        }
    }
}

static int check_cols(swgpu_ctx *ctx, const swgpu_decoded_cols *c) {
    if (!c || !c->meta || !c->token_hash || !c->name_hash || !c->event_date || !c->v0 || !c->v1 || !c->v2 ||
        !c->token_span || !c->name_span) {
        SW_SET_ERR(ctx, "decoded columns: the nine mandatory column pointers must be non-NULL");
        return SWGPU_E_INVALID;
    }
    return SWGPU_OK;
}

// small = true (batches of a few thousand records, where launches are what costs): the caller has sized and cleared the
// scratch (SW_SCRATCH_HDR + one word per 256 records); only the ordered pass is launched and it decodes the batch.
int sw_decode_pb_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint32_t max_rows,
                        const swgpu_decoded_cols *out, uint8_t *status, uint32_t *first_row, uint32_t *out_rows_dev,
                        uint32_t rec_base, const uint32_t *row_base, uint32_t *row_next, bool small) {
    if (!ctx) return SWGPU_E_INVALID;
    if (int rc = check_cols(ctx, out)) return rc;
    if (!rec_off || !first_row || !out_rows_dev || (n && !status)) {
        SW_SET_ERR(ctx, "swgpu_decode_pb_dev: NULL argument");
        return SWGPU_E_INVALID;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    if (n == 0) {
        SW_CUDA(ctx, cudaMemsetAsync(first_row, 0, sizeof(uint32_t), ctx->stream));
        SW_CUDA(ctx, cudaMemsetAsync(out_rows_dev, 0, sizeof(uint32_t), ctx->stream));
        return SWGPU_OK;
    }
    const uint32_t ntiles = (n + PB_THREADS - 1) / PB_THREADS;             // CTAs
    const size_t sbytes = SW_SCRATCH_HDR + (size_t)ntiles * 8;              // one scan state per tile
    if (!small) {
        if (int rc = sw_ensure_scratch(ctx, sbytes)) return rc;
        SW_CUDA(ctx, cudaMemsetAsync(ctx->scratch, 0, sbytes, ctx->stream));
    }
    PbArgs a;
    a.force_ordered = small ? 1u : 0u;
    a.bytes = bytes;
    a.rec_off = rec_off;
    a.n = n;
    a.max_rows = max_rows;
    a.cols = *out;
    a.status = status;
    a.first_row = first_row;
    a.out_rows = out_rows_dev;
    a.ticket = (uint32_t *)ctx->scratch;
    a.tile_state = (uint64_t *)((uint8_t *)ctx->scratch + SW_SCRATCH_HDR);
    a.rec_base = rec_base;
    a.row_base = row_base;
    a.row_next = row_next;
    a.not_dense = (uint32_t *)ctx->scratch + 1;
    const int smem = PB_WARPS * (PB_WSTAGE + 32);
    {
        // per device, once: the attribute is a property of the loaded function
        static std::mutex mu;
        static bool done[64] = {false};
        std::lock_guard<std::mutex> lk(mu);
        const int d = ctx->device & 63;
        if (!done[d]) {
            SW_CUDA(ctx, cudaFuncSetAttribute(pb_decode_ordered_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            SW_CUDA(ctx, cudaFuncSetAttribute(pb_decode_dense_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              PB_WARPS * 2 * (PD_STAGE + 32)));
            SW_CUDA(ctx, cudaFuncSetAttribute(pb_decode_dense_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              PB_WARPS * 2 * (PD_STAGE + 32)));
            done[d] = true;
        }
    }
    if (!small) {
        // optimistic: rows = records.  Pipelined kernel: PD_TILES warp tiles per warp.
        const int dsmem = PB_WARPS * 2 * (PD_STAGE + 32);
        const uint32_t nwt = (n + 31u) / 32u;
        const uint32_t per_cta = PB_WARPS * PD_TILES;
        pb_decode_dense_kernel<false><<<(nwt + per_cta - 1) / per_cta, PB_THREADS, dsmem, ctx->stream>>>(a);
    }
    SW_CUDA(ctx, cudaGetLastError());
    const uint32_t nredo = std::min<uint32_t>(ntiles, (uint32_t)ctx->sm_count * 4u);
    pb_decode_ordered_kernel<<<nredo, PB_THREADS, smem, ctx->stream>>>(a);  // exact redo iff not dense
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
}

// Fused decode + merge launch (swgpu_decode_merge[_dev]); the caller (merge.cu) prepared the table.
int sw_decode_merge_pb_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, const MgTable &t,
                              uint64_t seq_base, const uint64_t *id_hi, const uint64_t *id_lo, uint8_t *status,
                              uint32_t *out_rows_dev) {
    PbArgs a;
    memset(&a, 0, sizeof a);
    a.bytes = bytes;
    a.rec_off = rec_off;
    a.n = n;
    a.status = status;
    a.out_rows = out_rows_dev;
    a.mt = t;
    a.id_hi = id_hi;
    a.id_lo = id_lo;
    a.seq_base = seq_base;
    const int dsmem = PB_WARPS * 2 * (PD_STAGE + 32);
    {
        static std::mutex mu;
        static bool done[64] = {false};
        std::lock_guard<std::mutex> lk(mu);
        const int d = ctx->device & 63;
        if (!done[d]) {
            SW_CUDA(ctx, cudaFuncSetAttribute(pb_decode_dense_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, dsmem));
            done[d] = true;
        }
    }
    const uint32_t nwt = (n + 31u) / 32u;
    const uint32_t per_cta = PB_WARPS * PD_TILES;
    pb_decode_dense_kernel<true><<<(nwt + per_cta - 1) / per_cta, PB_THREADS, dsmem, ctx->stream>>>(a);
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
}

extern "C" int32_t swgpu_decode_pb_dev(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off,
                                       uint32_t n, uint32_t max_rows, const swgpu_decoded_cols *out,
                                       uint8_t *status, uint32_t *first_row, uint32_t *out_rows_dev) {
    return sw_decode_pb_launch(ctx, bytes, rec_off, n, max_rows, out, status, first_row, out_rows_dev, 0, nullptr,
                               nullptr, false);
}

// ---- host-pointer entry point: stage in, decode, stage out ------------------------------------------
int sw_stage_decode(swgpu_ctx *ctx, bool json, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                    uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status, uint32_t *first_row,
                    uint32_t *out_rows) {
    if (!ctx) return SWGPU_E_INVALID;
    if (int rc = check_cols(ctx, out)) return rc;
    if (!rec_off || !first_row || !out_rows || (n && !status) || (rec_off[n] && !bytes)) {
        SW_SET_ERR(ctx, "swgpu_decode: NULL argument");
        return SWGPU_E_INVALID;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    const uint64_t total = rec_off[n];
    const int NCOL = 13;
    if (int rc = sw_ensure(ctx, ctx->d_bytes, total + 64)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_off, (size_t)(n + 1) * 8)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_status, (size_t)n + 16)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_first, (size_t)(n + 2) * 4 + 16)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_cols, (size_t)max_rows * 8 * NCOL + 64)) return rc;
    cudaStream_t st = ctx->stream;
    if (total) SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_bytes.p, bytes, total, cudaMemcpyHostToDevice, st));
    SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_off.p, rec_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, st));
    swgpu_decoded_cols d;
    uint64_t *base = (uint64_t *)ctx->d_cols.p;
    uint64_t **dp = (uint64_t **)&d;
    uint64_t *const *hp = (uint64_t *const *)out;
    for (int c = 0; c < NCOL; c++) dp[c] = (c < 9 || hp[c]) ? base + (size_t)c * max_rows : nullptr;
    uint32_t *d_first = (uint32_t *)ctx->d_first.p;
    uint32_t *d_rows = d_first + (n + 1);
    int rc = json ? swgpu_decode_json_dev(ctx, (const uint8_t *)ctx->d_bytes.p, (const uint64_t *)ctx->d_off.p,
                                          n, max_rows, &d, (uint8_t *)ctx->d_status.p, d_first, d_rows)
                  : swgpu_decode_pb_dev(ctx, (const uint8_t *)ctx->d_bytes.p, (const uint64_t *)ctx->d_off.p, n,
                                        max_rows, &d, (uint8_t *)ctx->d_status.p, d_first, d_rows);
    if (rc) return rc;
    // one D2H for (first_row, out_rows), then the live prefix of each column
    SW_CUDA(ctx, cudaMemcpyAsync(first_row, d_first, (size_t)(n + 1) * 4, cudaMemcpyDeviceToHost, st));
    if (n) SW_CUDA(ctx, cudaMemcpyAsync(status, ctx->d_status.p, n, cudaMemcpyDeviceToHost, st));
    SW_CUDA(ctx, cudaStreamSynchronize(st));
    const uint32_t rows = first_row[n];
    *out_rows = rows;
    const uint32_t live = rows < max_rows ? rows : max_rows;
    if (live) {
        for (int c = 0; c < NCOL; c++)
            if (dp[c]) SW_CUDA(ctx, cudaMemcpyAsync(hp[c], dp[c], (size_t)live * 8, cudaMemcpyDeviceToHost, st));
        SW_CUDA(ctx, cudaStreamSynchronize(st));
    }
    if (rows > max_rows) {
        SW_SET_ERR(ctx, "decode produced %u rows but max_rows is %u", rows, max_rows);
        return SWGPU_E_CAPACITY;
    }
    return SWGPU_OK;
}

extern "C" int32_t swgpu_decode_pb(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                                   uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                                   uint32_t *first_row, uint32_t *out_rows) {
    return sw_stage_decode(ctx, false, bytes, rec_off, n, max_rows, out, status, first_row, out_rows);
}
