// encode.cu — the outbound payload of a decoded request, built on the device
// (reference: EventSourcesManager.handleDecodedEvent, service-event-sources/.../sources/EventSourcesManager.java:152-166,
//  forwards every decoded IDeviceEventCreateRequest as EventModelMarshaler.buildInboundEventPayloadMessage(payload),
//  sitewhere-grpc-event-management/.../EventModelMarshaler.java:93-115 = GInboundEventPayload.writeTo of
//  EventModelConverter.asGrpcInboundEventPayload :1878-1887 -> asGrpcDeviceEventCreateRequest :1621-1653 ->
//  asGrpcDevice{Measurement,Alert,Location}CreateRequest :399-406, :623-632, :770-784 and
//  createGrpcDeviceEventCreateRequest :196-211; schema device-event-model.proto:34-42, 45-49, 66-72, 75-80, 111-120,
//  307-312).
//
// One thread per decoded row, three launches (sizes -> scan of the 256-row tile totals -> bytes), so that the
// records come out back to back in row order with an offset table: the D2H leg of a batch is one contiguous,
// ready-to-produce buffer.  proto3 writeTo of protobuf-java 3.3.0: fields in number order, a scalar equal to its
// default is left out (a double when `value != 0D` is false: -0.0 too), a message field that was set is written
// even when empty.  A row gets a status instead of bytes when the reference would throw (a null where a setter or
// an unboxing needs a value, an event type that does not match the request class), when nothing is sent (null
// request), or when its bytes depend on Java-side state and the host shim has to build it (metadata map: HashMap
// order; strings with JSON escapes, coerced from other tokens, or holding a byte >= 0x80).
#include "swgpu_common.cuh"
#include "tile_scan.cuh"

#define EC_THREADS 256
#define EC_STAGE (32 * 1024)  // a tile whose records fit (128 bytes per row on average) is assembled in shared memory and
                              // copied out coalesced; 32 KiB keeps seven CTAs on an SM
#define EC_MAX_SOURCE 256

struct EcArgs {
    const uint8_t *bytes;
    const uint64_t *rec_off;
    swgpu_decoded_cols cols;
    uint32_t n;
    const uint32_t *n_rows_dev;
    int32_t format;
    int64_t now_ms;
    uint8_t *out;
    uint64_t out_cap;
    uint64_t *out_off;
    uint8_t *enc_status;
    uint64_t *out_bytes;
    uint32_t *lens;
    uint64_t *tile_tot, *tile_base;
    uint32_t source_len;
    uint8_t source[EC_MAX_SOURCE];
};

struct EcRow {
    const uint8_t *tok, *name, *msg, *orig, *alt;
    uint32_t tok_len, name_len, msg_len, orig_len, alt_len;
    uint32_t kind, flags, level, source;
    uint64_t date;
    double v0, v1, v2;
};

__device__ __forceinline__ uint32_t ec_vl(uint64_t v) {  // bytes of a varint
    return v < 0x80 ? 1u : (uint32_t)(70 - __clzll((long long)v)) / 7u;
}
__device__ __forceinline__ uint32_t ec_str(uint32_t len) { return len ? 1u + ec_vl(len) + len : 0u; }
// the aligned 32-bit words that hold s[0..n): word k of the stream, already shifted so that byte 0 is s[4k]
struct EcWords {
    const uint32_t *w;
    uint32_t sh, nw, cur, k;
    __device__ __forceinline__ void open(const uint8_t *s, uint32_t n) {
        const uintptr_t a = (uintptr_t)s;
        w = (const uint32_t *)(a & ~(uintptr_t)3);
        sh = (uint32_t)(a & 3) * 8;
        nw = ((uint32_t)(a & 3) + n + 3) >> 2;  // aligned words with at least one byte of the string
        k = 0;
        cur = n ? __ldg(w) : 0u;
    }
    __device__ __forceinline__ uint32_t next() {  // (a word past the last one that holds a byte of s is never read)
        k++;
        const uint32_t nx = k < nw ? __ldg(w + k) : 0u;
        const uint32_t v = __funnelshift_r(cur, nx, sh);
        cur = nx;
        return v;
    }
};
__device__ __forceinline__ bool ec_ascii(const uint8_t *s, uint32_t n) {
    if (n <= 28) {  // all the words of a short string at once (one round trip); bytes outside the string are masked off
        const uintptr_t a = (uintptr_t)s;
        const uint32_t *w = (const uint32_t *)(a & ~(uintptr_t)3);
        const uint32_t lead = (uint32_t)(a & 3), nw = (lead + n + 3) >> 2;
        uint32_t acc = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if ((uint32_t)k < nw) {
                uint32_t v = __ldg(w + k);
                const uint32_t lo = 4 * k, first = lead > lo ? lead - lo : 0u;              // string bytes of this word:
                const uint32_t last = min(4u, lead + n - lo);                               // [first, last)
                if (first) v &= ~((1u << (8 * first)) - 1u);
                if (last < 4) v &= (1u << (8 * last)) - 1u;
                acc |= v;
            }
        }
        return (acc & 0x80808080u) == 0;
    }
    EcWords ws;
    ws.open(s, n);
    uint32_t acc = 0;
    for (uint32_t i = 0; i < n; i += 4) {
        const uint32_t v = ws.next();
        acc |= n - i >= 4 ? v : v & ((1u << (8 * (n - i))) - 1u);
    }
    return (acc & 0x80808080u) == 0;
}

// loads row i and decides its status (see the header); lengths of the parts it will not write are zeroed.
// CHECK = false: the row is known to be SWGPU_ENC_OK (the emit pass): the strings are not scanned again
template <bool CHECK>
__device__ __forceinline__ uint32_t ec_load(const EcArgs &a, uint32_t i, EcRow &r) {
    const uint64_t m = __ldg(a.cols.meta + i);
    r.kind = SWGPU_META_KIND(m);
    r.flags = SWGPU_META_FLAGS(m);
    r.level = SWGPU_META_LEVEL(m);
    r.source = SWGPU_META_SOURCE(m);
    const uint8_t *rec = a.bytes + __ldg(a.rec_off + SWGPU_META_REC(m));
    auto span = [&](const uint64_t *col, const uint8_t *&p, uint32_t &len) {
        const uint64_t s = col ? __ldg(col + i) : 0ull;
        p = rec + (uint32_t)s;
        len = (uint32_t)(s >> 32);
    };
    span(a.cols.token_span, r.tok, r.tok_len);
    span(a.cols.name_span, r.name, r.name_len);
    span(a.cols.message_span, r.msg, r.msg_len);
    span(a.cols.originator_span, r.orig, r.orig_len);
    span(a.cols.alternate_id_span, r.alt, r.alt_len);
    r.v0 = __ldg(a.cols.v0 + i);
    r.v1 = __ldg(a.cols.v1 + i);
    r.v2 = __ldg(a.cols.v2 + i);
    // asGrpcDate: null -> 0; the protobuf decoder substitutes new Date() for a missing date, the JSON one leaves null
    r.date = (r.flags & SWGPU_F_HAS_DATE) ? (uint64_t)__ldg(a.cols.event_date + i)
                                          : (a.format == SWGPU_FMT_PROTOBUF ? (uint64_t)a.now_ms : 0ull);
    if (r.flags & SWGPU_F_NULL_REQUEST) return SWGPU_ENC_NOTHING;
    if (!SWGPU_META_HAS_ETYPE(m) || SWGPU_META_ETYPE(m) != r.kind || (r.flags & SWGPU_F_NULL_TOKEN)) return SWGPU_ENC_THROWS;
    if (r.kind == SWGPU_KIND_MEASUREMENT && (!(r.flags & SWGPU_F_HAS_NAME) || !(r.flags & SWGPU_F_HAS_V0))) return SWGPU_ENC_THROWS;
    if (r.kind == SWGPU_KIND_ALERT && (!(r.flags & SWGPU_F_HAS_NAME) || !(r.flags & SWGPU_F_HAS_MESSAGE) ||
                                       (r.flags & (SWGPU_F_NULL_LEVEL | SWGPU_F_NULL_SOURCE))))
        return SWGPU_ENC_THROWS;
    if (r.kind == SWGPU_KIND_LOCATION) r.name_len = 0;
    if (r.kind != SWGPU_KIND_ALERT) r.msg_len = 0;
    if (!(r.flags & SWGPU_F_HAS_ORIGINATOR)) r.orig_len = 0;
    if (!(r.flags & SWGPU_F_HAS_ALT_ID)) r.alt_len = 0;
    if (!CHECK) return SWGPU_ENC_OK;
    if ((r.flags & (SWGPU_F_HAS_METADATA | SWGPU_F_STR_ESCAPED | SWGPU_F_NAME_COERCED)) || !ec_ascii(r.tok, r.tok_len) ||
        !ec_ascii(r.name, r.name_len) || !ec_ascii(r.msg, r.msg_len) || !ec_ascii(r.orig, r.orig_len) ||
        !ec_ascii(r.alt, r.alt_len))
        return SWGPU_ENC_HANDBACK;
    return SWGPU_ENC_OK;
}

struct EcLens {
    uint32_t ev, body, any, total;
};
__device__ __forceinline__ EcLens ec_lens(const EcArgs &a, const EcRow &r) {
    EcLens l;
    l.ev = ((r.flags & SWGPU_F_HAS_ALT_ID) ? 1u + ec_vl(ec_str(r.alt_len)) + ec_str(r.alt_len) : 0u) +
           (r.date ? 1u + ec_vl(r.date) : 0u) + ((r.flags & SWGPU_F_UPDATE_STATE) ? 4u : 0u);
    const uint32_t evf = 1u + ec_vl(l.ev) + l.ev;
    if (r.kind == SWGPU_KIND_MEASUREMENT)
        l.body = ec_str(r.name_len) + (r.v0 != 0.0 ? 9u : 0u) + evf;
    else if (r.kind == SWGPU_KIND_ALERT)
        l.body = (r.source ? 2u : 0u) + (r.level ? 2u : 0u) + ec_str(r.name_len) + ec_str(r.msg_len) + evf;
    else
        l.body = ((r.flags & SWGPU_F_HAS_V0) ? 2u + (r.v0 != 0.0 ? 9u : 0u) : 0u) +
                 ((r.flags & SWGPU_F_HAS_V1) ? 2u + (r.v1 != 0.0 ? 9u : 0u) : 0u) +
                 ((r.flags & SWGPU_F_HAS_V2) ? 2u + (r.v2 != 0.0 ? 9u : 0u) : 0u) + evf;
    l.any = 1u + ec_vl(l.body) + l.body;
    l.total = ec_str(a.source_len) + ec_str(r.tok_len) +
              ((r.flags & SWGPU_F_HAS_ORIGINATOR) ? 1u + ec_vl(ec_str(r.orig_len)) + ec_str(r.orig_len) : 0u) + 1u +
              ec_vl(l.any) + l.any;
    return l;
}

struct EcWr {  // byte writer (generic address: shared-memory stage or global memory)
    uint8_t *p;
    __device__ __forceinline__ void b(uint32_t v) { *p++ = (uint8_t)v; }
    __device__ __forceinline__ void varint(uint64_t v) {
        while (v >= 0x80) {
            *p++ = (uint8_t)(v | 0x80);
            v >>= 7;
        }
        *p++ = (uint8_t)v;
    }
    __device__ __forceinline__ void f64(double v) {
        const uint64_t u = (uint64_t)__double_as_longlong(v);
#pragma unroll
        for (int k = 0; k < 8; k++) *p++ = (uint8_t)(u >> (8 * k));
    }
    __device__ __forceinline__ void str(uint32_t tag, const uint8_t *s, uint32_t len) {  // global source, 4 bytes per load
        if (!len) return;
        b(tag);
        varint(len);
        if (len <= 28) {
            // the usual string (a 21-byte token, a short name): all of its aligned words are asked for at once, so that
            // the row waits for one round trip, not for one per word
            const uintptr_t a = (uintptr_t)s;
            const uint32_t *w = (const uint32_t *)(a & ~(uintptr_t)3);
            const uint32_t sh = (uint32_t)(a & 3) * 8, nw = ((uint32_t)(a & 3) + len + 3) >> 2;
            uint32_t wb[9];
#pragma unroll
            for (int k = 0; k < 8; k++) wb[k] = (uint32_t)k < nw ? __ldg(w + k) : 0u;
            wb[8] = 0;
#pragma unroll
            for (int k = 0; k < 7; k++) {
                if ((uint32_t)(4 * k) < len) {
                    const uint32_t v = __funnelshift_r(wb[k], wb[k + 1], sh);
                    const uint32_t m = len - 4 * k;  // bytes of this word that belong to the string
                    p[4 * k] = (uint8_t)v;
                    if (m > 1) p[4 * k + 1] = (uint8_t)(v >> 8);
                    if (m > 2) p[4 * k + 2] = (uint8_t)(v >> 16);
                    if (m > 3) p[4 * k + 3] = (uint8_t)(v >> 24);
                }
            }
            p += len;
            return;
        }
        EcWords ws;
        ws.open(s, len);
        uint32_t i = 0;
        for (; i + 4 <= len; i += 4) {
            const uint32_t v = ws.next();
            p[0] = (uint8_t)v;
            p[1] = (uint8_t)(v >> 8);
            p[2] = (uint8_t)(v >> 16);
            p[3] = (uint8_t)(v >> 24);
            p += 4;
        }
        if (i < len) {
            uint32_t v = ws.next();
            for (; i < len; i++, v >>= 8) *p++ = (uint8_t)v;
        }
    }
    __device__ __forceinline__ void str_param(uint32_t tag, const uint8_t *s, uint32_t len) {  // kernel-parameter source
        if (!len) return;
        b(tag);
        varint(len);
        for (uint32_t i = 0; i < len; i++) *p++ = s[i];
    }
    __device__ __forceinline__ void opt_str(uint32_t tag, const uint8_t *s, uint32_t len) {
        b(tag);
        varint(ec_str(len));
        str(0x0A, s, len);
    }
    __device__ __forceinline__ void opt_f64(uint32_t tag, double v) {
        b(tag);
        if (v != 0.0) {
            b(9);
            b(0x09);
            f64(v);
        } else
            b(0);
    }
};

__device__ __forceinline__ void ec_write(const EcArgs &a, const EcRow &r, const EcLens &l, uint8_t *dst) {
    EcWr w{dst};
    w.str_param(0x0A, a.source, a.source_len);
    w.str(0x12, r.tok, r.tok_len);
    if (r.flags & SWGPU_F_HAS_ORIGINATOR) w.opt_str(0x1A, r.orig, r.orig_len);
    w.b(0x22);
    w.varint(l.any);
    w.b(r.kind == SWGPU_KIND_MEASUREMENT ? 0x0A : r.kind == SWGPU_KIND_ALERT ? 0x12 : 0x1A);
    w.varint(l.body);
    uint32_t ev_tag;
    if (r.kind == SWGPU_KIND_MEASUREMENT) {
        w.str(0x0A, r.name, r.name_len);
        if (r.v0 != 0.0) {
            w.b(0x11);
            w.f64(r.v0);
        }
        ev_tag = 0x1A;
    } else if (r.kind == SWGPU_KIND_ALERT) {
        if (r.source) {
            w.b(0x08);
            w.b(r.source);
        }
        if (r.level) {
            w.b(0x10);
            w.b(r.level);
        }
        w.str(0x1A, r.name, r.name_len);
        w.str(0x22, r.msg, r.msg_len);
        ev_tag = 0x2A;
    } else {
        if (r.flags & SWGPU_F_HAS_V0) w.opt_f64(0x0A, r.v0);
        if (r.flags & SWGPU_F_HAS_V1) w.opt_f64(0x12, r.v1);
        if (r.flags & SWGPU_F_HAS_V2) w.opt_f64(0x1A, r.v2);
        ev_tag = 0x22;
    }
    w.b(ev_tag);
    w.varint(l.ev);
    if (r.flags & SWGPU_F_HAS_ALT_ID) w.opt_str(0x0A, r.alt, r.alt_len);
    if (r.date) {
        w.b(0x28);
        w.varint(r.date);
    }
    if (r.flags & SWGPU_F_UPDATE_STATE) {
        w.b(0x32);
        w.b(2);
        w.b(0x08);
        w.b(1);
    }
}

__device__ __forceinline__ uint32_t ec_block_sum(uint32_t v, uint32_t *s_part) {
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = __reduce_add_sync(SW_FULL, v);
    if (lane == 0) s_part[warp] = v;
    __syncthreads();
    uint32_t t = 0;
#pragma unroll
    for (int k = 0; k < EC_THREADS / 32; k++) t += s_part[k];
    return t;
}

__global__ void __launch_bounds__(EC_THREADS) encode_size_kernel(EcArgs a) {
    __shared__ uint32_t s_part[EC_THREADS / 32];
    const uint32_t n_live = a.n_rows_dev ? min(__ldg(a.n_rows_dev), a.n) : a.n;
    const uint32_t i = blockIdx.x * EC_THREADS + threadIdx.x;
    uint32_t len = 0;
    if (i < n_live) {
        EcRow r;
        const uint32_t st = ec_load<true>(a, i, r);
        a.enc_status[i] = (uint8_t)st;
        if (st == SWGPU_ENC_OK) len = ec_lens(a, r).total;
    }
    if (i < a.n) a.lens[i] = len;
    const uint32_t tot = ec_block_sum(len, s_part);
    if (threadIdx.x == 0) a.tile_tot[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(EC_THREADS) encode_emit_kernel(EcArgs a) {
    extern __shared__ __align__(16) uint8_t stage[];  // EC_STAGE + 32
    __shared__ uint32_t s_w[EC_THREADS / 32];
    const uint32_t n_live = a.n_rows_dev ? min(__ldg(a.n_rows_dev), a.n) : a.n;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t i = blockIdx.x * EC_THREADS + tid;
    const uint32_t len = i < a.n ? a.lens[i] : 0u;
    // exclusive offset of this row inside the tile
    uint32_t inc = len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(SW_FULL, inc, o);
        if (lane >= (uint32_t)o) inc += u;
    }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    uint32_t wbase = 0, tile_len = 0;
#pragma unroll
    for (int k = 0; k < EC_THREADS / 32; k++) {
        if ((uint32_t)k < warp) wbase += s_w[k];
        tile_len += s_w[k];
    }
    const uint32_t excl = wbase + inc - len;
    const uint64_t tbase = a.tile_base[blockIdx.x];
    const uint64_t pos = tbase + excl;
    if (i < a.n) a.out_off[i] = pos;
    if (i == a.n - 1) a.out_off[a.n] = pos + len;
    if (i == 0) *a.out_bytes = a.tile_base[gridDim.x - 1] + a.tile_tot[gridDim.x - 1];
    // the whole tile goes out or nothing of it does (the caller sees the total and retries with room)
    const bool fits = tbase + tile_len <= a.out_cap;
    if (!fits || tile_len == 0) return;
    const uint32_t head = (uint32_t)((uintptr_t)(a.out + tbase) & 15);  // so that the copy-out is 16-byte aligned
    const bool staged = head + tile_len <= EC_STAGE;
    if (len && i < n_live) {
        EcRow r;
        ec_load<false>(a, i, r);
        const EcLens l = ec_lens(a, r);
        ec_write(a, r, l, staged ? stage + head + excl : a.out + pos);
    }
    if (!staged) return;
    __syncthreads();
    uint8_t *dst = a.out + tbase - head;  // 16-byte aligned
    const uint32_t end = head + tile_len;
    // first and last partial vectors byte by byte, the rest as 16-byte stores
    const uint32_t v0 = head ? 16u : 0u, v1 = end & ~15u;
    if (v1 > v0)
        for (uint32_t o = v0 + tid * 16; o < v1; o += EC_THREADS * 16) *(uint4 *)(dst + o) = *(const uint4 *)(stage + o);
    if (tid < 16) {
        const uint32_t o = tid;  // head vector
        if (head && o >= head && o < min(end, 16u)) dst[o] = stage[o];
        const uint32_t t = max(v1, v0) + tid;  // tail vector (or everything when the tile ends inside the head vector)
        if (t < end && t >= (head ? 16u : 0u)) dst[t] = stage[t];
    }
}

extern "C" int32_t swgpu_encode_inbound_dev(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off,
                                            const swgpu_decoded_cols *rows, uint32_t n_rows, const uint32_t *n_rows_dev,
                                            const uint8_t *source_id, uint32_t source_id_len, int64_t now_ms, uint8_t *out,
                                            uint64_t out_cap, uint64_t *out_off, uint8_t *enc_status, uint64_t *out_bytes_dev) {
    if (!ctx) return SWGPU_E_INVALID;
    if (!rec_off || !rows || !rows->meta || !rows->event_date || !rows->v0 || !rows->v1 || !rows->v2 || !rows->token_span ||
        !rows->name_span || !out_off || !out_bytes_dev || (n_rows && !enc_status) || (out_cap && !out) ||
        (source_id_len && !source_id) || (format != SWGPU_FMT_PROTOBUF && format != SWGPU_FMT_JSON)) {
        SW_SET_ERR(ctx, "swgpu_encode_inbound_dev: bad argument");
        return SWGPU_E_INVALID;
    }
    if (source_id_len > EC_MAX_SOURCE) {
        SW_SET_ERR(ctx, "swgpu_encode_inbound_dev: source id longer than %d bytes", EC_MAX_SOURCE);
        return SWGPU_E_INVALID;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    if (n_rows == 0) {
        SW_CUDA(ctx, cudaMemsetAsync(out_off, 0, sizeof(uint64_t), ctx->stream));
        SW_CUDA(ctx, cudaMemsetAsync(out_bytes_dev, 0, sizeof(uint64_t), ctx->stream));
        return SWGPU_OK;
    }
    const uint32_t ntiles = (n_rows + EC_THREADS - 1) / EC_THREADS;
    const size_t o_tot = SW_SCRATCH_HDR, o_base = o_tot + (size_t)ntiles * 8, o_lens = o_base + (size_t)ntiles * 8;
    if (int rc = sw_ensure_scratch(ctx, o_lens + (size_t)n_rows * 4)) return rc;
    EcArgs a;
    a.bytes = bytes;
    a.rec_off = rec_off;
    a.cols = *rows;
    a.n = n_rows;
    a.n_rows_dev = n_rows_dev;
    a.format = format;
    a.now_ms = now_ms;
    a.out = out;
    a.out_cap = out_cap;
    a.out_off = out_off;
    a.enc_status = enc_status;
    a.out_bytes = out_bytes_dev;
    a.tile_tot = (uint64_t *)((uint8_t *)ctx->scratch + o_tot);
    a.tile_base = (uint64_t *)((uint8_t *)ctx->scratch + o_base);
    a.lens = (uint32_t *)((uint8_t *)ctx->scratch + o_lens);
    a.source_len = source_id_len;
    memset(a.source, 0, sizeof a.source);
    if (source_id_len) memcpy(a.source, source_id, source_id_len);
    {
        static std::mutex mu;
        static bool done[64] = {false};
        std::lock_guard<std::mutex> lk(mu);
        const int d = ctx->device & 63;
        if (!done[d]) {
            SW_CUDA(ctx, cudaFuncSetAttribute(encode_emit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, EC_STAGE + 32));
            done[d] = true;
        }
    }
    encode_size_kernel<<<ntiles, EC_THREADS, 0, ctx->stream>>>(a);
    SW_CUDA(ctx, cudaGetLastError());
    if (ntiles <= ZS_MC_MAX_TILES)
        zone_scan_mc_kernel<<<(ntiles + ZS_THREADS - 1) / ZS_THREADS, ZS_THREADS, 0, ctx->stream>>>(a.tile_tot, a.tile_base, ntiles,
                                                                                                  out_bytes_dev);
    else
        zone_scan_kernel<<<1, ZS_THREADS, 0, ctx->stream>>>(a.tile_tot, a.tile_base, ntiles, out_bytes_dev);
    SW_CUDA(ctx, cudaGetLastError());
    encode_emit_kernel<<<ntiles, EC_THREADS, EC_STAGE + 32, ctx->stream>>>(a);
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
}
