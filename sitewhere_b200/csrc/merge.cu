// merge.cu — batched device-state last-writer merge
// (reference: service-device-state/src/main/java/com/sitewhere/devicestate/processing/
//  DeviceStateProcessingLogic.java:116-201 + persistence/DeviceStatePersistence.java:60-85: a sequential
//  fold in arrival order; lastLocationEventId := id, lastMeasurementEventIds[name] := id,
//  lastAlertEventIds[type] := id).
//
// Sort-free: the sequential "later event overwrites" is, per key, an argmax over the arrival
// sequence number.  The table and the per-row upsert are in merge_table.cuh; this file holds the
// kernel that folds rows read back from decoded columns (swgpu_merge[_dev]), the fix-up pass, table growth,
// snapshot / restore, and the host entry points.  The fused decode+merge kernels (swgpu_decode_merge[_dev]),
// which take the row's key bytes and payload straight from the decoder's registers and shared memory, are
// in decode_pb.cu / decode_json.cu.  HBM-bound, random access: one 64-byte read and 64 bytes of dirty sectors
// per event in the table.
#include <algorithm>

#include "merge_table.cuh"

#define MG_THREADS 256

struct MgState {
    MgSlot *slots = nullptr;
    uint64_t cap = 0;  // power of two
    uint64_t *arena = nullptr;
    uint64_t arena_cap = 0;  // words
    MgCounters *ctr = nullptr;
    MgCounters *h_ctr = nullptr;  // pinned read-back
    uint64_t keys_upper = 0;  // host upper bound on live keys
    swgpu_ctx::Buf fix;       // fix-up list of the current launch
};

struct MgArgs {
    MgTable t;
    const uint8_t *bytes;
    const uint64_t *rec_off;
    const uint64_t *meta, *token_hash, *name_hash, *token_span, *name_span;
    const int64_t *event_date;
    const double *v0;
    uint32_t n_rows;
    const uint32_t *n_rows_dev;
    uint64_t seq_base;
    const uint64_t *id_hi, *id_lo;
};

// 8 payload bytes starting at p (any alignment) as a little-endian word; bytes past `valid` are zero.
// Reads only the aligned words that hold valid bytes (the payload buffer is 8-byte aligned).
__device__ __forceinline__ uint64_t mg_payload_word(const uint8_t *p, uint32_t valid) {
    const uintptr_t a = (uintptr_t)p;
    const uint64_t *w = (const uint64_t *)(a & ~(uintptr_t)7);
    const uint32_t mis = (uint32_t)(a & 7);
    uint64_t v = __ldg(w) >> (8 * mis);
    if (mis && (8 - mis) < valid) v |= __ldg(w + 1) << (64 - 8 * mis);
    if (valid < 8) v &= (1ull << (8 * valid)) - 1;
    return v;
}
struct MgKeyGmem {  // key words from the payload in global memory
    const uint8_t *tok, *name;
    uint32_t tl, nl;
    __device__ __forceinline__ uint64_t word(uint32_t i) const {
        const uint32_t tw = (tl + 7u) >> 3;
        if (i <= tw) return mg_payload_word(tok + 8 * (i - 1), min(8u, tl - 8 * (i - 1)));
        const uint32_t j = i - 1 - tw;
        return mg_payload_word(name + 8 * j, min(8u, nl - 8 * j));
    }
};

__global__ void __launch_bounds__(MG_THREADS) merge_upsert_kernel(MgArgs a) {
    __shared__ MgLocalCnt s_cnt[MG_THREADS / 32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) s_cnt[warp].keys = s_cnt[warp].bytes = 0;
    __syncwarp();
    const uint32_t n_live = a.n_rows_dev ? min(__ldg(a.n_rows_dev), a.n_rows) : a.n_rows;
    const uint32_t ridx = blockIdx.x * MG_THREADS + threadIdx.x;
    if (ridx < n_live) {
        // newest rows first: CTAs start roughly in index order, so for a key with several rows in the batch the
        // winner tends to arrive first and the older rows lose their atomicMax at once
        const uint32_t r = n_live - 1 - ridx;
        const uint64_t m = __ldg(a.meta + r);
        const uint32_t kind = SWGPU_META_KIND(m), fl = SWGPU_META_FLAGS(m);
        // only Measurement / Location / Alert reach device state (DeviceStateProcessingLogic.java:122-127); a row
        // without a device token or with a null request never gets there (ES/EventSourcesManager.java:155)
        if (kind <= SWGPU_KIND_ALERT && !(fl & (SWGPU_F_NULL_TOKEN | SWGPU_F_NULL_REQUEST))) {
            const uint8_t *rec = a.bytes + __ldg(a.rec_off + SWGPU_META_REC(m));
            const uint64_t ts = __ldg(a.token_span + r), ns = __ldg(a.name_span + r);
            const uint32_t tl = SWGPU_SPAN_LEN(ts);
            const uint32_t nl = kind == SWGPU_KIND_LOCATION ? 0u : SWGPU_SPAN_LEN(ns);
            if (tl > 0xffffu || nl > 0xffffu) {
                atomicAdd(&a.t.ctr->key_too_long, 1ull);
            } else {
                const uint32_t kindbyte = mg_kindbyte(kind, fl);
                const uint64_t h = mg_mix(__ldg(a.token_hash + r),
                                          kind == SWGPU_KIND_LOCATION ? 0ull : __ldg(a.name_hash + r), kindbyte);
                MgKeyGmem key{rec + SWGPU_SPAN_OFF(ts), rec + SWGPU_SPAN_OFF(ns), tl, nl};
                mg_upsert(a.t, &s_cnt[warp], h, __ldg(a.token_hash + r), tl, nl, kindbyte, key, a.seq_base + r + 1, __ldg(a.id_hi + r),
                          __ldg(a.id_lo + r), (uint64_t)__ldg(a.event_date + r),
                          (uint64_t)__double_as_longlong(__ldg(a.v0 + r)));
            }
        }
    }
    __syncwarp();
    if (lane == 0) mg_flush_counts(a.t, &s_cnt[warp]);
}

// second pass, over the fix-up list only
__global__ void __launch_bounds__(MG_THREADS) merge_fixup_kernel(MgTable t) {
    unsigned long long n = t.ctr->n_fix;
    if (n > t.fix_cap) n = t.fix_cap;
    for (unsigned long long k = (unsigned long long)blockIdx.x * MG_THREADS + threadIdx.x; k < n;
         k += (unsigned long long)gridDim.x * MG_THREADS) {
        const MgFix f = t.fix[k];
        uint64_t *w = t.slots[f.slot].w;
        if (w[MG_W_BEST] != f.seq1) continue;  // a later event of this key won
        mg_store_payload(w, f.id_hi, f.id_lo, f.date, f.value);
    }
}

// re-insert live slots into a larger table (keys are distinct: no compare needed; tombstones are dropped)
__global__ void merge_rehash_kernel(const MgSlot *old_slots, uint64_t old_cap, MgSlot *slots, uint64_t mask) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= old_cap) return;
    const uint64_t tag = old_slots[i].w[MG_W_TAG];
    if (tag == 0 || tag == MG_TOMB) return;
    uint64_t s = (tag >> 3) & mask;
    for (;;) {
        const unsigned long long old = atomicCAS((unsigned long long *)slots[s].w, 0ull, (unsigned long long)tag);
        if (old == 0) break;
        s = (s + 1) & mask;
    }
    const uint64_t *src = old_slots[i].w;
    uint64_t *dst = slots[s].w;
#pragma unroll
    for (int k = 1; k < 16; k++) dst[k] = src[k];  // (the tag is already there)
}

void sw_state_free(swgpu_ctx *ctx) {
    MgState *st = (MgState *)ctx->state;
    if (!st) return;
    if (st->slots) cudaFree(st->slots);
    if (st->arena) cudaFree(st->arena);
    if (st->ctr) cudaFree(st->ctr);
    if (st->h_ctr) cudaFreeHost(st->h_ctr);
    if (st->fix.p) cudaFree(st->fix.p);
    delete st;
    ctx->state = nullptr;
}

static uint64_t pow2_at_least(uint64_t v) {
    uint64_t p = 1024;
    while (p < v) p <<= 1;
    return p;
}

static int mg_read_counters(swgpu_ctx *ctx, MgState *st, MgCounters *out) {
    SW_CUDA(ctx, cudaMemcpyAsync(st->h_ctr, st->ctr, sizeof(MgCounters), cudaMemcpyDeviceToHost, ctx->stream));
    SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *out = *st->h_ctr;
    return SWGPU_OK;
}

// make room for `max_keys` keys and `max_key_bytes` bytes of LONG keys (grows, never shrinks)
static int mg_reserve(swgpu_ctx *ctx, uint64_t max_keys, uint64_t max_key_bytes) {
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    MgState *st = (MgState *)ctx->state;
    if (!st) {
        st = new MgState();
        ctx->state = st;
        cudaError_t e = cudaMalloc(&st->ctr, sizeof(MgCounters));
        if (e == cudaSuccess) e = cudaMallocHost(&st->h_ctr, sizeof(MgCounters));
        if (e != cudaSuccess) {
            SW_SET_ERR(ctx, "cudaMalloc(state counters) failed: %s", cudaGetErrorString(e));
            return SWGPU_E_NOMEM;
        }
        SW_CUDA(ctx, cudaMemsetAsync(st->ctr, 0, sizeof(MgCounters), ctx->stream));
    }
    const uint64_t want_cap = pow2_at_least(max_keys * 2 + 16);
    if (want_cap > st->cap) {
        MgSlot *ns = nullptr;
        cudaError_t e = cudaMalloc(&ns, want_cap * sizeof(MgSlot));
        if (e != cudaSuccess) {
            SW_SET_ERR(ctx, "cudaMalloc(state table %llu slots) failed: %s", (unsigned long long)want_cap,
                       cudaGetErrorString(e));
            return SWGPU_E_NOMEM;
        }
        SW_CUDA(ctx, cudaMemsetAsync(ns, 0, want_cap * sizeof(MgSlot), ctx->stream));
        if (st->slots) {
            const uint64_t nb = (st->cap + 255) / 256;
            merge_rehash_kernel<<<(unsigned)nb, 256, 0, ctx->stream>>>(st->slots, st->cap, ns, want_cap - 1);
            SW_CUDA(ctx, cudaGetLastError());
            SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            SW_CUDA(ctx, cudaFree(st->slots));
        }
        st->slots = ns;
        st->cap = want_cap;
    }
    const uint64_t want_words = (max_key_bytes + 7) / 8;
    if (want_words > st->arena_cap || !st->arena) {
        const uint64_t want = want_words + want_words / 4 + 512;
        uint64_t *na = nullptr;
        cudaError_t e = cudaMalloc(&na, want * 8);
        if (e != cudaSuccess) {
            SW_SET_ERR(ctx, "cudaMalloc(key arena %llu) failed: %s", (unsigned long long)want * 8, cudaGetErrorString(e));
            return SWGPU_E_NOMEM;
        }
        if (st->arena) {
            MgCounters c;
            if (int rc = mg_read_counters(ctx, st, &c)) return rc;
            const uint64_t used = c.arena_top < st->arena_cap ? c.arena_top : st->arena_cap;
            if (used) SW_CUDA(ctx, cudaMemcpyAsync(na, st->arena, used * 8, cudaMemcpyDeviceToDevice, ctx->stream));
            SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            SW_CUDA(ctx, cudaFree(st->arena));
        }
        st->arena = na;
        st->arena_cap = want;
    }
    return SWGPU_OK;
}

extern "C" int32_t swgpu_state_reserve(swgpu_ctx *ctx, uint64_t max_keys, uint64_t max_key_bytes) {
    if (!ctx) return SWGPU_E_INVALID;
    return mg_reserve(ctx, max_keys, max_key_bytes);
}

// Before a launch that may add up to `new_keys` keys: grow the table if the host's upper bound says it could
// pass half full (the bound is tightened from the device counters first), size the fix-up list, fill `t`.
int sw_state_prepare(swgpu_ctx *ctx, uint64_t new_keys, uint64_t rows, uint64_t seq_floor, MgTable *t) {
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    MgState *st = (MgState *)ctx->state;
    if (ctx->opt_presence_now1) new_keys *= 2;  // every row may also add its device's presence row
    uint64_t ku = (st ? st->keys_upper : 0) + new_keys;
    if (!st || !st->arena || ku * 2 + 16 > st->cap) {
        if (st && st->ctr && st->slots) {
            MgCounters c;
            if (int rc = mg_read_counters(ctx, st, &c)) return rc;
            st->keys_upper = c.n_keys;
            ku = st->keys_upper + new_keys;
        }
        if (int rc = mg_reserve(ctx, ku, 0)) return rc;
        st = (MgState *)ctx->state;
    }
    st->keys_upper = ku;
    const uint64_t fix_cap = rows / 4 + 4096;
    if (int rc = sw_ensure(ctx, st->fix, (size_t)fix_cap * sizeof(MgFix))) return rc;
    t->slots = st->slots;
    t->mask = st->cap - 1;
    t->arena = st->arena;
    t->arena_cap = st->arena_cap;
    t->ctr = st->ctr;
    t->fix = (MgFix *)st->fix.p;
    t->fix_cap = fix_cap;
    t->seq_floor = seq_floor;
    t->presence_now1 = ctx->opt_presence_now1;
    SW_CUDA(ctx, cudaMemsetAsync(&st->ctr->n_fix, 0, sizeof(unsigned long long), ctx->stream));
    return SWGPU_OK;
}
int sw_state_fixup_launch(swgpu_ctx *ctx, const MgTable &t, uint64_t rows) {
    const uint64_t nb = (rows + MG_THREADS - 1) / MG_THREADS;
    merge_fixup_kernel<<<(unsigned)std::min<uint64_t>(std::max<uint64_t>(nb, 1), (uint64_t)ctx->sm_count * 4u), MG_THREADS, 0,
                         ctx->stream>>>(t);
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
}

// After a launch, on the host-synchronising entry points: read the counters, tighten the bounds, and say what to
// do.  Returns 0 = done, 1 = the caller should replay the launch (room was made: more arena, a larger table, or
// the fix-up list overflowed — a replay is idempotent and settles every payload), <0 = error.
int sw_state_after(swgpu_ctx *ctx, uint64_t rows, const char *who) {
    MgState *st = (MgState *)ctx->state;
    if (!st) return SWGPU_OK;
    MgCounters c;
    if (int rc = mg_read_counters(ctx, st, &c)) return rc;
    st->keys_upper = c.n_keys;
    if (c.key_too_long) {
        SW_CUDA(ctx, cudaMemsetAsync(&st->ctr->key_too_long, 0, sizeof(unsigned long long), ctx->stream));
        SW_SET_ERR(ctx, "%s: %llu rows have a device token or name longer than 65535 bytes and were not merged", who,
                   c.key_too_long);
        return SWGPU_E_CAPACITY;
    }
    if (!c.table_full && !c.arena_full && !c.fix_overflow) return SWGPU_OK;
    SW_CUDA(ctx, cudaMemsetAsync(&st->ctr->table_full, 0, 2 * sizeof(unsigned long long), ctx->stream));  // + arena_full
    SW_CUDA(ctx, cudaMemsetAsync(&st->ctr->fix_overflow, 0, sizeof(unsigned long long), ctx->stream));
    if (c.table_full) {
        if (int rc = mg_reserve(ctx, c.n_keys + rows, 0)) return rc;
    }
    if (c.arena_full) {
        if (int rc = mg_reserve(ctx, 0, (c.arena_top + c.arena_full + 4096) * 8)) return rc;
    }
    return 1;
}

int sw_merge_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n_records,
                    const swgpu_decoded_cols *rows, uint32_t n_rows, const uint32_t *n_rows_dev, uint64_t seq_base,
                    const uint64_t *id_hi, const uint64_t *id_lo) {
    (void)n_records;
    if (!ctx) return SWGPU_E_INVALID;
    if (n_rows == 0) return SWGPU_OK;
    if (!bytes || !rec_off || !rows || !rows->meta || !rows->token_hash || !rows->name_hash || !rows->token_span ||
        !rows->name_span || !rows->event_date || !rows->v0 || !id_hi || !id_lo) {
        SW_SET_ERR(ctx, "swgpu_merge_dev: NULL argument");
        return SWGPU_E_INVALID;
    }
    MgArgs a;
    if (int rc = sw_state_prepare(ctx, n_rows, n_rows, seq_base, &a.t)) return rc;
    a.bytes = bytes;
    a.rec_off = rec_off;
    a.meta = rows->meta;
    a.token_hash = rows->token_hash;
    a.name_hash = rows->name_hash;
    a.token_span = rows->token_span;
    a.name_span = rows->name_span;
    a.event_date = rows->event_date;
    a.v0 = rows->v0;
    a.n_rows = n_rows;
    a.n_rows_dev = n_rows_dev;
    a.seq_base = seq_base;
    a.id_hi = id_hi;
    a.id_lo = id_lo;
    const unsigned nb = (n_rows + MG_THREADS - 1) / MG_THREADS;
    merge_upsert_kernel<<<nb, MG_THREADS, 0, ctx->stream>>>(a);
    SW_CUDA(ctx, cudaGetLastError());
    return sw_state_fixup_launch(ctx, a.t, n_rows);
}

extern "C" int32_t swgpu_merge_dev(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off,
                                   uint32_t n_records, const swgpu_decoded_cols *rows, uint32_t n_rows,
                                   const uint32_t *n_rows_dev, uint64_t seq_base, const uint64_t *id_hi,
                                   const uint64_t *id_lo) {
    return sw_merge_launch(ctx, bytes, rec_off, n_records, rows, n_rows, n_rows_dev, seq_base, id_hi, id_lo);
}

extern "C" int32_t swgpu_merge(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n_records,
                               const swgpu_decoded_cols *rows, uint32_t n_rows, uint64_t seq_base,
                               const uint64_t *id_hi, const uint64_t *id_lo) {
    if (!ctx) return SWGPU_E_INVALID;
    if (n_rows == 0) return SWGPU_OK;
    if (!bytes || !rec_off || !rows || !id_hi || !id_lo) {
        SW_SET_ERR(ctx, "swgpu_merge: NULL argument");
        return SWGPU_E_INVALID;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    const uint64_t total = rec_off[n_records];
    if (int rc = sw_ensure(ctx, ctx->d_bytes, total + 64)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_off, (size_t)(n_records + 1) * 8)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_cols, (size_t)n_rows * 8 * 7 + 64)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_ids, (size_t)n_rows * 16 + 64)) return rc;
    SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_bytes.p, bytes, total, cudaMemcpyHostToDevice, s));
    SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_off.p, rec_off, (size_t)(n_records + 1) * 8, cudaMemcpyHostToDevice, s));
    uint64_t *base = (uint64_t *)ctx->d_cols.p;
    swgpu_decoded_cols d;
    memset(&d, 0, sizeof d);
    const void *src[7] = {rows->meta, rows->token_hash, rows->name_hash, rows->event_date,
                          rows->v0,   rows->token_span, rows->name_span};
    for (int c = 0; c < 7; c++) {
        if (!src[c]) {
            SW_SET_ERR(ctx, "swgpu_merge: missing column");
            return SWGPU_E_INVALID;
        }
        SW_CUDA(ctx, cudaMemcpyAsync(base + (size_t)c * n_rows, src[c], (size_t)n_rows * 8, cudaMemcpyHostToDevice, s));
    }
    d.meta = base;
    d.token_hash = base + (size_t)n_rows;
    d.name_hash = base + (size_t)2 * n_rows;
    d.event_date = (int64_t *)(base + (size_t)3 * n_rows);
    d.v0 = (double *)(base + (size_t)4 * n_rows);
    d.token_span = base + (size_t)5 * n_rows;
    d.name_span = base + (size_t)6 * n_rows;
    uint64_t *ids = (uint64_t *)ctx->d_ids.p;
    SW_CUDA(ctx, cudaMemcpyAsync(ids, id_hi, (size_t)n_rows * 8, cudaMemcpyHostToDevice, s));
    SW_CUDA(ctx, cudaMemcpyAsync(ids + n_rows, id_lo, (size_t)n_rows * 8, cudaMemcpyHostToDevice, s));
    for (int attempt = 0; attempt < 4; attempt++) {
        if (int rc = swgpu_merge_dev(ctx, (const uint8_t *)ctx->d_bytes.p, (const uint64_t *)ctx->d_off.p, n_records, &d, n_rows,
                                     nullptr, seq_base, ids, ids + n_rows))
            return rc;
        const int rc = sw_state_after(ctx, n_rows, "swgpu_merge");
        if (rc <= 0) return rc;
    }
    SW_SET_ERR(ctx, "swgpu_merge: the state table could not be grown to hold the batch");
    return SWGPU_E_CAPACITY;
}

// ---- fused decode + merge (kernels in decode_pb.cu / decode_json.cu) --------------------------------------------------
int sw_decode_merge_pb_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, const MgTable &t,
                              uint64_t seq_base, const uint64_t *id_hi, const uint64_t *id_lo, uint8_t *status,
                              uint32_t *out_rows_dev);
int sw_decode_merge_json_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, const MgTable &t,
                                uint64_t seq_base, const uint64_t *id_hi, const uint64_t *id_lo, uint8_t *status,
                                uint32_t *out_rows_dev);

int sw_decode_merge_enriched_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, const MgTable &t,
                                    uint64_t seq_base, uint8_t *status, uint32_t *out_rows_dev);

extern "C" int32_t swgpu_decode_merge_dev(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off,
                                          uint32_t n, uint64_t seq_base, const uint64_t *id_hi, const uint64_t *id_lo,
                                          uint8_t *status, uint32_t *out_rows_dev) {
    if (!ctx) return SWGPU_E_INVALID;
    const bool enriched = format == SWGPU_FMT_ENRICHED;  // (event ids come from the payload: id_hi / id_lo unused)
    if (!out_rows_dev || (format != SWGPU_FMT_PROTOBUF && format != SWGPU_FMT_JSON && !enriched) ||
        (n && (!rec_off || !bytes || !status || (!enriched && (!id_hi || !id_lo))))) {
        SW_SET_ERR(ctx, "swgpu_decode_merge_dev: bad argument");
        return SWGPU_E_INVALID;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    SW_CUDA(ctx, cudaMemsetAsync(out_rows_dev, 0, sizeof(uint32_t), ctx->stream));
    if (n == 0) return SWGPU_OK;
    MgTable t;
    if (int rc = sw_state_prepare(ctx, n, n, seq_base, &t)) return rc;
    const int rc = enriched ? sw_decode_merge_enriched_launch(ctx, bytes, rec_off, n, t, seq_base, status, out_rows_dev)
                   : format == SWGPU_FMT_JSON
                       ? sw_decode_merge_json_launch(ctx, bytes, rec_off, n, t, seq_base, id_hi, id_lo, status, out_rows_dev)
                       : sw_decode_merge_pb_launch(ctx, bytes, rec_off, n, t, seq_base, id_hi, id_lo, status, out_rows_dev);
    if (rc) return rc;
    return sw_state_fixup_launch(ctx, t, n);
}

extern "C" int32_t swgpu_decode_merge(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off,
                                      uint32_t n, uint64_t seq_base, const uint64_t *id_hi, const uint64_t *id_lo,
                                      uint8_t *status, uint32_t *out_rows) {
    if (!ctx) return SWGPU_E_INVALID;
    const bool enriched = format == SWGPU_FMT_ENRICHED;
    if (!rec_off || !out_rows || (format != SWGPU_FMT_PROTOBUF && format != SWGPU_FMT_JSON && !enriched) ||
        (n && ((!enriched && (!id_hi || !id_lo)) || (rec_off[n] && !bytes)))) {
        SW_SET_ERR(ctx, "swgpu_decode_merge: bad argument");
        return SWGPU_E_INVALID;
    }
    *out_rows = 0;
    if (n == 0) return SWGPU_OK;
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    const uint64_t off0 = rec_off[0], total = rec_off[n] - off0;
    if (int rc = sw_ensure(ctx, ctx->d_bytes, total + 64)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_off, (size_t)(n + 1) * 8)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_status, (size_t)n + 16)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_first, 64)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_ids, (size_t)n * 16 + 64)) return rc;
    if (int rc = sw_ensure_pinned(ctx, ctx->h_small, 64)) return rc;
    if (total) SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_bytes.p, bytes + off0, total, cudaMemcpyHostToDevice, s));
    SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_off.p, rec_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, s));
    uint64_t *ids = (uint64_t *)ctx->d_ids.p;
    if (!enriched) {
        SW_CUDA(ctx, cudaMemcpyAsync(ids, id_hi, (size_t)n * 8, cudaMemcpyHostToDevice, s));
        SW_CUDA(ctx, cudaMemcpyAsync(ids + n, id_lo, (size_t)n * 8, cudaMemcpyHostToDevice, s));
    }
    uint32_t *d_rows = (uint32_t *)ctx->d_first.p;
    // record offsets are absolute: bias the byte pointer so that bytes_dev[off] is the staged copy
    const uint8_t *biased = (const uint8_t *)ctx->d_bytes.p - off0;
    for (int attempt = 0; attempt < 4; attempt++) {
        if (int rc = swgpu_decode_merge_dev(ctx, format, biased, (const uint64_t *)ctx->d_off.p, n, seq_base, ids, ids + n,
                                            (uint8_t *)ctx->d_status.p, d_rows))
            return rc;
        if (status && attempt == 0) SW_CUDA(ctx, cudaMemcpyAsync(status, ctx->d_status.p, n, cudaMemcpyDeviceToHost, s));
        SW_CUDA(ctx, cudaMemcpyAsync(ctx->h_small.p, d_rows, 4, cudaMemcpyDeviceToHost, s));
        const int rc = sw_state_after(ctx, n, "swgpu_decode_merge");  // (synchronises: status and the row count have landed)
        *out_rows = *(const uint32_t *)ctx->h_small.p;
        if (rc <= 0) return rc;
    }
    SW_SET_ERR(ctx, "swgpu_decode_merge: the state table could not be grown to hold the batch");
    return SWGPU_E_CAPACITY;
}

// Number of live keys and bytes of their public form; reports (once) rows that a *_dev merge had to drop.
extern "C" int32_t swgpu_state_size(swgpu_ctx *ctx, uint64_t *n_keys, uint64_t *key_bytes) {
    if (!ctx) return SWGPU_E_INVALID;
    MgState *st = (MgState *)ctx->state;
    if (n_keys) *n_keys = 0;
    if (key_bytes) *key_bytes = 0;
    if (!st) return SWGPU_OK;
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    MgCounters c;
    if (int rc = mg_read_counters(ctx, st, &c)) return rc;
    if (n_keys) *n_keys = c.n_keys;
    if (key_bytes) *key_bytes = c.key_bytes;
    st->keys_upper = c.n_keys;
    if (c.table_full || c.arena_full || c.key_too_long || c.fix_overflow) {
        SW_CUDA(ctx, cudaMemsetAsync(&st->ctr->table_full, 0, 3 * sizeof(unsigned long long), ctx->stream));
        SW_CUDA(ctx, cudaMemsetAsync(&st->ctr->fix_overflow, 0, sizeof(unsigned long long), ctx->stream));
        SW_SET_ERR(ctx,
                   "device-resident merges since the last check dropped rows or left payloads unsettled (table_full=%llu "
                   "long_key_words_without_room=%llu key_too_long=%llu fixup_overflow=%llu): call swgpu_state_reserve and replay "
                   "those batches (a replay is idempotent)",
                   c.table_full, c.arena_full, c.key_too_long, c.fix_overflow);
        return SWGPU_E_CAPACITY;
    }
    return SWGPU_OK;
}

extern "C" int32_t swgpu_state_clear(swgpu_ctx *ctx) {
    if (!ctx) return SWGPU_E_INVALID;
    MgState *st = (MgState *)ctx->state;
    if (!st) return SWGPU_OK;
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    if (st->slots) SW_CUDA(ctx, cudaMemsetAsync(st->slots, 0, st->cap * sizeof(MgSlot), ctx->stream));
    SW_CUDA(ctx, cudaMemsetAsync(st->ctr, 0, sizeof(MgCounters), ctx->stream));
    st->keys_upper = 0;
    return SWGPU_OK;
}

// ---- presence scan ------------------------------------------------------------------------------------------------------
// DevicePresenceManager.PresenceChecker (service-device-state/.../presence/DevicePresenceManager.java:139-167): device
// states with lastInteractionDate <= now - interval (MongoDeviceStateManagement.searchDeviceStates :170-193) for
// which SendOnceNotificationStrategy permits an event (presenceMissingDate == null,
// PresenceNotificationStrategies.java:24-31) are reported and get presenceMissingDate := now.  One streaming pass
// over the table: sector 0 of every slot (tag, best, key header), the date word of the few that match.
struct MgPresRow {
    uint64_t last, key_pos;  // key_pos: byte offset | length << 40
};
__global__ void merge_presence_scan_kernel(MgSlot *slots, uint64_t cap, const uint64_t *arena, uint64_t cutoff1, uint64_t now_ms,
                                           MgPresRow *rows, uint64_t max_rows, uint8_t *keys, uint64_t max_key_bytes,
                                           unsigned long long *cursor /* [0] found, [1] key bytes, [2] reported */) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cap) return;
    uint64_t *w = slots[i].w;
    const ulonglong2 a0 = __ldg((const ulonglong2 *)w), a1 = __ldg((const ulonglong2 *)(w + 2));
    const uint64_t tag = a0.x, best = a0.y, hdr = a1.x;
    if (tag == 0 || tag == MG_TOMB || (uint32_t)((hdr >> 32) & 0xff) != MG_KIND_DEVICE) return;
    if (best == 0 || best > cutoff1) return;      // interacted after the cut-off
    if (w[MG_W_PAY + 2] != 0) return;             // already reported missing
    atomicAdd(cursor, 1ull);
    const uint32_t tl = (uint32_t)(hdr & 0xffff), kl = 3 + tl;
    const unsigned long long k = atomicAdd(cursor + 2, 1ull);
    if (k >= max_rows) {
        atomicAdd(cursor + 2, (unsigned long long)-1ll);
        return;  // no room in the caller's arrays: stays unreported, the next scan finds it again
    }
    const unsigned long long pos = atomicAdd(cursor + 1, (unsigned long long)kl);
    if (pos + kl > max_key_bytes) {
        atomicAdd(cursor + 1, (unsigned long long)(0ull - kl));
        atomicAdd(cursor + 2, (unsigned long long)-1ll);
        return;
    }
    w[MG_W_PAY + 2] = now_ms;  // presenceMissingDate
    rows[k].last = best - 1;
    rows[k].key_pos = pos | ((uint64_t)kl << 40);
    uint8_t *pk = keys + pos;
    pk[0] = (uint8_t)(tl & 0xff);
    pk[1] = (uint8_t)(tl >> 8);
    pk[2] = (uint8_t)MG_KIND_DEVICE;
    const bool lng = (hdr & MG_HDR_LONG) != 0;
    const uint64_t *aw = lng ? arena + w[MG_W_KEY0 + 1] : nullptr;
    for (uint32_t b = 0; b < tl; b++) {
        const uint32_t j = 1 + (b >> 3);
        const uint64_t kwd = lng ? aw[j - 1] : (j < 6 ? w[MG_W_KEY0 + j] : w[MG_W_KEY6 + j - 6]);
        pk[3 + b] = (uint8_t)(kwd >> (8 * (b & 7)));
    }
}

extern "C" int32_t swgpu_presence_scan(swgpu_ctx *ctx, int64_t now_ms, int64_t interval_ms, uint64_t max_devices,
                                       uint64_t max_key_bytes, int64_t *last_interaction, uint64_t *key_off, uint8_t *key_bytes,
                                       uint64_t *n_reported, uint64_t *n_found) {
    if (!ctx || !n_reported || now_ms <= 0 || interval_ms < 0 || (max_devices && (!last_interaction || !key_off || !key_bytes))) {
        if (ctx) SW_SET_ERR(ctx, "swgpu_presence_scan: bad argument");
        return SWGPU_E_INVALID;
    }
    *n_reported = 0;
    if (n_found) *n_found = 0;
    if (key_off) key_off[0] = 0;
    MgState *st = (MgState *)ctx->state;
    if (!st || !st->slots) return SWGPU_OK;
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    MgPresRow *d_rows = nullptr;
    uint8_t *d_keys = nullptr;
    unsigned long long *d_cur = nullptr, h_cur[3] = {0, 0, 0};
    cudaError_t e = cudaMalloc(&d_rows, (max_devices + 1) * sizeof(MgPresRow));
    if (e == cudaSuccess) e = cudaMalloc(&d_keys, max_key_bytes + 16);
    if (e == cudaSuccess) e = cudaMalloc(&d_cur, 32);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_cur, 0, 32, ctx->stream);
    if (e == cudaSuccess) {
        const uint64_t cutoff1 = now_ms > interval_ms ? (uint64_t)(now_ms - interval_ms) + 1 : 0;  // best <= cutoff + 1
        merge_presence_scan_kernel<<<(unsigned)((st->cap + 255) / 256), 256, 0, ctx->stream>>>(
            st->slots, st->cap, st->arena, cutoff1, (uint64_t)now_ms, d_rows, max_devices, d_keys, max_key_bytes, d_cur);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(h_cur, d_cur, 24, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    std::vector<MgPresRow> rows(h_cur[2]);
    std::vector<uint8_t> keys(h_cur[1] + 1);
    if (e == cudaSuccess && h_cur[2]) {
        e = cudaMemcpyAsync(rows.data(), d_rows, h_cur[2] * sizeof(MgPresRow), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(keys.data(), d_keys, h_cur[1], cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    }
    if (d_rows) cudaFree(d_rows);
    if (d_keys) cudaFree(d_keys);
    if (d_cur) cudaFree(d_cur);
    if (e != cudaSuccess) {
        SW_SET_ERR(ctx, "swgpu_presence_scan failed: %s", cudaGetErrorString(e));
        return e == cudaErrorMemoryAllocation ? SWGPU_E_NOMEM : SWGPU_E_CUDA;
    }
    // the reference's query sorts ascending by lastInteractionDate; ties by key bytes for a defined order
    std::sort(rows.begin(), rows.end(), [&](const MgPresRow &x, const MgPresRow &y) {
        if (x.last != y.last) return x.last < y.last;
        const uint64_t xp = x.key_pos & ((1ull << 40) - 1), xl = x.key_pos >> 40, yp = y.key_pos & ((1ull << 40) - 1), yl = y.key_pos >> 40;
        const int c = memcmp(keys.data() + xp, keys.data() + yp, (size_t)std::min(xl, yl));
        return c ? c < 0 : xl < yl;
    });
    uint64_t off = 0;
    for (uint64_t k = 0; k < rows.size(); k++) {
        last_interaction[k] = (int64_t)rows[k].last;
        key_off[k] = off;
        const uint64_t pos = rows[k].key_pos & ((1ull << 40) - 1), kl = rows[k].key_pos >> 40;
        memcpy(key_bytes + off, keys.data() + pos, kl);
        off += kl;
    }
    if (key_off) key_off[rows.size()] = off;
    *n_reported = rows.size();
    if (n_found) *n_found = h_cur[0];
    if (h_cur[0] > rows.size()) {
        SW_SET_ERR(ctx, "swgpu_presence_scan: %llu devices are missing, %llu fit the arrays and were reported; scan again for the rest",
                   h_cur[0], (unsigned long long)rows.size());
        return SWGPU_E_CAPACITY;
    }
    return SWGPU_OK;
}

// ---- snapshot / restore: checkpoint operations, not the hot path ---------------------------------------------------
// Snapshot: a kernel packs the live slots into dense rows and their keys, in the public form
// token_len(2 LE) | kind byte | token | name, into one byte buffer; only those are copied to the host.
struct MgSnapRow {
    uint64_t seq, id_hi, id_lo, date, value, key_pos;  // key_pos: byte offset | length << 40
};
__global__ void merge_snapshot_kernel(const MgSlot *slots, uint64_t cap, const uint64_t *arena, MgSnapRow *rows, uint8_t *keys,
                                      unsigned long long *cursor /* [0] rows, [1] key bytes */) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cap) return;
    const uint64_t *w = slots[i].w;
    const uint64_t tag = w[MG_W_TAG];
    if (tag == 0 || tag == MG_TOMB) return;
    const uint64_t hdr = w[MG_W_KEY0];
    const uint32_t tl = (uint32_t)(hdr & 0xffff), nl = (uint32_t)((hdr >> 16) & 0xffff);
    const uint32_t kl = 3 + tl + nl;
    const unsigned long long k = atomicAdd(cursor, 1ull), pos = atomicAdd(cursor + 1, (unsigned long long)kl);
    MgSnapRow r;
    r.seq = w[MG_W_BEST] - 1;
    r.id_hi = w[MG_W_PAY];
    r.id_lo = w[MG_W_PAY + 1];
    r.date = w[MG_W_PAY + 2];
    r.value = w[MG_W_PAY + 3];
    r.key_pos = pos | ((uint64_t)kl << 40);
    rows[k] = r;
    uint8_t *pk = keys + pos;
    pk[0] = (uint8_t)(tl & 0xff);
    pk[1] = (uint8_t)(tl >> 8);
    pk[2] = (uint8_t)(hdr >> 32);
    const bool lng = (hdr & MG_HDR_LONG) != 0;
    const uint64_t *aw = lng ? arena + w[MG_W_KEY0 + 1] : nullptr;
    const uint32_t tw = (tl + 7) >> 3;
    auto key_word = [&](uint32_t j) -> uint64_t {  // key word j >= 1
        if (lng) return aw[j - 1];
        return j < 6 ? w[MG_W_KEY0 + j] : w[MG_W_KEY6 + j - 6];
    };
    for (uint32_t b = 0; b < tl; b++) pk[3 + b] = (uint8_t)(key_word(1 + (b >> 3)) >> (8 * (b & 7)));
    for (uint32_t b = 0; b < nl; b++) pk[3 + tl + b] = (uint8_t)(key_word(1 + tw + (b >> 3)) >> (8 * (b & 7)));
}

extern "C" int32_t swgpu_state_snapshot(swgpu_ctx *ctx, uint64_t max_keys, uint64_t max_key_bytes,
                                        const swgpu_state_rows *out, uint64_t *n_keys) {
    if (!ctx || !out || !n_keys) return SWGPU_E_INVALID;
    *n_keys = 0;
    MgState *st = (MgState *)ctx->state;
    if (!st || !st->slots) {
        if (out->key_off) out->key_off[0] = 0;
        return SWGPU_OK;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    MgCounters c;
    if (int rc = mg_read_counters(ctx, st, &c)) return rc;
    if (c.n_keys > max_keys || c.key_bytes > max_key_bytes) {
        *n_keys = c.n_keys;
        SW_SET_ERR(ctx, "snapshot needs %llu keys / %llu key bytes", c.n_keys, c.key_bytes);
        return SWGPU_E_CAPACITY;
    }
    if (c.n_keys == 0) {
        if (out->key_off) out->key_off[0] = 0;
        return SWGPU_OK;
    }
    MgSnapRow *d_rows = nullptr;
    uint8_t *d_keys = nullptr;
    unsigned long long *d_cur = nullptr;
    cudaError_t e = cudaMalloc(&d_rows, c.n_keys * sizeof(MgSnapRow));
    if (e == cudaSuccess) e = cudaMalloc(&d_keys, c.key_bytes + 16);
    if (e == cudaSuccess) e = cudaMalloc(&d_cur, 16);
    std::vector<MgSnapRow> rows(c.n_keys);
    std::vector<uint8_t> keys(c.key_bytes + 16);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_cur, 0, 16, ctx->stream);
    if (e == cudaSuccess) {
        merge_snapshot_kernel<<<(unsigned)((st->cap + 255) / 256), 256, 0, ctx->stream>>>(st->slots, st->cap, st->arena, d_rows,
                                                                                          d_keys, d_cur);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(rows.data(), d_rows, c.n_keys * sizeof(MgSnapRow), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(keys.data(), d_keys, c.key_bytes, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (d_rows) cudaFree(d_rows);
    if (d_keys) cudaFree(d_keys);
    if (d_cur) cudaFree(d_cur);
    if (e != cudaSuccess) {
        SW_SET_ERR(ctx, "swgpu_state_snapshot failed: %s", cudaGetErrorString(e));
        return e == cudaErrorMemoryAllocation ? SWGPU_E_NOMEM : SWGPU_E_CUDA;
    }
    uint64_t off = 0;
    for (uint64_t k = 0; k < c.n_keys; k++) {
        const MgSnapRow &r = rows[k];
        out->seq[k] = r.seq;
        out->id_hi[k] = r.id_hi;
        out->id_lo[k] = r.id_lo;
        out->event_date[k] = (int64_t)r.date;
        memcpy(&out->value[k], &r.value, 8);
        out->key_off[k] = off;
        const uint64_t pos = r.key_pos & ((1ull << 40) - 1), kl = r.key_pos >> 40;
        memcpy(out->key_bytes + off, keys.data() + pos, kl);
        off += kl;
    }
    out->key_off[c.n_keys] = off;
    *n_keys = c.n_keys;
    return SWGPU_OK;
}

// Restore: the rows and their public-form keys are uploaded and inserted by the same upsert the merge uses.
struct MgKeyPublic {  // key words from a public-form key in global memory (byte loads: checkpoint path)
    const uint8_t *tok, *name;
    uint32_t tl, nl;
    __device__ __forceinline__ uint64_t piece(const uint8_t *p, uint32_t valid) const {
        uint64_t v = 0;
        for (uint32_t k = 0; k < valid; k++) v |= (uint64_t)p[k] << (8 * k);
        return v;
    }
    __device__ __forceinline__ uint64_t word(uint32_t i) const {
        const uint32_t tw = (tl + 7u) >> 3;
        if (i <= tw) return piece(tok + 8 * (i - 1), min(8u, tl - 8 * (i - 1)));
        const uint32_t j = i - 1 - tw;
        return piece(name + 8 * j, min(8u, nl - 8 * j));
    }
};
__global__ void __launch_bounds__(MG_THREADS) merge_restore_kernel(MgTable t, const uint64_t *seq, const uint64_t *id_hi,
                                                                   const uint64_t *id_lo, const int64_t *date, const double *value,
                                                                   const uint64_t *key_off, const uint8_t *key_bytes, uint64_t n) {
    __shared__ MgLocalCnt s_cnt[MG_THREADS / 32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) s_cnt[warp].keys = s_cnt[warp].bytes = 0;
    __syncwarp();
    const uint64_t k = (uint64_t)blockIdx.x * MG_THREADS + threadIdx.x;
    if (k < n) {
        const uint8_t *key = key_bytes + key_off[k];
        const uint32_t kl = (uint32_t)(key_off[k + 1] - key_off[k]);
        const uint32_t tl = key[0] | ((uint32_t)key[1] << 8), kindbyte = key[2], nl = kl - 3 - tl;
        const uint32_t kind = kindbyte & 0x7f;
        GmemBytes gb{key};
        const uint64_t th = sw_hash64(gb, 3, tl);
        const uint64_t nh = (kind == SWGPU_KIND_LOCATION || (kindbyte & 0x80)) ? 0ull : sw_hash64(gb, 3 + tl, nl);
        MgKeyPublic kr{key + 3, key + 3 + tl, tl, nl};
        mg_upsert(t, &s_cnt[warp], mg_mix(th, nh, kindbyte), th, tl, nl, kindbyte, kr, seq[k] + 1, id_hi[k], id_lo[k],
                  (uint64_t)date[k], (uint64_t)__double_as_longlong(value[k]));
    }
    __syncwarp();
    if (lane == 0) mg_flush_counts(t, &s_cnt[warp]);
}

extern "C" int32_t swgpu_state_restore(swgpu_ctx *ctx, const swgpu_state_rows *in, uint64_t n_keys) {
    if (!ctx || (n_keys && !in)) return SWGPU_E_INVALID;
    if (int rc = swgpu_state_clear(ctx)) return rc;
    if (n_keys == 0) return SWGPU_OK;
    uint64_t long_bytes = 0;
    for (uint64_t k = 0; k < n_keys; k++) {
        const uint64_t kl = in->key_off[k + 1] - in->key_off[k];
        const uint8_t *key = in->key_bytes + in->key_off[k];
        if (kl < 3 || 3 + (uint64_t)(key[0] | ((uint32_t)key[1] << 8)) > kl) {
            SW_SET_ERR(ctx, "swgpu_state_restore: malformed key %llu", (unsigned long long)k);
            return SWGPU_E_INVALID;
        }
        const uint32_t tl = key[0] | ((uint32_t)key[1] << 8);
        if (kl - 3 - tl > 0xffffull) {
            SW_SET_ERR(ctx, "swgpu_state_restore: key %llu too long", (unsigned long long)k);
            return SWGPU_E_INVALID;
        }
        const uint32_t nl = (uint32_t)(kl - 3 - tl);
        if (mg_key_words(tl, nl) > MG_KEY_INLINE_WORDS) long_bytes += 8ull * (mg_key_words(tl, nl) - 1);
    }
    if (int rc = mg_reserve(ctx, n_keys, long_bytes + 64)) return rc;
    MgTable t;
    if (int rc = sw_state_prepare(ctx, n_keys, n_keys, 0, &t)) return rc;
    t.seq_floor = ~0ull;  // keys are distinct: no row can follow another one of this launch
    t.presence_now1 = 0;  // (presence rows are among the snapshot's keys)
    const uint64_t kb = in->key_off[n_keys];
    uint8_t *d = nullptr;
    const size_t col = (size_t)n_keys * 8, total = col * 5 + (size_t)(n_keys + 1) * 8 + kb + 64;
    cudaError_t e = cudaMalloc(&d, total);
    if (e != cudaSuccess) {
        SW_SET_ERR(ctx, "swgpu_state_restore: cudaMalloc(%zu) failed: %s", total, cudaGetErrorString(e));
        return SWGPU_E_NOMEM;
    }
    const void *src[5] = {in->seq, in->id_hi, in->id_lo, in->event_date, in->value};
    for (int c = 0; c < 5 && e == cudaSuccess; c++)
        e = cudaMemcpyAsync(d + (size_t)c * col, src[c], col, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d + 5 * col, in->key_off, (size_t)(n_keys + 1) * 8, cudaMemcpyHostToDevice, ctx->stream);
    uint8_t *d_keys = d + 5 * col + (size_t)(n_keys + 1) * 8;
    if (e == cudaSuccess && kb) e = cudaMemcpyAsync(d_keys, in->key_bytes, kb, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) {
        merge_restore_kernel<<<(unsigned)((n_keys + MG_THREADS - 1) / MG_THREADS), MG_THREADS, 0, ctx->stream>>>(
            t, (const uint64_t *)d, (const uint64_t *)(d + col), (const uint64_t *)(d + 2 * col), (const int64_t *)(d + 3 * col),
            (const double *)(d + 4 * col), (const uint64_t *)(d + 5 * col), d_keys, n_keys);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d);
    if (e != cudaSuccess) {
        SW_SET_ERR(ctx, "swgpu_state_restore failed: %s", cudaGetErrorString(e));
        return SWGPU_E_CUDA;
    }
    const int rc = sw_state_after(ctx, n_keys, "swgpu_state_restore");
    if (rc > 0) {
        SW_SET_ERR(ctx, "swgpu_state_restore: the table could not hold the snapshot");
        return SWGPU_E_CAPACITY;
    }
    return rc;
}

// ---- alternate-id set --------------------------------------------------------------------
// AlternateIdDeduplicator.isDuplicate (service-event-sources/.../deduplicator/AlternateIdDeduplicator.java:42-57) asks
// the event store, per decoded request that carries an alternateId, whether an event with that id exists.  Here the
// context keeps the set of alternate ids it has been shown, in a second table of the device-state kind (key = the id's
// bytes, kind byte 0x7e, no name; `best` = a sequence number that DEcreases with arrival, so that the maximum is the
// first arrival).  Two passes of the same kernel over the rows: the first inserts / raises, the second reads who won.
#define DD_KIND 0x7eu
#define DD_SEQ_TOP (1ull << 62)
struct DdArgs {
    MgTable t;
    const uint8_t *bytes;
    const uint64_t *rec_off, *meta, *span;
    uint32_t n;
    const uint32_t *n_dev;
    uint64_t first_row;  // rows probed before this call
    uint8_t *dup;
};
struct DdKey {  // the id's bytes as key words
    const uint8_t *p;
    uint32_t len;
    __device__ __forceinline__ uint64_t word(uint32_t i) const { return mg_payload_word(p + 8 * (i - 1), min(8u, len - 8 * (i - 1))); }
};
template <int PASS>
__global__ void __launch_bounds__(MG_THREADS) dedup_kernel(DdArgs a) {
    __shared__ MgLocalCnt s_cnt[MG_THREADS / 32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) s_cnt[warp].keys = s_cnt[warp].bytes = 0;
    __syncwarp();
    const uint32_t n_live = a.n_dev ? min(__ldg(a.n_dev), a.n) : a.n;
    const uint32_t r = blockIdx.x * MG_THREADS + threadIdx.x;
    if (r < n_live) {
        const uint64_t m = __ldg(a.meta + r);
        const uint32_t fl = SWGPU_META_FLAGS(m);
        uint32_t verdict = SWGPU_DEDUP_NEW;
        if (fl & SWGPU_F_HAS_ALT_ID) {
            const uint64_t sp = __ldg(a.span + r);
            const uint32_t len = SWGPU_SPAN_LEN(sp);
            if ((fl & (SWGPU_F_STR_ESCAPED | SWGPU_F_NAME_COERCED)) || len > SWGPU_DEDUP_MAX_ID_BYTES) {
                // the id's characters are not the span's bytes, or the id does not fit a slot (the set keeps no arena:
                // a probe is asynchronous and cannot grow one): the store decides
                verdict = SWGPU_DEDUP_ASK;
            } else {
                const uint8_t *id = a.bytes + __ldg(a.rec_off + SWGPU_META_REC(m)) + SWGPU_SPAN_OFF(sp);
                DdKey key{id, len};
                const uint64_t ih = sw_hash64(GmemBytes{id}, 0, len);
                const uint64_t seq1 = DD_SEQ_TOP - (a.first_row + r);
                const unsigned long long old =
                    mg_upsert_one<true>(a.t, &s_cnt[warp], mg_mix(ih, 0ull, DD_KIND), len, 0u, DD_KIND, key, seq1, 0ull, 0ull, 0ull, 0ull);
                // second pass: `old` is the final maximum = the first arrival of this id
                verdict = old == MG_NOT_STORED ? SWGPU_DEDUP_ASK : (old > seq1 ? SWGPU_DEDUP_SEEN : SWGPU_DEDUP_NEW);
            }
        }
        if (PASS == 2) a.dup[r] = (uint8_t)verdict;
    }
    __syncwarp();
    if (lane == 0) mg_flush_counts(a.t, &s_cnt[warp]);
}

struct DdSwap {  // the state functions work on ctx->state: point it at the set for the duration of a call
    swgpu_ctx *c;
    explicit DdSwap(swgpu_ctx *ctx) : c(ctx) { std::swap(c->state, c->dedup_state); }
    ~DdSwap() { std::swap(c->state, c->dedup_state); }
};

extern "C" int32_t swgpu_dedup_probe_dev(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, const uint64_t *meta,
                                         const uint64_t *alternate_id_span, uint32_t n_rows, const uint32_t *n_rows_dev,
                                         uint8_t *dup) {
    if (!ctx) return SWGPU_E_INVALID;
    if (n_rows == 0) return SWGPU_OK;
    if (!bytes || !rec_off || !meta || !alternate_id_span || !dup) {
        SW_SET_ERR(ctx, "swgpu_dedup_probe_dev: NULL argument");
        return SWGPU_E_INVALID;
    }
    DdSwap swap(ctx);
    DdArgs a;
    if (int rc = sw_state_prepare(ctx, n_rows, n_rows, 0, &a.t)) return rc;
    a.t.presence_now1 = 0;
    a.bytes = bytes;
    a.rec_off = rec_off;
    a.meta = meta;
    a.span = alternate_id_span;
    a.n = n_rows;
    a.n_dev = n_rows_dev;
    a.first_row = ctx->dedup_rows;
    a.dup = dup;
    ctx->dedup_rows += n_rows;
    const unsigned nb = (n_rows + MG_THREADS - 1) / MG_THREADS;
    dedup_kernel<1><<<nb, MG_THREADS, 0, ctx->stream>>>(a);
    SW_CUDA(ctx, cudaGetLastError());
    dedup_kernel<2><<<nb, MG_THREADS, 0, ctx->stream>>>(a);
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
}

extern "C" int32_t swgpu_dedup_reserve(swgpu_ctx *ctx, uint64_t max_ids) {
    if (!ctx) return SWGPU_E_INVALID;
    DdSwap swap(ctx);
    return mg_reserve(ctx, max_ids, 0);
}

extern "C" int32_t swgpu_dedup_size(swgpu_ctx *ctx, uint64_t *n_ids) {
    if (!ctx || !n_ids) return SWGPU_E_INVALID;
    *n_ids = 0;
    DdSwap swap(ctx);
    MgState *st = (MgState *)ctx->state;
    if (!st || !st->ctr) return SWGPU_OK;
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    MgCounters c;
    if (int rc = mg_read_counters(ctx, st, &c)) return rc;
    *n_ids = c.n_keys;
    return SWGPU_OK;
}

extern "C" int32_t swgpu_dedup_clear(swgpu_ctx *ctx) {
    if (!ctx) return SWGPU_E_INVALID;
    DdSwap swap(ctx);
    ctx->dedup_rows = 0;
    return swgpu_state_clear(ctx);
}

// host form: `ids` = the alternate ids back to back, id_off[n + 1]; every row carries one.  Stages the ids, builds the
// two columns the device form reads (record index + "has an alternate id", span = the whole record), probes, waits.
__global__ void dedup_cols_kernel(const uint64_t *id_off, uint32_t n, uint64_t *meta, uint64_t *span) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    meta[i] = ((uint64_t)i << 32) | ((uint64_t)SWGPU_F_HAS_ALT_ID << 8);
    span[i] = (id_off[i + 1] - id_off[i]) << 32;
}
extern "C" int32_t swgpu_dedup_probe(swgpu_ctx *ctx, const uint8_t *ids, const uint64_t *id_off, uint32_t n, uint8_t *dup) {
    if (!ctx) return SWGPU_E_INVALID;
    if (n == 0) return SWGPU_OK;
    if (!id_off || !dup || (id_off[n] && !ids)) {
        SW_SET_ERR(ctx, "swgpu_dedup_probe: NULL argument");
        return SWGPU_E_INVALID;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    const uint64_t total = id_off[n];
    if (int rc = sw_ensure(ctx, ctx->d_bytes, total + 64)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_off, (size_t)(n + 1) * 8)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_cols, (size_t)n * 16 + 64)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_status, (size_t)n + 16)) return rc;
    cudaStream_t st = ctx->stream;
    if (total) SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_bytes.p, ids, total, cudaMemcpyHostToDevice, st));
    SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_off.p, id_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, st));
    uint64_t *meta = (uint64_t *)ctx->d_cols.p, *span = meta + n;
    dedup_cols_kernel<<<(n + 255) / 256, 256, 0, st>>>((const uint64_t *)ctx->d_off.p, n, meta, span);
    SW_CUDA(ctx, cudaGetLastError());
    if (int rc = swgpu_dedup_probe_dev(ctx, (const uint8_t *)ctx->d_bytes.p, (const uint64_t *)ctx->d_off.p, meta, span, n, nullptr,
                                       (uint8_t *)ctx->d_status.p))
        return rc;
    SW_CUDA(ctx, cudaMemcpyAsync(dup, ctx->d_status.p, n, cudaMemcpyDeviceToHost, st));
    SW_CUDA(ctx, cudaStreamSynchronize(st));
    return SWGPU_OK;
}
