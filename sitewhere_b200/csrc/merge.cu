// merge.cu — batched device-state last-writer merge
// (reference: service-device-state/src/main/java/com/sitewhere/devicestate/processing/
//  DeviceStateProcessingLogic.java:116-201 + persistence/DeviceStatePersistence.java:60-85: a sequential
//  fold in arrival order; lastLocationEventId := id, lastMeasurementEventIds[name] := id,
//  lastAlertEventIds[type] := id).
//
// Sort-free: the sequential "later event overwrites" is, per key, an argmax over the arrival
// sequence number.  A row finds/claims its key's slot in an open-addressing table (hash picks the
// bucket, the key BYTES decide equality), does atomicMax(slot.best, seq+1) and, if it raised the
// maximum, writes its payload (event id, date, value) at once; the rare rows that raised a maximum set
// by an older row of the same launch go on a fix-up list that a second, tiny kernel replays.  Rows
// are taken newest first, so most older rows of a key lose their atomicMax and leave.  HBM-bound: one
// 64-byte slot and one arena key per event, random access.
#include <algorithm>

#include "swgpu_common.cuh"

#define MG_THREADS 256
#define MG_PENDING 2ull

struct MgSlot {  // 64 bytes = two sectors: {tag,best,key_off,key_len} is what pass 1 touches
    uint64_t tag;   // 0 = empty; else (hash & ~3) | 1, bit1 = key bytes not yet published
    uint64_t best;  // seq + 1 of the current winner
    uint64_t key_off;
    uint32_t key_len, pad;
    uint64_t id_hi, id_lo;
    int64_t date;
    double value;
};
static_assert(sizeof(MgSlot) == 64, "MgSlot layout");

struct MgCounters {  // device
    unsigned long long n_keys, arena_top, table_full, arena_full, key_too_long;
    unsigned long long n_fix;  // rows on this launch's fix-up list (reset by the launcher)
};

struct MgState {
    MgSlot *slots = nullptr;
    uint64_t cap = 0;  // power of two
    uint8_t *arena = nullptr;
    uint64_t arena_cap = 0;
    MgCounters *ctr = nullptr;
    uint64_t keys_upper = 0;   // host upper bound on live keys
    uint64_t arena_upper = 0;  // host upper bound on arena bytes
    swgpu_ctx::Buf fix;        // fix-up list of the current launch (row + slot per entry)
};

__host__ __device__ __forceinline__ uint64_t mg_mix(uint64_t th, uint64_t nh, uint32_t kindbyte) {
    uint64_t h = th ^ (nh * 0x9e3779b97f4a7c15ull) ^ ((uint64_t)kindbyte << 56);
    h ^= h >> 33;
    h *= 0xff51afd7ed558ccdull;
    h ^= h >> 33;
    h *= 0xc4ceb9fe1a85ec53ull;
    h ^= h >> 33;
    return h;
}

struct MgArgs {
    MgSlot *slots;
    uint64_t mask;
    uint8_t *arena;
    uint64_t arena_cap;
    MgCounters *ctr;
    const uint8_t *bytes;
    const uint64_t *rec_off;
    const uint64_t *meta, *token_hash, *name_hash, *token_span, *name_span;
    const int64_t *event_date;
    const double *v0;
    uint32_t n_rows;
    const uint32_t *n_rows_dev;
    uint64_t seq_base;
    const uint64_t *id_hi, *id_lo;
    uint32_t ids_by_record;  // 1: id arrays are indexed by source record (every row of a record carries its id)
    uint32_t *fix_row;   // fix-up list: row ...
    uint64_t *fix_slot;  // ... and its slot
};


// Keys live in the arena as 8-byte words so that equality is a handful of 64-bit compares:
//   word 0            token length | name length << 16 | kind byte << 32
//   ceil(tl/8) words  token bytes, zero padded
//   ceil(nl/8) words  name bytes, zero padded
// (the snapshot converts to the public form  token_len(2) | kind | token | name).
__host__ __device__ __forceinline__ uint64_t mg_key_hdr(uint32_t tl, uint32_t nl, uint32_t kindbyte) {
    return (uint64_t)tl | ((uint64_t)nl << 16) | ((uint64_t)kindbyte << 32);
}
__host__ __device__ __forceinline__ uint32_t mg_key_bytes(uint32_t tl, uint32_t nl) {
    return 8u + ((tl + 7u) & ~7u) + ((nl + 7u) & ~7u);
}
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
__device__ __forceinline__ bool mg_key_equal(const uint64_t *arena_key, uint32_t key_len, uint32_t tl,
                                             uint32_t kindbyte, const uint8_t *tok, const uint8_t *name,
                                             uint32_t nl) {
    if (key_len != mg_key_bytes(tl, nl)) return false;
    // arena words were written during this launch by other SMs: read them at L2 (never from a stale L1 line)
    if (__ldcg(arena_key) != mg_key_hdr(tl, nl, kindbyte)) return false;
    uint32_t w = 1;
    for (uint32_t i = 0; i < tl; i += 8, w++)
        if (__ldcg(arena_key + w) != mg_payload_word(tok + i, min(8u, tl - i))) return false;
    for (uint32_t i = 0; i < nl; i += 8, w++)
        if (__ldcg(arena_key + w) != mg_payload_word(name + i, min(8u, nl - i))) return false;
    return true;
}
__device__ __forceinline__ uint64_t mg_ld_acquire_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(MG_THREADS) merge_upsert_kernel(MgArgs a) {
    const uint32_t n_live = a.n_rows_dev ? min(__ldg(a.n_rows_dev), a.n_rows) : a.n_rows;
    const uint32_t ridx = blockIdx.x * MG_THREADS + threadIdx.x;
    if (ridx >= n_live) return;
    // newest rows first: CTAs start roughly in index order, so for a key with several rows in the batch the
    // winner tends to arrive first and the older rows lose their atomicMax at once (no publish work for them)
    const uint32_t r = n_live - 1 - ridx;
    const uint64_t m = __ldg(a.meta + r);
    const uint32_t kind = SWGPU_META_KIND(m), fl = SWGPU_META_FLAGS(m);
    if (kind > SWGPU_KIND_ALERT || (fl & SWGPU_F_NULL_TOKEN)) return;
    const uint8_t *rec = a.bytes + __ldg(a.rec_off + SWGPU_META_REC(m));
    const uint64_t ts = __ldg(a.token_span + r), ns = __ldg(a.name_span + r);
    const uint32_t tl = SWGPU_SPAN_LEN(ts);
    const uint32_t nl = kind == SWGPU_KIND_LOCATION ? 0u : SWGPU_SPAN_LEN(ns);
    if (tl > 0xffffu || nl > 0xffffu) {
        atomicAdd(&a.ctr->key_too_long, 1ull);
        return;
    }
    const uint8_t *tok = rec + SWGPU_SPAN_OFF(ts), *name = rec + SWGPU_SPAN_OFF(ns);
    const uint32_t kindbyte = kind | ((kind != SWGPU_KIND_LOCATION && !(fl & SWGPU_F_HAS_NAME)) ? 0x80u : 0u);
    const uint64_t h = mg_mix(__ldg(a.token_hash + r), kind == SWGPU_KIND_LOCATION ? 0ull : __ldg(a.name_hash + r),
                              kindbyte);
    const uint64_t want = (h & ~3ull) | 1ull;
    uint64_t s = (h >> 2) & a.mask;
    // The row's key as arena words, fetched BEFORE the table is probed so that the payload loads and the
    // (random, DRAM-latency) slot load are in flight together.  Keys of up to 32 + 16 bytes stay in
    // registers; longer ones are compared word by word on demand.
    const bool fast = tl <= 32 && nl <= 16;
    const uint32_t tw = (tl + 7) >> 3, nw = (nl + 7) >> 3;
    uint64_t kw[7];
    kw[0] = mg_key_hdr(tl, nl, kindbyte);
    if (fast) {
#pragma unroll
        for (uint32_t i = 0; i < 4; i++) kw[1 + i] = i < tw ? mg_payload_word(tok + 8 * i, min(8u, tl - 8 * i)) : 0ull;
#pragma unroll
        for (uint32_t i = 0; i < 2; i++) kw[5 + i] = i < nw ? mg_payload_word(name + 8 * i, min(8u, nl - 8 * i)) : 0ull;
    }
    // the payload this row would store (most rows do: one row per key per batch is the common case)
    const uint32_t idx = a.ids_by_record ? SWGPU_META_REC(m) : r;
    const uint64_t v_hi = __ldg(a.id_hi + idx), v_lo = __ldg(a.id_lo + idx);
    const int64_t v_date = __ldg(a.event_date + r);
    const double v_val = __ldg(a.v0 + r);
    for (uint64_t probes = 0; probes <= a.mask; probes++, s = (s + 1) & a.mask) {
        MgSlot *sl = a.slots + s;
        uint64_t tag = mg_ld_acquire_u64(&sl->tag);  // acquire: a published tag makes the key words visible
        if (tag == 0) {
            unsigned long long old =
                atomicCAS((unsigned long long *)&sl->tag, 0ull, (unsigned long long)(want | MG_PENDING));
            if (old == 0) {
                // claimed: publish the key words, then clear the pending bit
                const uint32_t kl = mg_key_bytes(tl, nl);
                unsigned long long off = atomicAdd(&a.ctr->arena_top, (unsigned long long)kl);
                if (off + kl > a.arena_cap) {
                    atomicAdd(&a.ctr->arena_full, 1ull);
                    sl->key_len = 0xffffffffu;  // poisoned slot: never equal to any key
                    sl->key_off = 0;
                    __threadfence();
                    sw_st_volatile_u64(&sl->tag, want);
                    return;
                }
                uint64_t *k = (uint64_t *)(a.arena + off);
                uint32_t w = 0;
                k[w++] = mg_key_hdr(tl, nl, kindbyte);
                for (uint32_t i = 0; i < tl; i += 8) k[w++] = mg_payload_word(tok + i, min(8u, tl - i));
                for (uint32_t i = 0; i < nl; i += 8) k[w++] = mg_payload_word(name + i, min(8u, nl - i));
                sl->key_off = off;
                sl->key_len = kl;
                atomicAdd(&a.ctr->n_keys, 1ull);
                __threadfence();
                sw_st_volatile_u64(&sl->tag, want);
                break;
            }
            tag = old;
        }
        if ((tag & ~MG_PENDING) == want) {
            while (tag & MG_PENDING) tag = mg_ld_acquire_u64(&sl->tag);  // the claimer is publishing (rare)
            const uint32_t kl = __ldcg(&sl->key_len);
            const uint64_t ko = __ldcg(&sl->key_off);
            if (kl == mg_key_bytes(tl, nl)) {  // (a poisoned slot has key_len 0xffffffff)
                const uint64_t *ak = (const uint64_t *)(a.arena + ko);
                if (fast) {
                    // all arena words requested at once, then compared with the registers
                    uint64_t av[7];
#pragma unroll
                    for (uint32_t i = 0; i < 7; i++) {
                        const bool on = i == 0 || (i >= 1 && i - 1 < tw) || (i >= 5 && i - 5 < nw);
                        const uint32_t widx = i == 0 ? 0u : (i < 5 ? i : 1 + tw + (i - 5));
                        av[i] = on ? __ldcg(ak + widx) : 0ull;
                    }
                    bool eq = true;
#pragma unroll
                    for (uint32_t i = 0; i < 7; i++) eq = eq && av[i] == kw[i];
                    if (eq) break;
                } else if (mg_key_equal(ak, kl, tl, kindbyte, tok, name, nl))
                    break;
            }
        }
        if (probes == a.mask) {
            atomicAdd(&a.ctr->table_full, 1ull);
            return;
        }
    }
    // ---- last writer wins.  `best` = seq + 1 of the newest event seen for the key; the row that raises it owns
    // the slot's payload at that moment and writes it here, in the same pass.  Two rows of one launch can both
    // raise it (an older row got there first, then a newer one): their payload stores may land in either order,
    // so the newer row — which sees a previous value from this same launch — also puts itself on a fix-up
    // list, and a second, tiny kernel lets the final winner of each such key write once more after every store
    // of this kernel has landed.  With the newest rows first this is rare (same-key rows must be in flight
    // together): the list is a few percent of the rows at most, against a second pass over all of them.
    const unsigned long long mine = (unsigned long long)(a.seq_base + r + 1);
    const unsigned long long old = atomicMax((unsigned long long *)&a.slots[s].best, mine);
    if (old >= mine) return;  // an event at least as new is already there (also: replay of the same batch)
    MgSlot *sl = a.slots + s;
    sl->id_hi = v_hi;
    sl->id_lo = v_lo;
    sl->date = v_date;
    sl->value = v_val;
    if (old > a.seq_base) {  // the previous value was set by an older row of this launch
        const unsigned long long k = atomicAdd(&a.ctr->n_fix, 1ull);
        a.fix_row[k] = r;
        a.fix_slot[k] = s;
    }
}

// second pass, over the fix-up list only
__global__ void __launch_bounds__(MG_THREADS) merge_fixup_kernel(MgArgs a) {
    const unsigned long long n = a.ctr->n_fix;
    for (unsigned long long k = (unsigned long long)blockIdx.x * MG_THREADS + threadIdx.x; k < n;
         k += (unsigned long long)gridDim.x * MG_THREADS) {
        const uint32_t r = a.fix_row[k];
        MgSlot *sl = a.slots + a.fix_slot[k];
        if (sl->best != a.seq_base + r + 1) continue;  // a later event of this key won
        const uint32_t idx = a.ids_by_record ? SWGPU_META_REC(__ldg(a.meta + r)) : r;
        sl->id_hi = __ldg(a.id_hi + idx);
        sl->id_lo = __ldg(a.id_lo + idx);
        sl->date = __ldg(a.event_date + r);
        sl->value = __ldg(a.v0 + r);
    }
}

// re-insert live slots into a larger table (keys are distinct: no byte compare needed)
__global__ void merge_rehash_kernel(const MgSlot *old_slots, uint64_t old_cap, MgSlot *slots, uint64_t mask) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= old_cap) return;
    MgSlot v = old_slots[i];
    if (v.tag == 0) return;
    uint64_t s = (v.tag >> 2) & mask;
    for (;;) {
        unsigned long long old = atomicCAS((unsigned long long *)&slots[s].tag, 0ull, (unsigned long long)v.tag);
        if (old == 0) break;
        s = (s + 1) & mask;
    }
    MgSlot *d = slots + s;
    d->best = v.best;
    d->key_off = v.key_off;
    d->key_len = v.key_len;
    d->id_hi = v.id_hi;
    d->id_lo = v.id_lo;
    d->date = v.date;
    d->value = v.value;
}

void sw_state_free(swgpu_ctx *ctx) {
    MgState *st = (MgState *)ctx->state;
    if (!st) return;
    if (st->slots) cudaFree(st->slots);
    if (st->arena) cudaFree(st->arena);
    if (st->ctr) cudaFree(st->ctr);
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
    SW_CUDA(ctx, cudaMemcpyAsync(out, st->ctr, sizeof(MgCounters), cudaMemcpyDeviceToHost, ctx->stream));
    SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SWGPU_OK;
}

// make room for `max_keys` keys and `max_key_bytes` arena bytes (grows, never shrinks)
static int mg_reserve(swgpu_ctx *ctx, uint64_t max_keys, uint64_t max_key_bytes) {
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    MgState *st = (MgState *)ctx->state;
    if (!st) {
        st = new MgState();
        ctx->state = st;
        cudaError_t e = cudaMalloc(&st->ctr, sizeof(MgCounters));
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
    if (max_key_bytes > st->arena_cap) {
        const uint64_t want = max_key_bytes + max_key_bytes / 4 + 4096;
        uint8_t *na = nullptr;
        cudaError_t e = cudaMalloc(&na, want);
        if (e != cudaSuccess) {
            SW_SET_ERR(ctx, "cudaMalloc(key arena %llu) failed: %s", (unsigned long long)want, cudaGetErrorString(e));
            return SWGPU_E_NOMEM;
        }
        if (st->arena) {
            MgCounters c;
            if (int rc = mg_read_counters(ctx, st, &c)) return rc;
            uint64_t used = c.arena_top < st->arena_cap ? c.arena_top : st->arena_cap;
            if (used) SW_CUDA(ctx, cudaMemcpyAsync(na, st->arena, used, cudaMemcpyDeviceToDevice, ctx->stream));
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

int sw_merge_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n_records,
                    const swgpu_decoded_cols *rows, uint32_t n_rows, const uint32_t *n_rows_dev, uint64_t seq_base,
                    const uint64_t *id_hi, const uint64_t *id_lo, bool ids_by_record);

extern "C" int32_t swgpu_merge_dev(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off,
                                   uint32_t n_records, const swgpu_decoded_cols *rows, uint32_t n_rows,
                                   const uint32_t *n_rows_dev, uint64_t seq_base, const uint64_t *id_hi,
                                   const uint64_t *id_lo) {
    return sw_merge_launch(ctx, bytes, rec_off, n_records, rows, n_rows, n_rows_dev, seq_base, id_hi, id_lo, false);
}

int sw_merge_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n_records,
                    const swgpu_decoded_cols *rows, uint32_t n_rows, const uint32_t *n_rows_dev, uint64_t seq_base,
                    const uint64_t *id_hi, const uint64_t *id_lo, bool ids_by_record) {
    (void)n_records;
    if (!ctx) return SWGPU_E_INVALID;
    if (n_rows == 0) return SWGPU_OK;
    if (!bytes || !rec_off || !rows || !rows->meta || !rows->token_hash || !rows->name_hash || !rows->token_span ||
        !rows->name_span || !rows->event_date || !rows->v0 || !id_hi || !id_lo) {
        SW_SET_ERR(ctx, "swgpu_merge_dev: NULL argument");
        return SWGPU_E_INVALID;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    MgState *st = (MgState *)ctx->state;
    // capacity: every row could be a new key of up to (avg) 64 key bytes; tighten the host-side
    // upper bounds by reading the device counters only when they would force a growth
    {
        uint64_t ku = (st ? st->keys_upper : 0) + n_rows;
        uint64_t au = (st ? st->arena_upper : 0) + (uint64_t)n_rows * 64;
        if (!st || ku * 2 + 16 > st->cap || au > st->arena_cap) {
            if (st && st->ctr) {
                MgCounters c;
                if (int rc = mg_read_counters(ctx, st, &c)) return rc;
                st->keys_upper = c.n_keys;
                st->arena_upper = c.arena_top;
                ku = st->keys_upper + n_rows;
                au = st->arena_upper + (uint64_t)n_rows * 64;
            }
            if (int rc = mg_reserve(ctx, ku, au)) return rc;
            st = (MgState *)ctx->state;
        }
        st->keys_upper = ku;
        st->arena_upper = au;
    }
    if (int rc = sw_ensure(ctx, st->fix, (size_t)n_rows * 12 + 64)) return rc;
    MgArgs a;
    a.slots = st->slots;
    a.mask = st->cap - 1;
    a.arena = st->arena;
    a.arena_cap = st->arena_cap;
    a.ctr = st->ctr;
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
    a.ids_by_record = ids_by_record ? 1u : 0u;
    a.fix_slot = (uint64_t *)st->fix.p;
    a.fix_row = (uint32_t *)((uint8_t *)st->fix.p + (((size_t)n_rows * 8 + 63) & ~(size_t)63));
    const unsigned nb = (n_rows + MG_THREADS - 1) / MG_THREADS;
    SW_CUDA(ctx, cudaMemsetAsync(&st->ctr->n_fix, 0, sizeof(unsigned long long), ctx->stream));
    merge_upsert_kernel<<<nb, MG_THREADS, 0, ctx->stream>>>(a);
    SW_CUDA(ctx, cudaGetLastError());
    merge_fixup_kernel<<<std::min<unsigned>(nb, (unsigned)ctx->sm_count * 4u), MG_THREADS, 0, ctx->stream>>>(a);
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
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
    int rc = swgpu_merge_dev(ctx, (const uint8_t *)ctx->d_bytes.p, (const uint64_t *)ctx->d_off.p, n_records, &d,
                             n_rows, nullptr, seq_base, ids, ids + n_rows);
    if (rc) return rc;
    MgState *st = (MgState *)ctx->state;
    MgCounters c;
    if (int rc2 = mg_read_counters(ctx, st, &c)) return rc2;
    st->keys_upper = c.n_keys;
    st->arena_upper = c.arena_top < st->arena_cap ? c.arena_top : st->arena_cap;
    if (c.table_full || c.arena_full) {
        SW_SET_ERR(ctx, "state table capacity exceeded (table_full=%llu arena_full=%llu)", c.table_full, c.arena_full);
        return SWGPU_E_CAPACITY;
    }
    return SWGPU_OK;
}

// ---- fused host-buffer call of the device-state consumer: decode a poll() batch and fold it into the state,
// without the rows ever leaving the GPU (swgpu_decode_* followed by swgpu_merge would download the decoded
// columns and upload them again).
int sw_decode_pb_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint32_t max_rows,
                        const swgpu_decoded_cols *out, uint8_t *status, uint32_t *first_row, uint32_t *out_rows_dev,
                        uint32_t rec_base, const uint32_t *row_base, uint32_t *row_next);
int sw_decode_json_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint32_t max_rows,
                          const swgpu_decoded_cols *out, uint8_t *status, uint32_t *first_row, uint32_t *out_rows_dev,
                          uint32_t rec_base, const uint32_t *row_base, uint32_t *row_next);

extern "C" int32_t swgpu_decode_merge(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off,
                                      uint32_t n, uint64_t seq_base, const uint64_t *id_hi, const uint64_t *id_lo,
                                      uint8_t *status, uint32_t *out_rows) {
    if (!ctx) return SWGPU_E_INVALID;
    if (!rec_off || !out_rows || (format != SWGPU_FMT_PROTOBUF && format != SWGPU_FMT_JSON) ||
        (n && (!id_hi || !id_lo || (rec_off[n] && !bytes)))) {
        SW_SET_ERR(ctx, "swgpu_decode_merge: bad argument");
        return SWGPU_E_INVALID;
    }
    *out_rows = 0;
    if (n == 0) return SWGPU_OK;
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    const uint64_t total = rec_off[n];
    if (int rc = sw_ensure(ctx, ctx->d_bytes, total + 64)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_off, (size_t)(n + 1) * 8)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_status, (size_t)n + 16)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_first, (size_t)(n + 2) * 4 + 16)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_ids, (size_t)n * 16 + 64)) return rc;
    if (int rc = sw_ensure_pinned(ctx, ctx->h_small, 64)) return rc;
    if (total) SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_bytes.p, bytes, total, cudaMemcpyHostToDevice, s));
    SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_off.p, rec_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, s));
    uint64_t *ids = (uint64_t *)ctx->d_ids.p;
    SW_CUDA(ctx, cudaMemcpyAsync(ids, id_hi, (size_t)n * 8, cudaMemcpyHostToDevice, s));
    SW_CUDA(ctx, cudaMemcpyAsync(ids + n, id_lo, (size_t)n * 8, cudaMemcpyHostToDevice, s));
    uint32_t *d_first = (uint32_t *)ctx->d_first.p, *d_rows = d_first + (n + 1);
    uint32_t cap = n + 64, rows = 0;
    swgpu_decoded_cols d;
    for (int attempt = 0;; attempt++) {  // a DeviceMeasurements payload can fan out: second try with the exact size
        if (int rc = sw_ensure(ctx, ctx->d_cols, (size_t)cap * 8 * 9 + 64)) return rc;
        memset(&d, 0, sizeof d);
        uint64_t **dp = (uint64_t **)&d;
        for (int c = 0; c < 9; c++) dp[c] = (uint64_t *)ctx->d_cols.p + (size_t)c * cap;
        const int rc = format == SWGPU_FMT_JSON
                           ? sw_decode_json_launch(ctx, (const uint8_t *)ctx->d_bytes.p, (const uint64_t *)ctx->d_off.p, n, cap,
                                                   &d, (uint8_t *)ctx->d_status.p, d_first, d_rows, 0, nullptr, nullptr)
                           : sw_decode_pb_launch(ctx, (const uint8_t *)ctx->d_bytes.p, (const uint64_t *)ctx->d_off.p, n, cap,
                                                 &d, (uint8_t *)ctx->d_status.p, d_first, d_rows, 0, nullptr, nullptr);
        if (rc) return rc;
        SW_CUDA(ctx, cudaMemcpyAsync(ctx->h_small.p, d_rows, 4, cudaMemcpyDeviceToHost, s));
        SW_CUDA(ctx, cudaStreamSynchronize(s));
        rows = *(const uint32_t *)ctx->h_small.p;
        if (rows <= cap) break;
        if (attempt) {
            SW_SET_ERR(ctx, "swgpu_decode_merge: row count changed between two decodes of the same batch");
            return SWGPU_E_CUDA;
        }
        cap = rows;
    }
    if (status) SW_CUDA(ctx, cudaMemcpyAsync(status, ctx->d_status.p, n, cudaMemcpyDeviceToHost, s));
    *out_rows = rows;
    if (rows) {
        if (int rc = sw_merge_launch(ctx, (const uint8_t *)ctx->d_bytes.p, (const uint64_t *)ctx->d_off.p, n, &d, rows, nullptr,
                                     seq_base, ids, ids + n, true))
            return rc;
    }
    MgState *st = (MgState *)ctx->state;
    if (!st) {
        SW_CUDA(ctx, cudaStreamSynchronize(s));
        return SWGPU_OK;
    }
    MgCounters c;
    if (int rc2 = mg_read_counters(ctx, st, &c)) return rc2;  // (synchronises: `status` has landed too)
    st->keys_upper = c.n_keys;
    st->arena_upper = c.arena_top < st->arena_cap ? c.arena_top : st->arena_cap;
    if (c.table_full || c.arena_full) {
        SW_SET_ERR(ctx, "state table capacity exceeded (table_full=%llu arena_full=%llu)", c.table_full, c.arena_full);
        return SWGPU_E_CAPACITY;
    }
    return SWGPU_OK;
}

extern "C" int32_t swgpu_state_size(swgpu_ctx *ctx, uint64_t *n_keys, uint64_t *key_bytes) {
    if (!ctx) return SWGPU_E_INVALID;
    MgState *st = (MgState *)ctx->state;
    if (!st) {
        if (n_keys) *n_keys = 0;
        if (key_bytes) *key_bytes = 0;
        return SWGPU_OK;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    MgCounters c;
    if (int rc = mg_read_counters(ctx, st, &c)) return rc;
    if (n_keys) *n_keys = c.n_keys;
    if (key_bytes) *key_bytes = c.arena_top < st->arena_cap ? c.arena_top : st->arena_cap;
    if (c.table_full || c.arena_full) {
        SW_SET_ERR(ctx, "state table capacity exceeded (table_full=%llu arena_full=%llu)", c.table_full, c.arena_full);
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
    st->arena_upper = 0;
    return SWGPU_OK;
}

// Snapshot / restore are checkpoint operations, not the hot path: the table and arena are copied
// to the host and (de)compacted there.
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
    const uint64_t used = c.arena_top < st->arena_cap ? c.arena_top : st->arena_cap;
    if (c.n_keys > max_keys || used > max_key_bytes) {
        *n_keys = c.n_keys;
        SW_SET_ERR(ctx, "snapshot needs %llu keys / %llu key bytes", c.n_keys, (unsigned long long)used);
        return SWGPU_E_CAPACITY;
    }
    std::vector<MgSlot> slots(st->cap);
    std::vector<uint8_t> arena(used ? used : 1);
    SW_CUDA(ctx, cudaMemcpyAsync(slots.data(), st->slots, st->cap * sizeof(MgSlot), cudaMemcpyDeviceToHost, ctx->stream));
    if (used) SW_CUDA(ctx, cudaMemcpyAsync(arena.data(), st->arena, used, cudaMemcpyDeviceToHost, ctx->stream));
    SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    uint64_t k = 0, off = 0;
    for (uint64_t i = 0; i < st->cap; i++) {
        const MgSlot &s = slots[i];
        if (s.tag == 0 || s.key_len == 0xffffffffu) continue;
        out->seq[k] = s.best - 1;
        out->id_hi[k] = s.id_hi;
        out->id_lo[k] = s.id_lo;
        out->event_date[k] = s.date;
        out->value[k] = s.value;
        out->key_off[k] = off;
        // arena words -> public key form  token_len(2) | kind | token | name
        const uint8_t *ak = arena.data() + s.key_off;
        uint64_t hdr;
        memcpy(&hdr, ak, 8);
        const uint32_t tl = (uint32_t)(hdr & 0xffff), nl = (uint32_t)((hdr >> 16) & 0xffff);
        uint8_t *pk = out->key_bytes + off;
        pk[0] = (uint8_t)(tl & 0xff);
        pk[1] = (uint8_t)(tl >> 8);
        pk[2] = (uint8_t)(hdr >> 32);
        memcpy(pk + 3, ak + 8, tl);
        memcpy(pk + 3 + tl, ak + 8 + ((tl + 7u) & ~7u), nl);
        off += 3 + tl + nl;
        k++;
    }
    out->key_off[k] = off;
    *n_keys = k;
    return SWGPU_OK;
}

static uint64_t host_hash64(const uint8_t *d, uint32_t len) {  // sw_hash64 on the host
    uint32_t f = 0x811c9dc5u, i = 0, tail = 0;
    for (; i + 4 <= len; i += 4)
        f = sw_hash_hi_step(f, (uint32_t)d[i] | ((uint32_t)d[i + 1] << 8) | ((uint32_t)d[i + 2] << 16) | ((uint32_t)d[i + 3] << 24));
    for (uint32_t k = 0; i + k < len; k++) tail |= (uint32_t)d[i + k] << (8 * k);
    return ((uint64_t)sw_hash_hi_final(f, tail, len) << 32) | swgpu_murmur2(d, len);
}

extern "C" int32_t swgpu_state_restore(swgpu_ctx *ctx, const swgpu_state_rows *in, uint64_t n_keys) {
    if (!ctx || (n_keys && !in)) return SWGPU_E_INVALID;
    if (int rc = swgpu_state_clear(ctx)) return rc;
    if (n_keys == 0) return SWGPU_OK;
    // arena bytes in the internal word form: at most 8 (header) + 14 (padding) more than the public key
    const uint64_t kbytes = in->key_off[n_keys] + 24 * n_keys;
    if (int rc = mg_reserve(ctx, n_keys, kbytes)) return rc;
    MgState *st = (MgState *)ctx->state;
    std::vector<MgSlot> slots(st->cap);
    memset(slots.data(), 0, slots.size() * sizeof(MgSlot));
    std::vector<uint8_t> arena;
    arena.reserve(kbytes);
    const uint64_t mask = st->cap - 1;
    for (uint64_t k = 0; k < n_keys; k++) {
        const uint8_t *key = in->key_bytes + in->key_off[k];
        const uint32_t kl = (uint32_t)(in->key_off[k + 1] - in->key_off[k]);
        if (kl < 3) {
            SW_SET_ERR(ctx, "swgpu_state_restore: malformed key %llu", (unsigned long long)k);
            return SWGPU_E_INVALID;
        }
        const uint32_t tl = key[0] | ((uint32_t)key[1] << 8);
        const uint32_t kindbyte = key[2];
        if (3 + tl > kl) {
            SW_SET_ERR(ctx, "swgpu_state_restore: malformed key %llu", (unsigned long long)k);
            return SWGPU_E_INVALID;
        }
        const uint32_t nl = kl - 3 - tl;
        const uint32_t kind = kindbyte & 0x7f;
        const uint64_t th = host_hash64(key + 3, tl);
        const uint64_t nh = kind == SWGPU_KIND_LOCATION ? 0 : (((kindbyte & 0x80) && nl == 0) ? 0 : host_hash64(key + 3 + tl, nl));
        const uint64_t h = mg_mix(th, nh, kindbyte);
        uint64_t s = (h >> 2) & mask;
        while (slots[s].tag) s = (s + 1) & mask;
        MgSlot &d = slots[s];
        d.tag = (h & ~3ull) | 1ull;
        d.best = in->seq[k] + 1;
        if (tl > 0xffffu || nl > 0xffffu) {
            SW_SET_ERR(ctx, "swgpu_state_restore: key %llu too long", (unsigned long long)k);
            return SWGPU_E_INVALID;
        }
        d.key_off = arena.size();
        d.key_len = mg_key_bytes(tl, nl);
        arena.resize(arena.size() + d.key_len, 0);
        uint8_t *ak = arena.data() + d.key_off;
        const uint64_t hdr = mg_key_hdr(tl, nl, kindbyte);
        memcpy(ak, &hdr, 8);
        memcpy(ak + 8, key + 3, tl);
        memcpy(ak + 8 + ((tl + 7u) & ~7u), key + 3 + tl, nl);
        d.id_hi = in->id_hi[k];
        d.id_lo = in->id_lo[k];
        d.date = in->event_date[k];
        d.value = in->value[k];
    }
    MgCounters c;
    memset(&c, 0, sizeof c);
    c.n_keys = n_keys;
    c.arena_top = arena.size();
    SW_CUDA(ctx, cudaMemcpyAsync(st->slots, slots.data(), slots.size() * sizeof(MgSlot), cudaMemcpyHostToDevice, ctx->stream));
    SW_CUDA(ctx, cudaMemcpyAsync(st->arena, arena.data(), arena.size(), cudaMemcpyHostToDevice, ctx->stream));
    SW_CUDA(ctx, cudaMemcpyAsync(st->ctr, &c, sizeof c, cudaMemcpyHostToDevice, ctx->stream));
    SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    st->keys_upper = n_keys;
    st->arena_upper = kbytes;
    return SWGPU_OK;
}
