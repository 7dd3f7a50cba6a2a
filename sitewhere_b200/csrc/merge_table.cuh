// merge_table.cuh — the device-state table and the per-row upsert shared by the merge kernel (merge.cu, rows
// read back from the decoded columns) and the fused decode+merge kernels (decode_pb.cu / decode_json.cu, rows
// still in the decoder's registers and shared memory).
//
// Reference: the fold DeviceStateProcessingLogic.processDeviceStateEvent + merge*
// (service-device-state/.../processing/DeviceStateProcessingLogic.java:116-201) with
// DeviceStatePersistence.deviceStateUpdateLogic (.../persistence/DeviceStatePersistence.java:60-85): a sequential
// "later event overwrites" per key, i.e. per key an argmax over the arrival sequence number.
//
// Table: open addressing, linear probing, 128-byte slots laid out by 32-byte DRAM sector so that a row
// touches as few random sectors as the data allows:
//   sector 0   w0 tag | w1 best (seq + 1 of the winner) | key words 0-1
//   sector 1   key words 2-5
//   sector 2   winner's payload: id_hi, id_lo, event date, value     (written as one full sector: no fill read)
//   sector 3   key words 6-9 (only keys longer than 40 bytes of token + name reach it)
// A key is  word 0 = header (token length | name length << 16 | kind byte << 32 | bit 63)  then the token and
// the name, each zero-padded to 8-byte words: equality is a handful of 64-bit compares against registers.
// The common key (21-byte token, name of up to 16 bytes) lives in sectors 0-1, so an update is ONE random
// 64-byte read, one atomicMax on a line that is now in L2, and one 32-byte store; keys of more than ten words
// keep their words in an arena and the slot holds the offset.  Hashes only pick the bucket; bytes decide.
#pragma once
#include "swgpu_common.cuh"

#define MG_KEY_INLINE_WORDS 10u  // header + 9 words (72 bytes) of token and name
#define MG_PENDING 2ull          // tag bit: the claimer has not published the key words yet
#define MG_TOMB 4ull             // tag of a slot whose claimer could not store its (long) key: equal to no key
#define MG_MAX_PROBES 8192ull
#define MG_HDR_SET (1ull << 63)  // a header is never zero
#define MG_HDR_LONG (1ull << 62)  // key words are in the arena: word 1 = byte offset, word 2 = word count

struct __align__(128) MgSlot {
    uint64_t w[16];
};
#define MG_W_TAG 0
#define MG_W_BEST 1
#define MG_W_KEY0 2   // key words 0..5 -> w[2..7]
#define MG_W_PAY 8    // id_hi, id_lo, date, value -> w[8..11]
#define MG_W_KEY6 12  // key words 6..9 -> w[12..15]
static_assert(sizeof(MgSlot) == 128, "MgSlot layout");

struct MgCounters {  // device
    unsigned long long n_keys, key_bytes, arena_top, table_full;
    unsigned long long arena_full;  // 8-byte words of long keys that did not fit the arena
    unsigned long long key_too_long, n_fix, fix_overflow;
};
struct MgFix {  // a row that must write its payload once more after the launch (see mg_upsert)
    uint64_t slot, seq1, id_hi, id_lo, date, value;
};
struct MgTable {  // what a kernel needs
    MgSlot *slots;
    uint64_t mask;
    uint64_t *arena;  // 8-byte words
    uint64_t arena_cap;  // in words
    MgCounters *ctr;
    MgFix *fix;
    uint64_t fix_cap;
    uint64_t seq_floor;  // every sequence number of this launch is > seq_floor (stored values: seq + 1 > seq_floor)
    uint64_t presence_now1;  // 0 = no presence tracking; else the batch's timestamp (ms) + 1
};
#define MG_KIND_DEVICE 0x7fu  // kind byte of a device's presence row: best - 1 = lastInteractionDate, date word = presenceMissingDate

__host__ __device__ __forceinline__ uint64_t mg_mix(uint64_t th, uint64_t nh, uint32_t kindbyte) {
    uint64_t h = th ^ (nh * 0x9e3779b97f4a7c15ull) ^ ((uint64_t)kindbyte << 56);
    h ^= h >> 33;
    h *= 0xff51afd7ed558ccdull;
    h ^= h >> 33;
    h *= 0xc4ceb9fe1a85ec53ull;
    h ^= h >> 33;
    return h;
}
__host__ __device__ __forceinline__ uint64_t mg_key_hdr(uint32_t tl, uint32_t nl, uint32_t kindbyte) {
    return (uint64_t)tl | ((uint64_t)nl << 16) | ((uint64_t)kindbyte << 32) | MG_HDR_SET;
}
__host__ __device__ __forceinline__ uint32_t mg_key_words(uint32_t tl, uint32_t nl) {
    return 1u + ((tl + 7u) >> 3) + ((nl + 7u) >> 3);
}
// the state key's kind byte: kind, with bit 7 for a measurement / alert whose name is null
__host__ __device__ __forceinline__ uint32_t mg_kindbyte(uint32_t kind, uint32_t flags) {
    return kind | ((kind != SWGPU_KIND_LOCATION && !(flags & SWGPU_F_HAS_NAME)) ? 0x80u : 0u);
}

#ifdef __CUDACC__
__device__ __forceinline__ uint64_t mg_ld_acquire_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mg_st_release_u64(uint64_t *p, uint64_t v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ ulonglong2 mg_ldcg2(const uint64_t *p) { return __ldcg((const ulonglong2 *)p); }
// the winner's payload as ONE 256-bit store: a whole 32-byte sector, so L2 need not fetch it from DRAM first
// (two 16-byte stores each look like a partial write and cost a fill read: measured, tools/exp/randacc.cu)
__device__ __forceinline__ void mg_store_payload(uint64_t *w, uint64_t id_hi, uint64_t id_lo, uint64_t date, uint64_t value) {
    asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(w + MG_W_PAY), "l"(id_hi), "l"(id_lo), "l"(date), "l"(value)
                 : "memory");
}

// Key readers give word i of the row's key (i >= 1; word 0 is the header, held by the caller): the token's
// ceil(tl/8) words, then the name's.  `B` is one of the decoders' byte readers (shared / global memory).
template <typename B>
struct MgKeyBytes {
    B b;
    uint32_t tok_off, tl, name_off, nl;
    __device__ __forceinline__ uint64_t piece(uint32_t off, uint32_t valid) const {
        uint64_t v = 0;
        if (valid >= 8) return b.le64(off);
        for (uint32_t k = 0; k < valid; k++) v |= (uint64_t)b.byte(off + k) << (8 * k);
        return v;
    }
    __device__ __forceinline__ uint64_t word(uint32_t i) const {
        const uint32_t tw = (tl + 7u) >> 3;
        if (i <= tw) return piece(tok_off + 8 * (i - 1), min(8u, tl - 8 * (i - 1)));
        const uint32_t j = i - 1 - tw;
        return piece(name_off + 8 * j, min(8u, nl - 8 * j));
    }
};

// Upsert of one row: find or claim the key's slot, raise its `best`, write the payload if this row won.
//
// Last writer wins: `best` = seq + 1 of the newest event seen for the key; the row that raises it owns the
// slot's payload at that moment and writes it here, in the same pass.  Two rows of one launch can both raise it
// (an older row got there first, then a newer one): their payload stores may land in either order, so the newer
// row — which sees a previous value from this same launch — also puts itself on a fix-up list, and a second,
// tiny kernel (merge_fixup_kernel) lets the final winner of each such key write once more after every store of
// this launch has landed.  A row whose sequence number EQUALS the stored one rewrites the payload: that is a
// replay of the same batch (idempotent) or a later row of the same record beyond the per-record sequence range,
// handled by the same thread in program order.
//
// Call with any set of active lanes.  `local` = this warp's counters in shared memory (mg_flush_counts).
struct MgLocalCnt {
    uint32_t keys, bytes;
};
__device__ __forceinline__ void mg_flush_counts(const MgTable &t, MgLocalCnt *local) {  // one lane, after __syncwarp
    if (local->keys) {
        atomicAdd(&t.ctr->n_keys, (unsigned long long)local->keys);
        atomicAdd(&t.ctr->key_bytes, (unsigned long long)local->bytes);
        local->keys = local->bytes = 0;
    }
}
// One upsert.  DEVICE = true: the device's presence row (see mg_upsert): every writer stores the same values, so no
// fix-up, and a row that finds the batch's timestamp already there has nothing to do.
// Returns the key's `best` as it was before this row (0 for a new key), MG_NOT_STORED when the row found no room.
#define MG_NOT_STORED (~0ull)
template <bool DEVICE, typename KeyRd>
__device__ __forceinline__ unsigned long long mg_upsert_one(const MgTable &t, MgLocalCnt *local, uint64_t h, uint32_t tl,
                                                            uint32_t nl, uint32_t kindbyte, const KeyRd &key, uint64_t seq1,
                                                            uint64_t id_hi, uint64_t id_lo, uint64_t date, uint64_t value) {
    const uint32_t nw = mg_key_words(tl, nl);
    const bool inl = nw <= MG_KEY_INLINE_WORDS;
    const uint64_t hdr = mg_key_hdr(tl, nl, kindbyte) | (inl ? 0ull : MG_HDR_LONG);
    // key words 0..5 in registers (requested before the table is probed: their loads and the slot's random,
    // DRAM-latency load are in flight together); words 6..9 are read again on demand
    uint64_t kw[6];
    kw[0] = hdr;
    bool haszero = false;  // a zero word cannot tell "not written yet" from "written": such keys take the ordered path
#pragma unroll
    for (uint32_t i = 1; i < 6; i++) {
        kw[i] = (inl && i < nw) ? key.word(i) : 0ull;
        haszero |= (inl && i < nw && kw[i] == 0ull);
    }
    if (inl)
        for (uint32_t i = 6; i < nw; i++) haszero |= key.word(i) == 0ull;
    const uint64_t want = (h & ~7ull) | 1ull;
    uint64_t s = (h >> 3) & t.mask;
    bool found = false, inserted = false, dropped = false;
    // (the table is kept at most half full, so probe sequences are short; the bound only keeps a table that a
    //  device-resident caller overfilled from spinning: such rows are counted in table_full)
    const uint64_t max_probes = t.mask < MG_MAX_PROBES ? t.mask + 1 : MG_MAX_PROBES;
    for (uint64_t probes = 0; probes < max_probes; probes++, s = (s + 1) & t.mask) {
        uint64_t *w = t.slots[s].w;
        // optimistic read of sectors 0-1: tag, best and key words 0..5 in one round trip
        const ulonglong2 a0 = mg_ldcg2(w + 0), a1 = mg_ldcg2(w + 2), a2 = mg_ldcg2(w + 4), a3 = mg_ldcg2(w + 6);
        uint64_t tag = a0.x;
        if (tag == 0) {
            const unsigned long long old = atomicCAS((unsigned long long *)w, 0ull, (unsigned long long)(want | MG_PENDING));
            if (old == 0) {  // claimed: publish the key words, then clear the pending bit
                bool ok = true;
                if (inl) {
                    *(ulonglong2 *)(w + 2) = make_ulonglong2(kw[0], kw[1]);
                    *(ulonglong2 *)(w + 4) = make_ulonglong2(kw[2], kw[3]);
                    *(ulonglong2 *)(w + 6) = make_ulonglong2(kw[4], kw[5]);
                    for (uint32_t i = 6; i < nw; i++) w[MG_W_KEY6 + i - 6] = key.word(i);
                } else {
                    const unsigned long long off = atomicAdd(&t.ctr->arena_top, (unsigned long long)(nw - 1));
                    if (off + (nw - 1) > t.arena_cap) {
                        // no room for the key words: the reservation is handed back (an allocation that succeeds
                        // lies wholly below the end of the arena, one that fails reaches beyond it, so they never
                        // overlap), the words asked for are counted so that the host can grow the arena and replay
                        // the batch, and the slot becomes a tombstone (it cannot be freed: other keys may already
                        // have probed past it)
                        atomicAdd(&t.ctr->arena_top, (unsigned long long)(0ull - (unsigned long long)(nw - 1)));
                        atomicAdd(&t.ctr->arena_full, (unsigned long long)(nw - 1));
                        ok = false;
                    } else {
                        for (uint32_t i = 1; i < nw; i++) t.arena[off + i - 1] = key.word(i);
                        *(ulonglong2 *)(w + 2) = make_ulonglong2(hdr, (uint64_t)off);
                        w[4] = nw;
                    }
                }
                __threadfence();
                mg_st_release_u64(w, ok ? want : MG_TOMB);
                if (!ok) {
                    dropped = true;
                    break;
                }
                found = inserted = true;
                break;
            }
            tag = old;  // somebody else claimed it first: is it our key?
        }
        if ((tag & ~MG_PENDING) != want) continue;  // another key (or a tombstone)
        bool eq = false, settled = false;
        if (tag == want && a0.x == want && inl && !haszero) {
            // every word we hold is non-zero and a slot's key words only ever go from zero to their final
            // value: if all of them compare equal, they are the published key, whatever the order in which
            // the loads above were served
            eq = a1.x == kw[0] && a1.y == kw[1] && a2.x == kw[2] && a2.y == kw[3] && a3.x == kw[4] && a3.y == kw[5];
            for (uint32_t i = 6; eq && i < nw; i++) eq = __ldcg(w + MG_W_KEY6 + i - 6) == key.word(i);
            settled = eq;  // a mismatch may be a stale read: look again below
        }
        if (!settled) {
            // ordered re-read: wait for the claimer, then compare (acquire: the key words are visible)
            tag = mg_ld_acquire_u64(w);
            while (tag & MG_PENDING) tag = mg_ld_acquire_u64(w);
            eq = tag == want && __ldcg(w + MG_W_KEY0) == hdr;
            if (eq && inl) {
                for (uint32_t i = 1; eq && i < nw; i++)
                    eq = __ldcg(w + (i < 6 ? MG_W_KEY0 + i : MG_W_KEY6 + i - 6)) == key.word(i);
            } else if (eq) {
                const uint64_t off = __ldcg(w + MG_W_KEY0 + 1);
                eq = __ldcg(w + MG_W_KEY0 + 2) == nw;
                for (uint32_t i = 1; eq && i < nw; i++) eq = __ldcg(t.arena + off + i - 1) == key.word(i);
            }
        }
        if (eq) {
            found = true;
            break;
        }
    }
    if (inserted) {
        // key counters: summed per warp in shared memory, flushed once per warp by mg_flush_counts
        atomicAdd(&local->keys, 1u);
        atomicAdd(&local->bytes, 3u + tl + nl);
    }
    if (!found) {
        if (!dropped) atomicAdd(&t.ctr->table_full, 1ull);  // every slot probed
        return MG_NOT_STORED;
    }
    uint64_t *w = t.slots[s].w;
    const unsigned long long old = atomicMax((unsigned long long *)(w + MG_W_BEST), (unsigned long long)seq1);
    if (old > seq1) return old;                // a newer event is already there
    if (DEVICE && old == seq1) return old;     // the device row was touched by another row of this batch already
    mg_store_payload(w, id_hi, id_lo, date, value);
    if (!DEVICE && old > t.seq_floor && old != seq1) {  // the previous value was set by an older row of this launch
        const unsigned long long k = atomicAdd(&t.ctr->n_fix, 1ull);
        if (k < t.fix_cap) {
            MgFix f;
            f.slot = s;
            f.seq1 = seq1;
            f.id_hi = id_hi;
            f.id_lo = id_lo;
            f.date = date;
            f.value = value;
            t.fix[k] = f;
        } else
            atomicAdd(&t.ctr->fix_overflow, 1ull);
    }
    return old;
}
// the device row of a merged event, out of line: the hot path's registers are not shared with it
template <typename KeyRd>
__device__ __noinline__ void mg_touch_device(const MgTable &t, MgLocalCnt *local, uint64_t th, uint32_t tl, const KeyRd &key) {
    mg_upsert_one<true>(t, local, mg_mix(th, 0ull, MG_KIND_DEVICE), tl, 0u, MG_KIND_DEVICE, key, t.presence_now1, 0ull, 0ull, 0ull,
                        0ull);
}
// When the table tracks presence (t.presence_now1 != 0) every row also touches its device's row — key (token,
// MG_KIND_DEVICE, no name): best := the batch's timestamp + 1 (lastInteractionDate), payload cleared
// (presenceMissingDate := null), as DeviceStateProcessingLogic.processDeviceStateEvent does for every merged event
// (DeviceStateProcessingLogic.java:142-143).
template <typename KeyRd>
__device__ __forceinline__ void mg_upsert(const MgTable &t, MgLocalCnt *local, uint64_t h, uint64_t th, uint32_t tl, uint32_t nl,
                                          uint32_t kindbyte, const KeyRd &key, uint64_t seq1, uint64_t id_hi, uint64_t id_lo,
                                          uint64_t date, uint64_t value) {
    mg_upsert_one<false>(t, local, h, tl, nl, kindbyte, key, seq1, id_hi, id_lo, date, value);
    if (t.presence_now1) mg_touch_device(t, local, th, tl, key);
}
#endif  // __CUDACC__
