// pb_reader.cuh — protobuf wire-format reader shared by the device-payload decoder (decode_pb.cu: proto2,
// protobuf-java 3.3.0 lite, parseDelimitedFrom on a stream) and the enriched-event decoder (decode_enriched.cu:
// proto3, full runtime, parseFrom on a byte array).  The byte source `B` is one of swgpu_common.cuh's readers
// (a record staged in shared memory, or global memory).  The rules restated here are protobuf-java 3.3.0's
// (third party, recalled): readTag returns 0 at the end of input and rejects field number 0; varints are at most
// 10 bytes and 32-bit reads keep the low 32 bits; an unknown field is skipped by wire type, groups recursively
// with the end tag checked, END_GROUP stops the message, wire types 6 and 7 are invalid; readMessage reads a
// length (negative: error), pushes it as a limit (beyond the enclosing limit: error) and checks that the nested
// message did not stop on an END_GROUP tag.
#pragma once
#include "swgpu_common.cuh"

// Nesting of unknown-field groups followed before the record is declared malformed.  (protobuf-java recurses
// until the JVM stack ends; the test oracle uses the same bound.)
#define PB_MAX_GROUP_DEPTH 64

template <typename B>
struct PbReader {
    B b;
    uint32_t pos, end, limit;
    uint32_t last_tag;
    bool err;

    __device__ __forceinline__ bool at_end() const { return pos >= end || pos >= limit; }
    __device__ __forceinline__ uint32_t take() {
        if (at_end()) {
            err = true;  // truncatedMessage
            return 0;
        }
        return b.byte(pos++);
    }
    __device__ __forceinline__ uint64_t varint64() {
        uint64_t v = 0;
#pragma unroll 1
        for (int shift = 0; shift < 64; shift += 7) {
            uint32_t c = take();
            if (err) return 0;
            v |= (uint64_t)(c & 0x7f) << shift;
            if (!(c & 0x80)) return v;
        }
        err = true;  // malformedVarint
        return 0;
    }
    __device__ __forceinline__ uint32_t varint32() { return (uint32_t)varint64(); }
    __device__ __forceinline__ void skip(uint32_t n) {
        uint32_t lim = end < limit ? end : limit;
        if (n > lim - pos) {
            pos = lim;
            err = true;
            return;
        }
        pos += n;
    }
    __device__ __forceinline__ uint64_t fixed64() {
        uint32_t lim = end < limit ? end : limit;
        if (lim - pos < 8) {
            pos = lim;
            err = true;
            return 0;
        }
        uint64_t v = b.le64(pos);
        pos += 8;
        return v;
    }
    // readString / readBytes: span of the payload
    __device__ __forceinline__ void lenfield(uint32_t &off, uint32_t &len) {
        int32_t size = (int32_t)varint32();
        off = 0;
        len = 0;
        if (err) return;
        if (size < 0) {
            err = true;  // negativeSize
            return;
        }
        uint32_t start = pos;
        skip((uint32_t)size);
        if (err) return;
        off = start;
        len = (uint32_t)size;
    }
    __device__ __forceinline__ uint32_t read_tag() {
        if (at_end()) {
            last_tag = 0;
            return 0;
        }
        uint32_t t = varint32();
        if (err) return 0;
        last_tag = t;
        if ((t >> 3) == 0) {
            err = true;  // invalidTag
            return 0;
        }
        return t;
    }
    __device__ __forceinline__ void skip_scalar(uint32_t wt) {
        uint32_t o, l;
        switch (wt) {
        case 0: varint64(); break;
        case 1: skip(8); break;
        case 2: lenfield(o, l); break;
        case 5: skip(4); break;
        default: err = true;  // invalidWireType (6, 7)
        }
    }
    // generated-code default branch (GeneratedMessageLite.parseUnknownField +
    // UnknownFieldSetLite.mergeFieldFrom); false = END_GROUP, the message stops here.
    __device__ bool unknown(uint32_t tag) {
        uint32_t wt = tag & 7;
        if (wt == 4) return false;
        if (wt != 3) {
            skip_scalar(wt);
            return true;
        }
        uint32_t stack[PB_MAX_GROUP_DEPTH];
        int depth = 1;
        stack[0] = tag >> 3;
        while (depth > 0 && !err) {
            uint32_t t = read_tag();
            if (err) break;
            if (t == 0) {  // input ended inside the group: checkLastTagWas(END_GROUP) fails
                err = true;
                break;
            }
            uint32_t w = t & 7;
            if (w == 4) {
                if (t != ((stack[depth - 1] << 3) | 4)) err = true;
                depth--;
            } else if (w == 3) {
                if (depth >= PB_MAX_GROUP_DEPTH)
                    err = true;
                else
                    stack[depth++] = t >> 3;
            } else {
                skip_scalar(w);
            }
        }
        return true;
    }
    // readMessage prologue: length + pushLimit.  Returns the old limit.
    __device__ __forceinline__ uint32_t push_message() {
        int32_t length = (int32_t)varint32();
        uint32_t old = limit;
        if (err) return old;
        if (length < 0) {
            err = true;
            return old;
        }
        uint64_t nl = (uint64_t)pos + (uint32_t)length;
        if (nl > old) {
            err = true;
            return old;
        }
        limit = (uint32_t)nl;
        return old;
    }
    // Model.Metadata { required string name = 1; required string value = 2; }
    __device__ void metadata_entry() {
        uint32_t old = push_message();
        if (err) return;
        uint32_t has = 0, o, l;
        while (!err) {
            uint32_t t = read_tag();
            if (err || t == 0) break;
            if (t == 10) {
                lenfield(o, l);
                has |= 1;
            } else if (t == 18) {
                lenfield(o, l);
                has |= 2;
            } else if (!unknown(t))
                break;
        }
        if (!err && (last_tag != 0 || has != 3)) err = true;
        limit = old;
    }
    // Model.Measurement { required string measurementId = 1; required double measurementValue = 2; }
    __device__ void measurement(uint32_t &noff, uint32_t &nlen, uint64_t &vbits) {
        uint32_t old = push_message();
        if (err) return;
        uint32_t has = 0;
        while (!err) {
            uint32_t t = read_tag();
            if (err || t == 0) break;
            if (t == 10) {
                lenfield(noff, nlen);
                has |= 1;
            } else if (t == 17) {
                vbits = fixed64();
                has |= 2;
            } else if (!unknown(t))
                break;
        }
        if (!err && (last_tag != 0 || has != 3)) err = true;
        limit = old;
    }
    // AbstractParser.parseDelimitedFrom over the record stream: false = null (stream exhausted)
    __device__ __forceinline__ bool open_delimited(uint32_t &stream_pos, uint32_t total) {
        if (stream_pos >= total) return false;
        pos = stream_pos;
        end = total;
        limit = 0xffffffffu;
        err = false;
        last_tag = 0;
        int32_t size = (int32_t)varint32();  // readRawVarint32(firstByte, input)
        if (err) {
            stream_pos = total;
            return true;
        }
        uint32_t avail = total - pos;
        uint32_t takeb = size < 0 ? 0u : ((uint32_t)size < avail ? (uint32_t)size : avail);
        end = pos + takeb;
        uint64_t lim = (uint64_t)pos + 0x7fffffffull;  // currentLimit = Integer.MAX_VALUE
        limit = lim > 0xffffffffull ? 0xffffffffu : (uint32_t)lim;
        stream_pos = end;
        return true;
    }
};

