// ctx.cu — context lifetime, buffers, pinned staging, host-side Kafka partition helper.
#include <utility>

#include "swgpu_common.cuh"

static thread_local std::string g_create_err;

int sw_ensure(swgpu_ctx *ctx, swgpu_ctx::Buf &b, size_t bytes) {
    if (bytes <= b.cap) return SWGPU_OK;
    size_t want = bytes + bytes / 4 + 256;
    if (b.p) {
        SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        SW_CUDA(ctx, cudaFree(b.p));
        b.p = nullptr;
        b.cap = 0;
    }
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess) {
        SW_SET_ERR(ctx, "cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
        b.p = nullptr;
        return SWGPU_E_NOMEM;
    }
    b.cap = want;
    return SWGPU_OK;
}

int sw_ensure_pinned(swgpu_ctx *ctx, swgpu_ctx::Buf &b, size_t bytes) {
    if (bytes <= b.cap) return SWGPU_OK;
    size_t want = bytes + bytes / 4 + 256;
    if (b.p) {
        SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        SW_CUDA(ctx, cudaFreeHost(b.p));
        b.p = nullptr;
        b.cap = 0;
    }
    cudaError_t e = cudaMallocHost(&b.p, want);
    if (e != cudaSuccess) {
        SW_SET_ERR(ctx, "cudaMallocHost(%zu) failed: %s", want, cudaGetErrorString(e));
        b.p = nullptr;
        return SWGPU_E_NOMEM;
    }
    b.cap = want;
    return SWGPU_OK;
}

int sw_ensure_scratch(swgpu_ctx *ctx, size_t bytes) {
    if (bytes <= ctx->scratch_bytes) return SWGPU_OK;
    size_t want = bytes + bytes / 2 + 4096;
    if (ctx->scratch) {
        SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        SW_CUDA(ctx, cudaFree(ctx->scratch));
        ctx->scratch = nullptr;
        ctx->scratch_bytes = 0;
    }
    cudaError_t e = cudaMalloc(&ctx->scratch, want);
    if (e != cudaSuccess) {
        SW_SET_ERR(ctx, "cudaMalloc(scratch %zu) failed: %s", want, cudaGetErrorString(e));
        return SWGPU_E_NOMEM;
    }
    ctx->scratch_bytes = want;
    return SWGPU_OK;
}

extern "C" int32_t swgpu_abi_version(void) { return SWGPU_ABI_VERSION; }

extern "C" int32_t swgpu_create(int32_t device_ordinal, void *stream, swgpu_ctx **out_ctx) {
    if (!out_ctx) return SWGPU_E_INVALID;
    *out_ctx = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        g_create_err = std::string("no CUDA device: ") + cudaGetErrorString(e) +
                       " (libswgpu has no CPU path)";
        return SWGPU_E_CUDA;
    }
    if (device_ordinal < 0 || device_ordinal >= ndev) {
        g_create_err = "device ordinal out of range";
        return SWGPU_E_INVALID;
    }
    e = cudaSetDevice(device_ordinal);
    if (e != cudaSuccess) {
        g_create_err = std::string("cudaSetDevice: ") + cudaGetErrorString(e);
        return SWGPU_E_CUDA;
    }
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device_ordinal);
    if (e != cudaSuccess) {
        g_create_err = std::string("cudaGetDeviceProperties: ") + cudaGetErrorString(e);
        return SWGPU_E_CUDA;
    }
    if (prop.major != 10) {
        g_create_err = "libswgpu is built for sm_100a (B200) only; device is sm_" +
                       std::to_string(prop.major) + std::to_string(prop.minor);
        return SWGPU_E_CUDA;
    }
    swgpu_ctx *ctx = new swgpu_ctx();
    ctx->device = device_ordinal;
    ctx->sm_count = prop.multiProcessorCount > 0 ? prop.multiProcessorCount : 148;
    if (stream) {
        ctx->stream = (cudaStream_t)stream;
    } else {
        e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) {
            g_create_err = std::string("cudaStreamCreate: ") + cudaGetErrorString(e);
            delete ctx;
            return SWGPU_E_CUDA;
        }
        ctx->own_stream = true;
    }
    *out_ctx = ctx;
    return SWGPU_OK;
}

void sw_zone_free(swgpu_ctx *ctx);   // zone.cu
void sw_state_free(swgpu_ctx *ctx);  // merge.cu
void sw_pipeline_free(swgpu_ctx *ctx);  // pipeline.cu

extern "C" void swgpu_destroy(swgpu_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    swgpu_ctx::Buf *bufs[] = {&ctx->d_bytes, &ctx->d_off, &ctx->d_status, &ctx->d_first, &ctx->d_cols,
                              &ctx->d_pairs, &ctx->d_lon, &ctx->d_lat, &ctx->d_meta, &ctx->d_ids, &ctx->d_zwork};
    for (auto *b : bufs)
        if (b->p) cudaFree(b->p);
    if (ctx->h_small.p) cudaFreeHost(ctx->h_small.p);
    if (ctx->scratch) cudaFree(ctx->scratch);
    sw_pipeline_free(ctx);
    sw_zone_free(ctx);
    sw_state_free(ctx);
    std::swap(ctx->state, ctx->dedup_state);  // the alternate-id set is a table of the same kind
    sw_state_free(ctx);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" const char *swgpu_last_error(swgpu_ctx *ctx) {
    if (!ctx) return g_create_err.c_str();
    return ctx->err.c_str();
}

extern "C" void *swgpu_stream(swgpu_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

extern "C" int32_t swgpu_sync(swgpu_ctx *ctx) {
    if (!ctx) return SWGPU_E_INVALID;
    SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SWGPU_OK;
}

// Named integer options.  None of them changes a result; they select between equivalent code paths so that
// the tests can cover each one, or switch on work counters (see include/swgpu.h).
extern "C" int32_t swgpu_set_option(swgpu_ctx *ctx, const char *name, int64_t value) {
    if (!ctx || !name) return SWGPU_E_INVALID;
    if (!strcmp(name, "zone.force_ray")) {
        if (value < 0 || value > 4) {
            SW_SET_ERR(ctx, "swgpu_set_option: zone.force_ray takes 0 (choose) or 1..4 (+x, -x, +y, -y)");
            return SWGPU_E_INVALID;
        }
        ctx->opt_zone_force_ray = (uint32_t)value;
        return SWGPU_OK;
    }
    if (!strcmp(name, "zone.scan_sequential")) {
        ctx->opt_zone_scan_sequential = value ? 1u : 0u;
        return SWGPU_OK;
    }
    if (!strcmp(name, "state.presence_now_ms")) {
        if (value < 0) {
            SW_SET_ERR(ctx, "swgpu_set_option: state.presence_now_ms takes a time in ms since the epoch, or 0 to switch presence off");
            return SWGPU_E_INVALID;
        }
        ctx->opt_presence_now1 = value ? (uint64_t)value + 1 : 0;
        return SWGPU_OK;
    }
    if (!strcmp(name, "json.fast_path")) {
        ctx->opt_json_fast = value ? 1u : 0u;
        return SWGPU_OK;
    }
    if (!strcmp(name, "enriched.fast_path")) {
        ctx->opt_enriched_fast = value ? 1u : 0u;
        return SWGPU_OK;
    }
    if (!strcmp(name, "json.stats")) {
        ctx->opt_json_stats = value ? 1u : 0u;
        return SWGPU_OK;
    }
    if (!strcmp(name, "zone.stats")) {
        ctx->opt_zone_stats = value ? 1u : 0u;
        return SWGPU_OK;
    }
    SW_SET_ERR(ctx, "swgpu_set_option: unknown option '%s'", name);
    return SWGPU_E_INVALID;
}

extern "C" void *swgpu_alloc_pinned(size_t bytes) {
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) return nullptr;
    return p;
}
extern "C" void swgpu_free_pinned(void *p) {
    if (p) cudaFreeHost(p);
}

// Kafka Utils.murmur2 / DefaultPartitioner (kafka-clients 0.11.0.0), host side: used by the binding
// to route a token to its GPU shard exactly as Kafka routes it to a partition.
extern "C" uint32_t swgpu_murmur2(const uint8_t *data, uint32_t len) {
    const uint32_t m = 0x5bd1e995u;
    uint32_t h = 0x9747b28cu ^ len;
    uint32_t i = 0;
    for (; i + 4 <= len; i += 4) {
        uint32_t k;
        memcpy(&k, data + i, 4);  // little-endian hosts only (x86-64 / aarch64)
        k *= m;
        k ^= k >> 24;
        k *= m;
        h *= m;
        h ^= k;
    }
    uint32_t rem = len - i;
    if (rem) {
        uint32_t t = data[i];
        if (rem > 1) t |= (uint32_t)data[i + 1] << 8;
        if (rem > 2) t |= (uint32_t)data[i + 2] << 16;
        h ^= t;
        h *= m;
    }
    h ^= h >> 13;
    h *= m;
    h ^= h >> 15;
    return h;
}
extern "C" uint32_t swgpu_partition(const uint8_t *key, uint32_t len, uint32_t n_partitions) {
    if (n_partitions == 0) return 0;
    return (swgpu_murmur2(key, len) & 0x7fffffffu) % n_partitions;
}
