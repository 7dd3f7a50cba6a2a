// pipeline.cu — the fused host-buffer path: decode -> zone test (-> state merge) over a micro-batch that
// lives in HOST memory, as the Java binding sees it.  The batch is cut into chunks of records; a ring of
// three device slots lets the host->device copy of chunk i+1, the kernels of chunk i and the
// device->host copy of chunk i-1 run at the same time on three streams (PCIe is full duplex), so the
// call runs at the speed of the slower copy direction instead of the sum of copy + compute.
//
// Replaces, for a batch: IDeviceEventDecoder.decode per payload (ES/spi/IDeviceEventDecoder.java:39) followed by
// ZoneTestRuleProcessor.onLocation per decoded location (RP/processors/geospatial/ZoneTestRuleProcessor.java:47-63)
// and, when asked, DeviceStateProcessingLogic.process (DS/processing/DeviceStateProcessingLogic.java:65-67).
#include <condition_variable>
#include <deque>
#include <functional>
#include <map>
#include <thread>

#include "swgpu_common.cuh"

int sw_decode_pb_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint32_t max_rows,
                        const swgpu_decoded_cols *out, uint8_t *status, uint32_t *first_row, uint32_t *out_rows_dev,
                        uint32_t rec_base, const uint32_t *row_base, uint32_t *row_next, bool small);
int sw_decode_json_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint32_t max_rows,
                          const swgpu_decoded_cols *out, uint8_t *status, uint32_t *first_row, uint32_t *out_rows_dev,
                          uint32_t rec_base, const uint32_t *row_base, uint32_t *row_next, bool small);
int sw_decode_enriched_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint32_t max_rows,
                              const swgpu_decoded_cols *out, const swgpu_event_ids *ids, uint8_t *status, uint32_t *first_row,
                              uint32_t *out_rows_dev, uint32_t rec_base, const uint32_t *row_base, uint32_t *row_next,
                              bool small);
int sw_zone_launch(swgpu_ctx *ctx, const double *lon, const double *lat, const uint64_t *meta, uint32_t n,
                   const uint32_t *n_rows_dev, uint64_t max_pairs, uint64_t *pairs, uint64_t *n_pairs_dev,
                   const uint32_t *row_base);
int sw_zone_single_launch(swgpu_ctx *ctx, const double *lon, const double *lat, const uint64_t *meta, uint32_t n,
                          const uint32_t *n_rows_dev, uint64_t max_pairs, uint64_t *pairs, uint64_t *n_pairs_dev,
                          void *scratch);

#define PL_SMALL_MAX 16384u  // records: up to here a batch takes the few-launch path below
#define PL_SMALL_MAX_ROWS 65536u

// decode + zone test of one batch or chunk on the context's stream: decode, then test the decoded columns.  (Testing
// a location inside the protobuf decode kernel, while its row is in registers, was built and measured: 0.71 ms per
// 10 M locations against 0.57 for the two kernels — both are latency-bound and the fused kernel has a third of the
// zone kernel's resident warps)
// A small batch is bound by launches, not by bytes (a 64 KB poll: eight stream operations of 3-5 us each): it takes
// THREE — one clear of a scratch that serves both kernels, the decoder's ordered pass alone (exact placement by
// look-back; with a few dozen tiles nobody waits long), and the single-launch zone kernel.
static int pl_decode_zone(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                          uint32_t rows_cap, const swgpu_decoded_cols *d, uint8_t *status, uint32_t *first_row,
                          uint32_t *d_rows, uint32_t rec_base, const uint32_t *row_base, uint32_t *row_next,
                          uint64_t pairs_cap, uint64_t *pairs, uint64_t *d_pairs_n) {
    const bool small = n != 0 && n <= PL_SMALL_MAX && rows_cap <= PL_SMALL_MAX_ROWS && !rec_base && !row_base && !row_next &&
                       !ctx->opt_zone_stats;
    size_t dec_bytes = 0;
    if (small) {
        dec_bytes = (SW_SCRATCH_HDR + ((size_t)n / 128 + 2) * 8 + 255) / 256 * 256;
        const size_t zone_bytes = SW_SCRATCH_HDR + ((size_t)rows_cap / 256 + 2) * 8;
        if (int rc = sw_ensure_scratch(ctx, dec_bytes + zone_bytes)) return rc;
        SW_CUDA(ctx, cudaMemsetAsync(ctx->scratch, 0, dec_bytes + zone_bytes, ctx->stream));
    }
    int rc = format == SWGPU_FMT_ENRICHED
                 ? sw_decode_enriched_launch(ctx, bytes, rec_off, n, rows_cap, d, nullptr, status, first_row, d_rows, rec_base,
                                             row_base, row_next, small)
             : format == SWGPU_FMT_JSON
                 ? sw_decode_json_launch(ctx, bytes, rec_off, n, rows_cap, d, status, first_row, d_rows, rec_base, row_base,
                                         row_next, small)
                 : sw_decode_pb_launch(ctx, bytes, rec_off, n, rows_cap, d, status, first_row, d_rows, rec_base, row_base,
                                       row_next, small);
    if (rc) return rc;
    if (small)
        return sw_zone_single_launch(ctx, d->v1, d->v0, d->meta, rows_cap, d_rows, pairs_cap, pairs, d_pairs_n,
                                     (uint8_t *)ctx->scratch + dec_bytes);
    return sw_zone_launch(ctx, d->v1, d->v0, d->meta, rows_cap, d_rows, pairs_cap, pairs, d_pairs_n, row_base);
}

#define PL_SLOTS 3
#define PL_NCOL 13

struct PlSlot {
    swgpu_ctx::Buf bytes, off, cols, status, first, pairs, counts;  // device
    cudaEvent_t h2d_done = nullptr, compute_done = nullptr, d2h_done = nullptr;
    uint32_t rows_cap = 0;
    uint64_t pairs_cap = 0;
};
struct PlCounts {  // pinned mailbox per slot
    uint32_t rows;
    uint32_t pad;
    uint64_t pairs;
};
struct Pipeline {
    cudaStream_t s_in = nullptr, s_out = nullptr;
    PlSlot slot[PL_SLOTS];
    PlCounts *mail = nullptr;         // pinned + mapped, PL_SLOTS entries
    PlCounts *mail_dev = nullptr;     // the same memory as the device sees it
    uint32_t *row_totals = nullptr;   // device: running row total before each chunk
    size_t row_totals_cap = 0;
    // small-batch path: one staging block each way, host (pinned) and device
    swgpu_ctx::Buf sm_hin, sm_hout, sm_din, sm_dout;
    bool ready = false;
    // swgpu_submit_decode_zone_test: one worker thread per context runs the queued calls in order
    std::thread worker;
    std::mutex qmu;
    std::condition_variable qcv, dcv;
    std::deque<std::pair<uint64_t, std::function<int()>>> queue;
    std::map<uint64_t, int> done;
    uint64_t next_ticket = 1;
    bool stop = false;
};
static void pl_sync_streams(swgpu_ctx *ctx) {
    Pipeline *p = (Pipeline *)ctx->pipeline;
    cudaStreamSynchronize(ctx->stream);
    if (p && p->s_in) cudaStreamSynchronize(p->s_in);
    if (p && p->s_out) cudaStreamSynchronize(p->s_out);
}

static void pl_sync_streams(swgpu_ctx *ctx);
static int pl_ensure_buf(swgpu_ctx *ctx, swgpu_ctx::Buf &b, size_t bytes) {
    if (bytes <= b.cap) return SWGPU_OK;
    if (b.p) {
        // growth after start-up (swgpu_reserve_batch avoids it): wait for this context's three streams — not
        // for the device — before the old buffer goes.  cudaFree itself still serialises with the other contexts.
        pl_sync_streams(ctx);
        SW_CUDA(ctx, cudaFree(b.p));
        b.p = nullptr;
        b.cap = 0;
    }
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess) {
        SW_SET_ERR(ctx, "cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
        b.p = nullptr;
        return SWGPU_E_NOMEM;
    }
    b.cap = want;
    return SWGPU_OK;
}


// The chunk's row and alert counts go to the host through mapped pinned memory, written by this one-thread
// kernel, not through a device->host copy: a copy would queue on the download engine behind the previous
// chunk's result copies, and the host (which needs the counts to size this chunk's downloads) would
// learn them only after those finished — one idle gap on the engine per chunk.
__global__ void pl_publish_kernel(const uint32_t *rows, const uint64_t *pairs, PlCounts *mail) {
    mail->rows = *rows;
    mail->pad = 0;
    mail->pairs = *pairs;
    __threadfence_system();
}

static int pl_get(swgpu_ctx *ctx, Pipeline **out) {
    Pipeline *p = (Pipeline *)ctx->pipeline;
    if (!p) {
        p = new Pipeline();
        ctx->pipeline = p;
    }
    if (!p->ready) {
        SW_CUDA(ctx, cudaStreamCreateWithFlags(&p->s_in, cudaStreamNonBlocking));
        SW_CUDA(ctx, cudaStreamCreateWithFlags(&p->s_out, cudaStreamNonBlocking));
        for (int i = 0; i < PL_SLOTS; i++) {
            SW_CUDA(ctx, cudaEventCreateWithFlags(&p->slot[i].h2d_done, cudaEventDisableTiming));
            SW_CUDA(ctx, cudaEventCreateWithFlags(&p->slot[i].compute_done, cudaEventDisableTiming));
            SW_CUDA(ctx, cudaEventCreateWithFlags(&p->slot[i].d2h_done, cudaEventDisableTiming));
        }
        SW_CUDA(ctx, cudaHostAlloc((void **)&p->mail, sizeof(PlCounts) * PL_SLOTS, cudaHostAllocMapped));
        SW_CUDA(ctx, cudaHostGetDevicePointer((void **)&p->mail_dev, p->mail, 0));
        p->ready = true;
    }
    *out = p;
    return SWGPU_OK;
}

void sw_pipeline_free(swgpu_ctx *ctx) {
    Pipeline *p = (Pipeline *)ctx->pipeline;
    if (!p) return;
    if (p->worker.joinable()) {
        {
            std::lock_guard<std::mutex> lk(p->qmu);
            p->stop = true;
        }
        p->qcv.notify_all();
        p->worker.join();
    }
    for (int i = 0; i < PL_SLOTS; i++) {
        PlSlot &s = p->slot[i];
        swgpu_ctx::Buf *bufs[] = {&s.bytes, &s.off, &s.cols, &s.status, &s.first, &s.pairs, &s.counts};
        for (auto *b : bufs)
            if (b->p) cudaFree(b->p);
        if (s.h2d_done) cudaEventDestroy(s.h2d_done);
        if (s.compute_done) cudaEventDestroy(s.compute_done);
        if (s.d2h_done) cudaEventDestroy(s.d2h_done);
    }
    if (p->mail) cudaFreeHost(p->mail);
    if (p->row_totals) cudaFree(p->row_totals);
    if (p->sm_hin.p) cudaFreeHost(p->sm_hin.p);
    if (p->sm_hout.p) cudaFreeHost(p->sm_hout.p);
    if (p->sm_din.p) cudaFree(p->sm_din.p);
    if (p->sm_dout.p) cudaFree(p->sm_dout.p);
    if (p->s_in) cudaStreamDestroy(p->s_in);
    if (p->s_out) cudaStreamDestroy(p->s_out);
    delete p;
    ctx->pipeline = nullptr;
}

// ---- small batches (a Kafka poll of a few hundred KB).  The chunked ring above costs ~30 driver calls per
// chunk — about 0.1 ms on the calling thread, more than the GPU needs for a thousand events.  A batch of at
// most PL_SMALL_MAX records therefore goes through one staging block each way: the payload and its offsets are
// packed into pinned memory (a memcpy of tens of KB), one upload, the kernels, one download of every output
// (counts, first_row, status, all columns, alert pairs: laid out back to back on the device), one stream
// synchronise, and the results are copied out to the caller's arrays.  Returns 1 when the batch did not fit
// the block's row / alert capacity (fan-out, many Outside-condition alerts): the caller then takes the ring.
#define PL_ALIGN(x) (((x) + 255) / 256 * 256)

static int pl_small(swgpu_ctx *ctx, Pipeline *p, int32_t format, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                    uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status, uint32_t *first_row,
                    uint32_t *out_rows, uint64_t max_pairs, uint64_t *pairs, uint64_t *n_pairs) {
    const uint64_t off0 = rec_off[0], total = rec_off[n] - off0;
    const uint32_t rows_cap = n + 64;
    const uint64_t pairs_cap = (uint64_t)n * 2 + 1024;
    // in:  [rec_off (n+1) u64][payload]            out: [counts][first_row][status][13 columns][pairs]
    const size_t in_off_bytes = PL_ALIGN((size_t)(n + 1) * 8), in_bytes = in_off_bytes + PL_ALIGN(total + 64);
    const size_t o_first = 256, o_status = o_first + PL_ALIGN((size_t)(n + 2) * 4), o_cols = o_status + PL_ALIGN((size_t)n + 16);
    const size_t o_pairs = o_cols + PL_ALIGN((size_t)rows_cap * 8 * PL_NCOL), out_bytes = o_pairs + PL_ALIGN(pairs_cap * 8);
    if (int rc = sw_ensure_pinned(ctx, p->sm_hin, in_bytes)) return rc;
    if (int rc = sw_ensure_pinned(ctx, p->sm_hout, out_bytes)) return rc;
    if (int rc = sw_ensure(ctx, p->sm_din, in_bytes)) return rc;
    if (int rc = sw_ensure(ctx, p->sm_dout, out_bytes)) return rc;
    if (int rc = sw_ensure_scratch(ctx, SW_SCRATCH_HDR + ((size_t)n / 32 + 64) * 8)) return rc;
    uint8_t *hin = (uint8_t *)p->sm_hin.p, *hout = (uint8_t *)p->sm_hout.p;
    uint8_t *din = (uint8_t *)p->sm_din.p, *dout = (uint8_t *)p->sm_dout.p;
    memcpy(hin, rec_off, (size_t)(n + 1) * 8);
    if (total) memcpy(hin + in_off_bytes, bytes + off0, total);
    cudaStream_t st = ctx->stream;
    SW_CUDA(ctx, cudaMemcpyAsync(din, hin, in_off_bytes + total, cudaMemcpyHostToDevice, st));
    swgpu_decoded_cols d;
    uint64_t **dp = (uint64_t **)&d;
    for (int k = 0; k < PL_NCOL; k++) dp[k] = (uint64_t *)(dout + o_cols) + (size_t)k * rows_cap;
    uint32_t *d_rows = (uint32_t *)dout;
    uint64_t *d_pairs_n = (uint64_t *)(dout + 8);
    // record offsets are absolute: bias the byte pointer so that bytes_dev[off] is the staged copy
    const uint8_t *biased = din + in_off_bytes - off0;
    int rc = pl_decode_zone(ctx, format, biased, (const uint64_t *)din, n, rows_cap, &d, dout + o_status,
                            (uint32_t *)(dout + o_first), d_rows, 0, nullptr, nullptr, pairs_cap, (uint64_t *)(dout + o_pairs),
                            d_pairs_n);
    if (rc) return rc;
    SW_CUDA(ctx, cudaMemcpyAsync(hout, dout, out_bytes, cudaMemcpyDeviceToHost, st));
    SW_CUDA(ctx, cudaStreamSynchronize(st));
    const uint32_t rows = *(const uint32_t *)hout;
    const uint64_t np = *(const uint64_t *)(hout + 8);
    if (rows > rows_cap || np > pairs_cap) return 1;
    if (status) memcpy(status, hout + o_status, n);
    if (first_row) memcpy(first_row, hout + o_first, (size_t)(n + 1) * 4);
    const uint64_t take = rows < max_rows ? rows : max_rows;
    uint64_t *const *hcol = (uint64_t *const *)out;
    for (int k = 0; k < PL_NCOL; k++)
        if (hcol[k] && take) memcpy(hcol[k], hout + o_cols + (size_t)k * rows_cap * 8, take * 8);
    const uint64_t ptake = np < max_pairs ? np : max_pairs;
    if (ptake) memcpy(pairs, hout + o_pairs, ptake * 8);
    *out_rows = rows;
    *n_pairs = np;
    if (rows > max_rows || np > max_pairs) {
        SW_SET_ERR(ctx, "swgpu_decode_zone_test: %u rows / %llu alerts exceed max_rows %u / max_pairs %llu", rows,
                   (unsigned long long)np, max_rows, (unsigned long long)max_pairs);
        return SWGPU_E_CAPACITY;
    }
    return SWGPU_OK;
}

// Chunk size when the caller leaves it to us: about a dozen chunks per call.  Each chunk costs ~30 driver calls on the
// calling thread (~0.1 ms), so many small chunks make the host the bottleneck; a few large ones leave the first upload
// and the last download uncovered.  Measured on 10 M events: 64 K records per chunk 32.9 ms, 256 K 17.7 ms, 1 M
// 15.3 ms, 2 M 16.2 ms.
static uint32_t pl_chunk_records(uint32_t n) {
    const uint64_t want = ((uint64_t)n / 12 + 65535) / 65536 * 65536;
    return (uint32_t)(want < 131072 ? 131072 : want > 1048576 ? 1048576 : want);
}

// size the three ring slots for chunks of C records / chunk_bytes payload bytes / C alert pairs
static int pl_size_slots(swgpu_ctx *ctx, Pipeline *p, uint32_t C, uint64_t chunk_bytes, int nslots) {
    for (int i = 0; i < PL_SLOTS && i < nslots; i++) {
        PlSlot &s = p->slot[i];
        if (int rc = pl_ensure_buf(ctx, s.bytes, chunk_bytes + 64)) return rc;
        if (int rc = pl_ensure_buf(ctx, s.off, (size_t)(C + 1) * 8)) return rc;
        if (s.rows_cap < C) {
            if (int rc = pl_ensure_buf(ctx, s.cols, (size_t)C * 8 * PL_NCOL)) return rc;
            s.rows_cap = C;
        }
        if (int rc = pl_ensure_buf(ctx, s.status, (size_t)C + 16)) return rc;
        if (int rc = pl_ensure_buf(ctx, s.first, (size_t)(C + 2) * 4)) return rc;
        if (s.pairs_cap < C) {
            if (int rc = pl_ensure_buf(ctx, s.pairs, (size_t)C * 8)) return rc;
            s.pairs_cap = C;
        }
        if (int rc = pl_ensure_buf(ctx, s.counts, 64)) return rc;
    }
    return sw_ensure_scratch(ctx, SW_SCRATCH_HDR + ((size_t)C / 32 + 64) * 8);  // warp-tile scan states
}

// Allocate, at start-up, everything swgpu_decode_zone_test needs for batches of up to max_records records and
// max_bytes payload bytes, so that no cudaMalloc / cudaFree (which stall every stream of the device, the other
// contexts' too) happens once batches are flowing.  A batch that decodes to more rows than records (measurement
// fan-out) or fires more alerts than records still grows its slot.
extern "C" int32_t swgpu_reserve_batch(swgpu_ctx *ctx, uint32_t max_records, uint64_t max_bytes) {
    if (!ctx) return SWGPU_E_INVALID;
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    Pipeline *p;
    if (int rc = pl_get(ctx, &p)) return rc;
    if (max_records == 0) return SWGPU_OK;
    // the small-batch block (one staging block each way)
    {
        const uint32_t n = max_records < PL_SMALL_MAX ? max_records : PL_SMALL_MAX;
        const uint64_t total = max_records <= PL_SMALL_MAX ? max_bytes : (max_bytes / max_records + 1) * n * 2;
        const uint32_t rows_cap = n + 64;
        const uint64_t pairs_cap = (uint64_t)n * 2 + 1024;
        const size_t in_bytes = PL_ALIGN((size_t)(n + 1) * 8) + PL_ALIGN(total + 64);
        const size_t out_bytes = 256 + PL_ALIGN((size_t)(n + 2) * 4) + PL_ALIGN((size_t)n + 16) +
                                 PL_ALIGN((size_t)rows_cap * 8 * PL_NCOL) + PL_ALIGN(pairs_cap * 8);
        if (int rc = sw_ensure_pinned(ctx, p->sm_hin, in_bytes)) return rc;
        if (int rc = sw_ensure_pinned(ctx, p->sm_hout, out_bytes)) return rc;
        if (int rc = sw_ensure(ctx, p->sm_din, in_bytes)) return rc;
        if (int rc = sw_ensure(ctx, p->sm_dout, out_bytes)) return rc;
        if (int rc = sw_ensure(ctx, ctx->d_zwork, ((size_t)rows_cap * 8 + 255) / 256 * 256 + (size_t)(rows_cap / 256 + 2) * 16 + 64))
            return rc;
    }
    if (max_records > PL_SMALL_MAX) {
        const uint32_t C = pl_chunk_records(max_records);
        const uint32_t nchunks = (max_records + C - 1) / C;
        // a chunk's payload: its share of the batch with a margin for uneven record sizes
        const uint64_t chunk_bytes = nchunks <= 1 ? max_bytes : (uint64_t)((double)max_bytes / max_records * C * 1.25) + 65536;
        if (int rc = pl_size_slots(ctx, p, C, chunk_bytes, PL_SLOTS)) return rc;
        if (int rc = sw_ensure(ctx, ctx->d_zwork, ((size_t)C * 8 + 255) / 256 * 256 + (size_t)(C / 256 + 2) * 16 + 64)) return rc;
        if ((size_t)nchunks + 1 > p->row_totals_cap) {
            SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            if (p->row_totals) SW_CUDA(ctx, cudaFree(p->row_totals));
            p->row_totals_cap = (size_t)nchunks + 64;
            SW_CUDA(ctx, cudaMalloc((void **)&p->row_totals, sizeof(uint32_t) * p->row_totals_cap));
        }
    }
    return SWGPU_OK;
}

// Device buffers in, device buffers out, asynchronous on the context's stream: swgpu_decode_*_dev followed by
// swgpu_zone_test_dev on the decoded columns, in one call.
extern "C" int32_t swgpu_decode_zone_test_dev(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off,
                                              uint32_t n, uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                                              uint32_t *first_row, uint32_t *out_rows_dev, uint64_t max_pairs, uint64_t *pairs,
                                              uint64_t *n_pairs_dev) {
    if (!ctx) return SWGPU_E_INVALID;
    if (!rec_off || !out || !first_row || !out_rows_dev || !n_pairs_dev || (n && !status) || (max_pairs && !pairs) ||
        (format != SWGPU_FMT_PROTOBUF && format != SWGPU_FMT_JSON && format != SWGPU_FMT_ENRICHED)) {
        SW_SET_ERR(ctx, "swgpu_decode_zone_test_dev: bad argument");
        return SWGPU_E_INVALID;
    }
    if (!ctx->zone_blob) {
        SW_SET_ERR(ctx, "swgpu_decode_zone_test_dev: no zone table (call swgpu_set_zones first)");
        return SWGPU_E_STATE;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    return pl_decode_zone(ctx, format, bytes, rec_off, n, max_rows, out, status, first_row, out_rows_dev, 0, nullptr, nullptr,
                          max_pairs, pairs, n_pairs_dev);
}

extern "C" int32_t swgpu_decode_zone_test(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off,
                                          uint32_t n, uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                                          uint32_t *first_row, uint32_t *out_rows, uint64_t max_pairs, uint64_t *pairs,
                                          uint64_t *n_pairs, uint32_t chunk_records) {
    if (!ctx) return SWGPU_E_INVALID;
    if (!rec_off || !out || !out_rows || !n_pairs || (format != SWGPU_FMT_PROTOBUF && format != SWGPU_FMT_JSON && format != SWGPU_FMT_ENRICHED) ||
        (max_pairs && !pairs) || (n && rec_off[n] && !bytes)) {
        SW_SET_ERR(ctx, "swgpu_decode_zone_test: bad argument");
        return SWGPU_E_INVALID;
    }
    if (!ctx->zone_blob) {
        SW_SET_ERR(ctx, "swgpu_decode_zone_test: no zone table (call swgpu_set_zones first)");
        return SWGPU_E_STATE;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    *out_rows = 0;
    *n_pairs = 0;
    if (first_row) first_row[0] = 0;
    if (n == 0) return SWGPU_OK;
    Pipeline *p;
    if (int rc = pl_get(ctx, &p)) return rc;
    if (n <= PL_SMALL_MAX && (chunk_records == 0 || chunk_records >= n)) {
        const int rc = pl_small(ctx, p, format, bytes, rec_off, n, max_rows, out, status, first_row, out_rows, max_pairs,
                                pairs, n_pairs);
        if (rc != 1) return rc;
        *out_rows = 0;  // did not fit the small block: the ring below re-runs it with exact sizes
        *n_pairs = 0;
    }
    const uint32_t C = chunk_records ? chunk_records : pl_chunk_records(n);
    const uint32_t nchunks = (n + C - 1) / C;
    uint64_t *const *hcol = (uint64_t *const *)out;

    // size the slots for the largest chunk
    uint64_t max_chunk_bytes = 0;
    for (uint32_t c = 0; c < nchunks; c++) {
        const uint32_t a = c * C, b = (a + C < n) ? a + C : n;
        const uint64_t sz = rec_off[b] - rec_off[a];
        if (sz > max_chunk_bytes) max_chunk_bytes = sz;
    }
    // a DeviceMeasurements payload can fan out; the row capacity per chunk starts at the record
    // count and a chunk that overflows it is re-run with the exact size (rare)
    if (int rc = pl_size_slots(ctx, p, C, max_chunk_bytes, (int)nchunks)) return rc;
    // running row totals: entry k = rows before chunk k (device), ring of nchunks+1
    {
        if ((size_t)nchunks + 1 > p->row_totals_cap) {
            SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            if (p->row_totals) SW_CUDA(ctx, cudaFree(p->row_totals));
            p->row_totals_cap = (size_t)nchunks + 64;
            SW_CUDA(ctx, cudaMalloc((void **)&p->row_totals, sizeof(uint32_t) * p->row_totals_cap));
        }
        SW_CUDA(ctx, cudaMemsetAsync(p->row_totals, 0, sizeof(uint32_t) * ((size_t)nchunks + 1), ctx->stream));
    }
    cudaStream_t s_in = p->s_in, s_cmp = ctx->stream, s_out = p->s_out;
    uint64_t rows_done = 0, pairs_done = 0;
    bool rows_overflow = false, pairs_overflow = false;

    auto issue = [&](uint32_t c) -> int {  // H2D + kernels + count read-back for chunk c
        PlSlot &s = p->slot[c % PL_SLOTS];
        const uint32_t a = c * C, b = (a + C < n) ? a + C : n, nc = b - a;
        const uint64_t off0 = rec_off[a], sz = rec_off[b] - off0;
        // the slot's previous occupant must have been copied out
        SW_CUDA(ctx, cudaStreamWaitEvent(s_in, s.d2h_done, 0));
        if (sz) SW_CUDA(ctx, cudaMemcpyAsync(s.bytes.p, bytes + off0, sz, cudaMemcpyHostToDevice, s_in));
        SW_CUDA(ctx, cudaMemcpyAsync(s.off.p, rec_off + a, (size_t)(nc + 1) * 8, cudaMemcpyHostToDevice, s_in));
        SW_CUDA(ctx, cudaEventRecord(s.h2d_done, s_in));
        SW_CUDA(ctx, cudaStreamWaitEvent(s_cmp, s.h2d_done, 0));
        swgpu_decoded_cols d;
        uint64_t **dp = (uint64_t **)&d;
        for (int k = 0; k < PL_NCOL; k++) dp[k] = (uint64_t *)s.cols.p + (size_t)k * s.rows_cap;
        uint32_t *d_rows = (uint32_t *)s.counts.p;
        uint64_t *d_pairs_n = (uint64_t *)((uint8_t *)s.counts.p + 8);
        // record offsets are absolute: bias the byte pointer so that bytes_dev[off] is the chunk copy
        const uint8_t *biased = (const uint8_t *)s.bytes.p - off0;
        int rc = pl_decode_zone(ctx, format, biased, (const uint64_t *)s.off.p, nc, s.rows_cap, &d, (uint8_t *)s.status.p,
                                (uint32_t *)s.first.p, d_rows, a, p->row_totals + c, p->row_totals + c + 1, s.pairs_cap,
                                (uint64_t *)s.pairs.p, d_pairs_n);
        if (rc) return rc;
        pl_publish_kernel<<<1, 1, 0, s_cmp>>>(d_rows, d_pairs_n, p->mail_dev + c % PL_SLOTS);
        SW_CUDA(ctx, cudaGetLastError());
        SW_CUDA(ctx, cudaEventRecord(s.compute_done, s_cmp));
        return SWGPU_OK;
    };
    auto drain = [&](uint32_t c) -> int {  // wait for chunk c's kernels, then copy its results out
        PlSlot &s = p->slot[c % PL_SLOTS];
        const uint32_t a = c * C, b = (a + C < n) ? a + C : n, nc = b - a;
        SW_CUDA(ctx, cudaEventSynchronize(s.compute_done));
        const PlCounts cnt = p->mail[c % PL_SLOTS];
        if (cnt.rows > s.rows_cap || cnt.pairs > s.pairs_cap) {
            // fan-out or many alerts: grow this slot and re-run the chunk synchronously (rare)
            pl_sync_streams(ctx);
            if (cnt.rows > s.rows_cap) {
                if (int rc = pl_ensure_buf(ctx, s.cols, (size_t)cnt.rows * 8 * PL_NCOL)) return rc;
                s.rows_cap = cnt.rows;
            }
            if (cnt.pairs > s.pairs_cap) {
                if (int rc = pl_ensure_buf(ctx, s.pairs, (size_t)cnt.pairs * 8)) return rc;
                s.pairs_cap = cnt.pairs;
            }
            SW_CUDA(ctx, cudaEventRecord(s.d2h_done, s_out));
            if (int rc = issue(c)) return rc;
            SW_CUDA(ctx, cudaEventSynchronize(s.compute_done));
            const PlCounts c2 = p->mail[c % PL_SLOTS];
            if (c2.rows > s.rows_cap || c2.pairs > s.pairs_cap) {
                // the alert count depended on rows that were cut off the first time: once more
                pl_sync_streams(ctx);
                if (c2.pairs > s.pairs_cap) {
                    if (int rc = pl_ensure_buf(ctx, s.pairs, (size_t)c2.pairs * 8)) return rc;
                    s.pairs_cap = c2.pairs;
                }
                SW_CUDA(ctx, cudaEventRecord(s.d2h_done, s_out));
                if (int rc = issue(c)) return rc;
                SW_CUDA(ctx, cudaEventSynchronize(s.compute_done));
            }
        }
        const PlCounts fin = p->mail[c % PL_SLOTS];
        SW_CUDA(ctx, cudaStreamWaitEvent(s_out, s.compute_done, 0));
        if (status) SW_CUDA(ctx, cudaMemcpyAsync(status + a, s.status.p, nc, cudaMemcpyDeviceToHost, s_out));
        if (first_row)
            SW_CUDA(ctx, cudaMemcpyAsync(first_row + a, s.first.p, (size_t)(nc + 1) * 4, cudaMemcpyDeviceToHost, s_out));
        uint64_t take = fin.rows;
        if (rows_done + take > max_rows) {
            take = max_rows > rows_done ? max_rows - rows_done : 0;
            rows_overflow = true;
        }
        if (take)
            for (int k = 0; k < PL_NCOL; k++)
                if (hcol[k])
                    SW_CUDA(ctx, cudaMemcpyAsync(hcol[k] + rows_done, (uint64_t *)s.cols.p + (size_t)k * s.rows_cap, take * 8,
                                                 cudaMemcpyDeviceToHost, s_out));
        uint64_t ptake = fin.pairs;
        if (pairs_done + ptake > max_pairs) {
            ptake = max_pairs > pairs_done ? max_pairs - pairs_done : 0;
            pairs_overflow = true;
        }
        if (ptake) SW_CUDA(ctx, cudaMemcpyAsync(pairs + pairs_done, s.pairs.p, ptake * 8, cudaMemcpyDeviceToHost, s_out));
        SW_CUDA(ctx, cudaEventRecord(s.d2h_done, s_out));
        rows_done += fin.rows;
        pairs_done += fin.pairs;
        return SWGPU_OK;
    };

    // software pipeline: keep up to PL_SLOTS chunks in flight
    uint32_t issued = 0, drained = 0;
    while (drained < nchunks) {
        while (issued < nchunks && issued - drained < PL_SLOTS) {
            if (int rc = issue(issued)) return rc;
            issued++;
        }
        if (int rc = drain(drained)) return rc;
        drained++;
    }
    SW_CUDA(ctx, cudaStreamSynchronize(s_out));
    SW_CUDA(ctx, cudaStreamSynchronize(s_cmp));
    *out_rows = (uint32_t)rows_done;
    *n_pairs = pairs_done;
    if (rows_overflow || pairs_overflow) {
        SW_SET_ERR(ctx, "swgpu_decode_zone_test: %llu rows / %llu alerts exceed max_rows %u / max_pairs %llu",
                   (unsigned long long)rows_done, (unsigned long long)pairs_done, max_rows, (unsigned long long)max_pairs);
        return SWGPU_E_CAPACITY;
    }
    return SWGPU_OK;
}


// ---- asynchronous form -------------------------------------------------------------------------------------------------
// swgpu_submit_decode_zone_test queues the same call on the context's worker thread and returns a ticket at once;
// swgpu_wait blocks until that call is done and returns its status (the outputs are valid then).  The chunked ring
// is driven from the host (it needs each chunk's counts to size the downloads), so the asynchronous form runs it
// on a thread the library owns instead of the caller's: the binding's consumer thread can poll Kafka or stage the
// next batch meanwhile.  Tickets complete in submission order.  While tickets are pending no other call may be made
// on the context (calls on one context are serial); the caller's buffers belong to the library until swgpu_wait.
static void pl_worker(Pipeline *p) {
    for (;;) {
        std::pair<uint64_t, std::function<int()>> job;
        {
            std::unique_lock<std::mutex> lk(p->qmu);
            p->qcv.wait(lk, [&] { return p->stop || !p->queue.empty(); });
            if (p->queue.empty()) return;  // stop
            job = std::move(p->queue.front());
            p->queue.pop_front();
        }
        const int rc = job.second();
        {
            std::lock_guard<std::mutex> lk(p->qmu);
            p->done[job.first] = rc;
        }
        p->dcv.notify_all();
    }
}

extern "C" int32_t swgpu_submit_decode_zone_test(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off,
                                                 uint32_t n, uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                                                 uint32_t *first_row, uint32_t *out_rows, uint64_t max_pairs, uint64_t *pairs,
                                                 uint64_t *n_pairs, uint32_t chunk_records, uint64_t *ticket) {
    if (!ctx || !ticket || !out) {
        if (ctx) SW_SET_ERR(ctx, "swgpu_submit_decode_zone_test: NULL argument");
        return SWGPU_E_INVALID;
    }
    Pipeline *p;
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    if (int rc = pl_get(ctx, &p)) return rc;
    const swgpu_decoded_cols cols = *out;  // the struct is copied: only the arrays must outlive the call
    std::lock_guard<std::mutex> lk(p->qmu);
    if (!p->worker.joinable()) p->worker = std::thread(pl_worker, p);
    const uint64_t t = p->next_ticket++;
    p->queue.emplace_back(t, [=]() {
        return (int)swgpu_decode_zone_test(ctx, format, bytes, rec_off, n, max_rows, &cols, status, first_row, out_rows, max_pairs,
                                           pairs, n_pairs, chunk_records);
    });
    *ticket = t;
    p->qcv.notify_one();
    return SWGPU_OK;
}

extern "C" int32_t swgpu_wait(swgpu_ctx *ctx, uint64_t ticket) {
    if (!ctx) return SWGPU_E_INVALID;
    Pipeline *p = (Pipeline *)ctx->pipeline;
    if (!p || ticket == 0 || ticket >= p->next_ticket) {
        SW_SET_ERR(ctx, "swgpu_wait: unknown ticket %llu", (unsigned long long)ticket);
        return SWGPU_E_INVALID;
    }
    std::unique_lock<std::mutex> lk(p->qmu);
    p->dcv.wait(lk, [&] { return p->done.count(ticket) != 0; });
    const int rc = p->done[ticket];
    p->done.erase(ticket);
    return rc;
}
