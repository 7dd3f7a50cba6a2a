// decode_json.cu — batched JsonDeviceRequestDecoder.decode (placeholder until the kernel lands)
#include "swgpu_common.cuh"

int sw_stage_decode(swgpu_ctx *ctx, bool json, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                    uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status, uint32_t *first_row,
                    uint32_t *out_rows);

extern "C" int32_t swgpu_decode_json_dev(swgpu_ctx *ctx, const uint8_t *, const uint64_t *, uint32_t, uint32_t,
                                         const swgpu_decoded_cols *, uint8_t *, uint32_t *, uint32_t *) {
    SW_SET_ERR(ctx, "swgpu_decode_json: not built yet");
    return SWGPU_E_STATE;
}
extern "C" int32_t swgpu_decode_json(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                                     uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                                     uint32_t *first_row, uint32_t *out_rows) {
    return sw_stage_decode(ctx, true, bytes, rec_off, n, max_rows, out, status, first_row, out_rows);
}
