// avr_render_host.h — host helpers shared by avr_volpath.cu and avr_graph.cu.
#pragma once
#include "avr_internal.h"
#include "avr_volpath.cuh"

namespace avr {
// Validates an avr_render_desc, uploads its light/sensor tables and fills the kernel parameter block.
avr_status fill_render_params(avr_ctx *ctx, const avr_render_desc *d, RenderParams *rp, cudaStream_t stream);
// avr_filter_table -> the packed sampling tables of RenderParams::flt (host arithmetic in the reference's order).
avr_status pack_filter_table(avr_ctx *ctx, const avr_filter_table *t, const float radius[2], std::vector<float> *out);
// Device film scratch of the context, zeroed on `stream` (grown on demand).
avr_status film_scratch(avr_ctx *ctx, size_t bytes, cudaStream_t stream);
// film[i] += device film: pinned staging, the copy runs in chunks on ctx->stream while host threads add
// the chunks that have arrived.
avr_status film_add_to_host(avr_ctx *ctx, size_t bytes, double *film);
}  // namespace avr
