// avr_render_host.h — host helpers shared by avr_volpath.cu and avr_graph.cu.
#pragma once
#include "avr_internal.h"
#include "avr_volpath.cuh"

namespace avr {
// Validates an avr_render_desc, uploads its light/sensor tables and fills the kernel parameter block.
avr_status fill_render_params(avr_ctx *ctx, const avr_render_desc *d, RenderParams *rp, cudaStream_t stream);
}  // namespace avr
