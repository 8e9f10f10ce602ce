// Entry points of the bf16 tcgen05 inference path (fwd_bf16.cu).
#pragma once
#include "common.cuh"

namespace xf {
int packed_bytes_bf16(const xattn_config& c, int64_t* bytes);
int workspace_bytes_bf16(const xattn_config& c, int64_t* bytes);
int pack_weights_bf16(const xattn_config& c, const float* params, void* packed, cudaStream_t st);
// fork = false keeps the video branch on the caller's stream: no private stream, no events (xattn_forward_noalloc).
int forward_bf16(const xattn_config& c, const float* params, const void* packed, const void* xv, const void* xa, float* logits,
                 void* ws, cudaStream_t st, bool fork = true);
}  // namespace xf
