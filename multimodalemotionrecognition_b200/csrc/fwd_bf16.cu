#include "kernels_bf16.cuh"
namespace xf {
int packed_bytes_bf16(const xattn_config& c, int64_t* bytes) { *bytes = 16; return XATTN_OK; }
int workspace_bytes_bf16(const xattn_config& c, int64_t* bytes) { *bytes = 16; return XATTN_OK; }
int pack_weights_bf16(const xattn_config& c, const float* params, void* packed, cudaStream_t st) { XF_FAIL(XATTN_ERR_DTYPE, "bf16 path not built yet"); }
int forward_bf16(const xattn_config& c, const float* params, const void* packed, const void* xv, const void* xa, float* logits, void* ws, cudaStream_t st) { XF_FAIL(XATTN_ERR_DTYPE, "bf16 path not built yet"); }
}
