// Gather-form ("owner tile") splatting with the divide fused -- placeholder until the tiled kernel lands.
#include "vfi_common.cuh"

extern "C" int vfi_splat_fused_f32(const float*, const float*, const float*, float*, int, int, int, int, int, int, float, void*,
                                   size_t, void*) {
    VFI_UNSUPPORTED("vfi_splat_fused_f32: not built in this version; use vfi_splat_f32 + vfi_splat_normalize_f32");
}
