// tcgen05 / TMEM path of DCNv2 -- placeholder until the tensor-core kernel lands (see dcn.cu for the SIMT path).
#include "vfi_common.cuh"

namespace vfi {
struct DcnParams {
    int Cin, Cout, H, W, Ho, Wo, kh, kw, stride, pad, dil, WG, OG;
};
// 0 = unsupported, 1 = supported, 2 = supported and preferred by algo 0 (auto)
int dcn_tc_supported(const DcnParams&) { return 0; }
size_t dcn_tc_workspace_bytes(int, const DcnParams&) { return 0; }
int launch_dcn_tc(const float*, const float*, const float*, const float*, const float*, float*, int, const DcnParams&, void*,
                  size_t, cudaStream_t) {
    VFI_UNSUPPORTED("vfi_dcn_f32: tcgen05 path not built in this version");
}
}  // namespace vfi
