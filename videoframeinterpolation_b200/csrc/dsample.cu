// Grouped multi-point bilinear sampler, materialising form (API parity with the reference's
// _get_ref_feats: modules/deformable_attn.py:88-101, modules/madat.py:161-174).
//
// Write-dominated: 4*(C + 2*nG*nS + C*nS) bytes per pixel.  One thread per (pixel, group, sample): the
// coordinate replay and the four weights are computed once and reused for the C/nG channels of the
// group; a warp covers 32 consecutive pixels so the (B, C, nS, HW) stores are 128 B coalesced.
#include "sampler.cuh"

namespace vfi {

__global__ void __launch_bounds__(256) dsample_kernel(const float* __restrict__ feat, const float* __restrict__ coords,
                                                      const float* __restrict__ lin_x, const float* __restrict__ lin_y,
                                                      float* __restrict__ out, float* __restrict__ grid_out, int C, int H, int W,
                                                      int nG, int nS, FrameExt fe, int row0, int rows) {
    // coords / out / grid_out hold the output rows [row0, row0 + rows) only (the whole plane when rows == H); feat is whole
    const int64_t HW = (int64_t)H * W, HWs = (int64_t)rows * W;
    const int64_t pix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= HWs) return;
    const int gs = blockIdx.y;              // g * nS + s
    const int g = gs / nS, s = gs - g * nS;
    const int b = blockIdx.z;
    const int ys = (int)(pix / W), x = (int)(pix - (int64_t)ys * W), y = row0 + ys;
    const int gc = C / nG;
    const float* cp = coords + (((int64_t)b * nG * nS + gs) * 2) * HWs + pix;
    const TapGeom tg = sampler_geom(ld_stream(cp), ld_stream(cp + HWs), lin_x[x], lin_y[y], H, W, fe);
    if (grid_out != nullptr) {
        float2* gp = reinterpret_cast<float2*>(grid_out) + ((int64_t)b * nG * nS + gs) * HWs + pix;
        *gp = make_float2(tg.gx, tg.gy);
    }
    const float* fp = feat + ((int64_t)b * C + (int64_t)g * gc) * HW;
    float* op = out + (((int64_t)b * C + (int64_t)g * gc) * nS + s) * HWs + pix;
    for (int c = 0; c < gc; ++c) {
        st_stream(op, sample_plane(fp, tg, W));
        fp += HW;
        op += (int64_t)nS * HWs;
    }
}

}  // namespace vfi

extern "C" int vfi_dsample_rows_f32(const float* feat, const float* coords, const float* lin_x, const float* lin_y, float* out,
                                    float* grid_out, int B, int C, int H, int W, int nG, int nS, int row0, int rows,
                                    const int* frame_ext, void* stream) {
    using namespace vfi;
    if (B == 0 || rows == 0) return VFI_OK;
    VFI_REQUIRE(feat && coords && lin_x && lin_y && out, "vfi_dsample_rows_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1 && nG >= 1 && nS >= 1, "vfi_dsample_rows_f32: bad shape");
    VFI_REQUIRE(row0 >= 0 && rows >= 0 && row0 + rows <= H, "vfi_dsample_rows_f32: rows [%d, %d) outside 0..%d", row0, row0 + rows, H);
    VFI_REQUIRE(C % nG == 0, "vfi_dsample_rows_f32: C=%d not divisible by nG=%d", C, nG);
    VFI_REQUIRE((int64_t)nG * nS <= 65535 && B <= 65535, "vfi_dsample_rows_f32: nG*nS or B exceeds grid limits");
    const FrameExt fe = frame_ext_of(frame_ext, H, W);
    VFI_REQUIRE(frame_ext_ok(fe, H, W), "vfi_dsample_rows_f32: the crop does not lie inside the frame");
    dim3 grid((unsigned)ceil_div((int64_t)rows * W, 256), (unsigned)(nG * nS), (unsigned)B);
    dsample_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(feat, coords, lin_x, lin_y, out, grid_out, C, H, W, nG, nS,
                                                                        fe, row0, rows);
    VFI_CHECK_LAUNCH("dsample_kernel");
    return VFI_OK;
}

extern "C" int vfi_dsample_f32(const float* feat, const float* coords, const float* lin_x, const float* lin_y, float* out,
                               float* grid_out, int B, int C, int H, int W, int nG, int nS, const int* frame_ext, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(feat && coords && lin_x && lin_y && out, "vfi_dsample_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1 && nG >= 1 && nS >= 1, "vfi_dsample_f32: bad shape");
    VFI_REQUIRE(C % nG == 0, "vfi_dsample_f32: C=%d not divisible by nG=%d", C, nG);
    VFI_REQUIRE((int64_t)nG * nS <= 65535 && B <= 65535, "vfi_dsample_f32: nG*nS or B exceeds grid limits");
    const FrameExt fe = frame_ext_of(frame_ext, H, W);
    VFI_REQUIRE(frame_ext_ok(fe, H, W), "vfi_dsample_f32: the crop does not lie inside the frame");
    const int64_t HW = (int64_t)H * W;
    dim3 grid((unsigned)ceil_div(HW, 256), (unsigned)(nG * nS), (unsigned)B);
    dsample_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(feat, coords, lin_x, lin_y, out, grid_out, C, H, W, nG, nS, fe,
                                                                        0, H);
    VFI_CHECK_LAUNCH("dsample_kernel");
    return VFI_OK;
}
