// Flow upsampling between pyramid levels: out = res + mul * resize(x, scale_factor=2)
// Reference: BaseModel.resize (models/BaseModel.py:19-22: F.interpolate(bilinear, align_corners=True,
// recompute_scale_factor=False)) as used by modules/deformable_attn.py:127-131
//     up_ft = conv_res_flow(movement) + 2.0 * BaseModel.resize(ft, scale_factor=2.0)
// One pass instead of three (upsample, * 2.0, +); the arithmetic follows ATen's upsample_bilinear2d:
//     scale = (in - 1) / (out - 1);  src = scale * dst;  i0 = min(int(src), in - 1);  l1 = src - i0;  i1 = min(i0 + 1, in - 1)
//     v = l0y * (l0x * x[i0y][i0x] + l1x * x[i0y][i1x]) + l1y * (l0x * x[i1y][i0x] + l1x * x[i1y][i1x])
// align_corners=True makes the source position depend on the FRAME size, so a crop (4K tiles) passes the frame extent
// of the input level (frame_ext, include/vfi_b200.h) and samples exactly where the whole frame would.
#include "vfi_common.cuh"

namespace vfi {

struct Str4r {
    int64_t n, c, h, w;
};

__global__ void __launch_bounds__(256) resize2x_kernel(const float* __restrict__ x, Str4r sx, const float* __restrict__ res, Str4r sr,
                                                       float mul, float* __restrict__ out, int B, int C, int H, int W,
                                                       FrameExt fe) {
    const int Wo = 2 * W, Ho = 2 * H;
    const int64_t total = (int64_t)B * C * Ho * Wo;
    const float rw = fe.Wg > 1 ? (float)(fe.Wg - 1) / (float)(2 * fe.Wg - 1) : 0.0f;
    const float rh = fe.Hg > 1 ? (float)(fe.Hg - 1) / (float)(2 * fe.Hg - 1) : 0.0f;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int xo = (int)(idx % Wo);
        int64_t t = idx / Wo;
        const int yo = (int)(t % Ho);
        t /= Ho;
        const int c = (int)(t % C), b = (int)(t / C);
        const float srcx = __fmul_rn(rw, (float)(xo + 2 * fe.x0)), srcy = __fmul_rn(rh, (float)(yo + 2 * fe.y0));
        const int gx0 = min((int)srcx, fe.Wg - 1), gy0 = min((int)srcy, fe.Hg - 1);
        const float l1x = fminf(fmaxf(srcx - (float)gx0, 0.0f), 1.0f), l1y = fminf(fmaxf(srcy - (float)gy0, 0.0f), 1.0f);
        const float l0x = 1.0f - l1x, l0y = 1.0f - l1y;
        const int gx1 = min(gx0 + 1, fe.Wg - 1), gy1 = min(gy0 + 1, fe.Hg - 1);
        const int x0 = min(max(gx0 - fe.x0, 0), W - 1), x1 = min(max(gx1 - fe.x0, 0), W - 1);
        const int y0 = min(max(gy0 - fe.y0, 0), H - 1), y1 = min(max(gy1 - fe.y0, 0), H - 1);
        const float* p = x + b * sx.n + c * sx.c;
        const float v00 = __ldg(p + y0 * sx.h + x0 * sx.w), v01 = __ldg(p + y0 * sx.h + x1 * sx.w);
        const float v10 = __ldg(p + y1 * sx.h + x0 * sx.w), v11 = __ldg(p + y1 * sx.h + x1 * sx.w);
        const float top = __fadd_rn(__fmul_rn(l0x, v00), __fmul_rn(l1x, v01));
        const float bot = __fadd_rn(__fmul_rn(l0x, v10), __fmul_rn(l1x, v11));
        float v = __fmul_rn(mul, __fadd_rn(__fmul_rn(l0y, top), __fmul_rn(l1y, bot)));
        if (res != nullptr) v = __fadd_rn(__ldg(res + b * sr.n + c * sr.c + yo * sr.h + xo * sr.w), v);
        out[idx] = v;
    }
}

}  // namespace vfi

extern "C" int vfi_resize2x_f32(const float* x, const int64_t x_strides[4], const float* res, const int64_t res_strides[4], float mul,
                                float* out, int B, int C, int H, int W, const int* frame_ext, void* stream) {
    using namespace vfi;
    if (B == 0 || C == 0) return VFI_OK;
    VFI_REQUIRE(x && x_strides && out && (res == nullptr || res_strides), "vfi_resize2x_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 0 && H >= 1 && W >= 1, "vfi_resize2x_f32: bad shape");
    const FrameExt fe = frame_ext_of(frame_ext, H, W);
    VFI_REQUIRE(frame_ext_ok(fe, H, W), "vfi_resize2x_f32: the crop does not lie inside the frame");
    const Str4r sx{x_strides[0], x_strides[1], x_strides[2], x_strides[3]};
    Str4r sr{0, 0, 0, 0};
    if (res) sr = Str4r{res_strides[0], res_strides[1], res_strides[2], res_strides[3]};
    const int64_t total = (int64_t)B * C * 4 * H * W;
    const int blocks = (int)min64(ceil_div(total, 256), (int64_t)kNumSMs * 16);
    resize2x_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, sx, res, sr, mul, out, B, C, H, W, fe);
    VFI_CHECK_LAUNCH("resize2x_kernel");
    return VFI_OK;
}
