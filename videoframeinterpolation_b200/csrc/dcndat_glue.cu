// Glue of the DCNDAT block (models/archive/DCNDAT.py:150-157 + 69-70): one pass from the raw output of
// conv_res_offset_mask to the three tensors its consumers read, instead of ~12 elementwise / copy kernels:
//   raw (B, 3n, H, W), n = nG*nS: channels [0, 2n) are residual offsets, viewed (n, 2) = (x, y); [2n, 3n) the DCN mask
//   coords (B, n, 2, H, W)  = 2*tanh(raw[2j + {0,1}]) + ft[{0,1}]                 -> sampler / fused attention
//   dcn_off (B, 2, n, H, W) = coords.permute(0,2,1,3,4).flip(1): all y, then all x -> deform_conv2d offsets (line 69)
//   mask (B, n, H, W)       = sigmoid(raw[2n + j])                                  (line 70)
// Rounding follows the PyTorch composition op by op (tanhf, separate multiply and add; 1 / (1 + expf(-x))).
#include "vfi_common.cuh"

namespace vfi {

struct Str4g {
    int64_t n, c, h, w;
};

__global__ void __launch_bounds__(256) dcndat_coords_kernel(const float* __restrict__ raw, Str4g sr, const float* __restrict__ ft, Str4g sf,
                                                            float* __restrict__ coords, float* __restrict__ dcn_off,
                                                            float* __restrict__ mask, int n, int H, int W) {
    const int b = blockIdx.y;
    const int64_t HW = (int64_t)H * W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < HW; i += (int64_t)gridDim.x * blockDim.x) {
        const int y = (int)(i / W), x = (int)(i - (int64_t)y * W);
        const float* rp = raw + b * sr.n + y * sr.h + x * sr.w;
        const float* fp = ft + b * sf.n + y * sf.h + x * sf.w;
        const float fx = __ldg(fp), fy = __ldg(fp + sf.c);
        float* cp = coords + (int64_t)b * n * 2 * HW + i;
        float* dp = dcn_off + (int64_t)b * n * 2 * HW + i;
        float* mp = mask + (int64_t)b * n * HW + i;
        for (int j = 0; j < n; ++j) {
            const float cx = __fadd_rn(__fmul_rn(2.0f, tanhf(__ldg(rp + (2 * j) * sr.c))), fx);
            const float cy = __fadd_rn(__fmul_rn(2.0f, tanhf(__ldg(rp + (2 * j + 1) * sr.c))), fy);
            st_stream(cp + (int64_t)(2 * j) * HW, cx);
            st_stream(cp + (int64_t)(2 * j + 1) * HW, cy);
            st_stream(dp + (int64_t)j * HW, cy);
            st_stream(dp + (int64_t)(n + j) * HW, cx);
            const float m = __ldg(rp + (int64_t)(2 * n + j) * sr.c);
            st_stream(mp + (int64_t)j * HW, __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-m))));
        }
    }
}

}  // namespace vfi

extern "C" int vfi_dcndat_coords_f32(const float* raw, const int64_t raw_strides[4], const float* ft, const int64_t ft_strides[4],
                                     float* coords, float* dcn_offset, float* mask, int B, int n, int H, int W, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;
    VFI_REQUIRE(raw && raw_strides && ft && ft_strides && coords && dcn_offset && mask, "vfi_dcndat_coords_f32: null pointer");
    VFI_REQUIRE(B >= 1 && B <= 65535 && n >= 1 && H >= 1 && W >= 1, "vfi_dcndat_coords_f32: bad shape");
    const Str4g sr{raw_strides[0], raw_strides[1], raw_strides[2], raw_strides[3]};
    const Str4g sf{ft_strides[0], ft_strides[1], ft_strides[2], ft_strides[3]};
    const int64_t HW = (int64_t)H * W;
    const dim3 grid((unsigned)min64(ceil_div(HW, 256), (int64_t)kNumSMs * 8), (unsigned)B);
    dcndat_coords_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(raw, sr, ft, sf, coords, dcn_offset, mask, n, H, W);
    VFI_CHECK_LAUNCH("dcndat_coords_kernel");
    return VFI_OK;
}
