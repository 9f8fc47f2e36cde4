// Fused conv epilogue for the cuDNN conv stacks that feed / consume the hot path:
//   t = y + bias[c];  [t = res + t];  [t = t > 0 ? t : slope[c] * t];  [t = res + t]
// i.e. the bias add, the per-channel PReLU (modules/basic_blocks.py, modules/cnn_encoders.py) and the residual add
// that PyTorch issues as three generic strided elementwise kernels, in one vectorised pass (same fp32 operation
// order as the separate ops, so results are bit-identical to them).  NCHW and channels-last.
#include "vfi_common.cuh"

namespace vfi {

template <bool CL>
__global__ void __launch_bounds__(256) bias_act_kernel(const float* __restrict__ y, const float* __restrict__ bias,
                                                       const float* __restrict__ slope, const float* __restrict__ res,
                                                       float* __restrict__ out, int C, int64_t HW, int64_t total4, int res_mode) {
    // one thread = 4 consecutive elements: CL -> 4 channels of one pixel (C % 4 == 0); planar -> 4 pixels of one plane (HW % 4 == 0)
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t e = i * 4;
        float4 v = __ldg(reinterpret_cast<const float4*>(y) + i);
        float b[4], s[4];
        if (CL) {
            const int c = (int)(e % C);
            const float4 bb = bias ? __ldg(reinterpret_cast<const float4*>(bias + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
            b[0] = bb.x; b[1] = bb.y; b[2] = bb.z; b[3] = bb.w;
            if (slope) {
                const float4 ss = __ldg(reinterpret_cast<const float4*>(slope + c));
                s[0] = ss.x; s[1] = ss.y; s[2] = ss.z; s[3] = ss.w;
            }
        } else {
            const int c = (int)((e / HW) % C);
            b[0] = b[1] = b[2] = b[3] = bias ? __ldg(bias + c) : 0.0f;
            if (slope) s[0] = s[1] = s[2] = s[3] = __ldg(slope + c);
        }
        float t[4] = {v.x, v.y, v.z, v.w};
        float r[4] = {0.f, 0.f, 0.f, 0.f};
        if (res_mode) {
            const float4 rr = __ldg(reinterpret_cast<const float4*>(res) + i);
            r[0] = rr.x; r[1] = rr.y; r[2] = rr.z; r[3] = rr.w;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (bias) t[k] = __fadd_rn(t[k], b[k]);
            if (res_mode == 1) t[k] = __fadd_rn(r[k], t[k]);
            if (slope) t[k] = t[k] > 0.0f ? t[k] : __fmul_rn(s[k], t[k]);
            if (res_mode == 2) t[k] = __fadd_rn(r[k], t[k]);
        }
        reinterpret_cast<float4*>(out)[i] = make_float4(t[0], t[1], t[2], t[3]);
    }
}

__global__ void __launch_bounds__(256) bias_act_scalar_kernel(const float* __restrict__ y, const float* __restrict__ bias,
                                                              const float* __restrict__ slope, const float* __restrict__ res,
                                                              float* __restrict__ out, int C, int64_t HW, int64_t total, int cl,
                                                              int res_mode) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int c = cl ? (int)(e % C) : (int)((e / HW) % C);
        float t = y[e];
        if (bias) t = __fadd_rn(t, __ldg(bias + c));
        if (res_mode == 1) t = __fadd_rn(res[e], t);
        if (slope) t = t > 0.0f ? t : __fmul_rn(__ldg(slope + c), t);
        if (res_mode == 2) t = __fadd_rn(res[e], t);
        out[e] = t;
    }
}

// channels-last (pixel stride `ps` floats) -> planar NCHW through a shared-memory tile: the gather kernels (dattn.cu,
// dsample.cu) read planar features, the conv stacks produce channels-last ones.  Reads are contiguous pixel rows,
// writes are P consecutive pixels of one plane; the +1 row padding keeps the transposed reads conflict-free.
template <bool VEC4>
__global__ void __launch_bounds__(256) nhwc_to_nchw_kernel(const float* __restrict__ x, float* __restrict__ out, int C,
                                                           int64_t HW, int64_t ps, int P) {
    extern __shared__ float tr_tile[];      // [P][C + 1]
    const int b = blockIdx.y, ld = C + 1;
    const int64_t p0 = (int64_t)blockIdx.x * P;
    const int np = HW - p0 < (int64_t)P ? (int)(HW - p0) : P;
    const float* src = x + ((int64_t)b * HW + p0) * ps;
    if (VEC4) {
        const int c4 = C / 4;
        for (int i = threadIdx.x; i < np * c4; i += 256) {
            const int px = i / c4, c = (i - px * c4) * 4;
            const float4 v = __ldg(reinterpret_cast<const float4*>(src + (int64_t)px * ps + c));
            float* t = tr_tile + px * ld + c;
            t[0] = v.x; t[1] = v.y; t[2] = v.z; t[3] = v.w;
        }
    } else {
        for (int i = threadIdx.x; i < np * C; i += 256) {
            const int px = i / C, c = i - px * C;
            tr_tile[px * ld + c] = __ldg(src + (int64_t)px * ps + c);
        }
    }
    __syncthreads();
    float* dst = out + (int64_t)b * C * HW + p0;
    for (int i = threadIdx.x; i < C * P; i += 256) {
        const int c = i / P, px = i - c * P;
        if (px < np) dst[(int64_t)c * HW + px] = tr_tile[px * ld + c];
    }
}

}  // namespace vfi

extern "C" int vfi_bias_act_f32(const float* y, const float* bias, const float* slope, const float* residual, float* out, int B,
                                int C, int H, int W, int channels_last, int residual_mode, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(y && out, "vfi_bias_act_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1, "vfi_bias_act_f32: bad shape");
    VFI_REQUIRE(residual_mode >= 0 && residual_mode <= 2, "vfi_bias_act_f32: residual_mode %d", residual_mode);
    VFI_REQUIRE(residual_mode == 0 || residual != nullptr, "vfi_bias_act_f32: residual_mode %d needs a residual", residual_mode);
    if (B == 0) return VFI_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t HW = (int64_t)H * W, total = (int64_t)B * C * HW;
    auto al16 = [](const void* p) { return p == nullptr || (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    const bool vec = al16(y) && al16(out) && al16(residual) && (channels_last ? (C % 4 == 0 && al16(bias) && al16(slope)) : (HW % 4 == 0));
    if (vec) {
        const int64_t total4 = total / 4;
        const int blocks = (int)min64(ceil_div(total4, 256), (int64_t)kNumSMs * 16);
        if (channels_last) bias_act_kernel<true><<<blocks, 256, 0, st>>>(y, bias, slope, residual, out, C, HW, total4, residual_mode);
        else bias_act_kernel<false><<<blocks, 256, 0, st>>>(y, bias, slope, residual, out, C, HW, total4, residual_mode);
        VFI_CHECK_LAUNCH("bias_act_kernel");
    } else {
        const int blocks = (int)min64(ceil_div(total, 256), (int64_t)kNumSMs * 16);
        bias_act_scalar_kernel<<<blocks, 256, 0, st>>>(y, bias, slope, residual, out, C, HW, total, channels_last, residual_mode);
        VFI_CHECK_LAUNCH("bias_act_scalar_kernel");
    }
    return VFI_OK;
}

extern "C" int vfi_nhwc_to_nchw_f32(const float* x, float* out, int B, int C, int H, int W, int64_t x_pix_stride, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(x && out, "vfi_nhwc_to_nchw_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1 && x_pix_stride >= C, "vfi_nhwc_to_nchw_f32: bad shape");
    VFI_REQUIRE(B <= 65535, "vfi_nhwc_to_nchw_f32: batch %d too large", B);
    if (B == 0) return VFI_OK;
    const int64_t HW = (int64_t)H * W;
    int P = 128;
    while (P > 32 && (size_t)(C + 1) * P * 4 > 48 * 1024) P /= 2;
    const size_t smem = (size_t)(C + 1) * P * 4;
    VFI_REQUIRE(smem <= 48 * 1024, "vfi_nhwc_to_nchw_f32: %d channels do not fit a shared-memory tile", C);
    const dim3 grid((unsigned)ceil_div(HW, (int64_t)P), B);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool vec = C % 4 == 0 && x_pix_stride % 4 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
    if (vec) nhwc_to_nchw_kernel<true><<<grid, 256, smem, st>>>(x, out, C, HW, x_pix_stride, P);
    else nhwc_to_nchw_kernel<false><<<grid, 256, smem, st>>>(x, out, C, HW, x_pix_stride, P);
    VFI_CHECK_LAUNCH("nhwc_to_nchw_kernel");
    return VFI_OK;
}
