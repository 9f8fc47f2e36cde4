// Forward (summation) splatting with fused pre-scale, and the normalise pass.
// Reference: kernel text modules/softsplat.py:289-341, Python pre/post modules/warp.py:18-114.
//
// Scatter form (this file): one thread per SOURCE pixel; flow, metric and the four corner weights are
// computed once and reused for all C channels (the reference recomputes them per element); the weight
// channel is produced on the fly (no cat / exp / multiply passes).  Adds are fp32 red.global.
// The gather-form owner-tile kernel (no float atomics, fused divide) lives in splat_fused.cu.
#include "vfi_common.cuh"

namespace vfi {

__device__ __forceinline__ float splat_weight(int mode, const float* __restrict__ metric, const float* __restrict__ t,
                                              int b, int64_t pix) {
    switch (mode) {
        case VFI_SPLAT_W_LINEAR: return ld_stream(metric + pix);
        case VFI_SPLAT_W_SOFT: return expf(ld_stream(metric + pix));
        case VFI_SPLAT_W_SOFT_CLIP_T: return t[b] * expf(fminf(fmaxf(ld_stream(metric + pix), -20.0f), 20.0f));
        default: return 1.0f;
    }
}

struct SplatCorners {
    int x0, y0;
    float w_nw, w_ne, w_sw, w_se;
    bool valid;
};

// Same fp32 expressions as the reference kernel (modules/softsplat.py:303-323).
__device__ __forceinline__ SplatCorners splat_corners(int x, int y, float flow_x, float flow_y, int H, int W) {
    SplatCorners c;
    const float fx = (float)x + flow_x;
    const float fy = (float)y + flow_y;
    c.valid = isfinite(fx) && isfinite(fy) && fx > -1.0f && fy > -1.0f && fx < (float)W && fy < (float)H;
    // (outside (-1, W) x (-1, H) no corner lands in the frame; also keeps the int conversion defined)
    const float xf = floorf(c.valid ? fx : 0.0f), yf = floorf(c.valid ? fy : 0.0f);
    c.x0 = (int)xf;
    c.y0 = (int)yf;
    const float x1f = (float)(c.x0 + 1), y1f = (float)(c.y0 + 1);
    c.w_nw = (x1f - fx) * (y1f - fy);
    c.w_ne = (fx - xf) * (y1f - fy);
    c.w_sw = (x1f - fx) * (fy - yf);
    c.w_se = (fx - xf) * (fy - yf);
    return c;
}

__global__ void __launch_bounds__(256) splat_scatter_kernel(const float* __restrict__ in, const float* __restrict__ flow,
                                                            const float* __restrict__ metric, const float* __restrict__ t,
                                                            float* __restrict__ num, float* __restrict__ den, int B, int C,
                                                            int H, int W, int mode) {
    const int64_t HW = (int64_t)H * W;
    const int64_t total = (int64_t)B * HW;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int b = (int)(idx / HW);
        const int64_t pix = idx - (int64_t)b * HW;
        const int y = (int)(pix / W), x = (int)(pix - (int64_t)y * W);
        const float* fp = flow + (int64_t)b * 2 * HW + pix;
        const SplatCorners c = splat_corners(x, y, ld_stream(fp), ld_stream(fp + HW), H, W);
        if (!c.valid) continue;
        const float wm = splat_weight(mode, metric ? metric + (int64_t)b * HW : nullptr, t, b, pix);
        const bool in_x0 = c.x0 >= 0, in_x1 = c.x0 + 1 < W, in_y0 = c.y0 >= 0, in_y1 = c.y0 + 1 < H;
        const int64_t o_nw = (int64_t)c.y0 * W + c.x0;
        const bool k_nw = in_x0 && in_y0, k_ne = in_x1 && in_y0, k_sw = in_x0 && in_y1, k_se = in_x1 && in_y1;
        if (den != nullptr) {
            float* d = den + (int64_t)b * HW + o_nw;
            if (k_nw) atomicAdd(d, wm * c.w_nw);
            if (k_ne) atomicAdd(d + 1, wm * c.w_ne);
            if (k_sw) atomicAdd(d + W, wm * c.w_sw);
            if (k_se) atomicAdd(d + W + 1, wm * c.w_se);
        }
        const float* ip = in + (int64_t)b * C * HW + pix;
        float* op = num + (int64_t)b * C * HW + o_nw;
        for (int ch = 0; ch < C; ++ch) {
            const float v = ld_stream(ip) * wm;
            if (k_nw) atomicAdd(op, v * c.w_nw);
            if (k_ne) atomicAdd(op + 1, v * c.w_ne);
            if (k_sw) atomicAdd(op + W, v * c.w_sw);
            if (k_se) atomicAdd(op + W + 1, v * c.w_se);
            ip += HW;
            op += HW;
        }
    }
}

__device__ __forceinline__ float splat_divisor(float d, int eps_mode, float eps) {
    if (eps_mode == VFI_EPS_ADD) return d + eps;
    if (eps_mode == VFI_EPS_ZERO) return d == 0.0f ? 1.0f : d;
    return fmaxf(d, eps);
}

__global__ void __launch_bounds__(256) splat_normalize_kernel(const float* __restrict__ num, const float* __restrict__ den,
                                                              float* __restrict__ out, uint8_t* __restrict__ blank, int B,
                                                              int C, int64_t HW, int eps_mode, float eps, float blank_thr) {
    const int64_t total = (int64_t)B * HW;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int b = (int)(idx / HW);
        const int64_t pix = idx - (int64_t)b * HW;
        const float d = splat_divisor(den[idx], eps_mode, eps);
        if (blank) blank[idx] = d < blank_thr ? 1 : 0;
        const float* np = num + (int64_t)b * C * HW + pix;
        float* op = out + (int64_t)b * C * HW + pix;
        for (int ch = 0; ch < C; ++ch) op[(int64_t)ch * HW] = __fdiv_rn(np[(int64_t)ch * HW], d);
    }
}

int launch_splat_gather(const float* in, const float* flow, const float* metric, const float* t, float* out, float* den_io, int B, int C,
                        int H, int W, int weight_mode, int normalize, int eps_mode, float eps_value, int acc_mode, void* workspace,
                        cudaStream_t st, uint8_t* blank, float blank_thr);
bool splat_gather_applicable(int B, int C, int H, int W, const void* workspace, size_t workspace_bytes);

}  // namespace vfi

extern "C" int vfi_splat_f32(const float* in, const float* flow, const float* metric, const float* t, float* num,
                             float* den, int B, int C, int H, int W, int weight_mode, int accumulate, void* workspace,
                             size_t workspace_bytes, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(in && flow && num, "vfi_splat_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 0 && H >= 1 && W >= 1, "vfi_splat_f32: bad shape");
    VFI_REQUIRE(weight_mode >= VFI_SPLAT_W_NONE && weight_mode <= VFI_SPLAT_W_SOFT_CLIP_T, "vfi_splat_f32: bad weight_mode %d",
                weight_mode);
    if (weight_mode >= VFI_SPLAT_W_LINEAR) VFI_REQUIRE(metric != nullptr, "vfi_splat_f32: weight mode %d needs a metric", weight_mode);
    if (weight_mode == VFI_SPLAT_W_SOFT_CLIP_T) VFI_REQUIRE(t != nullptr, "vfi_splat_f32: SOFT_CLIP_T needs t");
    if (weight_mode != VFI_SPLAT_W_NONE) VFI_REQUIRE(den != nullptr, "vfi_splat_f32: weight mode %d needs den", weight_mode);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (B == 0) return VFI_OK;
    const int64_t HW = (int64_t)H * W;
    // a workspace selects the gather form (csrc/splat_fused.cu): link pass + one owner thread per target pixel that
    // stores / adds this source's sums -- no floating-point atomics, no zero fill
    if (splat_gather_applicable(B, C, H, W, workspace, workspace_bytes))
        return launch_splat_gather(in, flow, metric, t, num, weight_mode == VFI_SPLAT_W_NONE ? nullptr : den, B, C, H, W, weight_mode, 0,
                                   VFI_EPS_ADD, 0.0f, accumulate ? 1 : 0, workspace, st, nullptr, 0.0f);
    if (!accumulate) {
        if (C > 0) VFI_CUDA(cudaMemsetAsync(num, 0, sizeof(float) * (size_t)B * C * HW, st));
        if (den) VFI_CUDA(cudaMemsetAsync(den, 0, sizeof(float) * (size_t)B * HW, st));
    }
    const int blocks = (int)min64(ceil_div((int64_t)B * HW, 256), (int64_t)kNumSMs * 16);
    splat_scatter_kernel<<<blocks, 256, 0, st>>>(in, flow, metric, t, num, weight_mode == VFI_SPLAT_W_NONE ? nullptr : den, B, C, H,
                                                 W, weight_mode);
    VFI_CHECK_LAUNCH("splat_scatter_kernel");
    return VFI_OK;
}

extern "C" int vfi_splat_normalize_f32(const float* num, const float* den, float* out, uint8_t* blank_mask, int B, int C, int H,
                                       int W, int eps_mode, float eps_value, float blank_threshold, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(num && den && out, "vfi_splat_normalize_f32: null pointer");
    VFI_REQUIRE(eps_mode >= VFI_EPS_ADD && eps_mode <= VFI_EPS_CLIP, "vfi_splat_normalize_f32: bad eps_mode %d", eps_mode);
    if (B == 0) return VFI_OK;
    const int64_t HW = (int64_t)H * W;
    const int blocks = (int)min64(ceil_div((int64_t)B * HW, 256), (int64_t)kNumSMs * 16);
    splat_normalize_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(num, den, out, blank_mask, B, C, HW, eps_mode,
                                                                                  eps_value, blank_threshold);
    VFI_CHECK_LAUNCH("splat_normalize_kernel");
    return VFI_OK;
}
