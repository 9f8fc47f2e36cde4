// Frame I/O edge of the forward: what the reference does around model.inference with five to ten
// full-frame PyTorch passes -- uint8 -> float / 255 (evaluate.py:138-139), InputPadder replicate pad to a multiple of
// 16 (data/__init__.py:6-20), norm_w_rgb_mean (models/BaseModel.py:14-17: one mean over both padded frames per sample),
// and at the other end clamp(out + mean, 0, 1) (modules/generator.py:24) + unpad (+ the uint8 quantisation of a
// written frame) -- as two reductions-free passes plus one sum:
//   vfi_frame_prep    sum kernel (fp32 partials, one fp64 atomic per CTA) -> mean[b];
//                     normalise kernel: padded, mean-subtracted frames written channels-last with 4 channels (RGB0),
//                     the layout the encoder's first tensor-core conv reads (no separate channel-pad copy)
//   vfi_frame_finish  clamp(y + mean, 0, 1), crop, optional round-to-uint8
#include "vfi_common.cuh"

namespace vfi {

template <typename T>
__device__ __forceinline__ float load_unit(const T* p);
template <>
__device__ __forceinline__ float load_unit<float>(const float* p) { return __ldg(p); }
template <>
__device__ __forceinline__ float load_unit<uint8_t>(const uint8_t* p) { return __fdiv_rn((float)__ldg(p), 255.0f); }

struct FramePad {
    int H, W, Hp, Wp, top, left;
};

// sum over the PADDED frames (replicate padding repeats border pixels, and the reference's mean sees them)
template <typename T>
__global__ void __launch_bounds__(256) frame_sum_kernel(const T* __restrict__ x0, const T* __restrict__ x1, double* __restrict__ sums,
                                                        FramePad g) {
    const int b = blockIdx.y;
    const int64_t HW = (int64_t)g.H * g.W, n = (int64_t)g.Hp * g.Wp;
    float acc = 0.0f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int yp = (int)(i / g.Wp), xp = (int)(i - (int64_t)yp * g.Wp);
        const int y = min(max(yp - g.top, 0), g.H - 1), x = min(max(xp - g.left, 0), g.W - 1);
        const int64_t o = (int64_t)b * 3 * HW + (int64_t)y * g.W + x;
#pragma unroll
        for (int c = 0; c < 3; ++c) acc += load_unit<T>(x0 + o + c * HW) + load_unit<T>(x1 + o + c * HW);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    __shared__ float part[8];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += (double)part[w];
        atomicAdd(sums + b, s);
    }
}

template <typename T>
__global__ void __launch_bounds__(256) frame_norm_kernel(const T* __restrict__ x0, const T* __restrict__ x1, const double* __restrict__ sums,
                                                         float* __restrict__ out0, float* __restrict__ out1, float* __restrict__ mean_out,
                                                         FramePad g) {
    const int b = blockIdx.y;
    const int64_t HW = (int64_t)g.H * g.W, n = (int64_t)g.Hp * g.Wp;
    const float mean = (float)(sums[b] / (6.0 * (double)n));
    if (blockIdx.x == 0 && threadIdx.x == 0) mean_out[b] = mean;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int yp = (int)(i / g.Wp), xp = (int)(i - (int64_t)yp * g.Wp);
        const int y = min(max(yp - g.top, 0), g.H - 1), x = min(max(xp - g.left, 0), g.W - 1);
        const int64_t o = (int64_t)b * 3 * HW + (int64_t)y * g.W + x;
        float4 a, c;
        a.x = load_unit<T>(x0 + o) - mean; a.y = load_unit<T>(x0 + o + HW) - mean; a.z = load_unit<T>(x0 + o + 2 * HW) - mean; a.w = 0.0f;
        c.x = load_unit<T>(x1 + o) - mean; c.y = load_unit<T>(x1 + o + HW) - mean; c.z = load_unit<T>(x1 + o + 2 * HW) - mean; c.w = 0.0f;
        reinterpret_cast<float4*>(out0)[(int64_t)b * n + i] = a;
        reinterpret_cast<float4*>(out1)[(int64_t)b * n + i] = c;
    }
}

struct Str4f {
    int64_t n, c, h, w;
};

template <typename T>
__global__ void __launch_bounds__(256) frame_finish_kernel(const float* __restrict__ y, Str4f sy, const float* __restrict__ mean,
                                                           T* __restrict__ out, FramePad g) {
    const int b = blockIdx.y;
    const int64_t HW = (int64_t)g.H * g.W;
    const float m = __ldg(mean + b);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < HW; i += (int64_t)gridDim.x * blockDim.x) {
        const int yy = (int)(i / g.W), xx = (int)(i - (int64_t)yy * g.W);
        const float* p = y + b * sy.n + (int64_t)(yy + g.top) * sy.h + (int64_t)(xx + g.left) * sy.w;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float v = fminf(fmaxf(__fadd_rn(__ldg(p + c * sy.c), m), 0.0f), 1.0f);
            if (sizeof(T) == 1) out[((int64_t)b * 3 + c) * HW + i] = (T)__float2int_rn(v * 255.0f);
            else out[((int64_t)b * 3 + c) * HW + i] = (T)v;
        }
    }
}

static int frame_geom(FramePad& g, int B, int H, int W, int Hp, int Wp, int top, int left, const char* who) {
    VFI_REQUIRE(B >= 0 && H >= 1 && W >= 1 && Hp >= H && Wp >= W && top >= 0 && left >= 0 && top + H <= Hp && left + W <= Wp,
                "%s: bad frame geometry", who);
    VFI_REQUIRE(B <= 65535, "%s: batch %d too large", who, B);
    g.H = H; g.W = W; g.Hp = Hp; g.Wp = Wp; g.top = top; g.left = left;
    return VFI_OK;
}

}  // namespace vfi

extern "C" size_t vfi_frame_prep_workspace_bytes(int B) { return B > 0 ? (size_t)B * sizeof(double) : 0; }

extern "C" int vfi_frame_prep(const void* x0, const void* x1, int is_u8, float* out0, float* out1, float* mean, int B, int H, int W,
                              int Hp, int Wp, int pad_top, int pad_left, void* workspace, size_t workspace_bytes, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(x0 && x1 && out0 && out1 && mean, "vfi_frame_prep: null pointer");
    FramePad g;
    if (int rc = frame_geom(g, B, H, W, Hp, Wp, pad_top, pad_left, "vfi_frame_prep")) return rc;
    if (B == 0) return VFI_OK;
    VFI_REQUIRE(workspace && workspace_bytes >= vfi_frame_prep_workspace_bytes(B) && (reinterpret_cast<uintptr_t>(workspace) & 7) == 0,
                "vfi_frame_prep: workspace of %zu B (8-byte aligned) needed", vfi_frame_prep_workspace_bytes(B));
    VFI_REQUIRE((reinterpret_cast<uintptr_t>(out0) & 15) == 0 && (reinterpret_cast<uintptr_t>(out1) & 15) == 0,
                "vfi_frame_prep: outputs must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    double* sums = static_cast<double*>(workspace);
    VFI_CUDA(cudaMemsetAsync(sums, 0, sizeof(double) * B, st));
    const int64_t n = (int64_t)Hp * Wp;
    const dim3 grid((unsigned)min64(ceil_div(n, 256 * 4), (int64_t)kNumSMs * 4), (unsigned)B);
    if (is_u8) {
        frame_sum_kernel<uint8_t><<<grid, 256, 0, st>>>(static_cast<const uint8_t*>(x0), static_cast<const uint8_t*>(x1), sums, g);
        VFI_CHECK_LAUNCH("frame_sum_kernel");
        frame_norm_kernel<uint8_t><<<grid, 256, 0, st>>>(static_cast<const uint8_t*>(x0), static_cast<const uint8_t*>(x1), sums, out0, out1, mean, g);
    } else {
        frame_sum_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(x0), static_cast<const float*>(x1), sums, g);
        VFI_CHECK_LAUNCH("frame_sum_kernel");
        frame_norm_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(x0), static_cast<const float*>(x1), sums, out0, out1, mean, g);
    }
    VFI_CHECK_LAUNCH("frame_norm_kernel");
    return VFI_OK;
}

extern "C" int vfi_frame_finish(const float* y, const int64_t y_strides[4], const float* mean, void* out, int out_is_u8, int B, int H,
                                int W, int Hp, int Wp, int pad_top, int pad_left, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(y && y_strides && mean && out, "vfi_frame_finish: null pointer");
    FramePad g;
    if (int rc = frame_geom(g, B, H, W, Hp, Wp, pad_top, pad_left, "vfi_frame_finish")) return rc;
    if (B == 0) return VFI_OK;
    const Str4f sy{y_strides[0], y_strides[1], y_strides[2], y_strides[3]};
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const dim3 grid((unsigned)min64(ceil_div((int64_t)H * W, 256 * 2), (int64_t)kNumSMs * 8), (unsigned)B);
    if (out_is_u8) frame_finish_kernel<uint8_t><<<grid, 256, 0, st>>>(y, sy, mean, static_cast<uint8_t*>(out), g);
    else frame_finish_kernel<float><<<grid, 256, 0, st>>>(y, sy, mean, static_cast<float*>(out), g);
    VFI_CHECK_LAUNCH("frame_finish_kernel");
    return VFI_OK;
}
