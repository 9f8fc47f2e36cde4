// On-device validators: calculate_psnr / calculate_ssim of evaluate.py:15-84 in one pass.
// The reference's SSIM is an 11x11x11 Gaussian window (sigma 1.5) slid over the (channel, y, x) volume of each sample
// with replicate padding of 5 on every side -- five conv3d calls over padded copies plus ~15 elementwise passes.
// The window is separable, so here:
//   channel direction: replicate padding folds the 11 taps onto the C real channels -> a C-term mix per output channel
//   y / x directions:  a 42x42 halo tile of the five mixed fields (a, b, a*a, b*b, a*b) in shared memory, a
//                      horizontal and a vertical 11-tap pass, the SSIM map value and the block's partial sums
// One CTA = one 32x32 tile of one (sample, channel); partial sums (squared error, SSIM, contrast-structure) go to the
// workspace and a second kernel adds them in a fixed order, so the result is reproducible bit for bit.
#include <math.h>

#include "vfi_common.cuh"

namespace vfi {

constexpr int kWin = 11, kRad = 5, kMT = 32, kMH = kMT + 2 * kRad, kMHP = kMH + 1;   // tile, halo, padded row
constexpr int kMaxMetricC = 32;

struct Gauss11 {
    float g[kWin];
};

struct MetricsArgs {
    const float *a, *b;
    int C, H, W, tiles_x, tiles_y;
    float c1, c2;
    Gauss11 win;
    double* partial;   // [B][C][tiles][3]
};

__global__ void __launch_bounds__(256) ssim_psnr_kernel(MetricsArgs p) {
    extern __shared__ __align__(16) float sm[];
    float* f = sm;                              // [5][kMH][kMHP]  mixed fields over the halo tile
    float* h = f + 5 * kMH * kMHP;              // [5][kMH][kMT]   after the horizontal pass
    __shared__ float wc[kMaxMetricC];
    __shared__ double red[3][8];
    const int c = blockIdx.y, b = blockIdx.z;
    const int ty = blockIdx.x / p.tiles_x, tx = blockIdx.x - ty * p.tiles_x;
    const int y0 = ty * kMT, x0 = tx * kMT;
    const int64_t HW = (int64_t)p.H * p.W;
    // channel mix of output channel c: tap d reads channel clamp(c + d - 5)
    if (threadIdx.x < p.C) {
        float w = 0.0f;
        for (int d = 0; d < kWin; ++d)
            if (min(max(c + d - kRad, 0), p.C - 1) == (int)threadIdx.x) w += p.win.g[d];
        wc[threadIdx.x] = w;
    }
    __syncthreads();
    const float* pa = p.a + (int64_t)b * p.C * HW;
    const float* pb = p.b + (int64_t)b * p.C * HW;
    for (int i = threadIdx.x; i < kMH * kMH; i += blockDim.x) {
        const int yy = i / kMH, xx = i - yy * kMH;
        const int y = min(max(y0 + yy - kRad, 0), p.H - 1), x = min(max(x0 + xx - kRad, 0), p.W - 1);
        const int64_t o = (int64_t)y * p.W + x;
        float m1 = 0.0f, m2 = 0.0f, m11 = 0.0f, m22 = 0.0f, m12 = 0.0f;
        for (int k = 0; k < p.C; ++k) {
            const float w = wc[k];
            if (w == 0.0f) continue;
            const float va = __ldg(pa + k * HW + o), vb = __ldg(pb + k * HW + o);
            m1 = fmaf(w, va, m1);
            m2 = fmaf(w, vb, m2);
            m11 = fmaf(w, va * va, m11);
            m22 = fmaf(w, vb * vb, m22);
            m12 = fmaf(w, va * vb, m12);
        }
        const int s = yy * kMHP + xx;
        f[s] = m1;
        f[kMH * kMHP + s] = m2;
        f[2 * kMH * kMHP + s] = m11;
        f[3 * kMH * kMHP + s] = m22;
        f[4 * kMH * kMHP + s] = m12;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 5 * kMH * kMT; i += blockDim.x) {
        const int x = i % kMT, yy = (i / kMT) % kMH, fld = i / (kMT * kMH);
        const float* row = f + (fld * kMH + yy) * kMHP + x;
        float acc = 0.0f;
#pragma unroll
        for (int j = 0; j < kWin; ++j) acc = fmaf(p.win.g[j], row[j], acc);
        h[i] = acc;
    }
    __syncthreads();
    double sse = 0.0, ssim = 0.0, cs = 0.0;
    for (int i = threadIdx.x; i < kMT * kMT; i += blockDim.x) {
        const int yy = i / kMT, x = i - yy * kMT;
        if (y0 + yy >= p.H || x0 + x >= p.W) continue;
        float v[5];
#pragma unroll
        for (int fld = 0; fld < 5; ++fld) {
            const float* col = h + (fld * kMH + yy) * kMT + x;
            float acc = 0.0f;
#pragma unroll
            for (int j = 0; j < kWin; ++j) acc = fmaf(p.win.g[j], col[j * kMT], acc);
            v[fld] = acc;
        }
        // evaluate.py:53-70
        const float mu1_sq = v[0] * v[0], mu2_sq = v[1] * v[1], mu1_mu2 = v[0] * v[1];
        const float s1 = v[2] - mu1_sq, s2 = v[3] - mu2_sq, s12 = v[4] - mu1_mu2;
        const float v1 = 2.0f * s12 + p.c2, v2 = s1 + s2 + p.c2;
        cs += (double)(v1 / v2);
        ssim += (double)(((2.0f * mu1_mu2 + p.c1) * v1) / ((mu1_sq + mu2_sq + p.c1) * v2));
        const int64_t o = (int64_t)c * HW + (int64_t)(y0 + yy) * p.W + (x0 + x);
        const float d = __ldg(pa + o) - __ldg(pb + o);
        sse += (double)(d * d);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        sse += __shfl_xor_sync(0xffffffffu, sse, d);
        ssim += __shfl_xor_sync(0xffffffffu, ssim, d);
        cs += __shfl_xor_sync(0xffffffffu, cs, d);
    }
    if ((threadIdx.x & 31) == 0) {
        red[0][threadIdx.x >> 5] = sse;
        red[1][threadIdx.x >> 5] = ssim;
        red[2][threadIdx.x >> 5] = cs;
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += red[threadIdx.x][w];
        const int64_t blk = ((int64_t)b * p.C + c) * gridDim.x + blockIdx.x;
        p.partial[blk * 3 + threadIdx.x] = s;
    }
}

// out[b][k] = sum of the sample's partials in index order (fixed tree inside the block)
__global__ void __launch_bounds__(256) metrics_reduce_kernel(const double* __restrict__ partial, double* __restrict__ out, int64_t per_sample) {
    __shared__ double red[3][256];
    const int b = blockIdx.x;
    double s[3] = {0.0, 0.0, 0.0};
    for (int64_t i = threadIdx.x; i < per_sample; i += blockDim.x)
        for (int k = 0; k < 3; ++k) s[k] += partial[((int64_t)b * per_sample + i) * 3 + k];
    for (int k = 0; k < 3; ++k) red[k][threadIdx.x] = s[k];
    __syncthreads();
    for (int d = 128; d > 0; d >>= 1) {
        if ((int)threadIdx.x < d)
            for (int k = 0; k < 3; ++k) red[k][threadIdx.x] += red[k][threadIdx.x + d];
        __syncthreads();
    }
    if (threadIdx.x < 3) out[b * 3 + threadIdx.x] = red[threadIdx.x][0];
}

static int64_t metric_tiles(int H, int W) { return ceil_div(H, kMT) * ceil_div(W, kMT); }

}  // namespace vfi

extern "C" size_t vfi_frame_metrics_workspace_bytes(int B, int C, int H, int W) {
    if (B <= 0 || C <= 0 || H <= 0 || W <= 0) return 0;
    return (size_t)B * C * vfi::metric_tiles(H, W) * 3 * sizeof(double);
}

extern "C" int vfi_frame_metrics_f32(const float* img1, const float* img2, double* out, int B, int C, int H, int W, float val_range,
                                     void* workspace, size_t workspace_bytes, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;
    VFI_REQUIRE(img1 && img2 && out, "vfi_frame_metrics_f32: null pointer");
    VFI_REQUIRE(B >= 1 && B <= 65535 && C >= 1 && C <= kMaxMetricC, "vfi_frame_metrics_f32: need 1 <= B <= 65535, 1 <= C <= %d", kMaxMetricC);
    // evaluate.py:45-46 shrinks the window on frames smaller than 11 while still padding by 5: not reproduced
    if (H < kWin || W < kWin) VFI_UNSUPPORTED("vfi_frame_metrics_f32: frames smaller than the 11-tap window (H=%d W=%d)", H, W);
    VFI_REQUIRE(workspace && workspace_bytes >= vfi_frame_metrics_workspace_bytes(B, C, H, W) && (reinterpret_cast<uintptr_t>(workspace) & 7) == 0,
                "vfi_frame_metrics_f32: workspace of %zu B (8-byte aligned) needed", vfi_frame_metrics_workspace_bytes(B, C, H, W));
    MetricsArgs p;
    p.a = img1; p.b = img2; p.C = C; p.H = H; p.W = W;
    p.tiles_x = (int)ceil_div(W, kMT); p.tiles_y = (int)ceil_div(H, kMT);
    p.c1 = (0.01f * val_range) * (0.01f * val_range);
    p.c2 = (0.03f * val_range) * (0.03f * val_range);
    // gaussian(11, 1.5), evaluate.py:15-17: exp in double, stored as fp32, normalised in fp32
    float sum = 0.0f;
    for (int i = 0; i < kWin; ++i) {
        p.win.g[i] = (float)exp(-(double)((i - kWin / 2) * (i - kWin / 2)) / (2.0 * 1.5 * 1.5));
        sum += p.win.g[i];
    }
    for (int i = 0; i < kWin; ++i) p.win.g[i] /= sum;
    p.partial = static_cast<double*>(workspace);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t smem = (size_t)(5 * kMH * kMHP + 5 * kMH * kMT) * sizeof(float);
    VFI_CUDA(cudaFuncSetAttribute(ssim_psnr_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t tiles = metric_tiles(H, W);
    VFI_REQUIRE(tiles < ((int64_t)1 << 31), "vfi_frame_metrics_f32: frame too large");
    ssim_psnr_kernel<<<dim3((unsigned)tiles, (unsigned)C, (unsigned)B), 256, smem, st>>>(p);
    VFI_CHECK_LAUNCH("ssim_psnr_kernel");
    metrics_reduce_kernel<<<B, 256, 0, st>>>(p.partial, out, (int64_t)C * tiles);
    VFI_CHECK_LAUNCH("metrics_reduce_kernel");
    return VFI_OK;
}

// ---- allocation-free variant -------------------------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot allocate device
// memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The partial sums that
// vfi_frame_metrics_f32 keeps in a caller-supplied workspace live in a file-scope device array instead: 1 MiB, held by
// every GPU that loads the library until the process exits.  Every call on a device shares that array, so calls that
// overlap on the GPU (different streams, different host threads) race on it; calls on one stream are ordered and safe.
// Nothing in it has to start at zero: ssim_psnr_kernel overwrites every partial that metrics_reduce_kernel reads, so the
// previous call's contents are never seen.
namespace vfi {

constexpr size_t kStaticMetricsBytes = (size_t)1 << 20;
static __device__ __align__(256) unsigned char g_metrics_scratch[kStaticMetricsBytes];

// In host code `g_metrics_scratch` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
static int static_metrics_scratch(void** out) {
    constexpr int kMaxDev = 64;
    static void* ptrs[kMaxDev] = {};
    static bool have[kMaxDev] = {};
    int dev = -1;
    VFI_CUDA(cudaGetDevice(&dev));
    VFI_REQUIRE(dev >= 0 && dev < kMaxDev, "vfi_frame_metrics_static_f32: device ordinal %d out of range", dev);
    // benign race: two host threads may both query the same device the first time and store the same address
    if (!have[dev]) {
        void* p = nullptr;
        VFI_CUDA(cudaGetSymbolAddress(&p, g_metrics_scratch));
        VFI_REQUIRE(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0, "vfi_frame_metrics_static_f32: bad scratch address");
        ptrs[dev] = p;
        have[dev] = true;
    }
    *out = ptrs[dev];
    return VFI_OK;
}

}  // namespace vfi

extern "C" size_t vfi_frame_metrics_static_max_bytes(void) { return vfi::kStaticMetricsBytes; }

extern "C" int vfi_frame_metrics_static_f32(const float* img1, const float* img2, double* out, int B, int C, int H, int W,
                                            float val_range, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;
    VFI_REQUIRE(B >= 1 && C >= 1 && H >= 1 && W >= 1, "vfi_frame_metrics_static_f32: bad shape");
    const size_t need = vfi_frame_metrics_workspace_bytes(B, C, H, W);
    if (need > kStaticMetricsBytes)
        VFI_UNSUPPORTED("vfi_frame_metrics_static_f32: B=%d C=%d H=%d W=%d needs %zu B of scratch, the built-in array holds %zu B "
                        "(use vfi_frame_metrics_f32 with a caller-supplied workspace)", B, C, H, W, need, kStaticMetricsBytes);
    void* scratch = nullptr;
    const int rc = static_metrics_scratch(&scratch);
    if (rc) return rc;
    return vfi_frame_metrics_f32(img1, img2, out, B, C, H, W, val_range, scratch, kStaticMetricsBytes, stream);
}
