// Shared pieces of the DCNv2 kernels (SIMT and tcgen05 paths): parameters and torchvision's bilinear tap.
#pragma once
#include "vfi_common.cuh"

namespace vfi {

struct DcnParams {
    int Cin, Cout, H, W, Ho, Wo, kh, kw, stride, pad, dil, WG, OG;
};

struct DcnTap {
    int64_t o;  // offset of (h_low, w_low) in a plane
    int hl, wl; // h_low, w_low
    float w1, w2, w3, w4;
    bool k1, k2, k3, k4;
};

// torchvision bilinear_interpolate: 0 if h <= -1 || h >= H || w <= -1 || w >= W, else 4 corners, each
// dropped when outside.
__device__ __forceinline__ DcnTap dcn_tap(float h, float w, int H, int W) {
    DcnTap t;
    const bool inside = h > -1.0f && h < (float)H && w > -1.0f && w < (float)W;
    const float hc = inside ? h : 0.0f, wc = inside ? w : 0.0f;
    const float hl = floorf(hc), wl = floorf(wc);
    const int h_low = (int)hl, w_low = (int)wl;
    const float lh = hc - hl, lw = wc - wl;
    const float hh = 1.0f - lh, hw = 1.0f - lw;
    t.k1 = inside && h_low >= 0 && w_low >= 0;
    t.k2 = inside && h_low >= 0 && w_low + 1 <= W - 1;
    t.k3 = inside && h_low + 1 <= H - 1 && w_low >= 0;
    t.k4 = inside && h_low + 1 <= H - 1 && w_low + 1 <= W - 1;
    t.w1 = hh * hw;
    t.w2 = hh * lw;
    t.w3 = lh * hw;
    t.w4 = lh * lw;
    t.o = (int64_t)h_low * W + w_low;
    t.hl = h_low;
    t.wl = w_low;
    return t;
}

__device__ __forceinline__ float dcn_sample(const float* __restrict__ pl, const DcnTap& t, int W) {
    float v = 0.0f;
    if (t.k1) v = t.w1 * __ldg(pl);
    if (t.k2) v += t.w2 * __ldg(pl + 1);
    if (t.k3) v += t.w3 * __ldg(pl + W);
    if (t.k4) v += t.w4 * __ldg(pl + W + 1);
    return v;
}

}  // namespace vfi
