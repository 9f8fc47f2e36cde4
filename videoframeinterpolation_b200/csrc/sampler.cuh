// Coordinate replay + tap geometry of the reference's grouped sampler (shared by dsample.cu / dattn.cu).
#pragma once
#include "vfi_common.cuh"

namespace vfi {

struct TapGeom {
    int o_nw;       // offset of the north-west tap inside a plane (may be out of range when !k_nw); H*W < 2^31
    float w_nw, w_ne, w_sw, w_se;
    bool k_nw, k_ne, k_sw, k_se;
    float gx, gy;
};

// F.grid_sample(bilinear, padding_mode='zeros', align_corners=True) on the reference's grid:
//   gx = lin_x[x] + (dx / (W-1)) / 2    (sic: modules/deformable_attn.py:94)
//   gy = lin_y[y] + dy / ((H-1)/2)
// The tensor may be a crop (H x W at origin fe.x0, fe.y0) of a frame of fe.Wg x fe.Hg pixels: the grid is the
// FRAME's (lin_x / lin_y are slices of the frame's linspace vectors), so a crop samples exactly where the whole
// frame would; taps outside the crop read as zero (inside the frame they are data the crop does not hold: the
// caller's halo has to cover them).
__device__ __forceinline__ TapGeom sampler_geom(float dx, float dy, float lin_x, float lin_y, int H, int W, const FrameExt& fe) {
    TapGeom g;
    const float wm1 = (float)(fe.Wg - 1), hm1 = (float)(fe.Hg - 1);
    const float half_h = (float)((fe.Hg - 1.0) / 2.0);
    g.gx = __fadd_rn(lin_x, __fmul_rn(__fdiv_rn(dx, wm1), 0.5f));
    g.gy = __fadd_rn(lin_y, __fdiv_rn(dy, half_h));
    float ix = __fmul_rn(__fmul_rn(__fadd_rn(g.gx, 1.0f), 0.5f), wm1);
    float iy = __fmul_rn(__fmul_rn(__fadd_rn(g.gy, 1.0f), 0.5f), hm1);
    // anything beyond one pixel outside has no in-bounds corner; clamp so the int conversion is defined
    const bool finite = isfinite(ix) && isfinite(iy);
    ix = fminf(fmaxf(finite ? ix : -2.0f, -2.0f), (float)fe.Wg + 1.0f);
    iy = fminf(fmaxf(finite ? iy : -2.0f, -2.0f), (float)fe.Hg + 1.0f);
    const float xf = floorf(ix), yf = floorf(iy);
    const float tx = ix - xf, ty = iy - yf;
    const float ex = 1.0f - tx, ey = 1.0f - ty;
    const int x0 = (int)xf - fe.x0, y0 = (int)yf - fe.y0;      // crop coordinates
    const bool in_x0 = x0 >= 0 && x0 < W, in_x1 = x0 + 1 >= 0 && x0 + 1 < W;
    const bool in_y0 = y0 >= 0 && y0 < H, in_y1 = y0 + 1 >= 0 && y0 + 1 < H;
    g.k_nw = in_x0 && in_y0;
    g.k_ne = in_x1 && in_y0;
    g.k_sw = in_x0 && in_y1;
    g.k_se = in_x1 && in_y1;
    g.w_nw = ey * ex;
    g.w_ne = ey * tx;
    g.w_sw = ty * ex;
    g.w_se = ty * tx;
    g.o_nw = y0 * W + x0;
    return g;
}

__device__ __forceinline__ float sample_plane(const float* __restrict__ plane, const TapGeom& g, int W) {
    const float* p = plane + g.o_nw;
    float r = 0.0f;
    if (g.k_nw) r = __ldg(p) * g.w_nw;
    if (g.k_ne) r += __ldg(p + 1) * g.w_ne;
    if (g.k_sw) r += __ldg(p + W) * g.w_sw;
    if (g.k_se) r += __ldg(p + W + 1) * g.w_se;
    return r;
}

// Same taps, same arithmetic, for a loop over the planes of one group: the caller forms the byte pointers of the
// north / south tap rows of plane 0 once per sample; plane cc is one 32x32->64-bit multiply-add away (IMAD.WIDE).
__device__ __forceinline__ float sample_plane_rows(const char* __restrict__ north0, const char* __restrict__ south0,
                                                   int cc, int plane_bytes_n, int plane_bytes_s, const TapGeom& g) {
    const float *pn, *ps;
    asm("mad.wide.s32 %0, %1, %2, %3;" : "=l"(pn) : "r"(plane_bytes_n), "r"(cc), "l"(north0));
    asm("mad.wide.s32 %0, %1, %2, %3;" : "=l"(ps) : "r"(plane_bytes_s), "r"(cc), "l"(south0));
    float r = 0.0f;
    if (g.k_nw) r = __ldg(pn) * g.w_nw;
    if (g.k_ne) r += __ldg(pn + 1) * g.w_ne;
    if (g.k_sw) r += __ldg(ps) * g.w_sw;
    if (g.k_se) r += __ldg(ps + 1) * g.w_se;
    return r;
}

}  // namespace vfi
