/* CPU oracle, TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
 *
 * Plain-C restatement of the reference's summation-splat kernel, which exists only as a CUDA string
 * (reference: modules/softsplat.py:289-341).  Same fp32 expression order as the kernel: target =
 * (float)x + flow; skip when not finite (306-307); corner weights from int->float differences
 * (320-323); each corner bounds-tested on its own (325-339).  Serial, so the sum order is fixed
 * (source raster order), unlike the reference's atomics.
 *
 * Build: python oracle/build_oracle.py   ->  oracle/_build/libsplat_oracle.so
 */
#include <math.h>
#include <stdint.h>

static inline void deposit(float* plane, int64_t H, int64_t W, int64_t x, int64_t y, float v) {
    if (x >= 0 && x < W && y >= 0 && y < H) plane[y * W + x] += v;
}

/* in (N,C,H,W), flow (N,2,H,W), out (N,C,H,W) pre-zeroed; all contiguous fp32. */
void oracle_splat_sum_f32(const float* in, const float* flow, float* out,
                          int64_t N, int64_t C, int64_t H, int64_t W) {
    const int64_t HW = H * W;
    for (int64_t n = 0; n < N; ++n) {
        const float* fx_plane = flow + (n * 2 + 0) * HW;
        const float* fy_plane = flow + (n * 2 + 1) * HW;
        for (int64_t y = 0; y < H; ++y) {
            for (int64_t x = 0; x < W; ++x) {
                const float fx = (float)x + fx_plane[y * W + x];
                const float fy = (float)y + fy_plane[y * W + x];
                if (!isfinite(fx) || !isfinite(fy)) continue;
                if (fx < -2.0f || fy < -2.0f || fx > (float)W + 1.0f || fy > (float)H + 1.0f) continue; /* no corner can land */
                const int64_t x0 = (int64_t)floorf(fx), y0 = (int64_t)floorf(fy);
                const int64_t x1 = x0 + 1, y1 = y0 + 1;
                const float w_nw = ((float)x1 - fx) * ((float)y1 - fy);
                const float w_ne = (fx - (float)x0) * ((float)y1 - fy);
                const float w_sw = ((float)x1 - fx) * (fy - (float)y0);
                const float w_se = (fx - (float)x0) * (fy - (float)y0);
                for (int64_t c = 0; c < C; ++c) {
                    const float v = in[(n * C + c) * HW + y * W + x];
                    float* plane = out + (n * C + c) * HW;
                    deposit(plane, H, W, x0, y0, v * w_nw);
                    deposit(plane, H, W, x1, y0, v * w_ne);
                    deposit(plane, H, W, x0, y1, v * w_sw);
                    deposit(plane, H, W, x1, y1, v * w_se);
                }
            }
        }
    }
}
