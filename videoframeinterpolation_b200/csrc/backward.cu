// Backward passes of the two warps: what train.py reaches through autograd.
//   vfi_splat_bwd_f32  -- gradients of the summation splat wrt its input and its flow: the two kernels of
//                         softsplat_func.backward (modules/softsplat.py:364-533) in one pass per source pixel;
//   vfi_bwarp_bwd_f32  -- gradients of bwarp (modules/warp.py:7-15) wrt image and flow, i.e. what autograd derives
//                         for F.grid_sample(bilinear, padding_mode='border', align_corners=True) composed with the
//                         flow -> grid arithmetic (d grid / d flow * d pixel / d grid == 1).
// Planar NCHW fp32, contiguous; lanes = adjacent pixels (coalesced grad reads, neighbouring gathers).
#include "vfi_common.cuh"

namespace vfi {

// ---- summation splat backward -----------------------------------------------------------------------------------
// forward: out[c, q_k] += in[c, p] * w_k(p) for the 4 pixels q_k around p + flow(p)
//   grad_in[c, p]  = sum_k w_k * g[c, q_k]
//   grad_flow_x[p] = sum_c in[c, p] * sum_k dw_k/dfx * g[c, q_k]      (same for y)
__global__ void __launch_bounds__(256) splat_bwd_kernel(const float* __restrict__ in, const float* __restrict__ flow,
                                                        const float* __restrict__ gout, float* __restrict__ gin,
                                                        float* __restrict__ gflow, int C, int H, int W) {
    const int x = blockIdx.x * 32 + threadIdx.x;
    const int y = blockIdx.y * 8 + threadIdx.y;
    const int b = blockIdx.z;
    if (x >= W || y >= H) return;
    const int64_t HW = (int64_t)H * W, p = (int64_t)y * W + x;
    const float fx = (float)x + flow[((int64_t)b * 2 + 0) * HW + p];
    const float fy = (float)y + flow[((int64_t)b * 2 + 1) * HW + p];
    const bool fin = isfinite(fx) && isfinite(fy);      // the forward kernel skips such pixels: zero gradients
    const float cx = fin ? fminf(fmaxf(fx, -1.0e8f), 1.0e8f) : 0.0f, cy = fin ? fminf(fmaxf(fy, -1.0e8f), 1.0e8f) : 0.0f;
    const float x0f = floorf(cx), y0f = floorf(cy);
    const int x0 = (int)x0f, y0 = (int)y0f, x1 = x0 + 1, y1 = y0 + 1;
    const float ax = (x0f + 1.0f) - cx, bx = cx - x0f, ay = (y0f + 1.0f) - cy, by = cy - y0f;   // weights: ax*ay, bx*ay, ax*by, bx*by
    const bool inx0 = x0 >= 0 && x0 < W, inx1 = x1 >= 0 && x1 < W, iny0 = y0 >= 0 && y0 < H, iny1 = y1 >= 0 && y1 < H;
    const bool k_nw = fin && inx0 && iny0, k_ne = fin && inx1 && iny0, k_sw = fin && inx0 && iny1, k_se = fin && inx1 && iny1;
    const int64_t o_nw = (int64_t)y0 * W + x0, o_ne = o_nw + 1, o_sw = o_nw + W, o_se = o_sw + 1;
    float gfx = 0.0f, gfy = 0.0f;
    const float* gp = gout + (int64_t)b * C * HW;
    const float* ip = in + (int64_t)b * C * HW + p;
    float* gi = gin ? gin + (int64_t)b * C * HW + p : nullptr;
    for (int c = 0; c < C; ++c, gp += HW, ip += HW) {
        const float g_nw = k_nw ? __ldg(gp + o_nw) : 0.0f, g_ne = k_ne ? __ldg(gp + o_ne) : 0.0f;
        const float g_sw = k_sw ? __ldg(gp + o_sw) : 0.0f, g_se = k_se ? __ldg(gp + o_se) : 0.0f;
        if (gi) {
            gi[0] = ax * ay * g_nw + bx * ay * g_ne + ax * by * g_sw + bx * by * g_se;
            gi += HW;
        }
        if (gflow) {
            const float v = __ldg(ip);
            gfx += v * (ay * (g_ne - g_nw) + by * (g_se - g_sw));
            gfy += v * (ax * (g_sw - g_nw) + bx * (g_se - g_ne));
        }
    }
    if (gflow) {
        gflow[((int64_t)b * 2 + 0) * HW + p] = gfx;
        gflow[((int64_t)b * 2 + 1) * HW + p] = gfy;
    }
}

// ---- bwarp backward -----------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bwarp_bwd_kernel(const float* __restrict__ img, const float* __restrict__ flow,
                                                        const float* __restrict__ gout, const float* __restrict__ lin_x,
                                                        const float* __restrict__ lin_y, float* __restrict__ gimg,
                                                        float* __restrict__ gflow, int C, int H, int W) {
    const int x = blockIdx.x * 32 + threadIdx.x;
    const int y = blockIdx.y * 8 + threadIdx.y;
    const int b = blockIdx.z;
    if (x >= W || y >= H) return;
    const int64_t HW = (int64_t)H * W, p = (int64_t)y * W + x;
    const float fx = flow[((int64_t)b * 2 + 0) * HW + p], fy = flow[((int64_t)b * 2 + 1) * HW + p];
    const float half_w = (float)((W - 1.0) / 2.0), half_h = (float)((H - 1.0) / 2.0);
    const float wm1 = (float)(W - 1), hm1 = (float)(H - 1);
    // same coordinate replay as the forward kernel (bwarp.cu: bwarp_geom)
    const float ixr = replay_coord(lin_x[x], __fdiv_rn(fx, half_w), wm1);
    const float iyr = replay_coord(lin_y[y], __fdiv_rn(fy, half_h), hm1);
    const float ix = fminf(fmaxf(ixr, 0.0f), wm1), iy = fminf(fmaxf(iyr, 0.0f), hm1);
    // border padding: the clamp has zero slope outside [0, size-1] (ATen clip_coordinates_set_grad)
    const float mx = (ixr >= 0.0f && ixr <= wm1) ? 1.0f : 0.0f, my = (iyr >= 0.0f && iyr <= hm1) ? 1.0f : 0.0f;
    const float xf = floorf(ix), yf = floorf(iy);
    const float tx = ix - xf, ty = iy - yf, ex = 1.0f - tx, ey = 1.0f - ty;
    const int x0 = (int)xf, y0 = (int)yf;
    const bool hx = x0 + 1 <= W - 1, hy = y0 + 1 <= H - 1;        // the +1 taps exist (their weight is 0 otherwise)
    const int64_t o_nw = (int64_t)y0 * W + x0, o_ne = o_nw + (hx ? 1 : 0), o_sw = o_nw + (hy ? W : 0), o_se = o_sw + (hx ? 1 : 0);
    float gx = 0.0f, gy = 0.0f;
    const float* ip = img + (int64_t)b * C * HW;
    const float* gp = gout + (int64_t)b * C * HW + p;
    float* gi = gimg ? gimg + (int64_t)b * C * HW : nullptr;
    for (int c = 0; c < C; ++c, ip += HW, gp += HW) {
        const float g = __ldg(gp);
        if (gi) {
            atomicAdd(gi + o_nw, g * ey * ex);
            if (hx) atomicAdd(gi + o_ne, g * ey * tx);
            if (hy) atomicAdd(gi + o_sw, g * ty * ex);
            if (hx && hy) atomicAdd(gi + o_se, g * ty * tx);
            gi += HW;
        }
        if (gflow) {
            const float v_nw = __ldg(ip + o_nw), v_ne = hx ? __ldg(ip + o_ne) : 0.0f;
            const float v_sw = hy ? __ldg(ip + o_sw) : 0.0f, v_se = (hx && hy) ? __ldg(ip + o_se) : 0.0f;
            gx += g * ((v_ne - v_nw) * ey + (v_se - v_sw) * ty);
            gy += g * ((v_sw - v_nw) * ex + (v_se - v_ne) * tx);
        }
    }
    if (gflow) {
        // d(pixel coordinate)/d(flow) = ((size-1)/2) * (1 / ((size-1)/2)), formed as autograd forms it
        gflow[((int64_t)b * 2 + 0) * HW + p] = gx * mx * (wm1 * 0.5f) * (1.0f / half_w);
        gflow[((int64_t)b * 2 + 1) * HW + p] = gy * my * (hm1 * 0.5f) * (1.0f / half_h);
    }
}

// ---- grouped sampler backward (row f1): what autograd derives for modules/deformable_attn.py:88-101 --------------
// out[b, g*gc+c, s, p] = bilinear_zeros(feat[b, g*gc+c], ix, iy),  ix = ((lin_x + (dx/(W-1))/2 + 1)/2)*(W-1),
// iy = ((lin_y + dy/((H-1)/2) + 1)/2)*(H-1).  One thread per (pixel, group, sample) as in the forward kernel; the
// feature gradient is scattered with atomics (grad_feat zero-filled by the entry point), the coordinate gradient is
// d out/d ix = sum_c g_c ((v_ne - v_nw) ey + (v_se - v_sw) ty) with out-of-frame taps reading 0 (ATen
// grid_sampler_2d_backward, padding_mode = zeros), chained through the reference's own scale factors.
__global__ void __launch_bounds__(256) dsample_bwd_kernel(const float* __restrict__ feat, const float* __restrict__ coords,
                                                          const float* __restrict__ gout, const float* __restrict__ lin_x,
                                                          const float* __restrict__ lin_y, float* __restrict__ gfeat,
                                                          float* __restrict__ gcoords, int C, int H, int W, int nG, int nS,
                                                          int row0, int rows) {
    // coords / gout / gcoords hold the output rows [row0, row0 + rows) only (the whole plane when rows == H)
    const int64_t HW = (int64_t)H * W, HWs = (int64_t)rows * W;
    const int64_t pix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= HWs) return;
    const int gs = blockIdx.y, g = gs / nS, s = gs - g * nS, b = blockIdx.z;
    const int ys = (int)(pix / W), x = (int)(pix - (int64_t)ys * W), y = row0 + ys;
    const int gc = C / nG;
    const float* cp = coords + (((int64_t)b * nG * nS + gs) * 2) * HWs + pix;
    const float dx = __ldg(cp), dy = __ldg(cp + HWs);
    const float wm1 = (float)(W - 1), hm1 = (float)(H - 1), half_h = (float)((H - 1.0) / 2.0);
    const float gxn = __fadd_rn(lin_x[x], __fmul_rn(__fdiv_rn(dx, wm1), 0.5f));
    const float gyn = __fadd_rn(lin_y[y], __fdiv_rn(dy, half_h));
    float ix = __fmul_rn(__fmul_rn(__fadd_rn(gxn, 1.0f), 0.5f), wm1);
    float iy = __fmul_rn(__fmul_rn(__fadd_rn(gyn, 1.0f), 0.5f), hm1);
    const bool finite = isfinite(ix) && isfinite(iy);
    ix = fminf(fmaxf(finite ? ix : -2.0f, -2.0f), (float)W + 1.0f);
    iy = fminf(fmaxf(finite ? iy : -2.0f, -2.0f), (float)H + 1.0f);
    const float xf = floorf(ix), yf = floorf(iy);
    const float tx = ix - xf, ty = iy - yf, ex = 1.0f - tx, ey = 1.0f - ty;
    const int x0 = (int)xf, y0 = (int)yf;
    const bool in_x0 = x0 >= 0 && x0 < W, in_x1 = x0 + 1 >= 0 && x0 + 1 < W;
    const bool in_y0 = y0 >= 0 && y0 < H, in_y1 = y0 + 1 >= 0 && y0 + 1 < H;
    const bool k_nw = in_x0 && in_y0, k_ne = in_x1 && in_y0, k_sw = in_x0 && in_y1, k_se = in_x1 && in_y1;
    const int64_t o_nw = (int64_t)y0 * W + x0;
    const float* fp = feat + ((int64_t)b * C + (int64_t)g * gc) * HW;
    float* gf = gfeat ? gfeat + ((int64_t)b * C + (int64_t)g * gc) * HW : nullptr;
    const float* gp = gout + (((int64_t)b * C + (int64_t)g * gc) * nS + s) * HWs + pix;
    float gx = 0.0f, gy = 0.0f;
    for (int c = 0; c < gc; ++c) {
        const float go = __ldg(gp);
        if (gf) {
            if (k_nw) atomicAdd(gf + o_nw, go * ey * ex);
            if (k_ne) atomicAdd(gf + o_nw + 1, go * ey * tx);
            if (k_sw) atomicAdd(gf + o_nw + W, go * ty * ex);
            if (k_se) atomicAdd(gf + o_nw + W + 1, go * ty * tx);
            gf += HW;
        }
        if (gcoords) {
            const float v_nw = k_nw ? __ldg(fp + o_nw) : 0.0f, v_ne = k_ne ? __ldg(fp + o_nw + 1) : 0.0f;
            const float v_sw = k_sw ? __ldg(fp + o_nw + W) : 0.0f, v_se = k_se ? __ldg(fp + o_nw + W + 1) : 0.0f;
            gx += go * ((v_ne - v_nw) * ey + (v_se - v_sw) * ty);
            gy += go * ((v_sw - v_nw) * ex + (v_se - v_ne) * tx);
        }
        fp += HW;
        gp += (int64_t)nS * HWs;
    }
    if (gcoords) {
        float* gcp = gcoords + (((int64_t)b * nG * nS + gs) * 2) * HWs + pix;
        // unnormalise: d ix / d gx = (W-1)/2; then the reference's own (dx / (W-1)) / 2 and dy / ((H-1)/2)
        gcp[0] = finite ? __fmul_rn(__fdiv_rn(gx * (wm1 * 0.5f), wm1), 0.5f) : 0.0f;
        gcp[HWs] = finite ? __fdiv_rn(gy * (hm1 * 0.5f), half_h) : 0.0f;
    }
}

}  // namespace vfi

extern "C" int vfi_dsample_bwd_f32(const float* feat, const float* coords, const float* grad_out, const float* lin_x,
                                   const float* lin_y, float* grad_feat, float* grad_coords, int B, int C, int H, int W,
                                   int nG, int nS, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(feat && coords && grad_out && lin_x && lin_y, "vfi_dsample_bwd_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1 && nG >= 1 && nS >= 1 && C % nG == 0, "vfi_dsample_bwd_f32: bad shape");
    VFI_REQUIRE((int64_t)nG * nS <= 65535 && B <= 65535, "vfi_dsample_bwd_f32: nG*nS or B exceeds grid limits");
    if (B == 0 || (!grad_feat && !grad_coords)) return VFI_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (grad_feat) VFI_CUDA(cudaMemsetAsync(grad_feat, 0, sizeof(float) * (size_t)B * C * H * W, st));
    const int64_t HW = (int64_t)H * W;
    dim3 grid((unsigned)ceil_div(HW, 256), (unsigned)(nG * nS), (unsigned)B);
    dsample_bwd_kernel<<<grid, 256, 0, st>>>(feat, coords, grad_out, lin_x, lin_y, grad_feat, grad_coords, C, H, W, nG, nS, 0, H);
    VFI_CHECK_LAUNCH("dsample_bwd_kernel");
    return VFI_OK;
}

extern "C" int vfi_dsample_rows_bwd_f32(const float* feat, const float* coords, const float* grad_out, const float* lin_x,
                                        const float* lin_y, float* grad_feat, float* grad_coords, int B, int C, int H, int W,
                                        int nG, int nS, int row0, int rows, int accumulate, void* stream) {
    using namespace vfi;
    if (B == 0 || rows == 0) return VFI_OK;
    VFI_REQUIRE(feat && coords && grad_out && lin_x && lin_y, "vfi_dsample_rows_bwd_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1 && nG >= 1 && nS >= 1 && C % nG == 0, "vfi_dsample_rows_bwd_f32: bad shape");
    VFI_REQUIRE(row0 >= 0 && rows >= 0 && row0 + rows <= H, "vfi_dsample_rows_bwd_f32: rows [%d, %d) outside 0..%d", row0, row0 + rows, H);
    VFI_REQUIRE((int64_t)nG * nS <= 65535 && B <= 65535, "vfi_dsample_rows_bwd_f32: nG*nS or B exceeds grid limits");
    if (!grad_feat && !grad_coords) return VFI_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (grad_feat && !accumulate) VFI_CUDA(cudaMemsetAsync(grad_feat, 0, sizeof(float) * (size_t)B * C * H * W, st));
    dim3 grid((unsigned)ceil_div((int64_t)rows * W, 256), (unsigned)(nG * nS), (unsigned)B);
    dsample_bwd_kernel<<<grid, 256, 0, st>>>(feat, coords, grad_out, lin_x, lin_y, grad_feat, grad_coords, C, H, W, nG, nS, row0, rows);
    VFI_CHECK_LAUNCH("dsample_bwd_kernel");
    return VFI_OK;
}

extern "C" int vfi_splat_bwd_f32(const float* in, const float* flow, const float* grad_out, float* grad_in, float* grad_flow,
                                 int B, int C, int H, int W, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(in && flow && grad_out, "vfi_splat_bwd_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1, "vfi_splat_bwd_f32: bad shape");
    if (B == 0 || (!grad_in && !grad_flow)) return VFI_OK;
    VFI_REQUIRE(B <= 65535, "vfi_splat_bwd_f32: batch %d too large", B);
    const dim3 grid((W + 31) / 32, (H + 7) / 8, B);
    splat_bwd_kernel<<<grid, dim3(32, 8), 0, static_cast<cudaStream_t>(stream)>>>(in, flow, grad_out, grad_in, grad_flow, C, H, W);
    VFI_CHECK_LAUNCH("splat_bwd_kernel");
    return VFI_OK;
}

extern "C" int vfi_bwarp_bwd_f32(const float* img, const float* flow, const float* grad_out, const float* lin_x,
                                 const float* lin_y, float* grad_img, float* grad_flow, int B, int C, int H, int W, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(img && flow && grad_out && lin_x && lin_y, "vfi_bwarp_bwd_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1, "vfi_bwarp_bwd_f32: bad shape");
    if (B == 0 || (!grad_img && !grad_flow)) return VFI_OK;
    VFI_REQUIRE(B <= 65535, "vfi_bwarp_bwd_f32: batch %d too large", B);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (grad_img) VFI_CUDA(cudaMemsetAsync(grad_img, 0, sizeof(float) * (size_t)B * C * H * W, st));
    const dim3 grid((W + 31) / 32, (H + 7) / 8, B);
    bwarp_bwd_kernel<<<grid, dim3(32, 8), 0, st>>>(img, flow, grad_out, lin_x, lin_y, grad_img, grad_flow, C, H, W);
    VFI_CHECK_LAUNCH("bwarp_bwd_kernel");
    return VFI_OK;
}
