// Backward of the modulated deformable convolution: the gradients torchvision's
// deform_conv2d autograd gives the reference (call sites modules/dcnv2.py:60-66, models/archive/DCNDAT.py:71-77).
// Training only, so the torchvision decomposition is kept and only its two data-dependent steps are kernels:
//   vfi_dcn_im2col_f32      columns[b][c][k][p] = mask * bilinear(x[b][c], p + tap k + offset)      (for grad_weight)
//   vfi_dcn_col2im_f32      from grad_columns = W^T grad_out: grad_input (atomics), grad_offset, grad_mask
// The two dense contractions in between (grad_weight = grad_out . columns^T, grad_columns = W^T . grad_out, per weight
// group) are library GEMMs issued by the Python wrapper.  Sampling rules are torchvision's: bilinear_interpolate for
// values (0 beyond one pixel outside, per-corner bounds), get_coordinate_weight for the offset gradient.
#include "dcn_common.cuh"

namespace vfi {

struct DcnPos {
    float h, w, m;
};

__device__ __forceinline__ DcnPos dcn_pos(const DcnParams& P, const float* __restrict__ offset, const float* __restrict__ mask,
                                          int b, int og, int k, int oy, int ox) {
    const int K = P.kh * P.kw, i = k / P.kw, j = k - i * P.kw;
    const int64_t Po = (int64_t)P.Ho * P.Wo, p = (int64_t)oy * P.Wo + ox;
    const float* op = offset + ((int64_t)b * P.OG * 2 * K + (int64_t)(og * K + k) * 2) * Po + p;
    DcnPos r;
    r.h = (float)(oy * P.stride - P.pad + i * P.dil) + __ldg(op);
    r.w = (float)(ox * P.stride - P.pad + j * P.dil) + __ldg(op + Po);
    r.m = mask ? __ldg(mask + ((int64_t)b * P.OG * K + og * K + k) * Po + p) : 1.0f;
    return r;
}

// one thread per (output pixel, offset group, tap): loops over the channels of the offset group
__global__ void __launch_bounds__(256) dcn_im2col_kernel(const float* __restrict__ x, const float* __restrict__ offset,
                                                         const float* __restrict__ mask, float* __restrict__ columns, DcnParams P) {
    const int K = P.kh * P.kw;
    const int64_t Po = (int64_t)P.Ho * P.Wo, HW = (int64_t)P.H * P.W;
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= Po) return;
    const int og = blockIdx.y / K, k = blockIdx.y - og * K, b = blockIdx.z;
    const int oy = (int)(p / P.Wo), ox = (int)(p - (int64_t)oy * P.Wo);
    const DcnPos q = dcn_pos(P, offset, mask, b, og, k, oy, ox);
    const DcnTap t = dcn_tap(q.h, q.w, P.H, P.W);
    const int cpg = P.Cin / P.OG;
    for (int cc = 0; cc < cpg; ++cc) {
        const int c = og * cpg + cc;
        const float v = dcn_sample(x + ((int64_t)b * P.Cin + c) * HW + t.o, t, P.W);
        columns[(((int64_t)b * P.Cin + c) * K + k) * Po + p] = q.m * v;
    }
}

__global__ void __launch_bounds__(256) dcn_col2im_kernel(const float* __restrict__ x, const float* __restrict__ offset,
                                                         const float* __restrict__ mask, const float* __restrict__ gcol,
                                                         float* __restrict__ gx, float* __restrict__ goff, float* __restrict__ gmask,
                                                         DcnParams P) {
    const int K = P.kh * P.kw, H = P.H, W = P.W;
    const int64_t Po = (int64_t)P.Ho * P.Wo, HW = (int64_t)H * W;
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= Po) return;
    const int og = blockIdx.y / K, k = blockIdx.y - og * K, b = blockIdx.z;
    const int oy = (int)(p / P.Wo), ox = (int)(p - (int64_t)oy * P.Wo);
    const DcnPos q = dcn_pos(P, offset, mask, b, og, k, oy, ox);
    const DcnTap t = dcn_tap(q.h, q.w, H, W);                      // value taps (bilinear_interpolate)
    // get_coordinate_weight: plain floor, per-corner validity, no one-pixel early-out
    const float hc = fminf(fmaxf(q.h, -2.0f), (float)H + 1.0f), wc = fminf(fmaxf(q.w, -2.0f), (float)W + 1.0f);
    const bool fin = isfinite(q.h) && isfinite(q.w);
    const float yl = floorf(hc), xl = floorf(wc);
    const int y_l = (int)yl, x_l = (int)xl;
    const float dy = hc - yl, dx = wc - xl;
    const bool vyl = fin && y_l >= 0 && y_l < H, vyh = fin && y_l + 1 >= 0 && y_l + 1 < H;
    const bool vxl = fin && x_l >= 0 && x_l < W, vxh = fin && x_l + 1 >= 0 && x_l + 1 < W;
    const int64_t oc = (int64_t)y_l * W + x_l;
    const int cpg = P.Cin / P.OG;
    float g_h = 0.0f, g_w = 0.0f, g_m = 0.0f;
    for (int cc = 0; cc < cpg; ++cc) {
        const int c = og * cpg + cc;
        const float* pl = x + ((int64_t)b * P.Cin + c) * HW;
        const float gc = __ldg(gcol + (((int64_t)b * P.Cin + c) * K + k) * Po + p);
        if (goff) {
            const float v_yx = (vyl && vxl) ? __ldg(pl + oc) : 0.0f, v_yX = (vyl && vxh) ? __ldg(pl + oc + 1) : 0.0f;
            const float v_Yx = (vyh && vxl) ? __ldg(pl + oc + W) : 0.0f, v_YX = (vyh && vxh) ? __ldg(pl + oc + W + 1) : 0.0f;
            g_h += q.m * (dx * (v_YX - v_yX) + (1.0f - dx) * (v_Yx - v_yx)) * gc;
            g_w += q.m * (dy * (v_YX - v_Yx) + (1.0f - dy) * (v_yX - v_yx)) * gc;
        }
        if (gmask) g_m += gc * dcn_sample(pl + t.o, t, W);
        if (gx) {
            float* gp = gx + ((int64_t)b * P.Cin + c) * HW + t.o;
            const float s = q.m * gc;
            if (t.k1) atomicAdd(gp, s * t.w1);
            if (t.k2) atomicAdd(gp + 1, s * t.w2);
            if (t.k3) atomicAdd(gp + W, s * t.w3);
            if (t.k4) atomicAdd(gp + W + 1, s * t.w4);
        }
    }
    if (goff) {
        float* op = goff + ((int64_t)b * P.OG * 2 * K + (int64_t)(og * K + k) * 2) * Po + p;
        op[0] = g_h;
        op[Po] = g_w;
    }
    if (gmask) gmask[((int64_t)b * P.OG * K + og * K + k) * Po + p] = g_m;
}

static int dcn_bwd_params(DcnParams& P, int B, int Cin, int H, int W, int kh, int kw, int stride, int pad, int dil, int OG,
                          const char* who) {
    VFI_REQUIRE(B >= 0 && Cin >= 1 && H >= 1 && W >= 1 && kh >= 1 && kw >= 1 && stride >= 1 && dil >= 1 && pad >= 0 && OG >= 1 &&
                    Cin % OG == 0, "%s: bad shape", who);
    P.Cin = Cin; P.Cout = 0; P.H = H; P.W = W; P.kh = kh; P.kw = kw; P.stride = stride; P.pad = pad; P.dil = dil; P.WG = 1; P.OG = OG;
    P.Ho = (H + 2 * pad - (dil * (kh - 1) + 1)) / stride + 1;
    P.Wo = (W + 2 * pad - (dil * (kw - 1) + 1)) / stride + 1;
    VFI_REQUIRE(P.Ho >= 1 && P.Wo >= 1 && (int64_t)OG * kh * kw <= 65535 && B <= 65535, "%s: output / grid out of range", who);
    return VFI_OK;
}

}  // namespace vfi

extern "C" int vfi_dcn_im2col_f32(const float* x, const float* offset, const float* mask, float* columns, int B, int Cin, int H,
                                  int W, int kh, int kw, int stride, int pad, int dil, int n_offset_grps, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(x && offset && columns, "vfi_dcn_im2col_f32: null pointer");
    DcnParams P;
    if (int rc = dcn_bwd_params(P, B, Cin, H, W, kh, kw, stride, pad, dil, n_offset_grps, "vfi_dcn_im2col_f32")) return rc;
    if (B == 0) return VFI_OK;
    const dim3 grid((unsigned)ceil_div((int64_t)P.Ho * P.Wo, 256), (unsigned)(P.OG * kh * kw), (unsigned)B);
    dcn_im2col_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, offset, mask, columns, P);
    VFI_CHECK_LAUNCH("dcn_im2col_kernel");
    return VFI_OK;
}

extern "C" int vfi_dcn_col2im_f32(const float* x, const float* offset, const float* mask, const float* grad_columns,
                                  float* grad_x, float* grad_offset, float* grad_mask, int B, int Cin, int H, int W, int kh,
                                  int kw, int stride, int pad, int dil, int n_offset_grps, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(x && offset && grad_columns, "vfi_dcn_col2im_f32: null pointer");
    VFI_REQUIRE(grad_mask == nullptr || mask != nullptr, "vfi_dcn_col2im_f32: grad_mask without a mask");
    DcnParams P;
    if (int rc = dcn_bwd_params(P, B, Cin, H, W, kh, kw, stride, pad, dil, n_offset_grps, "vfi_dcn_col2im_f32")) return rc;
    if (B == 0 || (!grad_x && !grad_offset && !grad_mask)) return VFI_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (grad_x) VFI_CUDA(cudaMemsetAsync(grad_x, 0, sizeof(float) * (size_t)B * Cin * H * W, st));
    const dim3 grid((unsigned)ceil_div((int64_t)P.Ho * P.Wo, 256), (unsigned)(P.OG * kh * kw), (unsigned)B);
    dcn_col2im_kernel<<<grid, 256, 0, st>>>(x, offset, mask, grad_columns, grad_x, grad_offset, grad_mask, P);
    VFI_CHECK_LAUNCH("dcn_col2im_kernel");
    return VFI_OK;
}
