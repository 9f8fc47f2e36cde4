// Modulated deformable convolution (DCNv2), torchvision.ops.deform_conv2d semantics
// (reference call sites: modules/dcnv2.py:60-66, models/archive/DCNDAT.py:71-77).
//
// SIMT fp32 path (algo 1): fused gather + per-group contraction, no im2col buffer in HBM.
// CTA = 128 output pixels x one weight group; the group's weights sit in shared memory laid out
// [ci*K + tap][out] so one LDS.128 feeds four FMAs; a thread owns one pixel and accumulates the
// group's Cout/WG outputs in registers.  Offsets/mask are read once per (offset group, tap).
// HBM traffic = 4*(Cin + 3*OG*K + Cout) B/px (x is re-read through L1/L2 only).
#include "dcn_common.cuh"

namespace vfi {

template <int COGP>
__global__ void __launch_bounds__(128) dcn_simt_kernel(const float* __restrict__ x, const float* __restrict__ offset,
                                                       const float* __restrict__ mask, const float* __restrict__ weight,
                                                       const float* __restrict__ bias, float* __restrict__ out, DcnParams p,
                                                       int o_begin, int o_count) {
    extern __shared__ __align__(16) float w_s[];  // [cin_g*K][COGP]
    const int K = p.kh * p.kw;
    const int cin_g = p.Cin / p.WG, cog = p.Cout / p.WG, cpo = p.Cin / p.OG;
    const int g = blockIdx.y, b = blockIdx.z;
    // stage this group's weight slice: weight[(g*cog + o_begin + o), ci, tap] -> w_s[(ci*K+tap)*COGP + o]
    for (int i = threadIdx.x; i < cin_g * K * COGP; i += blockDim.x) {
        const int o = i % COGP, r = i / COGP;
        w_s[i] = o < o_count ? __ldg(weight + ((size_t)(g * cog + o_begin + o) * cin_g) * K + r) : 0.0f;
    }
    __syncthreads();
    const int64_t HWo = (int64_t)p.Ho * p.Wo, HW = (int64_t)p.H * p.W;
    const int64_t pix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= HWo) return;
    const int ho = (int)(pix / p.Wo), wo = (int)(pix - (int64_t)ho * p.Wo);
    const float base_h = (float)(ho * p.stride - p.pad), base_w = (float)(wo * p.stride - p.pad);

    float acc[COGP];
#pragma unroll
    for (int o = 0; o < COGP; ++o) acc[o] = 0.0f;

    const float* xb = x + ((int64_t)b * p.Cin + (int64_t)g * cin_g) * HW;
    const float* offb = offset + (int64_t)b * p.OG * 2 * K * HWo + pix;
    const float* maskb = mask ? mask + (int64_t)b * p.OG * K * HWo + pix : nullptr;

    for (int i = 0; i < p.kh; ++i) {
        for (int j = 0; j < p.kw; ++j) {
            const int tap = i * p.kw + j;
            int prev_og = -1;
            DcnTap t;
            float mval = 1.0f;
            for (int cl = 0; cl < cin_g; ++cl) {
                const int og = (g * cin_g + cl) / cpo;
                if (og != prev_og) {
                    prev_og = og;
                    const float dy = ld_stream(offb + ((int64_t)og * 2 * K + 2 * tap) * HWo);
                    const float dx = ld_stream(offb + ((int64_t)og * 2 * K + 2 * tap + 1) * HWo);
                    mval = maskb ? ld_stream(maskb + ((int64_t)og * K + tap) * HWo) : 1.0f;
                    t = dcn_tap(base_h + (float)(i * p.dil) + dy, base_w + (float)(j * p.dil) + dx, p.H, p.W);
                }
                float v = dcn_sample(xb + (int64_t)cl * HW + t.o, t, p.W);
                v *= mval;
                const float4* wr = reinterpret_cast<const float4*>(w_s + (size_t)(cl * K + tap) * COGP);
#pragma unroll
                for (int o4 = 0; o4 < COGP / 4; ++o4) {
                    const float4 wv = wr[o4];
                    acc[o4 * 4 + 0] = fmaf(wv.x, v, acc[o4 * 4 + 0]);
                    acc[o4 * 4 + 1] = fmaf(wv.y, v, acc[o4 * 4 + 1]);
                    acc[o4 * 4 + 2] = fmaf(wv.z, v, acc[o4 * 4 + 2]);
                    acc[o4 * 4 + 3] = fmaf(wv.w, v, acc[o4 * 4 + 3]);
                }
            }
        }
    }
    float* op = out + ((int64_t)b * p.Cout + (int64_t)g * cog + o_begin) * HWo + pix;
#pragma unroll
    for (int o = 0; o < COGP; ++o) {
        if (o < o_count) {
            const float bv = bias ? __ldg(bias + g * cog + o_begin + o) : 0.0f;
            st_stream(op + (int64_t)o * HWo, acc[o] + bv);
        }
    }
}

template <int COGP>
static int launch_dcn_simt(const float* x, const float* offset, const float* mask, const float* weight, const float* bias,
                           float* out, int B, const DcnParams& p, int o_begin, int o_count, cudaStream_t st) {
    const int K = p.kh * p.kw;
    const size_t smem = (size_t)(p.Cin / p.WG) * K * COGP * sizeof(float);
    VFI_REQUIRE(smem <= 200 * 1024, "vfi_dcn_f32: weight slice of %zu B does not fit shared memory", smem);
    VFI_CUDA(cudaFuncSetAttribute(dcn_simt_kernel<COGP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)ceil_div((int64_t)p.Ho * p.Wo, 128), (unsigned)p.WG, (unsigned)B);
    dcn_simt_kernel<COGP><<<grid, 128, smem, st>>>(x, offset, mask, weight, bias, out, p, o_begin, o_count);
    VFI_CHECK_LAUNCH("dcn_simt_kernel");
    return VFI_OK;
}

int dcn_tc_supported(const DcnParams& p);
int launch_dcn_tc(const float* x, const float* offset, const float* mask, const float* weight, const float* bias, float* out,
                  int B, const DcnParams& p, void* workspace, size_t workspace_bytes, cudaStream_t st);
size_t dcn_tc_workspace_bytes(int B, const DcnParams& p);
int dcn_win_supported(const DcnParams& p, const float* x);
int launch_dcn_win(const float* x, const float* offset, const float* mask, const float* weight, const float* bias, float* out, int B,
                   const DcnParams& p, cudaStream_t st);

}  // namespace vfi

static int dcn_fill_params(vfi::DcnParams& p, int Cin, int Cout, int H, int W, int kh, int kw, int stride, int pad, int dil,
                           int WG, int OG) {
    using namespace vfi;
    VFI_REQUIRE(Cin >= 1 && Cout >= 1 && H >= 1 && W >= 1 && kh >= 1 && kw >= 1 && stride >= 1 && pad >= 0 && dil >= 1,
                "vfi_dcn_f32: bad shape");
    VFI_REQUIRE(WG >= 1 && OG >= 1 && Cin % WG == 0 && Cout % WG == 0 && Cin % OG == 0,
                "vfi_dcn_f32: Cin=%d Cout=%d not divisible by weight groups %d / offset groups %d", Cin, Cout, WG, OG);
    p = DcnParams{Cin, Cout, H, W, 0, 0, kh, kw, stride, pad, dil, WG, OG};
    p.Ho = (H + 2 * pad - (dil * (kh - 1) + 1)) / stride + 1;
    p.Wo = (W + 2 * pad - (dil * (kw - 1) + 1)) / stride + 1;
    VFI_REQUIRE(p.Ho >= 1 && p.Wo >= 1, "vfi_dcn_f32: empty output");
    return VFI_OK;
}

extern "C" size_t vfi_dcn_workspace_bytes(int B, int Cin, int Cout, int H, int W, int kh, int kw, int n_weight_grps,
                                          int n_offset_grps, int algo) {
    vfi::DcnParams p;
    if (dcn_fill_params(p, Cin, Cout, H, W, kh, kw, 1, (kh - 1) / 2, 1, n_weight_grps, n_offset_grps) != VFI_OK) return 0;
    if (algo == 1 || algo == 3) return 0;
    return vfi::dcn_tc_workspace_bytes(B, p);
}

extern "C" int vfi_dcn_f32(const float* x, const float* offset, const float* mask, const float* weight, const float* bias,
                           float* out, int B, int Cin, int Cout, int H, int W, int kh, int kw, int stride, int pad, int dil,
                           int n_weight_grps, int n_offset_grps, int algo, void* workspace, size_t workspace_bytes,
                           void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(x && offset && weight && out, "vfi_dcn_f32: null pointer");
    VFI_REQUIRE(B >= 0 && B <= 65535, "vfi_dcn_f32: bad batch %d", B);
    DcnParams p;
    int rc = dcn_fill_params(p, Cin, Cout, H, W, kh, kw, stride, pad, dil, n_weight_grps, n_offset_grps);
    if (rc) return rc;
    if (B == 0) return VFI_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    VFI_REQUIRE(algo >= 0 && algo <= 3, "vfi_dcn_f32: bad algo %d", algo);
    // window form (dcn_win.cu): the default where it applies (3x3, stride 1, <= 16 channels per group, a plane large enough
    // to pay for the staging).  Taps outside the staged window go through a cooperative gather list, so frame-leaving
    // offsets cost a bounded extra pass instead of a global round trip per lane: BASELINE configs[2] 2.55 ms with no
    // outliers, 3.48 ms with SURVEY.md 8(d)'s 2 % (SIMT kernel: 3.66 / 3.67 ms); at 10 % outliers the SIMT kernel is the
    // faster one again (3.69 against 4.34 ms) -- algo 1 selects it.
    if (algo == 3 && !dcn_win_supported(p, x)) VFI_UNSUPPORTED("vfi_dcn_f32: window path does not cover this configuration");
    if (algo == 3 || (algo == 0 && dcn_tc_supported(p) != 2 && dcn_win_supported(p, x) && (int64_t)p.Ho * p.Wo >= 4096))
        return launch_dcn_win(x, offset, mask, weight, bias, out, B, p, st);
    if (algo == 2 && !dcn_tc_supported(p)) VFI_UNSUPPORTED("vfi_dcn_f32: tcgen05 path does not cover this configuration");
    if (algo == 2 || (algo == 0 && dcn_tc_supported(p) == 2))
        return launch_dcn_tc(x, offset, mask, weight, bias, out, B, p, workspace, workspace_bytes, st);
    const int cog = Cout / n_weight_grps;
    for (int o_begin = 0; o_begin < cog; o_begin += 16) {
        const int cnt = cog - o_begin < 16 ? cog - o_begin : 16;
        if (cnt <= 4) rc = launch_dcn_simt<4>(x, offset, mask, weight, bias, out, B, p, o_begin, cnt, st);
        else if (cnt <= 8) rc = launch_dcn_simt<8>(x, offset, mask, weight, bias, out, B, p, o_begin, cnt, st);
        else if (cnt <= 12) rc = launch_dcn_simt<12>(x, offset, mask, weight, bias, out, B, p, o_begin, cnt, st);
        else rc = launch_dcn_simt<16>(x, offset, mask, weight, bias, out, B, p, o_begin, cnt, st);
        if (rc) return rc;
    }
    return VFI_OK;
}
