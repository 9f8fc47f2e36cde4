// Fused deformable attention: offsets -> grouped bilinear sampling of both frames -> per-pixel multi-head
// softmax attention, without ever materialising the sampled kv tensor.
// Reference (unfused): modules/deformable_attn.py:81-101 (_get_ref_coords/_get_ref_feats) + 26-40 (Attn),
// and the madat flavour modules/madat.py:86-115, 154-174.
//
// Algebra (fp32 reassociation only): qp = Wq q + bq;  kt_h = scale * Wk_h^T qp_h;  cb_h = scale * qp_h.bk_h
//   logit[h,s] = kt_h . x_s + cb_h ;  a = softmax_s(logit) ;  out_h = Wv_h (sum_s a[h,s] x_s) + bv_h
// where x_s (C values) is slot s of the pixel: group g's channels sampled at group g's s-th point.
//
// Mapping: a TEAM of NG lanes owns one pixel; lane g of the team owns group g for sampling (its offsets,
// its geometry, its GC channels of x_s) and head g for the softmax.  Per slot the lane forms the NG partial
// dot products kt_h[block g] . x_s[block g]; a transposed butterfly over the team (NG-1 shuffles) leaves
// head h's logit in lane h.  The softmax shift is lazy (re-based only when a logit exceeds the current
// base by 30), so the common path per slot is 2*NG*GC FMAs + NG shuffles.  Requires nG == nH.
// HBM traffic per pixel: q, feat0, feat1, out (4C floats) + 2*2*nG*nS offsets + 4 flow values.
#include "sampler.cuh"

namespace vfi {

struct DattnArgs {
    const float *q, *feat[2], *offs[2], *ft[2];
    int64_t ft_sn, ft_sc, ft_sh, ft_sw;
    const float *Wq, *bq, *Wk, *bk, *Wv, *bv, *lin_x, *lin_y;
    float *out, *score[2];
    int B, H, W, nS;
    float offset_scale;
    int coords_final;
};

constexpr int kDattnThreads = 128;

template <int GC, int NG>
__global__ void __launch_bounds__(kDattnThreads) dattn_kernel(DattnArgs a) {
    constexpr int C = GC * NG;
    constexpr int CP = C | 1;                      // odd row stride: team lanes hit distinct banks
    constexpr int PPW = 32 / NG;                   // pixels per warp
    constexpr int PPB = (kDattnThreads / 32) * PPW;
    extern __shared__ float smem[];
    float* Wq_s = smem;
    float* Wk_s = Wq_s + C * CP;
    float* Wv_s = Wk_s + C * CP;
    for (int i = threadIdx.x; i < C * C; i += kDattnThreads) {
        const int r = i / C, c = i - r * C;
        Wq_s[r * CP + c] = __ldg(a.Wq + i);
        Wk_s[r * CP + c] = __ldg(a.Wk + i);
        Wv_s[r * CP + c] = __ldg(a.Wv + i);
    }
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane % NG, team = lane / NG, team_base = team * NG;
    const int64_t HW = (int64_t)a.H * a.W, total = (int64_t)a.B * HW;
    const float scale = rsqrtf((float)GC);
    const int nS = a.nS;

    for (int64_t tile = (int64_t)blockIdx.x * PPB; tile < total; tile += (int64_t)gridDim.x * PPB) {
        const int64_t gp_raw = tile + warp * PPW + team;
        const bool live = gp_raw < total;
        const int64_t gp = live ? gp_raw : total - 1;
        const int b = (int)(gp / HW);
        const int64_t pix = gp - (int64_t)b * HW;
        const int y = (int)(pix / a.W), x = (int)(pix - (int64_t)y * a.W);

        // ---- prologue: qp (my head), kt (all heads x my channel block), cb (my head) ----
        float qp[GC];
#pragma unroll
        for (int j = 0; j < GC; ++j) qp[j] = __ldg(a.bq + g * GC + j);
        {
            const float* qpix = a.q + (int64_t)b * C * HW + pix;
#pragma unroll 4
            for (int c = 0; c < C; ++c) {
                const float qv = __ldg(qpix + (int64_t)c * HW);
#pragma unroll
                for (int j = 0; j < GC; ++j) qp[j] = fmaf(Wq_s[(g * GC + j) * CP + c], qv, qp[j]);
            }
        }
        float cb = 0.0f;
#pragma unroll
        for (int j = 0; j < GC; ++j) cb = fmaf(qp[j], __ldg(a.bk + g * GC + j), cb);
        cb *= scale;
        float kt[NG][GC];
#pragma unroll
        for (int h = 0; h < NG; ++h) {
#pragma unroll
            for (int cc = 0; cc < GC; ++cc) kt[h][cc] = 0.0f;
#pragma unroll
            for (int j = 0; j < GC; ++j) {
                const float qh = __shfl_sync(0xffffffffu, qp[j], team_base + h) * scale;
                const float* wr = Wk_s + (h * GC + j) * CP + g * GC;
#pragma unroll
                for (int cc = 0; cc < GC; ++cc) kt[h][cc] = fmaf(qh, wr[cc], kt[h][cc]);
            }
        }

        float yv[NG][GC];
#pragma unroll
        for (int h = 0; h < NG; ++h)
#pragma unroll
            for (int cc = 0; cc < GC; ++cc) yv[h][cc] = 0.0f;
        float m_ref = -INFINITY, sum = 0.0f;
        const float lx = __ldg(a.lin_x + x), ly = __ldg(a.lin_y + y);

        for (int f = 0; f < 2; ++f) {
            float ftx = 0.0f, fty = 0.0f;
            if (!a.coords_final) {
                const float* fp = a.ft[f] + b * a.ft_sn + y * a.ft_sh + x * a.ft_sw;
                ftx = __ldg(fp);
                fty = __ldg(fp + a.ft_sc);
            }
            const float* offp = a.offs[f] + ((int64_t)b * NG * nS + (int64_t)g * nS) * 2 * HW + pix;
            const float* featp = a.feat[f] + ((int64_t)b * C + g * GC) * HW;
            for (int s = 0; s < nS; ++s) {
                float dx = ld_stream(offp + (int64_t)(2 * s) * HW);
                float dy = ld_stream(offp + (int64_t)(2 * s + 1) * HW);
                if (!a.coords_final) {
                    dx = __fadd_rn(__fmul_rn(a.offset_scale, tanhf(dx)), ftx);
                    dy = __fadd_rn(__fmul_rn(a.offset_scale, tanhf(dy)), fty);
                }
                const TapGeom tg = sampler_geom(dx, dy, lx, ly, a.H, a.W);
                float xs[GC];
#pragma unroll
                for (int cc = 0; cc < GC; ++cc) xs[cc] = sample_plane(featp + (int64_t)cc * HW, tg, a.W);
                float part[NG];
#pragma unroll
                for (int h = 0; h < NG; ++h) {
                    float d = 0.0f;
#pragma unroll
                    for (int cc = 0; cc < GC; ++cc) d = fmaf(kt[h][cc], xs[cc], d);
                    part[h] = d;
                }
                // transposed butterfly: afterwards part[0] of lane g is the full dot product of head g
#pragma unroll
                for (int off = NG / 2; off >= 1; off >>= 1) {
                    const bool upper = (g & off) != 0;
#pragma unroll
                    for (int i = 0; i < off; ++i) {
                        const float send = upper ? part[i] : part[i + off];
                        const float keep = upper ? part[i + off] : part[i];
                        part[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
                    }
                }
                const float l = part[0] + cb;
                if (a.score[f] != nullptr && live)
                    a.score[f][(((int64_t)b * NG + g) * nS + s) * HW + pix] = l;
                const float d = l - m_ref;
                const bool need = d > 30.0f;
                if (__any_sync(0xffffffffu, need)) {       // re-base the softmax shift (rare)
                    const float alpha = need ? expf(-d) : 1.0f;
                    if (need) {
                        m_ref = l;
                        sum *= alpha;
                    }
#pragma unroll
                    for (int h = 0; h < NG; ++h) {
                        const float ah = __shfl_sync(0xffffffffu, alpha, team_base + h);
#pragma unroll
                        for (int cc = 0; cc < GC; ++cc) yv[h][cc] *= ah;
                    }
                }
                const float p = expf(l - m_ref);
                sum += p;
#pragma unroll
                for (int h = 0; h < NG; ++h) {
                    const float ph = __shfl_sync(0xffffffffu, p, team_base + h);
#pragma unroll
                    for (int cc = 0; cc < GC; ++cc) yv[h][cc] = fmaf(ph, xs[cc], yv[h][cc]);
                }
            }
        }

        // ---- epilogue: out_h = Wv_h y_h / sum_h + bv_h, reduced over the team's channel blocks ----
        // reuse kt as the partial-output array: po[h][j] = sum_cc Wv[h*GC+j][g*GC+cc] * yv[h][cc]
#pragma unroll
        for (int h = 0; h < NG; ++h) {
#pragma unroll
            for (int j = 0; j < GC; ++j) {
                const float* wr = Wv_s + (h * GC + j) * CP + g * GC;
                float d = 0.0f;
#pragma unroll
                for (int cc = 0; cc < GC; ++cc) d = fmaf(wr[cc], yv[h][cc], d);
                kt[h][j] = d;
            }
        }
#pragma unroll
        for (int off = NG / 2; off >= 1; off >>= 1) {
            const bool upper = (g & off) != 0;
#pragma unroll
            for (int i = 0; i < off; ++i) {
#pragma unroll
                for (int j = 0; j < GC; ++j) {
                    const float send = upper ? kt[i][j] : kt[i + off][j];
                    const float keep = upper ? kt[i + off][j] : kt[i][j];
                    kt[i][j] = keep + __shfl_xor_sync(0xffffffffu, send, off);
                }
            }
        }
        if (live) {
            const float inv = 1.0f / sum;
            float* op = a.out + ((int64_t)b * C + g * GC) * HW + pix;
#pragma unroll
            for (int j = 0; j < GC; ++j) op[(int64_t)j * HW] = fmaf(kt[0][j], inv, __ldg(a.bv + g * GC + j));
        }
    }
}

int launch_softmax_samples(float* sc, int64_t R, int n, int64_t HW, cudaStream_t st);

template <int GC, int NG>
static int launch_dattn(const DattnArgs& a, cudaStream_t st) {
    constexpr int C = GC * NG;
    constexpr int PPB = (kDattnThreads / 32) * (32 / NG);
    const size_t smem = (size_t)3 * C * (C | 1) * sizeof(float);
    VFI_CUDA(cudaFuncSetAttribute(dattn_kernel<GC, NG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t tiles = ceil_div((int64_t)a.B * a.H * a.W, PPB);
    const int blocks = (int)min64(tiles, (int64_t)kNumSMs * 3);
    dattn_kernel<GC, NG><<<blocks, kDattnThreads, smem, st>>>(a);
    VFI_CHECK_LAUNCH("dattn_kernel");
    return VFI_OK;
}

}  // namespace vfi

extern "C" int vfi_dattn_f32(const float* q, const float* feat0, const float* feat1, const float* offs0, const float* offs1,
                             const float* ft0, const float* ft1, const int64_t ft_strides[4], const float* Wq, const float* bq,
                             const float* Wk, const float* bk, const float* Wv, const float* bv, const float* lin_x,
                             const float* lin_y, float* out, float* score0, float* score1, int B, int C, int H, int W, int nG,
                             int nS, int nH, float offset_scale, int coords_final, void* stream) {
    using namespace vfi;
    VFI_REQUIRE(q && feat0 && feat1 && offs0 && offs1 && Wq && bq && Wk && bk && Wv && bv && lin_x && lin_y && out,
                "vfi_dattn_f32: null pointer");
    VFI_REQUIRE(coords_final || (ft0 && ft1 && ft_strides), "vfi_dattn_f32: raw offsets need ft0/ft1 and their strides");
    VFI_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1 && nG >= 1 && nS >= 1, "vfi_dattn_f32: bad shape");
    VFI_REQUIRE((score0 == nullptr) == (score1 == nullptr), "vfi_dattn_f32: score0/score1 must both be given or both NULL");
    if (nG != nH || C % nG != 0) VFI_UNSUPPORTED("vfi_dattn_f32: fused kernel needs nG == nH dividing C (got C=%d nG=%d nH=%d)", C, nG, nH);
    if (B == 0) return VFI_OK;
    DattnArgs a;
    a.q = q;
    a.feat[0] = feat0; a.feat[1] = feat1;
    a.offs[0] = offs0; a.offs[1] = offs1;
    a.ft[0] = ft0; a.ft[1] = ft1;
    a.ft_sn = ft_strides ? ft_strides[0] : 0; a.ft_sc = ft_strides ? ft_strides[1] : 0;
    a.ft_sh = ft_strides ? ft_strides[2] : 0; a.ft_sw = ft_strides ? ft_strides[3] : 0;
    a.Wq = Wq; a.bq = bq; a.Wk = Wk; a.bk = bk; a.Wv = Wv; a.bv = bv;
    a.lin_x = lin_x; a.lin_y = lin_y;
    a.out = out; a.score[0] = score0; a.score[1] = score1;
    a.B = B; a.H = H; a.W = W; a.nS = nS;
    a.offset_scale = offset_scale;
    a.coords_final = coords_final;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int gc = C / nG;
    int rc;
#define VFI_DATTN_CASE(GC_, NG_) \
    if (gc == GC_ && nG == NG_) rc = launch_dattn<GC_, NG_>(a, st); else
    VFI_DATTN_CASE(9, 8)
    VFI_DATTN_CASE(18, 4)
    VFI_DATTN_CASE(8, 8)
    VFI_DATTN_CASE(16, 4)
    VFI_DATTN_CASE(4, 4)
    VFI_DATTN_CASE(2, 8)
    VFI_DATTN_CASE(8, 4)
    VFI_DATTN_CASE(3, 8)
    VFI_DATTN_CASE(6, 4)
    { VFI_UNSUPPORTED("vfi_dattn_f32: no instantiation for C/nG=%d, nG=%d", gc, nG); }
#undef VFI_DATTN_CASE
    if (rc) return rc;
    if (score0) {
        const int64_t HW = (int64_t)H * W;
        rc = launch_softmax_samples(score0, (int64_t)B * nH, nS, HW, st);
        if (rc) return rc;
        rc = launch_softmax_samples(score1, (int64_t)B * nH, nS, HW, st);
        if (rc) return rc;
    }
    return VFI_OK;
}
