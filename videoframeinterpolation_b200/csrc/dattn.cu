// Fused deformable attention: offsets -> grouped bilinear sampling of both frames -> per-pixel multi-head
// softmax attention, without ever materialising the sampled kv tensor.
// Reference (unfused): modules/deformable_attn.py:81-101 (_get_ref_coords/_get_ref_feats) + 26-40 (Attn),
// and the madat flavour modules/madat.py:86-115, 154-174.
//
// Algebra (fp32 reassociation only): qp = Wq q + bq;  kt_h = scale * Wk_h^T qp_h;  cb_h = scale * qp_h.bk_h
//   logit[h,s] = kt_h . x_s + cb_h ;  a = softmax_s(logit) ;  out_h = Wv_h (sum_s a[h,s] x_s) + bv_h
// where x_s (C values) is slot s of the pixel: group g's channels sampled at group g's s-th point.
//
// Mapping (v2).  The kernel is bound by the L1/LSU gather rate (4 taps x C x 2nS loads per pixel), so the
// mapping is chosen for the gathers: a WARP owns one group g for a tile of 32 adjacent pixels, i.e. the 32
// lanes of every tap load read neighbouring addresses of ONE feature plane (1-3 cache lines per request
// instead of the 8-10 of a lanes-across-groups mapping).  A CTA is NG warps = all groups of the same 32
// pixels.  Warp g also owns head g of the softmax.  Per chunk of SC slots:
//   A. every warp samples its group's GC channels for SC slots and forms the NG partial dot products
//      kt_h[block g] . x_s[block g]                       -> red[si][h][g][lane]      (shared memory)
//   C. warp h sums the NG partials of its head, updates the (lazily re-based) softmax state and
//      publishes p[h][si], alpha[h]                       -> pbuf / abuf             (shared memory)
//   D. every warp accumulates yv[h][cc] += p[h][si] * x[si][cc] for all heads.
// Epilogue: po[h][j] = Wv[h*GC+j][block g] . yv[h], reduced over g through shared memory by warp h.
// Requires nG == nH.  HBM traffic per pixel: q, feat0, feat1, out (4C floats) + 2*2*nG*nS offsets + 4 flow values.
#include <stdlib.h>

#include "dattn_common.cuh"

namespace vfi {

template <int GC, int NG, bool OCL>
__global__ void __launch_bounds__(DattnCfg<GC, NG>::THREADS, DattnCfg<GC, NG>::MIN_BLOCKS) dattn_kernel(DattnArgs a) {
    using Cfg = DattnCfg<GC, NG>;
    constexpr int C = Cfg::C, C4 = Cfg::C4, GCP = Cfg::GCP, SC = Cfg::SC, T = Cfg::THREADS;
    extern __shared__ __align__(16) float smem[];
    float* Wq_s = smem;                       // [C][C4]
    float* WkT_s = Wq_s + Cfg::WQ;            // [g][r][GCP] = Wk[r][g*GC + cc]
    float* WvT_s = WkT_s + Cfg::WT;           // [g][r][GCP] = Wv[r][g*GC + cc]
    float* uni = WvT_s + Cfg::WT;             // q_s/qp_s | red | epilogue exchange
    float* pbuf = uni + Cfg::UNI;             // [si][h/4][lane][h%4]  (128-bit reads of 4 heads)
    float* abuf = pbuf + Cfg::PB;             // [h/4][lane][h%4]
    int* rebase = reinterpret_cast<int*>(abuf + Cfg::AB);   // [chunk parity]: some head re-based its softmax shift
    static_assert(NG % 4 == 0, "head pairs / quads");
    constexpr int HP = NG / 2;                // head pairs: FFMA2 lanes = heads (2hp, 2hp+1)

    for (int i = threadIdx.x; i < C * C4; i += T) {
        const int r = i / C4, c = i - r * C4;
        Wq_s[i] = c < C ? __ldg(a.Wq + r * C + c) : 0.0f;
    }
    for (int i = threadIdx.x; i < NG * C * GCP; i += T) {
        const int cc = i % GCP, r = (i / GCP) % C, g = i / (GCP * C);
        const bool ok = cc < GC;
        WkT_s[i] = ok ? __ldg(a.Wk + r * C + g * GC + cc) : 0.0f;
        WvT_s[i] = ok ? __ldg(a.Wv + r * C + g * GC + cc) : 0.0f;
    }

    const int lane = threadIdx.x & 31;
    const int g = threadIdx.x >> 5;            // my group == my head
    const int64_t HW = (int64_t)a.H * a.W;
    const float scale = rsqrtf((float)GC);
    const int nS = a.nS;
    const int lx = lane & ((1 << a.tw_log2) - 1), ly = lane >> a.tw_log2;
    const int tiles_per_img = a.tiles_x * a.tiles_y;
    const int64_t n_tiles = (int64_t)a.B * tiles_per_img;
    float* q_s = uni;                // [C][32]
    float* qp_s = uni + C * 32;      // [C][32]

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int b = (int)(tile / tiles_per_img);
        const int trem = (int)(tile - (int64_t)b * tiles_per_img);
        const int ty = trem / a.tiles_x, tx = trem - ty * a.tiles_x;
        const int x_raw = (tx << a.tw_log2) + lx, y_raw = ty * (32 >> a.tw_log2) + ly;
        const bool live = x_raw < a.W && y_raw < a.H;
        const int x = min(x_raw, a.W - 1), y = min(y_raw, a.H - 1);
        const int64_t pix = (int64_t)y * a.W + x;

        __syncthreads();   // weights staged (first tile) / previous tile's epilogue reads done
        // ---- prologue ----
        {
            const int64_t q_sp = (a.qo_cl & 1) ? C : 1, q_sc = (a.qo_cl & 1) ? 1 : HW;
            const float* qpix = a.q + (int64_t)b * C * HW + pix * q_sp + (g * GC) * q_sc;
#pragma unroll
            for (int j = 0; j < GC; ++j) q_s[(g * GC + j) * 32 + lane] = __ldg(qpix + (int64_t)j * q_sc);
        }
        __syncthreads();
        float cb;
        {
            float qp[GC];
#pragma unroll
            for (int j = 0; j < GC; ++j) qp[j] = __ldg(a.bq + g * GC + j);
#pragma unroll 2
            for (int c4 = 0; c4 < C4; c4 += 4) {
                float qv[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) qv[k] = (c4 + k < C) ? q_s[(c4 + k) * 32 + lane] : 0.0f;
#pragma unroll
                for (int j = 0; j < GC; ++j) {
                    const float4 w = *reinterpret_cast<const float4*>(Wq_s + (g * GC + j) * C4 + c4);
                    qp[j] = fmaf(w.x, qv[0], qp[j]);
                    if (c4 + 1 < C) qp[j] = fmaf(w.y, qv[1], qp[j]);
                    if (c4 + 2 < C) qp[j] = fmaf(w.z, qv[2], qp[j]);
                    if (c4 + 3 < C) qp[j] = fmaf(w.w, qv[3], qp[j]);
                }
            }
            cb = 0.0f;
#pragma unroll
            for (int j = 0; j < GC; ++j) {
                cb = fmaf(qp[j], __ldg(a.bk + g * GC + j), cb);
                qp_s[(g * GC + j) * 32 + lane] = qp[j] * scale;
            }
            cb *= scale;
        }
        __syncthreads();
        f32x2 kt2[HP][GC];   // (kt[2hp][cc], kt[2hp+1][cc])
#pragma unroll
        for (int hp = 0; hp < HP; ++hp) {
            float ka[2][GCP];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int h = 2 * hp + e;
#pragma unroll
                for (int cc = 0; cc < GCP; ++cc) ka[e][cc] = 0.0f;
#pragma unroll
                for (int j = 0; j < GC; ++j) {
                    const float qh = qp_s[(h * GC + j) * 32 + lane];
                    const float* wr = WkT_s + (g * C + h * GC + j) * GCP;
#pragma unroll
                    for (int c4 = 0; c4 < GCP; c4 += 4) {
                        const float4 w = *reinterpret_cast<const float4*>(wr + c4);
                        ka[e][c4] = fmaf(qh, w.x, ka[e][c4]);
                        if (c4 + 1 < GC) ka[e][c4 + 1] = fmaf(qh, w.y, ka[e][c4 + 1]);
                        if (c4 + 2 < GC) ka[e][c4 + 2] = fmaf(qh, w.z, ka[e][c4 + 2]);
                        if (c4 + 3 < GC) ka[e][c4 + 3] = fmaf(qh, w.w, ka[e][c4 + 3]);
                    }
                }
            }
#pragma unroll
            for (int cc = 0; cc < GC; ++cc) kt2[hp][cc] = pack2(ka[0][cc], ka[1][cc]);
        }
        if (threadIdx.x == 0) rebase[0] = rebase[1] = 0;
        __syncthreads();   // q_s / qp_s (aliased with red) fully consumed

        f32x2 yv2[HP][GC];   // (yv[2hp][cc], yv[2hp+1][cc])
#pragma unroll
        for (int hp = 0; hp < HP; ++hp)
#pragma unroll
            for (int cc = 0; cc < GC; ++cc) yv2[hp][cc] = 0ull;
        int chunk = 0;
        float m_ref = -INFINITY, sum = 0.0f;
        const float lxv = __ldg(a.lin_x + x), lyv = __ldg(a.lin_y + y);
        float* red = uni;   // [si][h][g][lane]
        const int plane_bytes = (int)(HW * 4), row_bytes = a.W * 4;   // GC * HW * 4 < 2^31 checked by the host

        for (int f = 0; f < 2; ++f) {
            float ftx = 0.0f, fty = 0.0f;
            if (!a.coords_final) {
                const float* fp = a.ft[f] + b * a.ft_sn + y * a.ft_sh + x * a.ft_sw;
                ftx = __ldg(fp);
                fty = __ldg(fp + a.ft_sc);
            }
            const float* offp = a.offs[f] + ((int64_t)b * NG * nS + (int64_t)g * nS) * 2 * HW + pix;
            const float* featp = a.feat[f] + ((int64_t)b * C + g * GC) * HW;
            // offsets are the only HBM-streamed operand of the loop (2*nG*nS floats per pixel and frame): keep the
            // next chunk's loads in flight across the two barriers of the current chunk
            float odx[SC], ody[SC];
            // channels-last offsets: the 2*SC floats of a chunk are contiguous per pixel -> 128-bit loads when aligned
            const int nOff = NG * nS * 2;
            const float* offcl = a.offs[f] + ((int64_t)b * HW + pix) * nOff + (int64_t)g * nS * 2;
            const bool cl_vec = OCL && (nS % SC == 0) && ((reinterpret_cast<uintptr_t>(offcl) & 15) == 0) && (SC % 2 == 0);
            auto load_offs = [&](int s_first) {
                if (OCL) {
                    if (cl_vec) {
#pragma unroll
                        for (int k = 0; k < SC / 2; ++k) {
                            const float4 v = ld_stream4(offcl + 2 * s_first + 4 * k);
                            odx[2 * k] = v.x; ody[2 * k] = v.y; odx[2 * k + 1] = v.z; ody[2 * k + 1] = v.w;
                        }
                    } else {
#pragma unroll
                        for (int si = 0; si < SC; ++si) {
                            const int s = min(s_first + si, nS - 1);
                            odx[si] = ld_stream(offcl + 2 * s);
                            ody[si] = ld_stream(offcl + 2 * s + 1);
                        }
                    }
                } else {
#pragma unroll
                    for (int si = 0; si < SC; ++si) {
                        const int s = min(s_first + si, nS - 1);
                        odx[si] = ld_stream(offp + (int64_t)(2 * s) * HW);
                        ody[si] = ld_stream(offp + (int64_t)(2 * s + 1) * HW);
                    }
                }
            };
            load_offs(0);
            for (int s0 = 0; s0 < nS; s0 += SC) {
                float xs[SC][GC];
                // ---- A: sample + partial logits ----
#pragma unroll
                for (int si = 0; si < SC; ++si) {
                    float dx = odx[si], dy = ody[si];
                    if (a.obias != nullptr) {
                        const int s = min(s0 + si, nS - 1);
                        dx = __fadd_rn(dx, __ldg(a.obias + (g * nS + s) * 2));
                        dy = __fadd_rn(dy, __ldg(a.obias + (g * nS + s) * 2 + 1));
                    }
                    if (!a.coords_final) {
                        dx = __fadd_rn(__fmul_rn(a.offset_scale, tanhf(dx)), ftx);
                        dy = __fadd_rn(__fmul_rn(a.offset_scale, tanhf(dy)), fty);
                    }
                    const TapGeom tg = sampler_geom(dx, dy, lxv, lyv, a.H, a.W, a.fe);
                    const char* north0 = reinterpret_cast<const char*>(featp) + (int64_t)tg.o_nw * 4;
                    const char* south0 = north0 + row_bytes;
                    // a.zero == 0 at run time: makes the multiplicand sample-variant, otherwise ptxas hoists
                    // plane_bytes * cc out of the loop and pays a 64-bit add (2 instructions) per tap row
                    // (two distinct multiplicands: a shared product would again be split off into adds)
                    const int pb_n = plane_bytes + tg.o_nw * a.zero, pb_s = plane_bytes + (tg.o_nw + 1) * a.zero;
#pragma unroll
                    for (int cc = 0; cc < GC; ++cc) xs[si][cc] = sample_plane_rows(north0, south0, cc, pb_n, pb_s, tg);
#pragma unroll
                    for (int hp = 0; hp < HP; ++hp) {
                        f32x2 d2 = 0ull;
#pragma unroll
                        for (int cc = 0; cc < GC; ++cc) d2 = fma2s(kt2[hp][cc], xs[si][cc], d2);
                        float d0, d1;
                        unpack2(d2, d0, d1);
                        red[((si * NG + 2 * hp) * NG + g) * 32 + lane] = d0;
                        red[((si * NG + 2 * hp + 1) * NG + g) * 32 + lane] = d1;
                    }
                }
                if (s0 + SC < nS) load_offs(s0 + SC);
                const int par = chunk & 1;
                ++chunk;
                __syncthreads();
                // ---- C: head g: logits of this chunk, softmax state ----
                {
                    float l[SC];
                    float lmax = -INFINITY;
#pragma unroll
                    for (int si = 0; si < SC; ++si) {
                        float acc = red[((si * NG + g) * NG + 0) * 32 + lane];
#pragma unroll
                        for (int gg = 1; gg < NG; ++gg) acc += red[((si * NG + g) * NG + gg) * 32 + lane];
                        l[si] = acc + cb;
                        const bool valid = s0 + si < nS;
                        if (valid) {
                            lmax = fmaxf(lmax, l[si]);
                            if (a.score[f] != nullptr && live)
                                a.score[f][(((int64_t)b * NG + g) * nS + (s0 + si)) * HW + pix] = l[si];
                        }
                    }
                    float alpha = 1.0f;
                    if (lmax - m_ref > 30.0f) {          // lazy re-base of the softmax shift (first chunk: m_ref = -inf)
                        alpha = expf(m_ref - lmax);       // exp(-inf) = 0 on the first chunk
                        m_ref = lmax;
                        sum *= alpha;
                        rebase[par] = 1;                  // rare: tells phase D to read the alphas
                    }
                    abuf[((g >> 2) * 32 + lane) * 4 + (g & 3)] = alpha;
#pragma unroll
                    for (int si = 0; si < SC; ++si) {
                        const float p = (s0 + si < nS) ? expf(l[si] - m_ref) : 0.0f;
                        sum += p;
                        pbuf[((si * (NG / 4) + (g >> 2)) * 32 + lane) * 4 + (g & 3)] = p;
                    }
                }
                __syncthreads();
                // ---- D: weighted accumulation for all heads ----
                if (rebase[par] != 0) {       // CTA-uniform
#pragma unroll
                    for (int h4 = 0; h4 < NG / 4; ++h4) {
                        const float4 al = *reinterpret_cast<const float4*>(abuf + (h4 * 32 + lane) * 4);
                        const f32x2 a01 = pack2(al.x, al.y), a23 = pack2(al.z, al.w);
#pragma unroll
                        for (int cc = 0; cc < GC; ++cc) {
                            yv2[2 * h4][cc] = mul2(yv2[2 * h4][cc], a01);
                            yv2[2 * h4 + 1][cc] = mul2(yv2[2 * h4 + 1][cc], a23);
                        }
                    }
                }
                if (threadIdx.x == 0) rebase[par ^ 1] = 0;   // last read two barriers ago, next written after the next barrier
#pragma unroll
                for (int si = 0; si < SC; ++si) {
#pragma unroll
                    for (int h4 = 0; h4 < NG / 4; ++h4) {
                        const float4 pv = *reinterpret_cast<const float4*>(pbuf + ((si * (NG / 4) + h4) * 32 + lane) * 4);
                        const f32x2 p01 = pack2(pv.x, pv.y), p23 = pack2(pv.z, pv.w);
#pragma unroll
                        for (int cc = 0; cc < GC; ++cc) {
                            yv2[2 * h4][cc] = fma2s(p01, xs[si][cc], yv2[2 * h4][cc]);
                            yv2[2 * h4 + 1][cc] = fma2s(p23, xs[si][cc], yv2[2 * h4 + 1][cc]);
                        }
                    }
                }
            }
        }
        float yv[NG][GC];
#pragma unroll
        for (int hp = 0; hp < HP; ++hp)
#pragma unroll
            for (int cc = 0; cc < GC; ++cc) unpack2(yv2[hp][cc], yv[2 * hp][cc], yv[2 * hp + 1][cc]);

        // ---- epilogue: out_h = Wv_h y_h / sum_h + bv_h; y_h is spread over the NG warps ----
        const float inv = 1.0f / sum;
        float* epi = uni;   // [hh][j][g][lane]
#pragma unroll
        for (int r = 0; r < NG / Cfg::EPI_H; ++r) {
            __syncthreads();   // red / previous round fully read
#pragma unroll
            for (int hh = 0; hh < Cfg::EPI_H; ++hh) {
                const int h = r * Cfg::EPI_H + hh;
#pragma unroll
                for (int j = 0; j < GC; ++j) {
                    const float* wr = WvT_s + (g * C + h * GC + j) * GCP;
                    float d = 0.0f;
#pragma unroll
                    for (int c4 = 0; c4 < GCP; c4 += 4) {
                        const float4 w = *reinterpret_cast<const float4*>(wr + c4);
                        d = fmaf(w.x, yv[h][c4], d);
                        if (c4 + 1 < GC) d = fmaf(w.y, yv[h][c4 + 1], d);
                        if (c4 + 2 < GC) d = fmaf(w.z, yv[h][c4 + 2], d);
                        if (c4 + 3 < GC) d = fmaf(w.w, yv[h][c4 + 3], d);
                    }
                    epi[((hh * GC + j) * NG + g) * 32 + lane] = d;
                }
            }
            __syncthreads();
            if (g / Cfg::EPI_H == r) {
                const int hh = g - r * Cfg::EPI_H;
                const int64_t o_sp = (a.qo_cl & 2) ? C : 1, o_sc = (a.qo_cl & 2) ? 1 : HW;
                float* op = a.out + (int64_t)b * C * HW + pix * o_sp + (g * GC) * o_sc;
#pragma unroll
                for (int j = 0; j < GC; ++j) {
                    float acc = epi[((hh * GC + j) * NG + 0) * 32 + lane];
#pragma unroll
                    for (int gg = 1; gg < NG; ++gg) acc += epi[((hh * GC + j) * NG + gg) * 32 + lane];
                    if (live) st_stream(op + (int64_t)j * o_sc, fmaf(acc, inv, __ldg(a.bv + g * GC + j)));
                }
            }
        }
    }
}


int launch_softmax_samples(float* sc, int64_t R, int n, int64_t HW, cudaStream_t st);

template <int GC, int NG>
static int launch_dattn(DattnArgs a, cudaStream_t st) {
    using Cfg = DattnCfg<GC, NG>;
    const size_t smem = (size_t)Cfg::SMEM_FLOATS * sizeof(float);
    if (a.offs_cl) VFI_CUDA(cudaFuncSetAttribute(dattn_kernel<GC, NG, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else VFI_CUDA(cudaFuncSetAttribute(dattn_kernel<GC, NG, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // warp tile: 32x1 pixels of a row (fewest cache lines per tap request); narrow images use 16x2 / 8x4
    a.tw_log2 = a.W >= 32 ? 5 : (a.W >= 16 ? 4 : 3);
    static const int tw_env = [] {                       // VFI_DATTN_TW_LOG2=3|4|5 (profiling), read once
        const char* env = getenv("VFI_DATTN_TW_LOG2");
        return (env && env[0] >= '3' && env[0] <= '5') ? env[0] - '0' : 0;
    }();
    if (tw_env) a.tw_log2 = tw_env;
    a.tiles_x = (int)ceil_div(a.W, 1 << a.tw_log2);
    a.tiles_y = (int)ceil_div(a.H, 32 >> a.tw_log2);
    const int64_t tiles = (int64_t)a.B * a.tiles_x * a.tiles_y;
    const int blocks = (int)min64(tiles, (int64_t)kNumSMs * Cfg::MIN_BLOCKS);
    if (a.offs_cl) dattn_kernel<GC, NG, true><<<blocks, Cfg::THREADS, smem, st>>>(a);
    else dattn_kernel<GC, NG, false><<<blocks, Cfg::THREADS, smem, st>>>(a);
    VFI_CHECK_LAUNCH("dattn_kernel");
    return VFI_OK;
}

}  // namespace vfi

extern "C" int vfi_dattn_win_covers(int C, int nG, int nS, float offset_scale, int coords_final, int offs_channels_last) {
    return vfi::dattn_win_covers(C, nG, nS, offset_scale, coords_final, offs_channels_last) ? 1 : 0;
}

extern "C" int vfi_dattn_f32(const float* q, const float* feat0, const float* feat1, const float* offs0, const float* offs1,
                             const float* ft0, const float* ft1, const int64_t ft_strides[4], const float* Wq, const float* bq,
                             const float* Wk, const float* bk, const float* Wv, const float* bv, const float* lin_x,
                             const float* lin_y, float* out, float* score0, float* score1, int B, int C, int H, int W, int nG,
                             int nS, int nH, float offset_scale, int coords_final, const float* offs_bias, int offs_channels_last,
                             int q_channels_last, int out_channels_last, int feat_channels_last, const int* frame_ext,
                             void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(q && feat0 && feat1 && offs0 && offs1 && Wq && bq && Wk && bk && Wv && bv && lin_x && lin_y && out,
                "vfi_dattn_f32: null pointer");
    VFI_REQUIRE(coords_final || (ft0 && ft1 && ft_strides), "vfi_dattn_f32: raw offsets need ft0/ft1 and their strides");
    VFI_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1 && nG >= 1 && nS >= 1, "vfi_dattn_f32: bad shape");
    VFI_REQUIRE((score0 == nullptr) == (score1 == nullptr), "vfi_dattn_f32: score0/score1 must both be given or both NULL");
    if (nG != nH || C % nG != 0) VFI_UNSUPPORTED("vfi_dattn_f32: fused kernel needs nG == nH dividing C (got C=%d nG=%d nH=%d)", C, nG, nH);
    if ((int64_t)H * W * 4 * (C / nG) >= (int64_t)1 << 31) VFI_UNSUPPORTED("vfi_dattn_f32: a group's planes exceed 2 GiB (H=%d W=%d)", H, W);
    DattnArgs a;
    a.q = q;
    a.feat[0] = feat0; a.feat[1] = feat1;
    a.offs[0] = offs0; a.offs[1] = offs1;
    a.ft[0] = ft0; a.ft[1] = ft1;
    a.ft_sn = ft_strides ? ft_strides[0] : 0; a.ft_sc = ft_strides ? ft_strides[1] : 0;
    a.ft_sh = ft_strides ? ft_strides[2] : 0; a.ft_sw = ft_strides ? ft_strides[3] : 0;
    a.Wq = Wq; a.bq = bq; a.Wk = Wk; a.bk = bk; a.Wv = Wv; a.bv = bv;
    a.lin_x = lin_x; a.lin_y = lin_y;
    a.obias = offs_bias; a.offs_cl = offs_channels_last;
    a.qo_cl = (q_channels_last ? 1 : 0) | (out_channels_last ? 2 : 0);
    a.out = out; a.score[0] = score0; a.score[1] = score1;
    a.B = B; a.H = H; a.W = W; a.nS = nS;
    a.offset_scale = offset_scale;
    a.coords_final = coords_final;
    a.tw_log2 = 5; a.tiles_x = a.tiles_y = 0; a.zero = 0;
    a.fe = frame_ext_of(frame_ext, H, W);
    a.feat_cl = feat_channels_last ? 1 : 0;
    VFI_REQUIRE(frame_ext_ok(a.fe, H, W), "vfi_dattn_f32: the crop does not lie inside the frame");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int gc = C / nG;
    int rc;
    if (feat_channels_last) {
        // channels-last features: the shared-memory window kernel (dattn_win.cu); planar features: the gather kernel below
        rc = launch_dattn_win(a, C, nG, st);
        if (rc == VFI_ERR_UNSUPPORTED)
            VFI_UNSUPPORTED("vfi_dattn_f32: channels-last features are not covered for this shape (see vfi_dattn_win_covers)");
    } else {
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
    }
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
