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

#include "sampler.cuh"

namespace vfi {

struct DattnArgs {
    const float *q, *feat[2], *offs[2], *ft[2];
    int64_t ft_sn, ft_sc, ft_sh, ft_sw;
    const float *Wq, *bq, *Wk, *bk, *Wv, *bv, *lin_x, *lin_y;
    const float* obias;      // optional bias of the offset head (nG*nS*2), added before tanh
    int offs_cl;             // offsets are channels-last: (B, H, W, nG*nS*2)
    int qo_cl;               // bit 0: q is channels-last (B, H, W, C); bit 1: out is written channels-last
    float *out, *score[2];
    int B, H, W, nS;
    float offset_scale;
    int coords_final;
    int tw_log2;              // tile = (1 << tw_log2) x (32 >> tw_log2) pixels per warp
    int tiles_x, tiles_y;
    int zero;                 // always 0 (see the sampling loop)
};

constexpr int cmax(int a, int b) { return a > b ? a : b; }
constexpr int cmin(int a, int b) { return a < b ? a : b; }

template <int GC, int NG>
struct DattnCfg {
    static constexpr int C = GC * NG;
    static constexpr int C4 = (C + 3) & ~3;       // Wq row stride (16 B aligned rows: broadcast LDS.128)
    static constexpr int GCP = (GC + 3) & ~3;     // padded channel block
    static constexpr int SC = GC > 12 ? 2 : 4;    // slots per chunk (register budget: xs[SC][GC])
    static constexpr int THREADS = NG * 32;
    static constexpr int WQ = C * C4;
    static constexpr int WT = NG * C * GCP;       // Wk / Wv re-blocked [g][row][GCP]
    static constexpr int RED = SC * NG * NG * 32;
    static constexpr int EPI_CAP = 10240;         // floats (40 KB) for one epilogue round
    static constexpr int EPI_H_RAW = cmax(1, cmin(NG, EPI_CAP / (GC * NG * 32)));
    static constexpr int EPI_H = EPI_H_RAW >= 8 ? 8 : (EPI_H_RAW >= 4 ? 4 : (EPI_H_RAW >= 2 ? 2 : 1));
    static constexpr int EPI = EPI_H * GC * NG * 32;
    static constexpr int UNI = cmax(cmax(RED, 2 * C * 32), EPI);
    static constexpr int PB = NG * SC * 32;
    static constexpr int AB = NG * 32;
    static constexpr int SMEM_FLOATS = WQ + 2 * WT + UNI + PB + AB + 4;   // + rebase flags
    static constexpr int MIN_BLOCKS = NG <= 4 ? 2 : 1;
};

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
                    const TapGeom tg = sampler_geom(dx, dy, lxv, lyv, a.H, a.W);
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


// ---- v5: one barrier per chunk ----------------------------------------------------------------------------------
// Same mapping and algebra as dattn_kernel.  There, every chunk is A | barrier | C | barrier | D and the 250
// instructions of C sit alone between two barriers (18 % of the stall samples, profiles/r1r_dattn_ffma2_ncu_full.txt).
// Here the chunk is half as long (SC/2 slots), the sampled values, the partial logits and the softmax weights are
// double buffered, and the weighted accumulation of chunk c-1 is issued AFTER the barrier of chunk c:
//     A(c) | barrier | C(c), D(c-1), A(c+1) | barrier | C(c+1), D(c), A(c+2) | ...
// so a warp meets the others once per chunk and the gathers of the next chunk are in flight under the FMAs of D.
template <int GC, int NG, bool OCL>
__global__ void __launch_bounds__(DattnCfg<GC, NG>::THREADS, DattnCfg<GC, NG>::MIN_BLOCKS) dattn_pipe_kernel(DattnArgs a) {
    using Cfg = DattnCfg<GC, NG>;
    constexpr int C = Cfg::C, C4 = Cfg::C4, GCP = Cfg::GCP, T = Cfg::THREADS;
    constexpr int SC = Cfg::SC >= 2 ? Cfg::SC / 2 : 1;       // slots per chunk
    constexpr int HP = NG / 2;
    static_assert(NG % 4 == 0, "head pairs / quads");
    extern __shared__ __align__(16) float smem[];
    float* Wq_s = smem;
    float* WkT_s = Wq_s + Cfg::WQ;
    float* WvT_s = WkT_s + Cfg::WT;
    float* uni = WvT_s + Cfg::WT;             // q_s/qp_s | red[2] | epilogue exchange
    float* pbuf = uni + Cfg::UNI;             // [2][si][h/4][lane][h%4]
    float* abuf = pbuf + Cfg::PB;             // [2][h/4][lane][h%4]   (SMEM_FLOATS + AB floats, see launch)
    constexpr int RED1 = SC * NG * NG * 32, PB1 = SC * NG * 32, AB1 = NG * 32;
    static_assert(2 * RED1 <= Cfg::UNI && 2 * PB1 <= Cfg::PB, "double buffers fit the single-buffer regions");

    for (int i = threadIdx.x; i < C * C4; i += T) {
        const int r = i / C4, c = i - r * C4;
        Wq_s[i] = c < C ? __ldg(a.Wq + r * C + c) : 0.0f;
    }
    for (int i = threadIdx.x; i < NG * C * GCP; i += T) {
        const int cc = i % GCP, r = (i / GCP) % C, gg = i / (GCP * C);
        const bool ok = cc < GC;
        WkT_s[i] = ok ? __ldg(a.Wk + r * C + gg * GC + cc) : 0.0f;
        WvT_s[i] = ok ? __ldg(a.Wv + r * C + gg * GC + cc) : 0.0f;
    }

    const int lane = threadIdx.x & 31;
    const int g = threadIdx.x >> 5;
    const int64_t HW = (int64_t)a.H * a.W;
    const float scale = rsqrtf((float)GC);
    const int nS = a.nS;
    const int lx = lane & ((1 << a.tw_log2) - 1), ly = lane >> a.tw_log2;
    const int tiles_per_img = a.tiles_x * a.tiles_y;
    const int64_t n_tiles = (int64_t)a.B * tiles_per_img;
    float* q_s = uni;
    float* qp_s = uni + C * 32;
    const int cpf = (nS + SC - 1) / SC, nchunks = 2 * cpf;     // chunks per frame; the chunk index runs over both frames
    const int nOff = NG * nS * 2;
    const int plane_bytes = (int)(HW * 4), row_bytes = a.W * 4;

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int b = (int)(tile / tiles_per_img);
        const int trem = (int)(tile - (int64_t)b * tiles_per_img);
        const int ty = trem / a.tiles_x, tx = trem - ty * a.tiles_x;
        const int x_raw = (tx << a.tw_log2) + lx, y_raw = ty * (32 >> a.tw_log2) + ly;
        const bool live = x_raw < a.W && y_raw < a.H;
        const int x = min(x_raw, a.W - 1), y = min(y_raw, a.H - 1);
        const int64_t pix = (int64_t)y * a.W + x;

        __syncthreads();
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
        f32x2 kt2[HP][GC];
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
        __syncthreads();   // q_s / qp_s (aliased with red) fully consumed

        f32x2 yv2[HP][GC];
#pragma unroll
        for (int hp = 0; hp < HP; ++hp)
#pragma unroll
            for (int cc = 0; cc < GC; ++cc) yv2[hp][cc] = 0ull;
        float m_ref = -INFINITY, sum = 0.0f;
        const float lxv = __ldg(a.lin_x + x), lyv = __ldg(a.lin_y + y);
        float ftx[2] = {0.0f, 0.0f}, fty[2] = {0.0f, 0.0f};
        if (!a.coords_final) {
#pragma unroll
            for (int f = 0; f < 2; ++f) {
                const float* fp = a.ft[f] + b * a.ft_sn + y * a.ft_sh + x * a.ft_sw;
                ftx[f] = __ldg(fp);
                fty[f] = __ldg(fp + a.ft_sc);
            }
        }
        float odx[SC], ody[SC];
        auto load_offs = [&](int c) {
            const int f = c >= cpf ? 1 : 0, s_first = (c - f * cpf) * SC;
            if (OCL) {
                const float* offcl = a.offs[f] + ((int64_t)b * HW + pix) * nOff + (int64_t)g * nS * 2;
                const bool vec = (SC % 2 == 0) && (nS % SC == 0) && ((reinterpret_cast<uintptr_t>(offcl) & 15) == 0);
                if (vec) {
#pragma unroll
                    for (int k = 0; k < SC / 2; ++k) {
                        const float4 v = ld_stream4(offcl + 2 * s_first + 4 * k);
                        odx[2 * k] = v.x; ody[2 * k] = v.y;
                        if (2 * k + 1 < SC) { odx[2 * k + 1] = v.z; ody[2 * k + 1] = v.w; }
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
                const float* offp = a.offs[f] + ((int64_t)b * NG * nS + (int64_t)g * nS) * 2 * HW + pix;
#pragma unroll
                for (int si = 0; si < SC; ++si) {
                    const int s = min(s_first + si, nS - 1);
                    odx[si] = ld_stream(offp + (int64_t)(2 * s) * HW);
                    ody[si] = ld_stream(offp + (int64_t)(2 * s + 1) * HW);
                }
            }
        };
        // A: sample chunk c into xs, partial logits of all heads -> red[c & 1]; then prefetch the next chunk's offsets
        auto phase_a = [&](float (&xs)[SC][GC], int c) {
            const int f = c >= cpf ? 1 : 0, s0 = (c - f * cpf) * SC;
            const float* featp = a.feat[f] + ((int64_t)b * C + g * GC) * HW;
            float* red = uni + (c & 1) * RED1;
            const float fx = f ? ftx[1] : ftx[0], fy = f ? fty[1] : fty[0];
#pragma unroll
            for (int si = 0; si < SC; ++si) {
                float dx = odx[si], dy = ody[si];
                if (a.obias != nullptr) {
                    const int s = min(s0 + si, nS - 1);
                    dx = __fadd_rn(dx, __ldg(a.obias + (g * nS + s) * 2));
                    dy = __fadd_rn(dy, __ldg(a.obias + (g * nS + s) * 2 + 1));
                }
                if (!a.coords_final) {
                    dx = __fadd_rn(__fmul_rn(a.offset_scale, tanhf(dx)), fx);
                    dy = __fadd_rn(__fmul_rn(a.offset_scale, tanhf(dy)), fy);
                }
                const TapGeom tg = sampler_geom(dx, dy, lxv, lyv, a.H, a.W);
                const char* north0 = reinterpret_cast<const char*>(featp) + (int64_t)tg.o_nw * 4;
                const char* south0 = north0 + row_bytes;
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
            if (c + 1 < nchunks) load_offs(c + 1);
        };
        // C: head g: logits of chunk c, softmax state, weights -> pbuf[c & 1], re-base factor -> abuf[c & 1]
        auto phase_c = [&](int c) {
            const int f = c >= cpf ? 1 : 0, s0 = (c - f * cpf) * SC, par = c & 1;
            const float* red = uni + par * RED1;
            float l[SC];
            float lmax = -INFINITY;
#pragma unroll
            for (int si = 0; si < SC; ++si) {
                float acc = red[((si * NG + g) * NG + 0) * 32 + lane];
#pragma unroll
                for (int gg = 1; gg < NG; ++gg) acc += red[((si * NG + g) * NG + gg) * 32 + lane];
                l[si] = acc + cb;
                if (s0 + si < nS) {
                    lmax = fmaxf(lmax, l[si]);
                    if (a.score[f] != nullptr && live) a.score[f][(((int64_t)b * NG + g) * nS + (s0 + si)) * HW + pix] = l[si];
                }
            }
            float alpha = 1.0f;
            if (lmax - m_ref > 30.0f) {
                alpha = expf(m_ref - lmax);
                m_ref = lmax;
                sum *= alpha;
            }
            abuf[par * AB1 + ((g >> 2) * 32 + lane) * 4 + (g & 3)] = alpha;
#pragma unroll
            for (int si = 0; si < SC; ++si) {
                const float p = (s0 + si < nS) ? expf(l[si] - m_ref) : 0.0f;
                sum += p;
                pbuf[par * PB1 + ((si * (NG / 4) + (g >> 2)) * 32 + lane) * 4 + (g & 3)] = p;
            }
        };
        // D: weighted accumulation of a chunk whose weights were published before the last barrier
        auto phase_d = [&](const float (&xs)[SC][GC], int par) {
            float4 al[NG / 4];
            bool any = false;
#pragma unroll
            for (int h4 = 0; h4 < NG / 4; ++h4) {
                al[h4] = *reinterpret_cast<const float4*>(abuf + par * AB1 + (h4 * 32 + lane) * 4);
                any = any || al[h4].x != 1.0f || al[h4].y != 1.0f || al[h4].z != 1.0f || al[h4].w != 1.0f;
            }
            if (any) {     // rare (the running max grew by > 30)
#pragma unroll
                for (int h4 = 0; h4 < NG / 4; ++h4) {
                    const f32x2 a01 = pack2(al[h4].x, al[h4].y), a23 = pack2(al[h4].z, al[h4].w);
#pragma unroll
                    for (int cc = 0; cc < GC; ++cc) {
                        yv2[2 * h4][cc] = mul2(yv2[2 * h4][cc], a01);
                        yv2[2 * h4 + 1][cc] = mul2(yv2[2 * h4 + 1][cc], a23);
                    }
                }
            }
#pragma unroll
            for (int si = 0; si < SC; ++si) {
#pragma unroll
                for (int h4 = 0; h4 < NG / 4; ++h4) {
                    const float4 pv = *reinterpret_cast<const float4*>(pbuf + par * PB1 + ((si * (NG / 4) + h4) * 32 + lane) * 4);
                    const f32x2 p01 = pack2(pv.x, pv.y), p23 = pack2(pv.z, pv.w);
#pragma unroll
                    for (int cc = 0; cc < GC; ++cc) {
                        yv2[2 * h4][cc] = fma2s(p01, xs[si][cc], yv2[2 * h4][cc]);
                        yv2[2 * h4 + 1][cc] = fma2s(p23, xs[si][cc], yv2[2 * h4 + 1][cc]);
                    }
                }
            }
        };

        float xs_a[SC][GC], xs_b[SC][GC];
        load_offs(0);
        for (int c = 0; c < nchunks; c += 2) {          // nchunks is even
            phase_a(xs_a, c);
            __syncthreads();
            phase_c(c);
            if (c > 0) phase_d(xs_b, 1);
            phase_a(xs_b, c + 1);
            __syncthreads();
            phase_c(c + 1);
            phase_d(xs_a, 0);
        }
        __syncthreads();
        phase_d(xs_b, 1);

        float yv[NG][GC];
#pragma unroll
        for (int hp = 0; hp < HP; ++hp)
#pragma unroll
            for (int cc = 0; cc < GC; ++cc) unpack2(yv2[hp][cc], yv[2 * hp][cc], yv[2 * hp + 1][cc]);

        const float inv = 1.0f / sum;
        float* epi = uni;
#pragma unroll
        for (int r = 0; r < NG / Cfg::EPI_H; ++r) {
            __syncthreads();
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

template <int GC, int NG>
static int launch_dattn_pipe(DattnArgs a, cudaStream_t st) {
    using Cfg = DattnCfg<GC, NG>;
    const size_t smem = (size_t)(Cfg::SMEM_FLOATS + Cfg::AB) * sizeof(float);
    if (a.offs_cl) VFI_CUDA(cudaFuncSetAttribute(dattn_pipe_kernel<GC, NG, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else VFI_CUDA(cudaFuncSetAttribute(dattn_pipe_kernel<GC, NG, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    a.tw_log2 = a.W >= 32 ? 5 : (a.W >= 16 ? 4 : 3);
    const char* env = getenv("VFI_DATTN_TW_LOG2");
    if (env && env[0] >= '3' && env[0] <= '5') a.tw_log2 = env[0] - '0';
    a.tiles_x = (int)ceil_div(a.W, 1 << a.tw_log2);
    a.tiles_y = (int)ceil_div(a.H, 32 >> a.tw_log2);
    const int64_t tiles = (int64_t)a.B * a.tiles_x * a.tiles_y;
    const int blocks = (int)min64(tiles, (int64_t)kNumSMs * Cfg::MIN_BLOCKS);
    if (a.offs_cl) dattn_pipe_kernel<GC, NG, true><<<blocks, Cfg::THREADS, smem, st>>>(a);
    else dattn_pipe_kernel<GC, NG, false><<<blocks, Cfg::THREADS, smem, st>>>(a);
    VFI_CHECK_LAUNCH("dattn_pipe_kernel");
    return VFI_OK;
}

// ---- v4: channel-split mapping -------------------------------------------------------------------------------
// v2/v3 hold kt[NG][GC] and yv[NG][GC] per thread (144 registers at C = 72): 255 registers, 8 warps per SM, and
// the kernel is latency bound (ncu: issue slots 41-49 % busy, stalls spread over wait / scoreboard / barrier).
// Here a warp is 16 pixels x 2 channel halves of one group: lane = (pixel, half), each lane samples and
// accumulates only GCW = ceil(GC / 2) channels, so the per-thread state halves (<= 128 registers) and 16 warps fit
// on an SM.  The two halves of a pixel sit 16 lanes apart: the partial logits and the partial Wv products are
// combined with one shfl_xor instead of a second set of shared-memory partials; phase C splits the SC slots of a
// chunk between the halves and shares the running max the same way.  Everything else (algebra, lazy re-base,
// shared-memory layouts of the partial logits and softmax weights) is as in dattn_kernel above.
template <int GC, int NG>
struct DattnSplitCfg {
    static constexpr int C = GC * NG;
    static constexpr int C4 = (C + 3) & ~3;
    static constexpr int GCW = (GC + 1) / 2;           // channels per lane
    static constexpr int GCWP = (GCW + 3) & ~3;
    static constexpr int SC = GCW > 6 ? 2 : 4;         // slots per chunk (even: the halves split them in phase C)
    static constexpr int THREADS = NG * 64;
    static constexpr int WQ = C * C4;
    static constexpr int WT = NG * 2 * C * GCWP;       // Wk / Wv re-blocked [g][half][row][GCWP]
    static constexpr int RED = SC * NG * NG * 32;
    static constexpr int EPI_CAP = 10240;
    static constexpr int EPI_H_RAW = cmax(1, cmin(NG, EPI_CAP / (GC * NG * 32)));
    static constexpr int EPI_H = EPI_H_RAW >= 8 ? 8 : (EPI_H_RAW >= 4 ? 4 : (EPI_H_RAW >= 2 ? 2 : 1));
    static constexpr int EPI = EPI_H * GC * NG * 32;
    static constexpr int UNI = cmax(cmax(RED, 2 * C * 32), EPI);
    static constexpr int PB = NG * SC * 32;
    static constexpr int AB = NG * 32;
    static constexpr int SMEM_FLOATS = WQ + 2 * WT + UNI + PB + AB + 4;
    static constexpr int MIN_BLOCKS = NG <= 4 ? 2 : 1;
};

template <int GC, int NG, bool OCL>
__global__ void __launch_bounds__(DattnSplitCfg<GC, NG>::THREADS, DattnSplitCfg<GC, NG>::MIN_BLOCKS)
dattn_split_kernel(DattnArgs a) {
    using Cfg = DattnSplitCfg<GC, NG>;
    constexpr int C = Cfg::C, C4 = Cfg::C4, GCW = Cfg::GCW, GCWP = Cfg::GCWP, SC = Cfg::SC, T = Cfg::THREADS;
    constexpr int HP = NG / 2, SH = SC / 2;
    static_assert(NG % 4 == 0 && SC % 2 == 0, "head quads / slot halves");
    extern __shared__ __align__(16) float smem[];
    float* Wq_s = smem;                       // [C][C4]
    float* WkT_s = Wq_s + Cfg::WQ;            // [g][half][r][GCWP] = Wk[r][g*GC + half*GCW + cc] (0 beyond the group)
    float* WvT_s = WkT_s + Cfg::WT;
    float* uni = WvT_s + Cfg::WT;
    float* pbuf = uni + Cfg::UNI;             // [si][h/4][pixel][h%4]
    float* abuf = pbuf + Cfg::PB;             // [h/4][pixel][h%4]
    int* rebase = reinterpret_cast<int*>(abuf + Cfg::AB);

    for (int i = threadIdx.x; i < C * C4; i += T) {
        const int r = i / C4, c = i - r * C4;
        Wq_s[i] = c < C ? __ldg(a.Wq + r * C + c) : 0.0f;
    }
    for (int i = threadIdx.x; i < Cfg::WT; i += T) {
        const int cc = i % GCWP, r = (i / GCWP) % C, hf = (i / (GCWP * C)) & 1, gg = i / (GCWP * C * 2);
        const int ch = hf * GCW + cc;
        const bool ok = cc < GCW && ch < GC;
        WkT_s[i] = ok ? __ldg(a.Wk + r * C + gg * GC + ch) : 0.0f;
        WvT_s[i] = ok ? __ldg(a.Wv + r * C + gg * GC + ch) : 0.0f;
    }

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = warp >> 1;                   // my group == my head
    const int half = lane >> 4;                // my channel half of the group
    const int pt = ((warp & 1) << 4) | (lane & 15);   // my pixel of the 32-pixel tile
    const int ncw = half == 0 ? GCW : GC - GCW;       // channels that exist in my half
    const int ch0 = g * GC + half * GCW;
    const int64_t HW = (int64_t)a.H * a.W;
    const float scale = rsqrtf((float)GC);
    const int nS = a.nS;
    const int lx = pt & ((1 << a.tw_log2) - 1), ly = pt >> a.tw_log2;
    const int tiles_per_img = a.tiles_x * a.tiles_y;
    const int64_t n_tiles = (int64_t)a.B * tiles_per_img;
    float* q_s = uni;                // [C][32]
    float* qp_s = uni + C * 32;      // [C][32]
    const float* WkT_me = WkT_s + (size_t)(g * 2 + half) * C * GCWP;
    const float* WvT_me = WvT_s + (size_t)(g * 2 + half) * C * GCWP;

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int b = (int)(tile / tiles_per_img);
        const int trem = (int)(tile - (int64_t)b * tiles_per_img);
        const int ty = trem / a.tiles_x, tx = trem - ty * a.tiles_x;
        const int x_raw = (tx << a.tw_log2) + lx, y_raw = ty * (32 >> a.tw_log2) + ly;
        const bool live = x_raw < a.W && y_raw < a.H;
        const int x = min(x_raw, a.W - 1), y = min(y_raw, a.H - 1);
        const int64_t pix = (int64_t)y * a.W + x;

        __syncthreads();
        {
            const int64_t q_sp = (a.qo_cl & 1) ? C : 1, q_sc = (a.qo_cl & 1) ? 1 : HW;
            const float* qpix = a.q + (int64_t)b * C * HW + pix * q_sp + ch0 * q_sc;
#pragma unroll
            for (int j = 0; j < GCW; ++j)
                if (j < ncw) q_s[(ch0 + j) * 32 + pt] = __ldg(qpix + (int64_t)j * q_sc);
        }
        __syncthreads();
        float cb;
        {
            float qp[GCW];
#pragma unroll
            for (int j = 0; j < GCW; ++j) qp[j] = j < ncw ? __ldg(a.bq + ch0 + j) : 0.0f;
#pragma unroll 2
            for (int c4 = 0; c4 < C4; c4 += 4) {
                float qv[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) qv[k] = (c4 + k < C) ? q_s[(c4 + k) * 32 + pt] : 0.0f;
#pragma unroll
                for (int j = 0; j < GCW; ++j) {
                    const float4 w = *reinterpret_cast<const float4*>(Wq_s + min(ch0 + j, C - 1) * C4 + c4);
                    qp[j] = fmaf(w.x, qv[0], qp[j]);
                    if (c4 + 1 < C) qp[j] = fmaf(w.y, qv[1], qp[j]);
                    if (c4 + 2 < C) qp[j] = fmaf(w.z, qv[2], qp[j]);
                    if (c4 + 3 < C) qp[j] = fmaf(w.w, qv[3], qp[j]);
                }
            }
            cb = 0.0f;
#pragma unroll
            for (int j = 0; j < GCW; ++j) {
                if (j < ncw) {
                    cb = fmaf(qp[j], __ldg(a.bk + ch0 + j), cb);
                    qp_s[(ch0 + j) * 32 + pt] = qp[j] * scale;
                }
            }
            cb += __shfl_xor_sync(0xffffffffu, cb, 16);
            cb *= scale;
        }
        __syncthreads();
        f32x2 kt2[HP][GCW];
#pragma unroll
        for (int hp = 0; hp < HP; ++hp) {
            float ka[2][GCWP];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int h = 2 * hp + e;
#pragma unroll
                for (int cc = 0; cc < GCWP; ++cc) ka[e][cc] = 0.0f;
#pragma unroll
                for (int j = 0; j < GC; ++j) {
                    const float qh = qp_s[(h * GC + j) * 32 + pt];
                    const float* wr = WkT_me + (h * GC + j) * GCWP;
#pragma unroll
                    for (int c4 = 0; c4 < GCWP; c4 += 4) {
                        const float4 w = *reinterpret_cast<const float4*>(wr + c4);
                        ka[e][c4] = fmaf(qh, w.x, ka[e][c4]);
                        if (c4 + 1 < GCW) ka[e][c4 + 1] = fmaf(qh, w.y, ka[e][c4 + 1]);
                        if (c4 + 2 < GCW) ka[e][c4 + 2] = fmaf(qh, w.z, ka[e][c4 + 2]);
                        if (c4 + 3 < GCW) ka[e][c4 + 3] = fmaf(qh, w.w, ka[e][c4 + 3]);
                    }
                }
            }
#pragma unroll
            for (int cc = 0; cc < GCW; ++cc) kt2[hp][cc] = pack2(ka[0][cc], ka[1][cc]);
        }
        if (threadIdx.x == 0) rebase[0] = rebase[1] = 0;
        __syncthreads();

        f32x2 yv2[HP][GCW];
#pragma unroll
        for (int hp = 0; hp < HP; ++hp)
#pragma unroll
            for (int cc = 0; cc < GCW; ++cc) yv2[hp][cc] = 0ull;
        int chunk = 0;
        float m_ref = -INFINITY, sum = 0.0f;    // sum: my slots only (the halves are added at the end)
        const float lxv = __ldg(a.lin_x + x), lyv = __ldg(a.lin_y + y);
        float* red = uni;   // [si][h][g][pixel]
        const int plane_bytes = (int)(HW * 4), row_bytes = a.W * 4;

        for (int f = 0; f < 2; ++f) {
            float ftx = 0.0f, fty = 0.0f;
            if (!a.coords_final) {
                const float* fp = a.ft[f] + b * a.ft_sn + y * a.ft_sh + x * a.ft_sw;
                ftx = __ldg(fp);
                fty = __ldg(fp + a.ft_sc);
            }
            const float* offp = a.offs[f] + ((int64_t)b * NG * nS + (int64_t)g * nS) * 2 * HW + pix;
            const float* featp = a.feat[f] + ((int64_t)b * C + ch0) * HW;
            float odx[SC], ody[SC];
            const int nOff = NG * nS * 2;
            const float* offcl = a.offs[f] + ((int64_t)b * HW + pix) * nOff + (int64_t)g * nS * 2;
            const bool cl_vec = OCL && (nS % SC == 0) && ((reinterpret_cast<uintptr_t>(offcl) & 15) == 0);
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
                float xs[SC][GCW];
                // ---- A: sample my channels of SC slots, partial logits of all heads, halves combined by shuffle ----
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
                    TapGeom tg = sampler_geom(dx, dy, lxv, lyv, a.H, a.W);
                    const char* north0 = reinterpret_cast<const char*>(featp) + (int64_t)tg.o_nw * 4;
                    const char* south0 = north0 + row_bytes;
                    const int pb_n = plane_bytes + tg.o_nw * a.zero, pb_s = plane_bytes + (tg.o_nw + 1) * a.zero;
#pragma unroll
                    for (int cc = 0; cc < GCW; ++cc) {
                        if (GCW * 2 != GC && cc == GCW - 1) {   // the odd channel exists in half 0 only
                            const bool has = half == 0;
                            tg.k_nw = tg.k_nw && has; tg.k_ne = tg.k_ne && has; tg.k_sw = tg.k_sw && has; tg.k_se = tg.k_se && has;
                        }
                        xs[si][cc] = sample_plane_rows(north0, south0, cc, pb_n, pb_s, tg);
                    }
                    f32x2 d2[HP];
#pragma unroll
                    for (int hp = 0; hp < HP; ++hp) {
                        d2[hp] = 0ull;
#pragma unroll
                        for (int cc = 0; cc < GCW; ++cc) d2[hp] = fma2s(kt2[hp][cc], xs[si][cc], d2[hp]);
                    }
                    // half 0 finishes heads [0, NG/2), half 1 heads [NG/2, NG): send the other half's pairs, keep mine
#pragma unroll
                    for (int k = 0; k < HP / 2; ++k) {
                        const f32x2 mine = half ? d2[HP / 2 + k] : d2[k];
                        const f32x2 send = half ? d2[k] : d2[HP / 2 + k];
                        float s_lo, s_hi, m_lo, m_hi;
                        unpack2(send, s_lo, s_hi);
                        unpack2(mine, m_lo, m_hi);
                        m_lo += __shfl_xor_sync(0xffffffffu, s_lo, 16);
                        m_hi += __shfl_xor_sync(0xffffffffu, s_hi, 16);
                        const int h = half * (NG / 2) + 2 * k;
                        red[((si * NG + h) * NG + g) * 32 + pt] = m_lo;
                        red[((si * NG + h + 1) * NG + g) * 32 + pt] = m_hi;
                    }
                }
                if (s0 + SC < nS) load_offs(s0 + SC);
                const int par = chunk & 1;
                ++chunk;
                __syncthreads();
                // ---- C: head g, my pixel, my SH slots of the chunk; the running max is shared with the other half ----
                {
                    float l[SH];
                    float lmax = -INFINITY;
#pragma unroll
                    for (int k = 0; k < SH; ++k) {
                        const int si = half * SH + k;
                        float acc = red[((si * NG + g) * NG + 0) * 32 + pt];
#pragma unroll
                        for (int gg = 1; gg < NG; ++gg) acc += red[((si * NG + g) * NG + gg) * 32 + pt];
                        l[k] = acc + cb;
                        if (s0 + si < nS) {
                            lmax = fmaxf(lmax, l[k]);
                            if (a.score[f] != nullptr && live)
                                a.score[f][(((int64_t)b * NG + g) * nS + (s0 + si)) * HW + pix] = l[k];
                        }
                    }
                    lmax = fmaxf(lmax, __shfl_xor_sync(0xffffffffu, lmax, 16));
                    float alpha = 1.0f;
                    if (lmax - m_ref > 30.0f) {
                        alpha = expf(m_ref - lmax);
                        m_ref = lmax;
                        sum *= alpha;
                        rebase[par] = 1;
                    }
                    if (half == 0) abuf[((g >> 2) * 32 + pt) * 4 + (g & 3)] = alpha;
#pragma unroll
                    for (int k = 0; k < SH; ++k) {
                        const int si = half * SH + k;
                        const float p = (s0 + si < nS) ? expf(l[k] - m_ref) : 0.0f;
                        sum += p;
                        pbuf[((si * (NG / 4) + (g >> 2)) * 32 + pt) * 4 + (g & 3)] = p;
                    }
                }
                __syncthreads();
                // ---- D: weighted accumulation of my channels for all heads ----
                if (rebase[par] != 0) {
#pragma unroll
                    for (int h4 = 0; h4 < NG / 4; ++h4) {
                        const float4 al = *reinterpret_cast<const float4*>(abuf + (h4 * 32 + pt) * 4);
                        const f32x2 a01 = pack2(al.x, al.y), a23 = pack2(al.z, al.w);
#pragma unroll
                        for (int cc = 0; cc < GCW; ++cc) {
                            yv2[2 * h4][cc] = mul2(yv2[2 * h4][cc], a01);
                            yv2[2 * h4 + 1][cc] = mul2(yv2[2 * h4 + 1][cc], a23);
                        }
                    }
                }
                if (threadIdx.x == 0) rebase[par ^ 1] = 0;
#pragma unroll
                for (int si = 0; si < SC; ++si) {
#pragma unroll
                    for (int h4 = 0; h4 < NG / 4; ++h4) {
                        const float4 pv = *reinterpret_cast<const float4*>(pbuf + ((si * (NG / 4) + h4) * 32 + pt) * 4);
                        const f32x2 p01 = pack2(pv.x, pv.y), p23 = pack2(pv.z, pv.w);
#pragma unroll
                        for (int cc = 0; cc < GCW; ++cc) {
                            yv2[2 * h4][cc] = fma2s(p01, xs[si][cc], yv2[2 * h4][cc]);
                            yv2[2 * h4 + 1][cc] = fma2s(p23, xs[si][cc], yv2[2 * h4 + 1][cc]);
                        }
                    }
                }
            }
        }
        float yv[NG][GCW];
#pragma unroll
        for (int hp = 0; hp < HP; ++hp)
#pragma unroll
            for (int cc = 0; cc < GCW; ++cc) unpack2(yv2[hp][cc], yv[2 * hp][cc], yv[2 * hp + 1][cc]);
        sum += __shfl_xor_sync(0xffffffffu, sum, 16);

        // ---- epilogue: out_h = Wv_h y_h / sum_h + bv_h; y_h is spread over NG groups x 2 halves ----
        const float inv = 1.0f / sum;
        float* epi = uni;   // [hh][j][g][pixel]
#pragma unroll
        for (int r = 0; r < NG / Cfg::EPI_H; ++r) {
            __syncthreads();
#pragma unroll
            for (int hh = 0; hh < Cfg::EPI_H; ++hh) {
                const int h = r * Cfg::EPI_H + hh;
#pragma unroll
                for (int j = 0; j < GC; ++j) {
                    const float* wr = WvT_me + (h * GC + j) * GCWP;
                    float d = 0.0f;
#pragma unroll
                    for (int c4 = 0; c4 < GCWP; c4 += 4) {
                        const float4 w = *reinterpret_cast<const float4*>(wr + c4);
                        d = fmaf(w.x, yv[h][c4], d);
                        if (c4 + 1 < GCW) d = fmaf(w.y, yv[h][c4 + 1], d);
                        if (c4 + 2 < GCW) d = fmaf(w.z, yv[h][c4 + 2], d);
                        if (c4 + 3 < GCW) d = fmaf(w.w, yv[h][c4 + 3], d);
                    }
                    d += __shfl_xor_sync(0xffffffffu, d, 16);
                    if ((j & 1) == half) epi[((hh * GC + j) * NG + g) * 32 + pt] = d;
                }
            }
            __syncthreads();
            if (g / Cfg::EPI_H == r) {
                const int hh = g - r * Cfg::EPI_H;
                const int64_t o_sp = (a.qo_cl & 2) ? C : 1, o_sc = (a.qo_cl & 2) ? 1 : HW;
                float* op = a.out + (int64_t)b * C * HW + pix * o_sp + (g * GC) * o_sc;
#pragma unroll
                for (int j = 0; j < GC; ++j) {
                    if ((j & 1) != half) continue;
                    float acc = epi[((hh * GC + j) * NG + 0) * 32 + pt];
#pragma unroll
                    for (int gg = 1; gg < NG; ++gg) acc += epi[((hh * GC + j) * NG + gg) * 32 + pt];
                    if (live) st_stream(op + (int64_t)j * o_sc, fmaf(acc, inv, __ldg(a.bv + g * GC + j)));
                }
            }
        }
    }
}

template <int GC, int NG>
static int launch_dattn_split(DattnArgs a, cudaStream_t st) {
    using Cfg = DattnSplitCfg<GC, NG>;
    const size_t smem = (size_t)Cfg::SMEM_FLOATS * sizeof(float);
    if (a.offs_cl) VFI_CUDA(cudaFuncSetAttribute(dattn_split_kernel<GC, NG, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else VFI_CUDA(cudaFuncSetAttribute(dattn_split_kernel<GC, NG, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    a.tw_log2 = a.W >= 32 ? 5 : (a.W >= 16 ? 4 : 3);
    const char* env = getenv("VFI_DATTN_TW_LOG2");
    if (env && env[0] >= '3' && env[0] <= '5') a.tw_log2 = env[0] - '0';
    a.tiles_x = (int)ceil_div(a.W, 1 << a.tw_log2);
    a.tiles_y = (int)ceil_div(a.H, 32 >> a.tw_log2);
    const int64_t tiles = (int64_t)a.B * a.tiles_x * a.tiles_y;
    const int blocks = (int)min64(tiles, (int64_t)kNumSMs * Cfg::MIN_BLOCKS);
    if (a.offs_cl) dattn_split_kernel<GC, NG, true><<<blocks, Cfg::THREADS, smem, st>>>(a);
    else dattn_split_kernel<GC, NG, false><<<blocks, Cfg::THREADS, smem, st>>>(a);
    VFI_CHECK_LAUNCH("dattn_split_kernel");
    return VFI_OK;
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
    const char* env = getenv("VFI_DATTN_TW_LOG2");
    if (env && env[0] >= '3' && env[0] <= '5') a.tw_log2 = env[0] - '0';
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

extern "C" int vfi_dattn_f32(const float* q, const float* feat0, const float* feat1, const float* offs0, const float* offs1,
                             const float* ft0, const float* ft1, const int64_t ft_strides[4], const float* Wq, const float* bq,
                             const float* Wk, const float* bk, const float* Wv, const float* bv, const float* lin_x,
                             const float* lin_y, float* out, float* score0, float* score1, int B, int C, int H, int W, int nG,
                             int nS, int nH, float offset_scale, int coords_final, const float* offs_bias, int offs_channels_last,
                             int q_channels_last, int out_channels_last, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(q && feat0 && feat1 && offs0 && offs1 && Wq && bq && Wk && bk && Wv && bv && lin_x && lin_y && out,
                "vfi_dattn_f32: null pointer");
    VFI_REQUIRE(coords_final || (ft0 && ft1 && ft_strides), "vfi_dattn_f32: raw offsets need ft0/ft1 and their strides");
    VFI_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1 && nG >= 1 && nS >= 1, "vfi_dattn_f32: bad shape");
    VFI_REQUIRE((score0 == nullptr) == (score1 == nullptr), "vfi_dattn_f32: score0/score1 must both be given or both NULL");
    if (nG != nH || C % nG != 0) VFI_UNSUPPORTED("vfi_dattn_f32: fused kernel needs nG == nH dividing C (got C=%d nG=%d nH=%d)", C, nG, nH);
    if ((int64_t)H * W * 4 * (C / nG) >= (int64_t)1 << 31) VFI_UNSUPPORTED("vfi_dattn_f32: a group's planes exceed 2 GiB (H=%d W=%d)", H, W);
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
    a.obias = offs_bias; a.offs_cl = offs_channels_last;
    a.qo_cl = (q_channels_last ? 1 : 0) | (out_channels_last ? 2 : 0);
    a.out = out; a.score[0] = score0; a.score[1] = score1;
    a.B = B; a.H = H; a.W = W; a.nS = nS;
    a.offset_scale = offset_scale;
    a.coords_final = coords_final;
    a.tw_log2 = 5; a.tiles_x = a.tiles_y = 0; a.zero = 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int gc = C / nG;
    int rc;
    // "split" = dattn_split_kernel (16 warps per SM).  Measured slower than the default (12.1 vs 8.3 ms at the 1/2-res
    // DAT shape, profiles/r1w_dattn_split_ncu_full.txt): the per-lane geometry / addressing / predicate work does not
    // shrink with the channel count, so the instruction count per pixel grows 1.67x and eats the occupancy gain.
    const char* algo_env = getenv("VFI_DATTN_ALGO");
    const bool split = algo_env && algo_env[0] == 's';
    const bool pipe = algo_env && algo_env[0] == 'p';      // "pipe" = dattn_pipe_kernel (one barrier per chunk)
#define VFI_DATTN_CASE(GC_, NG_) \
    if (gc == GC_ && nG == NG_) rc = (split && GC_ >= 4) ? launch_dattn_split<GC_, NG_>(a, st) : (pipe ? launch_dattn_pipe<GC_, NG_>(a, st) : launch_dattn<GC_, NG_>(a, st)); else
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
