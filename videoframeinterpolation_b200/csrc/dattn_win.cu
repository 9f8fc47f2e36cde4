// Fused deformable attention, shared-memory WINDOW form (sm_100a: TMA + mbarrier).
// Reference (unfused): modules/deformable_attn.py:81-101 (_get_ref_coords / _get_ref_feats) + 26-40 (Attn).
//
// Same algebra, same warp mapping (warp = offset group = head, lane = pixel of an 8x4 tile) and the same phases
// A / C / D per chunk of samples as dattn.cu.  What changes is where the taps come from.  dattn.cu gathers every tap
// from global memory (9 planar LDG.32 per tap row, 4.9 sectors per request, 13 % of them L1 misses: the kernel is
// latency bound at 41 % issue utilisation, profiles/r2d_dattn_ncu_full.txt).  The residual offsets are bounded,
//      x displacement = (offset_scale * tanh(.) + ft_x) / 4     (the x/(W-1)/2 quirk, modules/deformable_attn.py:94)
//      y displacement =  offset_scale * tanh(.) + ft_y
// so all 2 nS samples of all pixels of a tile fall into a small window around tile + median flow.  Per (tile, frame)
// ONE 4-D TMA box load brings that window of the channels-last feature map into shared memory:
//      box {C + 4, Ww, Hw, 1} of the (B, H, W, C) tensor -> win[Hw][Ww][REC],  REC = C + 4 floats
// (the 4 extra "channels" are out of bounds in the innermost dimension, TMA zero-fills them; they make the record
// stride 19 x 16 B, odd, so the 128-bit tap reads of 8 neighbouring pixels hit 8 different bank quads).  Positions
// outside the feature map are zero-filled by TMA as well, which IS grid_sample's padding_mode='zeros': the fast path
// has no bounds predicates.  A tap is three LDS.128 (the group's GC = 9 channels start at float 9g: the 12 floats
// from the 16-byte aligned address below cover them; the warp-uniform shift 9g & 3 picks the 9) against nine LDG.32.
// A chunk whose warp has a sample outside the window (flow discontinuities) takes predicated global loads instead,
// so the result never depends on the window choice.
//
// Window pipeline: one buffer.  Thread 0 issues the load of (tile, frame 1) right after the barrier that ends the
// last sampling phase of (tile, frame 0), and the load of (next tile, frame 0) after the last sampling phase of
// frame 1 -- that one lands under phases C / D, the epilogue and the next prologue.
#include <stdlib.h>

#include "dattn_common.cuh"
#include "tma.cuh"
#include "umma.cuh"

namespace vfi {

struct WinGeom {
    int Ww, Hw;      // window size (positions)
    int mx, my;      // window origin = tile origin + floor(median displacement) - (mx, my)
    int row_bytes;   // Ww * REC * 4
    uint32_t box_bytes;
};

constexpr int kWinTW = 8, kWinTH = 4;   // pixel tile of a CTA (lane = pixel)

template <int GC, int NG, int SC>
struct WinCfg {
    static constexpr int C = GC * NG;
    static constexpr int C4 = (C + 3) & ~3;
    static constexpr int GCP = (GC + 3) & ~3;
    static constexpr int REC4 = ((C4 / 4) & 1) ? C4 / 4 : C4 / 4 + 1;   // record length in 16-byte units, odd
    static constexpr int REC = REC4 * 4;
    static constexpr int LDV = (((GC * (NG - 1)) & 3) + GC + 3) / 4;     // LDS.128 per tap covering any group's channels
    static constexpr int THREADS = NG * 32;
    static constexpr int WQ = C * C4;
    static constexpr int WT = NG * C * GCP;
    static constexpr int RED = SC * NG * NG * 32;
    static constexpr int UNI0 = cmax(RED, 2 * C * 32);
    static constexpr int EPI_H_RAW = cmax(1, cmin(NG, UNI0 / (GC * NG * 32)));
    static constexpr int EPI_H = EPI_H_RAW >= 8 ? 8 : (EPI_H_RAW >= 4 ? 4 : (EPI_H_RAW >= 2 ? 2 : 1));
    static constexpr int EPI = EPI_H * GC * NG * 32;
    static constexpr int UNI = cmax(UNI0, EPI);
    static constexpr int PB = NG * SC * 32;
    static constexpr int AB = NG * 32;
    static constexpr int MISC = 16;                                     // rebase[2], org[3][2], mbarrier, pad
    static constexpr int FIXED_FLOATS = WQ + 2 * WT + UNI + PB + AB + MISC;
    static_assert(LDV * 4 <= 12, "tap = at most three 128-bit loads");
};

__device__ __forceinline__ float warp_median32(float v, int lane) {
    int rank = 0;
#pragma unroll 8
    for (int j = 0; j < 32; ++j) {
        const float o = __shfl_sync(0xffffffffu, v, j);
        rank += (o < v || (o == v && j < lane)) ? 1 : 0;
    }
    const unsigned m = __ballot_sync(0xffffffffu, rank == 15);
    return __shfl_sync(0xffffffffu, v, __ffs(m) - 1);
}

// window origin (crop coordinates) of (tile, frame f): tile origin + floor(median displacement) - margin
__device__ __forceinline__ void window_origin(const DattnArgs& a, const WinGeom& w, int f, int b, int tx, int ty, int lane,
                                              int& ox, int& oy) {
    const int x = min(tx * kWinTW + (lane & (kWinTW - 1)), a.W - 1), y = min(ty * kWinTH + (lane / kWinTW), a.H - 1);
    const float* fp = a.ft[f] + b * a.ft_sn + y * a.ft_sh + x * a.ft_sw;
    float fx = __ldg(fp), fy = __ldg(fp + a.ft_sc);
    fx = isfinite(fx) ? fminf(fmaxf(fx, -1.0e6f), 1.0e6f) : 0.0f;
    fy = isfinite(fy) ? fminf(fmaxf(fy, -1.0e6f), 1.0e6f) : 0.0f;
    const float mx = warp_median32(fx, lane), my = warp_median32(fy, lane);
    ox = tx * kWinTW + (int)floorf(mx * 0.25f) - w.mx;
    oy = ty * kWinTH + (int)floorf(my) - w.my;
}

__device__ __forceinline__ void lds_v2u64(uint32_t addr, f32x2& a, f32x2& b) {
    asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "r"(addr));
}

template <int GC, int NG, int SC>
__global__ void __launch_bounds__(WinCfg<GC, NG, SC>::THREADS, 1)
dattn_win_kernel(const __grid_constant__ CUtensorMap tm0, const __grid_constant__ CUtensorMap tm1, DattnArgs a, WinGeom w) {
    using Cfg = WinCfg<GC, NG, SC>;
    constexpr int C = Cfg::C, C4 = Cfg::C4, GCP = Cfg::GCP, T = Cfg::THREADS, REC = Cfg::REC, LDV = Cfg::LDV;
    constexpr int HP = NG / 2;
    constexpr int OL = 4;                 // samples per offset load (32 B of a channels-last offset row)
    static_assert(NG % 4 == 0 && OL % SC == 0, "head quads; chunks per offset load");
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t* smem_al = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    float* win = reinterpret_cast<float*>(smem_al);                     // [Hw][Ww][REC]
    float* Wq_s = win + (size_t)w.Ww * w.Hw * REC;                      // [C][C4]
    float* WkT_s = Wq_s + Cfg::WQ;                                      // [g][r][GCP] = Wk[r][g*GC + cc]
    float* WvT_s = WkT_s + Cfg::WT;
    float* uni = WvT_s + Cfg::WT;                                       // q_s/qp_s | red | epilogue exchange
    float* pbuf = uni + Cfg::UNI;                                       // [si][h/4][lane][h%4]
    float* abuf = pbuf + Cfg::PB;                                       // [h/4][lane][h%4]
    int* rebase = reinterpret_cast<int*>(abuf + Cfg::AB);               // [chunk parity]
    int* org = rebase + 2;                                              // [3][2]: (tile,f0), (tile,f1), (next tile,f0)
    uint64_t* wbar = reinterpret_cast<uint64_t*>(rebase + 8);           // 8-byte aligned: every region is a multiple of 16 B

    const int lane = threadIdx.x & 31;
    const int g = threadIdx.x >> 5;            // my group == my head
    const int64_t HW = (int64_t)a.H * a.W;
    const float scale = rsqrtf((float)GC);
    const int nS = a.nS;
    const int lx = lane & (kWinTW - 1), ly = lane / kWinTW;
    const int tiles_per_img = a.tiles_x * a.tiles_y;
    const int64_t n_tiles = (int64_t)a.B * tiles_per_img;

    if (threadIdx.x == 0) {
        mbar_init(wbar, 1);
        fence_mbar_init();
        tma_prefetch_desc(&tm0);
        tma_prefetch_desc(&tm1);
    }
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
    __syncthreads();                           // mbarrier initialised before anybody arms it
    // first window of this CTA
    if ((int64_t)blockIdx.x < n_tiles && g == 0) {
        const int b = (int)(blockIdx.x / tiles_per_img);
        const int trem = (int)(blockIdx.x - (int64_t)b * tiles_per_img);
        int ox, oy;
        window_origin(a, w, 0, b, trem % a.tiles_x, trem / a.tiles_x, lane, ox, oy);
        if (lane == 0) {
            mbar_arrive_expect_tx(wbar, w.box_bytes);
            tma_load_4d(win, &tm0, wbar, 0, ox, oy, b);
        }
    }
    uint32_t wphase = 0;

    float* q_s = uni;                // [C][32]
    float* qp_s = uni + C * 32;      // [C][32]
    const float wm1 = (float)(a.fe.Wg - 1), hm1 = (float)(a.fe.Hg - 1);
    const float half_h = (float)((a.fe.Hg - 1.0) / 2.0);
    const float xmax = (float)a.fe.Wg + 1.0f, ymax = (float)a.fe.Hg + 1.0f;
    const int gbase = (g * GC) & ~3, gsh = (g * GC) & 3;                // aligned channel base of my group, shift inside it
    const uint32_t win_g = smem_u32(win) + gbase * 4;

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int b = (int)(tile / tiles_per_img);
        const int trem = (int)(tile - (int64_t)b * tiles_per_img);
        const int ty = trem / a.tiles_x, tx = trem - ty * a.tiles_x;
        const int x_raw = tx * kWinTW + lx, y_raw = ty * kWinTH + ly;
        const bool live = x_raw < a.W && y_raw < a.H;
        const int x = min(x_raw, a.W - 1), y = min(y_raw, a.H - 1);
        const int64_t pix = (int64_t)y * a.W + x;
        const int64_t tile_next = tile + gridDim.x;
        const bool has_next = tile_next < n_tiles;
        const int b_next = has_next ? (int)(tile_next / tiles_per_img) : 0;

        __syncthreads();   // previous tile's epilogue reads done; org[] free
        // ---- prologue ----
        {
            const int64_t q_sp = (a.qo_cl & 1) ? C : 1, q_sc = (a.qo_cl & 1) ? 1 : HW;
            const float* qpix = a.q + (int64_t)b * C * HW + pix * q_sp + (g * GC) * q_sc;
#pragma unroll
            for (int j = 0; j < GC; ++j) q_s[(g * GC + j) * 32 + lane] = __ldg(qpix + (int64_t)j * q_sc);
        }
        if (g < 2) {
            int ox, oy;
            window_origin(a, w, g, b, tx, ty, lane, ox, oy);
            if (lane == 0) { org[2 * g] = ox; org[2 * g + 1] = oy; }
        } else if (g == 2 && has_next) {
            const int trn = (int)(tile_next - (int64_t)b_next * tiles_per_img);
            int ox, oy;
            window_origin(a, w, 0, b_next, trn % a.tiles_x, trn / a.tiles_x, lane, ox, oy);
            if (lane == 0) { org[4] = ox; org[5] = oy; }
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
                    const float4 wv = *reinterpret_cast<const float4*>(Wq_s + (g * GC + j) * C4 + c4);
                    qp[j] = fmaf(wv.x, qv[0], qp[j]);
                    if (c4 + 1 < C) qp[j] = fmaf(wv.y, qv[1], qp[j]);
                    if (c4 + 2 < C) qp[j] = fmaf(wv.z, qv[2], qp[j]);
                    if (c4 + 3 < C) qp[j] = fmaf(wv.w, qv[3], qp[j]);
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
                        const float4 wv = *reinterpret_cast<const float4*>(wr + c4);
                        ka[e][c4] = fmaf(qh, wv.x, ka[e][c4]);
                        if (c4 + 1 < GC) ka[e][c4 + 1] = fmaf(qh, wv.y, ka[e][c4 + 1]);
                        if (c4 + 2 < GC) ka[e][c4 + 2] = fmaf(qh, wv.z, ka[e][c4 + 2]);
                        if (c4 + 3 < GC) ka[e][c4 + 3] = fmaf(qh, wv.w, ka[e][c4 + 3]);
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

        for (int f = 0; f < 2; ++f) {
            const float* fp = a.ft[f] + b * a.ft_sn + y * a.ft_sh + x * a.ft_sw;
            const float ftx = __ldg(fp), fty = __ldg(fp + a.ft_sc);
            const int ox = org[2 * f], oy = org[2 * f + 1];
            const int nOff = NG * nS * 2;
            const float* offcl = a.offs[f] + ((int64_t)b * HW + pix) * nOff + (int64_t)g * nS * 2;
            const float* featg = a.feat[f] + (int64_t)b * HW * C + gbase;      // channels-last, my group's aligned base
            float od[2 * OL];    // (dx, dy) of OL samples, prefetched one load group ahead
            auto load_offs = [&](int s_first) {
#pragma unroll
                for (int k = 0; k < OL / 2; ++k) {
                    const float4 v = ld_stream4(offcl + 2 * s_first + 4 * k);
                    od[4 * k] = v.x; od[4 * k + 1] = v.y; od[4 * k + 2] = v.z; od[4 * k + 3] = v.w;
                }
            };
            load_offs(0);
            mbar_wait(wbar, wphase);       // window (tile, f) has landed
            wphase ^= 1;
            for (int s00 = 0; s00 < nS; s00 += OL) {
                float oc[2 * OL];
#pragma unroll
                for (int k = 0; k < 2 * OL; ++k) oc[k] = od[k];
                if (s00 + OL < nS) load_offs(s00 + OL);
#pragma unroll
                for (int cg = 0; cg < OL / SC; ++cg) {
                    const int s0 = s00 + cg * SC;
                    float xs[SC][GC];
                    // ---- A: sample + partial logits ----
#pragma unroll
                    for (int si = 0; si < SC; ++si) {
                        float dx = oc[2 * (cg * SC + si)], dy = oc[2 * (cg * SC + si) + 1];
                        if (a.obias != nullptr) {
                            dx = __fadd_rn(dx, __ldg(a.obias + (g * nS + s0 + si) * 2));
                            dy = __fadd_rn(dy, __ldg(a.obias + (g * nS + s0 + si) * 2 + 1));
                        }
                        dx = __fadd_rn(__fmul_rn(a.offset_scale, tanhf(dx)), ftx);
                        dy = __fadd_rn(__fmul_rn(a.offset_scale, tanhf(dy)), fty);
                        // coordinate replay of sampler_geom (sampler.cuh), frame coordinates
                        const float gx = __fadd_rn(lxv, __fmul_rn(__fdiv_rn(dx, wm1), 0.5f));
                        const float gy = __fadd_rn(lyv, __fdiv_rn(dy, half_h));
                        float ix = __fmul_rn(__fmul_rn(__fadd_rn(gx, 1.0f), 0.5f), wm1);
                        float iy = __fmul_rn(__fmul_rn(__fadd_rn(gy, 1.0f), 0.5f), hm1);
                        const bool finite = isfinite(ix) && isfinite(iy);
                        ix = fminf(fmaxf(finite ? ix : -2.0f, -2.0f), xmax);
                        iy = fminf(fmaxf(finite ? iy : -2.0f, -2.0f), ymax);
                        const float xf = floorf(ix), yf = floorf(iy);
                        const float tx_ = ix - xf, ty_ = iy - yf;
                        const float ex = 1.0f - tx_, ey = 1.0f - ty_;
                        const float w_nw = ey * ex, w_ne = ey * tx_, w_sw = ty_ * ex, w_se = ty_ * tx_;
                        const int x0 = (int)xf - a.fe.x0, y0 = (int)yf - a.fe.y0;      // crop coordinates
                        const int wx = x0 - ox, wy = y0 - oy;                          // window coordinates
                        const bool in_win = (unsigned)wx <= (unsigned)(w.Ww - 2) && (unsigned)wy <= (unsigned)(w.Hw - 2);
                        f32x2 v[2 * LDV];
                        if (__all_sync(0xffffffffu, in_win)) {
                            const uint32_t an = win_g + (uint32_t)(wy * w.Ww + wx) * (REC * 4);
                            const uint32_t as = an + w.row_bytes;
                            f32x2 t[2 * LDV];
#pragma unroll
                            for (int k = 0; k < LDV; ++k) lds_v2u64(an + 16 * k, t[2 * k], t[2 * k + 1]);
#pragma unroll
                            for (int k = 0; k < 2 * LDV; ++k) v[k] = mul2s(t[k], w_nw);
#pragma unroll
                            for (int k = 0; k < LDV; ++k) lds_v2u64(an + REC * 4 + 16 * k, t[2 * k], t[2 * k + 1]);
#pragma unroll
                            for (int k = 0; k < 2 * LDV; ++k) v[k] = fma2s(t[k], w_ne, v[k]);
#pragma unroll
                            for (int k = 0; k < LDV; ++k) lds_v2u64(as + 16 * k, t[2 * k], t[2 * k + 1]);
#pragma unroll
                            for (int k = 0; k < 2 * LDV; ++k) v[k] = fma2s(t[k], w_sw, v[k]);
#pragma unroll
                            for (int k = 0; k < LDV; ++k) lds_v2u64(as + REC * 4 + 16 * k, t[2 * k], t[2 * k + 1]);
#pragma unroll
                            for (int k = 0; k < 2 * LDV; ++k) v[k] = fma2s(t[k], w_se, v[k]);
                        } else {
                            // some lane of the warp samples outside the window: every lane gathers from global memory
                            const bool in_x0 = (unsigned)x0 < (unsigned)a.W, in_x1 = (unsigned)(x0 + 1) < (unsigned)a.W;
                            const bool in_y0 = (unsigned)y0 < (unsigned)a.H, in_y1 = (unsigned)(y0 + 1) < (unsigned)a.H;
                            const float4* pn = reinterpret_cast<const float4*>(featg + ((int64_t)y0 * a.W + x0) * C);
                            const float4* ps = pn + (int64_t)a.W * (C / 4);
                            const float4 z4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
                            for (int k = 0; k < LDV; ++k) {
                                const float4 tnw = (in_x0 && in_y0) ? __ldg(pn + k) : z4;
                                const float4 tne = (in_x1 && in_y0) ? __ldg(pn + C / 4 + k) : z4;
                                const float4 tsw = (in_x0 && in_y1) ? __ldg(ps + k) : z4;
                                const float4 tse = (in_x1 && in_y1) ? __ldg(ps + C / 4 + k) : z4;
                                f32x2 r0 = mul2s(pack2(tnw.x, tnw.y), w_nw), r1 = mul2s(pack2(tnw.z, tnw.w), w_nw);
                                r0 = fma2s(pack2(tne.x, tne.y), w_ne, r0); r1 = fma2s(pack2(tne.z, tne.w), w_ne, r1);
                                r0 = fma2s(pack2(tsw.x, tsw.y), w_sw, r0); r1 = fma2s(pack2(tsw.z, tsw.w), w_sw, r1);
                                r0 = fma2s(pack2(tse.x, tse.y), w_se, r0); r1 = fma2s(pack2(tse.z, tse.w), w_se, r1);
                                v[2 * k] = r0; v[2 * k + 1] = r1;
                            }
                        }
                        float vf[4 * LDV];
#pragma unroll
                        for (int k = 0; k < 2 * LDV; ++k) unpack2(v[k], vf[2 * k], vf[2 * k + 1]);
                        // my group's GC channels start gsh floats into the aligned block (warp-uniform)
                        switch (gsh) {
                            case 0:
#pragma unroll
                                for (int cc = 0; cc < GC; ++cc) xs[si][cc] = vf[cc];
                                break;
                            case 1:
#pragma unroll
                                for (int cc = 0; cc < GC; ++cc) xs[si][cc] = vf[(cc + 1 < 4 * LDV) ? cc + 1 : 0];
                                break;
                            case 2:
#pragma unroll
                                for (int cc = 0; cc < GC; ++cc) xs[si][cc] = vf[(cc + 2 < 4 * LDV) ? cc + 2 : 0];
                                break;
                            default:
#pragma unroll
                                for (int cc = 0; cc < GC; ++cc) xs[si][cc] = vf[(cc + 3 < 4 * LDV) ? cc + 3 : 0];
                                break;
                        }
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
                    const int par = chunk & 1;
                    ++chunk;
                    __syncthreads();
                    // the window is dead after the last sampling phase of a frame: start the next load under C / D
                    if (threadIdx.x == 0 && s0 + SC >= nS) {
                        if (f == 0) {
                            mbar_arrive_expect_tx(wbar, w.box_bytes);
                            tma_load_4d(win, &tm1, wbar, 0, org[2], org[3], b);
                        } else if (has_next) {
                            mbar_arrive_expect_tx(wbar, w.box_bytes);
                            tma_load_4d(win, &tm0, wbar, 0, org[4], org[5], b_next);
                        }
                    }
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
                            lmax = fmaxf(lmax, l[si]);
                            if (a.score[f] != nullptr && live)
                                a.score[f][(((int64_t)b * NG + g) * nS + (s0 + si)) * HW + pix] = l[si];
                        }
                        float alpha = 1.0f;
                        if (lmax - m_ref > 30.0f) {          // lazy re-base of the softmax shift (first chunk: m_ref = -inf)
                            alpha = expf(m_ref - lmax);
                            m_ref = lmax;
                            sum *= alpha;
                            rebase[par] = 1;
                        }
                        abuf[((g >> 2) * 32 + lane) * 4 + (g & 3)] = alpha;
#pragma unroll
                        for (int si = 0; si < SC; ++si) {
                            const float p = expf(l[si] - m_ref);
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
                    if (threadIdx.x == 0) rebase[par ^ 1] = 0;
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
                        const float4 wv = *reinterpret_cast<const float4*>(wr + c4);
                        d = fmaf(wv.x, yv[h][c4], d);
                        if (c4 + 1 < GC) d = fmaf(wv.y, yv[h][c4 + 1], d);
                        if (c4 + 2 < GC) d = fmaf(wv.z, yv[h][c4 + 2], d);
                        if (c4 + 3 < GC) d = fmaf(wv.w, yv[h][c4 + 3], d);
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

// tensor maps of the two channels-last feature maps: box {REC, Ww, Hw, 1} of dims {C, W, H, B}
struct WinTmapEntry {
    const void* base;
    int B, H, W, C, Ww, Hw, rec;
    CUtensorMap map;
};

static int get_win_tmap(const float* feat, int B, int H, int W, int C, int rec, int Ww, int Hw, const CUtensorMap** out) {
    static thread_local WinTmapEntry cache[32];
    static thread_local int next = 0;
    for (int i = 0; i < 32; ++i) {
        const WinTmapEntry& e = cache[i];
        if (e.base == feat && e.B == B && e.H == H && e.W == W && e.C == C && e.Ww == Ww && e.Hw == Hw && e.rec == rec) {
            *out = &e.map;
            return VFI_OK;
        }
    }
    WinTmapEntry& e = cache[next];
    next = (next + 1) % 32;
    e.base = nullptr;
    const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)C * 4, (uint64_t)W * C * 4, (uint64_t)H * W * C * 4};
    const uint32_t box[4] = {(uint32_t)rec, (uint32_t)Ww, (uint32_t)Hw, 1};
    const int rc = make_tmap_f32_nd(&e.map, feat, 4, dims, strides, box, 0);
    if (rc) return rc;
    e.base = feat; e.B = B; e.H = H; e.W = W; e.C = C; e.Ww = Ww; e.Hw = Hw; e.rec = rec;
    *out = &e.map;
    return VFI_OK;
}

template <int GC, int NG, int SC>
static int launch_win(const DattnArgs& a_in, cudaStream_t st) {
    using Cfg = WinCfg<GC, NG, SC>;
    DattnArgs a = a_in;
    // reach of the residual offsets around pixel + flow (module comment): RX columns / RY rows on either side
    const int RX = (int)ceilf(a.offset_scale * 0.25f), RY = (int)ceilf(a.offset_scale);
    // needed: tile + reach + the east / south tap + one position of rounding safety on either side
    const int need_w = kWinTW + 2 * RX + 3, need_h = kWinTH + 2 * RY + 3;
    const size_t max_smem = 232448;   // 227 KB: the per-block opt-in limit of sm_100
    const size_t fixed = (size_t)Cfg::FIXED_FLOATS * 4 + 128 + 16;
    // slack for flow variation inside the tile: as much as fits, y first (x displacements are divided by 4)
    int sx = 0, sy = 0;
    auto fits = [&](int sx_, int sy_) {
        return fixed + (size_t)(need_w + 2 * sx_) * (need_h + 2 * sy_) * Cfg::REC * 4 <= max_smem;
    };
    if (!fits(0, 0)) return VFI_ERR_UNSUPPORTED;
    for (int it = 0; it < 8; ++it) {
        if (sy <= 2 * sx && sy < 4 && fits(sx, sy + 1)) ++sy;
        else if (sx < 2 && fits(sx + 1, sy)) ++sx;
        else if (sy < 4 && fits(sx, sy + 1)) ++sy;
        else break;
    }
    WinGeom w;
    w.Ww = need_w + 2 * sx;
    w.Hw = need_h + 2 * sy;
    if (w.Ww > 256 || w.Hw > 256) return VFI_ERR_UNSUPPORTED;
    w.mx = RX + 1 + sx;
    w.my = RY + 1 + sy;
    w.row_bytes = w.Ww * Cfg::REC * 4;
    w.box_bytes = (uint32_t)((size_t)w.Ww * w.Hw * Cfg::REC * 4);
    const size_t smem = fixed + w.box_bytes;
    const CUtensorMap *tm0 = nullptr, *tm1 = nullptr;
    int rc = get_win_tmap(a.feat[0], a.B, a.H, a.W, Cfg::C, Cfg::REC, w.Ww, w.Hw, &tm0);
    if (rc) return rc;
    rc = get_win_tmap(a.feat[1], a.B, a.H, a.W, Cfg::C, Cfg::REC, w.Ww, w.Hw, &tm1);
    if (rc) return rc;
    static thread_local size_t smem_set = 0;
    if (smem > smem_set) {
        VFI_CUDA(cudaFuncSetAttribute(dattn_win_kernel<GC, NG, SC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    a.tiles_x = (int)ceil_div(a.W, kWinTW);
    a.tiles_y = (int)ceil_div(a.H, kWinTH);
    const int64_t tiles = (int64_t)a.B * a.tiles_x * a.tiles_y;
    const int blocks = (int)min64(tiles, (int64_t)kNumSMs);
    dattn_win_kernel<GC, NG, SC><<<blocks, Cfg::THREADS, smem, st>>>(*tm0, *tm1, a, w);
    VFI_CHECK_LAUNCH("dattn_win_kernel");
    return VFI_OK;
}

bool dattn_win_covers(int C, int nG, int nS, float offset_scale, int coords_final, int offs_cl) {
    return C == 72 && nG == 8 && offs_cl && !coords_final && nS % 4 == 0 && offset_scale > 0.0f && offset_scale <= 16.0f;
}

int launch_dattn_win(const DattnArgs& a, int C, int nG, cudaStream_t st) {
    // covered: channels-last features and offsets, raw offsets (bounded reach), nS a multiple of 4, 16-byte aligned rows
    if (!a.feat_cl || !dattn_win_covers(C, nG, a.nS, a.offset_scale, a.coords_final, a.offs_cl)) return VFI_ERR_UNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(a.feat[0]) | reinterpret_cast<uintptr_t>(a.feat[1]) | reinterpret_cast<uintptr_t>(a.offs[0]) |
         reinterpret_cast<uintptr_t>(a.offs[1])) & 15)
        return VFI_ERR_UNSUPPORTED;
    if (C == 72 && nG == 8) return launch_win<9, 8, 2>(a, st);
    return VFI_ERR_UNSUPPORTED;
}

}  // namespace vfi
