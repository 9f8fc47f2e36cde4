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
    float r_wm1, r_half_h;   // RN(1 / (Wg - 1)), RN(1 / ((Hg - 1) / 2)): see div_rn
};

// x / d, correctly rounded, from r = RN(1 / d): q = RN(x r); e = x - q d (exact in one FMA); RN(q + e r).  Bit-identical to
// the IEEE division the reference's coordinate arithmetic performs (modules/deformable_attn.py:94-95) -- checked against
// numpy's float32 division on 2.9e8 operands for the divisors of every pyramid level, and on the B200 the kernel's
// output is bit-identical with either form (for infinities / NaN the result is non-finite either way, which is all the
// sampler looks at) -- in 3 instructions instead of ~10 + a branch.
__device__ __forceinline__ float div_rn(float x, float d, float r) {
    const float q = __fmul_rn(x, r);
    const float e = __fmaf_rn(-q, d, x);
    return __fmaf_rn(e, r, q);
}

constexpr int kWinTW = 8, kWinTH = 4;   // pixel tile of a CTA (8 warps x 4 horizontally adjacent pixels)

// Shared-memory plan (floats).  Row / block strides are padded so that the 8 lanes of a pixel (one per group, reading 8
// different rows / blocks with LDS.128) fall into 8 different bank quads.
template <int GC>
struct WinCfg {
    static constexpr int NG = 8;
    static constexpr int C = GC * NG;
    static constexpr int C4 = (C + 3) & ~3;
    static constexpr int GCP = (GC + 3) & ~3;
    static constexpr int REC4 = ((C4 / 4) & 1) ? C4 / 4 : C4 / 4 + 1;   // window record length in 16-byte units, odd
    static constexpr int REC = REC4 * 4;
    static constexpr int LDV = (((GC * (NG - 1)) & 3) + GC + 3) / 4;     // LDS.128 per tap covering any group's channels
    static constexpr int THREADS = 256;                                 // 8 warps x (4 pixels x 8 groups)
    static constexpr int WQS = C4 + 4;                                  // Wq row stride: GC * WQS = 12 (mod 32) for GC = 9
    static constexpr int WQ = C * WQS;
    static constexpr int WTB = C * GCP + 4;                             // Wk / Wv block stride: 4 (mod 32)
    static constexpr int WT = NG * WTB;
    static constexpr int OB = 512;                                      // offset-head bias, [s][g][2]
    static constexpr int MISC = 16;                                     // org[3][2], mbarrier, pad
    static constexpr int FIXED_FLOATS = WQ + 2 * WT + OB + MISC;
    static_assert(LDV * 4 <= 12, "tap = at most three 128-bit loads");
    static_assert((GC * WQS) % 32 == 12 || (GC * WQS) % 32 == 4 || (GC * WQS) % 32 == 20 || (GC * WQS) % 32 == 28, "Wq rows of the 8 groups conflict");
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

// reduce-scatter of 8 per-lane values over the 8 lanes of a pixel (lane & 7 = g): returns, in lane g, the sum over the
// 8 lanes of their d[g].  7 shuffles (xor 4, 2, 1), pairwise summation order.
__device__ __forceinline__ float reduce_scatter8(const float (&d)[8], int g) {
    const bool hi = (g & 4) != 0, mid = (g & 2) != 0, lo = (g & 1) != 0;
    float e[4], f[2];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float recv = __shfl_xor_sync(0xffffffffu, hi ? d[i] : d[i + 4], 4);
        e[i] = (hi ? d[i + 4] : d[i]) + recv;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const float recv = __shfl_xor_sync(0xffffffffu, mid ? e[i] : e[i + 2], 2);
        f[i] = (mid ? e[i + 2] : e[i]) + recv;
    }
    const float recv = __shfl_xor_sync(0xffffffffu, lo ? f[0] : f[1], 1);
    return (lo ? f[1] : f[0]) + recv;
}

// Mapping: lane = (pixel p = lane >> 3, group g = lane & 7): a warp owns 4 horizontally adjacent pixels and all 8 offset
// groups of them; lane (p, g) samples group g's GC channels and owns head g of pixel p.  The cross-group sums of the
// logits and the exchange of the softmax weights are warp shuffles among the 8 lanes of a pixel, so the sample loop has
// NO block barrier: the 8 warps of a CTA (8x4 pixel tile) drift freely and cover each other's latencies; they meet only
// to swap the window (once per frame).
template <int GC>
__global__ void __launch_bounds__(WinCfg<GC>::THREADS, 1)
dattn_win_kernel(const __grid_constant__ CUtensorMap tm0, const __grid_constant__ CUtensorMap tm1, DattnArgs a, WinGeom w) {
    using Cfg = WinCfg<GC>;
    constexpr int NG = 8, C = Cfg::C, GCP = Cfg::GCP, T = Cfg::THREADS, REC = Cfg::REC, LDV = Cfg::LDV;
    constexpr int WQS = Cfg::WQS, WTB = Cfg::WTB, HP = NG / 2;
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t* smem_al = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    float* win = reinterpret_cast<float*>(smem_al);                     // [Hw][Ww][REC]
    float* Wq_s = win + (size_t)w.Ww * w.Hw * REC;                      // [C][WQS]
    float* WkT_s = Wq_s + Cfg::WQ;                                      // [g][r][GCP] (+4 per g) = Wk[r][g*GC + cc]
    float* WvT_s = WkT_s + Cfg::WT;
    float* ob_s = WvT_s + Cfg::WT;                                      // [s][g][2] bias of the offset head
    int* org = reinterpret_cast<int*>(ob_s + Cfg::OB);                  // [3][2]: (tile,f0), (tile,f1), (next tile,f0)
    uint64_t* wbar = reinterpret_cast<uint64_t*>(org + 8);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane & 7, p = lane >> 3;
    const int pbase = lane & ~7;               // lane of (p, group 0)
    const int64_t HW = (int64_t)a.H * a.W;
    const float scale = rsqrtf((float)GC);
    const int nS = a.nS;
    const int tiles_per_img = a.tiles_x * a.tiles_y;
    const int64_t n_tiles = (int64_t)a.B * tiles_per_img;

    if (threadIdx.x == 0) {
        mbar_init(wbar, 1);
        fence_mbar_init();
        tma_prefetch_desc(&tm0);
        tma_prefetch_desc(&tm1);
    }
    for (int i = threadIdx.x; i < C * WQS; i += T) {
        const int r = i / WQS, c = i - r * WQS;
        Wq_s[i] = c < C ? __ldg(a.Wq + r * C + c) : 0.0f;
    }
    for (int i = threadIdx.x; i < NG * WTB; i += T) {
        const int gg = i / WTB, rem = i - gg * WTB;
        const int r = rem / GCP, cc = rem - r * GCP;
        const bool ok = r < C && cc < GC;
        WkT_s[i] = ok ? __ldg(a.Wk + r * C + gg * GC + cc) : 0.0f;
        WvT_s[i] = ok ? __ldg(a.Wv + r * C + gg * GC + cc) : 0.0f;
    }
    for (int i = threadIdx.x; i < NG * nS * 2; i += T) {
        const int xy = i & 1, gg = (i >> 1) & 7, s = i >> 4;
        ob_s[i] = a.obias != nullptr ? __ldg(a.obias + (gg * nS + s) * 2 + xy) : 0.0f;
    }
    __syncthreads();                           // mbarrier initialised before anybody arms it; weights staged
    if ((int64_t)blockIdx.x < n_tiles && warp == 0) {      // first window of this CTA
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

    const float wm1 = (float)(a.fe.Wg - 1), hm1 = (float)(a.fe.Hg - 1);
    const float half_h = (float)((a.fe.Hg - 1.0) / 2.0);
    const float xmax = (float)a.fe.Wg + 1.0f, ymax = (float)a.fe.Hg + 1.0f;
    const int gbase = (g * GC) & ~3, gsh = (g * GC) & 3;                // aligned channel base of my group, shift inside it
    const uint32_t win_g = smem_u32(win) + gbase * 4;
    const float* WkT_g = WkT_s + g * WTB;
    const float* WvT_g = WvT_s + g * WTB;

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int b = (int)(tile / tiles_per_img);
        const int trem = (int)(tile - (int64_t)b * tiles_per_img);
        const int ty = trem / a.tiles_x, tx = trem - ty * a.tiles_x;
        const int x_raw = tx * kWinTW + (warp & 1) * 4 + p, y_raw = ty * kWinTH + (warp >> 1);
        const bool live = x_raw < a.W && y_raw < a.H;
        const int x = min(x_raw, a.W - 1), y = min(y_raw, a.H - 1);
        const int64_t pix = (int64_t)y * a.W + x;
        const int64_t tile_next = tile + gridDim.x;
        const bool has_next = tile_next < n_tiles;
        const int b_next = has_next ? (int)(tile_next / tiles_per_img) : 0;

        // window origins of this tile's two frames and of the next tile's first frame (warps 0-2, one each)
        if (warp < 2) {
            int ox, oy;
            window_origin(a, w, warp, b, tx, ty, lane, ox, oy);
            if (lane == 0) { org[2 * warp] = ox; org[2 * warp + 1] = oy; }
        } else if (warp == 2 && has_next) {
            const int trn = (int)(tile_next - (int64_t)b_next * tiles_per_img);
            int ox, oy;
            window_origin(a, w, 0, b_next, trn % a.tiles_x, trn / a.tiles_x, lane, ox, oy);
            if (lane == 0) { org[4] = ox; org[5] = oy; }
        }

        // ---- prologue: qp_g = Wq_g q + bq_g (head g of pixel p), kt[h][my channels] = scale * Wk_h^T qp_h ----
        float qp[GC];
        float cb;
        {
#pragma unroll
            for (int j = 0; j < GC; ++j) qp[j] = __ldg(a.bq + g * GC + j);
            const float* wrow = Wq_s + (g * GC) * WQS;
            if (a.qo_cl & 1) {
                const float4* qv4 = reinterpret_cast<const float4*>(a.q + ((int64_t)b * HW + pix) * C);
#pragma unroll 3
                for (int c4 = 0; c4 < C / 4; ++c4) {
                    const float4 qv = __ldg(qv4 + c4);
#pragma unroll
                    for (int j = 0; j < GC; ++j) {
                        const float4 wv = *reinterpret_cast<const float4*>(wrow + j * WQS + 4 * c4);
                        qp[j] = fmaf(wv.x, qv.x, qp[j]);
                        qp[j] = fmaf(wv.y, qv.y, qp[j]);
                        qp[j] = fmaf(wv.z, qv.z, qp[j]);
                        qp[j] = fmaf(wv.w, qv.w, qp[j]);
                    }
                }
            } else {
                const float* qpl = a.q + (int64_t)b * C * HW + pix;
#pragma unroll 2
                for (int c = 0; c < C; ++c) {
                    const float qv = __ldg(qpl + (int64_t)c * HW);
#pragma unroll
                    for (int j = 0; j < GC; ++j) qp[j] = fmaf(wrow[j * WQS + c], qv, qp[j]);
                }
            }
            cb = 0.0f;
#pragma unroll
            for (int j = 0; j < GC; ++j) {
                cb = fmaf(qp[j], __ldg(a.bk + g * GC + j), cb);
                qp[j] *= scale;
            }
            cb *= scale;
        }
        f32x2 kt2[HP][GC];   // (kt[2hp][cc], kt[2hp+1][cc]) for my group's channels cc
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
                    const float qh = __shfl_sync(0xffffffffu, qp[j], pbase | h);      // qp_h[j] lives in lane (p, h)
                    const float* wr = WkT_g + (h * GC + j) * GCP;
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
        __syncthreads();   // org[] of this tile visible
        const int org_f1x = org[2], org_f1y = org[3], org_nx = org[4], org_ny = org[5];

        f32x2 yv2[HP][GC];   // (yv[2hp][cc], yv[2hp+1][cc]): weighted sums of my group's channels for every head
#pragma unroll
        for (int hp = 0; hp < HP; ++hp)
#pragma unroll
            for (int cc = 0; cc < GC; ++cc) yv2[hp][cc] = 0ull;
        float m_ref = -INFINITY, sum = 0.0f;      // softmax state of head g of pixel p
        const float lxv = __ldg(a.lin_x + x), lyv = __ldg(a.lin_y + y);

        for (int f = 0; f < 2; ++f) {
            const float* fp = a.ft[f] + b * a.ft_sn + y * a.ft_sh + x * a.ft_sw;
            const float ftx = __ldg(fp), fty = __ldg(fp + a.ft_sc);
            const int ox = f == 0 ? org[0] : org_f1x, oy = f == 0 ? org[1] : org_f1y;
            const float* offcl = a.offs[f] + ((int64_t)b * HW + pix) * (NG * nS * 2) + (int64_t)g * nS * 2;
            const float* featg = a.feat[f] + (int64_t)b * HW * C + gbase;      // channels-last, my group's aligned base
            // raw offsets of my group: 4 slots (32 B, one sector) per load group, loaded one group (two iterations of the
            // pair loop) ahead of their use
            float4 od_a = ld_stream4(offcl), od_b = ld_stream4(offcl + 4);
            mbar_wait(wbar, wphase);       // window (tile, f) has landed
            wphase ^= 1;
            float4 oc_a = od_a, oc_b = od_b;
            for (int s0 = 0; s0 < nS; s0 += 2) {
                if ((s0 & 2) == 0) {
                    oc_a = od_a;
                    oc_b = od_b;
                    if (s0 + 4 < nS) {
                        od_a = ld_stream4(offcl + 2 * (s0 + 4));
                        od_b = ld_stream4(offcl + 2 * (s0 + 4) + 4);
                    }
                }
                const float4 oc = (s0 & 2) ? oc_b : oc_a;
                float xs[2][GC];
                float lg[2];
#pragma unroll
                for (int si = 0; si < 2; ++si) {
                    // ---- sample my group's channels at slot s0 + si ----
                    const float2 ob = *reinterpret_cast<const float2*>(ob_s + ((s0 + si) * NG + g) * 2);
                    float dx = __fadd_rn(si == 0 ? oc.x : oc.z, ob.x), dy = __fadd_rn(si == 0 ? oc.y : oc.w, ob.y);
                    // libdevice tanhf: an SFU-only tanh (1 - 2 / (ex2 + 1), absolute error ~3e-7) was measured to push the
                    // output error from 7e-6 to 1.8e-5 of full scale on white-noise features, for 1 % of kernel time
                    dx = __fadd_rn(__fmul_rn(a.offset_scale, tanhf(dx)), ftx);
                    dy = __fadd_rn(__fmul_rn(a.offset_scale, tanhf(dy)), fty);
                    // coordinate replay of sampler_geom (sampler.cuh), frame coordinates
                    const float gx = __fadd_rn(lxv, __fmul_rn(div_rn(dx, wm1, w.r_wm1), 0.5f));
                    const float gy = __fadd_rn(lyv, div_rn(dy, half_h, w.r_half_h));
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
                    if (in_win) {
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
                        // this lane's sample lies outside the window (flow discontinuity): gather from global memory
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
                    // my group's GC channels start gsh = (9 g) & 3 = g & 3 floats into the aligned block
#pragma unroll
                    for (int cc = 0; cc < GC; ++cc) {
                        const float s01 = (gsh & 1) ? vf[(cc + 1) % (4 * LDV)] : vf[cc];
                        const float s23 = (gsh & 1) ? vf[(cc + 3) % (4 * LDV)] : vf[(cc + 2) % (4 * LDV)];
                        xs[si][cc] = (gsh & 2) ? s23 : s01;
                    }
                    // ---- partial logits of all heads over my channels, summed over the 8 groups of the pixel ----
                    float d[NG];
#pragma unroll
                    for (int hp = 0; hp < HP; ++hp) {
                        f32x2 d2 = 0ull;
#pragma unroll
                        for (int cc = 0; cc < GC; ++cc) d2 = fma2s(kt2[hp][cc], xs[si][cc], d2);
                        unpack2(d2, d[2 * hp], d[2 * hp + 1]);
                    }
                    lg[si] = reduce_scatter8(d, g) + cb;          // logit of head g at this slot
                }
                // ---- softmax state of head g (lazy re-base of the shift: only when the max grows by > 30) ----
                const float lmax = fmaxf(lg[0], lg[1]);
                float alpha = 1.0f;
                const bool reb = lmax - m_ref > 30.0f;            // first pair of the pixel: m_ref = -inf
                if (reb) {
                    alpha = expf(m_ref - lmax);                   // exp(-inf) = 0 on the first pair
                    m_ref = lmax;
                    sum *= alpha;
                }
                const float p0 = expf(lg[0] - m_ref), p1 = expf(lg[1] - m_ref);
                sum += p0;
                sum += p1;
                if (__any_sync(0xffffffffu, reb)) {               // rare after the first pair
#pragma unroll
                    for (int hp = 0; hp < HP; ++hp) {
                        const f32x2 a2 = pack2(__shfl_sync(0xffffffffu, alpha, pbase | (2 * hp)),
                                               __shfl_sync(0xffffffffu, alpha, pbase | (2 * hp + 1)));
#pragma unroll
                        for (int cc = 0; cc < GC; ++cc) yv2[hp][cc] = mul2(yv2[hp][cc], a2);
                    }
                }
                // ---- weighted accumulation: every lane needs the weights of all 8 heads of its pixel ----
#pragma unroll
                for (int hp = 0; hp < HP; ++hp) {
                    const f32x2 pa = pack2(__shfl_sync(0xffffffffu, p0, pbase | (2 * hp)), __shfl_sync(0xffffffffu, p0, pbase | (2 * hp + 1)));
                    const f32x2 pb = pack2(__shfl_sync(0xffffffffu, p1, pbase | (2 * hp)), __shfl_sync(0xffffffffu, p1, pbase | (2 * hp + 1)));
#pragma unroll
                    for (int cc = 0; cc < GC; ++cc) {
                        yv2[hp][cc] = fma2s(pa, xs[0][cc], yv2[hp][cc]);
                        yv2[hp][cc] = fma2s(pb, xs[1][cc], yv2[hp][cc]);
                    }
                }
            }
            __syncthreads();       // every warp is done with the window of (tile, f): start the next load
            if (threadIdx.x == 0) {
                if (f == 0) {
                    mbar_arrive_expect_tx(wbar, w.box_bytes);
                    tma_load_4d(win, &tm1, wbar, 0, org_f1x, org_f1y, b);
                } else if (has_next) {
                    mbar_arrive_expect_tx(wbar, w.box_bytes);
                    tma_load_4d(win, &tm0, wbar, 0, org_nx, org_ny, b_next);
                }
            }
        }

        // ---- epilogue: out_h = Wv_h (y_h / sum_h) + bv_h; y_h is spread over the 8 lanes of the pixel ----
        const float inv = 1.0f / sum;
        const int64_t o_sp = (a.qo_cl & 2) ? C : 1, o_sc = (a.qo_cl & 2) ? 1 : HW;
        float* op = a.out + (int64_t)b * C * HW + pix * o_sp + (g * GC) * o_sc;
#pragma unroll
        for (int j = 0; j < GC; ++j) {
            float d[NG];
#pragma unroll
            for (int hp = 0; hp < HP; ++hp) {
                // rows h * GC + j of Wv for h = 2hp, 2hp + 1, my group's columns
                const float* w0 = WvT_g + ((2 * hp) * GC + j) * GCP;
                const float* w1 = w0 + GC * GCP;
                f32x2 d2 = 0ull;
#pragma unroll
                for (int c4 = 0; c4 < GCP; c4 += 4) {
                    const float4 wa = *reinterpret_cast<const float4*>(w0 + c4);
                    const float4 wb = *reinterpret_cast<const float4*>(w1 + c4);
                    d2 = fma2(pack2(wa.x, wb.x), yv2[hp][c4], d2);
                    if (c4 + 1 < GC) d2 = fma2(pack2(wa.y, wb.y), yv2[hp][c4 + 1], d2);
                    if (c4 + 2 < GC) d2 = fma2(pack2(wa.z, wb.z), yv2[hp][c4 + 2], d2);
                    if (c4 + 3 < GC) d2 = fma2(pack2(wa.w, wb.w), yv2[hp][c4 + 3], d2);
                }
                unpack2(d2, d[2 * hp], d[2 * hp + 1]);
            }
            const float o = reduce_scatter8(d, g);              // (Wv y)[g * GC + j] of pixel p, in lane (p, g)
            if (live) st_stream(op + (int64_t)j * o_sc, fmaf(o, inv, __ldg(a.bv + g * GC + j)));
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

template <int GC>
static int launch_win(const DattnArgs& a_in, cudaStream_t st) {
    using Cfg = WinCfg<GC>;
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
    w.r_wm1 = (float)(1.0 / (double)(float)(a.fe.Wg - 1));
    w.r_half_h = (float)(1.0 / (double)(float)((a.fe.Hg - 1.0) / 2.0));
    const size_t smem = fixed + w.box_bytes;
    const CUtensorMap *tm0 = nullptr, *tm1 = nullptr;
    int rc = get_win_tmap(a.feat[0], a.B, a.H, a.W, Cfg::C, Cfg::REC, w.Ww, w.Hw, &tm0);
    if (rc) return rc;
    rc = get_win_tmap(a.feat[1], a.B, a.H, a.W, Cfg::C, Cfg::REC, w.Ww, w.Hw, &tm1);
    if (rc) return rc;
    static thread_local size_t smem_set = 0;
    if (smem > smem_set) {
        VFI_CUDA(cudaFuncSetAttribute(dattn_win_kernel<GC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    a.tiles_x = (int)ceil_div(a.W, kWinTW);
    a.tiles_y = (int)ceil_div(a.H, kWinTH);
    const int64_t tiles = (int64_t)a.B * a.tiles_x * a.tiles_y;
    const int blocks = (int)min64(tiles, (int64_t)kNumSMs);
    dattn_win_kernel<GC><<<blocks, Cfg::THREADS, smem, st>>>(*tm0, *tm1, a, w);
    VFI_CHECK_LAUNCH("dattn_win_kernel");
    return VFI_OK;
}

bool dattn_win_covers(int C, int nG, int nS, float offset_scale, int coords_final, int offs_cl) {
    return C == 72 && nG == 8 && offs_cl && !coords_final && nS % 4 == 0 && nS * nG * 2 <= 512 && offset_scale > 0.0f &&
           offset_scale <= 16.0f;
}

int launch_dattn_win(const DattnArgs& a, int C, int nG, cudaStream_t st) {
    // covered: channels-last features and offsets, raw offsets (bounded reach), nS a multiple of 4, 16-byte aligned rows
    if (!a.feat_cl || !dattn_win_covers(C, nG, a.nS, a.offset_scale, a.coords_final, a.offs_cl)) return VFI_ERR_UNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(a.feat[0]) | reinterpret_cast<uintptr_t>(a.feat[1]) | reinterpret_cast<uintptr_t>(a.offs[0]) |
         reinterpret_cast<uintptr_t>(a.offs[1])) & 15)
        return VFI_ERR_UNSUPPORTED;
    if (a.fe.Wg < 2 || a.fe.Hg < 2) return VFI_ERR_UNSUPPORTED;      // degenerate normalisation: the gather kernel handles it
    if (C == 72 && nG == 8) return launch_win<9>(a, st);
    return VFI_ERR_UNSUPPORTED;
}

}  // namespace vfi
