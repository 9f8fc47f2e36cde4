// Modulated deformable convolution (DCNv2), shared-memory WINDOW form: the hybrid of the two older kernels.
// torchvision.ops.deform_conv2d semantics (reference call sites: modules/dcnv2.py:60-66, models/archive/DCNDAT.py:71-77).
//
// dcn.cu gathers every bilinear corner from global memory (9 taps x 4 corners x Cin loads per pixel: LSU-bound, 12 % of
// HBM at BASELINE configs[2]); dcn_tc.cu stages the input window by TMA but then rewrites the sampled columns into UMMA
// slabs for a contraction that is only 72 MACs deep per sample (tensor pipe 1.2 % busy, bank conflicts on the slab
// writes).  Here the window staging is kept and the contraction stays on the FMA pipe:
//   CTA = one 32x8 tile of output pixels x ONE weight group (thread = pixel).
//   One 3-D TMA box brings the group's cin_g input planes of the window the tile's taps fall into (48x24 positions:
//   tile + 3x3 kernel + 7 px of reach around the tile's mean centre-tap offset) into shared memory, zero-filled
//   outside the frame -- which is torchvision's per-corner bounds rule, so the fast path needs no corner predicates.
//   While the box is in flight every thread loads its 9 x (dy, dx, mask) values (the only HBM-streamed operands).
//   Per tap: 4 LDS per channel (a warp = 32 neighbouring pixels of a row: consecutive addresses), the bilinear value,
//   then cog FMAs against the group's weights (shared memory, [tap][channel][out], broadcast LDS.128, packed FFMA2).
//   A pixel whose tap leaves the window (flow outliers) takes predicated global loads for that tap.
// Requires: 3x3 taps at most (kh * kw <= 9), stride 1, dilation 1, cin_g <= 16, cog <= 16, every weight group inside one
// offset group, W % 4 == 0 and a 16-byte aligned x (TMA).  Everything else stays on dcn.cu.
#include "dcn_common.cuh"
#include "tma.cuh"

namespace vfi {

constexpr int kDwTW = 32, kDwTH = 8;                 // output tile
constexpr int kDwHalo = 7;
constexpr int kDwWW = kDwTW + 2 + 2 * kDwHalo;       // 48
constexpr int kDwWH = kDwTH + 2 + 2 * kDwHalo;       // 24
constexpr int kDwThreads = kDwTW * kDwTH;            // 256
constexpr int kDwMaxE = 192;                         // out-of-window (pixel, tap) pairs served by the cooperative gather pass

template <int COGP>
__global__ void __launch_bounds__(kDwThreads, 3)
dcn_win_kernel(const __grid_constant__ CUtensorMap xmap, const float* __restrict__ x, const float* __restrict__ offset,
               const float* __restrict__ mask, const float* __restrict__ weight, const float* __restrict__ bias,
               float* __restrict__ out, DcnParams p, int tiles_x, int tiles_per_img) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t* smem_al = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    const int taps = p.kh * p.kw;
    const int cin_g = p.Cin / p.WG, cog = p.Cout / p.WG, cpo = p.Cin / p.OG;
    float* win = reinterpret_cast<float*>(smem_al);                      // [cin_g][kDwWH][kDwWW]
    float* w_s = win + (size_t)cin_g * kDwWH * kDwWW;                    // [tap][cl][COGP]
    uint64_t* bar = reinterpret_cast<uint64_t*>(w_s + (size_t)taps * cin_g * COGP);
    int* s_shift = reinterpret_cast<int*>(bar + 1);                      // [2] (y, x) window displacement, [2] = list length
    float* s_list = reinterpret_cast<float*>(s_shift + 4);               // [kDwMaxE][4]: (h, w, mask, -) of out-of-window taps
    float* s_val = s_list + kDwMaxE * 4;                                 // [kDwMaxE][16]: their masked bilinear values per channel

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = blockIdx.y;
    const int b = blockIdx.x / tiles_per_img;
    const int tile = blockIdx.x - b * tiles_per_img;
    const int ty = tile / tiles_x, tx = tile - ty * tiles_x;
    const int xo0 = tx * kDwTW, yo0 = ty * kDwTH;
    const int64_t HWo = (int64_t)p.Ho * p.Wo, HW = (int64_t)p.H * p.W;
    const int og = (g * cin_g) / cpo;                                    // the offset group of this weight group

    const bool live = xo0 + lane < p.Wo && yo0 + warp < p.Ho;
    const int wo = min(xo0 + lane, p.Wo - 1), ho = min(yo0 + warp, p.Ho - 1);
    const int64_t pix = (int64_t)ho * p.Wo + wo;
    const float* offb = offset + ((int64_t)b * p.OG + og) * 2 * taps * HWo + pix;
    const float* maskb = mask ? mask + ((int64_t)b * p.OG + og) * taps * HWo + pix : nullptr;

    // my taps' offsets and modulation: streamed once, all in flight while the window is loaded
    float dy[9], dx[9], mv[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) {
        if (t < taps) {
            dy[t] = ld_stream(offb + (int64_t)(2 * t) * HWo);
            dx[t] = ld_stream(offb + (int64_t)(2 * t + 1) * HWo);
            mv[t] = maskb ? ld_stream(maskb + (int64_t)t * HWo) : 1.0f;
        }
    }
    if (tid == 0) {
        mbar_init(bar, 1);
        fence_mbar_init();
        // window displacement: mean offset of the centre tap at four probe pixels of the tile.  Any value is correct
        // (taps outside the window are fetched from global memory); a good one keeps them rare.
        float sy = 0.0f, sx = 0.0f;
        const int ctap = taps / 2;
        for (int k = 0; k < 4; ++k) {
            const int px = min(xo0 + 4 + 8 * k, p.Wo - 1), py = min(yo0 + 1 + 2 * k, p.Ho - 1);
            const float* o = offset + ((int64_t)b * p.OG + og) * 2 * taps * HWo + (int64_t)py * p.Wo + px;
            const float oy = __ldg(o + (int64_t)(2 * ctap) * HWo), ox = __ldg(o + (int64_t)(2 * ctap + 1) * HWo);
            sy += isfinite(oy) ? fminf(fmaxf(oy, -1.0e6f), 1.0e6f) : 0.0f;
            sx += isfinite(ox) ? fminf(fmaxf(ox, -1.0e6f), 1.0e6f) : 0.0f;
        }
        const int shy = (int)rintf(sy * 0.25f), shx = (int)rintf(sx * 0.25f);
        s_shift[0] = shy;
        s_shift[1] = shx;
        s_shift[2] = 0;
        // window origin in input coordinates (x a multiple of 4: TMA's 16-byte rule for the innermost coordinate)
        const int wy0 = yo0 - p.pad + shy - kDwHalo;
        const int wx0 = (xo0 - p.pad + shx - kDwHalo) & ~3;
        tma_prefetch_desc(&xmap);
        mbar_arrive_expect_tx(bar, (uint32_t)(cin_g * kDwWH * kDwWW * 4));
        tma_load_3d(win, &xmap, bar, wx0, wy0, b * p.Cin + g * cin_g);
    }
    // this group's weights: weight[(g*cog + o), cl, tap] -> w_s[(tap*cin_g + cl)*COGP + o]
    for (int i = tid; i < taps * cin_g * COGP; i += kDwThreads) {
        const int o = i % COGP, r = i / COGP;
        const int cl = r % cin_g, t = r / cin_g;
        w_s[i] = o < cog ? __ldg(weight + ((size_t)(g * cog + o) * cin_g + cl) * taps + t) : 0.0f;
    }
    __syncthreads();
    const int wy0 = yo0 - p.pad + s_shift[0] - kDwHalo;
    const int wx0 = (xo0 - p.pad + s_shift[1] - kDwHalo) & ~3;
    const float base_h = (float)(ho - p.pad), base_w = (float)(wo - p.pad);
    const float* xb = x + ((int64_t)b * p.Cin + (int64_t)g * cin_g) * HW;

    f32x2 acc[COGP / 2];
#pragma unroll
    for (int o = 0; o < COGP / 2; ++o) acc[o] = 0ull;

    // ---- taps outside the staged window (flow outliers): collected once, gathered cooperatively ----
    // A lane that fetched such a tap itself would stall its whole warp on a global round trip per tap (measured: 8.9 ms
    // at BASELINE configs[2] with 2 % frame-leaving flows against 2.5 ms without them).  Instead every thread files its
    // out-of-window (pixel, tap) pairs in a shared list, all 256 threads gather list entry x channel items with every
    // load in flight at once (under the window's TMA round trip), and the main loop reads the values back from shared
    // memory.  eidx: list slot per tap, 8 bits each (0xFF = none / list full: that tap takes the in-thread path).
    uint32_t eidx[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
    {
        const float wy_lo = (float)wy0, wy_hi = (float)(wy0 + kDwWH - 1), wx_lo = (float)wx0, wx_hi = (float)(wx0 + kDwWW - 1);
#pragma unroll
        for (int t = 0; t < 9; ++t) {
            if (t < taps) {
                const int ki = t / p.kw, kj = t - ki * p.kw;
                const float h = base_h + (float)ki + dy[t], wq = base_w + (float)kj + dx[t];
                // cheap conservative test first: floor(h) in [wy0, wy0 + WH - 2] <=> wy0 <= h < wy0 + WH - 1
                const bool surely_in = h >= wy_lo && h < wy_hi && wq >= wx_lo && wq < wx_hi;
                const bool in_frame = h > -1.0f && h < (float)p.H && wq > -1.0f && wq < (float)p.W;
                if (!surely_in && in_frame) {
                    const int e = atomicAdd(&s_shift[2], 1);
                    if (e < kDwMaxE) {
                        s_list[4 * e] = h;
                        s_list[4 * e + 1] = wq;
                        s_list[4 * e + 2] = mv[t];
                        eidx[t >> 2] = (eidx[t >> 2] & ~(0xFFu << (8 * (t & 3)))) | ((uint32_t)e << (8 * (t & 3)));
                    }
                }
            }
        }
    }
    __syncthreads();
    {
        const int n_e = min(s_shift[2], kDwMaxE);                        // uniform over the CTA
        if (n_e > 0) {
            for (int item = tid; item < n_e * cin_g; item += kDwThreads) {
                const int e = item / cin_g, cl = item - e * cin_g;
                const DcnTap tp = dcn_tap(s_list[4 * e], s_list[4 * e + 1], p.H, p.W);
                s_val[e * 16 + cl] = dcn_sample(xb + (int64_t)cl * HW + tp.o, tp, p.W) * s_list[4 * e + 2];
            }
            __syncthreads();
        }
    }

    mbar_wait(bar, 0);
    const uint32_t win_u32 = smem_u32(win);
#pragma unroll
    for (int t = 0; t < 9; ++t) {
        if (t < taps) {
            const int ki = t / p.kw, kj = t - ki * p.kw;
            DcnTap tp = dcn_tap(base_h + (float)ki + dy[t], base_w + (float)kj + dx[t], p.H, p.W);
            const float m = mv[t];
            const int ly = tp.hl - wy0, lx = tp.wl - wx0;
            const bool in_win = ly >= 0 && ly + 1 < kDwWH && lx >= 0 && lx + 1 < kDwWW;
            const bool any_corner = tp.k1 || tp.k2 || tp.k3 || tp.k4;      // false: the sample point is outside the frame
            if (!any_corner) continue;                                      // contributes exactly 0
            const float* wt = w_s + (size_t)t * cin_g * COGP;
            if (in_win) {
                // zero fill outside the frame == torchvision's dropped corners: no corner predicates on this path
                const float w1 = tp.w1, w2 = tp.w2, w3 = tp.w3, w4 = tp.w4;
                const uint32_t ad = win_u32 + (uint32_t)((ly * kDwWW + lx) * 4);
#pragma unroll 4
                for (int cl = 0; cl < cin_g; ++cl) {
                    float v1, v2, v3, v4;
                    const uint32_t a = ad + (uint32_t)cl * (kDwWH * kDwWW * 4);
                    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v1) : "r"(a));
                    asm volatile("ld.shared.f32 %0, [%1+4];" : "=f"(v2) : "r"(a));
                    asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v3) : "r"(a), "n"(kDwWW * 4));
                    asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v4) : "r"(a), "n"(kDwWW * 4 + 4));
                    float v = w1 * v1;
                    v = fmaf(w2, v2, v);
                    v = fmaf(w3, v3, v);
                    v = fmaf(w4, v4, v);
                    v *= m;
                    const float4* wr = reinterpret_cast<const float4*>(wt + cl * COGP);
#pragma unroll
                    for (int o4 = 0; o4 < COGP / 4; ++o4) {
                        const float4 wv = wr[o4];
                        acc[2 * o4] = fma2s(pack2(wv.x, wv.y), v, acc[2 * o4]);
                        acc[2 * o4 + 1] = fma2s(pack2(wv.z, wv.w), v, acc[2 * o4 + 1]);
                    }
                }
            } else {
                const uint32_t e = (eidx[t >> 2] >> (8 * (t & 3))) & 0xFFu;
                if (e != 0xFFu) {
                    // gathered by the cooperative pass above: already masked
                    const float* sv = s_val + e * 16;
#pragma unroll 4
                    for (int cl = 0; cl < cin_g; ++cl) {
                        const float v = sv[cl];
                        const float4* wr = reinterpret_cast<const float4*>(wt + cl * COGP);
#pragma unroll
                        for (int o4 = 0; o4 < COGP / 4; ++o4) {
                            const float4 wv = wr[o4];
                            acc[2 * o4] = fma2s(pack2(wv.x, wv.y), v, acc[2 * o4]);
                            acc[2 * o4 + 1] = fma2s(pack2(wv.z, wv.w), v, acc[2 * o4 + 1]);
                        }
                    }
                } else {
                    // list full (more than kDwMaxE out-of-window taps in this tile): in-thread global gathers
                    for (int cl = 0; cl < cin_g; ++cl) {
                        const float v = dcn_sample(xb + (int64_t)cl * HW + tp.o, tp, p.W) * m;
                        const float4* wr = reinterpret_cast<const float4*>(wt + cl * COGP);
#pragma unroll
                        for (int o4 = 0; o4 < COGP / 4; ++o4) {
                            const float4 wv = wr[o4];
                            acc[2 * o4] = fma2s(pack2(wv.x, wv.y), v, acc[2 * o4]);
                            acc[2 * o4 + 1] = fma2s(pack2(wv.z, wv.w), v, acc[2 * o4 + 1]);
                        }
                    }
                }
            }
        }
    }
    if (live) {
        float* op = out + ((int64_t)b * p.Cout + (int64_t)g * cog) * HWo + pix;
#pragma unroll
        for (int o2 = 0; o2 < COGP / 2; ++o2) {
            float a0, a1;
            unpack2(acc[o2], a0, a1);
            if (2 * o2 < cog) st_stream(op + (int64_t)(2 * o2) * HWo, a0 + (bias ? __ldg(bias + g * cog + 2 * o2) : 0.0f));
            if (2 * o2 + 1 < cog) st_stream(op + (int64_t)(2 * o2 + 1) * HWo, a1 + (bias ? __ldg(bias + g * cog + 2 * o2 + 1) : 0.0f));
        }
    }
}

int dcn_win_supported(const DcnParams& p, const float* x) {
    const int cin_g = p.Cin / p.WG, cog = p.Cout / p.WG, cpo = p.Cin / p.OG;
    if (p.kh * p.kw > 9 || p.kh > 3 || p.kw > 3 || p.stride != 1 || p.dil != 1) return 0;
    if (cin_g > 16 || cog > 16 || (cpo % cin_g) != 0) return 0;
    if (p.W % 4 != 0 || (reinterpret_cast<uintptr_t>(x) & 15) != 0) return 0;
    if ((int64_t)p.Ho * p.Wo >= (1ll << 31)) return 0;
    return 1;
}

template <int COGP>
static int launch_dcn_win_t(const CUtensorMap& xmap, const float* x, const float* offset, const float* mask, const float* weight,
                            const float* bias, float* out, int B, const DcnParams& p, cudaStream_t st) {
    const int cin_g = p.Cin / p.WG, taps = p.kh * p.kw;
    const size_t smem = (size_t)cin_g * kDwWH * kDwWW * 4 + (size_t)taps * cin_g * COGP * 4 + 64 + (size_t)kDwMaxE * 20 * 4 + 128;
    VFI_CUDA(cudaFuncSetAttribute(dcn_win_kernel<COGP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int tiles_x = (int)ceil_div(p.Wo, kDwTW), tiles_y = (int)ceil_div(p.Ho, kDwTH);
    const int64_t gx = (int64_t)B * tiles_x * tiles_y;
    VFI_REQUIRE(gx < (1ll << 31) && p.WG <= 65535, "vfi_dcn_f32: grid too large");
    dcn_win_kernel<COGP><<<dim3((unsigned)gx, (unsigned)p.WG), kDwThreads, smem, st>>>(xmap, x, offset, mask, weight, bias, out, p,
                                                                                       tiles_x, tiles_x * tiles_y);
    VFI_CHECK_LAUNCH("dcn_win_kernel");
    return VFI_OK;
}

int launch_dcn_win(const float* x, const float* offset, const float* mask, const float* weight, const float* bias, float* out, int B,
                   const DcnParams& p, cudaStream_t st) {
    if ((int64_t)B * p.Cin >= (1ll << 31)) return VFI_ERR_UNSUPPORTED;
    CUtensorMap xmap;
    const int cin_g = p.Cin / p.WG, cog = p.Cout / p.WG;
    int rc = make_tmap_f32_3d(&xmap, x, (uint64_t)p.W, (uint64_t)p.H, (uint64_t)B * p.Cin, (uint64_t)p.W * 4, (uint64_t)p.H * p.W * 4,
                              kDwWW, kDwWH, (uint32_t)cin_g);
    if (rc) return rc;
    if (cog <= 8) return launch_dcn_win_t<8>(xmap, x, offset, mask, weight, bias, out, B, p, st);
    if (cog <= 12) return launch_dcn_win_t<12>(xmap, x, offset, mask, weight, bias, out, B, p, st);
    return launch_dcn_win_t<16>(xmap, x, offset, mask, weight, bias, out, B, p, st);
}

}  // namespace vfi
