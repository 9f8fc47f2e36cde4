// DCNv2 with the per-group contraction on the 5th-gen tensor cores (tcgen05.mma, accumulators in TMEM).
// torchvision.ops.deform_conv2d semantics (modules/dcnv2.py:60-66, models/archive/DCNDAT.py:71-77).
//
// torchvision materialises im2col columns [Cin*K, px] in HBM and runs one skinny SGEMM per weight group.
// Here a CTA owns 128 consecutive output pixels (the UMMA M dimension; TMEM lane = pixel).  For every weight
// group g and kernel tap, the four gather warps (thread = pixel) sample the group's channels (bilinear,
// zero outside, * mask) and write the 128 x CP slab straight into shared memory in the canonical K-major
// UMMA layout (no swizzle: 8-row x 16-byte core matrices), so the "im2col" tile only ever exists on chip;
// a fifth warp issues tcgen05.mma.kind::tf32 against the group's weights (N = 16: Cout/WG real rows, rest 0)
// accumulating into TMEM columns [16 g, 16 g + 16).  fp32 parity (1e-5) needs more than one TF32 pass:
// both operands are split hi = rna_tf32(v), lo = v - hi and three MMAs (hi*hi, hi*lo, lo*hi) are issued per
// k-step (error-compensated 3xTF32).  Slabs cycle through an mbarrier ring (full: 128 arrivals after
// fence.proxy.async; empty: tcgen05.commit), the epilogue reads TMEM with tcgen05.ld (lane = pixel, so the
// NCHW stores are coalesced) and adds the bias.  K ordering is tap-major (k' = tap*CP + channel): one tap =
// CP/8 MMA k-steps; the weights are re-laid once per call by a tiny prep kernel (workspace).
#include <stdlib.h>
#include <string.h>

#include "dcn_common.cuh"
#include "tma.cuh"

namespace vfi {

constexpr int kTcPix = 128;       // pixels per CTA = UMMA M
constexpr int kTcN = 16;          // UMMA N (padded out channels per group)
constexpr int kTcThreads = 192;   // 4 gather/epilogue warps + 1 MMA warp + 1 TMA warp
constexpr int kTcTW = 32, kTcTH = 4;                    // output tile (a warp = one row of 32 pixels)
constexpr int kWinW = 48, kWinH = 20, kWinHalo = 7;     // staged input window (stride 1): tile + kernel + halo

__device__ __forceinline__ uint32_t to_tf32_rna(float v) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return r;
}

// smem matrix descriptor, K-major, SWIZZLE_NONE: core matrix = 8 rows x 16 B; LBO = byte distance between the two
// 16-byte K chunks of one k-step, SBO = byte distance between 8-row groups (cute/arch/mma_sm100_desc.hpp).
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)((lbo >> 4) & 0x3FFFu) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFFu) << 32;
    d |= (uint64_t)1 << 46;   // descriptor version (Blackwell)
    return d;                 // base_offset 0, lbo_mode 0, layout_type 0 (no swizzle)
}

// instruction descriptor: D = F32, A = B = TF32, both K-major, N = 16, M = 128
constexpr uint32_t kIdescTf32 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kTcN >> 3) << 17) | ((uint32_t)(kTcPix >> 4) << 24);

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(kIdescTf32), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// weights (Cout, cin_g, taps) -> per group, tap, k-step: 16 rows x 8 floats in UMMA K-major core-matrix order,
// hi block then lo block.  Layout (floats): [g][part(hi,lo)][tap][ks][rowgroup 2][j 2][row 8][4]
template <int KS>
__global__ void dcn_tc_prep_kernel(const float* __restrict__ weight, float* __restrict__ wprep, int cin_g, int cog, int taps, int WG) {
    const int per_part = taps * KS * 128;          // floats per (group, part)
    const int total = WG * 2 * per_part;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int r = i;
        const int e = r % 4; r /= 4;
        const int row8 = r % 8; r /= 8;
        const int j = r % 2; r /= 2;
        const int rg = r % 2; r /= 2;
        const int ks = r % KS; r /= KS;
        const int tap = r % taps; r /= taps;
        const int part = r % 2; r /= 2;
        const int g = r;
        const int n = rg * 8 + row8;               // output channel within the group
        const int cl = ks * 8 + j * 4 + e;         // input channel within the group
        float v = 0.0f;
        if (n < cog && cl < cin_g) v = weight[((size_t)(g * cog + n) * cin_g + cl) * taps + tap];
        const float hi = __uint_as_float(to_tf32_rna(v));
        wprep[i] = part == 0 ? hi : v - hi;
    }
}

template <int KS>
__global__ void __launch_bounds__(kTcThreads, 2) dcn_tc_kernel(const float* __restrict__ x, const float* __restrict__ offset,
                                                            const float* __restrict__ mask, const float* __restrict__ wprep,
                                                            const float* __restrict__ bias, float* __restrict__ out, DcnParams p,
                                                            int tiles_x, int tiles_per_img, int b_bytes, int tmem_cols,
                                                            const __grid_constant__ CUtensorMap xmap, int use_win) {
    constexpr int CP = KS * 8;
    constexpr int NS = KS == 1 ? 4 : 3;
    constexpr int A_PART = 4096;                 // bytes of one (part, k-step) block: 16 row groups x 256 B
    constexpr int A_STAGE = 2 * KS * A_PART;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* a_ring = smem;
    uint8_t* b_buf = smem + NS * A_STAGE;
    uint64_t* full = reinterpret_cast<uint64_t*>(b_buf + 2 * b_bytes);
    uint64_t* empty = full + NS;
    uint64_t* done = empty + NS;
    uint64_t* win_full = done + 1;      // [2]
    uint64_t* win_empty = win_full + 2; // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(win_empty + 2);
    int* s_shift = reinterpret_cast<int*>(tmem_slot + 1);   // [2] window displacement (y, x)

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int taps = p.kh * p.kw;
    const int cin_g = p.Cin / p.WG, cog = p.Cout / p.WG, cpo = p.Cin / p.OG;
    const int b = blockIdx.x / tiles_per_img;
    const int tile = blockIdx.x - b * tiles_per_img;
    const int64_t HWo = (int64_t)p.Ho * p.Wo, HW = (int64_t)p.H * p.W;
    const int ty = tile / tiles_x, tx = tile - ty * tiles_x;
    const int xo0 = tx * kTcTW, yo0 = ty * kTcTH;
    const int win_bytes = kWinW * kWinH * cin_g * 4;
    uint8_t* win_buf = reinterpret_cast<uint8_t*>(s_shift + 2);
    win_buf += (128u - (smem_u32(win_buf) & 127u)) & 127u;      // TMA destination alignment

    if (tid == 0) {
        for (int s = 0; s < NS; ++s) {
            mbar_init(&full[s], kTcPix);
            mbar_init(&empty[s], 1);
        }
        mbar_init(done, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&win_full[s], 1);
            mbar_init(&win_empty[s], kTcPix);
        }
        fence_mbar_init();
        // window displacement: mean offset of the centre tap (offset group 0) at four probe pixels of the tile.
        // Any value is correct (taps outside the window are fetched from global memory); a good one keeps them rare.
        float sy = 0.0f, sx = 0.0f;
        const int ctap = taps / 2;
        for (int k = 0; k < 4; ++k) {
            const int px = min(xo0 + 4 + 8 * k, p.Wo - 1), py = min(yo0 + (k & 1) * 2 + 1, p.Ho - 1);
            const float* o = offset + (int64_t)b * p.OG * 2 * taps * HWo + (int64_t)py * p.Wo + px;
            const float dy = __ldg(o + (int64_t)(2 * ctap) * HWo), dx = __ldg(o + (int64_t)(2 * ctap + 1) * HWo);
            sy += isfinite(dy) ? fminf(fmaxf(dy, -1.0e6f), 1.0e6f) : 0.0f;
            sx += isfinite(dx) ? fminf(fmaxf(dx, -1.0e6f), 1.0e6f) : 0.0f;
        }
        s_shift[0] = (int)rintf(sy * 0.25f);
        s_shift[1] = (int)rintf(sx * 0.25f);
    }
    if (warp == 4) {   // one warp allocates (and later frees) the accumulator columns
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // window origin in input coordinates (x a multiple of 4: TMA's 16-byte rule for the innermost coordinate)
    const int wy0 = yo0 * p.stride - p.pad + s_shift[0] - kWinHalo;
    const int wx0 = (xo0 * p.stride - p.pad + s_shift[1] - kWinHalo) & ~3;

    if (warp < 4) {
        // ================= gather warps: thread = output pixel =================
        const int m = tid;
        const bool live = xo0 + lane < p.Wo && yo0 + warp < p.Ho;
        const int wo = min(xo0 + lane, p.Wo - 1), ho = min(yo0 + warp, p.Ho - 1);
        const int64_t pix = (int64_t)ho * p.Wo + wo;
        const float base_h = (float)(ho * p.stride - p.pad), base_w = (float)(wo * p.stride - p.pad);
        const float* offb = offset + (int64_t)b * p.OG * 2 * taps * HWo + pix;
        const float* maskb = mask ? mask + (int64_t)b * p.OG * taps * HWo + pix : nullptr;
        const uint32_t row_off = (uint32_t)((m >> 3) * 256 + (m & 7) * 16);
        int it = 0;
        // offsets / mask are the HBM-streamed operands: when a weight group sits inside one offset group (the usual
        // OG | WG case) the next tap's three values are requested one iteration ahead
        const bool uniform_og = (cpo % cin_g) == 0;
        // prefetch distance = one whole weight group (kMaxPf taps): registers hold the current group's values while the
        // next group's loads are in flight, so HBM latency never sits on the per-tap critical path
        constexpr int kMaxPf = 9;
        const bool pf = uniform_og && taps <= kMaxPf;
        float cdy[kMaxPf], cdx[kMaxPf], cm[kMaxPf], ndy[kMaxPf], ndx[kMaxPf], nm[kMaxPf];
        auto load_group = [&](int gg, float* dy, float* dx, float* mv) {
            const int og = (gg * cin_g) / cpo;
#pragma unroll
            for (int tt = 0; tt < kMaxPf; ++tt) {
                if (tt < taps) {
                    dy[tt] = ld_stream(offb + ((int64_t)og * 2 * taps + 2 * tt) * HWo);
                    dx[tt] = ld_stream(offb + ((int64_t)og * 2 * taps + 2 * tt + 1) * HWo);
                    mv[tt] = maskb ? ld_stream(maskb + ((int64_t)og * taps + tt) * HWo) : 1.0f;
                }
            }
        };
        if (pf) load_group(0, ndy, ndx, nm);
        for (int g = 0; g < p.WG; ++g) {
            {   // this group's re-laid weights (hi | lo) -> shared
                const float4* src = reinterpret_cast<const float4*>(wprep) + (size_t)g * (b_bytes / 16);
                float4* dst = reinterpret_cast<float4*>(b_buf + (g & 1) * b_bytes);
                for (int i = m; i < b_bytes / 16; i += kTcPix) dst[i] = __ldg(src + i);
            }
            const float* xb = x + ((int64_t)b * p.Cin + (int64_t)g * cin_g) * HW;
            if (pf) {
#pragma unroll
                for (int tt = 0; tt < kMaxPf; ++tt) {
                    cdy[tt] = ndy[tt];
                    cdx[tt] = ndx[tt];
                    cm[tt] = nm[tt];
                }
                if (g + 1 < p.WG) load_group(g + 1, ndy, ndx, nm);
            }
            const float* wb = reinterpret_cast<const float*>(win_buf + (size_t)(g & 1) * win_bytes);
            if (use_win) mbar_wait(&win_full[g & 1], (g >> 1) & 1);
#pragma unroll 1
            for (int tap = 0; tap < taps; ++tap, ++it) {
                const int stage = it % NS;
                const int ki = tap / p.kw, kj = tap - ki * p.kw;
                float v[CP];
                if (uniform_og) {
                    float dy, dx, mval;
                    if (pf) {   // select from registers without dynamic indexing
                        dy = cdy[0]; dx = cdx[0]; mval = cm[0];
#pragma unroll
                        for (int tt = 1; tt < kMaxPf; ++tt)
                            if (tap == tt) { dy = cdy[tt]; dx = cdx[tt]; mval = cm[tt]; }
                    } else {
                        const int og = (g * cin_g) / cpo;
                        dy = ld_stream(offb + ((int64_t)og * 2 * taps + 2 * tap) * HWo);
                        dx = ld_stream(offb + ((int64_t)og * 2 * taps + 2 * tap + 1) * HWo);
                        mval = maskb ? ld_stream(maskb + ((int64_t)og * taps + tap) * HWo) : 1.0f;
                    }
                    const DcnTap t = dcn_tap(base_h + (float)(ki * p.dil) + dy, base_w + (float)(kj * p.dil) + dx, p.H, p.W);
                    // taps inside the staged window come from shared memory (zero-filled outside the frame, and the
                    // per-corner predicates still apply); the rare pixel whose taps leave the window reads global memory
                    const int ly = t.hl - wy0, lx = t.wl - wx0;
                    const bool in_win = use_win && ly >= 0 && ly + 1 < kWinH && lx >= 0 && lx + 1 < kWinW;
                    if (in_win) {
                        const float* wp = wb + ly * kWinW + lx;
#pragma unroll
                        for (int cl = 0; cl < CP; ++cl) {
                            float r = 0.0f;
                            if (cl < cin_g) {
                                const float* q = wp + cl * (kWinW * kWinH);
                                if (t.k1) r = t.w1 * q[0];
                                if (t.k2) r += t.w2 * q[1];
                                if (t.k3) r += t.w3 * q[kWinW];
                                if (t.k4) r += t.w4 * q[kWinW + 1];
                                r *= mval;
                            }
                            v[cl] = r;
                        }
                    } else {
#pragma unroll
                        for (int cl = 0; cl < CP; ++cl) v[cl] = cl < cin_g ? dcn_sample(xb + (int64_t)cl * HW + t.o, t, p.W) * mval : 0.0f;
                    }
                } else {
                    int prev_og = -1;
                    DcnTap t;
                    float mval = 1.0f;
#pragma unroll
                    for (int cl = 0; cl < CP; ++cl) {
                        v[cl] = 0.0f;
                        if (cl < cin_g) {
                            const int og = (g * cin_g + cl) / cpo;
                            if (og != prev_og) {
                                prev_og = og;
                                const float dy = ld_stream(offb + ((int64_t)og * 2 * taps + 2 * tap) * HWo);
                                const float dx = ld_stream(offb + ((int64_t)og * 2 * taps + 2 * tap + 1) * HWo);
                                mval = maskb ? ld_stream(maskb + ((int64_t)og * taps + tap) * HWo) : 1.0f;
                                t = dcn_tap(base_h + (float)(ki * p.dil) + dy, base_w + (float)(kj * p.dil) + dx, p.H, p.W);
                            }
                            v[cl] = dcn_sample(xb + (int64_t)cl * HW + t.o, t, p.W) * mval;
                        }
                    }
                }
                if (it >= NS) mbar_wait(&empty[stage], ((it / NS) - 1) & 1);
                uint8_t* st = a_ring + stage * A_STAGE + row_off;
#pragma unroll
                for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        uint4 hi, lo;
                        const float* vv = v + ks * 8 + j * 4;
                        hi.x = to_tf32_rna(vv[0]); hi.y = to_tf32_rna(vv[1]); hi.z = to_tf32_rna(vv[2]); hi.w = to_tf32_rna(vv[3]);
                        lo.x = __float_as_uint(vv[0] - __uint_as_float(hi.x));
                        lo.y = __float_as_uint(vv[1] - __uint_as_float(hi.y));
                        lo.z = __float_as_uint(vv[2] - __uint_as_float(hi.z));
                        lo.w = __float_as_uint(vv[3] - __uint_as_float(hi.w));
                        *reinterpret_cast<uint4*>(st + (0 * KS + ks) * A_PART + j * 128) = hi;
                        *reinterpret_cast<uint4*>(st + (1 * KS + ks) * A_PART + j * 128) = lo;
                    }
                }
                fence_proxy_async_smem();     // generic-proxy stores (slab + weights) -> visible to the tensor core
                mbar_arrive(&full[stage]);
            }
            if (use_win) mbar_arrive(&win_empty[g & 1]);   // this thread is done reading the group's window
        }
        // ================= epilogue: TMEM -> registers -> + bias -> NCHW =================
        mbar_wait(done, 0);
        tc_fence_after();
        float* ob = out + (int64_t)b * p.Cout * HWo + pix;
        for (int g = 0; g < p.WG; ++g) {
            uint32_t r[16];
            const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(g * kTcN);
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                  "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (live) {
#pragma unroll
                for (int o = 0; o < 16; ++o) {
                    if (o < cog) {
                        const int oc = g * cog + o;
                        st_stream(ob + (int64_t)oc * HWo, __uint_as_float(r[o]) + (bias ? __ldg(bias + oc) : 0.0f));
                    }
                }
            }
        }
    } else if (warp == 4 && lane == 0) {
        // ================= MMA issuer (one thread) =================
        const uint32_t a_base = smem_u32(a_ring), b_base = smem_u32(b_buf);
        int it = 0;
        for (int g = 0; g < p.WG; ++g) {
            const uint32_t bg = b_base + (g & 1) * b_bytes;
            const uint32_t d = tmem_base + (uint32_t)(g * kTcN);
            for (int tap = 0; tap < taps; ++tap, ++it) {
                const int stage = it % NS;
                mbar_wait(&full[stage], (it / NS) & 1);
                tc_fence_after();
#pragma unroll
                for (int ks = 0; ks < KS; ++ks) {
                    const uint64_t a_hi = umma_desc(a_base + stage * A_STAGE + (0 * KS + ks) * A_PART, 128, 256);
                    const uint64_t a_lo = umma_desc(a_base + stage * A_STAGE + (1 * KS + ks) * A_PART, 128, 256);
                    const uint64_t b_hi = umma_desc(bg + ((0 * taps + tap) * KS + ks) * 512, 128, 256);
                    const uint64_t b_lo = umma_desc(bg + ((1 * taps + tap) * KS + ks) * 512, 128, 256);
                    umma_tf32(d, a_hi, b_hi, (tap > 0 || ks > 0) ? 1u : 0u);
                    umma_tf32(d, a_hi, b_lo, 1u);
                    umma_tf32(d, a_lo, b_hi, 1u);
                }
                umma_commit(&empty[stage]);    // slab free once these MMAs have read it
            }
        }
        umma_commit(done);
    } else if (warp == 5 && lane == 0 && use_win) {
        // ================= TMA producer: one window (cin_g planes) per weight group, double buffered =================
        tma_prefetch_desc(&xmap);
        for (int g = 0; g < p.WG; ++g) {
            const int s = g & 1;
            if (g >= 2) mbar_wait(&win_empty[s], ((g >> 1) - 1) & 1);
            mbar_arrive_expect_tx(&win_full[s], win_bytes);
            tma_load_3d(win_buf + (size_t)s * win_bytes, &xmap, &win_full[s], wx0, wy0, b * p.Cin + g * cin_g);
        }
    }
    __syncwarp();
    tc_fence_before();
    __syncthreads();
    if (warp == 4) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
    }
}

static int tc_ks(const DcnParams& p) { return (p.Cin / p.WG) <= 8 ? 1 : 2; }

static size_t tc_smem_bytes(const DcnParams& p) {
    const int ks = tc_ks(p), taps = p.kh * p.kw;
    return (size_t)(ks == 1 ? 4 : 3) * 2 * ks * 4096 + 2 * (size_t)(2 * taps * ks * 512) + 256 +
           2 * (size_t)kWinW * kWinH * (p.Cin / p.WG) * 4 + 128;
}

// 0 = unsupported, 1 = supported, 2 = supported and preferred by algo 0 (auto)
int dcn_tc_supported(const DcnParams& p) {
    const int cin_g = p.Cin / p.WG, cog = p.Cout / p.WG, taps = p.kh * p.kw;
    if (cin_g > 16 || cog > kTcN || p.WG * kTcN > 512 || taps > 49) return 0;
    if ((int64_t)p.Ho * p.Wo >= (1ll << 31)) return 0;
    if (tc_smem_bytes(p) > 220 * 1024) return 0;
    return 1;
}

size_t dcn_tc_workspace_bytes(int, const DcnParams& p) {
    if (!dcn_tc_supported(p)) return 0;
    return (size_t)p.WG * 2 * p.kh * p.kw * tc_ks(p) * 512 + 256;
}

int launch_dcn_tc(const float* x, const float* offset, const float* mask, const float* weight, const float* bias, float* out, int B,
                  const DcnParams& p, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    VFI_REQUIRE(workspace && workspace_bytes >= dcn_tc_workspace_bytes(B, p), "vfi_dcn_f32: tcgen05 path needs a workspace of %zu B",
                dcn_tc_workspace_bytes(B, p));
    VFI_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "vfi_dcn_f32: workspace must be 16-byte aligned");
    const int ks = tc_ks(p), taps = p.kh * p.kw;
    const int cin_g = p.Cin / p.WG, cog = p.Cout / p.WG;
    float* wprep = reinterpret_cast<float*>(workspace);
    const int total = p.WG * 2 * taps * ks * 128;
    const int pblocks = (total + 255) / 256;
    if (ks == 1) dcn_tc_prep_kernel<1><<<pblocks, 256, 0, st>>>(weight, wprep, cin_g, cog, taps, p.WG);
    else dcn_tc_prep_kernel<2><<<pblocks, 256, 0, st>>>(weight, wprep, cin_g, cog, taps, p.WG);
    VFI_CHECK_LAUNCH("dcn_tc_prep_kernel");
    const int b_bytes = 2 * taps * ks * 512;
    const size_t smem = tc_smem_bytes(p);
    int tmem_cols = 32;
    while (tmem_cols < p.WG * kTcN) tmem_cols *= 2;
    const int tiles_x = (int)ceil_div(p.Wo, kTcTW), tiles_y = (int)ceil_div(p.Ho, kTcTH);
    const int tiles_per_img = tiles_x * tiles_y;
    const int64_t grid = (int64_t)B * tiles_per_img;
    VFI_REQUIRE(grid < (1ll << 31), "vfi_dcn_f32: grid too large");
    // input window staged by TMA: stride 1, kernel extent <= 3, rows 16-byte aligned
    CUtensorMap xmap;
    memset(&xmap, 0, sizeof(xmap));
    int use_win = p.stride == 1 && (p.kh - 1) * p.dil <= 2 && (p.kw - 1) * p.dil <= 2 && p.W % 4 == 0 &&
                  (reinterpret_cast<uintptr_t>(x) & 15) == 0 && (int64_t)B * p.Cin < (1ll << 31);
    if (const char* e = getenv("VFI_DCN_WINDOW")) use_win = use_win && e[0] != '0';
    if (use_win) {
        int rc = make_tmap_f32_3d(&xmap, x, (uint64_t)p.W, (uint64_t)p.H, (uint64_t)B * p.Cin, (uint64_t)p.W * 4,
                                  (uint64_t)p.H * p.W * 4, kWinW, kWinH, (uint32_t)cin_g);
        if (rc) return rc;
    }
    if (ks == 1) {
        VFI_CUDA(cudaFuncSetAttribute(dcn_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        dcn_tc_kernel<1><<<(unsigned)grid, kTcThreads, smem, st>>>(x, offset, mask, wprep, bias, out, p, tiles_x, tiles_per_img, b_bytes,
                                                                  tmem_cols, xmap, use_win);
    } else {
        VFI_CUDA(cudaFuncSetAttribute(dcn_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        dcn_tc_kernel<2><<<(unsigned)grid, kTcThreads, smem, st>>>(x, offset, mask, wprep, bias, out, p, tiles_x, tiles_per_img, b_bytes,
                                                                  tmem_cols, xmap, use_win);
    }
    VFI_CHECK_LAUNCH("dcn_tc_kernel");
    return VFI_OK;
}

}  // namespace vfi
