// 3x3 / stride 1 / pad 1 convolution of the DAT conv stacks (the convs that feed and consume
// the hot path -- modules/basic_blocks.py:13-73, modules/cnn_encoders.py:6-42, modules/generator.py:7-24,
// modules/deformable_attn.py:56-64) as an implicit GEMM on the 5th-gen tensor cores, TF32 inputs / fp32 accumulate
// (the arithmetic PyTorch's default `cudnn.allow_tf32 = True` gives the reference), with the bias + residual + PReLU
// epilogue of epilogue.cu fused in.
//
// Persistent kernel, one CTA per SM, 6 warps:
//   warp 0  producer : per tile and per 8-channel chunk ONE 5-D TMA box {4 ch, 10 x, 18 y, 2 ch-quads, 1 image} of the
//                      NHWC input (zero-filled outside the image = the conv padding) into a ring of stages.  The box
//                      lands as [ch-quad][y][x][16 B], i.e. 8 horizontally adjacent pixels x 16 B = one UMMA core
//                      matrix, so the 9 taps of the 3x3 stencil are 9 *descriptor offsets* into the same stage --
//                      no im2col, every input byte is fetched once per tile (+ halo).
//   warp 1  MMA      : tcgen05.mma.cta_group::1.kind::tf32, M = 128 pixels (8 x 16 tile), N = Cout tile, K = 8 per
//                      instruction; 9 taps x Cin/8 chunks accumulate into one of two TMEM accumulators.  The whole
//                      weight tensor of the Cout tile (<= 187 KB, pre-rounded to TF32 and pre-arranged by
//                      conv3x3_prep_kernel) stays resident in shared memory for the life of the CTA.
//   warps 2-5 epilogue: tcgen05.ld (lane = pixel) -> bias, residual, PReLU -> global; overlaps the next tile's MMAs
//                      through the second accumulator.
#include <stdlib.h>

#include "umma.cuh"
#include "vfi_common.cuh"

namespace vfi {

constexpr int kCvTW = 8, kCvTH = 16;                    // output tile: an 8-pixel row segment = one 8-row core-matrix group
constexpr int kCvM = kCvTW * kCvTH;                     // 128 = UMMA M
constexpr int kCvWinW = kCvTW + 2, kCvWinH = kCvTH + 2;
constexpr int kCvStageBytes = 2 * kCvWinW * kCvWinH * 16;   // one 8-channel chunk of the window: 5760 B
constexpr int kCvThreads = 192;
constexpr int kCvMaxStages = 12;
constexpr int kCvSmemMax = 232448;                      // 227 KB opt-in limit
constexpr int kCvAccStride = 128;                       // TMEM columns between the two accumulators

struct ConvArgs {
    const float* wprep;   // this Cout tile: [tap 9][chunk][k-half 2][n NPAD][4] floats
    const float* bias;    // cn floats or null
    const float* slope;   // cn floats or null
    const float* res;     // residual (same layout family as out) or null
    float* out;
    int B, H, W, nchunk, cn;
    int res_mode;         // 0 none, 1 add before the activation, 2 add after it
    int vec;              // NHWC out/res with 16-byte aligned pixel rows: float4 path
    int64_t out_sb, out_sy, out_sx, out_sc;   // element strides of out
    int64_t res_sb, res_sy, res_sx, res_sc;
    int stages, tiles_x, tiles_y, n_tiles;
    int dbg;              // bring-up switches (VFI_CONV_DBG): 1 = no TMA loads, 2 = no epilogue stores, 4 = no MMAs
};

// weight (Cout, Cin, 3, 3) contiguous -> per Cout tile of T channels: [tap][chunk][k-half][n NPAD][4], RN-rounded to TF32
__global__ void conv3x3_prep_kernel(const float* __restrict__ w, float* __restrict__ wprep, int Cout, int Cin, int T, int NPAD,
                                    int n_nt) {
    const int nchunk = (Cin + 7) / 8;
    const int64_t per_tile = (int64_t)9 * nchunk * 2 * NPAD * 4;
    const int64_t total = per_tile * n_nt;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i;
        const int e = (int)(r % 4); r /= 4;
        const int n = (int)(r % NPAD); r /= NPAD;
        const int kh = (int)(r % 2); r /= 2;
        const int c = (int)(r % nchunk); r /= nchunk;
        const int tap = (int)(r % 9); r /= 9;
        const int nt = (int)r;
        const int co = nt * T + n, ci = c * 8 + kh * 4 + e;
        float v = 0.0f;
        if (n < T && co < Cout && ci < Cin) v = w[((int64_t)co * Cin + ci) * 9 + tap];
        uint32_t t;
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(t) : "f"(v));
        wprep[i] = __uint_as_float(t);
    }
}

__device__ __forceinline__ bool elect_one() {
    uint32_t p;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
    return p != 0;
}

template <int NPAD>
__global__ void __launch_bounds__(kCvThreads, 1) conv3x3_tc_kernel(const ConvArgs a, const __grid_constant__ CUtensorMap xmap) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t per_tap = (uint32_t)a.nchunk * NPAD * 32;
    const uint32_t wbytes = 9 * per_tap;
    uint8_t* b_s = smem;
    uint8_t* a_ring = b_s + wbytes;
    float* bias_s = reinterpret_cast<float*>(a_ring + a.stages * kCvStageBytes);
    float* slope_s = bias_s + NPAD;
    uint64_t* full = reinterpret_cast<uint64_t*>(slope_s + NPAD);
    uint64_t* empty = full + kCvMaxStages;
    uint64_t* wfull = empty + kCvMaxStages;
    uint64_t* tfull = wfull + 1;     // [2]
    uint64_t* tempty = tfull + 2;    // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < a.stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        mbar_init(wfull, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tfull[s], 1);
            mbar_init(&tempty[s], kCvM);
        }
        fence_mbar_init();
        tma_prefetch_desc(&xmap);
    }
    for (int i = tid; i < NPAD; i += kCvThreads) {
        bias_s[i] = (a.bias && i < a.cn) ? a.bias[i] : 0.0f;
        slope_s[i] = (a.slope && i < a.cn) ? a.slope[i] : 1.0f;
    }
    if (warp == 1) umma::tmem_alloc(tmem_slot, 2 * kCvAccStride);
    umma::fence_before();
    __syncthreads();
    umma::fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int tiles_per_img = a.tiles_x * a.tiles_y;

    if (warp == 0) {
        // ---- producer (whole warp converged, one elected lane issues): resident weights once, then the window ring ----
        if (elect_one()) {
            mbar_arrive_expect_tx(wfull, wbytes);
            for (int t = 0; t < 9; ++t)
                bulk_load_1d(b_s + t * per_tap, reinterpret_cast<const uint8_t*>(a.wprep) + (size_t)t * per_tap, per_tap, wfull);
        }
        uint32_t s = 0, ph = 0;
        for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
            const int b = tile / tiles_per_img, r = tile - b * tiles_per_img;
            const int ty = r / a.tiles_x, tx = r - ty * a.tiles_x;
            for (int c = 0; c < a.nchunk; ++c) {
                mbar_wait(&empty[s], ph ^ 1);
                if (elect_one()) {
                    if (a.dbg & 1) {
                        mbar_arrive(&full[s]);
                    } else {
                        mbar_arrive_expect_tx(&full[s], kCvStageBytes);
                        tma_load_5d(a_ring + s * kCvStageBytes, &xmap, &full[s], 0, tx * kCvTW - 1, ty * kCvTH - 1, 2 * c, b);
                    }
                }
                __syncwarp();
                if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ---- MMA issuer (whole warp converged; descriptors are warp-uniform, one elected lane issues) ----
        constexpr uint32_t idesc = umma::idesc_tf32(kCvM, NPAD);
        constexpr uint32_t lbo_a = kCvWinW * kCvWinH * 16, sbo_a = kCvWinW * 16;
        // descriptors differ only in the 14-bit start-address field: keep the invariant words and add offsets >> 4
        const uint64_t a_desc0 = umma::desc_kmajor(smem_u32(a_ring), (a.dbg & 16) ? 2048 : lbo_a, (a.dbg & 16) ? 128 : sbo_a);
        const uint64_t b_desc0 = umma::desc_kmajor(smem_u32(b_s), NPAD * 16, 128);
        mbar_wait(wfull, 0);
        uint32_t s = 0, ph = 0, i = 0;
        for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, ++i) {
            const uint32_t acc = i & 1;
            mbar_wait(&tempty[acc], ((i >> 1) & 1) ^ 1);
            umma::fence_after();
            const uint32_t d = tmem_base + acc * kCvAccStride;
            for (int c = 0; c < a.nchunk; ++c) {
                mbar_wait(&full[s], ph);
                umma::fence_after();
                if (elect_one()) {
                    const uint64_t ad0 = a_desc0 + (uint64_t)((s * kCvStageBytes) >> 4);
                    const uint64_t bd0 = b_desc0 + (uint64_t)((c * (NPAD * 32)) >> 4);
                    if (!(a.dbg & 4)) {
#pragma unroll
                        for (int tap = 0; tap < 9; ++tap) {
                            const uint64_t ad = ad0 + ((a.dbg & 32) ? (uint64_t)((tap / 3) * 8) : (uint64_t)(((tap / 3) * kCvWinW + (tap % 3))));
                            const uint64_t bd = bd0 + (uint64_t)((tap * per_tap) >> 4);
                            umma::mma_tf32((a.dbg & 8) ? tmem_base + (tap & 1) * kCvAccStride : d, ad, bd, idesc, tap == 0 ? (uint32_t)(c != 0) : 1u);
                        }
                    }
                    umma::commit(&empty[s]);
                    if (c == a.nchunk - 1) umma::commit(&tfull[acc]);
                }
                __syncwarp();
                if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1; }
            }
        }
    } else {
        // ---- epilogue warps: TMEM lane group = warp % 4; lane = pixel ----
        const int q = warp & 3;
        const int m = q * 32 + lane, ly = m >> 3, lx = m & 7;
        uint32_t i = 0;
        for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, ++i) {
            const int b = tile / tiles_per_img, r = tile - b * tiles_per_img;
            const int ty = r / a.tiles_x, tx = r - ty * a.tiles_x;
            const uint32_t acc = i & 1;
            mbar_wait(&tfull[acc], (i >> 1) & 1);
            __syncwarp();                  // ragged tiles leave lanes diverged; tcgen05.ld is .sync.aligned
            umma::fence_after();
            float v[NPAD];
            const uint32_t taddr = tmem_base + acc * kCvAccStride + ((uint32_t)(q * 32) << 16);
#pragma unroll
            for (int j = 0; j < NPAD / 8; ++j) umma::tmem_ld8(taddr + j * 8, v + 8 * j);
            umma::tmem_ld_wait();
            umma::fence_before();
            mbar_arrive(&tempty[acc]);     // accumulator is in registers: the MMA warp may overwrite it
            const int y = ty * kCvTH + ly, x = tx * kCvTW + lx;
            if (y >= a.H || x >= a.W || (a.dbg & 2)) continue;
            float* o = a.out + b * a.out_sb + y * a.out_sy + x * a.out_sx;
            const float* rp = a.res ? a.res + b * a.res_sb + y * a.res_sy + x * a.res_sx : nullptr;
            if (a.vec) {
#pragma unroll
                for (int j = 0; j < NPAD / 4; ++j) {
                    if (4 * j < a.cn) {
                        float4 rr = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (a.res_mode) rr = __ldg(reinterpret_cast<const float4*>(rp) + j);
                        const float rv[4] = {rr.x, rr.y, rr.z, rr.w};
                        float t[4];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            t[k] = __fadd_rn(v[4 * j + k], bias_s[4 * j + k]);
                            if (a.res_mode == 1) t[k] = __fadd_rn(rv[k], t[k]);
                            if (a.slope) t[k] = t[k] > 0.0f ? t[k] : __fmul_rn(slope_s[4 * j + k], t[k]);
                            if (a.res_mode == 2) t[k] = __fadd_rn(rv[k], t[k]);
                        }
                        reinterpret_cast<float4*>(o)[j] = make_float4(t[0], t[1], t[2], t[3]);
                    }
                }
            } else {
#pragma unroll
                for (int c = 0; c < NPAD; ++c) {
                    if (c < a.cn) {
                        float t = __fadd_rn(v[c], bias_s[c]);
                        const float rv = a.res_mode ? __ldg(rp + c * a.res_sc) : 0.0f;
                        if (a.res_mode == 1) t = __fadd_rn(rv, t);
                        if (a.slope) t = t > 0.0f ? t : __fmul_rn(slope_s[c], t);
                        if (a.res_mode == 2) t = __fadd_rn(rv, t);
                        o[c * a.out_sc] = t;
                    }
                }
            }
        }
    }
    umma::fence_before();
    __syncthreads();
    if (warp == 1) umma::tmem_dealloc(tmem_base, 2 * kCvAccStride);
}

// ---- host side -------------------------------------------------------------------------------------------------
static int conv_n_gran() {
    static int g = 0;
    if (!g) {
        const char* e = getenv("VFI_CONV_NGRAN");
        g = (e && atoi(e) == 8) ? 8 : 16;
    }
    return g;
}
struct ConvTiling {
    int T, NPAD, n_nt;
};
static ConvTiling conv_tiling(int Cout) {
    const int g = conv_n_gran();
    ConvTiling t;
    if (Cout <= 80) t.T = Cout;                         // one tile
    else if (g == 8 && Cout % 72 == 0) t.T = 72;
    else t.T = 64;
    const int sizes[6] = {16, 32, 48, 64, 72, 80};      // instantiated N (72 needs the 8-granular UMMA N)
    t.NPAD = 80;
    for (int i = 5; i >= 0; --i)
        if (sizes[i] >= t.T && (sizes[i] % g) == 0) t.NPAD = sizes[i];
    t.n_nt = (Cout + t.T - 1) / t.T;
    return t;
}
static size_t conv_smem_fixed(int NPAD) { return 1024 + 2 * NPAD * 4 + (2 * kCvMaxStages + 5) * 8 + 16; }

template <int NPAD>
static int conv_launch(const ConvArgs& a, const CUtensorMap& xmap, size_t smem, cudaStream_t st) {
    static bool attr = false;
    if (!attr) {
        VFI_CUDA(cudaFuncSetAttribute(conv3x3_tc_kernel<NPAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCvSmemMax));
        attr = true;
    }
    const int grid = a.n_tiles < kNumSMs ? a.n_tiles : kNumSMs;
    conv3x3_tc_kernel<NPAD><<<grid, kCvThreads, smem, st>>>(a, xmap);
    VFI_CHECK_LAUNCH("conv3x3_tc_kernel");
    return VFI_OK;
}

}  // namespace vfi

extern "C" size_t vfi_conv3x3_prep_bytes(int Cout, int Cin) {
    using namespace vfi;
    if (Cout < 1 || Cin < 4 || Cin % 4) return 0;
    const ConvTiling t = conv_tiling(Cout);
    return (size_t)t.n_nt * 9 * ((Cin + 7) / 8) * 2 * t.NPAD * 4 * sizeof(float);
}

extern "C" int vfi_conv3x3_prep_f32(const float* weight, float* wprep, int Cout, int Cin, void* stream) {
    using namespace vfi;
    VFI_REQUIRE(weight && wprep, "vfi_conv3x3_prep_f32: null pointer");
    VFI_REQUIRE(Cout >= 1 && Cin >= 4 && Cin % 4 == 0, "vfi_conv3x3_prep_f32: Cin %d must be a positive multiple of 4", Cin);
    const ConvTiling t = conv_tiling(Cout);
    const int64_t total = (int64_t)t.n_nt * 9 * ((Cin + 7) / 8) * 2 * t.NPAD * 4;
    const int blocks = (int)min64(ceil_div(total, 256), (int64_t)kNumSMs * 8);
    conv3x3_prep_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(weight, wprep, Cout, Cin, t.T, t.NPAD, t.n_nt);
    VFI_CHECK_LAUNCH("conv3x3_prep_kernel");
    return VFI_OK;
}

extern "C" int vfi_conv3x3_supported(int Cin, int x_pix_stride, int Cout) {
    using namespace vfi;
    if (Cin < 4 || Cin % 4 || x_pix_stride % 4 || x_pix_stride < Cin || Cout < 1) return 0;
    const ConvTiling t = conv_tiling(Cout);
    const size_t wbytes = (size_t)9 * ((Cin + 7) / 8) * t.NPAD * 32;
    return wbytes + conv_smem_fixed(t.NPAD) + 3 * (size_t)kCvStageBytes <= (size_t)kCvSmemMax;
}

extern "C" int vfi_conv3x3_f32(const float* x, const float* wprep, const float* bias, const float* slope, const float* residual,
                               float* out, int B, int H, int W, int Cin, int x_pix_stride, int Cout, int out_channels_last,
                               int out_c_total, int res_c_total, int residual_mode, void* stream) {
    using namespace vfi;
    VFI_REQUIRE(x && wprep && out, "vfi_conv3x3_f32: null pointer");
    VFI_REQUIRE(B >= 0 && H >= 1 && W >= 1, "vfi_conv3x3_f32: bad shape");
    VFI_REQUIRE(vfi_conv3x3_supported(Cin, x_pix_stride, Cout), "vfi_conv3x3_f32: unsupported Cin %d / pixel stride %d / Cout %d",
                Cin, x_pix_stride, Cout);
    VFI_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0, "vfi_conv3x3_f32: x must be 16-byte aligned");
    VFI_REQUIRE(residual_mode >= 0 && residual_mode <= 2 && (residual_mode == 0 || residual), "vfi_conv3x3_f32: residual_mode %d",
                residual_mode);
    VFI_REQUIRE(out_c_total >= Cout && (residual_mode == 0 || res_c_total >= Cout), "vfi_conv3x3_f32: channel totals");
    if (B == 0) return VFI_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const ConvTiling t = conv_tiling(Cout);
    CUtensorMap xmap;
    {
        const uint64_t dims[5] = {4, (uint64_t)W, (uint64_t)H, (uint64_t)(Cin / 4), (uint64_t)B};
        const uint64_t strides[4] = {(uint64_t)x_pix_stride * 4, (uint64_t)W * x_pix_stride * 4, 16,
                                     (uint64_t)H * W * x_pix_stride * 4};
        const uint32_t box[5] = {4, (uint32_t)kCvWinW, (uint32_t)kCvWinH, 2, 1};
        const int rc = make_tmap_f32_nd(&xmap, x, 5, dims, strides, box);
        if (rc != VFI_OK) return rc;
    }
    ConvArgs a = {};
    {
        const char* e = getenv("VFI_CONV_DBG");
        a.dbg = e ? atoi(e) : 0;
    }
    a.B = B; a.H = H; a.W = W; a.nchunk = (Cin + 7) / 8;   // a trailing half chunk is zero-filled by TMA (and by prep)
    a.res_mode = residual_mode;
    a.tiles_x = (W + kCvTW - 1) / kCvTW; a.tiles_y = (H + kCvTH - 1) / kCvTH;
    a.n_tiles = B * a.tiles_x * a.tiles_y;
    const int64_t HW = (int64_t)H * W;
    if (out_channels_last) {
        a.out_sb = HW * out_c_total; a.out_sy = (int64_t)W * out_c_total; a.out_sx = out_c_total; a.out_sc = 1;
        a.res_sb = HW * res_c_total; a.res_sy = (int64_t)W * res_c_total; a.res_sx = res_c_total; a.res_sc = 1;
    } else {
        a.out_sb = HW * out_c_total; a.out_sy = W; a.out_sx = 1; a.out_sc = HW;
        a.res_sb = HW * res_c_total; a.res_sy = W; a.res_sx = 1; a.res_sc = HW;
    }
    const size_t wbytes = (size_t)9 * a.nchunk * t.NPAD * 32;
    int stages = (int)(((size_t)kCvSmemMax - conv_smem_fixed(t.NPAD) - wbytes) / kCvStageBytes);
    if (stages > kCvMaxStages) stages = kCvMaxStages;
    a.stages = stages;
    const size_t smem = wbytes + (size_t)stages * kCvStageBytes + conv_smem_fixed(t.NPAD);
    for (int nt = 0; nt < t.n_nt; ++nt) {
        const int c0 = nt * t.T;
        a.cn = Cout - c0 < t.T ? Cout - c0 : t.T;
        a.wprep = wprep + (size_t)nt * (wbytes / 4);
        a.bias = bias ? bias + c0 : nullptr;
        a.slope = slope ? slope + c0 : nullptr;
        a.out = out + c0 * a.out_sc;
        a.res = residual ? residual + c0 * a.res_sc : nullptr;
        const bool al = ((reinterpret_cast<uintptr_t>(a.out) & 15) == 0) && (a.res == nullptr || (reinterpret_cast<uintptr_t>(a.res) & 15) == 0);
        a.vec = out_channels_last && al && a.cn % 4 == 0 && out_c_total % 4 == 0 && (residual_mode == 0 || res_c_total % 4 == 0);
        int rc;
        switch (t.NPAD) {
            case 16: rc = conv_launch<16>(a, xmap, smem, st); break;
            case 32: rc = conv_launch<32>(a, xmap, smem, st); break;
            case 48: rc = conv_launch<48>(a, xmap, smem, st); break;
            case 64: rc = conv_launch<64>(a, xmap, smem, st); break;
            case 72: rc = conv_launch<72>(a, xmap, smem, st); break;
            case 80: rc = conv_launch<80>(a, xmap, smem, st); break;
            default:
                set_error("vfi_conv3x3_f32: no kernel for a Cout tile padded to %d", t.NPAD);
                return VFI_ERR_UNSUPPORTED;
        }
        if (rc != VFI_OK) return rc;
    }
    return VFI_OK;
}
