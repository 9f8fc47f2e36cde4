// 3x3 / stride 1 / pad 1 convolution of the DAT conv stacks (the convs that feed and consume
// the hot path -- modules/basic_blocks.py:13-73, modules/cnn_encoders.py:6-42, modules/generator.py:7-24,
// modules/deformable_attn.py:56-64) as an implicit GEMM on the 5th-gen tensor cores, TF32 inputs / fp32 accumulate
// (the arithmetic PyTorch's default `cudnn.allow_tf32 = True` gives the reference), with the bias + residual + PReLU
// epilogue of epilogue.cu fused in.
//
// Formulation.  With shared-memory operands tcgen05.mma streams the M-side operand at ~64 B/clk, i.e. an
// M = 128, K = 8 (tf32) instruction costs >= 64 clk whatever N is (measured: 75 + 0.25 N clk), so N = Cout = 72 leaves
// the tensor core two thirds idle.  The three vertical taps are therefore concatenated along N:
//     P[dy][r, x, co] = sum_dx sum_ci X[r, x + dx, ci] * W[co, ci, dy, dx]         (one MMA chain, N = 3 * Cout)
//     out[y, x, co]   = P[-1][y - 1, x, co] + P[0][y, x, co] + P[+1][y + 1, x, co]
// The tile is 8 columns x 16 input rows stored COLUMN-major (M index = 16 * column + row), so the row shift of the
// second line is a +-1 lane shift inside a half-warp of the epilogue (two __shfl per value) and the first and last
// row of the tile are halo: 14 output rows per tile.  MMAs per tile: 3 (dx) x Cin/8 with N = 216 instead of
// 9 x Cin/8 with N = 72.
//
// Persistent kernel, one CTA per SM, 10 warps:
//   warp 0  producer : per tile and per 16-channel stage ONE 4-D TMA box {16 ch, 16 y, 10 x, 1 image} of the NHWC
//                      input (zero-filled outside the image = the conv padding; SWIZZLE_64B).  It lands as
//                      [x][y][64 B] = 160 pixel rows in UMMA's K-major 64-byte-swizzle layout, so the horizontal taps
//                      are descriptor offsets of 1024 B and the second k-step of a stage is +32 B -- no im2col.
//   warp 1  MMA      : tcgen05.mma.cta_group::1.kind::tf32, M = 128, N = 3 * (Cout tile), accumulating in one of two
//                      TMEM accumulators.  The weights of the Cout tile (<= 187 KB, RN-rounded to TF32 and arranged by
//                      conv3x3_prep_kernel) stay resident in shared memory for the life of the CTA.
//   warps 2-9 epilogue: tcgen05.ld (lane = pixel; two warps per TMEM lane quadrant, alternate 8-channel chunks) ->
//                      plane sum -> bias, residual, PReLU -> global (one 32-byte sector per thread and chunk);
//                      overlaps the next tile's MMAs through the second accumulator.
#include <stdlib.h>

#include "umma.cuh"
#include "vfi_common.cuh"

namespace vfi {

// set by vfi_conv3x3_prep_f32, consumed by the next conv launch of this host thread (see the PDL note at the launch)
static thread_local bool g_prep_pending = false;

constexpr int kCvTW = 8, kCvRows = 16;                  // tile: 8 columns x 16 input rows (M = 128, column-major)
constexpr int kCvTH = kCvRows - 2;                      // output rows per tile
constexpr int kCvM = kCvTW * kCvRows;
constexpr int kCvWinW = kCvTW + 2;
constexpr int kCvKSteps = 2;                            // 8-channel k-steps per stage (16 channels = one 64-byte swizzle row)
constexpr int kCvStageBytes = kCvWinW * kCvRows * 64;   // 160 pixel rows x 64 B = 10240 B
#ifndef VFI_CONV_EPI_WARPS
#define VFI_CONV_EPI_WARPS 12
#endif
constexpr int kCvEpiWarps = VFI_CONV_EPI_WARPS;         // a multiple of 4: kCvEpiWarps / 4 warps share each TMEM lane quadrant
constexpr int kCvThreads = 96 + 32 * kCvEpiWarps;       // producer warp, MMA warp, epilogue warps, store warp
constexpr int kCvMaxStages = 12;
constexpr int kCvSmemMax = 232448;                      // 227 KB opt-in limit
constexpr int kCvAccStride = 256;                       // TMEM columns between the two accumulators

#ifdef VFI_CONV_TRACE
// bring-up only (tools/build_variant.sh ... -DVFI_CONV_TRACE): clock64 samples of every role of CTA 0, printed at the end
__device__ long long g_tr[12][64];
__device__ long long g_ph[8][2];      // phase marks of CTA 0: (clock64, globaltimer ns)
__device__ __forceinline__ void cv_phase(int k) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        long long g;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
        g_ph[k][0] = clock64();
        g_ph[k][1] = g;
    }
}
#define CV_PHASE(k) cv_phase(k)
#define CV_TRACE(role, idx)                                                            \
    do {                                                                               \
        if (blockIdx.x == 0 && (idx) < 64) g_tr[role][idx] = clock64();                \
    } while (0)
#else
#define CV_TRACE(role, idx) do { } while (0)
#define CV_PHASE(k) do { } while (0)
#endif

__host__ __device__ static inline int conv_tile_pitch(int cn);

struct ConvArgs {
    const float* wprep;   // this Cout tile: [dx 3][chunk nchunkP][k-half 2][n = dy * NPAD + co][4] floats
    const float* bias;    // cn floats (element stride bias_stride) or null
    const float* slope;   // cn floats or null
    const float* res;     // residual (same layout family as out) or null
    float* out;
    int B, H, W, nstage_tile, nchunk, cn;   // nstage_tile = stages (16 channels) per tile, nchunk = 8-channel k-steps per tile
    int res_mode;         // 0 none, 1 add before the activation, 2 add after it
    int res_B;            // images in the residual: output image b adds residual image b % res_B (B unless broadcast)
    int vec;              // NHWC out/res: 1 = 16-byte aligned pixel rows (float4), 2 = 32-byte aligned (256-bit accesses)
    int64_t out_sb, out_sy, out_sx, out_sc;   // element strides of out
    int64_t res_sb, res_sy, res_sx, res_sc;
    int stages, tiles_x, tiles_y, n_tiles;
    int bias_stride;
    int staged;           // NHWC residual tiles are prefetched into shared memory by TMA (rmap), one tile ahead of the epilogue
    int ostg;             // NHWC output tiles leave through shared memory and one TMA store per tile (omap)
    int dbg;              // bring-up switches (VFI_CONV_DBG): 1 = no TMA loads, 2 = no epilogue stores, 4 = no MMAs, 8 = no epilogue work
};

// weight (Cout, Cin, 3, 3) contiguous -> per Cout tile of T channels:
// [dx][chunk nchunkP][k-half][n = dy * NPAD + co][4], RN-rounded to TF32, zero where co / ci are padding
// up2 > 1 (fused PixelShuffle): tiles are ordered [sub-position][channel tile] and row nn of a tile is conv channel
// (tile_c * T + nn) * up2 + sub, i.e. every tile produces all its channels for ONE of the up2 output pixels.
// ctas = 2: the N rows (dy * NPAD + co, padded to `ntot`) are split in two halves, one block per CTA of the pair:
// [tile][cta][dx][chunk][k-half][n_local ntot/ctas][4].
__global__ void conv3x3_prep_kernel(const float* __restrict__ w, float* __restrict__ wprep, int Cout, int Cin, int T, int NPAD,
                                    int n_nt, int nchunkP, int up2, int ntot, int ctas) {
    const int nh = ntot / ctas;
    const int64_t per_tile = (int64_t)ctas * 3 * nchunkP * 2 * nh * 4;
    const int64_t total = per_tile * n_nt;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i;
        const int e = (int)(r % 4); r /= 4;
        const int nl = (int)(r % nh); r /= nh;
        const int kh = (int)(r % 2); r /= 2;
        const int c = (int)(r % nchunkP); r /= nchunkP;
        const int dx = (int)(r % 3); r /= 3;
        const int cta = (int)(r % ctas); r /= ctas;
        const int nt = (int)r;
        const int n = cta * nh + nl;
        const int dy = n / NPAD, nn = n - dy * NPAD;
        const int n_ct = n_nt / up2, sub = nt / n_ct, tc = nt - sub * n_ct;
        const int cc = tc * T + nn;                      // channel after the shuffle (== conv channel when up2 == 1)
        const int co = cc * up2 + sub, ci = c * 8 + kh * 4 + e;
        float v = 0.0f;
        if (dy < 3 && nn < T && cc < Cout / up2 && ci < Cin) v = w[((int64_t)co * Cin + ci) * 9 + dy * 3 + dx];
        uint32_t t;
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(t) : "f"(v));
        wprep[i] = __uint_as_float(t);
    }
}

// 256-bit global accesses (sm_100): one full 32-byte sector per thread
__device__ __forceinline__ void ld_global_v8(const float* p, float* v) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "l"(p));
}
__device__ __forceinline__ void st_global_v8(float* p, const float* v) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7])
                 : "memory");
}

__device__ __forceinline__ bool elect_one() {
    uint32_t p;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
    return p != 0;
}

// CTAS = 2: the kernel runs as CTA pairs (cluster of 2, tcgen05 cta_group::2): one MMA covers the two tiles of the pair
// (M = 256), every CTA keeps only HALF of the weight rows in shared memory, and the tensor cores of each SM read
// half as many B bytes per tile -- the shared-memory operand feed is what bounds this kernel.
// RSTG: the residual tile [8 x][14 y][cn] of every output tile is fetched by the producer warp with one TMA box into one
// of two shared-memory buffers while the previous tile is still in its epilogue (a residual read straight from global
// memory in the epilogue exposes one HBM latency per 8-channel chunk: 157 vs 95 us for 72 -> 72 at 544x960).
// OSTG: the epilogue warps write their results into a shared-memory tile [8 x][14 y][cn] (over the residual tile when
// RSTG: same thread, same address) and a dedicated store warp sends it out with ONE TMA store per tile.  The direct path
// (lane = pixel, one 32-byte sector per thread and chunk) makes every warp store touch 32 lines: with the stores switched
// off the 72 -> 72 conv runs 84 -> 68 us, i.e. the LSU store path, not the tensor core, was the longest stage.
template <int NPAD, int CTAS, bool RSTG, bool OSTG>
__global__ void __launch_bounds__(kCvThreads, 1) conv3x3_tc_kernel(const ConvArgs a, const __grid_constant__ CUtensorMap xmap,
                                                                   const __grid_constant__ CUtensorMap rmap,
                                                                   const __grid_constant__ CUtensorMap omap) {
    constexpr int N = CTAS == 1 ? 3 * NPAD : (3 * NPAD + 15) / 16 * 16;    // UMMA N (cta_group::2 wants a multiple of 16)
    constexpr int NH = N / CTAS;                                          // B rows held by this CTA
    const uint32_t rank = CTAS == 2 ? cluster_ctarank() : 0u;
    const bool leader = rank == 0;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int nchunkP = a.nchunk;
    constexpr uint32_t slab = 2 * NH * 16;                      // one (dx, chunk) B slab of this CTA
    const uint32_t per_dx = (uint32_t)nchunkP * slab;
    const uint32_t wbytes = 3 * per_dx;
    uint8_t* b_s = smem;
    uint8_t* a_ring = b_s + ((wbytes + 1023u) & ~1023u);      // swizzled stages: 1024-byte aligned
    float* bias_s = reinterpret_cast<float*>(a_ring + a.stages * kCvStageBytes);
    float* slope_s = bias_s + NPAD;
    uint64_t* full = reinterpret_cast<uint64_t*>(slope_s + NPAD);
    uint64_t* empty = full + kCvMaxStages;
    uint64_t* wfull = empty + kCvMaxStages;
    uint64_t* tfull = wfull + 1;     // [2]
    uint64_t* tempty = tfull + 2;    // [2]
    uint64_t* pfull = tempty + 2;    // [stages] CTA pairs: "the peer's stage s is full", signalled by the peer's forwarding warp
    uint64_t* rfull = pfull + kCvMaxStages;   // [2] RSTG: residual tile landed in its buffer
    uint64_t* rempty = rfull + 2;             // [2] RSTG: the 8 epilogue warps are done with the buffer
    uint64_t* ofull = rempty + 2;             // [2] OSTG: every epilogue warp has written its part of the output tile
    uint64_t* oempty = ofull + 2;             // [2] OSTG without a residual: the TMA store has read the buffer
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(oempty + 2);
    // residual buffers: 2 x [8 x][14 y][cn] floats, 128-byte aligned, after the barriers
    float* rbuf = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(tmem_slot + 1) + 127) & ~(uintptr_t)127);
    const int rpitch = conv_tile_pitch(a.cn);
    const int rfloats = kCvTW * kCvTH * rpitch;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    CV_PHASE(0);
    if (tid == 0) {
        for (int s = 0; s < a.stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
            mbar_init(&pfull[s], 1);
        }
        mbar_init(wfull, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tfull[s], 1);
            mbar_init(&tempty[s], CTAS * kCvEpiWarps);  // one arrival per epilogue warp (of both CTAs, on the leader's barrier)
            mbar_init(&rfull[s], 1);
            mbar_init(&rempty[s], OSTG ? 1 : kCvEpiWarps);   // OSTG: released by the store warp once the tile has left
            mbar_init(&ofull[s], kCvEpiWarps);
            mbar_init(&oempty[s], 1);
        }
        fence_mbar_init();
        tma_prefetch_desc(&xmap);
        if (RSTG) tma_prefetch_desc(&rmap);
        if (OSTG) tma_prefetch_desc(&omap);
    }
    __syncthreads();      // barriers initialised before any thread of this CTA touches them
    if (warp == 1) {
        if (CTAS == 2) umma::tmem_alloc_2sm(tmem_slot, 2 * kCvAccStride);
        else umma::tmem_alloc(tmem_slot, 2 * kCvAccStride);
    }
    if (warp == 0 && elect_one()) {      // resident weights of this CTA
        mbar_arrive_expect_tx(wfull, wbytes);
        const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(a.wprep) + (size_t)rank * wbytes;
        for (int t = 0; t < 3; ++t) bulk_load_1d(b_s + t * per_dx, wsrc + (size_t)t * per_dx, per_dx, wfull);
    }
    // PDL: let the next conv of the stream start its own weight prologue on the SMs this grid frees first, and do not
    // read or write an activation before the previous kernel has completed and flushed
    CV_PHASE(1);
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    CV_PHASE(2);
    for (int i = tid; i < NPAD; i += kCvThreads) {
        bias_s[i] = (a.bias && i < a.cn) ? a.bias[(size_t)i * a.bias_stride] : 0.0f;
        slope_s[i] = (a.slope && i < a.cn) ? a.slope[i] : 1.0f;
    }
    umma::fence_before();
    if (CTAS == 2) {
        mbar_wait(wfull, 0);             // the leader's MMAs read the peer's weights too: both halves before the pair starts
        cluster_sync_all();
    } else {
        __syncthreads();
    }
    umma::fence_after();
    CV_PHASE(3);
    const uint32_t tmem_base = *tmem_slot;
    const int tiles_per_img = a.tiles_x * a.tiles_y;
    // work unit = CTAS horizontally consecutive tiles; a unit past the end gives this CTA an all-padding tile
    const int n_units = (a.n_tiles + CTAS - 1) / CTAS, unit0 = blockIdx.x / CTAS, unit_step = gridDim.x / CTAS;

    if (warp == 0) {
        // ---- producer (whole warp converged, one elected lane issues): resident weights once, then the window ring ----
        uint32_t s = 0, ph = 0, i = 0;
        for (int unit = unit0; unit < n_units; unit += unit_step, ++i) {
            const int tile = unit * CTAS + (int)rank;
            const int b = tile / tiles_per_img, r = tile - b * tiles_per_img;
            const int ty = r / a.tiles_x, tx = r - ty * a.tiles_x;
            if (RSTG && !OSTG) {
                // residual tile of this output tile into the buffer the epilogue released two tiles ago (rows / columns
                // / images outside the tensor read as 0 and still count their bytes)
                mbar_wait(&rempty[i & 1], ((i >> 1) & 1) ^ 1);
                if (elect_one()) {
                    mbar_arrive_expect_tx(&rfull[i & 1], (uint32_t)rfloats * 4);
                    tma_load_4d(rbuf + (i & 1) * rfloats, &rmap, &rfull[i & 1], 0, ty * kCvTH, tx * kCvTW, b % a.res_B);
                }
                __syncwarp();
            }
            for (int st = 0; st < a.nstage_tile; ++st) {
                mbar_wait(&empty[s], ph ^ 1);
                if (st == 0 && lane == 0) CV_TRACE(0, i);
                if (st == a.nstage_tile - 1 && lane == 0) CV_TRACE(1, i);
                if (elect_one()) {
                    // box {16 ch, 16 y, 10 x, 1 image} at (16 st, y0 - 1, x0 - 1, b); channels >= Cin and images >= B read as 0
                    // every CTA fills its own ring and signals its own barrier; in a pair the peer's warp 1 forwards
                    // "stage full" to the leader's MMA warp
                    if (a.dbg & 1) {
                        mbar_arrive(&full[s]);
                    } else {
                        mbar_arrive_expect_tx(&full[s], kCvStageBytes);
                        tma_load_4d(a_ring + s * kCvStageBytes, &xmap, &full[s], 16 * st, ty * kCvTH - 1, tx * kCvTW - 1, b);
                    }
                }
                __syncwarp();
                if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1 && leader) {
        // ---- MMA issuer (whole warp converged; descriptors are warp-uniform, one elected lane issues) ----
        constexpr uint32_t idesc = umma::idesc_tf32(kCvM * CTAS, N);
        // A: [x][y][64 B] rows in the 64-byte swizzle (8-row groups 512 B apart); B: [k-half][n][16 B], no swizzle
        const uint64_t a_desc0 = umma::desc_kmajor_sw64(smem_u32(a_ring), 512);
        const uint64_t b_desc0 = umma::desc_kmajor(smem_u32(b_s), NH * 16, 128);
        mbar_wait(wfull, 0);
        uint32_t s = 0, ph = 0, i = 0;
        for (int unit = unit0; unit < n_units; unit += unit_step, ++i) {
            const uint32_t acc = i & 1;
            if (lane == 0) CV_TRACE(2, i);
            mbar_wait(&tempty[acc], ((i >> 1) & 1) ^ 1);
            if (lane == 0) CV_TRACE(3, i);
            umma::fence_after();
            const uint32_t d = tmem_base + acc * kCvAccStride;
            for (int st = 0; st < a.nstage_tile; ++st) {
                mbar_wait(&full[s], ph);
                if (CTAS == 2) mbar_wait(&pfull[s], ph);
                umma::fence_after();
                if (elect_one()) {
                    const uint64_t ad0 = a_desc0 + (uint64_t)((s * kCvStageBytes) >> 4);
                    const uint64_t bd0 = b_desc0 + (uint64_t)((st * kCvKSteps * slab) >> 4);
                    if (!(a.dbg & 4)) {
#pragma unroll
                        for (int j = 0; j < kCvKSteps; ++j) {
                            if (st * kCvKSteps + j < a.nchunk) {      // the last stage of a tile may hold one k-step only
#pragma unroll
                                for (int dx = 0; dx < 3; ++dx) {
                                    const uint64_t ad = ad0 + (uint64_t)((j * 32 + dx * kCvRows * 64) >> 4);
                                    const uint64_t bd = bd0 + (uint64_t)((dx * per_dx + j * slab) >> 4);
                                    const uint32_t accum = (j == 0 && dx == 0) ? (uint32_t)(st != 0) : 1u;
                                    if (CTAS == 2) umma::mma_tf32_2sm(d, ad, bd, idesc, accum);
                                    else umma::mma_tf32(d, ad, bd, idesc, accum);
                                }
                            }
                        }
                    }
                    if (CTAS == 2) {
                        umma::commit_2sm(&empty[s], 0b11);
                        if (st == a.nstage_tile - 1) umma::commit_2sm(&tfull[acc], 0b11);
                    } else {
                        umma::commit(&empty[s]);
                        if (st == a.nstage_tile - 1) umma::commit(&tfull[acc]);
                    }
                }
                __syncwarp();
                if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1; }
            }
            if (lane == 0) CV_TRACE(4, i);
        }
    } else if (warp == 1) {
        // ---- peer CTA of a pair: forward "my stage s is full" to the leader's MMA warp ----
        uint32_t s = 0, ph = 0;
        for (int unit = unit0; unit < n_units; unit += unit_step) {
            for (int st = 0; st < a.nstage_tile; ++st) {
                mbar_wait(&full[s], ph);
                if (lane == 0) mbar_arrive_cluster(mapa_shared(smem_u32(&pfull[s]), 0));
                __syncwarp();
                if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 2 + kCvEpiWarps) {
        // ---- store warp (OSTG): one TMA store per finished output tile, then hand the buffer back ----
        if (OSTG) {
            // With a residual the same warp also fetches the residual tiles (two tiles ahead, into the buffer a store has
            // just left): the window producer never waits for an epilogue.
            auto fetch_residual = [&](const int unit, const uint32_t slot) {
                const int tile = unit * CTAS + (int)rank;
                const int b = tile / tiles_per_img, r = tile - b * tiles_per_img;
                const int ty = r / a.tiles_x, tx = r - ty * a.tiles_x;
                mbar_arrive_expect_tx(&rfull[slot], (uint32_t)rfloats * 4);     // (outside the tensor: zeros, bytes still count)
                tma_load_4d(rbuf + slot * rfloats, &rmap, &rfull[slot], 0, ty * kCvTH, tx * kCvTW, b % a.res_B);
            };
            if (RSTG && lane == 0) {
                if (unit0 < n_units) fetch_residual(unit0, 0);
                if (unit0 + unit_step < n_units) fetch_residual(unit0 + unit_step, 1);
            }
            uint32_t i = 0;
            for (int unit = unit0; unit < n_units; unit += unit_step, ++i) {
                const int tile = unit * CTAS + (int)rank;
                const int b = tile / tiles_per_img, r = tile - b * tiles_per_img;
                const int ty = r / a.tiles_x, tx = r - ty * a.tiles_x;
                mbar_wait(&ofull[i & 1], (i >> 1) & 1);
                if (lane == 0) {
                    if (tile < a.n_tiles && !(a.dbg & 2)) {
                        // rows / columns past the image edge are dropped by the TMA unit
                        tma_store_4d(&omap, rbuf + (i & 1) * rfloats, 0, ty * kCvTH, tx * kCvTW, b);
                        bulk_commit_group();
                        bulk_wait_group_read<0>();                       // the tile has left shared memory
                    }
                    CV_TRACE(9, i);
                    if (RSTG) {
                        if (unit + 2 * unit_step < n_units) fetch_residual(unit + 2 * unit_step, i & 1);
                    } else {
                        mbar_arrive(&oempty[i & 1]);                     // -> the epilogue warps may write the next tile
                    }
                }
                __syncwarp();
            }
            if (lane == 0) bulk_wait_group<0>();                         // global writes complete before the grid ends
        }
    } else if (warp >= 2) {
        // ---- epilogue warps: TMEM lane group = warp % 4; lane = pixel (column-major: m = 16 * column + row) ----
        const int q = warp & 3, sub = (warp - 2) >> 2;
        const int m = q * 32 + lane, lx = m >> 4, ly = m & 15;
        const uint32_t lead_tempty0 = CTAS == 2 ? mapa_shared(smem_u32(&tempty[0]), 0) : 0u;
        uint32_t i = 0;
        for (int unit = unit0; unit < n_units; unit += unit_step, ++i) {
            const int tile = unit * CTAS + (int)rank;
            const int b = tile / tiles_per_img, r = tile - b * tiles_per_img;
            const int ty = r / a.tiles_x, tx = r - ty * a.tiles_x;
            const uint32_t acc = i & 1;
            // my pixel's residual in the shared-memory tile [x][y][cn] (rows 1..14 of the 16-row window are outputs)
            float* rslot = rbuf + (i & 1) * rfloats + (lx * kCvTH + ly - 1) * rpitch;
            const bool rowok = ly >= 1 && ly <= kCvTH;       // rows 0 and 15 of the window are halo
            if (RSTG) mbar_wait(&rfull[i & 1], (i >> 1) & 1);
            else if (OSTG) mbar_wait(&oempty[i & 1], ((i >> 1) & 1) ^ 1);
            if (warp == 2 && lane == 0) CV_TRACE(5, i);
            mbar_wait(&tfull[acc], (i >> 1) & 1);
            if (warp == 2 && lane == 0) CV_TRACE(6, i);
            __syncwarp();                  // tcgen05.ld and the shuffles below are warp-collective
            umma::fence_after();
            const uint32_t taddr = tmem_base + acc * kCvAccStride + ((uint32_t)(q * 32) << 16);
            const int y = ty * kCvTH + ly - 1, x = tx * kCvTW + lx;
            const bool valid = ly >= 1 && ly <= kCvTH && y < a.H && x < a.W && tile < a.n_tiles && !(a.dbg & 2);
            float* o = a.out + b * a.out_sb + y * a.out_sy + x * a.out_sx;
            const float* rp = a.res ? a.res + (b % a.res_B) * a.res_sb + y * a.res_sy + x * a.res_sx : nullptr;
            // kCvEpiWarps / 4 warps share each TMEM lane quadrant and take every (kCvEpiWarps / 4)-th 8-channel chunk
            auto emit = [&](const float (&p0)[8], const float (&p1)[8], const float (&p2)[8], const int c0) {
                float t[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float up = __shfl_up_sync(0xffffffffu, p0[k], 1);      // plane dy=-1 of the row above
                    const float dn = __shfl_down_sync(0xffffffffu, p2[k], 1);    // plane dy=+1 of the row below
                    t[k] = __fadd_rn(__fadd_rn(up, p1[k]), dn);
                }
                if (OSTG) {
                    // through the shared-memory tile (the residual, when there is one, sits in the very slot)
                    if (rowok) {
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            if (c0 + 4 * h < a.cn) {
                                float4* sp = reinterpret_cast<float4*>(rslot + c0) + h;
                                float4 rr = make_float4(0.f, 0.f, 0.f, 0.f);
                                if (RSTG) rr = *sp;
                                const float rv[4] = {rr.x, rr.y, rr.z, rr.w};
                                // bias / slopes as two broadcast 128-bit loads (shared-memory wavefronts are what this
                                // kernel runs out of: the tensor core's operand reads share them)
                                const float4 b4 = *reinterpret_cast<const float4*>(bias_s + c0 + 4 * h);
                                const float4 s4 = *reinterpret_cast<const float4*>(slope_s + c0 + 4 * h);
                                const float bv[4] = {b4.x, b4.y, b4.z, b4.w}, sv[4] = {s4.x, s4.y, s4.z, s4.w};
                                float u[4];
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    float w = __fadd_rn(t[4 * h + k], bv[k]);
                                    if (RSTG && a.res_mode == 1) w = __fadd_rn(rv[k], w);
                                    if (a.slope) w = w > 0.0f ? w : __fmul_rn(sv[k], w);
                                    if (RSTG && a.res_mode == 2) w = __fadd_rn(rv[k], w);
                                    u[k] = w;
                                }
                                *sp = make_float4(u[0], u[1], u[2], u[3]);
                            }
                        }
                    }
                } else if (valid && a.vec) {
                    // NHWC: lane = pixel, 8 consecutive channels = one 32-byte sector per thread
                    float rv[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                    const bool full8 = c0 + 8 <= a.cn;
                    if (RSTG) {
#pragma unroll
                        for (int h = 0; h < 2; ++h)
                            if (c0 + 4 * h < a.cn) {
                                const float4 rr = reinterpret_cast<const float4*>(rslot + c0)[h];
                                rv[4 * h] = rr.x; rv[4 * h + 1] = rr.y; rv[4 * h + 2] = rr.z; rv[4 * h + 3] = rr.w;
                            }
                    } else if (a.res_mode) {
                        if (a.vec == 2 && full8) {
                            ld_global_v8(rp + c0, rv);
                        } else {
#pragma unroll
                            for (int h = 0; h < 2; ++h)
                                if (c0 + 4 * h < a.cn) {
                                    const float4 rr = __ldg(reinterpret_cast<const float4*>(rp + c0) + h);
                                    rv[4 * h] = rr.x; rv[4 * h + 1] = rr.y; rv[4 * h + 2] = rr.z; rv[4 * h + 3] = rr.w;
                                }
                        }
                    }
                    float u[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const int c = c0 + k;
                        float w = __fadd_rn(t[k], bias_s[c]);
                        if (a.res_mode == 1) w = __fadd_rn(rv[k], w);
                        if (a.slope) w = w > 0.0f ? w : __fmul_rn(slope_s[c], w);
                        if (a.res_mode == 2) w = __fadd_rn(rv[k], w);
                        u[k] = w;
                    }
                    if (a.vec == 2 && full8) {
                        st_global_v8(o + c0, u);
                    } else {
#pragma unroll
                        for (int h = 0; h < 2; ++h)
                            if (c0 + 4 * h < a.cn) reinterpret_cast<float4*>(o + c0)[h] = make_float4(u[4 * h], u[4 * h + 1], u[4 * h + 2], u[4 * h + 3]);
                    }
                } else if (valid) {
                    // any other layout (planar NCHW, unaligned channel counts): one scalar store per channel
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const int c = c0 + k;
                        if (c < a.cn) {
                            float u = __fadd_rn(t[k], bias_s[c]);
                            const float rv = a.res_mode ? __ldg(rp + c * a.res_sc) : 0.0f;
                            if (a.res_mode == 1) u = __fadd_rn(rv, u);
                            if (a.slope) u = u > 0.0f ? u : __fmul_rn(slope_s[c], u);
                            if (a.res_mode == 2) u = __fadd_rn(rv, u);
                            o[c * a.out_sc] = u;
                        }
                    }
                }
            };
            auto release_acc = [&]() {
                __syncwarp();
                umma::fence_before();
                if (lane == 0) {
                    if (CTAS == 2) mbar_arrive_cluster(lead_tempty0 + acc * 8);
                    else mbar_arrive(&tempty[acc]);
                }
            };
            // all of this warp's chunks are loaded first (<= 3 x 24 registers), the accumulator goes back to the MMA warp,
            // and only then the chunks are shuffled / written: the MMAs of tile i + 2 wait for TMEM reads, not for stores
            constexpr int cstep = 2 * kCvEpiWarps;
            constexpr int kMaxC = (NPAD / 8 + kCvEpiWarps / 4 - 1) / (kCvEpiWarps / 4);
            const int cbase = (a.dbg & 8) ? a.cn : 8 * sub;    // dbg 8: handshakes only, no epilogue work
            float pv[kMaxC][3][8];
#pragma unroll
            for (int j = 0; j < kMaxC; ++j) {
                const int c = cbase + j * cstep;
                if (c < a.cn) {
                    umma::tmem_ld8(taddr + c, pv[j][0]);
                    umma::tmem_ld8(taddr + NPAD + c, pv[j][1]);
                    umma::tmem_ld8(taddr + 2 * NPAD + c, pv[j][2]);
                }
            }
            umma::tmem_ld_wait();
            release_acc();
            if (warp == 2 && lane == 0) CV_TRACE(7, i);
#pragma unroll
            for (int j = 0; j < kMaxC; ++j) {
                const int c = cbase + j * cstep;
                if (c < a.cn) {
                    emit(pv[j][0], pv[j][1], pv[j][2], c);
                    __syncwarp();
                }
            }
            __syncwarp();
            if (warp == 2 && lane == 0) CV_TRACE(10, i);
            if (OSTG) {
                fence_proxy_async_smem();                            // my writes to the tile -> visible to the TMA store
                __syncwarp();
                if (lane == 0) mbar_arrive(&ofull[i & 1]);
                if (warp == 2 && lane == 0) CV_TRACE(8, i);
            } else if (RSTG && lane == 0) {
                mbar_arrive(&rempty[i & 1]);                         // (after the __syncwarp above: all lanes have read it)
            }
        }
    }
    CV_PHASE(4);
    umma::fence_before();
    if (CTAS == 2) cluster_sync_all();
    else __syncthreads();
    CV_PHASE(5);
#ifdef VFI_CONV_TRACE
    if (blockIdx.x == 0 && tid == 0) {
        printf("phases (clk / ns from entry): barriers+tmem %lld / %lld, pdl wait %lld / %lld, weights+sync %lld / %lld, thread0 loop end %lld / %lld, all roles done %lld / %lld; first tile go at %lld clk, tiles of this CTA: last traced index\n",
               g_ph[1][0] - g_ph[0][0], g_ph[1][1] - g_ph[0][1], g_ph[2][0] - g_ph[0][0], g_ph[2][1] - g_ph[0][1],
               g_ph[3][0] - g_ph[0][0], g_ph[3][1] - g_ph[0][1], g_ph[4][0] - g_ph[0][0], g_ph[4][1] - g_ph[0][1],
               g_ph[5][0] - g_ph[0][0], g_ph[5][1] - g_ph[0][1], g_tr[3][0] - g_ph[0][0]);
        for (int i = 0; i < 34; i += 3)
            printf("  tile %2d: M go %7lld  E done %7lld  S left %7lld (clk from entry)\n", i, g_tr[3][i] - g_ph[0][0], g_tr[8][i] - g_ph[0][0], g_tr[9][i] - g_ph[0][0]);
        const long long t0 = g_tr[3][8];
        for (int i = 8; i < 16; ++i)
            printf("tile %2d  P first %6lld last %6lld | M enter %6lld go %6lld issued %6lld | E enter %6lld tfull %6lld loaded %6lld emitted %6lld done %6lld | S left %6lld\n",
                   i, g_tr[0][i] - t0, g_tr[1][i] - t0, g_tr[2][i] - t0, g_tr[3][i] - t0, g_tr[4][i] - t0, g_tr[5][i] - t0,
                   g_tr[6][i] - t0, g_tr[7][i] - t0, g_tr[10][i] - t0, g_tr[8][i] - t0, g_tr[9][i] - t0);
    }
#endif
    if (warp == 1) {
        if (CTAS == 2) umma::tmem_dealloc_2sm(tmem_base, 2 * kCvAccStride);
        else umma::tmem_dealloc(tmem_base, 2 * kCvAccStride);
    }
}

// ---- host side -------------------------------------------------------------------------------------------------
static int conv_ctas() {      // CTAs per MMA: 2 = CTA pairs (tcgen05 cta_group::2); VFI_CONV_CTAS=1 keeps single-CTA MMAs
    static int c = 0;
    if (!c) {
        const char* e = getenv("VFI_CONV_CTAS");
        c = (e && atoi(e) == 1) ? 1 : 2;
    }
    return c;
}
struct ConvTiling {
    int T, NPAD, n_nt, nstage_tile, nchunkP, ctas, ntot;
    size_t wbytes;      // resident weights of one Cout tile IN ONE CTA
};
static ConvTiling conv_tiling(int Cout, int Cin) {
    ConvTiling t;
    t.ctas = conv_ctas();
    if (Cout <= 72) t.T = Cout;                         // one tile
    else if (Cout % 72 == 0) t.T = 72;
    else t.T = 64;
    const int sizes[5] = {16, 32, 48, 64, 72};          // instantiated Cout tiles (UMMA N = 3 * NPAD <= 216)
    t.NPAD = 72;
    for (int i = 4; i >= 0; --i)
        if (sizes[i] >= t.T) t.NPAD = sizes[i];
    t.n_nt = (Cout + t.T - 1) / t.T;
    const int nchunk = (Cin + 7) / 8;                   // a trailing half chunk is zero-filled by TMA (and by prep)
    t.nstage_tile = (nchunk + kCvKSteps - 1) / kCvKSteps;
    t.nchunkP = nchunk;
    t.ntot = t.ctas == 1 ? 3 * t.NPAD : (3 * t.NPAD + 15) / 16 * 16;
    t.wbytes = (size_t)3 * t.nchunkP * 2 * (t.ntot / t.ctas) * 16;
    return t;
}
static size_t conv_smem_fixed(int NPAD) { return 2048 + 2 * NPAD * 4 + (3 * kCvMaxStages + 13) * 8 + 16 + 128; }
// Pixel stride (floats) of the shared-memory residual / output tiles: the smallest P >= cn with P = 4 (mod 8), i.e. an odd
// number of 16-byte words, so that the 8 lanes of a quarter warp (8 consecutive pixels, 16 bytes each) hit 8 different
// bank groups (cn = 72 was a 2-way, cn = 64 an 8-way conflict).  The TMA boxes are P channels wide: the P - cn channels
// past the tensor's extent are zero-filled on loads and dropped on stores.
__host__ __device__ static inline int conv_tile_pitch(int cn) { return ((cn + 3) / 8) * 8 + 4; }
static size_t conv_rbuf_bytes(int cn) { return (size_t)2 * kCvTW * kCvTH * conv_tile_pitch(cn) * 4; }   // two tiles

template <int NPAD, int CTAS, bool RSTG, bool OSTG>
static int conv_launch(const ConvArgs& a, const CUtensorMap& xmap, const CUtensorMap& rmap, const CUtensorMap& omap, size_t smem,
                       cudaStream_t st) {
    static bool attr = false;
    if (!attr) {
        VFI_CUDA(cudaFuncSetAttribute(conv3x3_tc_kernel<NPAD, CTAS, RSTG, OSTG>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCvSmemMax));
        attr = true;
    }
    const int n_units = (a.n_tiles + CTAS - 1) / CTAS;
    const int grid = CTAS * (n_units < kNumSMs / CTAS ? n_units : kNumSMs / CTAS);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kCvThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CTAS;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    // programmatic dependent launch: this grid may become resident while the previous kernel of the stream drains its
    // last tiles; it loads its resident weights (constants) and then waits (griddepcontrol.wait) before touching any
    // activation.  The previous kernel releases it early only if it is one of ours (griddepcontrol.launch_dependents).
    static const bool pdl = [] { const char* e = getenv("VFI_CONV_PDL"); return !(e && e[0] == '0'); }();
    // the launch right after a weight preparation reads what that (PDL-unaware) kernel wrote before the wait: plain order
    const bool after_prep = g_prep_pending;
    g_prep_pending = false;
    if (pdl && !after_prep) {
        at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[1].val.programmaticStreamSerializationAllowed = 1;
        cfg.numAttrs = 2;
    }
    const cudaError_t e = cudaLaunchKernelEx(&cfg, conv3x3_tc_kernel<NPAD, CTAS, RSTG, OSTG>, a, xmap, rmap, omap);
    if (e != cudaSuccess) {
        set_error("conv3x3_tc_kernel: launch failed: %s", cudaGetErrorString(e));
        return VFI_ERR_CUDA;
    }
    VFI_CHECK_LAUNCH("conv3x3_tc_kernel");
    return VFI_OK;
}
template <int NPAD>
static int conv_launch_any(const ConvArgs& a, const CUtensorMap& xmap, const CUtensorMap& rmap, const CUtensorMap& omap, size_t smem,
                           int ctas, cudaStream_t st) {
    if (ctas == 2) {
        if (a.ostg) return a.staged ? conv_launch<NPAD, 2, true, true>(a, xmap, rmap, omap, smem, st)
                                    : conv_launch<NPAD, 2, false, true>(a, xmap, rmap, omap, smem, st);
        return a.staged ? conv_launch<NPAD, 2, true, false>(a, xmap, rmap, omap, smem, st)
                        : conv_launch<NPAD, 2, false, false>(a, xmap, rmap, omap, smem, st);
    }
    return conv_launch<NPAD, 1, false, false>(a, xmap, rmap, omap, smem, st);
}

}  // namespace vfi

static bool conv_upscale_ok(int Cout, int upscale) { return upscale == 1 || (upscale == 2 && Cout % 4 == 0); }

extern "C" size_t vfi_conv3x3_prep_bytes(int Cout, int Cin, int upscale) {
    using namespace vfi;
    if (Cout < 1 || Cin < 4 || Cin % 4 || !conv_upscale_ok(Cout, upscale)) return 0;
    const int up2 = upscale * upscale;
    const ConvTiling t = conv_tiling(Cout / up2, Cin);
    return (size_t)up2 * t.n_nt * t.ctas * t.wbytes;
}

extern "C" int vfi_conv3x3_prep_f32(const float* weight, float* wprep, int Cout, int Cin, int upscale, void* stream) {
    using namespace vfi;
    VFI_REQUIRE(weight && wprep, "vfi_conv3x3_prep_f32: null pointer");
    VFI_REQUIRE(Cout >= 1 && Cin >= 4 && Cin % 4 == 0, "vfi_conv3x3_prep_f32: Cin %d must be a positive multiple of 4", Cin);
    VFI_REQUIRE(conv_upscale_ok(Cout, upscale), "vfi_conv3x3_prep_f32: upscale %d with Cout %d", upscale, Cout);
    const int up2 = upscale * upscale;
    const ConvTiling t = conv_tiling(Cout / up2, Cin);
    const int64_t total = (int64_t)up2 * t.n_nt * t.ctas * (int64_t)(t.wbytes / 4);
    const int blocks = (int)min64(ceil_div(total, 256), (int64_t)kNumSMs * 8);
    conv3x3_prep_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(weight, wprep, Cout, Cin, t.T, t.NPAD, up2 * t.n_nt,
                                                                             t.nchunkP, up2, t.ntot, t.ctas);
    VFI_CHECK_LAUNCH("conv3x3_prep_kernel");
    vfi::g_prep_pending = true;
    return VFI_OK;
}

extern "C" int vfi_conv3x3_supported(int Cin, int x_pix_stride, int Cout, int upscale) {
    using namespace vfi;
    if (Cin < 4 || Cin % 4 || x_pix_stride % 4 || x_pix_stride < Cin || Cout < 1 || !conv_upscale_ok(Cout, upscale)) return 0;
    const ConvTiling t = conv_tiling(Cout / (upscale * upscale), Cin);
    return t.wbytes + conv_smem_fixed(t.NPAD) + 2 * (size_t)kCvStageBytes <= (size_t)kCvSmemMax;
}

extern "C" int vfi_conv3x3_f32(const float* x, const float* wprep, const float* bias, const float* slope, const float* residual,
                               float* out, int B, int H, int W, int Cin, int x_pix_stride, int Cout, int out_channels_last,
                               int out_c_total, int res_c_total, int residual_mode, int upscale, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(x && wprep && out, "vfi_conv3x3_f32: null pointer");
    VFI_REQUIRE(B >= 0 && H >= 1 && W >= 1, "vfi_conv3x3_f32: bad shape");
    VFI_REQUIRE(vfi_conv3x3_supported(Cin, x_pix_stride, Cout, upscale),
                "vfi_conv3x3_f32: unsupported Cin %d / pixel stride %d / Cout %d / upscale %d", Cin, x_pix_stride, Cout, upscale);
    VFI_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0, "vfi_conv3x3_f32: x must be 16-byte aligned");
    // bits 8.. of residual_mode: number of images in the residual when it is broadcast over the batch (0 = B)
    const int res_images = residual_mode >> 8;
    residual_mode &= 0xFF;
    VFI_REQUIRE(residual_mode >= 0 && residual_mode <= 2 && (residual_mode == 0 || residual), "vfi_conv3x3_f32: residual_mode %d",
                residual_mode);
    VFI_REQUIRE(res_images >= 0 && (res_images == 0 || (residual_mode != 0 && B % res_images == 0)),
                "vfi_conv3x3_f32: %d residual images do not divide the batch %d", res_images, B);
    VFI_REQUIRE(upscale == 1 || residual_mode == 0, "vfi_conv3x3_f32: a residual cannot be combined with the fused PixelShuffle");
    const int up = upscale, up2 = up * up, Cs = Cout / up2;      // Cs = channels after the shuffle
    VFI_REQUIRE(out_c_total >= Cs && (residual_mode == 0 || res_c_total >= Cout), "vfi_conv3x3_f32: channel totals");
    if (B == 0) return VFI_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const ConvTiling t = conv_tiling(Cs, Cin);
    CUtensorMap xmap;
    {
        // (channel, y, x, image): y before x so that a box lands column-major in shared memory; 64-byte swizzle
        const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)H, (uint64_t)W, (uint64_t)B};
        const uint64_t strides[3] = {(uint64_t)W * x_pix_stride * 4, (uint64_t)x_pix_stride * 4, (uint64_t)H * W * x_pix_stride * 4};
        const uint32_t box[4] = {16, (uint32_t)kCvRows, (uint32_t)kCvWinW, 1};
        const int rc = make_tmap_f32_nd(&xmap, x, 4, dims, strides, box, 64);
        if (rc != VFI_OK) return rc;
    }
    ConvArgs a = {};
    {
        const char* e = getenv("VFI_CONV_DBG");
        a.dbg = e ? atoi(e) : 0;
    }
    a.B = B; a.H = H; a.W = W;
    a.nstage_tile = t.nstage_tile;
    a.nchunk = t.nchunkP;
    a.res_mode = residual_mode;
    a.res_B = res_images ? res_images : B;
    a.bias_stride = up2;
    a.tiles_x = (W + kCvTW - 1) / kCvTW; a.tiles_y = (H + kCvTH - 1) / kCvTH;
    a.n_tiles = B * a.tiles_x * a.tiles_y;
    // element strides of the output seen from a conv pixel (y, x): with the fused PixelShuffle conv pixel (y, x) owns the
    // up x up output pixels (up*y + i, up*x + j) of a (B, Cs, up*H, up*W) tensor
    const int64_t Ho = (int64_t)H * up, Wo = (int64_t)W * up, HWo = Ho * Wo, HW = (int64_t)H * W;
    int64_t o_pix, o_row, o_c;      // strides of one OUTPUT pixel step, row step, channel step
    if (out_channels_last) {
        o_pix = out_c_total; o_row = Wo * out_c_total; o_c = 1;
        a.res_sb = HW * res_c_total; a.res_sy = (int64_t)W * res_c_total; a.res_sx = res_c_total; a.res_sc = 1;
    } else {
        o_pix = 1; o_row = Wo; o_c = HWo;
        a.res_sb = HW * res_c_total; a.res_sy = W; a.res_sx = 1; a.res_sc = HW;
    }
    a.out_sb = HWo * out_c_total; a.out_sy = o_row * up; a.out_sx = o_pix * up; a.out_sc = o_c;
    // VFI_CONV_RSTG=0: read the residual straight from global memory in the epilogue (A/B switch for the TMA-staged path)
    static const bool no_rstg = getenv("VFI_CONV_RSTG") && atoi(getenv("VFI_CONV_RSTG")) == 0;
    for (int sub = 0; sub < up2; ++sub) {
        for (int nt = 0; nt < t.n_nt; ++nt) {
            const int c0 = nt * t.T;                           // first channel (after the shuffle) of this tile
            a.cn = Cs - c0 < t.T ? Cs - c0 : t.T;
            a.wprep = wprep + (size_t)(sub * t.n_nt + nt) * t.ctas * (t.wbytes / 4);
            a.bias = bias ? bias + (size_t)c0 * up2 + sub : nullptr;
            a.slope = slope ? slope + c0 : nullptr;
            a.out = out + c0 * o_c + (sub / up) * o_row + (sub % up) * o_pix;
            a.res = residual ? residual + c0 * a.res_sc : nullptr;
            const bool al = ((reinterpret_cast<uintptr_t>(a.out) & 15) == 0) && (a.res == nullptr || (reinterpret_cast<uintptr_t>(a.res) & 15) == 0);
            a.vec = out_channels_last && al && a.cn % 4 == 0 && out_c_total % 4 == 0 && (residual_mode == 0 || res_c_total % 4 == 0);
            // vec 2: pixel rows of out / residual are 32-byte aligned too -> 256-bit accesses
            auto al32 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 31) == 0; };
            if (a.vec && al32(a.out) && out_c_total % 8 == 0 && (a.res == nullptr || (al32(a.res) && res_c_total % 8 == 0))) a.vec = 2;
            // residual through shared memory (CTA pairs leave room for it): two tile buffers filled by TMA, one tile ahead
            const size_t base = conv_smem_fixed(t.NPAD) + t.wbytes;
            a.staged = (t.ctas == 2 && a.vec && a.res && !no_rstg &&
                        base + conv_rbuf_bytes(a.cn) + 3 * (size_t)kCvStageBytes <= (size_t)kCvSmemMax) ? 1 : 0;
            // output through shared memory + TMA store (CTA pairs, NHWC, 16-byte aligned rows): the tile buffers are the
            // residual's when there is one (results written in place), else two buffers of the same size; a residual
            // that is not staged keeps the direct path.  VFI_CONV_OSTG=0: A/B switch
            static const bool no_ostg = getenv("VFI_CONV_OSTG") && atoi(getenv("VFI_CONV_OSTG")) == 0;
            a.ostg = (t.ctas == 2 && a.vec && !no_ostg && (a.staged || a.res == nullptr) &&
                      base + conv_rbuf_bytes(a.cn) + 3 * (size_t)kCvStageBytes <= (size_t)kCvSmemMax) ? 1 : 0;
            const size_t rb = (a.staged || a.ostg) ? conv_rbuf_bytes(a.cn) : 0;
            int stages = (int)(((size_t)kCvSmemMax - base - rb) / kCvStageBytes);
            if (stages > kCvMaxStages) stages = kCvMaxStages;
            if (const char* e = getenv("VFI_CONV_STAGES")) stages = atoi(e) >= 2 && atoi(e) < stages ? atoi(e) : stages;
            a.stages = stages;
            const size_t smem = base + rb + (size_t)stages * kCvStageBytes;
            CUtensorMap rmap = xmap, omap = xmap;      // placeholders when the direct paths run
            if (a.staged) {
                // (channel, y, x, image) view of this Cout tile of the residual; box = one output tile [8 x][14 y][cn]
                const uint64_t dims[4] = {(uint64_t)a.cn, (uint64_t)H, (uint64_t)W, (uint64_t)a.res_B};
                const uint32_t box[4] = {(uint32_t)conv_tile_pitch(a.cn), (uint32_t)kCvTH, (uint32_t)kCvTW, 1};
                const uint64_t sr[3] = {(uint64_t)a.res_sy * 4, (uint64_t)a.res_sx * 4, (uint64_t)a.res_sb * 4};
                const int rc2 = make_tmap_f32_nd(&rmap, a.res, 4, dims, sr, box, 0);
                if (rc2 != VFI_OK) return rc2;
            }
            if (a.ostg) {
                // the same view of the output, indexed by CONV pixel (with the fused PixelShuffle the strides step over
                // the other sub-pixels)
                const uint64_t dims[4] = {(uint64_t)a.cn, (uint64_t)H, (uint64_t)W, (uint64_t)B};
                const uint32_t box[4] = {(uint32_t)conv_tile_pitch(a.cn), (uint32_t)kCvTH, (uint32_t)kCvTW, 1};
                const uint64_t so[3] = {(uint64_t)a.out_sy * 4, (uint64_t)a.out_sx * 4, (uint64_t)a.out_sb * 4};
                const int rc3 = make_tmap_f32_nd(&omap, a.out, 4, dims, so, box, 0);
                if (rc3 != VFI_OK) return rc3;
            }
            int rc;
            switch (t.NPAD) {
                case 16: rc = conv_launch_any<16>(a, xmap, rmap, omap, smem, t.ctas, st); break;
                case 32: rc = conv_launch_any<32>(a, xmap, rmap, omap, smem, t.ctas, st); break;
                case 48: rc = conv_launch_any<48>(a, xmap, rmap, omap, smem, t.ctas, st); break;
                case 64: rc = conv_launch_any<64>(a, xmap, rmap, omap, smem, t.ctas, st); break;
                case 72: rc = conv_launch_any<72>(a, xmap, rmap, omap, smem, t.ctas, st); break;
                default:
                    set_error("vfi_conv3x3_f32: no kernel for a Cout tile padded to %d", t.NPAD);
                    return VFI_ERR_UNSUPPORTED;
            }
            if (rc != VFI_OK) return rc;
        }
    }
    return VFI_OK;
}
