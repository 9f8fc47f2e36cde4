// Backward bilinear warp with border clamp (reference: modules/warp.py:7-15).
//
// HBM-bound: 4*(2C+2) bytes per pixel.  Coordinates, corner indices and the four weights are computed
// once per pixel (flow read once) and reused for every channel; the channel loop keeps 4 channels x 4
// taps = 16 independent gathers in flight per thread.  Two layouts:
//   * planar (NCHW, stride_w == 1): a warp covers 32 consecutive x, so output stores are 128 B
//     coalesced and the taps of neighbouring lanes share L1 lines when the flow is locally smooth;
//   * channels-last (stride_c == 1): a thread owns 4 consecutive channels of one pixel, all taps and the
//     store are 128-bit.
#include <stdlib.h>

#include "tma.cuh"
#include "vfi_common.cuh"

namespace vfi {

struct Str4 {
    int64_t n, c, h, w;
};

struct BwarpGeom {
    int x0, x1, y0, y1;
    float w_nw, w_ne, w_sw, w_se;
};

// The tensor may be a crop (H x W at fe.x0, fe.y0) of a frame of fe.Wg x fe.Hg pixels (4K tiles): the
// normalisation and the border clamp are the FRAME's (lin_x / lin_y are slices of the frame's linspace vectors), so a
// crop samples exactly where the whole frame would; taps the crop does not hold clamp to the crop's edge (the
// caller's halo has to cover them).
__device__ __forceinline__ BwarpGeom bwarp_geom(float fx, float fy, float lin_x, float lin_y, int H, int W, const FrameExt& fe) {
    // reference op order: flow / ((W-1)/2) is a true fp32 division (torch CPU), then grid + flow_
    const float wm1 = (float)(fe.Wg - 1), hm1 = (float)(fe.Hg - 1);
    const float half_w = (float)((fe.Wg - 1.0) / 2.0), half_h = (float)((fe.Hg - 1.0) / 2.0);
    float ix = replay_coord(lin_x, __fdiv_rn(fx, half_w), wm1);
    float iy = replay_coord(lin_y, __fdiv_rn(fy, half_h), hm1);
    ix = fminf(fmaxf(ix, 0.0f), wm1);  // padding_mode='border'
    iy = fminf(fmaxf(iy, 0.0f), hm1);
    const float xf = floorf(ix), yf = floorf(iy);
    const float tx = ix - xf, ty = iy - yf;
    const float ex = 1.0f - tx, ey = 1.0f - ty;
    BwarpGeom g;
    const int x0 = (int)xf, y0 = (int)yf;
    const int x1 = min(x0 + 1, fe.Wg - 1);  // weight is exactly 0 whenever this clamps
    const int y1 = min(y0 + 1, fe.Hg - 1);
    g.x0 = min(max(x0 - fe.x0, 0), W - 1);   // crop coordinates (no-ops on a whole frame)
    g.x1 = min(max(x1 - fe.x0, 0), W - 1);
    g.y0 = min(max(y0 - fe.y0, 0), H - 1);
    g.y1 = min(max(y1 - fe.y0, 0), H - 1);
    g.w_nw = ey * ex;
    g.w_ne = ey * tx;
    g.w_sw = ty * ex;
    g.w_se = ty * tx;
    return g;
}

#ifndef VFI_BWARP_CUNROLL
#define VFI_BWARP_CUNROLL 4
#endif
constexpr int kCUnroll = VFI_BWARP_CUNROLL;

// planar / generic-stride kernel: thread = one pixel, loops over a chunk of channels
__global__ void __launch_bounds__(256) bwarp_planar_kernel(const float* __restrict__ img, const float* __restrict__ flow,
                                                           const float* __restrict__ lin_x,
                                                           const float* __restrict__ lin_y, float* __restrict__ out,
                                                           int C, int H, int W, int c_chunk, int n_chunks, Str4 si,
                                                           Str4 sf, Str4 so, FrameExt fe) {
    const int x = blockIdx.x * 32 + threadIdx.x;
    const int y = blockIdx.y * 8 + threadIdx.y;
    const int b = blockIdx.z / n_chunks;
    const int c_begin = (blockIdx.z % n_chunks) * c_chunk;
    // no early exit: the 8 rows of the CTA walk the planes in lockstep (barrier per channel group) so that row y's
    // south taps and row y+1's north taps -- the same sectors -- meet in the L1 (measured on the splat gather: 2x on
    // rough flows); out-of-frame threads compute on the clamped pixel and skip the store
    const bool live = x < W && y < H;
    const int xc = min(x, W - 1), yc = min(y, H - 1);
    const int c_end = min(C, c_begin + c_chunk);

    const float* fp = flow + b * sf.n + yc * sf.h + xc * sf.w;
    const float fx = ld_stream(fp), fy = ld_stream(fp + sf.c);
    const BwarpGeom g = bwarp_geom(fx, fy, lin_x[xc], lin_y[yc], H, W, fe);

    const int64_t o_nw = g.y0 * si.h + g.x0 * si.w, o_ne = g.y0 * si.h + g.x1 * si.w;
    const int64_t o_sw = g.y1 * si.h + g.x0 * si.w, o_se = g.y1 * si.h + g.x1 * si.w;
    const float* ip = img + b * si.n + c_begin * si.c;
    float* op = out + b * so.n + c_begin * so.c + yc * so.h + xc * so.w;

    int c = c_begin;
    for (; c + kCUnroll <= c_end; c += kCUnroll) {
        __syncthreads();
        float v[kCUnroll][4];
#pragma unroll
        for (int u = 0; u < kCUnroll; ++u) {
            const float* p = ip + u * si.c;
            v[u][0] = ld_gather(p + o_nw);
            v[u][1] = ld_gather(p + o_ne);
            v[u][2] = ld_gather(p + o_sw);
            v[u][3] = ld_gather(p + o_se);
        }
#pragma unroll
        for (int u = 0; u < kCUnroll; ++u) {
            const float r = v[u][0] * g.w_nw + v[u][1] * g.w_ne + v[u][2] * g.w_sw + v[u][3] * g.w_se;
            if (live) st_stream(op + u * so.c, r);
        }
        ip += kCUnroll * si.c;
        op += kCUnroll * so.c;
    }
    for (; c < c_end; ++c) {
        const float r = ld_gather(ip + o_nw) * g.w_nw + ld_gather(ip + o_ne) * g.w_ne + ld_gather(ip + o_sw) * g.w_sw +
                        ld_gather(ip + o_se) * g.w_se;
        if (live) st_stream(op, r);
        ip += si.c;
        op += so.c;
    }
}

// channels-last kernel: thread = (pixel, 4 consecutive channels); requires C % 4 == 0, stride_c == 1,
// 16-byte aligned bases and pixel strides.
__global__ void __launch_bounds__(256) bwarp_cl_kernel(const float* __restrict__ img, const float* __restrict__ flow,
                                                       const float* __restrict__ lin_x, const float* __restrict__ lin_y,
                                                       float* __restrict__ out, int B, int C, int H, int W, Str4 si,
                                                       Str4 sf, Str4 so, FrameExt fe, int flowcat) {
    const int cv = C >> 2;
    const int64_t total = (int64_t)B * H * W * cv;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int c4 = (int)(idx % cv) * 4;
        int64_t p = idx / cv;
        const int x = (int)(p % W);
        p /= W;
        const int y = (int)(p % H);
        const int b = (int)(p / H);
        const float* fp = flow + b * sf.n + y * sf.h + x * sf.w;
        const BwarpGeom g = bwarp_geom(__ldg(fp), __ldg(fp + sf.c), lin_x[x], lin_y[y], H, W, fe);
        const float* ip = img + b * si.n + c4;
        const float4 a = __ldg(reinterpret_cast<const float4*>(ip + g.y0 * si.h + g.x0 * si.w));
        const float4 bq = __ldg(reinterpret_cast<const float4*>(ip + g.y0 * si.h + g.x1 * si.w));
        const float4 c = __ldg(reinterpret_cast<const float4*>(ip + g.y1 * si.h + g.x0 * si.w));
        const float4 d = __ldg(reinterpret_cast<const float4*>(ip + g.y1 * si.h + g.x1 * si.w));
        float4 r;
        r.x = a.x * g.w_nw + bq.x * g.w_ne + c.x * g.w_sw + d.x * g.w_se;
        r.y = a.y * g.w_nw + bq.y * g.w_ne + c.y * g.w_sw + d.y * g.w_se;
        r.z = a.z * g.w_nw + bq.z * g.w_ne + c.z * g.w_sw + d.z * g.w_se;
        r.w = a.w * g.w_nw + bq.w * g.w_ne + c.w * g.w_sw + d.w * g.w_se;
        st_stream4(out + b * so.n + y * so.h + x * so.w + c4, r);
        // flowcat: the thread of the pixel's last channel quad also writes (flow_x, flow_y, 0, 0) behind the C warped channels
        if (flowcat && c4 == C - 4) st_stream4(out + b * so.n + y * so.h + x * so.w + C, make_float4(__ldg(fp), __ldg(fp + sf.c), 0.0f, 0.0f));
    }
}


// ---- tile kernel (contiguous NCHW): source window staged in shared memory by TMA -------------------------
// The direct-gather kernels above are bound by L1 wavefronts (4 unaligned 128 B tap requests per warp-output).
// Here a CTA owns a 64x16 output tile; the source window of the tile -- the tile displaced by the median flow of
// nine probe pixels, +-16 px in x and +-8 px in y -- is brought into shared memory one channel per pipeline
// stage with a 3-D TMA box load (out-of-frame parts are zero-filled and never read: taps are border-clamped),
// and the four taps become conflict-free LDS.  Pixels whose taps leave the window (flow outliers) fall back to
// global loads, so the result never depends on the window choice.  Geometry is computed once per pixel and
// reused for every channel.  Warps 0-7 consume, warp 8 is the TMA producer (full/empty mbarrier ring).
#ifndef VFI_BWARP_HVX
#define VFI_BWARP_HVX 16
#endif
#ifndef VFI_BWARP_HVY
#define VFI_BWARP_HVY 8
#endif
constexpr int kTW = 64, kTH = 16, kHVX = VFI_BWARP_HVX, kHVY = VFI_BWARP_HVY;
constexpr int kRW = kTW + 2 * kHVX, kRH = kTH + 2 * kHVY;   // 96 x 32 window
constexpr int kStageFloats = kRW * kRH;
#ifndef VFI_BWARP_STAGES
#define VFI_BWARP_STAGES 6
#endif
#ifndef VFI_BWARP_CTAS
#define VFI_BWARP_CTAS 2
#endif
constexpr int kStages = VFI_BWARP_STAGES;
constexpr int kTileThreads = 288;

struct TilePix {
    int o_s;        // window offset of the north-west tap
    int o_g;        // frame offset of the north-west tap
    int dxy;        // bit 0: x1 = x0 + 1, bit 1: y1 = y0 + 1, bit 2: taps inside the window, bit 3: pixel inside the frame
    float w_nw, w_ne, w_sw, w_se;
};

__device__ __forceinline__ float median9(float* v) {
    // tiny insertion sort (thread 0 only)
    for (int i = 1; i < 9; ++i) {
        const float k = v[i];
        int j = i - 1;
        while (j >= 0 && v[j] > k) {
            v[j + 1] = v[j];
            --j;
        }
        v[j + 1] = k;
    }
    return v[4];
}

__global__ void __launch_bounds__(kTileThreads, VFI_BWARP_CTAS) bwarp_tile_kernel(const __grid_constant__ CUtensorMap tmap,
                                                                  const float* __restrict__ img, const float* __restrict__ flow,
                                                                  const float* __restrict__ lin_x, const float* __restrict__ lin_y,
                                                                  float* __restrict__ out, int C, int H, int W, Str4 sf,
                                                                  int tiles_x, int tiles_y, int n_tiles_all, int c_chunk, FrameExt fe) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t* smem_al = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);   // TMA destinations need 128 B alignment
    float* stage = reinterpret_cast<float*>(smem_al);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_al + (size_t)kStages * kStageFloats * sizeof(float));
    uint64_t* empty = full + kStages;
    __shared__ float s_probe[18];
    __shared__ int s_shift[2];
    __shared__ int s_nout;
    __shared__ unsigned short s_outlist[kTW * kTH];   // tile-local ids of pixels whose taps leave the window

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tiles_per_img = tiles_x * tiles_y;
    // block order: all tiles of one channel chunk first, so that co-resident CTAs share windows in L2
    const int chunk = blockIdx.x / n_tiles_all;
    const int tile_id = blockIdx.x - chunk * n_tiles_all;
    const int b = tile_id / tiles_per_img;
    const int trem = tile_id - b * tiles_per_img;
    const int ty = trem / tiles_x, tx = trem - ty * tiles_x;
    const int x_t0 = tx * kTW, y_t0 = ty * kTH;
    const int c_begin = chunk * c_chunk, c_end = min(C, c_begin + c_chunk), n_c = c_end - c_begin;

    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 8);
        }
        fence_mbar_init();
    }
    // my 4 pixels' flow vectors: issued before the probe barrier so that the two global-memory latencies of the prologue
    // (probe flows -> window origin, pixel flows -> geometry) overlap instead of following each other
    float my_fx[4], my_fy[4];
    if (warp < 8) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int x = min(x_t0 + lane + 32 * (k & 1), W - 1), y = min(y_t0 + warp + 8 * (k >> 1), H - 1);
            const float* fp = flow + b * sf.n + y * sf.h + x * sf.w;
            my_fx[k] = ld_stream(fp);
            my_fy[k] = ld_stream(fp + sf.c);
        }
    }
    if (tid < 9) {   // 3x3 probe grid inside the tile
        const int px = min(x_t0 + (kTW / 6) * (1 + 2 * (tid % 3)), W - 1);
        const int py = min(y_t0 + (kTH / 6) * (1 + 2 * (tid / 3)), H - 1);
        const float* fp = flow + b * sf.n + py * sf.h + px * sf.w;
        const float fx = __ldg(fp), fy = __ldg(fp + sf.c);
        s_probe[tid] = isfinite(fx) ? fx : 0.0f;
        s_probe[9 + tid] = isfinite(fy) ? fy : 0.0f;
    }
    __syncthreads();
    if (tid == 0) {
        const float mx = median9(s_probe), my = median9(s_probe + 9);
        s_shift[0] = (int)rintf(fminf(fmaxf(mx, -1.0e6f), 1.0e6f));
        s_shift[1] = (int)rintf(fminf(fmaxf(my, -1.0e6f), 1.0e6f));
        s_nout = 0;
    }
    __syncthreads();
    // window origin, clamped so that it always overlaps the frame the taps are clamped to
    // (TMA: the innermost box coordinate must be a multiple of 16 bytes -- other values fault; `& ~3` floors)
    const int rx0 = min(max(x_t0 + s_shift[0] - kHVX, -kRW + 4), W - 1) & ~3;
    const int ry0 = min(max(y_t0 + s_shift[1] - kHVY, -kRH + 1), H - 1);
    const int64_t HW = (int64_t)H * W;

    if (warp == 8) {
        if (lane == 0) {
            tma_prefetch_desc(&tmap);
            for (int c = 0; c < n_c; ++c) {
                const int s = c % kStages;
                if (c >= kStages) mbar_wait(&empty[s], ((c / kStages) - 1) & 1);
                mbar_arrive_expect_tx(&full[s], kStageFloats * sizeof(float));
                tma_load_3d(stage + (size_t)s * kStageFloats, &tmap, &full[s], rx0, ry0, b * C + c_begin + c);
            }
        }
        return;
    }

    // ---- consumers: geometry of my 4 pixels (x = lane, lane+32; y = warp, warp+8) ----
    // Per pixel the plane loop needs ONE shared-memory address (the north-west tap; the other three are at the constant
    // offsets +4, +kRW*4, +kRW*4+4 bytes) and the four weights.  The east / south neighbours are always read, also
    // where the reference clamps them onto the border pixel itself (x1 = min(x0 + 1, W - 1)): their weight is exactly
    // 0 there and the window holds TMA's zero fill beyond the frame, so the product is an exact 0.
    uint32_t a_nw[4];
    float w_nw[4], w_ne[4], w_sw[4], w_se[4];
    bool ok[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int xr = x_t0 + lane + 32 * (k & 1), yr = y_t0 + warp + 8 * (k >> 1);
        const bool valid = xr < W && yr < H;
        const int x = min(xr, W - 1), y = min(yr, H - 1);
        const BwarpGeom g = bwarp_geom(my_fx[k], my_fy[k], __ldg(lin_x + x), __ldg(lin_y + y), H, W, fe);
        const int lx0 = g.x0 - rx0, ly0 = g.y0 - ry0;
        // a clamped neighbour (x1 == x0 or y1 == y0: frame border, or the edge of a crop) must carry weight 0 for the
        // fixed +1 offsets to be right; inside a crop that is not the case, those pixels take the tail pass
        const bool plus1 = (g.x1 == g.x0 + 1 || g.w_ne + g.w_se == 0.0f) && (g.y1 == g.y0 + 1 || g.w_sw + g.w_se == 0.0f);
        const bool inside = lx0 >= 0 && lx0 + 1 < kRW && ly0 >= 0 && ly0 + 1 < kRH && plus1;
        if (valid && !inside) s_outlist[atomicAdd(&s_nout, 1)] = (unsigned short)((yr - y_t0) * kTW + (xr - x_t0));
        a_nw[k] = (uint32_t)((ly0 * kRW + lx0) * 4);
        ok[k] = valid && inside;
        w_nw[k] = g.w_nw;
        w_ne[k] = g.w_ne;
        w_sw[k] = g.w_sw;
        w_se[k] = g.w_se;
    }
    // ---- pixels whose taps left the window: direct gathers, all channels, BEFORE the streamed part: the producer warp
    // is filling the ring meanwhile, so these gathers run under the first TMA round trip instead of after the last ----
    asm volatile("bar.sync 1, 256;" ::: "memory");   // consumers only (the producer warp is not part of it)
    const int n_out = s_nout;
    constexpr int kCh = 8;                             // channels per work item
    const int n_chunks = (n_c + kCh - 1) / kCh;
    for (int item = tid; item < n_out * n_chunks; item += 256) {   // empty when every tap is inside the window
        const int pid = s_outlist[item / n_chunks], c0 = c_begin + (item % n_chunks) * kCh;
        const int x = x_t0 + pid % kTW, y = y_t0 + pid / kTW;
        const float* fp = flow + b * sf.n + y * sf.h + x * sf.w;
        const BwarpGeom g = bwarp_geom(__ldg(fp), __ldg(fp + sf.c), __ldg(lin_x + x), __ldg(lin_y + y), H, W, fe);
        const int o_nw = g.y0 * W + g.x0, o_ne = g.y0 * W + g.x1, o_sw = g.y1 * W + g.x0, o_se = g.y1 * W + g.x1;
        const float* ip = img + ((int64_t)b * C + c0) * HW;
        float* op = out + ((int64_t)b * C + c0) * HW + (int64_t)y * W + x;
        const int cn = min(kCh, c_end - c0);
        float v[kCh][4];
#pragma unroll
        for (int u = 0; u < kCh; ++u) {
            if (u < cn) {
                const float* p = ip + (int64_t)u * HW;
                v[u][0] = ld_gather(p + o_nw);
                v[u][1] = ld_gather(p + o_ne);
                v[u][2] = ld_gather(p + o_sw);
                v[u][3] = ld_gather(p + o_se);
            }
        }
#pragma unroll
        for (int u = 0; u < kCh; ++u)
            if (u < cn) st_stream(op + (int64_t)u * HW, v[u][0] * g.w_nw + v[u][1] * g.w_ne + v[u][2] * g.w_sw + v[u][3] * g.w_se);
    }

    // output pointers of my two rows at plane c_begin (the second pixel of a row is 32 floats further)
    float* op0 = out + ((int64_t)b * C + c_begin) * HW + (int64_t)(y_t0 + warp) * W + (x_t0 + lane);
    float* op1 = op0 + (int64_t)8 * W;
    const uint32_t stage_u32 = smem_u32(stage);
    int s = 0;
    uint32_t parity = 0;
    for (int c = 0; c < n_c; ++c) {
        mbar_wait(&full[s], parity);
        const uint32_t sb = stage_u32 + (uint32_t)s * (kStageFloats * 4);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (ok[k]) {
                const uint32_t ad = sb + a_nw[k];
                float v0, v1, v2, v3;
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v0) : "r"(ad));
                asm volatile("ld.shared.f32 %0, [%1+4];" : "=f"(v1) : "r"(ad));
                asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v2) : "r"(ad), "n"(kRW * 4));
                asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v3) : "r"(ad), "n"(kRW * 4 + 4));
                const float r = v0 * w_nw[k] + v1 * w_ne[k] + v2 * w_sw[k] + v3 * w_se[k];
                st_stream(((k >> 1) ? op1 : op0) + 32 * (k & 1), r);
            }
        }
        op0 += HW;
        op1 += HW;
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
        if (++s == kStages) {
            s = 0;
            parity ^= 1;
        }
    }

}

struct TmapCacheEntry {
    const void* base;
    int BC, H, W;
    CUtensorMap map;
};

static int get_img_tmap(const float* img, int BC, int H, int W, const CUtensorMap** out) {
    static thread_local TmapCacheEntry cache[16];
    static thread_local int next = 0;
    for (int i = 0; i < 16; ++i)
        if (cache[i].base == img && cache[i].BC == BC && cache[i].H == H && cache[i].W == W) {
            *out = &cache[i].map;
            return VFI_OK;
        }
    TmapCacheEntry& e = cache[next];
    next = (next + 1) % 16;
    e.base = nullptr;
    int rc = make_tmap_f32_3d(&e.map, img, (uint64_t)W, (uint64_t)H, (uint64_t)BC, (uint64_t)W * 4, (uint64_t)H * W * 4, kRW, kRH, 1);
    if (rc) return rc;
    e.base = img;
    e.BC = BC;
    e.H = H;
    e.W = W;
    *out = &e.map;
    return VFI_OK;
}

}  // namespace vfi

extern "C" int vfi_bwarp_f32(const float* img, const float* flow, const float* lin_x, const float* lin_y, float* out,
                             int B, int C, int H, int W, const int64_t img_strides[4], const int64_t flow_strides[4],
                             const int64_t out_strides[4], const int* frame_ext, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(img && flow && lin_x && lin_y && out, "vfi_bwarp_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 0 && H >= 1 && W >= 1, "vfi_bwarp_f32: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
    VFI_REQUIRE(img_strides && flow_strides && out_strides, "vfi_bwarp_f32: null strides");
    if (B == 0 || C == 0) return VFI_OK;
    const Str4 si{img_strides[0], img_strides[1], img_strides[2], img_strides[3]};
    const Str4 sf{flow_strides[0], flow_strides[1], flow_strides[2], flow_strides[3]};
    const Str4 so{out_strides[0], out_strides[1], out_strides[2], out_strides[3]};
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const FrameExt fe = frame_ext_of(frame_ext, H, W);
    VFI_REQUIRE(frame_ext_ok(fe, H, W), "vfi_bwarp_f32: the crop does not lie inside the frame");

    auto aligned16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    const bool channels_last = si.c == 1 && so.c == 1 && C % 4 == 0 && C > 1 && aligned16(img) && aligned16(out) &&
                               si.w % 4 == 0 && si.h % 4 == 0 && si.n % 4 == 0 && so.w % 4 == 0 && so.h % 4 == 0 &&
                               so.n % 4 == 0;
    if (channels_last) {
        const int64_t total = (int64_t)B * H * W * (C / 4);
        const int blocks = (int)min64(ceil_div(total, 256), (int64_t)kNumSMs * 32);
        bwarp_cl_kernel<<<blocks, 256, 0, st>>>(img, flow, lin_x, lin_y, out, B, C, H, W, si, sf, so, fe, 0);
        VFI_CHECK_LAUNCH("bwarp_cl_kernel");
        return VFI_OK;
    }
    const bool nchw = si.w == 1 && si.h == W && si.c == (int64_t)H * W && si.n == (int64_t)C * H * W && so.w == 1 && so.h == W &&
                      so.c == (int64_t)H * W && so.n == (int64_t)C * H * W;
    const char* force = getenv("VFI_BWARP_ALGO");   // "gather" | "tile" (testing / profiling)
    // measured on B200 (profiles/): the tile kernel ties the gather kernel on smooth flows and is 1.5x faster on
    // rough ones once its per-CTA prologue is amortised over >= 8 channels
    const bool want_tile = force ? force[0] == 't' : ((int64_t)H * W >= 64 * 1024 && C >= 8);
    if (nchw && want_tile && W % 4 == 0 && aligned16(img) && (int64_t)B * C < (1ll << 31) && (int64_t)H * W < (1ll << 31)) {
        const CUtensorMap* tm = nullptr;
        int rc = get_img_tmap(img, B * C, H, W, &tm);
        if (rc) return rc;
        const int tiles_x = (int)ceil_div(W, kTW), tiles_y = (int)ceil_div(H, kTH);
        const size_t smem = (size_t)kStages * kStageFloats * sizeof(float) + 2 * kStages * sizeof(uint64_t) + 128;
        VFI_CUDA(cudaFuncSetAttribute(bwarp_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int c_chunk = 32;
        if (const char* e = getenv("VFI_BWARP_CCHUNK")) c_chunk = atoi(e) > 0 ? atoi(e) : c_chunk;
        const int n_tiles_all = tiles_x * tiles_y * B, n_cchunks = (int)ceil_div(C, c_chunk);
        bwarp_tile_kernel<<<(unsigned)(n_tiles_all * n_cchunks), kTileThreads, smem, st>>>(*tm, img, flow, lin_x, lin_y, out, C, H, W, sf,
                                                                                          tiles_x, tiles_y, n_tiles_all, c_chunk, fe);
        VFI_CHECK_LAUNCH("bwarp_tile_kernel");
        return VFI_OK;
    }
    // channel chunking: enough CTAs to fill the machine, flow re-read at most once per 16 channels
    const int64_t tiles = ceil_div(W, 32) * ceil_div(H, 8) * B;
    int n_chunks = 1;
    while (tiles * n_chunks < (int64_t)kNumSMs * 16 && ceil_div(C, n_chunks * 2) >= 16) n_chunks *= 2;
    int c_chunk = (int)ceil_div(C, n_chunks);
    c_chunk = (int)ceil_div(c_chunk, kCUnroll) * kCUnroll;
    n_chunks = (int)ceil_div(C, c_chunk);
    VFI_REQUIRE((int64_t)B * n_chunks <= 65535, "vfi_bwarp_f32: B*chunks=%lld exceeds grid.z", (long long)B * n_chunks);
    VFI_REQUIRE(ceil_div(H, 8) <= 65535, "vfi_bwarp_f32: H too large");
    dim3 grid((unsigned)ceil_div(W, 32), (unsigned)ceil_div(H, 8), (unsigned)(B * n_chunks));
    bwarp_planar_kernel<<<grid, dim3(32, 8), 0, st>>>(img, flow, lin_x, lin_y, out, C, H, W, c_chunk, n_chunks, si, sf, so, fe);
    VFI_CHECK_LAUNCH("bwarp_planar_kernel");
    return VFI_OK;
}

extern "C" int vfi_bwarp_flowcat_f32(const float* img, const float* flow, const float* lin_x, const float* lin_y, float* out,
                                     int B, int C, int H, int W, const int64_t img_strides[4], const int64_t flow_strides[4],
                                     const int64_t out_strides[4], const int* frame_ext, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;
    VFI_REQUIRE(img && flow && lin_x && lin_y && out && img_strides && flow_strides && out_strides, "vfi_bwarp_flowcat_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 4 && C % 4 == 0 && H >= 1 && W >= 1, "vfi_bwarp_flowcat_f32: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
    const Str4 si{img_strides[0], img_strides[1], img_strides[2], img_strides[3]};
    const Str4 sf{flow_strides[0], flow_strides[1], flow_strides[2], flow_strides[3]};
    const Str4 so{out_strides[0], out_strides[1], out_strides[2], out_strides[3]};
    const FrameExt fe = frame_ext_of(frame_ext, H, W);
    VFI_REQUIRE(frame_ext_ok(fe, H, W), "vfi_bwarp_flowcat_f32: the crop does not lie inside the frame");
    auto aligned16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    const bool channels_last = si.c == 1 && so.c == 1 && aligned16(img) && aligned16(out) && si.w % 4 == 0 && si.h % 4 == 0 &&
                               si.n % 4 == 0 && so.w % 4 == 0 && so.h % 4 == 0 && so.n % 4 == 0 && so.w >= C + 4;
    if (!channels_last) VFI_UNSUPPORTED("vfi_bwarp_flowcat_f32: channels-last img / out with 16-byte aligned pixels and >= C + 4 output channels per pixel");
    const int64_t total = (int64_t)B * H * W * (C / 4);
    const int blocks = (int)min64(ceil_div(total, 256), (int64_t)kNumSMs * 32);
    bwarp_cl_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(img, flow, lin_x, lin_y, out, B, C, H, W, si, sf, so, fe, 1);
    VFI_CHECK_LAUNCH("bwarp_cl_kernel");
    return VFI_OK;
}
