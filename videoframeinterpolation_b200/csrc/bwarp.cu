// Backward bilinear warp with border clamp (reference: modules/warp.py:7-15).
//
// HBM-bound: 4*(2C+2) bytes per pixel.  Coordinates, corner indices and the four weights are computed
// once per pixel (flow read once) and reused for every channel; the channel loop keeps 4 channels x 4
// taps = 16 independent gathers in flight per thread.  Two layouts:
//   * planar (NCHW, stride_w == 1): a warp covers 32 consecutive x, so output stores are 128 B
//     coalesced and the taps of neighbouring lanes share L1 lines when the flow is locally smooth;
//   * channels-last (stride_c == 1): a thread owns 4 consecutive channels of one pixel, all taps and the
//     store are 128-bit.
#include "vfi_common.cuh"

namespace vfi {

struct Str4 {
    int64_t n, c, h, w;
};

struct BwarpGeom {
    int x0, x1, y0, y1;
    float w_nw, w_ne, w_sw, w_se;
};

__device__ __forceinline__ BwarpGeom bwarp_geom(float fx, float fy, float lin_x, float lin_y, int H, int W,
                                                float half_w, float half_h) {
    // reference op order: flow / ((W-1)/2) is a true fp32 division (torch CPU), then grid + flow_
    const float wm1 = (float)(W - 1), hm1 = (float)(H - 1);
    float ix = replay_coord(lin_x, __fdiv_rn(fx, half_w), wm1);
    float iy = replay_coord(lin_y, __fdiv_rn(fy, half_h), hm1);
    ix = fminf(fmaxf(ix, 0.0f), wm1);  // padding_mode='border'
    iy = fminf(fmaxf(iy, 0.0f), hm1);
    const float xf = floorf(ix), yf = floorf(iy);
    const float tx = ix - xf, ty = iy - yf;
    const float ex = 1.0f - tx, ey = 1.0f - ty;
    BwarpGeom g;
    g.x0 = (int)xf;
    g.y0 = (int)yf;
    g.x1 = min(g.x0 + 1, W - 1);  // weight is exactly 0 whenever this clamps
    g.y1 = min(g.y0 + 1, H - 1);
    g.w_nw = ey * ex;
    g.w_ne = ey * tx;
    g.w_sw = ty * ex;
    g.w_se = ty * tx;
    return g;
}

constexpr int kCUnroll = 4;

// planar / generic-stride kernel: thread = one pixel, loops over a chunk of channels
__global__ void __launch_bounds__(256) bwarp_planar_kernel(const float* __restrict__ img, const float* __restrict__ flow,
                                                           const float* __restrict__ lin_x,
                                                           const float* __restrict__ lin_y, float* __restrict__ out,
                                                           int C, int H, int W, int c_chunk, int n_chunks, Str4 si,
                                                           Str4 sf, Str4 so) {
    const int x = blockIdx.x * 32 + threadIdx.x;
    const int y = blockIdx.y * 8 + threadIdx.y;
    const int b = blockIdx.z / n_chunks;
    const int c_begin = (blockIdx.z % n_chunks) * c_chunk;
    if (x >= W || y >= H) return;
    const int c_end = min(C, c_begin + c_chunk);

    const float* fp = flow + b * sf.n + y * sf.h + x * sf.w;
    const float fx = ld_stream(fp), fy = ld_stream(fp + sf.c);
    const float half_w = (float)((W - 1.0) / 2.0), half_h = (float)((H - 1.0) / 2.0);
    const BwarpGeom g = bwarp_geom(fx, fy, lin_x[x], lin_y[y], H, W, half_w, half_h);

    const int64_t o_nw = g.y0 * si.h + g.x0 * si.w, o_ne = g.y0 * si.h + g.x1 * si.w;
    const int64_t o_sw = g.y1 * si.h + g.x0 * si.w, o_se = g.y1 * si.h + g.x1 * si.w;
    const float* ip = img + b * si.n + c_begin * si.c;
    float* op = out + b * so.n + c_begin * so.c + y * so.h + x * so.w;

    int c = c_begin;
    for (; c + kCUnroll <= c_end; c += kCUnroll) {
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
            st_stream(op + u * so.c, r);
        }
        ip += kCUnroll * si.c;
        op += kCUnroll * so.c;
    }
    for (; c < c_end; ++c) {
        const float r = ld_gather(ip + o_nw) * g.w_nw + ld_gather(ip + o_ne) * g.w_ne + ld_gather(ip + o_sw) * g.w_sw +
                        ld_gather(ip + o_se) * g.w_se;
        st_stream(op, r);
        ip += si.c;
        op += so.c;
    }
}

// channels-last kernel: thread = (pixel, 4 consecutive channels); requires C % 4 == 0, stride_c == 1,
// 16-byte aligned bases and pixel strides.
__global__ void __launch_bounds__(256) bwarp_cl_kernel(const float* __restrict__ img, const float* __restrict__ flow,
                                                       const float* __restrict__ lin_x, const float* __restrict__ lin_y,
                                                       float* __restrict__ out, int B, int C, int H, int W, Str4 si,
                                                       Str4 sf, Str4 so) {
    const int cv = C >> 2;
    const int64_t total = (int64_t)B * H * W * cv;
    const float half_w = (float)((W - 1.0) / 2.0), half_h = (float)((H - 1.0) / 2.0);
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int c4 = (int)(idx % cv) * 4;
        int64_t p = idx / cv;
        const int x = (int)(p % W);
        p /= W;
        const int y = (int)(p % H);
        const int b = (int)(p / H);
        const float* fp = flow + b * sf.n + y * sf.h + x * sf.w;
        const BwarpGeom g = bwarp_geom(__ldg(fp), __ldg(fp + sf.c), lin_x[x], lin_y[y], H, W, half_w, half_h);
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
    }
}

}  // namespace vfi

extern "C" int vfi_bwarp_f32(const float* img, const float* flow, const float* lin_x, const float* lin_y, float* out,
                             int B, int C, int H, int W, const int64_t img_strides[4], const int64_t flow_strides[4],
                             const int64_t out_strides[4], void* stream) {
    using namespace vfi;
    VFI_REQUIRE(img && flow && lin_x && lin_y && out, "vfi_bwarp_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 0 && H >= 1 && W >= 1, "vfi_bwarp_f32: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
    VFI_REQUIRE(img_strides && flow_strides && out_strides, "vfi_bwarp_f32: null strides");
    if (B == 0 || C == 0) return VFI_OK;
    const Str4 si{img_strides[0], img_strides[1], img_strides[2], img_strides[3]};
    const Str4 sf{flow_strides[0], flow_strides[1], flow_strides[2], flow_strides[3]};
    const Str4 so{out_strides[0], out_strides[1], out_strides[2], out_strides[3]};
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    auto aligned16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    const bool channels_last = si.c == 1 && so.c == 1 && C % 4 == 0 && C > 1 && aligned16(img) && aligned16(out) &&
                               si.w % 4 == 0 && si.h % 4 == 0 && si.n % 4 == 0 && so.w % 4 == 0 && so.h % 4 == 0 &&
                               so.n % 4 == 0;
    if (channels_last) {
        const int64_t total = (int64_t)B * H * W * (C / 4);
        const int blocks = (int)min64(ceil_div(total, 256), (int64_t)kNumSMs * 32);
        bwarp_cl_kernel<<<blocks, 256, 0, st>>>(img, flow, lin_x, lin_y, out, B, C, H, W, si, sf, so);
        VFI_CHECK_LAUNCH("bwarp_cl_kernel");
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
    bwarp_planar_kernel<<<grid, dim3(32, 8), 0, st>>>(img, flow, lin_x, lin_y, out, C, H, W, c_chunk, n_chunks, si, sf, so);
    VFI_CHECK_LAUNCH("bwarp_planar_kernel");
    return VFI_OK;
}
