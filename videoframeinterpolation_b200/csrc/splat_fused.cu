// Gather-form forward splatting: no floating-point atomics, divide fused, no zero-fill / normalise passes.
// Reference semantics: modules/softsplat.py:289-341 (kernel text) + modules/warp.py:18-59 (pre / post).
//
// Why not the scatter form: the reference does 4 fp32 atomicAdds per (pixel, channel).  On sm_100a a global
// RED costs ~1.3 cycles per LANE and a shared-memory fp32 atomicAdd is a CAS loop (SASS: ATOMS.CAST.SPIN), so
// any per-channel atomic scheme is bound by the atomic units (measured: 0.55 TB/s at C = 64), not by HBM.
// Here the only atomics are 4 INTEGER operations per source pixel, independent of C:
//   K1 (link):   every source pixel computes its 4 corners and weights once.  Corner class k (0 = NW .. 3 = SE)
//                landing on output pixel q claims slot[q][k] with one atomicCAS; a second source of the same
//                class (converging flow) is pushed onto q's overflow list instead
//                  (node id = 4 * source_pixel + k, nodes[id] = {next, weight}, next = atomicExch(&head[q], id)).
//   K2 (gather): every output pixel reads its 4 class slots (two 128-bit loads) and up to 4 overflow nodes,
//                then for each channel sums w_k * in[c][p_k] -- gathers shaped like a backward warp: tap k of
//                neighbouring output pixels comes from neighbouring source pixels -- divides by the summed
//                weights per the eps rule and stores the result once.
// Lists have no capacity limit (the node pool is exactly 4 per source pixel); overflow beyond the 4 kept in
// registers is re-walked per channel.  Summation order is arbitrary, as it is in the reference.
// Workspace per source pixel: 32 B slots + 4 B head + 32 B nodes.
#include <stdlib.h>

#include "vfi_common.cuh"

namespace vfi {

struct SplatNode {
    int next;
    float w;
};

struct SplatCornersF {
    int x0, y0;
    float w_nw, w_ne, w_sw, w_se;
    bool valid;
};

// Same fp32 expressions as the reference kernel (modules/softsplat.py:303-323); identical to splat.cu.
__device__ __forceinline__ SplatCornersF splat_corners_f(int x, int y, float flow_x, float flow_y, int H, int W) {
    SplatCornersF c;
    const float fx = (float)x + flow_x;
    const float fy = (float)y + flow_y;
    c.valid = isfinite(fx) && isfinite(fy) && fx > -1.0f && fy > -1.0f && fx < (float)W && fy < (float)H;
    const float xf = floorf(c.valid ? fx : 0.0f), yf = floorf(c.valid ? fy : 0.0f);
    c.x0 = (int)xf;
    c.y0 = (int)yf;
    const float x1f = (float)(c.x0 + 1), y1f = (float)(c.y0 + 1);
    c.w_nw = (x1f - fx) * (y1f - fy);
    c.w_ne = (fx - xf) * (y1f - fy);
    c.w_sw = (x1f - fx) * (fy - yf);
    c.w_se = (fx - xf) * (fy - yf);
    return c;
}

__device__ __forceinline__ float splat_weight_f(int mode, const float* __restrict__ metric, const float* __restrict__ t, int b,
                                                int64_t idx) {
    switch (mode) {
        case VFI_SPLAT_W_LINEAR: return ld_stream(metric + idx);
        case VFI_SPLAT_W_SOFT: return expf(ld_stream(metric + idx));
        case VFI_SPLAT_W_SOFT_CLIP_T: return t[b] * expf(fminf(fmaxf(ld_stream(metric + idx), -20.0f), 20.0f));   // as splat.cu
        default: return 1.0f;
    }
}

// K1: one thread per source pixel
__global__ void __launch_bounds__(256) splat_link_kernel(const float* __restrict__ flow, const float* __restrict__ metric,
                                                         const float* __restrict__ t, int* __restrict__ slot_src, float* __restrict__ slot_w,
                                                         int* __restrict__ head, SplatNode* __restrict__ nodes, int B, int H,
                                                         int W, int mode) {
    const int64_t HW = (int64_t)H * W, total = (int64_t)B * HW;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int b = (int)(idx / HW);
        const int64_t pix = idx - (int64_t)b * HW;
        const int y = (int)(pix / W), x = (int)(pix - (int64_t)y * W);
        const float* fp = flow + (int64_t)b * 2 * HW + pix;
        const SplatCornersF c = splat_corners_f(x, y, ld_stream(fp), ld_stream(fp + HW), H, W);
        if (!c.valid) continue;
        const float wm = splat_weight_f(mode, metric, t, b, idx);
        const bool in_x0 = c.x0 >= 0, in_x1 = c.x0 + 1 < W, in_y0 = c.y0 >= 0, in_y1 = c.y0 + 1 < H;
        const int64_t q = (int64_t)b * HW + (int64_t)c.y0 * W + c.x0;     // output pixel of the north-west corner
        const int node0 = (int)(idx * 4);
        // the four claims are independent: issue all four atomics before looking at any result (the kernel is bound
        // by the atomic round trip, 4 in flight per thread instead of 1)
        const bool on[4] = {in_x0 && in_y0, in_x1 && in_y0, in_x0 && in_y1, in_x1 && in_y1};
        const int64_t qq[4] = {q, q + 1, q + W, q + W + 1};
        const float wk[4] = {wm * c.w_nw, wm * c.w_ne, wm * c.w_sw, wm * c.w_se};
        int prev[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) prev[k] = on[k] ? atomicCAS(slot_src + qq[k] * 4 + k, -1, (int)pix) : 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (!on[k]) continue;
            if (prev[k] == -1) slot_w[qq[k] * 4 + k] = wk[k];
            else nodes[node0 + k] = SplatNode{atomicExch(head + qq[k], node0 + k), wk[k]};
        }
    }
}

__device__ __forceinline__ float splat_divisor_f(float d, int eps_mode, float eps) {
    if (eps_mode == VFI_EPS_ADD) return d + eps;
    if (eps_mode == VFI_EPS_ZERO) return d == 0.0f ? 1.0f : d;
    return fmaxf(d, eps);
}

constexpr int kOvTaps = 4;    // overflow nodes kept in registers; longer lists take the re-walk path
constexpr int kChUnroll = 4;

// K2: one thread per output pixel (a warp = 32 consecutive x)
__global__ void __launch_bounds__(256) splat_gather_kernel(const float* __restrict__ in, const int* __restrict__ slot_src,
                                                           const float* __restrict__ slot_w, const int* __restrict__ head,
                                                           const SplatNode* __restrict__ nodes, float* __restrict__ out, int C,
                                                           int H, int W, int c_chunk, int n_chunks, int normalize, int eps_mode,
                                                           float eps, int lockstep, float* __restrict__ den_io, int acc_mode,
                                                           uint8_t* __restrict__ blank, float blank_thr) {
    const int x = blockIdx.x * 32 + threadIdx.x;
    const int y = blockIdx.y * 8 + threadIdx.y;
    const int b = blockIdx.z / n_chunks;
    const int c_begin = (blockIdx.z % n_chunks) * c_chunk;
    const bool live = x < W && y < H;
    if (!live && !lockstep) return;
    const int c_end = min(C, c_begin + c_chunk);
    const int64_t HW = (int64_t)H * W;
    const int64_t q = (int64_t)b * HW + (int64_t)min(y, H - 1) * W + min(x, W - 1);
    const int64_t base4 = (int64_t)b * HW * 4;   // node ids of this image start here

    constexpr int NT = 4 + kOvTaps;
    int src[NT];
    float wgt[NT];
    {
        const int4 s4 = __ldg(reinterpret_cast<const int4*>(slot_src) + q);
        const float4 w4 = __ldg(reinterpret_cast<const float4*>(slot_w) + q);
        src[0] = s4.x; src[1] = s4.y; src[2] = s4.z; src[3] = s4.w;
        wgt[0] = s4.x >= 0 ? w4.x : 0.0f;
        wgt[1] = s4.y >= 0 ? w4.y : 0.0f;
        wgt[2] = s4.z >= 0 ? w4.z : 0.0f;
        wgt[3] = s4.w >= 0 ? w4.w : 0.0f;
    }
    int node = __ldg(head + q);
#pragma unroll
    for (int k = 0; k < kOvTaps; ++k) {
        src[4 + k] = -1;
        wgt[4 + k] = 0.0f;
        if (node >= 0) {
            const int2 e = __ldg(reinterpret_cast<const int2*>(nodes) + node);
            src[4 + k] = (int)(((int64_t)node - base4) >> 2);
            wgt[4 + k] = __int_as_float(e.y);
            node = e.x;
        }
    }
    const int rest = node;                 // >= 0: more overflow nodes than registers
    // warp-uniform flags: does ANY lane of this warp hold a k-th overflow tap / a longer list
    unsigned ov_any = 0;
#pragma unroll
    for (int k = 0; k < kOvTaps; ++k) ov_any |= __any_sync(0xffffffffu, src[4 + k] >= 0) ? (1u << k) : 0u;
    const bool rest_any = __any_sync(0xffffffffu, rest >= 0);
    float den = 0.0f;
#pragma unroll
    for (int k = 0; k < NT; ++k) den += wgt[k];   // absent entries hold 0
    for (int nd = rest; nd >= 0;) {
        const int2 e = __ldg(reinterpret_cast<const int2*>(nodes) + nd);
        den += __int_as_float(e.y);
        nd = e.x;
    }
    // one reciprocal per pixel instead of an IEEE division per (pixel, channel): <= 1.5 ulp from the reference's
    // divide, far inside the 1e-4 splat tolerance (the summation order already differs from the atomics')
    // accumulate form (two-frame / multi-branch splat, modules/warp.py:68-114): this source's un-normalised sums are
    // stored (acc_mode 0) or added (acc_mode 1) to out / den_io; acc_mode 2 = last source: added, then divided by the
    // total weight in the same pass (den_io is only read).  Every (pixel, plane) has exactly one owner thread, so the
    // read-modify-write needs no atomic; den_io and the blank mask belong to the first channel chunk.
    const float den_fin = acc_mode == 2 ? den_io[q] + den : den;
    const float dv = splat_divisor_f(den_fin, eps_mode, eps);
    const float rdiv = normalize ? __frcp_rn(dv) : 1.0f;
    if (den_io != nullptr && c_begin == 0 && live && acc_mode != 2) den_io[q] = acc_mode ? den_io[q] + den : den;
    if (blank != nullptr && c_begin == 0 && live) blank[q] = dv < blank_thr ? 1 : 0;

    // byte offsets of the source pixels inside a plane (32-bit: H*W*4 < 2^32); plane bases are block-uniform, so a
    // tap is one IMAD.WIDE.U32 + one predicated LDG + one predicated FFMA
    unsigned so[NT];   // source pixel index (0 when the tap is absent; its loads are predicated off)
    bool on[NT];
#pragma unroll
    for (int k = 0; k < NT; ++k) {
        on[k] = src[k] >= 0;
        so[k] = on[k] ? (unsigned)src[k] : 0u;
    }
    const int64_t plane_bytes = HW * 4;
    const char* ibase = reinterpret_cast<const char*>(in) + ((int64_t)b * C + c_begin) * plane_bytes;
    char* obase = reinterpret_cast<char*>(out) + ((int64_t)b * C + c_begin) * plane_bytes;
    const unsigned oo = live ? (unsigned)(y * W + x) * 4u : 0u;
    auto tap = [](const char* plane, unsigned pixel) {
        const float* p;   // one IMAD.WIDE.U32 (left to itself the compiler re-loads the base and does a 3-term 64-bit add)
        asm("mad.wide.u32 %0, %1, 4, %2;" : "=l"(p) : "r"(pixel), "l"(plane));
        return ld_gather(p);
    };
    const int n_c = c_end - c_begin;
    int c = 0;
    for (; c + kChUnroll <= n_c; c += kChUnroll) {
        // the 8 rows of the CTA walk the planes in lockstep: row y's south taps and row y+1's north taps are the same
        // sectors, and they only meet in the L1 if the warps are on the same plane at the same time
        if (lockstep) __syncthreads();
        float acc[kChUnroll];
#pragma unroll
        for (int u = 0; u < kChUnroll; ++u) {
            const char* pl = ibase + (int64_t)(c + u) * plane_bytes;
            acc[u] = 0.0f;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (on[k]) acc[u] = fmaf(tap(pl, so[k]), wgt[k], acc[u]);
        }
        if (ov_any) {                               // warp-uniform
#pragma unroll
            for (int k = 0; k < kOvTaps; ++k) {
                if (ov_any & (1u << k)) {           // warp-uniform
#pragma unroll
                    for (int u = 0; u < kChUnroll; ++u)
                        if (on[4 + k]) acc[u] = fmaf(tap(ibase + (int64_t)(c + u) * plane_bytes, so[4 + k]), wgt[4 + k], acc[u]);
                }
            }
            if (rest_any) {
                for (int nd = rest; nd >= 0;) {
                    const int2 e = __ldg(reinterpret_cast<const int2*>(nodes) + nd);
                    const unsigned sp = (unsigned)(((int64_t)nd - base4) >> 2);
#pragma unroll
                    for (int u = 0; u < kChUnroll; ++u)
                        acc[u] = fmaf(tap(ibase + (int64_t)(c + u) * plane_bytes, sp), __int_as_float(e.y), acc[u]);
                    nd = e.x;
                }
            }
        }
#pragma unroll
        for (int u = 0; u < kChUnroll; ++u)
            if (live) {
                float* op = reinterpret_cast<float*>(obase + (int64_t)(c + u) * plane_bytes + oo);
                st_stream(op, acc_mode == 2 ? (*op + acc[u]) * rdiv : (acc_mode ? *op + acc[u] * rdiv : acc[u] * rdiv));
            }
    }
    for (; c < n_c; ++c) {
        const char* pl = ibase + (int64_t)c * plane_bytes;
        float acc = 0.0f;
#pragma unroll
        for (int k = 0; k < NT; ++k)
            if (on[k]) acc = fmaf(tap(pl, so[k]), wgt[k], acc);
        for (int nd = rest; nd >= 0;) {
            const int2 e = __ldg(reinterpret_cast<const int2*>(nodes) + nd);
            acc = fmaf(tap(pl, (unsigned)(((int64_t)nd - base4) >> 2)), __int_as_float(e.y), acc);
            nd = e.x;
        }
        if (live) {
            float* op = reinterpret_cast<float*>(obase + (int64_t)c * plane_bytes + oo);
            st_stream(op, acc_mode == 2 ? (*op + acc) * rdiv : (acc_mode ? *op + acc * rdiv : acc * rdiv));
        }
    }
}

}  // namespace vfi

extern "C" size_t vfi_splat_workspace_bytes(int B, int C, int H, int W) {
    (void)C;
    if (B <= 0 || H <= 0 || W <= 0) return 0;
    const size_t px = (size_t)B * H * W;
    return px * 4 * (sizeof(int) + sizeof(float)) + px * sizeof(int) + px * 4 * sizeof(vfi::SplatNode) + 256;
}

namespace vfi {

// link + gather for one source; shared by vfi_splat_fused_f32 (normalised, one source) and vfi_splat_f32's gather route
// (un-normalised sums stored or accumulated into num / den)
int launch_splat_gather(const float* in, const float* flow, const float* metric, const float* t, float* out, float* den_io, int B, int C,
                        int H, int W, int weight_mode, int normalize, int eps_mode, float eps_value, int acc_mode, void* workspace,
                        cudaStream_t st, uint8_t* blank, float blank_thr) {
    const int64_t px = (int64_t)B * H * W;
    // layout: slot_src[4 px] | head[px] (one memset of 0xFF covers both) | slot_w[4 px] | nodes[4 px]
    int* slot_src = reinterpret_cast<int*>(workspace);
    int* head = slot_src + px * 4;
    float* slot_w = reinterpret_cast<float*>(head + ((px + 3) & ~3ll));
    SplatNode* nodes = reinterpret_cast<SplatNode*>(slot_w + px * 4);
    VFI_CUDA(cudaMemsetAsync(slot_src, 0xFF, (size_t)px * 5 * sizeof(int), st));
    {
        const int blocks = (int)min64(ceil_div(px, 256), (int64_t)kNumSMs * 16);
        splat_link_kernel<<<blocks, 256, 0, st>>>(flow, metric, t, slot_src, slot_w, head, nodes, B, H, W, weight_mode);
        VFI_CHECK_LAUNCH("splat_link_kernel");
    }
    const int64_t tiles = ceil_div(W, 32) * ceil_div(H, 8) * B;
    int n_chunks = 1;
    while (tiles * n_chunks < (int64_t)kNumSMs * 16 && ceil_div(C, n_chunks * 2) >= 16) n_chunks *= 2;
    int c_chunk = (int)ceil_div(C, n_chunks);
    c_chunk = (int)ceil_div(c_chunk, kChUnroll) * kChUnroll;
    n_chunks = (int)ceil_div(C, c_chunk);
    VFI_REQUIRE((int64_t)B * n_chunks <= 65535 && ceil_div(H, 8) <= 65535, "splat (gather form): grid too large");
    dim3 grid((unsigned)ceil_div(W, 32), (unsigned)ceil_div(H, 8), (unsigned)(B * n_chunks));
    int lockstep = 1;
    if (const char* e = getenv("VFI_SPLAT_LOCKSTEP")) lockstep = atoi(e);
    splat_gather_kernel<<<grid, dim3(32, 8), 0, st>>>(in, slot_src, slot_w, head, nodes, out, C, H, W, c_chunk, n_chunks, normalize,
                                                      eps_mode, eps_value, lockstep, den_io, acc_mode, blank, blank_thr);
    VFI_CHECK_LAUNCH("splat_gather_kernel");
    return VFI_OK;
}

// gather route of vfi_splat_f32: 0 = not applicable (caller scatters), else done / error
bool splat_gather_applicable(int B, int C, int H, int W, const void* workspace, size_t workspace_bytes) {
    const int64_t px = (int64_t)B * H * W;
    return workspace != nullptr && C >= 1 && px * 4 < (1ll << 31) && (reinterpret_cast<uintptr_t>(workspace) & 15) == 0 &&
           workspace_bytes >= vfi_splat_workspace_bytes(B, C, H, W);
}

}  // namespace vfi

extern "C" int vfi_splat_fused_f32(const float* in, const float* flow, const float* metric, float* out, int B, int C, int H, int W,
                                   int weight_mode, int eps_mode, float eps_value, void* workspace, size_t workspace_bytes,
                                   void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(in && flow && out, "vfi_splat_fused_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 0 && H >= 1 && W >= 1, "vfi_splat_fused_f32: bad shape");
    VFI_REQUIRE(weight_mode >= VFI_SPLAT_W_NONE && weight_mode <= VFI_SPLAT_W_SOFT, "vfi_splat_fused_f32: bad weight_mode %d",
                weight_mode);
    VFI_REQUIRE(eps_mode >= VFI_EPS_ADD && eps_mode <= VFI_EPS_CLIP, "vfi_splat_fused_f32: bad eps_mode %d", eps_mode);
    if (weight_mode >= VFI_SPLAT_W_LINEAR) VFI_REQUIRE(metric != nullptr, "vfi_splat_fused_f32: weight mode %d needs a metric", weight_mode);
    if (B == 0 || C == 0) return VFI_OK;
    const int64_t px = (int64_t)B * H * W;
    if (px * 4 >= (1ll << 31)) VFI_UNSUPPORTED("vfi_splat_fused_f32: B*H*W = %lld needs 64-bit node ids; use vfi_splat_f32", (long long)px);
    VFI_REQUIRE(workspace && workspace_bytes >= vfi_splat_workspace_bytes(B, C, H, W), "vfi_splat_fused_f32: workspace of %zu B needed",
                vfi_splat_workspace_bytes(B, C, H, W));
    VFI_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "vfi_splat_fused_f32: workspace must be 16-byte aligned");
    return launch_splat_gather(in, flow, metric, nullptr, out, nullptr, B, C, H, W, weight_mode, weight_mode == VFI_SPLAT_W_NONE ? 0 : 1,
                               eps_mode, eps_value, 0, workspace, static_cast<cudaStream_t>(stream), nullptr, 0.0f);
}

extern "C" int vfi_splat_last_f32(const float* in, const float* flow, const float* metric, const float* t, float* num_out, const float* den,
                                  uint8_t* blank_mask, int B, int C, int H, int W, int weight_mode, int eps_mode, float eps_value,
                                  float blank_threshold, void* workspace, size_t workspace_bytes, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;
    VFI_REQUIRE(in && flow && num_out && den, "vfi_splat_last_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1, "vfi_splat_last_f32: bad shape");
    VFI_REQUIRE(weight_mode >= VFI_SPLAT_W_ONES && weight_mode <= VFI_SPLAT_W_SOFT_CLIP_T, "vfi_splat_last_f32: bad weight_mode %d", weight_mode);
    VFI_REQUIRE(eps_mode >= VFI_EPS_ADD && eps_mode <= VFI_EPS_CLIP, "vfi_splat_last_f32: bad eps_mode %d", eps_mode);
    if (weight_mode >= VFI_SPLAT_W_LINEAR) VFI_REQUIRE(metric != nullptr, "vfi_splat_last_f32: weight mode %d needs a metric", weight_mode);
    if (weight_mode == VFI_SPLAT_W_SOFT_CLIP_T) VFI_REQUIRE(t != nullptr, "vfi_splat_last_f32: SOFT_CLIP_T needs t");
    if (!splat_gather_applicable(B, C, H, W, workspace, workspace_bytes))
        VFI_UNSUPPORTED("vfi_splat_last_f32: needs a 16-byte aligned workspace of %zu B and B*H*W < 2^29 (else vfi_splat_f32 + vfi_splat_normalize_f32)",
                        vfi_splat_workspace_bytes(B, C, H, W));
    return launch_splat_gather(in, flow, metric, t, num_out, const_cast<float*>(den), B, C, H, W, weight_mode, 1, eps_mode, eps_value, 2,
                               workspace, static_cast<cudaStream_t>(stream), blank_mask, blank_threshold);
}
