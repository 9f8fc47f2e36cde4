// Shared helpers for the vfi_b200 kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/vfi_b200.h"

namespace vfi {

constexpr int kNumSMs = 148;  // B200

void set_error(const char* fmt, ...);
void note_launch();

#define VFI_REQUIRE(cond, ...)                         \
    do {                                               \
        if (!(cond)) {                                 \
            ::vfi::set_error(__VA_ARGS__);             \
            return VFI_ERR_INVALID_ARGUMENT;           \
        }                                              \
    } while (0)

#define VFI_UNSUPPORTED(...)                           \
    do {                                               \
        ::vfi::set_error(__VA_ARGS__);                 \
        return VFI_ERR_UNSUPPORTED;                    \
    } while (0)

// Check the launch that was just issued (no synchronisation).
#define VFI_CHECK_LAUNCH(name)                                                        \
    do {                                                                              \
        ::vfi::note_launch();                                                         \
        cudaError_t e__ = cudaGetLastError();                                         \
        if (e__ != cudaSuccess) {                                                     \
            ::vfi::set_error("%s: launch failed: %s", name, cudaGetErrorString(e__)); \
            return VFI_ERR_CUDA;                                                      \
        }                                                                             \
    } while (0)

#define VFI_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) {                                                        \
            ::vfi::set_error("%s failed: %s", #call, cudaGetErrorString(e__));           \
            return VFI_ERR_CUDA;                                                         \
        }                                                                                \
    } while (0)

// A tensor that is a crop of a larger frame (4K tiles): frame extent and the crop's origin in it.
struct FrameExt {
    int Wg, Hg, x0, y0;
};
// frame_ext = {Wg, Hg, x0, y0} (host ints) or NULL: the tensor is the whole frame
static inline FrameExt frame_ext_of(const int* frame_ext, int H, int W) {
    FrameExt fe;
    if (frame_ext) { fe.Wg = frame_ext[0]; fe.Hg = frame_ext[1]; fe.x0 = frame_ext[2]; fe.y0 = frame_ext[3]; }
    else { fe.Wg = W; fe.Hg = H; fe.x0 = 0; fe.y0 = 0; }
    return fe;
}
static inline bool frame_ext_ok(const FrameExt& fe, int H, int W) {
    return fe.x0 >= 0 && fe.y0 >= 0 && fe.Wg >= 1 && fe.Hg >= 1 && (int64_t)fe.x0 + W <= fe.Wg && (int64_t)fe.y0 + H <= fe.Hg;
}

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t min64(int64_t a, int64_t b) { return a < b ? a : b; }

// ---- device helpers ---------------------------------------------------------------------------

// streaming (read-once / write-once) accesses: keep them out of L1
__device__ __forceinline__ float ld_stream(const float* p) {
    float v;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ float4 ld_stream4(const float* p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_stream(float* p, float v) {
    asm volatile("st.global.cs.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void st_stream4(float* p, float4 v) {
    asm volatile("st.global.cs.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
// gather loads through the read-only path (cached in L1: neighbouring lanes share lines)
__device__ __forceinline__ float ld_gather(const float* p) { return __ldg(p); }

// ---- packed fp32 pairs (sm_100a FFMA2: two IEEE fp32 FMAs per issue slot; a scalar operand broadcasts) ----------
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
// pair * scalar + pair (ptxas folds the broadcast into the FFMA2 operand, no extra MOV)
__device__ __forceinline__ f32x2 fma2s(f32x2 a, float b, f32x2 c) { return fma2(a, pack2(b, b), c); }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 mul2s(f32x2 a, float b) {
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(pack2(b, b)));
    return d;
}

// Replay of the reference's normalise -> grid_sample unnormalise round trip, fp32, no FMA contraction
// (reference: modules/warp.py:9-13 + ATen grid_sampler unnormalize, align_corners=True).
//   g = lin + d / half_extent ;  i = ((g + 1) / 2) * (size - 1)
__device__ __forceinline__ float replay_coord(float lin, float d_scaled, float size_m1) {
    const float g = __fadd_rn(lin, d_scaled);
    return __fmul_rn(__fmul_rn(__fadd_rn(g, 1.0f), 0.5f), size_m1);
}

}  // namespace vfi
