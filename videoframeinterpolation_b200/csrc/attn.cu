// Per-pixel multi-head attention over a MATERIALISED sample tensor (API parity with the reference's
// Attn.forward / DeformAttnwMotion.forward: modules/deformable_attn.py:26-40, modules/madat.py:86-115).
// The fused sampler+attention kernel that never builds kv is in dattn.cu; this one exists so the
// reference's `attn(q, kv)` call keeps working.
//
// Algebra (exact up to fp32 reassociation): with qp = Wq q + bq,
//   logit[h,s] = scale * ( (Wk_h^T qp_h) . kv_s + qp_h . bk_h ),   out_h = Wv_h (sum_s a[h,s] kv_s) + bv_h
// so k and v are never formed: ~C/hc times fewer flops than projecting every sample.
// Thread = (pixel, head); a warp covers 32 consecutive pixels, so kv reads are 128 B coalesced.
#include "vfi_common.cuh"

namespace vfi {

constexpr int kAttnPix = 32;

// smem: kt[C][T] and yv[C][T] (T = threads per block, thread index fastest -> conflict free)
__global__ void attn_kernel(const float* __restrict__ q, const float* __restrict__ kv, const float* __restrict__ Wq,
                            const float* __restrict__ bq, const float* __restrict__ Wk, const float* __restrict__ bk,
                            const float* __restrict__ Wv, const float* __restrict__ bv, float* __restrict__ out,
                            float* __restrict__ score0, float* __restrict__ score1, int C, int64_t HW, int S, int nH) {
    extern __shared__ float smem[];
    const int T = blockDim.x * blockDim.y;
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    float* kt = smem + tid;            // kt[c * T]
    float* yv = smem + (size_t)C * T + tid;
    const int h = threadIdx.y;
    const int b = blockIdx.y;
    const int64_t pix = (int64_t)blockIdx.x * kAttnPix + threadIdx.x;
    if (pix >= HW) return;
    const int hc = C / nH;
    const float scale = rsqrtf((float)hc);
    const float* qpix = q + (int64_t)b * C * HW + pix;

    for (int c = 0; c < C; ++c) {
        kt[(size_t)c * T] = 0.0f;
        yv[(size_t)c * T] = 0.0f;
    }
    float cb = 0.0f;
    for (int j = 0; j < hc; ++j) {
        const int o = h * hc + j;
        float qp = __ldg(bq + o);
        for (int c = 0; c < C; ++c) qp = fmaf(__ldg(Wq + (size_t)o * C + c), qpix[(int64_t)c * HW], qp);
        cb = fmaf(qp, __ldg(bk + o), cb);
        for (int c = 0; c < C; ++c) kt[(size_t)c * T] = fmaf(qp, __ldg(Wk + (size_t)o * C + c), kt[(size_t)c * T]);
    }
    const float* kvp = kv + (int64_t)b * C * S * HW + pix;
    const int half = S / 2;
    float m = -INFINITY, sum = 0.0f;
    for (int s = 0; s < S; ++s) {
        const float* xs = kvp + (int64_t)s * HW;
        float l = cb;
        for (int c = 0; c < C; ++c) l = fmaf(kt[(size_t)c * T], __ldg(xs + (int64_t)c * S * HW), l);
        l *= scale;
        if (score0 != nullptr) {
            float* sp = (s < half ? score0 : score1) + (((int64_t)b * nH + h) * half + (s < half ? s : s - half)) * HW + pix;
            *sp = l;
        }
        if (l > m) {
            const float alpha = expf(m - l);   // exp(-inf) = 0 on the first sample
            sum *= alpha;
            for (int c = 0; c < C; ++c) yv[(size_t)c * T] *= alpha;
            m = l;
        }
        const float p = expf(l - m);
        sum += p;
        for (int c = 0; c < C; ++c) yv[(size_t)c * T] = fmaf(p, __ldg(xs + (int64_t)c * S * HW), yv[(size_t)c * T]);
    }
    const float inv = 1.0f / sum;
    float* op = out + (int64_t)b * C * HW + pix;
    for (int j = 0; j < hc; ++j) {
        const int o = h * hc + j;
        float acc = 0.0f;
        for (int c = 0; c < C; ++c) acc = fmaf(__ldg(Wv + (size_t)o * C + c), yv[(size_t)c * T], acc);
        op[(int64_t)o * HW] = fmaf(acc, inv, __ldg(bv + o));
    }
}

// in-place softmax over the sample axis of a (R, n, HW) tensor of raw logits
__global__ void __launch_bounds__(256) softmax_samples_kernel(float* __restrict__ sc, int64_t R, int n, int64_t HW) {
    const int64_t total = R * HW;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = idx / HW, pix = idx - r * HW;
        float* p = sc + r * n * HW + pix;
        float m = -INFINITY;
        for (int s = 0; s < n; ++s) m = fmaxf(m, p[(int64_t)s * HW]);
        float sum = 0.0f;
        for (int s = 0; s < n; ++s) sum += expf(p[(int64_t)s * HW] - m);
        const float inv = 1.0f / sum;
        for (int s = 0; s < n; ++s) p[(int64_t)s * HW] = expf(p[(int64_t)s * HW] - m) * inv;
    }
}

int launch_softmax_samples(float* sc, int64_t R, int n, int64_t HW, cudaStream_t st) {
    const int blocks = (int)min64(ceil_div(R * HW, 256), (int64_t)kNumSMs * 16);
    softmax_samples_kernel<<<blocks, 256, 0, st>>>(sc, R, n, HW);
    VFI_CHECK_LAUNCH("softmax_samples_kernel");
    return VFI_OK;
}

}  // namespace vfi

extern "C" int vfi_attn_f32(const float* q, const float* kv, const float* Wq, const float* bq, const float* Wk, const float* bk,
                            const float* Wv, const float* bv, float* out, float* score0, float* score1, int B, int C, int H,
                            int W, int S, int nH, void* stream) {
    using namespace vfi;
    if (B == 0) return VFI_OK;   // empty batch: nothing to do (an empty torch tensor has a NULL data pointer)
    VFI_REQUIRE(q && kv && Wq && bq && Wk && bk && Wv && bv && out, "vfi_attn_f32: null pointer");
    VFI_REQUIRE(B >= 0 && C >= 1 && H >= 1 && W >= 1 && S >= 1 && nH >= 1, "vfi_attn_f32: bad shape");
    VFI_REQUIRE(C % nH == 0, "vfi_attn_f32: C=%d not divisible by nH=%d", C, nH);
    VFI_REQUIRE((score0 == nullptr) == (score1 == nullptr), "vfi_attn_f32: score0/score1 must both be given or both NULL");
    if (score0) VFI_REQUIRE(S % 2 == 0, "vfi_attn_f32: per-frame scores need an even S");
    VFI_REQUIRE(nH <= 32, "vfi_attn_f32: nH=%d > 32", nH);
    if (B == 0) return VFI_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t HW = (int64_t)H * W;
    const int T = kAttnPix * nH;
    const size_t smem = (size_t)2 * C * T * sizeof(float);
    VFI_REQUIRE(smem <= 200 * 1024, "vfi_attn_f32: C=%d x nH=%d needs %zu B of shared memory", C, nH, smem);
    VFI_CUDA(cudaFuncSetAttribute(attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)ceil_div(HW, kAttnPix), (unsigned)B);
    attn_kernel<<<grid, dim3(kAttnPix, nH), smem, st>>>(q, kv, Wq, bq, Wk, bk, Wv, bv, out, score0, score1, C, HW, S, nH);
    VFI_CHECK_LAUNCH("attn_kernel");
    if (score0) {
        int rc = launch_softmax_samples(score0, (int64_t)B * nH, S / 2, HW, st);
        if (rc) return rc;
        rc = launch_softmax_samples(score1, (int64_t)B * nH, S / 2, HW, st);
        if (rc) return rc;
    }
    return VFI_OK;
}
