// Standalone bisect of the TMA path (debug tool; not part of the library).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../videoframeinterpolation_b200/csrc/tma.cuh"
using namespace vfi;
namespace vfi { void set_error(const char* fmt, ...) { printf("ERR %s\n", fmt); } void note_launch() {} }
constexpr int RW = 96, RH = 32;
__global__ void k(const __grid_constant__ CUtensorMap tmap, float* out, int mode, int x0, int y0, int z) {
    extern __shared__ __align__(128) uint8_t raw[];
    uint8_t* al = raw + ((128u - (smem_u32(raw) & 127u)) & 127u);
    float* st = (float*)al;
    uint64_t* bar = (uint64_t*)(al + RW * RH * 4);
    if (threadIdx.x == 0) { mbar_init(bar, 1); fence_mbar_init(); }
    __syncthreads();
    if (mode >= 1 && threadIdx.x == 0) tma_prefetch_desc(&tmap);
    if (mode >= 2 && threadIdx.x == 0) { mbar_arrive_expect_tx(bar, RW * RH * 4); }
    if (mode >= 3 && threadIdx.x == 0) { tma_load_3d(st, &tmap, bar, x0, y0, z); }
    if (mode >= 3) { mbar_wait(bar, 0); }
    __syncthreads();
    if (mode >= 3) for (int i = threadIdx.x; i < RW * RH; i += blockDim.x) out[i] = st[i];
}
int main(int argc, char** argv) {
    int W = 480, H = 270, BC = 3;
    std::vector<float> h((size_t)W * H * BC);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (float)(i % 1000);
    float *d, *o;
    cudaMalloc(&d, h.size() * 4); cudaMalloc(&o, RW * RH * 4);
    cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    CUtensorMap tm;
    int rc = make_tmap_f32_3d(&tm, d, W, H, BC, (uint64_t)W * 4, (uint64_t)W * H * 4, RW, RH, 1);
    printf("encode rc=%d\n", rc);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, RW * RH * 4 + 256);
    int X0 = atoi(argv[1]), Y0 = atoi(argv[2]);
    for (int mode = 3; mode <= 3; ++mode) {
        for (int neg = 1; neg < 2; ++neg) {
            k<<<1, 128, RW * RH * 4 + 256>>>(tm, o, mode, X0, Y0, 1);
            cudaError_t e = cudaDeviceSynchronize();
            printf("mode %d neg %d: %s\n", mode, neg, cudaGetErrorString(e));
            if (e != cudaSuccess) return 1;
            if (mode == 3) {
                std::vector<float> r(RW * RH);
                cudaMemcpy(r.data(), o, RW * RH * 4, cudaMemcpyDeviceToHost);
                int bad = 0;
                for (int y = 0; y < RH; ++y) for (int x = 0; x < RW; ++x) {
                    int gx = x + X0, gy = y + Y0;
                    float exp = (gx < 0 || gy < 0 || gx >= W || gy >= H) ? 0.f : h[(size_t)1 * W * H + (size_t)gy * W + gx];
                    if (r[y * RW + x] != exp) ++bad;
                }
                printf("  mismatches: %d\n", bad);
            }
        }
    }
    return 0;
}
