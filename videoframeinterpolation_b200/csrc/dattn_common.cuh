// Shared declarations of the fused deformable-attention kernels (dattn.cu: global-gather form, dattn_win.cu:
// shared-memory window form).
#pragma once
#include "sampler.cuh"

namespace vfi {

struct DattnArgs {
    const float *q, *feat[2], *offs[2], *ft[2];
    int64_t ft_sn, ft_sc, ft_sh, ft_sw;
    const float *Wq, *bq, *Wk, *bk, *Wv, *bv, *lin_x, *lin_y;
    const float* obias;      // optional bias of the offset head (nG*nS*2), added before tanh
    int offs_cl;             // offsets are channels-last: (B, H, W, nG*nS*2)
    int qo_cl;               // bit 0: q is channels-last (B, H, W, C); bit 1: out is written channels-last
    float *out, *score[2];
    int B, H, W, nS;
    float offset_scale;
    int coords_final;
    int tw_log2;              // tile = (1 << tw_log2) x (32 >> tw_log2) pixels per warp
    int tiles_x, tiles_y;
    int zero;                 // always 0 (see the sampling loop)
    FrameExt fe;              // frame extent + origin of this crop in it (whole frame: {W, H, 0, 0})
    int feat_cl;              // feat0 / feat1 are channels-last (B, H, W, C) -- window form only
};

constexpr int cmax(int a, int b) { return a > b ? a : b; }
constexpr int cmin(int a, int b) { return a < b ? a : b; }

template <int GC, int NG>
struct DattnCfg {
    static constexpr int C = GC * NG;
    static constexpr int C4 = (C + 3) & ~3;       // Wq row stride (16 B aligned rows: broadcast LDS.128)
    static constexpr int GCP = (GC + 3) & ~3;     // padded channel block
    static constexpr int SC = GC > 12 ? 2 : 4;    // slots per chunk (register budget: xs[SC][GC])
    static constexpr int THREADS = NG * 32;
    static constexpr int WQ = C * C4;
    static constexpr int WT = NG * C * GCP;       // Wk / Wv re-blocked [g][row][GCP]
    static constexpr int RED = SC * NG * NG * 32;
    static constexpr int EPI_CAP = 10240;         // floats (40 KB) for one epilogue round
    static constexpr int EPI_H_RAW = cmax(1, cmin(NG, EPI_CAP / (GC * NG * 32)));
    static constexpr int EPI_H = EPI_H_RAW >= 8 ? 8 : (EPI_H_RAW >= 4 ? 4 : (EPI_H_RAW >= 2 ? 2 : 1));
    static constexpr int EPI = EPI_H * GC * NG * 32;
    static constexpr int UNI = cmax(cmax(RED, 2 * C * 32), EPI);
    static constexpr int PB = NG * SC * 32;
    static constexpr int AB = NG * 32;
    static constexpr int SMEM_FLOATS = WQ + 2 * WT + UNI + PB + AB + 4;   // + rebase flags
    static constexpr int MIN_BLOCKS = NG <= 4 ? 2 : 1;
};

// window form (dattn_win.cu): VFI_OK when launched, VFI_ERR_UNSUPPORTED when the shape / layout is not covered
int launch_dattn_win(const DattnArgs& a, int C, int nG, cudaStream_t st);
bool dattn_win_covers(int C, int nG, int nS, float offset_scale, int coords_final, int offs_cl);

}  // namespace vfi
