/* vfi_b200 -- C ABI of the B200-native flow-warping / deformable-sampling library (libvfi_b200.so).
 *
 * The reference (Euiyeon-Kim/VideoFrameInterpolation) has no FFI layer: its boundary for this path is a
 * set of Python functions plus three library calls.  Each entry point below names the
 * reference interface it stands behind (file:line relative to the reference checkout).  The Python
 * drop-ins in videoframeinterpolation_b200/ bind these with ctypes; INTEGRATION.md shows the stub a
 * reference maintainer would add.
 *
 * Conventions
 *   - every function returns 0 (VFI_OK) or a vfi_status / cudaError-derived code; it never throws,
 *     never allocates or frees caller memory, never synchronises the device or the stream;
 *   - all data pointers are DEVICE pointers to fp32 unless said otherwise; strides are in ELEMENTS;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - scratch memory is supplied by the caller (`workspace`), sized by the matching *_workspace_bytes;
 *   - safe to capture in a CUDA graph;
 *   - `frame_ext` (vfi_bwarp_f32, vfi_dsample_f32, vfi_dattn_f32, vfi_resize2x_f32): NULL when the
 *     tensors are whole frames.  For one tile of a spatially split frame (BASELINE configs[4]) it points to four HOST
 *     ints {Wg, Hg, x0, y0}: the frame is Wg x Hg pixels at this tensor's resolution and the (H, W) tensors are the
 *     crop whose top-left pixel is (x0, y0).  lin_x / lin_y are then the slices [x0, x0+W) / [y0, y0+H) of the frame's
 *     linspace vectors, and normalisation, border clamp and zero padding are the FRAME's: a crop samples exactly where
 *     the whole frame would (modules/warp.py:9-14, modules/deformable_attn.py:91-99 on the whole frame).  Taps the
 *     crop does not hold are clamped to the crop (bwarp) or read as zero (samplers): the caller's halo must cover them.
 *     vfi_dcn_f32 needs no such argument: torchvision's offsets are absolute pixels, nothing is normalised by the frame.
 */
#ifndef VFI_B200_H_
#define VFI_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VFI_B200_VERSION 202 /* major*10000 + minor*100 + patch */

typedef enum vfi_status {
    VFI_OK = 0,
    VFI_ERR_INVALID_ARGUMENT = 1,
    VFI_ERR_UNSUPPORTED = 2,
    VFI_ERR_CUDA = 3
} vfi_status;

/* weight applied to each source pixel before it is splatted (modules/warp.py:26-33, 69-71) */
typedef enum vfi_splat_weight {
    VFI_SPLAT_W_NONE = 0,     /* 'sum'   : w = 1, no weight channel                                   */
    VFI_SPLAT_W_ONES = 1,     /* 'avg'   : w = 1, weight channel splatted                             */
    VFI_SPLAT_W_LINEAR = 2,   /* 'linear': w = metric                                                 */
    VFI_SPLAT_W_SOFT = 3,     /* 'soft'  : w = exp(metric)                                            */
    VFI_SPLAT_W_SOFT_CLIP_T = 4 /* two-frame / M2M: w = t[b] * exp(clip(metric, -20, 20))             */
} vfi_splat_weight;

/* how the splatted weight channel becomes the divisor (modules/warp.py:39-56, 76-83) */
typedef enum vfi_splat_eps {
    VFI_EPS_ADD = 0,  /* den + eps_value (default '' and '-addeps': eps_value = 1e-7)                   */
    VFI_EPS_ZERO = 1, /* '-zeroeps': den == 0 -> 1                                                      */
    VFI_EPS_CLIP = 2  /* '-clipeps': max(den, eps_value)                                                */
} vfi_splat_eps;

int vfi_version(void);
/* Human-readable description of the last non-zero status returned on the calling thread. */
const char* vfi_last_error_string(void);
/* 1 if the current device is compute capability 10.x (the only target this library is built for). */
int vfi_device_supported(void);

/* ---- a1: bwarp(img, flow) -- modules/warp.py:7-15 -------------------------------------------------
 * out[b,c,y,x] = bilinear(img[b,c], clamp(ix,0,W-1), clamp(iy,0,H-1)) where ix, iy replay the reference's
 * fp32 normalise -> F.grid_sample(align_corners=True, padding_mode='border') unnormalise round trip:
 *   gx = lin_x[x] + flow_x / ((W-1)/2);  ix = ((gx + 1) / 2) * (W-1)          (same for y)
 * lin_x (W floats) / lin_y (H floats) are torch.linspace(-1, 1, W/H) exactly as the reference builds them.
 * img/out may be NCHW-contiguous or channels-last (any strides with stride_w == 1 or stride_c == 1 run
 * the vectorised paths; other stridings run the generic path).  flow is (B,2,H,W) with arbitrary strides.
 */
int vfi_bwarp_f32(const float* img, const float* flow, const float* lin_x, const float* lin_y, float* out,
                  int B, int C, int H, int W,
                  const int64_t img_strides[4], const int64_t flow_strides[4], const int64_t out_strides[4],
                  const int* frame_ext, void* stream);

/* a1 + the channel concat of modules/deformable_attn.py:69 (`torch.cat((feat_t, bwarp(feat_x, ftx), ftx), dim=1)`, the input
 * of the movement extractor): channels-last only; writes the C warped channels of every pixel AND (flow_x, flow_y, 0, 0)
 * into channels [C, C + 4) of the same output pixel (out_strides[3] >= C + 4), i.e. warp, concat and the zero padding of
 * the 4-aligned NHWC layout in one pass.  VFI_ERR_UNSUPPORTED for other layouts (use vfi_bwarp_f32 + copies). */
int vfi_bwarp_flowcat_f32(const float* img, const float* flow, const float* lin_x, const float* lin_y, float* out,
                          int B, int C, int H, int W, const int64_t img_strides[4], const int64_t flow_strides[4],
                          const int64_t out_strides[4], const int* frame_ext, void* stream);

/* ---- a2-a5: forward splatting -- modules/softsplat.py:237-360, modules/warp.py:18-114 -------------
 * One call = pre-scale (weight mode) + summation splat of `in` (B,C,H,W) and of the weight channel.
 *   num (B,C,H,W) += splat(in * w);   den (B,1,H,W) += splat(w)   (den may be NULL for VFI_SPLAT_W_NONE)
 * accumulate == 0: num/den are zero-filled by this call first; accumulate != 0: added into (two-frame and
 * multi-branch callers, modules/warp.py:76-80, 106-111).  metric (B,1,H,W) or NULL; t (B floats) or NULL.
 * Non-finite targets are skipped and corners are bounds-tested one by one, as the reference kernel does
 * (modules/softsplat.py:306-307, 325-339).  All tensors contiguous NCHW here; the Python wrapper makes
 * non-contiguous views contiguous (the reference bakes strides into its NVRTC source instead).
 * workspace == NULL: scatter form (fp32 atomics).  A 16-byte aligned workspace of vfi_splat_workspace_bytes selects
 * the gather form of vfi_splat_fused_f32 below with the sums left un-normalised: a link pass, then one owner thread
 * per target pixel stores (accumulate == 0) or adds (accumulate != 0) this source's sums to num / den -- no
 * floating-point atomics and no zero fill; used by the two-frame / multi-branch callers from 8 channels up.
 */
int vfi_splat_f32(const float* in, const float* flow, const float* metric, const float* t,
                  float* num, float* den, int B, int C, int H, int W,
                  int weight_mode, int accumulate, void* workspace, size_t workspace_bytes, void* stream);
/* Bytes of scratch vfi_splat_fused_f32 needs for this shape (68 B per source pixel: 4 claim slots + 4 slot weights
 * + list head + 4 overflow nodes; independent of C).  vfi_splat_f32's scatter form needs none (workspace may be NULL). */
size_t vfi_splat_workspace_bytes(int B, int C, int H, int W);

/* out = num / f(den) per eps_mode; blank_mask (B,1,H,W) uint8, optional: f(den) < blank_threshold
 * (modules/warp.py:83,114 use 1e-5).  `out` may alias `num`. */
int vfi_splat_normalize_f32(const float* num, const float* den, float* out, uint8_t* blank_mask,
                            int B, int C, int H, int W, int eps_mode, float eps_value, float blank_threshold,
                            void* stream);

/* Last source of an accumulate sequence, gather route only (modules/warp.py:76-83, 106-114 in one pass less):
 *   num_out = (num_out + splat(in * w)) / f(den + splat(w)),  blank_mask = f(.) < blank_threshold (optional)
 * den (the sum of the earlier sources' weights) is read, not updated.  Needs the workspace of vfi_splat_workspace_bytes;
 * VFI_UNSUPPORTED otherwise (callers then use vfi_splat_f32(accumulate) + vfi_splat_normalize_f32). */
int vfi_splat_last_f32(const float* in, const float* flow, const float* metric, const float* t, float* num_out, const float* den,
                       uint8_t* blank_mask, int B, int C, int H, int W, int weight_mode, int eps_mode, float eps_value,
                       float blank_threshold, void* workspace, size_t workspace_bytes, void* stream);

/* Fused single-call form of softsplat(tenIn, tenFlow, tenMetric, strMode) for sum/avg/linear/soft in GATHER form:
 * a link pass records, per target pixel, which source pixels hit it (integer CAS / exchange only), then one thread
 * per target pixel sums its sources for every plane and divides in the same pass -- no floating-point atomics,
 * no zero-fill / normalise passes over HBM, bitwise reproducible.  Same results as vfi_splat_f32 +
 * vfi_splat_normalize_f32 up to fp32 summation order.  workspace: vfi_splat_workspace_bytes. */
int vfi_splat_fused_f32(const float* in, const float* flow, const float* metric, float* out,
                        int B, int C, int H, int W, int weight_mode, int eps_mode, float eps_value,
                        void* workspace, size_t workspace_bytes, void* stream);

/* ---- a8 / a10: grouped multi-point sampler -- modules/deformable_attn.py:88-101, modules/madat.py:161-174
 * feat (B,C,H,W) viewed as (B*nG, C/nG, H, W); coords (B, nG*nS, 2, H, W) = absolute displacement in pixels
 * (channel 0 = x, 1 = y).  out (B, C, nS, H*W).  grid_out (B, nG*nS, H*W, 2) optional (madat.py:174).
 * Coordinates replay the reference: gx = lin_x[x] + (dx / (W-1)) / 2   (sic, deformable_attn.py:94),
 * gy = lin_y[y] + dy / ((H-1)/2), F.grid_sample(align_corners=True, padding_mode='zeros').
 */
int vfi_dsample_f32(const float* feat, const float* coords, const float* lin_x, const float* lin_y,
                    float* out, float* grid_out, int B, int C, int H, int W, int nG, int nS, const int* frame_ext,
                    void* stream);

/* ---- a9 / a11: Attn.forward(q, kv) on a materialised kv -- modules/deformable_attn.py:26-40, madat.py:86-115
 * q (B,C,H,W) un-projected; kv (B,C,S,H*W); weights (C,C) row-major [out][in], biases (C).
 * out (B,C,H,W).  score0/score1 (B*nH, S/2, H*W) optional = softmax over each half of the scaled logits.
 */
int vfi_attn_f32(const float* q, const float* kv, const float* Wq, const float* bq, const float* Wk,
                 const float* bk, const float* Wv, const float* bv, float* out, float* score0, float* score1,
                 int B, int C, int H, int W, int S, int nH, void* stream);

/* ---- a8+a9 fused (both frames): sampler + projections + softmax + weighted sum, kv never materialised.
 * offs0/offs1 (B, nG*nS*2, H, W): if coords_final == 0 these are the RAW conv_res_offset outputs and the
 * kernel applies offset_scale * tanh(.) + ft (modules/deformable_attn.py:81-86); if coords_final != 0 they
 * are already absolute displacements and ft0/ft1 are ignored.  ft0/ft1 (B,2,H,W) with strides ft_strides.
 * offs_bias (nG*nS*2 floats or NULL) is added to the raw offsets first (the bias of conv_res_offset, so the conv can
 * run bias-free); offs_channels_last != 0: offs0/offs1 are laid out (B, H, W, nG*nS*2) as a channels-last conv writes them.
 * q_channels_last / out_channels_last != 0: q is read / out is written as (B, H, W, C).
 * feat_channels_last == 0: feat0 / feat1 are planar (B,C,H,W) and every tap is gathered from global memory (csrc/dattn.cu).
 * feat_channels_last != 0: feat0 / feat1 are (B, H, W, C) as the encoder convolutions write them and the kernel stages,
 *   per 8x4 pixel tile and frame, the window all samples fall into in shared memory with one TMA box load
 *   (csrc/dattn_win.cu; the reach of the residual offsets is bounded by offset_scale); covered shapes:
 *   vfi_dattn_win_covers (C = 72, nG = 8, raw channels-last offsets, nS % 4 == 0, offset_scale <= 16) -- others return
 *   VFI_ERR_UNSUPPORTED and the caller passes planar features.
 * Requires nG == nH (true for every block the reference instantiates; madat.py:121 asserts it).
 */
int vfi_dattn_win_covers(int C, int nG, int nS, float offset_scale, int coords_final, int offs_channels_last);
int vfi_dattn_f32(const float* q, const float* feat0, const float* feat1, const float* offs0, const float* offs1,
                  const float* ft0, const float* ft1, const int64_t ft_strides[4],
                  const float* Wq, const float* bq, const float* Wk, const float* bk, const float* Wv,
                  const float* bv, const float* lin_x, const float* lin_y, float* out, float* score0,
                  float* score1, int B, int C, int H, int W, int nG, int nS, int nH, float offset_scale,
                  int coords_final, const float* offs_bias, int offs_channels_last, int q_channels_last,
                  int out_channels_last, int feat_channels_last, const int* frame_ext, void* stream);

/* ---- a6 / a7: modulated deformable convolution -- torchvision.ops.deform_conv2d semantics
 * (call sites modules/dcnv2.py:60-66, models/archive/DCNDAT.py:71-77).
 * x (B,Cin,H,W); offset (B, OG*2*kh*kw, Ho, Wo) interleaved (dy,dx); mask (B, OG*kh*kw, Ho, Wo) or NULL;
 * weight (Cout, Cin/WG, kh, kw); bias (Cout) or NULL; out (B,Cout,Ho,Wo).  No im2col buffer in HBM.
 * algo: 0 = auto, 1 = SIMT fp32 FFMA with global gathers, 2 = tcgen05 3xTF32 (3x3/s1/p1/d1, Cin/WG in {8,9,16}),
 *       3 = SIMT fp32 FFMA on a TMA-staged shared-memory window (<= 3x3 taps, stride 1, dilation 1, W % 4 == 0,
 *       <= 16 channels per group); taps outside the window are gathered cooperatively through a per-tile list.
 *       auto picks 3 where it applies and the plane has >= 4096 pixels, else 1 (2 only for the widths it is faster at).
 */
int vfi_dcn_f32(const float* x, const float* offset, const float* mask, const float* weight, const float* bias,
                float* out, int B, int Cin, int Cout, int H, int W, int kh, int kw, int stride, int pad, int dil,
                int n_weight_grps, int n_offset_grps, int algo, void* workspace, size_t workspace_bytes,
                void* stream);
size_t vfi_dcn_workspace_bytes(int B, int Cin, int Cout, int H, int W, int kh, int kw, int n_weight_grps,
                               int n_offset_grps, int algo);

/* ---- conv epilogue (the PReLU / residual tails of modules/basic_blocks.py:13-73,
 * modules/cnn_encoders.py:6-42, modules/generator.py:7-24 around the cuDNN convolutions).
 *   t = y + bias[c];  residual_mode 1: t = residual + t;  slope != NULL: t = t > 0 ? t : slope[c] * t;
 *   residual_mode 2: t = residual + t.   y / residual / out are (B,C,H,W) contiguous NCHW (channels_last = 0) or
 * channels-last (channels_last = 1) and share one layout; out may alias y.  bias / slope (C) may be NULL.
 * Same fp32 operation order as the separate PyTorch ops (bit-identical to them). */
int vfi_bias_act_f32(const float* y, const float* bias, const float* slope, const float* residual, float* out,
                     int B, int C, int H, int W, int channels_last, int residual_mode, void* stream);

/* ---- layout change between the conv stacks (channels-last) and the gather kernels (planar): x is (B,H,W,C) with a
 * pixel stride of x_pix_stride >= C floats (a channel slice of a wider channels-last tensor is fine), out (B,C,H,W)
 * contiguous.  Replaces the `.contiguous()` PyTorch inserts when models/DAT.py:86-101 hands encoder features
 * (modules/cnn_encoders.py:6-42) to modules/deformable_attn.py:91-101. */
int vfi_nhwc_to_nchw_f32(const float* x, float* out, int B, int C, int H, int W, int64_t x_pix_stride, void* stream);

/* ---- 3x3 / stride 1 / pad 1 convolution on the tensor cores (nn.Conv2d(k=3, s=1, p=1) in
 * modules/basic_blocks.py:13-73, modules/cnn_encoders.py:6-42, modules/generator.py:7-24, and conv_res_offset of
 * modules/deformable_attn.py:56-64) with the epilogue of vfi_bias_act_f32 fused in.  TF32 inputs (weights rounded
 * to nearest, activations truncated by the tensor core), fp32 accumulation: the arithmetic class of PyTorch's default
 * cudnn.allow_tf32 = True.  Callers that need fp32 convolutions keep cuDNN.
 *   x        NHWC fp32: pixel (b,y,x) at x + ((b*H + y)*W + x) * x_pix_stride, Cin channels from there
 *            (a channel slice of a wider tensor is x_pix_stride > Cin); Cin % 4 == 0, x_pix_stride % 4 == 0.
 *   wprep    vfi_conv3x3_prep_f32(weight (Cout,Cin,3,3) contiguous) -> vfi_conv3x3_prep_bytes(Cout, Cin) bytes.
 *   out      out_channels_last = 1: NHWC with out_c_total channels per pixel; 0: NCHW (B, out_c_total, H, W);
 *            channels [0, Cout) are written.  residual (mode 1: before the activation, 2: after) has the layout of
 *            out with res_c_total channels.  bias / slope (Cout) may be NULL.  residual_mode may carry, in bits 8 and up,
 *            the number of images Br of the residual when it is shared by groups of the batch (B % Br == 0): output
 *            image b then adds residual image b % Br (0 = one residual image per output image).
 * vfi_conv3x3_supported: 1 when this (Cin, x_pix_stride, Cout) runs here (weights of a Cout tile must fit in shared
 * memory: Cin <= 72 for a 72-wide tile).
 * upscale = 2 fuses nn.PixelShuffle(2) behind the convolution (modules/generator.py:14-22: upconv1 -> pixel_shuffle ->
 * prelu1): out is (B, Cout/4, 2H, 2W), out[b, c, 2y+i, 2x+j] = act(conv[b, 4c + 2i + j, y, x] + bias[4c + 2i + j]) with
 * slope indexed by c (Cout/4 values); no residual.  wprep depends on upscale (same value in prep and conv calls). */
int vfi_conv3x3_supported(int Cin, int x_pix_stride, int Cout, int upscale);
size_t vfi_conv3x3_prep_bytes(int Cout, int Cin, int upscale);
int vfi_conv3x3_prep_f32(const float* weight, float* wprep, int Cout, int Cin, int upscale, void* stream);
int vfi_conv3x3_f32(const float* x, const float* wprep, const float* bias, const float* slope, const float* residual,
                    float* out, int B, int H, int W, int Cin, int x_pix_stride, int Cout, int out_channels_last,
                    int out_c_total, int res_c_total, int residual_mode, int upscale, void* stream);

/* ---- a13 / f1: backward passes of the two warps (what train.py reaches through autograd) ----------------------
 * vfi_splat_bwd_f32: gradients of the summation splat out = splat(in, flow) (softsplat_func.backward,
 * modules/softsplat.py:364-533): grad_in (B,C,H,W) and grad_flow (B,2,H,W), either may be NULL.  Non-finite targets
 * get zero gradients and corners are bounds-tested one by one, as in the forward kernel.
 * vfi_bwarp_bwd_f32: gradients of bwarp (modules/warp.py:7-15) = autograd of F.grid_sample(bilinear, border,
 * align_corners=True) through the flow -> grid arithmetic: grad_img (zero-filled by the call, then accumulated with
 * atomics) and grad_flow; either may be NULL.  All tensors NCHW-contiguous fp32. */
int vfi_splat_bwd_f32(const float* in, const float* flow, const float* grad_out, float* grad_in, float* grad_flow,
                      int B, int C, int H, int W, void* stream);
int vfi_bwarp_bwd_f32(const float* img, const float* flow, const float* grad_out, const float* lin_x, const float* lin_y,
                      float* grad_img, float* grad_flow, int B, int C, int H, int W, void* stream);

/* ---- f1: backward of the grouped sampler (vfi_dsample_f32) -- the gradients torch autograd derives for
 * modules/deformable_attn.py:88-101 (F.grid_sample bilinear / zeros / align_corners=True composed with the
 * reference's coordinate arithmetic, x displacement scaled by 1/(2(W-1)) as at line 94).
 * grad_out (B,C,nS,H*W); grad_feat (B,C,H,W) (zero-filled here, scattered with atomics) and/or grad_coords
 * (B,nG*nS,2,H,W); either may be NULL. */
int vfi_dsample_bwd_f32(const float* feat, const float* coords, const float* grad_out, const float* lin_x,
                        const float* lin_y, float* grad_feat, float* grad_coords, int B, int C, int H, int W,
                        int nG, int nS, void* stream);

/* ---- a8 / f1 over a window of OUTPUT rows: the sampler and its backward for the rows [row0, row0 + rows) of the
 * (H, W) output grid, sampling the WHOLE feature map.  coords (B, nG*nS, 2, rows, W), out / grad_out (B, C, nS, rows*W),
 * grad_coords (B, nG*nS, 2, rows, W) hold those rows only; feat / grad_feat are (B, C, H, W); lin_y is the frame's
 * whole vector.  accumulate != 0: grad_feat is added to, not zero-filled first.
 * The reference has no counterpart: these are the pieces of the memory-bounded backward of the fused attention
 * (videoframeinterpolation_b200/deformable_attn.py:_FusedAttnTrainFn), which recomputes the samples of
 * modules/deformable_attn.py:88-101 strip by strip instead of keeping the (B, C, 2 nS, H*W) kv tensor of
 * modules/deformable_attn.py:33-34 (9.6 GB per 1080p pair at 1/2 resolution) alive for autograd. */
int vfi_dsample_rows_f32(const float* feat, const float* coords, const float* lin_x, const float* lin_y, float* out,
                         float* grid_out, int B, int C, int H, int W, int nG, int nS, int row0, int rows,
                         const int* frame_ext, void* stream);
int vfi_dsample_rows_bwd_f32(const float* feat, const float* coords, const float* grad_out, const float* lin_x,
                             const float* lin_y, float* grad_feat, float* grad_coords, int B, int C, int H, int W,
                             int nG, int nS, int row0, int rows, int accumulate, void* stream);

/* ---- f1: backward of the modulated deformable convolution -- torchvision deform_conv2d autograd semantics
 * (modules/dcnv2.py:60-66, models/archive/DCNDAT.py:71-77).  Two data-dependent steps; the dense contractions between
 * them (grad_weight = grad_out . columns^T and grad_columns = W^T . grad_out per weight group) are library GEMMs in the
 * Python wrapper (training path only).
 *   columns / grad_columns (B, Cin, kh*kw, Ho*Wo);  grad_x (B,Cin,H,W) zero-filled here, scattered with atomics;
 *   grad_offset (B, OG*2*kh*kw, Ho, Wo);  grad_mask (B, OG*kh*kw, Ho, Wo);  any gradient pointer may be NULL. */
int vfi_dcn_im2col_f32(const float* x, const float* offset, const float* mask, float* columns, int B, int Cin, int H,
                       int W, int kh, int kw, int stride, int pad, int dil, int n_offset_grps, void* stream);
int vfi_dcn_col2im_f32(const float* x, const float* offset, const float* mask, const float* grad_columns,
                       float* grad_x, float* grad_offset, float* grad_mask, int B, int Cin, int H, int W, int kh,
                       int kw, int stride, int pad, int dil, int n_offset_grps, void* stream);

/* ---- f3: frame I/O edge of the forward -- evaluate.py:138-142,187-191 (uint8 -> float / 255, pad, unpad),
 * data/__init__.py:6-20 (InputPadder, replicate), models/BaseModel.py:14-17 (norm_w_rgb_mean over both padded frames),
 * modules/generator.py:24 (clamp(out + mean, 0, 1)).
 * vfi_frame_prep: x0, x1 (B,3,H,W) contiguous, fp32 in [0,1] or uint8 (is_u8) -> out0, out1 (B,Hp,Wp,4) channels-last
 *   fp32 = replicate-padded frame minus mean[b], 4th channel 0; mean (B).  workspace: vfi_frame_prep_workspace_bytes.
 * vfi_frame_finish: y (B,3,Hp,Wp) with element strides y_strides -> out (B,3,H,W) contiguous =
 *   clamp(y + mean[b], 0, 1) cropped at (pad_top, pad_left); out_is_u8: round(255 * .) as uint8. */
size_t vfi_frame_prep_workspace_bytes(int B);
int vfi_frame_prep(const void* x0, const void* x1, int is_u8, float* out0, float* out1, float* mean, int B, int H, int W,
                   int Hp, int Wp, int pad_top, int pad_left, void* workspace, size_t workspace_bytes, void* stream);
int vfi_frame_finish(const float* y, const int64_t y_strides[4], const float* mean, void* out, int out_is_u8, int B, int H,
                     int W, int Hp, int Wp, int pad_top, int pad_left, void* stream);

/* ---- flow upsampling between pyramid levels: out (B,C,2H,2W) contiguous = res + mul * resize(x, 2)
 * Replaces: BaseModel.resize (models/BaseModel.py:19-22, F.interpolate bilinear / align_corners=True) followed by
 * `* 2.0` and `conv_res_flow(..) +` at modules/deformable_attn.py:127-131 -- three elementwise passes in one, ATen's
 * upsample_bilinear2d arithmetic.  x (B,C,H,W) / res (B,C,2H,2W, optional) with element strides.  frame_ext refers
 * to x's level: align_corners=True makes the source position depend on the frame size (see the conventions above). */
int vfi_resize2x_f32(const float* x, const int64_t x_strides[4], const float* res, const int64_t res_strides[4], float mul,
                     float* out, int B, int C, int H, int W, const int* frame_ext, void* stream);

/* Glue of the DCNDAT block in one pass.  Replaces: models/archive/DCNDAT.py:150-157 (_get_ref_coords_mask: chunk,
 * 2*tanh, view, + flow) and :69-70 (offset.permute(0,2,1,3,4).flip(1).contiguous().view, sigmoid(mask)).
 * raw (B,3n,H,W) with element strides raw_strides, n = nG*nS; ft (B,2,H,W) with element strides ft_strides.
 * coords (B,n,2,H,W), dcn_offset (B,2n,H,W) [all y then all x, the order line 69 produces], mask (B,n,H,W): contiguous. */
int vfi_dcndat_coords_f32(const float* raw, const int64_t raw_strides[4], const float* ft, const int64_t ft_strides[4],
                          float* coords, float* dcn_offset, float* mask, int B, int n, int H, int W, void* stream);

/* On-device validators.  Replaces: evaluate.py:15-26 (gaussian / create_window_3d),
 * evaluate.py:29-79 (calculate_ssim: 11x11x11 Gaussian window over (channel, y, x), replicate padding 5, five
 * conv3d calls) and evaluate.py:82-84 (calculate_psnr).
 * img1, img2 (B,C,H,W) contiguous fp32; val_range = L of evaluate.py:30-42 (1 for frames in [0,1]).
 * out (B,3) fp64 per sample: [0] sum of squared differences, [1] sum of the SSIM map, [2] sum of v1/v2 (the
 *   contrast-structure term): the means of the reference are these sums / (C*H*W), psnr = -10 log10(sse / n).
 * Deterministic (fixed summation order).  H, W >= 11, C <= 32.  workspace: vfi_frame_metrics_workspace_bytes. */
size_t vfi_frame_metrics_workspace_bytes(int B, int C, int H, int W);
int vfi_frame_metrics_f32(const float* img1, const float* img2, double* out, int B, int C, int H, int W, float val_range,
                          void* workspace, size_t workspace_bytes, void* stream);

/* Allocation-free variant of vfi_frame_metrics_f32.  It has NO counterpart in the reference: it was added
 * for callers that cannot allocate device memory (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  Same inputs, outputs and results, but the partial sums live in a 1 MiB file-scope
 * device array (vfi_frame_metrics_static_max_bytes) that every GPU loading the library holds until exit; the call itself
 * allocates and frees nothing.  Shapes whose vfi_frame_metrics_workspace_bytes exceed the array (e.g. more than 7
 * 1080p RGB frame pairs per call) are refused with VFI_ERR_UNSUPPORTED.  All calls on a device share the array: calls
 * that overlap on the GPU (different streams / host threads) race; calls on one stream are safe.  Nothing needs clearing
 * between calls (every partial read is written first). */
size_t vfi_frame_metrics_static_max_bytes(void);
int vfi_frame_metrics_static_f32(const float* img1, const float* img2, double* out, int B, int C, int H, int W,
                                 float val_range, void* stream);

/* Number of kernel launches issued by this library on the calling thread since the last reset
 * (bench.py reports it as gpu_launches). */
int64_t vfi_launch_count(void);
void vfi_launch_count_reset(void);

#ifdef __cplusplus
}
#endif
#endif /* VFI_B200_H_ */
