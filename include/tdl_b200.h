/* tdl_b200 -- C ABI of the sm_100a native library (tensorflowdistributedlearning_b200/lib/libtdl_b200.so).
 *
 * Conventions
 *   - every function returns 0 on success; on failure the message is available from tdl_last_error()
 *   - pointers are DEVICE pointers unless a parameter name says host; `stream` is a cudaStream_t
 *   - activations are NHWC bf16, row-major, 16-byte aligned; C (channels) must be a multiple of 8
 *   - nothing here allocates device memory except tdl_symm_alloc (the symmetric heap)
 *
 * The reference (gf712/TensorflowDistributedLearning) has no native code; each group cites the reference
 * call site whose TensorFlow library kernel it replaces.
 */
#ifndef TDL_B200_H_
#define TDL_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- library ---- */
const char* tdl_last_error(void);
int tdl_version(void);
int tdl_num_sms(void);
uint32_t tdl_crc32c(const void* data, uint64_t n, uint32_t crc); /* host: CRC-32C for TF tensor-bundle checkpoints */

/* ---- convolution on tcgen05/TMEM/TMA (reference: slim.conv2d, core/resnet.py:79-85,130-144,164-166,449-492) ----
 * x [N,H,W,C], w [K, R*S, C] (OHWI), y [N,P,Q,K]; pad_h/pad_w are the top/left padding, P/Q the output size.
 * Optional fused epilogue: + bias[K] (fp32), + residual [N,P,Q,K] (bf16), ReLU. */
int tdl_conv2d_fprop_bf16(const void* x, const void* w, void* y, const float* bias, const void* residual, int N,
                          int H, int W, int C, int K, int R, int S, int stride, int dil, int pad_h, int pad_w, int P,
                          int Q, int relu, void* stream);
/* fprop that also emits per-CTA partial BatchNorm statistics of its bf16 output: stats_partial [tdl_num_sms()][2][K]
 * fp32 (written), *nparts_host = number of valid leading slots (host int); finish with tdl_bn_finalize_partials */
/* y = mask(conv(x, w) + residual): residual added on the tensor core, out_mask = [N*P*Q][K/8] bytes (1 bit per element,
 * layout of tdl_bn_apply's mask_out; may be NULL).  K % 32 == 0 when a mask is given. */
int tdl_conv2d_fprop_res_mask_bf16(const void* x, const void* w, void* y, const void* residual, const void* out_mask,
                                   int N, int H, int W, int C, int K, int R, int S, int stride, int dil, int pad_h,
                                   int pad_w, int P, int Q, void* stream);
int tdl_conv2d_set_pair_mode(int mode); /* 1 = CTA pairs (tcgen05 cta_group::2, M = 256 MMAs, half a weight tile per CTA) for Cout tiles of 128/256; also TDL_CONV_2CTA */
int tdl_conv2d_get_pair_mode(void);
int tdl_conv2d_set_halo_mode(int mode); /* experimental halo-reuse main loop for stride-1 RxS filters: 0 off (default), 1 on */
int tdl_conv2d_set_wgrad_pair_mode(int mode); /* weight gradient on CTA pairs (tcgen05 cta_group::2, 256 x 256 tile per pair) for Cout >= 256: 1 on, 0 off (default: faster alone, not inside the overlapped training step) */
int tdl_conv2d_set_wgrad_tuning(int wave_snap, float model_k); /* measurement hooks: whole-wave grids on/off, split-count model constant (< 0 = default 1.1) */
int tdl_conv2d_set_wgrad_bulk_mode(int mode); /* weight-gradient epilogue: 1 = bulk reductions (cp.reduce.async.bulk) from shared memory (default), 0 = per-thread red.global.add */
int tdl_conv2d_set_wgrad_row_mode(int mode); /* weight gradient of 64-channel 3-wide stride-1 filters: the taps of a filter row share one haloed input tile; 0 off, 1 on (default), 2 on with three narrow MMAs */
int tdl_conv2d_set_halo_res_mode(int mode); /* halo-reuse loop with shared-memory-resident weights where they fit (3x3 64->64): 1 on (default), 0 off */
int tdl_conv2d_fprop_stats_bf16(const void* x, const void* w, void* y, const float* bias, int N, int H, int W, int C,
                                int K, int R, int S, int stride, int dil, int pad_h, int pad_w, int P, int Q,
                                float* stats_partial, int* nparts_host, void* stream);
/* stride-1 convolution whose output pixel (p,q) is written to (p*y_step+y_off_h, q*y_step+y_off_w) of a
 * [N,y_full_h,y_full_w,K] tensor: one input-parity class of a strided convolution's data gradient per call */
int tdl_conv2d_fprop_strided_out_bf16(const void* x, const void* w, void* y, int N, int H, int W, int C, int K, int R,
                                      int S, int dil, int pad_h, int pad_w, int P, int Q, int y_full_h, int y_full_w,
                                      int y_step, int y_off_h, int y_off_w, void* stream);
/* same, with the R x S taps read out of a larger filter w [K, w_taps, C]: position (r,s) uses tap
 * tap_base + r*tap_sr + s*tap_ss (a parity class of a strided data gradient on the full flipped weight copy) */
int tdl_conv2d_fprop_strided_out_taps_bf16(const void* x, const void* w, void* y, int N, int H, int W, int C, int K,
                                           int R, int S, int w_taps, int tap_base, int tap_sr, int tap_ss, int dil,
                                           int pad_h, int pad_w, int P, int Q, int y_full_h, int y_full_w, int y_step,
                                           int y_off_h, int y_off_w, void* stream);
/* dw [K, R*S, C] fp32 is ACCUMULATED into (zero it first): dw += dy^T (*) x */
int tdl_conv2d_wgrad_bf16(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int K, int R, int S,
                          int stride, int dil, int pad_h, int pad_w, int P, int Q, void* stream);
/* fp32 master weights -> bf16 fprop copy [K,taps,C] and bf16 data-gradient copy [C,taps(flipped),K] (either may be NULL) */
int tdl_conv2d_weight_prep(const float* w, void* w_fprop, void* w_dgrad, int K, int taps, int C, void* stream);

/* ---- fused BatchNorm (+residual)(+ReLU) (reference: slim.batch_norm, core/resnet.py:71,126,242,373-387) ---- */
int64_t tdl_bn_scratch_floats(int64_t rows, int C);
int tdl_bn_fwd_stats(const void* x, int64_t rows, int C, float* scratch, int64_t scratch_floats, const float* gamma,
                     const float* beta, float* running_mean, float* running_var, float decay, float eps, float* mean,
                     float* invstd, float* scale, float* shift, void* stream);
/* No counterpart in the reference (an addition of this library).  Allocation-free variant of tdl_bn_fwd_stats
 * for callers that cannot supply scratch: the partial sums go to a fixed 1,048,576-float (4 MiB) array inside the
 * library (held by every GPU that loads it, until exit); larger shapes are refused with an error.  The first call
 * on a device looks up the array's device address (make it before any stream capture); all calls share the array,
 * so calls that can overlap on the GPU race.  Nothing in the array needs clearing between calls. */
int tdl_bn_fwd_stats_noalloc(const void* x, int64_t rows, int C, const float* gamma, const float* beta,
                             float* running_mean, float* running_var, float decay, float eps, float* mean,
                             float* invstd, float* scale, float* shift, void* stream);
int tdl_bn_finalize_partials(const float* partial, int nparts, int64_t rows, int C, const float* gamma,
                             const float* beta, float* running_mean, float* running_var, float decay, float eps,
                             float* mean, float* invstd, float* scale, float* shift, void* stream);
/* the same with mean_offset [C] (or NULL): a per-channel constant the producer left out of the tensor because this
 * normalisation removes it (a folded convolution bias); only the moving mean sees it */
int tdl_bn_finalize_partials_off(const float* partial, int nparts, int64_t rows, int C, const float* gamma,
                                 const float* beta, float* running_mean, float* running_var, float decay, float eps,
                                 float* mean, float* invstd, float* scale, float* shift, const float* mean_offset,
                                 void* stream);
int tdl_bn_infer_scale_shift(const float* gamma, const float* beta, const float* running_mean,
                             const float* running_var, float eps, float* scale, float* shift, int C, void* stream);
/* mask_out (may be NULL): [rows][C/8] bytes, one bit per element = "pre-ReLU value > 0".  The backward entry points
 * take either the activation z (bf16, z_is_bitmask = 0), that bit mask (z_is_bitmask = 1), or NULL (mask recomputed
 * from x, scale, shift -- only valid without a residual). */
int tdl_bn_apply(const void* x, const float* scale, const float* shift, const void* residual, void* y, void* mask_out,
                 int64_t rows, int C, int relu, void* stream);
int tdl_bn_bwd_reduce(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                      const float* shift, const float* mean, const float* invstd, int64_t rows, int C, int relu,
                      float* scratch, int64_t scratch_floats, float* dgamma, float* dbeta, float* acc_dgamma,
                      float* acc_dbeta, void* stream);
int tdl_bn_bwd_finalize_partials(const float* partial, int nparts, int C, float* dgamma, float* dbeta,
                                 float* acc_dgamma, float* acc_dbeta, void* stream);
/* leading-dimension / fused forms: y, dz may be channel slices of wider buffers (row pitch ld >= C);
 * tdl_bn_bwd_apply_ex also adds `addend` (a second gradient of the BatchNorm's input: the identity shortcut of a residual
 * unit) and, with mask_by_x, applies the ReLU backward of the input tensor itself (core/resnet.py:71-87,126-148) */
int tdl_bn_apply_ld(const void* x, const float* scale, const float* shift, const void* residual, void* y,
                    void* mask_out, int64_t rows, int C, int relu, int64_t ldy, void* stream);
int tdl_bn_bwd_reduce_ld(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                         const float* shift, const float* mean, const float* invstd, int64_t rows, int C, int relu,
                         int64_t ld_dz, float* scratch, int64_t scratch_floats, float* dgamma, float* dbeta,
                         float* acc_dgamma, float* acc_dbeta, void* stream);
int tdl_bn_bwd_apply_ld(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                        const float* shift, const float* mean, const float* invstd, const float* dgamma,
                        const float* dbeta, void* dx, void* dres, int64_t rows, int C, int relu, int64_t ld_dz,
                        void* stream);
int tdl_bn_bwd_apply_ex(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                        const float* shift, const float* mean, const float* invstd, const float* dgamma,
                        const float* dbeta, void* dx, void* dres, int64_t rows, int C, int relu, int64_t ld_dz,
                        const void* addend, int mask_by_x, void* stream);
/* one-launch forms with the statistics finalize folded in (thread-block clusters + DSMEM; measured slower, opt-in) and
 * the halves of the two-launch forms */
int tdl_bn_apply_fused(const void* x, const float* partial, int nparts, int64_t stat_rows, const float* gamma,
                       const float* beta, float* running_mean, float* running_var, float decay, float eps, float* mean,
                       float* invstd, float* scale, float* shift, const void* residual, void* y, void* mask_out,
                       int64_t rows, int C, int relu, void* stream);
int tdl_bn_partial_stats(const void* x, int64_t rows, int C, float* scratch, int64_t scratch_floats, int* nparts,
                         void* stream);
int tdl_bn_bwd_partial(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                       const float* shift, const float* mean, const float* invstd, int64_t rows, int C, int relu,
                       float* scratch, int64_t scratch_floats, int* nparts, void* stream);
int tdl_bn_bwd_partial_ld(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                          const float* shift, const float* mean, const float* invstd, int64_t rows, int C, int relu,
                          int64_t ld_dz, float* scratch, int64_t scratch_floats, int* nparts, void* stream);
int tdl_bn_bwd_apply_fused(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                           const float* shift, const float* mean, const float* invstd, const float* partial, int nparts,
                           float* dgamma, float* dbeta, float* acc_dgamma, float* acc_dbeta, void* dx, void* dres,
                           int64_t rows, int C, int relu, void* stream);
/* upsample into / out of a channel slice (row pitch ld >= C) */
int tdl_upsample_fwd_ld(const void* x, void* y, int N, int H, int W, int C, int OH, int OW, int64_t ldy, int is_bf16,
                        void* stream);
int tdl_upsample_bwd_ld(const void* dy, float* dx, int N, int H, int W, int C, int OH, int OW, int64_t ld_dy,
                        int is_bf16, void* stream);
int tdl_bn_bwd_apply(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                     const float* shift, const float* mean, const float* invstd, const float* dgamma,
                     const float* dbeta, void* dx, void* dres, int64_t rows, int C, int relu, void* stream);

/* ---- spatial ops (reference: core/resnet.py:241,246-248,464-465; core/layers.py:34-42,82-109) ---- */
/* global average pool folded into the 1x1 convolution that follows (the ASPP image-pooling branch, core/resnet.py:464-466):
 * y [N,K] bf16 = mean_hw(x [N,HW,C]) . w [K,C]^T, w = fp32 master weights; gap_out [N,C] fp32 is kept for the backward.
 * bwd: dx [N,HW,C] bf16 written, dw [K,C] fp32 accumulated into (either may be NULL) */
int tdl_gap_conv1x1_fwd(const void* x, const float* w, void* y, float* gap_out, int N, int HW, int C, int K, void* stream);
int tdl_gap_conv1x1_bwd(const void* dy, const float* w, const float* gap, void* dx, float* dw, int N, int HW, int C, int K,
                        void* stream);
/* 3x3 / stride-2 max-pool fused with the partial BatchNorm statistics of its output (the reference's pool1 ->
 * postnorm, core/resnet.py:241-242): stats_partial [*nparts][2][C], rows = tdl_maxpool_stats_rows(); 256 % (C/8) == 0 */
int tdl_maxpool_stats_rows(int N, int C, int P, int Q);
int tdl_maxpool_fwd_stats(const void* x, void* y, void* idx_u8, int N, int H, int W, int C, int P, int Q, int pad_h,
                          int pad_w, float* stats_partial, int* nparts, void* stream);
int tdl_maxpool_fwd(const void* x, void* y, void* idx_u8, int N, int H, int W, int C, int P, int Q, int k, int stride,
                    int pad_h, int pad_w, void* stream);
int tdl_maxpool_bwd(const void* dy, const void* idx_u8, void* dx, int N, int H, int W, int C, int P, int Q, int k,
                    int stride, int pad_h, int pad_w, void* stream);
int tdl_gap_fwd(const void* x, void* y, int N, int HW, int C, void* stream);
int tdl_gap_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream);
int tdl_depthwise_fwd(const void* x, const float* w, const float* bias, void* y, int N, int H, int W, int C, int P,
                      int Q, int R, int S, int stride, int dil, int pad_h, int pad_w, int relu, void* stream);
int tdl_depthwise_bwd(const void* x, const float* w, const void* dy, const void* y, void* dx, float* dw, float* dbias,
                      int N, int H, int W, int C, int P, int Q, int R, int S, int stride, int dil, int pad_h, int pad_w,
                      int relu, void* stream);
/* fast 3x3 depthwise (strip kernels): w [9,C] fp32; relu_in = ReLU on load, relu_out = ReLU on store;
 * mask (NULL ok) zeroes outputs where mask <= 0 -- the data gradient is this kernel on dy with flipped taps */
int tdl_depthwise3x3_fwd(const void* x, const float* w, const float* bias, const void* mask, void* y, int N, int H,
                         int W, int C, int P, int Q, int stride, int dil, int pad_h, int pad_w, int relu_in,
                         int relu_out, void* stream);
/* TMA-tiled variant of tdl_depthwise3x3_fwd (haloed input patch via one 4-D TMA box into swizzled smem);
 * C % 8 == 0, stride 1/2 with dilation 1, or stride 1 with dilation 2 */
int tdl_depthwise3x3_tma_fwd(const void* x, const float* w, const float* bias, const void* mask, void* y, int N,
                             int H, int W, int C, int P, int Q, int stride, int dil, int pad_h, int pad_w,
                             int relu_in, int relu_out, void* stream);
/* data gradient of a stride-1 3x3 depthwise layer (same map size in and out): the forward kernel on dy with the taps of
 * the layer's own [9][C] weights read in reverse order; mask (optional): zero dx where mask <= 0 (input ReLU);
 * addend (optional, dx-shaped): added after the mask (the gradient arriving through another consumer of the input) */
int tdl_depthwise3x3_tma_dgrad(const void* dy, const float* w, const void* mask, const void* addend, void* dx, int N,
                               int H, int W, int C, int dil, int pad_h, int pad_w, void* stream);
int tdl_depthwise3x3_tma_fwd_stats(const void* x, const float* w, const float* bias, void* y, int N, int H, int W, int C,
                                   int P, int Q, int stride, int dil, int pad_h, int pad_w, int relu_in, int relu_out,
                                   float* stats_partial /* [2*tdl_num_sms()][2][C] */, int* nparts_host, void* stream);
int tdl_depthwise3x3_dgrad_s2(const void* dy, const float* w, const void* xmask, void* dx, int N, int H, int W, int C,
                              int P, int Q, int pad_h, int pad_w, void* stream);
int tdl_depthwise3x3_tma_wgrad(const void* x, const void* dy, float* dw, float* dbias, int N, int H, int W, int C,
                               int P, int Q, int stride, int dil, int pad_h, int pad_w, int relu_in, void* stream);
int tdl_depthwise3x3_wgrad(const void* x, const void* dy, const void* ymask, float* dw, float* dbias, int N, int H,
                           int W, int C, int P, int Q, int stride, int dil, int pad_h, int pad_w, int relu_in,
                           void* stream);
/* First layer: stride-2 RxS convolution on 1..4 input channels (7x7 / 3x3 roots) on the tensor cores.  The image is
 * packed space-to-depth into two 8-channel planes [N][ceil(H/2)][2][ceil(W/2)][8]; the weights into [K][taps*16]
 * (tdl_conv_stem_weight_elems gives the element count).  K <= 64.  Optional fused BatchNorm partial statistics as in
 * tdl_conv2d_fprop_stats_bf16 (pass both stats pointers or neither). */
/* Every layer's weight re-layout in one launch.  descs: device array of nlayers 48-byte records
 * {const float* w; bf16* wf; bf16* wd (NULL = no data-gradient copy); int K, taps, C, ceil(C/64), ceil(K/64), 0} (C, K even; w 8-byte aligned);
 * block_start: device int[nlayers], prefix sums of ceil(C/64) * ceil(K/64) * taps; total_blocks: the full sum. */
int tdl_conv2d_weight_prep_batched(const void* descs, const int* block_start, int nlayers, int total_blocks,
                                   void* stream);
int tdl_conv_stem_pack(const void* x, void* planes, int N, int H, int W, int C, void* stream);
int tdl_conv_stem_weight_elems(int K, int R, int S, int pad_h, int pad_w);
int tdl_conv_stem_weight_prep(const float* w, void* w_packed, int K, int R, int S, int C, int pad_h, int pad_w,
                              void* stream);
int tdl_conv_stem_fprop(const void* planes, const void* w_packed, void* y, int N, int H, int W, int K, int R, int S,
                        int pad_h, int pad_w, int P, int Q, float* stats_partial, int* nparts_host, void* stream);
/* First-layer weight gradient from the packed image (tcgen05; one output row per GEMM K slab).  dwp_scratch holds
 * tdl_conv_stem_weight_elems() floats; dw is fp32 [K,R,S,C], overwritten or (accumulate != 0) added to.
 * One MMA per (row tap, plane) with overlapping column-tap groups; variant 0: M = 128, 1: M = 64. */
int tdl_conv_stem_wgrad(const void* planes, const void* dy, float* dwp_scratch, float* dw, int N, int H, int W, int C,
                        int K, int R, int S, int pad_h, int pad_w, int P, int Q, int accumulate, int variant,
                        void* stream);
/* Stride-1 data gradient that also reduces the BatchNorm-backward sums (sum dz, sum dz * xhat) of the tensor it
 * produces, in its epilogue; residual / out_mask optional (as tdl_conv2d_fprop_res_mask_bf16).  stats_partial is
 * [tdl_num_sms()][2][K]; feed it to tdl_bn_bwd_finalize_partials.  bn_scale / bn_shift (optional, not together with
 * out_mask): the sums see the ReLU mask fmaf(x, scale, shift) > 0, the stored gradient stays unmasked. */
int tdl_conv2d_dgrad_bnstats_bf16(const void* x, const void* w, void* y, const void* residual, const void* out_mask,
                                  const void* bn_x, const float* bn_mean, const float* bn_invstd, const float* bn_scale,
                                  const float* bn_shift, int N, int H, int W, int C, int K, int R, int S, int dil, int pad_h, int pad_w, int P, int Q,
                                  float* stats_partial, int* nparts_host, void* stream);
/* relu?(conv + bias? + residual) with the residual on the tensor core and BatchNorm partial statistics of the
 * stored output (stats_partial / nparts_host as in tdl_conv2d_fprop_stats_bf16). */
int tdl_conv2d_fprop_res_stats_bf16(const void* x, const void* w, void* y, const float* bias, const void* residual,
                                    int N, int H, int W, int C, int K, int R, int S, int stride, int dil, int pad_h,
                                    int pad_w, int P, int Q, int relu, float* stats_partial, int* nparts_host,
                                    void* stream);
/* Convolution with 1..4 output channels (decoder 3x3 -> 1 logit).  CUDA-core dot products; weight is the
 * fp32 master [K,R,S,C]; stride 1; dw / dbias are accumulated into. */
int tdl_conv2d_smallk_fwd(const void* x, const float* w, const float* bias, void* y, int N, int H, int W, int C, int K,
                          int R, int S, int dil, int pad_h, int pad_w, int P, int Q, void* stream);
int tdl_conv2d_smallk_bwd_data(const void* dy, const float* w, void* dx, int N, int H, int W, int C, int K, int R,
                               int S, int dil, int pad_h, int pad_w, int P, int Q, void* stream);
int tdl_conv2d_smallk_bwd_weight(const void* x, const void* dy, float* dw, float* dbias, int N, int H, int W, int C,
                                 int K, int R, int S, int dil, int pad_h, int pad_w, int P, int Q, void* stream);
int tdl_colsum_bf16(const void* dy, float* out, int64_t rows, int C, void* stream); /* out[C] += column sums (bias gradient) */
int tdl_upsample_fwd(const void* x, void* y, int N, int H, int W, int C, int OH, int OW, int is_bf16, void* stream);
int tdl_upsample_bwd(const void* dy, float* dx, int N, int H, int W, int C, int OH, int OW, int is_bf16, void* stream);
int tdl_u8_to_bf16_normalize(const void* x_u8, void* y_bf16, int64_t pixels, int C, int Cpad, const float* mean,
                             const float* inv_std, void* stream);

/* ---- batch augmentation on the GPU (reference: preprocessing/preprocessing.py:112-246, host-side there) ---- */
int tdl_augment_seg(const float* images, const float* masks, const float* transforms, const int* transposed,
                    float* out, float* out_mask, int B, int H, int W, int pad, float mean, float stddev,
                    int laplace_2d, void* stream);
/* same, with raw uint8 decoded inputs (input_is_u8: the kernel divides by 255) and an optional per-sample brightness
 * delta (preprocessing/preprocessing.py:169-170); this is the entry point the input pipeline calls per batch */
int tdl_augment_seg_ex(const void* images, const void* masks, int input_is_u8, const float* transforms,
                       const int* transposed, const float* brightness, float* out, float* out_mask, int B, int H,
                       int W, int pad, float mean, float stddev, int laplace_2d, void* stream);

/* ---- losses / metrics (reference: core/losses.py:5-92 pinned to CPU at model.py:391; core/metric.py:6-71) ---- */
int tdl_lovasz_hinge(const float* logits, const float* labels, int B, int P, float grad_scale, float* loss_out,
                     float* dlogits, void* stream);
int tdl_seg_metrics(const float* logits, const float* labels, int B, int P, float logit_threshold,
                    int reference_compat, float* iou_out, float* acc_out, void* stream);
int tdl_softmax_ce(const float* logits, const void* labels_i64, int rows, int classes, float grad_scale, int topk,
                   float* loss_rows, float* dlogits, float* top1, float* topk_out, void* stream);

/* ---- optimizers + device-resident schedule (reference: model.py:457-467) ---- */
int tdl_lr_schedule_step(void* global_step_i64, float base_lr, float decay_rate, float decay_steps, float beta1,
                         float beta2, float* lr_state, void* stream);
int tdl_sgd_momentum_flat(float* param, void* grad, float* mom, void* shadow_bf16, int64_t n, const float* lr_state,
                          float momentum, float weight_decay, float grad_scale, int grad_is_bf16, int zero_grad,
                          void* stream);
int tdl_adam_flat(float* param, void* grad, float* m1, float* m2, void* shadow_bf16, int64_t n, const float* lr_state,
                  float beta1, float beta2, float eps, float grad_scale, int grad_is_bf16, int zero_grad,
                  void* stream);
int tdl_cast_f32_to_bf16(const float* src, void* dst, int64_t n, void* stream);

/* BatchNorm (no activation) in front of a 1x1 convolution folded into the convolution: W' = W * scale (bf16),
 * b' = b + W shift; and the matching correction of the weight gradient computed on the un-normalised input
 * (dw = scale * dw + shift x dzsum, in place). */
int tdl_conv1x1_fold_affine(const float* w /* [K][C] */, const float* scale, const float* shift,
                            const float* bias_in /* or NULL */, void* wf_bf16, float* bias_out, int K, int C,
                            void* stream);
int tdl_conv1x1_fold_wgrad_fixup(float* dw /* [K][C] */, const float* scale, const float* shift,
                                 const float* dzsum /* [K], or NULL = all zero */, int K, int C, void* stream);

/* ---- symmetric heap on the CUDA VMM API with an NVLS multicast mapping (csrc/vmm_heap.cu) ----
 * One process per GPU.  Every tower: tdl_vmm_query -> pad the size to *granularity -> tdl_vmm_alloc (maps its own
 * physical allocation, returns a POSIX descriptor) -> pass the descriptor to every peer (UNIX socket, SCM_RIGHTS) ->
 * tdl_vmm_import each peer's.  Multicast: tower 0 tdl_vmm_mc_create, descriptor to all, others tdl_vmm_mc_import; all
 * tdl_vmm_mc_add_device; barrier; all tdl_vmm_mc_bind_map -> the address multimem.ld_reduce / multimem.st take.
 * Imported descriptors become the library's (closed by the matching free). */
int tdl_vmm_query(int world, int* vmm_ok, int* mc_ok, int64_t* granularity);
int tdl_vmm_alloc(int64_t padded_bytes, void** alloc_out, void** ptr_out, int* fd_out);
int tdl_vmm_import(int fd, int64_t padded_bytes, void** alloc_out, void** ptr_out);
int tdl_vmm_free(void* alloc);
int tdl_vmm_mc_create(int world, int64_t padded_bytes, void** mc_out, int* fd_out);
int tdl_vmm_mc_import(int fd, int64_t padded_bytes, void** mc_out);
int tdl_vmm_mc_add_device(void* mc);
int tdl_vmm_mc_bind_map(void* mc, void* alloc, void** mc_ptr_out);
int tdl_vmm_mc_free(void* mc);

/* ---- symmetric heap + P2P fused all-reduce/optimizer over NVLink (reference: MirroredStrategy NCCL all-reduce,
 *      model.py:115-119, Test.ipynb:200) ---- */
int tdl_symm_alloc(int64_t bytes, void** out);
int tdl_symm_free(void* p);
int tdl_ipc_export(void* p, void* handle64);
int tdl_ipc_open(const void* handle64, void** out);
int tdl_ipc_close(void* p);
int tdl_p2p_flag_bytes(void);
int tdl_p2p_comm_create(int world, int rank, void** grad_ptrs, void** param_ptrs, void** flag_ptrs, void* epoch_local,
                        int emulate, void** out_comm);
/* failure detection: how long a kernel waits for a peer's signal (seconds) before it records {rank, CTA, peer, epoch}
 * in a host-mapped status word and traps; tdl_p2p_comm_status reads the word back (out5 = timed_out, rank, cta, peer,
 * epoch) and works after the trap has poisoned the CUDA context */
int tdl_p2p_comm_set_timeout(void* comm, double seconds);
int tdl_p2p_comm_status(void* comm, int* out5);
/* protocol assertions for the signal pads (debug: TDL_P2P_DEBUG=1): every poll checks that a peer's epoch never goes
 * backwards and is at most one barrier ahead; `seen` = zeroed device array of 256 * 8 u32 (NULL = off) */
int tdl_p2p_comm_set_debug(void* comm, void* seen);
/* bf16 wire format: gradients accumulate in fp32 buckets but cross NVLink as bf16 (the cast is fused into the bucket
 * kernel: phase 1 packs + clears the fp32 bucket, phase 2 reduces the peers' bf16 copies in fp32 and runs the optimizer).
 * wire_ptrs[rank] = peer-mapped bf16 buffers with the gradient buckets' element offsets. */
int tdl_p2p_comm_set_wire(void* comm, void** wire_ptrs);
int tdl_p2p_fused_step_wire(void* comm, int64_t offset, int64_t n, int mode, int kind, int zero_grad, float* m1,
                            float* m2, const float* lr_state, float a, float b, float c, float grad_scale,
                            int max_blocks, int phases, void* stream);
int tdl_p2p_comm_set_multicast(void* comm, void* mc_grad, void* mc_param); /* NVLS multicast addresses, mode 3 */
int tdl_p2p_comm_destroy(void* comm);
int tdl_p2p_broadcast(void* comm, int src, int64_t offset, int64_t n, int max_blocks, void* stream); /* param mirror */
int tdl_p2p_fused_step(void* comm, int64_t offset, int64_t n, int mode, int kind, int grad_is_bf16, int zero_grad,
                       float* m1, float* m2, const float* lr_state, float a, float b, float c, float grad_scale,
                       int max_blocks, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TDL_B200_H_ */
