/* b2c.h - C-ABI of libb2c.so: hand-written sm_100a kernels for the cGAN patch-calibrator train step.
 *
 * The reference (1213ray/GAN-Calibrated-Semi-Supervised-Learning) has no FFI of its own: its hot path is Python calling
 * PyTorch ATen.  Each entry point below names the reference call site whose device work it replaces (paths relative
 * to the reference root).  Conventions: plain device pointers and sizes, no ownership transfer, no allocation, every
 * call enqueues on `stream` (a cudaStream_t) and returns 0 on success, else non-zero with b2c_last_error() set.
 * Activations: NHWC bf16 with an explicit pixel stride `ld*` in elements; parameters, statistics, losses: fp32.
 */
#ifndef B2C_H
#define B2C_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* b2c_last_error(void);
int b2c_abi_version(void); /* 2: epilogue scale, deterministic split-K / statistics reductions, batched spectral norm, flat Adam */
int b2c_zero(void* p, long long bytes, void* stream); /* cudaMemsetAsync: gradient buckets (optimizer.zero_grad, cgan_train_enhanced.py:305,346) */

/* ---- weight packing: fp32 master weights -> bf16 tensor-core operands.  inv_sigma (nullable) folds the spectral-norm
 * 1/sigma in (torch spectral_norm.py:111-112 via cgan/models.py:238); the train step passes NULL here, packs once per
 * optimizer step and applies 1/sigma as the `scale` of the GEMM epilogues instead ---- */
int b2c_pack_w_c(const float* w, const float* inv_sigma, void* out_bf16, int D0, int D1, void* stream);
int b2c_pack_w_t(const float* w, const float* inv_sigma, void* out_bf16, int D0, int D1, void* stream);
/* all stale packs of a network in one launch (after its optimizer step): HOST arrays of njobs entries, kind 0 = C-form,
 * 1 = T-form, un-scaled (inv_sigma = NULL above) */
int b2c_pack_w_multi(const void* const* w, void* const* out_bf16, const int* D0, const int* D1, const int* kind, int njobs,
                     void* stream);

/* ---- tcgen05 implicit-GEMM convolutions ----
 * conv4s2 : nn.Conv2d(k4,s2,p1) fprop (cgan/models.py:57,236) and ConvTranspose2d dgrad.   x dense (N,H,W,Cin).
 * convt4s2: nn.ConvTranspose2d(k4,s2,p1) fprop (cgan/models.py:72,113) and Conv2d dgrad.   y is (N,2H,2W,Cout).
 * wgrad4s2: weight gradient of both (autograd of the above; cgan/cgan_train_enhanced.py:330,366).
 * stats (nullable, int64 [N][Cout][2]) <- per-(image, channel) (sum, sum of squares) of the bf16 output, i.e. what
 * b2c_in_stats computes: the InstanceNorm2d that follows every such conv (cgan/models.py:59,73,114,241) gets its
 * statistics from the GEMM epilogue instead of re-reading the tensor.  Statistics are 64-bit FIXED-POINT sums, value =
 * int64 * 2^-30: integer atomics are order-independent, so a forward pass is bit-reproducible. */
/* scale (nullable device scalar): y = acc * scale + bias, i.e. W_eff = W * scale: the spectral-norm 1/sigma of THIS forward
 * instance without re-packing the weights. */
int b2c_conv4s2_bf16(const void* x, long long ldx, const void* wpack_c, const float* scale, const float* bias, void* y,
                     long long ldy, long long* stats, int N, int H, int W, int Cin, int Cout, void* stream);
int b2c_convt4s2_bf16(const void* x, long long ldx, const void* wpack_t, const float* scale, const float* bias, void* y,
                      long long ldy, long long* stats, const void* act, float slope, int N, int H, int W, int Cin, int Cout,
                      void* stream);
/* act (nullable, Cout = 64 only): dense bf16 (N,2H,2W,64) output of the LeakyReLU(slope) that feeds the conv whose data
 * gradient this call computes; y is multiplied by (act > 0 ? 1 : slope), i.e. LeakyReLU's backward (cgan/models.py:60,245)
 * is fused into the dgrad epilogue instead of a separate full-tensor pass. */
int b2c_set_f32_output(int on, void* stream); /* fp32 CHECK MODE for the two entry points above: while on, `y` is float*
                                               * (ldy in floats) and receives the unrounded fp32 accumulators; tests only */
/* wgrad4s2: dw[Cs][Cl][4][4] (=|+=) scale * sum_pix small (x) large.  ws: fp32 workspace of ws_floats floats; with
 * b2c_wgrad4s2_ws_floats(...) floats the split-K partials are reduced in a fixed order (deterministic), with less (>= 16*Cs*Cl)
 * by fp32 atomics.  scale: nullable device scalar (spectral norm: 1/sigma, so dw is the gradient of weight_orig up to the
 * rank-1 term).  wdot / part / nparts_host (nullable, together): part[0..*nparts_host) <- block partials of
 * <dw_unscaled, wdot> for b2c_sn_rank1 (part: room for 1024 floats; nparts_host: HOST int). */
long long b2c_wgrad4s2_ws_floats(int N, int Hs, int Ws, int Cs, int Cl, int sm_count);
int b2c_wgrad4s2_bf16(const void* small, long long lds, const void* large, long long ldl, float* ws, long long ws_floats,
                      float* dw, int accumulate, const float* scale, const float* wdot, float* part, int* nparts_host,
                      int N, int Hs, int Ws, int Cs, int Cl, int sm_count, void* stream);

/* ---- GeneratorSimpleRegressor (cgan/models.py:147-216, generator_type "simple"): Conv2d(k3,s1,p1) on the same tcgen05 tap
 * GEMM (9 taps), its weight packing (flip != 0: the data-gradient form), the first 3x3 conv on the NCHW fp32 patch,
 * MaxPool2d(2,2) forward / backward, and the regression head AvgPool -> Linear 512-256-64-4 -> Tanh -> * delta_scale ---- */
int b2c_pack_w_k3(const float* w, void* out_bf16, int Cout, int Cin, int flip, void* stream);
int b2c_conv3s1_bf16(const void* x, long long ldx, const void* wpack_k3, const float* scale, const float* bias, void* y,
                     long long ldy, long long* stats, int N, int H, int W, int Cin, int Cout, void* stream);
int b2c_conv3x3_first(const float* x, const float* w, const float* bias, void* y, int N, int H, int W, void* stream);
int b2c_maxpool2_fwd(const void* a, void* out, int N, int H, int W, int C, void* stream);
int b2c_maxpool2_bwd(const void* a, const void* g, void* ga, int N, int H, int W, int C, void* stream);
int b2c_simple_head_fwd(const void* a, int HW, const float* W1, const float* b1, const float* W2, const float* b2,
                        const float* W3, const float* b3, const uint8_t* m1, const uint8_t* m2, float scale, float* delta,
                        float* acts, int N, void* stream);

/* ---- GeneratorSimpleRegressor backward (training with generator_type: "simple", cgan/cgan_train_enhanced.py:26-31) ----
 * wgrad3s1: weight gradient of nn.Conv2d(Cl, Cs, 3, padding=1): small = grad of the conv output (N,H,W,Cs), large = conv input
 * (N,H,W,Cl), dense NHWC bf16; dw fp32 [Cs][Cl][3][3].  The input gradient is b2c_conv3s1_bf16 with b2c_pack_w_k3(flip = 1). */
long long b2c_wgrad3s1_ws_floats(int N, int H, int W, int Cs, int Cl, int sm_count);
int b2c_wgrad3s1_bf16(const void* small, const void* large, float* ws, long long ws_floats, float* dw, int accumulate, int N,
                      int H, int W, int Cs, int Cl, int sm_count, void* stream);
int b2c_conv3x3_first_wgrad(const float* x, const void* gz, float* dw, int N, int H, int W, void* stream);
int b2c_simple_head_bwd(const float* gdelta, const float* acts, const float* W1, const float* W2, const float* W3,
                        const uint8_t* m1, const uint8_t* m2, float scale, int HW, float* gv, void* ga, float* dW1, float* db1,
                        float* dW2, float* db2, float* dW3, float* db3, int accumulate, int N, void* stream);

/* ---- dataset indexing: CalibratorDataset._prepare_index / _hungarian_matching / _bbox_iou (cgan/dataset.py:57-70,128-205).
 * pred (P,4), gt (G,4): normalised (cx,cy,w,h) fp32, 16-byte aligned; gt_begin / gt_end (P): the row range in `gt` of the image
 * pred box i belongs to.  best_iou (P) <- max IoU over that range (the reference's fp32 arithmetic, bit-identical),
 * best_idx (P) <- index of its FIRST maximum within the range (torch.max(dim=1)), -1 for an empty range. */
int b2c_match_boxes(const float* pred, const float* gt, const int* gt_begin, const int* gt_end, float* best_iou,
                    int* best_idx, int P, void* stream);

/* ---- first layers, Cin = 3 / 6, NCHW fp32 in (cgan/models.py:103,245,257; interpolation cgan/losses.py:203-204) ---- */
int b2c_first_fprop(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha,
                    const float* w, const float* inv_sigma, const float* bias, void* y, int N, int H, int W,
                    float slope, void* stream);
/* first_fprop that also leaves the SIGN BITS of its output (one 64-bit word per output pixel, bit c = (y[c] > 0)), and the
 * Cout = 64 transposed conv / conv data gradient whose fused LeakyReLU backward (cgan/models.py:60,241 under autograd) reads those
 * bits instead of the activation tensor (8 instead of 128 bytes per output pixel in the epilogue). */
int b2c_first_fprop_bits(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha,
                         const float* w, const float* inv_sigma, const float* bias, void* y, int N, int H, int W,
                         float slope, unsigned long long* sign_bits, void* stream);
int b2c_convt4s2_signbits_bf16(const void* x, long long ldx, const void* wpack_t, const float* scale, void* y, long long ldy,
                               const unsigned long long* act_bits, float slope, int N, int H, int W, int Cin, int Cout,
                               void* stream);
int b2c_first_dgrad(const void* gz, const float* w, const float* inv_sigma, float* gx0, float* gx1, void* ws, int N,
                    int H, int W, void* stream); /* ws: >= 36864 bytes of scratch for the packed tap weights */
int b2c_first_wgrad(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha,
                    const void* gz, float* dw, float* db, const float* scale, const float* wdot, float* part,
                    int* nparts_host, int N, int H, int W, void* stream); /* dw/db ACCUMULATED into; scale/wdot/part as in wgrad4s2 */

/* ---- PatchGAN head Conv2d(512,1,k4,s1,p1) (cgan/models.py:252) ---- */
int b2c_head_fprop(const void* a, const float* w, float* out, int N, int h, int C, void* stream);
int b2c_head_dgrad(const float* gy, const float* w, void* ga, int N, int h, int C, void* stream);
int b2c_head_wgrad(const void* a, const float* gy, float* dw, int N, int h, int C, void* stream);

/* ---- InstanceNorm2d + LeakyReLU/ReLU + Dropout(0.5) (cgan/models.py:59-62,73-77,114,241-242) ---- */
int b2c_in_stats(const void* z, long long ldz, long long* zs, int N, int HW, int C, void* stream); /* zs: int64 [N][C][2], see above */
int b2c_in_act_fwd(const void* z, long long ldz, const long long* zs, const uint8_t* mask, float slope, void* o0,
                   long long ldo0, void* o1, long long ldo1, int N, int HW, int C, void* stream);
/* pooled_sums: planar int64 fixed-point [2][N][C] = (sum act(x_hat), sum act'(x_hat)) per plane (value = int64 * 2^-30) */
int b2c_in_act_pool_fwd(const void* z, long long ldz, const long long* zs, float slope, long long* pooled_sums, int N, int HW,
                        int C, void* stream);
/* g: full gradient tensor, or gvec: per-(n,c) broadcast gradient with (optionally) fwd_sums = the pooled_sums of the forward,
 * which makes the backward a single apply pass.  Planes of >= 64 pixels without a mask run as ONE cluster kernel (TMA ring,
 * plane sums exchanged through distributed shared memory, operands cross HBM once); ws: fp32 [N][C][2] scratch otherwise. */
int b2c_in_act_bwd(const void* g, long long ldg, const float* gvec, const long long* fwd_sums, const void* z, long long ldz,
                   const long long* zs, const uint8_t* mask, float slope, float* ws, void* gz, long long ldgz, int N, int HW,
                   int C, void* stream);
/* fp32 CHECK MODE of the pair above (north_star: <= 1e-4 per layer in an fp32 check mode; with b2c_set_f32_output a whole
 * conv -> InstanceNorm -> activation block runs on fp32 storage): z / g / out dense NHWC fp32.  Verification only. */
int b2c_in_act_fwd_f32(const float* z, float slope, float* out, int N, int HW, int C, void* stream);
int b2c_in_act_bwd_f32(const float* g, const float* z, float slope, float* gz, int N, int HW, int C, void* stream);
/* Allocation-free variant of b2c_in_act_bwd (full-gradient mode: g given; same result).  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that
 * forbid cudaMalloc).  The [N][C][2] fp32 scratch is a file-scope device array inside the library (1 MiB per GPU that loads
 * it, held until exit): N * C <= 131072 planes, larger calls are rejected with B2C_ERR_ARG.  The call allocates nothing (no
 * cudaMalloc / cudaFree / torch).  All calls share the array; it is cleared at the start of every call on `stream`; calls that
 * overlap on the GPU (other streams, concurrent graphs) race on it. */
int b2c_in_act_bwd_static(const void* g, long long ldg, const void* z, long long ldz, const long long* zs, const uint8_t* mask,
                          float slope, void* gz, long long ldgz, int N, int HW, int C, void* stream);
int b2c_in_act_bwd2(const void* h, long long ldh, const void* g, long long ldg, const void* z, long long ldz,
                    const long long* zs, const uint8_t* mask, float slope, float* ws, void* gg, long long ldgg, void* gx,
                    long long ldgx, int N, int HW, int C, void* stream);
int b2c_lrelu_bwd(const void* g, const void* a, void* out, long long n, float slope, void* stream);
int b2c_copy_strided(const void* src, long long lds, void* dst, long long ldd, long long npix, int C, void* stream);
/* out = a + b on NHWC channel slices (pixel strides lda/ldb/ldo): gradient accumulation of a U-Net skip tensor that feeds
   both the next down block and the concat (cgan/models.py:82,224-230), replacing autograd's strided add. */
int b2c_add_strided(const void* a, long long lda, const void* b, long long ldb, void* out, long long ldo, long long npix,
                    int C, void* stream);
int b2c_dropout_mask(uint8_t* mask, long long n, unsigned long long seed, unsigned long long offset,
                     const long long* offset_base, void* stream); /* stream position = offset + *offset_base (device, nullable) */
/* out[i] = U[0,1) from the same counter-based stream: the GP interpolation coefficients (torch.rand at cgan/losses.py:199) */
int b2c_uniform(float* out, long long n, unsigned long long seed, unsigned long long offset, const long long* offset_base,
                void* stream);

/* ---- spectral norm (torch spectral_norm.py:92-114); ws: K + ceil(K/32) + R floats of scratch (ceil(R/32)*K + R suffices for R > 32) ---- */
int b2c_sn_power_iter(const float* W, float* u, float* v, float* sig, float* ws, int R, int K, int train, void* stream);
/* b2c_sn_bwd: ws = min(592, ceil(R*K/1024)) floats of scratch (block partials of <dWe, W>, summed in a fixed order) */
int b2c_sn_bwd(const float* dWe, const float* W, const float* u, const float* v, const float* sig, float* ws, float* dW,
               int R, int K, int accumulate, void* stream);
/* All spectral-norm layers of one Discriminator forward (cgan/models.py:245-250) in three launches; arrays are HOST arrays of
 * `count` (<= 8) device pointers / sizes; u_snap / v_snap (nullable arrays): copies of the u / v this forward used. */
int b2c_sn_power_iter_multi(const float* const* W, float* const* u, float* const* v, float* const* sig, float* const* ws,
                            float* const* u_snap, float* const* v_snap, const int* R, const int* K, int count, int train,
                            void* stream);
/* Deferred rank-1 term of the spectral-norm backward, once per backward for all layers: grad[l] -= sum_i (<dWe_i, W> /
 * sigma_i^2) u_i v_i^T over the forward instances i of layer l (flat HOST arrays in layer order; part_i[nparts_i] = block
 * partials of <dWe_i, W> left by b2c_wgrad4s2_bf16 / b2c_first_wgrad). */
int b2c_sn_rank1(float* const* grad, const int* R, const int* K, const int* ninst, int nlayers, const float* const* u,
                 const float* const* v, const float* const* sig, const float* const* part, const int* nparts, void* stream);

/* ---- losses (cgan/losses.py:19-73,99-183,223-231; cgan/cgan_train_enhanced.py:327-328,362) ---- */
int b2c_apply_delta(const float* bbox, const float* delta, int training, float* out, const float* gout, float* gdelta,
                    int B, void* stream);
int b2c_eiou(const float* pred, const float* tgt, float eps, float* loss, const float* gscale, float* gpred,
             float* iou_out, int B, void* stream);
int b2c_rowsum(const float* x, long long n, int B, int square, float scale, float* out, void* stream);
int b2c_gp_sumsq(const float* g0, const float* g1, long long n, int B, float* part, int* splits_host, void* stream);
int b2c_gp_tail(const float* ss, int splits, float* gp, float* coef, int B, void* stream); /* ss: [B][splits] from b2c_gp_sumsq */
int b2c_map_mean(const float* x, long long n, float scale, float* out, void* stream); /* out[0] = scale * sum(x), one block */
/* d_loss = -(mean(real) - mean(fake)) + lambda_gp * gp in one launch (cgan_train_enhanced.py:327-328): out3 = (d_loss, wd, gp) */
int b2c_critic_loss(const float* real, long long n_real, const float* fake, long long n_fake, const float* gp,
                    float lambda_gp, float* out3, void* stream);
int b2c_critic_loss_bwd(const float* gscale, float* g_real, long long n_real, float* g_fake, long long n_fake, void* stream);
int b2c_scale_rows(const float* x, const float* coef, const float* gscale, float* out, long long n, int B, void* stream);
int b2c_fill_scaled(float* out, long long n, const float* gscale, float value, void* stream);

/* ---- refined patches / letterbox on the device: crop -> grey-pad to a square -> PIL-BICUBIC resize (bit-identical restatement
 * of Pillow's ImagingResample) -> ToTensor + Normalize(0.5, 0.5).  mode 0: CalibratorDataset._letterbox (cgan/dataset.py:103-124);
 * mode 1: get_refined_patch_batch (cgan/cgan_train_enhanced.py:37-137: hard-clamp delta update, [0.1,0.9]/[0.05,0.8] clamps,
 * "< 10 px -> pred_box" fallback); mode 2: inference crop_patch + letterbox (cgan/inference.py:51-66).  images: uint8 HWC RGB
 * images back to back (img_off/img_h/img_w per image), sample_img[B]: image of each sample, qmax >= largest image side. ---- */
long long b2c_refine_ws_bytes(int B, int S, int qmax);
int b2c_refine_patches(const uint8_t* images, const long long* img_off, const int* img_h, const int* img_w,
                       const int* sample_img, const float* pred_box, const float* delta, int mode, const float* fallback,
                       void* ws, long long ws_bytes, float* out, uint8_t* out_u8, float* box_out, int B, int S, int qmax,
                       void* stream);

/* ---- G head AvgPool -> Linear(64,4) -> Tanh -> *delta_scale (cgan/models.py:118-123,139-141) ---- */
int b2c_pool_fc_fwd(const long long* pooled_sums, const float* W, const float* bias, float scale, int HW, float* delta,
                    float* tanh_out, int N, int C, void* stream);
int b2c_pool_fc_bwd(const float* gdelta, const float* tanh_v, const long long* pooled_sums, const float* W, float scale,
                    int HW, float* dW, float* db, float* gvec, int N, int C, void* stream);

/* ---- clip_grad_norm_(1.0) + Adam (cgan/cgan_train_enhanced.py:256-257,331-332,368-369) ---- */
/* part[0..*nparts_host) <- per-block partials of sum g^2 (room for 1024 floats; every slot written, no atomics) */
int b2c_multi_sumsq(float* const* g, const long long* n, int count, float* part, int* nparts_host, void* stream);
/* g_eff = grad_scale * g (data parallel: 1/world after a SUM all-reduce); clip coefficient min(1, max_norm / (|g_eff| + 1e-6))
 * from the fixed-order sum of gnorm_part[nparts] (nullable); gnorm_sq_out[0] (nullable) <- |g_eff|^2; lr_dev (nullable device
 * scalar) overrides lr (ReduceLROnPlateau, cgan_train_enhanced.py:260-261,427-428, under a captured CUDA graph); step count =
 * step_index + *step_base (device, nullable).  128-bit vectorised when the tensors allow it. */
int b2c_multi_adam(float* const* p, float* const* g, float* const* m, float* const* v, const long long* n, int count,
                   const float* gnorm_part, int nparts, float* gnorm_sq_out, float grad_scale, float max_norm, float lr,
                   const float* lr_dev, float beta1, float beta2, float eps, int step_index, const long long* step_base,
                   void* stream);

#ifdef __cplusplus
}
#endif
#endif
