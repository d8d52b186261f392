/* zkb200.h -- C ABI of libzkb200.so: the hand-written sm_100a (B200) kernels behind the BinaryNet
 * train hot path of a zookeeper-configured Larq experiment.
 *
 * The reference (AdamHillier/zookeeper) has no FFI for this path: every operator below is a
 * Python call into TensorFlow / Larq made from examples/larq_experiment.py.  Each entry point cites
 * the reference call site (file:line under /root/reference) whose arithmetic it replaces.
 *
 * Conventions
 *   - Plain C types only.  All tensor pointers are DEVICE pointers owned by the caller (NHWC,
 *     channels innermost, 16-byte aligned); the library never allocates or frees caller-visible
 *     memory.  `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); every
 *     call only enqueues work and returns without synchronising.
 *   - Return value: 0 = ok; < 0 = argument error detected before any launch; > 0 = the
 *     cudaError_t / CUresult of the failing runtime call.  zkb_last_error_string() returns the
 *     message of the last failure on the calling thread.
 *   - "pm1 bf16" = bfloat16 holding exactly +1.0 or -1.0 (0x3F80 / 0xBF80): the output of larq's
 *     ste_sign.  Products of two such operands accumulate in f32 inside tensor memory and are
 *     exact integers for K < 2^24.
 *   - "hi/lo" = a float gradient split as hi = bf16(g), lo = bf16(g - hi); both are multiplied by
 *     the exact +-1 operand on tcgen05 and summed in f32 (about 16 mantissa bits).
 */
#ifndef ZKB200_H
#define ZKB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

int zkb_abi_version(void);   /* 2 */
const char* zkb_last_error_string(void);

/* ---- data pipeline --------------------------------------------------------------------------
 * Fused batch map: gather `index` from a resident uint8 pool, centre zero-pad to pad_size, random
 * crop OHxOW, random left-right flip, float32(x)/127.5 - 1 (true division), optional sign
 * binarisation, sparse label copy and optional one-hot.  Per-sample (ox, oy, flip) come from
 * `aug` ([B,3] int8) or, when NULL, from a counter-based generator keyed by (seed, step, sample).
 * training == 0: centre crop/pad only.  step_dev (optional, device memory) overrides `step`, so a
 * captured CUDA graph can advance the generator between replays.
 * Replaces: examples/larq_experiment.py:24-37 (PadCropAndFlip.input/output) mapped per example by
 * tf.data at :126-139 through zookeeper/tf/preprocessing.py:59-75. */
int zkb_preprocess_u8(const uint8_t* images, const long long* labels, const int* index,
                      const int8_t* aug, float* out, int* labels_out, float* onehot, int B, int IH,
                      int IW, int C, int OH, int OW, int pad_size, int training, int binarize,
                      int num_classes, unsigned long long seed, unsigned long long step,
                      const unsigned long long* step_dev, void* stream);

/* ---- bit-level formulation (north_star items 2, 3) --------------------------------------------
 * Sign + bit-pack: 32 consecutive values along K -> one uint32 (bit j = 1 iff value >= 0, i.e.
 * +1; sign(0) = sign(-0.0) = +1).  n must be a multiple of 32.  Warp-ballot kernel.
 * Replaces: larq ste_sign forward, "ste_sign" at examples/larq_experiment.py:53-54,65. */
int zkb_sign_pack(const void* x, void* packed_u32, long long n, int x_is_bf16, void* stream);
/* XNOR-popcount GEMM on CUDA cores: y[m][n] = 32*k_words - 2*popc(a[m] ^ b[n]) (int32).
 * Replaces: QuantDense matmul, examples/larq_experiment.py:95-99. */
int zkb_bgemm_popc(const void* a_packed, const void* b_packed, int* y, long long ldy, int M, int N,
                   int k_words, void* stream);

/* ---- first layer: float image (*) sign(W), 'valid' 3x3, Cin in {1,3} ----------------------------
 * wq_k_co / dw_k_co: [ks*ks*Cin][Cout] == HWIO flattened.  wgrad accumulates into dw (zero it).
 * Replaces: lq.layers.QuantConv2D without input quantizer, examples/larq_experiment.py:62-69. */
int zkb_conv_first_fwd(const float* x, const float* wq_k_co, float* y, int B, int H, int W, int Cin,
                       int Cout, int ksize, void* stream);
int zkb_conv_first_wgrad(const float* x, const float* dy, float* dw_k_co, int B, int H, int W,
                         int Cin, int Cout, int ksize, void* stream);

/* First layer fused with its BatchNormalization: the small conv is recomputed from the image
 * inside each pass, so its [B,OH,OW,Cout] float map is never written to memory.
 *   bn_stats  : per-channel sum / sumsq of conv(x, wq)          (then zkb_bn_finalize)
 *   bn_sign   : sign((conv - mean)*rstd + beta) as pm1 bf16
 *   bwd_reduce: g = grad w.r.t. that sign; sum(da), sum(da*xhat) with the ste mask |a| <= 1
 *               (then zkb_bn_bwd_finalize)
 *   bwd_wgrad : dW[k][co] += sum_pix x[pix+tap k] * rstd*(da - mean_da - xhat*mean_dax)
 * Replaces: examples/larq_experiment.py:62-70 (first QuantConv2D + BatchNormalization) and their
 *           autodiff in model.fit (:147-153). */
int zkb_conv_first_bn_stats(const float* x, const float* wq_k_co, int B, int H, int W, int Cin,
                            int Cout, int ksize, double* sum, double* sumsq, void* stream);
int zkb_conv_first_bn_sign(const float* x, const float* wq_k_co, int B, int H, int W, int Cin,
                           int Cout, int ksize, const float* mean, const float* rstd,
                           const float* beta, void* s_out_bf16, void* s_out_i8, void* stream);
int zkb_conv_first_bwd_reduce(const float* x, const float* wq_k_co, const float* g, int B, int H,
                              int W, int Cin, int Cout, int ksize, const float* mean,
                              const float* rstd, const float* beta, double* sum_da,
                              double* sum_dax, void* stream);
int zkb_conv_first_bwd_wgrad(const float* x, const float* wq_k_co, const float* g, int B, int H,
                             int W, int Cin, int Cout, int ksize, const float* mean,
                             const float* rstd, const float* beta, const float* mean_da,
                             const float* mean_dax, float* dw_k_co, void* stream);

/* ---- binary conv / dense on tcgen05 -------------------------------------------------------------
 * Forward: act_pm1 [NB,H,W,Cin] (conv) or [rows,K] (dense), pm1 bf16; weights as written by
 * zkb_prep_*_weights; y int16 (exact integer sums, |y| <= ks*ks*Cin).  'same' zero padding is
 * produced by TMA im2col out-of-bounds fill.  Cin % 64 == 0, K % 64 == 0 (wgrad: the
 * reduced operand's width an odd multiple of 64 is handled by M tiles of two 64-row units that may span taps).
 * Replaces: lq.layers.QuantConv2D(padding="same", ste_sign/ste_sign)
 *           examples/larq_experiment.py:71-73,76-78,80-82,85-87,89-91 and lq.layers.QuantDense :95,97,99. */
int zkb_bconv2d_fwd(const void* act_pm1, const void* wq_oc_k, void* y_i16, int NB, int H, int W,
                    int Cin, int Cout, int ksize, void* stream);
int zkb_bdense_fwd(const void* act_pm1, const void* wq_n_k, void* y_i16, long long ldy, int rows,
                   int K, int N, void* stream);
/* The same forward products on the int8 pipe (tcgen05 kind::i8, int32 accumulators): operands are
 * int8 holding +1 / -1 (0 in padded weight rows); half the operand bytes and twice the MMA rate of
 * the bf16 form, bit-identical y.  Cin % 128 == 0, K % 128 == 0.  This is the path the engine's
 * forward runs. */
int zkb_bconv2d_fwd_i8(const void* act_i8, const void* wq_oc_k_i8, void* y_i16, int NB, int H, int W,
                       int Cin, int Cout, int ksize, void* stream);
int zkb_bdense_fwd_i8(const void* act_i8, const void* wq_n_k_i8, void* y_i16, long long ldy, int rows,
                      int K, int N, void* stream);
/* The int8 forward that also leaves the BatchNorm batch statistics of its output: per channel
 * sum += sum of y, sumsq += sum of y*y over all rows (float64 accumulators holding exact integers,
 * zeroed by the caller) -- what zkb_bn_stats would compute from y_i16 in a second pass.  The
 * epilogue reads each staged int16 box back column-wise before the TMA store leaves. */
int zkb_bconv2d_fwd_i8_stats(const void* act_i8, const void* wq_oc_k_i8, void* y_i16, int NB, int H,
                             int W, int Cin, int Cout, int ksize, double* sum, double* sumsq,
                             void* stream);
int zkb_bdense_fwd_i8_stats(const void* act_i8, const void* wq_n_k_i8, void* y_i16, long long ldy,
                            int rows, int K, int N, double* sum, double* sumsq, void* stream);
/* The int8 conv forward of a layer that is followed by MaxPool2D(2, 2): the epilogue stages each
 * tile -- whole row pairs of the output map -- in shared memory and pools it there (first maximum
 * in row-major window order, as zkb_maxpool2_i16), so only the pooled int16 map p [NB,H/2,W/2,Cout],
 * the winning slot arg (uint8, same shape) and the pooled map's BatchNorm sums leave the kernel; the
 * full-resolution map (2 B written + 2 B read per element otherwise) never reaches memory.  Needs an
 * even map whose row pairs tile 256 MMA columns and Cout % 128 == 0: zkb_bconv2d_fwd_i8_pool_ok
 * returns 1 for such shapes (26x26, 30x30, 6x6, ...), 0 otherwise (13x13, 15x15, 7x7: those layers
 * keep zkb_bconv2d_fwd_i8 + zkb_maxpool2_i16).  Bit-identical to the two-kernel form.
 * Replaces: QuantConv2D + MaxPool2D, examples/larq_experiment.py:71-74, 85-92. */
int zkb_bconv2d_fwd_i8_pool_ok(int H, int W, int Cout);
int zkb_bconv2d_fwd_i8_pool(const void* act_i8, const void* wq_oc_k_i8, void* p_i16, uint8_t* arg_u8, int NB,
                            int H, int W, int Cin, int Cout, int ksize, double* sum, double* sumsq,
                            void* stream);
/* Inference (validation pass, examples/larq_experiment.py:133-139, 151-152): BatchNormalization with
 * moving statistics followed by ste_sign is a step function of the integer conv/dense output, so
 * the two fold into one threshold per output channel.  zkb_bn_sign_thresholds writes thr[c] = the
 * smallest integer y with (float(y) - mean[c]) * rstd[c] + beta[c] >= 0 (the float expression of
 * zkb_bn_apply_sign; INT_MAX / INT_MIN if never / always).  The *_sign forwards compare the int32
 * accumulator with thr in the epilogue and store int8 +1 / -1 ([NB,H,W,Cout] / [rows,lds]); no
 * int16 map is written.  Cout (N) must be a multiple of the tile width min(256, Cout).
 * zkb_maxpool2_sign_i8: 2x2/2 max-pool of such signs (= sign of the normalised pooled maximum,
 * because rstd > 0); C % 16 == 0. */
/* rstd[c] = rsqrt(moving_var[c] + eps): the inference scale of BatchNormalization. */
int zkb_bn_eval_rstd(const float* moving_var, float eps, int C, float* rstd, void* stream);
int zkb_bn_sign_thresholds(const float* mean, const float* rstd, const float* beta, int C,
                           int* thr_i32, void* stream);
int zkb_bconv2d_fwd_i8_sign(const void* act_i8, const void* wq_oc_k_i8, const int* thr, void* s_out_i8,
                            int NB, int H, int W, int Cin, int Cout, int ksize, void* stream);
int zkb_bdense_fwd_i8_sign(const void* act_i8, const void* wq_n_k_i8, const int* thr, void* s_out_i8,
                           long long lds, int rows, int K, int N, void* stream);
int zkb_maxpool2_sign_i8(const void* s_i8, int B, int H, int W, int C, void* out_i8, void* stream);
/* Backward w.r.t. the input (before the STE mask): dx f32 [NB,H,W,Cin] / [rows,K].
 * Replaces: TF autodiff of the layers above (model.fit, examples/larq_experiment.py:147-153). */
int zkb_bconv2d_dgrad(const void* dy_hi, const void* dy_lo, const void* wq_ic_k, void* dx_f32,
                      int NB, int H, int W, int Cin, int Cout, int ksize, void* stream);
int zkb_bdense_dgrad(const void* dy_hi, const void* dy_lo, const void* wq_k_n, void* dx_f32,
                     int rows, int K, int N, void* stream);
/* zkb_bconv2d_dgrad that also leaves the BatchNorm-backward sums of the layer below (what
 * zkb_bn_bwd_reduce would compute from dx_f32 in a second pass): x_i16 [NB,H,W,Cin] is that layer's
 * pre-BatchNorm integer map (after pooling, if it pools), mean/rstd/beta its statistics; per channel
 * sum_da += sum of da, sum_dax += sum of da*xhat with da = dx masked by |xhat + beta| <= 1.  The
 * epilogue reads each dx tile back from its shared-memory staging box column-wise, so the sums cost
 * one int16 read of x instead of an f32 read of dx plus x.  Cin % 128 == 0. */
int zkb_bconv2d_dgrad_bn(const void* dy_hi, const void* dy_lo, const void* wq_ic_k, void* dx_f32,
                         int NB, int H, int W, int Cin, int Cout, int ksize, const void* x_i16,
                         const float* mean, const float* rstd, const float* beta, double* sum_da,
                         double* sum_dax, void* stream);
/* The second conv's dgrad with the WHOLE backward of the first layer in its epilogue: the gradient
 * w.r.t. the first sign map ([NB,H,W,128] f32) is consumed from tensor memory and never stored.
 * Per output channel the epilogue recomputes the first conv (3x3 'valid', one input channel, from
 * x0 [NB,H0,W0] with H0 = H+2, W0 = W+2 and sign(W0) wq0_k_co [9][128]), BatchNorm (mean0, rstd0,
 * beta0) and the ste_sign mask, and adds into acc (float64 [11][128], zeroed by the caller):
 *   acc[0][c] = sum da     acc[1][c] = sum da * xhat     acc[2+k][c] = S1 = sum x_k * da   (9 taps k)
 * zkb_conv_first_gram adds the Gram matrix of the input taps over all output pixels into gram
 * (float64: G[9][9] = sum x_k x_k', then S2[9] = sum x_k; zeroed by the caller) -- the part of the
 * first layer that does not depend on the gradient.  zkb_conv_first_stats_from_gram adds the conv
 * output's per-channel sum / sum of squares (what zkb_bn_finalize consumes) computed from it, and
 * zkb_conv_first_bwd_finalize writes dbeta0 = dbeta_scale * sum da and ADDS the kernel gradient
 *   dW0[k][c] = rstd_c * (S1 - mean(da) * S2[k] - mean(da*xhat) * S3),
 *   S3 = sum x_k * xhat_c = rstd_c * (sum_k' w_k'c G[k][k'] - mean_c * S2[k])
 * into dw0_k_co; count is the number of pixels the two means run over (all replicas' under
 * synchronised BatchNorm, after an all-reduce of acc[0..1][*]).  Cin == 128, ksize == 3.
 * Replaces: TF autodiff of Conv2D -> BatchNormalization -> ste_sign input quantizer -> QuantConv2D
 * (examples/larq_experiment.py:62-78 under model.fit, :147-153). */
int zkb_bconv2d_dgrad_first_bwd(const void* dy_hi, const void* dy_lo, const void* wq_ic_k, int NB,
                                int H, int W, int Cin, int Cout, int ksize, const float* x0, int H0,
                                int W0, const float* wq0_k_co, const float* mean0, const float* rstd0,
                                const float* beta0, double* acc, void* stream);
int zkb_conv_first_gram(const float* x0, int NB, int H0, int W0, double* gram, void* stream);
int zkb_conv_first_stats_from_gram(const double* gram, const float* wq0_k_co, int Cout, double* sum,
                                   double* sumsq, void* stream);
int zkb_conv_first_bwd_finalize(const double* acc, const double* gram, const float* wq0_k_co,
                                const float* mean0, const float* rstd0, double count,
                                double dbeta_scale, float* dbeta0, float* dw0_k_co, int Cout,
                                void* stream);
/* Backward w.r.t. the latent kernel: dw f32 in HWIO / [K][ldw], ACCUMULATED into (zero it first;
 * partial tiles are combined with float atomics).  n_out <= N columns are written. */
int zkb_bconv2d_wgrad(const void* act_pm1, const void* dy_hi, const void* dy_lo, float* dw_hwio,
                      int NB, int H, int W, int Cin, int Cout, int ksize, void* stream);
int zkb_bdense_wgrad(const void* act_pm1, const void* dy_hi, const void* dy_lo, float* dw_k_n,
                     long long ldw, int rows, int K, int N, int n_out, void* stream);

/* ---- BatchNormalization(scale=False), training mode ---------------------------------------------
 * x: f32 or int16 [rows][C].  stats: per-channel sum / sum of squares into double accumulators
 * (caller zeroes them).  finalize: mean, rstd = rsqrt(biased var + eps); moving <- moving*momentum
 * + batch*(1-momentum), with the Bessel-corrected variance when bessel != 0 (TF fused 4-D kernel).
 * apply_sign: a = (x-mean)*rstd + beta; s = sign(a) as pm1 bf16 and/or int8 and/or a itself
 *             (any output pointer may be NULL).
 * C: any multiple of one 16-byte vector (8 int16 / 4 f32 channels) for stats, apply_sign and the
 * backward reduce / apply (f32 outputs or hi/lo): a block covers whole rows, maps wider than 256
 * vectors (4096-unit dense layers) are walked in column slices.  zkb_maxpool2_i16 and the pool
 * scatter of zkb_bn_bwd_apply (out_mode 2) take at most 256 vectors per row.
 * Replaces: tf.keras.layers.BatchNormalization(scale=False) examples/larq_experiment.py:70,75,79,
 *           84,88,93,96,98,100 and the ste_sign input quantizer of the following layer (:53). */
int zkb_bn_stats(const void* x, int x_is_i16, long long rows, int C, double* sum, double* sumsq,
                 void* stream);
int zkb_bn_finalize(const double* sum, const double* sumsq, double count, float eps, float momentum,
                    int bessel, int C, float* mean, float* rstd, float* moving_mean,
                    float* moving_var, void* stream);
int zkb_bn_apply_sign(const void* x, int x_is_i16, long long rows, int C, const float* mean,
                      const float* rstd, const float* beta, void* s_out_bf16, float* a_out,
                      void* s_out_i8, void* stream);
/* MaxPool2D(2,2) 'valid' on the integer conv output; first maximum in row-major window order wins
 * (arg = 0..3); also accumulates the BatchNorm statistics of the pooled map.
 * Replaces: tf.keras.layers.MaxPool2D examples/larq_experiment.py:74,83,92. */
int zkb_maxpool2_i16(const void* y, int B, int H, int W, int C, void* p_out, uint8_t* arg_out,
                     double* sum, double* sumsq, void* stream);
/* softmax + sparse categorical cross-entropy: loss_sum += sum_b -log softmax(z_b)[label_b],
 * correct += #argmax hits, dlogits = (softmax - onehot) * inv_global_batch (columns >= num_classes
 * up to ld are written as 0).
 * Replaces: Activation("softmax") + loss/metrics, examples/larq_experiment.py:101,117-118. */
int zkb_softmax_ce(const float* logits, int ld, const int* labels, int B, int num_classes,
                   float inv_global_batch, float* dlogits, double* loss_sum, int* correct,
                   void* stream);
/* BatchNorm backward.  g = gradient w.r.t. sign(a) (ste != 0: masked where |a| > 1, the ste_sign
 * backward) or w.r.t. a (ste == 0).  reduce: sum(da), sum(da*xhat); finalize: dbeta and the two
 * means (dbeta = dbeta_scale * sum(da): 1/world when the sums were all-reduced for SyncBN and the
 * gradient arena is summed across ranks afterwards); apply: dx = rstd*(da - mean(da) -
 * xhat*mean(da*xhat)) written as f32 (out_mode 0), as
 * hi/lo bf16 (1), or hi/lo scattered through the 2x2 pool argmax into the [B,H,W,C] map (2).
 * Replaces: TF autodiff of BatchNormalization / ste_sign / MaxPool2D (model.fit, :147-153). */
int zkb_bn_bwd_reduce(const float* g, const void* x, int x_is_i16, long long rows, int C,
                      const float* mean, const float* rstd, const float* beta, int ste,
                      double* sum_da, double* sum_dax, void* stream);
int zkb_bn_bwd_finalize(const double* sum_da, const double* sum_dax, double count,
                        double dbeta_scale, int C, float* dbeta, float* mean_da, float* mean_dax,
                        void* stream);
int zkb_bn_bwd_apply(const float* g, const void* x, int x_is_i16, long long rows, int C,
                     const float* mean, const float* rstd, const float* beta, const float* mean_da,
                     const float* mean_dax, int ste, int out_mode, float* out_f32, void* out_hi,
                     void* out_lo, const uint8_t* arg, int H, int W, void* stream);

/* ---- weights and optimizer ----------------------------------------------------------------------
 * sign(W) in the operand layouts of the GEMM engines (kernel_quantizer="ste_sign",
 * examples/larq_experiment.py:54,65): conv HWIO f32 -> fwd [Cout][tap][Cin], dgrad
 * [Cin][flipped tap][Cout]; dense [K][N] f32 -> fwd [Npad][K], dgrad [K][Npad] (pad = 0);
 * w_fwd_i8 (optional) receives the forward layout as int8 for the kind::i8 GEMM. */
int zkb_prep_conv_weights(const float* w_hwio, int ksize, int Cin, int Cout, void* w_fwd,
                          void* w_dgrad, void* w_fwd_i8, void* stream);
int zkb_prep_dense_weights(const float* w_k_n, int K, int N, int Npad, void* w_fwd, void* w_dgrad,
                           void* w_fwd_i8, void* stream);
int zkb_sign_f32(const float* w, float* out, long long n, void* stream);
/* Adam in the operation order of TF's ApplyAdam (lr_t = lr*sqrt(1-b2^t)/(1-b1^t) computed by the
 * caller) followed, for kernels (is_kernel != 0), by the ste_sign gradient mask |w| <= 1 and the
 * clip of the latent weight to [-1, 1].  lr_t_dev (optional, device memory) overrides lr_t (for
 * captured CUDA graphs).
 * Replaces: tf.keras.optimizers.Adam examples/larq_experiment.py:120-122 and
 *           kernel_constraint="weight_clip" :55,66. */
int zkb_adam_clip(float* w, const float* g, float* m, float* v, long long n, float lr_t,
                  float beta1, float beta2, float eps, float grad_scale, int is_kernel,
                  const float* lr_t_dev, void* stream);

/* ---- Bi-Real-Net style binary residual units ---------------
 * Not in the reference tree: BASELINE.json configs[3] and [4] name larq-zoo models, which the
 * reference reaches through the same hook as BinaryNet (a @factory model built from larq layers,
 * examples/larq_experiment.py:41-103, quantizer names in the kwhparams dict :52-57).  Semantics are
 * the published larq / larq-zoo ones, restated in oracle/bireal_oracle.py.  A unit is
 *   h_next = BatchNorm(QuantConv2D(h; input_quantizer="approx_sign",
 *                                  kernel_quantizer="magnitude_aware_sign")) + h
 * The binary contraction runs on the entry points above (sign(h) through zkb_bn_apply_sign with
 * mean 0 / rstd 1 / beta 0; forward zkb_bconv2d_fwd_i8_stats; backward zkb_bconv2d_dgrad / _wgrad
 * on dy = alpha * dz, which makes the wgrad output the latent-weight gradient directly).
 *
 * zkb_weight_abs_mean: alpha[c] = mean over the K = ks*ks*Cin rows of |W[k][c]| (magnitude_aware_sign's
 *   per-output-channel scale; its gradient is stopped, as in larq).
 * zkb_bn_gain_finalize: BatchNormalization WITH scale over z = alpha*y from the integer sums of y:
 *   mean_y; rs_bwd = alpha*rstd_z (xhat = (y - mean_y)*rs_bwd); gain = alpha*gamma*rstd_z; moving
 *   statistics of z (bessel: unbiased variance into the moving variance, TF's fused kernel).
 * zkb_bn_gain_add: out = (y - mean_y)*gain + beta (+ shortcut), f32.
 * zkb_bn_gain_bwd_apply: dy = gain*(d - mean_d - xhat*mean_dx) as the bf16 hi/lo pair; the sums
 *   behind mean_d / mean_dx (and dbeta, dgamma) come from zkb_bn_bwd_reduce(ste = 0, rstd = rs_bwd).
 * zkb_approx_sign_bwd_add: out = d_short + ds*(2 - 2|h|)*[|h| <= 1] (approx_sign's backward, plus
 *   the gradient arriving over the shortcut; d_short may be NULL).  n % 4 == 0. */
int zkb_weight_abs_mean(const float* w_k_co, long long K, int Cout, float* alpha, void* stream);
int zkb_bn_gain_finalize(const double* sum, const double* sumsq, double count, const float* alpha,
                         const float* gamma, float eps, float momentum, int bessel, int C,
                         float* mean_y, float* rs_bwd, float* gain, float* moving_mean,
                         float* moving_var, void* stream);
int zkb_bn_gain_add(const void* y_i16, long long rows, int C, const float* mean_y, const float* gain,
                    const float* beta, const float* shortcut, float* out, void* stream);
int zkb_bn_gain_bwd_apply(const float* d, const void* y_i16, long long rows, int C, const float* mean_y,
                          const float* rs_bwd, const float* gain, const float* mean_d,
                          const float* mean_dx, void* out_hi, void* out_lo, void* stream);
int zkb_approx_sign_bwd_add(const float* ds, const float* h, const float* d_short, float* out,
                            long long n, void* stream);
/* The float side of a down-sampling unit's shortcut (AveragePooling2D(2, 2) -> Conv2D 1x1 without
 * bias -> BatchNormalization): the pool and its backward, and the gain BatchNorm pair for a FLOAT
 * input map (alpha = 1: pass ones to zkb_bn_gain_finalize; statistics from zkb_bn_stats(x_is_i16 = 0),
 * backward sums from zkb_bn_bwd_reduce(x_is_i16 = 0, ste = 0)).  The 1x1 conv itself is
 * zkb_dense_f32_fwd / _bwd over the [pixels, C] view of the map. */
int zkb_avgpool2_f32_fwd(const float* x, int B, int H, int W, int C, float* out, void* stream);
int zkb_avgpool2_f32_bwd(const float* dout, int B, int H, int W, int C, float* dx, void* stream);
int zkb_bn_gain_add_f32(const float* x, long long rows, int C, const float* mean, const float* gain,
                        const float* beta, const float* shortcut, float* out, void* stream);
int zkb_bn_gain_bwd_apply_f32(const float* d, const float* x, long long rows, int C, const float* mean,
                              const float* rs_bwd, const float* gain, const float* mean_d,
                              const float* mean_dx, float* out, void* stream);
/* Head of the larq-zoo models: global average pooling over the HW positions of [B,HW,C] (C % 4 == 0)
 * and a float Dense layer y[M][ldy] = x[M][K] w[K][N] + bias (bias may be NULL); the backward
 * WRITES dw[K][N], dbias[N] (optional) and dx[M][K] (optional). */
int zkb_global_avgpool_fwd(const float* h, int B, int HW, int C, float* out, void* stream);
int zkb_global_avgpool_bwd(const float* dout, int B, int HW, int C, float* dh, void* stream);
int zkb_dense_f32_fwd(const float* x, const float* w_k_n, const float* bias, float* y, int M, int K,
                      int N, int ldy, void* stream);
int zkb_dense_f32_bwd(const float* x, const float* w_k_n, const float* dy, int lddy, float* dx,
                      float* dw_k_n, float* dbias, int M, int K, int N, void* stream);
/* Stride-2 binary 'same' convolutions (the down-sampling units): TensorFlow's padding rule --
 * out = ceil(in/2), total padding max((out-1)*2 + ksize - in, 0), the smaller half first -- goes into
 * the im2col tensor map (traversal stride 2, asymmetric corners), so the forward and the weight
 * gradient are the same GEMM kernels reading a strided window.  y [NB,ceil(H/2),ceil(W/2),Cout];
 * act / dw as for the stride-1 entry points.  The input gradient is the stride-1 zkb_bconv2d_dgrad
 * over the zero-inserted gradient map [NB,H,W,Cout] that zkb_dilate2_hi_lo writes. */
int zkb_bconv2d_fwd_i8_s2(const void* act_i8, const void* wq_oc_k_i8, void* y_i16, int NB, int H, int W,
                          int Cin, int Cout, int ksize, void* stream);
/* The same on bf16 +-1 operands (exact: products accumulate in f32 below 2^24), for channel counts
 * that are multiples of 64 but not of 128 -- the 64-channel first stage of Bi-Real-Net-18 -- which
 * the int8 forward (128 channels per shared-memory row) cannot read.  The stride-1 forward of such
 * layers is zkb_bconv2d_fwd; zkb_bconv2d_dgrad / _wgrad / _wgrad_s2 accept Cin, Cout % 64 == 0. */
int zkb_bconv2d_fwd_s2(const void* act_pm1, const void* wq_oc_k, void* y_i16, int NB, int H, int W, int Cin,
                       int Cout, int ksize, void* stream);
int zkb_bconv2d_wgrad_s2(const void* act_pm1, const void* dy_hi, const void* dy_lo, float* dw_hwio, int NB,
                         int H, int W, int Cin, int Cout, int ksize, void* stream);
int zkb_dilate2_hi_lo(const void* dy_hi, const void* dy_lo, int B, int H, int W, int C, int ksize,
                      void* out_hi, void* out_lo, void* stream);
/* MaxPool2D(pool_size=3, strides=2, padding="same") of the stem, float NHWC, C % 4 == 0:
 * out [B,ceil(H/2),ceil(W/2),C]; arg (same shape, uint8) = winning window slot kh*3+kw, first
 * maximum in row-major order; the backward gathers dout through arg into dx [B,H,W,C]. */
int zkb_maxpool3s2_f32_fwd(const float* x, int B, int H, int W, int C, float* out, uint8_t* arg,
                           void* stream);
int zkb_maxpool3s2_f32_bwd(const float* dout, const uint8_t* arg, int B, int H, int W, int C, float* dx,
                           void* stream);
/* The same pool for either padding and for the INTEGER conv outputs BinaryAlexNet pools between a
 * QuantConv2D and its BatchNormalization (larq-zoo; MaxPool2D(pool_size=3, strides=2), "valid":
 * out = (in - 3) / 2 + 1): x / out int16 (x_is_i16 = 1) or f32, same != 0 for "same" padding.  The
 * backward gathers (every input pixel sums the windows it won; overlapping windows need no atomics)
 * and writes either dx_f32 or the bf16 pair dx_hi / dx_lo the backward GEMMs read -- pass the other
 * as NULL. */
int zkb_maxpool3s2_fwd(const void* x, int x_is_i16, int B, int H, int W, int C, int same, void* out,
                       uint8_t* arg, void* stream);
int zkb_maxpool3s2_bwd(const float* dout, const uint8_t* arg, int B, int H, int W, int C, int same,
                       float* dx_f32, void* dx_hi, void* dx_lo, void* stream);
/* Float stem convolutions on the 3-channel image x f32 [B,H,W,3] with TensorFlow 'same' padding:
 * Conv2D(Cout, 7, strides=2) of Bi-Real-Net-18 and the 11x11 / stride-4 first layer of BinaryAlexNet
 * (pass the quantised kernel for a QuantConv2D).  w / dw: f32 HWIO [ksize,ksize,3,Cout], Cout % 64 == 0;
 * y / dy: f32 [B,ceil(H/stride),ceil(W/stride),Cout].  (ksize, stride) is (7, 2) or (11, 4).  The
 * weight gradient is ACCUMULATED into dw (the caller zeroes it); there is no input gradient.
 * FP32 FMA kernels on shared-memory patches (K = 147 / 363 on 12-byte pixels is not a TMA operand). */
int zkb_conv_stem_f32_fwd(const float* x, const float* w_hwio, float* y, int B, int H, int W, int Cin,
                          int Cout, int ksize, int stride, void* stream);
int zkb_conv_stem_f32_wgrad(const float* x, const float* dy, float* dw_hwio, int B, int H, int W, int Cin,
                            int Cout, int ksize, int stride, void* stream);

/* ---- for callers that cannot allocate device memory ----------------------------------------------
 * (No counterpart in the reference: this entry point exists for hosts of this library, not for
 * parity.)  BatchNorm(train) + sign in one call -- statistics, moving-average update, normalise,
 * sign -- i.e. zkb_bn_stats + zkb_bn_finalize + zkb_bn_apply_sign, but the per-channel accumulators
 * and statistics those need live in a fixed array compiled into the library instead of buffers the
 * caller provides.  Meant for code that is handed input/output buffers only and must not call
 * cudaMalloc itself: an operator callback inside a runtime that owns the device allocator, or a
 * loop that cannot take the device-wide synchronisation cudaMalloc/cudaFree imply.  The call
 * allocates nothing.
 *   - Limit: C <= zkb_bn_train_sign_static_max_channels(x_is_i16): 2048 channels for int16 input,
 *     1024 for f32 input, and C must be 8 (int16) or 4 (f32) times a power of two -- each thread of
 *     the 256-thread blocks owns one 16-byte vector of channels.  Anything else is refused with a
 *     negative status and a message naming the limit; nothing is launched.
 *   - Every GPU that loads the library holds the array (48 KiB) until the process exits.
 *   - The first call on each device looks up the array's device address (that device must be
 *     current); make that first call before starting any stream capture.
 *   - All calls on a GPU share the array.  It is cleared on entry, so back-to-back calls on one
 *     stream are fine; calls that may overlap on the GPU (other streams, other threads) race. */
int zkb_bn_train_sign_static_max_channels(int x_is_i16);
int zkb_bn_train_sign_static(const void* x, int x_is_i16, long long rows, int C, const float* beta,
                             float eps, float momentum, int bessel, float* moving_mean,
                             float* moving_var, void* s_out_bf16, float* a_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ZKB200_H */
