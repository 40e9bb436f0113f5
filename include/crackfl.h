/* crackfl C ABI: every sm_100a kernel of the framework behind a plain `extern "C"` launcher.
 *
 * Conventions
 *   - activations: NHWC bf16 (`void*`), channel counts multiples of 8;
 *   - parameters / gradients / optimiser state: fp32 views into the flat arena (Keras layouts);
 *   - every launcher enqueues on `stream` and returns 0 or a cudaError_t value; nothing synchronises;
 *   - no launcher allocates: all scratch is passed in by the caller.
 * The "ref:" notes name the reference site (file:line under the upstream repo) each kernel replaces.
 */
#ifndef CRACKFL_H_
#define CRACKFL_H_

#include <cuda_runtime.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* BatchNorm finalize folded into the consumer kernel (the *_fin launchers): scale/shift are derived from the
 * sum / sum-of-squares accumulators on the fly; block 0 publishes bn_out = [scale|shift|mean|invstd] and updates
 * the moving statistics.  Saves one launch per BN layer on the critical path. */
typedef struct {
  const float* stats;  /* [sum C][sumsq C] */
  const float* gamma;
  const float* beta;
  float* run_mean;
  float* run_var;
  float* bn_out;       /* [4C] */
  float count, eps, momentum;
} crackfl_bn_fin;

/* ---- entry block ------------------------------------------------------------------------- */
/* ref: client_fit_model.py:100 Conv2D(32,3,strides=2,"same") + :43 x/255.  uint8 NHWC in, bf16 out. */
int crackfl_entry_conv_fwd(const void* x_u8, const float* w, const float* bias, void* y, int B, int H, int W, int C,
                           cudaStream_t stream);
int crackfl_entry_conv_wgrad(const void* x_u8, const void* dy, float* dw, int B, int H, int W, int C, cudaStream_t stream);
/* second generation: col[B*H/2*W/2][32] bf16 = the /255-normalised 3x3 stride-2 patches (k = (kh*3+kw)*3+ci, k >= 27 zero);
 * the entry conv and its weight gradient then run as tcgen05 GEMMs over col (crackfl_conv_tc_fwd with K = 32,
 * crackfl_wgrad_tc_ws_cv with Cin = 32, cin_valid = 27). */
int crackfl_entry_im2col(const void* x_u8, void* col, int B, int H, int W, cudaStream_t stream);

/* ---- batch normalisation (ref: client_fit_model.py:101,110,114,130,134; eps 1e-3, momentum .99) ---- */
/* stats = [sum C][sumsq C]; bn = [scale C][shift C][mean C][invstd C]; sums = [sum dz C][sum dz*(y-mean) C] */
int crackfl_bn_stats(const void* y, float* stats, long P, int C, cudaStream_t stream);
int crackfl_bn_finalize(const float* stats, const float* gamma, const float* beta, float* run_mean, float* run_var,
                        float* bn, int C, float count, float eps, float momentum, cudaStream_t stream);
int crackfl_bn_eval_prepare(const float* gamma, const float* beta, const float* run_mean, const float* run_var, float* bn,
                            int C, float eps, cudaStream_t stream);
/* every BN layer's inference table in one launch: offsets are element offsets into the fp32 parameter arena / the bn table */
typedef struct {
  long gamma_off, beta_off, mean_off, var_off, bn_off;
  int C, pad_;
} crackfl_bn_eval_desc;
int crackfl_bn_eval_prepare_all(const float* arena, float* bn_params, const crackfl_bn_eval_desc* descs_dev, int n, float eps,
                                cudaStream_t stream);
int crackfl_bn_act(const void* y, const float* bn, void* out, void* even_or_null, int B, int H, int W, int C, int relu,
                   cudaStream_t stream);
/* mode: 0 dz=g, 1 dz=g*1[bn(y)>0], 2 dz=maxpool-backward gather of g through `arg` */
int crackfl_bn_bwd_reduce(const void* g, const void* arg, const void* y, const float* bn, float* sums, int B, int H, int W,
                          int C, int mode, cudaStream_t stream);
int crackfl_bn_bwd_apply(const void* g, const void* arg, const void* y, const float* bn, const float* sums, void* dy,
                         float* dgamma, float* dbeta, int B, int H, int W, int C, int mode, cudaStream_t stream);

/* Training-mode BatchNorm(+ReLU) forward for callers that cannot allocate device memory (a plug-in inside a
 * process whose device heap belongs to another framework, a fixed-budget step function, recovery code after an
 * out-of-memory error).  Same result as crackfl_bn_stats + crackfl_bn_finalize + crackfl_bn_act, but the 6*C
 * floats of scratch come from a fixed array inside the library (48 KiB per GPU, held until exit) instead of the
 * caller.  C <= CRACKFL_NOALLOC_MAX_CHANNELS, larger C is refused.  Calls share that array: do not let two calls
 * overlap on one GPU.  The first call per device resolves the array's address; make it before any stream capture.
 * No counterpart in the reference (it never manages device memory itself). */
#define CRACKFL_NOALLOC_MAX_CHANNELS 2048
int crackfl_bn_train_fwd_noalloc(const void* y, const float* gamma, const float* beta, float* run_mean, float* run_var,
                                 void* out, int B, int H, int W, int C, int relu, float eps, float momentum,
                                 cudaStream_t stream);
int crackfl_noalloc_max_channels(void);

int crackfl_bn_act_fin(const void* y, const crackfl_bn_fin* fin, void* out, void* even_or_null, int B, int H, int W, int C,
                       int relu, cudaStream_t stream);

/* ---- depthwise 3x3 (ref: SeparableConv2D, client_fit_model.py:109,113) -------------------- */
/* mode 0: a = relu(x); mode 1: a = relu(x*scale+shift) */
int crackfl_dw3x3_fwd(const void* x, const float* w, const float* bn, void* out, int B, int H, int W, int C, int mode,
                      cudaStream_t stream);
int crackfl_dw3x3_bwd(const void* g, const void* src, const float* w, const float* bn, const void* resid_t, void* dz,
                      float* dwk, int B, int H, int W, int C, int mode, int resid, cudaStream_t stream);

/* the same backward split in two launches (input gradient / weight gradient) for concurrent streams; W % 4 == 0 */
int crackfl_dw3x3_dgrad(const void* g, const void* src, const float* w, const float* bn, const void* resid_t, void* dz, int B,
                        int H, int W, int C, int mode, int resid, cudaStream_t stream);
int crackfl_dw3x3_wgrad(const void* g, const void* src, const float* bn, float* dwk, int B, int H, int W, int C, int mode,
                        cudaStream_t stream);

int crackfl_dw3x3_fwd_fin(const void* x, const float* w, const crackfl_bn_fin* fin, void* out, int B, int H, int W, int C,
                          cudaStream_t stream); /* a = relu(bn(x)), W % 4 == 0 */

/* second generation: TMA-staged halo tiles (csrc/dw_tma.cu).  Forward: src = NULL, flip = 0, act_mode 0 relu(x) /
 * 1 relu(bn(x)) (bn pointer or fused-finalize struct).  Input gradient: x = g, act_mode = 2, flip = 1, src = forward
 * input (src_mode 0: mask = src > 0, 1: mask = bn(src) > 0), resid as crackfl_dw3x3_bwd.  C in {32, 64, 128, ...}. */
int crackfl_dw3x3_tma(const void* x, const float* w, const float* bn, const crackfl_bn_fin* fin_or_null, void* out,
                      const void* src_or_null, const float* src_bn, const void* resid_t, int B, int H, int W, int C,
                      int act_mode, int flip, int src_mode, int resid, cudaStream_t stream);
int crackfl_dw3x3_wgrad_tma(const void* g, const void* src, const float* bn, float* dwk, int B, int H, int W, int C, int mode,
                            cudaStream_t stream);
/* third generation: input gradient and weight gradient in one pass over g and src (warps 0-7 / 8-15 of a 512-thread CTA) */
int crackfl_dw3x3_bwd_tma(const void* g, const void* src, const float* w, const float* src_bn, const void* resid_t, void* dz,
                          float* dwk, int B, int H, int W, int C, int src_mode, int resid, cudaStream_t stream);
/* the same + the BN-backward reductions of the BN behind src (red_bn = [scale][shift][mean][invstd], red_sums pre-zeroed) */
int crackfl_dw3x3_bwd_tma_bn(const void* g, const void* src, const float* w, const float* src_bn, const void* resid_t, void* dz,
                             float* dwk, int B, int H, int W, int C, int src_mode, int resid, const float* red_bn,
                             float* red_sums, cudaStream_t stream);

/* ---- block tails --------------------------------------------------------------------------- */
/* ref: client_fit_model.py:116-123 MaxPooling2D(3,2,"same") + add(residual) */
int crackfl_pool_add_fwd(const void* y, const float* bn, const void* res, void* out, void* even_or_null,
                         void* relu_out_or_null, void* argmax_u8, int B, int H, int W, int C, cudaStream_t stream);
/* ref: client_fit_model.py:136-142 UpSampling2D(2) x2 + Conv2D 1x1 + add, evaluated at low resolution */
int crackfl_dec_combine_fwd(const void* y, const float* bn, const void* rl, void* s, void* a_next_or_null, int B, int H,
                            int W, int C, int rl_up, cudaStream_t stream);
int crackfl_pool_add_fwd_fin(const void* y, const crackfl_bn_fin* fin, const void* res, void* out, void* even_or_null,
                             void* relu_out_or_null, void* argmax_u8, int B, int H, int W, int C, cudaStream_t stream);
int crackfl_dec_combine_fwd_fin(const void* y, const crackfl_bn_fin* fin, const void* rl, void* s, void* a_next_or_null, int B,
                                int H, int W, int C, int rl_up, cudaStream_t stream);
/* max-pool backward: scatter-free gather of g [B,H/2,W/2,C] through the arg-max map into dz [B,H,W,C] */
int crackfl_pool_bwd(const void* g, const void* argmax_u8, void* dz, int B, int H, int W, int C, cudaStream_t stream);
/* the same plus the BN-backward reductions of the layer below (sums = [sum dz C][sum dz*(y-mean) C], pre-zeroed), one pass */
int crackfl_pool_bwd_bn(const void* g, const void* argmax_u8, const void* y, const float* bn, void* dz, float* sums, int B,
                        int H, int W, int C, cudaStream_t stream);
int crackfl_down2sum(const void* in, const void* mask_or_null, const void* add_or_null, void* out, int B, int Ho, int Wo,
                     int C, cudaStream_t stream);
/* crackfl_down2sum + the two launches that follow it in the decoder's backward pass: sums[c] += sum out, sums[C + c] +=
   sum out * (y - mean[c]) (= crackfl_bn_bwd_reduce(out, y, bn, mode 0); sums pre-zeroed) and, when lowres != NULL,
   lowres = 2x2 sum of out (= crackfl_down2sum(out)).  Ho, Wo even. */
int crackfl_down2sum_bn(const void* in, const void* mask_or_null, const void* add_or_null, void* out, void* lowres_or_null,
                        const void* y, const float* bn, float* sums, int B, int Ho, int Wo, int C, cudaStream_t stream);
int crackfl_relu_bwd_add(const void* g, const void* x, const void* add_or_null, void* out, long numel, cudaStream_t stream);
int crackfl_colsum(const void* g, float* out, long P, int C, cudaStream_t stream);

/* ---- head: Conv2D(1,1x1,sigmoid) + binary_crossentropy (+Dice ext.) + accuracy (ref: :145,157) ---- */
/* acc = [bce_sum, correct, intersection, prob_sum, mask_sum, pixel_count] (accumulated; caller zeroes)
   bce_clip = 0: cross-entropy from the logit, the form Keras 2 `fit` uses for a sigmoid output layer (the sigmoid's input is
   recovered, tf.nn.sigmoid_cross_entropy_with_logits): gradient sigmoid(z)-y for every z.
   bce_clip = 1: probabilities clipped to [1e-7, 1-1e-7] with no gradient outside the clip (backend fall-back / Keras 3). */
int crackfl_head_fwd(const void* s, const float* w, const float* bias, const void* mask_u8_or_null, float* acc,
                     float* prob_or_null, int B, int Hl, int Wl, int C, int bce_clip, cudaStream_t stream);
int crackfl_head_bwd(const void* s, const float* w, const float* bias, const void* mask_u8, const float* sums, void* ds,
                     float* dw, float* db, int B, int Hl, int Wl, int C, int use_dice, float dice_smooth, float loss_scale,
                     int bce_clip, cudaStream_t stream);
/* plain-BCE training step: forward metrics (acc_out[6], pre-zeroed) + backward in one pass over s */
int crackfl_head_fwd_bwd(const void* s, const float* w, const float* bias, const void* mask_u8, float* acc_out, void* ds,
                         float* dw, float* db, int B, int Hl, int Wl, int C, float loss_scale, int bce_clip,
                         cudaStream_t stream);

/* ---- GEMM-shaped convolutions ---------------------------------------------------------------- */
/* 3x3 conv whose input is activated inside the kernel (K1: ReLU -> Conv2DTranspose -> BN of client_fit_model.py:128-134
 * without a BN-apply pass): xmode 1 = relu(X), 2 = relu(scale*X + shift) with scale/shift from bn[2*Cin] or derived from
 * the statistics in *fin (block 0 publishes table + moving statistics).  side_or_null: the activated input is also
 * written (B,H,W,Cin).  diag != 0: depthwise 3x3 on the tensor cores (client_fit_model.py:109,113), Wk = block-diagonal
 * tiles from crackfl_dw_diag_prep, N == Cin.  H >= 16, W >= 8. */
int crackfl_conv_tc_xf(const void* X, const void* Wk, const float* bias_or_null, const void* add_or_null, void* Y,
                       float* stats_or_null, int B, int H, int W, int Cin, int N, int xmode, const float* bn_or_null,
                       const crackfl_bn_fin* fin_or_null, void* side_or_null, int diag, cudaStream_t stream);
/* K2: input gradient of a 3x3 conv with the BatchNorm backward of its input folded in (ref: the backward of
 * BatchNormalization -> [ReLU] in front of a Conv2DTranspose, client_fit_model.py:128-134): dy = BN-backward(dz, y) is formed
 * inside the kernel from bn = [scale|shift|mean|invstd] and sums = [sum dz | sum dz*(y-mean)] (crackfl_bn_bwd_reduce), written to
 * dy_out for the weight gradient, and dgamma / dbeta are written too.  Replaces crackfl_bn_bwd_apply + crackfl_conv_tc_fwd. */
int crackfl_conv_tc_bnbwd(const void* dz, const void* y, const void* Wk, void* Y, int B, int H, int W, int Cin, int N,
                          const float* bn, const float* sums, int relu_mask, void* dy_out, float* dgamma, float* dbeta,
                          cudaStream_t stream);
typedef struct {
  long src_off, dst_off; /* element offsets: depthwise kernel (3,3,C,1) in the fp32 arena / tiles in the bf16 operand buffer */
  int C, kc, flip, pad_;  /* kc = 64 (C % 64 == 0) or 32; flip: taps reversed (input-gradient form) */
} crackfl_dw_diag_desc;
int crackfl_dw_diag_prep(const float* arena, void* ops, const crackfl_dw_diag_desc* descs_dev, int num_desc, cudaStream_t stream);

/* CUDA-core implicit GEMM (cross-check + odd shapes).  Wk[N][taps*Cin] bf16, taps in {1,9}. */
int crackfl_conv_simt_fwd(const void* X, const void* Wk, const float* bias_or_null, const void* add_or_null, void* Y, int B,
                          int H, int W, int Cin, int N, int taps, cudaStream_t stream);
int crackfl_wgrad_simt(const void* X, const void* dY, float* dW, int B, int H, int W, int Cin, int Cout, int taps, long s_tap,
                       long s_ci, long s_co, cudaStream_t stream);
/* tcgen05 / TMEM / TMA implicit GEMM (ref: Conv2DTranspose :129,133, SeparableConv2D pointwise :109,113,
 * residual Conv2D :119,140 and their dgrad).  stats_or_null: fused BN sum/sumsq epilogue. */
int crackfl_conv_tc_fwd(const void* X, const void* Wk, const float* bias_or_null, const void* add_or_null, void* Y,
                        float* stats_or_null, int B, int H, int W, int Cin, int N, int taps, cudaStream_t stream);
/* tcgen05 weight gradient with MN-major operands, split over pixels, fp32 atomics into the arena. */
int crackfl_wgrad_tc(const void* X, const void* dY, float* dW, int B, int H, int W, int Cin, int Cout, int taps, long s_tap,
                     long s_ci, long s_co, cudaStream_t stream);
/* same with a caller-provided workspace (>= groups*ntiles*splits*128*BN floats, 32 MB covers every layer of the
 * model): per-CTA partial tiles + a reduction pass instead of contended fp32 atomics (about 2x faster). */
int crackfl_wgrad_tc_ws(const void* X, const void* dY, float* dW, int B, int H, int W, int Cin, int Cout, int taps,
                        long s_tap, long s_ci, long s_co, float* workspace, long workspace_floats, cudaStream_t stream);
/* as crackfl_wgrad_tc_ws; gradient rows of the zero-padded input channels ci >= cin_valid are not written */
int crackfl_wgrad_tc_ws_cv(const void* X, const void* dY, float* dW, int B, int H, int W, int Cin, int Cout, int taps,
                           long s_tap, long s_ci, long s_co, float* workspace, long workspace_floats, int cin_valid,
                           cudaStream_t stream);
/* deferred split reduction: every wgrad GEMM of the step writes its per-CTA partial tiles into its own slice of one
 * workspace and ONE launch at the end of the backward pass reduces them all into the gradient arena. */
typedef struct {
  long ws_off;  /* float offset of the layer's partial tiles in the workspace (caller fills) */
  long dw_off;  /* float offset of the gradient tensor (incl. any tap base) in the gradient arena (caller fills) */
  long s_tap, s_ci, s_co;
  long need;    /* workspace floats the layer uses */
  int groups, ntn, splits, BN, cw, mblk, total_blocks, cchunks, cin_valid, zl;
} crackfl_wgrad_red_desc;
int crackfl_wgrad_plan(int B, int H, int W, int Cin, int Cout, int taps, long s_tap, long s_ci, long s_co, int cin_valid,
                       crackfl_wgrad_red_desc* desc_host);
int crackfl_wgrad_tc_deferred(const void* X, const void* dY, int B, int H, int W, int Cin, int Cout, int taps, float* ws_layer,
                              long ws_floats, cudaStream_t stream);
int crackfl_wgrad_reduce_all(const crackfl_wgrad_red_desc* descs_dev, int n, const float* ws_base, float* dw_base,
                             cudaStream_t stream);
int crackfl_tc_selftest_supported(void);

/* ---- optimiser / weight re-layout ------------------------------------------------------------ */
typedef struct {
  long src_off; /* element offset into the fp32 arena (already includes any flip base) */
  long dst_off; /* element offset into the bf16 operand buffer */
  int N, T, K;  /* out[dst + (n*T + t)*K + k] = arena[src + n*s_n + t*s_t + k*s_k] */
  int pad_; /* row stride of the operand when K is padded up (0: = K); pad columns are never written */
  long s_n, s_t, s_k;
} crackfl_prep_desc;
int crackfl_weight_prep(const float* arena, void* ops_bf16, const crackfl_prep_desc* descs_dev, int num_desc,
                        cudaStream_t stream);
/* same result, one 32x32 tile per block: tile_prefix_dev[i] = number of 32x32 tiles (ceil(N/32) * ceil(K/32) * T) of the
   operands before i, tile_prefix_dev[num_desc] = total_tiles; num_desc <= 128 */
int crackfl_weight_prep_flat(const float* arena, void* ops_bf16, const crackfl_prep_desc* descs_dev, const int* tile_prefix_dev,
                             int num_desc, int total_tiles, cudaStream_t stream);
/* ref: client_fit_model.py:157 optimizer="Adam" (Keras form, eps 1e-7); also zeroes g */
int crackfl_adam_step(float* p, float* g, float* m, float* v, int* step_dev, long n, float lr, float b1, float b2, float eps,
                      cudaStream_t stream);
/* Adam over a sub-range, launched before the step's crackfl_adam_step_tail (which covers the rest and ticks the counter):
   t = *step + 1, the counter is not written.  n % 4 == 0, pointers 16-byte aligned. */
int crackfl_adam_part(float* p, float* g, float* m, float* v, const int* step_dev, long n, float lr, float b1, float b2, float eps,
                      cudaStream_t stream);
int crackfl_f32_to_bf16(const float* in, void* out, long n, cudaStream_t stream);
/* total[i] += step[i], i < n <= 32 (per-epoch metric accumulation of the per-step loss/accuracy sums) */
/* Adam + step-count tick + metric accumulation (metrics_total[i] += step_acc[i]) in one launch; *tail_counter == 0 on entry.
   zero_base != NULL: the last block also saves step_acc to last_acc[nacc] and clears zero_n floats at zero_base (the
   per-step scratch region), so the step needs no fill launch. */
int crackfl_adam_step_tail(float* p, float* g, float* m, float* v, int* step_dev, long n, float lr, float b1, float b2, float eps,
                           unsigned* tail_counter, float* metrics_total, const float* step_acc, int nacc, float* last_acc,
                           float* zero_base, int zero_n, cudaStream_t stream);
int crackfl_accum_small(float* total, const float* step, int n, cudaStream_t stream);
/* staging of a client update as a bf16 delta against the global model (ref: fl_client.py:61-65 upload) */
int crackfl_publish_delta(const float* local, const float* global, void* out_bf16, long n, cudaStream_t stream);

/* ---- synthetic data + measurement --------------------------------------------------------------- */
/* ref: dataset of client_fit_model.py:58-59 is not shipped; crack-like uint8 images (B,H,W,3) + masks (B,H,W) */
int crackfl_synth_cracks(void* img_u8, void* mask_u8, int B, int H, int W, unsigned seed, cudaStream_t stream);
/* ref: Predict/contour, client_fit_model.py:209-221 + test/Segmentation2.py:118-124 (device half):
 * out4[b] = {foreground area, boundary pixels, component seeds, sum of probabilities} at threshold (127) */
int crackfl_measure_masks(const float* prob, float* out4, int B, int H, int W, int threshold, cudaStream_t stream);

/* ---- fused federated aggregation (ref: fl_server.py:92-102 + the gRPC gather/broadcast/poll) ---- */
#define CRACKFL_MAX_CLIENTS 128
#define CRACKFL_MAX_OUT 16
#define CRACKFL_MAX_RANKS 16
typedef struct {
  const void* client[CRACKFL_MAX_CLIENTS]; /* update arenas, local or peer-mapped */
  float coef[CRACKFL_MAX_CLIENTS];         /* n_k / sum n (or 1/K) */
  void* out[CRACKFL_MAX_OUT];              /* destination arenas for the new global model */
  int num_clients, num_out;
  int payload_bf16, out_bf16;
  int delta;      /* payload is w_k - w_global */
  int server_opt; /* 0 FedAvg, 1 FedAdam */
  float eta, beta1, beta2, tau;
  float* master;            /* fp32 global model (NULL: pure weighted sum, e.g. per-GPU pre-reduce) */
  float* m;                 /* FedAdam moments */
  float* v;
  const uint8_t* trainable; /* per-element mask for FedAdam (BN moving stats are plain-averaged) */
  float* mc_out;            /* NVLS multicast address of the destination arena (fp32) or NULL */
  long begin, end;          /* this rank's shard [begin,end), multiples of 4 elements */
  uint32_t* pads[CRACKFL_MAX_RANKS]; /* signal pads of every rank: [2][CRACKFL_MAX_RANKS] u32 */
  int rank, world;
  uint32_t epoch;           /* 1,2,3,... monotonically per call */
  int* error_flag;          /* set non-zero on spin timeout */
  uint32_t* local_sync;     /* [2] u32 in local memory, zero-initialised once */
  const void* mc_in;        /* NVLS multicast address of the (pre-scaled) partial arena: multimem.ld_reduce pull, or NULL */
  int unroll;               /* independent 16-byte loads per client per thread: 4 (default when 0) or 8 */
  int load_mode;            /* 0: ld.relaxed.sys (default), 1: ld.global.nc.L1::no_allocate (updates are immutable during the launch) */
  long long timeout_cycles; /* spin limit of the cross-GPU barriers in SM clocks; 0 = default (~60 s) */
} crackfl_fedagg_params;
int crackfl_fedagg(const crackfl_fedagg_params* p, int grid_blocks, cudaStream_t stream);
int crackfl_ipc_export(void* dev_ptr, void* handle_out64);
int crackfl_ipc_open(const void* handle64, void** dev_ptr_out);
int crackfl_ipc_close(void* dev_ptr);
int crackfl_enable_peer_access(int peer_device);
/* whole-allocation device buffers (zero-filled) whose base address can be exported with crackfl_ipc_export */
int crackfl_dev_malloc(void** out, size_t bytes);
int crackfl_dev_free(void* p);

#ifdef __cplusplus
}
#endif
#endif /* CRACKFL_H_ */
