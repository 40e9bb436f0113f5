// Elementwise / reduction kernels of the U-Net train step (NHWC bf16, fp32 math).
// Replaces the implicit TF ops O1,O2,O4-O6,O8,O10,O12-O14 and their backward (O16) of
// client_fit_model.py:100-145,157.
// Every kernel is HBM-bound; the design goal is one pass per tensor with the neighbouring
// pointwise work (BN-apply, ReLU, residual add, nearest upsample, /255) fused in.
#include <cstdio>
#include <cstdlib>

#include <mutex>
#include <unordered_map>
#include "common.cuh"
#include "../include/crackfl.h"

using namespace ck;

namespace {

constexpr int TPB = 256;

// BatchNorm "finalize" folded into the consumer of the normalised tensor: the consumer derives scale/shift for its
// own channel vector straight from the sum / sum-of-squares accumulators (a handful of flops, once per thread) and
// block 0 publishes [scale, shift, mean, invstd] for the backward pass and updates the moving statistics.  This
// removes one tiny launch (and its dependency bubble) per BN layer from the critical path of the step.
typedef crackfl_bn_fin BnFin;

// Flat vector index -> (pixel, w, h, b).  Channel-vector counts are powers of two and so are the spatial sizes of the
// 128 / 256-pixel configurations: shifts and masks then replace three divide sequences (~25 instructions each, or ~100 for
// the 64-bit ones the loss head used) per element; any other size takes the division path.
struct IdxDec {
  int lc, lw, lh;  // log2 of c8, W, H or -1
};
__device__ __forceinline__ int ck_log2_or_neg(int v) { return (v > 0 && (v & (v - 1)) == 0) ? (__ffs(v) - 1) : -1; }
__device__ __forceinline__ IdxDec make_idxdec(int c8, int Wd, int Hd) {
  IdxDec d;
  d.lc = ck_log2_or_neg(c8); d.lw = ck_log2_or_neg(Wd); d.lh = ck_log2_or_neg(Hd);
  if (d.lc < 0 || d.lw < 0 || d.lh < 0) d.lc = -1;
  return d;
}
__device__ __forceinline__ void idx_dec(const IdxDec& d, int g, int c8, int Wd, int Hd, int& pix, int& w, int& h, int& b) {
  if (d.lc >= 0) {
    pix = g >> d.lc;
    w = pix & (Wd - 1);
    const int q = pix >> d.lw;
    h = q & (Hd - 1);
    b = q >> d.lh;
  } else {
    pix = g / c8;
    w = pix % Wd;
    const int q = pix / Wd;
    h = q % Hd;
    b = q / Hd;
  }
}

__device__ __forceinline__ void bn_coeffs(const BnFin& f, const float* __restrict__ bn, int C, int cv, float sc[8],
                                          float sh[8]) {
  if (f.stats == nullptr) {
    load8f(bn + cv * 8, sc);
    load8f(bn + C + cv * 8, sh);
    return;
  }
  float s1[8], s2[8], g[8], b[8];
  load8f(f.stats + cv * 8, s1); load8f(f.stats + C + cv * 8, s2);
  load8f(f.gamma + cv * 8, g); load8f(f.beta + cv * 8, b);
  const float inv = 1.0f / f.count;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float mean = s1[j] * inv;
    const float var = fmaxf(s2[j] * inv - mean * mean, 0.f);
    const float invstd = rsqrtf(var + f.eps);
    sc[j] = g[j] * invstd;
    sh[j] = b[j] - mean * sc[j];
  }
}

__device__ __forceinline__ void bn_publish(const BnFin& f, int C) {
  if (f.stats == nullptr || blockIdx.x != 0) return;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float mean = f.stats[c] / f.count;
    const float var = fmaxf(f.stats[C + c] / f.count - mean * mean, 0.f);
    const float invstd = rsqrtf(var + f.eps);
    const float scl = f.gamma[c] * invstd;
    f.bn_out[c] = scl;
    f.bn_out[C + c] = f.beta[c] - mean * scl;
    f.bn_out[2 * C + c] = mean;
    f.bn_out[3 * C + c] = invstd;
    const float unb = f.count > 1.f ? var * (f.count / (f.count - 1.f)) : var;
    f.run_mean[c] = f.run_mean[c] * f.momentum + mean * (1.f - f.momentum);
    f.run_var[c] = f.run_var[c] * f.momentum + unb * (1.f - f.momentum);
  }
}

// ---------------------------------------------------------------------------------------
// K4 entry conv 3x3 s2, TF-"same" (pad bottom/right only), uint8 input, /255 fused.  K=27 is too
// small for the tensor cores to pay; CUDA-core kernel, weights in shared memory.
__global__ void __launch_bounds__(TPB) entry_conv_fwd_k(const uint8_t* __restrict__ x, const float* __restrict__ w,
                                                        const float* __restrict__ bias, __nv_bfloat16* __restrict__ y,
                                                        int B, int H, int W, int C) {
  ck::pdl_enter();
  extern __shared__ float sw[];  // [27][C] + bias[C]
  for (int i = threadIdx.x; i < 27 * C; i += blockDim.x) sw[i] = w[i] * (1.0f / 255.0f);
  for (int i = threadIdx.x; i < C; i += blockDim.x) sw[27 * C + i] = bias[i];
  __syncthreads();
  const int Ho = H >> 1, Wo = W >> 1, c8 = C >> 3;
  const long total = (long)B * Ho * Wo * c8;
  for (long g = (long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long)gridDim.x * blockDim.x) {
    const int cv = (int)(g % c8);
    long p = g / c8;
    const int ow = (int)(p % Wo); p /= Wo;
    const int oh = (int)(p % Ho);
    const int b = (int)(p / Ho);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = sw[27 * C + cv * 8 + j];
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int ih = 2 * oh + kh;
      if (ih >= H) continue;
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int iw = 2 * ow + kw;
        if (iw >= W) continue;
        const uint8_t* px = x + ((long)(b * H + ih) * W + iw) * 3;
#pragma unroll
        for (int ci = 0; ci < 3; ++ci) {
          const float v = (float)px[ci];
          const float* wr = sw + ((kh * 3 + kw) * 3 + ci) * C + cv * 8;
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] = fmaf(v, wr[j], acc[j]);
        }
      }
    }
    stg16(y + g * 8, pack8(acc));
  }
}

// wgrad of the entry conv: dW[27][C] += sum_p patch(p)[27] * dy[p][C].  blockDim = (C/8, 27).
__global__ void entry_conv_wgrad_k(const uint8_t* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                   float* __restrict__ dw, int B, int H, int W, int C, int pix_per_block) {
  ck::pdl_enter();
  const int Ho = H >> 1, Wo = W >> 1, c8 = C >> 3;
  const int cv = threadIdx.x, k = threadIdx.y;
  const int kh = k / 9, kw = (k / 3) % 3, ci = k % 3;
  const long P = (long)B * Ho * Wo;
  const long p0 = (long)blockIdx.x * pix_per_block;
  const long p1 = min(P, p0 + pix_per_block);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (long p = p0; p < p1; ++p) {
    const int ow = (int)(p % Wo);
    const int oh = (int)((p / Wo) % Ho);
    const int b = (int)(p / ((long)Wo * Ho));
    const int ih = 2 * oh + kh, iw = 2 * ow + kw;
    if (ih >= H || iw >= W) continue;
    const float v = (float)x[((long)(b * H + ih) * W + iw) * 3 + ci] * (1.0f / 255.0f);
    float g[8];
    unpack8(ldg16(dy + (p * c8 + cv) * 8), g);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = fmaf(v, g[j], acc[j]);
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) atomicAdd(dw + (long)k * C + cv * 8 + j, acc[j]);
}

// ---------------------------------------------------------------------------------------
// BN statistics (forward): per-channel sum and sum of squares.  stats = [sum C][sumsq C].
__global__ void __launch_bounds__(TPB) bn_stats_k(const __nv_bfloat16* __restrict__ y, float* __restrict__ stats,
                                                  long P, int C, int pix_per_block) {
  ck::pdl_enter();
  __shared__ float sm[TPB * 8];
  const int c8 = C >> 3, lanes = TPB / c8;
  const int cv = threadIdx.x % c8, lane = threadIdx.x / c8;
  const long p0 = (long)blockIdx.x * pix_per_block;
  const long p1 = min(P, p0 + pix_per_block);
  float s[8] = {0, 0, 0, 0, 0, 0, 0, 0}, q[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (long p = p0 + lane; p < p1; p += lanes) {
    float f[8];
    unpack8(ldg16(y + (p * c8 + cv) * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j] += f[j]; q[j] = fmaf(f[j], f[j], q[j]); }
  }
  block_channel_atomic(s, c8, sm, stats);
  block_channel_atomic(q, c8, sm, stats + C);
}

// stats -> (scale, shift, mean, invstd) and the moving-average update (momentum 0.99, eps 1e-3,
// unbiased variance into the running estimate as TF's fused batch norm does).
// bn = [scale C][shift C][mean C][invstd C]
__global__ void bn_finalize_k(const float* __restrict__ stats, const float* __restrict__ gamma,
                              const float* __restrict__ beta, float* __restrict__ run_mean,
                              float* __restrict__ run_var, float* __restrict__ bn, int C, float count, float eps,
                              float momentum) {
  ck::pdl_enter();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float mean = stats[c] / count;
  float var = stats[C + c] / count - mean * mean;
  var = fmaxf(var, 0.f);
  const float invstd = rsqrtf(var + eps);
  const float sc = gamma[c] * invstd;
  bn[c] = sc;
  bn[C + c] = beta[c] - mean * sc;
  bn[2 * C + c] = mean;
  bn[3 * C + c] = invstd;
  const float unb = count > 1.f ? var * (count / (count - 1.f)) : var;
  run_mean[c] = run_mean[c] * momentum + mean * (1.f - momentum);
  run_var[c] = run_var[c] * momentum + unb * (1.f - momentum);
}

// inference: scale/shift from the moving statistics
__global__ void bn_eval_prepare_k(const float* __restrict__ gamma, const float* __restrict__ beta,
                                  const float* __restrict__ run_mean, const float* __restrict__ run_var,
                                  float* __restrict__ bn, int C, float eps) {
  ck::pdl_enter();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float invstd = rsqrtf(run_var[c] + eps);
  const float sc = gamma[c] * invstd;
  bn[c] = sc;
  bn[C + c] = beta[c] - run_mean[c] * sc;
  bn[2 * C + c] = run_mean[c];
  bn[3 * C + c] = invstd;
}

// the same for every BN layer of the model in ONE launch (blockIdx.y = layer): the eval / predict forward used to spend 15
// launches of ~4.5 us on these tables
__global__ void bn_eval_prepare_all_k(const float* __restrict__ arena, float* __restrict__ bn_params,
                                      const crackfl_bn_eval_desc* __restrict__ descs, float eps) {
  ck::pdl_enter();
  const crackfl_bn_eval_desc d = descs[blockIdx.y];
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < d.C; c += gridDim.x * blockDim.x) {
    const float rm = arena[d.mean_off + c];
    const float invstd = rsqrtf(arena[d.var_off + c] + eps);
    const float sc = arena[d.gamma_off + c] * invstd;
    float* bn = bn_params + d.bn_off;
    bn[c] = sc;
    bn[d.C + c] = arena[d.beta_off + c] - rm * sc;
    bn[2 * d.C + c] = rm;
    bn[3 * d.C + c] = invstd;
  }
}

// out = [relu](y*scale+shift); optional compact copy of the even pixels (input of the 1x1 s2 residual conv)
__global__ void __launch_bounds__(TPB, 4) bn_act_k(const __nv_bfloat16* __restrict__ y, const float* __restrict__ bn,
                                                   __nv_bfloat16* __restrict__ out, __nv_bfloat16* __restrict__ even,
                                                   int B, int H, int W, int C, int relu, const BnFin fin) {
  ck::pdl_enter();
  const int c8 = C >> 3;
  const int total = (int)B * H * W * c8;
  const int stride = (int)gridDim.x * blockDim.x;  // multiple of c8: a thread keeps its channel vector
  const int g0 = (int)blockIdx.x * blockDim.x + threadIdx.x;
  const int cv = (int)(g0 % c8);
  float sc[8], sh[8];
  bn_coeffs(fin, bn, C, cv, sc, sh);
  bn_publish(fin, C);
  for (int g = g0; g < total; g += 2 * stride) {
    const int g2 = g + stride;
    const bool has2 = g2 < total;
    float f[2][8];
    unpack8(ldg16(y + g * 8), f[0]);
    if (has2) unpack8(ldg16(y + g2 * 8), f[1]);
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !has2) break;
      const int gi = u ? g2 : g;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        f[u][j] = fmaf(f[u][j], sc[j], sh[j]);
        if (relu) f[u][j] = fmaxf(f[u][j], 0.f);
      }
      const uint4 v = pack8(f[u]);
      stg16(out + gi * 8, v);
      if (even) {
        int p, w, h, b;
        idx_dec(make_idxdec(c8, W, H), gi, c8, W, H, p, w, h, b);
        if (!(h & 1) && !(w & 1))
          stg16(even + ((((int)b * (H >> 1) + (h >> 1)) * (W >> 1) + (w >> 1)) * c8 + cv) * 8, v);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// K5 depthwise 3x3 s1 same.  mode 0: a = relu(x); mode 1: a = relu(x*scale+shift) (BN-apply+ReLU of
// the producer fused into the load).  Zero padding applies to `a`.
__device__ __forceinline__ void act_load(const __nv_bfloat16* __restrict__ src, long vec_idx, int mode,
                                         const float sc[8], const float sh[8], float a[8]) {
  unpack8(ldg16(src + vec_idx * 8), a);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (mode == 1) a[j] = fmaf(a[j], sc[j], sh[j]);
    a[j] = fmaxf(a[j], 0.f);
  }
}

__global__ void __launch_bounds__(TPB) dw3x3_fwd_k(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w,
                                                   const float* __restrict__ bn, __nv_bfloat16* __restrict__ out,
                                                   int B, int H, int W, int C, int mode) {
  ck::pdl_enter();
  const int c8 = C >> 3;
  const long total = (long)B * H * W * c8;
  for (long g = (long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long)gridDim.x * blockDim.x) {
    const int cv = (int)(g % c8);
    long p = g / c8;
    const int ww = (int)(p % W); p /= W;
    const int hh = (int)(p % H);
    const int b = (int)(p / H);
    float sc[8], sh[8], acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (mode == 1) { load8f(bn + cv * 8, sc); load8f(bn + C + cv * 8, sh); }
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int ih = hh + kh - 1;
      if (ih < 0 || ih >= H) continue;
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int iw = ww + kw - 1;
        if (iw < 0 || iw >= W) continue;
        float a[8], wk[8];
        act_load(x, (((long)b * H + ih) * W + iw) * c8 + cv, mode, sc, sh, a);
        load8f(w + (kh * 3 + kw) * C + cv * 8, wk);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(a[j], wk[j], acc[j]);
      }
    }
    stg16(out + g * 8, pack8(acc));
  }
}

// Backward of the depthwise conv w.r.t. its (activated) input and its weights.
//   da[p] = sum_tap g[p - off(tap)] * w[tap];   dz[p] = 1[a[p] > 0] * da[p]  (+ residual term)
//   dw[tap][c] += sum_o a[o + off(tap)][c] * g[o][c]
// resid: 0 none; 1 dz += T[h/2,w/2] on even pixels (outside the mask); 2 same but inside the mask.
__global__ void __launch_bounds__(TPB) dw3x3_bwd_k(const __nv_bfloat16* __restrict__ g, const __nv_bfloat16* __restrict__ src,
                                                   const float* __restrict__ w, const float* __restrict__ bn,
                                                   const __nv_bfloat16* __restrict__ resid_t,
                                                   __nv_bfloat16* __restrict__ dz, float* __restrict__ dwk, int B, int H,
                                                   int W, int C, int mode, int resid, int pix_per_block) {
  ck::pdl_enter();
  __shared__ float sm[TPB * 8];
  const int c8 = C >> 3, lanes = TPB / c8;
  const int cv = threadIdx.x % c8, lane = threadIdx.x / c8;
  const long P = (long)B * H * W;
  const long p0 = (long)blockIdx.x * pix_per_block;
  const long p1 = min(P, p0 + pix_per_block);
  float sc[8], sh[8], wk[9][8], dacc[9][8];
  if (mode == 1) { load8f(bn + cv * 8, sc); load8f(bn + C + cv * 8, sh); }
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    load8f(w + t * C + cv * 8, wk[t]);
#pragma unroll
    for (int j = 0; j < 8; ++j) dacc[t][j] = 0.f;
  }
  for (long p = p0 + lane; p < p1; p += lanes) {
    const int ww = (int)(p % W);
    const int hh = (int)((p / W) % H);
    const int b = (int)(p / ((long)W * H));
    float gc[8], ac[8], da[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    unpack8(ldg16(g + (p * c8 + cv) * 8), gc);
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int t = kh * 3 + kw;
        // weight gradient: a at (hh+kh-1, ww+kw-1) times g at (hh, ww)
        const int ih = hh + kh - 1, iw = ww + kw - 1;
        if (ih >= 0 && ih < H && iw >= 0 && iw < W) {
          float a[8];
          act_load(src, (((long)b * H + ih) * W + iw) * c8 + cv, mode, sc, sh, a);
#pragma unroll
          for (int j = 0; j < 8; ++j) dacc[t][j] = fmaf(a[j], gc[j], dacc[t][j]);
          if (t == 4) {
#pragma unroll
            for (int j = 0; j < 8; ++j) ac[j] = a[j];
          }
        }
        // input gradient: g at (hh-kh+1, ww-kw+1) times w[tap]
        const int oh = hh - kh + 1, ow = ww - kw + 1;
        if (oh >= 0 && oh < H && ow >= 0 && ow < W) {
          float gg[8];
          unpack8(ldg16(g + ((((long)b * H + oh) * W + ow) * c8 + cv) * 8), gg);
#pragma unroll
          for (int j = 0; j < 8; ++j) da[j] = fmaf(gg[j], wk[t][j], da[j]);
        }
      }
    }
    float r[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (resid && !(hh & 1) && !(ww & 1))
      unpack8(ldg16(resid_t + ((((long)b * (H >> 1) + (hh >> 1)) * (W >> 1) + (ww >> 1)) * c8 + cv) * 8), r);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float m = ac[j] > 0.f ? 1.f : 0.f;
      da[j] = (resid == 2) ? m * (da[j] + r[j]) : m * da[j] + r[j];
    }
    stg16(dz + (p * c8 + cv) * 8, pack8(da));
  }
#pragma unroll
  for (int t = 0; t < 9; ++t) block_channel_atomic(dacc[t], c8, sm, dwk + t * C);
}

// ---------------------------------------------------------------------------------------
// K7 encoder block tail: x_next = maxpool3x3s2_same(bn2(y2)) + residual; argmax kept for backward.
__global__ void __launch_bounds__(TPB) pool_add_fwd_k(const __nv_bfloat16* __restrict__ y, const float* __restrict__ bn,
                                                      const __nv_bfloat16* __restrict__ res, __nv_bfloat16* __restrict__ out,
                                                      __nv_bfloat16* __restrict__ even, __nv_bfloat16* __restrict__ relu_out,
                                                      uint8_t* __restrict__ arg, int B, int H, int W, int C, const BnFin fin) {
  ck::pdl_enter();
  const int c8 = C >> 3, Ho = H >> 1, Wo = W >> 1;
  const int total = (int)B * Ho * Wo * c8;
  const int g0 = (int)blockIdx.x * blockDim.x + threadIdx.x;
  const int cv = (int)(g0 % c8);
  float sc[8], sh[8];
  bn_coeffs(fin, bn, C, cv, sc, sh);
  bn_publish(fin, C);
  const IdxDec dec = make_idxdec(c8, Wo, Ho);
  for (int g = g0; g < total; g += (int)gridDim.x * blockDim.x) {
    int p, ow, oh, b;
    idx_dec(dec, g, c8, Wo, Ho, p, ow, oh, b);
    float best[8];
    int bi[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; bi[j] = 0; }
    // all nine window loads (and the residual) are issued before the first compare: the first version loaded, compared,
    // loaded ... behind `continue`s, one dependent DRAM round trip per tap (ncu: long-scoreboard stall on every unpack).
    // Taps that fall off the bottom / right edge re-read the window's first pixel and are skipped in the compare.
    uint4 win[9];
    const int ybase = (((int)b * H + 2 * oh) * W + 2 * ow) * c8 + cv;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const int kh = t / 3, kw = t - kh * 3;
      const bool ok = 2 * oh + kh < H && 2 * ow + kw < W;
      win[t] = ldg16(y + (ok ? ybase + (kh * W + kw) * c8 : ybase) * 8);
    }
    const uint4 rraw = ldg16(res + g * 8);
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const int kh = t / 3, kw = t - kh * 3;
      if (2 * oh + kh >= H || 2 * ow + kw >= W) continue;
      float f[8];
      unpack8(win[t], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float z = fmaf(f[j], sc[j], sh[j]);
        if (z > best[j]) { best[j] = z; bi[j] = t; }
      }
    }
    float r[8];
    unpack8(rraw, r);
#pragma unroll
    for (int j = 0; j < 8; ++j) best[j] += r[j];
    const uint4 v = pack8(best);
    stg16(out + g * 8, v);
    if (arg) {
      uint2 a;
      a.x = bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24);
      a.y = bi[4] | (bi[5] << 8) | (bi[6] << 16) | (bi[7] << 24);
      *reinterpret_cast<uint2*>(arg + g * 8) = a;
    }
    if (even && !(oh & 1) && !(ow & 1))
      stg16(even + ((((int)b * (Ho >> 1) + (oh >> 1)) * (Wo >> 1) + (ow >> 1)) * c8 + cv) * 8, v);
    if (relu_out) {
#pragma unroll
      for (int j = 0; j < 8; ++j) best[j] = fmaxf(best[j], 0.f);
      stg16(relu_out + g * 8, pack8(best));
    }
  }
}

// K8 decoder block tail: s = bn2(y2) + (up2 ? upsample2(rl) : rl);  a_next = relu(upsample2(s)).
__global__ void __launch_bounds__(TPB) dec_combine_fwd_k(const __nv_bfloat16* __restrict__ y, const float* __restrict__ bn,
                                                         const __nv_bfloat16* __restrict__ rl, __nv_bfloat16* __restrict__ s,
                                                         __nv_bfloat16* __restrict__ a_next, int B, int H, int W, int C,
                                                         int rl_up, const BnFin fin) {
  ck::pdl_enter();
  const int c8 = C >> 3;
  const int total = (int)B * H * W * c8;
  const int g0 = (int)blockIdx.x * blockDim.x + threadIdx.x;
  const int cv = (int)(g0 % c8);
  float sc[8], sh[8];
  bn_coeffs(fin, bn, C, cv, sc, sh);
  bn_publish(fin, C);
  const IdxDec dec = make_idxdec(c8, W, H);
  for (int g = g0; g < total; g += (int)gridDim.x * blockDim.x) {
    int p, w, h, b;
    idx_dec(dec, g, c8, W, H, p, w, h, b);
    float f[8], r[8];
    unpack8(ldg16(y + g * 8), f);
    const int ri = rl_up ? ((((int)b * (H >> 1) + (h >> 1)) * (W >> 1) + (w >> 1)) * c8 + cv) : g;
    unpack8(ldg16(rl + ri * 8), r);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = fmaf(f[j], sc[j], sh[j]) + r[j];
    stg16(s + g * 8, pack8(f));
    if (a_next) {
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
      const uint4 v = pack8(f);
      const int W2 = 2 * W;
      const int base = (((int)b * 2 * H + 2 * h) * W2 + 2 * w) * c8 + cv;
      stg16(a_next + base * 8, v);
      stg16(a_next + (base + c8) * 8, v);
      stg16(a_next + (base + W2 * c8) * 8, v);
      stg16(a_next + (base + W2 * c8 + c8) * 8, v);
    }
  }
}

// out[q] = (mask ? 1[mask[q] > 0] : 1) * sum_{2x2} in[2q+d]  (+ add[q])     backward of nearest upsample
__global__ void __launch_bounds__(TPB) down2sum_k(const __nv_bfloat16* __restrict__ in, const __nv_bfloat16* __restrict__ mask,
                                                  const __nv_bfloat16* __restrict__ add, __nv_bfloat16* __restrict__ out,
                                                  int B, int Ho, int Wo, int C) {
  ck::pdl_enter();
  const int c8 = C >> 3;
  const int total = (int)B * Ho * Wo * c8;
  const IdxDec dec = make_idxdec(c8, Wo, Ho);
  for (int g = (int)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (int)gridDim.x * blockDim.x) {
    int p, w, h, b;
    idx_dec(dec, g, c8, Wo, Ho, p, w, h, b);
    const int cv = g - p * c8;
    const int W2 = 2 * Wo;
    const int base = (((int)b * 2 * Ho + 2 * h) * W2 + 2 * w) * c8 + cv;
    float a0[8], a1[8], a2[8], a3[8], o[8];
    unpack8(ldg16(in + base * 8), a0);
    unpack8(ldg16(in + (base + c8) * 8), a1);
    unpack8(ldg16(in + (base + W2 * c8) * 8), a2);
    unpack8(ldg16(in + (base + W2 * c8 + c8) * 8), a3);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = (a0[j] + a1[j]) + (a2[j] + a3[j]);
    if (mask) {
      float m[8];
      unpack8(ldg16(mask + g * 8), m);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = m[j] > 0.f ? o[j] : 0.f;
    }
    if (add) {
      float t[8];
      unpack8(ldg16(add + g * 8), t);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] += t[j];
    }
    stg16(out + g * 8, pack8(o));
  }
}

// down2sum_k for a 2x2 block of output pixels per thread, plus what the next two launches of the decoder's backward pass
// would do with that output: the BN-backward reductions of the layer whose output gradient it is (sum out and
// sum out * (y - mean), as bn_bwd_reduce_k<0>) and, optionally, the next down2sum of it (lowres = 2x2 sum of out).
// Everything downstream sees the bf16-rounded out values, exactly as the separate launches would.
__global__ void __launch_bounds__(TPB, 3) down2sum_bn_k(const __nv_bfloat16* __restrict__ in, const __nv_bfloat16* __restrict__ mask,
                                                        const __nv_bfloat16* __restrict__ add, __nv_bfloat16* __restrict__ out,
                                                        __nv_bfloat16* __restrict__ lowres, const __nv_bfloat16* __restrict__ y,
                                                        const float* __restrict__ bn, float* __restrict__ sums, int B, int Ho,
                                                        int Wo, int C, int items_per_block) {
  ck::pdl_enter();
  __shared__ float sm[TPB * 8];
  const int c8 = C >> 3, Hq = Ho >> 1, Wq = Wo >> 1, lanes = TPB / c8;
  const int cv = threadIdx.x % c8, lane = threadIdx.x / c8;
  float mu[8], s1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  load8f(bn + 2 * C + cv * 8, mu);
  const int P2 = B * Hq * Wq;
  const int p0 = blockIdx.x * items_per_block;
  const int p1 = min(P2, p0 + items_per_block);
  const IdxDec pdec = make_idxdec(1, Wq, Hq);
  const int W2 = 2 * Wo;
  for (int p = p0 + lane; p < p1; p += lanes) {
    int pp_, j, i, b;
    idx_dec(pdec, p, 1, Wq, Hq, pp_, j, i, b);
    float part[2][8];  // (q0 + q1) and (q2 + q3) of the rounded outputs: down2sum_k's summation order
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int h = 2 * i + (q >> 1), w = 2 * j + (q & 1);
      const int g = ((b * Ho + h) * Wo + w) * c8 + cv;
      const int base = ((b * 2 * Ho + 2 * h) * W2 + 2 * w) * c8 + cv;
      const uint4 v0 = ldg16(in + (size_t)base * 8), v1 = ldg16(in + (size_t)(base + c8) * 8);
      const uint4 v2 = ldg16(in + (size_t)(base + W2 * c8) * 8), v3 = ldg16(in + (size_t)(base + W2 * c8 + c8) * 8);
      const uint4 yv4 = ldg16(y + (size_t)g * 8);
      uint4 m4 = make_uint4(0u, 0u, 0u, 0u), t4 = make_uint4(0u, 0u, 0u, 0u);
      if (mask) m4 = ldg16(mask + (size_t)g * 8);
      if (add) t4 = ldg16(add + (size_t)g * 8);
      float a0[8], a1[8], a2[8], a3[8], o[8], yv[8];
      unpack8(v0, a0); unpack8(v1, a1); unpack8(v2, a2); unpack8(v3, a3); unpack8(yv4, yv);
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = (a0[e] + a1[e]) + (a2[e] + a3[e]);
      if (mask) {
        float m[8];
        unpack8(m4, m);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = m[e] > 0.f ? o[e] : 0.f;
      }
      if (add) {
        float t[8];
        unpack8(t4, t);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += t[e];
      }
      const uint4 pk = pack8(o);
      float dr[8];
      unpack8(pk, dr);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        s1[e] += dr[e];
        s2[e] = fmaf(dr[e], yv[e] - mu[e], s2[e]);
        part[q >> 1][e] = (q & 1) ? part[q >> 1][e] + dr[e] : dr[e];
      }
      stg16(out + (size_t)g * 8, pk);
    }
    if (lowres) {
      float lo[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) lo[e] = part[0][e] + part[1][e];
      stg16(lowres + (size_t)(((b * Hq + i) * Wq + j) * c8 + cv) * 8, pack8(lo));
    }
  }
  block_channel_atomic(s1, c8, sm, sums);
  block_channel_atomic(s2, c8, sm, sums + C);
}

// out = 1[x > 0] * g + add      (first decoder block: gradient into the bottleneck tensor)
__global__ void __launch_bounds__(TPB) relu_bwd_add_k(const __nv_bfloat16* __restrict__ g, const __nv_bfloat16* __restrict__ x,
                                                      const __nv_bfloat16* __restrict__ add, __nv_bfloat16* __restrict__ out,
                                                      long nvec) {
  ck::pdl_enter();
  for (int i = (int)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int)gridDim.x * blockDim.x) {
    float a[8], m[8], t[8];
    unpack8(ldg16(g + i * 8), a);
    unpack8(ldg16(x + i * 8), m);
    if (add) unpack8(ldg16(add + i * 8), t);
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = (m[j] > 0.f ? a[j] : 0.f) + (add ? t[j] : 0.f);
    stg16(out + i * 8, pack8(a));
  }
}

// ---------------------------------------------------------------------------------------
// BN backward.  dz (gradient w.r.t. the BN output) comes from one of three sources:
//   0 DIRECT  dz = g[p]
//   1 RELU    dz = g[p] * 1[bn(y[p]) > 0]
//   2 POOL    dz[p] = sum over the (<=4) pooling windows q covering p of gq[q] * [argmax[q] == offset(p,q)]
template <int MODE>
__device__ __forceinline__ void load_dz(const __nv_bfloat16* __restrict__ g, const uint8_t* __restrict__ arg, long p, int cv,
                                        int c8, int H, int W, const float yv[8], const float sc[8], const float sh[8],
                                        float dz[8]) {
  if (MODE == 0 || MODE == 1) {
    unpack8(ldg16(g + (p * c8 + cv) * 8), dz);
    if (MODE == 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j) dz[j] = fmaf(yv[j], sc[j], sh[j]) > 0.f ? dz[j] : 0.f;
    }
  } else {
    int pp_, w, h, b;
    idx_dec(make_idxdec(1, W, H), (int)p, 1, W, H, pp_, w, h, b);
    const int Ho = H >> 1, Wo = W >> 1;
#pragma unroll
    for (int j = 0; j < 8; ++j) dz[j] = 0.f;
    const int qh0 = (h & 1) ? (h >> 1) : (h >> 1) - 1, qh1 = h >> 1;
    const int qw0 = (w & 1) ? (w >> 1) : (w >> 1) - 1, qw1 = w >> 1;
    for (int qh = max(qh0, 0); qh <= min(qh1, Ho - 1); ++qh)
      for (int qw = max(qw0, 0); qw <= min(qw1, Wo - 1); ++qw) {
        const int q = ((b * Ho + qh) * Wo + qw) * c8 + cv;
        const int off = (h - 2 * qh) * 3 + (w - 2 * qw);
        const uint2 a = __ldg(reinterpret_cast<const uint2*>(arg + q * 8));
        float gq[8];
        unpack8(ldg16(g + q * 8), gq);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int idx = ((j < 4 ? a.x : a.y) >> ((j & 3) * 8)) & 0xff;
          if (idx == off) dz[j] += gq[j];
        }
      }
  }
}

// sums = [sum dz  C][sum dz*(y-mean)  C]   (centred moment: no cancellation; apply multiplies by invstd)
template <int MODE>
__global__ void __launch_bounds__(TPB, MODE == 2 ? 2 : (MODE == 0 ? 4 : 3)) bn_bwd_reduce_k(const __nv_bfloat16* __restrict__ g, const uint8_t* __restrict__ arg,
                                                       const __nv_bfloat16* __restrict__ y, const float* __restrict__ bn,
                                                       float* __restrict__ sums, long P, int H, int W, int C,
                                                       int pix_per_block) {
  ck::pdl_enter();
  __shared__ float sm[TPB * 8];
  const int c8 = C >> 3, lanes = TPB / c8;
  const int cv = threadIdx.x % c8, lane = threadIdx.x / c8;
  float sc[8], sh[8], mu[8], s1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (MODE == 1) { load8f(bn + cv * 8, sc); load8f(bn + C + cv * 8, sh); }
  load8f(bn + 2 * C + cv * 8, mu);
  const int p0 = (int)blockIdx.x * pix_per_block;
  const int p1 = min((int)P, p0 + pix_per_block);
  for (int p = p0 + lane; p < p1; p += 2 * lanes) {
    const int q = p + lanes;
    const bool has2 = q < p1;
    float yv[2][8], dz[2][8];
    unpack8(ldg16(y + (p * c8 + cv) * 8), yv[0]);
    if (has2) unpack8(ldg16(y + (q * c8 + cv) * 8), yv[1]);
    load_dz<MODE>(g, arg, p, cv, c8, H, W, yv[0], sc, sh, dz[0]);
    if (has2) load_dz<MODE>(g, arg, q, cv, c8, H, W, yv[1], sc, sh, dz[1]);
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !has2) break;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s1[j] += dz[u][j];
        s2[j] = fmaf(dz[u][j], yv[u][j] - mu[j], s2[j]);
      }
    }
  }
  block_channel_atomic(s1, c8, sm, sums);
  block_channel_atomic(s2, c8, sm, sums + C);
}

// dy = scale * (dz - mean(dz) - xhat * mean(dz*xhat)), folded into dy = A*dz + Bc*y + Cc per channel;
// block 0 also emits dgamma = sum dz*xhat and dbeta = sum dz.
template <int MODE>
__global__ void __launch_bounds__(TPB, MODE == 2 ? 2 : (MODE == 0 ? 4 : 3)) bn_bwd_apply_k(const __nv_bfloat16* __restrict__ g, const uint8_t* __restrict__ arg,
                                                      const __nv_bfloat16* __restrict__ y, const float* __restrict__ bn,
                                                      const float* __restrict__ sums, __nv_bfloat16* __restrict__ dy,
                                                      float* __restrict__ dgamma, float* __restrict__ dbeta, long P, int H,
                                                      int W, int C) {
  ck::pdl_enter();
  const int c8 = C >> 3;
  const float invn = 1.0f / (float)P;
  if (blockIdx.x == 0) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      dbeta[c] = sums[c];
      dgamma[c] = bn[3 * C + c] * sums[C + c];
    }
  }
  const int total = P * c8;
  const int stride = (int)gridDim.x * blockDim.x;  // multiple of c8: a thread keeps its channel vector
  const int g0 = (int)blockIdx.x * blockDim.x + threadIdx.x;
  const int cv = (int)(g0 % c8);
  float sc[8], sh[8], A[8], Bc[8], Cc[8];
  {
    float mu[8], is[8], a1[8], a2[8];
    load8f(bn + cv * 8, sc); load8f(bn + C + cv * 8, sh);
    load8f(bn + 2 * C + cv * 8, mu); load8f(bn + 3 * C + cv * 8, is);
    load8f(sums + cv * 8, a1); load8f(sums + C + cv * 8, a2);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float m1 = a1[j] * invn;                               // mean(dz)
      const float m2 = is[j] * a2[j] * invn;                       // mean(dz * xhat)
      A[j] = sc[j];
      Bc[j] = -sc[j] * is[j] * m2;
      Cc[j] = -sc[j] * m1 - Bc[j] * mu[j];
    }
  }
  for (int gi = g0; gi < total; gi += 2 * stride) {
    const int gj = gi + stride;
    const bool has2 = gj < total;
    float yv[2][8], dz[2][8];
    unpack8(ldg16(y + gi * 8), yv[0]);
    if (has2) unpack8(ldg16(y + gj * 8), yv[1]);
    // (DIRECT / RELU sources only need the flat index back: the pixel index is computed with a shift, the division is
    //  kept for the pooled source whose shapes need not be powers of two)
    const int lc8 = ck_log2_or_neg(c8);
    load_dz<MODE>(g, arg, lc8 >= 0 ? (gi >> lc8) : gi / c8, cv, c8, H, W, yv[0], sc, sh, dz[0]);
    if (has2) load_dz<MODE>(g, arg, lc8 >= 0 ? (gj >> lc8) : gj / c8, cv, c8, H, W, yv[1], sc, sh, dz[1]);
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !has2) break;
#pragma unroll
      for (int j = 0; j < 8; ++j) dz[u][j] = fmaf(A[j], dz[u][j], fmaf(Bc[j], yv[u][j], Cc[j]));
      stg16(dy + (u ? gj : gi) * 8, pack8(dz[u]));
    }
  }
}

// max-pool backward materialised once: dz[p] = sum over the windows covering p of g[q] * [argmax[q] == offset(p,q)].
// The BN2 reduce + apply kernels then run in their cheap DIRECT mode instead of decoding the arg-max twice.
__global__ void __launch_bounds__(TPB, 3) pool_bwd_k(const __nv_bfloat16* __restrict__ g, const uint8_t* __restrict__ arg,
                                                     __nv_bfloat16* __restrict__ dz, long P, int H, int W, int C) {
  ck::pdl_enter();
  const int c8 = C >> 3;
  const int total = P * c8;
  float unused[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int gi = (int)blockIdx.x * blockDim.x + threadIdx.x; gi < total; gi += (int)gridDim.x * blockDim.x) {
    float v[8];
    load_dz<2>(g, arg, gi / c8, (int)(gi % c8), c8, H, W, unused, unused, unused, v);
    stg16(dz + gi * 8, pack8(v));
  }
}

// max-pool backward + the BN-backward reductions of the layer below it, in one pass (second version of the pair
// pool_bwd_k + bn_bwd_reduce_k<0>): a thread owns one 2x2 block of input pixels x 8 channels, reads the (<= 4) pooling
// windows that cover it once (the per-pixel version read 9 windows per 2x2 block), writes the four dz vectors and
// accumulates sum dz and sum dz * (y - mean) on the way.
__global__ void __launch_bounds__(TPB, 2) pool_bwd_bn_k(const __nv_bfloat16* __restrict__ g, const uint8_t* __restrict__ arg,
                                                        const __nv_bfloat16* __restrict__ y, const float* __restrict__ bn,
                                                        __nv_bfloat16* __restrict__ dz, float* __restrict__ sums, int B, int H,
                                                        int W, int C, int items_per_block) {
  ck::pdl_enter();
  __shared__ float sm[TPB * 8];
  const int c8 = C >> 3, Ho = H >> 1, Wo = W >> 1, lanes = TPB / c8;
  const int cv = threadIdx.x % c8, lane = threadIdx.x / c8;
  float mu[8], s1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  load8f(bn + 2 * C + cv * 8, mu);
  const int P2 = B * Ho * Wo;
  const int p0 = blockIdx.x * items_per_block;
  const int p1 = min(P2, p0 + items_per_block);
  const IdxDec pdec = make_idxdec(1, Wo, Ho);
  for (int p = p0 + lane; p < p1; p += lanes) {
    int pp_, j, i, b;
    idx_dec(pdec, p, 1, Wo, Ho, pp_, j, i, b);
    float d[4][8];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int e = 0; e < 8; ++e) d[q][e] = 0.f;
    // window (i + di, j + dj), di, dj in {0, -1}; pixel (r, c) of the 2x2 block sits at offset (r - 2di)*3 + (c - 2dj)
    // the four windows' arg-max bytes and gradients, and the four y vectors of the 2x2 block, are all requested before
    // the first use (the first version loaded and consumed them one by one behind `continue`s: one dependent DRAM round
    // trip per window and per pixel - ncu: long-scoreboard on the first unpack after every load)
    uint2 aw[4];
    uint4 gw[4], yw[4];
#pragma unroll
    for (int wi = 0; wi < 4; ++wi) {
      const int qi = max(i - (wi >> 1), 0), qj = max(j - (wi & 1), 0);  // clamped: an off-image window is skipped below
      const int qv = ((b * Ho + qi) * Wo + qj) * c8 + cv;
      aw[wi] = __ldg(reinterpret_cast<const uint2*>(arg + (size_t)qv * 8));
      gw[wi] = ldg16(g + (size_t)qv * 8);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q)
      yw[q] = ldg16(y + (((size_t)(b * H + 2 * i + (q >> 1)) * W + 2 * j + (q & 1)) * c8 + cv) * 8);
#pragma unroll
    for (int wi = 0; wi < 4; ++wi) {
      const int di = -(wi >> 1), dj = -(wi & 1);
      const int qi = i + di, qj = j + dj;
      if (qi < 0 || qj < 0) continue;
      const uint2 a = aw[wi];
      float gq[8];
      unpack8(gw[wi], gq);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int r = q >> 1, c = q & 1;
        const int rr = r - 2 * di, cc = c - 2 * dj;  // position inside the 3x3 window
        if (rr > 2 || cc > 2) continue;             // compile-time: this window does not cover that pixel
        const int off = rr * 3 + cc;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int idx = ((e < 4 ? a.x : a.y) >> ((e & 3) * 8)) & 0xff;
          if (idx == off) d[q][e] += gq[e];
        }
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int h = 2 * i + (q >> 1), w = 2 * j + (q & 1);
      const size_t v = ((size_t)(b * H + h) * W + w) * c8 + cv;
      float yv[8];
      unpack8(yw[q], yv);
      // the reductions see the bf16-rounded dz, exactly what the apply pass will read back
      const uint4 pk = pack8(d[q]);
      float dr[8];
      unpack8(pk, dr);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        s1[e] += dr[e];
        s2[e] = fmaf(dr[e], yv[e] - mu[e], s2[e]);
      }
      stg16(dz + v * 8, pk);
    }
  }
  block_channel_atomic(s1, c8, sm, sums);
  block_channel_atomic(s2, c8, sm, sums + C);
}

// column sum of a [P][C] bf16 matrix into fp32 (bias gradient of the residual convs)
__global__ void __launch_bounds__(TPB) colsum_k(const __nv_bfloat16* __restrict__ g, float* __restrict__ out, long P, int C,
                                                int pix_per_block) {
  ck::pdl_enter();
  __shared__ float sm[TPB * 8];
  const int c8 = C >> 3, lanes = TPB / c8;
  const int cv = threadIdx.x % c8, lane = threadIdx.x / c8;
  float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const int p0 = (int)blockIdx.x * pix_per_block;
  const int p1 = min((int)P, p0 + pix_per_block);
  for (int p = p0 + lane; p < p1; p += lanes) {
    float f[8];
    unpack8(ldg16(g + (p * c8 + cv) * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] += f[j];
  }
  block_channel_atomic(s, c8, sm, out);
}

// ---------------------------------------------------------------------------------------
// K9 head.  The final Up2x commutes with the 1x1 conv and the sigmoid, so the logit is computed at
// the decoder's last low resolution (Hl = H/2) and compared against the 2x2 full-resolution mask
// pixels it covers: the (B,H,W,32) tensor of client_fit_model.py:136-141 is never materialised.
// acc = [bce_sum, correct, inter, psum, ysum, pixel_count]
// Thread mapping: thread = (pixel, channel-vector); the c8 threads of one pixel are adjacent lanes of a
// warp and reduce the 1x1-conv dot product with shuffles (c8 must divide 32).
// Binary cross-entropy of one logit against y in {0,1}.
//   clip == 0 (default): the form Keras 2 trains with.  `fit` on a model whose output layer is
//     Conv2D(..., activation="sigmoid") (client_fit_model.py:145,157) does not see probabilities: the sigmoid caches its
//     input (`output._keras_logits`; in graph mode the op type "Sigmoid" is traced back) and backend.binary_crossentropy
//     switches to tf.nn.sigmoid_cross_entropy_with_logits: max(z,0) - z*y + log(1+exp(-|z|)), gradient sigmoid(z)-y for
//     EVERY z.
//   clip == 1: the fall-back of the same function (and Keras 3): probabilities clipped to [1e-7, 1-1e-7], and
//     clip_by_value passes no gradient outside the clip - once |z| > 16.1 everywhere the step is dead for good.
__device__ __forceinline__ float bce_logit(float z, float yv) {
  return fmaxf(z, 0.f) - z * yv + __logf(1.f + __expf(-fabsf(z)));
}

__device__ __forceinline__ float head_logit(const float f[8], const float* sw, int cv, int c8) {
  float l = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) l = fmaf(f[j], sw[cv * 8 + j], l);
  for (int o = c8 >> 1; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
  return l;
}

__global__ void __launch_bounds__(TPB) head_fwd_k(const __nv_bfloat16* __restrict__ s, const float* __restrict__ w,
                                                  const float* __restrict__ bias, const uint8_t* __restrict__ mask,
                                                  float* __restrict__ acc, float* __restrict__ prob, int B, int Hl, int Wl,
                                                  int C, int bce_clip) {
  ck::pdl_enter();
  extern __shared__ float sw[];
  for (int i = threadIdx.x; i < C; i += blockDim.x) sw[i] = w[i];
  __syncthreads();
  const int c8 = C >> 3;
  const long P = (long)B * Hl * Wl;
  const long total = P * c8;
  const long stride = (long)gridDim.x * blockDim.x;  // multiple of 32 -> whole pixels stay inside a warp
  const int Wf = 2 * Wl, Hf = 2 * Hl;
  float a[5] = {0, 0, 0, 0, 0};
  const long iters = (total + stride - 1) / stride;
  const IdxDec hdec = make_idxdec(c8, Wl, Hl);  // (total fits 32 bits: checked by the launcher)
  for (long it = 0; it < iters; ++it) {
    const long g = it * stride + (long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = g < total;
    const long gg = valid ? g : 0;
    int p_, wl, hl, b_;
    idx_dec(hdec, (int)gg, c8, Wl, Hl, p_, wl, hl, b_);
    const int cv = (int)gg - p_ * c8;
    const long p = p_, b = b_;
    float f[8];
    unpack8(ldg16(s + gg * 8), f);
    const float logit = head_logit(f, sw, cv, c8) + bias[0];
    if (!valid || cv != 0) continue;
    const float pr = 1.f / (1.f + __expf(-logit));
    const long base = (b * Hf + 2 * hl) * Wf + 2 * wl;
    if (prob) {
      prob[base] = pr; prob[base + 1] = pr; prob[base + Wf] = pr; prob[base + Wf + 1] = pr;
    }
    if (mask) {
      const float pc = fminf(fmaxf(pr, 1e-7f), 1.f - 1e-7f);
      const float lp = bce_clip ? -__logf(pc) : bce_logit(logit, 1.f);
      const float lq = bce_clip ? -__logf(1.f - pc) : bce_logit(logit, 0.f);
#pragma unroll
      for (int d = 0; d < 4; ++d) {
        const float yv = mask[base + (d >> 1) * Wf + (d & 1)] ? 1.f : 0.f;
        a[0] += yv * lp + (1.f - yv) * lq;
        a[1] += ((pr > 0.5f) == (yv > 0.5f)) ? 1.f : 0.f;
        a[2] += pr * yv;
        a[3] += pr;
        a[4] += yv;
      }
    }
  }
  if (mask) {
    __shared__ float red[5][TPB / 32];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      const float v = warp_sum(a[k]);
      if ((threadIdx.x & 31) == 0) red[k][threadIdx.x >> 5] = v;
    }
    __syncthreads();
    if (threadIdx.x < 5) {
      float v = 0.f;
      for (int i = 0; i < TPB / 32; ++i) v += red[threadIdx.x][i];
      atomicAdd(acc + threadIdx.x, v);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(acc + 5, (float)(P * 4));
  }
}

// gradient of mean-BCE (+ optional soft Dice) w.r.t. the low-res logit, then ds = dlogit * w,
// dw += dlogit * s, db += dlogit.  `sums` = the forward accumulators of THIS batch (needed by Dice).
__global__ void __launch_bounds__(TPB) head_bwd_k(const __nv_bfloat16* __restrict__ s, const float* __restrict__ w,
                                                  const float* __restrict__ bias, const uint8_t* __restrict__ mask,
                                                  const float* __restrict__ sums, __nv_bfloat16* __restrict__ ds,
                                                  float* __restrict__ dw, float* __restrict__ db, int B, int Hl, int Wl, int C,
                                                  int use_dice, float dice_smooth, float loss_scale,
                                                  float* __restrict__ acc_out, int bce_clip) {
  ck::pdl_enter();
  extern __shared__ float sh[];  // w[C] | reduction scratch [TPB*8]
  float* sw = sh;
  float* sm = sh + C;
  for (int i = threadIdx.x; i < C; i += blockDim.x) sw[i] = w[i];
  __syncthreads();
  const int c8 = C >> 3;
  const long P = (long)B * Hl * Wl;
  const long total = P * c8;
  const long stride = (long)gridDim.x * blockDim.x;
  const int Wf = 2 * Wl, Hf = 2 * Hl;
  const float invn = 1.0f / (float)(P * 4);
  float dnum = 0.f, dden = 1.f;
  if (use_dice) {
    dden = sums[3] + sums[4] + dice_smooth;
    dnum = 2.f * sums[2] + dice_smooth;
  }
  float dwacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  float dbl = 0.f;
  float am[5] = {0, 0, 0, 0, 0};  // acc_out != NULL (plain BCE): the forward's loss / metric sums, so head_fwd_k is not launched
  const long iters = (total + stride - 1) / stride;
  const IdxDec hdec = make_idxdec(c8, Wl, Hl);  // (total fits 32 bits: checked by the launcher)
  for (long it = 0; it < iters; ++it) {
    const long g = it * stride + (long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = g < total;
    const long gg = valid ? g : 0;
    int p_, wl, hl, b_;
    idx_dec(hdec, (int)gg, c8, Wl, Hl, p_, wl, hl, b_);
    const int cv = (int)gg - p_ * c8;
    const long p = p_, b = b_;
    float f[8], o[8];
    unpack8(ldg16(s + gg * 8), f);
    const float logit = head_logit(f, sw, cv, c8) + bias[0];
    if (!valid) continue;
    const float pr = 1.f / (1.f + __expf(-logit));
    const long base = (b * Hf + 2 * hl) * Wf + 2 * wl;
    const bool inside = !bce_clip || ((pr >= 1e-7f) && (pr <= 1.f - 1e-7f));  // clip_by_value passes no gradient outside
    float dl = 0.f;
    const bool metrics = acc_out != nullptr && cv == 0;
    float lp = 0.f, lq = 0.f;
    if (metrics) {
      const float pc = fminf(fmaxf(pr, 1e-7f), 1.f - 1e-7f);
      lp = bce_clip ? -__logf(pc) : bce_logit(logit, 1.f);
      lq = bce_clip ? -__logf(1.f - pc) : bce_logit(logit, 0.f);
    }
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      const float yv = mask[base + (d >> 1) * Wf + (d & 1)] ? 1.f : 0.f;
      if (inside) dl += (pr - yv) * invn;
      if (use_dice) dl += -(2.f * yv * dden - dnum) / (dden * dden) * pr * (1.f - pr);
      if (metrics) {
        am[0] += yv * lp + (1.f - yv) * lq;
        am[1] += ((pr > 0.5f) == (yv > 0.5f)) ? 1.f : 0.f;
        am[2] += pr * yv;
        am[3] += pr;
        am[4] += yv;
      }
    }
    dl *= loss_scale;
    if (cv == 0) dbl += dl;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      o[j] = dl * sw[cv * 8 + j];
      dwacc[j] = fmaf(dl, f[j], dwacc[j]);
    }
    stg16(ds + gg * 8, pack8(o));
  }
  // thread t always owns channel vector (t % c8) because stride and blockDim are multiples of c8
  block_channel_atomic(dwacc, c8, sm, dw);
  dbl = warp_sum(dbl);
  if ((threadIdx.x & 31) == 0) atomicAdd(db, dbl);
  if (acc_out) {
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      const float v = warp_sum(am[k]);
      if ((threadIdx.x & 31) == 0) sm[k * (TPB / 32) + (threadIdx.x >> 5)] = v;
    }
    __syncthreads();
    if (threadIdx.x < 5) {
      float v = 0.f;
      for (int i = 0; i < TPB / 32; ++i) v += sm[threadIdx.x * (TPB / 32) + i];
      atomicAdd(acc_out + threadIdx.x, v);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(acc_out + 5, (float)(P * 4));
  }
}


// ---------------------------------------------------------------------------------------
// v2 depthwise kernels: the naive versions above issue 9 (fwd) / 18 (bwd) 16-byte L1 loads per output vector and
// are L1-bandwidth bound (128 B/clk/SM).  Here a thread owns a strip of S=4 consecutive pixels of one row for one
// channel vector and slides a 3 x (S+2) register window over it: 4.5 loads (and BN+ReLU evaluations) per output.
constexpr int DWS = 4;

template <bool ACT>
__device__ __forceinline__ void load_row_window(const __nv_bfloat16* __restrict__ src, long row_base_vec, int w0, int W,
                                                int c8, int cv, int mode, const float sc[8], const float sh[8],
                                                float win[DWS + 2][8]) {
#pragma unroll
  for (int i = 0; i < DWS + 2; ++i) {
    const int iw = w0 - 1 + i;
    if (iw >= 0 && iw < W) {
      if (ACT) act_load(src, row_base_vec + (long)iw * c8 + cv, mode, sc, sh, win[i]);
      else unpack8(ldg16(src + (row_base_vec + (long)iw * c8 + cv) * 8), win[i]);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) win[i][j] = 0.f;
    }
  }
}

__global__ void __launch_bounds__(TPB, 2) dw3x3_fwd_v2_k(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w,
                                                         const float* __restrict__ bn, __nv_bfloat16* __restrict__ out,
                                                         int B, int H, int W, int C, int mode, const BnFin fin) {
  ck::pdl_enter();
  const int c8 = C >> 3, WS = W / DWS;
  const long total = (long)B * H * WS * c8;
  const long g0 = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const int cv = (int)(g0 % c8);  // grid stride is a multiple of c8: the channel vector is loop invariant
  float sc[8], sh[8];
  if (mode == 1) bn_coeffs(fin, bn, C, cv, sc, sh);
  bn_publish(fin, C);
  for (long g = g0; g < total; g += (long)gridDim.x * blockDim.x) {
    long s = g / c8;
    const int ws = (int)(s % WS); s /= WS;
    const int hh = (int)(s % H);
    const long b = s / H;
    const int w0 = ws * DWS;
    float acc[DWS][8];
#pragma unroll
    for (int i = 0; i < DWS; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int ih = hh + r - 1;
      if (ih < 0 || ih >= H) continue;
      float win[DWS + 2][8], wk[3][8];
      load_row_window<true>(x, (b * H + ih) * (long)W * c8, w0, W, c8, cv, mode, sc, sh, win);
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) load8f(w + (r * 3 + kw) * C + cv * 8, wk[kw]);
#pragma unroll
      for (int i = 0; i < DWS; ++i)
#pragma unroll
        for (int kw = 0; kw < 3; ++kw)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(win[i + kw][j], wk[kw][j], acc[i][j]);
    }
    const long obase = ((b * H + hh) * (long)W + w0) * c8 + cv;
#pragma unroll
    for (int i = 0; i < DWS; ++i) stg16(out + (obase + (long)i * c8) * 8, pack8(acc[i]));
  }
}

// input gradient: da[p] = sum_tap g[p - off(tap)] * w[tap]; dz = mask * da (+ residual), see dw3x3_bwd_k
__global__ void __launch_bounds__(TPB, 2) dw3x3_dgrad_v2_k(const __nv_bfloat16* __restrict__ g, const __nv_bfloat16* __restrict__ src,
                                                           const float* __restrict__ w, const float* __restrict__ bn,
                                                           const __nv_bfloat16* __restrict__ resid_t,
                                                           __nv_bfloat16* __restrict__ dz, int B, int H, int W, int C, int mode,
                                                           int resid) {
  ck::pdl_enter();
  const int c8 = C >> 3, WS = W / DWS;
  const long total = (long)B * H * WS * c8;
  for (long gi = (long)blockIdx.x * blockDim.x + threadIdx.x; gi < total; gi += (long)gridDim.x * blockDim.x) {
    const int cv = (int)(gi % c8);
    long s = gi / c8;
    const int ws = (int)(s % WS); s /= WS;
    const int hh = (int)(s % H);
    const long b = s / H;
    const int w0 = ws * DWS;
    float sc[8], sh[8], acc[DWS][8];
    if (mode == 1) { load8f(bn + cv * 8, sc); load8f(bn + C + cv * 8, sh); }
#pragma unroll
    for (int i = 0; i < DWS; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int oh = hh + r - 1;  // row of g; it used tap kh = 2 - r to reach row hh
      if (oh < 0 || oh >= H) continue;
      float win[DWS + 2][8], wk[3][8];
      load_row_window<false>(g, (b * H + oh) * (long)W * c8, w0, W, c8, cv, 0, sc, sh, win);
#pragma unroll
      for (int c = 0; c < 3; ++c) load8f(w + ((2 - r) * 3 + (2 - c)) * C + cv * 8, wk[c]);
#pragma unroll
      for (int i = 0; i < DWS; ++i)
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(win[i + c][j], wk[c][j], acc[i][j]);
    }
    const long obase = ((b * H + hh) * (long)W + w0) * c8 + cv;
#pragma unroll
    for (int i = 0; i < DWS; ++i) {
      float a[8], r8[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      act_load(src, obase + (long)i * c8, mode, sc, sh, a);
      const int ww = w0 + i;
      if (resid && !(hh & 1) && !(ww & 1))
        unpack8(ldg16(resid_t + (((b * (H >> 1) + (hh >> 1)) * (long)(W >> 1) + (ww >> 1)) * c8 + cv) * 8), r8);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float m = a[j] > 0.f ? 1.f : 0.f;
        acc[i][j] = (resid == 2) ? m * (acc[i][j] + r8[j]) : m * acc[i][j] + r8[j];
      }
      stg16(dz + (obase + (long)i * c8) * 8, pack8(acc[i]));
    }
  }
}

// weight gradient: dw[tap][c] += sum_o a[o + off(tap)][c] * g[o][c]
__global__ void __launch_bounds__(TPB, 1) dw3x3_wgrad_v2_k(const __nv_bfloat16* __restrict__ g, const __nv_bfloat16* __restrict__ src,
                                                           const float* __restrict__ bn, float* __restrict__ dwk, int B, int H,
                                                           int W, int C, int mode, int strips_per_block) {
  ck::pdl_enter();
  __shared__ float sm[TPB * 8];
  const int c8 = C >> 3, WS = W / DWS, lanes = TPB / c8;
  const int cv = threadIdx.x % c8, lane = threadIdx.x / c8;
  const long S = (long)B * H * WS;
  const long s0 = (long)blockIdx.x * strips_per_block;
  const long s1 = min(S, s0 + strips_per_block);
  float sc[8], sh[8], dacc[9][8];
  if (mode == 1) { load8f(bn + cv * 8, sc); load8f(bn + C + cv * 8, sh); }
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int j = 0; j < 8; ++j) dacc[t][j] = 0.f;
  for (long s = s0 + lane; s < s1; s += lanes) {
    const int ws = (int)(s % WS);
    const int hh = (int)((s / WS) % H);
    const long b = s / ((long)WS * H);
    const int w0 = ws * DWS;
    float gc[DWS][8];
    const long obase = ((b * H + hh) * (long)W + w0) * c8 + cv;
#pragma unroll
    for (int i = 0; i < DWS; ++i) unpack8(ldg16(g + (obase + (long)i * c8) * 8), gc[i]);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int ih = hh + r - 1;
      if (ih < 0 || ih >= H) continue;
      float win[DWS + 2][8];
      load_row_window<true>(src, (b * H + ih) * (long)W * c8, w0, W, c8, cv, mode, sc, sh, win);
#pragma unroll
      for (int kw = 0; kw < 3; ++kw)
#pragma unroll
        for (int i = 0; i < DWS; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) dacc[r * 3 + kw][j] = fmaf(win[i + kw][j], gc[i][j], dacc[r * 3 + kw][j]);
    }
  }
#pragma unroll
  for (int t = 0; t < 9; ++t) block_channel_atomic(dacc[t], c8, sm, dwk + t * C);
}


// v3 depthwise weight gradient: the three tap rows are split over threadIdx.y, so a thread keeps 3x8 accumulators
// instead of 9x8 (the v2 kernel needed 212 registers and ran one 8-warp block per SM).  blockDim = (128, 3).
__global__ void __launch_bounds__(384, 1) dw3x3_wgrad_v3_k(const __nv_bfloat16* __restrict__ g, const __nv_bfloat16* __restrict__ src,
                                                           const float* __restrict__ bn, float* __restrict__ dwk, int B, int H,
                                                           int W, int C, int mode, int strips_per_block) {
  ck::pdl_enter();
  __shared__ float sm[384 * 8];
  const int c8 = C >> 3, WS = W / DWS, lanes = 128 / c8;
  const int tx = threadIdx.x, r = threadIdx.y;
  const int cv = tx % c8, lane = tx / c8;
  const long S = (long)B * H * WS;
  const long s0 = (long)blockIdx.x * strips_per_block;
  const long s1 = min(S, s0 + strips_per_block);
  float sc[8], sh[8], dacc[3][8];
  if (mode == 1) { load8f(bn + cv * 8, sc); load8f(bn + C + cv * 8, sh); }
#pragma unroll
  for (int t = 0; t < 3; ++t)
#pragma unroll
    for (int j = 0; j < 8; ++j) dacc[t][j] = 0.f;
  for (long s = s0 + lane; s < s1; s += lanes) {
    const int ws = (int)(s % WS);
    const int hh = (int)((s / WS) % H);
    const long b = s / ((long)WS * H);
    const int w0 = ws * DWS;
    const int ih = hh + r - 1;
    if (ih < 0 || ih >= H) continue;
    float gc[DWS][8], win[DWS + 2][8];
    const long obase = ((b * H + hh) * (long)W + w0) * c8 + cv;
#pragma unroll
    for (int i = 0; i < DWS; ++i) unpack8(ldg16(g + (obase + (long)i * c8) * 8), gc[i]);
    load_row_window<true>(src, (b * H + ih) * (long)W * c8, w0, W, c8, cv, mode, sc, sh, win);
#pragma unroll
    for (int kw = 0; kw < 3; ++kw)
#pragma unroll
      for (int i = 0; i < DWS; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) dacc[kw][j] = fmaf(win[i + kw][j], gc[i][j], dacc[kw][j]);
  }
  const int tid = r * 128 + tx;
  for (int kw = 0; kw < 3; ++kw) {
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 8; ++j) sm[tid * 8 + j] = dacc[kw][j];
    __syncthreads();
    for (int i = tid; i < 3 * c8 * 8; i += 384) {
      const int plane = i / (c8 * 8), rem = i % (c8 * 8);
      const int cvv = rem >> 3, j = rem & 7;
      float v = 0.f;
      for (int l = 0; l < lanes; ++l) v += sm[((plane * 128) + l * c8 + cvv) * 8 + j];
      atomicAdd(dwk + (long)(plane * 3 + kw) * C + cvv * 8 + j, v);
    }
  }
}

// v2 entry-conv weight gradient: block = 256 threads = (C channels) x (256/C k-groups); 128-pixel chunks staged in smem.
__global__ void __launch_bounds__(TPB) entry_conv_wgrad_v2_k(const uint8_t* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                                             float* __restrict__ dw, int B, int H, int W, int C,
                                                             int pix_per_block) {
  ck::pdl_enter();
  extern __shared__ float smw[];  // patch[128][28] | dy[128][C]
  float* sp = smw;
  float* sd = smw + 128 * 28;
  const int Ho = H >> 1, Wo = W >> 1, c8 = C >> 3;
  const long P = (long)B * Ho * Wo;
  const long p0 = (long)blockIdx.x * pix_per_block;
  const long p1 = min(P, p0 + pix_per_block);
  const int G = TPB / C;                 // k-groups
  const int c = threadIdx.x % C, kg = threadIdx.x / C;
  float acc[7] = {0, 0, 0, 0, 0, 0, 0};  // k = kg + i*G, i < ceil(27/G) <= 7 (C <= 64)
  for (long pc = p0; pc < p1; pc += 128) {
    const int np = (int)min((long)128, p1 - pc);
    __syncthreads();
    if (threadIdx.x < 128) {
      const int i = threadIdx.x;
      float* row = sp + i * 28;
      if (i < np) {
        const long p = pc + i;
        const int ow = (int)(p % Wo);
        const int oh = (int)((p / Wo) % Ho);
        const long b = p / ((long)Wo * Ho);
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
          const int ih = 2 * oh + kh;
#pragma unroll
          for (int kw = 0; kw < 3; ++kw) {
            const int iw = 2 * ow + kw;
            const bool ok = ih < H && iw < W;
            const uint8_t* px = x + ((b * H + ih) * (long)W + iw) * 3;
#pragma unroll
            for (int ci = 0; ci < 3; ++ci) row[(kh * 3 + kw) * 3 + ci] = ok ? (float)px[ci] * (1.0f / 255.0f) : 0.f;
          }
        }
      } else {
        for (int k = 0; k < 27; ++k) row[k] = 0.f;
      }
    }
    for (int v = threadIdx.x; v < 128 * c8; v += TPB) {
      const int i = v / c8, cvv = v % c8;
      float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      if (i < np) unpack8(ldg16(dy + ((pc + i) * c8 + cvv) * 8), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) sd[i * C + cvv * 8 + j] = f[j];
    }
    __syncthreads();
    for (int i = 0; i < 128; ++i) {
      const float d = sd[i * C + c];
      const float* row = sp + i * 28;
#pragma unroll
      for (int q = 0; q < 7; ++q) {
        const int k = kg + q * G;
        if (k < 27) acc[q] = fmaf(row[k], d, acc[q]);
      }
    }
  }
#pragma unroll
  for (int q = 0; q < 7; ++q) {
    const int k = kg + q * G;
    if (k < 27) atomicAdd(dw + (long)k * C + c, acc[q]);
  }
}

// Entry conv as a tensor-core GEMM (second generation of K4): the 3x3 stride-2 patches of the uint8 image are written out
// once as a bf16 matrix col[pixel][32] (k = (kh*3+kw)*3+ci for k < 27, zeros for k = 27..31, /255 applied; TF "same"
// padding = bottom/right zeros), and both the forward conv and its weight gradient become K=32 / Cin=32 tcgen05 GEMMs
// over that matrix (the CUDA-core kernels above spend 37 us / 103 us on what is 5 + 10 us of memory traffic).
// ref: client_fit_model.py:100 Conv2D(32,3,strides=2,"same"), :43 x/255
__global__ void __launch_bounds__(TPB) entry_im2col_k(const uint8_t* __restrict__ x, __nv_bfloat16* __restrict__ col, int B, int H,
                                                      int W) {
  ck::pdl_enter();
  const int Ho = H >> 1, Wo = W >> 1;
  const int total = (int)B * Ho * Wo * 4;
  const IdxDec dec = make_idxdec(4, Wo, Ho);
  for (int i = (int)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int)gridDim.x * blockDim.x) {
    const int q = (int)(i & 3);
    int p, ow, oh, b;
    idx_dec(dec, i, 4, Wo, Ho, p, ow, oh, b);
    float f[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = q * 8 + j;
      const int tap = k / 3, ci = k - tap * 3;
      const int kh = tap / 3, kw = tap - kh * 3;
      const int ih = 2 * oh + kh, iw = 2 * ow + kw;
      const bool ok = k < 27 && ih < H && iw < W;
      f[j] = ok ? (float)__ldg(x + ((b * H + ih) * (int)W + iw) * 3 + ci) * (1.0f / 255.0f) : 0.f;
    }
    stg16(col + p * 32 + q * 8, pack8(f));
  }
}

// Grid of a grid-stride streaming kernel capped at what is RESIDENT at once: a kernel that fits three blocks per SM and is
// launched with 4 x 148 runs a second wave of 148 lone blocks over the last quarter of the tensor (8 warps per SM).
// CRACKFL_EW_OCC=0 restores the fixed 4 x 148.
inline int blocks_per_sm_of(const void* kernel, int want) {
  static int mode = -1;
  if (mode < 0) { const char* e = getenv("CRACKFL_EW_OCC"); mode = e ? atoi(e) : 1; }
  if (!mode) return want;
  static std::mutex mu;
  static std::unordered_map<const void*, int> cache;
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(kernel);
  if (it == cache.end()) {
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, TPB, 0) != cudaSuccess || occ < 1) occ = want;
    it = cache.emplace(kernel, occ).first;
  }
  return it->second < want ? it->second : want;
}

inline int grid_for(long total, int cap = 0) {
  // blocks per SM of the grid-stride streaming kernels.  Measured on the full step (ms): 2: 1.354, 3: 1.325, 4: 1.311,
  // 6: 1.315, 8: 1.315, 12: 1.336, 16: 1.347, 24: 1.375, 48: 1.404 - about one resident wave is best (no tail of
  // partially filled waves, and room on the SMs for the kernels of the side streams).
  static int env_cap = -1;
  if (env_cap < 0) { const char* e = getenv("CRACKFL_EW_BLOCKS_PER_SM"); env_cap = e ? atoi(e) : 4; if (env_cap < 1) env_cap = 4; }
  if (cap == 0) cap = 148 * env_cap;
  long b = (total + TPB - 1) / TPB;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

inline int ppb_for(long P, int C) {
  // pixels per block for the reduction kernels: 2 blocks per SM measured best (fewer contended atomics per channel), >= one pass of all lanes
  const int lanes = TPB / (C >> 3);
  static int waves = -1;
  if (waves < 0) { const char* e = getenv("CRACKFL_REDUCE_BLOCKS_PER_SM"); waves = e ? atoi(e) : 2; if (waves < 1) waves = 1; }
  long target = (P + 148L * waves - 1) / (148L * waves);
  long ppb = ((target + lanes - 1) / lanes) * lanes;
  if (ppb < lanes * 4) ppb = lanes * 4;
  return (int)ppb;
}

}  // namespace

// the streaming kernels index with 32-bit integers (64-bit div/mod per element cost more than the memory traffic)
#define CK_FITS32(elems) do { if ((long)(elems) >= (1L << 31)) return (int)cudaErrorInvalidValue; } while (0)

extern "C" {

int crackfl_entry_conv_fwd(const void* x_u8, const float* w, const float* bias, void* y, int B, int H, int W, int C,
                           cudaStream_t st) {
  const long total = (long)B * (H / 2) * (W / 2) * (C / 8);
  ck::launch(entry_conv_fwd_k, grid_for(total), TPB, (27 * C + C) * sizeof(float), st, 
      (const uint8_t*)x_u8, w, bias, (__nv_bfloat16*)y, B, H, W, C);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_entry_im2col(const void* x_u8, void* col, int B, int H, int W, cudaStream_t st) {
  CK_FITS32((long)B * H * W * 16);
  const long total = (long)B * (H / 2) * (W / 2) * 4;
  ck::launch(entry_im2col_k, grid_for(total), TPB, 0, st, (const uint8_t*)x_u8, (__nv_bfloat16*)col, B, H, W);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_entry_conv_wgrad(const void* x_u8, const void* dy, float* dw, int B, int H, int W, int C, cudaStream_t st) {
  const long P = (long)B * (H / 2) * (W / 2);
  if (C <= 64 && TPB % C == 0) {
    const int ppb2 = 512;
    ck::launch(entry_conv_wgrad_v2_k, (int)((P + ppb2 - 1) / ppb2), TPB, (128 * 28 + 128 * C) * sizeof(float), st, 
        (const uint8_t*)x_u8, (const __nv_bfloat16*)dy, dw, B, H, W, C, ppb2);
    CK_LAUNCH_CHECK();
    return 0;
  }
  const int ppb = 512;
  dim3 blk(C / 8, 27);
  if (blk.x * blk.y > 1024) return (int)cudaErrorInvalidValue;
  ck::launch(entry_conv_wgrad_k, (int)((P + ppb - 1) / ppb), blk, 0, st, (const uint8_t*)x_u8, (const __nv_bfloat16*)dy, dw, B, H, W,
                                                                C, ppb);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_bn_stats(const void* y, float* stats, long P, int C, cudaStream_t st) {
  if (C % 8 || TPB % (C / 8)) return (int)cudaErrorInvalidValue;
  const int ppb = ppb_for(P, C);
  ck::launch(bn_stats_k, (int)((P + ppb - 1) / ppb), TPB, 0, st, (const __nv_bfloat16*)y, stats, P, C, ppb);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_bn_finalize(const float* stats, const float* gamma, const float* beta, float* run_mean, float* run_var,
                        float* bn, int C, float count, float eps, float momentum, cudaStream_t st) {
  ck::launch(bn_finalize_k, (C + 127) / 128, 128, 0, st, stats, gamma, beta, run_mean, run_var, bn, C, count, eps, momentum);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_bn_eval_prepare(const float* gamma, const float* beta, const float* run_mean, const float* run_var, float* bn,
                            int C, float eps, cudaStream_t st) {
  ck::launch(bn_eval_prepare_k, (C + 127) / 128, 128, 0, st, gamma, beta, run_mean, run_var, bn, C, eps);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_bn_eval_prepare_all(const float* arena, float* bn_params, const crackfl_bn_eval_desc* descs_dev, int n, float eps,
                                cudaStream_t st) {
  if (n <= 0) return 0;
  dim3 grid(2, n);
  ck::launch(bn_eval_prepare_all_k, grid, 128, 0, st, arena, bn_params, descs_dev, eps);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_bn_act(const void* y, const float* bn, void* out, void* even, int B, int H, int W, int C, int relu,
                   cudaStream_t st) {
  CK_FITS32((long)B * H * W * C);
  const long total = (long)B * H * W * (C / 8);
  ck::launch(bn_act_k, grid_for(total), TPB, 0, st, (const __nv_bfloat16*)y, bn, (__nv_bfloat16*)out, (__nv_bfloat16*)even, B, H, W,
                                            C, relu, BnFin{});
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_dw3x3_fwd(const void* x, const float* w, const float* bn, void* out, int B, int H, int W, int C, int mode,
                      cudaStream_t st) {
  const long total = (long)B * H * W * (C / 8);
  if (W % DWS == 0) {
    ck::launch(dw3x3_fwd_v2_k, grid_for(total / DWS, 148 * 8), TPB, 0, st, (const __nv_bfloat16*)x, w, bn, (__nv_bfloat16*)out, B, H, W,
                                                                  C, mode, BnFin{});
    CK_LAUNCH_CHECK();
    return 0;
  }
  ck::launch(dw3x3_fwd_k, grid_for(total), TPB, 0, st, (const __nv_bfloat16*)x, w, bn, (__nv_bfloat16*)out, B, H, W, C, mode);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_dw3x3_bwd(const void* g, const void* src, const float* w, const float* bn, const void* resid_t, void* dz,
                      float* dwk, int B, int H, int W, int C, int mode, int resid, cudaStream_t st) {
  if (C % 8 || TPB % (C / 8)) return (int)cudaErrorInvalidValue;
  const long P = (long)B * H * W;
  if (W % DWS == 0) {
    const long strips = P / DWS;
    ck::launch(dw3x3_dgrad_v2_k, grid_for(strips * (C / 8), 148 * 8), TPB, 0, st, 
        (const __nv_bfloat16*)g, (const __nv_bfloat16*)src, w, bn, (const __nv_bfloat16*)resid_t, (__nv_bfloat16*)dz, B, H, W, C,
        mode, resid);
    CK_LAUNCH_CHECK();
    const int lanes = TPB / (C / 8);
    long spb = (strips + 148 * 4 - 1) / (148 * 4);
    spb = ((spb + lanes - 1) / lanes) * lanes;
    if (spb < lanes) spb = lanes;
    ck::launch(dw3x3_wgrad_v2_k, (int)((strips + spb - 1) / spb), TPB, 0, st, (const __nv_bfloat16*)g, (const __nv_bfloat16*)src, bn,
                                                                      dwk, B, H, W, C, mode, (int)spb);
    CK_LAUNCH_CHECK();
    return 0;
  }
  const int ppb = ppb_for(P, C);
  ck::launch(dw3x3_bwd_k, (int)((P + ppb - 1) / ppb), TPB, 0, st, (const __nv_bfloat16*)g, (const __nv_bfloat16*)src, w, bn,
                                                          (const __nv_bfloat16*)resid_t, (__nv_bfloat16*)dz, dwk, B, H, W,
                                                          C, mode, resid, ppb);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_pool_add_fwd(const void* y, const float* bn, const void* res, void* out, void* even, void* relu_out, void* arg,
                         int B, int H, int W, int C, cudaStream_t st) {
  CK_FITS32((long)B * H * W * C);
  const long total = (long)B * (H / 2) * (W / 2) * (C / 8);
  ck::launch(pool_add_fwd_k, grid_for(total, 148 * blocks_per_sm_of((const void*)pool_add_fwd_k, 4)), TPB, 0, st, (const __nv_bfloat16*)y, bn, (const __nv_bfloat16*)res,
                                                  (__nv_bfloat16*)out, (__nv_bfloat16*)even, (__nv_bfloat16*)relu_out,
                                                  (uint8_t*)arg, B, H, W, C, BnFin{});
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_dec_combine_fwd(const void* y, const float* bn, const void* rl, void* s, void* a_next, int B, int H, int W,
                            int C, int rl_up, cudaStream_t st) {
  CK_FITS32((long)B * H * W * C * 4);
  const long total = (long)B * H * W * (C / 8);
  ck::launch(dec_combine_fwd_k, grid_for(total), TPB, 0, st, (const __nv_bfloat16*)y, bn, (const __nv_bfloat16*)rl,
                                                     (__nv_bfloat16*)s, (__nv_bfloat16*)a_next, B, H, W, C, rl_up, BnFin{});
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_down2sum(const void* in, const void* mask, const void* add, void* out, int B, int Ho, int Wo, int C,
                     cudaStream_t st) {
  CK_FITS32((long)B * Ho * Wo * C * 4);
  const long total = (long)B * Ho * Wo * (C / 8);
  ck::launch(down2sum_k, grid_for(total), TPB, 0, st, (const __nv_bfloat16*)in, (const __nv_bfloat16*)mask,
                                              (const __nv_bfloat16*)add, (__nv_bfloat16*)out, B, Ho, Wo, C);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_down2sum_bn(const void* in, const void* mask, const void* add, void* out, void* lowres, const void* y, const float* bn,
                        float* sums, int B, int Ho, int Wo, int C, cudaStream_t st) {
  CK_FITS32((long)B * Ho * Wo * C * 4);
  if ((Ho | Wo) & 1 || C % 8 || TPB % (C >> 3) || !y || !bn || !sums) return (int)cudaErrorInvalidValue;
  const int P2 = B * (Ho / 2) * (Wo / 2), lanes = TPB / (C >> 3);
  static int per_sm = -1;  // tuning knob
  if (per_sm < 0) { const char* e = getenv("CRACKFL_D2S_BLOCKS_PER_SM"); per_sm = e ? atoi(e) : 3; if (per_sm < 1) per_sm = 3; }
  int ipb = (P2 + 148 * per_sm - 1) / (148 * per_sm);
  ipb = (ipb + lanes - 1) / lanes * lanes;
  ck::launch(down2sum_bn_k, (P2 + ipb - 1) / ipb, TPB, 0, st, (const __nv_bfloat16*)in, (const __nv_bfloat16*)mask,
             (const __nv_bfloat16*)add, (__nv_bfloat16*)out, (__nv_bfloat16*)lowres, (const __nv_bfloat16*)y, bn, sums, B, Ho, Wo,
             C, ipb);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_relu_bwd_add(const void* g, const void* x, const void* add, void* out, long numel, cudaStream_t st) {
  CK_FITS32(numel);
  const long nvec = numel / 8;
  ck::launch(relu_bwd_add_k, grid_for(nvec), TPB, 0, st, (const __nv_bfloat16*)g, (const __nv_bfloat16*)x,
                                                 (const __nv_bfloat16*)add, (__nv_bfloat16*)out, nvec);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_bn_bwd_reduce(const void* g, const void* arg, const void* y, const float* bn, float* sums, int B, int H, int W,
                          int C, int mode, cudaStream_t st) {
  CK_FITS32((long)B * H * W * C);
  if (C % 8 || TPB % (C / 8)) return (int)cudaErrorInvalidValue;
  const long P = (long)B * H * W;
  const int ppb = ppb_for(P, C);
  const int grid = (int)((P + ppb - 1) / ppb);
  const __nv_bfloat16* gp = (const __nv_bfloat16*)g;
  const __nv_bfloat16* yp = (const __nv_bfloat16*)y;
  const uint8_t* ap = (const uint8_t*)arg;
  if (mode == 0) ck::launch(bn_bwd_reduce_k<0>, grid, TPB, 0, st, gp, ap, yp, bn, sums, P, H, W, C, ppb);
  else if (mode == 1) ck::launch(bn_bwd_reduce_k<1>, grid, TPB, 0, st, gp, ap, yp, bn, sums, P, H, W, C, ppb);
  else ck::launch(bn_bwd_reduce_k<2>, grid, TPB, 0, st, gp, ap, yp, bn, sums, P, H, W, C, ppb);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_bn_bwd_apply(const void* g, const void* arg, const void* y, const float* bn, const float* sums, void* dy,
                         float* dgamma, float* dbeta, int B, int H, int W, int C, int mode, cudaStream_t st) {
  CK_FITS32((long)B * H * W * C);
  const long P = (long)B * H * W;
  const void* kfn = mode == 0 ? (const void*)bn_bwd_apply_k<0> : mode == 1 ? (const void*)bn_bwd_apply_k<1> : (const void*)bn_bwd_apply_k<2>;
  const int grid = grid_for(P * (C / 8), 148 * blocks_per_sm_of(kfn, 4));
  const __nv_bfloat16* gp = (const __nv_bfloat16*)g;
  const __nv_bfloat16* yp = (const __nv_bfloat16*)y;
  const uint8_t* ap = (const uint8_t*)arg;
  __nv_bfloat16* dp = (__nv_bfloat16*)dy;
  if (mode == 0) ck::launch(bn_bwd_apply_k<0>, grid, TPB, 0, st, gp, ap, yp, bn, sums, dp, dgamma, dbeta, P, H, W, C);
  else if (mode == 1) ck::launch(bn_bwd_apply_k<1>, grid, TPB, 0, st, gp, ap, yp, bn, sums, dp, dgamma, dbeta, P, H, W, C);
  else ck::launch(bn_bwd_apply_k<2>, grid, TPB, 0, st, gp, ap, yp, bn, sums, dp, dgamma, dbeta, P, H, W, C);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_colsum(const void* g, float* out, long P, int C, cudaStream_t st) {
  CK_FITS32(P * C);
  if (C % 8 || TPB % (C / 8)) return (int)cudaErrorInvalidValue;
  const int ppb = ppb_for(P, C);
  ck::launch(colsum_k, (int)((P + ppb - 1) / ppb), TPB, 0, st, (const __nv_bfloat16*)g, out, P, C, ppb);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_head_fwd(const void* s, const float* w, const float* bias, const void* mask_u8, float* acc, float* prob, int B,
                     int Hl, int Wl, int C, int bce_clip, cudaStream_t st) {
  CK_FITS32((long)B * Hl * Wl * C);
  const long P = (long)B * Hl * Wl;
  ck::launch(head_fwd_k, grid_for(P * (C / 8), 148 * 4), TPB, C * sizeof(float), st, (const __nv_bfloat16*)s, w, bias,
                                                                   (const uint8_t*)mask_u8, acc, prob, B, Hl, Wl, C, bce_clip);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_head_bwd(const void* s, const float* w, const float* bias, const void* mask_u8, const float* sums, void* ds,
                     float* dw, float* db, int B, int Hl, int Wl, int C, int use_dice, float dice_smooth, float loss_scale,
                     int bce_clip, cudaStream_t st) {
  CK_FITS32((long)B * Hl * Wl * C);
  const long P = (long)B * Hl * Wl;
  ck::launch(head_bwd_k, grid_for(P * (C / 8), 148 * 4), TPB, (C + TPB * 8) * sizeof(float), st, 
      (const __nv_bfloat16*)s, w, bias, (const uint8_t*)mask_u8, sums, (__nv_bfloat16*)ds, dw, db, B, Hl, Wl, C, use_dice,
      dice_smooth, loss_scale, (float*)nullptr, bce_clip);
  CK_LAUNCH_CHECK();
  return 0;
}

// plain-BCE training: forward metrics (loss / accuracy / dice sums into acc_out[6]) and the backward in ONE pass over s
// (the BCE gradient needs nothing global but 1/N); with Dice the two-kernel sequence head_fwd -> head_bwd is required
int crackfl_head_fwd_bwd(const void* s, const float* w, const float* bias, const void* mask_u8, float* acc_out, void* ds,
                         float* dw, float* db, int B, int Hl, int Wl, int C, float loss_scale, int bce_clip,
                         cudaStream_t st) {
  CK_FITS32((long)B * Hl * Wl * C);
  const long P = (long)B * Hl * Wl;
  ck::launch(head_bwd_k, grid_for(P * (C / 8), 148 * 4), TPB, (C + TPB * 8) * sizeof(float), st,
      (const __nv_bfloat16*)s, w, bias, (const uint8_t*)mask_u8, (const float*)nullptr, (__nv_bfloat16*)ds, dw, db, B, Hl, Wl, C,
      0, 1.0f, loss_scale, acc_out, bce_clip);
  CK_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"

extern "C" {

// split backward of the depthwise conv so that the weight gradient can run on a side stream
int crackfl_dw3x3_dgrad(const void* g, const void* src, const float* w, const float* bn, const void* resid_t, void* dz, int B,
                        int H, int W, int C, int mode, int resid, cudaStream_t st) {
  if (W % DWS) return (int)cudaErrorInvalidValue;
  const long strips = (long)B * H * W / DWS;
  ck::launch(dw3x3_dgrad_v2_k, grid_for(strips * (C / 8), 148 * 8), TPB, 0, st, 
      (const __nv_bfloat16*)g, (const __nv_bfloat16*)src, w, bn, (const __nv_bfloat16*)resid_t, (__nv_bfloat16*)dz, B, H, W, C, mode,
      resid);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_dw3x3_wgrad(const void* g, const void* src, const float* bn, float* dwk, int B, int H, int W, int C, int mode,
                        cudaStream_t st) {
  if (W % DWS || C % 8 || 128 % (C / 8)) return (int)cudaErrorInvalidValue;
  const long strips = (long)B * H * W / DWS;
  const int lanes = 128 / (C / 8);
  long spb = (strips + 148 * 2 - 1) / (148 * 2);
  spb = ((spb + lanes - 1) / lanes) * lanes;
  if (spb < lanes) spb = lanes;
  ck::launch(dw3x3_wgrad_v3_k, (int)((strips + spb - 1) / spb), dim3(128, 3), 0, st, (const __nv_bfloat16*)g, (const __nv_bfloat16*)src,
                                                                             bn, dwk, B, H, W, C, mode, (int)spb);
  CK_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"

extern "C" {

int crackfl_bn_act_fin(const void* y, const crackfl_bn_fin* fin, void* out, void* even, int B, int H, int W, int C, int relu,
                       cudaStream_t st) {
  CK_FITS32((long)B * H * W * C);
  const long total = (long)B * H * W * (C / 8);
  ck::launch(bn_act_k, grid_for(total), TPB, 0, st, (const __nv_bfloat16*)y, nullptr, (__nv_bfloat16*)out, (__nv_bfloat16*)even, B, H, W,
                                            C, relu, *fin);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_dw3x3_fwd_fin(const void* x, const float* w, const crackfl_bn_fin* fin, void* out, int B, int H, int W, int C,
                          cudaStream_t st) {
  if (W % DWS) return (int)cudaErrorInvalidValue;
  const long total = (long)B * H * W * (C / 8);
  ck::launch(dw3x3_fwd_v2_k, grid_for(total / DWS, 148 * 8), TPB, 0, st, (const __nv_bfloat16*)x, w, nullptr, (__nv_bfloat16*)out, B, H,
                                                                W, C, 1, *fin);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_pool_add_fwd_fin(const void* y, const crackfl_bn_fin* fin, const void* res, void* out, void* even, void* relu_out,
                             void* arg, int B, int H, int W, int C, cudaStream_t st) {
  CK_FITS32((long)B * H * W * C);
  const long total = (long)B * (H / 2) * (W / 2) * (C / 8);
  ck::launch(pool_add_fwd_k, grid_for(total, 148 * blocks_per_sm_of((const void*)pool_add_fwd_k, 4)), TPB, 0, st, (const __nv_bfloat16*)y, nullptr, (const __nv_bfloat16*)res,
                                                  (__nv_bfloat16*)out, (__nv_bfloat16*)even, (__nv_bfloat16*)relu_out,
                                                  (uint8_t*)arg, B, H, W, C, *fin);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_dec_combine_fwd_fin(const void* y, const crackfl_bn_fin* fin, const void* rl, void* s, void* a_next, int B, int H,
                                int W, int C, int rl_up, cudaStream_t st) {
  CK_FITS32((long)B * H * W * C * 4);
  const long total = (long)B * H * W * (C / 8);
  ck::launch(dec_combine_fwd_k, grid_for(total), TPB, 0, st, (const __nv_bfloat16*)y, nullptr, (const __nv_bfloat16*)rl,
                                                     (__nv_bfloat16*)s, (__nv_bfloat16*)a_next, B, H, W, C, rl_up, *fin);
  CK_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"

extern "C" {

int crackfl_pool_bwd_bn(const void* g, const void* arg, const void* y, const float* bn, void* dz, float* sums, int B, int H,
                        int W, int C, cudaStream_t st) {
  CK_FITS32((long)B * H * W * C);
  if ((H | W) & 1 || TPB % (C >> 3)) return (int)cudaErrorInvalidValue;
  const int P2 = B * (H / 2) * (W / 2), lanes = TPB / (C >> 3);
  static int per_sm = -1;  // tuning knob
  if (per_sm < 0) { const char* e = getenv("CRACKFL_POOLBWD_BLOCKS_PER_SM"); per_sm = e ? atoi(e) : 2; if (per_sm < 1) per_sm = 2; }
  int ipb = (P2 + 148 * per_sm - 1) / (148 * per_sm);
  ipb = (ipb + lanes - 1) / lanes * lanes;
  ck::launch(pool_bwd_bn_k, (P2 + ipb - 1) / ipb, TPB, 0, st, (const __nv_bfloat16*)g, (const uint8_t*)arg,
             (const __nv_bfloat16*)y, bn, (__nv_bfloat16*)dz, sums, B, H, W, C, ipb);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_pool_bwd(const void* g, const void* arg, void* dz, int B, int H, int W, int C, cudaStream_t st) {
  CK_FITS32((long)B * H * W * C);
  const long P = (long)B * H * W;
  ck::launch(pool_bwd_k, grid_for(P * (C / 8)), TPB, 0, st, (const __nv_bfloat16*)g, (const uint8_t*)arg, (__nv_bfloat16*)dz, P, H, W, C);
  CK_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------
// Allocation-free training-mode BatchNorm(+ReLU) forward.
//
// Who needs this: code that drives this library from a place where it cannot obtain device memory of its own.
// Examples are a plug-in running inside a host process that has handed the whole device heap to another
// framework's pool, a step function that is only allowed to touch buffers it was given at start-up (fixed
// memory budget, no allocator in the hot path), and clean-up or recovery code that runs after an
// out-of-memory condition, when asking for more memory is the one thing that cannot work.  The ordinary path
// (crackfl_bn_stats + crackfl_bn_finalize + crackfl_bn_act) expects the caller to bring 6*C floats of scratch
// for the per-channel sums and the scale/shift table; here that scratch lives in the library image instead.
//
// The scratch is ONE fixed array per GPU (every GPU that loads the library carries it until the process ends):
// 6 * CRACKFL_NOALLOC_MAX_CHANNELS floats = 48 KiB.  Consequences the caller has to know:
//   * channel counts above CRACKFL_NOALLOC_MAX_CHANNELS are refused (message on stderr, cudaErrorInvalidValue);
//   * all calls on a GPU share the array, and it keeps whatever the previous call left in it.  The part that
//     must start at zero (the sums) is cleared on the stream at the start of every call; two calls whose GPU
//     work overlaps in time (different streams, no ordering between them) race on it and give wrong results;
//   * the first call on each device looks the array's device address up (cudaGetSymbolAddress, with that
//     device current).  Make that first call during start-up, before any stream capture is begun.
// Nothing in here calls cudaMalloc, cudaFree or any routine that allocates.
static __device__ __align__(256) float g_noalloc_bn_scratch[6 * CRACKFL_NOALLOC_MAX_CHANNELS];

namespace {
constexpr int kMaxDevices = 64;
float* g_noalloc_ptr[kMaxDevices] = {nullptr};

// In host code `g_noalloc_bn_scratch` is only a handle for the device symbol, not an address a kernel can use.
// The per-device cache below is a plain array: two host threads that make their first call on the same device at the
// same moment may both do the look-up; they store the same value, so the only cost is the repeated look-up.
int noalloc_scratch_for_current_device(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kMaxDevices) return (int)cudaErrorInvalidDevice;
  if (g_noalloc_ptr[dev] == nullptr) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_noalloc_bn_scratch);
    if (e != cudaSuccess) return (int)e;
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace

extern "C" {

int crackfl_noalloc_max_channels(void) { return CRACKFL_NOALLOC_MAX_CHANNELS; }

int crackfl_bn_train_fwd_noalloc(const void* y, const float* gamma, const float* beta, float* run_mean, float* run_var,
                                 void* out, int B, int H, int W, int C, int relu, float eps, float momentum,
                                 cudaStream_t st) {
  if (C > CRACKFL_NOALLOC_MAX_CHANNELS) {
    fprintf(stderr,
            "crackfl_bn_train_fwd_noalloc: C=%d exceeds the built-in scratch (max %d channels); use crackfl_bn_stats/"
            "crackfl_bn_finalize/crackfl_bn_act with caller-provided scratch instead\n",
            C, CRACKFL_NOALLOC_MAX_CHANNELS);
    return (int)cudaErrorInvalidValue;
  }
  if (C < 8 || C % 8 || TPB % (C / 8) || B < 1 || H < 1 || W < 1) {
    fprintf(stderr, "crackfl_bn_train_fwd_noalloc: unsupported shape B=%d H=%d W=%d C=%d (C must be a multiple of 8 "
                    "with C/8 dividing %d)\n", B, H, W, C, TPB);
    return (int)cudaErrorInvalidValue;
  }
  CK_FITS32((long)B * H * W * C);  // bn_act_k indexes with 32-bit ints, as crackfl_bn_act does
  float* scratch = nullptr;
  int rc = noalloc_scratch_for_current_device(&scratch);
  if (rc) return rc;
  float* stats = scratch;                                  // [sum C][sumsq C]   must start at zero
  float* bn = scratch + 2 * CRACKFL_NOALLOC_MAX_CHANNELS;  // [scale][shift][mean][invstd], fully rewritten
  cudaError_t e = cudaMemsetAsync(stats, 0, 2 * (size_t)C * sizeof(float), st);
  if (e != cudaSuccess) return (int)e;
  const long P = (long)B * H * W;
  const int ppb = ppb_for(P, C);
  ck::launch(bn_stats_k, (int)((P + ppb - 1) / ppb), TPB, 0, st, (const __nv_bfloat16*)y, stats, P, C, ppb);
  CK_LAUNCH_CHECK();
  ck::launch(bn_finalize_k, (C + 127) / 128, 128, 0, st, stats, gamma, beta, run_mean, run_var, bn, C, (float)P, eps, momentum);
  CK_LAUNCH_CHECK();
  ck::launch(bn_act_k, grid_for(P * (C / 8)), TPB, 0, st, (const __nv_bfloat16*)y, bn, (__nv_bfloat16*)out, nullptr, B, H, W, C,
                                                 relu, BnFin{});
  CK_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
