// CUDA-core implicit-GEMM convolution and weight-gradient kernels.
//
// These are the *reference implementation inside the library*: simple shared-memory tiled
// kernels used (a) as the on-device cross-check for the tcgen05 kernels in conv_tc.cu and
// (b) for shapes the tensor-core path does not take.  Same operand layouts as the tcgen05 path:
//   activations  X [B,H,W,Cin]  bf16 NHWC
//   weights      Wk[N][taps*Cin] bf16, K index = tap*Cin + ci  ("K-major"), taps = 1 or 9
//   output       Y [B,H,W,N]    bf16
// A 3x3 stride-1 "same" conv, Keras Conv2DTranspose(stride 1, same) (flipped taps / swapped
// channels, client_fit_model.py:129,133), their dgrad, and all 1x1 convs share this kernel; the
// weight-prep kernel (optim.cu) lays the taps out so that tap t always reads pixel p + (t/3-1, t%3-1).
#include "common.cuh"
#include "../include/crackfl.h"

using namespace ck;

namespace {

constexpr int BM = 64, BN = 64, BK = 32;

__global__ void __launch_bounds__(256) conv_simt_fwd_k(const __nv_bfloat16* __restrict__ X, const __nv_bfloat16* __restrict__ Wk,
                                                       const float* __restrict__ bias, const __nv_bfloat16* __restrict__ addt,
                                                       __nv_bfloat16* __restrict__ Y, int B, int H, int W, int Cin, int N,
                                                       int taps) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const long M = (long)B * H * W;
  const long m0 = (long)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int tid = threadIdx.x;
  const int tx = tid % 16, ty = tid / 16;  // 16 x 16 threads, 4 x 4 outputs each
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  // loader mapping: row = tid / 4 (0..63), 8-element chunk = tid % 4
  const int lr = tid >> 2, lc = tid & 3;
  const long pm = m0 + lr;
  int pb = 0, ph = 0, pw = 0;
  const bool prow = pm < M;
  if (prow) {
    pw = (int)(pm % W);
    ph = (int)((pm / W) % H);
    pb = (int)(pm / ((long)W * H));
  }
  const int Ktot = taps * Cin;
  for (int t = 0; t < taps; ++t) {
    const int dh = taps == 9 ? t / 3 - 1 : 0, dw = taps == 9 ? t % 3 - 1 : 0;
    const int ih = ph + dh, iw = pw + dw;
    const bool inb = prow && ih >= 0 && ih < H && iw >= 0 && iw < W;
    for (int c0 = 0; c0 < Cin; c0 += BK) {
      float a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, b[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      if (inb) unpack8(ldg16(X + (((long)pb * H + ih) * W + iw) * Cin + c0 + lc * 8), a);
      if (n0 + lr < N) unpack8(ldg16(Wk + (long)(n0 + lr) * Ktot + t * Cin + c0 + lc * 8), b);
      __syncthreads();
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        As[lc * 8 + j][lr] = a[j];
        Bs[lc * 8 + j][lr] = b[j];
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < BK; ++k) {
        float av[4], bv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) av[i] = As[k][ty * 4 + i];
#pragma unroll
        for (int j = 0; j < 4; ++j) bv[j] = Bs[k][tx * 4 + j];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long m = m0 + ty * 4 + i;
    if (m >= M) continue;
    const int n = n0 + tx * 4;
    if (n >= N) continue;
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = acc[i][j] + (bias ? bias[n + j] : 0.f);
    if (addt) {
      const uint2 t = __ldg(reinterpret_cast<const uint2*>(addt + m * N + n));
      o[0] += __uint_as_float(t.x << 16); o[1] += __uint_as_float(t.x & 0xffff0000u);
      o[2] += __uint_as_float(t.y << 16); o[3] += __uint_as_float(t.y & 0xffff0000u);
    }
    uint2 v;
    v.x = pack2(o[0], o[1]);
    v.y = pack2(o[2], o[3]);
    *reinterpret_cast<uint2*>(Y + m * N + n) = v;
  }
}

// dW[t][ci][co] += sum_p X[p + off(t)][ci] * dY[p][co]; written with arbitrary strides so it lands
// directly in the Keras-layout gradient arena (HWIO for Conv2D, (kh,kw,out,in) flipped for ConvT).
constexpr int WT = 32, WP = 32;
__global__ void __launch_bounds__(256) wgrad_simt_k(const __nv_bfloat16* __restrict__ X, const __nv_bfloat16* __restrict__ dY,
                                                    float* __restrict__ dW, int B, int H, int W, int Cin, int Cout, int taps,
                                                    long s_tap, long s_ci, long s_co, int pix_per_block) {
  __shared__ float Xs[WP][WT + 1];
  __shared__ float Ys[WP][WT + 1];
  const int nci = Cin / WT, nco = Cout / WT;
  int tile = blockIdx.x;
  const int co0 = (tile % nco) * WT; tile /= nco;
  const int ci0 = (tile % nci) * WT; tile /= nci;
  const int t = tile;
  const int dh = taps == 9 ? t / 3 - 1 : 0, dw = taps == 9 ? t % 3 - 1 : 0;
  const long M = (long)B * H * W;
  const long p0 = (long)blockIdx.y * pix_per_block;
  const long p1 = min(M, p0 + pix_per_block);
  const int tid = threadIdx.x;
  const int ti = tid / 16, tj = tid % 16;  // 2x2 outputs: ci = ti*2+{0,1}, co = tj*2+{0,1}
  const int lr = tid >> 2, lc = tid & 3;   // loader: pixel row lr (0..63 -> use 0..31 twice), chunk lc
  float acc[2][2] = {{0, 0}, {0, 0}};
  for (long pb0 = p0; pb0 < p1; pb0 += WP) {
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const int r = lr & 31;
    const long p = pb0 + r;
    const bool isx = lr < 32;
    if (p < p1) {
      if (isx) {
        const int pw = (int)(p % W);
        const int ph = (int)((p / W) % H);
        const long pbb = p / ((long)W * H);
        const int ih = ph + dh, iw = pw + dw;
        if (ih >= 0 && ih < H && iw >= 0 && iw < W)
          unpack8(ldg16(X + ((pbb * H + ih) * W + iw) * Cin + ci0 + lc * 8), a);
      } else {
        unpack8(ldg16(dY + p * Cout + co0 + lc * 8), a);
      }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 8; ++j) (isx ? Xs : Ys)[r][lc * 8 + j] = a[j];
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < WP; ++k) {
      const float x0 = Xs[k][ti * 2], x1 = Xs[k][ti * 2 + 1];
      const float y0 = Ys[k][tj * 2], y1 = Ys[k][tj * 2 + 1];
      acc[0][0] = fmaf(x0, y0, acc[0][0]); acc[0][1] = fmaf(x0, y1, acc[0][1]);
      acc[1][0] = fmaf(x1, y0, acc[1][0]); acc[1][1] = fmaf(x1, y1, acc[1][1]);
    }
  }
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j)
      atomicAdd(dW + t * s_tap + (long)(ci0 + ti * 2 + i) * s_ci + (long)(co0 + tj * 2 + j) * s_co, acc[i][j]);
}

}  // namespace

extern "C" {

int crackfl_conv_simt_fwd(const void* X, const void* Wk, const float* bias, const void* add, void* Y, int B, int H, int W,
                          int Cin, int N, int taps, cudaStream_t st) {
  if (Cin % 32 || N % 4 || (taps != 1 && taps != 9)) return (int)cudaErrorInvalidValue;
  const long M = (long)B * H * W;
  dim3 grid((unsigned)((M + BM - 1) / BM), (N + BN - 1) / BN);
  conv_simt_fwd_k<<<grid, 256, 0, st>>>((const __nv_bfloat16*)X, (const __nv_bfloat16*)Wk, bias, (const __nv_bfloat16*)add,
                                        (__nv_bfloat16*)Y, B, H, W, Cin, N, taps);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_wgrad_simt(const void* X, const void* dY, float* dW, int B, int H, int W, int Cin, int Cout, int taps,
                       long s_tap, long s_ci, long s_co, cudaStream_t st) {
  if (Cin % 32 || Cout % 32 || (taps != 1 && taps != 9)) return (int)cudaErrorInvalidValue;
  const long M = (long)B * H * W;
  const int tiles = (Cin / WT) * (Cout / WT) * taps;
  long splits = (148L * 8 + tiles - 1) / tiles;
  const long maxsplits = (M + 255) / 256;
  if (splits > maxsplits) splits = maxsplits;
  if (splits < 1) splits = 1;
  long ppb = (M + splits - 1) / splits;
  ppb = (ppb + WP - 1) / WP * WP;
  dim3 grid(tiles, (unsigned)((M + ppb - 1) / ppb));
  wgrad_simt_k<<<grid, 256, 0, st>>>((const __nv_bfloat16*)X, (const __nv_bfloat16*)dY, dW, B, H, W, Cin, Cout, taps, s_tap,
                                     s_ci, s_co, (int)ppb);
  CK_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
