// Depthwise 3x3 kernels with TMA-staged halo tiles (second generation).
//
// ncu showed the register-strip kernels in elementwise.cu latency-bound (20 % occupancy, ~1 TB/s): every thread
// chases its own 18 global loads.  Here one elected thread per CTA issues a single bulk TMA copy of a
// (TY+2) x 18-pixel x CB-channel halo tile (out-of-bounds pixels arrive as zeros) into a double-buffered shared
// memory ring, and the 256 threads only do shared-memory reads + FMAs: thread = (4 channels, one column), sliding a
// 3x3 register window down TY/RG rows (3.4 LDS.64 per output instead of 9 global loads), so the HBM stream is one
// asynchronous bulk read per tile and the next tile's copy overlaps this tile's arithmetic.
//   dw_stencil_k : forward  (x -> relu / relu(bn) -> 3x3 depthwise)  and, with flipped taps, the input gradient
//                  (g -> 3x3^T -> * relu-mask of the forward input (+ residual-path gradient)).
//   dw_wgrad_k   : weight gradient  dW[t][c] = sum_p a[p + off(t)][c] * g[p][c]  (a from a halo tile, g from a plain tile).
// ref: SeparableConv2D depthwise stage, client_fit_model.py:109,113.
#include "common.cuh"
#include "tc_common.cuh"
#include "../include/crackfl.h"

using namespace ck;

namespace {

constexpr int TX = 16;      // tile columns
constexpr int HW = TX + 2;  // halo width
typedef crackfl_bn_fin BnFin;

struct DwArgs {
  const float* w;               // [9][C]
  const float* bn;              // [scale C][shift C]... for act_mode 1 (when fin.stats == 0)
  BnFin fin;                    // optional fused finalize (forward, act_mode 1)
  __nv_bfloat16* out;
  const __nv_bfloat16* src;     // dgrad: forward input, gives the relu mask
  const float* src_bn;          // dgrad: bn of src when src_mode == 1
  const __nv_bfloat16* resid_t; // dgrad: residual-path gradient at half resolution
  float* dwk;                   // wgrad output [9][C]
  int B, H, W, C, CB;
  int act_mode;                 // 0 relu(x), 1 relu(x*scale+shift), 2 identity
  int flip;                     // use w[8 - t] (input gradient)
  int src_mode, resid;          // dgrad epilogue (see dw3x3_bwd_k in elementwise.cu)
  int tiles_x, tiles_y, total_tiles;
};

__device__ __forceinline__ void load4(const uint8_t* smem_row, int off_bytes, float f[4]) {
  const uint2 v = *reinterpret_cast<const uint2*>(smem_row + off_bytes);
  f[0] = __uint_as_float(v.x << 16); f[1] = __uint_as_float(v.x & 0xffff0000u);
  f[2] = __uint_as_float(v.y << 16); f[3] = __uint_as_float(v.y & 0xffff0000u);
}

__device__ __forceinline__ void coeffs4(const BnFin& f, const float* __restrict__ bn, int C, int c, float sc[4], float sh[4]) {
  if (f.stats == nullptr) {
#pragma unroll
    for (int j = 0; j < 4; ++j) { sc[j] = bn[c + j]; sh[j] = bn[C + c + j]; }
    return;
  }
  const float inv = 1.0f / f.count;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float mean = f.stats[c + j] * inv;
    const float var = fmaxf(f.stats[C + c + j] * inv - mean * mean, 0.f);
    const float invstd = rsqrtf(var + f.eps);
    sc[j] = f.gamma[c + j] * invstd;
    sh[j] = f.beta[c + j] - mean * sc[j];
  }
}

__device__ __forceinline__ void fin_publish(const BnFin& f, int C) {
  if (f.stats == nullptr || blockIdx.x != 0) return;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float mean = f.stats[c] / f.count;
    const float var = fmaxf(f.stats[C + c] / f.count - mean * mean, 0.f);
    const float invstd = rsqrtf(var + f.eps);
    const float scl = f.gamma[c] * invstd;
    f.bn_out[c] = scl; f.bn_out[C + c] = f.beta[c] - mean * scl;
    f.bn_out[2 * C + c] = mean; f.bn_out[3 * C + c] = invstd;
    const float unb = f.count > 1.f ? var * (f.count / (f.count - 1.f)) : var;
    f.run_mean[c] = f.run_mean[c] * f.momentum + mean * (1.f - f.momentum);
    f.run_var[c] = f.run_var[c] * f.momentum + unb * (1.f - f.momentum);
  }
}

// TY = tile rows; threads: (c4 = CB/4) x 16 columns x RG row groups = 256 (RG = 1 for CB = 64, 2 for CB = 32)
template <int TY, int RG>
__global__ void __launch_bounds__(256, 2) dw_stencil_k(const __grid_constant__ CUtensorMap tmX, const DwArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const int CB = a.CB, rowb = CB * 2;
  const uint32_t stage_bytes = (uint32_t)((TY + 2) * HW * rowb);
  const uint32_t bar0 = base + 2u * stage_bytes;
  const int nc4 = CB >> 2;
  constexpr int rows = TY / RG;
  const int c4 = threadIdx.x % nc4, col = (threadIdx.x / nc4) % TX, rg = threadIdx.x / (nc4 * TX);
  const int spatial = a.tiles_x * a.tiles_y * a.B;

  if (threadIdx.x == 0) {
    mbar_init(bar0, 1);
    mbar_init(bar0 + 8u, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&tmX);
  }
  pdl_enter();  // barrier init / descriptor prefetch above overlap the previous kernel's tail
  fin_publish(a.fin, a.C);
  __syncthreads();

  auto issue = [&](int tile, int stage) {
    const int cblk = tile / spatial;
    int sp = tile % spatial;
    const int tx_i = sp % a.tiles_x; sp /= a.tiles_x;
    const int ty_i = sp % a.tiles_y;
    const int b = sp / a.tiles_y;
    mbar_expect_tx(bar0 + 8u * stage, stage_bytes);
    tma_load_4d(base + stage * stage_bytes, &tmX, bar0 + 8u * stage, cblk * CB, tx_i * TX - 1, ty_i * TY - 1, b);
  };
  if (threadIdx.x == 0 && (int)blockIdx.x < a.total_tiles) issue(blockIdx.x, 0);

  int cur_cblk = -1;
  float wk[9][4], sc[4], sh[4], msc[4], msh[4];
  int it = 0;
  for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
    const int stage = it & 1;
    const int nxt = tile + gridDim.x;
    if (threadIdx.x == 0 && nxt < a.total_tiles) issue(nxt, stage ^ 1);  // its last readers passed the barrier below
    const int cblk = tile / spatial;
    int sp = tile % spatial;
    const int tx_i = sp % a.tiles_x; sp /= a.tiles_x;
    const int ty_i = sp % a.tiles_y;
    const int b = sp / a.tiles_y;
    const int cg = cblk * CB + c4 * 4;  // first global channel of this thread
    if (cblk != cur_cblk) {
      cur_cblk = cblk;
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(a.w + (a.flip ? 8 - t : t) * a.C + cg));
        wk[t][0] = v.x; wk[t][1] = v.y; wk[t][2] = v.z; wk[t][3] = v.w;
      }
      if (a.act_mode == 1) coeffs4(a.fin, a.bn, a.C, cg, sc, sh);
      if (a.src && a.src_mode == 1) {
#pragma unroll
        for (int j = 0; j < 4; ++j) { msc[j] = a.src_bn[cg + j]; msh[j] = a.src_bn[a.C + cg + j]; }
      }
    }
    mbar_wait(bar0 + 8u * stage, (uint32_t)(it >> 1) & 1u);
    const uint8_t* tile_s = gbase + stage * stage_bytes;
    const int w0 = tx_i * TX, h0 = ty_i * TY + rg * rows;
    const int gw = w0 + col;
    float win[3][3][4];
#pragma unroll
    for (int r = 0; r < rows + 2; ++r) {
      const int gh = h0 - 1 + r;
      const uint8_t* rowp = tile_s + (size_t)((rg * rows + r) * HW) * rowb + c4 * 8;
      // slide: row r goes into slot r % 3
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        float v[4];
        load4(rowp, (col + kx) * rowb, v);
        const bool inside = gh >= 0 && gh < a.H && (gw - 1 + kx) >= 0 && (gw - 1 + kx) < a.W;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float x = v[j];
          if (a.act_mode == 1) x = fmaf(x, sc[j], sh[j]);
          if (a.act_mode != 2) x = fmaxf(x, 0.f);
          win[r % 3][kx][j] = inside ? x : 0.f;
        }
      }
      if (r >= 2) {
        const int oh = h0 + r - 2;
        if (oh < a.H && gw < a.W) {
          float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int kh = 0; kh < 3; ++kh)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx)
#pragma unroll
              for (int j = 0; j < 4; ++j) acc[j] = fmaf(win[(r - 2 + kh) % 3][kx][j], wk[kh * 3 + kx][j], acc[j]);
          const long o = (((long)b * a.H + oh) * a.W + gw) * a.C + cg;
          if (a.src) {
            // input-gradient epilogue: relu mask of the forward input, residual-path gradient on even pixels
            const uint2 sv = __ldg(reinterpret_cast<const uint2*>(a.src + o));
            float s4[4] = {__uint_as_float(sv.x << 16), __uint_as_float(sv.x & 0xffff0000u), __uint_as_float(sv.y << 16),
                           __uint_as_float(sv.y & 0xffff0000u)};
            float r4[4] = {0.f, 0.f, 0.f, 0.f};
            if (a.resid && !(oh & 1) && !(gw & 1)) {
              const uint2 rv = __ldg(reinterpret_cast<const uint2*>(
                  a.resid_t + (((long)b * (a.H >> 1) + (oh >> 1)) * (a.W >> 1) + (gw >> 1)) * a.C + cg));
              r4[0] = __uint_as_float(rv.x << 16); r4[1] = __uint_as_float(rv.x & 0xffff0000u);
              r4[2] = __uint_as_float(rv.y << 16); r4[3] = __uint_as_float(rv.y & 0xffff0000u);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float pre = a.src_mode == 1 ? fmaf(s4[j], msc[j], msh[j]) : s4[j];
              const float m = pre > 0.f ? 1.f : 0.f;
              acc[j] = (a.resid == 2) ? m * (acc[j] + r4[j]) : m * acc[j] + r4[j];
            }
          }
          uint2 pk;
          pk.x = pack2(acc[0], acc[1]);
          pk.y = pack2(acc[2], acc[3]);
          *reinterpret_cast<uint2*>(a.out + o) = pk;
        }
      }
    }
    __syncthreads();  // every thread is done with this stage before it is refilled two iterations from now
  }
}

// weight gradient: a-halo tile + g tile per stage
template <int TY, int RG>
__global__ void __launch_bounds__(256, 2) dw_wgrad_k(const __grid_constant__ CUtensorMap tmA,
                                                     const __grid_constant__ CUtensorMap tmG, const DwArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const int CB = a.CB, rowb = CB * 2;
  const uint32_t a_bytes = (uint32_t)((TY + 2) * HW * rowb), g_bytes = (uint32_t)(TY * TX * rowb);
  const uint32_t stage_bytes = a_bytes + g_bytes;
  const uint32_t bar0 = base + 2u * stage_bytes;
  float* red = reinterpret_cast<float*>(gbase + 2u * stage_bytes + 64);  // [256][4] floats
  const int nc4 = CB >> 2;
  constexpr int rows = TY / RG;
  const int c4 = threadIdx.x % nc4, col = (threadIdx.x / nc4) % TX, rg = threadIdx.x / (nc4 * TX);
  const int spatial = a.tiles_x * a.tiles_y * a.B;

  if (threadIdx.x == 0) {
    mbar_init(bar0, 1);
    mbar_init(bar0 + 8u, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmG);
  }
  __syncthreads();
  pdl_enter();

  auto issue = [&](int tile, int stage) {
    const int cblk = tile / spatial;
    int sp = tile % spatial;
    const int tx_i = sp % a.tiles_x; sp /= a.tiles_x;
    const int ty_i = sp % a.tiles_y;
    const int b = sp / a.tiles_y;
    mbar_expect_tx(bar0 + 8u * stage, stage_bytes);
    tma_load_4d(base + stage * stage_bytes, &tmA, bar0 + 8u * stage, cblk * CB, tx_i * TX - 1, ty_i * TY - 1, b);
    tma_load_4d(base + stage * stage_bytes + a_bytes, &tmG, bar0 + 8u * stage, cblk * CB, tx_i * TX, ty_i * TY, b);
  };
  if (threadIdx.x == 0 && (int)blockIdx.x < a.total_tiles) issue(blockIdx.x, 0);

  float acc[9][4], sc[4], sh[4];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[t][j] = 0.f;
  int cur_cblk = -1;

  auto flush = [&](int cblk) {
    // reduce the 16 columns x RG row groups that share (c4) and add into dwk[t][cblk*CB + c4*4 + j]
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      __syncthreads();
#pragma unroll
      for (int j = 0; j < 4; ++j) red[threadIdx.x * 4 + j] = acc[t][j];
      __syncthreads();
      if (threadIdx.x < nc4 * 4) {
        const int cc = threadIdx.x >> 2, j = threadIdx.x & 3;
        float v = 0.f;
        for (int q = 0; q < 256 / nc4; ++q) v += red[(q * nc4 + cc) * 4 + j];
        atomicAdd(a.dwk + (long)t * a.C + cblk * CB + cc * 4 + j, v);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[t][j] = 0.f;
    }
  };

  int it = 0;
  for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
    const int stage = it & 1;
    const int nxt = tile + gridDim.x;
    if (threadIdx.x == 0 && nxt < a.total_tiles) issue(nxt, stage ^ 1);
    const int cblk = tile / spatial;
    int sp = tile % spatial;
    const int tx_i = sp % a.tiles_x; sp /= a.tiles_x;
    const int ty_i = sp % a.tiles_y;
    const int cg = cblk * CB + c4 * 4;
    if (cblk != cur_cblk) {
      if (cur_cblk >= 0) flush(cur_cblk);
      cur_cblk = cblk;
      if (a.act_mode == 1) coeffs4(a.fin, a.bn, a.C, cg, sc, sh);
    }
    mbar_wait(bar0 + 8u * stage, (uint32_t)(it >> 1) & 1u);
    const uint8_t* tile_a = gbase + stage * stage_bytes;
    const uint8_t* tile_g = tile_a + a_bytes;
    const int w0 = tx_i * TX, h0 = ty_i * TY + rg * rows;
    const int gw = w0 + col;
    float win[3][3][4];
#pragma unroll
    for (int r = 0; r < rows + 2; ++r) {
      const int gh = h0 - 1 + r;
      const uint8_t* rowp = tile_a + (size_t)((rg * rows + r) * HW) * rowb + c4 * 8;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        float v[4];
        load4(rowp, (col + kx) * rowb, v);
        const bool inside = gh >= 0 && gh < a.H && (gw - 1 + kx) >= 0 && (gw - 1 + kx) < a.W;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float x = v[j];
          if (a.act_mode == 1) x = fmaf(x, sc[j], sh[j]);
          x = fmaxf(x, 0.f);
          win[r % 3][kx][j] = inside ? x : 0.f;
        }
      }
      if (r >= 2) {
        float g4[4];  // out-of-image g pixels are TMA zero fill: they contribute nothing
        load4(tile_g + (size_t)((rg * rows + r - 2) * TX + col) * rowb + c4 * 8, 0, g4);
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
          for (int kx = 0; kx < 3; ++kx)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[kh * 3 + kx][j] = fmaf(win[(r - 2 + kh) % 3][kx][j], g4[j], acc[kh * 3 + kx][j]);
      }
    }
    __syncthreads();
  }
  if (cur_cblk >= 0) flush(cur_cblk);
}

int g_sms = 0;
int num_sms() {
  if (g_sms <= 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_sms <= 0) g_sms = 148;
  }
  return g_sms;
}

bool shape_ok(int C) { return C % 32 == 0 && (C <= 64 || C % 64 == 0); }

}  // namespace

extern "C" {

// forward (src == NULL) or input gradient (src != NULL, flip = 1, act_mode = 2)
int crackfl_dw3x3_tma(const void* x, const float* w, const float* bn, const crackfl_bn_fin* fin, void* out, const void* src,
                      const float* src_bn, const void* resid_t, int B, int H, int W, int C, int act_mode, int flip,
                      int src_mode, int resid, cudaStream_t st) {
  if (!shape_ok(C)) return (int)cudaErrorInvalidValue;
  constexpr int TY = 16;
  DwArgs a;
  memset(&a, 0, sizeof(a));
  a.w = w; a.bn = bn; if (fin) a.fin = *fin;
  a.out = (__nv_bfloat16*)out; a.src = (const __nv_bfloat16*)src; a.src_bn = src_bn; a.resid_t = (const __nv_bfloat16*)resid_t;
  a.B = B; a.H = H; a.W = W; a.C = C; a.CB = C >= 64 ? 64 : 32;
  a.act_mode = act_mode; a.flip = flip; a.src_mode = src_mode; a.resid = resid;
  a.tiles_x = (W + TX - 1) / TX; a.tiles_y = (H + TY - 1) / TY;
  a.total_tiles = a.tiles_x * a.tiles_y * B * (C / a.CB);
  CUtensorMap tm;
  const int d[4] = {C, W, H, B}, bx[4] = {a.CB, HW, TY + 2, 1};
  int rc = ck::get_tensor_map(x, 4, d, bx, 0, &tm);
  if (rc) return rc;
  const size_t smem = 2 * (size_t)(TY + 2) * HW * a.CB * 2 + 128 + 64;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(dw_stencil_k<TY, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dw_stencil_k<TY, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  int grid = 2 * num_sms();
  if (grid > a.total_tiles) grid = a.total_tiles;
  if (a.CB == 64) ck::launch(dw_stencil_k<TY, 1>, grid, 256, smem, st, tm, a);
  else ck::launch(dw_stencil_k<TY, 2>, grid, 256, smem, st, tm, a);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_dw3x3_wgrad_tma(const void* g, const void* src, const float* bn, float* dwk, int B, int H, int W, int C, int mode,
                            cudaStream_t st) {
  if (!shape_ok(C)) return (int)cudaErrorInvalidValue;
  constexpr int TY = 8;
  DwArgs a;
  memset(&a, 0, sizeof(a));
  a.bn = bn; a.dwk = dwk;
  a.B = B; a.H = H; a.W = W; a.C = C; a.CB = C >= 64 ? 64 : 32;
  a.act_mode = mode;
  a.tiles_x = (W + TX - 1) / TX; a.tiles_y = (H + TY - 1) / TY;
  a.total_tiles = a.tiles_x * a.tiles_y * B * (C / a.CB);
  CUtensorMap tmA, tmG;
  const int d[4] = {C, W, H, B}, ba[4] = {a.CB, HW, TY + 2, 1}, bg[4] = {a.CB, TX, TY, 1};
  int rc = ck::get_tensor_map(src, 4, d, ba, 0, &tmA);
  if (rc) return rc;
  rc = ck::get_tensor_map(g, 4, d, bg, 0, &tmG);
  if (rc) return rc;
  const size_t smem = 2 * ((size_t)(TY + 2) * HW + TY * TX) * a.CB * 2 + 128 + 64 + 256 * 4 * sizeof(float);
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(dw_wgrad_k<TY, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dw_wgrad_k<TY, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  int grid = 2 * num_sms();
  if (grid > a.total_tiles) grid = a.total_tiles;
  if (a.CB == 64) ck::launch(dw_wgrad_k<TY, 1>, grid, 256, smem, st, tmA, tmG, a);
  else ck::launch(dw_wgrad_k<TY, 2>, grid, 256, smem, st, tmA, tmG, a);
  CK_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
