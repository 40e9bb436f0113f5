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
  unsigned long long m_sp, m_tx, m_ty;  // ceil(2^40 / d) for d = tiles per channel block, tiles_x, tiles_y (tile -> coordinates without divide sequences)
  int nstage;                   // depth of the TMA ring (2..4)
  const float* rbn;             // dw_bwd_k: BN ([scale][shift][mean][invstd]) whose backward reductions ride along
  float* rsums;                 //           [sum dz C][sum dz*(y-mean) C], pre-zeroed (NULL: off)
};

__host__ __device__ __forceinline__ unsigned long long dw_magic(int d) { return ((1ull << 40) + (unsigned long long)d - 1) / (unsigned long long)d; }
__device__ __forceinline__ int dw_fdiv(int x, unsigned long long m) { return (int)(((unsigned long long)(unsigned)x * m) >> 40); }

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

// Second generation of the stencil kernel.  ncu on the first one (profiles/ncu_dw_stencil.md): 2344 warp instructions per
// 16x16x64 tile, issue-bound at 52 % with `no_instruction` the top stall (an 18-row fully unrolled body with the
// activation inlined three times per element outgrew the 32 KB instruction cache).  Here
//   * the activation (relu / relu(bn)) is applied ONCE per element, in place in the shared-memory halo tile, before the
//     stencil (cvt.rn.relu.bf16x2 packs and clamps in one instruction);
//   * the stencil accumulates channel pairs with FFMA2 (18 instead of 36 FMAs per output quad);
//   * the input-gradient variant gets the forward input (relu mask) as a second TMA tile instead of a dependent global
//     load per output row, and prefetches the residual-path gradient before it waits for the tile.
// TY = tile rows; threads: (c4 = CB/4) x 16 columns x RG row groups = 256 (RG = 1 for CB = 64, 2 for CB = 32)
template <int TY, int RG, bool DGRAD>
__global__ void __launch_bounds__(256, 2) dw_stencil_k(const __grid_constant__ CUtensorMap tmX,
                                                       const __grid_constant__ CUtensorMap tmS, const DwArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const int CB = a.CB, rowb = CB * 2;
  const uint32_t halo_bytes = (uint32_t)((TY + 2) * HW * rowb);
  const uint32_t stage_bytes = halo_bytes + (DGRAD ? (uint32_t)(TY * TX * rowb) : 0u);
  // ring of NS stages: measured, with the original 2 stages (one tile in flight while one is processed) these kernels
  // sat at 25-30 % of the HBM rate however few instructions they executed - latency-bound on the bulk copies
  const int NS = a.nstage;
  const uint32_t bar0 = base + (uint32_t)NS * stage_bytes;
  float* coef = reinterpret_cast<float*>(gbase + (uint32_t)NS * stage_bytes + 64);  // [scale CB][shift CB] of the current channel block
  const int nc4 = CB >> 2;
  constexpr int rows = TY / RG;
  const int c4 = threadIdx.x % nc4, col = (threadIdx.x / nc4) % TX, rg = threadIdx.x / (nc4 * TX);
  const int spatial = a.tiles_x * a.tiles_y * a.B;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NS; ++i) mbar_init(bar0 + 8u * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&tmX);
    if (DGRAD) tma_prefetch_desc(&tmS);
  }
  pdl_enter();  // barrier init / descriptor prefetch above overlap the previous kernel's tail
  fin_publish(a.fin, a.C);
  __syncthreads();

  auto issue = [&](int tile, int stage) {
    const int cblk = dw_fdiv(tile, a.m_sp);
    const int sp = tile - cblk * spatial;
    const int q_ = dw_fdiv(sp, a.m_tx);
    const int tx_i = sp - q_ * a.tiles_x;
    const int b = dw_fdiv(q_, a.m_ty);
    const int ty_i = q_ - b * a.tiles_y;
    mbar_expect_tx(bar0 + 8u * stage, stage_bytes);
    tma_load_4d(base + stage * stage_bytes, &tmX, bar0 + 8u * stage, cblk * CB, tx_i * TX - 1, ty_i * TY - 1, b);
    if (DGRAD) tma_load_4d(base + stage * stage_bytes + halo_bytes, &tmS, bar0 + 8u * stage, cblk * CB, tx_i * TX, ty_i * TY, b);
  };
  if (threadIdx.x == 0)
    for (int i = 0; i < NS - 1; ++i)
      if ((int)(blockIdx.x + i * gridDim.x) < a.total_tiles) issue(blockIdx.x + i * gridDim.x, i);

  int cur_cblk = -1;
  f2_t wk[9][2], msc[2], msh[2];
  msc[0] = msc[1] = msh[0] = msh[1] = 0ull;
  int it = 0, stage = 0;
  uint32_t par = 0;
  for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
    {
      // refill the stage consumed in the previous iteration (its last readers passed the barrier at the loop tail)
      const int nxt = tile + (NS - 1) * gridDim.x;
      const int ns = stage == 0 ? NS - 1 : stage - 1;
      if (threadIdx.x == 0 && nxt < a.total_tiles) issue(nxt, ns);
    }
    const int cblk = dw_fdiv(tile, a.m_sp);
    const int sp = tile - cblk * spatial;
    const int q_ = dw_fdiv(sp, a.m_tx);
    const int tx_i = sp - q_ * a.tiles_x;
    const int b = dw_fdiv(q_, a.m_ty);
    const int ty_i = q_ - b * a.tiles_y;
    const int cg = cblk * CB + c4 * 4;  // first global channel of this thread
    if (cblk != cur_cblk) {             // uniform over the CTA
      cur_cblk = cblk;
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(a.w + (a.flip ? 8 - t : t) * a.C + cg));
        wk[t][0] = pk2(v.x, v.y); wk[t][1] = pk2(v.z, v.w);
      }
      if (!DGRAD && a.act_mode == 1) {
        if (threadIdx.x < nc4) {
          float sc[4], sh[4];
          coeffs4(a.fin, a.bn, a.C, cblk * CB + threadIdx.x * 4, sc, sh);
#pragma unroll
          for (int j = 0; j < 4; ++j) { coef[threadIdx.x * 4 + j] = sc[j]; coef[CB + threadIdx.x * 4 + j] = sh[j]; }
        }
        __syncthreads();
      }
      if (DGRAD && a.src_mode == 1) {
        const float4 s4 = __ldg(reinterpret_cast<const float4*>(a.src_bn + cg));
        const float4 h4 = __ldg(reinterpret_cast<const float4*>(a.src_bn + a.C + cg));
        msc[0] = pk2(s4.x, s4.y); msc[1] = pk2(s4.z, s4.w);
        msh[0] = pk2(h4.x, h4.y); msh[1] = pk2(h4.z, h4.w);
      }
    }
    const int w0 = tx_i * TX, ty0 = ty_i * TY, h0 = ty0 + rg * rows;
    const int gw = w0 + col;
    // residual-path gradient (half resolution, even pixels only): in flight while we wait for the tile
    uint2 rres[(rows + 1) / 2];
    if (DGRAD) {
#pragma unroll
      for (int i = 0; i < (rows + 1) / 2; ++i) {
        rres[i] = make_uint2(0u, 0u);
        const int oh = h0 + 2 * i;  // h0 is even
        if (a.resid && !(col & 1) && oh < a.H && gw < a.W)
          rres[i] = __ldg(reinterpret_cast<const uint2*>(
              a.resid_t + (((long)b * (a.H >> 1) + (oh >> 1)) * (a.W >> 1) + (gw >> 1)) * a.C + cg));
      }
    }
    mbar_wait(bar0 + 8u * stage, par);
    uint8_t* tile_s = gbase + stage * stage_bytes;

    if (!DGRAD && a.act_mode != 2) {
      // activation pre-pass, in place; out-of-image pixels (TMA zero fill) must stay zero under relu(bn(.))
      const int c8n = CB >> 3;
      const int nchunk = (TY + 2) * HW * c8n;
      const bool interior = w0 >= 1 && w0 + TX < a.W && ty0 >= 1 && ty0 + TY < a.H;
      if (a.act_mode == 0) {
        for (int idx = threadIdx.x; idx < nchunk; idx += 256) {
          uint4* q = reinterpret_cast<uint4*>(tile_s) + idx;
          uint4 v = *q;
          v.x = relu_bf2(v.x); v.y = relu_bf2(v.y); v.z = relu_bf2(v.z); v.w = relu_bf2(v.w);
          *q = v;
        }
      } else {
        const int c8 = threadIdx.x % c8n;  // 256 % c8n == 0: the same channel octet on every trip
        f2_t scp[4], shp[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          scp[j] = *reinterpret_cast<const f2_t*>(coef + c8 * 8 + 2 * j);
          shp[j] = *reinterpret_cast<const f2_t*>(coef + CB + c8 * 8 + 2 * j);
        }
        for (int idx = threadIdx.x; idx < nchunk; idx += 256) {
          uint4* q = reinterpret_cast<uint4*>(tile_s) + idx;
          uint4 v = *q;
          v.x = f2_to_bf2_relu(ffma2(bf2_to_f2(v.x), scp[0], shp[0]));
          v.y = f2_to_bf2_relu(ffma2(bf2_to_f2(v.y), scp[1], shp[1]));
          v.z = f2_to_bf2_relu(ffma2(bf2_to_f2(v.z), scp[2], shp[2]));
          v.w = f2_to_bf2_relu(ffma2(bf2_to_f2(v.w), scp[3], shp[3]));
          if (!interior) {
            const int pi = idx >> (c8n == 8 ? 3 : 2);  // c8n = CB / 8 is 8 or 4
            const int hr = pi / HW, hc = pi - hr * HW;
            const int gh = ty0 - 1 + hr, gx = w0 - 1 + hc;
            if (gh < 0 || gh >= a.H || gx < 0 || gx >= a.W) v = make_uint4(0u, 0u, 0u, 0u);
          }
          *q = v;
        }
      }
      __syncthreads();
    }

    f2_t win[3][3][2];
    const long out_rs = (long)a.W * a.C;  // elements per image row
    __nv_bfloat16* const out_p = a.out + (((long)b * a.H + h0) * a.W + gw) * a.C + cg;
    const uint8_t* const row0 = tile_s + (size_t)(((rg * rows) * HW + col) * rowb + c4 * 8);
    const uint8_t* const src0 = tile_s + halo_bytes + (size_t)(((rg * rows) * TX + col) * rowb + c4 * 8);
#pragma unroll
    for (int r = 0; r < rows + 2; ++r) {
      const uint8_t* rowp = row0 + r * (HW * rowb);
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const uint2 v = *reinterpret_cast<const uint2*>(rowp + kx * rowb);
        win[r % 3][kx][0] = bf2_to_f2(v.x);
        win[r % 3][kx][1] = bf2_to_f2(v.y);
      }
      if (r >= 2) {
        const int oh = h0 + r - 2;
        f2_t acc0 = fmul2(win[(r - 2) % 3][0][0], wk[0][0]), acc1 = fmul2(win[(r - 2) % 3][0][1], wk[0][1]);
#pragma unroll
        for (int t = 1; t < 9; ++t) {
          acc0 = ffma2(win[(r - 2 + t / 3) % 3][t % 3][0], wk[t][0], acc0);
          acc1 = ffma2(win[(r - 2 + t / 3) % 3][t % 3][1], wk[t][1], acc1);
        }
        float o4[4];
        upk2(acc0, o4[0], o4[1]);
        upk2(acc1, o4[2], o4[3]);
        if (DGRAD) {
          // input-gradient epilogue: relu mask of the forward input, residual-path gradient on even pixels
          const uint2 sv = *reinterpret_cast<const uint2*>(src0 + (r - 2) * (TX * rowb));
          f2_t p0 = bf2_to_f2(sv.x), p1 = bf2_to_f2(sv.y);
          if (a.src_mode == 1) { p0 = ffma2(p0, msc[0], msh[0]); p1 = ffma2(p1, msc[1], msh[1]); }
          float pre[4];
          upk2(p0, pre[0], pre[1]);
          upk2(p1, pre[2], pre[3]);
          float r4[4] = {0.f, 0.f, 0.f, 0.f};
          if (((r - 2) & 1) == 0) {
            const uint2 rv = rres[(r - 2) / 2];
            r4[0] = __uint_as_float(rv.x << 16); r4[1] = __uint_as_float(rv.x & 0xffff0000u);
            r4[2] = __uint_as_float(rv.y << 16); r4[3] = __uint_as_float(rv.y & 0xffff0000u);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const bool m = pre[j] > 0.f;
            o4[j] = (a.resid == 2) ? (m ? o4[j] + r4[j] : 0.f) : ((m ? o4[j] : 0.f) + r4[j]);
          }
        }
        if (oh < a.H && gw < a.W) {
          uint2 pk;
          pk.x = pack2(o4[0], o4[1]);
          pk.y = pack2(o4[2], o4[3]);
          *reinterpret_cast<uint2*>(out_p + (r - 2) * out_rs) = pk;
        }
      }
    }
    // generic-proxy writes (pre-pass) are ordered before the TMA refill of this stage two iterations from now
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();  // every thread is done with this stage before it is refilled
    if (++stage == NS) { stage = 0; par ^= 1u; }
  }
}

// weight gradient: a-halo tile + g tile per stage; same structure as the stencil kernel (in-place activation pre-pass,
// FFMA2 channel pairs): dW[t][c] accumulates in 18 packed registers per thread over all tiles of a channel block.
template <int TY, int RG>
__global__ void __launch_bounds__(256, 2) dw_wgrad_k(const __grid_constant__ CUtensorMap tmA,
                                                     const __grid_constant__ CUtensorMap tmG, const DwArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const int CB = a.CB, rowb = CB * 2;
  const uint32_t a_bytes = (uint32_t)((TY + 2) * HW * rowb), g_bytes = (uint32_t)(TY * TX * rowb);
  const uint32_t stage_bytes = a_bytes + g_bytes;
  const int NS = a.nstage;
  const uint32_t bar0 = base + (uint32_t)NS * stage_bytes;
  float* red = reinterpret_cast<float*>(gbase + (uint32_t)NS * stage_bytes + 64);  // [256][4] floats
  float* coef = red + 256 * 4;                                            // [scale CB][shift CB]
  const int nc4 = CB >> 2;
  constexpr int rows = TY / RG;
  const int c4 = threadIdx.x % nc4, col = (threadIdx.x / nc4) % TX, rg = threadIdx.x / (nc4 * TX);
  const int spatial = a.tiles_x * a.tiles_y * a.B;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NS; ++i) mbar_init(bar0 + 8u * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmG);
  }
  __syncthreads();
  pdl_enter();

  auto issue = [&](int tile, int stage) {
    const int cblk = dw_fdiv(tile, a.m_sp);
    const int sp = tile - cblk * spatial;
    const int q_ = dw_fdiv(sp, a.m_tx);
    const int tx_i = sp - q_ * a.tiles_x;
    const int b = dw_fdiv(q_, a.m_ty);
    const int ty_i = q_ - b * a.tiles_y;
    mbar_expect_tx(bar0 + 8u * stage, stage_bytes);
    tma_load_4d(base + stage * stage_bytes, &tmA, bar0 + 8u * stage, cblk * CB, tx_i * TX - 1, ty_i * TY - 1, b);
    tma_load_4d(base + stage * stage_bytes + a_bytes, &tmG, bar0 + 8u * stage, cblk * CB, tx_i * TX, ty_i * TY, b);
  };
  if (threadIdx.x == 0)
    for (int i = 0; i < NS - 1; ++i)
      if ((int)(blockIdx.x + i * gridDim.x) < a.total_tiles) issue(blockIdx.x + i * gridDim.x, i);

  f2_t acc[9][2];
#pragma unroll
  for (int t = 0; t < 9; ++t) { acc[t][0] = 0ull; acc[t][1] = 0ull; }
  int cur_cblk = -1;

  auto flush = [&](int cblk) {
    // reduce the 16 columns x RG row groups that share (c4) and add into dwk[t][cblk*CB + c4*4 + j]
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      __syncthreads();
      float f0, f1, f2, f3;
      upk2(acc[t][0], f0, f1);
      upk2(acc[t][1], f2, f3);
      *reinterpret_cast<float4*>(red + threadIdx.x * 4) = make_float4(f0, f1, f2, f3);
      __syncthreads();
      if (threadIdx.x < nc4 * 4) {
        const int cc = threadIdx.x >> 2, j = threadIdx.x & 3;
        float v = 0.f;
        for (int q = 0; q < 256 / nc4; ++q) v += red[(q * nc4 + cc) * 4 + j];
        atomicAdd(a.dwk + (long)t * a.C + cblk * CB + cc * 4 + j, v);
      }
      acc[t][0] = 0ull; acc[t][1] = 0ull;
    }
  };

  int it = 0, stage = 0;
  uint32_t par = 0;
  for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
    {
      const int nxt = tile + (NS - 1) * gridDim.x;
      const int ns = stage == 0 ? NS - 1 : stage - 1;
      if (threadIdx.x == 0 && nxt < a.total_tiles) issue(nxt, ns);
    }
    const int cblk = dw_fdiv(tile, a.m_sp);
    const int sp = tile - cblk * spatial;
    const int q_ = dw_fdiv(sp, a.m_tx);
    const int tx_i = sp - q_ * a.tiles_x;
    const int ty_i = q_ - dw_fdiv(q_, a.m_ty) * a.tiles_y;
    if (cblk != cur_cblk) {  // uniform over the CTA
      if (cur_cblk >= 0) flush(cur_cblk);
      cur_cblk = cblk;
      if (a.act_mode == 1) {
        __syncthreads();  // the previous tile's pre-pass readers of coef are done
        if (threadIdx.x < nc4) {
          float sc[4], sh[4];
          coeffs4(a.fin, a.bn, a.C, cblk * CB + threadIdx.x * 4, sc, sh);
#pragma unroll
          for (int j = 0; j < 4; ++j) { coef[threadIdx.x * 4 + j] = sc[j]; coef[CB + threadIdx.x * 4 + j] = sh[j]; }
        }
        __syncthreads();
      }
    }
    mbar_wait(bar0 + 8u * stage, par);
    uint8_t* tile_a = gbase + stage * stage_bytes;
    const uint8_t* tile_g = tile_a + a_bytes;
    const int w0 = tx_i * TX, ty0 = ty_i * TY;

    {  // activation pre-pass, in place (relu / relu(bn)); out-of-image pixels stay zero
      const int c8n = CB >> 3;
      const int nchunk = (TY + 2) * HW * c8n;
      if (a.act_mode == 0) {
        for (int idx = threadIdx.x; idx < nchunk; idx += 256) {
          uint4* q = reinterpret_cast<uint4*>(tile_a) + idx;
          uint4 v = *q;
          v.x = relu_bf2(v.x); v.y = relu_bf2(v.y); v.z = relu_bf2(v.z); v.w = relu_bf2(v.w);
          *q = v;
        }
      } else {
        const bool interior = w0 >= 1 && w0 + TX < a.W && ty0 >= 1 && ty0 + TY < a.H;
        const int c8 = threadIdx.x % c8n;
        f2_t scp[4], shp[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          scp[j] = *reinterpret_cast<const f2_t*>(coef + c8 * 8 + 2 * j);
          shp[j] = *reinterpret_cast<const f2_t*>(coef + CB + c8 * 8 + 2 * j);
        }
        for (int idx = threadIdx.x; idx < nchunk; idx += 256) {
          uint4* q = reinterpret_cast<uint4*>(tile_a) + idx;
          uint4 v = *q;
          v.x = f2_to_bf2_relu(ffma2(bf2_to_f2(v.x), scp[0], shp[0]));
          v.y = f2_to_bf2_relu(ffma2(bf2_to_f2(v.y), scp[1], shp[1]));
          v.z = f2_to_bf2_relu(ffma2(bf2_to_f2(v.z), scp[2], shp[2]));
          v.w = f2_to_bf2_relu(ffma2(bf2_to_f2(v.w), scp[3], shp[3]));
          if (!interior) {
            const int pi = idx >> (c8n == 8 ? 3 : 2);  // c8n = CB / 8 is 8 or 4
            const int hr = pi / HW, hc = pi - hr * HW;
            const int gh = ty0 - 1 + hr, gx = w0 - 1 + hc;
            if (gh < 0 || gh >= a.H || gx < 0 || gx >= a.W) v = make_uint4(0u, 0u, 0u, 0u);
          }
          *q = v;
        }
      }
      __syncthreads();
    }

    f2_t win[3][3][2];
    const uint8_t* const row0 = tile_a + (size_t)(((rg * rows) * HW + col) * rowb + c4 * 8);
    const uint8_t* const g0 = tile_g + (size_t)(((rg * rows) * TX + col) * rowb + c4 * 8);
#pragma unroll
    for (int r = 0; r < rows + 2; ++r) {
      const uint8_t* rowp = row0 + r * (HW * rowb);
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const uint2 v = *reinterpret_cast<const uint2*>(rowp + kx * rowb);
        win[r % 3][kx][0] = bf2_to_f2(v.x);
        win[r % 3][kx][1] = bf2_to_f2(v.y);
      }
      if (r >= 2) {
        // out-of-image g pixels are TMA zero fill: they contribute nothing
        const uint2 gv = *reinterpret_cast<const uint2*>(g0 + (r - 2) * (TX * rowb));
        const f2_t ga = bf2_to_f2(gv.x), gb = bf2_to_f2(gv.y);
#pragma unroll
        for (int t = 0; t < 9; ++t) {
          acc[t][0] = ffma2(win[(r - 2 + t / 3) % 3][t % 3][0], ga, acc[t][0]);
          acc[t][1] = ffma2(win[(r - 2 + t / 3) % 3][t % 3][1], gb, acc[t][1]);
        }
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (++stage == NS) { stage = 0; par ^= 1u; }
  }
  if (cur_cblk >= 0) flush(cur_cblk);
}

// Third generation of the forward stencil.  ncu on dw_stencil_k (64 channels, 128x128): the in-place activation pre-pass
// is 28 % of the instructions and 41 % of the stall samples (its barrier, dependent LDS -> STS round trips, the
// coefficient hand-off through shared memory).  Here the activation is applied to the three words a thread loads per
// window row; the halo comes through a tensor map that delivers out-of-image pixels as NaN, so relu(scale * x + shift)
// is exactly 0 there (max(NaN, 0) = 0) with no border masks; the channel block is a compile-time constant; every thread
// computes the BN coefficients of its own four channels; a CTA owns a contiguous range of tiles.
template <int TY, int RG, int AMODE>
__global__ void __launch_bounds__(256, 2) dw_fwd2_k(const __grid_constant__ CUtensorMap tmX, const DwArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  constexpr int CB = 64 / RG, rowb = CB * 2, nc4 = CB >> 2;
  constexpr uint32_t stage_bytes = (uint32_t)((TY + 2) * HW * rowb);
  const int NS = a.nstage;
  const uint32_t bar0 = base + (uint32_t)NS * stage_bytes;
  constexpr int rows = TY / RG;
  const int c4 = threadIdx.x % nc4, col = (threadIdx.x / nc4) % TX, rg = threadIdx.x / (nc4 * TX);
  const int spatial = a.tiles_x * a.tiles_y * a.B;
  const int t_lo = (int)(((long)blockIdx.x * a.total_tiles) / gridDim.x);
  const int t_hi = (int)(((long)(blockIdx.x + 1) * a.total_tiles) / gridDim.x);

  if (threadIdx.x == 0) {
    for (int i = 0; i < NS; ++i) mbar_init(bar0 + 8u * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&tmX);
  }
  // the stencil weights are parameters (last written by the previous step's Adam, long complete): the first channel
  // block's nine float4 loads go out BEFORE the wait on the previous kernel (ncu: ~8 % of the samples of the whole kernel
  // sat on that first dependent load)
  f2_t wk[9][2], sc[2] = {0ull, 0ull}, sh[2] = {0ull, 0ull};
  int w_cblk = -1;
  if (t_lo < t_hi) {
    w_cblk = dw_fdiv(t_lo, a.m_sp);
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(a.w + (a.flip ? 8 - t : t) * a.C + w_cblk * CB + c4 * 4));
      wk[t][0] = pk2(v.x, v.y); wk[t][1] = pk2(v.z, v.w);
    }
  }
  pdl_enter();
  fin_publish(a.fin, a.C);
  __syncthreads();

  auto issue = [&](int tile, int stage) {
    const int cblk = dw_fdiv(tile, a.m_sp);
    const int sp = tile - cblk * spatial;
    const int q_ = dw_fdiv(sp, a.m_tx);
    const int tx_i = sp - q_ * a.tiles_x;
    const int b = dw_fdiv(q_, a.m_ty);
    const int ty_i = q_ - b * a.tiles_y;
    mbar_expect_tx(bar0 + 8u * stage, stage_bytes);
    tma_load_4d(base + stage * stage_bytes, &tmX, bar0 + 8u * stage, cblk * CB, tx_i * TX - 1, ty_i * TY - 1, b);
  };
  if (threadIdx.x == 0)
    for (int i = 0; i < NS - 1; ++i)
      if (t_lo + i < t_hi) issue(t_lo + i, i);

  int cur_cblk = -1;
  int stage = 0;
  uint32_t par = 0;
  for (int tile = t_lo; tile < t_hi; ++tile) {
    {
      // refill the stage consumed in the previous iteration (its last readers passed the barrier at the loop tail)
      const int nxt = tile + (NS - 1);
      const int ns = stage == 0 ? NS - 1 : stage - 1;
      if (threadIdx.x == 0 && nxt < t_hi) issue(nxt, ns);
    }
    const int cblk = dw_fdiv(tile, a.m_sp);
    const int sp = tile - cblk * spatial;
    const int q_ = dw_fdiv(sp, a.m_tx);
    const int tx_i = sp - q_ * a.tiles_x;
    const int b = dw_fdiv(q_, a.m_ty);
    const int ty_i = q_ - b * a.tiles_y;
    const int cg = cblk * CB + c4 * 4;  // first global channel of this thread
    if (cblk != cur_cblk) {             // uniform over the CTA
      cur_cblk = cblk;
      if (cblk != w_cblk) {
        w_cblk = cblk;
#pragma unroll
        for (int t = 0; t < 9; ++t) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(a.w + (a.flip ? 8 - t : t) * a.C + cg));
          wk[t][0] = pk2(v.x, v.y); wk[t][1] = pk2(v.z, v.w);
        }
      }
      if (AMODE == 1) {
        float s_[4], h_[4];
        coeffs4(a.fin, a.bn, a.C, cg, s_, h_);
        sc[0] = pk2(s_[0], s_[1]); sc[1] = pk2(s_[2], s_[3]);
        sh[0] = pk2(h_[0], h_[1]); sh[1] = pk2(h_[2], h_[3]);
      }
    }
    const int w0 = tx_i * TX, ty0 = ty_i * TY, h0 = ty0 + rg * rows;
    const int gw = w0 + col;
    mbar_wait(bar0 + 8u * stage, par);
    const uint8_t* tile_s = gbase + stage * stage_bytes;

    f2_t win[3][3][2];
    const long out_rs = (long)a.W * a.C;  // elements per image row
    __nv_bfloat16* const out_p = a.out + (((long)b * a.H + h0) * a.W + gw) * a.C + cg;
    const uint8_t* const row0 = tile_s + (uint32_t)(((rg * rows) * HW + col) * rowb + c4 * 8);
#pragma unroll
    for (int r = 0; r < rows + 2; ++r) {
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        uint2 v = *reinterpret_cast<const uint2*>(row0 + (r * HW + kx) * rowb);
        if (AMODE == 0) {
          win[r % 3][kx][0] = bf2_to_f2(relu_bf2(v.x));
          win[r % 3][kx][1] = bf2_to_f2(relu_bf2(v.y));
        } else if (AMODE == 1) {
          float l0, h0_, l1, h1;
          upk2(ffma2(bf2_to_f2(v.x), sc[0], sh[0]), l0, h0_);
          upk2(ffma2(bf2_to_f2(v.y), sc[1], sh[1]), l1, h1);
          win[r % 3][kx][0] = pk2(fmaxf(l0, 0.f), fmaxf(h0_, 0.f));
          win[r % 3][kx][1] = pk2(fmaxf(l1, 0.f), fmaxf(h1, 0.f));
        } else {
          win[r % 3][kx][0] = bf2_to_f2(v.x);
          win[r % 3][kx][1] = bf2_to_f2(v.y);
        }
      }
      if (r >= 2) {
        const int oh = h0 + r - 2;
        f2_t p0[3], p1[3];  // one partial sum per window row: three independent chains per channel pair
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
          p0[ky] = fmul2(win[(r - 2 + ky) % 3][0][0], wk[ky * 3][0]);
          p1[ky] = fmul2(win[(r - 2 + ky) % 3][0][1], wk[ky * 3][1]);
#pragma unroll
          for (int kx = 1; kx < 3; ++kx) {
            p0[ky] = ffma2(win[(r - 2 + ky) % 3][kx][0], wk[ky * 3 + kx][0], p0[ky]);
            p1[ky] = ffma2(win[(r - 2 + ky) % 3][kx][1], wk[ky * 3 + kx][1], p1[ky]);
          }
        }
        float o4[4];
        upk2(fadd2(fadd2(p0[0], p0[1]), p0[2]), o4[0], o4[1]);
        upk2(fadd2(fadd2(p1[0], p1[1]), p1[2]), o4[2], o4[3]);
        if (oh < a.H && gw < a.W) {
          uint2 pk;
          pk.x = pack2(o4[0], o4[1]);
          pk.y = pack2(o4[2], o4[3]);
          *reinterpret_cast<uint2*>(out_p + (r - 2) * out_rs) = pk;
        }
      }
    }
    __syncthreads();  // every thread is done reading this stage before it is refilled
    if (++stage == NS) { stage = 0; par ^= 1u; }
  }
}

// Depthwise backward in ONE pass (third generation): the gradient tile g (with halo) and the forward-input tile src (with
// halo) are staged once; after an in-place activation pre-pass on src (a = relu(src) or relu(bn(src)))
//   warps 0-7   compute the input gradient  dz = (3x3^T g) * [a > 0] (+ residual-path gradient) and,
//   warps 8-15  accumulate the weight gradient dW[t][c] += a[p + off(t)][c] * g[p][c] in registers,
// so g and src are read once instead of twice (what-if: the separate weight-gradient kernels cost 134 us of a 1.45 ms
// step even on their own stream) and one launch per layer goes away.  The relu mask is taken from the activated tile
// (relu(z) > 0 <=> z > 0).  512 threads x <= 128 registers: one CTA per SM, two stages of 2 x (TY+2) x 18 x CB.
template <int TY, int RG>
__global__ void __launch_bounds__(512, 1) dw_bwd_k(const __grid_constant__ CUtensorMap tmG,
                                                   const __grid_constant__ CUtensorMap tmS, const DwArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const int CB = a.CB, rowb = CB * 2;
  const uint32_t halo_bytes = (uint32_t)((TY + 2) * HW * rowb);
  const uint32_t stage_bytes = 2u * halo_bytes;  // [g halo][src halo]
  const uint32_t bar0 = base + 2u * stage_bytes;
  float* red = reinterpret_cast<float*>(gbase + 2u * stage_bytes + 64);  // [256][4] floats
  float* coef = red + 256 * 4;                                            // [scale CB][shift CB]
  float* red0 = coef + 2 * 64;                                            // [256][4] floats, role 0 (BN reductions)
  const int nc4 = CB >> 2;
  constexpr int rows = TY / RG;
  const int role = threadIdx.x >> 8, tl = threadIdx.x & 255;
  const int c4 = tl % nc4, col = (tl / nc4) % TX, rg = tl / (nc4 * TX);
  const int spatial = a.tiles_x * a.tiles_y * a.B;

  if (threadIdx.x == 0) {
    mbar_init(bar0, 1);
    mbar_init(bar0 + 8u, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&tmG);
    tma_prefetch_desc(&tmS);
  }
  pdl_enter();
  __syncthreads();

  auto issue = [&](int tile, int stage) {
    const int cblk = dw_fdiv(tile, a.m_sp);
    const int sp = tile - cblk * spatial;
    const int q_ = dw_fdiv(sp, a.m_tx);
    const int tx_i = sp - q_ * a.tiles_x;
    const int b = dw_fdiv(q_, a.m_ty);
    const int ty_i = q_ - b * a.tiles_y;
    mbar_expect_tx(bar0 + 8u * stage, stage_bytes);
    tma_load_4d(base + stage * stage_bytes, &tmG, bar0 + 8u * stage, cblk * CB, tx_i * TX - 1, ty_i * TY - 1, b);
    tma_load_4d(base + stage * stage_bytes + halo_bytes, &tmS, bar0 + 8u * stage, cblk * CB, tx_i * TX - 1, ty_i * TY - 1, b);
  };
  if (threadIdx.x == 0 && (int)blockIdx.x < a.total_tiles) issue(blockIdx.x, 0);

  // role 0: flipped taps; role 1: weight-gradient accumulators (same registers, disjoint warps)
  f2_t wacc[9][2];
#pragma unroll
  for (int t = 0; t < 9; ++t) { wacc[t][0] = 0ull; wacc[t][1] = 0ull; }
  int cur_cblk = -1;

  auto flush = [&](int cblk) {  // role 1 only (named barrier 1, 256 threads)
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      asm volatile("bar.sync 1, 256;" ::: "memory");
      float f0, f1, f2, f3;
      upk2(wacc[t][0], f0, f1);
      upk2(wacc[t][1], f2, f3);
      *reinterpret_cast<float4*>(red + tl * 4) = make_float4(f0, f1, f2, f3);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (tl < nc4 * 4) {
        const int cc = tl >> 2, j = tl & 3;
        float v = 0.f;
        for (int q = 0; q < 256 / nc4; ++q) v += red[(q * nc4 + cc) * 4 + j];
        atomicAdd(a.dwk + (long)t * a.C + cblk * CB + cc * 4 + j, v);
      }
      wacc[t][0] = 0ull; wacc[t][1] = 0ull;
    }
  };

  // BN-backward reductions of the layer below (role 0): sum dz and sum dz * a, a = the activated tile.  Where the
  // mask is on, a = scale * y + shift, so  sum dz * (y - mean) = (sum dz * a - beta * sum dz) / scale,  beta = shift + mean * scale.
  f2_t bs1[2] = {0ull, 0ull}, bs2[2] = {0ull, 0ull};
  auto flush_bn = [&](int cblk) {  // role 0 only (named barrier 2, 256 threads)
#pragma unroll
    for (int which = 0; which < 2; ++which) {
      asm volatile("bar.sync 2, 256;" ::: "memory");
      float f0, f1, f2, f3;
      upk2(which ? bs2[0] : bs1[0], f0, f1);
      upk2(which ? bs2[1] : bs1[1], f2, f3);
      *reinterpret_cast<float4*>(red0 + tl * 4) = make_float4(f0, f1, f2, f3);
      asm volatile("bar.sync 2, 256;" ::: "memory");
      if (tl < nc4 * 4) {
        const int cc = tl >> 2, j = tl & 3;
        float v = 0.f;
        for (int q = 0; q < 256 / nc4; ++q) v += red0[(q * nc4 + cc) * 4 + j];
        red0[1024 + which * 64 + tl] = v;  // [which][CB] staging behind the partials
      }
    }
    asm volatile("bar.sync 2, 256;" ::: "memory");
    if (tl < nc4 * 4) {
      const int c = cblk * CB + tl;
      const float S1 = red0[1024 + tl], S2 = red0[1024 + 64 + tl];
      const float sc = a.rbn[c], sh = a.rbn[a.C + c], mu = a.rbn[2 * a.C + c];
      const float beta = sh + mu * sc;
      atomicAdd(a.rsums + c, S1);
      atomicAdd(a.rsums + a.C + c, sc != 0.f ? (S2 - beta * S1) / sc : 0.f);
    }
    bs1[0] = bs1[1] = bs2[0] = bs2[1] = 0ull;
  };

  int it = 0;
  for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
    const int stage = it & 1;
    const int nxt = tile + gridDim.x;
    if (threadIdx.x == 0 && nxt < a.total_tiles) issue(nxt, stage ^ 1);  // its last readers passed the barrier at the loop tail
    const int cblk = dw_fdiv(tile, a.m_sp);
    const int sp = tile - cblk * spatial;
    const int q_ = dw_fdiv(sp, a.m_tx);
    const int tx_i = sp - q_ * a.tiles_x;
    const int b = dw_fdiv(q_, a.m_ty);
    const int ty_i = q_ - b * a.tiles_y;
    const int cg = cblk * CB + c4 * 4;
    if (cblk != cur_cblk) {  // uniform over the CTA
      if (role == 1 && cur_cblk >= 0) flush(cur_cblk);
      if (role == 0 && cur_cblk >= 0 && a.rsums) flush_bn(cur_cblk);
      cur_cblk = cblk;
      if (role == 0) {
#pragma unroll
        for (int t = 0; t < 9; ++t) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(a.w + (8 - t) * a.C + cg));
          wacc[t][0] = pk2(v.x, v.y); wacc[t][1] = pk2(v.z, v.w);
        }
      }
      if (a.src_mode == 1) {
        __syncthreads();  // the previous tile's pre-pass readers of coef are done
        if (threadIdx.x < nc4) {
          const float4 s4 = __ldg(reinterpret_cast<const float4*>(a.src_bn + cblk * CB + threadIdx.x * 4));
          const float4 h4 = __ldg(reinterpret_cast<const float4*>(a.src_bn + a.C + cblk * CB + threadIdx.x * 4));
          *reinterpret_cast<float4*>(coef + threadIdx.x * 4) = s4;
          *reinterpret_cast<float4*>(coef + CB + threadIdx.x * 4) = h4;
        }
        __syncthreads();
      }
    }
    const int w0 = tx_i * TX, ty0 = ty_i * TY, h0 = ty0 + rg * rows;
    const int gw = w0 + col;
    uint2 rres[(rows + 1) / 2];
    if (role == 0) {
#pragma unroll
      for (int i = 0; i < (rows + 1) / 2; ++i) {
        rres[i] = make_uint2(0u, 0u);
        const int oh = h0 + 2 * i;  // h0 is even
        if (a.resid && !(col & 1) && oh < a.H && gw < a.W)
          rres[i] = __ldg(reinterpret_cast<const uint2*>(
              a.resid_t + (((long)b * (a.H >> 1) + (oh >> 1)) * (a.W >> 1) + (gw >> 1)) * a.C + cg));
      }
    }
    mbar_wait(bar0 + 8u * stage, (uint32_t)(it >> 1) & 1u);
    uint8_t* tile_g = gbase + stage * stage_bytes;
    uint8_t* tile_a = tile_g + halo_bytes;

    {  // activation pre-pass on the src halo, in place, all 512 threads; out-of-image pixels stay zero
      const int c8n = CB >> 3;
      const int nchunk = (TY + 2) * HW * c8n;
      if (a.src_mode == 0) {
        for (int idx = threadIdx.x; idx < nchunk; idx += 512) {
          uint4* q = reinterpret_cast<uint4*>(tile_a) + idx;
          uint4 v = *q;
          v.x = relu_bf2(v.x); v.y = relu_bf2(v.y); v.z = relu_bf2(v.z); v.w = relu_bf2(v.w);
          *q = v;
        }
      } else {
        const bool interior = w0 >= 1 && w0 + TX < a.W && ty0 >= 1 && ty0 + TY < a.H;
        const int c8 = threadIdx.x % c8n;
        f2_t scp[4], shp[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          scp[j] = *reinterpret_cast<const f2_t*>(coef + c8 * 8 + 2 * j);
          shp[j] = *reinterpret_cast<const f2_t*>(coef + CB + c8 * 8 + 2 * j);
        }
        for (int idx = threadIdx.x; idx < nchunk; idx += 512) {
          uint4* q = reinterpret_cast<uint4*>(tile_a) + idx;
          uint4 v = *q;
          v.x = f2_to_bf2_relu(ffma2(bf2_to_f2(v.x), scp[0], shp[0]));
          v.y = f2_to_bf2_relu(ffma2(bf2_to_f2(v.y), scp[1], shp[1]));
          v.z = f2_to_bf2_relu(ffma2(bf2_to_f2(v.z), scp[2], shp[2]));
          v.w = f2_to_bf2_relu(ffma2(bf2_to_f2(v.w), scp[3], shp[3]));
          if (!interior) {
            const int pi = idx >> (c8n == 8 ? 3 : 2);  // c8n = CB / 8 is 8 or 4
            const int hr = pi / HW, hc = pi - hr * HW;
            const int gh = ty0 - 1 + hr, gx = w0 - 1 + hc;
            if (gh < 0 || gh >= a.H || gx < 0 || gx >= a.W) v = make_uint4(0u, 0u, 0u, 0u);
          }
          *q = v;
        }
      }
      __syncthreads();
    }

    f2_t win[3][3][2];
    const size_t toff = (size_t)(((rg * rows) * HW + col) * rowb + c4 * 8);
    if (role == 0) {
      // ---- input gradient
      const long out_rs = (long)a.W * a.C;
      __nv_bfloat16* const out_p = a.out + (((long)b * a.H + h0) * a.W + gw) * a.C + cg;
      const uint8_t* const row0 = tile_g + toff;
      const uint8_t* const act0 = tile_a + toff + (HW + 1) * rowb;  // centre pixel of output row 0
#pragma unroll
      for (int r = 0; r < rows + 2; ++r) {
        const uint8_t* rowp = row0 + r * (HW * rowb);
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const uint2 v = *reinterpret_cast<const uint2*>(rowp + kx * rowb);
          win[r % 3][kx][0] = bf2_to_f2(v.x);
          win[r % 3][kx][1] = bf2_to_f2(v.y);
        }
        if (r >= 2) {
          const int oh = h0 + r - 2;
          f2_t acc0 = fmul2(win[(r - 2) % 3][0][0], wacc[0][0]), acc1 = fmul2(win[(r - 2) % 3][0][1], wacc[0][1]);
#pragma unroll
          for (int t = 1; t < 9; ++t) {
            acc0 = ffma2(win[(r - 2 + t / 3) % 3][t % 3][0], wacc[t][0], acc0);
            acc1 = ffma2(win[(r - 2 + t / 3) % 3][t % 3][1], wacc[t][1], acc1);
          }
          float o4[4];
          upk2(acc0, o4[0], o4[1]);
          upk2(acc1, o4[2], o4[3]);
          const uint2 sv = *reinterpret_cast<const uint2*>(act0 + (r - 2) * (HW * rowb));
          // activated values are >= 0: a bf16 is positive iff its 15 magnitude bits are non-zero
          const bool m0 = (sv.x & 0x7fffu) != 0, m1 = (sv.x & 0x7fff0000u) != 0;
          const bool m2 = (sv.y & 0x7fffu) != 0, m3 = (sv.y & 0x7fff0000u) != 0;
          float r4[4] = {0.f, 0.f, 0.f, 0.f};
          if (((r - 2) & 1) == 0) {
            const uint2 rv = rres[(r - 2) / 2];
            r4[0] = __uint_as_float(rv.x << 16); r4[1] = __uint_as_float(rv.x & 0xffff0000u);
            r4[2] = __uint_as_float(rv.y << 16); r4[3] = __uint_as_float(rv.y & 0xffff0000u);
          }
          const bool mm[4] = {m0, m1, m2, m3};
#pragma unroll
          for (int j = 0; j < 4; ++j)
            o4[j] = (a.resid == 2) ? (mm[j] ? o4[j] + r4[j] : 0.f) : ((mm[j] ? o4[j] : 0.f) + r4[j]);
          uint2 pk;
          pk.x = pack2(o4[0], o4[1]);
          pk.y = pack2(o4[2], o4[3]);
          if (a.rsums) {  // sums over the bf16-rounded dz, exactly what the apply pass reads back
            const f2_t d0 = bf2_to_f2(pk.x), d1 = bf2_to_f2(pk.y);
            bs1[0] = fadd2(bs1[0], d0); bs1[1] = fadd2(bs1[1], d1);
            bs2[0] = ffma2(d0, bf2_to_f2(sv.x), bs2[0]); bs2[1] = ffma2(d1, bf2_to_f2(sv.y), bs2[1]);
          }
          if (oh < a.H && gw < a.W) *reinterpret_cast<uint2*>(out_p + (r - 2) * out_rs) = pk;
        }
      }
    } else {
      // ---- weight gradient
      const uint8_t* const row0 = tile_a + toff;
      const uint8_t* const g0 = tile_g + toff + (HW + 1) * rowb;  // centre pixel of output row 0 (out-of-image g = zero fill)
#pragma unroll
      for (int r = 0; r < rows + 2; ++r) {
        const uint8_t* rowp = row0 + r * (HW * rowb);
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const uint2 v = *reinterpret_cast<const uint2*>(rowp + kx * rowb);
          win[r % 3][kx][0] = bf2_to_f2(v.x);
          win[r % 3][kx][1] = bf2_to_f2(v.y);
        }
        if (r >= 2) {
          const uint2 gv = *reinterpret_cast<const uint2*>(g0 + (r - 2) * (HW * rowb));
          const f2_t ga = bf2_to_f2(gv.x), gb = bf2_to_f2(gv.y);
#pragma unroll
          for (int t = 0; t < 9; ++t) {
            wacc[t][0] = ffma2(win[(r - 2 + t / 3) % 3][t % 3][0], ga, wacc[t][0]);
            wacc[t][1] = ffma2(win[(r - 2 + t / 3) % 3][t % 3][1], gb, wacc[t][1]);
          }
        }
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();  // every thread is done with this stage before it is refilled
  }
  if (role == 1 && cur_cblk >= 0) flush(cur_cblk);
  if (role == 0 && cur_cblk >= 0 && a.rsums) flush_bn(cur_cblk);
}

// Fourth generation of the one-pass depthwise backward.  ncu on dw_bwd_k (profiles/ncu_step_all_kernels_part2.md): issue
// slots 35 % busy with `barrier` the top stall - the input-gradient warps execute twice the instructions of the
// weight-gradient warps, so half the CTA waits at the end-of-tile barrier - then `no_instruction` (two different unrolled
// bodies share the instruction cache).  And with tiles dealt out round-robin a CTA changed channel block, and paid the
// 18-barrier weight-gradient flush, after almost every tile of the 32x32 / 64x64 layers.  Here
//   * every thread does BOTH jobs for (2 channels, one column, TY/RG rows): the g window it slides for the input
//     gradient holds the centre value the weight gradient needs and the activation window holds the mask / BN-sum value
//     the input gradient needs, so all 16 warps run one body of equal length (halved per-thread register arrays keep it
//     under 128 registers);
//   * a CTA owns a CONTIGUOUS range of tiles (channel-block-major order), so it flushes once or twice per launch, through
//     a [slot][11][CB] staging area summed by a column of threads instead of tap-by-tap barriers;
//   * (ncu on the first version of this kernel: the in-place activation pre-pass over the src halo was 25 % of all
//     instructions and 28 % of the stall samples, ~80 instructions per 16-byte chunk with its barrier) the activation is
//     applied to the three window values a thread loads per row instead - relu on the packed word, or FFMA2 + max for
//     relu(bn); the src halo comes through a tensor map that delivers out-of-image pixels as NaN, so both give an exact
//     zero there (max(NaN, 0) = 0) with no border masks - and the channel block is a compile-time constant
//     (CB = 64 / RG), so every shared-memory offset is an immediate.
template <int TY, int RG, int SMODE>
__global__ void __launch_bounds__(512, 1) dw_bwd2_k(const __grid_constant__ CUtensorMap tmG,
                                                    const __grid_constant__ CUtensorMap tmS, const DwArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  constexpr int CB = 64 / RG, rowb = CB * 2, nc2 = CB >> 1;
  constexpr uint32_t halo_bytes = (uint32_t)((TY + 2) * HW * rowb);
  constexpr uint32_t stage_bytes = 2u * halo_bytes;  // [g halo][src halo]
  const uint32_t bar0 = base + 2u * stage_bytes;
  float* red = reinterpret_cast<float*>(gbase + 2u * stage_bytes + 64);  // [512 / nc2 slots][11][CB]
  constexpr int rows = TY / RG;
  constexpr int nslot = 512 / nc2;
  const int tl = threadIdx.x;
  const int c2 = tl % nc2, col = (tl / nc2) % TX, rg = tl / (nc2 * TX), slot = tl / nc2;
  const int spatial = a.tiles_x * a.tiles_y * a.B;
  const int t_lo = (int)(((long)blockIdx.x * a.total_tiles) / gridDim.x);
  const int t_hi = (int)(((long)(blockIdx.x + 1) * a.total_tiles) / gridDim.x);

  if (threadIdx.x == 0) {
    mbar_init(bar0, 1);
    mbar_init(bar0 + 8u, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&tmG);
    tma_prefetch_desc(&tmS);
  }
  // the (flipped) stencil weights of the first channel block are fetched BEFORE the wait on the previous kernel: they are
  // parameters, last written by the previous step's Adam (ncu: 8 % of the kernel's samples sat on that first load)
  f2_t wk[9], wacc[9];
  int w_cblk = -1;
#pragma unroll
  for (int t = 0; t < 9; ++t) { wk[t] = 0ull; wacc[t] = 0ull; }
  if (t_lo < t_hi) {
    w_cblk = dw_fdiv(t_lo, a.m_sp);
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float2 v = __ldg(reinterpret_cast<const float2*>(a.w + (8 - t) * a.C + w_cblk * CB + c2 * 2));
      wk[t] = pk2(v.x, v.y);
    }
  }
  pdl_enter();
  __syncthreads();

  auto issue = [&](int tile, int stage) {
    const int cblk = dw_fdiv(tile, a.m_sp);
    const int sp = tile - cblk * spatial;
    const int q_ = dw_fdiv(sp, a.m_tx);
    const int tx_i = sp - q_ * a.tiles_x;
    const int b = dw_fdiv(q_, a.m_ty);
    const int ty_i = q_ - b * a.tiles_y;
    mbar_expect_tx(bar0 + 8u * stage, stage_bytes);
    tma_load_4d(base + stage * stage_bytes, &tmG, bar0 + 8u * stage, cblk * CB, tx_i * TX - 1, ty_i * TY - 1, b);
    tma_load_4d(base + stage * stage_bytes + halo_bytes, &tmS, bar0 + 8u * stage, cblk * CB, tx_i * TX - 1, ty_i * TY - 1, b);
  };
  if (threadIdx.x == 0 && t_lo < t_hi) issue(t_lo, 0);

  f2_t bs1 = 0ull, bs2 = 0ull;  // BN-backward reductions of the layer below: sum dz and sum dz * a (see dw_bwd_k)
  f2_t sc2 = 0ull, sh2 = 0ull;  // relu(bn) source: scale / shift of this thread's two channels
  int cur_cblk = -1;

  auto flush = [&](int cblk) {  // all 512 threads
    float* mine = red + slot * 11 * CB + c2 * 2;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      float lo, hi;
      upk2(wacc[t], lo, hi);
      *reinterpret_cast<float2*>(mine + t * CB) = make_float2(lo, hi);
      wacc[t] = 0ull;
    }
    {
      float lo, hi;
      upk2(bs1, lo, hi);
      *reinterpret_cast<float2*>(mine + 9 * CB) = make_float2(lo, hi);
      upk2(bs2, lo, hi);
      *reinterpret_cast<float2*>(mine + 10 * CB) = make_float2(lo, hi);
      bs1 = bs2 = 0ull;
    }
    __syncthreads();
    for (int i = tl; i < 11 * CB; i += 512) {
      float v = 0.f;
#pragma unroll 8
      for (int q = 0; q < nslot; ++q) v += red[q * 11 * CB + i];
      red[i] = v;  // slot 0, element i: read and written by this thread only
    }
    __syncthreads();
    for (int i = tl; i < 9 * CB; i += 512) {
      const int t = i / CB, c = i - t * CB;
      atomicAdd(a.dwk + (long)t * a.C + cblk * CB + c, red[i]);
    }
    if (a.rsums && tl < CB) {
      const int c = cblk * CB + tl;
      const float S1 = red[9 * CB + tl], S2 = red[10 * CB + tl];
      const float sc = a.rbn[c], sh = a.rbn[a.C + c], mu = a.rbn[2 * a.C + c];
      const float beta = sh + mu * sc;
      atomicAdd(a.rsums + c, S1);
      atomicAdd(a.rsums + a.C + c, sc != 0.f ? (S2 - beta * S1) / sc : 0.f);
    }
    __syncthreads();
  };

  int it = 0;
  for (int tile = t_lo; tile < t_hi; ++tile, ++it) {
    const int stage = it & 1;
    if (threadIdx.x == 0 && tile + 1 < t_hi) issue(tile + 1, stage ^ 1);  // its last readers passed the barrier at the loop tail
    const int cblk = dw_fdiv(tile, a.m_sp);
    const int sp = tile - cblk * spatial;
    const int q_ = dw_fdiv(sp, a.m_tx);
    const int tx_i = sp - q_ * a.tiles_x;
    const int b = dw_fdiv(q_, a.m_ty);
    const int ty_i = q_ - b * a.tiles_y;
    const int cg = cblk * CB + c2 * 2;
    if (cblk != cur_cblk) {  // uniform over the CTA
      if (cur_cblk >= 0) flush(cur_cblk);
      cur_cblk = cblk;
      if (cblk != w_cblk) {
        w_cblk = cblk;
#pragma unroll
        for (int t = 0; t < 9; ++t) {
          const float2 v = __ldg(reinterpret_cast<const float2*>(a.w + (8 - t) * a.C + cg));
          wk[t] = pk2(v.x, v.y);
        }
      }
      if (SMODE == 1) {
        const float2 s_ = __ldg(reinterpret_cast<const float2*>(a.src_bn + cg));
        const float2 h_ = __ldg(reinterpret_cast<const float2*>(a.src_bn + a.C + cg));
        sc2 = pk2(s_.x, s_.y); sh2 = pk2(h_.x, h_.y);
      }
    }
    const int w0 = tx_i * TX, ty0 = ty_i * TY, h0 = ty0 + rg * rows;
    const int gw = w0 + col;
    uint32_t rres[(rows + 1) / 2];
#pragma unroll
    for (int i = 0; i < (rows + 1) / 2; ++i) {
      rres[i] = 0u;
      const int oh = h0 + 2 * i;  // h0 is even
      if (a.resid && !(col & 1) && oh < a.H && gw < a.W)
        rres[i] = __ldg(reinterpret_cast<const uint32_t*>(
            a.resid_t + (((long)b * (a.H >> 1) + (oh >> 1)) * (a.W >> 1) + (gw >> 1)) * a.C + cg));
    }
    mbar_wait(bar0 + 8u * stage, (uint32_t)(it >> 1) & 1u);
    const uint8_t* tile_g = gbase + stage * stage_bytes;
    const uint8_t* tile_a = tile_g + halo_bytes;

    f2_t wg[3][3], wa[3][3];
    const uint32_t toff = (uint32_t)(((rg * rows) * HW + col) * rowb + c2 * 4);
    const long out_rs = (long)a.W * a.C;
    __nv_bfloat16* const out_p = a.out + (((long)b * a.H + h0) * a.W + gw) * a.C + cg;
    const uint8_t* const g_row0 = tile_g + toff;
    const uint8_t* const a_row0 = tile_a + toff;
#pragma unroll
    for (int r = 0; r < rows + 2; ++r) {
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        wg[r % 3][kx] = bf2_to_f2(*reinterpret_cast<const uint32_t*>(g_row0 + (r * HW + kx) * rowb));
        const uint32_t av = *reinterpret_cast<const uint32_t*>(a_row0 + (r * HW + kx) * rowb);
        if (SMODE == 0) {
          wa[r % 3][kx] = bf2_to_f2(relu_bf2(av));
        } else {
          const f2_t y_ = ffma2(bf2_to_f2(av), sc2, sh2);  // out-of-image pixels arrive as NaN: max(NaN, 0) = 0
          float lo, hi;
          upk2(y_, lo, hi);
          wa[r % 3][kx] = pk2(fmaxf(lo, 0.f), fmaxf(hi, 0.f));
        }
      }
      if (r >= 2) {
        const int oh = h0 + r - 2;
        // input gradient: three independent partial sums (one per window row) instead of one 9-deep chain
        f2_t p[3];
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
          p[ky] = fmul2(wg[(r - 2 + ky) % 3][0], wk[ky * 3]);
          p[ky] = ffma2(wg[(r - 2 + ky) % 3][1], wk[ky * 3 + 1], p[ky]);
          p[ky] = ffma2(wg[(r - 2 + ky) % 3][2], wk[ky * 3 + 2], p[ky]);
        }
        const f2_t acc = fadd2(fadd2(p[0], p[1]), p[2]);
        // weight gradient: the centre of the g window is g at this output pixel (out-of-image g = TMA zero fill)
        const f2_t gc = wg[(r - 1) % 3][1];
#pragma unroll
        for (int t = 0; t < 9; ++t) wacc[t] = ffma2(wa[(r - 2 + t / 3) % 3][t % 3], gc, wacc[t]);
        const f2_t ac = wa[(r - 1) % 3][1];  // activated forward input at this pixel: > 0 <=> the relu passed it
        float o0, o1, a0, a1;
        upk2(acc, o0, o1);
        upk2(ac, a0, a1);
        if (((r - 2) & 1) == 0) {  // (compile time: the residual-path gradient lives on even rows)
          const uint32_t rv = rres[(r - 2) / 2];
          const float r0 = __uint_as_float(rv << 16), r1 = __uint_as_float(rv & 0xffff0000u);
          if (a.resid == 2) {
            o0 = a0 > 0.f ? o0 + r0 : 0.f;
            o1 = a1 > 0.f ? o1 + r1 : 0.f;
          } else {
            o0 = (a0 > 0.f ? o0 : 0.f) + r0;
            o1 = (a1 > 0.f ? o1 : 0.f) + r1;
          }
        } else {
          o0 = a0 > 0.f ? o0 : 0.f;
          o1 = a1 > 0.f ? o1 : 0.f;
        }
        const bool inimg = oh < a.H && gw < a.W;
        const uint32_t pk = inimg ? pack2(o0, o1) : 0u;
        if (a.rsums) {  // sums over the bf16-rounded dz, exactly what the apply pass reads back
          const f2_t d = bf2_to_f2(pk);
          bs1 = fadd2(bs1, d);
          bs2 = ffma2(d, ac, bs2);
        }
        if (inimg) *reinterpret_cast<uint32_t*>(out_p + (r - 2) * out_rs) = pk;
      }
    }
    __syncthreads();  // every thread is done reading this stage before it is refilled
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

int dw_per_sm() {  // CTAs per SM of the persistent depthwise kernels (tuning knob; 2 = what fits)
  static int v = -1;
  if (v < 0) { const char* e = getenv("CRACKFL_DW_CTAS_PER_SM"); v = e ? atoi(e) : 2; if (v < 1 || v > 2) v = 2; }
  return v;
}

bool shape_ok(int C) { return C % 32 == 0 && (C <= 64 || C % 64 == 0); }

// channel block and ring depth: as many stages (<= 4) as fit twice per SM (2 x <= 112 KB); CRACKFL_DW_CB /
// CRACKFL_DW_STAGES override for tuning.  Measured (bench/dw_bench.py, 128x128x64): ring depth 2, 3 or 4 makes no
// difference (the kernels are bound by the shared FMA/ALU pipe, ncu math_pipe_throttle, not by copy latency) and
// 32-channel blocks are 10-25 % slower than 64 (twice the per-tile overhead), so 64-channel blocks x 2 stages stay.
void pick_ring(DwArgs& a, size_t stage_bytes_per_channel, size_t extra) {
  static int env_cb = -1, env_ns = -1;
  if (env_cb < 0) { const char* e = getenv("CRACKFL_DW_CB"); env_cb = e ? atoi(e) : 0; }
  if (env_ns < 0) { const char* e = getenv("CRACKFL_DW_STAGES"); env_ns = e ? atoi(e) : 0; }
  a.CB = (a.C >= 64 && env_cb != 32) ? 64 : 32;
  int ns = (int)((112 * 1024 - extra) / (stage_bytes_per_channel * a.CB));
  if (ns > 4) ns = 4;
  if (env_ns >= 2 && env_ns <= ns) ns = env_ns;
  if (ns < 2) ns = 2;
  a.nstage = ns;
}

// forward (src == NULL) or input gradient (src != NULL, flip = 1, act_mode = 2)
template <int TY, bool DGRAD>
static int launch_stencil(const void* x, DwArgs& a, cudaStream_t st) {
  const size_t extra = 64 + 2 * 64 * sizeof(float) + 128;
  pick_ring(a, ((size_t)(TY + 2) * HW + (DGRAD ? (size_t)TY * TX : 0)) * 2, extra);
  a.tiles_x = (a.W + TX - 1) / TX; a.tiles_y = (a.H + TY - 1) / TY;
  a.total_tiles = a.tiles_x * a.tiles_y * a.B * (a.C / a.CB);
  a.m_sp = dw_magic(a.tiles_x * a.tiles_y * a.B); a.m_tx = dw_magic(a.tiles_x); a.m_ty = dw_magic(a.tiles_y);
  CUtensorMap tm, tmS;
  const int d[4] = {a.C, a.W, a.H, a.B}, bx[4] = {a.CB, HW, TY + 2, 1}, bs[4] = {a.CB, TX, TY, 1};
  int rc = ck::get_tensor_map(x, 4, d, bx, 0, &tm);
  if (rc) return rc;
  rc = ck::get_tensor_map(DGRAD ? (const void*)a.src : x, 4, d, DGRAD ? bs : bx, 0, &tmS);
  if (rc) return rc;
  const size_t stage = (size_t)(TY + 2) * HW * a.CB * 2 + (DGRAD ? (size_t)TY * TX * a.CB * 2 : 0);
  const size_t smem = (size_t)a.nstage * stage + extra;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(dw_stencil_k<TY, 1, DGRAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(dw_stencil_k<TY, 2, DGRAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  int grid = dw_per_sm() * num_sms();
  if (grid > a.total_tiles) grid = a.total_tiles;
  if (a.CB == 64) ck::launch(dw_stencil_k<TY, 1, DGRAD>, grid, 256, smem, st, tm, tmS, a);
  else ck::launch(dw_stencil_k<TY, 2, DGRAD>, grid, 256, smem, st, tm, tmS, a);
  CK_LAUNCH_CHECK();
  return 0;
}

}  // namespace

extern "C" {

int crackfl_dw3x3_tma(const void* x, const float* w, const float* bn, const crackfl_bn_fin* fin, void* out, const void* src,
                      const float* src_bn, const void* resid_t, int B, int H, int W, int C, int act_mode, int flip,
                      int src_mode, int resid, cudaStream_t st) {
  if (!shape_ok(C)) return (int)cudaErrorInvalidValue;
  DwArgs a;
  memset(&a, 0, sizeof(a));
  a.w = w; a.bn = bn; if (fin) a.fin = *fin;
  a.out = (__nv_bfloat16*)out; a.src = (const __nv_bfloat16*)src; a.src_bn = src_bn; a.resid_t = (const __nv_bfloat16*)resid_t;
  a.B = B; a.H = H; a.W = W; a.C = C; a.CB = C >= 64 ? 64 : 32;
  a.act_mode = act_mode; a.flip = flip; a.src_mode = src_mode; a.resid = resid;
  if (src) {
    if (act_mode != 2) return (int)cudaErrorInvalidValue;  // the gradient variant takes g as is
    return launch_stencil<8, true>(x, a, st);
  }
  static int gen = -1;  // CRACKFL_DW_FWD=2: the pre-pass kernel (dw_stencil_k); default: activation on load (dw_fwd2_k)
  if (gen < 0) { const char* e = getenv("CRACKFL_DW_FWD"); gen = e ? atoi(e) : 3; }
  if (gen == 2) return launch_stencil<16, false>(x, a, st);
  constexpr int TY = 16;
  const size_t extra = 64 + 128;
  pick_ring(a, (size_t)(TY + 2) * HW * 2, extra);
  a.tiles_x = (a.W + TX - 1) / TX; a.tiles_y = (a.H + TY - 1) / TY;
  a.total_tiles = a.tiles_x * a.tiles_y * a.B * (a.C / a.CB);
  a.m_sp = dw_magic(a.tiles_x * a.tiles_y * a.B); a.m_tx = dw_magic(a.tiles_x); a.m_ty = dw_magic(a.tiles_y);
  CUtensorMap tm;
  const int d[4] = {a.C, a.W, a.H, a.B}, bx[4] = {a.CB, HW, TY + 2, 1};
  // relu / relu(bn) sources: out-of-image pixels arrive as NaN and leave the activation as exact zeros
  int rc = ck::get_tensor_map(x, 4, d, bx, act_mode == 2 ? 0 : ck::CK_MAP_OOB_NAN, &tm);
  if (rc) return rc;
  const size_t smem = (size_t)a.nstage * (TY + 2) * HW * a.CB * 2 + extra;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaSuccess;
#define CK_DWF_ATTR(RG_, AM_) if (e == cudaSuccess) e = cudaFuncSetAttribute(dw_fwd2_k<TY, RG_, AM_>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
    CK_DWF_ATTR(1, 0) CK_DWF_ATTR(1, 1) CK_DWF_ATTR(1, 2) CK_DWF_ATTR(2, 0) CK_DWF_ATTR(2, 1) CK_DWF_ATTR(2, 2)
#undef CK_DWF_ATTR
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  int grid = dw_per_sm() * num_sms();
  if (grid > a.total_tiles) grid = a.total_tiles;
#define CK_DWF_GO(RG_, AM_) ck::launch(dw_fwd2_k<TY, RG_, AM_>, grid, 256, smem, st, tm, a)
  if (a.CB == 64) { if (act_mode == 0) CK_DWF_GO(1, 0); else if (act_mode == 1) CK_DWF_GO(1, 1); else CK_DWF_GO(1, 2); }
  else { if (act_mode == 0) CK_DWF_GO(2, 0); else if (act_mode == 1) CK_DWF_GO(2, 1); else CK_DWF_GO(2, 2); }
#undef CK_DWF_GO
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
  a.B = B; a.H = H; a.W = W; a.C = C;
  a.act_mode = mode;
  const size_t extra = 64 + 256 * 4 * sizeof(float) + 2 * 64 * sizeof(float) + 128;
  pick_ring(a, ((size_t)(TY + 2) * HW + (size_t)TY * TX) * 2, extra);
  a.tiles_x = (W + TX - 1) / TX; a.tiles_y = (H + TY - 1) / TY;
  a.total_tiles = a.tiles_x * a.tiles_y * B * (C / a.CB);
  a.m_sp = dw_magic(a.tiles_x * a.tiles_y * a.B); a.m_tx = dw_magic(a.tiles_x); a.m_ty = dw_magic(a.tiles_y);
  CUtensorMap tmA, tmG;
  const int d[4] = {C, W, H, B}, ba[4] = {a.CB, HW, TY + 2, 1}, bg[4] = {a.CB, TX, TY, 1};
  int rc = ck::get_tensor_map(src, 4, d, ba, 0, &tmA);
  if (rc) return rc;
  rc = ck::get_tensor_map(g, 4, d, bg, 0, &tmG);
  if (rc) return rc;
  const size_t smem = (size_t)a.nstage * ((size_t)(TY + 2) * HW + TY * TX) * a.CB * 2 + extra;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(dw_wgrad_k<TY, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dw_wgrad_k<TY, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  int grid = dw_per_sm() * num_sms();
  if (grid > a.total_tiles) grid = a.total_tiles;
  if (a.CB == 64) ck::launch(dw_wgrad_k<TY, 1>, grid, 256, smem, st, tmA, tmG, a);
  else ck::launch(dw_wgrad_k<TY, 2>, grid, 256, smem, st, tmA, tmG, a);
  CK_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"

namespace {
// dw_bwd2_k with TY tile rows (a: everything but the tile geometry filled in)
template <int TY>
int launch_bwd2(const void* g, const void* src, DwArgs& a, cudaStream_t st) {
  a.tiles_x = (a.W + TX - 1) / TX; a.tiles_y = (a.H + TY - 1) / TY;
  a.total_tiles = a.tiles_x * a.tiles_y * a.B * (a.C / a.CB);
  a.m_sp = dw_magic(a.tiles_x * a.tiles_y * a.B); a.m_tx = dw_magic(a.tiles_x); a.m_ty = dw_magic(a.tiles_y);
  CUtensorMap tmG, tmS;
  const int d[4] = {a.C, a.W, a.H, a.B}, bx[4] = {a.CB, HW, TY + 2, 1};
  int rc = ck::get_tensor_map(g, 4, d, bx, 0, &tmG);
  if (rc) return rc;
  rc = ck::get_tensor_map(src, 4, d, bx, ck::CK_MAP_OOB_NAN, &tmS);
  if (rc) return rc;
  const size_t smem = 2 * 2 * (size_t)(TY + 2) * HW * a.CB * 2 + 64 + (size_t)(512 / (a.CB / 2)) * 11 * a.CB * sizeof(float) + 128;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(dw_bwd2_k<TY, 1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dw_bwd2_k<TY, 1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dw_bwd2_k<TY, 2, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dw_bwd2_k<TY, 2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  int grid = num_sms();
  if (grid > a.total_tiles) grid = a.total_tiles;
  if (a.CB == 64) {
    if (a.src_mode == 1) ck::launch(dw_bwd2_k<TY, 1, 1>, grid, 512, smem, st, tmG, tmS, a);
    else ck::launch(dw_bwd2_k<TY, 1, 0>, grid, 512, smem, st, tmG, tmS, a);
  } else {
    if (a.src_mode == 1) ck::launch(dw_bwd2_k<TY, 2, 1>, grid, 512, smem, st, tmG, tmS, a);
    else ck::launch(dw_bwd2_k<TY, 2, 0>, grid, 512, smem, st, tmG, tmS, a);
  }
  CK_LAUNCH_CHECK();
  return 0;
}
}  // namespace

extern "C" {

// dz (input gradient) and dW (weight gradient) of the depthwise 3x3 in one pass; arguments as crackfl_dw3x3_tma (gradient
// variant) + crackfl_dw3x3_wgrad_tma
int crackfl_dw3x3_bwd_tma_bn(const void* g, const void* src, const float* w, const float* src_bn, const void* resid_t, void* dz,
                             float* dwk, int B, int H, int W, int C, int src_mode, int resid, const float* red_bn,
                             float* red_sums, cudaStream_t st);

int crackfl_dw3x3_bwd_tma(const void* g, const void* src, const float* w, const float* src_bn, const void* resid_t, void* dz,
                          float* dwk, int B, int H, int W, int C, int src_mode, int resid, cudaStream_t st) {
  return crackfl_dw3x3_bwd_tma_bn(g, src, w, src_bn, resid_t, dz, dwk, B, H, W, C, src_mode, resid, nullptr, nullptr, st);
}

// + the BN-backward reductions of the BN that produced src's activation (red_bn = its [scale][shift][mean][invstd],
// red_sums = [sum dz C][sum dz*(y-mean) C], pre-zeroed): saves the separate crackfl_bn_bwd_reduce pass over dz and y
int crackfl_dw3x3_bwd_tma_bn(const void* g, const void* src, const float* w, const float* src_bn, const void* resid_t, void* dz,
                             float* dwk, int B, int H, int W, int C, int src_mode, int resid, const float* red_bn,
                             float* red_sums, cudaStream_t st) {
  if (!shape_ok(C)) return (int)cudaErrorInvalidValue;
  constexpr int TY = 8;
  DwArgs a;
  memset(&a, 0, sizeof(a));
  a.w = w; a.src_bn = src_bn; a.out = (__nv_bfloat16*)dz; a.src = (const __nv_bfloat16*)src;
  a.resid_t = (const __nv_bfloat16*)resid_t; a.dwk = dwk;
  a.B = B; a.H = H; a.W = W; a.C = C; a.CB = C >= 64 ? 64 : 32;
  a.act_mode = 2; a.flip = 1; a.src_mode = src_mode; a.resid = resid; a.nstage = 2;
  a.rbn = red_bn; a.rsums = red_bn ? red_sums : nullptr;
  a.tiles_x = (W + TX - 1) / TX; a.tiles_y = (H + TY - 1) / TY;
  a.total_tiles = a.tiles_x * a.tiles_y * B * (C / a.CB);
  a.m_sp = dw_magic(a.tiles_x * a.tiles_y * a.B); a.m_tx = dw_magic(a.tiles_x); a.m_ty = dw_magic(a.tiles_y);
  CUtensorMap tmG, tmS;
  const int d[4] = {C, W, H, B}, bx[4] = {a.CB, HW, TY + 2, 1};
  int rc = ck::get_tensor_map(g, 4, d, bx, 0, &tmG);
  if (rc) return rc;
  rc = ck::get_tensor_map(src, 4, d, bx, 0, &tmS);
  if (rc) return rc;
  const size_t smem = 2 * 2 * (size_t)(TY + 2) * HW * a.CB * 2 + 64 + 256 * 4 * sizeof(float) + 2 * 64 * sizeof(float) +
                      (256 * 4 + 2 * 64) * sizeof(float) + 128;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(dw_bwd_k<TY, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dw_bwd_k<TY, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  int grid = num_sms();
  if (grid > a.total_tiles) grid = a.total_tiles;
  static int gen = -1;  // CRACKFL_DW_BWD=3: the role-split kernel; default: every thread does both gradients (dw_bwd2_k)
  if (gen < 0) { const char* e = getenv("CRACKFL_DW_BWD"); gen = e ? atoi(e) : 4; }
  if (gen != 3) {
    // tile rows per channel block width (CRACKFL_DW_BWD_TY64 / _TY32 = 8 | 16): 16 halves the halo rows per output row
    static int ty64 = -1, ty32 = -1;
    if (ty64 < 0) { const char* e = getenv("CRACKFL_DW_BWD_TY64"); ty64 = e ? atoi(e) : 16; }
    if (ty32 < 0) { const char* e = getenv("CRACKFL_DW_BWD_TY32"); ty32 = e ? atoi(e) : 16; }
    const int ty = a.CB == 64 ? ty64 : ty32;
    return (ty == 16 && H >= 16) ? launch_bwd2<16>(g, src, a, st) : launch_bwd2<8>(g, src, a, st);
  }
  if (a.CB == 64) ck::launch(dw_bwd_k<TY, 1>, grid, 512, smem, st, tmG, tmS, a);
  else ck::launch(dw_bwd_k<TY, 2>, grid, 512, smem, st, tmG, tmS, a);
  CK_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
