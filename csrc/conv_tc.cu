// tcgen05 / TMEM / TMA implicit-GEMM convolution kernels for sm_100a  (SURVEY section 8 rows K1-K3).
//
//   conv_tc_fwd : Y[p][n] = sum_{t,c} X[p + off(t)][c] * Wk[n][t*Cin + c]  (+bias, +add, BN sum/sumsq epilogue)
//                 3x3 stride-1 "same" conv == Keras Conv2DTranspose(3, stride 1, same) with flipped taps
//                 (client_fit_model.py:129,133), every 1x1 conv (pointwise of SeparableConv2D :109,113 and
//                 the residual projections :119,140) and, with the transposed operand, all their dgrads.
//   wgrad_tc    : dW[t][ci][co] += sum_p X[p + off(t)][ci] * dY[p][co]      (fp32 atomics into the Keras-layout arena)
//
// Structure (one CTA = one 128-pixel x BN-channel output tile, 192 threads):
//   warp 0   : TMA producer.  The 128 pixels are a TB x TH x TW box of the NHWC tensor; for tap t the box is
//              simply shifted by (dh,dw) and TMA's out-of-bounds zero fill IS the "same" padding, so the
//              im2col matrix is never built.  A and B tiles land in shared memory in the 128B/64B-swizzled
//              K-major layout that tcgen05 reads directly.  Multi-stage ring, mbarrier full/empty pairs.
//   warp 1   : allocates TMEM, then one elected lane issues tcgen05.mma (M=128, N=BN, K=16) per 32 bytes of K,
//              accumulating in TMEM; tcgen05.commit releases smem stages and finally signals the epilogue.
//   warps 2-5: epilogue.  tcgen05.ld 32 lanes x 32 columns per warp, +bias, +residual, bf16 pack, 64-byte
//              contiguous stores per thread, and (optionally) the BatchNorm per-channel sum / sum-of-squares of
//              the *stored* values via a 31-shuffle recursive-halving transpose-reduce + one atomic per column.
// wgrad uses the same TMA boxes but feeds them as MN-major operands (the pixel axis is K), pairs/quadruples
// (tap, channel-chunk) blocks to fill M=128 and splits the pixel axis over CTAs.
#include <cuda.h>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <string>

#include "common.cuh"
#include "tc_common.cuh"
#include "../include/crackfl.h"

namespace {

// Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
//   [0,14) start>>4 | [16,30) LBO>>4 | [32,46) SBO>>4 | [46,48) version=1 | [49,52) base_offset | [61,64) layout
// layout: 2 = SWIZZLE_128B, 4 = SWIZZLE_64B.
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}

// Instruction descriptor for kind::f16, bf16 x bf16 -> fp32 (cute::UMMA::InstrDescriptor bit layout)
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

struct TileGeom {
  int TW, TH, TB;           // pixel box, TW*TH*TB == 128
  int tiles_w, tiles_h, tiles_b;
  // ceil(2^40 / d) for d = tiles_w, tiles_h, tiles per n-tile: x / d == (x * m) >> 40 for x * d < 2^40.  Every role of
  // the persistent kernels decomposes its tile index once per tile; with runtime divisors that was three divide
  // sequences (~25 instructions each) per thread per tile, ~10 % of all instructions of a pointwise GEMM launch
  // (ncu source page, conv_tc_fwd_v2_k 64->64 @128^2: 11.7 M warp instructions, MUFU.RCP in every code region).
  unsigned long long m_w, m_h, m_ntm;
};

__host__ __device__ __forceinline__ unsigned long long fd_magic(int d) { return ((1ull << 40) + (unsigned long long)d - 1) / (unsigned long long)d; }
__device__ __forceinline__ int fd_div(int x, unsigned long long m) { return (int)(((unsigned long long)(unsigned)x * m) >> 40); }
// tile -> (n-tile, w-tile, h-tile, b-tile) without divide sequences
__device__ __forceinline__ void tile_coords(int tile, int n_tiles_m, const TileGeom& g, int& nt, int& tw_i, int& th_i, int& tb_i) {
  nt = fd_div(tile, g.m_ntm);
  int tm_ = tile - nt * n_tiles_m;
  int q = fd_div(tm_, g.m_w);
  tw_i = tm_ - q * g.tiles_w;
  tb_i = fd_div(q, g.m_h);
  th_i = q - tb_i * g.tiles_h;
}

struct ConvArgs {
  const float* bias;
  const __nv_bfloat16* add;
  __nv_bfloat16* Y;
  float* stats;
  int B, H, W, Cin, N, taps;
  int BN, kc, stages;
  uint32_t layout, sbo;     // swizzle mode of A/B tiles (2: 128B, 4: 64B) and its 8-row group stride
  TileGeom g;
};

// Column sums of a 32(lanes) x 32(columns) register tile via recursive halving: after the 5 exchange steps
// lane L holds the total of column L.  31 shuffles instead of 160.
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    const bool up = (lane & o) != 0;
#pragma unroll
    for (int j = 0; j < o; ++j) {
      const float send = up ? v[j] : v[j + o];
      const float keep = up ? v[j + o] : v[j];
      v[j] = keep + __shfl_xor_sync(0xffffffffu, send, o);
    }
  }
  return v[0];
}

__global__ void __launch_bounds__(192) conv_tc_fwd_k(const __grid_constant__ CUtensorMap tmA,
                                                     const __grid_constant__ CUtensorMap tmB, const ConvArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_bytes = 128u * a.kc * 2u, b_bytes = (uint32_t)a.BN * a.kc * 2u;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t bar_base = base + a.stages * stage_bytes;  // full[s], empty[s], tmem_full, tmem_ptr
  const uint32_t full0 = bar_base, empty0 = bar_base + 8u * a.stages, tfull = bar_base + 16u * a.stages;
  const uint32_t tptr = tfull + 8u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // tile coordinates
  int tile = blockIdx.x;
  const int tw_i = tile % a.g.tiles_w; tile /= a.g.tiles_w;
  const int th_i = tile % a.g.tiles_h; tile /= a.g.tiles_h;
  const int tb_i = tile;
  const int w0 = tw_i * a.g.TW, h0 = th_i * a.g.TH, b0 = tb_i * a.g.TB;
  const int n0 = blockIdx.y * a.BN;
  const int kchunks = a.Cin / a.kc;
  const int iters = a.taps * kchunks;

  if (threadIdx.x == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(full0 + 8u * s, 1);
      mbar_init(empty0 + 8u * s, 1);
    }
    mbar_init(tfull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) tmem_alloc(tptr, (uint32_t)(a.BN < 32 ? 32 : a.BN));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  ck::pdl_enter();  // set-up above overlapped the previous kernel's tail; global memory is touched only from here on
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tptr));

  if (warp == 0) {
    if (lane == 0) {
      for (int it = 0; it < iters; ++it) {
        const int s = it % a.stages;
        const uint32_t ph = (uint32_t)(it / a.stages) & 1u;
        mbar_wait(empty0 + 8u * s, ph ^ 1u);
        const int t = it / kchunks, c = it % kchunks;
        const int dh = a.taps == 9 ? t / 3 - 1 : 0, dw = a.taps == 9 ? t % 3 - 1 : 0;
        const uint32_t sa = base + s * stage_bytes, sb = sa + a_bytes;
        mbar_expect_tx(full0 + 8u * s, stage_bytes);
        tma_load_4d(sa, &tmA, full0 + 8u * s, c * a.kc, w0 + dw, h0 + dh, b0);
        tma_load_2d(sb, &tmB, full0 + 8u * s, t * a.Cin + c * a.kc, n0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc(128, a.BN, 0, 0);
      for (int it = 0; it < iters; ++it) {
        const int s = it % a.stages;
        const uint32_t ph = (uint32_t)(it / a.stages) & 1u;
        mbar_wait(full0 + 8u * s, ph);
        tc_fence_after();
        const uint32_t sa = base + s * stage_bytes, sb = sa + a_bytes;
        const uint64_t ad = make_desc(sa, 16, a.sbo, a.layout);
        const uint64_t bd = make_desc(sb, 16, a.sbo, a.layout);
        for (int k = 0; k < a.kc / 16; ++k)  // +32 bytes of K per MMA: start-address field += 2
          umma_f16(tmem_base, ad + 2u * k, bd + 2u * k, idesc, (it > 0 || k > 0) ? 1u : 0u);
        umma_commit(empty0 + 8u * s);
      }
      umma_commit(tfull);
    }
  } else {
    // ---- epilogue warps: TMEM lane quarter = warp % 4
    const int q = warp & 3;
    const int r = q * 32 + lane;  // accumulator row == pixel index inside the box (tw fastest)
    const int tw = r % a.g.TW, th = (r / a.g.TW) % a.g.TH, tb = r / (a.g.TW * a.g.TH);
    const int pw = w0 + tw, ph_ = h0 + th, pb = b0 + tb;
    const bool valid = pw < a.W && ph_ < a.H && pb < a.B;
    const long pix = ((long)pb * a.H + ph_) * a.W + pw;
    mbar_wait(tfull, 0);
    tc_fence_after();
    for (int c0 = 0; c0 < a.BN; c0 += 32) {
      uint32_t raw[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, raw);
      float v[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(raw[j]);
      if (a.bias) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 bb = __ldg(reinterpret_cast<const float4*>(a.bias + n0 + c0 + j));
          v[j] += bb.x; v[j + 1] += bb.y; v[j + 2] += bb.z; v[j + 3] += bb.w;
        }
      }
      if (a.add && valid) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          float t8[8];
          ck::unpack8(ck::ldg16(a.add + pix * a.N + n0 + c0 + j), t8);
#pragma unroll
          for (int e = 0; e < 8; ++e) v[j + e] += t8[e];
        }
      }
      if (valid) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) ck::stg16(a.Y + pix * a.N + n0 + c0 + j, ck::pack8(&v[j]));
      }
      if (a.stats) {
        float s1[32], s2[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          // statistics of the value as stored (bf16-rounded), zero for masked pixels
          const float x = valid ? __bfloat162float(__float2bfloat16_rn(v[j])) : 0.f;
          s1[j] = x;
          s2[j] = x * x;
        }
        const float t1 = warp_colsum32(s1, lane);
        const float t2 = warp_colsum32(s2, lane);
        atomicAdd(a.stats + n0 + c0 + lane, t1);
        atomicAdd(a.stats + a.N + n0 + c0 + lane, t2);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)(a.BN < 32 ? 32 : a.BN));
  }
}

// ----------------------------------------------------------------------------------------------------------
// v2: persistent CTAs, double-buffered TMEM accumulators, TMA-store epilogue.
// The v1 kernel above pays its prologue (barrier init, descriptor fetch, TMEM alloc) and an un-overlapped epilogue
// once per 128-pixel tile; for the thin 1x1 / small-K layers of this network that is most of the tile's life.
// Here a CTA loops over tiles (m fastest, so consecutive tiles reuse the weight tile in L2), the producer runs
// ahead across tile boundaries, the MMA warp alternates between two TMEM accumulator stages, and the epilogue
// warps drain stage s (tcgen05.ld -> bias/add/stats -> bf16 -> swizzled smem -> one TMA bulk store, which also
// clips out-of-bounds pixels) while the tensor core already fills stage s^1.
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* tm, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(tm)), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
// 8 epilogue warps: thin-K layers (1x1 convs, K = 32..64) are epilogue-bound, so two warps share each TMEM lane
// quarter and split the accumulator columns between them.
constexpr int EPI_WARPS = 8;
constexpr int EPI_THREADS = EPI_WARPS * 32;
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

struct ConvArgs2 {
  const float* bias;
  const __nv_bfloat16* add;
  float* stats;
  int B, H, W, Cin, N, taps;
  int BN, kc, stages;
  int sub_w;                 // channels per output sub-tile (64 with 128B swizzle, 32 with 64B swizzle)
  int n_tiles_m, n_tiles_n;  // tile counts
  uint32_t layout, sbo;
  int dbg_noflush;           // profiling aid (CRACKFL_DBG_NOFLUSH=1): skip the statistics atomics
  TileGeom g;
};

__global__ void __launch_bounds__(64 + EPI_THREADS) conv_tc_fwd_v2_k(const __grid_constant__ CUtensorMap tmA,
                                                        const __grid_constant__ CUtensorMap tmB,
                                                        const __grid_constant__ CUtensorMap tmY, const ConvArgs2 a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_bytes = 128u * a.kc * 2u, b_bytes = (uint32_t)a.BN * a.kc * 2u;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t out_bytes = 128u * (uint32_t)a.BN * 2u;               // one staging buffer
  const uint32_t out0 = base + a.stages * stage_bytes;                  // 2 staging buffers
  const uint32_t bar_base = out0 + 2u * out_bytes;
  const uint32_t full0 = bar_base, empty0 = bar_base + 8u * a.stages;
  const uint32_t tfull0 = bar_base + 16u * a.stages, tempty0 = tfull0 + 16u;
  const uint32_t tptr = tempty0 + 16u;
  const uint32_t stat_smem = tptr + 16u;  // EPI_WARPS x 2 x 128 floats
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int kchunks = a.Cin / a.kc;
  const int iters = a.taps * kchunks;
  const int total_tiles = a.n_tiles_m * a.n_tiles_n;
  const uint32_t ncols = (uint32_t)(2 * a.BN < 32 ? 32 : 2 * a.BN);

  if (threadIdx.x == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(full0 + 8u * s, 1);
      mbar_init(empty0 + 8u * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull0 + 8u * s, 1);
      mbar_init(tempty0 + 8u * s, EPI_WARPS);  // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmY);
  }
  if (warp == 1) tmem_alloc(tptr, ncols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  ck::pdl_enter();  // set-up above overlapped the previous kernel's tail; global memory is touched only from here on
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tptr));

  if (warp == 0) {
    // ---- producer: uniform control flow, one elected lane issues (see tc_common.cuh)
    uint32_t it_g = 0;  // global k-iteration counter across tiles (ring position)
    uint32_t s = 0, ph = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      int nt_, tw_i, th_i, tm_;
      tile_coords(tile, a.n_tiles_m, a.g, nt_, tw_i, th_i, tm_);
      const int n0 = nt_ * a.BN;
      const int w0 = tw_i * a.g.TW, h0 = th_i * a.g.TH, b0 = tm_ * a.g.TB;
      int t = 0, c = 0;
      for (int it = 0; it < iters; ++it, ++it_g) {
        mbar_wait(empty0 + 8u * s, ph ^ 1u);
        const int dh = a.taps == 9 ? t / 3 - 1 : 0, dw = a.taps == 9 ? t % 3 - 1 : 0;
        const uint32_t sa = base + s * stage_bytes, sb = sa + a_bytes;
        if (elect_one()) {
          mbar_expect_tx(full0 + 8u * s, stage_bytes);
          tma_load_4d(sa, &tmA, full0 + 8u * s, c * a.kc, w0 + dw, h0 + dh, b0);
          tma_load_2d(sb, &tmB, full0 + 8u * s, t * a.Cin + c * a.kc, n0);
        }
        __syncwarp();
        if (++c == kchunks) { c = 0; ++t; }
        if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ---- MMA issue: uniform control flow, {lo, hi} descriptor words, one elected lane fires the instructions
    const uint32_t idesc = make_idesc(128, a.BN, 0, 0);
    const uint32_t dhi = desc_hi(a.sbo, a.layout);
    const bool k64 = a.kc == 64;
    uint32_t tcount = 0, s = 0, ph = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
      const uint32_t as = tcount & 1u, aph = (tcount >> 1) & 1u;
      mbar_wait(tempty0 + 8u * as, aph ^ 1u);  // epilogue has drained this accumulator stage
      tc_fence_after();
      const uint32_t tacc = tmem_base + as * (uint32_t)a.BN;
      for (int it = 0; it < iters; ++it) {
        mbar_wait(full0 + 8u * s, ph);
        tc_fence_after();
        const uint32_t sa = base + s * stage_bytes;
        const uint32_t alo = desc_lo(sa, 16), blo = desc_lo(sa + a_bytes, 16);
        if (elect_one()) {
          umma_f16_w(tacc, alo, dhi, blo, dhi, idesc, it > 0 ? 1u : 0u);
          umma_f16_w(tacc, alo + 2u, dhi, blo + 2u, dhi, idesc, 1u);
          if (k64) {
            umma_f16_w(tacc, alo + 4u, dhi, blo + 4u, dhi, idesc, 1u);
            umma_f16_w(tacc, alo + 6u, dhi, blo + 6u, dhi, idesc, 1u);
          }
          umma_commit(empty0 + 8u * s);
        }
        __syncwarp();
        if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1u; }
      }
      if (elect_one()) umma_commit(tfull0 + 8u * as);
      __syncwarp();
    }
  } else {
    const int q = warp & 3;          // TMEM lane quarter this warp may read
    const int eg = (warp - 2) >> 2;  // column group: chunks eg, eg+2, ...
    const int ew = eg * 4 + q;       // 0..7
    const int r = q * 32 + lane;
    const int tw = r % a.g.TW, th = (r / a.g.TW) % a.g.TH, tb = r / (a.g.TW * a.g.TH);
    const int et = threadIdx.x - 64;  // 0..255 among the epilogue threads
    const int sub_sh = a.sub_w == 64 ? 6 : 5;
    const int nsub = a.BN >> sub_sh;
    const uint32_t sub_bytes = 128u * (uint32_t)a.sub_w * 2u;
    const uint32_t row_bytes = (uint32_t)a.sub_w * 2u;
    // BN statistics are accumulated per CTA across all its tiles of one n-tile (registers: lane <-> column) and
    // flushed with 2*BN atomics per CTA instead of 8*BN per tile: the per-tile version serialised thousands of
    // same-address atomics in L2 and doubled the kernel time.
    // Column sums come from the staged bf16 tile in shared memory (second version; the first transposed 2 x 32 registers
    // through 62 warp shuffles per chunk and cost 8-11 us on the 128x128-pixel layers): thread = one bf16x2 word column x
    // one group of rows, packed fp32 accumulators that live across all tiles of the CTA's current n-tile.
    f2_t cs1 = 0ull, cs2 = 0ull;
    const int wpr = a.BN >> 1;                 // bf16x2 words per tile row
    const int s_wcol = et % wpr, s_rgrp = et / wpr, s_groups = EPI_THREADS / wpr, s_rows = 128 / s_groups;
    int cur_n0 = -1;
    auto flush_stats = [&](int n0f) {
      // per-thread partials -> smem -> one atomic per column
      float* red = reinterpret_cast<float*>(smem_raw + (stat_smem - smem_u32(smem_raw)));
      epi_bar_sync();
      float x0, x1, q0, q1;
      upk2(cs1, x0, x1);
      upk2(cs2, q0, q1);
      red[(s_rgrp * 2 + 0) * a.BN + 2 * s_wcol] = x0; red[(s_rgrp * 2 + 0) * a.BN + 2 * s_wcol + 1] = x1;
      red[(s_rgrp * 2 + 1) * a.BN + 2 * s_wcol] = q0; red[(s_rgrp * 2 + 1) * a.BN + 2 * s_wcol + 1] = q1;
      cs1 = 0ull; cs2 = 0ull;
      epi_bar_sync();
      for (int i = et; i < 2 * a.BN; i += EPI_THREADS) {
        const int which = i / a.BN, col = i % a.BN;
        float v = 0.f;
        for (int g = 0; g < s_groups; ++g) v += red[(g * 2 + which) * a.BN + col];
        if (!a.dbg_noflush) atomicAdd(a.stats + which * a.N + n0f + col, v);
      }
    };
    // bias vector into shared memory once (ordered by the first epi_bar_sync below): no global-memory latency between
    // the accumulator becoming ready and the tile being stored
    float* sbias = reinterpret_cast<float*>(smem_raw + (stat_smem + 4096u - smem_u32(smem_raw)));
    if (a.bias)
      for (int i = et; i < a.N; i += EPI_THREADS) sbias[i] = __ldg(a.bias + i);
    uint32_t tcount = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
      int nt_, tw_i, th_i, tm_;
      tile_coords(tile, a.n_tiles_m, a.g, nt_, tw_i, th_i, tm_);
      const int n0 = nt_ * a.BN;
      if (a.stats && n0 != cur_n0) {
        if (cur_n0 >= 0) flush_stats(cur_n0);
        cur_n0 = n0;
      }
      const int w0 = tw_i * a.g.TW, h0 = th_i * a.g.TH, b0 = tm_ * a.g.TB;
      const int pw = w0 + tw, ph_ = h0 + th, pb = b0 + tb;
      const bool valid = pw < a.W && ph_ < a.H && pb < a.B;
      const long pix = ((long)pb * a.H + ph_) * a.W + pw;
      const uint32_t as = tcount & 1u, aph = (tcount >> 1) & 1u;
      const uint32_t obuf = out0 + as * out_bytes;
      // the TMA store that last read this staging buffer (two tiles ago) must have finished reading it
      if (et == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      epi_bar_sync();
      mbar_wait(tfull0 + 8u * as, aph);
      tc_fence_after();
      for (int c0 = eg * 32; c0 < a.BN; c0 += 64) {
        uint32_t raw[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + as * (uint32_t)a.BN + (uint32_t)c0, raw);
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(raw[j]);
        if (a.bias) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 bb = *reinterpret_cast<const float4*>(sbias + n0 + c0 + j);
            v[j] += bb.x; v[j + 1] += bb.y; v[j + 2] += bb.z; v[j + 3] += bb.w;
          }
        }
        if (a.add && valid) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            float t8[8];
            ck::unpack8(ck::ldg16(a.add + pix * a.N + n0 + c0 + j), t8);
#pragma unroll
            for (int e = 0; e < 8; ++e) v[j + e] += t8[e];
          }
        }
        if (a.stats && !valid) {  // out-of-image rows are clipped by the TMA store but must not enter the statistics
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
        }
        // bf16 pack + swizzled staging store (row r of sub-tile c0 / sub_w)
        const int sub = c0 >> sub_sh, cin_sub = c0 & (a.sub_w - 1);  // sub_w is 64 or 32: no divide sequence per chunk
        const uint32_t rowaddr = obuf + sub * sub_bytes + (uint32_t)r * row_bytes;
        const uint32_t xorv = (a.sub_w == 64) ? (uint32_t)(r & 7) : (uint32_t)((r >> 1) & 3);
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          const uint4 pk = ck::pack8(&v[j]);
          const uint32_t chunk = (uint32_t)((cin_sub + j) >> 3);
          const uint32_t dst = rowaddr + ((chunk ^ xorv) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(pk.x), "r"(pk.y), "r"(pk.z), "r"(pk.w)
                       : "memory");
        }
      }
      // accumulator stage fully read: hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty0 + 8u * as);
      // make the staging writes visible to the async proxy, then one thread stores the tile
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      epi_bar_sync();
      if (et == 0) {
        for (int sub = 0; sub < nsub; ++sub)
          tma_store_4d(&tmY, obuf + sub * sub_bytes, n0 + sub * a.sub_w, w0, h0, b0);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      if (a.stats) {
        // the staged tile is complete (barrier above) and stays untouched until two tiles from now
        const int col = 2 * s_wcol;
        const int sub = col >> sub_sh, cin_sub = col & (a.sub_w - 1);
        const uint32_t chunk = (uint32_t)(cin_sub >> 3), within = (uint32_t)(cin_sub & 7) * 2u;
        const uint32_t colbase = obuf + sub * sub_bytes + within;
        // rows of a thread start at a multiple of 8 and s_rows (= BN / 4) is a multiple of 8: the swizzle term of row
        // r0 + i only depends on i & 7, so the eight chunk offsets are tile-invariant and the loop body is load + unpack +
        // two packed adds (the first version recomputed the swizzled address per row: ~13 instructions per row, as many
        // instructions per tile as the rest of the epilogue - ncu source page of conv_tc_fwd_v2_k)
        uint32_t raddr = colbase + (uint32_t)(s_rgrp * s_rows) * row_bytes;
        for (int rr = 0; rr < s_rows; rr += 8) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t xv = (a.sub_w == 64) ? (uint32_t)i : (uint32_t)(i >> 1);
            uint32_t w;
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(raddr + (uint32_t)i * row_bytes + ((chunk ^ xv) << 4)));
            const f2_t x = bf2_to_f2(w);
            cs1 = fadd2(cs1, x);
            cs2 = ffma2(x, x, cs2);
          }
          raddr += 8u * row_bytes;
        }
      }
    }
    if (a.stats && cur_n0 >= 0) flush_stats(cur_n0);
    // only the shared-memory reads of the last stores must finish before the CTA retires; the global writes are
    // complete at grid end like any other store
    if (et == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, ncols);
  }
}

// ----------------------------------------------------------------------------------------------------------
// v3 ("halo"): 3x3 convs with Cin % 64 == 0.  v1/v2 reload the shifted 128-pixel box for each of the 9 taps, so the
// L2 -> shared-memory traffic of the activations is 9x the tensor; ncu showed the kernel bound by exactly that.  Here
// ONE (16+2) x (8+2)-pixel halo tile per 64-channel chunk is loaded and the nine taps are nine shared-memory matrix
// descriptors into it (start address shifted by (kh*10 + kw) pixels, 8-row group stride = one halo row), i.e. a real
// implicit GEMM: activation traffic drops from 9x to 1.4x.  Weight tiles stream through their own ring.
struct ConvArgs3 {
  const float* bias;
  const __nv_bfloat16* add;
  float* stats;
  int B, H, W, Cin, N, taps;
  int BN, kc, stages;
  int sub_w;
  int n_tiles_m, n_tiles_n;
  uint32_t layout, sbo;
  int base_off;              // set the descriptor's base_offset field from the start address
  int dbg_noflush;
  int astages;               // halo-tile ring depth (2..4)
  int resident;              // all 9 * Cin/kc weight tiles stay in shared memory for the CTA's lifetime (stages == that count)
  uint32_t a_stage, a_bytes; // halo stage pitch (1024-aligned) / payload bytes (180 rows of kc channels)
  TileGeom g;
};

// "resident" mode (small layers: 9 * Cin * N * 2 bytes <= ~80 KB, one n-tile): the weight tiles are loaded ONCE per CTA
// onto a single barrier; per tile the producer only streams halo tiles and the issuing thread fires its 9 * kc/16 MMAs
// back to back (no per-tap barrier round trip, no re-streaming of the weights from L2 for every 128-pixel tile, which
// for N = 32 was more traffic than the activations).  kc = 32 (Cin = 32) uses the SWIZZLE_64B flavour of the same trick.
__global__ void __launch_bounds__(64 + EPI_THREADS) conv_tc_fwd_v3_k(const __grid_constant__ CUtensorMap tmA,
                                                        const __grid_constant__ CUtensorMap tmB,
                                                        const __grid_constant__ CUtensorMap tmY, const ConvArgs3 a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // A ring: 2 halo tiles of (TH+2) x (TW+2) = 18 x 10 pixels x 64 channels (23040 B, padded to 23552);
  // B ring: a.stages weight tiles [BN][64] (one per (chunk, tap)); then 2 output staging buffers.
  constexpr uint32_t HALO_W = 10u;
  const uint32_t A_STAGE = a.a_stage, A_BYTES = a.a_bytes, row_b = (uint32_t)a.kc * 2u;
  const uint32_t b_bytes = (uint32_t)a.BN * row_b;
  const uint32_t bring = base + (uint32_t)a.astages * A_STAGE;
  const uint32_t out_bytes = 128u * (uint32_t)a.BN * 2u;
  const uint32_t out0 = bring + a.stages * b_bytes;
  const uint32_t bar_base = out0 + 2u * out_bytes;
  const uint32_t fullA0 = bar_base, emptyA0 = bar_base + 32u;
  const uint32_t full0 = bar_base + 64u, empty0 = full0 + 8u * a.stages;
  const uint32_t tfull0 = empty0 + 8u * a.stages, tempty0 = tfull0 + 16u;
  const uint32_t tptr = tempty0 + 16u;
  const uint32_t stat_smem = tptr + 16u;  // EPI_WARPS x 2 x 128 floats
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int kchunks = a.Cin / a.kc;
  const int total_tiles = a.n_tiles_m * a.n_tiles_n;
  const uint32_t ncols = (uint32_t)(2 * a.BN < 32 ? 32 : 2 * a.BN);

  if (threadIdx.x == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(full0 + 8u * s, 1);
      mbar_init(empty0 + 8u * s, 1);
    }
    for (int s = 0; s < a.astages; ++s) {
      mbar_init(fullA0 + 8u * s, 1);
      mbar_init(emptyA0 + 8u * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull0 + 8u * s, 1);
      mbar_init(tempty0 + 8u * s, EPI_WARPS);  // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmY);
  }
  if (warp == 1) tmem_alloc(tptr, ncols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  ck::pdl_enter();  // set-up above overlapped the previous kernel's tail; global memory is touched only from here on
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tptr));

  if (warp == 0) {
    // ---- producer: all lanes walk the schedule (uniform control flow), one elected lane issues
    uint32_t ra_s = 0, ra_ph = 0, rb_s = 0, rb_ph = 0;  // ring positions and phases: halo tiles / weight tiles
    if (a.resident && (int)blockIdx.x < total_tiles) {
      if (elect_one()) {
        mbar_expect_tx(full0, (uint32_t)a.stages * b_bytes);
        for (int c = 0; c < kchunks; ++c)
          for (int t = 0; t < 9; ++t)
            tma_load_2d(bring + (uint32_t)(c * 9 + t) * b_bytes, &tmB, full0, t * a.Cin + c * a.kc, 0);
      }
      __syncwarp();
    }
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      int nt_, tw_i, th_i, tm_;
      tile_coords(tile, a.n_tiles_m, a.g, nt_, tw_i, th_i, tm_);
      const int n0 = nt_ * a.BN;
      const int w0 = tw_i * a.g.TW, h0 = th_i * a.g.TH, b0 = tm_ * a.g.TB;
      for (int c = 0; c < kchunks; ++c) {
        const uint32_t sa = ra_s, pha = ra_ph;  // running ring position (no division on the issue path)
        if (++ra_s == (uint32_t)a.astages) { ra_s = 0; ra_ph ^= 1u; }
        mbar_wait(emptyA0 + 8u * sa, pha ^ 1u);
        if (elect_one()) {
          mbar_expect_tx(fullA0 + 8u * sa, A_BYTES);
          tma_load_4d(base + sa * A_STAGE, &tmA, fullA0 + 8u * sa, c * a.kc, w0 - 1, h0 - 1, b0);  // halo: zero fill = padding
        }
        __syncwarp();
        if (a.resident) continue;
        for (int t = 0; t < 9; ++t) {
          const uint32_t s = rb_s, ph = rb_ph;
          if (++rb_s == (uint32_t)a.stages) { rb_s = 0; rb_ph ^= 1u; }
          mbar_wait(empty0 + 8u * s, ph ^ 1u);
          if (elect_one()) {
            mbar_expect_tx(full0 + 8u * s, b_bytes);
            tma_load_2d(bring + s * b_bytes, &tmB, full0 + 8u * s, t * a.Cin + c * a.kc, n0);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ---- MMA issue: uniform control flow, descriptors in {lo, hi} words, one elected lane fires the instructions
    const uint32_t idesc = make_idesc(128, a.BN, 0, 0);
    const uint32_t ahi = desc_hi(HALO_W * row_b, a.layout), bhi = desc_hi(a.sbo, a.layout);
    const bool k64 = a.kc == 64;
    // Issue path without divisions: ring positions are running counters, the descriptor words of a tap are the words of
    // the stage plus constants (the first version recomputed stage = counter % depth with two divide sequences per TAP:
    // ~95 SASS instructions per four MMAs, which made the single issuing warp - not the tensor pipe, not TMA - the
    // bottleneck of every halo-kernel launch; ncu source page of conv_tc_xf_k, profiles/ncu_xf_issue.md)
    const uint32_t rb16 = row_b >> 4, bstep16 = b_bytes >> 4;
    const uint32_t blo_res0 = desc_lo(bring, 16);
    uint32_t ra_s = 0, ra_ph = 0, rb_s = 0, rb_ph = 0, tcount = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
      const uint32_t as = tcount & 1u, aph = (tcount >> 1) & 1u;
      mbar_wait(tempty0 + 8u * as, aph ^ 1u);
      tc_fence_after();
      const uint32_t tacc = tmem_base + as * (uint32_t)a.BN;
      if (a.resident && tcount == 0) {
        mbar_wait(full0, 0u);  // the resident weight tiles have landed
        tc_fence_after();
      }
      for (int c = 0; c < kchunks; ++c) {
        const uint32_t sa = ra_s, pha = ra_ph;
        if (++ra_s == (uint32_t)a.astages) { ra_s = 0; ra_ph ^= 1u; }
        mbar_wait(fullA0 + 8u * sa, pha);
        tc_fence_after();
        const uint32_t alo_base = desc_lo(base + sa * A_STAGE, 16);
        const uint32_t blo_c = blo_res0 + (uint32_t)(c * 9) * bstep16;
#pragma unroll
        for (int t = 0; t < 9; ++t) {
          uint32_t s = 0;
          if (!a.resident) {
            s = rb_s;
            const uint32_t ph = rb_ph;
            if (++rb_s == (uint32_t)a.stages) { rb_s = 0; rb_ph ^= 1u; }
            mbar_wait(full0 + 8u * s, ph);
            tc_fence_after();
          }
          // tap (kh,kw): output pixel (th,tw) reads halo pixel (th+kh, tw+kw).  One 8-row swizzle group = the 8
          // pixels of one tile row, groups HALO_W * row_b bytes apart: the shifted window is still a canonical
          // K-major swizzled operand, it just starts (kh*HALO_W + kw) rows into the halo tile.
          const uint32_t alo = alo_base + (uint32_t)((t / 3) * HALO_W + (t % 3)) * rb16;
          const uint32_t blo = a.resident ? blo_c + (uint32_t)t * bstep16 : blo_res0 + s * bstep16;
          const uint32_t first = (c > 0 || t > 0) ? 1u : 0u;
          if (elect_one()) {
            umma_f16_w(tacc, alo, ahi, blo, bhi, idesc, first);
            umma_f16_w(tacc, alo + 2u, ahi, blo + 2u, bhi, idesc, 1u);
            if (k64) {
              umma_f16_w(tacc, alo + 4u, ahi, blo + 4u, bhi, idesc, 1u);
              umma_f16_w(tacc, alo + 6u, ahi, blo + 6u, bhi, idesc, 1u);
            }
            if (!a.resident) umma_commit(empty0 + 8u * s);
          }
          __syncwarp();
        }
        if (elect_one()) umma_commit(emptyA0 + 8u * sa);
        __syncwarp();
      }
      if (elect_one()) umma_commit(tfull0 + 8u * as);
      __syncwarp();
    }
  } else {
    const int q = warp & 3;          // TMEM lane quarter this warp may read
    const int eg = (warp - 2) >> 2;  // column group: chunks eg, eg+2, ...
    const int ew = eg * 4 + q;       // 0..7
    const int r = q * 32 + lane;
    const int tw = r % a.g.TW, th = (r / a.g.TW) % a.g.TH, tb = r / (a.g.TW * a.g.TH);
    const int et = threadIdx.x - 64;  // 0..255 among the epilogue threads
    const int sub_sh = a.sub_w == 64 ? 6 : 5;
    const int nsub = a.BN >> sub_sh;
    const uint32_t sub_bytes = 128u * (uint32_t)a.sub_w * 2u;
    const uint32_t row_bytes = (uint32_t)a.sub_w * 2u;
    // BN statistics are accumulated per CTA across all its tiles of one n-tile (registers: lane <-> column) and
    // flushed with 2*BN atomics per CTA instead of 8*BN per tile: the per-tile version serialised thousands of
    // same-address atomics in L2 and doubled the kernel time.
    // Column sums come from the staged bf16 tile in shared memory (second version; the first transposed 2 x 32 registers
    // through 62 warp shuffles per chunk and cost 8-11 us on the 128x128-pixel layers): thread = one bf16x2 word column x
    // one group of rows, packed fp32 accumulators that live across all tiles of the CTA's current n-tile.
    f2_t cs1 = 0ull, cs2 = 0ull;
    const int wpr = a.BN >> 1;                 // bf16x2 words per tile row
    const int s_wcol = et % wpr, s_rgrp = et / wpr, s_groups = EPI_THREADS / wpr, s_rows = 128 / s_groups;
    int cur_n0 = -1;
    auto flush_stats = [&](int n0f) {
      // per-thread partials -> smem -> one atomic per column
      float* red = reinterpret_cast<float*>(smem_raw + (stat_smem - smem_u32(smem_raw)));
      epi_bar_sync();
      float x0, x1, q0, q1;
      upk2(cs1, x0, x1);
      upk2(cs2, q0, q1);
      red[(s_rgrp * 2 + 0) * a.BN + 2 * s_wcol] = x0; red[(s_rgrp * 2 + 0) * a.BN + 2 * s_wcol + 1] = x1;
      red[(s_rgrp * 2 + 1) * a.BN + 2 * s_wcol] = q0; red[(s_rgrp * 2 + 1) * a.BN + 2 * s_wcol + 1] = q1;
      cs1 = 0ull; cs2 = 0ull;
      epi_bar_sync();
      for (int i = et; i < 2 * a.BN; i += EPI_THREADS) {
        const int which = i / a.BN, col = i % a.BN;
        float v = 0.f;
        for (int g = 0; g < s_groups; ++g) v += red[(g * 2 + which) * a.BN + col];
        if (!a.dbg_noflush) atomicAdd(a.stats + which * a.N + n0f + col, v);
      }
    };
    // bias vector into shared memory once (ordered by the first epi_bar_sync below): no global-memory latency between
    // the accumulator becoming ready and the tile being stored
    float* sbias = reinterpret_cast<float*>(smem_raw + (stat_smem + 4096u - smem_u32(smem_raw)));
    if (a.bias)
      for (int i = et; i < a.N; i += EPI_THREADS) sbias[i] = __ldg(a.bias + i);
    uint32_t tcount = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
      int nt_, tw_i, th_i, tm_;
      tile_coords(tile, a.n_tiles_m, a.g, nt_, tw_i, th_i, tm_);
      const int n0 = nt_ * a.BN;
      if (a.stats && n0 != cur_n0) {
        if (cur_n0 >= 0) flush_stats(cur_n0);
        cur_n0 = n0;
      }
      const int w0 = tw_i * a.g.TW, h0 = th_i * a.g.TH, b0 = tm_ * a.g.TB;
      const int pw = w0 + tw, ph_ = h0 + th, pb = b0 + tb;
      const bool valid = pw < a.W && ph_ < a.H && pb < a.B;
      const long pix = ((long)pb * a.H + ph_) * a.W + pw;
      const uint32_t as = tcount & 1u, aph = (tcount >> 1) & 1u;
      const uint32_t obuf = out0 + as * out_bytes;
      // the TMA store that last read this staging buffer (two tiles ago) must have finished reading it
      if (et == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      epi_bar_sync();
      mbar_wait(tfull0 + 8u * as, aph);
      tc_fence_after();
      for (int c0 = eg * 32; c0 < a.BN; c0 += 64) {
        uint32_t raw[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + as * (uint32_t)a.BN + (uint32_t)c0, raw);
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(raw[j]);
        if (a.bias) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 bb = *reinterpret_cast<const float4*>(sbias + n0 + c0 + j);
            v[j] += bb.x; v[j + 1] += bb.y; v[j + 2] += bb.z; v[j + 3] += bb.w;
          }
        }
        if (a.add && valid) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            float t8[8];
            ck::unpack8(ck::ldg16(a.add + pix * a.N + n0 + c0 + j), t8);
#pragma unroll
            for (int e = 0; e < 8; ++e) v[j + e] += t8[e];
          }
        }
        if (a.stats && !valid) {  // out-of-image rows are clipped by the TMA store but must not enter the statistics
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
        }
        // bf16 pack + swizzled staging store (row r of sub-tile c0 / sub_w)
        const int sub = c0 >> sub_sh, cin_sub = c0 & (a.sub_w - 1);  // sub_w is 64 or 32: no divide sequence per chunk
        const uint32_t rowaddr = obuf + sub * sub_bytes + (uint32_t)r * row_bytes;
        const uint32_t xorv = (a.sub_w == 64) ? (uint32_t)(r & 7) : (uint32_t)((r >> 1) & 3);
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          const uint4 pk = ck::pack8(&v[j]);
          const uint32_t chunk = (uint32_t)((cin_sub + j) >> 3);
          const uint32_t dst = rowaddr + ((chunk ^ xorv) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(pk.x), "r"(pk.y), "r"(pk.z), "r"(pk.w)
                       : "memory");
        }
      }
      // accumulator stage fully read: hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty0 + 8u * as);
      // make the staging writes visible to the async proxy, then one thread stores the tile
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      epi_bar_sync();
      if (et == 0) {
        for (int sub = 0; sub < nsub; ++sub)
          tma_store_4d(&tmY, obuf + sub * sub_bytes, n0 + sub * a.sub_w, w0, h0, b0);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      if (a.stats) {
        // the staged tile is complete (barrier above) and stays untouched until two tiles from now
        const int col = 2 * s_wcol;
        const int sub = col >> sub_sh, cin_sub = col & (a.sub_w - 1);
        const uint32_t chunk = (uint32_t)(cin_sub >> 3), within = (uint32_t)(cin_sub & 7) * 2u;
        const uint32_t colbase = obuf + sub * sub_bytes + within;
        // rows of a thread start at a multiple of 8 and s_rows (= BN / 4) is a multiple of 8: the swizzle term of row
        // r0 + i only depends on i & 7, so the eight chunk offsets are tile-invariant and the loop body is load + unpack +
        // two packed adds (the first version recomputed the swizzled address per row: ~13 instructions per row, as many
        // instructions per tile as the rest of the epilogue - ncu source page of conv_tc_fwd_v2_k)
        uint32_t raddr = colbase + (uint32_t)(s_rgrp * s_rows) * row_bytes;
        for (int rr = 0; rr < s_rows; rr += 8) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t xv = (a.sub_w == 64) ? (uint32_t)i : (uint32_t)(i >> 1);
            uint32_t w;
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(raddr + (uint32_t)i * row_bytes + ((chunk ^ xv) << 4)));
            const f2_t x = bf2_to_f2(w);
            cs1 = fadd2(cs1, x);
            cs2 = ffma2(x, x, cs2);
          }
          raddr += 8u * row_bytes;
        }
      }
    }
    if (a.stats && cur_n0 >= 0) flush_stats(cur_n0);
    // only the shared-memory reads of the last stores must finish before the CTA retires; the global writes are
    // complete at grid end like any other store
    if (et == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, ncols);
  }
}

// ----------------------------------------------------------------------------------------------------------
struct ConvArgsX {
  const float* bias;
  const __nv_bfloat16* add;
  float* stats;
  int B, H, W, Cin, N, taps;
  int BN, kc, stages;
  int sub_w;
  int n_tiles_m, n_tiles_n;
  uint32_t layout, sbo;
  int base_off;              // set the descriptor's base_offset field from the start address
  int dbg_noflush;
  int astages;               // halo-tile ring depth (2..4)
  int resident;              // all 9 * Cin/kc weight tiles stay in shared memory for the CTA's lifetime (stages == that count)
  uint32_t a_stage, a_bytes; // halo stage pitch (1024-aligned) / payload bytes (180 rows of kc channels)
  TileGeom g;
  // transform stage
  int xmode;                 // 1: relu(x), 2: relu(scale[c] * x + shift[c]) with zero outside the image
  int diag;                  // depthwise: n-tile == channel chunk, block-diagonal weight tiles, one chunk per CTA
  const float* bn;           // [scale C | shift C] (xmode 2, fin.stats == NULL)
  crackfl_bn_fin fin;        // or: derive them from the statistics here and let block 0 publish them
  __nv_bfloat16* side;       // optional: the transformed activation is also written out (B,H,W,Cin), by the n-tile-0 CTAs
  // xmode 4 / 5 (K2): the input is BN-backward-applied on the fly, dy = A*dz + B*y + C per channel (5: dz masked by
  // bn(y) > 0 first); dz comes through tmA, y through tmA2 into the second half of every halo stage
  const float* sums;         // [sum dz C | sum dz*(y-mean) C] of the BN-backward reduce pass
  float count;               // pixels per channel
  float* dgamma;             // block 0 writes the BN parameter gradients, as bn_bwd_apply does
  float* dbeta;
  uint32_t a_pitch;          // bytes per halo ring slot: a_stage, or 2 * a_stage with two inputs
};

// "xf" variant of the halo kernel: four extra warps TRANSFORM every halo tile in shared memory between the TMA load and
// the MMAs - ReLU, or BatchNorm-apply + ReLU with the scale/shift derived from the statistics (K1: the
// ReLU -> Conv2DTranspose -> BN chain of client_fit_model.py:128-134 without a separate BN-apply pass) - re-zeroing the
// pixels that lie outside the image ("same" padding pads the ACTIVATION with zeros, and relu(shift) is not zero).
// With `diag` the weight tiles are the 64x64 (32x32) block-diagonal expansion of a depthwise 3x3 kernel, so the depthwise
// half of a SeparableConv2D (client_fit_model.py:109,113) runs on the tensor cores: 36 MMAs of 128x64x16 per 128-pixel
// tile (~0.6 us) against ~0.9 us of HBM time for the tile, instead of an issue-bound CUDA-core stencil.
// One CTA per SM (576 threads), everything else as in conv_tc_fwd_v3_k below.
//
// "resident" mode (small layers: 9 * Cin * N * 2 bytes <= ~80 KB, one n-tile): the weight tiles are loaded ONCE per CTA
// onto a single barrier; per tile the producer only streams halo tiles and the issuing thread fires its 9 * kc/16 MMAs
// back to back (no per-tap barrier round trip, no re-streaming of the weights from L2 for every 128-pixel tile, which
// for N = 32 was more traffic than the activations).  kc = 32 (Cin = 32) uses the SWIZZLE_64B flavour of the same trick.
// Two instantiations: one input (ReLU / BN-apply prologue: 8 transform warps, 576 threads) and two inputs (BN-backward
// prologue: 6 transform warps, 512 threads and 128 registers each - with 8 warps the coefficient registers spilled).
template <bool TWO_IN>
__global__ void __launch_bounds__(64 + EPI_THREADS + (TWO_IN ? 6 : 8) * 32, 1) conv_tc_xf_k(const __grid_constant__ CUtensorMap tmA,
                                                        const __grid_constant__ CUtensorMap tmB,
                                                        const __grid_constant__ CUtensorMap tmY,
                                                        const __grid_constant__ CUtensorMap tmA2, const ConvArgsX a) {
  constexpr int XF_WARPS = TWO_IN ? 6 : 8;
  constexpr int XF_THREADS = XF_WARPS * 32;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // A ring: 2 halo tiles of (TH+2) x (TW+2) = 18 x 10 pixels x 64 channels (23040 B, padded to 23552);
  // B ring: a.stages weight tiles [BN][64] (one per (chunk, tap)); then 2 output staging buffers.
  constexpr uint32_t HALO_W = 10u;
  const uint32_t A_STAGE = a.a_pitch, A_BYTES = a.a_bytes, row_b = (uint32_t)a.kc * 2u;
  constexpr bool two_in = TWO_IN;
  const uint32_t b_bytes = (uint32_t)a.BN * row_b;
  const uint32_t bring = base + (uint32_t)a.astages * A_STAGE;
  const uint32_t out_bytes = 128u * (uint32_t)a.BN * 2u;
  const uint32_t out0 = bring + a.stages * b_bytes;
  const uint32_t bar_base = out0 + 2u * out_bytes;
  const uint32_t fullA0 = bar_base, emptyA0 = bar_base + 32u, xfA0 = bar_base + 64u;
  const uint32_t full0 = bar_base + 96u, empty0 = full0 + 8u * a.stages;
  const uint32_t tfull0 = empty0 + 8u * a.stages, tempty0 = tfull0 + 16u;
  const uint32_t tptr = tempty0 + 16u;
  const uint32_t stat_smem = tptr + 16u;  // EPI_WARPS x 2 x 128 floats
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int kchunks = a.diag ? 1 : a.Cin / a.kc;
  const int nch = a.Cin / a.kc;
  // schedule: tile k of this CTA.  Plain: tile = blockIdx.x + k * gridDim.x over (m fastest, n).  diag: the CTA keeps ONE
  // channel chunk (its weight tiles stay resident) and walks the m-tiles with stride gridDim.x / nch.
  const int my_chunk = a.diag ? (int)(blockIdx.x % nch) : 0;
  const int t_first = a.diag ? (int)(blockIdx.x / nch) : (int)blockIdx.x;
  const int t_step = a.diag ? (int)(gridDim.x / nch) : (int)gridDim.x;
  const int total_tiles = a.diag ? a.n_tiles_m : a.n_tiles_m * a.n_tiles_n;
  const uint32_t ncols = (uint32_t)(2 * a.BN < 32 ? 32 : 2 * a.BN);

  if (threadIdx.x == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(full0 + 8u * s, 1);
      mbar_init(empty0 + 8u * s, 1);
    }
    for (int s = 0; s < a.astages; ++s) {
      mbar_init(fullA0 + 8u * s, 1);
      mbar_init(emptyA0 + 8u * s, 1);
      mbar_init(xfA0 + 8u * s, TWO_IN ? 6 : 8);  // one arrival per transform warp
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull0 + 8u * s, 1);
      mbar_init(tempty0 + 8u * s, EPI_WARPS);  // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmY);
  }
  if (warp == 1) tmem_alloc(tptr, ncols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  ck::pdl_enter();  // set-up above overlapped the previous kernel's tail; global memory is touched only from here on
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tptr));
  // BN scale / shift of the INPUT channels (xmode 2): from the published table, or straight from the statistics of the
  // producing conv (finalize folded in, as in the *_fin elementwise kernels; block 0 publishes table + moving statistics)
  float* xf_tab = reinterpret_cast<float*>(smem_raw + (stat_smem + 5120u - smem_u32(smem_raw)));  // [scale Cin | shift Cin]
  if (a.xmode >= 4) {
    // BN backward: dy = sc*is... per channel:  m1 = mean(dz), m2 = is * mean(dz * (y - mu));  dy = sc * (dz - m1 - xhat * m2)
    //            = A * dz + B * y + C  with A = sc, B = -sc * is * m2, C = -sc * m1 - B * mu       (bn = [sc | sh | mu | is])
    const float invn = 1.0f / a.count;
    for (int c = threadIdx.x; c < a.Cin; c += blockDim.x) {
      const float sc = a.bn[c], sh = a.bn[a.Cin + c], mu = a.bn[2 * a.Cin + c], is = a.bn[3 * a.Cin + c];
      const float s1 = a.sums[c], s2 = a.sums[a.Cin + c];
      const float m1 = s1 * invn, m2 = is * s2 * invn;
      const float Bc = -sc * is * m2;
      xf_tab[c] = sc; xf_tab[a.Cin + c] = Bc; xf_tab[2 * a.Cin + c] = -sc * m1 - Bc * mu;
      xf_tab[3 * a.Cin + c] = sc; xf_tab[4 * a.Cin + c] = sh;  // forward BN (mask of xmode 5)
      if (blockIdx.x == 0 && a.dgamma) { a.dbeta[c] = s1; a.dgamma[c] = is * s2; }
    }
    __syncthreads();
  }
  if (a.xmode == 2) {
    for (int c = threadIdx.x; c < a.Cin; c += blockDim.x) {
      float scl, sft;
      if (a.fin.stats) {
        const float mean = a.fin.stats[c] / a.fin.count;
        const float var = fmaxf(a.fin.stats[a.Cin + c] / a.fin.count - mean * mean, 0.f);
        const float invstd = rsqrtf(var + a.fin.eps);
        scl = a.fin.gamma[c] * invstd;
        sft = a.fin.beta[c] - mean * scl;
        if (blockIdx.x == 0) {
          a.fin.bn_out[c] = scl; a.fin.bn_out[a.Cin + c] = sft; a.fin.bn_out[2 * a.Cin + c] = mean; a.fin.bn_out[3 * a.Cin + c] = invstd;
          const float unb = a.fin.count > 1.f ? var * (a.fin.count / (a.fin.count - 1.f)) : var;
          a.fin.run_mean[c] = a.fin.run_mean[c] * a.fin.momentum + mean * (1.f - a.fin.momentum);
          a.fin.run_var[c] = a.fin.run_var[c] * a.fin.momentum + unb * (1.f - a.fin.momentum);
        }
      } else {
        scl = a.bn[c]; sft = a.bn[a.Cin + c];
      }
      xf_tab[c] = scl; xf_tab[a.Cin + c] = sft;
    }
    __syncthreads();
  }

  if (warp == 0) {
    // ---- producer: all lanes walk the schedule (uniform control flow), one elected lane issues
    uint32_t ra_s = 0, ra_ph = 0, rb_s = 0, rb_ph = 0;  // ring positions and phases: halo tiles / weight tiles
    if (a.resident && t_first < total_tiles) {
      if (elect_one()) {
        mbar_expect_tx(full0, (uint32_t)a.stages * b_bytes);
        for (int c = 0; c < kchunks; ++c)
          for (int t = 0; t < 9; ++t)
            tma_load_2d(bring + (uint32_t)(c * 9 + t) * b_bytes, &tmB, full0, a.diag ? 0 : t * a.Cin + c * a.kc,
                        a.diag ? (my_chunk * 9 + t) * a.kc : 0);
      }
      __syncwarp();
    }
    for (int tile = t_first; tile < total_tiles; tile += t_step) {
      int nt_, tw_i, th_i, tm_;
      tile_coords(tile, a.n_tiles_m, a.g, nt_, tw_i, th_i, tm_);
      const int n0 = a.diag ? my_chunk * a.kc : nt_ * a.BN;
      const int w0 = tw_i * a.g.TW, h0 = th_i * a.g.TH, b0 = tm_ * a.g.TB;
      for (int c = 0; c < kchunks; ++c) {
        const uint32_t sa = ra_s, pha = ra_ph;  // running ring position (no division on the issue path)
        if (++ra_s == (uint32_t)a.astages) { ra_s = 0; ra_ph ^= 1u; }
        mbar_wait(emptyA0 + 8u * sa, pha ^ 1u);
        if (elect_one()) {
          mbar_expect_tx(fullA0 + 8u * sa, two_in ? 2u * A_BYTES : A_BYTES);
          tma_load_4d(base + sa * A_STAGE, &tmA, fullA0 + 8u * sa, (a.diag ? my_chunk : c) * a.kc, w0 - 1, h0 - 1, b0);  // halo: zero fill = padding
          if (two_in) tma_load_4d(base + sa * A_STAGE + a.a_stage, &tmA2, fullA0 + 8u * sa, c * a.kc, w0 - 1, h0 - 1, b0);
        }
        __syncwarp();
        if (a.resident) continue;
        for (int t = 0; t < 9; ++t) {
          const uint32_t s = rb_s, ph = rb_ph;
          if (++rb_s == (uint32_t)a.stages) { rb_s = 0; rb_ph ^= 1u; }
          mbar_wait(empty0 + 8u * s, ph ^ 1u);
          if (elect_one()) {
            mbar_expect_tx(full0 + 8u * s, b_bytes);
            tma_load_2d(bring + s * b_bytes, &tmB, full0 + 8u * s, t * a.Cin + c * a.kc, n0);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ---- MMA issue: uniform control flow, descriptors in {lo, hi} words, one elected lane fires the instructions
    const uint32_t idesc = make_idesc(128, a.BN, 0, 0);
    const uint32_t ahi = desc_hi(HALO_W * row_b, a.layout), bhi = desc_hi(a.sbo, a.layout);
    const bool k64 = a.kc == 64;
    // Issue path without divisions: ring positions are running counters, the descriptor words of a tap are the words of
    // the stage plus constants (the first version recomputed stage = counter % depth with two divide sequences per TAP:
    // ~95 SASS instructions per four MMAs, which made the single issuing warp - not the tensor pipe, not TMA - the
    // bottleneck of every halo-kernel launch; ncu source page of conv_tc_xf_k, profiles/ncu_xf_issue.md)
    const uint32_t rb16 = row_b >> 4, bstep16 = b_bytes >> 4;
    const uint32_t blo_res0 = desc_lo(bring, 16);
    uint32_t ra_s = 0, ra_ph = 0, rb_s = 0, rb_ph = 0, tcount = 0;
    for (int tile = t_first; tile < total_tiles; tile += t_step, ++tcount) {
      const uint32_t as = tcount & 1u, aph = (tcount >> 1) & 1u;
      mbar_wait(tempty0 + 8u * as, aph ^ 1u);
      tc_fence_after();
      const uint32_t tacc = tmem_base + as * (uint32_t)a.BN;
      if (a.resident && tcount == 0) {
        mbar_wait(full0, 0u);  // the resident weight tiles have landed
        tc_fence_after();
      }
      for (int c = 0; c < kchunks; ++c) {
        const uint32_t sa = ra_s, pha = ra_ph;
        if (++ra_s == (uint32_t)a.astages) { ra_s = 0; ra_ph ^= 1u; }
        mbar_wait(xfA0 + 8u * sa, pha);  // the transform warps have rewritten the tile
        tc_fence_after();
        const uint32_t alo_base = desc_lo(base + sa * A_STAGE, 16);
        const uint32_t blo_c = blo_res0 + (uint32_t)(c * 9) * bstep16;
#pragma unroll
        for (int t = 0; t < 9; ++t) {
          uint32_t s = 0;
          if (!a.resident) {
            s = rb_s;
            const uint32_t ph = rb_ph;
            if (++rb_s == (uint32_t)a.stages) { rb_s = 0; rb_ph ^= 1u; }
            mbar_wait(full0 + 8u * s, ph);
            tc_fence_after();
          }
          // tap (kh,kw): output pixel (th,tw) reads halo pixel (th+kh, tw+kw).  One 8-row swizzle group = the 8
          // pixels of one tile row, groups HALO_W * row_b bytes apart: the shifted window is still a canonical
          // K-major swizzled operand, it just starts (kh*HALO_W + kw) rows into the halo tile.
          const uint32_t alo = alo_base + (uint32_t)((t / 3) * HALO_W + (t % 3)) * rb16;
          const uint32_t blo = a.resident ? blo_c + (uint32_t)t * bstep16 : blo_res0 + s * bstep16;
          const uint32_t first = (c > 0 || t > 0) ? 1u : 0u;
          if (elect_one()) {
            umma_f16_w(tacc, alo, ahi, blo, bhi, idesc, first);
            umma_f16_w(tacc, alo + 2u, ahi, blo + 2u, bhi, idesc, 1u);
            if (k64) {
              umma_f16_w(tacc, alo + 4u, ahi, blo + 4u, bhi, idesc, 1u);
              umma_f16_w(tacc, alo + 6u, ahi, blo + 6u, bhi, idesc, 1u);
            }
            if (!a.resident) umma_commit(empty0 + 8u * s);
          }
          __syncwarp();
        }
        if (elect_one()) umma_commit(emptyA0 + 8u * sa);
        __syncwarp();
      }
      if (elect_one()) umma_commit(tfull0 + 8u * as);
      __syncwarp();
    }
  } else if (warp >= 2 + EPI_WARPS) {
    // ---- transform warps: halo tile in shared memory -> activation, in place, between TMA and MMA.
    // Thread = one 16-byte chunk COLUMN of the tile (j) and every (XF_THREADS / cpr)-th pixel row: the rows a thread
    // visits are congruent modulo the swizzle period, so the logical channel group behind its physical chunk slot - and
    // with it the 8 scale / shift values - is fixed for the whole kernel and lives in registers.  (First version: 128
    // threads, chunk index and table look-up per element; one latency-bound warp per scheduler made the stage 4 us per
    // tile and the kernel 1.6x slower than the CUDA-core stencil it replaces.)
    const int xt = threadIdx.x - (64 + EPI_THREADS);  // 0..XF_THREADS-1
    const int lane_ = threadIdx.x & 31;
    const int cpr = a.kc >> 3;                        // 16-byte chunks per halo pixel (8 or 4)
    const int j = xt & (cpr - 1), prow = xt / cpr;    // chunk column, first row
    const int rstep = XF_THREADS / cpr;               // 24 / 32 (kc 64) or 48 / 64 (kc 32) rows per pass: multiples of the swizzle period
    // SWIZZLE_128B: slot j of row p holds chunk j ^ (p & 7); SWIZZLE_64B (64-byte rows): j ^ ((p >> 1) & 3)
    const int lch = a.kc == 64 ? (j ^ (prow & 7)) : (j ^ ((prow >> 1) & 3));
    const uint32_t toff = (uint32_t)prow * row_b + (uint32_t)j * 16u;
    // per-thread row table, fixed for the whole kernel: halo row / column of the k-th row this thread visits, whether it
    // exists (p < 180) and whether it is an interior (= output) pixel; the tile loop only adds the tile origin
    constexpr int XF_IT = TWO_IN ? 8 : 6;  // ceil(180 / rows per pass): 24 rows with 192 threads, 32 with 256 (kc 64)
    uint32_t live = 0, inter = 0;
#pragma unroll
    for (int k = 0; k < XF_IT; ++k) {
      const int p = prow + k * rstep;
      const int hr = (p * 205) >> 11, hc = p - hr * 10;  // p / 10, exact for p < 1024 (kept out of registers: recomputed below)
      if (p < 180) live |= 1u << k;
      if (p < 180 && hr >= 1 && hr <= a.g.TH && hc >= 1 && hc <= a.g.TW) inter |= 1u << k;
    }
    const uint32_t kstride = (uint32_t)rstep * row_b;
    const long pix_c = (long)a.Cin;
    uint32_t ra_s = 0, ra_ph = 0;
    for (int tile = t_first; tile < total_tiles; tile += t_step) {
      int nt_, tw_i, th_i, tm_;
      tile_coords(tile, a.n_tiles_m, a.g, nt_, tw_i, th_i, tm_);
      const int n0 = a.diag ? my_chunk * a.kc : nt_ * a.BN;
      const int w0 = tw_i * a.g.TW - 1, h0 = th_i * a.g.TH - 1, b0 = tm_ * a.g.TB;  // origin of the HALO tile
      const bool do_side = a.side != nullptr && (a.diag || n0 == 0);
      // a tile whose halo lies completely inside the image needs no per-pixel test (all but the border tiles)
      const bool all_in = h0 >= 0 && h0 + 18 <= a.H && w0 >= 0 && w0 + 10 <= a.W;
      for (int c = 0; c < kchunks; ++c) {
        const uint32_t sa = ra_s, pha = ra_ph;
        if (++ra_s == (uint32_t)a.astages) { ra_s = 0; ra_ph ^= 1u; }
        const int cbase = (a.diag ? my_chunk : c) * a.kc + lch * 8;
        f2_t sc2[4], sh2[4], cb2[4], cc2[4];
        if constexpr (TWO_IN) {
          const float* tA = xf_tab + cbase;
          const int Cn = a.Cin;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            sc2[e] = pk2(tA[2 * e], tA[2 * e + 1]);                         // A
            cb2[e] = pk2(tA[Cn + 2 * e], tA[Cn + 2 * e + 1]);               // B
            cc2[e] = pk2(tA[2 * Cn + 2 * e], tA[2 * Cn + 2 * e + 1]);       // C
            sh2[e] = 0ull;
          }
        }
        if (a.xmode == 2) {
          const float4 s0 = *reinterpret_cast<const float4*>(xf_tab + cbase), s1 = *reinterpret_cast<const float4*>(xf_tab + cbase + 4);
          const float4 t0 = *reinterpret_cast<const float4*>(xf_tab + a.Cin + cbase);
          const float4 t1 = *reinterpret_cast<const float4*>(xf_tab + a.Cin + cbase + 4);
          sc2[0] = pk2(s0.x, s0.y); sc2[1] = pk2(s0.z, s0.w); sc2[2] = pk2(s1.x, s1.y); sc2[3] = pk2(s1.z, s1.w);
          sh2[0] = pk2(t0.x, t0.y); sh2[1] = pk2(t0.z, t0.w); sh2[2] = pk2(t1.x, t1.y); sh2[3] = pk2(t1.z, t1.w);
        }
        __nv_bfloat16* sbase = a.side + (((long)b0 * a.H + h0) * a.W + w0) * pix_c + cbase;  // halo pixel (0,0), this channel group
        mbar_wait(fullA0 + 8u * sa, pha);
        if (a.xmode == 3) {
          __syncwarp();
          if (lane_ == 0) mbar_arrive(xfA0 + 8u * sa);
          continue;
        }
        const uint32_t tbase = base + sa * A_STAGE + toff;
        // four rows at a time: their loads are issued together, then transformed and stored (all eight at once spilled)
#pragma unroll
        constexpr int XF_B = TWO_IN ? 4 : XF_IT;
#pragma unroll
        for (int kb = 0; kb < XF_IT; kb += XF_B) {
        uint32_t v[XF_IT][4];
#pragma unroll
        for (int k = kb; k < kb + XF_B; ++k)
          if (live >> k & 1u)
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v[k][0]), "=r"(v[k][1]), "=r"(v[k][2]), "=r"(v[k][3])
                         : "r"(tbase + (uint32_t)k * kstride));
#pragma unroll
        for (int k = kb; k < kb + XF_B; ++k) {
          if (!(live >> k & 1u)) continue;
          const int p_ = prow + k * rstep;
          const int hr_ = (p_ * 205) >> 11, hc_ = p_ - hr_ * 10;
          bool inside = true;
          if (!all_in) {
            const int ih = h0 + hr_, iw = w0 + hc_;
            inside = ih >= 0 && ih < a.H && iw >= 0 && iw < a.W;
          }
          if constexpr (TWO_IN) {
            uint32_t y4[4];
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(y4[0]), "=r"(y4[1]), "=r"(y4[2]), "=r"(y4[3])
                         : "r"(tbase + a.a_stage + (uint32_t)k * kstride));
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const f2_t yv = bf2_to_f2(y4[e]);
              f2_t dzv = bf2_to_f2(v[k][e]);
              if (a.xmode == 5) {  // dz passes only where the forward activation was positive
                float z0, z1, d0, d1;
                // (forward scale / shift stay in shared memory: keeping them in registers too made the kernel spill)
                const f2_t ms = *reinterpret_cast<const f2_t*>(xf_tab + 3 * a.Cin + cbase + 2 * e);
                const f2_t mt = *reinterpret_cast<const f2_t*>(xf_tab + 4 * a.Cin + cbase + 2 * e);
                upk2(ffma2(yv, ms, mt), z0, z1);
                upk2(dzv, d0, d1);
                dzv = pk2(z0 > 0.f ? d0 : 0.f, z1 > 0.f ? d1 : 0.f);
              }
              const f2_t o = ffma2(sc2[e], dzv, ffma2(cb2[e], yv, cc2[e]));
              float o0, o1;
              upk2(o, o0, o1);
              v[k][e] = inside ? ck::pack2(o0, o1) : 0u;  // zero outside the image: the padding of dy, not C
            }
          } else if (a.xmode == 1) {
#pragma unroll
            for (int e = 0; e < 4; ++e) v[k][e] = relu_bf2(v[k][e]);  // zero fill stays zero
          } else if (inside) {
#pragma unroll
            for (int e = 0; e < 4; ++e) v[k][e] = f2_to_bf2_relu(ffma2(bf2_to_f2(v[k][e]), sc2[e], sh2[e]));
          } else {
#pragma unroll
            for (int e = 0; e < 4; ++e) v[k][e] = 0u;  // "same" padding pads the activation: relu(shift) must not leak in
          }
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(tbase + (uint32_t)k * kstride), "r"(v[k][0]), "r"(v[k][1]),
                       "r"(v[k][2]), "r"(v[k][3]) : "memory");
          if (do_side && inside && (inter >> k & 1u)) {
            uint4 o; o.x = v[k][0]; o.y = v[k][1]; o.z = v[k][2]; o.w = v[k][3];
            ck::stg16(sbase + ((long)hr_ * a.W + hc_) * pix_c, o);
          }
        }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the MMA's reads
        __syncwarp();
        if (lane_ == 0) mbar_arrive(xfA0 + 8u * sa);
      }
    }
  } else {
    const int q = warp & 3;          // TMEM lane quarter this warp may read
    const int eg = (warp - 2) >> 2;  // column group: chunks eg, eg+2, ...
    const int ew = eg * 4 + q;       // 0..7
    const int r = q * 32 + lane;
    const int tw = r % a.g.TW, th = (r / a.g.TW) % a.g.TH, tb = r / (a.g.TW * a.g.TH);
    const int et = threadIdx.x - 64;  // 0..255 among the epilogue threads
    const int sub_sh = a.sub_w == 64 ? 6 : 5;
    const int nsub = a.BN >> sub_sh;
    const uint32_t sub_bytes = 128u * (uint32_t)a.sub_w * 2u;
    const uint32_t row_bytes = (uint32_t)a.sub_w * 2u;
    // BN statistics are accumulated per CTA across all its tiles of one n-tile (registers: lane <-> column) and
    // flushed with 2*BN atomics per CTA instead of 8*BN per tile: the per-tile version serialised thousands of
    // same-address atomics in L2 and doubled the kernel time.
    // Column sums come from the staged bf16 tile in shared memory (second version; the first transposed 2 x 32 registers
    // through 62 warp shuffles per chunk and cost 8-11 us on the 128x128-pixel layers): thread = one bf16x2 word column x
    // one group of rows, packed fp32 accumulators that live across all tiles of the CTA's current n-tile.
    f2_t cs1 = 0ull, cs2 = 0ull;
    const int wpr = a.BN >> 1;                 // bf16x2 words per tile row
    const int s_wcol = et % wpr, s_rgrp = et / wpr, s_groups = EPI_THREADS / wpr, s_rows = 128 / s_groups;
    int cur_n0 = -1;
    auto flush_stats = [&](int n0f) {
      // per-thread partials -> smem -> one atomic per column
      float* red = reinterpret_cast<float*>(smem_raw + (stat_smem - smem_u32(smem_raw)));
      epi_bar_sync();
      float x0, x1, q0, q1;
      upk2(cs1, x0, x1);
      upk2(cs2, q0, q1);
      red[(s_rgrp * 2 + 0) * a.BN + 2 * s_wcol] = x0; red[(s_rgrp * 2 + 0) * a.BN + 2 * s_wcol + 1] = x1;
      red[(s_rgrp * 2 + 1) * a.BN + 2 * s_wcol] = q0; red[(s_rgrp * 2 + 1) * a.BN + 2 * s_wcol + 1] = q1;
      cs1 = 0ull; cs2 = 0ull;
      epi_bar_sync();
      for (int i = et; i < 2 * a.BN; i += EPI_THREADS) {
        const int which = i / a.BN, col = i % a.BN;
        float v = 0.f;
        for (int g = 0; g < s_groups; ++g) v += red[(g * 2 + which) * a.BN + col];
        if (!a.dbg_noflush) atomicAdd(a.stats + which * a.N + n0f + col, v);
      }
    };
    // bias vector into shared memory once (ordered by the first epi_bar_sync below): no global-memory latency between
    // the accumulator becoming ready and the tile being stored
    float* sbias = reinterpret_cast<float*>(smem_raw + (stat_smem + 4096u - smem_u32(smem_raw)));
    if (a.bias)
      for (int i = et; i < a.N; i += EPI_THREADS) sbias[i] = __ldg(a.bias + i);
    uint32_t tcount = 0;
    for (int tile = t_first; tile < total_tiles; tile += t_step, ++tcount) {
      int nt_, tw_i, th_i, tm_;
      tile_coords(tile, a.n_tiles_m, a.g, nt_, tw_i, th_i, tm_);
      const int n0 = a.diag ? my_chunk * a.kc : nt_ * a.BN;
      if (a.stats && n0 != cur_n0) {
        if (cur_n0 >= 0) flush_stats(cur_n0);
        cur_n0 = n0;
      }
      const int w0 = tw_i * a.g.TW, h0 = th_i * a.g.TH, b0 = tm_ * a.g.TB;
      const int pw = w0 + tw, ph_ = h0 + th, pb = b0 + tb;
      const bool valid = pw < a.W && ph_ < a.H && pb < a.B;
      const long pix = ((long)pb * a.H + ph_) * a.W + pw;
      const uint32_t as = tcount & 1u, aph = (tcount >> 1) & 1u;
      const uint32_t obuf = out0 + as * out_bytes;
      // the TMA store that last read this staging buffer (two tiles ago) must have finished reading it
      if (et == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      epi_bar_sync();
      mbar_wait(tfull0 + 8u * as, aph);
      tc_fence_after();
      for (int c0 = eg * 32; c0 < a.BN; c0 += 64) {
        uint32_t raw[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + as * (uint32_t)a.BN + (uint32_t)c0, raw);
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(raw[j]);
        if (a.bias) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 bb = *reinterpret_cast<const float4*>(sbias + n0 + c0 + j);
            v[j] += bb.x; v[j + 1] += bb.y; v[j + 2] += bb.z; v[j + 3] += bb.w;
          }
        }
        if (a.add && valid) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            float t8[8];
            ck::unpack8(ck::ldg16(a.add + pix * a.N + n0 + c0 + j), t8);
#pragma unroll
            for (int e = 0; e < 8; ++e) v[j + e] += t8[e];
          }
        }
        if (a.stats && !valid) {  // out-of-image rows are clipped by the TMA store but must not enter the statistics
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
        }
        // bf16 pack + swizzled staging store (row r of sub-tile c0 / sub_w)
        const int sub = c0 >> sub_sh, cin_sub = c0 & (a.sub_w - 1);  // sub_w is 64 or 32: no divide sequence per chunk
        const uint32_t rowaddr = obuf + sub * sub_bytes + (uint32_t)r * row_bytes;
        const uint32_t xorv = (a.sub_w == 64) ? (uint32_t)(r & 7) : (uint32_t)((r >> 1) & 3);
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          const uint4 pk = ck::pack8(&v[j]);
          const uint32_t chunk = (uint32_t)((cin_sub + j) >> 3);
          const uint32_t dst = rowaddr + ((chunk ^ xorv) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(pk.x), "r"(pk.y), "r"(pk.z), "r"(pk.w)
                       : "memory");
        }
      }
      // accumulator stage fully read: hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty0 + 8u * as);
      // make the staging writes visible to the async proxy, then one thread stores the tile
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      epi_bar_sync();
      if (et == 0) {
        for (int sub = 0; sub < nsub; ++sub)
          tma_store_4d(&tmY, obuf + sub * sub_bytes, n0 + sub * a.sub_w, w0, h0, b0);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      if (a.stats) {
        // the staged tile is complete (barrier above) and stays untouched until two tiles from now
        const int col = 2 * s_wcol;
        const int sub = col >> sub_sh, cin_sub = col & (a.sub_w - 1);
        const uint32_t chunk = (uint32_t)(cin_sub >> 3), within = (uint32_t)(cin_sub & 7) * 2u;
        const uint32_t colbase = obuf + sub * sub_bytes + within;
        // rows of a thread start at a multiple of 8 and s_rows (= BN / 4) is a multiple of 8: the swizzle term of row
        // r0 + i only depends on i & 7, so the eight chunk offsets are tile-invariant and the loop body is load + unpack +
        // two packed adds (the first version recomputed the swizzled address per row: ~13 instructions per row, as many
        // instructions per tile as the rest of the epilogue - ncu source page of conv_tc_fwd_v2_k)
        uint32_t raddr = colbase + (uint32_t)(s_rgrp * s_rows) * row_bytes;
        for (int rr = 0; rr < s_rows; rr += 8) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t xv = (a.sub_w == 64) ? (uint32_t)i : (uint32_t)(i >> 1);
            uint32_t w;
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(raddr + (uint32_t)i * row_bytes + ((chunk ^ xv) << 4)));
            const f2_t x = bf2_to_f2(w);
            cs1 = fadd2(cs1, x);
            cs2 = ffma2(x, x, cs2);
          }
          raddr += 8u * row_bytes;
        }
      }
    }
    if (a.stats && cur_n0 >= 0) flush_stats(cur_n0);
    // only the shared-memory reads of the last stores must finish before the CTA retires; the global writes are
    // complete at grid end like any other store
    if (et == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, ncols);
  }
}


// ----------------------------------------------------------------------------------------------------------
struct WgradArgs {
  float* dW;
  long s_tap, s_ci, s_co;
  int B, H, W, Cin, Cout, taps;
  int cw, nw, BN, mblk, nblk;     // channel width of one A box / one B box; boxes per operand
  int total_blocks, ntiles, tiles_per_split, stages;
  int dbg_noatomic;  // profiling aid: skip the epilogue atomics
  float* ws;         // per-CTA partial tiles [grid][128][BN] (NULL: fp32 atomics straight into dW)
  int cluster;       // > 1: the `cluster` CTAs along the split axis sum their tiles through distributed shared memory
  int cin_valid;
  uint32_t layoutA, sboA, layoutB, sboB;
  TileGeom g;
};

__global__ void __launch_bounds__(192) wgrad_tc_k(const __grid_constant__ CUtensorMap tmX,
                                                  const __grid_constant__ CUtensorMap tmDY, const WgradArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t ablk_bytes = 128u * a.cw * 2u, bblk_bytes = 128u * a.nw * 2u;
  // a layer with a single (tap, chunk) block only stages that block (the other M-blocks alias it, see lboA below)
  const uint32_t a_bytes = ablk_bytes * (a.total_blocks == 1 ? 1u : (uint32_t)a.mblk), b_bytes = bblk_bytes * a.nblk;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t bar_base = base + a.stages * stage_bytes;
  const uint32_t full0 = bar_base, empty0 = bar_base + 8u * a.stages, tfull = bar_base + 16u * a.stages;
  const uint32_t tptr = tfull + 8u;
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int group = blockIdx.x;
  const int n0 = blockIdx.y * a.BN;
  const int tile_begin = blockIdx.z * a.tiles_per_split;
  const int tile_end = min(a.ntiles, tile_begin + a.tiles_per_split);
  const int iters = tile_end - tile_begin;
  const int cchunks = a.Cin / a.cw;
  const uint32_t ncols = (uint32_t)(a.BN < 32 ? 32 : a.BN);
  // a group with a single real (tap, chunk) block does not load padding copies: the descriptor's block stride
  // (LBO) is set to 0 so that every M-block of the MMA aliases block 0; the duplicate rows are discarded.
  const int nvalid = min(a.mblk, a.total_blocks - group * a.mblk);
  const int nload = nvalid == 1 ? 1 : a.mblk;
  const uint32_t lboA = nvalid == 1 ? 0u : ablk_bytes;

  if (threadIdx.x == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(full0 + 8u * s, 1);
      mbar_init(empty0 + 8u * s, 1);
    }
    mbar_init(tfull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmDY);
  }
  if (warp == 1) tmem_alloc(tptr, ncols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  ck::pdl_enter();  // set-up above overlapped the previous kernel's tail; global memory is touched only from here on
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tptr));

  if (iters <= 0 && a.ws && warp >= 2) {
    const long cta = ((long)blockIdx.x * gridDim.y + blockIdx.y) * gridDim.z + blockIdx.z;
    const int r = (warp & 3) * 32 + lane;
    for (int c = 0; c < a.BN; ++c) a.ws[(cta * 128 + r) * a.BN + c] = 0.f;
  }
  // cluster mode: fp32 tile [128][BN + 4] in shared memory (the pipeline stages are idle by then)
  float* ctile = reinterpret_cast<float*>(smem_raw + (base - smem_u32(smem_raw)));
  const int cpitch = a.BN + 4;  // keeps rows 16-byte aligned for the vector reads of the reduction
  if (iters <= 0 && a.cluster > 1 && warp >= 2) {
    const int r = (warp & 3) * 32 + lane;
    for (int c = 0; c < a.BN; ++c) ctile[r * cpitch + c] = 0.f;
  }
  if (iters > 0) {
    if (warp == 0) {
      // ---- producer: uniform control flow, one elected lane issues (see tc_common.cuh)
      uint32_t s = 0, ph = 0;
      // (tap, chunk) of this CTA's M blocks: fixed for the whole pixel range, so the divisions are done once, not per stage
      int m_c[4], m_dh[4], m_dw[4];
#pragma unroll
      for (int m = 0; m < 4; ++m) {
        const int blk = min(group * a.mblk + m, a.total_blocks - 1);
        const int t = blk / cchunks;
        m_c[m] = (blk - t * cchunks) * a.cw;
        m_dh[m] = a.taps == 9 ? t / 3 - 1 : 0;
        m_dw[m] = a.taps == 9 ? t % 3 - 1 : 0;
      }
      for (int it = 0; it < iters; ++it) {
        mbar_wait(empty0 + 8u * s, ph ^ 1u);
        int tile = tile_begin + it;
        const int q_ = fd_div(tile, a.g.m_w);
        const int tw_i = tile - q_ * a.g.tiles_w;
        tile = fd_div(q_, a.g.m_h);
        const int th_i = q_ - tile * a.g.tiles_h;
        const int w0 = tw_i * a.g.TW, h0 = th_i * a.g.TH, b0 = tile * a.g.TB;
        const uint32_t sa = base + s * stage_bytes, sb = sa + a_bytes;
        if (elect_one()) {
          mbar_expect_tx(full0 + 8u * s, (uint32_t)nload * ablk_bytes + b_bytes);
#pragma unroll
          for (int m = 0; m < 4; ++m)
            if (m < nload) tma_load_4d(sa + m * ablk_bytes, &tmX, full0 + 8u * s, m_c[m], w0 + m_dw[m], h0 + m_dh[m], b0);
          for (int nb = 0; nb < a.nblk; ++nb)
            tma_load_4d(sb + nb * bblk_bytes, &tmDY, full0 + 8u * s, n0 + nb * a.nw, w0, h0, b0);
        }
        __syncwarp();
        if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1u; }
      }
    } else if (warp == 1) {
      // ---- MMA issue: uniform control flow, {lo, hi} descriptor words, one elected lane fires the instructions
      const uint32_t idesc = make_idesc(128, a.BN, 1, 1);
      const uint32_t ahi = desc_hi(a.sboA, a.layoutA), bhi = desc_hi(a.sboB, a.layoutB);
      const uint32_t astep = (uint32_t)(16 * a.cw * 2) >> 4, bstep = (uint32_t)(16 * a.nw * 2) >> 4;  // 16 pixels (K) per MMA
      uint32_t s = 0, ph = 0;
      for (int it = 0; it < iters; ++it) {
        mbar_wait(full0 + 8u * s, ph);
        tc_fence_after();
        const uint32_t sa = base + s * stage_bytes;
        const uint32_t alo = desc_lo(sa, lboA), blo = desc_lo(sa + a_bytes, bblk_bytes);
        if (elect_one()) {
          umma_f16_w(tmem_base, alo, ahi, blo, bhi, idesc, it > 0 ? 1u : 0u);
#pragma unroll
          for (uint32_t k = 1; k < 8; ++k) umma_f16_w(tmem_base, alo + k * astep, ahi, blo + k * bstep, bhi, idesc, 1u);
          umma_commit(empty0 + 8u * s);
        }
        __syncwarp();
        if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1u; }
      }
      if (elect_one()) umma_commit(tfull);
      __syncwarp();
    } else {
      const int q = warp & 3;
      const int r = q * 32 + lane;  // accumulator row = (block m, channel within block)
      const int m = r / a.cw;
      const int blk = group * a.mblk + m;
      const bool valid = blk < a.total_blocks;
      const int bc = valid ? blk : 0;
      const int t = bc / cchunks, ci = (bc % cchunks) * a.cw + (r % a.cw);
      float* dst = a.dW + (long)t * a.s_tap + (long)ci * a.s_ci;
      mbar_wait(tfull, 0);
      tc_fence_after();
      for (int c0 = 0; c0 < a.BN; c0 += 32) {
        uint32_t raw[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, raw);
        if (a.cluster > 1) {
#pragma unroll
          for (int j = 0; j < 32; ++j) ctile[r * cpitch + c0 + j] = __uint_as_float(raw[j]);  // lanes -> distinct banks
        } else if (a.ws) {
          // contention-free: this CTA's partial tile goes to its own slab, wgrad_reduce_k sums the slabs
          const long cta = ((long)blockIdx.x * gridDim.y + blockIdx.y) * gridDim.z + blockIdx.z;
          float4* wp = reinterpret_cast<float4*>(a.ws + (cta * 128 + r) * a.BN + c0);
          if (valid)  // padding blocks are never read back
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            wp[j >> 2] = make_float4(__uint_as_float(raw[j]), __uint_as_float(raw[j + 1]), __uint_as_float(raw[j + 2]),
                                     __uint_as_float(raw[j + 3]));
        } else if (valid && ci < a.cin_valid && !a.dbg_noatomic) {
          // no workspace: this CTA's partial tile is added straight into the gradient arena.  A thread owns one row
          // and 32 consecutive columns; with the columns contiguous in memory that is 8 vector reductions
          // (red.global.add.v4.f32) instead of 32 scalar ones, and no reduction launch behind the GEMM.
          float* d = dst + (long)(n0 + c0) * a.s_co;
          if (a.s_co == 1 && (reinterpret_cast<uintptr_t>(d) & 15) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(d + j), "f"(__uint_as_float(raw[j])),
                           "f"(__uint_as_float(raw[j + 1])), "f"(__uint_as_float(raw[j + 2])),
                           "f"(__uint_as_float(raw[j + 3])) : "memory");
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) atomicAdd(d + (long)j * a.s_co, __uint_as_float(raw[j]));
          }
        }
      }
    }
  }
  if (a.cluster > 1) {
    // Split-K reduction inside the cluster: every CTA has its 128 x BN partial tile in shared memory; CTA `rank` sums
    // rows [rank * 128/CS, (rank+1) * 128/CS) over all peers through distributed shared memory and adds them to the
    // gradient arena (lanes along the output-channel axis).  Replaces the workspace round trip + one reduction launch
    // per layer; the contention on an output element drops from #splits to #splits / CS atomics.
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (warp >= 2) {
      uint32_t rank;
      asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
      const int CS = a.cluster, rows_per = 128 / CS;
      const int et = threadIdx.x - 64;  // 0..127
      const uint32_t tile_s = smem_u32(ctile);
      uint32_t peer[8];  // this tile's base address in every peer's shared memory
#pragma unroll
      for (int p = 0; p < 8; ++p) {
        peer[p] = 0;
        if (p < CS) asm("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(peer[p]) : "r"(tile_s), "r"(p));
      }
      const int bn4 = a.BN >> 2;
      for (int idx = et; idx < rows_per * bn4; idx += 128) {
        const int rr = idx / bn4, c = (idx - rr * bn4) * 4;
        const int row = (int)rank * rows_per + rr;
        const int blk = group * a.mblk + row / a.cw;
        const int ci = (blk % cchunks) * a.cw + (row % a.cw);
        if (blk >= a.total_blocks || ci >= a.cin_valid) continue;
        const uint32_t off = (uint32_t)(row * cpitch + c) * 4u;
        float4 v[8];
#pragma unroll
        for (int p = 0; p < 8; ++p) {  // independent loads, all in flight together (ordered by the cluster barrier above)
          v[p] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (p < CS)
            asm("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
                : "=f"(v[p].x), "=f"(v[p].y), "=f"(v[p].z), "=f"(v[p].w) : "r"(peer[p] + off));
        }
        float4 sum = v[0];
#pragma unroll
        for (int p = 1; p < 8; ++p) { sum.x += v[p].x; sum.y += v[p].y; sum.z += v[p].z; sum.w += v[p].w; }
        const int t = blk / cchunks;
        float* d = a.dW + (long)t * a.s_tap + (long)ci * a.s_ci + (long)(n0 + c) * a.s_co;
        if (a.s_co == 1 && (reinterpret_cast<uintptr_t>(d) & 15) == 0) {
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(d), "f"(sum.x), "f"(sum.y), "f"(sum.z), "f"(sum.w)
                       : "memory");
        } else {
          atomicAdd(d, sum.x); atomicAdd(d + a.s_co, sum.y); atomicAdd(d + 2 * a.s_co, sum.z); atomicAdd(d + 3 * a.s_co, sum.w);
        }
      }
    }
    // nobody leaves (and frees its shared memory) while a peer may still read it
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, ncols);
  }
}

// ----------------------------------------------------------------------------------------------------------
// Halo weight gradient for the 3x3 layers.  wgrad_tc_k gives every (tap, channel-chunk) group its own CTAs, each of
// which reloads the shifted activation box AND the dY tile: for a 64 -> 64 layer that is 9 x 3 tensor reads out of L2
// (ncu: the MMA warp waits on TMA, tensor pipe 21 %, L2 the busiest unit).  Here a CTA owns one input-channel chunk x one
// 64-column dY tile x a range of pixel tiles and, per 128-pixel tile, loads ONE (16+2) x (8+2) halo tile of X and one dY
// tile; the nine taps are nine MN-major windows into the halo tile (start shifted by (kh*10 + kw) pixel rows, the 8-pixel
// K groups 10 rows apart, consecutive kw taps as consecutive M blocks one row apart), accumulating into 6 (64-channel
// chunks) or 3 (32-channel chunks) TMEM tiles of 128 x BN that live for the CTA's whole pixel range.  The last M block of
// a group may be a non-existent tap kw = 3: its rows are computed and dropped.  Partial tiles are added to the Keras-layout
// gradient arena with vector reductions.
struct WgradHaloArgs {
  float* dW;
  long s_tap, s_ci, s_co;
  int B, H, W, Cin, Cout;
  int cw, BN;                 // channel chunk of X (64 / 32), dY columns per CTA (64 / 32)
  int ntiles, tiles_per_split, stages, groups, khsplit;
  uint32_t layoutA, layoutB, sboB;
  uint32_t a_stage, a_bytes, b_bytes;
  TileGeom g;
};

__global__ void __launch_bounds__(192) wgrad_halo_k(const __grid_constant__ CUtensorMap tmX,
                                                    const __grid_constant__ CUtensorMap tmDY, const WgradHaloArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t stage_bytes = a.a_stage + a.b_bytes;  // both multiples of 1024
  const uint32_t bar_base = base + a.stages * stage_bytes;
  const uint32_t full0 = bar_base, empty0 = bar_base + 8u * a.stages, tfull = bar_base + 16u * a.stages;
  const uint32_t tptr = tfull + 8u;
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  // khsplit: one kernel row (kh) per CTA - a third of the accumulators and of the atomics per CTA, the halo tile is loaded
  // by three CTAs instead of one
  const int chunk = blockIdx.x;
  const int n0 = (a.khsplit ? (int)blockIdx.y / 3 : (int)blockIdx.y) * a.BN;
  const int kh_only = a.khsplit ? (int)blockIdx.y % 3 : -1;
  const int my_groups = a.khsplit ? a.groups / 3 : a.groups;
  const int tile_begin = blockIdx.z * a.tiles_per_split;
  const int tile_end = min(a.ntiles, tile_begin + a.tiles_per_split);
  const int iters = tile_end - tile_begin;
  const uint32_t row_b = (uint32_t)a.cw * 2u;
  uint32_t ncols = 32;
  while (ncols < (uint32_t)(my_groups * a.BN)) ncols <<= 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(full0 + 8u * s, 1);
      mbar_init(empty0 + 8u * s, 1);
    }
    mbar_init(tfull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmDY);
  }
  if (warp == 1) tmem_alloc(tptr, ncols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  ck::pdl_enter();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tptr));

  if (iters > 0) {
    if (warp == 0) {
      uint32_t s = 0, ph = 0;
      for (int it = 0; it < iters; ++it) {
        mbar_wait(empty0 + 8u * s, ph ^ 1u);
        int tile = tile_begin + it;
        const int q_ = fd_div(tile, a.g.m_w);
        const int tw_i = tile - q_ * a.g.tiles_w;
        tile = fd_div(q_, a.g.m_h);
        const int th_i = q_ - tile * a.g.tiles_h;
        const int w0 = tw_i * a.g.TW, h0 = th_i * a.g.TH, b0 = tile;
        const uint32_t sa = base + s * stage_bytes;
        if (elect_one()) {
          mbar_expect_tx(full0 + 8u * s, a.a_bytes + a.b_bytes);
          tma_load_4d(sa, &tmX, full0 + 8u * s, chunk * a.cw, w0 - 1, h0 - 1, b0);       // halo: zero fill = padding
          tma_load_4d(sa + a.a_stage, &tmDY, full0 + 8u * s, n0, w0, h0, b0);
        }
        __syncwarp();
        if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1u; }
      }
    } else if (warp == 1) {
      const uint32_t idesc = make_idesc(128, a.BN, 1, 1);
      // A: M blocks (consecutive kw taps) one pixel row apart, 8-pixel K groups one HALO row (10 pixels) apart
      const uint32_t ahi = desc_hi(10u * row_b, a.layoutA), bhi = desc_hi(a.sboB, a.layoutB);
      const uint32_t rb16 = row_b >> 4;
      const uint32_t akstep = 20u * rb16;                                   // 16 pixels of K = 2 tile rows = 20 halo rows
      const uint32_t bkstep = (16u * (uint32_t)a.BN * 2u) >> 4;            // 16 dY rows
      const int gpk = a.groups / 3;                                         // groups per kh: 2 (cw 64) or 1 (cw 32)
      uint32_t s = 0, ph = 0;
      for (int it = 0; it < iters; ++it) {
        mbar_wait(full0 + 8u * s, ph);
        tc_fence_after();
        const uint32_t sa = base + s * stage_bytes;
        const uint32_t alo0 = desc_lo(sa, row_b), blo0 = desc_lo(sa + a.a_stage, 0);
        if (elect_one()) {
          for (int g = 0; g < my_groups; ++g) {
            const int kh = kh_only >= 0 ? kh_only : g / gpk, kw0 = (kh_only >= 0 ? g : g - kh * gpk) * 2;
            const uint32_t alo = alo0 + (uint32_t)(kh * 10 + kw0) * rb16;
            const uint32_t tacc = tmem_base + (uint32_t)(g * a.BN);
#pragma unroll
            for (uint32_t k = 0; k < 8; ++k)
              umma_f16_w(tacc, alo + k * akstep, ahi, blo0 + k * bkstep, bhi, idesc, (it > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(empty0 + 8u * s);
        }
        __syncwarp();
        if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1u; }
      }
      if (elect_one()) umma_commit(tfull);
      __syncwarp();
    } else {
      const int q = warp & 3;
      const int r = q * 32 + lane;           // accumulator row = (M block, channel within the chunk)
      const int blk = r / a.cw, ci = chunk * a.cw + (r - blk * a.cw);
      const int gpk = a.groups / 3;
      mbar_wait(tfull, 0);
      tc_fence_after();
      for (int g = 0; g < my_groups; ++g) {
        const int kh = kh_only >= 0 ? kh_only : g / gpk, kw = (kh_only >= 0 ? g : g - kh * gpk) * 2 + blk;
        const bool valid = kw < 3;           // the M block past the last tap is scratch
        float* dst = a.dW + (long)(kh * 3 + kw) * a.s_tap + (long)ci * a.s_ci;
        for (int c0 = 0; c0 < a.BN; c0 += 32) {
          uint32_t raw[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(g * a.BN + c0), raw);
          if (!valid) continue;
          float* d = dst + (long)(n0 + c0) * a.s_co;
          if (a.s_co == 1 && (reinterpret_cast<uintptr_t>(d) & 15) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(d + j), "f"(__uint_as_float(raw[j])),
                           "f"(__uint_as_float(raw[j + 1])), "f"(__uint_as_float(raw[j + 2])),
                           "f"(__uint_as_float(raw[j + 3])) : "memory");
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) atomicAdd(d + (long)j * a.s_co, __uint_as_float(raw[j]));
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, ncols);
  }
}

// Sum the per-CTA partial tiles of wgrad_tc_k over the pixel splits and add the result into the Keras-layout
// gradient arena (one writer per element, no atomics).  A block is (256 / zl) float4 columns x zl split lanes: every
// lane strides over the splits with four loads in flight, the lanes meet in shared memory, lane 0 writes.  (The first
// version walked all splits serially per thread: 16-64 blocks x up to 293 dependent-latency loads, 13-29 us a launch.)
__device__ __forceinline__ void wgrad_reduce_body(const float* __restrict__ ws, float* __restrict__ dW, int groups, int ntn,
                                                  int splits, int BN, int cw, int mblk, int total_blocks, int cchunks,
                                                  long s_tap, long s_ci, long s_co, int zl, int cin_valid, int bid, int nblk,
                                                  float4* sm) {
  const long per_tile = 128L * BN;
  const long total4 = (long)groups * ntn * per_tile / 4;
  const int vpb = 256 / zl;
  const int v = threadIdx.x % vpb, lane = threadIdx.x / vpb;
  for (long base = (long)bid * vpb; base < total4; base += (long)nblk * vpb) {
    const long i = (base + v) * 4;
    bool live = base + v < total4;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const long gt = live ? i / per_tile : 0;  // g * ntn + nt
    const long in_tile = i - gt * per_tile;
    {
      // rows of padding blocks / padded input channels are neither read nor written (a 32-channel 1x1 layer fills only a
      // quarter of the 128-row tile)
      const int r_ = (int)(in_tile / BN);
      const int blk_ = (int)(gt / ntn) * mblk + r_ / cw;
      live = live && blk_ < total_blocks && (blk_ % cchunks) * cw + (r_ % cw) < cin_valid;
    }
    if (live) {
      const float4* p = reinterpret_cast<const float4*>(ws + (gt * splits) * per_tile + in_tile);
      const long stride4 = per_tile / 4;
      int z = lane;
      for (; z + 3 * zl < splits; z += 4 * zl) {
        const float4 a0 = __ldg(p + (long)z * stride4), a1 = __ldg(p + (long)(z + zl) * stride4);
        const float4 a2 = __ldg(p + (long)(z + 2 * zl) * stride4), a3 = __ldg(p + (long)(z + 3 * zl) * stride4);
        acc.x += (a0.x + a1.x) + (a2.x + a3.x); acc.y += (a0.y + a1.y) + (a2.y + a3.y);
        acc.z += (a0.z + a1.z) + (a2.z + a3.z); acc.w += (a0.w + a1.w) + (a2.w + a3.w);
      }
      for (; z < splits; z += zl) {
        const float4 a0 = __ldg(p + (long)z * stride4);
        acc.x += a0.x; acc.y += a0.y; acc.z += a0.z; acc.w += a0.w;
      }
    }
    if (zl > 1) {
      __syncthreads();
      sm[threadIdx.x] = acc;
      __syncthreads();
      if (lane == 0)
        for (int l = 1; l < zl; ++l) {
          const float4 o = sm[l * vpb + v];
          acc.x += o.x; acc.y += o.y; acc.z += o.z; acc.w += o.w;
        }
    }
    if (!live || lane != 0) continue;
    const int c = (int)(in_tile % BN);
    const int r = (int)(in_tile / BN);
    const int nt = (int)(gt % ntn), g = (int)(gt / ntn);
    const int blk = g * mblk + r / cw;
    if (blk >= total_blocks) continue;
    const int t = blk / cchunks, ci = (blk % cchunks) * cw + (r % cw);
    if (ci >= cin_valid) continue;  // zero-padded input channels (entry conv: 27 of 32)
    float* d = dW + (long)t * s_tap + (long)ci * s_ci + (long)(nt * BN + c) * s_co;
    d[0] += acc.x; d[s_co] += acc.y; d[2 * s_co] += acc.z; d[3 * s_co] += acc.w;
  }
}

__global__ void __launch_bounds__(256) wgrad_reduce_k(const float* __restrict__ ws, float* __restrict__ dW, int groups, int ntn,
                                                      int splits, int BN, int cw, int mblk, int total_blocks, int cchunks,
                                                      long s_tap, long s_ci, long s_co, int zl, int cin_valid) {
  ck::pdl_enter();
  __shared__ float4 sm[256];
  wgrad_reduce_body(ws, dW, groups, ntn, splits, BN, cw, mblk, total_blocks, cchunks, s_tap, s_ci, s_co, zl, cin_valid,
                    blockIdx.x, gridDim.x, sm);
}

// every layer's split reduction in ONE launch at the end of the backward pass (blockIdx.y = layer): the per-layer
// reduction launches were 22 x 5-12 us of mostly fixed cost
__global__ void __launch_bounds__(256) wgrad_reduce_all_k(const crackfl_wgrad_red_desc* __restrict__ descs,
                                                          const float* __restrict__ ws_base, float* __restrict__ dw_base) {
  ck::pdl_enter();
  __shared__ float4 sm[256];
  const crackfl_wgrad_red_desc d = descs[blockIdx.y];
  wgrad_reduce_body(ws_base + d.ws_off, dw_base + d.dw_off, d.groups, d.ntn, d.splits, d.BN, d.cw, d.mblk, d.total_blocks,
                    d.cchunks, d.s_tap, d.s_ci, d.s_co, d.zl, d.cin_valid, blockIdx.x, gridDim.x, sm);
}

// ----------------------------------------------------------------------------------------------------------
// host side: tensor-map cache + geometry
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  });
  return fn;
}

struct MapKey {
  const void* ptr;
  int d[4];
  int box[4];
  int swz;
  bool operator==(const MapKey& o) const { return memcmp(this, &o, sizeof(MapKey)) == 0; }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    std::string s(reinterpret_cast<const char*>(&k), sizeof(MapKey));
    return std::hash<std::string>()(s);
  }
};
std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_maps;
std::mutex g_maps_mu;

// bf16 tensor with `rank` dims (innermost first) tightly packed; box per dim; swizzle 128B or 64B.
}  // namespace

// swz_bytes | CK_MAP_OOB_NAN: out-of-bounds elements are delivered as NaN instead of zero (for consumers that apply
// relu(scale * x + shift) on load: max(NaN, 0) = 0, where a zero would have become relu(shift))
int ck::get_tensor_map(const void* ptr, int rank, const int* dims, const int* box, int swz_bytes, CUtensorMap* out) {
  const bool oob_nan = (swz_bytes & CK_MAP_OOB_NAN) != 0;
  swz_bytes &= ~CK_MAP_OOB_NAN;
  MapKey k;
  memset(&k, 0, sizeof(k));
  k.ptr = ptr;
  for (int i = 0; i < rank; ++i) { k.d[i] = dims[i]; k.box[i] = box[i]; }
  k.swz = swz_bytes * 8 + rank + (oob_nan ? (1 << 20) : 0);
  std::lock_guard<std::mutex> lk(g_maps_mu);
  auto it = g_maps.find(k);
  if (it != g_maps.end()) { *out = it->second; return 0; }
  PFN_encodeTiled enc = get_encode();
  if (!enc) return (int)cudaErrorNotSupported;
  cuuint64_t gdim[4], gstr[3];
  cuuint32_t bx[4], es[4];
  cuuint64_t stride = 2;
  for (int i = 0; i < rank; ++i) {
    gdim[i] = (cuuint64_t)dims[i];
    bx[i] = (cuuint32_t)box[i];
    es[i] = 1;
    stride *= (cuuint64_t)dims[i];
    if (i < rank - 1) gstr[i] = stride;
  }
  CUtensorMap tm;
  CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(ptr), gdim, gstr, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swz_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (swz_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE),
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   oob_nan ? CU_TENSOR_MAP_FLOAT_OOB_FILL_NAN_REQUEST_ZERO_FMA : CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return 10000 + (int)r;
  g_maps[k] = tm;
  *out = tm;
  return 0;
}

namespace {
inline int get_map(const void* ptr, int rank, const int* dims, const int* box, int swz_bytes, CUtensorMap* out) {
  return ck::get_tensor_map(ptr, rank, dims, box, swz_bytes, out);
}

int pick(int extent, int maxv) {  // power of two in {4..maxv} minimising padded waste, larger wins ties
  int best = 4;
  long bestw = -1;
  for (int v = maxv; v >= 4; v >>= 1) {
    const long padded = (long)((extent + v - 1) / v) * v;
    if (bestw < 0 || padded < bestw) { bestw = padded; best = v; }
  }
  return best;
}

TileGeom make_geom(int B, int H, int W) {
  TileGeom g;
  g.TW = pick(W, 16);
  int rest = 128 / g.TW;
  g.TH = 1;
  if (H >= 2) {
    int v = pick(H, rest < 4 ? 4 : rest);
    if (v > rest) v = rest;
    g.TH = v;
  }
  g.TB = 128 / (g.TW * g.TH);
  g.tiles_w = (W + g.TW - 1) / g.TW;
  g.tiles_h = (H + g.TH - 1) / g.TH;
  g.tiles_b = (B + g.TB - 1) / g.TB;
  g.m_w = fd_magic(g.tiles_w); g.m_h = fd_magic(g.tiles_h); g.m_ntm = fd_magic(g.tiles_w * g.tiles_h * g.tiles_b);
  return g;
}

}  // namespace

extern "C" {

int crackfl_tc_selftest_supported(void) { return get_encode() ? 1 : 0; }

static int conv_tc_fwd_v1(const void* X, const void* Wk, const float* bias, const void* add, void* Y, float* stats, int B, int H,
                          int W, int Cin, int N, int taps, cudaStream_t st) {
  if (Cin % 32 || N % 32 || (taps != 1 && taps != 9)) return (int)cudaErrorInvalidValue;
  ConvArgs a;
  a.bias = bias; a.add = (const __nv_bfloat16*)add; a.Y = (__nv_bfloat16*)Y; a.stats = stats;
  a.B = B; a.H = H; a.W = W; a.Cin = Cin; a.N = N; a.taps = taps;
  a.BN = N > 128 ? 128 : N;
  if (N % a.BN) return (int)cudaErrorInvalidValue;
  a.kc = (Cin % 64 == 0) ? 64 : 32;
  a.layout = a.kc == 64 ? 2u : 4u;
  a.sbo = a.kc == 64 ? 1024u : 512u;
  a.g = make_geom(B, H, W);
  const int iters = taps * (Cin / a.kc);
  const uint32_t stage_bytes = 128u * a.kc * 2u + (uint32_t)a.BN * a.kc * 2u;
  int stages = (int)(98304u / stage_bytes);
  if (stages > 6) stages = 6;
  if (stages > iters) stages = iters;
  if (stages < 1) stages = 1;
  a.stages = stages;
  const size_t smem = (size_t)stages * stage_bytes + 1024 + 256;
  CUtensorMap tmA, tmB;
  const int dA[4] = {Cin, W, H, B}, bA[4] = {a.kc, a.g.TW, a.g.TH, a.g.TB};
  int rc = get_map(X, 4, dA, bA, a.kc * 2, &tmA);
  if (rc) return rc;
  const int dB[2] = {taps * Cin, N}, bB[2] = {a.kc, a.BN};
  rc = get_map(Wk, 2, dB, bB, a.kc * 2, &tmB);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_tc_fwd_k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  dim3 grid(a.g.tiles_w * a.g.tiles_h * a.g.tiles_b, N / a.BN);
  ck::launch(conv_tc_fwd_k, grid, 192, smem, st, tmA, tmB, a);
  CK_LAUNCH_CHECK();
  return 0;
}

static int g_num_sms = 0;
static int wred_cap() {  // blocks per SM of the split-reduction kernel (tuning knob)
  static int v = -1;
  // measured on the full step (ms): 1: 1.283, 2: 1.287, 3: 1.289, 4: 1.297, 8: 1.312, 16: 1.347.  With the engine's
  // one-descriptor reduction launches (fixed grid) the knob acts through the number of split lanes per block it allows:
  // few lanes with long load chains beat many lanes + a shared-memory tree next to the following GEMM
  if (v < 0) { const char* e = getenv("CRACKFL_WRED_BLOCKS_PER_SM"); v = e ? atoi(e) : 1; if (v < 1) v = 1; }
  return v;
}
static int dbg_noflush() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("CRACKFL_DBG_NOFLUSH"); v = e ? atoi(e) : 0; }
  return v;
}

int crackfl_conv_tc_fwd(const void* X, const void* Wk, const float* bias, const void* add, void* Y, float* stats, int B, int H,
                        int W, int Cin, int N, int taps, cudaStream_t st) {
  static int use_v1 = -1;
  if (use_v1 < 0) {
    const char* e = getenv("CRACKFL_TC_V1");
    use_v1 = (e && e[0] == '1') ? 1 : 0;
  }
  if (use_v1) return conv_tc_fwd_v1(X, Wk, bias, add, Y, stats, B, H, W, Cin, N, taps, st);
  if (Cin % 32 || N % 32 || (taps != 1 && taps != 9)) return (int)cudaErrorInvalidValue;
  static int use_halo = -1, halo_base_off = 0;  // measured: tcgen05 applies the swizzle XOR on absolute smem address bits, so a window that starts
  // mid-atom needs base_offset = 0 (setting it from the start address gives wrong results)
  if (use_halo < 0) {
    const char* e = getenv("CRACKFL_TC_HALO");
    use_halo = e ? atoi(e) : 1;
    const char* e2 = getenv("CRACKFL_HALO_BASEOFF");
    if (e2) halo_base_off = atoi(e2);
  }
  // measured (profiles/conv_bench_r1.json): the halo kernel wins when the N tile is 128 wide (d0/d1 and the 64->128
  // dgrads); with narrower N tiles the per-tap barrier round trips of its single producer/issuer threads dominate and
  // the two-CTA-per-SM v2 kernel is faster.  CRACKFL_TC_HALO=2 forces it for every eligible layer.
  const int kc3 = (Cin % 64 == 0) ? 64 : 32;
  int BN3 = N > 128 ? 128 : N;
  {
    // tuning knob.  Measured: 64-wide n-tiles for the small-M layers (fewer tiles than SMs at 128) do not help (d0 23 ->
    // 22 us) or hurt (d1 24 -> 33 us): those layers are bound by the L2 -> shared-memory weight stream, not by the MMA chain
    static int dbg_bn = -1;
    if (dbg_bn < 0) { const char* e = getenv("CRACKFL_V3_BN"); dbg_bn = e ? atoi(e) : 0; }
    if (dbg_bn > 0 && N % dbg_bn == 0 && dbg_bn <= BN3) BN3 = dbg_bn;
  }
  const bool halo_ok = taps == 9 && Cin % kc3 == 0 && H >= 16 && W >= 8;
  // resident-weight mode: every (chunk, tap) weight tile of the layer fits next to the halo ring
  const uint32_t wbytes = 9u * (uint32_t)Cin * (uint32_t)BN3 * 2u;
  static int use_res = -1;
  if (use_res < 0) { const char* e = getenv("CRACKFL_TC_RESIDENT"); use_res = e ? atoi(e) : 1; }
  const bool resident = halo_ok && use_res && N == BN3 && wbytes <= 81920u;
  if (halo_ok && (use_halo == 2 || (use_halo == 1 && (N >= 128 || resident)))) {
    if (kc3 == 32 && !resident) goto v2_path;  // the streaming variant is only set up for 64-channel chunks
    ConvArgs3 a;
    a.bias = bias; a.add = (const __nv_bfloat16*)add; a.stats = stats;
    a.B = B; a.H = H; a.W = W; a.Cin = Cin; a.N = N; a.taps = taps;
    a.BN = BN3;
    if (N % a.BN) return (int)cudaErrorInvalidValue;
    a.kc = kc3; a.layout = kc3 == 64 ? 2u : 4u; a.sbo = kc3 == 64 ? 1024u : 512u;
    a.sub_w = (a.BN % 64 == 0) ? 64 : 32;
    a.base_off = halo_base_off;
    a.dbg_noflush = dbg_noflush();
    a.resident = resident ? 1 : 0;
    a.a_bytes = 180u * (uint32_t)kc3 * 2u;
    a.a_stage = (a.a_bytes + 1023u) & ~1023u;
    a.g.TW = 8; a.g.TH = 16; a.g.TB = 1;
    a.g.tiles_w = (W + 7) / 8; a.g.tiles_h = (H + 15) / 16; a.g.tiles_b = B;
    a.n_tiles_m = a.g.tiles_w * a.g.tiles_h * a.g.tiles_b;
    a.g.m_w = fd_magic(a.g.tiles_w); a.g.m_h = fd_magic(a.g.tiles_h); a.g.m_ntm = fd_magic(a.n_tiles_m);
  a.g.m_w = fd_magic(a.g.tiles_w); a.g.m_h = fd_magic(a.g.tiles_h); a.g.m_ntm = fd_magic(a.n_tiles_m);
    a.n_tiles_n = N / a.BN;
    const uint32_t b_bytes = (uint32_t)a.BN * (uint32_t)kc3 * 2u;
    const uint32_t out_bytes = 2u * 128u * (uint32_t)a.BN * 2u;
    const uint32_t fixed = out_bytes + 10240u;  // staging + barriers + statistics scratch + alignment slack
    int per_sm = 1;
    int stages = 0;
    if (resident) {
      stages = 9 * (Cin / kc3);
      // two CTAs per SM (<= 113 KB each) when the weights plus >= 2 halo stages fit, else one CTA with a deeper ring
      const uint32_t lim2 = 114688u, lim1 = 220u * 1024u;
      if (wbytes + fixed + 2u * a.a_stage <= lim2) {
        per_sm = 2;
        a.astages = (int)((lim2 - wbytes - fixed) / a.a_stage);
      } else {
        a.astages = (int)((lim1 - wbytes - fixed) / a.a_stage);
      }
      if (a.astages > 4) a.astages = 4;
      {
        static int dbg_persm = -1, dbg_ast = -1;  // tuning knobs
        if (dbg_persm < 0) { const char* e = getenv("CRACKFL_V3_PERSM"); dbg_persm = e ? atoi(e) : 0; }
        if (dbg_ast < 0) { const char* e = getenv("CRACKFL_V3_ASTAGES"); dbg_ast = e ? atoi(e) : 0; }
        if (dbg_persm == 1 && per_sm == 2) {
          per_sm = 1;
          a.astages = (int)((lim1 - wbytes - fixed) / a.a_stage);
          if (a.astages > 4) a.astages = 4;
        }
        if (dbg_ast >= 2 && dbg_ast <= a.astages) a.astages = dbg_ast;
      }
      if (a.astages < 2) return (int)cudaErrorInvalidValue;
    } else {
      a.astages = 2;
      // two CTAs per SM when >= 3 weight stages still fit in ~104 KB (small BN), else one CTA with a deep ring
      if (2u * a.a_stage + out_bytes + 3u * b_bytes + 10240u <= 106496u) {
        per_sm = 2;
        stages = (int)((106496u - 10240u - 2u * a.a_stage - out_bytes) / b_bytes);
      } else {
        stages = (int)((200704u - 2u * a.a_stage - out_bytes) / b_bytes);
      }
      {
        static int cap = -1;  // tuning knob: depth of the streamed weight ring
        if (cap < 0) { const char* e = getenv("CRACKFL_V3_BSTAGES"); cap = e ? atoi(e) : 9; if (cap < 2 || cap > 24) cap = 9; }
        if (stages > cap) stages = cap;
      }
      if (stages < 2) stages = 2;
    }
    a.stages = stages;
    const size_t smem = (size_t)a.astages * a.a_stage + (size_t)stages * b_bytes + out_bytes + 1024 + 512 + 8192;
    CUtensorMap tmA, tmB, tmY;
    const int dA[4] = {Cin, W, H, B}, bA[4] = {kc3, 10, 18, 1};
    int rc = get_map(X, 4, dA, bA, kc3 * 2, &tmA);
    if (rc) return rc;
    const int dB[2] = {taps * Cin, N}, bB[2] = {kc3, a.BN};
    rc = get_map(Wk, 2, dB, bB, kc3 * 2, &tmB);
    if (rc) return rc;
    const int dY[4] = {N, W, H, B}, bY[4] = {a.sub_w, 8, 16, 1};
    rc = get_map(Y, 4, dY, bY, a.sub_w * 2, &tmY);
    if (rc) return rc;
    static bool attr3 = false;
    if (!attr3) {
      cudaError_t e = cudaFuncSetAttribute(conv_tc_fwd_v3_k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
      if (e != cudaSuccess) return (int)e;
      int dev = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
      if (g_num_sms <= 0) g_num_sms = 148;
      attr3 = true;
    }
    const int total = a.n_tiles_m * a.n_tiles_n;
    int grid = per_sm * g_num_sms;
    if (grid > total) grid = total;
    ck::launch(conv_tc_fwd_v3_k, grid, 64 + EPI_THREADS, smem, st, tmA, tmB, tmY, a);
    CK_LAUNCH_CHECK();
    return 0;
  }
v2_path:
  ConvArgs2 a;
  a.dbg_noflush = dbg_noflush();
  a.bias = bias; a.add = (const __nv_bfloat16*)add; a.stats = stats;
  a.B = B; a.H = H; a.W = W; a.Cin = Cin; a.N = N; a.taps = taps;
  a.BN = N > 128 ? 128 : N;
  if (N % a.BN) return (int)cudaErrorInvalidValue;
  a.kc = (Cin % 64 == 0) ? 64 : 32;
  a.layout = a.kc == 64 ? 2u : 4u;
  a.sbo = a.kc == 64 ? 1024u : 512u;
  a.sub_w = (a.BN % 64 == 0) ? 64 : 32;
  a.g = make_geom(B, H, W);
  a.n_tiles_m = a.g.tiles_w * a.g.tiles_h * a.g.tiles_b;
  a.g.m_w = fd_magic(a.g.tiles_w); a.g.m_h = fd_magic(a.g.tiles_h); a.g.m_ntm = fd_magic(a.n_tiles_m);
  a.n_tiles_n = N / a.BN;
  const int iters = taps * (Cin / a.kc);
  const uint32_t stage_bytes = 128u * a.kc * 2u + (uint32_t)a.BN * a.kc * 2u;
  const uint32_t out_bytes = 2u * 128u * (uint32_t)a.BN * 2u;
  int stages = (int)((102400u - out_bytes) / stage_bytes);  // ~104 KB per CTA -> two CTAs per SM
  if (stages > 8) stages = 8;
  if (stages < 2) stages = 2;
  a.stages = stages;
  (void)iters;
  const size_t smem = (size_t)stages * stage_bytes + out_bytes + 1024 + 256 + 8192;
  CUtensorMap tmA, tmB, tmY;
  const int dA[4] = {Cin, W, H, B}, bA[4] = {a.kc, a.g.TW, a.g.TH, a.g.TB};
  int rc = get_map(X, 4, dA, bA, a.kc * 2, &tmA);
  if (rc) return rc;
  const int dB[2] = {taps * Cin, N}, bB[2] = {a.kc, a.BN};
  rc = get_map(Wk, 2, dB, bB, a.kc * 2, &tmB);
  if (rc) return rc;
  const int dY[4] = {N, W, H, B}, bY[4] = {a.sub_w, a.g.TW, a.g.TH, a.g.TB};
  rc = get_map(Y, 4, dY, bY, a.sub_w * 2, &tmY);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_tc_fwd_v2_k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return (int)e;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
    attr_set = true;
  }
  const int total = a.n_tiles_m * a.n_tiles_n;
  int grid = 2 * g_num_sms;
  if (grid > total) grid = total;
  ck::launch(conv_tc_fwd_v2_k, grid, 64 + EPI_THREADS, smem, st, tmA, tmB, tmY, a);
  CK_LAUNCH_CHECK();
  return 0;
}

// 3x3 conv (or, with diag != 0, depthwise 3x3 through block-diagonal weight tiles) whose INPUT is activated on the fly
// inside the kernel: xmode 1 = relu(X), xmode 2 = relu(scale * X + shift) with scale/shift from `bn` ([scale C | shift C])
// or, when fin != NULL, derived from the BN statistics (block 0 then also publishes the table and the moving
// statistics, like the *_fin elementwise launchers).  side != NULL: the activated input is also written out (the weight
// gradient of this layer needs it).  diag: Wk is the [(C/kc)*9*kc][kc] matrix written by crackfl_dw_diag_prep, N == Cin.
static int conv_tc_xf_impl(const void* X, const void* X2, const void* Wk, const float* bias, const void* add, void* Y,
                           float* stats, int B, int H, int W, int Cin, int N, int xmode, const float* bn,
                           const crackfl_bn_fin* fin, void* side, int diag, const float* sums, float count, float* dgamma,
                           float* dbeta, cudaStream_t st) {
  if (Cin % 32 || N % 32 || H < 16 || W < 8 || xmode < 1 || xmode > 5) return (int)cudaErrorInvalidValue;  // 3: profiling aid, tiles pass through untouched
  if (xmode == 2 && !bn && !fin) return (int)cudaErrorInvalidValue;
  if (xmode >= 4 && (!X2 || !bn || !sums || diag || count <= 0.f)) return (int)cudaErrorInvalidValue;
  if (diag && N != Cin) return (int)cudaErrorInvalidValue;
  const int kc = (Cin % 64 == 0) ? 64 : 32;
  if (Cin % kc) return (int)cudaErrorInvalidValue;
  ConvArgsX a;
  memset(&a, 0, sizeof(a));
  a.bias = bias; a.add = (const __nv_bfloat16*)add; a.stats = stats;
  a.B = B; a.H = H; a.W = W; a.Cin = Cin; a.N = N; a.taps = 9;
  a.BN = diag ? kc : (N > 128 ? 128 : N);
  if (N % a.BN) return (int)cudaErrorInvalidValue;
  a.kc = kc; a.layout = kc == 64 ? 2u : 4u; a.sbo = kc == 64 ? 1024u : 512u;
  a.sub_w = (a.BN % 64 == 0) ? 64 : 32;
  a.dbg_noflush = dbg_noflush();
  a.xmode = xmode; a.diag = diag ? 1 : 0; a.bn = bn; a.side = (__nv_bfloat16*)side;
  if (fin) a.fin = *fin;
  a.a_bytes = 180u * (uint32_t)kc * 2u;
  a.a_stage = (a.a_bytes + 1023u) & ~1023u;
  a.a_pitch = xmode >= 4 ? 2u * a.a_stage : a.a_stage;
  a.sums = sums; a.count = count; a.dgamma = dgamma; a.dbeta = dbeta;
  a.g.TW = 8; a.g.TH = 16; a.g.TB = 1;
  a.g.tiles_w = (W + 7) / 8; a.g.tiles_h = (H + 15) / 16; a.g.tiles_b = B;
  a.n_tiles_m = a.g.tiles_w * a.g.tiles_h * a.g.tiles_b;
  a.g.m_w = fd_magic(a.g.tiles_w); a.g.m_h = fd_magic(a.g.tiles_h); a.g.m_ntm = fd_magic(a.n_tiles_m);
  a.n_tiles_n = diag ? 1 : N / a.BN;
  const uint32_t b_bytes = (uint32_t)a.BN * (uint32_t)kc * 2u;
  const uint32_t out_bytes = 2u * 128u * (uint32_t)a.BN * 2u;
  const uint32_t fixed = out_bytes + 16384u;  // staging + barriers + statistics scratch + bias + coefficient tables + slack
  const uint32_t lim1 = 220u * 1024u;
  const uint32_t wbytes = diag ? 9u * b_bytes : 9u * (uint32_t)Cin * (uint32_t)a.BN * 2u;
  const bool resident = diag || (N == a.BN && wbytes + fixed + 2u * a.a_pitch <= lim1);
  a.resident = resident ? 1 : 0;
  int stages;
  if (resident) {
    stages = diag ? 9 : 9 * (Cin / kc);
    a.astages = (int)((lim1 - wbytes - fixed) / a.a_pitch);
    if (a.astages > 4) a.astages = 4;
    {
      static int dbg_ast = -1;  // tuning knob
      if (dbg_ast < 0) { const char* e = getenv("CRACKFL_XF_ASTAGES"); dbg_ast = e ? atoi(e) : 0; }
      if (dbg_ast >= 1 && dbg_ast <= a.astages) a.astages = dbg_ast;
    }
    if (a.astages < 1) return (int)cudaErrorInvalidValue;
  } else {
    if (kc != 64) return (int)cudaErrorInvalidValue;  // the streaming variant is only set up for 64-channel chunks
    a.astages = xmode >= 4 ? 2 : 3;
    stages = (int)((lim1 - fixed - (uint32_t)a.astages * a.a_pitch) / b_bytes);
    if (stages > 9) stages = 9;
    if (stages < 2) return (int)cudaErrorInvalidValue;
  }
  a.stages = stages;
  const size_t smem = (size_t)a.astages * a.a_pitch + (size_t)stages * b_bytes + out_bytes + 1024 + 512 + 8192 + 2048 + 4096;
  CUtensorMap tmA, tmB, tmY, tmA2;
  const int dA[4] = {Cin, W, H, B}, bA[4] = {kc, 10, 18, 1};
  int rc = get_map(X, 4, dA, bA, kc * 2, &tmA);
  if (rc) return rc;
  rc = get_map(X2 ? X2 : X, 4, dA, bA, kc * 2, &tmA2);
  if (rc) return rc;
  if (diag) {
    const int dB[2] = {kc, (Cin / kc) * 9 * kc}, bB[2] = {kc, kc};
    rc = get_map(Wk, 2, dB, bB, kc * 2, &tmB);
  } else {
    const int dB[2] = {9 * Cin, N}, bB[2] = {kc, a.BN};
    rc = get_map(Wk, 2, dB, bB, kc * 2, &tmB);
  }
  if (rc) return rc;
  const int dY[4] = {N, W, H, B}, bY[4] = {a.sub_w, 8, 16, 1};
  rc = get_map(Y, 4, dY, bY, a.sub_w * 2, &tmY);
  if (rc) return rc;
  static bool attrx = false;
  if (!attrx) {
    cudaError_t e = cudaFuncSetAttribute(conv_tc_xf_k<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_tc_xf_k<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return (int)e;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
    attrx = true;
  }
  int grid = g_num_sms;
  if (diag) {
    const int nch = Cin / kc;
    const int total = a.n_tiles_m * nch;
    if (grid > total) grid = total;
    grid = grid / nch * nch;  // every CTA keeps one channel chunk: the grid is a multiple of the chunk count
    if (grid < nch) grid = nch;
  } else {
    const int total = a.n_tiles_m * a.n_tiles_n;
    if (grid > total) grid = total;
  }
  if (xmode >= 4) ck::launch(conv_tc_xf_k<true>, grid, 64 + EPI_THREADS + 6 * 32, smem, st, tmA, tmB, tmY, tmA2, a);
  else ck::launch(conv_tc_xf_k<false>, grid, 64 + EPI_THREADS + 8 * 32, smem, st, tmA, tmB, tmY, tmA2, a);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_conv_tc_xf(const void* X, const void* Wk, const float* bias, const void* add, void* Y, float* stats, int B, int H,
                       int W, int Cin, int N, int xmode, const float* bn, const crackfl_bn_fin* fin, void* side, int diag,
                       cudaStream_t st) {
  if (xmode > 3) return (int)cudaErrorInvalidValue;
  return conv_tc_xf_impl(X, nullptr, Wk, bias, add, Y, stats, B, H, W, Cin, N, xmode, bn, fin, side, diag, nullptr, 0.f, nullptr,
                         nullptr, st);
}

// K2: 3x3 conv (the input gradient of a decoder conv) whose input dy is formed inside the kernel from the gradient dz w.r.t.
// the BN output and the BN input y:  dy = BN-backward(dz, y) per channel from `bn` ([scale|shift|mean|invstd]) and the
// reduce pass's `sums` ([sum dz | sum dz*(y-mean)]); relu_mask != 0 masks dz by bn(y) > 0 first (the BN + ReLU pair).
// dy is written to dy_out (the layer's weight gradient reads it) and block 0 writes dgamma / dbeta: the launch replaces
// crackfl_bn_bwd_apply + crackfl_conv_tc_fwd.
int crackfl_conv_tc_bnbwd(const void* dz, const void* y, const void* Wk, void* Y, int B, int H, int W, int Cin, int N,
                          const float* bn, const float* sums, int relu_mask, void* dy_out, float* dgamma, float* dbeta,
                          cudaStream_t st) {
  return conv_tc_xf_impl(dz, y, Wk, nullptr, nullptr, Y, nullptr, B, H, W, Cin, N, relu_mask ? 5 : 4, bn, nullptr, dy_out, 0, sums,
                         (float)((long)B * H * W), dgamma, dbeta, st);
}

// cin_valid < Cin: the last Cin - cin_valid input channels are zero padding; their gradient rows are not written
// (workspace path only)
// mode 0: launch + per-layer reduction; 1: launch only, fill *desc for crackfl_wgrad_reduce_all; 2: plan only (no launch)
static int wgrad_halo(const void* X, const void* dY, float* dW, int B, int H, int W, int Cin, int Cout, long s_tap, long s_ci,
                      long s_co, cudaStream_t st) {
  WgradHaloArgs a;
  memset(&a, 0, sizeof(a));
  a.dW = dW; a.s_tap = s_tap; a.s_ci = s_ci; a.s_co = s_co;
  a.B = B; a.H = H; a.W = W; a.Cin = Cin; a.Cout = Cout;
  a.cw = (Cin % 64 == 0) ? 64 : 32;
  a.BN = (Cout % 64 == 0) ? 64 : 32;
  a.groups = a.cw == 64 ? 6 : 3;
  a.layoutA = a.cw == 64 ? 2u : 4u;
  a.layoutB = a.BN == 64 ? 2u : 4u; a.sboB = a.BN == 64 ? 1024u : 512u;
  a.a_bytes = 180u * (uint32_t)a.cw * 2u;
  a.a_stage = (a.a_bytes + 1023u) & ~1023u;   // >= 4 spare pixel rows: the scratch tap kw = 3 may read past the tile
  a.b_bytes = 128u * (uint32_t)a.BN * 2u;
  a.g.TW = 8; a.g.TH = 16; a.g.TB = 1;
  a.g.tiles_w = (W + 7) / 8; a.g.tiles_h = (H + 15) / 16; a.g.tiles_b = B;
  a.ntiles = a.g.tiles_w * a.g.tiles_h * a.g.tiles_b;
  a.g.m_w = fd_magic(a.g.tiles_w); a.g.m_h = fd_magic(a.g.tiles_h); a.g.m_ntm = fd_magic(a.ntiles);
  const int chunks = Cin / a.cw, ntn = Cout / a.BN;
  const uint32_t stage_bytes = a.a_stage + a.b_bytes;
  static int khs = -1;
  if (khs < 0) { const char* e = getenv("CRACKFL_WGRAD_HALO_KH"); khs = e ? atoi(e) : 1; }
  a.khsplit = khs ? 1 : 0;
  const int kmul = a.khsplit ? 3 : 1;
  int per_sm = (2u * 2u * stage_bytes + 4096u <= 220u * 1024u && (a.groups / kmul) * a.BN <= 256) ? 2 : 1;
  int stages = (int)(((per_sm == 2 ? 110u : 220u) * 1024u - 2048u) / stage_bytes);
  if (stages > 5) stages = 5;
  if (stages < 2) { per_sm = 1; stages = (int)((220u * 1024u - 2048u) / stage_bytes); }
  static int fillh = -1;
  if (fillh < 0) { const char* e = getenv("CRACKFL_WGRAD_HALO_FILL"); fillh = e ? atoi(e) : 50; if (fillh < 10 || fillh > 200) fillh = 50; }
  int splits = (148 * per_sm * fillh / 100) / (chunks * ntn * kmul);
  if (splits > a.ntiles) splits = a.ntiles;
  if (splits < 1) splits = 1;
  a.tiles_per_split = (a.ntiles + splits - 1) / splits;
  splits = (a.ntiles + a.tiles_per_split - 1) / a.tiles_per_split;
  if (stages > a.tiles_per_split) stages = a.tiles_per_split;
  if (stages < 1) stages = 1;
  a.stages = stages;
  CUtensorMap tmX, tmDY;
  const int dX[4] = {Cin, W, H, B}, bX[4] = {a.cw, 10, 18, 1};
  int rc = get_map(X, 4, dX, bX, a.cw * 2, &tmX);
  if (rc) return rc;
  const int dD[4] = {Cout, W, H, B}, bD[4] = {a.BN, 8, 16, 1};
  rc = get_map(dY, 4, dD, bD, a.BN * 2, &tmDY);
  if (rc) return rc;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_halo_k, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr = true;
  }
  const size_t smem = (size_t)stages * stage_bytes + 1024 + 256;
  dim3 grid(chunks, ntn * kmul, splits);
  ck::launch(wgrad_halo_k, grid, 192, smem, st, tmX, tmDY, a);
  CK_LAUNCH_CHECK();
  return 0;
}

static int wgrad_impl(const void* X, const void* dY, float* dW, int B, int H, int W, int Cin, int Cout, int taps,
                      long s_tap, long s_ci, long s_co, float* workspace, long workspace_floats, int cin_valid,
                      int mode, crackfl_wgrad_red_desc* desc, cudaStream_t st) {
  if (Cin % 32 || Cout % 32 || (taps != 1 && taps != 9)) return (int)cudaErrorInvalidValue;
  {
    static int use_halo_w = -1;
    if (use_halo_w < 0) { const char* e = getenv("CRACKFL_WGRAD_HALO"); use_halo_w = e ? atoi(e) : 1; }
    // measured per layer (bench/wgrad_bench.py, us, per-(tap,chunk) kernel -> halo kernel with one kernel row per CTA and half
    // the CTA slots): 16^2 x256: 17.8 -> 18.0, 32^2 x128: 17.1 -> 18.8, 64^2 x64: 20.4 -> 19.9, 128^2 x32: 32.1 -> 24.5,
    // 128^2 32->64: 34.0 -> 26.3.  Both are bound by the split-K reduction (the halo kernel's CTAs each carry 3 taps), so the
    // halo kernel is used where the pixel axis is long (>= 1024 tiles of 128 pixels); CRACKFL_WGRAD_HALO=2 forces it.
    const long px_tiles = (long)B * ((H + 15) / 16) * ((W + 7) / 8);
    if (use_halo_w && taps == 9 && mode == 0 && !workspace && H >= 16 && W >= 8 && (cin_valid <= 0 || cin_valid == Cin) &&
        (px_tiles >= 1024 || use_halo_w >= 2))
      return wgrad_halo(X, dY, dW, B, H, W, Cin, Cout, s_tap, s_ci, s_co, st);
  }
  WgradArgs a;
  a.dW = dW; a.s_tap = s_tap; a.s_ci = s_ci; a.s_co = s_co;
  a.B = B; a.H = H; a.W = W; a.Cin = Cin; a.Cout = Cout; a.taps = taps;
  a.cw = (Cin % 64 == 0) ? 64 : 32;
  a.nw = (Cout % 64 == 0) ? 64 : 32;
  a.BN = Cout > 128 ? 128 : Cout;
  if (Cout % a.BN) return (int)cudaErrorInvalidValue;
  a.mblk = 128 / a.cw;
  a.nblk = a.BN / a.nw;
  a.layoutA = a.cw == 64 ? 2u : 4u; a.sboA = a.cw == 64 ? 1024u : 512u;
  a.layoutB = a.nw == 64 ? 2u : 4u; a.sboB = a.nw == 64 ? 1024u : 512u;
  a.g = make_geom(B, H, W);
  a.total_blocks = taps * (Cin / a.cw);
  const int groups = (a.total_blocks + a.mblk - 1) / a.mblk;
  const int ntn = Cout / a.BN;
  a.ntiles = a.g.tiles_w * a.g.tiles_h * a.g.tiles_b;
  // one wave of CTAs: two per SM when two pipelines of >= 2 stages fit in shared memory, else one per SM with
  // up to 3 stages; few, long-running CTAs keep the fp32 atomics of the epilogue (one per output element per
  // CTA) from serialising in L2
  const uint32_t stage_bytes = (a.total_blocks == 1 ? 128u * (uint32_t)a.cw * 2u : 128u * 128u * 2u) + 128u * (uint32_t)a.BN * 2u;
  int per_sm = (2u * stage_bytes <= 106496u) ? 2 : 1;
  {
    static int dbg_persm = -1, dbg_noat = -1;
    if (dbg_persm < 0) { const char* e = getenv("CRACKFL_WGRAD_PERSM"); dbg_persm = e ? atoi(e) : 0; }
    if (dbg_noat < 0) { const char* e = getenv("CRACKFL_WGRAD_NOATOMIC"); dbg_noat = e ? atoi(e) : 0; }
    if (dbg_persm == 1) per_sm = 1;
    a.dbg_noatomic = dbg_noat;
  }
  int stages = per_sm == 2 ? (int)(106496u / stage_bytes) : (int)(200704u / stage_bytes);
  if (stages > 6) stages = 6;
  static int fill = -1;  // tuning knob: percentage of the 148 * per_sm CTA slots the split count may fill
  if (fill < 0) { const char* e = getenv("CRACKFL_WGRAD_FILL"); fill = e ? atoi(e) : 0; if (fill < 10 || fill > 100) fill = 0; }
  // default: every slot when the partial tiles go to a workspace (one reduction launch sums them), half of them when they
  // are added to the arena with vector atomics (fewer, longer CTAs = fewer contended reductions per element and more room
  // for the main chain; measured on the full step at the 1.03 ms build: 1.029 / 1.025 / 1.028 / 1.027 / 1.031 / 1.068 ms
  // at 40 / 50 / 60 / 67 / 75 / 90 %)
  const int fill_eff = fill ? fill : ((!workspace && mode == 0) ? 50 : 100);
  int splits = (148 * per_sm * fill_eff / 100) / (groups * ntn);
  if (splits > a.ntiles) splits = a.ntiles;
  if (splits < 1) splits = 1;
  a.tiles_per_split = (a.ntiles + splits - 1) / splits;
  splits = (a.ntiles + a.tiles_per_split - 1) / a.tiles_per_split;
  if (stages > a.tiles_per_split) stages = a.tiles_per_split;
  if (stages < 1) stages = 1;
  a.stages = stages;
  // optional in-cluster split reduction (CRACKFL_WGRAD_CLUSTER=2|4|8): CS CTAs along the split axis sum their tiles
  // through distributed shared memory and add them to dW with vector atomics, no workspace, no reduction launch.
  // Measured on the full step: 1.421 / 1.433 / 1.442 ms for CS = 2 / 4 / 8 against 1.412 ms for the workspace +
  // reduction-kernel path (the 37-150-way contended atomics and the cluster barriers cost more than the launches they
  // save), so it stays off by default.
  static int use_cluster = -1;
  if (use_cluster < 0) { const char* e = getenv("CRACKFL_WGRAD_CLUSTER"); use_cluster = e ? atoi(e) : 0; }
  int CS = 1;
  if (use_cluster > 1 && mode == 0) {
    CS = use_cluster > 8 ? 8 : use_cluster;
    while (CS > 1 && (splits < CS || (CS & (CS - 1)))) CS >>= 1;
    if (CS > 1) {
      splits = splits / CS * CS;
      a.tiles_per_split = (a.ntiles + splits - 1) / splits;
      if (stages > a.tiles_per_split) stages = a.tiles_per_split;
      if (stages < 1) stages = 1;
      a.stages = stages;
    }
  }
  a.cluster = CS;
  a.cin_valid = cin_valid > 0 ? cin_valid : Cin;
  size_t smem = (size_t)stages * stage_bytes + 1024 + 256;
  if (CS > 1 && smem < (size_t)128 * (a.BN + 4) * 4 + 1024 + 256) smem = (size_t)128 * (a.BN + 4) * 4 + 1024 + 256;
  const long need = (long)groups * ntn * splits * 128L * a.BN;
  int zl = 1;
  {
    const long total4 = (long)groups * ntn * 128L * a.BN / 4;
    while (zl * 2 <= splits && zl < 32 && total4 * (zl * 2) / 256 <= 148 * wred_cap()) zl *= 2;  // split lanes per block
  }
  if (desc) {
    desc->s_tap = s_tap; desc->s_ci = s_ci; desc->s_co = s_co; desc->need = need;
    desc->groups = groups; desc->ntn = ntn; desc->splits = splits; desc->BN = a.BN; desc->cw = a.cw; desc->mblk = a.mblk;
    desc->total_blocks = a.total_blocks; desc->cchunks = Cin / a.cw; desc->cin_valid = cin_valid > 0 ? cin_valid : Cin;
    desc->zl = zl;
  }
  if (mode == 2) return 0;
  CUtensorMap tmX, tmDY;
  const int dX[4] = {Cin, W, H, B}, bX[4] = {a.cw, a.g.TW, a.g.TH, a.g.TB};
  int rc = get_map(X, 4, dX, bX, a.cw * 2, &tmX);
  if (rc) return rc;
  const int dD[4] = {Cout, W, H, B}, bD[4] = {a.nw, a.g.TW, a.g.TH, a.g.TB};
  rc = get_map(dY, 4, dD, bD, a.nw * 2, &tmDY);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_tc_k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  dim3 grid(groups, ntn, splits);
  a.ws = (workspace && workspace_floats >= need) ? workspace : nullptr;
  if (CS > 1) {
    a.ws = nullptr;
    cudaError_t e = ck::launch_cluster_z(wgrad_tc_k, grid, 192, smem, st, CS, tmX, tmDY, a);
    if (e != cudaSuccess) return (int)e;
    CK_LAUNCH_CHECK();
    return 0;
  }
  if (!a.ws && mode == 1) return (int)cudaErrorInvalidValue;
  ck::launch(wgrad_tc_k, grid, 192, smem, st, tmX, tmDY, a);
  CK_LAUNCH_CHECK();
  if (a.ws && mode == 0) {
    const long total4 = (long)groups * ntn * 128L * a.BN / 4;
    long blocks = (total4 * zl + 255) / 256;
    if (blocks > 148 * wred_cap()) blocks = 148 * wred_cap();
    ck::launch(wgrad_reduce_k, (int)blocks, 256, 0, st, a.ws, dW, groups, ntn, splits, a.BN, a.cw, a.mblk, a.total_blocks, Cin / a.cw,
               s_tap, s_ci, s_co, zl, cin_valid > 0 ? cin_valid : Cin);
    CK_LAUNCH_CHECK();
  }
  return 0;
}

int crackfl_wgrad_tc_ws_cv(const void* X, const void* dY, float* dW, int B, int H, int W, int Cin, int Cout, int taps,
                           long s_tap, long s_ci, long s_co, float* workspace, long workspace_floats, int cin_valid,
                           cudaStream_t st) {
  return wgrad_impl(X, dY, dW, B, H, W, Cin, Cout, taps, s_tap, s_ci, s_co, workspace, workspace_floats, cin_valid, 0, nullptr, st);
}

int crackfl_wgrad_plan(int B, int H, int W, int Cin, int Cout, int taps, long s_tap, long s_ci, long s_co, int cin_valid,
                       crackfl_wgrad_red_desc* desc) {
  return wgrad_impl(nullptr, nullptr, nullptr, B, H, W, Cin, Cout, taps, s_tap, s_ci, s_co, nullptr, 0, cin_valid, 2, desc, 0);
}

int crackfl_wgrad_tc_deferred(const void* X, const void* dY, int B, int H, int W, int Cin, int Cout, int taps, float* ws_layer,
                              long ws_floats, cudaStream_t st) {
  return wgrad_impl(X, dY, nullptr, B, H, W, Cin, Cout, taps, 0, 0, 0, ws_layer, ws_floats, 0, 1, nullptr, st);
}

int crackfl_wgrad_reduce_all(const crackfl_wgrad_red_desc* descs_dev, int n, const float* ws_base, float* dw_base,
                             cudaStream_t st) {
  if (n <= 0) return 0;
  static int gx = -1;
  if (gx < 0) { const char* e = getenv("CRACKFL_WRED_GRID"); gx = e ? atoi(e) : 592; }
  dim3 grid(gx, n);
  ck::launch(wgrad_reduce_all_k, grid, 256, 0, st, descs_dev, ws_base, dw_base);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_wgrad_tc_ws(const void* X, const void* dY, float* dW, int B, int H, int W, int Cin, int Cout, int taps,
                        long s_tap, long s_ci, long s_co, float* workspace, long workspace_floats, cudaStream_t st) {
  return crackfl_wgrad_tc_ws_cv(X, dY, dW, B, H, W, Cin, Cout, taps, s_tap, s_ci, s_co, workspace, workspace_floats, 0, st);
}

int crackfl_wgrad_tc(const void* X, const void* dY, float* dW, int B, int H, int W, int Cin, int Cout, int taps, long s_tap,
                     long s_ci, long s_co, cudaStream_t st) {
  return crackfl_wgrad_tc_ws(X, dY, dW, B, H, W, Cin, Cout, taps, s_tap, s_ci, s_co, nullptr, 0, st);
}

}  // extern "C"
