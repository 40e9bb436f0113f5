// Optimiser-side kernels over the flat parameter arena (SURVEY section 8 rows K10 and the weight re-layout).
//   * Keras-form Adam (client_fit_model.py:157 optimizer="Adam"): theta -= lr_t * m / (sqrt(v) + eps),
//     lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t), eps = 1e-7 outside the bias correction; zero-grad fused.
//   * weight_prep: fp32 Keras-layout master weights -> bf16 K-major GEMM operands for conv_tc/conv_simt,
//     one table-driven launch for every operand of the model.
#include "common.cuh"
#include "../include/crackfl.h"

namespace {

__global__ void adam_tick_k(int* __restrict__ step) {
  ck::pdl_enter(); *step += 1; }

// tail != NULL: the step counter tick and the metric accumulation ride along (two tiny launches less per step): every
// block uses t = *step + 1, block 0 adds the step's metric sums to the running totals, and the LAST block to finish
// (counter tail[0], zeroed with the step's scratch region) publishes the new step count.
__global__ void __launch_bounds__(256) adam_k(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m,
                                              float* __restrict__ v, int* __restrict__ step, long n4, float lr,
                                              float b1, float b2, float eps, unsigned* __restrict__ tail,
                                              float* __restrict__ total, const float* __restrict__ acc, int nacc,
                                              float* __restrict__ last_acc, float* __restrict__ zbase, int zn, int tplus) {
  ck::pdl_enter();
  __shared__ float s_lrt;
  __shared__ int s_t;
  __shared__ int s_last;
  if (threadIdx.x == 0) {
    s_t = *step + ((tail || tplus) ? 1 : 0);
    const float t = (float)s_t;
    s_lrt = lr * sqrtf(1.f - powf(b2, t)) / (1.f - powf(b1, t));
  }
  __syncthreads();
  if (tail && blockIdx.x == 0 && (int)threadIdx.x < nacc) total[threadIdx.x] += acc[threadIdx.x];
  const float lrt = s_lrt;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = reinterpret_cast<float4*>(g)[i];
    float4 mv = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
#define CK_ADAM1(c)                                  \
  mv.c = b1 * mv.c + (1.f - b1) * gv.c;              \
  vv.c = b2 * vv.c + (1.f - b2) * gv.c * gv.c;       \
  pv.c -= lrt * mv.c / (sqrtf(vv.c) + eps);
    CK_ADAM1(x) CK_ADAM1(y) CK_ADAM1(z) CK_ADAM1(w)
#undef CK_ADAM1
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(m)[i] = mv;
    reinterpret_cast<float4*>(v)[i] = vv;
    reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (tail) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      s_last = (atomicAdd(tail, 1u) == gridDim.x - 1);
      if (s_last) *step = s_t;  // every block has read the old count by now
    }
    if (zbase) {
      // the last block to finish leaves the per-step scratch region (BN sums, loss accumulators, this counter) zeroed
      // for the next step - no separate fill launch - after saving the step's loss/metric sums for the host to read
      __syncthreads();
      if (s_last) {
        __threadfence();
        if ((int)threadIdx.x < nacc) last_acc[threadIdx.x] = __ldcg(acc + threadIdx.x);
        __syncthreads();
        for (int i = threadIdx.x; i < zn; i += blockDim.x) zbase[i] = 0.f;
      }
    }
  }
}

// out[dst + (n*T + t)*Kd + k] = bf16(arena[src + n*s_n + t*s_t + k*s_k]), Kd = pad_ ? pad_ : K.  Exactly one of s_n / s_k is 1 for every
// operand of the model; when it is s_n (the source is contiguous along n but the operand wants k contiguous) the
// 32x32 tile goes through shared memory so that both the fp32 reads and the bf16 writes are coalesced.
__global__ void __launch_bounds__(256) weight_prep_k(const float* __restrict__ arena, __nv_bfloat16* __restrict__ ops,
                                                     const crackfl_prep_desc* __restrict__ descs) {
  ck::pdl_enter();
  __shared__ float tile[32][33];
  const crackfl_prep_desc d = descs[blockIdx.y];
  const int tn = (d.N + 31) / 32, tk = (d.K + 31) / 32;
  const int Kd = d.pad_ ? d.pad_ : d.K;  // row stride of the operand (K padded up, the pad columns stay zero)
  const int tiles = tn * tk * d.T;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  for (int tl = blockIdx.x; tl < tiles; tl += gridDim.x) {
    const int kt = tl % tk, nt = (tl / tk) % tn, t = tl / (tk * tn);
    const int n0 = nt * 32, k0 = kt * 32;
    if (d.s_n == 1 && d.s_k != 1) {
      // read with n fastest, write with k fastest
      for (int r = ty; r < 32; r += 8) {
        const int n = n0 + tx, k = k0 + r;
        tile[r][tx] = (n < d.N && k < d.K) ? arena[d.src_off + (long)n * d.s_n + (long)t * d.s_t + (long)k * d.s_k] : 0.f;
      }
      __syncthreads();
      for (int r = ty; r < 32; r += 8) {
        const int n = n0 + r, k = k0 + tx;
        if (n < d.N && k < d.K) ops[d.dst_off + ((long)n * d.T + t) * Kd + k] = __float2bfloat16_rn(tile[tx][r]);
      }
      __syncthreads();
    } else {
      for (int r = ty; r < 32; r += 8) {
        const int n = n0 + r, k = k0 + tx;
        if (n < d.N && k < d.K)
          ops[d.dst_off + ((long)n * d.T + t) * Kd + k] =
              __float2bfloat16_rn(arena[d.src_off + (long)n * d.s_n + (long)t * d.s_t + (long)k * d.s_k]);
      }
    }
  }
}

// One 32x32 tile per block over ALL operands (tile_prefix[i] = first flat tile of operand i, tile_prefix[num] = total):
// weight_prep_k walks up to 12 tiles per block with two barriers each for the 256x9x256 decoder kernels while most of its
// 48 x 43 blocks find nothing to do; here every tile of the step is in flight at once.
__global__ void __launch_bounds__(256) weight_prep_flat_k(const float* __restrict__ arena, __nv_bfloat16* __restrict__ ops,
                                                          const crackfl_prep_desc* __restrict__ descs,
                                                          const int* __restrict__ tile_prefix, int num_desc) {
  ck::pdl_enter();
  __shared__ float tile[32][33];
  __shared__ int pref[129];
  for (int i = threadIdx.x; i <= num_desc; i += 256) pref[i] = tile_prefix[i];
  __syncthreads();
  int lo = 0, hi = num_desc - 1;  // last operand whose first tile is <= blockIdx.x
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (pref[mid] <= (int)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const crackfl_prep_desc d = descs[lo];
  const int tl = (int)blockIdx.x - pref[lo];
  const int tn = (d.N + 31) / 32, tk = (d.K + 31) / 32;
  const int Kd = d.pad_ ? d.pad_ : d.K;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const int kt = tl % tk, nt = (tl / tk) % tn, t = tl / (tk * tn);
  const int n0 = nt * 32, k0 = kt * 32;
  const float* src = arena + d.src_off + (long)t * d.s_t;
  __nv_bfloat16* dst = ops + d.dst_off + (long)t * Kd;
  if (d.s_n == 1 && d.s_k != 1) {  // read with n fastest, write with k fastest
    float v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int n = n0 + tx, k = k0 + ty + 8 * i;
      v[i] = (n < d.N && k < d.K) ? src[n + (long)k * d.s_k] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) tile[ty + 8 * i][tx] = v[i];
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int n = n0 + ty + 8 * i, k = k0 + tx;
      if (n < d.N && k < d.K) dst[(long)n * d.T * Kd + k] = __float2bfloat16_rn(tile[tx][ty + 8 * i]);
    }
  } else {
    float v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int n = n0 + ty + 8 * i, k = k0 + tx;
      v[i] = (n < d.N && k < d.K) ? src[(long)n * d.s_n + (long)k * d.s_k] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int n = n0 + ty + 8 * i, k = k0 + tx;
      if (n < d.N && k < d.K) dst[(long)n * d.T * Kd + k] = __float2bfloat16_rn(v[i]);
    }
  }
}

__global__ void __launch_bounds__(256) f32_to_bf16_k(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long n) {
  ck::pdl_enter();
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    out[i] = __float2bfloat16_rn(in[i]);
}

// K16 staging: client update (fp32, just copied H2D or just trained) -> bf16 delta against the global model
__global__ void __launch_bounds__(256) publish_delta_k(const float* __restrict__ local, const float* __restrict__ global,
                                                       __nv_bfloat16* __restrict__ out, long n) {
  ck::pdl_enter();
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    out[i] = __float2bfloat16_rn(local[i] - global[i]);
}

// Depthwise 3x3 kernels (Keras (3,3,C,1), fp32) -> block-diagonal bf16 weight tiles for the tensor-core depthwise path
// (conv_tc_xf_k, diag): out[((chunk*9 + t)*kc + n)*kc + k] = (n == k) ? w[tp*C + chunk*kc + n] : 0, tp = flip ? 8-t : t.
// Only the diagonal is written; the buffer is zeroed once when it is allocated.
__global__ void __launch_bounds__(256) dw_diag_prep_k(const float* __restrict__ arena, __nv_bfloat16* __restrict__ ops,
                                                      const crackfl_dw_diag_desc* __restrict__ descs) {
  ck::pdl_enter();
  const crackfl_dw_diag_desc d = descs[blockIdx.y];
  const int total = 9 * d.C;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int t = i / d.C, c = i - t * d.C;
    const int chunk = c / d.kc, n = c - chunk * d.kc;
    const int tp = d.flip ? 8 - t : t;
    ops[d.dst_off + ((long)(chunk * 9 + t) * d.kc + n) * d.kc + n] = __float2bfloat16_rn(arena[d.src_off + (long)tp * d.C + c]);
  }
}

__global__ void accum_k(float* __restrict__ total, const float* __restrict__ step, int n) {
  ck::pdl_enter();
  if (threadIdx.x < n) total[threadIdx.x] += step[threadIdx.x];
}

}  // namespace

static int opt_cap() {  // blocks per SM of the flat-arena kernels (tuning knob)
  static int v = -1;
  if (v < 0) { const char* e = getenv("CRACKFL_OPT_BLOCKS_PER_SM"); v = e ? atoi(e) : 8; if (v < 1) v = 8; }
  return v;
}

extern "C" {

int crackfl_accum_small(float* total, const float* step, int n, cudaStream_t st) {
  if (n > 32) return (int)cudaErrorInvalidValue;
  ck::launch(accum_k, 1, 32, 0, st, total, step, n);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_publish_delta(const float* local, const float* global, void* out_bf16, long n, cudaStream_t st) {
  long blocks = (n + 255) / 256;
  if (blocks > 148 * opt_cap()) blocks = 148 * opt_cap();
  ck::launch(publish_delta_k, (int)blocks, 256, 0, st, local, global, (__nv_bfloat16*)out_bf16, n);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_adam_step(float* p, float* g, float* m, float* v, int* step, long n, float lr, float b1, float b2, float eps,
                      cudaStream_t st) {
  if (n % 4) return (int)cudaErrorInvalidValue;
  ck::launch(adam_tick_k, 1, 1, 0, st, step);
  const long n4 = n / 4;
  long blocks = (n4 + 255) / 256;
  if (blocks > 148 * opt_cap()) blocks = 148 * opt_cap();
  ck::launch(adam_k, (int)blocks, 256, 0, st, p, g, m, v, step, n4, lr, b1, b2, eps, (unsigned*)nullptr, (float*)nullptr,
             (const float*)nullptr, 0, (float*)nullptr, (float*)nullptr, 0, 0);
  CK_LAUNCH_CHECK();
  return 0;
}

// Adam over a sub-range of the arena BEFORE the step's crackfl_adam_step_tail launch (which covers the rest and publishes
// the new step count): uses t = *step + 1 and leaves the counter alone.  The engine updates the decoder's parameters
// this way as soon as the decoder's gradients are complete, next to the encoder's backward pass.
int crackfl_adam_part(float* p, float* g, float* m, float* v, const int* step, long n, float lr, float b1, float b2, float eps,
                      cudaStream_t st) {
  if (n % 4 || n <= 0) return (int)cudaErrorInvalidValue;
  const long n4 = n / 4;
  long blocks = (n4 + 255) / 256;
  if (blocks > 148 * opt_cap()) blocks = 148 * opt_cap();
  ck::launch(adam_k, (int)blocks, 256, 0, st, p, g, m, v, const_cast<int*>(step), n4, lr, b1, b2, eps, (unsigned*)nullptr,
             (float*)nullptr, (const float*)nullptr, 0, (float*)nullptr, (float*)nullptr, 0, 1);
  CK_LAUNCH_CHECK();
  return 0;
}

// Adam + step tick + metric accumulation in one launch; tail_counter must be zero on entry (it lives in the per-step
// scratch region).  With zero_base != NULL the launch also ends the step: step_acc is saved to last_acc[nacc] and the
// zero_n floats at zero_base (the whole per-step scratch region, counter included) are cleared for the next step.
int crackfl_adam_step_tail(float* p, float* g, float* m, float* v, int* step, long n, float lr, float b1, float b2, float eps,
                           unsigned* tail_counter, float* metrics_total, const float* step_acc, int nacc, float* last_acc,
                           float* zero_base, int zero_n, cudaStream_t st) {
  if (n % 4 || nacc > 32 || !tail_counter || (zero_base && !last_acc)) return (int)cudaErrorInvalidValue;
  const long n4 = n / 4;
  long blocks = (n4 + 255) / 256;
  if (blocks > 148 * opt_cap()) blocks = 148 * opt_cap();
  ck::launch(adam_k, (int)blocks, 256, 0, st, p, g, m, v, step, n4, lr, b1, b2, eps, tail_counter, metrics_total, step_acc, nacc,
             last_acc, zero_base, zero_n, 1);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_weight_prep(const float* arena, void* ops, const crackfl_prep_desc* descs_dev, int num_desc, cudaStream_t st) {
  static int gx = -1;  // tuning knob; 192: <= 3 tiles of 32x32 per block for the largest operand (256 x 9 x 256)
  if (gx < 0) { const char* e = getenv("CRACKFL_PREP_GRID_X"); gx = e ? atoi(e) : 48; if (gx < 1) gx = 48; }  // measured: 48: 1.279, 96: 1.282, 192: 1.286 ms/step (it overlaps the entry kernels on the auxiliary stream)
  dim3 grid(gx, num_desc);
  ck::launch(weight_prep_k, grid, 256, 0, st, arena, (__nv_bfloat16*)ops, descs_dev);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_weight_prep_flat(const float* arena, void* ops, const crackfl_prep_desc* descs_dev, const int* tile_prefix_dev,
                             int num_desc, int total_tiles, cudaStream_t st) {
  if (num_desc < 1 || num_desc > 128 || total_tiles < 1) return (int)cudaErrorInvalidValue;
  ck::launch(weight_prep_flat_k, total_tiles, 256, 0, st, arena, (__nv_bfloat16*)ops, descs_dev, tile_prefix_dev, num_desc);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_dw_diag_prep(const float* arena, void* ops, const crackfl_dw_diag_desc* descs_dev, int num_desc, cudaStream_t st) {
  if (num_desc <= 0) return 0;
  dim3 grid(9, num_desc);
  ck::launch(dw_diag_prep_k, grid, 256, 0, st, arena, (__nv_bfloat16*)ops, descs_dev);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_f32_to_bf16(const float* in, void* out, long n, cudaStream_t st) {
  long blocks = (n + 255) / 256;
  if (blocks > 148 * opt_cap()) blocks = 148 * opt_cap();
  ck::launch(f32_to_bf16_k, (int)blocks, 256, 0, st, in, (__nv_bfloat16*)out, n);
  CK_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
