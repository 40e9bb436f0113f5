// Shared device helpers for the crackfl sm_100a kernels.
// Activations are NHWC bf16 with C a multiple of 8; the unit of access is one uint4 (8 x bf16).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#define CK_LAUNCH_CHECK() do { cudaError_t e__ = cudaGetLastError(); if (e__ != cudaSuccess) return (int)e__; } while (0)

namespace ck {

// ---- programmatic dependent launch (PDL) ------------------------------------------------------------------------------
// Every kernel of the training step is launched with programmaticStreamSerialization and starts with
// pdl_wait(); pdl_trigger();  The wait blocks until the preceding kernel in the stream has completed and flushed, so no
// global memory is touched early; what overlaps with the predecessor's tail is the launch latency, CTA scheduling and the
// per-CTA set-up before the wait (barrier init, TMEM allocation, descriptor prefetch).  The trigger lets the NEXT kernel be
// scheduled as soon as all CTAs of this one are resident.  Because every kernel waits before it exits, completion order
// stays transitive (kernel n+2 cannot pass its wait before kernel n has finished).  CRACKFL_PDL=0 turns the attribute off;
// the two instructions are then no-ops.  Measured on the full step: wait-then-trigger 1.28 ms, trigger-then-wait 1.30 ms,
// wait without an explicit trigger (dependents launch only when the grid drains) 1.34 ms.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_enter() { pdl_wait(); pdl_trigger(); }

inline int pdl_enabled() {
  static int on = -1;
  if (on < 0) { const char* e = getenv("CRACKFL_PDL"); on = e ? (atoi(e) != 0) : 1; }
  return on;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = pdl_enabled();
  cfg.attrs = at; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// same, as thread-block clusters of (1, 1, cluster_z) CTAs (cluster_z must divide grid.z)
template <typename... KArgs, typename... Args>
inline cudaError_t launch_cluster_z(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster_z,
                                    Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = pdl_enabled();
  at[1].id = cudaLaunchAttributeClusterDimension;
  at[1].val.clusterDim.x = 1; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = (unsigned)cluster_z;
  cfg.attrs = at; cfg.numAttrs = 2;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}


__device__ __forceinline__ void unpack8(const uint4& v, float f[8]) {
  f[0] = __uint_as_float(v.x << 16); f[1] = __uint_as_float(v.x & 0xffff0000u);
  f[2] = __uint_as_float(v.y << 16); f[3] = __uint_as_float(v.y & 0xffff0000u);
  f[4] = __uint_as_float(v.z << 16); f[5] = __uint_as_float(v.z & 0xffff0000u);
  f[6] = __uint_as_float(v.w << 16); f[7] = __uint_as_float(v.w & 0xffff0000u);
}

__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

__device__ __forceinline__ uint4 pack8(const float f[8]) {
  uint4 v;
  v.x = pack2(f[0], f[1]); v.y = pack2(f[2], f[3]); v.z = pack2(f[4], f[5]); v.w = pack2(f[6], f[7]);
  return v;
}

__device__ __forceinline__ uint4 ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void stg16(void* p, const uint4& v) { *reinterpret_cast<uint4*>(p) = v; }

__device__ __forceinline__ void load8f(const float* __restrict__ p, float f[8]) {
  float4 a = __ldg(reinterpret_cast<const float4*>(p));
  float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-level reduction of per-thread partial sums laid out as (pixel-lane, channel-vector):
// thread t owns channel vector (t % c8) and pixel lane (t / c8).  All lanes of a channel vector
// are summed through shared memory and lane 0 issues one atomicAdd per channel.
// `sm` must hold blockDim.x * 8 floats.
__device__ __forceinline__ void block_channel_atomic(const float acc[8], int c8, float* sm, float* __restrict__ out) {
  const int t = threadIdx.x;
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 8; ++j) sm[t * 8 + j] = acc[j];
  __syncthreads();
  const int lanes = blockDim.x / c8;
  for (int i = t; i < c8 * 8; i += blockDim.x) {
    const int cv = i >> 3, j = i & 7;
    float s = 0.f;
    for (int l = 0; l < lanes; ++l) s += sm[(l * c8 + cv) * 8 + j];
    atomicAdd(out + cv * 8 + j, s);
  }
}

}  // namespace ck
