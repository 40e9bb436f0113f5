// Fused federated aggregation: ONE kernel that
//   (1) waits for every rank's update to be published      (signal pads, st.release.sys / ld.acquire.sys)
//   (2) for its shard of the flat arena pulls every client's update with 128-bit loads
//       (local HBM or NVLink peer memory: the pointers decide), scales by the per-client coefficient
//       n_k / sum n, casts bf16->fp32 and accumulates in fp32 registers
//   (3) applies the server update on the fp32 master shard: FedAvg (w = sum) or FedAdam
//   (4) stores the new global shard straight into every destination arena (each client's live
//       parameter arena on every GPU = "broadcast fused with the weight reload"), either as peer
//       stores or as ONE multimem.st to an NVLS multicast address (the switch replicates)
//   (5) signals completion to the peers and waits for theirs.
// It replaces the reference's gather (N gRPC uploads, fl_client.py:61-65 -> fl_server.py:176-179),
// NumPy reduce (fl_server.py:92-102), N unicast downloads (fl_server.py:190,201) and the 20 s poll
// barrier (fl_client.py:136-152).  No host round trip, no NCCL call on this path.
//
// With world == 1 and K local client buffers the same kernel is the per-GPU pre-reduce (K14) and
// the single-GPU emulation of a multi-rank round used by the 1-GPU tests.
#include "common.cuh"
#include "../include/crackfl.h"

namespace {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_nc_f4(const float* p) {
  float4 r;
  asm volatile("ld.global.relaxed.sys.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ld_nc_u2(const void* p) {
  uint2 r;
  asm volatile("ld.global.relaxed.sys.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ float4 ld_stream_f4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ld_stream_u2(const void* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// NVLS: one load that the NVSwitch turns into a reduction over every GPU's replica of the address
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc) : "memory");
  return r;
}
// 4 bf16 per load, accumulated in fp32 inside the switch
__device__ __forceinline__ uint2 multimem_ld_reduce_bf16x4(const void* mc) {
  uint2 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v2.bf16x2 {%0,%1}, [%2];"
               : "=r"(r.x), "=r"(r.y) : "l"(mc) : "memory");
  return r;
}

// Spin limit when the caller passes timeout_cycles == 0: ~60 s at 1.9 GHz.  A dead rank still cannot hang the box,
// and ranks that reach the round boundary seconds apart (uneven client data, K clients per GPU, a JPEG-bound
// loader) no longer trip it; callers that rendezvous on the host first can pass something much shorter.
constexpr long long DEFAULT_TIMEOUT_CYCLES = 120000000000LL;

__device__ __forceinline__ long long spin_limit(const crackfl_fedagg_params& P) {
  return P.timeout_cycles > 0 ? P.timeout_cycles : DEFAULT_TIMEOUT_CYCLES;
}

// Cross-GPU barrier over the signal pads.  pad layout per rank: [phase(2)][world] u32, monotonically
// increasing epoch values (no reset race).  One thread per peer.
__device__ void pad_barrier(const crackfl_fedagg_params& P, int phase) {
  const int j = threadIdx.x;
  if (j < P.world) {
    st_release_sys(P.pads[j] + phase * CRACKFL_MAX_RANKS + P.rank, P.epoch);
    const uint32_t* mine = P.pads[P.rank] + phase * CRACKFL_MAX_RANKS + j;
    const long long t0 = clock64(), lim = spin_limit(P);
    while (ld_acquire_sys(mine) < P.epoch) {
      if (clock64() - t0 > lim) {
        atomicExch(P.error_flag, 1 + phase);
        break;
      }
    }
  }
}

template <bool PAYLOAD_BF16, bool OUT_BF16, int U>
__global__ void __launch_bounds__(256) fedagg_k(const crackfl_fedagg_params P) {
  if (P.world > 1) {
    // (1) enter: block 0 does the cross-GPU handshake and then releases the local grid
    if (blockIdx.x == 0) {
      pad_barrier(P, 0);
      __syncthreads();
      if (threadIdx.x == 0) st_release_gpu(P.local_sync, P.epoch);
    } else if (threadIdx.x == 0) {
      const long long t0 = clock64(), lim = 2 * spin_limit(P);
      while (ld_acquire_gpu(P.local_sync) < P.epoch) {
        if (clock64() - t0 > lim) { atomicExch(P.error_flag, 3); break; }
      }
    }
    __syncthreads();
  }

  const long n4_begin = P.begin >> 2, n4_end = P.end >> 2;
  // U independent 16-byte loads per client per thread: NVLink latency is ~2 us, keep bytes in flight
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i0 = n4_begin + (long)blockIdx.x * blockDim.x + threadIdx.x; i0 < n4_end; i0 += stride * U) {
    float4 acc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (P.mc_in) {
      // NVLS pull: the partials were pre-scaled by their coefficients, the switch adds the G replicas
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long i = i0 + u * stride;
        if (i >= n4_end) continue;
        if (PAYLOAD_BF16) {
          const uint2 w4 = multimem_ld_reduce_bf16x4(reinterpret_cast<const uint2*>(P.mc_in) + i);
          const uint32_t lo = w4.x, hi = w4.y;
          acc[u].x = __uint_as_float(lo << 16); acc[u].y = __uint_as_float(lo & 0xffff0000u);
          acc[u].z = __uint_as_float(hi << 16); acc[u].w = __uint_as_float(hi & 0xffff0000u);
        } else {
          acc[u] = multimem_ld_reduce_f4(reinterpret_cast<const float*>(P.mc_in) + i * 4);
        }
      }
    } else
    for (int k = 0; k < P.num_clients; ++k) {
      const float c = P.coef[k];
      if (PAYLOAD_BF16) {
        uint2 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long i = i0 + u * stride;
          v[u] = i < n4_end ? (P.load_mode ? ld_stream_u2(reinterpret_cast<const uint2*>(P.client[k]) + i)
                                           : ld_nc_u2(reinterpret_cast<const uint2*>(P.client[k]) + i))
                             : make_uint2(0u, 0u);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          acc[u].x = fmaf(c, __uint_as_float(v[u].x << 16), acc[u].x);
          acc[u].y = fmaf(c, __uint_as_float(v[u].x & 0xffff0000u), acc[u].y);
          acc[u].z = fmaf(c, __uint_as_float(v[u].y << 16), acc[u].z);
          acc[u].w = fmaf(c, __uint_as_float(v[u].y & 0xffff0000u), acc[u].w);
        }
      } else {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long i = i0 + u * stride;
          v[u] = i < n4_end ? (P.load_mode ? ld_stream_f4(reinterpret_cast<const float*>(P.client[k]) + i * 4)
                                           : ld_nc_f4(reinterpret_cast<const float*>(P.client[k]) + i * 4))
                             : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          acc[u].x = fmaf(c, v[u].x, acc[u].x); acc[u].y = fmaf(c, v[u].y, acc[u].y);
          acc[u].z = fmaf(c, v[u].z, acc[u].z); acc[u].w = fmaf(c, v[u].w, acc[u].w);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long i = i0 + u * stride;
      if (i >= n4_end) break;
      float4 nw = acc[u];
      if (P.master) {
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
        if (P.delta || P.server_opt == 1) w = reinterpret_cast<float4*>(P.master)[i];  // plain FedAvg overwrites it
        if (P.delta) { nw.x += w.x; nw.y += w.y; nw.z += w.z; nw.w += w.w; }
        if (P.server_opt == 1) {
          float4 m = reinterpret_cast<float4*>(P.m)[i];
          float4 v = reinterpret_cast<float4*>(P.v)[i];
          const uchar4 tr = P.trainable ? reinterpret_cast<const uchar4*>(P.trainable)[i] : make_uchar4(1, 1, 1, 1);
#define CK_FEDADAM1(c)                                                   \
  if (tr.c) {                                                            \
    const float d = nw.c - w.c;                                          \
    m.c = P.beta1 * m.c + (1.f - P.beta1) * d;                           \
    v.c = P.beta2 * v.c + (1.f - P.beta2) * d * d;                       \
    nw.c = w.c + P.eta * m.c / (sqrtf(v.c) + P.tau);                     \
  }
          CK_FEDADAM1(x) CK_FEDADAM1(y) CK_FEDADAM1(z) CK_FEDADAM1(w)
#undef CK_FEDADAM1
          reinterpret_cast<float4*>(P.m)[i] = m;
          reinterpret_cast<float4*>(P.v)[i] = v;
        }
        reinterpret_cast<float4*>(P.master)[i] = nw;
      }
      if (P.mc_out) {
        multimem_st_f4(P.mc_out + i * 4, nw);
      } else if (OUT_BF16) {
        uint2 o;
        o.x = ck::pack2(nw.x, nw.y);
        o.y = ck::pack2(nw.z, nw.w);
        for (int r = 0; r < P.num_out; ++r) reinterpret_cast<uint2*>(P.out[r])[i] = o;
      } else {
        for (int r = 0; r < P.num_out; ++r) reinterpret_cast<float4*>(P.out[r])[i] = nw;
      }
    }
  }

  if (P.world > 1) {
    // (5) exit: every block's stores are fenced system-wide, the last block to arrive handshakes
    __threadfence_system();
    __syncthreads();
    __shared__ int s_last;
    // arrival counter of THIS launch: the last block resets it, so a round that timed out, or a different grid size
    // on the next call, leaves nothing behind (every block gets here, also after a spin time-out)
    if (threadIdx.x == 0) {
      const unsigned prev = atomicAdd(P.local_sync + 1, 1u);
      s_last = (prev + 1 == gridDim.x);
      if (s_last) atomicExch(P.local_sync + 1, 0u);
    }
    __syncthreads();
    if (s_last) pad_barrier(P, 1);
  }
}

}  // namespace

extern "C" {

int crackfl_fedagg(const crackfl_fedagg_params* p, int grid_blocks, cudaStream_t st) {
  if ((p->num_clients < 1 && !p->mc_in) || p->num_clients > CRACKFL_MAX_CLIENTS || p->num_out > CRACKFL_MAX_OUT ||
      p->world > CRACKFL_MAX_RANKS || (p->begin & 3) || (p->end & 3) || p->end < p->begin)
    return (int)cudaErrorInvalidValue;
  if (p->mc_out && p->out_bf16) return (int)cudaErrorInvalidValue;
  long n4 = (p->end - p->begin) / 4;
  long blocks = (n4 + 256 * 4 - 1) / (256 * 4);
  if (p->unroll == 8) blocks = (n4 + 256 * 8 - 1) / (256 * 8);
  if (blocks > grid_blocks) blocks = grid_blocks;
  if (blocks < 1) blocks = 1;
  // (every block of a multi-rank launch spins on block 0's release, so all of them must be resident at once:
  //  grid_blocks is at most 8 blocks of 256 threads per SM; small shards simply use fewer blocks)
  const int g = (int)blocks;
  const bool u8 = p->unroll == 8;
  if (p->payload_bf16) {
    if (p->out_bf16) { if (u8) fedagg_k<true, true, 8><<<g, 256, 0, st>>>(*p); else fedagg_k<true, true, 4><<<g, 256, 0, st>>>(*p); }
    else { if (u8) fedagg_k<true, false, 8><<<g, 256, 0, st>>>(*p); else fedagg_k<true, false, 4><<<g, 256, 0, st>>>(*p); }
  } else {
    if (p->out_bf16) { if (u8) fedagg_k<false, true, 8><<<g, 256, 0, st>>>(*p); else fedagg_k<false, true, 4><<<g, 256, 0, st>>>(*p); }
    else { if (u8) fedagg_k<false, false, 8><<<g, 256, 0, st>>>(*p); else fedagg_k<false, false, 4><<<g, 256, 0, st>>>(*p); }
  }
  CK_LAUNCH_CHECK();
  return 0;
}

// ---- peer-memory plumbing (legacy CUDA IPC: works between the per-GPU processes of one box) ----
int crackfl_ipc_export(void* dev_ptr, void* handle_out64) {
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, dev_ptr);
  if (e != cudaSuccess) return (int)e;
  memcpy(handle_out64, &h, sizeof(h));
  return 0;
}

int crackfl_ipc_open(const void* handle64, void** dev_ptr_out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  cudaError_t e = cudaIpcOpenMemHandle(dev_ptr_out, h, cudaIpcMemLazyEnablePeerAccess);
  return (int)e;
}

int crackfl_ipc_close(void* dev_ptr) { return (int)cudaIpcCloseMemHandle(dev_ptr); }

int crackfl_dev_malloc(void** out, size_t bytes) {
  cudaError_t e = cudaMalloc(out, bytes);
  if (e == cudaSuccess) e = cudaMemset(*out, 0, bytes);
  return (int)e;
}

int crackfl_dev_free(void* p) { return (int)cudaFree(p); }

int crackfl_enable_peer_access(int peer_device) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return 0; }
  return (int)e;
}

}  // extern "C"
