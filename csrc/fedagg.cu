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


// ----------------------------------------------------------------------------------------------------------------
// Bulk-copy (TMA) variant of the cross-GPU launch, fp32 payload, up to 8 ranks.
//
// fedagg_k keeps each thread in a load -> wait -> add -> store sequence: a thread's remote reads and its remote writes
// never overlap, and with the few iterations a thread makes at 8-256 MB the whole grid moves through "everybody reads"
// and "everybody writes" phases, so the two directions of the links take turns (bench/p2p_probe.py: read + write
// together reach ~500 GB/s each against 750 / 710 GB/s alone).  Here the transfers are asynchronous bulk copies:
//   warp 0      producer: per 4 KB chunk of the shard, one cp.async.bulk per rank from that rank's arena (peer memory
//               over NVLink or local HBM) into a shared-memory stage, completion counted on the stage's mbarrier;
//   warps 1-4   consumers: sum the `world` copies of the chunk with their coefficients, apply the server update
//               (FedAvg / FedAdam on the local fp32 master shard), write the result to an output buffer in shared
//               memory, and one thread issues a bulk store of it to every rank's arena.
// Loads of the next stages and stores of the previous chunks are in flight while a chunk is being summed, so both
// link directions stay busy, and no thread ever blocks on a remote access.  Barriers, time-outs and the exit protocol
// are those of fedagg_k.  (No cluster multicast anywhere: every load names exactly one source.)
constexpr int TMA_CH = 1024;           // floats per chunk and rank (4 KB)
constexpr int TMA_CONSUMERS = 128;     // 4 warps x 2 float4 each = one chunk
constexpr int TMA_MAX_WORLD = 8;

__device__ __forceinline__ uint32_t fa_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void fa_mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fa_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void fa_mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool fa_mbar_wait(uint32_t bar, uint32_t parity, long long limit) {
  uint32_t done = 0;
  const long long t0 = clock64();
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) return true;
    if (clock64() - t0 > limit) return false;
  }
}
__device__ __forceinline__ void fa_bulk_load(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_smem), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void fa_bulk_store(void* dst, uint32_t src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src_smem), "r"(bytes) : "memory");
}

__global__ void __launch_bounds__(32 + TMA_CONSUMERS) fedagg_tma_k(const crackfl_fedagg_params P, int nstages) {
  extern __shared__ __align__(128) uint8_t fa_smem[];
  const uint32_t sbase = (fa_smem_u32(fa_smem) + 127u) & ~127u;
  const uint32_t stage_bytes = (uint32_t)P.world * TMA_CH * 4u;
  const uint32_t out0 = sbase + (uint32_t)nstages * stage_bytes;       // 2 output buffers of one chunk
  const uint32_t bar0 = out0 + 2u * TMA_CH * 4u;                       // full[nstages] | empty[nstages]
  const int warp = threadIdx.x >> 5;
  const long long lim = spin_limit(P);

  if (threadIdx.x == 0) {
    for (int s = 0; s < nstages; ++s) {
      fa_mbar_init(bar0 + 8u * s, 1);
      fa_mbar_init(bar0 + 8u * (nstages + s), 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (P.world > 1) {
    if (blockIdx.x == 0) {
      pad_barrier(P, 0);
      __syncthreads();
      if (threadIdx.x == 0) st_release_gpu(P.local_sync, P.epoch);
    } else if (threadIdx.x == 0) {
      const long long t0 = clock64();
      while (ld_acquire_gpu(P.local_sync) < P.epoch) {
        if (clock64() - t0 > 2 * lim) { atomicExch(P.error_flag, 3); break; }
      }
    }
  }
  __syncthreads();

  const long begin = P.begin, end = P.end;
  const long nchunks = (end - begin + TMA_CH - 1) / TMA_CH;

  if (warp == 0) {
    if ((threadIdx.x & 31) == 0) {
      // the updates were written through the generic proxy (by other GPUs, ordered by the pad barrier above); the
      // bulk copies read through the async proxy
      asm volatile("fence.proxy.async;" ::: "memory");
      uint32_t it = 0;
      for (long c = blockIdx.x; c < nchunks; c += gridDim.x, ++it) {
        const uint32_t s = it % (uint32_t)nstages, ph = (it / (uint32_t)nstages) & 1u;
        if (!fa_mbar_wait(bar0 + 8u * (nstages + s), ph ^ 1u, lim)) { atomicExch(P.error_flag, 4); break; }
        const long e0 = begin + c * TMA_CH;
        const uint32_t bytes = (uint32_t)(min((long)TMA_CH, end - e0) * 4);
        fa_mbar_expect_tx(bar0 + 8u * s, bytes * (uint32_t)P.world);
        for (int r = 0; r < P.world; ++r)
          fa_bulk_load(sbase + s * stage_bytes + (uint32_t)r * TMA_CH * 4u,
                       reinterpret_cast<const float*>(P.client[r]) + e0, bytes, bar0 + 8u * s);
      }
    }
  } else {
    const int t = threadIdx.x - 32;  // 0..127
    float cf[TMA_MAX_WORLD];
#pragma unroll
    for (int r = 0; r < TMA_MAX_WORLD; ++r) cf[r] = r < P.world ? P.coef[r] : 0.f;
    uint32_t it = 0;
    for (long c = blockIdx.x; c < nchunks; c += gridDim.x, ++it) {
      const uint32_t s = it % (uint32_t)nstages, ph = (it / (uint32_t)nstages) & 1u;
      const long e0 = begin + c * TMA_CH;
      const int n4 = (int)(min((long)TMA_CH, end - e0) >> 2);
      const uint32_t obuf = out0 + (it & 1u) * TMA_CH * 4u;
      // the bulk stores that read this output buffer two chunks ago must have finished reading it
      if (t == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      if (!fa_mbar_wait(bar0 + 8u * s, ph, lim)) { atomicExch(P.error_flag, 5); break; }
      asm volatile("bar.sync 1, %0;" ::"n"(TMA_CONSUMERS) : "memory");
      const uint32_t st0 = sbase + s * stage_bytes;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int i4 = t + h * TMA_CONSUMERS;
        if (i4 < n4) {
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int r = 0; r < TMA_MAX_WORLD; ++r) {
            if (r < P.world) {
              float4 v;
              asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                           : "r"(st0 + (uint32_t)r * TMA_CH * 4u + (uint32_t)i4 * 16u));
              acc.x = fmaf(cf[r], v.x, acc.x); acc.y = fmaf(cf[r], v.y, acc.y);
              acc.z = fmaf(cf[r], v.z, acc.z); acc.w = fmaf(cf[r], v.w, acc.w);
            }
          }
          const long i = (e0 >> 2) + i4;
          float4 nw = acc;
          if (P.master) {
            if (P.server_opt == 1) {
              const float4 w = reinterpret_cast<const float4*>(P.master)[i];
              float4 m = reinterpret_cast<float4*>(P.m)[i];
              float4 v = reinterpret_cast<float4*>(P.v)[i];
              const uchar4 tr = P.trainable ? reinterpret_cast<const uchar4*>(P.trainable)[i] : make_uchar4(1, 1, 1, 1);
#define CK_FEDADAM2(c)                                                   \
  if (tr.c) {                                                            \
    const float d = nw.c - w.c;                                          \
    m.c = P.beta1 * m.c + (1.f - P.beta1) * d;                           \
    v.c = P.beta2 * v.c + (1.f - P.beta2) * d * d;                       \
    nw.c = w.c + P.eta * m.c / (sqrtf(v.c) + P.tau);                     \
  }
              CK_FEDADAM2(x) CK_FEDADAM2(y) CK_FEDADAM2(z) CK_FEDADAM2(w)
#undef CK_FEDADAM2
              reinterpret_cast<float4*>(P.m)[i] = m;
              reinterpret_cast<float4*>(P.v)[i] = v;
            }
            reinterpret_cast<float4*>(P.master)[i] = nw;
          }
          asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(obuf + (uint32_t)i4 * 16u), "f"(nw.x), "f"(nw.y),
                       "f"(nw.z), "f"(nw.w) : "memory");
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      asm volatile("bar.sync 1, %0;" ::"n"(TMA_CONSUMERS) : "memory");
      if (t == 0) {
        fa_mbar_arrive(bar0 + 8u * (nstages + s));  // every consumer has read the stage: the producer may refill it
        for (int r = 0; r < P.num_out; ++r)
          fa_bulk_store(reinterpret_cast<float*>(P.out[r]) + e0, obuf, (uint32_t)n4 * 16u);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
    if (t == 0) {
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // the stores themselves, not only their reads
      asm volatile("fence.proxy.async;" ::: "memory");
    }
  }

  if (P.world > 1) {
    __threadfence_system();
    __syncthreads();
    __shared__ int s_last;
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
  {
    // bulk-copy variant: cross-GPU launches with fp32 payload and plain peer stores (CRACKFL_AGG_TMA=0: always fedagg_k)
    static int use_tma = -1, tma_grid = -1, tma_stages = -1;
    // measured (aligned ranks, fp32, fraction of bytes-per-direction / 900 GB/s at 64 MB / 256 MB / 1 GB):
    //   8 GPUs: bulk copies 0.60 / 0.71 / 0.75, fedagg_k 0.49 / 0.60 / 0.68;  2 GPUs: 0.55 / 0.68 / 0.72 against 0.56 / 0.72 / 0.76
    // so the bulk-copy kernel is the default from 4 ranks up (CRACKFL_AGG_TMA=1 forces it, =0 disables it)
    if (use_tma < 0) { const char* e = getenv("CRACKFL_AGG_TMA"); use_tma = e ? atoi(e) : -2; }
    if (tma_grid < 0) { const char* e = getenv("CRACKFL_AGG_TMA_GRID"); tma_grid = e ? atoi(e) : 148; if (tma_grid < 1) tma_grid = 148; }
    if (tma_stages < 0) { const char* e = getenv("CRACKFL_AGG_TMA_STAGES"); tma_stages = e ? atoi(e) : 0; }
    const long nel = p->end - p->begin;
    const bool want_tma = use_tma == -2 ? p->world >= 4 : use_tma > 0;
    if (want_tma && p->world > 1 && p->world <= TMA_MAX_WORLD && p->num_clients == p->world && p->num_out == p->world &&
        !p->payload_bf16 && !p->out_bf16 && !p->delta && !p->mc_in && !p->mc_out && nel > 0) {
      const long nchunks = (nel + TMA_CH - 1) / TMA_CH;
      const int stage_bytes = p->world * TMA_CH * 4;
      int nst = tma_stages > 0 ? tma_stages : 160 * 1024 / stage_bytes;
      if (nst > 12) nst = 12;
      if (nst < 2) nst = 2;
      const size_t smem = (size_t)nst * stage_bytes + 2 * TMA_CH * 4 + 16 * nst + 256;
      static bool attr = false;
      if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(fedagg_tma_k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return (int)e;
        attr = true;
      }
      int g = (int)(nchunks < tma_grid ? nchunks : tma_grid);
      fedagg_tma_k<<<g, 32 + TMA_CONSUMERS, smem, st>>>(*p, nst);
      CK_LAUNCH_CHECK();
      return 0;
    }
  }
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
