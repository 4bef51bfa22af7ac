// NVLink / NVSwitch collectives written as ordinary CUDA kernels over symmetric memory
// (every rank maps every peer's buffer; optionally an NVLS multicast mapping of the same
// buffer).  No NCCL on these paths.
//
//   b200_allreduce  - DDP bucket all-reduce, in place in the symmetric gradient arena, fused
//                     with the 1/world scale: two-shot, rank r reduces slice r with
//                     multimem.ld_reduce (in-switch reduction) and broadcasts it with
//                     multimem.st; falls back to peer loads/stores when no multicast mapping
//                     exists.  (reference: the DDP reducer behind main-ddp.py:55,124)
//   b200_rs_adamw   - FSDP reduce-scatter fused with the sharded AdamW update: the owner of a
//                     shard reduces it straight out of the peers' gradient buffers, scales,
//                     updates its fp32 master shard and moments and writes the bf16 shard the
//                     peers will gather.  (reference: FSDP post-backward reduce-scatter +
//                     optimizer step, main-fsdp.py:138-140)
//   b200_allgather  - pulls every peer's shard of a unit into the local full buffer (used as
//                     the prefetch path; the fused path lives in the GEMM, ag_gemm.cu).
//   b200_barrier    - cross-GPU barrier on the flag words.
//
// Cross-GPU synchronisation: per (channel, peer) 32-bit flag words in a symmetric flag
// buffer, self-resetting CAS protocol (signal = CAS 0->1 on the peer's word with release.sys,
// wait = CAS 1->0 on the own word with acquire.sys), so the kernels can be replayed from a
// CUDA graph with no host-side epoch.  Every spin is bounded: a dead peer produces an error
// code (b200_device_error), not a wedged GPU.
#include "../include/b200gpt.h"
#include "common.cuh"

namespace {

constexpr int MAX_WORLD = 16;
enum : int { ERR_COMM_SIGNAL = 301, ERR_COMM_WAIT = 302 };
#ifndef B200_COMM_TIMEOUT_CYCLES
#define B200_COMM_TIMEOUT_CYCLES 40000000000LL  // ~20 s: ranks may be skewed by whole steps
#endif

// bound of every cross-GPU spin, in SM cycles; b200_set_comm_timeout_ms changes it for the current device (tests
// shorten it to prove that a missing peer is reported instead of hanging)
__device__ long long g_comm_timeout_cycles = B200_COMM_TIMEOUT_CYCLES;

struct CommCtx {
  void* const* bufs;        // device array [world] of peer base pointers (data buffer)
  void* mc;                 // multicast base pointer of the same buffer, or nullptr
  uint32_t* const* flags;   // device array [world] of peer flag-buffer pointers
  int rank, world;
  int channel_base;
  int* err_word;
};

__device__ __forceinline__ uint32_t cas_release_sys(uint32_t* addr, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.release.sys.global.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val) : "memory");
  return old;
}
__device__ __forceinline__ uint32_t cas_acquire_sys(uint32_t* addr, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.acquire.sys.global.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val) : "memory");
  return old;
}

// Block-level barrier across all ranks on channel (channel_base + blockIdx.x).
// Returns false (for every thread of the block) if a peer did not show up in time.
__device__ bool block_barrier(const CommCtx& c, int* s_fail) {
  __syncthreads();
  if (threadIdx.x == 0) *s_fail = 0;
  __syncthreads();
  if (threadIdx.x < c.world) {
    const int peer = threadIdx.x;
    const int ch = c.channel_base + blockIdx.x;
    uint32_t* remote = c.flags[peer] + ch * MAX_WORLD + c.rank;
    uint32_t* local = c.flags[c.rank] + ch * MAX_WORLD + peer;
    const long long t0 = clock64();
    const long long limit = g_comm_timeout_cycles;
    bool ok = true;
    while (cas_release_sys(remote, 0u, 1u) != 0u) {
      if (clock64() - t0 > limit) { ok = false; atomicCAS(c.err_word, 0, ERR_COMM_SIGNAL); break; }
    }
    while (ok && cas_acquire_sys(local, 1u, 0u) != 1u) {
      if (clock64() - t0 > limit) { ok = false; atomicCAS(c.err_word, 0, ERR_COMM_WAIT); break; }
    }
    if (!ok) *s_fail = 1;
  }
  __syncthreads();
  return *s_fail == 0;
}

__device__ __forceinline__ float4 mc_ld_reduce_f32x4(const float* mc_addr) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc_addr) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_f32x4(float* mc_addr, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               ::"l"(mc_addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 ld_sys_f32x4(const float* addr) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void st_sys_f32x4(float* addr, float4 v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_sys_b32x4(const void* addr) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(addr) : "memory");
  return v;
}

// slice of [0, nvec) owned by `part` of `parts` (in float4 units)
__device__ __forceinline__ void part_range(int64_t nvec, int parts, int part, int64_t& lo, int64_t& hi) {
  const int64_t per = (nvec + parts - 1) / parts;
  lo = per * part;
  hi = lo + per < nvec ? lo + per : nvec;
  if (lo > nvec) lo = nvec;
}

// ------------------------------------------------------------------------------------------
template <bool USE_MC>
__global__ void __launch_bounds__(512)
allreduce_kernel(const CommCtx c, int64_t off, int64_t n, float scale) {
  __shared__ int s_fail;
  if (!block_barrier(c, &s_fail)) return;          // every rank's bucket is complete
  const int64_t nvec = n >> 2;
  int64_t lo, hi;
  part_range(nvec, c.world, c.rank, lo, hi);
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  constexpr int U = 8;  // independent 16-byte transactions in flight per thread (NVLink latency ~2 us)
  if (USE_MC) {
    float* mc = reinterpret_cast<float*>(c.mc) + off;
    for (int64_t i0 = lo + blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i0 < hi; i0 += stride * U) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + u * stride;
        if (i < hi) v[u] = mc_ld_reduce_f32x4(mc + 4 * i);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + u * stride;
        if (i < hi) {
          v[u].x *= scale; v[u].y *= scale; v[u].z *= scale; v[u].w *= scale;
          mc_st_f32x4(mc + 4 * i, v[u]);
        }
      }
    }
  } else {
    for (int64_t i0 = lo + blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i0 < hi; i0 += stride * 4) {
      float4 acc[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
      for (int p = 0; p < c.world; ++p) {
        const int peer = (c.rank + p) % c.world;
        const float* src = reinterpret_cast<const float*>(c.bufs[peer]) + off;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int64_t i = i0 + u * stride;
          if (i < hi) {
            const float4 v = ld_sys_f32x4(src + 4 * i);
            acc[u].x += v.x; acc[u].y += v.y; acc[u].z += v.z; acc[u].w += v.w;
          }
        }
      }
#pragma unroll 1
      for (int p = 0; p < c.world; ++p) {
        const int peer = (c.rank + p) % c.world;
        float* dst = reinterpret_cast<float*>(c.bufs[peer]) + off;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int64_t i = i0 + u * stride;
          if (i < hi)
            st_sys_f32x4(dst + 4 * i, make_float4(acc[u].x * scale, acc[u].y * scale, acc[u].z * scale, acc[u].w * scale));
        }
      }
    }
  }
  __threadfence_system();
  block_barrier(c, &s_fail);                       // every rank's slice has landed everywhere
}

// ------------------------------------------------------------------------------------------
// All-reduce with bf16 on the wire (fp32 accumulation in the switch): halves the NVLink bytes of a bucket.
// The fp32 gradients stay where they are; `stage` is a symmetric bf16 buffer of the same geometry.
//   phase 0  cast own gradients fp32 -> bf16 into the local stage
//   phase 1  rank r reduces slice r:  multimem.ld_reduce.acc::f32 (bf16x2) -> * scale -> multimem.st (bf16x2)
//   phase 2  stage (now the reduced values, identical on every rank) -> fp32 gradients
// Work is laid out so that a block only ever touches, in every phase, sub-part `blockIdx.x` of every rank's
// slice: the per-block cross-GPU barriers between the phases are then sufficient (no grid-wide barrier).
__device__ __forceinline__ uint4 mc_ld_reduce_bf16x8(const void* mc_addr) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc_addr) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_bf16x8(void* mc_addr, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1, %2, %3, %4};"
               ::"l"(mc_addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__global__ void __launch_bounds__(512)
allreduce_bf16_kernel(const CommCtx c, float* __restrict__ grad, int64_t off, int64_t n, float scale) {
  __shared__ int s_fail;
  bf16* stage = reinterpret_cast<bf16*>(c.bufs[c.rank]) + off;
  bf16* stage_mc = reinterpret_cast<bf16*>(c.mc) + off;
  float* g = grad + off;
  const int64_t nvec = n >> 3;                       // 8 elements = 16 bytes of bf16
  const int64_t per_rank = (nvec + c.world - 1) / c.world;
  const int64_t per_blk = (per_rank + gridDim.x - 1) / gridDim.x;
  // this block's sub-part of rank r's slice: [lo(r), hi(r))
  auto span = [&](int r, int64_t& lo, int64_t& hi) {
    const int64_t s0 = per_rank * r, s1 = s0 + per_rank < nvec ? s0 + per_rank : nvec;
    lo = s0 + per_blk * blockIdx.x;
    hi = lo + per_blk < s1 ? lo + per_blk : s1;
    if (lo > s1) lo = s1;
  };
  for (int r = 0; r < c.world; ++r) {                 // phase 0
    int64_t lo, hi;
    span(r, lo, hi);
    for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
      const float4 a = reinterpret_cast<const float4*>(g)[2 * i], b = reinterpret_cast<const float4*>(g)[2 * i + 1];
      reinterpret_cast<uint4*>(stage)[i] = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w),
                                                      pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
    }
  }
  __threadfence_system();
  if (!block_barrier(c, &s_fail)) return;
  {                                                   // phase 1
    int64_t lo, hi;
    span(c.rank, lo, hi);
    constexpr int U = 8;
    for (int64_t i0 = lo + threadIdx.x; i0 < hi; i0 += static_cast<int64_t>(blockDim.x) * U) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + static_cast<int64_t>(u) * blockDim.x;
        if (i < hi) v[u] = mc_ld_reduce_bf16x8(stage_mc + 8 * i);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + static_cast<int64_t>(u) * blockDim.x;
        if (i < hi) {
          uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const float2 f = unpack_bf16x2(w[t]);
            w[t] = pack_bf16x2(f.x * scale, f.y * scale);
          }
          mc_st_bf16x8(stage_mc + 8 * i, make_uint4(w[0], w[1], w[2], w[3]));
        }
      }
    }
  }
  __threadfence_system();
  if (!block_barrier(c, &s_fail)) return;
  for (int r = 0; r < c.world; ++r) {                 // phase 2
    int64_t lo, hi;
    span(r, lo, hi);
    for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
      const uint4 u = reinterpret_cast<const uint4*>(stage)[i];
      const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
      reinterpret_cast<float4*>(g)[2 * i] = make_float4(a.x, a.y, b.x, b.y);
      reinterpret_cast<float4*>(g)[2 * i + 1] = make_float4(cc.x, cc.y, d.x, d.y);
    }
  }
}

// ------------------------------------------------------------------------------------------
// reduce-scatter fused with sharded AdamW.  Shard = elements [shard_off, shard_off + shard_n) of the
// symmetric gradient buffer; p/m/v are the local fp32 shard arrays (index 0 = shard_off);
// p_bf16 (nullable) is the local bf16 shard the peers gather from.
struct AdamArgs {
  float* p; float* m; float* v; bf16* p_bf16;
  const float* hyper;   // {lr, step}
  float beta1, beta2, eps, wd;
};

template <bool USE_MC>
__global__ void __launch_bounds__(512)
rs_adamw_kernel(const CommCtx c, int64_t shard_off, int64_t shard_n, float scale, const AdamArgs a,
                int end_barrier) {
  __shared__ int s_fail;
  if (!block_barrier(c, &s_fail)) return;
  const float lr = a.hyper[0], step = a.hyper[1];
  const float bc1 = 1.0f - powf(a.beta1, step);
  const float bc2_rsqrt = rsqrtf(1.0f - powf(a.beta2, step));
  const float step_size = lr / bc1, decay = 1.0f - lr * a.wd;
  const int64_t nvec = shard_n >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  constexpr int U = 4;
  for (int64_t i0 = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i0 < nvec; i0 += stride * U) {
    float4 gq[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = i0 + u * stride;
      gq[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < nvec) {
        if (USE_MC) {
          gq[u] = mc_ld_reduce_f32x4(reinterpret_cast<const float*>(c.mc) + shard_off + 4 * i);
        } else {
#pragma unroll 1
          for (int p = 0; p < c.world; ++p) {
            const int peer = (c.rank + p) % c.world;
            const float4 v = ld_sys_f32x4(reinterpret_cast<const float*>(c.bufs[peer]) + shard_off + 4 * i);
            gq[u].x += v.x; gq[u].y += v.y; gq[u].z += v.z; gq[u].w += v.w;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = i0 + u * stride;
      if (i >= nvec) continue;
      float4 g = gq[u];
      float4 pv = reinterpret_cast<float4*>(a.p)[i];
      float4 mv = reinterpret_cast<float4*>(a.m)[i];
      float4 vv = reinterpret_cast<float4*>(a.v)[i];
      float* pp = &pv.x; float* gp = &g.x; float* mp = &mv.x; float* vp = &vv.x;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float gj = gp[j] * scale;
        if (isfinite(gj)) {
          pp[j] *= decay;
          mp[j] = a.beta1 * mp[j] + (1.0f - a.beta1) * gj;
          vp[j] = a.beta2 * vp[j] + (1.0f - a.beta2) * gj * gj;
          pp[j] -= step_size * mp[j] / (sqrtf(vp[j]) * bc2_rsqrt + a.eps);
        } else {
          atomicAdd(const_cast<float*>(a.hyper) + 2, 1.0f);   // skipped element, see adamw_kernel
        }
      }
      reinterpret_cast<float4*>(a.p)[i] = pv;
      reinterpret_cast<float4*>(a.m)[i] = mv;
      reinterpret_cast<float4*>(a.v)[i] = vv;
      if (a.p_bf16 != nullptr)
        reinterpret_cast<uint2*>(a.p_bf16)[i] = make_uint2(pack_bf16x2(pv.x, pv.y), pack_bf16x2(pv.z, pv.w));
    }
  }
  if (end_barrier) {
    __threadfence_system();
    block_barrier(c, &s_fail);
  }
}

// ------------------------------------------------------------------------------------------
// reduce-scatter alone: out[i] = scale * sum over ranks of grad[shard_off + i].  Meant for a SMALL grid
// (<= 16 blocks): with NVLS the switch does the additions and one multimem.ld_reduce brings 16 reduced
// bytes, so a handful of blocks with 8 loads in flight per thread saturate what a training step needs,
// and the compute kernels keep their SMs.  The optimizer then runs on `out` as an ordinary kernel.
template <bool USE_MC>
__global__ void __launch_bounds__(512)
reduce_scatter_kernel(const CommCtx c, int64_t shard_off, int64_t shard_n, float scale, float* __restrict__ out,
                      int end_barrier) {
  __shared__ int s_fail;
  if (!block_barrier(c, &s_fail)) return;           // every rank's gradients of this range are complete
  const int64_t nvec = shard_n >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  constexpr int U = 8;
  for (int64_t i0 = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i0 < nvec; i0 += stride * U) {
    float4 g[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = i0 + u * stride;
      g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < nvec) {
        if (USE_MC) {
          g[u] = mc_ld_reduce_f32x4(reinterpret_cast<const float*>(c.mc) + shard_off + 4 * i);
        } else {
#pragma unroll 1
          for (int p = 0; p < c.world; ++p) {
            const int peer = (c.rank + p) % c.world;
            const float4 v = ld_sys_f32x4(reinterpret_cast<const float*>(c.bufs[peer]) + shard_off + 4 * i);
            g[u].x += v.x; g[u].y += v.y; g[u].z += v.z; g[u].w += v.w;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = i0 + u * stride;
      if (i < nvec)
        reinterpret_cast<float4*>(out)[i] = make_float4(g[u].x * scale, g[u].y * scale, g[u].z * scale, g[u].w * scale);
    }
  }
  if (end_barrier) {   // nobody overwrites its gradients (next step) while a peer still reads them
    __threadfence_system();
    block_barrier(c, &s_fail);
  }
}

// ------------------------------------------------------------------------------------------
// all-gather (pull): dst[p * shard_bytes + i] = peer p's src shard, 16-byte vectors.
__global__ void __launch_bounds__(512)
allgather_kernel(const CommCtx c, int64_t src_off_bytes, int64_t shard_bytes, uint8_t* dst,
                 int start_barrier) {
  __shared__ int s_fail;
  if (start_barrier && !block_barrier(c, &s_fail)) return;
  const int64_t nvec = shard_bytes >> 4;
  const int64_t total = nvec * c.world;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i0 = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i0 < total; i0 += stride * 4) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t i = i0 + u * stride;
      if (i < total) {
        const int peer = (c.rank + static_cast<int>(i / nvec)) % c.world;   // local shard first, then ring order
        v[u] = ld_sys_b32x4(reinterpret_cast<const uint8_t*>(c.bufs[peer]) + src_off_bytes + 16 * (i % nvec));
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t i = i0 + u * stride;
      if (i < total) {
        const int peer = (c.rank + static_cast<int>(i / nvec)) % c.world;
        reinterpret_cast<uint4*>(dst + static_cast<int64_t>(peer) * shard_bytes)[i % nvec] = v[u];
      }
    }
  }
}

// busy-wait of a given length on the stream (race-stress tests skew the ranks with it)
__global__ void spin_kernel(long long ns) {
  long long t0, t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  do {
    __nanosleep(200);
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  } while (t - t0 < ns);
}

__global__ void barrier_kernel(const CommCtx c) {
  __shared__ int s_fail;
  __threadfence_system();
  block_barrier(c, &s_fail);
}

int make_ctx(CommCtx* c, const void* bufs_dev, void* mc, const void* flags_dev, int rank, int world,
             int channel_base, int nblocks) {
  B200_REQUIRE(world >= 1 && world <= MAX_WORLD, "comm: world %d not supported (max %d)", world, MAX_WORLD);
  B200_REQUIRE(rank >= 0 && rank < world, "comm: bad rank %d/%d", rank, world);
  B200_REQUIRE(nblocks >= 1 && channel_base >= 0 && channel_base + nblocks <= 256, "comm: bad channel range");
  c->bufs = reinterpret_cast<void* const*>(bufs_dev);
  c->mc = mc;
  c->flags = reinterpret_cast<uint32_t* const*>(flags_dev);
  c->rank = rank; c->world = world; c->channel_base = channel_base;
  c->err_word = b200_error_word();
  B200_REQUIRE(c->err_word != nullptr, "comm: no CUDA device");
  return 0;
}

}  // namespace

extern "C" {

int b200_allreduce(const void* bufs_dev, void* mc_ptr, const void* flags_dev, int rank, int world,
                   int64_t offset_elems, int64_t n_elems, float scale, int nblocks, int channel_base,
                   void* stream) {
  B200_REQUIRE(n_elems % 4 == 0 && offset_elems % 4 == 0, "allreduce: offset/count must be multiples of 4");
  if (n_elems == 0) return 0;
  CommCtx c;
  int rc = make_ctx(&c, bufs_dev, mc_ptr, flags_dev, rank, world, channel_base, nblocks);
  if (rc != 0) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (mc_ptr != nullptr) allreduce_kernel<true><<<nblocks, 512, 0, st>>>(c, offset_elems, n_elems, scale);
  else allreduce_kernel<false><<<nblocks, 512, 0, st>>>(c, offset_elems, n_elems, scale);
  B200_LAUNCH_CHECK();
  return 0;
}

// Bucket all-reduce with bf16 on the wire (the "bf16/fp32 cast" variant of the DDP reducer, main-ddp.py:55,124):
// grad (local fp32 arena) elements [offset, offset + n) are summed over the ranks, scaled, and written back; the
// values travel as bf16 through `stage` (symmetric bf16 buffer with a multicast mapping, same element offsets),
// accumulated in fp32 inside the switch.  The result carries one bf16 rounding of every rank's addend and one of
// the sum.  Needs NVLS (stage_mc_ptr != NULL); callers fall back to b200_allreduce otherwise.
int b200_allreduce_bf16(float* grad, const void* stage_bufs_dev, void* stage_mc_ptr, const void* flags_dev, int rank,
                        int world, int64_t offset_elems, int64_t n_elems, float scale, int nblocks, int channel_base,
                        void* stream) {
  B200_REQUIRE(n_elems % 8 == 0 && offset_elems % 8 == 0, "allreduce_bf16: offset/count must be multiples of 8");
  B200_REQUIRE(stage_mc_ptr != nullptr, "allreduce_bf16: needs a multicast mapping of the staging buffer");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(grad) & 15) == 0, "allreduce_bf16: gradient arena must be 16-byte aligned");
  if (n_elems == 0) return 0;
  CommCtx c;
  int rc = make_ctx(&c, stage_bufs_dev, stage_mc_ptr, flags_dev, rank, world, channel_base, nblocks);
  if (rc != 0) return rc;
  allreduce_bf16_kernel<<<nblocks, 512, 0, static_cast<cudaStream_t>(stream)>>>(c, grad, offset_elems, n_elems, scale);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_rs_adamw(const void* grad_bufs_dev, void* grad_mc_ptr, const void* flags_dev, int rank, int world,
                  int64_t shard_off, int64_t shard_n, float scale, float* p, float* m, float* v,
                  void* p_bf16, const float* hyper, float beta1, float beta2, float eps, float wd,
                  int end_barrier, int nblocks, int channel_base, void* stream) {
  B200_REQUIRE(shard_n % 4 == 0 && shard_off % 4 == 0, "rs_adamw: offset/count must be multiples of 4");
  if (shard_n == 0) return 0;
  CommCtx c;
  int rc = make_ctx(&c, grad_bufs_dev, grad_mc_ptr, flags_dev, rank, world, channel_base, nblocks);
  if (rc != 0) return rc;
  AdamArgs a{p, m, v, reinterpret_cast<bf16*>(p_bf16), hyper, beta1, beta2, eps, wd};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (grad_mc_ptr != nullptr)
    rs_adamw_kernel<true><<<nblocks, 512, 0, st>>>(c, shard_off, shard_n, scale, a, end_barrier);
  else
    rs_adamw_kernel<false><<<nblocks, 512, 0, st>>>(c, shard_off, shard_n, scale, a, end_barrier);
  B200_LAUNCH_CHECK();
  return 0;
}

// Reduce-scatter of this rank's shard into a local fp32 buffer (FSDP post-backward reduce-scatter,
// main-fsdp.py:138-140), for callers that run the optimizer as a separate kernel: small grid, no AdamW
// traffic inside the kernel that spins on the peers.
int b200_reduce_scatter(const void* grad_bufs_dev, void* grad_mc_ptr, const void* flags_dev, int rank, int world,
                        int64_t shard_off, int64_t shard_n, float scale, float* out, int end_barrier, int nblocks,
                        int channel_base, void* stream) {
  B200_REQUIRE(shard_n % 4 == 0 && shard_off % 4 == 0, "reduce_scatter: offset/count must be multiples of 4");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0, "reduce_scatter: output must be 16-byte aligned");
  if (shard_n == 0) return 0;
  CommCtx c;
  int rc = make_ctx(&c, grad_bufs_dev, grad_mc_ptr, flags_dev, rank, world, channel_base, nblocks);
  if (rc != 0) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (grad_mc_ptr != nullptr)
    reduce_scatter_kernel<true><<<nblocks, 512, 0, st>>>(c, shard_off, shard_n, scale, out, end_barrier);
  else
    reduce_scatter_kernel<false><<<nblocks, 512, 0, st>>>(c, shard_off, shard_n, scale, out, end_barrier);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_allgather(const void* src_bufs_dev, const void* flags_dev, int rank, int world,
                   int64_t src_off_bytes, int64_t shard_bytes, void* dst, int start_barrier, int nblocks,
                   int channel_base, void* stream) {
  B200_REQUIRE(shard_bytes % 16 == 0 && src_off_bytes % 16 == 0, "allgather: sizes must be multiples of 16 bytes");
  if (shard_bytes == 0) return 0;
  CommCtx c;
  int rc = make_ctx(&c, src_bufs_dev, nullptr, flags_dev, rank, world, channel_base, nblocks);
  if (rc != 0) return rc;
  allgather_kernel<<<nblocks, 512, 0, static_cast<cudaStream_t>(stream)>>>(
      c, src_off_bytes, shard_bytes, reinterpret_cast<uint8_t*>(dst), start_barrier);
  B200_LAUNCH_CHECK();
  return 0;
}

// Test aids.  b200_spin: keep the stream busy for `ns` nanoseconds (skews ranks against each other).
// b200_set_comm_timeout_ms: bound of every cross-GPU wait on the current device (default ~20 s); b200_reset_flags
// clears a flag buffer after a deliberately provoked timeout left its words half-signalled.
int b200_spin(int64_t ns, void* stream) {
  spin_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<long long>(ns));
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_set_comm_timeout_ms(int ms) {
  B200_REQUIRE(ms > 0, "comm timeout must be positive");
  int dev = 0, khz = 0;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_CHECK_CUDA(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
  const long long cycles = static_cast<long long>(ms) * (khz > 0 ? khz : 1965000);
  B200_CHECK_CUDA(cudaMemcpyToSymbol(g_comm_timeout_cycles, &cycles, sizeof(cycles)));
  return 0;
}

int b200_barrier(const void* flags_dev, int rank, int world, int channel, void* stream) {
  CommCtx c;
  int rc = make_ctx(&c, nullptr, nullptr, flags_dev, rank, world, channel, 1);
  if (rc != 0) return rc;
  barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(c);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
