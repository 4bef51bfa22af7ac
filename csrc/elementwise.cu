// Memory-bound kernels of the GPT training step: fused token+position embedding (fwd/bwd),
// LayerNorm fp32 -> bf16 (fwd/bwd with dgamma/dbeta), fused multi-tensor AdamW over the flat
// arena (writes the bf16 compute copy in the same pass), ReLU-backward + bias-gradient, and the
// in-place cross-entropy forward+backward over a bf16 logits chunk (loss, argmax accuracy and
// dlogits in one pass), plus the small kernels around the softmax-epilogue LM-head GEMM.  All are
// vectorised (16-byte accesses) and sized against the measured HBM roofline.  The column reductions
// (bias / gamma / beta gradients) use a caller-provided scratch of 16 replica rows and a ticket
// counter; b200_relu_bwd_static at the end of the file carries that scratch inside the library.
//
// Reference call sites: models/gpt.py:185 (embeddings), :119,122,217 (LayerNorm), :35,38
// (ReLU), main-single.py:42,99-101 (AdamW + GradScaler), :95-96,127-131 (CE + accuracy).
#include <math.h>

#include <mutex>

#include "../include/b200gpt.h"
#include "common.cuh"

namespace {

// ------------------------------------------------------------------------------------------
// embedding
// ------------------------------------------------------------------------------------------
__global__ void embedding_fwd_kernel(const int64_t* __restrict__ ids, const int64_t* __restrict__ pos,
                                     const float* __restrict__ tok_w, const float* __restrict__ pos_w,
                                     float* __restrict__ out, int n_tokens, int dim, int vocab,
                                     int max_pos) {
  const int vec_per_row = dim >> 2;
  const int64_t total = static_cast<int64_t>(n_tokens) * vec_per_row;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int t = static_cast<int>(i / vec_per_row);
    const int c = static_cast<int>(i % vec_per_row);
    int64_t id = ids[t], ps = pos[t];
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    ps = ps < 0 ? 0 : (ps >= max_pos ? max_pos - 1 : ps);
    const float4 a = __ldg(reinterpret_cast<const float4*>(tok_w + id * dim) + c);
    const float4 b = __ldg(reinterpret_cast<const float4*>(pos_w + ps * dim) + c);
    reinterpret_cast<float4*>(out + static_cast<int64_t>(t) * dim)[c] =
        make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
  }
}

__device__ __forceinline__ void red_add_f32x4(float* addr, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

__global__ void embedding_bwd_kernel(const int64_t* __restrict__ ids, const int64_t* __restrict__ pos,
                                     const float* __restrict__ dout, float* __restrict__ dtok,
                                     float* __restrict__ dpos, int n_tokens, int dim, int vocab,
                                     int max_pos) {
  const int vec_per_row = dim >> 2;
  const int64_t total = static_cast<int64_t>(n_tokens) * vec_per_row;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int t = static_cast<int>(i / vec_per_row);
    const int c = static_cast<int>(i % vec_per_row);
    int64_t id = ids[t], ps = pos[t];
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    ps = ps < 0 ? 0 : (ps >= max_pos ? max_pos - 1 : ps);
    const float4 g = __ldg(reinterpret_cast<const float4*>(dout + static_cast<int64_t>(t) * dim) + c);
    red_add_f32x4(dtok + id * dim + 4 * c, g);
    red_add_f32x4(dpos + ps * dim + 4 * c, g);
  }
}

// ------------------------------------------------------------------------------------------
// Column reductions (bias / gamma / beta gradients) without same-address atomic pile-ups: every
// block adds its partial column sums into one of REPLICAS replica rows of a scratch buffer (so a
// cache line sees 1/REPLICAS of the atomics), then takes a ticket; the block that draws the last
// ticket folds the replicas into the real gradient and leaves scratch and ticket counter zeroed
// for the next launch (so the kernels replay inside a CUDA graph).
// scratch: [REPLICAS][ncols] floats, zero on entry;  counter: one int, zero on entry.
// Launches that share one scratch buffer must not overlap on the GPU.
// ------------------------------------------------------------------------------------------
constexpr int REPLICAS = 16;

__device__ __forceinline__ void finish_column_sums(float* __restrict__ scratch, int* __restrict__ counter,
                                                   int ncols, int total_blocks, float* __restrict__ out_a,
                                                   float* __restrict__ out_b, int split,
                                                   float* __restrict__ out_c = nullptr, bool three = false) {
  __shared__ int s_last;
  __threadfence();
  __syncthreads();
  const int tid = threadIdx.y * blockDim.x + threadIdx.x;
  const int nthreads = blockDim.x * blockDim.y;
  if (tid == 0) s_last = (atomicAdd(counter, 1) == total_blocks - 1) ? 1 : 0;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int c = tid; c < ncols; c += nthreads) {
    float s = 0.f;
#pragma unroll
    for (int r = 0; r < REPLICAS; ++r) {
      s += __ldcg(scratch + r * ncols + c);
      scratch[r * ncols + c] = 0.f;
    }
    if (c < split) out_a[c] += s;
    else if (!three || c < 2 * split) out_b[c - split] += s;
    else if (out_c != nullptr) out_c[c - 2 * split] += s;   // three equally wide outputs, the third optional
  }
  if (tid == 0) *counter = 0;
}

// ------------------------------------------------------------------------------------------
// LayerNorm: one warp per row, NV float4 per lane held in registers (dim <= NV*128)
// ------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(256)
layernorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                     const float* __restrict__ beta, bf16* __restrict__ y, float* __restrict__ mean_out,
                     float* __restrict__ rstd_out, int rows, int dim, float eps) {
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int nvec = dim >> 2;
  const float inv_dim = 1.0f / static_cast<float>(dim);
  for (int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < rows;
       row += gridDim.x * warps_per_block) {
    const float4* xr = reinterpret_cast<const float4*>(x + static_cast<int64_t>(row) * dim);
    float4 v[NV];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      v[i] = c < nvec ? xr[c] : make_float4(0.f, 0.f, 0.f, 0.f);
      s += v[i].x + v[i].y + v[i].z + v[i].w;
    }
    const float mean = warp_sum(s) * inv_dim;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) {
        const float a = v[i].x - mean, b = v[i].y - mean, cc = v[i].z - mean, d = v[i].w - mean;
        sq += a * a + b * b + cc * cc + d * d;
      }
    }
    const float rstd = rsqrtf(warp_sum(sq) * inv_dim + eps);
    uint2* yr = reinterpret_cast<uint2*>(y + static_cast<int64_t>(row) * dim);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) {
        const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + c);
        const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + c);
        const float o0 = (v[i].x - mean) * rstd * g.x + b.x;
        const float o1 = (v[i].y - mean) * rstd * g.y + b.y;
        const float o2 = (v[i].z - mean) * rstd * g.z + b.z;
        const float o3 = (v[i].w - mean) * rstd * g.w + b.w;
        yr[c] = make_uint2(pack_bf16x2(o0, o1), pack_bf16x2(o2, o3));
      }
    }
    if (lane == 0) {
      mean_out[row] = mean;
      rstd_out[row] = rstd;
    }
  }
}

// Narrow rows (dim <= 256): one warp handles R rows per iteration with every load of the R rows
// issued before the first use (the kernel is latency bound: a row is only 1.5 KB), values stay in
// registers for the single pass, and the dres (residual-branch gradient) add is folded in.
// FUSE: the kernel also does the first step of the backward of the layer that PRODUCED x (a Linear or MLP whose
// epilogue added the residual, reference models/gpt.py:33-41,105-123): dz = dx * act'(mask_src) as bf16 (the operand
// of that layer's dgrad / wgrad GEMMs) and dbias_z += column sums of dz, so no separate pass re-reads dx.
template <int NV, int R, bool FUSE = false>
__global__ void __launch_bounds__(256, R == 1 ? 3 : 2)   // <= 128 (R = 1: 80) registers: 16 (24) warps per SM
layernorm_bwd_small_kernel(const bf16* __restrict__ dy, const float* __restrict__ x,
                           const float* __restrict__ gamma, const float* __restrict__ mean_in,
                           const float* __restrict__ rstd_in, const float* __restrict__ dres,
                           float* __restrict__ dx, bf16* __restrict__ dx_bf16, float* __restrict__ dgamma,
                           float* __restrict__ dbeta, int rows, int dim, float* __restrict__ scratch,
                           int* __restrict__ counter, const bf16* __restrict__ mask_src = nullptr, int act = 0,
                           float* __restrict__ dbias_z = nullptr) {
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int warps_per_block = blockDim.x >> 5;
  const int nvec = dim >> 2;
  const float inv_dim = 1.0f / static_cast<float>(dim);
  float4 g4[NV], dg[NV], db[NV];
  [[maybe_unused]] float4 dbz[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + 32 * i;
    g4[i] = c < nvec ? __ldg(reinterpret_cast<const float4*>(gamma) + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    dg[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    db[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if constexpr (FUSE) dbz[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const int row_stride = gridDim.x * warps_per_block * R;
  for (int row0 = (blockIdx.x * warps_per_block + warp) * R; row0 < rows; row0 += row_stride) {
    float4 xv[R][NV], rv[R][NV];
    uint2 dv[R][NV];
    [[maybe_unused]] uint2 mv[R][NV];   // FUSE: the producer's activation mask, fetched with everything else
    float mean[R], rstd[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int row = row0 + r;
      const bool ok = row < rows;
      mean[r] = ok ? mean_in[row] : 0.f;
      rstd[r] = ok ? rstd_in[row] : 0.f;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int c = lane + 32 * i;
        const bool okc = ok && c < nvec;
        const size_t off = static_cast<size_t>(ok ? row : 0) * dim;
        xv[r][i] = okc ? reinterpret_cast<const float4*>(x + off)[c] : make_float4(0.f, 0.f, 0.f, 0.f);
        dv[r][i] = okc ? reinterpret_cast<const uint2*>(dy + off)[c] : make_uint2(0u, 0u);
        rv[r][i] = (okc && dres != nullptr) ? reinterpret_cast<const float4*>(dres + off)[c]
                                            : make_float4(0.f, 0.f, 0.f, 0.f);
        if constexpr (FUSE)
          mv[r][i] = (okc && mask_src != nullptr) ? reinterpret_cast<const uint2*>(mask_src + off)[c]
                                                  : make_uint2(0x3f803f80u, 0x3f803f80u);   // 1.0: keeps everything
      }
    }
    float c1[R], c2[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      c1[r] = 0.f; c2[r] = 0.f;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const float2 d01 = unpack_bf16x2(dv[r][i].x), d23 = unpack_bf16x2(dv[r][i].y);
        const float h0 = (xv[r][i].x - mean[r]) * rstd[r], h1 = (xv[r][i].y - mean[r]) * rstd[r],
                    h2 = (xv[r][i].z - mean[r]) * rstd[r], h3 = (xv[r][i].w - mean[r]) * rstd[r];
        const float a0 = d01.x * g4[i].x, a1 = d01.y * g4[i].y, a2 = d23.x * g4[i].z, a3 = d23.y * g4[i].w;
        c1[r] += a0 + a1 + a2 + a3;
        c2[r] += a0 * h0 + a1 * h1 + a2 * h2 + a3 * h3;
        dg[i].x += d01.x * h0; dg[i].y += d01.y * h1; dg[i].z += d23.x * h2; dg[i].w += d23.y * h3;
        db[i].x += d01.x; db[i].y += d01.y; db[i].z += d23.x; db[i].w += d23.y;
        xv[r][i] = make_float4(h0, h1, h2, h3);   // keep xhat
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        c1[r] += __shfl_xor_sync(0xffffffffu, c1[r], o);
        c2[r] += __shfl_xor_sync(0xffffffffu, c2[r], o);
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int row = row0 + r;
      if (row >= rows) continue;
      const float m1 = c1[r] * inv_dim, m2 = c2[r] * inv_dim;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int c = lane + 32 * i;
        if (c < nvec) {
          const float2 d01 = unpack_bf16x2(dv[r][i].x), d23 = unpack_bf16x2(dv[r][i].y);
          float4 o;
          o.x = (d01.x * g4[i].x - m1 - xv[r][i].x * m2) * rstd[r] + rv[r][i].x;
          o.y = (d01.y * g4[i].y - m1 - xv[r][i].y * m2) * rstd[r] + rv[r][i].y;
          o.z = (d23.x * g4[i].z - m1 - xv[r][i].z * m2) * rstd[r] + rv[r][i].z;
          o.w = (d23.y * g4[i].w - m1 - xv[r][i].w * m2) * rstd[r] + rv[r][i].w;
          reinterpret_cast<float4*>(dx + static_cast<size_t>(row) * dim)[c] = o;
          if constexpr (FUSE) {
            if (mask_src != nullptr) {
              const float2 m01 = unpack_bf16x2(mv[r][i].x), m23 = unpack_bf16x2(mv[r][i].y);
              if (act >= 2) {   // GELU: mask_src is the pre-activation
                o.x *= gelu_grad(m01.x, act); o.y *= gelu_grad(m01.y, act);
                o.z *= gelu_grad(m23.x, act); o.w *= gelu_grad(m23.y, act);
              } else {          // ReLU: mask_src is the activated output
                if (!(m01.x > 0.f)) o.x = 0.f;
                if (!(m01.y > 0.f)) o.y = 0.f;
                if (!(m23.x > 0.f)) o.z = 0.f;
                if (!(m23.y > 0.f)) o.w = 0.f;
              }
            }
            dbz[i].x += o.x; dbz[i].y += o.y; dbz[i].z += o.z; dbz[i].w += o.w;
          }
          if (dx_bf16 != nullptr)
            reinterpret_cast<uint2*>(dx_bf16 + static_cast<size_t>(row) * dim)[c] =
                make_uint2(pack_bf16x2(o.x, o.y), pack_bf16x2(o.z, o.w));
        }
      }
    }
  }
  constexpr int NOUT = FUSE ? 3 : 2;
  extern __shared__ float red[];  // [warps][NOUT][dim]
  float* my = red + static_cast<size_t>(warp) * NOUT * dim;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) {
      reinterpret_cast<float4*>(my)[c] = dg[i];
      reinterpret_cast<float4*>(my + dim)[c] = db[i];
      if constexpr (FUSE) reinterpret_cast<float4*>(my + 2 * dim)[c] = dbz[i];
    }
  }
  __syncthreads();
  float* rep = scratch + static_cast<size_t>(blockIdx.x % REPLICAS) * NOUT * dim;
  for (int j = threadIdx.x; j < NOUT * dim; j += blockDim.x) {
    float sacc = 0.f;
    for (int w = 0; w < warps_per_block; ++w) sacc += red[static_cast<size_t>(w) * NOUT * dim + j];
    atomicAdd(rep + j, sacc);
  }
  finish_column_sums(scratch, counter, NOUT * dim, gridDim.x, dgamma, dbeta, dim, dbias_z, FUSE);
}

template <int NV, bool DX_ONLY = false>
__global__ void __launch_bounds__(256)
layernorm_bwd_kernel(const bf16* __restrict__ dy, const float* __restrict__ x,
                     const float* __restrict__ gamma, const float* __restrict__ mean_in,
                     const float* __restrict__ rstd_in, const float* __restrict__ dres,
                     float* __restrict__ dx, bf16* __restrict__ dx_bf16, float* __restrict__ dgamma,
                     float* __restrict__ dbeta, int rows, int dim, float* __restrict__ scratch,
                     int* __restrict__ counter, const bf16* __restrict__ mask_src = nullptr, int act = 0) {
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int warps_per_block = blockDim.x >> 5;
  const int nvec = dim >> 2;
  const float inv_dim = 1.0f / static_cast<float>(dim);
  const float4* g4p = reinterpret_cast<const float4*>(gamma);
  float4 dg[DX_ONLY ? 1 : NV], db[DX_ONLY ? 1 : NV];
#pragma unroll
  for (int i = 0; i < (DX_ONLY ? 1 : NV); ++i) {
    dg[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    db[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (int row = blockIdx.x * warps_per_block + warp; row < rows; row += gridDim.x * warps_per_block) {
    const float4* xr = reinterpret_cast<const float4*>(x + static_cast<int64_t>(row) * dim);
    const uint2* dyr = reinterpret_cast<const uint2*>(dy + static_cast<int64_t>(row) * dim);
    const float mean = mean_in[row], rstd = rstd_in[row];
    float c1 = 0.f, c2 = 0.f;
    // pass 1: row statistics of dy*gamma and the dgamma/dbeta partials
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) {
        const float4 xv = xr[c];
        const uint2 u = dyr[c];
        const float4 g = __ldg(g4p + c);
        const float2 d01 = unpack_bf16x2(u.x), d23 = unpack_bf16x2(u.y);
        const float h0 = (xv.x - mean) * rstd, h1 = (xv.y - mean) * rstd, h2 = (xv.z - mean) * rstd,
                    h3 = (xv.w - mean) * rstd;
        const float a0 = d01.x * g.x, a1 = d01.y * g.y, a2 = d23.x * g.z, a3 = d23.y * g.w;
        c1 += a0 + a1 + a2 + a3;
        c2 += a0 * h0 + a1 * h1 + a2 * h2 + a3 * h3;
        if constexpr (!DX_ONLY) {
          dg[i].x += d01.x * h0; dg[i].y += d01.y * h1; dg[i].z += d23.x * h2; dg[i].w += d23.y * h3;
          db[i].x += d01.x; db[i].y += d01.y; db[i].z += d23.x; db[i].w += d23.y;
        }
      }
    }
    c1 = warp_sum(c1) * inv_dim;
    c2 = warp_sum(c2) * inv_dim;
    // pass 2: dx (the row is re-read from L1; keeps the register footprint independent of dim)
    float4* dxr = reinterpret_cast<float4*>(dx + static_cast<int64_t>(row) * dim);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) {
        const float4 xv = xr[c];
        const uint2 u = dyr[c];
        const float4 g = __ldg(g4p + c);
        const float2 d01 = unpack_bf16x2(u.x), d23 = unpack_bf16x2(u.y);
        float4 o;
        o.x = (d01.x * g.x - c1 - (xv.x - mean) * rstd * c2) * rstd;
        o.y = (d01.y * g.y - c1 - (xv.y - mean) * rstd * c2) * rstd;
        o.z = (d23.x * g.z - c1 - (xv.z - mean) * rstd * c2) * rstd;
        o.w = (d23.y * g.w - c1 - (xv.w - mean) * rstd * c2) * rstd;
        if (dres != nullptr) {
          const float4 r = reinterpret_cast<const float4*>(dres + static_cast<int64_t>(row) * dim)[c];
          o.x += r.x; o.y += r.y; o.z += r.z; o.w += r.w;
        }
        dxr[c] = o;
        if (mask_src != nullptr) {   // dx_bf16 becomes dz = dx * act'(mask_src) of the layer that produced x
          const uint2 m = reinterpret_cast<const uint2*>(mask_src + static_cast<int64_t>(row) * dim)[c];
          const float2 m01 = unpack_bf16x2(m.x), m23 = unpack_bf16x2(m.y);
          if (act >= 2) {
            o.x *= gelu_grad(m01.x, act); o.y *= gelu_grad(m01.y, act);
            o.z *= gelu_grad(m23.x, act); o.w *= gelu_grad(m23.y, act);
          } else {
            if (!(m01.x > 0.f)) o.x = 0.f;
            if (!(m01.y > 0.f)) o.y = 0.f;
            if (!(m23.x > 0.f)) o.z = 0.f;
            if (!(m23.y > 0.f)) o.w = 0.f;
          }
        }
        if (dx_bf16 != nullptr)
          reinterpret_cast<uint2*>(dx_bf16 + static_cast<int64_t>(row) * dim)[c] =
              make_uint2(pack_bf16x2(o.x, o.y), pack_bf16x2(o.z, o.w));
      }
    }
  }
  if constexpr (DX_ONLY) return;
  // block reduction of the per-warp dgamma/dbeta partials, then one atomic per column per block
  extern __shared__ float red[];  // [warps][2][dim]
  float* my = red + static_cast<size_t>(warp) * 2 * dim;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) {
      reinterpret_cast<float4*>(my)[c] = dg[i];
      reinterpret_cast<float4*>(my + dim)[c] = db[i];
    }
  }
  __syncthreads();
  float* rep = scratch + static_cast<size_t>(blockIdx.x % REPLICAS) * 2 * dim;
  for (int j = threadIdx.x; j < 2 * dim; j += blockDim.x) {
    float s = 0.f;
    for (int w = 0; w < warps_per_block; ++w) s += red[static_cast<size_t>(w) * 2 * dim + j];
    atomicAdd(rep + j, s);
  }
  finish_column_sums(scratch, counter, 2 * dim, gridDim.x, dgamma, dbeta, dim);
}

// dgamma / dbeta for wide rows: block (32 column groups of 4, 8 row lanes) walks a slab of rows,
// recomputing xhat from the saved statistics; partial sums go through the replica scratch.
constexpr int LN_DG_ROWS = 128;
__global__ void __launch_bounds__(256)
layernorm_bwd_dgdb_kernel(const bf16* __restrict__ dy, const float* __restrict__ x,
                          const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
                          float* __restrict__ dgamma, float* __restrict__ dbeta, int rows, int dim,
                          float* __restrict__ scratch, int* __restrict__ counter,
                          const bf16* __restrict__ dz = nullptr, float* __restrict__ dbias_z = nullptr) {
  // dz / dbias_z (optional): third column sum, over the bf16 dz the dx kernel wrote (fused backward hand-off)
  const int cg = blockIdx.x * 32 + threadIdx.x;
  const int col = cg * 4;
  const bool col_ok = col < dim;
  const bool three = dz != nullptr;
  const int nout = three ? 3 : 2;
  float4 g = make_float4(0.f, 0.f, 0.f, 0.f), b = make_float4(0.f, 0.f, 0.f, 0.f), z = make_float4(0.f, 0.f, 0.f, 0.f);
  const int r0 = blockIdx.y * LN_DG_ROWS;
  if (col_ok) {
#pragma unroll 4
    for (int r = r0 + threadIdx.y; r < min(r0 + LN_DG_ROWS, rows); r += 8) {
      const float4 xv = *reinterpret_cast<const float4*>(x + static_cast<size_t>(r) * dim + col);
      const uint2 u = *reinterpret_cast<const uint2*>(dy + static_cast<size_t>(r) * dim + col);
      const float mean = mean_in[r], rstd = rstd_in[r];
      const float2 d01 = unpack_bf16x2(u.x), d23 = unpack_bf16x2(u.y);
      g.x += d01.x * (xv.x - mean) * rstd; g.y += d01.y * (xv.y - mean) * rstd;
      g.z += d23.x * (xv.z - mean) * rstd; g.w += d23.y * (xv.w - mean) * rstd;
      b.x += d01.x; b.y += d01.y; b.z += d23.x; b.w += d23.y;
      if (three) {
        const uint2 w = *reinterpret_cast<const uint2*>(dz + static_cast<size_t>(r) * dim + col);
        const float2 z01 = unpack_bf16x2(w.x), z23 = unpack_bf16x2(w.y);
        z.x += z01.x; z.y += z01.y; z.z += z23.x; z.w += z23.y;
      }
    }
  }
  __shared__ float red[8][32][13];
  red[threadIdx.y][threadIdx.x][0] = g.x; red[threadIdx.y][threadIdx.x][1] = g.y;
  red[threadIdx.y][threadIdx.x][2] = g.z; red[threadIdx.y][threadIdx.x][3] = g.w;
  red[threadIdx.y][threadIdx.x][4] = b.x; red[threadIdx.y][threadIdx.x][5] = b.y;
  red[threadIdx.y][threadIdx.x][6] = b.z; red[threadIdx.y][threadIdx.x][7] = b.w;
  red[threadIdx.y][threadIdx.x][8] = z.x; red[threadIdx.y][threadIdx.x][9] = z.y;
  red[threadIdx.y][threadIdx.x][10] = z.z; red[threadIdx.y][threadIdx.x][11] = z.w;
  __syncthreads();
  if (threadIdx.y == 0 && col_ok) {
    float* rep = scratch + static_cast<size_t>(blockIdx.y % REPLICAS) * nout * dim;
#pragma unroll
    for (int j = 0; j < 12; ++j) {
      if (j >= 4 * nout) break;
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) s += red[w][threadIdx.x][j];
      atomicAdd(rep + (j >> 2) * dim + col + (j & 3), s);
    }
  }
  finish_column_sums(scratch, counter, nout * dim, gridDim.x * gridDim.y, dgamma, dbeta, dim, dbias_z, three);
}

// ------------------------------------------------------------------------------------------
// AdamW (decoupled weight decay), torch.optim.AdamW maths; hyper = {lr, step, ...} on device
// ------------------------------------------------------------------------------------------
__global__ void adamw_tick_kernel(float* hyper) { hyper[1] += 1.0f; }

__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
             float* __restrict__ v, bf16* __restrict__ pb, int64_t n, const float* __restrict__ hyper,
             float beta1, float beta2, float eps, float wd) {
  const float lr = hyper[0];
  const float step = hyper[1];
  const float bc1 = 1.0f - powf(beta1, step);
  const float bc2_rsqrt = rsqrtf(1.0f - powf(beta2, step));
  const float step_size = lr / bc1;
  const float decay = 1.0f - lr * wd;
  const int64_t nvec = n >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < nvec; i += stride) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = __ldg(reinterpret_cast<const float4*>(g) + i);
    float4 mv = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float* pp = &pv.x; const float* gp = &gv.x; float* mp = &mv.x; float* vp = &vv.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gj = gp[j];
      if (isfinite(gj)) {
        pp[j] *= decay;
        mp[j] = beta1 * mp[j] + (1.0f - beta1) * gj;
        vp[j] = beta2 * vp[j] + (1.0f - beta2) * gj * gj;
        const float denom = sqrtf(vp[j]) * bc2_rsqrt + eps;
        pp[j] -= step_size * mp[j] / denom;
      } else {
        // A non-finite gradient element leaves its parameter and moments untouched and is counted in hyper[2].
        // Deliberate deviation from the reference's GradScaler (main-single.py:99-101), which drops the WHOLE
        // step on any inf/nan after a host round trip: here the finite elements still update, nothing syncs.
        atomicAdd(const_cast<float*>(hyper) + 2, 1.0f);
      }
    }
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(m)[i] = mv;
    reinterpret_cast<float4*>(v)[i] = vv;
    if (pb != nullptr)
      reinterpret_cast<uint2*>(pb)[i] = make_uint2(pack_bf16x2(pv.x, pv.y), pack_bf16x2(pv.z, pv.w));
  }
  // tail (n not a multiple of 4)
  for (int64_t i = (nvec << 2) + blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += stride) {
    const float gj = g[i];
    if (!isfinite(gj)) atomicAdd(const_cast<float*>(hyper) + 2, 1.0f);
    if (isfinite(gj)) {
      float pj = p[i] * decay;
      const float mj = beta1 * m[i] + (1.0f - beta1) * gj;
      const float vj = beta2 * v[i] + (1.0f - beta2) * gj * gj;
      pj -= step_size * mj / (sqrtf(vj) * bc2_rsqrt + eps);
      p[i] = pj; m[i] = mj; v[i] = vj;
    }
    if (pb != nullptr) pb[i] = __float2bfloat16_rn(p[i]);
  }
}

__global__ void __launch_bounds__(256)
cast_bf16_kernel(const float* __restrict__ in, bf16* __restrict__ out, int64_t n) {
  const int64_t nvec = n >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < nvec; i += stride) {
    const float4 f = __ldg(reinterpret_cast<const float4*>(in) + i);
    reinterpret_cast<uint2*>(out)[i] = make_uint2(pack_bf16x2(f.x, f.y), pack_bf16x2(f.z, f.w));
  }
  for (int64_t i = (nvec << 2) + blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += stride)
    out[i] = __float2bfloat16_rn(in[i]);
}

// ------------------------------------------------------------------------------------------
// dz = dy * (y_act > 0) -> bf16 ; dbias += column sums of dz.  block (32, 8): 32 column groups of
// 8 columns x 8 row lanes, 64 rows per block.
// ------------------------------------------------------------------------------------------
constexpr int RELU_ROWS = 64;  // rows per block (8 row lanes x 8)

template <bool DY_FP32>
__global__ void __launch_bounds__(256)
relu_bwd_kernel(const void* __restrict__ dy_, const bf16* __restrict__ y_act, bf16* __restrict__ dz,
                float* __restrict__ dbias, int rows, int cols, int ld, float* __restrict__ scratch,
                int* __restrict__ counter, int act = 1) {
  // act 1: y_act is the ReLU output (mask = y > 0); act 2 / 3: y_act is the GELU pre-activation (factor gelu'(z))
  const int cg = blockIdx.x * 32 + threadIdx.x;  // column group (8 columns)
  const int col = cg * 8;
  const bool col_ok = col < cols;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const int r0 = blockIdx.y * RELU_ROWS;
  if (col_ok) {
#pragma unroll 2
    for (int r = r0 + threadIdx.y; r < min(r0 + RELU_ROWS, rows); r += 8) {
      const size_t off = static_cast<size_t>(r) * ld + col;
      float v[8];
      if (DY_FP32) {
        const float* dy = reinterpret_cast<const float*>(dy_) + off;
        const float4 a = *reinterpret_cast<const float4*>(dy);
        const float4 b = *reinterpret_cast<const float4*>(dy + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
      } else {
        const uint4 u = *reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(dy_) + off);
        const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z),
                     d = unpack_bf16x2(u.w);
        v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
      }
      if (y_act != nullptr) {
        const uint4 u = *reinterpret_cast<const uint4*>(y_act + off);
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const float2 f = unpack_bf16x2(w[t]);
          if (act >= 2) {
            v[2 * t] *= gelu_grad(f.x, act);
            v[2 * t + 1] *= gelu_grad(f.y, act);
          } else {
            if (!(f.x > 0.f)) v[2 * t] = 0.f;
            if (!(f.y > 0.f)) v[2 * t + 1] = 0.f;
          }
        }
      }
      if (dz != nullptr)
        *reinterpret_cast<uint4*>(dz + off) =
            make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                       pack_bf16x2(v[6], v[7]));
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
  }
  if (dbias == nullptr) return;
  __shared__ float red[8][32][9];
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.y][threadIdx.x][j] = acc[j];
  __syncthreads();
  if (threadIdx.y == 0 && col_ok) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) s += red[w][threadIdx.x][j];
      atomicAdd(scratch + (blockIdx.y % REPLICAS) * cols + col + j, s);
    }
  }
  finish_column_sums(scratch, counter, cols, gridDim.x * gridDim.y, dbias, dbias, cols);
}

// ------------------------------------------------------------------------------------------
// cross entropy
// ------------------------------------------------------------------------------------------
__global__ void count_valid_kernel(const int64_t* __restrict__ targets, float* __restrict__ stats,
                                   int ignore_index, int64_t n) {
  __shared__ int sm[32];
  int c = 0;
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) c += (targets[i] != ignore_index) ? 1 : 0;
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < (blockDim.x >> 5); ++w) t += sm[w];
    stats[0] = static_cast<float>(t);
    stats[1] = 0.f;
    stats[2] = 0.f;
    stats[3] = 0.f;
  }
}

constexpr int CE_THREADS = 512;

__device__ __forceinline__ void cp_async_16(uint32_t smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}

// One block per row.  The logits row (bf16, <= 113 KB) is staged in shared memory with cp.async (no
// register staging, so two rows per SM overlap each other's phases), reduced (packed-bf16 max,
// sum-exp) and - when write_grad - overwritten in place in global memory with
// dlogits = (softmax - onehot) * scale.  exp() is evaluated once per element: p = exp(l - max) is
// written back to the staged row as bf16 and reused by the gradient pass.
__global__ void __launch_bounds__(CE_THREADS, 2)
ce_fwd_bwd_kernel(bf16* __restrict__ logits, const int64_t* __restrict__ targets,
                  float* __restrict__ stats, int vocab, int ld, int ignore_index, int write_grad,
                  float grad_scale) {
  extern __shared__ uint4 srow[];
  const int row = blockIdx.x;
  bf16* lr = logits + static_cast<size_t>(row) * ld;
  const int64_t tgt64 = targets[row];
  const bool valid = tgt64 != ignore_index;
  const int tgt = static_cast<int>(tgt64);
  const int nvec = ld >> 3;
  __shared__ float s_f[CE_THREADS / 32];
  __shared__ float s_bcast[2];

  if (!valid) {  // ignored row: no loss, zero gradient
    if (write_grad)
      for (int c = threadIdx.x; c < nvec; c += CE_THREADS)
        reinterpret_cast<uint4*>(lr)[c] = make_uint4(0u, 0u, 0u, 0u);
    return;
  }
  const uint32_t sbase = smem_u32(srow);
  for (int c = threadIdx.x; c < nvec; c += CE_THREADS) cp_async_16(sbase + c * 16, reinterpret_cast<const uint4*>(lr) + c);
  asm volatile("cp.async.commit_group;" ::: "memory");
  const float tgt_logit = __bfloat162float(lr[tgt]);
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();

  // pad columns [vocab, ld) -> -inf in the staged row (each thread patches the vectors it copied
  // itself), so the passes below run unpredicated and the pad gradient comes out as exact zeros
  for (int c = vocab >> 3; c < nvec; ++c) {
    if (c % CE_THREADS != static_cast<int>(threadIdx.x)) continue;
    unsigned short* e = reinterpret_cast<unsigned short*>(srow + c);
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (c * 8 + j >= vocab) e[j] = 0xFF80u;   // bf16 -inf
  }

  // pass 1: row max with packed bf16x2 max (one instruction per two logits)
  __nv_bfloat162 m2 = __floats2bfloat162_rn(-INFINITY, -INFINITY);
  for (int c = threadIdx.x; c < nvec; c += CE_THREADS) {
    const uint4 u = srow[c];
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int t = 0; t < 4; ++t) m2 = __hmax2(m2, *reinterpret_cast<const __nv_bfloat162*>(&w[t]));
  }
  float mx = fmaxf(__low2float(m2), __high2float(m2));
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0) s_f[warp] = mx;
  __syncthreads();
  if (warp == 0) {
    float t2 = lane < CE_THREADS / 32 ? s_f[lane] : -INFINITY;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t2 = fmaxf(t2, __shfl_xor_sync(0xffffffffu, t2, o));
    if (lane == 0) s_bcast[0] = t2;
  }
  __syncthreads();
  const float row_max = s_bcast[0];
  // prediction correct <=> the target holds the max and no lower index ties it (torch.argmax
  // returns the first maximum).  Only rows whose target logit equals the max pay for the scan.
  bool correct = false;
  if (tgt_logit == row_max) {           // block-uniform
    int earlier = 0;
    const int cend = tgt >> 3;
    for (int c = threadIdx.x; c <= cend; c += CE_THREADS) {
      const bf16* e = reinterpret_cast<const bf16*>(srow + c);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        earlier |= (c * 8 + j < tgt && __bfloat162float(e[j]) == row_max) ? 1 : 0;
    }
    correct = __syncthreads_or(earlier) == 0;
  }
  const float kLog2e = 1.4426950408889634f;
  const float mscaled = row_max * kLog2e;

  // pass 2: p = exp(l - max), summed in fp32 and written back to the staged row as bf16
  float sum0 = 0.f, sum1 = 0.f;
  for (int c = threadIdx.x; c < nvec; c += CE_THREADS) {
    const uint4 u = srow[c];
    uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const float2 f = unpack_bf16x2(w[t]);
      const float p0 = exp2f(fmaf(f.x, kLog2e, -mscaled));
      const float p1 = exp2f(fmaf(f.y, kLog2e, -mscaled));
      sum0 += p0;
      sum1 += p1;
      w[t] = pack_bf16x2(p0, p1);
    }
    srow[c] = make_uint4(w[0], w[1], w[2], w[3]);   // same thread re-reads it below: no sync needed
  }
  float sum = warp_sum(sum0 + sum1);
  if (lane == 0) s_f[warp] = sum;
  __syncthreads();
  if (warp == 0) {
    float s2 = lane < CE_THREADS / 32 ? s_f[lane] : 0.f;
    s2 = warp_sum(s2);
    if (lane == 0) s_bcast[1] = s2;
  }
  __syncthreads();
  const float total = s_bcast[1];
  if (threadIdx.x == 0) {
    const float lse = row_max + logf(total);
    atomicAdd(stats + 1, lse - tgt_logit);
    if (correct) atomicAdd(stats + 2, 1.0f);
  }
  if (!write_grad) return;
  const float n_valid = fmaxf(stats[0], 1.0f);
  const float scale = grad_scale / n_valid;
  const float pscale = scale / total;
  // pass 3: dlogits = p * scale / sum, streamed straight to global; the target column is fixed up
  // afterwards by the thread that stored its vector (program order on one address)
  for (int c = threadIdx.x; c < nvec; c += CE_THREADS) {
    const uint4 u = srow[c];
    uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const float2 f = unpack_bf16x2(w[t]);
      w[t] = pack_bf16x2(f.x * pscale, f.y * pscale);
    }
    reinterpret_cast<uint4*>(lr)[c] = make_uint4(w[0], w[1], w[2], w[3]);
  }
  if ((tgt >> 3) % CE_THREADS == static_cast<int>(threadIdx.x)) {
    const float pt = __bfloat162float(reinterpret_cast<const bf16*>(srow)[tgt]);
    lr[tgt] = __float2bfloat16(pt * pscale - scale);
  }
}

// ------------------------------------------------------------------------------------------
// LM head around the softmax-epilogue GEMM (b200_lmhead_gemm_bf16, csrc/gemm_tcgen05.cu):
//   begin   count the valid targets, zero the statistics, rotate the gap estimate
//   norms   l2 norm of every activation row; largest l2 norm of a weight row
//   [GEMM]  E = exp(logit - shift), (sum, max) partials per row and column span, target logits
//   finish  per row: total sum / max -> loss, accuracy, row scale of the gradient, E[target] -= sum (the
//           "- onehot" of softmax - onehot), scaled activations for the weight-gradient GEMM; rows whose
//           maximum fell outside the shift's window are recomputed exactly, here, by their warp.
// lmstate (4 floats that live with the weight): [0] wmax  [1] gap used by this call  [2] smallest gap seen
// by this call (the next call's estimate)  [3] rows recomputed by this call.
// ------------------------------------------------------------------------------------------
__global__ void lm_begin_kernel(const int64_t* __restrict__ targets, float* __restrict__ stats,
                                float* __restrict__ lmstate, int ignore_index, int64_t n) {
  __shared__ int sm[32];
  int c = 0;
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) c += (targets[i] != ignore_index) ? 1 : 0;
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < (blockDim.x >> 5); ++w) t += sm[w];
    stats[0] = static_cast<float>(t);
    stats[1] = 0.f;
    stats[2] = 0.f;
    stats[3] = 0.f;
    const float seen = lmstate[2];
    if (isfinite(seen) && seen >= 0.f) lmstate[1] = seen;   // no valid row last time: keep the estimate
    lmstate[0] = 0.f;
    lmstate[2] = INFINITY;
    lmstate[3] = 0.f;
  }
}

__device__ __forceinline__ float dot8_bf16(const uint4 a, const uint4 b, float acc) {
  const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const float2 fa = unpack_bf16x2(aw[t]), fb = unpack_bf16x2(bw[t]);
    acc = fmaf(fa.x, fb.x, acc);
    acc = fmaf(fa.y, fb.y, acc);
  }
  return acc;
}

// one warp per row: rows [0, M) are activations (norm -> xnorm), rows [M, M + V) weights (max norm -> lmstate[0])
__global__ void __launch_bounds__(256)
lm_norms_kernel(const bf16* __restrict__ X, int ldx, const bf16* __restrict__ W, int ldw, int M, int V, int D,
                float* __restrict__ xnorm, float* __restrict__ lmstate) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  float wmax = 0.f;
  for (int r = warp; r < M + V; r += nwarps) {
    const bf16* src = r < M ? X + static_cast<size_t>(r) * ldx : W + static_cast<size_t>(r - M) * ldw;
    float acc = 0.f;
    for (int d = lane * 8; d < D; d += 256) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(src + d));
      acc = dot8_bf16(v, v, acc);
    }
    acc = sqrtf(warp_sum(acc));
    if (r < M) {
      if (lane == 0) xnorm[r] = acc;
    } else {
      wmax = fmaxf(wmax, acc);
    }
  }
  if (lane == 0 && wmax > 0.f) atomicMax(reinterpret_cast<int*>(lmstate), __float_as_int(wmax));   // positive floats order like ints
}

constexpr int LMF_ROWS = 64;   // rows per block of lm_finish_kernel

__global__ void __launch_bounds__(256)
lm_finish_kernel(const float2* __restrict__ part, int num_parts, int M, const float* __restrict__ xnorm,
                 const float* __restrict__ tgt_val, const int64_t* __restrict__ targets, int ignore_index,
                 float* __restrict__ stats, float* __restrict__ lmstate, bf16* __restrict__ E, int lde, int V,
                 const bf16* __restrict__ X, int ldx, int D, const bf16* __restrict__ W, int ldw,
                 bf16* __restrict__ Xs, float* __restrict__ rscale, float grad_scale, int write_grad) {
  __shared__ float sh_sum[4][LMF_ROWS], sh_max[4][LMF_ROWS];
  __shared__ float sh_red[3][8];
  const int row0 = blockIdx.x * LMF_ROWS;
  {   // phase 1: total (sum, max) of each row; consecutive threads read consecutive rows of one span
    const int r = threadIdx.x & (LMF_ROWS - 1), kg = threadIdx.x / LMF_ROWS;
    float s = 0.f, m = -INFINITY;
    if (row0 + r < M) {
      for (int k = kg; k < num_parts; k += 4) {
        const float2 pr = __ldg(part + static_cast<size_t>(k) * M + row0 + r);
        s += pr.x;
        m = fmaxf(m, pr.y);
      }
    }
    sh_sum[kg][r] = s;
    sh_max[kg][r] = m;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float wmax = lmstate[0], gap_used = lmstate[1];
  const float n_valid = fmaxf(stats[0], 1.0f);
  constexpr float kLog2e = 1.4426950408889634f;
  float loss_acc = 0.f, corr_acc = 0.f, gap_acc = INFINITY, bad_acc = 0.f;
  for (int rr = warp; rr < LMF_ROWS; rr += 8) {
    const int row = row0 + rr;
    if (row >= M) break;
    float s = (sh_sum[0][rr] + sh_sum[1][rr]) + (sh_sum[2][rr] + sh_sum[3][rr]);
    float m = fmaxf(fmaxf(sh_max[0][rr], sh_max[1][rr]), fmaxf(sh_max[2][rr], sh_max[3][rr]));
    const float xn = xnorm[row];
    const float ubound = __fmul_rn(__fmul_rn(xn, wmax), 1.0002f);
    float shift = lm_row_shift(xn, wmax, gap_used);
    const long long t64 = targets[row];
    const bool valid = t64 != ignore_index && t64 >= 0 && t64 < V;
    const int tgt = valid ? static_cast<int>(t64) : -1;
    float tl = valid ? tgt_val[row] : 0.f;
    const float a = m - shift;
    const bool good = isfinite(s) && s > 0.f && a >= kLmWindowLo && a <= kLmWindowHi;
    // NaN in the row (NaN activations or weights) cannot be repaired by recomputing it: let it through, the
    // loss comes out NaN like the reference's
    const bool hopeless = s != s || m != m || m == INFINITY;
    if (!good && !hopeless) {
      // the shift estimate missed this row (or the row holds non-finite values): redo it exactly with
      // shift = row max.  Every lane takes its own vocabulary entries; x is read through L1, W through L2.
      const bf16* xr = X + static_cast<size_t>(row) * ldx;
      float mx = -INFINITY;
      for (int j = lane; j < V; j += 32) {
        const bf16* wr = W + static_cast<size_t>(j) * ldw;
        float acc = 0.f;
        for (int d = 0; d < D; d += 8)
          acc = dot8_bf16(__ldg(reinterpret_cast<const uint4*>(xr + d)), __ldg(reinterpret_cast<const uint4*>(wr + d)), acc);
        mx = fmaxf(mx, acc);
      }
      mx = warp_max(mx);
      float ssum = 0.f, tfound = -INFINITY;
      for (int j = lane; j < V; j += 32) {
        const bf16* wr = W + static_cast<size_t>(j) * ldw;
        float acc = 0.f;
        for (int d = 0; d < D; d += 8)
          acc = dot8_bf16(__ldg(reinterpret_cast<const uint4*>(xr + d)), __ldg(reinterpret_cast<const uint4*>(wr + d)), acc);
        const float e = fast_exp2((acc - mx) * kLog2e);
        ssum += e;
        if (j == tgt) tfound = acc;
        if (write_grad) E[static_cast<size_t>(row) * lde + j] = __float2bfloat16(e);
      }
      s = warp_sum(ssum);
      tfound = warp_max(tfound);
      m = mx;
      shift = mx;
      if (valid) tl = tfound;
      bad_acc += 1.f;
      __syncwarp();
    }
    if (valid) {
      loss_acc += (shift + logf(s)) - tl;
      corr_acc += (tl >= m) ? 1.f : 0.f;
      if (isfinite(m)) gap_acc = fminf(gap_acc, fmaxf(ubound - m, 0.f));
    }
    if (write_grad) {
      const float r = valid ? grad_scale / (n_valid * s) : 0.f;
      if (lane == 0) {
        rscale[row] = r;
        // softmax - onehot: (E[t] - sum) / sum = p[t] - 1
        if (valid) E[static_cast<size_t>(row) * lde + tgt] = __float2bfloat16(fast_exp2((tl - shift) * kLog2e) - s);
      }
      const bf16* xr = X + static_cast<size_t>(row) * ldx;
      bf16* xs = Xs + static_cast<size_t>(row) * D;
      for (int d = lane * 8; d < D; d += 256) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(xr + d));
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        uint32_t o[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const float2 f = unpack_bf16x2(w[t]);
          o[t] = pack_bf16x2(f.x * r, f.y * r);
        }
        *reinterpret_cast<uint4*>(xs + d) = make_uint4(o[0], o[1], o[2], o[3]);
      }
    }
  }
  if (lane == 0) {
    sh_red[0][warp] = loss_acc;
    sh_red[1][warp] = corr_acc;
    sh_red[2][warp] = gap_acc;
  }
  if (lane == 0 && bad_acc > 0.f) atomicAdd(lmstate + 3, bad_acc);
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, c = 0.f, g = INFINITY;
    for (int w = 0; w < 8; ++w) {
      l += sh_red[0][w];
      c += sh_red[1][w];
      g = fminf(g, sh_red[2][w]);
    }
    atomicAdd(stats + 1, l);
    if (c > 0.f) atomicAdd(stats + 2, c);
    if (isfinite(g)) atomicMin(reinterpret_cast<int*>(lmstate + 2), __float_as_int(g));
  }
}

inline int grid_for(int64_t work_items, int threads, int max_blocks_per_sm = 8) {
  int64_t blocks = ceil_div64(work_items, threads);
  const int64_t cap = static_cast<int64_t>(b200_num_sms()) * max_blocks_per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

}  // namespace

extern "C" {

int b200_embedding_fwd(const int64_t* ids, const int64_t* pos, const float* tok_w, const float* pos_w,
                       float* out, int n_tokens, int dim, int vocab, int max_pos, void* stream) {
  B200_REQUIRE(dim % 4 == 0, "embedding: dim %d must be a multiple of 4", dim);
  if (n_tokens == 0) return 0;
  const int grid = grid_for(static_cast<int64_t>(n_tokens) * (dim / 4), 256);
  embedding_fwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(ids, pos, tok_w, pos_w, out,
                                                                           n_tokens, dim, vocab, max_pos);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_embedding_bwd(const int64_t* ids, const int64_t* pos, const float* dout, float* dtok,
                       float* dpos, int n_tokens, int dim, int vocab, int max_pos, void* stream) {
  B200_REQUIRE(dim % 4 == 0, "embedding: dim %d must be a multiple of 4", dim);
  if (n_tokens == 0) return 0;
  const int grid = grid_for(static_cast<int64_t>(n_tokens) * (dim / 4), 256);
  embedding_bwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(ids, pos, dout, dtok, dpos,
                                                                           n_tokens, dim, vocab, max_pos);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_layernorm_fwd(const float* x, const float* gamma, const float* beta, void* y, float* mean,
                       float* rstd, int rows, int dim, float eps, void* stream) {
  B200_REQUIRE(dim % 4 == 0 && dim <= 4096, "layernorm: dim %d must be a multiple of 4 and <= 4096", dim);
  if (rows == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = grid_for(static_cast<int64_t>(rows) * 32, 256);
  bf16* yb = reinterpret_cast<bf16*>(y);
  if (dim <= 256) layernorm_fwd_kernel<2><<<grid, 256, 0, st>>>(x, gamma, beta, yb, mean, rstd, rows, dim, eps);
  else if (dim <= 1024) layernorm_fwd_kernel<8><<<grid, 256, 0, st>>>(x, gamma, beta, yb, mean, rstd, rows, dim, eps);
  else if (dim <= 2048) layernorm_fwd_kernel<16><<<grid, 256, 0, st>>>(x, gamma, beta, yb, mean, rstd, rows, dim, eps);
  else layernorm_fwd_kernel<32><<<grid, 256, 0, st>>>(x, gamma, beta, yb, mean, rstd, rows, dim, eps);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_layernorm_bwd(const void* dy, const float* x, const float* gamma, const float* mean,
                       const float* rstd, const float* dres, float* dx, void* dx_bf16, float* dgamma,
                       float* dbeta, int rows, int dim, float* scratch, int* counter, void* stream) {
  return b200_layernorm_bwd_fused(dy, x, gamma, mean, rstd, dres, dx, dx_bf16, dgamma, dbeta, rows, dim, scratch,
                                  counter, nullptr, 0, nullptr, 0, stream);
}

// fuse != 0: dz_out (the dx_bf16 slot) = bf16(dx * act'(mask_src)) and dbias_z += column sums of dz - the first step of
// the backward of the Linear / MLP whose residual epilogue produced x (mask_src NULL: no activation; dbias_z NULL:
// no bias).  scratch: 16 x 3*dim floats.
int b200_layernorm_bwd_fused(const void* dy, const float* x, const float* gamma, const float* mean,
                             const float* rstd, const float* dres, float* dx, void* dx_bf16, float* dgamma,
                             float* dbeta, int rows, int dim, float* scratch, int* counter, const void* mask_src,
                             int act, float* dbias_z, int fuse, void* stream) {
  B200_REQUIRE(dim % 4 == 0 && dim <= 4096, "layernorm: dim %d must be a multiple of 4 and <= 4096", dim);
  B200_REQUIRE(scratch != nullptr && counter != nullptr, "layernorm_bwd: scratch (16 x 3*dim floats) and counter required");
  B200_REQUIRE(!fuse || dx_bf16 != nullptr, "layernorm_bwd: the fused hand-off needs the bf16 output buffer");
  if (rows == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int grid = b200_num_sms() * 4;
  const int max_grid = ceil_div(rows, 8);
  if (grid > max_grid) grid = max_grid;
  const size_t smem = static_cast<size_t>(8) * (fuse ? 3 : 2) * dim * sizeof(float);
  const bf16* dyb = reinterpret_cast<const bf16*>(dy);
  bf16* dxb = reinterpret_cast<bf16*>(dx_bf16);
  const bf16* msk = reinterpret_cast<const bf16*>(mask_src);
#define LN_BWD(NV)                                                                                  \
  do {                                                                                              \
    if (smem > 48 * 1024)                                                                           \
      B200_CHECK_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel<NV>,                                \
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    layernorm_bwd_kernel<NV><<<grid, 256, smem, st>>>(dyb, x, gamma, mean, rstd, dres, dx, dxb,     \
                                                      dgamma, dbeta, rows, dim, scratch, counter);  \
  } while (0)
  if (dim <= 256) {
    static const int rows_per_iter0 = []() { const char* e = getenv("B200GPT_LN_ROWS"); return e ? atoi(e) : 1; }();
    int g = b200_num_sms() * (rows_per_iter0 == 1 && dim > 128 ? 3 : 2);
    const int mg = ceil_div(rows, 8 * 4);
    if (g > mg) g = mg;
    if (g < 1) g = 1;
    // rows a warp keeps in flight per iteration for 129..256 columns (B200GPT_LN_ROWS = 1 / 2 / 3 / 4)
    // (measured in the default training step: 1 row - 80 registers, 24 warps per SM - 4.19 ms, 2 rows 4.22 ms,
    // 3 rows 4.27 ms, 4 rows - which spills - 4.36 ms)
    static const int rows_per_iter = []() { const char* e = getenv("B200GPT_LN_ROWS"); return e ? atoi(e) : 1; }();
#define LN_SMALL(NV, R, F)                                                                                          \
  layernorm_bwd_small_kernel<NV, R, F><<<g, 256, smem, st>>>(dyb, x, gamma, mean, rstd, dres, dx, dxb, dgamma, dbeta, \
                                                             rows, dim, scratch, counter, msk, act, dbias_z)
    if (fuse) {
      if (dim <= 128) LN_SMALL(1, 4, true);
      else if (rows_per_iter == 1) LN_SMALL(2, 1, true);
      else if (rows_per_iter == 2) LN_SMALL(2, 2, true);
      else if (rows_per_iter == 3) LN_SMALL(2, 3, true);
      else LN_SMALL(2, 4, true);
    } else {
      if (dim <= 128) LN_SMALL(1, 4, false);
      else if (rows_per_iter == 1) LN_SMALL(2, 1, false);
      else if (rows_per_iter == 2) LN_SMALL(2, 2, false);
      else if (rows_per_iter == 3) LN_SMALL(2, 3, false);
      else LN_SMALL(2, 4, false);
    }
#undef LN_SMALL
  } else if (dim <= 1024 && !fuse) LN_BWD(8);
  else {
    // wide rows (and the fused hand-off above 256 columns): dx by the warp-per-row kernel without the column
    // accumulators (they would need 128+ registers per lane), dgamma / dbeta (/ dbias_z) by a column-parallel pass
    const int gdx = (rows + 7) / 8 < b200_num_sms() * 8 ? (rows + 7) / 8 : b200_num_sms() * 8;
    if (dim <= 1024)
      layernorm_bwd_kernel<8, true><<<gdx, 256, 0, st>>>(dyb, x, gamma, mean, rstd, dres, dx, dxb, dgamma, dbeta,
                                                         rows, dim, scratch, counter, msk, act);
    else if (dim <= 2048)
      layernorm_bwd_kernel<16, true><<<gdx, 256, 0, st>>>(dyb, x, gamma, mean, rstd, dres, dx, dxb, dgamma, dbeta,
                                                          rows, dim, scratch, counter, msk, act);
    else
      layernorm_bwd_kernel<32, true><<<gdx, 256, 0, st>>>(dyb, x, gamma, mean, rstd, dres, dx, dxb, dgamma, dbeta,
                                                          rows, dim, scratch, counter, msk, act);
    B200_LAUNCH_CHECK();
    dim3 g2(ceil_div(dim / 4, 32), ceil_div(rows, LN_DG_ROWS)), b2(32, 8);
    layernorm_bwd_dgdb_kernel<<<g2, b2, 0, st>>>(dyb, x, mean, rstd, dgamma, dbeta, rows, dim, scratch, counter,
                                                 (fuse && dbias_z != nullptr) ? dxb : nullptr, dbias_z);
  }
#undef LN_BWD
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_adamw_tick(float* hyper, void* stream) {
  adamw_tick_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(hyper);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_adamw(float* p, const float* g, float* m, float* v, void* p_bf16, int64_t n,
               const float* hyper, float beta1, float beta2, float eps, float wd, void* stream) {
  if (n == 0) return 0;
  B200_REQUIRE((reinterpret_cast<uintptr_t>(p) & 15) == 0 && (reinterpret_cast<uintptr_t>(g) & 15) == 0 &&
                   (reinterpret_cast<uintptr_t>(m) & 15) == 0 && (reinterpret_cast<uintptr_t>(v) & 15) == 0 &&
                   (reinterpret_cast<uintptr_t>(p_bf16) & 7) == 0,
               "adamw: buffers must be 16-byte aligned");
  const int grid = grid_for(n / 4 + 1, 256, 16);
  adamw_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      p, g, m, v, reinterpret_cast<bf16*>(p_bf16), n, hyper, beta1, beta2, eps, wd);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_cast_bf16(const float* in, void* out, int64_t n, void* stream) {
  if (n == 0) return 0;
  B200_REQUIRE((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 7) == 0,
               "cast: buffers must be 16/8-byte aligned");
  const int grid = grid_for(n / 4 + 1, 256, 16);
  cast_bf16_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(in, reinterpret_cast<bf16*>(out), n);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_relu_bwd(const void* dy, const void* y_act, void* dz, float* dbias, int rows, int cols,
                  int dy_fp32, float* scratch, int* counter, void* stream) {
  B200_REQUIRE(cols % 8 == 0, "relu_bwd: cols %d must be a multiple of 8", cols);
  B200_REQUIRE(dbias == nullptr || (scratch != nullptr && counter != nullptr),
               "relu_bwd: bias gradient needs scratch (16 x cols floats) and a counter");
  if (rows == 0) return 0;
  dim3 grid(ceil_div(cols / 8, 32), ceil_div(rows, RELU_ROWS)), block(32, 8);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dy_fp32)
    relu_bwd_kernel<true><<<grid, block, 0, st>>>(dy, reinterpret_cast<const bf16*>(y_act),
                                                  reinterpret_cast<bf16*>(dz), dbias, rows, cols, cols, scratch, counter);
  else
    relu_bwd_kernel<false><<<grid, block, 0, st>>>(dy, reinterpret_cast<const bf16*>(y_act),
                                                   reinterpret_cast<bf16*>(dz), dbias, rows, cols, cols, scratch,
                                                   counter);
  B200_LAUNCH_CHECK();
  return 0;
}

// dz = dy * act'(.) as bf16, dbias += column sums of dz: act 1 = ReLU (y_or_z = activated output), 2 / 3 = GELU erf /
// tanh form (y_or_z = pre-activation).  b200_relu_bwd is the act = 1 case.  (reference models/gpt.py:14-15,35,38)
int b200_act_bwd(const void* dy, const void* y_or_z, void* dz, float* dbias, int rows, int cols, int dy_fp32, int act,
                 float* scratch, int* counter, void* stream) {
  B200_REQUIRE(cols % 8 == 0, "act_bwd: cols %d must be a multiple of 8", cols);
  B200_REQUIRE(act >= 1 && act <= 3, "act_bwd: activation %d not supported (1 relu, 2 gelu, 3 gelu-tanh)", act);
  B200_REQUIRE(dbias == nullptr || (scratch != nullptr && counter != nullptr),
               "act_bwd: bias gradient needs scratch (16 x cols floats) and a counter");
  if (rows == 0) return 0;
  dim3 grid(ceil_div(cols / 8, 32), ceil_div(rows, RELU_ROWS)), block(32, 8);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dy_fp32)
    relu_bwd_kernel<true><<<grid, block, 0, st>>>(dy, reinterpret_cast<const bf16*>(y_or_z), reinterpret_cast<bf16*>(dz),
                                                  dbias, rows, cols, cols, scratch, counter, act);
  else
    relu_bwd_kernel<false><<<grid, block, 0, st>>>(dy, reinterpret_cast<const bf16*>(y_or_z), reinterpret_cast<bf16*>(dz),
                                                   dbias, rows, cols, cols, scratch, counter, act);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_colsum_bf16(const void* x, float* out_accum, int rows, int cols, int ld, float* scratch,
                     int* counter, void* stream) {
  B200_REQUIRE(scratch != nullptr && counter != nullptr, "colsum: scratch and counter required");
  B200_REQUIRE(cols % 8 == 0 && ld % 8 == 0, "colsum: cols/ld must be multiples of 8");
  if (rows == 0) return 0;
  dim3 grid(ceil_div(cols / 8, 32), ceil_div(rows, RELU_ROWS)), block(32, 8);
  relu_bwd_kernel<false><<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(
      x, nullptr, nullptr, out_accum, rows, cols, ld, scratch, counter);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_count_valid(const int64_t* targets, float* stats, int ignore_index, int64_t n, void* stream) {
  count_valid_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(targets, stats, ignore_index, n);
  B200_LAUNCH_CHECK();
  return 0;
}

// LM head, step 1 of 4 (reference main-single.py:95-96 F.cross_entropy(ignore_index=-100)): stats[0] = number of
// targets != ignore_index, stats[1..3] = 0, and the per-weight state `lmstate` (4 floats, zero-initialised once by
// the caller and then left alone) is rotated for this call.
int b200_lmhead_begin(const int64_t* targets, float* stats, float* lmstate, int ignore_index, int64_t n,
                      void* stream) {
  B200_REQUIRE(stats != nullptr && lmstate != nullptr, "lmhead_begin: stats and lmstate are required");
  lm_begin_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(targets, stats, lmstate, ignore_index, n);
  B200_LAUNCH_CHECK();
  return 0;
}

// step 2: xnorm[m] = |X[m,:]|, lmstate[0] = max_v |W[v,:]| (bf16 operands, fp32 math)
int b200_lmhead_norms(const void* X, int ldx, const void* W, int ldw, int M, int V, int D, float* xnorm,
                      float* lmstate, void* stream) {
  B200_REQUIRE(D % 8 == 0 && ldx % 8 == 0 && ldw % 8 == 0, "lmhead_norms: D / ldx / ldw must be multiples of 8");
  B200_REQUIRE(((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(W)) & 15) == 0,
               "lmhead_norms: operands must be 16-byte aligned");
  const int grid = grid_for(static_cast<int64_t>(M + V) * 32, 256, 8);
  lm_norms_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(X), ldx, reinterpret_cast<const bf16*>(W), ldw, M, V, D, xnorm, lmstate);
  B200_LAUNCH_CHECK();
  return 0;
}

// step 4 (step 3 is b200_lmhead_gemm_bf16): stats[1] += sum of row losses, stats[2] += rows whose target holds the
// row maximum; with write_grad: rscale[m] = grad_scale / (n_valid * sum_m) (0 for ignored rows),
// E[m, target] -= sum_m, Xs[m,:] = rscale[m] * X[m,:], so that afterwards
//   dX = diag(rscale) . (E . W)      (b200_gemm_rowscale_bf16)      dW += E^T . Xs      (b200_gemm_bf16).
int b200_lmhead_finish(const void* part, int num_parts, int M, const float* xnorm, const float* tgt_val,
                       const int64_t* targets, int ignore_index, float* stats, float* lmstate, void* E, int lde,
                       int V, const void* X, int ldx, int D, const void* W, int ldw, void* Xs, float* rscale,
                       float grad_scale, int write_grad, void* stream) {
  B200_REQUIRE(D % 8 == 0 && ldx % 8 == 0 && ldw % 8 == 0, "lmhead_finish: D / ldx / ldw must be multiples of 8");
  B200_REQUIRE(!write_grad || (E != nullptr && Xs != nullptr && rscale != nullptr),
               "lmhead_finish: gradient outputs missing");
  if (M == 0) return 0;
  lm_finish_kernel<<<ceil_div(M, LMF_ROWS), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float2*>(part), num_parts, M, xnorm, tgt_val, targets, ignore_index, stats, lmstate,
      reinterpret_cast<bf16*>(E), lde, V, reinterpret_cast<const bf16*>(X), ldx, D, reinterpret_cast<const bf16*>(W),
      ldw, reinterpret_cast<bf16*>(Xs), rscale, grad_scale, write_grad);
  B200_LAUNCH_CHECK();
  return 0;
}

int b200_ce_fwd_bwd(void* logits, const int64_t* targets, float* stats, int rows, int vocab, int ld,
                    int ignore_index, int write_grad, float grad_scale, void* stream) {
  B200_REQUIRE(ld % 8 == 0 && ld >= vocab, "ce: ld %d must be a multiple of 8 and >= vocab %d", ld, vocab);
  const size_t smem = static_cast<size_t>(ld) * 2;
  B200_REQUIRE(smem <= 113 * 1024, "ce: vocab %d too large for the shared-memory row (max %d)", vocab, 113 * 512);
  B200_REQUIRE((reinterpret_cast<uintptr_t>(logits) & 15) == 0, "ce: logits must be 16-byte aligned");
  if (rows == 0) return 0;
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!configured[dev & 63]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(ce_fwd_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 113 * 1024));
    B200_CHECK_CUDA(cudaFuncSetAttribute(ce_fwd_bwd_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    configured[dev & 63] = true;
  }
  ce_fwd_bwd_kernel<<<rows, CE_THREADS, smem, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<bf16*>(logits), targets, stats, vocab, ld, ignore_index, write_grad, grad_scale);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// b200_relu_bwd_static: the ReLU-backward + bias-gradient kernel for callers that have no device
// memory of their own to lend.  b200_relu_bwd needs a scratch buffer (16 replica rows of column
// sums plus a ticket counter) from the caller; here that scratch is a fixed array that is part of
// the library image, so the call needs nothing but its inputs and outputs.
// ------------------------------------------------------------------------------------------
namespace {

constexpr int STATIC_MAX_COLS = 4096;  // widest matrix the built-in scratch can serve

struct StaticColsumScratch {
  float replicas[REPLICAS * STATIC_MAX_COLS];  // 256 KiB
  int counter;
  int pad[63];
};
// Lives in every GPU's memory from the moment the library is loaded on it until the process exits.
static __device__ __align__(256) StaticColsumScratch g_static_colsum_scratch;

// Device address of the array for each device, resolved on first use.  (In host code the symbol
// name itself is only a handle - it must go through cudaGetSymbolAddress, with the target device
// current, to become a pointer a kernel can use.)
StaticColsumScratch* g_static_scratch_ptr[64] = {nullptr};
std::mutex g_static_scratch_mu;

int static_scratch_for_current_device(StaticColsumScratch** out) {
  int dev = -1;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < 64, "relu_bwd_static: device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_scratch_mu);
  if (g_static_scratch_ptr[dev] == nullptr) {
    StaticColsumScratch* ptr = nullptr;
    B200_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_static_colsum_scratch));
    B200_REQUIRE(ptr != nullptr, "relu_bwd_static: could not resolve the scratch array on device %d", dev);
    g_static_scratch_ptr[dev] = ptr;
  }
  *out = g_static_scratch_ptr[dev];
  return 0;
}

}  // namespace

extern "C" int b200_relu_bwd_static(const void* dy, const void* y_act, void* dz, float* dbias, int rows,
                                    int cols, int dy_fp32, void* stream) {
  B200_REQUIRE(cols % 8 == 0, "relu_bwd_static: cols %d must be a multiple of 8", cols);
  B200_REQUIRE(cols <= STATIC_MAX_COLS,
               "relu_bwd_static: cols %d exceeds the built-in scratch (max %d columns); use b200_relu_bwd "
               "with a caller-provided scratch buffer instead", cols, STATIC_MAX_COLS);
  if (rows == 0) return 0;
  StaticColsumScratch* sc = nullptr;
  int rc = static_scratch_for_current_device(&sc);  // first call on a device resolves the address
  if (rc != 0) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // The array keeps whatever the previous call (or an interrupted one) left behind: the replica
  // rows this call will use and the ticket counter have to start from zero.
  B200_CHECK_CUDA(cudaMemsetAsync(sc->replicas, 0, sizeof(float) * REPLICAS * static_cast<size_t>(cols), st));
  B200_CHECK_CUDA(cudaMemsetAsync(&sc->counter, 0, sizeof(int), st));
  dim3 grid(ceil_div(cols / 8, 32), ceil_div(rows, RELU_ROWS)), block(32, 8);
  if (dy_fp32)
    relu_bwd_kernel<true><<<grid, block, 0, st>>>(dy, reinterpret_cast<const bf16*>(y_act),
                                                  reinterpret_cast<bf16*>(dz), dbias, rows, cols, cols,
                                                  sc->replicas, &sc->counter);
  else
    relu_bwd_kernel<false><<<grid, block, 0, st>>>(dy, reinterpret_cast<const bf16*>(y_act),
                                                   reinterpret_cast<bf16*>(dz), dbias, rows, cols, cols,
                                                   sc->replicas, &sc->counter);
  B200_LAUNCH_CHECK();
  return 0;
}
