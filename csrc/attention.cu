// Causal flash attention (forward + backward) on tcgen05 for sm_100a.
//
// Reference semantics: models/gpt.py:75-100  softmax(q k^T / sqrt(d) + causal(-1e9) [+ key
// padding]) v, with the S x S score matrix never written to memory and no host-built mask.
// Input is the packed projection output qkv[B, T, 3, H, D] (q | k | v sections of width H*D).
//
// Tiling: 128 query rows x 128 keys per step, head_dim D in {32, 64, 128}, any T (tails are
// zero-filled / masked).  All five backward GEMMs and both forward GEMMs run on tcgen05.mma
// (M = 128, fp32 accumulators in TMEM); softmax runs on the 128 threads of the CTA, one TMEM
// lane (= one score row) per thread.
//
// Shared-memory operand layout (one layout serves every use): a [128 x C] bf16 tile is
// stored as 16-byte chunks  offset(row, col) = (col/8) * 2048 + row * 16 + (col%8) * 2,
// i.e. the un-swizzled UMMA "core matrix" layout (8 rows x 16 B contiguous).  Read as a
// K-major operand (rows = M/N, cols = K):  LBO = 2048 (next 8 columns), SBO = 128 (next 8
// rows).  Read as an MN-major operand (cols = M/N, rows = K) - which is how P^T, dS^T, dO^T,
// Q^T, K^T and V^T are consumed - LBO = 128 (next 8 rows = next K group), SBO = 2048 (next 8
// columns = next MN group).  No tile is ever transposed.
//
// Forward, per CTA = (batch, head, 128-query tile): for each key tile up to the diagonal:
//   S = Q K^T (TMEM) -> online softmax in registers -> P (bf16, back into TMEM) -> PV = P V (A operand
//   from TMEM) ->
//   O = O * alpha + PV in registers.  Writes O (bf16) and LSE (fp32).
// Backward, per CTA = (batch, head, 128-key tile): loops over the query tiles at or below the
// diagonal; S and dP = dO V^T into TMEM, P and dS = P * (dP - delta) * scale to smem, then
// dV += P^T dO, dK += dS^T Q (TMEM accumulators across the loop) and dQ = dS K, which is
// reduced across key tiles with fp32 vector atomics into a scratch buffer (converted to bf16
// by a small second kernel).  delta = rowsum(dO * O) comes from a pre-pass.
#include <stdlib.h>

#include "../include/b200gpt.h"
#include "common.cuh"

namespace {

constexpr int TQ = 128;   // query rows per tile (= UMMA M = TMEM lanes)
constexpr int TK = 128;   // keys per tile
constexpr int ATTN_THREADS = 128;
constexpr uint32_t CHUNK_STRIDE = TQ * 16;  // 2048 B between 8-column groups
constexpr float kLog2e = 1.4426950408889634f;

enum : int { ERR_ATTN_FWD = 201, ERR_ATTN_BWD = 202 };

// no-swizzle descriptors (layout type 0)
__device__ __forceinline__ uint64_t desc_nosw(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// K-major view of a tile (rows = M/N): 16 K-columns per MMA = two 8-column chunks
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t tile, int kk) {
  return desc_nosw(tile + kk * 2 * CHUNK_STRIDE, CHUNK_STRIDE, 128);
}
// MN-major view (cols = M/N, rows = K): 16 K-rows per MMA = two 8-row groups (256 B)
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t tile, int kk) {
  return desc_nosw(tile + kk * 256, 128, CHUNK_STRIDE);
}

// global [rows x D] (row stride in elements) -> smem tile, rows >= nvalid zero-filled.
// Asynchronous copies (cp.async, 16 B each, zero-fill through src-size 0): a thread fires all its
// copies back to back and the caller waits once with load_tiles_wait(), so a tile costs one
// global-memory latency instead of one per 16-byte chunk.
template <int D, int NT = ATTN_THREADS>
__device__ __forceinline__ void load_tile(uint32_t smem_dst, const bf16* __restrict__ g, int nvalid,
                                          size_t row_stride) {
  constexpr int CPR = D / 8;  // 16-byte chunks per row
#pragma unroll
  for (int c = threadIdx.x; c < TQ * CPR; c += NT) {
    const int row = c / CPR, kc = c % CPR;
    const bool ok = row < nvalid;
    const bf16* src = g + static_cast<size_t>(ok ? row : 0) * row_stride + kc * 8;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_dst + kc * CHUNK_STRIDE + row * 16),
                 "l"(src), "r"(ok ? 16 : 0)
                 : "memory");
  }
}
__device__ __forceinline__ void load_tiles_wait() {
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d)
               : "memory");
}


// Bit j of the result is set when key (key0 + j) is visible to query row q: inside the sequence,
// not above the causal diagonal (only checked on diagonal tiles) and not a padding key.  Built once
// per 32-key chunk (one predicated byte load per lane + a ballot), so the per-element softmax
// code is a bit test and a select - no branches.
__device__ __forceinline__ uint32_t keep_bits(int key0, int q, bool diag, bool row_ok, int T,
                                              const uint8_t* __restrict__ mask_row) {
  int limit = T - key0;                       // keys < T
  if (diag) limit = min(limit, q + 1 - key0);  // keys <= q
  uint32_t bits = limit >= 32 ? 0xffffffffu : (limit <= 0 ? 0u : ((1u << limit) - 1u));
  if (mask_row != nullptr) {
    const int lane = threadIdx.x & 31;
    const int key = key0 + lane;
    const bool pad = key < T ? (mask_row[key] != 0) : true;
    bits &= ~__ballot_sync(0xffffffffu, pad);
  }
  return row_ok ? bits : 0u;
}

struct AttnParams {
  const bf16* qkv;     // [B, T, 3*H*D]
  bf16* out;           // [B, T, H*D]
  float* lse;          // [B, H, T]
  const bf16* dout;    // [B, T, H*D]
  bf16* dqkv;          // [B, T, 3*H*D]
  float* delta;        // [B, H, T]
  float* dq_acc;       // [B, T, H*D] fp32 scratch
  const uint8_t* key_mask;  // [B, T] 1 = padding, nullable
  int B, T, H;
  int tile_in_z;       // backward: key-tile index comes from blockIdx.z (see attn_bwd_kernel)
  float scale;
  int* err_word;
  long long* dbg;  // optional per-CTA phase timestamps (32 slots), bring-up only
};

long long* g_attn_dbg = nullptr;

__device__ __forceinline__ long long attn_global_ns() {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// Phase stamps (tools/attn_phases.py, tools/attn_bwd_phases.py) cost registers in kernels that sit at their
// occupancy cap (3.7 us per backward launch), so they are only compiled into the bring-up build
// (B200_GEMM_TRACE=1 python -m csrc.build).
#ifdef B200_GEMM_TRACE
#define ADBG(slot)                                                                                    \
  do {                                                                                                \
    if (p.dbg != nullptr && threadIdx.x == 0)                                                         \
      p.dbg[((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 32 + (slot)] = clock64(); \
  } while (0)
#define ADBG_NS(slot)                                                                                 \
  do {                                                                                                \
    if (p.dbg != nullptr && threadIdx.x == 0)                                                         \
      p.dbg[((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 32 + (slot)] = attn_global_ns(); \
  } while (0)
#else
#define ADBG(slot) do { } while (0)
#define ADBG_NS(slot) do { } while (0)
#endif

// ------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(ATTN_THREADS)
attn_fwd_kernel(const AttnParams p) {
  constexpr int TILE_BYTES = TQ * D * 2;
  // Tensor memory: S = Q K^T in columns [0,128) (fp32).  Each thread turns its S row into P (bf16, two keys
  // per 32-bit column) and stores it back over the columns it has already consumed, [0,64); the PV MMAs
  // take P straight from tensor memory (A operand in TMEM) and write to [PV_COL, PV_COL + D).  P never
  // touches shared memory, and for D <= 64 a CTA needs 128 columns and 3 * 128 * D * 2 bytes of smem, so
  // four CTAs fit an SM.
  constexpr uint32_t PV_COL = (D <= 64) ? 64 : 128;
  constexpr uint32_t TMEM_COLS = (D <= 64) ? 128 : 256;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;
  const uint32_t sQ = base, sK = sQ + TILE_BYTES, sV = sK + TILE_BYTES;
  const uint32_t bar = sV + TILE_BYTES;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(gen + 3 * TILE_BYTES + 8);
  volatile int* abort_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);

  const int tid = threadIdx.x, warp = tid >> 5;
  // grid = (H, B, query tiles): z is the slowest launch dimension, so the late query tiles (most key tiles)
  // of every (batch, head) are scheduled first and the short ones fill the tail
  const int qt = gridDim.z - 1 - blockIdx.z, h = blockIdx.x, b = blockIdx.y;
  const int HD = p.H * D;
  const size_t row_stride = static_cast<size_t>(3) * HD;
  const bf16* qkv_b = p.qkv + static_cast<size_t>(b) * p.T * row_stride;
  const int q0 = qt * TQ;
  const int q = q0 + tid;  // this thread's query row
  ADBG(0);
  ADBG_NS(24);

  if (tid == 0) {
    *abort_flag = 0;
    mbar_init(bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(tmem_ptr_smem), TMEM_COLS);
    tmem_relinquish();
  }
  // Q is waited for together with the first K/V tile (one global-memory latency for all three)
  load_tile<D>(sQ, qkv_b + static_cast<size_t>(q0) * row_stride + h * D, min(TQ, p.T - q0), row_stride);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *tmem_ptr_smem;
  const uint32_t t_lane = tmem + (static_cast<uint32_t>(warp * 32) << 16);
  const uint32_t idesc_s = umma_idesc_bf16(TQ, TK, 0, 0);
  const uint32_t idesc_pv = umma_idesc_bf16(TQ, D, 0, 1);
  const float sl2 = p.scale * kLog2e;
  const uint8_t* mask_row = p.key_mask != nullptr ? p.key_mask + static_cast<size_t>(b) * p.T : nullptr;
  ADBG(1);

  float o[D];
#pragma unroll
  for (int i = 0; i < D; ++i) o[i] = 0.f;
  float m_run = -INFINITY, l_run = 0.f;
  uint32_t phase = 0;
  bool ok = true;

  for (int kt = 0; kt <= qt && ok; ++kt) {
    const int k0 = kt * TK;
    __syncthreads();  // previous iteration's MMAs are complete (we waited); smem is reusable
    load_tile<D>(sK, qkv_b + static_cast<size_t>(k0) * row_stride + HD + h * D, min(TK, p.T - k0), row_stride);
    load_tile<D>(sV, qkv_b + static_cast<size_t>(k0) * row_stride + 2 * HD + h * D, min(TK, p.T - k0), row_stride);
    load_tiles_wait();
    fence_proxy_async_smem();
    __syncthreads();
    if (kt < 2) ADBG(2 + 6 * kt);
    if (tid == 0) {
      tcgen05_fence_after();
#pragma unroll
      for (int kk = 0; kk < D / 16; ++kk)
        umma_f16_ss(tmem, desc_kmajor(sQ, kk), desc_kmajor(sK, kk), idesc_s, kk > 0 ? 1u : 0u);
      umma_commit(bar);
    }
    if (!mbar_wait(bar, phase, abort_flag, p.err_word, ERR_ATTN_FWD)) { ok = false; }
    phase ^= 1;
    if (!__syncthreads_and(ok)) { ok = false; break; }
    tcgen05_fence_after();
    if (kt < 2) ADBG(3 + 6 * kt);

    // pass A: row max of the masked, scaled scores
    const bool diag = (kt == qt);
    float m_tile = -INFINITY;
#pragma unroll 1
    for (int c = 0; c < TK / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(t_lane + c * 32, r);
      tmem_ld_wait();
      const uint32_t kb = keep_bits(k0 + c * 32, q, diag, true, p.T, mask_row);
#pragma unroll
      for (int j = 0; j < 32; ++j)
        m_tile = fmaxf(m_tile, ((kb >> j) & 1u) ? __uint_as_float(r[j]) : -INFINITY);
    }
    if (kt < 2) ADBG(4 + 6 * kt);
    const float m_new = fmaxf(m_run, m_tile * p.scale);  // scale > 0
    const float m_safe = (m_new == -INFINITY) ? 0.f : m_new;
    const float alpha = fast_exp2((m_run - m_safe) * kLog2e);  // m_run = -inf -> 0
    const float m_l2 = m_safe * kLog2e;
    // pass B: p = exp(s - m), row sum, P (bf16) back into tensor memory
    float l_tile = 0.f;
#pragma unroll 1
    for (int c = 0; c < TK / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(t_lane + c * 32, r);
      tmem_ld_wait();
      uint32_t pk[16];
      const uint32_t kb = keep_bits(k0 + c * 32, q, diag, true, p.T, mask_row);
#pragma unroll
      for (int j = 0; j < 32; j += 2) {
        float pv[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const float e = fast_exp2(fmaf(__uint_as_float(r[j + u]), sl2, -m_l2));
          pv[u] = ((kb >> (j + u)) & 1u) ? e : 0.f;
        }
        // sum what the tensor core will actually see (bf16-rounded), like the reference bf16 path
        const uint32_t packed = pack_bf16x2(pv[0], pv[1]);
        const float2 back = unpack_bf16x2(packed);
        l_tile += back.x + back.y;
        pk[j >> 1] = packed;
      }
      tmem_st_32x32b_x16(t_lane + c * 16, pk);   // keys [32c, 32c+32) -> columns [16c, 16c+16): already consumed
    }
    tmem_st_wait();
    l_run = l_run * alpha + l_tile;
    m_run = m_new;
    tcgen05_fence_before();
    __syncthreads();
    if (kt < 2) ADBG(5 + 6 * kt);
    if (tid == 0) {
      tcgen05_fence_after();
#pragma unroll
      for (int kk = 0; kk < TK / 16; ++kk)
        umma_f16_ts(tmem + PV_COL, tmem + kk * 8, desc_mnmajor(sV, kk), idesc_pv, kk > 0 ? 1u : 0u);
      umma_commit(bar);
    }
    if (!mbar_wait(bar, phase, abort_flag, p.err_word, ERR_ATTN_FWD)) { ok = false; }
    phase ^= 1;
    if (!__syncthreads_and(ok)) { ok = false; break; }
    tcgen05_fence_after();
    if (kt < 2) ADBG(6 + 6 * kt);
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(t_lane + PV_COL + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) o[c * 32 + j] = o[c * 32 + j] * alpha + __uint_as_float(r[j]);
    }
    tcgen05_fence_before();
    if (kt < 2) ADBG(7 + 6 * kt);
  }

  if (ok && q < p.T) {
    const float inv_l = l_run > 0.f ? 1.0f / l_run : 0.f;
    bf16* orow = p.out + (static_cast<size_t>(b) * p.T + q) * HD + h * D;
#pragma unroll
    for (int j = 0; j < D; j += 8) {
      *reinterpret_cast<uint4*>(orow + j) =
          make_uint4(pack_bf16x2(o[j] * inv_l, o[j + 1] * inv_l), pack_bf16x2(o[j + 2] * inv_l, o[j + 3] * inv_l),
                     pack_bf16x2(o[j + 4] * inv_l, o[j + 5] * inv_l), pack_bf16x2(o[j + 6] * inv_l, o[j + 7] * inv_l));
    }
    p.lse[(static_cast<size_t>(b) * p.H + h) * p.T + q] = l_run > 0.f ? m_run + logf(l_run) : -INFINITY;
  }
  ADBG(20);
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc(tmem, TMEM_COLS);
  }
  ADBG_NS(25);
}

// ------------------------------------------------------------------------------------------
// forward, warp-specialised and software-pipelined (long sequences: >= 4 key tiles per head)
// ------------------------------------------------------------------------------------------
// One CTA per (batch, head, 128-query tile), 192 threads:
//   warp 0   loader: one lane issues TMA loads (4-D tensor map over qkv: box {8 elements, 128 rows, D/8 column
//            groups, 1 batch} lands as the un-swizzled core-matrix layout in one instruction, rows past T are
//            zero-filled); Q once, K / V tiles through a 2-stage ring
//   warp 1   MMA issuer (one lane): S(j+1) = Q K(j+1)^T is issued before it waits for P(j), so the tensor pipe
//            works on the next tile's scores while the softmax warps are busy; then O += P(j) V(j) with P read
//            from tensor memory
//   warps 2-5  softmax: one score row per thread, the whole 128-key row in registers (one pass: max, exp2,
//            row sum), P written back as bf16 over the S columns it came from
// Tensor memory: S double-buffered [0,128) [128,256), O accumulates in [256, 256 + D) across all key tiles.  The
// running maximum is only raised when a tile exceeds it by more than 2^8 (then O and the row sum are rescaled
// in tensor memory, which needs the previous PV to have retired) - rescaling is rare after the first tiles.
enum : int { ERR_ATTN_WS = 203, ERR_ATTN_SMEM_ALIGN = 204 };

// wait / phase accounting of the warp-specialised kernels (trace build only: B200_GEMM_TRACE=1 python -m csrc.build;
// tools/attn_ws_phases.py): cycles a role spent in each wait and in its own work, dumped per CTA into p.dbg
#ifdef B200_GEMM_TRACE
#define WS_NOW() clock64()
#define WS_T0() _ws_t0 = clock64()
#define WS_ACC(v) (v) += clock64() - _ws_t0
#define WS_DUMP(slot, v)                                                                                         \
  do {                                                                                                           \
    if (p.dbg != nullptr && lane == 0)                                                                           \
      p.dbg[((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 32 + (slot)] = (long long)(v);   \
  } while (0)
#else
#define WS_NOW() 0LL
#define WS_T0() do { } while (0)
#define WS_ACC(v) do { } while (0)
#define WS_DUMP(slot, v) do { } while (0)
#endif
constexpr int ATTN_WS_THREADS = 192;

// Tile layout of the warp-specialised kernels: a [128 x D] bf16 tile is D / 64 blocks of [128 rows x 64 columns]
// (16 KB: rows of 128 bytes whose 16-byte chunks are XOR-swizzled with row % 8) - exactly what ONE TMA box
// {64 columns, 128 rows} with SWIZZLE_128B writes, so a tile costs D / 64 bulk tensor loads of full 128-byte rows
// (the un-swizzled core-matrix layout of the per-tile kernels would cost 16 x as many 16-byte rows).  The same block
// serves as a K-major operand (MN = rows, K = columns) and as an MN-major operand (K = rows, MN = columns).
constexpr uint32_t SW_BLK = 16384;
__device__ __forceinline__ uint64_t desc_sw_k(uint32_t tile, int kk) {    // 16 K-columns per MMA, 4 MMAs per block
  return umma_desc_at(umma_smem_desc_base(16, 1024), tile + (kk >> 2) * SW_BLK + (kk & 3) * 32);
}
__device__ __forceinline__ uint64_t desc_sw_mn(uint32_t tile, int kk) {   // 16 K-rows per MMA; 64-wide MN groups one block apart
  return umma_desc_at(umma_smem_desc_base(SW_BLK, 1024), tile + kk * 2048);
}
// one tile = D / 64 TMA boxes; coordinates: (column, row in the sequence, batch)
template <int D>
__device__ __forceinline__ void tma_load_tile(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int col0, int t0, int b) {
#pragma unroll
  for (int i = 0; i < D / 64; ++i) tma_load_3d(dst + i * SW_BLK, tm, bar, col0 + 64 * i, t0, b);
}

template <int D>
__global__ void __launch_bounds__(ATTN_WS_THREADS, 1)
attn_fwd_ws_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnParams p) {
  constexpr int TILE_BYTES = TQ * D * 2;
  constexpr uint32_t TMEM_COLS = 512;
  constexpr uint32_t C_O = 256;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);      // swizzled blocks need 1024-byte alignment (checked below)
  // K and V travel through separate NST-deep rings: a K slot is free again as soon as the S MMAs that read it have
  // retired, a V slot only after its PV - with one shared 2-deep ring the load of tile j + 2 could not start before
  // PV(j) and its whole latency sat in front of QK(j + 2)
  constexpr int NST = 3;
  const uint32_t sQ = base, sK = sQ + TILE_BYTES, sV = sK + NST * TILE_BYTES;
  const uint32_t bars = sV + NST * TILE_BYTES;
  const uint32_t q_full = bars, k_full = bars + 8, k_empty = bars + 32, v_full = bars + 56, v_empty = bars + 80,
                 s_full = bars + 104, p_ready = bars + 120, o_done = bars + 136;
  uint8_t* gen = smem_raw;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(gen + (1 + 2 * NST) * TILE_BYTES + 144);
  volatile int* abort_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);
  if ((base & 1023u) != 0u) {
    if (threadIdx.x == 0) atomicCAS(p.err_word, 0, ERR_ATTN_SMEM_ALIGN);
    return;
  }

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = gridDim.z - 1 - blockIdx.z, h = blockIdx.x, b = blockIdx.y;   // long query tiles first
  const int HD = p.H * D;
  const int q0 = qt * TQ;
  const int n_kt = qt + 1;

  if (threadIdx.x == 0) {
    *abort_flag = 0;
    mbar_init(q_full, 1);
    for (int i = 0; i < NST; ++i) {
      mbar_init(k_full + 8 * i, 1);
      mbar_init(k_empty + 8 * i, 1);
      mbar_init(v_full + 8 * i, 1);
      mbar_init(v_empty + 8 * i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(s_full + 8 * i, 1);
      mbar_init(p_ready + 8 * i, 4);
    }
    mbar_init(o_done, 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmQKV);
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_ptr_smem), TMEM_COLS);
    tmem_relinquish();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *tmem_ptr_smem;

  if (warp == 0) {
    // ===================== loader =====================
    if (lane == 0) {
      mbar_arrive_expect_tx(q_full, TILE_BYTES);
      tma_load_tile<D>(sQ, &tmQKV, q_full, h * D, q0, b);
      for (int j = 0; j < n_kt; ++j) {
        const int st = j % NST;
        const uint32_t use = static_cast<uint32_t>(j / NST);
        if (j >= NST && !mbar_wait(k_empty + 8 * st, (use - 1) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) break;
        mbar_arrive_expect_tx(k_full + 8 * st, TILE_BYTES);
        tma_load_tile<D>(sK + st * TILE_BYTES, &tmQKV, k_full + 8 * st, HD + h * D, j * TK, b);
        if (j >= NST && !mbar_wait(v_empty + 8 * st, (use - 1) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) break;
        mbar_arrive_expect_tx(v_full + 8 * st, TILE_BYTES);
        tma_load_tile<D>(sV + st * TILE_BYTES, &tmQKV, v_full + 8 * st, 2 * HD + h * D, j * TK, b);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t idesc_s = umma_idesc_bf16(TQ, TK, 0, 0);
    const uint32_t idesc_pv = umma_idesc_bf16(TQ, D, 0, 1);
    bool ok = mbar_wait(q_full, 0, abort_flag, p.err_word, ERR_ATTN_WS);
    auto issue_qk = [&](int j) {   // S[j & 1] = Q K(j)^T
      const int s = j & 1, st = j % NST;
      if (!mbar_wait(k_full + 8 * st, (j / NST) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) return false;
      tcgen05_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16_ss(tmem + s * 128, desc_sw_k(sQ, kk), desc_sw_k(sK + st * TILE_BYTES, kk), idesc_s, kk > 0 ? 1u : 0u);
        umma_commit(k_empty + 8 * st);
        umma_commit(s_full + 8 * s);
      }
      __syncwarp();
      return true;
    };
    if (ok) ok = issue_qk(0);
    for (int j = 0; j < n_kt && ok; ++j) {
      const int s = j & 1, st = j % NST;
      if (j + 1 < n_kt) ok = issue_qk(j + 1);
      if (!ok || !mbar_wait(p_ready + 8 * s, (j >> 1) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
      if (!mbar_wait(v_full + 8 * st, (j / NST) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
      tcgen05_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int kk = 0; kk < TK / 16; ++kk)   // O (+)= P(j) V(j), P from tensor memory
          umma_f16_ts(tmem + C_O, tmem + s * 128 + kk * 8, desc_sw_mn(sV + st * TILE_BYTES, kk), idesc_pv,
                      (j > 0 || kk > 0) ? 1u : 0u);
        umma_commit(v_empty + 8 * st);
        umma_commit(o_done);
      }
      __syncwarp();
    }
  } else {
    // ===================== softmax (4 warps, one row per thread) =====================
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int q = q0 + row;
    const uint32_t t_lane = tmem + (static_cast<uint32_t>(quarter * 32) << 16);
    const float sl2 = p.scale * kLog2e;
    const uint8_t* mask_row = p.key_mask != nullptr ? p.key_mask + static_cast<size_t>(b) * p.T : nullptr;
    float m_used = -INFINITY, l_run = 0.f;
    bool ok = true;
    for (int j = 0; j < n_kt; ++j) {
      const int s = j & 1;
      if (!mbar_wait(s_full + 8 * s, (j >> 1) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
      tcgen05_fence_after();
      uint32_t r[128];
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld_32x32b_x32(t_lane + s * 128 + c * 32, r + c * 32);
      tmem_ld_wait();
      const bool diag = (j == qt);
      const bool masked_tile = diag || mask_row != nullptr;
      uint32_t kb[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
      if (masked_tile) {
#pragma unroll
        for (int c = 0; c < 4; ++c) kb[c] = keep_bits(j * TK + c * 32, q, diag, true, p.T, mask_row);
      }
      // (masked and unmasked tiles run separate straight-line loops: a per-element branch or select on a run-time
      // flag would cut the unrolled code into two-element pieces and leave the single warp of each scheduler
      // waiting out every MUFU latency)
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      if (!masked_tile) {
#pragma unroll
        for (int i = 0; i < 128; ++i) mx[i & 3] = fmaxf(mx[i & 3], __uint_as_float(r[i]));
      } else {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            mx[i & 3] = fmaxf(mx[i & 3], ((kb[c] >> i) & 1u) ? __uint_as_float(r[c * 32 + i]) : -INFINITY);
        }
      }
      float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      m_tile *= sl2;                                    // log2 domain (scale > 0)
      // raise the reference maximum only when this tile exceeds it by 2^8 (or there is none yet)
      const bool raise = m_tile > m_used + 8.0f || (m_used == -INFINITY && m_tile > -INFINITY);
      // o_done is a parity barrier: a waiter must never fall two phases behind it.  Every iteration j >= 1
      // therefore observes phase j - 1 (PV(j - 1) retired) exactly once, and does so BEFORE it reports P(j) - PV(j)
      // cannot be issued earlier, so the barrier is at phase j - 1 or j when it is tested.  By then that MMA has
      // long finished (it was issued when this tile's softmax began), so the wait costs nothing.
      bool seen_prev = false;
      if (__any_sync(0xffffffffu, raise)) {
        const float m_new = raise ? m_tile : m_used;
        const float alpha = (m_used == -INFINITY) ? 0.f : fast_exp2(m_used - m_new);   // 1 for rows that keep theirs
        if (j > 0) {
          // O of the previous tiles is rescaled in tensor memory: PV(j - 1) must have retired
          if (!mbar_wait(o_done, (j - 1) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
          seen_prev = true;
          tcgen05_fence_after();
#pragma unroll
          for (int c = 0; c < D / 16; ++c) {
            uint32_t o[16];
            tmem_ld_32x32b_x16(t_lane + C_O + c * 16, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32b_x16(t_lane + C_O + c * 16, o);
          }
        }
        l_run *= alpha;
        m_used = m_new;
      }
      const float m_sub = (m_used == -INFINITY) ? 0.f : m_used;
      float ls[4] = {0.f, 0.f, 0.f, 0.f};
      if (!masked_tile) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t pk[16];
          float e[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) e[i] = fast_exp2(fmaf(__uint_as_float(r[c * 32 + i]), sl2, -m_sub));
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            pk[i >> 1] = pack_bf16x2(e[i], e[i + 1]);
            ls[(i >> 1) & 3] += e[i] + e[i + 1];
          }
          tmem_st_32x32b_x16(t_lane + s * 128 + c * 16, pk);   // keys [32c, 32c + 32) -> columns [16c, 16c + 16)
        }
      } else {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t pk[16];
          float e[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float x = fast_exp2(fmaf(__uint_as_float(r[c * 32 + i]), sl2, -m_sub));
            e[i] = ((kb[c] >> i) & 1u) ? x : 0.f;
          }
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            pk[i >> 1] = pack_bf16x2(e[i], e[i + 1]);
            ls[(i >> 1) & 3] += e[i] + e[i + 1];
          }
          tmem_st_32x32b_x16(t_lane + s * 128 + c * 16, pk);
        }
      }
      const float l_tile = (ls[0] + ls[1]) + (ls[2] + ls[3]);
      l_run += l_tile;
      tmem_st_wait();
      if (j > 0 && !seen_prev && !mbar_wait(o_done, (j - 1) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_ready + 8 * s);
    }
    if (ok && mbar_wait(o_done, (n_kt - 1) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) {
      tcgen05_fence_after();
      const float inv_l = l_run > 0.f ? 1.0f / l_run : 0.f;
      bf16* orow = p.out + (static_cast<size_t>(b) * p.T + (q < p.T ? q : 0)) * HD + h * D;
#pragma unroll
      for (int c = 0; c < D / 16; ++c) {
        uint32_t o[16];
        tmem_ld_32x32b_x16(t_lane + C_O + c * 16, o);
        tmem_ld_wait();
        if (q < p.T) {
#pragma unroll
          for (int i = 0; i < 16; i += 8)
            *reinterpret_cast<uint4*>(orow + c * 16 + i) = make_uint4(
                pack_bf16x2(__uint_as_float(o[i]) * inv_l, __uint_as_float(o[i + 1]) * inv_l),
                pack_bf16x2(__uint_as_float(o[i + 2]) * inv_l, __uint_as_float(o[i + 3]) * inv_l),
                pack_bf16x2(__uint_as_float(o[i + 4]) * inv_l, __uint_as_float(o[i + 5]) * inv_l),
                pack_bf16x2(__uint_as_float(o[i + 6]) * inv_l, __uint_as_float(o[i + 7]) * inv_l));
        }
      }
      if (q < p.T)
        p.lse[(static_cast<size_t>(b) * p.H + h) * p.T + q] =
            l_run > 0.f ? m_used * (1.0f / kLog2e) + logf(l_run) : -INFINITY;
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(256)
attn_delta_kernel(const bf16* __restrict__ out, const bf16* __restrict__ dout, float* __restrict__ delta,
                  int B, int T, int H, float* __restrict__ dq_zero = nullptr) {
  // dq_zero (optional): the fp32 dQ accumulator [B, T, H*D] the backward kernel reduces into - cleared here instead
  // of by a separate memset launch
  const int64_t total = static_cast<int64_t>(B) * T * H;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int h = static_cast<int>(i % H);
    const int64_t bt = i / H;
    const int t = static_cast<int>(bt % T);
    const int b = static_cast<int>(bt / T);
    const uint4* o4 = reinterpret_cast<const uint4*>(out + (bt * H + h) * D);
    const uint4* d4 = reinterpret_cast<const uint4*>(dout + (bt * H + h) * D);
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < D / 8; ++c) {
      const uint4 a = __ldg(o4 + c), g = __ldg(d4 + c);
      const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float2 x = unpack_bf16x2(aw[u]), y = unpack_bf16x2(gw[u]);
        s += x.x * y.x + x.y * y.y;
      }
    }
    delta[(static_cast<int64_t>(b) * H + h) * T + t] = s;
  }
  if (dq_zero != nullptr) {   // flat, fully coalesced 16-byte stores
    float4* z = reinterpret_cast<float4*>(dq_zero);
    const int64_t nz = total * (D / 4);
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < nz;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x)
      z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// 256 threads: threads t and t + 128 share score row t (same TMEM lane) and split its 128 keys - and the
// D output columns of dQ / dK / dV - in halves.  The softmax statistics are inputs here (lse, delta), so the
// two halves never have to talk to each other, and the exp / pack / store work per thread is halved.
constexpr int ATTN_BWD_THREADS = 256;

template <int D>
__global__ void __launch_bounds__(ATTN_BWD_THREADS, (128 + 2 * D) <= 256 ? 2 : 1)
attn_bwd_kernel(const AttnParams p) {
  constexpr int TILE_BYTES = TQ * D * 2;
  constexpr int SS_BYTES = TQ * TK * 2;
  // TMEM columns.  D <= 64 ("compact"): S, then dP, then dQ reuse [0,128) one after the other and
  // dV / dK follow at 128 / 128+D - 256 columns, so two CTAs share an SM and hide each other's
  // MMA / softmax latencies.  D = 128: S/dQ [0,128) dP [128,256) dV [256,384) dK [384,512).
  constexpr bool COMPACT = (128 + 2 * D) <= 256;
  constexpr uint32_t TMEM_COLS = COMPACT ? 256 : 512;
  constexpr uint32_t C_S = 0, C_DP = COMPACT ? 0 : 128, C_DV = COMPACT ? 128 : 256, C_DK = C_DV + D;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;
  const uint32_t sK = base, sV = sK + TILE_BYTES, sQ = sV + TILE_BYTES, sdO = sQ + TILE_BYTES;
  const uint32_t sP = sdO + TILE_BYTES, sdS = sP + SS_BYTES;
  const uint32_t bar = sdS + SS_BYTES;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(gen + 4 * TILE_BYTES + 2 * SS_BYTES + 8);
  volatile int* abort_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);

  const int tid = threadIdx.x & 127, warp = (threadIdx.x >> 5) & 3;   // score row / TMEM lane quarter
  const int half = threadIdx.x >> 7;                                   // which half of the keys / columns
  constexpr int CH = TK / 32 / 2;                                      // 32-key chunks per thread
  // short sequences: grid = (H, B, key tiles), so the early key tiles (most query tiles to visit) are all
  // scheduled first; long ones: grid = (key tiles, H, B), the key tiles of one head run together and share
  // its Q / dO in L2
  const bool kz = p.tile_in_z != 0;
  const int kt = kz ? blockIdx.z : blockIdx.x, h = kz ? blockIdx.x : blockIdx.y, b = kz ? blockIdx.y : blockIdx.z;
  const int HD = p.H * D;
  const size_t row_stride = static_cast<size_t>(3) * HD;
  const bf16* qkv_b = p.qkv + static_cast<size_t>(b) * p.T * row_stride;
  const bf16* dout_b = p.dout + static_cast<size_t>(b) * p.T * HD;
  const int k0 = kt * TK;
  const int nq_tiles = (p.T + TQ - 1) / TQ;
  ADBG(0);
  ADBG_NS(24);

  if (threadIdx.x == 0) {
    *abort_flag = 0;
    mbar_init(bar, 1);
    fence_barrier_init();
  }
  if (threadIdx.x < 32) {
    tmem_alloc(smem_u32(tmem_ptr_smem), TMEM_COLS);
    tmem_relinquish();
  }
  load_tile<D, ATTN_BWD_THREADS>(sK, qkv_b + static_cast<size_t>(k0) * row_stride + HD + h * D, min(TK, p.T - k0), row_stride);
  load_tile<D, ATTN_BWD_THREADS>(sV, qkv_b + static_cast<size_t>(k0) * row_stride + 2 * HD + h * D, min(TK, p.T - k0), row_stride);
  // the first query tile's Q / dO ride on the same global-memory latency as K / V
  load_tile<D, ATTN_BWD_THREADS>(sQ, qkv_b + static_cast<size_t>(k0) * row_stride + h * D, min(TQ, p.T - k0), row_stride);
  load_tile<D, ATTN_BWD_THREADS>(sdO, dout_b + static_cast<size_t>(k0) * HD + h * D, min(TQ, p.T - k0), HD);
  load_tiles_wait();
  fence_proxy_async_smem();
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *tmem_ptr_smem;
  const uint32_t t_lane = tmem + (static_cast<uint32_t>(warp * 32) << 16);
  const uint32_t idesc_ss = umma_idesc_bf16(TQ, TK, 0, 0);    // S, dP: K-major x K-major, N = 128
  const uint32_t idesc_tt = umma_idesc_bf16(TK, D, 1, 1);     // dV, dK: MN-major x MN-major, N = D
  const uint32_t idesc_dq = umma_idesc_bf16(TQ, D, 0, 1);     // dQ: K-major x MN-major, N = D
  const float sl2 = p.scale * kLog2e;
  const uint8_t* mask_row = p.key_mask != nullptr ? p.key_mask + static_cast<size_t>(b) * p.T : nullptr;
  ADBG(1);

  uint32_t phase = 0;
  bool ok = true;
  for (int qt = kt; qt < nq_tiles && ok; ++qt) {
    [[maybe_unused]] const int it = qt - kt;   // bring-up stamps cover the first two iterations
    const int q0 = qt * TQ;
    const int q = q0 + tid;
    const bool q_ok = q < p.T;
    if (qt > kt) {   // (the diagonal tile was loaded with K / V above)
      __syncthreads();
      load_tile<D, ATTN_BWD_THREADS>(sQ, qkv_b + static_cast<size_t>(q0) * row_stride + h * D, min(TQ, p.T - q0), row_stride);
      load_tile<D, ATTN_BWD_THREADS>(sdO, dout_b + static_cast<size_t>(q0) * HD + h * D, min(TQ, p.T - q0), HD);
      load_tiles_wait();
      fence_proxy_async_smem();
      __syncthreads();
    }
    if (it < 2) ADBG(2 + 8 * it);
    if (threadIdx.x == 0) {
      tcgen05_fence_after();
#pragma unroll
      for (int kk = 0; kk < D / 16; ++kk)
        umma_f16_ss(tmem + C_S, desc_kmajor(sQ, kk), desc_kmajor(sK, kk), idesc_ss, kk > 0 ? 1u : 0u);
      if constexpr (!COMPACT) {
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16_ss(tmem + C_DP, desc_kmajor(sdO, kk), desc_kmajor(sV, kk), idesc_ss, kk > 0 ? 1u : 0u);
      }
      umma_commit(bar);
    }
    const size_t stat_idx = (static_cast<size_t>(b) * p.H + h) * p.T + (q_ok ? q : 0);
    const float lse_l2 = q_ok ? p.lse[stat_idx] * kLog2e : 0.f;
    const float delta = q_ok ? p.delta[stat_idx] : 0.f;
    if (!mbar_wait(bar, phase, abort_flag, p.err_word, ERR_ATTN_BWD)) { ok = false; }
    phase ^= 1;
    if (!__syncthreads_and(ok)) { ok = false; break; }
    tcgen05_fence_after();
    if (it < 2) ADBG(3 + 8 * it);

    const bool diag = (qt == kt);
    if constexpr (COMPACT) {
      // P from S (kept packed in registers and written to smem), then dP into the same columns
      uint32_t pkeep[CH * 16];
#pragma unroll
      for (int cc = 0; cc < CH; ++cc) {
        const int c = half * CH + cc;
        uint32_t rs[32];
        tmem_ld_32x32b_x32(t_lane + C_S + c * 32, rs);
        tmem_ld_wait();
        const uint32_t kb = keep_bits(k0 + c * 32, q, diag, q_ok, p.T, mask_row);
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          const float e0 = fast_exp2(fmaf(__uint_as_float(rs[j]), sl2, -lse_l2));
          const float e1 = fast_exp2(fmaf(__uint_as_float(rs[j + 1]), sl2, -lse_l2));
          pkeep[cc * 16 + (j >> 1)] = pack_bf16x2(((kb >> j) & 1u) ? e0 : 0.f, ((kb >> (j + 1)) & 1u) ? e1 : 0.f);
        }
#pragma unroll
        for (int g = 0; g < 4; ++g)
          st_shared_v4(sP + (c * 4 + g) * CHUNK_STRIDE + tid * 16, pkeep[cc * 16 + 4 * g], pkeep[cc * 16 + 4 * g + 1],
                       pkeep[cc * 16 + 4 * g + 2], pkeep[cc * 16 + 4 * g + 3]);
      }
      tcgen05_fence_before();
      __syncthreads();                     // every thread has read S: dP may overwrite the columns
      if (it < 2) ADBG(4 + 8 * it);
      if (threadIdx.x == 0) {
        tcgen05_fence_after();
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16_ss(tmem + C_DP, desc_kmajor(sdO, kk), desc_kmajor(sV, kk), idesc_ss, kk > 0 ? 1u : 0u);
        umma_commit(bar);
      }
      if (!mbar_wait(bar, phase, abort_flag, p.err_word, ERR_ATTN_BWD)) { ok = false; }
      phase ^= 1;
      if (!__syncthreads_and(ok)) { ok = false; break; }
      tcgen05_fence_after();
      if (it < 2) ADBG(5 + 8 * it);
#pragma unroll
      for (int cc = 0; cc < CH; ++cc) {
        const int c = half * CH + cc;
        uint32_t rd[32], dk[16];
        tmem_ld_32x32b_x32(t_lane + C_DP + c * 32, rd);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          const float2 pr = unpack_bf16x2(pkeep[cc * 16 + (j >> 1)]);
          dk[j >> 1] = pack_bf16x2(pr.x * (__uint_as_float(rd[j]) - delta) * p.scale,
                                   pr.y * (__uint_as_float(rd[j + 1]) - delta) * p.scale);
        }
#pragma unroll
        for (int g = 0; g < 4; ++g)
          st_shared_v4(sdS + (c * 4 + g) * CHUNK_STRIDE + tid * 16, dk[4 * g], dk[4 * g + 1], dk[4 * g + 2], dk[4 * g + 3]);
      }
    } else {
#pragma unroll 1
      for (int cc = 0; cc < CH; ++cc) {
        const int c = half * CH + cc;
        uint32_t rs[32], rd[32];
        tmem_ld_32x32b_x32(t_lane + C_S + c * 32, rs);
        tmem_ld_32x32b_x32(t_lane + C_DP + c * 32, rd);
        tmem_ld_wait();
        uint32_t pk[16], dk[16];
        const uint32_t kb = keep_bits(k0 + c * 32, q, diag, q_ok, p.T, mask_row);
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          float pv[2], dsv[2];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const float e = fast_exp2(fmaf(__uint_as_float(rs[j + u]), sl2, -lse_l2));
            const float pr = ((kb >> (j + u)) & 1u) ? e : 0.f;
            pv[u] = pr;
            dsv[u] = pr * (__uint_as_float(rd[j + u]) - delta) * p.scale;
          }
          pk[j >> 1] = pack_bf16x2(pv[0], pv[1]);
          dk[j >> 1] = pack_bf16x2(dsv[0], dsv[1]);
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const uint32_t off = (c * 4 + g) * CHUNK_STRIDE + tid * 16;
          st_shared_v4(sP + off, pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
          st_shared_v4(sdS + off, dk[4 * g], dk[4 * g + 1], dk[4 * g + 2], dk[4 * g + 3]);
        }
      }
    }
    tcgen05_fence_before();
    fence_proxy_async_smem();
    __syncthreads();
    if (it < 2) ADBG(6 + 8 * it);
    if (threadIdx.x == 0) {
      tcgen05_fence_after();
      const uint32_t acc0 = (qt > kt) ? 1u : 0u;
#pragma unroll
      for (int kk = 0; kk < TQ / 16; ++kk)  // dV[key, d] += P^T dO   (reduction over the 128 queries)
        umma_f16_ss(tmem + C_DV, desc_mnmajor(sP, kk), desc_mnmajor(sdO, kk), idesc_tt, (kk > 0) ? 1u : acc0);
#pragma unroll
      for (int kk = 0; kk < TQ / 16; ++kk)  // dK[key, d] += dS^T Q
        umma_f16_ss(tmem + C_DK, desc_mnmajor(sdS, kk), desc_mnmajor(sQ, kk), idesc_tt, (kk > 0) ? 1u : acc0);
#pragma unroll
      for (int kk = 0; kk < TK / 16; ++kk)  // dQ[q, d] = dS K          (reduction over the 128 keys)
        umma_f16_ss(tmem + C_S, desc_kmajor(sdS, kk), desc_mnmajor(sK, kk), idesc_dq, kk > 0 ? 1u : 0u);
      umma_commit(bar);
    }
    if (!mbar_wait(bar, phase, abort_flag, p.err_word, ERR_ATTN_BWD)) { ok = false; }
    phase ^= 1;
    if (!__syncthreads_and(ok)) { ok = false; break; }
    tcgen05_fence_after();
    if (it < 2) ADBG(7 + 8 * it);
    float* dq_row = p.dq_acc + (static_cast<size_t>(b) * p.T + (q_ok ? q : 0)) * HD + h * D;
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {   // this thread's D / 2 columns, 16 at a time
      const int col = half * (D / 2) + c * 16;
      uint32_t r[16];
      tmem_ld_32x32b_x16(t_lane + C_S + col, r);
      tmem_ld_wait();
      if (q_ok) {
#pragma unroll
        for (int j = 0; j < 16; j += 4)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dq_row + col + j),
                       "f"(__uint_as_float(r[j])), "f"(__uint_as_float(r[j + 1])),
                       "f"(__uint_as_float(r[j + 2])), "f"(__uint_as_float(r[j + 3]))
                       : "memory");
      }
    }
    tcgen05_fence_before();
    if (it < 2) ADBG(8 + 8 * it);
  }
  ADBG(20);

  // dK / dV for this key tile: row = key
  const int key = k0 + tid;
  if (ok) {
    bf16* dk_row = p.dqkv + (static_cast<size_t>(b) * p.T + (key < p.T ? key : 0)) * row_stride + HD + h * D;
    bf16* dv_row = dk_row + HD;
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {   // this thread's D / 2 columns, 16 at a time
      const int col = half * (D / 2) + c * 16;
      uint32_t rk[16], rv[16];
      tmem_ld_32x32b_x16(t_lane + C_DK + col, rk);
      tmem_ld_32x32b_x16(t_lane + C_DV + col, rv);
      tmem_ld_wait();
      if (key < p.T) {
#pragma unroll
        for (int j = 0; j < 16; j += 8) {
          *reinterpret_cast<uint4*>(dk_row + col + j) = make_uint4(
              pack_bf16x2(__uint_as_float(rk[j]), __uint_as_float(rk[j + 1])),
              pack_bf16x2(__uint_as_float(rk[j + 2]), __uint_as_float(rk[j + 3])),
              pack_bf16x2(__uint_as_float(rk[j + 4]), __uint_as_float(rk[j + 5])),
              pack_bf16x2(__uint_as_float(rk[j + 6]), __uint_as_float(rk[j + 7])));
          *reinterpret_cast<uint4*>(dv_row + col + j) = make_uint4(
              pack_bf16x2(__uint_as_float(rv[j]), __uint_as_float(rv[j + 1])),
              pack_bf16x2(__uint_as_float(rv[j + 2]), __uint_as_float(rv[j + 3])),
              pack_bf16x2(__uint_as_float(rv[j + 4]), __uint_as_float(rv[j + 5])),
              pack_bf16x2(__uint_as_float(rv[j + 6]), __uint_as_float(rv[j + 7])));
        }
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) {
    tcgen05_fence_after();
    tmem_dealloc(tmem, TMEM_COLS);
  }
  ADBG(21);
  ADBG_NS(25);
}

// ------------------------------------------------------------------------------------------
// backward, warp-specialised and TMA-fed (long sequences)
// ------------------------------------------------------------------------------------------
// Same decomposition as attn_bwd_kernel - one CTA per (batch, head, 128-key tile) walking the query tiles at or
// below the diagonal - but computed TRANSPOSED so that the probabilities never leave tensor memory:
//   S^T  = K Q^T          dP^T = V dO^T                          (TMEM lanes = keys, columns = queries)
//   P^T  = exp2(S^T * scale*log2e - lse[q]),  dS^T = P^T (dP^T - delta[q]) * scale     -> bf16, back into TMEM
//   dV  += P^T dO         dK  += dS^T Q                          (A operand from tensor memory)
//   dQ   = dS K                                                  (dS^T also goes to smem once: MN-major A operand)
// 320 threads: warp 0 loader (TMA; K / V once, Q / dO double-buffered so the next query tile arrives under the
// current one), warp 1 MMA issuer, warps 2-9 compute (two threads per key row, half of the query columns each) -
// they also drain dQ into the fp32 accumulator with vector reductions.
// Tensor memory (512 columns): A [0,128) S^T -> P^T;  B [128,256) dP^T -> dS^T -> dQ (so the next S^T MMA can run
// while dQ is still being drained);  dV [256,256+D);  dK after it.
// Each half-row thread writes its packed bf16 results over the first half of the columns it read itself
// (queries [64h, 64h+64) -> columns 64h + [0,32)), so the two threads of a row never touch each other's inputs.
constexpr int ATTN_BWD_WS_THREADS = 320;

// WITH_DQ = false: dK / dV only (dQ comes from attn_bwd_dq_kernel): no dS^T copy in shared memory, no dQ MMA, no
// drain - the next tile's MMAs follow the dK MMA directly.
template <int D, bool WITH_DQ>
__global__ void __launch_bounds__(ATTN_BWD_WS_THREADS, 1)
attn_bwd_ws_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                   const __grid_constant__ CUtensorMap tmDQ, const AttnParams p) {
  constexpr int TILE_BYTES = TQ * D * 2;
  constexpr int SS_BYTES = TQ * TK * 2;
  constexpr uint32_t TMEM_COLS = 512;
  constexpr uint32_t C_A = 0, C_B = 128, C_DV = 256, C_DK = 256 + D;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);      // swizzled blocks need 1024-byte alignment (checked below)
  const uint32_t sK = base, sV = sK + TILE_BYTES, sQ = sV + TILE_BYTES, sdO = sQ + 2 * TILE_BYTES;
  const uint32_t sdS = sdO + 2 * TILE_BYTES, sStat = sdS + SS_BYTES;       // sStat: [2][lse2 128 | delta 128] floats
  const uint32_t bars = sStat + 2048;
  const uint32_t kv_full = bars, qdo_full = bars + 8, qdo_empty = bars + 24, sdp_full = bars + 40, ps_ready = bars + 48,
                 dq_full = bars + 56, a_free = bars + 64, stat_bar = bars + 72, drain_bar = bars + 80;
  uint8_t* gen = smem_raw;
  float* stat = reinterpret_cast<float*>(gen + 6 * TILE_BYTES + SS_BYTES);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(gen + 6 * TILE_BYTES + SS_BYTES + 2048 + 88);
  volatile int* abort_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);
  if ((base & 1023u) != 0u) {
    if (threadIdx.x == 0) atomicCAS(p.err_word, 0, ERR_ATTN_SMEM_ALIGN);
    return;
  }

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;     // early key tiles (most query tiles) first
  const int HD = p.H * D;
  const int k0 = kt * TK;
  const int nq_tiles = (p.T + TQ - 1) / TQ;
  const int n_it = nq_tiles - kt;

  if (threadIdx.x == 0) {
    *abort_flag = 0;
    mbar_init(kv_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(qdo_full + 8 * i, 1);
      mbar_init(qdo_empty + 8 * i, 1);
    }
    mbar_init(sdp_full, 1);
    mbar_init(ps_ready, 8);
    mbar_init(dq_full, 1);
    mbar_init(a_free, 8);
    mbar_init(stat_bar, 256);
    mbar_init(drain_bar, 8);
    fence_barrier_init();
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmDO);
    if (WITH_DQ) tma_prefetch_desc(&tmDQ);
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_ptr_smem), TMEM_COLS);
    tmem_relinquish();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *tmem_ptr_smem;

  if (warp == 0) {
    // ===================== loader =====================
    if (lane == 0) {
      mbar_arrive_expect_tx(kv_full, 2 * TILE_BYTES);
      tma_load_tile<D>(sK, &tmQKV, kv_full, HD + h * D, k0, b);
      tma_load_tile<D>(sV, &tmQKV, kv_full, 2 * HD + h * D, k0, b);
      auto load_qdo = [&](int it) {
        const int s = it & 1;
        mbar_arrive_expect_tx(qdo_full + 8 * s, 2 * TILE_BYTES);
        tma_load_tile<D>(sQ + s * TILE_BYTES, &tmQKV, qdo_full + 8 * s, h * D, (kt + it) * TQ, b);
        tma_load_tile<D>(sdO + s * TILE_BYTES, &tmDO, qdo_full + 8 * s, h * D, (kt + it) * TQ, b);
      };
      if constexpr (!WITH_DQ) {
        for (int it = 0; it < n_it; ++it) {
          if (it >= 2 && !mbar_wait(qdo_empty + 8 * (it & 1), ((it >> 1) - 1) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) break;
          load_qdo(it);
        }
      } else {
        // with dQ this thread also hands every drained dQ tile (staged by the compute warps in the Q / dO slot of
        // its iteration) to the TMA engine as reduce-adds, and refills the slot once the engine has read it - the
        // wait for that read would otherwise sit in a compute warp and hold up the next tile's math
        load_qdo(0);
        if (n_it > 1) load_qdo(1);
        for (int it = 0; it < n_it; ++it) {
          const int s = it & 1;
          if (!mbar_wait(drain_bar, it & 1, abort_flag, p.err_word, ERR_ATTN_WS)) break;
          if (p.tile_in_z != 2) {
#pragma unroll
            for (int bx = 0; bx < D / 32; ++bx) {
              const uint32_t src = ((bx * 32 >= D / 2) ? sdO : sQ) + s * TILE_BYTES + ((bx * 32) % (D / 2) >> 5) * SW_BLK;
              tma_reduce_add_3d(&tmDQ, src, h * D + bx * 32, (kt + it) * TQ, b);
            }
            tma_store_commit();
            tma_store_wait_read0();
          }
          if (it + 2 < n_it) load_qdo(it + 2);
        }
        tma_store_wait_all();     // the reduce-adds have landed before the kernel ends
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t idesc_st = umma_idesc_bf16(TK, TQ, 0, 0);     // S^T, dP^T: K-major x K-major, N = 128 queries
    const uint32_t idesc_tv = umma_idesc_bf16(TK, D, 0, 1);      // dV, dK: A from TMEM, B MN-major
    const uint32_t idesc_dq = umma_idesc_bf16(TQ, D, 1, 1);      // dQ: both operands MN-major
    bool ok = mbar_wait(kv_full, 0, abort_flag, p.err_word, ERR_ATTN_WS);
    [[maybe_unused]] long long w_qdo = 0, w_afree = 0, w_ps = 0, t_begin = WS_NOW(), _ws_t0 = 0;
    for (int it = 0; it < n_it && ok; ++it) {
      const int s = it & 1;
      const uint32_t q_s = sQ + s * TILE_BYTES, do_s = sdO + s * TILE_BYTES;
      WS_T0();
      if (!mbar_wait(qdo_full + 8 * s, (it >> 1) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
      WS_ACC(w_qdo);
      tcgen05_fence_after();
      if (lane == 0) {    // S^T goes to region A, which the previous iteration's dV MMA has finished reading (in order)
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16_ss(tmem + C_A, desc_sw_k(sK, kk), desc_sw_k(q_s, kk), idesc_st, kk > 0 ? 1u : 0u);
      }
      __syncwarp();
      // dP^T goes to region B, where the previous dQ sits until the compute warps have drained it
      WS_T0();
      if (WITH_DQ && it > 0 && !mbar_wait(a_free, (it - 1) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
      WS_ACC(w_afree);
      tcgen05_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16_ss(tmem + C_B, desc_sw_k(sV, kk), desc_sw_k(do_s, kk), idesc_st, kk > 0 ? 1u : 0u);
        umma_commit(sdp_full);
      }
      __syncwarp();
      WS_T0();
      if (!mbar_wait(ps_ready, it & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
      WS_ACC(w_ps);
      tcgen05_fence_after();
      if (lane == 0) {
        const uint32_t acc0 = it > 0 ? 1u : 0u;
#pragma unroll
        for (int kk = 0; kk < TQ / 16; ++kk) {   // reduction over the 128 queries; query 16 kk -> column 8 (kk % 4) + 64 (kk / 4)
          const uint32_t col = static_cast<uint32_t>((kk & 3) * 8 + (kk >> 2) * 64);
          umma_f16_ts(tmem + C_DV, tmem + C_A + col, desc_sw_mn(do_s, kk), idesc_tv, kk > 0 ? 1u : acc0);
        }
#pragma unroll
        for (int kk = 0; kk < TQ / 16; ++kk) {
          const uint32_t col = static_cast<uint32_t>((kk & 3) * 8 + (kk >> 2) * 64);
          umma_f16_ts(tmem + C_DK, tmem + C_B + col, desc_sw_mn(q_s, kk), idesc_tv, kk > 0 ? 1u : acc0);
        }
        if constexpr (WITH_DQ) {
#pragma unroll
          for (int kk = 0; kk < TK / 16; ++kk)     // dQ[q, d] = sum over keys dS[q, key] K[key, d]
            umma_f16_ss(tmem + C_B, desc_sw_mn(sdS, kk), desc_sw_mn(sK, kk), idesc_dq, kk > 0 ? 1u : 0u);
        }
        if constexpr (!WITH_DQ) umma_commit(qdo_empty + 8 * s);   // (with dQ: the drain stages dQ in this Q / dO slot first)
        if (WITH_DQ || it == n_it - 1) umma_commit(dq_full);
      }
      __syncwarp();
    }
    WS_DUMP(0, w_qdo); WS_DUMP(1, w_afree); WS_DUMP(2, w_ps); WS_DUMP(3, WS_NOW() - t_begin); WS_DUMP(4, n_it);
  } else {
    // ===================== compute (8 warps) =====================
    const int quarter = warp & 3, half = (warp - 2) >> 2;
    const int krow = quarter * 32 + lane;              // key row of the tile = TMEM lane
    const int key = k0 + krow;
    const int ctid = threadIdx.x - 64;                 // 0 .. 255
    const uint32_t t_lane = tmem + (static_cast<uint32_t>(quarter * 32) << 16);
    const float sl2 = p.scale * kLog2e;
    const bool key_ok = key < p.T && !(p.key_mask != nullptr && p.key_mask[static_cast<size_t>(b) * p.T + key] != 0);
    const size_t stat_base = (static_cast<size_t>(b) * p.H + h) * p.T;
    bool ok = true;
    [[maybe_unused]] long long w_stat = 0, w_sdp = 0, w_dq = 0, t_comp = 0, t_drain = 0, _ws_t0 = 0;
    // lse (log2 units, +inf for rows past T: their probabilities vanish) and delta of a query tile are staged one
    // iteration ahead (two buffers), off the critical path
    auto stage_stats = [&](int it_) {
      const int qq = (kt + it_) * TQ + (ctid & 127);
      float v;
      if (ctid < 128) v = qq < p.T ? p.lse[stat_base + qq] * kLog2e : INFINITY;
      else v = qq < p.T ? p.delta[stat_base + qq] : 0.f;
      stat[(it_ & 1) * 256 + ctid] = v;
      mbar_arrive(stat_bar);
    };
    stage_stats(0);
    for (int it = 0; it < n_it; ++it) {
      const int s = it & 1;
      const int q0 = (kt + it) * TQ;
      WS_T0();
      if (!mbar_wait(stat_bar, it & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
      WS_ACC(w_stat);
      WS_T0();
      if (!mbar_wait(sdp_full, it & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
      WS_ACC(w_sdp);
      WS_T0();
      tcgen05_fence_after();
      const bool diag = it == 0;
      // two passes of 32 query columns each; inside a pass everything is straight-line code over independent elements
      // (two warps per scheduler cannot hide a MUFU / LDS latency behind a branchy loop)
#pragma unroll
      for (int sub = 0; sub < 2; ++sub) {
        const int cbase = half * 64 + sub * 32;          // first query column of this pass
        uint32_t rs[32], rd[32];
        tmem_ld_32x32b_x32(t_lane + C_A + cbase, rs);
        tmem_ld_32x32b_x32(t_lane + C_B + cbase, rd);
        float lv[32], dv[32];
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          const float4 l4 = reinterpret_cast<const float4*>(stat + s * 256 + cbase)[g];
          const float4 d4 = reinterpret_cast<const float4*>(stat + s * 256 + 128 + cbase)[g];
          lv[4 * g] = l4.x; lv[4 * g + 1] = l4.y; lv[4 * g + 2] = l4.z; lv[4 * g + 3] = l4.w;
          dv[4 * g] = d4.x; dv[4 * g + 1] = d4.y; dv[4 * g + 2] = d4.z; dv[4 * g + 3] = d4.w;
        }
        // keys this row may attend from: all 32 queries, except above the diagonal (first iteration) / padded keys
        uint32_t kb = 0xffffffffu;
        if (diag) {
          const int first = key - (q0 + cbase);          // queries q0 + cbase + i with i >= first see this key
          kb = first <= 0 ? 0xffffffffu : (first >= 32 ? 0u : (0xffffffffu << first));
        }
        if (!key_ok) kb = 0u;
        tmem_ld_wait();
        float e[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) e[i] = fast_exp2(fmaf(__uint_as_float(rs[i]), sl2, -lv[i]));
        if (__any_sync(0xffffffffu, kb != 0xffffffffu)) {
#pragma unroll
          for (int i = 0; i < 32; ++i) e[i] = ((kb >> i) & 1u) ? e[i] : 0.f;
        }
        uint32_t pk[16], dk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const float d0 = e[i] * ((__uint_as_float(rd[i]) - dv[i]) * p.scale);
          const float d1 = e[i + 1] * ((__uint_as_float(rd[i + 1]) - dv[i + 1]) * p.scale);
          pk[i >> 1] = pack_bf16x2(e[i], e[i + 1]);
          dk[i >> 1] = pack_bf16x2(d0, d1);
        }
        tmem_st_32x32b_x16(t_lane + C_A + half * 64 + sub * 16, pk);
        tmem_st_32x32b_x16(t_lane + C_B + half * 64 + sub * 16, dk);
        // dS^T row of this key into smem: queries [cbase, cbase + 32) = block `half`, four swizzled 16-byte chunks
        if constexpr (WITH_DQ) {
#pragma unroll
          for (int g = 0; g < 4; ++g)
            st_shared_v4(sdS + half * SW_BLK + krow * 128 + (((sub * 4 + g) ^ (krow & 7)) << 4), dk[4 * g], dk[4 * g + 1],
                         dk[4 * g + 2], dk[4 * g + 3]);
        }
      }
      tmem_st_wait();
      fence_proxy_async_smem();
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(ps_ready);
      if (it + 1 < n_it) stage_stats(it + 1);
      WS_ACC(t_comp);
      if constexpr (!WITH_DQ) continue;
      // dQ of this (query tile, key tile) pair: lanes are query rows now
      WS_T0();
      if (!mbar_wait(dq_full, it & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
      WS_ACC(w_dq);
      WS_T0();
      tcgen05_fence_after();
      // dQ tile -> fp32 in the Q / dO slot of this iteration (its MMAs have retired), laid out as 128B-swizzled
      // [128 rows x 32 columns] boxes, then ONE thread hands it to the TMA engine as reduce-adds into the fp32
      // accumulator: full 128-byte lines instead of 32 row-divergent 16-byte reductions per warp instruction
      // (which kept the LSU busy for ~4400 cycles per tile pair), and asynchronous
      {
        const uint32_t stage = (half ? sdO : sQ) + s * TILE_BYTES;
#pragma unroll
        for (int c = 0; c < D / 32; ++c) {       // this thread's D / 2 columns, 16 at a time
          const int col = c * 16;                // within the half
          uint32_t r[16];
          tmem_ld_32x32b_x16(t_lane + C_B + half * (D / 2) + col, r);
          tmem_ld_wait();
          const uint32_t box = stage + (col >> 5) * SW_BLK + krow * 128;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int chunk = ((col & 31) >> 2) + j;          // 16-byte chunk (4 floats) within the 128-byte row
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(box + ((chunk ^ (krow & 7)) << 4)),
                         "r"(r[4 * j]), "r"(r[4 * j + 1]), "r"(r[4 * j + 2]), "r"(r[4 * j + 3]) : "memory");
          }
        }
        tcgen05_fence_before();
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(a_free);                   // region B has been read: the next dP^T MMA may overwrite it
          mbar_arrive(drain_bar);
        }
      }
      WS_ACC(t_drain);
    }
    if (warp == 2) { WS_DUMP(8, w_stat); WS_DUMP(9, w_sdp); WS_DUMP(10, w_dq); WS_DUMP(11, t_comp); WS_DUMP(12, t_drain); }
    if (!WITH_DQ && ok) ok = mbar_wait(dq_full, 0, abort_flag, p.err_word, ERR_ATTN_WS);   // the one commit after the last MMAs
    if (ok) {   // dK / dV of this key tile (complete since the last dq_full)
      tcgen05_fence_after();
      bf16* dk_row = p.dqkv + (static_cast<size_t>(b) * p.T + (key < p.T ? key : 0)) * (3 * static_cast<size_t>(HD)) + HD + h * D;
      bf16* dv_row = dk_row + HD;
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        const int col = half * (D / 2) + c * 16;
        uint32_t rk[16], rv[16];
        tmem_ld_32x32b_x16(t_lane + C_DK + col, rk);
        tmem_ld_32x32b_x16(t_lane + C_DV + col, rv);
        tmem_ld_wait();
        if (key < p.T) {
#pragma unroll
          for (int j = 0; j < 16; j += 8) {
            *reinterpret_cast<uint4*>(dk_row + col + j) = make_uint4(
                pack_bf16x2(__uint_as_float(rk[j]), __uint_as_float(rk[j + 1])),
                pack_bf16x2(__uint_as_float(rk[j + 2]), __uint_as_float(rk[j + 3])),
                pack_bf16x2(__uint_as_float(rk[j + 4]), __uint_as_float(rk[j + 5])),
                pack_bf16x2(__uint_as_float(rk[j + 6]), __uint_as_float(rk[j + 7])));
            *reinterpret_cast<uint4*>(dv_row + col + j) = make_uint4(
                pack_bf16x2(__uint_as_float(rv[j]), __uint_as_float(rv[j + 1])),
                pack_bf16x2(__uint_as_float(rv[j + 2]), __uint_as_float(rv[j + 3])),
                pack_bf16x2(__uint_as_float(rv[j + 4]), __uint_as_float(rv[j + 5])),
                pack_bf16x2(__uint_as_float(rv[j + 6]), __uint_as_float(rv[j + 7])));
          }
        }
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------
// backward, dQ by the query tile's own CTA (no atomics): one CTA per (batch, head, 128-query tile) walks the key
// tiles up to the diagonal like the forward does -  S = Q K^T, dP = dO V^T, dS = P (dP - delta) * scale written back
// as bf16 over the S columns, dQ += dS K with the A operand from tensor memory - and writes its dQ rows once, as
// bf16, straight into dqkv.  Costs two extra MMAs and a second exp per tile pair compared with reducing dQ out of
// the dK / dV kernel, but removes 64 KB of fp32 read-modify-write per pair (a third of that kernel's time), the
// accumulator memset and the fp32 -> bf16 conversion pass.
// 192 threads: loader (Q, dO once; K, V through a 2-deep ring), MMA issuer, four compute warps (one query row per
// thread: lse and delta are per-thread scalars here).  TMEM: S / dS [0,128), dP [128,256), dQ [256, 256 + D).
template <int D>
__global__ void __launch_bounds__(ATTN_WS_THREADS, 1)
attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                   const AttnParams p) {
  constexpr int TILE_BYTES = TQ * D * 2;
  constexpr uint32_t TMEM_COLS = 512;
  constexpr uint32_t C_S = 0, C_DP = 128, C_DQ = 256;
  constexpr int NST = 2;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  const uint32_t sQ = base, sdO = sQ + TILE_BYTES, sK = sdO + TILE_BYTES, sV = sK + NST * TILE_BYTES;
  const uint32_t bars = sV + NST * TILE_BYTES;
  const uint32_t q_full = bars, kv_full = bars + 8, kv_empty = bars + 24, sdp_full = bars + 40, ds_ready = bars + 48,
                 dq_done = bars + 56;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem_raw + (2 + 2 * NST) * TILE_BYTES + 64);
  volatile int* abort_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);
  if ((base & 1023u) != 0u) {
    if (threadIdx.x == 0) atomicCAS(p.err_word, 0, ERR_ATTN_SMEM_ALIGN);
    return;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = gridDim.z - 1 - blockIdx.z, h = blockIdx.x, b = blockIdx.y;   // long query tiles first
  const int HD = p.H * D;
  const int q0 = qt * TQ;
  const int n_kt = qt + 1;

  if (threadIdx.x == 0) {
    *abort_flag = 0;
    mbar_init(q_full, 1);
    for (int i = 0; i < NST; ++i) {
      mbar_init(kv_full + 8 * i, 1);
      mbar_init(kv_empty + 8 * i, 1);
    }
    mbar_init(sdp_full, 1);
    mbar_init(ds_ready, 4);
    mbar_init(dq_done, 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmDO);
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_ptr_smem), TMEM_COLS);
    tmem_relinquish();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *tmem_ptr_smem;

  if (warp == 0) {
    // ===================== loader =====================
    if (lane == 0) {
      mbar_arrive_expect_tx(q_full, 2 * TILE_BYTES);
      tma_load_tile<D>(sQ, &tmQKV, q_full, h * D, q0, b);
      tma_load_tile<D>(sdO, &tmDO, q_full, h * D, q0, b);
      for (int j = 0; j < n_kt; ++j) {
        const int st = j % NST;
        if (j >= NST && !mbar_wait(kv_empty + 8 * st, (j / NST - 1) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) break;
        mbar_arrive_expect_tx(kv_full + 8 * st, 2 * TILE_BYTES);
        tma_load_tile<D>(sK + st * TILE_BYTES, &tmQKV, kv_full + 8 * st, HD + h * D, j * TK, b);
        tma_load_tile<D>(sV + st * TILE_BYTES, &tmQKV, kv_full + 8 * st, 2 * HD + h * D, j * TK, b);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t idesc_s = umma_idesc_bf16(TQ, TK, 0, 0);      // S, dP: K-major x K-major, N = 128 keys
    const uint32_t idesc_dq = umma_idesc_bf16(TQ, D, 0, 1);      // dQ: A from TMEM, B = K tile MN-major
    bool ok = mbar_wait(q_full, 0, abort_flag, p.err_word, ERR_ATTN_WS);
    for (int j = 0; j < n_kt && ok; ++j) {
      const int st = j % NST;
      if (!mbar_wait(kv_full + 8 * st, (j / NST) & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
      tcgen05_fence_after();
      if (lane == 0) {     // (S / dP of tile j overwrite dS of tile j - 1: its dQ MMA was issued before, in order)
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16_ss(tmem + C_S, desc_sw_k(sQ, kk), desc_sw_k(sK + st * TILE_BYTES, kk), idesc_s, kk > 0 ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_f16_ss(tmem + C_DP, desc_sw_k(sdO, kk), desc_sw_k(sV + st * TILE_BYTES, kk), idesc_s, kk > 0 ? 1u : 0u);
        umma_commit(sdp_full);
      }
      __syncwarp();
      if (!mbar_wait(ds_ready, j & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
      tcgen05_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int kk = 0; kk < TK / 16; ++kk)   // dQ (+)= dS K, reduction over the 128 keys
          umma_f16_ts(tmem + C_DQ, tmem + C_S + kk * 8, desc_sw_mn(sK + st * TILE_BYTES, kk), idesc_dq,
                      (j > 0 || kk > 0) ? 1u : 0u);
        umma_commit(kv_empty + 8 * st);
        if (j == n_kt - 1) umma_commit(dq_done);
      }
      __syncwarp();
    }
  } else {
    // ===================== compute (4 warps, one query row per thread) =====================
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int q = q0 + row;
    const bool q_ok = q < p.T;
    const uint32_t t_lane = tmem + (static_cast<uint32_t>(quarter * 32) << 16);
    const float sl2 = p.scale * kLog2e;
    const uint8_t* mask_row = p.key_mask != nullptr ? p.key_mask + static_cast<size_t>(b) * p.T : nullptr;
    const size_t stat_idx = (static_cast<size_t>(b) * p.H + h) * p.T + (q_ok ? q : 0);
    const float lse2 = q_ok ? p.lse[stat_idx] * kLog2e : INFINITY;      // +inf: rows past T contribute nothing
    const float delta = q_ok ? p.delta[stat_idx] : 0.f;
    bool ok = true;
    for (int j = 0; j < n_kt; ++j) {
      if (!mbar_wait(sdp_full, j & 1, abort_flag, p.err_word, ERR_ATTN_WS)) { ok = false; break; }
      tcgen05_fence_after();
      const bool diag = (j == qt);
      const bool masked_tile = diag || mask_row != nullptr;
#pragma unroll
      for (int c = 0; c < 4; ++c) {      // 32 keys per pass, straight-line code inside
        uint32_t rs[32], rd[32];
        tmem_ld_32x32b_x32(t_lane + C_S + c * 32, rs);
        tmem_ld_32x32b_x32(t_lane + C_DP + c * 32, rd);
        uint32_t kb = 0xffffffffu;
        if (masked_tile) kb = keep_bits(j * TK + c * 32, q, diag, q_ok, p.T, mask_row);
        tmem_ld_wait();
        float e[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) e[i] = fast_exp2(fmaf(__uint_as_float(rs[i]), sl2, -lse2));
        if (masked_tile) {
#pragma unroll
          for (int i = 0; i < 32; ++i) e[i] = ((kb >> i) & 1u) ? e[i] : 0.f;
        }
        uint32_t dk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2)
          dk[i >> 1] = pack_bf16x2(e[i] * ((__uint_as_float(rd[i]) - delta) * p.scale),
                                   e[i + 1] * ((__uint_as_float(rd[i + 1]) - delta) * p.scale));
        // keys [32c, 32c + 32) -> columns [16c, 16c + 16) of the S region: already read by this thread
        tmem_st_32x32b_x16(t_lane + C_S + c * 16, dk);
      }
      tmem_st_wait();
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(ds_ready);
    }
    if (ok && mbar_wait(dq_done, 0, abort_flag, p.err_word, ERR_ATTN_WS)) {
      tcgen05_fence_after();
      bf16* dq_row = p.dqkv + (static_cast<size_t>(b) * p.T + (q_ok ? q : 0)) * (3 * static_cast<size_t>(HD)) + h * D;
#pragma unroll
      for (int c = 0; c < D / 16; ++c) {
        uint32_t o[16];
        tmem_ld_32x32b_x16(t_lane + C_DQ + c * 16, o);
        tmem_ld_wait();
        if (q_ok) {
#pragma unroll
          for (int i = 0; i < 16; i += 8)
            *reinterpret_cast<uint4*>(dq_row + c * 16 + i) = make_uint4(
                pack_bf16x2(__uint_as_float(o[i]), __uint_as_float(o[i + 1])),
                pack_bf16x2(__uint_as_float(o[i + 2]), __uint_as_float(o[i + 3])),
                pack_bf16x2(__uint_as_float(o[i + 4]), __uint_as_float(o[i + 5])),
                pack_bf16x2(__uint_as_float(o[i + 6]), __uint_as_float(o[i + 7])));
        }
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem, TMEM_COLS);
  }
}

// dq fp32 scratch [B*T, H*D] -> bf16 q section of dqkv [B*T, 3*H*D]
__global__ void __launch_bounds__(256)
attn_dq_convert_kernel(const float* __restrict__ dq_acc, bf16* __restrict__ dqkv, int64_t rows, int HD) {
  const int vec_per_row = HD >> 2;
  const int64_t total = rows * vec_per_row;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / vec_per_row;
    const int c = static_cast<int>(i % vec_per_row);
    const float4 f = __ldg(reinterpret_cast<const float4*>(dq_acc + r * HD) + c);
    reinterpret_cast<uint2*>(dqkv + r * 3 * HD)[c] = make_uint2(pack_bf16x2(f.x, f.y), pack_bf16x2(f.z, f.w));
  }
}

template <int D>
int fwd_launch(const AttnParams& p, cudaStream_t st) {
  constexpr int smem = 3 * TQ * D * 2 + 8 + 16 + 128;
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!configured[dev & 63]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured[dev & 63] = true;
  }
  dim3 grid(p.H, p.B, ceil_div(p.T, TQ));
  attn_fwd_kernel<D><<<grid, ATTN_THREADS, smem, st>>>(p);
  B200_LAUNCH_CHECK();
  return 0;
}

// 3-D view (columns, T, B) of qkv [B, T, 3*H*D] / dO [B, T, H*D] for the TMA loads of the warp-specialised kernels:
// a box {64 columns, 128 rows, 1} with the 128-byte swizzle is one block of a tile (rows past T are zero-filled).
int make_qkv_tmap(CUtensorMap* out, const bf16* base, int B, int T, int cols, int D) {
  (void)D;
  const uint64_t dims[3] = {static_cast<uint64_t>(cols), static_cast<uint64_t>(T), static_cast<uint64_t>(B)};
  const uint64_t strides[2] = {static_cast<uint64_t>(cols) * 2, static_cast<uint64_t>(T) * cols * 2};
  const uint32_t box[3] = {64, TQ, 1};
  return b200_make_tmap_nd(out, base, 3, dims, strides, box, 2, 1);
}

// B200GPT_ATTN_WS=0 / 1 forces the per-tile (v1) / warp-specialised forward; unset: by sequence length
int g_attn_ws = []() { const char* e = getenv("B200GPT_ATTN_WS"); return e == nullptr ? -1 : (e[0] == '0' ? 0 : 1); }();

// B200GPT_ATTN_DQ=atomic: dQ reduced out of the dK / dV kernel with fp32 vector reductions (+ memset + conversion
// pass) instead of the query-owned dQ kernel
bool g_attn_dq_kernel = []() { const char* e = getenv("B200GPT_ATTN_DQ"); return e != nullptr && e[0] == 'k'; }();
int g_attn_bwd_ws = []() { const char* e = getenv("B200GPT_ATTN_BWD_WS"); return e == nullptr ? -1 : (e[0] == '0' ? 0 : 1); }();

template <int D>
int fwd_ws_launch(const AttnParams& p, cudaStream_t st) {
  constexpr int smem = 7 * TQ * D * 2 + 144 + 16;
  static_assert(smem <= 227 * 1024, "attention forward (warp-specialised): shared memory budget exceeded");
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!configured[dev & 63]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_ws_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured[dev & 63] = true;
  }
  CUtensorMap tm;
  int rc = make_qkv_tmap(&tm, p.qkv, p.B, p.T, 3 * p.H * D, D);
  if (rc != 0) return rc;
  dim3 grid(p.H, p.B, ceil_div(p.T, TQ));
  attn_fwd_ws_kernel<D><<<grid, ATTN_WS_THREADS, smem, st>>>(tm, p);
  B200_LAUNCH_CHECK();
  return 0;
}

template <int D>
int bwd_launch(const AttnParams& p, cudaStream_t st, int mode) {
  constexpr int smem = 4 * TQ * D * 2 + 2 * TQ * TK * 2 + 8 + 16 + 128;
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!configured[dev & 63]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured[dev & 63] = true;
  }
  const int64_t rows = static_cast<int64_t>(p.B) * p.T;
  const int HD = p.H * D;
  const int nkt = ceil_div(p.T, TK);
  // mode: -1 automatic, 0 per-tile kernel, 1 warp-specialised + dQ by fp32 reductions, 2 warp-specialised dK / dV + dQ kernel
  const bool ws = D >= 64 && (mode >= 0 ? mode >= 1 : (g_attn_bwd_ws >= 0 ? g_attn_bwd_ws == 1 : nkt >= 4));
  const bool dq_kernel = mode >= 0 ? mode == 2 : g_attn_dq_kernel;
  bool launched = false;
  if constexpr (D >= 64) if (ws && dq_kernel) {
    // dK / dV by key-tile CTAs, dQ by query-tile CTAs: no atomics, no accumulator memset, no conversion pass
    constexpr int smem_kv = 6 * TQ * D * 2 + TQ * TK * 2 + 2048 + 88 + 16;
    constexpr int smem_dq = 6 * TQ * D * 2 + 64 + 16;
    static bool configured_kv[64] = {false};
    if (!configured_kv[dev & 63]) {
      B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_ws_kernel<D, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_kv));
      B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_dq_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dq));
      configured_kv[dev & 63] = true;
    }
    {
      const int64_t total = rows * p.H;
      int grid = static_cast<int>((total + 255) / 256);
      if (grid > b200_num_sms() * 16) grid = b200_num_sms() * 16;
      attn_delta_kernel<D><<<grid, 256, 0, st>>>(p.out, p.dout, p.delta, p.B, p.T, p.H);
      B200_LAUNCH_CHECK();
    }
    CUtensorMap tq, tdo;
    int rc = make_qkv_tmap(&tq, p.qkv, p.B, p.T, 3 * HD, D);
    if (rc != 0) return rc;
    rc = make_qkv_tmap(&tdo, p.dout, p.B, p.T, HD, D);
    if (rc != 0) return rc;
    const int nkt2 = ceil_div(p.T, TK);
    attn_bwd_ws_kernel<D, false><<<dim3(nkt2, p.H, p.B), ATTN_BWD_WS_THREADS, smem_kv, st>>>(tq, tdo, tq, p);
    B200_LAUNCH_CHECK();
    attn_bwd_dq_kernel<D><<<dim3(p.H, p.B, ceil_div(p.T, TQ)), ATTN_WS_THREADS, smem_dq, st>>>(tq, tdo, p);
    B200_LAUNCH_CHECK();
    return 0;
  }
  {   // delta = rowsum(dO * O), and the dQ accumulator cleared in the same pass
    const int64_t total = rows * p.H;
    int grid = static_cast<int>((total + 255) / 256);
    if (grid > b200_num_sms() * 16) grid = b200_num_sms() * 16;
    attn_delta_kernel<D><<<grid, 256, 0, st>>>(p.out, p.dout, p.delta, p.B, p.T, p.H, p.dq_acc);
    B200_LAUNCH_CHECK();
  }
  if constexpr (D >= 64) if (ws) {
    launched = true;
    constexpr int smem_ws = 6 * TQ * D * 2 + TQ * TK * 2 + 2048 + 88 + 16;
    static_assert(smem_ws <= 227 * 1024, "attention backward (warp-specialised): shared memory budget exceeded");
    static bool configured_ws[64] = {false};
    if (!configured_ws[dev & 63]) {
      B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_ws_kernel<D, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_ws));
      configured_ws[dev & 63] = true;
    }
    CUtensorMap tq, tdo;
    int rc = make_qkv_tmap(&tq, p.qkv, p.B, p.T, 3 * HD, D);
    if (rc != 0) return rc;
    rc = make_qkv_tmap(&tdo, p.dout, p.B, p.T, HD, D);
    if (rc != 0) return rc;
    AttnParams pw = p;
    pw.tile_in_z = getenv("B200GPT_ATTN_NO_DQ") != nullptr ? 2 : 0;
    CUtensorMap tdq;   // fp32 dQ accumulator [B, T, H*D] as (columns, T, B): box {32 columns = 128 bytes, 128 rows}
    {
      const uint64_t dims[3] = {static_cast<uint64_t>(HD), static_cast<uint64_t>(p.T), static_cast<uint64_t>(p.B)};
      const uint64_t strides[2] = {static_cast<uint64_t>(HD) * 4, static_cast<uint64_t>(p.T) * HD * 4};
      const uint32_t box[3] = {32, TQ, 1};
      rc = b200_make_tmap_nd(&tdq, p.dq_acc, 3, dims, strides, box, 4, 1);
      if (rc != 0) return rc;
    }
    attn_bwd_ws_kernel<D, true><<<dim3(nkt, p.H, p.B), ATTN_BWD_WS_THREADS, smem_ws, st>>>(tq, tdo, tdq, pw);
    B200_LAUNCH_CHECK();
  }
  if (!launched) {
    AttnParams pb = p;
    pb.tile_in_z = nkt <= 2 ? 1 : 0;
    dim3 grid = pb.tile_in_z ? dim3(p.H, p.B, nkt) : dim3(nkt, p.H, p.B);
    attn_bwd_kernel<D><<<grid, ATTN_BWD_THREADS, smem, st>>>(pb);
    B200_LAUNCH_CHECK();
  }
  {
    const int64_t total = rows * (HD / 4);
    int grid2 = static_cast<int>((total + 255) / 256);
    if (grid2 > b200_num_sms() * 16) grid2 = b200_num_sms() * 16;
    attn_dq_convert_kernel<<<grid2, 256, 0, st>>>(p.dq_acc, p.dqkv, rows, HD);
    B200_LAUNCH_CHECK();
  }
  return 0;
}

}  // namespace

extern "C" {

int b200_attn_set_debug_buffer(void* buf) {
  g_attn_dbg = reinterpret_cast<long long*>(buf);
  return 0;
}

static int attn_fwd_impl(const void* qkv, void* out, float* lse, int B, int T, int H, int D, float scale,
                         const void* key_mask, int force_ws, void* stream) {
  B200_REQUIRE(D == 32 || D == 64 || D == 128, "attention: head_dim %d not supported (32, 64, 128)", D);
  B200_REQUIRE(B > 0 && T > 0 && H > 0 && B <= 65535 && H <= 65535, "attention: bad shape");
  AttnParams p{};
  p.qkv = reinterpret_cast<const bf16*>(qkv);
  p.out = reinterpret_cast<bf16*>(out);
  p.lse = lse;
  p.key_mask = reinterpret_cast<const uint8_t*>(key_mask);
  p.B = B; p.T = T; p.H = H; p.scale = scale;
  p.err_word = b200_error_word();
  p.dbg = g_attn_dbg;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool ws = force_ws >= 0 ? force_ws == 1 : (g_attn_ws >= 0 ? g_attn_ws == 1 : (T > 3 * TQ && D >= 64));
  if (ws && D >= 64) {
    if (D == 64) return fwd_ws_launch<64>(p, st);
    return fwd_ws_launch<128>(p, st);
  }
  if (D == 32) return fwd_launch<32>(p, st);
  if (D == 64) return fwd_launch<64>(p, st);
  return fwd_launch<128>(p, st);
}

int b200_attn_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int D, float scale,
                  const void* key_mask, void* stream) {
  return attn_fwd_impl(qkv, out, lse, B, T, H, D, scale, key_mask, -1, stream);
}

// same with the kernel chosen by the caller: ws = 0 per-tile kernel, 1 warp-specialised kernel, -1 automatic
int b200_attn_fwd_select(const void* qkv, void* out, float* lse, int B, int T, int H, int D, float scale,
                         const void* key_mask, int ws, void* stream) {
  return attn_fwd_impl(qkv, out, lse, B, T, H, D, scale, key_mask, ws, stream);
}

static int attn_bwd_impl(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                         float* delta, float* dq_acc, int B, int T, int H, int D, float scale,
                         const void* key_mask, int mode, void* stream) {
  B200_REQUIRE(D == 32 || D == 64 || D == 128, "attention: head_dim %d not supported (32, 64, 128)", D);
  B200_REQUIRE(B > 0 && T > 0 && H > 0 && B <= 65535 && H <= 65535, "attention: bad shape");
  AttnParams p{};
  p.qkv = reinterpret_cast<const bf16*>(qkv);
  p.out = const_cast<bf16*>(reinterpret_cast<const bf16*>(out));
  p.dout = reinterpret_cast<const bf16*>(dout);
  p.lse = const_cast<float*>(lse);
  p.dqkv = reinterpret_cast<bf16*>(dqkv);
  p.delta = delta;
  p.dq_acc = dq_acc;
  p.key_mask = reinterpret_cast<const uint8_t*>(key_mask);
  p.B = B; p.T = T; p.H = H; p.scale = scale;
  p.err_word = b200_error_word();
  p.dbg = g_attn_dbg;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (D == 32) return bwd_launch<32>(p, st, mode);
  if (D == 64) return bwd_launch<64>(p, st, mode);
  return bwd_launch<128>(p, st, mode);
}

int b200_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                  float* delta, float* dq_acc, int B, int T, int H, int D, float scale,
                  const void* key_mask, void* stream) {
  return attn_bwd_impl(qkv, out, dout, lse, dqkv, delta, dq_acc, B, T, H, D, scale, key_mask, -1, stream);
}

// same with the kernels chosen by the caller: mode 0 per-tile kernel, 1 warp-specialised with dQ reduced by fp32 vector
// reductions, 2 warp-specialised dK / dV kernel + query-owned dQ kernel (no atomics; dq_acc unused), -1 automatic
int b200_attn_bwd_select(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                         float* delta, float* dq_acc, int B, int T, int H, int D, float scale,
                         const void* key_mask, int mode, void* stream) {
  return attn_bwd_impl(qkv, out, dout, lse, dqkv, delta, dq_acc, B, T, H, D, scale, key_mask, mode, stream);
}

}  // extern "C"
