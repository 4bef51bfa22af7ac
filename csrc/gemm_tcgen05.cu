// Persistent, warp-specialised bf16 GEMM for sm_100a: TMA -> 128B-swizzled smem ring ->
// tcgen05.mma (fp32 accumulators double-buffered in TMEM) -> tcgen05.ld epilogue with fused bias /
// ReLU / ReLU-mask / fp32 residual / bias-gradient column sums / cast / split-K reduction.  One
// kernel serves forward (K-major x K-major), dgrad (K-major x MN-major) and wgrad (MN-major x
// MN-major): operand major-ness only changes the TMA box shape and the UMMA descriptors, so no
// tensor is ever transposed in memory.
//
//   C[M,N] (+)= A[M,K] . B[N,K]^T        K = reduction extent
//   K-major  operand: memory is [rows][K]  (row stride ld)   -> one TMA box {64 k, rows}
//   MN-major operand: memory is [K][rows]  (row stride ld)   -> rows/64 TMA boxes {64 rows, 64 k}
//
// Two tile engines (template parameter CG):
//   CG = 2  CTA pair (cluster of 2, the two SMs of a TPC) per 256 x 256 tile, tcgen05 cta_group::2:
//           each CTA stages its 128 rows of A and half of B, the leader issues M = 256 MMAs and
//           multicast commits.  320 threads: warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer
//           (leader only), warps 2..9 epilogue (two per TMEM lane quarter, half of the columns each).
//   CG = 1  one CTA per 128 x 256 (or 128 x 128, 2 CTAs/SM) tile, cta_group::1.  224 threads: warp 0
//           TMA producer, warp 1 MMA issuer, warps 2..5 epilogue, warp 6 gather warp of the fused
//           FSDP all-gather (b200_ag_gemm_bf16).
// All waits are bounded (common.cuh: mbar_wait) so a protocol bug surfaces as an error code, never
// as a hung GPU.
//
// Serves reference call sites models/gpt.py:60-64,71-73,102 (projections), :29-38 (MLP),
// :219,231 (LM head) and every backward GEMM autograd derives from them (SURVEY K3,K5-K9).
#include <stdlib.h>

#include <mutex>
#include <unordered_map>

#include "../include/b200gpt.h"
#include "common.cuh"

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int GEMM_THREADS = 224;  // CG = 1: TMA producer, MMA issuer, 4 epilogue warps, gather warp (fused all-gather)
// CTA pairs run 8 epilogue warps (two per TMEM lane quarter, each taking half of the tile's columns): with
// K = 256 a tile's MMAs last ~2 K cycles, less than 4 warps need for a bias/mask/residual epilogue
template <int CG> constexpr int epilogue_warps() { return CG == 2 ? 8 : 4; }
// AG: the kernel carries the gather warp of the fused FSDP all-gather.  Single CTAs always do (warp 6); CTA pairs
// only in their AG instantiation (warp 10), which pays for the bounce buffers with one ring stage.
template <int CG, bool AG = false> constexpr int gemm_threads() {
  return CG == 2 ? 32 * (2 + epilogue_warps<2>() + (AG ? 1 : 0)) : GEMM_THREADS;
}
constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KB

struct GemmParams {
  int M, N, K;
  int ldc, ldr;
  int num_m_tiles, num_n_tiles;
  int k_splits, kb_total, kb_per_split;
  uint32_t idesc;
  uint64_t a_desc_base, b_desc_base;
  uint32_t a_kstep, b_kstep;  // smem byte advance per UMMA_K
  int a_mn, b_mn;
  const float* bias;
  const float* residual;
  void* C;
  bf16* aux_out;
  const bf16* mask_src;
  float* colsum;  // EPI_CSUM: colsum[n] += sum over rows of the (bf16-rounded) output - the bias gradient of the layer that produced mask_src
  int act, out_fp32, accumulate;
  int tma_store;  // 1: epilogue stages tiles in smem and writes them with TMA (coalesced, clipped)
  int n_fastest;  // tile order: 0 = consecutive CTAs walk down M (share the B tile), 1 = walk along N (share A)
  long long* dbg;  // optional per-CTA phase timestamps (16 slots per CTA), bring-up only
  int* err_word;
  // Fused all-gather of the B operand (FSDP, b200_ag_gemm_bf16): before its B tiles are read, the
  // kernel itself pulls bytes [ag_off, ag_off + ag_bytes) of a flat parameter group, sharded evenly
  // over the ranks, out of the owning peers' shard buffers (NVLink peer loads) into the local
  // full buffer that the B tensor map points at.  ag_bytes == 0 disables it.
  const uint8_t* const* ag_peers;  // device array [world]: base of each rank's shard of the group
  uint8_t* ag_full;                // local full (un-sharded) buffer of the group
  long long ag_off, ag_bytes, ag_shard_bytes;
  int* ag_sync;                    // [1] exited-CTA tickets, [2 + j] slices landed for n-tile j (all zero between launches)
  int ag_tile_bytes;               // bytes of W per n-tile (BN * K * 2), a multiple of AG_SLICE
  // EPI_RSCALE: row_a[m] multiplies output row m.  EPI_SMAX (LM head, see b200_lmhead_gemm_bf16): row_a[m] is
  // the l2 norm of A's row m, wmax[0] the largest l2 norm of a B row, wmax[1] the gap estimate; the epilogue
  // stores exp(c[m,n] - lm_row_shift(...)) and leaves (sum, max) partials per row and column span in part.
  const float* row_a;
  const float* wmax;
  float2* part;                    // [num_parts][M]
  const int64_t* targets;          // [M] column whose raw value is copied to tgt_val[m] (out of range: none)
  float* tgt_val;
};

// Grouped launch (small weight-gradient GEMMs of one transformer block): up to MAX_GROUP independent problems with
// the same operand majors, epilogue and K split share one persistent grid; a tile index selects its problem.
constexpr int MAX_GROUP = 4;
struct GroupDesc {
  int tile_end;                    // exclusive prefix sum of tiles (m tiles x n tiles x k splits)
  int M, N;
  int num_m_tiles, num_n_tiles;
  int k_splits, kb_total, kb_per_split;
};
struct alignas(64) GroupParams {
  CUtensorMap a[MAX_GROUP], b[MAX_GROUP], c[MAX_GROUP];
  GroupDesc g[MAX_GROUP];
  int n, total_tiles;
};
struct NoGroup { int unused; };
template <bool GRP> using GroupArg = std::conditional_t<GRP, GroupParams, NoGroup>;
struct TileMap {
  int m0, n0, kb0, kb1, M, N;      // m0 / n0: first row / column of the (CTA pair's) tile
  const CUtensorMap *a, *b, *c;
};

enum : int { ERR_GEMM_PRODUCER = 101, ERR_GEMM_MMA = 102, ERR_GEMM_EPILOGUE = 103, ERR_GEMM_SMEM_ALIGN = 104,
             ERR_GEMM_AG_WAIT = 105 };


// one 32-row x 128-byte swizzled staging tile per epilogue warp (x NSTG buffers)
template <int CG> constexpr int staging_bytes() { return epilogue_warps<CG>() * 4096; }
// staging buffers per epilogue warp: with two, a warp fills one while the TMA engine still reads the other
// (CTA pairs hide that wait behind the second warp of the lane quarter; a pair kernel with 5 operand stages and
// no gather buffers has room for a second 32 KB set, which the write-bound LM-head GEMM uses: its exp2 work
// then runs while the stores of the previous tile drain)
template <int BN, int CG, int STAGES = 6, bool AG = false> constexpr int staging_bufs() {
  return (CG == 1 && BN == 128) ? 2 : (CG == 2 && STAGES <= 5 && !AG) ? 2 : 1;
}
// Gather warp of the fused all-gather: slices of ag_slice() bytes travel peer shard -> smem bounce buffer -> local
// full buffer with bulk copies.  Single CTAs: one 16 KB buffer.  CTA pairs (AG instantiation): four 8 KB buffers,
// three loads in flight while the fourth buffer drains.
template <int CG> constexpr int ag_slice() { return CG == 2 ? 8192 : 16384; }
template <int CG> constexpr int ag_nbuf() { return CG == 2 ? 4 : 1; }
template <int CG, bool AG> constexpr int ag_bounce_bytes() { return CG == 2 ? (AG ? ag_slice<2>() * ag_nbuf<2>() : 0) : ag_slice<1>(); }
template <int N> __device__ __forceinline__ void bulk_wait_group() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// Wait accounting (tools/gemm_phases.py) costs ~1 % of a training step even when the debug buffer is
// null, so it is only compiled in with -DB200_GEMM_TRACE (B200_GEMM_TRACE=1 python -m csrc.build).
#ifdef B200_GEMM_TRACE
#define TRACE_ON(p) ((p).dbg != nullptr)
#else
#define TRACE_ON(p) false
#endif

__device__ __forceinline__ long long global_ns() {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define DBG_STAMP_NS(slot)                                                                \
  do {                                                                                    \
    if (p.dbg != nullptr) p.dbg[blockIdx.x * 32 + (slot)] = global_ns();                  \
  } while (0)
#define DBG_STAMP(slot)                                                                   \
  do {                                                                                    \
    if (p.dbg != nullptr) p.dbg[blockIdx.x * 32 + (slot)] = clock64();                    \
  } while (0)

enum : int { EPI_BIAS = 1, EPI_RELU = 2, EPI_AUX = 4, EPI_MASK = 8, EPI_RES = 16, EPI_F32 = 32, EPI_ACC = 64,
             EPI_CSUM = 128, EPI_SMAX = 256, EPI_NOSTORE = 512, EPI_RSCALE = 1024, EPI_GELU = 2048 };

// CG = 1: one CTA per 128 x BN tile (tcgen05 cta_group::1).
// CG = 2: a CTA pair (cluster of 2 = the two SMs of a TPC) per 256 x BN tile (cta_group::2): each CTA
//         stages its own 128 rows of A and HALF of the B tile, the leader's single thread issues
//         M = 256 MMAs that read both CTAs' shared memory and write both CTAs' TMEM.  Per FLOP a CTA
//         pulls 1/3 less operand data out of L2, which is what bounds the short-K GEMMs here.
template <int BN, int STAGES, int MIN_CTAS, int EPI, int CG, bool AG = false, bool GRP = false>
__global__ void __launch_bounds__(gemm_threads<CG, AG>(), MIN_CTAS)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, const GemmParams p,
                    const __grid_constant__ GroupArg<GRP> gp) {
  constexpr int B_ROWS = BN / CG;                 // rows of the B tile staged by this CTA
  constexpr int B_STAGE_BYTES = B_ROWS * BK * 2;
  constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  constexpr uint32_t TMEM_COLS = 2 * BN;
  constexpr int NSTG = staging_bufs<BN, CG, STAGES, AG>();
  constexpr int STAGING_BYTES = staging_bytes<CG>();
  constexpr int EPW = epilogue_warps<CG>();

  // 128B-swizzled TMA/UMMA tiles need 1024-byte alignment; the dynamic smem window of a kernel
  // without static smem starts 1024-aligned (checked below, reported through the error word).
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = smem_u32(smem_raw);
  const uint32_t sA = smem_base;
  const uint32_t sB = smem_base + STAGES * A_STAGE_BYTES;
  const uint32_t sStage = smem_base + STAGES * STAGE_BYTES;
  const uint32_t sBounce = sStage + NSTG * STAGING_BYTES;
  const uint32_t bars = sBounce + ag_bounce_bytes<CG, AG>();  // 8-byte mbarriers
  const uint32_t full_bar = bars;
  const uint32_t empty_bar = bars + 8 * STAGES;
  const uint32_t tfull_bar = bars + 16 * STAGES;
  const uint32_t tempty_bar = tfull_bar + 16;
  const uint32_t gather_bar = tempty_bar + 16;   // 4 barriers (one per bounce buffer)
  uint32_t* tmem_ptr_smem =
      reinterpret_cast<uint32_t*>(smem_raw + STAGES * STAGE_BYTES + NSTG * STAGING_BYTES + ag_bounce_bytes<CG, AG>() +
                                  16 * STAGES + 64);
  volatile int* abort_flag = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);
  if ((smem_base & 1023u) != 0u) {
    if (threadIdx.x == 0) atomicCAS(p.err_word, 0, ERR_GEMM_SMEM_ALIGN);
    return;
  }

  const int warp_idx = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int cta_rank = (CG == 2) ? static_cast<int>(cluster_ctarank()) : 0;
  const int unit = blockIdx.x / CG;               // CTA (CG = 1) or CTA pair (CG = 2) index
  const int num_units = gridDim.x / CG;

  if (threadIdx.x == 0) {
    DBG_STAMP(0);
    DBG_STAMP_NS(12);
    *abort_flag = 0;
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar + 8 * s, 1);
      mbar_init(empty_bar + 8 * s, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(tfull_bar + 8 * i, 1);
      mbar_init(tempty_bar + 8 * i, EPW * CG);   // the epilogue warps of both CTAs of a pair
    }
    for (int i = 0; i < 4; ++i) mbar_init(gather_bar + 8 * i, 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
  }
  if (warp_idx == 1) {
    if constexpr (CG == 2) {
      tmem_alloc_2sm(smem_u32(tmem_ptr_smem), TMEM_COLS);
      tmem_relinquish_2sm();
    } else {
      tmem_alloc(smem_u32(tmem_ptr_smem), TMEM_COLS);
      tmem_relinquish();
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if constexpr (CG == 2) cluster_sync_all();   // the peer's barriers exist before anything arrives on them
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  if (threadIdx.x == 0) DBG_STAMP(1);
  // Programmatic dependent launch: everything above (barrier init, TMEM allocation, descriptor
  // prefetch) may run while the previous kernel on the stream is still draining; from here on
  // global memory written by it is read, so wait for it.  Our own dependents may start their
  // prologue as soon as this grid has begun.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");

  int total_tiles = p.num_m_tiles * p.num_n_tiles * p.k_splits;
  if constexpr (GRP) total_tiles = gp.total_tiles;
  // tile index -> problem, tile origin (of the CTA pair), k-block range
  auto map_tile = [&](int tile) -> TileMap {
    TileMap t;
    if constexpr (GRP) {
      int g = 0;
      while (g + 1 < gp.n && tile >= gp.g[g].tile_end) ++g;
      const GroupDesc& d = gp.g[g];
      const int local = tile - (g > 0 ? gp.g[g - 1].tile_end : 0);
      const int split = local % d.k_splits;
      const int t2 = local / d.k_splits;
      t.m0 = (t2 % d.num_m_tiles) * (BM * CG);
      t.n0 = (t2 / d.num_m_tiles) * BN;
      t.kb0 = split * d.kb_per_split;
      t.kb1 = min(t.kb0 + d.kb_per_split, d.kb_total);
      t.M = d.M; t.N = d.N;
      t.a = &gp.a[g]; t.b = &gp.b[g]; t.c = &gp.c[g];
    } else {
      const int split = tile % p.k_splits;
      const int t2 = tile / p.k_splits;
      t.m0 = (p.n_fastest ? t2 / p.num_n_tiles : t2 % p.num_m_tiles) * (BM * CG);
      t.n0 = (p.n_fastest ? t2 % p.num_n_tiles : t2 / p.num_m_tiles) * BN;
      t.kb0 = split * p.kb_per_split;
      t.kb1 = min(t.kb0 + p.kb_per_split, p.kb_total);
      t.M = p.M; t.N = p.N;
      t.a = &tmA; t.b = &tmB; t.c = &tmC;
    }
    return t;
  };

  if (warp_idx == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      bool ok = true;
      int ag_tile_seen = -1;
      long long w_empty = 0;   // bring-up: cycles this producer spent waiting for a free smem slot
      for (int tile = unit; tile < total_tiles && ok; tile += num_units) {
        const TileMap tm = map_tile(tile);
        const int m0 = tm.m0 + cta_rank * BM;
        const int n0 = tm.n0;
        const int nb0 = n0 + cta_rank * B_ROWS;   // first B row staged by this CTA
        const int kb0 = tm.kb0, kb1 = tm.kb1;
        for (int kb = kb0; kb < kb1; ++kb) {
          const long long tw0 = TRACE_ON(p) ? clock64() : 0;
          if (!mbar_wait(empty_bar + 8 * stage, phase ^ 1, abort_flag, p.err_word, ERR_GEMM_PRODUCER)) {
            ok = false;
            break;
          }
          if (TRACE_ON(p)) w_empty += clock64() - tw0;
          const uint32_t fb = full_bar + 8 * stage;
          // CG = 2: the bytes of both CTAs are reported to the leader's barrier, armed by the leader
          if (CG == 1 || cta_rank == 0) mbar_arrive_expect_tx(fb, CG * STAGE_BYTES);
          const int k0 = kb * BK;
          const uint32_t a_dst = sA + stage * A_STAGE_BYTES;
          const uint32_t b_dst = sB + stage * B_STAGE_BYTES;
          auto load = [&](uint32_t dst, const CUtensorMap* tm, int c0, int c1) {
            if constexpr (CG == 2) tma_load_2d_2sm(dst, tm, fb, c0, c1);
            else tma_load_2d(dst, tm, fb, c0, c1);
          };
          if (!p.a_mn) {
            load(a_dst, tm.a, k0, m0);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) load(a_dst + j * 8192, tm.a, m0 + 64 * j, k0);
          }
          if ((CG == 1 || AG) && p.ag_bytes > 0 && ag_tile_seen != n0) {
            // the gather warps of all CTAs pull W from the peers tile by tile: wait for this n-tile
            constexpr int AG_SLICE = ag_slice<CG>();
            const int j = n0 / BN;
            const long long left = p.ag_bytes - static_cast<long long>(j) * p.ag_tile_bytes;
            const int expect = static_cast<int>(((left < p.ag_tile_bytes ? left : p.ag_tile_bytes) + AG_SLICE - 1) / AG_SLICE);
            const long long t0 = clock64();
            int seen = 0;
            for (;;) {
              asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(p.ag_sync + 2 + j) : "memory");
              if (seen >= expect) break;
              if (clock64() - t0 > B200_WAIT_TIMEOUT_CYCLES) {
                *abort_flag = 1;
                atomicCAS(p.err_word, 0, ERR_GEMM_AG_WAIT);
                break;
              }
            }
            asm volatile("fence.proxy.async;" ::: "memory");  // peers' bytes were written through the async proxy of other SMs
            ag_tile_seen = n0;
          }
          if (!p.b_mn) {
            load(b_dst, tm.b, k0, nb0);
          } else {
#pragma unroll
            for (int j = 0; j < B_ROWS / 64; ++j) load(b_dst + j * 8192, tm.b, nb0 + 64 * j, k0);
          }
          if (kb == kb0 && tile == unit) DBG_STAMP(2);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
      DBG_STAMP(10);
      if (TRACE_ON(p)) p.dbg[blockIdx.x * 32 + 17] = w_empty;
    }
    __syncwarp();  // lanes 1..31 must not reach the CTA barrier before lane 0 (bar.sync counts warps)
  } else if (warp_idx == 1) {
    // ===================== MMA issuer =====================
    uint32_t stage = 0, phase = 0;
    int lt = 0;
    bool ok = true;
    long long w_full = 0, w_tempty = 0;   // bring-up: cycles waiting for operands / for a drained accumulator
    for (int tile = unit; tile < total_tiles && ok && (CG == 1 || cta_rank == 0); tile += num_units, ++lt) {
      const TileMap tm = map_tile(tile);
      const int kb0 = tm.kb0, kb1 = tm.kb1;
      const uint32_t as = lt & 1, aph = (lt >> 1) & 1;
      long long tw0 = TRACE_ON(p) ? clock64() : 0;
      if (!mbar_wait(tempty_bar + 8 * as, aph ^ 1, abort_flag, p.err_word, ERR_GEMM_MMA)) break;
      if (TRACE_ON(p)) w_tempty += clock64() - tw0;
      tcgen05_fence_after();
      const uint32_t tmem_d = tmem_base + as * BN;
      for (int kb = kb0; kb < kb1; ++kb) {
        tw0 = TRACE_ON(p) ? clock64() : 0;
        if (!mbar_wait(full_bar + 8 * stage, phase, abort_flag, p.err_word, ERR_GEMM_MMA)) {
          ok = false;
          break;
        }
        if (TRACE_ON(p)) w_full += clock64() - tw0;
        tcgen05_fence_after();
        if (lane == 0 && kb == kb0 && lt == 0) DBG_STAMP(3);
        if (lane == 0) {
          const uint32_t a_src = sA + stage * A_STAGE_BYTES;
          const uint32_t b_src = sB + stage * B_STAGE_BYTES;
#pragma unroll
          for (int kk = 0; kk < BK / UMMA_K; ++kk) {
            const uint64_t ad = umma_desc_at(p.a_desc_base, a_src + kk * p.a_kstep);
            const uint64_t bd = umma_desc_at(p.b_desc_base, b_src + kk * p.b_kstep);
            if constexpr (CG == 2) umma_f16_ss_2sm(tmem_d, ad, bd, p.idesc, (kb > kb0 || kk > 0) ? 1u : 0u);
            else umma_f16_ss(tmem_d, ad, bd, p.idesc, (kb > kb0 || kk > 0) ? 1u : 0u);
          }
          // frees the smem slot (in both CTAs of a pair) when these MMAs retire
          if constexpr (CG == 2) umma_commit_2sm(empty_bar + 8 * stage);
          else umma_commit(empty_bar + 8 * stage);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (ok && lane == 0) {  // accumulator ready for the epilogue (of both CTAs of a pair)
        if constexpr (CG == 2) umma_commit_2sm(tfull_bar + 8 * as);
        else umma_commit(tfull_bar + 8 * as);
      }
      if (lane == 0 && lt == 0) DBG_STAMP(4);
      __syncwarp();
    }
    if (TRACE_ON(p) && lane == 0 && (CG == 1 || cta_rank == 0)) {
      p.dbg[blockIdx.x * 32 + 14] = w_full;
      p.dbg[blockIdx.x * 32 + 15] = w_tempty;
    }
  } else if ((CG == 1 && warp_idx == 6) || (CG == 2 && AG && warp_idx == 2 + EPW)) {
    // ===================== gather warp (fused all-gather of the B operand) =====================
    // One lane moves this CTA's share of W with the bulk-copy engine: slices travel peer shard -> smem bounce
    // buffer -> local full buffer, round-robin over the CTAs in n-tile order so the first tiles land first.
    // With NBUF > 1 buffers NBUF - 1 loads are in flight while one buffer drains.  (Plain bulk loads: a TMA
    // multicast load must only ever read this GPU's own memory.)
    if (lane == 0 && p.ag_bytes > 0) {
      constexpr int AG_SLICE = ag_slice<CG>();
      constexpr int NBUF = ag_nbuf<CG>();
      constexpr int AHEAD = NBUF > 1 ? NBUF - 1 : 1;
      const long long nslices = (p.ag_bytes + AG_SLICE - 1) / AG_SLICE;
      const long long n_mine = nslices > blockIdx.x ? (nslices - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
      auto slice_bytes = [&](long long k) {
        const long long byte = (blockIdx.x + k * gridDim.x) * static_cast<long long>(AG_SLICE);
        return p.ag_bytes - byte < AG_SLICE ? p.ag_bytes - byte : static_cast<long long>(AG_SLICE);
      };
      auto issue_load = [&](long long k) {   // a slice may straddle a shard boundary: one bulk copy per owner
        long long byte = (blockIdx.x + k * gridDim.x) * static_cast<long long>(AG_SLICE);
        long long left = slice_bytes(k);
        const uint32_t buf = sBounce + static_cast<uint32_t>(k % NBUF) * AG_SLICE;
        const uint32_t bar = gather_bar + 8 * static_cast<uint32_t>(k % NBUF);
        mbar_arrive_expect_tx(bar, static_cast<uint32_t>(left));
        uint32_t done = 0;
        while (left > 0) {
          const long long gbyte = p.ag_off + byte;
          const int peer = static_cast<int>(gbyte / p.ag_shard_bytes);
          const long long in_shard = gbyte - peer * p.ag_shard_bytes;
          long long seg = p.ag_shard_bytes - in_shard;
          if (seg > left) seg = left;
          bulk_load_g2s(buf + done, p.ag_peers[peer] + in_shard, static_cast<uint32_t>(seg), bar);
          byte += seg; left -= seg; done += static_cast<uint32_t>(seg);
        }
      };
      auto announce = [&](long long k) {     // slice k is in the full buffer: count it for its n-tile
        const long long byte = (blockIdx.x + k * gridDim.x) * static_cast<long long>(AG_SLICE);
        asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(p.ag_sync + 2 + static_cast<int>(byte / p.ag_tile_bytes)) : "memory");
      };
      bool ok = true;
      for (long long k = 0; k < AHEAD && k < n_mine; ++k) issue_load(k);
      for (long long k = 0; k < n_mine && ok; ++k) {
        const uint32_t b = static_cast<uint32_t>(k % NBUF);
        if (!mbar_wait(gather_bar + 8 * b, static_cast<uint32_t>((k / NBUF) & 1), abort_flag, p.err_word, ERR_GEMM_AG_WAIT)) {
          ok = false;
          break;
        }
        const long long byte = (blockIdx.x + k * gridDim.x) * static_cast<long long>(AG_SLICE);
        bulk_store_s2g(p.ag_full + p.ag_off + byte, sBounce + b * AG_SLICE, static_cast<uint32_t>(slice_bytes(k)));
        tma_store_commit();
        if constexpr (NBUF > 1) {
          bulk_wait_group<1>();              // everything but the store just issued is in global memory:
          if (k > 0) announce(k - 1);        // slice k - 1 can be announced and its buffer refilled
          if (k + AHEAD < n_mine) issue_load(k + AHEAD);
        } else {
          bulk_wait_group<0>();
          announce(k);
          if (k + 1 < n_mine) issue_load(k + 1);
        }
      }
      if constexpr (NBUF > 1) {
        bulk_wait_group<0>();
        if (ok && n_mine > 0) announce(n_mine - 1);
      }
    }
    __syncwarp();
  } else {
    // ===================== epilogue (4 warps; 8 in a CTA pair) =====================
    // TMEM -> registers -> (bias / ReLU / aux / mask / residual, selected at compile time by EPI) ->
    // 128B-swizzled smem tile -> TMA store (or TMA reduce-add).  Global traffic is whole 128-byte
    // rows, M/N tails are clipped by the tensor map, and the residual / mask operands of chunk
    // c+1 are prefetched under chunk c.
    constexpr bool E_BIAS = (EPI & EPI_BIAS) != 0, E_RELU = (EPI & EPI_RELU) != 0, E_AUX = (EPI & EPI_AUX) != 0,
                   E_MASK = (EPI & EPI_MASK) != 0, E_RES = (EPI & EPI_RES) != 0, E_F32 = (EPI & EPI_F32) != 0,
                   E_ACC = (EPI & EPI_ACC) != 0, E_CSUM = (EPI & EPI_CSUM) != 0, E_SMAX = (EPI & EPI_SMAX) != 0,
                   E_NOSTORE = (EPI & EPI_NOSTORE) != 0, E_RSCALE = (EPI & EPI_RSCALE) != 0,
                   E_GELU = (EPI & EPI_GELU) != 0;
    constexpr float kLog2e = 1.4426950408889634f;
    const int q = warp_idx & 3;  // TMEM lane quarter this warp may access
    const int ew = warp_idx - 2;  // epilogue warp index
    constexpr int NCH = (BN / 32) / (EPW / 4);   // 32-column chunks per warp
    const int cbase = (ew >> 2) * NCH;           // first chunk of this warp (second warp of a quarter: upper half)
    const uint32_t stage_base = sStage + ew * 4096;
    const int swz = lane & 7;
    uint32_t nstore = 0;  // TMA stores issued by this warp (selects the staging buffer)
    int lt = 0;
    long long w_tfull = 0, w_stage = 0;   // bring-up: cycles waiting for the MMAs / for a free staging buffer
    for (int tile = unit; tile < total_tiles; tile += num_units, ++lt) {
      const TileMap tm = map_tile(tile);
      const int m0 = tm.m0 + cta_rank * BM;
      const int n0 = tm.n0 + cbase * 32;  // this warp's first column
      const uint32_t as = lt & 1, aph = (lt >> 1) & 1;
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < tm.M;
      const size_t row_c = static_cast<size_t>(row_ok ? row : 0);
      [[maybe_unused]] const float* res_row = E_RES ? p.residual + row_c * p.ldr : nullptr;
      [[maybe_unused]] const bf16* msk_row = E_MASK ? p.mask_src + row_c * p.ldc : nullptr;
      [[maybe_unused]] float4 nres[8];
      [[maybe_unused]] uint4 nmsk[4];
      [[maybe_unused]] float nbias = 0.f;
      // EPI_RSCALE: this row's scale.  EPI_SMAX: this row's shift in log2 units, its target column and the
      // running (sum of stored values, max of raw values) over the columns this warp owns in the tile
      [[maybe_unused]] float rowa = 0.f;
      [[maybe_unused]] int tgt = -1;
      [[maybe_unused]] float s_sum[4] = {0.f, 0.f, 0.f, 0.f};
      [[maybe_unused]] float s_max4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};   // independent chains
      if constexpr (E_RSCALE) rowa = row_ok ? __ldg(p.row_a + row) : 0.f;
      if constexpr (E_SMAX) {
        rowa = lm_row_shift(row_ok ? __ldg(p.row_a + row) : 0.f, __ldg(p.wmax), __ldg(p.wmax + 1)) * kLog2e;
        if (row_ok && p.targets != nullptr) {
          const long long t64 = p.targets[row];
          tgt = (t64 >= 0 && t64 < p.N) ? static_cast<int>(t64) : -1;
        }
      }
      // prefetch the epilogue operands of chunk 0 while the MMAs of this tile are still running
      if constexpr (E_BIAS) nbias = (n0 + lane < p.N) ? __ldg(p.bias + n0 + lane) : 0.f;
      if constexpr (E_RES) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          nres[j] = (row_ok && n0 < p.N) ? __ldg(reinterpret_cast<const float4*>(res_row + n0) + j) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      if constexpr (E_MASK) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          nmsk[j] = (row_ok && n0 < p.N) ? __ldg(reinterpret_cast<const uint4*>(msk_row + n0) + j) : make_uint4(0u, 0u, 0u, 0u);
      }
      const long long tw0 = TRACE_ON(p) ? clock64() : 0;
      if (!mbar_wait(tfull_bar + 8 * as, aph, abort_flag, p.err_word, ERR_GEMM_EPILOGUE)) break;
      if (TRACE_ON(p)) w_tfull += clock64() - tw0;
      tcgen05_fence_after();
      if (warp_idx == 2 && lane == 0 && lt == 0) DBG_STAMP(5);
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN + cbase * 32;
      uint32_t rr[2][32];  // TMEM loads are software-pipelined: chunk c+1 is in flight under chunk c
      tmem_ld_32x32b_x32(taddr, rr[0]);
      if constexpr (E_SMAX) {
        // LM head.  exp2 runs on the 16-lane/clk SFU pipe (8 cycles per warp instruction) and everything that
        // consumes its results queues up behind it.  The loop is skewed by one chunk: iteration c holds exp2 of
        // chunk c AND the sums / bf16 packing / staging stores of chunk c-1 in one branch-free block, independent
        // of each other, so ptxas interleaves SFU and ALU work (582 -> 561 us for the 16320 x 50257 x 256 head;
        // starting the second warp of a lane quarter 150-900 cycles late on top of that changed nothing).  Rare
        // cases (vocabulary tail, target column) are handled on the raw values before the block; waits and the
        // TMA store sit between blocks.
        float ex[2][32];
#pragma unroll
        for (int c = 0; c <= NCH; ++c) {
          const int col0 = n0 + c * 32;
          uint32_t* r = rr[c & 1];
          if (c < NCH) {
            tmem_ld_wait_dep(r);
            if (c + 1 < NCH) tmem_ld_32x32b_x32(taddr + (c + 1) * 32, rr[(c + 1) & 1]);
            if (c == NCH - 1) {  // accumulator fully in registers: hand the TMEM buffer back early
              tcgen05_fence_before();
              __syncwarp();
              if (lane == 0) {
                if constexpr (CG == 2) mbar_arrive_leader(tempty_bar + 8 * as);
                else mbar_arrive(tempty_bar + 8 * as);
              }
            }
            if (col0 + 32 > p.N) {   // columns past N (zero-filled B rows) must not count
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (col0 + j >= p.N) r[j] = 0xff800000u;   // -inf
            }
            if (tgt >= col0 && tgt < col0 + 32) {   // rare: this thread holds its row's target column
              uint32_t tv = 0u;
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (col0 + j == tgt) tv = r[j];
              p.tgt_val[row] = __uint_as_float(tv);
            }
          }
          [[maybe_unused]] const uint32_t stage_buf = stage_base + (nstore % NSTG) * STAGING_BYTES;
          if constexpr (!E_NOSTORE) {
            if (c >= 1 && ((c - 1) & 1) == 0) {   // the store issued NSTG stores ago must be done reading this buffer
              const long long ts0 = TRACE_ON(p) ? clock64() : 0;
              if (lane == 0) {
                if constexpr (NSTG == 2) tma_store_wait_read1();
                else tma_store_wait_read0();
              }
              __syncwarp();
              if (TRACE_ON(p)) w_stage += clock64() - ts0;
            }
          }
          // ---- the block: exp2 of chunk c, sums and staging of chunk c-1 ----
          if (c < NCH) {
            float* e = ex[c & 1];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float v = __uint_as_float(r[j]);
              s_max4[j & 3] = fmaxf(s_max4[j & 3], v);
              e[j] = fast_exp2(fmaf(v, kLog2e, -rowa));
            }
          }
          if (c >= 1) {
            const float* e = ex[(c - 1) & 1];
#pragma unroll
            for (int j = 0; j < 32; ++j) s_sum[j & 3] += e[j];
            if constexpr (!E_NOSTORE) {
              const uint32_t my_row = stage_buf + lane * 128;
              const int half = ((c - 1) & 1) * 4;
#pragma unroll
              for (int j = 0; j < 4; ++j)
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + (((half + j) ^ swz) << 4)),
                             "r"(pack_bf16x2(e[8 * j], e[8 * j + 1])), "r"(pack_bf16x2(e[8 * j + 2], e[8 * j + 3])),
                             "r"(pack_bf16x2(e[8 * j + 4], e[8 * j + 5])), "r"(pack_bf16x2(e[8 * j + 6], e[8 * j + 7]))
                             : "memory");
            }
          }
          // ---- between blocks: two chunks fill one 32 x 64 bf16 staging tile ----
          if constexpr (!E_NOSTORE) {
            if (c >= 1 && ((c - 1) & 1) == 1) {
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) {
                const int sc0 = col0 - 64;   // first column of chunk c-2
                if (sc0 < p.N && m0 + q * 32 < p.M) tma_store_2d(&tmC, stage_buf, sc0, m0 + q * 32);
                tma_store_commit();
              }
              ++nstore;
            }
          }
        }
      } else {
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        uint32_t* r = rr[c & 1];
        [[maybe_unused]] float4 res[8];
        [[maybe_unused]] uint4 msk[4];
        [[maybe_unused]] const float biasv = nbias;
        const int col0 = n0 + c * 32;
        if constexpr (E_RES) {
#pragma unroll
          for (int j = 0; j < 8; ++j) res[j] = nres[j];
        }
        if constexpr (E_MASK) {
#pragma unroll
          for (int j = 0; j < 4; ++j) msk[j] = nmsk[j];
        }
        if (c + 1 < NCH) {   // host guarantees N % 32 == 0 whenever res / mask / aux are used
          const int ncol0 = col0 + 32;
          if constexpr (E_BIAS) nbias = (ncol0 + lane < p.N) ? __ldg(p.bias + ncol0 + lane) : 0.f;
          if constexpr (E_RES) {
            if (row_ok && ncol0 < p.N) {
#pragma unroll
              for (int j = 0; j < 8; ++j) nres[j] = __ldg(reinterpret_cast<const float4*>(res_row + ncol0) + j);
            }
          }
          if constexpr (E_MASK) {
            if (row_ok && ncol0 < p.N) {
#pragma unroll
              for (int j = 0; j < 4; ++j) nmsk[j] = __ldg(reinterpret_cast<const uint4*>(msk_row + ncol0) + j);
            }
          }
        }
        tmem_ld_wait_dep(r);
        if (c + 1 < NCH) tmem_ld_32x32b_x32(taddr + (c + 1) * 32, rr[(c + 1) & 1]);
        if (c == NCH - 1) {  // accumulator fully in registers: hand the TMEM buffer back early
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) {
            if constexpr (CG == 2) mbar_arrive_leader(tempty_bar + 8 * as);
            else mbar_arrive(tempty_bar + 8 * as);
          }
        }
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
        if constexpr (E_BIAS) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += __shfl_sync(0xffffffffu, biasv, j);
        }
        auto store_aux = [&]() {
          if (row_ok && col0 < p.N) {
            bf16* dst = p.aux_out + static_cast<size_t>(row) * p.ldc + col0;
#pragma unroll
            for (int j = 0; j < 32; j += 8)
              *reinterpret_cast<uint4*>(dst + j) =
                  make_uint4(pack_bf16x2(v[j], v[j + 1]), pack_bf16x2(v[j + 2], v[j + 3]),
                             pack_bf16x2(v[j + 4], v[j + 5]), pack_bf16x2(v[j + 6], v[j + 7]));
          }
        };
        if constexpr (E_RELU) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.0f);
        }
        if constexpr (E_GELU && !E_MASK) {
          if constexpr (E_AUX) store_aux();   // GELU: the backward pass needs the pre-activation
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = gelu_fwd(v[j], p.act);
        }
        if constexpr (E_AUX && !(E_GELU && !E_MASK)) store_aux();   // ReLU: copy of the activated output (its own mask)
        if constexpr (E_MASK) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t w[4] = {msk[j].x, msk[j].y, msk[j].z, msk[j].w};
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              const float2 f = unpack_bf16x2(w[t]);
              if constexpr (E_GELU) {   // mask_src holds the pre-activation: multiply by gelu'(z)
                v[8 * j + 2 * t] *= gelu_grad(f.x, p.act);
                v[8 * j + 2 * t + 1] *= gelu_grad(f.y, p.act);
              } else {
                if (!(f.x > 0.0f)) v[8 * j + 2 * t] = 0.0f;
                if (!(f.y > 0.0f)) v[8 * j + 2 * t + 1] = 0.0f;
              }
            }
          }
        }
        if constexpr (E_RES) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            v[4 * j] += res[j].x; v[4 * j + 1] += res[j].y; v[4 * j + 2] += res[j].z; v[4 * j + 3] += res[j].w;
          }
        }
        if constexpr (E_RSCALE) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] *= rowa;
        }
        if constexpr (!E_NOSTORE) {
        // stage: fp32 -> one 32x32 tile per chunk; bf16 -> two chunks fill one 32x64 tile
        constexpr bool kStoreEvery = E_F32;
        const bool first_half = kStoreEvery || (c & 1) == 0;
        const uint32_t stage_buf = stage_base + (nstore % NSTG) * STAGING_BYTES;
        const uint32_t my_row = stage_buf + lane * 128;
        if (first_half) {  // the store issued NSTG stores ago must be done reading this buffer
          const long long ts0 = TRACE_ON(p) ? clock64() : 0;
          if (lane == 0) {
            if constexpr (NSTG == 2) tma_store_wait_read1();
            else tma_store_wait_read0();
          }
          __syncwarp();
          if (TRACE_ON(p)) w_stage += clock64() - ts0;
        }
        if constexpr (E_F32) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + ((j ^ swz) << 4)),
                         "f"(v[4 * j]), "f"(v[4 * j + 1]), "f"(v[4 * j + 2]), "f"(v[4 * j + 3]) : "memory");
        } else {
          const int half = (c & 1) * 4;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + (((half + j) ^ swz) << 4)),
                         "r"(pack_bf16x2(v[8 * j], v[8 * j + 1])), "r"(pack_bf16x2(v[8 * j + 2], v[8 * j + 3])),
                         "r"(pack_bf16x2(v[8 * j + 4], v[8 * j + 5])), "r"(pack_bf16x2(v[8 * j + 6], v[8 * j + 7]))
                         : "memory");
        }
        if (kStoreEvery || (c & 1) == 1) {
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            const int sc0 = E_F32 ? col0 : col0 - 32;
            if (sc0 < tm.N && m0 + q * 32 < tm.M) {
              if constexpr (E_ACC) tma_reduce_add_2d(tm.c, stage_buf, sc0, m0 + q * 32);
              else tma_store_2d(tm.c, stage_buf, sc0, m0 + q * 32);
            }
            tma_store_commit();
          }
          ++nstore;
        }
        }  // !E_NOSTORE
        if constexpr (E_CSUM) {
          // column sums of this warp's 32 x 32 block by a butterfly transpose-reduce (31 shuffles): lane l
          // ends up with the sum of column l over the warp's 32 rows (rows past M are exact zeros: their
          // A rows were zero-filled by TMA), then one fire-and-forget reduction per lane
          if constexpr (!E_F32) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __bfloat162float(__float2bfloat16_rn(v[j]));   // what the consumer of the bf16 output sees
          }
#pragma unroll
          for (int sft = 16; sft >= 1; sft >>= 1) {
            const bool up = (lane & sft) != 0;
#pragma unroll
            for (int j = 0; j < sft; ++j) {
              const float send = up ? v[j] : v[j + sft];
              const float keep = up ? v[j + sft] : v[j];
              v[j] = keep + __shfl_xor_sync(0xffffffffu, send, sft);
            }
          }
          if (col0 + lane < p.N) atomicAdd(p.colsum + col0 + lane, v[0]);
        }
      }
      }  // generic epilogue loop
      if constexpr (E_SMAX) {   // one (sum, max) partial per row and per NCH * 32-column span, coalesced over the warp
        if (row_ok)
          p.part[static_cast<size_t>(n0 / (NCH * 32)) * p.M + row] =
              make_float2((s_sum[0] + s_sum[1]) + (s_sum[2] + s_sum[3]),
                          fmaxf(fmaxf(s_max4[0], s_max4[1]), fmaxf(s_max4[2], s_max4[3])));
      }
    }
    if (warp_idx == 2 && lane == 0) DBG_STAMP(6);
    if (TRACE_ON(p) && warp_idx == 2 && lane == 0) {
      p.dbg[blockIdx.x * 32 + 16] = w_tfull;
      p.dbg[blockIdx.x * 32 + 18] = w_stage;
    }
    if (lane == 0) tma_store_wait_all();
    __syncwarp();
    if (warp_idx == 2 && lane == 0) DBG_STAMP(7);
  }

  tcgen05_fence_before();
  __syncthreads();
  if (p.ag_bytes > 0 && threadIdx.x == 0) {
    // the last CTA to leave knows every producer has passed its wait: leave the counters at zero
    if (atomicAdd(p.ag_sync + 1, 1) == static_cast<int>(gridDim.x) - 1) {
      for (int j = 0; j < p.num_n_tiles; ++j) p.ag_sync[2 + j] = 0;
      __threadfence();
      p.ag_sync[1] = 0;
    }
  }
  if (threadIdx.x == 0) DBG_STAMP(8);
  if constexpr (CG == 2) cluster_sync_all();   // nobody frees TMEM / leaves while the peer may still touch it
  if (warp_idx == 1) {
    tcgen05_fence_after();
    if constexpr (CG == 2) tmem_dealloc_2sm(tmem_base, TMEM_COLS);
    else tmem_dealloc(tmem_base, TMEM_COLS);
    if (lane == 0) DBG_STAMP(9);
    if (lane == 0) DBG_STAMP_NS(13);
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
struct TmapKey {
  const void* base;
  uint64_t inner, outer, ld;
  uint32_t bi, bo;
  int dev, esz;
  bool operator==(const TmapKey& o) const {
    return base == o.base && inner == o.inner && outer == o.outer && ld == o.ld && bi == o.bi &&
           bo == o.bo && dev == o.dev && esz == o.esz;
  }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    size_t h = reinterpret_cast<size_t>(k.base);
    auto mix = [&h](uint64_t v) { h ^= v + 0x9e3779b97f4a7c15ULL + (h << 6) + (h >> 2); };
    mix(k.inner); mix(k.outer); mix(k.ld); mix(k.bi); mix(k.bo); mix((uint64_t)k.dev); mix((uint64_t)k.esz);
    return h;
  }
};

int get_tmap(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t ld,
             uint32_t bi, uint32_t bo, int esz = 2) {
  static std::mutex mu;
  static std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> cache;
  int dev = 0;
  cudaGetDevice(&dev);
  TmapKey key{base, inner, outer, ld, bi, bo, dev, esz};
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) {
      *out = it->second;
      return 0;
    }
  }
  int rc = b200_make_tmap_2d(out, base, inner, outer, ld, bi, bo, esz);
  if (rc != 0) return rc;
  std::lock_guard<std::mutex> g(mu);
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, *out);
  return 0;
}

struct AgArgs {
  const void* peers_dev;
  void* full_base;
  long long off, bytes, shard_bytes;
  int* sync;
  int max_tiles;
};

long long* g_gemm_dbg = nullptr;
// programmatic dependent launch of the GEMM kernels: measured gain 0.4 % of a step, so opt-in (B200GPT_PDL=1)
bool g_gemm_pdl = []() { const char* e = getenv("B200GPT_PDL"); return e != nullptr && e[0] == '1'; }();

// CTA-pair (cta_group::2) GEMMs are the default for 256-wide tiles; B200GPT_GEMM_2SM=0 turns them off
bool g_gemm_2sm = []() { const char* e = getenv("B200GPT_GEMM_2SM"); return e == nullptr || e[0] != '0'; }();

// extra epilogue operands of the LM-head kernels
struct EpiExtra {
  const float* row_a;
  const float* wmax;
  float2* part;
  const int64_t* targets;
  float* tgt_val;
  int smax, nostore, rscale;
  int* num_parts_out;   // host: number of column spans (rows of `part`) the launch wrote
  int max_parts;
};

struct GemmOverrides {
  int a_lbo, a_sbo, a_kstep, b_lbo, b_sbo, b_kstep;  // < 0 => default
  int force_bn, force_splits;
  int no_tma_store;
};

template <int BN, int STAGES, int MIN_CTAS, int EPI, int CG, bool AG = false, bool GRP = false>
int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const GemmParams& p,
           int grid, cudaStream_t stream, const GroupArg<GRP>& gp = GroupArg<GRP>{}) {
  constexpr int smem = STAGES * (A_STAGE_BYTES + (BN / CG) * BK * 2) + staging_bufs<BN, CG, STAGES, AG>() * staging_bytes<CG>() +
                       ag_bounce_bytes<CG, AG>() + 16 * STAGES + 64 + 16;
  static_assert(smem <= 227 * 1024, "gemm: shared memory budget exceeded");
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!configured[dev & 63]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, STAGES, MIN_CTAS, EPI, CG, AG, GRP>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, STAGES, MIN_CTAS, EPI, CG, AG, GRP>,
                                         cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    configured[dev & 63] = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(gemm_threads<CG, AG>());
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (g_gemm_pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (CG == 2) {   // CTA pair: a cluster of two CTAs is placed on the two SMs of one TPC
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 2;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  B200_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_tcgen05_kernel<BN, STAGES, MIN_CTAS, EPI, CG, AG, GRP>, tmA, tmB, tmC, p, gp));
  return 0;
}

// the epilogue combinations the framework uses (one pruned kernel each)
template <int BN, int STAGES, int MIN_CTAS, int CG>
int launch_epi(int epi, const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
               const GemmParams& p, int grid, cudaStream_t st) {
  switch (epi) {
#define B200_EPI_CASE(E) case (E): return launch<BN, STAGES, MIN_CTAS, (E), CG>(tmA, tmB, tmC, p, grid, st);
    B200_EPI_CASE(0)                                                   // plain bf16 (QKV, dgrad, logits)
    B200_EPI_CASE(EPI_BIAS)                                            // + bias
    B200_EPI_CASE(EPI_BIAS | EPI_RELU)                                 // MLP up
    B200_EPI_CASE(EPI_MASK)                                            // dgrad fused with the ReLU mask
    B200_EPI_CASE(EPI_MASK | EPI_CSUM)                                 // ... and with the bias gradient of that layer
    B200_EPI_CASE(EPI_F32)                                             // plain fp32
    B200_EPI_CASE(EPI_F32 | EPI_ACC)                                   // wgrad / split-K (TMA reduce-add)
    B200_EPI_CASE(EPI_F32 | EPI_BIAS)
    B200_EPI_CASE(EPI_F32 | EPI_BIAS | EPI_RES)                        // attention out-proj + residual
    B200_EPI_CASE(EPI_F32 | EPI_RES)
    B200_EPI_CASE(EPI_F32 | EPI_BIAS | EPI_RELU | EPI_RES)             // MLP down + residual (inference)
    B200_EPI_CASE(EPI_F32 | EPI_BIAS | EPI_RELU | EPI_AUX | EPI_RES)   // MLP down + residual (training)
    B200_EPI_CASE(EPI_SMAX)                                            // LM head: exp(logit - shift) + row statistics
    B200_EPI_CASE(EPI_SMAX | EPI_NOSTORE)                              // ... statistics only (evaluation)
    B200_EPI_CASE(EPI_RSCALE)                                          // per-row scale (LM head dX)
    B200_EPI_CASE(EPI_BIAS | EPI_GELU)                                 // GELU MLP up (inference)
    B200_EPI_CASE(EPI_BIAS | EPI_GELU | EPI_AUX)                       // GELU MLP up (training: + pre-activation)
    B200_EPI_CASE(EPI_F32 | EPI_BIAS | EPI_GELU | EPI_RES)             // GELU MLP down + residual (inference)
    B200_EPI_CASE(EPI_F32 | EPI_BIAS | EPI_GELU | EPI_AUX | EPI_RES)   // GELU MLP down + residual (training)
    B200_EPI_CASE(EPI_MASK | EPI_GELU)                                 // dgrad times gelu'(pre-activation)
    B200_EPI_CASE(EPI_MASK | EPI_GELU | EPI_CSUM)                      // ... and the bias gradient
#undef B200_EPI_CASE
    default:
      return b200_set_error(-1, "gemm: epilogue combination 0x%x is not instantiated", epi);
  }
}

// CTA pairs with the gather warp (fused FSDP all-gather): the epilogues of the forward GEMMs that consume a weight
// first - QKV, attention out-projection, MLP up / down (ReLU and GELU)
int launch_epi_ag_pair(int epi, const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                       const GemmParams& p, int grid, cudaStream_t st) {
  switch (epi) {
#define B200_EPI_CASE(E) case (E): return launch<256, 5, 1, (E), 2, true>(tmA, tmB, tmC, p, grid, st);
    B200_EPI_CASE(0)
    B200_EPI_CASE(EPI_BIAS | EPI_RELU)
    B200_EPI_CASE(EPI_F32 | EPI_BIAS | EPI_RES)
    B200_EPI_CASE(EPI_F32 | EPI_BIAS | EPI_RELU | EPI_RES)
    B200_EPI_CASE(EPI_F32 | EPI_BIAS | EPI_RELU | EPI_AUX | EPI_RES)
    B200_EPI_CASE(EPI_BIAS | EPI_GELU)
    B200_EPI_CASE(EPI_BIAS | EPI_GELU | EPI_AUX)
    B200_EPI_CASE(EPI_F32 | EPI_BIAS | EPI_GELU | EPI_RES)
    B200_EPI_CASE(EPI_F32 | EPI_BIAS | EPI_GELU | EPI_AUX | EPI_RES)
#undef B200_EPI_CASE
    default:
      return -100;   // not instantiated: the caller falls back to single-CTA tiles
  }
}
constexpr bool ag_pair_epilogue(int epi) {
  return epi == 0 || epi == (EPI_BIAS | EPI_RELU) || epi == (EPI_F32 | EPI_BIAS | EPI_RES) ||
         epi == (EPI_F32 | EPI_BIAS | EPI_RELU | EPI_RES) || epi == (EPI_F32 | EPI_BIAS | EPI_RELU | EPI_AUX | EPI_RES) ||
         epi == (EPI_BIAS | EPI_GELU) || epi == (EPI_BIAS | EPI_GELU | EPI_AUX) ||
         epi == (EPI_F32 | EPI_BIAS | EPI_GELU | EPI_RES) || epi == (EPI_F32 | EPI_BIAS | EPI_GELU | EPI_AUX | EPI_RES);
}
// B200GPT_AG_2SM=0: fused all-gather on single-CTA tiles only (the first version of the kernel)
bool g_ag_2sm = []() { const char* e = getenv("B200GPT_AG_2SM"); return e == nullptr || e[0] != '0'; }();

// B200GPT_GEMM_STAGING2=0: keep the 6-stage / one-staging-set pair kernel for every epilogue
bool g_two_staging = []() { const char* e = getenv("B200GPT_GEMM_STAGING2"); return e == nullptr || e[0] != '0'; }();

int gemm_impl(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
              int a_mn, int b_mn, const float* bias, const float* residual, int ldr, int act,
              int out_fp32, int accumulate, void* aux_out, const void* mask_src, int cfg,
              const GemmOverrides* ov, void* stream, const AgArgs* ag = nullptr, float* colsum = nullptr,
              const EpiExtra* ex = nullptr) {
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem %dx%dx%d", M, N, K);
  B200_REQUIRE(!accumulate || out_fp32, "gemm: accumulate needs an fp32 output");
  B200_REQUIRE(act >= 0 && act <= 3, "gemm: activation %d not supported (0 none, 1 relu, 2 gelu, 3 gelu-tanh)", act);

  int bn = (cfg & 0x3ff);
  if (ov && ov->force_bn > 0) bn = ov->force_bn;
  if (bn == 0) bn = (N >= 256) ? 256 : 128;  // 128x256 tiles read 25% less operand data from L2 per FLOP
  B200_REQUIRE(bn == 128 || bn == 256, "gemm: tile N %d not supported", bn);

  // CTA pairs (cta_group::2, 256 x 256 tiles): cfg bit 10 forces them on, bit 11 off
  const bool fused_ag = ag != nullptr && ag->bytes > 0;
  const int epi_bits = (bias ? EPI_BIAS : 0) | (act == 1 ? EPI_RELU : 0) | (act >= 2 ? EPI_GELU : 0) |
                       (aux_out ? EPI_AUX : 0) | (mask_src ? EPI_MASK : 0) | (residual ? EPI_RES : 0) |
                       (out_fp32 ? EPI_F32 : 0) | (accumulate ? EPI_ACC : 0) | (colsum ? EPI_CSUM : 0);
  int cg = (bn == 256 && M > BM && g_gemm_2sm && (!fused_ag || (g_ag_2sm && ag_pair_epilogue(epi_bits)))) ? 2 : 1;
  if (cfg & 0x400) cg = 2;
  if (cfg & 0x800) cg = 1;
  B200_REQUIRE(cg == 1 || (bn == 256 && (!fused_ag || ag_pair_epilogue(epi_bits))),
               "gemm: CTA pairs need 256-wide tiles (and, with a fused all-gather, one of the forward epilogues)");

  GemmParams p{};
  p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.ldr = ldr;
  p.num_m_tiles = ceil_div(M, BM * cg);
  p.num_n_tiles = ceil_div(N, bn);
  p.kb_total = ceil_div(K, BK);
  const int tiles = p.num_m_tiles * p.num_n_tiles;
  const int sms = b200_num_sms();
  int splits = 1;
  const bool plain = (bias == nullptr && residual == nullptr && act == 0 && aux_out == nullptr &&
                      mask_src == nullptr);
  B200_REQUIRE(mask_src == nullptr || act != 1, "gemm: a mask epilogue takes act 0 (ReLU mask) or 2 / 3 (GELU derivative)");
  const int slots = sms * ((bn == 128) ? 2 : 1) / cg;
  if (accumulate && out_fp32 && plain && tiles < slots) {
    splits = slots / tiles;
    const int max_splits = p.kb_total / 4 > 0 ? p.kb_total / 4 : 1;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
  }
  if (ov && ov->force_splits > 0) splits = ov->force_splits;
  B200_REQUIRE(splits == 1 || (accumulate && plain), "gemm: split-K needs a plain accumulating epilogue");
  p.kb_per_split = ceil_div(p.kb_total, splits);
  p.k_splits = ceil_div(p.kb_total, p.kb_per_split);

  p.idesc = umma_idesc_bf16(BM * cg, bn, a_mn, b_mn);
  // K-major SW128: 8-row groups 1024 B apart (SBO), LBO unused (1); K advance = 32 B in the row.
  // MN-major SW128: 64-wide MN groups one TMA box (8192 B) apart (LBO), 8-k groups 1024 B apart (SBO);
  //                 K advance of 16 = two k groups = 2048 B.
  uint32_t a_lbo = a_mn ? 8192 : 16, a_sbo = 1024, a_kstep = a_mn ? 2048 : 32;
  uint32_t b_lbo = b_mn ? 8192 : 16, b_sbo = 1024, b_kstep = b_mn ? 2048 : 32;
  if (ov) {
    if (ov->a_lbo >= 0) a_lbo = ov->a_lbo;
    if (ov->a_sbo >= 0) a_sbo = ov->a_sbo;
    if (ov->a_kstep >= 0) a_kstep = ov->a_kstep;
    if (ov->b_lbo >= 0) b_lbo = ov->b_lbo;
    if (ov->b_sbo >= 0) b_sbo = ov->b_sbo;
    if (ov->b_kstep >= 0) b_kstep = ov->b_kstep;
  }
  p.a_desc_base = umma_smem_desc_base(a_lbo, a_sbo);
  p.b_desc_base = umma_smem_desc_base(b_lbo, b_sbo);
  p.a_kstep = a_kstep; p.b_kstep = b_kstep;
  p.a_mn = a_mn; p.b_mn = b_mn;
  p.bias = bias; p.residual = residual; p.C = C;
  p.aux_out = reinterpret_cast<bf16*>(aux_out);
  p.mask_src = reinterpret_cast<const bf16*>(mask_src);
  p.colsum = colsum;
  p.act = act; p.out_fp32 = out_fp32; p.accumulate = accumulate;
  p.err_word = b200_error_word();
  p.dbg = g_gemm_dbg;
  const bool nostore = ex != nullptr && ex->nostore;
  if (ex != nullptr) {
    B200_REQUIRE(plain && !out_fp32 && !accumulate && colsum == nullptr && splits == 1,
                 "gemm: softmax / row-scale epilogues take a plain bf16 output");
    p.row_a = ex->row_a; p.wmax = ex->wmax; p.part = ex->part; p.targets = ex->targets; p.tgt_val = ex->tgt_val;
    if (ex->smax) {
      const int span = (cg == 2) ? bn / 2 : bn;   // columns one epilogue warp owns in a tile
      const int parts = p.num_n_tiles * (bn / span);
      B200_REQUIRE(parts <= ex->max_parts, "lmhead_gemm: %d column spans exceed the partial buffer (%d)", parts, ex->max_parts);
      B200_REQUIRE(ex->row_a && ex->wmax && ex->part && ex->tgt_val, "lmhead_gemm: missing statistics buffers");
      if (ex->num_parts_out) *ex->num_parts_out = parts;
    } else {
      B200_REQUIRE(ex->row_a != nullptr, "gemm: row scale vector missing");
    }
  }
  if (ag != nullptr && ag->bytes > 0) {
    B200_REQUIRE(!b_mn && splits == 1, "ag_gemm: B must be a K-major weight and the GEMM un-split");
    B200_REQUIRE(ag->off % 16 == 0 && ag->bytes % 16 == 0 && ag->shard_bytes % 16 == 0 && ag->shard_bytes > 0,
                 "ag_gemm: byte ranges must be multiples of 16");
    B200_REQUIRE(reinterpret_cast<const uint8_t*>(B) == reinterpret_cast<const uint8_t*>(ag->full_base) + ag->off,
                 "ag_gemm: B must point at offset ag_off of the full buffer");
    p.ag_peers = reinterpret_cast<const uint8_t* const*>(ag->peers_dev);
    p.ag_full = reinterpret_cast<uint8_t*>(ag->full_base);
    p.ag_off = ag->off; p.ag_bytes = ag->bytes; p.ag_shard_bytes = ag->shard_bytes;
    p.ag_sync = ag->sync;
    p.ag_tile_bytes = bn * K * 2;
    const int slice = cg == 2 ? ag_slice<2>() : ag_slice<1>();
    B200_REQUIRE(p.ag_tile_bytes % slice == 0 && ldb == K && p.num_n_tiles <= ag->max_tiles,
                 "ag_gemm: needs a contiguous weight with (tile N x K x 2) %% %d == 0 and <= %d n-tiles (N=%d K=%d)",
                 slice, ag->max_tiles, N, K);
  }
  B200_REQUIRE(p.err_word != nullptr, "gemm: no CUDA device");

  CUtensorMap tmA, tmB;
  int rc;
  if (!a_mn) rc = get_tmap(&tmA, A, K, M, lda, BK, BM);
  else       rc = get_tmap(&tmA, A, M, K, lda, 64, BK);
  if (rc != 0) return rc;
  if (!b_mn) rc = get_tmap(&tmB, B, K, N, ldb, BK, bn / cg);
  else       rc = get_tmap(&tmB, B, N, K, ldb, 64, BK);
  if (rc != 0) return rc;

  // output through TMA when the layout allows it (16-byte aligned base and row stride)
  CUtensorMap tmC = tmA;
  const int esz = out_fp32 ? 4 : 2;
  B200_REQUIRE(nostore || ((reinterpret_cast<uintptr_t>(C) & 15) == 0 && (static_cast<int64_t>(ldc) * esz) % 16 == 0),
               "gemm: output must be 16-byte aligned with a 16-byte multiple row stride (ldc=%d)", ldc);
  if (residual != nullptr || mask_src != nullptr || aux_out != nullptr)
    B200_REQUIRE(N % 32 == 0 && ldc % 8 == 0 && (residual == nullptr || ldr % 4 == 0) &&
                     (reinterpret_cast<uintptr_t>(residual) & 15) == 0 &&
                     (reinterpret_cast<uintptr_t>(mask_src) & 15) == 0 &&
                     (reinterpret_cast<uintptr_t>(aux_out) & 15) == 0,
                 "gemm: residual / mask / aux epilogues need N %% 32 == 0 and 16-byte aligned rows (N=%d)", N);
  p.tma_store = 1;
  p.n_fastest = (cfg & 0x1000) && !fused_ag ? 1 : 0;
  if (!nostore) {
    rc = get_tmap(&tmC, C, N, M, ldc, out_fp32 ? 32 : 64, 32, esz);
    if (rc != 0) return rc;
  }
  const int total = tiles * p.k_splits;
  const int ctas_per_sm = (bn == 128) ? 2 : 1;
  const int cap = sms * ctas_per_sm / cg;
  const int grid = (total < cap ? total : cap) * cg;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int epi = (bias ? EPI_BIAS : 0) | (act == 1 ? EPI_RELU : 0) | (act >= 2 ? EPI_GELU : 0) | (aux_out ? EPI_AUX : 0) |
                  (mask_src ? EPI_MASK : 0) | (residual ? EPI_RES : 0) | (out_fp32 ? EPI_F32 : 0) |
                  (accumulate ? EPI_ACC : 0) | (colsum ? EPI_CSUM : 0) | ((ex && ex->smax) ? EPI_SMAX : 0) |
                  (nostore ? EPI_NOSTORE : 0) | ((ex && ex->rscale) ? EPI_RSCALE : 0);
  if (bn == 128) return launch_epi<128, 2, 2, 1>(epi, tmA, tmB, tmC, p, grid, st);
  if (cg == 2 && fused_ag) return launch_epi_ag_pair(epi, tmA, tmB, tmC, p, grid, st);
  if (cg == 2 && g_two_staging) {
    // LM head: 5 operand stages leave room for a second staging set per epilogue warp (the plain logits GEMM
    // measured 4 % slower that way and keeps 6 stages)
    if (epi == EPI_SMAX) return launch<256, 5, 1, EPI_SMAX, 2>(tmA, tmB, tmC, p, grid, st);
  }
  if (cg == 2) return launch_epi<256, 6, 1, 2>(epi, tmA, tmB, tmC, p, grid, st);
  return launch_epi<256, 4, 1, 1>(epi, tmA, tmB, tmC, p, grid, st);
}

}  // namespace

extern "C" {

int b200_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb,
                   int ldc, int a_mn_major, int b_mn_major, const float* bias, const float* residual,
                   int ldr, int act, int out_fp32, int accumulate, void* aux_out,
                   const void* mask_src, int cfg, void* stream) {
  return gemm_impl(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, bias, residual, ldr, act,
                   out_fp32, accumulate, aux_out, mask_src, cfg, nullptr, stream);
}

// GEMM with the all-gather of its weight operand fused in (FSDP forward): W = B[N,K] (K-major) sits at
// byte offset ag_off of a flat parameter group whose full copy lives at full_base; the group is
// sharded evenly over `world` ranks (shard_bytes each) and peers_dev is a device array of every
// rank's shard base pointer (symmetric memory).  The kernel pulls the weight's bytes over NVLink
// while its prologue and A loads run, then computes C = A . W^T with the usual epilogue.
// ag_sync: 2 + ag_sync_tiles ints in device memory (ag_sync_tiles >= number of n-tiles), zero on entry and on exit.  All peers' shards must be final
// (b200_rs_adamw ends with a cross-GPU barrier for exactly that).
int b200_ag_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                      const float* bias, const float* residual, int ldr, int act, int out_fp32, void* aux_out,
                      const void* peers_dev, void* full_base, int64_t ag_off, int64_t ag_bytes,
                      int64_t shard_bytes, int* ag_sync, int ag_sync_tiles, void* stream) {
  AgArgs ag{peers_dev, full_base, static_cast<long long>(ag_off), static_cast<long long>(ag_bytes),
            static_cast<long long>(shard_bytes), ag_sync, ag_sync_tiles};
  return gemm_impl(A, B, C, M, N, K, lda, ldb, ldc, 0, 0, bias, residual, ldr, act, out_fp32, 0, aux_out, nullptr,
                   0, nullptr, stream, &ag);
}

// Masked dgrad with the bias gradient fused in (MLP backward, reference models/gpt.py:33-41):
//   C[M,N] (bf16) = (A[M,K] . B) * (mask_src > 0)     and     colsum[n] += sum_m C[m,n]
// so the [tokens, 4*dim] activation gradient is written once and never re-read for its column sums.
int b200_gemm_mask_colsum_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                               int b_mn_major, const void* mask_src, float* colsum, int cfg, void* stream) {
  B200_REQUIRE(mask_src != nullptr && colsum != nullptr, "gemm_mask_colsum: mask_src and colsum are required");
  return gemm_impl(A, B, C, M, N, K, lda, ldb, ldc, 0, b_mn_major, nullptr, nullptr, 0, 0, 0, 0, nullptr, mask_src,
                   cfg, nullptr, stream, nullptr, colsum);
}

// LM head, forward (reference models/gpt.py:219,231 + F.cross_entropy main-single.py:95-96): the logits
// GEMM whose epilogue turns each logit into E = exp(logit - shift[m]) straight out of tensor memory, where
// shift[m] = lm_row_shift(xnorm[m], lmstate[0], lmstate[1]) is fixed before the kernel starts (see common.cuh),
// so the softmax needs no second pass over the row: the kernel stores E as bf16 (or nothing at all when
// E_out is null: evaluation) and leaves, per row and per column span, (sum of E, max logit) in part[span][M],
// plus the raw target logit in tgt_val[m].  b200_lmhead_finish turns these into loss / accuracy / the
// gradient's row scale.  *num_parts (host int) receives the number of spans written (<= max_parts).
int b200_lmhead_gemm_bf16(const void* X, const void* W, void* E_out, int M, int V, int K, int ldx, int ldw, int lde,
                          const float* xnorm, const float* lmstate, void* part, int max_parts, int* num_parts,
                          const int64_t* targets, float* tgt_val, int cfg, void* stream) {
  EpiExtra ex{xnorm, lmstate, reinterpret_cast<float2*>(part), targets, tgt_val, 1, E_out == nullptr ? 1 : 0, 0,
              num_parts, max_parts};
  return gemm_impl(X, W, E_out, M, V, K, ldx, ldw, lde, 0, 0, nullptr, nullptr, 0, 0, 0, 0, nullptr, nullptr, cfg,
                   nullptr, stream, nullptr, nullptr, &ex);
}

// C[M,N] (bf16) = diag(row_scale) . (A . op(B)): plain GEMM with a per-row factor in the epilogue (LM head dX:
// the softmax normalisation 1/sum and the loss scale are per-row constants that commute with the GEMM).
int b200_gemm_rowscale_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                            int b_mn_major, const float* row_scale, int cfg, void* stream) {
  EpiExtra ex{row_scale, nullptr, nullptr, nullptr, nullptr, 0, 0, 1, nullptr, 0};
  return gemm_impl(A, B, C, M, N, K, lda, ldb, ldc, 0, b_mn_major, nullptr, nullptr, 0, 0, 0, 0, nullptr, nullptr, cfg,
                   nullptr, stream, nullptr, nullptr, &ex);
}

// Same with the activation selectable: act 0 / 1 = ReLU (mask_src = activated output), 2 / 3 = GELU erf / tanh form
// (mask_src = pre-activation, the product is taken with gelu'(mask_src)).  Reference models/gpt.py:14-15 takes any
// activation callable; F.gelu is the other one this library fuses.
int b200_gemm_dact_colsum_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                               int b_mn_major, const void* mask_src, float* colsum, int act, int cfg, void* stream) {
  B200_REQUIRE(mask_src != nullptr, "gemm_dact_colsum: mask_src is required");
  return gemm_impl(A, B, C, M, N, K, lda, ldb, ldc, 0, b_mn_major, nullptr, nullptr, 0, act >= 2 ? act : 0, 0, 0, nullptr,
                   mask_src, cfg, nullptr, stream, nullptr, colsum);
}

int b200_gemm_set_debug_buffer(void* buf) {
  g_gemm_dbg = reinterpret_cast<long long*>(buf);
  return 0;
}

// Grouped weight gradients: C_i[M_i, N_i] (fp32) += A_i^T . B_i for up to 4 problems in ONE launch, A_i stored
// [K, M_i] and B_i stored [K, N_i] (both MN-major: dZ and X of a Linear layer as they sit in memory, K = tokens).
// The problems share one persistent grid of CTA pairs and one K split, so the fixed cost of a launch (cluster
// scheduling, tensor-memory allocation, pipeline fill, the reduce-add tail) is paid once per transformer block
// instead of once per weight.  Returns -100 when CTA pairs are switched off (the caller launches one by one).
int b200_gemm_wgrad_group(int n, const void* const* A, const void* const* B, void* const* C, const int* M,
                          const int* N, int K, const int* lda, const int* ldb, const int* ldc, void* stream) {
  B200_REQUIRE(n >= 1 && n <= MAX_GROUP && K > 0, "gemm_wgrad_group: 1..%d problems (got %d)", MAX_GROUP, n);
  if (!g_gemm_2sm) return -100;
  constexpr int CG = 2, BN = 256;
  GroupParams gp{};
  GemmParams p{};
  const int kb_total = ceil_div(K, BK);
  int base_tiles = 0;
  for (int i = 0; i < n; ++i) {
    B200_REQUIRE(M[i] > 0 && N[i] > 0, "gemm_wgrad_group: empty problem %d", i);
    base_tiles += ceil_div(M[i], BM * CG) * ceil_div(N[i], BN);
  }
  const int slots = b200_num_sms() / CG;
  int splits = slots / base_tiles;
  const int max_splits = kb_total / 4 > 0 ? kb_total / 4 : 1;
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  const int kb_per_split = ceil_div(kb_total, splits);
  const int k_splits = ceil_div(kb_total, kb_per_split);
  int end = 0;
  for (int i = 0; i < n; ++i) {
    B200_REQUIRE((reinterpret_cast<uintptr_t>(C[i]) & 15) == 0 && (static_cast<int64_t>(ldc[i]) * 4) % 16 == 0,
                 "gemm_wgrad_group: output %d must be 16-byte aligned with a 16-byte multiple row stride", i);
    GroupDesc& d = gp.g[i];
    d.M = M[i]; d.N = N[i];
    d.num_m_tiles = ceil_div(M[i], BM * CG);
    d.num_n_tiles = ceil_div(N[i], BN);
    d.k_splits = k_splits; d.kb_total = kb_total; d.kb_per_split = kb_per_split;
    end += d.num_m_tiles * d.num_n_tiles * k_splits;
    d.tile_end = end;
    int rc = get_tmap(&gp.a[i], A[i], M[i], K, lda[i], 64, BK);
    if (rc != 0) return rc;
    rc = get_tmap(&gp.b[i], B[i], N[i], K, ldb[i], 64, BK);
    if (rc != 0) return rc;
    rc = get_tmap(&gp.c[i], C[i], N[i], M[i], ldc[i], 32, 32, 4);
    if (rc != 0) return rc;
  }
  gp.n = n; gp.total_tiles = end;
  p.M = M[0]; p.N = N[0]; p.K = K; p.ldc = ldc[0];
  p.num_m_tiles = gp.g[0].num_m_tiles; p.num_n_tiles = gp.g[0].num_n_tiles;
  p.kb_total = kb_total; p.kb_per_split = kb_per_split; p.k_splits = k_splits;
  p.idesc = umma_idesc_bf16(BM * CG, BN, 1, 1);
  p.a_desc_base = umma_smem_desc_base(8192, 1024);
  p.b_desc_base = umma_smem_desc_base(8192, 1024);
  p.a_kstep = 2048; p.b_kstep = 2048;
  p.a_mn = 1; p.b_mn = 1;
  p.C = C[0];
  p.out_fp32 = 1; p.accumulate = 1;
  p.err_word = b200_error_word();
  p.dbg = g_gemm_dbg;
  p.tma_store = 1;
  B200_REQUIRE(p.err_word != nullptr, "gemm: no CUDA device");
  const int grid = (end < slots ? end : slots) * CG;
  return launch<256, 6, 1, EPI_F32 | EPI_ACC, 2, false, true>(gp.a[0], gp.b[0], gp.c[0], p, grid,
                                                              static_cast<cudaStream_t>(stream), gp);
}

// Bring-up / regression entry: same GEMM with the UMMA descriptor fields overridable, so one GPU
// session can check alternative encodings (values < 0 keep the defaults).  fp32 output, no epilogue.
int b200_gemm_debug(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb,
                    int ldc, int a_mn_major, int b_mn_major, int a_lbo, int a_sbo, int a_kstep,
                    int b_lbo, int b_sbo, int b_kstep, int force_bn, int force_splits, void* stream) {
  GemmOverrides ov{a_lbo, a_sbo, a_kstep, b_lbo, b_sbo, b_kstep, force_bn & 0xffff, force_splits, (force_bn >> 16) & 1};
  return gemm_impl(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, nullptr, nullptr, 0, 0, 1,
                   force_splits > 1 ? 1 : 0, nullptr, nullptr, ((force_bn >> 17) & 1) ? 0x400 : 0x800, &ov, stream);
}

}  // extern "C"
