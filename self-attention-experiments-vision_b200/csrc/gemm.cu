// Persistent warp-specialised bf16 GEMM for sm_100a: TMA -> 128B-swizzled smem ring -> tcgen05.mma
// (accumulators double-buffered in TMEM) -> fused epilogue -> swizzled smem slab -> TMA store.
// One CTA (or a cta_group::2 pair on 256-row tiles) per SM, 608 threads:
//   warp 0       TMA producer   } whole warp walks the loop, one elect.sync lane issues: warp-uniform
//   warp 1       MMA issuer     } control flow keeps descriptors in uniform registers (no per-MMA
//                               } vote / R2UR waterfall); warp 1 also owns the TMEM allocation
//   warps 2-17   epilogue: every 128-row x 128-byte output slab is produced by all 16 warps (four per
//                TMEM sub-partition, a quarter of the slab's columns each): tcgen05.ld -> bias / GELU /
//                dGELU / LayerScale / drop-path / residual -> smem slab -> fence.proxy.async ->
//                bar.arrive on the slot's named barrier.  They never wait for a store or each other.
//   warp 18      store warp: bar.sync on the slot's barrier -> one TMA store (coalesced, clipped at
//                the M/N edges) -> wait_group.read -> slot handed back (slot_free mbarrier, or the
//                TMA load of the next epilogue input into it).  Epilogue *inputs* (fp32 residual
//                stream, saved GELU pre-activation) land in the very slab buffer they are combined
//                in, four slabs ahead.
// Operands may be K-major or MN-major (selected by the smem + instruction descriptors), so
// forward (X W), dX (dY W^T) and dW (X^T dY) all run here without any transposed copies.  dW is
// split over K (tokens) and the partial tiles are combined with TMA reduce-add (fp32) in L2.
#include <stdlib.h>

#include <mutex>

#include "common.cuh"
#include "host.h"

namespace sav {

constexpr int BM = 128;  // UMMA M
constexpr int BK = 64;   // 64 bf16 = one 128-byte swizzle row
constexpr int kEpiWarps = 16;                        // four per TMEM sub-partition / scheduler
constexpr int kColParts = kEpiWarps / 4;             // column parts of a slab (one per warp of a sub-partition)
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kStoreWarp = 2 + kEpiWarps;          // issues the epilogue's TMA stores / input loads
constexpr int kGemmThreads = 64 + kEpiThreads + 32;
constexpr int SLAB_BYTES = BM * 128;        // 128 rows x 128 B
constexpr int kOutBytes = 4 * SLAB_BYTES;   // epilogue slab ring: 4 x 16 KB (or 2 x 32 KB)

enum EpiMode { EPI_PLAIN = 0, EPI_GELU2 = 1, EPI_F32 = 2, EPI_DGELU = 3, EPI_ACCUM = 4 };

struct GemmParams {
  int64_t M, N, K;
  int m_blocks, n_blocks, split_k, k_blocks, k_per_split;
  int epi;          // EpiMode
  int act;          // EPI_PLAIN: 1 = tanh-GELU on the single output
  int has_aux_in;   // EPI_F32: fp32 residual present; EPI_DGELU: always
  const float* bias;
  const float* col_scale;
  const float* row_scale;
  int64_t rows_per_scale;
  float* colsum;    // EPI_DGELU only: column sums of the (fp32) result accumulated here, or null
  // LayerNorm fused into the fp32-stream epilogue (EPI_F32 + residual): ln_tiles = N / BN (1 or 2)
  // N-tiles make a complete row and are processed back to back by the same CTA (pair); 0 = off
  int ln_tiles;
  const float* ln_gamma;
  const float* ln_beta;
  float* ln_mean;
  float* ln_rstd;
  float ln_eps;
  int batch_inner;  // batched (BATCH=true) only: size of the inner batch dim (heads)
  int nbatch;       // total batches (1 when not batched)
  long long* dbg;   // development only: timeline buffer (tests/tools/gemm_timeline.py) or null
  // shared-memory split between the operand pipeline and the epilogue slab ring (the sum is fixed
  // per tile shape): short-K GEMMs with an epilogue that streams as many bytes as the operands
  // (FC2 dX with its pre-activation input, FC1 with two outputs, the LayerNorm mode) trade pipeline
  // stages they do not need for more 16 KB slab slots in flight
  int stages, nslot;
  int debug;        // development only (SAV_GEMM_DEBUG): 1 = skip TMA loads, 2 = skip MMAs
};

constexpr int kSmemBudget = 232448;  // 227 KB opt-in limit per CTA

template <int BN, bool B_MN, int CTAS>
struct GemmCfg {
  static constexpr int kBCols = BN / CTAS;  // B columns (n) staged by this CTA
  static constexpr int kABytes = BM * BK * 2;
  // MN-major B is staged as whole 64-column swizzle atoms (a 96-column half over-fetches 32)
  static constexpr int kBBoxes = (kBCols + 63) / 64;
  static constexpr int kBBytes = B_MN ? kBBoxes * (BK * 128) : kBCols * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  // after the stages and the slab ring: 256 B of barriers, 2 KB of per-tile column-sum accumulators
  // (2 x BN floats, EPI_DGELU with colsum).  The dynamic segment is declared __align__(1024), so
  // no alignment slack is budgeted.
  static constexpr int kBarBytes = 320;
  static constexpr int kTailBytes = kBarBytes + 2048;
  static constexpr int kFit = (kSmemBudget - kOutBytes - kTailBytes) / kStageBytes;
  static constexpr int kStages = kFit > 6 ? 6 : kFit;
  static constexpr int kSmem = kStages * kStageBytes + kOutBytes + kTailBytes;
  static constexpr uint32_t kTmemCols = 512;  // kAcc accumulators of BN fp32 columns
  // accumulator ring: 2 x 256 / 2 x 192 / 4 x 128 columns.  Four matter for the fused LayerNorm,
  // where a row block keeps its N / BN accumulators until the row statistics are known: with
  // N = 384 = 3 x 128 the fourth lets the next row block's MMAs start under that epilogue.
  static constexpr int kAcc = BN <= 128 ? 4 : 2;
  static_assert(kStageBytes % 1024 == 0, "stage must keep 1024 B alignment");
  static_assert(kStages >= 3, "too few stages");
};

__device__ __forceinline__ void named_bar_arrive(int id, int count) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* m, const void* src, int c0,
                                                  int c1) {
  asm volatile(
      "cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::
          "l"(reinterpret_cast<uint64_t>(m)),
      "r"(smem_u32(src)), "r"(c0), "r"(c1)
      : "memory");
}

// tanh-GELU with the hardware tanh (MUFU): the result is rounded to bf16 right after, so the
// 2^-11 relative error of tanh.approx is below the output rounding.
__device__ __forceinline__ float gelu_fast(float x) {
  const float u = x * (0.7978845608028654f + 0.035677408136300125f * x * x);
  return 0.5f * x * (1.f + tanh_fast(u));
}
__device__ __forceinline__ float dgelu_fast(float x) {
  const float x2 = x * x;
  const float u = x * (0.7978845608028654f + 0.035677408136300125f * x2);
  const float t = tanh_fast(u);
  const float du = 0.7978845608028654f + 0.10703222440890037f * x2;
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * du;
}
// The same two functions on element pairs through the packed fp32x2 pipe: the epilogue of the MLP
// GEMMs is FP32-issue bound (two warps per scheduler), so halving the instruction count is what
// shortens it.  Agrees with the scalar versions to fp32 rounding (results are rounded to bf16).
__device__ __forceinline__ void gelu_fast2(float& a, float& b) {
  const uint64_t c0 = f2_pack(0.7978845608028654f, 0.7978845608028654f);
  const uint64_t c1 = f2_pack(0.035677408136300125f, 0.035677408136300125f);
  const uint64_t hf = f2_pack(0.5f, 0.5f);
  const uint64_t x = f2_pack(a, b);
  const uint64_t u = f2_mul(x, f2_fma(c1, f2_mul(x, x), c0));
  float ua, ub;
  f2_unpack(u, ua, ub);
  const uint64_t hx = f2_mul(hf, x);
  f2_unpack(f2_fma(hx, f2_pack(tanh_fast(ua), tanh_fast(ub)), hx), a, b);
}
// (va, vb) *= gelu'(xa, xb)
__device__ __forceinline__ void dgelu_mul2(float& va, float& vb, float xa, float xb) {
  const uint64_t c0 = f2_pack(0.7978845608028654f, 0.7978845608028654f);
  const uint64_t c1 = f2_pack(0.035677408136300125f, 0.035677408136300125f);
  const uint64_t c2 = f2_pack(0.10703222440890037f, 0.10703222440890037f);
  const uint64_t hf = f2_pack(0.5f, 0.5f), one = f2_pack(1.f, 1.f), neg1 = f2_pack(-1.f, -1.f);
  const uint64_t x = f2_pack(xa, xb);
  const uint64_t x2 = f2_mul(x, x);
  const uint64_t u = f2_mul(x, f2_fma(c1, x2, c0));
  float ua, ub;
  f2_unpack(u, ua, ub);
  const uint64_t t = f2_pack(tanh_fast(ua), tanh_fast(ub));
  const uint64_t du = f2_fma(c2, x2, c0);
  const uint64_t sech2 = f2_fma(f2_mul(neg1, t), t, one);              // 1 - t^2
  const uint64_t lhs = f2_fma(hf, t, hf);                              // 0.5 (1 + t)
  const uint64_t g = f2_fma(f2_mul(f2_mul(hf, x), sech2), du, lhs);
  f2_unpack(f2_mul(f2_pack(va, vb), g), va, vb);
}

// bias / LayerScale / drop-path on NV consecutive columns starting at `col`
template <int NV>
__device__ __forceinline__ void epi_affine(const GemmParams& p, float (&v)[NV], int col, float rs,
                                           const float4* bias_pre = nullptr) {
  if (col + NV <= p.N) {
    if (p.bias) {
#pragma unroll
      for (int j = 0; j < NV; j += 4) {
        const float4 b = bias_pre ? bias_pre[j / 4]
                                  : __ldg(reinterpret_cast<const float4*>(p.bias + col + j));
        v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
      }
    }
    if (p.col_scale) {
#pragma unroll
      for (int j = 0; j < NV; j += 4) {
        const float4 s = __ldg(reinterpret_cast<const float4*>(p.col_scale + col + j));
        v[j] *= s.x; v[j + 1] *= s.y; v[j + 2] *= s.z; v[j + 3] *= s.w;
      }
    }
  } else {  // N edge: columns >= N are clipped by the TMA store, only guard the loads
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      if (col + j < p.N) {
        if (p.bias) v[j] += p.bias[col + j];
        if (p.col_scale) v[j] *= p.col_scale[col + j];
      }
    }
  }
  if (p.row_scale) {
#pragma unroll
    for (int j = 0; j < NV; ++j) v[j] *= rs;
  }
}

// 16-byte chunk `c` (0..7) of row `r` of a [128 x 128 B] slab laid out for a SWIZZLE_128B TMA box
__device__ __forceinline__ uint32_t slab_off(int r, int c) {
  return (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4));
}
template <int NV>
__device__ __forceinline__ void st_bf16(uint8_t* slab, int r, int cb, const float (&v)[NV]) {
#pragma unroll
  for (int g = 0; g < NV / 8; ++g)
    *reinterpret_cast<uint4*>(slab + slab_off(r, cb + g)) =
        make_uint4(pack_bf16(v[g * 8 + 0], v[g * 8 + 1]), pack_bf16(v[g * 8 + 2], v[g * 8 + 3]),
                   pack_bf16(v[g * 8 + 4], v[g * 8 + 5]), pack_bf16(v[g * 8 + 6], v[g * 8 + 7]));
}

// Development instrumentation, compiled in only with -DSAV_GEMM_DEV (SAV_GEMM_DEV=1 python build.py):
//  * SAV_GEMM_DEBUG bits switch parts of the kernel off (see launch_gemm) for limit experiments;
//  * the timeline writes (event, clock64) pairs of CTA 0, 256 slots per role.
// The production build carries neither: the epilogue is instruction-issue bound.
#ifdef SAV_GEMM_DEV
#define GEMM_DBG(bits) (p.debug & (bits))
#define GEMM_STAMP(role, ev)                                         \
  do {                                                               \
    if (p.dbg && blockIdx.x == 0 && dbg_n < 255 && STAMP_LANE) {     \
      p.dbg[(role) * 512 + 2 * dbg_n] = (ev);                        \
      p.dbg[(role) * 512 + 2 * dbg_n + 1] = clock64();               \
      ++dbg_n;                                                       \
    }                                                                \
  } while (0)
#else
#define GEMM_DBG(bits) false
#define GEMM_STAMP(role, ev) do { (void)dbg_n; } while (0)
#endif

#define STAMP_LANE true

template <int BN, bool A_MN, bool B_MN, int CTAS, bool BATCH>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
            const __grid_constant__ CUtensorMap tma_d, const __grid_constant__ CUtensorMap tma_x,
            const __grid_constant__ CUtensorMap tma_y, const GemmParams p) {
  using Cfg = GemmCfg<BN, B_MN, CTAS>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;   // SW128 tiles need a 1024-byte aligned base (checked below)
  const int nst = p.stages;                                    // <= Cfg::kStages
  uint8_t* smem_out = smem + nst * Cfg::kStageBytes;           // 1024-aligned (stage bytes % 1024 == 0)
  uint8_t* smem_tail = smem + Cfg::kStages * Cfg::kStageBytes + kOutBytes;
  float* s_colsum = reinterpret_cast<float*>(smem_tail + Cfg::kBarBytes);   // [2][BN]
  // LayerNorm mode: the slab ring runs with one slot less and the last holds the per-row partial
  // statistics, [4 column parts][128 rows][sum, sum of squares] -- plain stores combined in a fixed
  // order, so the normalised output is reproducible run to run (no floating-point atomics)
  float* s_stats = reinterpret_cast<float*>(smem_out + (p.nslot - 1) * SLAB_BYTES);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_tail);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 4;
  uint64_t* aux_bar = tempty_bar + 4;  // [8]
  uint64_t* slot_free = aux_bar + 8;  // [8] slab slot has been read by its TMA store
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(slot_free + 8);
  static_assert((2 * Cfg::kStages + 8 + 16) * 8 + 4 <= Cfg::kBarBytes, "barrier block");

  // warp index through a shuffle so that the compiler knows the role branches are warp-uniform
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    if (smem_u32(smem) & 1023u) __trap();
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    tma_prefetch_desc(&tma_d);
    if (p.epi == EPI_GELU2 || p.has_aux_in) tma_prefetch_desc(&tma_x);
    if (p.ln_tiles) tma_prefetch_desc(&tma_y);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 4; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], kEpiWarps * CTAS);  // one arrival per epilogue warp (both CTAs)
    }
    for (int s = 0; s < 8; ++s) { mbar_init(&aux_bar[s], 1); mbar_init(&slot_free[s], 1); }
    fence_barrier_init();
  }
  if (warp == 1) {
    if (CTAS == 2) tmem_alloc_pair<Cfg::kTmemCols>(tmem_slot);
    else tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  }
  if (p.colsum)
    for (int i = threadIdx.x; i < 2 * BN; i += blockDim.x) s_colsum[i] = 0.f;
  tc_fence_before();
  __syncthreads();
  if (CTAS == 2) cluster_sync_all();  // the peer's barriers exist before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // a "tile" is BM*CTAS rows x BN columns, owned by one CTA (pair); rank = CTA inside the pair
  const int rank = CTAS == 2 ? (int)cluster_ctarank() : 0;
  const int unit = (int)blockIdx.x / CTAS;          // which CTA (pair) of the grid
  const int units = (int)gridDim.x / CTAS;
  const int tiles_mn = p.m_blocks * p.n_blocks;
  const int tiles_per_batch = tiles_mn * p.split_k;
  const int num_tiles = tiles_per_batch * (BATCH ? p.nbatch : 1);
  // Local tile sequence of this CTA (pair).  Normally tiles unit, unit + units, ...; with the fused
  // LayerNorm the ln_tiles N-tiles of a row block are consecutive on the same CTA (pair), so that
  // the epilogue sees complete rows.
  const int lt = p.ln_tiles;
  const int my_tiles_all = lt ? lt * max(0, (p.m_blocks - unit + units - 1) / units)
                              : (num_tiles - unit + units - 1) / units;
  auto tile_at = [&](int itl) { return lt ? (unit + (itl / lt) * units) * lt + (itl % lt) : unit + itl * units; };

  if (warp == 0) {
    // TMA producer: whole warp walks the loop, one elected lane issues (see the MMA warp)
    {
#undef STAMP_LANE
#define STAMP_LANE (lane == 0)
      int stage = 0;
      uint32_t phase = 0;
      int dbg_n = 0;
      for (int itl = 0; itl < my_tiles_all; ++itl) {
        const int tile = tile_at(itl);
        GEMM_STAMP(0, 1);
        const int z = BATCH ? tile / tiles_per_batch : 0;
        const int zi = BATCH ? z % p.batch_inner : 0, zo = BATCH ? z / p.batch_inner : 0;
        const int tl = tile - z * tiles_per_batch;
        const int split = tl / tiles_mn;
        const int mn = tl - split * tiles_mn;
        const int m_idx = (mn / p.n_blocks) * (BM * CTAS) + rank * BM;
        const int n_idx = (mn % p.n_blocks) * BN + rank * Cfg::kBCols;
        const int kb0 = split * p.k_per_split;
        const int kb1 = min(kb0 + p.k_per_split, p.k_blocks);
        for (int kb = kb0; kb < kb1; ++kb) {
          if (!(GEMM_DBG(512))) mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kABytes;
          // the leader's full barrier collects the bytes of both CTAs' loads
          if (GEMM_DBG(1)) {  // MMA-only rate: operands are whatever the smem holds
            if (rank == 0 && elect_one()) mbar_arrive(&full_bar[stage]);
            __syncwarp();
            if (++stage == nst) { stage = 0; phase ^= 1; }
            continue;
          }
          const int k_idx = kb * BK;
          if (elect_one()) {
          if (rank == 0) mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes * CTAS);
          auto load = [&](void* dst, const CUtensorMap* m, int c0, int c1) {
            if (BATCH) tma_load_4d(dst, m, &full_bar[stage], c0, c1, zi, zo);
            else if (CTAS == 2) tma_load_2d_pair(dst, m, &full_bar[stage], c0, c1);
            else tma_load_2d(dst, m, &full_bar[stage], c0, c1);
          };
          if (!A_MN) {
            load(sa, &tma_a, k_idx, m_idx);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) load(sa + j * (BK * 128), &tma_a, m_idx + j * 64, k_idx);
          }
          if (!B_MN) {
            load(sb, &tma_b, k_idx, n_idx);
          } else {
#pragma unroll
            for (int j = 0; j < Cfg::kBBoxes; ++j)
              load(sb + j * (BK * 128), &tma_b, n_idx + j * 64, k_idx);
          }
          }
          __syncwarp();
          if (++stage == nst) { stage = 0; phase ^= 1; }
        }
      }
#undef STAMP_LANE
#define STAMP_LANE true
    }
  } else if (warp == 1) {
    // The whole warp walks the loop (every lane polls the barriers) and one elected lane issues:
    // with warp-uniform control flow and operands ptxas keeps the descriptors in uniform registers.
    // Issuing from an `if (lane == 0)` region instead wraps every UTCHMMA in a vote / R2UR
    // waterfall loop that costs ~150 cycles per MMA -- more than a 128x192x16 MMA takes.
    if (rank == 0) {
#undef STAMP_LANE
#define STAMP_LANE (lane == 0)
      constexpr uint32_t idesc = make_idesc_bf16(BM * CTAS, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      int dbg_n = 0;
      for (; it < my_tiles_all; ++it) {
        const int tile = tile_at(it);
        GEMM_STAMP(1, 10);
        const int split = (tile % tiles_per_batch) / tiles_mn;
        const int kb0 = split * p.k_per_split;
        const int kb1 = min(kb0 + p.k_per_split, p.k_blocks);
        const int acc = it % Cfg::kAcc;
        const uint32_t acc_phase = (it / Cfg::kAcc) & 1;
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        GEMM_STAMP(1, 11);
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          // debug bits 256 / 512 (with bit 1 only): free-running issue without the full-barrier
          // wait / without the per-k-block commit, to separate hand-shake cost from MMA pacing
          if (!(GEMM_DBG(256))) mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (kb == kb0) GEMM_STAMP(1, 12);
          const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint32_t sb = sa + Cfg::kABytes;
          // descriptors as (low, high) words: within a stage only the 14-bit start-address field
          // moves (K-major +32 B, MN-major +2048 B per 16-deep step)
          const uint64_t da0 = A_MN ? make_smem_desc(sa, BK * 128, 1024) : make_smem_desc(sa, 16, 1024);
          const uint64_t db0 = B_MN ? make_smem_desc(sb, BK * 128, 1024) : make_smem_desc(sb, 16, 1024);
          const uint32_t a_lo = (uint32_t)da0, a_hi = (uint32_t)(da0 >> 32);
          const uint32_t b_lo = (uint32_t)db0, b_hi = (uint32_t)(db0 >> 32);
          constexpr uint32_t a_step = A_MN ? (2048 >> 4) : (32 >> 4);
          constexpr uint32_t b_step = B_MN ? (2048 >> 4) : (32 >> 4);
          if (elect_one()) {
            // debug bit 128: issue every MMA twice (timing experiment: per-MMA vs per-k-block cost)
            for (int rep = 0; rep < ((GEMM_DBG(128)) ? 2 : 1); ++rep)
              if (!(GEMM_DBG(2))) {
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) {
                  const uint32_t acc_flag = (kb > kb0 || k > 0 || rep > 0) ? 1u : 0u;
                  if (CTAS == 2)
                    umma_f16_pair_lh(d_tmem, a_lo + k * a_step, a_hi, b_lo + k * b_step, b_hi, idesc, acc_flag);
                  else
                    umma_f16_lh(d_tmem, a_lo + k * a_step, a_hi, b_lo + k * b_step, b_hi, idesc, acc_flag);
                }
              }
            if (!(GEMM_DBG(512))) {
              if (CTAS == 2) umma_commit_pair(&empty_bar[stage]);
              else umma_commit(&empty_bar[stage]);
            }
          }
          __syncwarp();
          if (++stage == nst) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) {
          if (CTAS == 2) umma_commit_pair(&tfull_bar[acc]);
          else umma_commit(&tfull_bar[acc]);
        }
        __syncwarp();
        GEMM_STAMP(1, 13);
      }
#undef STAMP_LANE
#define STAMP_LANE true
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9)
    // and its store warp (warp 10).  The eight epilogue warps turn accumulator slabs into output
    // slabs in shared memory and only ARRIVE on the slab's named barrier; the store warp waits
    // there, issues the TMA store, and hands the slot back (slot_free mbarrier, or the arrival of
    // the next epilogue input loaded into it) once the store has read it.  No epilogue warp ever
    // waits for a TMA issue or for the other warps.
    const int ew = warp - 2;
    const int sub = warp & 3;                 // TMEM sub-partition this warp may read
    const int part = (ew >> 2) & (kColParts - 1);   // which column part of the slab
    const int r = sub * 32 + lane;            // row inside the tile (TMEM lane)
    const bool leader = (ew == 0 && lane == 0);
    const bool f32 = (p.epi == EPI_F32 || p.epi == EPI_ACCUM);
    const int SC = f32 ? 32 : 64;             // accumulator columns per slab
    const int slabs_per_tile = BN / SC;
    const bool two_out = (p.epi == EPI_GELU2);
    const int nslot = two_out ? p.nslot / 2 : (lt ? p.nslot - 1 : p.nslot);
    const int slot_bytes = two_out ? 2 * SLAB_BYTES : SLAB_BYTES;
    const bool aux_in = p.has_aux_in != 0;
    const int my_tiles = my_tiles_all;
    // LayerNorm mode: per row block, ln_tiles * BN/32 fp32 slabs (the stream) then ln_tiles * BN/64
    // bf16 slabs (the normalised operand)
    constexpr int S32 = BN / 32, S64 = BN / 64;
    const uint32_t total_slabs = lt ? (uint32_t)(my_tiles / lt) * (uint32_t)(lt * (S32 + S64))
                                    : (uint32_t)my_tiles * slabs_per_tile;
    constexpr int kSlabBarThreads = kEpiThreads + 32;

    if (warp == kStoreWarp) {
#undef STAMP_LANE
#define STAMP_LANE (lane == 0)
      int dbg_n = 0;
      // global coordinates of slab sg of this CTA
      // LayerNorm mode: slab sg of this CTA -> (row block, is it a bf16 slab of the normalised output,
      // first column)
      constexpr int kLnPer32 = BN / 32, kLnPer64 = BN / 64;
      auto ln_slab = [&](uint32_t sg, bool& is_y, int& c0, int& row) {
        const int per = lt * (kLnPer32 + kLnPer64);
        const int kb = (int)(sg / per), pos = (int)(sg % per);
        row = (unit + kb * units) * (BM * CTAS) + rank * BM;
        if (pos < lt * kLnPer32) { is_y = false; c0 = (pos / kLnPer32) * BN + (pos % kLnPer32) * 32; }
        else { const int q = pos - lt * kLnPer32; is_y = true; c0 = (q / kLnPer64) * BN + (q % kLnPer64) * 64; }
      };
      auto slab_coords = [&](uint32_t sg, int& c0, int& row, int& zi, int& zo) {
        if (lt) { bool is_y; ln_slab(sg, is_y, c0, row); zi = zo = 0; return; }
        const int tile = unit + (int)(sg / slabs_per_tile) * units;
        const int mn = tile % tiles_mn;
        c0 = (mn % p.n_blocks) * BN + (int)(sg % slabs_per_tile) * SC;
        row = (mn / p.n_blocks) * (BM * CTAS) + rank * BM;
        const int z = BATCH ? tile / tiles_per_batch : 0;
        zi = BATCH ? z % p.batch_inner : 0;
        zo = BATCH ? z / p.batch_inner : 0;
      };
      auto issue_aux = [&](uint32_t sg) {  // elected lane: TMA-load the epilogue input of slab sg
        int c0, row, zi, zo;
        slab_coords(sg, c0, row, zi, zo);
        const int slot = sg % nslot;
        mbar_expect_tx(&aux_bar[slot], SLAB_BYTES);
        tma_load_2d(smem_out + slot * slot_bytes, &tma_x, &aux_bar[slot], c0, row);
      };
      // hand the slot of slab t back: either directly, or by loading the next input into it
      auto recycle = [&](uint32_t t) {
        if (lt) {
          // mixed ring: the fp32 stream slabs take an input (the residual), the bf16 slabs do not;
          // every slot hand-over is signalled on slot_free, inputs additionally on aux_bar
          if (t + nslot < total_slabs) {
            bool is_y; int c0, row;
            ln_slab(t + nslot, is_y, c0, row);
            if (!is_y) issue_aux(t + nslot);
          }
          mbar_arrive(&slot_free[t % nslot]);
        } else if (aux_in) {
          if (t + nslot < total_slabs) issue_aux(t + nslot);
        } else {
          mbar_arrive(&slot_free[t % nslot]);
        }
      };
      if (aux_in && elect_one())
        for (uint32_t sg = 0; sg < (uint32_t)nslot && sg < total_slabs; ++sg) issue_aux(sg);
      __syncwarp();
      // stores allowed to be still reading when the next one is issued
      const uint32_t lag = nslot >= 3 ? 1 : 0;
      for (uint32_t sg = 0; sg < total_slabs; ++sg) {
        const int slot = sg % nslot;
        named_bar_sync(1 + slot, kSlabBarThreads);
        GEMM_STAMP(3, 26);
        if (elect_one()) {
          int c0, row, zi, zo;
          slab_coords(sg, c0, row, zi, zo);
          uint8_t* slab = smem_out + slot * slot_bytes;
          if (lt) {
            bool is_y; int c0y, rowy;
            ln_slab(sg, is_y, c0y, rowy);
            tma_store_2d(is_y ? &tma_y : &tma_d, slab, c0, row);
          } else if (p.epi == EPI_ACCUM) {
            tma_reduce_add_2d(&tma_d, slab, c0, row);
          } else if (GEMM_DBG(16)) {   // development: no global stores
          } else if (two_out) {
            tma_store_2d(&tma_x, slab, c0, row);
            tma_store_2d(&tma_d, slab + SLAB_BYTES, c0, row);
          } else if (BATCH) {
            tma_store_4d(&tma_d, slab, c0, row, zi, zo);
          } else {
            tma_store_2d(&tma_d, slab, c0, row);
          }
          tma_store_commit();
          GEMM_STAMP(3, 28);
          if (lag == 0) {
            tma_store_wait_read<0>();
            recycle(sg);
          } else {
            tma_store_wait_read<1>();
            if (sg >= 1) recycle(sg - 1);
          }
          GEMM_STAMP(3, 27);
        }
        __syncwarp();
        if (p.colsum && (sg % slabs_per_tile) == (uint32_t)(slabs_per_tile - 1)) {
          // Last slab of a tile: the barrier above means all sixteen epilogue warps have added
          // their column sums of this tile.  Flush the tile's accumulator (buffer = tile parity) to
          // the global sums and clear it.  The epilogue warps never wait for this: they can only
          // touch this buffer again two tiles later, behind slab slots this warp hands back after
          // the flush (program order here).
          const int itl = (int)(sg / slabs_per_tile);
          const int tile = unit + itl * units;
          const int n0 = ((tile % tiles_mn) % p.n_blocks) * BN;
          float* acc_s = s_colsum + (itl & 1) * BN;
#pragma unroll
          for (int c = lane; c < BN; c += 32) {
            const float vsum = acc_s[c];
            acc_s[c] = 0.f;
            if (n0 + c < p.N) atomicAdd(p.colsum + n0 + c, vsum);
          }
          __syncwarp();
        }
      }
      if (elect_one()) {
        tma_store_wait<0>();
      }
      __syncwarp();
#undef STAMP_LANE
#define STAMP_LANE true
    } else {

    // this warp has read its part of accumulator `a` out of TMEM: one arrival per warp on the
    // (leader's) tempty barrier
    auto release_acc = [&](int a) {
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CTAS == 2) mbar_arrive_leader(&tempty_bar[a]);
        else mbar_arrive(&tempty_bar[a]);
      }
    };
    int m_idx = 0, n_idx = 0, acc = 0, zi = 0, zo = 0;
    float rs = 1.f;
    int dbg_n = leader ? 0 : 1000;
    // bias of the NEXT slab is fetched while this one is stored, so the (L2-latency) loads never
    // sit between the TMEM read and the shared-memory writes
    constexpr int CB = 64 / kColParts;   // columns per thread, bf16 slabs (64 accumulator columns)
    constexpr int CF = 32 / kColParts;   // columns per thread, fp32 slabs (32 accumulator columns)
    float4 bias_pre[CB / 4];
    auto tile_col0 = [&](int it_) { return (((unit + it_ * units) % tiles_mn) % p.n_blocks) * BN; };
    auto prefetch_bias = [&](int it_n, int j_n, int n_idx_n) {  // slab (it_n, j_n) of this CTA
      if (f32 || !p.bias || it_n >= my_tiles) return;
      const int col_n = n_idx_n + j_n * SC + part * CB;
      if (col_n + CB <= p.N) {
#pragma unroll
        for (int q = 0; q < CB / 4; ++q)
          bias_pre[q] = __ldg(reinterpret_cast<const float4*>(p.bias + col_n + 4 * q));
      }
    };
    if (lt) {
      // ------------------------------------------------------------------ fused LayerNorm
      // x_new = acc + bias + residual goes out as the fp32 stream (as in EPI_F32) AND stays in TMEM
      // (written back over the accumulator); row sums are per-thread (thread = row) over this
      // warp's column quarter and are combined across the four warps of a sub-partition in shared
      // memory.  Second pass: the normalised bf16 operand (vit.py:19,26) straight out of TMEM.
      const uint32_t lane_off = (uint32_t)(sub * 32) << 16;
      const int nblk = my_tiles / lt;
      int slot = -1;
      uint32_t slot_phase = 1, aux_par = 0;
      auto next_slot = [&]() { if (++slot == nslot) { slot = 0; slot_phase ^= 1; } };
      for (int kb = 0; kb < nblk; ++kb) {
        const int m_idx = (unit + kb * units) * (BM * CTAS) + rank * BM;
        float rsum = 0.f, rsq = 0.f;
        for (int n = 0; n < lt; ++n) {
          const int itn = kb * lt + n, acc = itn % Cfg::kAcc;
          mbar_wait(&tfull_bar[acc], (itn / Cfg::kAcc) & 1);
          tc_fence_after();
          for (int j = 0; j < S32; ++j) {
            next_slot();
            uint8_t* slab = smem_out + slot * SLAB_BYTES;
            const uint32_t taddr = tmem_base + lane_off + acc * BN + j * 32 + part * CF;
            const int col = n * BN + j * 32 + part * CF;
            mbar_wait(&slot_free[slot], slot_phase);
            uint32_t rr[CF];
            tmem_ld8(taddr, rr);
            tmem_ld_wait();
            float v[CF];
#pragma unroll
            for (int i = 0; i < CF; ++i) v[i] = __uint_as_float(rr[i]);
            epi_affine<CF>(p, v, col, 1.f);
            mbar_wait(&aux_bar[slot], (aux_par >> slot) & 1u);
            aux_par ^= 1u << slot;
#pragma unroll
            for (int g = 0; g < 2; ++g) {
              const float4 a = *reinterpret_cast<const float4*>(slab + slab_off(r, part * 2 + g));
              v[g * 4] += a.x; v[g * 4 + 1] += a.y; v[g * 4 + 2] += a.z; v[g * 4 + 3] += a.w;
            }
#pragma unroll
            for (int i = 0; i < CF; ++i) { rsum += v[i]; rsq += v[i] * v[i]; rr[i] = __float_as_uint(v[i]); }
            tmem_st8(taddr, rr);
#pragma unroll
            for (int g = 0; g < 2; ++g)
              *reinterpret_cast<float4*>(slab + slab_off(r, part * 2 + g)) =
                  make_float4(v[g * 4], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
            fence_proxy_async_smem();
            named_bar_arrive(1 + slot, kSlabBarThreads);
          }
        }
        tmem_st_wait();
        // Single-buffered: a thread reaches this store for row block kb+1 only after its first
        // pass has cycled through every ring slot, i.e. after the store warp has shipped the last
        // second-pass slab of block kb -- which needs every epilogue thread's arrival, made after
        // that thread read the statistics of block kb.
        *reinterpret_cast<float2*>(s_stats + 2 * (part * BM + r)) = make_float2(rsum, rsq);
        tc_fence_before();
        named_bar_sync(10, kEpiThreads);
        tc_fence_after();
        float tot = 0.f, tot2 = 0.f;
#pragma unroll
        for (int q = 0; q < kColParts; ++q) {
          const float2 s2 = *reinterpret_cast<const float2*>(s_stats + 2 * (q * BM + r));
          tot += s2.x; tot2 += s2.y;
        }
        const float inv_n = 1.f / (float)p.N;
        const float mean = tot * inv_n;
        const float rstd = rsqrtf(fmaxf(tot2 * inv_n - mean * mean, 0.f) + p.ln_eps);
        if (part == 0 && (int64_t)m_idx + r < p.M) {
          p.ln_mean[m_idx + r] = mean;
          p.ln_rstd[m_idx + r] = rstd;
        }
        for (int n = 0; n < lt; ++n) {
          const int acc = (kb * lt + n) % Cfg::kAcc;
          uint8_t* slab = nullptr;
          for (int j = 0; j < S32; ++j) {
            if (!(j & 1)) {
              next_slot();
              slab = smem_out + slot * SLAB_BYTES;
              mbar_wait(&slot_free[slot], slot_phase);
            }
            uint32_t rr[CF];
            tmem_ld8(tmem_base + lane_off + acc * BN + j * 32 + part * CF, rr);
            tmem_ld_wait();
            if (j == S32 - 1) release_acc(acc);
            const int col = n * BN + j * 32 + part * CF;
            const float4 g0 = __ldg(reinterpret_cast<const float4*>(p.ln_gamma + col));
            const float4 g1 = __ldg(reinterpret_cast<const float4*>(p.ln_gamma + col + 4));
            const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.ln_beta + col));
            const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.ln_beta + col + 4));
            const float gm[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
            const float bt[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
            float y[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) y[i] = (__uint_as_float(rr[i]) - mean) * rstd * gm[i] + bt[i];
            *reinterpret_cast<uint4*>(slab + slab_off(r, (j & 1) * 4 + part)) =
                make_uint4(pack_bf16(y[0], y[1]), pack_bf16(y[2], y[3]), pack_bf16(y[4], y[5]),
                           pack_bf16(y[6], y[7]));
            if (j & 1) {
              fence_proxy_async_smem();
              named_bar_arrive(1 + slot, kSlabBarThreads);
            }
          }
        }
      }
    } else {
    prefetch_bias(0, 0, tile_col0(0));
    int it = 0, j = -1, slot = -1;
    uint32_t slot_phase = 1;          // parity of the slot ring's current round, inverted
    for (uint32_t sg = 0; sg < total_slabs; ++sg) {
      if (++j == slabs_per_tile) { j = 0; ++it; }
      if (++slot == nslot) { slot = 0; slot_phase ^= 1; }
      if (j == 0) {
        const int tile = unit + it * units;
        const int mn = tile % tiles_mn;
        if (BATCH) {
          const int z = tile / tiles_per_batch;
          zi = z % p.batch_inner;
          zo = z / p.batch_inner;
        }
        m_idx = (mn / p.n_blocks) * (BM * CTAS) + rank * BM;
        n_idx = (mn % p.n_blocks) * BN;
        acc = it % Cfg::kAcc;
        GEMM_STAMP(2, 20);
        mbar_wait(&tfull_bar[acc], (it / Cfg::kAcc) & 1);
        tc_fence_after();
        GEMM_STAMP(2, 21);
        if (p.row_scale) {
          const int64_t row = (int64_t)m_idx + r;
          rs = row < p.M ? p.row_scale[row / p.rows_per_scale] : 0.f;
        }
      }
      uint8_t* slab = smem_out + slot * slot_bytes;
      const uint32_t taddr = tmem_base + ((uint32_t)(sub * 32) << 16) + acc * BN + j * SC + part * (SC / kColParts);
      const int col = n_idx + j * SC + part * (SC / kColParts);
      // the slot's previous store has read it (with an epilogue input, its arrival implies that)
      if (!aux_in) mbar_wait(&slot_free[slot], slot_phase);
      constexpr int CH = 2;   // 16-byte chunks per thread and slab row (CB bf16 or CF floats = 32 B)
      static_assert(CB * 2 == CH * 16 && CF * 4 == CH * 16, "epilogue warp layout");
      if (!f32) {
        uint32_t rr[CB];
        if (GEMM_DBG(8)) {   // development: no TMEM read
#pragma unroll
          for (int i = 0; i < CB; ++i) rr[i] = 0x3f800000u;
        } else {
          tmem_ld16(taddr, rr);
          tmem_ld_wait();
        }
        GEMM_STAMP(2, 23);
        if (j == slabs_per_tile - 1) release_acc(acc);
        float v[CB];
#pragma unroll
        for (int i = 0; i < CB; ++i) v[i] = __uint_as_float(rr[i]);
        if (p.epi == EPI_DGELU) {
          mbar_wait(&aux_bar[slot], slot_phase ^ 1);
#pragma unroll
          for (int g = 0; g < CH; ++g) {
            const uint4 a = *reinterpret_cast<const uint4*>(slab + slab_off(r, part * CH + g));
            const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float2 h = unpack_bf16(w[q]);
              dgelu_mul2(v[g * 8 + 2 * q], v[g * 8 + 2 * q + 1], h.x, h.y);
            }
          }
          if (p.colsum) {
            // Column sums of this slab's 16 columns over the warp's 32 rows (the bias gradient of
            // the Dense whose pre-activation gradient this GEMM produces): a transposing butterfly
            // -- every step halves the columns a lane carries and doubles the rows they cover -- so
            // 16 shuffles instead of 16 x 5, then one shared-memory add per column and warp.
            // Rows >= M contribute 0 (zero-filled operands).
            static_assert(CB == 16, "butterfly below is written for 16 columns per thread");
            float t8[8], t4[4], t2[2], t1;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const float send = (lane & 16) ? v[k] : v[k + 8];
              const float keep = (lane & 16) ? v[k + 8] : v[k];
              t8[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float send = (lane & 8) ? t8[k] : t8[k + 4];
              const float keep = (lane & 8) ? t8[k + 4] : t8[k];
              t4[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
            }
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              const float send = (lane & 4) ? t4[k] : t4[k + 2];
              const float keep = (lane & 4) ? t4[k + 2] : t4[k];
              t2[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
            }
            {
              const float send = (lane & 2) ? t2[0] : t2[1];
              const float keep = (lane & 2) ? t2[1] : t2[0];
              t1 = keep + __shfl_xor_sync(0xffffffffu, send, 2);
            }
            t1 += __shfl_xor_sync(0xffffffffu, t1, 1);
            const int cc = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
            if (!(lane & 1)) atomicAdd(&s_colsum[(it & 1) * BN + j * SC + part * CB + cc], t1);
          }
          st_bf16<CB>(slab, r, part * CH, v);
        } else {
          epi_affine<CB>(p, v, col, rs, bias_pre);
          if (j + 1 < slabs_per_tile) prefetch_bias(it, j + 1, n_idx);
          else prefetch_bias(it + 1, 0, tile_col0(it + 1));
          if (two_out) {
            st_bf16<CB>(slab, r, part * CH, v);  // pre-activation (saved for backward)
            if (!(GEMM_DBG(4))) {
#pragma unroll
              for (int i = 0; i < CB; i += 2) gelu_fast2(v[i], v[i + 1]);
            }
            st_bf16<CB>(slab + SLAB_BYTES, r, part * CH, v);
          } else {
            if (p.act == 1) {
#pragma unroll
              for (int i = 0; i < CB; i += 2) gelu_fast2(v[i], v[i + 1]);
            }
            st_bf16<CB>(slab, r, part * CH, v);
          }
        }
      } else {
        uint32_t rr[CF];
        tmem_ld8(taddr, rr);
        tmem_ld_wait();
        if (j == slabs_per_tile - 1) release_acc(acc);
        float v[CF];
#pragma unroll
        for (int i = 0; i < CF; ++i) v[i] = __uint_as_float(rr[i]);
        if (p.epi == EPI_F32) epi_affine<CF>(p, v, col, rs);
        if (aux_in) {
          mbar_wait(&aux_bar[slot], slot_phase ^ 1);
#pragma unroll
          for (int g = 0; g < CH; ++g) {
            const float4 a = *reinterpret_cast<const float4*>(slab + slab_off(r, part * CH + g));
            v[g * 4] += a.x; v[g * 4 + 1] += a.y; v[g * 4 + 2] += a.z; v[g * 4 + 3] += a.w;
          }
        }
#pragma unroll
        for (int g = 0; g < CH; ++g)
          *reinterpret_cast<float4*>(slab + slab_off(r, part * CH + g)) =
              make_float4(v[g * 4], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
      }
      GEMM_STAMP(2, 24);
      if (!(GEMM_DBG(32))) fence_proxy_async_smem();
      GEMM_STAMP(2, 25);
      named_bar_arrive(1 + slot, kSlabBarThreads);   // never blocks: the store warp does the waiting
    }
    }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CTAS == 2) cluster_sync_all();  // the peer may still read this CTA's smem / signal its barriers
  if (warp == 1) {
    tc_fence_after();
    if (CTAS == 2) tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base);
    else tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

// ---- column sums of a bf16 [M,N] matrix (bias gradients; ff.py:24,27 reverse-mode).
// grid = (N/256 column groups, row splits); a lane owns 8 adjacent columns (one 16-byte load per
// row, a warp reads 512 contiguous bytes), 8 row-lanes per block, smem reduce, one red.add per
// column per block (out is an accumulator).
__global__ void __launch_bounds__(256)
colsum_kernel(const bf16* __restrict__ X, int64_t M, int64_t N, int64_t ldx, float* __restrict__ out) {
  __shared__ float red[8][32][8];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t col = (int64_t)blockIdx.x * 256 + tx * 8;
  float acc[8] = {};
  if (col < N) {
    const int64_t step = (int64_t)gridDim.y * 8;
    int64_t r = (int64_t)blockIdx.y * 8 + ty;
    // two independent loads in flight per thread
    for (; r + step < M; r += 2 * step) {
      const uint4 u0 = *reinterpret_cast<const uint4*>(X + r * ldx + col);
      const uint4 u1 = *reinterpret_cast<const uint4*>(X + (r + step) * ldx + col);
      const uint32_t w0[4] = {u0.x, u0.y, u0.z, u0.w}, w1[4] = {u1.x, u1.y, u1.z, u1.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 a = unpack_bf16(w0[j]), b = unpack_bf16(w1[j]);
        acc[2 * j] += a.x + b.x;
        acc[2 * j + 1] += a.y + b.y;
      }
    }
    if (r < M) {
      const uint4 u0 = *reinterpret_cast<const uint4*>(X + r * ldx + col);
      const uint32_t w0[4] = {u0.x, u0.y, u0.z, u0.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 a = unpack_bf16(w0[j]);
        acc[2 * j] += a.x;
        acc[2 * j + 1] += a.y;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[ty][tx][j] = acc[j];
  __syncthreads();
  // thread (ty, tx) finishes column tx*8 + ty
  if (col < N) {
    float v = 0.f;
#pragma unroll
    for (int y = 0; y < 8; ++y) v += red[y][tx][ty];
    atomicAdd(out + col + ty, v);
  }
}

// ---- two-pass column sums.  The one-kernel version above ends in one red.add per column and
// block: with ~600 row splits that is ~600 same-address atomics per column, serialised in L2
// (N = 384: 56 us for 39 MB).  Here every block stores its partial row; a second kernel folds them.
// A block reads whole rows of a column group (width chosen so that lanes x rows-per-pass fills
// the 256 threads: N = 384 -> one group, 5 rows per pass; 1536 -> three groups of 512, 4 rows), so
// each pass is one contiguous chunk of memory and the blocks of the grid sweep the matrix like a
// grid-stride copy; 8 independent 16-byte loads in flight per thread.
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const bf16* __restrict__ X, int64_t M, int64_t N, int64_t ldx,
                      float* __restrict__ part, int W) {
  __shared__ float red[256 * 8];
  const int64_t c0 = (int64_t)blockIdx.x * W;   // column group of W <= 2048 columns
  const int width = (int)((N - c0) < W ? (N - c0) : W);
  const int lanes = width / 8;          // threads per row
  const int rpp = 256 / lanes;          // rows per pass
  const int rl = threadIdx.x / lanes, cl = threadIdx.x - rl * lanes;
  float acc[8] = {};
  if (rl < rpp) {
    const bf16* px = X + c0 + cl * 8;
    const int64_t step = (int64_t)gridDim.y * rpp;
    int64_t r = (int64_t)blockIdx.y * rpp + rl;
    auto add = [&](const uint4& u) {
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 a = unpack_bf16(w[j]);
        acc[2 * j] += a.x;
        acc[2 * j + 1] += a.y;
      }
    };
    // batches of 8 loads, all issued before the first add (volatile asm keeps ptxas from
    // interleaving them with the adds, which left two or three in flight); rows past the end are
    // predicated off and contribute zeros, so there is no serial tail
    for (; r < M; r += 8 * step) {
      uint4 u[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int64_t ri = r + i * step;
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
            "mov.u32 %0, 0;\n\tmov.u32 %1, 0;\n\tmov.u32 %2, 0;\n\tmov.u32 %3, 0;\n\t"
            "@p ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];\n\t}\n"
            : "=&r"(u[i].x), "=&r"(u[i].y), "=&r"(u[i].z), "=&r"(u[i].w)
            : "l"(px + ri * ldx), "r"((int)(ri < M)));
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) add(u[i]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.x * 8 + j] = acc[j];   // [rl][cl*8 + j], row pitch lanes*8
  __syncthreads();
  for (int c = threadIdx.x; c < width; c += 256) {
    float v = 0.f;
    for (int y = 0; y < rpp; ++y) v += red[y * lanes * 8 + c];
    part[(int64_t)blockIdx.y * N + c0 + c] = v;
  }
}
// out[n] (+)= sum_y part[y][n];  block = 32 columns x 32 row lanes, four loads in flight per thread
// (the grid is only N/32 blocks, so the partials must be walked wide, not deep)
__global__ void __launch_bounds__(1024)
colsum_finish_kernel(const float* __restrict__ part, int64_t Y, int64_t N, float* __restrict__ out,
                     int accumulate) {
  __shared__ float red[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t n = (int64_t)blockIdx.x * 32 + tx;
  float a[4] = {};
  if (n < N) {
    for (int64_t y = ty; y < Y; y += 128) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int64_t yi = y + 32 * i;
        a[i] += yi < Y ? part[yi * N + n] : 0.f;
      }
    }
  }
  red[ty][tx] = (a[0] + a[1]) + (a[2] + a[3]);
  __syncthreads();
  if (ty == 0 && n < N) {
    float v = accumulate ? out[n] : 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) v += red[j][tx];
    out[n] = v;
  }
}

__global__ void fill_kernel(float* p, float v, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    p[i] = v;
}
__global__ void fill2d_kernel(float* p, float v, int64_t rows, int64_t cols, int64_t ld) {
  const int64_t total = rows * cols;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x)
    p[(i / cols) * ld + (i % cols)] = v;
}

static int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

static int pick_bn_batched(int64_t N) { return N <= 64 ? 64 : 128; }

// Development overrides (tile width, CTA pairing, column-sum group width, debug switches) are read
// from the environment ONLY in the -DSAV_GEMM_DEV build (SAV_GEMM_DEV=1 python build.py); the
// production library has no tuning back door.
static int dev_env(const char* name) {
#ifdef SAV_GEMM_DEV
  const char* e = getenv(name);
  return e ? atoi(e) : 0;
#else
  (void)name;
  return 0;
#endif
}

static int pick_bn(int64_t N) {
  static const int forced = dev_env("SAV_GEMM_BN");
  if (forced == 128 || forced == 192 || forced == 256) return forced;
  if (N % 256 == 0) return 256;
  if (N % 192 == 0) return 192;
  if (N > 128 && N % 128 != 0 && N <= 192) return 192;
  return 128;
}

static int plan_split(const sav_gemm_args* a, int bn, int bm) {
  if (a->split_k > 0) return a->split_k;
  // auto (split_k == 0): only for fp32 outputs without epilogue, fill the machine
  const int64_t tiles = ((a->M + bm - 1) / bm) * ((a->N + bn - 1) / bn);
  const int64_t kb = (a->K + BK - 1) / BK;
  const int64_t units = num_sms() / (bm / BM);
  if (!a->d_fp32 || tiles >= units || kb < 16) return 1;
  int64_t s = units / tiles;
  if (s > kb / 8) s = kb / 8;
  return (int)(s < 1 ? 1 : s);
}

template <int BN, bool A_MN, bool B_MN, int CTAS, bool BATCH = false>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td,
                       const CUtensorMap& tx, const CUtensorMap& ty, const GemmParams& p,
                       cudaStream_t st) {
  using Cfg = GemmCfg<BN, B_MN, CTAS>;
  auto kern = gemm_kernel<BN, A_MN, B_MN, CTAS, BATCH>;
  static bool attr_set = false;  // benign race: idempotent
  if (!attr_set) {
    cudaError_t e =
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmem);
    if (e != cudaSuccess) return set_error(SAV_ERR_CUDA, "gemm smem attr: %s", cudaGetErrorString(e));
    attr_set = true;
  }
  // (fused LayerNorm: a work unit is a whole row block = ln_tiles N-tiles)
  const int tiles = p.ln_tiles ? p.m_blocks : p.m_blocks * p.n_blocks * p.split_k * (BATCH ? p.nbatch : 1);
  const int units = num_sms() / CTAS;
  const int grid = (tiles < units ? tiles : units) * CTAS;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::kSmem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CTAS;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  // operand stages vs slab slots (see GemmParams::stages)
  GemmParams q = p;
  q.stages = Cfg::kStages;
  q.nslot = 4;
  {
    static const int f_stages = dev_env("SAV_GEMM_STAGES"), f_slots = dev_env("SAV_GEMM_SLOTS");
    const bool streaming_epi = p.has_aux_in || p.epi == EPI_GELU2 || p.ln_tiles;
    // measured (profiles/r02aa_split_sweep.txt): with K = 384 one stage less and two slabs more is
    // 2-9 % faster for FC1 / out-projection / FC2 dX; with K = 1536 the pipeline needs its stages
    if (streaming_epi && p.k_per_split <= 8 && Cfg::kStages >= 5) q.stages = Cfg::kStages - 1;
    if (streaming_epi && f_stages >= 3 && f_stages <= Cfg::kStages) q.stages = f_stages;
    const int room = ((Cfg::kStages - q.stages) * Cfg::kStageBytes + kOutBytes) / SLAB_BYTES;
    // LayerNorm mode: at most five ring slots + the statistics slot (the single-buffered statistics
    // rely on a row block spanning more slabs than the ring holds)
    const int cap = p.ln_tiles ? 6 : 8;
    int ns = room < cap ? room : cap;
    if (f_slots >= 2 && f_slots < ns) ns = f_slots;
    if (p.epi == EPI_GELU2) ns &= ~1;
    q.nslot = q.stages == Cfg::kStages ? 4 : ns;
  }
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, tb, td, tx, ty, q);
  if (e != cudaSuccess) return set_error(SAV_ERR_CUDA, "sav_gemm launch: %s", cudaGetErrorString(e));
  count_launch();
  return check_launch("sav_gemm");
}

// CTA pairs pay off when the pair tile (256 rows) is mostly full; SAV_GEMM_CTAS=1|2 overrides.
static int pick_ctas(int64_t M) {
  static const int forced = dev_env("SAV_GEMM_CTAS");
  if (forced == 1 || forced == 2) return forced;
  const int64_t pair_rows = ((M + 2 * BM - 1) / (2 * BM)) * (2 * BM);
  const int64_t single_rows = ((M + BM - 1) / BM) * BM;
  return pair_rows * 8 <= single_rows * 9 ? 2 : 1;  // <= 12.5 % extra padded rows
}

}  // namespace sav

using namespace sav;

// N-tiles per row when the LayerNorm that follows can be fused into the fp32-stream epilogue: the
// row must be complete within the accumulator ring (4 x 128 columns; 192 as one tile).
static int ln_plan(int64_t N, int* bn) {
  if (N > 0 && N % 128 == 0 && N <= 512) { *bn = 128; return (int)(N / 128); }
  if (N == 192) { *bn = 192; return 1; }
  *bn = 0;
  return 0;
}

extern "C" int32_t sav_gemm_ln_tiles(int64_t N) {
  int bn;
  return ln_plan(N, &bn);
}

extern "C" int64_t sav_gemm_workspace_bytes(const sav_gemm_args* a) {
  (void)a;
  return 0;  // split-K partials are combined with TMA reduce-add directly in D
}

extern "C" int sav_gemm(const sav_gemm_args* a, void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!a || a->M <= 0 || a->N <= 0 || a->K <= 0)
    return set_error(SAV_ERR_SHAPE, "sav_gemm: M,N,K must be positive");
  if (a->N % 8 || a->lda % 8 || a->ldb % 8 || a->ldd % (a->d_fp32 ? 4 : 8) ||
      (a->residual && a->ldr % 4))
    return set_error(SAV_ERR_ALIGN, "sav_gemm: N, lda, ldb, ldd, ldr must be multiples of 8");
  if (!aligned16(a->A) || !aligned16(a->B) || !aligned16(a->D) || !aligned16(a->bias) ||
      !aligned16(a->aux_in) || !aligned16(a->aux_out) || !aligned16(a->residual) ||
      !aligned16(a->col_scale))
    return set_error(SAV_ERR_ALIGN, "sav_gemm: pointers must be 16 B aligned");
  int bn = pick_bn(a->N);
  if (a->ln_out && !ln_plan(a->N, &bn))
    return set_error(SAV_ERR_ARG, "sav_gemm: fused LayerNorm needs N = 192 or a multiple of 128 up to 512 "
                                  "(sav_gemm_ln_tiles)");
  const int ctas = pick_ctas(a->M);
  const int bm = BM * ctas;
  const int split = plan_split(a, bn, bm);
  const bool accum = a->accumulate || split > 1;
  const bool any_affine = a->bias || a->col_scale || a->row_scale;
  if (accum && (!a->d_fp32 || any_affine || a->act || a->aux_in || a->aux_out || a->residual))
    return set_error(SAV_ERR_ARG, "sav_gemm: accumulate / split_k need d_fp32 and no epilogue");
  if (a->d_fp32 && (a->act || a->aux_in || a->aux_out))
    return set_error(SAV_ERR_ARG, "sav_gemm: activation epilogues need a bf16 output");
  if (a->residual && !a->d_fp32)
    return set_error(SAV_ERR_ARG, "sav_gemm: residual (fp32 residual stream) needs d_fp32");
  if (a->aux_in && (a->act || a->aux_out || any_affine))
    return set_error(SAV_ERR_ARG, "sav_gemm: aux_in (dGELU) cannot be combined with bias/act/scales");
  if (a->aux_out && a->act != 1)
    return set_error(SAV_ERR_ARG, "sav_gemm: aux_out is the pre-activation of act=1");

  GemmParams p;
  p.M = a->M; p.N = a->N; p.K = a->K;
  p.m_blocks = (int)((a->M + bm - 1) / bm);
  p.n_blocks = (int)((a->N + bn - 1) / bn);
  p.k_blocks = (int)((a->K + BK - 1) / BK);
  p.k_per_split = (p.k_blocks + split - 1) / split;
  p.split_k = (p.k_blocks + p.k_per_split - 1) / p.k_per_split;
  p.epi = accum ? EPI_ACCUM : a->d_fp32 ? EPI_F32 : a->aux_in ? EPI_DGELU : a->aux_out ? EPI_GELU2 : EPI_PLAIN;
  p.act = a->act;
  p.has_aux_in = (a->aux_in || a->residual) ? 1 : 0;
  p.bias = a->bias; p.col_scale = a->col_scale; p.row_scale = a->row_scale;
  p.rows_per_scale = a->rows_per_scale > 0 ? a->rows_per_scale : 1;
  p.colsum = a->dgelu_colsum;
  if (p.colsum && p.epi != EPI_DGELU)
    return set_error(SAV_ERR_ARG, "sav_gemm: dgelu_colsum needs the aux_in (dGELU) epilogue");
  p.ln_tiles = 0;
  p.ln_gamma = p.ln_beta = nullptr; p.ln_mean = p.ln_rstd = nullptr; p.ln_eps = 0.f;
  if (a->ln_out) {
    const int ltl = sav_gemm_ln_tiles(a->N);
    if (!ltl || p.epi != EPI_F32 || !a->residual || a->col_scale || a->row_scale || split > 1)
      return set_error(SAV_ERR_ARG, "sav_gemm: fused LayerNorm needs the fp32 residual epilogue without "
                                    "scales, unsplit");
    if (!a->ln_gamma || !a->ln_beta || !a->ln_mean || !a->ln_rstd || a->ldy % 8 || !aligned16(a->ln_out) ||
        !aligned16(a->ln_gamma) || !aligned16(a->ln_beta))
      return set_error(SAV_ERR_ARG, "sav_gemm: fused LayerNorm needs gamma, beta, mean, rstd, ldy %% 8");
    p.ln_tiles = ltl;
    p.ln_gamma = a->ln_gamma; p.ln_beta = a->ln_beta; p.ln_mean = a->ln_mean; p.ln_rstd = a->ln_rstd;
    p.ln_eps = a->ln_eps;
  }
  {
    static const int dbg = dev_env("SAV_GEMM_DEBUG");
    p.debug = dbg;
  }
  p.batch_inner = 1; p.nbatch = 1;
  p.dbg = reinterpret_cast<long long*>(sav_debug_buffer());

  CUtensorMap ta, tb, td, tx;
  int rc;
  {
    uint64_t dims[2], str[2];
    uint32_t box[2];
    if (!a->a_mn) { dims[0] = a->K; dims[1] = a->M; box[0] = BK; box[1] = BM; }
    else          { dims[0] = a->M; dims[1] = a->K; box[0] = 64; box[1] = BK; }
    str[0] = 0; str[1] = (uint64_t)a->lda * 2;
    if ((rc = make_tmap_bf16(&ta, a->A, 2, dims, str, box))) return rc;
    if (!a->b_mn) { dims[0] = a->K; dims[1] = a->N; box[0] = BK; box[1] = bn / ctas; }
    else          { dims[0] = a->N; dims[1] = a->K; box[0] = 64; box[1] = BK; }
    str[1] = (uint64_t)a->ldb * 2;
    if ((rc = make_tmap_bf16(&tb, a->B, 2, dims, str, box))) return rc;
  }
  const int es = a->d_fp32 ? 4 : 2;
  if ((rc = make_tmap_2d(&td, a->D, es, a->N, a->M, a->ldd, 128 / es, BM))) return rc;
  tx = td;
  if (a->aux_out) rc = make_tmap_2d(&tx, a->aux_out, 2, a->N, a->M, a->ldd, 64, BM);
  else if (a->aux_in) rc = make_tmap_2d(&tx, a->aux_in, 2, a->N, a->M, a->ldd, 64, BM);
  else if (a->residual) rc = make_tmap_2d(&tx, a->residual, 4, a->N, a->M, a->ldr, 32, BM);
  if (rc) return rc;
  CUtensorMap ty = td;
  if (p.ln_tiles && (rc = make_tmap_2d(&ty, a->ln_out, 2, a->N, a->M, a->ldy, 64, BM))) return rc;

  if (split > 1 && !a->accumulate) {  // reduce-add target must start from zero
    const int64_t total = a->M * a->N;
    int blocks = (int)((total + 255) / 256 < num_sms() * 8 ? (total + 255) / 256 : num_sms() * 8);
    fill2d_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<float*>(a->D), 0.f, a->M, a->N, a->ldd);
    count_launch();
  }
#define SAV_GEMM_CASE(BNv, AM, BMj)                                                  \
  if (bn == BNv && (a->a_mn != 0) == AM && (a->b_mn != 0) == BMj)                     \
    rc = ctas == 2 ? launch_gemm<BNv, AM, BMj, 2>(ta, tb, td, tx, ty, p, st)          \
                   : launch_gemm<BNv, AM, BMj, 1>(ta, tb, td, tx, ty, p, st);
  rc = SAV_ERR_ARG;
  SAV_GEMM_CASE(128, false, false) SAV_GEMM_CASE(128, false, true)
  SAV_GEMM_CASE(128, true, false)  SAV_GEMM_CASE(128, true, true)
  SAV_GEMM_CASE(192, false, false) SAV_GEMM_CASE(192, false, true)
  SAV_GEMM_CASE(192, true, false)  SAV_GEMM_CASE(192, true, true)
  SAV_GEMM_CASE(256, false, false) SAV_GEMM_CASE(256, false, true)
  SAV_GEMM_CASE(256, true, false)  SAV_GEMM_CASE(256, true, true)
#undef SAV_GEMM_CASE
  return rc;
}


// ---- batched GEMM over (outer, inner) batch dims with independent strides: the materialised-
// logits attention path (talking heads, class attention, 2-D relative logits, head dims != 64).
extern "C" int sav_bgemm(const sav_bgemm_args* a, void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!a || a->M <= 0 || a->N <= 0 || a->K <= 0 || a->batch_outer <= 0 || a->batch_inner <= 0)
    return set_error(SAV_ERR_SHAPE, "sav_bgemm: M,N,K and batch dims must be positive");
  if (a->batch_outer * a->batch_inner > (1 << 22))
    return set_error(SAV_ERR_SHAPE, "sav_bgemm: too many batches");
  const int bn = pick_bn_batched(a->N);
  GemmParams p = {};
  p.M = a->M; p.N = a->N; p.K = a->K;
  p.m_blocks = (int)((a->M + BM - 1) / BM);
  p.n_blocks = (int)((a->N + bn - 1) / bn);
  p.k_blocks = (int)((a->K + BK - 1) / BK);
  p.k_per_split = p.k_blocks;
  p.split_k = 1;
  p.epi = a->d_fp32 ? EPI_F32 : EPI_PLAIN;
  p.rows_per_scale = 1;
  p.dbg = nullptr;
  p.batch_inner = (int)a->batch_inner;
  p.nbatch = (int)(a->batch_inner * a->batch_outer);
  CUtensorMap ta, tb, td;
  int rc;
  uint64_t dims[4], str[4];
  uint32_t box[4] = {0, 0, 1, 1};
  dims[2] = a->batch_inner; dims[3] = a->batch_outer;
  str[0] = 0;
  if (!a->a_mn) { dims[0] = a->K; dims[1] = a->M; box[0] = BK; box[1] = BM; }
  else          { dims[0] = a->M; dims[1] = a->K; box[0] = 64; box[1] = BK; }
  str[1] = (uint64_t)a->lda * 2; str[2] = (uint64_t)a->a_stride_inner * 2; str[3] = (uint64_t)a->a_stride_outer * 2;
  if ((rc = make_tmap_nd(&ta, a->A, 2, 4, dims, str, box))) return rc;
  if (!a->b_mn) { dims[0] = a->K; dims[1] = a->N; box[0] = BK; box[1] = bn; }
  else          { dims[0] = a->N; dims[1] = a->K; box[0] = 64; box[1] = BK; }
  str[1] = (uint64_t)a->ldb * 2; str[2] = (uint64_t)a->b_stride_inner * 2; str[3] = (uint64_t)a->b_stride_outer * 2;
  if ((rc = make_tmap_nd(&tb, a->B, 2, 4, dims, str, box))) return rc;
  const int es = a->d_fp32 ? 4 : 2;
  dims[0] = a->N; dims[1] = a->M; box[0] = 128 / es; box[1] = BM;
  str[1] = (uint64_t)a->ldd * es; str[2] = (uint64_t)a->d_stride_inner * es; str[3] = (uint64_t)a->d_stride_outer * es;
  if ((rc = make_tmap_nd(&td, a->D, es, 4, dims, str, box))) return rc;
#define SAV_BGEMM_CASE(BNv, AM, BMj)                                  \
  if (bn == BNv && (a->a_mn != 0) == AM && (a->b_mn != 0) == BMj)     \
    rc = launch_gemm<BNv, AM, BMj, 1, true>(ta, tb, td, td, td, p, st);
  rc = SAV_ERR_ARG;
  SAV_BGEMM_CASE(64, false, false) SAV_BGEMM_CASE(64, false, true)
  SAV_BGEMM_CASE(64, true, false)  SAV_BGEMM_CASE(64, true, true)
  SAV_BGEMM_CASE(128, false, false) SAV_BGEMM_CASE(128, false, true)
  SAV_BGEMM_CASE(128, true, false)  SAV_BGEMM_CASE(128, true, true)
#undef SAV_BGEMM_CASE
  return rc;
}

extern "C" int sav_fill_f32(float* ptr, float v, int64_t n, void* stream) {
  if (n <= 0) return SAV_OK;
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  fill_kernel<<<blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(ptr, v, n);
  count_launch();
  return check_launch("sav_fill_f32");
}

// column groups: N = gx * W with W % 8 == 0 when possible (else 2048-wide groups with a narrower
// last one), gx picked to keep the most threads busy.  Power-of-two widths are avoided: blocks of
// the same rows then start 256 B / 512 B / 1 KB apart and camp on the same HBM channels (measured,
// N = 1536: W = 512 / 256 / 128 -> 42-45 us, W = 384 / 192 -> 33 us; profiles/r01j_colsum_w.log).
static void colsum_plan(int64_t N, int64_t* gx_out, int64_t* W_out) {
  const int64_t g0 = (N + 2047) / 2048;
  int64_t best_g = g0, best_w = N < 2048 ? N : 2048, best_active = 0;
  for (int64_t g = g0; g < g0 + 8; ++g) {
    if (N % (g * 8) != 0) continue;
    const int64_t w = N / g;
    if ((w & (w - 1)) == 0) continue;
    const int64_t lanes = w / 8, active = lanes * (256 / lanes);
    if (active > best_active) { best_active = active; best_g = g; best_w = w; }
  }
  static const int forced = dev_env("SAV_COLSUM_W");   // development override: group width
  if (forced >= 8 && forced <= 2048 && forced % 8 == 0 && N % forced == 0) {
    best_g = N / forced;
    best_w = forced;
  }
  *gx_out = best_g;
  *W_out = best_w;
}
static int64_t colsum_splits(int64_t M, int64_t N) {
  int64_t gx, W;
  colsum_plan(N, &gx, &W);
  int64_t want = (num_sms() * 8 + gx - 1) / gx;
  const int64_t rpp = 256 / (W / 8);
  const int64_t maxy = (M + rpp * 16 - 1) / (rpp * 16);   // at least two full load batches per thread
  if (want > maxy) want = maxy;
  return want < 1 ? 1 : want;
}
extern "C" int64_t sav_colsum_workspace_bytes(int64_t M, int64_t N) {
  if (M <= 0 || N <= 0) return 0;
  return colsum_splits(M, N) * N * (int64_t)sizeof(float);
}
extern "C" int sav_colsum_bf16_ws(const void* X, int64_t M, int64_t N, int64_t ldx, float* out,
                                  int32_t accumulate, void* workspace, int64_t workspace_bytes,
                                  void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (M <= 0 || N <= 0 || N % 8 || ldx % 8 || !aligned16(X))
    return set_error(SAV_ERR_SHAPE, "sav_colsum: need N, ldx multiples of 8 and a 16 B aligned base");
  if (!workspace || !aligned16(workspace) || workspace_bytes < sav_colsum_workspace_bytes(M, N))
    return set_error(SAV_ERR_ARG, "sav_colsum_bf16_ws: workspace missing, misaligned or too small");
  const int64_t Y = colsum_splits(M, N);
  int64_t gx, W;
  colsum_plan(N, &gx, &W);
  dim3 grid((unsigned)gx, (unsigned)Y);
  float* part = reinterpret_cast<float*>(workspace);
  colsum_partial_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const bf16*>(X), M, N, ldx, part, (int)W);
  count_launch();
  int rc = check_launch("sav_colsum_bf16_ws");
  if (rc) return rc;
  colsum_finish_kernel<<<(unsigned)((N + 31) / 32), 1024, 0, st>>>(part, Y, N, out, accumulate);
  count_launch();
  return check_launch("sav_colsum_bf16_ws");
}

extern "C" int sav_colsum_bf16(const void* X, int64_t M, int64_t N, int64_t ldx, float* out,
                               int32_t accumulate, void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (M <= 0 || N <= 0 || N % 8 || ldx % 8 || !aligned16(X))
    return set_error(SAV_ERR_SHAPE, "sav_colsum: need N, ldx multiples of 8 and a 16 B aligned base");
  if (!accumulate) {
    int rc = sav_fill_f32(out, 0.f, N, stream);
    if (rc) return rc;
  }
  dim3 grid((unsigned)((N + 255) / 256), 1);
  int64_t want = (num_sms() * 8 + grid.x - 1) / grid.x;
  int64_t maxy = (M + 7) / 8;
  grid.y = (unsigned)(want < maxy ? (want < 1 ? 1 : want) : maxy);
  colsum_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const bf16*>(X), M, N, ldx, out);
  count_launch();
  return check_launch("sav_colsum_bf16");
}

// ---- Column sums without any allocation or caller-provided workspace.
// No counterpart in the reference; added for callers that
// cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid
// it).  The row-split partial sums of sav_colsum_bf16_ws go to a file-scope device array instead of a
// caller workspace.  Every GPU that loads the library holds the 4 MiB until exit.  All calls share
// the array: two calls that overlap on the GPU (different streams, or graph branches) race on it.
// The array keeps the previous call's contents, which is harmless here: the partial kernel writes
// every word [0, splits * N) before the finishing kernel reads it, so nothing has to start at zero.
constexpr int64_t kColsumStaticBytes = 4ll << 20;
static __device__ __align__(256) unsigned char g_colsum_static_scratch[kColsumStaticBytes];

// In host code `g_colsum_static_scratch` is a host-side stand-in for the symbol, not a device
// address: the address a kernel can use comes from cudaGetSymbolAddress, taken once per device with
// that device current (set-up, cached; not an allocation).
static int colsum_static_scratch(float** out_ptr) {
  constexpr int kMaxDevices = 64;
  static void* cached[kMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices)
    return set_error(SAV_ERR_CUDA, "sav_colsum_bf16_static: cudaGetDevice: %s (device %d)",
                     cudaGetErrorString(e), dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!cached[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_colsum_static_scratch);
    if (e != cudaSuccess || !p)
      return set_error(SAV_ERR_CUDA, "sav_colsum_bf16_static: cudaGetSymbolAddress: %s",
                       cudaGetErrorString(e));
    cached[dev] = p;
  }
  *out_ptr = reinterpret_cast<float*>(cached[dev]);
  return SAV_OK;
}

extern "C" int64_t sav_colsum_static_max_bytes(void) { return kColsumStaticBytes; }

extern "C" int sav_colsum_bf16_static(const void* X, int64_t M, int64_t N, int64_t ldx, float* out,
                                      int32_t accumulate, void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!X || !out || M <= 0 || N <= 0 || N % 8 || ldx % 8 || !aligned16(X))
    return set_error(SAV_ERR_SHAPE, "sav_colsum: need N, ldx multiples of 8 and a 16 B aligned base");
  const int64_t need = sav_colsum_workspace_bytes(M, N);
  if (need > kColsumStaticBytes)
    return set_error(SAV_ERR_SHAPE,
                     "sav_colsum_bf16_static: M=%lld, N=%lld needs %lld bytes of scratch, the built-in "
                     "array holds %lld; use sav_colsum_bf16_ws with a caller workspace",
                     (long long)M, (long long)N, (long long)need, (long long)kColsumStaticBytes);
  float* part = nullptr;
  if (int rc = colsum_static_scratch(&part)) return rc;
  const int64_t Y = colsum_splits(M, N);
  int64_t gx, W;
  colsum_plan(N, &gx, &W);
  dim3 grid((unsigned)gx, (unsigned)Y);
  colsum_partial_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const bf16*>(X), M, N, ldx, part, (int)W);
  count_launch();
  if (int rc = check_launch("sav_colsum_bf16_static")) return rc;
  colsum_finish_kernel<<<(unsigned)((N + 31) / 32), 1024, 0, st>>>(part, Y, N, out, accumulate);
  count_launch();
  return check_launch("sav_colsum_bf16_static");
}
