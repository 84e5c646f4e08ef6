// Fused attention backward for sm_100a, any Nq / Nk (reverse mode of attention.py:39-58):
//   S = q k^T * scale, P = exp(S - lse), dP = dO v^T, dS = P o (dP - delta) * scale,
//   dV = P^T dO, dK = dS^T q, dQ = dS k.
// One persistent CTA owns a work item (b, h, query group): the group's one or two 128-row query tiles
// (Q, dO) stay resident in shared memory, ALL key tiles stream through a 2-stage TMA ring, and
// S / P / dP / dS never leave the SM.  dQ accumulates in TMEM across the key tiles and is stored once;
// dK / dV accumulate in TMEM across the group's query tiles.  Up to 256 queries there is one group per
// (b, h) and dK / dV are plain stores; longer sequences (ViT-B/16 384^2: 577 tokens = groups of 256,
// 256, 65; CvT: 3136 x 784) have G = ceil(Nq / 256) groups per (b, h) whose dK / dV partial tiles are
// combined by the TMA unit itself: cp.reduce.async.bulk.tensor .add of fp32 half tiles into a
// zero-filled fp32 workspace [2][B, Nk, H * 64], converted to bf16 once at the end.  (Adding the
// partials in bf16 straight into the gradients was tried first: 0.1 ms faster at 577 tokens, but the
// order-dependent roundings of up to 13 partials moved CvT's worst parity tensor by +-0.5e-2 from
// run to run.)  In fp32 the arrival order changes the sum by ~1e-7; the bf16 rounding happens once.
//
// The transposed problem is computed (S^T = K Q^T, dP^T = V dO^T: TMEM lane = key row), so P^T / dS^T
// rows come out K-major for dV += P^T dO, dK += dS^T Q, and the same dS^T buffer read MN-major is
// the A operand of dQ += dS K.  The unit of pipelining is a CHUNK = (key tile, 64 query columns):
// S^T / dP^T chunks are double-buffered in TMEM, so the tensor pipe works on chunk n+1 / n+2 while
// the math warps are on chunk n, and nothing ever waits for an epilogue.
//
// 640 threads, one CTA per SM, every role free-running behind mbarriers:
//   warp 0       TMA producer (Q_i, dO_i per item; K_j, V_j ring)
//   warp 1       MMA issuer A: S^T and dP^T chunks (runs up to two chunks ahead of the math)
//   warp 2       MMA issuer B: dV, dK per chunk, dQ per query tile -- every gradient MMA is issued
//                by ONE thread, so a single tcgen05.commit covers all readers of P^T / dS^T
//   warp 3       TMEM allocator; per-item statistics (-lse*log2e, -delta) prefetched into registers
//                and handed to the math warps through shared memory
//   warps 4-11   math: thread = (key row, 32 of the chunk's 64 query columns):
//                p = exp2(s*c - lse), ds = p (dp - delta), packed f32x2 math, bf16 P^T / dS^T into
//                128B-swizzled shared memory
//   warps 12-15  epilogue: dV_j / dK_j after each key tile, dQ_0 / dQ_1 after each item, out of
//                TMEM -> bf16 -> own staging buffers -> TMA store.  The math warps never drain.
//   warps 16-19  delta = rowsum(dO o (O + O_lo)) of the NEXT item (no separate kernel, no second
//                pass over dO): straight from global, one item ahead, handed to warp 3 through the
//                global workspace.  Registers are redistributed with setmaxnreg: 56 for the
//                producer / issuer / statistics warpgroup, 80 for the epilogue and delta warpgroups,
//                128 for the two math warpgroups (640 x 96 at launch).
// TMEM (512 columns): S^T[2] 64+64 | dP^T[2] 64+64 | dV 64 | dK 64 | dQ_0 64 | dQ_1 64.
// The softmax scale of dK / dQ is applied once at read-out.
#include "common.cuh"
#include "host.h"

namespace sav {

constexpr int BT = 128;                 // tile edge (keys per tile, queries per tile)
constexpr int BD = 64;                  // head dim
constexpr int BC = 64;                  // query columns per chunk
constexpr int BT_BYTES = BT * BD * 2;   // 16 KB: one [128 x 64] bf16 tile, 128-byte rows
constexpr int kBwdThreads = 640;
constexpr float B_LOG2E = 1.4426950408889634f;

struct AttnBwd2Params {
  const float* lse;
  float* delta;            // workspace [B, H, Nq]: rows -> columns hand-over inside warp 3
  const bf16* o;           // forward output, its bf16 rounding remainder (may be NULL) and dO:
  const bf16* o_lo;        //   [B, Nq, H, 64] views, element strides o_bs / o_rs
  const bf16* d_o;
  long long o_bs, o_rs;
  int B, H, Nq, Nk;
  int n_items;   // B * H * G
  int nkt;       // key tiles (any number)
  int G;         // query groups of <= 256 rows per (b, h); G > 1: dK / dV tiles are reduce-added
  int nqt;       // query tiles of the only group when G == 1 (1 or 2)
  float scale, scale_log2;
  long long* dbg;
};

// Development timeline (compiled in only with -DSAV_ATTN_DEV: SAV_ATTN_DEV=1 python build.py):
// (event, clock64) pairs of CTA 0, 256 slots per role, read by tests/tools/attn_timeline.py --bwd.
#ifdef SAV_ATTN_DEV
#define BWD_STAMP(role, ev)                                                       \
  do {                                                                            \
    if (p.dbg && blockIdx.x == 0 && dbg_on && dbg_n < 255) {                      \
      p.dbg[(role) * 512 + 2 * dbg_n] = (ev);                                     \
      p.dbg[(role) * 512 + 2 * dbg_n + 1] = clock64();                            \
      ++dbg_n;                                                                    \
    }                                                                             \
  } while (0)
#else
#define BWD_STAMP(role, ev) do { } while (0)
#endif

// byte offset of 16-byte group `g16` (0..7) of row `r` inside a [rows x 128 B] SW128 tile
__device__ __forceinline__ uint32_t bsw128(int r, int g16) {
  return (uint32_t)(r * 128 + ((g16 ^ (r & 7)) << 4));
}

// shared-memory map (bytes from the 1024-aligned base)
constexpr int kOffQ = 0;                        // [i] 2 x 16 KB
constexpr int kOffdO = 2 * BT_BYTES;            // [i] 2 x 16 KB
constexpr int kOffKV = 4 * BT_BYTES;            // [stage] { K 16 KB, V 16 KB }
constexpr int kOffPt = 8 * BT_BYTES;            // P^T  [cc] 2 x [128 keys x 64 q]
constexpr int kOffdSt = 10 * BT_BYTES;          // dS^T [cc] 2 x [128 keys x 64 q] (contiguous: one
                                                //      MN-major [128 keys x 128 q] operand for dQ)
constexpr int kOffStage = 12 * BT_BYTES;        // epilogue staging, 2 x 16 KB
constexpr int kOffStats = 14 * BT_BYTES;        // float[256] -lse*log2e | float[256] -delta
constexpr int kOffBars = kOffStats + 2 * 256 * 4;
constexpr int kNumBars = 28;   // 25 used
constexpr int kBwd2Smem = kOffBars + kNumBars * 8 + 16;
static_assert(kBwd2Smem <= 232448, "attention backward exceeds the 227 KB shared-memory limit");

__device__ __forceinline__ int round16(int x) { return (x + 15) & ~15; }

// work item -> (b, h, first query row, query rows, query tiles); the groups of one (b, h) are
// neighbours in the item order, so they run at the same time and share K / V through L2
struct BwdItem {
  int b, h, bh, q0, nq, nqt;
};
template <bool kGroups>
__device__ __forceinline__ BwdItem bwd_item(const AttnBwd2Params& p, int item) {
  BwdItem it;
  if (!kGroups) {   // one group per (b, h): nothing to decode (the DeiT / CaiT shapes)
    it.bh = item;
    it.h = item % p.H;
    it.b = item / p.H;
    it.q0 = 0;
    it.nq = p.Nq;
    it.nqt = p.nqt;      // (a kernel parameter: a value computed here is hoisted and spilled)
    return it;
  }
  const int qg = item % p.G;
  it.bh = item / p.G;
  it.h = it.bh % p.H;
  it.b = it.bh / p.H;
  it.q0 = qg * 2 * BT;
  it.nq = min(2 * BT, p.Nq - it.q0);
  it.nqt = (it.nq + BT - 1) / BT;
  return it;
}

__device__ __forceinline__ void tma_reduce_add_3d(const CUtensorMap* m, const void* src, int c0, int c1,
                                                  int c2) {
  asm volatile(
      "cp.reduce.async.bulk.tensor.3d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
          reinterpret_cast<uint64_t>(m)),
      "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

__device__ __forceinline__ void pack_rows_to_stage(uint8_t* stage, int r, const uint32_t (&pk)[32]) {
#pragma unroll
  for (int g = 0; g < 8; ++g)
    *reinterpret_cast<uint4*>(stage + bsw128(r, g)) =
        make_uint4(pk[g * 4], pk[g * 4 + 1], pk[g * 4 + 2], pk[g * 4 + 3]);
}

// 64 fp32 TMEM columns of this thread's lane -> 32 packed bf16x2 registers, scaled by f
__device__ __forceinline__ void tmem_row64_to_bf16(uint32_t taddr, float f, uint32_t (&pk)[32]) {
  uint32_t r0[32];
  tmem_ld32(taddr, r0);
  tmem_ld_wait();
#pragma unroll
  for (int e = 0; e < 16; ++e)
    pk[e] = pack_bf16(__uint_as_float(r0[2 * e]) * f, __uint_as_float(r0[2 * e + 1]) * f);
  tmem_ld32(taddr + 32, r0);
  tmem_ld_wait();
#pragma unroll
  for (int e = 0; e < 16; ++e)
    pk[16 + e] = pack_bf16(__uint_as_float(r0[2 * e]) * f, __uint_as_float(r0[2 * e + 1]) * f);
}

template <bool kGroups>
__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd3_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                 const __grid_constant__ CUtensorMap tv, const __grid_constant__ CUtensorMap tdo,
                 const __grid_constant__ CUtensorMap tdq, const __grid_constant__ CUtensorMap tdk,
                 const __grid_constant__ CUtensorMap tdv, const AttnBwd2Params p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sQ = smem + kOffQ;
  uint8_t* sdO = smem + kOffdO;
  uint8_t* sKV = smem + kOffKV;
  uint8_t* sPt = smem + kOffPt;
  uint8_t* sdSt = smem + kOffdSt;
  uint8_t* sStage = smem + kOffStage;
  float* s_nl2 = reinterpret_cast<float*>(smem + kOffStats);   // [256] -lse * log2(e)  (-inf beyond Nq)
  float* s_nd = s_nl2 + 2 * BT;                                 // [256] -delta
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBars);
  uint64_t* qdo_full = bars;          // [i]      TMA
  uint64_t* qdo_empty = bars + 2;     // [i]      2 commits (issuer A, issuer B)
  uint64_t* kv_full = bars + 4;       // [stage]  TMA
  uint64_t* kv_empty = bars + 6;      // [stage]  2 commits
  uint64_t* s_full = bars + 8;        // [buf]    commit by A: S^T / dP^T chunk in TMEM
  uint64_t* s_free = bars + 10;       // [buf]    8 math warps: TMEM chunk read out
  uint64_t* pds_full = bars + 12;     // [cc]     8 math warps: P^T / dS^T chunk in shared memory
  uint64_t* pds_free = bars + 14;     //          commit by B after a query tile's dQ MMAs
  uint64_t* dkv_full = bars + 15;     //          commit by B after a key tile's last dV / dK
  uint64_t* dkv_free = bars + 16;     //          4 epilogue warps
  uint64_t* dq_full = bars + 17;      //          commit by B after an item's last dQ
  uint64_t* dq_free = bars + 18;      //          4 epilogue warps
  uint64_t* st_full = bars + 19;      //          32 lanes of warp 3: statistics of the item
  uint64_t* st_empty = bars + 20;     //          8 math warps
  uint64_t* dl_ready = bars + 21;     // [item&1] 128 delta threads: delta of the item is in global
  uint64_t* dl_free = bars + 23;      // [item&1] warp 3 has taken it (bounds the delta warps' run-ahead)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + kNumBars);

  // warp index through a shuffle: the role branches are then provably warp-uniform for ptxas
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const int my_items = (p.n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
#ifdef SAV_ATTN_DEV
  int dbg_n = 0;
  const bool dbg_on = (lane == 0);
#endif

  if (tid == 0) {
    if (smem_u32(smem) & 1023u) __trap();   // SW128 tiles need a 1024-byte aligned base
    tma_prefetch_desc(&tq); tma_prefetch_desc(&tk); tma_prefetch_desc(&tv); tma_prefetch_desc(&tdo);
    tma_prefetch_desc(&tdq); tma_prefetch_desc(&tdk); tma_prefetch_desc(&tdv);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&qdo_full[i], 1); mbar_init(&qdo_empty[i], 2);
      mbar_init(&kv_full[i], 1); mbar_init(&kv_empty[i], 2);
      mbar_init(&s_full[i], 1); mbar_init(&s_free[i], 8);
      mbar_init(&pds_full[i], 8);
    }
    mbar_init(pds_free, 1);
    mbar_init(dkv_full, 1); mbar_init(dkv_free, 4);
    mbar_init(dq_full, 1); mbar_init(dq_free, 4);
    mbar_init(st_full, 32); mbar_init(st_empty, 8);
    mbar_init(&dl_ready[0], 128); mbar_init(&dl_ready[1], 128);
    mbar_init(&dl_free[0], 1); mbar_init(&dl_free[1], 1);
    fence_barrier_init();
  }
  if (warp == 3) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tdP = tmem + 128, tdV = tmem + 256, tdK = tmem + 320, tdQ = tmem + 384;
  constexpr uint32_t TILE16 = BT_BYTES >> 4;  // descriptor address units (16 B)
  auto lo = [](uint64_t d) { return (uint32_t)d; };
  auto hi = [](uint64_t d) { return (uint32_t)(d >> 32); };

  // Register budget: the CTA's pool is what it was launched with, 640 threads x 96 = 61 440.
  // Warpgroup limits after redistribution: 56 (producer / issuers / statistics) + 2 x 128 (math) +
  // 80 (epilogue) + 80 (delta) = 472 x 128 threads = 60 416 <= pool, so both setmaxnreg.inc can be
  // satisfied (a sum above the pool deadlocks the second one).  The code of a role has to be
  // DOMINATED by its warpgroup's setmaxnreg for ptxas to allocate against the new limit (a
  // setmaxnreg.dec that merely precedes the role dispatch caps the whole kernel), hence the nesting.
  if (warp < 4) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t gk = 0, quse[2] = {0, 0};   // quse[i]: how often query buffer i has been filled
      for (int itl = 0; itl < my_items; ++itl) {
        const BwdItem it = bwd_item<kGroups>(p, blockIdx.x + itl * gridDim.x);
        const int h = it.h, b = it.b;
        auto load_qdo = [&](int i) {
          BWD_STAMP(0, 1);
          mbar_wait_sleep(&qdo_empty[i], (quse[i] & 1) ^ 1);
          ++quse[i];
          BWD_STAMP(0, 2);
          mbar_expect_tx(&qdo_full[i], 2 * BT_BYTES);
          tma_load_3d(sQ + i * BT_BYTES, &tq, &qdo_full[i], h * BD, it.q0 + i * BT, b);
          tma_load_3d(sdO + i * BT_BYTES, &tdo, &qdo_full[i], h * BD, it.q0 + i * BT, b);
        };
        auto load_kv = [&](int j) {
          const int st = gk & 1;
          BWD_STAMP(0, 3);
          mbar_wait_sleep(&kv_empty[st], ((gk >> 1) & 1) ^ 1);
          BWD_STAMP(0, 4);
          mbar_expect_tx(&kv_full[st], 2 * BT_BYTES);
          tma_load_3d(sKV + st * 2 * BT_BYTES, &tk, &kv_full[st], h * BD, j * BT, b);
          tma_load_3d(sKV + st * 2 * BT_BYTES + BT_BYTES, &tv, &kv_full[st], h * BD, j * BT, b);
          ++gk;
        };
        load_qdo(0);
        load_kv(0);
        if (it.nqt > 1) load_qdo(1);
        for (int j = 1; j < p.nkt; ++j) load_kv(j);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer A: S^T, dP^T
    // whole warp walks the loop, one elected lane issues: warp-uniform control flow keeps the
    // descriptors in uniform registers (no per-MMA vote / R2UR waterfall)
    const uint64_t dQ0 = make_smem_desc(smem_u32(sQ), 16, 1024);            // K-major B of S^T
    const uint64_t ddO0 = make_smem_desc(smem_u32(sdO), 16, 1024);          // K-major B of dP^T
    const uint64_t dK0 = make_smem_desc(smem_u32(sKV), 16, 1024);           // K-major A of S^T
    const uint64_t dV0 = make_smem_desc(smem_u32(sKV + BT_BYTES), 16, 1024);
    uint32_t n = 0, gk = 0, quse[2] = {0, 0};
    for (int itl = 0; itl < my_items; ++itl) {
      const BwdItem it = bwd_item<kGroups>(p, blockIdx.x + itl * gridDim.x);
      for (int j = 0; j < p.nkt; ++j, ++gk) {
        const int st = gk & 1;
        BWD_STAMP(1, 10);
        mbar_wait_sleep(&kv_full[st], (gk >> 1) & 1);
        BWD_STAMP(1, 11);
        const uint32_t a_k = lo(dK0) + st * 2 * TILE16, a_v = lo(dV0) + st * 2 * TILE16;
        for (int i = 0; i < it.nqt; ++i) {
          if (j == 0) {
            BWD_STAMP(1, 14);
            mbar_wait_sleep(&qdo_full[i], quse[i] & 1);
            ++quse[i];
            BWD_STAMP(1, 15);
          }
          const int nq_cur = round16(min(BT, it.nq - i * BT));
          const int nch = (nq_cur + BC - 1) / BC;
          for (int cc = 0; cc < nch; ++cc, ++n) {
            const int ncols = min(BC, nq_cur - cc * BC);
            const uint32_t buf = n & 1, use = n >> 1;
            BWD_STAMP(1, 12);
            if (use > 0) mbar_wait_sleep(&s_free[buf], (use - 1) & 1);
            BWD_STAMP(1, 13);
            tc_fence_after();
            const uint32_t id_s = make_idesc_bf16(BT, ncols, false, false);
            // rows [cc*64, cc*64 + ncols) of the K-major Q_i / dO_i tiles: 64 rows = 8 KB
            const uint32_t b_q = lo(dQ0) + i * TILE16 + cc * (TILE16 / 2);
            const uint32_t b_do = lo(ddO0) + i * TILE16 + cc * (TILE16 / 2);
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < BD / 16; ++k)
                umma_f16_lh(tS + buf * BC, a_k + k * 2, hi(dK0), b_q + k * 2, hi(dQ0), id_s, k > 0);
#pragma unroll
              for (int k = 0; k < BD / 16; ++k)
                umma_f16_lh(tdP + buf * BC, a_v + k * 2, hi(dV0), b_do + k * 2, hi(ddO0), id_s, k > 0);
              umma_commit(&s_full[buf]);
              if (i == it.nqt - 1 && cc == nch - 1) umma_commit(&kv_empty[st]);
              if (j == p.nkt - 1 && cc == nch - 1) umma_commit(&qdo_empty[i]);
            }
            __syncwarp();
          }
        }
      }
    }
  } else if (warp == 2) {
    // ------------------------------------------------------------------ MMA issuer B: dV, dK, dQ
    const uint64_t dQm0 = make_smem_desc(smem_u32(sQ), BT_BYTES, 1024);     // MN-major B of dK
    const uint64_t ddOm0 = make_smem_desc(smem_u32(sdO), BT_BYTES, 1024);   // MN-major B of dV
    const uint64_t dKm0 = make_smem_desc(smem_u32(sKV), BT_BYTES, 1024);    // MN-major B of dQ
    const uint64_t dPt0 = make_smem_desc(smem_u32(sPt), 16, 1024);          // K-major A of dV
    const uint64_t ddSt0 = make_smem_desc(smem_u32(sdSt), 16, 1024);        // K-major A of dK
    const uint64_t ddSm0 = make_smem_desc(smem_u32(sdSt), BT_BYTES, 1024);  // MN-major A of dQ
    constexpr uint32_t id_kmn = make_idesc_bf16(BT, BD, false, true);
    constexpr uint32_t id_mnmn = make_idesc_bf16(BT, BD, true, true);
    uint32_t gk = 0, use_cc[2] = {0, 0};
    for (int itl = 0; itl < my_items; ++itl) {
      const BwdItem it = bwd_item<kGroups>(p, blockIdx.x + itl * gridDim.x);
      for (int j = 0; j < p.nkt; ++j, ++gk) {
        const int st = gk & 1;
        const int nk_steps = round16(min(BT, p.Nk - j * BT)) >> 4;
        const uint32_t b_k = lo(dKm0) + st * 2 * TILE16;
        for (int i = 0; i < it.nqt; ++i) {
          const int nq_cur = round16(min(BT, it.nq - i * BT));
          const int nch = (nq_cur + BC - 1) / BC;
          for (int cc = 0; cc < nch; ++cc) {
            const int ksteps = min(BC, nq_cur - cc * BC) >> 4;
            BWD_STAMP(2, 20);
            mbar_wait_sleep(&pds_full[cc], use_cc[cc] & 1);
            BWD_STAMP(2, 21);
            ++use_cc[cc];
            const bool first_of_keytile = (i == 0 && cc == 0);
            const bool last_of_tile = (cc == nch - 1);
            if (first_of_keytile && gk > 0) mbar_wait_sleep(dkv_free, (gk - 1) & 1);
            if (last_of_tile && j == 0 && i == 0 && itl > 0) mbar_wait_sleep(dq_free, (itl - 1) & 1);
            BWD_STAMP(2, 22);
            tc_fence_after();
            // dV[keys,d] += P^T[keys,q] dO_i[q,d];  dK[keys,d] += dS^T[keys,q] Q_i[q,d]  (K = q)
            const uint32_t a_p = lo(dPt0) + cc * TILE16, a_ds = lo(ddSt0) + cc * TILE16;
            const uint32_t b_do = lo(ddOm0) + i * TILE16 + cc * 4 * 128;
            const uint32_t b_q = lo(dQm0) + i * TILE16 + cc * 4 * 128;
            const uint32_t acc0 = first_of_keytile ? 0u : 1u;
            if (elect_one()) {
#pragma unroll
              for (int ks = 0; ks < 4; ++ks)
                if (ks < ksteps)
                  umma_f16_lh(tdV, a_p + ks * 2, hi(dPt0), b_do + ks * 128, hi(ddOm0), id_kmn,
                              ks > 0 ? 1u : acc0);
#pragma unroll
              for (int ks = 0; ks < 4; ++ks)
                if (ks < ksteps)
                  umma_f16_lh(tdK, a_ds + ks * 2, hi(ddSt0), b_q + ks * 128, hi(dQm0), id_kmn,
                              ks > 0 ? 1u : acc0);
              if (last_of_tile) {
                // dQ_i[q,d] += dS[q,keys] K_j[keys,d]  (K = keys; A = both dS^T chunks, MN-major)
#pragma unroll
                for (int ks = 0; ks < 8; ++ks)
                  if (ks < nk_steps)
                    umma_f16_lh(tdQ + i * BD, lo(ddSm0) + ks * 128, hi(ddSm0), b_k + ks * 128,
                                hi(dKm0), id_mnmn, ks > 0 ? 1u : (j > 0 ? 1u : 0u));
                umma_commit(pds_free);
                if (i == it.nqt - 1) {
                  umma_commit(dkv_full);
                  umma_commit(&kv_empty[st]);
                }
                if (j == p.nkt - 1) umma_commit(&qdo_empty[i]);
                if (j == p.nkt - 1 && i == it.nqt - 1) umma_commit(dq_full);
              }
            }
            __syncwarp();
          }
        }
      }
    }
  } else if (warp == 3) {
    // ------------------------------------------------------------------ per-item statistics
    // 8 query columns per lane.  delta comes from the delta warps through the global workspace;
    // the loads of item itl are issued before waiting for the math warps to release the buffer,
    // so only a shared-memory store sits on the critical path.
    for (int itl = 0; itl < my_items; ++itl) {
      const BwdItem it = bwd_item<kGroups>(p, blockIdx.x + itl * gridDim.x);
      const int64_t base = (int64_t)it.bh * p.Nq + it.q0;   // lse / delta are [B, H, Nq]
      float nl[8], nd[8];
      BWD_STAMP(6, 50);
      mbar_wait_sleep(&dl_ready[itl & 1], (itl >> 1) & 1);
      BWD_STAMP(6, 51);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int c = lane * 8 + e;
        nl[e] = (c < it.nq) ? -p.lse[base + c] * B_LOG2E : -INFINITY;
        nd[e] = (c < it.nq) ? -__ldcg(p.delta + base + c) : 0.f;
      }
      // (the values are consumed by the stores below, i.e. the loads have completed by then)
      BWD_STAMP(6, 52);
      if (itl > 0) mbar_wait_sleep(st_empty, (itl - 1) & 1);
      BWD_STAMP(6, 53);
#pragma unroll
      for (int e = 0; e < 8; e += 4) {
        *reinterpret_cast<float4*>(s_nl2 + lane * 8 + e) = make_float4(nl[e], nl[e + 1], nl[e + 2], nl[e + 3]);
        *reinterpret_cast<float4*>(s_nd + lane * 8 + e) = make_float4(nd[e], nd[e + 1], nd[e + 2], nd[e + 3]);
      }
      mbar_arrive(st_full);
      __syncwarp();
      if (lane == 0) mbar_arrive(&dl_free[itl & 1]);
    }
  }
  } else if (warp < 12) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 128;");
    // ------------------------------------------------------------------ math warps
    const int sub = warp & 3;              // TMEM sub-partition
    const int hf = (warp - 4) >> 2;        // which 32 of the chunk's 64 query columns
    const int r = sub * 32 + lane;         // key row inside the tile
    const uint32_t lane_off = (uint32_t)(sub * 32) << 16;
    const uint64_t sc2 = f2_pack(p.scale_log2, p.scale_log2);
    uint32_t n = 0, tiles = 0;
#ifdef SAV_ATTN_DEV
    const int mrole = (warp == 4) ? 3 : (warp == 8 ? 4 : 7);
#endif
    for (int itl = 0; itl < my_items; ++itl) {
      const BwdItem it = bwd_item<kGroups>(p, blockIdx.x + itl * gridDim.x);
      BWD_STAMP(mrole, 35);
      mbar_wait(st_full, itl & 1);
      BWD_STAMP(mrole, 36);
      for (int j = 0; j < p.nkt; ++j) {
        // key rows at or beyond the (16-rounded) key count of this tile are never read: as K of
        // the dQ MMAs they are cut by nk_steps, as M rows of dV / dK they are clipped by the store
        const bool rows_live = sub * 32 < round16(min(BT, p.Nk - j * BT));
        for (int i = 0; i < it.nqt; ++i, ++tiles) {
          const int nq_cur = round16(min(BT, it.nq - i * BT));
          const int nch = (nq_cur + BC - 1) / BC;
          for (int cc = 0; cc < nch; ++cc, ++n) {
            const int ncols = min(BC, nq_cur - cc * BC);
            const uint32_t buf = n & 1, use = n >> 1;
            const bool live = rows_live && hf * 32 < ncols;
            uint32_t pk_p[16], pk_ds[16];
            BWD_STAMP(mrole, 30);
            mbar_wait(&s_full[buf], use & 1);
            BWD_STAMP(mrole, 31);
            tc_fence_after();
            if (live) {
              uint32_t rs[32], rp[32];
              tmem_ld32(tS + buf * BC + lane_off + hf * 32, rs);
              tmem_ld32(tdP + buf * BC + lane_off + hf * 32, rp);
              tmem_ld_wait();
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(&s_free[buf]);   // TMEM chunk may be overwritten
              const float* nl = s_nl2 + i * BT + cc * BC + hf * 32;
              const float* nd = s_nd + i * BT + cc * BC + hf * 32;
#pragma unroll
              for (int e = 0; e < 32; e += 4) {
                const float4 l4 = *reinterpret_cast<const float4*>(nl + e);
                const float4 d4 = *reinterpret_cast<const float4*>(nd + e);
                float t0, t1, t2, t3;
                f2_unpack(f2_fma(f2_pack_u(rs[e], rs[e + 1]), sc2, f2_pack(l4.x, l4.y)), t0, t1);
                f2_unpack(f2_fma(f2_pack_u(rs[e + 2], rs[e + 3]), sc2, f2_pack(l4.z, l4.w)), t2, t3);
                const float p0 = ex2_fast(t0), p1 = ex2_fast(t1), p2 = ex2_fast(t2), p3 = ex2_fast(t3);
                float s0, s1, s2, s3;
                f2_unpack(f2_mul(f2_pack(p0, p1), f2_add(f2_pack_u(rp[e], rp[e + 1]), f2_pack(d4.x, d4.y))), s0, s1);
                f2_unpack(f2_mul(f2_pack(p2, p3), f2_add(f2_pack_u(rp[e + 2], rp[e + 3]), f2_pack(d4.z, d4.w))), s2, s3);
                pk_p[e >> 1] = pack_bf16(p0, p1);
                pk_p[(e >> 1) + 1] = pack_bf16(p2, p3);
                pk_ds[e >> 1] = pack_bf16(s0, s1);
                pk_ds[(e >> 1) + 1] = pack_bf16(s2, s3);
              }
            } else {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(&s_free[buf]);
            }
            // the first chunk of a query tile reuses buffers the previous tile's dV / dK / dQ MMAs
            // read: wait for issuer B's commit behind that tile's dQ
            BWD_STAMP(mrole, 32);
            if (cc == 0 && tiles > 0) mbar_wait(pds_free, (tiles - 1) & 1);
            BWD_STAMP(mrole, 33);
            if (live) {
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                const uint32_t off = cc * BT_BYTES + bsw128(r, hf * 4 + g);
                *reinterpret_cast<uint4*>(sPt + off) =
                    make_uint4(pk_p[g * 4], pk_p[g * 4 + 1], pk_p[g * 4 + 2], pk_p[g * 4 + 3]);
                *reinterpret_cast<uint4*>(sdSt + off) =
                    make_uint4(pk_ds[g * 4], pk_ds[g * 4 + 1], pk_ds[g * 4 + 2], pk_ds[g * 4 + 3]);
              }
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(&pds_full[cc]);
            BWD_STAMP(mrole, 34);
          }
        }
      }
      // this item's statistics are no longer needed
      __syncwarp();
      if (lane == 0) mbar_arrive(st_empty);
    }
  } else if (warp < 16) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 80;");
    // ------------------------------------------------------------------ epilogue warps (12..15)
    const int sub = warp & 3;
    const int r = sub * 32 + lane;         // key row (dV / dK) or query row (dQ) of the tile
    const int et = tid - 384;              // 0..127
    const bool leader = (et == 0);
    const uint32_t lane_off = (uint32_t)(sub * 32) << 16;
    uint32_t gk = 0, nstore = 0;
    // one [128 x 64] bf16 tile: registers -> staging buffer (nstore & 1) -> TMA store.  The buffer
    // is reused every second store: the leader keeps at most one store group reading.
    auto store_tile = [&](const CUtensorMap* map, const uint32_t (&pk)[32], int c0, int c1, int c2) {
      uint8_t* stage = sStage + (nstore & 1) * BT_BYTES;
      if (leader) tma_store_wait_read<1>();
      named_barrier_sync(2, 128);
      pack_rows_to_stage(stage, r, pk);
      fence_proxy_async_smem();
      named_barrier_sync(2, 128);
      if (leader) {
        tma_store_3d(map, stage, c0, c1, c2);
        tma_store_commit();
      }
      ++nstore;
    };
    // partial tile of one query group: 64 fp32 TMEM columns -> two [128 x 32] fp32 half tiles
    // (128-byte rows, same swizzle) -> reduce-add into the fp32 workspace
    auto reduce_tile_f32 = [&](const CUtensorMap* map, uint32_t taddr, float f, int c0, int c1, int c2,
                               bool release) {
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        uint32_t r0[32];
        tmem_ld32(taddr + half * 32, r0);
        tmem_ld_wait();
        if (release && half == 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(dkv_free);      // next key tile may overwrite dV / dK
        }
        uint8_t* stage = sStage + (nstore & 1) * BT_BYTES;
        if (leader) tma_store_wait_read<1>();
        named_barrier_sync(2, 128);
#pragma unroll
        for (int gi = 0; gi < 8; ++gi)
          *reinterpret_cast<float4*>(stage + bsw128(r, gi)) =
              make_float4(__uint_as_float(r0[gi * 4]) * f, __uint_as_float(r0[gi * 4 + 1]) * f,
                          __uint_as_float(r0[gi * 4 + 2]) * f, __uint_as_float(r0[gi * 4 + 3]) * f);
        fence_proxy_async_smem();
        named_barrier_sync(2, 128);
        if (leader) {
          tma_reduce_add_3d(map, stage, c0 + half * 32, c1, c2);
          tma_store_commit();
        }
        ++nstore;
      }
    };
    for (int itl = 0; itl < my_items; ++itl) {
      const BwdItem it = bwd_item<kGroups>(p, blockIdx.x + itl * gridDim.x);
      const int h = it.h, b = it.b;
      for (int j = 0; j < p.nkt; ++j, ++gk) {
        uint32_t pk[32];
        BWD_STAMP(5, 42);
        mbar_wait(dkv_full, gk & 1);
        BWD_STAMP(5, 43);
        tc_fence_after();
        if (kGroups) {
          reduce_tile_f32(&tdv, tdV + lane_off, 1.f, h * BD, j * BT, b, false);
          reduce_tile_f32(&tdk, tdK + lane_off, p.scale, h * BD, j * BT, b, true);
        } else {
          tmem_row64_to_bf16(tdV + lane_off, 1.f, pk);
          store_tile(&tdv, pk, h * BD, j * BT, b);   // key rows >= Nk are clipped by the map
          tmem_row64_to_bf16(tdK + lane_off, p.scale, pk);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(dkv_free);        // next key tile may overwrite dV / dK
          store_tile(&tdk, pk, h * BD, j * BT, b);
        }
        BWD_STAMP(5, 44);
      }
      BWD_STAMP(5, 45);
      mbar_wait(dq_full, itl & 1);
      BWD_STAMP(5, 46);
      tc_fence_after();
      {
        uint32_t pk[32];
        tmem_row64_to_bf16(tdQ + lane_off, p.scale, pk);
        if (it.nqt == 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(dq_free);       // next item may overwrite dQ
        }
        store_tile(&tdq, pk, h * BD, it.q0, b);   // rows >= Nq clipped
        if (it.nqt > 1) {
          tmem_row64_to_bf16(tdQ + BD + lane_off, p.scale, pk);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(dq_free);
          store_tile(&tdq, pk, h * BD, it.q0 + BT, b);
        }
      }
    }
    if (leader) tma_store_wait<0>();
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 80;");
    // ------------------------------------------------------------------ delta warps (16..19)
    // delta[q] = sum_d dO[q,d] (O + O_lo)[q,d] of the item AFTER the one the math warps are on:
    // 8 lanes per 128-byte row, 16 rows per pass of the 128 threads, the loads of 6 passes in
    // flight.  Nothing waits for these warps except the statistics hand-over of the next item.
    const int dt = tid - 512;
    const int part = dt & 7, row0 = dt >> 3;
    for (int itl = 0; itl < my_items; ++itl) {
      const BwdItem it = bwd_item<kGroups>(p, blockIdx.x + itl * gridDim.x);
      const int h = it.h, b = it.b;
      const int64_t base = (int64_t)b * p.o_bs + (int64_t)it.q0 * p.o_rs + h * BD + part * 8;
      float* dout = p.delta + (int64_t)it.bh * p.Nq + it.q0;
      if (itl >= 2) mbar_wait_sleep(&dl_free[itl & 1], ((itl - 2) >> 1) & 1);
      constexpr int PB = 4;
      for (int q0 = 0; q0 < it.nq; q0 += 16 * PB) {
        uint4 a[PB], l[PB], g[PB];
#pragma unroll
        for (int u = 0; u < PB; ++u) {
          const int q = q0 + u * 16 + row0;
          if (q < it.nq) {
            const int64_t off = base + (int64_t)q * p.o_rs;
            a[u] = __ldcg(reinterpret_cast<const uint4*>(p.o + off));
            g[u] = __ldcg(reinterpret_cast<const uint4*>(p.d_o + off));
            l[u] = p.o_lo ? __ldcg(reinterpret_cast<const uint4*>(p.o_lo + off)) : make_uint4(0, 0, 0, 0);
          }
        }
#pragma unroll
        for (int u = 0; u < PB; ++u) {
          const int q = q0 + u * 16 + row0;
          float sum = 0.f;
          if (q < it.nq) {
            const uint32_t aw[4] = {a[u].x, a[u].y, a[u].z, a[u].w}, gw[4] = {g[u].x, g[u].y, g[u].z, g[u].w};
            const uint32_t lw[4] = {l[u].x, l[u].y, l[u].z, l[u].w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float2 af = unpack_bf16(aw[k]), gf = unpack_bf16(gw[k]), lf = unpack_bf16(lw[k]);
              sum += (af.x + lf.x) * gf.x + (af.y + lf.y) * gf.y;
            }
          }
          sum += __shfl_xor_sync(0xffffffffu, sum, 1);
          sum += __shfl_xor_sync(0xffffffffu, sum, 2);
          sum += __shfl_xor_sync(0xffffffffu, sum, 4);
          if (q < it.nq && part == 0) dout[q] = sum;
        }
      }
      mbar_arrive(&dl_ready[itl & 1]);   // release: this thread's global stores are ordered before
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 3) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

static int make_bnhd_map128(CUtensorMap* m, const void* base, int64_t H, int64_t N, int64_t B,
                            int64_t bs, int64_t rs) {
  uint64_t dims[3] = {(uint64_t)(H * BD), (uint64_t)N, (uint64_t)B};
  uint64_t str[3] = {0, (uint64_t)rs * 2, (uint64_t)bs * 2};
  uint32_t box[3] = {BD, BT, 1};
  return make_tmap_bf16(m, base, 3, dims, str, box);
}

// fp32 accumulator [B, N, H*64] (contiguous) -> bf16 view with element strides bs / rs
__global__ void acc_to_bnhd_kernel(const float* __restrict__ acc, bf16* __restrict__ base, int64_t total8,
                                   int N, int row8, int64_t bs, int64_t rs) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total8;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % row8);
    const int64_t t = i / row8;
    const int n = (int)(t % N);
    const int64_t b = t / N;
    const float4 lo = __ldcs(reinterpret_cast<const float4*>(acc + i * 8));
    const float4 hi = __ldcs(reinterpret_cast<const float4*>(acc + i * 8) + 1);
    *reinterpret_cast<uint4*>(base + b * bs + (int64_t)n * rs + c * 8) =
        make_uint4(pack_bf16(lo.x, lo.y), pack_bf16(lo.z, lo.w), pack_bf16(hi.x, hi.y), pack_bf16(hi.z, hi.w));
  }
}

static int acc_to_bnhd(const float* acc, void* base, int64_t B, int64_t N, int64_t H, int64_t bs, int64_t rs,
                       cudaStream_t st) {
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (bs % 8) || (rs % 8))
    return set_error(SAV_ERR_ALIGN, "sav_attn_bwd: dk / dv need 16-byte aligned rows");
  const int row8 = (int)(H * BD / 8);
  const int64_t total8 = B * N * row8;
  const int blocks = (int)((total8 + 255) / 256 < 148 * 8 ? (total8 + 255) / 256 : 148 * 8);
  acc_to_bnhd_kernel<<<blocks, 256, 0, st>>>(acc, reinterpret_cast<bf16*>(base), total8, (int)N, row8, bs, rs);
  count_launch();
  return check_launch("sav_attn_bwd (dk / dv conversion)");
}

// sav_attn_bwd for every shape (called from attn.cu)
int launch_attn_bwd2(const sav_attn_bwd_args* a, cudaStream_t st) {
  CUtensorMap tq, tk, tv, tdo, tdq, tdk, tdv;
  int rc;
  if ((rc = make_bnhd_map128(&tq, a->q, a->H, a->Nq, a->B, a->q_batch_stride, a->q_row_stride))) return rc;
  if ((rc = make_bnhd_map128(&tk, a->k, a->H, a->Nk, a->B, a->k_batch_stride, a->k_row_stride))) return rc;
  if ((rc = make_bnhd_map128(&tv, a->v, a->H, a->Nk, a->B, a->v_batch_stride, a->v_row_stride))) return rc;
  if ((rc = make_bnhd_map128(&tdo, a->d_o, a->H, a->Nq, a->B, a->o_batch_stride, a->o_row_stride))) return rc;
  if ((rc = make_bnhd_map128(&tdq, a->dq, a->H, a->Nq, a->B, a->q_batch_stride, a->q_row_stride))) return rc;
  const int n_groups = (int)((a->Nq + 2 * BT - 1) / (2 * BT));
  const int64_t acc_elems = a->B * a->Nk * a->H * BD;           // one fp32 accumulator (dK; dV follows)
  if (n_groups > 1) {
    // several query groups per (b, h): fp32 half tiles [128 x 32] reduce-added into the workspace
    if (!a->dq_accum)
      return set_error(SAV_ERR_ARG, "sav_attn_bwd: Nq > 256 needs the fp32 workspace (2 * B * Nk * H * 64 floats)");
    uint64_t dims[3] = {(uint64_t)(a->H * BD), (uint64_t)a->Nk, (uint64_t)a->B};
    uint64_t str[3] = {0, (uint64_t)(a->H * BD) * 4, (uint64_t)(a->Nk * a->H * BD) * 4};
    uint32_t box[3] = {32, BT, 1};
    if ((rc = make_tmap_nd(&tdk, a->dq_accum, 4, 3, dims, str, box))) return rc;
    if ((rc = make_tmap_nd(&tdv, a->dq_accum + acc_elems, 4, 3, dims, str, box))) return rc;
  } else {
    if ((rc = make_bnhd_map128(&tdk, a->dk, a->H, a->Nk, a->B, a->k_batch_stride, a->k_row_stride))) return rc;
    if ((rc = make_bnhd_map128(&tdv, a->dv, a->H, a->Nk, a->B, a->v_batch_stride, a->v_row_stride))) return rc;
  }
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(attn_bwd3_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwd2Smem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(attn_bwd3_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwd2Smem);
    if (e != cudaSuccess) return set_error(SAV_ERR_CUDA, "attn bwd2 smem attr: %s", cudaGetErrorString(e));
    attr = true;
  }
  AttnBwd2Params p;
  p.lse = a->lse; p.delta = a->delta;
  p.o = reinterpret_cast<const bf16*>(a->o); p.o_lo = reinterpret_cast<const bf16*>(a->o_lo);
  p.d_o = reinterpret_cast<const bf16*>(a->d_o);
  p.o_bs = a->o_batch_stride; p.o_rs = a->o_row_stride;
  p.B = (int)a->B; p.H = (int)a->H; p.Nq = (int)a->Nq; p.Nk = (int)a->Nk;
  p.G = (p.Nq + 2 * BT - 1) / (2 * BT);
  p.n_items = p.B * p.H * p.G;
  p.nkt = (p.Nk + BT - 1) / BT;
  p.nqt = (p.Nq + BT - 1) / BT;
  if (p.G > 1 && (rc = sav_fill_f32(a->dq_accum, 0.f, 2 * acc_elems, st))) return rc;
  p.scale = a->scale; p.scale_log2 = a->scale * B_LOG2E;
  p.dbg = reinterpret_cast<long long*>(sav_debug_buffer());
  int sms = 148, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = p.n_items < sms ? p.n_items : sms;
  if (p.G > 1) attn_bwd3_kernel<true><<<grid, kBwdThreads, kBwd2Smem, st>>>(tq, tk, tv, tdo, tdq, tdk, tdv, p);
  else attn_bwd3_kernel<false><<<grid, kBwdThreads, kBwd2Smem, st>>>(tq, tk, tv, tdo, tdq, tdk, tdv, p);
  count_launch();
  if ((rc = check_launch("sav_attn_bwd (resident-Q path)"))) return rc;
  if (p.G > 1) {
    if ((rc = acc_to_bnhd(a->dq_accum, a->dk, a->B, a->Nk, a->H, a->k_batch_stride, a->k_row_stride, st))) return rc;
    if ((rc = acc_to_bnhd(a->dq_accum + acc_elems, a->dv, a->B, a->Nk, a->H, a->v_batch_stride, a->v_row_stride, st)))
      return rc;
  }
  return SAV_OK;
}

}  // namespace sav
