// Fused attention backward for sm_100a, Nq <= 256 (reverse mode of attention.py:39-58):
//   S = q k^T * scale, P = exp(S - lse), dP = dO v^T, dS = P o (dP - delta) * scale,
//   dV = P^T dO, dK = dS^T q, dQ = dS k.
// S / P / dP / dS never leave the SM and, unlike the generic kernel in attn.cu, neither do the
// partial dQ sums: one persistent CTA owns a whole (b, h) -- both 128-row query tiles stay resident
// in shared memory while the key tiles stream through a 2-stage TMA ring -- so dQ accumulates in
// TMEM across key tiles (no fp32 red.add to HBM, no dq-convert pass), dK/dV accumulate in TMEM
// across the query tiles.
//
// 384 threads, one CTA per SM:
//   warp 0      TMA producer (Q_i, dO_i per item; K_j, V_j ring)
//   warp 1      MMA issuer: S^T = K_j Q_i^T and dP^T = V_j dO_i^T (the transposed problem: TMEM lane =
//               key row, so P^T / dS^T rows come out K-major for dV += P^T dO_i, dK += dS^T Q_i, and
//               the same dS^T buffer read MN-major is the A operand of dQ_i += dS K_j)
//   warp 2      second MMA issuer (dV, dK); warp 1 issues S^T, dP^T and dQ
//   warp 3      TMEM allocator
//   warps 4-11  compute: thread = (key row, half of the tile's query columns):
//               p = exp2(s*c - lse), ds = p (dp - delta) with packed f32x2 math, bf16 P^T / dS^T into
//               swizzled shared memory; the S^T/dP^T MMAs of the next (j, i) pair are issued as soon
//               as this pair's TMEM reads are done, so tensor work overlaps the exp2 phase.
// TMEM (512 columns): S^T [0,128) | dP^T [128,256) | dV [256,320) | dK [320,384) | dQ_0 [384,448) |
//                     dQ_1 [448,512).  The softmax scale of dK / dQ is applied once at read-out.
#include "common.cuh"
#include "host.h"

namespace sav {

constexpr int BT = 128;                 // tile edge (keys per tile, queries per tile)
constexpr int BD = 64;                  // head dim
constexpr int BT_BYTES = BT * BD * 2;   // 16 KB: one [128 x 64] bf16 tile, 128-byte rows
constexpr int kBwdThreads = 384;
constexpr float B_LOG2E = 1.4426950408889634f;

struct AttnBwd2Params {
  const float* lse;
  const float* delta;
  int B, H, Nq, Nk;
  int n_items;   // B * H
  int nkt;       // key tiles
  int nqt;       // query tiles (1 or 2)
  float scale, scale_log2;
  long long* dbg;
};

#define SAV_RECB(role, ev)                                                        \
  do {                                                                            \
    if (p.dbg && blockIdx.x == 0 && dbg_n < 255) {                                \
      p.dbg[(role) * 512 + 2 * dbg_n] = (ev);                                     \
      p.dbg[(role) * 512 + 2 * dbg_n + 1] = clock64();                            \
      ++dbg_n;                                                                    \
    }                                                                             \
  } while (0)

// byte offset of 16-byte group `g16` (0..7) of row `r` inside a [rows x 128 B] SW128 tile
__device__ __forceinline__ uint32_t bsw128(int r, int g16) {
  return (uint32_t)(r * 128 + ((g16 ^ (r & 7)) << 4));
}

__device__ __forceinline__ uint64_t lds64(const float* p) {
  return *reinterpret_cast<const uint64_t*>(p);
}

__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd2_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                 const __grid_constant__ CUtensorMap tv, const __grid_constant__ CUtensorMap tdo,
                 const __grid_constant__ CUtensorMap tdq, const __grid_constant__ CUtensorMap tdk,
                 const __grid_constant__ CUtensorMap tdv, const AttnBwd2Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sQ = smem;                   // [i] 2 x 16 KB
  uint8_t* sdO = sQ + 2 * BT_BYTES;     // [i] 2 x 16 KB
  uint8_t* sKV = sdO + 2 * BT_BYTES;    // [stage] { K 16 KB, V 16 KB }
  uint8_t* sPt = sKV + 4 * BT_BYTES;    // [128 keys x 128 q] bf16: 2 chunks of [128 x 64]
  uint8_t* sdSt = sPt + 2 * BT_BYTES;   // same
  float* s_nl2 = reinterpret_cast<float*>(sdSt + 2 * BT_BYTES);  // [256] -lse * log2(e)  (-inf beyond Nq)
  float* s_nd = s_nl2 + 2 * BT;                                   // [256] -delta
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_nd + 2 * BT);
  uint64_t* qdo_full = bars;        // [i]
  uint64_t* qdo_empty = bars + 2;   // [i]
  uint64_t* kv_full = bars + 4;     // [stage]
  uint64_t* kv_empty = bars + 6;    // [stage]
  uint64_t* st_full = bars + 8;
  uint64_t* st_free = bars + 9;     // 256 arrivals
  uint64_t* pds_full = bars + 10;   // 256 arrivals
  uint64_t* pds_free = bars + 11;
  uint64_t* dkv_full = bars + 12;
  uint64_t* dkv_free = bars + 13;   // 256 arrivals
  uint64_t* dq_full = bars + 14;
  uint64_t* dq_free = bars + 15;    // 256 arrivals
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  // warp index through a shuffle: the role branches are then provably warp-uniform for ptxas
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const int ppi = p.nkt * p.nqt;  // (key tile, query tile) pairs per item
  const int my_items = (p.n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int total_pairs = my_items * ppi;

  if (tid == 0) {
    tma_prefetch_desc(&tq); tma_prefetch_desc(&tk); tma_prefetch_desc(&tv); tma_prefetch_desc(&tdo);
    tma_prefetch_desc(&tdq); tma_prefetch_desc(&tdk); tma_prefetch_desc(&tdv);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&qdo_full[i], 1); mbar_init(&qdo_empty[i], 2);
      mbar_init(&kv_full[i], 1); mbar_init(&kv_empty[i], 1);
    }
    mbar_init(st_full, 1); mbar_init(st_free, 256);
    mbar_init(pds_full, 256); mbar_init(pds_free, 2);
    mbar_init(dkv_full, 1); mbar_init(dkv_free, 256);
    mbar_init(dq_full, 1); mbar_init(dq_free, 256);
    fence_barrier_init();
  }
  if (warp == 3) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tSt = tmem, tdPt = tmem + 128, tdV = tmem + 256, tdK = tmem + 320, tdQ = tmem + 384;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t gk = 0;
      for (int itl = 0; itl < my_items; ++itl) {
        const int item = blockIdx.x + itl * gridDim.x;
        const int h = item % p.H, b = item / p.H;
        auto load_qdo = [&](int i) {
          mbar_wait_sleep(&qdo_empty[i], (itl & 1) ^ 1);
          mbar_expect_tx(&qdo_full[i], 2 * BT_BYTES);
          tma_load_3d(sQ + i * BT_BYTES, &tq, &qdo_full[i], h * BD, i * BT, b);
          tma_load_3d(sdO + i * BT_BYTES, &tdo, &qdo_full[i], h * BD, i * BT, b);
        };
        auto load_kv = [&](int j) {
          const int st = gk & 1;
          mbar_wait_sleep(&kv_empty[st], ((gk >> 1) & 1) ^ 1);
          mbar_expect_tx(&kv_full[st], 2 * BT_BYTES);
          tma_load_3d(sKV + st * 2 * BT_BYTES, &tk, &kv_full[st], h * BD, j * BT, b);
          tma_load_3d(sKV + st * 2 * BT_BYTES + BT_BYTES, &tv, &kv_full[st], h * BD, j * BT, b);
          ++gk;
        };
        load_qdo(0);
        load_kv(0);
        if (p.nqt > 1) load_qdo(1);
        for (int j = 1; j < p.nkt; ++j) load_kv(j);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    // whole warp walks the loop, one elected lane issues: warp-uniform control flow keeps the
    // descriptors in uniform registers (no per-MMA vote / R2UR waterfall)
    {
      const uint64_t dQ0 = make_smem_desc(smem_u32(sQ), 16, 1024);            // K-major B of S^T
      const uint64_t ddO0 = make_smem_desc(smem_u32(sdO), 16, 1024);          // K-major B of dP^T
      const uint64_t dQm0 = make_smem_desc(smem_u32(sQ), BT_BYTES, 1024);     // MN-major B of dK
      const uint64_t ddOm0 = make_smem_desc(smem_u32(sdO), BT_BYTES, 1024);   // MN-major B of dV
      const uint64_t dK0 = make_smem_desc(smem_u32(sKV), 16, 1024);           // K-major A of S^T
      const uint64_t dV0 = make_smem_desc(smem_u32(sKV + BT_BYTES), 16, 1024);
      const uint64_t dKm0 = make_smem_desc(smem_u32(sKV), BT_BYTES, 1024);    // MN-major B of dQ
      const uint64_t dPt0 = make_smem_desc(smem_u32(sPt), 16, 1024);          // K-major A of dV
      const uint64_t ddSt0 = make_smem_desc(smem_u32(sdSt), 16, 1024);        // K-major A of dK
      const uint64_t ddSm0 = make_smem_desc(smem_u32(sdSt), BT_BYTES, 1024);  // MN-major A of dQ
      constexpr uint32_t id_kmn = make_idesc_bf16(BT, BD, false, true);
      constexpr uint32_t id_mnmn = make_idesc_bf16(BT, BD, true, true);
      constexpr uint32_t TILE16 = BT_BYTES >> 4;  // descriptor address units (16 B)
      int dbg_n = lane == 0 ? 0 : 1000;

      auto lo = [](uint64_t d) { return (uint32_t)d; };
      auto hi = [](uint64_t d) { return (uint32_t)(d >> 32); };
      // kv tile bookkeeping is derived from the linear pair index q of this CTA
      auto issue_sdp = [&](int q) {
        const int itl = q / ppi, rem = q - itl * ppi;
        const int j = rem / p.nqt, i = rem - j * p.nqt;
        const uint32_t gk = (uint32_t)(itl * p.nkt + j);
        const int st = gk & 1;
        if (i == 0) mbar_wait_sleep(&kv_full[st], (gk >> 1) & 1);
        if (j == 0) mbar_wait_sleep(&qdo_full[i], itl & 1);
        SAV_RECB(1, 40);
        if (q > 0) mbar_wait_sleep(st_free, (q - 1) & 1);
        SAV_RECB(1, 41);
        tc_fence_after();
        const int nq_cur = (min(BT, p.Nq - i * BT) + 15) & ~15;
        const uint32_t id_s = make_idesc_bf16(BT, nq_cur, false, false);
        const uint32_t a_k = lo(dK0) + st * 2 * TILE16, a_v = lo(dV0) + st * 2 * TILE16;
        const uint32_t b_q = lo(dQ0) + i * TILE16, b_do = lo(ddO0) + i * TILE16;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BD / 16; ++k)
            umma_f16_lh(tSt, a_k + k * 2, hi(dK0), b_q + k * 2, hi(dQ0), id_s, k > 0);
#pragma unroll
          for (int k = 0; k < BD / 16; ++k)
            umma_f16_lh(tdPt, a_v + k * 2, hi(dV0), b_do + k * 2, hi(ddO0), id_s, k > 0);
          umma_commit(st_full);
        }
        __syncwarp();
        SAV_RECB(1, 42);
      };
      auto issue_grads = [&](int q) {
        const int itl = q / ppi, rem = q - itl * ppi;
        const int j = rem / p.nqt, i = rem - j * p.nqt;
        const uint32_t gk = (uint32_t)(itl * p.nkt + j);
        const int st = gk & 1;
        SAV_RECB(1, 43);
        mbar_wait_sleep(pds_full, q & 1);
        SAV_RECB(1, 44);
        if (j == 0 && i == 0 && itl > 0) mbar_wait_sleep(dq_free, (itl - 1) & 1);
        tc_fence_after();
        const int nk_steps = ((min(BT, p.Nk - j * BT) + 15) & ~15) >> 4;
        const uint32_t b_k = lo(dKm0) + st * 2 * TILE16;
        const uint32_t acc_j = j > 0;
        // dQ_i[q,d] += dS[q,keys] K_j[keys,d]   (K = keys; A = the dS^T buffer read MN-major)
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            if (ks < nk_steps)
              umma_f16_lh(tdQ + i * BD, lo(ddSm0) + ks * 128, hi(ddSm0), b_k + ks * 128, hi(dKm0),
                          id_mnmn, ks > 0 ? 1u : acc_j);
          umma_commit(pds_free);
          if (i == p.nqt - 1) umma_commit(&kv_empty[st]);
          if (j == p.nkt - 1) umma_commit(&qdo_empty[i]);
          if (rem == ppi - 1) umma_commit(dq_full);
        }
        __syncwarp();
        SAV_RECB(1, 45);
      };
      if (total_pairs > 0) issue_sdp(0);
      for (int q = 0; q < total_pairs; ++q) {
        // S^T / dP^T of the next pair go out as soon as pair q's TMEM reads are done -- unless the
        // next pair starts a new item whose Q tile still waits for THIS pair's dK MMAs (nqt == 1)
        const bool more = q + 1 < total_pairs;
        const bool early = more && !(p.nqt == 1 && (q + 1) % ppi == 0);
        if (early) issue_sdp(q + 1);
        issue_grads(q);
        if (more && !early) issue_sdp(q + 1);
      }
    }
  }
  else if (warp == 2) {
    // ------------------------------------------------------------------ second MMA issuer: dV, dK
    // (two issuing warps: the waits of one chain overlap the issue of the other)
    {
      const uint64_t dQm0 = make_smem_desc(smem_u32(sQ), BT_BYTES, 1024);     // MN-major B of dK
      const uint64_t ddOm0 = make_smem_desc(smem_u32(sdO), BT_BYTES, 1024);   // MN-major B of dV
      const uint64_t dPt0 = make_smem_desc(smem_u32(sPt), 16, 1024);          // K-major A of dV
      const uint64_t ddSt0 = make_smem_desc(smem_u32(sdSt), 16, 1024);        // K-major A of dK
      constexpr uint32_t id_kmn = make_idesc_bf16(BT, BD, false, true);
      constexpr uint32_t TILE16 = BT_BYTES >> 4;
      auto lo = [](uint64_t d) { return (uint32_t)d; };
      auto hi = [](uint64_t d) { return (uint32_t)(d >> 32); };
      for (int q = 0; q < total_pairs; ++q) {
        const int itl = q / ppi, rem = q - itl * ppi;
        const int j = rem / p.nqt, i = rem - j * p.nqt;
        const uint32_t gk = (uint32_t)(itl * p.nkt + j);
        mbar_wait_sleep(pds_full, q & 1);
        if (i == 0 && gk > 0) mbar_wait_sleep(dkv_free, (gk - 1) & 1);
        tc_fence_after();
        const int nq_steps = ((min(BT, p.Nq - i * BT) + 15) & ~15) >> 4;
        const uint32_t b_do = lo(ddOm0) + i * TILE16, b_q = lo(dQm0) + i * TILE16;
        const uint32_t acc_i = i > 0;
        // dV[keys,d] += P^T[keys,q] dO_i[q,d];  dK[keys,d] += dS^T[keys,q] Q_i[q,d]   (K = queries)
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            if (ks < nq_steps)
              umma_f16_lh(tdV, lo(dPt0) + (ks >> 2) * TILE16 + (ks & 3) * 2, hi(dPt0), b_do + ks * 128,
                          hi(ddOm0), id_kmn, ks > 0 ? 1u : acc_i);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            if (ks < nq_steps)
              umma_f16_lh(tdK, lo(ddSt0) + (ks >> 2) * TILE16 + (ks & 3) * 2, hi(ddSt0), b_q + ks * 128,
                          hi(dQm0), id_kmn, ks > 0 ? 1u : acc_i);
          umma_commit(pds_free);
          if (i == p.nqt - 1) umma_commit(dkv_full);
          if (j == p.nkt - 1) umma_commit(&qdo_empty[i]);
        }
        __syncwarp();
      }
    }
  }
  else if (warp >= 4) {
    // ------------------------------------------------------------------ compute warps
    const int sub = warp & 3;              // TMEM sub-partition
    const int hf = (warp - 4) >> 2;        // which 64 query columns of the pair / which output
    const int r = sub * 32 + lane;         // key row inside the tile (and q row for the dQ read-out)
    const int ct = tid - 128;              // 0..255
    const bool leader = (ct == 0);
    const uint32_t lane_off = (uint32_t)(sub * 32) << 16;
    const uint64_t sc2 = f2_pack(p.scale_log2, p.scale_log2);
    uint32_t gk = 0;
    int q = 0;
    int dbg_n = (lane == 0 && sub == 0) ? 0 : 1000;
    for (int itl = 0; itl < my_items; ++itl) {
      const int item = blockIdx.x + itl * gridDim.x;
      const int h = item % p.H, b = item / p.H;
      // per-query statistics of this (b, h): -lse*log2(e) and -delta, -inf / 0 beyond Nq
      named_barrier_sync(1, 256);  // everyone is done with the previous item's values
      {
        const int64_t si = ((int64_t)b * p.H + h) * p.Nq + ct;
        s_nl2[ct] = (ct < p.Nq) ? -p.lse[si] * B_LOG2E : -INFINITY;
        s_nd[ct] = (ct < p.Nq) ? -p.delta[si] : 0.f;
      }
      named_barrier_sync(1, 256);
      for (int j = 0; j < p.nkt; ++j, ++gk) {
        for (int i = 0; i < p.nqt; ++i, ++q) {
          const int nq_cur = (min(BT, p.Nq - i * BT) + 15) & ~15;
          const int ncols = max(0, min(64, nq_cur - hf * 64));  // my share of the query columns
          SAV_RECB(3 + hf, 50);
          mbar_wait(st_full, q & 1);
          SAV_RECB(3 + hf, 51);
          tc_fence_after();
          uint32_t pk_p[32], pk_ds[32];
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            const int c0 = half * 32;
            if (c0 < ncols) {
              uint32_t rs[32], rp[32];
              tmem_ld32(tSt + lane_off + hf * 64 + c0, rs);
              tmem_ld32(tdPt + lane_off + hf * 64 + c0, rp);
              tmem_ld_wait();
              const float* nl = s_nl2 + i * BT + hf * 64 + c0;
              const float* nd = s_nd + i * BT + hf * 64 + c0;
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
                pk_p[(c0 + e) >> 1] = pack_bf16(p0, p1);
                pk_p[((c0 + e) >> 1) + 1] = pack_bf16(p2, p3);
                pk_ds[(c0 + e) >> 1] = pack_bf16(s0, s1);
                pk_ds[((c0 + e) >> 1) + 1] = pack_bf16(s2, s3);
              }
            }
          }
          tc_fence_before();
          SAV_RECB(3 + hf, 52);
          mbar_arrive(st_free);  // S^T / dP^T of the next pair may overwrite TMEM
          if (q > 0) mbar_wait(pds_free, (q - 1) & 1);  // previous pair's MMAs are done with smem
          SAV_RECB(3 + hf, 53);
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            if (half * 32 < ncols) {
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                const uint32_t off = bsw128(r, half * 4 + g);
                const int k4 = half * 16 + g * 4;
                *reinterpret_cast<uint4*>(sPt + hf * BT_BYTES + off) =
                    make_uint4(pk_p[k4], pk_p[k4 + 1], pk_p[k4 + 2], pk_p[k4 + 3]);
                *reinterpret_cast<uint4*>(sdSt + hf * BT_BYTES + off) =
                    make_uint4(pk_ds[k4], pk_ds[k4 + 1], pk_ds[k4 + 2], pk_ds[k4 + 3]);
              }
            }
          }
          fence_proxy_async_smem();
          SAV_RECB(3 + hf, 54);
          mbar_arrive(pds_full);
        }
        // ---- key tile j finished: dV_j (warps 4-7) and dK_j (warps 8-11) out of TMEM
        mbar_wait(dkv_full, gk & 1);
        SAV_RECB(3 + hf, 55);
        tc_fence_after();
        {
          uint32_t r0[32], r1[32];
          const uint32_t src = (hf ? tdK : tdV) + lane_off;
          tmem_ld32(src, r0);
          tmem_ld32(src + 32, r1);
          tmem_ld_wait();
          tc_fence_before();
          mbar_arrive(dkv_free);
          const float f = hf ? p.scale : 1.f;
          uint8_t* stage = (hf ? sdSt : sPt);  // all MMAs reading these buffers have completed
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            *reinterpret_cast<uint4*>(stage + bsw128(r, g)) = make_uint4(
                pack_bf16(__uint_as_float(r0[g * 8 + 0]) * f, __uint_as_float(r0[g * 8 + 1]) * f),
                pack_bf16(__uint_as_float(r0[g * 8 + 2]) * f, __uint_as_float(r0[g * 8 + 3]) * f),
                pack_bf16(__uint_as_float(r0[g * 8 + 4]) * f, __uint_as_float(r0[g * 8 + 5]) * f),
                pack_bf16(__uint_as_float(r0[g * 8 + 6]) * f, __uint_as_float(r0[g * 8 + 7]) * f));
            *reinterpret_cast<uint4*>(stage + bsw128(r, 4 + g)) = make_uint4(
                pack_bf16(__uint_as_float(r1[g * 8 + 0]) * f, __uint_as_float(r1[g * 8 + 1]) * f),
                pack_bf16(__uint_as_float(r1[g * 8 + 2]) * f, __uint_as_float(r1[g * 8 + 3]) * f),
                pack_bf16(__uint_as_float(r1[g * 8 + 4]) * f, __uint_as_float(r1[g * 8 + 5]) * f),
                pack_bf16(__uint_as_float(r1[g * 8 + 6]) * f, __uint_as_float(r1[g * 8 + 7]) * f));
          }
        }
        fence_proxy_async_smem();
        named_barrier_sync(1, 256);
        if (leader) {
          tma_store_3d(&tdv, sPt, h * BD, j * BT, b);   // key rows >= Nk are clipped by the map
          tma_store_3d(&tdk, sdSt, h * BD, j * BT, b);
          tma_store_commit();
          tma_store_wait_read<0>();
        }
        named_barrier_sync(1, 256);  // the staging buffers may be overwritten again
        SAV_RECB(3 + hf, 56);
      }
      // ---- item finished: dQ_0 (warps 4-7) and dQ_1 (warps 8-11) out of TMEM
      mbar_wait(dq_full, itl & 1);
      SAV_RECB(3 + hf, 57);
      tc_fence_after();
      if (hf < p.nqt) {
        uint32_t r0[32], r1[32];
        const uint32_t src = tdQ + hf * BD + lane_off;
        tmem_ld32(src, r0);
        tmem_ld32(src + 32, r1);
        tmem_ld_wait();
        const float f = p.scale;
        uint8_t* stage = sPt + hf * BT_BYTES;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          *reinterpret_cast<uint4*>(stage + bsw128(r, g)) = make_uint4(
              pack_bf16(__uint_as_float(r0[g * 8 + 0]) * f, __uint_as_float(r0[g * 8 + 1]) * f),
              pack_bf16(__uint_as_float(r0[g * 8 + 2]) * f, __uint_as_float(r0[g * 8 + 3]) * f),
              pack_bf16(__uint_as_float(r0[g * 8 + 4]) * f, __uint_as_float(r0[g * 8 + 5]) * f),
              pack_bf16(__uint_as_float(r0[g * 8 + 6]) * f, __uint_as_float(r0[g * 8 + 7]) * f));
          *reinterpret_cast<uint4*>(stage + bsw128(r, 4 + g)) = make_uint4(
              pack_bf16(__uint_as_float(r1[g * 8 + 0]) * f, __uint_as_float(r1[g * 8 + 1]) * f),
              pack_bf16(__uint_as_float(r1[g * 8 + 2]) * f, __uint_as_float(r1[g * 8 + 3]) * f),
              pack_bf16(__uint_as_float(r1[g * 8 + 4]) * f, __uint_as_float(r1[g * 8 + 5]) * f),
              pack_bf16(__uint_as_float(r1[g * 8 + 6]) * f, __uint_as_float(r1[g * 8 + 7]) * f));
        }
      }
      tc_fence_before();
      mbar_arrive(dq_free);
      fence_proxy_async_smem();
      named_barrier_sync(1, 256);
      if (leader) {
        for (int i = 0; i < p.nqt; ++i)
          tma_store_3d(&tdq, sPt + i * BT_BYTES, h * BD, i * BT, b);  // rows >= Nq clipped
        tma_store_commit();
        tma_store_wait_read<0>();
      }
      named_barrier_sync(1, 256);
    }
    if (leader) tma_store_wait<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 3) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

// delta[b,h,q] = sum_d dO[b,q,h,d] * (O + O_lo)[b,q,h,d]: 8 lanes per (b,q,h) row, 16-byte loads
__global__ void attn_delta2_kernel(const bf16* __restrict__ o, const bf16* __restrict__ o_lo,
                                   const bf16* __restrict__ d_o, float* __restrict__ delta, int B,
                                   int H, int Nq, int64_t o_bs, int64_t o_rs) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t rowh = t >> 3;  // (b, q, h) index, h fastest
  const int part = (int)(t & 7);
  const bool ok = rowh < (int64_t)B * Nq * H;
  float s = 0.f;
  if (ok) {
    const int hh = (int)(rowh % H);
    const int q = (int)((rowh / H) % Nq);
    const int b = (int)(rowh / ((int64_t)H * Nq));
    const int64_t off = (int64_t)b * o_bs + (int64_t)q * o_rs + hh * BD + part * 8;
    const uint4 a = *reinterpret_cast<const uint4*>(o + off);
    const uint4 g = *reinterpret_cast<const uint4*>(d_o + off);
    uint4 l = make_uint4(0, 0, 0, 0);
    if (o_lo) l = *reinterpret_cast<const uint4*>(o_lo + off);
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, gw[4] = {g.x, g.y, g.z, g.w}, lw[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 af = unpack_bf16(aw[k]), gf = unpack_bf16(gw[k]), lf = unpack_bf16(lw[k]);
      s += (af.x + lf.x) * gf.x + (af.y + lf.y) * gf.y;
    }
  }
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  if (ok && part == 0) {
    const int hh = (int)(rowh % H);
    const int q = (int)((rowh / H) % Nq);
    const int b = (int)(rowh / ((int64_t)H * Nq));
    delta[((int64_t)b * H + hh) * Nq + q] = s;
  }
}

constexpr int kBwd2Smem = 12 * BT_BYTES + 4 * BT * 4 + 1024 + 256;

static int make_bnhd_map128(CUtensorMap* m, const void* base, int64_t H, int64_t N, int64_t B,
                            int64_t bs, int64_t rs) {
  uint64_t dims[3] = {(uint64_t)(H * BD), (uint64_t)N, (uint64_t)B};
  uint64_t str[3] = {0, (uint64_t)rs * 2, (uint64_t)bs * 2};
  uint32_t box[3] = {BD, BT, 1};
  return make_tmap_bf16(m, base, 3, dims, str, box);
}

// Nq <= 256 path of sav_attn_bwd (called from attn.cu)
int launch_attn_bwd2(const sav_attn_bwd_args* a, cudaStream_t st) {
  CUtensorMap tq, tk, tv, tdo, tdq, tdk, tdv;
  int rc;
  if ((rc = make_bnhd_map128(&tq, a->q, a->H, a->Nq, a->B, a->q_batch_stride, a->q_row_stride))) return rc;
  if ((rc = make_bnhd_map128(&tk, a->k, a->H, a->Nk, a->B, a->k_batch_stride, a->k_row_stride))) return rc;
  if ((rc = make_bnhd_map128(&tv, a->v, a->H, a->Nk, a->B, a->v_batch_stride, a->v_row_stride))) return rc;
  if ((rc = make_bnhd_map128(&tdo, a->d_o, a->H, a->Nq, a->B, a->o_batch_stride, a->o_row_stride))) return rc;
  if ((rc = make_bnhd_map128(&tdq, a->dq, a->H, a->Nq, a->B, a->q_batch_stride, a->q_row_stride))) return rc;
  if ((rc = make_bnhd_map128(&tdk, a->dk, a->H, a->Nk, a->B, a->k_batch_stride, a->k_row_stride))) return rc;
  if ((rc = make_bnhd_map128(&tdv, a->dv, a->H, a->Nk, a->B, a->v_batch_stride, a->v_row_stride))) return rc;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(attn_bwd2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwd2Smem);
    if (e != cudaSuccess) return set_error(SAV_ERR_CUDA, "attn bwd2 smem attr: %s", cudaGetErrorString(e));
    attr = true;
  }
  const int64_t rows8 = a->B * a->Nq * a->H * 8;
  attn_delta2_kernel<<<(unsigned)((rows8 + 255) / 256), 256, 0, st>>>(
      reinterpret_cast<const bf16*>(a->o), reinterpret_cast<const bf16*>(a->o_lo),
      reinterpret_cast<const bf16*>(a->d_o), a->delta, (int)a->B, (int)a->H, (int)a->Nq,
      a->o_batch_stride, a->o_row_stride);
  count_launch();
  AttnBwd2Params p;
  p.lse = a->lse; p.delta = a->delta;
  p.B = (int)a->B; p.H = (int)a->H; p.Nq = (int)a->Nq; p.Nk = (int)a->Nk;
  p.n_items = p.B * p.H;
  p.nkt = (p.Nk + BT - 1) / BT;
  p.nqt = (p.Nq + BT - 1) / BT;
  p.scale = a->scale; p.scale_log2 = a->scale * B_LOG2E;
  p.dbg = reinterpret_cast<long long*>(sav_debug_buffer());
  int sms = 148, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = p.n_items < sms ? p.n_items : sms;
  attn_bwd2_kernel<<<grid, kBwdThreads, kBwd2Smem, st>>>(tq, tk, tv, tdo, tdq, tdk, tdv, p);
  count_launch();
  return check_launch("sav_attn_bwd (resident-Q path)");
}

}  // namespace sav
