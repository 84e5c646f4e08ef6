// Fused attention forward for sm_100a (attention.py:39-58): S = (q/sqrt(d)) k^T, softmax over
// keys, O = P v.  S, P and the per-block O never leave the SM.
//
// Persistent, warp-specialised, one CTA per SM, 512 threads (384 when Nk > 256):
//   warp 0        TMA producer: Q tiles (double-buffered per softmax group) and a 2-stage ring of
//                 K/V blocks of up to 256 keys, through rank-3 tensor maps over the strided
//                 [B, N, H, d] views of the fused QKV buffer (out-of-range tokens zero-fill)
//   warps 1, 2    MMA issuers, one per softmax group (one elected lane each): S = Q K^T into the
//                 group's TMEM region, then O_blk = P V with P read straight from TMEM
//   warp 3        TMEM allocator
//   warps 4-7     softmax group 0: q-tile 2i   of every work item } thread t owns query row t:
//   warps 8-11    softmax group 1: q-tile 2i+1 of every work item } fp32 online softmax out of TMEM
//   warps 12-15   epilogue (single-key-block shapes, Nk <= 256): O out of TMEM -> 1/l -> swizzled
//                 smem slab -> TMA store of O and of its bf16 remainder O_lo, lse; the softmax
//                 warps never wait for P V and go straight to the next tile
// A work item is (b, h, pair of 128-row q-tiles); both groups share the K/V blocks in smem and run
// unsynchronised, so one group's exp2 (MUFU -- the bound of this kernel at d = 64) fills the other
// group's MMA / TMEM round trips.  P (bf16) overwrites S in TMEM and is the A operand of the PV MMA
// (no shared-memory round trip); the block's O lands in the dead tail of the S region.
// TMEM region of a group (256 columns): S fp32 [0,256) -> P bf16x2 [0,128) -> O_blk fp32 [128,192).
#include "common.cuh"
#include "host.h"

namespace sav {

constexpr int FQ = 128;            // query rows per tile (UMMA M)
constexpr int FK = 256;            // keys per block (max UMMA N, one TMA box)
constexpr int FD = 64;             // head dim
constexpr int FQ_BYTES = FQ * FD * 2;   // 16 KB
constexpr int FKV_BYTES = FK * FD * 2;  // 32 KB
constexpr float F_LOG2E = 1.4426950408889634f;
constexpr float F_LN2 = 0.6931471805599453f;

struct AttnFwd2Params {
  bf16* o;
  bf16* o_lo;
  float* lse;
  int B, H, Nq, Nk;
  int q_pairs;      // ceil(ceil(Nq/128) / 2)
  int n_items;      // B * H * q_pairs
  int nkb;          // ceil(Nk / 256)
  int64_t o_bs, o_rs;
  float scale_log2;  // scale * log2(e)
  long long* dbg;    // optional timeline buffer (development): [role][512] (event, clock) pairs of CTA 0
};

#define SAV_REC(role, ev)                                                         \
  do {                                                                            \
    if (p.dbg && blockIdx.x == 0 && dbg_n < 255) {                                \
      p.dbg[(role) * 512 + 2 * dbg_n] = (ev);                                     \
      p.dbg[(role) * 512 + 2 * dbg_n + 1] = clock64();                            \
      ++dbg_n;                                                                    \
    }                                                                             \
  } while (0)

// One key block of the online softmax for the 32 query rows of this warp (thread = row):
// S (fp32, TMEM columns [0, n_cur)) -> P (bf16 pairs, TMEM columns [0, n_cur/2)).
// Returns alpha = 2^(m_old - m_new); updates m_run / l_run.
__device__ __forceinline__ float softmax_block(uint32_t region, int n_valid, int n_cur,
                                               float scale_log2, float& m_run, float& l_run) {
  const int n_full = n_valid & ~31;  // columns covered by unmasked 32-wide chunks
  // Both passes are software-pipelined over two register buffers: the tcgen05.ld of chunk i+1 is
  // in flight while chunk i is processed (tcgen05.wait::ld waits for ALL outstanding loads, so
  // the order is wait -> issue next -> compute).
  uint32_t ra[32], rb[32];
  // pass 1: row max over the valid keys (4 independent FMNMX3 chains)
  float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
  auto max_chunk = [&](const uint32_t (&r)[32]) {
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
      mx0 = fmax3(mx0, __uint_as_float(r[i]), __uint_as_float(r[i + 1]));
      mx1 = fmax3(mx1, __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
      mx2 = fmax3(mx2, __uint_as_float(r[i + 4]), __uint_as_float(r[i + 5]));
      mx3 = fmax3(mx3, __uint_as_float(r[i + 6]), __uint_as_float(r[i + 7]));
    }
  };
  if (n_full > 0) tmem_ld32(region, ra);
  for (int c0 = 0; c0 < n_full; c0 += 64) {
    tmem_ld_wait();
    if (c0 + 32 < n_full) tmem_ld32(region + c0 + 32, rb);
    max_chunk(ra);
    if (c0 + 32 < n_full) {
      tmem_ld_wait();
      if (c0 + 64 < n_full) tmem_ld32(region + c0 + 64, ra);
      max_chunk(rb);
    }
  }
  float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
  for (int c0 = n_full; c0 < n_cur; c0 += 16) {  // masked tail, 16 columns at a time
    uint32_t r[16];
    tmem_ld16(region + c0, r);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 16; ++i)
      if (c0 + i < n_valid) mx = fmaxf(mx, __uint_as_float(r[i]));
  }
  if (n_full > 0) tmem_ld32(region, ra);  // first chunk of pass 2
  const float m_new = fmaxf(m_run, mx * scale_log2);
  const float alpha = ex2_fast(m_run - m_new);  // 0 on the first block
  // pass 2: p = exp2(s*c - m), row sum, bf16 pairs back into TMEM over the dead S columns
  // (packed f32x2 FMA / ADD: 2.5 issue slots per element, so the MUFU is the bound)
  const uint64_t sc2 = f2_pack(scale_log2, scale_log2);
  const uint64_t nm2 = f2_pack(-m_new, -m_new);
  uint64_t sum2a = f2_pack(0.f, 0.f), sum2b = f2_pack(0.f, 0.f);
  auto exp_chunk = [&](const uint32_t (&r)[32], int c0) {
    uint32_t pk[16];
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      float t0, t1, t2, t3;
      f2_unpack(f2_fma(f2_pack_u(r[i], r[i + 1]), sc2, nm2), t0, t1);
      f2_unpack(f2_fma(f2_pack_u(r[i + 2], r[i + 3]), sc2, nm2), t2, t3);
      const float e0 = ex2_fast(t0), e1 = ex2_fast(t1), e2 = ex2_fast(t2), e3 = ex2_fast(t3);
      sum2a = f2_add(sum2a, f2_pack(e0, e1));
      sum2b = f2_add(sum2b, f2_pack(e2, e3));
      pk[i >> 1] = pack_bf16(e0, e1);
      pk[(i >> 1) + 1] = pack_bf16(e2, e3);
    }
    tmem_st16(region + (c0 >> 1), pk);
  };
  for (int c0 = 0; c0 < n_full; c0 += 64) {
    tmem_ld_wait();
    if (c0 + 32 < n_full) tmem_ld32(region + c0 + 32, rb);
    exp_chunk(ra, c0);
    if (c0 + 32 < n_full) {
      tmem_ld_wait();
      if (c0 + 64 < n_full) tmem_ld32(region + c0 + 64, ra);
      exp_chunk(rb, c0 + 32);
    }
  }
  const uint64_t sum2 = f2_add(sum2a, sum2b);
  float psum = 0.f;
  for (int c0 = n_full; c0 < n_cur; c0 += 16) {
    uint32_t r[16];
    tmem_ld16(region + c0, r);
    tmem_ld_wait();
    uint32_t pk[8];
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
      float e0 = ex2_fast(__uint_as_float(r[i]) * scale_log2 - m_new);
      float e1 = ex2_fast(__uint_as_float(r[i + 1]) * scale_log2 - m_new);
      e0 = (c0 + i < n_valid) ? e0 : 0.f;
      e1 = (c0 + i + 1 < n_valid) ? e1 : 0.f;
      psum += e0 + e1;
      pk[i >> 1] = pack_bf16(e0, e1);
    }
    tmem_st8(region + (c0 >> 1), pk);
  }
  float s0, s1;
  f2_unpack(sum2, s0, s1);
  psum += s0 + s1;
  tmem_st_wait();
  l_run = l_run * alpha + psum;
  m_run = m_new;
  return alpha;
}

// 16-byte chunk `c` (0..7) of row `r` of a [128 x 128 B] slab laid out for a SWIZZLE_128B TMA box
__device__ __forceinline__ uint32_t fslab_off(int r, int c) {
  return (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4));
}

template <bool SINGLE>
__global__ void __launch_bounds__(SINGLE ? 512 : 384, 1)
attn_fwd2_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                 const __grid_constant__ CUtensorMap tv, const __grid_constant__ CUtensorMap to,
                 const __grid_constant__ CUtensorMap tlo, const AttnFwd2Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sQ = smem;                        // [group][buf] 4 x 16 KB
  uint8_t* sKV = sQ + 4 * FQ_BYTES;          // [stage] { K 32 KB, V 32 KB }
  uint8_t* sOut = sKV + 4 * FKV_BYTES;       // SINGLE: O slab 16 KB, O_lo slab 16 KB
  float* s_stat = reinterpret_cast<float*>(sOut + (SINGLE ? 2 * FQ_BYTES : 0));  // [group][128] row sums l
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_stat + 2 * FQ);
  uint64_t* q_full = bars;        // [g*2+buf]
  uint64_t* q_empty = bars + 4;   // [g*2+buf]
  uint64_t* kv_full = bars + 8;   // [stage]
  uint64_t* kv_empty = bars + 10; // [stage]  (2 arrivals: one per MMA issuer)
  uint64_t* s_full = bars + 12;   // [g]
  uint64_t* p_full = bars + 14;   // [g]  (128 arrivals)
  uint64_t* o_full = bars + 16;   // [g]
  uint64_t* s_free = bars + 18;   // [g]  (128 arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);

  // warp index through a shuffle: the role branches are then provably warp-uniform for ptxas
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const int q_tiles = (p.Nq + FQ - 1) / FQ;

  if (tid == 0) {
    tma_prefetch_desc(&tq);
    tma_prefetch_desc(&tk);
    tma_prefetch_desc(&tv);
    if (SINGLE) { tma_prefetch_desc(&to); tma_prefetch_desc(&tlo); }
    for (int i = 0; i < 4; ++i) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 2);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 128);
      mbar_init(&o_full[i], 1);
      mbar_init(&s_free[i], 128);
    }
    fence_barrier_init();
  }
  if (warp == 3) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int dbg_n = 0;
      uint32_t uq[2] = {0, 0};  // Q tiles loaded so far per group (buffer = uq & 1)
      uint32_t g = 0;           // K/V blocks loaded so far (stage = g & 1)
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
        const int qp = item % p.q_pairs;
        const int bh = item / p.q_pairs;
        const int h = bh % p.H, b = bh / p.H;
        for (int w = 0; w < 2; ++w) {
          const int qt = qp * 2 + w;
          if (qt < q_tiles) {
            const int qb = w * 2 + (uq[w] & 1);
            mbar_wait_sleep(&q_empty[qb], ((uq[w] >> 1) & 1) ^ 1);
            mbar_expect_tx(&q_full[qb], FQ_BYTES);
            tma_load_3d(sQ + qb * FQ_BYTES, &tq, &q_full[qb], h * FD, qt * FQ, b);
            ++uq[w];
          }
        }
        for (int j = 0; j < p.nkb; ++j, ++g) {
          const int st = g & 1;
          SAV_REC(0, 1);
          mbar_wait_sleep(&kv_empty[st], ((g >> 1) & 1) ^ 1);
          SAV_REC(0, 2);
          mbar_expect_tx(&kv_full[st], 2 * FKV_BYTES);
          tma_load_3d(sKV + st * 2 * FKV_BYTES, &tk, &kv_full[st], h * FD, j * FK, b);
          tma_load_3d(sKV + st * 2 * FKV_BYTES + FKV_BYTES, &tv, &kv_full[st], h * FD, j * FK, b);
        }
      }
    }
  } else if (warp == 1 || warp == 2) {
    // ------------------------------------------------------------------ MMA issuer of group w
    // All 32 lanes walk the loop and poll the barriers; one elected lane issues.  With warp-uniform
    // control flow the descriptors live in uniform registers; issuing from `if (lane == 0)` wraps
    // every UTCHMMA in a vote / R2UR waterfall (~150 cycles per MMA, several times its run time).
    {
      const int w = warp - 1;
      int dbg_n = lane == 0 ? 0 : 1000;
      const uint32_t region = tmem + w * 256;
      uint32_t uq = 0, g = 0, c = 0;  // Q tiles used, kv blocks seen, blocks processed by this group
      // descriptors are built once; per MMA only the 14-bit start-address field advances
      const uint64_t dq0 = make_smem_desc(smem_u32(sQ + (w * 2) * FQ_BYTES), 16, 1024);
      const uint64_t dk0 = make_smem_desc(smem_u32(sKV), 16, 1024);
      const uint64_t dv0 = make_smem_desc(smem_u32(sKV + FKV_BYTES), FQ_BYTES, 1024);
      constexpr uint32_t id_pv = make_idesc_bf16(FQ, FD, false, true);  // B = V is MN-major
      // One-time stagger: group 1 starts half a period late (when group 0 has finished its first
      // softmax), so that afterwards one group's exp2 phase overlaps the other group's MMA / TMEM
      // round trips instead of both groups idling and computing in lockstep.
      if (w == 1 && q_tiles > 1) mbar_wait_sleep(&s_free[0], 0);
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
        const int qp = item % p.q_pairs;
        const bool active = (qp * 2 + w) < q_tiles;
        const int qb = w * 2 + (uq & 1);
        const uint32_t q_phase = (uq >> 1) & 1;
        const uint64_t dq = dq0 + (uint64_t)((uq & 1) * (FQ_BYTES >> 4));
        if (active) ++uq;
        for (int j = 0; j < p.nkb; ++j, ++g) {
          const int st = g & 1;
          SAV_REC(1 + w, 10);
          mbar_wait_sleep(&kv_full[st], (g >> 1) & 1);
          SAV_REC(1 + w, 11);
          if (!active) {
            if (elect_one()) mbar_arrive(&kv_empty[st]);
            __syncwarp();
            continue;
          }
          if (j == 0) mbar_wait_sleep(&q_full[qb], q_phase);
          mbar_wait_sleep(&s_free[w], (c & 1) ^ 1);
          SAV_REC(1 + w, 12);
          tc_fence_after();
          const int n_valid = min(FK, p.Nk - j * FK);
          const int n_cur = (n_valid + 15) & ~15;
          const uint64_t dk = dk0 + (uint64_t)(st * ((2 * FKV_BYTES) >> 4));
          const uint64_t dv = dv0 + (uint64_t)(st * ((2 * FKV_BYTES) >> 4));
          const uint32_t id_s = make_idesc_bf16(FQ, n_cur, false, false);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < FD / 16; ++k)  // +32 B per 16-element K step
              umma_f16_lh(region, (uint32_t)dq + k * 2, (uint32_t)(dq >> 32), (uint32_t)dk + k * 2,
                          (uint32_t)(dk >> 32), id_s, k > 0);
            umma_commit(&s_full[w]);
            if (j == p.nkb - 1) umma_commit(&q_empty[qb]);
          }
          __syncwarp();
          SAV_REC(1 + w, 13);
          mbar_wait_sleep(&p_full[w], c & 1);
          SAV_REC(1 + w, 14);
          tc_fence_after();
          const int ksteps = n_cur >> 4;
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < FK / 16; ++ks)  // +8 TMEM columns of P, +2048 B of V per 16 keys
              if (ks < ksteps)
                umma_f16_ts_lh(region + 128, region + ks * 8, (uint32_t)dv + ks * 128, (uint32_t)(dv >> 32),
                               id_pv, ks > 0);
            umma_commit(&o_full[w]);
            umma_commit(&kv_empty[st]);
          }
          __syncwarp();
          SAV_REC(1 + w, 15);
          ++c;
        }
      }
    }
  } else if (warp >= 4 && warp < 12) {
    // ------------------------------------------------------------------ softmax groups
    const int w = (warp - 4) >> 2;
    const int row_t = (warp & 3) * 32 + lane;  // query row inside the tile == TMEM lane
    const uint32_t region = tmem + w * 256 + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t c = 0;
    int dbg_n = (lane == 0 && (warp & 3) == 0) ? 0 : 1000;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
      const int qp = item % p.q_pairs;
      const int qt = qp * 2 + w;
      if (qt >= q_tiles) continue;
      const int bh = item / p.q_pairs;
      const int h = bh % p.H, b = bh / p.H;
      // warps whose 32 rows are all beyond Nq still run the barrier protocol but skip the math
      const bool warp_live = qt * FQ + (warp & 3) * 32 < p.Nq;
      float m_run = -INFINITY, l_run = 0.f;
      if (SINGLE) {
        SAV_REC(3 + w, 20);
        mbar_wait(&s_full[w], c & 1);
        SAV_REC(3 + w, 21);
        tc_fence_after();
        if (warp_live) {
          const int n_cur = (p.Nk + 15) & ~15;
          softmax_block(region, p.Nk, n_cur, p.scale_log2, m_run, l_run);
          s_stat[w * FQ + row_t] = l_run;
          const int row = qt * FQ + row_t;
          if (row < p.Nq) p.lse[((int64_t)b * p.H + h) * p.Nq + row] = m_run * F_LN2 + logf(l_run);
        }
        tc_fence_before();
        SAV_REC(3 + w, 23);
        mbar_arrive(&p_full[w]);
        ++c;
        continue;  // the epilogue warps take O from here
      }
      float o_acc[FD];
#pragma unroll
      for (int i = 0; i < FD; ++i) o_acc[i] = 0.f;
      for (int j = 0; j < p.nkb; ++j, ++c) {
        const int n_valid = min(FK, p.Nk - j * FK);
        const int n_cur = (n_valid + 15) & ~15;
        mbar_wait(&s_full[w], c & 1);
        tc_fence_after();
        float alpha = 1.f;
        if (warp_live) alpha = softmax_block(region, n_valid, n_cur, p.scale_log2, m_run, l_run);
        tc_fence_before();
        mbar_arrive(&p_full[w]);
        mbar_wait(&o_full[w], c & 1);
        tc_fence_after();
        if (warp_live) {
#pragma unroll
          for (int c0 = 0; c0 < FD; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(region + 128 + c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o_acc[c0 + i] = o_acc[c0 + i] * alpha + __uint_as_float(r[i]);
          }
        }
        tc_fence_before();
        mbar_arrive(&s_free[w]);
      }
      const int row = qt * FQ + row_t;
      if (row < p.Nq) {
        const float inv = 1.f / l_run;
        const int64_t off = (int64_t)b * p.o_bs + (int64_t)row * p.o_rs + h * FD;
        bf16* dst = p.o + off;
#pragma unroll
        for (int i = 0; i < FD; ++i) o_acc[i] *= inv;
#pragma unroll
        for (int g8 = 0; g8 < FD / 8; ++g8)
          *reinterpret_cast<uint4*>(dst + g8 * 8) = make_uint4(
              pack_bf16(o_acc[g8 * 8 + 0], o_acc[g8 * 8 + 1]), pack_bf16(o_acc[g8 * 8 + 2], o_acc[g8 * 8 + 3]),
              pack_bf16(o_acc[g8 * 8 + 4], o_acc[g8 * 8 + 5]), pack_bf16(o_acc[g8 * 8 + 6], o_acc[g8 * 8 + 7]));
        if (p.o_lo) {
          bf16* dlo = p.o_lo + off;
#pragma unroll
          for (int i = 0; i < FD; ++i) o_acc[i] -= __bfloat162float(__float2bfloat16(o_acc[i]));
#pragma unroll
          for (int g8 = 0; g8 < FD / 8; ++g8)
            *reinterpret_cast<uint4*>(dlo + g8 * 8) = make_uint4(
                pack_bf16(o_acc[g8 * 8 + 0], o_acc[g8 * 8 + 1]), pack_bf16(o_acc[g8 * 8 + 2], o_acc[g8 * 8 + 3]),
                pack_bf16(o_acc[g8 * 8 + 4], o_acc[g8 * 8 + 5]), pack_bf16(o_acc[g8 * 8 + 6], o_acc[g8 * 8 + 7]));
        }
        p.lse[((int64_t)b * p.H + h) * p.Nq + row] = m_run * F_LN2 + logf(l_run);
      }
    }
  }
  else if (SINGLE && warp >= 12) {
    // ------------------------------------------------------------------ epilogue (Nk <= 256)
    const int row_t = (warp & 3) * 32 + lane;
    const bool leader = (warp == 12 && lane == 0);
    uint32_t cg[2] = {0, 0};
    int dbg_n = (lane == 0 && warp == 12) ? 0 : 1000;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
      const int qp = item % p.q_pairs;
      const int bh = item / p.q_pairs;
      const int h = bh % p.H, b = bh / p.H;
      for (int w = 0; w < 2; ++w) {
        const int qt = qp * 2 + w;
        if (qt >= q_tiles) continue;
        const uint32_t region = tmem + w * 256 + ((uint32_t)((warp & 3) * 32) << 16);
        mbar_wait(&o_full[w], cg[w] & 1);
        ++cg[w];
        SAV_REC(5, 30 + w);
        tc_fence_after();
        uint32_t r0[32], r1[32];
        tmem_ld32(region + 128, r0);
        tmem_ld32(region + 160, r1);
        tmem_ld_wait();
        const float l_row = s_stat[w * FQ + row_t];
        tc_fence_before();
        mbar_arrive(&s_free[w]);  // region (and s_stat) of group w may be overwritten from here
        const float inv = 1.f / l_row;
        // the previous tile's TMA stores must have finished reading the slabs
        if (leader) tma_store_wait_read<0>();
        named_barrier_sync(2, 128);
        // O -> bf16 slab; its rounding remainder (o - bf16(o)) -> second slab: lets backward form
        // delta = rowsum(dO * O) from a ~16-bit mantissa O (the bf16-rounded O alone costs ~0.5e-2
        // of the 2e-2 gradient parity budget)
        auto emit = [&](const uint32_t (&r)[32], int chunk0) {
#pragma unroll
          for (int g8 = 0; g8 < 4; ++g8) {
            float v[8], lo[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              v[i] = __uint_as_float(r[g8 * 8 + i]) * inv;
              lo[i] = v[i] - __bfloat162float(__float2bfloat16(v[i]));
            }
            *reinterpret_cast<uint4*>(sOut + fslab_off(row_t, chunk0 + g8)) =
                make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]),
                           pack_bf16(v[6], v[7]));
            if (p.o_lo)
              *reinterpret_cast<uint4*>(sOut + FQ_BYTES + fslab_off(row_t, chunk0 + g8)) =
                  make_uint4(pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]), pack_bf16(lo[4], lo[5]),
                             pack_bf16(lo[6], lo[7]));
          }
        };
        emit(r0, 0);
        emit(r1, 4);
        fence_proxy_async_smem();
        named_barrier_sync(2, 128);
        if (leader) {
          tma_store_3d(&to, sOut, h * FD, qt * FQ, b);   // rows >= Nq are clipped by the tensor map
          if (p.o_lo) tma_store_3d(&tlo, sOut + FQ_BYTES, h * FD, qt * FQ, b);
          tma_store_commit();
        }
        SAV_REC(5, 32);
      }
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

constexpr int kFwd2SmemMulti = 4 * FQ_BYTES + 4 * FKV_BYTES + 2 * FQ * 4 + 1024 + 256;
static_assert(4 * FQ_BYTES + 4 * FKV_BYTES + 2 * FQ * 4 + 1024 + 256 + 2 * FQ_BYTES <= 232448, "smem budget");
constexpr int kFwd2SmemSingle = kFwd2SmemMulti + 2 * FQ_BYTES;

static int make_bnhd_map(CUtensorMap* m, const void* base, int64_t H, int64_t N, int64_t B,
                         int64_t bs, int64_t rs, uint32_t box_rows) {
  uint64_t dims[3] = {(uint64_t)(H * FD), (uint64_t)N, (uint64_t)B};
  uint64_t str[3] = {0, (uint64_t)rs * 2, (uint64_t)bs * 2};
  uint32_t box[3] = {FD, box_rows, 1};
  return make_tmap_bf16(m, base, 3, dims, str, box);
}

}  // namespace sav

using namespace sav;

extern "C" int sav_attn_fwd(const sav_attn_fwd_args* a, void* stream) {
  if (!a || a->D != FD) return set_error(SAV_ERR_SHAPE, "sav_attn_fwd: head dim must be 64");
  if (a->B <= 0 || a->H <= 0 || a->Nq <= 0 || a->Nk <= 0)
    return set_error(SAV_ERR_SHAPE, "sav_attn_fwd: empty shape");
  if (a->o_row_stride % 8 || a->o_batch_stride % 8 || !aligned16(a->o) || !aligned16(a->o_lo))
    return set_error(SAV_ERR_ALIGN, "sav_attn_fwd: output strides/pointer alignment");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  CUtensorMap tq, tk, tv, to, tlo;
  int rc;
  if ((rc = make_bnhd_map(&tq, a->q, a->H, a->Nq, a->B, a->q_batch_stride, a->q_row_stride, FQ))) return rc;
  if ((rc = make_bnhd_map(&tk, a->k, a->H, a->Nk, a->B, a->k_batch_stride, a->k_row_stride, FK))) return rc;
  if ((rc = make_bnhd_map(&tv, a->v, a->H, a->Nk, a->B, a->v_batch_stride, a->v_row_stride, FK))) return rc;
  if ((rc = make_bnhd_map(&to, a->o, a->H, a->Nq, a->B, a->o_batch_stride, a->o_row_stride, FQ))) return rc;
  tlo = to;
  if (a->o_lo &&
      (rc = make_bnhd_map(&tlo, a->o_lo, a->H, a->Nq, a->B, a->o_batch_stride, a->o_row_stride, FQ)))
    return rc;
  const bool single = a->Nk <= FK;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(attn_fwd2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwd2SmemSingle);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(attn_fwd2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwd2SmemMulti);
    if (e != cudaSuccess) return set_error(SAV_ERR_CUDA, "attn fwd smem attr: %s", cudaGetErrorString(e));
    attr = true;
  }
  AttnFwd2Params p;
  p.o = reinterpret_cast<bf16*>(a->o); p.o_lo = reinterpret_cast<bf16*>(a->o_lo); p.lse = a->lse;
  p.B = (int)a->B; p.H = (int)a->H; p.Nq = (int)a->Nq; p.Nk = (int)a->Nk;
  const int q_tiles = (p.Nq + FQ - 1) / FQ;
  p.q_pairs = (q_tiles + 1) / 2;
  p.n_items = p.B * p.H * p.q_pairs;
  p.nkb = (p.Nk + FK - 1) / FK;
  p.o_bs = a->o_batch_stride; p.o_rs = a->o_row_stride;
  p.scale_log2 = a->scale * F_LOG2E;
  p.dbg = reinterpret_cast<long long*>(sav_debug_buffer());
  int sms = 148, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = p.n_items < sms ? p.n_items : sms;
  if (single)
    attn_fwd2_kernel<true><<<grid, 512, kFwd2SmemSingle, st>>>(tq, tk, tv, to, tlo, p);
  else
    attn_fwd2_kernel<false><<<grid, 384, kFwd2SmemMulti, st>>>(tq, tk, tv, to, tlo, p);
  count_launch();
  return check_launch("sav_attn_fwd");
}
