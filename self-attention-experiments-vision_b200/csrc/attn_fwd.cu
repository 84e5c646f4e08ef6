// Fused attention forward for sm_100a (attention.py:39-58): S = (q/sqrt(d)) k^T, softmax over
// keys, O = P v.  S, P and the per-block O never leave the SM.
//
// Persistent, warp-specialised, one CTA per SM, 384 threads:
//   warp 0        TMA producer: Q tiles (double-buffered per softmax group) and a 2-stage ring of
//                 K/V blocks of up to 256 keys, through rank-3 tensor maps over the strided
//                 [B, N, H, d] views of the fused QKV buffer (out-of-range tokens zero-fill)
//   warps 1, 2    MMA issuers, one per softmax group (one elected lane each): S = Q K^T into the
//                 group's TMEM region, then O_blk = P V with P read straight from TMEM
//   warp 3        TMEM allocator
//   warps 4-7     softmax group 0: q-tile 2i   of every work item } thread t owns query row t:
//   warps 8-11    softmax group 1: q-tile 2i+1 of every work item } fp32 online softmax out of TMEM
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
constexpr int kFwdThreads = 384;
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
};

__global__ void __launch_bounds__(kFwdThreads, 1)
attn_fwd2_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                 const __grid_constant__ CUtensorMap tv, const AttnFwd2Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sQ = smem;                        // [group][buf] 4 x 16 KB
  uint8_t* sKV = sQ + 4 * FQ_BYTES;          // [stage] { K 32 KB, V 32 KB }
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + 4 * FKV_BYTES);
  uint64_t* q_full = bars;        // [g*2+buf]
  uint64_t* q_empty = bars + 4;   // [g*2+buf]
  uint64_t* kv_full = bars + 8;   // [stage]
  uint64_t* kv_empty = bars + 10; // [stage]  (2 arrivals: one per MMA issuer)
  uint64_t* s_full = bars + 12;   // [g]
  uint64_t* p_full = bars + 14;   // [g]  (128 arrivals)
  uint64_t* o_full = bars + 16;   // [g]
  uint64_t* s_free = bars + 18;   // [g]  (128 arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q_tiles = (p.Nq + FQ - 1) / FQ;

  if (tid == 0) {
    tma_prefetch_desc(&tq);
    tma_prefetch_desc(&tk);
    tma_prefetch_desc(&tv);
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
            mbar_wait(&q_empty[qb], ((uq[w] >> 1) & 1) ^ 1);
            mbar_expect_tx(&q_full[qb], FQ_BYTES);
            tma_load_3d(sQ + qb * FQ_BYTES, &tq, &q_full[qb], h * FD, qt * FQ, b);
            ++uq[w];
          }
        }
        for (int j = 0; j < p.nkb; ++j, ++g) {
          const int st = g & 1;
          mbar_wait(&kv_empty[st], ((g >> 1) & 1) ^ 1);
          mbar_expect_tx(&kv_full[st], 2 * FKV_BYTES);
          tma_load_3d(sKV + st * 2 * FKV_BYTES, &tk, &kv_full[st], h * FD, j * FK, b);
          tma_load_3d(sKV + st * 2 * FKV_BYTES + FKV_BYTES, &tv, &kv_full[st], h * FD, j * FK, b);
        }
      }
    }
  } else if (warp == 1 || warp == 2) {
    // ------------------------------------------------------------------ MMA issuer of group w
    if (lane == 0) {
      const int w = warp - 1;
      const uint32_t region = tmem + w * 256;
      uint32_t uq = 0, g = 0, c = 0;  // Q tiles used, kv blocks seen, blocks processed by this group
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
        const int qp = item % p.q_pairs;
        const bool active = (qp * 2 + w) < q_tiles;
        const int qb = w * 2 + (uq & 1);
        const uint32_t q_phase = (uq >> 1) & 1;
        if (active) ++uq;
        for (int j = 0; j < p.nkb; ++j, ++g) {
          const int st = g & 1;
          mbar_wait(&kv_full[st], (g >> 1) & 1);
          if (!active) { mbar_arrive(&kv_empty[st]); continue; }
          if (j == 0) mbar_wait(&q_full[qb], q_phase);
          mbar_wait(&s_free[w], (c & 1) ^ 1);
          tc_fence_after();
          const int n_valid = min(FK, p.Nk - j * FK);
          const int n_cur = (n_valid + 15) & ~15;
          const uint32_t aq = smem_u32(sQ + qb * FQ_BYTES);
          const uint32_t bk = smem_u32(sKV + st * 2 * FKV_BYTES);
          const uint32_t bv = bk + FKV_BYTES;
          const uint32_t id_s = make_idesc_bf16(FQ, n_cur, false, false);
#pragma unroll
          for (int k = 0; k < FD / 16; ++k)
            umma_f16(region, make_smem_desc(aq + k * 32, 16, 1024),
                     make_smem_desc(bk + k * 32, 16, 1024), id_s, k > 0);
          umma_commit(&s_full[w]);
          if (j == p.nkb - 1) umma_commit(&q_empty[qb]);
          mbar_wait(&p_full[w], c & 1);
          tc_fence_after();
          const uint32_t id_pv = make_idesc_bf16(FQ, FD, false, true);  // B = V is MN-major
          for (int ks = 0; ks < n_cur / 16; ++ks)
            umma_f16_ts(region + 128, region + ks * 8, make_smem_desc(bv + ks * 2048, FQ_BYTES, 1024),
                        id_pv, ks > 0);
          umma_commit(&o_full[w]);
          umma_commit(&kv_empty[st]);
          ++c;
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ softmax groups
    const int w = (warp - 4) >> 2;
    const int row_t = (warp & 3) * 32 + lane;  // query row inside the tile == TMEM lane
    const uint32_t region = tmem + w * 256 + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t c = 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
      const int qp = item % p.q_pairs;
      const int qt = qp * 2 + w;
      if (qt >= q_tiles) continue;
      const int bh = item / p.q_pairs;
      const int h = bh % p.H, b = bh / p.H;
      // warps whose 32 rows are all beyond Nq still run the barrier protocol but skip the math
      const bool warp_live = qt * FQ + (warp & 3) * 32 < p.Nq;
      float m_run = -INFINITY, l_run = 0.f;
      float o_acc[FD];
#pragma unroll
      for (int i = 0; i < FD; ++i) o_acc[i] = 0.f;
      for (int j = 0; j < p.nkb; ++j, ++c) {
        const int n_valid = min(FK, p.Nk - j * FK);
        const int n_cur = (n_valid + 15) & ~15;
        mbar_wait(&s_full[w], c & 1);
        tc_fence_after();
        float alpha = 1.f;
        if (warp_live) {
          // pass 1: row max over the valid keys
          float mx = -INFINITY;
          for (int c0 = 0; c0 < n_cur; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(region + c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (c0 + i < n_valid) mx = fmaxf(mx, __uint_as_float(r[i]));
          }
          const float m_new = fmaxf(m_run, mx * p.scale_log2);
          alpha = ex2_fast(m_run - m_new);  // 0 on the first block
          // pass 2: p = exp2(s*c - m), row sum, bf16 pairs back into TMEM over the dead S columns
          float psum = 0.f;
          for (int c0 = 0; c0 < n_cur; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(region + c0, r);
            tmem_ld_wait();
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              float e0 = ex2_fast(__uint_as_float(r[i]) * p.scale_log2 - m_new);
              float e1 = ex2_fast(__uint_as_float(r[i + 1]) * p.scale_log2 - m_new);
              e0 = (c0 + i < n_valid) ? e0 : 0.f;
              e1 = (c0 + i + 1 < n_valid) ? e1 : 0.f;
              psum += e0 + e1;
              pk[i >> 1] = pack_bf16(e0, e1);
            }
            tmem_st16(region + (c0 >> 1), pk);
          }
          tmem_st_wait();
          l_run = l_run * alpha + psum;
          m_run = m_new;
        }
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
          // low half of O (o - bf16(o)): lets backward form delta = rowsum(dO * O) from a ~16-bit
          // mantissa O; the bf16-rounded O alone costs ~0.5e-2 of the 2e-2 gradient parity budget
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
  tc_fence_before();
  __syncthreads();
  if (warp == 3) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

constexpr int kFwd2Smem = 4 * FQ_BYTES + 4 * FKV_BYTES + 1024 + 256;

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
  CUtensorMap tq, tk, tv;
  int rc;
  if ((rc = make_bnhd_map(&tq, a->q, a->H, a->Nq, a->B, a->q_batch_stride, a->q_row_stride, FQ))) return rc;
  if ((rc = make_bnhd_map(&tk, a->k, a->H, a->Nk, a->B, a->k_batch_stride, a->k_row_stride, FK))) return rc;
  if ((rc = make_bnhd_map(&tv, a->v, a->H, a->Nk, a->B, a->v_batch_stride, a->v_row_stride, FK))) return rc;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(attn_fwd2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwd2Smem);
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
  int sms = 148, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = p.n_items < sms ? p.n_items : sms;
  attn_fwd2_kernel<<<grid, kFwdThreads, kFwd2Smem, st>>>(tq, tk, tv, p);
  count_launch();
  return check_launch("sav_attn_fwd");
}
