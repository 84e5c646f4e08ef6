// Fused attention core for sm_100a (attention.py:39-58): S = (q/sqrt(d)) k^T, softmax over keys,
// O = P v, with S/P never leaving the SM.  TMA (3-D tensor maps over the strided [B,N,H,D] views
// of the fused QKV buffer, out-of-range tokens zero-filled) -> tcgen05.mma with accumulators in
// TMEM -> one thread per query row does the fp32 online softmax straight out of TMEM.
// Backward recomputes S^T/P^T per (key block, query block) and forms dV, dK (TMEM-resident across
// the query loop) and dQ (fp32 red.add across key blocks).  Head dim D = 64.
#include "common.cuh"
#include "host.h"

namespace sav {

constexpr int AT = 128;        // query/key block (UMMA M and max UMMA N)
constexpr int AD = 64;         // head dim
constexpr int TILE_B = AT * AD * 2;  // 16 KB: one [128 x 64] bf16 tile, 128-byte rows
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

// byte offset of 16-byte group `g16` (0..7) of row `r` inside a [rows x 128 B] SW128 tile
__device__ __forceinline__ uint32_t sw128(int r, int g16) {
  return (uint32_t)(r * 128 + ((g16 ^ (r & 7)) << 4));
}

// ------------------------------------------------------------------------------- backward
struct AttnBwdParams {
  const float* lse;
  const float* delta;
  float* dq_accum;  // fp32 [B,Nq,H,D]
  bf16* dk;
  bf16* dv;
  int B, H, Nq, Nk;
  int64_t k_bs, k_rs, v_bs, v_rs;
  float scale, scale_log2;
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b),
               "f"(c), "f"(d)
               : "memory");
}

// grid (k_tiles, H, B); 128 threads; thread t owns KEY row t of the block: the kernel works on
// S^T = K Q^T and dP^T = V dO^T so that P^T / dS^T rows are written K-major for the dV / dK MMAs
// and the very same dS^T buffer is the MN-major A operand of dQ = dS K.
__global__ void __launch_bounds__(128)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                const __grid_constant__ CUtensorMap tv, const __grid_constant__ CUtensorMap tdo,
                const AttnBwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sK = smem;                  // 16 KB
  uint8_t* sV = sK + TILE_B;           // 16 KB
  uint8_t* sQ = sV + TILE_B;           // 2 x 16 KB
  uint8_t* sdO = sQ + 2 * TILE_B;      // 2 x 16 KB
  uint8_t* sPt = sdO + 2 * TILE_B;     // 2 chunks x 16 KB: [128 keys] x [64 q] per chunk
  uint8_t* sdSt = sPt + 2 * TILE_B;    // 2 chunks x 16 KB
  float* s_lse = reinterpret_cast<float*>(sdSt + 2 * TILE_B);  // [128] (log2 domain)
  float* s_delta = s_lse + AT;                                   // [128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_delta + AT);
  uint64_t* kv_full = bars;
  uint64_t* qdo_full = bars + 1;  // [2]
  uint64_t* s_full = bars + 3;
  uint64_t* mma_done = bars + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5);

  // warp index through a shuffle: `warp == 0` branches are then provably warp-uniform
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int kt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int nq_blk = (p.Nq + AT - 1) / AT;

  if (tid == 0) {
    tma_prefetch_desc(&tq);
    tma_prefetch_desc(&tk);
    tma_prefetch_desc(&tv);
    tma_prefetch_desc(&tdo);
    mbar_init(kv_full, 1);
    mbar_init(&qdo_full[0], 1);
    mbar_init(&qdo_full[1], 1);
    mbar_init(s_full, 1);
    mbar_init(mma_done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tSt = tmem, tdPt = tmem + 128, tdV = tmem + 256, tdK = tmem + 320,
                 tdQ = tmem + 384;
  const uint32_t lane_off = (uint32_t)(warp * 32) << 16;

  if (tid == 0) {
    mbar_expect_tx(kv_full, 2 * TILE_B);
    tma_load_3d(sK, &tk, kv_full, h * AD, kt * AT, b);
    tma_load_3d(sV, &tv, kv_full, h * AD, kt * AT, b);
    mbar_expect_tx(&qdo_full[0], 2 * TILE_B);
    tma_load_3d(sQ, &tq, &qdo_full[0], h * AD, 0, b);
    tma_load_3d(sdO, &tdo, &qdo_full[0], h * AD, 0, b);
  }

  for (int i = 0; i < nq_blk; ++i) {
    const int buf = i & 1;
    const int q0 = i * AT;
    const int nq_valid = min(AT, p.Nq - q0);
    const int nq_cur = (nq_valid + 15) & ~15;
    // per-query statistics of this block (plain loads; visible after the __syncthreads below)
    {
      const int q = q0 + tid;
      const int64_t si = ((int64_t)b * p.H + h) * p.Nq + q;
      s_lse[tid] = (q < p.Nq) ? p.lse[si] * LOG2E : INFINITY;
      s_delta[tid] = (q < p.Nq) ? p.delta[si] : 0.f;
    }
    // warp 0 waits as a whole and one elected lane issues: uniform control flow keeps the MMA
    // descriptors in uniform registers (no per-instruction vote / R2UR waterfall)
    if (warp == 0) {
      if (i + 1 < nq_blk && elect_one()) {
        const int nb = (i + 1) & 1;
        mbar_expect_tx(&qdo_full[nb], 2 * TILE_B);
        tma_load_3d(sQ + nb * TILE_B, &tq, &qdo_full[nb], h * AD, (i + 1) * AT, b);
        tma_load_3d(sdO + nb * TILE_B, &tdo, &qdo_full[nb], h * AD, (i + 1) * AT, b);
      }
      __syncwarp();
      if (i == 0) mbar_wait(kv_full, 0);
      mbar_wait(&qdo_full[buf], (i >> 1) & 1);
      tc_fence_after();
      const uint32_t idesc = make_idesc_bf16(AT, nq_cur, false, false);
      const uint32_t ak = smem_u32(sK), av = smem_u32(sV);
      const uint32_t bq = smem_u32(sQ + buf * TILE_B), bdo = smem_u32(sdO + buf * TILE_B);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < AD / 16; ++k)
          umma_f16(tSt, make_smem_desc(ak + k * 32, 16, 1024), make_smem_desc(bq + k * 32, 16, 1024),
                   idesc, k > 0);
#pragma unroll
        for (int k = 0; k < AD / 16; ++k)
          umma_f16(tdPt, make_smem_desc(av + k * 32, 16, 1024),
                   make_smem_desc(bdo + k * 32, 16, 1024), idesc, k > 0);
        umma_commit(s_full);
      }
      __syncwarp();
    }
    __syncthreads();  // s_lse / s_delta visible; also reconverges warp 0
    mbar_wait(s_full, i & 1);
    tc_fence_after();

    for (int c = 0; c < nq_cur; c += 32) {
      uint32_t rs[32], rp[32];
      tmem_ld32(tSt + lane_off + c, rs);
      tmem_ld32(tdPt + lane_off + c, rp);
      tmem_ld_wait();
      float pt[32], ds[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float pv = exp2f(__uint_as_float(rs[j]) * p.scale_log2 - s_lse[c + j]);
        pt[j] = (c + j < nq_valid) ? pv : 0.f;
        ds[j] = pt[j] * (__uint_as_float(rp[j]) - s_delta[c + j]) * p.scale;
      }
      uint8_t* cp = sPt + (c >> 6) * TILE_B;
      uint8_t* cs = sdSt + (c >> 6) * TILE_B;
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const uint32_t off = sw128(tid, ((c & 63) >> 3) + g);
        *reinterpret_cast<uint4*>(cp + off) =
            make_uint4(pack_bf16(pt[g * 8 + 0], pt[g * 8 + 1]), pack_bf16(pt[g * 8 + 2], pt[g * 8 + 3]),
                       pack_bf16(pt[g * 8 + 4], pt[g * 8 + 5]), pack_bf16(pt[g * 8 + 6], pt[g * 8 + 7]));
        *reinterpret_cast<uint4*>(cs + off) =
            make_uint4(pack_bf16(ds[g * 8 + 0], ds[g * 8 + 1]), pack_bf16(ds[g * 8 + 2], ds[g * 8 + 3]),
                       pack_bf16(ds[g * 8 + 4], ds[g * 8 + 5]), pack_bf16(ds[g * 8 + 6], ds[g * 8 + 7]));
      }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
      tc_fence_after();
      const uint32_t apt = smem_u32(sPt), ads = smem_u32(sdSt);
      const uint32_t bq = smem_u32(sQ + buf * TILE_B), bdo = smem_u32(sdO + buf * TILE_B);
      const uint32_t bk = smem_u32(sK);
      const uint32_t id_kmn = make_idesc_bf16(AT, AD, false, true);  // A K-major, B MN-major
      const uint32_t id_mnmn = make_idesc_bf16(AT, AD, true, true);  // A MN-major, B MN-major
      const int ksteps = nq_cur / 16;
      if (elect_one()) {
        // dV[keys,d] += P^T[keys,q] dO[q,d] ; dK[keys,d] += dS^T[keys,q] Q[q,d]   (K = queries)
        for (int ks = 0; ks < ksteps; ++ks)
          umma_f16(tdV, make_smem_desc(apt + (ks >> 2) * TILE_B + (ks & 3) * 32, 16, 1024),
                   make_smem_desc(bdo + ks * 2048, TILE_B, 1024), id_kmn, (i > 0 || ks > 0));
        for (int ks = 0; ks < ksteps; ++ks)
          umma_f16(tdK, make_smem_desc(ads + (ks >> 2) * TILE_B + (ks & 3) * 32, 16, 1024),
                   make_smem_desc(bq + ks * 2048, TILE_B, 1024), id_kmn, (i > 0 || ks > 0));
        // dQ[q,d] = dS[q,keys] K[keys,d]   (K = keys; A = dS^T buffer read MN-major)
        for (int ks = 0; ks < AT / 16; ++ks)
          umma_f16(tdQ, make_smem_desc(ads + ks * 2048, TILE_B, 1024),
                   make_smem_desc(bk + ks * 2048, TILE_B, 1024), id_mnmn, ks > 0);
        umma_commit(mma_done);
      }
    }
    __syncwarp();
    mbar_wait(mma_done, i & 1);
    tc_fence_after();
    {
      const int q = q0 + tid;  // TMEM lane = query row of dQ
      float* dst = p.dq_accum + (((int64_t)b * p.Nq + q) * p.H + h) * AD;
#pragma unroll
      for (int c = 0; c < AD; c += 32) {
        uint32_t r[32];
        tmem_ld32(tdQ + lane_off + c, r);
        tmem_ld_wait();
        if (q < p.Nq) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            red_add_v4(dst + c + j, __uint_as_float(r[j]), __uint_as_float(r[j + 1]),
                       __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
        }
      }
    }
    tc_fence_before();
    __syncthreads();  // s_lse/s_delta, sPt/sdSt and TMEM S/dP/dQ are reused next iteration
  }

  const int key = kt * AT + tid;
#pragma unroll
  for (int which = 0; which < 2; ++which) {
    const uint32_t src = which ? tdK : tdV;
    bf16* base = which ? (p.dk + (int64_t)b * p.k_bs + (int64_t)key * p.k_rs + h * AD)
                       : (p.dv + (int64_t)b * p.v_bs + (int64_t)key * p.v_rs + h * AD);
#pragma unroll
    for (int c = 0; c < AD; c += 32) {
      uint32_t r[32];
      tmem_ld32(src + lane_off + c, r);
      tmem_ld_wait();
      if (key < p.Nk) {
#pragma unroll
        for (int g = 0; g < 4; ++g)
          *reinterpret_cast<uint4*>(base + c + g * 8) = make_uint4(
              pack_bf16(__uint_as_float(r[g * 8 + 0]), __uint_as_float(r[g * 8 + 1])),
              pack_bf16(__uint_as_float(r[g * 8 + 2]), __uint_as_float(r[g * 8 + 3])),
              pack_bf16(__uint_as_float(r[g * 8 + 4]), __uint_as_float(r[g * 8 + 5])),
              pack_bf16(__uint_as_float(r[g * 8 + 6]), __uint_as_float(r[g * 8 + 7])));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

// delta[b,h,q] = sum_d dO[b,q,h,d] * O[b,q,h,d]   (warp per (b,q): lanes over H*D/2 pairs)
__global__ void attn_delta_kernel(const bf16* __restrict__ o, const bf16* __restrict__ o_lo,
                                  const bf16* __restrict__ d_o, float* __restrict__ delta, int B,
                                  int H, int Nq, int64_t o_bs, int64_t o_rs) {
  const int lane = threadIdx.x & 31;
  const int64_t w = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (w >= (int64_t)B * Nq * H) return;
  const int hh = (int)(w % H);
  const int q = (int)((w / H) % Nq);
  const int b = (int)(w / ((int64_t)H * Nq));
  const int64_t off = (int64_t)b * o_bs + (int64_t)q * o_rs + hh * AD + lane * 2;
  float2 a = unpack_bf16(*reinterpret_cast<const uint32_t*>(o + off));
  if (o_lo) {
    const float2 l = unpack_bf16(*reinterpret_cast<const uint32_t*>(o_lo + off));
    a.x += l.x;
    a.y += l.y;
  }
  const float2 g = unpack_bf16(*reinterpret_cast<const uint32_t*>(d_o + off));
  float s = a.x * g.x + a.y * g.y;
#pragma unroll
  for (int k = 16; k > 0; k >>= 1) s += __shfl_xor_sync(0xffffffffu, s, k);
  if (lane == 0) delta[((int64_t)b * H + hh) * Nq + q] = s;
}

// dq (bf16, strided [B,Nq,H,D] view) = bf16(dq_accum fp32 [B,Nq,H,D])
__global__ void attn_dq_store_kernel(const float* __restrict__ acc, bf16* __restrict__ dq,
                                     int64_t total4, int Nq, int HD, int64_t q_bs, int64_t q_rs) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total4;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = i * 4;
    const int c = (int)(e % HD);
    const int64_t r = e / HD;
    const int q = (int)(r % Nq);
    const int64_t b = r / Nq;
    const float4 v = *reinterpret_cast<const float4*>(acc + e);
    *reinterpret_cast<uint2*>(dq + b * q_bs + (int64_t)q * q_rs + c) =
        make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
  }
}

static int make_qkv_map(CUtensorMap* m, const void* base, int64_t H, int64_t N, int64_t B,
                        int64_t bs, int64_t rs) {
  uint64_t dims[3] = {(uint64_t)(H * AD), (uint64_t)N, (uint64_t)B};
  uint64_t str[3] = {0, (uint64_t)rs * 2, (uint64_t)bs * 2};
  uint32_t box[3] = {AD, AT, 1};
  return make_tmap_bf16(m, base, 3, dims, str, box);
}

constexpr int kBwdSmem = 10 * TILE_B + 2 * AT * 4 + 1024 + 128;

}  // namespace sav

using namespace sav;

extern "C" int sav_attn_bwd(const sav_attn_bwd_args* a, void* stream) {
  if (!a || a->D != AD) return set_error(SAV_ERR_SHAPE, "sav_attn_bwd: head dim must be 64");
  if (a->B <= 0 || a->H <= 0 || a->Nq <= 0 || a->Nk <= 0)
    return set_error(SAV_ERR_SHAPE, "sav_attn_bwd: empty shape");
  if (!a->delta) return set_error(SAV_ERR_ARG, "sav_attn_bwd: delta workspace required");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // every shape runs the resident-query kernel of attn_bwd.cu (query groups of <= 256 rows).  The
  // older key-tile-per-CTA kernel below (fp32 dQ accumulation + a conversion pass; 921 us against
  // the new path on ViT-B/16 384^2) is kept for A/B runs of the development build only.
#ifndef SAV_ATTN_DEV
  return launch_attn_bwd2(a, st);
#else
  if (a->Nq <= 256 || !getenv("SAV_ATTN_BWD_GENERIC")) return launch_attn_bwd2(a, st);
#endif
  if (!a->dq_accum) return set_error(SAV_ERR_ARG, "sav_attn_bwd: dq_accum workspace required for Nq > 256");
  CUtensorMap tq, tk, tv, tdo;
  int rc;
  if ((rc = make_qkv_map(&tq, a->q, a->H, a->Nq, a->B, a->q_batch_stride, a->q_row_stride))) return rc;
  if ((rc = make_qkv_map(&tk, a->k, a->H, a->Nk, a->B, a->k_batch_stride, a->k_row_stride))) return rc;
  if ((rc = make_qkv_map(&tv, a->v, a->H, a->Nk, a->B, a->v_batch_stride, a->v_row_stride))) return rc;
  if ((rc = make_qkv_map(&tdo, a->d_o, a->H, a->Nq, a->B, a->o_batch_stride, a->o_row_stride))) return rc;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem);
    if (e != cudaSuccess) return set_error(SAV_ERR_CUDA, "attn bwd smem attr: %s", cudaGetErrorString(e));
    attr = true;
  }
  const int64_t rows = a->B * a->Nq * a->H;
  attn_delta_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(
      reinterpret_cast<const bf16*>(a->o), reinterpret_cast<const bf16*>(a->o_lo),
      reinterpret_cast<const bf16*>(a->d_o), a->delta,
      (int)a->B, (int)a->H, (int)a->Nq, a->o_batch_stride, a->o_row_stride);
  count_launch();
  const int64_t nacc = a->B * a->Nq * a->H * AD;
  if ((rc = sav_fill_f32(a->dq_accum, 0.f, nacc, stream))) return rc;
  AttnBwdParams p;
  p.lse = a->lse; p.delta = a->delta; p.dq_accum = a->dq_accum;
  p.dk = reinterpret_cast<bf16*>(a->dk); p.dv = reinterpret_cast<bf16*>(a->dv);
  p.B = (int)a->B; p.H = (int)a->H; p.Nq = (int)a->Nq; p.Nk = (int)a->Nk;
  p.k_bs = a->k_batch_stride; p.k_rs = a->k_row_stride;
  p.v_bs = a->v_batch_stride; p.v_rs = a->v_row_stride;
  p.scale = a->scale; p.scale_log2 = a->scale * LOG2E;
  dim3 grid((unsigned)((a->Nk + AT - 1) / AT), (unsigned)a->H, (unsigned)a->B);
  attn_bwd_kernel<<<grid, 128, kBwdSmem, st>>>(tq, tk, tv, tdo, p);
  count_launch();
  if ((rc = check_launch("sav_attn_bwd"))) return rc;
  const int64_t total4 = nacc / 4;
  int blocks = (int)((total4 + 255) / 256 < 148 * 8 ? (total4 + 255) / 256 : 148 * 8);
  attn_dq_store_kernel<<<blocks, 256, 0, st>>>(a->dq_accum, reinterpret_cast<bf16*>(a->dq), total4,
                                               (int)a->Nq, (int)(a->H * AD), a->q_batch_stride,
                                               a->q_row_stride);
  count_launch();
  return check_launch("sav_attn_bwd dq store");
}
