// Persistent warp-specialised bf16 GEMM for sm_100a: TMA -> 128B-swizzled smem ring -> tcgen05.mma
// (accumulators double-buffered in TMEM) -> fused epilogue.  One CTA per SM, 192 threads:
//   warp 0     TMA producer (one elected lane)
//   warp 1     MMA issuer (one lane) + TMEM allocator
//   warps 2-5  epilogue: tcgen05.ld -> bias / GELU / dGELU / LayerScale / drop-path / residual -> HBM
// Operands may be K-major or MN-major (selected by the smem + instruction descriptors), so
// forward (X W), dX (dY W^T) and dW (X^T dY) all run here without any transposed copies.
#include "common.cuh"
#include "host.h"

namespace sav {

constexpr int BM = 128;  // UMMA M
constexpr int BK = 64;   // 64 bf16 = one 128-byte swizzle row

struct GemmParams {
  int64_t M, N, K;
  int m_blocks, n_blocks, split_k, k_blocks, k_per_split;
  void* D;
  int64_t ldd;
  int d_fp32, accumulate;
  const float* bias;
  int act;
  bf16* aux_out;
  const bf16* aux_in;
  const float* col_scale;
  const float* row_scale;
  int64_t rows_per_scale;
  const bf16* residual;
  int64_t ldr;
  float* workspace;
};

template <int BN>
struct GemmCfg {
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kSmem = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
  static constexpr uint32_t kTmemCols = 2 * BN;  // 256 or 512, power of two
};

__device__ __forceinline__ void epilogue_group8(const GemmParams& p, float (&v)[8], int64_t row,
                                                int col, float rs, void* Dbase, int64_t ldd) {
  if (p.bias) {
    const float4 b0 = *reinterpret_cast<const float4*>(p.bias + col);
    const float4 b1 = *reinterpret_cast<const float4*>(p.bias + col + 4);
    v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
    v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
  }
  if (p.aux_out) {
    uint4 o = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]),
                         pack_bf16(v[6], v[7]));
    *reinterpret_cast<uint4*>(p.aux_out + row * p.ldd + col) = o;
  }
  if (p.act == 1) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = gelu_tanh(v[j]);
  }
  if (p.aux_in) {
    const uint4 a = *reinterpret_cast<const uint4*>(p.aux_in + row * p.ldd + col);
    const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float2 h = unpack_bf16(w[j]);
      v[2 * j] *= dgelu_tanh(h.x);
      v[2 * j + 1] *= dgelu_tanh(h.y);
    }
  }
  if (p.col_scale) {
    const float4 s0 = *reinterpret_cast<const float4*>(p.col_scale + col);
    const float4 s1 = *reinterpret_cast<const float4*>(p.col_scale + col + 4);
    v[0] *= s0.x; v[1] *= s0.y; v[2] *= s0.z; v[3] *= s0.w;
    v[4] *= s1.x; v[5] *= s1.y; v[6] *= s1.z; v[7] *= s1.w;
  }
  if (p.row_scale) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] *= rs;
  }
  if (p.residual) {
    const uint4 a = *reinterpret_cast<const uint4*>(p.residual + row * p.ldr + col);
    const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float2 h = unpack_bf16(w[j]);
      v[2 * j] += h.x;
      v[2 * j + 1] += h.y;
    }
  }
  if (p.d_fp32) {
    float* d = reinterpret_cast<float*>(Dbase) + row * ldd + col;
    float4 o0 = make_float4(v[0], v[1], v[2], v[3]);
    float4 o1 = make_float4(v[4], v[5], v[6], v[7]);
    if (p.accumulate && p.split_k == 1) {
      const float4 c0 = *reinterpret_cast<const float4*>(d);
      const float4 c1 = *reinterpret_cast<const float4*>(d + 4);
      o0.x += c0.x; o0.y += c0.y; o0.z += c0.z; o0.w += c0.w;
      o1.x += c1.x; o1.y += c1.y; o1.z += c1.z; o1.w += c1.w;
    }
    *reinterpret_cast<float4*>(d) = o0;
    *reinterpret_cast<float4*>(d + 4) = o1;
  } else {
    bf16* d = reinterpret_cast<bf16*>(Dbase) + row * ldd + col;
    *reinterpret_cast<uint4*>(d) = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]),
                                              pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
  }
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(192, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
            const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 128);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int tiles_mn = p.m_blocks * p.n_blocks;
  const int num_tiles = tiles_mn * p.split_k;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int split = tile / tiles_mn;
        const int mn = tile - split * tiles_mn;
        const int m_idx = (mn / p.n_blocks) * BM;
        const int n_idx = (mn % p.n_blocks) * BN;
        const int kb0 = split * p.k_per_split;
        const int kb1 = min(kb0 + p.k_per_split, p.k_blocks);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kABytes;
          mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          const int k_idx = kb * BK;
          if (!A_MN) {
            tma_load_2d(sa, &tma_a, &full_bar[stage], k_idx, m_idx);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_2d(sa + j * (BK * 128), &tma_a, &full_bar[stage], m_idx + j * 64, k_idx);
          }
          if (!B_MN) {
            tma_load_2d(sb, &tma_b, &full_bar[stage], k_idx, n_idx);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(sb + j * (BK * 128), &tma_b, &full_bar[stage], n_idx + j * 64, k_idx);
          }
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int split = tile / tiles_mn;
        const int kb0 = split * p.k_per_split;
        const int kb1 = min(kb0 + p.k_per_split, p.k_blocks);
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint32_t sb = sa + Cfg::kABytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = A_MN ? make_smem_desc(sa + k * 2048, BK * 128, 1024)
                                     : make_smem_desc(sa + k * 32, 16, 1024);
            const uint64_t db = B_MN ? make_smem_desc(sb + k * 2048, BK * 128, 1024)
                                     : make_smem_desc(sb + k * 32, 16, 1024);
            umma_f16(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull_bar[acc]);
      }
    }
  } else {
    const int sub = warp & 3;  // TMEM sub-partition of this warp
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int split = tile / tiles_mn;
      const int mn = tile - split * tiles_mn;
      const int m_idx = (mn / p.n_blocks) * BM;
      const int n_idx = (mn % p.n_blocks) * BN;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const int64_t row = m_idx + sub * 32 + lane;
      const bool row_ok = row < p.M;
      float rs = 1.f;
      if (p.row_scale && row_ok) rs = p.row_scale[row / p.rows_per_scale];
      void* Dbase = p.D;
      int64_t ldd = p.ldd;
      if (p.split_k > 1) {
        Dbase = p.workspace + (int64_t)split * p.M * p.N;
        ldd = p.N;
      }
      const uint32_t taddr = tmem_base + ((uint32_t)(sub * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        if (n_idx + c * 32 >= p.N) break;  // warp-uniform
        uint32_t r[32];
        tmem_ld32(taddr + c * 32, r);
        tmem_ld_wait();
        if (row_ok) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const int col = n_idx + c * 32 + g * 8;
            if (col < p.N) {
              float v[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[g * 8 + j]);
              epilogue_group8(p, v, row, col, rs, Dbase, ldd);
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tempty_bar[acc]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

// ---- split-K reduction: D (+)= sum_s ws[s]  (fp32, vectorised, HBM-bound)
__global__ void splitk_reduce_kernel(const float* __restrict__ ws, float* __restrict__ D,
                                     int64_t M, int64_t N, int64_t ldd, int splits,
                                     int accumulate) {
  const int64_t n4 = N / 4;
  const int64_t total = M * n4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / n4, c = (i - r * n4) * 4;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int s = 0; s < splits; ++s) {
      const float4 v = *reinterpret_cast<const float4*>(ws + ((int64_t)s * M + r) * N + c);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    float* d = D + r * ldd + c;
    if (accumulate) {
      const float4 o = *reinterpret_cast<const float4*>(d);
      acc.x += o.x; acc.y += o.y; acc.z += o.z; acc.w += o.w;
    }
    *reinterpret_cast<float4*>(d) = acc;
  }
}

// ---- column sums of a bf16 [M,N] matrix (bias gradients). grid = (N/64 col-groups, row-splits);
// each thread owns 2 adjacent columns, 8 row-lanes per block, smem reduce, one atomicAdd per
// column per block.  `out` must be zeroed by the caller path (done in the launcher).
__global__ void colsum_kernel(const bf16* __restrict__ X, int64_t M, int64_t N, int64_t ldx,
                              float* __restrict__ out) {
  __shared__ float2 red[8][32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t col = (int64_t)blockIdx.x * 64 + tx * 2;
  float2 acc = make_float2(0.f, 0.f);
  if (col < N) {
    for (int64_t r = (int64_t)blockIdx.y * 8 + ty; r < M; r += (int64_t)gridDim.y * 8) {
      const uint32_t u = *reinterpret_cast<const uint32_t*>(X + r * ldx + col);
      const float2 f = unpack_bf16(u);
      acc.x += f.x;
      acc.y += f.y;
    }
  }
  red[ty][tx] = acc;
  __syncthreads();
  if (ty == 0 && col < N) {
    for (int j = 1; j < 8; ++j) { acc.x += red[j][tx].x; acc.y += red[j][tx].y; }
    atomicAdd(out + col, acc.x);
    atomicAdd(out + col + 1, acc.y);
  }
}

__global__ void fill_kernel(float* p, float v, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    p[i] = v;
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

static int pick_bn(int64_t N) { return (N % 256 == 0) ? 256 : 128; }

static int plan_split(const sav_gemm_args* a, int bn) {
  if (a->split_k > 0) return a->split_k;
  // auto (split_k == 0): only for fp32 outputs without epilogue, fill the machine
  const int64_t tiles = ((a->M + BM - 1) / BM) * ((a->N + bn - 1) / bn);
  const int64_t kb = (a->K + BK - 1) / BK;
  if (!a->d_fp32 || tiles >= num_sms() || kb < 16) return 1;
  int64_t s = num_sms() / tiles;
  if (s > kb / 8) s = kb / 8;
  return (int)(s < 1 ? 1 : s);
}

template <int BN, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p,
                       cudaStream_t st) {
  using Cfg = GemmCfg<BN>;
  auto kern = gemm_kernel<BN, A_MN, B_MN>;
  static bool attr_set = false;  // benign race: idempotent
  if (!attr_set) {
    cudaError_t e =
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmem);
    if (e != cudaSuccess) return set_error(SAV_ERR_CUDA, "gemm smem attr: %s", cudaGetErrorString(e));
    attr_set = true;
  }
  const int tiles = p.m_blocks * p.n_blocks * p.split_k;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  kern<<<grid, 192, Cfg::kSmem, st>>>(ta, tb, p);
  count_launch();
  return check_launch("sav_gemm");
}

}  // namespace sav

using namespace sav;

extern "C" int64_t sav_gemm_workspace_bytes(const sav_gemm_args* a) {
  const int s = plan_split(a, pick_bn(a->N));
  return s > 1 ? (int64_t)s * a->M * a->N * 4 : 0;
}

extern "C" int sav_gemm(const sav_gemm_args* a, void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!a || a->M <= 0 || a->N <= 0 || a->K <= 0)
    return set_error(SAV_ERR_SHAPE, "sav_gemm: M,N,K must be positive");
  if (a->N % 8 || a->lda % 8 || a->ldb % 8 || a->ldd % (a->d_fp32 ? 4 : 8) ||
      (a->residual && a->ldr % 8))
    return set_error(SAV_ERR_ALIGN, "sav_gemm: N, lda, ldb, ldd, ldr must be multiples of 8");
  if (!aligned16(a->A) || !aligned16(a->B) || !aligned16(a->D) || !aligned16(a->bias) ||
      !aligned16(a->aux_in) || !aligned16(a->aux_out) || !aligned16(a->residual) ||
      !aligned16(a->col_scale))
    return set_error(SAV_ERR_ALIGN, "sav_gemm: pointers must be 16 B aligned");
  const int bn = pick_bn(a->N);
  const int split = plan_split(a, bn);
  if (split > 1) {
    if (!a->d_fp32 || a->bias || a->act || a->aux_in || a->aux_out || a->col_scale ||
        a->row_scale || a->residual)
      return set_error(SAV_ERR_ARG, "sav_gemm: split_k>1 needs d_fp32 and no epilogue");
    if (!a->workspace || a->workspace_bytes < (int64_t)split * a->M * a->N * 4)
      return set_error(SAV_ERR_ARG, "sav_gemm: workspace too small for split_k=%d", split);
  }
  if (a->accumulate && !a->d_fp32) return set_error(SAV_ERR_ARG, "sav_gemm: accumulate needs d_fp32");

  GemmParams p;
  p.M = a->M; p.N = a->N; p.K = a->K;
  p.m_blocks = (int)((a->M + BM - 1) / BM);
  p.n_blocks = (int)((a->N + bn - 1) / bn);
  p.k_blocks = (int)((a->K + BK - 1) / BK);
  p.k_per_split = (p.k_blocks + split - 1) / split;
  p.split_k = (p.k_blocks + p.k_per_split - 1) / p.k_per_split;
  p.D = a->D; p.ldd = a->ldd; p.d_fp32 = a->d_fp32; p.accumulate = a->accumulate;
  p.bias = a->bias; p.act = a->act;
  p.aux_out = reinterpret_cast<bf16*>(a->aux_out);
  p.aux_in = reinterpret_cast<const bf16*>(a->aux_in);
  p.col_scale = a->col_scale; p.row_scale = a->row_scale;
  p.rows_per_scale = a->rows_per_scale > 0 ? a->rows_per_scale : 1;
  p.residual = reinterpret_cast<const bf16*>(a->residual); p.ldr = a->ldr;
  p.workspace = reinterpret_cast<float*>(a->workspace);

  CUtensorMap ta, tb;
  int rc;
  {
    uint64_t dims[2], str[2];
    uint32_t box[2];
    if (!a->a_mn) { dims[0] = a->K; dims[1] = a->M; box[0] = BK; box[1] = BM; }
    else          { dims[0] = a->M; dims[1] = a->K; box[0] = 64; box[1] = BK; }
    str[0] = 0; str[1] = (uint64_t)a->lda * 2;
    if ((rc = make_tmap_bf16(&ta, a->A, 2, dims, str, box))) return rc;
    if (!a->b_mn) { dims[0] = a->K; dims[1] = a->N; box[0] = BK; box[1] = bn; }
    else          { dims[0] = a->N; dims[1] = a->K; box[0] = 64; box[1] = BK; }
    str[1] = (uint64_t)a->ldb * 2;
    if ((rc = make_tmap_bf16(&tb, a->B, 2, dims, str, box))) return rc;
  }
#define SAV_GEMM_CASE(BNv, AM, BMj)                                                  \
  if (bn == BNv && (a->a_mn != 0) == AM && (a->b_mn != 0) == BMj)                     \
    rc = launch_gemm<BNv, AM, BMj>(ta, tb, p, st);
  rc = SAV_ERR_ARG;
  SAV_GEMM_CASE(128, false, false) SAV_GEMM_CASE(128, false, true)
  SAV_GEMM_CASE(128, true, false)  SAV_GEMM_CASE(128, true, true)
  SAV_GEMM_CASE(256, false, false) SAV_GEMM_CASE(256, false, true)
  SAV_GEMM_CASE(256, true, false)  SAV_GEMM_CASE(256, true, true)
#undef SAV_GEMM_CASE
  if (rc) return rc;
  if (p.split_k > 1) {
    const int64_t total = a->M * (a->N / 4);
    int blocks = (int)((total + 255) / 256);
    if (blocks > num_sms() * 8) blocks = num_sms() * 8;
    splitk_reduce_kernel<<<blocks, 256, 0, st>>>(p.workspace, reinterpret_cast<float*>(a->D),
                                                 a->M, a->N, a->ldd, p.split_k, a->accumulate);
    count_launch();
    return check_launch("sav_gemm split-K reduce");
  }
  return SAV_OK;
}

extern "C" int sav_fill_f32(float* ptr, float v, int64_t n, void* stream) {
  if (n <= 0) return SAV_OK;
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  fill_kernel<<<blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(ptr, v, n);
  count_launch();
  return check_launch("sav_fill_f32");
}

extern "C" int sav_colsum_bf16(const void* X, int64_t M, int64_t N, int64_t ldx, float* out,
                               int32_t accumulate, void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (M <= 0 || N <= 0 || N % 2 || ldx % 2) return set_error(SAV_ERR_SHAPE, "sav_colsum: bad shape");
  if (!accumulate) {
    int rc = sav_fill_f32(out, 0.f, N, stream);
    if (rc) return rc;
  }
  dim3 grid((unsigned)((N + 63) / 64), 1);
  int64_t want = (num_sms() * 4 + grid.x - 1) / grid.x;
  int64_t maxy = (M + 7) / 8;
  grid.y = (unsigned)(want < maxy ? (want < 1 ? 1 : want) : maxy);
  colsum_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const bf16*>(X), M, N, ldx, out);
  count_launch();
  return check_launch("sav_colsum_bf16");
}
