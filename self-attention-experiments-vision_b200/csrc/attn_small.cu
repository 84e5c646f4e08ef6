// Attention over very short sequences with very narrow heads: the INNER transformer of TNT
// (models/tnt.py:60-69: SelfAttentionBlock over the 16 "pixel" tokens of one patch, 24 or 40
// channels in 4 heads => head dim 6 or 10, batch = images x 196 patches).  attention.py:39-58 per
// (sequence, head) is 16 x 16 logits over a 6-wide contraction -- far below any tensor-core tile and
// not even 16-byte aligned per head -- so this path is plain FP32 CUDA-core work and HBM-bound:
// one warp owns one (sequence, head), stages q, k, v (and dO) in shared memory, a lane owns one
// query row (forward, dq) and then one key row (dk, dv).  Nothing is saved for the backward pass:
// P is recomputed from q, k (N <= 32 keys, D <= 16).  Sequences of up to 16 tokens (TNT: exactly 16)
// use the NM = 16 instantiation: two (sequence, head) items per warp, one per half-warp, with the
// token loops, register arrays and shared-memory tiles sized for 16 -- the 32-token layout left half
// of every warp idle and ran the 12 544 x 4 items of a 64-image TNT-S step at 260 / 592 us
// (forward / backward) against a 10 / 25 us HBM floor (profiles/r02ay_timeline_tnt_s_patch16.txt).
#include "common.cuh"
#include "host.h"

namespace sav {

constexpr int SM_N = 32;      // max tokens per sequence
constexpr int SM_D = 16;      // max head dim
constexpr int SM_WARPS = 2;   // warps per block

template <int NM>              // NM = 16 or 32: token capacity of one item's tiles
struct SmallAttnSmem {
  float q[NM][SM_D + 1], k[NM][SM_D + 1], v[NM][SM_D + 1], g[NM][SM_D + 1];
  float p[NM][NM + 1], ds[NM][NM + 1];
};

// row i of softmax(scale q k^T): p[j] for j < N (0 beyond)
template <int NM>
__device__ __forceinline__ void small_softmax_row(const SmallAttnSmem<NM>& s, int i, int N, int D, float scale,
                                                  float (&p)[NM]) {
  float mx = -INFINITY;
#pragma unroll
  for (int j = 0; j < NM; ++j) {
    float a = 0.f;
    if (j < N) {
      for (int d = 0; d < D; ++d) a += s.q[i][d] * s.k[j][d];
      a *= scale;                                   // attention.py:39: q / sqrt(d) before the product
      mx = fmaxf(mx, a);
    }
    p[j] = a;
  }
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < NM; ++j) {
    p[j] = j < N ? __expf(p[j] - mx) : 0.f;         // attention.py:48
    sum += p[j];
  }
  const float inv = 1.f / sum;
#pragma unroll
  for (int j = 0; j < NM; ++j) p[j] *= inv;
}

template <bool BWD, int NM>
__global__ void __launch_bounds__(SM_WARPS * 32)
attn_small_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ d_o, bf16* __restrict__ out,
                  int64_t items, int N_, int H, int D, float scale) {
  constexpr int IPW = 32 / NM;                       // items per warp (one per NM-lane group)
  __shared__ SmallAttnSmem<NM> smem[SM_WARPS * IPW];
  const int warp = threadIdx.x >> 5, wl = threadIdx.x & 31;
  const int sub = wl / NM, lane = wl % NM;           // `lane` = position inside the item's lane group
  const int64_t item = ((int64_t)blockIdx.x * SM_WARPS + warp) * IPW + sub;
  // a lane group without an item walks the code with N = 0 (the warp-wide syncs stay convergent)
  const int N = item < items ? N_ : 0;
  SmallAttnSmem<NM>& s = smem[warp * IPW + sub];
  const int64_t b = item < items ? item / H : 0;
  const int h = (int)(item % H);
  const int64_t row_pitch = (int64_t)3 * H * D;     // one token of [B, N, 3, H, D]
  const bf16* base = qkv + b * N * row_pitch + (int64_t)h * D;
  for (int idx = lane; idx < N * D; idx += NM) {
    const int n = idx / D, d = idx - n * D;
    const bf16* tok = base + (int64_t)n * row_pitch + d;
    s.q[n][d] = __bfloat162float(tok[0]);
    s.k[n][d] = __bfloat162float(tok[(int64_t)H * D]);
    s.v[n][d] = __bfloat162float(tok[(int64_t)2 * H * D]);
    if (BWD) s.g[n][d] = __bfloat162float(d_o[((b * N + n) * H + h) * D + d]);
  }
  __syncwarp();
  float p[NM];
  if (lane < N) {
    const int i = lane;
    small_softmax_row<NM>(s, i, N, D, scale, p);
    if (!BWD) {
      bf16* o = out + ((b * N + i) * H + h) * D;     // [B, N, H, D]
      for (int d = 0; d < D; ++d) {
        float a = 0.f;
#pragma unroll
        for (int j = 0; j < NM; ++j)
          if (j < N) a += p[j] * s.v[j][d];          // attention.py:57
        o[d] = __float2bfloat16(a);
      }
    } else {
      float delta = 0.f;
      float dp[NM];
#pragma unroll
      for (int j = 0; j < NM; ++j) {
        float a = 0.f;
        if (j < N)
          for (int d = 0; d < D; ++d) a += s.g[i][d] * s.v[j][d];
        dp[j] = a;
        delta += p[j] * a;
      }
#pragma unroll
      for (int j = 0; j < NM; ++j) {
        if (j < N) {
          s.p[i][j] = p[j];
          s.ds[i][j] = p[j] * (dp[j] - delta);
        }
      }
      bf16* dq = out + (b * N + i) * row_pitch + (int64_t)h * D;
      for (int d = 0; d < D; ++d) {
        float a = 0.f;
#pragma unroll
        for (int j = 0; j < NM; ++j)
          if (j < N) a += p[j] * (dp[j] - delta) * s.k[j][d];
        dq[d] = __float2bfloat16(a * scale);
      }
    }
  }
  if (BWD) {
    __syncwarp();
    if (lane < N) {
      const int j = lane;
      bf16* dk = out + (b * N + j) * row_pitch + (int64_t)(H + h) * D;
      bf16* dv = dk + (int64_t)H * D;
      for (int d = 0; d < D; ++d) {
        float ak = 0.f, av = 0.f;
        for (int i = 0; i < N; ++i) {
          ak += s.ds[i][j] * s.q[i][d];
          av += s.p[i][j] * s.g[i][d];
        }
        dk[d] = __float2bfloat16(ak * scale);
        dv[d] = __float2bfloat16(av);
      }
    }
  }
}

static int small_check(const char* who, const void* a, const void* b, int64_t B, int64_t N, int64_t H,
                       int64_t D) {
  if (!a || !b || B <= 0 || H <= 0) return set_error(SAV_ERR_ARG, "%s: null pointer or empty batch", who);
  if (N <= 0 || N > SM_N || D <= 0 || D > SM_D)
    return set_error(SAV_ERR_SHAPE, "%s: needs N <= %d tokens and head dim <= %d (N=%lld, D=%lld)", who,
                     SM_N, SM_D, (long long)N, (long long)D);
  return SAV_OK;
}

}  // namespace sav

using namespace sav;

extern "C" int sav_attn_small_fwd(const void* qkv, void* o, int64_t B, int64_t N, int64_t H, int64_t D,
                                  float scale, void* stream) {
  if (int rc = small_check("sav_attn_small_fwd", qkv, o, B, N, H, D)) return rc;
  const int64_t items = B * H;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (N <= 16)
    attn_small_kernel<false, 16><<<(unsigned)((items + 2 * SM_WARPS - 1) / (2 * SM_WARPS)), SM_WARPS * 32, 0, st>>>(
        reinterpret_cast<const bf16*>(qkv), nullptr, reinterpret_cast<bf16*>(o), items, (int)N, (int)H, (int)D, scale);
  else
    attn_small_kernel<false, 32><<<(unsigned)((items + SM_WARPS - 1) / SM_WARPS), SM_WARPS * 32, 0, st>>>(
        reinterpret_cast<const bf16*>(qkv), nullptr, reinterpret_cast<bf16*>(o), items, (int)N, (int)H, (int)D, scale);
  count_launch();
  return check_launch("sav_attn_small_fwd");
}

extern "C" int sav_attn_small_bwd(const void* qkv, const void* d_o, void* dqkv, int64_t B, int64_t N,
                                  int64_t H, int64_t D, float scale, void* stream) {
  if (int rc = small_check("sav_attn_small_bwd", qkv, dqkv, B, N, H, D)) return rc;
  if (!d_o) return set_error(SAV_ERR_ARG, "sav_attn_small_bwd: d_o is null");
  const int64_t items = B * H;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (N <= 16)
    attn_small_kernel<true, 16><<<(unsigned)((items + 2 * SM_WARPS - 1) / (2 * SM_WARPS)), SM_WARPS * 32, 0, st>>>(
        reinterpret_cast<const bf16*>(qkv), reinterpret_cast<const bf16*>(d_o), reinterpret_cast<bf16*>(dqkv),
        items, (int)N, (int)H, (int)D, scale);
  else
    attn_small_kernel<true, 32><<<(unsigned)((items + SM_WARPS - 1) / SM_WARPS), SM_WARPS * 32, 0, st>>>(
        reinterpret_cast<const bf16*>(qkv), reinterpret_cast<const bf16*>(d_o), reinterpret_cast<bf16*>(dqkv),
        items, (int)N, (int)H, (int)D, scale);
  count_launch();
  return check_launch("sav_attn_small_bwd");
}
