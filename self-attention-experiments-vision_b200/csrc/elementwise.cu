// HBM-bound helpers around the residual stream that no GEMM epilogue can absorb:
//   * reverse mode of the LayerScale / drop-path epilogue (layerscale.py:20-23,
//     stochastic_depth.py:16-27): dy_branch = bf16(g * col_scale[n] * row_scale[m]) from the fp32
//     stream gradient, and the LayerScale gradient recovered from the already-computed weight
//     gradient (no saved activation):  with u = xW + b and y = u * cs,
//         d cs[n] = sum_m g[m,n] rs[m] u[m,n] = (sum_k W[k,n] dW[k,n] + b[n] db[n]) / cs[n]
//     because dW = x^T (g cs rs) and db = colsum(g cs rs);
//   * strided fp32 row copies / accumulations (class-token concat of cait.py:97,175 and its
//     reverse mode, fan-in of fp32 stream gradients).
// All vectorised 16-byte accesses, grid sized in multiples of the SM count.
#include "common.cuh"
#include "host.h"

namespace sav {

__global__ void scale_rows_cols_kernel(const float* __restrict__ g, int64_t ldg,
                                       const float* __restrict__ cs, const float* __restrict__ rs,
                                       int64_t rows_per_scale, bf16* __restrict__ out, int64_t ldo,
                                       int64_t rows, int cols4) {
  const int64_t total = rows * cols4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / cols4;
    const int c = (int)(i - r * cols4) * 4;
    float4 v = *reinterpret_cast<const float4*>(g + r * ldg + c);
    if (cs) {
      const float4 s = __ldg(reinterpret_cast<const float4*>(cs + c));
      v.x *= s.x; v.y *= s.y; v.z *= s.z; v.w *= s.w;
    }
    if (rs) {
      const float f = __ldg(rs + r / rows_per_scale);
      v.x *= f; v.y *= f; v.z *= f; v.w *= f;
    }
    *reinterpret_cast<uint2*>(out + r * ldo + c) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
  }
}

// out[n] += (sum_k W[k,n] G[k,n] + b[n] gb[n]) / cs[n].  Block (x, y) owns 32 columns and the
// kLsRows-row slice y of K (8 row lanes); slices combine with one fp32 atomicAdd per column, like the
// split-K partials of dW itself.  (One block per 32 columns over the whole of K left 12 blocks on
// 148 SMs walking 192 dependent loads each: 30-43 us per call, 52 calls per CaiT-S24 step.)
constexpr int kLsRows = 64;
__global__ void layerscale_grad_kernel(const float* __restrict__ W, int64_t ldw,
                                       const float* __restrict__ G, int64_t ldg, int64_t K, int64_t N,
                                       const float* __restrict__ b, const float* __restrict__ gb,
                                       const float* __restrict__ cs, float* __restrict__ out) {
  __shared__ float red[8][32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t n = (int64_t)blockIdx.x * 32 + tx;
  const int64_t k0 = (int64_t)blockIdx.y * kLsRows;
  const int64_t k1 = k0 + kLsRows < K ? k0 + kLsRows : K;
  float acc = 0.f;
  if (n < N) {
#pragma unroll 8
    for (int64_t k = k0 + ty; k < k1; k += 8) acc = fmaf(W[k * ldw + n], G[k * ldg + n], acc);
  }
  red[ty][tx] = acc;
  __syncthreads();
  if (ty == 0 && n < N) {
    for (int j = 1; j < 8; ++j) acc += red[j][tx];
    if (b && blockIdx.y == 0) acc = fmaf(b[n], gb[n], acc);
    const float s = cs[n];
    if (s != 0.f) atomicAdd(out + n, acc / s);
  }
}

__global__ void copy2d_f32_kernel(const float* __restrict__ src, int64_t lds, float* __restrict__ dst,
                                  int64_t ldd, int64_t rows, int64_t cols4, int accumulate) {
  const int64_t total = rows * cols4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / cols4;
    const int64_t c = (i - r * cols4) * 4;
    float4 v = *reinterpret_cast<const float4*>(src + r * lds + c);
    float4* d = reinterpret_cast<float4*>(dst + r * ldd + c);
    if (accumulate) {
      const float4 o = *d;
      v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
    }
    *d = v;
  }
}

// ---- rotary position embedding (position_embed.py:8-45).  The reference's frequency table is
// inv_freq[i] = 1 / (10e4^(2i) / dim) evaluated in float32 (position_embed.py:29: 10e4 = 1e5, the
// power overflows to inf for 2i >= 8, so those frequencies are exactly 0);  theta[t,i] = t inv_freq[i];
//   y[2i]   = x[2i] cos(theta) - x[2i+1] sin(theta)
//   y[2i+1] = x[2i+1] cos(theta) + x[2i] sin(theta)          (x cos + rotate_every_two(x) sin)
// `inverse` rotates by -theta: the reverse mode (the rotation is orthogonal).
__device__ __forceinline__ float rotary_inv_freq(int i, int dim) {
  const float pw = powf(1e5f, (float)(2 * i));
  return 1.f / (pw / (float)dim);
}

template <typename T>
__global__ void rotary_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t outer, int64_t n,
                              int64_t inner, int d, float sign) {
  const int half = d / 2;
  const int64_t total = outer * n * inner * half;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(idx % half);
    const int64_t row = idx / half;              // (outer, t, inner)
    const int64_t t = (row / inner) % n;
    float sn, cs;
    sincosf((float)t * rotary_inv_freq(i, d), &sn, &cs);
    sn *= sign;
    float a, b;
    if (sizeof(T) == 2) {
      const float2 v = unpack_bf16(reinterpret_cast<const uint32_t*>(x)[idx]);
      a = v.x; b = v.y;
      reinterpret_cast<uint32_t*>(y)[idx] = pack_bf16(a * cs - b * sn, b * cs + a * sn);
    } else {
      const float2 v = reinterpret_cast<const float2*>(x)[idx];
      a = v.x; b = v.y;
      reinterpret_cast<float2*>(y)[idx] = make_float2(a * cs - b * sn, b * cs + a * sn);
    }
  }
}

// FixedPositionalEmbedding (position_embed.py:23-34): emb[t, :] = concat(freqs[t], freqs[t]),
// freqs[t, i] = t * inv_freq[i]; fp32 [n, dim].
__global__ void rotary_table_kernel(float* __restrict__ emb, int64_t n, int dim) {
  const int half = dim / 2;
  const int64_t total = n * half;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(idx % half);
    const int64_t t = idx / half;
    const float f = (float)t * rotary_inv_freq(i, dim);
    emb[t * dim + i] = f;
    emb[t * dim + half + i] = f;
  }
}

// dst fp32 [rows, cols] (row pitch ldd) (+)= src bf16 [rows, cols] (row pitch lds): the bf16 gradient of
// a row selection handed back to the fp32 stream (x[:, 0] of cvt.py:165)
__global__ void cast2d_bf16_f32_kernel(const bf16* __restrict__ src, int64_t lds, float* __restrict__ dst,
                                       int64_t ldd, int64_t rows, int64_t cols2, int accumulate) {
  const int64_t total = rows * cols2;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / cols2;
    const int64_t c = (i - r * cols2) * 2;
    float2 v = unpack_bf16(*reinterpret_cast<const uint32_t*>(src + r * lds + c));
    float2* d = reinterpret_cast<float2*>(dst + r * ldd + c);
    if (accumulate) { const float2 o = *d; v.x += o.x; v.y += o.y; }
    *d = v;
  }
}

static int ew_grid(int64_t work, int threads) {
  int64_t b = (work + threads - 1) / threads;
  const int64_t cap = 148 * 8;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

}  // namespace sav

using namespace sav;

extern "C" int sav_scale_rows_cols(const float* g, int64_t ldg, const float* col_scale,
                                   const float* row_scale, int64_t rows_per_scale, void* out,
                                   int64_t ldo, int64_t rows, int64_t cols, void* stream) {
  if (rows <= 0 || cols <= 0) return SAV_OK;
  if (cols % 4 || ldg % 4 || ldo % 4 || !aligned16(g) || !aligned16(col_scale) ||
      (reinterpret_cast<uintptr_t>(out) & 7))
    return set_error(SAV_ERR_ALIGN, "sav_scale_rows_cols: cols/ld %% 4 and 16 B alignment required");
  scale_rows_cols_kernel<<<ew_grid(rows * (cols / 4), 256), 256, 0,
                           reinterpret_cast<cudaStream_t>(stream)>>>(
      g, ldg, col_scale, row_scale, rows_per_scale > 0 ? rows_per_scale : 1,
      reinterpret_cast<bf16*>(out), ldo, rows, (int)(cols / 4));
  count_launch();
  return check_launch("sav_scale_rows_cols");
}

extern "C" int sav_layerscale_grad(const float* W, int64_t ldw, const float* dW, int64_t ldg,
                                   int64_t K, int64_t N, const float* bias, const float* dbias,
                                   const float* col_scale, float* dcol_scale, void* stream) {
  if (K <= 0 || N <= 0) return set_error(SAV_ERR_SHAPE, "sav_layerscale_grad: bad shape");
  if (!W || !dW || !col_scale || !dcol_scale || ((bias == nullptr) != (dbias == nullptr)))
    return set_error(SAV_ERR_ARG, "sav_layerscale_grad: W, dW, col_scale, dcol_scale required");
  layerscale_grad_kernel<<<dim3((unsigned)((N + 31) / 32), (unsigned)((K + kLsRows - 1) / kLsRows)), 256, 0,
                           reinterpret_cast<cudaStream_t>(stream)>>>(
      W, ldw, dW, ldg, K, N, bias, dbias, col_scale, dcol_scale);
  count_launch();
  return check_launch("sav_layerscale_grad");
}

extern "C" int sav_copy2d_f32(const float* src, int64_t lds, float* dst, int64_t ldd, int64_t rows,
                              int64_t cols, int32_t accumulate, void* stream) {
  if (rows <= 0 || cols <= 0) return SAV_OK;
  if (cols % 4 || lds % 4 || ldd % 4 || !aligned16(src) || !aligned16(dst))
    return set_error(SAV_ERR_ALIGN, "sav_copy2d_f32: cols/ld %% 4 and 16 B alignment required");
  copy2d_f32_kernel<<<ew_grid(rows * (cols / 4), 256), 256, 0,
                      reinterpret_cast<cudaStream_t>(stream)>>>(src, lds, dst, ldd, rows, cols / 4,
                                                                accumulate);
  count_launch();
  return check_launch("sav_copy2d_f32");
}

extern "C" int sav_rotary(const void* x, void* y, int64_t outer, int64_t n, int64_t inner, int64_t d,
                          int32_t is_fp32, int32_t inverse, void* stream) {
  if (outer <= 0 || n <= 0 || inner <= 0 || d <= 0) return SAV_OK;
  if (d % 2) return set_error(SAV_ERR_SHAPE, "sav_rotary: the rotated dimension must be even");
  if ((reinterpret_cast<uintptr_t>(x) & 7) || (reinterpret_cast<uintptr_t>(y) & 7))
    return set_error(SAV_ERR_ALIGN, "sav_rotary: 8 B alignment required");
  const int64_t work = outer * n * inner * (d / 2);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const float sign = inverse ? -1.f : 1.f;
  if (is_fp32)
    rotary_kernel<float><<<ew_grid(work, 256), 256, 0, st>>>(
        reinterpret_cast<const float*>(x), reinterpret_cast<float*>(y), outer, n, inner, (int)d, sign);
  else
    rotary_kernel<bf16><<<ew_grid(work, 256), 256, 0, st>>>(
        reinterpret_cast<const bf16*>(x), reinterpret_cast<bf16*>(y), outer, n, inner, (int)d, sign);
  count_launch();
  return check_launch("sav_rotary");
}

extern "C" int sav_rotary_table(float* emb, int64_t n, int64_t dim, void* stream) {
  if (n <= 0 || dim <= 0) return SAV_OK;
  if (dim % 2) return set_error(SAV_ERR_SHAPE, "sav_rotary_table: dim must be even");
  rotary_table_kernel<<<ew_grid(n * (dim / 2), 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      emb, n, (int)dim);
  count_launch();
  return check_launch("sav_rotary_table");
}

extern "C" int sav_cast2d_bf16_f32(const void* src, int64_t lds, float* dst, int64_t ldd, int64_t rows,
                                   int64_t cols, int32_t accumulate, void* stream) {
  if (rows <= 0 || cols <= 0) return SAV_OK;
  if (cols % 2 || lds % 2 || ldd % 2 || (reinterpret_cast<uintptr_t>(src) & 3) ||
      (reinterpret_cast<uintptr_t>(dst) & 7))
    return set_error(SAV_ERR_ALIGN, "sav_cast2d_bf16_f32: cols/ld %% 2 and 4/8 B alignment required");
  cast2d_bf16_f32_kernel<<<ew_grid(rows * (cols / 2), 256), 256, 0,
                           reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(src), lds, dst, ldd, rows, cols / 2, accumulate);
  count_launch();
  return check_launch("sav_cast2d_bf16_f32");
}

// ---- batched transpose with zero padding: out[b, c, r] = in[b, r, c] (+ residual[b, c, r]) for
// r < R, c < C; columns R .. out_cols-1 of every output row are written as zeros.  The token-mixing
// MLP of MLP-Mixer (mlp_mixer.py:18-23: '... l d -> ... d l', FFBlock over the token axis, and
// back) runs its two Dense layers as ordinary GEMMs over [B*C, L] rows; L = 196 is not a multiple
// of 8 (the GEMM's pitch unit), so the transposed activations and weights live in buffers padded to
// the next multiple of 8 whose pad columns are exact zeros.
namespace sav {
__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(bf16 v) { return __bfloat162float(v); }
__device__ __forceinline__ void from_f32(float& d, float v) { d = v; }
__device__ __forceinline__ void from_f32(bf16& d, float v) { d = __float2bfloat16_rn(v); }

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(256)
transpose_pad_kernel(const TIn* __restrict__ in, int64_t ld_in, TOut* __restrict__ out, int64_t ld_out,
                     const float* __restrict__ res, int64_t ld_res, int64_t R, int64_t C,
                     int64_t out_cols) {
  __shared__ float tile[32][33];
  const int64_t b = blockIdx.z;
  const int64_t r0 = (int64_t)blockIdx.x * 32, c0 = (int64_t)blockIdx.y * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;   // 32 x 8
  const TIn* src = in + b * R * ld_in;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t r = r0 + ty + i * 8, c = c0 + tx;
    float v = 0.f;
    if (r < R && c < C) v = to_f32(src[r * ld_in + c]);
    tile[ty + i * 8][tx] = v;
  }
  __syncthreads();
  TOut* dst = out + b * C * ld_out;
  const float* rs = res ? res + b * C * ld_res : nullptr;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t c = c0 + ty + i * 8, r = r0 + tx;
    if (c < C && r < out_cols) {
      float v = tile[tx][ty + i * 8];
      if (rs && r < R) v += rs[c * ld_res + r];
      from_f32(dst[c * ld_out + r], v);
    }
  }
}

// 64 x 64 tiles with two adjacent elements per access (bf16x2 = 4 bytes, fp32x2 = 8 bytes) on both
// sides: the 32 x 32 / one-element kernel above moves 64 bytes per warp row for bf16 and ran the
// activation transposes of Mixer-B/16 (38.5 MB each way) at half the HBM rate
// (profiles/r02ay_timeline_mixer_b_patch16.txt).  Needs even pitches and aligned bases.
__device__ __forceinline__ float2 ld2(const float* p) { return *reinterpret_cast<const float2*>(p); }
__device__ __forceinline__ float2 ld2(const bf16* p) { return unpack_bf16(*reinterpret_cast<const uint32_t*>(p)); }
__device__ __forceinline__ void st2(float* p, float a, float b) { *reinterpret_cast<float2*>(p) = make_float2(a, b); }
__device__ __forceinline__ void st2(bf16* p, float a, float b) { *reinterpret_cast<uint32_t*>(p) = pack_bf16(a, b); }

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(256)
transpose_pad64_kernel(const TIn* __restrict__ in, int64_t ld_in, TOut* __restrict__ out, int64_t ld_out,
                       const float* __restrict__ res, int64_t ld_res, int64_t R, int64_t C,
                       int64_t out_cols) {
  __shared__ float tile[64][65];
  const int64_t b = blockIdx.z;
  const int64_t r0 = (int64_t)blockIdx.x * 64, c0 = (int64_t)blockIdx.y * 64;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  const TIn* src = in + b * R * ld_in;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int rl = ty + i * 8;
    const int64_t r = r0 + rl, c = c0 + 2 * tx;
    float2 v = make_float2(0.f, 0.f);
    if (r < R) {
      if (c + 1 < C) v = ld2(src + r * ld_in + c);
      else if (c < C) v.x = to_f32(src[r * ld_in + c]);
    }
    tile[rl][2 * tx] = v.x;
    tile[rl][2 * tx + 1] = v.y;
  }
  __syncthreads();
  TOut* dst = out + b * C * ld_out;
  const float* rs = res ? res + b * C * ld_res : nullptr;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int cl = ty + i * 8;
    const int64_t c = c0 + cl, r = r0 + 2 * tx;
    if (c < C && r < out_cols) {
      float v0 = tile[2 * tx][cl], v1 = tile[2 * tx + 1][cl];
      if (rs) {
        if (r + 1 < R) { const float2 q = ld2(rs + c * ld_res + r); v0 += q.x; v1 += q.y; }
        else if (r < R) v0 += rs[c * ld_res + r];
      }
      if (r + 1 < out_cols) st2(dst + c * ld_out + r, v0, v1);
      else from_f32(dst[c * ld_out + r], v0);
    }
  }
}
}  // namespace sav

extern "C" int sav_transpose_pad(const void* in, int32_t in_fp32, int64_t ld_in, void* out,
                                 int32_t out_fp32, int64_t ld_out, const float* residual,
                                 int64_t ld_res, int64_t B, int64_t R, int64_t C, int64_t out_cols,
                                 void* stream) {
  using namespace sav;
  if (B <= 0 || R <= 0 || C <= 0) return SAV_OK;
  if (!in || !out || out_cols < R || ld_out < out_cols || ld_in < C || (residual && ld_res < R) ||
      B > 65535 || (C + 31) / 32 > 65535)
    return set_error(SAV_ERR_SHAPE, "sav_transpose_pad: R <= out_cols <= ld_out, C <= ld_in, B <= 65535");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // two elements per access when every pitch is even and the bases are aligned for the pair
  const uintptr_t am_in = in_fp32 ? 7 : 3, am_out = out_fp32 ? 7 : 3;
  if (R >= 32 && C >= 32 && ld_in % 2 == 0 && ld_out % 2 == 0 && (!residual || ld_res % 2 == 0) &&
      !(reinterpret_cast<uintptr_t>(in) & am_in) && !(reinterpret_cast<uintptr_t>(out) & am_out) &&
      !(reinterpret_cast<uintptr_t>(residual) & 7) && (R * ld_in) % 2 == 0 && (C * ld_out) % 2 == 0 &&
      (C * ld_res) % 2 == 0) {
    const dim3 grid64((unsigned)((out_cols + 63) / 64), (unsigned)((C + 63) / 64), (unsigned)B);
#define SAV_TP64(TI, TO)                                                                          \
  transpose_pad64_kernel<TI, TO><<<grid64, 256, 0, st>>>(reinterpret_cast<const TI*>(in), ld_in,   \
      reinterpret_cast<TO*>(out), ld_out, residual, ld_res, R, C, out_cols)
    if (in_fp32 && out_fp32) SAV_TP64(float, float);
    else if (in_fp32) SAV_TP64(float, bf16);
    else if (out_fp32) SAV_TP64(bf16, float);
    else SAV_TP64(bf16, bf16);
#undef SAV_TP64
    count_launch();
    return check_launch("sav_transpose_pad");
  }
  const dim3 grid((unsigned)((out_cols + 31) / 32), (unsigned)((C + 31) / 32), (unsigned)B), block(32, 8);
#define SAV_TP(TI, TO)                                                                          \
  transpose_pad_kernel<TI, TO><<<grid, block, 0, st>>>(reinterpret_cast<const TI*>(in), ld_in,  \
      reinterpret_cast<TO*>(out), ld_out, residual, ld_res, R, C, out_cols)
  if (in_fp32 && out_fp32) SAV_TP(float, float);
  else if (in_fp32) SAV_TP(float, bf16);
  else if (out_fp32) SAV_TP(bf16, float);
  else SAV_TP(bf16, bf16);
#undef SAV_TP
  count_launch();
  return check_launch("sav_transpose_pad");
}
