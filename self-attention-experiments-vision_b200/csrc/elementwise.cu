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

// out[n] += (sum_k W[k,n] G[k,n] + b[n] gb[n]) / cs[n];  one block per 32 columns, 8 row lanes
__global__ void layerscale_grad_kernel(const float* __restrict__ W, int64_t ldw,
                                       const float* __restrict__ G, int64_t ldg, int64_t K, int64_t N,
                                       const float* __restrict__ b, const float* __restrict__ gb,
                                       const float* __restrict__ cs, float* __restrict__ out) {
  __shared__ float red[8][32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t n = (int64_t)blockIdx.x * 32 + tx;
  float acc = 0.f;
  if (n < N)
    for (int64_t k = ty; k < K; k += 8) acc = fmaf(W[k * ldw + n], G[k * ldg + n], acc);
  red[ty][tx] = acc;
  __syncthreads();
  if (ty == 0 && n < N) {
    for (int j = 1; j < 8; ++j) acc += red[j][tx];
    if (b) acc = fmaf(b[n], gb[n], acc);
    const float s = cs[n];
    if (s != 0.f) out[n] += acc / s;
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
  layerscale_grad_kernel<<<(int)((N + 31) / 32), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
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
