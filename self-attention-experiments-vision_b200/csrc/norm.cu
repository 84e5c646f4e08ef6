// LayerNorm forward / backward: warp-per-row, 16-byte vector loads, fp32 statistics via warp
// shuffles.  HBM-roofline kernels (algorithmic bytes: fwd 2*rows*C*2 B, bwd 3*rows*C*2 B).
// Semantics follow Flax nn.LayerNorm as used at vit.py:19,26,57: var = E[x^2] - E[x]^2,
// y = (x - mean) * rsqrt(var + eps) * scale + bias, computed in fp32, stored bf16.
#include "common.cuh"
#include "host.h"

namespace sav {

constexpr int LN_MAX_CHUNKS = 4;  // 16-byte chunks per lane -> C <= 32*4*8 = 1024

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void load8(const bf16* p, float (&f)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 t = unpack_bf16(w[j]);
    f[2 * j] = t.x;
    f[2 * j + 1] = t.y;
  }
}
__device__ __forceinline__ void store8(bf16* p, const float (&f)[8]) {
  *reinterpret_cast<uint4*>(p) = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]),
                                            pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

template <int CHUNKS>
__global__ void __launch_bounds__(256)
ln_fwd_kernel(const bf16* __restrict__ x, const float* __restrict__ scale,
              const float* __restrict__ bias, bf16* __restrict__ y, float* __restrict__ mean_out,
              float* __restrict__ rstd_out, int64_t rows, int C, float eps) {
  const int lane = threadIdx.x & 31;
  const int64_t warp_global = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int64_t warps_total = (int64_t)gridDim.x * 8;
  const int nchunk = C / 8;
  for (int64_t row = warp_global; row < rows; row += warps_total) {
    float v[CHUNKS][8];
    float s = 0.f, ss = 0.f;
#pragma unroll
    for (int i = 0; i < CHUNKS; ++i) {
      const int ch = lane + i * 32;
      if (ch < nchunk) {
        load8(x + row * C + ch * 8, v[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) { s += v[i][j]; ss += v[i][j] * v[i][j]; }
      }
    }
    s = warp_sum(s);
    ss = warp_sum(ss);
    const float mean = s / C;
    const float var = fmaxf(ss / C - mean * mean, 0.f);
    const float rstd = rsqrtf(var + eps);
    if (lane == 0) { mean_out[row] = mean; rstd_out[row] = rstd; }
#pragma unroll
    for (int i = 0; i < CHUNKS; ++i) {
      const int ch = lane + i * 32;
      if (ch < nchunk) {
        float o[8];
        const float4 g0 = *reinterpret_cast<const float4*>(scale + ch * 8);
        const float4 g1 = *reinterpret_cast<const float4*>(scale + ch * 8 + 4);
        const float4 b0 = *reinterpret_cast<const float4*>(bias + ch * 8);
        const float4 b1 = *reinterpret_cast<const float4*>(bias + ch * 8 + 4);
        const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * g[j] + b[j];
        store8(y + row * C + ch * 8, o);
      }
    }
  }
}

// dx = rstd * (g - mean_c(g) - xhat * mean_c(g * xhat)),  g = dy * scale   (+ dx_add)
// per-block partial dscale/dbias -> partials[block][2][C]
template <int CHUNKS>
__global__ void __launch_bounds__(256)
ln_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x,
              const float* __restrict__ scale, const float* __restrict__ mean_in,
              const float* __restrict__ rstd_in, const bf16* __restrict__ dx_add,
              bf16* __restrict__ dx, float* __restrict__ partials, int64_t rows, int C) {
  extern __shared__ float sacc[];  // [2][C]
  const int lane = threadIdx.x & 31;
  const int64_t warp_global = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int64_t warps_total = (int64_t)gridDim.x * 8;
  const int nchunk = C / 8;
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) sacc[i] = 0.f;
  __syncthreads();
  float ds[CHUNKS][8], db[CHUNKS][8], g[CHUNKS][8];
#pragma unroll
  for (int i = 0; i < CHUNKS; ++i) {
    const int ch = lane + i * 32;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      ds[i][j] = 0.f;
      db[i][j] = 0.f;
      g[i][j] = (ch < nchunk) ? scale[ch * 8 + j] : 0.f;
    }
  }
  for (int64_t row = warp_global; row < rows; row += warps_total) {
    const float mean = mean_in[row], rstd = rstd_in[row];
    float xh[CHUNKS][8], gy[CHUNKS][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < CHUNKS; ++i) {
      const int ch = lane + i * 32;
      if (ch < nchunk) {
        float xv[8], dv[8];
        load8(x + row * C + ch * 8, xv);
        load8(dy + row * C + ch * 8, dv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (xv[j] - mean) * rstd;
          gy[i][j] = dv[j] * g[i][j];
          s1 += gy[i][j];
          s2 += gy[i][j] * xh[i][j];
          ds[i][j] += dv[j] * xh[i][j];
          db[i][j] += dv[j];
        }
      }
    }
    s1 = warp_sum(s1) / C;
    s2 = warp_sum(s2) / C;
#pragma unroll
    for (int i = 0; i < CHUNKS; ++i) {
      const int ch = lane + i * 32;
      if (ch < nchunk) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (gy[i][j] - s1 - xh[i][j] * s2);
        if (dx_add) {
          float a[8];
          load8(dx_add + row * C + ch * 8, a);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += a[j];
        }
        store8(dx + row * C + ch * 8, o);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < CHUNKS; ++i) {
    const int ch = lane + i * 32;
    if (ch < nchunk) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        atomicAdd(&sacc[ch * 8 + j], ds[i][j]);
        atomicAdd(&sacc[C + ch * 8 + j], db[i][j]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x)
    partials[(int64_t)blockIdx.x * 2 * C + i] = sacc[i];
}

__global__ void ln_bwd_reduce_kernel(const float* __restrict__ partials, int nblocks, int C,
                                     float* __restrict__ dscale, float* __restrict__ dbias,
                                     int accumulate) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;  // over 2*C
  if (i >= 2 * C) return;
  float s = 0.f;
  for (int b = 0; b < nblocks; ++b) s += partials[(int64_t)b * 2 * C + i];
  float* out = (i < C) ? (dscale + i) : (dbias + (i - C));
  *out = accumulate ? (*out + s) : s;
}

static int ln_grid(int64_t rows) {
  int64_t g = (rows + 7) / 8;
  const int64_t cap = 148 * 4;
  return (int)(g < cap ? (g < 1 ? 1 : g) : cap);
}

}  // namespace sav

using namespace sav;

extern "C" int sav_layernorm_fwd(const void* x, const float* scale, const float* bias, void* y,
                                 float* mean, float* rstd, int64_t rows, int64_t C, float eps,
                                 void* stream) {
  if (rows <= 0 || C <= 0 || C % 8 || C > 32 * LN_MAX_CHUNKS * 8)
    return set_error(SAV_ERR_SHAPE, "sav_layernorm_fwd: need C %% 8 == 0 and C <= 1024 (C=%lld)", (long long)C);
  if (!aligned16(x) || !aligned16(y) || !aligned16(scale) || !aligned16(bias))
    return set_error(SAV_ERR_ALIGN, "sav_layernorm_fwd: pointers must be 16 B aligned");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = ln_grid(rows);
  const int chunks = (int)((C / 8 + 31) / 32);
#define LN_FWD(CH)                                                                              \
  ln_fwd_kernel<CH><<<grid, 256, 0, st>>>(reinterpret_cast<const bf16*>(x), scale, bias,        \
                                           reinterpret_cast<bf16*>(y), mean, rstd, rows, (int)C, eps)
  switch (chunks) {
    case 1: LN_FWD(1); break;
    case 2: LN_FWD(2); break;
    case 3: LN_FWD(3); break;
    default: LN_FWD(4); break;
  }
#undef LN_FWD
  count_launch();
  return check_launch("sav_layernorm_fwd");
}

extern "C" int64_t sav_layernorm_bwd_workspace_bytes(int64_t rows, int64_t C) {
  return (int64_t)ln_grid(rows) * 2 * C * 4;
}

extern "C" int sav_layernorm_bwd(const void* dy, const void* x, const float* scale,
                                 const float* mean, const float* rstd, const void* dx_add,
                                 void* dx, float* dscale, float* dbias, float* partials,
                                 int64_t rows, int64_t C, int32_t accumulate, void* stream) {
  if (rows <= 0 || C <= 0 || C % 8 || C > 32 * LN_MAX_CHUNKS * 8)
    return set_error(SAV_ERR_SHAPE, "sav_layernorm_bwd: need C %% 8 == 0 and C <= 1024");
  if (!aligned16(dy) || !aligned16(x) || !aligned16(dx) || !aligned16(dx_add))
    return set_error(SAV_ERR_ALIGN, "sav_layernorm_bwd: pointers must be 16 B aligned");
  if (!partials) return set_error(SAV_ERR_ARG, "sav_layernorm_bwd: workspace required");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = ln_grid(rows);
  const int chunks = (int)((C / 8 + 31) / 32);
  const size_t sm = 2 * C * sizeof(float);
#define LN_BWD(CH)                                                                             \
  ln_bwd_kernel<CH><<<grid, 256, sm, st>>>(                                                    \
      reinterpret_cast<const bf16*>(dy), reinterpret_cast<const bf16*>(x), scale, mean, rstd,  \
      reinterpret_cast<const bf16*>(dx_add), reinterpret_cast<bf16*>(dx), partials, rows, (int)C)
  switch (chunks) {
    case 1: LN_BWD(1); break;
    case 2: LN_BWD(2); break;
    case 3: LN_BWD(3); break;
    default: LN_BWD(4); break;
  }
#undef LN_BWD
  count_launch();
  int rc = check_launch("sav_layernorm_bwd");
  if (rc) return rc;
  ln_bwd_reduce_kernel<<<(int)((2 * C + 255) / 256), 256, 0, st>>>(partials, grid, (int)C, dscale,
                                                                  dbias, accumulate);
  count_launch();
  return check_launch("sav_layernorm_bwd reduce");
}
