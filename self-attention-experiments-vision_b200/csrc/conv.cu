// CvT convolutional projection (cvt_attention.py:12-40): depthwise 3x3 conv (stride s, SAME, no
// bias) -> BatchNorm (batch statistics, momentum / running averages are host state) -> 1x1 conv.
// The 1x1 conv is a GEMM (gemm.cu); this file holds the channel-last, HBM-bound pieces around it:
//   dwconv3x3_fwd      x fp32 [B,H,W,C] (the residual stream: CvT has no LayerNorm before the
//                      attention, cvt.py:50-61) -> y bf16 [B,Ho,Wo,C], and per-channel sum / sum of
//                      squares of the fp32 results (BatchNorm statistics come for free)
//   bn_finalize        mean, rstd from the sums (biased variance, flax nn.BatchNorm)
//   bn_apply           z = (y - mean) rstd gamma + beta -> bf16 (the GEMM A operand)
//   bn_bwd_reduce      dbeta = sum dz, dgamma = sum dz yhat          (per channel, red.add)
//   bn_bwd_apply       dy = gamma rstd (dz - dbeta/n - yhat dgamma/n) -> bf16
//   dwconv3x3_bwd      dx (fp32, optionally accumulated: q, k and v all read the same input) and
//                      dw [3,3,C] (red.add) in one pass over dy
// Threads own 4 consecutive channels of one output pixel: 16-byte fp32 / 8-byte bf16 accesses,
// coalesced along C; per-channel reductions go registers -> shared -> one red.add per block.
#include "common.cuh"
#include "host.h"

namespace sav {

struct ConvGeom {
  int B, H, W, C, Ho, Wo, stride, pad_t, pad_l;
};

__device__ __forceinline__ void ld_bf16x4(const bf16* p, float (&v)[4]) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void st_bf16x4(bf16* p, const float (&v)[4]) {
  *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]));
}

// block = (C/4 channel groups [<= 128 threads in x]) x (pixels in y); per-channel partial sums
// are reduced over threadIdx.y in shared memory.
template <int NV>
__device__ __forceinline__ void block_channel_add(float (&v)[NV][4], float* const* dst, int c0,
                                                  float* smem, bool cok) {
  // smem: [blockDim.y][blockDim.x][NV*4]
  const int tx = threadIdx.x, ty = threadIdx.y;
  float* mine = smem + ((size_t)ty * blockDim.x + tx) * (NV * 4);
#pragma unroll
  for (int n = 0; n < NV; ++n)
#pragma unroll
    for (int j = 0; j < 4; ++j) mine[n * 4 + j] = v[n][j];
  __syncthreads();
  if (ty == 0 && cok) {   // (every thread reaches the barrier above: half-valid warps exist)
    for (int y = 1; y < (int)blockDim.y; ++y) {
      const float* o = smem + ((size_t)y * blockDim.x + tx) * (NV * 4);
#pragma unroll
      for (int n = 0; n < NV; ++n)
#pragma unroll
        for (int j = 0; j < 4; ++j) v[n][j] += o[n * 4 + j];
    }
#pragma unroll
    for (int n = 0; n < NV; ++n)
#pragma unroll
      for (int j = 0; j < 4; ++j) atomicAdd(dst[n] + c0 + j, v[n][j]);
  }
}

__global__ void dwconv3x3_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                     bf16* __restrict__ y, float* __restrict__ sum,
                                     float* __restrict__ sumsq, const ConvGeom g) {
  extern __shared__ float smem[];
  const int cg = blockIdx.x * blockDim.x + threadIdx.x;   // channel group
  const int c0 = cg * 4;
  const bool cok = c0 < g.C;
  float wt[9][4];
  if (cok) {
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float4 f = __ldg(reinterpret_cast<const float4*>(w + (size_t)t * g.C + c0));
      wt[t][0] = f.x; wt[t][1] = f.y; wt[t][2] = f.z; wt[t][3] = f.w;
    }
  }
  float acc[2][4] = {};
  const int64_t npix = (int64_t)g.B * g.Ho * g.Wo;
  for (int64_t pix = (int64_t)blockIdx.y * blockDim.y + threadIdx.y; pix < npix && cok;
       pix += (int64_t)gridDim.y * blockDim.y) {
    const int ox = (int)(pix % g.Wo);
    const int oy = (int)((pix / g.Wo) % g.Ho);
    const int64_t b = pix / ((int64_t)g.Wo * g.Ho);
    float o[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int iy = oy * g.stride + ky - g.pad_t;
      if (iy < 0 || iy >= g.H) continue;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int ix = ox * g.stride + kx - g.pad_l;
        if (ix < 0 || ix >= g.W) continue;
        const float4 v = *reinterpret_cast<const float4*>(x + ((b * g.H + iy) * g.W + ix) * g.C + c0);
        const float* ww = wt[ky * 3 + kx];
        o[0] = fmaf(v.x, ww[0], o[0]); o[1] = fmaf(v.y, ww[1], o[1]);
        o[2] = fmaf(v.z, ww[2], o[2]); o[3] = fmaf(v.w, ww[3], o[3]);
      }
    }
    st_bf16x4(y + pix * g.C + c0, o);
#pragma unroll
    for (int j = 0; j < 4; ++j) { acc[0][j] += o[j]; acc[1][j] = fmaf(o[j], o[j], acc[1][j]); }
  }
  float* dst[2] = {sum, sumsq};
  block_channel_add<2>(acc, dst, c0, smem, cok);
}

__global__ void bn_finalize_kernel(const float* __restrict__ sum, const float* __restrict__ sumsq,
                                   float inv_n, float eps, float* __restrict__ mean,
                                   float* __restrict__ rstd, float* __restrict__ var, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float m = sum[c] * inv_n;
  const float v = fmaxf(sumsq[c] * inv_n - m * m, 0.f);
  mean[c] = m;
  rstd[c] = rsqrtf(v + eps);
  if (var) var[c] = v;
}

__global__ void bn_apply_kernel(const bf16* __restrict__ y, const float* __restrict__ mean,
                                const float* __restrict__ rstd, const float* __restrict__ gamma,
                                const float* __restrict__ beta, bf16* __restrict__ z, int64_t rows,
                                int C4) {
  const int64_t total = rows * C4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c0 = (int)(i % C4) * 4;
    const int64_t off = (i / C4) * (C4 * 4) + c0;
    float v[4];
    ld_bf16x4(y + off, v);
    const float4 m = __ldg(reinterpret_cast<const float4*>(mean + c0));
    const float4 r = __ldg(reinterpret_cast<const float4*>(rstd + c0));
    const float4 ga = __ldg(reinterpret_cast<const float4*>(gamma + c0));
    const float4 be = __ldg(reinterpret_cast<const float4*>(beta + c0));
    v[0] = (v[0] - m.x) * r.x * ga.x + be.x; v[1] = (v[1] - m.y) * r.y * ga.y + be.y;
    v[2] = (v[2] - m.z) * r.z * ga.z + be.z; v[3] = (v[3] - m.w) * r.w * ga.w + be.w;
    st_bf16x4(z + off, v);
  }
}

__global__ void bn_bwd_reduce_kernel(const bf16* __restrict__ dz, const bf16* __restrict__ y,
                                     const float* __restrict__ mean, const float* __restrict__ rstd,
                                     float* __restrict__ dgamma, float* __restrict__ dbeta,
                                     int64_t rows, int C) {
  extern __shared__ float smem[];
  const int c0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const bool cok = c0 < C;
  float acc[2][4] = {};
  if (cok) {
    const float4 m = __ldg(reinterpret_cast<const float4*>(mean + c0));
    const float4 r = __ldg(reinterpret_cast<const float4*>(rstd + c0));
    const float mm[4] = {m.x, m.y, m.z, m.w}, rr[4] = {r.x, r.y, r.z, r.w};
    for (int64_t row = (int64_t)blockIdx.y * blockDim.y + threadIdx.y; row < rows;
         row += (int64_t)gridDim.y * blockDim.y) {
      float d[4], v[4];
      ld_bf16x4(dz + row * C + c0, d);
      ld_bf16x4(y + row * C + c0, v);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[0][j] = fmaf(d[j], (v[j] - mm[j]) * rr[j], acc[0][j]);
        acc[1][j] += d[j];
      }
    }
  }
  float* dst[2] = {dgamma, dbeta};
  block_channel_add<2>(acc, dst, c0, smem, cok);
}

// sums[0] = sum dz yhat, sums[1] = sum dz of THIS call (not the accumulated parameter gradient)
__global__ void bn_bwd_apply_kernel(const bf16* __restrict__ dz, const bf16* __restrict__ y,
                                    const float* __restrict__ mean, const float* __restrict__ rstd,
                                    const float* __restrict__ gamma, const float* __restrict__ s_dg,
                                    const float* __restrict__ s_db, float inv_n,
                                    bf16* __restrict__ dy, int64_t rows, int C4) {
  const int64_t total = rows * C4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c0 = (int)(i % C4) * 4;
    const int64_t off = (i / C4) * (C4 * 4) + c0;
    float d[4], v[4], o[4];
    ld_bf16x4(dz + off, d);
    ld_bf16x4(y + off, v);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float r = __ldg(rstd + c0 + j);
      const float yh = (v[j] - __ldg(mean + c0 + j)) * r;
      o[j] = __ldg(gamma + c0 + j) * r *
             (d[j] - __ldg(s_db + c0 + j) * inv_n - yh * __ldg(s_dg + c0 + j) * inv_n);
    }
    st_bf16x4(dy + off, o);
  }
}

// one thread per INPUT pixel x 4 channels: gathers the (at most 9) output pixels that read it
// (dx), and -- in the same pass -- its contribution x * dy to the 9 weight gradients.
__global__ void dwconv3x3_bwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                     const bf16* __restrict__ dy, float* __restrict__ dx,
                                     float* __restrict__ dw, int accumulate, const ConvGeom g) {
  extern __shared__ float smem[];
  const int c0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const bool cok = c0 < g.C;
  float wt[9][4], gw[9][4] = {};
  if (cok) {
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float4 f = __ldg(reinterpret_cast<const float4*>(w + (size_t)t * g.C + c0));
      wt[t][0] = f.x; wt[t][1] = f.y; wt[t][2] = f.z; wt[t][3] = f.w;
    }
  }
  const int64_t npix = (int64_t)g.B * g.H * g.W;
  for (int64_t pix = (int64_t)blockIdx.y * blockDim.y + threadIdx.y; pix < npix && cok;
       pix += (int64_t)gridDim.y * blockDim.y) {
    const int ix = (int)(pix % g.W);
    const int iy = (int)((pix / g.W) % g.H);
    const int64_t b = pix / ((int64_t)g.W * g.H);
    const float4 xv4 = *reinterpret_cast<const float4*>(x + pix * g.C + c0);
    const float xv[4] = {xv4.x, xv4.y, xv4.z, xv4.w};
    float o[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int ny = iy + g.pad_t - ky;            // = oy * stride
      if (ny < 0 || ny % g.stride) continue;
      const int oy = ny / g.stride;
      if (oy >= g.Ho) continue;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int nx = ix + g.pad_l - kx;
        if (nx < 0 || nx % g.stride) continue;
        const int ox = nx / g.stride;
        if (ox >= g.Wo) continue;
        float d[4];
        ld_bf16x4(dy + ((b * g.Ho + oy) * g.Wo + ox) * g.C + c0, d);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          o[j] = fmaf(d[j], wt[ky * 3 + kx][j], o[j]);
          gw[ky * 3 + kx][j] = fmaf(d[j], xv[j], gw[ky * 3 + kx][j]);
        }
      }
    }
    float4* dst = reinterpret_cast<float4*>(dx + pix * g.C + c0);
    float4 r = make_float4(o[0], o[1], o[2], o[3]);
    if (accumulate) {
      const float4 p = *dst;
      r.x += p.x; r.y += p.y; r.z += p.z; r.w += p.w;
    }
    *dst = r;
  }
  float* dst[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) dst[t] = dw + (size_t)t * g.C;
  block_channel_add<9>(gw, dst, c0, smem, cok);
}

__global__ void axpy2_kernel(const float* a, float* da, const float* b, float* db, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { da[i] += a[i]; db[i] += b[i]; }
}

static int same_pad(int in, int stride, int* out) {
  *out = (in + stride - 1) / stride;
  int total = (*out - 1) * stride + 3 - in;
  if (total < 0) total = 0;
  return total / 2;   // jax / flax 'SAME': the extra element goes to the high side
}

static int conv_geom(ConvGeom& g, int64_t B, int64_t H, int64_t W, int64_t C, int stride,
                     const char* who) {
  if (B <= 0 || H <= 0 || W <= 0 || C <= 0 || C % 4 || stride < 1 || stride > 2)
    return set_error(SAV_ERR_SHAPE, "%s: need C %% 4 == 0 and stride 1 or 2", who);
  g.B = (int)B; g.H = (int)H; g.W = (int)W; g.C = (int)C; g.stride = stride;
  g.pad_t = same_pad(g.H, stride, &g.Ho);
  g.pad_l = same_pad(g.W, stride, &g.Wo);
  return SAV_OK;
}

// block (tx = channel groups, ty = pixels), grid.y capped so that per-block atomics stay few
static void conv_launch_shape(int C, int64_t npix, dim3& grid, dim3& block) {
  const int groups = C / 4;
  const int tx = groups < 32 ? groups : 32;
  const int ty = 256 / tx;
  block = dim3(tx, ty);
  int64_t gy = (npix + ty - 1) / ty;
  const int64_t cap = 148 * 4;
  grid = dim3((groups + tx - 1) / tx, (unsigned)(gy < cap ? (gy < 1 ? 1 : gy) : cap));
}

static int ew_blocks(int64_t work) {
  int64_t b = (work + 255) / 256;
  const int64_t cap = 148 * 8;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

}  // namespace sav

using namespace sav;

extern "C" int sav_dwconv3x3_out_hw(int64_t H, int64_t W, int32_t stride, int64_t* Ho, int64_t* Wo) {
  int ho, wo;
  same_pad((int)H, stride, &ho);
  same_pad((int)W, stride, &wo);
  *Ho = ho; *Wo = wo;
  return SAV_OK;
}

extern "C" int sav_dwconv3x3_fwd(const float* x, const float* w, void* y, float* sum, float* sumsq,
                                 int64_t B, int64_t H, int64_t W, int64_t C, int32_t stride,
                                 void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  ConvGeom g;
  int rc = conv_geom(g, B, H, W, C, stride, "sav_dwconv3x3_fwd");
  if (rc) return rc;
  if (!aligned16(x) || !aligned16(w) || (reinterpret_cast<uintptr_t>(y) & 7))
    return set_error(SAV_ERR_ALIGN, "sav_dwconv3x3_fwd: alignment");
  if ((rc = sav_fill_f32(sum, 0.f, C, stream)) || (rc = sav_fill_f32(sumsq, 0.f, C, stream))) return rc;
  dim3 grid, block;
  conv_launch_shape(g.C, (int64_t)g.B * g.Ho * g.Wo, grid, block);
  const size_t sm = (size_t)block.x * block.y * 8 * sizeof(float);
  dwconv3x3_fwd_kernel<<<grid, block, sm, st>>>(x, w, reinterpret_cast<bf16*>(y), sum, sumsq, g);
  count_launch();
  return check_launch("sav_dwconv3x3_fwd");
}

extern "C" int sav_batchnorm_fwd(const void* y, const float* sum, const float* sumsq,
                                 const float* gamma, const float* beta, float eps, void* z,
                                 float* mean, float* rstd, float* var, int64_t rows, int64_t C,
                                 void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (rows <= 0 || C <= 0 || C % 4) return set_error(SAV_ERR_SHAPE, "sav_batchnorm_fwd: C %% 4");
  bn_finalize_kernel<<<(int)((C + 127) / 128), 128, 0, st>>>(sum, sumsq, 1.f / (float)rows, eps, mean,
                                                             rstd, var, (int)C);
  count_launch();
  bn_apply_kernel<<<ew_blocks(rows * (C / 4)), 256, 0, st>>>(
      reinterpret_cast<const bf16*>(y), mean, rstd, gamma, beta, reinterpret_cast<bf16*>(z), rows,
      (int)(C / 4));
  count_launch();
  return check_launch("sav_batchnorm_fwd");
}

extern "C" int sav_bn_finalize(const float* sum, const float* sumsq, int64_t rows, float eps, float* mean,
                               float* rstd, float* var, int64_t C, void* stream) {
  if (rows <= 0 || C <= 0) return set_error(SAV_ERR_SHAPE, "sav_bn_finalize: bad shape");
  bn_finalize_kernel<<<(int)((C + 127) / 128), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      sum, sumsq, 1.f / (float)rows, eps, mean, rstd, var, (int)C);
  count_launch();
  return check_launch("sav_bn_finalize");
}

extern "C" int sav_batchnorm_bwd(const void* dz, const void* y, const float* mean, const float* rstd,
                                 const float* gamma, float* dgamma, float* dbeta, float* scratch,
                                 void* dy, int64_t rows, int64_t C, void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (rows <= 0 || C <= 0 || C % 4) return set_error(SAV_ERR_SHAPE, "sav_batchnorm_bwd: C %% 4");
  int rc = sav_fill_f32(scratch, 0.f, 2 * C, stream);
  if (rc) return rc;
  dim3 grid, block;
  conv_launch_shape((int)C, rows, grid, block);
  const size_t sm = (size_t)block.x * block.y * 8 * sizeof(float);
  bn_bwd_reduce_kernel<<<grid, block, sm, st>>>(reinterpret_cast<const bf16*>(dz),
                                                reinterpret_cast<const bf16*>(y), mean, rstd, scratch,
                                                scratch + C, rows, (int)C);
  count_launch();
  bn_bwd_apply_kernel<<<ew_blocks(rows * (C / 4)), 256, 0, st>>>(
      reinterpret_cast<const bf16*>(dz), reinterpret_cast<const bf16*>(y), mean, rstd, gamma, scratch,
      scratch + C, 1.f / (float)rows, reinterpret_cast<bf16*>(dy), rows, (int)(C / 4));
  count_launch();
  // parameter gradients accumulate (the arena is zeroed once per step)
  axpy2_kernel<<<(int)((C + 255) / 256), 256, 0, st>>>(scratch, dgamma, scratch + C, dbeta, (int)C);
  count_launch();
  return check_launch("sav_batchnorm_bwd");
}

extern "C" int sav_dwconv3x3_bwd(const float* x, const float* w, const void* dy, float* dx, float* dw,
                                 int64_t B, int64_t H, int64_t W, int64_t C, int32_t stride,
                                 int32_t accumulate_dx, void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  ConvGeom g;
  int rc = conv_geom(g, B, H, W, C, stride, "sav_dwconv3x3_bwd");
  if (rc) return rc;
  if (!aligned16(x) || !aligned16(w) || !aligned16(dx) || (reinterpret_cast<uintptr_t>(dy) & 7))
    return set_error(SAV_ERR_ALIGN, "sav_dwconv3x3_bwd: alignment");
  dim3 grid, block;
  conv_launch_shape(g.C, (int64_t)g.B * g.H * g.W, grid, block);
  const size_t sm = (size_t)block.x * block.y * 36 * sizeof(float);
  if (sm > 48 * 1024) {  // 256 threads x 36 floats = 36 KB: fits; guard against shape changes
    return set_error(SAV_ERR_SHAPE, "sav_dwconv3x3_bwd: shared memory");
  }
  dwconv3x3_bwd_kernel<<<grid, block, sm, st>>>(x, w, reinterpret_cast<const bf16*>(dy), dx, dw,
                                                accumulate_dx, g);
  count_launch();
  return check_launch("sav_dwconv3x3_bwd");
}
