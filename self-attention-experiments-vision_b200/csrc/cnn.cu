// Channel-last (NHWC, bf16) pieces of the BoTNet conv stages (botnet.py:17-67,202-331) around the
// tcgen05 GEMMs (1x1 convs directly, 3x3 / 7x7 through im2col, conv2d.cu).  All HBM-bound, fp32
// arithmetic, bf16 storage, two channels per thread (4-byte accesses, coalesced along C):
//   bn_act          z = act(BatchNorm(y))                      botnet.py:44-50 (norm + activation_fn)
//   bn_act_bwd      reduce (dbeta, dgamma) + apply, act' recomputed from y (nothing saved but y)
//   gate_res_act    out = act(residual + x * gate[b, c])        botnet.py:54-67 (SE gate, skip, act)
//   gate_res_act_bwd dgate[b,c] = sum_hw dpre x;  dx = dpre gate;  dresidual = dpre
//   spatial_mean    [B, HW, C] -> [B, C] and its reverse mode   jnp.mean(.., (1, 2)), SE squeeze
//   maxpool 3x3/2   nn.max_pool(.., (3,3), strides (2,2), 'SAME') with the arg-max kept (1 byte)
//   act             swish / sigmoid / relu on small tensors (SE hidden layer, SE gate)
#include "common.cuh"
#include "host.h"

namespace sav {

enum { ACT_NONE = 0, ACT_SWISH = 1, ACT_SIGMOID = 2, ACT_RELU = 3, ACT_GELU = 4 };

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }
__device__ __forceinline__ float act_apply(float x, int kind) {
  if (kind == ACT_SWISH) return x * sigmoidf_(x);
  if (kind == ACT_SIGMOID) return sigmoidf_(x);
  if (kind == ACT_RELU) return fmaxf(x, 0.f);
  if (kind == ACT_GELU) return gelu_tanh(x);      // jax.nn.gelu default (leff.py:13,43,55,61)
  return x;
}
// d act(x) / dx
__device__ __forceinline__ float act_grad(float x, int kind) {
  if (kind == ACT_SWISH) { const float s = sigmoidf_(x); return s * (1.f + x * (1.f - s)); }
  if (kind == ACT_SIGMOID) { const float s = sigmoidf_(x); return s * (1.f - s); }
  if (kind == ACT_RELU) return x > 0.f ? 1.f : 0.f;
  if (kind == ACT_GELU) return dgelu_tanh(x);
  return 1.f;
}

__global__ void act_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, int64_t n2, int kind) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n2; i += (int64_t)gridDim.x * blockDim.x) {
    const float2 v = unpack_bf16(reinterpret_cast<const uint32_t*>(x)[i]);
    reinterpret_cast<uint32_t*>(y)[i] = pack_bf16(act_apply(v.x, kind), act_apply(v.y, kind));
  }
}
__global__ void act_bwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dy,
                               bf16* __restrict__ dx, int64_t n2, int kind) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n2; i += (int64_t)gridDim.x * blockDim.x) {
    const float2 v = unpack_bf16(reinterpret_cast<const uint32_t*>(x)[i]);
    const float2 g = unpack_bf16(reinterpret_cast<const uint32_t*>(dy)[i]);
    reinterpret_cast<uint32_t*>(dx)[i] = pack_bf16(g.x * act_grad(v.x, kind), g.y * act_grad(v.y, kind));
  }
}

// z = act((y - mean) rstd gamma + beta)
__global__ void bn_act_fwd_kernel(const bf16* __restrict__ y, const float* __restrict__ mean,
                                  const float* __restrict__ rstd, const float* __restrict__ gamma,
                                  const float* __restrict__ beta, bf16* __restrict__ z, int64_t rows,
                                  int C2, int kind) {
  const int64_t total = rows * C2;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C2) * 2;
    const float2 v = unpack_bf16(reinterpret_cast<const uint32_t*>(y)[i]);
    const float a = (v.x - mean[c]) * rstd[c] * gamma[c] + beta[c];
    const float b = (v.y - mean[c + 1]) * rstd[c + 1] * gamma[c + 1] + beta[c + 1];
    reinterpret_cast<uint32_t*>(z)[i] = pack_bf16(act_apply(a, kind), act_apply(b, kind));
  }
}

// per channel: s1 = sum dbn, s2 = sum dbn yhat with dbn = dz act'(bn); block = 64 channel pairs x 4
// row lanes (same shape as bn_stats_kernel)
__global__ void __launch_bounds__(256)
bn_act_bwd_reduce_kernel(const bf16* __restrict__ dz, const bf16* __restrict__ y,
                         const float* __restrict__ mean, const float* __restrict__ rstd,
                         const float* __restrict__ gamma, const float* __restrict__ beta, int64_t rows,
                         int C, int kind, float* __restrict__ s1, float* __restrict__ s2) {
  __shared__ float red[4][64][4];
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  const int c = (blockIdx.x * 64 + tx) * 2;
  float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
  if (c < C) {
    const float m0 = mean[c], m1 = mean[c + 1], r0 = rstd[c], r1 = rstd[c + 1];
    const float g0 = gamma[c], g1 = gamma[c + 1], e0 = beta[c], e1 = beta[c + 1];
    for (int64_t r = (int64_t)blockIdx.y * 4 + ty; r < rows; r += (int64_t)gridDim.y * 4) {
      const float2 v = unpack_bf16(*reinterpret_cast<const uint32_t*>(y + r * C + c));
      const float2 g = unpack_bf16(*reinterpret_cast<const uint32_t*>(dz + r * C + c));
      const float h0 = (v.x - m0) * r0, h1 = (v.y - m1) * r1;
      const float d0 = g.x * act_grad(h0 * g0 + e0, kind), d1 = g.y * act_grad(h1 * g1 + e1, kind);
      a0 += d0; a1 += d1; b0 += d0 * h0; b1 += d1 * h1;
    }
  }
  red[ty][tx][0] = a0; red[ty][tx][1] = a1; red[ty][tx][2] = b0; red[ty][tx][3] = b1;
  __syncthreads();
  if (ty == 0 && c < C) {
    for (int k = 1; k < 4; ++k) {
      a0 += red[k][tx][0]; a1 += red[k][tx][1]; b0 += red[k][tx][2]; b1 += red[k][tx][3];
    }
    atomicAdd(s1 + c, a0); atomicAdd(s1 + c + 1, a1);
    atomicAdd(s2 + c, b0); atomicAdd(s2 + c + 1, b1);
  }
}

// dy = gamma rstd (dbn - s1/n - yhat s2/n)
__global__ void bn_act_bwd_apply_kernel(const bf16* __restrict__ dz, const bf16* __restrict__ y,
                                        const float* __restrict__ mean, const float* __restrict__ rstd,
                                        const float* __restrict__ gamma, const float* __restrict__ beta,
                                        const float* __restrict__ s1, const float* __restrict__ s2,
                                        float inv_n, bf16* __restrict__ dy, int64_t rows, int C2, int kind) {
  const int64_t total = rows * C2;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C2) * 2;
    const float2 v = unpack_bf16(reinterpret_cast<const uint32_t*>(y)[i]);
    const float2 g = unpack_bf16(reinterpret_cast<const uint32_t*>(dz)[i]);
    const float h0 = (v.x - mean[c]) * rstd[c], h1 = (v.y - mean[c + 1]) * rstd[c + 1];
    const float d0 = g.x * act_grad(h0 * gamma[c] + beta[c], kind);
    const float d1 = g.y * act_grad(h1 * gamma[c + 1] + beta[c + 1], kind);
    reinterpret_cast<uint32_t*>(dy)[i] =
        pack_bf16(gamma[c] * rstd[c] * (d0 - s1[c] * inv_n - h0 * s2[c] * inv_n),
                  gamma[c + 1] * rstd[c + 1] * (d1 - s1[c + 1] * inv_n - h1 * s2[c + 1] * inv_n));
  }
}

__global__ void axpy2b_kernel(const float* __restrict__ a, float* __restrict__ da,
                              const float* __restrict__ b, float* __restrict__ db, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { da[i] += a[i]; db[i] += b[i]; }
}

// out = act(residual + x * gate[b, c])   (gate / residual optional)
__global__ void gate_res_act_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ gate,
                                        const bf16* __restrict__ res, bf16* __restrict__ out,
                                        int64_t rows, int64_t rows_per_sample, int C2, int kind) {
  const int64_t total = rows * C2;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / C2;
    const int c2 = (int)(i - r * C2);
    float2 v = unpack_bf16(reinterpret_cast<const uint32_t*>(x)[i]);
    if (gate) {
      const float2 s = unpack_bf16(reinterpret_cast<const uint32_t*>(gate)[(r / rows_per_sample) * C2 + c2]);
      v.x *= s.x; v.y *= s.y;
    }
    if (res) {
      const float2 q = unpack_bf16(reinterpret_cast<const uint32_t*>(res)[i]);
      v.x += q.x; v.y += q.y;
    }
    reinterpret_cast<uint32_t*>(out)[i] = pack_bf16(act_apply(v.x, kind), act_apply(v.y, kind));
  }
}

// dpre = dout act'(pre), pre = residual + x gate;  dx = dpre gate, dres = dpre,
// dgate[b,c] += sum_rows-of-sample dpre x   (fp32 accumulator [B, C], zeroed by the launcher)
// grid = (channel-pair blocks of 64, sample, row chunks); 4 row lanes per block
__global__ void __launch_bounds__(256)
gate_res_act_bwd_kernel(const bf16* __restrict__ dout, const bf16* __restrict__ x,
                        const bf16* __restrict__ gate, const bf16* __restrict__ res,
                        bf16* __restrict__ dx, bf16* __restrict__ dres, float* __restrict__ dgate,
                        int64_t rows_per_sample, int C, int kind) {
  __shared__ float red[4][64][2];
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  const int c = (blockIdx.x * 64 + tx) * 2;
  const int64_t b = blockIdx.y;
  float a0 = 0.f, a1 = 0.f;
  if (c < C) {
    float2 s = make_float2(1.f, 1.f);
    if (gate) s = unpack_bf16(*reinterpret_cast<const uint32_t*>(gate + b * C + c));
    for (int64_t rr = (int64_t)blockIdx.z * 4 + ty; rr < rows_per_sample; rr += (int64_t)gridDim.z * 4) {
      const int64_t off = (b * rows_per_sample + rr) * C + c;
      const float2 v = unpack_bf16(*reinterpret_cast<const uint32_t*>(x + off));
      const float2 g = unpack_bf16(*reinterpret_cast<const uint32_t*>(dout + off));
      float p0 = v.x * s.x, p1 = v.y * s.y;
      if (res) {
        const float2 q = unpack_bf16(*reinterpret_cast<const uint32_t*>(res + off));
        p0 += q.x; p1 += q.y;
      }
      const float d0 = g.x * act_grad(p0, kind), d1 = g.y * act_grad(p1, kind);
      *reinterpret_cast<uint32_t*>(dx + off) = pack_bf16(d0 * s.x, d1 * s.y);
      if (dres) *reinterpret_cast<uint32_t*>(dres + off) = pack_bf16(d0, d1);
      a0 += d0 * v.x; a1 += d1 * v.y;
    }
  }
  if (dgate) {
    red[ty][tx][0] = a0; red[ty][tx][1] = a1;
    __syncthreads();
    if (ty == 0 && c < C) {
      for (int k = 1; k < 4; ++k) { a0 += red[k][tx][0]; a1 += red[k][tx][1]; }
      atomicAdd(dgate + b * C + c, a0);
      atomicAdd(dgate + b * C + c + 1, a1);
    }
  }
}

// out[b, c] = mean over the sample's rows;  grid = (channel-pair blocks, sample)
__global__ void __launch_bounds__(256)
spatial_mean_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ out, int64_t rows_per_sample, int C) {
  __shared__ float red[4][64][2];
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  const int c = (blockIdx.x * 64 + tx) * 2;
  const int64_t b = blockIdx.y;
  float a0 = 0.f, a1 = 0.f;
  if (c < C)
    for (int64_t rr = ty; rr < rows_per_sample; rr += 4) {
      const float2 v = unpack_bf16(*reinterpret_cast<const uint32_t*>(x + (b * rows_per_sample + rr) * C + c));
      a0 += v.x; a1 += v.y;
    }
  red[ty][tx][0] = a0; red[ty][tx][1] = a1;
  __syncthreads();
  if (ty == 0 && c < C) {
    for (int k = 1; k < 4; ++k) { a0 += red[k][tx][0]; a1 += red[k][tx][1]; }
    const float inv = 1.f / (float)rows_per_sample;
    *reinterpret_cast<uint32_t*>(out + b * C + c) = pack_bf16(a0 * inv, a1 * inv);
  }
}
__global__ void spatial_mean_bwd_kernel(const bf16* __restrict__ dout, bf16* __restrict__ dx, int64_t rows,
                                        int64_t rows_per_sample, int C2) {
  const int64_t total = rows * C2;
  const float inv = 1.f / (float)rows_per_sample;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / C2;
    const int c2 = (int)(i - r * C2);
    const float2 g = unpack_bf16(reinterpret_cast<const uint32_t*>(dout)[(r / rows_per_sample) * C2 + c2]);
    reinterpret_cast<uint32_t*>(dx)[i] = pack_bf16(g.x * inv, g.y * inv);
  }
}

struct PoolGeom { int B, H, W, C, Ho, Wo, pt, pl; };

// nn.max_pool((3,3), strides (2,2), 'SAME'): -inf padding; the first maximum in window order wins
__global__ void maxpool_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y,
                                   uint8_t* __restrict__ arg, PoolGeom g) {
  const int64_t total = (int64_t)g.B * g.Ho * g.Wo * g.C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % g.C);
    int64_t t = i / g.C;
    const int wo = (int)(t % g.Wo);
    t /= g.Wo;
    const int ho = (int)(t % g.Ho);
    const int b = (int)(t / g.Ho);
    float best = -INFINITY;
    int bi = 0;
    for (int ky = 0; ky < 3; ++ky) {
      const int h = ho * 2 - g.pt + ky;
      if (h < 0 || h >= g.H) continue;
      for (int kx = 0; kx < 3; ++kx) {
        const int w = wo * 2 - g.pl + kx;
        if (w < 0 || w >= g.W) continue;
        const float v = __bfloat162float(x[(((int64_t)b * g.H + h) * g.W + w) * g.C + c]);
        if (v > best) { best = v; bi = ky * 3 + kx; }
      }
    }
    y[i] = __float2bfloat16(best);
    arg[i] = (uint8_t)bi;
  }
}
// gather form: every input element sums the gradients of the <= 4 windows whose arg-max it is
__global__ void maxpool_bwd_kernel(const bf16* __restrict__ dy, const uint8_t* __restrict__ arg,
                                   bf16* __restrict__ dx, PoolGeom g) {
  const int64_t total = (int64_t)g.B * g.H * g.W * g.C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % g.C);
    int64_t t = i / g.C;
    const int w = (int)(t % g.W);
    t /= g.W;
    const int h = (int)(t % g.H);
    const int b = (int)(t / g.H);
    float acc = 0.f;
    for (int ky = 0; ky < 3; ++ky) {
      const int hn = h + g.pt - ky;
      if (hn < 0 || (hn & 1)) continue;
      const int ho = hn >> 1;
      if (ho >= g.Ho) continue;
      for (int kx = 0; kx < 3; ++kx) {
        const int wn = w + g.pl - kx;
        if (wn < 0 || (wn & 1)) continue;
        const int wo = wn >> 1;
        if (wo >= g.Wo) continue;
        const int64_t o = (((int64_t)b * g.Ho + ho) * g.Wo + wo) * g.C + c;
        if (arg[o] == ky * 3 + kx) acc += __bfloat162float(dy[o]);
      }
    }
    dx[i] = __float2bfloat16(acc);
  }
}

static int cnn_blocks(int64_t work) {
  int64_t b = (work + 255) / 256;
  const int64_t cap = 148 * 16;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}
static bool al4(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 3) == 0; }

}  // namespace sav

using namespace sav;

extern "C" int sav_act_fwd(const void* x, void* y, int64_t n, int32_t kind, void* stream) {
  if (n <= 0) return SAV_OK;
  if (n % 2 || !al4(x) || !al4(y)) return set_error(SAV_ERR_ALIGN, "sav_act_fwd: n %% 2, 4 B alignment");
  act_fwd_kernel<<<cnn_blocks(n / 2), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(x), reinterpret_cast<bf16*>(y), n / 2, kind);
  count_launch();
  return check_launch("sav_act_fwd");
}

extern "C" int sav_act_bwd(const void* x, const void* dy, void* dx, int64_t n, int32_t kind, void* stream) {
  if (n <= 0) return SAV_OK;
  if (n % 2 || !al4(x) || !al4(dy) || !al4(dx)) return set_error(SAV_ERR_ALIGN, "sav_act_bwd: n %% 2, 4 B alignment");
  act_bwd_kernel<<<cnn_blocks(n / 2), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(x), reinterpret_cast<const bf16*>(dy), reinterpret_cast<bf16*>(dx), n / 2, kind);
  count_launch();
  return check_launch("sav_act_bwd");
}

extern "C" int sav_bn_act_fwd(const void* y, const float* mean, const float* rstd, const float* gamma,
                              const float* beta, void* z, int64_t rows, int64_t C, int32_t kind, void* stream) {
  if (rows <= 0 || C <= 0 || C % 2) return set_error(SAV_ERR_SHAPE, "sav_bn_act_fwd: C %% 2");
  bn_act_fwd_kernel<<<cnn_blocks(rows * (C / 2)), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(y), mean, rstd, gamma, beta, reinterpret_cast<bf16*>(z), rows, (int)(C / 2), kind);
  count_launch();
  return check_launch("sav_bn_act_fwd");
}

extern "C" int sav_bn_act_bwd(const void* dz, const void* y, const float* mean, const float* rstd,
                              const float* gamma, const float* beta, float* dgamma, float* dbeta,
                              float* scratch, void* dy, int64_t rows, int64_t C, int32_t kind, void* stream) {
  if (rows <= 0 || C <= 0 || C % 2) return set_error(SAV_ERR_SHAPE, "sav_bn_act_bwd: C %% 2");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int rc = sav_fill_f32(scratch, 0.f, 2 * C, stream);
  if (rc) return rc;
  const int gx = (int)((C / 2 + 63) / 64);
  int64_t gy = (rows + 63) / 64;
  const int64_t cap = (148 * 8 + gx - 1) / gx;
  if (gy > cap) gy = cap;
  if (gy < 1) gy = 1;
  bn_act_bwd_reduce_kernel<<<dim3(gx, (unsigned)gy), 256, 0, st>>>(
      reinterpret_cast<const bf16*>(dz), reinterpret_cast<const bf16*>(y), mean, rstd, gamma, beta, rows,
      (int)C, kind, scratch, scratch + C);
  count_launch();
  bn_act_bwd_apply_kernel<<<cnn_blocks(rows * (C / 2)), 256, 0, st>>>(
      reinterpret_cast<const bf16*>(dz), reinterpret_cast<const bf16*>(y), mean, rstd, gamma, beta, scratch,
      scratch + C, 1.f / (float)rows, reinterpret_cast<bf16*>(dy), rows, (int)(C / 2), kind);
  count_launch();
  // dbeta += s1, dgamma += s2 (parameter gradients accumulate: the arena is zeroed once per step)
  axpy2b_kernel<<<(int)((C + 255) / 256), 256, 0, st>>>(scratch, dbeta, scratch + C, dgamma, (int)C);
  count_launch();
  return check_launch("sav_bn_act_bwd");
}

extern "C" int sav_gate_res_act_fwd(const void* x, const void* gate, const void* residual, void* out,
                                    int64_t B, int64_t rows_per_sample, int64_t C, int32_t kind, void* stream) {
  if (B <= 0 || rows_per_sample <= 0 || C <= 0 || C % 2) return set_error(SAV_ERR_SHAPE, "sav_gate_res_act_fwd: C %% 2");
  gate_res_act_fwd_kernel<<<cnn_blocks(B * rows_per_sample * (C / 2)), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(x), reinterpret_cast<const bf16*>(gate), reinterpret_cast<const bf16*>(residual),
      reinterpret_cast<bf16*>(out), B * rows_per_sample, rows_per_sample, (int)(C / 2), kind);
  count_launch();
  return check_launch("sav_gate_res_act_fwd");
}

extern "C" int sav_gate_res_act_bwd(const void* dout, const void* x, const void* gate, const void* residual,
                                    void* dx, void* dresidual, float* dgate, int64_t B, int64_t rows_per_sample,
                                    int64_t C, int32_t kind, void* stream) {
  if (B <= 0 || rows_per_sample <= 0 || C <= 0 || C % 2) return set_error(SAV_ERR_SHAPE, "sav_gate_res_act_bwd: C %% 2");
  if ((gate == nullptr) != (dgate == nullptr)) return set_error(SAV_ERR_ARG, "sav_gate_res_act_bwd: gate and dgate go together");
  if (B > 65535) return set_error(SAV_ERR_SHAPE, "sav_gate_res_act_bwd: batch > 65535");
  int rc;
  if (dgate && (rc = sav_fill_f32(dgate, 0.f, B * C, stream))) return rc;
  const int gx = (int)((C / 2 + 63) / 64);
  int64_t gz = (rows_per_sample + 63) / 64;
  const int64_t cap = (148 * 8 + gx * B - 1) / (gx * B);
  if (gz > cap) gz = cap;
  if (gz < 1) gz = 1;
  if (gz > 65535) gz = 65535;
  gate_res_act_bwd_kernel<<<dim3(gx, (unsigned)B, (unsigned)gz), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(dout), reinterpret_cast<const bf16*>(x), reinterpret_cast<const bf16*>(gate),
      reinterpret_cast<const bf16*>(residual), reinterpret_cast<bf16*>(dx), reinterpret_cast<bf16*>(dresidual), dgate,
      rows_per_sample, (int)C, kind);
  count_launch();
  return check_launch("sav_gate_res_act_bwd");
}

extern "C" int sav_spatial_mean_fwd(const void* x, void* out, int64_t B, int64_t rows_per_sample, int64_t C, void* stream) {
  if (B <= 0 || rows_per_sample <= 0 || C <= 0 || C % 2 || B > 65535) return set_error(SAV_ERR_SHAPE, "sav_spatial_mean_fwd");
  spatial_mean_fwd_kernel<<<dim3((unsigned)((C / 2 + 63) / 64), (unsigned)B), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(x), reinterpret_cast<bf16*>(out), rows_per_sample, (int)C);
  count_launch();
  return check_launch("sav_spatial_mean_fwd");
}

extern "C" int sav_spatial_mean_bwd(const void* dout, void* dx, int64_t B, int64_t rows_per_sample, int64_t C, void* stream) {
  if (B <= 0 || rows_per_sample <= 0 || C <= 0 || C % 2) return set_error(SAV_ERR_SHAPE, "sav_spatial_mean_bwd");
  spatial_mean_bwd_kernel<<<cnn_blocks(B * rows_per_sample * (C / 2)), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(dout), reinterpret_cast<bf16*>(dx), B * rows_per_sample, rows_per_sample, (int)(C / 2));
  count_launch();
  return check_launch("sav_spatial_mean_bwd");
}

static int pool_geom(PoolGeom& g, int64_t B, int64_t H, int64_t W, int64_t C, bool valid = false) {
  if (B <= 0 || H <= 0 || W <= 0 || C <= 0) return set_error(SAV_ERR_SHAPE, "maxpool: bad shape");
  g.B = (int)B; g.H = (int)H; g.W = (int)W; g.C = (int)C;
  if (valid) {   // padding 'VALID' (flax nn.max_pool default, image_to_token.py:33-37): no padding
    if (H < 3 || W < 3) return set_error(SAV_ERR_SHAPE, "maxpool VALID: H, W >= 3");
    g.Ho = (int)((H - 3) / 2 + 1); g.Wo = (int)((W - 3) / 2 + 1); g.pt = g.pl = 0;
    return SAV_OK;
  }
  int64_t o; int32_t lo;
  sav_conv_same_geom(H, 3, 2, &o, &lo); g.Ho = (int)o; g.pt = lo;
  sav_conv_same_geom(W, 3, 2, &o, &lo); g.Wo = (int)o; g.pl = lo;
  return SAV_OK;
}

extern "C" int sav_maxpool3x3s2_fwd(const void* x, void* y, void* argmax, int64_t B, int64_t H, int64_t W,
                                    int64_t C, void* stream) {
  PoolGeom g;
  int rc = pool_geom(g, B, H, W, C);
  if (rc) return rc;
  maxpool_fwd_kernel<<<cnn_blocks((int64_t)g.B * g.Ho * g.Wo * g.C), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(x), reinterpret_cast<bf16*>(y), reinterpret_cast<uint8_t*>(argmax), g);
  count_launch();
  return check_launch("sav_maxpool3x3s2_fwd");
}

extern "C" int sav_maxpool3x3s2_bwd(const void* dy, const void* argmax, void* dx, int64_t B, int64_t H,
                                    int64_t W, int64_t C, void* stream) {
  PoolGeom g;
  int rc = pool_geom(g, B, H, W, C);
  if (rc) return rc;
  maxpool_bwd_kernel<<<cnn_blocks((int64_t)g.B * g.H * g.W * g.C), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(dy), reinterpret_cast<const uint8_t*>(argmax), reinterpret_cast<bf16*>(dx), g);
  count_launch();
  return check_launch("sav_maxpool3x3s2_bwd");
}

// nn.max_pool(x, (3,3), strides=(2,2)) with flax's default 'VALID' padding (CeiT's Image2TokenBlock,
// image_to_token.py:33-37): the same kernels with no padding and Ho = (H - 3) / 2 + 1.
extern "C" int sav_maxpool3x3s2_valid_fwd(const void* x, void* y, void* argmax, int64_t B, int64_t H,
                                          int64_t W, int64_t C, void* stream) {
  PoolGeom g;
  int rc = pool_geom(g, B, H, W, C, true);
  if (rc) return rc;
  maxpool_fwd_kernel<<<cnn_blocks((int64_t)g.B * g.Ho * g.Wo * g.C), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(x), reinterpret_cast<bf16*>(y), reinterpret_cast<uint8_t*>(argmax), g);
  count_launch();
  return check_launch("sav_maxpool3x3s2_valid_fwd");
}

extern "C" int sav_maxpool3x3s2_valid_bwd(const void* dy, const void* argmax, void* dx, int64_t B, int64_t H,
                                          int64_t W, int64_t C, void* stream) {
  PoolGeom g;
  int rc = pool_geom(g, B, H, W, C, true);
  if (rc) return rc;
  maxpool_bwd_kernel<<<cnn_blocks((int64_t)g.B * g.H * g.W * g.C), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(dy), reinterpret_cast<const uint8_t*>(argmax), reinterpret_cast<bf16*>(dx), g);
  count_launch();
  return check_launch("sav_maxpool3x3s2_valid_bwd");
}

// 'b (h ph) (w pw) c -> b (h w) (ph pw c)' of a bf16 ACTIVATION [B,H,W,C] (image_to_token.py:38-43:
// the pooled feature map cut into 4x4 patches) and its inverse (the gradient's way back): a pure
// permutation, two bf16 elements per thread (C even, so a pair never straddles a pixel).
namespace sav {
__global__ void patchify_act_kernel(const bf16* __restrict__ src, bf16* __restrict__ dst, int64_t total2,
                                    int H, int W, int C2, int ph, int pw, int inverse) {
  const int gw = W / pw;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total2; i += (int64_t)gridDim.x * blockDim.x) {
    // i indexes the [B,H,W,C/2] image layout
    const int c2 = (int)(i % C2);
    int64_t t = i / C2;
    const int x = (int)(t % W); t /= W;
    const int y = (int)(t % H);
    const int64_t b = t / H;
    const int hy = y / ph, py = y - hy * ph, wx = x / pw, px = x - wx * pw;
    const int64_t tok = (b * (H / ph) + hy) * gw + wx;
    const int64_t j = (tok * ph * pw + py * pw + px) * C2 + c2;     // [B, tokens, ph, pw, C/2]
    if (inverse) reinterpret_cast<uint32_t*>(dst)[i] = reinterpret_cast<const uint32_t*>(src)[j];
    else reinterpret_cast<uint32_t*>(dst)[j] = reinterpret_cast<const uint32_t*>(src)[i];
  }
}
}  // namespace sav

extern "C" int sav_patchify_act(const void* src, void* dst, int64_t B, int64_t H, int64_t W, int64_t C,
                                int64_t ph, int64_t pw, int32_t inverse, void* stream) {
  if (B <= 0 || !src || !dst || ph <= 0 || pw <= 0 || H % ph || W % pw || C % 2)
    return set_error(SAV_ERR_SHAPE, "sav_patchify_act: H, W multiples of the patch shape, C even");
  const int64_t total2 = B * H * W * (C / 2);
  patchify_act_kernel<<<cnn_blocks(total2), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(src), reinterpret_cast<bf16*>(dst), total2, (int)H, (int)W, (int)(C / 2),
      (int)ph, (int)pw, inverse);
  count_launch();
  return check_launch("sav_patchify_act");
}
