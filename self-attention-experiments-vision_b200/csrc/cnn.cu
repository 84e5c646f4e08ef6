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
// TXP = channel pairs per block (64, or 32 for C <= 64 so that no lane idles on the widest tensors)
template <int TXP>
__global__ void __launch_bounds__(256)
bn_act_bwd_reduce_kernel(const bf16* __restrict__ dz, const bf16* __restrict__ y,
                         const float* __restrict__ mean, const float* __restrict__ rstd,
                         const float* __restrict__ gamma, const float* __restrict__ beta, int64_t rows,
                         int C, int kind, float* __restrict__ s1, float* __restrict__ s2) {
  constexpr int TYP = 256 / TXP;
  __shared__ float red[TYP][TXP][4];
  const int tx = threadIdx.x % TXP, ty = threadIdx.x / TXP;
  const int c = (blockIdx.x * TXP + tx) * 2;
  float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
  if (c < C) {
    const float m0 = mean[c], m1 = mean[c + 1], r0 = rstd[c], r1 = rstd[c + 1];
    const float g0 = gamma[c], g1 = gamma[c + 1], e0 = beta[c], e1 = beta[c + 1];
    for (int64_t r = (int64_t)blockIdx.y * TYP + ty; r < rows; r += (int64_t)gridDim.y * TYP) {
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
    for (int k = 1; k < TYP; ++k) {
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

// ---- eight channels (16 bytes) per thread for C % 8 == 0 and 16-byte aligned tensors.  The 4-byte
// kernels above leave these memory-bound passes at a third of the HBM rate on B200
// (profiles/r02az_timeline_botnet_t3.txt: BatchNorm + activation backward 393 us on the 103 MB stem
// tensor against an 85 us floor); they stay as the fallback for odd channel counts.
__device__ __forceinline__ void ld8(const bf16* p, float (&f)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 v = unpack_bf16(w[e]);
    f[2 * e] = v.x;
    f[2 * e + 1] = v.y;
  }
}
__device__ __forceinline__ void st8(bf16* p, const float (&f)[8]) {
  *reinterpret_cast<uint4*>(p) = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]),
                                            pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}
__device__ __forceinline__ void ldp8(const float* p, float (&f)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

__global__ void bn_act_fwd_vec8_kernel(const bf16* __restrict__ y, const float* __restrict__ mean,
                                       const float* __restrict__ rstd, const float* __restrict__ gamma,
                                       const float* __restrict__ beta, bf16* __restrict__ z, int64_t rows,
                                       int C8, int kind) {
  // the launcher makes gridDim.x * blockDim.x a multiple of C8: a thread keeps its channel octet, so
  // the per-channel parameters are loaded once, folded to z = act(y * sc + sh)
  const int64_t total = rows * C8;
  const int64_t i0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int c = (int)(i0 % C8) * 8;
  float m[8], r[8], g[8], b[8];
  ldp8(mean + c, m); ldp8(rstd + c, r); ldp8(gamma + c, g); ldp8(beta + c, b);
#pragma unroll
  for (int e = 0; e < 8; ++e) { r[e] *= g[e]; b[e] -= m[e] * r[e]; }
  for (int64_t i = i0; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    float v[8];
    ld8(y + i * 8, v);
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = act_apply(fmaf(v[e], r[e], b[e]), kind);
    st8(z + i * 8, v);
  }
}

__global__ void bn_act_bwd_apply_vec8_kernel(const bf16* __restrict__ dz, const bf16* __restrict__ y,
                                             const float* __restrict__ mean, const float* __restrict__ rstd,
                                             const float* __restrict__ gamma, const float* __restrict__ beta,
                                             const float* __restrict__ s1, const float* __restrict__ s2,
                                             float inv_n, bf16* __restrict__ dy, int64_t rows, int C8, int kind) {
  const int64_t total = rows * C8;
  const int64_t i0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int c = (int)(i0 % C8) * 8;      // fixed per thread (thread count is a multiple of C8)
  float m[8], r[8], g[8], b[8], a1[8], a2[8];
  ldp8(mean + c, m); ldp8(rstd + c, r); ldp8(gamma + c, g); ldp8(beta + c, b);
  ldp8(s1 + c, a1); ldp8(s2 + c, a2);
  for (int64_t i = i0; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    float v[8], d[8];
    ld8(y + i * 8, v);
    ld8(dz + i * 8, d);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float h = (v[e] - m[e]) * r[e];
      const float dd = d[e] * act_grad(h * g[e] + b[e], kind);
      v[e] = g[e] * r[e] * (dd - a1[e] * inv_n - h * a2[e] * inv_n);
    }
    st8(dy + i * 8, v);
  }
}

// per-channel reductions over the rows, eight channels per thread: block = TX channel octets x
// (256 / TX) row lanes; partials meet by shuffle inside the warp, across warps in shared memory, one
// atomicAdd per channel and block.  kMode 0: (sum y, sum y^2) = batch statistics;
// kMode 1: (sum dbn, sum dbn yhat) with dbn = dz act'(bn(y)).
template <int kMode>
__global__ void __launch_bounds__(256)
bn_reduce_vec8_kernel(const bf16* __restrict__ y, const bf16* __restrict__ dz, const float* __restrict__ mean,
                      const float* __restrict__ rstd, const float* __restrict__ gamma,
                      const float* __restrict__ beta, int64_t rows, int C8, int TX, int kind,
                      float* __restrict__ o1, float* __restrict__ o2) {
  __shared__ float red[8][8][16];
  const int tx = threadIdx.x % TX, ty = threadIdx.x / TX, TY = 256 / TX;
  const int oct = blockIdx.x * TX + tx;
  float a[8], b[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) { a[e] = 0.f; b[e] = 0.f; }
  if (oct < C8) {
    const int c = oct * 8;
    float m[8], r[8], g[8], bt[8];
    if (kMode == 1) { ldp8(mean + c, m); ldp8(rstd + c, r); ldp8(gamma + c, g); ldp8(beta + c, bt); }
    for (int64_t row = (int64_t)blockIdx.y * TY + ty; row < rows; row += (int64_t)gridDim.y * TY) {
      float v[8];
      ld8(y + (row * C8 + oct) * 8, v);
      if (kMode == 0) {
#pragma unroll
        for (int e = 0; e < 8; ++e) { a[e] += v[e]; b[e] += v[e] * v[e]; }
      } else {
        float d[8];
        ld8(dz + (row * C8 + oct) * 8, d);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float h = (v[e] - m[e]) * r[e];
          const float dd = d[e] * act_grad(h * g[e] + bt[e], kind);
          a[e] += dd;
          b[e] += dd * h;
        }
      }
    }
  }
  // lanes of equal tx inside the warp (TX divides 32)
  for (int off = TX; off < 32; off <<= 1) {
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      a[e] += __shfl_xor_sync(0xffffffffu, a[e], off);
      b[e] += __shfl_xor_sync(0xffffffffu, b[e], off);
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane < TX) {
#pragma unroll
    for (int e = 0; e < 8; ++e) { red[warp][lane][e] = a[e]; red[warp][lane][8 + e] = b[e]; }
  }
  __syncthreads();
  if (threadIdx.x < TX * 16) {
    const int x = threadIdx.x / 16, q = threadIdx.x % 16;
    const int oc = blockIdx.x * TX + x;
    if (oc < C8) {
      float sum = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) sum += red[w][x][q];
      atomicAdd((q < 8 ? o1 : o2) + oc * 8 + (q & 7), sum);
    }
  }
}

__global__ void gate_res_act_fwd_vec8_kernel(const bf16* __restrict__ x, const bf16* __restrict__ gate,
                                             const bf16* __restrict__ res, bf16* __restrict__ out,
                                             int64_t rows, int64_t rows_per_sample, int C8, int kind) {
  const int64_t total = rows * C8;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / C8;
    const int c8 = (int)(i - r * C8);
    float v[8];
    ld8(x + i * 8, v);
    if (gate) {
      float s[8];
      ld8(gate + ((r / rows_per_sample) * C8 + c8) * 8, s);
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] *= s[e];
    }
    if (res) {
      float q[8];
      ld8(res + i * 8, q);
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] += q[e];
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = act_apply(v[e], kind);
    st8(out + i * 8, v);
  }
}

static bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static void launch_bn_act_bwd_reduce(const void* dz, const void* y, const float* mean, const float* rstd,
                                     const float* gamma, const float* beta, int64_t rows, int64_t C, int kind,
                                     float* scratch, cudaStream_t st) {
  const int txp = C <= 64 ? 32 : 64;
  const int gx = (int)((C / 2 + txp - 1) / txp);
  int64_t gy = (rows + 63) / 64;
  const int64_t cap = (148 * 8 + gx - 1) / gx;
  if (gy > cap) gy = cap;
  if (gy < 1) gy = 1;
  if (txp == 32)
    bn_act_bwd_reduce_kernel<32><<<dim3(gx, (unsigned)gy), 256, 0, st>>>(
        reinterpret_cast<const bf16*>(dz), reinterpret_cast<const bf16*>(y), mean, rstd, gamma, beta, rows,
        (int)C, kind, scratch, scratch + C);
  else
    bn_act_bwd_reduce_kernel<64><<<dim3(gx, (unsigned)gy), 256, 0, st>>>(
        reinterpret_cast<const bf16*>(dz), reinterpret_cast<const bf16*>(y), mean, rstd, gamma, beta, rows,
        (int)C, kind, scratch, scratch + C);
}
// blocks of 256 threads for `work` octets such that the thread count is a multiple of C8
static int vec8_blocks(int64_t work, int C8) {
  int64_t b = (work + 255) / 256;
  const int64_t cap = 148 * 16;
  if (b > cap) b = cap;
  int gcd = C8, x = 256;
  while (x) { const int t = gcd % x; gcd = x; x = t; }
  const int mult = C8 / gcd;
  b = (b + mult - 1) / mult * mult;
  return (int)(b < mult ? mult : b);
}
// launch geometry of bn_reduce_vec8_kernel
static void reduce8_geom(int64_t rows, int C8, int& TX, dim3& grid) {
  TX = C8 >= 8 ? 8 : (C8 >= 4 ? 4 : (C8 >= 2 ? 2 : 1));
  const int gx = (C8 + TX - 1) / TX;
  const int TY = 256 / TX;
  int64_t gy = (rows + 4 * TY - 1) / (4 * TY);      // >= 4 rows per thread
  const int64_t cap = (148 * 8 + gx - 1) / gx;
  if (gy > cap) gy = cap;
  if (gy < 1) gy = 1;
  grid = dim3((unsigned)gx, (unsigned)gy);
}

// batch statistics entry for conv2d.cu's sav_bn_stats (sum / sumsq zeroed by the caller)
bool bn_stats_vec8(const void* y, int64_t rows, int64_t C, float* sum, float* sumsq, cudaStream_t st) {
  if (C % 8 || !al16(y)) return false;
  int TX;
  dim3 grid;
  reduce8_geom(rows, (int)(C / 8), TX, grid);
  bn_reduce_vec8_kernel<0><<<grid, 256, 0, st>>>(reinterpret_cast<const bf16*>(y), nullptr, nullptr, nullptr,
                                                 nullptr, nullptr, rows, (int)(C / 8), TX, 0, sum, sumsq);
  return true;
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
  if (C % 8 == 0 && al16(y) && al16(z) && al16(mean) && al16(rstd) && al16(gamma) && al16(beta)) {
    bn_act_fwd_vec8_kernel<<<vec8_blocks(rows * (C / 8), (int)(C / 8)), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const bf16*>(y), mean, rstd, gamma, beta, reinterpret_cast<bf16*>(z), rows, (int)(C / 8), kind);
    count_launch();
    return check_launch("sav_bn_act_fwd");
  }
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
  if (C % 8 == 0 && al16(dz) && al16(y) && al16(dy) && al16(mean) && al16(rstd) && al16(gamma) && al16(beta) &&
      al16(scratch)) {
    // (the per-channel sums stay on the 4-byte kernel: the eight-channel reduction measured slower
    // here -- 16 accumulators and 32 parameter registers per thread cost more occupancy than the wider
    // loads return, profiles/r02bb_timeline_botnet_t3.txt)
    launch_bn_act_bwd_reduce(dz, y, mean, rstd, gamma, beta, rows, C, kind, scratch, st);
    count_launch();
    bn_act_bwd_apply_vec8_kernel<<<vec8_blocks(rows * (C / 8), (int)(C / 8)), 256, 0, st>>>(
        reinterpret_cast<const bf16*>(dz), reinterpret_cast<const bf16*>(y), mean, rstd, gamma, beta, scratch,
        scratch + C, 1.f / (float)rows, reinterpret_cast<bf16*>(dy), rows, (int)(C / 8), kind);
    count_launch();
  } else {
    launch_bn_act_bwd_reduce(dz, y, mean, rstd, gamma, beta, rows, C, kind, scratch, st);
    count_launch();
    bn_act_bwd_apply_kernel<<<cnn_blocks(rows * (C / 2)), 256, 0, st>>>(
        reinterpret_cast<const bf16*>(dz), reinterpret_cast<const bf16*>(y), mean, rstd, gamma, beta, scratch,
        scratch + C, 1.f / (float)rows, reinterpret_cast<bf16*>(dy), rows, (int)(C / 2), kind);
    count_launch();
  }
  // dbeta += s1, dgamma += s2 (parameter gradients accumulate: the arena is zeroed once per step)
  axpy2b_kernel<<<(int)((C + 255) / 256), 256, 0, st>>>(scratch, dbeta, scratch + C, dgamma, (int)C);
  count_launch();
  return check_launch("sav_bn_act_bwd");
}

extern "C" int sav_gate_res_act_fwd(const void* x, const void* gate, const void* residual, void* out,
                                    int64_t B, int64_t rows_per_sample, int64_t C, int32_t kind, void* stream) {
  if (B <= 0 || rows_per_sample <= 0 || C <= 0 || C % 2) return set_error(SAV_ERR_SHAPE, "sav_gate_res_act_fwd: C %% 2");
  if (C % 8 == 0 && al16(x) && al16(out) && (!gate || al16(gate)) && (!residual || al16(residual))) {
    gate_res_act_fwd_vec8_kernel<<<cnn_blocks(B * rows_per_sample * (C / 8)), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const bf16*>(x), reinterpret_cast<const bf16*>(gate), reinterpret_cast<const bf16*>(residual),
        reinterpret_cast<bf16*>(out), B * rows_per_sample, rows_per_sample, (int)(C / 8), kind);
    count_launch();
    return check_launch("sav_gate_res_act_fwd");
  }
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
