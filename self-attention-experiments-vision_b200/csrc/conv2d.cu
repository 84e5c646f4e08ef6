// Dense (non-depthwise) convolutions of the model stems and conv stages as im2col + the tcgen05
// GEMM of gemm.cu (nn.Conv at cvt.py:27-31 -- the CvT token embedding, 7x7 stride 4 / 3x3 stride 2,
// SAME, with bias -- and botnet.py:291-305,47-64 -- ResNet stem / bottleneck 1x1, 3x3 convs):
//   im2col   x [B,H,W,Cin] (fp32 or bf16, arbitrary element strides: also the reference loader's
//            [H,W,C,N] image layout, train.py:80) -> bf16 patch rows [B*Ho*Wo, ld] with the
//            feature order (kh, kw, cin) of a Flax kernel [kh,kw,cin,cout] flattened; zero padding;
//            the pad columns K..ld (row pitch kept a multiple of 8 for TMA) are zero
//   col2im   reverse mode: dcols bf16 [B*Ho*Wo, ld] -> dx fp32 or bf16 [B,H,W,Cin] in GATHER form
//            (each input element sums the <= ceil(k/s)^2 patch entries that read it: no atomics)
// plus the standalone BatchNorm pieces the conv stages need beyond conv.cu: per-channel batch
// statistics of a bf16 [rows, C] tensor, the running-average update (momentum) and the eval-mode
// apply from running averages (cvt_attention.py:33-36, botnet.py:286-290).
#include "common.cuh"
#include "host.h"

namespace sav {

// cnn.cu: eight-channels-per-thread batch statistics (false: shape / alignment not eligible)
bool bn_stats_vec8(const void* y, int64_t rows, int64_t C, float* sum, float* sumsq, cudaStream_t st);

struct Im2colGeom {
  int B, H, W, C, Ho, Wo, kh, kw, sh, sw, pt, pl;
  int64_t xs_b, xs_h, xs_w, xs_c;   // element strides of x
  int64_t ld;                        // row pitch of cols (elements)
};

template <typename TX>
__global__ void im2col_kernel(const TX* __restrict__ x, bf16* __restrict__ cols, Im2colGeom g) {
  const int K = g.kh * g.kw * g.C;
  const int64_t total = (int64_t)g.B * g.Ho * g.Wo * g.ld;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int k = (int)(idx % g.ld);
    const int64_t row = idx / g.ld;
    float v = 0.f;
    if (k < K) {
      const int c = k % g.C;
      const int kk = k / g.C;
      const int kx = kk % g.kw, ky = kk / g.kw;
      const int wo = (int)(row % g.Wo);
      const int64_t t = row / g.Wo;
      const int ho = (int)(t % g.Ho);
      const int b = (int)(t / g.Ho);
      const int h = ho * g.sh - g.pt + ky, w = wo * g.sw - g.pl + kx;
      if (h >= 0 && h < g.H && w >= 0 && w < g.W)
        v = (float)x[b * g.xs_b + h * g.xs_h + w * g.xs_w + c * g.xs_c];
    }
    cols[idx] = __float2bfloat16(v);
  }
}

// bf16 activations with channels innermost, C % 8 == 0: eight channels (16 bytes) per thread.  The
// index arithmetic (six divisions) is paid once per 16-byte access instead of once per element --
// the scalar kernel above ran the 173 M-element patch matrix of CeiT's 3x3 x 1536-channel LeFF
// convolution at 0.3 TB/s (1.2 ms against a 0.09 ms write floor, profiles/r02ay_timeline_ceit_s.txt).
__global__ void im2col_vec8_kernel(const bf16* __restrict__ x, bf16* __restrict__ cols, Im2colGeom g) {
  const int K8 = g.kh * g.kw * g.C / 8;
  const int ld8 = (int)(g.ld / 8);
  const int64_t total = (int64_t)g.B * g.Ho * g.Wo * ld8;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int k8 = (int)(idx % ld8);
    const int64_t row = idx / ld8;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (k8 < K8) {
      const int k = k8 * 8;
      const int c = k % g.C;
      const int kk = k / g.C;
      const int kx = kk % g.kw, ky = kk / g.kw;
      const int wo = (int)(row % g.Wo);
      const int64_t t = row / g.Wo;
      const int ho = (int)(t % g.Ho);
      const int b = (int)(t / g.Ho);
      const int h = ho * g.sh - g.pt + ky, w = wo * g.sw - g.pl + kx;
      if (h >= 0 && h < g.H && w >= 0 && w < g.W)
        v = __ldg(reinterpret_cast<const uint4*>(x + b * g.xs_b + h * g.xs_h + w * g.xs_w + c));
    }
    *reinterpret_cast<uint4*>(cols + row * g.ld + (int64_t)k8 * 8) = v;
  }
}

// gather form of the reverse mode, eight channels per thread (C % 8 == 0)
template <typename TD>
__global__ void col2im_vec8_kernel(const bf16* __restrict__ cols, TD* __restrict__ dx, Im2colGeom g,
                                   int accumulate) {
  const int C8 = g.C / 8;
  const int64_t total = (int64_t)g.B * g.H * g.W * C8;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % C8) * 8;
    int64_t t = idx / C8;
    const int w = (int)(t % g.W);
    t /= g.W;
    const int h = (int)(t % g.H);
    const int b = (int)(t / g.H);
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
    for (int ky = 0; ky < g.kh; ++ky) {
      const int hn = h + g.pt - ky;
      if (hn < 0 || hn % g.sh) continue;
      const int ho = hn / g.sh;
      if (ho >= g.Ho) continue;
      for (int kx = 0; kx < g.kw; ++kx) {
        const int wn = w + g.pl - kx;
        if (wn < 0 || wn % g.sw) continue;
        const int wo = wn / g.sw;
        if (wo >= g.Wo) continue;
        const int64_t row = ((int64_t)b * g.Ho + ho) * g.Wo + wo;
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(cols + row * g.ld + (ky * g.kw + kx) * g.C + c));
        const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = unpack_bf16(u[e]);
          acc[2 * e] += f.x;
          acc[2 * e + 1] += f.y;
        }
      }
    }
    TD* d = dx + idx * 8;
    if (accumulate) {
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += (float)d[e];
    }
    if (sizeof(TD) == 2) {
      *reinterpret_cast<uint4*>(d) = make_uint4(pack_bf16(acc[0], acc[1]), pack_bf16(acc[2], acc[3]),
                                                pack_bf16(acc[4], acc[5]), pack_bf16(acc[6], acc[7]));
    } else {
      reinterpret_cast<float4*>(d)[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
      reinterpret_cast<float4*>(d)[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
    }
  }
}

template <typename TD>
__global__ void col2im_kernel(const bf16* __restrict__ cols, TD* __restrict__ dx, Im2colGeom g,
                              int accumulate) {
  const int64_t total = (int64_t)g.B * g.H * g.W * g.C;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % g.C);
    int64_t t = idx / g.C;
    const int w = (int)(t % g.W);
    t /= g.W;
    const int h = (int)(t % g.H);
    const int b = (int)(t / g.H);
    float acc = 0.f;
    for (int ky = 0; ky < g.kh; ++ky) {
      const int hn = h + g.pt - ky;
      if (hn < 0 || hn % g.sh) continue;
      const int ho = hn / g.sh;
      if (ho >= g.Ho) continue;
      for (int kx = 0; kx < g.kw; ++kx) {
        const int wn = w + g.pl - kx;
        if (wn < 0 || wn % g.sw) continue;
        const int wo = wn / g.sw;
        if (wo >= g.Wo) continue;
        const int64_t row = ((int64_t)b * g.Ho + ho) * g.Wo + wo;
        acc += __bfloat162float(cols[row * g.ld + (ky * g.kw + kx) * g.C + c]);
      }
    }
    if (accumulate) acc += (float)dx[idx];
    dx[idx] = (TD)acc;
  }
}

// per-channel sum / sum of squares of a bf16 [rows, C] tensor: block = 64 channel pairs x 4 row
// lanes, grid-stride over rows, shared-memory reduce, one red.add per channel per block
__global__ void __launch_bounds__(256)
bn_stats_kernel(const bf16* __restrict__ y, int64_t rows, int C, float* __restrict__ sum,
                float* __restrict__ sumsq) {
  __shared__ float red[4][64][4];
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  const int c = (blockIdx.x * 64 + tx) * 2;
  float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
  if (c < C) {
    for (int64_t r = (int64_t)blockIdx.y * 4 + ty; r < rows; r += (int64_t)gridDim.y * 4) {
      const float2 v = unpack_bf16(*reinterpret_cast<const uint32_t*>(y + r * C + c));
      s0 += v.x; s1 += v.y; q0 += v.x * v.x; q1 += v.y * v.y;
    }
  }
  red[ty][tx][0] = s0; red[ty][tx][1] = s1; red[ty][tx][2] = q0; red[ty][tx][3] = q1;
  __syncthreads();
  if (ty == 0 && c < C) {
    for (int k = 1; k < 4; ++k) {
      s0 += red[k][tx][0]; s1 += red[k][tx][1]; q0 += red[k][tx][2]; q1 += red[k][tx][3];
    }
    atomicAdd(sum + c, s0); atomicAdd(sum + c + 1, s1);
    atomicAdd(sumsq + c, q0); atomicAdd(sumsq + c + 1, q1);
  }
}

// flax nn.BatchNorm running averages: ra = momentum * ra + (1 - momentum) * batch statistic
__global__ void bn_running_kernel(const float* __restrict__ mean, const float* __restrict__ var,
                                  float* __restrict__ ra_mean, float* __restrict__ ra_var,
                                  float momentum, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) {
    ra_mean[c] = momentum * ra_mean[c] + (1.f - momentum) * mean[c];
    ra_var[c] = momentum * ra_var[c] + (1.f - momentum) * var[c];
  }
}

// eval mode: z = (y - ra_mean) rsqrt(ra_var + eps) gamma + beta
__global__ void bn_eval_kernel(const bf16* __restrict__ y, const float* __restrict__ ra_mean,
                               const float* __restrict__ ra_var, const float* __restrict__ gamma,
                               const float* __restrict__ beta, float eps, bf16* __restrict__ z,
                               int64_t rows, int C2) {
  const int64_t total = rows * C2;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C2) * 2;
    const float2 v = unpack_bf16(reinterpret_cast<const uint32_t*>(y)[i]);
    const float r0 = rsqrtf(ra_var[c] + eps), r1 = rsqrtf(ra_var[c + 1] + eps);
    reinterpret_cast<uint32_t*>(z)[i] = pack_bf16((v.x - ra_mean[c]) * r0 * gamma[c] + beta[c],
                                                   (v.y - ra_mean[c + 1]) * r1 * gamma[c + 1] + beta[c + 1]);
  }
}

static int c2_blocks(int64_t work) {
  int64_t b = (work + 255) / 256;
  const int64_t cap = 148 * 16;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

static int fill_geom(Im2colGeom& g, const sav_conv2d_geom* a, const char* who) {
  if (a->B <= 0 || a->H <= 0 || a->W <= 0 || a->C <= 0 || a->kh <= 0 || a->kw <= 0 || a->sh <= 0 ||
      a->sw <= 0 || a->Ho <= 0 || a->Wo <= 0)
    return set_error(SAV_ERR_SHAPE, "%s: bad geometry", who);
  if (a->ld < (int64_t)a->kh * a->kw * a->C || a->ld % 8)
    return set_error(SAV_ERR_SHAPE, "%s: ld must be >= kh*kw*C and a multiple of 8", who);
  g.B = (int)a->B; g.H = (int)a->H; g.W = (int)a->W; g.C = (int)a->C;
  g.Ho = (int)a->Ho; g.Wo = (int)a->Wo; g.kh = a->kh; g.kw = a->kw; g.sh = a->sh; g.sw = a->sw;
  g.pt = a->pad_top; g.pl = a->pad_left;
  g.xs_b = a->xs_b; g.xs_h = a->xs_h; g.xs_w = a->xs_w; g.xs_c = a->xs_c;
  g.ld = a->ld;
  return SAV_OK;
}

}  // namespace sav

using namespace sav;

extern "C" int sav_conv_same_geom(int64_t in, int32_t k, int32_t stride, int64_t* out,
                                  int32_t* pad_lo) {
  if (in <= 0 || k <= 0 || stride <= 0) return set_error(SAV_ERR_SHAPE, "sav_conv_same_geom");
  const int64_t o = (in + stride - 1) / stride;
  int64_t total = (o - 1) * stride + k - in;
  if (total < 0) total = 0;
  *out = o;
  *pad_lo = (int32_t)(total / 2);   // XLA / flax SAME: the extra element goes to the high side
  return SAV_OK;
}

extern "C" int sav_im2col(const void* x, int32_t x_is_fp32, void* cols, const sav_conv2d_geom* a,
                          void* stream) {
  Im2colGeom g;
  int rc = fill_geom(g, a, "sav_im2col");
  if (rc) return rc;
  const int64_t work = (int64_t)g.B * g.Ho * g.Wo * g.ld;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool vec8 = !x_is_fp32 && g.C % 8 == 0 && g.xs_c == 1 && g.xs_w % 8 == 0 && g.xs_h % 8 == 0 &&
                    g.xs_b % 8 == 0 && !(reinterpret_cast<uintptr_t>(x) & 15) &&
                    !(reinterpret_cast<uintptr_t>(cols) & 15);
  if (vec8)
    im2col_vec8_kernel<<<c2_blocks(work / 8), 256, 0, st>>>(reinterpret_cast<const bf16*>(x),
                                                           reinterpret_cast<bf16*>(cols), g);
  else if (x_is_fp32)
    im2col_kernel<float><<<c2_blocks(work), 256, 0, st>>>(reinterpret_cast<const float*>(x),
                                                          reinterpret_cast<bf16*>(cols), g);
  else
    im2col_kernel<bf16><<<c2_blocks(work), 256, 0, st>>>(reinterpret_cast<const bf16*>(x),
                                                         reinterpret_cast<bf16*>(cols), g);
  count_launch();
  return check_launch("sav_im2col");
}

extern "C" int sav_col2im(const void* dcols, void* dx, int32_t dx_is_fp32, int32_t accumulate,
                          const sav_conv2d_geom* a, void* stream) {
  Im2colGeom g;
  int rc = fill_geom(g, a, "sav_col2im");
  if (rc) return rc;
  if (g.xs_c != 1 || g.xs_w != g.C || g.xs_h != (int64_t)g.W * g.C || g.xs_b != (int64_t)g.H * g.W * g.C)
    return set_error(SAV_ERR_SHAPE, "sav_col2im: dx must be contiguous [B,H,W,C]");
  const int64_t work = (int64_t)g.B * g.H * g.W * g.C;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool vec8 = g.C % 8 == 0 && !(reinterpret_cast<uintptr_t>(dcols) & 15) &&
                    !(reinterpret_cast<uintptr_t>(dx) & 31);
  if (vec8 && dx_is_fp32)
    col2im_vec8_kernel<float><<<c2_blocks(work / 8), 256, 0, st>>>(reinterpret_cast<const bf16*>(dcols),
                                                                  reinterpret_cast<float*>(dx), g, accumulate);
  else if (vec8)
    col2im_vec8_kernel<bf16><<<c2_blocks(work / 8), 256, 0, st>>>(reinterpret_cast<const bf16*>(dcols),
                                                                 reinterpret_cast<bf16*>(dx), g, accumulate);
  else if (dx_is_fp32)
    col2im_kernel<float><<<c2_blocks(work), 256, 0, st>>>(reinterpret_cast<const bf16*>(dcols),
                                                          reinterpret_cast<float*>(dx), g, accumulate);
  else
    col2im_kernel<bf16><<<c2_blocks(work), 256, 0, st>>>(reinterpret_cast<const bf16*>(dcols),
                                                         reinterpret_cast<bf16*>(dx), g, accumulate);
  count_launch();
  return check_launch("sav_col2im");
}

extern "C" int sav_bn_stats(const void* y, int64_t rows, int64_t C, float* sum, float* sumsq,
                            void* stream) {
  if (rows <= 0 || C <= 0 || C % 2) return set_error(SAV_ERR_SHAPE, "sav_bn_stats: C %% 2");
  int rc;
  if ((rc = sav_fill_f32(sum, 0.f, C, stream)) || (rc = sav_fill_f32(sumsq, 0.f, C, stream))) return rc;
  if (!(reinterpret_cast<uintptr_t>(sum) & 15) && !(reinterpret_cast<uintptr_t>(sumsq) & 15) &&
      bn_stats_vec8(y, rows, C, sum, sumsq, reinterpret_cast<cudaStream_t>(stream))) {
    count_launch();
    return check_launch("sav_bn_stats");
  }
  const int gx = (int)((C / 2 + 63) / 64);
  int64_t gy = (rows + 63) / 64;
  const int64_t cap = (148 * 8 + gx - 1) / gx;
  if (gy > cap) gy = cap;
  if (gy < 1) gy = 1;
  bn_stats_kernel<<<dim3(gx, (unsigned)gy), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(y), rows, (int)C, sum, sumsq);
  count_launch();
  return check_launch("sav_bn_stats");
}

extern "C" int sav_bn_update_running(const float* mean, const float* var, float* ra_mean,
                                     float* ra_var, float momentum, int64_t C, void* stream) {
  if (C <= 0) return SAV_OK;
  bn_running_kernel<<<(int)((C + 127) / 128), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      mean, var, ra_mean, ra_var, momentum, (int)C);
  count_launch();
  return check_launch("sav_bn_update_running");
}

extern "C" int sav_batchnorm_eval(const void* y, const float* ra_mean, const float* ra_var,
                                  const float* gamma, const float* beta, float eps, void* z,
                                  int64_t rows, int64_t C, void* stream) {
  if (rows <= 0 || C <= 0 || C % 2) return set_error(SAV_ERR_SHAPE, "sav_batchnorm_eval: C %% 2");
  bn_eval_kernel<<<c2_blocks(rows * (C / 2)), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(y), ra_mean, ra_var, gamma, beta, eps,
      reinterpret_cast<bf16*>(z), rows, (int)(C / 2));
  count_launch();
  return check_launch("sav_batchnorm_eval");
}
