// Loss head (train.py:83-91) and optimizer chain (train.py:25-27,100; simple_train.py:25-27,90)
// as HBM-bound single-pass kernels over the flat parameter arena (28 B/param for the Adam step).
#include "common.cuh"
#include "host.h"

namespace sav {

__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float wmax(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// warp per sample.  target = (1-a)*onehot + a/classes (optax.smooth_labels);
// loss_b = -sum target*log_softmax;  dlogits = (softmax - target) * grad_scale.
__global__ void __launch_bounds__(128)
xent_kernel(const bf16* __restrict__ logits, int64_t ld, const int32_t* __restrict__ labels,
            const int32_t* __restrict__ mix_labels, const float* __restrict__ ratio,
            int B, int classes, float alpha, float grad_scale, bf16* __restrict__ dlogits,
            float* __restrict__ loss_out) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (b >= B) return;
  const bf16* row = logits + (int64_t)b * ld;
  float mx = -INFINITY;
  for (int c = lane; c < classes; c += 32) mx = fmaxf(mx, __bfloat162float(row[c]));
  mx = wmax(mx);
  float se = 0.f, sl = 0.f;
  int arg = 0x7fffffff;
  for (int c = lane; c < classes; c += 32) {
    const float l = __bfloat162float(row[c]);
    se += __expf(l - mx);
    sl += l;
    if (l == mx && c < arg) arg = c;
  }
  se = wsum(se);
  sl = wsum(sl);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) arg = min(arg, __shfl_xor_sync(0xffffffffu, arg, o));
  const float lse = mx + logf(se);
  const int lab = labels[b];
  const float l_lab = __bfloat162float(row[lab]);
  // mixup / cutmix targets (train.py:84-87): y = r*onehot(lab) + (1-r)*onehot(lab2)
  const int lab2 = mix_labels ? mix_labels[b] : lab;
  const float r = mix_labels ? ratio[b] : 1.f;
  const float l_lab2 = __bfloat162float(row[lab2]);
  const float w1 = (1.f - alpha) * r, w2 = (1.f - alpha) * (1.f - r);
  // -sum_c t_c (l_c - lse) = lse - w1 l_lab - w2 l_lab2 - a/classes * sum_c l_c
  const float loss = lse - w1 * l_lab - w2 * l_lab2 - alpha / classes * sl;
  const float inv = 1.f / se;
  for (int c = lane; c < classes; c += 32) {
    const float l = __bfloat162float(row[c]);
    const float t = (c == lab ? w1 : 0.f) + (c == lab2 ? w2 : 0.f) + alpha / classes;
    dlogits[(int64_t)b * ld + c] = __float2bfloat16((__expf(l - mx) * inv - t) * grad_scale);
  }
  if (lane == 0) {
    atomicAdd(loss_out, loss);
    if (arg == lab) atomicAdd(loss_out + 1, 1.f);
  }
}

__global__ void cast_kernel(const float* __restrict__ s, bf16* __restrict__ d, int64_t n) {
  const int64_t n4 = n / 4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4;
       i += (int64_t)gridDim.x * blockDim.x) {
    const float4 v = *reinterpret_cast<const float4*>(s + i * 4);
    *reinterpret_cast<uint2*>(d + i * 4) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const int64_t i = n4 * 4 + threadIdx.x;
    d[i] = __float2bfloat16(s[i]);
  }
}

__global__ void sumsq_kernel(const float* __restrict__ g, int64_t n, float* __restrict__ out) {
  float acc = 0.f;
  const int64_t n4 = n / 4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4;
       i += (int64_t)gridDim.x * blockDim.x) {
    const float4 v = *reinterpret_cast<const float4*>(g + i * 4);
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const float v = g[n4 * 4 + threadIdx.x];
    acc += v * v;
  }
  acc = wsum(acc);
  __shared__ float red[8];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += red[i];
    atomicAdd(out, s);
  }
}

// clip_by_global_norm -> scale_by_adam -> additive_weight_decay -> scale(-lr) -> apply_updates
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                            float* __restrict__ m, float* __restrict__ v, bf16* __restrict__ pb,
                            int64_t n, const float* __restrict__ sumsq, float max_norm, float lr,
                            float b1, float b2, float omb1, float omb2, float eps, float wd,
                            float bc1, float bc2, const float* __restrict__ hyper) {
  if (hyper) { lr = hyper[0]; bc1 = hyper[1]; bc2 = hyper[2]; }
  float clip = 1.f;
  if (max_norm > 0.f && sumsq) {
    const float gn = sqrtf(sumsq[0]);
    if (gn >= max_norm) clip = max_norm / gn;  // optax: g * max_norm / ||g|| when ||g|| >= max_norm
  }
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const float gi = g[i] * clip;
    // optax update_moment: (1 - decay) * g^order + decay * moment, with (1 - decay) formed in
    // double from the Python float and THEN rounded (omb = fp32(1 - decay), not 1.f - fp32(decay))
    const float mi = omb1 * gi + b1 * m[i];
    const float vi = omb2 * (gi * gi) + b2 * v[i];
    m[i] = mi;
    v[i] = vi;
    const float pi = p[i];
    const float upd = (mi / bc1) / (sqrtf(vi / bc2) + eps) + wd * pi;
    const float np = pi - lr * upd;
    p[i] = np;
    if (pb) pb[i] = __float2bfloat16(np);
  }
}

// utils.py:20-31 topk_correct: jnp.argsort is ascending and stable, the top-k are the last k
// indices, so among equal logits the larger class index ranks higher.  rank = number of classes
// ranked above the label; correct_k = rank < k.  Warp per example.
__global__ void __launch_bounds__(128)
topk_kernel(const bf16* __restrict__ logits, int64_t ld, const int32_t* __restrict__ labels, int B,
            int classes, int nk, int4 ks, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (b >= B) return;
  const bf16* row = logits + (int64_t)b * ld;
  const int lab = labels[b];
  const float l_lab = __bfloat162float(row[lab]);
  int above = 0;
  for (int c = lane; c < classes; c += 32) {
    const float l = __bfloat162float(row[c]);
    above += (l > l_lab || (l == l_lab && c > lab)) ? 1 : 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) above += __shfl_xor_sync(0xffffffffu, above, o);
  if (lane == 0) {
    const int kk[4] = {ks.x, ks.y, ks.z, ks.w};
    for (int i = 0; i < nk; ++i) out[(int64_t)i * B + b] = above < kk[i] ? 1.f : 0.f;
  }
}

__global__ void add_kernel(const bf16* __restrict__ a, const bf16* __restrict__ b,
                           bf16* __restrict__ y, int64_t n8) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n8;
       i += (int64_t)gridDim.x * blockDim.x) {
    const uint4 ua = reinterpret_cast<const uint4*>(a)[i];
    const uint4 ub = reinterpret_cast<const uint4*>(b)[i];
    const uint32_t wa[4] = {ua.x, ua.y, ua.z, ua.w}, wb[4] = {ub.x, ub.y, ub.z, ub.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 fa = unpack_bf16(wa[j]), fb = unpack_bf16(wb[j]);
      o[j] = pack_bf16(fa.x + fb.x, fa.y + fb.y);
    }
    reinterpret_cast<uint4*>(y)[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

__global__ void copy2d_kernel(const bf16* __restrict__ src, int64_t lds, bf16* __restrict__ dst,
                              int64_t ldd, int64_t rows, int cols8) {
  const int64_t total = rows * cols8;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / cols8;
    const int c = (int)(i - r * cols8) * 8;
    *reinterpret_cast<uint4*>(dst + r * ldd + c) = *reinterpret_cast<const uint4*>(src + r * lds + c);
  }
}

// optax.warmup_cosine_decay_schedule (train.py:215-220) evaluated at `count`.
__device__ __forceinline__ float warmup_cosine(float init_v, float peak, int64_t warm, int64_t decay,
                                               float end_v, int64_t count) {
  if (count < warm) {
    // linear_schedule(init, peak, warm)
    const float frac = 1.f - (float)count / (float)warm;
    return (init_v - peak) * frac + peak;
  }
  // cosine_decay_schedule(peak, decay - warm, alpha = end / peak) at count - warm
  const int64_t D = decay - warm > 0 ? decay - warm : 1;
  int64_t c = count - warm;
  if (c > D) c = D;
  const float cosv = 0.5f * (1.f + cospif((float)c / (float)D));
  const float alpha = peak != 0.f ? end_v / peak : 0.f;
  return peak * ((1.f - alpha) * cosv + alpha);
}

// One thread: advances the device-resident step counter and writes {lr, 1-b1^t, 1-b2^t} for the
// Adam kernel that follows it IN THE SAME (possibly graph-replayed) stream, so no host value is
// read at execution time.  optax.scale_by_schedule evaluates the schedule at the count BEFORE the
// increment (first update uses schedule(0)); scale_by_adam bias-corrects with count + 1.
__global__ void schedule_kernel(int64_t* __restrict__ step, float* __restrict__ hyper, int mode,
                                float init_v, float peak, int64_t warm, int64_t decay, float end_v,
                                float b1, float b2) {
  if (threadIdx.x || blockIdx.x) return;
  const int64_t count = step[0];
  const int64_t t = count + 1;
  step[0] = t;
  if (mode == 0) hyper[0] = peak;
  else if (mode == 1) hyper[0] = warmup_cosine(init_v, peak, warm, decay, end_v, count);
  // mode 2: hyper[0] was written by the host (arbitrary Python schedule)
  hyper[1] = 1.f - powf(b1, (float)t);
  hyper[2] = 1.f - powf(b2, (float)t);
}

// Philox4x32-10 counter-based generator: key = seed, counter = (sample, call index, step lo, hi).
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}

// stochastic_depth.py:16-27: factor[b] = floor(keep + U[b]) / keep, U uniform in [0,1).
// The step is read from the device counter (if given) so that graph replays resample.
__global__ void drop_path_kernel(uint64_t seed, const int64_t* __restrict__ step_dev,
                                 int64_t step_host, int64_t call, int B, float keep,
                                 float* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const int64_t step = step_dev ? step_dev[0] : step_host;
  const uint4 r = philox4x32_10(make_uint4((uint32_t)b, (uint32_t)call, (uint32_t)step,
                                           (uint32_t)((uint64_t)step >> 32)),
                                make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
  const float u = (float)(r.x >> 8) * (1.f / 16777216.f);
  out[b] = floorf(keep + u) / keep;
}

static int grid_for(int64_t work, int threads) {
  int64_t b = (work + threads - 1) / threads;
  const int64_t cap = 148 * 8;
  return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

}  // namespace sav

using namespace sav;

extern "C" int sav_softmax_xent_mix(const void* logits, int64_t ld, const int32_t* labels,
                                    const int32_t* mix_labels, const float* ratio, int64_t B,
                                    int64_t classes, float label_smoothing, float grad_scale,
                                    void* dlogits, float* loss_out, void* stream) {
  if (B <= 0 || classes <= 0) return set_error(SAV_ERR_SHAPE, "sav_softmax_xent: bad shape");
  if ((mix_labels == nullptr) != (ratio == nullptr))
    return set_error(SAV_ERR_ARG, "sav_softmax_xent_mix: mix_labels and ratio go together");
  int rc = sav_fill_f32(loss_out, 0.f, 2, stream);
  if (rc) return rc;
  xent_kernel<<<(int)((B + 3) / 4), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(logits), ld, labels, mix_labels, ratio, (int)B, (int)classes,
      label_smoothing, grad_scale, reinterpret_cast<bf16*>(dlogits), loss_out);
  count_launch();
  return check_launch("sav_softmax_xent");
}

extern "C" int sav_softmax_xent(const void* logits, int64_t ld, const int32_t* labels, int64_t B,
                                int64_t classes, float label_smoothing, float grad_scale,
                                void* dlogits, float* loss_out, void* stream) {
  return sav_softmax_xent_mix(logits, ld, labels, nullptr, nullptr, B, classes, label_smoothing,
                              grad_scale, dlogits, loss_out, stream);
}

extern "C" int sav_schedule_step(int64_t* step, float* hyper, int32_t mode, float init_value,
                                 float peak_value, int64_t warmup_steps, int64_t decay_steps,
                                 float end_value, double b1, double b2, void* stream) {
  if (!step || !hyper) return set_error(SAV_ERR_ARG, "sav_schedule_step: null step / hyper");
  if (mode < 0 || mode > 2) return set_error(SAV_ERR_ARG, "sav_schedule_step: mode in {0,1,2}");
  if (mode == 1 && warmup_steps < 0) return set_error(SAV_ERR_ARG, "sav_schedule_step: warmup < 0");
  schedule_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      step, hyper, mode, init_value, peak_value, warmup_steps, decay_steps, end_value, (float)b1,
      (float)b2);
  count_launch();
  return check_launch("sav_schedule_step");
}

extern "C" int sav_drop_path_factors(uint64_t seed, const int64_t* step_dev, int64_t step_host,
                                     int64_t call_index, int64_t B, float keep, float* out,
                                     void* stream) {
  if (B <= 0) return SAV_OK;
  if (!(keep > 0.f && keep <= 1.f)) return set_error(SAV_ERR_ARG, "sav_drop_path_factors: keep in (0,1]");
  drop_path_kernel<<<(int)((B + 127) / 128), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      seed, step_dev, step_host, call_index, (int)B, keep, out);
  count_launch();
  return check_launch("sav_drop_path_factors");
}

extern "C" int sav_topk_correct(const void* logits, int64_t ld, const int32_t* labels, int64_t B,
                                int64_t classes, int32_t nk, int32_t k0, int32_t k1, int32_t k2,
                                int32_t k3, float* out, void* stream) {
  if (B <= 0 || classes <= 0 || nk < 1 || nk > 4)
    return set_error(SAV_ERR_SHAPE, "sav_topk_correct: bad shape (1 <= nk <= 4)");
  topk_kernel<<<(int)((B + 3) / 4), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(logits), ld, labels, (int)B, (int)classes, nk,
      make_int4(k0, k1, k2, k3), out);
  count_launch();
  return check_launch("sav_topk_correct");
}

extern "C" int sav_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream) {
  if (n <= 0) return SAV_OK;
  if (!aligned16(src) || (reinterpret_cast<uintptr_t>(dst) & 7))
    return set_error(SAV_ERR_ALIGN, "sav_cast_f32_bf16: alignment");
  cast_kernel<<<grid_for(n / 4 + 1, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      src, reinterpret_cast<bf16*>(dst), n);
  count_launch();
  return check_launch("sav_cast_f32_bf16");
}

extern "C" int sav_sumsq(const float* g, int64_t n, float* out, void* stream) {
  if (n <= 0) return SAV_OK;
  if (!aligned16(g)) return set_error(SAV_ERR_ALIGN, "sav_sumsq: alignment");
  sumsq_kernel<<<grid_for(n / 4 + 1, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(g, n, out);
  count_launch();
  return check_launch("sav_sumsq");
}

extern "C" int sav_adam_step(float* p, const float* g, float* m, float* v, void* p_bf16,
                             int64_t n, const float* sumsq, float max_norm, float lr, double b1d,
                             double b2d, float eps, float wd, int64_t t, const float* hyper,
                             void* stream) {
  const float b1 = (float)b1d, b2 = (float)b2d;
  const float omb1 = (float)(1.0 - b1d), omb2 = (float)(1.0 - b2d);
  if (n <= 0) return SAV_OK;
  if (t < 1 && !hyper) return set_error(SAV_ERR_ARG, "sav_adam_step: step count t starts at 1");
  const float bc1 = 1.f - powf(b1, (float)t), bc2 = 1.f - powf(b2, (float)t);
  adam_kernel<<<grid_for(n, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      p, g, m, v, reinterpret_cast<bf16*>(p_bf16), n, sumsq, max_norm, lr, b1, b2, omb1, omb2, eps, wd,
      bc1, bc2, hyper);
  count_launch();
  return check_launch("sav_adam_step");
}

extern "C" int sav_add_bf16(const void* a, const void* b, void* y, int64_t n, void* stream) {
  if (n <= 0) return SAV_OK;
  if (n % 8 || !aligned16(a) || !aligned16(b) || !aligned16(y))
    return set_error(SAV_ERR_ALIGN, "sav_add_bf16: n %% 8 and 16 B alignment required");
  add_kernel<<<grid_for(n / 8, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(a), reinterpret_cast<const bf16*>(b),
      reinterpret_cast<bf16*>(y), n / 8);
  count_launch();
  return check_launch("sav_add_bf16");
}

extern "C" int sav_copy2d_bf16(const void* src, int64_t lds, void* dst, int64_t ldd, int64_t rows,
                               int64_t cols, void* stream) {
  if (rows <= 0 || cols <= 0) return SAV_OK;
  if (cols % 8 || lds % 8 || ldd % 8 || !aligned16(src) || !aligned16(dst))
    return set_error(SAV_ERR_ALIGN, "sav_copy2d_bf16: cols/ld %% 8 and 16 B alignment required");
  copy2d_kernel<<<grid_for(rows * (cols / 8), 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(src), lds, reinterpret_cast<bf16*>(dst), ldd, rows, (int)(cols / 8));
  count_launch();
  return check_launch("sav_copy2d_bf16");
}
