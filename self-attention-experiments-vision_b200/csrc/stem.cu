// Patch-embedding stem data movement (HBM-bound): HWCN fp32 images -> token-major bf16 patches
// (train.py:80-81 + patch_embed.py:19-22), cls/pos-embed add (vit.py:81-85,
// position_embed.py:54-56) -> fp32 residual stream, and its reverse-mode.
#include "common.cuh"
#include "host.h"

namespace sav {

// images [H,W,C,N] fp32 (N contiguous).  One block: fixed image row h, fixed patch column gx,
// 64 consecutive n.  Reads pw*C rows of 64 contiguous floats; writes, per n, pw*C contiguous bf16.
// Algorithmic bytes: 4 (read) + 2 (write) per pixel-channel.
__global__ void __launch_bounds__(256)
patchify_hwcn_kernel(const float* __restrict__ img, bf16* __restrict__ out, int N, int H, int W,
                     int C, int ph, int pw) {
  extern __shared__ float tile[];  // [pw*C][65]
  const int gw = W / pw, gh = H / ph;
  const int gx = blockIdx.x % gw;
  const int h = blockIdx.x / gw;
  const int n0 = blockIdx.y * 64;
  const int F = pw * C;  // contiguous (w-in-patch, c) run
  const int64_t in_base = ((int64_t)h * W + (int64_t)gx * pw) * C * N;
  for (int i = threadIdx.x; i < F * 64; i += blockDim.x) {
    const int f = i >> 6, n = i & 63;
    tile[f * 65 + n] = (n0 + n < N) ? img[in_base + (int64_t)f * N + n0 + n] : 0.f;
  }
  __syncthreads();
  const int gy = h / ph, py = h % ph;
  const int K = ph * pw * C;
  for (int i = threadIdx.x; i < 64 * (F / 2); i += blockDim.x) {
    const int n = i / (F / 2), f2 = (i % (F / 2)) * 2;
    if (n0 + n < N) {
      const int64_t row = ((int64_t)(n0 + n) * gh + gy) * gw + gx;
      const uint32_t v = pack_bf16(tile[f2 * 65 + n], tile[(f2 + 1) * 65 + n]);
      *reinterpret_cast<uint32_t*>(out + row * K + (int64_t)py * F + f2) = v;
    }
  }
}

// images [N,H,W,C] fp32: each output row segment (py) is pw*C contiguous input floats.
__global__ void patchify_nhwc_kernel(const float* __restrict__ img, bf16* __restrict__ out,
                                     int64_t total2, int H, int W, int C, int ph, int pw) {
  const int gw = W / pw, gh = H / ph;
  const int F = pw * C, K = ph * F;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total2;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = i * 2;
    const int64_t row = e / K;
    const int k = (int)(e - row * K);
    const int py = k / F, f = k % F;
    const int gx = (int)(row % gw);
    const int gy = (int)((row / gw) % gh);
    const int64_t n = row / ((int64_t)gw * gh);
    const float2 v = *reinterpret_cast<const float2*>(
        img + ((n * H + (int64_t)gy * ph + py) * W + (int64_t)gx * pw) * C + f);
    *reinterpret_cast<uint32_t*>(out + e) = pack_bf16(v.x, v.y);
  }
}

// x[b, has_cls + i, :] = tok[b, i, :] + pos[has_cls + i, :];  x[b, 0, :] = cls + pos[0]
__global__ void embed_fwd_kernel(const bf16* __restrict__ tok, const float* __restrict__ cls,
                                 const float* __restrict__ pos, float* __restrict__ x, int64_t B,
                                 int n_tok, int C, int has_cls) {
  const int NT = n_tok + has_cls;
  const int64_t total2 = B * NT * (int64_t)C / 2;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total2;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = i * 2;
    const int c = (int)(e % C);
    const int64_t r = e / C;
    const int t = (int)(r % NT);
    const int64_t b = r / NT;
    float2 v;
    if (has_cls && t == 0) {
      v = make_float2(cls[c], cls[c + 1]);
    } else {
      v = unpack_bf16(*reinterpret_cast<const uint32_t*>(tok + (b * n_tok + (t - has_cls)) * C + c));
    }
    v.x += pos[(int64_t)t * C + c];
    v.y += pos[(int64_t)t * C + c + 1];
    *reinterpret_cast<float2*>(x + e) = v;
  }
}

// thread per (t, c-pair): loops over b; writes dpos[t,c], dcls (t==0), copies dtok.
__global__ void embed_bwd_kernel(const float* __restrict__ dx, bf16* __restrict__ dtok,
                                 float* __restrict__ dcls, float* __restrict__ dpos, int64_t B,
                                 int n_tok, int C, int has_cls, int accumulate) {
  const int NT = n_tok + has_cls;
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= (int64_t)NT * C / 2) return;
  const int c = (int)((i * 2) % C);
  const int t = (int)((i * 2) / C);
  float2 acc = make_float2(0.f, 0.f);
  for (int64_t b = 0; b < B; ++b) {
    const float2 f = *reinterpret_cast<const float2*>(dx + (b * NT + t) * C + c);
    acc.x += f.x;
    acc.y += f.y;
    if (dtok && t >= has_cls)
      *reinterpret_cast<uint32_t*>(dtok + (b * n_tok + (t - has_cls)) * C + c) = pack_bf16(f.x, f.y);
  }
  float* p = dpos + (int64_t)t * C + c;
  if (accumulate) { p[0] += acc.x; p[1] += acc.y; } else { p[0] = acc.x; p[1] = acc.y; }
  if (has_cls && t == 0 && dcls) {
    if (accumulate) { dcls[c] += acc.x; dcls[c + 1] += acc.y; }
    else { dcls[c] = acc.x; dcls[c + 1] = acc.y; }
  }
}

}  // namespace sav

using namespace sav;

extern "C" int sav_patchify(const float* images, void* patches, int64_t N, int64_t H, int64_t W,
                            int64_t C, int64_t ph, int64_t pw, int32_t hwcn, void* stream) {
  if (N <= 0 || H % ph || W % pw || (pw * C) % 2)
    return set_error(SAV_ERR_SHAPE, "sav_patchify: H,W must be multiples of the patch shape");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (hwcn) {
    dim3 grid((unsigned)(H * (W / pw)), (unsigned)((N + 63) / 64));
    const size_t sm = (size_t)pw * C * 65 * sizeof(float);
    if (sm > 48 * 1024) return set_error(SAV_ERR_SHAPE, "sav_patchify: pw*C too large");
    patchify_hwcn_kernel<<<grid, 256, sm, st>>>(images, reinterpret_cast<bf16*>(patches), (int)N,
                                                (int)H, (int)W, (int)C, (int)ph, (int)pw);
  } else {
    const int64_t total2 = N * H * W * C / 2;
    int blocks = (int)((total2 + 255) / 256 < 148 * 16 ? (total2 + 255) / 256 : 148 * 16);
    patchify_nhwc_kernel<<<blocks, 256, 0, st>>>(images, reinterpret_cast<bf16*>(patches), total2,
                                                 (int)H, (int)W, (int)C, (int)ph, (int)pw);
  }
  count_launch();
  return check_launch("sav_patchify");
}

extern "C" int sav_embed_fwd(const void* tok, const float* cls, const float* pos, void* x,
                             int64_t B, int64_t n_tok, int64_t C, int32_t has_cls, void* stream) {
  if (B <= 0 || n_tok <= 0 || C % 2) return set_error(SAV_ERR_SHAPE, "sav_embed_fwd: bad shape");
  const int64_t total2 = B * (n_tok + has_cls) * C / 2;
  int blocks = (int)((total2 + 255) / 256 < 148 * 16 ? (total2 + 255) / 256 : 148 * 16);
  embed_fwd_kernel<<<blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const bf16*>(tok), cls, pos, reinterpret_cast<float*>(x), B, (int)n_tok,
      (int)C, has_cls);
  count_launch();
  return check_launch("sav_embed_fwd");
}

extern "C" int sav_embed_bwd(const void* dx, void* dtok, float* dcls, float* dpos, int64_t B,
                             int64_t n_tok, int64_t C, int32_t has_cls, int32_t accumulate,
                             void* stream) {
  if (B <= 0 || n_tok <= 0 || C % 2) return set_error(SAV_ERR_SHAPE, "sav_embed_bwd: bad shape");
  const int64_t work = (n_tok + has_cls) * C / 2;
  embed_bwd_kernel<<<(int)((work + 127) / 128), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float*>(dx), reinterpret_cast<bf16*>(dtok), dcls, dpos, B, (int)n_tok,
      (int)C, has_cls, accumulate);
  count_launch();
  return check_launch("sav_embed_bwd");
}
