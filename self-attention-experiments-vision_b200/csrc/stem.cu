// Patch-embedding stem data movement (HBM-bound): HWCN fp32 images -> token-major bf16 patches
// (train.py:80-81 + patch_embed.py:19-22), cls/pos-embed add (vit.py:81-85,
// position_embed.py:54-56) -> fp32 residual stream, and its reverse-mode.
#include "common.cuh"
#include "host.h"

namespace sav {

// images [H,W,C,N] fp32 (N contiguous).  One block: PY consecutive image rows of one patch row,
// one patch column gx, 32 consecutive n.  Reads 128-byte runs (32 floats along N) for each of the
// PY * pw*C (row, w-in-patch, c) features; writes, per n, PY * pw*C CONTIGUOUS bf16 (the features of
// consecutive rows of a patch are adjacent in the (ph pw c) order of patch_embed.py:19-22), i.e.
// 384-byte runs at PY = 4 instead of the 96-byte runs of a one-row tile.
// Algorithmic bytes: 4 (read) + 2 (write) per pixel-channel.
template <int PY>
__global__ void __launch_bounds__(256)
patchify_hwcn_kernel(const float* __restrict__ img, bf16* __restrict__ out, int N, int H, int W,
                     int C, int ph, int pw) {
  extern __shared__ float tile[];  // [PY * pw*C][33]
  const int gw = W / pw, gh = H / ph;
  const int rows_per_patch = ph / PY;            // row groups per patch row
  const int gx = blockIdx.x % gw;
  const int hg = blockIdx.x / gw;                // row group index over the image
  const int gy = hg / rows_per_patch, py0 = (hg % rows_per_patch) * PY;
  const int n0 = blockIdx.y * 32;
  const int F = pw * C;                          // contiguous (w-in-patch, c) run of one image row
  const int FT = PY * F;
  if ((N & 3) == 0) {
    // 16-byte loads along N, several per thread in flight (the loop is latency-bound otherwise)
#pragma unroll 6
    for (int i = threadIdx.x; i < FT * 8; i += blockDim.x) {
      const int ft = i >> 3, n = (i & 7) * 4;
      const int r = ft / F, f = ft - r * F;
      const int h = gy * ph + py0 + r;
      const int64_t in = (((int64_t)h * W + (int64_t)gx * pw) * C + f) * N + n0 + n;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (n0 + n < N) v = __ldcs(reinterpret_cast<const float4*>(img + in));
      float* t = tile + ft * 33 + n;
      t[0] = v.x; t[1] = v.y; t[2] = v.z; t[3] = v.w;
    }
  } else {
    for (int i = threadIdx.x; i < FT * 32; i += blockDim.x) {
      const int ft = i >> 5, n = i & 31;
      const int r = ft / F, f = ft - r * F;
      const int h = gy * ph + py0 + r;
      const int64_t in = (((int64_t)h * W + (int64_t)gx * pw) * C + f) * N + n0 + n;
      tile[ft * 33 + n] = (n0 + n < N) ? img[in] : 0.f;
    }
  }
  __syncthreads();
  const int K = ph * F;
  for (int i = threadIdx.x; i < 32 * (FT / 2); i += blockDim.x) {
    const int n = i / (FT / 2), f2 = (i - n * (FT / 2)) * 2;
    if (n0 + n < N) {
      const int64_t row = ((int64_t)(n0 + n) * gh + gy) * gw + gx;
      const uint32_t v = pack_bf16(tile[f2 * 33 + n], tile[(f2 + 1) * 33 + n]);
      *reinterpret_cast<uint32_t*>(out + row * K + (int64_t)py0 * F + f2) = v;
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

// same, 8 channels per thread (16-byte bf16 load, two 16-byte fp32 stores); C % 8 == 0
__global__ void embed_fwd8_kernel(const bf16* __restrict__ tok, const float* __restrict__ cls,
                                  const float* __restrict__ pos, float* __restrict__ x, int64_t B,
                                  int n_tok, int C8, int has_cls) {
  const int NT = n_tok + has_cls;
  const int64_t total = B * NT * (int64_t)C8;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C8) * 8;
    const int64_t r = i / C8;
    const int t = (int)(r % NT);
    const int64_t b = r / NT;
    float v[8];
    if (has_cls && t == 0) {
      const float4 a = *reinterpret_cast<const float4*>(cls + c), d = *reinterpret_cast<const float4*>(cls + c + 4);
      v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = d.x; v[5] = d.y; v[6] = d.z; v[7] = d.w;
    } else {
      const uint4 u = *reinterpret_cast<const uint4*>(tok + (b * n_tok + (t - has_cls)) * (int64_t)(C8 * 8) + c);
      const float2 a = unpack_bf16(u.x), d = unpack_bf16(u.y), e = unpack_bf16(u.z), f = unpack_bf16(u.w);
      v[0] = a.x; v[1] = a.y; v[2] = d.x; v[3] = d.y; v[4] = e.x; v[5] = e.y; v[6] = f.x; v[7] = f.y;
    }
    const float4 p0 = __ldg(reinterpret_cast<const float4*>(pos + (int64_t)t * C8 * 8 + c));
    const float4 p1 = __ldg(reinterpret_cast<const float4*>(pos + (int64_t)t * C8 * 8 + c + 4));
    float* o = x + (r * C8) * 8 + c;
    *reinterpret_cast<float4*>(o) = make_float4(v[0] + p0.x, v[1] + p0.y, v[2] + p0.z, v[3] + p0.w);
    *reinterpret_cast<float4*>(o + 4) = make_float4(v[4] + p1.x, v[5] + p1.y, v[6] + p1.z, v[7] + p1.w);
  }
}

// thread per (t, c-pair): loops over b; writes dpos[t,c], dcls (t==0), copies dtok.  When there are
// few (t, c) pairs and many sequences (TNT's pixel stream: 16 x 40 features, 12 544 sequences per
// 64 images -- 320 threads looped over all of them in 2.9 ms, profiles/r02ay_timeline_tnt_s_patch16.txt)
// blockIdx.y splits the batch and the slices combine with atomicAdd (accumulate mode only).
__global__ void embed_bwd_kernel(const float* __restrict__ dx, bf16* __restrict__ dtok,
                                 float* __restrict__ dcls, float* __restrict__ dpos, int64_t Bfull,
                                 int n_tok, int C, int has_cls, int accumulate, int64_t chunk) {
  const int NT = n_tok + has_cls;
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= (int64_t)NT * C / 2) return;
  const int c = (int)((i * 2) % C);
  const int t = (int)((i * 2) / C);
  float2 acc = make_float2(0.f, 0.f);
  int64_t b = (int64_t)blockIdx.y * chunk;
  const int64_t B = b + chunk < Bfull ? b + chunk : Bfull;
  for (; b + 8 <= B; b += 8) {   // eight independent loads in flight per thread
    float2 f[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) f[u] = *reinterpret_cast<const float2*>(dx + ((b + u) * NT + t) * C + c);
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      acc.x += f[u].x;
      acc.y += f[u].y;
      if (dtok && t >= has_cls)
        *reinterpret_cast<uint32_t*>(dtok + ((b + u) * n_tok + (t - has_cls)) * C + c) = pack_bf16(f[u].x, f[u].y);
    }
  }
  for (; b < B; ++b) {
    const float2 f = *reinterpret_cast<const float2*>(dx + (b * NT + t) * C + c);
    acc.x += f.x;
    acc.y += f.y;
    if (dtok && t >= has_cls)
      *reinterpret_cast<uint32_t*>(dtok + (b * n_tok + (t - has_cls)) * C + c) = pack_bf16(f.x, f.y);
  }
  float* p = dpos + (int64_t)t * C + c;
  if (gridDim.y > 1) {
    atomicAdd(p, acc.x);
    atomicAdd(p + 1, acc.y);
    if (has_cls && t == 0 && dcls) { atomicAdd(dcls + c, acc.x); atomicAdd(dcls + c + 1, acc.y); }
    return;
  }
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
    int PY = ph % 4 == 0 ? 4 : (ph % 2 == 0 ? 2 : 1);
    while (PY > 1 && (size_t)PY * pw * C * 33 * sizeof(float) > 48 * 1024) PY /= 2;   // 32-wide patches
    dim3 grid((unsigned)((H / PY) * (W / pw)), (unsigned)((N + 31) / 32));
    const size_t sm = (size_t)PY * pw * C * 33 * sizeof(float);
    if (sm > 48 * 1024) return set_error(SAV_ERR_SHAPE, "sav_patchify: pw*C too large");
    bf16* o = reinterpret_cast<bf16*>(patches);
    if (PY == 4) patchify_hwcn_kernel<4><<<grid, 256, sm, st>>>(images, o, (int)N, (int)H, (int)W, (int)C, (int)ph, (int)pw);
    else if (PY == 2) patchify_hwcn_kernel<2><<<grid, 256, sm, st>>>(images, o, (int)N, (int)H, (int)W, (int)C, (int)ph, (int)pw);
    else patchify_hwcn_kernel<1><<<grid, 256, sm, st>>>(images, o, (int)N, (int)H, (int)W, (int)C, (int)ph, (int)pw);
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
  if (C % 8 == 0 && aligned16(tok) && aligned16(pos) && aligned16(x) && (!has_cls || aligned16(cls))) {
    const int64_t total8 = B * (n_tok + has_cls) * (C / 8);
    const int b8 = (int)((total8 + 255) / 256 < 148 * 16 ? (total8 + 255) / 256 : 148 * 16);
    embed_fwd8_kernel<<<b8, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const bf16*>(tok), cls, pos, reinterpret_cast<float*>(x), B, (int)n_tok,
        (int)(C / 8), has_cls);
  } else {
    embed_fwd_kernel<<<blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const bf16*>(tok), cls, pos, reinterpret_cast<float*>(x), B, (int)n_tok,
        (int)C, has_cls);
  }
  count_launch();
  return check_launch("sav_embed_fwd");
}

extern "C" int sav_embed_bwd(const void* dx, void* dtok, float* dcls, float* dpos, int64_t B,
                             int64_t n_tok, int64_t C, int32_t has_cls, int32_t accumulate,
                             void* stream) {
  if (B <= 0 || n_tok <= 0 || C % 2) return set_error(SAV_ERR_SHAPE, "sav_embed_bwd: bad shape");
  const int64_t work = (n_tok + has_cls) * C / 2;
  // enough threads to fill the machine: split the batch when (tokens x channels) alone cannot
  int64_t split = 1;
  if (accumulate && work < 148 * 1024) {
    split = (148 * 2048 + work - 1) / work;
    if (split > B / 8) split = B / 8;
    if (split < 1) split = 1;
    if (split > 65535) split = 65535;
  }
  const int64_t chunk = (B + split - 1) / split;
  split = (B + chunk - 1) / chunk;
  embed_bwd_kernel<<<dim3((unsigned)((work + 127) / 128), (unsigned)split), 128, 0,
                     reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float*>(dx), reinterpret_cast<bf16*>(dtok), dcls, dpos, B, (int)n_tok,
      (int)C, has_cls, accumulate, chunk);
  count_launch();
  return check_launch("sav_embed_bwd");
}

// ---- TNT pixel embedding input (models/tnt.py:19-26): 'b (h p1) (w p2) c -> (b h w) p1 p2 c' then
// 'n (p1 t1) (p2 t2) c -> n (p1 p2) (c t1 t2)': every ph x pw patch becomes (ph/th)*(pw/tw) "pixel"
// tokens of C*th*tw features, feature order (c, t1, t2) with t2 fastest.  fp32 image in ([N,H,W,C]
// or the loader's [H,W,C,N], train.py:80), bf16 GEMM operand out; a pure gather (images need no
// gradient).
namespace sav {
__global__ void __launch_bounds__(256)
pixel_patchify_kernel(const float* __restrict__ img, bf16* __restrict__ out, int64_t total, int N, int H,
                      int W, int C, int ph, int pw, int th, int tw, int64_t sy, int64_t sx, int64_t sc,
                      int64_t sb) {
  const int gw = W / pw, gh = H / ph, q1 = ph / th, q2 = pw / tw;
  const int feat = C * th * tw;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = idx;
    const int f = (int)(r % feat); r /= feat;
    const int pos = (int)(r % (q1 * q2)); r /= (q1 * q2);
    const int w = (int)(r % gw); r /= gw;
    const int h = (int)(r % gh); r /= gh;
    const int b = (int)r;
    const int t2 = f % tw, t1 = (f / tw) % th, c = f / (tw * th);
    const int p2 = pos % q2, p1 = pos / q2;
    const int y = h * ph + p1 * th + t1, x = w * pw + p2 * tw + t2;
    out[idx] = __float2bfloat16(img[y * sy + x * sx + c * sc + b * sb]);
  }
}
}  // namespace sav

extern "C" int sav_pixel_patchify(const float* images, void* out, int64_t N, int64_t H, int64_t W,
                                  int64_t C, int64_t ph, int64_t pw, int64_t th, int64_t tw, int32_t hwcn,
                                  void* stream) {
  using namespace sav;
  if (N <= 0 || !images || !out || ph <= 0 || pw <= 0 || th <= 0 || tw <= 0 || H % ph || W % pw ||
      ph % th || pw % tw)
    return set_error(SAV_ERR_SHAPE, "sav_pixel_patchify: H,W multiples of the patch, patch of the sub-patch");
  const int64_t total = N * H * W * C;
  int64_t sy, sx, sc, sb;
  if (hwcn) { sb = 1; sc = N; sx = C * N; sy = W * C * N; }
  else { sc = 1; sx = C; sy = W * C; sb = H * W * C; }
  const int64_t blocks = (total + 255) / 256;
  pixel_patchify_kernel<<<(unsigned)(blocks < 148 * 32 ? blocks : 148 * 32), 256, 0,
                          reinterpret_cast<cudaStream_t>(stream)>>>(
      images, reinterpret_cast<bf16*>(out), total, (int)N, (int)H, (int)W, (int)C, (int)ph, (int)pw,
      (int)th, (int)tw, sy, sx, sc, sb);
  count_launch();
  return check_launch("sav_pixel_patchify");
}
