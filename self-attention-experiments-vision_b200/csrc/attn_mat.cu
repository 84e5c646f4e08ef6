// Row kernels of the materialised-logits attention path.  The fused TMEM kernel
// (attn_fwd.cu / attn_bwd.cu) covers plain attention with head dim 64; the variants that couple
// heads or add position logits run as  batched GEMM (S = q k^T, sav_bgemm)  ->  THIS FILE  ->
// batched GEMM (O = P v):
//   * talking heads (attention.py:44-52, talking_heads.py:10-13): H x H mixing of the logits before
//     the softmax and of the probabilities after it -- all H heads of one (image, query) row are
//     handled by one warp, so both mixes, the softmax and (backward) both transform gradients are
//     one pass over the logits;
//   * BoTNet 2-D relative position logits (botnet.py:70-141): the per-query tables q.r^W, q.r^H
//     (one skinny GEMM) are gathered into the logits here, and their gradient is scattered back;
//   * class attention (cait.py:10-15, Nq = 1) and head dims other than 64 need nothing special.
// One warp per row.  Rows of up to 224 keys with at most 8 heads (every CaiT-XXS / XS / S layer) run
// on the tensor cores: both H x H mixes, the backward's dp / dS products and the two transform
// gradients are TF32 mma.sync tiles over 16-key blocks (see "talking heads entirely on the tensor
// cores" below).  Everything else (12 / 16 heads, longer rows, relative logits, independent heads)
// uses the CUDA-core kernels: lanes stride over the keys (coalesced 128-byte accesses per head), fp32
// statistics by warp shuffle, three passes over L1/L2-resident rows (max+sum, delta, outputs).
#include "common.cuh"
#include "host.h"

namespace sav {

constexpr float kLog2e = 1.4426950408889634f;
constexpr int kRowWarps = 4;

struct SoftArgs {
  const float* S;     // [B,H,Nq,ld] raw q.k logits
  bf16* P;            // fwd: [B,H,Nq,ld] probabilities (post-mix)
  const float* dP;    // bwd: [B,H,Nq,ld] gradient of the post-mix probabilities
  bf16* dS;           // bwd: [B,H,Nq,ld] gradient of the raw logits
  const float* mix_pre;   // [H,H] (h -> i) or null
  const float* mix_post;  // [H,H] (i -> j) or null
  float* dmix_pre;
  float* dmix_post;
  const float* rel;   // [B,Nq,H,ldr] or null
  float* drel;
  int B, H, Nq, Nk, ld, ldr, rel_w, rel_h;
  float scale;
};

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// relative-logit column indices of key k for query q on a rel_w x rel_h grid (botnet.py:78-141:
// abs[i,j] = rel[i, j-i+l-1] along each axis)
__device__ __forceinline__ void rel_index(const SoftArgs& a, int q, int k, int& iw, int& ih) {
  const int qx = q % a.rel_w, qy = q / a.rel_w;
  const int kx = k % a.rel_w, ky = k / a.rel_w;
  iw = kx - qx + a.rel_w - 1;
  ih = (2 * a.rel_w - 1) + ky - qy + a.rel_h - 1;
}

// ------------------------------------------------------------------------------ independent heads
// row = (b, h, q)
__global__ void __launch_bounds__(kRowWarps * 32)
softmax_rows_fwd_kernel(const SoftArgs a) {
  const int lane = threadIdx.x & 31;
  const int64_t rows = (int64_t)a.B * a.H * a.Nq;
  const float sc = a.scale * kLog2e;
  for (int64_t row = (int64_t)blockIdx.x * kRowWarps + (threadIdx.x >> 5); row < rows;
       row += (int64_t)gridDim.x * kRowWarps) {
    const int q = (int)(row % a.Nq);
    const int h = (int)((row / a.Nq) % a.H);
    const int64_t b = row / ((int64_t)a.Nq * a.H);
    const float* s = a.S + row * a.ld;
    const float* r = a.rel ? a.rel + ((b * a.Nq + q) * a.H + h) * (int64_t)a.ldr : nullptr;
    auto logit = [&](int k) {
      float t = s[k];
      if (r) { int iw, ih; rel_index(a, q, k, iw, ih); t += r[iw] + r[ih]; }
      return t * sc;
    };
    float m = -INFINITY;
    for (int k = lane; k < a.Nk; k += 32) m = fmaxf(m, logit(k));
    m = warp_max(m);
    float l = 0.f;
    for (int k = lane; k < a.Nk; k += 32) l += ex2_fast(logit(k) - m);
    l = warp_sum(l);
    const float inv = 1.f / l;
    bf16* p = a.P + row * a.ld;
    for (int k = lane; k < a.ld; k += 32)
      p[k] = __float2bfloat16(k < a.Nk ? ex2_fast(logit(k) - m) * inv : 0.f);
  }
}

__global__ void __launch_bounds__(kRowWarps * 32)
softmax_rows_bwd_kernel(const SoftArgs a) {
  extern __shared__ float sm_rel[];  // [kRowWarps][ldr]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* dr_s = sm_rel + warp * a.ldr;
  const int64_t rows = (int64_t)a.B * a.H * a.Nq;
  const float sc = a.scale * kLog2e;
  for (int64_t row = (int64_t)blockIdx.x * kRowWarps + warp; row < rows;
       row += (int64_t)gridDim.x * kRowWarps) {
    const int q = (int)(row % a.Nq);
    const int h = (int)((row / a.Nq) % a.H);
    const int64_t b = row / ((int64_t)a.Nq * a.H);
    const float* s = a.S + row * a.ld;
    const float* dp = a.dP + row * a.ld;
    const int64_t roff = ((b * a.Nq + q) * a.H + h) * (int64_t)a.ldr;
    const float* r = a.rel ? a.rel + roff : nullptr;
    auto logit = [&](int k) {
      float t = s[k];
      if (r) { int iw, ih; rel_index(a, q, k, iw, ih); t += r[iw] + r[ih]; }
      return t * sc;
    };
    float m = -INFINITY;
    for (int k = lane; k < a.Nk; k += 32) m = fmaxf(m, logit(k));
    m = warp_max(m);
    float l = 0.f, dl = 0.f;
    for (int k = lane; k < a.Nk; k += 32) {
      const float e = ex2_fast(logit(k) - m);
      l += e;
      dl += e * dp[k];
    }
    l = warp_sum(l);
    dl = warp_sum(dl);
    const float inv = 1.f / l;
    const float delta = dl * inv;
    if (r) {
      for (int j = lane; j < a.ldr; j += 32) dr_s[j] = 0.f;
      __syncwarp();
    }
    bf16* ds = a.dS + row * a.ld;
    for (int k = lane; k < a.ld; k += 32) {
      float g = 0.f;
      if (k < a.Nk) {
        const float pk = ex2_fast(logit(k) - m) * inv;
        g = pk * (dp[k] - delta) * a.scale;
        if (r) {
          int iw, ih;
          rel_index(a, q, k, iw, ih);
          atomicAdd(&dr_s[iw], g);
          atomicAdd(&dr_s[ih], g);
        }
      }
      ds[k] = __float2bfloat16(g);
    }
    if (r) {
      __syncwarp();
      for (int j = lane; j < a.ldr; j += 32) a.drel[roff + j] = dr_s[j];
      __syncwarp();
    }
  }
}

// ------------------------------------------------------------------------------ talking heads
// row = (b, q); one warp owns all H heads of the row.
template <int H>
__device__ __forceinline__ void th_logits(const SoftArgs& a, const float* s0, int64_t hs, int k,
                                          const float* t1, float sc, float (&s)[H], float (&t)[H]) {
#pragma unroll
  for (int h = 0; h < H; ++h) s[h] = s0[h * hs + k];
#pragma unroll
  for (int i = 0; i < H; ++i) {
    float acc = 0.f;
#pragma unroll
    for (int h = 0; h < H; ++h) acc = fmaf(t1[h * H + i], s[h], acc);
    t[i] = acc * sc;
  }
}

template <int H>
__device__ __forceinline__ void th_stats(const SoftArgs& a, const float* s0, int64_t hs, int lane,
                                         const float* t1, float sc, float (&m)[H], float (&inv)[H]) {
  float s[H], t[H];
#pragma unroll
  for (int i = 0; i < H; ++i) m[i] = -INFINITY;
  for (int k = lane; k < a.Nk; k += 32) {
    th_logits<H>(a, s0, hs, k, t1, sc, s, t);
#pragma unroll
    for (int i = 0; i < H; ++i) m[i] = fmaxf(m[i], t[i]);
  }
#pragma unroll
  for (int i = 0; i < H; ++i) { m[i] = warp_max(m[i]); inv[i] = 0.f; }
  for (int k = lane; k < a.Nk; k += 32) {
    th_logits<H>(a, s0, hs, k, t1, sc, s, t);
#pragma unroll
    for (int i = 0; i < H; ++i) inv[i] += ex2_fast(t[i] - m[i]);
  }
#pragma unroll
  for (int i = 0; i < H; ++i) inv[i] = 1.f / warp_sum(inv[i]);
}

template <int H>
__global__ void __launch_bounds__(kRowWarps * 32)
thsoftmax_fwd_kernel(const SoftArgs a) {
  __shared__ float t1[H * H], t2[H * H];
  for (int i = threadIdx.x; i < H * H; i += blockDim.x) {
    t1[i] = a.mix_pre[i];
    t2[i] = a.mix_post[i];
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int64_t rows = (int64_t)a.B * a.Nq;
  const int64_t hs = (int64_t)a.Nq * a.ld;  // head stride
  const float sc = a.scale * kLog2e;
  for (int64_t row = (int64_t)blockIdx.x * kRowWarps + (threadIdx.x >> 5); row < rows;
       row += (int64_t)gridDim.x * kRowWarps) {
    const int q = (int)(row % a.Nq);
    const int64_t b = row / a.Nq;
    const int64_t off = (b * a.H * a.Nq + q) * (int64_t)a.ld;
    const float* s0 = a.S + off;
    bf16* p0 = a.P + off;
    float m[H], inv[H];
    th_stats<H>(a, s0, hs, lane, t1, sc, m, inv);
    for (int k = lane; k < a.ld; k += 32) {
      float s[H], t[H], pr[H];
      if (k < a.Nk) {
        th_logits<H>(a, s0, hs, k, t1, sc, s, t);
#pragma unroll
        for (int i = 0; i < H; ++i) pr[i] = ex2_fast(t[i] - m[i]) * inv[i];
      } else {
#pragma unroll
        for (int i = 0; i < H; ++i) pr[i] = 0.f;
      }
#pragma unroll
      for (int j = 0; j < H; ++j) {
        float acc = 0.f;
#pragma unroll
        for (int i = 0; i < H; ++i) acc = fmaf(t2[i * H + j], pr[i], acc);
        p0[j * hs + k] = __float2bfloat16(acc);
      }
    }
  }
}

template <int H>
__global__ void __launch_bounds__(kRowWarps * 32)
thsoftmax_bwd_kernel(const SoftArgs a) {
  __shared__ float t1[H * H], t2[H * H];
  __shared__ float red1[H * H], red2[H * H];
  for (int i = threadIdx.x; i < H * H; i += blockDim.x) {
    t1[i] = a.mix_pre[i];
    t2[i] = a.mix_post[i];
    red1[i] = 0.f;
    red2[i] = 0.f;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int64_t rows = (int64_t)a.B * a.Nq;
  const int64_t hs = (int64_t)a.Nq * a.ld;
  const float sc = a.scale * kLog2e;
  float g1[H * H], g2[H * H];  // lane-local partial sums of d mix_pre[h,i], d mix_post[i,j]
#pragma unroll
  for (int i = 0; i < H * H; ++i) { g1[i] = 0.f; g2[i] = 0.f; }
  for (int64_t row = (int64_t)blockIdx.x * kRowWarps + (threadIdx.x >> 5); row < rows;
       row += (int64_t)gridDim.x * kRowWarps) {
    const int q = (int)(row % a.Nq);
    const int64_t b = row / a.Nq;
    const int64_t off = (b * a.H * a.Nq + q) * (int64_t)a.ld;
    const float* s0 = a.S + off;
    const float* d0 = a.dP + off;
    bf16* ds0 = a.dS + off;
    float m[H], inv[H], delta[H];
    th_stats<H>(a, s0, hs, lane, t1, sc, m, inv);
#pragma unroll
    for (int i = 0; i < H; ++i) delta[i] = 0.f;
    for (int k = lane; k < a.Nk; k += 32) {
      float s[H], t[H], dpo[H];
      th_logits<H>(a, s0, hs, k, t1, sc, s, t);
#pragma unroll
      for (int j = 0; j < H; ++j) dpo[j] = d0[j * hs + k];
#pragma unroll
      for (int i = 0; i < H; ++i) {
        const float pi = ex2_fast(t[i] - m[i]) * inv[i];
        float dpi = 0.f;
#pragma unroll
        for (int j = 0; j < H; ++j) {
          dpi = fmaf(t2[i * H + j], dpo[j], dpi);
          g2[i * H + j] = fmaf(pi, dpo[j], g2[i * H + j]);
        }
        delta[i] = fmaf(pi, dpi, delta[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < H; ++i) delta[i] = warp_sum(delta[i]);
    for (int k = lane; k < a.ld; k += 32) {
      float dt[H], s[H], t[H];
      if (k < a.Nk) {
        float dpo[H];
        th_logits<H>(a, s0, hs, k, t1, sc, s, t);
#pragma unroll
        for (int j = 0; j < H; ++j) dpo[j] = d0[j * hs + k];
#pragma unroll
        for (int i = 0; i < H; ++i) {
          const float pi = ex2_fast(t[i] - m[i]) * inv[i];
          float dpi = 0.f;
#pragma unroll
          for (int j = 0; j < H; ++j) dpi = fmaf(t2[i * H + j], dpo[j], dpi);
          dt[i] = pi * (dpi - delta[i]) * a.scale;   // d(logit_i) * scale
        }
      } else {
#pragma unroll
        for (int i = 0; i < H; ++i) { dt[i] = 0.f; s[i] = 0.f; }
      }
#pragma unroll
      for (int h = 0; h < H; ++h) {
        float acc = 0.f;
#pragma unroll
        for (int i = 0; i < H; ++i) {
          acc = fmaf(t1[h * H + i], dt[i], acc);
          g1[h * H + i] = fmaf(s[h], dt[i], g1[h * H + i]);
        }
        ds0[h * hs + k] = __float2bfloat16(acc);
      }
    }
  }
  // lane -> warp -> block -> global
#pragma unroll
  for (int i = 0; i < H * H; ++i) {
    const float v1 = warp_sum(g1[i]);
    const float v2 = warp_sum(g2[i]);
    if (lane == 0) {
      atomicAdd(&red1[i], v1);
      atomicAdd(&red2[i], v2);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < H * H; i += blockDim.x) {
    atomicAdd(a.dmix_pre + i, red1[i]);
    atomicAdd(a.dmix_post + i, red2[i]);
  }
}

// ---- many heads (H = 12, 16: CaiT-M).  The H x H gradient accumulators of the kernel above do not
// fit in registers (2 * 256 floats per lane), so the backward runs H / HS times, each launch
// accumulating the HS columns [I0, I0 + HS) of d mix_pre and the matching rows of d mix_post;
// the launch with I0 == 0 also writes dS.  Everything else is recomputed per launch -- this path
// is about coverage of the 16-head constructors, not speed.
template <int H, int HS, int I0>
__global__ void __launch_bounds__(kRowWarps * 32)
thsoftmax_bwd_slice_kernel(const SoftArgs a) {
  __shared__ float t1[H * H], t2[H * H];
  __shared__ float red1[H * HS], red2[HS * H];
  for (int i = threadIdx.x; i < H * H; i += blockDim.x) {
    t1[i] = a.mix_pre[i];
    t2[i] = a.mix_post[i];
  }
  for (int i = threadIdx.x; i < H * HS; i += blockDim.x) { red1[i] = 0.f; red2[i] = 0.f; }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int64_t rows = (int64_t)a.B * a.Nq;
  const int64_t hs = (int64_t)a.Nq * a.ld;
  const float sc = a.scale * kLog2e;
  float g1[H * HS], g2[HS * H];   // d mix_pre[h, I0+ii], d mix_post[I0+ii, j]
#pragma unroll
  for (int i = 0; i < H * HS; ++i) { g1[i] = 0.f; g2[i] = 0.f; }
  for (int64_t row = (int64_t)blockIdx.x * kRowWarps + (threadIdx.x >> 5); row < rows;
       row += (int64_t)gridDim.x * kRowWarps) {
    const int q = (int)(row % a.Nq);
    const int64_t b = row / a.Nq;
    const int64_t off = (b * a.H * a.Nq + q) * (int64_t)a.ld;
    const float* s0 = a.S + off;
    const float* d0 = a.dP + off;
    bf16* ds0 = a.dS + off;
    float m[H], inv[H], delta[H];
    th_stats<H>(a, s0, hs, lane, t1, sc, m, inv);
#pragma unroll
    for (int i = 0; i < H; ++i) delta[i] = 0.f;
    for (int k = lane; k < a.Nk; k += 32) {
      float s[H], t[H], dpo[H];
      th_logits<H>(a, s0, hs, k, t1, sc, s, t);
#pragma unroll
      for (int j = 0; j < H; ++j) dpo[j] = d0[j * hs + k];
#pragma unroll
      for (int i = 0; i < H; ++i) {
        const float pi = ex2_fast(t[i] - m[i]) * inv[i];
        float dpi = 0.f;
#pragma unroll
        for (int j = 0; j < H; ++j) {
          dpi = fmaf(t2[i * H + j], dpo[j], dpi);
          if (i >= I0 && i < I0 + HS) g2[(i - I0) * H + j] = fmaf(pi, dpo[j], g2[(i - I0) * H + j]);
        }
        delta[i] = fmaf(pi, dpi, delta[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < H; ++i) delta[i] = warp_sum(delta[i]);
    for (int k = lane; k < (I0 == 0 ? a.ld : a.Nk); k += 32) {
      float dt[H], s[H], t[H];
      if (k < a.Nk) {
        float dpo[H];
        th_logits<H>(a, s0, hs, k, t1, sc, s, t);
#pragma unroll
        for (int j = 0; j < H; ++j) dpo[j] = d0[j * hs + k];
#pragma unroll
        for (int i = 0; i < H; ++i) {
          const float pi = ex2_fast(t[i] - m[i]) * inv[i];
          float dpi = 0.f;
#pragma unroll
          for (int j = 0; j < H; ++j) dpi = fmaf(t2[i * H + j], dpo[j], dpi);
          dt[i] = pi * (dpi - delta[i]) * a.scale;
        }
      } else {
#pragma unroll
        for (int i = 0; i < H; ++i) { dt[i] = 0.f; s[i] = 0.f; }
      }
#pragma unroll
      for (int h = 0; h < H; ++h) {
        float acc = 0.f;
#pragma unroll
        for (int i = 0; i < H; ++i) {
          acc = fmaf(t1[h * H + i], dt[i], acc);
          if (i >= I0 && i < I0 + HS) g1[h * HS + (i - I0)] = fmaf(s[h], dt[i], g1[h * HS + (i - I0)]);
        }
        if (I0 == 0) ds0[h * hs + k] = __float2bfloat16(acc);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < H * HS; ++i) {
    const float v1 = warp_sum(g1[i]);
    const float v2 = warp_sum(g2[i]);
    if (lane == 0) {
      atomicAdd(&red1[i], v1);
      atomicAdd(&red2[i], v2);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < H * HS; i += blockDim.x) {
    atomicAdd(a.dmix_pre + (i / HS) * H + I0 + (i % HS), red1[i]);     // [h][I0 + ii]
    atomicAdd(a.dmix_post + (I0 + i / H) * H + (i % H), red2[i]);      // [I0 + ii][j]
  }
}
template <int H, int HS, int I0>
static void launch_th_bwd_slices(const SoftArgs& a, int grid, cudaStream_t st) {
  thsoftmax_bwd_slice_kernel<H, HS, I0><<<grid, kRowWarps * 32, 0, st>>>(a);
  count_launch();
  if constexpr (I0 + HS < H) launch_th_bwd_slices<H, HS, I0 + HS>(a, grid, st);
}

// ---- rows of up to 32 * kKPL = 224 keys (CaiT: 196 / 197) have dedicated kernels below
constexpr int kKPL = 7;

// 4-byte asynchronous global -> shared copy: a row's dP is requested up front without holding a
// register per element
__device__ __forceinline__ void cp_async4(float* dst, const float* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.wait_all;" ::: "memory");
}

// ---- talking heads entirely on the tensor cores (H <= 8, rows of <= 224 keys).
// Both head mixes are [keys x 8] x [8 x 8] products, so a warp walks its row in blocks of 16 keys
// with mma.sync.m16n8k8 (TF32): M = 16 keys, K = N = heads.  Lane (g, t) = (lane / 4, lane % 4) loads
// the A fragment of the raw logits straight from global memory (heads t, t + 4 at keys g, g + 8: each
// group of eight lanes reads one 32-byte sector per head) -- no shared-memory staging -- and the
// accumulator fragment it gets back (keys g, g + 8 x heads 2t, 2t + 1) IS the A fragment of the next
// product once the contraction index is renumbered (column t <-> head 2t, column t + 4 <-> head 2t + 1)
// and the constant B fragment is permuted the same way.  The softmax statistics of a head are a
// register reduction plus three shuffles over g.  The pre-softmax mix uses the error-compensated
// 3 x TF32 split (hi * hi + lo * hi + hi * lo): logits keep fp32 accuracy; the post-softmax mix of
// probabilities in [0, 1] is rounded to bf16 right after, one TF32 product is enough.
__device__ __forceinline__ uint32_t tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void mma_tf32_m16n8k8(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                                 uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
      "{%0, %1, %2, %3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// c += A * B with A (fp32 values) split into TF32 hi + lo and B given as (hi, lo)
__device__ __forceinline__ void mma_3xtf32(float (&c)[4], const float (&av)[4], uint32_t bh0, uint32_t bh1,
                                           uint32_t bl0, uint32_t bl1) {
  uint32_t ah[4], al[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    ah[e] = tf32_rna(av[e]);
    al[e] = tf32_rna(av[e] - __uint_as_float(ah[e]));
  }
  mma_tf32_m16n8k8(c, al[0], al[1], al[2], al[3], bh0, bh1);
  mma_tf32_m16n8k8(c, ah[0], ah[1], ah[2], ah[3], bl0, bl1);
  mma_tf32_m16n8k8(c, ah[0], ah[1], ah[2], ah[3], bh0, bh1);
}

constexpr int kTcBlocks = 2 * kKPL;    // 16-key blocks of a row (<= 224 keys)

// mixed, scaled logits of the whole row as accumulator fragments: tm[blk] = (keys g, g + 8) x (heads
// 2t, 2t + 1) of block blk; -inf at keys >= Nk
template <int H>
__device__ __forceinline__ void th_tc_logits(const SoftArgs& a, const float* s0, int64_t hs, int g, int t,
                                             uint32_t bh0, uint32_t bh1, uint32_t bl0, uint32_t bl1,
                                             float (&tm)[kTcBlocks][4]) {
  // all loads of the row first (one DRAM round trip per row, not one per block), parked in the
  // registers that receive the products
  const float* sa = s0 + (t < H ? t : 0) * hs + g;            // head t,     key g of block 0
  const float* sb = s0 + (t + 4 < H ? t + 4 : 0) * hs + g;    // head t + 4
  const int nk_g = a.Nk - g;                                   // key 16 blk + g (+ 8) < Nk  <=>  16 blk (+ 8) < nk_g
#pragma unroll
  for (int blk = 0; blk < kTcBlocks; ++blk) {
    tm[blk][0] = (t < H && 16 * blk < nk_g) ? __ldg(sa + 16 * blk) : 0.f;
    tm[blk][1] = (t < H && 16 * blk + 8 < nk_g) ? __ldg(sa + 16 * blk + 8) : 0.f;
    tm[blk][2] = (t + 4 < H && 16 * blk < nk_g) ? __ldg(sb + 16 * blk) : 0.f;
    tm[blk][3] = (t + 4 < H && 16 * blk + 8 < nk_g) ? __ldg(sb + 16 * blk + 8) : 0.f;
  }
#pragma unroll
  for (int blk = 0; blk < kTcBlocks; ++blk) {
    if (16 * blk < a.Nk) {                         // warp-uniform
      const float av[4] = {tm[blk][0], tm[blk][1], tm[blk][2], tm[blk][3]};
      float c[4] = {0.f, 0.f, 0.f, 0.f};
      mma_3xtf32(c, av, bh0, bh1, bl0, bl1);
      tm[blk][0] = (16 * blk < nk_g) ? c[0] : -INFINITY;
      tm[blk][1] = (16 * blk < nk_g) ? c[1] : -INFINITY;
      tm[blk][2] = (16 * blk + 8 < nk_g) ? c[2] : -INFINITY;
      tm[blk][3] = (16 * blk + 8 < nk_g) ? c[3] : -INFINITY;
    } else {
      tm[blk][0] = tm[blk][1] = tm[blk][2] = tm[blk][3] = -INFINITY;
    }
  }
}

// max / sum over the keys of this lane's two heads (all lanes of equal t agree afterwards);
// tm becomes exp2(tm - max)
__device__ __forceinline__ void th_tc_softmax(float (&tm)[kTcBlocks][4], float (&inv)[2]) {
  float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
  for (int blk = 0; blk < kTcBlocks; ++blk) {
    m0 = fmaxf(m0, fmaxf(tm[blk][0], tm[blk][2]));
    m1 = fmaxf(m1, fmaxf(tm[blk][1], tm[blk][3]));
  }
#pragma unroll
  for (int o = 4; o < 32; o <<= 1) {
    m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, o));
    m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, o));
  }
  float s0 = 0.f, s1 = 0.f;
#pragma unroll
  for (int blk = 0; blk < kTcBlocks; ++blk) {
    tm[blk][0] = ex2_fast(tm[blk][0] - m0);   // exp2(-inf) = 0 for the padded keys
    tm[blk][2] = ex2_fast(tm[blk][2] - m0);
    tm[blk][1] = ex2_fast(tm[blk][1] - m1);
    tm[blk][3] = ex2_fast(tm[blk][3] - m1);
    s0 += tm[blk][0] + tm[blk][2];
    s1 += tm[blk][1] + tm[blk][3];
  }
#pragma unroll
  for (int o = 4; o < 32; o <<= 1) {
    s0 += __shfl_xor_sync(0xffffffffu, s0, o);
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
  }
  inv[0] = 1.f / s0;
  inv[1] = 1.f / s1;
}

template <int H>
__global__ void __launch_bounds__(kRowWarps * 32)
thsoftmax_fwd_tc_kernel(const SoftArgs a) {
  static_assert(H <= 8, "one m16n8k8 tile holds the H x H transform");
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int64_t rows = (int64_t)a.B * a.Nq;
  const int64_t hs = (int64_t)a.Nq * a.ld;
  const float sc = a.scale * kLog2e;
  // B of the pre-softmax mix: [k = h][n = i] = T1[h][i] * scale * log2(e), lane holds (h = t, t + 4; i = g)
  const float b1a = (t < H && g < H) ? a.mix_pre[t * H + g] * sc : 0.f;
  const float b1b = (t + 4 < H && g < H) ? a.mix_pre[(t + 4) * H + g] * sc : 0.f;
  const uint32_t bh0 = tf32_rna(b1a), bh1 = tf32_rna(b1b);
  const uint32_t bl0 = tf32_rna(b1a - __uint_as_float(bh0)), bl1 = tf32_rna(b1b - __uint_as_float(bh1));
  // B of the post-softmax mix with the contraction index renumbered (k = t <-> i = 2t, k = t + 4 <-> 2t + 1):
  // [i][n = j] = T2[i][j], lane holds (i = 2t, 2t + 1; j = g)
  const uint32_t b2a = tf32_rna((2 * t < H && g < H) ? a.mix_post[(2 * t) * H + g] : 0.f);
  const uint32_t b2b = tf32_rna((2 * t + 1 < H && g < H) ? a.mix_post[(2 * t + 1) * H + g] : 0.f);
  for (int64_t row = (int64_t)blockIdx.x * kRowWarps + (threadIdx.x >> 5); row < rows;
       row += (int64_t)gridDim.x * kRowWarps) {
    const int q = (int)(row % a.Nq);
    const int64_t b = row / a.Nq;
    const int64_t off = (b * a.H * a.Nq + q) * (int64_t)a.ld;
    const float* s0 = a.S + off;
    bf16* p0 = a.P + off;
    bf16* pa = p0 + (2 * t < H ? 2 * t : 0) * hs + g;          // output head 2t,     key g of block 0
    bf16* pb = p0 + (2 * t + 1 < H ? 2 * t + 1 : 0) * hs + g;  // output head 2t + 1
    const int ld_g = a.ld - g;
    float tm[kTcBlocks][4], inv[2];
    th_tc_logits<H>(a, s0, hs, g, t, bh0, bh1, bl0, bl1, tm);
    th_tc_softmax(tm, inv);
#pragma unroll
    for (int blk = 0; blk < kTcBlocks; ++blk) {
      if (16 * blk < a.ld) {                       // warp-uniform
        float c[4] = {0.f, 0.f, 0.f, 0.f};
        // A = probabilities: (key g, col t <-> head 2t), (key g + 8, col t), (key g, col t + 4 <-> 2t + 1), ...
        mma_tf32_m16n8k8(c, tf32_rna(tm[blk][0] * inv[0]), tf32_rna(tm[blk][2] * inv[0]),
                         tf32_rna(tm[blk][1] * inv[1]), tf32_rna(tm[blk][3] * inv[1]), b2a, b2b);
        // c = mixed probabilities: keys g, g + 8 x output heads 2t, 2t + 1
        if (2 * t < H) {
          if (16 * blk < ld_g) pa[16 * blk] = __float2bfloat16(c[0]);
          if (16 * blk + 8 < ld_g) pa[16 * blk + 8] = __float2bfloat16(c[2]);
        }
        if (2 * t + 1 < H) {
          if (16 * blk < ld_g) pb[16 * blk] = __float2bfloat16(c[1]);
          if (16 * blk + 8 < ld_g) pb[16 * blk + 8] = __float2bfloat16(c[3]);
        }
      }
    }
  }
}

// Backward of the same design.  Per row: raw dP goes straight to shared memory with cp.async (issued
// first, it lands while the logits are recomputed); the probabilities stay in accumulator-fragment
// registers.  The two transform gradients contract over the KEYS, which sit on the wrong fragment
// axis (rows), so p -- and later dt -- take one trip through shared memory as [head][key] rows of
// kTcRow = 228 floats (228 = 4 mod 32: conflict-free both for the fragment-layout stores, bank
// 8t + g, and for the [head g][key 8m + t] operand loads, bank 4g + t).  The softmax correction
// needs no second pass: delta_i = sum_k p_i[k] dp_i[k] = sum_j T2[i][j] (sum_k p_i[k] dP_j[k]) is
// read off the row's own d mix_post tile.  dp = T2 dP and dS = T1 dt are one MMA each per 16 keys,
// chained through registers like the forward.
constexpr int kTcRow = 228;

template <int H>
__global__ void __launch_bounds__(kRowWarps * 32, 3)
thsoftmax_bwd_tc_kernel(const SoftArgs a) {
  static_assert(H <= 8, "one m16n8k8 tile holds the H x H transform");
  extern __shared__ float park[];   // [kRowWarps][2][8][kTcRow]: p / dt rows, raw dP rows
  __shared__ float red1[64], red2[64];
  for (int i = threadIdx.x; i < 64; i += blockDim.x) { red1[i] = 0.f; red2[i] = 0.f; }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  float* bufP = park + (size_t)warp * 2 * 8 * kTcRow;
  float* bufD = bufP + 8 * kTcRow;
  const int64_t rows = (int64_t)a.B * a.Nq;
  const int64_t hs = (int64_t)a.Nq * a.ld;
  const float sc = a.scale * kLog2e;
  auto T1 = [&](int h, int i) { return (h < H && i < H) ? a.mix_pre[h * H + i] : 0.f; };
  auto T2 = [&](int i, int j) { return (i < H && j < H) ? a.mix_post[i * H + j] : 0.f; };
  // pre-softmax mix, 3 x TF32: B[k = h][n = i] = T1[h][i] * scale * log2(e); lane: h = t, t + 4; i = g
  const float b1a = T1(t, g) * sc, b1b = T1(t + 4, g) * sc;
  const uint32_t bh0 = tf32_rna(b1a), bh1 = tf32_rna(b1b);
  const uint32_t bl0 = tf32_rna(b1a - __uint_as_float(bh0)), bl1 = tf32_rna(b1b - __uint_as_float(bh1));
  // dp^T[keys x i] = dP^T[keys x j] T2^T: B[k = j][n = i] = T2[i][j]; lane: j = t, t + 4; i = g
  const uint32_t bd0 = tf32_rna(T2(g, t)), bd1 = tf32_rna(T2(g, t + 4));
  // dS^T[keys x h] = dt[keys x i] T1^T with i renumbered (k = t <-> 2t, k = t + 4 <-> 2t + 1):
  // B[i][n = h] = T1[h][i]; lane: i = 2t, 2t + 1; h = g
  const uint32_t bs0 = tf32_rna(T1(g, 2 * t)), bs1 = tf32_rna(T1(g, 2 * t + 1));
  // delta from the row's d mix_post tile: lane holds [i = g][j = 2t, 2t + 1]
  const float w0 = T2(g, 2 * t), w1 = T2(g, 2 * t + 1);
  float c1[2][4], c2[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) { c1[0][e] = 0.f; c1[1][e] = 0.f; c2[e] = 0.f; }
  for (int64_t row = (int64_t)blockIdx.x * kRowWarps + warp; row < rows;
       row += (int64_t)gridDim.x * kRowWarps) {
    const int q = (int)(row % a.Nq);
    const int64_t b = row / a.Nq;
    const int64_t off = (b * a.H * a.Nq + q) * (int64_t)a.ld;
    const float* s0 = a.S + off;
    const float* d0 = a.dP + off;
    bf16* ds0 = a.dS + off;
    bf16* dsa = ds0 + (2 * t < H ? 2 * t : 0) * hs + g;
    bf16* dsb = ds0 + (2 * t + 1 < H ? 2 * t + 1 : 0) * hs + g;
    const int ld_g = a.ld - g, nk_g = a.Nk - g;
    const float* da = d0 + (t < H ? t : 0) * hs + g;            // dP head t,     key g of block 0
    const float* db = d0 + (t + 4 < H ? t + 4 : 0) * hs + g;    // dP head t + 4
    float* bda = bufD + t * kTcRow + g;
    float* bdb = bufD + (t + 4) * kTcRow + g;
    // (a) raw dP -> bufD[j][key], zeros beyond Nk (the MMAs read whole blocks); lane: j = t, t + 4; keys g, g + 8
#pragma unroll
    for (int blk = 0; blk < kTcBlocks; ++blk) {
      if (16 * blk < a.Nk) {
        if (t < H && 16 * blk < nk_g) cp_async4(bda + 16 * blk, da + 16 * blk);
        else bda[16 * blk] = 0.f;
        if (t < H && 16 * blk + 8 < nk_g) cp_async4(bda + 16 * blk + 8, da + 16 * blk + 8);
        else bda[16 * blk + 8] = 0.f;
        if (t + 4 < H && 16 * blk < nk_g) cp_async4(bdb + 16 * blk, db + 16 * blk);
        else bdb[16 * blk] = 0.f;
        if (t + 4 < H && 16 * blk + 8 < nk_g) cp_async4(bdb + 16 * blk + 8, db + 16 * blk + 8);
        else bdb[16 * blk + 8] = 0.f;
      }
    }
    // (b) probabilities in fragment registers: tm[blk] = keys (g, g + 8) x heads (2t, 2t + 1)
    float tm[kTcBlocks][4], inv[2];
    th_tc_logits<H>(a, s0, hs, g, t, bh0, bh1, bl0, bl1, tm);
    th_tc_softmax(tm, inv);
    // (c) p -> bufP[i][key]
#pragma unroll
    for (int blk = 0; blk < kTcBlocks; ++blk) {
      if (16 * blk < a.Nk) {
        const int k0 = 16 * blk + g;
        tm[blk][0] *= inv[0]; tm[blk][2] *= inv[0];
        tm[blk][1] *= inv[1]; tm[blk][3] *= inv[1];
        bufP[(2 * t) * kTcRow + k0] = tm[blk][0];
        bufP[(2 * t + 1) * kTcRow + k0] = tm[blk][1];
        bufP[(2 * t) * kTcRow + k0 + 8] = tm[blk][2];
        bufP[(2 * t + 1) * kTcRow + k0 + 8] = tm[blk][3];
      }
    }
    cp_async_wait_all();
    __syncwarp();
    // (e) this row's d mix_post tile: r2[i][j] = sum_k p_i[k] dP_j[k]; A = p [i = g][keys t, t + 4],
    //     B = dP [keys t, t + 4][j = g]; two chains
    float r2[2][4];
#pragma unroll
    for (int e = 0; e < 4; ++e) { r2[0][e] = 0.f; r2[1][e] = 0.f; }
#pragma unroll
    for (int mb = 0; mb < 2 * kTcBlocks; ++mb) {
      if (8 * mb < a.Nk) {                          // warp-uniform
        const int o = g * kTcRow + 8 * mb + t;
        mma_tf32_m16n8k8(r2[mb & 1], __float_as_uint(bufP[o]), 0u, __float_as_uint(bufP[o + 4]), 0u,
                         __float_as_uint(bufD[o]), __float_as_uint(bufD[o + 4]));
      }
    }
    const float g2a = r2[0][0] + r2[1][0], g2b = r2[0][1] + r2[1][1];
    c2[0] += g2a;
    c2[1] += g2b;
    // (f) delta_i for i = g on every lane of group g, then for this lane's heads 2t, 2t + 1
    float dl = w0 * g2a + w1 * g2b;
    dl += __shfl_xor_sync(0xffffffffu, dl, 1);
    dl += __shfl_xor_sync(0xffffffffu, dl, 2);
    const float dl0 = __shfl_sync(0xffffffffu, dl, 8 * t), dl1 = __shfl_sync(0xffffffffu, dl, 8 * t + 4);
    // (g) per 16 keys: dp = T2 dP, dt = p (dp - delta) scale, dS = T1 dt
#pragma unroll
    for (int blk = 0; blk < kTcBlocks; ++blk) {
      if (16 * blk < a.ld) {                        // warp-uniform
        float dp[4] = {0.f, 0.f, 0.f, 0.f};
        mma_tf32_m16n8k8(dp, __float_as_uint(bda[16 * blk]), __float_as_uint(bda[16 * blk + 8]),
                         __float_as_uint(bdb[16 * blk]), __float_as_uint(bdb[16 * blk + 8]), bd0, bd1);
        // dp: keys (g, g + 8) x heads (2t, 2t + 1), like tm
        tm[blk][0] = tm[blk][0] * (dp[0] - dl0) * a.scale;
        tm[blk][1] = tm[blk][1] * (dp[1] - dl1) * a.scale;
        tm[blk][2] = tm[blk][2] * (dp[2] - dl0) * a.scale;
        tm[blk][3] = tm[blk][3] * (dp[3] - dl1) * a.scale;
        float ds[4] = {0.f, 0.f, 0.f, 0.f};
        mma_tf32_m16n8k8(ds, tf32_rna(tm[blk][0]), tf32_rna(tm[blk][2]), tf32_rna(tm[blk][1]),
                         tf32_rna(tm[blk][3]), bs0, bs1);
        // ds: keys (g, g + 8) x raw heads (2t, 2t + 1)
        if (2 * t < H) {
          if (16 * blk < ld_g) dsa[16 * blk] = __float2bfloat16(ds[0]);
          if (16 * blk + 8 < ld_g) dsa[16 * blk + 8] = __float2bfloat16(ds[2]);
        }
        if (2 * t + 1 < H) {
          if (16 * blk < ld_g) dsb[16 * blk] = __float2bfloat16(ds[1]);
          if (16 * blk + 8 < ld_g) dsb[16 * blk + 8] = __float2bfloat16(ds[3]);
        }
      }
    }
    __syncwarp();        // every lane has read p from bufP
    // (h) dt -> bufP[i][key]
#pragma unroll
    for (int blk = 0; blk < kTcBlocks; ++blk) {
      if (16 * blk < a.Nk) {
        const int k0 = 16 * blk + g;
        bufP[(2 * t) * kTcRow + k0] = tm[blk][0];
        bufP[(2 * t + 1) * kTcRow + k0] = tm[blk][1];
        bufP[(2 * t) * kTcRow + k0 + 8] = tm[blk][2];
        bufP[(2 * t + 1) * kTcRow + k0 + 8] = tm[blk][3];
      }
    }
    __syncwarp();
    // (i) d mix_pre[h][i] += sum_k s_h[k] dt_i[k]: A = raw S [h = g][keys t, t + 4] (L2 resident, the
    //     loads of 32 keys issued together), B = dt [keys t, t + 4][i = g]
#pragma unroll
    for (int kk = 0; kk < kKPL; ++kk) {
      if (32 * kk < a.Nk) {                         // warp-uniform
        const float* sg = s0 + (g < H ? g : 0) * hs + 32 * kk + t;
        float a0[4], a2[4];
#pragma unroll
        for (int mb = 0; mb < 4; ++mb) {
          const int k0 = 32 * kk + 8 * mb + t;
          a0[mb] = (g < H && k0 < a.Nk) ? __ldg(sg + 8 * mb) : 0.f;
          a2[mb] = (g < H && k0 + 4 < a.Nk) ? __ldg(sg + 8 * mb + 4) : 0.f;
        }
#pragma unroll
        for (int mb = 0; mb < 4; ++mb) {
          if (32 * kk + 8 * mb < a.Nk) {            // warp-uniform
            const int o = g * kTcRow + 32 * kk + 8 * mb + t;
            mma_tf32_m16n8k8(c1[mb & 1], __float_as_uint(a0[mb]), 0u, __float_as_uint(a2[mb]), 0u,
                             __float_as_uint(bufP[o]), __float_as_uint(bufP[o + 4]));
          }
        }
      }
    }
    __syncwarp();   // the next row overwrites bufP / bufD
  }
  // lane (g, t) holds entries [g][2t], [g][2t + 1] of both gradients
  if (g < H) {
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int col = 2 * t + e;
      if (col < H) {
        atomicAdd(&red1[g * H + col], c1[0][e] + c1[1][e]);
        atomicAdd(&red2[g * H + col], c2[e]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < H * H; i += blockDim.x) {
    atomicAdd(a.dmix_pre + i, red1[i]);
    atomicAdd(a.dmix_post + i, red2[i]);
  }
}

template <int H>
static int launch_th_bwd_tc(const SoftArgs& a, cudaStream_t st) {
  const size_t sm = (size_t)kRowWarps * 2 * 8 * kTcRow * sizeof(float);
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(thsoftmax_bwd_tc_kernel<H>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    if (e != cudaSuccess) return set_error(SAV_ERR_CUDA, "thsoftmax smem attr: %s", cudaGetErrorString(e));
    attr_set = true;
  }
  // three resident blocks per SM (3 x 58 KB of shared memory); every block ends with 2 H^2 atomics
  const int64_t blocks = ((int64_t)a.B * a.Nq + kRowWarps - 1) / kRowWarps;
  const int grid = (int)(blocks < 148 * 3 ? blocks : 148 * 3);
  thsoftmax_bwd_tc_kernel<H><<<grid, kRowWarps * 32, sm, st>>>(a);
  return SAV_OK;
}

static int soft_grid(int64_t rows) {
  int64_t blocks = (rows + kRowWarps - 1) / kRowWarps;
  const int64_t cap = 148 * 8;
  return (int)(blocks < cap ? blocks : cap);
}

static int fill_args(const sav_thsoftmax_args* x, SoftArgs& a, const char* who) {
  if (!x || x->B <= 0 || x->H <= 0 || x->Nq <= 0 || x->Nk <= 0 || x->ld < x->Nk)
    return set_error(SAV_ERR_SHAPE, "%s: bad shape", who);
  if ((x->mix_pre == nullptr) != (x->mix_post == nullptr))
    return set_error(SAV_ERR_ARG, "%s: mix_pre and mix_post come together", who);
  if (x->mix_pre && x->rel)
    return set_error(SAV_ERR_ARG, "%s: talking heads and relative logits are not combined", who);
  if (x->rel && ((int64_t)x->rel_w * x->rel_h != x->Nk || x->Nq != x->Nk ||
                 x->ldr < 2 * x->rel_w - 1 + 2 * x->rel_h - 1))
    return set_error(SAV_ERR_SHAPE, "%s: relative logits need Nq == Nk == rel_w*rel_h", who);
  a.S = x->S; a.P = reinterpret_cast<bf16*>(x->P); a.dP = x->dP;
  a.dS = reinterpret_cast<bf16*>(x->dS);
  a.mix_pre = x->mix_pre; a.mix_post = x->mix_post;
  a.dmix_pre = x->dmix_pre; a.dmix_post = x->dmix_post;
  a.rel = x->rel; a.drel = x->drel;
  a.B = (int)x->B; a.H = (int)x->H; a.Nq = (int)x->Nq; a.Nk = (int)x->Nk; a.ld = (int)x->ld;
  a.ldr = (int)x->ldr; a.rel_w = x->rel_w; a.rel_h = x->rel_h;
  a.scale = x->scale;
  return SAV_OK;
}

}  // namespace sav

using namespace sav;

extern "C" int sav_thsoftmax_fwd(const sav_thsoftmax_args* x, void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  SoftArgs a;
  int rc = fill_args(x, a, "sav_thsoftmax_fwd");
  if (rc) return rc;
  if (!a.S || !a.P) return set_error(SAV_ERR_ARG, "sav_thsoftmax_fwd: S and P required");
  if (!a.mix_pre) {
    softmax_rows_fwd_kernel<<<soft_grid((int64_t)a.B * a.H * a.Nq), kRowWarps * 32, 0, st>>>(a);
  } else {
    const int grid = soft_grid((int64_t)a.B * a.Nq);
    const bool res = a.ld <= 32 * kKPL;   // the row fits the register-resident variant
    switch (a.H) {
#define SAV_TH_CASE(Hv)                                                              \
  case Hv:                                                                           \
    if (res) thsoftmax_fwd_tc_kernel<Hv><<<grid, kRowWarps * 32, 0, st>>>(a);        \
    else thsoftmax_fwd_kernel<Hv><<<grid, kRowWarps * 32, 0, st>>>(a);               \
    break;
      SAV_TH_CASE(2) SAV_TH_CASE(3) SAV_TH_CASE(4) SAV_TH_CASE(6) SAV_TH_CASE(8)
#undef SAV_TH_CASE
      case 12: thsoftmax_fwd_kernel<12><<<grid, kRowWarps * 32, 0, st>>>(a); break;
      case 16: thsoftmax_fwd_kernel<16><<<grid, kRowWarps * 32, 0, st>>>(a); break;
      default:
        return set_error(SAV_ERR_SHAPE, "sav_thsoftmax_fwd: talking heads built for H in {2,3,4,6,8,12,16}");
    }
  }
  count_launch();
  return check_launch("sav_thsoftmax_fwd");
}

extern "C" int sav_thsoftmax_bwd(const sav_thsoftmax_args* x, void* stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  SoftArgs a;
  int rc = fill_args(x, a, "sav_thsoftmax_bwd");
  if (rc) return rc;
  if (!a.S || !a.dP || !a.dS) return set_error(SAV_ERR_ARG, "sav_thsoftmax_bwd: S, dP, dS required");
  if (a.rel && !a.drel) return set_error(SAV_ERR_ARG, "sav_thsoftmax_bwd: drel required with rel");
  if (!a.mix_pre) {
    const size_t sm = a.rel ? (size_t)kRowWarps * a.ldr * sizeof(float) : 0;
    if (sm > 48 * 1024) return set_error(SAV_ERR_SHAPE, "sav_thsoftmax_bwd: relative table too wide");
    softmax_rows_bwd_kernel<<<soft_grid((int64_t)a.B * a.H * a.Nq), kRowWarps * 32, sm, st>>>(a);
  } else {
    if (!a.dmix_pre || !a.dmix_post)
      return set_error(SAV_ERR_ARG, "sav_thsoftmax_bwd: dmix_pre / dmix_post required");
    // few, fat blocks: every block ends with H*H atomics per transform
    int64_t blocks = ((int64_t)a.B * a.Nq + kRowWarps - 1) / kRowWarps;
    const int grid = (int)(blocks < 148 * 4 ? blocks : 148 * 4);
    const bool res = a.ld <= 32 * kKPL;
    switch (a.H) {
#define SAV_TH_CASE(Hv)                                                              \
  case Hv:                                                                           \
    if (res) { if ((rc = launch_th_bwd_tc<Hv>(a, st))) return rc; }                  \
    else thsoftmax_bwd_kernel<Hv><<<grid, kRowWarps * 32, 0, st>>>(a);               \
    break;
      SAV_TH_CASE(2) SAV_TH_CASE(3) SAV_TH_CASE(4) SAV_TH_CASE(6) SAV_TH_CASE(8)
#undef SAV_TH_CASE
      case 12: launch_th_bwd_slices<12, 3, 0>(a, grid, st); break;
      case 16: launch_th_bwd_slices<16, 2, 0>(a, grid, st); break;
      default:
        return set_error(SAV_ERR_SHAPE, "sav_thsoftmax_bwd: talking heads built for H in {2,3,4,6,8,12,16}");
    }
  }
  count_launch();
  return check_launch("sav_thsoftmax_bwd");
}
