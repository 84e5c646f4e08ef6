// LayerNorm forward / backward over the fp32 residual stream: warp-per-row, 16/32-byte vector
// loads, fp32 statistics via warp shuffles.  HBM-roofline kernels.  Algorithmic bytes per element:
//   fwd  4 (x fp32) + 2 (y bf16)                                   = 6 B
//   bwd  2 (dy) + 4 (x) + 4 (dx_add) + 4 (dx) + 2 (dx bf16 copy)   = 16 B
// Semantics follow Flax nn.LayerNorm as used at vit.py:19,26,57: var = E[x^2] - E[x]^2,
// y = (x - mean) * rsqrt(var + eps) * scale + bias, computed in fp32, stored bf16.
// The residual stream and its gradient stay fp32 (DESIGN.md "precision"): a bf16 stream costs
// 1.5e-2 of the 2e-2 parity budget on its own (tests/tools/bf16_error_budget.py).
#include "common.cuh"
#include "host.h"

namespace sav {

constexpr int LN_MAX_CHUNKS = 4;  // 8-element chunks per lane -> C <= 32*4*8 = 1024

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
__device__ __forceinline__ void load8(const float* p, float (&f)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
  f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ void store8(bf16* p, const float (&f)[8]) {
  *reinterpret_cast<uint4*>(p) = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]),
                                            pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}
__device__ __forceinline__ void store8(float* p, const float (&f)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(f[4], f[5], f[6], f[7]);
}

template <int CHUNKS, typename TY>
__global__ void __launch_bounds__(256)
ln_fwd_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ scale,
              const float* __restrict__ bias, TY* __restrict__ y, int64_t ldy,
              float* __restrict__ mean_out, float* __restrict__ rstd_out, int64_t rows, int C,
              float eps) {
  const int lane = threadIdx.x & 31;
  const int64_t warp_global = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int64_t warps_total = (int64_t)gridDim.x * 8;
  const int nchunk = C / 8;
  float g[CHUNKS][8], b[CHUNKS][8];
#pragma unroll
  for (int i = 0; i < CHUNKS; ++i) {
    const int ch = lane + i * 32;
    if (ch < nchunk) {
      load8(scale + ch * 8, g[i]);
      load8(bias + ch * 8, b[i]);
    }
  }
  for (int64_t row = warp_global; row < rows; row += warps_total) {
    float v[CHUNKS][8];
    float s = 0.f, ss = 0.f;
#pragma unroll
    for (int i = 0; i < CHUNKS; ++i) {
      const int ch = lane + i * 32;
      if (ch < nchunk) {
        load8(x + row * ldx + ch * 8, v[i]);
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
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * g[i][j] + b[i][j];
        store8(y + row * ldy + ch * 8, o);
      }
    }
  }
}

// Forward for C <= 512 with balanced 16-byte pieces (4 columns per lane per piece: C = 384 is three
// pieces per lane, where 8-column chunks leave half the lanes with one chunk and half with two).
// Two rows per warp iteration keep twice the bytes in flight.
template <int PIECES>
__global__ void __launch_bounds__(256, 4)
ln_fwd_pieces_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ scale,
                     const float* __restrict__ bias, bf16* __restrict__ y, int64_t ldy,
                     float* __restrict__ mean_out, float* __restrict__ rstd_out, int64_t rows, int C,
                     float eps) {
  const int lane = threadIdx.x & 31;
  const int64_t warp_global = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int64_t warps_total = (int64_t)gridDim.x * 8;
  const int npiece = C / 4;
  float4 g[PIECES], b[PIECES];
#pragma unroll
  for (int i = 0; i < PIECES; ++i) {
    const int pc = lane + i * 32;
    g[i] = b[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (pc < npiece) {
      g[i] = *reinterpret_cast<const float4*>(scale + pc * 4);
      b[i] = *reinterpret_cast<const float4*>(bias + pc * 4);
    }
  }
  const float invC = 1.f / (float)C;
  for (int64_t row = warp_global * 2; row < rows; row += warps_total * 2) {
    const bool two = row + 1 < rows;
    float4 v0[PIECES], v1[PIECES];
#pragma unroll
    for (int i = 0; i < PIECES; ++i) {
      const int pc = lane + i * 32;
      v0[i] = v1[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (pc < npiece) {
        v0[i] = *reinterpret_cast<const float4*>(x + row * ldx + pc * 4);
        if (two) v1[i] = *reinterpret_cast<const float4*>(x + (row + 1) * ldx + pc * 4);
      }
    }
    float s0 = 0.f, q0 = 0.f, s1 = 0.f, q1 = 0.f;
#pragma unroll
    for (int i = 0; i < PIECES; ++i) {
      s0 += (v0[i].x + v0[i].y) + (v0[i].z + v0[i].w);
      q0 += (v0[i].x * v0[i].x + v0[i].y * v0[i].y) + (v0[i].z * v0[i].z + v0[i].w * v0[i].w);
      s1 += (v1[i].x + v1[i].y) + (v1[i].z + v1[i].w);
      q1 += (v1[i].x * v1[i].x + v1[i].y * v1[i].y) + (v1[i].z * v1[i].z + v1[i].w * v1[i].w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s0 += __shfl_xor_sync(0xffffffffu, s0, o);
      q0 += __shfl_xor_sync(0xffffffffu, q0, o);
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      q1 += __shfl_xor_sync(0xffffffffu, q1, o);
    }
    const float m0 = s0 * invC, m1 = s1 * invC;
    const float r0 = rsqrtf(fmaxf(q0 * invC - m0 * m0, 0.f) + eps);
    const float r1 = rsqrtf(fmaxf(q1 * invC - m1 * m1, 0.f) + eps);
    if (lane == 0) {
      mean_out[row] = m0; rstd_out[row] = r0;
      if (two) { mean_out[row + 1] = m1; rstd_out[row + 1] = r1; }
    }
#pragma unroll
    for (int i = 0; i < PIECES; ++i) {
      const int pc = lane + i * 32;
      if (pc < npiece) {
        *reinterpret_cast<uint2*>(y + row * ldy + pc * 4) = make_uint2(
            pack_bf16((v0[i].x - m0) * r0 * g[i].x + b[i].x, (v0[i].y - m0) * r0 * g[i].y + b[i].y),
            pack_bf16((v0[i].z - m0) * r0 * g[i].z + b[i].z, (v0[i].w - m0) * r0 * g[i].w + b[i].w));
        if (two)
          *reinterpret_cast<uint2*>(y + (row + 1) * ldy + pc * 4) = make_uint2(
              pack_bf16((v1[i].x - m1) * r1 * g[i].x + b[i].x, (v1[i].y - m1) * r1 * g[i].y + b[i].y),
              pack_bf16((v1[i].z - m1) * r1 * g[i].z + b[i].z, (v1[i].w - m1) * r1 * g[i].w + b[i].w));
      }
    }
  }
}

// dx = rstd * (g - mean_c(g) - xhat * mean_c(g * xhat)),  g = dy * scale   (+ dx_add)
// dscale / dbias: per-warp register partials -> shared-memory reduce -> one red.add per column
// per block into the (zero-initialised or accumulating) gradient arena.
template <int CHUNKS, typename TDY>
__global__ void __launch_bounds__(256, CHUNKS <= 2 ? 2 : 1)
ln_bwd_kernel(const TDY* __restrict__ dy, int64_t ldy, const float* __restrict__ x, int64_t ldx,
              const float* __restrict__ scale, const float* __restrict__ mean_in,
              const float* __restrict__ rstd_in, const float* __restrict__ dx_add,
              float* __restrict__ dx, bf16* __restrict__ dx16, float* __restrict__ dscale,
              float* __restrict__ dbias, float* __restrict__ dsum, int64_t rows, int C) {
  extern __shared__ float sacc[];  // [2][C] (+ [C] column sums of the bf16 dx copy when dsum)
  const int lane = threadIdx.x & 31;
  const int64_t warp_global = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int64_t warps_total = (int64_t)gridDim.x * 8;
  const int nchunk = C / 8;
  const int nacc = dsum ? 3 * C : 2 * C;
  for (int i = threadIdx.x; i < nacc; i += blockDim.x) sacc[i] = 0.f;
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
    float xh[CHUNKS][8], gy[CHUNKS][8], add[CHUNKS][8];
    float s1 = 0.f, s2 = 0.f;
    // all global loads of the row are issued before the first reduction (one latency per row)
    if (dx_add) {
#pragma unroll
      for (int i = 0; i < CHUNKS; ++i) {
        const int ch = lane + i * 32;
        if (ch < nchunk) load8(dx_add + row * ldx + ch * 8, add[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < CHUNKS; ++i) {
      const int ch = lane + i * 32;
      if (ch < nchunk) {
        float xv[8], dv[8];
        load8(x + row * ldx + ch * 8, xv);
        load8(dy + row * ldy + ch * 8, dv);
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
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += add[i][j];
        }
        store8(dx + row * ldx + ch * 8, o);
        if (dx16) store8(dx16 + row * ldx + ch * 8, o);
        if (dsum) {
          // column sums of the bf16 copy = bias gradient of the Linear that produced this stream
          // (its dY is exactly dx16): shared-memory adds, the accumulator registers are spoken for
#pragma unroll
          for (int j = 0; j < 8; ++j)
            atomicAdd(&sacc[2 * C + ch * 8 + j], __bfloat162float(__float2bfloat16(o[j])));
        }
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
  for (int i = threadIdx.x; i < nacc; i += blockDim.x) {
    float* out = (i < C) ? (dscale + i) : (i < 2 * C ? dbias + (i - C) : dsum + (i - 2 * C));
    atomicAdd(out, sacc[i]);
  }
}

// Lean backward for C <= 512 (every DeiT / CaiT / CvT width): a lane owns 16-byte pieces (4
// columns) lane, lane+32, ...; C = 384 is exactly three pieces per lane.  dy stays packed (bf16x4 in
// two registers) across the row reduction and g = dy*scale is recomputed after it, which keeps
// the kernel at <= 85 registers = three 256-thread blocks per SM: the kernel is bound by bytes in
// flight (one row = 3.8 KB per warp), not by arithmetic.
// DSUM: also accumulate the column sums of the bf16 copy dx16 -- the bias gradient of the Linear
// that wrote this stream (its dY IS dx16) -- in REGISTER accumulators (the shared-memory version
// of round 1 cost more than the separate 14 us column-sum pass it removed).
template <int PIECES, bool DSUM>
__global__ void __launch_bounds__(256, (PIECES <= 3 && !DSUM) ? 3 : 2)
ln_bwd_pieces_kernel(const bf16* __restrict__ dy, int64_t ldy, const float* __restrict__ x,
                     int64_t ldx, const float* __restrict__ scale, const float* __restrict__ mean_in,
                     const float* __restrict__ rstd_in, const float* __restrict__ dx_add,
                     float* __restrict__ dx, bf16* __restrict__ dx16, float* __restrict__ dscale,
                     float* __restrict__ dbias, float* __restrict__ dsum, int64_t rows, int C) {
  extern __shared__ float sacc[];  // [2 or 3][C]
  const int lane = threadIdx.x & 31;
  const int64_t warp_global = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int64_t warps_total = (int64_t)gridDim.x * 8;
  const int npiece = C / 4;
  const int nacc = (DSUM ? 3 : 2) * C;
  for (int i = threadIdx.x; i < nacc; i += blockDim.x) sacc[i] = 0.f;
  __syncthreads();
  float4 ds[PIECES], db[PIECES], g[PIECES], dc[DSUM ? PIECES : 1];
#pragma unroll
  for (int i = 0; i < PIECES; ++i) {
    const int pc = lane + i * 32;
    ds[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    db[i] = ds[i];
    if (DSUM) dc[i] = ds[i];
    g[i] = pc < npiece ? *reinterpret_cast<const float4*>(scale + pc * 4) : ds[i];
  }
  for (int64_t row = warp_global; row < rows; row += warps_total) {
    const float mean = mean_in[row], rstd = rstd_in[row];
    float4 xh[PIECES], add[PIECES];
    uint2 dvp[PIECES];
    // all global loads of the row are issued before the first reduction (one latency per row)
#pragma unroll
    for (int i = 0; i < PIECES; ++i) {
      const int pc = lane + i * 32;
      if (pc < npiece) {
        if (dx_add) add[i] = *reinterpret_cast<const float4*>(dx_add + row * ldx + pc * 4);
        xh[i] = *reinterpret_cast<const float4*>(x + row * ldx + pc * 4);
        dvp[i] = *reinterpret_cast<const uint2*>(dy + row * ldy + pc * 4);
      }
    }
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < PIECES; ++i) {
      const int pc = lane + i * 32;
      if (pc < npiece) {
        const float2 d01 = unpack_bf16(dvp[i].x), d23 = unpack_bf16(dvp[i].y);
        xh[i].x = (xh[i].x - mean) * rstd; xh[i].y = (xh[i].y - mean) * rstd;
        xh[i].z = (xh[i].z - mean) * rstd; xh[i].w = (xh[i].w - mean) * rstd;
        const float g0 = d01.x * g[i].x, g1 = d01.y * g[i].y, g2 = d23.x * g[i].z, g3 = d23.y * g[i].w;
        s1 += (g0 + g1) + (g2 + g3);
        s2 += (g0 * xh[i].x + g1 * xh[i].y) + (g2 * xh[i].z + g3 * xh[i].w);
        ds[i].x += d01.x * xh[i].x; ds[i].y += d01.y * xh[i].y;
        ds[i].z += d23.x * xh[i].z; ds[i].w += d23.y * xh[i].w;
        db[i].x += d01.x; db[i].y += d01.y; db[i].z += d23.x; db[i].w += d23.y;
      }
    }
    s1 = warp_sum(s1) / (float)C;
    s2 = warp_sum(s2) / (float)C;
#pragma unroll
    for (int i = 0; i < PIECES; ++i) {
      const int pc = lane + i * 32;
      if (pc < npiece) {
        const float2 d01 = unpack_bf16(dvp[i].x), d23 = unpack_bf16(dvp[i].y);
        float4 o;
        o.x = rstd * (d01.x * g[i].x - s1 - xh[i].x * s2);
        o.y = rstd * (d01.y * g[i].y - s1 - xh[i].y * s2);
        o.z = rstd * (d23.x * g[i].z - s1 - xh[i].z * s2);
        o.w = rstd * (d23.y * g[i].w - s1 - xh[i].w * s2);
        if (dx_add) { o.x += add[i].x; o.y += add[i].y; o.z += add[i].z; o.w += add[i].w; }
        *reinterpret_cast<float4*>(dx + row * ldx + pc * 4) = o;
        if (dx16) {
          const uint32_t p01 = pack_bf16(o.x, o.y), p23 = pack_bf16(o.z, o.w);
          *reinterpret_cast<uint2*>(dx16 + row * ldx + pc * 4) = make_uint2(p01, p23);
          if (DSUM) {
            const float2 r01 = unpack_bf16(p01), r23 = unpack_bf16(p23);
            dc[i].x += r01.x; dc[i].y += r01.y; dc[i].z += r23.x; dc[i].w += r23.y;
          }
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < PIECES; ++i) {
    const int pc = lane + i * 32;
    if (pc < npiece) {
      float* a = sacc + pc * 4;
      atomicAdd(a, ds[i].x); atomicAdd(a + 1, ds[i].y); atomicAdd(a + 2, ds[i].z); atomicAdd(a + 3, ds[i].w);
      a += C;
      atomicAdd(a, db[i].x); atomicAdd(a + 1, db[i].y); atomicAdd(a + 2, db[i].z); atomicAdd(a + 3, db[i].w);
      if (DSUM) {
        a += C;
        atomicAdd(a, dc[i].x); atomicAdd(a + 1, dc[i].y); atomicAdd(a + 2, dc[i].z); atomicAdd(a + 3, dc[i].w);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nacc; i += blockDim.x) {
    float* out = (i < C) ? (dscale + i) : (i < 2 * C ? dbias + (i - C) : dsum + (i - 2 * C));
    atomicAdd(out, sacc[i]);
  }
}

// Rows of 513 .. 1024 columns (ViT-B / Mixer-B 768, TNT-S 640, ViT-L 1024): the same lean kernel with
// a warp PAIR per row, each warp holding half of the row's pieces, so the register budget -- and the
// three (two with column sums) resident blocks per SM -- of the C <= 512 kernel carries over.  The
// chunk kernel above keeps 1 block per SM at these widths and adds the column sums with
// shared-memory atomics per element: 2.4 TB/s (profiles/r02ay_timeline_mixer_b_patch16.txt).
template <int PIECES, bool DSUM>
__global__ void __launch_bounds__(256, (PIECES <= 3 && !DSUM) ? 3 : 2)
ln_bwd_pieces2_kernel(const bf16* __restrict__ dy, int64_t ldy, const float* __restrict__ x,
                     int64_t ldx, const float* __restrict__ scale, const float* __restrict__ mean_in,
                     const float* __restrict__ rstd_in, const float* __restrict__ dx_add,
                     float* __restrict__ dx, bf16* __restrict__ dx16, float* __restrict__ dscale,
                     float* __restrict__ dbias, float* __restrict__ dsum, int64_t rows, int C) {
  extern __shared__ float sacc[];  // [2 or 3][C]
  __shared__ float xchg[4][2][2][2];   // [pair][row parity][half][s1, s2]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int pair = warp >> 1, half = warp & 1;
  const int pbase = half * PIECES * 32;                 // first piece of this warp's half row
  const int64_t warp_global = (int64_t)blockIdx.x * 4 + pair;      // a warp PAIR owns a row
  const int64_t warps_total = (int64_t)gridDim.x * 4;
  const int npiece = C / 4;
  uint32_t parity = 0;
  const int nacc = (DSUM ? 3 : 2) * C;
  for (int i = threadIdx.x; i < nacc; i += blockDim.x) sacc[i] = 0.f;
  __syncthreads();
  float4 ds[PIECES], db[PIECES], g[PIECES], dc[DSUM ? PIECES : 1];
#pragma unroll
  for (int i = 0; i < PIECES; ++i) {
    const int pc = pbase + lane + i * 32;
    ds[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    db[i] = ds[i];
    if (DSUM) dc[i] = ds[i];
    g[i] = pc < npiece ? *reinterpret_cast<const float4*>(scale + pc * 4) : ds[i];
  }
  for (int64_t row = warp_global; row < rows; row += warps_total) {
    const float mean = mean_in[row], rstd = rstd_in[row];
    float4 xh[PIECES], add[PIECES];
    uint2 dvp[PIECES];
    // all global loads of the row are issued before the first reduction (one latency per row)
#pragma unroll
    for (int i = 0; i < PIECES; ++i) {
      const int pc = pbase + lane + i * 32;
      if (pc < npiece) {
        if (dx_add) add[i] = *reinterpret_cast<const float4*>(dx_add + row * ldx + pc * 4);
        xh[i] = *reinterpret_cast<const float4*>(x + row * ldx + pc * 4);
        dvp[i] = *reinterpret_cast<const uint2*>(dy + row * ldy + pc * 4);
      }
    }
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < PIECES; ++i) {
      const int pc = pbase + lane + i * 32;
      if (pc < npiece) {
        const float2 d01 = unpack_bf16(dvp[i].x), d23 = unpack_bf16(dvp[i].y);
        xh[i].x = (xh[i].x - mean) * rstd; xh[i].y = (xh[i].y - mean) * rstd;
        xh[i].z = (xh[i].z - mean) * rstd; xh[i].w = (xh[i].w - mean) * rstd;
        const float g0 = d01.x * g[i].x, g1 = d01.y * g[i].y, g2 = d23.x * g[i].z, g3 = d23.y * g[i].w;
        s1 += (g0 + g1) + (g2 + g3);
        s2 += (g0 * xh[i].x + g1 * xh[i].y) + (g2 * xh[i].z + g3 * xh[i].w);
        ds[i].x += d01.x * xh[i].x; ds[i].y += d01.y * xh[i].y;
        ds[i].z += d23.x * xh[i].z; ds[i].w += d23.y * xh[i].w;
        db[i].x += d01.x; db[i].y += d01.y; db[i].z += d23.x; db[i].w += d23.y;
      }
    }
    // the two halves of the row meet in shared memory (slots alternate with the row parity, so one
    // 64-thread barrier per row is enough)
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    if (lane == 0) { xchg[pair][parity][half][0] = s1; xchg[pair][parity][half][1] = s2; }
    named_barrier_sync(1 + pair, 64);
    s1 = (xchg[pair][parity][0][0] + xchg[pair][parity][1][0]) / (float)C;
    s2 = (xchg[pair][parity][0][1] + xchg[pair][parity][1][1]) / (float)C;
    parity ^= 1;
#pragma unroll
    for (int i = 0; i < PIECES; ++i) {
      const int pc = pbase + lane + i * 32;
      if (pc < npiece) {
        const float2 d01 = unpack_bf16(dvp[i].x), d23 = unpack_bf16(dvp[i].y);
        float4 o;
        o.x = rstd * (d01.x * g[i].x - s1 - xh[i].x * s2);
        o.y = rstd * (d01.y * g[i].y - s1 - xh[i].y * s2);
        o.z = rstd * (d23.x * g[i].z - s1 - xh[i].z * s2);
        o.w = rstd * (d23.y * g[i].w - s1 - xh[i].w * s2);
        if (dx_add) { o.x += add[i].x; o.y += add[i].y; o.z += add[i].z; o.w += add[i].w; }
        *reinterpret_cast<float4*>(dx + row * ldx + pc * 4) = o;
        if (dx16) {
          const uint32_t p01 = pack_bf16(o.x, o.y), p23 = pack_bf16(o.z, o.w);
          *reinterpret_cast<uint2*>(dx16 + row * ldx + pc * 4) = make_uint2(p01, p23);
          if (DSUM) {
            const float2 r01 = unpack_bf16(p01), r23 = unpack_bf16(p23);
            dc[i].x += r01.x; dc[i].y += r01.y; dc[i].z += r23.x; dc[i].w += r23.y;
          }
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < PIECES; ++i) {
    const int pc = pbase + lane + i * 32;
    if (pc < npiece) {
      float* a = sacc + pc * 4;
      atomicAdd(a, ds[i].x); atomicAdd(a + 1, ds[i].y); atomicAdd(a + 2, ds[i].z); atomicAdd(a + 3, ds[i].w);
      a += C;
      atomicAdd(a, db[i].x); atomicAdd(a + 1, db[i].y); atomicAdd(a + 2, db[i].z); atomicAdd(a + 3, db[i].w);
      if (DSUM) {
        a += C;
        atomicAdd(a, dc[i].x); atomicAdd(a + 1, dc[i].y); atomicAdd(a + 2, dc[i].z); atomicAdd(a + 3, dc[i].w);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nacc; i += blockDim.x) {
    float* out = (i < C) ? (dscale + i) : (i < 2 * C ? dbias + (i - C) : dsum + (i - 2 * C));
    atomicAdd(out, sacc[i]);
  }
}

static int ln_grid(int64_t rows) {
  int64_t g = (rows + 7) / 8;
  const int64_t cap = 148 * 4;
  return (int)(g < cap ? (g < 1 ? 1 : g) : cap);
}

}  // namespace sav

using namespace sav;

extern "C" int sav_layernorm_fwd(const float* x, int64_t ldx, const float* scale,
                                 const float* bias, void* y, int64_t ldy, float* mean,
                                 float* rstd, int64_t rows, int64_t C, float eps, void* stream) {
  if (rows <= 0 || C <= 0 || C % 8 || C > 32 * LN_MAX_CHUNKS * 8 || ldx % 4 || ldy % 8)
    return set_error(SAV_ERR_SHAPE, "sav_layernorm_fwd: need C %% 8 == 0 and C <= 1024 (C=%lld)", (long long)C);
  if (!aligned16(x) || !aligned16(y) || !aligned16(scale) || !aligned16(bias))
    return set_error(SAV_ERR_ALIGN, "sav_layernorm_fwd: pointers must be 16 B aligned");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = ln_grid(rows);
  const int chunks = (int)((C / 8 + 31) / 32);
  // (no row-count threshold: the same kernel -- the same summation order -- for every batch size,
  // so per-sample results do not depend on how a batch is sharded over ranks)
  if (C <= 512) {
    const int pieces = (int)((C / 4 + 31) / 32);
    const int64_t want = (rows + 15) / 16;
    const int grid_p = (int)(want < 148 * 8 ? want : 148 * 8);
#define LN_FWDP(PC)                                                                              \
  ln_fwd_pieces_kernel<PC><<<grid_p, 256, 0, st>>>(x, ldx, scale, bias, reinterpret_cast<bf16*>(y), \
                                                    ldy, mean, rstd, rows, (int)C, eps)
    switch (pieces) {
      case 1: LN_FWDP(1); break;
      case 2: LN_FWDP(2); break;
      case 3: LN_FWDP(3); break;
      default: LN_FWDP(4); break;
    }
#undef LN_FWDP
    count_launch();
    return check_launch("sav_layernorm_fwd");
  }
#define LN_FWD(CH)                                                                          \
  ln_fwd_kernel<CH, bf16><<<grid, 256, 0, st>>>(x, ldx, scale, bias, reinterpret_cast<bf16*>(y), ldy, \
                                                 mean, rstd, rows, (int)C, eps)
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

extern "C" int sav_layernorm_bwd_colsum(const void* dy, int64_t ldy, const float* x, int64_t ldx,
                                        const float* scale, const float* mean, const float* rstd,
                                        const float* dx_add, float* dx, void* dx_bf16,
                                        float* dscale, float* dbias, float* dx_bf16_colsum,
                                        int64_t rows, int64_t C, void* stream) {
  if (dx_bf16_colsum && !dx_bf16)
    return set_error(SAV_ERR_ARG, "sav_layernorm_bwd_colsum: column sums need the bf16 copy");
  if (rows <= 0 || C <= 0 || C % 8 || C > 32 * LN_MAX_CHUNKS * 8 || ldx % 8 || ldy % 8)
    return set_error(SAV_ERR_SHAPE, "sav_layernorm_bwd: need C %% 8 == 0 and C <= 1024");
  if (!aligned16(dy) || !aligned16(x) || !aligned16(dx) || !aligned16(dx_add) || !aligned16(dx_bf16))
    return set_error(SAV_ERR_ALIGN, "sav_layernorm_bwd: pointers must be 16 B aligned");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = ln_grid(rows);
  const int chunks = (int)((C / 8 + 31) / 32);
  float* dsum = dx_bf16_colsum;
  const size_t sm = (dsum ? 3 : 2) * C * sizeof(float);
  if (C <= 512 && dx_bf16) {
    const int pieces = (int)((C / 4 + 31) / 32);
    const int grid_p = (int)((rows + 7) / 8 < 148 * 6 ? ((rows + 7) / 8 < 1 ? 1 : (rows + 7) / 8) : 148 * 6);
#define LN_BWDP(PC, DS)                                                                          \
  ln_bwd_pieces_kernel<PC, DS><<<grid_p, 256, sm, st>>>(reinterpret_cast<const bf16*>(dy), ldy, x, ldx, \
      scale, mean, rstd, dx_add, dx, reinterpret_cast<bf16*>(dx_bf16), dscale, dbias, dsum, rows, (int)C)
    if (dsum) {
      switch (pieces) {
        case 1: LN_BWDP(1, true); break;
        case 2: LN_BWDP(2, true); break;
        case 3: LN_BWDP(3, true); break;
        default: LN_BWDP(4, true); break;
      }
    } else {
      switch (pieces) {
        case 1: LN_BWDP(1, false); break;
        case 2: LN_BWDP(2, false); break;
        case 3: LN_BWDP(3, false); break;
        default: LN_BWDP(4, false); break;
      }
    }
#undef LN_BWDP
    count_launch();
    return check_launch("sav_layernorm_bwd");
  }
  if (C > 512 && dx_bf16) {
    const int pieces = (int)((C / 4 + 63) / 64);       // per warp of the pair
    const int grid_p = (int)((rows + 3) / 4 < 148 * 6 ? ((rows + 3) / 4 < 1 ? 1 : (rows + 3) / 4) : 148 * 6);
#define LN_BWDP2(PC, DS)                                                                         \
  ln_bwd_pieces2_kernel<PC, DS><<<grid_p, 256, sm, st>>>(reinterpret_cast<const bf16*>(dy), ldy, x, ldx, \
      scale, mean, rstd, dx_add, dx, reinterpret_cast<bf16*>(dx_bf16), dscale, dbias, dsum, rows, (int)C)
    if (dsum) {
      if (pieces <= 3) LN_BWDP2(3, true); else LN_BWDP2(4, true);
    } else {
      if (pieces <= 3) LN_BWDP2(3, false); else LN_BWDP2(4, false);
    }
#undef LN_BWDP2
    count_launch();
    return check_launch("sav_layernorm_bwd");
  }
#define LN_BWD(CH)                                                                            \
  ln_bwd_kernel<CH, bf16><<<grid, 256, sm, st>>>(reinterpret_cast<const bf16*>(dy), ldy, x, ldx, scale, \
                                                  mean, rstd, dx_add, dx,                            \
                                                  reinterpret_cast<bf16*>(dx_bf16), dscale, dbias, dsum, rows, (int)C)
  switch (chunks) {
    case 1: LN_BWD(1); break;
    case 2: LN_BWD(2); break;
    case 3: LN_BWD(3); break;
    default: LN_BWD(4); break;
  }
#undef LN_BWD
  count_launch();
  return check_launch("sav_layernorm_bwd");
}

extern "C" int sav_layernorm_bwd(const void* dy, int64_t ldy, const float* x, int64_t ldx,
                                 const float* scale, const float* mean, const float* rstd,
                                 const float* dx_add, float* dx, void* dx_bf16, float* dscale,
                                 float* dbias, int64_t rows, int64_t C, void* stream) {
  return sav_layernorm_bwd_colsum(dy, ldy, x, ldx, scale, mean, rstd, dx_add, dx, dx_bf16, dscale,
                                  dbias, nullptr, rows, C, stream);
}

// ---- fp32-in / fp32-out flavour: the LayerNorm that closes the CvT conv token embedding
// (cvt.py:34) normalises INTO the residual stream, so its output and the gradient coming back are
// the fp32 stream, not a bf16 GEMM operand.  Three launches per CvT step: the generic kernels.
extern "C" int sav_layernorm_fwd_f32(const float* x, int64_t ldx, const float* scale,
                                     const float* bias, float* y, int64_t ldy, float* mean,
                                     float* rstd, int64_t rows, int64_t C, float eps, void* stream) {
  if (rows <= 0 || C <= 0 || C % 8 || C > 32 * LN_MAX_CHUNKS * 8 || ldx % 4 || ldy % 4)
    return set_error(SAV_ERR_SHAPE, "sav_layernorm_fwd_f32: need C %% 8 == 0 and C <= 1024");
  if (!aligned16(x) || !aligned16(y) || !aligned16(scale) || !aligned16(bias))
    return set_error(SAV_ERR_ALIGN, "sav_layernorm_fwd_f32: pointers must be 16 B aligned");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = ln_grid(rows);
  const int chunks = (int)((C / 8 + 31) / 32);
#define LN_FWD32(CH) \
  ln_fwd_kernel<CH, float><<<grid, 256, 0, st>>>(x, ldx, scale, bias, y, ldy, mean, rstd, rows, (int)C, eps)
  switch (chunks) {
    case 1: LN_FWD32(1); break;
    case 2: LN_FWD32(2); break;
    case 3: LN_FWD32(3); break;
    default: LN_FWD32(4); break;
  }
#undef LN_FWD32
  count_launch();
  return check_launch("sav_layernorm_fwd_f32");
}

extern "C" int sav_layernorm_bwd_f32(const float* dy, int64_t ldy, const float* x, int64_t ldx,
                                     const float* scale, const float* mean, const float* rstd,
                                     const float* dx_add, float* dx, void* dx_bf16, float* dscale,
                                     float* dbias, int64_t rows, int64_t C, void* stream) {
  if (rows <= 0 || C <= 0 || C % 8 || C > 32 * LN_MAX_CHUNKS * 8 || ldx % 8 || ldy % 4)
    return set_error(SAV_ERR_SHAPE, "sav_layernorm_bwd_f32: need C %% 8 == 0 and C <= 1024");
  if (!aligned16(dy) || !aligned16(x) || !aligned16(dx) || !aligned16(dx_add) || !aligned16(dx_bf16))
    return set_error(SAV_ERR_ALIGN, "sav_layernorm_bwd_f32: pointers must be 16 B aligned");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = ln_grid(rows);
  const int chunks = (int)((C / 8 + 31) / 32);
  const size_t sm = 2 * C * sizeof(float);
#define LN_BWD32(CH)                                                                           \
  ln_bwd_kernel<CH, float><<<grid, 256, sm, st>>>(dy, ldy, x, ldx, scale, mean, rstd, dx_add, dx, \
                                                   reinterpret_cast<bf16*>(dx_bf16), dscale, dbias, nullptr, rows, (int)C)
  switch (chunks) {
    case 1: LN_BWD32(1); break;
    case 2: LN_BWD32(2); break;
    case 3: LN_BWD32(3); break;
    default: LN_BWD32(4); break;
  }
#undef LN_BWD32
  count_launch();
  return check_launch("sav_layernorm_bwd_f32");
}
