// Gradient exchange of the data-parallel step (train.py:96: lax.pmean over the devices) as ONE
// kernel over NVLink / NVSwitch peer memory instead of a library collective.
//
// Every rank's gradient arena lives in symmetric memory (same size on every rank, every rank holds
// a device pointer to every peer's copy, plus one multicast address that targets all copies at
// once).  Two-shot all-reduce, in place:
//   barrier A   block b of every rank has started  => every rank's gradient kernels have finished
//   reduce      rank r owns 1/world of the range: with multicast, `multimem.ld_reduce.add.v4.f32`
//               pulls the sum of the 16-byte vector across all ranks from the switch (NVLS) and
//               `multimem.st` pushes it back to every copy; without multicast the same with peer
//               loads in rank order and peer stores
//   barrier B   block b of every rank has written its part everywhere (release / acquire)
// One rank computes each element and broadcasts it, so all ranks end with bit-identical sums.
// The barriers are per block index (block b only talks to block b of the peers, through CAS on
// signal words that flip 0 -> 1 -> 0, so the kernel is re-launchable and CUDA-graph replayable
// without a host reset) and need no co-residency of the blocks of one GPU.
#include "common.cuh"
#include "host.h"

namespace sav {

__device__ __forceinline__ void dp_send_signal_relaxed(uint32_t* addr) {
  uint32_t old;
  do {
    asm volatile("atom.global.relaxed.sys.cas.b32 %0, [%1], 0, 1;" : "=r"(old) : "l"(addr) : "memory");
  } while (old != 0u);
}
__device__ __forceinline__ void dp_wait_signal_relaxed(uint32_t* addr) {
  uint32_t old;
  do {
    asm volatile("atom.global.relaxed.sys.cas.b32 %0, [%1], 1, 0;" : "=r"(old) : "l"(addr) : "memory");
  } while (old != 1u);
}
__device__ __forceinline__ void dp_send_signal_release(uint32_t* addr) {
  uint32_t old;
  do {
    asm volatile("atom.global.release.sys.cas.b32 %0, [%1], 0, 1;" : "=r"(old) : "l"(addr) : "memory");
  } while (old != 0u);
}
__device__ __forceinline__ void dp_wait_signal_acquire(uint32_t* addr) {
  uint32_t old;
  do {
    asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], 1, 0;" : "=r"(old) : "l"(addr) : "memory");
  } while (old != 1u);
}

// block b of this rank <-> block b of every peer.  Signal word (b, sender) of a rank's pad is
// written by `sender` and consumed by the owner.
template <bool ORDERED>
__device__ __forceinline__ void dp_sync_blocks(uint32_t* const* pads, int rank, int world) {
  if ((int)threadIdx.x < world) {
    const int peer = threadIdx.x;
    uint32_t* put = pads[peer] + (size_t)blockIdx.x * world + rank;
    uint32_t* get = pads[rank] + (size_t)blockIdx.x * world + peer;
    if (ORDERED) {
      dp_send_signal_release(put);
      dp_wait_signal_acquire(get);
    } else {
      dp_send_signal_relaxed(put);
      dp_wait_signal_relaxed(get);
    }
  }
}

template <bool MULTICAST>
__global__ void __launch_bounds__(512)
dp_allreduce_kernel(float* mc, float* const* __restrict__ peers, uint32_t* const* __restrict__ pads,
                    int64_t off_vec, int64_t nvec, int rank, int world) {
  dp_sync_blocks<false>(pads, rank, world);
  __syncthreads();
  const int64_t per = (nvec + world - 1) / world;
  const int64_t lo = (int64_t)rank * per;
  const int64_t hi = lo + per < nvec ? lo + per : nvec;
  // four independent 16-byte vectors per thread in flight (a switch round trip is several
  // microseconds: the bytes in flight, not the instruction count, set the rate)
  constexpr int U = 4;
  const int64_t step = (int64_t)gridDim.x * blockDim.x * U;
  for (int64_t base = lo + (int64_t)blockIdx.x * blockDim.x * U + threadIdx.x; base < hi; base += step) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = base + (int64_t)u * blockDim.x;
      if (i >= hi) break;
      const int64_t e = (off_vec + i) * 4;
      if (MULTICAST) {
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w)
                     : "l"(mc + e)
                     : "memory");
      } else {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int p = 0; p < world; ++p) {      // fixed rank order: the same sum whoever computes it
          float4 t;
          asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w)
                       : "l"(peers[p] + e)
                       : "memory");
          acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
        }
        v[u] = acc;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = base + (int64_t)u * blockDim.x;
      if (i >= hi) break;
      const int64_t e = (off_vec + i) * 4;
      if (MULTICAST) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc + e),
                     "f"(v[u].x), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w)
                     : "memory");
      } else {
        for (int p = 0; p < world; ++p)
          asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(peers[p] + e),
                       "f"(v[u].x), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w)
                       : "memory");
      }
    }
  }
  __syncthreads();
  dp_sync_blocks<true>(pads, rank, world);
}

}  // namespace sav

using namespace sav;

extern "C" int sav_dp_allreduce_sum_f32(uint64_t multicast_ptr, const uint64_t* peer_ptrs_dev,
                                        const uint64_t* signal_pad_ptrs_dev, int64_t signal_pad_bytes,
                                        int64_t offset, int64_t n, int32_t rank, int32_t world,
                                        int32_t blocks, void* stream) {
  if (n <= 0) return SAV_OK;
  if (world < 2 || world > 32 || rank < 0 || rank >= world || !peer_ptrs_dev || !signal_pad_ptrs_dev)
    return set_error(SAV_ERR_ARG, "sav_dp_allreduce_sum_f32: 2 <= world <= 32, peer and signal tables required");
  if (offset % 4 || n % 4 || (multicast_ptr & 15))
    return set_error(SAV_ERR_ALIGN, "sav_dp_allreduce_sum_f32: offset, n %% 4 floats, 16 B aligned buffers");
  if (blocks <= 0) blocks = 148;   // it runs alone (after the backward pass): one CTA per B200 SM
  const int64_t max_blocks = signal_pad_bytes / (4 * (int64_t)world);
  if (blocks > max_blocks) blocks = (int32_t)max_blocks;
  if (blocks <= 0) return set_error(SAV_ERR_ARG, "sav_dp_allreduce_sum_f32: signal pad too small");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  float* const* peers = reinterpret_cast<float* const*>(peer_ptrs_dev);
  uint32_t* const* pads = reinterpret_cast<uint32_t* const*>(signal_pad_ptrs_dev);
  if (multicast_ptr)
    dp_allreduce_kernel<true><<<blocks, 512, 0, st>>>(reinterpret_cast<float*>(multicast_ptr), peers, pads,
                                                      offset / 4, n / 4, rank, world);
  else
    dp_allreduce_kernel<false><<<blocks, 512, 0, st>>>(nullptr, peers, pads, offset / 4, n / 4, rank, world);
  count_launch();
  return check_launch("sav_dp_allreduce_sum_f32");
}
