// Host-side helpers shared by the launchers: thread-local error string, launch counter,
// TMA tensor-map encoding through the driver entry point (no link-time libcuda dependency).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/sav_abi.h"

namespace sav {

int set_error(int code, const char* fmt, ...);
void count_launch(int n = 1);
int check_launch(const char* what);

// Encode a rank-`rank` bf16 tensor map with 128-byte swizzle.
//   dims[i], box[i] in elements (dim 0 innermost, contiguous); strides_bytes[i] for i>=1
// Returns 0 or a negative sav_status.
int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                   const uint64_t* strides_bytes, const uint32_t* box);

// Same for 2-byte (bf16) or 4-byte (fp32) elements.
int make_tmap_nd(CUtensorMap* out, const void* base, int elem_bytes, int rank, const uint64_t* dims,
                 const uint64_t* strides_bytes, const uint32_t* box);

// 2-D row-major tensor [rows, cols] (elem_bytes 2 = bf16, 4 = fp32), row pitch ld elements,
// box [box_rows, box_cols] with box_cols * elem_bytes == 128 (one swizzle row).
int make_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, int64_t cols, int64_t rows,
                 int64_t ld, uint32_t box_cols, uint32_t box_rows);

// development timeline buffer (NULL unless sav_set_debug_buffer was called)
void* sav_debug_buffer();

// attention backward, Nq <= 256 path (attn_bwd.cu)
int launch_attn_bwd2(const sav_attn_bwd_args* a, cudaStream_t st);

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace sav
