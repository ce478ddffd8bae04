// Host-side TMA tensor-map construction (driver entry point resolved at run time; no link-time libcuda dependency).
#pragma once
#include <cuda.h>
#include <stdint.h>

// 2D row-major matrix [rows, cols] of bf16 (elem_bytes 2) or fp32 (elem_bytes 4) with row pitch `pitch_bytes`;
// box = box_cols x box_rows, 128B swizzle, out-of-bounds elements read as zero.  box_cols * elem_bytes must be 128.
int vda_make_tmap_2d(CUtensorMap* out, const void* base, uint64_t cols, uint64_t rows, uint64_t pitch_bytes,
                     uint32_t box_cols, uint32_t box_rows, int elem_bytes = 2);
