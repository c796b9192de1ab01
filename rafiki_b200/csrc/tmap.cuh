// Host helper: build a CUtensorMap (TMA descriptor) without linking libcuda -- the encoder is
// fetched from the driver through the runtime (cudaGetDriverEntryPoint).
#pragma once
#include <cuda.h>

#include "rfk_common.cuh"

namespace rfk {

// 2-D row-major uint8 tensor [rows, cols] with row pitch `pitch` bytes; box = box_rows x box_cols
// bytes, out-of-range elements read as zero.  swizzle128 (box_cols must be 128): 16-byte chunk c of
// box row r lands at chunk position c ^ (r & 7) of its 128-byte shared-memory row.
int make_tmap_u8_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t pitch,
                    uint32_t box_rows, uint32_t box_cols, bool swizzle128 = false);

// 2-D row-major tensor of 16-bit elements [rows, cols] with row pitch `pitch_bytes`; box = box_rows x
// 64 elements (128 bytes) written to shared memory with the 128-byte swizzle tcgen05.mma expects.
int make_tmap_16bit_sw128(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t pitch_bytes,
                          uint32_t box_rows);

}  // namespace rfk
