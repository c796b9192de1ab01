// Host-side packers that lay weights out exactly as the tensor cores read them.
#pragma once
#include <stdint.h>
#include <string.h>

#include <vector>

namespace rfk {

// src: [rows, ld] 16-bit elements (bf16 or fp16 bits) row-major with ld a multiple of 64 (K zero-padded).  Result: for each
// 64-wide K block ("atom") a tile of `rows` 128-byte rows in the K-major / 128-byte-swizzle
// layout described in sm100.cuh: chunk c of row r lives at chunk position c ^ (r & 7).
// rows must be a multiple of 8.  Output size: (ld / 64) * rows * 128 bytes.
inline void pack_kmajor_sw128(const uint16_t* src, int rows, int ld, std::vector<uint8_t>* out) {
  const int atoms = ld / 64;
  out->assign((size_t)atoms * rows * 128, 0);
  for (int a = 0; a < atoms; ++a)
    for (int r = 0; r < rows; ++r)
      for (int c = 0; c < 8; ++c) {
        const uint16_t* s = src + (size_t)r * ld + a * 64 + c * 8;
        uint8_t* d = out->data() + ((size_t)a * rows + r) * 128 + (size_t)((c ^ (r & 7)) * 16);
        memcpy(d, s, 16);
      }
}

}  // namespace rfk
