// pcv_internal.h — declarations shared between the translation units of libplaincv_b200.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/plaincv_b200.h"

namespace pcv {
int num_sms();
int make_tmap_2d_bf16(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows);
int get_tmap_2d_bf16(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows);
int get_tmap_2d_f32(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows);
void count_launch(int n = 1);
inline int cuda_rc() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? PCV_OK : static_cast<int>(e);
}
}  // namespace pcv
