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
// Programmatic dependent launch (PDL): a kernel launched through launch_pdl may start — up to its pdl_wait() — while
// its predecessor in the stream is still draining, so its prologue (barrier set-up, tensor-memory allocation, descriptor
// prefetch) and its launch latency leave the critical path.  Only kernels that execute pdl_wait() before their first
// global-memory access may be launched this way.  Off unless pcv_set_pdl(1) / env PCV_PDL=1 (measured: no gain).
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
inline int cuda_rc() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? PCV_OK : static_cast<int>(e);
}
}  // namespace pcv
