// ubench_tmem.cu — tensor-memory read bandwidth of one SM (sm_100a): W warps each loop over tcgen05.ld.32x32b.x32 of
// their own 32-lane quarter (warp w -> lanes 32*(w%4)), 4 KB per instruction.  Prints bytes per clock per SM for
// W = 1, 2, 4, 8, 16 warps, with and without 32 MUFU.EX2 per load (the softmax's ratio), and for two CTAs per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_tmem ubench_tmem.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ float ex2f(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__device__ __forceinline__ void ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}

template <int MUFU, int COLS>
__global__ void __launch_bounds__(512) k(long long* cyc, float* sink, int iters) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tptr)), "n"(COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t base = tptr + (static_cast<uint32_t>((warp & 3) * 32) << 16);
  float acc = threadIdx.x * 1e-6f;
  uint32_t a[32], b[32];
  __syncthreads();
  const long long t0 = clock64();
  ld32(base, a);
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int it = 0; it < iters; ++it) {
    // software-pipelined like the softmax: the next load is in flight while this chunk is "used"
    ld32(base + ((it & 1) ? 0 : 32) % (COLS - 31), b);
    if (MUFU) {
#pragma unroll
      for (int i = 0; i < 32; ++i) acc += ex2f(__uint_as_float(a[i]) * 1e-30f);
    } else {
      acc += __uint_as_float(a[it & 31]);
    }
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) a[i] = b[i];
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tptr), "n"(COLS));
}

template <int MUFU, int COLS>
static void run(int warps, int ctas_per_sm, const char* tag) {
  long long* cyc; float* sink;
  const int nblk = 148 * ctas_per_sm, iters = 2000;
  cudaMalloc(&cyc, nblk * sizeof(long long));
  cudaMalloc(&sink, sizeof(float) * nblk * 512);
  k<MUFU, COLS><<<nblk, warps * 32>>>(cyc, sink, iters);
  k<MUFU, COLS><<<nblk, warps * 32>>>(cyc, sink, iters);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[8];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  const double bytes = 4096.0 * iters * warps * ctas_per_sm;
  printf("%-28s warps/CTA %2d  CTAs/SM %d : %8lld cycles  -> %6.1f B/clk/SM  (%5.1f clk per 4 KB load per warp)%s\n", tag, warps,
         ctas_per_sm, h[0], bytes / h[0], (double)h[0] / iters, e == cudaSuccess ? "" : cudaGetErrorString(e));
  cudaFree(cyc); cudaFree(sink);
}

int main() {
  for (int w : {1, 2, 4, 8, 16}) run<0, 512>(w, 1, "tcgen05.ld x32 only");
  for (int w : {1, 2, 4, 8, 16}) run<1, 512>(w, 1, "ld x32 + 32 MUFU.EX2");
  for (int w : {4, 8}) run<0, 256>(w, 2, "ld only, 256 cols");
  for (int w : {4, 8}) run<1, 256>(w, 2, "ld + 32 MUFU, 256 cols");
  return 0;
}
