// ubench_tmem2.cu — tcgen05.ld shapes: bytes per clock per SM with 4 warps (one per lane quarter).
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

#define R4(a, o) "=r"(a[o]), "=r"(a[o + 1]), "=r"(a[o + 2]), "=r"(a[o + 3])
#define R16(a, o) R4(a, o), R4(a, o + 4), R4(a, o + 8), R4(a, o + 12)
#define R32(a, o) R16(a, o), R16(a, o + 16)
#define P4(o) "%" #o
// MODE 0: 32x32b.x32 (32 regs, 4 KB/warp)   MODE 1: 32x32b.x64 (64 regs, 8 KB)   MODE 2: 16x256b.x4 (16 regs: 16 lanes x 128 B = 2 KB)
// MODE 3: 16x256b.x8 (32 regs, 4 KB)        MODE 4: 16x128b.x8 (16 regs, 2 KB)     MODE 5: 32x32b.x16 (16 regs, 2 KB)   MODE 6: 16x64b.x16 (16 regs 2KB)
template <int MODE>
__device__ __forceinline__ int ld(uint32_t t, uint32_t* r) {
  if (MODE == 0) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : R32(r, 0) : "r"(t) : "memory");
    return 4096;
  } else if (MODE == 1) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x64.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,"
                 "%32,%33,%34,%35,%36,%37,%38,%39,%40,%41,%42,%43,%44,%45,%46,%47,%48,%49,%50,%51,%52,%53,%54,%55,%56,%57,%58,%59,%60,%61,%62,%63}, [%64];"
                 : R32(r, 0), R32(r, 32) : "r"(t) : "memory");
    return 8192;
  } else if (MODE == 2) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : R16(r, 0) : "r"(t) : "memory");
    return 2048;
  } else if (MODE == 3) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : R32(r, 0) : "r"(t) : "memory");
    return 4096;
  } else if (MODE == 4) {
    asm volatile("tcgen05.ld.sync.aligned.16x128b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : R16(r, 0) : "r"(t) : "memory");
    return 2048;
  } else if (MODE == 5) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : R16(r, 0) : "r"(t) : "memory");
    return 2048;
  } else {
    asm volatile("tcgen05.ld.sync.aligned.16x64b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : R16(r, 0) : "r"(t) : "memory");
    return 1024 * 2;
  }
}

template <int MODE, int DEPTH>
__global__ void __launch_bounds__(512) k(long long* cyc, float* sink, int iters, int* bytes_out) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tptr)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t base = tptr + (static_cast<uint32_t>((warp & 3) * 32) << 16);
  float acc = 0.f;
  uint32_t a[DEPTH][64];
  int bytes = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int dd = 0; dd < DEPTH; ++dd) bytes = ld<MODE>(base + dd * 64, a[dd]);   // DEPTH loads in flight, then one wait
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int dd = 0; dd < DEPTH; ++dd) acc += __uint_as_float(a[dd][it & 15]);
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) { cyc[blockIdx.x] = t1 - t0; *bytes_out = bytes * DEPTH; }
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tptr), "n"(512));
}

template <int MODE, int DEPTH>
static void run(int warps, const char* tag) {
  long long* cyc; float* sink; int* bo;
  const int iters = 2000;
  cudaMalloc(&cyc, 148 * sizeof(long long)); cudaMalloc(&sink, sizeof(float) * 148 * 512); cudaMalloc(&bo, 4);
  k<MODE, DEPTH><<<148, warps * 32>>>(cyc, sink, iters, bo);
  k<MODE, DEPTH><<<148, warps * 32>>>(cyc, sink, iters, bo);
  cudaError_t e = cudaDeviceSynchronize();
  long long h; int b;
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&b, bo, 4, cudaMemcpyDeviceToHost);
  printf("%-22s depth %d warps %2d : %6.1f B/clk/SM  (%6.1f clk per iteration)%s\n", tag, DEPTH, warps,
         (double)b * iters * warps / h, (double)h / iters, e == cudaSuccess ? "" : cudaGetErrorString(e));
  cudaFree(cyc); cudaFree(sink); cudaFree(bo);
}

int main() {
  for (int w : {1, 4, 8}) {
    run<0, 1>(w, "32x32b.x32");  run<0, 2>(w, "32x32b.x32");
    run<1, 1>(w, "32x32b.x64");
    run<5, 2>(w, "32x32b.x16");
    run<2, 2>(w, "16x256b.x4");
    run<3, 1>(w, "16x256b.x8");  run<3, 2>(w, "16x256b.x8");
    run<4, 2>(w, "16x128b.x8");
    run<6, 2>(w, "16x64b.x16");
  }
  return 0;
}
