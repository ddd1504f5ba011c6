// ubench_pipes.cu — issue-rate microbenchmark for the instructions of the attention softmax (sm_100a).
// Prints cycles per warp-instruction per SM sub-partition with 2 warps resident per sub-partition (256 threads/CTA,
// one CTA per SM), the occupancy of the softmax warpgroups.   nvcc -arch=sm_100a -O3 -o ubench_pipes ubench_pipes.cu
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ unsigned pack(float a, float b) {
  unsigned r;
  asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}

template <int MODE>
__global__ void k(float* out, long long* cyc, int iters) {
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = threadIdx.x * 1e-3f + i * 0.01f;
  unsigned acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
      if (MODE == 0) { v[i] = ex2f(v[i]); v[i + 1] = ex2f(v[i + 1]); }                       // 2 MUFU
      if (MODE == 1) { acc ^= pack(v[i], v[i + 1]); v[i] += 1.f; }                            // 1 F2FP (+1 FADD)
      if (MODE == 2) {                                                                      // 2 MUFU + 1 F2FP
        const float a = ex2f(v[i]), b = ex2f(v[i + 1]);
        acc ^= pack(a, b);
        v[i] = a; v[i + 1] = b;
      }
      if (MODE == 3) {                                                                      // FFMA2
        float2 x = __ffma2_rn(make_float2(v[i], v[i + 1]), make_float2(1.0001f, 0.9999f), make_float2(0.1f, 0.2f));
        v[i] = x.x; v[i + 1] = x.y;
      }
      if (MODE == 4) { v[i] = fmaxf(v[i], fmaxf(v[i + 1], v[(i + 2) & 15])); }                // FMNMX3
      if (MODE == 5) {                                                                      // the softmax mix per pair
        float2 x = __ffma2_rn(make_float2(v[i], v[i + 1]), make_float2(0.18f, 0.18f), make_float2(-0.1f, -0.1f));
        const float a = ex2f(x.x), b = ex2f(x.y);
        float2 s = __fadd2_rn(make_float2(v[(i + 2) & 15], v[(i + 3) & 15]), make_float2(a, b));
        acc ^= pack(a, b);
        v[i] = s.x; v[i + 1] = fmaxf(s.y, fmaxf(a, b));
      }
      if (MODE == 6) {  // exp2 on the FMA pipe: round-to-nearest split + degree-3 polynomial + exponent insert
        float2 x = make_float2(v[i], v[i + 1]);
        const float2 magic = make_float2(12582912.f, 12582912.f);
        float2 xr = __fadd2_rn(x, magic);
        float2 n = __fadd2_rn(xr, make_float2(-12582912.f, -12582912.f));
        float2 f = __fadd2_rn(x, make_float2(-n.x, -n.y));
        float2 p = __ffma2_rn(f, make_float2(0.0555f, 0.0555f), make_float2(0.2402f, 0.2402f));
        p = __ffma2_rn(p, f, make_float2(0.6931f, 0.6931f));
        p = __ffma2_rn(p, f, make_float2(1.f, 1.f));
        v[i] = __int_as_float(__float_as_int(p.x) + (__float_as_int(xr.x) << 23));
        v[i + 1] = __int_as_float(__float_as_int(p.y) + (__float_as_int(xr.y) << 23));
      }
    }
  }
  const long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

template <int MODE>
static void run(const char* name, double instr_per_pair) {
  float* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 256 * 4);
  cudaMalloc(&cyc, 8);
  const int iters = 2000;
  k<MODE><<<148, 256>>>(out, cyc, iters);
  k<MODE><<<148, 256>>>(out, cyc, iters);
  long long h = 0;
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  const double pairs = double(iters) * 8;  // per thread
  // 2 warps per SMSP share the pipes: cycles per (warp x pair) on one SMSP
  printf("%-34s %8.2f cyc per warp-pair of elements (2 warps/SMSP)  -> %6.2f cyc per warp-instr\n", name,
         h / pairs / 2.0, h / pairs / 2.0 / instr_per_pair);
  cudaFree(out);
  cudaFree(cyc);
}

int main() {
  run<0>("MUFU.EX2 x2", 2);
  run<1>("F2FP.BF16.PACK_AB (+FADD)", 2);
  run<2>("MUFU.EX2 x2 + F2FP", 3);
  run<3>("FFMA2", 1);
  run<4>("FMNMX3", 1);
  run<5>("softmax mix (FFMA2,2xEX2,FADD2,F2FP,FMNMX3)", 6);
  run<6>("poly exp2 x2 on FMA/ALU pipes", 11);
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return 0;
}
