// test_gemm.cu — standalone check of pcv_gemm_bf16 against a host fp32 reference (no torch needed).
// Build: make -C plaincv_b200/csrc test_gemm ; run on a B200: ./plaincv_b200/csrc/test_gemm
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../include/plaincv_b200.h"

static float bf(float x) { return __bfloat162float(__float2bfloat16(x)); }
static uint32_t rng_state = 12345;
static float frand() {
  rng_state = rng_state * 1664525u + 1013904223u;
  return ((rng_state >> 8) & 0xFFFF) / 65536.0f - 0.5f;
}
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2);} } while (0)

// A(m,k): K-major -> a[m*lda+k]; MN-major -> a[k*lda+m]
static int run_case(int M, int N, int K, int a_mn, int b_mn, int epi, int splits, bool timeit) {
  const int lda = a_mn ? ((M + 7) / 8 * 8) : ((K + 7) / 8 * 8);
  const int ldb = b_mn ? ((N + 7) / 8 * 8) : ((K + 7) / 8 * 8);
  const int ldc = N + 8;
  const size_t a_rows = a_mn ? K : M, b_rows = b_mn ? K : N;
  std::vector<__nv_bfloat16> ha(a_rows * lda), hb(b_rows * ldb);
  std::vector<float> fa(ha.size()), fb(hb.size());
  for (size_t i = 0; i < ha.size(); ++i) { fa[i] = bf(frand()); ha[i] = __float2bfloat16(fa[i]); }
  for (size_t i = 0; i < hb.size(); ++i) { fb[i] = bf(frand()); hb[i] = __float2bfloat16(fb[i]); }
  const size_t c_bytes = (size_t)M * ldc * (epi == PCV_EPI_BF16 ? 2 : 4);
  void *da, *db, *dc;
  CK(cudaMalloc(&da, ha.size() * 2)); CK(cudaMalloc(&db, hb.size() * 2)); CK(cudaMalloc(&dc, c_bytes));
  CK(cudaMemcpy(da, ha.data(), ha.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice));
  std::vector<float> c0((size_t)M * ldc, 0.f);
  if (epi == PCV_EPI_F32_ATOMIC) {
    for (auto& v : c0) v = frand();
    CK(cudaMemcpy(dc, c0.data(), c_bytes, cudaMemcpyHostToDevice));
  } else {
    CK(cudaMemset(dc, 0xFF, c_bytes));
  }
  const float alpha = 0.5f;
  int rc = pcv_gemm_bf16(da, db, dc, M, N, K, lda, ldb, ldc, a_mn, b_mn, epi, splits, alpha, nullptr);
  if (rc) { printf("launch rc=%d\n", rc); return 1; }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("kernel error: %s\n", cudaGetErrorString(e)); exit(3); }
  std::vector<float> got((size_t)M * ldc);
  if (epi == PCV_EPI_BF16) {
    std::vector<__nv_bfloat16> t((size_t)M * ldc);
    CK(cudaMemcpy(t.data(), dc, c_bytes, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < t.size(); ++i) got[i] = __bfloat162float(t[i]);
  } else {
    CK(cudaMemcpy(got.data(), dc, c_bytes, cudaMemcpyDeviceToHost));
  }
  // reference on a sample of rows/cols (all if small)
  double max_err = 0; int bad = 0; long checked = 0;
  const int mstep = M > 512 ? 37 : 1, nstep = N > 512 ? 29 : 1;
  for (int m = 0; m < M; m += mstep)
    for (int n = 0; n < N; n += nstep) {
      double acc = 0;
      for (int k = 0; k < K; ++k) {
        const float av = a_mn ? fa[(size_t)k * lda + m] : fa[(size_t)m * lda + k];
        const float bv = b_mn ? fb[(size_t)k * ldb + n] : fb[(size_t)n * ldb + k];
        acc += (double)av * bv;
      }
      double want = alpha * acc + (epi == PCV_EPI_F32_ATOMIC ? c0[(size_t)m * ldc + n] : 0.0);
      double g = got[(size_t)m * ldc + n];
      double tol = (epi == PCV_EPI_BF16 ? 1e-2 * fabs(want) + 2e-2 : 1e-4 * fabs(want) + 1e-3) ;
      double err = fabs(g - want);
      if (!(err <= tol)) { if (bad < 5) printf("   mismatch m=%d n=%d got=%f want=%f\n", m, n, g, want); ++bad; }
      if (err > max_err) max_err = err;
      ++checked;
    }
  printf("[%s] M=%d N=%d K=%d a_mn=%d b_mn=%d epi=%d splits=%d  checked=%ld bad=%d max_err=%.5f\n", bad ? "FAIL" : "ok  ",
         M, N, K, a_mn, b_mn, epi, splits, checked, bad, max_err);
  if (timeit && !bad) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) pcv_gemm_bf16(da, db, dc, M, N, K, lda, ldb, ldc, a_mn, b_mn, epi, splits, alpha, nullptr);
    cudaEventRecord(e0);
    const int iters = 20;
    for (int i = 0; i < iters; ++i) pcv_gemm_bf16(da, db, dc, M, N, K, lda, ldb, ldc, a_mn, b_mn, epi, splits, alpha, nullptr);
    cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= iters;
    printf("      time %.3f ms  -> %.1f TFLOP/s\n", ms, 2.0 * M * N * K / ms * 1e-9);
  }
  cudaFree(da); cudaFree(db); cudaFree(dc);
  return bad ? 1 : 0;
}

int main(int argc, char** argv) {
  int fails = 0;
  const bool quick = argc > 1 && atoi(argv[1]) == 1;
  const bool notime = argc > 1 && atoi(argv[1]) == 2;
  for (int mode = 1; mode <= 2; ++mode) {  // 1: single-CTA tiles, 2: CTA-pair (cta_group::2) tiles
    pcv_gemm_set_mode(mode);
    printf("---- gemm mode %d (%s)\n", mode, mode == 1 ? "single-CTA 128xBN" : "CTA pair 256x256");
    // smallest first: one tile, one k-block, K-major/K-major
    fails += run_case(128, 128, 64, 0, 0, PCV_EPI_F32, 1, false);
    fails += run_case(256, 256, 64, 0, 0, PCV_EPI_F32, 1, false);
    fails += run_case(128, 128, 256, 0, 0, PCV_EPI_F32, 1, false);
    fails += run_case(256, 256, 64, 0, 1, PCV_EPI_F32, 1, false);
    fails += run_case(256, 256, 64, 1, 0, PCV_EPI_F32, 1, false);
    fails += run_case(256, 256, 64, 1, 1, PCV_EPI_F32, 1, false);
    if (quick) continue;
    for (int a_mn = 0; a_mn < 2; ++a_mn)
      for (int b_mn = 0; b_mn < 2; ++b_mn) {
        fails += run_case(256, 384, 192, a_mn, b_mn, PCV_EPI_BF16, 1, false);
        fails += run_case(200, 136, 72, a_mn, b_mn, PCV_EPI_F32, 1, false);       // ragged M/N/K tails
        fails += run_case(1000, 2736, 1000, a_mn, b_mn, PCV_EPI_BF16, 1, false);
        fails += run_case(384, 520, 2000, a_mn, b_mn, PCV_EPI_F32_ATOMIC, 4, false);
        fails += run_case(3000, 5472, 520, a_mn, b_mn, PCV_EPI_BF16, 1, false);   // several tiles per CTA (pair)
      }
  }
  if (quick) { printf("%s (%d failing cases)\n", fails ? "GEMM TEST FAILED" : "GEMM TEST PASSED", fails); return fails; }
  // persistent path + timing at the 420M shapes, both tilings
  for (int mode = 1; mode <= 2 && !notime; ++mode) {
    pcv_gemm_set_mode(mode);
    printf("---- timing, gemm mode %d\n", mode);
    fails += run_case(16384, 3072, 1024, 0, 1, PCV_EPI_BF16, 1, true);     // fwd  QKV
    fails += run_case(16384, 1024, 3072, 0, 0, PCV_EPI_BF16, 1, true);     // dgrad QKV
    fails += run_case(1024, 3072, 16384, 1, 1, PCV_EPI_F32_ATOMIC, mode == 1 ? 4 : 3, true);  // wgrad QKV
    fails += run_case(16384, 1024, 1024, 0, 1, PCV_EPI_BF16, 1, true);     // fwd out-proj
    fails += run_case(16384, 5472, 1024, 0, 1, PCV_EPI_BF16, 1, true);     // fwd gate/up
    fails += run_case(16384, 1024, 2736, 0, 1, PCV_EPI_BF16, 1, true);     // fwd down
    fails += run_case(1024, 5472, 16384, 1, 1, PCV_EPI_F32_ATOMIC, mode == 1 ? 4 : 2, true);  // wgrad gate/up
    fails += run_case(16384, 50280, 1024, 0, 1, PCV_EPI_BF16, 1, true);    // LM head fwd
    fails += run_case(16384, 1024, 50280, 0, 0, PCV_EPI_BF16, 1, true);    // LM head dgrad
    fails += run_case(8192, 8192, 8192, 0, 0, PCV_EPI_BF16, 1, true);      // the cuBLAS peak shape
  }
  pcv_gemm_set_mode(0);
  printf("%s (%d failing cases)\n", fails ? "GEMM TEST FAILED" : "GEMM TEST PASSED", fails);
  return fails;
}
