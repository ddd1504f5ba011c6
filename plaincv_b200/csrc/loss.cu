// loss.cu — softmax cross-entropy over bf16 logits with integer labels, producing the per-row loss, the first-max
// argmax accuracy and (in the same pass over the row, held in shared memory) d(logits) for the LM-head backward.
// Restates _loss_and_acc (train_lm.py:181-186): fp32 logsumexp(logits) - logits[label], mean over all tokens.
// HBM-bound: one read of the bf16 row + one write of the bf16 gradient row.
#include "sm100.cuh"
#include "pcv_internal.h"
#include <float.h>

namespace pcv {

constexpr int kCeThreads = 512;
constexpr float kLog2eF = 1.4426950408889634f;
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct MaxIdx {
  float m;
  int i;
};
__device__ __forceinline__ MaxIdx better(MaxIdx a, MaxIdx b) {  // larger value wins; ties -> smaller index
  return (b.m > a.m || (b.m == a.m && b.i < a.i)) ? b : a;
}

// One block per row.  If kCache, the row is staged in shared memory so HBM sees it exactly once.
template <bool kCache>
__global__ void __launch_bounds__(kCeThreads)
softmax_ce_kernel(const __nv_bfloat16* __restrict__ logits, const int32_t* __restrict__ labels,
                  __nv_bfloat16* __restrict__ dlogits, float* __restrict__ row_loss, float* __restrict__ row_correct,
                  int V, int ldl, float grad_scale) {
  extern __shared__ uint4 row_cache[];  // ldl/8 vectors when kCache
  __shared__ float s_m[kCeThreads / 32], s_s[kCeThreads / 32];
  __shared__ int s_i[kCeThreads / 32];
  __shared__ float s_bcast[2];
  const int row = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const __nv_bfloat16* lrow = logits + static_cast<long>(row) * ldl;
  const int nvec = ldl / 8;
  const int label = labels[row];

  // pass 1: online max / sum-exp, argmax (first max)
  float m = -FLT_MAX, s = 0.f;
  MaxIdx best{-FLT_MAX, 0x7fffffff};
  constexpr int kUnroll = 4;  // loads of one batch are all in flight before the first is consumed
  for (int v0 = tid; v0 < nvec; v0 += kCeThreads * kUnroll) {
    uint4 raws[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int v = v0 + u * kCeThreads;
      raws[u] = (v < nvec) ? __ldg(reinterpret_cast<const uint4*>(lrow) + v) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int v = v0 + u * kCeThreads;
      if (v >= nvec) break;
      const uint4 raw = raws[u];
      if (kCache) row_cache[v] = raw;
      const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
      float f[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 t = unpack_bf16x2(w[j]);
        f[2 * j] = t.x;
        f[2 * j + 1] = t.y;
      }
      if (v * 8 >= V) continue;  // a vector of nothing but pad columns (the pad may be wider than one vector)
      if (v * 8 + 8 > V) {       // ragged last vector: neutralise the pad columns (it holds at least one real logit,
                                 // so m is finite below and exp2(-FLT_MAX * log2e - mb) = 0)
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (v * 8 + j >= V) f[j] = -FLT_MAX;
      }
      const float vm = fmaxf(fmaxf(fmaxf(f[0], f[1]), fmaxf(f[2], f[3])), fmaxf(fmaxf(f[4], f[5]), fmaxf(f[6], f[7])));
      if (vm > best.m) {  // rare after the first few vectors; first max wins inside the vector
        best.m = vm;
#pragma unroll
        for (int j = 7; j >= 0; --j)
          if (f[j] == vm) best.i = v * 8 + j;
      }
      if (vm > m) { s *= ex2_approx((m - vm) * kLog2eF); m = vm; }
      const float mb = m * kLog2eF;
      float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
      for (int j = 0; j < 8; j += 2) {
        acc0 += ex2_approx(fmaf(f[j], kLog2eF, -mb));      // pad columns: exp2(-huge) = 0
        acc1 += ex2_approx(fmaf(f[j + 1], kLog2eF, -mb));
      }
      s += acc0 + acc1;
    }
  }
  // warp reduce
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, m, o), os = __shfl_xor_sync(0xffffffffu, s, o);
    const float nm = fmaxf(m, om);
    s = s * __expf(m - nm) + os * __expf(om - nm);
    m = nm;
    MaxIdx ob{__shfl_xor_sync(0xffffffffu, best.m, o), __shfl_xor_sync(0xffffffffu, best.i, o)};
    best = better(best, ob);
  }
  if (lane == 0) { s_m[warp] = m; s_s[warp] = s; s_i[warp] = best.i; }
  __syncthreads();
  if (warp == 0) {
    float bm = (lane < kCeThreads / 32) ? s_m[lane] : -FLT_MAX;
    float bs = (lane < kCeThreads / 32) ? s_s[lane] : 0.f;
    // the per-warp argmax carries the warp max as its value
    MaxIdx bb{bm, (lane < kCeThreads / 32) ? s_i[lane] : 0x7fffffff};
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, bm, o), os = __shfl_xor_sync(0xffffffffu, bs, o);
      const float nm = fmaxf(bm, om);
      bs = bs * __expf(bm - nm) + os * __expf(om - nm);
      bm = nm;
      MaxIdx ob{__shfl_xor_sync(0xffffffffu, bb.m, o), __shfl_xor_sync(0xffffffffu, bb.i, o)};
      bb = better(bb, ob);
    }
    if (lane == 0) {
      const float lse = bm + logf(bs);
      const float picked = __bfloat162float(lrow[label]);
      row_loss[row] = lse - picked;
      row_correct[row] = (bb.i == label) ? 1.f : 0.f;
      s_bcast[0] = lse;
    }
  }
  if (dlogits == nullptr) return;
  __syncthreads();
  const float lse = s_bcast[0];
  __nv_bfloat16* drow = dlogits + static_cast<long>(row) * ldl;
  for (int v = tid; v < nvec; v += kCeThreads) {
    const uint4 raw = kCache ? row_cache[v] : *(reinterpret_cast<const uint4*>(lrow) + v);
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
    float g[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 t = unpack_bf16x2(w[j]);
      g[2 * j] = t.x;
      g[2 * j + 1] = t.y;
    }
    const float lse2 = lse * kLog2eF;
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] = ex2_approx(fmaf(g[j], kLog2eF, -lse2));
    if (v * 8 + 8 > V) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (v * 8 + j >= V) g[j] = 0.f;
    }
    if (v == (label >> 3)) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (j == (label & 7)) g[j] -= 1.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] *= grad_scale;
    uint4 o;
    o.x = pack_bf16x2(g[0], g[1]); o.y = pack_bf16x2(g[2], g[3]);
    o.z = pack_bf16x2(g[4], g[5]); o.w = pack_bf16x2(g[6], g[7]);
    *(reinterpret_cast<uint4*>(drow) + v) = o;
  }
}

// Second half of the cross-entropy that starts in the LM-head GEMM epilogue (pcv_gemm_bf16_ce): per (row, 128-column
// slot) the epilogue left (max, sum exp(logit - max)) of the bf16-rounded logits, slot-major [slots][M].  Three small
// steps, none with a per-row serial prologue (a one-block-per-row form — merge, barrier, stream — ran at 2.0 TB/s: two
// 62-register CTAs per SM could not hide the merge's dependent loads and its barrier in front of every row):
//   ce_merge_kernel : one THREAD per row walks the slots (coalesced: slot-major) -> (lse, row max) into slot 0 of the
//                     row, argmax candidate = INT_MAX
//   ce_apply_kernel : grid-stride over all 16-byte vectors of logits[M][ldl]: d(logits) = (exp(logit - lse) - onehot) *
//                     scale written in place, ONE exponential per logit; elements equal to the row max race for the
//                     first-max argmax with atomicMin; the thread that holds the label column writes the row's loss
//   ce_finish_kernel: row_correct = (argmax == label)
__global__ void __launch_bounds__(256)
ce_merge_kernel(float2* __restrict__ ce_ms, int n_slots, int M, int* __restrict__ arg) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= M) return;
  float m = -3.0e38f, sum = 0.f;
  for (int sl = 0; sl < n_slots; ++sl) {
    const float2 ms = ce_ms[static_cast<size_t>(sl) * M + row];
    const float nm = fmaxf(m, ms.x);
    sum = sum * ex2_approx((m - nm) * kLog2eF) + ms.y * ex2_approx((ms.x - nm) * kLog2eF);
    m = nm;
  }
  ce_ms[row] = make_float2(m + logf(sum), m);  // slot 0 of this row has been consumed: (lse, row max) take its place
  arg[row] = 0x7fffffff;
}

__global__ void __launch_bounds__(256)
ce_apply_kernel(const __nv_bfloat16* logits, const int32_t* __restrict__ labels, const float2* __restrict__ lse_max,
                __nv_bfloat16* dlogits, float* __restrict__ row_loss, int* __restrict__ arg, int M, int V, int ldl,
                float grad_scale) {
  // work item = (row, chunk of kUnroll * 256 vectors): the row's constants and all index arithmetic are per item, the
  // per-vector work is load, 8 exponentials, store
  constexpr int kUnroll = 4;
  constexpr int kChunk = kUnroll * 256;
  const int nvec = ldl / 8;
  const int nchunk = (nvec + kChunk - 1) / kChunk;
  const long items = static_cast<long>(M) * nchunk;
  for (long it = blockIdx.x; it < items; it += gridDim.x) {
    const int row = static_cast<int>(it / nchunk);
    const int c = static_cast<int>(it - static_cast<long>(row) * nchunk);
    const uint4* src = reinterpret_cast<const uint4*>(logits + static_cast<long>(row) * ldl);
    uint4* dst = reinterpret_cast<uint4*>(dlogits + static_cast<long>(row) * ldl);
    const int vbase = c * kChunk + threadIdx.x;
    uint4 raws[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int v = vbase + u * 256;
      raws[u] = (v < nvec) ? src[v] : make_uint4(0, 0, 0, 0);
    }
    const float2 lm = __ldg(lse_max + row);
    const int label = __ldg(labels + row);
    const float lse2 = lm.x * kLog2eF;
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int v = vbase + u * 256;
      if (v >= nvec) continue;  // (not `break`: it would keep raws[] dynamically indexed, i.e. in local memory)
      const uint32_t w[4] = {raws[u].x, raws[u].y, raws[u].z, raws[u].w};
      float f[8], g[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 t = unpack_bf16x2(w[j]);
        f[2 * j] = t.x;
        f[2 * j + 1] = t.y;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] = ex2_approx(fmaf(f[j], kLog2eF, -lse2));
      if (v * 8 + 8 > V) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (v * 8 + j >= V) { g[j] = 0.f; f[j] = -3.0e38f; }
      }
      const float vm = fmaxf(fmaxf(fmaxf(f[0], f[1]), fmaxf(f[2], f[3])), fmaxf(fmaxf(f[4], f[5]), fmaxf(f[6], f[7])));
      if (vm >= lm.y) {  // holds a row maximum: candidate for the first-max argmax
        int a = 0x7fffffff;
#pragma unroll
        for (int j = 7; j >= 0; --j)
          if (f[j] >= lm.y) a = v * 8 + j;
        atomicMin(arg + row, a);
      }
      // one-hot and label pick as selects over all eight lanes (an `if (j == label & 7)` body is turned into a
      // dynamically indexed access by the compiler, which puts f[] and g[] into local memory)
      const int lj = (v == (label >> 3)) ? (label & 7) : -1;
      float picked = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const bool hit = (j == lj);
        picked += hit ? f[j] : 0.f;
        g[j] = (g[j] - (hit ? 1.f : 0.f)) * grad_scale;
      }
      if (lj >= 0) row_loss[row] = lm.x - picked;
      uint4 o;
      o.x = pack_bf16x2(g[0], g[1]); o.y = pack_bf16x2(g[2], g[3]);
      o.z = pack_bf16x2(g[4], g[5]); o.w = pack_bf16x2(g[6], g[7]);
      dst[v] = o;
    }
  }
}

__global__ void __launch_bounds__(256)
ce_finish_kernel(const int* __restrict__ arg, const int32_t* __restrict__ labels, float* __restrict__ row_correct, int M) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row < M) {
    const int a = arg[row];  // (row_correct aliases arg: read before write, one thread per element)
    row_correct[row] = (a == labels[row]) ? 1.f : 0.f;
  }
}

__global__ void __launch_bounds__(1024) mean2_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                      float* __restrict__ out, int M) {
  __shared__ double sa[32], sb[32];
  double xa = 0, xb = 0;
  for (int i = threadIdx.x; i < M; i += blockDim.x) { xa += a[i]; xb += b[i]; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    xa += __shfl_xor_sync(0xffffffffu, xa, o);
    xb += __shfl_xor_sync(0xffffffffu, xb, o);
  }
  if ((threadIdx.x & 31) == 0) { sa[threadIdx.x >> 5] = xa; sb[threadIdx.x >> 5] = xb; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ta = 0, tb = 0;
    for (int w = 0; w < (blockDim.x >> 5); ++w) { ta += sa[w]; tb += sb[w]; }
    out[0] = static_cast<float>(ta / M);
    out[1] = static_cast<float>(tb / M);
  }
}

}  // namespace pcv

using namespace pcv;

extern "C" int pcv_softmax_ce(const void* logits, const int32_t* labels, void* dlogits, float* row_loss,
                              float* row_correct, int M, int V, int ldl, float grad_scale, void* stream) {
  if (!logits || !labels || !row_loss || !row_correct || M <= 0 || V <= 0 || V > ldl) return PCV_ERR_ARG;
  if (ldl % 8 || (reinterpret_cast<uintptr_t>(logits) & 15) || (dlogits && (reinterpret_cast<uintptr_t>(dlogits) & 15)))
    return PCV_ERR_ALIGN;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t row_bytes = static_cast<size_t>(ldl) * 2;
  if (row_bytes <= 200 * 1024) {
    auto kern = softmax_ce_kernel<true>;
    static bool attr = false;
    if (!attr) {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      if (e != cudaSuccess) return static_cast<int>(e);
      attr = true;
    }
    kern<<<M, kCeThreads, row_bytes, st>>>(static_cast<const __nv_bfloat16*>(logits), labels,
                                           static_cast<__nv_bfloat16*>(dlogits), row_loss, row_correct, V, ldl, grad_scale);
  } else {
    softmax_ce_kernel<false><<<M, kCeThreads, 0, st>>>(static_cast<const __nv_bfloat16*>(logits), labels,
                                                       static_cast<__nv_bfloat16*>(dlogits), row_loss, row_correct, V,
                                                       ldl, grad_scale);
  }
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_ce_from_stats(const void* logits, const int32_t* labels, void* ce_ms, int n_slots, void* dlogits,
                                 float* row_loss, float* row_correct, int M, int V, int ldl, float grad_scale, void* stream) {
  if (!logits || !labels || !ce_ms || !dlogits || !row_loss || !row_correct || M <= 0 || V <= 0 || V > ldl || n_slots <= 0)
    return PCV_ERR_ARG;
  if (n_slots != (ldl + 127) / 128) return PCV_ERR_ARG;
  if (ldl % 8 || (reinterpret_cast<uintptr_t>(logits) & 15) || (reinterpret_cast<uintptr_t>(dlogits) & 15) ||
      (reinterpret_cast<uintptr_t>(ce_ms) & 7) || (reinterpret_cast<uintptr_t>(row_correct) & 3))
    return PCV_ERR_ALIGN;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int* arg = reinterpret_cast<int*>(row_correct);  // the argmax candidates live in row_correct until the last step
  float2* ms = static_cast<float2*>(ce_ms);
  const int rb = (M + 255) / 256;
  ce_merge_kernel<<<rb, 256, 0, st>>>(ms, n_slots, M, arg);
  const long items = static_cast<long>(M) * ((ldl / 8 + 1023) / 1024);
  long blocks = items;
  const long cap = static_cast<long>(num_sms()) * 8;
  if (blocks > cap) blocks = cap;
  ce_apply_kernel<<<static_cast<int>(blocks), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(logits), labels, ms,
                                                            static_cast<__nv_bfloat16*>(dlogits), row_loss, arg, M, V, ldl,
                                                            grad_scale);
  ce_finish_kernel<<<rb, 256, 0, st>>>(arg, labels, row_correct, M);
  count_launch(3);
  return cuda_rc();
}

extern "C" int pcv_mean2(const float* row_loss, const float* row_correct, float* out, int M, void* stream) {
  if (!row_loss || !row_correct || !out || M <= 0) return PCV_ERR_ARG;
  mean2_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(row_loss, row_correct, out, M);
  count_launch();
  return cuda_rc();
}
