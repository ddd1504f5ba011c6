// optim.cu — the step right after the hot path: global-norm clip + AdamW over the model's flat
// fp32 buffers, with the bf16 compute copy of the parameters written in the same pass.
//
// Replaces, in the reference:
//   _clip_grads                      train_lm.py:173-178   scale = min(1, max_norm / (global_norm(g) + 1e-6)), g *= scale
//   apply_grads                      train_lm.py:316-353   updates = tx.update(g, opt_state, params); params += updates
//   optax.adamw(lr, b1, b2, eps, wd) optim/factory.py:193-205  (optax 0.2.6: scale_by_adam -> add_decayed_weights ->
//                                    scale_by_learning_rate; no decay mask, eps_root = 0, bias correction with t = count+1)
// and flax's per-call promote_dtype of every kernel (the bf16 copy).  Both kernels are HBM-bound streams:
//   sum of squares : 4 B read per parameter
//   update         : 16 B read (p, g, m, v) + 14 B written (p, m, v fp32, p bf16) per parameter
#include <cuda_bf16.h>
#include "pcv_internal.h"

namespace pcv {

constexpr int kOptThreads = 256;

__device__ __forceinline__ float block_sum(float v, float* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  float t = 0.f;
  if (warp == 0) {
    t = lane < (blockDim.x >> 5) ? sh[lane] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  return t;  // valid in warp 0
}

// partials[blockIdx.x] = sum of g[i]^2 over this block's grid-stride share.  One partial per block and a fixed
// reduction order inside the block: the norm is bit-reproducible for a given grid.
__global__ void __launch_bounds__(kOptThreads)
sumsq_partials_kernel(const float* __restrict__ g, long n4, long n, float* __restrict__ partials) {
  __shared__ float sh[kOptThreads / 32];
  float acc = 0.f;
  const float4* g4 = reinterpret_cast<const float4*>(g);
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const float4 a = __ldg(g4 + i);
    acc += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
  }
  if (blockIdx.x == 0) {  // tail (n not a multiple of 4)
    for (long i = n4 * 4 + threadIdx.x; i < n; i += blockDim.x) acc += g[i] * g[i];
  }
  const float t = block_sum(acc, sh);
  if (threadIdx.x == 0) partials[blockIdx.x] = t;
}

struct AdamArgs {
  float max_norm;  // <= 0: no clipping
  float lr, b1, b2, eps, wd;
  float omb1, omb2;        // 1 - b1, 1 - b2 evaluated in double on the host (1.f - 0.999f is off by 5e-5 relative)
  float inv_bc1, inv_bc2;  // 1 / (1 - b1^t), 1 / (1 - b2^t)
};

__device__ __forceinline__ void adam_elt(float& p, float g, float& m, float& v, const AdamArgs& a, float scale) {
  g *= scale;
  m = a.b1 * m + a.omb1 * g;
  v = a.b2 * v + a.omb2 * g * g;
  const float mh = m * a.inv_bc1, vh = v * a.inv_bc2;
  const float u = mh / (sqrtf(vh) + a.eps) + a.wd * p;
  p -= a.lr * u;
}

// Every block first folds the partial sums of squares (at most a few hundred floats, L2-resident) into the clip
// scale, then streams its share: p, m, v updated in place, the bf16 copy written next to them.
__global__ void __launch_bounds__(kOptThreads)
adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
             __nv_bfloat16* __restrict__ pb, long n4, long n, const float* __restrict__ partials, int n_partials,
             float* __restrict__ norm_out, AdamArgs a) {
  __shared__ float sh[kOptThreads / 32];
  __shared__ float s_scale;
  float scale = 1.f;
  if (partials != nullptr) {
    float acc = 0.f;
    for (int i = threadIdx.x; i < n_partials; i += blockDim.x) acc += partials[i];
    const float t = block_sum(acc, sh);
    if (threadIdx.x == 0) {
      const float norm = sqrtf(t);
      s_scale = a.max_norm > 0.f ? fminf(1.f, a.max_norm / (norm + 1e-6f)) : 1.f;
      if (blockIdx.x == 0 && norm_out != nullptr) *norm_out = norm;
    }
    __syncthreads();
    scale = s_scale;
  }
  float4* p4 = reinterpret_cast<float4*>(p);
  const float4* g4 = reinterpret_cast<const float4*>(g);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  uint2* pb2 = reinterpret_cast<uint2*>(pb);
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    float4 pp = p4[i], mm = m4[i], vv = v4[i];
    const float4 gg = __ldg(g4 + i);
    adam_elt(pp.x, gg.x, mm.x, vv.x, a, scale);
    adam_elt(pp.y, gg.y, mm.y, vv.y, a, scale);
    adam_elt(pp.z, gg.z, mm.z, vv.z, a, scale);
    adam_elt(pp.w, gg.w, mm.w, vv.w, a, scale);
    p4[i] = pp;
    m4[i] = mm;
    v4[i] = vv;
    if (pb != nullptr) {
      const __nv_bfloat162 lo = __floats2bfloat162_rn(pp.x, pp.y), hi = __floats2bfloat162_rn(pp.z, pp.w);
      uint2 o;
      o.x = *reinterpret_cast<const uint32_t*>(&lo);
      o.y = *reinterpret_cast<const uint32_t*>(&hi);
      pb2[i] = o;
    }
  }
  if (blockIdx.x == 0) {
    for (long i = n4 * 4 + threadIdx.x; i < n; i += blockDim.x) {
      float pp = p[i], mm = m[i], vv = v[i];
      adam_elt(pp, g[i], mm, vv, a, scale);
      p[i] = pp;
      m[i] = mm;
      v[i] = vv;
      if (pb != nullptr) pb[i] = __float2bfloat16_rn(pp);
    }
  }
}

// Gradient buckets on their way to / from the all-reduce: fp32 -> bf16 (round to nearest even) and back.
__global__ void __launch_bounds__(kOptThreads)
cast_f32_to_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long n4, long n) {
  const float4* s4 = reinterpret_cast<const float4*>(src);
  uint2* d2 = reinterpret_cast<uint2*>(dst);
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const float4 a = __ldg(s4 + i);
    const __nv_bfloat162 lo = __floats2bfloat162_rn(a.x, a.y), hi = __floats2bfloat162_rn(a.z, a.w);
    uint2 o;
    o.x = *reinterpret_cast<const uint32_t*>(&lo);
    o.y = *reinterpret_cast<const uint32_t*>(&hi);
    d2[i] = o;
  }
  if (blockIdx.x == 0)
    for (long i = n4 * 4 + threadIdx.x; i < n; i += blockDim.x) dst[i] = __float2bfloat16_rn(src[i]);
}
__global__ void __launch_bounds__(kOptThreads)
cast_bf16_to_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, long n4, long n) {
  const uint2* s2 = reinterpret_cast<const uint2*>(src);
  float4* d4 = reinterpret_cast<float4*>(dst);
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const uint2 a = __ldg(s2 + i);
    d4[i] = make_float4(__uint_as_float(a.x << 16), __uint_as_float(a.x & 0xffff0000u), __uint_as_float(a.y << 16),
                        __uint_as_float(a.y & 0xffff0000u));
  }
  if (blockIdx.x == 0)
    for (long i = n4 * 4 + threadIdx.x; i < n; i += blockDim.x) dst[i] = __bfloat162float(src[i]);
}

static bool al16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; }

}  // namespace pcv

extern "C" int pcv_sumsq_partials(const float* g, long long n, float* partials, int n_partials, void* stream) {
  using namespace pcv;
  if (!g || !partials || n <= 0 || n_partials <= 0 || n_partials > 65535) return PCV_ERR_ARG;
  if (!al16(g)) return PCV_ERR_ALIGN;
  sumsq_partials_kernel<<<n_partials, kOptThreads, 0, static_cast<cudaStream_t>(stream)>>>(g, n / 4, n, partials);
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_adamw_step(float* p, const float* g, float* m, float* v, void* p_bf16, long long n,
                              const float* partials, int n_partials, float* norm_out, double max_norm, double lr,
                              double b1, double b2, double eps, double weight_decay, int step, void* stream) {
  using namespace pcv;
  if (!p || !g || !m || !v || n <= 0 || step < 1) return PCV_ERR_ARG;
  if (max_norm > 0.0 && (!partials || n_partials <= 0)) return PCV_ERR_ARG;
  if (!al16(p) || !al16(g) || !al16(m) || !al16(v) || (p_bf16 && (reinterpret_cast<uintptr_t>(p_bf16) & 7)))
    return PCV_ERR_ALIGN;
  AdamArgs a;
  // hyper-parameters arrive as doubles: 1 - b2 and the bias corrections are formed before rounding to fp32
  a.max_norm = static_cast<float>(max_norm);
  a.lr = static_cast<float>(lr); a.b1 = static_cast<float>(b1); a.b2 = static_cast<float>(b2);
  a.eps = static_cast<float>(eps); a.wd = static_cast<float>(weight_decay);
  a.omb1 = static_cast<float>(1.0 - b1);
  a.omb2 = static_cast<float>(1.0 - b2);
  a.inv_bc1 = static_cast<float>(1.0 / (1.0 - pow(b1, step)));
  a.inv_bc2 = static_cast<float>(1.0 / (1.0 - pow(b2, step)));
  const int grid = num_sms() * 8;
  adamw_kernel<<<grid, kOptThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      p, g, m, v, static_cast<__nv_bfloat16*>(p_bf16), n / 4, n, partials, partials ? n_partials : 0, norm_out, a);
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_cast_f32_bf16(const void* src, void* dst, long long n, int to_bf16, void* stream) {
  using namespace pcv;
  if (!src || !dst || n <= 0) return PCV_ERR_ARG;
  // the vector body needs 16-byte aligned fp32 and 8-byte aligned bf16 addresses
  const void* f = to_bf16 ? src : dst;
  const void* b = to_bf16 ? dst : src;
  if (!al16(f) || (reinterpret_cast<uintptr_t>(b) & 7)) return PCV_ERR_ALIGN;
  long blocks = (n / 4 + kOptThreads - 1) / kOptThreads;
  const long cap = static_cast<long>(num_sms()) * 8;
  const int grid = static_cast<int>(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (to_bf16)
    cast_f32_to_bf16_kernel<<<grid, kOptThreads, 0, st>>>(static_cast<const float*>(src),
                                                          static_cast<__nv_bfloat16*>(dst), n / 4, n);
  else
    cast_bf16_to_f32_kernel<<<grid, kOptThreads, 0, st>>>(static_cast<const __nv_bfloat16*>(src),
                                                          static_cast<float*>(dst), n / 4, n);
  count_launch();
  return cuda_rc();
}
