// elementwise.cu — the HBM-bound pieces of the training step: parameter cast, embedding gather / scatter-add,
// residual-add + RMSNorm forward / backward, SwiGLU forward / backward.  All of them are vectorised (16-byte)
// coalesced kernels with warp-shuffle reductions; their roofline is HBM bandwidth (see DESIGN.md for the
// algorithmic bytes of each).
#include "sm100.cuh"
#include "pcv_internal.h"
#include <cstdlib>

namespace pcv {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// bf16 pair -> fp32 pair in two ALU ops (shift, mask): the same values as __bfloat1622float2, which costs three
__device__ __forceinline__ float2 unpack2_fast(uint32_t v) {
  return make_float2(__uint_as_float(v << 16), __uint_as_float(v & 0xffff0000u));
}

#ifndef PCV_RMS_BWD_PREFETCH
#define PCV_RMS_BWD_PREFETCH 1  // rows ahead (per warp) that the RMSNorm backward prefetches into L2; 0 = off
#endif
// asynchronous bulk prefetch of `bytes` (a multiple of 16, 16-byte aligned) of global memory into L2
__device__ __forceinline__ void l2_prefetch_bulk(const void* p, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

struct bf16x8 {
  uint4 raw;
  __device__ __forceinline__ void unpack(float* f) const {
    float2 a = unpack_bf16x2(raw.x), b = unpack_bf16x2(raw.y), c = unpack_bf16x2(raw.z), d = unpack_bf16x2(raw.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
  }
  __device__ __forceinline__ void pack(const float* f) {
    raw.x = pack_bf16x2(f[0], f[1]); raw.y = pack_bf16x2(f[2], f[3]);
    raw.z = pack_bf16x2(f[4], f[5]); raw.w = pack_bf16x2(f[6], f[7]);
  }
};

// ------------------------------------------------------------------------------------------------ cast + pad
__global__ void cast_pad_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int rows, int cols,
                                int src_ld, int dst_ld, bool vec_ok) {
  const int vpr = dst_ld / 8;
  const long total = static_cast<long>(rows) * vpr;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(i / vpr), c = static_cast<int>(i % vpr) * 8;
    const float* s = src + static_cast<long>(r) * src_ld + c;
    float f[8];
    if (vec_ok && c + 8 <= cols) {
      const float4 a = *reinterpret_cast<const float4*>(s), b = *reinterpret_cast<const float4*>(s + 4);
      f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = (c + j < cols) ? s[j] : 0.f;
    }
    bf16x8 o;
    o.pack(f);
    *reinterpret_cast<uint4*>(dst + static_cast<long>(r) * dst_ld + c) = o.raw;
  }
}

// ------------------------------------------------------------------------------------------------ embedding
__global__ void embed_fwd_kernel(const int32_t* __restrict__ ids, const __nv_bfloat16* __restrict__ table,
                                 __nv_bfloat16* __restrict__ x, int M, int d) {
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  const int vpr = d / 8;
  for (int m = blockIdx.x * warps_per_block + (threadIdx.x >> 5); m < M; m += gridDim.x * warps_per_block) {
    const uint4* src = reinterpret_cast<const uint4*>(table + static_cast<long>(ids[m]) * d);
    uint4* dst = reinterpret_cast<uint4*>(x + static_cast<long>(m) * d);
    for (int v = lane; v < vpr; v += 32) dst[v] = __ldg(src + v);
  }
}

__global__ void embed_bwd_kernel(const int32_t* __restrict__ ids, const __nv_bfloat16* __restrict__ dx,
                                 float* __restrict__ dE, int M, int d, int ldE) {
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  const int vpr = d / 8;
  for (int m = blockIdx.x * warps_per_block + (threadIdx.x >> 5); m < M; m += gridDim.x * warps_per_block) {
    const uint4* src = reinterpret_cast<const uint4*>(dx + static_cast<long>(m) * d);
    float* dst = dE + static_cast<long>(ids[m]) * ldE;
    for (int v = lane; v < vpr; v += 32) {
      bf16x8 in;
      in.raw = __ldg(src + v);
      float f[8];
      in.unpack(f);
      float* p = dst + v * 8;
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(f[0]), "f"(f[1]), "f"(f[2]), "f"(f[3]) : "memory");
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p + 4), "f"(f[4]), "f"(f[5]), "f"(f[6]), "f"(f[7]) : "memory");
    }
  }
}

// ------------------------------------------------------------------------------------------------ RMSNorm + residual
// One warp per row; each lane keeps VPL 16-byte vectors of the row in registers (d <= VPL*256).
template <int VPL>
__global__ void __launch_bounds__(256)
rmsnorm_residual_fwd_kernel(const __nv_bfloat16* __restrict__ res_in, const __nv_bfloat16* __restrict__ delta,
                            const float* __restrict__ scale, __nv_bfloat16* __restrict__ res_out,
                            __nv_bfloat16* __restrict__ y, float* __restrict__ rstd_out, int M, int d, float eps) {
  pdl_wait();
  pdl_trigger();
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= M) return;
  const int vpr = d / 8;
  const long base = static_cast<long>(row) * d;
  // every load of the row is in flight before the first use (the interleaved form left the second vector's loads
  // behind the first vector's store: 58 -> 37 us on the tangent twin of this kernel)
  bf16x8 ra[VPL], rb[VPL];
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int v = lane + 32 * i;
    if (v < vpr) {
      ra[i].raw = *reinterpret_cast<const uint4*>(res_in + base + v * 8);
      if (delta != nullptr) rb[i].raw = *reinterpret_cast<const uint4*>(delta + base + v * 8);
    }
  }
  float x[VPL][8];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int v = lane + 32 * i;
    if (v < vpr) {
      ra[i].unpack(x[i]);
      if (delta != nullptr) {
        float t[8];
        rb[i].unpack(t);
#pragma unroll
        for (int j = 0; j < 8; ++j) x[i][j] += t[j];
        // the residual stream is bf16 in the reference (x = x + f(...) in the compute dtype): round, then norm
        bf16x8 r;
        r.pack(x[i]);
        *reinterpret_cast<uint4*>(res_out + base + v * 8) = r.raw;
        r.unpack(x[i]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += x[i][j] * x[i][j];
    }
  }
  ss = warp_sum(ss);
  const float rstd = rsqrtf(ss / static_cast<float>(d) + eps);
  if (lane == 0) rstd_out[row] = rstd;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int v = lane + 32 * i;
    if (v < vpr) {
      const float4 s0 = __ldg(reinterpret_cast<const float4*>(scale + v * 8));
      const float4 s1 = __ldg(reinterpret_cast<const float4*>(scale + v * 8 + 4));
      const float sc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = x[i][j] * (rstd * sc[j]);
      bf16x8 r;
      r.pack(o);
      *reinterpret_cast<uint4*>(y + base + v * 8) = r.raw;
    }
  }
}

// Backward: persistent over rows so that each block reduces its dscale partial once.  Register diet (bf16 inputs stay
// packed, the scale is re-read from L1) so that two 256-thread blocks fit per SM and every load of a row — dy, x and the
// incoming residual gradient — is in flight before the first use.
template <int VPL>
__global__ void __launch_bounds__(256, 2)
rmsnorm_residual_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ dres_out,
                            const __nv_bfloat16* __restrict__ xres, const float* __restrict__ scale,
                            const float* __restrict__ rstd_in, __nv_bfloat16* __restrict__ dres_in,
                            float* __restrict__ dscale, int M, int d) {
  extern __shared__ float sm_dscale[];  // [warps][d]
  pdl_wait();
  pdl_trigger();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int vpr = d / 8;
  // fp32x2 math throughout (FMUL2 / FFMA2 halve the issue slots: the scalar form of this loop ran ~700 instructions per
  // row, 46 % issue utilisation at 4.6 TB/s against the forward kernel's 6.4 TB/s on the same bytes); dy * scale is kept
  // from the first pass, so the second pass unpacks only x and the incoming residual gradient.
  float2 ds2[VPL][4];
#pragma unroll
  for (int i = 0; i < VPL; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) ds2[i][j] = make_float2(0.f, 0.f);
  const float inv_d = 1.f / static_cast<float>(d);
  for (int row = blockIdx.x * nwarps + warp; row < M; row += gridDim.x * nwarps) {
    const long base = static_cast<long>(row) * d;
    bf16x8 rdy[VPL], rx[VPL], rres[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int v = lane + 32 * i;
      if (v < vpr) {
        rdy[i].raw = *reinterpret_cast<const uint4*>(dy + base + v * 8);
        rx[i].raw = *reinterpret_cast<const uint4*>(xres + base + v * 8);
        if (dres_out != nullptr) rres[i].raw = *reinterpret_cast<const uint4*>(dres_out + base + v * 8);
      }
    }
#if PCV_RMS_BWD_PREFETCH
    // this warp's next row starts its way from HBM to L2 now, under the math of this one: a warp walks ~7 rows one after
    // the other, each a full load -> reduce -> store round trip, and it is that latency, not the instruction count,
    // that held the kernel at 4.7 TB/s
    {
      const int nrow = row + PCV_RMS_BWD_PREFETCH * gridDim.x * nwarps;
      if (lane < 3 && nrow < M) {
        const __nv_bfloat16* src = lane == 0 ? dy : (lane == 1 ? xres : dres_out);
        if (src != nullptr) l2_prefetch_bulk(src + static_cast<long>(nrow) * d, static_cast<uint32_t>(d) * 2u);
      }
    }
#endif
    const float rstd = rstd_in[row];
    const float2 r2 = make_float2(rstd, rstd);
    float2 dot2 = make_float2(0.f, 0.f);
    float2 t2[VPL][4];  // dy * scale
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int v = lane + 32 * i;
      if (v < vpr) {
        const float4 s0 = __ldg(reinterpret_cast<const float4*>(scale + v * 8));
        const float4 s1 = __ldg(reinterpret_cast<const float4*>(scale + v * 8 + 4));
        const float2 sc[4] = {make_float2(s0.x, s0.y), make_float2(s0.z, s0.w), make_float2(s1.x, s1.y),
                              make_float2(s1.z, s1.w)};
        const uint32_t wdy[4] = {rdy[i].raw.x, rdy[i].raw.y, rdy[i].raw.z, rdy[i].raw.w};
        const uint32_t wx[4] = {rx[i].raw.x, rx[i].raw.y, rx[i].raw.z, rx[i].raw.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 g = unpack2_fast(wdy[j]);
          const float2 xh = __fmul2_rn(unpack2_fast(wx[j]), r2);
          ds2[i][j] = __ffma2_rn(g, xh, ds2[i][j]);
          t2[i][j] = __fmul2_rn(g, sc[j]);
          dot2 = __ffma2_rn(t2[i][j], xh, dot2);
        }
      }
    }
    const float dot = warp_sum(dot2.x + dot2.y) * inv_d;
    // dx = rstd * (dy * scale - xhat * dot) = (dy * scale) * rstd + x * (-rstd^2 * dot)   [xhat = x * rstd]
    const float kx = -rstd * rstd * dot;
    const float2 kx2 = make_float2(kx, kx);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int v = lane + 32 * i;
      if (v < vpr) {
        const uint32_t wx[4] = {rx[i].raw.x, rx[i].raw.y, rx[i].raw.z, rx[i].raw.w};
        const uint32_t wr[4] = {rres[i].raw.x, rres[i].raw.y, rres[i].raw.z, rres[i].raw.w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float2 acc = __fmul2_rn(t2[i][j], r2);
          if (dres_out != nullptr) acc = __fadd2_rn(acc, unpack2_fast(wr[j]));
          acc = __ffma2_rn(unpack2_fast(wx[j]), kx2, acc);
          o[j] = pack_bf16x2(acc.x, acc.y);
        }
        *reinterpret_cast<uint4*>(dres_in + base + v * 8) = make_uint4(o[0], o[1], o[2], o[3]);
      }
    }
  }
  // block reduction of the dscale partials, then one atomic per column per block
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int v = lane + 32 * i;
    if (v < vpr) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        sm_dscale[warp * d + v * 8 + 2 * j] = ds2[i][j].x;
        sm_dscale[warp * d + v * 8 + 2 * j + 1] = ds2[i][j].y;
      }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < d; c += blockDim.x) {
    float s = 0.f;
    for (int w = 0; w < nwarps; ++w) s += sm_dscale[w * d + c];
    atomicAdd(dscale + c, s);
  }
}

// ------------------------------------------------------------------------------------------------ SwiGLU
__global__ void swiglu_fwd_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ h, int M, int hid) {
  const int vpr = hid / 8;
  const long total = static_cast<long>(M) * vpr;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const long r = i / vpr;
    const int c = static_cast<int>(i % vpr) * 8;
    bf16x8 g, u;
    g.raw = *reinterpret_cast<const uint4*>(gu + r * 2 * hid + c);
    u.raw = *reinterpret_cast<const uint4*>(gu + r * 2 * hid + hid + c);
    float fg[8], fu[8], o[8];
    g.unpack(fg);
    u.unpack(fu);
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
      const float2 h2 = swiglu_fwd_pair(make_float2(fg[j], fg[j + 1]), make_float2(fu[j], fu[j + 1]));
      o[j] = h2.x;
      o[j + 1] = h2.y;
    }
    bf16x8 w;
    w.pack(o);
    *reinterpret_cast<uint4*>(h + r * hid + c) = w.raw;
  }
}

__global__ void swiglu_bwd_kernel(const __nv_bfloat16* __restrict__ gu, const __nv_bfloat16* __restrict__ dh,
                                  __nv_bfloat16* __restrict__ dgu, int M, int hid) {
  const int vpr = hid / 8;
  const long total = static_cast<long>(M) * vpr;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const long r = i / vpr;
    const int c = static_cast<int>(i % vpr) * 8;
    bf16x8 g, u, d;
    g.raw = *reinterpret_cast<const uint4*>(gu + r * 2 * hid + c);
    u.raw = *reinterpret_cast<const uint4*>(gu + r * 2 * hid + hid + c);
    d.raw = *reinterpret_cast<const uint4*>(dh + r * hid + c);
    float fg[8], fu[8], fd[8], og[8], ou[8];
    g.unpack(fg);
    u.unpack(fu);
    d.unpack(fd);
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
      float2 dg, du;
      swiglu_bwd_pair(make_float2(fg[j], fg[j + 1]), make_float2(fu[j], fu[j + 1]), make_float2(fd[j], fd[j + 1]), dg, du);
      og[j] = dg.x; og[j + 1] = dg.y;
      ou[j] = du.x; ou[j + 1] = du.y;
    }
    bf16x8 wg, wu;
    wg.pack(og);
    wu.pack(ou);
    *reinterpret_cast<uint4*>(dgu + r * 2 * hid + c) = wg.raw;
    *reinterpret_cast<uint4*>(dgu + r * 2 * hid + hid + c) = wu.raw;
  }
}

// ---- un-gated MLP activations: MLP = silu(fc1 x) (transformer.py:70-97), MLPReluSquared =
// relu(fc1 x)^2 (transformer.py:138-165).  mode 0: silu, 1: relu^2.  fp32 math on the bf16 Dense output, bf16 result.
template <int MODE>
__device__ __forceinline__ float act_fwd_elt(float x) {
  if constexpr (MODE == 0) return x * sigmoid_fast(x);
  const float r = fmaxf(x, 0.f);
  return r * r;
}
template <int MODE>
__device__ __forceinline__ float act_bwd_elt(float x, float dh) {
  if constexpr (MODE == 0) {
    const float s = sigmoid_fast(x);
    return dh * (s + x * s * (1.f - s));
  }
  return dh * 2.f * fmaxf(x, 0.f);
}
template <int MODE, bool BWD>
__global__ void act_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dh,
                           __nv_bfloat16* __restrict__ out, long nvec) {
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    bf16x8 a, d;
    a.raw = reinterpret_cast<const uint4*>(x)[i];
    float fa[8], fd[8], o[8];
    a.unpack(fa);
    if constexpr (BWD) {
      d.raw = reinterpret_cast<const uint4*>(dh)[i];
      d.unpack(fd);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = act_bwd_elt<MODE>(fa[j], fd[j]);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = act_fwd_elt<MODE>(fa[j]);
    }
    a.pack(o);
    reinterpret_cast<uint4*>(out)[i] = a.raw;
  }
}

__global__ void add_bf16_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                                __nv_bfloat16* __restrict__ out, long nvec) {
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    bf16x8 x, y;
    x.raw = reinterpret_cast<const uint4*>(a)[i];
    y.raw = reinterpret_cast<const uint4*>(b)[i];
    float fx[8], fy[8];
    x.unpack(fx);
    y.unpack(fy);
#pragma unroll
    for (int j = 0; j < 8; ++j) fx[j] += fy[j];
    x.pack(fx);
    reinterpret_cast<uint4*>(out)[i] = x.raw;
  }
}

static inline int ew_grid(long work_items, int threads) {
  long blocks = (work_items + threads - 1) / threads;
  const long cap = static_cast<long>(num_sms()) * 16;
  return static_cast<int>(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}
static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace pcv

using namespace pcv;

extern "C" int pcv_cast_pad_bf16(const float* src, void* dst, int rows, int cols, int src_ld, int dst_ld, void* stream) {
  if (!src || !dst || rows <= 0 || cols <= 0 || cols > dst_ld || cols > src_ld) return PCV_ERR_ARG;
  if (dst_ld % 8 || !aligned16(dst)) return PCV_ERR_ALIGN;
  const bool vec_ok = (src_ld % 4 == 0) && aligned16(src);
  const long total = static_cast<long>(rows) * (dst_ld / 8);
  cast_pad_kernel<<<ew_grid(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      src, static_cast<__nv_bfloat16*>(dst), rows, cols, src_ld, dst_ld, vec_ok);
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_embed_fwd(const int32_t* ids, const void* table_bf16, void* x, int M, int d, void* stream) {
  if (!ids || !table_bf16 || !x || M <= 0 || d <= 0) return PCV_ERR_ARG;
  if (d % 8 || !aligned16(table_bf16) || !aligned16(x)) return PCV_ERR_ALIGN;
  embed_fwd_kernel<<<ew_grid(static_cast<long>(M) * 32, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      ids, static_cast<const __nv_bfloat16*>(table_bf16), static_cast<__nv_bfloat16*>(x), M, d);
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_embed_bwd(const int32_t* ids, const void* dx, float* dE, int M, int d, int ldE, void* stream) {
  if (!ids || !dx || !dE || M <= 0 || d <= 0) return PCV_ERR_ARG;
  if (d % 8 || ldE % 4 || !aligned16(dx) || !aligned16(dE)) return PCV_ERR_ALIGN;
  embed_bwd_kernel<<<ew_grid(static_cast<long>(M) * 32, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      ids, static_cast<const __nv_bfloat16*>(dx), dE, M, d, ldE);
  count_launch();
  return cuda_rc();
}

template <int VPL>
static int launch_rms_fwd(const void* res_in, const void* delta, const float* scale, void* res_out, void* y, float* rstd,
                          int M, int d, float eps, cudaStream_t st) {
  const int rows_per_block = 8;
  cudaError_t le = launch_pdl(rmsnorm_residual_fwd_kernel<VPL>, dim3((M + rows_per_block - 1) / rows_per_block), dim3(256), 0, st,
                              static_cast<const __nv_bfloat16*>(res_in), static_cast<const __nv_bfloat16*>(delta), scale,
                              static_cast<__nv_bfloat16*>(res_out), static_cast<__nv_bfloat16*>(y), rstd, M, d, eps);
  count_launch();
  return le == cudaSuccess ? cuda_rc() : static_cast<int>(le);
}

extern "C" int pcv_rmsnorm_residual_fwd(const void* res_in, const void* delta, const float* scale, void* res_out,
                                        void* y, float* rstd, int M, int d, float eps, void* stream) {
  if (!res_in || !scale || !y || !rstd || M <= 0 || d <= 0 || d > 2048) return PCV_ERR_ARG;
  if (delta && !res_out) return PCV_ERR_ARG;
  if (d % 8 || !aligned16(res_in) || !aligned16(y) || !aligned16(scale) || (delta && !aligned16(delta)) ||
      (res_out && !aligned16(res_out)))
    return PCV_ERR_ALIGN;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int vpl = (d / 8 + 31) / 32;
  switch (vpl) {
    case 1: return launch_rms_fwd<1>(res_in, delta, scale, res_out, y, rstd, M, d, eps, st);
    case 2: return launch_rms_fwd<2>(res_in, delta, scale, res_out, y, rstd, M, d, eps, st);
    case 3: return launch_rms_fwd<3>(res_in, delta, scale, res_out, y, rstd, M, d, eps, st);
    case 4: return launch_rms_fwd<4>(res_in, delta, scale, res_out, y, rstd, M, d, eps, st);
    default: return launch_rms_fwd<8>(res_in, delta, scale, res_out, y, rstd, M, d, eps, st);
  }
}

template <int VPL>
static int launch_rms_bwd(const void* dy, const void* dres_out, const void* res_out, const float* scale,
                          const float* rstd, void* dres_in, float* dscale, int M, int d, cudaStream_t st) {
  const int threads = 256, nwarps = threads / 32;
  int grid = num_sms() * 2;
  const int max_grid = (M + nwarps - 1) / nwarps;
  if (grid > max_grid) grid = max_grid;
  const size_t smem = static_cast<size_t>(nwarps) * d * sizeof(float);
  auto kern = rmsnorm_residual_bwd_kernel<VPL>;
  if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  cudaError_t le = launch_pdl(kern, dim3(grid), dim3(threads), smem, st, static_cast<const __nv_bfloat16*>(dy),
                              static_cast<const __nv_bfloat16*>(dres_out), static_cast<const __nv_bfloat16*>(res_out), scale, rstd,
                              static_cast<__nv_bfloat16*>(dres_in), dscale, M, d);
  count_launch();
  return le == cudaSuccess ? cuda_rc() : static_cast<int>(le);
}

extern "C" int pcv_rmsnorm_residual_bwd(const void* dy, const void* dres_out, const void* res_out, const float* scale,
                                        const float* rstd, void* dres_in, float* dscale, int M, int d, void* stream) {
  if (!dy || !res_out || !scale || !rstd || !dres_in || !dscale || M <= 0 || d <= 0 || d > 2048) return PCV_ERR_ARG;
  if (d % 8 || !aligned16(dy) || !aligned16(res_out) || !aligned16(dres_in) || (dres_out && !aligned16(dres_out)))
    return PCV_ERR_ALIGN;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int vpl = (d / 8 + 31) / 32;
  switch (vpl) {
    case 1: return launch_rms_bwd<1>(dy, dres_out, res_out, scale, rstd, dres_in, dscale, M, d, st);
    case 2: return launch_rms_bwd<2>(dy, dres_out, res_out, scale, rstd, dres_in, dscale, M, d, st);
    case 3: return launch_rms_bwd<3>(dy, dres_out, res_out, scale, rstd, dres_in, dscale, M, d, st);
    case 4: return launch_rms_bwd<4>(dy, dres_out, res_out, scale, rstd, dres_in, dscale, M, d, st);
    default: return launch_rms_bwd<8>(dy, dres_out, res_out, scale, rstd, dres_in, dscale, M, d, st);
  }
}

extern "C" int pcv_swiglu_fwd(const void* gu, void* h, int M, int hid, void* stream) {
  if (!gu || !h || M <= 0 || hid <= 0) return PCV_ERR_ARG;
  if (hid % 8 || !aligned16(gu) || !aligned16(h)) return PCV_ERR_ALIGN;
  const long total = static_cast<long>(M) * (hid / 8);
  swiglu_fwd_kernel<<<ew_grid(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(gu), static_cast<__nv_bfloat16*>(h), M, hid);
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_swiglu_bwd(const void* gu, const void* dh, void* dgu, int M, int hid, void* stream) {
  if (!gu || !dh || !dgu || M <= 0 || hid <= 0) return PCV_ERR_ARG;
  if (hid % 8 || !aligned16(gu) || !aligned16(dh) || !aligned16(dgu)) return PCV_ERR_ALIGN;
  const long total = static_cast<long>(M) * (hid / 8);
  swiglu_bwd_kernel<<<ew_grid(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(gu), static_cast<const __nv_bfloat16*>(dh), static_cast<__nv_bfloat16*>(dgu), M,
      hid);
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_act_fwd(const void* x, void* h, long long n, int mode, void* stream) {
  if (!x || !h || n <= 0 || mode < 0 || mode > 1) return PCV_ERR_ARG;
  if (n % 8 || !aligned16(x) || !aligned16(h)) return PCV_ERR_ALIGN;
  const int grid = ew_grid(n / 8, 256);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
  __nv_bfloat16* hp = static_cast<__nv_bfloat16*>(h);
  if (mode == 0) act_kernel<0, false><<<grid, 256, 0, st>>>(xp, nullptr, hp, n / 8);
  else act_kernel<1, false><<<grid, 256, 0, st>>>(xp, nullptr, hp, n / 8);
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_act_bwd(const void* x, const void* dh, void* dx, long long n, int mode, void* stream) {
  if (!x || !dh || !dx || n <= 0 || mode < 0 || mode > 1) return PCV_ERR_ARG;
  if (n % 8 || !aligned16(x) || !aligned16(dh) || !aligned16(dx)) return PCV_ERR_ALIGN;
  const int grid = ew_grid(n / 8, 256);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
  const __nv_bfloat16* dp = static_cast<const __nv_bfloat16*>(dh);
  __nv_bfloat16* op = static_cast<__nv_bfloat16*>(dx);
  if (mode == 0) act_kernel<0, true><<<grid, 256, 0, st>>>(xp, dp, op, n / 8);
  else act_kernel<1, true><<<grid, 256, 0, st>>>(xp, dp, op, n / 8);
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_add_bf16(const void* a, const void* b, void* out, long long n, void* stream) {
  if (!a || !b || !out || n <= 0) return PCV_ERR_ARG;
  if (n % 8 || !aligned16(a) || !aligned16(b) || !aligned16(out)) return PCV_ERR_ALIGN;
  add_bf16_kernel<<<ew_grid(n / 8, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(a), static_cast<const __nv_bfloat16*>(b), static_cast<__nv_bfloat16*>(out), n / 8);
  count_launch();
  return cuda_rc();
}
