// curvature.cu — forward-mode (tangent) kernels for the generalised Gauss-Newton matvec of the LM.
//
// The reference builds GGN.v = J^T H (J v) with jax.jvp / jax.vjp around the same Transformer.__call__
// (optim/ggn_utils.py:69-153) and applies the softmax cross-entropy Hessian in closed form
// (softmax_cross_entropy_hessian_vec, optim/ggn_utils.py:19-53).  J^T is the backward path that already exists; this
// file holds what J v needs beyond plain GEMMs (every Dense tangent is two launches of the tcgen05 GEMM):
//
//   pcv_rmsnorm_residual_jvp   tangent of residual add + nn.RMSNorm            (transformer.py:41-47, 293, 334)
//   pcv_swiglu_jvp             tangent of silu(gate) * up                      (transformer.py:126)
//   pcv_act_jvp                tangent of silu / relu^2 (MLP, MLPReluSquared)  (transformer.py:70-97, 138-165)
//   pcv_attn_jvp               tangent of jax.nn.dot_product_attention         (transformer.py:233-240), tcgen05/TMEM
//   pcv_ce_hessian_vec         (diag(p) - p p^T) v / n_rows                    (optim/ggn_utils.py:19-53, :122-124)
//
// RoPE is linear, so its tangent is pcv_rope_qk applied to the tangent of the QKV projection.
#include "sm100.cuh"
#include "pcv_internal.h"
#include <float.h>
#include <math.h>

namespace pcv {
namespace {

constexpr float kLog2eC = 1.4426950408889634f;
constexpr float kScaleC = 0.125f;  // 1/sqrt(64)
constexpr int kTile = 128 * 128;   // bytes of one 128 x 64 bf16 tile

__device__ __forceinline__ float ex2c(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float warp_sum_c(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ void unpack8(uint4 raw, float* f) {
  const float2 a = unpack_bf16x2(raw.x), b = unpack_bf16x2(raw.y), c = unpack_bf16x2(raw.z), d = unpack_bf16x2(raw.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 o;
  o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
  o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
  return o;
}

// ------------------------------------------------------------------------------------------------ RMSNorm tangent
// y = x * r * s, r = rsqrt(mean(x^2) + eps)  =>  ty = r * s * (tx - xh * mean(xh * tx)) + xh * ts,  xh = x * r.
// x = the primal residual stream after the add (saved by the forward), tx = t_res + t_delta rounded to bf16 like the
// primal add.  One warp per row, 16-byte loads, fp32 statistics.
template <int VPL>
__global__ void __launch_bounds__(256)
rmsnorm_residual_jvp_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ rstd_in,
                            const float* __restrict__ scale, const float* __restrict__ tscale,
                            const __nv_bfloat16* __restrict__ tres_in, const __nv_bfloat16* __restrict__ tdelta,
                            __nv_bfloat16* __restrict__ tres_out, __nv_bfloat16* __restrict__ ty, int M, int d) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= M) return;
  const int vpr = d / 8;
  const long base = static_cast<long>(row) * d;
  const float rstd = rstd_in[row];
  // every load of the row is issued before the first use
  uint4 rx[VPL], rt[VPL], rd[VPL];
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int v = lane + 32 * i;
    if (v < vpr) {
      rx[i] = *reinterpret_cast<const uint4*>(x + base + v * 8);
      rt[i] = *reinterpret_cast<const uint4*>(tres_in + base + v * 8);
      if (tdelta != nullptr) rd[i] = *reinterpret_cast<const uint4*>(tdelta + base + v * 8);
    }
  }
  float xh[VPL][8], tx[VPL][8];
  float dot = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int v = lane + 32 * i;
    if (v < vpr) {
      unpack8(rx[i], xh[i]);
      unpack8(rt[i], tx[i]);
      if (tdelta != nullptr) {
        float t[8];
        unpack8(rd[i], t);
#pragma unroll
        for (int j = 0; j < 8; ++j) tx[i][j] += t[j];
        const uint4 r = pack8(tx[i]);
        *reinterpret_cast<uint4*>(tres_out + base + v * 8) = r;
        unpack8(r, tx[i]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        xh[i][j] *= rstd;
        dot += xh[i][j] * tx[i][j];
      }
    }
  }
  dot = warp_sum_c(dot) / static_cast<float>(d);
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int v = lane + 32 * i;
    if (v < vpr) {
      const float4 s0 = __ldg(reinterpret_cast<const float4*>(scale + v * 8));
      const float4 s1 = __ldg(reinterpret_cast<const float4*>(scale + v * 8 + 4));
      const float4 t0 = __ldg(reinterpret_cast<const float4*>(tscale + v * 8));
      const float4 t1 = __ldg(reinterpret_cast<const float4*>(tscale + v * 8 + 4));
      const float sc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
      const float ts[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = rstd * sc[j] * (tx[i][j] - xh[i][j] * dot) + xh[i][j] * ts[j];
      *reinterpret_cast<uint4*>(ty + base + v * 8) = pack8(o);
    }
  }
}

// ------------------------------------------------------------------------------------------------ activation tangents
__device__ __forceinline__ float sigmoid_c(float g) { return 1.f / (1.f + __expf(-g)); }

// th = silu'(g) * tg * u + silu(g) * tu      (gu, tgu: [M][2*hid] as [gate | up]; th: [M][hid])
__global__ void swiglu_jvp_kernel(const __nv_bfloat16* __restrict__ gu, const __nv_bfloat16* __restrict__ tgu,
                                  __nv_bfloat16* __restrict__ th, int M, int hid) {
  const int vpr = hid / 8;
  const long total = static_cast<long>(M) * vpr;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const long r = i / vpr;
    const int c = static_cast<int>(i % vpr) * 8;
    float g[8], u[8], tg[8], tu[8], o[8];
    unpack8(*reinterpret_cast<const uint4*>(gu + r * 2 * hid + c), g);
    unpack8(*reinterpret_cast<const uint4*>(gu + r * 2 * hid + hid + c), u);
    unpack8(*reinterpret_cast<const uint4*>(tgu + r * 2 * hid + c), tg);
    unpack8(*reinterpret_cast<const uint4*>(tgu + r * 2 * hid + hid + c), tu);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sg = sigmoid_c(g[j]);
      const float silu = g[j] * sg;
      const float dsilu = sg * (1.f + g[j] * (1.f - sg));
      o[j] = dsilu * tg[j] * u[j] + silu * tu[j];
    }
    *reinterpret_cast<uint4*>(th + r * hid + c) = pack8(o);
  }
}

// mode 0: th = silu'(x) * tx; mode 1: th = 2 * relu(x) * tx
__global__ void act_jvp_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ tx,
                               __nv_bfloat16* __restrict__ th, long nvec, int mode) {
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    float a[8], t[8], o[8];
    unpack8(*(reinterpret_cast<const uint4*>(x) + i), a);
    unpack8(*(reinterpret_cast<const uint4*>(tx) + i), t);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (mode == 0) {
        const float sg = sigmoid_c(a[j]);
        o[j] = sg * (1.f + a[j] * (1.f - sg)) * t[j];
      } else {
        o[j] = 2.f * fmaxf(a[j], 0.f) * t[j];
      }
    }
    *(reinterpret_cast<uint4*>(th) + i) = pack8(o);
  }
}

// ------------------------------------------------------------------------------------------------ CE Hessian-vector
// hv = (p * v - p * sum(p * v)) * out_scale per row, p = softmax(fp32(logits)); pad columns (>= V) get 0.
// One block per row; the bf16 row (<= ~100 KB) is re-read from L1/L2 in the later passes.
constexpr int kHvThreads = 512;

__device__ __forceinline__ float block_reduce(float v, float* sh, bool is_max) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float w = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_max ? fmaxf(v, w) : v + w;
  }
  __syncthreads();  // sh may still be read from the previous reduction
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  float r = is_max ? -FLT_MAX : 0.f;
  for (int w = 0; w < kHvThreads / 32; ++w) r = is_max ? fmaxf(r, sh[w]) : r + sh[w];
  return r;
}

// kCache: the logits row is staged in shared memory by the first pass, so the later passes re-read only `vec`.
template <bool kCache>
__global__ void __launch_bounds__(kHvThreads)
ce_hessian_vec_kernel(const __nv_bfloat16* __restrict__ logits, const __nv_bfloat16* vec, __nv_bfloat16* hv, int V,
                      int ldl, float out_scale) {  // hv may be vec itself (each thread rewrites only what it read)
  extern __shared__ uint4 hv_row_cache[];
  __shared__ float sh[kHvThreads / 32];
  const int tid = threadIdx.x;
  const long off = static_cast<long>(blockIdx.x) * ldl;
  const uint4* lrow = reinterpret_cast<const uint4*>(logits + off);
  const uint4* vrow = reinterpret_cast<const uint4*>(vec + off);
  uint4* orow = reinterpret_cast<uint4*>(hv + off);
  const int nvec = ldl / 8;
  float m = -FLT_MAX;
  for (int v = tid; v < nvec; v += kHvThreads) {
    const uint4 raw = __ldg(lrow + v);
    if (kCache) hv_row_cache[v] = raw;
    float f[8];
    unpack8(raw, f);
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (v * 8 + j < V) m = fmaxf(m, f[j]);
  }
  m = block_reduce(m, sh, true);  // (its barriers also publish the row cache)
  const float mb = m * kLog2eC;
  float se = 0.f, sev = 0.f;
  for (int v = tid; v < nvec; v += kHvThreads) {
    float f[8], t[8];
    unpack8(kCache ? hv_row_cache[v] : __ldg(lrow + v), f);
    unpack8(vrow[v], t);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (v * 8 + j < V) {
        const float e = ex2c(fmaf(f[j], kLog2eC, -mb));
        se += e;
        sev += e * t[j];
      }
    }
  }
  se = block_reduce(se, sh, false);
  sev = block_reduce(sev, sh, false);
  const float inv = 1.f / se;
  const float inner = sev * inv;  // p^T v
  for (int v = tid; v < nvec; v += kHvThreads) {
    float f[8], t[8], o[8];
    unpack8(kCache ? hv_row_cache[v] : __ldg(lrow + v), f);
    unpack8(vrow[v], t);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float p = ex2c(fmaf(f[j], kLog2eC, -mb)) * inv;
      o[j] = (v * 8 + j < V) ? p * (t[j] - inner) * out_scale : 0.f;
    }
    orow[v] = pack8(o);
  }
}

// ------------------------------------------------------------------------------------------------ attention tangent
// O = P V, P = softmax(S), S = scale * Q K^T (causal / document window).  With lse known from the forward pass P needs
// no running maximum, and
//     tO = P tV + (P * tS) V - delta * O,   tS = scale * (tQ K^T + Q tK^T),   delta_i = sum_j P_ij tS_ij,
// so one flash-style pass accumulates tO' = P tV + W V (W = P * tS) in tensor memory and delta in a register; the
// rank-one correction with the saved forward output O is applied when the tile is written.
//
// CTA (320 threads) = one (batch, head, 128-query tile):
//   warps 0-7 : two warpgroups, each one thread per query row and half of the tile's 128 key columns: S and tS from
//               TMEM -> P, W (bf16 pairs) back into TMEM, partial delta (exchanged through shared memory at the end);
//               the epilogue splits the 64 output dims the same way
//   warp 8    : TMA producer (Q, tQ once; K, tK, V, tV ring of 2 stages) + TMEM allocation
//   warp 9    : tcgen05.mma issuer: S = Q.K^T; tS = tQ.K^T + Q.tK^T; tO' += P.tV + W.V (A operands from TMEM)
// TMEM: S [0,128) tS [128,256) tO' [256,320) P [320,384) W [384,448).
constexpr int kJvpThreads = 320;
struct AttnJvpSmem {
  static constexpr int kQ = 0;                        // Q, tQ
  static constexpr int kRing = kQ + 2 * kTile;        // 2 stages x (K, tK, V, tV)
  static constexpr int kBar = kRing + 8 * kTile;
  static constexpr int kDelta = kBar + 256;           // [2][128] fp32 partial deltas
  static constexpr int kBytes = kDelta + 1024 + 1024; // + alignment slack
};

__global__ void __launch_bounds__(kJvpThreads, 1)
attn_jvp_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_tqkv,
                const int32_t* __restrict__ doc_start, const __nv_bfloat16* __restrict__ out,
                const float* __restrict__ lse, __nv_bfloat16* __restrict__ tout, int T, int NH) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = sbase + AttnJvpSmem::kQ, sTQ = sQ + kTile, sRing = sbase + AttnJvpSmem::kRing,
                 sBar = sbase + AttnJvpSmem::kBar;
  auto sK = [&](int s) { return sRing + (4 * s + 0) * kTile; };
  auto sTK = [&](int s) { return sRing + (4 * s + 1) * kTile; };
  auto sV = [&](int s) { return sRing + (4 * s + 2) * kTile; };
  auto sTV = [&](int s) { return sRing + (4 * s + 3) * kTile; };
  const uint32_t bQ = sBar;                                   // (tx)
  auto bKfull = [&](int s) { return sBar + 8u * (1 + s); };   // [2] (tx)
  auto bVfull = [&](int s) { return sBar + 8u * (3 + s); };   // [2] (tx)
  auto bEmpty = [&](int s) { return sBar + 8u * (5 + s); };   // [2] (1, commit)
  const uint32_t bSfull = sBar + 8u * 7;                      // (1, commit)
  const uint32_t bPfull = sBar + 8u * 8;                      // (256)
  const uint32_t bPVdone = sBar + 8u * 9;                     // (1, commit)
  const uint32_t tmem_ptr_smem = sBar + 8u * 10;
  volatile uint32_t* tmem_ptr_generic =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_ptr_smem - smem_u32(smem_raw)));

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // launch order (x fastest, z slowest): one sequence after the other — its K, tK, V, tV (16.8 MB at 16 heads x 2048)
  // stay in L2 while its query tiles sweep them (all sequences at once: 134 MB at C3, 3.5x the DRAM reads) — and within
  // a sequence the longest key loops first
  const int qt = gridDim.y - 1 - blockIdx.y;
  const int h = blockIdx.x, b = blockIdx.z;
  const int d = NH * 64;
  const int q0 = qt * 128;
  const long seq_row0 = static_cast<long>(b) * T;
  const int row0 = static_cast<int>(seq_row0);
  const int jhi = qt;
  const int jlo = doc_start ? (doc_start[seq_row0 + q0] / 128) : 0;
  const int n_it = jhi - jlo + 1;

  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tm_qkv);
    tma_prefetch_desc(&tm_tqkv);
    mbar_init(bQ, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(bKfull(s), 1);
      mbar_init(bVfull(s), 1);
      mbar_init(bEmpty(s), 1);
    }
    mbar_init(bSfull, 1);
    mbar_init(bPfull, 256);
    mbar_init(bPVdone, 1);
    fence_mbar_init();
  }
  if (warp == 8) {
    __syncwarp();
    tmem_alloc<512>(tmem_ptr_smem);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_generic;

  if (warp == 8) {
    // ============================== TMA producer ==============================
    if (lane == 0) {
      mbar_expect_tx(bQ, 2 * kTile);
      tma_load_2d(sQ, &tm_qkv, bQ, h * 64, row0 + q0);
      tma_load_2d(sTQ, &tm_tqkv, bQ, h * 64, row0 + q0);
      for (int idx = 0; idx < n_it; ++idx) {
        const int j = jlo + idx;
        const int slot = idx & 1;
        mbar_wait(bEmpty(slot), ((idx >> 1) & 1) ^ 1u);
        mbar_expect_tx(bKfull(slot), 2 * kTile);
        tma_load_2d(sK(slot), &tm_qkv, bKfull(slot), d + h * 64, row0 + j * 128);
        tma_load_2d(sTK(slot), &tm_tqkv, bKfull(slot), d + h * 64, row0 + j * 128);
        mbar_expect_tx(bVfull(slot), 2 * kTile);
        tma_load_2d(sV(slot), &tm_qkv, bVfull(slot), 2 * d + h * 64, row0 + j * 128);
        tma_load_2d(sTV(slot), &tm_tqkv, bVfull(slot), 2 * d + h * 64, row0 + j * 128);
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // ============================== MMA issuer ==============================
    const uint32_t lead = elect_one_sync();
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, false, false);
    constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, false, true);
    auto issue_S = [&](int slot) {
#pragma unroll
      for (int k = 0; k < 4; ++k)  // S = Q.K^T
        umma_bf16_ss_p(tb, umma_smem_desc(sQ + k * 32, 16, 1024), umma_smem_desc(sK(slot) + k * 32, 16, 1024), idesc_s,
                       k > 0 ? 1u : 0u, lead);
#pragma unroll
      for (int k = 0; k < 4; ++k)  // tS = tQ.K^T ...
        umma_bf16_ss_p(tb + 128, umma_smem_desc(sTQ + k * 32, 16, 1024), umma_smem_desc(sK(slot) + k * 32, 16, 1024),
                       idesc_s, k > 0 ? 1u : 0u, lead);
#pragma unroll
      for (int k = 0; k < 4; ++k)  // ... + Q.tK^T
        umma_bf16_ss_p(tb + 128, umma_smem_desc(sQ + k * 32, 16, 1024), umma_smem_desc(sTK(slot) + k * 32, 16, 1024),
                       idesc_s, 1u, lead);
      tc_commit_p(bSfull, lead);
    };
    auto issue_PV = [&](int slot, bool accumulate) {
#pragma unroll
      for (int k = 0; k < 8; ++k)  // tO' (+)= P.tV   (A = P from tensor memory, 8 columns per 16 keys)
        umma_bf16_ts_p(tb + 256, tb + 320 + k * 8, umma_smem_desc(sTV(slot) + k * 2048, kTile, 1024), idesc_pv,
                       (accumulate || k > 0) ? 1u : 0u, lead);
#pragma unroll
      for (int k = 0; k < 8; ++k)  // tO' += W.V
        umma_bf16_ts_p(tb + 256, tb + 384 + k * 8, umma_smem_desc(sV(slot) + k * 2048, kTile, 1024), idesc_pv, 1u, lead);
      tc_commit_p(bPVdone, lead);
    };
    mbar_wait(bQ, 0);
    mbar_wait(bKfull(0), 0);
    tc_fence_after();
    issue_S(0);
    for (int idx = 0; idx < n_it; ++idx) {
      const int slot = idx & 1;
      mbar_wait(bPfull, idx & 1);  // P, W of this tile are in tensor memory; S and tS have been consumed
      tc_fence_after();
      if (idx + 1 < n_it) {
        mbar_wait(bKfull(slot ^ 1), ((idx + 1) >> 1) & 1);
        tc_fence_after();
        issue_S(slot ^ 1);
      }
      mbar_wait(bVfull(slot), (idx >> 1) & 1);
      tc_fence_after();
      issue_PV(slot, idx > 0);
      tc_commit_p(bEmpty(slot), lead);  // K, tK (read by this tile's S) and V, tV are free once these complete
    }
    __syncwarp();
  } else {
    // ============================== softmax rows ==============================
    const int wg = warp >> 2;   // key columns [64 wg, 64 wg + 64) of every tile, output dims [32 wg, 32 wg + 32)
    const int r = tid & 127;
    const int t = q0 + r;
    const int ds_t = doc_start ? doc_start[seq_row0 + t] : 0;
    const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tS = tmem_base + lane_off, tTS = tS + 128, tO = tS + 256, tP = tS + 320, tW = tS + 384;
    constexpr float c2 = kScaleC * kLog2eC;
    const float lse2 = lse[(static_cast<long>(b) * NH + h) * T + t] * kLog2eC;
    float delta = 0.f;
    for (int idx = 0; idx < n_it; ++idx) {
      const int key0 = (jlo + idx) * 128;
      mbar_wait(bSfull, idx & 1);
      tc_fence_after();
      const int lo0 = ds_t - key0, hi0 = t - key0;  // visible key window of this row, relative to the tile
#pragma unroll 1
      for (int cc = 0; cc < 2; ++cc) {
        const int c = 2 * wg + cc;
        uint32_t sr[32], tr[32];
        tmem_ld_32x32b_x32(tS + c * 32, sr);
        tmem_ld_32x32b_x32(tTS + c * 32, tr);
        tmem_ld_wait();
        uint32_t pk[16], pw[16];
        const int lo = lo0 - c * 32, hi = hi0 - c * 32;
#pragma unroll
        for (int k = 0; k < 32; k += 2) {
          float e0 = ex2c(fmaf(__uint_as_float(sr[k]), c2, -lse2));
          float e1 = ex2c(fmaf(__uint_as_float(sr[k + 1]), c2, -lse2));
          e0 = (k >= lo && k <= hi) ? e0 : 0.f;
          e1 = (k + 1 >= lo && k + 1 <= hi) ? e1 : 0.f;
          const float w0 = e0 * (__uint_as_float(tr[k]) * kScaleC);
          const float w1 = e1 * (__uint_as_float(tr[k + 1]) * kScaleC);
          delta += w0 + w1;
          pk[k >> 1] = pack_bf16x2(e0, e1);
          pw[k >> 1] = pack_bf16x2(w0, w1);
        }
        if (cc == 0 && idx > 0) {  // the previous tile's P.tV / W.V still read P and W until this fires
          mbar_wait(bPVdone, (idx - 1) & 1);
          tc_fence_after();
        }
        tmem_st_32x32b_x16(tP + c * 16, pk);
        tmem_st_32x32b_x16(tW + c * 16, pw);
      }
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(bPfull);
    }
    // ---- the two halves of a row's delta meet in shared memory
    float* sdelta = reinterpret_cast<float*>(smem_raw + (sbase + AttnJvpSmem::kDelta - smem_u32(smem_raw)));
    sdelta[wg * 128 + r] = delta;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    delta += sdelta[(wg ^ 1) * 128 + r];
    // ---- epilogue: tO = tO' - delta * O, this warpgroup's 32 of the 64 output dims
    mbar_wait(bPVdone, (n_it - 1) & 1);
    tc_fence_after();
    uint32_t o0[32];
    tmem_ld_32x32b_x32(tO + 32 * wg, o0);
    tmem_ld_wait();
    const long grow = (seq_row0 + t) * d + h * 64 + 32 * wg;
    const uint4* osrc = reinterpret_cast<const uint4*>(out + grow);
    uint4* odst = reinterpret_cast<uint4*>(tout + grow);
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      float of[8], res[8];
      unpack8(__ldg(osrc + v), of);
#pragma unroll
      for (int j = 0; j < 8; ++j) res[j] = __uint_as_float(o0[v * 8 + j]) - delta * of[j];
      odst[v] = pack8(res);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace
}  // namespace pcv

using namespace pcv;

static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static inline int ew_blocks(long items, int threads) {
  long blocks = (items + threads - 1) / threads;
  const long cap = static_cast<long>(num_sms()) * 16;
  if (blocks > cap) blocks = cap;
  return static_cast<int>(blocks < 1 ? 1 : blocks);
}

template <int VPL>
static int launch_rms_jvp(const void* x, const float* rstd, const float* scale, const float* tscale, const void* tres_in,
                          const void* tdelta, void* tres_out, void* ty, int M, int d, cudaStream_t st) {
  rmsnorm_residual_jvp_kernel<VPL><<<(M + 7) / 8, 256, 0, st>>>(
      static_cast<const __nv_bfloat16*>(x), rstd, scale, tscale, static_cast<const __nv_bfloat16*>(tres_in),
      static_cast<const __nv_bfloat16*>(tdelta), static_cast<__nv_bfloat16*>(tres_out), static_cast<__nv_bfloat16*>(ty),
      M, d);
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_rmsnorm_residual_jvp(const void* x, const float* rstd, const float* scale, const float* tscale,
                                        const void* tres_in, const void* tdelta, void* tres_out, void* ty, int M, int d,
                                        void* stream) {
  if (!x || !rstd || !scale || !tscale || !tres_in || !ty || M <= 0 || d <= 0 || d > 2048) return PCV_ERR_ARG;
  if (tdelta && !tres_out) return PCV_ERR_ARG;
  if (d % 8 || !al16(x) || !al16(tres_in) || !al16(ty) || !al16(scale) || !al16(tscale) || (tdelta && !al16(tdelta)) ||
      (tres_out && !al16(tres_out)))
    return PCV_ERR_ALIGN;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch ((d / 8 + 31) / 32) {
    case 1: return launch_rms_jvp<1>(x, rstd, scale, tscale, tres_in, tdelta, tres_out, ty, M, d, st);
    case 2: return launch_rms_jvp<2>(x, rstd, scale, tscale, tres_in, tdelta, tres_out, ty, M, d, st);
    case 3: return launch_rms_jvp<3>(x, rstd, scale, tscale, tres_in, tdelta, tres_out, ty, M, d, st);
    case 4: return launch_rms_jvp<4>(x, rstd, scale, tscale, tres_in, tdelta, tres_out, ty, M, d, st);
    default: return launch_rms_jvp<8>(x, rstd, scale, tscale, tres_in, tdelta, tres_out, ty, M, d, st);
  }
}

extern "C" int pcv_swiglu_jvp(const void* gu, const void* tgu, void* th, int M, int hid, void* stream) {
  if (!gu || !tgu || !th || M <= 0 || hid <= 0) return PCV_ERR_ARG;
  if (hid % 8 || !al16(gu) || !al16(tgu) || !al16(th)) return PCV_ERR_ALIGN;
  const long total = static_cast<long>(M) * (hid / 8);
  swiglu_jvp_kernel<<<ew_blocks(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(gu), static_cast<const __nv_bfloat16*>(tgu), static_cast<__nv_bfloat16*>(th), M,
      hid);
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_act_jvp(const void* x, const void* tx, void* th, long long n, int mode, void* stream) {
  if (!x || !tx || !th || n <= 0 || (mode != 0 && mode != 1)) return PCV_ERR_ARG;
  if (n % 8 || !al16(x) || !al16(tx) || !al16(th)) return PCV_ERR_ALIGN;
  act_jvp_kernel<<<ew_blocks(n / 8, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(tx), static_cast<__nv_bfloat16*>(th), n / 8,
      mode);
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_ce_hessian_vec(const void* logits, const void* vec, void* hv, int M, int V, int ldl, float out_scale,
                                  void* stream) {
  if (!logits || !vec || !hv || M <= 0 || V <= 0 || V > ldl) return PCV_ERR_ARG;
  if (ldl % 8 || !al16(logits) || !al16(vec) || !al16(hv)) return PCV_ERR_ALIGN;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t row_bytes = static_cast<size_t>(ldl) * 2;
  if (row_bytes <= 110 * 1024) {  // two blocks per SM keep their rows in shared memory
    auto kern = ce_hessian_vec_kernel<true>;
    static bool attr = false;
    if (!attr) {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024);
      if (e != cudaSuccess) return static_cast<int>(e);
      attr = true;
    }
    kern<<<M, kHvThreads, row_bytes, st>>>(static_cast<const __nv_bfloat16*>(logits),
                                           static_cast<const __nv_bfloat16*>(vec), static_cast<__nv_bfloat16*>(hv), V,
                                           ldl, out_scale);
  } else {
    ce_hessian_vec_kernel<false><<<M, kHvThreads, 0, st>>>(static_cast<const __nv_bfloat16*>(logits),
                                                           static_cast<const __nv_bfloat16*>(vec),
                                                           static_cast<__nv_bfloat16*>(hv), V, ldl, out_scale);
  }
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_attn_jvp(const void* qkv, const void* tqkv, const int32_t* doc_start, const void* out,
                            const float* lse, void* tout, int B, int T, int n_heads, void* stream) {
  if (!qkv || !tqkv || !out || !lse || !tout || B <= 0 || T <= 0 || n_heads <= 0 || T % 128) return PCV_ERR_ARG;
  if (!al16(out) || !al16(tout)) return PCV_ERR_ALIGN;
  const int d = n_heads * 64;
  const uint64_t BT = static_cast<uint64_t>(B) * T;
  CUtensorMap tm, tm_t;
  int rc = get_tmap_2d_bf16(&tm, qkv, BT, 3 * d, 3 * d, 128);
  if (rc) return rc;
  rc = get_tmap_2d_bf16(&tm_t, tqkv, BT, 3 * d, 3 * d, 128);
  if (rc) return rc;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(attn_jvp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnJvpSmem::kBytes);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr = true;
  }
  dim3 grid(n_heads, T / 128, B);
  attn_jvp_kernel<<<grid, kJvpThreads, AttnJvpSmem::kBytes, static_cast<cudaStream_t>(stream)>>>(
      tm, tm_t, doc_start, static_cast<const __nv_bfloat16*>(out), lse, static_cast<__nv_bfloat16*>(tout), T, n_heads);
  count_launch();
  return cuda_rc();
}
