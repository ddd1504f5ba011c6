// gemm.cu — persistent, warp-specialised tcgen05/TMEM GEMM for sm_100a, fed by TMA.
//
// Replaces every `flax.linen.Dense` contraction on the reference's training hot path
// (reference: models/LM/transformer.py:110-134 GLU, :194-201 w_qkv, :246-253 w_out, :398-405 lm_head) and the
// dgrad / wgrad contractions XLA derives from them (train_lm.py:267 value_and_grad).
//
//   C[M,N] (=|+=)  sum_k A(m,k) * B(n,k)       bf16 operands, fp32 accumulation in tensor memory
//
// Operand storage (either operand, independently):
//   K-major  : stored [rows = M or N][K]  (contraction contiguous)        -> TMA box {64 k, rows}
//   MN-major : stored [K][cols = M or N]  (output dimension contiguous)   -> TMA boxes {64 mn, 64 k}
// With the reference's (in,out) `kernel` layout this gives, without any transposed weight copy:
//   forward  Y = X W        : A = X  K-major, B = W  MN-major
//   dgrad    dX = dY W^T    : A = dY K-major, B = W  K-major
//   wgrad    dW += X^T dY   : A = X  MN-major, B = dY MN-major, fp32 red.add epilogue (split-K)
//
// Structure (one CTA per SM, 256 threads):
//   warp 0   : TMA producer (one elected lane)      smem ring of kStages {A,B} tiles, mbarrier full/empty
//   warp 1   : MMA issuer (one elected lane)        tcgen05.mma 128 x BN x 16, accumulators double-buffered in TMEM
//   warp 2   : TMEM allocator
//   warps 4-7: epilogue                             tcgen05.ld -> registers -> global (bf16 / fp32 / fp32 red.add)
#include "sm100.cuh"
#include "pcv_internal.h"

#include <atomic>
#include <cstdlib>
#include <mutex>
#include <unordered_map>

namespace pcv {

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kGemmThreads = 256;
constexpr int kPairThreads = 384;  // CTA-pair kernel: producer, MMA, allocator, idle + 8 epilogue warps

template <int BN>
struct GemmCfg {
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr int kABytes = kBM * kBK * 2;  // 16 KB
  static constexpr int kBBytes = BN * kBK * 2;   // 32 / 16 KB
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarBytes = 256;
  static constexpr int kSmemBytes = kStages * kStageBytes + kBarBytes + 1024;  // +1024 for manual alignment
  static constexpr uint32_t kTmemCols = 2 * BN;                              // 512 or 256
};

struct GemmParams {
  void* C;
  int M, N, K;
  int ldc;
  int m_tiles, n_tiles, splits, kb_per_split, kb_total;
  float alpha;
  // EPI_BF16_ROPE only: interleaved-pair rotation of output columns [0, rope_cols) at position (row % T)
  const float* cos_sin;
  int T, rope_cols;
  // EPI_SWIGLU / EPI_DSWIGLU only (CTA-pair kernel): second tensor of the fused epilogue and the column distance
  // between the gate and the up half of a [M][2*hid] activation
  void* aux;
  int ld_aux, half_off;
  // CTA-pair kernel, tail of the persistent schedule: the tiles of the last, partial wave (index >= full_tiles) are cut
  // into two 256 x 128 halves when that lets every pair work on it (narrow = 1), instead of leaving pairs idle
  int full_tiles, narrow;
  // CTA-pair kernel: 1 = the launch order of the grid is the work queue (one cluster launched per work item; a running
  // pair takes its own blockIdx first and then the launches it cancels through cluster launch control), 0 = every pair
  // walks the fixed sequence pair, pair + npairs, ...
  int dynamic;
  // EPI_BF16_CE only (LM head fused with the first half of softmax cross-entropy): per (row, 128-column slot) the
  // maximum of the bf16-rounded logits and the sum of exp(logit - max), slot-major [slots][M]; columns >= ce_valid are pad
  float2* ce_ms;
  int ce_valid;
#ifdef PCV_MEASURE
  // -DPCV_MEASURE builds only (never the shipped library): per-phase cycle counters of the epilogue warps
  // (pcv_gemm_set_profile) and switches that leave out operand loads / epilogue stores (pcv_gemm_set_mode(100 + flags))
  unsigned int* prof;
  int dbg;
#endif
};
#ifdef PCV_MEASURE
#define PCV_DBG(p, bits) (((p).dbg & (bits)) != 0)
#define PCV_PROF(p) ((p).prof != nullptr)
#else
#define PCV_DBG(p, bits) false
#define PCV_PROF(p) false
#endif
constexpr int EPI_BF16_ROPE = 3;  // internal: PCV_EPI_BF16 + RoPE (pcv_gemm_bf16_rope)
constexpr int EPI_SWIGLU = 4;     // internal: [g | u] = X W_gu written to C, h = silu(g) * u written to aux
constexpr int EPI_DSWIGLU = 5;    // internal: dh = dM W_down^T kept in registers, [dg | du] written to C, gu read from aux
constexpr int EPI_BF16_CE = 6;    // internal: PCV_EPI_BF16 + per-(row, 128-column slot) softmax partials (pcv_gemm_bf16_ce)

// Work item -> (m tile, n tile, k split).  Tiles are visited in column-major order inside groups of kGroupM tile
// rows, so the CTAs that run concurrently share a few A tiles AND a few B tiles (L2 reuse for wide outputs such as
// the LM head, where a plain row-major order would stream all of B for every tile row).
constexpr int kGroupM = 16;
__device__ __forceinline__ void decode_work(const GemmParams& p, int w, int& m_blk, int& n_blk, int& split) {
  const int per_split = p.m_tiles * p.n_tiles;
  split = w / per_split;
  const int t = w - split * per_split;
  const int per_group = kGroupM * p.n_tiles;
  const int g = t / per_group;
  const int first_m = g * kGroupM;
  const int gm = min(kGroupM, p.m_tiles - first_m);
  const int r = t - g * per_group;
  m_blk = first_m + r % gm;
  n_blk = r / gm;
}


__device__ __forceinline__ unsigned int clk32() {
  unsigned int c;
  asm volatile("mov.u32 %0, %%clock;" : "=r"(c));
  return c;
}

// CTA-pair schedule: work index -> tile, k split and (for the halved tail tiles) which 128-column half.
struct PairWork {
  int m_blk, n_blk, split, ncol0, narrow;
};
__device__ __forceinline__ PairWork decode_pair_work(const GemmParams& p, int w) {
  PairWork r;
  r.ncol0 = 0;
  r.narrow = 0;
  if (p.narrow && w >= p.full_tiles) {
    const int u = w - p.full_tiles;
    r.narrow = 1;
    r.ncol0 = (u & 1) * 128;
    w = p.full_tiles + (u >> 1);
  }
  decode_work(p, w, r.m_blk, r.n_blk, r.split);
  return r;
}
__device__ __forceinline__ int pair_total_work(const GemmParams& p) {
  const int tiles = p.m_tiles * p.n_tiles * p.splits;
  return p.narrow ? p.full_tiles + 2 * (tiles - p.full_tiles) : tiles;
}

// One thread's 32 consecutive accumulator columns of one output row -> global memory.
template <int EPI>
__device__ __forceinline__ void store_row_chunk(const GemmParams& p, int row, int col0, const uint32_t* r) {
  if constexpr (EPI == EPI_BF16_ROPE) {
    // q and k columns: rotate the 16 (even, odd) pairs of this chunk in fp32, before the bf16 rounding
    // (apply_rotary_emb_complex_like, models/LM/embedding.py:28-67); v columns pass through.
    __nv_bfloat16* crow = reinterpret_cast<__nv_bfloat16*>(p.C) + static_cast<size_t>(row) * p.ldc + col0;
    const bool rot = col0 < p.rope_cols;
    const float4* cs = reinterpret_cast<const float4*>(p.cos_sin + (static_cast<size_t>(row % p.T) * 32 + ((col0 & 63) >> 1)) * 2);
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      if (col0 + v * 8 + 8 <= p.N) {
        float x[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) x[k] = __uint_as_float(r[v * 8 + k]) * p.alpha;
        if (rot) {
          const float4 c0 = __ldg(cs + 2 * v), c1 = __ldg(cs + 2 * v + 1);  // (cos, sin) x 2 pairs each
          const float a0 = x[0], b0 = x[1], a1 = x[2], b1 = x[3], a2 = x[4], b2 = x[5], a3 = x[6], b3 = x[7];
          x[0] = a0 * c0.x - b0 * c0.y; x[1] = b0 * c0.x + a0 * c0.y;
          x[2] = a1 * c0.z - b1 * c0.w; x[3] = b1 * c0.z + a1 * c0.w;
          x[4] = a2 * c1.x - b2 * c1.y; x[5] = b2 * c1.x + a2 * c1.y;
          x[6] = a3 * c1.z - b3 * c1.w; x[7] = b3 * c1.z + a3 * c1.w;
        }
        uint4 o;
        o.x = pack_bf16x2(x[0], x[1]);
        o.y = pack_bf16x2(x[2], x[3]);
        o.z = pack_bf16x2(x[4], x[5]);
        o.w = pack_bf16x2(x[6], x[7]);
        *reinterpret_cast<uint4*>(crow + v * 8) = o;
      }
    }
  } else if constexpr (EPI == PCV_EPI_BF16) {
    __nv_bfloat16* crow = reinterpret_cast<__nv_bfloat16*>(p.C) + static_cast<size_t>(row) * p.ldc + col0;
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      if (col0 + v * 8 + 8 <= p.N) {
        uint4 o;
        o.x = pack_bf16x2(__uint_as_float(r[v * 8 + 0]) * p.alpha, __uint_as_float(r[v * 8 + 1]) * p.alpha);
        o.y = pack_bf16x2(__uint_as_float(r[v * 8 + 2]) * p.alpha, __uint_as_float(r[v * 8 + 3]) * p.alpha);
        o.z = pack_bf16x2(__uint_as_float(r[v * 8 + 4]) * p.alpha, __uint_as_float(r[v * 8 + 5]) * p.alpha);
        o.w = pack_bf16x2(__uint_as_float(r[v * 8 + 6]) * p.alpha, __uint_as_float(r[v * 8 + 7]) * p.alpha);
        *reinterpret_cast<uint4*>(crow + v * 8) = o;
      }
    }
  } else {
    float* crow = reinterpret_cast<float*>(p.C) + static_cast<size_t>(row) * p.ldc + col0;
#pragma unroll
    for (int v = 0; v < 8; ++v) {
      if (col0 + v * 4 + 4 <= p.N) {
        const float x0 = __uint_as_float(r[v * 4 + 0]) * p.alpha, x1 = __uint_as_float(r[v * 4 + 1]) * p.alpha;
        const float x2 = __uint_as_float(r[v * 4 + 2]) * p.alpha, x3 = __uint_as_float(r[v * 4 + 3]) * p.alpha;
        if constexpr (EPI == PCV_EPI_F32_ATOMIC) {
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(crow + v * 4), "f"(x0), "f"(x1),
                       "f"(x2), "f"(x3)
                       : "memory");
        } else {
          *reinterpret_cast<float4*>(crow + v * 4) = make_float4(x0, x1, x2, x3);
        }
      }
    }
  }
}

// ---- CTA-pair epilogue helpers: 32 accumulator columns of one row (fp32 bits in r[32]) -> 4 x 16 B of bf16.
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void pack32_bf16(const uint32_t* r, float alpha, uint4* q) {
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    q[v].x = pack_bf16x2(__uint_as_float(r[v * 8 + 0]) * alpha, __uint_as_float(r[v * 8 + 1]) * alpha);
    q[v].y = pack_bf16x2(__uint_as_float(r[v * 8 + 2]) * alpha, __uint_as_float(r[v * 8 + 3]) * alpha);
    q[v].z = pack_bf16x2(__uint_as_float(r[v * 8 + 4]) * alpha, __uint_as_float(r[v * 8 + 5]) * alpha);
    q[v].w = pack_bf16x2(__uint_as_float(r[v * 8 + 6]) * alpha, __uint_as_float(r[v * 8 + 7]) * alpha);
  }
}
__device__ __forceinline__ void pack32_f32(const uint32_t* r, float alpha, uint4* q) {  // 32 fp32 -> 8 x 16 B
#pragma unroll
  for (int v = 0; v < 8; ++v) {
    q[v].x = __float_as_uint(__uint_as_float(r[v * 4 + 0]) * alpha);
    q[v].y = __float_as_uint(__uint_as_float(r[v * 4 + 1]) * alpha);
    q[v].z = __float_as_uint(__uint_as_float(r[v * 4 + 2]) * alpha);
    q[v].w = __float_as_uint(__uint_as_float(r[v * 4 + 3]) * alpha);
  }
}
// q / k columns of the fused QKV projection: rotate the 16 (even, odd) pairs of a 32-column chunk in fp32, before the
// bf16 rounding (apply_rotary_emb_complex_like, models/LM/embedding.py:28-67); cs = 8 x (cos, sin, cos, sin).
__device__ __forceinline__ void rope32_bf16(const uint32_t* r, const float4* cs, bool rot, float alpha, uint4* q) {
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    float x[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] = __uint_as_float(r[v * 8 + k]) * alpha;
    if (rot) {
      const float4 c0 = cs[2 * v], c1 = cs[2 * v + 1];
      const float a0 = x[0], b0 = x[1], a1 = x[2], b1 = x[3], a2 = x[4], b2 = x[5], a3 = x[6], b3 = x[7];
      x[0] = a0 * c0.x - b0 * c0.y; x[1] = b0 * c0.x + a0 * c0.y;
      x[2] = a1 * c0.z - b1 * c0.w; x[3] = b1 * c0.z + a1 * c0.w;
      x[4] = a2 * c1.x - b2 * c1.y; x[5] = b2 * c1.x + a2 * c1.y;
      x[6] = a3 * c1.z - b3 * c1.w; x[7] = b3 * c1.z + a3 * c1.w;
    }
    q[v].x = pack_bf16x2(x[0], x[1]);
    q[v].y = pack_bf16x2(x[2], x[3]);
    q[v].z = pack_bf16x2(x[4], x[5]);
    q[v].w = pack_bf16x2(x[6], x[7]);
  }
}
// Fused SwiGLU (GLU, models/LM/transformer.py:110-134).  Forward: g and u are rounded to bf16 first (they are Dense
// outputs in the reference) and h = silu(g) * u is computed from the rounded values, exactly as pcv_swiglu_fwd does.
__device__ __forceinline__ void swiglu32(const uint32_t* g, const uint32_t* u, uint4* qg, uint4* qu, uint4* qh) {
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    uint32_t* gw = reinterpret_cast<uint32_t*>(&qg[v]);
    uint32_t* uw = reinterpret_cast<uint32_t*>(&qu[v]);
    uint32_t* hw = reinterpret_cast<uint32_t*>(&qh[v]);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      gw[k] = pack_bf16x2(__uint_as_float(g[v * 8 + 2 * k]), __uint_as_float(g[v * 8 + 2 * k + 1]));
      uw[k] = pack_bf16x2(__uint_as_float(u[v * 8 + 2 * k]), __uint_as_float(u[v * 8 + 2 * k + 1]));
      const float2 fh = swiglu_fwd_pair(unpack_bf16x2(gw[k]), unpack_bf16x2(uw[k]));
      hw[k] = pack_bf16x2(fh.x, fh.y);
    }
  }
}
// Backward: dh (32 accumulator columns, rounded to bf16 as the reference's transposed Dense output is) with the saved
// gate / up values in qg / qu, which are overwritten by dg / du; the arithmetic is pcv_swiglu_bwd's.
__device__ __forceinline__ void dswiglu32(const uint32_t* r, float alpha, uint4* qg, uint4* qu) {
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    uint32_t* gw = reinterpret_cast<uint32_t*>(&qg[v]);
    uint32_t* uw = reinterpret_cast<uint32_t*>(&qu[v]);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 acc = __fmul2_rn(make_float2(__uint_as_float(r[v * 8 + 2 * k]), __uint_as_float(r[v * 8 + 2 * k + 1])),
                                    make_float2(alpha, alpha));
      const float2 fd = unpack_bf16x2(pack_bf16x2(acc.x, acc.y));
      float2 dg, du;
      swiglu_bwd_pair(unpack_bf16x2(gw[k]), unpack_bf16x2(uw[k]), fd, dg, du);
      gw[k] = pack_bf16x2(dg.x, dg.y);
      uw[k] = pack_bf16x2(du.x, du.y);
    }
  }
}

// Softmax partials of one row's 64 columns, from the bf16-ROUNDED logits q (8 x 16 bytes) — the values the loss is
// defined on (the reference rounds the Dense output to bf16 before _loss_and_acc promotes it to fp32).  Columns
// [0, n_valid) are real, the rest pad.  Merges (m, s) = (max, sum exp(x - max)) into the running pair.
__device__ __forceinline__ void ce_slab_stats(const uint4* q, int n_valid, float& m_run, float& s_run) {
  constexpr float kL2e = 1.4426950408889634f;
  float x[64];
#pragma unroll
  for (int v = 0; v < 8; ++v) {
    const uint32_t w4[4] = {q[v].x, q[v].y, q[v].z, q[v].w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 t = unpack_bf16x2(w4[k]);
      x[v * 8 + 2 * k] = t.x;
      x[v * 8 + 2 * k + 1] = t.y;
    }
  }
  if (n_valid < 64) {
#pragma unroll
    for (int j = 0; j < 64; ++j)
      if (j >= n_valid) x[j] = -3.0e38f;
  }
  float m = x[0];
#pragma unroll
  for (int j = 1; j < 64; ++j) m = fmaxf(m, x[j]);
  const float m_new = fmaxf(m_run, m);
  const float mb = m_new * kL2e;
  float s0 = 0.f, s1 = 0.f;
#pragma unroll
  for (int j = 0; j < 64; j += 2) {
    float e0, e1;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(fmaf(x[j], kL2e, -mb)));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(fmaf(x[j + 1], kL2e, -mb)));
    s0 += e0;
    s1 += e1;
  }
  float resc;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(resc) : "f"((m_run - m_new) * kL2e));  // 0 when m_run = -3e38 (first slab)
  s_run = s_run * resc + s0 + s1;
  m_run = m_new;
}

template <int BN, bool A_MN, bool B_MN, int EPI>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                 const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smem_a = smem_base;
  const uint32_t smem_b = smem_base + Cfg::kStages * Cfg::kABytes;
  const uint32_t bar_base = smem_base + Cfg::kStages * Cfg::kStageBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (Cfg::kStages + s); };
  auto tmem_full_bar = [&](int s) { return bar_base + 8u * (2 * Cfg::kStages + s); };
  auto tmem_empty_bar = [&](int s) { return bar_base + 8u * (2 * Cfg::kStages + 2 + s); };
  const uint32_t tmem_ptr_smem = bar_base + 8u * (2 * Cfg::kStages + 4);
  volatile uint32_t* tmem_ptr_generic =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_ptr_smem - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tmem_full_bar(s), 1);
      mbar_init(tmem_empty_bar(s), 4);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc<Cfg::kTmemCols>(tmem_ptr_smem);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_generic;
  pdl_wait();     // everything above ran under the previous kernel's tail; from here on its results are read
  pdl_trigger();  // (after the wait: a dependent that starts early then knows every kernel before this one has completed)

  const int total_work = p.m_tiles * p.n_tiles * p.splits;

  if (warp == 0) {
    if (lane == 0) {
      // ===================== TMA producer =====================
      int stage = 0;
      uint32_t phase = 0;
      for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
        int m_blk, n_blk, split;
        decode_work(p, w, m_blk, n_blk, split);
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.kb_total);
        const int m0 = m_blk * kBM, n0 = n_blk * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          mbar_expect_tx(full_bar(stage), Cfg::kStageBytes);
          const uint32_t sa = smem_a + stage * Cfg::kABytes;
          const uint32_t sb = smem_b + stage * Cfg::kBBytes;
          const int k0 = kb * kBK;
          if constexpr (!A_MN) {
            tma_load_2d(sa, &tma_a, full_bar(stage), k0, m0);
          } else {
#pragma unroll
            for (int i = 0; i < kBM / 64; ++i) tma_load_2d(sa + i * (64 * kBK * 2), &tma_a, full_bar(stage), m0 + 64 * i, k0);
          }
          if constexpr (!B_MN) {
            tma_load_2d(sb, &tma_b, full_bar(stage), k0, n0);
          } else {
#pragma unroll
            for (int i = 0; i < BN / 64; ++i) tma_load_2d(sb + i * (64 * kBK * 2), &tma_b, full_bar(stage), n0 + 64 * i, k0);
          }
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      // ===================== MMA issuer =====================
      constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN, A_MN, B_MN);
      // per-UMMA_K(16) advance inside a stage: K-major 32 B along the swizzled row; MN-major 16 rows of 128 B
      constexpr uint32_t a_kstep = A_MN ? 16 * 128 : 32;
      constexpr uint32_t b_kstep = B_MN ? 16 * 128 : 32;
      constexpr uint32_t a_lbo = A_MN ? 64 * kBK * 2 : 16, b_lbo = B_MN ? 64 * kBK * 2 : 16;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int w = blockIdx.x; w < total_work; w += gridDim.x, ++it) {
        const int split = w / (p.n_tiles * p.m_tiles);
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.kb_total);
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(tmem_empty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_a + stage * Cfg::kABytes;
          const uint32_t sb = smem_b + stage * Cfg::kBBytes;
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            const uint64_t da = umma_smem_desc(sa + k * a_kstep, a_lbo, 1024);
            const uint64_t db = umma_smem_desc(sb + k * b_kstep, b_lbo, 1024);
            umma_bf16_ss(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          tc_commit(empty_bar(stage));  // smem slot reusable once these MMAs have read it
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1u; }
        }
        tc_commit(tmem_full_bar(acc));  // accumulator complete -> epilogue
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int ew = warp & 3;  // TMEM lane quarter this warp may access
    int it = 0;
    for (int w = blockIdx.x; w < total_work; w += gridDim.x, ++it) {
      int m_blk, n_blk, split_unused;
      decode_work(p, w, m_blk, n_blk, split_unused);
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(tmem_full_bar(acc), acc_phase);
      tc_fence_after();
      const int row = m_blk * kBM + ew * 32 + lane;
      const int n0 = n_blk * BN;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(taddr + c * 32, r);
        tmem_ld_wait();
        const int col0 = n0 + c * 32;
        if (row < p.M && col0 < p.N) store_row_chunk<EPI>(p, row, col0, r);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tmem_empty_bar(acc));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}


// ------------------------------------------------------------------------------------------------ CTA-pair kernel
// Two CTAs (a cluster of 2 = one TPC) cooperate on a 256 x 256 output tile with tcgen05 cta_group::2:
//   CTA r stages rows [r*128, r*128+128) of the A tile and rows [r*128, r*128+128) of the B tile (half of N each);
//   one thread of the leader CTA issues 256 x 256 x 16 MMAs that read A/B from both CTAs' shared memory and
//   accumulate into both CTAs' tensor memory (128 lanes each); each CTA's epilogue warps drain their own half.
// Per k-block a CTA pulls 32 KB for 2*128*256*64 flop: 128 flop per L2 byte against 85 for the single-CTA
// 128 x 256 tile — the single-CTA kernel saturates the L2 -> SM fabric (~12 TB/s) near 1.1 PFLOP/s.
#ifndef PCV_SWIGLU_BOXES
#define PCV_SWIGLU_BOXES 1
#endif
#ifndef PCV_DSWIGLU_STAGES
#define PCV_DSWIGLU_STAGES 4
#endif
#ifndef PCV_GEMM_DYNAMIC_DEFAULT
#define PCV_GEMM_DYNAMIC_DEFAULT 0
#endif
constexpr bool pair_epi_fused(int epi) { return epi == EPI_SWIGLU || epi == EPI_DSWIGLU; }
template <int EPI>
struct PairCfgT {
  static constexpr bool kFused = pair_epi_fused(EPI);  // SwiGLU forward / backward in the epilogue
  static constexpr int kBN = 256;
  static constexpr int kTileM = 256;
  // SwiGLU forward keeps all six stages and one box per epilogue warp: its epilogue has slack (the warps wait for the
  // accumulator ~40 % of the time), while the ring feeds a mainloop that already runs at the L2 -> SM limit and has
  // to absorb the epilogue's store traffic (6 stages / 1 box: 142 us, 5 stages / 2 boxes: 150 us at the 420M shape).
  // SwiGLU backward gives up two stages for three boxes: the saved gate and up slabs arrive by TMA while the previous
  // slab is still being worked on, and dg / du leave through the third
  static constexpr int kStages = EPI == EPI_DSWIGLU ? PCV_DSWIGLU_STAGES : (kFused ? (PCV_SWIGLU_BOXES == 1 ? 6 : 5) : 6);
  static constexpr int kABytes = 128 * kBK * 2;  // this CTA's 128 rows of A
  static constexpr int kBBytes = 128 * kBK * 2;  // this CTA's 128 of the 256 B rows
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarBytes = 1024;  // [0, 256) pipeline barriers, [256, 384) work-queue barriers, [512, 768) answers
  static constexpr int kWorkSlots = 16;   // ring of cluster-launch-control answers (see next_work)
  static constexpr int kEpiWarps = 8;
  static constexpr int kEpiBoxBytes = 32 * 128;  // 32 rows x 128 B (one 128B-swizzled TMA box)
  static constexpr int kEpiBoxes = EPI == EPI_DSWIGLU ? 3 : (kFused ? PCV_SWIGLU_BOXES : 1);
  static constexpr int kEpiBufBytes = kEpiBoxes * kEpiBoxBytes;  // per epilogue warp
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiWarps * kEpiBufBytes + kBarBytes + 1024;
  static constexpr uint32_t kTmemCols = 512;  // two 256-column accumulators
};
using PairCfg = PairCfgT<0>;
static_assert(PairCfgT<0>::kSmemBytes <= 232448 && PairCfgT<EPI_SWIGLU>::kSmemBytes <= 232448 &&
                  PairCfgT<EPI_DSWIGLU>::kSmemBytes <= 232448,
              "CTA-pair GEMM shared memory exceeds 227 KB");

template <bool A_MN, bool B_MN, int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPairThreads, 1)
gemm_bf16_pair_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                      const __grid_constant__ CUtensorMap tma_bn, const __grid_constant__ CUtensorMap tma_c,
                      const __grid_constant__ CUtensorMap tma_c2, const __grid_constant__ CUtensorMap tma_x,
                      const GemmParams p) {
  using Cfg = PairCfgT<EPI>;
  constexpr int BN = Cfg::kBN;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smem_a = smem_base;
  const uint32_t smem_b = smem_base + Cfg::kStages * Cfg::kABytes;
  const uint32_t smem_epi = smem_base + Cfg::kStages * Cfg::kStageBytes;  // 1024-aligned: stages are 32 KB
  const uint32_t bar_base = smem_epi + Cfg::kEpiWarps * Cfg::kEpiBufBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (Cfg::kStages + s); };
  auto tmem_full_bar = [&](int s) { return bar_base + 8u * (2 * Cfg::kStages + s); };
  auto tmem_empty_bar = [&](int s) { return bar_base + 8u * (2 * Cfg::kStages + 2 + s); };
  const uint32_t tmem_ptr_smem = bar_base + 8u * (2 * Cfg::kStages + 4);
  volatile uint32_t* tmem_ptr_generic =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_ptr_smem - smem_u32(smem_raw)));
  auto work_bar = [&](int k) { return bar_base + 256u + 8u * (k & (Cfg::kWorkSlots - 1)); };    // (1; tx: 16 bytes)
  auto work_slot = [&](int k) { return bar_base + 512u + 16u * (k & (Cfg::kWorkSlots - 1)); };  // 16-byte answers

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1;
  const int npairs = gridDim.x >> 1;
  // The k-th work item of this pair (k = 0: its own blockIdx).  Static schedule: pair + k * npairs.  Dynamic schedule
  // (p.dynamic): the answer to the cluster-launch-control request the producer issued while it loaded item k - 1 — the
  // index of a cluster whose launch this pair cancelled, or "nothing left".  Every role of both CTAs reads the answers
  // itself; they sit in a ring of kWorkSlots (item k -> slot k mod 16; the barrier's phase counts turns of the ring).
  // The producer is at most kStages items ahead of the MMA issuer (one k-block per item at worst), the MMA issuer at
  // most two accumulators ahead of the slowest epilogue warp of either CTA, so no reader is ever 16 items behind the
  // request that reuses its slot.
  auto next_work = [&](int k_next, int w_cur, int& w) -> bool {
    if (!p.dynamic) {
      w = w_cur + npairs;
      return w < pair_total_work(p);
    }
    mbar_wait(work_bar(k_next), static_cast<uint32_t>((k_next - 1) / Cfg::kWorkSlots) & 1u);
    int x, y, z;
    if (!clc_query(work_slot(k_next), x, y, z)) return false;
    w = x >> 1;
    return true;
  };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    tma_prefetch_desc(&tma_bn);
    tma_prefetch_desc(&tma_c);
    tma_prefetch_desc(&tma_c2);
    tma_prefetch_desc(&tma_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(full_bar(s), 1);   // leader: its producer's arrive.expect_tx (bytes of BOTH CTAs)
      mbar_init(empty_bar(s), 1);  // multicast tcgen05.commit
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tmem_full_bar(s), 1);   // multicast tcgen05.commit
      mbar_init(tmem_empty_bar(s), 16);  // leader: 8 epilogue warps of each CTA
    }
    if constexpr (EPI == EPI_DSWIGLU) {  // one per epilogue warp: its gate / up slab has landed
      for (int s = 0; s < Cfg::kEpiWarps; ++s) mbar_init(bar_base + 8u * (2 * Cfg::kStages + 5 + s), 1);
    }
    for (int s = 0; s < Cfg::kWorkSlots; ++s) mbar_init(work_bar(s), 1);  // this CTA's producer: arrive.expect_tx(16)
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_pair<Cfg::kTmemCols>(tmem_ptr_smem);
  }
  tc_fence_before();
  cluster_sync_all();  // barriers of both CTAs initialised, both TMEM allocations done
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_generic;
  pdl_wait();     // everything above ran under the previous kernel's tail; from here on its results are read
  pdl_trigger();  // (after the wait: a dependent that starts early then knows every kernel before this one has completed)

  const int total_work = pair_total_work(p);

  if (warp == 0) {
    if (lane == 0) {
      // ===================== TMA producer (both CTAs; completion lands on the leader's barrier) =====================
      int stage = 0;
      uint32_t phase = 0;
      int w = pair;
      for (int k = 0; w < total_work; ++k) {
        if (p.dynamic) {
          // ask for the item after this one now (the answer to the request for THIS item was a success, so another
          // request is allowed): it is there long before any role needs it.  Both CTAs arm their own barrier; the
          // leader's request is answered in both CTAs' shared memory
          mbar_expect_tx(work_bar(k + 1), 16);
          if (rank == 0) clc_try_cancel_cluster(work_slot(k + 1), work_bar(k + 1));
        }
        const PairWork wk = decode_pair_work(p, w);
        const int kb0 = wk.split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.kb_total);
        const int m0 = wk.m_blk * Cfg::kTileM + static_cast<int>(rank) * 128;
        // this CTA's share of the B tile: 128 of 256 rows, or 64 of 128 for a halved tail tile.  Fused SwiGLU forward:
        // the tile is 128 gate columns (staged by CTA 0) next to the matching 128 up columns (staged by CTA 1), so
        // every epilogue thread holds g and u of the same hidden unit
        const int n0 = (EPI == EPI_SWIGLU) ? wk.n_blk * 128 + static_cast<int>(rank) * p.half_off
                       : wk.narrow         ? wk.n_blk * BN + wk.ncol0 + static_cast<int>(rank) * 64
                                           : wk.n_blk * BN + static_cast<int>(rank) * 128;
        uint32_t tx = wk.narrow ? 2 * (Cfg::kABytes + Cfg::kBBytes / 2) : 2 * Cfg::kStageBytes;
        // measurement only: leave out the B (64) / A (128) operand loads to see how much of the mainloop is the L2 -> SM feed
        const bool skip_b = PCV_DBG(p, 64), skip_a = PCV_DBG(p, 128) && !skip_b;  // never both: nothing would arm the barrier
        if (skip_b) tx -= wk.narrow ? Cfg::kBBytes : 2 * Cfg::kBBytes;
        if (skip_a) tx -= 2 * Cfg::kABytes;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          if (rank == 0) mbar_expect_tx(full_bar(stage), tx);
          const uint32_t bar = mapa_shared(full_bar(stage), 0);
          const uint32_t sa = smem_a + stage * Cfg::kABytes;
          const uint32_t sb = smem_b + stage * Cfg::kBBytes;
          const int k0 = kb * kBK;
          if (skip_a) {  // (with both skipped the leader's arrive.expect_tx of 0 bytes completes the phase by itself)
          } else if constexpr (!A_MN) {
            tma_load_2d_pair(sa, &tma_a, bar, k0, m0);
          } else {
#pragma unroll
            for (int i = 0; i < 2; ++i) tma_load_2d_pair(sa + i * (64 * kBK * 2), &tma_a, bar, m0 + 64 * i, k0);
          }
          if (skip_b) {
          } else if constexpr (!B_MN) {
            if (wk.narrow) tma_load_2d_pair(sb, &tma_bn, bar, k0, n0);
            else tma_load_2d_pair(sb, &tma_b, bar, k0, n0);
          } else {
            tma_load_2d_pair(sb, &tma_b, bar, n0, k0);
            if (!wk.narrow) tma_load_2d_pair(sb + 64 * kBK * 2, &tma_b, bar, n0 + 64, k0);
          }
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1u; }
        }
        if (!next_work(k + 1, w, w)) break;
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      // ===================== MMA issuer (leader CTA only) =====================
      constexpr uint32_t idesc_full = umma_idesc_bf16(256, BN, A_MN, B_MN);
      constexpr uint32_t idesc_half = umma_idesc_bf16(256, BN / 2, A_MN, B_MN);
      constexpr uint32_t a_kstep = A_MN ? 16 * 128 : 32;
      constexpr uint32_t b_kstep = B_MN ? 16 * 128 : 32;
      constexpr uint32_t a_lbo = A_MN ? 64 * kBK * 2 : 16, b_lbo = B_MN ? 64 * kBK * 2 : 16;
      int stage = 0;
      uint32_t phase = 0;
      int w = pair;
      for (int it = 0; w < total_work; ++it) {
        const PairWork wk = decode_pair_work(p, w);
        const int kb0 = wk.split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.kb_total);
        const uint32_t idesc = wk.narrow ? idesc_half : idesc_full;
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(tmem_empty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_a + stage * Cfg::kABytes;
          const uint32_t sb = smem_b + stage * Cfg::kBBytes;
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            const uint64_t da = umma_smem_desc(sa + k * a_kstep, a_lbo, 1024);
            const uint64_t db = umma_smem_desc(sb + k * b_kstep, b_lbo, 1024);
            umma_bf16_ss_pair(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          tc_commit_pair(empty_bar(stage), 3);  // frees this smem slot in both CTAs
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1u; }
        }
        tc_commit_pair(tmem_full_bar(acc), 3);  // accumulator complete -> both CTAs' epilogues
        if (!next_work(it + 1, w, w)) break;
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ===================== epilogue (each CTA drains its own 128 accumulator rows) =====================
    // 8 warps: warp (ew, hh) owns TMEM lanes [32 ew, 32 ew + 32) (its row quarter) and column half hh of the tile.
    // Results leave through a per-warp 32 x 128 B staging box in shared memory and one TMA store (fp32 reduce-add for
    // the wgrad epilogue) per box: whole 128-byte lines reach L2 instead of one 16-byte piece per row per instruction.
    const int ew = warp & 3;
    const int hh = (warp - 4) >> 2;
    const uint32_t buf = smem_epi + static_cast<uint32_t>(warp - 4) * Cfg::kEpiBufBytes;
    const uint32_t sw = lane & 7;
    // with two boxes per warp (fused SwiGLU forward) consecutive puts alternate between them: a box is rewritten only
    // once the store issued two puts earlier has read it, so one store is always in flight behind the math
    uint32_t put_box = 0;
    unsigned int pf0 = 0, pf1 = 0, pf2 = 0, pf3 = 0, pt = 0;  // profile: wait accumulator | wait operands | math | stores
    const bool prof = PCV_PROF(p);
    const unsigned int pstart = clk32();
    auto put = [&](const CUtensorMap* m, const uint4* q, int c0, int r0) {
      if (PCV_DBG(p, 1)) return;
      const uint32_t bx = buf + put_box * Cfg::kEpiBoxBytes;
      if (lane == 0) tma_store_wait_read<Cfg::kEpiBoxes - 1>();  // the store that last used this box has read it
      __syncwarp();
#pragma unroll
      for (int v = 0; v < 8; ++v) st_shared_v4(bx + lane * 128 + ((v ^ sw) << 4), q[v]);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0 && !PCV_DBG(p, 32)) {
        if constexpr (EPI == PCV_EPI_F32_ATOMIC) tma_reduce_add_2d(m, bx, c0, r0);
        else tma_store_2d(m, bx, c0, r0);
        tma_store_commit();
      }
      if constexpr (Cfg::kEpiBoxes == 2) put_box ^= 1u;
    };
    if constexpr (EPI == EPI_DSWIGLU) {
      // SwiGLU backward in the epilogue of the fc2 dgrad GEMM.  Per warp and tile: two slabs of 32 rows x 64 columns.
      // The saved gate and up values of a slab arrive by TMA in the warp's two boxes (requested as soon as the previous
      // slab's stores have read the boxes, i.e. while the MMAs of this tile still run), every lane reads its own row,
      // turns (dh, g, u) into (dg, du) in place and the two boxes leave again by TMA store.
      const uint32_t box_g = buf, box_u = buf + Cfg::kEpiBoxBytes, box_o = buf + 2 * Cfg::kEpiBoxBytes;
      const uint32_t row_g = box_g + lane * 128, row_u = box_u + lane * 128, row_o = box_o + lane * 128;
      const uint32_t ld_bar = bar_base + 8u * (2 * Cfg::kStages + 5 + (warp - 4));
      uint32_t ld_phase = 0;
      // slab (w, sl) of this warp: first column / first row, and whether any of it lies inside the matrix
      auto coords = [&](int w, int sl, int& c, int& r) -> bool {
        const PairWork wk = decode_pair_work(p, w);
        r = wk.m_blk * Cfg::kTileM + static_cast<int>(rank) * 128 + ew * 32;
        c = wk.n_blk * BN + hh * 128 + sl * 64;
        return c < p.N && r < p.M;
      };
      // first slab at or after slab sl of this pair's k-th work item (= w) that exists; false: no work is left
      auto seek = [&](int& k, int& w, int& sl, int& c, int& r) -> bool {
        while (true) {
          if (coords(w, sl, c, r)) return true;
          if (sl == 0) sl = 1;
          else {
            sl = 0;
            if (!next_work(k + 1, w, w)) return false;
            ++k;
          }
        }
      };
      auto request = [&](int c, int r) {  // lane 0
        if (PCV_DBG(p, 8)) return;
        mbar_expect_tx(ld_bar, 2 * Cfg::kEpiBoxBytes);
        tma_load_2d(box_g, &tma_x, ld_bar, c, r);
        tma_load_2d(box_u, &tma_bn, ld_bar, c, r);
      };
      int nk = 0, nw = pair, nsl = 0, nc = 0, nr = 0;  // the slab whose gate / up values are on their way
      if (seek(nk, nw, nsl, nc, nr) && lane == 0) request(nc, nr);
      int w = pair;
      for (int it = 0; w < total_work; ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + acc * BN + hh * 128;
        int c0, c1, row0;
        const bool has0 = coords(w, 0, c0, row0);
        const bool has1 = coords(w, 1, c1, row0);
        if (prof) pt = clk32();
        mbar_wait(tmem_full_bar(acc), acc_phase);
        tc_fence_after();
        if (prof) { const unsigned int t = clk32(); pf0 += t - pt; pt = t; }
        auto release = [&]() {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(mapa_shared(tmem_empty_bar(acc), 0));
        };
        if (!has0 && !has1) release();
#pragma unroll
        for (int sl = 0; sl < 2; ++sl) {
          if (sl == 0 ? !has0 : !has1) continue;
          const bool last = (sl == 1) || !has1;
          uint4 pg[8], pu[8];
          if (!PCV_DBG(p, 8)) {
            mbar_wait(ld_bar, ld_phase);
            ld_phase ^= 1u;
          }
          if (prof) { const unsigned int t = clk32(); pf1 += t - pt; pt = t; }
#pragma unroll
          for (int v = 0; v < 8; ++v) {
            pg[v] = ld_shared_v4(row_g + ((v ^ sw) << 4));
            pu[v] = ld_shared_v4(row_u + ((v ^ sw) << 4));
          }
          __syncwarp();
          // both input boxes are in registers: the slab after this one can be on its way during the math
          nk = it;
          nw = w;
          nsl = sl;
          bool more = true;
          if (nsl == 0) nsl = 1;
          else {
            nsl = 0;
            more = next_work(nk + 1, nw, nw);
            ++nk;
          }
          if (more && seek(nk, nw, nsl, nc, nr) && lane == 0) request(nc, nr);
          if (prof) { const unsigned int t = clk32(); pf1 += t - pt; pt = t; }  // lds + issue of the next slab's loads
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            uint32_t r[32];
            tmem_ld_32x32b_x32(taddr + sl * 64 + half * 32, r);
            tmem_ld_wait();
            if (last && half == 1) release();
            if (PCV_DBG(p, 4)) {
              pack32_bf16(r, p.alpha, pg + 4 * half);
              pack32_bf16(r, 2.f, pu + 4 * half);
            } else {
              dswiglu32(r, p.alpha, pg + 4 * half, pu + 4 * half);
            }
          }
          if (prof) { const unsigned int t = clk32(); pf2 += t - pt; pt = t; }
          // dg, then du, through the third box
#pragma unroll
          for (int o = 0; o < 2; ++o) {
            if (lane == 0) tma_store_wait_read<0>();  // the previous store has read the box
            __syncwarp();
#pragma unroll
            for (int v = 0; v < 8; ++v) st_shared_v4(row_o + ((v ^ sw) << 4), o == 0 ? pg[v] : pu[v]);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && !PCV_DBG(p, 1)) {
              tma_store_2d(o == 0 ? &tma_c : &tma_c2, box_o, sl == 0 ? c0 : c1, row0);
              tma_store_commit();
            }
          }
          if (prof) { const unsigned int t = clk32(); pf3 += t - pt; pt = t; }
        }
        if (!next_work(it + 1, w, w)) break;
      }
    } else {
    int w = pair;
    for (int it = 0; w < total_work; ++it) {
      const PairWork wk = decode_pair_work(p, w);
      const int m_blk = wk.m_blk, n_blk = wk.n_blk;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      const int row0 = m_blk * Cfg::kTileM + static_cast<int>(rank) * 128 + ew * 32;  // first row of this warp's boxes
      // this warp's columns: half of a full tile (two 64-column slabs), a quarter of a halved tail tile (one slab)
      const int nslab = wk.narrow ? 1 : 2;
      const int tcol = wk.narrow ? hh * 64 : hh * 128;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + acc * BN + tcol;
      auto release = [&]() {  // this warp has read its part of the accumulator: hand it back to the MMA issuer
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(mapa_shared(tmem_empty_bar(acc), 0));
      };
      if constexpr (EPI == EPI_SWIGLU) {
        // accumulator columns [0,128) = gate, [128,256) = up of hidden units [n_blk*128, +128); this warp: 64 of them
        const int n0 = n_blk * 128 + hh * 64;
        const uint32_t tg = taddr - tcol + hh * 64;
        if (prof) pt = clk32();
        mbar_wait(tmem_full_bar(acc), acc_phase);
        tc_fence_after();
        if (prof) { const unsigned int t = clk32(); pf0 += t - pt; pt = t; }
        uint4 qg[8], qu[8], qh[8];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t g[32], u[32];
          tmem_ld_32x32b_x32(tg + c * 32, g);
          tmem_ld_32x32b_x32(tg + 128 + c * 32, u);
          tmem_ld_wait();
          if (c == 1) release();
          if (PCV_DBG(p, 4)) {
            pack32_bf16(g, 1.f, qg + 4 * c);
            pack32_bf16(u, 1.f, qu + 4 * c);
            pack32_bf16(u, 2.f, qh + 4 * c);
          } else {
            swiglu32(g, u, qg + 4 * c, qu + 4 * c, qh + 4 * c);
          }
        }
        if (prof) { const unsigned int t = clk32(); pf2 += t - pt; pt = t; }
        if (n0 < p.N) {
          put(&tma_c, qg, n0, row0);
          put(&tma_c2, qu, n0, row0);
          if (!PCV_DBG(p, 2)) put(&tma_x, qh, n0, row0);
        }
        if (prof) { const unsigned int t = clk32(); pf3 += t - pt; pt = t; }
      } else if constexpr (EPI == EPI_BF16_ROPE || EPI == PCV_EPI_BF16 || EPI == EPI_BF16_CE) {
        const int n0 = n_blk * BN + wk.ncol0 + tcol;
        float ce_m = -3.0e38f, ce_s = 0.f;  // EPI_BF16_CE: this row's partials over the warp's 128 columns
        // (cos, sin) of this row's position for the 32 pairs of a head: the same for every head of the tile, so they
        // are fetched once per tile, before the accumulator is waited for
        const bool rot = (EPI == EPI_BF16_ROPE) && n0 < p.rope_cols;
        float4 cs[16];
        if constexpr (EPI == EPI_BF16_ROPE) {
          if (rot) {
            const int row = min(row0 + lane, p.M - 1);
            const float4* src = reinterpret_cast<const float4*>(p.cos_sin + static_cast<size_t>(row % p.T) * 64);
#pragma unroll
            for (int v = 0; v < 16; ++v) cs[v] = __ldg(src + v);
          }
        }
        if (prof) pt = clk32();
        mbar_wait(tmem_full_bar(acc), acc_phase);
        tc_fence_after();
        if (prof) { const unsigned int t = clk32(); pf0 += t - pt; pt = t; }
#pragma unroll
        for (int sl = 0; sl < 2; ++sl) {
          if (sl < nslab) {
            uint32_t r0[32], r1[32];
            tmem_ld_32x32b_x32(taddr + sl * 64, r0);
            tmem_ld_32x32b_x32(taddr + sl * 64 + 32, r1);
            tmem_ld_wait();
            if (sl == nslab - 1) release();
            uint4 q[8];
            if constexpr (EPI == EPI_BF16_ROPE) {
              rope32_bf16(r0, cs, rot, p.alpha, q);
              rope32_bf16(r1, cs + 8, rot, p.alpha, q + 4);
            } else {
              pack32_bf16(r0, p.alpha, q);
              pack32_bf16(r1, p.alpha, q + 4);
            }
            if (prof) { const unsigned int t = clk32(); pf2 += t - pt; pt = t; }
            if (n0 + sl * 64 < p.N) put(&tma_c, q, n0 + sl * 64, row0);
            if constexpr (EPI == EPI_BF16_CE) {
              const int nv = p.ce_valid - (n0 + sl * 64);
              if (nv > 0) ce_slab_stats(q, nv, ce_m, ce_s);
            }
            if (prof) { const unsigned int t = clk32(); pf3 += t - pt; pt = t; }
          }
        }
        if constexpr (EPI == EPI_BF16_CE) {
          // slot = this warp's 128-column stripe of the row; slot-major so that a warp's 32 rows store 256 contiguous bytes
          const int row = row0 + lane;
          if (n0 < p.N && row < p.M) p.ce_ms[static_cast<size_t>(n0 >> 7) * p.M + row] = make_float2(ce_m, ce_s);
        }
      } else {  // fp32 store / fp32 reduce-add: boxes of 32 columns
        const int n0 = n_blk * BN + wk.ncol0 + tcol;
        mbar_wait(tmem_full_bar(acc), acc_phase);
        tc_fence_after();
#pragma unroll
        for (int sl = 0; sl < 4; ++sl) {
          if (sl < 2 * nslab) {
            uint32_t r[32];
            tmem_ld_32x32b_x32(taddr + sl * 32, r);
            tmem_ld_wait();
            if (sl == 2 * nslab - 1) release();
            uint4 q[8];
            pack32_f32(r, p.alpha, q);
            if (n0 + sl * 32 < p.N) put(&tma_c, q, n0 + sl * 32, row0);
          }
        }
      }
      if (!next_work(it + 1, w, w)) break;
    }
    }  // EPI != EPI_DSWIGLU
#ifdef PCV_MEASURE
    if (prof && lane == 0) {
      unsigned int* dst = p.prof + (static_cast<size_t>(blockIdx.x) * 8 + (warp - 4)) * 8;
      dst[0] = pf0; dst[1] = pf1; dst[2] = pf2; dst[3] = pf3; dst[4] = clk32() - pstart;
    }
#else
    (void)pf0; (void)pf1; (void)pf2; (void)pf3; (void)pstart;
#endif
    if (lane == 0) tma_store_wait<0>();  // every box has left shared memory and reached global memory
    __syncwarp();
  }

  tc_fence_before();
  cluster_sync_all();  // the peer's MMAs/TMEM reads are finished before either CTA releases TMEM or exits
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------------ host side

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

// 2-D row-major [rows][cols] (ld elements between rows), 128B-swizzled box {128 bytes, box_rows}; bf16 or fp32.
static int make_tmap_2d(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                        bool f32) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) return PCV_ERR_DRIVER;
  const uint64_t esz = f32 ? 4 : 2;
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (ld * esz) % 16) return PCV_ERR_ALIGN;
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * esz};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(128 / esz), box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                   const_cast<void*>(ptr), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? PCV_OK : PCV_ERR_TMAP;
}
int make_tmap_2d_bf16(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
  return make_tmap_2d(out, ptr, rows, cols, ld, box_rows, false);
}

namespace {
struct TmapKey {
  const void* ptr;
  uint64_t rows, cols, ld;
  uint32_t box_rows;
  bool f32;
  bool operator==(const TmapKey& o) const {
    return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows && f32 == o.f32;
  }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    size_t h = reinterpret_cast<size_t>(k.ptr);
    h = h * 1000003u ^ k.rows;
    h = h * 1000003u ^ k.cols;
    h = h * 1000003u ^ k.ld;
    h = h * 1000003u ^ (k.box_rows * 2u + (k.f32 ? 1u : 0u));
    return h;
  }
};
std::mutex g_tmap_mu;
std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> g_tmap_cache;
}  // namespace

// Descriptors depend only on (pointer, shape, box): cache them, the training step re-uses the same buffers.
static int get_tmap_2d(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                       bool f32) {
  TmapKey key{ptr, rows, cols, ld, box_rows, f32};
  {
    std::lock_guard<std::mutex> g(g_tmap_mu);
    auto it = g_tmap_cache.find(key);
    if (it != g_tmap_cache.end()) {
      *out = it->second;
      return PCV_OK;
    }
  }
  int rc = make_tmap_2d(out, ptr, rows, cols, ld, box_rows, f32);
  if (rc != PCV_OK) return rc;
  std::lock_guard<std::mutex> g(g_tmap_mu);
  if (g_tmap_cache.size() > 65536) g_tmap_cache.clear();
  g_tmap_cache.emplace(key, *out);
  return PCV_OK;
}
int get_tmap_2d_bf16(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
  return get_tmap_2d(out, ptr, rows, cols, ld, box_rows, false);
}
int get_tmap_2d_f32(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
  return get_tmap_2d(out, ptr, rows, cols, ld, box_rows, true);
}

int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  }
  return n;
}

template <int BN, bool A_MN, bool B_MN, int EPI>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid, cudaStream_t st) {
  using Cfg = GemmCfg<BN>;
  auto kern = gemm_bf16_kernel<BN, A_MN, B_MN, EPI>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_set = true;
  }
  cudaError_t e = launch_pdl(kern, dim3(grid), dim3(kGemmThreads), Cfg::kSmemBytes, st, ta, tb, p);
  if (e == cudaSuccess) e = cudaGetLastError();
  return e == cudaSuccess ? PCV_OK : static_cast<int>(e);
}

template <int BN, bool A_MN, bool B_MN>
static int dispatch_epi(int epi, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                        cudaStream_t st) {
  switch (epi) {
    case PCV_EPI_BF16: return launch_gemm<BN, A_MN, B_MN, PCV_EPI_BF16>(ta, tb, p, grid, st);
    case PCV_EPI_F32_ATOMIC: return launch_gemm<BN, A_MN, B_MN, PCV_EPI_F32_ATOMIC>(ta, tb, p, grid, st);
    case PCV_EPI_F32: return launch_gemm<BN, A_MN, B_MN, PCV_EPI_F32>(ta, tb, p, grid, st);
    case EPI_BF16_ROPE:
      if constexpr (!A_MN && B_MN) return launch_gemm<BN, A_MN, B_MN, EPI_BF16_ROPE>(ta, tb, p, grid, st);
      else return PCV_ERR_ARG;
  }
  return PCV_ERR_ARG;
}

template <int BN>
static int dispatch_major(int a_mn, int b_mn, int epi, const CUtensorMap& ta, const CUtensorMap& tb,
                          const GemmParams& p, int grid, cudaStream_t st) {
  if (!a_mn && !b_mn) return dispatch_epi<BN, false, false>(epi, ta, tb, p, grid, st);
  if (!a_mn && b_mn) return dispatch_epi<BN, false, true>(epi, ta, tb, p, grid, st);
  if (a_mn && !b_mn) return dispatch_epi<BN, true, false>(epi, ta, tb, p, grid, st);
  return dispatch_epi<BN, true, true>(epi, ta, tb, p, grid, st);
}


// Output tensor maps of the CTA-pair kernel (boxes of 32 rows x 128 bytes): c = C (or the gate / dg half), c2 = the up /
// du half, x = h of the fused SwiGLU forward.  Unused ones repeat c.
struct PairOut {
  CUtensorMap c, c2, x;
  CUtensorMap bn;  // B with a 64-row box (K-major B, halved tail tiles); repeats tb when unused
};

template <bool A_MN, bool B_MN, int EPI>
static int launch_gemm_pair(const CUtensorMap& ta, const CUtensorMap& tb, const PairOut& o, const GemmParams& p,
                            int grid, cudaStream_t st) {
  auto kern = gemm_bf16_pair_kernel<A_MN, B_MN, EPI>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, PairCfgT<EPI>::kSmemBytes);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_set = true;
  }
  // __cluster_dims__(2,1,1): grid is even
  cudaError_t e = launch_pdl(kern, dim3(grid), dim3(kPairThreads), PairCfgT<EPI>::kSmemBytes, st, ta, tb, o.bn, o.c, o.c2, o.x, p);
  if (e == cudaSuccess) e = cudaGetLastError();
  return e == cudaSuccess ? PCV_OK : static_cast<int>(e);
}

template <bool A_MN, bool B_MN>
static int dispatch_epi_pair(int epi, const CUtensorMap& ta, const CUtensorMap& tb, const PairOut& o,
                             const GemmParams& p, int grid, cudaStream_t st) {
  switch (epi) {
    case PCV_EPI_BF16: return launch_gemm_pair<A_MN, B_MN, PCV_EPI_BF16>(ta, tb, o, p, grid, st);
    case PCV_EPI_F32_ATOMIC: return launch_gemm_pair<A_MN, B_MN, PCV_EPI_F32_ATOMIC>(ta, tb, o, p, grid, st);
    case PCV_EPI_F32: return launch_gemm_pair<A_MN, B_MN, PCV_EPI_F32>(ta, tb, o, p, grid, st);
    case EPI_BF16_ROPE:
      if constexpr (!A_MN && B_MN) return launch_gemm_pair<A_MN, B_MN, EPI_BF16_ROPE>(ta, tb, o, p, grid, st);
      else return PCV_ERR_ARG;
    case EPI_SWIGLU:
      if constexpr (!A_MN && B_MN) return launch_gemm_pair<A_MN, B_MN, EPI_SWIGLU>(ta, tb, o, p, grid, st);
      else return PCV_ERR_ARG;
    case EPI_DSWIGLU:
      if constexpr (!A_MN && !B_MN) return launch_gemm_pair<A_MN, B_MN, EPI_DSWIGLU>(ta, tb, o, p, grid, st);
      else return PCV_ERR_ARG;
    case EPI_BF16_CE:
      if constexpr (!A_MN) return launch_gemm_pair<A_MN, B_MN, EPI_BF16_CE>(ta, tb, o, p, grid, st);
      else return PCV_ERR_ARG;
  }
  return PCV_ERR_ARG;
}

// Tile schedule of the CTA-pair kernel (pcv_gemm_set_mode(5 / 6) or env PCV_GEMM_DYNAMIC=0 / 1, read once): 0 = static
// persistent (one pair per TPC, tiles dealt round-robin), 1 = dynamic (one cluster LAUNCHED per work item; running pairs
// cancel the launches behind them and take their tiles — pairs that start late or lose their SMs to another kernel,
// e.g. an NCCL all-reduce or the side-stream wgrad GEMM, simply take fewer tiles instead of holding up a whole wave).
static std::atomic<int> g_gemm_dynamic{-1};
static int gemm_dynamic() {
  int v = g_gemm_dynamic.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("PCV_GEMM_DYNAMIC");
    v = e ? (atoi(e) != 0 ? 1 : 0) : PCV_GEMM_DYNAMIC_DEFAULT;
    g_gemm_dynamic.store(v, std::memory_order_relaxed);
  }
  return v;
}

static int dispatch_major_pair(int a_mn, int b_mn, int epi, const CUtensorMap& ta, const CUtensorMap& tb,
                               const PairOut& o, const GemmParams& p_in, int grid, cudaStream_t st) {
  GemmParams p = p_in;
  p.dynamic = gemm_dynamic();
  if (p.dynamic) {
    const long tiles = static_cast<long>(p.m_tiles) * p.n_tiles * p.splits;
    const long work = p.narrow ? p.full_tiles + 2 * (tiles - p.full_tiles) : tiles;
    if (work > 0x3fffffffL) return PCV_ERR_ARG;
    grid = static_cast<int>(2 * work);
  }
  if (!a_mn && !b_mn) return dispatch_epi_pair<false, false>(epi, ta, tb, o, p, grid, st);
  if (!a_mn && b_mn) return dispatch_epi_pair<false, true>(epi, ta, tb, o, p, grid, st);
  if (a_mn && !b_mn) return dispatch_epi_pair<true, false>(epi, ta, tb, o, p, grid, st);
  return dispatch_epi_pair<true, true>(epi, ta, tb, o, p, grid, st);
}


// Split-K factor for the accumulate-into-fp32 (wgrad) GEMMs: the one that fills whole waves of CTA pairs best while
// keeping >= 16 k-blocks per split; ties go to the smaller factor (less red.add traffic).
static int auto_splits(int M, int N, int K) {
  const int pairs = num_sms() / 2;
  const long tiles = static_cast<long>((M + 255) / 256) * ((N + 255) / 256);
  const int kb = (K + kBK - 1) / kBK;
  int best = 1;
  double best_eff = 0.0;
  for (int s = 1; s <= 16; ++s) {
    if (s > 1 && kb / s < 16) break;
    const long work = tiles * s;
    const long waves = (work + pairs - 1) / pairs;
    const double eff = static_cast<double>(work) / static_cast<double>(waves * pairs);
    if (eff > best_eff + 0.02) { best_eff = eff; best = s; }
  }
  return best;
}

// Process-wide tiling override (tests and A/B runs; every setting computes the same product): relaxed atomics.
static std::atomic<int> g_gemm_mode{0};  // 0 auto, 1 single-CTA tiles only, 2 CTA-pair tiles whenever legal
static std::atomic<int> g_gemm_tail{1};  // halve the tiles of a sparse last wave (pcv_gemm_set_mode(3 / 4): off / on)
static std::atomic<long long> g_pair_launches{0};
#ifdef PCV_MEASURE
static int g_gemm_dbg = 0;
static unsigned int* g_gemm_prof = nullptr;
#endif

}  // namespace pcv

#ifdef PCV_MEASURE
extern "C" int pcv_gemm_set_profile(void* dev_buf) {
  pcv::g_gemm_prof = static_cast<unsigned int*>(dev_buf);
  return PCV_OK;
}
#endif

extern "C" long long pcv_gemm_pair_launch_count(void) { return pcv::g_pair_launches.load(std::memory_order_relaxed); }

extern "C" int pcv_gemm_set_mode(int mode) {
#ifdef PCV_MEASURE
  if (mode >= 100 && mode < 356) { pcv::g_gemm_dbg = mode - 100; return PCV_OK; }
#endif
  if (mode < 0 || mode > 6) return PCV_ERR_ARG;
  if (mode >= 5) pcv::g_gemm_dynamic.store(mode == 6, std::memory_order_relaxed);
  else if (mode >= 3) pcv::g_gemm_tail.store(mode == 4, std::memory_order_relaxed);
  else pcv::g_gemm_mode.store(mode, std::memory_order_relaxed);
  return PCV_OK;
}

static int gemm_impl(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                     int a_mn_major, int b_mn_major, int epilogue, int splits, float alpha, const float* cos_sin, int T,
                     int rope_cols, void* stream, float2* ce_ms = nullptr, int ce_valid = 0) {
  using namespace pcv;
  if (M <= 0 || N <= 0 || K <= 0 || !A || !B || !C) return PCV_ERR_ARG;
  if (splits < 0) return PCV_ERR_ARG;
  if (splits == 0) splits = (epilogue == PCV_EPI_F32_ATOMIC) ? auto_splits(M, N, K) : 1;
  if (epilogue != PCV_EPI_F32_ATOMIC && splits != 1) return PCV_ERR_ARG;
  const int vec = (epilogue == PCV_EPI_BF16 || epilogue == EPI_BF16_ROPE || epilogue == EPI_BF16_CE) ? 8 : 4;
  if (N % vec || ldc % vec || (reinterpret_cast<uintptr_t>(C) & 15)) return PCV_ERR_ALIGN;

  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int kb_total = (K + kBK - 1) / kBK;
  // CTA-pair 256 x 256 tiles whenever they can occupy every SM pair at least once; otherwise single-CTA tiles.
  {
    const int m2 = (M + 255) / 256, n2 = (N + 255) / 256;
    const int kps = (kb_total + splits - 1) / splits;
    const int sp = (kb_total + kps - 1) / kps;
    const long t2 = static_cast<long>(m2) * n2 * sp;
    const int mode = g_gemm_mode.load(std::memory_order_relaxed);
    // (the fused cross-entropy partials exist in the CTA-pair epilogue only)
    const bool pair = mode == 2 || epilogue == EPI_BF16_CE || (mode == 0 && t2 >= num_sms() / 2);
    if (pair) {
      CUtensorMap ta, tb;
      int rc;
      if (!a_mn_major) rc = get_tmap_2d_bf16(&ta, A, M, K, lda, 128);
      else rc = get_tmap_2d_bf16(&ta, A, K, M, lda, 64);
      if (rc) return rc;
      if (!b_mn_major) rc = get_tmap_2d_bf16(&tb, B, N, K, ldb, 128);
      else rc = get_tmap_2d_bf16(&tb, B, K, N, ldb, 64);
      if (rc) return rc;
      GemmParams p;
      p.C = C;
      p.M = M; p.N = N; p.K = K; p.ldc = ldc;
      p.m_tiles = m2; p.n_tiles = n2;
      p.kb_total = kb_total; p.kb_per_split = kps; p.splits = sp;
      p.alpha = alpha;
      p.cos_sin = cos_sin; p.T = T; p.rope_cols = rope_cols;
      p.aux = nullptr; p.ld_aux = 0; p.half_off = 0;
      PairOut o;
      p.ce_ms = ce_ms;
      p.ce_valid = ce_valid;
      if (epilogue == PCV_EPI_BF16 || epilogue == EPI_BF16_ROPE || epilogue == EPI_BF16_CE) rc = get_tmap_2d_bf16(&o.c, C, M, N, ldc, 32);
      else rc = get_tmap_2d_f32(&o.c, C, M, N, ldc, 32);
      if (rc) return rc;
      o.c2 = o.c;
      o.x = o.c;
      o.bn = tb;
      // tail of the schedule: if the last wave would occupy at most half of the pairs, run its tiles as 256 x 128 halves
      p.full_tiles = static_cast<int>(t2);
      p.narrow = 0;
#ifdef PCV_MEASURE
      p.dbg = g_gemm_dbg;
      p.prof = g_gemm_prof;
#endif
      {
        const long np = num_sms() / 2;
        const long rem = t2 % np;
        // (not with the fused cross-entropy partials: their slots are the 128-column stripes of full tiles)
        if (g_gemm_tail.load(std::memory_order_relaxed) && sp == 1 && epilogue != PCV_EPI_F32_ATOMIC && epilogue != EPI_BF16_CE &&
            t2 > np && rem > 0 && 2 * rem <= np) {
          p.full_tiles = static_cast<int>(t2 - rem);
          p.narrow = 1;
          if (!b_mn_major) {
            rc = get_tmap_2d_bf16(&o.bn, B, N, K, ldb, 64);
            if (rc) return rc;
          }
        }
      }
      const long pairs = t2 < num_sms() / 2 ? t2 : num_sms() / 2;
      count_launch();
      g_pair_launches.fetch_add(1, std::memory_order_relaxed);
      return dispatch_major_pair(a_mn_major, b_mn_major, epilogue, ta, tb, o, p, static_cast<int>(2 * pairs), st);
    }
  }

  // tile-N choice: 256 unless the problem cannot fill the machine with 128x256 tiles
  const int m_tiles = (M + kBM - 1) / kBM;
  int BN = 256;
  {
    const long t256 = static_cast<long>(m_tiles) * ((N + 255) / 256) * splits;
    if (N <= 128 || t256 < 2L * num_sms()) BN = 128;
  }
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn_major) rc = get_tmap_2d_bf16(&ta, A, M, K, lda, kBM);
  else rc = get_tmap_2d_bf16(&ta, A, K, M, lda, 64);
  if (rc) return rc;
  if (!b_mn_major) rc = get_tmap_2d_bf16(&tb, B, N, K, ldb, BN);
  else rc = get_tmap_2d_bf16(&tb, B, K, N, ldb, 64);
  if (rc) return rc;

  GemmParams p;
  p.C = C;
  p.M = M; p.N = N; p.K = K; p.ldc = ldc;
  p.m_tiles = m_tiles;
  p.n_tiles = (N + BN - 1) / BN;
  p.kb_total = kb_total;
  p.kb_per_split = (p.kb_total + splits - 1) / splits;
  p.splits = (p.kb_total + p.kb_per_split - 1) / p.kb_per_split;
  p.alpha = alpha;
  p.cos_sin = cos_sin; p.T = T; p.rope_cols = rope_cols;
  p.aux = nullptr; p.ld_aux = 0; p.half_off = 0;
  p.ce_ms = nullptr; p.ce_valid = 0;
  const long total = static_cast<long>(p.m_tiles) * p.n_tiles * p.splits;
  const int grid = static_cast<int>(total < num_sms() ? total : num_sms());
  count_launch();
  if (BN == 256) return dispatch_major<256>(a_mn_major, b_mn_major, epilogue, ta, tb, p, grid, st);
  return dispatch_major<128>(a_mn_major, b_mn_major, epilogue, ta, tb, p, grid, st);
}

extern "C" int pcv_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                             int a_mn_major, int b_mn_major, int epilogue, int splits, float alpha, void* stream) {
  if (epilogue < PCV_EPI_BF16 || epilogue > PCV_EPI_F32) return PCV_ERR_ARG;
  return gemm_impl(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, epilogue, splits, alpha, nullptr, 1, 0, stream);
}

extern "C" int pcv_gemm_bf16_rope(const void* X, const void* W, void* Y, int M, int N, int K, int ldx, int ldw, int ldy,
                                  const float* cos_sin, int T, int rope_cols, void* stream) {
  if (!cos_sin || T <= 0 || rope_cols < 0 || rope_cols > N || rope_cols % 64 || M % T) return PCV_ERR_ARG;
  return gemm_impl(X, W, Y, M, N, K, ldx, ldw, ldy, 0, 1, pcv::EPI_BF16_ROPE, 1, 1.0f, cos_sin, T, rope_cols, stream);
}

extern "C" int pcv_gemm_bf16_ce(const void* X, const void* W, void* logits, int M, int N, int K, int ldx, int ldw, int ldl,
                                int w_mn_major, int n_valid, void* ce_ms, void* stream) {
  if (!ce_ms || n_valid <= 0 || n_valid > N) return PCV_ERR_ARG;
  if (reinterpret_cast<uintptr_t>(ce_ms) & 7) return PCV_ERR_ALIGN;
  return gemm_impl(X, W, logits, M, N, K, ldx, ldw, ldl, 0, w_mn_major, pcv::EPI_BF16_CE, 1, 1.0f, nullptr, 1, 0, stream,
                   static_cast<float2*>(ce_ms), n_valid);
}

// Fused SwiGLU GEMMs: always CTA-pair tiles (the gate | up pairing of the forward needs the two B halves of a pair).
static int gemm_swiglu_impl(bool bwd, const void* A, const void* W, void* C, void* aux, int M, int hid, int K, int lda,
                            int ldw, int ldc, int ld_aux, float alpha, void* stream) {
  using namespace pcv;
  if (M <= 0 || hid <= 0 || K <= 0 || !A || !W || !C || !aux) return PCV_ERR_ARG;
  if (hid % 8 || ldc % 8 || ld_aux % 8 || (reinterpret_cast<uintptr_t>(C) & 15) || (reinterpret_cast<uintptr_t>(aux) & 15))
    return PCV_ERR_ALIGN;
  CUtensorMap ta, tb;
  int rc = get_tmap_2d_bf16(&ta, A, M, K, lda, 128);
  if (rc) return rc;
  // forward: W_gu stored [K][2*hid] (MN-major B, boxes of 64 columns); backward: W_down stored [hid][K] (K-major B)
  if (!bwd) rc = get_tmap_2d_bf16(&tb, W, K, 2 * static_cast<uint64_t>(hid), ldw, 64);
  else rc = get_tmap_2d_bf16(&tb, W, hid, K, ldw, 128);
  if (rc) return rc;
  GemmParams p;
  p.C = C;
  p.M = M; p.N = hid; p.K = K; p.ldc = ldc;
  p.m_tiles = (M + 255) / 256;
  p.n_tiles = bwd ? (hid + 255) / 256 : (hid + 127) / 128;
  p.kb_total = (K + kBK - 1) / kBK; p.kb_per_split = p.kb_total; p.splits = 1;
  p.alpha = alpha;
  p.cos_sin = nullptr; p.T = 1; p.rope_cols = 0;
  p.aux = aux; p.ld_aux = ld_aux; p.half_off = hid;
  p.ce_ms = nullptr; p.ce_valid = 0;
  // outputs: the gate (dg) and the up (du) half of C = [M][2*hid] as two maps clipped at hid columns, and h
  PairOut o;
  __nv_bfloat16* cb = static_cast<__nv_bfloat16*>(C);
  rc = get_tmap_2d_bf16(&o.c, cb, M, hid, ldc, 32);
  if (rc) return rc;
  rc = get_tmap_2d_bf16(&o.c2, cb + hid, M, hid, ldc, 32);
  if (rc) return rc;
  // forward: x = h (output).  backward: x / bn = the gate / up half of the saved gu (inputs of the epilogue)
  rc = get_tmap_2d_bf16(&o.x, aux, M, hid, ld_aux, 32);
  if (rc) return rc;
  if (bwd) rc = get_tmap_2d_bf16(&o.bn, static_cast<__nv_bfloat16*>(aux) + hid, M, hid, ld_aux, 32);
  else o.bn = tb;
  if (rc) return rc;
  p.full_tiles = p.m_tiles * p.n_tiles;
  p.narrow = 0;
#ifdef PCV_MEASURE
  p.dbg = g_gemm_dbg;
  p.prof = g_gemm_prof;
#endif
  const long t2 = static_cast<long>(p.m_tiles) * p.n_tiles;
  const long pairs = t2 < num_sms() / 2 ? t2 : num_sms() / 2;
  count_launch();
  g_pair_launches.fetch_add(1, std::memory_order_relaxed);
  return dispatch_major_pair(0, bwd ? 0 : 1, bwd ? EPI_DSWIGLU : EPI_SWIGLU, ta, tb, o, p, static_cast<int>(2 * pairs),
                             static_cast<cudaStream_t>(stream));
}

extern "C" int pcv_gemm_bf16_swiglu(const void* X, const void* W_gu, void* gu, void* h, int M, int hid, int K, int ldx,
                                    int ldw, int ldgu, int ldh, void* stream) {
  if (ldgu < 2 * hid || ldh < hid || ldw < 2 * hid) return PCV_ERR_ARG;
  return gemm_swiglu_impl(false, X, W_gu, gu, h, M, hid, K, ldx, ldw, ldgu, ldh, 1.0f, stream);
}

extern "C" int pcv_gemm_bf16_dswiglu(const void* dY, const void* W_down, const void* gu, void* dgu, int M, int hid,
                                     int K, int lddy, int ldw, int ldgu, void* stream) {
  if (ldgu < 2 * hid || ldw < K) return PCV_ERR_ARG;
  return gemm_swiglu_impl(true, dY, W_down, dgu, const_cast<void*>(gu), M, hid, K, lddy, ldw, ldgu, ldgu, 1.0f, stream);
}
