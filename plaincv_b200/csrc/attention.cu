// attention.cu — flash-style causal / document-masked attention forward and backward on tcgen05 + TMEM for sm_100a,
// head_dim = 64.  q and k arrive ALREADY ROTATED: the interleaved-pair RoPE of the reference is applied by the QKV GEMM
// epilogue (pcv_gemm_bf16_rope) or by pcv_rope_qk below; the backward kernels apply its transpose to dQ and dK.
//
// Replaces, on the reference's hot path (all XLA-generated there, O(T^2) memory):
//   jax.nn.dot_product_attention       models/LM/transformer.py:233-240  (fp32 logits/softmax, bf16 P before P.V)
//   the bool [B,T,T] block-diagonal mask of train_lm.py:107-131 / data_prep_utils.py:14-43 — consumed here as
//   doc_start[b*T+t] (first position of t's document); the [T,T] mask is never built.
//   apply_rotary_emb_complex_like      models/LM/embedding.py:28-67  (stand-alone form: rope_qk_kernel)
// and their autodiff transposes (train_lm.py:267).
//
// Layout: qkv[B*T][3d] bf16, columns [q | k | v], each third head-major [N][64].  out[B*T][d] bf16.
// lse[B][N][T] fp32 (natural log).  Tiles are 128 rows x 64 columns = 128 x 128 B, loaded by TMA and 128B-swizzled in
// shared memory: K-major operand for Q.K^T, and the *same bytes* serve as the MN-major operand of P.V / dS^T.Q / P^T.dO.
//
// Forward (attn_fwd_kernel, 192 threads, TWO CTAs per SM): one CTA = one (batch, head, 128-query tile).
//   warps 0-3 : softmax warpgroup, one thread per query row; 64-key half tiles, S double-buffered in tensor memory, one
//               pass per half tile with a lazily moved exponent reference; P goes back to tensor memory as bf16
//   warp 4    : TMA producer (Q; K / V rings of 2) + tensor-memory allocation;   warp 5 : tcgen05.mma issuer
//   TMEM (256 columns per CTA): S0 [0,64) S1 [64,128) O [128,192) P0 [192,224) P1 [224,256).
//
// Backward (attn_bwd_kernel, 512 threads, persistent): a CTA takes (batch, head, 128-key tile) items — its own
// blockIdx first, then launches it cancels through cluster launch control — and loops over the query tiles that see
// the key tile:  S = Q_i.K^T, dP = dO_i.V^T -> P = exp2(S*c - lse), dS = P*(dP - delta)*scale;
//   dV += P^T.dO_i, dK += dS^T.Q_i (accumulated in tensor memory over the loop), dQ_i = dS.K (TMA reduce-add into fp32).
//   warps 0-7 : compute (row r, 64 key columns each);  warps 8-11 : everything that leaves tensor memory (dQ per tile,
//   dK / dV per item);  warp 12 : TMA producer + work stealing;  warp 13 : tcgen05.mma issuer
//   TMEM: S [0,128) dP [128,256) dV [256,320) dK [320,384) dQ [384,448) dS(bf16 pairs) [448,512).
// attn_bwd_prep_kernel (delta = rowsum(dO * O), zero-fill of the dQ accumulator) runs before it,
// attn_bwd_dq_post_kernel (un-rotation + bf16 of dQ) after it.
#include "sm100.cuh"
#include "pcv_internal.h"
#include <math.h>
#include <mutex>

namespace pcv {

#ifdef PCV_TRACE
// Debug-only timeline (clock64 at phase boundaries of CTA (0,0,0)); compiled in with -DPCV_TRACE, never in the product build.
__device__ long long g_trace[8192];
__device__ long long g_cta_trace[4096 * 4];  // per backward CTA: globaltimer at entry / exit, SM id, iterations
__device__ __forceinline__ long long gtimer() {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define TRACE(slot, idx) do { if (trace_on && (idx) < 1024) g_trace[(slot) * 1024 + (idx)] = clock64(); } while (0)
#else
#define TRACE(slot, idx) do { } while (0)
#endif

constexpr int kAttnThreads = 192;     // forward: softmax warpgroup + producer + MMA issuer
constexpr int kAttnBwdThreads = 512;  // backward: 2 compute warpgroups, 1 dQ warpgroup, producer + MMA warps
constexpr int kTileBytes = 128 * 128;  // 128 rows x 64 bf16
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kSoftmaxScale = 0.125f;  // 1/sqrt(64)

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// 2^x on the FMA / ALU pipes instead of MUFU (the forward kernel is bound by the 16 exponentials per clock an SM's
// MUFU units deliver, so a share of them is evaluated here): x is clamped to [-126, inf), split as x = n + f with
// n = round(x) taken from the low mantissa bits of x + 1.5 * 2^23, f in [-0.5, 0.5], 2^f by a degree-4 polynomial
// (max rel. error 4e-6 — two orders below the bf16 rounding of P) and n added into the exponent field.
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -126.f);
  const float xr = x + 12582912.f;               // 1.5 * 2^23: the integer part lands in the low mantissa bits
  const float f = x - (xr - 12582912.f);
  float p = fmaf(f, 9.6181291e-3f, 5.5504109e-2f);  // Taylor/minimax of 2^f around 0 on [-0.5, 0.5]
  p = fmaf(p, f, 2.4022651e-1f);
  p = fmaf(p, f, 6.9314718e-1f);
  p = fmaf(p, f, 1.0f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(xr) << 23));
}
#ifndef PCV_ATTN_POLY
#define PCV_ATTN_POLY 0  // of every 8 exponentials in the forward softmax, this many take the polynomial
#endif
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// tcgen05.wait::ld that also names the 32 destination registers of the load it completes, so no use of them can be
// scheduled ahead of it (the loads below are issued one chunk ahead of their use)
__device__ __forceinline__ void tmem_ld_wait_for(uint32_t* r) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                 "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                 "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                 "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :
               : "memory");
}
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_global_nc_v4(const void* p) {
  uint4 v;
  asm("ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Rotate the 4 interleaved pairs held in one 16-byte chunk.  cs0 = (cos0, sin0, cos1, sin1), cs1 = (cos2, sin2, cos3, sin3).
// sign = +1: forward rotation (embedding.py:58-59); sign = -1: its transpose (gradient).
__device__ __forceinline__ uint4 rope_chunk(uint4 raw, float4 cs0, float4 cs1, float sign) {
  const float2 p0 = unpack_bf16x2(raw.x), p1 = unpack_bf16x2(raw.y), p2 = unpack_bf16x2(raw.z), p3 = unpack_bf16x2(raw.w);
  uint4 o;
  o.x = pack_bf16x2(p0.x * cs0.x - sign * p0.y * cs0.y, p0.y * cs0.x + sign * p0.x * cs0.y);
  o.y = pack_bf16x2(p1.x * cs0.z - sign * p1.y * cs0.w, p1.y * cs0.z + sign * p1.x * cs0.w);
  o.z = pack_bf16x2(p2.x * cs1.x - sign * p2.y * cs1.y, p2.y * cs1.x + sign * p2.x * cs1.y);
  o.w = pack_bf16x2(p3.x * cs1.z - sign * p3.y * cs1.w, p3.y * cs1.z + sign * p3.x * cs1.w);
  return o;
}

// ---- softmax chunk helpers (32 score columns held by one thread).  MASK variants take the valid column window
// [lo, hi] relative to the chunk start; the unmasked variants are branch- and select-free.
template <bool MASK>
__device__ __forceinline__ float chunk_rowmax(const uint32_t* sr, int lo, int hi, float rmax) {
#pragma unroll
  for (int k = 0; k < 32; ++k) {
    float v = __uint_as_float(sr[k]);
    if (MASK) v = (k >= lo && k <= hi) ? v : -INFINITY;
    rmax = fmaxf(rmax, v);
  }
  return rmax;
}
// p = exp2(s*c2 - mb) -> packed bf16 pairs; returns the fp32 row-sum contribution
template <bool MASK>
__device__ __forceinline__ float chunk_exp_pack(const uint32_t* sr, float c2, float mb, int lo, int hi, uint32_t* packed) {
  float s0 = 0.f, s1 = 0.f;
#pragma unroll
  for (int k = 0; k < 32; k += 2) {
    float e0 = ex2f(fmaf(__uint_as_float(sr[k]), c2, -mb));
    float e1 = ex2f(fmaf(__uint_as_float(sr[k + 1]), c2, -mb));
    if (MASK) {
      e0 = (k >= lo && k <= hi) ? e0 : 0.f;
      e1 = (k + 1 >= lo && k + 1 <= hi) ? e1 : 0.f;
    }
    s0 += e0;
    s1 += e1;
    packed[k >> 1] = pack_bf16x2(e0, e1);
  }
  return s0 + s1;
}

// backward: p = exp2(s*c2 - lse2), ds = p * (dp*scale - delta*scale), both packed to bf16 pairs (fp32x2 math)
template <bool MASK>
__device__ __forceinline__ void chunk_p_ds(const uint32_t* sr, const uint32_t* pr, float c2, float lse2, float dlt_s,
                                           int lo, int hi, uint32_t* pk_p, uint32_t* pk_ds) {
  const float2 c2v = make_float2(c2, c2), lv = make_float2(-lse2, -lse2);
  const float2 scv = make_float2(kSoftmaxScale, kSoftmaxScale), dv = make_float2(-dlt_s, -dlt_s);
#pragma unroll
  for (int k = 0; k < 32; k += 2) {
    const float2 x = __ffma2_rn(make_float2(__uint_as_float(sr[k]), __uint_as_float(sr[k + 1])), c2v, lv);
    float2 e = make_float2(ex2f(x.x), ex2f(x.y));
    if (MASK) {
      e.x = (k >= lo && k <= hi) ? e.x : 0.f;
      e.y = (k + 1 >= lo && k + 1 <= hi) ? e.y : 0.f;
    }
    const float2 tt = __ffma2_rn(make_float2(__uint_as_float(pr[k]), __uint_as_float(pr[k + 1])), scv, dv);
    const float2 g = __fmul2_rn(e, tt);
    pk_p[k >> 1] = pack_bf16x2(e.x, e.y);
    pk_ds[k >> 1] = pack_bf16x2(g.x, g.y);
  }
}

// ================================================================================================ forward
struct AttnFwdSmem {
  static constexpr int kSlots = 2;                          // K / V ring depth
  static constexpr int kQ = 0;                              // 1 tile (doubles as the staging tile of O at the end)
  static constexpr int kK = kQ + kTileBytes;                // kSlots tiles
  static constexpr int kV = kK + kSlots * kTileBytes;       // kSlots tiles
  static constexpr int kBar = kV + kSlots * kTileBytes;     // barriers
  static constexpr int kBytes = kBar + 256 + 1024;
};
// two CTAs per SM: 2 x (dynamic + 1 KB system reservation) must fit in 227 KB
static_assert(2 * (AttnFwdSmem::kBytes + 1024) <= 232448, "attn_fwd: two CTAs must fit in one SM's shared memory");

// 32 score columns of one row: p = exp2(s*c2 - mb) (0 outside [lo, hi] when MASK), packed to bf16 pairs.
// Accumulates the row sum into `sum2` (two partial sums) and, when TRACK, the largest valid score into `smax`.
// Packed fp32x2 FMA / ADD halve the issue slots next to the one MUFU op per element.
template <bool MASK, bool TRACK>
__device__ __forceinline__ void chunk_softmax(const uint32_t* sr, float c2, float mb, int lo, int hi, uint32_t* packed,
                                              float2& sum2, float& smax) {
  const float2 c2v = make_float2(c2, c2), mbv = make_float2(-mb, -mb);
#pragma unroll
  for (int k = 0; k < 32; k += 2) {
    const float s0 = __uint_as_float(sr[k]), s1 = __uint_as_float(sr[k + 1]);
    const float2 x = __ffma2_rn(make_float2(s0, s1), c2v, mbv);
    // elements (k mod 8) < PCV_ATTN_POLY take the FMA-pipe polynomial, the rest MUFU.EX2
    float e0 = ((k & 7) < PCV_ATTN_POLY) ? ex2_poly(x.x) : ex2f(x.x);
    float e1 = (((k + 1) & 7) < PCV_ATTN_POLY) ? ex2_poly(x.y) : ex2f(x.y);
    if (MASK) {
      const bool v0 = (k >= lo && k <= hi), v1 = (k + 1 >= lo && k + 1 <= hi);
      e0 = v0 ? e0 : 0.f;
      e1 = v1 ? e1 : 0.f;
      if (TRACK) smax = fmaxf(smax, fmaxf(v0 ? s0 : -INFINITY, v1 ? s1 : -INFINITY));
    } else if (TRACK) {
      smax = fmaxf(smax, fmaxf(s0, s1));
    }
    sum2 = __fadd2_rn(sum2, make_float2(e0, e1));
    packed[k >> 1] = pack_bf16x2(e0, e1);
  }
}

// Forward CTA (192 threads) = one (batch, head, 128-query tile); TWO CTAs are resident per SM (256 of the 512
// tensor-memory columns and 82 KB of shared memory each), so one CTA's prologue (barrier set-up, tensor-memory
// allocation, first loads) and epilogue run under the other CTA's exponentials (the former 256-query CTA spent 3.5 us
// of fixed cost per CTA, 18 % of the kernel, overlapped with nothing).
// Keys are processed in HALF tiles of 64 with the scores double-buffered in tensor memory: S(h+1) = Q.K^T is computed
// while the warpgroup exponentiates S(h), so the softmax warps never wait for the tensor pipe (with one 128-key S
// buffer they idled ~500 of every 2700 cycles, and two CTAs with equal periods stay phase-locked: the exponent units
// (MUFU, the bound of this kernel: 16 / clock / SM) were 74 % busy in steady state).
// q, k in `qkv` are already rotated (RoPE is applied by the QKV GEMM epilogue, pcv_gemm_bf16_rope, or by pcv_rope_qk):
//   warps 0-3 : softmax warpgroup, one thread per query row.  One pass over S per half tile: the exponent reference is
//               the row max of the row's first visible half tile and is only moved if a later row sum shows that a
//               score exceeded it by ~2^64 (everything is floating point: a stale reference costs no precision), so
//               P.V accumulates in tensor memory across key tiles with no per-tile rescale or read-back.  The
//               tensor-memory load of score chunk 1 is in flight while chunk 0 is exponentiated.
//   warp 4    : TMA producer (Q once, K / V rings of 2 x 128-key tiles) + tensor-memory allocation
//   warp 5    : tcgen05.mma issuer: S(h) = Q.K_h^T (128 x 64 x 64), O += P(h).V_h (128 x 64 x 64)
// TMEM (256 columns): S0 [0,64) S1 [64,128) O [128,192) P0 [192,224) P1 [224,256) (P: bf16 pairs).
__global__ void __launch_bounds__(kAttnThreads, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_out, const int32_t* __restrict__ doc_start,
                __nv_bfloat16* __restrict__ out, float* __restrict__ lse, int T, int NH) {
  extern __shared__ uint8_t smem_raw[];
  constexpr int kSlots = AttnFwdSmem::kSlots;
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = sbase + AttnFwdSmem::kQ, sK = sbase + AttnFwdSmem::kK, sV = sbase + AttnFwdSmem::kV,
                 sBar = sbase + AttnFwdSmem::kBar;
  const uint32_t bQ = sBar;                                                // (tx)
  auto bKfull = [&](int s) { return sBar + 8u * (1 + s); };                // [kSlots] (tx)
  auto bVfull = [&](int s) { return sBar + 8u * (1 + kSlots + s); };       // [kSlots] (tx)
  auto bKempty = [&](int s) { return sBar + 8u * (1 + 2 * kSlots + s); };  // [kSlots] (1, commit)
  auto bVempty = [&](int s) { return sBar + 8u * (1 + 3 * kSlots + s); };  // [kSlots] (1, commit)
  auto bSfull = [&](int u) { return sBar + 8u * (1 + 4 * kSlots + u); };   // [2] (1, commit)
  auto bPfull = [&](int u) { return sBar + 8u * (3 + 4 * kSlots + u); };   // [2] (128)
  auto bPVdone = [&](int u) { return sBar + 8u * (5 + 4 * kSlots + u); };  // [2] (1, commit)
  const uint32_t tmem_ptr_smem = sBar + 8u * (7 + 4 * kSlots);
  volatile uint32_t* tmem_ptr_generic = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_ptr_smem - smem_u32(smem_raw)));

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#ifdef PCV_TRACE
  const bool trace_on = (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) && ((threadIdx.x & 127) == 0 || threadIdx.x == 160);
  const int cta_lin = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  if (tid == 0 && cta_lin < 4096) {
    unsigned int smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    g_cta_trace[cta_lin * 4 + 0] = gtimer();
    g_cta_trace[cta_lin * 4 + 2] = smid;
  }
#endif
  // launch order = blockIdx.x fastest, z slowest: every (head, batch) pair of the heaviest (latest) query tile first,
  // the lightest tiles last, so that the SMs run dry within one short CTA of each other
  const int n_qt = gridDim.z;
  const int qt = n_qt - 1 - blockIdx.z;
  const int h = blockIdx.x, b = blockIdx.y;
  const int d = NH * 64;
  const int q0 = qt * 128;
  const long seq_row0 = static_cast<long>(b) * T;

  // key tiles [jlo, jhi] visible to this query tile: causal, and not before the first row's document
  const int jhi = qt;
  const int jlo = doc_start ? (doc_start[seq_row0 + q0] / 128) : 0;
  const int n_it = jhi - jlo + 1;
  const int n_ht = 2 * n_it;  // half tiles of 64 keys

  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tm_qkv);
    tma_prefetch_desc(&tm_out);
    mbar_init(bQ, 1);
    for (int s = 0; s < kSlots; ++s) {
      mbar_init(bKfull(s), 1);
      mbar_init(bVfull(s), 1);
      mbar_init(bKempty(s), 1);
      mbar_init(bVempty(s), 1);
    }
    for (int u = 0; u < 2; ++u) {
      mbar_init(bSfull(u), 1);
      mbar_init(bPfull(u), 128);
      mbar_init(bPVdone(u), 1);
    }
    fence_mbar_init();
  }
  if (warp == 4) {
    __syncwarp();
    tmem_alloc<256>(tmem_ptr_smem);
  }
  // programmatic dependent launch: everything above may have run under the tail of the kernel that produces qkv
  pdl_wait();
  pdl_trigger();
  if (warp == 4 && lane == 0) {
    // Q and the first K / V tile are requested right away: their latency runs under the block-wide barrier below
    const int row0 = static_cast<int>(seq_row0);
    mbar_expect_tx(bQ, kTileBytes);
    tma_load_2d(sQ, &tm_qkv, bQ, h * 64, row0 + q0);
    mbar_expect_tx(bKfull(0), kTileBytes);
    tma_load_2d(sK, &tm_qkv, bKfull(0), d + h * 64, row0 + jlo * 128);
    mbar_expect_tx(bVfull(0), kTileBytes);
    tma_load_2d(sV, &tm_qkv, bVfull(0), 2 * d + h * 64, row0 + jlo * 128);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_generic;

  if (warp == 4) {
    // ============================== TMA producer ==============================
    if (lane == 0) {
      const int row0 = static_cast<int>(seq_row0);
      int idx = 1;
      for (int j = jlo + 1; j <= jhi; ++j, ++idx) {
        const int slot = idx % kSlots;
        const uint32_t par = (idx / kSlots) & 1;
        TRACE(6, idx * 3 + 0);
        mbar_wait(bKempty(slot), par ^ 1u);
        mbar_expect_tx(bKfull(slot), kTileBytes);
        tma_load_2d(sK + slot * kTileBytes, &tm_qkv, bKfull(slot), d + h * 64, row0 + j * 128);
        TRACE(6, idx * 3 + 1);
        mbar_wait(bVempty(slot), par ^ 1u);
        mbar_expect_tx(bVfull(slot), kTileBytes);
        tma_load_2d(sV + slot * kTileBytes, &tm_qkv, bVfull(slot), 2 * d + h * 64, row0 + j * 128);
        TRACE(6, idx * 3 + 2);
      }
    }
    __syncwarp();
  } else if (warp == 5) {
    // ============================== MMA issuer ==============================
    {
      const uint32_t lead = elect_one_sync();
      const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 64, false, false);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, false, true);
      // S(hh) = Q . K[64 keys of half tile hh]^T -> score buffer hh & 1
      auto issue_S = [&](int hh) {
        const int i = hh >> 1, sub = hh & 1, slot = i % kSlots;
        if (sub == 0) {
          mbar_wait(bKfull(slot), (i / kSlots) & 1);
          tc_fence_after();
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint64_t da = umma_smem_desc(sQ + k * 32, 16, 1024);
          const uint64_t db = umma_smem_desc(sK + slot * kTileBytes + sub * 8192 + k * 32, 16, 1024);
          umma_bf16_ss_p(tb + (hh & 1) * 64, da, db, idesc_s, k > 0 ? 1u : 0u, lead);
        }
        tc_commit_p(bSfull(hh & 1), lead);
        if (sub == 1) tc_commit_p(bKempty(slot), lead);  // both halves of this K tile have been read
      };
      // O += P(hh) . V[64 keys of half tile hh]
      auto issue_PV = [&](int hh) {
        const int i = hh >> 1, sub = hh & 1, slot = i % kSlots;
        if (sub == 0) {
          mbar_wait(bVfull(slot), (i / kSlots) & 1);
          tc_fence_after();
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          // A = P straight from tensor memory (bf16 pairs, 8 columns per 16 keys); B = V (MN-major: one 128-byte
          // row per key).  P never touches shared memory.
          const uint64_t db = umma_smem_desc(sV + slot * kTileBytes + sub * 8192 + k * 2048, kTileBytes, 1024);
          umma_bf16_ts_p(tb + 128, tb + 192 + (hh & 1) * 32 + k * 8, db, idesc_pv, (hh > 0 || k > 0) ? 1u : 0u, lead);
        }
        tc_commit_p(bPVdone(hh & 1), lead);
        if (sub == 1) tc_commit_p(bVempty(slot), lead);
      };
      mbar_wait(bQ, 0);
      tc_fence_after();
      issue_S(0);
      issue_S(1);
      for (int hh = 0; hh < n_ht; ++hh) {
        TRACE(4, hh * 2 + 0);
        mbar_wait(bPfull(hh & 1), (hh >> 1) & 1);  // P(hh) is in tensor memory and S buffer hh & 1 has been consumed
        tc_fence_after();
        TRACE(4, hh * 2 + 1);
        issue_PV(hh);
        if (hh + 2 < n_ht) issue_S(hh + 2);
      }
    }
    __syncwarp();
  } else {
    // ============================== softmax warpgroup ==============================
    const int r = tid;
    const int t = q0 + r;
    const int ds_t = doc_start ? doc_start[seq_row0 + min(t, T - 1)] : 0;
    const int ds_hi = doc_start ? doc_start[seq_row0 + min(q0 + 127, T - 1)] : 0;
    const uint32_t lane_off = static_cast<uint32_t>(warp * 32) << 16;
    const uint32_t tO = tmem_base + lane_off + 128;
    constexpr float c2 = kSoftmaxScale * kLog2e;
    constexpr float kSumLimit = 1.8446744e19f;  // 2^64: a row sum beyond it means a score left the reference's range
    float ref = -INFINITY, l_run = 0.f;
    for (int hh = 0; hh < n_ht; ++hh) {
      const int u = hh & 1;
      const uint32_t tS = tmem_base + lane_off + u * 64;
      const uint32_t tP = tmem_base + lane_off + 192 + u * 32;
      const int key0 = (jlo + (hh >> 1)) * 128 + u * 64;
      TRACE(0, hh * 4 + 0);
      mbar_wait(bSfull(u), (hh >> 1) & 1);
      tc_fence_after();
      TRACE(0, hh * 4 + 1);
      const bool need_mask = (key0 + 63 > q0) || (key0 < ds_hi);
      const int hi0 = t - key0, lo0 = ds_t - key0;  // valid key window relative to the half tile's first key
      // row max of this half tile (masked): the reference of rows that have not seen a visible key yet, and the new
      // reference when a row sum overflows the range of the old one
      auto row_max = [&]() {
        float rmax = -INFINITY;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          uint32_t sr[32];
          tmem_ld_32x32b_x32(tS + c * 32, sr);
          tmem_ld_wait();
          rmax = need_mask ? chunk_rowmax<true>(sr, lo0 - c * 32, hi0 - c * 32, rmax)
                           : chunk_rowmax<false>(sr, 0, 0, rmax);
        }
        return rmax;
      };
      const bool first = __any_sync(0xffffffffu, ref == -INFINITY);
      if (first) {
        const float rmax = row_max();
        if (ref == -INFINITY) ref = rmax;
      }
      TRACE(0, hh * 4 + 2);
      bool pv_waited = (hh < 2);
      bool check = true;  // (rows that just took their reference from this half tile pass trivially)
      float2 psum;
      while (true) {
        const float mb = (ref == -INFINITY ? 0.f : ref) * c2;
        psum = make_float2(0.f, 0.f);
        float unused_max = 0.f;
        // the load of chunk 1 is in flight while chunk 0 is exponentiated
        uint32_t s0[32], s1[32], pk[16];
        tmem_ld_32x32b_x32(tS, s0);
        tmem_ld_wait_for(s0);
        tmem_ld_32x32b_x32(tS + 32, s1);
        if (need_mask) chunk_softmax<true, false>(s0, c2, mb, lo0, hi0, pk, psum, unused_max);
        else chunk_softmax<false, false>(s0, c2, mb, 0, 0, pk, psum, unused_max);
        if (!pv_waited) {  // P buffer u is still being read by P.V of half tile hh - 2 until this fires
          mbar_wait(bPVdone(u), ((hh >> 1) - 1) & 1);
          tc_fence_after();
          pv_waited = true;
        }
        tmem_st_32x32b_x16(tP, pk);
        tmem_ld_wait_for(s1);
        if (need_mask) chunk_softmax<true, false>(s1, c2, mb, lo0 - 32, hi0 - 32, pk, psum, unused_max);
        else chunk_softmax<false, false>(s1, c2, mb, 0, 0, pk, psum, unused_max);
        tmem_st_32x32b_x16(tP + 16, pk);
        if (!check) break;
        const bool bad = !(psum.x + psum.y <= kSumLimit);
        if (!__any_sync(0xffffffffu, bad)) break;
        // a score far beyond the reference: move the reference to this half tile's row max, rescale what has been
        // accumulated (every P.V issued so far must have completed first), redo the half tile
        const float ref_new = fmaxf(ref, row_max());
        const float alpha = ex2f((ref - ref_new) * c2);
        if (hh >= 1) mbar_wait(bPVdone((hh - 1) & 1), ((hh - 1) >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t orr[16];
          tmem_ld_32x32b_x16(tO + c * 16, orr);
          tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < 16; ++k) orr[k] = __float_as_uint(__uint_as_float(orr[k]) * alpha);
          tmem_st_32x32b_x16(tO + c * 16, orr);
        }
        tmem_st_wait();
        l_run *= alpha;
        ref = ref_new;
        check = false;
      }
      l_run += psum.x + psum.y;
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(bPfull(u));
      TRACE(0, hh * 4 + 3);
    }
    // ---- epilogue: O is complete in TMEM; normalise, store
    {
      mbar_wait(bPVdone((n_ht - 1) & 1), ((n_ht - 1) >> 1) & 1);  // MMAs complete in order: this is the last one
      tc_fence_after();
      const float inv_l = 1.f / l_run;
      uint32_t o0[32], o1[32];
      tmem_ld_32x32b_x32(tO, o0);
      tmem_ld_32x32b_x32(tO + 32, o1);
      tmem_ld_wait();
      // the 128 x 64 output tile leaves through the Q tile (all S MMAs are long done) and one TMA store: whole
      // 128-byte rows, nothing for the CTA's last instructions to wait on but the read of the tile
      const uint32_t orow = sQ + r * 128;
#pragma unroll
      for (int v = 0; v < 4; ++v) {
        uint4 o;
        o.x = pack_bf16x2(__uint_as_float(o0[v * 8 + 0]) * inv_l, __uint_as_float(o0[v * 8 + 1]) * inv_l);
        o.y = pack_bf16x2(__uint_as_float(o0[v * 8 + 2]) * inv_l, __uint_as_float(o0[v * 8 + 3]) * inv_l);
        o.z = pack_bf16x2(__uint_as_float(o0[v * 8 + 4]) * inv_l, __uint_as_float(o0[v * 8 + 5]) * inv_l);
        o.w = pack_bf16x2(__uint_as_float(o0[v * 8 + 6]) * inv_l, __uint_as_float(o0[v * 8 + 7]) * inv_l);
        st_shared_v4(orow + ((v ^ (r & 7)) << 4), o);
      }
#pragma unroll
      for (int v = 0; v < 4; ++v) {
        uint4 o;
        o.x = pack_bf16x2(__uint_as_float(o1[v * 8 + 0]) * inv_l, __uint_as_float(o1[v * 8 + 1]) * inv_l);
        o.y = pack_bf16x2(__uint_as_float(o1[v * 8 + 2]) * inv_l, __uint_as_float(o1[v * 8 + 3]) * inv_l);
        o.z = pack_bf16x2(__uint_as_float(o1[v * 8 + 4]) * inv_l, __uint_as_float(o1[v * 8 + 5]) * inv_l);
        o.w = pack_bf16x2(__uint_as_float(o1[v * 8 + 6]) * inv_l, __uint_as_float(o1[v * 8 + 7]) * inv_l);
        st_shared_v4(orow + (((4 + v) ^ (r & 7)) << 4), o);
      }
      fence_proxy_async_smem();
      named_bar_sync(1, 128);
      if (r == 0) {
        tma_store_2d(&tm_out, sQ, h * 64, static_cast<int>(seq_row0) + q0);
        tma_store_commit();
      }
      if (t < T) lse[(static_cast<long>(b) * NH + h) * T + t] = ref * kSoftmaxScale + logf(l_run);
      if (r == 0) tma_store_wait_read<0>();  // shared memory is given up when the CTA exits
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc<256>(tmem_base);
  }
#ifdef PCV_TRACE
  if (tid == 0 && cta_lin < 4096) {
    g_cta_trace[cta_lin * 4 + 1] = gtimer();
    g_cta_trace[cta_lin * 4 + 3] = n_it;
  }
#endif
}

// ================================================================================================ backward
// delta[b][h][t] = sum_c dout[b,t,h,c] * out[b,t,h,c]   (one warp per (token, head) pair would waste lanes: 8 lanes each)
// and, in the same pass over [B*T][d], the zero-fill of the fp32 dQ accumulator the main kernel reduce-adds into (a
// separate 64 MB memset node per layer cost 20 us at the 420M shape; here its stores ride under the loads).
__global__ void attn_bwd_prep_kernel(const __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ dout,
                                     float* __restrict__ delta, float* __restrict__ dq_acc, int BT, int T, int NH) {
  pdl_wait();
  pdl_trigger();
  const long gid = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x;
  const long item = gid >> 3;  // (row, head)
  const int c = static_cast<int>(gid & 7);
  if (item >= static_cast<long>(BT) * NH) return;
  const long row = item / NH;
  const int h = static_cast<int>(item % NH);
  const long off = row * (NH * 64) + h * 64 + c * 8;
  const uint4 a = ld_global_nc_v4(out + off), g = ld_global_nc_v4(dout + off);
  *reinterpret_cast<float4*>(dq_acc + off) = make_float4(0.f, 0.f, 0.f, 0.f);
  *reinterpret_cast<float4*>(dq_acc + off + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
  const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, gw[4] = {g.x, g.y, g.z, g.w};
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 x = unpack_bf16x2(aw[k]), y = unpack_bf16x2(gw[k]);
    s += x.x * y.x + x.y * y.y;
  }
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  if (c == 0) {
    const long bb = row / T;
    const int t = static_cast<int>(row % T);
    delta[(bb * NH + h) * T + t] = s;
  }
}

struct AttnBwdSmem {
  static constexpr int kK = 0;                      // 2 buffers: K_j rotated, of this item and (prefetched) of the next
  static constexpr int kV = kK + 2 * kTileBytes;    // V_j
  static constexpr int kStages = 3;                 // Q_i / dO_i ring
  static constexpr int kQ = kV + kTileBytes;        // kStages x rotated Q_i
  static constexpr int kdO = kQ + kStages * kTileBytes;
  static constexpr int kP = kdO + kStages * kTileBytes;  // 128 x 128 bf16
  static constexpr int kdS = kP + 2 * kTileBytes;   // 128 x 128 bf16
  static constexpr int kOut = kdS + 2 * kTileBytes;  // one staging tile of the dQ warpgroup: fp32 128 x 32 or bf16 128 x 64
  static constexpr int kBar = kOut + kTileBytes;    // 256 bytes of barriers, 64 of work-stealing answers
  static constexpr int kBytes = kBar + 512 + 1024;
};
static_assert(AttnBwdSmem::kBytes <= 232448, "attn_bwd shared memory");

template <uint32_t N>
__device__ __forceinline__ void reg_alloc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <uint32_t N>
__device__ __forceinline__ void reg_dealloc() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

// Backward CTA (512 threads), PERSISTENT: a CTA works on one (batch, head, 128-key tile) item after the other.  Its
// first item is its own blockIdx; further ones are taken from the CTAs of the grid that have not been launched yet
// (cluster launch control: the hardware cancels such a CTA and hands its blockIdx over), so the grid's launch order —
// per sequence, longest loops first — is the work queue and nothing is scheduled statically.  Tensor memory, barriers
// and descriptors are set up once per CTA, and across an item boundary the next item's K / V / Q / dO loads and its
// first S / dP MMAs run while the compute warps still write the finished item's dK / dV (one CTA per item spent
// 6.2k cycles before its first tile and 4.9k after its last: 27 % of the kernel at the 420M shape).
//   warps 0-7  : compute — thread (r, half) owns query row r, 64 of the 128 key columns of a tile:
//                S, dP (tensor memory) -> P = exp2(S*c - lse), dS = P*(dP - delta)*scale -> shared memory (P, dS: the
//                MN-major A operands of dV += P^T.dO, dK += dS^T.Q) and tensor memory (dS: the A operand of dQ = dS.K);
//                after an item's last tile: dK (un-rotated) and dV -> bf16 -> shared memory -> TMA store
//   warps 8-11 : dQ warpgroup — dQ of every tile from tensor memory through an fp32 staging tile and a TMA reduce-add
//   warp 12    : TMA producer (K, V per item; Q / dO ring of 3; the rotation-table rows of the key tile take one more
//                turn of the ring) and the work-stealing requests
//   warp 13    : tcgen05.mma issuer
// TMEM: S [0,128) dP [128,256) dV [256,320) dK [320,384) dQ [384,448) dS(bf16 pairs) [448,512).
// Barrier phases: `g` counts tiles over the CTA's whole life, `ring` counts turns of the Q / dO ring, `k` counts items.
__global__ void __launch_bounds__(kAttnBwdThreads, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_do,
                const __grid_constant__ CUtensorMap tm_dq, const __grid_constant__ CUtensorMap tm_cs,
                const __grid_constant__ CUtensorMap tm_dqkv, const float* __restrict__ cos_sin,
                const int32_t* __restrict__ doc_start, const float* __restrict__ lse, const float* __restrict__ delta,
                float* __restrict__ dq_acc, __nv_bfloat16* __restrict__ dqkv, int T, int NH) {
  extern __shared__ uint8_t smem_raw[];
  constexpr int kStages = AttnBwdSmem::kStages;
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sK0 = sbase + AttnBwdSmem::kK, sV = sbase + AttnBwdSmem::kV, sQ = sbase + AttnBwdSmem::kQ,
                 sdO = sbase + AttnBwdSmem::kdO, sP = sbase + AttnBwdSmem::kP, sdS = sbase + AttnBwdSmem::kdS,
                 sOut = sbase + AttnBwdSmem::kOut, sBar = sbase + AttnBwdSmem::kBar;
  const uint32_t bVfull = sBar;                                   // (tx: V)                            per item
  auto bQdOfull = [&](int s) { return sBar + 8u * (1 + s); };     // [3] (tx: Q_i and dO_i)             per ring turn
  auto bStageEmpty = [&](int s) { return sBar + 8u * (4 + s); };  // [3] (1: commit, or the dK/dV tail's arrive)
  const uint32_t bSdPfull = sBar + 8u * 8;                        // (1, commit)                        per tile
  const uint32_t bPdSfull = sBar + 8u * 9;                        // (256)                              per tile
  const uint32_t bdQfull = sBar + 8u * 10;                        // (1, commit)                        per tile
  const uint32_t bdQempty = sBar + 8u * 11;                       // (128)                              per tile
  const uint32_t bAccFull = sBar + 8u * 12;                       // (1, commit) dK/dV complete          per item
  const uint32_t tmem_ptr_smem = sBar + 8u * 13;
  const uint32_t bSdPfree = sBar + 8u * 14;                       // (256) S / dP of a tile are in registers
  const uint32_t bPfree = sBar + 8u * 15;                         // (1, commit) dV MMAs done: P in shared memory is free
  const uint32_t bVfree = sBar + 8u * 16;                         // (1, commit) the item's last dP MMA has read V
  const uint32_t bAccFree = sBar + 8u * 17;                       // (128) dK / dV have left tensor memory, per item
  auto bKfull = [&](int u) { return sBar + 8u * (18 + u); };      // [2] (tx: K) item k uses buffer k & 1
  // answers of the work-stealing requests: item k' -> slot k' & 3.  Four slots, because the dQ warpgroup reads the answer
  // for item k + 1 only after the dK / dV tail of item k, by when the producer may already have asked for item k + 3
  auto bItem = [&](int u) { return sBar + 8u * (20 + u); };       // [4] (tx: 16 bytes)
  auto bKfree = [&](int u) { return sBar + 8u * (24 + u); };      // [2] (1, commit) every MMA of the item on K buffer u is done
  auto sItem = [&](int u) { return sBar + 256u + 16u * u; };      // [4] x 16 bytes, 16-byte aligned
  volatile uint32_t* tmem_ptr_generic = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_ptr_smem - smem_u32(smem_raw)));

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#ifdef PCV_TRACE
  const bool trace_on = (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) && ((threadIdx.x & 127) == 0 || threadIdx.x == 416);
  const int cta_lin = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  if (tid == 0 && cta_lin < 4096) {
    unsigned int smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    g_cta_trace[cta_lin * 4 + 0] = gtimer();
    g_cta_trace[cta_lin * 4 + 2] = smid;
  }
  int trace_items = 0;
#endif
  const int d = NH * 64;
  const int n_qt = (T + 127) / 128;

  // An item = (h, j, b) in blockIdx terms.  Launch order (x fastest, z slowest): per sequence, key tile 0 of every head
  // first (all query tiles visible: the longest loop), the one-tile loops last.
  struct Item {
    int h, j, b, i_lo, n_it, key0;
    long seq_row0;
  };
  auto make_item = [&](int x, int y, int z) {
    Item w;
    w.h = x; w.j = y; w.b = z;
    w.seq_row0 = static_cast<long>(z) * T;
    w.key0 = y * 128;
    // query tiles that can see this key tile: i >= j (causal) and doc_start[first row of tile i] <= last key of tile j
    w.i_lo = y;
    int i_hi = n_qt - 1;
    if (doc_start) {
      const int key_last = min(w.key0 + 127, T - 1);
      // doc_start is non-decreasing in t: binary search the last tile whose first row's document starts <= key_last
      int lo = y, hi = n_qt - 1;
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (doc_start[w.seq_row0 + mid * 128] <= key_last) lo = mid; else hi = mid - 1;
      }
      i_hi = lo;
    }
    w.n_it = i_hi - w.i_lo + 1;  // >= 1: tile j always sees itself
    return w;
  };
  // item k + 1 of this CTA: the answer to the request issued while item k was loaded (every role reads it itself)
  auto next_item = [&](int k_next, Item& w) -> bool {
    mbar_wait(bItem(k_next & 3), ((k_next - 1) >> 2) & 1);  // slot u: items u, u + 4, ... (item 0 is never asked for)
    int x, y, z;
    if (!clc_query(sItem(k_next & 3), x, y, z)) return false;
    w = make_item(x, y, z);
    return true;
  };
  Item first = make_item(blockIdx.x, blockIdx.y, blockIdx.z);

  if (warp == 12 && lane == 0) {
    tma_prefetch_desc(&tm_qkv);
    tma_prefetch_desc(&tm_do);
    tma_prefetch_desc(&tm_dq);
    tma_prefetch_desc(&tm_cs);
    tma_prefetch_desc(&tm_dqkv);
    mbar_init(bVfull, 1);
    mbar_init(bKfull(0), 1);
    mbar_init(bKfull(1), 1);
    mbar_init(bVfree, 1);
    mbar_init(bKfree(0), 1);
    mbar_init(bKfree(1), 1);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bQdOfull(s), 1);
      mbar_init(bStageEmpty(s), 1);
    }
    mbar_init(bSdPfull, 1);
    mbar_init(bPdSfull, 256);
    mbar_init(bSdPfree, 256);
    mbar_init(bPfree, 1);
    mbar_init(bdQfull, 1);
    mbar_init(bdQempty, 128);
    mbar_init(bAccFull, 1);
    mbar_init(bAccFree, 128);
    for (int u = 0; u < 4; ++u) mbar_init(bItem(u), 1);
    fence_mbar_init();
  }
  if (warp == 12) {
    __syncwarp();
    tmem_alloc<512>(tmem_ptr_smem);
  }
  // programmatic dependent launch: everything above may have run under the tail of the preceding kernel (the prep pass)
  pdl_wait();
  pdl_trigger();
  if (warp == 12 && lane == 0) {
    // K, V and the first Q / dO tile of the first item are requested right away: their latency runs under the
    // block-wide barrier below
    const int row0 = static_cast<int>(first.seq_row0);
    mbar_expect_tx(bKfull(0), kTileBytes);
    tma_load_2d(sK0, &tm_qkv, bKfull(0), d + first.h * 64, row0 + first.key0);
    mbar_expect_tx(bVfull, kTileBytes);
    tma_load_2d(sV, &tm_qkv, bVfull, 2 * d + first.h * 64, row0 + first.key0);
    mbar_expect_tx(bQdOfull(0), 2 * kTileBytes);
    tma_load_2d(sQ, &tm_qkv, bQdOfull(0), first.h * 64, row0 + first.i_lo * 128);
    tma_load_2d(sdO, &tm_do, bQdOfull(0), first.h * 64, row0 + first.i_lo * 128);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_generic;
  const uint32_t tS = tmem_base, tdP = tmem_base + 128, tdV = tmem_base + 256, tdK = tmem_base + 320, tdQ = tmem_base + 384,
                 tdS = tmem_base + 448;  // dS as bf16 pairs: the A operand of dQ = dS.K, read from tensor memory

  if (warp >= 12) {
    reg_dealloc<56>();  // warpgroup 3: producer, MMA issuer, two idle warps (8 x 160 + 4 x 120 + 4 x 56 registers per lane: 63488 of 65536)
  }
  if (warp == 12) {
    // ============================== TMA producer (q, k in qkv are already rotated) ==============================
    if (lane == 0) {
      Item w = first;
      int ring = 0;
      for (int k = 0;; ++k) {
        const int row0 = static_cast<int>(w.seq_row0);
        if (k > 0) {
          // This point is reached while the previous item still has ~3 tiles to go (the Q / dO ring holds the producer
          // back), so both loads are prefetches.  K goes to the other K buffer, last read by item k - 2 ...
          // (a barrier per buffer: its next phase cannot complete before this very load has landed)
          if (k > 1) mbar_wait(bKfree(k & 1), ((k - 2) >> 1) & 1);
          mbar_expect_tx(bKfull(k & 1), kTileBytes);
          tma_load_2d(sK0 + (k & 1) * kTileBytes, &tm_qkv, bKfull(k & 1), d + w.h * 64, row0 + w.key0);
          // ... V to the single V buffer as soon as the previous item's last dP MMA has read it
          mbar_wait(bVfree, (k - 1) & 1);
          mbar_expect_tx(bVfull, kTileBytes);
          tma_load_2d(sV, &tm_qkv, bVfull, 2 * d + w.h * 64, row0 + w.key0);
        }
        // ask for the item after this one now: the answer is there long before it is needed
        mbar_expect_tx(bItem((k + 1) & 3), 16);
        clc_try_cancel(sItem((k + 1) & 3), bItem((k + 1) & 3));
        for (int it = (k == 0 ? 1 : 0); it < w.n_it; ++it) {
          const int i = w.i_lo + it;
          const int rr = ring + it;
          const int stage = rr % kStages;
          mbar_wait(bStageEmpty(stage), ((rr / kStages) & 1) ^ 1u);
          mbar_expect_tx(bQdOfull(stage), 2 * kTileBytes);
          tma_load_2d(sQ + stage * kTileBytes, &tm_qkv, bQdOfull(stage), w.h * 64, row0 + i * 128);
          tma_load_2d(sdO + stage * kTileBytes, &tm_do, bQdOfull(stage), w.h * 64, row0 + i * 128);
        }
        ring += w.n_it;
        {
          // one more turn of the ring: (cos, sin) of the 128 key rows, for the transposed rotation of dK at the end —
          // fetched two tiles ahead of its use instead of from global memory when every SM's tail is waiting on L2
          const int stage = ring % kStages;
          mbar_wait(bStageEmpty(stage), ((ring / kStages) & 1) ^ 1u);
          // (signalled on the stage's own full barrier, like every turn of the ring, so that its phases keep counting turns)
          mbar_expect_tx(bQdOfull(stage), 2 * kTileBytes);
          tma_load_2d(sQ + stage * kTileBytes, &tm_cs, bQdOfull(stage), 0, w.key0);     // pairs 0..15  (compute half 0)
          tma_load_2d(sdO + stage * kTileBytes, &tm_cs, bQdOfull(stage), 32, w.key0);   // pairs 16..31 (compute half 1)
          ++ring;
        }
        if (!next_item(k + 1, w)) break;
      }
    }
    __syncwarp();
  } else if (warp == 13) {
    // ============================== MMA issuer ==============================
    {
      const uint32_t lead = elect_one_sync();
      const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t tS = tb, tdP = tb + 128, tdV = tb + 256, tdK = tb + 320, tdQ = tb + 384, tdS = tb + 448;
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, false, false);   // S, dP
      constexpr uint32_t idesc_kv = umma_idesc_bf16(128, 64, true, true);     // dV = P^T.dO, dK = dS^T.Q
      constexpr uint32_t idesc_dq = umma_idesc_bf16(128, 64, false, true);    // dQ = dS.K
      uint32_t sK = sK0;  // this item's K buffer
      auto issue_SdP = [&](int stage, uint32_t k_buf) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t da = umma_smem_desc(sQ + stage * kTileBytes + kk * 32, 16, 1024);
          const uint64_t db = umma_smem_desc(k_buf + kk * 32, 16, 1024);
          umma_bf16_ss_p(tS, da, db, idesc_s, kk > 0 ? 1u : 0u, lead);
        }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t da = umma_smem_desc(sdO + stage * kTileBytes + kk * 32, 16, 1024);
          const uint64_t db = umma_smem_desc(sV + kk * 32, 16, 1024);
          umma_bf16_ss_p(tdP, da, db, idesc_s, kk > 0 ? 1u : 0u, lead);
        }
        tc_commit_p(bSdPfull, lead);
      };
      Item w = first, wn = first;
      bool first_issued = false;  // the item's first S / dP were already issued during the previous item's last tile
      int g = 0, ring = 0;
      for (int k = 0;; ++k) {
        sK = sK0 + (k & 1) * kTileBytes;
        if (!first_issued) {
          // first S / dP of the item: its operands have landed and the previous tile's S / dP are in registers
          mbar_wait(bKfull(k & 1), (k >> 1) & 1);
          mbar_wait(bVfull, k & 1);
          const int stage = ring % kStages;
          mbar_wait(bQdOfull(stage), (ring / kStages) & 1);
          if (g > 0) mbar_wait(bSdPfree, (g - 1) & 1);
          tc_fence_after();
          issue_SdP(stage, sK);
          if (w.n_it == 1) tc_commit_p(bVfree, lead);  // the item's last (only) dP: V may be replaced by the next item's
        }
        first_issued = false;
        bool have_next = false, next_valid = false;
        for (int it = 0; it < w.n_it; ++it, ++g) {
          const int rr = ring + it;
          const int stage = rr % kStages;
          TRACE(4, it * 2 + 0);
          // next tile's S and dP first: they are what the compute warps wait for.  They may overwrite S / dP as soon as
          // every compute thread holds its part of this tile in registers — about half of the tile's math still to go,
          // which is what hides the latency of these MMAs
          if (it + 1 < w.n_it) {
            const int s1 = (rr + 1) % kStages;
            mbar_wait(bQdOfull(s1), ((rr + 1) / kStages) & 1);
            mbar_wait(bSdPfree, g & 1);
            tc_fence_after();
            TRACE(5, it * 4 + 0);
            issue_SdP(s1, sK);
            if (it + 2 == w.n_it) tc_commit_p(bVfree, lead);  // that was the item's last dP
            TRACE(5, it * 4 + 1);
          } else {
            // last tile of the item: the "next tile" is the first one of the CTA's next item.  Its K was prefetched
            // into the other buffer, its V follows the dP MMA issued one tile ago; if both (and Q / dO) are there by
            // now, S / dP of the next item are computed under this tile's math as well — otherwise after this item
            have_next = true;
            next_valid = next_item(k + 1, wn);
            if (next_valid) {
              const int turn = ring + w.n_it + 1;  // (+1: this item's rotation-table turn)
              const int sn = turn % kStages;
              bool ready = false;
              for (int spin = 0; spin < 16 && !ready; ++spin)
                ready = mbar_try_wait(bVfull, (k + 1) & 1) && mbar_try_wait(bKfull((k + 1) & 1), ((k + 1) >> 1) & 1) &&
                        mbar_try_wait(bQdOfull(sn), (turn / kStages) & 1);
              ready = __shfl_sync(0xffffffffu, ready ? 1 : 0, 0) != 0;  // (one decision for the warp)
              if (ready) {
                mbar_wait(bSdPfree, g & 1);
                tc_fence_after();
                issue_SdP(sn, sK0 + ((k + 1) & 1) * kTileBytes);
                if (wn.n_it == 1) tc_commit_p(bVfree, lead);
                first_issued = true;
              }
            }
          }
          mbar_wait(bPdSfull, g & 1);  // P / dS of this tile are in shared memory and tensor memory
          tc_fence_after();
          TRACE(4, it * 2 + 1);
          mbar_wait(bdQempty, (g & 1) ^ 1u);  // previous tile's dQ has been read out of TMEM
          // a new item's dV / dK start from zero: the previous item's must have been read out of tensor memory
          if (it == 0 && k > 0) mbar_wait(bAccFree, (k - 1) & 1);
          tc_fence_after();
          TRACE(5, it * 4 + 2);
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {  // contraction over the 128 query rows, 16 per step
            // A = P^T (MN-major view of P[q][key]: two 64-key atoms 16 KB apart); B = dO (MN-major)
            const uint64_t da = umma_smem_desc(sP + kk * 2048, kTileBytes, 1024);
            const uint64_t db = umma_smem_desc(sdO + stage * kTileBytes + kk * 2048, kTileBytes, 1024);
            umma_bf16_ss_p(tdV, da, db, idesc_kv, (it > 0 || kk > 0) ? 1u : 0u, lead);
          }
          tc_commit_p(bPfree, lead);  // P may be overwritten while dK and dQ are still being computed
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint64_t da = umma_smem_desc(sdS + kk * 2048, kTileBytes, 1024);
            const uint64_t db = umma_smem_desc(sQ + stage * kTileBytes + kk * 2048, kTileBytes, 1024);
            umma_bf16_ss_p(tdK, da, db, idesc_kv, (it > 0 || kk > 0) ? 1u : 0u, lead);
          }
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {  // contraction over the 128 keys
            const uint64_t db = umma_smem_desc(sK + kk * 2048, kTileBytes, 1024);
            umma_bf16_ts_p(tdQ, tdS + kk * 8, db, idesc_dq, kk > 0 ? 1u : 0u, lead);
          }
          tc_commit_p(bdQfull, lead);
          tc_commit_p(bStageEmpty(stage), lead);
          TRACE(5, it * 4 + 3);
        }
        tc_commit_p(bAccFull, lead);
        tc_commit_p(bKfree(k & 1), lead);
        ring += w.n_it + 1;  // (+1: the rotation-table turn of the ring)
        if (!have_next) next_valid = next_item(k + 1, wn);
        if (!next_valid) break;
        w = wn;
      }
    }
    __syncwarp();
  } else if (warp >= 8 && warp < 12) {
    // ============================== dQ warpgroup (128 threads: one row each) ==============================
    // Everything that leaves tensor memory for global memory, off the compute warps' critical path:
    //   per tile : dQ -> fp32 staging tile (two 32-column halves, one after the other) -> TMA reduce-add into dq_acc
    //   per item : dV, dK (the transpose of the rotation applied) -> bf16 staging tile -> TMA store — while the compute
    //              warps and the MMA issuer are already working on the CTA's next item
    reg_dealloc<120>();  // (an explicit limit: ptxas would otherwise compile this branch under warpgroup 3's 56)
    const int r = tid & 127;
    const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t srow = sOut + r * 128;
    // one staging tile: wait until the TMA that last read it is done, fill it (row r, eight 16-byte pieces), hand over
    auto stage_rows = [&](const uint32_t* v32) {
      if (r == 0) tma_store_wait_read<0>();
      named_bar_sync(1, 128);
#pragma unroll
      for (int v = 0; v < 8; ++v)
        st_shared_v4(srow + ((v ^ (r & 7)) << 4), make_uint4(v32[v * 4], v32[v * 4 + 1], v32[v * 4 + 2], v32[v * 4 + 3]));
      fence_proxy_async_smem();
      named_bar_sync(1, 128);
    };
    Item w = first;
    int g = 0, ring = 0;
    for (int k = 0;; ++k) {
      for (int it = 0; it < w.n_it; ++it, ++g) {
        mbar_wait(bdQfull, g & 1);
        tc_fence_after();
        uint32_t q0r[32], q1r[32];
        tmem_ld_32x32b_x32(tdQ + lane_off, q0r);
        tmem_ld_32x32b_x32(tdQ + lane_off + 32, q1r);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(bdQempty);
        const int row0 = static_cast<int>(w.seq_row0) + (w.i_lo + it) * 128;
        stage_rows(q0r);
        if (r == 0) {
          tma_reduce_add_2d(&tm_dq, sOut, w.h * 64, row0);
          tma_store_commit();
        }
        stage_rows(q1r);
        if (r == 0) {
          tma_reduce_add_2d(&tm_dq, sOut, w.h * 64 + 32, row0);
          tma_store_commit();
        }
      }
      // ---- the item's dV and dK
      {
        const int cs_turn = ring + w.n_it;
        const int cs_stage = cs_turn % kStages;
        const int row0 = static_cast<int>(w.seq_row0) + w.key0;
        mbar_wait(bAccFull, k & 1);
        tc_fence_after();
        uint32_t a[32], o[32];  // (one 32-column row in flight at a time: this warpgroup has 128 registers per thread)
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          tmem_ld_32x32b_x32(tdV + lane_off + hf * 32, a);
          tmem_ld_wait();
#pragma unroll
          for (int c = 0; c < 16; ++c) o[hf * 16 + c] = pack_bf16x2(__uint_as_float(a[2 * c]), __uint_as_float(a[2 * c + 1]));
        }
        stage_rows(o);
        if (r == 0) {
          tma_store_2d(&tm_dqkv, sOut, 2 * d + w.h * 64, row0);
          tma_store_commit();
        }
        // (cos, sin) of key row r: the ring turn after the item's last Q / dO tile, 128B-swizzled rows of 32 floats
        // (pairs 0..15 in the Q slot, pairs 16..31 in the dO slot); dK gets the transpose of the rotation in fp32
        mbar_wait(bQdOfull(cs_stage), (cs_turn / kStages) & 1);
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          tmem_ld_32x32b_x32(tdK + lane_off + hf * 32, a);
          tmem_ld_wait();
          if (hf == 1) {
            tc_fence_before();
            mbar_arrive(bAccFree);  // the next item's dV / dK MMAs may overwrite the accumulators
          }
          const uint32_t cs_row = (hf == 0 ? sQ : sdO) + cs_stage * kTileBytes + r * 128;
#pragma unroll
          for (int v = 0; v < 4; ++v) {
            const uint4 u0 = ld_shared_v4(cs_row + (((2 * v) ^ (r & 7)) << 4));
            const uint4 u1 = ld_shared_v4(cs_row + (((2 * v + 1) ^ (r & 7)) << 4));
            const float4 c0 = make_float4(__uint_as_float(u0.x), __uint_as_float(u0.y), __uint_as_float(u0.z), __uint_as_float(u0.w));
            const float4 c1 = make_float4(__uint_as_float(u1.x), __uint_as_float(u1.y), __uint_as_float(u1.z), __uint_as_float(u1.w));
            const float g0 = __uint_as_float(a[v * 8 + 0]), g1 = __uint_as_float(a[v * 8 + 1]);
            const float g2 = __uint_as_float(a[v * 8 + 2]), g3 = __uint_as_float(a[v * 8 + 3]);
            const float g4 = __uint_as_float(a[v * 8 + 4]), g5 = __uint_as_float(a[v * 8 + 5]);
            const float g6 = __uint_as_float(a[v * 8 + 6]), g7 = __uint_as_float(a[v * 8 + 7]);
            o[hf * 16 + v * 4 + 0] = pack_bf16x2(g0 * c0.x + g1 * c0.y, g1 * c0.x - g0 * c0.y);
            o[hf * 16 + v * 4 + 1] = pack_bf16x2(g2 * c0.z + g3 * c0.w, g3 * c0.z - g2 * c0.w);
            o[hf * 16 + v * 4 + 2] = pack_bf16x2(g4 * c1.x + g5 * c1.y, g5 * c1.x - g4 * c1.y);
            o[hf * 16 + v * 4 + 3] = pack_bf16x2(g6 * c1.z + g7 * c1.w, g7 * c1.z - g6 * c1.w);
          }
        }
        stage_rows(o);  // (its barriers also order every thread's rotation-table reads before the arrive below)
        if (r == 0) {
          mbar_arrive(bStageEmpty(cs_stage));  // the ring slot of the rotation table is free
          tma_store_2d(&tm_dqkv, sOut, d + w.h * 64, row0);
          tma_store_commit();
        }
      }
      ring += w.n_it + 1;
      if (!next_item(k + 1, w)) break;
    }
    TRACE(7, 0);
    if (r == 0) tma_store_wait<0>();
    TRACE(7, 1);
  } else if (warp < 8) {
    // ============================== compute warps (256 threads: row r, column half `half`) ==============================
    reg_alloc<160>();
    const int r = tid & 127;
    const int half = tid >> 7;
    const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
    constexpr float c2 = kSoftmaxScale * kLog2e;
    Item w = first;
    int g = 0;
    for (int k = 0;; ++k) {
      const int b = w.b, h = w.h, i_lo = w.i_lo, n_it = w.n_it, key0 = w.key0;
      const long seq_row0 = w.seq_row0;
      // per-row statistics of a query tile; those of tile it + 1 are requested while tile it is being worked on, so
      // their global-memory latency stays off the loop's critical path
      float n_lse = 0.f, n_dlt = 0.f;
      int n_ds = 0, n_dshi = 0;
      auto load_stats = [&](int it) {
        const int i = i_lo + it;
        const int t = min(i * 128 + r, T - 1);
        const long stat_idx = (static_cast<long>(b) * NH + h) * T + t;
        n_lse = __ldg(lse + stat_idx);
        n_dlt = __ldg(delta + stat_idx);
        if (doc_start) {
          n_ds = __ldg(doc_start + seq_row0 + t);
          n_dshi = __ldg(doc_start + seq_row0 + min(i * 128 + 127, T - 1));
        }
      };
      load_stats(0);
      for (int it = 0; it < n_it; ++it, ++g) {
        const int i = i_lo + it;
        const int t = i * 128 + r;
        const float lse2 = n_lse * kLog2e;
        const float dlt_s = n_dlt * kSoftmaxScale;
        const int ds_t = n_ds;
        const int ds_hi_of_tile = n_dshi;
        const bool need_mask = (key0 + 127 > i * 128) || (key0 < ds_hi_of_tile);
        TRACE(half, it * 4 + 0);
        mbar_wait(bSdPfull, g & 1);
        tc_fence_after();
        if (it + 1 < n_it) load_stats(it + 1);
        TRACE(half, it * 4 + 1);
        // the math for this thread's 64 columns stays in registers ...
        uint32_t pk_p[2][16], pk_ds[2][16];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const int col0 = half * 64 + c * 32;
          uint32_t sr[32], pr[32];
          tmem_ld_32x32b_x32(tS + lane_off + col0, sr);
          tmem_ld_32x32b_x32(tdP + lane_off + col0, pr);
          tmem_ld_wait();
          if (c == 1) {  // this thread's share of S and dP has left tensor memory
            tc_fence_before();
            mbar_arrive(bSdPfree);
          }
          if (need_mask) chunk_p_ds<true>(sr, pr, c2, lse2, dlt_s, ds_t - key0 - col0, t - key0 - col0, pk_p[c], pk_ds[c]);
          else chunk_p_ds<false>(sr, pr, c2, lse2, dlt_s, 0, 0, pk_p[c], pk_ds[c]);
          // P leaves for shared memory chunk by chunk, under the math of the next chunk: its buffer is free as soon as
          // the previous tile's dV MMAs are done (dS has to wait for dK and dQ as well)
          if (c == 0 && g > 0) mbar_wait(bPfree, (g - 1) & 1);
          {
            const uint32_t rowoff = half * kTileBytes + r * 128;
#pragma unroll
            for (int v = 0; v < 4; ++v) {
              const uint32_t off = rowoff + (((c * 4 + v) ^ (r & 7)) << 4);
              st_shared_v4(sP + off, make_uint4(pk_p[c][v * 4], pk_p[c][v * 4 + 1], pk_p[c][v * 4 + 2], pk_p[c][v * 4 + 3]));
            }
          }
        }
        // ... until the previous tile's dV/dK/dQ MMAs (which read P and dS from shared memory) have completed
        TRACE(half, it * 4 + 2);
        if (g > 0) mbar_wait(bdQfull, (g - 1) & 1);
        TRACE(half, it * 4 + 3);
        {
          const uint32_t rowoff = half * kTileBytes + r * 128;
#pragma unroll
          for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int v = 0; v < 4; ++v) {
              const uint32_t off = rowoff + (((c * 4 + v) ^ (r & 7)) << 4);
              st_shared_v4(sdS + off, make_uint4(pk_ds[c][v * 4], pk_ds[c][v * 4 + 1], pk_ds[c][v * 4 + 2], pk_ds[c][v * 4 + 3]));
            }
        }
        TRACE(2 + half, it * 4 + 0);
        // dS a second time, into tensor memory (32 columns of bf16 pairs per thread), for dQ = dS.K
        tmem_st_32x32b_x16(tdS + lane_off + half * 32, pk_ds[0]);
        tmem_st_32x32b_x16(tdS + lane_off + half * 32 + 16, pk_ds[1]);
        tmem_st_wait();
        TRACE(2 + half, it * 4 + 1);
        tc_fence_before();
        fence_proxy_async_smem();
        TRACE(2 + half, it * 4 + 2);
        mbar_arrive(bPdSfull);
        TRACE(2 + half, it * 4 + 3);
      }
      // (the item's dK / dV leave through the dQ warpgroup: these warps go straight on to the next item)
#ifdef PCV_TRACE
      ++trace_items;
#endif
      if (!next_item(k + 1, w)) break;
    }
  }
  tc_fence_before();
  __syncthreads();
  TRACE(6, 3);
  if (warp == 12) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
#ifdef PCV_TRACE
  TRACE(6, 4);
  if (tid == 0 && cta_lin < 4096) {
    g_cta_trace[cta_lin * 4 + 1] = gtimer();
    g_cta_trace[cta_lin * 4 + 3] = trace_items;
  }
#endif
}

// dqkv[:, q columns] = bf16( R(-theta) . dq_acc )    (dq_acc fp32 [B*T][d])
__global__ void attn_bwd_dq_post_kernel(const float* __restrict__ dq_acc, const float* __restrict__ cos_sin,
                                        __nv_bfloat16* __restrict__ dqkv, long BT, int T, int d) {
  pdl_wait();
  pdl_trigger();
  const int vpr = d / 8;
  const long total = BT * vpr;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const long row = i / vpr;
    const int col = static_cast<int>(i % vpr) * 8;
    const int t = static_cast<int>(row % T);
    const int fc = (col & 63) >> 1;  // first frequency index of this chunk
    const float4 a = *reinterpret_cast<const float4*>(dq_acc + row * d + col);
    const float4 bq = *reinterpret_cast<const float4*>(dq_acc + row * d + col + 4);
    const float4* cs = reinterpret_cast<const float4*>(cos_sin + (static_cast<long>(t) * 32 + fc) * 2);
    const float4 c0 = __ldg(cs), c1 = __ldg(cs + 1);
    uint4 o;
    o.x = pack_bf16x2(a.x * c0.x + a.y * c0.y, a.y * c0.x - a.x * c0.y);
    o.y = pack_bf16x2(a.z * c0.z + a.w * c0.w, a.w * c0.z - a.z * c0.w);
    o.z = pack_bf16x2(bq.x * c1.x + bq.y * c1.y, bq.y * c1.x - bq.x * c1.y);
    o.w = pack_bf16x2(bq.z * c1.z + bq.w * c1.w, bq.w * c1.z - bq.z * c1.w);
    *reinterpret_cast<uint4*>(dqkv + row * 3 * d + col) = o;
  }
}


// In-place interleaved-pair RoPE of the q and k columns of qkv[B*T][3d] (models/LM/embedding.py:28-67); the model
// path fuses this into the QKV GEMM epilogue instead (pcv_gemm_bf16_rope).
__global__ void rope_qk_kernel(__nv_bfloat16* __restrict__ qkv, const float* __restrict__ cos_sin, long BT, int T, int d) {
  const int vpr = 2 * d / 8;  // 16-byte chunks in the q|k part of a row
  const long total = BT * vpr;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const long row = i / vpr;
    const int col = static_cast<int>(i % vpr) * 8;
    const int t = static_cast<int>(row % T);
    const int fc = (col & 63) >> 1;
    uint4* ptr = reinterpret_cast<uint4*>(qkv + row * 3 * d + col);
    const float4* cs = reinterpret_cast<const float4*>(cos_sin + (static_cast<long>(t) * 32 + fc) * 2);
    *ptr = rope_chunk(*ptr, __ldg(cs), __ldg(cs + 1), 1.f);
  }
}

}  // namespace pcv

using namespace pcv;

static int attn_check(int B, int T, int NH) {
  if (B <= 0 || T <= 0 || NH <= 0) return PCV_ERR_ARG;
  if (T % 128) return PCV_ERR_ARG;  // v1: whole 128-row tiles
  return PCV_OK;
}

#ifdef PCV_TRACE
extern "C" int pcv_debug_read_trace(long long* host, int n) {
  return static_cast<int>(cudaMemcpyFromSymbol(host, pcv::g_trace, sizeof(long long) * n));
}
extern "C" int pcv_debug_clear_cta_trace(void) {
  void* p = nullptr;
  cudaGetSymbolAddress(&p, pcv::g_cta_trace);
  return static_cast<int>(cudaMemset(p, 0, sizeof(long long) * 4096 * 4));
}
extern "C" int pcv_debug_read_cta_trace(long long* host, int n) {
  return static_cast<int>(cudaMemcpyFromSymbol(host, pcv::g_cta_trace, sizeof(long long) * n));
}
#endif

extern "C" int pcv_rope_qk(void* qkv, const float* cos_sin, int B, int T, int n_heads, void* stream) {
  if (!qkv || !cos_sin) return PCV_ERR_ARG;
  int rc = attn_check(B, T, n_heads);
  if (rc) return rc;
  const int d = n_heads * 64;
  const long BT = static_cast<long>(B) * T;
  const long total = BT * (2 * d / 8);
  long blocks = (total + 255) / 256;
  const long cap = static_cast<long>(num_sms()) * 16;
  if (blocks > cap) blocks = cap;
  rope_qk_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<__nv_bfloat16*>(qkv), cos_sin, BT, T, d);
  count_launch();
  return cuda_rc();
}

extern "C" int pcv_attn_fwd(const void* qkv, const int32_t* doc_start, void* out, float* lse, int B, int T, int n_heads,
                            void* stream) {
  if (!qkv || !out || !lse) return PCV_ERR_ARG;
  int rc = attn_check(B, T, n_heads);
  if (rc) return rc;
  const int d = n_heads * 64;
  CUtensorMap tm;
  rc = get_tmap_2d_bf16(&tm, qkv, static_cast<uint64_t>(B) * T, 3 * d, 3 * d, 128);
  if (rc) return rc;
  CUtensorMap tm_out;
  rc = get_tmap_2d_bf16(&tm_out, out, static_cast<uint64_t>(B) * T, d, d, 128);
  if (rc) return rc;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnFwdSmem::kBytes);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr = true;
  }
  dim3 grid(n_heads, B, T / 128);
  cudaError_t le = launch_pdl(attn_fwd_kernel, grid, dim3(kAttnThreads), AttnFwdSmem::kBytes, static_cast<cudaStream_t>(stream),
                              tm, tm_out, doc_start, static_cast<__nv_bfloat16*>(out), lse, T, n_heads);
  count_launch();
  return le == cudaSuccess ? cuda_rc() : static_cast<int>(le);
}

extern "C" int pcv_attn_bwd(const void* qkv, const float* cos_sin, const int32_t* doc_start, const void* out,
                            const void* dout, const float* lse, float* delta, float* dq_acc, void* dqkv, int B, int T,
                            int n_heads, void* stream) {
  if (!qkv || !cos_sin || !out || !dout || !lse || !delta || !dq_acc || !dqkv) return PCV_ERR_ARG;
  int rc = attn_check(B, T, n_heads);
  if (rc) return rc;
  const int d = n_heads * 64;
  const long BT = static_cast<long>(B) * T;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CUtensorMap tm_qkv, tm_do;
  rc = get_tmap_2d_bf16(&tm_qkv, qkv, BT, 3 * d, 3 * d, 128);
  if (rc) return rc;
  rc = get_tmap_2d_bf16(&tm_do, dout, BT, d, d, 128);
  if (rc) return rc;
  CUtensorMap tm_dq;
  rc = get_tmap_2d_f32(&tm_dq, dq_acc, BT, d, d, 128);
  if (rc) return rc;
  CUtensorMap tm_cs, tm_dqkv;
  rc = get_tmap_2d_f32(&tm_cs, cos_sin, T, 64, 64, 128);  // rows = positions, 32 x (cos, sin); boxes of 32 floats
  if (rc) return rc;
  rc = get_tmap_2d_bf16(&tm_dqkv, dqkv, BT, 3 * d, 3 * d, 128);
  if (rc) return rc;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnBwdSmem::kBytes);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr = true;
  }
  {
    const long threads = BT * n_heads * 8;
    cudaError_t le = launch_pdl(attn_bwd_prep_kernel, dim3(static_cast<unsigned>((threads + 255) / 256)), dim3(256), 0, st,
                                static_cast<const __nv_bfloat16*>(out), static_cast<const __nv_bfloat16*>(dout), delta, dq_acc,
                                static_cast<int>(BT), T, n_heads);
    if (le != cudaSuccess) return static_cast<int>(le);
  }
  dim3 grid(n_heads, T / 128, B);
  {
    cudaError_t le = launch_pdl(attn_bwd_kernel, grid, dim3(kAttnBwdThreads), AttnBwdSmem::kBytes, st, tm_qkv, tm_do, tm_dq, tm_cs,
                                tm_dqkv, cos_sin, doc_start, lse, static_cast<const float*>(delta), dq_acc,
                                static_cast<__nv_bfloat16*>(dqkv), T, n_heads);
    if (le != cudaSuccess) return static_cast<int>(le);
  }
  {
    const long total = BT * (d / 8);
    long blocks = (total + 255) / 256;
    const long cap = static_cast<long>(num_sms()) * 16;
    if (blocks > cap) blocks = cap;
    cudaError_t le = launch_pdl(attn_bwd_dq_post_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), 0, st,
                                static_cast<const float*>(dq_acc), cos_sin, static_cast<__nv_bfloat16*>(dqkv), BT, T, d);
    if (le != cudaSuccess) return static_cast<int>(le);
  }
  count_launch(3);
  return cuda_rc();
}

// ------------------------------------------------------------------------------------------------ no-allocation variant
// pcv_attn_bwd_static: pcv_attn_bwd for callers that cannot allocate device memory (code under CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  The two scratch tensors of the backward pass — the fp32 dQ
// accumulator [B*T][d] and delta [B][N][T] — live in a file-scope device array instead of caller-supplied buffers.
// No counterpart in the reference.
//   * Limit: B*T*d <= 4,194,304 floats (e.g. 4096 tokens at d = 1024) — 16.25 MiB of device memory, held by every GPU
//     that loads the library until the process exits.  Larger problems get PCV_ERR_SCRATCH.
//   * Every call shares the array and finds the previous call's contents in it: the dQ accumulator is zero-filled by the
//     prep kernel of each call before the main kernel reduce-adds into it, delta is overwritten in full.  Two calls that
//     overlap on the GPU (different streams, no dependency) RACE on it: serialise them.
//   * The entry point calls no cudaMalloc / cudaFree / allocating library routine.
namespace pcv {
constexpr size_t kBwdStaticDqFloats = size_t(4) << 20;       // dQ accumulator: B*T*d floats
constexpr size_t kBwdStaticDeltaFloats = size_t(64) << 10;   // delta: B*N*T = B*T*d / 64 floats
static __device__ __align__(256) float g_attn_bwd_scratch[kBwdStaticDqFloats + kBwdStaticDeltaFloats];

// In host code `g_attn_bwd_scratch` is only a host-side stand-in for the symbol, not a device address: the address the
// kernels need is asked for once per device, with that device current.
static int attn_bwd_scratch_ptr(float** out) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static float* ptrs[kMaxDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return static_cast<int>(e);
  if (dev < 0 || dev >= kMaxDevices) return PCV_ERR_ARG;
  std::lock_guard<std::mutex> g(mu);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_attn_bwd_scratch);
    if (e != cudaSuccess) return static_cast<int>(e);
    if (p == nullptr || (reinterpret_cast<uintptr_t>(p) & 255)) return PCV_ERR_ALIGN;
    ptrs[dev] = static_cast<float*>(p);
  }
  *out = ptrs[dev];
  return PCV_OK;
}
}  // namespace pcv

extern "C" int pcv_attn_bwd_static(const void* qkv, const float* cos_sin, const int32_t* doc_start, const void* out,
                                   const void* dout, const float* lse, void* dqkv, int B, int T, int n_heads,
                                   void* stream) {
  int rc = attn_check(B, T, n_heads);
  if (rc) return rc;
  const size_t dq_floats = static_cast<size_t>(B) * T * n_heads * 64;
  const size_t delta_floats = static_cast<size_t>(B) * T * n_heads;
  if (dq_floats > pcv::kBwdStaticDqFloats || delta_floats > pcv::kBwdStaticDeltaFloats) return PCV_ERR_SCRATCH;
  float* base = nullptr;
  rc = pcv::attn_bwd_scratch_ptr(&base);
  if (rc) return rc;
  // dq_acc at the start of the array (TMA reduce-add target: 256-byte aligned), delta behind its maximum extent
  return pcv_attn_bwd(qkv, cos_sin, doc_start, out, dout, lse, base + pcv::kBwdStaticDqFloats, base, dqkv, B, T, n_heads,
                      stream);
}
