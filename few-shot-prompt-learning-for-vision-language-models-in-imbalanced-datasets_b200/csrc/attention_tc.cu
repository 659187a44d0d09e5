// tcgen05 attention for the bf16 fast mode: head_dim 64, 32 < L <= 256 (vision 197/201 tokens, dense text 77).
//
// One CTA per (sequence, head).  Q, K, V (and dO) tiles are fetched by TMA (3-D tensor map over
// [n_seq][L][3D], rows past L zero-filled) into 128B-swizzled shared memory and never touched by a thread.
// Every contraction runs on the 5th-gen tensor cores with the accumulator in TMEM:
//   forward   S = Q K^T  (SS MMA, 128 x Lp x 64)  ->  one thread per query row does the softmax straight out of
//             TMEM (no shuffles), writes P as packed bf16 back INTO TMEM  ->  O = P V  (TS MMA: A from TMEM,
//             B = V as an MN-major shared-memory operand)  ->  O / rowsum, log-sum-exp.
//   backward  two orientations, both recomputing P from the saved log-sum-exp, no atomics, deterministic:
//             queries on lanes:  S = Q K^T, dP = dO V^T  ->  dS = P (dP - delta)  ->  dQ = dS K
//             keys on lanes:     S^T = K Q^T, dP^T = V dO^T  ->  P^T, dS^T  ->  dV = P^T dO, dK = dS^T Q
// Two CTAs share an SM in the forward (256 TMEM columns each) so one head's softmax overlaps the other's MMAs.
// Reference: nn.MultiheadAttention core called from PromptSRC/clip/model.py:183/227/285, causal mask
// model.py:592-598.
#include <cuda.h>
#include <stdlib.h>
#include "common.cuh"

namespace psrc {
namespace atc {

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr int kTile = 128 * 128;  // bytes of one 128-row x 64-dim bf16 tile
constexpr uint32_t kSpinLimit = 1u << 27;

__device__ __forceinline__ void wait_bounded(uint64_t* bar, uint32_t parity) {
  for (uint32_t i = 0; !mbar_try_wait(bar, parity); ++i) {
    if (i > kSpinLimit) __trap();
  }
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
// K-major operand ([rows][64] bf16, 128-byte rows, SWIZZLE_128B): 8-row atoms 1024 B apart
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(1024u >> 4) << 32) | ((uint64_t)1 << 46) |
         ((uint64_t)2 << 61);
}
// MN-major B operand: the same [k rows][64 n] tile read with N contiguous; 8-k-row groups 1024 B apart (SBO),
// a single 64-element atom along N (LBO unused)
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(1024u >> 4) << 16) | ((uint64_t)(1024u >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
constexpr uint32_t kIdescBase = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(128 >> 4) << 24);  // f32 acc, bf16 a/b, M=128
__device__ __forceinline__ uint32_t idesc_kk(int n) { return kIdescBase | ((uint32_t)(n >> 3) << 17); }
constexpr uint32_t kIdescMn64 = kIdescBase | (1u << 16) | ((uint32_t)(64 >> 3) << 17);  // B MN-major, N = 64


// ---- softmax helpers: W columns of one TMEM lane per call, four independent max / sum chains (a single chain of
//      ~200 dependent FMNMX / FADD is latency-bound with one softmax warp per scheduler), no per-element predicates
//      unless the chunk straddles the row's valid length
template <int W, bool MASK>
__device__ __forceinline__ void chunk_max(const uint32_t (&r)[W], int c0, int ncol, float (&m)[4]) {
#pragma unroll
  for (int i = 0; i < W; ++i) {
    float v = __uint_as_float(r[i]);
    if (MASK && c0 + i >= ncol) v = -INFINITY;
    m[i & 3] = fmaxf(m[i & 3], v);
  }
}
template <int W, bool MASK>
__device__ __forceinline__ void chunk_exp(const uint32_t (&r)[W], uint32_t (&pk)[W / 2], int c0, int ncol,
                                          float scale_log2, float ms, float (&sum)[4]) {
#pragma unroll
  for (int i = 0; i < W / 2; ++i) {
    float p0 = ex2(fmaf(__uint_as_float(r[2 * i]), scale_log2, -ms));
    float p1 = ex2(fmaf(__uint_as_float(r[2 * i + 1]), scale_log2, -ms));
    if (MASK) {
      if (c0 + 2 * i >= ncol) p0 = 0.f;
      if (c0 + 2 * i + 1 >= ncol) p1 = 0.f;
    }
    sum[(2 * i) & 3] += p0;
    sum[(2 * i + 1) & 3] += p1;
    pk[i] = pack2(p0, p1);
  }
}

// ------------------------------------------------------------------------------------------- forward
constexpr int kFwdThreads = 160;  // warps 0-3: softmax + epilogue (one TMEM lane each); warp 4: TMEM/TMA/MMA
constexpr uint32_t kFwdO = 128;   // TMEM column of the O accumulator (S at 0, P packed over S)

template <bool kCausal>
__global__ void __launch_bounds__(kFwdThreads)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmap_qkv, __nv_bfloat16* __restrict__ out,
                   float* __restrict__ lse, int L, int H, float scale_log2) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int n_mt = (L + 127) >> 7;
  const int Lp = (L + 15) & ~15;
  uint8_t* Qs = smem;
  uint8_t* Ks = Qs + n_mt * kTile;
  uint8_t* Vs = Ks + n_mt * kTile;
  uint64_t* bars = reinterpret_cast<uint64_t*>(Vs + n_mt * kTile);
  uint64_t *bar_load = bars, *bar_s = bars + 1, *bar_p = bars + 2, *bar_o = bars + 3, *bar_e = bars + 4;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bars + 5);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.x % H, seq = blockIdx.x / H;   // linear grid: no 65 535 limit on sequences
  const int D = H * 64;

  if (warp == 4) {
    if (lane == 0) {
      tma_prefetch_desc(&tmap_qkv);
      mbar_init(bar_load, 1);
      mbar_init(bar_s, 1);
      mbar_init(bar_p, 4);
      mbar_init(bar_o, 1);
      mbar_init(bar_e, 4);
      fence_mbar_init();
      mbar_expect_tx(bar_load, 3 * n_mt * kTile);   // the loads do not depend on the TMEM allocation
      for (int t = 0; t < n_mt; ++t) {
        tma_load_3d(&tmap_qkv, bar_load, Qs + t * kTile, h * 64, t * 128, seq);
        tma_load_3d(&tmap_qkv, bar_load, Ks + t * kTile, D + h * 64, t * 128, seq);
        tma_load_3d(&tmap_qkv, bar_load, Vs + t * kTile, 2 * D + h * 64, t * 128, seq);
      }
    }
    __syncwarp();
    tmem_alloc(tslot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = *tslot;

  if (warp == 4) {
    if (lane == 0) {
      wait_bounded(bar_load, 0);
      const uint32_t idesc_s = idesc_kk(Lp);
      const uint32_t ks = smem_u32(Ks), vs = smem_u32(Vs);
      for (int t = 0; t < n_mt; ++t) {
        if (t > 0) {
          wait_bounded(bar_e, (t - 1) & 1);
          tc_fence_after();
        }
        const uint32_t qs = smem_u32(Qs + t * kTile);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(tb, desc_kmajor(qs + k * 32), desc_kmajor(ks + k * 32), idesc_s, k > 0);
        umma_commit(bar_s);
        wait_bounded(bar_p, t & 1);
        tc_fence_after();
        for (int k = 0; k < (Lp >> 4); ++k)
          umma_bf16_ts(tb + kFwdO, tb + 8 * k, desc_mnmajor(vs + k * 2048), kIdescMn64, k > 0);
        umma_commit(bar_o);
      }
    }
  } else {
    const uint32_t tl = tb + ((uint32_t)(warp * 32) << 16);
    for (int t = 0; t < n_mt; ++t) {
      const int row = t * 128 + warp * 32 + lane;
      wait_bounded(bar_s, t & 1);
      tc_fence_after();
      int ncol = L;
      if (kCausal) ncol = min(L, row + 1);
      // columns any lane of this warp needs (warp-uniform loop bound)
      int cmax = Lp;
      if (kCausal) cmax = min(Lp, (t * 128 + warp * 32 + 32 + 15) & ~15);
      float ms, sum;
      if (!kCausal) {
        // ---- fast path (vision tower): every row sees all L keys; only the last chunk straddles L
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        int c0 = 0;
        for (; c0 + 64 <= Lp; c0 += 64) {          // two loads in flight per wait
          uint32_t r0[32], r1[32];
          tmem_ld_32x32(tl + c0, r0);
          tmem_ld_32x32(tl + c0 + 32, r1);
          tmem_ld_wait();
          if (c0 + 64 <= L) {
            chunk_max<32, false>(r0, c0, L, mx);
            chunk_max<32, false>(r1, c0 + 32, L, mx);
          } else {
            chunk_max<32, true>(r0, c0, L, mx);
            chunk_max<32, true>(r1, c0 + 32, L, mx);
          }
        }
        if (c0 + 32 <= Lp) {
          uint32_t r0[32];
          tmem_ld_32x32(tl + c0, r0);
          tmem_ld_wait();
          if (c0 + 32 <= L) chunk_max<32, false>(r0, c0, L, mx); else chunk_max<32, true>(r0, c0, L, mx);
          c0 += 32;
        }
        if (c0 < Lp) {
          uint32_t r0[16];
          tmem_ld_32x16(tl + c0, r0);
          tmem_ld_wait();
          chunk_max<16, true>(r0, c0, L, mx);
        }
        ms = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * scale_log2;
        float sm4[4] = {0.f, 0.f, 0.f, 0.f};
        c0 = 0;
        for (; c0 + 32 <= Lp; c0 += 32) {
          uint32_t r0[32], pk[16];
          tmem_ld_32x32(tl + c0, r0);
          tmem_ld_wait();
          if (c0 + 32 <= L) chunk_exp<32, false>(r0, pk, c0, L, scale_log2, ms, sm4);
          else chunk_exp<32, true>(r0, pk, c0, L, scale_log2, ms, sm4);
          tmem_st_32x16(tl + (c0 >> 1), pk);
        }
        if (c0 < Lp) {
          uint32_t r0[16], pk[8];
          tmem_ld_32x16(tl + c0, r0);
          tmem_ld_wait();
          chunk_exp<16, true>(r0, pk, c0, L, scale_log2, ms, sm4);
          tmem_st_32x8(tl + (c0 >> 1), pk);
        }
        sum = (sm4[0] + sm4[1]) + (sm4[2] + sm4[3]);
      } else {
      // ---- pass 1: row maximum
      float m = -INFINITY;
      for (int c0 = 0; c0 < cmax; c0 += 32) {
        if (c0 + 32 <= cmax) {
          uint32_t r[32];
          tmem_ld_32x32(tl + c0, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c0 + i < ncol) m = fmaxf(m, __uint_as_float(r[i]));
        } else {
          uint32_t r[16];
          tmem_ld_32x16(tl + c0, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (c0 + i < ncol) m = fmaxf(m, __uint_as_float(r[i]));
        }
      }
      ms = m * scale_log2;
      // ---- pass 2: P = exp2(S*c - m*c) as packed bf16 over the S columns, row sum in fp32
      sum = 0.f;
      for (int c0 = 0; c0 < Lp; c0 += 32) {
        if (c0 + 32 <= Lp) {
          uint32_t r[32], pk[16];
          if (c0 < cmax) {
            tmem_ld_32x32(tl + c0, r);
            tmem_ld_wait();
          }
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float p0 = 0.f, p1 = 0.f;
            if (c0 + 2 * i < ncol) p0 = ex2(fmaf(__uint_as_float(r[2 * i]), scale_log2, -ms));
            if (c0 + 2 * i + 1 < ncol) p1 = ex2(fmaf(__uint_as_float(r[2 * i + 1]), scale_log2, -ms));
            sum += p0 + p1;
            pk[i] = pack2(p0, p1);
          }
          tmem_st_32x16(tl + (c0 >> 1), pk);
        } else {
          uint32_t r[16], pk[8];
          if (c0 < cmax) {
            tmem_ld_32x16(tl + c0, r);
            tmem_ld_wait();
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float p0 = 0.f, p1 = 0.f;
            if (c0 + 2 * i < ncol) p0 = ex2(fmaf(__uint_as_float(r[2 * i]), scale_log2, -ms));
            if (c0 + 2 * i + 1 < ncol) p1 = ex2(fmaf(__uint_as_float(r[2 * i + 1]), scale_log2, -ms));
            sum += p0 + p1;
            pk[i] = pack2(p0, p1);
          }
          tmem_st_32x8(tl + (c0 >> 1), pk);
        }
      }
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p);
      // ---- epilogue: O / sum, log-sum-exp
      wait_bounded(bar_o, t & 1);
      tc_fence_after();
      uint32_t o0[32], o1[32];
      tmem_ld_32x32(tl + kFwdO, o0);
      tmem_ld_32x32(tl + kFwdO + 32, o1);
      tmem_ld_wait();
      // O is in registers: hand the TMEM tile back to the MMA warp before the global stores
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_e);
      if (row < L) {
        const float inv = 1.f / sum;
        uint4* dst = reinterpret_cast<uint4*>(out + ((long long)seq * L + row) * D + h * 64);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 u;
          u.x = pack2(__uint_as_float(o0[8 * j]) * inv, __uint_as_float(o0[8 * j + 1]) * inv);
          u.y = pack2(__uint_as_float(o0[8 * j + 2]) * inv, __uint_as_float(o0[8 * j + 3]) * inv);
          u.z = pack2(__uint_as_float(o0[8 * j + 4]) * inv, __uint_as_float(o0[8 * j + 5]) * inv);
          u.w = pack2(__uint_as_float(o0[8 * j + 6]) * inv, __uint_as_float(o0[8 * j + 7]) * inv);
          dst[j] = u;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 u;
          u.x = pack2(__uint_as_float(o1[8 * j]) * inv, __uint_as_float(o1[8 * j + 1]) * inv);
          u.y = pack2(__uint_as_float(o1[8 * j + 2]) * inv, __uint_as_float(o1[8 * j + 3]) * inv);
          u.z = pack2(__uint_as_float(o1[8 * j + 4]) * inv, __uint_as_float(o1[8 * j + 5]) * inv);
          u.w = pack2(__uint_as_float(o1[8 * j + 6]) * inv, __uint_as_float(o1[8 * j + 7]) * inv);
          dst[4 + j] = u;
        }
        lse[((long long)seq * H + h) * L + row] = (ms + log2f(sum)) * kLn2;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 4) tmem_dealloc(tb, 256);
}

// ------------------------------------------------------------------------------------------- backward
constexpr int kBwdThreads = 288;  // warps 0-7: elementwise + epilogue (lane quarter = w & 3, column half = w >> 2)
constexpr uint32_t kColS = 0, kColDP = 256;      // fp32 score / dP accumulators (Lp <= 208... 256 columns each)
// 64-column output accumulators are placed over score columns that are dead once P / dS are packed: the first
// warp's range [0, sp) keeps packed values in [0, sp/2), the second's [sp, Lp) in [sp, sp + (Lp - sp)/2)
__device__ __forceinline__ uint32_t bwd_out_col(int Lp, int sp) {
  return (uint32_t)(sp >= 128 ? (sp >> 1) : sp + ((Lp - sp) >> 1));
}

__device__ __forceinline__ void store_row64(__nv_bfloat16* dst, const uint32_t (&a)[32], const uint32_t (&b)[32],
                                            float s) {
  uint4* d = reinterpret_cast<uint4*>(dst);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint4 u;
    u.x = pack2(__uint_as_float(a[8 * j]) * s, __uint_as_float(a[8 * j + 1]) * s);
    u.y = pack2(__uint_as_float(a[8 * j + 2]) * s, __uint_as_float(a[8 * j + 3]) * s);
    u.z = pack2(__uint_as_float(a[8 * j + 4]) * s, __uint_as_float(a[8 * j + 5]) * s);
    u.w = pack2(__uint_as_float(a[8 * j + 6]) * s, __uint_as_float(a[8 * j + 7]) * s);
    d[j] = u;
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint4 u;
    u.x = pack2(__uint_as_float(b[8 * j]) * s, __uint_as_float(b[8 * j + 1]) * s);
    u.y = pack2(__uint_as_float(b[8 * j + 2]) * s, __uint_as_float(b[8 * j + 3]) * s);
    u.z = pack2(__uint_as_float(b[8 * j + 4]) * s, __uint_as_float(b[8 * j + 5]) * s);
    u.w = pack2(__uint_as_float(b[8 * j + 6]) * s, __uint_as_float(b[8 * j + 7]) * s);
    d[4 + j] = u;
  }
}
__device__ __forceinline__ void store_row32(__nv_bfloat16* dst, const uint32_t (&a)[32], float s) {
  uint4* d = reinterpret_cast<uint4*>(dst);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint4 u;
    u.x = pack2(__uint_as_float(a[8 * j]) * s, __uint_as_float(a[8 * j + 1]) * s);
    u.y = pack2(__uint_as_float(a[8 * j + 2]) * s, __uint_as_float(a[8 * j + 3]) * s);
    u.z = pack2(__uint_as_float(a[8 * j + 4]) * s, __uint_as_float(a[8 * j + 5]) * s);
    u.w = pack2(__uint_as_float(a[8 * j + 6]) * s, __uint_as_float(a[8 * j + 7]) * s);
    d[j] = u;
  }
}

// One 16*NW-column chunk of the recomputation.  kKeysOnLanes = false: this thread's lane is query `row`,
// columns are keys; true: lane is key `row`, columns are queries (per-column lse/delta from shared memory).
// Writes dS (and P when kKeysOnLanes) as packed bf16 over the first half of the consumed fp32 columns.
template <int NC, bool kKeysOnLanes, bool kCausal, bool kMaskCols>
__device__ __forceinline__ void bwd_chunk(uint32_t t_s, uint32_t t_dp, uint32_t t_ps, uint32_t t_ds, int c0,
                                          int row, int L, float scale_log2, float lse2_row, float delta_row,
                                          const float* lse2_s, const float* delta_s) {
  uint32_t s[NC], dp[NC], pk_p[NC / 2], pk_ds[NC / 2];
  if (NC == 32) {
    tmem_ld_32x32(t_s + c0, reinterpret_cast<uint32_t(&)[32]>(s));
    tmem_ld_32x32(t_dp + c0, reinterpret_cast<uint32_t(&)[32]>(dp));
  } else {
    tmem_ld_32x16(t_s + c0, reinterpret_cast<uint32_t(&)[16]>(s));
    tmem_ld_32x16(t_dp + c0, reinterpret_cast<uint32_t(&)[16]>(dp));
  }
  // per-column statistics (keys on lanes) as 16-byte shared-memory loads, issued under the TMEM load latency
  float l2c[kKeysOnLanes ? NC : 1], dlc[kKeysOnLanes ? NC : 1];
  if (kKeysOnLanes) {
#pragma unroll
    for (int i = 0; i < NC; i += 4) {
      const float4 a = *reinterpret_cast<const float4*>(lse2_s + c0 + i);
      const float4 b = *reinterpret_cast<const float4*>(delta_s + c0 + i);
      l2c[i] = a.x; l2c[i + 1] = a.y; l2c[i + 2] = a.z; l2c[i + 3] = a.w;
      dlc[i] = b.x; dlc[i + 1] = b.y; dlc[i + 2] = b.z; dlc[i + 3] = b.w;
    }
  }
  tmem_ld_wait();
  float pv[2], dv[2];
#pragma unroll
  for (int i = 0; i < NC; ++i) {
    const int col = c0 + i;
    const float l2 = kKeysOnLanes ? l2c[kKeysOnLanes ? i : 0] : lse2_row;
    const float dl = kKeysOnLanes ? dlc[kKeysOnLanes ? i : 0] : delta_row;
    // non-causal: padded QUERIES carry lse2 = 1e30 (P = 0) and padded lanes are never stored, so only padded KEY
    // columns of the queries-on-lanes orientation need a mask, and only in the chunk that straddles L
    bool valid = true;
    if (kCausal) {
      valid = kKeysOnLanes ? (row < L && col < L && row <= col) : (col < L && col <= row);
    } else if (kMaskCols && !kKeysOnLanes) {
      valid = col < L;
    }
    float p = ex2(fmaf(__uint_as_float(s[i]), scale_log2, -l2));
    float d = p * (__uint_as_float(dp[i]) - dl);
    if (kCausal || (kMaskCols && !kKeysOnLanes)) {
      p = valid ? p : 0.f;
      d = valid ? d : 0.f;
    }
    pv[i & 1] = p;
    dv[i & 1] = d;
    if (i & 1) {
      pk_p[i >> 1] = pack2(pv[0], pv[1]);
      pk_ds[i >> 1] = pack2(dv[0], dv[1]);
    }
  }
  if (NC == 32) {
    if (kKeysOnLanes) tmem_st_32x16(t_ps + (c0 >> 1), reinterpret_cast<const uint32_t(&)[16]>(pk_p));
    tmem_st_32x16(t_ds + (c0 >> 1), reinterpret_cast<const uint32_t(&)[16]>(pk_ds));
  } else {
    if (kKeysOnLanes) tmem_st_32x8(t_ps + (c0 >> 1), pk_p);
    tmem_st_32x8(t_ds + (c0 >> 1), pk_ds);
  }
}

template <bool kCausal>
__global__ void __launch_bounds__(kBwdThreads)
attn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_do,
                   const __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ dout,
                   const float* __restrict__ lse, __nv_bfloat16* __restrict__ dqkv, int L, int H, float scale,
                   float scale_log2) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int n_mt = (L + 127) >> 7;
  const int Lp = (L + 15) & ~15;
  const int sp = ((Lp >> 1) + 15) & ~15;  // column split between the two warps of a lane quarter
  uint8_t* Qs = smem;
  uint8_t* Ks = Qs + n_mt * kTile;
  uint8_t* Vs = Ks + n_mt * kTile;
  uint8_t* dOs = Vs + n_mt * kTile;
  float* lse2_s = reinterpret_cast<float*>(dOs + n_mt * kTile);
  float* delta_s = lse2_s + 256;
  uint64_t* bars = reinterpret_cast<uint64_t*>(delta_s + 256);
  uint64_t *bar_load = bars, *bar_s = bars + 1, *bar_p = bars + 2, *bar_o = bars + 3, *bar_e = bars + 4;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bars + 5);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.x % H, seq = blockIdx.x / H;   // linear grid: no 65 535 limit on sequences
  const int D = H * 64;
  const long long ld3 = 3LL * D;

  if (warp == 8) {
    if (lane == 0) {
      tma_prefetch_desc(&tmap_qkv);
      tma_prefetch_desc(&tmap_do);
      mbar_init(bar_load, 1);
      mbar_init(bar_s, 1);
      mbar_init(bar_p, 8);
      mbar_init(bar_o, 1);
      mbar_init(bar_e, 8);
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tslot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = *tslot;
  const int n_phase = 2 * n_mt;
  const uint32_t kColOutA = kColS + bwd_out_col(Lp, sp), kColOutB = kColDP + bwd_out_col(Lp, sp);

  if (warp == 8) {
    if (lane == 0) {
      mbar_expect_tx(bar_load, 4 * n_mt * kTile);
      for (int t = 0; t < n_mt; ++t) {
        tma_load_3d(&tmap_qkv, bar_load, Qs + t * kTile, h * 64, t * 128, seq);
        tma_load_3d(&tmap_qkv, bar_load, Ks + t * kTile, D + h * 64, t * 128, seq);
        tma_load_3d(&tmap_qkv, bar_load, Vs + t * kTile, 2 * D + h * 64, t * 128, seq);
        tma_load_3d(&tmap_do, bar_load, dOs + t * kTile, h * 64, t * 128, seq);
      }
      wait_bounded(bar_load, 0);
      const uint32_t idesc_s = idesc_kk(Lp);
      const uint32_t qs = smem_u32(Qs), ks = smem_u32(Ks), vs = smem_u32(Vs), ds = smem_u32(dOs);
      const int ksteps = Lp >> 4, ksplit = sp >> 4;
      for (int ph = 0; ph < n_phase; ++ph) {
        const bool qlanes = ph < n_mt;
        const int t = qlanes ? ph : ph - n_mt;
        if (ph > 0) {
          wait_bounded(bar_e, (ph - 1) & 1);
          tc_fence_after();
        }
        // scores and dP for this 128-lane tile
        const uint32_t a_s = (qlanes ? qs : ks) + t * kTile, b_s = qlanes ? ks : qs;
        const uint32_t a_dp = (qlanes ? ds : vs) + t * kTile, b_dp = qlanes ? vs : ds;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(tb + kColS, desc_kmajor(a_s + k * 32), desc_kmajor(b_s + k * 32), idesc_s, k > 0);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(tb + kColDP, desc_kmajor(a_dp + k * 32), desc_kmajor(b_dp + k * 32), idesc_s, k > 0);
        umma_commit(bar_s);
        wait_bounded(bar_p, ph & 1);
        tc_fence_after();
        // packed bf16 operands: k-steps [0, ksplit) at column 8k, k-steps [ksplit, ksteps) at sp + 8(k - ksplit)
        for (int k = 0; k < ksteps; ++k) {
          const uint32_t acol = k < ksplit ? 8 * k : sp + 8 * (k - ksplit);
          if (qlanes) {
            umma_bf16_ts(tb + kColOutA, tb + kColS + acol, desc_mnmajor(ks + k * 2048), kIdescMn64, k > 0);   // dQ = dS K
          } else {
            umma_bf16_ts(tb + kColOutA, tb + kColS + acol, desc_mnmajor(ds + k * 2048), kIdescMn64, k > 0);   // dV = P^T dO
            umma_bf16_ts(tb + kColOutB, tb + kColDP + acol, desc_mnmajor(qs + k * 2048), kIdescMn64, k > 0);  // dK = dS^T Q
          }
        }
        umma_commit(bar_o);
      }
    }
  } else {
    const int q = warp & 3, hh = warp >> 2;
    const int r = q * 32 + lane;
    const uint32_t tl = tb + ((uint32_t)(q * 32) << 16);
    // ---- prologue: delta_i = <dO_i, O_i>, lse * log2(e)
    {
      const int i = warp * 32 + lane;  // 0..255
      float acc = 0.f, l2 = 1e30f;     // padding rows: exp2(s - 1e30) = 0
      if (i < L) {
        const uint4* po = reinterpret_cast<const uint4*>(out + ((long long)seq * L + i) * D + h * 64);
        const uint4* pd = reinterpret_cast<const uint4*>(dout + ((long long)seq * L + i) * D + h * 64);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const uint4 a = po[c], b = pd[c];
          const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&a);
          const __nv_bfloat162* hb = reinterpret_cast<const __nv_bfloat162*>(&b);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 fa = __bfloat1622float2(ha[j]), fb = __bfloat1622float2(hb[j]);
            acc = fmaf(fa.x, fb.x, acc);
            acc = fmaf(fa.y, fb.y, acc);
          }
        }
        l2 = lse[((long long)seq * H + h) * L + i] * kLog2e;
      }
      delta_s[i] = acc;
      lse2_s[i] = l2;
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");  // the eight elementwise warps only

    for (int ph = 0; ph < n_phase; ++ph) {
      const bool qlanes = ph < n_mt;
      const int t = qlanes ? ph : ph - n_mt;
      const int row = t * 128 + r;
      wait_bounded(bar_s, ph & 1);
      tc_fence_after();
      const int c_lo = hh ? sp : 0, c_hi = hh ? Lp : sp;
      const uint32_t t_s = tl + kColS, t_dp = tl + kColDP;
      // packed outputs start at the beginning of this warp's own column range
      const uint32_t t_ps = tl + kColS + c_lo - (c_lo >> 1), t_ds_q = t_ps, t_ds_k = tl + kColDP + c_lo - (c_lo >> 1);
      const float l2r = lse2_s[row < 256 ? row : 0], dlr = delta_s[row < 256 ? row : 0];
      for (int c0 = c_lo; c0 < c_hi; c0 += 32) {
        const bool wide = c0 + 32 <= c_hi;
        const bool straddles = c0 + (wide ? 32 : 16) > L;   // chunk contains padded key columns
        if (qlanes) {
          if (wide) {
            if (straddles) bwd_chunk<32, false, kCausal, true>(t_s, t_dp, t_ps, t_ds_q, c0, row, L, scale_log2, l2r, dlr, lse2_s, delta_s);
            else bwd_chunk<32, false, kCausal, false>(t_s, t_dp, t_ps, t_ds_q, c0, row, L, scale_log2, l2r, dlr, lse2_s, delta_s);
          } else {
            if (straddles) bwd_chunk<16, false, kCausal, true>(t_s, t_dp, t_ps, t_ds_q, c0, row, L, scale_log2, l2r, dlr, lse2_s, delta_s);
            else bwd_chunk<16, false, kCausal, false>(t_s, t_dp, t_ps, t_ds_q, c0, row, L, scale_log2, l2r, dlr, lse2_s, delta_s);
          }
        } else {
          if (wide)
            bwd_chunk<32, true, kCausal, false>(t_s, t_dp, t_ps, t_ds_k, c0, row, L, scale_log2, l2r, dlr, lse2_s, delta_s);
          else
            bwd_chunk<16, true, kCausal, false>(t_s, t_dp, t_ps, t_ds_k, c0, row, L, scale_log2, l2r, dlr, lse2_s, delta_s);
        }
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p);
      // ---- epilogue
      wait_bounded(bar_o, ph & 1);
      tc_fence_after();
      __nv_bfloat16* drow = dqkv + ((long long)seq * L + row) * ld3 + h * 64;
      // the accumulators are copied to registers, the TMEM columns handed back to the MMA warp (next phase's
      // S / dP MMAs start), and only then do the global stores run
      if (qlanes) {  // dQ: each warp of the pair writes 32 of the 64 dims
        uint32_t a[32];
        tmem_ld_32x32(tl + kColOutA + hh * 32, a);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_e);
        if (row < L) store_row32(drow + hh * 32, a, scale);
      } else {       // hh = 0: dV, hh = 1: dK (scaled)
        uint32_t a[32], b[32];
        const uint32_t col = hh ? kColOutB : kColOutA;
        tmem_ld_32x32(tl + col, a);
        tmem_ld_32x32(tl + col + 32, b);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_e);
        if (row < L) store_row64(drow + (hh ? D : 2 * D), a, b, hh ? scale : 1.f);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 8) tmem_dealloc(tb, 512);
}

// ------------------------------------------------------------------------------------------- backward, pipelined
// Same mathematics and orientations as attn_bwd_tc_kernel, one CTA per (sequence, head), but the columns of every
// 128-lane phase are walked in chunks of <= 64 through TWO TMEM chunk buffers (S_c | dP_c, 128 columns each) and two
// output buffers (out A | out B), so that the MMA warp computes the scores of chunk i+1 and the output products of
// chunk i-1 while the eight elementwise warps -- which are bound by the 64 B/clk TMEM read port -- work on chunk i,
// and the epilogue of a phase is deferred until after the first chunk of the next one.  tcgen05.mma issue order
// keeps chunk i-1's output MMAs (A = packed dS / P inside the chunk buffer) ahead of chunk i+1's score MMAs that
// overwrite the same buffer.
constexpr uint32_t kP3Buf = 128, kP3DP = 64, kP3Out = 256, kP3OutB = 64;
constexpr int kBwd3Threads = 320;   // warps 0-7 elementwise + epilogue, warp 8: TMEM / TMA / MMA issuer A, warp 9: MMA issuer B
#ifdef PSRC_ATTN_TIMELINE   // dev instrumentation (tests/attn_timeline.cu): clock64 stamps of one (sequence, head)
__device__ long long g_tl[2][64];
#define TL(who, idx) do { if (blockIdx.x == gridDim.x / 2 + 3) g_tl[who][idx] = clock64(); } while (0)
#else
#define TL(who, idx) do { } while (0)
#endif

template <bool kCausal>
__global__ void __launch_bounds__(kBwd3Threads)
attn_bwd_tc3_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_do,
                    const __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ dout,
                    const float* __restrict__ lse, __nv_bfloat16* __restrict__ dqkv, int L, int H, float scale,
                    float scale_log2) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int n_mt = (L + 127) >> 7;
  const int n_phase = 2 * n_mt;
  const int Lp = (L + 15) & ~15;
  const int n_grp = Lp >> 4;                      // 16-column groups
  const int n_ch = (n_grp + 3) >> 2;              // chunks of at most 4 groups, sizes as even as possible
  const int g_base = n_grp / n_ch, g_rem = n_grp % n_ch;
  const int n_step = n_phase * n_ch;
  uint8_t* Qs = smem;
  uint8_t* Ks = Qs + n_mt * kTile;
  uint8_t* Vs = Ks + n_mt * kTile;
  uint8_t* dOs = Vs + n_mt * kTile;
  float* lse2_s = reinterpret_cast<float*>(dOs + n_mt * kTile);
  float* delta_s = lse2_s + 256;
  uint64_t* bars = reinterpret_cast<uint64_t*>(delta_s + 256);
  uint64_t *bar_load = bars, *bar_s = bars + 1, *bar_p = bars + 3, *bar_o = bars + 5, *bar_e = bars + 7;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bars + 9);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.x % H, seq = blockIdx.x / H;   // linear grid: no 65 535 limit on sequences
  const int D = H * 64;
  const long long ld3 = 3LL * D;

  if (warp == 8) {
    if (lane == 0) {
      TL(0, 0);
      tma_prefetch_desc(&tmap_qkv);
      tma_prefetch_desc(&tmap_do);
      mbar_init(bar_load, 1);
      for (int i = 0; i < 2; ++i) {
        mbar_init(bar_s + i, 2);   // one tcgen05.commit from each of the two issuing threads
        mbar_init(bar_p + i, 8);
        mbar_init(bar_o + i, 2);
        mbar_init(bar_e + i, 8);
      }
      fence_mbar_init();
      mbar_expect_tx(bar_load, 4 * n_mt * kTile);   // the loads do not depend on the TMEM allocation
      for (int t = 0; t < n_mt; ++t) {
        tma_load_3d(&tmap_qkv, bar_load, Qs + t * kTile, h * 64, t * 128, seq);
        tma_load_3d(&tmap_qkv, bar_load, Ks + t * kTile, D + h * 64, t * 128, seq);
        tma_load_3d(&tmap_do, bar_load, dOs + t * kTile, h * 64, t * 128, seq);
        tma_load_3d(&tmap_qkv, bar_load, Vs + t * kTile, 2 * D + h * 64, t * 128, seq);
      }
    }
    __syncwarp();
    tmem_alloc(tslot, 512);
    tmem_relinquish();
  }
  // ---- prologue, part 1: the global loads behind delta_i = <dO_i, O_i> need neither TMEM nor the barriers -- they are
  //      issued before the CTA-wide sync and consumed after it
  uint4 pro_a[8], pro_b[8];
  float pro_l2 = 1e30f;              // padding rows: exp2(s - 1e30) = 0
  const int pro_i = warp * 32 + lane;
  if (warp < 8 && pro_i < L) {
    const uint4* po = reinterpret_cast<const uint4*>(out + ((long long)seq * L + pro_i) * D + h * 64);
    const uint4* pd = reinterpret_cast<const uint4*>(dout + ((long long)seq * L + pro_i) * D + h * 64);
#pragma unroll
    for (int c = 0; c < 8; ++c) { pro_a[c] = po[c]; pro_b[c] = pd[c]; }
    pro_l2 = lse[((long long)seq * H + h) * L + pro_i] * kLog2e;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = *tslot;

  if (warp >= 8) {
    // Two issuing threads (warp 8: everything that touches the S half of a chunk buffer -- score MMAs and the output
    // MMAs whose A operand is packed there (dQ / dV); warp 9: the dP half -- dP MMAs and dK).  Measured: one thread
    // issues a narrow tcgen05.mma every ~100-200 clk, far above the tensor-pipe floor (32 clk at N = 64), so a single
    // issuer was the bottleneck of the pipelined loop.  Each thread's own issue order covers its write-after-read
    // hazards (the next scores overwrite the columns its previous output MMAs read).
    if (lane == 0) {
      const bool second = warp == 9;
      if (!second) TL(0, 1);
      wait_bounded(bar_load, 0);
      if (!second) TL(0, 2);
      const uint32_t qs = smem_u32(Qs), ks = smem_u32(Ks), vs = smem_u32(Vs), ds = smem_u32(dOs);
      auto issue_sd = [&](int ph, int c, int b) {
        const bool ql = ph < n_mt;
        const int t = ql ? ph : ph - n_mt;
        const int ng = g_base + (c < g_rem ? 1 : 0);
        const uint32_t roff = (uint32_t)(c * g_base + min(c, g_rem)) * 2048;   // 16 rows x 128 B per group
        const uint32_t idesc = idesc_kk(ng * 16);
        const uint32_t a_op = (second ? (ql ? ds : vs) : (ql ? qs : ks)) + t * kTile;
        const uint32_t b_op = (second ? (ql ? vs : ds) : (ql ? ks : qs)) + roff;
        const uint32_t tS = tb + (uint32_t)b * kP3Buf + (second ? kP3DP : 0u);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(tS, desc_kmajor(a_op + k * 32), desc_kmajor(b_op + k * 32), idesc, k > 0);
        umma_commit(bar_s + b);
      };
      issue_sd(0, 0, 0);
      int ph = 0, c = 0;          // current step
      int ph_n = 0, c_n = 0;      // next step
      for (int i = 0; i < n_step; ++i) {
        if (++c_n == n_ch) { c_n = 0; ++ph_n; }
        if (i + 1 < n_step) issue_sd(ph_n, c_n, (i + 1) & 1);
        if (!second) TL(0, 22 + i);
        wait_bounded(bar_p + (i & 1), (i >> 1) & 1);
        tc_fence_after();
        if (!second) TL(0, 3 + i);
        if (c == 0 && ph >= 2) {   // output buffer ph & 1 was last read by the epilogue of phase ph - 2
          wait_bounded(bar_e + (ph & 1), ((ph - 2) >> 1) & 1);
          tc_fence_after();
        }
        const bool ql = ph < n_mt;
        const int ng = g_base + (c < g_rem ? 1 : 0);
        const uint32_t roff = (uint32_t)(c * g_base + min(c, g_rem)) * 2048;
        const uint32_t tS = tb + (uint32_t)(i & 1) * kP3Buf;
        const uint32_t tO = tb + kP3Out + (uint32_t)(ph & 1) * 128;
        if (!second) {
          const uint32_t bop = (ql ? ks : ds) + roff;            // dQ += dS K   |   dV += P^T dO
          for (int k = 0; k < ng; ++k)
            umma_bf16_ts(tO, tS + (uint32_t)((k >> 1) * 32 + (k & 1) * 8), desc_mnmajor(bop + k * 2048), kIdescMn64,
                         (c > 0 || k > 0) ? 1u : 0u);
        } else if (!ql) {
          for (int k = 0; k < ng; ++k)                             // dK += dS^T Q
            umma_bf16_ts(tO + kP3OutB, tS + kP3DP + (uint32_t)((k >> 1) * 32 + (k & 1) * 8),
                         desc_mnmajor(qs + roff + k * 2048), kIdescMn64, (c > 0 || k > 0) ? 1u : 0u);
        }
        if (c == n_ch - 1) umma_commit(bar_o + (ph & 1));
        if (!second) TL(0, 40 + i);
        if (++c == n_ch) { c = 0; ++ph; }
      }
      if (!second) TL(0, 20);
    }
  } else {
    const int q = warp & 3, hh = warp >> 2;
    const int r = q * 32 + lane;
    const uint32_t tl = tb + ((uint32_t)(q * 32) << 16);
    {  // prologue, part 2
      float acc = 0.f;
      if (pro_i < L) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&pro_a[c]);
          const __nv_bfloat162* hb = reinterpret_cast<const __nv_bfloat162*>(&pro_b[c]);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 fa = __bfloat1622float2(ha[j]), fb = __bfloat1622float2(hb[j]);
            acc = fmaf(fa.x, fb.x, acc);
            acc = fmaf(fa.y, fb.y, acc);
          }
        }
      }
      delta_s[pro_i] = acc;
      lse2_s[pro_i] = pro_l2;
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");  // the eight elementwise warps only
    if (threadIdx.x == 0) TL(1, 0);

    // accumulators of phase p -> registers -> global; the TMEM buffer is handed back before the stores
    auto epilogue = [&](int p) {
      const bool ql = p < n_mt;
      const int row = (ql ? p : p - n_mt) * 128 + r;
      wait_bounded(bar_o + (p & 1), (p >> 1) & 1);
      tc_fence_after();
      const uint32_t tO = tl + kP3Out + (uint32_t)(p & 1) * 128;
      __nv_bfloat16* drow = dqkv + ((long long)seq * L + row) * ld3 + h * 64;
      if (ql) {  // dQ: each warp of the pair writes 32 of the 64 dims
        uint32_t a[32];
        tmem_ld_32x32(tO + hh * 32, a);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_e + (p & 1));
        if (row < L) store_row32(drow + hh * 32, a, scale);
      } else {   // hh = 0: dV, hh = 1: dK (scaled)
        uint32_t a[32], b[32];
        const uint32_t col = hh ? kP3OutB : 0;
        tmem_ld_32x32(tO + col, a);
        tmem_ld_32x32(tO + col + 32, b);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_e + (p & 1));
        if (row < L) store_row64(drow + (hh ? D : 2 * D), a, b, hh ? scale : 1.f);
      }
      if (threadIdx.x == 0) TL(1, 40 + p);
    };

    int ph = 0, c = 0;
    for (int i = 0; i < n_step; ++i) {
      const bool ql = ph < n_mt;
      const int row = (ql ? ph : ph - n_mt) * 128 + r;
      const int ng = g_base + (c < g_rem ? 1 : 0);
      const int wc = ng * 16;
      const int loc = hh * 32;                       // this warp's first column inside the chunk
      const float l2r = lse2_s[row < 256 ? row : 0], dlr = delta_s[row < 256 ? row : 0];
      wait_bounded(bar_s + (i & 1), (i >> 1) & 1);
      tc_fence_after();
      if (threadIdx.x == 0) TL(1, 1 + 2 * i);
      if (loc < wc) {
        const int c0 = (c * g_base + min(c, g_rem)) * 16 + loc;   // logical column (key or query index)
        const bool wide = loc + 32 <= wc;
        const bool straddles = c0 + (wide ? 32 : 16) > L;
        // bwd_chunk addresses TMEM as base + c0 (fp32) and base + c0/2 (packed): bias the bases to buffer-local columns
        const uint32_t tbuf = tl + (uint32_t)(i & 1) * kP3Buf + (uint32_t)loc;
        const uint32_t t_s = tbuf - (uint32_t)c0, t_dp = tbuf + kP3DP - (uint32_t)c0;
        const uint32_t t_ps = tbuf - (uint32_t)(c0 >> 1), t_dk = tbuf + kP3DP - (uint32_t)(c0 >> 1);
        if (ql) {
          if (wide) {
            if (straddles) bwd_chunk<32, false, kCausal, true>(t_s, t_dp, t_ps, t_ps, c0, row, L, scale_log2, l2r, dlr, lse2_s, delta_s);
            else bwd_chunk<32, false, kCausal, false>(t_s, t_dp, t_ps, t_ps, c0, row, L, scale_log2, l2r, dlr, lse2_s, delta_s);
          } else {
            if (straddles) bwd_chunk<16, false, kCausal, true>(t_s, t_dp, t_ps, t_ps, c0, row, L, scale_log2, l2r, dlr, lse2_s, delta_s);
            else bwd_chunk<16, false, kCausal, false>(t_s, t_dp, t_ps, t_ps, c0, row, L, scale_log2, l2r, dlr, lse2_s, delta_s);
          }
        } else {
          if (wide)
            bwd_chunk<32, true, kCausal, false>(t_s, t_dp, t_ps, t_dk, c0, row, L, scale_log2, l2r, dlr, lse2_s, delta_s);
          else
            bwd_chunk<16, true, kCausal, false>(t_s, t_dp, t_ps, t_dk, c0, row, L, scale_log2, l2r, dlr, lse2_s, delta_s);
        }
        tmem_st_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p + (i & 1));
      if (threadIdx.x == 0) TL(1, 2 + 2 * i);
      if (c == 0 && ph > 0) epilogue(ph - 1);       // deferred: the next phase's first chunk is already with the MMA warp
      if (++c == n_ch) { c = 0; ++ph; }
    }
    epilogue(n_phase - 1);
    if (threadIdx.x == 0) TL(1, 50);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 8) tmem_dealloc(tb, 512);
}

// ---------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}
// [n_seq][L][width] bf16 (row stride = width elements) -> boxes of 128 rows x 64 columns of one sequence
static int make_map3(CUtensorMap* map, const void* base, long long n_seq, long long L, long long width) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return PSRC_ERR_CUDA;
  }
  cuuint64_t gdim[3] = {(cuuint64_t)width, (cuuint64_t)L, (cuuint64_t)n_seq};
  cuuint64_t gstride[2] = {(cuuint64_t)width * 2, (cuuint64_t)L * width * 2};
  cuuint32_t box[3] = {64, 128, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (attention) failed: CUresult %d (n_seq=%lld L=%lld width=%lld)", (int)r, n_seq,
              L, width);
    return PSRC_ERR_CUDA;
  }
  return PSRC_OK;
}

template <typename K>
static int set_smem(K kernel, size_t bytes) {
  PSRC_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return PSRC_OK;
}

}  // namespace atc

bool attn_tc_supported(int L) { return L > 32 && L <= 256; }

int attn_fwd_tc(const void* qkv, void* out, float* lse, int n_seq, int L, int H, int causal, cudaStream_t s) {
  using namespace atc;
  const int n_mt = (L + 127) >> 7;
  const size_t smem = (size_t)3 * n_mt * kTile + 64 + 1024;
  CUtensorMap map;
  int rc = make_map3(&map, qkv, n_seq, L, 3LL * H * 64);
  if (rc) return rc;
  dim3 grid((unsigned)H * (unsigned)n_seq);
  const float scale_log2 = 0.125f * kLog2e;
  __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
  if (causal) {
    if ((rc = set_smem(attn_fwd_tc_kernel<true>, smem))) return rc;
    attn_fwd_tc_kernel<true><<<grid, kFwdThreads, smem, s>>>(map, o, lse, L, H, scale_log2);
  } else {
    if ((rc = set_smem(attn_fwd_tc_kernel<false>, smem))) return rc;
    attn_fwd_tc_kernel<false><<<grid, kFwdThreads, smem, s>>>(map, o, lse, L, H, scale_log2);
  }
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

int attn_bwd_tc(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, int n_seq, int L,
                int H, int causal, cudaStream_t s) {
  using namespace atc;
  const int n_mt = (L + 127) >> 7;
  CUtensorMap map_qkv, map_do;
  int rc = make_map3(&map_qkv, qkv, n_seq, L, 3LL * H * 64);
  if (rc) return rc;
  if ((rc = make_map3(&map_do, dout, n_seq, L, (long long)H * 64))) return rc;
  const float scale = 0.125f, scale_log2 = 0.125f * kLog2e;
  const __nv_bfloat16* o = reinterpret_cast<const __nv_bfloat16*>(out);
  const __nv_bfloat16* d = reinterpret_cast<const __nv_bfloat16*>(dout);
  __nv_bfloat16* dq = reinterpret_cast<__nv_bfloat16*>(dqkv);
  // PSRC_ATTN_BWD = 1: serial kernel; 3: chunk-pipelined kernel; unset: by shape
  static const int version = getenv("PSRC_ATTN_BWD") ? atoi(getenv("PSRC_ATTN_BWD")) : 0;
  // measured on B200 (tests/attn_bench.py): L = 201: 249 vs 255 us, L = 77 causal: 478 vs 467 us -> pipelined kernel
  // when there are two lane tiles (16 steps to pipeline), serial kernel otherwise
  if (version == 3 || (version == 0 && n_mt >= 2)) {
    const size_t smem3 = (size_t)4 * n_mt * kTile + 2 * 256 * sizeof(float) + 128 + 1024;
    dim3 grid3((unsigned)H * (unsigned)n_seq);
    if (causal) {
      if ((rc = set_smem(attn_bwd_tc3_kernel<true>, smem3))) return rc;
      attn_bwd_tc3_kernel<true><<<grid3, kBwd3Threads, smem3, s>>>(map_qkv, map_do, o, d, lse, dq, L, H, scale, scale_log2);
    } else {
      if ((rc = set_smem(attn_bwd_tc3_kernel<false>, smem3))) return rc;
      attn_bwd_tc3_kernel<false><<<grid3, kBwd3Threads, smem3, s>>>(map_qkv, map_do, o, d, lse, dq, L, H, scale, scale_log2);
    }
    PSRC_CHECK_LAUNCH();
    return PSRC_OK;
  }
  const size_t smem = (size_t)4 * n_mt * kTile + 2 * 256 * sizeof(float) + 64 + 1024;
  dim3 grid((unsigned)H * (unsigned)n_seq);
  if (causal) {
    if ((rc = set_smem(attn_bwd_tc_kernel<true>, smem))) return rc;
    attn_bwd_tc_kernel<true><<<grid, kBwdThreads, smem, s>>>(map_qkv, map_do, o, d, lse, dq, L, H, scale, scale_log2);
  } else {
    if ((rc = set_smem(attn_bwd_tc_kernel<false>, smem))) return rc;
    attn_bwd_tc_kernel<false><<<grid, kBwdThreads, smem, s>>>(map_qkv, map_do, o, d, lse, dq, L, H, scale, scale_log2);
  }
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

}  // namespace psrc
