// Tensor-core attention for very short sequences (L <= 16, head_dim 64, bf16): the truncated text tower
// (SURVEY F12).  One warp owns one (sequence, head) problem padded to a single 16 x 16 score tile:
//   forward : S = Q K^T (8 mma.m16n8k16), softmax in the accumulator fragments, O = P V (8 mma)
//   backward: S, dP = dO V^T, dS = P o (dP - delta) with delta_i = sum_j P_ij dP_ij (= <dO_i, O_i>, so O is never
//             read), dQ = dS K, dK = dS^T Q, dV = P^T dO -- 40 mma; P / dS cross to the transposed operand through a
//             512-byte shared tile and ldmatrix.trans.
// Q/K/V/dO tiles are staged with coalesced 16-byte loads into padded shared memory (row stride 144 B: conflict-free
// ldmatrix), results leave through the same tiles with 16-byte stores.  The kernel is HBM-bound: 8 KB (fwd 6 KB) in
// flight per warp.  Reference: nn.MultiheadAttention core called from PromptSRC/clip/model.py:183/227/285 with the
// causal mask of model.py:592-598.
#include "common.cuh"

namespace psrc {
namespace asm2 {

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr int kLd = 72;    // tile row stride in elements (64 + 8 pad)
constexpr int kXLd = 24;   // P / dS tile row stride (16 + 8 pad)
constexpr int kWarps = 4;

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// F16 = true: fp16 forward operands (q, k, v, P, O in IEEE half; 3 more mantissa bits than bf16 at the same rate)
template <bool F16 = false>
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  if constexpr (F16) {
    __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  } else {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  }
}
__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], const __nv_bfloat16* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_u32(p)));
}
__device__ __forceinline__ void ldsm4t(uint32_t (&r)[4], const __nv_bfloat16* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_u32(p)));
}
template <bool F16 = false>
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  if constexpr (F16) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  } else {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
}
// a staged [rows][kLd] tile of fp16 values -> bf16 in place (each lane converts the 16-byte chunks it staged).  The
// backward's gradient contractions pair q / k / v with bf16 gradients; mma.sync wants one 16-bit format per instruction
template <int ROWS>
__device__ __forceinline__ void tile_f16_to_bf16(__nv_bfloat16* tile, int lane) {
#pragma unroll
  for (int i = 0; i < ROWS / 4; ++i) {
    const int idx = lane + 32 * i, r = idx >> 3, c = idx & 7;
    uint4 u = *reinterpret_cast<uint4*>(tile + r * 72 + c * 8);
    uint32_t* w = reinterpret_cast<uint32_t*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = __half22float2(*reinterpret_cast<__half2*>(&w[j]));
      __nv_bfloat162 b = __floats2bfloat162_rn(f.x, f.y);
      w[j] = *reinterpret_cast<uint32_t*>(&b);
    }
    *reinterpret_cast<uint4*>(tile + r * 72 + c * 8) = u;
  }
}

// stage L rows of 64 bf16 (row stride `ld` elements in global) into a [16][kLd] tile, zero-filling rows >= L
__device__ __forceinline__ void stage_tile(__nv_bfloat16* tile, const __nv_bfloat16* __restrict__ src, long long ld,
                                           int L, int lane) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int idx = lane + 32 * i, r = idx >> 3, c = idx & 7;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (r < L) v = __ldg(reinterpret_cast<const uint4*>(src + (long long)r * ld) + c);
    *reinterpret_cast<uint4*>(tile + r * kLd + c * 8) = v;
  }
}
// the inverse: rows < L of a [16][kLd] tile to global with 16-byte stores
__device__ __forceinline__ void unstage_tile(const __nv_bfloat16* tile, __nv_bfloat16* __restrict__ dst, long long ld,
                                             int L, int lane) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int idx = lane + 32 * i, r = idx >> 3, c = idx & 7;
    if (r < L) *(reinterpret_cast<uint4*>(dst + (long long)r * ld) + c) =
        *reinterpret_cast<const uint4*>(tile + r * kLd + c * 8);
  }
}
// accumulator fragments of a [16 x 64] result (8 n-tiles) -> bf16 tile, scaled
template <bool F16 = false>
__device__ __forceinline__ void frags_to_tile(__nv_bfloat16* tile, const float (&c)[8][4], float s0, float s1,
                                              int g, int t) {
#pragma unroll
  for (int n = 0; n < 8; ++n) {
    *reinterpret_cast<uint32_t*>(tile + g * kLd + n * 8 + 2 * t) = pack_bf16<F16>(c[n][0] * s0, c[n][1] * s0);
    *reinterpret_cast<uint32_t*>(tile + (g + 8) * kLd + n * 8 + 2 * t) = pack_bf16<F16>(c[n][2] * s1, c[n][3] * s1);
  }
}

// C[16 x 16] = A[16 x 64] . B[16 x 64]^T, both tiles row-major [row][64]  (Q K^T, dO V^T)
template <bool F16 = false>
__device__ __forceinline__ void mma_abt(float (&c)[2][4], const __nv_bfloat16* A, const __nv_bfloat16* B, int lane) {
  const int arow = (lane & 7) + ((lane >> 3) & 1) * 8, acol = (lane >> 4) * 8;
  const int brow = lane & 7, bcol = (lane >> 3) * 8;
#pragma unroll
  for (int n = 0; n < 2; ++n)
#pragma unroll
    for (int i = 0; i < 4; ++i) c[n][i] = 0.f;
#pragma unroll
  for (int kk = 0; kk < 4; kk += 2) {
    uint32_t a0[4], a1[4], b[2][4];
    ldsm4(a0, A + arow * kLd + kk * 16 + acol);
    ldsm4(a1, A + arow * kLd + (kk + 1) * 16 + acol);
    ldsm4(b[0], B + brow * kLd + kk * 16 + bcol);          // n-tile 0, k-steps kk, kk+1
    ldsm4(b[1], B + (8 + brow) * kLd + kk * 16 + bcol);    // n-tile 1
#pragma unroll
    for (int n = 0; n < 2; ++n) {
      mma16816<F16>(c[n], a0, b[n][0], b[n][1]);
      mma16816<F16>(c[n], a1, b[n][2], b[n][3]);
    }
  }
}
// C[16 x 64] = A[16 x 16] . B[16 x 64], A given as a register fragment, B tile row-major [k][64]  (P V, dS K, ...)
template <bool F16 = false>
__device__ __forceinline__ void mma_ab(float (&c)[8][4], const uint32_t (&a)[4], const __nv_bfloat16* B, int lane) {
  const int brow = (lane & 7) + ((lane >> 3) & 1) * 8, bcol = (lane >> 4) * 8;
#pragma unroll
  for (int n = 0; n < 8; n += 2) {
    uint32_t b[4];
    ldsm4t(b, B + brow * kLd + n * 8 + bcol);
#pragma unroll
    for (int i = 0; i < 4; ++i) c[n][i] = c[n + 1][i] = 0.f;
    mma16816<F16>(c[n], a, b[0], b[1]);
    mma16816<F16>(c[n + 1], a, b[2], b[3]);
  }
}

template <bool kCausal, bool kF16>
__global__ void __launch_bounds__(kWarps * 32)
attn_small_mma_fwd_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out,
                          float* __restrict__ lse, int n_prob, int L, int H) {
  __shared__ __align__(16) __nv_bfloat16 sm[kWarps][3][16 * kLd];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pr = blockIdx.x * kWarps + warp;
  if (pr >= n_prob) return;
  const int seq = pr / H, h = pr % H, D = H * 64;
  const long long ld = 3LL * D;
  const __nv_bfloat16* base = qkv + (long long)seq * L * ld + h * 64;
  __nv_bfloat16 *Qs = sm[warp][0], *Ks = sm[warp][1], *Vs = sm[warp][2];
  stage_tile(Qs, base, ld, L, lane);
  stage_tile(Ks, base + D, ld, L, lane);
  stage_tile(Vs, base + 2 * D, ld, L, lane);
  __syncwarp();
  const int g = lane >> 2, t = lane & 3;
  float s[2][4];
  mma_abt<kF16>(s, Qs, Ks, lane);
  // scores -> log2 domain, mask (key >= L, causal key > query)
  float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
  for (int n = 0; n < 2; ++n)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int row = g + (i >> 1) * 8, col = n * 8 + 2 * t + (i & 1);
      const bool ok = col < L && (!kCausal || col <= row);
      s[n][i] = ok ? s[n][i] * (0.125f * kLog2e) : -INFINITY;
      if (i < 2) m0 = fmaxf(m0, s[n][i]); else m1 = fmaxf(m1, s[n][i]);
    }
  m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1));
  m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
  m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1));
  m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
  float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
  for (int n = 0; n < 2; ++n)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      s[n][i] = ex2(s[n][i] - ((i < 2) ? m0 : m1));      // exp2(-inf) = 0 for masked entries
      if (i < 2) sum0 += s[n][i]; else sum1 += s[n][i];
    }
  sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1);
  sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
  sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1);
  sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
  uint32_t p[4] = {pack_bf16<kF16>(s[0][0], s[0][1]), pack_bf16<kF16>(s[0][2], s[0][3]), pack_bf16<kF16>(s[1][0], s[1][1]),
                   pack_bf16<kF16>(s[1][2], s[1][3])};
  float o[8][4];
  mma_ab<kF16>(o, p, Vs, lane);
  __syncwarp();                       // every lane is done reading Qs (mma_abt) before it is reused for O
  frags_to_tile<kF16>(Qs, o, 1.f / sum0, 1.f / sum1, g, t);
  __syncwarp();
  unstage_tile(Qs, out + (long long)seq * L * D + h * 64, (long long)D, L, lane);
  if (t == 0) {
    float* l = lse + ((long long)seq * H + h) * L;
    if (g < L) l[g] = (m0 + log2f(sum0)) * kLn2;
    if (g + 8 < L) l[g + 8] = (m1 + log2f(sum1)) * kLn2;
  }
}

struct BwdSmem {
  __nv_bfloat16 q[16 * kLd], k[16 * kLd], v[16 * kLd], d[16 * kLd];
  __nv_bfloat16 p[16 * kXLd], ds[16 * kXLd];
};

template <bool kCausal, bool kF16>
__global__ void __launch_bounds__(kWarps * 32)
attn_small_mma_bwd_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ dout,
                          const float* __restrict__ lse, __nv_bfloat16* __restrict__ dqkv, int n_prob, int L, int H) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pr = blockIdx.x * kWarps + warp;
  if (pr >= n_prob) return;
  BwdSmem& S = reinterpret_cast<BwdSmem*>(smem_raw)[warp];
  const int seq = pr / H, h = pr % H, D = H * 64;
  const long long ld = 3LL * D;
  const __nv_bfloat16* base = qkv + (long long)seq * L * ld + h * 64;
  stage_tile(S.q, base, ld, L, lane);
  stage_tile(S.k, base + D, ld, L, lane);
  stage_tile(S.v, base + 2 * D, ld, L, lane);
  stage_tile(S.d, dout + (long long)seq * L * D + h * 64, (long long)D, L, lane);
  const int g = lane >> 2, t = lane & 3;
  const float* lrow = lse + ((long long)seq * H + h) * L;
  const float l0 = (g < L) ? lrow[g] * kLog2e : 0.f, l1 = (g + 8 < L) ? lrow[g + 8] * kLog2e : 0.f;
  __syncwarp();
  float s[2][4], dp[2][4];
  mma_abt<kF16>(s, S.q, S.k, lane);         // the scores exactly as the forward computed them
  if (kF16) {                                // gradient contractions run in bf16 (dO, dS are bf16)
    __syncwarp();
    tile_f16_to_bf16<16>(S.q, lane);
    tile_f16_to_bf16<16>(S.k, lane);
    tile_f16_to_bf16<16>(S.v, lane);
    __syncwarp();
  }
  mma_abt(dp, S.d, S.v, lane);
  float del0 = 0.f, del1 = 0.f;
#pragma unroll
  for (int n = 0; n < 2; ++n)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int row = g + (i >> 1) * 8, col = n * 8 + 2 * t + (i & 1);
      const bool ok = row < L && col < L && (!kCausal || col <= row);
      s[n][i] = ok ? ex2(s[n][i] * (0.125f * kLog2e) - ((i < 2) ? l0 : l1)) : 0.f;   // P_ij
      if (i < 2) del0 = fmaf(s[n][i], dp[n][i], del0); else del1 = fmaf(s[n][i], dp[n][i], del1);
    }
  del0 += __shfl_xor_sync(0xffffffffu, del0, 1);
  del0 += __shfl_xor_sync(0xffffffffu, del0, 2);
  del1 += __shfl_xor_sync(0xffffffffu, del1, 1);
  del1 += __shfl_xor_sync(0xffffffffu, del1, 2);
  uint32_t dsf[4];
#pragma unroll
  for (int n = 0; n < 2; ++n) {
    const float d0 = s[n][0] * (dp[n][0] - del0), d1 = s[n][1] * (dp[n][1] - del0);
    const float d2 = s[n][2] * (dp[n][2] - del1), d3 = s[n][3] * (dp[n][3] - del1);
    dsf[2 * n] = pack_bf16(d0, d1);
    dsf[2 * n + 1] = pack_bf16(d2, d3);
    *reinterpret_cast<uint32_t*>(S.ds + g * kXLd + n * 8 + 2 * t) = dsf[2 * n];
    *reinterpret_cast<uint32_t*>(S.ds + (g + 8) * kXLd + n * 8 + 2 * t) = dsf[2 * n + 1];
    *reinterpret_cast<uint32_t*>(S.p + g * kXLd + n * 8 + 2 * t) = pack_bf16(s[n][0], s[n][1]);
    *reinterpret_cast<uint32_t*>(S.p + (g + 8) * kXLd + n * 8 + 2 * t) = pack_bf16(s[n][2], s[n][3]);
  }
  __syncwarp();
  // transposed operand fragments: A = X^T with X[q][k] row-major
  uint32_t dst_[4], pt_[4];
  {
    const int xr = (lane & 7) + (lane >> 4) * 8, xc = ((lane >> 3) & 1) * 8;
    ldsm4t(dst_, S.ds + xr * kXLd + xc);
    ldsm4t(pt_, S.p + xr * kXLd + xc);
  }
  float acc[8][4];
  __nv_bfloat16* drow = dqkv + (long long)seq * L * ld + h * 64;
  // dV = P^T dO  (reads S.d), staged out through S.v (V is no longer needed)
  mma_ab(acc, pt_, S.d, lane);
  frags_to_tile(S.v, acc, 1.f, 1.f, g, t);
  // dQ = 0.125 dS K  (reads S.k), staged out through S.d (every lane has finished reading S.d only after a sync)
  mma_ab(acc, dsf, S.k, lane);
  __syncwarp();
  frags_to_tile(S.d, acc, 0.125f, 0.125f, g, t);
  // dK = 0.125 dS^T Q  (reads S.q), staged out through S.k
  mma_ab(acc, dst_, S.q, lane);
  __syncwarp();
  frags_to_tile(S.k, acc, 0.125f, 0.125f, g, t);
  __syncwarp();
  unstage_tile(S.d, drow, ld, L, lane);
  unstage_tile(S.k, drow + D, ld, L, lane);
  unstage_tile(S.v, drow + 2 * D, ld, L, lane);
}


// ------------------------------------------------------------------------------------------- 16 < L <= 32
// Two 16-row query tiles and two 16-key tiles per problem (class names of 8+ tokens, e.g. real ImageNet prompts).
// Same building blocks; the score tile of one query tile is 16 x 32 (four n-tiles), P / dS live in a 32 x 32 shared
// tile, dK / dV accumulate over the query tiles.
constexpr int kR2 = 32;     // rows per staged matrix
constexpr int kXLd2 = 40;   // P / dS tile row stride (32 + 8 pad): 80-byte rows, conflict-free ldmatrix

__device__ __forceinline__ void stage_tile32(__nv_bfloat16* tile, const __nv_bfloat16* __restrict__ src, long long ld,
                                             int L, int lane) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int idx = lane + 32 * i, r = idx >> 3, c = idx & 7;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (r < L) v = __ldg(reinterpret_cast<const uint4*>(src + (long long)r * ld) + c);
    *reinterpret_cast<uint4*>(tile + r * kLd + c * 8) = v;
  }
}
// rows [row0, row0 + 16) of a staged tile to global rows row0.. (only rows < L)
__device__ __forceinline__ void unstage_rows16(const __nv_bfloat16* tile16, __nv_bfloat16* __restrict__ dst, long long ld,
                                               int row0, int L, int lane) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int idx = lane + 32 * i, r = idx >> 3, c = idx & 7;
    if (row0 + r < L) *(reinterpret_cast<uint4*>(dst + (long long)(row0 + r) * ld) + c) =
        *reinterpret_cast<const uint4*>(tile16 + r * kLd + c * 8);
  }
}
// C[16 x 8*NT] = A[16 x 64] . B[8*NT x 64]^T  (A, B: pointers to the first row of the operand inside a staged tile)
template <int NT, bool F16 = false>
__device__ __forceinline__ void mma_abt_n(float (&c)[NT][4], const __nv_bfloat16* A, const __nv_bfloat16* B, int lane) {
  const int arow = (lane & 7) + ((lane >> 3) & 1) * 8, acol = (lane >> 4) * 8;
  const int brow = lane & 7, bcol = (lane >> 3) * 8;
#pragma unroll
  for (int n = 0; n < NT; ++n)
#pragma unroll
    for (int i = 0; i < 4; ++i) c[n][i] = 0.f;
#pragma unroll
  for (int kk = 0; kk < 4; kk += 2) {
    uint32_t a0[4], a1[4];
    ldsm4(a0, A + arow * kLd + kk * 16 + acol);
    ldsm4(a1, A + arow * kLd + (kk + 1) * 16 + acol);
#pragma unroll
    for (int n = 0; n < NT; ++n) {
      uint32_t b[4];
      ldsm4(b, B + (8 * n + brow) * kLd + kk * 16 + bcol);
      mma16816<F16>(c[n], a0, b[0], b[1]);
      mma16816<F16>(c[n], a1, b[2], b[3]);
    }
  }
}
// C[16 x 64] += A[16 x 16] . B[16 x 64]  (B: pointer to the first of the 16 k-rows)
template <bool F16 = false>
__device__ __forceinline__ void mma_ab_acc(float (&c)[8][4], const uint32_t (&a)[4], const __nv_bfloat16* B, int lane) {
  const int brow = (lane & 7) + ((lane >> 3) & 1) * 8, bcol = (lane >> 4) * 8;
#pragma unroll
  for (int n = 0; n < 8; n += 2) {
    uint32_t b[4];
    ldsm4t(b, B + brow * kLd + n * 8 + bcol);
    mma16816<F16>(c[n], a, b[0], b[1]);
    mma16816<F16>(c[n + 1], a, b[2], b[3]);
  }
}
__device__ __forceinline__ void zero_acc(float (&c)[8][4]) {
#pragma unroll
  for (int n = 0; n < 8; ++n)
#pragma unroll
    for (int i = 0; i < 4; ++i) c[n][i] = 0.f;
}
__device__ __forceinline__ float quad_max(float v) {
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
  return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
__device__ __forceinline__ float quad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v + __shfl_xor_sync(0xffffffffu, v, 2);
}

constexpr int kWarps2 = 4;
struct FwdSmem2 {
  __nv_bfloat16 q[kR2 * kLd], k[kR2 * kLd], v[kR2 * kLd];
};

template <bool kCausal, bool kF16>
__global__ void __launch_bounds__(kWarps2 * 32)
attn_small_mma2_fwd_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out,
                           float* __restrict__ lse, int n_prob, int L, int H) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pr = blockIdx.x * kWarps2 + warp;
  if (pr >= n_prob) return;
  FwdSmem2& S = reinterpret_cast<FwdSmem2*>(smem_raw)[warp];
  const int seq = pr / H, h = pr % H, D = H * 64;
  const long long ld = 3LL * D;
  const __nv_bfloat16* base = qkv + (long long)seq * L * ld + h * 64;
  stage_tile32(S.q, base, ld, L, lane);
  stage_tile32(S.k, base + D, ld, L, lane);
  stage_tile32(S.v, base + 2 * D, ld, L, lane);
  __syncwarp();
  const int g = lane >> 2, t = lane & 3;
  const int n_qt = (L + 15) >> 4;
#pragma unroll 1
  for (int qt = 0; qt < n_qt; ++qt) {
    float s[4][4];
    mma_abt_n<4, kF16>(s, S.q + qt * 16 * kLd, S.k, lane);
    float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = qt * 16 + g + (i >> 1) * 8, col = n * 8 + 2 * t + (i & 1);
        const bool ok = col < L && (!kCausal || col <= row);
        s[n][i] = ok ? s[n][i] * (0.125f * kLog2e) : -INFINITY;
        if (i < 2) m0 = fmaxf(m0, s[n][i]); else m1 = fmaxf(m1, s[n][i]);
      }
    m0 = quad_max(m0);
    m1 = quad_max(m1);
    float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        s[n][i] = ex2(s[n][i] - ((i < 2) ? m0 : m1));
        if (i < 2) sum0 += s[n][i]; else sum1 += s[n][i];
      }
    sum0 = quad_sum(sum0);
    sum1 = quad_sum(sum1);
    float o[8][4];
    zero_acc(o);
#pragma unroll
    for (int kt = 0; kt < 2; ++kt) {
      const uint32_t p[4] = {pack_bf16<kF16>(s[2 * kt][0], s[2 * kt][1]), pack_bf16<kF16>(s[2 * kt][2], s[2 * kt][3]),
                             pack_bf16<kF16>(s[2 * kt + 1][0], s[2 * kt + 1][1]), pack_bf16<kF16>(s[2 * kt + 1][2], s[2 * kt + 1][3])};
      mma_ab_acc<kF16>(o, p, S.v + kt * 16 * kLd, lane);
    }
    __syncwarp();                       // all lanes have read Q rows of this tile; they become the O staging rows
    frags_to_tile<kF16>(S.q + qt * 16 * kLd, o, 1.f / sum0, 1.f / sum1, g, t);
    if (t == 0) {
      float* l = lse + ((long long)seq * H + h) * L;
      const int r0 = qt * 16 + g;
      if (r0 < L) l[r0] = (m0 + log2f(sum0)) * kLn2;
      if (r0 + 8 < L) l[r0 + 8] = (m1 + log2f(sum1)) * kLn2;
    }
  }
  __syncwarp();
  __nv_bfloat16* orow = out + (long long)seq * L * D + h * 64;
  unstage_rows16(S.q, orow, (long long)D, 0, L, lane);
  unstage_rows16(S.q + 16 * kLd, orow, (long long)D, 16, L, lane);
}

struct BwdSmem2 {
  __nv_bfloat16 q[kR2 * kLd], k[kR2 * kLd], v[kR2 * kLd], d[kR2 * kLd];
  __nv_bfloat16 p[kR2 * kXLd2], ds[kR2 * kXLd2];
  __nv_bfloat16 stg[16 * kLd];
};

template <bool kCausal, bool kF16>
__global__ void __launch_bounds__(kWarps2 * 32)
attn_small_mma2_bwd_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ dout,
                           const float* __restrict__ lse, __nv_bfloat16* __restrict__ dqkv, int n_prob, int L, int H) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pr = blockIdx.x * kWarps2 + warp;
  if (pr >= n_prob) return;
  BwdSmem2& S = reinterpret_cast<BwdSmem2*>(smem_raw)[warp];
  const int seq = pr / H, h = pr % H, D = H * 64;
  const long long ld = 3LL * D;
  const __nv_bfloat16* base = qkv + (long long)seq * L * ld + h * 64;
  stage_tile32(S.q, base, ld, L, lane);
  stage_tile32(S.k, base + D, ld, L, lane);
  stage_tile32(S.v, base + 2 * D, ld, L, lane);
  stage_tile32(S.d, dout + (long long)seq * L * D + h * 64, (long long)D, L, lane);
  const int g = lane >> 2, t = lane & 3;
  const float* lrow = lse + ((long long)seq * H + h) * L;
  __nv_bfloat16* drow = dqkv + (long long)seq * L * ld + h * 64;
  const int n_t = (L + 15) >> 4;          // query tiles = key tiles
  __syncwarp();
  // fp16 forward operands: the scores of both query tiles from the fp16 q / k (as the forward computed them), then
  // q / k / v become bf16 in place for the gradient contractions
  float s_all[2][4][4];
  if (kF16) {
#pragma unroll
    for (int qt = 0; qt < 2; ++qt) mma_abt_n<4, true>(s_all[qt], S.q + qt * 16 * kLd, S.k, lane);
    __syncwarp();
    tile_f16_to_bf16<32>(S.q, lane);
    tile_f16_to_bf16<32>(S.k, lane);
    tile_f16_to_bf16<32>(S.v, lane);
    __syncwarp();
  }
  // ---- pass 1 (per query tile): P, dS into the shared tiles; dQ
#pragma unroll(kF16 ? 2 : 1)
  for (int qt = 0; qt < 2; ++qt) {
    float s[4][4], dp[4][4];
    if (kF16) {
#pragma unroll
      for (int n = 0; n < 4; ++n)
#pragma unroll
        for (int i = 0; i < 4; ++i) s[n][i] = s_all[qt][n][i];
    } else {
      mma_abt_n<4>(s, S.q + qt * 16 * kLd, S.k, lane);
    }
    mma_abt_n<4>(dp, S.d + qt * 16 * kLd, S.v, lane);
    const int r0 = qt * 16 + g;
    const float l0 = (r0 < L) ? lrow[r0] * kLog2e : 0.f, l1 = (r0 + 8 < L) ? lrow[r0 + 8] * kLog2e : 0.f;
    float del0 = 0.f, del1 = 0.f;
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = r0 + (i >> 1) * 8, col = n * 8 + 2 * t + (i & 1);
        const bool ok = row < L && col < L && (!kCausal || col <= row);
        s[n][i] = ok ? ex2(s[n][i] * (0.125f * kLog2e) - ((i < 2) ? l0 : l1)) : 0.f;
        if (i < 2) del0 = fmaf(s[n][i], dp[n][i], del0); else del1 = fmaf(s[n][i], dp[n][i], del1);
      }
    del0 = quad_sum(del0);
    del1 = quad_sum(del1);
    uint32_t dsf[8];
#pragma unroll
    for (int n = 0; n < 4; ++n) {
      dsf[2 * n] = pack_bf16(s[n][0] * (dp[n][0] - del0), s[n][1] * (dp[n][1] - del0));
      dsf[2 * n + 1] = pack_bf16(s[n][2] * (dp[n][2] - del1), s[n][3] * (dp[n][3] - del1));
      const int c = n * 8 + 2 * t;
      *reinterpret_cast<uint32_t*>(S.ds + (qt * 16 + g) * kXLd2 + c) = dsf[2 * n];
      *reinterpret_cast<uint32_t*>(S.ds + (qt * 16 + g + 8) * kXLd2 + c) = dsf[2 * n + 1];
      *reinterpret_cast<uint32_t*>(S.p + (qt * 16 + g) * kXLd2 + c) = pack_bf16(s[n][0], s[n][1]);
      *reinterpret_cast<uint32_t*>(S.p + (qt * 16 + g + 8) * kXLd2 + c) = pack_bf16(s[n][2], s[n][3]);
    }
    if (qt < n_t) {
      float acc[8][4];
      zero_acc(acc);
#pragma unroll
      for (int kt = 0; kt < 2; ++kt) {
        const uint32_t a[4] = {dsf[4 * kt], dsf[4 * kt + 1], dsf[4 * kt + 2], dsf[4 * kt + 3]};
        mma_ab_acc(acc, a, S.k + kt * 16 * kLd, lane);       // dQ_qt += dS[qt][kt] K_kt
      }
      __syncwarp();                                          // previous unstage of the staging tile is complete
      frags_to_tile(S.stg, acc, 0.125f, 0.125f, g, t);
      __syncwarp();
      unstage_rows16(S.stg, drow, ld, qt * 16, L, lane);
    }
  }
  __syncwarp();
  // ---- pass 2 (per key tile): dV = P^T dO, dK = 0.125 dS^T Q, accumulated over the query tiles
  const int xr = (lane & 7) + (lane >> 4) * 8, xc = ((lane >> 3) & 1) * 8;
#pragma unroll 1
  for (int kt = 0; kt < n_t; ++kt) {
#pragma unroll 1
    for (int which = 0; which < 2; ++which) {              // 0: dV (P, dO)   1: dK (dS, Q)
      const __nv_bfloat16* X = which ? S.ds : S.p;
      const __nv_bfloat16* Bm = which ? S.q : S.d;
      float acc[8][4];
      zero_acc(acc);
#pragma unroll
      for (int qt = 0; qt < 2; ++qt) {
        if (kCausal && qt < kt) continue;                  // P[qt][kt] = 0 above the diagonal
        uint32_t a[4];
        ldsm4t(a, X + (qt * 16 + xr) * kXLd2 + kt * 16 + xc);
        mma_ab_acc(acc, a, Bm + qt * 16 * kLd, lane);
      }
      __syncwarp();
      const float sc = which ? 0.125f : 1.f;
      frags_to_tile(S.stg, acc, sc, sc, g, t);
      __syncwarp();
      unstage_rows16(S.stg, drow + (which ? D : 2 * D), ld, kt * 16, L, lane);
    }
  }
}

}  // namespace asm2

bool attn_small_mma_supported(int L) { return L >= 1 && L <= 32; }

int attn_fwd_small_mma(const void* qkv, void* out, float* lse, int dtype, int n_seq, int L, int H, int causal,
                       cudaStream_t s) {
  using namespace asm2;
  const int n_prob = n_seq * H;
  const int grid = (n_prob + kWarps - 1) / kWarps;
  const __nv_bfloat16* q = reinterpret_cast<const __nv_bfloat16*>(qkv);   // 16-bit storage; the format is kF16
  __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
  const bool f16 = dtype == PSRC_F16;
#define PSRC_SM_PICK(KERNEL, ...)                                  \
  do {                                                             \
    if (causal && f16) KERNEL<true, true> __VA_ARGS__;             \
    else if (causal) KERNEL<true, false> __VA_ARGS__;              \
    else if (f16) KERNEL<false, true> __VA_ARGS__;                 \
    else KERNEL<false, false> __VA_ARGS__;                         \
  } while (0)
#define PSRC_SM_ATTR(KERNEL, BYTES)                                                                              \
  do {                                                                                                           \
    PSRC_CHECK_CUDA(cudaFuncSetAttribute(KERNEL<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BYTES));   \
    PSRC_CHECK_CUDA(cudaFuncSetAttribute(KERNEL<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, BYTES));  \
    PSRC_CHECK_CUDA(cudaFuncSetAttribute(KERNEL<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BYTES));  \
    PSRC_CHECK_CUDA(cudaFuncSetAttribute(KERNEL<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, BYTES)); \
  } while (0)
  if (L > 16) {
    const int bytes = kWarps2 * (int)sizeof(FwdSmem2);
    static bool attr2 = false;
    if (!attr2) {
      PSRC_SM_ATTR(attn_small_mma2_fwd_kernel, bytes);
      attr2 = true;
    }
    const int grid2 = (n_prob + kWarps2 - 1) / kWarps2;
    PSRC_SM_PICK(attn_small_mma2_fwd_kernel, <<<grid2, kWarps2 * 32, bytes, s>>>(q, o, lse, n_prob, L, H));
    PSRC_CHECK_LAUNCH();
    return PSRC_OK;
  }
  PSRC_SM_PICK(attn_small_mma_fwd_kernel, <<<grid, kWarps * 32, 0, s>>>(q, o, lse, n_prob, L, H));
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

// qkv_dtype: PSRC_BF16 or PSRC_F16 (forward operands); dout / dqkv are always bf16
int attn_bwd_small_mma(const void* qkv, const void* dout, const float* lse, void* dqkv, int qkv_dtype, int n_seq, int L,
                       int H, int causal, cudaStream_t s) {
  using namespace asm2;
  const int n_prob = n_seq * H;
  const bool f16 = qkv_dtype == PSRC_F16;
  const __nv_bfloat16* q = reinterpret_cast<const __nv_bfloat16*>(qkv);
  const __nv_bfloat16* d = reinterpret_cast<const __nv_bfloat16*>(dout);
  __nv_bfloat16* dq = reinterpret_cast<__nv_bfloat16*>(dqkv);
  if (L > 16) {
    const int bytes2 = kWarps2 * (int)sizeof(BwdSmem2);
    static bool attr2 = false;
    if (!attr2) {
      PSRC_SM_ATTR(attn_small_mma2_bwd_kernel, bytes2);
      attr2 = true;
    }
    const int grid2 = (n_prob + kWarps2 - 1) / kWarps2;
    PSRC_SM_PICK(attn_small_mma2_bwd_kernel, <<<grid2, kWarps2 * 32, bytes2, s>>>(q, d, lse, dq, n_prob, L, H));
    PSRC_CHECK_LAUNCH();
    return PSRC_OK;
  }
  const int grid = (n_prob + kWarps - 1) / kWarps;
  const int bytes = kWarps * (int)sizeof(BwdSmem);
  static bool attr_set = false;
  if (!attr_set) {
    PSRC_SM_ATTR(attn_small_mma_bwd_kernel, bytes);
    attr_set = true;
  }
  PSRC_SM_PICK(attn_small_mma_bwd_kernel, <<<grid, kWarps * 32, bytes, s>>>(q, d, lse, dq, n_prob, L, H));
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
#undef PSRC_SM_PICK
#undef PSRC_SM_ATTR
}

}  // namespace psrc
