// Tensor-core attention for the bf16 fast mode (head_dim 64, whole sequence resident on chip).
// One CTA per (sequence, head); K/V (and Q/dO for the backward) live in XOR-swizzled shared memory;
// each warp owns 16-row tiles and runs flash-style online softmax over 64-key blocks with
// mma.sync.m16n8k16 (bf16 in, fp32 accumulate).  The backward is atomic-free and deterministic:
// pass 1 (warp per 16-key tile) produces dK/dV, pass 2 (warp per 16-query tile) produces dQ, both
// recomputing P from the saved log-sum-exp.  Sequences here are 77..259 tokens, so a whole head
// fits one CTA ("long context: not applicable").
// Reference: nn.MultiheadAttention core called from PromptSRC/clip/model.py:183/227/285, causal
// mask model.py:592-598.
#include "common.cuh"

namespace psrc {

constexpr int kAT = 128;  // threads per CTA (4 warps)
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
// byte offset of element (row, dim) in a [rows][64] bf16 tile with 128-byte rows, 16-byte chunks XOR-swizzled
__device__ __forceinline__ uint32_t sw_off(int row, int dim) {
  return (uint32_t)(row * 128 + ((((dim >> 3) ^ (row & 7)) << 4) | ((dim & 7) << 1)));
}
__device__ __forceinline__ uint32_t lds32(const uint8_t* tile, int row, int dim) {
  return *reinterpret_cast<const uint32_t*>(tile + sw_off(row, dim));
}
// B-operand fragments for two adjacent 8-wide n-tiles (n0, n0+8) of a row-major [k][n] tile, k-step k0..k0+15
__device__ __forceinline__ void ldsm_x4_trans(const uint8_t* tile, int k0, int n0, int lane, uint32_t (&b)[4]) {
  const int mi = lane >> 3;
  const int row = k0 + ((mi & 1) << 3) + (lane & 7);
  const int dim = n0 + ((mi >> 1) << 3);
  const uint32_t addr = smem_u32(tile + sw_off(row, dim));
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3])
               : "r"(addr));
}
// A-operand fragments (16 rows x 64 dims -> 4 k-steps) of a row-major swizzled tile
__device__ __forceinline__ void load_a_frags(const uint8_t* tile, int r0, int g, int t, uint32_t (&a)[4][4]) {
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
    a[ks][0] = lds32(tile, r0 + g, 16 * ks + 2 * t);
    a[ks][1] = lds32(tile, r0 + g + 8, 16 * ks + 2 * t);
    a[ks][2] = lds32(tile, r0 + g, 16 * ks + 8 + 2 * t);
    a[ks][3] = lds32(tile, r0 + g + 8, 16 * ks + 8 + 2 * t);
  }
}
// global [L][64] slice (row stride ld elements) -> swizzled smem tile of Lp rows (rows >= L zero)
__device__ __forceinline__ void stage_tile(uint8_t* tile, const __nv_bfloat16* src, long long ld, int L, int Lp) {
  for (int idx = threadIdx.x; idx < Lp * 8; idx += blockDim.x) {
    const int r = idx >> 3, c = idx & 7;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (r < L) v = *reinterpret_cast<const uint4*>(src + (long long)r * ld + c * 8);
    *reinterpret_cast<uint4*>(tile + r * 128 + ((c ^ (r & 7)) << 4)) = v;
  }
}
__device__ __forceinline__ float quad_max(float v) {
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
  return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
__device__ __forceinline__ float quad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v + __shfl_xor_sync(0xffffffffu, v, 2);
}

template <bool kCausal>
__global__ void __launch_bounds__(kAT)
attn_fwd_mma_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out,
                    float* __restrict__ lse, int L, int H, float scale_log2) {
  extern __shared__ __align__(128) uint8_t smem_attn[];
  const int Lp = (L + 15) & ~15;
  uint8_t* Ks = smem_attn;
  uint8_t* Vs = smem_attn + (size_t)Lp * 128;
  const int h = blockIdx.x % H, seq = blockIdx.x / H;   // linear grid: no 65 535 limit on sequences
  const int D = H * 64;
  const long long ld = 3LL * D;
  const __nv_bfloat16* base = qkv + (long long)seq * L * ld + h * 64;
  stage_tile(Ks, base + D, ld, L, Lp);
  stage_tile(Vs, base + 2 * D, ld, L, Lp);
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int n_kb_all = (Lp + 63) >> 6;
  for (int qt = warp; qt < (Lp >> 4); qt += kAT / 32) {
    const int r0 = qt * 16;
    const int rowA = r0 + g, rowB = r0 + g + 8;
    uint32_t qa[4][4];
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      const __nv_bfloat16* pa = base + (long long)rowA * ld + 16 * ks + 2 * t;
      const __nv_bfloat16* pb = base + (long long)rowB * ld + 16 * ks + 2 * t;
      qa[ks][0] = rowA < L ? *reinterpret_cast<const uint32_t*>(pa) : 0u;
      qa[ks][1] = rowB < L ? *reinterpret_cast<const uint32_t*>(pb) : 0u;
      qa[ks][2] = rowA < L ? *reinterpret_cast<const uint32_t*>(pa + 8) : 0u;
      qa[ks][3] = rowB < L ? *reinterpret_cast<const uint32_t*>(pb + 8) : 0u;
    }
    float o[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    int n_kb = n_kb_all;
    if (kCausal) n_kb = min(n_kb_all, ((r0 + 15) >> 6) + 1);
    for (int kb = 0; kb < n_kb; ++kb) {
      float s[8][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
        const int key0 = kb * 64 + nt * 8;
        if (key0 < Lp) {
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            const uint32_t b0 = lds32(Ks, key0 + g, 16 * ks + 2 * t);
            const uint32_t b1 = lds32(Ks, key0 + g, 16 * ks + 8 + 2 * t);
            mma16816(s[nt], qa[ks], b0, b1);
          }
        }
      }
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int col = kb * 64 + nt * 8 + 2 * t + (e & 1);
          const int row = (e & 2) ? rowB : rowA;
          const bool valid = col < L && (!kCausal || col <= row);
          const float v = valid ? s[nt][e] * scale_log2 : -INFINITY;
          s[nt][e] = v;
          if (e & 2) mx1 = fmaxf(mx1, v); else mx0 = fmaxf(mx0, v);
        }
      }
      mx0 = quad_max(mx0);
      mx1 = quad_max(mx1);
      const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
      const float ms0 = (mn0 == -INFINITY) ? 0.f : mn0, ms1 = (mn1 == -INFINITY) ? 0.f : mn1;
      const float c0 = exp2f(m0 - ms0), c1 = exp2f(m1 - ms1);
      float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        s[nt][0] = exp2f(s[nt][0] - ms0);
        s[nt][1] = exp2f(s[nt][1] - ms0);
        s[nt][2] = exp2f(s[nt][2] - ms1);
        s[nt][3] = exp2f(s[nt][3] - ms1);
        ps0 += s[nt][0] + s[nt][1];
        ps1 += s[nt][2] + s[nt][3];
      }
      l0 = l0 * c0 + ps0;
      l1 = l1 * c1 + ps1;
      m0 = mn0;
      m1 = mn1;
#pragma unroll
      for (int nd = 0; nd < 8; ++nd) {
        o[nd][0] *= c0; o[nd][1] *= c0; o[nd][2] *= c1; o[nd][3] *= c1;
      }
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const int k0 = kb * 64 + kk * 16;
        if (k0 < Lp) {
          uint32_t pa[4];
          pa[0] = pack_bf16(s[2 * kk][0], s[2 * kk][1]);
          pa[1] = pack_bf16(s[2 * kk][2], s[2 * kk][3]);
          pa[2] = pack_bf16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
          pa[3] = pack_bf16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
          for (int ndp = 0; ndp < 4; ++ndp) {
            uint32_t b[4];
            ldsm_x4_trans(Vs, k0, ndp * 16, lane, b);
            mma16816(o[2 * ndp], pa, b[0], b[1]);
            mma16816(o[2 * ndp + 1], pa, b[2], b[3]);
          }
        }
      }
    }
    l0 = quad_sum(l0);
    l1 = quad_sum(l1);
    const float inv0 = 1.f / l0, inv1 = 1.f / l1;
    if (rowA < L) {
      __nv_bfloat16* op = out + ((long long)seq * L + rowA) * D + h * 64 + 2 * t;
#pragma unroll
      for (int nd = 0; nd < 8; ++nd)
        *reinterpret_cast<__nv_bfloat162*>(op + nd * 8) = __floats2bfloat162_rn(o[nd][0] * inv0, o[nd][1] * inv0);
      if (t == 0) lse[((long long)seq * H + h) * L + rowA] = (m0 + log2f(l0)) * kLn2;
    }
    if (rowB < L) {
      __nv_bfloat16* op = out + ((long long)seq * L + rowB) * D + h * 64 + 2 * t;
#pragma unroll
      for (int nd = 0; nd < 8; ++nd)
        *reinterpret_cast<__nv_bfloat162*>(op + nd * 8) = __floats2bfloat162_rn(o[nd][2] * inv1, o[nd][3] * inv1);
      if (t == 0) lse[((long long)seq * H + h) * L + rowB] = (m1 + log2f(l1)) * kLn2;
    }
  }
}

// ------------------------------------------------------------------------------------ backward
template <bool kCausal>
__global__ void __launch_bounds__(kAT)
attn_bwd_mma_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ out,
                    const __nv_bfloat16* __restrict__ dout, const float* __restrict__ lse,
                    __nv_bfloat16* __restrict__ dqkv, int L, int H, float scale, float scale_log2) {
  extern __shared__ __align__(128) uint8_t smem_attn[];
  const int Lp = (L + 15) & ~15;
  uint8_t* Qs = smem_attn;
  uint8_t* Ks = Qs + (size_t)Lp * 128;
  uint8_t* Vs = Ks + (size_t)Lp * 128;
  uint8_t* dOs = Vs + (size_t)Lp * 128;
  float* lse2_s = reinterpret_cast<float*>(dOs + (size_t)Lp * 128);  // lse * log2(e)
  float* delta_s = lse2_s + Lp;
  const int h = blockIdx.x % H, seq = blockIdx.x / H;   // linear grid: no 65 535 limit on sequences
  const int D = H * 64;
  const long long ld = 3LL * D;
  const __nv_bfloat16* base = qkv + (long long)seq * L * ld + h * 64;
  const __nv_bfloat16* obase = out + (long long)seq * L * D + h * 64;
  const __nv_bfloat16* dobase = dout + (long long)seq * L * D + h * 64;
  __nv_bfloat16* dbase = dqkv + (long long)seq * L * ld + h * 64;
  stage_tile(Qs, base, ld, L, Lp);
  stage_tile(Ks, base + D, ld, L, Lp);
  stage_tile(Vs, base + 2 * D, ld, L, Lp);
  stage_tile(dOs, dobase, D, L, Lp);
  for (int i = threadIdx.x; i < Lp; i += blockDim.x) {
    float acc = 0.f, l2 = 0.f;
    if (i < L) {
      const uint4* po = reinterpret_cast<const uint4*>(obase + (long long)i * D);
      const uint4* pd = reinterpret_cast<const uint4*>(dobase + (long long)i * D);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        uint4 a = po[c], b = pd[c];
        const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&a);
        const __nv_bfloat162* hb = reinterpret_cast<const __nv_bfloat162*>(&b);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float2 fa = __bfloat1622float2(ha[j]), fb = __bfloat1622float2(hb[j]);
          acc = fmaf(fa.x, fb.x, acc);
          acc = fmaf(fa.y, fb.y, acc);
        }
      }
      l2 = lse[((long long)seq * H + h) * L + i] * kLog2e;
    }
    delta_s[i] = acc;
    lse2_s[i] = l2;
  }
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int n_tiles = Lp >> 4;
  const int n_blk = (Lp + 31) >> 5;  // 32-wide column blocks

  // ---------------- pass 1: warp per 16-key tile -> dK, dV
  for (int kt = warp; kt < n_tiles; kt += kAT / 32) {
    const int k0 = kt * 16;
    const int keyA = k0 + g, keyB = k0 + g + 8;
    uint32_t ka[4][4], va[4][4];
    load_a_frags(Ks, k0, g, t, ka);
    load_a_frags(Vs, k0, g, t, va);
    float dk[8][4], dv[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      dk[i][0] = dk[i][1] = dk[i][2] = dk[i][3] = 0.f;
      dv[i][0] = dv[i][1] = dv[i][2] = dv[i][3] = 0.f;
    }
    const int qb0 = kCausal ? (k0 >> 5) : 0;
    for (int qb = qb0; qb < n_blk; ++qb) {
      float st[4][4], dp[4][4];
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        st[nt][0] = st[nt][1] = st[nt][2] = st[nt][3] = 0.f;
        dp[nt][0] = dp[nt][1] = dp[nt][2] = dp[nt][3] = 0.f;
        const int q0 = qb * 32 + nt * 8;
        if (q0 < Lp) {
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            mma16816(st[nt], ka[ks], lds32(Qs, q0 + g, 16 * ks + 2 * t), lds32(Qs, q0 + g, 16 * ks + 8 + 2 * t));
            mma16816(dp[nt], va[ks], lds32(dOs, q0 + g, 16 * ks + 2 * t), lds32(dOs, q0 + g, 16 * ks + 8 + 2 * t));
          }
        }
      }
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int qi = qb * 32 + nt * 8 + 2 * t + (e & 1);
          const int key = (e & 2) ? keyB : keyA;
          const bool valid = key < L && qi < L && (!kCausal || key <= qi);
          float p = 0.f, ds = 0.f;
          if (valid) {
            p = exp2f(st[nt][e] * scale_log2 - lse2_s[qi]);
            ds = p * (dp[nt][e] - delta_s[qi]);
          }
          st[nt][e] = p;
          dp[nt][e] = ds;
        }
      }
#pragma unroll
      for (int kk = 0; kk < 2; ++kk) {
        const int q0 = qb * 32 + kk * 16;
        if (q0 < Lp) {
          uint32_t pa[4], da[4];
          pa[0] = pack_bf16(st[2 * kk][0], st[2 * kk][1]);
          pa[1] = pack_bf16(st[2 * kk][2], st[2 * kk][3]);
          pa[2] = pack_bf16(st[2 * kk + 1][0], st[2 * kk + 1][1]);
          pa[3] = pack_bf16(st[2 * kk + 1][2], st[2 * kk + 1][3]);
          da[0] = pack_bf16(dp[2 * kk][0], dp[2 * kk][1]);
          da[1] = pack_bf16(dp[2 * kk][2], dp[2 * kk][3]);
          da[2] = pack_bf16(dp[2 * kk + 1][0], dp[2 * kk + 1][1]);
          da[3] = pack_bf16(dp[2 * kk + 1][2], dp[2 * kk + 1][3]);
#pragma unroll
          for (int ndp = 0; ndp < 4; ++ndp) {
            uint32_t b[4];
            ldsm_x4_trans(dOs, q0, ndp * 16, lane, b);
            mma16816(dv[2 * ndp], pa, b[0], b[1]);
            mma16816(dv[2 * ndp + 1], pa, b[2], b[3]);
            ldsm_x4_trans(Qs, q0, ndp * 16, lane, b);
            mma16816(dk[2 * ndp], da, b[0], b[1]);
            mma16816(dk[2 * ndp + 1], da, b[2], b[3]);
          }
        }
      }
    }
    if (keyA < L) {
      __nv_bfloat16* pk = dbase + (long long)keyA * ld + D + 2 * t;
      __nv_bfloat16* pv = dbase + (long long)keyA * ld + 2 * D + 2 * t;
#pragma unroll
      for (int nd = 0; nd < 8; ++nd) {
        *reinterpret_cast<__nv_bfloat162*>(pk + nd * 8) = __floats2bfloat162_rn(dk[nd][0] * scale, dk[nd][1] * scale);
        *reinterpret_cast<__nv_bfloat162*>(pv + nd * 8) = __floats2bfloat162_rn(dv[nd][0], dv[nd][1]);
      }
    }
    if (keyB < L) {
      __nv_bfloat16* pk = dbase + (long long)keyB * ld + D + 2 * t;
      __nv_bfloat16* pv = dbase + (long long)keyB * ld + 2 * D + 2 * t;
#pragma unroll
      for (int nd = 0; nd < 8; ++nd) {
        *reinterpret_cast<__nv_bfloat162*>(pk + nd * 8) = __floats2bfloat162_rn(dk[nd][2] * scale, dk[nd][3] * scale);
        *reinterpret_cast<__nv_bfloat162*>(pv + nd * 8) = __floats2bfloat162_rn(dv[nd][2], dv[nd][3]);
      }
    }
  }

  // ---------------- pass 2: warp per 16-query tile -> dQ
  for (int qt = warp; qt < n_tiles; qt += kAT / 32) {
    const int r0 = qt * 16;
    const int rowA = r0 + g, rowB = r0 + g + 8;
    uint32_t qa[4][4], doa[4][4];
    load_a_frags(Qs, r0, g, t, qa);
    load_a_frags(dOs, r0, g, t, doa);
    const float lA = lse2_s[rowA], lB = lse2_s[rowB];
    const float dA = delta_s[rowA], dB = delta_s[rowB];
    float dq[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) dq[i][0] = dq[i][1] = dq[i][2] = dq[i][3] = 0.f;
    int n_kb = n_blk;
    if (kCausal) n_kb = min(n_blk, ((r0 + 15) >> 5) + 1);
    for (int kb = 0; kb < n_kb; ++kb) {
      float s[4][4], dp[4][4];
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
        dp[nt][0] = dp[nt][1] = dp[nt][2] = dp[nt][3] = 0.f;
        const int key0 = kb * 32 + nt * 8;
        if (key0 < Lp) {
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            mma16816(s[nt], qa[ks], lds32(Ks, key0 + g, 16 * ks + 2 * t), lds32(Ks, key0 + g, 16 * ks + 8 + 2 * t));
            mma16816(dp[nt], doa[ks], lds32(Vs, key0 + g, 16 * ks + 2 * t), lds32(Vs, key0 + g, 16 * ks + 8 + 2 * t));
          }
        }
      }
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int col = kb * 32 + nt * 8 + 2 * t + (e & 1);
          const int row = (e & 2) ? rowB : rowA;
          const bool valid = col < L && row < L && (!kCausal || col <= row);
          float ds = 0.f;
          if (valid) {
            const float p = exp2f(s[nt][e] * scale_log2 - ((e & 2) ? lB : lA));
            ds = p * (dp[nt][e] - ((e & 2) ? dB : dA));
          }
          s[nt][e] = ds;
        }
      }
#pragma unroll
      for (int kk = 0; kk < 2; ++kk) {
        const int k0 = kb * 32 + kk * 16;
        if (k0 < Lp) {
          uint32_t da[4];
          da[0] = pack_bf16(s[2 * kk][0], s[2 * kk][1]);
          da[1] = pack_bf16(s[2 * kk][2], s[2 * kk][3]);
          da[2] = pack_bf16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
          da[3] = pack_bf16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
          for (int ndp = 0; ndp < 4; ++ndp) {
            uint32_t b[4];
            ldsm_x4_trans(Ks, k0, ndp * 16, lane, b);
            mma16816(dq[2 * ndp], da, b[0], b[1]);
            mma16816(dq[2 * ndp + 1], da, b[2], b[3]);
          }
        }
      }
    }
    if (rowA < L) {
      __nv_bfloat16* p = dbase + (long long)rowA * ld + 2 * t;
#pragma unroll
      for (int nd = 0; nd < 8; ++nd)
        *reinterpret_cast<__nv_bfloat162*>(p + nd * 8) = __floats2bfloat162_rn(dq[nd][0] * scale, dq[nd][1] * scale);
    }
    if (rowB < L) {
      __nv_bfloat16* p = dbase + (long long)rowB * ld + 2 * t;
#pragma unroll
      for (int nd = 0; nd < 8; ++nd)
        *reinterpret_cast<__nv_bfloat162*>(p + nd * 8) = __floats2bfloat162_rn(dq[nd][2] * scale, dq[nd][3] * scale);
    }
  }
}

template <typename K>
static int set_smem_attr(K kernel, size_t bytes) {
  if (bytes > 48 * 1024)
    PSRC_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return PSRC_OK;
}

int attn_fwd_mma(const void* qkv, void* out, float* lse, int n_seq, int L, int H, int causal, cudaStream_t s) {
  const int Lp = (L + 15) & ~15;
  const size_t smem = (size_t)2 * Lp * 128;
  if (smem > 227 * 1024) {
    set_error("attn_fwd (tensor-core): L=%d too long for shared memory", L);
    return PSRC_ERR_UNSUPPORTED;
  }
  dim3 grid((unsigned)H * (unsigned)n_seq);
  const float scale_log2 = 0.125f * kLog2e;
  const __nv_bfloat16* q = reinterpret_cast<const __nv_bfloat16*>(qkv);
  __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
  int rc;
  if (causal) {
    if ((rc = set_smem_attr(attn_fwd_mma_kernel<true>, smem))) return rc;
    attn_fwd_mma_kernel<true><<<grid, kAT, smem, s>>>(q, o, lse, L, H, scale_log2);
  } else {
    if ((rc = set_smem_attr(attn_fwd_mma_kernel<false>, smem))) return rc;
    attn_fwd_mma_kernel<false><<<grid, kAT, smem, s>>>(q, o, lse, L, H, scale_log2);
  }
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

int attn_bwd_mma(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, int n_seq,
                 int L, int H, int causal, cudaStream_t s) {
  const int Lp = (L + 15) & ~15;
  const size_t smem = (size_t)4 * Lp * 128 + (size_t)2 * Lp * sizeof(float);
  if (smem > 227 * 1024) {
    set_error("attn_bwd (tensor-core): L=%d too long for shared memory", L);
    return PSRC_ERR_UNSUPPORTED;
  }
  dim3 grid((unsigned)H * (unsigned)n_seq);
  const float scale = 0.125f, scale_log2 = 0.125f * kLog2e;
  const __nv_bfloat16* q = reinterpret_cast<const __nv_bfloat16*>(qkv);
  const __nv_bfloat16* o = reinterpret_cast<const __nv_bfloat16*>(out);
  const __nv_bfloat16* d = reinterpret_cast<const __nv_bfloat16*>(dout);
  __nv_bfloat16* dq = reinterpret_cast<__nv_bfloat16*>(dqkv);
  int rc;
  if (causal) {
    if ((rc = set_smem_attr(attn_bwd_mma_kernel<true>, smem))) return rc;
    attn_bwd_mma_kernel<true><<<grid, kAT, smem, s>>>(q, o, d, lse, dq, L, H, scale, scale_log2);
  } else {
    if ((rc = set_smem_attr(attn_bwd_mma_kernel<false>, smem))) return rc;
    attn_bwd_mma_kernel<false><<<grid, kAT, smem, s>>>(q, o, d, lse, dq, L, H, scale, scale_log2);
  }
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

}  // namespace psrc
