// One-query attention for the LAST transformer block: only one row per sequence leaves the tower (the class token,
// ln_post(x[:, 0]) -- PromptSRC/clip/model.py:426; the EOT row of the text tower, trainers/promptsrc.py:64-68), so
// the attention core of that block is needed for ONE query per (sequence, head): scores against every key, one
// softmax row, one output row; in the backward the gradient enters through that row only, so dK / dV of every key
// come from one row of dS / P and dQ is zero everywhere else.  Row-wise the arithmetic is the full kernel's.
//
// One warp per (sequence, head).  A K / V row (64 dims) is split into 16-byte chunks over 8 lanes (16 for fp32), so a
// warp-wide access covers 4 (2) whole rows: coalesced.  Scores are reduced over a row's lanes with shuffles; the
// forward keeps an online softmax per row group and merges the groups at the end; the backward keeps P_j and
// dP_j of the one query row in shared memory between its two passes.  HBM-bound: the forward reads
// K and V of every token once (79 MB at cfg 3: B = 128, L = 201, D = 768), the backward reads them again and writes
// the whole dqkv tensor (zeros for the dQ rows of the other tokens).
#include <type_traits>
#include "common.cuh"

namespace psrc {
namespace arow {

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr int kWarps = 4;
constexpr int kMaxKeysPerLane = 9;    // L <= 288 (ViT-L/14: 257 + prompts)

template <typename T> struct Vec16;   // 16 bytes of T as floats
template <> struct Vec16<float> {
  static constexpr int kN = 4;
  __device__ static void load(const float* p, float* v) {
    const float4 f = *reinterpret_cast<const float4*>(p);
    v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
  }
  __device__ static void store(float* p, const float* v) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <> struct Vec16<__nv_bfloat16> {
  static constexpr int kN = 8;
  __device__ static void load(const __nv_bfloat16* p, float* v) {
    const uint4 u = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __bfloat1622float2(h[i]);
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
  }
  __device__ static void store(__nv_bfloat16* p, const float* v) {
    uint4 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = u;
  }
};
template <> struct Vec16<__half> {
  static constexpr int kN = 8;
  __device__ static void load(const __half* p, float* v) {
    const uint4 u = *reinterpret_cast<const uint4*>(p);
    const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(h[i]);
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
  }
  __device__ static void store(__half* p, const float* v) {
    uint4 u;
    __half2* h = reinterpret_cast<__half2*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = u;
  }
};

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Lane layout: a 64-dim row of T is kLPR = 64 / kN sixteen-byte chunks; lane l owns chunk (l % kLPR) of the key
// (l / kLPR) of each group of kRPI = 32 / kLPR keys, so one warp-wide 16-byte access covers kRPI whole rows
// (four 128-byte lines for the 16-bit formats) -- fully coalesced.
template <typename T> struct Lay {
  static constexpr int kN = Vec16<T>::kN;
  static constexpr int kLPR = 64 / kN;
  static constexpr int kRPI = 32 / kLPR;
};
template <int LPR>
__device__ __forceinline__ float row_sum(float v) {      // sum over the kLPR lanes that share a row
#pragma unroll
  for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
template <int N>
__device__ __forceinline__ float dotN(const float (&a)[N], const float (&b)[N]) {
  float s = 0.f;
#pragma unroll
  for (int e = 0; e < N; ++e) s = fmaf(a[e], b[e], s);
  return s;
}

// One CTA of kWarps warps per (sequence, head): warp w owns a contiguous quarter of the keys (a chain of only
// ceil(L / 4 / (kU * kRPI)) DRAM round trips per warp, ~40 resident warps per SM), partial results meet in shared memory.
__device__ __forceinline__ void key_range(int nkeys, int rpi, int warp, int& lo, int& hi) {
  const int per = ((nkeys + kWarps - 1) / kWarps + rpi - 1) / rpi * rpi;
  lo = min(nkeys, warp * per);
  hi = min(nkeys, lo + per);
}

// T: storage type of qkv / out.  The query of sequence s is row (idx ? idx[s] : 0); causal: keys <= query row.
// Online softmax per row group (running max / sum / output chunk), merged over the row groups and the warps.
// raw 16-byte loads kept in their own registers until used: unpacking at the load site makes the compiler reuse the
// destination registers and thereby serialise the loads of one trip (seen in the SASS: one DRAM round trip per pair)
__device__ __forceinline__ uint4 ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
template <typename T>
__device__ __forceinline__ void unpack16(const uint4& u, float* v) {
  if constexpr (sizeof(T) == 4) {
    v[0] = __uint_as_float(u.x); v[1] = __uint_as_float(u.y); v[2] = __uint_as_float(u.z); v[3] = __uint_as_float(u.w);
  } else {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f;
      if constexpr (std::is_same<T, __half>::value) f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
      else f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[i]));
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kWarps * 32)
attn_row_fwd_kernel(const T* __restrict__ qkv, T* __restrict__ out, float* __restrict__ lse,
                    const int32_t* __restrict__ idx, int L, int H, int causal) {
  constexpr int N = Lay<T>::kN, LPR = Lay<T>::kLPR, RPI = Lay<T>::kRPI;
  __shared__ float part[kWarps][2 + 64];          // per warp: m, l, 64 output sums
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int seq = blockIdx.x / H, h = blockIdx.x % H, D = H * 64;
  const long long ld = 3LL * D;
  const int qrow = idx ? idx[seq] : 0;
  const int nkeys = causal ? (qrow + 1) : L;
  const int c = lane % LPR, r = lane / LPR;
  const T* base = qkv + (long long)seq * L * ld + h * 64 + c * N;
  float q[N];
  Vec16<T>::load(base + (long long)qrow * ld, q);
  float m = -INFINITY, l = 0.f, acc[N];
#pragma unroll
  for (int e = 0; e < N; ++e) acc[e] = 0.f;
  int klo, khi;
  key_range(nkeys, RPI, warp, klo, khi);
  constexpr int kU = 4;                           // row groups per trip: 2 * kU sixteen-byte loads in flight per lane
  for (int j0 = klo; j0 < khi; j0 += kU * RPI) {
    uint4 kr[kU], vr[kU];
    float sc[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int j = min(j0 + u * RPI + r, khi - 1);          // clamped: always a valid row, masked below
      kr[u] = ldg16(base + (long long)j * ld + D);
      vr[u] = ldg16(base + (long long)j * ld + 2 * D);
    }
    float mn = m;
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      float k[N];
      unpack16<T>(kr[u], k);
      sc[u] = row_sum<LPR>(dotN<N>(k, q)) * (0.125f * kLog2e);
      if (j0 + u * RPI + r >= khi) sc[u] = -INFINITY;
      mn = fmaxf(mn, sc[u]);
    }
    if (mn > -INFINITY) {                         // (a group that has seen no key yet keeps m = -inf, l = 0)
      const float a = ex2(m - mn);
      l *= a;
#pragma unroll
      for (int e = 0; e < N; ++e) acc[e] *= a;
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const float p = ex2(sc[u] - mn);                      // 0 for masked rows (sc = -inf)
        float v[N];
        unpack16<T>(vr[u], v);
        l += p;
#pragma unroll
        for (int e = 0; e < N; ++e) acc[e] = fmaf(p, v[e], acc[e]);
      }
      m = mn;
    }
  }
  // merge the row groups of this warp (lanes with equal chunk index c, different r) ...
#pragma unroll
  for (int o = LPR; o < 32; o <<= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o), l2 = __shfl_xor_sync(0xffffffffu, l, o);
    const float mn = fmaxf(m, m2);
    const float a = (m > -INFINITY) ? ex2(m - mn) : 0.f, a2 = (m2 > -INFINITY) ? ex2(m2 - mn) : 0.f;
    l = l * a + l2 * a2;
#pragma unroll
    for (int e = 0; e < N; ++e) acc[e] = acc[e] * a + __shfl_xor_sync(0xffffffffu, acc[e], o) * a2;
    m = mn;
  }
  if (r == 0) {
#pragma unroll
    for (int e = 0; e < N; ++e) part[warp][2 + c * N + e] = acc[e];
    if (c == 0) {
      part[warp][0] = m;
      part[warp][1] = l;
    }
  }
  __syncthreads();
  // ... and the warps: warp 0, two output dims per lane
  if (warp == 0) {
    float mt = -INFINITY;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) mt = fmaxf(mt, part[w][0]);
    float lt = 0.f, o0 = 0.f, o1 = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const float a = (part[w][0] > -INFINITY) ? ex2(part[w][0] - mt) : 0.f;
      lt = fmaf(part[w][1], a, lt);
      o0 = fmaf(part[w][2 + 2 * lane], a, o0);
      o1 = fmaf(part[w][2 + 2 * lane + 1], a, o1);
    }
    const float inv = 1.f / lt;
    T* orow = out + ((long long)seq * L + qrow) * D + h * 64;
    orow[2 * lane] = from_f32<T>(o0 * inv);
    orow[2 * lane + 1] = from_f32<T>(o1 * inv);
    if (lane == 0) lse[((long long)seq * H + h) * L + qrow] = (mt + log2f(lt)) * kLn2;
  }
}

// T: qkv storage; G: dout / dqkv storage (same element size as T).  dqkv is fully written (zeros where nothing flows).
template <typename T, typename G>
__global__ void __launch_bounds__(kWarps * 32)
attn_row_bwd_kernel(const T* __restrict__ qkv, const G* __restrict__ dout, const float* __restrict__ lse,
                    G* __restrict__ dqkv, const int32_t* __restrict__ idx, int L, int H, int causal) {
  static_assert(sizeof(T) == sizeof(G), "one lane layout for operands and gradients");
  constexpr int N = Lay<T>::kN, LPR = Lay<T>::kLPR, RPI = Lay<T>::kRPI;
  __shared__ float ps[32 * kMaxKeysPerLane], dps[32 * kMaxKeysPerLane];   // P_j and dP_j of the query row
  __shared__ float part[kWarps][1 + 64];                                  // per warp: delta part, dQ part
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int seq = blockIdx.x / H, h = blockIdx.x % H, D = H * 64;
  const long long ld = 3LL * D;
  const int qrow = idx ? idx[seq] : 0;
  const int nkeys = causal ? (qrow + 1) : L;
  const int c = lane % LPR, r = lane / LPR;
  const T* base = qkv + (long long)seq * L * ld + h * 64 + c * N;
  G* dbase = dqkv + (long long)seq * L * ld + h * 64 + c * N;
  const float lse2 = lse[((long long)seq * H + h) * L + qrow] * kLog2e;
  float q[N], go[N];
  Vec16<T>::load(base + (long long)qrow * ld, q);
  Vec16<G>::load(dout + ((long long)seq * L + qrow) * D + h * 64 + c * N, go);
  int klo, khi;
  key_range(nkeys, RPI, warp, klo, khi);
  // ---- pass 1: P_j, dP_j = <dO, V_j>, delta = sum_j P_j dP_j
  float delta = 0.f;
  constexpr int kU = 4;                           // row groups per trip: 2 * kU loads in flight per lane
  for (int j0 = klo; j0 < khi; j0 += kU * RPI) {
    uint4 kr[kU], vr[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int j = min(j0 + u * RPI + r, khi - 1);          // clamped: always a valid row, masked below
      kr[u] = ldg16(base + (long long)j * ld + D);
      vr[u] = ldg16(base + (long long)j * ld + 2 * D);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int j = j0 + u * RPI + r;
      float k[N], v[N];
      unpack16<T>(kr[u], k);
      unpack16<T>(vr[u], v);
      const float sc = row_sum<LPR>(dotN<N>(k, q)), dp = row_sum<LPR>(dotN<N>(v, go));
      if (j < khi && c == 0) {
        const float p = ex2(sc * (0.125f * kLog2e) - lse2);
        ps[j] = p;
        dps[j] = dp;
        delta = fmaf(p, dp, delta);
      }
    }
  }
  delta = warp_sum(delta);
  if (lane == 0) part[warp][0] = delta;
  __syncthreads();
  delta = 0.f;
#pragma unroll
  for (int w = 0; w < kWarps; ++w) delta += part[w][0];
  // ---- pass 2 over this warp's quarter of ALL rows: dK_j = dS_j q / 8, dV_j = P_j dO, zero dQ rows;
  //      dQ of the query row = sum_j dS_j K_j / 8
  int rlo, rhi;
  key_range(L, RPI, warp, rlo, rhi);
  float acc[N], zero[N];
#pragma unroll
  for (int e = 0; e < N; ++e) acc[e] = zero[e] = 0.f;
  for (int j0 = rlo; j0 < rhi; j0 += kU * RPI) {
    uint4 kr[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int j = min(j0 + u * RPI + r, rhi - 1);
      kr[u] = ldg16(base + (long long)j * ld + D);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int j = j0 + u * RPI + r;
      if (j >= rhi) continue;
      G* drow = dbase + (long long)j * ld;
      if (j < nkeys) {
        const float p = ps[j], ds = p * (dps[j] - delta);
        float t[N], k[N];
        unpack16<T>(kr[u], k);
#pragma unroll
        for (int e = 0; e < N; ++e) {
          acc[e] = fmaf(ds, k[e], acc[e]);
          t[e] = 0.125f * ds * q[e];
        }
        Vec16<G>::store(drow + D, t);
#pragma unroll
        for (int e = 0; e < N; ++e) t[e] = p * go[e];
        Vec16<G>::store(drow + 2 * D, t);
      } else {
        Vec16<G>::store(drow + D, zero);
        Vec16<G>::store(drow + 2 * D, zero);
      }
      if (j != qrow) Vec16<G>::store(drow, zero);
    }
  }
#pragma unroll
  for (int o = LPR; o < 32; o <<= 1)
#pragma unroll
    for (int e = 0; e < N; ++e) acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], o);
  if (r == 0) {
#pragma unroll
    for (int e = 0; e < N; ++e) part[warp][1 + c * N + e] = acc[e];
  }
  __syncthreads();
  if (warp == 0) {
    float d0 = 0.f, d1 = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      d0 += part[w][1 + 2 * lane];
      d1 += part[w][1 + 2 * lane + 1];
    }
    G* dq = dqkv + ((long long)seq * L + qrow) * ld + h * 64;
    dq[2 * lane] = from_f32<G>(0.125f * d0);
    dq[2 * lane + 1] = from_f32<G>(0.125f * d1);
  }
}

}  // namespace arow
}  // namespace psrc

using namespace psrc;

static bool al16r(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int psrc_attn_row_fwd(const void* qkv, void* out, float* lse, int dtype, const int32_t* idx, int n_seq,
                                 int L, int H, int causal, psrc_stream_t stream) {
  if (n_seq == 0) return PSRC_OK;   // (empty tensors have NULL storage)
  PSRC_CHECK_ARG(qkv && out && lse, "psrc_attn_row_fwd: null pointer");
  PSRC_CHECK_ARG(n_seq >= 0 && L > 0 && L <= 32 * arow::kMaxKeysPerLane && H > 0,
                 "psrc_attn_row_fwd: bad shape n_seq=%d L=%d H=%d (L <= %d)", n_seq, L, H, 32 * arow::kMaxKeysPerLane);
  PSRC_CHECK_ARG(dtype >= PSRC_F32 && dtype <= PSRC_F16 && al16r(qkv) && al16r(out), "psrc_attn_row_fwd: dtype / alignment");
  if (n_seq == 0) return PSRC_OK;
  const int grid = n_seq * H;
  cudaStream_t s = as_stream(stream);
  if (dtype == PSRC_F32)
    arow::attn_row_fwd_kernel<float><<<grid, arow::kWarps * 32, 0, s>>>(
        reinterpret_cast<const float*>(qkv), reinterpret_cast<float*>(out), lse, idx, L, H, causal);
  else if (dtype == PSRC_BF16)
    arow::attn_row_fwd_kernel<__nv_bfloat16><<<grid, arow::kWarps * 32, 0, s>>>(
        reinterpret_cast<const __nv_bfloat16*>(qkv), reinterpret_cast<__nv_bfloat16*>(out), lse, idx, L, H,
        causal);
  else
    arow::attn_row_fwd_kernel<__half><<<grid, arow::kWarps * 32, 0, s>>>(
        reinterpret_cast<const __half*>(qkv), reinterpret_cast<__half*>(out), lse, idx, L, H, causal);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

extern "C" int psrc_attn_row_bwd(const void* qkv, const void* dout, const float* lse, void* dqkv, int dtype,
                                 int grad_dtype, const int32_t* idx, int n_seq, int L, int H, int causal,
                                 psrc_stream_t stream) {
  if (n_seq == 0) return PSRC_OK;
  PSRC_CHECK_ARG(qkv && dout && lse && dqkv, "psrc_attn_row_bwd: null pointer");
  PSRC_CHECK_ARG(n_seq >= 0 && L > 0 && L <= 32 * arow::kMaxKeysPerLane && H > 0,
                 "psrc_attn_row_bwd: bad shape n_seq=%d L=%d H=%d (L <= %d)", n_seq, L, H, 32 * arow::kMaxKeysPerLane);
  PSRC_CHECK_ARG(al16r(qkv) && al16r(dout) && al16r(dqkv), "psrc_attn_row_bwd: 16-byte alignment");
  const bool ok = (dtype == PSRC_F32 && grad_dtype == PSRC_F32) ||
                  ((dtype == PSRC_BF16 || dtype == PSRC_F16) && grad_dtype == PSRC_BF16);
  PSRC_CHECK_ARG(ok, "psrc_attn_row_bwd: (qkv, gradient) dtypes must be (f32, f32), (bf16, bf16) or (f16, bf16)");
  if (n_seq == 0) return PSRC_OK;
  const int grid = n_seq * H;
  cudaStream_t s = as_stream(stream);
  if (dtype == PSRC_F32)
    arow::attn_row_bwd_kernel<float, float><<<grid, arow::kWarps * 32, 0, s>>>(
        reinterpret_cast<const float*>(qkv), reinterpret_cast<const float*>(dout), lse,
        reinterpret_cast<float*>(dqkv), idx, L, H, causal);
  else if (dtype == PSRC_BF16)
    arow::attn_row_bwd_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, arow::kWarps * 32, 0, s>>>(
        reinterpret_cast<const __nv_bfloat16*>(qkv), reinterpret_cast<const __nv_bfloat16*>(dout), lse,
        reinterpret_cast<__nv_bfloat16*>(dqkv), idx, L, H, causal);
  else
    arow::attn_row_bwd_kernel<__half, __nv_bfloat16><<<grid, arow::kWarps * 32, 0, s>>>(
        reinterpret_cast<const __half*>(qkv), reinterpret_cast<const __nv_bfloat16*>(dout), lse,
        reinterpret_cast<__nv_bfloat16*>(dqkv), idx, L, H, causal);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}
