// LayerNorm forward / input-gradient, one warp per row, 16-byte coalesced accesses, values held
// in registers between the statistics and the normalisation (one HBM read + one write per row).
// Mirrors PromptSRC/clip/model.py:153-159 (fp32 LayerNorm, eps 1e-5, biased variance).  gamma/beta
// are frozen in every trainer on this path (trainers/promptsrc.py:238-247), so there is no
// dgamma/dbeta kernel.  Also the gathered variants for ln_post(x[:,0]) (model.py:426) and
// ln_final at the EOT row (trainers/promptsrc.py:64-68).
#include "common.cuh"

namespace psrc {

constexpr int kLnWarps = 8;
constexpr int kLnBwdWarps = 2;

template <typename T>
__device__ __forceinline__ void store4(T* p, float4 v);
template <>
__device__ __forceinline__ void store4<float>(float* p, float4 v) {
  *reinterpret_cast<float4*>(p) = v;
}
template <>
__device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, float4 v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = u;
}
template <>
__device__ __forceinline__ void store4<__half>(__half* p, float4 v) {
  __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = u;
}
template <typename T>
__device__ __forceinline__ float4 load4(const T* p);
template <>
__device__ __forceinline__ float4 load4<float>(const float* p) {
  return *reinterpret_cast<const float4*>(p);
}
template <>
__device__ __forceinline__ float4 load4<__nv_bfloat16>(const __nv_bfloat16* p) {
  uint2 u = *reinterpret_cast<const uint2*>(p);
  float2 a = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u.x));
  float2 b = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u.y));
  return make_float4(a.x, a.y, b.x, b.y);
}

// NV = ceil(D / 128): float4 slots per lane
template <int NV>
__device__ __forceinline__ void row_stats(const float4 (&v)[NV], int lane, int D, float eps, float& mean,
                                          float& rstd) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i)
    if ((lane + 32 * i) * 4 < D) s += v[i].x + v[i].y + v[i].z + v[i].w;
  s = warp_sum(s);
  mean = s / (float)D;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i)
    if ((lane + 32 * i) * 4 < D) {
      float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
      q += a * a + b * b + c * c + d * d;
    }
  q = warp_sum(q);
  rstd = rsqrtf(q / (float)D + eps);
}

// idx == nullptr and L == 0: plain row-wise LN (in row r -> out row r)
// otherwise: out row s <- LN(x[s*L + (idx ? idx[s] : 0)])
template <typename T, int NV>
__global__ void __launch_bounds__(kLnWarps * 32)
ln_fwd_kernel(const float* __restrict__ x, const int32_t* __restrict__ idx, int L,
              const float* __restrict__ gamma, const float* __restrict__ beta, T* __restrict__ y,
              float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int D, float eps) {
  const int lane = threadIdx.x & 31;
  const int r = blockIdx.x * kLnWarps + (threadIdx.x >> 5);
  if (r >= rows) return;
  long long src = r;
  if (L > 0) src = (long long)r * L + (idx ? idx[r] : 0);
  const float* xr = x + src * D;
  float4 v[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    int c = (lane + 32 * i) * 4;
    v[i] = (c < D) ? load4<float>(xr + c) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  float mean, rstd;
  row_stats<NV>(v, lane, D, eps, mean, rstd);
  if (lane == 0) {
    if (mean_out) mean_out[r] = mean;
    if (rstd_out) rstd_out[r] = rstd;
  }
  T* yr = y + (long long)r * D;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    int c = (lane + 32 * i) * 4;
    if (c < D) {
      float4 g = load4<float>(gamma + c), b = load4<float>(beta + c);
      float4 o;
      o.x = (v[i].x - mean) * rstd * g.x + b.x;
      o.y = (v[i].y - mean) * rstd * g.y + b.y;
      o.z = (v[i].z - mean) * rstd * g.z + b.z;
      o.w = (v[i].w - mean) * rstd * g.w + b.w;
      store4<T>(yr + c, o);
    }
  }
}

// dx_out[r] = (dx_in ? dx_in[r] : 0) + rstd * (g*dy - mean(g*dy) - xhat * mean(g*dy*xhat))
// gathered form (L > 0): dy row s applies to x row s*L+idx[s]; every other row of dx_out is zeroed.
template <typename T, typename TA, int NV, int W>
__global__ void __launch_bounds__(W * 32)
ln_bwd_kernel(const T* __restrict__ dy, const float* __restrict__ x, const int32_t* __restrict__ idx, int L,
              const float* __restrict__ gamma, const float* __restrict__ mean_in,
              const float* __restrict__ rstd_in, const float* dx_in, float* dx_out, TA* __restrict__ dx_act,
              int rows, int D) {
  const int lane = threadIdx.x & 31;
  const int r = blockIdx.x * W + (threadIdx.x >> 5);  // row of x / dx
  if (r >= rows) return;
  int s = r;  // row of dy / mean / rstd
  if (L > 0) {
    s = r / L;
    const int pos = r - s * L;
    const int want = idx ? idx[s] : 0;
    if (pos != want) {
      float* o = dx_out + (long long)r * D;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        int c = (lane + 32 * i) * 4;
        if (c < D) {
          store4<float>(o + c, make_float4(0.f, 0.f, 0.f, 0.f));
          if (dx_act) store4<TA>(dx_act + (long long)r * D + c, make_float4(0.f, 0.f, 0.f, 0.f));
        }
      }
      return;
    }
  }
  const float mean = mean_in[s], rstd = rstd_in[s];
  const float* xr = x + (long long)r * D;
  const T* dyr = dy + (long long)s * D;
  float4 xh[NV], gd[NV], res[NV];
  float s1 = 0.f, s2 = 0.f;
  // the residual-gradient row is fetched up front with the other operands: dx_out may alias dx_in, so a load placed
  // after the first store below could not be hoisted by the compiler and every slot would pay a full DRAM round trip
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    int c = (lane + 32 * i) * 4;
    res[i] = (dx_in && c < D) ? load4<float>(dx_in + (long long)r * D + c) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    int c = (lane + 32 * i) * 4;
    if (c < D) {
      float4 xv = load4<float>(xr + c), d = load4<T>(dyr + c), g = load4<float>(gamma + c);
      xh[i] = make_float4((xv.x - mean) * rstd, (xv.y - mean) * rstd, (xv.z - mean) * rstd, (xv.w - mean) * rstd);
      gd[i] = make_float4(d.x * g.x, d.y * g.y, d.z * g.z, d.w * g.w);
      s1 += gd[i].x + gd[i].y + gd[i].z + gd[i].w;
      s2 += gd[i].x * xh[i].x + gd[i].y * xh[i].y + gd[i].z * xh[i].z + gd[i].w * xh[i].w;
    } else {
      xh[i] = gd[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  s1 = warp_sum(s1) / (float)D;
  s2 = warp_sum(s2) / (float)D;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    int c = (lane + 32 * i) * 4;
    if (c < D) {
      float4 o;
      o.x = rstd * (gd[i].x - s1 - xh[i].x * s2);
      o.y = rstd * (gd[i].y - s1 - xh[i].y * s2);
      o.z = rstd * (gd[i].z - s1 - xh[i].z * s2);
      o.w = rstd * (gd[i].w - s1 - xh[i].w * s2);
      o.x += res[i].x; o.y += res[i].y; o.z += res[i].z; o.w += res[i].w;
      store4<float>(dx_out + (long long)r * D + c, o);
      if (dx_act) store4<TA>(dx_act + (long long)r * D + c, o);
    }
  }
}

template <typename T>
static int launch_fwd(const float* x, const int32_t* idx, int L, const float* g, const float* b, void* y,
                      float* mean, float* rstd, int rows, int D, float eps, cudaStream_t s) {
  const int nv = (D + 127) / 128;
  const int grid = (rows + kLnWarps - 1) / kLnWarps;
  T* yt = reinterpret_cast<T*>(y);
#define LN_CASE(NV)                                                                                   \
  case NV:                                                                                            \
    ln_fwd_kernel<T, NV><<<grid, kLnWarps * 32, 0, s>>>(x, idx, L, g, b, yt, mean, rstd, rows, D, eps); \
    break;
  switch (nv) {
    LN_CASE(1) LN_CASE(2) LN_CASE(3) LN_CASE(4) LN_CASE(5) LN_CASE(6) LN_CASE(7) LN_CASE(8)
    default:
      set_error("layernorm: D=%d not supported (max 1024)", D);
      return PSRC_ERR_UNSUPPORTED;
  }
#undef LN_CASE
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

template <typename T, typename TA>
static int launch_bwd(const void* dy, const float* x, const int32_t* idx, int L, const float* g,
                      const float* mean, const float* rstd, const float* dx_in, float* dx_out, void* dx_act,
                      int rows, int D, cudaStream_t s) {
  const int nv = (D + 127) / 128;
  const T* dyt = reinterpret_cast<const T*>(dy);
  TA* at = reinterpret_cast<TA*>(dx_act);
  // two rows per 64-thread block: at 92 registers per thread an 8-warp block leaves only 16 warps per SM and their
  // load / reduce / store phases line up; small blocks measured 67.6 us vs 77.8 us (25 728 x 768, cold L2)
#define LN_CASE(NV)                                                                                         \
  case NV:                                                                                                  \
    ln_bwd_kernel<T, TA, NV, kLnBwdWarps><<<(rows + kLnBwdWarps - 1) / kLnBwdWarps, kLnBwdWarps * 32, 0, s>>>( \
        dyt, x, idx, L, g, mean, rstd, dx_in, dx_out, at, rows, D);                                         \
    break;
  switch (nv) {
    LN_CASE(1) LN_CASE(2) LN_CASE(3) LN_CASE(4) LN_CASE(5) LN_CASE(6) LN_CASE(7) LN_CASE(8)
    default:
      set_error("layernorm: D=%d not supported (max 1024)", D);
      return PSRC_ERR_UNSUPPORTED;
  }
#undef LN_CASE
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

static int bwd_dispatch(const void* dy, int dy_dtype, const float* x, const int32_t* idx, int L,
                        const float* g, const float* mean, const float* rstd, const float* dx_in,
                        float* dx_out, void* dx_act, int act_dtype, int rows, int D, cudaStream_t s) {
  const bool act_bf16 = dx_act && act_dtype == PSRC_BF16;
  if (dy_dtype == PSRC_BF16) {
    if (act_bf16)
      return launch_bwd<__nv_bfloat16, __nv_bfloat16>(dy, x, idx, L, g, mean, rstd, dx_in, dx_out, dx_act, rows, D, s);
    return launch_bwd<__nv_bfloat16, float>(dy, x, idx, L, g, mean, rstd, dx_in, dx_out, dx_act, rows, D, s);
  }
  if (act_bf16)
    return launch_bwd<float, __nv_bfloat16>(dy, x, idx, L, g, mean, rstd, dx_in, dx_out, dx_act, rows, D, s);
  return launch_bwd<float, float>(dy, x, idx, L, g, mean, rstd, dx_in, dx_out, dx_act, rows, D, s);
}

}  // namespace psrc

using namespace psrc;

#define CHECK_LN_COMMON(name)                                                                   \
  PSRC_CHECK_ARG(D > 0 && D % 4 == 0 && D <= 1024, name ": D=%d must be a multiple of 4, <= 1024", D); \
  PSRC_CHECK_ARG(rows >= 0, name ": negative rows");

extern "C" int psrc_layernorm_fwd(const float* x, const float* gamma, const float* beta, void* y,
                                  int out_dtype, float* mean, float* rstd, int rows, int D, float eps,
                                  psrc_stream_t stream) {
  CHECK_LN_COMMON("psrc_layernorm_fwd");
  if (rows == 0) return PSRC_OK;   // (empty tensors have NULL storage)
  PSRC_CHECK_ARG(x && gamma && beta && y, "psrc_layernorm_fwd: null pointer");
  if (out_dtype == PSRC_F16)
    return launch_fwd<__half>(x, nullptr, 0, gamma, beta, y, mean, rstd, rows, D, eps, as_stream(stream));
  if (out_dtype == PSRC_BF16)
    return launch_fwd<__nv_bfloat16>(x, nullptr, 0, gamma, beta, y, mean, rstd, rows, D, eps, as_stream(stream));
  return launch_fwd<float>(x, nullptr, 0, gamma, beta, y, mean, rstd, rows, D, eps, as_stream(stream));
}

extern "C" int psrc_layernorm_bwd(const void* dy, int dy_dtype, const float* x, const float* gamma,
                                  const float* mean, const float* rstd, const float* dx_in, float* dx_out,
                                  void* dx_act, int act_dtype, int rows, int D, psrc_stream_t stream) {
  CHECK_LN_COMMON("psrc_layernorm_bwd");
  if (rows == 0) return PSRC_OK;
  PSRC_CHECK_ARG(dy && x && gamma && mean && rstd && dx_out, "psrc_layernorm_bwd: null pointer");
  return bwd_dispatch(dy, dy_dtype, x, nullptr, 0, gamma, mean, rstd, dx_in, dx_out, dx_act, act_dtype, rows, D,
                      as_stream(stream));
}

extern "C" int psrc_gather_ln_fwd(const float* x, const int32_t* idx, const float* gamma, const float* beta,
                                  void* y, int out_dtype, float* mean, float* rstd, int n_seq, int L, int D,
                                  float eps, psrc_stream_t stream) {
  const int rows = n_seq;
  CHECK_LN_COMMON("psrc_gather_ln_fwd");
  if (rows == 0) return PSRC_OK;
  PSRC_CHECK_ARG(x && gamma && beta && y && L > 0, "psrc_gather_ln_fwd: bad arguments");
  if (out_dtype == PSRC_F16)
    return launch_fwd<__half>(x, idx, L, gamma, beta, y, mean, rstd, rows, D, eps, as_stream(stream));
  if (out_dtype == PSRC_BF16)
    return launch_fwd<__nv_bfloat16>(x, idx, L, gamma, beta, y, mean, rstd, rows, D, eps, as_stream(stream));
  return launch_fwd<float>(x, idx, L, gamma, beta, y, mean, rstd, rows, D, eps, as_stream(stream));
}

extern "C" int psrc_gather_ln_bwd(const void* dy, int dy_dtype, const float* x, const int32_t* idx,
                                  const float* gamma, const float* mean, const float* rstd,
                                  const float* dx_in, float* dx, void* dx_act, int act_dtype, int n_seq, int L,
                                  int D, psrc_stream_t stream) {
  const int rows = n_seq * L;
  CHECK_LN_COMMON("psrc_gather_ln_bwd");
  if (rows == 0) return PSRC_OK;
  PSRC_CHECK_ARG(dy && x && gamma && mean && rstd && dx && L > 0, "psrc_gather_ln_bwd: bad arguments");
  return bwd_dispatch(dy, dy_dtype, x, idx, L, gamma, mean, rstd, dx_in, dx, dx_act, act_dtype, rows, D,
                      as_stream(stream));
}
