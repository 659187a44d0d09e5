// CUDA-core attention (validation mode, any activation dtype, fp32 math throughout).
// One CTA per (sequence, head); K/V (forward) resp. K/V then Q/dO (backward) staged in shared
// memory as fp32; one thread per query row (forward, dQ) or key row (dK, dV).  No atomics, fixed
// summation order.  Mirrors nn.MultiheadAttention's scaled-dot-product core as called from
// PromptSRC/clip/model.py:183/227/285 with the causal mask of model.py:592-598 for text.
#include "common.cuh"

namespace psrc {

constexpr int kHD = 64;
constexpr int kAttnThreads = 128;

template <typename T>
__device__ __forceinline__ void load_row64(const T* p, float (&v)[kHD]) {
#pragma unroll
  for (int d = 0; d < kHD; ++d) v[d] = to_f32(p[d]);
}
__device__ __forceinline__ float dot64(const float (&a)[kHD], const float* __restrict__ s) {
  float acc = 0.f;
#pragma unroll
  for (int d = 0; d < kHD; d += 4) {
    float4 k = *reinterpret_cast<const float4*>(s + d);
    acc = fmaf(a[d], k.x, acc);
    acc = fmaf(a[d + 1], k.y, acc);
    acc = fmaf(a[d + 2], k.z, acc);
    acc = fmaf(a[d + 3], k.w, acc);
  }
  return acc;
}

template <typename T>
__global__ void __launch_bounds__(kAttnThreads)
attn_fwd_simt_kernel(const T* __restrict__ qkv, T* __restrict__ out, float* __restrict__ lse, int L, int H,
                     int causal, float scale) {
  extern __shared__ float sm[];
  float* Ks = sm;
  float* Vs = sm + (size_t)L * kHD;
  const int h = blockIdx.x % H, seq = blockIdx.x / H;   // linear grid: no 65 535 limit on sequences
  const int D = H * kHD, ld = 3 * D;
  const T* base = qkv + (long long)seq * L * ld + h * kHD;
  for (int idx = threadIdx.x; idx < L * kHD; idx += blockDim.x) {
    const int r = idx / kHD, d = idx % kHD;
    Ks[idx] = to_f32(base[(long long)r * ld + D + d]);
    Vs[idx] = to_f32(base[(long long)r * ld + 2 * D + d]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < L; i += blockDim.x) {
    float q[kHD], o[kHD];
    load_row64(base + (long long)i * ld, q);
#pragma unroll
    for (int d = 0; d < kHD; ++d) { q[d] *= scale; o[d] = 0.f; }
    float m = -INFINITY, l = 0.f;
    const int jmax = causal ? i : L - 1;
    for (int j = 0; j <= jmax; ++j) {
      const float s = dot64(q, Ks + j * kHD);
      const float mn = fmaxf(m, s);
      const float corr = expf(m - mn);
      const float p = expf(s - mn);
      l = l * corr + p;
      const float* v = Vs + j * kHD;
#pragma unroll
      for (int d = 0; d < kHD; d += 4) {
        float4 vv = *reinterpret_cast<const float4*>(v + d);
        o[d] = fmaf(p, vv.x, o[d] * corr);
        o[d + 1] = fmaf(p, vv.y, o[d + 1] * corr);
        o[d + 2] = fmaf(p, vv.z, o[d + 2] * corr);
        o[d + 3] = fmaf(p, vv.w, o[d + 3] * corr);
      }
      m = mn;
    }
    const float inv = 1.f / l;
    T* op = out + ((long long)seq * L + i) * D + h * kHD;
#pragma unroll
    for (int d = 0; d < kHD; ++d) op[d] = from_f32<T>(o[d] * inv);
    lse[((long long)seq * H + h) * L + i] = m + logf(l);
  }
}

template <typename T>
__global__ void __launch_bounds__(kAttnThreads)
attn_bwd_simt_kernel(const T* __restrict__ qkv, const T* __restrict__ out, const T* __restrict__ dout,
                     const float* __restrict__ lse, T* __restrict__ dqkv, int L, int H, int causal, float scale) {
  extern __shared__ float sm[];
  float* bufA = sm;                          // K, then Q
  float* bufB = sm + (size_t)L * kHD;        // V, then dO
  float* lse_s = sm + (size_t)2 * L * kHD;
  float* delta_s = lse_s + L;
  const int h = blockIdx.x % H, seq = blockIdx.x / H;   // linear grid: no 65 535 limit on sequences
  const int D = H * kHD, ld = 3 * D;
  const T* base = qkv + (long long)seq * L * ld + h * kHD;
  const T* obase = out + (long long)seq * L * D + h * kHD;
  const T* dobase = dout + (long long)seq * L * D + h * kHD;
  T* dbase = dqkv + (long long)seq * L * ld + h * kHD;

  for (int idx = threadIdx.x; idx < L * kHD; idx += blockDim.x) {
    const int r = idx / kHD, d = idx % kHD;
    bufA[idx] = to_f32(base[(long long)r * ld + D + d]);
    bufB[idx] = to_f32(base[(long long)r * ld + 2 * D + d]);
  }
  for (int i = threadIdx.x; i < L; i += blockDim.x) {
    float acc = 0.f;
    for (int d = 0; d < kHD; ++d) acc += to_f32(obase[(long long)i * D + d]) * to_f32(dobase[(long long)i * D + d]);
    delta_s[i] = acc;
    lse_s[i] = lse[((long long)seq * H + h) * L + i];
  }
  __syncthreads();

  // ---- pass A: dQ_i = scale * sum_j dS_ij K_j, two halves of 32 dims to bound registers
  for (int i = threadIdx.x; i < L; i += blockDim.x) {
    float q[kHD], dO[kHD];
    load_row64(base + (long long)i * ld, q);
    load_row64(dobase + (long long)i * D, dO);
    const float li = lse_s[i], di = delta_s[i];
    const int jmax = causal ? i : L - 1;
    for (int half = 0; half < 2; ++half) {
      float dq[32];
#pragma unroll
      for (int d = 0; d < 32; ++d) dq[d] = 0.f;
      for (int j = 0; j <= jmax; ++j) {
        const float s = scale * dot64(q, bufA + j * kHD);
        const float p = expf(s - li);
        const float dp = dot64(dO, bufB + j * kHD);
        const float ds = p * (dp - di);
        const float* k = bufA + j * kHD + half * 32;
#pragma unroll
        for (int d = 0; d < 32; ++d) dq[d] = fmaf(ds, k[d], dq[d]);
      }
      T* dqp = dbase + (long long)i * ld + half * 32;
#pragma unroll
      for (int d = 0; d < 32; ++d) dqp[d] = from_f32<T>(dq[d] * scale);
    }
  }
  __syncthreads();
  // ---- restage: bufA <- Q, bufB <- dO
  for (int idx = threadIdx.x; idx < L * kHD; idx += blockDim.x) {
    const int r = idx / kHD, d = idx % kHD;
    bufA[idx] = to_f32(base[(long long)r * ld + d]);
    bufB[idx] = to_f32(dobase[(long long)r * D + d]);
  }
  __syncthreads();
  // ---- pass B: per key j: dV_j = sum_i P_ij dO_i ; dK_j = scale * sum_i dS_ij Q_i
  for (int j = threadIdx.x; j < L; j += blockDim.x) {
    float k[kHD], v[kHD];
    load_row64(base + (long long)j * ld + D, k);
    load_row64(base + (long long)j * ld + 2 * D, v);
    const int i0 = causal ? j : 0;
    for (int half = 0; half < 2; ++half) {
      float dk[32], dv[32];
#pragma unroll
      for (int d = 0; d < 32; ++d) { dk[d] = 0.f; dv[d] = 0.f; }
      for (int i = i0; i < L; ++i) {
        const float s = scale * dot64(k, bufA + i * kHD);
        const float p = expf(s - lse_s[i]);
        const float dp = dot64(v, bufB + i * kHD);
        const float ds = p * (dp - delta_s[i]);
        const float* qi = bufA + i * kHD + half * 32;
        const float* doi = bufB + i * kHD + half * 32;
#pragma unroll
        for (int d = 0; d < 32; ++d) {
          dk[d] = fmaf(ds, qi[d], dk[d]);
          dv[d] = fmaf(p, doi[d], dv[d]);
        }
      }
      T* dkp = dbase + (long long)j * ld + D + half * 32;
      T* dvp = dbase + (long long)j * ld + 2 * D + half * 32;
#pragma unroll
      for (int d = 0; d < 32; ++d) {
        dkp[d] = from_f32<T>(dk[d] * scale);
        dvp[d] = from_f32<T>(dv[d]);
      }
    }
  }
}

template <typename K>
static int set_smem(K kernel, size_t bytes) {
  if (bytes > 48 * 1024) {
    PSRC_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  }
  return PSRC_OK;
}

int attn_fwd_simt(const void* qkv, void* out, float* lse, int dtype, int n_seq, int L, int H, int causal,
                  cudaStream_t s) {
  const size_t smem = (size_t)2 * L * kHD * sizeof(float);
  if (smem > 227 * 1024) {
    set_error("attn_fwd (CUDA-core): L=%d too long for shared memory", L);
    return PSRC_ERR_UNSUPPORTED;
  }
  dim3 grid((unsigned)H * (unsigned)n_seq);
  const float scale = 0.125f;  // 1/sqrt(64)
  int rc;
  if (dtype == PSRC_BF16) {
    if ((rc = set_smem(attn_fwd_simt_kernel<__nv_bfloat16>, smem))) return rc;
    attn_fwd_simt_kernel<__nv_bfloat16><<<grid, kAttnThreads, smem, s>>>(
        reinterpret_cast<const __nv_bfloat16*>(qkv), reinterpret_cast<__nv_bfloat16*>(out), lse, L, H, causal, scale);
  } else {
    if ((rc = set_smem(attn_fwd_simt_kernel<float>, smem))) return rc;
    attn_fwd_simt_kernel<float><<<grid, kAttnThreads, smem, s>>>(
        reinterpret_cast<const float*>(qkv), reinterpret_cast<float*>(out), lse, L, H, causal, scale);
  }
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

int attn_bwd_simt(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, int dtype,
                  int n_seq, int L, int H, int causal, cudaStream_t s) {
  const size_t smem = ((size_t)2 * L * kHD + 2 * L) * sizeof(float);
  if (smem > 227 * 1024) {
    set_error("attn_bwd (CUDA-core): L=%d too long for shared memory", L);
    return PSRC_ERR_UNSUPPORTED;
  }
  dim3 grid((unsigned)H * (unsigned)n_seq);
  const float scale = 0.125f;
  int rc;
  if (dtype == PSRC_BF16) {
    if ((rc = set_smem(attn_bwd_simt_kernel<__nv_bfloat16>, smem))) return rc;
    attn_bwd_simt_kernel<__nv_bfloat16><<<grid, kAttnThreads, smem, s>>>(
        reinterpret_cast<const __nv_bfloat16*>(qkv), reinterpret_cast<const __nv_bfloat16*>(out),
        reinterpret_cast<const __nv_bfloat16*>(dout), lse, reinterpret_cast<__nv_bfloat16*>(dqkv), L, H, causal, scale);
  } else {
    if ((rc = set_smem(attn_bwd_simt_kernel<float>, smem))) return rc;
    attn_bwd_simt_kernel<float><<<grid, kAttnThreads, smem, s>>>(
        reinterpret_cast<const float*>(qkv), reinterpret_cast<const float*>(out),
        reinterpret_cast<const float*>(dout), lse, reinterpret_cast<float*>(dqkv), L, H, causal, scale);
  }
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

}  // namespace psrc
