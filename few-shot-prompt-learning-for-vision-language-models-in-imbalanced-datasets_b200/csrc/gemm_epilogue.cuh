// Fused GEMM epilogue shared by the tcgen05 and the CUDA-core kernels:
//   y = act(alpha*acc + bias[n]) (+ residual[m,n]);  optional second output = pre-activation.
#pragma once
#include "common.cuh"

namespace psrc {

struct EpiParams {
  int M, N;
  const float* bias;
  int act;
  const void* aux; long long ldaux; int aux_dtype;
  const float* residual; long long ldres;
  void* out; long long ldo; int out_dtype;
  void* out2; long long ldo2; int out2_dtype;
  float alpha;
  int vec_ok;  // all row strides / base pointers allow 16-byte vector access and N % 8 == 0
};

__device__ __forceinline__ void load_row_chunk(const void* base, int dtype, long long ld, int row,
                                               int col0, float (&v)[32]) {
  if (dtype == PSRC_F16) {
    const uint4* p = reinterpret_cast<const uint4*>(reinterpret_cast<const __half*>(base) + (long long)row * ld + col0);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4 u = p[i];
      const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 f = __half22float2(h[j]);
        v[i * 8 + j * 2] = f.x;
        v[i * 8 + j * 2 + 1] = f.y;
      }
    }
  } else if (dtype == PSRC_BF16) {
    const uint4* p = reinterpret_cast<const uint4*>(
        reinterpret_cast<const __nv_bfloat16*>(base) + (long long)row * ld + col0);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4 u = p[i];
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 f = __bfloat1622float2(h[j]);
        v[i * 8 + j * 2] = f.x;
        v[i * 8 + j * 2 + 1] = f.y;
      }
    }
  } else {
    const float4* p =
        reinterpret_cast<const float4*>(reinterpret_cast<const float*>(base) + (long long)row * ld + col0);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float4 f = p[i];
      v[i * 4] = f.x; v[i * 4 + 1] = f.y; v[i * 4 + 2] = f.z; v[i * 4 + 3] = f.w;
    }
  }
}

__device__ __forceinline__ void store_row_chunk(void* base, int dtype, long long ld, int row, int col0,
                                                const float (&v)[32]) {
  if (dtype == PSRC_F16) {
    uint4* p = reinterpret_cast<uint4*>(reinterpret_cast<__half*>(base) + (long long)row * ld + col0);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4 u;
      __half2* h = reinterpret_cast<__half2*>(&u);
#pragma unroll
      for (int j = 0; j < 4; ++j) h[j] = __floats2half2_rn(v[i * 8 + j * 2], v[i * 8 + j * 2 + 1]);
      p[i] = u;
    }
  } else if (dtype == PSRC_BF16) {
    uint4* p = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(base) + (long long)row * ld + col0);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4 u;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
      for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(v[i * 8 + j * 2], v[i * 8 + j * 2 + 1]);
      p[i] = u;
    }
  } else {
    float4* p = reinterpret_cast<float4*>(reinterpret_cast<float*>(base) + (long long)row * ld + col0);
#pragma unroll
    for (int i = 0; i < 8; ++i) p[i] = make_float4(v[i * 4], v[i * 4 + 1], v[i * 4 + 2], v[i * 4 + 3]);
  }
}

__device__ __forceinline__ float load_elem(const void* base, int dtype, long long idx) {
  if (dtype == PSRC_F16) return __half2float(reinterpret_cast<const __half*>(base)[idx]);
  return dtype == PSRC_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[idx])
                            : reinterpret_cast<const float*>(base)[idx];
}
__device__ __forceinline__ void store_elem(void* base, int dtype, long long idx, float v) {
  if (dtype == PSRC_F16) reinterpret_cast<__half*>(base)[idx] = __float2half_rn(v);
  else if (dtype == PSRC_BF16) reinterpret_cast<__nv_bfloat16*>(base)[idx] = __float2bfloat16_rn(v);
  else reinterpret_cast<float*>(base)[idx] = v;
}

// Fused epilogue for one thread's 32 consecutive columns of one row.
__device__ __forceinline__ void epilogue_chunk(const EpiParams& p, int row, int col0, float (&v)[32]) {
  const bool full = p.vec_ok && (col0 + 32 <= p.N);
  if (full) {
    if (p.bias) {
      const float4* b = reinterpret_cast<const float4*>(p.bias + col0);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float4 f = __ldg(b + i);
        v[i * 4] = fmaf(p.alpha, v[i * 4], f.x);
        v[i * 4 + 1] = fmaf(p.alpha, v[i * 4 + 1], f.y);
        v[i * 4 + 2] = fmaf(p.alpha, v[i * 4 + 2], f.z);
        v[i * 4 + 3] = fmaf(p.alpha, v[i * 4 + 3], f.w);
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] *= p.alpha;
    }
    if (p.act == PSRC_ACT_QUICKGELU) {
      if (p.out2) store_row_chunk(p.out2, p.out2_dtype, p.ldo2, row, col0, v);
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = quickgelu(v[i]);
    } else if (p.act == PSRC_ACT_QUICKGELU_GRAD) {
      float u[32];
      load_row_chunk(p.aux, p.aux_dtype, p.ldaux, row, col0, u);
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] *= quickgelu_grad(u[i]);
    }
    if (p.residual) {
      float r[32];
      load_row_chunk(p.residual, PSRC_F32, p.ldres, row, col0, r);
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += r[i];
    }
    store_row_chunk(p.out, p.out_dtype, p.ldo, row, col0, v);
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      int col = col0 + i;
      if (col >= p.N) continue;
      float x = p.alpha * v[i] + (p.bias ? __ldg(p.bias + col) : 0.f);
      if (p.act == PSRC_ACT_QUICKGELU) {
        if (p.out2) store_elem(p.out2, p.out2_dtype, (long long)row * p.ldo2 + col, x);
        x = quickgelu(x);
      } else if (p.act == PSRC_ACT_QUICKGELU_GRAD) {
        x *= quickgelu_grad(load_elem(p.aux, p.aux_dtype, (long long)row * p.ldaux + col));
      }
      if (p.residual) x += p.residual[(long long)row * p.ldres + col];
      store_elem(p.out, p.out_dtype, (long long)row * p.ldo + col, x);
    }
  }
}


// scalar form used by the CUDA-core kernel
__device__ __forceinline__ void epilogue_elem(const EpiParams& p, int row, int col, float acc) {
  float x = p.alpha * acc + (p.bias ? __ldg(p.bias + col) : 0.f);
  if (p.act == PSRC_ACT_QUICKGELU) {
    if (p.out2) store_elem(p.out2, p.out2_dtype, (long long)row * p.ldo2 + col, x);
    x = quickgelu(x);
  } else if (p.act == PSRC_ACT_QUICKGELU_GRAD) {
    x *= quickgelu_grad(load_elem(p.aux, p.aux_dtype, (long long)row * p.ldaux + col));
  }
  if (p.residual) x += p.residual[(long long)row * p.ldres + col];
  store_elem(p.out, p.out_dtype, (long long)row * p.ldo + col, x);
}

void fill_epi(const psrc_gemm_args& a, EpiParams& e);
bool gemm_tc_supported(const psrc_gemm_args& a);
int gemm_tc1(const psrc_gemm_args& a, cudaStream_t stream);   // single-CTA kernel (gemm_tc1.cu)
bool gemm_pair_supported(const psrc_gemm_args& a);
int gemm_pair(const psrc_gemm_args& a, cudaStream_t stream);  // CTA-pair kernel (gemm_pair.cu)

}  // namespace psrc
