// CUDA-core GEMMs: (1) the fp32 validation-mode kernel behind psrc_gemm (same fused epilogue as
// the tcgen05 kernel, any dtype/shape), (2) a strided fp32 GEMM for the tiny [B,C] cosine-logit
// contractions and their two gradients (PromptSRC/trainers/promptsrc.py:193,201).
#include "gemm_epilogue.cuh"

namespace psrc {

constexpr int kTM = 64, kTN = 64, kTK = 16;

// out = epi(A[M,K] . W[N,K]^T); 256 threads, each a 4x4 micro-tile of a 64x64 block tile
template <typename T>
__global__ void __launch_bounds__(256)
gemm_simt_kernel(const T* __restrict__ A, long long lda, const T* __restrict__ W, long long ldw, int K,
                 EpiParams epi) {
  __shared__ float sA[kTK][kTM + 4];
  __shared__ float sW[kTK][kTN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * kTM, n0 = blockIdx.x * kTN;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  // each thread loads 4 elements of A and 4 of W per k-step: row = tid/4 (0..63), k = (tid%4)*4..+3
  const int lr = tid >> 2, lk = (tid & 3) * 4;
  for (int k0 = 0; k0 < K; k0 += kTK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int k = k0 + lk + i;
      int m = m0 + lr, n = n0 + lr;
      sA[lk + i][lr] = (m < epi.M && k < K) ? to_f32(A[(long long)m * lda + k]) : 0.f;
      sW[lk + i][lr] = (n < epi.N && k < K) ? to_f32(W[(long long)n * ldw + k]) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kTK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = sA[k][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = sW[k][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int m = m0 + ty * 4 + i;
    if (m >= epi.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n < epi.N) epilogue_elem(epi, m, n, acc[i][j]);
    }
  }
}

// C[M,N] = alpha * A . B^T + beta * C with arbitrary element strides.  T x T block tile (T = 64: 4x4 per thread, T = 32:
// 2x2 -- the logit contractions are 128 x 1000 x 512 and smaller, so the small tile is what fills the SMs), 16-deep
// k-tiles, the next k-tile's global loads in flight (registers) while the current one is multiplied.
template <int T>
__global__ void __launch_bounds__(256)
gemm_f32_strided_kernel(int M, int N, int K, const float* __restrict__ A, long long a_sm, long long a_sk,
                        const float* __restrict__ B, long long b_sn, long long b_sk, float* __restrict__ C,
                        long long ldc, float alpha, float beta) {
  constexpr int MT = T / 16;             // micro-tile edge
  constexpr int NL = T * kTK / 256;      // elements of each operand tile loaded per thread
  __shared__ float sA[kTK][T + 4];
  __shared__ float sB[kTK][T + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * T, n0 = blockIdx.x * T;
  float acc[MT][MT];
#pragma unroll
  for (int i = 0; i < MT; ++i)
#pragma unroll
    for (int j = 0; j < MT; ++j) acc[i][j] = 0.f;
  // Two load mappings so that the contiguous global dimension maps to consecutive threads.
  const bool a_kfast = (a_sk == 1), b_kfast = (b_sk == 1);
  float ra[NL], rb[NL];
  auto fetch = [&](int k0) {
#pragma unroll
    for (int i = 0; i < NL; ++i) {
      const int e = tid + i * 256;  // over a T x 16 tile
      int r, k;
      if (a_kfast) { r = e >> 4; k = e & 15; } else { r = e % T; k = e / T; }
      const int m = m0 + r;
      int kk = k0 + k;
      ra[i] = (m < M && kk < K) ? A[(long long)m * a_sm + (long long)kk * a_sk] : 0.f;
      if (b_kfast) { r = e >> 4; k = e & 15; } else { r = e % T; k = e / T; }
      const int n = n0 + r;
      kk = k0 + k;
      rb[i] = (n < N && kk < K) ? B[(long long)n * b_sn + (long long)kk * b_sk] : 0.f;
    }
  };
  fetch(0);
  for (int k0 = 0; k0 < K; k0 += kTK) {
#pragma unroll
    for (int i = 0; i < NL; ++i) {
      const int e = tid + i * 256;
      int r, k;
      if (a_kfast) { r = e >> 4; k = e & 15; } else { r = e % T; k = e / T; }
      sA[k][r] = ra[i];
      if (b_kfast) { r = e >> 4; k = e & 15; } else { r = e % T; k = e / T; }
      sB[k][r] = rb[i];
    }
    __syncthreads();
    if (k0 + kTK < K) fetch(k0 + kTK);
#pragma unroll
    for (int k = 0; k < kTK; ++k) {
      float a[MT], b[MT];
#pragma unroll
      for (int i = 0; i < MT; ++i) a[i] = sA[k][ty * MT + i];
#pragma unroll
      for (int j = 0; j < MT; ++j) b[j] = sB[k][tx * MT + j];
#pragma unroll
      for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < MT; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < MT; ++i) {
    const int m = m0 + ty * MT + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < MT; ++j) {
      const int n = n0 + tx * MT + j;
      if (n >= N) continue;
      float* c = C + (long long)m * ldc + n;
      float v = alpha * acc[i][j];
      if (beta != 0.f) v += beta * (*c);
      *c = v;
    }
  }
}

static int gemm_simt(const psrc_gemm_args& a, cudaStream_t stream) {
  EpiParams epi;
  fill_epi(a, epi);
  dim3 grid((a.N + kTN - 1) / kTN, (a.M + kTM - 1) / kTM);
  if (a.ab_dtype == PSRC_F16) {
    gemm_simt_kernel<__half><<<grid, 256, 0, stream>>>(
        reinterpret_cast<const __half*>(a.A), a.lda, reinterpret_cast<const __half*>(a.W), a.ldw, a.K, epi);
  } else if (a.ab_dtype == PSRC_BF16) {
    gemm_simt_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(
        reinterpret_cast<const __nv_bfloat16*>(a.A), a.lda, reinterpret_cast<const __nv_bfloat16*>(a.W), a.ldw,
        a.K, epi);
  } else {
    gemm_simt_kernel<float><<<grid, 256, 0, stream>>>(reinterpret_cast<const float*>(a.A), a.lda,
                                                      reinterpret_cast<const float*>(a.W), a.ldw, a.K, epi);
  }
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

}  // namespace psrc

using namespace psrc;

extern "C" int psrc_gemm(const psrc_gemm_args* args, psrc_stream_t stream) {
  PSRC_CHECK_ARG(args != nullptr, "psrc_gemm: null args");
  const psrc_gemm_args& a = *args;
  PSRC_CHECK_ARG(a.M >= 0 && a.N > 0 && a.K > 0, "psrc_gemm: bad shape M=%d N=%d K=%d", a.M, a.N, a.K);
  if (a.M == 0) return PSRC_OK;   // an empty operand has no storage (NULL): nothing to do, not an error
  PSRC_CHECK_ARG(a.A && a.W && a.out, "psrc_gemm: null A/W/out");
  PSRC_CHECK_ARG(a.ab_dtype >= PSRC_F32 && a.ab_dtype <= PSRC_F16, "psrc_gemm: bad ab_dtype %d", a.ab_dtype);
  PSRC_CHECK_ARG(a.out_dtype >= PSRC_F32 && a.out_dtype <= PSRC_F16, "psrc_gemm: bad out_dtype");
  PSRC_CHECK_ARG(a.act >= PSRC_ACT_NONE && a.act <= PSRC_ACT_QUICKGELU_GRAD, "psrc_gemm: bad act %d", a.act);
  PSRC_CHECK_ARG(a.act != PSRC_ACT_QUICKGELU_GRAD || a.aux != nullptr, "psrc_gemm: QUICKGELU_GRAD needs aux");
  PSRC_CHECK_ARG(a.lda >= a.K && a.ldw >= a.K && a.ldo >= a.N, "psrc_gemm: leading dimension too small");
  if (a.M == 0) return PSRC_OK;
  int backend = a.backend;
  if (backend == PSRC_GEMM_AUTO) backend = gemm_tc_supported(a) ? PSRC_GEMM_TCGEN05 : PSRC_GEMM_SIMT;
  if (backend == PSRC_GEMM_TCGEN05 || backend == PSRC_GEMM_TCGEN05_1CTA) {
    PSRC_CHECK_ARG(gemm_tc_supported(a),
                   "psrc_gemm: tcgen05 backend needs bf16 / fp16 A/W, K%%8==0, lda/ldw%%8==0, 16B-aligned pointers");
    if (backend == PSRC_GEMM_TCGEN05 && gemm_pair_supported(a)) return gemm_pair(a, as_stream(stream));
    return gemm_tc1(a, as_stream(stream));
  }
  return gemm_simt(a, as_stream(stream));
}

extern "C" int psrc_gemm_f32_strided(int M, int N, int K, const float* A, int64_t a_sm, int64_t a_sk,
                                     const float* B, int64_t b_sn, int64_t b_sk, float* C, int64_t ldc,
                                     float alpha, float beta, psrc_stream_t stream) {
  PSRC_CHECK_ARG(M >= 0 && N >= 0 && K > 0 && A && B && C, "psrc_gemm_f32_strided: bad arguments");
  if (M == 0 || N == 0) return PSRC_OK;
  // 64 x 64 tiles only when they alone give every SM a block
  if ((long long)((N + 63) / 64) * ((M + 63) / 64) >= 148) {
    dim3 grid((N + 63) / 64, (M + 63) / 64);
    gemm_f32_strided_kernel<64><<<grid, 256, 0, as_stream(stream)>>>(M, N, K, A, a_sm, a_sk, B, b_sn, b_sk, C, ldc,
                                                                     alpha, beta);
  } else {
    dim3 grid((N + 31) / 32, (M + 31) / 32);
    gemm_f32_strided_kernel<32><<<grid, 256, 0, as_stream(stream)>>>(M, N, K, A, a_sm, a_sk, B, b_sn, b_sk, C, ldc,
                                                                     alpha, beta);
  }
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}
