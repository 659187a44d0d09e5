// The only TRAINABLE dense layers on the path: MaPLe's text->vision coupling Linears
// (PromptSRC/trainers/maple.py:145, :155-156, :199-204) and CoCoOp's meta-net
// (PromptSRC/trainers/cocoop.py:139-143, :187).  They act on a handful of rows (n_ctx = 2 prompt
// tokens, or one feature row per image), so they are latency-bound fp32 CUDA-core kernels, batched
// over `G` independent (X, W, b) groups so MaPLe's 9 projections are one launch.
//   forward : Y[g] = X[g] . W[g]^T + b[g]   (optionally ReLU)
//   backward: dW[g] = dY[g]^T . X[g];  db[g] = sum_r dY[g];  dX[g] (+)= dY[g] . W[g]
// plus the CoCoOp per-image logits (cocoop.py:249-252) and the split of the grouped context
// gradient into d ctx / d bias (autograd of cocoop.py:187-188).
#include "common.cuh"

namespace psrc {

// one warp per output element (g, r, n): lanes stride over K (coalesced X and W rows)
__global__ void small_linear_fwd_kernel(const float* __restrict__ X, const float* __restrict__ W,
                                        const float* __restrict__ b, float* __restrict__ Y, int G, int R, int N,
                                        int K, int relu) {
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long total = (long long)G * R * N;
  if (warp >= total) return;
  const int n = (int)(warp % N);
  const int r = (int)((warp / N) % R);
  const int g = (int)(warp / ((long long)N * R));
  const float* x = X + ((long long)g * R + r) * K;
  const float* w = W + ((long long)g * N + n) * K;
  float acc = 0.f;
  for (int k = lane; k < K; k += 32) acc = fmaf(x[k], w[k], acc);
  acc = warp_sum(acc);
  if (lane == 0) {
    acc += b ? b[(long long)g * N + n] : 0.f;
    if (relu) acc = fmaxf(acc, 0.f);
    Y[((long long)g * R + r) * N + n] = acc;
  }
}

// one thread per weight element (g, n, k): dW = sum_r dY[g,r,n] * X[g,r,k]; the k == 0 thread also writes db
__global__ void small_linear_wgrad_kernel(const float* __restrict__ X, const float* __restrict__ dY,
                                          const float* __restrict__ Yrelu, float* __restrict__ dW,
                                          float* __restrict__ db, int G, int R, int N, int K) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)G * N * K;
  if (i >= total) return;
  const int k = (int)(i % K);
  const int n = (int)((i / K) % N);
  const int g = (int)(i / ((long long)K * N));
  float acc = 0.f, accb = 0.f;
  for (int r = 0; r < R; ++r) {
    const long long row = (long long)g * R + r;
    float d = dY[row * N + n];
    if (Yrelu && !(Yrelu[row * N + n] > 0.f)) d = 0.f;
    acc = fmaf(d, X[row * K + k], acc);
    accb += d;
  }
  dW[i] = acc;
  if (k == 0 && db) db[(long long)g * N + n] = accb;
}

// one thread per input element (g, r, k): dX (+)= sum_n dY[g,r,n] * W[g,n,k]  (W reads coalesced over k)
__global__ void small_linear_dgrad_kernel(const float* __restrict__ W, const float* __restrict__ dY,
                                          const float* __restrict__ Yrelu, float* __restrict__ dX, int G, int R,
                                          int N, int K, int accumulate) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)G * R * K;
  if (i >= total) return;
  const int k = (int)(i % K);
  const long long row = i / K;
  const int g = (int)(row / R);
  const float* w = W + (long long)g * N * K + k;
  const float* dy = dY + row * N;
  const float* yr = Yrelu ? Yrelu + row * N : nullptr;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  int n = 0;
  for (; n + 4 <= N; n += 4) {
    float d0 = dy[n], d1 = dy[n + 1], d2 = dy[n + 2], d3 = dy[n + 3];
    if (yr) {
      d0 = yr[n] > 0.f ? d0 : 0.f;
      d1 = yr[n + 1] > 0.f ? d1 : 0.f;
      d2 = yr[n + 2] > 0.f ? d2 : 0.f;
      d3 = yr[n + 3] > 0.f ? d3 : 0.f;
    }
    a0 = fmaf(d0, w[(long long)n * K], a0);
    a1 = fmaf(d1, w[(long long)(n + 1) * K], a1);
    a2 = fmaf(d2, w[(long long)(n + 2) * K], a2);
    a3 = fmaf(d3, w[(long long)(n + 3) * K], a3);
  }
  for (; n < N; ++n) {
    float d = dy[n];
    if (yr && !(yr[n] > 0.f)) d = 0.f;
    a0 = fmaf(d, w[(long long)n * K], a0);
  }
  const float acc = (a0 + a1) + (a2 + a3);
  dX[i] = accumulate ? dX[i] + acc : acc;
}

// CoCoOp logits (cocoop.py:249-252): logits[b,c] = scale * <f_hat[b,:], t_hat[b,c,:]>; one warp per (b, c)
__global__ void group_logits_fwd_kernel(const float* __restrict__ f_hat, const float* __restrict__ t_hat,
                                        float* __restrict__ logits, int B, int C, int E, float scale) {
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (warp >= (long long)B * C) return;
  const int b = (int)(warp / C);
  const float* f = f_hat + (long long)b * E;
  const float* t = t_hat + warp * E;
  float acc = 0.f;
  for (int e = lane; e < E; e += 32) acc = fmaf(f[e], t[e], acc);
  acc = warp_sum(acc);
  if (lane == 0) logits[warp] = scale * acc;
}

// d t_hat[b,c,:] = scale * dlogits[b,c] * f_hat[b,:]   (the image tower is frozen: no d f_hat needed)
__global__ void group_logits_bwd_kernel(const float* __restrict__ f_hat, const float* __restrict__ d_logits,
                                        float* __restrict__ d_that, int B, int C, int E, float scale) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)B * C * E;
  if (i >= total) return;
  const int e = (int)(i % E);
  const long long bc = i / E;
  const int b = (int)(bc / C);
  d_that[i] = scale * d_logits[bc] * f_hat[(long long)b * E + e];
}

// gg [G, n_ctx, D] (context-row gradient summed over the classes of image g) ->
//   d ctx[j,:] = sum_g gg[g,j,:]     d bias[g,:] = sum_j gg[g,j,:]      (ctx_shifted = ctx + bias, cocoop.py:187-188)
__global__ void ctx_bias_grad_kernel(const float* __restrict__ gg, float* __restrict__ dctx,
                                     float* __restrict__ dbias, int G, int n_ctx, int D) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int nc = n_ctx * D, nb = G * D;
  if (i < nc) {
    float acc = 0.f;
    for (int g = 0; g < G; ++g) acc += gg[(long long)g * nc + i];
    dctx[i] = acc;
  } else if (i < nc + nb) {
    const int k = i - nc, g = k / D, d = k % D;
    float acc = 0.f;
    for (int j = 0; j < n_ctx; ++j) acc += gg[((long long)g * n_ctx + j) * D + d];
    dbias[k] = acc;
  }
}

// CoOp CLASS_TOKEN_POSITION middle / front (coop.py:278-293): the prompt of class s is the 'end' layout with its rows
// permuted.  forward: dst[s,p,:] = src[s,perm[s,p],:] (+ add[p,:]);  inverse (the backward): dst[s,perm[s,p],:] = src[s,p,:]
__global__ void permute_rows_kernel(const float4* __restrict__ src, const int32_t* __restrict__ perm,
                                    const float4* __restrict__ add, float4* __restrict__ dst, int n_seq, int L, int D4,
                                    int inverse) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)n_seq * L * D4) return;
  const int d = (int)(i % D4);
  const long long row = i / D4;
  const int p = (int)(row % L);
  const long long s = row / L;
  const int q = perm[row];
  if (inverse) {
    dst[(s * L + q) * D4 + d] = src[i];
  } else {
    float4 v = src[(s * L + q) * D4 + d];
    if (add) {
      const float4 a = add[(long long)p * D4 + d];
      v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
    }
    dst[i] = v;
  }
}

}  // namespace psrc

using namespace psrc;

extern "C" int psrc_permute_rows(const float* src, const int32_t* perm, const float* add, float* dst, int n_seq, int L,
                                 int D, int inverse, psrc_stream_t stream) {
  PSRC_CHECK_ARG(src && perm && dst && src != dst && n_seq > 0 && L > 0 && D > 0 && D % 4 == 0,
                 "psrc_permute_rows: bad arguments");
  const long long n = (long long)n_seq * L * (D / 4);
  permute_rows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, as_stream(stream)>>>(
      reinterpret_cast<const float4*>(src), perm, reinterpret_cast<const float4*>(add),
      reinterpret_cast<float4*>(dst), n_seq, L, D / 4, inverse);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

static inline unsigned blocks_for(long long n, int block) { return (unsigned)((n + block - 1) / block); }

extern "C" int psrc_small_linear_fwd(const float* X, const float* W, const float* b, float* Y, int G, int R, int N,
                                     int K, int relu, psrc_stream_t stream) {
  PSRC_CHECK_ARG(X && W && Y && G > 0 && R > 0 && N > 0 && K > 0, "psrc_small_linear_fwd: bad arguments");
  const long long warps = (long long)G * R * N;
  small_linear_fwd_kernel<<<blocks_for(warps * 32, 256), 256, 0, as_stream(stream)>>>(X, W, b, Y, G, R, N, K, relu);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

extern "C" int psrc_small_linear_bwd(const float* X, const float* W, const float* dY, const float* Y_relu, float* dX,
                                     int accumulate_dX, float* dW, float* db, int G, int R, int N, int K,
                                     psrc_stream_t stream) {
  PSRC_CHECK_ARG(X && W && dY && G > 0 && R > 0 && N > 0 && K > 0, "psrc_small_linear_bwd: bad arguments");
  cudaStream_t s = as_stream(stream);
  if (dW) {
    small_linear_wgrad_kernel<<<blocks_for((long long)G * N * K, 256), 256, 0, s>>>(X, dY, Y_relu, dW, db, G, R, N, K);
    PSRC_CHECK_LAUNCH();
  }
  if (dX) {
    small_linear_dgrad_kernel<<<blocks_for((long long)G * R * K, 128), 128, 0, s>>>(W, dY, Y_relu, dX, G, R, N, K,
                                                                                   accumulate_dX);
    PSRC_CHECK_LAUNCH();
  }
  return PSRC_OK;
}

extern "C" int psrc_group_logits_fwd(const float* f_hat, const float* t_hat, float* logits, int B, int C, int E,
                                     float scale, psrc_stream_t stream) {
  PSRC_CHECK_ARG(f_hat && t_hat && logits && B > 0 && C > 0 && E > 0, "psrc_group_logits_fwd: bad arguments");
  group_logits_fwd_kernel<<<blocks_for((long long)B * C * 32, 256), 256, 0, as_stream(stream)>>>(f_hat, t_hat, logits,
                                                                                                B, C, E, scale);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

extern "C" int psrc_group_logits_bwd(const float* f_hat, const float* d_logits, float* d_that, int B, int C, int E,
                                     float scale, psrc_stream_t stream) {
  PSRC_CHECK_ARG(f_hat && d_logits && d_that && B > 0 && C > 0 && E > 0, "psrc_group_logits_bwd: bad arguments");
  group_logits_bwd_kernel<<<blocks_for((long long)B * C * E, 256), 256, 0, as_stream(stream)>>>(f_hat, d_logits,
                                                                                               d_that, B, C, E, scale);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

extern "C" int psrc_ctx_bias_grad(const float* gg, float* dctx, float* dbias, int G, int n_ctx, int D,
                                  psrc_stream_t stream) {
  PSRC_CHECK_ARG(gg && dctx && dbias && G > 0 && n_ctx > 0 && D > 0, "psrc_ctx_bias_grad: bad arguments");
  ctx_bias_grad_kernel<<<blocks_for((long long)(n_ctx + G) * D, 256), 256, 0, as_stream(stream)>>>(gg, dctx, dbias, G,
                                                                                                  n_ctx, D);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}
