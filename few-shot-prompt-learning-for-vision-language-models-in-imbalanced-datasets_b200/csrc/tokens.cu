// Token-stream assembly kernels (all HBM-bound, vectorised, coalesced):
//   im2col for the patch-embedding GEMM, vision token assembly fused with ln_pre
//   (PromptSRC/clip/model.py:401-420), text prompt assembly + positional embedding
//   (trainers/promptsrc.py:144-168, :60; coop.py:259-296 "end"; cocoop.py:173-198),
//   deep-prompt replacement and its gradient (model.py:232-252).
#include <cooperative_groups.h>
#include <type_traits>
#include "common.cuh"

namespace psrc {

constexpr int kPgSplit = 8;   // CTAs per cluster splitting the sequences of one prompt-gradient reduction

// image [B,3,R,R] fp32 -> patches [B*G*G, 3*P*P]; two elements per thread (P is even)
template <typename T>
__global__ void im2col_kernel(const float* __restrict__ img, T* __restrict__ out, int B, int R, int P,
                              long long total2) {
  const int G = R / P, PP = P * P, K = 3 * PP;
  long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; e < total2; e += stride) {
    const long long el = e * 2;
    const int col = (int)(el % K);
    const long long row = el / K;
    const int gx = (int)(row % G), gy = (int)((row / G) % G), b = (int)(row / ((long long)G * G));
    const int c = col / PP, rem = col - c * PP, ky = rem / P, kx = rem - ky * P;
    const float2 v = *reinterpret_cast<const float2*>(
        img + (((long long)(b * 3 + c) * R + gy * P + ky) * R + gx * P + kx));
    if constexpr (std::is_same<T, __half>::value) {
      *reinterpret_cast<__half2*>(out + el) = __floats2half2_rn(v.x, v.y);
    } else if constexpr (sizeof(T) == 2) {
      *reinterpret_cast<__nv_bfloat162*>(out + el) = __floats2bfloat162_rn(v.x, v.y);
    } else {
      *reinterpret_cast<float2*>(out + el) = v;
    }
  }
}

// P % 8 == 0 (ViT-B/16, ViT-B/32), 16-bit output: eight elements per thread -- two 16-byte loads, one 16-byte store
template <typename T>
__global__ void im2col8_kernel(const float* __restrict__ img, T* __restrict__ out, int B, int R, int P,
                               long long total8) {
  const int G = R / P, PP = P * P, K = 3 * PP;
  long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; e < total8; e += stride) {
    const long long el = e * 8;
    const int col = (int)(el % K);
    const long long row = el / K;
    const int gx = (int)(row % G), gy = (int)((row / G) % G), b = (int)(row / ((long long)G * G));
    const int c = col / PP, rem = col - c * PP, ky = rem / P, kx = rem - ky * P;
    const float4* src = reinterpret_cast<const float4*>(
        img + (((long long)(b * 3 + c) * R + gy * P + ky) * R + gx * P + kx));
    const float4 v0 = __ldg(src), v1 = __ldg(src + 1);
    uint4 u;
    if constexpr (std::is_same<T, __half>::value) {
      __half2* h = reinterpret_cast<__half2*>(&u);
      h[0] = __floats2half2_rn(v0.x, v0.y);
      h[1] = __floats2half2_rn(v0.z, v0.w);
      h[2] = __floats2half2_rn(v1.x, v1.y);
      h[3] = __floats2half2_rn(v1.z, v1.w);
    } else {
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
      h[0] = __floats2bfloat162_rn(v0.x, v0.y);
      h[1] = __floats2bfloat162_rn(v0.z, v0.w);
      h[2] = __floats2bfloat162_rn(v1.x, v1.y);
      h[3] = __floats2bfloat162_rn(v1.z, v1.w);
    }
    *reinterpret_cast<uint4*>(out + el) = u;
  }
}

// One warp per token row: build the pre-LN row, LayerNorm it, write the fp32 residual stream.
template <typename T, int NV>
__global__ void __launch_bounds__(256)
assemble_vision_kernel(const T* __restrict__ pe, const float* __restrict__ cls, const float* __restrict__ pos,
                       const float* __restrict__ vpt, const float* __restrict__ g, const float* __restrict__ bta,
                       float* __restrict__ x, int B, int n_patches, int n_ctx, int D, float eps) {
  const int L = 1 + n_patches + n_ctx;
  const int lane = threadIdx.x & 31;
  const long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= (long long)B * L) return;
  const int b = (int)(r / L), l = (int)(r - (long long)b * L);
  float4 v[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (lane + 32 * i) * 4;
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < D) {
      if (l == 0) {
        float4 a = *reinterpret_cast<const float4*>(cls + c), p = *reinterpret_cast<const float4*>(pos + c);
        t = make_float4(a.x + p.x, a.y + p.y, a.z + p.z, a.w + p.w);
      } else if (l <= n_patches) {
        const T* src = pe + ((long long)b * n_patches + (l - 1)) * D + c;
        float4 a;
        if constexpr (std::is_same<T, __half>::value) {
          uint2 u = *reinterpret_cast<const uint2*>(src);
          float2 lo = __half22float2(*reinterpret_cast<__half2*>(&u.x));
          float2 hi = __half22float2(*reinterpret_cast<__half2*>(&u.y));
          a = make_float4(lo.x, lo.y, hi.x, hi.y);
        } else if constexpr (sizeof(T) == 2) {
          uint2 u = *reinterpret_cast<const uint2*>(src);
          float2 lo = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u.x));
          float2 hi = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u.y));
          a = make_float4(lo.x, lo.y, hi.x, hi.y);
        } else {
          a = *reinterpret_cast<const float4*>(src);
        }
        float4 p = *reinterpret_cast<const float4*>(pos + (long long)l * D + c);
        t = make_float4(a.x + p.x, a.y + p.y, a.z + p.z, a.w + p.w);
      } else {
        float4 a = *reinterpret_cast<const float4*>(vpt + (long long)(l - 1 - n_patches) * D + c);
        t = make_float4(fp16_round(a.x), fp16_round(a.y), fp16_round(a.z), fp16_round(a.w));
      }
    }
    v[i] = t;
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) s += v[i].x + v[i].y + v[i].z + v[i].w;
  const float mean = warp_sum(s) / (float)D;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i)
    if ((lane + 32 * i) * 4 < D) {
      float a = v[i].x - mean, b2 = v[i].y - mean, c2 = v[i].z - mean, d2 = v[i].w - mean;
      q += a * a + b2 * b2 + c2 * c2 + d2 * d2;
    }
  const float rstd = rsqrtf(warp_sum(q) / (float)D + eps);
  float* xr = x + r * D;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (lane + 32 * i) * 4;
    if (c < D) {
      float4 gg = *reinterpret_cast<const float4*>(g + c), bb = *reinterpret_cast<const float4*>(bta + c);
      float4 o;
      o.x = (v[i].x - mean) * rstd * gg.x + bb.x;
      o.y = (v[i].y - mean) * rstd * gg.y + bb.y;
      o.z = (v[i].z - mean) * rstd * gg.z + bb.z;
      o.w = (v[i].w - mean) * rstd * gg.w + bb.w;
      *reinterpret_cast<float4*>(xr + c) = o;
    }
  }
}

// x[(g*C + c), l, :] = src + pos[l];  one float4 per thread
__global__ void assemble_text_kernel(const float* __restrict__ prefix, const float* __restrict__ suffix,
                                     int suffix_len, const float* __restrict__ ctx, int ctx_per_class,
                                     const float* __restrict__ ctx_bias, const float* __restrict__ pos,
                                     float* __restrict__ x, int n_groups, int C, int L, int n_ctx, int D,
                                     long long total4) {
  const int D4 = D / 4;
  long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; e < total4; e += stride) {
    const int d = (int)(e % D4) * 4;
    const long long row = e / D4;
    const int l = (int)(row % L);
    const long long seq = row / L;
    const int c = (int)(seq % C), grp = (int)(seq / C);
    float4 a;
    if (l == 0) {
      a = *reinterpret_cast<const float4*>(prefix + (long long)c * D + d);
    } else if (l <= n_ctx) {
      const float* src = ctx_per_class ? ctx + ((long long)c * n_ctx + (l - 1)) * D : ctx + (long long)(l - 1) * D;
      a = *reinterpret_cast<const float4*>(src + d);
      if (ctx_bias) {
        float4 bb = *reinterpret_cast<const float4*>(ctx_bias + (long long)grp * D + d);
        a.x += bb.x; a.y += bb.y; a.z += bb.z; a.w += bb.w;
      }
    } else {
      a = *reinterpret_cast<const float4*>(suffix + ((long long)c * suffix_len + (l - 1 - n_ctx)) * D + d);
    }
    float4 p = *reinterpret_cast<const float4*>(pos + (long long)l * D + d);
    *reinterpret_cast<float4*>(x + row * D + d) = make_float4(a.x + p.x, a.y + p.y, a.z + p.z, a.w + p.w);
  }
}

__global__ void insert_prompt_kernel(float* __restrict__ x, const float* __restrict__ prompt, int n_seq, int L,
                                     int start, int n_ctx, int D, int round_fp16, long long total4) {
  const int D4 = D / 4;
  long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; e < total4; e += stride) {
    const int d = (int)(e % D4) * 4;
    const long long t = e / D4;
    const int j = (int)(t % n_ctx);
    const long long seq = t / n_ctx;
    float4 a = *reinterpret_cast<const float4*>(prompt + (long long)j * D + d);
    if (round_fp16) a = make_float4(fp16_round(a.x), fp16_round(a.y), fp16_round(a.z), fp16_round(a.w));
    *reinterpret_cast<float4*>(x + (seq * L + start + j) * D + d) = a;
  }
}

// grad[j, d4] (+)= sum over sequences of dx[seq, start+j, d4]; block (64, 8): x over d4, y over sequences.
// Deterministic (fixed summation order), then optionally zeroes the rows it consumed.
__global__ void __cluster_dims__(1, 1, kPgSplit) __launch_bounds__(512)
prompt_grad_kernel(float* __restrict__ dx, __nv_bfloat16* __restrict__ dx_bf16, float* __restrict__ dx_f32copy,
                   float* __restrict__ grad, int n_seq, int L, int start, int n_ctx, int D, int zero_rows,
                   int accumulate, int round_fp16) {
  // grid (D/256, n_ctx, kPgSplit): the kPgSplit CTAs of a cluster split the sequences; partial sums meet in the
  // rank-0 CTA through distributed shared memory in a fixed order (deterministic, no atomics, no scratch)
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ float4 red[8][64];
  __shared__ float4 part[64];
  const int d4 = blockIdx.x * 64 + threadIdx.x;
  const int j = blockIdx.y;
  const int z = blockIdx.z;
  const int D4 = D / 4;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (d4 < D4) {
    // four independent loads in flight per thread before any of the zeroing stores (which the compiler must keep
    // ordered behind loads through the same base pointer)
    constexpr int kU = 4;
    const int step = 8 * kPgSplit;
    for (int seq0 = z * 8 + threadIdx.y; seq0 < n_seq; seq0 += kU * step) {
      float4 v[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int seq = seq0 + u * step;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (seq < n_seq) v[u] = *(reinterpret_cast<const float4*>(dx + ((long long)seq * L + start + j) * D) + d4);
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int seq = seq0 + u * step;
        if (seq >= n_seq) break;
        float4 w = v[u];
        if (round_fp16) w = make_float4(fp16_round(w.x), fp16_round(w.y), fp16_round(w.z), fp16_round(w.w));
        s.x += w.x; s.y += w.y; s.z += w.z; s.w += w.w;
        if (zero_rows) {
          const long long off = ((long long)seq * L + start + j) * D + (long long)d4 * 4;
          *reinterpret_cast<float4*>(dx + off) = make_float4(0.f, 0.f, 0.f, 0.f);
          if (dx_bf16) *reinterpret_cast<uint2*>(dx_bf16 + off) = make_uint2(0u, 0u);
          if (dx_f32copy) *reinterpret_cast<float4*>(dx_f32copy + off) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
    }
  }
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0) {
    float4 t = red[0][threadIdx.x];
#pragma unroll
    for (int y = 1; y < 8; ++y) {
      float4 v = red[y][threadIdx.x];
      t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
    }
    part[threadIdx.x] = t;
  }
  cluster.sync();
  if (z == 0 && threadIdx.y == 0 && d4 < D4) {
    float4 t = part[threadIdx.x];
    for (int rk = 1; rk < kPgSplit; ++rk) {
      const float4 v = cluster.map_shared_rank(part, rk)[threadIdx.x];
      t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
    }
    float4* g = reinterpret_cast<float4*>(grad + (long long)j * D) + d4;
    if (accumulate) {
      float4 o = *g;
      t.x += o.x; t.y += o.y; t.z += o.z; t.w += o.w;
    }
    *g = t;
  }
  cluster.sync();   // peers' shared memory must stay alive until rank 0 has read it
}

static int grid_for(long long n, int block) {
  long long g = (n + block - 1) / block;
  if (g > 148LL * 16) g = 148LL * 16;
  return (int)(g < 1 ? 1 : g);
}

}  // namespace psrc

using namespace psrc;

extern "C" int psrc_im2col_patches(const float* image, void* patches, int out_dtype, int B, int R, int P,
                                   psrc_stream_t stream) {
  PSRC_CHECK_ARG(image && patches && B >= 0 && R > 0 && P > 0 && R % P == 0 && P % 2 == 0,
                 "psrc_im2col_patches: bad arguments (R=%d P=%d)", R, P);
  if (B == 0) return PSRC_OK;
  const int G = R / P;
  const long long total2 = (long long)B * G * G * 3 * P * P / 2;
  const bool al16 = ((reinterpret_cast<uintptr_t>(image) | reinterpret_cast<uintptr_t>(patches)) & 15) == 0;
  if (out_dtype == PSRC_F16 && P % 8 == 0 && R % 4 == 0 && al16)
    im2col8_kernel<__half><<<grid_for(total2 / 4, 256), 256, 0, as_stream(stream)>>>(
        image, reinterpret_cast<__half*>(patches), B, R, P, total2 / 4);
  else if (out_dtype == PSRC_F16)
    im2col_kernel<__half><<<grid_for(total2, 256), 256, 0, as_stream(stream)>>>(
        image, reinterpret_cast<__half*>(patches), B, R, P, total2);
  else if (out_dtype == PSRC_BF16 && P % 8 == 0 && R % 4 == 0 && al16)
    im2col8_kernel<__nv_bfloat16><<<grid_for(total2 / 4, 256), 256, 0, as_stream(stream)>>>(
        image, reinterpret_cast<__nv_bfloat16*>(patches), B, R, P, total2 / 4);
  else if (out_dtype == PSRC_BF16)
    im2col_kernel<__nv_bfloat16><<<grid_for(total2, 256), 256, 0, as_stream(stream)>>>(
        image, reinterpret_cast<__nv_bfloat16*>(patches), B, R, P, total2);
  else
    im2col_kernel<float><<<grid_for(total2, 256), 256, 0, as_stream(stream)>>>(
        image, reinterpret_cast<float*>(patches), B, R, P, total2);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

template <typename T>
static int launch_assemble_vision(const void* pe, const float* cls, const float* pos, const float* vpt,
                                  const float* g, const float* b, float* x, int B, int n_patches, int n_ctx,
                                  int D, float eps, cudaStream_t s) {
  const long long rows = (long long)B * (1 + n_patches + n_ctx);
  const int grid = (int)((rows + 7) / 8);
  const T* pet = reinterpret_cast<const T*>(pe);
#define AV_CASE(NV)                                                                                          \
  case NV:                                                                                                   \
    assemble_vision_kernel<T, NV><<<grid, 256, 0, s>>>(pet, cls, pos, vpt, g, b, x, B, n_patches, n_ctx, D, eps); \
    break;
  switch ((D + 127) / 128) {
    AV_CASE(1) AV_CASE(2) AV_CASE(3) AV_CASE(4) AV_CASE(5) AV_CASE(6) AV_CASE(7) AV_CASE(8)
    default:
      set_error("assemble_vision_tokens: D=%d not supported", D);
      return PSRC_ERR_UNSUPPORTED;
  }
#undef AV_CASE
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

extern "C" int psrc_assemble_vision_tokens(const void* patch_embed, int pe_dtype, const float* cls,
                                           const float* pos, const float* vpt, const float* ln_g,
                                           const float* ln_b, float* x, int B, int n_patches, int n_ctx, int D,
                                           float eps, psrc_stream_t stream) {
  PSRC_CHECK_ARG(patch_embed && cls && pos && ln_g && ln_b && x, "psrc_assemble_vision_tokens: null pointer");
  PSRC_CHECK_ARG(n_ctx == 0 || vpt, "psrc_assemble_vision_tokens: n_ctx > 0 needs vpt");
  PSRC_CHECK_ARG(D > 0 && D % 4 == 0 && D <= 1024 && n_patches > 0 && n_ctx >= 0 && B >= 0,
                 "psrc_assemble_vision_tokens: bad shape");
  if (B == 0) return PSRC_OK;
  if (pe_dtype == PSRC_F16)
    return launch_assemble_vision<__half>(patch_embed, cls, pos, vpt, ln_g, ln_b, x, B, n_patches, n_ctx, D, eps,
                                          as_stream(stream));
  if (pe_dtype == PSRC_BF16)
    return launch_assemble_vision<__nv_bfloat16>(patch_embed, cls, pos, vpt, ln_g, ln_b, x, B, n_patches, n_ctx, D,
                                                 eps, as_stream(stream));
  return launch_assemble_vision<float>(patch_embed, cls, pos, vpt, ln_g, ln_b, x, B, n_patches, n_ctx, D, eps,
                                       as_stream(stream));
}

extern "C" int psrc_assemble_text_prompts(const float* prefix, const float* suffix, int suffix_len,
                                          const float* ctx, int ctx_per_class, const float* ctx_bias,
                                          int n_groups, const float* pos, float* x, int C, int L, int n_ctx,
                                          int D, psrc_stream_t stream) {
  PSRC_CHECK_ARG(prefix && suffix && ctx && pos && x, "psrc_assemble_text_prompts: null pointer");
  PSRC_CHECK_ARG(D > 0 && D % 4 == 0 && C >= 0 && n_groups >= 1 && L >= 1 && n_ctx >= 0 &&
                     L - 1 - n_ctx <= suffix_len,
                 "psrc_assemble_text_prompts: bad shape (L=%d n_ctx=%d suffix_len=%d)", L, n_ctx, suffix_len);
  const long long total4 = (long long)n_groups * C * L * (D / 4);
  if (total4 == 0) return PSRC_OK;
  assemble_text_kernel<<<grid_for(total4, 256), 256, 0, as_stream(stream)>>>(
      prefix, suffix, suffix_len, ctx, ctx_per_class, ctx_bias, pos, x, n_groups, C, L, n_ctx, D, total4);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

extern "C" int psrc_insert_prompt(float* x, const float* prompt, int n_seq, int L, int start, int n_ctx, int D,
                                  int round_fp16, psrc_stream_t stream) {
  PSRC_CHECK_ARG(x && prompt && D > 0 && D % 4 == 0 && start >= 0 && n_ctx >= 0 && start + n_ctx <= L,
                 "psrc_insert_prompt: bad arguments");
  const long long total4 = (long long)n_seq * n_ctx * (D / 4);
  if (total4 <= 0) return PSRC_OK;
  insert_prompt_kernel<<<grid_for(total4, 256), 256, 0, as_stream(stream)>>>(x, prompt, n_seq, L, start, n_ctx, D,
                                                                            round_fp16, total4);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

extern "C" int psrc_prompt_grad(float* dx, void* dx_act, int act_dtype, float* grad, int n_seq, int L, int start,
                                int n_ctx, int D, int zero_rows, int accumulate, int round_fp16,
                                psrc_stream_t stream) {
  PSRC_CHECK_ARG(dx && grad && D > 0 && D % 4 == 0 && start >= 0 && n_ctx >= 0 && start + n_ctx <= L && n_seq >= 0,
                 "psrc_prompt_grad: bad arguments");
  if (n_ctx == 0) return PSRC_OK;
  dim3 grid((D / 4 + 63) / 64, n_ctx, kPgSplit), block(64, 8);
  __nv_bfloat16* a16 = (dx_act && act_dtype == PSRC_BF16) ? reinterpret_cast<__nv_bfloat16*>(dx_act) : nullptr;
  float* a32 = (dx_act && act_dtype == PSRC_F32 && dx_act != (void*)dx) ? reinterpret_cast<float*>(dx_act) : nullptr;
  prompt_grad_kernel<<<grid, block, 0, as_stream(stream)>>>(dx, a16, a32, grad, n_seq, L, start, n_ctx, D,
                                                            zero_rows, accumulate, round_fp16);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}
