// On-device input pipeline: uint8 HWC images -> CLIP-normalised fp32 [B,3,R,R].
//
// Replaces, for a whole batch in one launch, the per-image PIL / torchvision chain that
// Dassl's `build_transform` composes for the PromptSRC configs
// (Dassl.pytorch/dassl/data/transforms/transforms.py:176-326; yaml INPUT.TRANSFORMS =
// ["random_resized_crop", "random_flip", "normalize"], INTERPOLATION bicubic):
//   train: RandomResizedCrop(size, scale, interpolation) -> RandomHorizontalFlip -> ToTensor -> Normalize
//   test : Resize(max(size)) -> CenterCrop(size) -> ToTensor -> Normalize   (:329-357)
// The random draws stay on the host (psrc_b200/input_pipeline.py makes the very torch RNG calls
// torchvision makes, so the parameter / index / label stream is bit-identical); this kernel does the
// arithmetic: antialiased separable resampling (PIL's `precompute_coeffs` geometry: support scaled by the
// down-scaling factor, window clamped to the sampled image, weights normalised), optional rounding to
// uint8 like the PIL image the reference hands to ToTensor, horizontal flip, /255, (v - mean) / std.
//
// One block = one image x kRows output rows, one thread per output column.  The column weights of the
// block (<= max_taps per column) are built once in shared memory, tap-major so a warp reads consecutive
// words; row weights are per-row scalars recomputed by every thread (<= max_taps evaluations per row).
// HBM-bound in principle (B*3*R*R*4 bytes out), in practice bound by the byte gathers through L1.
#include "common.cuh"

namespace psrc {

constexpr int kPreRows = 8;

__device__ __forceinline__ float resample_filter(float x, int bicubic) {
  x = fabsf(x);
  if (!bicubic) return x < 1.f ? 1.f - x : 0.f;            // PIL bilinear_filter
  const float a = -0.5f;                                    // PIL bicubic_filter (Keys, a = -0.5)
  if (x < 1.f) return ((a + 2.f) * x - (a + 3.f)) * x * x + 1.f;
  if (x < 2.f) return (((x - 5.f) * x + 8.f) * x - 4.f) * a;
  return 0.f;
}

// window of source indices [lo, lo + n) and the filter argument scale for output index `o`
__device__ __forceinline__ void resample_window(int o, float scale, float off, int size, float support_base,
                                                int& lo, int& n, float& center, float& inv) {
  const float fs = fmaxf(scale, 1.f);
  const float support = support_base * fs;
  center = ((float)o + off + 0.5f) * scale;
  lo = max((int)(center - support + 0.5f), 0);
  const int hi = min((int)(center + support + 0.5f), size);
  n = max(hi - lo, 0);
  inv = 1.f / fs;
}

__global__ void __launch_bounds__(256)
crop_resize_normalize_kernel(const uint8_t* __restrict__ src, long long image_stride, int row_stride,
                             const float* __restrict__ geom_f, const int* __restrict__ geom_i,
                             float* __restrict__ out, int R, int bicubic, int quantize, int max_taps,
                             float m0, float m1, float m2, float s0, float s1, float s2) {
  extern __shared__ float wx[];                     // [max_taps][R] tap-major column weights (normalised)
  __shared__ int xlo_s[256], xn_s[256];
  const int b = blockIdx.y, y0 = blockIdx.x * kPreRows, x = threadIdx.x;
  const float sx = geom_f[b * 4 + 0], sy = geom_f[b * 4 + 1], ox = geom_f[b * 4 + 2], oy = geom_f[b * 4 + 3];
  const int left = geom_i[b * 5 + 0], top = geom_i[b * 5 + 1], w = geom_i[b * 5 + 2], h = geom_i[b * 5 + 3];
  const int flip = geom_i[b * 5 + 4];
  const float support_base = bicubic ? 2.f : 1.f;
  if (x < R) {
    int lo, n;
    float c, inv;
    resample_window(x, sx, ox, w, support_base, lo, n, c, inv);
    n = min(n, max_taps);
    float tot = 0.f;
    for (int k = 0; k < n; ++k) {
      const float wv = resample_filter(((float)(k + lo) - c + 0.5f) * inv, bicubic);
      wx[k * R + x] = wv;
      tot += wv;
    }
    const float r = tot != 0.f ? 1.f / tot : 0.f;
    for (int k = 0; k < n; ++k) wx[k * R + x] *= r;
    xlo_s[x] = lo;
    xn_s[x] = n;
  }
  __syncthreads();
  if (x >= R) return;
  const int xlo = xlo_s[x], xn = xn_s[x];
  const uint8_t* img = src + (long long)b * image_stride + (long long)top * row_stride + (long long)left * 3;
  const int xo = flip ? (R - 1 - x) : x;             // RandomHorizontalFlip acts on the resized crop
  const long long plane = (long long)R * R;
  float* ob = out + (long long)b * 3 * plane;
  for (int yy = 0; yy < kPreRows; ++yy) {
    const int y = y0 + yy;
    if (y >= R) break;
    int ylo, yn;
    float cy, invy;
    resample_window(y, sy, oy, h, support_base, ylo, yn, cy, invy);
    yn = min(yn, max_taps);
    float toty = 0.f;
    for (int j = 0; j < yn; ++j) toty += resample_filter(((float)(j + ylo) - cy + 0.5f) * invy, bicubic);
    const float ry = toty != 0.f ? 1.f / toty : 0.f;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f;
    for (int j = 0; j < yn; ++j) {
      const float wy = resample_filter(((float)(j + ylo) - cy + 0.5f) * invy, bicubic) * ry;
      const uint8_t* row = img + (long long)(ylo + j) * row_stride + (long long)xlo * 3;
      float r0 = 0.f, r1 = 0.f, r2 = 0.f;
      for (int k = 0; k < xn; ++k) {
        const float wv = wx[k * R + x];
        r0 = fmaf(wv, (float)row[3 * k + 0], r0);
        r1 = fmaf(wv, (float)row[3 * k + 1], r1);
        r2 = fmaf(wv, (float)row[3 * k + 2], r2);
      }
      if (quantize == 2) {     // PIL resamples in two passes and stores the horizontal pass as uint8
        r0 = fminf(fmaxf(rintf(r0), 0.f), 255.f);
        r1 = fminf(fmaxf(rintf(r1), 0.f), 255.f);
        r2 = fminf(fmaxf(rintf(r2), 0.f), 255.f);
      }
      a0 = fmaf(wy, r0, a0);
      a1 = fmaf(wy, r1, a1);
      a2 = fmaf(wy, r2, a2);
    }
    if (quantize) {            // the PIL image handed to ToTensor is uint8
      a0 = fminf(fmaxf(rintf(a0), 0.f), 255.f);
      a1 = fminf(fmaxf(rintf(a1), 0.f), 255.f);
      a2 = fminf(fmaxf(rintf(a2), 0.f), 255.f);
    }
    const long long o = (long long)y * R + xo;
    ob[o] = (a0 * (1.f / 255.f) - m0) / s0;          // ToTensor, Normalize (transforms.py:196)
    ob[plane + o] = (a1 * (1.f / 255.f) - m1) / s1;
    ob[2 * plane + o] = (a2 * (1.f / 255.f) - m2) / s2;
  }
}

}  // namespace psrc

using namespace psrc;

extern "C" int psrc_crop_resize_normalize(const uint8_t* src, int64_t image_stride, int32_t row_stride,
                                          const float* geom_f, const int32_t* geom_i, float* out, int32_t B,
                                          int32_t R, int32_t interpolation, int32_t quantize, int32_t max_taps,
                                          const float* mean_host, const float* std_host, psrc_stream_t stream) {
  PSRC_CHECK_ARG(src && geom_f && geom_i && out && mean_host && std_host, "psrc_crop_resize_normalize: null pointer");
  PSRC_CHECK_ARG(B >= 0 && R > 0 && R <= 256 && (interpolation == 0 || interpolation == 1) && quantize >= 0 &&
                     quantize <= 2 && max_taps > 0 && row_stride > 0,
                 "psrc_crop_resize_normalize: bad arguments (B=%d R=%d interpolation=%d max_taps=%d)", B, R,
                 interpolation, max_taps);
  PSRC_CHECK_ARG(std_host[0] != 0.f && std_host[1] != 0.f && std_host[2] != 0.f, "zero pixel std");
  if (B == 0) return PSRC_OK;
  const size_t smem = (size_t)max_taps * R * sizeof(float);
  PSRC_CHECK_ARG(smem <= 200 * 1024, "psrc_crop_resize_normalize: %d taps x %d columns exceed shared memory "
                                     "(reduce the image on the host first)", max_taps, R);
  if (smem > 48 * 1024)
    PSRC_CHECK_CUDA(cudaFuncSetAttribute(crop_resize_normalize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem));
  dim3 grid((R + kPreRows - 1) / kPreRows, B);
  crop_resize_normalize_kernel<<<grid, 256, smem, as_stream(stream)>>>(
      src, image_stride, row_stride, geom_f, geom_i, out, R, interpolation, quantize, max_taps, mean_host[0],
      mean_host[1], mean_host[2], std_host[0], std_host[1], std_host[2]);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}
