// Library plumbing (error string, launch counter) and the small vectorised utility kernels:
// dtype casts and the fused momentum-SGD step on the flat prompt-parameter buffer.
#include <atomic>
#include <string.h>
#include "common.cuh"

namespace psrc {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

template <typename TS, typename TD>
__global__ void cast_kernel(const TS* __restrict__ src, TD* __restrict__ dst, long long n) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) dst[i] = from_f32<TD>(to_f32(src[i]));
}

// 4 elements per thread, 16-byte accesses on the fp32 side
__global__ void cast_f32_to_bf16_vec4(const float4* __restrict__ src, uint2* __restrict__ dst, long long n4) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n4; i += stride) {
    float4 f = src[i];
    __nv_bfloat162 a = __floats2bfloat162_rn(f.x, f.y), b = __floats2bfloat162_rn(f.z, f.w);
    uint2 u;
    u.x = *reinterpret_cast<uint32_t*>(&a);
    u.y = *reinterpret_cast<uint32_t*>(&b);
    dst[i] = u;
  }
}

// torch.optim.SGD update rule (Dassl.pytorch/dassl/optim/optimizer.py:105-113 builds it)
__global__ void sgd_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ buf,
                           const float* __restrict__ lr_dev, float momentum, float dampening, float wd,
                           int nesterov, int first, long long n) {
  const float lr = *lr_dev;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    float w = p[i];
    float d = g[i] + wd * w;
    if (momentum != 0.f) {
      float b = first ? d : momentum * buf[i] + (1.f - dampening) * d;
      buf[i] = b;
      d = nesterov ? d + momentum * b : b;
    }
    p[i] = w - lr * d;
  }
}

__global__ void axpy_kernel(float* __restrict__ y, const float* __restrict__ x, float a, long long n) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) y[i] = fmaf(a, x[i], y[i]);
}

}  // namespace psrc

using namespace psrc;

extern "C" const char* psrc_last_error(void) { return g_err; }
extern "C" int psrc_version(void) { return 100; }
extern "C" long long psrc_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

static int grid_for(long long n, int block) {
  long long g = (n + block - 1) / block;
  if (g > 148LL * 16) g = 148LL * 16;  // grid-stride beyond 16 CTAs per SM
  if (g < 1) g = 1;
  return (int)g;
}

extern "C" int psrc_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n,
                         psrc_stream_t stream) {
  PSRC_CHECK_ARG(src && dst && n >= 0, "psrc_cast: bad arguments");
  if (n == 0) return PSRC_OK;
  cudaStream_t s = as_stream(stream);
  if (src_dtype == PSRC_F32 && dst_dtype == PSRC_BF16) {
    if (n % 4 == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 7) == 0) {
      cast_f32_to_bf16_vec4<<<grid_for(n / 4, 256), 256, 0, s>>>(reinterpret_cast<const float4*>(src),
                                                                  reinterpret_cast<uint2*>(dst), n / 4);
    } else {
      cast_kernel<float, __nv_bfloat16><<<grid_for(n, 256), 256, 0, s>>>(
          reinterpret_cast<const float*>(src), reinterpret_cast<__nv_bfloat16*>(dst), n);
    }
  } else if (src_dtype == PSRC_BF16 && dst_dtype == PSRC_F32) {
    cast_kernel<__nv_bfloat16, float><<<grid_for(n, 256), 256, 0, s>>>(
        reinterpret_cast<const __nv_bfloat16*>(src), reinterpret_cast<float*>(dst), n);
  } else if (src_dtype == PSRC_F32 && dst_dtype == PSRC_F32) {
    cast_kernel<float, float><<<grid_for(n, 256), 256, 0, s>>>(reinterpret_cast<const float*>(src),
                                                               reinterpret_cast<float*>(dst), n);
  } else if (src_dtype == PSRC_BF16 && dst_dtype == PSRC_BF16) {
    cast_kernel<__nv_bfloat16, __nv_bfloat16><<<grid_for(n, 256), 256, 0, s>>>(
        reinterpret_cast<const __nv_bfloat16*>(src), reinterpret_cast<__nv_bfloat16*>(dst), n);
  } else {
    PSRC_CHECK_ARG(false, "psrc_cast: unsupported dtype pair %d -> %d", src_dtype, dst_dtype);
  }
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

extern "C" int psrc_sgd_step(float* p, const float* g, float* buf, const float* lr_dev, float momentum,
                             float dampening, float weight_decay, int nesterov, int first_step, int64_t n,
                             psrc_stream_t stream) {
  PSRC_CHECK_ARG(p && g && lr_dev && n >= 0, "psrc_sgd_step: bad arguments");
  PSRC_CHECK_ARG(momentum == 0.f || buf != nullptr, "psrc_sgd_step: momentum needs a buffer");
  if (n == 0) return PSRC_OK;
  sgd_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(p, g, buf, lr_dev, momentum, dampening,
                                                              weight_decay, nesterov, first_step, n);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

extern "C" int psrc_axpy(float* y, const float* x, float a, int64_t n, psrc_stream_t stream) {
  PSRC_CHECK_ARG(y && x && n >= 0, "psrc_axpy: bad arguments");
  if (n == 0) return PSRC_OK;
  axpy_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(y, x, a, n);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}
