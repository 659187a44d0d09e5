// Feature normalisation and the fused PromptSRC objective with its gradient.
//   l2norm fwd/bwd      : trainers/promptsrc.py:190-191, :196 (x / x.norm(dim=-1, keepdim=True))
//   psrc_loss_fwd_bwd   : trainers/promptsrc.py:204 (CE), :304-324 (L1 text, L1 image, KL logits);
//                         focal CE: trainers/coop.py:146-163, promptsrc_pets_focal.py:93-117.
// One pass over each [B,C] logits row produces its loss partials and d(loss)/d(logits).  All
// reductions use a fixed order (deterministic); every rank of a class/batch-sharded job emits its
// additive share of the single-device loss (divisors are the GLOBAL batch / class counts).
#include <mutex>
#include "common.cuh"

namespace psrc {

constexpr int kL1Blocks = 148;

// one warp per row
__global__ void __launch_bounds__(256)
l2norm_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, float* __restrict__ inv_norm, int rows,
                  int D) {
  const int lane = threadIdx.x & 31;
  const int r = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= rows) return;
  const float* xr = x + (long long)r * D;
  float s = 0.f;
  for (int c = lane; c < D; c += 32) {
    float v = xr[c];
    s += v * v;
  }
  s = warp_sum(s);
  const float inv = 1.f / sqrtf(s);
  if (lane == 0 && inv_norm) inv_norm[r] = inv;
  float* yr = y + (long long)r * D;
  for (int c = lane; c < D; c += 32) yr[c] = xr[c] * inv;
}

__global__ void __launch_bounds__(256)
l2norm_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y, const float* __restrict__ inv_norm,
                  float* __restrict__ dx, int rows, int D) {
  const int lane = threadIdx.x & 31;
  const int r = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= rows) return;
  const float* dyr = dy + (long long)r * D;
  const float* yr = y + (long long)r * D;
  float dot = 0.f;
  for (int c = lane; c < D; c += 32) dot += dyr[c] * yr[c];
  dot = warp_sum(dot);
  const float inv = inv_norm[r];
  float* dxr = dx + (long long)r * D;
  for (int c = lane; c < D; c += 32) dxr[c] = (dyr[c] - yr[c] * dot) * inv;
}

// One CTA per logits row.
__global__ void __launch_bounds__(256)
loss_rows_kernel(psrc_loss_args a) {
  __shared__ float red[32];
  const int b = blockIdx.x;
  const int C = a.C;
  const float* lg = a.logits + (long long)b * C;
  const float* zl = a.zs_logits ? a.zs_logits + (long long)b * C : nullptr;
  float mx = -INFINITY, mz = -INFINITY;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    mx = fmaxf(mx, lg[c]);
    if (zl) mz = fmaxf(mz, zl[c]);
  }
  mx = block_max(mx, red);
  if (zl) mz = block_max(mz, red);
  float sx = 0.f, sz = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    sx += expf(lg[c] - mx);
    if (zl) sz += expf(zl[c] - mz);
  }
  sx = block_sum(sx, red);
  if (zl) sz = block_sum(sz, red);
  const float lse_x = mx + logf(sx);
  const float lse_z = zl ? mz + logf(sz) : 0.f;

  const int y = a.labels[b];
  const float ce = lse_x - lg[y];  // -log p_y
  // d(ce-term)/d(logits_c) = coef * (softmax_c - [c == y])
  float ce_loss, coef;
  if (a.focal_alpha) {
    const float al = a.focal_alpha[y];
    const float pt = expf(-ce);
    const float om = 1.f - pt;
    const float g = a.focal_gamma;
    const float omg = (g == 2.f) ? om * om : powf(om, g);
    const float omg1 = (g == 2.f) ? om : powf(om, g - 1.f);
    ce_loss = al * omg * ce;
    coef = al * (g * omg1 * pt * ce + omg);
  } else {
    ce_loss = ce;
    coef = 1.f;
  }
  const float inv_B = 1.f / (float)a.B_global;
  const float kl_scale = zl ? a.w_kl / ((float)a.B_global * (float)C) : 0.f;
  float kl = 0.f;
  float* dl = a.d_logits + (long long)b * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float p = lg[c] - lse_x;  // log-prob
    const float sm = expf(p);
    float d = coef * inv_B * (sm - (c == y ? 1.f : 0.f));
    if (zl) {
      const float q = zl[c] - lse_z;
      const float eq = expf(q);
      kl += eq * (q - p);
      d += kl_scale * (sm - eq);
    }
    dl[c] = d;
  }
  kl = block_sum(kl, red);
  if (threadIdx.x == 0) {
    a.row_scratch[b] = ce_loss * inv_B;
    a.row_scratch[a.B + b] = kl * kl_scale;
  }
}

// w * mean|a - b| over [rows_global, E]; this rank handles `n` elements; d = w*sign(a-b)/(rows_global*E)
__global__ void __launch_bounds__(256)
l1_kernel(const float* __restrict__ x, const float* __restrict__ t, float* __restrict__ d, long long n, float scale,
          float* __restrict__ partial) {
  __shared__ float red[32];
  float s = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float df = x[i] - t[i];
    s += fabsf(df);
    d[i] = df > 0.f ? scale : (df < 0.f ? -scale : 0.f);
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) partial[blockIdx.x] = s * scale;
}

__global__ void __launch_bounds__(256)
loss_finalize_kernel(const float* __restrict__ scratch, int B, int has_text, int has_image, float* __restrict__ losses) {
  __shared__ float red[32];
  float ce = 0.f, kl = 0.f, lt = 0.f, li = 0.f;
  for (int i = threadIdx.x; i < B; i += blockDim.x) {
    ce += scratch[i];
    kl += scratch[B + i];
  }
  const float* pt = scratch + 2 * B;
  const float* pi = pt + kL1Blocks;
  for (int i = threadIdx.x; i < kL1Blocks; i += blockDim.x) {
    if (has_text) lt += pt[i];
    if (has_image) li += pi[i];
  }
  ce = block_sum(ce, red);
  kl = block_sum(kl, red);
  lt = block_sum(lt, red);
  li = block_sum(li, red);
  if (threadIdx.x == 0) {
    losses[0] = ce; losses[1] = kl; losses[2] = lt; losses[3] = li;
  }
}

}  // namespace psrc

using namespace psrc;

extern "C" int psrc_l2norm_fwd(const float* x, float* y, float* inv_norm, int rows, int D, psrc_stream_t stream) {
  PSRC_CHECK_ARG(x && y && rows >= 0 && D > 0, "psrc_l2norm_fwd: bad arguments");
  if (rows == 0) return PSRC_OK;
  l2norm_fwd_kernel<<<(rows + 7) / 8, 256, 0, as_stream(stream)>>>(x, y, inv_norm, rows, D);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

extern "C" int psrc_l2norm_bwd(const float* dy, const float* y, const float* inv_norm, float* dx, int rows, int D,
                               psrc_stream_t stream) {
  PSRC_CHECK_ARG(dy && y && inv_norm && dx && rows >= 0 && D > 0, "psrc_l2norm_bwd: bad arguments");
  if (rows == 0) return PSRC_OK;
  l2norm_bwd_kernel<<<(rows + 7) / 8, 256, 0, as_stream(stream)>>>(dy, y, inv_norm, dx, rows, D);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

extern "C" int psrc_loss_fwd_bwd(const psrc_loss_args* args, psrc_stream_t stream) {
  PSRC_CHECK_ARG(args != nullptr, "psrc_loss_fwd_bwd: null args");
  const psrc_loss_args& a = *args;
  PSRC_CHECK_ARG(a.B > 0 && a.C > 0 && a.B_global >= a.B, "psrc_loss_fwd_bwd: bad B/C");
  PSRC_CHECK_ARG(a.logits && a.labels && a.d_logits && a.losses && a.row_scratch, "psrc_loss_fwd_bwd: null pointer");
  cudaStream_t s = as_stream(stream);
  loss_rows_kernel<<<a.B, 256, 0, s>>>(a);
  PSRC_CHECK_LAUNCH();
  const int has_text = (a.t_hat && a.t_zs && a.d_that && a.Ct > 0) ? 1 : 0;
  const int has_image = (a.f_hat && a.f_zs && a.d_fhat) ? 1 : 0;
  float* part_t = a.row_scratch + 2 * a.B;
  float* part_i = part_t + kL1Blocks;
  if (has_text) {
    PSRC_CHECK_ARG(a.Ct_global >= a.Ct && a.E > 0, "psrc_loss_fwd_bwd: bad Ct/E");
    const float scale = a.w_text / ((float)a.Ct_global * (float)a.E);
    l1_kernel<<<kL1Blocks, 256, 0, s>>>(a.t_hat, a.t_zs, a.d_that, (long long)a.Ct * a.E, scale, part_t);
    PSRC_CHECK_LAUNCH();
  }
  if (has_image) {
    PSRC_CHECK_ARG(a.E > 0, "psrc_loss_fwd_bwd: bad E");
    const float scale = a.w_image / ((float)a.B_global * (float)a.E);
    l1_kernel<<<kL1Blocks, 256, 0, s>>>(a.f_hat, a.f_zs, a.d_fhat, (long long)a.B * a.E, scale, part_i);
    PSRC_CHECK_LAUNCH();
  }
  loss_finalize_kernel<<<1, 256, 0, s>>>(a.row_scratch, a.B, has_text, has_image, a.losses);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

// ------------------------------------------------------------------------------------------------
// psrc_loss_fwd_bwd_noalloc -- the same objective for callers that cannot hand over a workspace.
//
// psrc_loss_fwd_bwd needs a few KB of global memory (`row_scratch`) between its kernels.  Some hosts
// of this library have no way to provide it: a plugin that is only handed the framework's input and
// output buffers and is forbidden from calling cudaMalloc itself (a serving runtime that has already
// pre-reserved all of the device's memory, a callback running while another thread captures a CUDA
// graph -- cudaMalloc is illegal there --, or a real-time loop that may not take the allocator's
// locks).  For them the workspace lives in the library image: a file-scope __device__ array that each
// GPU gets when it loads this module and keeps until the process exits (34 KB per GPU).
//
// `g_loss_scratch` below is, in HOST code, only a handle for the symbol -- not a device address.  The
// address differs per GPU and is looked up with cudaGetSymbolAddress once per device (with that
// device current) and remembered; the kernels get the looked-up pointer through args.row_scratch.
namespace psrc {

constexpr int kNoAllocMaxRows = 4096;                                    // B limit of the entry point
constexpr int kNoAllocFloats = 2 * kNoAllocMaxRows + 2 * kL1Blocks;      // what loss_finalize_kernel reads
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) float g_loss_scratch[kNoAllocFloats];

static float* g_loss_scratch_addr[kNoAllocMaxDevices];  // device address of g_loss_scratch, per GPU ordinal
static std::mutex g_loss_scratch_mu;

// Device address of the static workspace on the CURRENT device (first use per device does the lookup).
static int loss_scratch_for_current_device(float** out) {
  int dev = -1;
  PSRC_CHECK_CUDA(cudaGetDevice(&dev));
  PSRC_CHECK_ARG(dev >= 0 && dev < kNoAllocMaxDevices,
                 "psrc_loss_fwd_bwd_noalloc: device ordinal %d is beyond the %d this entry point keeps addresses for",
                 dev, kNoAllocMaxDevices);
  std::lock_guard<std::mutex> lock(g_loss_scratch_mu);
  if (g_loss_scratch_addr[dev] == nullptr) {
    float* p = nullptr;
    PSRC_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_loss_scratch));
    PSRC_CHECK_ARG(p != nullptr, "psrc_loss_fwd_bwd_noalloc: cudaGetSymbolAddress returned a null address");
    g_loss_scratch_addr[dev] = p;
  }
  *out = g_loss_scratch_addr[dev];
  return PSRC_OK;
}

}  // namespace psrc

extern "C" int psrc_loss_fwd_bwd_noalloc(const psrc_loss_args* args, psrc_stream_t stream) {
  PSRC_CHECK_ARG(args != nullptr, "psrc_loss_fwd_bwd_noalloc: null args");
  PSRC_CHECK_ARG(args->B > 0 && args->C > 0, "psrc_loss_fwd_bwd_noalloc: bad B/C");
  PSRC_CHECK_ARG(args->B <= kNoAllocMaxRows,
                 "psrc_loss_fwd_bwd_noalloc: B = %d logits rows exceed the built-in workspace (at most %d rows per "
                 "call); split the batch or use psrc_loss_fwd_bwd with a caller-owned row_scratch",
                 args->B, kNoAllocMaxRows);
  float* scratch = nullptr;
  const int rc = loss_scratch_for_current_device(&scratch);
  if (rc != PSRC_OK) return rc;
  // Nothing in the workspace has to start at zero: every slot loss_finalize_kernel reads ([0, 2B) and,
  // per enabled L1 term, its 148 partials) is written earlier in this same call, so what the previous
  // call left behind is never observed.
  psrc_loss_args with_scratch = *args;   // the caller's row_scratch (if any) is ignored
  with_scratch.row_scratch = scratch;
  return psrc_loss_fwd_bwd(&with_scratch, stream);
}
