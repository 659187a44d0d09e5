// Attention for very short sequences (L <= 32, head_dim 64): the truncated text tower (the prompts
// "a photo of a <name>." end at position 8..24, so the text transformer runs L = max(EOT)+1 positions).
//
// At L = 11 a (sequence, head) problem is an 11 x 11 score matrix -- far below one tensor-core tile -- and the
// layer has n_seq * H = 8000 of them, so the kernel is organised for occupancy, not for MMA throughput:
// one (sequence, head) per HALF warp when L <= 16 (two per warp), one per warp otherwise; a lane owns one
// query row (forward, dQ) resp. one key row (dK, dV); K / V / Q / dO rows are staged in shared memory in
// their 16-bit storage type with coalesced 16-byte loads and read back as warp-wide broadcasts; all
// arithmetic is fp32; the backward keeps P and dS in a padded shared-memory tile so each is computed once.
// No atomics, fixed summation order.  Reference: nn.MultiheadAttention core called from
// PromptSRC/clip/model.py:183/227/285 with the causal mask of model.py:592-598.
#include "common.cuh"

namespace psrc {
namespace asm_ {

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr int kRow = 64;  // head dim

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <typename T> struct Cvt;
template <> struct Cvt<__nv_bfloat16> {
  // bf16 -> fp32 is a 16-bit shift: two integer-pipe ops per pair instead of two conversion-pipe CVTs (the
  // kernel is otherwise bound by the quarter-rate conversion unit)
  static __device__ __forceinline__ void unpack8(const uint4& u, float* f) {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
    }
  }
  static __device__ __forceinline__ uint32_t pack2(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  }
};
template <> struct Cvt<__half> {
  static __device__ __forceinline__ void unpack8(const uint4& u, float* f) {
    const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 v = __half22float2(h[i]);
      f[2 * i] = v.x;
      f[2 * i + 1] = v.y;
    }
  }
  static __device__ __forceinline__ uint32_t pack2(float a, float b) {
    __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  }
};

// dot of 64 fp32 registers with one 128-byte shared-memory row (broadcast read)
template <typename T>
__device__ __forceinline__ float dot_row(const float (&a)[kRow], const uint4* __restrict__ row) {
  float acc[8];   // eight independent FMA chains (a single chain of 64 is latency-bound at this occupancy)
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    float f[8];
    Cvt<T>::unpack8(row[c], f);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = fmaf(a[8 * c + i], f[i], acc[i]);
  }
  return ((acc[0] + acc[1]) + (acc[2] + acc[3])) + ((acc[4] + acc[5]) + (acc[6] + acc[7]));
}
// acc[64] += w * row
template <typename T>
__device__ __forceinline__ void axpy_row(float (&acc)[kRow], float w, const uint4* __restrict__ row) {
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    float f[8];
    Cvt<T>::unpack8(row[c], f);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[8 * c + i] = fmaf(w, f[i], acc[8 * c + i]);
  }
}
template <typename T>
__device__ __forceinline__ void load_row(const T* __restrict__ g, float (&v)[kRow], float scale) {
  const uint4* p = reinterpret_cast<const uint4*>(g);
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    float f[8];
    Cvt<T>::unpack8(__ldg(p + c), f);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[8 * c + i] = f[i] * scale;
  }
}
template <typename T>
__device__ __forceinline__ void store_row(T* __restrict__ g, const float (&v)[kRow], float scale) {
  uint4* p = reinterpret_cast<uint4*>(g);
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    uint4 u;
    u.x = Cvt<T>::pack2(v[8 * c] * scale, v[8 * c + 1] * scale);
    u.y = Cvt<T>::pack2(v[8 * c + 2] * scale, v[8 * c + 3] * scale);
    u.z = Cvt<T>::pack2(v[8 * c + 4] * scale, v[8 * c + 5] * scale);
    u.w = Cvt<T>::pack2(v[8 * c + 6] * scale, v[8 * c + 7] * scale);
    p[c] = u;
  }
}

// cooperative copy of L rows (64 elements, 128 B each) of one matrix of one (seq, head) into shared memory
template <typename T>
__device__ __forceinline__ void stage_rows(uint4* __restrict__ dst, const T* __restrict__ src, long long ld, int L,
                                           int lane) {
  for (int idx = lane; idx < L * 8; idx += 32) {
    const int r = idx >> 3, c = idx & 7;
    dst[idx] = __ldg(reinterpret_cast<const uint4*>(src + (long long)r * ld) + c);
  }
}

constexpr int kFwdWarps = 4;

// LP = rows per problem slot (16: two problems per warp, 32: one)
template <typename T, int LP, bool kCausal>
__global__ void __launch_bounds__(kFwdWarps * 32)
attn_small_fwd_kernel(const T* __restrict__ qkv, T* __restrict__ out, float* __restrict__ lse, int n_prob, int L,
                      int H) {
  constexpr int G = 32 / LP;
  __shared__ uint4 sm[kFwdWarps][G][2][LP * 8];  // [warp][problem][K|V][rows * 8 chunks]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sub = lane / LP, r = lane % LP;
  const int D = H * kRow;
  const long long ld = 3LL * D;
  const int prob0 = (blockIdx.x * kFwdWarps + warp) * G;
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const int pr = prob0 + g;
    if (pr < n_prob) {
      const int seq = pr / H, h = pr % H;
      const T* base = qkv + (long long)seq * L * ld + h * kRow;
      stage_rows(sm[warp][g][0], base + D, ld, L, lane);
      stage_rows(sm[warp][g][1], base + 2 * D, ld, L, lane);
    }
  }
  __syncwarp();
  const int pr = prob0 + sub;
  if (pr >= n_prob || r >= L) return;
  const int seq = pr / H, h = pr % H;
  float q[kRow];
  load_row(qkv + ((long long)seq * L + r) * ld + h * kRow, q, 0.125f * kLog2e);
  const uint4* Ks = sm[warp][sub][0];
  const uint4* Vs = sm[warp][sub][1];
  const int jmax = kCausal ? r : L - 1;
  float s[LP];
  float m = -INFINITY;
#pragma unroll
  for (int j = 0; j < LP; ++j) {
    s[j] = -INFINITY;
    if (j <= jmax) {
      s[j] = dot_row<T>(q, Ks + j * 8);
      m = fmaxf(m, s[j]);
    }
  }
  float sum = 0.f;
  float o[kRow];
#pragma unroll
  for (int i = 0; i < kRow; ++i) o[i] = 0.f;
#pragma unroll
  for (int j = 0; j < LP; ++j) {
    if (j <= jmax) {
      const float p = ex2(s[j] - m);
      sum += p;
      axpy_row<T>(o, p, Vs + j * 8);
    }
  }
  store_row(out + ((long long)seq * L + r) * D + h * kRow, o, 1.f / sum);
  lse[((long long)seq * H + h) * L + r] = (m + log2f(sum)) * kLn2;
}

constexpr int kBwdWarps = 2;

template <int LP>
struct BwdSmem {
  static constexpr int G = 32 / LP;
  static constexpr int kMatBytes = LP * 128;                 // one staged matrix
  static constexpr int kTileFloats = LP * (LP + 1);          // padded P / dS tile
  static constexpr int kProbBytes = 4 * kMatBytes + 2 * kTileFloats * 4;
  static constexpr int kWarpBytes = G * kProbBytes;
  static constexpr int kBytes = kBwdWarps * kWarpBytes;
};

// TA: storage type of q/k/v/out (forward activations); TG: storage type of dout / dqkv (gradients)
template <typename TA, typename TG, int LP, bool kCausal>
__global__ void __launch_bounds__(kBwdWarps * 32)
attn_small_bwd_kernel(const TA* __restrict__ qkv, const TA* __restrict__ out, const TG* __restrict__ dout,
                      const float* __restrict__ lse, TG* __restrict__ dqkv, int n_prob, int L, int H) {
  using S = BwdSmem<LP>;
  constexpr int G = S::G;
  extern __shared__ uint4 smem_dyn[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sub = lane / LP, r = lane % LP;
  const int D = H * kRow;
  const long long ld = 3LL * D;
  uint8_t* wbase = reinterpret_cast<uint8_t*>(smem_dyn) + warp * S::kWarpBytes;
  const int prob0 = (blockIdx.x * kBwdWarps + warp) * G;
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const int pr = prob0 + g;
    if (pr < n_prob) {
      const int seq = pr / H, h = pr % H;
      uint4* m0 = reinterpret_cast<uint4*>(wbase + g * S::kProbBytes);
      const TA* base = qkv + (long long)seq * L * ld + h * kRow;
      stage_rows(m0, base, ld, L, lane);                                       // Q
      stage_rows(m0 + LP * 8, base + D, ld, L, lane);                          // K
      stage_rows(m0 + 2 * LP * 8, base + 2 * D, ld, L, lane);                  // V
      stage_rows(m0 + 3 * LP * 8, dout + (long long)seq * L * D + h * kRow, (long long)D, L, lane);  // dO
    }
  }
  __syncwarp();
  const int pr = prob0 + sub;
  const bool active = pr < n_prob && r < L;
  const int seq = active ? pr / H : 0, h = active ? pr % H : 0;
  uint8_t* pbase = wbase + sub * S::kProbBytes;
  const uint4* Qs = reinterpret_cast<const uint4*>(pbase);
  const uint4* Ks = Qs + LP * 8;
  const uint4* Vs = Ks + LP * 8;
  const uint4* dOs = Vs + LP * 8;
  float* Pt = reinterpret_cast<float*>(pbase + 4 * S::kMatBytes);   // [i][j], row stride LP + 1
  float* dSt = Pt + S::kTileFloats;

  // ---- pass A (lane = query row i): P_ij, dS_ij, dQ_i
  if (active) {
    float a[kRow];
    float pv[LP];
    const int jmax = kCausal ? r : L - 1;
    {  // scores -> probabilities
      // q from shared memory (row r of Qs): conflict-free enough (each lane its own 128-byte row, 16-byte reads)
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        float f[8];
        Cvt<TA>::unpack8(Qs[r * 8 + c], f);
#pragma unroll
        for (int i = 0; i < 8; ++i) a[8 * c + i] = f[i] * (0.125f * kLog2e);
      }
      const float l2 = lse[((long long)seq * H + h) * L + r] * kLog2e;
#pragma unroll
      for (int j = 0; j < LP; ++j) {
        pv[j] = 0.f;
        if (j <= jmax) pv[j] = ex2(dot_row<TA>(a, Ks + j * 8) - l2);
      }
    }
    // dO_i (fp32 registers, reusing `a`) and delta_i = <dO_i, O_i>
    float delta = 0.f;
    {
      float ov[kRow];
      load_row(out + ((long long)seq * L + r) * D + h * kRow, ov, 1.f);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        float f[8];
        Cvt<TG>::unpack8(dOs[r * 8 + c], f);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          a[8 * c + i] = f[i];
          delta = fmaf(f[i], ov[8 * c + i], delta);
        }
      }
    }
    float dq[kRow];
#pragma unroll
    for (int i = 0; i < kRow; ++i) dq[i] = 0.f;
#pragma unroll
    for (int j = 0; j < LP; ++j) {
      float ds = 0.f;
      if (j <= jmax) {
        ds = pv[j] * (dot_row<TA>(a, Vs + j * 8) - delta);
        axpy_row<TA>(dq, ds, Ks + j * 8);
      }
      if (j < L) {
        Pt[r * (LP + 1) + j] = pv[j];
        dSt[r * (LP + 1) + j] = ds;
      }
    }
    store_row(dqkv + ((long long)seq * L + r) * ld + h * kRow, dq, 0.125f);
  }
  __syncwarp();
  // ---- pass B (lane = key row j): dV_j = sum_i P_ij dO_i ; dK_j = scale * sum_i dS_ij Q_i
  if (active) {
    const int i0 = kCausal ? r : 0;
    float acc[kRow];
#pragma unroll
    for (int i = 0; i < kRow; ++i) acc[i] = 0.f;
    for (int i = i0; i < L; ++i) axpy_row<TG>(acc, Pt[i * (LP + 1) + r], dOs + i * 8);
    store_row(dqkv + ((long long)seq * L + r) * ld + 2 * D + h * kRow, acc, 1.f);
#pragma unroll
    for (int i = 0; i < kRow; ++i) acc[i] = 0.f;
    for (int i = i0; i < L; ++i) axpy_row<TA>(acc, dSt[i * (LP + 1) + r], Qs + i * 8);
    store_row(dqkv + ((long long)seq * L + r) * ld + D + h * kRow, acc, 0.125f);
  }
}

}  // namespace asm_

bool attn_small_supported(int L) { return L >= 1 && L <= 32; }

template <typename T>
static int fwd_launch(const void* qkv, void* out, float* lse, int n_seq, int L, int H, int causal, cudaStream_t s) {
  using namespace asm_;
  const int n_prob = n_seq * H;
  const T* q = reinterpret_cast<const T*>(qkv);
  T* o = reinterpret_cast<T*>(out);
  if (L <= 16) {
    const int grid = (n_prob + kFwdWarps * 2 - 1) / (kFwdWarps * 2);
    if (causal) attn_small_fwd_kernel<T, 16, true><<<grid, kFwdWarps * 32, 0, s>>>(q, o, lse, n_prob, L, H);
    else attn_small_fwd_kernel<T, 16, false><<<grid, kFwdWarps * 32, 0, s>>>(q, o, lse, n_prob, L, H);
  } else {
    const int grid = (n_prob + kFwdWarps - 1) / kFwdWarps;
    if (causal) attn_small_fwd_kernel<T, 32, true><<<grid, kFwdWarps * 32, 0, s>>>(q, o, lse, n_prob, L, H);
    else attn_small_fwd_kernel<T, 32, false><<<grid, kFwdWarps * 32, 0, s>>>(q, o, lse, n_prob, L, H);
  }
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

template <typename TA, typename TG, int LP, bool kCausal>
static int bwd_launch_one(const TA* qkv, const TA* out, const TG* dout, const float* lse, TG* dqkv, int n_prob, int L,
                          int H, cudaStream_t s) {
  using namespace asm_;
  using S = BwdSmem<LP>;
  auto kernel = attn_small_bwd_kernel<TA, TG, LP, kCausal>;
  static bool attr_set = false;
  if (!attr_set) {
    PSRC_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kBytes));
    attr_set = true;
  }
  const int per_cta = kBwdWarps * S::G;
  kernel<<<(n_prob + per_cta - 1) / per_cta, kBwdWarps * 32, S::kBytes, s>>>(qkv, out, dout, lse, dqkv, n_prob, L, H);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

template <typename TA, typename TG>
static int bwd_launch(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, int n_seq,
                      int L, int H, int causal, cudaStream_t s) {
  const int n_prob = n_seq * H;
  const TA* q = reinterpret_cast<const TA*>(qkv);
  const TA* o = reinterpret_cast<const TA*>(out);
  const TG* d = reinterpret_cast<const TG*>(dout);
  TG* dq = reinterpret_cast<TG*>(dqkv);
  if (L <= 16) {
    return causal ? bwd_launch_one<TA, TG, 16, true>(q, o, d, lse, dq, n_prob, L, H, s)
                  : bwd_launch_one<TA, TG, 16, false>(q, o, d, lse, dq, n_prob, L, H, s);
  }
  return causal ? bwd_launch_one<TA, TG, 32, true>(q, o, d, lse, dq, n_prob, L, H, s)
                : bwd_launch_one<TA, TG, 32, false>(q, o, d, lse, dq, n_prob, L, H, s);
}

// act_dtype: q/k/v/out storage; grad_dtype: dout/dqkv storage (PSRC_BF16 or PSRC_F16)
int attn_fwd_small(const void* qkv, void* out, float* lse, int act_dtype, int n_seq, int L, int H, int causal,
                   cudaStream_t s) {
  if (act_dtype == PSRC_F16) return fwd_launch<__half>(qkv, out, lse, n_seq, L, H, causal, s);
  return fwd_launch<__nv_bfloat16>(qkv, out, lse, n_seq, L, H, causal, s);
}

int attn_bwd_small(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, int act_dtype,
                   int grad_dtype, int n_seq, int L, int H, int causal, cudaStream_t s) {
  if (act_dtype == PSRC_F16 && grad_dtype == PSRC_BF16)
    return bwd_launch<__half, __nv_bfloat16>(qkv, out, dout, lse, dqkv, n_seq, L, H, causal, s);
  if (act_dtype == PSRC_F16 && grad_dtype == PSRC_F16)
    return bwd_launch<__half, __half>(qkv, out, dout, lse, dqkv, n_seq, L, H, causal, s);
  if (act_dtype == PSRC_BF16 && grad_dtype == PSRC_BF16)
    return bwd_launch<__nv_bfloat16, __nv_bfloat16>(qkv, out, dout, lse, dqkv, n_seq, L, H, causal, s);
  set_error("attn_bwd_small: unsupported dtype pair act=%d grad=%d", act_dtype, grad_dtype);
  return PSRC_ERR_UNSUPPORTED;
}

}  // namespace psrc
