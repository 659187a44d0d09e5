// Single-CTA tcgen05 GEMM: general fallback for epilogue combinations / shapes the CTA-pair kernel
// (gemm_pair.cu) does not cover, and the A/B reference it is measured against.
// Dense contraction  out = act(alpha * A . W^T + bias) + residual  on the 5th-gen tensor cores.
//
// Persistent, warp-specialised sm_100a kernel: one CTA per SM loops over 128 x BN output
// tiles; warp 0 feeds a ring of shared-memory stages with TMA (128B-swizzled K-major boxes of
// A and W), one elected thread of warp 1 issues tcgen05.mma (UMMA 128 x BN x 16, bf16 inputs,
// fp32 accumulators in TMEM, two accumulator stages so the epilogue of tile i overlaps the
// MMAs of tile i+1), warps 4..7 drain TMEM with tcgen05.ld and apply the fused epilogue.
// Replaces the nn.Linear / MHA-projection / conv1 / `@ proj` call sites listed in
// include/psrc_b200.h (PromptSRC/clip/model.py:171-176,183,376,428; trainers/promptsrc.py:68).
#include <cuda.h>
#include <algorithm>
#include "gemm_epilogue.cuh"

namespace psrc {

constexpr int kBM = 128;       // tile rows  (= UMMA M, one TMEM lane per row)
constexpr int kBK = 64;        // K per stage: 64 bf16 = 128 B = one swizzle atom row
constexpr int kUmmaK = 16;     // K per tcgen05.mma for 16-bit inputs
constexpr int kGemmThreads = 256;
constexpr uint32_t kSpinLimit = 1u << 26;

__device__ __forceinline__ void mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
  for (uint32_t i = 0; !mbar_try_wait(bar, parity); ++i) {
    if (i > kSpinLimit) __trap();  // a lost arrival must not hang the GPU box
  }
}

__device__ __forceinline__ uint64_t make_smem_desc_sw128(uint32_t saddr) {
  // K-major, SWIZZLE_128B canonical layout: rows of 128 B, 8-row atoms 1024 B apart (SBO),
  // LBO unused (one atom along K), descriptor version 1 (Blackwell), base_offset 0.
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)(1024u >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

template <int BN, int STAGES>
struct GemmSmem {
  static constexpr int kABytes = kBM * kBK * 2;
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarrierOffset = STAGES * kStageBytes;
  // full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2], tmem base address
  static constexpr int kTotalBytes = kBarrierOffset + (2 * STAGES + 4) * 8 + 16;
  static constexpr int kDynBytes = kTotalBytes + 1024;  // slack for manual 1024 B alignment
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
               int K, int tiles_m, int tiles_n, int ab_f16, EpiParams epi) {
  using S = GemmSmem<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarrierOffset);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_kb = (K + kBK - 1) / kBK;
  const int num_tiles = tiles_m * tiles_n;
  constexpr uint32_t kTmemCols = 2 * BN;  // 256 or 512: power of two

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_w);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 4);  // one arrival per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_base_slot, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m_blk = tile / tiles_n, n_blk = tile % tiles_n;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait_bounded(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * S::kStageBytes;
          uint8_t* sb = sa + S::kABytes;
          mbar_expect_tx(&full_bar[stage], S::kStageBytes);
          tma_load_2d(&tmap_a, &full_bar[stage], sa, kb * kBK, m_blk * kBM);
          tma_load_2d(&tmap_w, &full_bar[stage], sb, kb * kBK, n_blk * BN);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      // instruction descriptor: D=f32, A=B=bf16, both K-major, N = BN, M = 128
      const uint32_t idesc = (1u << 4) | (ab_f16 ? 0u : ((1u << 7) | (1u << 10))) | ((uint32_t)(BN >> 3) << 17) |
                                 ((uint32_t)(kBM >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait_bounded(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait_bounded(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * S::kStageBytes);
          const uint32_t sb = sa + S::kABytes;
#pragma unroll
          for (int k = 0; k < kBK / kUmmaK; ++k) {
            const uint64_t da = make_smem_desc_sw128(sa + k * kUmmaK * 2);
            const uint64_t db = make_smem_desc_sw128(sb + k * kUmmaK * 2);
            umma_bf16(d_tmem, da, db, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);  // smem stage reusable once these MMAs retire
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tmem_full[as]);  // accumulator ready for the epilogue
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue warps (TMEM -> regs -> global) =====================
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    int as = 0;
    uint32_t aphase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m_blk = tile / tiles_n, n_blk = tile % tiles_n;
      mbar_wait_bounded(&tmem_full[as], aphase);
      tc_fence_after();
      const int row = m_blk * kBM + q * 32 + lane;
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN);
#pragma unroll 1
      for (int c = 0; c < BN; c += 32) {
        const int col0 = n_blk * BN + c;
        if (col0 >= epi.N) break;  // warp-uniform
        uint32_t r[32];
        tmem_ld_32x32(t_row + (uint32_t)c, r);
        tmem_ld_wait();
        if (row < epi.M) {
          float v[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
          epilogue_chunk(epi, row, col0, v);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[as]);
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

// ---------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// rows x K bf16 row-major matrix (row stride ld elements) -> boxes of box_rows x 64, 128B swizzle
static int make_tmap(CUtensorMap* map, const void* base, long long rows, long long K, long long ld,
                     int box_rows, int f16) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return PSRC_ERR_CUDA;
  }
  cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                  const_cast<void*>(base), gdim, gstride, box,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed: CUresult %d (rows=%lld K=%lld ld=%lld box=%d)", (int)r, rows,
              K, ld, box_rows);
    return PSRC_ERR_CUDA;
  }
  return PSRC_OK;
}

static int g_num_sms = 0;
static int num_sms() {
  if (g_num_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  return g_num_sms;
}

template <int BN, int STAGES>
static int launch_tc(const psrc_gemm_args& a, const EpiParams& epi, cudaStream_t stream) {
  using S = GemmSmem<BN, STAGES>;
  static bool attr_set = false;
  if (!attr_set) {
    PSRC_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN, STAGES>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, S::kDynBytes));
    attr_set = true;
  }
  CUtensorMap ta, tw;
  const int f16 = a.ab_dtype == PSRC_F16;
  int rc = make_tmap(&ta, a.A, a.M, a.K, a.lda, kBM, f16);
  if (rc) return rc;
  rc = make_tmap(&tw, a.W, a.N, a.K, a.ldw, BN, f16);
  if (rc) return rc;
  const int tiles_m = (a.M + kBM - 1) / kBM, tiles_n = (a.N + BN - 1) / BN;
  const int grid = std::min(tiles_m * tiles_n, num_sms());
  gemm_tc_kernel<BN, STAGES><<<grid, kGemmThreads, S::kDynBytes, stream>>>(ta, tw, a.K, tiles_m, tiles_n, f16, epi);
  PSRC_CHECK_LAUNCH();
  return PSRC_OK;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

bool gemm_tc_supported(const psrc_gemm_args& a) {
  return (a.ab_dtype == PSRC_BF16 || a.ab_dtype == PSRC_F16) && a.K % 8 == 0 && a.lda % 8 == 0 && a.ldw % 8 == 0 && aligned16(a.A) &&
         aligned16(a.W) && a.M > 0 && a.N > 0 && a.K > 0;
}

void fill_epi(const psrc_gemm_args& a, EpiParams& e) {
  e.M = a.M; e.N = a.N;
  e.bias = a.bias; e.act = a.act;
  e.aux = a.aux; e.ldaux = a.ldaux; e.aux_dtype = a.aux_dtype;
  e.residual = a.residual; e.ldres = a.ldres;
  e.out = a.out; e.ldo = a.ldo; e.out_dtype = a.out_dtype;
  e.out2 = a.out2; e.ldo2 = a.ldo2; e.out2_dtype = a.out2_dtype;
  e.alpha = a.alpha;
  auto ok = [](const void* p, long long ld, int dtype) {
    if (!p) return true;
    int per16 = dtype == PSRC_F32 ? 4 : 8;
    return aligned16(p) && (ld % per16 == 0);
  };
  e.vec_ok = (a.N % 8 == 0) && ok(a.bias, 4, PSRC_F32) && ok(a.aux, a.ldaux, a.aux_dtype) &&
             ok(a.residual, a.ldres, PSRC_F32) && ok(a.out, a.ldo, a.out_dtype) &&
             ok(a.out2, a.ldo2, a.out2_dtype);
}

int gemm_tc1(const psrc_gemm_args& a, cudaStream_t stream) {
  EpiParams epi;
  fill_epi(a, epi);
  // 128x256 tiles when they still fill the machine at least ~2x, else 128x128
  const long long tiles256 = (long long)((a.M + kBM - 1) / kBM) * ((a.N + 255) / 256);
  if (a.N > 128 && tiles256 >= 2LL * num_sms()) return launch_tc<256, 4>(a, epi, stream);
  return launch_tc<128, 6>(a, epi, stream);
}

}  // namespace psrc
