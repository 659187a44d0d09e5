// Dense contraction  out = act(alpha * A . W^T + bias) (+ residual)  on CTA PAIRS (tcgen05 cta_group::2).
//
// One persistent cluster of two CTAs per SM pair (TPC) loops over 256 x BN output tiles.  Each CTA stages
// its own 128 rows of A and its own BN/2 rows of W with TMA (128B-swizzled K-major boxes, 64 K-elements per
// stage) -- so every operand byte is fetched from L2 once per pair and each SM reads only half of W from
// shared memory per MMA -- and the leader CTA's elected thread issues tcgen05.mma.cta_group::2
// (UMMA 256 x BN x 16, bf16 in, fp32 accumulators in the TMEM of both SMs, two accumulator stages).
// Epilogue (both CTAs, warps 4..11: two warps per TMEM lane quarter taking the even / odd 32-column chunks, one
// TMEM lane = one output row per thread):
//   tcgen05.ld 32 columns -> registers -> fused bias / QuickGELU / QuickGELU' * aux / + residual ->
//   bank-conflict-free swizzled shared-memory staging -> TMA bulk tensor store (coalesced, asynchronous).
// The fp32 residual / bf16 aux operands are prefetched by a dedicated warp with TMA into a ring of
// shared-memory slots, so the epilogue never issues an uncoalesced global access.
// Replaces the nn.Linear / MHA in/out-projection / conv1 / `@ proj` call sites listed in
// include/psrc_b200.h (PromptSRC/clip/model.py:171-176,183,376,428; trainers/promptsrc.py:68) and their
// autograd input-gradients.
#include <cuda.h>
#include <algorithm>
#include <stdlib.h>
#include "gemm_epilogue.cuh"

namespace psrc {

namespace pair {

constexpr int kBMcta = 128;     // rows per CTA (UMMA M = 256 over the pair)
constexpr int kBK = 64;         // K per stage: 64 bf16 = 128 B = one swizzle atom row
constexpr int kUmmaK = 16;
constexpr int kEpiWarps = 8;    // two epilogue warps per TMEM lane quarter: even / odd 32-column chunks
constexpr int kThreads = (4 + kEpiWarps) * 32;
constexpr int kInSlots = 4;     // ring of prefetched residual / aux chunks (128 rows x 32 columns each)
constexpr int kOutSlotBytes = 32 * 128;   // per epilogue warp: 32 rows x 32 fp32 (or bf16 out + bf16 out2)
constexpr uint32_t kSpinLimit = 1u << 27;

// kResidualInPlace: out aliases the residual (forward-only towers update the residual stream in place): the epilogue
// stages alpha*acc + bias like kNone and lets the TMA store ADD it to global memory -- no residual bytes cross the SM
enum Mode { kNone = 0, kGelu = 1, kGeluGrad = 2, kResidual = 3, kResidualInPlace = 4 };

__device__ __forceinline__ void wait_bounded(uint64_t* bar, uint32_t parity) {
  for (uint32_t i = 0; !mbar_try_wait(bar, parity); ++i) {
    if (i > kSpinLimit) __trap();  // a lost arrival must not hang the GPU box
  }
}

__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
  // K-major, SWIZZLE_128B canonical layout: 128-byte rows, 8-row atoms 1024 B apart (SBO), version 1
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)(1024u >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

template <int BN, int STAGES, int IN_SLOTS, int IN_SLOT_BYTES>
struct Smem {
  static constexpr int kABytes = kBMcta * kBK * 2;        // 16 KB
  static constexpr int kBBytes = (BN / 2) * kBK * 2;      // this CTA's half of the W tile
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kInOffset = STAGES * kStageBytes;
  static constexpr int kInSlotBytes = IN_SLOT_BYTES;
  static constexpr int kOutOffset = kInOffset + IN_SLOTS * IN_SLOT_BYTES;
  static constexpr int kBarOffset = kOutOffset + kEpiWarps * kOutSlotBytes;   // one staging slot per epilogue warp
  // full[STAGES] empty[STAGES] tmem_full[2] tmem_empty[2] in_full[IN_SLOTS] in_empty[IN_SLOTS] + tmem slot
  static constexpr int kNumBars = 2 * STAGES + 4 + 2 * IN_SLOTS;
  static constexpr int kTotal = kBarOffset + kNumBars * 8 + 16;
  static constexpr int kDyn = kTotal + 1024;               // slack for manual 1024 B alignment
};

struct Params {
  int M, N, K;
  int tiles_m, tiles_n;   // tiles of 256 x BN
  const float* bias;
  float alpha;
  int has_out2;
  // 16-bit formats (0 = bf16, 1 = fp16): A / W operands, the 16-bit output(s), the QuickGELU' aux operand.  fp16
  // forward operands carry 3 more mantissa bits than bf16 at the same tensor-core rate ("fp16 forward")
  int ab_f16, out_f16, out2_f16, aux_f16;
  // Tail split (wave quantisation): work items [0, n_full) are whole 256 x BN tiles; every remaining tile is cut
  // along N into `split` items of 256 x (BN / split), so the last, partially filled round of the persistent loop
  // occupies split x as many CTA pairs for a fraction of a tile time each.  split = 1: no tail items.
  int n_full, split, items;
};

struct Item {
  int m_blk, col0, bn;
};
template <int BN>
__device__ __forceinline__ Item decode_item(int w, const Params& p) {
  Item it;
  int tile = w, sub = 0;
  it.bn = BN;
  if (w >= p.n_full) {
    const int t = w - p.n_full;
    tile = p.n_full + t / p.split;
    sub = t % p.split;
    it.bn = BN / p.split;
  }
  it.m_blk = tile / p.tiles_n;
  it.col0 = (tile % p.tiles_n) * BN + sub * it.bn;
  return it;
}

__device__ __forceinline__ uint32_t pack16(float lo, float hi, int f16) {
  if (f16) {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack16(uint32_t u, int f16) {
  if (f16) return __half22float2(*reinterpret_cast<__half2*>(&u));
  return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u));
}

// 16-byte chunk j of row r inside a staging box whose rows are RB bytes (128 -> SWIZZLE_128B, 64 -> SWIZZLE_64B)
template <int RB>
__device__ __forceinline__ uint32_t sw_chunk(int r, int j) {
  if (RB == 128) return (uint32_t)(r * 128 + ((j ^ (r & 7)) << 4));
  return (uint32_t)(r * 64 + ((j ^ ((r >> 1) & 3)) << 4));
}

template <int BN, int STAGES, int MODE, int OUT_BYTES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                 const __grid_constant__ CUtensorMap tmap_w_tail, const __grid_constant__ CUtensorMap tmap_out,
                 const __grid_constant__ CUtensorMap tmap_out2, const __grid_constant__ CUtensorMap tmap_in, Params p) {
  constexpr bool kHasIn = (MODE == kGeluGrad || MODE == kResidual);
  constexpr int kIn = kHasIn ? kInSlots : 0;   // epilogues without a second operand give the ring's 64 KB to the pipeline
  using S = Smem<BN, STAGES, kIn, (MODE == kResidual) ? 128 * 128 : 128 * 64>;   // fp32 residual / bf16 aux chunk
  constexpr int kInRowBytes = (MODE == kResidual) ? 128 : 64;
  constexpr int kOutRowBytes = 32 * OUT_BYTES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarOffset);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* in_full = tmem_empty + 2;
  uint64_t* in_empty = in_full + kIn;
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(in_empty + kIn);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();        // 0 = MMA leader
  const int pair_id = blockIdx.x >> 1;
  const int n_pairs = gridDim.x >> 1;
  const int num_kb = (p.K + kBK - 1) / kBK;
  constexpr uint32_t kTmemCols = 2 * BN > 256 ? 512 : 256;   // two accumulator stages, power-of-two allocation

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_w);
    if (p.split > 1) tma_prefetch_desc(&tmap_w_tail);
    tma_prefetch_desc(&tmap_out);
    if (MODE == kGelu) tma_prefetch_desc(&tmap_out2);
    if (kHasIn) tma_prefetch_desc(&tmap_in);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);    // leader's producer arms it; bytes arrive from both CTAs
      mbar_init(&empty_bar[s], 1);   // multicast tcgen05.commit
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);   // multicast tcgen05.commit
      mbar_init(&tmem_empty[s], 2 * kEpiWarps);  // epilogue warps x 2 CTAs (leader's copy is the one used)
    }
    for (int s = 0; s < kIn; ++s) {
      mbar_init(&in_full[s], 1);
      mbar_init(&in_empty[s], 4);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_pair(tmem_base_slot, kTmemCols);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();   // barrier inits + TMEM allocation of BOTH CTAs visible before any cross-CTA traffic
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs, one thread each) =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = pair_id; w < p.items; w += n_pairs) {
        const Item it = decode_item<BN>(w, p);
        const int row_a = it.m_blk * 256 + (int)rank * kBMcta;
        const int row_w = it.col0 + (int)rank * (it.bn / 2);
        const CUtensorMap* tw = it.bn == BN ? &tmap_w : &tmap_w_tail;   // box of bn / 2 rows of W per CTA
        const uint32_t bytes = 2u * (uint32_t)(S::kABytes + (it.bn / 2) * kBK * 2);
        for (int kb = 0; kb < num_kb; ++kb) {
          wait_bounded(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * S::kStageBytes;
          uint8_t* sb = sa + S::kABytes;
          if (rank == 0) mbar_expect_tx(&full_bar[stage], bytes);
          tma_load_2d_pair(&tmap_a, &full_bar[stage], sa, kb * kBK, row_a);
          tma_load_2d_pair(tw, &full_bar[stage], sb, kb * kBK, row_w);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA, one thread) =====================
    if (rank == 0 && lane == 0) {
      // instruction descriptor: D=f32, A=B=bf16 (format 1) or fp16 (format 0), both K-major, N = BN, M = 256 (pair)
      const uint32_t idesc0 = (1u << 4) | (p.ab_f16 ? 0u : ((1u << 7) | (1u << 10))) | ((uint32_t)(256 >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int w = pair_id; w < p.items; w += n_pairs) {
        const uint32_t idesc = idesc0 | ((uint32_t)(decode_item<BN>(w, p).bn >> 3) << 17);   // N = bn
        wait_bounded(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          wait_bounded(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * S::kStageBytes);
          const uint32_t sb = sa + S::kABytes;
#pragma unroll
          for (int k = 0; k < kBK / kUmmaK; ++k) {
            const uint64_t da = smem_desc_sw128(sa + k * kUmmaK * 2);
            const uint64_t db = smem_desc_sw128(sb + k * kUmmaK * 2);
            umma_bf16_pair(d_tmem, da, db, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit_pair(&empty_bar[stage]);   // both CTAs' smem slot reusable once these MMAs retire
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit_pair(&tmem_full[as]);        // accumulators ready in both CTAs
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else if (warp == 3) {
    // ===================== epilogue operand prefetch (residual / aux), both CTAs =====================
    if (kHasIn && lane == 0) {
      int slot = 0;
      uint32_t phase = 0;
      for (int w = pair_id; w < p.items; w += n_pairs) {
        const Item it = decode_item<BN>(w, p);
        const int row0 = it.m_blk * 256 + (int)rank * kBMcta;
        for (int c = 0; c < it.bn / 32; ++c) {
          const int col0 = it.col0 + c * 32;
          if (col0 >= p.N) break;
          wait_bounded(&in_empty[slot], phase ^ 1);
          mbar_expect_tx(&in_full[slot], 128 * kInRowBytes);
          tma_load_2d(&tmap_in, &in_full[slot], smem + S::kInOffset + slot * S::kInSlotBytes, col0, row0);
          if (++slot == kIn) { slot = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue warps (TMEM -> regs -> swizzled smem -> TMA store) =====================
    const int q = warp & 3;                 // TMEM lane quarter this warp may access
    const int par = (warp - 4) >> 2;        // this warp takes the chunks with c % 2 == par
    const int r_cta = q * 32 + lane;        // row inside this CTA's 128-row half
    uint8_t* oslot = smem + S::kOutOffset + (warp - 4) * kOutSlotBytes;
    int as = 0;
    uint32_t aphase = 0;
    uint32_t gbase = 0;                     // chunks the prefetch warp has issued before this tile (ring position)
    for (int w = pair_id; w < p.items; w += n_pairs) {
      const Item it = decode_item<BN>(w, p);
      const int row0 = it.m_blk * 256 + (int)rank * kBMcta + q * 32;   // first row of this warp's 32-row box
      wait_bounded(&tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN);
      const int tile_chunks = max(0, min(it.bn / 32, (p.N - it.col0 + 31) >> 5));
#pragma unroll 1
      for (int c = par; c < tile_chunks; c += 2) {
        const int col0 = it.col0 + c * 32;
        uint32_t r[32];
        tmem_ld_32x32(t_row + (uint32_t)(c * 32), r);
        float v[32];
        tmem_ld_wait();
        if (p.bias) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
            if (col0 + 4 * i < p.N) b = __ldg(reinterpret_cast<const float4*>(p.bias + col0) + i);  // N % 8 == 0
            v[4 * i] = fmaf(p.alpha, __uint_as_float(r[4 * i]), b.x);
            v[4 * i + 1] = fmaf(p.alpha, __uint_as_float(r[4 * i + 1]), b.y);
            v[4 * i + 2] = fmaf(p.alpha, __uint_as_float(r[4 * i + 2]), b.z);
            v[4 * i + 3] = fmaf(p.alpha, __uint_as_float(r[4 * i + 3]), b.w);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = p.alpha * __uint_as_float(r[i]);
        }
        if (kHasIn) {
          const uint32_t g = gbase + (uint32_t)c;              // position of this chunk in the prefetch ring
          const int islot = (int)(g % (kIn > 0 ? kIn : 1));
          wait_bounded(&in_full[islot], (g / (kIn > 0 ? kIn : 1)) & 1u);
          const uint8_t* in = smem + S::kInOffset + islot * S::kInSlotBytes;
          if (MODE == kResidual) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 f = *reinterpret_cast<const float4*>(in + sw_chunk<128>(r_cta, j));
              v[4 * j] += f.x; v[4 * j + 1] += f.y; v[4 * j + 2] += f.z; v[4 * j + 3] += f.w;
            }
          } else {  // QuickGELU'(aux) * acc
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint4 u = *reinterpret_cast<const uint4*>(in + sw_chunk<64>(r_cta, j));
              const uint32_t* h = reinterpret_cast<const uint32_t*>(&u);
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float2 f = unpack16(h[e], p.aux_f16);
                v[8 * j + 2 * e] *= quickgelu_grad_fast(f.x);
                v[8 * j + 2 * e + 1] *= quickgelu_grad_fast(f.y);
              }
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&in_empty[islot]);
        }
        // this warp's staging slot must have been read by the TMA store of its previous chunk
        if (lane == 0) bulk_wait_read<0>();
        __syncwarp();
        if (MODE == kGelu) {
          if (p.has_out2) {  // pre-activation copy for the backward (second half of the slot)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint4 u;
              uint32_t* h = reinterpret_cast<uint32_t*>(&u);
#pragma unroll
              for (int e = 0; e < 4; ++e) h[e] = pack16(v[8 * j + 2 * e], v[8 * j + 2 * e + 1], p.out2_f16);
              *reinterpret_cast<uint4*>(oslot + 2048 + sw_chunk<64>(lane, j)) = u;
            }
          }
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = quickgelu_fast(v[i]);
        }
        if (OUT_BYTES == 4) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(oslot + sw_chunk<128>(lane, j)) =
                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 u;
            uint32_t* h = reinterpret_cast<uint32_t*>(&u);
#pragma unroll
            for (int e = 0; e < 4; ++e) h[e] = pack16(v[8 * j + 2 * e], v[8 * j + 2 * e + 1], p.out_f16);
            *reinterpret_cast<uint4*>(oslot + sw_chunk<kOutRowBytes>(lane, j)) = u;
          }
        }
        fence_async_proxy();   // generic-proxy smem writes -> visible to the TMA (async proxy)
        __syncwarp();
        if (lane == 0) {
          if (MODE == kResidualInPlace) tma_reduce_add_2d(&tmap_out, oslot, col0, row0);
          else tma_store_2d(&tmap_out, oslot, col0, row0);
          if (MODE == kGelu && p.has_out2) tma_store_2d(&tmap_out2, oslot + 2048, col0, row0);
          bulk_commit();
        }
      }
      gbase += (uint32_t)tile_chunks;
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_leader(&tmem_empty[as]);   // release this accumulator stage to the leader's MMA
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    if (lane == 0) bulk_wait_read<0>();   // staging smem must outlive the last stores' reads
  }

  tc_fence_before();
  cluster_sync_all();   // no CTA of the pair exits (or frees TMEM) while the other may still signal it
  tc_fence_after();
  if (warp == 2) tmem_dealloc_pair(tmem_base, kTmemCols);
}

// ---------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

// rows x cols row-major matrix (row stride ld elements) -> boxes of box_rows x box_cols
static int make_map(CUtensorMap* map, const void* base, int dtype, long long rows, long long cols, long long ld,
                    int box_cols, int box_rows, CUtensorMapSwizzle sw) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return PSRC_ERR_CUDA;
  }
  const int esz = dtype == PSRC_F32 ? 4 : 2;
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * esz};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  const CUtensorMapDataType tdt = dtype == PSRC_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                  : dtype == PSRC_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUresult r = fn(map, tdt, 2,
                  const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed: CUresult %d (rows=%lld cols=%lld ld=%lld box=%dx%d esz=%d)", (int)r,
              rows, cols, ld, box_rows, box_cols, esz);
    return PSRC_ERR_CUDA;
  }
  return PSRC_OK;
}

static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// How finely to cut the tiles of the last, partially filled round (rem = tiles % pairs of them) along N.  A 256 x bn
// item re-reads the whole 256-row A tile, so its cost is not bn / BN of a tile: the shared-memory-bandwidth model of
// DESIGN 4 (A and W both written by TMA and read by the MMA) gives (32 + 2 * 16 * bn / 256) / 64 of a full tile per
// k-block: 0.75 at bn = 128, 0.625 at bn = 64, plus the fixed per-item prologue / epilogue.  Pick the split whose
// extra rounds cost least; PSRC_PAIR_SPLIT = 1 / 2 / 4 forces it (tuning knob).
static int tail_split(int tiles, int pairs, int BN) {
  static const int force = getenv("PSRC_PAIR_SPLIT") ? atoi(getenv("PSRC_PAIR_SPLIT")) : 0;
  const int rem = tiles % pairs;
  if (rem == 0) return 1;
  if (force == 1 || force == 2 || force == 4) return (BN / force >= 64) ? force : 1;
  const float cost[3] = {1.0f, 0.80f, 0.70f};
  int best = 1;
  float best_cost = 1.0f;   // split 1: one more full round
  for (int i = 1; i < 3; ++i) {
    const int sp = 1 << i;
    if (BN / sp < 64) break;
    const float c = (float)((rem * sp + pairs - 1) / pairs) * cost[i];
    if (c < best_cost - 0.05f) {
      best_cost = c;
      best = sp;
    }
  }
  return best;
}

template <int BN, int STAGES, int MODE, int OUT_BYTES>
static int launch(const psrc_gemm_args& a, cudaStream_t stream) {
  using S = Smem<BN, STAGES, (MODE == kGeluGrad || MODE == kResidual) ? kInSlots : 0, (MODE == kResidual) ? 128 * 128 : 128 * 64>;
  static_assert(S::kDyn <= 227 * 1024, "shared memory budget exceeded");
  auto kernel = gemm_pair_kernel<BN, STAGES, MODE, OUT_BYTES>;
  static bool attr_set = false;
  if (!attr_set) {
    PSRC_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kDyn));
    attr_set = true;
  }
  Params p;
  p.M = a.M; p.N = a.N; p.K = a.K;
  p.tiles_m = (a.M + 255) / 256;
  p.tiles_n = (a.N + BN - 1) / BN;
  const int tiles = p.tiles_m * p.tiles_n, max_pairs = sm_count() / 2;
  p.split = tail_split(tiles, max_pairs, BN);
  const int rem = tiles % max_pairs;
  p.n_full = p.split > 1 ? tiles - rem : tiles;
  p.items = p.n_full + (tiles - p.n_full) * p.split;
  CUtensorMap ta, tw, twt, to, to2, tin;
  int rc;
  if ((rc = make_map(&ta, a.A, a.ab_dtype, a.M, a.K, a.lda, kBK, kBMcta, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = make_map(&tw, a.W, a.ab_dtype, a.N, a.K, a.ldw, kBK, BN / 2, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  twt = tw;
  if (p.split > 1)
    if ((rc = make_map(&twt, a.W, a.ab_dtype, a.N, a.K, a.ldw, kBK, BN / p.split / 2, CU_TENSOR_MAP_SWIZZLE_128B)))
      return rc;
  const CUtensorMapSwizzle osw = OUT_BYTES == 4 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
  if ((rc = make_map(&to, a.out, a.out_dtype, a.M, a.N, a.ldo, 32, 32, osw))) return rc;
  to2 = to;
  tin = to;
  if (MODE == kGelu && a.out2)
    if ((rc = make_map(&to2, a.out2, a.out2_dtype, a.M, a.N, a.ldo2, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B))) return rc;
  if (MODE == kResidual)
    if ((rc = make_map(&tin, a.residual, PSRC_F32, a.M, a.N, a.ldres, 32, 128, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if (MODE == kGeluGrad)
    if ((rc = make_map(&tin, a.aux, a.aux_dtype, a.M, a.N, a.ldaux, 32, 128, CU_TENSOR_MAP_SWIZZLE_64B))) return rc;
  p.bias = a.bias;
  p.alpha = a.alpha;
  p.has_out2 = (MODE == kGelu && a.out2) ? 1 : 0;
  p.ab_f16 = a.ab_dtype == PSRC_F16;
  p.out_f16 = a.out_dtype == PSRC_F16;
  p.out2_f16 = a.out2_dtype == PSRC_F16;
  p.aux_f16 = a.aux_dtype == PSRC_F16;
  const int pairs = std::min(p.items, max_pairs);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = S::kDyn;
  cfg.stream = stream;
  PSRC_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kernel, ta, tw, twt, to, to2, tin, p));
  count_launch();
  return PSRC_OK;
}

template <int MODE, int OUT_BYTES>
static int launch_bn(const psrc_gemm_args& a, cudaStream_t stream) {
  // 256 x 256 tiles always (measured: UMMA 256x128 is shared-memory-bandwidth bound, 830 vs 1230 TFLOP/s on the
  // QKV shape); 256 x 128 only when N itself is that narrow.  PSRC_PAIR_BN is a tuning knob.
  static const int force_bn = getenv("PSRC_PAIR_BN") ? atoi(getenv("PSRC_PAIR_BN")) : 0;
  if (force_bn == 128 || (force_bn == 0 && a.N <= 128)) return launch<128, 5, MODE, OUT_BYTES>(a, stream);
  // 256 x 192 tiles were measured for the N = 768 shapes (303 tiles = 4.09 -> 5 rounds on 74 pairs): 105 vs 98 us at
  // K = 3072 -- the narrower tile costs 0.9 of the wide one, so the extra round is not recovered.
  // epilogues without a residual / aux operand: the 64 KB of the prefetch ring become two more pipeline stages
  static const int deep = getenv("PSRC_PAIR_DEEP") ? atoi(getenv("PSRC_PAIR_DEEP")) : 1;
  if constexpr (MODE == kNone || MODE == kGelu || MODE == kResidualInPlace) {
    if (deep) return launch<256, 6, MODE, OUT_BYTES>(a, stream);
  }
  if constexpr (MODE == kGeluGrad) {   // bf16 aux chunks are half the size of fp32 residual chunks: one more stage
    if (deep) return launch<256, 5, MODE, OUT_BYTES>(a, stream);
  }
  return launch<256, 4, MODE, OUT_BYTES>(a, stream);
}

}  // namespace pair

static bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// which fused-epilogue combinations the CTA-pair kernel implements (everything the training step launches)
static bool is16(int dtype) { return dtype == PSRC_BF16 || dtype == PSRC_F16; }

static int pair_mode(const psrc_gemm_args& a) {
  if (a.act == PSRC_ACT_NONE && !a.residual && !a.out2) return pair::kNone;
  if (a.act == PSRC_ACT_QUICKGELU && !a.residual && is16(a.out_dtype) && (!a.out2 || is16(a.out2_dtype)))
    return pair::kGelu;
  if (a.act == PSRC_ACT_QUICKGELU_GRAD && !a.residual && !a.out2 && is16(a.aux_dtype) && is16(a.out_dtype))
    return pair::kGeluGrad;
  if (a.act == PSRC_ACT_NONE && a.residual && !a.out2 && a.out_dtype == PSRC_F32)
    return (a.residual == a.out && a.ldres == a.ldo) ? pair::kResidualInPlace : pair::kResidual;
  return -1;
}

bool gemm_pair_supported(const psrc_gemm_args& a) {
  if (!gemm_tc_supported(a) || pair_mode(a) < 0) return false;
  auto ok = [](const void* p, long long ld, int dtype) {
    if (!p) return true;
    const int per16 = dtype == PSRC_F32 ? 4 : 8;
    return al16(p) && (ld % per16 == 0);
  };
  return a.N % 8 == 0 && ok(a.bias, 4, PSRC_F32) && ok(a.aux, a.ldaux, a.aux_dtype) &&
         ok(a.residual, a.ldres, PSRC_F32) && ok(a.out, a.ldo, a.out_dtype) && ok(a.out2, a.ldo2, a.out2_dtype);
}

int gemm_pair(const psrc_gemm_args& a, cudaStream_t stream) {
  switch (pair_mode(a)) {
    case pair::kNone:
      return a.out_dtype == PSRC_F32 ? pair::launch_bn<pair::kNone, 4>(a, stream)
                                     : pair::launch_bn<pair::kNone, 2>(a, stream);
    case pair::kGelu: return pair::launch_bn<pair::kGelu, 2>(a, stream);
    case pair::kGeluGrad: return pair::launch_bn<pair::kGeluGrad, 2>(a, stream);
    case pair::kResidual: return pair::launch_bn<pair::kResidual, 4>(a, stream);
    case pair::kResidualInPlace: return pair::launch_bn<pair::kResidualInPlace, 4>(a, stream);
  }
  set_error("gemm_pair: unsupported epilogue combination");
  return PSRC_ERR_UNSUPPORTED;
}

}  // namespace psrc
