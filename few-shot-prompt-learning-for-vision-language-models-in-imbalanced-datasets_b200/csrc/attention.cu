// C-ABI entry points of the attention core; picks the tcgen05 kernels (bf16 fast mode, 32 < L <= 256), the
// mma.sync kernels (bf16, very short or very long sequences) or the CUDA-core kernels (fp32 validation mode).  See include/psrc_b200.h for the contract.
#include "common.cuh"

namespace psrc {
int attn_fwd_simt(const void* qkv, void* out, float* lse, int dtype, int n_seq, int L, int H, int causal,
                  cudaStream_t s);
int attn_bwd_simt(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, int dtype,
                  int n_seq, int L, int H, int causal, cudaStream_t s);
int attn_fwd_mma(const void* qkv, void* out, float* lse, int n_seq, int L, int H, int causal, cudaStream_t s);
int attn_bwd_mma(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, int n_seq,
                 int L, int H, int causal, cudaStream_t s);
bool attn_tc_supported(int L);
bool attn_small_mma_supported(int L);
int attn_fwd_small_mma(const void* qkv, void* out, float* lse, int dtype, int n_seq, int L, int H, int causal,
                       cudaStream_t s);
int attn_bwd_small_mma(const void* qkv, const void* dout, const float* lse, void* dqkv, int qkv_dtype, int n_seq, int L,
                       int H, int causal, cudaStream_t s);
int attn_fwd_tc(const void* qkv, void* out, float* lse, int n_seq, int L, int H, int causal, cudaStream_t s);
int attn_bwd_tc(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, int n_seq,
                int L, int H, int causal, cudaStream_t s);
}  // namespace psrc

using namespace psrc;

static bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
// bf16 fast mode: tcgen05 for 32 < L <= 256, one-warp mma.sync tiles for L <= 32 (one 16 x 16 score tile up to 16,
// 2 x 2 tiles up to 32), flash-style mma.sync otherwise
static int auto_backend(int L) {
  if (attn_tc_supported(L)) return 3;
  return attn_small_mma_supported(L) ? 5 : 2;
}

extern "C" int psrc_attn_fwd(const void* qkv, void* out, float* lse, int dtype, int n_seq, int L, int H,
                             int causal, int backend, psrc_stream_t stream) {
  PSRC_CHECK_ARG(n_seq >= 0 && L > 0 && H > 0, "psrc_attn_fwd: bad shape n_seq=%d L=%d H=%d", n_seq, L, H);
  if (n_seq == 0) return PSRC_OK;   // (empty tensors have NULL storage)
  PSRC_CHECK_ARG(qkv && out && lse, "psrc_attn_fwd: null pointer");
  PSRC_CHECK_ARG(dtype >= PSRC_F32 && dtype <= PSRC_F16, "psrc_attn_fwd: bad dtype");
  if (n_seq == 0) return PSRC_OK;
  if (dtype == PSRC_F16) {   // fp16 forward operands: the short-sequence tile kernels only (the truncated text tower)
    PSRC_CHECK_ARG((backend == 0 || backend == 5) && al16(qkv) && al16(out) && attn_small_mma_supported(L),
                   "psrc_attn_fwd: fp16 needs the mma.sync tile backend, aligned pointers and L <= 32 (L=%d)", L);
    return attn_fwd_small_mma(qkv, out, lse, dtype, n_seq, L, H, causal, as_stream(stream));
  }
  if (backend == 0)
    backend = (dtype == PSRC_BF16 && al16(qkv) && al16(out)) ? auto_backend(L) : 1;
  if (backend == 5) {
    PSRC_CHECK_ARG(dtype == PSRC_BF16 && al16(qkv) && al16(out) && attn_small_mma_supported(L),
                   "psrc_attn_fwd: mma.sync tile backend needs aligned bf16 and L <= 32 (L=%d)", L);
    return attn_fwd_small_mma(qkv, out, lse, dtype, n_seq, L, H, causal, as_stream(stream));
  }
  PSRC_CHECK_ARG(backend != 4, "psrc_attn_fwd: backend 4 (CUDA-core short-sequence kernel) was removed; use 5");
  if (backend == 2 || backend == 3) {
    PSRC_CHECK_ARG(dtype == PSRC_BF16 && al16(qkv) && al16(out), "psrc_attn_fwd: tensor-core backend needs aligned bf16");
    if (backend == 3) {
      PSRC_CHECK_ARG(attn_tc_supported(L), "psrc_attn_fwd: tcgen05 backend needs 32 < L <= 256 (L=%d)", L);
      return attn_fwd_tc(qkv, out, lse, n_seq, L, H, causal, as_stream(stream));
    }
    return attn_fwd_mma(qkv, out, lse, n_seq, L, H, causal, as_stream(stream));
  }
  return attn_fwd_simt(qkv, out, lse, dtype, n_seq, L, H, causal, as_stream(stream));
}

extern "C" int psrc_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                             int dtype, int n_seq, int L, int H, int causal, int backend, psrc_stream_t stream) {
  PSRC_CHECK_ARG(n_seq >= 0 && L > 0 && H > 0, "psrc_attn_bwd: bad shape n_seq=%d L=%d H=%d", n_seq, L, H);
  if (n_seq == 0) return PSRC_OK;
  PSRC_CHECK_ARG(qkv && out && dout && lse && dqkv, "psrc_attn_bwd: null pointer");
  PSRC_CHECK_ARG(dtype >= PSRC_F32 && dtype <= PSRC_F16, "psrc_attn_bwd: bad dtype");
  if (n_seq == 0) return PSRC_OK;
  const bool ok16 = al16(qkv) && al16(out) && al16(dout) && al16(dqkv);
  if (dtype == PSRC_F16) {   // qkv / out are fp16 (forward operands); dout / dqkv are bf16
    PSRC_CHECK_ARG((backend == 0 || backend == 5) && ok16 && attn_small_mma_supported(L),
                   "psrc_attn_bwd: fp16 needs the mma.sync tile backend, aligned pointers and L <= 32 (L=%d)", L);
    return attn_bwd_small_mma(qkv, dout, lse, dqkv, dtype, n_seq, L, H, causal, as_stream(stream));
  }
  if (backend == 0)
    backend = (dtype == PSRC_BF16 && ok16) ? auto_backend(L) : 1;
  if (backend == 5) {
    PSRC_CHECK_ARG(dtype == PSRC_BF16 && ok16 && attn_small_mma_supported(L),
                   "psrc_attn_bwd: mma.sync tile backend needs aligned bf16 and L <= 32 (L=%d)", L);
    return attn_bwd_small_mma(qkv, dout, lse, dqkv, dtype, n_seq, L, H, causal, as_stream(stream));
  }
  PSRC_CHECK_ARG(backend != 4, "psrc_attn_bwd: backend 4 (CUDA-core short-sequence kernel) was removed; use 5");
  if (backend == 2 || backend == 3) {
    PSRC_CHECK_ARG(dtype == PSRC_BF16 && ok16, "psrc_attn_bwd: tensor-core backend needs aligned bf16");
    if (backend == 3) {
      PSRC_CHECK_ARG(attn_tc_supported(L), "psrc_attn_bwd: tcgen05 backend needs 32 < L <= 256 (L=%d)", L);
      return attn_bwd_tc(qkv, out, dout, lse, dqkv, n_seq, L, H, causal, as_stream(stream));
    }
    return attn_bwd_mma(qkv, out, dout, lse, dqkv, n_seq, L, H, causal, as_stream(stream));
  }
  return attn_bwd_simt(qkv, out, dout, lse, dqkv, dtype, n_seq, L, H, causal, as_stream(stream));
}
