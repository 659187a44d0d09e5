// Dev tool (GPU box): clock64 timeline of one (sequence, head) of the chunk-pipelined attention backward at the cfg-3 vision shape.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DPSRC_ATTN_TIMELINE -Iinclude -Ifew-shot-prompt-learning-for-vision-language-models-in-imbalanced-datasets_b200/csrc \
//        tests/attn_timeline.cu few-shot-prompt-learning-for-vision-language-models-in-imbalanced-datasets_b200/csrc/core.cu -o devbin/attn_timeline -lcuda
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "attention_tc.cu"

int main() {
  const int n_seq = 128, L = 201, H = 12, D = H * 64;
  const size_t n_qkv = (size_t)n_seq * L * 3 * D, n_o = (size_t)n_seq * L * D;
  std::vector<__nv_bfloat16> h(n_qkv);
  srand(1);
  for (auto& v : h) v = __float2bfloat16((rand() / (float)RAND_MAX - 0.5f) * 2.f);
  __nv_bfloat16 *qkv, *out, *dout, *dqkv;
  float* lse;
  cudaMalloc(&qkv, n_qkv * 2); cudaMalloc(&dqkv, n_qkv * 2); cudaMalloc(&out, n_o * 2); cudaMalloc(&dout, n_o * 2);
  cudaMalloc(&lse, (size_t)n_seq * H * L * 4);
  cudaMemcpy(qkv, h.data(), n_qkv * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dout, h.data(), n_o * 2, cudaMemcpyHostToDevice);
  if (psrc::attn_fwd_tc(qkv, out, lse, n_seq, L, H, 0, 0)) { printf("fwd failed\n"); return 1; }
  for (int i = 0; i < 3; ++i)
    if (psrc::attn_bwd_tc(qkv, out, dout, lse, dqkv, n_seq, L, H, 0, 0)) { printf("bwd failed\n"); return 1; }
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("sync failed\n"); return 1; }
  long long tl[2][64];
  cudaMemcpyFromSymbol(tl, psrc::atc::g_tl, sizeof(tl));
  const long long t0 = tl[0][0];
  printf("mma : post-alloc-sync %lld, loaded %lld, done %lld clk\n", tl[0][1] - t0, tl[0][2] - t0, tl[0][20] - t0);
  printf("elem: prologue done %lld, end %lld clk\n", tl[1][0] - t0, tl[1][50] - t0);
  for (int i = 0; i < 16; ++i)
    printf("  step %2d: S seen %7lld  P arrived %7lld (elementwise %5lld)  P seen by mma %7lld | mma: SD(i+1) issued %7lld  out(i) issued %7lld\n",
           i, tl[1][1 + 2 * i] - t0, tl[1][2 + 2 * i] - t0, tl[1][2 + 2 * i] - tl[1][1 + 2 * i], tl[0][3 + i] - t0,
           tl[0][22 + i] - t0, tl[0][40 + i] - t0);
  for (int p = 0; p < 4; ++p) printf("  epilogue %d stores issued %7lld\n", p, tl[1][40 + p] - t0);
  return 0;
}
