// K10 embed: token (+ learned position) embedding gather.
//
// The reference shards the table over ranks, remaps out-of-shard ids to a zero row and
// all-reduces [B,S,H] (SURVEY.md G1 + N2; layers.py:182-214; BigCode does it twice,
// gpt_bigcode_modeling.py:696-697).  With 180 GB of HBM per GPU the tables (<= 0.8 GB)
// are simply replicated, which removes 1-2 collectives per forward; the sharded + reduce
// variant remains available through TensorParallelEmbedding for API parity.
#include "common.cuh"

namespace {
template <typename T>
__global__ void __launch_bounds__(128) embed_kernel(const int* __restrict__ tokens, const int* __restrict__ positions,
                                                    const T* __restrict__ wte, const T* __restrict__ wpe,
                                                    T* __restrict__ out, int H, int vocab, int max_pos, int pos_offset,
                                                    int64_t ldo) {
  pdl_launch_dependents();
  pdl_wait();
  const int t = blockIdx.x;
  int tok = tokens[t];
  tok = min(max(tok, 0), vocab - 1);
  const T* src = wte + (int64_t)tok * H;
  const T* psrc = nullptr;
  if (wpe != nullptr) {
    int p = positions[t] + pos_offset;
    p = min(max(p, 0), max_pos - 1);
    psrc = wpe + (int64_t)p * H;
  }
  for (int c = threadIdx.x * 8; c < H; c += blockDim.x * 8) {
    uint4 raw = *reinterpret_cast<const uint4*>(src + c);
    if (psrc != nullptr) {
      float a[8], b[8];
      unpack8<T>(raw, a);
      unpack8<T>(*reinterpret_cast<const uint4*>(psrc + c), b);
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] += b[i];
      raw = pack8<T>(a);
    }
    *reinterpret_cast<uint4*>(out + (int64_t)t * ldo + c) = raw;
  }
}
}  // namespace

LLMSS_API int llmss_embed(const int* tokens, const int* positions, const void* wte, const void* wpe, void* out, int T,
                          int H, int vocab, int max_pos, int pos_offset, int64_t ldo, int dtype,
                          cudaStream_t stream) {
  LLMSS_REQUIRE(H % 8 == 0 && ldo % 8 == 0, "embed: H and ldo must be multiples of 8");
  LLMSS_REQUIRE(dtype == LLMSS_F16 || dtype == LLMSS_BF16, "embed: dtype must be f16/bf16");
  if (T == 0) return 0;
  if (dtype == LLMSS_F16)
    LLMSS_CHECK_CUDA(launch_pdl(embed_kernel<__half>, dim3(T), dim3(128), 0, stream, tokens, positions, (const __half*)wte, (const __half*)wpe, (__half*)out, H, vocab, max_pos, pos_offset, ldo));
  else
    LLMSS_CHECK_CUDA(launch_pdl(embed_kernel<__nv_bfloat16>, dim3(T), dim3(128), 0, stream, tokens, positions, (const __nv_bfloat16*)wte, (const __nv_bfloat16*)wpe, (__nv_bfloat16*)out, H, vocab, max_pos, pos_offset, ldo));
  LLMSS_CHECK_CUDA(cudaGetLastError());
  return 0;
}
