// K7 rope_kv_write: rotary embedding on q and k (in place in the packed QKV activation) and
// append of k, v into the paged KV cache, one launch.
//
// Replaces the reference's ~10 ATen kernels for RoPE (gptj_modeling.py:26-47,
// 206-224: gather sin/cos by position, rotate_every_two, cat) and the per-step torch.cat
// KV append (G6; gptj_modeling.py:229-233, gpt_bigcode_modeling.py:288-289), which copies
// the whole cache every token.  Styles: GPT-J interleaved pairs on the first rotary_dim
// dims, NeoX/Llama half rotation, or none (GPT-2 / OPT / BigCode learned positions).
// cos/sin are evaluated in fp32 from the position (no table, no gather).
#include "common.cuh"

namespace {

// grid = (tokens, ceil(items / 128)): one 16-byte vector of [q | k | v] per thread, so a decode step
// (64 tokens) still spreads over hundreds of CTAs and pays one memory round trip instead of a
// latency-bound loop.  q vectors outside the rotary range are untouched and not even read.
template <typename T>
__global__ void __launch_bounds__(128) rope_kv_write_kernel(
    T* __restrict__ qkv, const int* __restrict__ positions, const int* __restrict__ slot_mapping,
    T* __restrict__ k_cache, T* __restrict__ v_cache, int64_t ld, int hq, int hkv, int d, int rotary_dim,
    int style, float log2_base, int q_off, int k_off, int v_off, int block_size) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float s_cos[128], s_sin[128];  // rotary_dim / 2 <= 128 frequencies
  const int t = blockIdx.x;
  const int slot = slot_mapping[t];
  T* row = qkv + (int64_t)t * ld;
  const int vec_per_head = d >> 3;
  const int nq = hq * vec_per_head, nk = hkv * vec_per_head;
  const int half = rotary_dim >> 1;
  const int it0 = blockIdx.y * 128;
  const int it = it0 + threadIdx.x;

  // cos/sin depend on (token, frequency) only: evaluated once per CTA that holds q / k vectors
  if (style != LLMSS_ROPE_NONE && it0 < nq + nk) {
    const float pos = (float)positions[t];
    const float inv_rd = 1.0f / (float)(rotary_dim > 0 ? rotary_dim : 1);
    for (int i = threadIdx.x; i < half; i += blockDim.x) {
      const float inv_freq = exp2f(-2.0f * (float)i * inv_rd * log2_base);
      sincosf(pos * inv_freq, &s_sin[i], &s_cos[i]);
    }
    __syncthreads();
  }
  int64_t cache_base = 0;
  if (slot >= 0) cache_base = ((int64_t)(slot / block_size) * hkv) * block_size * d + (int64_t)(slot % block_size) * d;

  if (it >= nq + nk) {
    // v heads: straight copy into the cache
    const int j = it - nq - nk;
    if (j < nk && slot >= 0) {
      const int h = j / vec_per_head, c = (j % vec_per_head) << 3;
      *reinterpret_cast<uint4*>(v_cache + cache_base + (int64_t)h * block_size * d + c) =
          *reinterpret_cast<const uint4*>(row + v_off + h * d + c);
    }
    return;
  }
  // q and k heads: items [0, nq) are q vectors, [nq, nq+nk) are k vectors
  const bool is_k = it >= nq;
  const int j = is_k ? it - nq : it;
  const int h = j / vec_per_head, c = (j % vec_per_head) << 3;  // column within head
  T* p = row + (is_k ? k_off : q_off) + h * d + c;
  T* kc = k_cache + cache_base + (int64_t)h * block_size * d + c;
  const bool to_cache = is_k && slot >= 0;
  float f[8];
  if (style == LLMSS_ROPE_GPTJ && c < rotary_dim) {
    unpack8<T>(*reinterpret_cast<const uint4*>(p), f);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int pair = (c >> 1) + i;  // frequency index
      const float sn = s_sin[pair], cs = s_cos[pair];
      const float x1 = f[2 * i], x2 = f[2 * i + 1];
      f[2 * i] = x1 * cs - x2 * sn;
      f[2 * i + 1] = x2 * cs + x1 * sn;
    }
    const uint4 pf = pack8<T>(f);
    *reinterpret_cast<uint4*>(p) = pf;
    if (to_cache) *reinterpret_cast<uint4*>(kc) = pf;
  } else if (style == LLMSS_ROPE_NEOX && c < rotary_dim) {
    if (c < half) {  // this thread also owns the second-half partner vector
      float g[8];
      unpack8<T>(*reinterpret_cast<const uint4*>(p), f);
      unpack8<T>(*reinterpret_cast<const uint4*>(p + half), g);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float sn = s_sin[c + i], cs = s_cos[c + i];
        const float x1 = f[i], x2 = g[i];
        f[i] = x1 * cs - x2 * sn;
        g[i] = x2 * cs + x1 * sn;
      }
      const uint4 pf = pack8<T>(f), pg = pack8<T>(g);
      *reinterpret_cast<uint4*>(p) = pf;
      *reinterpret_cast<uint4*>(p + half) = pg;
      if (to_cache) {
        *reinterpret_cast<uint4*>(kc) = pf;
        *reinterpret_cast<uint4*>(kc + half) = pg;
      }
    }
  } else if (to_cache) {
    *reinterpret_cast<uint4*>(kc) = *reinterpret_cast<const uint4*>(p);
  }
}

}  // namespace

LLMSS_API int llmss_rope_kv_write(void* qkv, const int* positions, const int* slot_mapping, void* k_cache,
                                  void* v_cache, int T, int64_t ld, int hq, int hkv, int d, int rotary_dim, int style,
                                  float base, int q_off, int k_off, int v_off, int block_size, int dtype,
                                  cudaStream_t stream) {
  LLMSS_REQUIRE(d % 8 == 0 && ld % 8 == 0 && q_off % 8 == 0 && k_off % 8 == 0 && v_off % 8 == 0,
                "rope_kv_write: head_dim/offsets/stride must be multiples of 8");
  LLMSS_REQUIRE(style == LLMSS_ROPE_NONE || (rotary_dim > 0 && rotary_dim <= d && rotary_dim <= 256 && rotary_dim % 16 == 0),
                "rope_kv_write: rotary_dim=%d must be a multiple of 16 and <= head_dim", rotary_dim);
  LLMSS_REQUIRE(dtype == LLMSS_F16 || dtype == LLMSS_BF16, "rope_kv_write: dtype must be f16/bf16");
  if (T == 0) return 0;
  const float log2_base = log2f(base);
  const int chunks = ((hq + 2 * hkv) * (d / 8) + 127) / 128;
  if (dtype == LLMSS_F16)
    LLMSS_CHECK_CUDA(launch_pdl(rope_kv_write_kernel<__half>, dim3(T, chunks), dim3(128), 0, stream, (__half*)qkv, positions, slot_mapping, (__half*)k_cache, (__half*)v_cache, ld, hq, hkv, d, rotary_dim, style, log2_base, q_off, k_off, v_off, block_size));
  else
    LLMSS_CHECK_CUDA(launch_pdl(rope_kv_write_kernel<__nv_bfloat16>, dim3(T, chunks), dim3(128), 0, stream, (__nv_bfloat16*)qkv, positions, slot_mapping, (__nv_bfloat16*)k_cache, (__nv_bfloat16*)v_cache, ld, hq, hkv, d, rotary_dim, style, log2_base, q_off, k_off, v_off, block_size));
  LLMSS_CHECK_CUDA(cudaGetLastError());
  return 0;
}
