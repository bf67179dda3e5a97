// K9 attn_decode_paged: single-query attention against the paged KV cache, split-KV.
//
// Replaces the reference's decode-step attention (fp32 upcast, cuBLAS
// batched QK^T, where-mask, softmax, cuBLAS PV on a [B,h,1,T] score tensor, plus the
// G17 mask buffers and the H2D mask copy of quirk Q18).  Decode attention is pure
// KV-bandwidth: every cached byte is read once per step, so the kernel is organised
// around 16-byte streaming loads (ld.global.nc, no L1 allocation) with 2*U loads in
// flight per lane, fp32 online softmax, and a split-KV grid so that small batches still
// fill 148 SMs.  One CTA = (split, kv head x q-head tile, sequence); GQA/MQA heads that
// share a KV head are processed from a single KV read (GT q heads per CTA).
//
// Layouts: q/out rows are strided views into the packed activation (ldq/ldo elements);
// caches are [num_blocks, hkv, block_size, D]; partials live in caller-provided scratch.
#include "common.cuh"

namespace {

constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;
constexpr int kStages = 4;  // cp.async ring depth per warp
constexpr int kRingBytes(int U) { return kWarps * kStages * U * 2 * 32 * 16; }

template <typename T, int D, int GT, int U>
__global__ void __launch_bounds__(kThreads) attn_decode_kernel(
    const T* __restrict__ q, int64_t ldq, const T* __restrict__ k_cache, const T* __restrict__ v_cache,
    const int* __restrict__ block_tables, int max_blocks, const int* __restrict__ seq_lens, T* __restrict__ out,
    int64_t ldo, float* __restrict__ part_o, float* __restrict__ part_ml, int hq, int hkv, int block_size,
    float scale_log2, int num_splits) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int LPT = D / 8;     // lanes per token row
  constexpr int TPW = 32 / LPT;  // tokens per warp-wide load
  static_assert(LPT >= 1 && LPT <= 32 && (32 % LPT) == 0, "head_dim must be 8..256, power of two");

  __shared__ float sm_acc[kWarps][GT][D];
  __shared__ float sm_ml[kWarps][GT][2];

  const int split = blockIdx.x, b = blockIdx.z;
  const int G = hq / hkv;
  const int g_tiles = (G + GT - 1) / GT;
  const int hk = blockIdx.y / g_tiles, g0 = (blockIdx.y % g_tiles) * GT;
  const int gn = min(GT, G - g0);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int grp = lane / LPT, lig = lane % LPT;  // token group within the warp / lane in group

  const int L = seq_lens[b];
  const int chunk = (L + num_splits - 1) / num_splits;
  const int t_begin = split * chunk;
  const int t_end = min(L, t_begin + chunk);

  // q fragment (pre-scaled into the log2 domain)
  float qf[GT][8];
#pragma unroll
  for (int g = 0; g < GT; ++g) {
    if (g < gn) {
      unpack8<T>(*reinterpret_cast<const uint4*>(q + (int64_t)b * ldq + (int64_t)(hk * G + g0 + g) * D + lig * 8), qf[g]);
#pragma unroll
      for (int i = 0; i < 8; ++i) qf[g][i] *= scale_log2;
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) qf[g][i] = 0.f;
    }
  }

  float m[GT], l[GT], acc[GT][8];
#pragma unroll
  for (int g = 0; g < GT; ++g) {
    m[g] = -INFINITY;
    l[g] = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[g][i] = 0.f;
  }

  const int* bt = block_tables + (int64_t)b * max_blocks;
  const int64_t head_stride = (int64_t)block_size * D;
  constexpr int TOK_PER_ITER = TPW * U;

  // cp.async ring: every lane stages its own 16-byte K and V slices for the next kStages-1
  // iterations in shared memory (no registers held by loads in flight, no cross-lane sharing, so a
  // per-thread cp.async.wait_group is the only synchronisation).  ~12 KB of requests stay in
  // flight per warp for the whole stream.
  extern __shared__ uint4 ring_raw[];
  uint4* ring = ring_raw + (size_t)warp * (kStages * U * 2 * 32) + lane;  // [stage][u][k|v][lane]
  auto issue = [&](int t0, int stage) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int t = t0 + u * TPW + grp;
      const bool ok = t < t_end;
      const int tt = ok ? t : t_begin;  // any valid token: with src-size 0 nothing is read, zeros are written
      const int blk = __ldg(bt + tt / block_size);
      const int64_t off = ((int64_t)blk * hkv + hk) * head_stride + (int64_t)(tt % block_size) * D + lig * 8;
      const uint32_t dst = smem_u32(ring + (size_t)((stage * U + u) * 2) * 32);
      const int nbytes = ok ? 16 : 0;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(k_cache + off), "r"(nbytes) : "memory");
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + 512), "l"(v_cache + off), "r"(nbytes) : "memory");
    }
  };
  const int t_first = t_begin + warp * TOK_PER_ITER;
  const int t_stride = kWarps * TOK_PER_ITER;
#pragma unroll
  for (int st = 0; st < kStages - 1; ++st) {
    if (t_first + st * t_stride < t_end) issue(t_first + st * t_stride, st);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  int stage = 0;
  for (int t0 = t_first; t0 < t_end; t0 += t_stride) {
    {
      const int t_pre = t0 + (kStages - 1) * t_stride;
      int st_pre = stage + kStages - 1;
      if (st_pre >= kStages) st_pre -= kStages;
      if (t_pre < t_end) issue(t_pre, st_pre);
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group %0;" ::"n"(kStages - 1) : "memory");
    }
    uint4 kraw[U], vraw[U];
    bool valid[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      valid[u] = (t0 + u * TPW + grp) < t_end;
      kraw[u] = ring[(size_t)((stage * U + u) * 2) * 32];
      vraw[u] = ring[(size_t)((stage * U + u) * 2 + 1) * 32];
    }
    if (++stage == kStages) stage = 0;
    float s[U][GT];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float kf[8];
      unpack8<T>(kraw[u], kf);
#pragma unroll
      for (int g = 0; g < GT; ++g) {
        float p = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) p = fmaf(qf[g][i], kf[i], p);
        s[u][g] = p;
      }
    }
    // reduce partial dot products across the LPT lanes of each token group
#pragma unroll
    for (int o = LPT / 2; o > 0; o >>= 1) {
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int g = 0; g < GT; ++g) s[u][g] += __shfl_xor_sync(0xffffffffu, s[u][g], o);
    }
#pragma unroll
    for (int g = 0; g < GT; ++g) {
      float mx = m[g];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!valid[u]) s[u][g] = -INFINITY;
        mx = fmaxf(mx, s[u][g]);
      }
      const float corr = (mx == -INFINITY) ? 1.f : exp2f(m[g] - mx);
      m[g] = mx;
      l[g] *= corr;
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[g][i] *= corr;
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const float p = valid[u] ? exp2f(s[u][g] - mx) : 0.f;
        s[u][g] = p;
        l[g] += p;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float vf[8];
      unpack8<T>(vraw[u], vf);
#pragma unroll
      for (int g = 0; g < GT; ++g)
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[g][i] = fmaf(s[u][g], vf[i], acc[g][i]);
    }
  }

  // merge the TPW token groups of the warp
#pragma unroll
  for (int o = LPT; o < 32; o <<= 1) {
#pragma unroll
    for (int g = 0; g < GT; ++g) {
      const float mo = __shfl_xor_sync(0xffffffffu, m[g], o);
      const float lo = __shfl_xor_sync(0xffffffffu, l[g], o);
      const float mn = fmaxf(m[g], mo);
      const float a = (m[g] == -INFINITY) ? 0.f : exp2f(m[g] - mn);
      const float c = (mo == -INFINITY) ? 0.f : exp2f(mo - mn);
      l[g] = l[g] * a + lo * c;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float ao = __shfl_xor_sync(0xffffffffu, acc[g][i], o);
        acc[g][i] = acc[g][i] * a + ao * c;
      }
      m[g] = mn;
    }
  }
  if (grp == 0) {
#pragma unroll
    for (int g = 0; g < GT; ++g) {
#pragma unroll
      for (int i = 0; i < 8; ++i) sm_acc[warp][g][lig * 8 + i] = acc[g][i];
      if (lig == 0) {
        sm_ml[warp][g][0] = m[g];
        sm_ml[warp][g][1] = l[g];
      }
    }
  }
  __syncthreads();

  for (int it = threadIdx.x; it < gn * D; it += kThreads) {
    const int g = it / D, c = it % D;
    float mn = -INFINITY;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) mn = fmaxf(mn, sm_ml[w][g][0]);
    float lt = 0.f, at = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const float mw = sm_ml[w][g][0];
      const float sc = (mw == -INFINITY) ? 0.f : exp2f(mw - mn);
      lt += sm_ml[w][g][1] * sc;
      at += sm_acc[w][g][c] * sc;
    }
    const int head = hk * G + g0 + g;
    if (num_splits == 1) {
      out[(int64_t)b * ldo + (int64_t)head * D + c] = from_f32<T>(lt > 0.f ? at / lt : 0.f);
    } else {
      const int64_t pidx = ((int64_t)b * hq + head) * num_splits + split;
      part_o[pidx * D + c] = at;
      if (c == 0) {
        part_ml[pidx * 2] = mn;
        part_ml[pidx * 2 + 1] = lt;
      }
    }
  }
}

template <typename T, int D>
__global__ void attn_decode_combine_kernel(const float* __restrict__ part_o, const float* __restrict__ part_ml,
                                           T* __restrict__ out, int64_t ldo, int hq, int num_splits) {
  pdl_launch_dependents();
  pdl_wait();
  const int head = blockIdx.x, b = blockIdx.y;
  const int64_t base = ((int64_t)b * hq + head) * num_splits;
  float mn = -INFINITY;
  for (int s = 0; s < num_splits; ++s) mn = fmaxf(mn, part_ml[(base + s) * 2]);
  for (int c = threadIdx.x; c < D; c += blockDim.x) {
    float lt = 0.f, at = 0.f;
    for (int s = 0; s < num_splits; ++s) {
      const float ms = part_ml[(base + s) * 2];
      const float sc = (ms == -INFINITY) ? 0.f : exp2f(ms - mn);
      lt += part_ml[(base + s) * 2 + 1] * sc;
      at += part_o[(base + s) * D + c] * sc;
    }
    out[(int64_t)b * ldo + (int64_t)head * D + c] = from_f32<T>(lt > 0.f ? at / lt : 0.f);
  }
}

template <typename T, int D, int GT>
int launch_decode(const void* q, int64_t ldq, const void* kc, const void* vc, const int* bt, int max_blocks,
                  const int* seq_lens, void* out, int64_t ldo, float* part_o, float* part_ml, int B, int hq, int hkv,
                  int block_size, float scale, int num_splits, cudaStream_t stream) {
  constexpr int U = (D >= 256) ? 4 : 4;
  const int G = hq / hkv;
  const int g_tiles = (G + GT - 1) / GT;
  dim3 grid(num_splits, hkv * g_tiles, B);
  const float scale_log2 = scale * 1.4426950408889634f;
  static PerDeviceSmem configured;  // per device: the attribute does not carry over to a second GPU of the process
  LLMSS_CHECK_CUDA(configured.ensure(attn_decode_kernel<T, D, GT, U>, kRingBytes(U)));
  attn_decode_kernel<T, D, GT, U><<<grid, kThreads, kRingBytes(U), stream>>>(
      (const T*)q, ldq, (const T*)kc, (const T*)vc, bt, max_blocks, seq_lens, (T*)out, ldo, part_o, part_ml, hq, hkv,
      block_size, scale_log2, num_splits);
  LLMSS_CHECK_CUDA(cudaGetLastError());
  if (num_splits > 1) {
    attn_decode_combine_kernel<T, D><<<dim3(hq, B), (D < 128 ? D : 128), 0, stream>>>(part_o, part_ml, (T*)out, ldo, hq,
                                                                                    num_splits);
    LLMSS_CHECK_CUDA(cudaGetLastError());
  }
  return 0;
}

template <typename T, int D>
int dispatch_gt(int G, const void* q, int64_t ldq, const void* kc, const void* vc, const int* bt, int max_blocks,
                const int* seq_lens, void* out, int64_t ldo, float* part_o, float* part_ml, int B, int hq, int hkv,
                int block_size, float scale, int num_splits, cudaStream_t stream) {
#define LLMSS_GT(GT_)                                                                                              \
  return launch_decode<T, D, GT_>(q, ldq, kc, vc, bt, max_blocks, seq_lens, out, ldo, part_o, part_ml, B, hq, hkv, \
                                  block_size, scale, num_splits, stream)
  if (G == 1) LLMSS_GT(1);
  if (G == 2) LLMSS_GT(2);
  if (D <= 128) LLMSS_GT(4);
  LLMSS_GT(2);
#undef LLMSS_GT
}

}  // namespace

// Split-KV factor that fills whole waves: CTAs = B * hkv * g_tiles * splits should be close to (from
// below) a multiple of the resident-CTA capacity, with >= 64 tokens per split.
LLMSS_API int llmss_attn_decode_pick_splits(int B, int hq, int hkv, int d, int max_kv, int dtype) {
  const int G = hq / (hkv > 0 ? hkv : 1);
  const int GT = (G == 1) ? 1 : (G == 2) ? 2 : (d <= 128 ? 4 : 2);
  const int items = B * hkv * ((G + GT - 1) / GT);
  if (items <= 0) return 1;
  int occ = 4, sms = 148, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
#define LLMSS_OCC(T_, D_, GT_)                                                                                      \
  do {                                                                                                            \
    cudaFuncSetAttribute(attn_decode_kernel<T_, D_, GT_, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kRingBytes(4)); \
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, attn_decode_kernel<T_, D_, GT_, 4>, kThreads, kRingBytes(4));     \
  } while (0)
  if (dtype == LLMSS_F16) {
    if (d == 64) { if (GT == 1) LLMSS_OCC(__half, 64, 1); else if (GT == 2) LLMSS_OCC(__half, 64, 2); else LLMSS_OCC(__half, 64, 4); }
    else if (d == 128) { if (GT == 1) LLMSS_OCC(__half, 128, 1); else if (GT == 2) LLMSS_OCC(__half, 128, 2); else LLMSS_OCC(__half, 128, 4); }
    else { if (GT == 1) LLMSS_OCC(__half, 256, 1); else LLMSS_OCC(__half, 256, 2); }
  } else {
    if (d == 64) { if (GT == 1) LLMSS_OCC(__nv_bfloat16, 64, 1); else if (GT == 2) LLMSS_OCC(__nv_bfloat16, 64, 2); else LLMSS_OCC(__nv_bfloat16, 64, 4); }
    else if (d == 128) { if (GT == 1) LLMSS_OCC(__nv_bfloat16, 128, 1); else if (GT == 2) LLMSS_OCC(__nv_bfloat16, 128, 2); else LLMSS_OCC(__nv_bfloat16, 128, 4); }
    else { if (GT == 1) LLMSS_OCC(__nv_bfloat16, 256, 1); else LLMSS_OCC(__nv_bfloat16, 256, 2); }
  }
#undef LLMSS_OCC
  if (occ < 1) occ = 1;
  // Heuristic: the split that fills whole waves of resident CTAs (>= 64 tokens per split), with a small
  // penalty per extra split for the partial write + combine pass.  The engine refines this by timing the
  // candidates on the real shape when it captures a decode graph (ModelRunner.autotune_decode_splits).
  const double cap = (double)sms * occ;
  const int max_s = max_kv / 64 > 1 ? (max_kv / 64 < 32 ? max_kv / 64 : 32) : 1;
  int best = 1;
  double best_eff = 0.0;
  for (int sp = 1; sp <= max_s; ++sp) {
    const double waves = items * (double)sp / cap;
    const double full = waves < 1.0 ? 1.0 : (double)(long long)(waves + 0.999999);
    const double eff = waves / full - 0.004 * (sp - 1);
    if (eff > best_eff + 1e-9) { best_eff = eff; best = sp; }
  }
  return best;
}

// Merge the split-KV partials ([B, hq, splits, d] + [B, hq, splits, 2]) into `out`; shared with the
// tensor-core decode kernel (attn_decode_tc.cu).
LLMSS_API int llmss_attn_decode_combine(const float* part_o, const float* part_ml, void* out, int64_t ldo, int B, int hq,
                                        int d, int num_splits, int dtype, cudaStream_t stream) {
  LLMSS_REQUIRE(d == 64 || d == 128 || d == 256, "attn_decode_combine: head_dim=%d unsupported", d);
  if (B == 0) return 0;
  const dim3 grid(hq, B);
  const int threads = d < 128 ? d : 128;
#define LLMSS_C(T_, D_) \
  LLMSS_CHECK_CUDA(launch_pdl(attn_decode_combine_kernel<T_, D_>, grid, dim3(threads), 0, stream, part_o, part_ml, (T_*)out, ldo, hq, num_splits))
  if (dtype == LLMSS_F16) {
    if (d == 64) LLMSS_C(__half, 64); else if (d == 128) LLMSS_C(__half, 128); else LLMSS_C(__half, 256);
  } else if (dtype == LLMSS_BF16) {
    if (d == 64) LLMSS_C(__nv_bfloat16, 64); else if (d == 128) LLMSS_C(__nv_bfloat16, 128); else LLMSS_C(__nv_bfloat16, 256);
  } else {
    LLMSS_REQUIRE(false, "attn_decode_combine: dtype must be f16/bf16");
  }
#undef LLMSS_C
  LLMSS_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Scratch (floats) needed when num_splits > 1.
LLMSS_API int64_t llmss_attn_decode_scratch_floats(int B, int hq, int d, int num_splits) {
  return num_splits > 1 ? (int64_t)B * hq * num_splits * (d + 2) : 0;
}

LLMSS_API int llmss_attn_decode(const void* q, int64_t ldq, const void* k_cache, const void* v_cache,
                                const int* block_tables, int max_blocks, const int* seq_lens, void* out, int64_t ldo,
                                float* scratch, int B, int hq, int hkv, int d, int block_size, float scale,
                                int num_splits, int dtype, cudaStream_t stream) {
  LLMSS_REQUIRE(hq > 0 && hkv > 0, "attn_decode: head counts must be positive (hq=%d hkv=%d)", hq, hkv);
  LLMSS_REQUIRE(hq % hkv == 0, "attn_decode: hq=%d not a multiple of hkv=%d", hq, hkv);
  LLMSS_REQUIRE(d == 64 || d == 128 || d == 256, "attn_decode: head_dim=%d unsupported (64/128/256)", d);
  LLMSS_REQUIRE(dtype == LLMSS_F16 || dtype == LLMSS_BF16, "attn_decode: dtype must be f16/bf16");
  LLMSS_REQUIRE(num_splits >= 1 && (num_splits == 1 || scratch != nullptr), "attn_decode: scratch required for splits");
  LLMSS_REQUIRE(ldq % 8 == 0 && ldo % 8 == 0, "attn_decode: row strides must be multiples of 8");
  if (B == 0) return 0;
  float* part_o = scratch;
  float* part_ml = scratch ? scratch + (int64_t)B * hq * num_splits * d : nullptr;
  const int G = hq / hkv;
#define LLMSS_D(T_, D_)                                                                                           \
  return dispatch_gt<T_, D_>(G, q, ldq, k_cache, v_cache, block_tables, max_blocks, seq_lens, out, ldo, part_o, \
                             part_ml, B, hq, hkv, block_size, scale, num_splits, stream)
  if (dtype == LLMSS_F16) {
    if (d == 64) LLMSS_D(__half, 64);
    if (d == 128) LLMSS_D(__half, 128);
    LLMSS_D(__half, 256);
  } else {
    if (d == 64) LLMSS_D(__nv_bfloat16, 64);
    if (d == 128) LLMSS_D(__nv_bfloat16, 128);
    LLMSS_D(__nv_bfloat16, 256);
  }
#undef LLMSS_D
}

// ---------------------------------------------------------------------------------------
// llmss_attn_decode_static: the same split-KV decode attention, for callers that have no way
// to give us scratch memory and must not trigger an allocation.
//
// Who that is: (a) a server that has already committed all of HBM to weights and KV-cache
// pages, where one more cudaMalloc can fail or fragment the pool; (b) a latency-critical decode
// loop, because cudaMalloc / cudaFree synchronise the whole device and would stall every stream;
// (c) a plain C / C++ host (no torch allocator behind it) that only owns its input and output
// buffers.  For them the split-KV partials live in a fixed array that is part of the library
// image itself: every GPU that loads libllmss_b200.so carries these 8 MiB until the process
// exits, whether or not this entry point is ever called.
//
// Rules of use:
//   * Limit: B * hq * num_splits * (d + 2) floats must fit in kStaticScratchFloats (8 MiB);
//     larger requests are refused with an error, nothing is launched.  num_splits == 1 needs no
//     scratch and is always accepted.
//   * The first call on each device looks up the array's device address (one
//     cudaGetSymbolAddress, which may also make the driver load the module).  Make that first
//     call during start-up, before any stream capture begins on that device.
//   * All calls on a device share the one array and it keeps whatever the previous call left
//     there.  Nothing in it has to start at zero: every slot the combine pass reads is written by
//     the first pass of the same call (empty splits write m = -inf, l = 0), so no clearing is
//     done.  Two calls whose kernels overlap on the GPU (different streams, or two host threads)
//     race on the array and corrupt each other; serialise them on one stream.
//   * No cudaMalloc, cudaFree, or allocating library routine is called on any path.
// ---------------------------------------------------------------------------------------
namespace {

constexpr size_t kStaticScratchFloats = (size_t)2 * 1024 * 1024;  // 8 MiB per GPU
static __device__ __align__(256) float g_static_scratch[kStaticScratchFloats];

constexpr int kMaxDevices = 64;
float* g_static_scratch_addr[kMaxDevices] = {};  // host-side cache of the per-device address

// The name `g_static_scratch` in host code is only a handle for the runtime, not a device
// pointer; the usable address has to come from cudaGetSymbolAddress with the device current.
int static_scratch_for_current_device(float** out) {
  int dev = -1;
  LLMSS_CHECK_CUDA(cudaGetDevice(&dev));
  LLMSS_REQUIRE(dev >= 0 && dev < kMaxDevices, "attn_decode_static: device ordinal %d out of range", dev);
  float* p = __atomic_load_n(&g_static_scratch_addr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    void* sym = nullptr;
    LLMSS_CHECK_CUDA(cudaGetSymbolAddress(&sym, g_static_scratch));
    LLMSS_REQUIRE(sym != nullptr, "attn_decode_static: cudaGetSymbolAddress returned a null address");
    p = static_cast<float*>(sym);
    __atomic_store_n(&g_static_scratch_addr[dev], p, __ATOMIC_RELEASE);  // same value from any thread
  }
  *out = p;
  return 0;
}

}  // namespace

LLMSS_API int64_t llmss_attn_decode_static_max_floats() { return (int64_t)kStaticScratchFloats; }

LLMSS_API int llmss_attn_decode_static(const void* q, int64_t ldq, const void* k_cache, const void* v_cache,
                                       const int* block_tables, int max_blocks, const int* seq_lens, void* out,
                                       int64_t ldo, int B, int hq, int hkv, int d, int block_size, float scale,
                                       int num_splits, int dtype, cudaStream_t stream) {
  LLMSS_REQUIRE(B >= 0 && hq > 0 && hkv > 0 && d > 0 && num_splits >= 1,
                "attn_decode_static: bad sizes B=%d hq=%d hkv=%d d=%d splits=%d", B, hq, hkv, d, num_splits);
  const int64_t need = llmss_attn_decode_scratch_floats(B, hq, d, num_splits);
  LLMSS_REQUIRE(need <= (int64_t)kStaticScratchFloats,
                "attn_decode_static: B=%d x hq=%d x splits=%d x (d=%d + 2) = %lld floats of scratch exceed the built-in "
                "%lld (8 MiB); lower num_splits or use llmss_attn_decode with caller-provided scratch",
                B, hq, num_splits, d, (long long)need, (long long)kStaticScratchFloats);
  float* scratch = nullptr;
  if (num_splits > 1) {
    if (int e = static_scratch_for_current_device(&scratch)) return e;
  }
  return llmss_attn_decode(q, ldq, k_cache, v_cache, block_tables, max_blocks, seq_lens, out, ldo, scratch, B, hq, hkv,
                           d, block_size, scale, num_splits, dtype, stream);
}
