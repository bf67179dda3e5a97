// K5 / K1 / K2: tcgen05 GEMM for every linear layer, with the tensor-parallel all-reduce
// fused into the epilogue over NVLink peer memory.
//
//   out[M, N] = epilogue( x[M, K] @ w[N, K]^T )          (w in HF nn.Linear layout)
//
// Replaces the reference's cuBLAS `F.linear` call sites (SURVEY.md G3, G4, G9, G10, G12,
// G14; layers.py:62,123) and, for row-parallel layers, the separate blocking NCCL
// all-reduce that follows them (N1; layers.py:175-179).
//
// Design (B200-first, not a cuBLAS clone):
//  * swap-AB: the *weight* tile is the UMMA A operand (M=128 output features per tile) and
//    the activations are the B operand (N = 16..256 tokens).  Decode batches (1..256 tokens)
//    therefore use full-width tensor-core instructions with no padding waste, and the
//    accumulator sits in TMEM as [feature lane][token column] so bias is a per-thread scalar.
//  * TMA (cp.async.bulk.tensor, 128B swizzle) feeds a multi-stage mbarrier ring; the TMA and MMA
//    warps run their loops converged and issue from one elected lane (elect_one: operands stay in
//    uniform registers); accumulators are double-buffered in TMEM so the epilogue of one work
//    segment overlaps the main loop of the next.
//  * scheduling, chosen on the host (llmss_gemm_plan_query2):
//      - <= 74 tiles: cluster split-K kernel (gemm_cluster_kernel, DSMEM reduction);
//      - a grid that owns whole tiles exists near the SM count (tiles <= SMs, or tiles = 2 x grid):
//        that grid, no partial tiles;
//      - otherwise stream-K: the (tile x k-block) space is cut into gridDim.x equal contiguous
//        ranges; tiles touched by several CTAs are reduced through an L2-resident fp32 workspace by
//        the last arriver in CTA order (deterministic); >= 4 tiles per CTA: ranges cut at tile bounds.
//  * fused collectives (llmss_gemm_ex): all-reduce / reduce-scatter push finished tiles as
//    self-validating 16-byte vectors into the peers' Lamport area and sum them in rank order
//    (bit-identical on all ranks; tile i is on the wire while tile i+1 computes); all-gather mode
//    stages the ranks' activation shards through the bulk area while the own shard already computes.
//  * epilogues: bias, activation (optionally only for features >= act_start: GPT-J's fused
//    QKV+fc_in GEMM), SwiGLU gating on interleaved (gate, up) rows, residual add, fp32 out --
//    template-specialised per (flavour, activation), residual / partial loads prefetched per chunk.
#include "common.cuh"
#include "symm.cuh"

#include <cuda.h>
#include <stdlib.h>
#include <string.h>

int64_t llmss_nvls_min_bytes();     // comm.cu
int64_t llmss_twoshot_min_bytes();  // comm.cu

namespace {

constexpr int BM = 128;  // output features per tile (UMMA M)
constexpr int BK = 64;   // k-block: 64 x 16-bit = one 128-byte swizzle row
constexpr int kNumThreads = 192;  // warp 0 TMA, warp 1 MMA (+TMEM alloc), warps 2..5 epilogue

template <int BN>
struct Cfg {
  static constexpr int W_BYTES = BM * BK * 2;
  static constexpr int X_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = W_BYTES + X_BYTES;
  static constexpr int STAGES_RAW = (196 * 1024) / STAGE_BYTES;
  static constexpr int STAGES = STAGES_RAW > 12 ? 12 : STAGES_RAW;
  static constexpr int TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128 : (2 * BN <= 256) ? 256 : 512;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 512 /*barriers*/;
};

// Cluster split-K kernel: for token tiles up to 128 wide the CTAs PUSH their partial slices into the owner's shared
// memory (st.shared::cluster, posted, hidden behind the cluster barrier) instead of the owner pulling them afterwards
// (ld.shared::cluster: a ~1 us round trip on the critical path of a ~10 us kernel).  The receive area cannot alias
// the TMA ring -- a fast peer would write into stages that are still landing -- so the ring gives up that much.
template <int BN>
struct CCfg {
  static constexpr bool PUSH = BN <= 128;
  static constexpr int W_BYTES = BM * BK * 2;
  static constexpr int X_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = W_BYTES + X_BYTES;
  static constexpr int RECV_BYTES = PUSH ? (BM * BN * 4 + 8 * 32 * 32) : 0;   // one fp32 tile + per-CTA rounding slack
  static constexpr int STAGES_RAW = (196 * 1024 - RECV_BYTES) / STAGE_BYTES;
  static constexpr int STAGES = STAGES_RAW > 12 ? 12 : STAGES_RAW;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 512 /*barriers*/ + RECV_BYTES;
};

// all-gather mode: tensor maps over the gathered [M, K] activation in this rank's own bulk buffer, one per epoch
// parity (built on the host, copied to a caller-provided device scratch per call).
struct AgMaps {
  CUtensorMap m[2];
};

struct GemmArgs {
  int M, N, K;
  int tiles_n, tiles_m, kb_total;
  int dtype;  // LLMSS_F16 / LLMSS_BF16
  const void* bias;
  const void* residual;
  int64_t ldr;
  void* out;
  int64_t ldo;
  int out_f32;
  int act, act_start, gated;
  int epi;                // EPI_* flavour, derived from the flags above on the host
  int aligned;            // stream-K kernel: cut CTA ranges at tile boundaries (no partial tiles)
  float* sk_ws;           // stream-K partial slots: [2 * grid][BN][128] fp32 (first / last segment of a CTA)
  uint32_t* sk_counters;  // [tiles], zero between launches
  int ar;                 // fused collective over ws: 0 none, 1 all-reduce, 2 reduce-scatter (token rows)
  int rs_rows;            // ar == 2 / ag: token rows owned by each rank (M / nranks, multiple of BN)
  int ag;                 // x rows are sharded over the ranks' symmetric buffers (all-gather fused into the loads)
  const void* ag_src;     // ag: this rank's [rs_rows, K] shard (row stride ag_ld), staged into the symmetric buffer
  int64_t ag_ld;
  const CUtensorMap* ag_maps;  // ag: device array [2 parities] of maps over the gathered [M, K] activation (local)
  int tm_rot;                  // token-tile rotation: with ag, every rank starts on the tiles of its own shard
  unsigned long long* dbg;  // optional per-CTA phase timestamps [grid][8] (globaltimer ns), see llmss_gemm_set_debug
  SymmWs ws;
};
#define LLMSS_STAMP(slot_)                                                                   \
  do {                                                                                       \
    if (a.dbg != nullptr) a.dbg[(size_t)blockIdx.x * 8 + (slot_)] = globaltimer_ns();        \
  } while (0)

// ----- stream-K work decomposition (identical on every rank) -----
struct Segment {
  int tile, kb0, kb1;  // k-block range [kb0, kb1) of `tile`
  int first_cta, last_cta;  // CTAs contributing to this tile
};
__device__ __forceinline__ int64_t cta_start(int g, int64_t total, int G) { return ((int64_t)g * total) / G; }
__device__ __forceinline__ int cta_of(int64_t w, int64_t total, int G) { return (int)((((w + 1) * G) - 1) / total); }

struct SegIter {
  int64_t w, w_end, total;
  int kb_total, G;
  // aligned: CTA ranges are cut at tile boundaries (pure data-parallel, no partial tiles) -- used when
  // there are many tiles per CTA, where the one-tile imbalance is cheaper than the stream-K fix-up.
  __device__ SegIter(int g, int G_, int tiles, int kb_total_, bool aligned = false) : kb_total(kb_total_), G(G_) {
    total = (int64_t)tiles * kb_total_;
    if (aligned) {
      w = (((int64_t)g * tiles) / G_) * kb_total_;
      w_end = (((int64_t)(g + 1) * tiles) / G_) * kb_total_;
    } else {
      w = cta_start(g, total, G_);
      w_end = cta_start(g + 1, total, G_);
    }
  }
  __device__ bool next(Segment& s) {
    if (w >= w_end) return false;
    s.tile = (int)(w / kb_total);
    s.kb0 = (int)(w % kb_total);
    const int64_t tile_end = (int64_t)(s.tile + 1) * kb_total;
    const int64_t e = tile_end < w_end ? tile_end : w_end;
    s.kb1 = s.kb0 + (int)(e - w);
    s.first_cta = cta_of((int64_t)s.tile * kb_total, total, G);
    s.last_cta = cta_of(tile_end - 1, total, G);
    w = e;
    return true;
  }
};

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
__device__ __forceinline__ float4 ld_cg_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}


// ---- bias / activation / gating / residual + transposed store of 8 token columns of one feature ----
// The epilogue flavour is decided once on the host (GemmArgs::epi) and dispatched once per group, so
// the per-element code is a handful of instructions (an earlier fully generic per-element version
// was ~140 instructions per output and made the tail of the kernel instruction-bound).
enum { EPI_PLAIN = 0, EPI_ACT = 1, EPI_GATED = 2, EPI_F32 = 3 };

__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float act_fast(float x, int act) {
  if (act == LLMSS_ACT_GELU_TANH) return 0.5f * x * (1.0f + tanh_fast(0.7978845608028654f * (x + 0.044715f * x * x * x)));
  if (act == LLMSS_ACT_RELU) return fmaxf(x, 0.0f);
  if (act == LLMSS_ACT_SILU) return __fdividef(x, 1.0f + __expf(-x));
  if (act == LLMSS_ACT_GELU_ERF) return 0.5f * x * (1.0f + erff(x * 0.7071067811865475f));
  return x;
}

// Residual values of one 8-token group, fetched ahead of the accumulator so that their L2 latency is
// paid once per chunk of groups instead of once per group (the epilogue of the last tile is exposed).
template <typename T>
__device__ __forceinline__ void epilogue_load_res8(const GemmArgs& a, float (&rv)[8], int n, int mb, int lane, int m_limit) {
  const int rows = m_limit - mb;
  const int nn = (a.epi == EPI_GATED) ? (n >> 1) : n;
  const bool mine = (a.epi == EPI_GATED) ? ((lane & 1) == 0 && n < a.N) : (n < a.N);
#pragma unroll
  for (int j = 0; j < 8; ++j) rv[j] = 0.f;
  if (a.residual == nullptr || !mine) return;
  const T* r = reinterpret_cast<const T*>(a.residual) + (int64_t)mb * a.ldr + nn;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (j < rows) rv[j] = to_f32(r[(int64_t)j * a.ldr]);
}

// The epilogue of a tile runs once or twice per CTA, so every instruction it executes is an instruction-
// cache miss: what matters is how much code is *executed*, not how much exists.  The flavour (plain /
// activation / gated / fp32) and the activation are therefore template parameters selected once per chunk
// of CH groups; inside, the code is straight-line and branch-free (a runtime `switch` per element made
// the GELU epilogue of the fused QKV+up GEMM cost 25 us of a 70 us kernel).
constexpr int ACT_RUNTIME = -1;  // rare combinations: decided per element by act_fast

template <int ACT>
__device__ __forceinline__ float act_t(float x, int runtime_act) {
  if constexpr (ACT == LLMSS_ACT_NONE) return x;
  else if constexpr (ACT == LLMSS_ACT_GELU_TANH) return 0.5f * x * (1.0f + tanh_fast(0.7978845608028654f * (x + 0.044715f * x * x * x)));
  else if constexpr (ACT == LLMSS_ACT_SILU) return __fdividef(x, 1.0f + __expf(-x));
  else return act_fast(x, runtime_act);
}

template <typename T, int CH, int EPI, int ACT>
__device__ __forceinline__ void store_chunk_t(const GemmArgs& a, const float (&v)[CH][8], const float (&rv)[CH][8], int n,
                                              int mb0, int lane, float bval, int m_limit) {
  const bool n_ok = n < a.N;
  const bool do_act = n >= a.act_start;  // per lane; tiles are 128 features wide, act_start usually is too
#pragma unroll
  for (int u = 0; u < CH; ++u) {
    const int mb = mb0 + u * 8;
    const int rows = m_limit - mb;  // valid token rows in this group (may be <= 0 or >= 8)
    if constexpr (EPI == EPI_GATED) {
      // rows of w interleave (gate, up): even lane holds gate_j, odd lane up_j of output feature n/2
      T* o = reinterpret_cast<T*>(a.out) + (int64_t)mb * a.ldo + (n >> 1);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float y = v[u][j] + bval;
        const float other = __shfl_xor_sync(0xffffffffu, y, 1);
        if ((lane & 1) == 0 && n_ok && j < rows) o[(int64_t)j * a.ldo] = from_f32<T>(act_t<ACT>(y, a.act) * other + rv[u][j]);
      }
    } else if constexpr (EPI == EPI_F32) {
      float* o = reinterpret_cast<float*>(a.out) + (int64_t)mb * a.ldo + n;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float x = v[u][j] + bval;
        const float y = (ACT == LLMSS_ACT_NONE || !do_act) ? x : act_t<ACT>(x, a.act);
        if (n_ok && j < rows) o[(int64_t)j * a.ldo] = y + rv[u][j];
      }
    } else {
      T* o = reinterpret_cast<T*>(a.out) + (int64_t)mb * a.ldo + n;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float x = v[u][j] + bval;
        const float y = (ACT == LLMSS_ACT_NONE || !do_act) ? x : act_t<ACT>(x, a.act);
        if (n_ok && j < rows) o[(int64_t)j * a.ldo] = from_f32<T>(y + rv[u][j]);
      }
    }
  }
}

template <typename T, int CH>
__device__ __forceinline__ void store_chunk(const GemmArgs& a, const float (&v)[CH][8], const float (&rv)[CH][8], int n,
                                            int mb0, int lane, float bval, int m_limit) {
#define LLMSS_SC(EPI_, ACT_) store_chunk_t<T, CH, EPI_, ACT_>(a, v, rv, n, mb0, lane, bval, m_limit)
  if (a.epi == EPI_PLAIN) {
    LLMSS_SC(EPI_PLAIN, LLMSS_ACT_NONE);
  } else if (a.epi == EPI_ACT) {
    if (a.act == LLMSS_ACT_GELU_TANH) LLMSS_SC(EPI_ACT, LLMSS_ACT_GELU_TANH);
    else if (a.act == LLMSS_ACT_SILU) LLMSS_SC(EPI_ACT, LLMSS_ACT_SILU);
    else LLMSS_SC(EPI_ACT, ACT_RUNTIME);
  } else if (a.epi == EPI_GATED) {
    if (a.act == LLMSS_ACT_SILU) LLMSS_SC(EPI_GATED, LLMSS_ACT_SILU);
    else if (a.act == LLMSS_ACT_GELU_TANH) LLMSS_SC(EPI_GATED, LLMSS_ACT_GELU_TANH);
    else LLMSS_SC(EPI_GATED, ACT_RUNTIME);
  } else {
    if (a.act == LLMSS_ACT_NONE) LLMSS_SC(EPI_F32, LLMSS_ACT_NONE);
    else LLMSS_SC(EPI_F32, ACT_RUNTIME);
  }
#undef LLMSS_SC
}

// Plain flavour (bias / residual only), all CH*8 token rows and the feature in range: one pointer walked by the row
// stride, no per-element predicate or address arithmetic -- ~4 instructions per output instead of ~25 (the generic
// path above made the epilogue of a 128 x 256 tile instruction-bound: 11 us, ncu source page of gemm_pair_kernel).
template <typename T, int CH>
__device__ __forceinline__ void store_chunk_plain_full(const GemmArgs& a, const float (&v)[CH][8], int n, int mb0, float bval) {
  T* o = reinterpret_cast<T*>(a.out) + (int64_t)mb0 * a.ldo + n;
  if (a.residual != nullptr) {
    const T* r = reinterpret_cast<const T*>(a.residual) + (int64_t)mb0 * a.ldr + n;
    float rv[CH][8];
#pragma unroll
    for (int u = 0; u < CH; ++u)
#pragma unroll
      for (int j = 0; j < 8; ++j) { rv[u][j] = to_f32(*r); r += a.ldr; }
#pragma unroll
    for (int u = 0; u < CH; ++u)
#pragma unroll
      for (int j = 0; j < 8; ++j) { *o = from_f32<T>(v[u][j] + bval + rv[u][j]); o += a.ldo; }
  } else {
#pragma unroll
    for (int u = 0; u < CH; ++u)
#pragma unroll
      for (int j = 0; j < 8; ++j) { *o = from_f32<T>(v[u][j] + bval); o += a.ldo; }
  }
}

// The same for the activation and gated flavours (the whole warp's 32 features in range: the gated form shuffles).
template <typename T, int CH, int EPI, int ACT>
__device__ __forceinline__ void store_chunk_full_t(const GemmArgs& a, const float (&v)[CH][8], int n, int mb0, int lane, float bval) {
  const bool has_res = a.residual != nullptr;
  if constexpr (EPI == EPI_GATED) {
    // rows of w interleave (gate, up): even lane holds gate_j, odd lane up_j of output feature n/2
    T* o = reinterpret_cast<T*>(a.out) + (int64_t)mb0 * a.ldo + (n >> 1);
    const T* r = reinterpret_cast<const T*>(a.residual) + (int64_t)mb0 * a.ldr + (n >> 1);
    const bool writer = (lane & 1) == 0;
#pragma unroll
    for (int u = 0; u < CH; ++u)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float y = v[u][j] + bval;
        const float other = __shfl_xor_sync(0xffffffffu, y, 1);
        if (writer) *o = from_f32<T>(act_t<ACT>(y, a.act) * other + (has_res ? to_f32(*r) : 0.f));
        o += a.ldo;
        r += a.ldr;
      }
  } else {
    T* o = reinterpret_cast<T*>(a.out) + (int64_t)mb0 * a.ldo + n;
    const T* r = reinterpret_cast<const T*>(a.residual) + (int64_t)mb0 * a.ldr + n;
    const bool do_act = n >= a.act_start;
#pragma unroll
    for (int u = 0; u < CH; ++u)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float x = v[u][j] + bval;
        const float y = do_act ? act_t<ACT>(x, a.act) : x;
        *o = from_f32<T>(y + (has_res ? to_f32(*r) : 0.f));
        o += a.ldo;
        r += a.ldr;
      }
  }
}
// Dispatch of the full-chunk stores; false = flavour without a fast form (fp32 output): use store_chunk.
template <typename T, int CH>
__device__ __forceinline__ bool store_chunk_full(const GemmArgs& a, const float (&v)[CH][8], int n, int mb0, int lane, float bval) {
  if (a.epi == EPI_PLAIN) {
    store_chunk_plain_full<T, CH>(a, v, n, mb0, bval);
  } else if (a.epi == EPI_ACT) {
    if (a.act == LLMSS_ACT_GELU_TANH) store_chunk_full_t<T, CH, EPI_ACT, LLMSS_ACT_GELU_TANH>(a, v, n, mb0, lane, bval);
    else if (a.act == LLMSS_ACT_SILU) store_chunk_full_t<T, CH, EPI_ACT, LLMSS_ACT_SILU>(a, v, n, mb0, lane, bval);
    else store_chunk_full_t<T, CH, EPI_ACT, ACT_RUNTIME>(a, v, n, mb0, lane, bval);
  } else if (a.epi == EPI_GATED) {
    if (a.act == LLMSS_ACT_SILU) store_chunk_full_t<T, CH, EPI_GATED, LLMSS_ACT_SILU>(a, v, n, mb0, lane, bval);
    else if (a.act == LLMSS_ACT_GELU_TANH) store_chunk_full_t<T, CH, EPI_GATED, LLMSS_ACT_GELU_TANH>(a, v, n, mb0, lane, bval);
    else store_chunk_full_t<T, CH, EPI_GATED, ACT_RUNTIME>(a, v, n, mb0, lane, bval);
  } else {
    return false;
  }
  return true;
}

// ---- thread-block-cluster helpers (split-K kernel) ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// execution-only rendezvous of the cluster (no memory ordering: no MEMBAR / cache invalidation in the SASS)
__device__ __forceinline__ void cluster_sync_relaxed() {
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_shared(uint32_t local_smem_addr, uint32_t cta_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(cta_rank));
  return r;
}
__device__ __forceinline__ float4 ld_dsmem_f4(uint32_t cluster_addr) {
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(cluster_addr)
               : "memory");
  return v;
}

// COLL = false: no collective code at all (TP=1, column-parallel GEMMs, LM head): a shorter once-executed epilogue.
template <typename T, int BN, bool COLL>
__global__ void __launch_bounds__(kNumThreads, 1)
gemm_swapab_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                   const __grid_constant__ GemmArgs a) {
  using C = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + C::STAGES;
  uint64_t* tfull_bar = bars + 2 * C::STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  volatile int* s_flag = reinterpret_cast<volatile int*>(tmem_slot + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int G = gridDim.x, g = blockIdx.x;
  const int tiles = a.tiles_n * a.tiles_m;
  pdl_launch_dependents();
  if (threadIdx.x == 0) LLMSS_STAMP(0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_w);
    tma_prefetch_desc(&tmap_x);
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 4);  // one arrive per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, C::TMEM_COLS);
    tmem_relinquish();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // everything below reads what earlier kernels of the stream wrote (x, workspaces, epoch)
  uint32_t epoch = 0;
  if constexpr (COLL) { if (a.ar || a.ag) epoch = *a.ws.epoch + 1; }
  if (threadIdx.x == 0) LLMSS_STAMP(1);

  if (warp == 0) {
    // ===================== TMA producer (whole warp walks the schedule, one elected lane issues) ==========
    {
      const uint64_t pol_w = l2_policy_evict_first();  // weights are streamed once
      const uint64_t pol_x = l2_policy_evict_last();   // activations are re-read by every tile
      int stage = 0;
      uint32_t phase = 0;
      SegIter it(g, G, tiles, a.kb_total, a.aligned != 0);
      Segment sg;
      uint32_t acquired = 0;  // ag: bit q set once rank q's shard is known to be staged
      while (it.next(sg)) {
        const int tn = sg.tile % a.tiles_n, tm = (sg.tile / a.tiles_n + a.tm_rot) % a.tiles_m;
        const void* xmap = &tmap_x;
        int xrow = tm * BN;
        if (COLL && a.ag && xrow / a.rs_rows == a.ws.rank) {
          xrow -= a.ws.rank * a.rs_rows;  // own shard: tmap_x covers it in place, no waiting
        } else if (COLL && a.ag) {
          const int q = xrow / a.rs_rows;  // rank whose shard holds these token rows
          if (!(acquired & (1u << q))) {
            wait_flag_ge(symm_flag(a.ws, a.ws.rank, 1, 0) + q, epoch, 11);
            asm volatile("fence.proxy.async.global;" ::: "memory");  // peers' generic stores -> our TMA reads
            acquired |= 1u << q;
          }
          xmap = a.ag_maps + (epoch & 1);  // the gathered [M, K] activation in this rank's own bulk buffer
        }
        for (int kb = sg.kb0; kb < sg.kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1, 10);
          uint8_t* sw = smem + stage * C::STAGE_BYTES;
          if (elect_one()) {
            mbar_arrive_expect_tx(&full_bar[stage], C::STAGE_BYTES);
            tma_load_2d(sw, &tmap_w, &full_bar[stage], kb * BK, tn * BM, pol_w);
            tma_load_2d(sw + C::W_BYTES, xmap, &full_bar[stage], kb * BK, xrow, pol_x);
          }
          __syncwarp();
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (whole warp runs the loop, one elected lane issues) =================
    {
      const uint32_t idesc = umma_idesc_f16(a.dtype == LLMSS_BF16 ? 1 : 0, BM, BN);
      const uint32_t smem_base = smem_u32(smem);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      SegIter it(g, G, tiles, a.kb_total, a.aligned != 0);
      Segment sg;
      while (it.next(sg)) {
        mbar_wait(&tempty_bar[as], aphase ^ 1, 20);
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
        for (int kb = sg.kb0; kb < sg.kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase, 21);
          if (lane == 0 && a.dbg != nullptr && a.dbg[(size_t)blockIdx.x * 8 + 3] == 0) LLMSS_STAMP(3);
          tcgen05_fence_after();
          const uint32_t sw = smem_base + stage * C::STAGE_BYTES;
          const uint32_t sx = sw + C::W_BYTES;
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              umma_f16_ss(d_tmem, umma_desc_k_sw128(sw + k * 32), umma_desc_k_sw128(sx + k * 32), idesc,
                          (kb > sg.kb0 || k > 0) ? 1u : 0u);
            }
            umma_commit(&empty_bar[stage]);  // frees the smem stage when these MMAs retire
          }
          __syncwarp();
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) umma_commit(&tfull_bar[as]);  // accumulator ready for the epilogue
        __syncwarp();
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
      if (lane == 0) LLMSS_STAMP(4);
    }
  } else {
    // ===================== epilogue (4 warps = 128 TMEM lanes) =====================
    // The accumulator is consumed in groups of 8 token columns through *rolled* loops: the
    // epilogue runs once per segment, so its code must stay small enough not to be
    // instruction-fetch bound (a fully unrolled 256-column epilogue is ~20k SASS instructions).
    const int quarter = warp & 3;            // TMEM lane quarter this warp may access
    const int f = quarter * 32 + lane;       // feature within the tile
    const int epi_tid = threadIdx.x - 64;
    int as = 0;
    uint32_t aphase = 0;
    SegIter it(g, G, tiles, a.kb_total, a.aligned != 0);
    Segment sg;
    const T* bias = reinterpret_cast<const T*>(a.bias);
    constexpr int NG = BN / 8;               // column groups per tile
    int seg_idx = -1;
    const size_t ar_slot_bytes = (size_t)(a.ar == 2 ? tiles / a.ws.nranks : tiles) * BM * BN * sizeof(T);
    const int par = epoch & 1;

    if (COLL && a.ag) {
      // ---- fused all-gather (K3): while the TMA / MMA warps already work on the token tiles of this
      // rank's own shard, the (still idle) epilogue warps push that shard into every rank's gathered
      // activation buffer over NVLink; the last CTA to finish raises the "shard landed" flags.  Readers
      // then fetch tiles from their *local* copy (L2-cacheable; peer memory would bypass L2).
      const int kvec = a.K / 8;
      const int64_t nvec = (int64_t)a.rs_rows * kvec;
      const int64_t base = (int64_t)a.ws.rank * nvec;
      for (int64_t v = (int64_t)blockIdx.x * 128 + epi_tid; v < nvec; v += (int64_t)gridDim.x * 128) {
        const int64_t r = v / kvec, c = v % kvec;
        const uint4 val = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(a.ag_src) + r * a.ag_ld + c * 8);
        if (a.ws.mc_base != nullptr) {  // NVLS: one multicast store lands in every rank's gathered buffer
          mc_st_v4(symm_mc_buf<uint4>(a.ws, par) + base + v, val);
          continue;
        }
#pragma unroll 1
        for (int p = 0; p < a.ws.nranks; ++p) {
          const int peer = (a.ws.rank + p) % a.ws.nranks;
          symm_buf<uint4>(a.ws, peer, par)[base + v] = val;
        }
      }
      __threadfence_system();
      epi_bar();
      if (epi_tid == 0) {
        if (atomicAdd(a.ws.ticket + 1, 1u) == gridDim.x - 1) {
          a.ws.ticket[1] = 0;
          __threadfence_system();
          for (int r = 0; r < a.ws.nranks; ++r) st_release_sys(symm_flag(a.ws, r, 1, 0) + a.ws.rank, epoch);
        }
      }
    }

    while (it.next(sg)) {
      ++seg_idx;
      const int tn = sg.tile % a.tiles_n, tm = (sg.tile / a.tiles_n + a.tm_rot) % a.tiles_m;
      const int n = tn * BM + f;
      const int m0 = tm * BN;
      const bool n_ok = n < a.N;
      const bool full_tile = (sg.kb0 == 0 && sg.kb1 == a.kb_total);
      const uint32_t t_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(as * BN);
      const float bval = (bias != nullptr && n_ok) ? to_f32(bias[n]) : 0.f;

      mbar_wait(&tfull_bar[as], aphase, 30);
      tcgen05_fence_after();
      if (epi_tid == 0) LLMSS_STAMP(5);

      bool have_final = full_tile;
      if (!full_tile) {
        // ---- stream-K: publish my partial, last arriver reduces ----
        // partial layout [BN/4][128 features][4 cols]: one coalesced 16-byte store per 4 columns
        float4* part = reinterpret_cast<float4*>(a.sk_ws + (size_t)(2 * g + (seg_idx > 0 ? 1 : 0)) * (BM * BN)) + f;
#pragma unroll 1
        for (int cg = 0; cg < NG; ++cg) {
          uint32_t r[8];
          tmem_ld_32x32b_x8(t_addr + cg * 8, r);
          tmem_ld_wait();
          part[(cg * 2 + 0) * BM] = make_float4(__uint_as_float(r[0]), __uint_as_float(r[1]), __uint_as_float(r[2]), __uint_as_float(r[3]));
          part[(cg * 2 + 1) * BM] = make_float4(__uint_as_float(r[4]), __uint_as_float(r[5]), __uint_as_float(r[6]), __uint_as_float(r[7]));
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[as]);
        __threadfence();
        epi_bar();
        if (epi_tid == 0) {
          const int contributors = sg.last_cta - sg.first_cta + 1;
          const uint32_t old = atomicAdd(&a.sk_counters[sg.tile], 1u);
          const int last = (old == (uint32_t)(contributors - 1));
          if (last) a.sk_counters[sg.tile] = 0;
          *s_flag = last;
        }
        epi_bar();
        have_final = (*s_flag != 0);
        epi_bar();  // s_flag may be rewritten by the next segment
        if (have_final) __threadfence();
        if (epi_tid == 0) LLMSS_STAMP(2);
      }

      if (have_final) {
        // slot of the first contributor: a tile is the *first* segment of every contributor
        // except (possibly) the first one, whose range may start in an earlier tile
        const int first_slot = 2 * sg.first_cta +
            ((cta_start(sg.first_cta, it.total, G) != (int64_t)sg.tile * a.kb_total) ? 1 : 0);
        // pass 0: produce the rank-local tile (from TMEM or from the stream-K partials) and either
        //         store it (no AR) or push it to every rank's exchange buffer (AR);
        // pass 1: (AR only) after the per-tile handshake, sum the ranks' tiles and store.
        // reduce-scatter: the tile goes only to the rank that owns its token rows, which alone runs pass 1
        const int owner = (COLL && a.ar == 2) ? m0 / a.rs_rows : -1;
        const int n_pass = (COLL && (a.ar == 1 || (a.ar == 2 && owner == a.ws.rank))) ? 2 : 1;
        // exchange-buffer tile index: global for all-reduce, compact (tiles owned by one rank) for reduce-scatter
        const int xtile = (COLL && a.ar == 2) ? (tm - owner * (a.rs_rows / BN)) * a.tiles_n + tn : sg.tile;
#pragma unroll 1
        for (int pass = 0; pass < n_pass; ++pass) {
          if (pass == 1 && full_tile) {
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty_bar[as]);
          }
          // Groups of 8 token columns are processed CH at a time: all global loads of a chunk (residual,
          // stream-K partials) are issued before any of them is consumed, so the exposed tail pays one
          // L2 round trip per chunk and contributor instead of one per group.
          constexpr int CH = NG < 4 ? NG : 4;
          const bool storing = (pass == n_pass - 1);
          const int mrow0 = (COLL && a.ar == 2) ? m0 - a.ws.rank * a.rs_rows : m0;
          const int mlim = (COLL && a.ar == 2) ? a.rs_rows : a.M;
#pragma unroll 1
          for (int cg0 = 0; cg0 < NG; cg0 += CH) {
            float v[CH][8], rv[CH][8];
            if (storing && a.residual != nullptr) {
#pragma unroll
              for (int u = 0; u < CH; ++u) epilogue_load_res8<T>(a, rv[u], n, mrow0 + (cg0 + u) * 8, lane, mlim);
            } else {
#pragma unroll
              for (int u = 0; u < CH; ++u)
#pragma unroll
                for (int j = 0; j < 8; ++j) rv[u][j] = 0.f;
            }
            if (pass == 0) {
              if (full_tile) {
#pragma unroll
                for (int u = 0; u < CH; ++u) {
                  uint32_t r[8];
                  tmem_ld_32x32b_x8(t_addr + (cg0 + u) * 8, r);
                  tmem_ld_wait();
#pragma unroll
                  for (int j = 0; j < 8; ++j) v[u][j] = __uint_as_float(r[j]);
                }
              } else {
#pragma unroll
                for (int u = 0; u < CH; ++u)
#pragma unroll
                  for (int j = 0; j < 8; ++j) v[u][j] = 0.f;
                // contributors are summed in CTA order (deterministic)
                for (int cta = sg.first_cta; cta <= sg.last_cta; ++cta) {
                  const int slot = (cta == sg.first_cta) ? first_slot : 2 * cta;
                  const float4* p = reinterpret_cast<const float4*>(a.sk_ws + (size_t)slot * (BM * BN)) + f;
                  float4 t[CH][2];
#pragma unroll
                  for (int u = 0; u < CH; ++u) {
                    t[u][0] = ld_cg_f4(p + ((cg0 + u) * 2 + 0) * BM);
                    t[u][1] = ld_cg_f4(p + ((cg0 + u) * 2 + 1) * BM);
                  }
#pragma unroll
                  for (int u = 0; u < CH; ++u) {
                    v[u][0] += t[u][0].x; v[u][1] += t[u][0].y; v[u][2] += t[u][0].z; v[u][3] += t[u][0].w;
                    v[u][4] += t[u][1].x; v[u][5] += t[u][1].y; v[u][6] += t[u][1].z; v[u][7] += t[u][1].w;
                  }
                }
              }
              if (COLL && a.ar) {
#pragma unroll
                for (int u = 0; u < CH; ++u) {
                  const size_t xoff = ((size_t)xtile * NG + cg0 + u) * (BM * 16) + (size_t)f * 16;  // exchange-buffer offset
                  // round to T, one 16-byte vector per thread per group, tile-blocked + coalesced
                  const uint4 vec = lam_sanitize<2>(pack8<T>(v[u]));
                  if (a.ar == 2) {
                    *reinterpret_cast<uint4*>(symm_lam<uint8_t>(a.ws, owner, par) + (size_t)a.ws.rank * ar_slot_bytes + xoff) = vec;
                  } else {
#pragma unroll 1
                    for (int p = 0; p < a.ws.nranks; ++p) {
                      const int peer = (a.ws.rank + p) % a.ws.nranks;
                      *reinterpret_cast<uint4*>(symm_lam<uint8_t>(a.ws, peer, par) + (size_t)a.ws.rank * ar_slot_bytes + xoff) = vec;
                    }
                  }
                }
                continue;
              }
            } else if constexpr (COLL) {
              uint8_t* mine = symm_lam<uint8_t>(a.ws, a.ws.rank, par);
#pragma unroll
              for (int u = 0; u < CH; ++u) {
                const size_t xoff = ((size_t)xtile * NG + cg0 + u) * (BM * 16) + (size_t)f * 16;
#pragma unroll
                for (int j = 0; j < 8; ++j) v[u][j] = 0.f;
                // self-validating vectors: wait for all ranks' copies together, re-arm the slots (no flags, no fences)
                uint4* slot[kSymmMaxRanks];
                uint4 got[kSymmMaxRanks];
#pragma unroll
                for (int r = 0; r < kSymmMaxRanks; ++r)
                  slot[r] = reinterpret_cast<uint4*>(mine + (size_t)(r < a.ws.nranks ? r : 0) * ar_slot_bytes + xoff);
                lam_take_batch<kSymmMaxRanks>(slot, a.ws.nranks, got, 31);
#pragma unroll
                for (int r = 0; r < kSymmMaxRanks; ++r) {  // rank order: bit-identical on every rank
                  if (r < a.ws.nranks) {
                    float t8[8];
                    unpack8<T>(got[r], t8);
#pragma unroll
                    for (int j = 0; j < 8; ++j) v[u][j] += t8[j];
                  }
                }
              }
            }
            // all-reduce / reduce-scatter epilogues are plain (bias + residual); all-gather -> GEMM keeps every flavour
            if (COLL && a.ar) store_chunk_t<T, CH, EPI_PLAIN, LLMSS_ACT_NONE>(a, v, rv, n, mrow0 + cg0 * 8, lane, bval, mlim);
            else store_chunk<T, CH>(a, v, rv, n, mrow0 + cg0 * 8, lane, bval, mlim);
          }
        }
        if (full_tile && n_pass == 1) {
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty_bar[as]);
        }
      }
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    if (epi_tid == 0) LLMSS_STAMP(6);
  }

  // ---- teardown ----
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
  if constexpr (COLL) { if (a.ar || a.ag) symm_finish(a.ws, epoch); }
  if (threadIdx.x == 0) LLMSS_STAMP(7);
}


// ======================================================================================
// Cluster split-K variant for shapes with few output tiles (decode GEMMs at TP >= 2, every
// row-parallel GEMM: N = hidden -> 32 tiles).  A thread-block cluster of S CTAs shares one
// tile: CTA r streams k-blocks [r*kb/S, (r+1)*kb/S), parks its fp32 accumulator in its own
// shared memory (the drained TMA ring), and after one cluster barrier each CTA reduces 1/S of
// the tile straight out of its peers' shared memory (DSMEM, fixed CTA order -> deterministic)
// and runs the epilogue / fused all-reduce for that slice.  No global workspace, no atomics,
// no fences: the fix-up costs ~1 us instead of several L2 round trips.
// ======================================================================================
// Cluster split-K: sum of the S CTAs' fp32 partials of up to two (column group, feature) items of the tile.  All
// DSMEM loads of both items are in flight before the first is consumed -- the reduce is bound by the DSMEM round
// trip, not by its bandwidth -- and the sum runs in CTA order (deterministic).
__device__ __forceinline__ void dsmem_reduce_pair(uint32_t stage_u32, int S, int item0, int item1, bool two,
                                                  float (&v0)[8], float (&v1)[8], int rank_stride = 1, int rank_base = 0) {
  const uint32_t off0 = (uint32_t)((((item0 / BM) * 2) * BM + item0 % BM) * 16);
  const uint32_t off1 = (uint32_t)((((item1 / BM) * 2) * BM + item1 % BM) * 16);
  float4 lo0[8], hi0[8], lo1[8], hi1[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    if (q < S) {
      const uint32_t b0 = mapa_shared(stage_u32 + off0, (uint32_t)(q * rank_stride + rank_base));
      lo0[q] = ld_dsmem_f4(b0);
      hi0[q] = ld_dsmem_f4(b0 + BM * 16);
      if (two) {
        const uint32_t b1 = mapa_shared(stage_u32 + off1, (uint32_t)(q * rank_stride + rank_base));
        lo1[q] = ld_dsmem_f4(b1);
        hi1[q] = ld_dsmem_f4(b1 + BM * 16);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) v0[j] = v1[j] = 0.f;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    if (q < S) {
      v0[0] += lo0[q].x; v0[1] += lo0[q].y; v0[2] += lo0[q].z; v0[3] += lo0[q].w;
      v0[4] += hi0[q].x; v0[5] += hi0[q].y; v0[6] += hi0[q].z; v0[7] += hi0[q].w;
      if (two) {
        v1[0] += lo1[q].x; v1[1] += lo1[q].y; v1[2] += lo1[q].z; v1[3] += lo1[q].w;
        v1[4] += hi1[q].x; v1[5] += hi1[q].y; v1[6] += hi1[q].z; v1[7] += hi1[q].w;
      }
    }
  }
}

// Push mode: the S partials of an item already sit in this CTA's receive area ([source CTA][half][item] float4 planes).
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void recv_reduce_pair(uint32_t recv_u32, int per, int S, int li0, int li1, bool two,
                                                 float (&v0)[8], float (&v1)[8]) {
#pragma unroll
  for (int j = 0; j < 8; ++j) v0[j] = v1[j] = 0.f;
  if (!two) li1 = li0;
#pragma unroll 1
  for (int q = 0; q < S; ++q) {  // CTA order: deterministic; a rolled loop keeps the once-executed epilogue short
    const uint32_t base = recv_u32 + (uint32_t)(q * 2 * per) * 16;
    const float4 lo0 = lds_f4(base + li0 * 16), hi0 = lds_f4(base + (per + li0) * 16);
    const float4 lo1 = lds_f4(base + li1 * 16), hi1 = lds_f4(base + (per + li1) * 16);
    v0[0] += lo0.x; v0[1] += lo0.y; v0[2] += lo0.z; v0[3] += lo0.w;
    v0[4] += hi0.x; v0[5] += hi0.y; v0[6] += hi0.z; v0[7] += hi0.w;
    v1[0] += lo1.x; v1[1] += lo1.y; v1[2] += lo1.z; v1[3] += lo1.w;
    v1[4] += hi1.x; v1[5] += hi1.y; v1[6] += hi1.z; v1[7] += hi1.w;
  }
}
__device__ __forceinline__ void st_dsmem_f4(uint32_t cluster_addr, const float4& v) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(cluster_addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// AR = false: the instantiation without any collective code (column-parallel GEMMs, TP=1): its epilogue -- executed
// once per CTA, so every instruction-cache line it touches is a miss -- stays compact.
template <typename T, int BN, bool AR>
__global__ void __launch_bounds__(kNumThreads, 1)
gemm_cluster_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                    const __grid_constant__ GemmArgs a) {
  using C = CCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + C::STAGES;
  uint64_t* tfull_bar = bars + 2 * C::STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull_bar + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = (int)cluster_nctarank(), r = (int)cluster_ctarank();
  const int tile = blockIdx.x / S;
  const int tn = tile % a.tiles_n, tm = tile / a.tiles_n;
  const int kb0 = (int)(((int64_t)r * a.kb_total) / S), kb1 = (int)(((int64_t)(r + 1) * a.kb_total) / S);
  constexpr int NG = BN / 8;
  constexpr int TCOLS = BN < 32 ? 32 : BN;
  pdl_launch_dependents();
  if (threadIdx.x == 0) LLMSS_STAMP(0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_w);
    tma_prefetch_desc(&tmap_x);
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tfull_bar[0], 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, TCOLS);
    tmem_relinquish();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  // push mode: a CTA may only be written to once it is known to be running.  Everybody arrives here and waits later,
  // just before its first remote store (epilogue warps) or after its loop (TMA / MMA warps): the barrier costs nothing
  // on the critical path.
  if constexpr (C::PUSH) asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  float4* staging = reinterpret_cast<float4*>(smem);  // pull mode: [NG*2][128] float4, reuses the drained TMA ring
  float4* recv = reinterpret_cast<float4*>(smem + C::STAGES * C::STAGE_BYTES + 512);  // push mode: never aliases the ring
  const int total_items_c = NG * BM;
  const int per_c = ((total_items_c + S - 1) / S + 31) / 32 * 32;   // items owned by each CTA of the cluster
  pdl_wait();  // everything below reads what earlier kernels of the stream wrote (x, workspaces, epoch)
  uint32_t epoch = 0;
  if constexpr (AR) epoch = *a.ws.epoch + 1;
  if (threadIdx.x == 0) LLMSS_STAMP(1);

  if (warp == 0) {
    // whole warp walks the K range, one elected lane issues (see elect_one in common.cuh)
    const uint64_t pol_w = l2_policy_evict_first();
    const uint64_t pol_x = l2_policy_evict_last();
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = kb0; kb < kb1; ++kb) {
      mbar_wait(&empty_bar[stage], phase ^ 1, 40);
      uint8_t* sw = smem + stage * C::STAGE_BYTES;
      if (elect_one()) {
        mbar_arrive_expect_tx(&full_bar[stage], C::STAGE_BYTES);
        tma_load_2d(sw, &tmap_w, &full_bar[stage], kb * BK, tn * BM, pol_w);
        tma_load_2d(sw + C::W_BYTES, &tmap_x, &full_bar[stage], kb * BK, tm * BN, pol_x);
      }
      __syncwarp();
      if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
    }
    if constexpr (C::PUSH) asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  } else if (warp == 1) {
    const uint32_t idesc = umma_idesc_f16(a.dtype == LLMSS_BF16 ? 1 : 0, BM, BN);
    const uint32_t smem_base = smem_u32(smem);
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = kb0; kb < kb1; ++kb) {
      mbar_wait(&full_bar[stage], phase, 41);
      if (lane == 0 && kb == kb0) LLMSS_STAMP(3);
      tcgen05_fence_after();
      const uint32_t sw = smem_base + stage * C::STAGE_BYTES;
      const uint32_t sx = sw + C::W_BYTES;
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < BK / 16; ++k)
          umma_f16_ss(tmem_base, umma_desc_k_sw128(sw + k * 32), umma_desc_k_sw128(sx + k * 32), idesc,
                      (kb > kb0 || k > 0) ? 1u : 0u);
        umma_commit(&empty_bar[stage]);
      }
      __syncwarp();
      if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
    }
    if (elect_one()) umma_commit(&tfull_bar[0]);
    __syncwarp();
    if (lane == 0) LLMSS_STAMP(4);
    if constexpr (C::PUSH) asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  } else {
    if constexpr (C::PUSH) asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");  // all peers are running
    // park the accumulator in shared memory: [(cg*2 + half)][feature] float4, conflict-free
    const int quarter = warp & 3;
    const int f = quarter * 32 + lane;
    const uint32_t t_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    mbar_wait(&tfull_bar[0], 0, 42);  // all MMAs retired => every TMA stage has landed and been consumed
    tcgen05_fence_after();
    if (threadIdx.x == 64) LLMSS_STAMP(5);
    const uint32_t recv_u32 = smem_u32(recv);
#pragma unroll 1
    for (int cg = 0; cg < NG; ++cg) {
      uint32_t q[8];
      tmem_ld_32x32b_x8(t_addr + cg * 8, q);
      tmem_ld_wait();
      const float4 lo = make_float4(__uint_as_float(q[0]), __uint_as_float(q[1]), __uint_as_float(q[2]), __uint_as_float(q[3]));
      const float4 hi = make_float4(__uint_as_float(q[4]), __uint_as_float(q[5]), __uint_as_float(q[6]), __uint_as_float(q[7]));
      if constexpr (C::PUSH) {
        // hand the 8 token columns of this feature to the CTA that owns the item (posted remote shared-memory stores)
        const int item = cg * BM + f;
        const int owner = item / per_c, li = item - owner * per_c;
        const uint32_t dst = mapa_shared(recv_u32 + (uint32_t)(((r * 2) * per_c + li) * 16), (uint32_t)owner);
        st_dsmem_f4(dst, lo);
        st_dsmem_f4(dst + (uint32_t)per_c * 16, hi);
      } else {
        staging[(cg * 2 + 0) * BM + f] = lo;
        staging[(cg * 2 + 1) * BM + f] = hi;
      }
    }
  }
  __syncwarp();
  cluster_sync_all();  // every CTA's partial tile is now visible cluster-wide
  if (threadIdx.x == 64) LLMSS_STAMP(2);

  if (warp >= 2) {
    const int epi_tid = threadIdx.x - 64;
    // (column group, feature) items owned by this CTA: warp-granular, the last CTA may own fewer
    const int total_items = NG * BM;
    const int per = ((total_items + S - 1) / S + 31) / 32 * 32;
    const int my_items = max(0, min(per, total_items - r * per));
    const T* bias = reinterpret_cast<const T*>(a.bias);
    const int tiles = a.tiles_n * a.tiles_m;
    const size_t ar_slot_bytes = (size_t)tiles * BM * BN * sizeof(T);
    const int par = epoch & 1;
    const uint32_t stage_u32 = smem_u32(staging);
    if (AR && a.ar == 3) {
      // ---- NVLS all-reduce (K1, in-switch): (A) the reduced bf16/fp16 slice goes to this rank's OWN bulk buffer
      // (local stores, nothing on the wire) and one multicast flag store tells every rank that this CTA slot
      // of this rank is staged; (B) rank q sums the q-th share of the slot with multimem.ld_reduce -- the
      // switch reads the t copies and adds them in fp32 -- and broadcasts the result with one multicast store
      // into every rank's Lamport area; (C) each rank takes its slice from its own Lamport area (the vector is
      // its own arrival flag), adds bias / residual and stores.  Per rank and layer the wire carries ~2 x the
      // message instead of (t-1) x, and every rank ends up with the same bits (the owner's sum).
      uint8_t* part = symm_buf<uint8_t>(a.ws, a.ws.rank, par);
      const size_t tile_off = (size_t)tile * NG * BM * 16;
#pragma unroll 1
      for (int it0 = epi_tid; it0 < my_items; it0 += 256) {
        const bool two = it0 + 128 < my_items;
        const int item0 = r * per + it0, item1 = two ? item0 + 128 : item0;
        float va[8], vb[8];
        if constexpr (C::PUSH) recv_reduce_pair(smem_u32(recv), per, S, it0, it0 + 128, two, va, vb);
        else dsmem_reduce_pair(stage_u32, S, item0, item1, two, va, vb);
        *reinterpret_cast<uint4*>(part + tile_off + (size_t)item0 * 16) = pack8<T>(va);
        if (two) *reinterpret_cast<uint4*>(part + tile_off + (size_t)item1 * 16) = pack8<T>(vb);
      }
      epi_bar();
      if (epi_tid == 0) {
        __threadfence_system();
        mc_st_release_u32(symm_mc_flag(a.ws, 0, blockIdx.x) + a.ws.rank, epoch);
      }
      if (epi_tid < a.ws.nranks) wait_flag_ge(symm_flag(a.ws, a.ws.rank, 0, blockIdx.x) + epi_tid, epoch, 44);
      epi_bar();
      if (threadIdx.x == 64) LLMSS_STAMP(2);
      {
        const int sub = (my_items + a.ws.nranks - 1) / a.ws.nranks;
        const uint8_t* mc_part = symm_mc_buf<uint8_t>(a.ws, par);
        uint8_t* mc_res = symm_mc_lam<uint8_t>(a.ws, par);
#pragma unroll 1
        for (int i = epi_tid; i < sub; i += 128) {
          const int li = a.ws.rank * sub + i;
          if (li < my_items) {
            const size_t xoff = tile_off + (size_t)(r * per + li) * 16;
            mc_st_v4(mc_res + xoff, lam_sanitize<2>(mc_ld_reduce<T>(mc_part + xoff)));
          }
        }
      }
      uint8_t* mine = symm_lam<uint8_t>(a.ws, a.ws.rank, par);
#pragma unroll 1
      for (int it = epi_tid; it < my_items; it += 128) {
        const int item = r * per + it;
        const int cg = item / BM, f = item % BM;
        const int n = tn * BM + f;
        float v[1][8], rv[1][8];
        epilogue_load_res8<T>(a, rv[0], n, tm * BN + cg * 8, lane, a.M);  // ahead of the wait
        unpack8<T>(lam_take(reinterpret_cast<uint4*>(mine + tile_off + (size_t)item * 16), 45), v[0]);
        const float bval = (bias != nullptr && n < a.N) ? to_f32(bias[n]) : 0.f;
        store_chunk_t<T, 1, EPI_PLAIN, LLMSS_ACT_NONE>(a, v, rv, n, tm * BN + cg * 8, lane, bval, a.M);
      }
    }
    if (AR && a.ar == 4) {
      // ---- two-shot all-reduce without flags (K1 for t >= 4): (A) every vector of this CTA's slice goes to the ONE
      // rank that owns it (1/t of the slice each) as a self-validating Lamport vector; (B) the owner sums the t
      // contributions in rank order and broadcasts the result -- one multicast store through the NVSwitch when
      // the heap is multicast-bound, else t peer stores; (C) every rank takes its slice of results, adds bias /
      // residual and stores.  (t-1)/t + 1/t of the message leaves each GPU instead of (t-1) x, both hops are
      // validated by the data itself (no flag, no fence), and all ranks hold the owner's bits.
      const int t = a.ws.nranks, me = a.ws.rank;
      const int sub = (my_items + t - 1) / t;
      const size_t tile_off = (size_t)tile * NG * BM * 16;
#pragma unroll 1
      for (int it0 = epi_tid; it0 < my_items; it0 += 256) {
        const bool two = it0 + 128 < my_items;
        const int item0 = r * per + it0, item1 = two ? item0 + 128 : item0;
        float va[8], vb[8];
        if constexpr (C::PUSH) recv_reduce_pair(smem_u32(recv), per, S, it0, it0 + 128, two, va, vb);
        else dsmem_reduce_pair(stage_u32, S, item0, item1, two, va, vb);
        *reinterpret_cast<uint4*>(symm_lam<uint8_t>(a.ws, it0 / sub, par) + (size_t)me * ar_slot_bytes + tile_off +
                                  (size_t)item0 * 16) = lam_sanitize<2>(pack8<T>(va));
        if (two)
          *reinterpret_cast<uint4*>(symm_lam<uint8_t>(a.ws, (it0 + 128) / sub, par) + (size_t)me * ar_slot_bytes + tile_off +
                                    (size_t)item1 * 16) = lam_sanitize<2>(pack8<T>(vb));
      }
      if (threadIdx.x == 64) LLMSS_STAMP(2);
      uint8_t* mine = symm_lam<uint8_t>(a.ws, me, par);
      const size_t res_off = (size_t)t * ar_slot_bytes;
#pragma unroll 1
      for (int i = epi_tid; i < sub; i += 128) {
        const int li = me * sub + i;
        if (li >= my_items) break;
        const size_t xoff = tile_off + (size_t)(r * per + li) * 16;
        uint4* slot[kSymmMaxRanks];
        uint4 got[kSymmMaxRanks];
#pragma unroll
        for (int q = 0; q < kSymmMaxRanks; ++q)
          slot[q] = reinterpret_cast<uint4*>(mine + (size_t)(q < t ? q : 0) * ar_slot_bytes + xoff);
        lam_take_batch<kSymmMaxRanks>(slot, t, got, 46);  // all t contributions polled together
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
        for (int q = 0; q < kSymmMaxRanks; ++q) {  // rank order
          if (q < t) {
            float t8[8];
            unpack8<T>(got[q], t8);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += t8[j];
          }
        }
        const uint4 vec = lam_sanitize<2>(pack8<T>(acc));
        if (a.ws.mc_base != nullptr) {
          mc_st_v4(symm_mc_lam<uint8_t>(a.ws, par) + res_off + xoff, vec);
        } else {
#pragma unroll 1
          for (int p = 0; p < t; ++p)
            *reinterpret_cast<uint4*>(symm_lam<uint8_t>(a.ws, (me + p) % t, par) + res_off + xoff) = vec;
        }
      }
      // results: up to four items per thread are awaited together
#pragma unroll 1
      for (int it0 = epi_tid; it0 < my_items; it0 += 4 * 128) {
        uint4* slot[4];
        uint4 got[4];
        float rv[4][8];
        int m = 0;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int it = it0 + u * 128;
          const int item = r * per + (it < my_items ? it : it0);
          slot[u] = reinterpret_cast<uint4*>(mine + res_off + tile_off + (size_t)item * 16);
          if (it < my_items) {
            m = u + 1;
            epilogue_load_res8<T>(a, rv[u], tn * BM + item % BM, tm * BN + (item / BM) * 8, lane, a.M);  // ahead of the wait
          }
        }
        lam_take_batch<4>(slot, m, got, 47);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (u < m) {
            const int item = r * per + it0 + u * 128;
            const int cg = item / BM, f = item % BM;
            const int n = tn * BM + f;
            float v[1][8], rv1[1][8];
            unpack8<T>(got[u], v[0]);
#pragma unroll
            for (int j = 0; j < 8; ++j) rv1[0][j] = rv[u][j];
            const float bval = (bias != nullptr && n < a.N) ? to_f32(bias[n]) : 0.f;
            store_chunk_t<T, 1, EPI_PLAIN, LLMSS_ACT_NONE>(a, v, rv1, n, tm * BN + cg * 8, lane, bval, a.M);  // fused collectives: plain epilogue only
          }
        }
      }
    }
    const int n_pass = !AR ? 1 : (a.ar >= 3 ? 0 : (a.ar ? 2 : 1));
#pragma unroll 1
    for (int pass = 0; pass < n_pass; ++pass) {
#pragma unroll 1
      for (int it0 = epi_tid; it0 < my_items; it0 += 256) {
        const bool two = it0 + 128 < my_items;
        float vv[2][8];
        if (pass == 0) {
          if constexpr (C::PUSH) recv_reduce_pair(smem_u32(recv), per, S, it0, it0 + 128, two, vv[0], vv[1]);
          else dsmem_reduce_pair(stage_u32, S, r * per + it0, r * per + (two ? it0 + 128 : it0), two, vv[0], vv[1]);
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          if (u == 1 && !two) break;
          const int item = r * per + it0 + u * 128;
          const int cg = item / BM, f = item % BM;
          const int n = tn * BM + f;
          const size_t xoff = ((size_t)tile * NG + cg) * (BM * 16) + (size_t)f * 16;
          float v[1][8], rv[1][8];
          if (pass == n_pass - 1) epilogue_load_res8<T>(a, rv[0], n, tm * BN + cg * 8, lane, a.M);  // ahead of the wait
          if (pass == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[0][j] = vv[u][j];
            if (AR && a.ar) {
              const uint4 vec = lam_sanitize<2>(pack8<T>(v[0]));
#pragma unroll 1
              for (int p = 0; p < a.ws.nranks; ++p) {
                const int peer = (a.ws.rank + p) % a.ws.nranks;
                *reinterpret_cast<uint4*>(symm_lam<uint8_t>(a.ws, peer, par) + (size_t)a.ws.rank * ar_slot_bytes + xoff) = vec;
              }
              continue;
            }
          } else if constexpr (AR) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[0][j] = 0.f;
            uint8_t* mine = symm_lam<uint8_t>(a.ws, a.ws.rank, par);
            uint4* slot[kSymmMaxRanks];
            uint4 got[kSymmMaxRanks];
#pragma unroll
            for (int q = 0; q < kSymmMaxRanks; ++q)
              slot[q] = reinterpret_cast<uint4*>(mine + (size_t)(q < a.ws.nranks ? q : 0) * ar_slot_bytes + xoff);
            lam_take_batch<kSymmMaxRanks>(slot, a.ws.nranks, got, 43);  // polled together, not one round trip per rank
#pragma unroll
            for (int q = 0; q < kSymmMaxRanks; ++q) {  // rank order: bit-identical on every rank
              if (q < a.ws.nranks) {
                float t8[8];
                unpack8<T>(got[q], t8);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[0][j] += t8[j];
              }
            }
          }
          const float bval = (bias != nullptr && n < a.N) ? to_f32(bias[n]) : 0.f;
          if constexpr (AR) store_chunk_t<T, 1, EPI_PLAIN, LLMSS_ACT_NONE>(a, v, rv, n, tm * BN + cg * 8, lane, bval, a.M);
          else store_chunk<T, 1>(a, v, rv, n, tm * BN + cg * 8, lane, bval, a.M);
        }
      }
    }
  }
  if (threadIdx.x == 64) LLMSS_STAMP(6);
  __syncwarp();
  cluster_sync_all();  // nobody may exit while a peer is still reading its shared memory

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, TCOLS);
  }
  if constexpr (AR) symm_finish(a.ws, epoch);
  if (threadIdx.x == 0) LLMSS_STAMP(7);
}

// ======================================================================================
// CTA-pair kernel (tcgen05 cta_group::2) for token dimensions above 128.
//
// With 128-feature tiles every CTA re-reads the whole 256-token activation tile per k-block: 32 KB of activations from
// L2 for 16 KB of weights, and from M = 256 upwards the L2 -> SM traffic and the bytes a 4-stage ring can keep in
// flight (not HBM, not the tensor pipe) bound the GEMM.  Here two CTAs on the two SMs of a TPC share one
// 256 x 256 x 16 MMA: each CTA loads its own 128 weight rows and HALF of the token rows (both at the same shared-memory
// offsets), the even CTA issues `tcgen05.mma.cta_group::2`, the tensor cores of both SMs read both halves of the
// activation tile, and each CTA's tensor memory receives the accumulator of its own 128 features.  Activation traffic per
// weight byte halves and the ring holds 6 stages instead of 4 (measured: 313 ns per k-block of 256 x 256 x 64, against
// ~410 ns for two single-CTA tiles at M = 8192).
//
// Two schedules, chosen on the host:
//  * persistent (clusters of 2, S = 1): every pair walks tiles p, p + P, p + 2P, ... with TWO accumulators of 256 columns
//    in each CTA's tensor memory, so the epilogue of one tile (straight out of tensor memory: bias / activation / gating
//    / residual / store) runs under the main loop of the next;
//  * split-K (clusters of 2 x S, one tile per cluster, few-tile shapes): pair p = rank / 2 streams the p-th share of K,
//    parity h = rank % 2 selects the feature half; the S partial tiles of one parity are reduced out of the peers' shared
//    memory like in the cluster kernel above (pull mode, CTA order, deterministic), each CTA finishing 1/S of its tile.
//
// Protocol: every CTA's TMA warp fills its own ring and arms its own `full` barrier.  The odd CTA's warp 1 forwards
// each `full` completion to the even CTA's `peer_full` barrier (remote mbarrier arrive); the even CTA's warp 1 waits
// for both before it issues, and frees the stage in BOTH CTAs with one multicast commit.  Accumulator hand-over: one
// multicast commit raises `tfull` in both CTAs; the epilogue warps of both CTAs arrive on the EVEN CTA's `tempty`
// (8 arrivals: 4 local, 4 remote) when they have drained a buffer.  All loads are plain per-CTA TMA loads of this
// GPU's own operands.
// ======================================================================================
struct PCfg2 {
  static constexpr int BN = 256;                 // token columns of the pair's accumulator
  static constexpr int XH = BN / 2;              // token rows each CTA loads
  static constexpr int W_BYTES = BM * BK * 2;
  static constexpr int X_BYTES = XH * BK * 2;
  static constexpr int STAGE_BYTES = W_BYTES + X_BYTES;
  static constexpr int STAGES = 6;               // 192 KB ring; the parked fp32 tile of the split-K schedule (128 KB) reuses it
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 512 /*barriers*/;
};

template <typename T>
__global__ void __launch_bounds__(kNumThreads, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                 const __grid_constant__ GemmArgs a) {
  using C = PCfg2;
  constexpr int BN = C::BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);
  uint64_t* full_bar = bars;                       // own ring stage landed (TMA transaction bytes)
  uint64_t* empty_bar = bars + C::STAGES;          // stage consumed (multicast commit of the even CTA)
  uint64_t* peer_full_bar = bars + 2 * C::STAGES;  // even CTA only: the odd CTA's stage landed
  uint64_t* tfull_bar = bars + 3 * C::STAGES;      // [2] accumulator complete (multicast commit)
  uint64_t* tempty_bar = tfull_bar + 2;            // [2] even CTA only: accumulator drained by both CTAs' epilogues
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int CS = (int)cluster_nctarank(), r = (int)cluster_ctarank();
  const int S = CS >> 1, p = r >> 1, h = r & 1;
  const int pair_tiles_n = (a.tiles_n + 1) >> 1;
  const int n_tiles = pair_tiles_n * a.tiles_m;
  const int n_clusters = (int)gridDim.x / CS, cluster_id = (int)blockIdx.x / CS;   // S > 1: n_clusters == n_tiles
  const int kb0 = (int)(((int64_t)p * a.kb_total) / S), kb1 = (int)(((int64_t)(p + 1) * a.kb_total) / S);
  constexpr int NG = BN / 8;
  constexpr int TCOLS = 2 * BN;
  // bring-up / profiling switch (LLMSS_GEMM_PAIR_DEBUG; results are WRONG unless 0): 2 = no MMAs (load pipeline
  // alone), 3 = no loads (MMA issue rate alone)
  const int dbg_mode = a.tm_rot;
  pdl_launch_dependents();
  if (threadIdx.x == 0) LLMSS_STAMP(0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_w);
    tma_prefetch_desc(&tmap_x);
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
      mbar_init(&peer_full_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 8);   // one arrive per epilogue warp of either CTA
    }
    fence_mbar_init();
  }
  if (warp == 1) {   // warp 1 of BOTH CTAs: the pair allocates together and gets the same columns
    tmem_alloc_pair(tmem_slot, TCOLS);
    tmem_relinquish_pair();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  cluster_sync_relaxed();  // barriers (fence.mbarrier_init above) and tensor memory of both CTAs exist before anything arrives remotely
  const uint32_t tmem_base = *tmem_slot;
  float4* staging = reinterpret_cast<float4*>(smem);  // split-K: [NG*2][128] float4, reuses the drained TMA ring
  pdl_wait();
  if (threadIdx.x == 0) LLMSS_STAMP(1);

  if (warp == 0) {
    const uint64_t pol_w = l2_policy_evict_first();
    const uint64_t pol_x = l2_policy_evict_last();
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = cluster_id; tile < n_tiles && dbg_mode != 3; tile += n_clusters) {
      // token tile fastest: the pairs running at the same time share ONE weight tile (streamed from HBM once, L2 hits
      // for the others) and sweep the activation, which stays L2-resident across the weight tiles
      const int tn = (tile / a.tiles_m) * 2 + h;        // this CTA's 128-feature tile (may lie past N: zero-filled, unstored)
      const int xrow = (tile % a.tiles_m) * BN + (h ^ (a.aligned & 1)) * C::XH;   // a.aligned: bring-up switch for the half order
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1, 60);
        uint8_t* sw = smem + stage * C::STAGE_BYTES;
        if (elect_one()) {
          mbar_arrive_expect_tx(&full_bar[stage], C::STAGE_BYTES);
          tma_load_2d(sw, &tmap_w, &full_bar[stage], kb * BK, tn * BM, pol_w);
          tma_load_2d(sw + C::W_BYTES, &tmap_x, &full_bar[stage], kb * BK, xrow, pol_x);
        }
        __syncwarp();
        if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1 && h == 0) {
    // ---- issuer (even CTA): one MMA of 256 features x 256 tokens x 16 per instruction, over both CTAs' stages
    const uint32_t idesc = umma_idesc_f16(a.dtype == LLMSS_BF16 ? 1 : 0, 2 * BM, BN);
    const uint32_t smem_base = smem_u32(smem);
    const uint16_t pair_mask = (uint16_t)(3u << (r & ~1));
    int stage = 0, as = 0;
    uint32_t phase = 0, aphase = 0;
    for (int tile = cluster_id; tile < n_tiles; tile += n_clusters) {
      mbar_wait(&tempty_bar[as], aphase ^ 1, 65);   // both CTAs' epilogues have drained this accumulator
      tcgen05_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
      for (int kb = kb0; kb < kb1; ++kb) {
        if (dbg_mode != 3) {
          mbar_wait(&full_bar[stage], phase, 61);
          mbar_wait(&peer_full_bar[stage], phase, 62);
        }
        if (lane == 0 && a.dbg != nullptr && a.dbg[(size_t)blockIdx.x * 8 + 3] == 0) LLMSS_STAMP(3);
        tcgen05_fence_after();
        const uint32_t sw = smem_base + stage * C::STAGE_BYTES;
        const uint32_t sx = sw + C::W_BYTES;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            if (dbg_mode != 2)
              umma_f16_ss_pair(d_tmem, umma_desc_k_sw128(sw + k * 32), umma_desc_k_sw128(sx + k * 32), idesc,
                               (kb > kb0 || k > 0) ? 1u : 0u);
          umma_commit_pair(&empty_bar[stage], pair_mask);   // frees this stage in both CTAs
        }
        __syncwarp();
        if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
      }
      if (elect_one()) umma_commit_pair(&tfull_bar[as], pair_mask);
      __syncwarp();
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    if (lane == 0) LLMSS_STAMP(4);
  } else if (warp == 1) {
    // ---- forwarder (odd CTA): tell the issuer that this CTA's half of the stage has landed
    int stage = 0;
    uint32_t phase = 0;
    const uint32_t leader = (uint32_t)(r & ~1);
    for (int tile = cluster_id; tile < n_tiles && dbg_mode != 3; tile += n_clusters) {
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase, 63);
        if (elect_one()) mbar_arrive_remote(mapa_shared(smem_u32(&peer_full_bar[stage]), leader));
        __syncwarp();
        if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    const int quarter = warp & 3;
    const int f = quarter * 32 + lane;
    const T* bias = reinterpret_cast<const T*>(a.bias);
    int as = 0;
    uint32_t aphase = 0;
    const uint32_t tempty_leader = mapa_shared(smem_u32(&tempty_bar[0]), (uint32_t)(r & ~1));
    for (int tile = cluster_id; tile < n_tiles; tile += n_clusters) {
      const int tn = (tile / a.tiles_m) * 2 + h, tm = tile % a.tiles_m;
      const int n = tn * BM + f, m0 = tm * BN;
      const uint32_t t_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(as * BN);
      mbar_wait(&tfull_bar[as], aphase, 64);  // all MMAs of the tile retired => its stages of BOTH CTAs have been consumed
      tcgen05_fence_after();
      if (threadIdx.x == 64) LLMSS_STAMP(5);
      if (S == 1) {
        // ---- straight out of tensor memory, CH groups of 8 token columns at a time (residuals fetched ahead)
        constexpr int CH = 4;
        const float bval = (bias != nullptr && n < a.N) ? to_f32(bias[n]) : 0.f;
        // fast stores need the warp's 32 features in range (warp-uniform: the gated flavour shuffles) and a flavour that
        // has one; the row test below is per chunk
        const bool plain_ok = a.epi != EPI_F32 && tn * BM + quarter * 32 + 32 <= a.N;
#pragma unroll 1
        for (int cg0 = 0; cg0 < NG; cg0 += CH) {
          float v[CH][8], rv[CH][8];
          const bool fast = plain_ok && m0 + (cg0 + CH) * 8 <= a.M;
          if (!fast) {
#pragma unroll
            for (int u = 0; u < CH; ++u) epilogue_load_res8<T>(a, rv[u], n, m0 + (cg0 + u) * 8, lane, a.M);
          }
          {
            // one 32-column load and ONE wait per chunk: a tensor-memory round trip is ~0.3 us, and paying it per 8
            // columns made the epilogue of a 128 x 256 tile 11 us (measured) instead of ~3
            uint32_t q[32];
            tmem_ld_32x32b_x32(t_addr + cg0 * 8, q);
            tmem_ld_wait();
#pragma unroll
            for (int u = 0; u < CH; ++u)
#pragma unroll
              for (int j = 0; j < 8; ++j) v[u][j] = __uint_as_float(q[u * 8 + j]);
          }
          if (cg0 + CH >= NG) {   // accumulator fully read: hand the buffer back before the stores
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_remote(tempty_leader + (uint32_t)(as * 8));
          }
          if (fast) store_chunk_full<T, CH>(a, v, n, m0 + cg0 * 8, lane, bval);
          else store_chunk<T, CH>(a, v, rv, n, m0 + cg0 * 8, lane, bval, a.M);
        }
      } else {
        // ---- split-K: park the partial accumulator in shared memory: [(cg*2 + half)][feature] float4, conflict-free
#pragma unroll 1
        for (int cg0 = 0; cg0 < NG; cg0 += 4) {
          uint32_t q[32];
          tmem_ld_32x32b_x32(t_addr + cg0 * 8, q);
          tmem_ld_wait();
#pragma unroll
          for (int u = 0; u < 8; ++u)
            staging[((cg0 + (u >> 1)) * 2 + (u & 1)) * BM + f] = make_float4(
                __uint_as_float(q[u * 4 + 0]), __uint_as_float(q[u * 4 + 1]), __uint_as_float(q[u * 4 + 2]), __uint_as_float(q[u * 4 + 3]));
        }
      }
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
  }
  if (S > 1) {   // cluster-uniform: one tile per cluster, partials parked above
    tcgen05_fence_before();
    __syncwarp();
    cluster_sync_all();  // every CTA's partial tile is parked and visible cluster-wide
    if (threadIdx.x == 64) LLMSS_STAMP(2);
    if (warp >= 2) {
      const int tile = cluster_id;
      const int tn = (tile / a.tiles_m) * 2 + h, tm = tile % a.tiles_m;
      const int epi_tid = threadIdx.x - 64;
      const int total_items = NG * BM;
      const int per = ((total_items + S - 1) / S + 31) / 32 * 32;   // (column group, feature) items finished by each pair
      const int my_items = max(0, min(per, total_items - p * per));
      const T* bias = reinterpret_cast<const T*>(a.bias);
      const uint32_t stage_u32 = smem_u32(staging);
#pragma unroll 1
      for (int it0 = epi_tid; it0 < my_items; it0 += 256) {
        const bool two = it0 + 128 < my_items;
        float vv[2][8];
        dsmem_reduce_pair(stage_u32, S, p * per + it0, p * per + (two ? it0 + 128 : it0), two, vv[0], vv[1], 2, h);
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          if (u == 1 && !two) break;
          const int item = p * per + it0 + u * 128;
          const int cg = item / BM, f = item % BM;
          const int n = tn * BM + f;
          float v[1][8], rv[1][8];
          epilogue_load_res8<T>(a, rv[0], n, tm * BN + cg * 8, lane, a.M);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[0][j] = vv[u][j];
          const float bval = (bias != nullptr && n < a.N) ? to_f32(bias[n]) : 0.f;
          store_chunk<T, 1>(a, v, rv, n, tm * BN + cg * 8, lane, bval, a.M);
        }
      }
    }
  }
  if (threadIdx.x == 64) LLMSS_STAMP(6);
  tcgen05_fence_before();
  __syncwarp();
  cluster_sync_relaxed();  // nobody may exit (or free tensor memory) while a peer still reads its shared memory / issues into it
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc_pair(tmem_base, TCOLS);
  }
  if (threadIdx.x == 0) LLMSS_STAMP(7);
}

// ------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  }
  return fn;
}

// 2-D row-major [rows, cols] 16-bit tensor with row stride `ld` elements; box = [box_rows, 64 cols], 128B swizzle.
// cuTensorMapEncodeTiled costs a few microseconds of host time and every eager launch (prefill: 2-4 GEMMs x layers)
// needs two maps; weights and the static activation buffers recur, so each host thread keeps the last 256 encodings.
struct TmapKey {
  const void* ptr;
  int64_t rows, cols, ld;
  int dtype, box_rows;
  bool operator==(const TmapKey& o) const {
    return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && dtype == o.dtype && box_rows == o.box_rows;
  }
};
struct TmapCache {
  static constexpr int N = 256;
  TmapKey key[N];
  CUtensorMap map[N];
  bool used[N] = {false};
};
int make_tmap_uncached(CUtensorMap* map, const void* ptr, int dtype, int64_t rows, int64_t cols, int64_t ld, int box_rows);
int make_tmap(CUtensorMap* map, const void* ptr, int dtype, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
  static thread_local TmapCache* cache = new TmapCache();
  const TmapKey k{ptr, rows, cols, ld, dtype, box_rows};
  uint64_t h = (uint64_t)(uintptr_t)ptr * 0x9E3779B97F4A7C15ull ^ (uint64_t)rows * 0xC2B2AE3D27D4EB4Full ^
               (uint64_t)cols * 0x165667B19E3779F9ull ^ (uint64_t)(ld * 31 + box_rows * 7 + dtype);
  const int slot = (int)((h >> 40) % TmapCache::N);
  if (cache->used[slot] && cache->key[slot] == k) {
    *map = cache->map[slot];
    return 0;
  }
  if (int e = make_tmap_uncached(map, ptr, dtype, rows, cols, ld, box_rows)) return e;
  cache->key[slot] = k;
  cache->map[slot] = *map;
  cache->used[slot] = true;
  return 0;
}
int make_tmap_uncached(CUtensorMap* map, const void* ptr, int dtype, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
  PFN_encodeTiled enc = get_encode();
  LLMSS_REQUIRE(enc != nullptr, "gemm: cuTensorMapEncodeTiled entry point not available");
  LLMSS_REQUIRE(((uintptr_t)ptr % 16) == 0 && (ld * 2) % 16 == 0, "gemm: operand base / row stride must be 16-byte aligned");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, dtype == LLMSS_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                   const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  LLMSS_REQUIRE(r == CUDA_SUCCESS, "gemm: cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld cols=%lld ld=%lld)",
                (int)r, (long long)rows, (long long)cols, (long long)ld);
  return 0;
}

unsigned long long* g_dbg = nullptr;
int g_num_sms = 0;
int num_sms() {
  if (g_num_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  return g_num_sms;
}

template <typename T, int BN, bool COLL>
int launch_coll(const CUtensorMap& tw, const CUtensorMap& tx, const GemmArgs& a, int grid, cudaStream_t stream) {
  using C = Cfg<BN>;
  static PerDeviceSmem configured;
  auto kern = gemm_swapab_kernel<T, BN, COLL>;
  LLMSS_CHECK_CUDA(configured.ensure(kern, C::SMEM_BYTES));
  LLMSS_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(kNumThreads), (size_t)C::SMEM_BYTES, stream, tw, tx, a));
  LLMSS_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <typename T, int BN>
int launch(const CUtensorMap& tw, const CUtensorMap& tx, const GemmArgs& a, int grid, cudaStream_t stream) {
  return (a.ar || a.ag) ? launch_coll<T, BN, true>(tw, tx, a, grid, stream) : launch_coll<T, BN, false>(tw, tx, a, grid, stream);
}

template <typename T, int BN, bool AR>
int launch_cluster_ar(const CUtensorMap& tw, const CUtensorMap& tx, const GemmArgs& a, int tiles, int split,
                      cudaStream_t stream) {
  using C = CCfg<BN>;
  static PerDeviceSmem configured;
  auto kern = gemm_cluster_kernel<T, BN, AR>;
  LLMSS_CHECK_CUDA(configured.ensure(kern, C::SMEM_BYTES));
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(tiles * split);
  cfg.blockDim = dim3(kNumThreads);
  cfg.dynamicSmemBytes = C::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = split;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = llmss_pdl_enabled() ? 2 : 1;
  LLMSS_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, tw, tx, a));
  return 0;
}

template <typename T, int BN>
int launch_cluster(const CUtensorMap& tw, const CUtensorMap& tx, const GemmArgs& a, int tiles, int split,
                   cudaStream_t stream) {
  return a.ar ? launch_cluster_ar<T, BN, true>(tw, tx, a, tiles, split, stream)
              : launch_cluster_ar<T, BN, false>(tw, tx, a, tiles, split, stream);
}

template <typename T>
int launch_pair(const CUtensorMap& tw, const CUtensorMap& tx, const GemmArgs& a, int clusters, int split,
                cudaStream_t stream) {
  static PerDeviceSmem configured;
  auto kern = gemm_pair_kernel<T>;
  LLMSS_CHECK_CUDA(configured.ensure(kern, PCfg2::SMEM_BYTES));
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(clusters * 2 * split);
  cfg.blockDim = dim3(kNumThreads);
  cfg.dynamicSmemBytes = PCfg2::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2 * split;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = llmss_pdl_enabled() ? 2 : 1;
  LLMSS_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, tw, tx, a));
  return 0;
}

// CTA-pair kernel: -1 = environment (LLMSS_GEMM_PAIR: off | auto (default) | 1 | 2 | 4), 0 = off, 1 / 2 / 4 = pairs per
// cluster (1 = persistent, 2 / 4 = split-K depth), 8 = auto.  llmss_gemm_set_pair() overrides the environment (A/B measurements inside one process).
int g_pair_mode = -1;
int pair_mode() {
  if (g_pair_mode >= 0) return g_pair_mode;
  static int env_mode = -1;
  if (env_mode < 0) {
    const char* e = getenv("LLMSS_GEMM_PAIR");
    env_mode = 8;  // default: automatic (pick_pair_split)
    if (e != nullptr) env_mode = strcmp(e, "off") == 0 ? 0 : strcmp(e, "auto") == 0 ? 8 : atoi(e);
    if (env_mode != 0 && env_mode != 1 && env_mode != 2 && env_mode != 4 && env_mode != 8) env_mode = 8;
  }
  return env_mode;
}
// Pairs per cluster for a plain GEMM with a 256-token tile, 0 = keep the single-CTA kernels.  Deepest split-K that
// still fits the machine in one wave and leaves every pair at least two k-blocks.
int pick_pair_split(int M, int N, int kb_total, int sms) {
  const int mode = pair_mode();
  if (mode == 0 || M <= 128) return 0;
  const int pair_tiles = ((N + 2 * BM - 1) / (2 * BM)) * ((M + 255) / 256);
  if (mode == 1 || mode == 2 || mode == 4) return mode <= kb_total ? mode : 1;
  // automatic (profiles/bench_gemm_pair.md): persistent pairs from 38 pair tiles upwards (N >= 9728 at M <= 256, every
  // prefill GEMM); 24..37 pair tiles: two pairs per cluster split K, which fills the machine (4 CTAs per tile) and beat
  // both the persistent form and the single-CTA kernels on every measured shape; below that the single-CTA cluster
  // kernel with its deeper split wins.
  if (pair_tiles >= 38) return 1;
  if (pair_tiles >= 24 && pair_tiles * 4 <= sms && kb_total >= 4) return 2;
  return 0;
}

// Largest cluster size (2..8) for which all `tiles` clusters are co-resident (one wave) and every CTA still
// streams >= 2 k-blocks; 0 = use the stream-K persistent kernel.  LLMSS_GEMM_MODE=streamk|cluster forces one.
template <typename T, int BN>
int max_active_clusters(int split) {
  static int cache[9] = {-1, -1, -1, -1, -1, -1, -1, -1, -1};
  if (cache[split] >= 0) return cache[split];
  using C = CCfg<BN>;
  auto kern = gemm_cluster_kernel<T, BN, false>;  // both instantiations have the same launch footprint
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(split * 64);
  cfg.blockDim = dim3(kNumThreads);
  cfg.dynamicSmemBytes = C::SMEM_BYTES;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = split;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) {
    cudaGetLastError();
    n = 0;
  }
  cache[split] = n;
  return n;
}

int clusters_for(int dtype, int bn, int split) {
#define LLMSS_MAC(T_)                                          \
  switch (bn) {                                                \
    case 16: return max_active_clusters<T_, 16>(split);        \
    case 32: return max_active_clusters<T_, 32>(split);        \
    case 64: return max_active_clusters<T_, 64>(split);        \
    case 128: return max_active_clusters<T_, 128>(split);      \
    default: return max_active_clusters<T_, 256>(split);       \
  }
  if (dtype == LLMSS_F16) { LLMSS_MAC(__half) }
  LLMSS_MAC(__nv_bfloat16)
#undef LLMSS_MAC
}

int pick_cluster_split(int tiles, int kb_total, int bn, int sms, int dtype) {
  static int forced = -1;
  static int min_kb = -1;  // k-blocks (64 x 16-bit columns, 16 KB of weights) every CTA of a split must at least stream
  if (forced < 0) {
    const char* e = getenv("LLMSS_GEMM_MODE");
    forced = (e == nullptr) ? 0 : (strcmp(e, "streamk") == 0 ? 1 : (strcmp(e, "cluster") == 0 ? 2 : 0));
    const char* m = getenv("LLMSS_GEMM_MIN_KB_PER_CTA");
    min_kb = (m != nullptr) ? atoi(m) : 2;
    if (min_kb < 1) min_kb = 1;
  }
  if (forced == 1) return 0;
  // Splitting K over a cluster divides the weight stream AND the epilogue (each CTA finishes 1/S of the tile) at the
  // price of a cluster barrier and a DSMEM exchange.  Measured on the TP=8 per-rank shapes (M = 16..128, K = 512..2560,
  // scripts/bench_gemm_small.py): the deepest split that leaves two k-blocks per CTA wins everywhere -- 8.2 us vs
  // 12.0 us for one CTA per tile at M=64, K=512 -- because the epilogue, not the stream, is the long pole.  A "cluster"
  // of one (K of one or two k-blocks) still runs this kernel, so the fused all-reduce epilogues stay available.
  for (int s = 8; s >= 1; --s) {
    if (s > 1 && kb_total / s < min_kb) continue;
    if (forced != 2 && (tiles * s > sms || (s > 1 && tiles > clusters_for(dtype, bn, s)))) continue;
    // stream-K uses every SM; take the tile-per-CTA / cluster kernel only if it occupies a fair share of them, or if
    // the whole GEMM is so small that the stream-K fix-up would dominate
    if (forced != 2 && tiles * s * 10 < sms * 7 && (int64_t)tiles * kb_total > 4 * sms) return 0;
    return s;
  }
  return 0;
}

int pick_bn(int M) {
  static int max_bn = -1;  // LLMSS_GEMM_MAX_BN caps the token-tile width (A/B aid: 128 halves the epilogue of a CTA at M = 256)
  if (max_bn < 0) {
    const char* e = getenv("LLMSS_GEMM_MAX_BN");
    max_bn = (e != nullptr) ? atoi(e) : 256;
    if (max_bn != 16 && max_bn != 32 && max_bn != 64 && max_bn != 128) max_bn = 256;
  }
  const int bn = M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : M <= 128 ? 128 : 256;
  return bn < max_bn ? bn : max_bn;
}

}  // namespace

// Profiling aid: when non-null, every CTA of the following llmss_gemm launches writes 8 globaltimer
// stamps (entry, prologue done, last TMA issued, first stage landed, last MMA issued, first
// accumulator ready, epilogue done, exit) to buf[cta*8 .. cta*8+7].  buf must be zeroed by the caller.
LLMSS_API int llmss_gemm_set_pair(int mode) {
  g_pair_mode = mode;
  return 0;
}
LLMSS_API int llmss_gemm_set_debug(unsigned long long* buf) {
  g_dbg = buf;
  return 0;
}

// Plan of a GEMM launch (so the caller can size the workspaces before capture).
struct llmss_gemm_plan {
  int bn, tiles, kb_total, grid;  // grid: CTAs launched (stream-K) or tiles * split (cluster)
  int split;                // 0 = stream-K kernel, else cluster split-K factor
  int64_t sk_ws_floats;     // stream-K partial workspace (floats)
  int64_t sk_counters;      // uint32 counters (must be zero before first use)
  int64_t ar_bytes;         // symmetric buffer bytes needed per parity when ar=1
};

LLMSS_API int llmss_gemm_plan_query2(int M, int N, int K, int nranks, int elem_bytes, int max_ctas, int dtype,
                                     llmss_gemm_plan* p);
LLMSS_API int llmss_gemm_plan_query(int M, int N, int K, int nranks, int elem_bytes, int max_ctas, llmss_gemm_plan* p) {
  return llmss_gemm_plan_query2(M, N, K, nranks, elem_bytes, max_ctas, LLMSS_BF16, p);
}
static int gemm_plan_impl(int M, int N, int K, int nranks, int elem_bytes, int max_ctas, int dtype, int force_bn,
                          llmss_gemm_plan* p);
LLMSS_API int llmss_gemm_plan_query2(int M, int N, int K, int nranks, int elem_bytes, int max_ctas, int dtype,
                                     llmss_gemm_plan* p) {
  return gemm_plan_impl(M, N, K, nranks, elem_bytes, max_ctas, dtype, 0, p);
}
// force_bn > 0: token-tile width imposed by the reduce-scatter / all-gather modes (an argument, not a global: two host
// threads may plan GEMMs at the same time)
static int gemm_plan_impl(int M, int N, int K, int nranks, int elem_bytes, int max_ctas, int dtype, int force_bn,
                          llmss_gemm_plan* p) {
  const int bn = force_bn > 0 ? force_bn : pick_bn(M);
  const int tiles_n = (N + BM - 1) / BM, tiles_m = (M + bn - 1) / bn;
  const int tiles = tiles_n * tiles_m;
  const int kb_total = (K + BK - 1) / BK;
  const int64_t work = (int64_t)tiles * kb_total;
  int sms = max_ctas > 0 ? max_ctas : num_sms();
  // at least ~4 k-blocks (64 KB of weights) per CTA so the stream-K fix-up stays a small fraction
  int64_t grid = work / 4;
  if (grid < tiles && tiles <= sms) grid = tiles;
  if (grid < 1) grid = 1;
  if (grid > sms) grid = sms;
  if (grid > work) grid = work;
  p->bn = bn;
  p->tiles = tiles;
  p->kb_total = kb_total;
  p->split = pick_cluster_split(tiles, kb_total, bn, sms, dtype);
  if (p->split > 0) {
    grid = (int64_t)tiles * p->split;
  } else if (grid == sms) {
    // Prefer a grid that owns whole tiles when one exists close to the SM count: no partial tiles, so no
    // stream-K dump / ticket / fix-up tail.  Measured (M=64, K=4096, cold weights): N=14336 31.9 -> 27.3 us,
    // N=12288 31.4 -> 25.5, N=16384 34.5 -> 29.8, N=11008 29.8 -> 25.0, N=28672 46.5 -> 45.2; a grid that
    // leaves some CTAs one tile short (N=50432: 394 tiles on 132 CTAs) does not pay.
    if (tiles <= sms) {
      if (tiles * 100 >= sms * 55) grid = tiles;
    } else {
      const int per = (tiles + sms - 1) / sms;
      const int g = (tiles + per - 1) / per;
      if (per <= 2 && tiles % g == 0 && g * 100 >= sms * 70) grid = g;
    }
  }
  p->grid = (int)grid;
  p->sk_ws_floats = p->split > 0 ? 0 : (int64_t)2 * grid * BM * bn;
  p->sk_counters = tiles;
  p->ar_bytes = (int64_t)nranks * tiles * BM * bn * elem_bytes;
  return 0;
}

// out = epilogue(x @ w^T).  x: [M,K] row stride ldx; w: [N,K] row stride ldw; both 16-bit `dtype`.
// gated: rows of w interleave (gate, up); out has N/2 columns.  ar: fused all-reduce over `ws`
// (bias/residual are applied once, after the reduction).  sk_ws / sk_counters: see plan query.
LLMSS_API int llmss_gemm_ex(const void* x, int64_t ldx, const void* w, int64_t ldw, const void* bias,
                            const void* residual, int64_t ldr, void* out, int64_t ldo, int M, int N, int K, int dtype,
                            int out_f32, int act, int act_start, int gated, float* sk_ws, uint32_t* sk_counters,
                            SymmWs* ws, int max_ctas, int collective, void* ag_maps_dev, cudaStream_t stream);

LLMSS_API int llmss_gemm(const void* x, int64_t ldx, const void* w, int64_t ldw, const void* bias,
                         const void* residual, int64_t ldr, void* out, int64_t ldo, int M, int N, int K, int dtype,
                         int out_f32, int act, int act_start, int gated, float* sk_ws, uint32_t* sk_counters,
                         SymmWs* ws, int max_ctas, cudaStream_t stream) {
  return llmss_gemm_ex(x, ldx, w, ldw, bias, residual, ldr, out, ldo, M, N, K, dtype, out_f32, act, act_start, gated,
                       sk_ws, sk_counters, ws, max_ctas, (ws != nullptr && ws->nranks > 1) ? 1 : 0, nullptr, stream);
}

// collective: 0 none | 1 GEMM -> all-reduce | 2 GEMM -> reduce-scatter over token rows (out / residual are this
// rank's [M/t, N] shard) | 3 all-gather -> GEMM (x is this rank's [M/t, K] shard; M is the gathered row count;
// ag_maps_dev: 64-byte aligned device scratch of 2 * 128 bytes for the tensor maps of the gathered activation).
LLMSS_API int llmss_gemm_ex(const void* x, int64_t ldx, const void* w, int64_t ldw, const void* bias,
                            const void* residual, int64_t ldr, void* out, int64_t ldo, int M, int N, int K, int dtype,
                            int out_f32, int act, int act_start, int gated, float* sk_ws, uint32_t* sk_counters,
                            SymmWs* ws, int max_ctas, int collective, void* ag_maps_dev, cudaStream_t stream) {
  LLMSS_REQUIRE(dtype == LLMSS_F16 || dtype == LLMSS_BF16, "gemm: dtype must be f16/bf16");
  LLMSS_REQUIRE(M > 0 && N > 0 && K > 0 && K % 8 == 0, "gemm: bad shape M=%d N=%d K=%d (K must be a multiple of 8)", M, N, K);
  LLMSS_REQUIRE(!gated || (N % 2 == 0), "gemm: gated epilogue needs even N");
  LLMSS_REQUIRE(!(gated && out_f32), "gemm: gated epilogue with fp32 output is not supported");
  if (ws == nullptr || ws->nranks <= 1) collective = 0;
  // collective 4 / 5 / 6 pin the all-reduce algorithm: 4 = NVLS in-switch reduction, 5 = one-shot Lamport push,
  // 6 = two-shot Lamport (reduce-scatter over peer stores + multicast broadcast)
  const bool ar = collective == 1 || (collective >= 4 && collective <= 6), rs = collective == 2, ag = collective == 3;
  LLMSS_REQUIRE(!((ar || rs) && (gated || act != LLMSS_ACT_NONE || out_f32)),
                "gemm: fused all-reduce / reduce-scatter support bias/residual epilogues only");
  llmss_gemm_plan p;
  int rs_rows = 0, force_bn = 0;
  if (rs || ag) {
    // token rows are owned by ranks in contiguous blocks; tiles must not straddle two owners, so the
    // token tile is the largest power of two (16..256) that divides the rows per rank
    LLMSS_REQUIRE(M % ws->nranks == 0, "gemm: M=%d is not divisible by %d ranks", M, ws->nranks);
    rs_rows = M / ws->nranks;
    int bn = 256;
    while (bn > 16 && rs_rows % bn != 0) bn >>= 1;
    LLMSS_REQUIRE(rs_rows % bn == 0, "gemm: rows per rank (%d) must be a multiple of 16", rs_rows);
    force_bn = bn;
  }
  gemm_plan_impl(M, N, K, (ar || rs) ? ws->nranks : 1, 2, max_ctas, dtype, force_bn, &p);
  if (rs || ag) {
    p.split = 0;  // these modes live in the persistent stream-K kernel
    int sms = max_ctas > 0 ? max_ctas : num_sms();
    int64_t work = (int64_t)p.tiles * p.kb_total, grid = work / 4;
    if (grid < p.tiles && p.tiles <= sms) grid = p.tiles;
    if (grid < 1) grid = 1;
    if (grid > sms) grid = sms;
    p.grid = (int)grid;
  }
  static const int aligned_min = getenv("LLMSS_GEMM_ALIGNED_MIN") ? atoi(getenv("LLMSS_GEMM_ALIGNED_MIN")) : 4;
  const bool aligned = p.split == 0 && p.tiles >= aligned_min * p.grid;    // many tiles per CTA: plain data-parallel
  const bool need_sk = p.split == 0 && !aligned && (p.tiles % p.grid) != 0;  // some CTA range cuts through a tile
  LLMSS_REQUIRE(!need_sk || (sk_ws != nullptr && sk_counters != nullptr), "gemm: stream-K workspace required for this shape");
  int ar_mode = 1;  // kernel-side GemmArgs::ar: 1 one-shot push, 3 NVLS (flag + multimem.ld_reduce), 4 two-shot Lamport
  if (ar) {
    LLMSS_REQUIRE(p.tiles * (p.split > 0 ? p.split : 1) <= kSymmMaxSlots,
                  "gemm: %d tiles exceed the %d flag slots of the symmetric workspace", p.tiles, kSymmMaxSlots);
    // in-switch reduction: cluster kernel only (one CTA slot per flag), message staged in the bulk area and
    // returned through one Lamport slot set; chosen from llmss_nvls_min_bytes() of padded message upwards
    const size_t msg = (size_t)p.tiles * BM * p.bn * 2;
    const bool nvls_ok = ws->mc_base != nullptr && p.split > 0 && msg <= ws->buf_bytes && msg <= ws->lam_bytes;
    const bool two_ok = p.split > 0 && (size_t)(ws->nranks + 1) * msg <= ws->lam_bytes;
    LLMSS_REQUIRE(collective != 4 || nvls_ok, "gemm: NVLS all-reduce unavailable for this shape / workspace (multicast %s, split %d)",
                  ws->mc_base != nullptr ? "bound" : "not bound", p.split);
    LLMSS_REQUIRE(collective != 6 || two_ok, "gemm: two-shot all-reduce epilogue unavailable for this shape / workspace (split %d)", p.split);
    // automatic choice: the one-shot push is one hop but (t-1) x the bytes, the two-shot exchange two hops and ~1 x:
    // from 4 ranks and LLMSS_TWOSHOT_MIN_BYTES upwards the two-shot form wins (profiles/bench_gemm_ar.md)
    if (collective == 4) ar_mode = 3;
    else if (collective == 6) ar_mode = 4;
    else if (collective == 1 && two_ok && ((ws->nranks >= 4 && (int64_t)M * N * 2 >= llmss_twoshot_min_bytes()) ||
                                          (int64_t)M * N * 2 >= (1 << 20))) ar_mode = 4;
    LLMSS_REQUIRE(ar_mode != 1 || (size_t)p.ar_bytes <= ws->lam_bytes, "gemm: fused all-reduce needs %lld bytes of Lamport buffer (have %zu)",
                  (long long)p.ar_bytes, ws->lam_bytes);
  }
  if (rs) {
    const size_t need = (size_t)p.tiles * BM * p.bn * 2;  // nranks sources x (tiles / nranks) owned tiles
    LLMSS_REQUIRE(need <= ws->lam_bytes, "gemm: fused reduce-scatter needs %zu bytes of Lamport buffer (have %zu)", need,
                  ws->lam_bytes);
  }
  if (collective == 0) {
    // token tiles of 129..256 rows: two CTAs share one 256 x 256 MMA, each loading half of the activation tile
    const int ps = pick_pair_split(M, N, p.kb_total, max_ctas > 0 ? max_ctas : num_sms());
    if (ps > 0) {
      CUtensorMap pw, px;
      if (int e = make_tmap(&pw, w, dtype, N, K, ldw, BM)) return e;
      if (int e = make_tmap(&px, x, dtype, M, K, ldx, PCfg2::XH)) return e;
      GemmArgs a;
      memset(&a, 0, sizeof(a));
      a.M = M; a.N = N; a.K = K;
      a.tiles_n = (N + BM - 1) / BM;
      a.tiles_m = (M + PCfg2::BN - 1) / PCfg2::BN;
      a.kb_total = p.kb_total;
      a.dtype = dtype;
      a.bias = bias; a.residual = residual; a.ldr = ldr;
      a.out = out; a.ldo = ldo; a.out_f32 = out_f32;
      a.act = act; a.act_start = act_start; a.gated = gated;
      a.epi = gated ? EPI_GATED : (out_f32 ? EPI_F32 : (act != LLMSS_ACT_NONE ? EPI_ACT : EPI_PLAIN));
      static const int flip = getenv("LLMSS_GEMM_PAIR_FLIP") ? atoi(getenv("LLMSS_GEMM_PAIR_FLIP")) : 0;
      a.aligned = flip;
      { const char* d = getenv("LLMSS_GEMM_PAIR_DEBUG"); a.tm_rot = d ? atoi(d) : 0; }
      a.dbg = g_dbg;
      const int pair_tiles = ((a.tiles_n + 1) / 2) * a.tiles_m;
      // split-K: one tile per cluster; otherwise persistent pairs, as many as the machine holds (balanced over the tiles)
      int clusters = pair_tiles;
      if (ps == 1) {
        int max_pairs = (max_ctas > 0 ? max_ctas : num_sms()) / 2;
        if (max_pairs < 1) max_pairs = 1;
        const int rounds = (pair_tiles + max_pairs - 1) / max_pairs;
        clusters = (pair_tiles + rounds - 1) / rounds;
      }
      return dtype == LLMSS_F16 ? launch_pair<__half>(pw, px, a, clusters, ps, stream)
                                : launch_pair<__nv_bfloat16>(pw, px, a, clusters, ps, stream);
    }
  }
  CUtensorMap tw, tx;
  AgMaps ag_maps;  // all-gather mode only
  memset(&tx, 0, sizeof(tx));
  memset(&ag_maps, 0, sizeof(ag_maps));
  if (int e = make_tmap(&tw, w, dtype, N, K, ldw, BM)) return e;
  if (ag) {
    // x is this rank's [rs_rows, K] shard; every rank stages its shard in its symmetric bulk buffer and the
    // kernel's TMA producer pulls token tiles from the owners (peer HBM over NVLink).
    LLMSS_REQUIRE(ag_maps_dev != nullptr && ((uintptr_t)ag_maps_dev % 64) == 0, "gemm: all-gather mode needs a 64-byte aligned map scratch");
    LLMSS_REQUIRE((size_t)M * K * 2 <= ws->buf_bytes, "gemm: gathered activation of %zu bytes exceeds the symmetric bulk buffer (%zu)",
                  (size_t)M * K * 2, ws->buf_bytes);
    for (int par = 0; par < 2; ++par) {
      const char* base = ws->peer_base[ws->rank] + kSymmDataOffset + 2 * ws->lam_bytes + (size_t)par * ws->buf_bytes;
      if (int e = make_tmap(&ag_maps.m[par], base, dtype, M, K, K, p.bn)) return e;
    }
    // the maps live in device memory (copied per call from this stack frame: pageable source, so the copy is staged
    // synchronously and the launch is NOT capturable in a CUDA graph in this mode)
    LLMSS_CHECK_CUDA(cudaMemcpyAsync(ag_maps_dev, &ag_maps, sizeof(ag_maps), cudaMemcpyHostToDevice, stream));
    if (int e = make_tmap(&tx, x, dtype, rs_rows, K, ldx, p.bn)) return e;  // this rank's own shard, read in place
  } else {
    if (int e = make_tmap(&tx, x, dtype, M, K, ldx, p.bn)) return e;
  }
  GemmArgs a;
  memset(&a, 0, sizeof(a));
  a.M = M; a.N = N; a.K = K;
  a.tiles_n = (N + BM - 1) / BM;
  a.tiles_m = (M + p.bn - 1) / p.bn;
  a.kb_total = p.kb_total;
  a.dtype = dtype;
  a.bias = bias; a.residual = residual; a.ldr = ldr;
  a.out = out; a.ldo = ldo; a.out_f32 = out_f32;
  a.act = act; a.act_start = act_start; a.gated = gated;
  a.aligned = aligned ? 1 : 0;
  a.epi = gated ? EPI_GATED : (out_f32 ? EPI_F32 : (act != LLMSS_ACT_NONE ? EPI_ACT : EPI_PLAIN));
  a.sk_ws = sk_ws; a.sk_counters = sk_counters;
  a.ar = ar ? ar_mode : (rs ? 2 : 0);
  a.rs_rows = rs_rows;
  a.ag = ag ? 1 : 0;
  a.ag_src = x;
  a.ag_ld = ldx;
  a.ag_maps = reinterpret_cast<const CUtensorMap*>(ag_maps_dev);
  a.tm_rot = ag ? ws->rank * (rs_rows / p.bn) : 0;
  a.dbg = g_dbg;
  if (collective) a.ws = *ws;
#define LLMSS_LAUNCH(T_, BN_) \
  return p.split > 0 ? launch_cluster<T_, BN_>(tw, tx, a, p.tiles, p.split, stream) : launch<T_, BN_>(tw, tx, a, p.grid, stream)
  if (dtype == LLMSS_F16) {
    switch (p.bn) {
      case 16: LLMSS_LAUNCH(__half, 16);
      case 32: LLMSS_LAUNCH(__half, 32);
      case 64: LLMSS_LAUNCH(__half, 64);
      case 128: LLMSS_LAUNCH(__half, 128);
      default: LLMSS_LAUNCH(__half, 256);
    }
  } else {
    switch (p.bn) {
      case 16: LLMSS_LAUNCH(__nv_bfloat16, 16);
      case 32: LLMSS_LAUNCH(__nv_bfloat16, 32);
      case 64: LLMSS_LAUNCH(__nv_bfloat16, 64);
      case 128: LLMSS_LAUNCH(__nv_bfloat16, 128);
      default: LLMSS_LAUNCH(__nv_bfloat16, 256);
    }
  }
#undef LLMSS_LAUNCH
}
