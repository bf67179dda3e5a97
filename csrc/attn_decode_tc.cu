// K9b attn_decode_tc: paged split-KV decode attention on the 5th-generation tensor cores.
//
// The CUDA-core decode kernel (attn_decode.cu) spends ~60 instructions per cached token per query head,
// which is fine for MHA but hopeless for MQA (StarCoder: 48 q heads share one KV head; ncu: 3 % of HBM
// peak).  Here the q heads of a group are the rows of a tcgen05.mma, so the KV pages are read once and
// the per-token work moves to the tensor pipe:  S = Q_group K^T  and  O += P V  as in the prefill
// kernel, with K / V pages fetched by TMA straight from the paged cache ([block][kv head][16][d] ->
// one 16-row box per page) through the block table.
//
// Persistent: one CTA per SM walks (sequence, kv head, split) items round-robin; the TMA producer warp
// runs ahead across item boundaries through a 3..8 stage K/V ring, so the HBM stream never drains while
// the softmax warps finish an item.  Q is double-buffered and stored compactly (8*ceil(G/8) rows): the
// MMA still reads 128 rows, the extra rows are whatever follows in shared memory and only feed
// accumulator rows nobody reads.  Softmax warps whose 32 rows are all >= G do not run at all.
//
// Contract note: keys past a sequence's length are masked to -inf before the exponential, but the V rows of a
// partially filled page still enter the PV MMA multiplied by P = 0, so cache pages must hold finite values
// (0 * NaN = NaN).  The engine zero-initialises the pool and pages only ever hold real K/V, so this holds.
//
// Replaces SURVEY.md G8 (`baddbmm` + TorchScript softmax + `bmm` on a replicated KV tensor,
// gpt_bigcode_modeling.py:170-246) and G7 for grouped models.
#include "common.cuh"

#include <cuda.h>
#include <stdlib.h>
#include <string.h>

namespace {

constexpr int BM = 128;  // UMMA M (q-head rows; rows >= G are don't-care)
// keys per tile: 64 (= one 64-token page) for head_dim 128 / 256, 128 for head_dim 64 so that a tile is
// still 32 KB of K + V and the per-tile costs (barriers, MMA issue, softmax hand-off) stay amortised
__host__ __device__ constexpr int tile_keys(int d) { return d == 64 ? 128 : 64; }
constexpr int kThreads = 224;  // warp 0 TMA, warp 1 QK issue, warps 2-5 softmax, warp 6 PV issue
constexpr int kMaxStages = 8;

struct DArgs {
  const int* block_tables;
  const int* seq_lens;
  int max_blocks;
  void* out;
  int64_t ldo;
  float* part_o;
  float* part_ml;
  int hq, hkv, G;
  int q_box_rows;    // rows fetched per Q TMA op
  int q_copies;      // TMA ops per chunk: the same rows land at q_copies row offsets (physical replication)
  int q_alloc_rows;  // rows of one Q buffer in shared memory
  int q_sbo;         // stride between 8-row groups in the Q descriptors (0: every group re-reads rows 0..7)
  int ws;            // G <= 8, 64-token pages, 64-key tiles: weight-stationary 32-row MMAs (tcgen05.mma.ws), whose
                     // accumulator layout puts key / head-dim columns [q*N/4, (q+1)*N/4) into sub-partition q
  int block_size;
  int num_splits;
  int n_items;
  int stages;
  int B;
  int bn;         // keys per tile = tile_keys(head_dim)
  int balanced;   // 1: flat partition of all key tiles over the grid (ignores num_splits)
  int max_parts;  // balanced: partial slots per (sequence, q head)
  float scale_log2;
  int dtype;
};

struct Item {
  int b, kvh, split, L, t_begin, t_end, n_tiles;
  int direct;   // this unit covers the whole sequence: write the normalised output, no partial
  int part;     // partial slot within the (sequence, head) row
  int pstride;  // slots per (sequence, head)
};

constexpr int kMinQuota = 4;  // balanced mode: no CTA gets fewer key tiles than this (bounds the parts per item)

__device__ __forceinline__ int balanced_quota(int total_tiles, int grid) {
  int active = (total_tiles + kMinQuota - 1) / kMinQuota;
  active = active < grid ? active : grid;
  active = active > 1 ? active : 1;
  return (total_tiles + active - 1) / active > 1 ? (total_tiles + active - 1) / active : 1;
}

__device__ __forceinline__ Item decode_item(const DArgs& a, int it) {
  Item r;
  r.split = it % a.num_splits;
  const int t = it / a.num_splits;
  r.kvh = t % a.hkv;
  r.b = t / a.hkv;
  r.L = a.seq_lens[r.b];
  int chunk = (r.L + a.num_splits - 1) / a.num_splits;
  chunk = (chunk + a.bn - 1) / a.bn * a.bn;  // whole key tiles per split
  r.t_begin = r.split * chunk;
  r.t_end = min(r.L, r.t_begin + chunk);
  r.n_tiles = r.t_end > r.t_begin ? (r.t_end - r.t_begin + a.bn - 1) / a.bn : 0;
  r.direct = a.num_splits == 1;
  r.part = r.split;
  r.pstride = a.num_splits;
  return r;
}

// Work-unit iterator shared by all warp roles (every role walks the same deterministic sequence).
//  * split mode: units are (sequence, kv head, split) items, round-robin over the CTAs.
//  * balanced mode: all key tiles of the batch, ordered (sequence, kv head, tile), are cut into equal
//    contiguous ranges, one per CTA ("stream-K" for attention): a CTA's range may start and end in the
//    middle of an item, those pieces are written as partials and merged by the combine kernel.  Removes
//    the wave quantisation of items / SMs (e.g. 128 items on 148 SMs at TP=8).
struct Sched {
  const DArgs& a;
  int it;                             // split mode cursor
  int t, t_stop, b, kvh, j, quota;    // balanced mode cursor: flat tile index, (sequence, kv head, tile)
  __device__ __forceinline__ int ntiles(int i) const { return (a.seq_lens[i] + a.bn - 1) / a.bn; }
  __device__ Sched(const DArgs& a_) : a(a_), it(blockIdx.x), t(0), t_stop(0), b(0), kvh(0), j(0), quota(1) {
    if (!a.balanced) return;
    int total = 0;
    for (int i = 0; i < a.B; ++i) total += ntiles(i);
    total *= a.hkv;
    quota = balanced_quota(total, gridDim.x);
    const long long t0 = (long long)blockIdx.x * quota;
    t = t0 < total ? (int)t0 : total;
    t_stop = t + quota < total ? t + quota : total;
    int rem = t;
    while (b < a.B) {
      const int span = ntiles(b) * a.hkv;
      if (rem < span) break;
      rem -= span;
      ++b;
    }
    if (b < a.B) {
      const int n = ntiles(b);
      kvh = rem / n;
      j = rem % n;
    }
  }
  __device__ bool next(Item& r) {
    if (!a.balanced) {
      if (it >= a.n_items) return false;
      r = decode_item(a, it);
      it += gridDim.x;
      return true;
    }
    if (t >= t_stop) return false;
    int n = ntiles(b);
    while (n == 0) n = ntiles(++b);  // t < total: a non-empty sequence follows
    const int take = (n - j) < (t_stop - t) ? (n - j) : (t_stop - t);
    r.b = b;
    r.kvh = kvh;
    r.split = 0;
    r.L = a.seq_lens[b];
    r.t_begin = j * a.bn;
    r.t_end = r.L < (j + take) * a.bn ? r.L : (j + take) * a.bn;
    r.n_tiles = take;
    r.direct = take == n;
    r.part = (int)blockIdx.x - (t - j) / quota;  // CTAs that share this item are consecutive
    r.pstride = a.max_parts;
    t += take;
    j += take;
    if (j == n) {
      j = 0;
      if (++kvh == a.hkv) { kvh = 0; ++b; }
    }
    return true;
  }
};

// MN-major (N contiguous) 128B-swizzled operand: rows are K (keys), 64 elements of N per 128-byte row;
// LBO = distance between 64-element N chunks, SBO = distance between 8-row K groups.
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__host__ __device__ constexpr uint32_t idesc_f16(int fmt, int M, int N, int b_mn_major) {
  return (1u << 4) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ float ex2_approx(float x) {  // 2^x, one MUFU; 2^-inf = 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <typename T>
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  if constexpr (std::is_same<T, __half>::value) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  } else {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
}


__device__ __forceinline__ void tma_load_3d(void* smem_dst, const void* map, uint64_t* bar, int c0, int c1, int c2,
                                            uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(smem_u32(smem_dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "l"(policy)
      : "memory");
}

__device__ __forceinline__ uint64_t umma_desc_k_sw128_sbo(uint32_t smem_addr, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// tcgen05.mma.ws (weight-stationary form, M = 32 / 64 / 128, N = 64 / 128 / 256).  For M = 32 element (m, n) of D lives
// in TMEM lane m + 32 * (n / (N/4)), column n % (N/4).
__device__ __forceinline__ void umma_f16_ss_ws(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.ws.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ void softmax_bar(int threads) { asm volatile("bar.sync 1, %0;" ::"r"(threads) : "memory"); }

// R = how many times the q-head rows are replicated down the 128 MMA rows.  A softmax warp can only read
// the 32 TMEM lanes of its own SM sub-partition, so with G <= 32 heads one warp (one sub-partition's
// MUFU) would do all the exponentials.  Replicating the rows R times lets the warps of copy c handle the
// key columns [c * 64 / R, (c + 1) * 64 / R) of the same heads: the row max is agreed through shared
// memory once per tile, P rows are assembled in shared memory by all copies, l is summed at the end.
template <typename T, int D, int R>
__global__ void __launch_bounds__(kThreads, 1)
attn_decode_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                      const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ DArgs a) {
  constexpr int DSUB = D / 64;
  constexpr int BN = tile_keys(D);
  constexpr int P_BYTES = BM * BN * 2;
  constexpr int KV_BYTES = BN * D * 2;  // one K (or V) tile
  constexpr int TMEM_COLS = (2 * BN + D <= 256) ? 256 : 512;  // S0 at column 0, S1 at BN, O at 2*BN
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int S = a.stages;
  const int q_bytes = a.q_alloc_rows * D * 2;
  const int q_chunk = a.q_alloc_rows * 128;  // bytes between the 64-column chunks of a Q buffer
  uint8_t* sQ = smem;                    // 2 buffers
  uint8_t* sP = sQ + 2 * q_bytes;
  uint8_t* sK = sP + 2 * P_BYTES;        // S stages (P is double-buffered)
  uint8_t* sV = sK + S * KV_BYTES;       // S stages
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + S * KV_BYTES);
  uint64_t* bar_qfull = bars;            // [2]
  uint64_t* bar_qempty = bars + 2;       // [2]
  uint64_t* bar_s = bars + 4;            // [2]
  uint64_t* bar_p = bars + 6;            // [2] one per P buffer
  uint64_t* bar_o = bars + 8;            // [2] PV done, one per P buffer
  uint64_t* bar_sfree = bars + 10;       // [2] softmax has copied S(g) out of tensor memory
  uint64_t* bar_kfull = bars + 12;
  uint64_t* bar_vfull = bar_kfull + kMaxStages;
  uint64_t* bar_kempty = bar_vfull + kMaxStages;
  uint64_t* bar_vempty = bar_kempty + kMaxStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_vempty + kMaxStages);
  float* s_mx = reinterpret_cast<float*>(tmem_slot + 4);  // [2][4][32] per-tile local row max of every warp
  float* s_l = s_mx + 256;                                // [4][32] per-item partial row sums

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int G = a.G;
  pdl_launch_dependents();
  constexpr int QPC = 4 / R;  // softmax warps (TMEM lane quarters) per copy
  int n_live_warps = 0;
  for (int q = 0; q < 4; ++q) n_live_warps += ((q % QPC) * 32 < G) ? 1 : 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_q);
    tma_prefetch_desc(&tmap_k);
    tma_prefetch_desc(&tmap_v);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar_qfull[i], 1);
      mbar_init(&bar_qempty[i], 1);
      mbar_init(&bar_s[i], 1);
      mbar_init(&bar_p[i], n_live_warps);
      mbar_init(&bar_o[i], 1);
      mbar_init(&bar_sfree[i], n_live_warps);
    }
    for (int s = 0; s < S; ++s) {
      mbar_init(&bar_kfull[s], 1);
      mbar_init(&bar_vfull[s], 1);
      mbar_init(&bar_kempty[s], 1);
      mbar_init(&bar_vempty[s], 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tO = tmem_base + 2 * BN;
  pdl_wait();  // q, the KV cache and the step metadata come from earlier kernels of the stream

  if (warp == 0) {
    // ===================== TMA producer: runs ahead across items through the K/V ring =====================
    const uint64_t pol_q = l2_policy_evict_last();
    const uint64_t pol_kv = l2_policy_evict_first();  // every cached byte is read once per step
    const int pages = BN / a.block_size;
    int stage = 0, n = 0;
    uint32_t phase = 0;
    Sched sch(a);
    Item I;
    while (sch.next(I)) {
      if (I.n_tiles == 0) continue;
      const int qb = n & 1;
      mbar_wait(&bar_qempty[qb], ((n >> 1) & 1) ^ 1, 40);
      if (elect_one()) {
        mbar_arrive_expect_tx(&bar_qfull[qb], (uint32_t)(a.q_copies * a.q_box_rows * D * 2));
        for (int cc = 0; cc < a.q_copies; ++cc)
#pragma unroll
          for (int c = 0; c < DSUB; ++c)
            tma_load_3d(sQ + qb * q_bytes + c * q_chunk + cc * (a.q_box_rows * 128), &tmap_q, &bar_qfull[qb], c * 64,
                        I.kvh * G, I.b, pol_q);
      }
      // Block-table entries are fetched 32 pages at a time (one coalesced load per warp, the next group
      // prefetched while this one is issued); each lane keeps the cache row of one page.
      const int* bt = a.block_tables + (int64_t)I.b * a.max_blocks;
      const int bs = a.block_size;
      const int last_page = (I.L - 1) / bs;
      const int page_begin = I.t_begin / bs;
      const int tiles_per_group = 32 / pages;
      auto load_rows = [&](int group) {
        const int pg = page_begin + group * 32 + lane;
        // pages past the end of the sequence are masked later: fetch the reserved block 0 instead
        const int blk = (pg <= last_page) ? __ldg(bt + pg) : 0;
        return (blk * a.hkv + I.kvh) * bs;
      };
      int cur = load_rows(0), nxt = 0;
      for (int j = 0; j < I.n_tiles; ++j) {
        const int tg = j % tiles_per_group;
        if (tg == 0) {
          if (j > 0) cur = nxt;
          if (j + tiles_per_group < I.n_tiles) nxt = load_rows(j / tiles_per_group + 1);
        }
        mbar_wait(&bar_kempty[stage], phase ^ 1, 50);
        if (elect_one()) mbar_arrive_expect_tx(&bar_kfull[stage], KV_BYTES);
        __syncwarp();
        for (int p = 0; p < pages; ++p) {
          const int row = __shfl_sync(0xffffffffu, cur, tg * pages + p);
          if (elect_one())
            tma_load_3d(sK + stage * KV_BYTES + p * (bs * D * 2), &tmap_k, &bar_kfull[stage], 0, row, 0, pol_kv);
        }
        mbar_wait(&bar_vempty[stage], phase ^ 1, 51);
        if (elect_one()) mbar_arrive_expect_tx(&bar_vfull[stage], KV_BYTES);
        __syncwarp();
        for (int p = 0; p < pages; ++p) {
          const int row = __shfl_sync(0xffffffffu, cur, tg * pages + p);
          if (elect_one())
            tma_load_3d(sV + stage * KV_BYTES + p * (bs * D * 2), &tmap_v, &bar_vfull[stage], 0, row, 0, pol_kv);
        }
        if (++stage == S) { stage = 0; phase ^= 1; }
      }
      ++n;
    }
  } else if (warp == 1) {
    // ===================== QK issuer: S(g) = Q K(g)^T as soon as K(g) has landed and S buffer g&1 is free.
    // The whole warp runs the loop, one elected lane issues (see elect_one).  QK and PV are issued by
    // different warps so that S runs two tiles ahead of the softmax instead of queueing behind PV.
    const int fmt = a.dtype == LLMSS_BF16 ? 1 : 0;
    const int bs = a.block_size, pages = BN / bs;
    // G <= 8 (q_sbo == 0: every 8-row group re-reads the same q rows): a 64-row MMA is enough -- its rows
    // 16q..16q+15 land in lanes 0..15 of sub-partition q, so each softmax warp still finds heads 0..G-1 in
    // its lanes 0..G-1 -- and it halves the tensor-pipe work (and power) spent on replicated rows.
    const int mma_m = (a.q_sbo == 0) ? 64 : BM;
    const uint32_t idesc_s = idesc_f16(fmt, mma_m, bs, 0);  // one MMA group per page: N = block_size keys
    const uint32_t k_base = smem_u32(sK), q_base = smem_u32(sQ);
    const uint32_t page_bytes = bs * D * 2, chunk_bytes = bs * 128;
    int kst = 0, n = 0;
    uint32_t kph = 0, g = 0;
    Sched sch(a);
    Item I;
    while (sch.next(I)) {
      if (I.n_tiles == 0) continue;
      const int qb = n & 1;
      const uint32_t q_addr = q_base + qb * q_bytes;
      mbar_wait(&bar_qfull[qb], (n >> 1) & 1, 60);
      for (int j = 0; j < I.n_tiles; ++j, ++g) {
        if (g >= 2) mbar_wait(&bar_sfree[g & 1], ((g >> 1) & 1) ^ 1, 64);
        mbar_wait(&bar_kfull[kst], kph, 61);
        tcgen05_fence_after();
        const uint32_t k_addr = k_base + kst * KV_BYTES;
        if (a.ws) {
          if (elect_one()) {  // one 64-key page per tile: S[32 x 64], 16 TMEM columns per buffer
            const uint32_t idesc_ws = idesc_f16(fmt, 32, BN, 0);
#pragma unroll
            for (int c = 0; c < DSUB; ++c)
#pragma unroll
              for (int k = 0; k < 4; ++k)
                umma_f16_ss_ws(tS + (g & 1) * 16, umma_desc_k_sw128_sbo(q_addr + c * q_chunk + k * 32, a.q_sbo),
                               umma_desc_k_sw128(k_addr + c * chunk_bytes + k * 32), idesc_ws, (c | k) ? 1u : 0u);
            umma_commit(&bar_s[g & 1]);
            umma_commit(&bar_kempty[kst]);
            if (j + 1 == I.n_tiles) umma_commit(&bar_qempty[qb]);
          }
        } else if (elect_one()) {
          for (int p = 0; p < pages; ++p)
#pragma unroll
            for (int c = 0; c < DSUB; ++c)
#pragma unroll
              for (int k = 0; k < 4; ++k)
                umma_f16_ss(tS + (g & 1) * BN + p * bs, umma_desc_k_sw128_sbo(q_addr + c * q_chunk + k * 32, a.q_sbo),
                            umma_desc_k_sw128(k_addr + p * page_bytes + c * chunk_bytes + k * 32), idesc_s,
                            (c | k) ? 1u : 0u);
          umma_commit(&bar_s[g & 1]);
          umma_commit(&bar_kempty[kst]);
          if (j + 1 == I.n_tiles) umma_commit(&bar_qempty[qb]);
        }
        __syncwarp();
        if (++kst == S) { kst = 0; kph ^= 1; }
      }
      ++n;
    }
  } else if (warp == 6) {
    // ===================== PV issuer: O += P(g) V(g) once the softmax warps have published P(g) ==========
    const int fmt = a.dtype == LLMSS_BF16 ? 1 : 0;
    const int bs = a.block_size;
    const uint32_t idesc_o = idesc_f16(fmt, (a.q_sbo == 0) ? 64 : BM, D, 1);  // see the QK issuer: 64 rows when G <= 8
    const uint32_t p_addr = smem_u32(sP), v_base = smem_u32(sV);
    const uint32_t page_bytes = bs * D * 2, chunk_bytes = bs * 128;
    int vst = 0;
    uint32_t vph = 0, g = 0;
    Sched sch(a);
    Item I;
    while (sch.next(I)) {
      for (int j = 0; j < I.n_tiles; ++j, ++g) {
        mbar_wait(&bar_p[g & 1], (g >> 1) & 1, 62);
        mbar_wait(&bar_vfull[vst], vph, 63);
        tcgen05_fence_after();
        const uint32_t v_addr = v_base + vst * KV_BYTES;
        if (a.ws) {
          if (elect_one()) {  // O[32 x D]: head-dim columns [q*D/4, (q+1)*D/4) land in sub-partition q
            const uint32_t idesc_ws = idesc_f16(fmt, 32, D, 1);
#pragma unroll
            for (int k = 0; k < BN / 16; ++k)
              umma_f16_ss_ws(tO, umma_desc_k_sw128(p_addr + (g & 1) * P_BYTES + k * 32),
                             umma_desc_mn_sw128(v_addr + (k * 16) * 128, chunk_bytes), idesc_ws, (j > 0 || k > 0) ? 1u : 0u);
            umma_commit(&bar_o[g & 1]);
            umma_commit(&bar_vempty[vst]);
          }
        } else if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BN / 16; ++k) {  // 16 keys per step; a step never straddles a page (block_size >= 16)
            const int pg = (k * 16) / bs, ro = (k * 16) % bs;
            umma_f16_ss(tO, umma_desc_k_sw128(p_addr + (g & 1) * P_BYTES + (k >> 2) * (BM * 128) + (k & 3) * 32),
                        umma_desc_mn_sw128(v_addr + pg * page_bytes + ro * 128, chunk_bytes), idesc_o,
                        (j > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&bar_o[g & 1]);
          umma_commit(&bar_vempty[vst]);
        }
        __syncwarp();
        if (++vst == S) { vst = 0; vph ^= 1; }
      }
    }
  } else if (warp < 6 && ((warp & 3) % QPC) * 32 < G) {
    // ===================== softmax: copy `cidx` of row `row`, key columns [c0, c0 + CW) =====================
    constexpr int CW = BN / R;  // key columns per thread
    const int quarter = warp & 3;
    const int cidx = quarter / QPC;
    const int row = (quarter % QPC) * 32 + lane;  // q head within the group == P / O row
    const bool live = row < G;
    const bool owner = cidx == 0;                 // copy 0 owns the accumulator rows (rescale + output)
    const int c0 = cidx * CW;
    const int n_threads = n_live_warps * 32;
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    uint32_t g = 0;
    Sched sch(a);
    Item I;
    while (sch.next(I)) {
      const int head = I.kvh * G + row;
      const int64_t pidx = ((int64_t)I.b * a.hq + head) * I.pstride + I.part;
      if (I.n_tiles == 0) {  // empty split (or empty sequence): neutral partial / zero output
        if (live && owner) {
          if (I.direct) {
            T* orow = reinterpret_cast<T*>(a.out) + (int64_t)I.b * a.ldo + (int64_t)head * D;
            for (int c = 0; c < D; c += 8) *reinterpret_cast<uint4*>(orow + c) = make_uint4(0, 0, 0, 0);
          } else {
            for (int c = 0; c < D; c += 4) *reinterpret_cast<float4*>(a.part_o + pidx * D + c) = make_float4(0, 0, 0, 0);
            a.part_ml[pidx * 2] = -INFINITY;
            a.part_ml[pidx * 2 + 1] = 0.f;
          }
        }
        continue;
      }
      float m_run = -INFINITY, l_run = 0.f;
      for (int j = 0; j < I.n_tiles; ++j, ++g) {
        mbar_wait(&bar_s[g & 1], (g >> 1) & 1, 70);
        tcgen05_fence_after();
        const int kbase = I.t_begin + j * BN + c0;
        float sc[CW];
        if constexpr (CW == 16) {
          uint32_t r[16];
          // ws: this sub-partition already holds exactly its 16 key columns, at columns [0, 16) of the buffer
          tmem_ld_32x32b_x16(a.ws ? tS + (g & 1) * 16 + lane_off : tS + (g & 1) * BN + lane_off + c0, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) sc[i] = __uint_as_float(r[i]);
        } else {
#pragma unroll
          for (int c = 0; c < CW / 32; ++c) {
            uint32_t r[32];
            tmem_ld_32x32b_x32(tS + (g & 1) * BN + lane_off + c0 + c * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) sc[c * 32 + i] = __uint_as_float(r[i]);
          }
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_sfree[g & 1]);  // S(g) is in registers: the buffer can take QK(g+2)
        if (kbase + CW > I.t_end) {
#pragma unroll
          for (int i = 0; i < CW; ++i)
            if (kbase + i >= I.t_end) sc[i] = -INFINITY;
        }
        float m_tile = -INFINITY;
#pragma unroll
        for (int i = 0; i < CW; ++i) m_tile = fmaxf(m_tile, sc[i]);
        if constexpr (R > 1) {  // agree on the row max across the copies
          float* mx = s_mx + (g & 1) * 128;
          mx[quarter * 32 + lane] = m_tile;
          softmax_bar(n_threads);
#pragma unroll
          for (int cc = 0; cc < R; ++cc) m_tile = fmaxf(m_tile, mx[(cc * QPC + quarter % QPC) * 32 + lane]);
        }
        // Lazy rescale: keep the old running max unless the tile exceeds it by more than 2^8, so the
        // accumulator is rarely touched and softmax(g+1) does not wait for PV(g).
        const float m_tile_s = m_tile * a.scale_log2;
        const bool grow = (j == 0) || (m_tile_s > m_run + 8.0f);
        const float m_new = grow ? fmaxf(m_run, m_tile_s) : m_run;
        float alpha = (j == 0 || !grow) ? 1.0f : ex2_approx(m_run - m_new);
        if (!live) alpha = 1.0f;  // don't-care rows may hold NaN; keep them out of the warp vote
        m_run = m_new;
        const float neg_m = (m_new == -INFINITY) ? 0.f : -m_new;
        if (g >= 2) mbar_wait(&bar_o[g & 1], ((g >> 1) & 1) ^ 1, 71);  // PV(g-2) has released this P buffer
        float l_tile = 0.f;
        uint8_t* sub = sP + (g & 1) * P_BYTES + row * 128;
#pragma unroll
        for (int c = 0; c < CW / 8; ++c) {
          uint32_t pk[4];
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            const float p0 = ex2_approx(fmaf(sc[c * 8 + i], a.scale_log2, neg_m));
            const float p1 = ex2_approx(fmaf(sc[c * 8 + i + 1], a.scale_log2, neg_m));
            l_tile += p0 + p1;
            pk[i >> 1] = pack2<T>(p0, p1);
          }
          const int gc = c0 / 8 + c;  // 16-byte chunk of the P row: 64-key sub-tile gc / 8, swizzled slot within it
          const int chunk16 = (gc & 7) ^ (row & 7);
          if (live) *reinterpret_cast<uint4*>(sub + (gc >> 3) * (BM * 128) + chunk16 * 16) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        }
        // ws: the accumulator's head-dim columns are spread over the four sub-partitions, every warp owns D/4
        const int o_cols = a.ws ? D / 4 : D;
        if ((owner || a.ws) && j > 0 && __any_sync(0xffffffffu, alpha != 1.0f)) {
          mbar_wait(&bar_o[(g - 1) & 1], ((g - 1) >> 1) & 1, 73);  // PV(g-1) done: O is stable
          tcgen05_fence_after();
#pragma unroll 1
          for (int c = 0; c < o_cols / 32; ++c) {
            uint32_t r[32];
            tmem_ld_32x32b_x32(tO + lane_off + c * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
            tmem_st_32x32b_x32(tO + lane_off + c * 32, r);
          }
          tmem_st_wait();
        }
        l_run = l_run * alpha + l_tile;
        tcgen05_fence_before();
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_p[g & 1]);
      }
      if constexpr (R > 1) {  // total row sum = sum over the copies' column slices
        s_l[quarter * 32 + lane] = l_run;
        softmax_bar(n_threads);
        if (owner || a.ws) {
          l_run = 0.f;
#pragma unroll
          for (int cc = 0; cc < R; ++cc) l_run += s_l[(cc * QPC + quarter % QPC) * 32 + lane];
        }
        softmax_bar(n_threads);  // s_l may be rewritten by the next item
      }
      if (!owner && !a.ws) continue;
      mbar_wait(&bar_o[(g - 1) & 1], ((g - 1) >> 1) & 1, 72);
      tcgen05_fence_after();
      const float inv_l = l_run > 0.f ? 1.0f / l_run : 0.f;
      const int e_cols = a.ws ? D / 4 : D;            // head-dim columns this warp holds ...
      const int e_base = a.ws ? quarter * (D / 4) : 0;  // ... starting at this head-dim column
#pragma unroll 1
      for (int c = 0; c < e_cols / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tO + lane_off + c * 32, r);
        tmem_ld_wait();
        if (!live) continue;
        if (I.direct) {
          T* orow = reinterpret_cast<T*>(a.out) + (int64_t)I.b * a.ldo + (int64_t)head * D + e_base;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float f8[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) f8[i] = __uint_as_float(r[q * 8 + i]) * inv_l;
            *reinterpret_cast<uint4*>(orow + c * 32 + q * 8) = pack8<T>(f8);
          }
        } else {
          float4* dst = reinterpret_cast<float4*>(a.part_o + pidx * D + e_base + c * 32);
#pragma unroll
          for (int q = 0; q < 8; ++q)
            dst[q] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]), __uint_as_float(r[4 * q + 2]),
                                 __uint_as_float(r[4 * q + 3]));
        }
      }
      if (live && !I.direct && owner) {
        a.part_ml[pidx * 2] = m_run;
        a.part_ml[pidx * 2 + 1] = l_run;
      }
      tcgen05_fence_before();  // O reads retire before the next item's first PV overwrites the accumulator
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// Balanced mode: merge the pieces of the items that were cut at a CTA boundary (and zero the rows of empty
// sequences).  Recomputes the same partition as Sched from seq_lens.
template <typename T, int D>
__global__ void attn_decode_tc_balanced_combine_kernel(const float* __restrict__ part_o, const float* __restrict__ part_ml,
                                                       T* __restrict__ out, int64_t ldo, const int* __restrict__ seq_lens,
                                                       int B, int hq, int hkv, int max_parts, int grid_main, int bn) {
  pdl_launch_dependents();
  pdl_wait();
  const int head = blockIdx.x, b = blockIdx.y;
  const int kvh = head / (hq / hkv);
  int total = 0, pre = 0;
  for (int i = 0; i < B; ++i) {
    const int n = (seq_lens[i] + bn - 1) / bn;
    if (i < b) pre += n;
    total += n;
  }
  total *= hkv;
  const int n = (seq_lens[b] + bn - 1) / bn;
  T* orow = out + (int64_t)b * ldo + (int64_t)head * D;
  if (n == 0) {
    for (int c = threadIdx.x; c < D; c += blockDim.x) orow[c] = from_f32<T>(0.f);
    return;
  }
  const int quota = balanced_quota(total, grid_main);
  const int start = pre * hkv + kvh * n;
  const int first = start / quota, last = (start + n - 1) / quota;
  const int parts = last - first + 1;
  if (parts == 1) return;  // written directly by the attention kernel
  const int64_t base = ((int64_t)b * hq + head) * max_parts;
  float mn = -INFINITY;
  for (int s = 0; s < parts; ++s) mn = fmaxf(mn, part_ml[(base + s) * 2]);
  for (int c = threadIdx.x; c < D; c += blockDim.x) {
    float lt = 0.f, at = 0.f;
    for (int s = 0; s < parts; ++s) {
      const float ms = part_ml[(base + s) * 2];
      const float sc = (ms == -INFINITY) ? 0.f : exp2f(ms - mn);
      lt += part_ml[(base + s) * 2 + 1] * sc;
      at += part_o[(base + s) * D + c] * sc;
    }
    orow[c] = from_f32<T>(lt > 0.f ? at / lt : 0.f);
  }
}

typedef CUresult (*PFN_encodeTiledD)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                     const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                     CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
PFN_encodeTiledD get_encode_d() {
  static PFN_encodeTiledD fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiledD)p;
  }
  return fn;
}

// Upper bound of the pieces one (sequence, head) item can be cut into in balanced mode.
int balanced_max_parts(int max_blocks, int block_size, int bn) {
  const int max_tiles = (max_blocks * block_size + bn - 1) / bn;
  return (max_tiles + kMinQuota - 2) / kMinQuota + 1;
}

int sm_count() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  return sms;
}

constexpr int kSmemBudget = 227 * 1024;

int plan_stages(int q_alloc_rows, int d, int* smem_bytes) {
  const int bn = tile_keys(d);
  const int fixed = 1024 /*alignment*/ + 2048 /*barriers + softmax exchange*/ + 2 * q_alloc_rows * d * 2 + 2 * (BM * bn * 2);
  const int per_stage = 2 * bn * d * 2;
  int stages = (kSmemBudget - fixed) / per_stage;
  if (stages > kMaxStages) stages = kMaxStages;
  *smem_bytes = fixed + stages * per_stage;
  return stages;
}

template <typename T, int D, int R>
int launch_decode_tc(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const DArgs& a, int smem,
                     cudaStream_t stream) {
  static PerDeviceSmem configured;
  auto kern = attn_decode_tc_kernel<T, D, R>;
  LLMSS_CHECK_CUDA(configured.ensure(kern, smem));
  const int grid = (a.balanced || a.n_items >= sm_count()) ? sm_count() : a.n_items;
  LLMSS_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(kThreads), (size_t)smem, stream, tq, tk, tv, a));
  return 0;
}

}  // namespace

// Scratch (floats) of the balanced mode (num_splits = 0).
LLMSS_API int64_t llmss_attn_decode_tc_balanced_scratch_floats(int B, int hq, int d, int max_blocks, int block_size) {
  return (int64_t)B * hq * balanced_max_parts(max_blocks, block_size, tile_keys(d)) * (d + 2);
}

extern "C" int llmss_attn_decode_combine(const float* part_o, const float* part_ml, void* out, int64_t ldo, int B, int hq,
                                         int d, int num_splits, int dtype, cudaStream_t stream);

// Split factor for the persistent tensor-core kernel: CTAs walk items round-robin, so the cost of a
// candidate is rounds * (tiles per item + per-item overhead), plus the combine pass when it splits.
LLMSS_API int llmss_attn_decode_tc_pick_splits(int B, int hkv, int d, int max_kv) {
  const int BN = tile_keys(d);
  const int sms = sm_count();
  const int items = B * hkv;
  if (items <= 0) return 1;
  const int tiles = (max_kv + BN - 1) / BN;
  int max_s = tiles / 2;
  if (max_s < 1) max_s = 1;
  if (max_s > 32) max_s = 32;
  int best = 1;
  double best_cost = 1e30;
  for (int sp = 1; sp <= max_s; ++sp) {
    const int rounds = (items * sp + sms - 1) / sms;
    const int per = (tiles + sp - 1) / sp;
    const double cost = rounds * (per + 1.5) + (sp > 1 ? 2.0 + 0.15 * sp : 0.0);
    if (cost < best_cost - 1e-9) { best_cost = cost; best = sp; }
  }
  return best;
}

// Tensor-core (tcgen05) paged decode attention: the q heads of a KV group are MMA rows.  Same contract
// as llmss_attn_decode; `scratch` holds the split-KV partials.  Requires block_size in {16, 32, 64}, head_dim in
// {64, 128, 256}, hq / hkv <= 128.
LLMSS_API int llmss_attn_decode_tc(const void* q, int64_t ldq, const void* k_cache, const void* v_cache,
                                   const int* block_tables, int max_blocks, const int* seq_lens, void* out,
                                   int64_t ldo, float* scratch, int B, int hq, int hkv, int d, int block_size,
                                   int num_blocks, float scale, int num_splits, int dtype, cudaStream_t stream) {
  LLMSS_REQUIRE(dtype == LLMSS_F16 || dtype == LLMSS_BF16, "attn_decode_tc: dtype must be f16/bf16");
  LLMSS_REQUIRE(d == 64 || d == 128 || d == 256, "attn_decode_tc: head_dim=%d unsupported", d);
  LLMSS_REQUIRE(hq > 0 && hkv > 0, "attn_decode_tc: head counts must be positive (hq=%d hkv=%d)", hq, hkv);
  LLMSS_REQUIRE(hq % hkv == 0 && hq / hkv <= 128, "attn_decode_tc: bad head grouping hq=%d hkv=%d", hq, hkv);
  LLMSS_REQUIRE(block_size == 16 || block_size == 32 || block_size == 64, "attn_decode_tc: block_size must be 16/32/64");
  LLMSS_REQUIRE(num_splits >= 0 && (num_splits == 1 || scratch != nullptr), "attn_decode_tc: scratch required for splits");
  const int balanced = num_splits == 0 ? 1 : 0;  // flat partition of all key tiles over the SMs
  const int max_parts = balanced_max_parts(max_blocks, block_size, tile_keys(d));
  LLMSS_REQUIRE(ldq % 8 == 0 && ldo % 8 == 0 && ((uintptr_t)q % 16) == 0, "attn_decode_tc: q must be 16-byte aligned");
  if (B == 0) return 0;
  PFN_encodeTiledD enc = get_encode_d();
  LLMSS_REQUIRE(enc != nullptr, "attn_decode_tc: cuTensorMapEncodeTiled entry point not available");
  const int G = hq / hkv;
  // Row replication plan (see the kernel comment): R copies of the group's heads down the 128 MMA rows.
  int R, q_box_rows, q_copies, q_alloc_rows, q_sbo;
  if (G <= 8) { R = 4; q_box_rows = 8; q_copies = 1; q_alloc_rows = 8; q_sbo = 0; }
  else if (G <= 32) { R = 4; q_box_rows = 32; q_copies = 4; q_alloc_rows = 128; q_sbo = 1024; }
  else if (G <= 64) { R = 2; q_box_rows = 64; q_copies = 2; q_alloc_rows = 128; q_sbo = 1024; }
  else { R = 1; q_box_rows = (G + 7) / 8 * 8; q_copies = 1; q_alloc_rows = q_box_rows; q_sbo = 1024; }
  int smem = 0;
  const int stages = plan_stages(q_alloc_rows, d, &smem);
  LLMSS_REQUIRE(stages >= 2, "attn_decode_tc: shared memory too small for hq/hkv=%d d=%d", G, d);
  const CUtensorMapDataType dt = dtype == LLMSS_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  CUtensorMap tq, tk, tv;
  {
    // q viewed as [B][hq][d]: box = [1][q_box_rows][64]
    cuuint64_t dims[3] = {(cuuint64_t)d, (cuuint64_t)hq, (cuuint64_t)B};
    cuuint64_t strides[2] = {(cuuint64_t)d * 2, (cuuint64_t)ldq * 2};
    cuuint32_t box[3] = {64, (cuuint32_t)q_box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(&tq, dt, 3, const_cast<void*>(q), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    LLMSS_REQUIRE(r == CUDA_SUCCESS, "attn_decode_tc: q tensor map failed (CUresult %d)", (int)r);
  }
  for (int which = 0; which < 2; ++which) {
    // cache viewed as [d/64 chunks][num_blocks * hkv * block_size rows][64]: one box = one page of one head,
    // landing in shared memory chunk-major ([chunk][block_size rows][128 B]), which is what the MMA
    // descriptors of a page expect -- a single TMA op per page whatever the head size.
    cuuint64_t dims[3] = {64, (cuuint64_t)num_blocks * hkv * block_size, (cuuint64_t)(d / 64)};
    cuuint64_t strides[2] = {(cuuint64_t)d * 2, 128};
    cuuint32_t box[3] = {64, (cuuint32_t)block_size, (cuuint32_t)(d / 64)};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(which == 0 ? &tk : &tv, dt, 3, const_cast<void*>(which == 0 ? k_cache : v_cache), dims, strides, box,
                     estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    LLMSS_REQUIRE(r == CUDA_SUCCESS, "attn_decode_tc: cache tensor map failed (CUresult %d)", (int)r);
  }
  DArgs a;
  memset(&a, 0, sizeof(a));
  a.block_tables = block_tables;
  a.seq_lens = seq_lens;
  a.max_blocks = max_blocks;
  a.out = out;
  a.ldo = ldo;
  a.part_o = scratch;
  a.part_ml = scratch ? scratch + (int64_t)B * hq * (balanced ? max_parts : num_splits) * d : nullptr;
  a.B = B;
  a.bn = tile_keys(d);
  a.balanced = balanced;
  a.max_parts = max_parts;
  a.hq = hq;
  a.hkv = hkv;
  a.G = G;
  a.q_box_rows = q_box_rows;
  a.q_copies = q_copies;
  a.q_alloc_rows = q_alloc_rows;
  a.q_sbo = q_sbo;
  {
    const char* e = getenv("LLMSS_ATTN_WS");  // LLMSS_ATTN_WS=0 falls back to the 64-row tcgen05.mma form
    a.ws = (!(e != nullptr && e[0] == '0') && G <= 8 && block_size == 64 && tile_keys(d) == 64) ? 1 : 0;
  }
  a.block_size = block_size;
  a.num_splits = num_splits;
  a.n_items = B * hkv * num_splits;
  a.stages = stages;
  a.scale_log2 = scale * 1.4426950408889634f;
  a.dtype = dtype;
  int rc = 0;
#define LLMSS_DT(T_, D_)                                                         \
  do {                                                                           \
    if (R == 4) rc = launch_decode_tc<T_, D_, 4>(tq, tk, tv, a, smem, stream);      \
    else if (R == 2) rc = launch_decode_tc<T_, D_, 2>(tq, tk, tv, a, smem, stream); \
    else rc = launch_decode_tc<T_, D_, 1>(tq, tk, tv, a, smem, stream);             \
  } while (0)
  if (dtype == LLMSS_F16) {
    if (d == 64) LLMSS_DT(__half, 64);
    else if (d == 128) LLMSS_DT(__half, 128);
    else LLMSS_DT(__half, 256);
  } else {
    if (d == 64) LLMSS_DT(__nv_bfloat16, 64);
    else if (d == 128) LLMSS_DT(__nv_bfloat16, 128);
    else LLMSS_DT(__nv_bfloat16, 256);
  }
#undef LLMSS_DT
  if (rc != 0 || num_splits == 1) return rc;
  if (balanced) {
    const dim3 cgrid(hq, B);
    const int threads = d < 128 ? d : 128;
#define LLMSS_BC(T_, D_)                                                                                             \
  LLMSS_CHECK_CUDA(launch_pdl(attn_decode_tc_balanced_combine_kernel<T_, D_>, cgrid, dim3(threads), 0, stream, \
                              (const float*)a.part_o, (const float*)a.part_ml, (T_*)out, ldo, seq_lens, B, hq, hkv,  \
                              max_parts, sm_count(), a.bn))
    if (dtype == LLMSS_F16) {
      if (d == 64) LLMSS_BC(__half, 64); else if (d == 128) LLMSS_BC(__half, 128); else LLMSS_BC(__half, 256);
    } else {
      if (d == 64) LLMSS_BC(__nv_bfloat16, 64); else if (d == 128) LLMSS_BC(__nv_bfloat16, 128); else LLMSS_BC(__nv_bfloat16, 256);
    }
#undef LLMSS_BC
    return 0;
  }
  return llmss_attn_decode_combine(a.part_o, a.part_ml, out, ldo, B, hq, d, num_splits, dtype, stream);
}
