// K8 attn_prefill: causal varlen flash-attention forward on the 5th-generation tensor cores.
//
// Replaces the reference's prefill attention (cuBLAS QK^T into a materialised
// [B,h,S,T] fp32 score tensor, where-mask from a [max_pos,max_pos] bool buffer, softmax, cuBLAS PV;
// gptj_modeling.py:128-169, gpt_bigcode_modeling.py:170-246) with one kernel that never
// materialises S^2 data:
//
//   one CTA = 128 queries of one (sequence, q head); it walks the causal range of KV tiles.
//   warp 0      TMA producer: Q once, then K_j / V_j tiles (2-stage ring), 128B-swizzled
//   warp 1      issues  S_j = Q K_j^T  (A,B K-major) -> TMEM [128 x BN] fp32, double-buffered: S runs up
//               to two tiles ahead of the softmax (gated only by "softmax has copied S out of TMEM")
//   warp 6      issues  O += P_j V_j   (B = V, MN-major) -> TMEM [128 x D] fp32, as soon as P_j is published
//   warps 2-5   one thread per query row: tcgen05.ld its S row, scale + causal mask, online softmax in
//               fp32 (no shuffles: the row is thread-private), write P_j (bf16/fp16) into one of two
//               shared-memory buffers in the UMMA K-major swizzled layout.  The running max is lazy: it
//               only moves when a tile exceeds it by 2^8, so O is rarely rescaled and softmax_{j+1} does
//               not wait for PV_j -- the tensor pipe (QK_{j+2}, PV_j) and the MUFU pipe overlap.
//   All tcgen05 / TMA instructions are issued by one elected lane of a converged warp (elect_one).
//   q, k, v are column ranges of the packed QKV activation (row stride ld*), so the kernel reads
//   the GEMM output in place; `out` may alias q (the fused GPT-J block writes the context over q).
//   GQA/MQA: kv head = q head / (hq / hkv).
#include "common.cuh"

#include <cuda.h>
#include <string.h>

#ifndef LLMSS_PREFILL_P_TMEM
#define LLMSS_PREFILL_P_TMEM 1
#endif
#ifndef LLMSS_PREFILL_FMA_EXP_EVERY
#define LLMSS_PREFILL_FMA_EXP_EVERY 2   // n > 0: every n-th pair computes one of its two exps on the FMA pipe (1 in 4 at n = 2:
                                        // +2 % at head_dim 128 once P moved to tensor memory; no gain before that)
#endif

namespace {

constexpr int BM = 128;  // queries per CTA
constexpr int kThreads = 224;

// D = 256: one fat CTA per SM (64-key tiles, P double-buffered).  D = 128 / 64: 64-key tiles and ONE P buffer
// shrink the CTA to <= 112 KB of shared memory and 256 TMEM columns, so TWO CTAs share an SM: eight softmax
// warps (two per scheduler) keep the MUFU pipe and the FMA pipe busy at the same time -- with four warps and
// 128-key tiles the kernel was bound by 128 exponentials + ~8 ALU instructions per key per thread, issued from a
// single warp per scheduler (0.55-0.64x of the library flash kernel).  The second CTA's softmax also covers the
// wait for "PV_{j-1} has read P" that a single P buffer introduces.
template <int D>
struct PCfg {
  static constexpr int BN = 64;                              // keys per tile
  static constexpr int KV_STAGES = 2;
  // D <= 128: P never touches shared memory -- the softmax writes it (packed 16-bit pairs, tcgen05.st) over the S
  // columns it has just read, and the PV MMA takes its A operand from tensor memory.  S is double-buffered, so P is
  // too, for free; the QK of tile j+2 then has to wait for PV_j instead of for "S_j is in registers".
  static constexpr bool P_TMEM = (D < 256) && (LLMSS_PREFILL_P_TMEM != 0);
  static constexpr int P_BUFS = P_TMEM ? 0 : ((D >= 256) ? 2 : 1);
  static constexpr int MIN_CTAS = (D >= 256) ? 1 : 2;        // CTAs per SM the launch bounds ask for
  static constexpr int FMA_EXP_EVERY = (D >= 256) ? 0 : LLMSS_PREFILL_FMA_EXP_EVERY;  // every n-th PAIR computes one of its two exps on the FMA pipe (0 = none)
  static constexpr int Q_BYTES = BM * D * 2;
  static constexpr int K_BYTES = BN * D * 2;
  static constexpr int V_BYTES = BN * D * 2;
  static constexpr int P_BYTES = BM * BN * 2;
  // two CTAs per SM leave (228 KB - 2 x 1 KB reserved) / 2 = 115,712 bytes each: the tiles (114,688 at D = 128) plus the
  // barriers leave 768 bytes for aligning the tile base to the 1 KiB swizzle period (the kernel traps if the
  // dynamic segment starts at an address that needs more)
  static constexpr int ALIGN_SLACK = (D >= 256) ? 1024 : 768;
  static constexpr int SMEM_BYTES = Q_BYTES + KV_STAGES * (K_BYTES + V_BYTES) + P_BUFS * P_BYTES + ALIGN_SLACK + 192;
  static constexpr int TMEM_COLS = (2 * BN + D <= 256) ? 256 : 512;  // S0 at column 0, S1 at BN, O at 2*BN
};

struct PArgs {
  const int* cu_seqlens;
  // paged mode (queries over existing context): keys / values come from the paged cache through the block table
  const int* seq_lens;       // [num_seqs] cached length INCLUDING this step's tokens
  const int* block_tables;   // [num_seqs, max_blocks]
  int max_blocks, page;      // page = tokens per cache block (16 / 32 / 64)
  void* out;
  int64_t ldo;
  int hq, hkv;
  int q_col, k_col, v_col;  // first column of q / k / v inside the packed row (elements)
  float scale_log2;
  int dtype;
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

// 2^x on the FMA pipe (no MUFU): round-to-nearest split x = n + r, r in [-0.5, 0.5], cubic for 2^r (relative error
// 1.1e-4, far below the bf16 / fp16 rounding of P), exponent added into the float's bits.  A quarter of the
// exponentials of the softmax go this way: at head_dim 128 / 64 the MUFU pipe (16 ex2 per clock and SM) needs as
// many cycles per tile as the tensor pipe, so the two cannot both run at full rate unless some of the exps move.
__device__ __forceinline__ float ex2_fma(float x) {
  x = fmaxf(x, -125.0f);
  const float t = x + 12582912.0f;                 // 1.5 * 2^23: the integer part lands in the low mantissa bits
  const int n = __float_as_int(t) - 0x4B400000;
  const float r = x - (t - 12582912.0f);
  float p = fmaf(r, 0.05459282f, 0.24221784f);   // least-squares cubic on [-0.5, 0.5]: max relative error 1.1e-4
  p = fmaf(r, p, 0.6933686f);
  p = fmaf(r, p, 1.0f);
  return __int_as_float(__float_as_int(p) + (n << 23));
}

// D[tmem] (+)= A[tmem] * B[smem desc]: the A operand (M x 16, 16-bit, two K elements per 32-bit column, row m in
// lane m) is read from tensor memory.
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

// Packed fp32x2 arithmetic (one issue slot for two lanes): the softmax warps are bound by instruction issue
// (ncu: issue 50 %, tensor pipe 45 %, ~13 instructions per key and thread before this), so the scale-and-shift
// and the row-sum run as FFMA2 / FADD2.
__device__ __forceinline__ void ffma2(float& d0, float& d1, float a0, float a1, float b0, float b1, float c0, float c1) {
  asm("{\n\t.reg .b64 a, b, c, d;\n\tmov.b64 a, {%2, %3};\n\tmov.b64 b, {%4, %5};\n\tmov.b64 c, {%6, %7};\n\t"
      "fma.rn.f32x2 d, a, b, c;\n\tmov.b64 {%0, %1}, d;\n\t}"
      : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(b0), "f"(b1), "f"(c0), "f"(c1));
}
__device__ __forceinline__ void fadd2(float& d0, float& d1, float a0, float a1, float b0, float b1) {
  asm("{\n\t.reg .b64 a, b, d;\n\tmov.b64 a, {%2, %3};\n\tmov.b64 b, {%4, %5};\n\tadd.rn.f32x2 d, a, b;\n\tmov.b64 {%0, %1}, d;\n\t}"
      : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
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

// PAGED = false: fresh prefill, K / V are column ranges of the packed activation (tmap).
// PAGED = true : multi-token queries over prior context (chunked prefill, prefix reuse; the reference handles any
//                S with a past through the same dense path, gptj_modeling.py:128-169, 229-233): K / V tiles are
//                assembled page by page from the paged cache (tmap_k / tmap_v, one TMA box per page and 64-column
//                chunk, addressed through the block table), which already holds this step's keys (rope_kv_write
//                runs first).  Query i of a sequence sits at position ctx + i, ctx = seq_len - q_len.
template <typename T, int D, bool PAGED>
__global__ void __launch_bounds__(kThreads, PCfg<D>::MIN_CTAS)
attn_prefill_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_k,
                    const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ PArgs a) {
  using C = PCfg<D>;
  constexpr int BN = C::BN;
  constexpr int DSUB = D / 64;   // 64-column (128-byte) sub-tiles along head_dim
  constexpr int NSUB = BN / 64;  // 64-key sub-tiles of P along the key dimension
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  if (smem - smem_raw > C::ALIGN_SLACK) {
    if (threadIdx.x == 0) printf("[llmss_b200] attn_prefill: dynamic shared memory base needs %d bytes of alignment slack\n", (int)(smem - smem_raw));
    __trap();
  }
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + C::Q_BYTES;
  uint8_t* sV = sK + C::KV_STAGES * C::K_BYTES;
  uint8_t* sP = sV + C::KV_STAGES * C::V_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + C::P_BUFS * C::P_BYTES);
  uint64_t* bar_q = bars;
  uint64_t* bar_kfull = bars + 1;       // [KV_STAGES]
  uint64_t* bar_vfull = bars + 3;
  uint64_t* bar_kempty = bars + 5;
  uint64_t* bar_vempty = bars + 7;
  uint64_t* bar_s = bars + 9;           // [2] S_j ready in TMEM buffer j & 1 (tcgen05.commit)
  uint64_t* bar_p = bars + 11;          // [2] P_j written into buffer j & 1 (4 warp arrivals)
  uint64_t* bar_o = bars + 13;          // [2] PV_j retired (tcgen05.commit): P buffer j & 1 free, O stable
  uint64_t* bar_sfree = bars + 15;      // [2] softmax_j has copied S_j out of TMEM buffer j & 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 17);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int seq = blockIdx.z, head = blockIdx.y;
  pdl_launch_dependents();
  pdl_wait();
  const int seq_start = a.cu_seqlens[seq];
  const int len = a.cu_seqlens[seq + 1] - seq_start;
  // heaviest (longest causal range) query tiles are scheduled first
  const int q0 = (gridDim.x - 1 - blockIdx.x) * BM;  // first query (position within the sequence) of this CTA
  if (q0 >= len) return;           // uniform for the whole CTA
  const int kvh = head / (a.hq / a.hkv);
  const int ctx = PAGED ? a.seq_lens[seq] - len : 0;   // tokens already in the cache before this step's queries
  const int kv_end = min(ctx + len, ctx + q0 + BM);     // causal: keys [0, kv_end)
  const int n_tiles = (kv_end + BN - 1) / BN;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap);
    if (PAGED) {
      tma_prefetch_desc(&tmap_k);
      tma_prefetch_desc(&tmap_v);
    }
    mbar_init(bar_q, 1);
    for (int s = 0; s < C::KV_STAGES; ++s) {
      mbar_init(&bar_kfull[s], 1);
      mbar_init(&bar_vfull[s], 1);
      mbar_init(&bar_kempty[s], 1);
      mbar_init(&bar_vempty[s], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar_s[i], 1);
      mbar_init(&bar_p[i], 4);
      mbar_init(&bar_o[i], 1);
      mbar_init(&bar_sfree[i], 4);
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
  const uint32_t tS = tmem_base, tO = tmem_base + 2 * BN;  // S buffer b lives at tS + b * BN

  if (warp == 0) {
    // ===================== TMA producer (whole warp runs the loop, one elected lane issues) ===============
    {
      const uint64_t pol = l2_policy_evict_last();
      if (elect_one()) {
        mbar_arrive_expect_tx(bar_q, C::Q_BYTES);
        // the tensor-map box is [BN rows x 64 cols]; a 128-row Q sub-tile is BM / BN stacked boxes
        // (the 128B-swizzle pattern repeats every 8 rows, so stacked boxes form one contiguous tile)
#pragma unroll
        for (int c = 0; c < DSUB; ++c)
#pragma unroll
          for (int h = 0; h < BM / BN; ++h)
            tma_load_2d(sQ + c * (BM * 128) + h * (BN * 128), &tmap, bar_q, a.q_col + head * D + c * 64,
                        seq_start + q0 + h * BN, pol);
      }
      __syncwarp();
      int stage = 0;
      uint32_t phase = 0;
      for (int j = 0; j < n_tiles; ++j) {
        const int row = seq_start + j * BN;
        mbar_wait(&bar_kempty[stage], phase ^ 1, 50);
        if (elect_one()) {
          mbar_arrive_expect_tx(&bar_kfull[stage], C::K_BYTES);
          if constexpr (PAGED) {
            // BN / page boxes per 64-column chunk; 8-row multiples keep the 128B-swizzle pattern contiguous
            const int* bt = a.block_tables + (int64_t)seq * a.max_blocks;
            for (int pg = 0; pg < BN / a.page; ++pg) {
              const int bi = min(a.max_blocks - 1, (j * BN) / a.page + pg);
              const int crow = (bt[bi] * a.hkv + kvh) * a.page;
#pragma unroll
              for (int c = 0; c < DSUB; ++c)
                tma_load_2d(sK + stage * C::K_BYTES + c * (BN * 128) + pg * a.page * 128, &tmap_k, &bar_kfull[stage], c * 64, crow, pol);
            }
          } else {
#pragma unroll
            for (int c = 0; c < DSUB; ++c)
              tma_load_2d(sK + stage * C::K_BYTES + c * (BN * 128), &tmap, &bar_kfull[stage], a.k_col + kvh * D + c * 64, row, pol);
          }
        }
        __syncwarp();
        mbar_wait(&bar_vempty[stage], phase ^ 1, 51);
        if (elect_one()) {
          mbar_arrive_expect_tx(&bar_vfull[stage], C::V_BYTES);
          if constexpr (PAGED) {
            const int* bt = a.block_tables + (int64_t)seq * a.max_blocks;
            for (int pg = 0; pg < BN / a.page; ++pg) {
              const int bi = min(a.max_blocks - 1, (j * BN) / a.page + pg);
              const int crow = (bt[bi] * a.hkv + kvh) * a.page;
#pragma unroll
              for (int c = 0; c < DSUB; ++c)
                tma_load_2d(sV + stage * C::V_BYTES + c * (BN * 128) + pg * a.page * 128, &tmap_v, &bar_vfull[stage], c * 64, crow, pol);
            }
          } else {
#pragma unroll
            for (int c = 0; c < DSUB; ++c)
              tma_load_2d(sV + stage * C::V_BYTES + c * (BN * 128), &tmap, &bar_vfull[stage], a.v_col + kvh * D + c * 64, row, pol);
          }
        }
        __syncwarp();
        if (++stage == C::KV_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== QK issuer (whole warp runs the loop, one elected lane issues) =================
    const int fmt = a.dtype == LLMSS_BF16 ? 1 : 0;
    const uint32_t idesc_s = idesc_f16(fmt, BM, BN, 0);  // S = Q K^T : both operands K-major
    const uint32_t q_addr = smem_u32(sQ), k_base = smem_u32(sK);
    mbar_wait(bar_q, 0, 60);
    int stage = 0;
    uint32_t phase = 0;
    for (int j = 0; j < n_tiles; ++j) {
      if constexpr (C::P_TMEM) {
        if (j >= 2) mbar_wait(&bar_o[j & 1], ((j >> 1) & 1) ^ 1, 64);    // PV_{j-2} has read P_{j-2} out of this S buffer
      } else {
        if (j >= 2) mbar_wait(&bar_sfree[j & 1], ((j >> 1) & 1) ^ 1, 64);  // softmax_{j-2} has left this S buffer
      }
      mbar_wait(&bar_kfull[stage], phase, 61);
      tcgen05_fence_after();
      const uint32_t k_addr = k_base + stage * C::K_BYTES;
      if (elect_one()) {
#pragma unroll
        for (int c = 0; c < DSUB; ++c)
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16_ss(tS + (j & 1) * BN, umma_desc_k_sw128(q_addr + c * (BM * 128) + k * 32),
                        umma_desc_k_sw128(k_addr + c * (BN * 128) + k * 32), idesc_s, (c | k) ? 1u : 0u);
        umma_commit(&bar_s[j & 1]);
        umma_commit(&bar_kempty[stage]);
      }
      __syncwarp();
      if (++stage == C::KV_STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 6) {
    // ===================== PV issuer =====================
    const int fmt = a.dtype == LLMSS_BF16 ? 1 : 0;
    const uint32_t idesc_o = idesc_f16(fmt, BM, D, 1);   // O += P V  : B (= V) is MN-major
    const uint32_t p_addr = smem_u32(sP), v_base = smem_u32(sV);
    int stage = 0;
    uint32_t phase = 0;
    for (int j = 0; j < n_tiles; ++j) {
      mbar_wait(&bar_p[j & 1], (j >> 1) & 1, 62);
      mbar_wait(&bar_vfull[stage], phase, 63);
      tcgen05_fence_after();
      const uint32_t v_addr = v_base + stage * C::V_BYTES;
      const uint32_t pj = p_addr + (C::P_BUFS == 2 ? (j & 1) : 0) * C::P_BYTES;
      if (elect_one()) {
#pragma unroll
        for (int c = 0; c < NSUB; ++c)
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if constexpr (C::P_TMEM)   // 16 keys = 8 packed columns per step, P_j sits at the start of S buffer j & 1
              umma_f16_ts(tO, tS + (j & 1) * BN + (c * 4 + k) * 8, umma_desc_mn_sw128(v_addr + (c * 64 + k * 16) * 128, BN * 128),
                          idesc_o, (j > 0 || c > 0 || k > 0) ? 1u : 0u);
            else
              umma_f16_ss(tO, umma_desc_k_sw128(pj + c * (BM * 128) + k * 32),
                          umma_desc_mn_sw128(v_addr + (c * 64 + k * 16) * 128, BN * 128), idesc_o,
                          (j > 0 || c > 0 || k > 0) ? 1u : 0u);
          }
        umma_commit(&bar_o[j & 1]);
        umma_commit(&bar_vempty[stage]);
      }
      __syncwarp();
      if (++stage == C::KV_STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp < 6) {
    // ===================== softmax / correction / epilogue: one thread per query row =====================
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;         // query row within the tile == TMEM lane
    const int qrow = q0 + row;                   // query index within this step's tokens of the sequence
    const int qpos = ctx + qrow;                 // absolute position (== key index of the same token)
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    float m_run = -INFINITY, l_run = 0.f;
    for (int j = 0; j < n_tiles; ++j) {
      mbar_wait(&bar_s[j & 1], (j >> 1) & 1, 70);
      tcgen05_fence_after();
      const int kbase = j * BN;
      const bool diag = (kbase + BN - 1) > ctx + q0;   // some key of this tile may lie above the diagonal
      // the whole S row lives in registers (one TMEM read): raw scores, masked on diagonal tiles
      float sc[BN];
      {
        // both 32-column loads are in flight before the single wait (a TMEM load has a few hundred cycles of latency)
        uint32_t r[BN / 32][32];
#pragma unroll
        for (int c = 0; c < BN / 32; ++c) tmem_ld_32x32b_x32(tS + (j & 1) * BN + lane_off + c * 32, r[c]);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < BN / 32; ++c)
#pragma unroll
          for (int i = 0; i < 32; ++i) sc[c * 32 + i] = __uint_as_float(r[c][i]);
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_sfree[j & 1]);  // S_j is in registers: the buffer can take QK_{j+2}
      if (diag) {
#pragma unroll
        for (int i = 0; i < BN; ++i)
          if (kbase + i > qpos) sc[i] = -INFINITY;
      }
      // four independent max chains (two schedulers' worth of latency hiding comes from ILP, not from warps)
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < BN; i += 8) {
#pragma unroll
        for (int u = 0; u < 4; ++u) mx[u] = fmaxf(mx[u], fmaxf(sc[i + 2 * u], sc[i + 2 * u + 1]));
      }
      const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      // Lazy running max (scale > 0: max commutes with scaling): keep the old reference unless this tile
      // exceeds it by more than 2^8 -- p stays <= 256, far inside fp32 / bf16 / fp16 range.
      const float m_tile_s = m_tile * a.scale_log2;
      const bool grow = (j == 0) || (m_tile_s > m_run + 8.0f);
      const float m_new = grow ? fmaxf(m_run, m_tile_s) : m_run;
      const float alpha = (j == 0 || !grow) ? 1.0f : ex2_approx(m_run - m_new);
      m_run = m_new;
      // p = 2^(s*scale - m): one FFMA + one MUFU per element; P -> shared memory (K-major, chunk ^= row & 7)
      const float neg_m = (m_new == -INFINITY) ? 0.f : -m_new;
      if constexpr (C::P_TMEM) {
        // nothing to wait for: QK_j was only issued after PV_{j-2} had finished with this buffer
      } else if constexpr (C::P_BUFS == 2) {
        if (j >= 2) mbar_wait(&bar_o[j & 1], ((j >> 1) & 1) ^ 1, 71);  // PV_{j-2} has released this P buffer
      } else {
        if (j >= 1) mbar_wait(&bar_o[(j - 1) & 1], ((j - 1) >> 1) & 1, 71);  // single buffer: PV_{j-1} has read it
      }
      float la[4] = {0.f, 0.f, 0.f, 0.f}, lb[4] = {0.f, 0.f, 0.f, 0.f};  // four independent packed sum chains
      uint8_t* pbuf = sP + (C::P_BUFS == 2 ? (j & 1) : 0) * C::P_BYTES;
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float e0, e1;
          ffma2(e0, e1, sc[c * 32 + i], sc[c * 32 + i + 1], a.scale_log2, a.scale_log2, neg_m, neg_m);
          const float p0 = ex2_approx(e0);
          const float p1 = (C::FMA_EXP_EVERY > 0 && ((i >> 1) % (C::FMA_EXP_EVERY > 0 ? C::FMA_EXP_EVERY : 1)) == 0) ? ex2_fma(e1) : ex2_approx(e1);
          fadd2(la[(i >> 1) & 3], lb[(i >> 1) & 3], la[(i >> 1) & 3], lb[(i >> 1) & 3], p0, p1);
          pk[i >> 1] = pack2<T>(p0, p1);
        }
        if constexpr (C::P_TMEM) {
          // 32 keys -> 16 packed columns of this row's lane, over the S columns that are already in registers
          asm volatile(
              "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
              "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(tS + (j & 1) * BN + lane_off + c * 16),
              "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]), "r"(pk[4]), "r"(pk[5]), "r"(pk[6]), "r"(pk[7]), "r"(pk[8]),
              "r"(pk[9]), "r"(pk[10]), "r"(pk[11]), "r"(pk[12]), "r"(pk[13]), "r"(pk[14]), "r"(pk[15])
              : "memory");
        } else {
          uint8_t* sub = pbuf + (c >> 1) * (BM * 128) + row * 128;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int chunk = ((c & 1) * 4 + q) ^ (row & 7);
            *reinterpret_cast<uint4*>(sub + chunk * 16) = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
          }
        }
      }
      if constexpr (C::P_TMEM) tmem_st_wait();
      const float l_tile = ((la[0] + lb[0]) + (la[1] + lb[1])) + ((la[2] + lb[2]) + (la[3] + lb[3]));
      if (j > 0 && __any_sync(0xffffffffu, alpha != 1.0f)) {
        mbar_wait(&bar_o[(j - 1) & 1], ((j - 1) >> 1) & 1, 73);  // PV_{j-1} retired: O is stable
        tcgen05_fence_after();
#pragma unroll 1
        for (int c = 0; c < D / 32; ++c) {
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
      fence_proxy_async_smem();  // generic-proxy writes of P must be visible to the tensor-core (async) proxy
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_p[j & 1]);
    }
    // epilogue: O / l -> out
    mbar_wait(&bar_o[(n_tiles - 1) & 1], ((n_tiles - 1) >> 1) & 1, 72);
    tcgen05_fence_after();
    const float inv_l = l_run > 0.f ? 1.0f / l_run : 0.f;
    T* orow = reinterpret_cast<T*>(a.out) + (int64_t)(seq_start + qrow) * a.ldo + (int64_t)head * D;
#pragma unroll 1
    for (int c = 0; c < D / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(tO + lane_off + c * 32, r);
      tmem_ld_wait();
      if (qrow < len) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float f8[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) f8[i] = __uint_as_float(r[q * 8 + i]) * inv_l;
          *reinterpret_cast<uint4*>(orow + c * 32 + q * 8) = pack8<T>(f8);
        }
      }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

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

template <typename T, int D, bool PAGED>
int launch_prefill(const CUtensorMap& tm, const CUtensorMap& tk, const CUtensorMap& tv, const PArgs& a, int num_seqs,
                   int max_seqlen, cudaStream_t stream) {
  using C = PCfg<D>;
  static PerDeviceSmem configured;
  auto kern = attn_prefill_kernel<T, D, PAGED>;
  LLMSS_CHECK_CUDA(configured.ensure(kern, C::SMEM_BYTES));
  dim3 grid((max_seqlen + BM - 1) / BM, a.hq, num_seqs);
  LLMSS_CHECK_CUDA(launch_pdl(kern, grid, dim3(kThreads), (size_t)C::SMEM_BYTES, stream, tm, tk, tv, a));
  return 0;
}

int encode_2d(CUtensorMap* tm, const void* base, int dtype, int64_t cols, int64_t rows, int64_t ld, int box_rows) {
  PFN_encodeTiled enc = get_encode();
  LLMSS_REQUIRE(enc != nullptr, "attn_prefill: cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, dtype == LLMSS_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                   const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  LLMSS_REQUIRE(r == CUDA_SUCCESS, "attn_prefill: cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

}  // namespace

// q / k / v: pointers to the first q / k / v element of row 0 inside ONE packed activation
// (same base allocation, same row stride ld); out may alias q.  cu_seqlens: [num_seqs + 1] int32.
LLMSS_API int llmss_attn_prefill(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                 void* out, int64_t ldo, const int* cu_seqlens, int num_seqs, int max_seqlen,
                                 int total_tokens, int hq, int hkv, int d, float scale, int dtype,
                                 cudaStream_t stream) {
  LLMSS_REQUIRE(dtype == LLMSS_F16 || dtype == LLMSS_BF16, "attn_prefill: dtype must be f16/bf16");
  LLMSS_REQUIRE(d == 64 || d == 128 || d == 256, "attn_prefill: head_dim=%d unsupported (64/128/256)", d);
  LLMSS_REQUIRE(hq > 0 && hkv > 0, "attn_prefill: head counts must be positive (hq=%d hkv=%d)", hq, hkv);
  LLMSS_REQUIRE(hq % hkv == 0, "attn_prefill: hq=%d not a multiple of hkv=%d", hq, hkv);
  LLMSS_REQUIRE(ldq == ldk && ldk == ldv, "attn_prefill: q, k, v must share one row stride (packed QKV activation)");
  LLMSS_REQUIRE(ldo % 8 == 0 && ((uintptr_t)out % 16) == 0, "attn_prefill: output must be 16-byte aligned");
  if (num_seqs == 0 || total_tokens == 0) return 0;
  // one tensor map over the packed rows, based at the lowest of the three pointers
  const char* base = (const char*)q;
  if ((const char*)k < base) base = (const char*)k;
  if ((const char*)v < base) base = (const char*)v;
  const int64_t q_col = ((const char*)q - base) / 2, k_col = ((const char*)k - base) / 2, v_col = ((const char*)v - base) / 2;
  int64_t width = q_col + (int64_t)hq * d;
  if (k_col + (int64_t)hkv * d > width) width = k_col + (int64_t)hkv * d;
  if (v_col + (int64_t)hkv * d > width) width = v_col + (int64_t)hkv * d;
  LLMSS_REQUIRE(width <= ldq, "attn_prefill: q/k/v column ranges exceed the row stride");
  LLMSS_REQUIRE(((uintptr_t)base % 16) == 0 && (ldq * 2) % 16 == 0 && q_col % 8 == 0 && k_col % 8 == 0 && v_col % 8 == 0,
                "attn_prefill: packed activation must be 16-byte aligned");
  PFN_encodeTiled enc = get_encode();
  LLMSS_REQUIRE(enc != nullptr, "attn_prefill: cuTensorMapEncodeTiled entry point not available");
  CUtensorMap tm;
  cuuint64_t dims[2] = {(cuuint64_t)width, (cuuint64_t)total_tokens};
  cuuint64_t strides[1] = {(cuuint64_t)ldq * 2};
  const int bn = 64;
  // one box shape serves Q (128 rows) and K/V (BN rows) only if they agree; use two loads of BN rows for Q when BN < 128
  cuuint32_t box[2] = {64, (cuuint32_t)bn};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(&tm, dtype == LLMSS_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                   (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  LLMSS_REQUIRE(r == CUDA_SUCCESS, "attn_prefill: cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  PArgs a;
  memset(&a, 0, sizeof(a));
  a.cu_seqlens = cu_seqlens;
  a.out = out;
  a.ldo = ldo;
  a.hq = hq;
  a.hkv = hkv;
  a.q_col = (int)q_col;
  a.k_col = (int)k_col;
  a.v_col = (int)v_col;
  a.scale_log2 = scale * 1.4426950408889634f;
  a.dtype = dtype;
#define LLMSS_PF(T_, D_) return launch_prefill<T_, D_, false>(tm, tm, tm, a, num_seqs, max_seqlen, stream)
  if (dtype == LLMSS_F16) {
    if (d == 64) LLMSS_PF(__half, 64);
    if (d == 128) LLMSS_PF(__half, 128);
    LLMSS_PF(__half, 256);
  } else {
    if (d == 64) LLMSS_PF(__nv_bfloat16, 64);
    if (d == 128) LLMSS_PF(__nv_bfloat16, 128);
    LLMSS_PF(__nv_bfloat16, 256);
  }
#undef LLMSS_PF
}

// Multi-token queries over prior context (chunked prefill / prefix reuse / speculative verification).
// q: [total_q, >= hq * d] query rows of the packed activation (row stride ldq); K / V: the paged caches
// [num_blocks, hkv, block_size, d] which already contain this step's keys and values; seq_lens[i] = cached length of
// sequence i INCLUDING its cu_q[i+1] - cu_q[i] new tokens; out may alias q.  block_size in {16, 32, 64}.
LLMSS_API int llmss_attn_prefill_paged(const void* q, int64_t ldq, const void* k_cache, const void* v_cache,
                                       const int* block_tables, int max_blocks, const int* seq_lens, void* out, int64_t ldo,
                                       const int* cu_q, int num_seqs, int max_q, int total_q, int hq, int hkv, int d,
                                       int block_size, int num_blocks, float scale, int dtype, cudaStream_t stream) {
  LLMSS_REQUIRE(dtype == LLMSS_F16 || dtype == LLMSS_BF16, "attn_prefill_paged: dtype must be f16/bf16");
  LLMSS_REQUIRE(d == 64 || d == 128 || d == 256, "attn_prefill_paged: head_dim=%d unsupported (64/128/256)", d);
  LLMSS_REQUIRE(hq > 0 && hkv > 0 && hq % hkv == 0, "attn_prefill_paged: bad head grouping hq=%d hkv=%d", hq, hkv);
  LLMSS_REQUIRE(block_size == 16 || block_size == 32 || block_size == 64, "attn_prefill_paged: block_size=%d unsupported (16/32/64)", block_size);
  LLMSS_REQUIRE(ldo % 8 == 0 && ((uintptr_t)out % 16) == 0 && ((uintptr_t)q % 16) == 0 && (ldq * 2) % 16 == 0,
                "attn_prefill_paged: q / out must be 16-byte aligned");
  LLMSS_REQUIRE((int64_t)hq * d <= ldq, "attn_prefill_paged: q columns exceed the row stride");
  if (num_seqs == 0 || total_q == 0) return 0;
  const int bn = 64;
  CUtensorMap tm, tk, tv;
  if (int e = encode_2d(&tm, q, dtype, (int64_t)hq * d, total_q, ldq, bn)) return e;
  const int64_t crow = (int64_t)num_blocks * hkv * block_size;
  if (int e = encode_2d(&tk, k_cache, dtype, d, crow, d, block_size)) return e;
  if (int e = encode_2d(&tv, v_cache, dtype, d, crow, d, block_size)) return e;
  PArgs a;
  memset(&a, 0, sizeof(a));
  a.cu_seqlens = cu_q;
  a.seq_lens = seq_lens;
  a.block_tables = block_tables;
  a.max_blocks = max_blocks;
  a.page = block_size;
  a.out = out;
  a.ldo = ldo;
  a.hq = hq;
  a.hkv = hkv;
  a.q_col = 0;
  a.scale_log2 = scale * 1.4426950408889634f;
  a.dtype = dtype;
#define LLMSS_PP(T_, D_) return launch_prefill<T_, D_, true>(tm, tk, tv, a, num_seqs, max_q, stream)
  if (dtype == LLMSS_F16) {
    if (d == 64) LLMSS_PP(__half, 64);
    if (d == 128) LLMSS_PP(__half, 128);
    LLMSS_PP(__half, 256);
  } else {
    if (d == 64) LLMSS_PP(__nv_bfloat16, 64);
    if (d == 128) LLMSS_PP(__nv_bfloat16, 128);
    LLMSS_PP(__nv_bfloat16, 256);
  }
#undef LLMSS_PP
}
