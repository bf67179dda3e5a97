/* llmss_b200 -- C ABI of the hand-written sm_100a kernel library (libllmss_b200.so).
 *
 * Every entry point returns 0 on success; on failure it returns a non-zero code and
 * llmss_last_error() describes the problem.  Pointers are device pointers unless noted.
 * All launches go to the given stream; nothing here synchronises the device except the
 * symmetric-memory set-up calls.  dtype: 0 = fp16, 1 = bf16, 2 = fp32.
 *
 * Each declaration cites the reference (jongwon-jay-lee/llmss) code it replaces; see
 * SURVEY.md sections 2.5 / 2.6 for the N-number / G-number labels.
 */
#ifndef LLMSS_B200_H_
#define LLMSS_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* llmss_stream_t;
typedef struct SymmWs llmss_symm_t;

enum { LLMSS_DT_F16 = 0, LLMSS_DT_BF16 = 1, LLMSS_DT_F32 = 2 };

const char* llmss_last_error(void);
int llmss_version(void);

/* ---- symmetric memory (replaces c10d/NCCL bring-up for the hot path; dist.py:44-73) ---- */
int llmss_symm_alloc(size_t buf_bytes, size_t lam_bytes, void** ptr); /* flags + Lamport area + bulk area */
int llmss_symm_free(void* ptr);
int llmss_ipc_get_handle(void* ptr, unsigned char* out64);
int llmss_ipc_open_handle(const unsigned char* in64, void** ptr);
int llmss_ipc_close(void* ptr);
size_t llmss_symm_region_bytes(size_t buf_bytes, size_t lam_bytes);
int llmss_symm_create(int rank, int nranks, void* const* peer_bases, size_t buf_bytes, size_t lam_bytes,
                      llmss_symm_t** out);
int llmss_symm_destroy(llmss_symm_t* ws);
int llmss_symm_epoch(llmss_symm_t* ws, uint32_t* out_host);

/* ---- symmetric heap on the CUDA VMM API + NVLS multicast (csrc/vmm.cu) ----
 * Each rank: llmss_vmm_padded_size (identical on all ranks) -> llmss_vmm_alloc -> llmss_symm_init_region on
 * llmss_vmm_ptr -> llmss_vmm_export_fd once per peer, descriptor sent with llmss_fdx_send / received with
 * llmss_fdx_recv (abstract Unix datagram sockets, SCM_RIGHTS) -> llmss_vmm_import.  Optional NVLS: rank 0
 * llmss_mc_create + descriptor to every peer (llmss_mc_import), all llmss_mc_add_device, barrier, all
 * llmss_mc_bind_and_map, barrier; llmss_vmm_mc_ptr is then the multicast mapping (a store lands on every GPU,
 * multimem.ld_reduce sums all copies in the switch).  llmss_symm_create2 takes the peer mappings and that pointer
 * (null: P2P only).  The reference has none of this: its buffers belong to NCCL (dist.py:44-73). */
typedef struct llmss_vmm_region llmss_vmm_region;
int llmss_vmm_caps(int* caps); /* bit 0: VMM + POSIX-fd handles, bit 1: multicast objects */
int llmss_vmm_padded_size(size_t bytes, int nranks, int want_multicast, size_t* padded);
int llmss_vmm_alloc(size_t padded, llmss_vmm_region** out);
void* llmss_vmm_ptr(llmss_vmm_region* r);
void* llmss_vmm_mc_ptr(llmss_vmm_region* r);
size_t llmss_vmm_bytes(llmss_vmm_region* r);
int llmss_vmm_export_fd(llmss_vmm_region* r, int* fd);
int llmss_vmm_import(int fd, size_t padded, void** ptr); /* closes fd */
int llmss_vmm_unmap(void* ptr, size_t padded);
int llmss_vmm_free(llmss_vmm_region* r);
int llmss_mc_create(int nranks, size_t padded, unsigned long long* mc, int* fd);
int llmss_mc_import(int fd, unsigned long long* mc);
int llmss_mc_add_device(unsigned long long mc);
int llmss_mc_bind_and_map(unsigned long long mc, llmss_vmm_region* r);
int llmss_mc_release(unsigned long long mc);
int llmss_fdx_open(const char* name, int* sock);
int llmss_fdx_send(int sock, const char* peer_name, int tag, int fd);
int llmss_fdx_recv(int sock, int timeout_ms, int* tag, int* fd);
int llmss_fdx_close(int sock);
int llmss_fd_close(int fd);
int llmss_symm_init_region(void* ptr, size_t buf_bytes, size_t lam_bytes); /* zero flags, arm the Lamport area */
int llmss_symm_create2(int rank, int nranks, void* const* peer_bases, void* mc_base, size_t buf_bytes, size_t lam_bytes,
                       llmss_symm_t** out);
int llmss_symm_has_multicast(llmss_symm_t* ws);
/* watchdog self test: waits on `flag` (a device uint32 nobody raises to `want`); the kernel traps after the 4 s spin
 * limit and the context is lost -- call it from a throw-away process only */
int llmss_debug_wait_flag(const uint32_t* flag, uint32_t want, llmss_stream_t stream);

/* ---- K4 collectives over NVLink peer memory (N1/N2: layers.py:178,213; N3: layers.py:133) ---- */
/* algo: 0 auto | 1 one-shot Lamport push | 2 two-shot with flags | 3 NVLS (stage locally, multimem.ld_reduce of the owned
 * 1/t, multimem.st broadcast) | 4 two-shot on Lamport vectors (reduce-scatter by peer stores, multicast broadcast) */
int llmss_allreduce(llmss_symm_t* ws, const void* in, void* out, int64_t n, int dtype, int algo,
                    llmss_stream_t stream);
int llmss_allgather(llmss_symm_t* ws, const void* in, void* out, int64_t bytes, llmss_stream_t stream);
/* gather along the last dimension: in [rows][row_bytes] -> out [rows][nranks * row_bytes] (N3: the vocab-parallel
 * logits all-gather + cat of layers.py:131-135 in one pass) */
int llmss_allgather_cols(llmss_symm_t* ws, const void* in, void* out, int64_t rows, int64_t row_bytes,
                         llmss_stream_t stream);

/* ---- K5/K1/K2 tcgen05 GEMM (+ fused all-reduce epilogue) (G3,G4,G9,G10,G12,G14; layers.py:62,175-179) ---- */
typedef struct llmss_gemm_plan {
  int bn, tiles, kb_total, grid;
  int split; /* 0 = stream-K persistent kernel, 1..8 = thread-block-cluster split-K factor (1: one CTA per tile) */
  int64_t sk_ws_floats;
  int64_t sk_counters;
  int64_t ar_bytes;
} llmss_gemm_plan;
/* CTA-pair kernel (tcgen05 cta_group::2) for token tiles of 129..256 rows: -1 = environment (LLMSS_GEMM_PAIR =
 * off | auto | 1 | 2 | 4), 0 = off, 1 / 2 / 4 = pairs per cluster (split-K depth), 8 = automatic */
int llmss_gemm_set_pair(int mode);
int llmss_gemm_set_debug(unsigned long long* per_cta_stamps); /* profiling aid: 8 globaltimer stamps per CTA */
int llmss_gemm_plan_query(int M, int N, int K, int nranks, int elem_bytes, int max_ctas, llmss_gemm_plan* p);
int llmss_gemm_plan_query2(int M, int N, int K, int nranks, int elem_bytes, int max_ctas, int dtype, llmss_gemm_plan* p);
int llmss_gemm(const void* x, int64_t ldx, const void* w, int64_t ldw, const void* bias, const void* residual,
               int64_t ldr, void* out, int64_t ldo, int M, int N, int K, int dtype, int out_f32, int act,
               int act_start, int gated, float* sk_ws, uint32_t* sk_counters, llmss_symm_t* ws, int max_ctas,
               llmss_stream_t stream);

/* collective: 0 none | 1 GEMM -> all-reduce (exchange chosen by size and rank count) | 2 GEMM -> reduce-scatter over
 * token rows (K2) | 3 all-gather -> GEMM (K3: x is this rank's row shard; shards are staged into every rank's own
 * buffer -- one multicast store each when the heap is multicast-bound -- and TMA-read from local memory) |
 * 4 / 5 / 6 pin the all-reduce exchange: in-switch multimem.ld_reduce / one-shot push / two-shot Lamport. */
int llmss_gemm_ex(const void* x, int64_t ldx, const void* w, int64_t ldw, const void* bias, const void* residual,
                  int64_t ldr, void* out, int64_t ldo, int M, int N, int K, int dtype, int out_f32, int act,
                  int act_start, int gated, float* sk_ws, uint32_t* sk_counters, llmss_symm_t* ws, int max_ctas,
                  int collective, void* ag_maps_dev, llmss_stream_t stream);

/* ---- K6 fused residual add + LayerNorm / RMSNorm (G2, G13, G18; layers.py:12-36,217-256) ---- */
int llmss_norm_residual(const void* x, const void* x2, const void* residual_in, void* residual_out,
                        const void* weight, const void* bias, void* out, int M, int H, int64_t ldx, int64_t ldx2,
                        int64_t ldr, int64_t ldo, float eps, int kind, int dtype, llmss_stream_t stream);

/* ---- K7 RoPE + paged KV-cache append (G5, G6; gptj_modeling.py:26-47,206-233) ---- */
int llmss_rope_kv_write(void* qkv, const int* positions, const int* slot_mapping, void* k_cache, void* v_cache,
                        int T, int64_t ld, int hq, int hkv, int d, int rotary_dim, int style, float base, int q_off,
                        int k_off, int v_off, int block_size, int dtype, llmss_stream_t stream);

/* ---- K9 paged split-KV decode attention (G7, G8, G17; gptj_modeling.py:128-169) ---- */
/* Tensor-core (tcgen05) paged decode attention for grouped / multi-query models (hq / hkv >= 2): the
 * q heads that share a KV head are the rows of one MMA, so each KV page is read once per group
 * (reference: baddbmm/softmax/bmm on a replicated KV tensor, gpt_bigcode_modeling.py:170-246).
 * Same contract as llmss_attn_decode plus num_blocks (pages in the cache); block_size must be 16.
 * Scratch size: llmss_attn_decode_scratch_floats. */
int llmss_attn_decode_tc(const void* q, int64_t ldq, const void* k_cache, const void* v_cache,
                         const int* block_tables, int max_blocks, const int* seq_lens, void* out,
                         int64_t ldo, float* scratch, int B, int hq, int hkv, int d, int block_size,
                         int num_blocks, float scale, int num_splits, int dtype, llmss_stream_t stream);
int llmss_attn_decode_tc_pick_splits(int B, int hkv, int d, int max_kv);
/* num_splits = 0 selects the balanced mode: all key tiles of the batch are cut into equal contiguous ranges,
 * one per SM (no wave quantisation of items / SMs); its scratch is sized by the function below. */
int64_t llmss_attn_decode_tc_balanced_scratch_floats(int B, int hq, int d, int max_blocks, int block_size);
/* Merge split-KV partials (part_o [B,hq,splits,d], part_ml [B,hq,splits,2]) into out. */
int llmss_attn_decode_combine(const float* part_o, const float* part_ml, void* out, int64_t ldo, int B,
                              int hq, int d, int num_splits, int dtype, llmss_stream_t stream);
int64_t llmss_attn_decode_scratch_floats(int B, int hq, int d, int num_splits);
int llmss_attn_decode_pick_splits(int B, int hq, int hkv, int d, int max_kv, int dtype); /* wave-filling split factor */
int llmss_attn_decode(const void* q, int64_t ldq, const void* k_cache, const void* v_cache,
                      const int* block_tables, int max_blocks, const int* seq_lens, void* out, int64_t ldo,
                      float* scratch, int B, int hq, int hkv, int d, int block_size, float scale, int num_splits,
                      int dtype, llmss_stream_t stream);

/* ---- K8 causal varlen prefill attention on tcgen05 (G7, G8, G17; gpt_bigcode_modeling.py:170-246) ---- */
int llmss_attn_prefill(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv, void* out,
                       int64_t ldo, const int* cu_seqlens, int num_seqs, int max_seqlen, int total_tokens, int hq,
                       int hkv, int d, float scale, int dtype, llmss_stream_t stream);

/* Multi-token queries over prior context (chunked prefill, prefix reuse; gptj_modeling.py:128-169 with a past,
 * 229-233): same tcgen05 kernel, K / V tiles gathered page by page from the paged cache (which already holds this
 * step's keys); seq_lens[i] includes the cu_q[i+1] - cu_q[i] new tokens; block_size 16 / 32 / 64; out may alias q. */
int llmss_attn_prefill_paged(const void* q, int64_t ldq, const void* k_cache, const void* v_cache,
                             const int* block_tables, int max_blocks, const int* seq_lens, void* out, int64_t ldo,
                             const int* cu_q, int num_seqs, int max_q, int total_q, int hq, int hkv, int d,
                             int block_size, int num_blocks, float scale, int dtype, llmss_stream_t stream);

/* No counterpart in the reference (it has no native code and allocates freely through torch).
 * Same operation as llmss_attn_decode for hosts that cannot allocate: servers whose HBM is fully
 * committed to weights + KV pages, decode loops that cannot afford the device-wide sync of
 * cudaMalloc/cudaFree, and plain C hosts with no allocator to lend us scratch.  The split-KV
 * partials live in an 8 MiB array built into the library (held by every GPU that loads it).
 * Refuses B*hq*num_splits*(d+2) floats > llmss_attn_decode_static_max_floats().  The first call
 * on a device resolves the array's address: make it at start-up, before any stream capture.
 * Calls share the array; kernels of two calls that overlap on the GPU race. */
int64_t llmss_attn_decode_static_max_floats(void);
int llmss_attn_decode_static(const void* q, int64_t ldq, const void* k_cache, const void* v_cache,
                             const int* block_tables, int max_blocks, const int* seq_lens, void* out, int64_t ldo,
                             int B, int hq, int hkv, int d, int block_size, float scale, int num_splits, int dtype,
                             llmss_stream_t stream);

/* ---- K10 embedding gather (G1; layers.py:182-214) ---- */
int llmss_embed(const int* tokens, const int* positions, const void* wte, const void* wpe, void* out, int T, int H,
                int vocab, int max_pos, int pos_offset, int64_t ldo, int dtype, llmss_stream_t stream);

/* ---- K11 (greedy case) vocab-parallel argmax: per-rank candidate, gathered with llmss_allgather_cols (16 bytes
 * per row and rank), merged on every rank -- no [B, V] logits all-gather (generate.py:111, layers.py:131-135) ---- */
int llmss_argmax_local(const float* logits, int64_t ld, int B, int n_valid, int idx_offset, void* cand,
                       llmss_stream_t stream);
int llmss_argmax_merge(const void* cand_all, int nranks, int B, int* out_tokens, long long* out_tokens64,
                       llmss_stream_t stream);

/* ---- K11 (general case) vocab-parallel top-k / top-p / temperature sampling (generate.py:109-126, layers.py:131-135):
 * step 1 on every rank: one record of llmss_vp_record_floats(ks) floats per row -- [max, sum-exp, count, sum-exp at the row's temperature,
 * vals[ks], idx[ks]] -- from its fp32 shard; gather the records with llmss_allgather_cols; step 2 on every rank: merge + draw
 * (identical tokens everywhere, rng_state advanced).  Exact for 1 <= top_k <= ks; nranks * ks <= 1024. ---- */
int llmss_vp_record_floats(int ks);
int llmss_vp_summary(const float* logits, int64_t ld, int B, int n_valid, int idx_offset, const int* greedy,
                     const int* top_k, const float* temperature, float* recs, int ks, llmss_stream_t stream);
/* rows drawn from the whole distribution (top_k = 0, top_p = 1 -- what the reference's loop effectively samples, quirk Q1):
 * after the record gather every rank runs llmss_vp_multinomial_pick (the rank whose CDF interval holds the draw names the
 * token, all others answer -1; does not advance the RNG), `cand` ([B][4] int32) is gathered with llmss_allgather_cols,
 * llmss_vp_merge_sample runs for all rows and llmss_vp_multinomial_merge overwrites the tokens of those rows. */
int llmss_vp_multinomial_pick(const float* logits, int64_t ld, int B, int n_valid, int idx_offset, const float* recs_all,
                              int nranks, int rank, int ks, const int* greedy, const float* temperature, const int* top_k,
                              const float* top_p, const uint32_t* rng_state, int* cand, llmss_stream_t stream);
int llmss_vp_multinomial_merge(const int* cand_all, int nranks, int B, int* out_tokens, long long* out_tokens64,
                               llmss_stream_t stream);
int llmss_vp_merge_sample(const float* recs_all, int nranks, int ks, int B, int vocab, const int* greedy,
                          const float* temperature, const int* top_k, const float* top_p, uint32_t* rng_state,
                          int* out_tokens, long long* out_tokens64, llmss_stream_t stream);

/* ---- K11 fused sampler (G15, G16, N6; generate.py:109-144) ----
 * rng_state: [B][2] = (request seed, tokens sampled so far); the draw depends on nothing else (not on the row).
 * top_k = 0 with top_p < 1: the nucleus is searched among the 1024 most probable tokens (a wider nucleus is cut there);
 * top_k = 0 with top_p = 1 samples the whole distribution. */
int llmss_sample(const float* logits, int64_t ld, int B, int vocab, const int* greedy, const float* temperature,
                 const int* top_k, const float* top_p, uint32_t* rng_state, int* out_tokens,
                 long long* out_tokens64, llmss_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* LLMSS_B200_H_ */
