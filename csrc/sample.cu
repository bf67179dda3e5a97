// K11 (sampling half): fused greedy / top-p / top-k / temperature sampler over fp32 logits.
//
// Replaces the reference's rank-0-only sampling (SURVEY.md G15/G16 + N6: argmax or
// softmax + HF warpers + multinomial on rank 0, then a broadcast of the token,
// generate.py:109-144) with one launch that every rank runs on identical gathered logits
// and an identical counter-based RNG, so all ranks agree without a collective.  It
// implements the *intended* warper semantics (quirk Q1: the reference never applies
// them): top-p on the raw distribution, then top-k, then temperature, then one draw.
//
// One CTA per row.  Passes over the row (L2-resident): max/argmax, sum-exp, a 4-pass
// 8-bit radix select for the k-th largest logit, candidate gather (<= 1024), in-block
// bitonic sort (value desc, index asc), prefix-mass scan for top-p, inverse-CDF draw.
// top_k == 0 is treated as k = 1024 when top_p < 1 (top-p is evaluated over the 1024 most
// probable tokens); top_k == 0 && top_p >= 1 is a plain multinomial over the whole row.
#include "common.cuh"

namespace {

constexpr int kThreads = 1024;
constexpr int kMaxCand = 1024;

__device__ __forceinline__ uint32_t fkey(float x) {  // order-preserving float -> uint
  uint32_t u = __float_as_uint(x);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ uint32_t fmix32(uint32_t x) {
  x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
  return x;
}
// Keyed on the request's (seed, step) only -- not on the batch row -- so a request draws the same numbers
// whichever slot of the running batch it occupies (the engine derives the seed from the request id when the
// caller gives none).
__device__ __forceinline__ float rng_uniform(uint32_t seed, uint32_t step) {
  uint32_t x = 0x9E3779B1u + step * 0x85EBCA77u + seed * 0xC2B2AE3Du;
  x = fmix32(fmix32(x));
  return (float)(x >> 8) * (1.0f / 16777216.0f);
}

struct MaxIdx { float v; int i; };
__device__ __forceinline__ MaxIdx better(MaxIdx a, MaxIdx b) {
  return (b.v > a.v || (b.v == a.v && b.i < a.i)) ? b : a;
}

__device__ __forceinline__ float block_sum_f(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = (l < kThreads / 32) ? red[l] : 0.f;
  t = warp_sum(t);
  return t;
}

// ---- building blocks shared by the single-GPU sampler and the vocab-parallel one (all threads of a 1024-thread CTA
// call them; the shared arrays are the caller's) ----

// Key (fkey order) of the k-th largest of x[0..n): 4-pass 8-bit radix select.
__device__ __forceinline__ uint32_t block_radix_kth(const float* __restrict__ x, int n, int k, int* hist,
                                                    uint32_t* s_prefix, int* s_krem) {
  const int tid = threadIdx.x;
  if (tid == 0) { *s_prefix = 0; *s_krem = k; }
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 24 - 8 * pass;
    if (tid < 256) hist[tid] = 0;
    __syncthreads();
    const uint32_t prefix = *s_prefix;
    const uint32_t pmask = (pass == 0) ? 0u : (0xffffffffu << (shift + 8));
    for (int i = tid; i < n; i += kThreads) {
      const uint32_t key = fkey(x[i]);
      if ((key & pmask) == prefix) atomicAdd(&hist[(key >> shift) & 0xff], 1);
    }
    __syncthreads();
    if (tid == 0) {
      int rem = *s_krem, bin = 255;
      for (; bin > 0; --bin) {
        if (hist[bin] >= rem) break;
        rem -= hist[bin];
      }
      *s_krem = rem;
      *s_prefix = prefix | ((uint32_t)bin << shift);
    }
    __syncthreads();
  }
  return *s_prefix;
}

// In-block bitonic sort of (cval, cidx)[0..count): value descending, index ascending.
__device__ __forceinline__ void block_sort_candidates(float* cval, int* cidx, int count) {
  const int tid = threadIdx.x;
  int nsort = 1;
  while (nsort < count) nsort <<= 1;
  for (int i = tid; i < nsort; i += kThreads)
    if (i >= count) { cval[i] = -INFINITY; cidx[i] = 0x7fffffff; }
  __syncthreads();
  for (int size = 2; size <= nsort; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      const int i = tid;
      if (i < nsort) {
        const int j = i ^ stride;
        if (j > i) {
          const bool desc_block = ((i & size) == 0);  // this block sorted "better first"
          const float vi = cval[i], vj = cval[j];
          const int ii = cidx[i], ij = cidx[j];
          const bool i_before_j = (vi > vj) || (vi == vj && ii < ij);
          if (i_before_j != desc_block) { cval[i] = vj; cval[j] = vi; cidx[i] = ij; cidx[j] = ii; }
        }
      }
      __syncthreads();
    }
  }
}

// Every x[i] whose key is >= thr_key (at most kMaxCand of them), sorted; indices are offset by idx_offset.
__device__ __forceinline__ int block_gather_sorted(const float* __restrict__ x, int n, uint32_t thr_key, int idx_offset,
                                                   float* cval, int* cidx, int* s_count) {
  const int tid = threadIdx.x;
  if (tid == 0) *s_count = 0;
  __syncthreads();
  for (int i = tid; i < n; i += kThreads) {
    const float v = x[i];
    if (fkey(v) >= thr_key) {
      const int slot = atomicAdd(s_count, 1);
      if (slot < kMaxCand) { cval[slot] = v; cidx[slot] = i + idx_offset; }
    }
  }
  __syncthreads();
  const int count = min(*s_count, kMaxCand);
  block_sort_candidates(cval, cidx, count);
  return count;
}

// Warp 0: nucleus (top-p on the raw distribution, Z = its normaliser) over the sorted candidates, temperature
// weights, one inverse-CDF draw.  Returns the position of the chosen candidate.
__device__ __forceinline__ int warp_draw_from_sorted(const float* cval, int count, float xmax, float Z, float tp,
                                                     float inv_t, float u01) {
  const int tid = threadIdx.x;
  // n_p = number of sorted positions whose preceding mass is < top_p
  float run = 0.f;  // mass before the current 32-chunk
  int n_p = 0;
  for (int base = 0; base < count; base += 32) {
    const int i = base + tid;
    const float p = (i < count) ? __expf(cval[i] - xmax) / Z : 0.f;
    float incl = p;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      float n = __shfl_up_sync(0xffffffffu, incl, o);
      if (tid >= o) incl += n;
    }
    const bool ok = (i < count) && ((run + incl - p) < tp);
    const unsigned mask = __ballot_sync(0xffffffffu, ok);
    n_p += __popc(mask);
    run += __shfl_sync(0xffffffffu, incl, 31);
    if (mask != 0xffffffffu) break;
  }
  n_p = max(n_p, 1);
  const float thr_p = cval[n_p - 1];
  // keep = prefix with value >= thr_p (ties stay), already limited to >= k-th largest
  int n_keep = 0;
  for (int base = 0; base < count; base += 32) {
    const int i = base + tid;
    const unsigned mask = __ballot_sync(0xffffffffu, (i < count) && (cval[i] >= thr_p));
    n_keep += __popc(mask);
    if (mask != 0xffffffffu) break;
  }
  // total weight at temperature
  float tot = 0.f;
  for (int i = tid; i < n_keep; i += 32) tot += __expf((cval[i] - xmax) * inv_t);
  tot = warp_sum(tot);
  const float target = u01 * tot;
  float before = 0.f;
  int pick = n_keep - 1;
  bool found = false;
  for (int base = 0; base < n_keep && !found; base += 32) {
    const int i = base + tid;
    const float w = (i < n_keep) ? __expf((cval[i] - xmax) * inv_t) : 0.f;
    float incl = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      float n = __shfl_up_sync(0xffffffffu, incl, o);
      if (tid >= o) incl += n;
    }
    const bool hit = (i < n_keep) && (target < before + incl);
    const unsigned mask = __ballot_sync(0xffffffffu, hit);
    if (mask) { pick = base + __ffs(mask) - 1; found = true; }
    before += __shfl_sync(0xffffffffu, incl, 31);
  }
  return pick;
}

__global__ void __launch_bounds__(kThreads) sample_kernel(
    const float* __restrict__ logits, int64_t ld, int vocab, const int* __restrict__ greedy,
    const float* __restrict__ temperature, const int* __restrict__ top_k, const float* __restrict__ top_p,
    uint32_t* __restrict__ rng_state, int* __restrict__ out_tokens, long long* __restrict__ out_tokens64) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  __shared__ int redi[32];
  __shared__ int hist[256];
  __shared__ float cval[kMaxCand];
  __shared__ int cidx[kMaxCand];
  __shared__ int s_count;
  __shared__ uint32_t s_prefix;
  __shared__ int s_krem;
  __shared__ int s_pick;
  __shared__ float s_f;

  const int row = blockIdx.x;
  const float* x = logits + (int64_t)row * ld;
  const int tid = threadIdx.x;

  // ---- pass 1: max / argmax (first index on ties)
  MaxIdx best{-INFINITY, 0x7fffffff};
  for (int i = tid; i < vocab; i += kThreads) best = better(best, MaxIdx{x[i], i});
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MaxIdx other{__shfl_xor_sync(0xffffffffu, best.v, o), __shfl_xor_sync(0xffffffffu, best.i, o)};
    best = better(best, other);
  }
  if ((tid & 31) == 0) { red[tid >> 5] = best.v; redi[tid >> 5] = best.i; }
  __syncthreads();
  if (tid < 32) {
    MaxIdx b{red[tid], redi[tid]};
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      MaxIdx other{__shfl_xor_sync(0xffffffffu, b.v, o), __shfl_xor_sync(0xffffffffu, b.i, o)};
      b = better(b, other);
    }
    if (tid == 0) { s_f = b.v; s_pick = b.i; }
  }
  // every thread reads the counter before the barrier, thread 0 advances it after: no thread can see step + 1
  const uint32_t seed = rng_state[2 * row], step = rng_state[2 * row + 1];
  __syncthreads();
  const float xmax = s_f;
  const int amax = s_pick;
  if (tid == 0) rng_state[2 * row + 1] = step + 1;
  if (greedy[row]) {
    if (tid == 0) {
      if (out_tokens) out_tokens[row] = amax;
      if (out_tokens64) out_tokens64[row] = amax;
    }
    return;
  }
  const float temp = temperature[row];
  const float inv_t = 1.0f / temp;
  const float tp = top_p[row];
  int k = top_k[row];
  const float u01 = rng_uniform(seed, step);

  if (k <= 0 && tp >= 1.0f) {
    // ---- plain multinomial over the whole row, inverse CDF in index order
    float part = 0.f;
    const int per = (vocab + kThreads - 1) / kThreads;
    const int i0 = tid * per, i1 = min(vocab, i0 + per);
    for (int i = i0; i < i1; ++i) part += __expf((x[i] - xmax) * inv_t);
    // block exclusive scan of `part`
    float incl = part;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      float n = __shfl_up_sync(0xffffffffu, incl, o);
      if ((tid & 31) >= o) incl += n;
    }
    __syncthreads();
    if ((tid & 31) == 31) red[tid >> 5] = incl;
    __syncthreads();
    if (tid < 32) {
      float w = red[tid], wi = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        float n = __shfl_up_sync(0xffffffffu, wi, o);
        if (tid >= o) wi += n;
      }
      red[tid] = wi - w;  // exclusive warp offsets
      if (tid == 31) s_f = wi;
    }
    if (tid == 0) s_count = kThreads;  // lowest thread whose cumulative weight exceeds the target
    __syncthreads();
    const float total = s_f;
    const float target = u01 * total;
    // Gap-free ownership: thread t owns [cum(t-1), cum(t)) with ONE expression for cum (its inclusive prefix), so
    // neighbouring intervals share their end points exactly; the owner is the lowest thread with target < cum.
    const float cum = red[tid >> 5] + incl;
    if (target < cum) atomicMin(&s_count, tid);
    __syncthreads();
    if (tid == s_count) {
      float c = cum - part;
      int pick = -1, last_pos = -1;
      for (int i = i0; i < i1; ++i) {
        const float w = __expf((x[i] - xmax) * inv_t);
        c += w;
        if (w > 0.f) last_pos = i;
        if (target < c) { pick = i; break; }
      }
      // rounding (the walk re-adds the weights in another order than the scan) can leave the target just past
      // the last element: take the thread's last token of non-zero weight
      s_pick = pick >= 0 ? pick : (last_pos >= 0 ? last_pos : amax);
    } else if (tid == 0 && s_count == kThreads) {
      s_pick = amax;  // target >= total (u01 * total rounded up): any positive-probability token; take the mode
    }
    __syncthreads();
    if (tid == 0) {
      if (out_tokens) out_tokens[row] = s_pick;
      if (out_tokens64) out_tokens64[row] = s_pick;
    }
    return;
  }

  // ---- sum-exp at temperature 1 (top-p is applied to the raw distribution)
  float part = 0.f;
  for (int i = tid; i < vocab; i += kThreads) part += __expf(x[i] - xmax);
  const float Z = block_sum_f(part, red);

  // ---- k-th largest logit, candidates >= it, sorted (value desc, index asc)
  if (k <= 0 || k > kMaxCand) k = kMaxCand;
  if (k > vocab) k = vocab;
  const uint32_t thr_key = block_radix_kth(x, vocab, k, hist, &s_prefix, &s_krem);
  const int count = block_gather_sorted(x, vocab, thr_key, 0, cval, cidx, &s_count);
  // ---- top-p over the sorted candidates, temperature weights, inverse-CDF draw (thread 0..31)
  if (tid < 32) {
    const int pick = warp_draw_from_sorted(cval, count, xmax, Z, tp, inv_t, u01);
    if (tid == 0) {
      const int tok = cidx[pick];
      if (out_tokens) out_tokens[row] = tok;
      if (out_tokens64) out_tokens64[row] = tok;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Vocab-parallel greedy decoding (K11, greedy case): instead of all-gathering [B, V] fp32 logits, every
// rank reduces its own vocab shard to one (value, global index) candidate per row, the candidates
// (16 bytes per row and rank) are gathered, and every rank picks the winner -- same token on all ranks,
// ties resolved towards the lowest index exactly like the single-GPU sampler.
struct Cand { float v; int idx; int pad0, pad1; };

__global__ void __launch_bounds__(256) argmax_local_kernel(const float* __restrict__ logits, int64_t ld, int n_valid,
                                                          int idx_offset, Cand* __restrict__ cand) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float sv[8];
  __shared__ int si[8];
  const int row = blockIdx.x;
  const float* x = logits + (int64_t)row * ld;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int i = threadIdx.x; i < n_valid; i += blockDim.x) {
    const float v = x[i];
    if (v > best) { best = v; bi = i; }  // increasing i per thread: first occurrence wins
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w)
      if (sv[w] > best || (sv[w] == best && si[w] < bi)) { best = sv[w]; bi = si[w]; }
    Cand c;
    c.v = best;
    c.idx = (bi == 0x7fffffff) ? 0x7fffffff : bi + idx_offset;  // empty shard: never wins a tie
    c.pad0 = c.pad1 = 0;
    cand[row] = c;
  }
}

__global__ void argmax_merge_kernel(const Cand* __restrict__ cand, int nranks, int B, int* __restrict__ out32,
                                    long long* __restrict__ out64) {
  pdl_launch_dependents();
  pdl_wait();
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= B) return;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int r = 0; r < nranks; ++r) {
    const Cand c = cand[(int64_t)row * nranks + r];
    if (c.v > best || (c.v == best && c.idx < bi)) { best = c.v; bi = c.idx; }
  }
  if (bi == 0x7fffffff) bi = 0;
  if (out32) out32[row] = bi;
  if (out64) out64[row] = bi;
}

// ---------------------------------------------------------------------------------------------
// Vocab-parallel sampling (K11, general case): top-k / top-p / temperature without gathering [B, V] logits.
// Each rank condenses its vocab shard to a fixed-size record per row -- local max, local sum-exp (relative to that
// max) and its local top-k candidates, ties at the k-th value included, sorted -- the records are gathered
// (1 KiB per row and rank instead of V/t floats), and every rank merges them identically: global max and
// normaliser from the (max, sum-exp) pairs, union of the candidates sorted (value desc, index asc), global k-th
// value, nucleus, temperature, one draw with the request's counter RNG.  An element >= the global k-th value is
// >= its own rank's local k-th value, so the union is complete and the result equals sample_kernel's on the
// gathered logits (oracle: ops/reference.py:sample_vocab_parallel).
// Record layout (floats): [max, sumexp, count(int), sumexp at the row's temperature (top_k = 0 rows), vals[KS], idx[KS](int)].
__global__ void __launch_bounds__(kThreads) vp_summary_kernel(const float* __restrict__ logits, int64_t ld, int n_valid,
                                                              int idx_offset, const int* __restrict__ greedy,
                                                              const int* __restrict__ top_k,
                                                              const float* __restrict__ temperature,
                                                              float* __restrict__ recs, int KS) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  __shared__ int hist[256];
  __shared__ float cval[kMaxCand];
  __shared__ int cidx[kMaxCand];
  __shared__ int s_count;
  __shared__ uint32_t s_prefix;
  __shared__ int s_krem;
  __shared__ float s_f;
  const int row = blockIdx.x, tid = threadIdx.x;
  const float* x = logits + (int64_t)row * ld;
  float* rec = recs + (int64_t)row * (4 + 2 * KS);
  if (n_valid <= 0) {  // this rank's shard is all padding
    if (tid == 0) { rec[0] = -INFINITY; rec[1] = 0.f; reinterpret_cast<int*>(rec)[2] = 0; rec[3] = 0.f; }
    return;
  }
  float best = -INFINITY;
  for (int i = tid; i < n_valid; i += kThreads) best = fmaxf(best, x[i]);
  best = warp_max(best);
  if ((tid & 31) == 0) red[tid >> 5] = best;
  __syncthreads();
  if (tid < 32) {
    float b = warp_max(red[tid]);
    if (tid == 0) s_f = b;
  }
  __syncthreads();
  const float xmax = s_f;
  float part = 0.f;
  for (int i = tid; i < n_valid; i += kThreads) part += __expf(x[i] - xmax);
  const float sumexp = block_sum_f(part, red);
  // rows sampled from the whole distribution (top_k = 0) also need the weight sum at their temperature
  float sumexp_t = 0.f;
  if (!greedy[row] && top_k[row] <= 0 && temperature != nullptr) {
    const float inv_t = 1.0f / temperature[row];
    float pt = 0.f;
    for (int i = tid; i < n_valid; i += kThreads) pt += __expf((x[i] - xmax) * inv_t);
    sumexp_t = block_sum_f(pt, red);
  }
  int k = greedy[row] ? 1 : top_k[row];
  if (k <= 0 || k > KS) k = KS;
  if (k > n_valid) k = n_valid;
  const uint32_t thr_key = block_radix_kth(x, n_valid, k, hist, &s_prefix, &s_krem);
  int count = block_gather_sorted(x, n_valid, thr_key, idx_offset, cval, cidx, &s_count);
  count = min(count, KS);  // more than KS - k + 1 exact ties at the k-th value: the highest indices are dropped
  if (tid == 0) { rec[0] = xmax; rec[1] = sumexp; reinterpret_cast<int*>(rec)[2] = count; rec[3] = sumexp_t; }
  for (int i = tid; i < count; i += kThreads) {
    rec[4 + i] = cval[i];
    reinterpret_cast<int*>(rec)[4 + KS + i] = cidx[i];
  }
}

__global__ void __launch_bounds__(kThreads) vp_merge_sample_kernel(
    const float* __restrict__ recs, int nranks, int KS, int vocab, const int* __restrict__ greedy,
    const float* __restrict__ temperature, const int* __restrict__ top_k, const float* __restrict__ top_p,
    uint32_t* __restrict__ rng_state, int* __restrict__ out_tokens, long long* __restrict__ out_tokens64) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float cval[kMaxCand];
  __shared__ int cidx[kMaxCand];
  __shared__ float s_max[8], s_sum[8];
  __shared__ int s_off[9];
  const int row = blockIdx.x, tid = threadIdx.x;
  const int rf = 4 + 2 * KS;
  const float* base = recs + (int64_t)row * nranks * rf;
  if (tid == 0) {
    int off = 0;
    for (int r = 0; r < nranks; ++r) {
      s_max[r] = base[(int64_t)r * rf];
      s_sum[r] = base[(int64_t)r * rf + 1];
      s_off[r] = off;
      off = min(kMaxCand, off + min(KS, max(0, reinterpret_cast<const int*>(base + (int64_t)r * rf)[2])));
    }
    s_off[nranks] = off;
  }
  // every thread reads the counter before the barrier, thread 0 advances it after
  const uint32_t seed = rng_state[2 * row], step = rng_state[2 * row + 1];
  __syncthreads();
  if (tid == 0) rng_state[2 * row + 1] = step + 1;
  float M = -INFINITY;
  for (int r = 0; r < nranks; ++r) M = fmaxf(M, s_max[r]);
  float Z = 0.f;
  for (int r = 0; r < nranks; ++r)
    if (s_sum[r] > 0.f) Z += s_sum[r] * __expf(s_max[r] - M);  // rank order: identical on every rank
  const int n_union = s_off[nranks];
  for (int r = 0; r < nranks; ++r) {
    const int o = s_off[r], c = s_off[r + 1] - o;
    const float* rv = base + (int64_t)r * rf + 4;
    const int* ri = reinterpret_cast<const int*>(base + (int64_t)r * rf) + 4 + KS;
    for (int i = tid; i < c; i += kThreads) { cval[o + i] = rv[i]; cidx[o + i] = ri[i]; }
  }
  __syncthreads();
  block_sort_candidates(cval, cidx, n_union);
  if (greedy[row] || n_union <= 1) {
    if (tid == 0) {
      const int tok = n_union > 0 ? cidx[0] : 0;
      if (out_tokens) out_tokens[row] = tok;
      if (out_tokens64) out_tokens64[row] = tok;
    }
    return;
  }
  int k = top_k[row];
  if (k <= 0 || k > KS) k = KS;
  k = min(min(k, vocab), n_union);
  const float thr_k = cval[k - 1];
  const int count = __syncthreads_count(tid < n_union && cval[tid] >= thr_k);  // sorted: a prefix, ties included
  if (tid < 32) {
    const float u01 = rng_uniform(seed, step);
    const int pick = warp_draw_from_sorted(cval, count, M, Z, top_p[row], 1.0f / temperature[row], u01);
    if (tid == 0) {
      const int tok = cidx[pick];
      if (out_tokens) out_tokens[row] = tok;
      if (out_tokens64) out_tokens64[row] = tok;
    }
  }
}

// ---- vocab-parallel draw from the WHOLE distribution (top_k = 0, top_p = 1: what the reference's sampling loop
// effectively does, quirk Q1).  From the gathered record headers every rank knows the global max and every rank's
// weight sum at the row's temperature, hence the total and each rank's CDF interval in rank (= index) order.  The
// one rank whose interval holds u * total walks its own shard for the token; 16 bytes per row are gathered and
// the owner's answer replaces the row's token.  Runs BEFORE vp_merge_sample (which advances the RNG counter).
__global__ void __launch_bounds__(kThreads) vp_multinomial_pick_kernel(
    const float* __restrict__ logits, int64_t ld, int n_valid, int idx_offset, const float* __restrict__ recs_all,
    int nranks, int rank, int KS, const int* __restrict__ greedy, const float* __restrict__ temperature,
    const int* __restrict__ top_k, const float* __restrict__ top_p, const uint32_t* __restrict__ rng_state,
    int* __restrict__ cand /* [B][4] */) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  __shared__ int s_owner_thread;
  __shared__ int s_pick;
  __shared__ float s_f;
  const int row = blockIdx.x, tid = threadIdx.x;
  int* my = cand + (int64_t)row * 4;
  const bool plain = !greedy[row] && top_k[row] <= 0 && top_p[row] >= 1.0f;
  if (!plain) {
    if (tid == 0) { my[0] = -1; my[1] = my[2] = my[3] = 0; }
    return;
  }
  const int rf = 4 + 2 * KS;
  const float* base = recs_all + (int64_t)row * nranks * rf;
  const float inv_t = 1.0f / temperature[row];
  float M = -INFINITY;
  for (int r = 0; r < nranks; ++r) M = fmaxf(M, base[(int64_t)r * rf]);
  float lower = 0.f, mine_w = 0.f, total = 0.f;
  int last_nonempty = -1;
  for (int r = 0; r < nranks; ++r) {  // rank order = index order; identical arithmetic on every rank
    const float s_r = base[(int64_t)r * rf + 3];
    const float w = s_r > 0.f ? s_r * __expf((base[(int64_t)r * rf] - M) * inv_t) : 0.f;
    if (r == rank) { lower = total; mine_w = w; }
    if (w > 0.f) last_nonempty = r;
    total += w;
  }
  const float u01 = rng_uniform(rng_state[2 * row], rng_state[2 * row + 1]);
  const float target = u01 * total;
  // owner = first rank whose upper bound exceeds the target; the last non-empty rank also takes target >= total
  int owner = last_nonempty;
  {
    float acc = 0.f;
    for (int r = 0; r < nranks; ++r) {
      const float s_r = base[(int64_t)r * rf + 3];
      const float w = s_r > 0.f ? s_r * __expf((base[(int64_t)r * rf] - M) * inv_t) : 0.f;
      acc += w;
      if (w > 0.f && target < acc) { owner = r; break; }
    }
  }
  if (owner != rank || n_valid <= 0) {
    if (tid == 0) { my[0] = -1; my[1] = my[2] = my[3] = 0; }
    return;
  }
  // local units: weights relative to this rank's own maximum
  const float xmax = base[(int64_t)rank * rf];
  const float s_local = base[(int64_t)rank * rf + 3];
  const float t_local = fminf(fmaxf((target - lower) / fmaxf(mine_w, 1e-30f), 0.f), 1.f) * s_local;
  const float* x = logits + (int64_t)row * ld;
  const int per = (n_valid + kThreads - 1) / kThreads;
  const int i0 = tid * per, i1 = min(n_valid, i0 + per);
  float part = 0.f;
  for (int i = i0; i < i1; ++i) part += __expf((x[i] - xmax) * inv_t);
  float incl = part;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float n = __shfl_up_sync(0xffffffffu, incl, o);
    if ((tid & 31) >= o) incl += n;
  }
  if ((tid & 31) == 31) red[tid >> 5] = incl;
  if (tid == 0) { s_owner_thread = kThreads; s_pick = -1; }
  __syncthreads();
  if (tid < 32) {
    float w = red[tid], wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      float n = __shfl_up_sync(0xffffffffu, wi, o);
      if (tid >= o) wi += n;
    }
    red[tid] = wi - w;
    if (tid == 31) s_f = wi;
  }
  __syncthreads();
  const float cum = red[tid >> 5] + incl;
  if (part > 0.f && t_local < cum) atomicMin(&s_owner_thread, tid);
  __syncthreads();
  if (tid == s_owner_thread) {
    float c = cum - part;
    int pick = -1, last_pos = -1;
    for (int i = i0; i < i1; ++i) {
      const float w = __expf((x[i] - xmax) * inv_t);
      c += w;
      if (w > 0.f) last_pos = i;
      if (t_local < c) { pick = i; break; }
    }
    s_pick = pick >= 0 ? pick : last_pos;
  }
  __syncthreads();
  if (s_owner_thread == kThreads) {
    // rounding left the target past this shard's total: the last token of positive weight
    int lp = -1;
    for (int i = i1 - 1; i >= i0; --i)
      if (__expf((x[i] - xmax) * inv_t) > 0.f) { lp = i; break; }
    if (lp >= 0) atomicMax(&s_pick, lp);
    __syncthreads();
  }
  if (tid == 0) {
    my[0] = (s_pick >= 0 ? s_pick : 0) + idx_offset;
    my[1] = my[2] = my[3] = 0;
  }
}

__global__ void vp_multinomial_merge_kernel(const int* __restrict__ cand_all, int nranks, int B, int* __restrict__ out32,
                                            long long* __restrict__ out64) {
  pdl_launch_dependents();
  pdl_wait();
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= B) return;
  int tok = -1;
  for (int r = 0; r < nranks; ++r) tok = max(tok, cand_all[((int64_t)row * nranks + r) * 4]);
  if (tok < 0) return;  // not a whole-distribution row: keep what vp_merge_sample wrote
  if (out32) out32[row] = tok;
  if (out64) out64[row] = tok;
}

}  // namespace

// cand: [B] 16-byte records (this rank's best logit of each row and its global vocabulary index).
LLMSS_API int llmss_argmax_local(const float* logits, int64_t ld, int B, int n_valid, int idx_offset, void* cand,
                                 cudaStream_t stream) {
  LLMSS_REQUIRE(n_valid >= 0 && ld >= n_valid, "argmax_local: bad shard width");
  if (B == 0) return 0;
  LLMSS_CHECK_CUDA(launch_pdl(argmax_local_kernel, dim3(B), dim3(256), 0, stream, logits, ld, n_valid, idx_offset,
                              (Cand*)cand));
  return 0;
}

// cand_all: [B][nranks] records (the column-gathered candidates); writes the winning token of every row.
LLMSS_API int llmss_argmax_merge(const void* cand_all, int nranks, int B, int* out_tokens, long long* out_tokens64,
                                 cudaStream_t stream) {
  if (B == 0) return 0;
  LLMSS_CHECK_CUDA(launch_pdl(argmax_merge_kernel, dim3((B + 127) / 128), dim3(128), 0, stream, (const Cand*)cand_all,
                              nranks, B, out_tokens, out_tokens64));
  return 0;
}

// Vocab-parallel sampling, step 1 (every rank, on its own fp32 shard [B, >= n_valid] whose column 0 is global token
// idx_offset): one record of llmss_vp_record_floats(ks) floats per row into `recs`.  ks = candidates kept per rank
// (even, so records are 16-byte multiples; 1 <= top_k <= ks is exact; nranks * ks <= 1024).
LLMSS_API int llmss_vp_record_floats(int ks) { return 4 + 2 * ks; }
LLMSS_API int llmss_vp_summary(const float* logits, int64_t ld, int B, int n_valid, int idx_offset, const int* greedy,
                               const int* top_k, const float* temperature, float* recs, int ks, cudaStream_t stream) {
  LLMSS_REQUIRE(ks >= 2 && ks <= kMaxCand && ks % 2 == 0, "vp_summary: ks=%d must be even and within [2, %d]", ks, kMaxCand);
  LLMSS_REQUIRE(n_valid >= 0 && ld >= n_valid, "vp_summary: bad shard width");
  if (B == 0) return 0;
  LLMSS_CHECK_CUDA(launch_pdl(vp_summary_kernel, dim3(B), dim3(kThreads), 0, stream, logits, ld, n_valid, idx_offset,
                              greedy, top_k, temperature, recs, ks));
  return 0;
}
// Step 2 (every rank, on the column-gathered records [B][nranks][record]): merge + sample; advances rng_state.
LLMSS_API int llmss_vp_merge_sample(const float* recs_all, int nranks, int ks, int B, int vocab, const int* greedy,
                                    const float* temperature, const int* top_k, const float* top_p,
                                    uint32_t* rng_state, int* out_tokens, long long* out_tokens64, cudaStream_t stream) {
  LLMSS_REQUIRE(nranks >= 1 && nranks <= 8 && nranks * ks <= kMaxCand, "vp_merge_sample: %d ranks x %d candidates exceed %d",
                nranks, ks, kMaxCand);
  if (B == 0) return 0;
  LLMSS_CHECK_CUDA(launch_pdl(vp_merge_sample_kernel, dim3(B), dim3(kThreads), 0, stream, recs_all, nranks, ks, vocab,
                              greedy, temperature, top_k, top_p, rng_state, out_tokens, out_tokens64));
  return 0;
}

// Whole-distribution rows (top_k = 0, top_p = 1) of a vocab-parallel batch.  Call order per step: llmss_vp_summary ->
// gather records -> llmss_vp_multinomial_pick (every rank, on its shard + the gathered records; does NOT advance the RNG)
// -> gather `cand` ([B][4] int32, 16 bytes per row) with llmss_allgather_cols -> llmss_vp_merge_sample (all rows, advances
// the RNG) -> llmss_vp_multinomial_merge (overwrites the tokens of the whole-distribution rows).
LLMSS_API int llmss_vp_multinomial_pick(const float* logits, int64_t ld, int B, int n_valid, int idx_offset,
                                        const float* recs_all, int nranks, int rank, int ks, const int* greedy,
                                        const float* temperature, const int* top_k, const float* top_p,
                                        const uint32_t* rng_state, int* cand, cudaStream_t stream) {
  LLMSS_REQUIRE(nranks >= 1 && nranks <= 8 && rank >= 0 && rank < nranks, "vp_multinomial_pick: bad rank %d of %d", rank, nranks);
  if (B == 0) return 0;
  LLMSS_CHECK_CUDA(launch_pdl(vp_multinomial_pick_kernel, dim3(B), dim3(kThreads), 0, stream, logits, ld, n_valid, idx_offset,
                              recs_all, nranks, rank, ks, greedy, temperature, top_k, top_p, rng_state, cand));
  return 0;
}
LLMSS_API int llmss_vp_multinomial_merge(const int* cand_all, int nranks, int B, int* out_tokens, long long* out_tokens64,
                                         cudaStream_t stream) {
  if (B == 0) return 0;
  LLMSS_CHECK_CUDA(launch_pdl(vp_multinomial_merge_kernel, dim3((B + 127) / 128), dim3(128), 0, stream, cand_all, nranks, B,
                              out_tokens, out_tokens64));
  return 0;
}

LLMSS_API int llmss_sample(const float* logits, int64_t ld, int B, int vocab, const int* greedy,
                           const float* temperature, const int* top_k, const float* top_p, uint32_t* rng_state,
                           int* out_tokens, long long* out_tokens64, cudaStream_t stream) {
  LLMSS_REQUIRE(vocab > 0 && ld >= vocab, "sample: bad vocab/ld");
  if (B == 0) return 0;
  LLMSS_CHECK_CUDA(launch_pdl(sample_kernel, dim3(B), dim3(kThreads), 0, stream, logits, ld, vocab, greedy, temperature, top_k, top_p, rng_state, out_tokens, out_tokens64));
  LLMSS_CHECK_CUDA(cudaGetLastError());
  return 0;
}
