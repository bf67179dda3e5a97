// comm runtime + K4 collectives over NVLink 5 / NVSwitch peer memory.
//
// The reference issues every collective as a separate blocking NCCL launch through c10d
// (SURVEY.md sections 2.4/2.5; layers.py:125,133,178,213).  Here each rank owns one
// symmetric region (cudaMalloc + CUDA IPC, mapped into every peer once at start-up) and
// the collectives are plain kernels that store straight into peer memory and synchronise
// with system-scope release/acquire flags.  No host involvement, fixed addresses ->
// CUDA-graph capturable.  The same region/protocol is used from the epilogue of the fused
// GEMM -> all-reduce kernel (gemm_tcgen05.cu).
//
// Protocol (shared by all kernels on a workspace):
//   * epoch: device-local counter; a call uses e = *epoch + 1, the last CTA to finish
//     (ticket counter) publishes *epoch = e, so the next kernel in stream order sees it.
//   * data buffers are double-buffered by epoch parity.  A peer can only run one call
//     ahead (it cannot finish call n+1 without this rank's call-n+1 pushes), so the buffer
//     of call n is never overwritten before call n has been fully consumed.
//   * flag[slot][src_rank] on the destination holds the epoch of the latest push for that
//     CTA slot; consumers wait for (flag - e) >= 0.  Every CTA pushes before it waits and
//     grids never exceed one resident wave, so there is no cross-GPU deadlock.
//   * bounded spins (common.cuh) trap instead of hanging if a peer dies.
#include "common.cuh"
#include "symm.cuh"

#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

// ---------------------------------------------------------------- error string
static thread_local char g_err[1024] = "";
void llmss_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
LLMSS_API const char* llmss_last_error() { return g_err; }
LLMSS_API int llmss_version() { return 100; }

// ---------------------------------------------------------------- symmetric memory plumbing
bool llmss_pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    // Off by default: measured at TP=1 and TP=2 (GPT-J-6B decode, CUDA graph) it changes nothing --
    // every hot kernel is one fat persistent CTA per SM, so the next kernel's CTAs cannot become
    // resident before the previous ones retire anyway.  LLMSS_PDL=1 turns the launch attribute on.
    const char* e = getenv("LLMSS_PDL");
    v = (e != nullptr && e[0] == '1') ? 1 : 0;
  }
  return v != 0;
}
namespace {
__global__ void fill_u32_kernel(uint32_t* p, size_t n, uint32_t v) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}
}  // namespace

// One region per rank: flags (zeroed) + Lamport area (filled with the sentinel) + bulk area.
LLMSS_API int llmss_symm_alloc(size_t buf_bytes, size_t lam_bytes, void** ptr) {
  const size_t bytes = kSymmDataOffset + 2 * lam_bytes + 2 * buf_bytes;
  LLMSS_REQUIRE(lam_bytes % 16 == 0 && buf_bytes % 16 == 0, "symm_alloc: sizes must be multiples of 16");
  LLMSS_CHECK_CUDA(cudaMalloc(ptr, bytes));
  LLMSS_CHECK_CUDA(cudaMemset(*ptr, 0, bytes));
  if (lam_bytes > 0) {
    fill_u32_kernel<<<256, 256>>>(reinterpret_cast<uint32_t*>((char*)*ptr + kSymmDataOffset), 2 * lam_bytes / 4, kLamSentinel);
    LLMSS_CHECK_CUDA(cudaGetLastError());
  }
  LLMSS_CHECK_CUDA(cudaDeviceSynchronize());
  return 0;
}
// Same initial state for a region that was created elsewhere (vmm.cu: cuMemCreate + map).
LLMSS_API int llmss_symm_init_region(void* ptr, size_t buf_bytes, size_t lam_bytes) {
  const size_t bytes = kSymmDataOffset + 2 * lam_bytes + 2 * buf_bytes;
  LLMSS_REQUIRE(lam_bytes % 16 == 0 && buf_bytes % 16 == 0, "symm_init_region: sizes must be multiples of 16");
  LLMSS_CHECK_CUDA(cudaMemset(ptr, 0, bytes));
  if (lam_bytes > 0) {
    fill_u32_kernel<<<256, 256>>>(reinterpret_cast<uint32_t*>((char*)ptr + kSymmDataOffset), 2 * lam_bytes / 4, kLamSentinel);
    LLMSS_CHECK_CUDA(cudaGetLastError());
  }
  LLMSS_CHECK_CUDA(cudaDeviceSynchronize());
  return 0;
}
LLMSS_API int llmss_symm_free(void* ptr) {
  LLMSS_CHECK_CUDA(cudaFree(ptr));
  return 0;
}
LLMSS_API int llmss_ipc_get_handle(void* ptr, unsigned char* out64) {
  cudaIpcMemHandle_t h;
  LLMSS_CHECK_CUDA(cudaIpcGetMemHandle(&h, ptr));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(out64, &h, 64);
  return 0;
}
LLMSS_API int llmss_ipc_open_handle(const unsigned char* in64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, in64, 64);
  LLMSS_CHECK_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
LLMSS_API int llmss_ipc_close(void* ptr) {
  LLMSS_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}
LLMSS_API size_t llmss_symm_region_bytes(size_t buf_bytes, size_t lam_bytes) {
  return kSymmDataOffset + 2 * lam_bytes + 2 * buf_bytes;
}

// Build the by-value descriptor the kernels take.  `peer_bases[r]` is the mapping of rank
// r's region in this process (peer_bases[rank] = the local allocation).
// `mc_base`: the NVLS multicast mapping of the region (vmm.cu), or null when the heap is P2P-only.
LLMSS_API int llmss_symm_create2(int rank, int nranks, void* const* peer_bases, void* mc_base, size_t buf_bytes,
                                  size_t lam_bytes, SymmWs** out);
LLMSS_API int llmss_symm_create(int rank, int nranks, void* const* peer_bases, size_t buf_bytes, size_t lam_bytes,
                                 SymmWs** out) {
  return llmss_symm_create2(rank, nranks, peer_bases, nullptr, buf_bytes, lam_bytes, out);
}
// Failure-detection self test: waits for a flag nobody will ever raise, so the device watchdog (SpinGuard,
// common.cuh) must end the kernel with a trap after LLMSS_SPIN_TIMEOUT_NS instead of spinning for ever.  The CUDA
// context is unusable afterwards: run it in a process of its own (tests/test_gpu_watchdog.py).
__global__ void debug_wait_flag_kernel(const uint32_t* flag, uint32_t want) { wait_flag_ge(flag, want, 99); }
LLMSS_API int llmss_debug_wait_flag(const uint32_t* flag, uint32_t want, cudaStream_t stream) {
  debug_wait_flag_kernel<<<1, 1, 0, stream>>>(flag, want);
  LLMSS_CHECK_CUDA(cudaGetLastError());
  return 0;
}
LLMSS_API int llmss_symm_has_multicast(SymmWs* ws) { return (ws != nullptr && ws->mc_base != nullptr) ? 1 : 0; }
LLMSS_API int llmss_symm_create2(int rank, int nranks, void* const* peer_bases, void* mc_base, size_t buf_bytes,
                                  size_t lam_bytes, SymmWs** out) {
  LLMSS_REQUIRE(nranks >= 1 && nranks <= kSymmMaxRanks, "symm_create: nranks=%d unsupported (max %d)", nranks,
                kSymmMaxRanks);
  SymmWs* ws = new SymmWs();
  memset(ws, 0, sizeof(*ws));
  for (int r = 0; r < nranks; ++r) ws->peer_base[r] = (char*)peer_bases[r];
  ws->mc_base = (char*)mc_base;
  ws->rank = rank;
  ws->nranks = nranks;
  ws->buf_bytes = buf_bytes;
  ws->lam_bytes = lam_bytes;
  uint32_t* ctl = nullptr;
  LLMSS_CHECK_CUDA(cudaMalloc(&ctl, 256));
  LLMSS_CHECK_CUDA(cudaMemset(ctl, 0, 256));
  ws->epoch = ctl;
  ws->ticket = ctl + 32;
  LLMSS_CHECK_CUDA(cudaDeviceSynchronize());
  *out = ws;
  return 0;
}
LLMSS_API int llmss_symm_destroy(SymmWs* ws) {
  if (ws) {
    cudaFree(ws->epoch);
    delete ws;
  }
  return 0;
}
LLMSS_API int llmss_symm_epoch(SymmWs* ws, uint32_t* out) {
  LLMSS_CHECK_CUDA(cudaMemcpy(out, ws->epoch, 4, cudaMemcpyDeviceToHost));
  return 0;
}

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ uint4 ld_cg_v4(const void* p) {  // L2-only load: peers write these lines
  uint4 r;
  asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

template <typename T>
__device__ __forceinline__ void accum8(float (&acc)[8], const uint4& raw) {
  float f[8];
  unpack8<T>(raw, f);
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] += f[i];
}
template <>
__device__ __forceinline__ void accum8<float>(float (&acc)[8], const uint4& raw) {
  acc[0] += __uint_as_float(raw.x);
  acc[1] += __uint_as_float(raw.y);
  acc[2] += __uint_as_float(raw.z);
  acc[3] += __uint_as_float(raw.w);
}

// ---- one-shot all-reduce: every rank pushes its whole input to every peer, then sums locally.
// Latency-optimal for decode-size messages: self-validating vectors (symm.cuh), so the whole
// collective is one NVLink one-way trip -- no flags, no fences, no block barriers.
template <typename T>
__global__ void __launch_bounds__(kThreads) allreduce_oneshot_kernel(SymmWs ws, const uint4* __restrict__ in,
                                                                     uint4* __restrict__ out, int64_t nvec) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int EB = std::is_same<T, float>::value ? 4 : 2;
  const uint32_t e = *ws.epoch + 1;
  const int par = e & 1;
  const int64_t per_cta = (nvec + gridDim.x - 1) / gridDim.x;
  const int64_t v0 = (int64_t)blockIdx.x * per_cta, v1 = min(nvec, v0 + per_cta);

  // push my chunk into slot [rank] of every peer's buffer
  for (int64_t v = v0 + threadIdx.x; v < v1; v += kThreads) {
    const uint4 val = lam_sanitize<EB>(in[v]);
#pragma unroll 1
    for (int p = 1; p < ws.nranks; ++p) {
      const int peer = (ws.rank + p) % ws.nranks;
      symm_lam<uint4>(ws, peer, par)[(int64_t)ws.rank * nvec + v] = val;
    }
  }
  uint4* mine = symm_lam<uint4>(ws, ws.rank, par);
  for (int64_t v = v0 + threadIdx.x; v < v1; v += kThreads) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    // the peers' vectors are polled together: one L2 round trip after the last arrival instead of one per peer
    uint4* slot[kSymmMaxRanks];
    uint4 got[kSymmMaxRanks];
    int m = 0;
    for (int r = 0; r < ws.nranks; ++r)
      if (r != ws.rank) slot[m++] = mine + (int64_t)r * nvec + v;
    for (int i = m; i < kSymmMaxRanks; ++i) slot[i] = slot[0];
    lam_take_batch<kSymmMaxRanks>(slot, m, got, 1);
    m = 0;
    for (int r = 0; r < ws.nranks; ++r) {  // fixed order -> bit-identical on all ranks
      const uint4 raw = (r == ws.rank) ? lam_sanitize<EB>(in[v]) : got[m++];
      accum8<T>(acc, raw);
    }
    uint4 o;
    if constexpr (std::is_same<T, float>::value) {
      o = make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
    } else {
      o = pack8<T>(acc);
    }
    out[v] = o;
  }
  symm_finish(ws, e);
}

// ---- two-shot all-reduce: reduce-scatter (each rank owns nvec/nranks vectors) + all-gather.
// Per-GPU NVLink traffic 2*(t-1)/t * bytes instead of (t-1) * bytes: the bandwidth regime.
template <typename T>
__global__ void __launch_bounds__(kThreads) allreduce_twoshot_kernel(SymmWs ws, const uint4* __restrict__ in,
                                                                     uint4* __restrict__ out, int64_t nvec) {
  pdl_launch_dependents();
  pdl_wait();
  const uint32_t e = *ws.epoch + 1;
  const int par = e & 1;
  const int n = ws.nranks;
  const int64_t chunk = nvec / n;  // host guarantees divisibility
  const int64_t per_cta = (chunk + gridDim.x - 1) / gridDim.x;
  const int64_t c0 = (int64_t)blockIdx.x * per_cta, c1 = min(chunk, c0 + per_cta);
  // buffer layout per parity: [n][chunk] partials (phase 1)  then  [n][chunk] results (phase 2)

  // phase 1: send sub-chunk of chunk `peer` to its owner
  for (int64_t c = c0 + threadIdx.x; c < c1; c += kThreads) {
#pragma unroll 1
    for (int p = 1; p < n; ++p) {
      const int peer = (ws.rank + p) % n;
      symm_buf<uint4>(ws, peer, par)[(int64_t)ws.rank * chunk + c] = in[(int64_t)peer * chunk + c];
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) __threadfence_system();
  __syncthreads();
  if (threadIdx.x < n && threadIdx.x != ws.rank) st_release_sys(symm_flag(ws, threadIdx.x, 0, blockIdx.x) + ws.rank, e);
  if (threadIdx.x < n && threadIdx.x != ws.rank) wait_flag_ge(symm_flag(ws, ws.rank, 0, blockIdx.x) + threadIdx.x, e, 2);
  __syncthreads();

  // phase 2: reduce my chunk, broadcast the result
  const uint4* mine = symm_buf<uint4>(ws, ws.rank, par);
  for (int64_t c = c0 + threadIdx.x; c < c1; c += kThreads) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int r = 0; r < n; ++r) {
      const uint4 raw = (r == ws.rank) ? in[(int64_t)ws.rank * chunk + c] : ld_cg_v4(mine + (int64_t)r * chunk + c);
      accum8<T>(acc, raw);
    }
    uint4 o;
    if constexpr (std::is_same<T, float>::value) {
      o = make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
    } else {
      o = pack8<T>(acc);
    }
    out[(int64_t)ws.rank * chunk + c] = o;
#pragma unroll 1
    for (int p = 1; p < n; ++p) {
      const int peer = (ws.rank + p) % n;
      symm_buf<uint4>(ws, peer, par)[(int64_t)(n + ws.rank) * chunk + c] = o;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) __threadfence_system();
  __syncthreads();
  if (threadIdx.x < n && threadIdx.x != ws.rank) st_release_sys(symm_flag(ws, threadIdx.x, 1, blockIdx.x) + ws.rank, e);
  if (threadIdx.x < n && threadIdx.x != ws.rank) wait_flag_ge(symm_flag(ws, ws.rank, 1, blockIdx.x) + threadIdx.x, e, 3);
  __syncthreads();

  // phase 3: collect the other owners' results
  for (int64_t c = c0 + threadIdx.x; c < c1; c += kThreads) {
#pragma unroll 1
    for (int p = 1; p < n; ++p) {
      const int r = (ws.rank + p) % n;
      out[(int64_t)r * chunk + c] = ld_cg_v4(mine + (int64_t)(n + r) * chunk + c);
    }
  }
  symm_finish(ws, e);
}


// ---- NVLS all-reduce: reduce-scatter and all-gather both happen inside the NVSwitch.
// Every rank stages its input in its own bulk buffer (local stores), raises one multicast flag, then
// reduces 1/t of each CTA's range with multimem.ld_reduce (the switch reads the t copies and returns
// their sum: fp32 accumulation, one rounding) and broadcasts the result with a multicast store.
// Per-GPU NVLink traffic: (t-1)/t * bytes out + bytes/t in for the reduction, bytes/t out + (t-1)/t in
// for the broadcast -- a t-th of the one-shot push, half of the P2P two-shot -- and only two one-way
// trips plus one round trip of latency.  LAM: results travel as self-validating vectors through the
// Lamport area (no second flag / fence); otherwise through the bulk area behind a second flag.
template <typename T, bool LAM>
__global__ void __launch_bounds__(kThreads) allreduce_nvls_kernel(SymmWs ws, const uint4* __restrict__ in,
                                                                  uint4* __restrict__ out, int64_t nvec) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int EB = std::is_same<T, float>::value ? 4 : 2;
  const uint32_t e = *ws.epoch + 1;
  const int par = e & 1;
  const int n = ws.nranks;
  const int64_t per_cta = (nvec + gridDim.x - 1) / gridDim.x;
  const int64_t v0 = min(nvec, (int64_t)blockIdx.x * per_cta), v1 = min(nvec, v0 + per_cta);
  uint4* part = symm_buf<uint4>(ws, ws.rank, par);
  constexpr int U = 4;
  for (int64_t vb = v0 + threadIdx.x; vb < v1; vb += (int64_t)kThreads * U) {
    uint4 val[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (vb + (int64_t)u * kThreads < v1) val[u] = in[vb + (int64_t)u * kThreads];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (vb + (int64_t)u * kThreads < v1) part[vb + (int64_t)u * kThreads] = val[u];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    mc_st_release_u32(symm_mc_flag(ws, 0, blockIdx.x) + ws.rank, e);  // one store, seen by every rank
  }
  if (threadIdx.x < n) wait_flag_ge(symm_flag(ws, ws.rank, 0, blockIdx.x) + threadIdx.x, e, 5);
  __syncthreads();

  const int64_t sub = (v1 - v0 + n - 1) / n;
  const int64_t b0 = min(v1, v0 + ws.rank * sub), b1 = min(v1, b0 + sub);
  const uint4* mc_part = symm_mc_buf<uint4>(ws, par);
  uint4* mc_res = LAM ? symm_mc_lam<uint4>(ws, par) : symm_mc_buf<uint4>(ws, par) + nvec;
  for (int64_t vb = b0 + threadIdx.x; vb < b1; vb += (int64_t)kThreads * U) {
    uint4 val[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (vb + (int64_t)u * kThreads < b1) val[u] = mc_ld_reduce<T>(mc_part + vb + (int64_t)u * kThreads);
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (vb + (int64_t)u * kThreads < b1) mc_st_v4(mc_res + vb + (int64_t)u * kThreads, LAM ? lam_sanitize<EB>(val[u]) : val[u]);
  }
  if constexpr (LAM) {
    uint4* mine = symm_lam<uint4>(ws, ws.rank, par);
  for (int64_t vb = v0 + threadIdx.x; vb < v1; vb += (int64_t)kThreads * 4) {
    uint4* slot[4];
    uint4 got[4];
    int m = 0;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t v = vb + (int64_t)u * kThreads;
      slot[u] = mine + (v < v1 ? v : vb);
      if (v < v1) m = u + 1;
    }
    lam_take_batch<4>(slot, m, got, 6);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (u < m) out[vb + (int64_t)u * kThreads] = got[u];
  }
  } else {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      mc_st_release_u32(symm_mc_flag(ws, 1, blockIdx.x) + ws.rank, e);
    }
    if (threadIdx.x < n) wait_flag_ge(symm_flag(ws, ws.rank, 1, blockIdx.x) + threadIdx.x, e, 7);
    __syncthreads();
    const uint4* res = symm_buf<uint4>(ws, ws.rank, par) + nvec;
    for (int64_t vb = v0 + threadIdx.x; vb < v1; vb += (int64_t)kThreads * U) {
      uint4 val[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (vb + (int64_t)u * kThreads < v1) val[u] = ld_cg_v4(res + vb + (int64_t)u * kThreads);
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (vb + (int64_t)u * kThreads < v1) out[vb + (int64_t)u * kThreads] = val[u];
    }
  }
  symm_finish(ws, e);
}

// ---- two-shot all-reduce on self-validating vectors (no flags, no fences): reduce-scatter by peer stores into the
// owner's Lamport inbox, owner sums in rank order and broadcasts (one multicast store through the switch when the
// heap is multicast-bound).  Two one-way trips, ~1 x the message per GPU on the wire: the mid-size regime
// (>= 4 ranks, ~100 KB .. a few MB) between the one-shot push and the bandwidth kernels.
template <typename T>
__global__ void __launch_bounds__(kThreads) allreduce_twoshot_lam_kernel(SymmWs ws, const uint4* __restrict__ in,
                                                                         uint4* __restrict__ out, int64_t nvec) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int EB = std::is_same<T, float>::value ? 4 : 2;
  const uint32_t e = *ws.epoch + 1;
  const int par = e & 1;
  const int n = ws.nranks, me = ws.rank;
  const int64_t per_cta = (nvec + gridDim.x - 1) / gridDim.x;
  const int64_t v0 = min(nvec, (int64_t)blockIdx.x * per_cta), v1 = min(nvec, v0 + per_cta);
  const int64_t sub = max((int64_t)1, (v1 - v0 + n - 1) / n);
  for (int64_t v = v0 + threadIdx.x; v < v1; v += kThreads)
    symm_lam<uint4>(ws, (int)((v - v0) / sub), par)[(int64_t)me * nvec + v] = lam_sanitize<EB>(in[v]);
  uint4* mine = symm_lam<uint4>(ws, me, par);
  const int64_t b0 = min(v1, v0 + me * sub), b1 = min(v1, b0 + sub);
  for (int64_t v = b0 + threadIdx.x; v < b1; v += kThreads) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint4* slot[kSymmMaxRanks];
    uint4 got[kSymmMaxRanks];
    for (int r = 0; r < kSymmMaxRanks; ++r) slot[r] = mine + (int64_t)(r < n ? r : 0) * nvec + v;
    lam_take_batch<kSymmMaxRanks>(slot, n, got, 8);
    for (int r = 0; r < n; ++r) accum8<T>(acc, got[r]);  // rank order
    uint4 o;
    if constexpr (std::is_same<T, float>::value) {
      o = make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
    } else {
      o = pack8<T>(acc);
    }
    o = lam_sanitize<EB>(o);
    if (ws.mc_base != nullptr) {
      mc_st_v4(symm_mc_lam<uint4>(ws, par) + (int64_t)n * nvec + v, o);
    } else {
#pragma unroll 1
      for (int p = 0; p < n; ++p) symm_lam<uint4>(ws, (me + p) % n, par)[(int64_t)n * nvec + v] = o;
    }
  }
  for (int64_t vb = v0 + threadIdx.x; vb < v1; vb += (int64_t)kThreads * 4) {
    uint4* slot[4];
    uint4 got[4];
    int m = 0;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t v = vb + (int64_t)u * kThreads;
      slot[u] = mine + (int64_t)n * nvec + (v < v1 ? v : vb);
      if (v < v1) m = u + 1;
    }
    lam_take_batch<4>(slot, m, got, 9);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (u < m) out[vb + (int64_t)u * kThreads] = got[u];
  }
  symm_finish(ws, e);
}

// ---- all-gather: out[r] = rank r's input (used for the vocab-parallel logits, reference N3)
// Input: [rows][vpr] 16-byte vectors (contiguous).  Output: [rows][nranks * vpr], rank r's block at column
// r * vpr -- i.e. the concatenation along the last dimension, written in place (rows = 1 gives the plain
// [nranks][n] gather).  Four vectors are in flight per thread in both phases.
__global__ void __launch_bounds__(kThreads) allgather_kernel(SymmWs ws, const uint4* __restrict__ in,
                                                             uint4* __restrict__ out, int64_t nvec, int64_t vpr) {
  pdl_launch_dependents();
  pdl_wait();
  const uint32_t e = *ws.epoch + 1;
  const int par = e & 1;
  const int64_t per_cta = (nvec + gridDim.x - 1) / gridDim.x;
  const int64_t v0 = (int64_t)blockIdx.x * per_cta, v1 = min(nvec, v0 + per_cta);
  const int64_t pitch = vpr * ws.nranks;
  constexpr int U = 4;
  for (int64_t vb = v0 + threadIdx.x; vb < v1; vb += (int64_t)kThreads * U) {
    uint4 val[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t v = vb + (int64_t)u * kThreads;
      if (v < v1) val[u] = in[v];
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t v = vb + (int64_t)u * kThreads;
      if (v >= v1) continue;
      out[(v / vpr) * pitch + (int64_t)ws.rank * vpr + v % vpr] = val[u];
      if (ws.mc_base != nullptr) {  // NVLS: one store, replicated by the switch into every rank's buffer
        mc_st_v4(symm_mc_buf<uint4>(ws, par) + (int64_t)ws.rank * nvec + v, val[u]);
        continue;
      }
#pragma unroll 1
      for (int p = 1; p < ws.nranks; ++p) {
        const int peer = (ws.rank + p) % ws.nranks;
        symm_buf<uint4>(ws, peer, par)[(int64_t)ws.rank * nvec + v] = val[u];
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) __threadfence_system();
  __syncthreads();
  if (threadIdx.x < ws.nranks && threadIdx.x != ws.rank)
    st_release_sys(symm_flag(ws, threadIdx.x, 0, blockIdx.x) + ws.rank, e);
  if (threadIdx.x < ws.nranks && threadIdx.x != ws.rank)
    wait_flag_ge(symm_flag(ws, ws.rank, 0, blockIdx.x) + threadIdx.x, e, 4);
  __syncthreads();
  const uint4* mine = symm_buf<uint4>(ws, ws.rank, par);
#pragma unroll 1
  for (int p = 1; p < ws.nranks; ++p) {
    const int r = (ws.rank + p) % ws.nranks;
    for (int64_t vb = v0 + threadIdx.x; vb < v1; vb += (int64_t)kThreads * U) {
      uint4 val[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t v = vb + (int64_t)u * kThreads;
        if (v < v1) val[u] = ld_cg_v4(mine + (int64_t)r * nvec + v);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t v = vb + (int64_t)u * kThreads;
        if (v < v1) out[(v / vpr) * pitch + (int64_t)r * vpr + v % vpr] = val[u];
      }
    }
  }
  symm_finish(ws, e);
}

int pick_grid(int64_t nvec, int max_ctas) {
  int64_t g = (nvec + kThreads - 1) / kThreads;  // one 16-byte vector per thread per pass
  if (g < 1) g = 1;
  if (g > max_ctas) g = max_ctas;
  return (int)g;
}

}  // namespace

// Message size (bytes) from which the automatic choice prefers the in-switch (NVLS) reduction over the one-shot
// push; LLMSS_NVLS_MIN_BYTES overrides it, LLMSS_NVLS=0 disables the multicast paths altogether.
int64_t llmss_nvls_min_bytes() {
  static int64_t v = -1;
  if (v < 0) {
    const char* off = getenv("LLMSS_NVLS");
    const char* e = getenv("LLMSS_NVLS_MIN_BYTES");
    if (off != nullptr && off[0] == '0') v = INT64_MAX;
    else v = (e != nullptr) ? atoll(e) : (int64_t)(4 << 20);
  }
  return v;
}

int64_t llmss_twoshot_min_bytes() {
  static int64_t v = -1;
  if (v < 0) {
    const char* e = getenv("LLMSS_TWOSHOT_MIN_BYTES");
    v = (e != nullptr) ? atoll(e) : (int64_t)(256 << 10);
  }
  return v;
}

// algo: 0 = auto, 1 = one-shot, 2 = two-shot (flags), 3 = NVLS (multimem reduce + broadcast), 4 = two-shot on Lamport
// vectors (+ multicast broadcast).  `n` = element count.
LLMSS_API int llmss_allreduce(SymmWs* ws, const void* in, void* out, int64_t n, int dtype, int algo,
                              cudaStream_t stream) {
  const int esz = (dtype == LLMSS_F32) ? 4 : 2;
  const int64_t bytes = n * esz;
  LLMSS_REQUIRE(bytes % 16 == 0, "allreduce: byte size %lld must be a multiple of 16", (long long)bytes);
  LLMSS_REQUIRE(((uintptr_t)in % 16 == 0) && ((uintptr_t)out % 16 == 0), "allreduce: pointers must be 16-byte aligned");
  const int64_t nvec = bytes / 16;
  if (ws->nranks == 1) {
    if (in != out) LLMSS_CHECK_CUDA(cudaMemcpyAsync(out, in, bytes, cudaMemcpyDeviceToDevice, stream));
    return 0;
  }
  const bool nvls_can_lam = (size_t)bytes <= ws->buf_bytes && (size_t)bytes <= ws->lam_bytes;
  const bool nvls_can_flag = (size_t)bytes * 2 <= ws->buf_bytes;
  const bool nvls_fits = ws->mc_base != nullptr && (nvls_can_lam || nvls_can_flag);
  const bool two_lam_fits = (size_t)bytes * (ws->nranks + 1) <= ws->lam_bytes;
  if (algo == 0 && nvls_fits && bytes >= llmss_nvls_min_bytes()) algo = 3;
  if (algo == 0 && two_lam_fits && ((ws->nranks >= 4 && bytes >= llmss_twoshot_min_bytes()) || bytes > (int64_t)(1 << 20))) algo = 4;
  if (algo == 4) {
    LLMSS_REQUIRE(two_lam_fits, "allreduce: message of %lld bytes needs %lld bytes of Lamport buffer (have %zu)",
                  (long long)bytes, (long long)bytes * (ws->nranks + 1), ws->lam_bytes);
    const int grid = pick_grid(nvec, 128);
    SymmWs w = *ws;
    if (dtype == LLMSS_F16) LLMSS_CHECK_CUDA(launch_pdl(allreduce_twoshot_lam_kernel<__half>, dim3(grid), dim3(kThreads), 0, stream, w, (const uint4*)in, (uint4*)out, nvec));
    else if (dtype == LLMSS_BF16) LLMSS_CHECK_CUDA(launch_pdl(allreduce_twoshot_lam_kernel<__nv_bfloat16>, dim3(grid), dim3(kThreads), 0, stream, w, (const uint4*)in, (uint4*)out, nvec));
    else LLMSS_CHECK_CUDA(launch_pdl(allreduce_twoshot_lam_kernel<float>, dim3(grid), dim3(kThreads), 0, stream, w, (const uint4*)in, (uint4*)out, nvec));
    LLMSS_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  if (algo == 0)
    algo = ((bytes <= (int64_t)(1 << 20) && (size_t)(bytes * ws->nranks) <= ws->lam_bytes) || nvec % ws->nranks != 0) ? 1 : 2;
  if (algo == 2 && nvec % ws->nranks != 0) algo = 1;
  if (algo == 3) {
    LLMSS_REQUIRE(ws->mc_base != nullptr, "allreduce: the NVLS algorithm needs a multicast-bound workspace");
    LLMSS_REQUIRE(nvls_fits, "allreduce: message of %lld bytes does not fit the workspace for the NVLS algorithm", (long long)bytes);
    const bool lam = nvls_can_lam && (bytes <= (int64_t)(4 << 20) || !nvls_can_flag);
    const int grid = pick_grid((nvec + 3) / 4, 144);
    SymmWs w = *ws;
#define LLMSS_NVLS_AR(T_)                                                                                              \
    LLMSS_CHECK_CUDA(lam ? launch_pdl(allreduce_nvls_kernel<T_, true>, dim3(grid), dim3(kThreads), 0, stream, w,        \
                                      (const uint4*)in, (uint4*)out, nvec)                                             \
                         : launch_pdl(allreduce_nvls_kernel<T_, false>, dim3(grid), dim3(kThreads), 0, stream, w,       \
                                      (const uint4*)in, (uint4*)out, nvec))
    if (dtype == LLMSS_F16) LLMSS_NVLS_AR(__half);
    else if (dtype == LLMSS_BF16) LLMSS_NVLS_AR(__nv_bfloat16);
    else LLMSS_NVLS_AR(float);
#undef LLMSS_NVLS_AR
    LLMSS_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  const int64_t need = (algo == 1) ? bytes * ws->nranks : bytes * 2;
  const size_t have = (algo == 1) ? ws->lam_bytes : ws->buf_bytes;
  LLMSS_REQUIRE((size_t)need <= have, "allreduce: message of %lld bytes needs %lld bytes of workspace (have %zu)",
                (long long)bytes, (long long)need, have);
  SymmWs w = *ws;
  if (algo == 1) {
    const int grid = pick_grid(nvec, 128);
    if (dtype == LLMSS_F16) LLMSS_CHECK_CUDA(launch_pdl(allreduce_oneshot_kernel<__half>, dim3(grid), dim3(kThreads), 0, stream, w, (const uint4*)in, (uint4*)out, nvec));
    else if (dtype == LLMSS_BF16) LLMSS_CHECK_CUDA(launch_pdl(allreduce_oneshot_kernel<__nv_bfloat16>, dim3(grid), dim3(kThreads), 0, stream, w, (const uint4*)in, (uint4*)out, nvec));
    else LLMSS_CHECK_CUDA(launch_pdl(allreduce_oneshot_kernel<float>, dim3(grid), dim3(kThreads), 0, stream, w, (const uint4*)in, (uint4*)out, nvec));
  } else {
    const int grid = pick_grid(nvec / ws->nranks, 144);
    if (dtype == LLMSS_F16) LLMSS_CHECK_CUDA(launch_pdl(allreduce_twoshot_kernel<__half>, dim3(grid), dim3(kThreads), 0, stream, w, (const uint4*)in, (uint4*)out, nvec));
    else if (dtype == LLMSS_BF16) LLMSS_CHECK_CUDA(launch_pdl(allreduce_twoshot_kernel<__nv_bfloat16>, dim3(grid), dim3(kThreads), 0, stream, w, (const uint4*)in, (uint4*)out, nvec));
    else LLMSS_CHECK_CUDA(launch_pdl(allreduce_twoshot_kernel<float>, dim3(grid), dim3(kThreads), 0, stream, w, (const uint4*)in, (uint4*)out, nvec));
  }
  LLMSS_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Gather along the last dimension: in is [rows][row_bytes], out is [rows][nranks * row_bytes] (rank r's block at
// byte column r * row_bytes) -- torch.cat(all_gather(x), dim=-1) without the intermediate copy (vocab-parallel
// logits).  rows = 1 is the plain gather: out holds nranks * bytes, slot r receives rank r's input.
LLMSS_API int llmss_allgather_cols(SymmWs* ws, const void* in, void* out, int64_t rows, int64_t row_bytes,
                                   cudaStream_t stream) {
  const int64_t bytes = rows * row_bytes;
  LLMSS_REQUIRE(row_bytes % 16 == 0, "allgather: row size %lld must be a multiple of 16 bytes", (long long)row_bytes);
  LLMSS_REQUIRE(((uintptr_t)in % 16 == 0) && ((uintptr_t)out % 16 == 0), "allgather: pointers must be 16-byte aligned");
  if (bytes == 0) return 0;
  if (ws->nranks == 1) {
    if (in != out) LLMSS_CHECK_CUDA(cudaMemcpyAsync(out, in, bytes, cudaMemcpyDeviceToDevice, stream));
    return 0;
  }
  LLMSS_REQUIRE((size_t)(bytes * ws->nranks) <= ws->buf_bytes, "allgather: %lld bytes x %d ranks exceeds workspace %zu",
                (long long)bytes, ws->nranks, ws->buf_bytes);
  const int64_t nvec = bytes / 16;
  LLMSS_CHECK_CUDA(launch_pdl(allgather_kernel, dim3(pick_grid((nvec + 3) / 4, 128)), dim3(kThreads), 0, stream, *ws,
                              (const uint4*)in, (uint4*)out, nvec, row_bytes / 16));
  return 0;
}

// out must hold nranks * bytes; slot r receives rank r's input.
LLMSS_API int llmss_allgather(SymmWs* ws, const void* in, void* out, int64_t bytes, cudaStream_t stream) {
  LLMSS_REQUIRE(bytes % 16 == 0, "allgather: byte size %lld must be a multiple of 16", (long long)bytes);
  LLMSS_REQUIRE(((uintptr_t)in % 16 == 0) && ((uintptr_t)out % 16 == 0), "allgather: pointers must be 16-byte aligned");
  if (ws->nranks == 1) {
    if (in != out) LLMSS_CHECK_CUDA(cudaMemcpyAsync(out, in, bytes, cudaMemcpyDeviceToDevice, stream));
    return 0;
  }
  LLMSS_REQUIRE((size_t)(bytes * ws->nranks) <= ws->buf_bytes, "allgather: %lld bytes x %d ranks exceeds workspace %zu",
                (long long)bytes, ws->nranks, ws->buf_bytes);
  const int64_t nvec = bytes / 16;
  LLMSS_CHECK_CUDA(launch_pdl(allgather_kernel, dim3(pick_grid((nvec + 3) / 4, 128)), dim3(kThreads), 0, stream, *ws,
                              (const uint4*)in, (uint4*)out, nvec, nvec));
  LLMSS_CHECK_CUDA(cudaGetLastError());
  return 0;
}
