// Shared device/host helpers for the llmss_b200 sm_100a kernels.
//
// Everything here is hand-written PTX wrappers for the Blackwell execution model:
// mbarrier pipelines, TMA (cp.async.bulk.tensor), tcgen05 (alloc / mma / commit / ld),
// system-scope acquire/release for the NVLink peer-memory protocols, and a bounded-spin
// watchdog so a lost peer or a broken pipeline traps instead of hanging the GPU
// (the reference's only failure detection is the 60 s c10d timeout).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <utility>
#include <type_traits>

#define LLMSS_API extern "C" __attribute__((visibility("default")))

enum { LLMSS_F16 = 0, LLMSS_BF16 = 1, LLMSS_F32 = 2 };
enum { LLMSS_ACT_NONE = 0, LLMSS_ACT_GELU_TANH = 1, LLMSS_ACT_GELU_ERF = 2, LLMSS_ACT_RELU = 3, LLMSS_ACT_SILU = 4 };
enum { LLMSS_NORM_LAYER = 0, LLMSS_NORM_RMS = 1 };
enum { LLMSS_ROPE_NONE = 0, LLMSS_ROPE_GPTJ = 1, LLMSS_ROPE_NEOX = 2 };

// ---------------------------------------------------------------- host error plumbing
void llmss_set_error(const char* fmt, ...);
#define LLMSS_CHECK_CUDA(expr)                                                         \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess) {                                                           \
      llmss_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr,               \
                      cudaGetErrorString(_e));                                         \
      return (int)_e;                                                                  \
    }                                                                                  \
  } while (0)
#define LLMSS_REQUIRE(cond, ...)                                                       \
  do {                                                                                 \
    if (!(cond)) {                                                                     \
      llmss_set_error(__VA_ARGS__);                                                    \
      return -1;                                                                       \
    }                                                                                  \
  } while (0)

// cudaFuncSetAttribute applies to the CURRENT device only: a process that drives several GPUs has to repeat it on
// each.  One of these per kernel instantiation remembers, per device ordinal, the dynamic shared memory size
// already granted.
struct PerDeviceSmem {
  int granted[64] = {0};
  template <typename K>
  cudaError_t ensure(K kern, int bytes) {
    int d = 0;
    cudaError_t e = cudaGetDevice(&d);
    if (e != cudaSuccess) return e;
    d &= 63;
    if (granted[d] >= bytes) return cudaSuccess;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) granted[d] = bytes;
    return e;
  }
};

// ---------------------------------------------------------------- dtype helpers
template <typename T> struct Vec2;
template <> struct Vec2<__half> { using type = __half2; };
template <> struct Vec2<__nv_bfloat16> { using type = __nv_bfloat162; };

__device__ __forceinline__ float to_f32(__half v) { return __half2float(v); }
__device__ __forceinline__ float to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_f32(float v) { return v; }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }

// 8 x 16-bit elements <-> 8 floats through one 128-bit access
template <typename T>
__device__ __forceinline__ void unpack8(const uint4& raw, float (&f)[8]) {
  using T2 = typename Vec2<T>::type;
  const T2* p = reinterpret_cast<const T2*>(&raw);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t;
    if constexpr (sizeof(T) == 2 && std::is_same<T, __half>::value) t = __half22float2(p[i]);
    else t = __bfloat1622float2(p[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
template <typename T>
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  using T2 = typename Vec2<T>::type;
  uint4 raw;
  T2* p = reinterpret_cast<T2*>(&raw);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if constexpr (std::is_same<T, __half>::value) p[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
    else p[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  }
  return raw;
}

__device__ __forceinline__ float act_apply(float x, int act) {
  switch (act) {
    case LLMSS_ACT_GELU_TANH: {
      const float k = 0.7978845608028654f;  // sqrt(2/pi)
      float u = k * (x + 0.044715f * x * x * x);
      return 0.5f * x * (1.0f + tanhf(u));
    }
    case LLMSS_ACT_GELU_ERF: return 0.5f * x * (1.0f + erff(x * 0.7071067811865475f));
    case LLMSS_ACT_RELU: return fmaxf(x, 0.0f);
    case LLMSS_ACT_SILU: return x / (1.0f + __expf(-x));
    default: return x;
  }
}

// ---------------------------------------------------------------- warp / block reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---------------------------------------------------------------- programmatic dependent launch
// Every kernel of the library starts with pdl_launch_dependents() (the next kernel of the stream may be
// scheduled as soon as all our CTAs are resident) and calls pdl_wait() before it touches anything the
// previous kernel may have written: barrier init, TMEM allocation, descriptor prefetch and the launch
// latency itself then overlap the tail of the previous kernel.  Both are no-ops for a plain launch.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

bool llmss_pdl_enabled();  // comm.cu: LLMSS_PDL=1 enables the launch attribute (default off: no measured gain)

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = llmss_pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

// ---------------------------------------------------------------- watchdog
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#ifndef LLMSS_SPIN_TIMEOUT_NS
#define LLMSS_SPIN_TIMEOUT_NS 4000000000ull  // 4 s: far above any legitimate wait
#endif
struct SpinGuard {
  uint64_t t0 = 0;
  uint32_t n = 0;
  // returns false once the deadline has passed
  __device__ __forceinline__ bool ok() {
    if ((++n & 0x3fffu) == 0) {
      uint64_t now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > LLMSS_SPIN_TIMEOUT_NS) return false;
    }
    return true;
  }
};
static __device__ __noinline__ void spin_timeout_trap(const char* what, int a, int b) {
  printf("[llmss_b200] device watchdog: %s timed out (block %d,%d,%d thread %d; %d %d)\n", what,
         blockIdx.x, blockIdx.y, blockIdx.z, threadIdx.x, a, b);
  __trap();
}

// ---------------------------------------------------------------- shared-memory addresses
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, int tag = 0) {
  SpinGuard g;
  while (!mbar_try_wait(bar, parity)) {
    if (!g.ok()) spin_timeout_trap("mbarrier wait", tag, (int)parity);
  }
}

// One lane of a converged warp.  Issuing tcgen05 / TMA instructions under this predicate (instead of a
// long `if (lane == 0)` region) keeps the surrounding loop state and address arithmetic warp-uniform, so
// ptxas feeds UTCHMMA / UTMALDG from uniform registers; inside a divergent single-lane region it emits a
// vector->uniform "waterfall" (ELECT / R2UR.BROADCAST / BRA.U.ANY, ~18 instructions) per instruction.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

// ---------------------------------------------------------------- TMA
__device__ __forceinline__ void tma_prefetch_desc(const void* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
// 2-D tile load: coordinates are (c0 = innermost/contiguous, c1 = row)
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* map, uint64_t* bar, int c0, int c1,
                                            uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}

// ---------------------------------------------------------------- tcgen05 / TMEM
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {  // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {  // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; single thread issues.
__device__ __forceinline__ void umma_f16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on an mbarrier when all previously issued tcgen05.mma of this thread have completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// ---- CTA-pair forms (cta_group::2): the two CTAs of a cluster whose ranks differ in the last bit share one MMA.
// Each CTA holds half of the A rows and half of the B rows at the SAME shared-memory offsets, the even CTA issues,
// and each CTA's tensor memory receives the accumulator rows of its own A half.
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_dst, uint32_t ncols) {  // same warp id in both CTAs
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish_pair() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {  // same warp id in both CTAs
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_f16_ss_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                 uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on the mbarrier at this shared-memory offset in every CTA of `cta_mask` (bit = rank in the cluster) once the
// MMAs issued so far have completed.
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar, uint16_t cta_mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"(cta_mask)
               : "memory");
}
// Arrive on an mbarrier of another CTA of the cluster.  Default semantics (release at CTA scope), the form CUTLASS uses
// for its cross-CTA pipeline barriers: a release at CLUSTER scope compiles to MEMBAR.ALL.GPU (~1 us), and what the
// arrival announces here -- TMA data that landed in this CTA's shared memory -- was written and is read by the async
// proxy, ordered by the mbarrier's own transaction count, not by this thread's memory operations.
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// 32 lanes x 32 consecutive fp32 columns: thread t of the warp gets lane (base_lane + t).
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32b_x8(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// UMMA shared-memory descriptor for a K-major, 128B-swizzled tile whose rows are 128 bytes
// (64 x 16-bit) apart: 8-row swizzle atoms are 1024 B apart (SBO), LBO unused (=1),
// descriptor version 1 (sm_100), layout_type 2 = SWIZZLE_128B.
__device__ __forceinline__ uint64_t umma_desc_k_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);  // [0,14)  start address
  d |= (uint64_t)1 << 16;                        // [16,30) leading byte offset (ignored for SW128 K-major)
  d |= (uint64_t)(1024 >> 4) << 32;              // [32,46) stride byte offset
  d |= (uint64_t)1 << 46;                        // [46,48) version = 1
  d |= (uint64_t)2 << 61;                        // [61,64) SWIZZLE_128B
  return d;
}
// Instruction descriptor, kind::f16: fp32 accumulate, A/B both K-major, dtype 0 = f16 / 1 = bf16.
__host__ __device__ constexpr uint32_t umma_idesc_f16(int ab_fmt, int M, int N) {
  return (1u << 4)                      // c_format = F32
         | ((uint32_t)ab_fmt << 7)      // a_format
         | ((uint32_t)ab_fmt << 10)     // b_format
         | (0u << 15) | (0u << 16)      // a_major, b_major = K
         | ((uint32_t)(N >> 3) << 17)   // n_dim
         | ((uint32_t)(M >> 4) << 24);  // m_dim
}

// ---------------------------------------------------------------- system-scope flags (NVLink peers)
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void wait_flag_ge(const uint32_t* p, uint32_t want, int tag) {
  SpinGuard g;
  while ((int32_t)(ld_acquire_sys(p) - want) < 0) {
    if (!g.ok()) spin_timeout_trap("peer flag wait", tag, (int)want);
  }
}
