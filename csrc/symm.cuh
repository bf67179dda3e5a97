// Symmetric-memory workspace descriptor shared by comm.cu and the fused GEMM epilogue.
#pragma once
#include "common.cuh"

constexpr int kSymmMaxRanks = 8;
constexpr int kSymmMaxSlots = 2048;                                     // CTA slots per flag set
constexpr size_t kSymmFlagSetBytes = (size_t)kSymmMaxSlots * kSymmMaxRanks * 4;  // 64 KiB
constexpr int kSymmFlagSets = 2;
constexpr size_t kSymmDataOffset = kSymmFlagSetBytes * kSymmFlagSets;   // data buffers start here

// Passed to kernels by value (fixed addresses -> graph capturable).
struct SymmWs {
  char* peer_base[kSymmMaxRanks];  // this process' mapping of every rank's region
  char* mc_base;                   // NVLS multicast mapping of the same region on all ranks (null: P2P only)
  uint32_t* epoch;                 // device-local: epoch of the last completed call
  uint32_t* ticket;                // device-local: CTAs finished in the current call
  size_t buf_bytes;                // bytes per parity buffer of the flag-synchronised (bulk) area
  size_t lam_bytes;                // bytes per parity buffer of the self-validating (Lamport) area
  int rank, nranks;
};

// region layout:  [flag set 0][flag set 1][lamport parity 0][lamport parity 1][bulk parity 0][bulk parity 1]
template <typename V>
__device__ __forceinline__ V* symm_buf(const SymmWs& ws, int rank, int parity) {
  return reinterpret_cast<V*>(ws.peer_base[rank] + kSymmDataOffset + 2 * ws.lam_bytes + (size_t)parity * ws.buf_bytes);
}
template <typename V>
__device__ __forceinline__ V* symm_lam(const SymmWs& ws, int rank, int parity) {
  return reinterpret_cast<V*>(ws.peer_base[rank] + kSymmDataOffset + (size_t)parity * ws.lam_bytes);
}

// ---- self-validating ("Lamport") exchange for latency-critical messages ----------------------
// The Lamport area is kept filled with a sentinel word that a sender never emits (16-bit payloads:
// -0.0 halves are rewritten to +0.0; fp32 payloads: that exact denormal is rewritten to 0).  A
// receiver spins on the 16-byte vector itself until none of its words is the sentinel, consumes it
// and puts the sentinel back.  No flags, no fences, one NVLink one-way latency.  Every pushed vector
// is read (and re-armed) by exactly one thread; parity double-buffering keeps call n+1's pushes away
// from call n's vectors (a peer can be at most one call ahead, see comm.cu).
constexpr uint32_t kLamSentinel = 0x80008000u;
template <int ELEM_BYTES>
__device__ __forceinline__ uint32_t lam_sanitize_word(uint32_t w) {
  if constexpr (ELEM_BYTES == 2) {
    if ((w & 0x0000FFFFu) == 0x00008000u) w &= 0xFFFF0000u;
    if ((w & 0xFFFF0000u) == 0x80000000u) w &= 0x0000FFFFu;
    return w;
  } else {
    return w == kLamSentinel ? 0u : w;
  }
}
template <int ELEM_BYTES>
__device__ __forceinline__ uint4 lam_sanitize(uint4 v) {
  v.x = lam_sanitize_word<ELEM_BYTES>(v.x);
  v.y = lam_sanitize_word<ELEM_BYTES>(v.y);
  v.z = lam_sanitize_word<ELEM_BYTES>(v.z);
  v.w = lam_sanitize_word<ELEM_BYTES>(v.w);
  return v;
}
__device__ __forceinline__ uint4 ld_relaxed_sys_v4(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}
// Wait for a vector pushed by a peer, re-arm the slot, return the payload.
__device__ __forceinline__ uint4 lam_take(uint4* slot, int tag) {
  SpinGuard g;
  uint4 v = ld_relaxed_sys_v4(slot);
  while (v.x == kLamSentinel || v.y == kLamSentinel || v.z == kLamSentinel || v.w == kLamSentinel) {
    if (!g.ok()) spin_timeout_trap("lamport vector wait", tag, 0);
    v = ld_relaxed_sys_v4(slot);
  }
  *slot = make_uint4(kLamSentinel, kLamSentinel, kLamSentinel, kLamSentinel);
  return v;
}
// ---- NVLS (NVSwitch multicast) accessors: same offsets as symm_buf / symm_lam / symm_flag, but through the
// multicast mapping: a store lands in every rank's copy, multimem.ld_reduce sums all copies in the switch.
template <typename V>
__device__ __forceinline__ V* symm_mc_buf(const SymmWs& ws, int parity) {
  return reinterpret_cast<V*>(ws.mc_base + kSymmDataOffset + 2 * ws.lam_bytes + (size_t)parity * ws.buf_bytes);
}
template <typename V>
__device__ __forceinline__ V* symm_mc_lam(const SymmWs& ws, int parity) {
  return reinterpret_cast<V*>(ws.mc_base + kSymmDataOffset + (size_t)parity * ws.lam_bytes);
}
__device__ __forceinline__ uint32_t* symm_mc_flag(const SymmWs& ws, int set, int slot) {
  return reinterpret_cast<uint32_t*>(ws.mc_base + (size_t)set * kSymmFlagSetBytes) + (size_t)slot * kSymmMaxRanks;
}
// Sum of the 16-byte vector at the same offset of every rank's copy (8 x 16-bit lanes, fp32 accumulation in
// the switch, one rounding) or 4 x fp32.
template <typename T>
__device__ __forceinline__ uint4 mc_ld_reduce(const void* mc_addr) {
  uint4 r;
  if constexpr (std::is_same<T, __nv_bfloat16>::value) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc_addr) : "memory");
  } else if constexpr (std::is_same<T, __half>::value) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc_addr) : "memory");
  } else {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc_addr) : "memory");
  }
  return r;
}
// Broadcast store: one instruction, the switch replicates it into every rank's copy.
__device__ __forceinline__ void mc_st_v4(void* mc_addr, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               ::"l"(mc_addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void mc_st_release_u32(uint32_t* mc_addr, uint32_t v) {
  asm volatile("multimem.st.release.sys.global.b32 [%0], %1;" ::"l"(mc_addr), "r"(v) : "memory");
}

// The same for up to N vectors at once: all loads are issued before any is examined, and only the vectors that
// are still missing are polled again.  A relaxed system-scope load is a full L2 round trip (~0.7 us) even when the
// data is already there, so taking the t contributions of an all-reduce one after the other costs t round trips;
// taken together they cost one after the last arrival.  slot[i] for i >= n is ignored (alias any valid slot).
template <int N>
__device__ __forceinline__ void lam_take_batch(uint4* const (&slot)[N], int n, uint4 (&v)[N], int tag) {
  uint32_t pending = (n >= 32) ? 0xffffffffu : ((1u << n) - 1u);
  SpinGuard g;
  while (true) {
#pragma unroll
    for (int i = 0; i < N; ++i)
      if ((pending >> i) & 1u) v[i] = ld_relaxed_sys_v4(slot[i]);
#pragma unroll
    for (int i = 0; i < N; ++i) {
      if (((pending >> i) & 1u) &&
          !(v[i].x == kLamSentinel || v[i].y == kLamSentinel || v[i].z == kLamSentinel || v[i].w == kLamSentinel))
        pending &= ~(1u << i);
    }
    if (pending == 0) break;
    if (!g.ok()) spin_timeout_trap("lamport vector wait", tag, (int)pending);
  }
#pragma unroll
  for (int i = 0; i < N; ++i)
    if (i < n) *slot[i] = make_uint4(kLamSentinel, kLamSentinel, kLamSentinel, kLamSentinel);
}
// &flag[set][slot][0] inside rank `rank`'s region; index by source rank.
__device__ __forceinline__ uint32_t* symm_flag(const SymmWs& ws, int rank, int set, int slot) {
  return reinterpret_cast<uint32_t*>(ws.peer_base[rank] + (size_t)set * kSymmFlagSetBytes) +
         (size_t)slot * kSymmMaxRanks;
}
// Last CTA out publishes the epoch (all CTAs read *epoch at their start, before this can run).
__device__ __forceinline__ void symm_finish(const SymmWs& ws, uint32_t e) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t total = gridDim.x * gridDim.y * gridDim.z;
    __threadfence();
    if (atomicAdd(ws.ticket, 1u) == total - 1) {
      *ws.ticket = 0;
      __threadfence();
      *ws.epoch = e;
    }
  }
}
