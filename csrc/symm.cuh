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
  uint32_t* epoch;                 // device-local: epoch of the last completed call
  uint32_t* ticket;                // device-local: CTAs finished in the current call
  size_t buf_bytes;                // bytes per parity buffer
  int rank, nranks;
};

// region layout:  [flag set 0][flag set 1][data parity 0 : buf_bytes][data parity 1 : buf_bytes]
template <typename V>
__device__ __forceinline__ V* symm_buf(const SymmWs& ws, int rank, int parity) {
  return reinterpret_cast<V*>(ws.peer_base[rank] + kSymmDataOffset + (size_t)parity * ws.buf_bytes);
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
