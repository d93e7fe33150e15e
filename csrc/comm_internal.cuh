// Shared definitions of the symmetric-memory context (comm.cu, comm_vmm.cu).
#pragma once
#include "common.cuh"

#include <vector>

namespace tadnn {

constexpr int kMaxRanks = TADNN_COMM_MAX_RANKS;
constexpr int kChannels = TADNN_COMM_CHANNELS;
constexpr int kMaxBlocks = TADNN_COMM_MAX_BLOCKS;
constexpr size_t kPadBytes = static_cast<size_t>(kChannels) * kMaxBlocks * kMaxRanks * sizeof(uint32_t);
constexpr unsigned long long kSpinTimeoutNs = 20ull * 1000ull * 1000ull * 1000ull;  // 20 s

struct CommDev {
  uint8_t* heap[kMaxRanks];
  uint32_t* pad[kMaxRanks];
  uint8_t* mc_heap;  // multicast mapping of the heaps (NVLS) or nullptr
  int rank;          // -1: single-launch emulation, rank = blockIdx.y
  int world;
};

}  // namespace tadnn

struct tadnn_comm {
  tadnn::CommDev dev;
  int rank = 0, world = 1;
  int device = 0;
  unsigned long long heap_bytes = 0;
  void* base = nullptr;  // own allocation (heap + pad); nullptr if attached
  bool owns = false;
  bool local_set = false;  // created by tadnn_comm_create_local
  std::vector<void*> opened;  // cudaIpcOpenMemHandle results to close
  // VMM / multicast backend (comm_vmm.cu)
  bool vmm = false;
  unsigned long long alloc_bytes = 0;     // heap + pad rounded up to the allocation granularity
  unsigned long long mem_handle = 0;      // CUmemGenericAllocationHandle of the own heap
  unsigned long long mc_handle = 0;       // CUmemGenericAllocationHandle of the multicast object
  std::vector<unsigned long long> peer_handles;  // imported peer allocations
  std::vector<void*> peer_vas;                   // their mappings
  void* mc_va = nullptr;
};

