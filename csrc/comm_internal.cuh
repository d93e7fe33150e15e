// Shared definitions of the symmetric-memory context (comm.cu, comm_vmm.cu).
#pragma once
#include "common.cuh"

#include <vector>

namespace tadnn {

constexpr int kMaxRanks = TADNN_COMM_MAX_RANKS;
constexpr int kChannels = TADNN_COMM_CHANNELS;
constexpr int kMaxBlocks = TADNN_COMM_MAX_BLOCKS;
constexpr size_t kPadBytes = static_cast<size_t>(kChannels) * kMaxBlocks * kMaxRanks * sizeof(uint32_t);
constexpr unsigned long long kDefaultTimeoutNs = 120ull * 1000ull * 1000ull * 1000ull;  // 120 s; tadnn_comm_set_timeout

struct CommDev {
  uint8_t* heap[kMaxRanks];
  uint32_t* pad[kMaxRanks];
  uint8_t* mc_heap;  // multicast mapping of the heaps (NVLS) or nullptr
  int rank;          // -1: single-launch emulation, rank = blockIdx.y
  int world;
  // Failure reporting (no __trap: a trap poisons the CUDA context of every survivor).  A barrier that
  // waited longer than timeout_ns records WHO it was waiting for in `status` (a word in mapped pinned
  // host memory the host reads without a CUDA call) and sets `poison` (device memory); every later
  // barrier on this context sees the poison after a few hundred spins and falls through, so the
  // stream drains quickly and the host can raise an error that names the missing rank.
  unsigned long long timeout_ns;
  uint32_t* poison;  // device word, 0 = healthy
  uint32_t* status;  // host-mapped word: 0, or 1 + (waiting rank << 20 | awaited peer << 12 | channel << 4 | phase)
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
  // failure reporting
  uint32_t* status_host = nullptr;  // cudaHostAlloc (mapped); dev.status is its device alias
  void* poison_dev = nullptr;       // cudaMalloc, 16 bytes
};

namespace tadnn {
int comm_init_status(tadnn_comm* c);     // allocates status/poison words and fills c->dev (comm.cu)
void comm_free_status(tadnn_comm* c);
}

