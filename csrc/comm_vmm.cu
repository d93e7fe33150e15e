// VMM + multicast backend of the symmetric-memory context: the heap is a cuMemCreate allocation
// exported as a POSIX file descriptor (peers import and map it), and -- where the driver exposes
// NVLS -- a multicast object is created over all ranks' heaps and mapped, which is what the
// multimem.ld_reduce / multimem.st all-reduce kernel (K3, comm.cu) addresses.
//
// File descriptors travel between the rank processes over AF_UNIX sockets (SCM_RIGHTS); that
// exchange is done by the Python side (parallel/symm.py), this file only produces/consumes fds.
// The driver API is reached through cudaGetDriverEntryPoint, so libtadnn.so has no link-time
// dependency on libcuda (it must load on a machine without a GPU for the build check).
#include "comm_internal.cuh"

#include <unistd.h>

namespace tadnn {
namespace {

struct Drv {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemGetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*McCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*McAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*McBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*McGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*DeviceGet)(CUdevice*, int);
  bool ok = false;
};

template <typename F>
bool load(F& fn, const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) return false;
  fn = reinterpret_cast<F>(p);
  return true;
}

Drv& drv() {
  static Drv d = [] {  // C++11 magic static: initialised once, thread-safe
    Drv d;
    d.ok = load(d.MemCreate, "cuMemCreate") && load(d.MemRelease, "cuMemRelease") &&
           load(d.MemAddressReserve, "cuMemAddressReserve") && load(d.MemAddressFree, "cuMemAddressFree") &&
           load(d.MemMap, "cuMemMap") && load(d.MemUnmap, "cuMemUnmap") && load(d.MemSetAccess, "cuMemSetAccess") &&
           load(d.MemExport, "cuMemExportToShareableHandle") && load(d.MemImport, "cuMemImportFromShareableHandle") &&
           load(d.MemGetGranularity, "cuMemGetAllocationGranularity") && load(d.McCreate, "cuMulticastCreate") &&
           load(d.McAddDevice, "cuMulticastAddDevice") && load(d.McBindMem, "cuMulticastBindMem") &&
           load(d.McGetGranularity, "cuMulticastGetGranularity") && load(d.DeviceGetAttribute, "cuDeviceGetAttribute") &&
           load(d.DeviceGet, "cuDeviceGet");
    return d;
  }();
  return d;
}

#define TADNN_CHECK_CU(expr)                                                                      \
  do {                                                                                            \
    CUresult _r = (expr);                                                                         \
    if (_r != CUDA_SUCCESS) {                                                                     \
      ::tadnn::set_error("%s:%d: %s -> CUresult %d", __FILE__, __LINE__, #expr, static_cast<int>(_r)); \
      return 1;                                                                                   \
    }                                                                                             \
  } while (0)

CUmemAllocationProp heap_prop(int device) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

// Reserve a VA range, map `h` into it read-write for `device`.  On failure nothing stays reserved or
// mapped and *va is 0.
int map_rw(Drv& d, CUdeviceptr* va, size_t bytes, size_t gran, CUmemGenericAllocationHandle h, int device) {
  *va = 0;
  CUdeviceptr p = 0;
  TADNN_CHECK_CU(d.MemAddressReserve(&p, bytes, gran, 0, 0));
  CUresult r = d.MemMap(p, bytes, 0, h, 0);
  if (r != CUDA_SUCCESS) {
    d.MemAddressFree(p, bytes);
    set_error("%s:%d: cuMemMap -> CUresult %d", __FILE__, __LINE__, static_cast<int>(r));
    return 1;
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d.MemSetAccess(p, bytes, &acc, 1);
  if (r != CUDA_SUCCESS) {
    d.MemUnmap(p, bytes);
    d.MemAddressFree(p, bytes);
    set_error("%s:%d: cuMemSetAccess -> CUresult %d", __FILE__, __LINE__, static_cast<int>(r));
    return 1;
  }
  *va = p;
  return 0;
}

}  // namespace
}  // namespace tadnn

using namespace tadnn;

extern "C" int tadnn_multicast_supported(int* out) {
  TADNN_REQUIRE(out != nullptr, "multicast_supported: null out");
  *out = 0;
  Drv& d = drv();
  if (!d.ok) return 0;
  int dev = 0;
  TADNN_CHECK_CUDA(cudaGetDevice(&dev));
  TADNN_CHECK_CUDA(cudaFree(nullptr));  // make sure the primary context exists
  CUdevice cud;
  TADNN_CHECK_CU(d.DeviceGet(&cud, dev));
  int v = 0;
  if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cud) == CUDA_SUCCESS) *out = v;
  return 0;
}

static int create_vmm_into(tadnn_comm* c, int rank, int world, unsigned long long heap_bytes) {
  Drv& d = drv();
  c->rank = rank; c->world = world; c->vmm = true; c->owns = true;
  TADNN_CHECK_CUDA(cudaGetDevice(&c->device));
  heap_bytes = (heap_bytes + 1023ull) & ~1023ull;
  c->heap_bytes = heap_bytes;
  CUmemAllocationProp prop = heap_prop(c->device);
  size_t gran = 0;
  TADNN_CHECK_CU(d.MemGetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  int mc = 0;
  tadnn_multicast_supported(&mc);
  if (mc) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = static_cast<unsigned>(world);
    mp.size = heap_bytes + kPadBytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (d.McGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran) gran = mg;
  }
  c->alloc_bytes = ((heap_bytes + kPadBytes + gran - 1) / gran) * gran;
  CUmemGenericAllocationHandle h;
  TADNN_CHECK_CU(d.MemCreate(&h, c->alloc_bytes, &prop, 0));
  c->mem_handle = h;
  CUdeviceptr va = 0;
  if (int rc = map_rw(d, &va, c->alloc_bytes, gran, h, c->device)) return rc;
  c->base = reinterpret_cast<void*>(va);
  TADNN_CHECK_CUDA(cudaMemset(static_cast<uint8_t*>(c->base) + heap_bytes, 0, kPadBytes));
  TADNN_CHECK_CUDA(cudaDeviceSynchronize());
  memset(&c->dev, 0, sizeof(c->dev));
  c->dev.rank = rank; c->dev.world = world;
  c->dev.heap[rank] = static_cast<uint8_t*>(c->base);
  c->dev.pad[rank] = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(c->base) + heap_bytes);
  c->peer_handles.assign(world, 0);
  c->peer_vas.assign(world, nullptr);
  return comm_init_status(c);
}

extern "C" int tadnn_comm_create_vmm(int rank, int world, unsigned long long heap_bytes, tadnn_comm** out) {
  TADNN_REQUIRE(out != nullptr, "comm_create_vmm: null out");
  *out = nullptr;
  TADNN_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, "comm_create_vmm: bad rank/world %d/%d", rank, world);
  TADNN_REQUIRE(drv().ok, "comm_create_vmm: the CUDA driver does not export the VMM/multicast entry points");
  TADNN_CHECK_CUDA(cudaFree(nullptr));
  tadnn_comm* c = new tadnn_comm();
  if (int rc = create_vmm_into(c, rank, world, heap_bytes)) {
    tadnn_comm_vmm_release(c);  // unmaps / releases whatever was set up before the failure
    comm_free_status(c);
    delete c;
    return rc;
  }
  *out = c;
  return 0;
}

extern "C" int tadnn_comm_export_fd(tadnn_comm* c, int* fd) {
  TADNN_REQUIRE(c && c->vmm && fd, "comm_export_fd: not a VMM context");
  int f = -1;
  TADNN_CHECK_CU(drv().MemExport(&f, c->mem_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd = f;
  return 0;
}

extern "C" int tadnn_comm_import_peer_fd(tadnn_comm* c, int peer, int fd) {
  TADNN_REQUIRE(c && c->vmm && peer >= 0 && peer < c->world && peer != c->rank, "comm_import_peer_fd: bad argument");
  Drv& d = drv();
  CUmemGenericAllocationHandle h;
  const CUresult ir = d.MemImport(&h, reinterpret_cast<void*>(static_cast<intptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  close(fd);  // this call owns the descriptor; the driver keeps its own reference after a successful import
  TADNN_REQUIRE(ir == CUDA_SUCCESS, "comm_import_peer_fd: cuMemImportFromShareableHandle -> CUresult %d", static_cast<int>(ir));
  c->peer_handles[peer] = h;  // recorded first so that vmm_release frees it if the mapping fails
  CUmemAllocationProp prop = heap_prop(c->device);
  size_t gran = 0;
  TADNN_CHECK_CU(d.MemGetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  CUdeviceptr va = 0;
  if (int rc = map_rw(d, &va, c->alloc_bytes, gran, h, c->device)) return rc;
  c->peer_vas[peer] = reinterpret_cast<void*>(va);
  c->dev.heap[peer] = reinterpret_cast<uint8_t*>(va);
  c->dev.pad[peer] = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(va) + c->heap_bytes);
  return 0;
}

// Rank 0 creates the multicast object and exports it; everybody else imports it.
extern "C" int tadnn_comm_mc_create(tadnn_comm* c, int* fd) {
  TADNN_REQUIRE(c && c->vmm && fd, "comm_mc_create: not a VMM context");
  Drv& d = drv();
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = static_cast<unsigned>(c->world);
  mp.size = c->alloc_bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  TADNN_CHECK_CU(d.McCreate(&h, &mp));
  c->mc_handle = h;
  int f = -1;
  TADNN_CHECK_CU(d.MemExport(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd = f;
  return 0;
}

extern "C" int tadnn_comm_mc_import(tadnn_comm* c, int fd) {
  TADNN_REQUIRE(c && c->vmm, "comm_mc_import: not a VMM context");
  CUmemGenericAllocationHandle h;
  const CUresult ir = drv().MemImport(&h, reinterpret_cast<void*>(static_cast<intptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  close(fd);
  TADNN_REQUIRE(ir == CUDA_SUCCESS, "comm_mc_import: cuMemImportFromShareableHandle -> CUresult %d", static_cast<int>(ir));
  c->mc_handle = h;
  return 0;
}

// Step 1 (all ranks): add this device to the multicast team.  A host barrier must follow.
extern "C" int tadnn_comm_mc_add_device(tadnn_comm* c) {
  TADNN_REQUIRE(c && c->vmm && c->mc_handle, "comm_mc_add_device: no multicast object");
  Drv& d = drv();
  CUdevice cud;
  TADNN_CHECK_CU(d.DeviceGet(&cud, c->device));
  TADNN_CHECK_CU(d.McAddDevice(c->mc_handle, cud));
  return 0;
}

// Step 2 (all ranks, after every rank added its device): bind the heap and map the multicast VA.
extern "C" int tadnn_comm_mc_bind(tadnn_comm* c) {
  TADNN_REQUIRE(c && c->vmm && c->mc_handle, "comm_mc_bind: no multicast object");
  Drv& d = drv();
  TADNN_CHECK_CU(d.McBindMem(c->mc_handle, 0, c->mem_handle, 0, c->alloc_bytes, 0));
  CUmemAllocationProp prop = heap_prop(c->device);
  size_t gran = 0;
  TADNN_CHECK_CU(d.MemGetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  CUdeviceptr va = 0;
  if (int rc = map_rw(d, &va, c->alloc_bytes, gran, c->mc_handle, c->device)) return rc;
  c->mc_va = reinterpret_cast<void*>(va);
  c->dev.mc_heap = reinterpret_cast<uint8_t*>(va);
  return 0;
}

extern "C" int tadnn_comm_vmm_release(tadnn_comm* c) {
  if (c == nullptr || !c->vmm) return 0;
  Drv& d = drv();
  if (!d.ok) return 0;
  if (c->mc_va) { d.MemUnmap(reinterpret_cast<CUdeviceptr>(c->mc_va), c->alloc_bytes); d.MemAddressFree(reinterpret_cast<CUdeviceptr>(c->mc_va), c->alloc_bytes); }
  if (c->mc_handle) d.MemRelease(c->mc_handle);
  for (size_t r = 0; r < c->peer_vas.size(); ++r) {
    if (c->peer_vas[r]) { d.MemUnmap(reinterpret_cast<CUdeviceptr>(c->peer_vas[r]), c->alloc_bytes); d.MemAddressFree(reinterpret_cast<CUdeviceptr>(c->peer_vas[r]), c->alloc_bytes); }
    if (c->peer_handles[r]) d.MemRelease(c->peer_handles[r]);
  }
  if (c->base) { d.MemUnmap(reinterpret_cast<CUdeviceptr>(c->base), c->alloc_bytes); d.MemAddressFree(reinterpret_cast<CUdeviceptr>(c->base), c->alloc_bytes); }
  if (c->mem_handle) d.MemRelease(c->mem_handle);
  return 0;
}
