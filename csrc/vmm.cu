// Symmetric heap on the CUDA virtual-memory-management API + NVLS multicast binding.
//
// The reference brings up an NCCL communicator (dist.py:44-73) and lets NCCL own every buffer.
// Here each rank creates ONE physical allocation with cuMemCreate (POSIX file-descriptor handle
// type), hands the descriptor to every peer of the node over a Unix-domain socket (SCM_RIGHTS),
// maps all peers' allocations into its own address space, and -- when the driver and the fabric
// permit it -- binds the allocations of all ranks to one NVSwitch multicast object
// (cuMulticastCreate / AddDevice / BindMem).  A store to the multicast mapping lands in every
// GPU's copy; `multimem.ld_reduce` on it returns the sum over all copies, reduced inside the
// switch.  comm.cu and gemm_tcgen05.cu use both mappings from inside their kernels.
//
// cudaMalloc + cudaIpc (comm.cu: llmss_symm_alloc / llmss_ipc_*) remains as the fall-back when
// the VMM path is refused (no fd handle support, no permission to pass descriptors): memory
// from cudaMalloc cannot be bound to a multicast object, so that path is P2P-only.
//
// All driver entry points are resolved with cudaGetDriverEntryPoint: the library does not link
// libcuda, so it also loads (and reports "unsupported") on machines without a driver.
#include "common.cuh"

#include <cuda.h>
#include <errno.h>
#include <poll.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

namespace {

struct Driver {
  decltype(&cuMemCreate) memCreate = nullptr;
  decltype(&cuMemRelease) memRelease = nullptr;
  decltype(&cuMemAddressReserve) addrReserve = nullptr;
  decltype(&cuMemAddressFree) addrFree = nullptr;
  decltype(&cuMemMap) memMap = nullptr;
  decltype(&cuMemUnmap) memUnmap = nullptr;
  decltype(&cuMemSetAccess) setAccess = nullptr;
  decltype(&cuMemExportToShareableHandle) exportHandle = nullptr;
  decltype(&cuMemImportFromShareableHandle) importHandle = nullptr;
  decltype(&cuMemGetAllocationGranularity) memGranularity = nullptr;
  decltype(&cuMulticastCreate) mcCreate = nullptr;
  decltype(&cuMulticastAddDevice) mcAddDevice = nullptr;
  decltype(&cuMulticastBindMem) mcBindMem = nullptr;
  decltype(&cuMulticastUnbind) mcUnbind = nullptr;
  decltype(&cuMulticastGetGranularity) mcGranularity = nullptr;
  decltype(&cuDeviceGet) deviceGet = nullptr;
  decltype(&cuDeviceGetAttribute) deviceGetAttribute = nullptr;
  decltype(&cuGetErrorString) errorString = nullptr;
  bool ok = false;
};

template <typename F>
bool resolve(const char* name, F* fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
      p == nullptr) {
    cudaGetLastError();
    return false;
  }
  *fn = reinterpret_cast<F>(p);
  return true;
}

Driver& drv() {
  static Driver d;
  static bool tried = false;
  if (!tried) {
    tried = true;
    bool ok = true;
    ok &= resolve("cuMemCreate", &d.memCreate);
    ok &= resolve("cuMemRelease", &d.memRelease);
    ok &= resolve("cuMemAddressReserve", &d.addrReserve);
    ok &= resolve("cuMemAddressFree", &d.addrFree);
    ok &= resolve("cuMemMap", &d.memMap);
    ok &= resolve("cuMemUnmap", &d.memUnmap);
    ok &= resolve("cuMemSetAccess", &d.setAccess);
    ok &= resolve("cuMemExportToShareableHandle", &d.exportHandle);
    ok &= resolve("cuMemImportFromShareableHandle", &d.importHandle);
    ok &= resolve("cuMemGetAllocationGranularity", &d.memGranularity);
    ok &= resolve("cuDeviceGet", &d.deviceGet);
    ok &= resolve("cuDeviceGetAttribute", &d.deviceGetAttribute);
    ok &= resolve("cuGetErrorString", &d.errorString);
    d.ok = ok;
    // multicast entry points are optional (older drivers): their absence only disables NVLS
    resolve("cuMulticastCreate", &d.mcCreate);
    resolve("cuMulticastAddDevice", &d.mcAddDevice);
    resolve("cuMulticastBindMem", &d.mcBindMem);
    resolve("cuMulticastUnbind", &d.mcUnbind);
    resolve("cuMulticastGetGranularity", &d.mcGranularity);
  }
  return d;
}

const char* cu_err(CUresult r) {
  const char* s = nullptr;
  if (drv().errorString != nullptr && drv().errorString(r, &s) == CUDA_SUCCESS && s != nullptr) return s;
  return "unknown CUresult";
}

#define LLMSS_CHECK_CU(expr)                                                                             \
  do {                                                                                                   \
    CUresult _r = (expr);                                                                                \
    if (_r != CUDA_SUCCESS) {                                                                            \
      llmss_set_error("%s:%d: %s failed: %s (CUresult %d)", __FILE__, __LINE__, #expr, cu_err(_r), (int)_r); \
      return (int)_r;                                                                                    \
    }                                                                                                    \
  } while (0)

int current_device(CUdevice* dev, int* ordinal) {
  int o = 0;
  LLMSS_CHECK_CUDA(cudaGetDevice(&o));
  LLMSS_CHECK_CUDA(cudaFree(nullptr));  // make sure the primary context exists and is current
  LLMSS_CHECK_CU(drv().deviceGet(dev, o));
  if (ordinal) *ordinal = o;
  return 0;
}

CUmemAllocationProp alloc_prop(int ordinal) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = ordinal;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

// Reserve an address range, map `h` into it and grant the current device read/write access.
int map_handle(CUmemGenericAllocationHandle h, size_t bytes, size_t align, int ordinal, CUdeviceptr* out) {
  CUdeviceptr va = 0;
  LLMSS_CHECK_CU(drv().addrReserve(&va, bytes, align, 0, 0));
  CUresult r = drv().memMap(va, bytes, 0, h, 0);
  if (r != CUDA_SUCCESS) {
    drv().addrFree(va, bytes);
    llmss_set_error("vmm: cuMemMap of %zu bytes failed: %s (CUresult %d)", bytes, cu_err(r), (int)r);
    return (int)r;
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = ordinal;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = drv().setAccess(va, bytes, &acc, 1);
  if (r != CUDA_SUCCESS) {
    drv().memUnmap(va, bytes);
    drv().addrFree(va, bytes);
    llmss_set_error("vmm: cuMemSetAccess failed: %s (CUresult %d)", cu_err(r), (int)r);
    return (int)r;
  }
  *out = va;
  return 0;
}

}  // namespace

// One rank's share of the symmetric heap.
struct llmss_vmm_region {
  CUmemGenericAllocationHandle handle;
  CUdeviceptr va;
  size_t bytes;  // padded to the allocation (and, when multicast is wanted, the multicast) granularity
  int ordinal;
  // multicast (NVLS) state: 0 / null when not bound
  CUmemGenericAllocationHandle mc_handle;
  CUdeviceptr mc_va;
  int mc_bound;
};

// caps bit 0: VMM allocations with POSIX-fd handles; bit 1: multicast objects (NVLS).
LLMSS_API int llmss_vmm_caps(int* caps) {
  *caps = 0;
  if (!drv().ok) return 0;
  CUdevice dev;
  int ordinal = 0;
  if (int e = current_device(&dev, &ordinal)) return e;
  int v = 0;
  if (drv().deviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev) == CUDA_SUCCESS && v)
    *caps |= 1;
  v = 0;
  if (drv().mcCreate != nullptr && drv().mcBindMem != nullptr &&
      drv().deviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) == CUDA_SUCCESS && v)
    *caps |= 2;
  return 0;
}

// Size a region must be padded to so that it can be created, mapped and (nranks > 1, want_mc) bound to a multicast
// object of `nranks` devices.
LLMSS_API int llmss_vmm_padded_size(size_t bytes, int nranks, int want_mc, size_t* padded) {
  LLMSS_REQUIRE(drv().ok, "vmm: CUDA driver VMM entry points are not available");
  CUdevice dev;
  int ordinal = 0;
  if (int e = current_device(&dev, &ordinal)) return e;
  CUmemAllocationProp prop = alloc_prop(ordinal);
  size_t gran = 0;
  LLMSS_CHECK_CU(drv().memGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (want_mc && nranks > 1 && drv().mcGranularity != nullptr) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)nranks;
    mp.size = bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (drv().mcGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran) gran = mg;
  }
  if (gran == 0) gran = 2u << 20;
  *padded = (bytes + gran - 1) / gran * gran;
  return 0;
}

// Create + map this rank's region (`padded` from llmss_vmm_padded_size; identical on every rank).
LLMSS_API int llmss_vmm_alloc(size_t padded, llmss_vmm_region** out) {
  LLMSS_REQUIRE(drv().ok, "vmm: CUDA driver VMM entry points are not available");
  CUdevice dev;
  int ordinal = 0;
  if (int e = current_device(&dev, &ordinal)) return e;
  CUmemAllocationProp prop = alloc_prop(ordinal);
  llmss_vmm_region* r = new llmss_vmm_region();
  memset(r, 0, sizeof(*r));
  r->bytes = padded;
  r->ordinal = ordinal;
  CUresult cr = drv().memCreate(&r->handle, padded, &prop, 0);
  if (cr != CUDA_SUCCESS) {
    delete r;
    llmss_set_error("vmm: cuMemCreate(%zu bytes, posix-fd handle) failed: %s (CUresult %d)", padded, cu_err(cr), (int)cr);
    return (int)cr;
  }
  if (int e = map_handle(r->handle, padded, 0, ordinal, &r->va)) {
    drv().memRelease(r->handle);
    delete r;
    return e;
  }
  *out = r;
  return 0;
}
LLMSS_API void* llmss_vmm_ptr(llmss_vmm_region* r) { return reinterpret_cast<void*>(r->va); }
LLMSS_API void* llmss_vmm_mc_ptr(llmss_vmm_region* r) { return r->mc_bound ? reinterpret_cast<void*>(r->mc_va) : nullptr; }
LLMSS_API size_t llmss_vmm_bytes(llmss_vmm_region* r) { return r->bytes; }

// A file descriptor that names the allocation; the caller sends it to the peers and closes it.
LLMSS_API int llmss_vmm_export_fd(llmss_vmm_region* r, int* fd) {
  int f = -1;
  LLMSS_CHECK_CU(drv().exportHandle(&f, r->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd = f;
  return 0;
}

// Map a peer's region (descriptor received from that peer) into this process; the descriptor is closed.
LLMSS_API int llmss_vmm_import(int fd, size_t padded, void** ptr) {
  CUdevice dev;
  int ordinal = 0;
  if (int e = current_device(&dev, &ordinal)) return e;
  CUmemGenericAllocationHandle h;
  CUresult cr = drv().importHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  close(fd);
  if (cr != CUDA_SUCCESS) {
    llmss_set_error("vmm: cuMemImportFromShareableHandle failed: %s (CUresult %d)", cu_err(cr), (int)cr);
    return (int)cr;
  }
  CUdeviceptr va = 0;
  int e = map_handle(h, padded, 0, ordinal, &va);
  drv().memRelease(h);  // the mapping keeps the allocation alive
  if (e) return e;
  *ptr = reinterpret_cast<void*>(va);
  return 0;
}
LLMSS_API int llmss_vmm_unmap(void* ptr, size_t padded) {
  if (ptr == nullptr) return 0;
  LLMSS_CHECK_CU(drv().memUnmap((CUdeviceptr)ptr, padded));
  LLMSS_CHECK_CU(drv().addrFree((CUdeviceptr)ptr, padded));
  return 0;
}
LLMSS_API int llmss_vmm_free(llmss_vmm_region* r) {
  if (r == nullptr) return 0;
  if (r->mc_bound) {
    drv().memUnmap(r->mc_va, r->bytes);
    drv().addrFree(r->mc_va, r->bytes);
    CUdevice dev;
    if (drv().mcUnbind != nullptr && drv().deviceGet(&dev, r->ordinal) == CUDA_SUCCESS) drv().mcUnbind(r->mc_handle, dev, 0, r->bytes);
    drv().memRelease(r->mc_handle);
  }
  drv().memUnmap(r->va, r->bytes);
  drv().addrFree(r->va, r->bytes);
  drv().memRelease(r->handle);
  delete r;
  return 0;
}

// ---------------------------------------------------------------- NVLS multicast object
// Sequence (every step is collective over the ranks; the caller places a barrier between them):
//   rank 0: llmss_mc_create -> fd to every peer;  peers: llmss_mc_import
//   all:    llmss_mc_add_device            | barrier: every device added before anything is bound
//   all:    llmss_mc_bind_and_map          | barrier: every rank bound before the first multimem access
LLMSS_API int llmss_mc_create(int nranks, size_t padded, unsigned long long* mc, int* fd) {
  LLMSS_REQUIRE(drv().mcCreate != nullptr, "vmm: this driver has no multicast entry points");
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)nranks;
  mp.size = padded;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  LLMSS_CHECK_CU(drv().mcCreate(&h, &mp));
  int f = -1;
  CUresult cr = drv().exportHandle(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (cr != CUDA_SUCCESS) {
    drv().memRelease(h);
    llmss_set_error("vmm: exporting the multicast object failed: %s (CUresult %d)", cu_err(cr), (int)cr);
    return (int)cr;
  }
  *mc = (unsigned long long)h;
  *fd = f;
  return 0;
}
LLMSS_API int llmss_mc_import(int fd, unsigned long long* mc) {
  CUmemGenericAllocationHandle h;
  CUresult cr = drv().importHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  close(fd);
  if (cr != CUDA_SUCCESS) {
    llmss_set_error("vmm: importing the multicast object failed: %s (CUresult %d)", cu_err(cr), (int)cr);
    return (int)cr;
  }
  *mc = (unsigned long long)h;
  return 0;
}
LLMSS_API int llmss_mc_add_device(unsigned long long mc) {
  CUdevice dev;
  if (int e = current_device(&dev, nullptr)) return e;
  LLMSS_CHECK_CU(drv().mcAddDevice((CUmemGenericAllocationHandle)mc, dev));
  return 0;
}
LLMSS_API int llmss_mc_bind_and_map(unsigned long long mc, llmss_vmm_region* r) {
  LLMSS_CHECK_CU(drv().mcBindMem((CUmemGenericAllocationHandle)mc, 0, r->handle, 0, r->bytes, 0));
  CUdeviceptr va = 0;
  if (int e = map_handle((CUmemGenericAllocationHandle)mc, r->bytes, 0, r->ordinal, &va)) return e;
  r->mc_handle = (CUmemGenericAllocationHandle)mc;
  r->mc_va = va;
  r->mc_bound = 1;
  return 0;
}
LLMSS_API int llmss_mc_release(unsigned long long mc) {  // a rank that gives up before binding
  if (mc != 0 && drv().memRelease != nullptr) drv().memRelease((CUmemGenericAllocationHandle)mc);
  return 0;
}

// ---------------------------------------------------------------- descriptor exchange (SCM_RIGHTS)
// Every rank binds one datagram socket in the abstract namespace (no file-system entry, vanishes with the
// process) under a job-unique name and sends `(tag, fd)` messages to its peers' names.
namespace {
int make_addr(const char* name, sockaddr_un* addr, socklen_t* len) {
  const size_t n = strlen(name);
  LLMSS_REQUIRE(n > 0 && n + 1 < sizeof(addr->sun_path), "fdx: socket name '%s' is empty or too long", name);
  memset(addr, 0, sizeof(*addr));
  addr->sun_family = AF_UNIX;
  addr->sun_path[0] = '\0';
  memcpy(addr->sun_path + 1, name, n);
  *len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
  return 0;
}
}  // namespace

LLMSS_API int llmss_fdx_open(const char* name, int* sock) {
  sockaddr_un addr;
  socklen_t len;
  if (int e = make_addr(name, &addr, &len)) return e;
  const int s = socket(AF_UNIX, SOCK_DGRAM | SOCK_CLOEXEC, 0);
  LLMSS_REQUIRE(s >= 0, "fdx: socket() failed: %s", strerror(errno));
  if (bind(s, reinterpret_cast<sockaddr*>(&addr), len) != 0) {
    llmss_set_error("fdx: bind('%s') failed: %s", name, strerror(errno));
    close(s);
    return -1;
  }
  *sock = s;
  return 0;
}
LLMSS_API int llmss_fdx_send(int sock, const char* peer_name, int tag, int fd) {
  sockaddr_un addr;
  socklen_t len;
  if (int e = make_addr(peer_name, &addr, &len)) return e;
  int payload = tag;
  iovec iov = {&payload, sizeof(payload)};
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  msg.msg_name = &addr;
  msg.msg_namelen = len;
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  for (int attempt = 0; attempt < 200; ++attempt) {  // the peer's queue may be full for a moment
    if (sendmsg(sock, &msg, 0) >= 0) return 0;
    if (errno != EAGAIN && errno != ENOBUFS && errno != EINTR && errno != ECONNREFUSED) break;
    usleep(10000);
  }
  llmss_set_error("fdx: sendmsg to '%s' failed: %s", peer_name, strerror(errno));
  return -1;
}
LLMSS_API int llmss_fdx_recv(int sock, int timeout_ms, int* tag, int* fd) {
  pollfd p = {sock, POLLIN, 0};
  int pr;
  do {
    pr = poll(&p, 1, timeout_ms);
  } while (pr < 0 && errno == EINTR);
  LLMSS_REQUIRE(pr > 0, "fdx: no descriptor arrived within %d ms", timeout_ms);
  int payload = 0;
  iovec iov = {&payload, sizeof(payload)};
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  const ssize_t n = recvmsg(sock, &msg, MSG_CMSG_CLOEXEC);
  LLMSS_REQUIRE(n == (ssize_t)sizeof(payload), "fdx: recvmsg failed: %s", n < 0 ? strerror(errno) : "short message");
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  LLMSS_REQUIRE(c != nullptr && c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS,
                "fdx: message carried no descriptor (was it truncated?)");
  int f = -1;
  memcpy(&f, CMSG_DATA(c), sizeof(int));
  *tag = payload;
  *fd = f;
  return 0;
}
LLMSS_API int llmss_fdx_close(int sock) {
  if (sock >= 0) close(sock);
  return 0;
}
LLMSS_API int llmss_fd_close(int fd) {
  if (fd >= 0) close(fd);
  return 0;
}
