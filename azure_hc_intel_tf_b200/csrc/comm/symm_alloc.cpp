// symm_alloc.cpp — symmetric device memory on the CUDA virtual-memory-management API, with an NVLS multicast mapping.
//
// Role in the reference stack: communicator / transport set-up — MPI_Init and the UCX (`rc_x,sm,self`) or libfabric
// endpoints that `mpirun -mca pml ucx ...` / `mpiexec.hydra -genv I_MPI_FABRICS shm:ofi` bring up
// (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:85-95, run-tf-sing-libfabric-intelmpi.sh:86-101). Here "bringing the
// transport up" means: every rank creates one physical allocation (cuMemCreate), exports it as a POSIX file
// descriptor, passes that descriptor to every peer over a Unix-domain socket (SCM_RIGHTS), and maps every peer's
// allocation into its own address space; rank 0 additionally creates a multicast object spanning all GPUs of the job,
// every rank binds its allocation to it, and the multicast address becomes the operand of multimem.ld_reduce /
// multimem.st in csrc/comm/allreduce.cu. Nothing here depends on torch.
#include <cuda.h>
#include <cuda_runtime.h>
#include <errno.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <cstdio>
#include <vector>

#include "b200comm.h"
#include "common.cuh"

namespace {

// The driver API is reached through cudaGetDriverEntryPoint, so the library has no link-time dependency on libcuda and
// still loads (for its host-side planning functions and the build check) on a machine without a driver.
struct Drv {
#define B200_DRV_FN(name) decltype(&::name) name = nullptr;
  B200_DRV_FN(cuDeviceGet)
  B200_DRV_FN(cuDeviceGetAttribute)
  B200_DRV_FN(cuGetErrorString)
  B200_DRV_FN(cuMemAddressFree)
  B200_DRV_FN(cuMemAddressReserve)
  B200_DRV_FN(cuMemCreate)
  B200_DRV_FN(cuMemExportToShareableHandle)
  B200_DRV_FN(cuMemGetAllocationGranularity)
  B200_DRV_FN(cuMemImportFromShareableHandle)
  B200_DRV_FN(cuMemMap)
  B200_DRV_FN(cuMemRelease)
  B200_DRV_FN(cuMemSetAccess)
  B200_DRV_FN(cuMemUnmap)
  B200_DRV_FN(cuMulticastAddDevice)
  B200_DRV_FN(cuMulticastBindMem)
  B200_DRV_FN(cuMulticastCreate)
  B200_DRV_FN(cuMulticastGetGranularity)
  B200_DRV_FN(cuMulticastUnbind)
#undef B200_DRV_FN
  bool ok = false;
};

Drv load_driver() {
  Drv d;
  bool ok = true;
  auto get = [&ok](const char* sym) -> void* {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(sym, &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
      ok = false;
      return nullptr;
    }
    return fn;
  };
  d.cuDeviceGet = reinterpret_cast<decltype(d.cuDeviceGet)>(get("cuDeviceGet"));
  d.cuDeviceGetAttribute = reinterpret_cast<decltype(d.cuDeviceGetAttribute)>(get("cuDeviceGetAttribute"));
  d.cuGetErrorString = reinterpret_cast<decltype(d.cuGetErrorString)>(get("cuGetErrorString"));
  d.cuMemAddressFree = reinterpret_cast<decltype(d.cuMemAddressFree)>(get("cuMemAddressFree"));
  d.cuMemAddressReserve = reinterpret_cast<decltype(d.cuMemAddressReserve)>(get("cuMemAddressReserve"));
  d.cuMemCreate = reinterpret_cast<decltype(d.cuMemCreate)>(get("cuMemCreate"));
  d.cuMemExportToShareableHandle = reinterpret_cast<decltype(d.cuMemExportToShareableHandle)>(get("cuMemExportToShareableHandle"));
  d.cuMemGetAllocationGranularity = reinterpret_cast<decltype(d.cuMemGetAllocationGranularity)>(get("cuMemGetAllocationGranularity"));
  d.cuMemImportFromShareableHandle = reinterpret_cast<decltype(d.cuMemImportFromShareableHandle)>(get("cuMemImportFromShareableHandle"));
  d.cuMemMap = reinterpret_cast<decltype(d.cuMemMap)>(get("cuMemMap"));
  d.cuMemRelease = reinterpret_cast<decltype(d.cuMemRelease)>(get("cuMemRelease"));
  d.cuMemSetAccess = reinterpret_cast<decltype(d.cuMemSetAccess)>(get("cuMemSetAccess"));
  d.cuMemUnmap = reinterpret_cast<decltype(d.cuMemUnmap)>(get("cuMemUnmap"));
  d.cuMulticastAddDevice = reinterpret_cast<decltype(d.cuMulticastAddDevice)>(get("cuMulticastAddDevice"));
  d.cuMulticastBindMem = reinterpret_cast<decltype(d.cuMulticastBindMem)>(get("cuMulticastBindMem"));
  d.cuMulticastCreate = reinterpret_cast<decltype(d.cuMulticastCreate)>(get("cuMulticastCreate"));
  d.cuMulticastGetGranularity = reinterpret_cast<decltype(d.cuMulticastGetGranularity)>(get("cuMulticastGetGranularity"));
  d.cuMulticastUnbind = reinterpret_cast<decltype(d.cuMulticastUnbind)>(get("cuMulticastUnbind"));
  d.ok = ok;
  return d;
}

Drv& drv() {
  static Drv d = load_driver();
  return d;
}

#define B200_NEED_DRIVER()                                                                    \
  do {                                                                                        \
    if (!drv().ok) {                                                                          \
      ::b200::set_error("the CUDA driver API is not available (no driver on this machine?)"); \
      return -2;                                                                              \
    }                                                                                         \
  } while (0)

#define B200_CU(expr)                                                                         \
  do {                                                                                        \
    CUresult _r = (expr);                                                                     \
    if (_r != CUDA_SUCCESS) {                                                                 \
      const char* _s = nullptr;                                                               \
      drv().cuGetErrorString(_r, &_s);                                                              \
      ::b200::set_error("%s failed: %s (%s:%d)", #expr, _s ? _s : "?", __FILE__, __LINE__);   \
      return -2;                                                                              \
    }                                                                                         \
  } while (0)

struct VmmBuf {
  int device = 0, world = 0, rank = 0;
  size_t bytes = 0;                       // padded to the allocation (and multicast) granularity
  CUmemGenericAllocationHandle local = 0;
  std::vector<CUmemGenericAllocationHandle> peer_handles;   // imported (own rank: the local handle)
  std::vector<CUdeviceptr> ptrs;                            // mapped address of every rank's allocation
  CUmemGenericAllocationHandle mc = 0;
  CUdeviceptr mc_ptr = 0;
  bool mc_bound = false;
};

CUmemAccessDesc access_for(int device) {
  CUmemAccessDesc a{};
  a.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  a.location.id = device;
  a.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  return a;
}

int map_handle(CUmemGenericAllocationHandle h, size_t bytes, int device, CUdeviceptr* out) {
  CUdeviceptr va = 0;
  B200_CU(drv().cuMemAddressReserve(&va, bytes, 0, 0, 0));
  B200_CU(drv().cuMemMap(va, bytes, 0, h, 0));
  CUmemAccessDesc a = access_for(device);
  B200_CU(drv().cuMemSetAccess(va, bytes, &a, 1));
  *out = va;
  return 0;
}

sockaddr_un abstract_addr(const char* name, socklen_t* len) {
  sockaddr_un a{};
  a.sun_family = AF_UNIX;
  const size_t n = strnlen(name, sizeof(a.sun_path) - 2);
  a.sun_path[0] = '\0';                    // abstract namespace: no file system entry, gone with the process
  memcpy(a.sun_path + 1, name, n);
  *len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
  return a;
}

}  // namespace

extern "C" {

int b200_vmm_supported(int32_t device, int32_t* multicast_ok) {
  B200_NEED_DRIVER();
  int vmm = 0, fd = 0, mc = 0;
  CUdevice dev;
  B200_CUDA(cudaFree(nullptr));            // make sure the primary context exists
  B200_CU(drv().cuDeviceGet(&dev, device));
  B200_CU(drv().cuDeviceGetAttribute(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev));
  B200_CU(drv().cuDeviceGetAttribute(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev));
  B200_CU(drv().cuDeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev));
  if (multicast_ok) *multicast_ok = mc;
  return (vmm && fd) ? 1 : 0;
}

/* ---- file-descriptor exchange over abstract Unix-domain datagram sockets ------------------------------------ */

int b200_fd_listen(const char* name) {
  int s = socket(AF_UNIX, SOCK_DGRAM, 0);
  if (s < 0) {
    ::b200::set_error("fd_listen: socket: %s", strerror(errno));
    return -1;
  }
  socklen_t len;
  sockaddr_un a = abstract_addr(name, &len);
  if (bind(s, reinterpret_cast<sockaddr*>(&a), len) != 0) {
    ::b200::set_error("fd_listen: bind(%s): %s", name, strerror(errno));
    close(s);
    return -1;
  }
  return s;
}

int b200_fd_send(int sock, const char* peer_name, int fd, int32_t tag) {
  socklen_t len;
  sockaddr_un a = abstract_addr(peer_name, &len);
  msghdr msg{};
  iovec iov{&tag, sizeof tag};
  alignas(cmsghdr) char ctl[CMSG_SPACE(sizeof(int))] = {};
  msg.msg_name = &a;
  msg.msg_namelen = len;
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctl;
  msg.msg_controllen = sizeof ctl;
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  for (int tries = 0; tries < 2000; ++tries) {          // the peer may not have bound its socket yet
    if (sendmsg(sock, &msg, 0) >= 0) return 0;
    if (errno != ECONNREFUSED && errno != ENOENT && errno != EAGAIN) break;
    usleep(5000);
  }
  ::b200::set_error("fd_send to %s: %s", peer_name, strerror(errno));
  return -1;
}

int b200_fd_recv(int sock, int32_t* tag_out) {
  msghdr msg{};
  int32_t tag = -1;
  iovec iov{&tag, sizeof tag};
  alignas(cmsghdr) char ctl[CMSG_SPACE(sizeof(int))] = {};
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctl;
  msg.msg_controllen = sizeof ctl;
  timeval tv{120, 0};
  setsockopt(sock, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);
  if (recvmsg(sock, &msg, 0) < 0) {
    ::b200::set_error("fd_recv: %s", strerror(errno));
    return -1;
  }
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (!c || c->cmsg_type != SCM_RIGHTS) {
    ::b200::set_error("fd_recv: message carried no descriptor");
    return -1;
  }
  int fd = -1;
  memcpy(&fd, CMSG_DATA(c), sizeof(int));
  if (tag_out) *tag_out = tag;
  return fd;
}

int b200_fd_close(int fd) { return fd >= 0 ? close(fd) : 0; }

/* ---- the allocation ------------------------------------------------------------------------------------------- */

int b200_vmm_create(int64_t bytes, int32_t world, int32_t rank, int32_t want_multicast, void** handle_out, int32_t* fd_out,
                    int64_t* padded_bytes) {
  B200_NEED_DRIVER();
  B200_CHECK_ARG(bytes > 0 && world >= 1 && world <= B200_COMM_MAX_RANKS && rank >= 0 && rank < world && handle_out && fd_out,
                 "vmm_create: bad arguments");
  int device = 0;
  B200_CUDA(cudaGetDevice(&device));
  B200_CUDA(cudaFree(nullptr));
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  B200_CU(drv().cuMemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (want_multicast) {
    CUmulticastObjectProp mp{};
    mp.numDevices = (unsigned)world;
    mp.size = (size_t)bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    B200_CU(drv().cuMulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    if (mg > gran) gran = mg;
  }
  auto* b = new VmmBuf();
  b->device = device;
  b->world = world;
  b->rank = rank;
  b->bytes = ((size_t)bytes + gran - 1) / gran * gran;
  b->peer_handles.assign(world, 0);
  b->ptrs.assign(world, 0);
  CUresult r = drv().cuMemCreate(&b->local, b->bytes, &prop, 0);
  if (r != CUDA_SUCCESS) {
    delete b;
    const char* s = nullptr;
    drv().cuGetErrorString(r, &s);
    ::b200::set_error("vmm_create: drv().cuMemCreate(%zu bytes): %s", (size_t)bytes, s ? s : "?");
    return -2;
  }
  b->peer_handles[rank] = b->local;
  if (int e = map_handle(b->local, b->bytes, device, &b->ptrs[rank])) return e;
  B200_CUDA(cudaMemset(reinterpret_cast<void*>(b->ptrs[rank]), 0, b->bytes));
  int fd = -1;
  B200_CU(drv().cuMemExportToShareableHandle(&fd, b->local, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd_out = fd;
  *handle_out = b;
  if (padded_bytes) *padded_bytes = (int64_t)b->bytes;
  return 0;
}

int b200_vmm_map_peer(void* handle, int32_t peer, int32_t fd) {
  B200_NEED_DRIVER();
  auto* b = static_cast<VmmBuf*>(handle);
  B200_CHECK_ARG(b && peer >= 0 && peer < b->world && peer != b->rank && fd >= 0, "vmm_map_peer: bad arguments");
  B200_CU(drv().cuMemImportFromShareableHandle(&b->peer_handles[peer], reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                         CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  return map_handle(b->peer_handles[peer], b->bytes, b->device, &b->ptrs[peer]);
}

uint64_t b200_vmm_ptr(void* handle, int32_t rank) {
  auto* b = static_cast<VmmBuf*>(handle);
  return (b && rank >= 0 && rank < b->world) ? (uint64_t)b->ptrs[rank] : 0;
}

/* Multicast, three collective steps (a barrier between each): rank 0 creates the object and exports a descriptor that
 * the others import; EVERY rank adds its device; only then may anyone bind memory; then the object is mapped. */
int b200_vmm_mc_create(void* handle, int32_t* fd_out) {
  B200_NEED_DRIVER();
  auto* b = static_cast<VmmBuf*>(handle);
  B200_CHECK_ARG(b && fd_out, "vmm_mc_create: bad arguments");
  CUmulticastObjectProp mp{};
  mp.numDevices = (unsigned)b->world;
  mp.size = b->bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  B200_CU(drv().cuMulticastCreate(&b->mc, &mp));
  int fd = -1;
  B200_CU(drv().cuMemExportToShareableHandle(&fd, b->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd_out = fd;
  return 0;
}

int b200_vmm_mc_import(void* handle, int32_t fd) {
  B200_NEED_DRIVER();
  auto* b = static_cast<VmmBuf*>(handle);
  B200_CHECK_ARG(b && fd >= 0, "vmm_mc_import: bad arguments");
  B200_CU(drv().cuMemImportFromShareableHandle(&b->mc, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                         CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  return 0;
}

int b200_vmm_mc_add_device(void* handle) {
  B200_NEED_DRIVER();
  auto* b = static_cast<VmmBuf*>(handle);
  B200_CHECK_ARG(b && b->mc, "vmm_mc_add_device: no multicast object");
  CUdevice dev;
  B200_CU(drv().cuDeviceGet(&dev, b->device));
  B200_CU(drv().cuMulticastAddDevice(b->mc, dev));
  return 0;
}

int b200_vmm_mc_bind(void* handle) {
  B200_NEED_DRIVER();
  auto* b = static_cast<VmmBuf*>(handle);
  B200_CHECK_ARG(b && b->mc, "vmm_mc_bind: no multicast object");
  B200_CU(drv().cuMulticastBindMem(b->mc, 0, b->local, 0, b->bytes, 0));
  b->mc_bound = true;
  return map_handle(b->mc, b->bytes, b->device, &b->mc_ptr);
}

uint64_t b200_vmm_mc_ptr(void* handle) {
  auto* b = static_cast<VmmBuf*>(handle);
  return b ? (uint64_t)b->mc_ptr : 0;
}

int b200_vmm_free(void* handle) {
  auto* b = static_cast<VmmBuf*>(handle);
  if (!b) return 0;
  if (!drv().ok) return -2;
  cudaDeviceSynchronize();
  if (b->mc_ptr) {
    drv().cuMemUnmap(b->mc_ptr, b->bytes);
    drv().cuMemAddressFree(b->mc_ptr, b->bytes);
  }
  if (b->mc) {
    if (b->mc_bound) {
      CUdevice dev;
      if (drv().cuDeviceGet(&dev, b->device) == CUDA_SUCCESS) drv().cuMulticastUnbind(b->mc, dev, 0, b->bytes);
    }
    drv().cuMemRelease(b->mc);
  }
  for (int r = 0; r < b->world; ++r) {
    if (b->ptrs[r]) {
      drv().cuMemUnmap(b->ptrs[r], b->bytes);
      drv().cuMemAddressFree(b->ptrs[r], b->bytes);
    }
    if (b->peer_handles[r]) drv().cuMemRelease(b->peer_handles[r]);
  }
  delete b;
  return 0;
}

}  // extern "C"
