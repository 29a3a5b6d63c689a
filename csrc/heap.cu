// Symmetric heap plumbing — SURVEY W3 / §5.1.
// Replaces libipc's socket transport for tensors (and the CUDA-IPC hint at
// examples/mnist.lua:93): each rank creates one physical allocation with the CUDA virtual
// memory management API, exports it as a POSIX file descriptor, passes the fd to the other
// local ranks over a unix-domain socket (SCM_RIGHTS), and maps every peer's allocation into its
// own address space.  When the GPUs sit behind an NVSwitch with multicast support the same
// physical memory is also bound to a multicast object (NVLS).  cudaIpc* is the fallback when
// the VMM path is unavailable.
//
// The driver API is reached through cudaGetDriverEntryPoint so the library has no link-time
// dependency on libcuda.so (the build machine has none).
#include <cuda.h>
#include <cuda_runtime.h>
#include <errno.h>
#include <poll.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

#include "dl_host.h"

namespace {

template <typename Fn>
Fn drv(const char* name) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
    return nullptr;
  return reinterpret_cast<Fn>(fn);
}

const char* drv_err(CUresult r) {
  typedef CUresult (*F)(CUresult, const char**);
  static F f = drv<F>("cuGetErrorString");
  const char* s = nullptr;
  if (f && f(r, &s) == CUDA_SUCCESS && s) return s;
  return "unknown driver error";
}

#define DL_DRV_FN(var, type, name)                       \
  static type var = drv<type>(name);                     \
  if (!var) {                                            \
    dl_set_error("driver entry point %s not found", name); \
    return DL_EUNSUPPORTED;                              \
  }

#define DL_DRV(expr)                                                                     \
  do {                                                                                   \
    CUresult _r = (expr);                                                                \
    if (_r != CUDA_SUCCESS) {                                                            \
      dl_set_error("%s:%d %s -> %s (%d)", __FILE__, __LINE__, #expr, drv_err(_r), (int)_r); \
      return DL_ECUDA;                                                                   \
    }                                                                                    \
  } while (0)

int ensure_ctx(int dev) {
  DL_CUDA(cudaSetDevice(dev));
  DL_CUDA(cudaFree(0));
  return DL_OK;
}

CUmemAllocationProp alloc_prop(int dev) {
  CUmemAllocationProp p;
  memset(&p, 0, sizeof(p));
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = dev;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

typedef CUresult (*cuDeviceGetAttribute_t)(int*, CUdevice_attribute, CUdevice);
typedef CUresult (*cuMemGetAllocationGranularity_t)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
typedef CUresult (*cuMemCreate_t)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
typedef CUresult (*cuMemExportToShareableHandle_t)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
typedef CUresult (*cuMemImportFromShareableHandle_t)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
typedef CUresult (*cuMemAddressReserve_t)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
typedef CUresult (*cuMemMap_t)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
typedef CUresult (*cuMemSetAccess_t)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
typedef CUresult (*cuMemUnmap_t)(CUdeviceptr, size_t);
typedef CUresult (*cuMemAddressFree_t)(CUdeviceptr, size_t);
typedef CUresult (*cuMemRelease_t)(CUmemGenericAllocationHandle);
typedef CUresult (*cuMulticastCreate_t)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
typedef CUresult (*cuMulticastAddDevice_t)(CUmemGenericAllocationHandle, CUdevice);
typedef CUresult (*cuMulticastBindMem_t)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
typedef CUresult (*cuMulticastUnbind_t)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
typedef CUresult (*cuMulticastGetGranularity_t)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);

CUmulticastObjectProp mc_prop(int ndev, size_t bytes) {
  CUmulticastObjectProp p;
  memset(&p, 0, sizeof(p));
  p.numDevices = static_cast<unsigned int>(ndev);
  p.size = bytes;
  p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  p.flags = 0;
  return p;
}

long long now_ms() {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return static_cast<long long>(ts.tv_sec) * 1000 + ts.tv_nsec / 1000000;
}

// abstract-namespace unix socket address: no file to clean up, scoped to the network namespace
socklen_t make_addr(const char* name, struct sockaddr_un* addr) {
  memset(addr, 0, sizeof(*addr));
  addr->sun_family = AF_UNIX;
  size_t len = strlen(name);
  if (len > sizeof(addr->sun_path) - 2) len = sizeof(addr->sun_path) - 2;
  addr->sun_path[0] = '\0';
  memcpy(addr->sun_path + 1, name, len);
  return static_cast<socklen_t>(offsetof(struct sockaddr_un, sun_path) + 1 + len);
}

}  // namespace

extern "C" int dl_device_caps(int dev, int* vmm, int* fd_handles, int* multicast) {
  int rc = ensure_ctx(dev);
  if (rc) return rc;
  DL_DRV_FN(getattr, cuDeviceGetAttribute_t, "cuDeviceGetAttribute");
  int v = 0, f = 0, m = 0;
  DL_DRV(getattr(&v, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev));
  DL_DRV(getattr(&f, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev));
  if (getattr(&m, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) m = 0;
  if (vmm) *vmm = v;
  if (fd_handles) *fd_handles = f;
  if (multicast) *multicast = m;
  return DL_OK;
}

extern "C" int dl_vmm_granularity(int dev, size_t* gran) {
  int rc = ensure_ctx(dev);
  if (rc) return rc;
  DL_DRV_FN(fn, cuMemGetAllocationGranularity_t, "cuMemGetAllocationGranularity");
  CUmemAllocationProp p = alloc_prop(dev);
  DL_DRV(fn(gran, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  return DL_OK;
}

extern "C" int dl_vmm_alloc(int dev, size_t bytes, uint64_t* handle, int* fd) {
  int rc = ensure_ctx(dev);
  if (rc) return rc;
  DL_DRV_FN(create, cuMemCreate_t, "cuMemCreate");
  DL_DRV_FN(exp, cuMemExportToShareableHandle_t, "cuMemExportToShareableHandle");
  CUmemAllocationProp p = alloc_prop(dev);
  CUmemGenericAllocationHandle h;
  DL_DRV(create(&h, bytes, &p, 0));
  int out_fd = -1;
  if (fd) {
    CUresult r = exp(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
      dl_set_error("cuMemExportToShareableHandle -> %s (%d)", drv_err(r), (int)r);
      static cuMemRelease_t rel = drv<cuMemRelease_t>("cuMemRelease");
      if (rel) rel(h);
      return DL_ECUDA;
    }
    *fd = out_fd;
  }
  *handle = static_cast<uint64_t>(h);
  return DL_OK;
}

extern "C" int dl_vmm_import(int fd, uint64_t* handle) {
  DL_DRV_FN(imp, cuMemImportFromShareableHandle_t, "cuMemImportFromShareableHandle");
  CUmemGenericAllocationHandle h;
  DL_DRV(imp(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *handle = static_cast<uint64_t>(h);
  return DL_OK;
}

extern "C" int dl_vmm_map(int dev, uint64_t handle, size_t bytes, void** ptr) {
  int rc = ensure_ctx(dev);
  if (rc) return rc;
  DL_DRV_FN(reserve, cuMemAddressReserve_t, "cuMemAddressReserve");
  DL_DRV_FN(map, cuMemMap_t, "cuMemMap");
  DL_DRV_FN(access, cuMemSetAccess_t, "cuMemSetAccess");
  DL_DRV_FN(afree, cuMemAddressFree_t, "cuMemAddressFree");
  DL_DRV_FN(unmap, cuMemUnmap_t, "cuMemUnmap");
  CUdeviceptr va = 0;
  DL_DRV(reserve(&va, bytes, 0, 0, 0));
  CUresult r = map(va, bytes, 0, static_cast<CUmemGenericAllocationHandle>(handle), 0);
  if (r != CUDA_SUCCESS) {
    dl_set_error("cuMemMap -> %s (%d)", drv_err(r), (int)r);
    afree(va, bytes);
    return DL_ECUDA;
  }
  CUmemAccessDesc d;
  memset(&d, 0, sizeof(d));
  d.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  d.location.id = dev;
  d.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = access(va, bytes, &d, 1);
  if (r != CUDA_SUCCESS) {
    dl_set_error("cuMemSetAccess -> %s (%d)", drv_err(r), (int)r);
    unmap(va, bytes);
    afree(va, bytes);
    return DL_ECUDA;
  }
  *ptr = reinterpret_cast<void*>(va);
  return DL_OK;
}

extern "C" int dl_vmm_unmap(void* ptr, size_t bytes) {
  DL_DRV_FN(unmap, cuMemUnmap_t, "cuMemUnmap");
  DL_DRV_FN(afree, cuMemAddressFree_t, "cuMemAddressFree");
  DL_DRV(unmap(reinterpret_cast<CUdeviceptr>(ptr), bytes));
  DL_DRV(afree(reinterpret_cast<CUdeviceptr>(ptr), bytes));
  return DL_OK;
}

extern "C" int dl_vmm_release(uint64_t handle) {
  DL_DRV_FN(rel, cuMemRelease_t, "cuMemRelease");
  DL_DRV(rel(static_cast<CUmemGenericAllocationHandle>(handle)));
  return DL_OK;
}

extern "C" int dl_mc_granularity(int ndev, size_t bytes, size_t* gran) {
  DL_DRV_FN(fn, cuMulticastGetGranularity_t, "cuMulticastGetGranularity");
  CUmulticastObjectProp p = mc_prop(ndev, bytes);
  DL_DRV(fn(gran, &p, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  return DL_OK;
}

extern "C" int dl_mc_create(int ndev, size_t bytes, uint64_t* handle, int* fd) {
  DL_DRV_FN(create, cuMulticastCreate_t, "cuMulticastCreate");
  DL_DRV_FN(exp, cuMemExportToShareableHandle_t, "cuMemExportToShareableHandle");
  CUmulticastObjectProp p = mc_prop(ndev, bytes);
  CUmemGenericAllocationHandle h;
  DL_DRV(create(&h, &p));
  int out_fd = -1;
  DL_DRV(exp(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *handle = static_cast<uint64_t>(h);
  *fd = out_fd;
  return DL_OK;
}

extern "C" int dl_mc_add_device(uint64_t handle, int dev) {
  int rc = ensure_ctx(dev);
  if (rc) return rc;
  DL_DRV_FN(add, cuMulticastAddDevice_t, "cuMulticastAddDevice");
  DL_DRV(add(static_cast<CUmemGenericAllocationHandle>(handle), dev));
  return DL_OK;
}

extern "C" int dl_mc_bind(uint64_t mc_handle, uint64_t mem_handle, size_t bytes) {
  DL_DRV_FN(bind, cuMulticastBindMem_t, "cuMulticastBindMem");
  DL_DRV(bind(static_cast<CUmemGenericAllocationHandle>(mc_handle), 0, static_cast<CUmemGenericAllocationHandle>(mem_handle), 0,
              bytes, 0));
  return DL_OK;
}

extern "C" int dl_mc_unbind(uint64_t mc_handle, int dev, size_t bytes) {
  DL_DRV_FN(unbind, cuMulticastUnbind_t, "cuMulticastUnbind");
  DL_DRV(unbind(static_cast<CUmemGenericAllocationHandle>(mc_handle), dev, 0, bytes));
  return DL_OK;
}

// ---- cudaIpc fallback -------------------------------------------------------------------------
extern "C" int dl_ipc_alloc(int dev, size_t bytes, void** ptr, void* handle64) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle size");
  DL_CUDA(cudaSetDevice(dev));
  DL_CUDA(cudaMalloc(ptr, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, *ptr);
  if (e != cudaSuccess) {
    dl_set_error("cudaIpcGetMemHandle -> %s", cudaGetErrorString(e));
    cudaFree(*ptr);
    return DL_ECUDA;
  }
  memcpy(handle64, &h, 64);
  return DL_OK;
}

extern "C" int dl_ipc_open(int dev, const void* handle64, void** ptr) {
  DL_CUDA(cudaSetDevice(dev));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  DL_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return DL_OK;
}

extern "C" int dl_ipc_close(void* ptr) {
  DL_CUDA(cudaIpcCloseMemHandle(ptr));
  return DL_OK;
}

extern "C" int dl_ipc_free(void* ptr) {
  DL_CUDA(cudaFree(ptr));
  return DL_OK;
}

extern "C" int dl_can_access_peer(int dev, int peer, int* can) {
  DL_CUDA(cudaDeviceCanAccessPeer(can, dev, peer));
  return DL_OK;
}

extern "C" int dl_enable_peer(int dev, int peer) {
  DL_CUDA(cudaSetDevice(dev));
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return DL_OK;
  }
  DL_CUDA(e);
  return DL_OK;
}

// ---- fd passing over unix-domain sockets --------------------------------------------------------
extern "C" int dl_uds_listen(const char* name, int* sock) {
  int s = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
  if (s < 0) {
    dl_set_error("socket: %s", strerror(errno));
    return DL_ESYS;
  }
  struct sockaddr_un addr;
  socklen_t len = make_addr(name, &addr);
  if (bind(s, reinterpret_cast<struct sockaddr*>(&addr), len) != 0 || listen(s, 64) != 0) {
    dl_set_error("bind/listen '%s': %s", name, strerror(errno));
    close(s);
    return DL_ESYS;
  }
  *sock = s;
  return DL_OK;
}

extern "C" int dl_uds_send_fd(const char* name, int fd, int tag, int timeout_ms) {
  const long long deadline = now_ms() + (timeout_ms > 0 ? timeout_ms : 30000);
  struct sockaddr_un addr;
  socklen_t len = make_addr(name, &addr);
  int s = -1;
  while (true) {
    s = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (s < 0) {
      dl_set_error("socket: %s", strerror(errno));
      return DL_ESYS;
    }
    if (connect(s, reinterpret_cast<struct sockaddr*>(&addr), len) == 0) break;
    close(s);
    if (now_ms() > deadline) {
      dl_set_error("connect '%s': %s", name, strerror(errno));
      return DL_ETIMEOUT;
    }
    usleep(2000);
  }
  int payload[2] = {tag, 0};
  struct iovec iov = {payload, sizeof(payload)};
  char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  struct msghdr msg;
  memset(&msg, 0, sizeof(msg));
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  struct cmsghdr* cm = CMSG_FIRSTHDR(&msg);
  cm->cmsg_level = SOL_SOCKET;
  cm->cmsg_type = SCM_RIGHTS;
  cm->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(cm), &fd, sizeof(int));
  ssize_t n = sendmsg(s, &msg, 0);
  if (n != static_cast<ssize_t>(sizeof(payload))) {
    dl_set_error("sendmsg '%s': %s", name, strerror(errno));
    close(s);
    return DL_ESYS;
  }
  // wait for the receiver's 1-byte ack so the fd is not closed before it was received
  char ack = 0;
  struct pollfd pfd = {s, POLLIN, 0};
  if (poll(&pfd, 1, timeout_ms > 0 ? timeout_ms : 30000) > 0) (void)!read(s, &ack, 1);
  close(s);
  return DL_OK;
}

extern "C" int dl_uds_recv_fd(int sock, int* fd, int* tag, int timeout_ms) {
  struct pollfd pfd = {sock, POLLIN, 0};
  int pr = poll(&pfd, 1, timeout_ms > 0 ? timeout_ms : 30000);
  if (pr <= 0) {
    dl_set_error("timed out waiting for a peer to send a memory handle");
    return DL_ETIMEOUT;
  }
  int s = accept4(sock, nullptr, nullptr, SOCK_CLOEXEC);
  if (s < 0) {
    dl_set_error("accept: %s", strerror(errno));
    return DL_ESYS;
  }
  int payload[2] = {0, 0};
  struct iovec iov = {payload, sizeof(payload)};
  char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  struct msghdr msg;
  memset(&msg, 0, sizeof(msg));
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  ssize_t n = recvmsg(s, &msg, MSG_WAITALL);
  if (n != static_cast<ssize_t>(sizeof(payload))) {
    dl_set_error("recvmsg: %s", n < 0 ? strerror(errno) : "short read");
    close(s);
    return DL_ESYS;
  }
  struct cmsghdr* cm = CMSG_FIRSTHDR(&msg);
  if (!cm || cm->cmsg_level != SOL_SOCKET || cm->cmsg_type != SCM_RIGHTS) {
    dl_set_error("recvmsg: no file descriptor attached");
    close(s);
    return DL_ESYS;
  }
  memcpy(fd, CMSG_DATA(cm), sizeof(int));
  *tag = payload[0];
  char ack = 1;
  (void)!write(s, &ack, 1);
  close(s);
  return DL_OK;
}

extern "C" int dl_uds_close(int sock) {
  close(sock);
  return DL_OK;
}

extern "C" int dl_close_fd(int fd) {
  close(fd);
  return DL_OK;
}
