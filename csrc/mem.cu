// Peer-memory manager (SURVEY §7.2 "mem"): CUDA VMM allocations shared between the worker /
// coordinator processes of one NVSwitch box, and NVLS multicast objects on top of them.
//
// Replaces the reference's IPC medium - checkpoint files on disk and base64 strings in protobuf
// (src/client.py:19-23, src/server.py:55-58,159-160) - with memory that every process maps:
//   * fedb_mem_create      cuMemCreate (POSIX-fd shareable, pinned on one device) + reserve + map
//   * fedb_mem_export_fd   fd for SCM_RIGHTS transfer; the physical memory outlives the exporter
//                          for as long as an importer holds it (what failover relies on)
//   * fedb_mem_import_fd   map a peer's allocation into this process and grant this device access
//   * fedb_mc_*            multicast object spanning one allocation per device: multimem.st through
//                          its mapping lands in every member, multimem.ld_reduce sums across them
//   * fedb_fd_serve/fetch  the fd exchange itself, over a unix domain socket
// Driver entry points are resolved through cudaGetDriverEntryPoint, so the library has no link-time
// dependency on libcuda and loads on a box without a driver (the CPU build check).
#include "common.cuh"
#include "../include/fedb200.h"
#include <cuda.h>
#include <errno.h>
#include <fcntl.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

thread_local char g_fedb_err[512] = {0};

int fedb_fail(int code, const char* file, int line, const char* what) {
  snprintf(g_fedb_err, sizeof(g_fedb_err), "%s:%d: %s (code %d)", file, line, what ? what : "error", code);
  return code ? code : -1;
}

extern "C" const char* fedb_last_error(void) { return g_fedb_err; }
extern "C" int fedb_version(void) { return 100; }

namespace {

template <typename Fn>
int resolve(const char* name, Fn* out) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q);
  if (e != cudaSuccess || fn == nullptr || q != cudaDriverEntryPointSuccess)
    return fedb_fail(e ? (int)e : -2, __FILE__, __LINE__, name);
  *out = reinterpret_cast<Fn>(fn);
  return 0;
}

#define DRV(name, ...)                                                              \
  do {                                                                              \
    static decltype(&name) fn_ = nullptr;                                           \
    if (!fn_) { if (int rc_ = resolve(#name, &fn_)) return rc_; }                   \
    CUresult r_ = fn_(__VA_ARGS__);                                                 \
    if (r_ != CUDA_SUCCESS) return fedb_fail((int)r_, __FILE__, __LINE__, #name);   \
  } while (0)

CUmemAllocationProp device_prop(int device) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

int map_rw(CUmemGenericAllocationHandle h, uint64_t bytes, int device, uint64_t align, uint64_t* out_ptr) {
  CUdeviceptr va = 0;
  DRV(cuMemAddressReserve, &va, bytes, align, 0, 0);
  DRV(cuMemMap, va, bytes, 0, h, 0);
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  DRV(cuMemSetAccess, va, bytes, &acc, 1);
  *out_ptr = (uint64_t)va;
  return 0;
}

}  // namespace

extern "C" int fedb_mem_supported(int device, int* vmm, int* posix_fd, int* multicast) {
  FEDB_CHECK(cudaFree(0));   // make sure a context exists
  CUdevice dev;
  DRV(cuDeviceGet, &dev, device);
  int v = 0;
  DRV(cuDeviceGetAttribute, &v, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev);
  if (vmm) *vmm = v;
  DRV(cuDeviceGetAttribute, &v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev);
  if (posix_fd) *posix_fd = v;
  DRV(cuDeviceGetAttribute, &v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  if (multicast) *multicast = v;
  return 0;
}

extern "C" int fedb_mem_create(uint64_t bytes, int device, fedb_mem_t* out) {
  FEDB_REQUIRE(out != nullptr && bytes > 0, "fedb_mem_create: bad arguments");
  FEDB_CHECK(cudaSetDevice(device));
  FEDB_CHECK(cudaFree(0));
  CUmemAllocationProp prop = device_prop(device);
  size_t gran = 0;
  DRV(cuMemGetAllocationGranularity, &gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  if (gran < (2u << 20)) gran = 2u << 20;
  // a multicast object must be a multiple of ITS granularity, and it spans the whole allocation:
  // pad to the larger of the two so that fedb_mc_create(bytes) succeeds for every arena size
  {
    CUdevice dev;
    DRV(cuDeviceGet, &dev, device);
    int mc = 0;
    DRV(cuDeviceGetAttribute, &mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
    for (int it = 0; mc && it < 4; ++it) {   // the granularity may depend on the (padded) size: iterate to a fixpoint
      CUmulticastObjectProp mp;
      memset(&mp, 0, sizeof(mp));
      mp.numDevices = 2;
      mp.size = (bytes + gran - 1) / gran * gran;
      mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
      size_t mgran = 0;
      DRV(cuMulticastGetGranularity, &mgran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED);
      if (mgran <= gran || mgran % gran != 0) {
        if (mgran > gran) gran = mgran;
        else break;
      } else {
        gran = mgran;
      }
    }
  }
  const uint64_t padded = (bytes + gran - 1) / gran * gran;
  CUmemGenericAllocationHandle h;
  DRV(cuMemCreate, &h, padded, &prop, 0);
  uint64_t ptr = 0;
  if (int rc = map_rw(h, padded, device, gran, &ptr)) return rc;
  FEDB_CHECK(cudaMemset((void*)ptr, 0, padded));
  out->handle = (uint64_t)h;
  out->ptr = ptr;
  out->bytes = padded;
  out->device = device;
  out->fd = -1;
  return 0;
}

extern "C" int fedb_mem_export_fd(fedb_mem_t* mem) {
  FEDB_REQUIRE(mem != nullptr, "fedb_mem_export_fd: null");
  if (mem->fd >= 0) return 0;
  int fd = -1;
  DRV(cuMemExportToShareableHandle, (void*)&fd, (CUmemGenericAllocationHandle)mem->handle,
      CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  mem->fd = fd;
  return 0;
}

extern "C" int fedb_mem_import_fd(int fd, uint64_t bytes, int device, fedb_mem_t* out) {
  FEDB_REQUIRE(out != nullptr && fd >= 0, "fedb_mem_import_fd: bad arguments");
  FEDB_CHECK(cudaSetDevice(device));
  FEDB_CHECK(cudaFree(0));
  CUmemGenericAllocationHandle h;
  DRV(cuMemImportFromShareableHandle, &h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  uint64_t ptr = 0;
  if (int rc = map_rw(h, bytes, device, 2u << 20, &ptr)) return rc;
  out->handle = (uint64_t)h;
  out->ptr = ptr;
  out->bytes = bytes;
  out->device = -1;
  out->fd = -1;
  return 0;
}

extern "C" int fedb_mem_release(fedb_mem_t* mem) {
  if (!mem || !mem->ptr) return 0;
  DRV(cuMemUnmap, (CUdeviceptr)mem->ptr, mem->bytes);
  DRV(cuMemAddressFree, (CUdeviceptr)mem->ptr, mem->bytes);
  DRV(cuMemRelease, (CUmemGenericAllocationHandle)mem->handle);
  if (mem->fd >= 0) close(mem->fd);
  mem->ptr = 0;
  mem->fd = -1;
  return 0;
}

// ------------------------------------------------------------------------------- multicast (NVLS)
extern "C" int fedb_mc_create(uint64_t bytes, int n_devices, fedb_mem_t* out_mc) {
  FEDB_REQUIRE(out_mc != nullptr && n_devices >= 1, "fedb_mc_create: bad arguments");
  CUmulticastObjectProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.numDevices = (unsigned)n_devices;
  prop.size = bytes;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  DRV(cuMulticastGetGranularity, &gran, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED);
  FEDB_REQUIRE(bytes % gran == 0, "fedb_mc_create: size is not a multiple of the multicast granularity");
  CUmemGenericAllocationHandle h;
  DRV(cuMulticastCreate, &h, &prop);
  int fd = -1;
  DRV(cuMemExportToShareableHandle, (void*)&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  out_mc->handle = (uint64_t)h;
  out_mc->ptr = 0;
  out_mc->bytes = bytes;
  out_mc->device = -1;
  out_mc->fd = fd;
  return 0;
}

extern "C" int fedb_mc_import_fd(int fd, uint64_t bytes, fedb_mem_t* out_mc) {
  FEDB_REQUIRE(out_mc != nullptr && fd >= 0, "fedb_mc_import_fd: bad arguments");
  CUmemGenericAllocationHandle h;
  DRV(cuMemImportFromShareableHandle, &h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  out_mc->handle = (uint64_t)h;
  out_mc->ptr = 0;
  out_mc->bytes = bytes;
  out_mc->device = -1;
  out_mc->fd = -1;
  return 0;
}

extern "C" int fedb_mc_add_device(fedb_mem_t* mc, int device) {
  FEDB_REQUIRE(mc != nullptr, "fedb_mc_add_device: null");
  CUdevice dev;
  DRV(cuDeviceGet, &dev, device);
  DRV(cuMulticastAddDevice, (CUmemGenericAllocationHandle)mc->handle, dev);
  return 0;
}

extern "C" int fedb_mc_bind_and_map(fedb_mem_t* mc, const fedb_mem_t* local_mem, int device) {
  FEDB_REQUIRE(mc && local_mem, "fedb_mc_bind_and_map: null");
  FEDB_REQUIRE(local_mem->bytes >= mc->bytes, "fedb_mc_bind_and_map: local allocation smaller than the multicast object");
  // every device must have been added (by every rank) before the first bind: callers barrier between
  DRV(cuMulticastBindMem, (CUmemGenericAllocationHandle)mc->handle, 0,
      (CUmemGenericAllocationHandle)local_mem->handle, 0, mc->bytes, 0);
  uint64_t ptr = 0;
  if (int rc = map_rw((CUmemGenericAllocationHandle)mc->handle, mc->bytes, device, 2u << 20, &ptr)) return rc;
  mc->ptr = ptr;
  return 0;
}

// ------------------------------------------------------------------------------- fd exchange
namespace {

int send_fds(int sock, const int* fds, int n) {
  char ctrl[CMSG_SPACE(sizeof(int) * 32)];
  memset(ctrl, 0, sizeof(ctrl));
  int32_t count = n;
  struct iovec io = {&count, sizeof(count)};
  struct msghdr msg;
  memset(&msg, 0, sizeof(msg));
  msg.msg_iov = &io;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = CMSG_SPACE(sizeof(int) * n);
  struct cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int) * n);
  memcpy(CMSG_DATA(c), fds, sizeof(int) * n);
  return sendmsg(sock, &msg, 0) == (ssize_t)sizeof(count) ? 0 : -1;
}

int recv_fds(int sock, int* fds, int n) {
  char ctrl[CMSG_SPACE(sizeof(int) * 32)];
  memset(ctrl, 0, sizeof(ctrl));
  int32_t count = 0;
  struct iovec io = {&count, sizeof(count)};
  struct msghdr msg;
  memset(&msg, 0, sizeof(msg));
  msg.msg_iov = &io;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  if (recvmsg(sock, &msg, 0) != (ssize_t)sizeof(count)) return -1;
  struct cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (!c || c->cmsg_level != SOL_SOCKET || c->cmsg_type != SCM_RIGHTS || count != n) return -1;
  memcpy(fds, CMSG_DATA(c), sizeof(int) * n);
  return 0;
}

long long now_ms() {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (long long)ts.tv_sec * 1000 + ts.tv_nsec / 1000000;
}

}  // namespace

extern "C" int fedb_fd_serve(const char* path, const int* fds, int n_fds, int n_clients, int timeout_ms) {
  FEDB_REQUIRE(path && fds && n_fds > 0 && n_fds <= 32, "fedb_fd_serve: bad arguments");
  int srv = socket(AF_UNIX, SOCK_STREAM, 0);
  FEDB_REQUIRE(srv >= 0, "fedb_fd_serve: socket()");
  struct sockaddr_un addr;
  memset(&addr, 0, sizeof(addr));
  addr.sun_family = AF_UNIX;
  strncpy(addr.sun_path, path, sizeof(addr.sun_path) - 1);
  unlink(path);
  if (bind(srv, (struct sockaddr*)&addr, sizeof(addr)) != 0 || listen(srv, 64) != 0) {
    close(srv);
    return fedb_fail(errno, __FILE__, __LINE__, "fedb_fd_serve: bind/listen");
  }
  chmod(path, 0600);             // the descriptors give read/write access to this process's GPU memory: owner only
  const long long deadline = now_ms() + timeout_ms;
  int served = 0, rc = 0;
  while (served < n_clients) {
    struct pollfd pfd = {srv, POLLIN, 0};
    long long left = deadline - now_ms();
    if (left <= 0 || poll(&pfd, 1, (int)left) <= 0) { rc = fedb_fail(-3, __FILE__, __LINE__, "fedb_fd_serve: timeout"); break; }
    int c = accept(srv, nullptr, nullptr);
    if (c < 0) continue;
    // only processes of the same user (or root) get the descriptors, whatever the mode of the socket's directory
    struct ucred cred;
    socklen_t clen = sizeof(cred);
    if (getsockopt(c, SOL_SOCKET, SO_PEERCRED, &cred, &clen) != 0 || (cred.uid != geteuid() && cred.uid != 0)) {
      close(c);
      continue;
    }
    if (send_fds(c, fds, n_fds) != 0) rc = fedb_fail(errno, __FILE__, __LINE__, "fedb_fd_serve: sendmsg");
    // wait (bounded) until the client has the fds before closing: a client that connects and never acknowledges must
    // not wedge a persistent server
    struct pollfd ack_fd = {c, POLLIN, 0};
    if (poll(&ack_fd, 1, 5000) > 0) { char ack = 0; (void)!read(c, &ack, 1); }
    close(c);
    ++served;
  }
  close(srv);
  unlink(path);
  return rc;
}

extern "C" int fedb_fd_fetch(const char* path, int* out_fds, int n_fds, int timeout_ms) {
  FEDB_REQUIRE(path && out_fds && n_fds > 0 && n_fds <= 32, "fedb_fd_fetch: bad arguments");
  const long long deadline = now_ms() + timeout_ms;
  while (true) {
    int s = socket(AF_UNIX, SOCK_STREAM, 0);
    FEDB_REQUIRE(s >= 0, "fedb_fd_fetch: socket()");
    struct sockaddr_un addr;
    memset(&addr, 0, sizeof(addr));
    addr.sun_family = AF_UNIX;
    strncpy(addr.sun_path, path, sizeof(addr.sun_path) - 1);
    if (connect(s, (struct sockaddr*)&addr, sizeof(addr)) == 0) {
      int rc = recv_fds(s, out_fds, n_fds);
      char ack = 1;
      (void)!write(s, &ack, 1);
      close(s);
      if (rc != 0) return fedb_fail(errno, __FILE__, __LINE__, "fedb_fd_fetch: recvmsg");
      return 0;
    }
    close(s);
    if (now_ms() > deadline) return fedb_fail(-3, __FILE__, __LINE__, "fedb_fd_fetch: timeout waiting for the peer socket");
    usleep(20000);
  }
}

// Asynchronous copy between any two of {pinned host, device, peer device} on `stream`.  Exists so that the step
// engine can put the H2D copies of a round's inputs INSIDE the captured round graph (memcpy nodes on a forked
// branch, steps depending on their own chunk only) without going through the framework allocator's bookkeeping.
extern "C" int fedb_copy_async(void* dst, const void* src, uint64_t bytes, void* stream) {
  if (bytes == 0) return 0;
  FEDB_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, (cudaStream_t)stream));
  return 0;
}

// Zero-fill as a memset node of the stream / graph (no kernel launch): gradients of a stride-2 1x1 data gradient, counters.
extern "C" int fedb_zero_async(void* dst, uint64_t bytes, void* stream) {
  if (bytes == 0) return 0;
  FEDB_CHECK(cudaMemsetAsync(dst, 0, bytes, (cudaStream_t)stream));
  return 0;
}
