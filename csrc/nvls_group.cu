// NVLink multicast (NVLS) group: one multicast object that every GPU of the job binds a buffer of its own into.
//
// A store to the multicast address is delivered by NVSwitch to the bound buffer of every GPU (`multimem.st`), a load
// with reduction returns the sum of all of them, added inside the switch (`multimem.ld_reduce`): the two operations the
// NVLS exchange variant is made of (exchange.cuh, variant 2). What has to happen on the host before that:
//
//   open  rank 0 creates the multicast object (cuMulticastCreate, sized to the recommended granularity), exports it as a
//         POSIX file descriptor and hands a copy to every other rank over an abstract-namespace Unix socket
//         (SCM_RIGHTS: descriptors only mean something inside a process, the kernel has to duplicate them across);
//         the others import it; everybody adds its device (cuMulticastAddDevice).
//   ----- a barrier of the caller's control plane: binding is only legal once ALL devices are in -----
//   map   everybody allocates physical memory (cuMemCreate), binds it (cuMulticastBindMem) and maps it twice: once as
//         ordinary device memory (the local view) and once through the multicast object (the all-GPU view).
//   ----- a second barrier before the first multimem instruction: a store to an unbound member is lost -----
//
// The driver API is reached through cudaGetDriverEntryPoint, so the library has no link dependency on libcuda.
// Every blocking socket call carries the caller's timeout; nothing in here can wait for ever.
// No counterpart in the reference, whose only collective is the rank-0 star of data_parallelism_train.py:226-244.
#include <cuda.h>
#include <cuda_runtime.h>
#include <errno.h>
#include <poll.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "../include/dnnb200.h"
#include "common.cuh"

namespace dnnb200 {
namespace {

struct DriverApi {
  CUresult (*DeviceGet)(CUdevice*, int) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                               unsigned long long) = nullptr;
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
                                         unsigned long long) = nullptr;
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  bool ok = false;
};

template <typename Fn>
bool entry(const char* name, Fn* fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
    cudaGetLastError();
    return false;
  }
  *fn = reinterpret_cast<Fn>(p);
  return true;
}

const DriverApi& driver() {
  static DriverApi api = [] {
    DriverApi a;
    a.ok = entry("cuDeviceGet", &a.DeviceGet) && entry("cuDeviceGetAttribute", &a.DeviceGetAttribute) &&
           entry("cuMulticastCreate", &a.MulticastCreate) && entry("cuMulticastGetGranularity", &a.MulticastGetGranularity) &&
           entry("cuMulticastAddDevice", &a.MulticastAddDevice) && entry("cuMulticastBindMem", &a.MulticastBindMem) &&
           entry("cuMulticastUnbind", &a.MulticastUnbind) && entry("cuMemCreate", &a.MemCreate) &&
           entry("cuMemRelease", &a.MemRelease) && entry("cuMemGetAllocationGranularity", &a.MemGetAllocationGranularity) &&
           entry("cuMemExportToShareableHandle", &a.MemExportToShareableHandle) &&
           entry("cuMemImportFromShareableHandle", &a.MemImportFromShareableHandle) &&
           entry("cuMemAddressReserve", &a.MemAddressReserve) && entry("cuMemAddressFree", &a.MemAddressFree) &&
           entry("cuMemMap", &a.MemMap) && entry("cuMemUnmap", &a.MemUnmap) && entry("cuMemSetAccess", &a.MemSetAccess);
    return a;
  }();
  return api;
}

#define DNN_CU_CHECK(expr)                                                                              \
  do {                                                                                                  \
    const CUresult cu_r_ = (expr);                                                                      \
    if (cu_r_ != CUDA_SUCCESS) {                                                                        \
      ::dnnb200::set_last_error("%s failed at %s:%d with CUresult %d", #expr, __FILE__, __LINE__, (int)cu_r_); \
      return ::dnnb200::kErrCuda;                                                                       \
    }                                                                                                   \
  } while (0)

// The calls below work on the group's device and leave the caller's current device as they found it.
struct DeviceScope {
  int prev = -1;
  explicit DeviceScope(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; cudaGetLastError(); }
    if (prev != dev) cudaSetDevice(dev);
  }
  ~DeviceScope() {
    if (prev >= 0) cudaSetDevice(prev);
  }
  DeviceScope(const DeviceScope&) = delete;
  DeviceScope& operator=(const DeviceScope&) = delete;
};

int64_t now_ms() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (int64_t)ts.tv_sec * 1000 + ts.tv_nsec / 1000000;
}

// abstract namespace (leading NUL): no file to clean up, gone with the last descriptor
socklen_t abstract_addr(sockaddr_un* a, const std::string& name) {
  memset(a, 0, sizeof(*a));
  a->sun_family = AF_UNIX;
  const size_t n = name.size() < sizeof(a->sun_path) - 2 ? name.size() : sizeof(a->sun_path) - 2;
  memcpy(a->sun_path + 1, name.data(), n);
  return (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
}

bool wait_fd(int fd, short events, int64_t deadline) {
  for (;;) {
    const int64_t left = deadline - now_ms();
    if (left <= 0) return false;
    pollfd p = {fd, events, 0};
    const int r = poll(&p, 1, (int)(left > 1000 ? 1000 : left));
    if (r > 0) return (p.revents & events) != 0;
    if (r < 0 && errno != EINTR) return false;
  }
}

// one message: `word` as payload and, when fd >= 0, that descriptor as SCM_RIGHTS ancillary data
bool send_word_fd(int sock, uint64_t word, int fd) {
  iovec iov = {&word, sizeof(word)};
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  if (fd >= 0) {
    memset(ctrl, 0, sizeof(ctrl));
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    cmsghdr* c = CMSG_FIRSTHDR(&msg);
    c->cmsg_level = SOL_SOCKET;
    c->cmsg_type = SCM_RIGHTS;
    c->cmsg_len = CMSG_LEN(sizeof(int));
    memcpy(CMSG_DATA(c), &fd, sizeof(int));
  }
  for (;;) {
    const ssize_t r = sendmsg(sock, &msg, MSG_NOSIGNAL);
    if (r == (ssize_t)sizeof(word)) return true;
    if (r < 0 && errno == EINTR) continue;
    return false;
  }
}

bool recv_word_fd(int sock, uint64_t* word, int* fd, int64_t deadline) {
  if (!wait_fd(sock, POLLIN, deadline)) return false;
  iovec iov = {word, sizeof(*word)};
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  ssize_t r;
  do { r = recvmsg(sock, &msg, MSG_CMSG_CLOEXEC); } while (r < 0 && errno == EINTR);
  if (r != (ssize_t)sizeof(*word)) return false;
  if (fd) {
    *fd = -1;
    for (cmsghdr* c = CMSG_FIRSTHDR(&msg); c; c = CMSG_NXTHDR(&msg, c))
      if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) memcpy(fd, CMSG_DATA(c), sizeof(int));
  }
  return true;
}


// Rank 0 hands (word, descriptor) to every other rank; the others return what they got. 0 = ok.
int share_from_root(int rank, int world, const std::string& name, uint64_t* word, int* fd, int64_t deadline) {
  sockaddr_un addr;
  const socklen_t alen = abstract_addr(&addr, name);
  if (rank == 0) {
    const int ls = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    bool ok = ls >= 0 && bind(ls, (sockaddr*)&addr, alen) == 0 && listen(ls, world) == 0;
    std::vector<bool> served((size_t)world, false);
    for (int left = world - 1; ok && left > 0;) {
      if (!wait_fd(ls, POLLIN, deadline)) { ok = false; break; }
      const int c = accept4(ls, nullptr, nullptr, SOCK_CLOEXEC);
      if (c < 0) { if (errno == EINTR || errno == EAGAIN) continue; ok = false; break; }
      uint64_t who = ~0ull;
      if (recv_word_fd(c, &who, nullptr, deadline) && who > 0 && who < (uint64_t)world && !served[(size_t)who] &&
          send_word_fd(c, *word, *fd)) {
        served[(size_t)who] = true;
        --left;
      }
      close(c);
    }
    if (ls >= 0) close(ls);
    if (!ok) { set_last_error("share_fd: rank 0 could not serve every rank in time (errno %d)", errno); return kErrTimeout; }
    return kOk;
  }
  int s = -1;
  bool connected = false;
  while (!connected && now_ms() < deadline) {         // rank 0 may not be listening yet
    s = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (s < 0) break;
    if (connect(s, (sockaddr*)&addr, alen) == 0) { connected = true; break; }
    close(s);
    s = -1;
    usleep(2000);
  }
  int got = -1;
  const bool ok = connected && send_word_fd(s, (uint64_t)rank, -1) && recv_word_fd(s, word, &got, deadline) && got >= 0;
  if (s >= 0) close(s);
  if (!ok) { if (got >= 0) close(got); set_last_error("share_fd: rank %d did not get the descriptor from rank 0", rank); return kErrTimeout; }
  *fd = got;
  return kOk;
}

}  // namespace
}  // namespace dnnb200

using namespace dnnb200;

struct dnnb200_nvls {
  int rank = 0, world = 0, device = 0;
  CUdevice cu_dev = 0;
  size_t bytes = 0;                        // rounded up to the multicast granularity
  CUmemGenericAllocationHandle mc = 0, mem = 0;
  bool have_mc = false, have_mem = false, bound = false;
  CUdeviceptr local_va = 0, mc_va = 0;
  bool local_mapped = false, mc_mapped = false;
};

extern "C" {

int dnnb200_share_fd(int rank, int world, const char* name, unsigned long long* word, int* fd, int timeout_ms) {
  DNN_ARG_CHECK(name && name[0] && word && fd, "share_fd: null argument");
  DNN_ARG_CHECK(world >= 1 && rank >= 0 && rank < world, "share_fd: bad rank/world");
  DNN_ARG_CHECK(rank != 0 || *fd >= 0, "share_fd: rank 0 must bring the descriptor");
  if (world == 1) return kOk;
  uint64_t w = *word;
  const int rc = share_from_root(rank, world, std::string("dnnb200-fd-") + name, &w, fd,
                                 now_ms() + (timeout_ms > 0 ? timeout_ms : 30000));
  *word = w;
  return rc;
}

int dnnb200_nvls_supported(int device) {
  const DriverApi& d = driver();
  if (!d.ok) return 0;
  CUdevice dev;
  int v = 0;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) { cudaGetLastError(); return 0; }
  DeviceScope scope(device);
  if (cudaFree(nullptr) != cudaSuccess) { cudaGetLastError(); return 0; }
  if (d.DeviceGet(&dev, device) != CUDA_SUCCESS) return 0;
  if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return 0;
  return v != 0;
}

int dnnb200_nvls_open(dnnb200_nvls** out, int rank, int world, int device, size_t bytes, const char* name,
                      int timeout_ms) {
  DNN_ARG_CHECK(out && name && name[0], "nvls_open: null argument");
  DNN_ARG_CHECK(world >= 1 && rank >= 0 && rank < world && bytes > 0, "nvls_open: bad rank/world/bytes");
  *out = nullptr;
  const DriverApi& d = driver();
  if (!d.ok) { set_last_error("nvls_open: this driver has no cuMulticast* entry points"); return kErrState; }
  int count = 0;
  DNN_CUDA_CHECK(cudaGetDeviceCount(&count));
  DNN_ARG_CHECK(device >= 0 && device < count, "nvls_open: no device %d", device);
  DeviceScope scope(device);
  DNN_CUDA_CHECK(cudaFree(nullptr));                       // make sure the primary context exists and is current
  dnnb200_nvls* g = new dnnb200_nvls;
  g->rank = rank; g->world = world; g->device = device;
  auto fail = [&](int rc) { dnnb200_nvls_close(g); return rc; };
  if (d.DeviceGet(&g->cu_dev, device) != CUDA_SUCCESS) { set_last_error("nvls_open: cuDeviceGet failed"); return fail(kErrCuda); }
  int supported = 0;
  d.DeviceGetAttribute(&supported, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, g->cu_dev);
  if (!supported) { set_last_error("nvls_open: device %d does not support NVLink multicast", device); return fail(kErrState); }

  // Every rank computes the same size (same world, same request, same kind of GPU): object and bound buffers agree.
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)world;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  mp.size = bytes;
  size_t gran_mc = 0, gran_mem = 0;
  CUresult r = d.MulticastGetGranularity(&gran_mc, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED);
  if (r != CUDA_SUCCESS || gran_mc == 0) { set_last_error("nvls_open: cuMulticastGetGranularity failed (%d)", (int)r); return fail(kErrCuda); }
  CUmemAllocationProp ap;
  memset(&ap, 0, sizeof(ap));
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  r = d.MemGetAllocationGranularity(&gran_mem, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  if (r != CUDA_SUCCESS || gran_mem == 0) { set_last_error("nvls_open: cuMemGetAllocationGranularity failed (%d)", (int)r); return fail(kErrCuda); }
  const size_t gran = gran_mc > gran_mem ? gran_mc : gran_mem;
  g->bytes = (bytes + gran - 1) / gran * gran;
  mp.size = g->bytes;

  const int64_t deadline = now_ms() + (timeout_ms > 0 ? timeout_ms : 30000);
  const std::string sock_name = std::string("dnnb200-nvls-") + name;
  if (rank == 0) {
    r = d.MulticastCreate(&g->mc, &mp);
    if (r != CUDA_SUCCESS) { set_last_error("nvls_open: cuMulticastCreate failed (%d)", (int)r); return fail(kErrCuda); }
    g->have_mc = true;
    if (world > 1) {
      int fd = -1;
      r = d.MemExportToShareableHandle(&fd, g->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
      if (r != CUDA_SUCCESS || fd < 0) { set_last_error("nvls_open: export of the multicast handle failed (%d)", (int)r); return fail(kErrCuda); }
      uint64_t size_word = (uint64_t)g->bytes;
      const int rc = share_from_root(0, world, sock_name, &size_word, &fd, deadline);
      close(fd);
      if (rc) return fail(rc);
    }
  } else {
    uint64_t size_word = 0;
    int fd = -1;
    const int rc = share_from_root(rank, world, sock_name, &size_word, &fd, deadline);
    if (rc) return fail(rc);
    if (size_word != (uint64_t)g->bytes) { close(fd); set_last_error("nvls_open: ranks disagree on the size (%llu vs %llu)", (unsigned long long)size_word, (unsigned long long)g->bytes); return fail(kErrState); }
    r = d.MemImportFromShareableHandle(&g->mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    close(fd);
    if (r != CUDA_SUCCESS) { set_last_error("nvls_open: import of the multicast handle failed (%d)", (int)r); return fail(kErrCuda); }
    g->have_mc = true;
  }
  r = d.MulticastAddDevice(g->mc, g->cu_dev);
  if (r != CUDA_SUCCESS) { set_last_error("nvls_open: cuMulticastAddDevice failed (%d)", (int)r); return fail(kErrCuda); }
  *out = g;
  return kOk;
}

int dnnb200_nvls_map(dnnb200_nvls* g, float** local, float** multicast, size_t* bytes) {
  DNN_ARG_CHECK(g && local && multicast, "nvls_map: null argument");
  DNN_ARG_CHECK(g->have_mc && !g->bound, "nvls_map: group is not open, or already mapped");
  const DriverApi& d = driver();
  DeviceScope scope(g->device);
  CUmemAllocationProp ap;
  memset(&ap, 0, sizeof(ap));
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = g->device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  DNN_CU_CHECK(d.MemCreate(&g->mem, g->bytes, &ap, 0));
  g->have_mem = true;
  DNN_CU_CHECK(d.MulticastBindMem(g->mc, 0, g->mem, 0, g->bytes, 0));
  g->bound = true;
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = g->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  DNN_CU_CHECK(d.MemAddressReserve(&g->local_va, g->bytes, 0, 0, 0));
  DNN_CU_CHECK(d.MemMap(g->local_va, g->bytes, 0, g->mem, 0));
  g->local_mapped = true;
  DNN_CU_CHECK(d.MemSetAccess(g->local_va, g->bytes, &acc, 1));
  DNN_CU_CHECK(d.MemAddressReserve(&g->mc_va, g->bytes, 0, 0, 0));
  DNN_CU_CHECK(d.MemMap(g->mc_va, g->bytes, 0, g->mc, 0));
  g->mc_mapped = true;
  DNN_CU_CHECK(d.MemSetAccess(g->mc_va, g->bytes, &acc, 1));
  DNN_CUDA_CHECK(cudaMemset((void*)g->local_va, 0, g->bytes));   // the exchange's flags start at zero
  DNN_CUDA_CHECK(cudaDeviceSynchronize());
  *local = (float*)g->local_va;
  *multicast = (float*)g->mc_va;
  if (bytes) *bytes = g->bytes;
  return kOk;
}

int dnnb200_nvls_close(dnnb200_nvls* g) {
  if (!g) return kOk;
  const DriverApi& d = driver();
  if (d.ok) {
    DeviceScope scope(g->device);
    cudaDeviceSynchronize();
    if (g->mc_mapped) d.MemUnmap(g->mc_va, g->bytes);
    if (g->mc_va) d.MemAddressFree(g->mc_va, g->bytes);
    if (g->local_mapped) d.MemUnmap(g->local_va, g->bytes);
    if (g->local_va) d.MemAddressFree(g->local_va, g->bytes);
    if (g->bound) d.MulticastUnbind(g->mc, g->cu_dev, 0, g->bytes);
    if (g->have_mem) d.MemRelease(g->mem);
    if (g->have_mc) d.MemRelease(g->mc);
    cudaGetLastError();
  }
  delete g;
  return kOk;
}

}  // extern "C"
