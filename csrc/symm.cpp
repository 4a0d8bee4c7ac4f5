// Symmetric memory bootstrap: one CUDA VMM allocation per rank, shared with every peer of the
// node as a POSIX file descriptor passed over unix-domain sockets (SCM_RIGHTS), mapped into every rank's address
// space, and optionally bound to an NVSwitch multicast object for NVLS (multimem.*) kernels.
// This is the plumbing the reference gets from init_process_group + NCCL (src/trainer.py:59); nothing here is
// on the hot path.  The driver API is resolved at run time (no link dependency on libcuda).
#include <cuda.h>
#include <cuda_runtime.h>

#include <cerrno>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include <sys/socket.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

#include "../include/mlt_b200.h"

namespace mlt {
void set_error(const char* fmt, ...);
}
using mlt::set_error;

namespace {

struct Driver {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemGetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*McCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*McAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*McBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*McGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGet)(CUdevice*, int);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  bool ok = false;
  std::string missing;
};

template <typename F>
bool resolve(F& fn, const char* name, Driver& d) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
    d.missing = name;
    return false;
  }
  fn = reinterpret_cast<F>(p);
  return true;
}

Driver& driver() {
  static Driver d;
  static bool tried = false;
  if (!tried) {
    tried = true;
    cudaFree(nullptr);   // make sure the runtime has a context
    d.ok = resolve(d.MemCreate, "cuMemCreate", d) && resolve(d.MemRelease, "cuMemRelease", d) &&
           resolve(d.MemExport, "cuMemExportToShareableHandle", d) && resolve(d.MemImport, "cuMemImportFromShareableHandle", d) &&
           resolve(d.MemAddressReserve, "cuMemAddressReserve", d) && resolve(d.MemAddressFree, "cuMemAddressFree", d) &&
           resolve(d.MemMap, "cuMemMap", d) && resolve(d.MemUnmap, "cuMemUnmap", d) && resolve(d.MemSetAccess, "cuMemSetAccess", d) &&
           resolve(d.MemGetGranularity, "cuMemGetAllocationGranularity", d) && resolve(d.McCreate, "cuMulticastCreate", d) &&
           resolve(d.McAddDevice, "cuMulticastAddDevice", d) && resolve(d.McBindMem, "cuMulticastBindMem", d) &&
           resolve(d.McGetGranularity, "cuMulticastGetGranularity", d) && resolve(d.DeviceGet, "cuDeviceGet", d) &&
           resolve(d.DeviceGetAttribute, "cuDeviceGetAttribute", d);
  }
  return d;
}

#define DRV(call, what)                                                        \
  do {                                                                         \
    CUresult _r = (call);                                                      \
    if (_r != CUDA_SUCCESS) {                                                  \
      set_error("symm: %s failed with CUresult %d (%s:%d)", what, (int)_r, __FILE__, __LINE__); \
      return 1;                                                                \
    }                                                                          \
  } while (0)

struct Symm {
  int rank = 0, world = 1, device = 0;
  size_t bytes = 0;                        // rounded to granularity
  std::string tag;
  CUmemGenericAllocationHandle local = 0;
  int local_fd = -1;
  int listen_fd = -1;
  std::vector<CUmemGenericAllocationHandle> peer_handles;
  std::vector<CUdeviceptr> peer_va;
  // multicast
  CUmemGenericAllocationHandle mc = 0;
  int mc_fd = -1;
  CUdeviceptr mc_va = 0;
  bool mc_ready = false;
};

std::string sock_name(const std::string& tag, int rank) { return "mlt-symm-" + tag + "-" + std::to_string(rank); }

int make_listener(const std::string& name, int backlog) {
  int fd = socket(AF_UNIX, SOCK_STREAM, 0);
  if (fd < 0) return -1;
  sockaddr_un addr;
  std::memset(&addr, 0, sizeof(addr));
  addr.sun_family = AF_UNIX;
  // abstract namespace: sun_path[0] == '\0', no file to clean up
  std::memcpy(addr.sun_path + 1, name.c_str(), name.size());
  socklen_t len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + name.size());
  if (bind(fd, (sockaddr*)&addr, len) != 0 || listen(fd, backlog) != 0) {
    close(fd);
    return -1;
  }
  return fd;
}

int connect_to(const std::string& name, double timeout_s) {
  timespec t0;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (;;) {
    int fd = socket(AF_UNIX, SOCK_STREAM, 0);
    if (fd < 0) return -1;
    sockaddr_un addr;
    std::memset(&addr, 0, sizeof(addr));
    addr.sun_family = AF_UNIX;
    std::memcpy(addr.sun_path + 1, name.c_str(), name.size());
    socklen_t len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + name.size());
    if (connect(fd, (sockaddr*)&addr, len) == 0) return fd;
    close(fd);
    timespec t1;
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if ((t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec) > timeout_s) return -1;
    usleep(2000);
  }
}

// message = {int sender_rank, int kind} + one fd as ancillary data
int send_fd(int sock, int fd, int sender, int kind) {
  int payload[2] = {sender, kind};
  iovec iov{payload, sizeof(payload)};
  char ctrl[CMSG_SPACE(sizeof(int))];
  std::memset(ctrl, 0, sizeof(ctrl));
  msghdr msg{};
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  std::memcpy(CMSG_DATA(c), &fd, sizeof(int));
  return sendmsg(sock, &msg, 0) == (ssize_t)sizeof(payload) ? 0 : -1;
}

int recv_fd(int sock, int* fd, int* sender, int* kind) {
  int payload[2] = {-1, -1};
  iovec iov{payload, sizeof(payload)};
  char ctrl[CMSG_SPACE(sizeof(int))];
  std::memset(ctrl, 0, sizeof(ctrl));
  msghdr msg{};
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  if (recvmsg(sock, &msg, MSG_WAITALL) != (ssize_t)sizeof(payload)) return -1;
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (!c || c->cmsg_level != SOL_SOCKET || c->cmsg_type != SCM_RIGHTS) return -1;
  std::memcpy(fd, CMSG_DATA(c), sizeof(int));
  *sender = payload[0];
  *kind = payload[1];
  return 0;
}

int map_handle(Driver& d, CUmemGenericAllocationHandle h, size_t bytes, int device, CUdeviceptr* va) {
  DRV(d.MemAddressReserve(va, bytes, 0, 0, 0), "cuMemAddressReserve");
  DRV(d.MemMap(*va, bytes, 0, h, 0), "cuMemMap");
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  DRV(d.MemSetAccess(*va, bytes, &acc, 1), "cuMemSetAccess");
  return 0;
}

}  // namespace

extern "C" void* mlt_symm_create(int rank, int world, int device, long long bytes, const char* job_tag) {
  Driver& d = driver();
  if (!d.ok) {
    set_error("symm: driver entry point %s is unavailable", d.missing.c_str());
    return nullptr;
  }
  if (world < 1 || world > 8 || rank < 0 || rank >= world || bytes <= 0) {
    set_error("symm: bad arguments (rank %d, world %d, bytes %lld)", rank, world, bytes);
    return nullptr;
  }
  if (cudaSetDevice(device) != cudaSuccess) {
    set_error("symm: cudaSetDevice(%d) failed", device);
    return nullptr;
  }
  cudaFree(nullptr);
  Symm* s = new Symm();
  s->rank = rank; s->world = world; s->device = device; s->tag = job_tag ? job_tag : "job";
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  if (d.MemGetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || gran == 0) gran = 2 << 20;
  // multicast binding wants its own (usually larger) granularity: round to the larger of the two up front
  CUmulticastObjectProp mp{};
  mp.numDevices = (unsigned)world; mp.size = (size_t)bytes; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t mgran = 0;
  if (world > 1 && d.McGetGranularity(&mgran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mgran > gran) gran = mgran;
  s->bytes = ((size_t)bytes + gran - 1) / gran * gran;
  CUresult r = d.MemCreate(&s->local, s->bytes, &prop, 0);
  if (r != CUDA_SUCCESS) {
    set_error("symm: cuMemCreate(%zu bytes) failed with CUresult %d", s->bytes, (int)r);
    delete s;
    return nullptr;
  }
  s->peer_handles.assign(world, 0);
  s->peer_va.assign(world, 0);
  s->peer_handles[rank] = s->local;
  if (map_handle(d, s->local, s->bytes, device, &s->peer_va[rank]) != 0) { delete s; return nullptr; }
  if (cudaMemset((void*)s->peer_va[rank], 0, s->bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
    set_error("symm: zero-fill of the arena failed");
    delete s;
    return nullptr;
  }
  if (world > 1) {
    if (d.MemExport(&s->local_fd, s->local, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) != CUDA_SUCCESS) {
      set_error("symm: cuMemExportToShareableHandle failed");
      delete s;
      return nullptr;
    }
    s->listen_fd = make_listener(sock_name(s->tag, rank), 4 * world);
    if (s->listen_fd < 0) {
      set_error("symm: cannot listen on abstract socket %s: %s", sock_name(s->tag, rank).c_str(), strerror(errno));
      delete s;
      return nullptr;
    }
  }
  return s;
}

// Send my fd to every peer, receive theirs, import and map.  All ranks must have returned from mlt_symm_create.
extern "C" int mlt_symm_exchange(void* handle) {
  Symm* s = (Symm*)handle;
  Driver& d = driver();
  if (s->world == 1) return 0;
  for (int p = 0; p < s->world; ++p) {
    if (p == s->rank) continue;
    int c = connect_to(sock_name(s->tag, p), 60.0);
    if (c < 0) { set_error("symm: rank %d cannot reach rank %d's socket", s->rank, p); return 1; }
    int rc = send_fd(c, s->local_fd, s->rank, 0);
    close(c);
    if (rc != 0) { set_error("symm: rank %d failed to send its handle to rank %d", s->rank, p); return 1; }
  }
  for (int i = 0; i < s->world - 1; ++i) {
    int c = accept(s->listen_fd, nullptr, nullptr);
    if (c < 0) { set_error("symm: accept failed: %s", strerror(errno)); return 1; }
    int fd = -1, sender = -1, kind = -1;
    int rc = recv_fd(c, &fd, &sender, &kind);
    close(c);
    if (rc != 0 || sender < 0 || sender >= s->world || kind != 0) { set_error("symm: bad handle message (sender %d kind %d)", sender, kind); return 1; }
    DRV(d.MemImport(&s->peer_handles[sender], (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR), "cuMemImportFromShareableHandle");
    close(fd);
    if (map_handle(d, s->peer_handles[sender], s->bytes, s->device, &s->peer_va[sender]) != 0) return 1;
  }
  return 0;
}

// phase 0: rank 0 creates the multicast object and ships its fd; everyone imports it and adds its device.
// phase 1: (after a barrier: all devices added) bind the local arena and map the multicast address.
// phase 2: nothing device-side; marks the mapping usable (kept as a phase so callers barrier after the binds).
extern "C" int mlt_symm_multicast_setup(void* handle, int phase) {
  Symm* s = (Symm*)handle;
  Driver& d = driver();
  if (s->world == 1) return -1;
  CUdevice dev;
  DRV(d.DeviceGet(&dev, s->device), "cuDeviceGet");
  if (phase == 0) {
    int supported = 0;
    if (d.DeviceGetAttribute(&supported, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) supported = 0;
    CUmulticastObjectProp mp{};
    mp.numDevices = (unsigned)s->world;
    mp.size = s->bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    if (s->rank == 0) {
      // rank 0 ALWAYS tells every peer the outcome (kind 1 = handle attached, kind 2 = no multicast), so nobody blocks
      int fd = -1;
      if (supported) {
        CUresult r = d.McCreate(&s->mc, &mp);
        if (r != CUDA_SUCCESS) { set_error("symm: cuMulticastCreate failed with CUresult %d", (int)r); s->mc = 0; }
        if (s->mc && d.MemExport(&fd, s->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) != CUDA_SUCCESS) fd = -1;
      }
      for (int p = 1; p < s->world; ++p) {
        int c = connect_to(sock_name(s->tag, p), 60.0);
        if (c < 0) { set_error("symm: rank 0 cannot reach rank %d for the multicast handle", p); return 1; }
        int rc = send_fd(c, fd >= 0 ? fd : s->local_fd, 0, fd >= 0 ? 1 : 2);
        close(c);
        if (rc != 0) { set_error("symm: failed to send the multicast handle to rank %d", p); return 1; }
      }
      if (fd >= 0) close(fd);
      if (fd < 0 || !s->mc) return -1;
    } else {
      int c = accept(s->listen_fd, nullptr, nullptr);
      if (c < 0) { set_error("symm: accept (multicast) failed: %s", strerror(errno)); return 1; }
      int fd = -1, sender = -1, kind = -1;
      int rc = recv_fd(c, &fd, &sender, &kind);
      close(c);
      if (rc != 0) { set_error("symm: bad multicast handle message"); return 1; }
      if (kind != 1 || !supported) { close(fd); return -1; }
      DRV(d.MemImport(&s->mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR), "cuMemImportFromShareableHandle(multicast)");
      close(fd);
    }
    DRV(d.McAddDevice(s->mc, dev), "cuMulticastAddDevice");
    return 0;
  }
  if (phase == 1) {
    if (!s->mc) return -1;
    DRV(d.McBindMem(s->mc, 0, s->local, 0, s->bytes, 0), "cuMulticastBindMem");
    if (map_handle(d, s->mc, s->bytes, s->device, &s->mc_va) != 0) return 1;
    return 0;
  }
  if (phase == 2) {
    s->mc_ready = s->mc_va != 0;
    return s->mc_ready ? 0 : -1;
  }
  set_error("symm: unknown multicast phase %d", phase);
  return 1;
}

extern "C" void* mlt_symm_peer_ptr(void* handle, int peer) {
  Symm* s = (Symm*)handle;
  if (!s || peer < 0 || peer >= s->world) return nullptr;
  return (void*)s->peer_va[peer];
}

extern "C" void* mlt_symm_multicast_ptr(void* handle) {
  Symm* s = (Symm*)handle;
  return (s && s->mc_ready) ? (void*)s->mc_va : nullptr;
}

extern "C" long long mlt_symm_bytes(void* handle) { return handle ? (long long)((Symm*)handle)->bytes : 0; }

extern "C" int mlt_symm_destroy(void* handle) {
  Symm* s = (Symm*)handle;
  if (!s) return 0;
  Driver& d = driver();
  cudaDeviceSynchronize();
  if (s->mc_va) { d.MemUnmap(s->mc_va, s->bytes); d.MemAddressFree(s->mc_va, s->bytes); }
  if (s->mc) d.MemRelease(s->mc);
  for (int p = 0; p < s->world; ++p) {
    if (s->peer_va[p]) { d.MemUnmap(s->peer_va[p], s->bytes); d.MemAddressFree(s->peer_va[p], s->bytes); }
    if (s->peer_handles[p]) d.MemRelease(s->peer_handles[p]);
  }
  if (s->local_fd >= 0) close(s->local_fd);
  if (s->listen_fd >= 0) close(s->listen_fd);
  delete s;
  return 0;
}
