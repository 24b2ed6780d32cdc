// symm_heap.cc — symmetric heap over CUDA VMM (+ optional NVLS multicast) with
// a cudaIpc fallback. Native replacement for what the reference gets from
// NVSHMEM/nvshmem4py: nvshmem.core.tensor / get_peer_tensor / free_tensor
// (python/triton_dist/utils.py:114-142), nvshmemx_mc_ptr
// (backends/nvidia/language/cuda/libnvshmem_device.py:216-236) and the UID
// bootstrap (utils.py:99-110). SURVEY.md components N17/N18.
//
// Design: every rank cuMemCreate()s one physical allocation, exports it as a
// POSIX file descriptor and hands the descriptor to every other rank of the
// node over abstract-namespace UNIX sockets (SCM_RIGHTS). Each rank then
// reserves ONE virtual range of world*bytes and maps rank p's allocation at
// offset p*bytes, so peer translation is plain arithmetic. A multicast object
// bound to all allocations gives the NVLS alias used by multimem.* kernels.
#include <errno.h>
#include <fcntl.h>
#include <stddef.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

#include <string>
#include <thread>
#include <vector>

#include "../../include/tdist.h"
#include "driver.h"

namespace tdist {

struct Heap {
  int rank = 0, world = 1, device = 0;
  int mode = 0;  // 0 local cudaMalloc (world==1 fallback), 1 VMM, 2 cudaIpc
  size_t bytes = 0;
  std::string job;
  // VMM
  CUdeviceptr va = 0;
  std::vector<CUmemGenericAllocationHandle> handles;
  // multicast
  CUmemGenericAllocationHandle mc_handle = 0;
  CUdeviceptr mc_va = 0;
  bool mc_added = false, mc_mapped = false, mc_bound = false;
  // generic
  void* base[TDIST_MAX_RANKS] = {nullptr};
};

// ------------------------------------------------------------------ unix-socket fd passing
static std::string sock_name(const std::string& job, int tag, int rank) {
  return "tdist-" + job + "-" + std::to_string(tag) + "-" + std::to_string(rank);
}

static int make_listener(const std::string& name) {
  int s = socket(AF_UNIX, SOCK_STREAM, 0);
  if (s < 0) return -1;
  sockaddr_un addr;
  memset(&addr, 0, sizeof(addr));
  addr.sun_family = AF_UNIX;
  // abstract namespace: sun_path[0] == 0, no filesystem entry to clean up
  size_t n = name.size() < sizeof(addr.sun_path) - 2 ? name.size() : sizeof(addr.sun_path) - 2;
  memcpy(addr.sun_path + 1, name.data(), n);
  socklen_t len = static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + n);
  if (bind(s, reinterpret_cast<sockaddr*>(&addr), len) != 0 || listen(s, 64) != 0) {
    close(s);
    return -1;
  }
  return s;
}

static int connect_retry(const std::string& name, double timeout_s) {
  sockaddr_un addr;
  memset(&addr, 0, sizeof(addr));
  addr.sun_family = AF_UNIX;
  size_t n = name.size() < sizeof(addr.sun_path) - 2 ? name.size() : sizeof(addr.sun_path) - 2;
  memcpy(addr.sun_path + 1, name.data(), n);
  socklen_t len = static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + n);
  timespec t0;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (;;) {
    int s = socket(AF_UNIX, SOCK_STREAM, 0);
    if (s < 0) return -1;
    if (connect(s, reinterpret_cast<sockaddr*>(&addr), len) == 0) return s;
    close(s);
    timespec t1;
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if ((t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec) > timeout_s) return -1;
    usleep(2000);
  }
}

static bool send_fd(int sock, int fd, int src_rank) {
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  iovec iov;
  iov.iov_base = &src_rank;
  iov.iov_len = sizeof(src_rank);
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  return sendmsg(sock, &msg, 0) == static_cast<ssize_t>(sizeof(src_rank));
}

static bool recv_fd(int sock, int* fd, int* src_rank) {
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  iovec iov;
  iov.iov_base = src_rank;
  iov.iov_len = sizeof(int);
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  if (recvmsg(sock, &msg, MSG_WAITALL) != static_cast<ssize_t>(sizeof(int))) return false;
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (!c || c->cmsg_level != SOL_SOCKET || c->cmsg_type != SCM_RIGHTS) return false;
  memcpy(fd, CMSG_DATA(c), sizeof(int));
  return true;
}

// Receive `expect` descriptors on `listener`; slot index = sender rank.
static bool recv_many(int listener, int expect, std::vector<int>* fds, double timeout_s) {
  timeval tv;
  tv.tv_sec = static_cast<long>(timeout_s);
  tv.tv_usec = 0;
  setsockopt(listener, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
  for (int i = 0; i < expect; ++i) {
    int c = accept(listener, nullptr, nullptr);
    if (c < 0) return false;
    setsockopt(c, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
    int fd = -1, src = -1;
    bool ok = recv_fd(c, &fd, &src);
    close(c);
    if (!ok || src < 0 || src >= static_cast<int>(fds->size())) return false;
    (*fds)[src] = fd;
  }
  return true;
}

// all-gather of one fd per rank
static bool allgather_fds(const std::string& job, int tag, int rank, int world, int my_fd, std::vector<int>* out) {
  out->assign(world, -1);
  (*out)[rank] = my_fd;
  if (world == 1) return true;
  int listener = make_listener(sock_name(job, tag, rank));
  if (listener < 0) return false;
  bool recv_ok = false;
  std::thread rx([&] { recv_ok = recv_many(listener, world - 1, out, 120.0); });
  bool send_ok = true;
  for (int i = 1; i < world; ++i) {
    int peer = (rank + i) % world;
    int s = connect_retry(sock_name(job, tag, peer), 120.0);
    if (s < 0 || !send_fd(s, my_fd, rank)) send_ok = false;
    if (s >= 0) close(s);
  }
  rx.join();
  close(listener);
  return send_ok && recv_ok;
}

// root sends one fd to everybody else
static bool bcast_fd(const std::string& job, int tag, int rank, int world, int root, int* fd_inout) {
  if (world == 1) return true;
  if (rank == root) {
    bool ok = true;
    for (int p = 0; p < world; ++p) {
      if (p == root) continue;
      int s = connect_retry(sock_name(job, tag, p), 120.0);
      if (s < 0 || !send_fd(s, *fd_inout, rank)) ok = false;
      if (s >= 0) close(s);
    }
    return ok;
  }
  int listener = make_listener(sock_name(job, tag, rank));
  if (listener < 0) return false;
  std::vector<int> fds(world, -1);
  bool ok = recv_many(listener, 1, &fds, 120.0);
  close(listener);
  if (ok) *fd_inout = fds[root];
  return ok;
}

static size_t round_up(size_t x, size_t g) { return (x + g - 1) / g * g; }

static int heap_create_vmm(Heap* h) {
  const DriverApi& d = driver();
  if (!d.ok) {
    set_error("CUDA driver not available: " + d.err);
    return -1;
  }
  CUdevice dev;
  TDIST_CU_TRY(d.cuDeviceGet(&dev, h->device));
  int vmm = 0, fd_ok = 0;
  TDIST_CU_TRY(d.cuDeviceGetAttribute(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev));
  TDIST_CU_TRY(
      d.cuDeviceGetAttribute(&fd_ok, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev));
  if (!vmm || !fd_ok) {
    set_error("VMM / POSIX fd export not supported on this device");
    return -1;
  }
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = h->device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  TDIST_CU_TRY(d.cuMemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  int mc_supported = 0;
  d.cuDeviceGetAttribute(&mc_supported, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  if (mc_supported && d.cuMulticastGetGranularity) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = h->world;
    mp.size = h->bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (d.cuMulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_MINIMUM) == CUDA_SUCCESS && mg > gran)
      gran = mg;
  }
  h->bytes = round_up(h->bytes, gran);

  h->handles.assign(h->world, 0);
  TDIST_CU_TRY(d.cuMemCreate(&h->handles[h->rank], h->bytes, &prop, 0));
  int my_fd = -1;
  if (h->world > 1) {
    TDIST_CU_TRY(d.cuMemExportToShareableHandle(&my_fd, h->handles[h->rank],
                                                CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  }
  std::vector<int> fds;
  if (!allgather_fds(h->job, 0, h->rank, h->world, my_fd, &fds)) {
    set_error("symmetric heap: fd exchange over unix sockets failed");
    return -1;
  }
  for (int p = 0; p < h->world; ++p) {
    if (p == h->rank) continue;
    TDIST_CU_TRY(d.cuMemImportFromShareableHandle(&h->handles[p], reinterpret_cast<void*>(static_cast<intptr_t>(fds[p])),
                                                  CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    close(fds[p]);
  }
  if (my_fd >= 0) close(my_fd);

  TDIST_CU_TRY(d.cuMemAddressReserve(&h->va, h->bytes * h->world, gran, 0, 0));
  for (int p = 0; p < h->world; ++p) {
    TDIST_CU_TRY(d.cuMemMap(h->va + static_cast<size_t>(p) * h->bytes, h->bytes, 0, h->handles[p], 0));
    h->base[p] = reinterpret_cast<void*>(h->va + static_cast<size_t>(p) * h->bytes);
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = h->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  TDIST_CU_TRY(d.cuMemSetAccess(h->va, h->bytes * h->world, &acc, 1));
  h->mode = 1;
  return 0;
}

}  // namespace tdist

using tdist::Heap;

extern "C" {

const char* tdist_last_error(void) { return tdist::last_error(); }

static void release_vmm(Heap* h);

int tdist_heap_create(int rank, int world, int device, size_t bytes, const char* job_id, tdist_heap** out) {
  if (world < 1 || world > TDIST_MAX_RANKS || rank < 0 || rank >= world) {
    tdist::set_error("tdist_heap_create: bad rank/world");
    return -1;
  }
  TDIST_RT_TRY(cudaSetDevice(device));
  TDIST_RT_TRY(cudaFree(0));
  Heap* h = new Heap();
  h->rank = rank;
  h->world = world;
  h->device = device;
  h->bytes = bytes;
  h->job = job_id ? job_id : "job";
  if (tdist::heap_create_vmm(h) != 0) {
    release_vmm(h);
    delete h;
    return -1;
  }
  TDIST_RT_TRY(cudaMemset(h->base[rank], 0, h->bytes));
  TDIST_RT_TRY(cudaDeviceSynchronize());
  *out = reinterpret_cast<tdist_heap*>(h);
  return 0;
}

// cudaIpc fallback, step 1: allocate locally and return the 64-byte handle.
int tdist_heap_create_ipc(int rank, int world, int device, size_t bytes, tdist_heap** out, void* handle64) {
  if (world < 1 || world > TDIST_MAX_RANKS || rank < 0 || rank >= world) {
    tdist::set_error("tdist_heap_create_ipc: bad rank/world");
    return -1;
  }
  TDIST_RT_TRY(cudaSetDevice(device));
  Heap* h = new Heap();
  h->rank = rank;
  h->world = world;
  h->device = device;
  h->bytes = tdist::round_up(bytes, size_t(2) << 20);
  h->mode = world == 1 ? 0 : 2;
  void* p = nullptr;
  TDIST_RT_TRY(cudaMalloc(&p, h->bytes));
  TDIST_RT_TRY(cudaMemset(p, 0, h->bytes));
  TDIST_RT_TRY(cudaDeviceSynchronize());
  h->base[rank] = p;
  if (world > 1) {
    cudaIpcMemHandle_t ih;
    TDIST_RT_TRY(cudaIpcGetMemHandle(&ih, p));
    static_assert(sizeof(ih) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle64, &ih, 64);
  } else if (handle64) {
    memset(handle64, 0, 64);
  }
  *out = reinterpret_cast<tdist_heap*>(h);
  return 0;
}

// cudaIpc fallback, step 2: open all peers (all_handles = world * 64 bytes, rank order).
int tdist_heap_open_ipc(tdist_heap* hh, const void* all_handles) {
  Heap* h = reinterpret_cast<Heap*>(hh);
  for (int p = 0; p < h->world; ++p) {
    if (p == h->rank) continue;
    cudaIpcMemHandle_t ih;
    memcpy(&ih, static_cast<const char*>(all_handles) + 64 * p, 64);
    void* ptr = nullptr;
    TDIST_RT_TRY(cudaIpcOpenMemHandle(&ptr, ih, cudaIpcMemLazyEnablePeerAccess));
    h->base[p] = ptr;
  }
  return 0;
}

// Multicast, phase A (collective): root creates the object and ships its fd,
// everybody adds its device. Callers must barrier between phase A and B.
int tdist_heap_mc_phase_a(tdist_heap* hh) {
  Heap* h = reinterpret_cast<Heap*>(hh);
  const tdist::DriverApi& d = tdist::driver();
  if (h->mode != 1 || !d.cuMulticastCreate) {
    tdist::set_error("multicast needs the VMM heap and a multicast-capable driver");
    return -1;
  }
  CUdevice dev;
  TDIST_CU_TRY(d.cuDeviceGet(&dev, h->device));
  int mc_supported = 0;
  TDIST_CU_TRY(d.cuDeviceGetAttribute(&mc_supported, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev));
  int fd = -1;
  int create_failed = 0;
  if (h->rank == 0) {
    if (mc_supported) {
      CUmulticastObjectProp mp;
      memset(&mp, 0, sizeof(mp));
      mp.numDevices = h->world;
      mp.size = h->bytes;
      mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
      CUresult r = d.cuMulticastCreate(&h->mc_handle, &mp);
      if (r != CUDA_SUCCESS) {
        tdist::set_error("cuMulticastCreate failed: " + tdist::cu_err(r));
        create_failed = 1;
      } else if (h->world > 1) {
        r = d.cuMemExportToShareableHandle(&fd, h->mc_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
        if (r != CUDA_SUCCESS) create_failed = 1;
      }
    } else {
      create_failed = 1;
    }
    // On failure ship /dev/null so peers do not block; they detect it on import.
    if (create_failed && h->world > 1) fd = open("/dev/null", O_RDONLY);
  }
  if (!tdist::bcast_fd(h->job, 1, h->rank, h->world, 0, &fd)) {
    tdist::set_error("multicast: fd broadcast failed");
    return -1;
  }
  if (h->rank != 0) {
    if (!mc_supported) {
      close(fd);
      tdist::set_error("multicast not supported on this device");
      return -1;
    }
    CUresult r = d.cuMemImportFromShareableHandle(&h->mc_handle, reinterpret_cast<void*>(static_cast<intptr_t>(fd)),
                                                  CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    close(fd);
    if (r != CUDA_SUCCESS) {
      tdist::set_error("multicast import failed: " + tdist::cu_err(r));
      return -1;
    }
  } else {
    if (fd >= 0) close(fd);
    if (create_failed) return -1;
  }
  TDIST_CU_TRY(d.cuMulticastAddDevice(h->mc_handle, dev));
  h->mc_added = true;
  return 0;
}

int tdist_heap_mc_phase_b(tdist_heap* hh) {
  Heap* h = reinterpret_cast<Heap*>(hh);
  const tdist::DriverApi& d = tdist::driver();
  if (!h->mc_added) {
    tdist::set_error("multicast phase A did not complete");
    return -1;
  }
  TDIST_CU_TRY(d.cuMulticastBindMem(h->mc_handle, 0, h->handles[h->rank], 0, h->bytes, 0));
  h->mc_bound = true;
  TDIST_CU_TRY(d.cuMemAddressReserve(&h->mc_va, h->bytes, 0, 0, 0));
  TDIST_CU_TRY(d.cuMemMap(h->mc_va, h->bytes, 0, h->mc_handle, 0));
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = h->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  TDIST_CU_TRY(d.cuMemSetAccess(h->mc_va, h->bytes, &acc, 1));
  h->mc_mapped = true;
  return 0;
}

void* tdist_heap_base(tdist_heap* hh, int peer) { return reinterpret_cast<Heap*>(hh)->base[peer]; }
void* tdist_heap_mc_base(tdist_heap* hh) {
  Heap* h = reinterpret_cast<Heap*>(hh);
  return h->mc_mapped ? reinterpret_cast<void*>(h->mc_va) : nullptr;
}
size_t tdist_heap_bytes(tdist_heap* hh) { return reinterpret_cast<Heap*>(hh)->bytes; }
int tdist_heap_mode(tdist_heap* hh) { return reinterpret_cast<Heap*>(hh)->mode; }

int tdist_heap_fill_ctx(tdist_heap* hh, tdist_symm_ctx* out) {
  Heap* h = reinterpret_cast<Heap*>(hh);
  memset(out, 0, sizeof(*out));
  for (int p = 0; p < h->world; ++p) out->base[p] = static_cast<char*>(h->base[p]);
  out->mc_base = h->mc_mapped ? reinterpret_cast<char*>(h->mc_va) : nullptr;
  out->bytes = h->bytes;
  out->rank = h->rank;
  out->world = h->world;
  return 0;
}

// Undo whatever part of a VMM heap exists (also called for a heap whose creation failed half way: nothing may stay
// mapped or pinned, because the caller then falls back to a second, cudaIpc-backed heap of the same size).
static void release_vmm(Heap* h) {
  const tdist::DriverApi& d = tdist::driver();
  if (!d.ok) return;
  if (h->mc_mapped) {
    d.cuMemUnmap(h->mc_va, h->bytes);
    d.cuMemAddressFree(h->mc_va, h->bytes);
    h->mc_mapped = false;
  }
  if (h->mc_handle) {
    // unbind this device's memory from the multicast object before the handles go away
    if (h->mc_bound && d.cuMulticastUnbind) d.cuMulticastUnbind(h->mc_handle, h->device, 0, h->bytes);
    d.cuMemRelease(h->mc_handle);
    h->mc_handle = 0;
  }
  for (int p = 0; p < static_cast<int>(h->handles.size()); ++p) {
    if (h->va && h->base[p]) d.cuMemUnmap(h->va + static_cast<size_t>(p) * h->bytes, h->bytes);
    if (h->handles[p]) d.cuMemRelease(h->handles[p]);
    h->handles[p] = 0;
    h->base[p] = nullptr;
  }
  if (h->va) d.cuMemAddressFree(h->va, h->bytes * h->world);
  h->va = 0;
}

int tdist_heap_destroy(tdist_heap* hh) {
  Heap* h = reinterpret_cast<Heap*>(hh);
  cudaDeviceSynchronize();
  if (h->mode == 1) {
    release_vmm(h);
  } else {
    for (int p = 0; p < h->world; ++p) {
      if (p == h->rank) continue;
      if (h->base[p]) cudaIpcCloseMemHandle(h->base[p]);
    }
    if (h->base[h->rank]) cudaFree(h->base[h->rank]);
  }
  delete h;
  return 0;
}

// ------------------------------------------------------------------ stream memory ops (N20)
// cuStreamWriteValue32/WaitValue32 equivalents of the reference's
// kernels/nvidia/common_ops.py:347-389 (set_signal / wait_eq).
int tdist_stream_write_value32(void* stream, void* addr, unsigned int value) {
  const tdist::DriverApi& d = tdist::driver();
  if (!d.ok) return -1;
  TDIST_CU_TRY(d.cuStreamWriteValue32(static_cast<CUstream>(stream), reinterpret_cast<CUdeviceptr>(addr), value, 0));
  return 0;
}
int tdist_stream_wait_value32(void* stream, void* addr, unsigned int value, int cmp_geq) {
  const tdist::DriverApi& d = tdist::driver();
  if (!d.ok) return -1;
  TDIST_CU_TRY(d.cuStreamWaitValue32(static_cast<CUstream>(stream), reinterpret_cast<CUdeviceptr>(addr), value,
                                     cmp_geq ? CU_STREAM_WAIT_VALUE_GEQ : CU_STREAM_WAIT_VALUE_EQ));
  return 0;
}
int tdist_stream_write_value64(void* stream, void* addr, unsigned long long value) {
  const tdist::DriverApi& d = tdist::driver();
  if (!d.ok) return -1;
  TDIST_CU_TRY(d.cuStreamWriteValue64(static_cast<CUstream>(stream), reinterpret_cast<CUdeviceptr>(addr), value, 0));
  return 0;
}
int tdist_stream_wait_value64(void* stream, void* addr, unsigned long long value, int cmp_geq) {
  const tdist::DriverApi& d = tdist::driver();
  if (!d.ok) return -1;
  TDIST_CU_TRY(d.cuStreamWaitValue64(static_cast<CUstream>(stream), reinterpret_cast<CUdeviceptr>(addr), value,
                                     cmp_geq ? CU_STREAM_WAIT_VALUE_GEQ : CU_STREAM_WAIT_VALUE_EQ));
  return 0;
}

}  // extern "C"
