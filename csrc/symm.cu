// Symmetric memory for one-process-per-GPU jobs on an NVSwitch node (W3).
//
// The reference reaches peer GPUs only through NCCL (dist.init_process_group('nccl'),
// mnist_distributed.py:50).  Here every rank maps every peer's gradient buckets and flag
// pages into its own address space so that OUR kernels (csrc/allreduce.cu) can load/store
// peer memory directly over NVLink, and binds the same buffers to an NVLS multicast object
// so `multimem.ld_reduce` / `multimem.st` let the NVSwitch reduce and replicate:
//
//   cuMemCreate(POSIX_FD) -> export fd -> fds exchanged over AF_UNIX sockets (SCM_RIGHTS)
//   -> cuMemImportFromShareableHandle -> cuMemAddressReserve / cuMemMap / cuMemSetAccess
//   cuMulticastCreate (rank 0) -> fd to all -> cuMulticastAddDevice -> cuMulticastBindMem
//   -> map the multicast handle.
//
// All driver entry points are resolved at run time (cudaGetDriverEntryPoint), so the
// library loads on a CPU-only box.  Every call here is a collective over the group and is
// cold-path (bring-up / bucket allocation).
#include "common.h"
#include <cuda.h>

#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>
#include <poll.h>

#include <chrono>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace {

struct Drv {
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
  CUresult (*McBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                        unsigned long long);
  CUresult (*McGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGet)(CUdevice*, int);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*GetErrorString)(CUresult, const char**);
  bool ok = false;
};

template <typename F>
bool resolve(F& fn, const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
    fn = nullptr;
    return false;
  }
  fn = reinterpret_cast<F>(p);
  return true;
}

Drv& drv() {
  static Drv d;
  static bool tried = false;
  if (!tried) {
    tried = true;
    bool ok = true;
    ok &= resolve(d.MemCreate, "cuMemCreate");
    ok &= resolve(d.MemRelease, "cuMemRelease");
    ok &= resolve(d.MemExport, "cuMemExportToShareableHandle");
    ok &= resolve(d.MemImport, "cuMemImportFromShareableHandle");
    ok &= resolve(d.MemAddressReserve, "cuMemAddressReserve");
    ok &= resolve(d.MemAddressFree, "cuMemAddressFree");
    ok &= resolve(d.MemMap, "cuMemMap");
    ok &= resolve(d.MemUnmap, "cuMemUnmap");
    ok &= resolve(d.MemSetAccess, "cuMemSetAccess");
    ok &= resolve(d.MemGetGranularity, "cuMemGetAllocationGranularity");
    ok &= resolve(d.DeviceGet, "cuDeviceGet");
    ok &= resolve(d.DeviceGetAttribute, "cuDeviceGetAttribute");
    ok &= resolve(d.GetErrorString, "cuGetErrorString");
    // multicast entry points are optional (P2P-only fallback)
    resolve(d.McCreate, "cuMulticastCreate");
    resolve(d.McAddDevice, "cuMulticastAddDevice");
    resolve(d.McBindMem, "cuMulticastBindMem");
    resolve(d.McGetGranularity, "cuMulticastGetGranularity");
    d.ok = ok;
  }
  return d;
}

const char* cu_err(CUresult r) {
  const char* s = nullptr;
  if (drv().GetErrorString) drv().GetErrorString(r, &s);
  return s ? s : "unknown CUresult";
}
#define B200_CU(expr)                                                                         \
  do {                                                                                        \
    CUresult _r = (expr);                                                                     \
    if (_r != CUDA_SUCCESS) return b200_fail((int)_r, "%s failed: %s (%s:%d)", #expr, cu_err(_r), __FILE__, __LINE__); \
  } while (0)

// ----------------------------------------------------------- fd exchange ------
// Abstract-namespace AF_UNIX sockets, one listener per rank.  Lower ranks are connected to,
// higher ranks are accepted from; a unix stream socket buffers the first message before
// accept(), so the exchange cannot deadlock regardless of arrival order.
struct Comm {
  int rank = 0, world = 1, device = 0;
  std::string tag;
  int listen_fd = -1;
  std::vector<int> peer_sock;   // connected stream socket to every peer (-1 for self)
};

sockaddr_un make_addr(const std::string& name, socklen_t* len) {
  sockaddr_un a{};
  a.sun_family = AF_UNIX;
  a.sun_path[0] = '\0';
  strncpy(a.sun_path + 1, name.c_str(), sizeof(a.sun_path) - 2);
  *len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + name.size());
  return a;
}

int send_fd(int sock, int fd, int32_t tag) {
  msghdr msg{};
  iovec iov{&tag, sizeof(tag)};
  char ctl[CMSG_SPACE(sizeof(int))];
  memset(ctl, 0, sizeof(ctl));
  msg.msg_iov = &iov; msg.msg_iovlen = 1;
  if (fd >= 0) {
    msg.msg_control = ctl; msg.msg_controllen = sizeof(ctl);
    cmsghdr* c = CMSG_FIRSTHDR(&msg);
    c->cmsg_level = SOL_SOCKET; c->cmsg_type = SCM_RIGHTS; c->cmsg_len = CMSG_LEN(sizeof(int));
    memcpy(CMSG_DATA(c), &fd, sizeof(int));
  }
  return sendmsg(sock, &msg, MSG_NOSIGNAL) == (ssize_t)sizeof(tag) ? 0 : -1;
}
int recv_fd(int sock, int* fd, int32_t* tag, int timeout_ms) {
  pollfd p{sock, POLLIN, 0};
  if (poll(&p, 1, timeout_ms) <= 0) return -2;
  msghdr msg{};
  iovec iov{tag, sizeof(*tag)};
  char ctl[CMSG_SPACE(sizeof(int))];
  msg.msg_iov = &iov; msg.msg_iovlen = 1; msg.msg_control = ctl; msg.msg_controllen = sizeof(ctl);
  if (recvmsg(sock, &msg, 0) != (ssize_t)sizeof(*tag)) return -1;
  *fd = -1;
  for (cmsghdr* c = CMSG_FIRSTHDR(&msg); c; c = CMSG_NXTHDR(&msg, c))
    if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) memcpy(fd, CMSG_DATA(c), sizeof(int));
  return 0;
}

int comm_connect(Comm* c, int timeout_ms) {
  c->peer_sock.assign(c->world, -1);
  if (c->world == 1) return 0;
  socklen_t len;
  c->listen_fd = socket(AF_UNIX, SOCK_STREAM, 0);
  if (c->listen_fd < 0) return b200_fail(-1, "symm: socket(AF_UNIX) failed: %s", strerror(errno));
  sockaddr_un me = make_addr(c->tag + "_" + std::to_string(c->rank), &len);
  if (bind(c->listen_fd, (sockaddr*)&me, len) != 0 || listen(c->listen_fd, 64) != 0)
    return b200_fail(-1, "symm: cannot bind/listen on abstract socket %s_%d: %s", c->tag.c_str(), c->rank, strerror(errno));
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  for (int p = 0; p < c->rank; ++p) {          // connect to every lower rank (retry until it listens)
    for (;;) {
      int s = socket(AF_UNIX, SOCK_STREAM, 0);
      sockaddr_un a = make_addr(c->tag + "_" + std::to_string(p), &len);
      if (connect(s, (sockaddr*)&a, len) == 0) {
        int32_t hello = c->rank;
        if (send(s, &hello, sizeof(hello), MSG_NOSIGNAL) != (ssize_t)sizeof(hello)) { close(s); return b200_fail(-1, "symm: hello to rank %d failed", p); }
        c->peer_sock[p] = s;
        break;
      }
      close(s);
      if (std::chrono::steady_clock::now() > deadline)
        return b200_fail(-1, "symm: rank %d could not connect to rank %d within %d ms", c->rank, p, timeout_ms);
      std::this_thread::sleep_for(std::chrono::milliseconds(5));
    }
  }
  for (int k = c->rank + 1; k < c->world; ++k) {   // accept every higher rank, in any order
    pollfd pf{c->listen_fd, POLLIN, 0};
    int left = (int)std::chrono::duration_cast<std::chrono::milliseconds>(deadline - std::chrono::steady_clock::now()).count();
    if (left < 1 || poll(&pf, 1, left) <= 0)
      return b200_fail(-1, "symm: rank %d timed out waiting for higher ranks to connect", c->rank);
    int s = accept(c->listen_fd, nullptr, nullptr);
    int32_t who = -1;
    if (s < 0 || recv(s, &who, sizeof(who), MSG_WAITALL) != (ssize_t)sizeof(who) || who <= c->rank || who >= c->world)
      return b200_fail(-1, "symm: bad hello on rank %d", c->rank);
    c->peer_sock[who] = s;
  }
  return 0;
}

// all-to-all exchange of one fd per rank: out[p] = rank p's fd (out[rank] = my_fd, not dup'ed)
int exchange_fds(Comm* c, int my_fd, std::vector<int>& out, int timeout_ms) {
  out.assign(c->world, -1);
  out[c->rank] = my_fd;
  for (int p = 0; p < c->world; ++p)
    if (p != c->rank && send_fd(c->peer_sock[p], my_fd, c->rank) != 0)
      return b200_fail(-1, "symm: sending fd to rank %d failed: %s", p, strerror(errno));
  for (int p = 0; p < c->world; ++p) {
    if (p == c->rank) continue;
    int32_t tag = -1;
    int rc = recv_fd(c->peer_sock[p], &out[p], &tag, timeout_ms);
    if (rc != 0 || out[p] < 0 || tag != p)
      return b200_fail(-1, "symm: receiving fd from rank %d failed (rc=%d)", p, rc);
  }
  return 0;
}
// broadcast one fd from `root`
int bcast_fd(Comm* c, int root, int* fd, int32_t* status, int timeout_ms) {
  if (c->rank == root) {
    for (int p = 0; p < c->world; ++p)
      if (p != root && send_fd(c->peer_sock[p], *fd, *status) != 0)
        return b200_fail(-1, "symm: broadcasting fd to rank %d failed", p);
    return 0;
  }
  int rc = recv_fd(c->peer_sock[root], fd, status, timeout_ms);
  if (rc != 0) return b200_fail(-1, "symm: receiving broadcast fd from rank %d failed (rc=%d)", root, rc);
  return 0;
}
// socket barrier (all-to-all one int); cold path only
int sock_barrier(Comm* c, int32_t value, int32_t* min_out, int timeout_ms) {
  int32_t mn = value;
  for (int p = 0; p < c->world; ++p)
    if (p != c->rank && send(c->peer_sock[p], &value, sizeof(value), MSG_NOSIGNAL) != (ssize_t)sizeof(value))
      return b200_fail(-1, "symm: barrier send to rank %d failed", p);
  for (int p = 0; p < c->world; ++p) {
    if (p == c->rank) continue;
    pollfd pf{c->peer_sock[p], POLLIN, 0};
    int32_t v;
    if (poll(&pf, 1, timeout_ms) <= 0 || recv(c->peer_sock[p], &v, sizeof(v), MSG_WAITALL) != (ssize_t)sizeof(v))
      return b200_fail(-1, "symm: barrier recv from rank %d failed/timed out", p);
    if (v < mn) mn = v;
  }
  if (min_out) *min_out = mn;
  return 0;
}

struct Alloc {
  size_t bytes = 0;                      // padded size
  CUmemGenericAllocationHandle local = 0;
  std::vector<CUmemGenericAllocationHandle> imported;
  std::vector<CUdeviceptr> ptrs;         // [world]; ptrs[rank] = local mapping
  CUmemGenericAllocationHandle mc = 0;
  CUdeviceptr mc_ptr = 0;
  void** dev_table = nullptr;            // device copy of ptrs (for kernels)
};

}  // namespace

// ------------------------------------------------------------------ C ABI ------
B200_API void* b200_symm_comm_create(const char* tag, int rank, int world, int device, int timeout_ms) {
  if (!drv().ok) { b200_fail(-3, "symm: CUDA driver VMM entry points are not available"); return nullptr; }
  Comm* c = new Comm();
  c->rank = rank; c->world = world; c->device = device; c->tag = tag;
  if (cudaSetDevice(device) != cudaSuccess) { b200_fail(-3, "symm: cudaSetDevice(%d) failed", device); delete c; return nullptr; }
  cudaFree(0);
  if (comm_connect(c, timeout_ms) != 0) { delete c; return nullptr; }
  return c;
}

B200_API void b200_symm_comm_destroy(void* h) {
  if (!h) return;
  Comm* c = static_cast<Comm*>(h);
  for (int s : c->peer_sock) if (s >= 0) close(s);
  if (c->listen_fd >= 0) close(c->listen_fd);
  delete c;
}

// 1 if this device (and the driver) can do NVLS multicast.
B200_API int b200_symm_multicast_supported(int device) {
  if (!drv().ok || !drv().McCreate) return 0;
  CUdevice dev;
  int v = 0;
  if (drv().DeviceGet(&dev, device) != CUDA_SUCCESS) return 0;
  if (drv().DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return 0;
  return v;
}

// Collective.  Allocates `bytes` (rounded up) on every rank, maps all peers, optionally binds
// an NVLS multicast object.  Outputs: ptrs_out[world] (device VAs valid in THIS process),
// *mc_out (0 when multicast is unavailable), *padded_out.
B200_API void* b200_symm_alloc(void* comm, size_t bytes, int want_multicast, void** ptrs_out, void** mc_out,
                               size_t* padded_out, void** dev_table_out, int timeout_ms) {
  Comm* c = static_cast<Comm*>(comm);
  Drv& d = drv();
  Alloc* a = new Alloc();
  auto bail = [&](int) -> void* { delete a; return nullptr; };
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = c->device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  if (d.MemGetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || gran == 0) gran = 2u << 20;
  bool mc_ok = want_multicast && c->world > 1 && d.McCreate && b200_symm_multicast_supported(c->device);
  CUmulticastObjectProp mprop{};
  if (mc_ok) {
    mprop.numDevices = (unsigned)c->world;
    mprop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    mprop.size = bytes;
    size_t mg = 0;
    if (d.McGetGranularity(&mg, &mprop, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran) gran = mg;
  }
  a->bytes = (bytes + gran - 1) / gran * gran;
  mprop.size = a->bytes;

  auto fail_cu = [&](CUresult r, const char* what) -> void* {
    b200_fail((int)r, "symm_alloc: %s failed: %s", what, cu_err(r));
    return bail(0);
  };
  CUresult r = d.MemCreate(&a->local, a->bytes, &prop, 0);
  if (r != CUDA_SUCCESS) return fail_cu(r, "cuMemCreate");
  int my_fd = -1;
  r = d.MemExport(&my_fd, a->local, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) return fail_cu(r, "cuMemExportToShareableHandle");
  std::vector<int> fds;
  if (exchange_fds(c, my_fd, fds, timeout_ms) != 0) return bail(0);

  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = c->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  a->ptrs.assign(c->world, 0);
  a->imported.assign(c->world, 0);
  for (int p = 0; p < c->world; ++p) {
    CUmemGenericAllocationHandle h = a->local;
    if (p != c->rank) {
      r = d.MemImport(&h, (void*)(uintptr_t)fds[p], CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
      close(fds[p]);
      if (r != CUDA_SUCCESS) return fail_cu(r, "cuMemImportFromShareableHandle (is peer access / IPC permitted?)");
      a->imported[p] = h;
    }
    CUdeviceptr va = 0;
    r = d.MemAddressReserve(&va, a->bytes, gran, 0, 0);
    if (r != CUDA_SUCCESS) return fail_cu(r, "cuMemAddressReserve");
    r = d.MemMap(va, a->bytes, 0, h, 0);
    if (r != CUDA_SUCCESS) return fail_cu(r, "cuMemMap");
    r = d.MemSetAccess(va, a->bytes, &acc, 1);
    if (r != CUDA_SUCCESS) return fail_cu(r, "cuMemSetAccess (peer mapping)");
    a->ptrs[p] = va;
  }
  close(my_fd);
  if (cudaMemset((void*)a->ptrs[c->rank], 0, a->bytes) != cudaSuccess) { b200_fail(-1, "symm_alloc: memset failed"); return bail(0); }

  // ---- NVLS multicast (best effort, agreed on by all ranks) ------------------------
  int32_t status = mc_ok ? 1 : 0, agreed = 0;
  int mc_fd = -1;
  if (mc_ok && c->rank == 0) {
    r = d.McCreate(&a->mc, &mprop);
    if (r == CUDA_SUCCESS) r = d.MemExport(&mc_fd, a->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) { status = 0; a->mc = 0; mc_fd = -1; }
  }
  if (c->world > 1) {
    if (sock_barrier(c, status, &agreed, timeout_ms) != 0) return bail(0);        // everyone capable + rank0 created?
    if (agreed) {
      int32_t st = 1;
      if (bcast_fd(c, 0, &mc_fd, &st, timeout_ms) != 0) return bail(0);
      int32_t good = 1;
      if (c->rank != 0) {
        r = d.MemImport(&a->mc, (void*)(uintptr_t)mc_fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
        if (r != CUDA_SUCCESS) { good = 0; a->mc = 0; }
      }
      if (mc_fd >= 0) close(mc_fd);
      CUdevice dev;
      d.DeviceGet(&dev, c->device);
      if (good && d.McAddDevice(a->mc, dev) != CUDA_SUCCESS) good = 0;
      int32_t all_good = 0;
      if (sock_barrier(c, good, &all_good, timeout_ms) != 0) return bail(0);       // all devices added before any bind
      if (all_good) {
        if (d.McBindMem(a->mc, 0, a->local, 0, a->bytes, 0) != CUDA_SUCCESS) good = 0;
        CUdeviceptr va = 0;
        if (good && d.MemAddressReserve(&va, a->bytes, gran, 0, 0) != CUDA_SUCCESS) good = 0;
        if (good && d.MemMap(va, a->bytes, 0, a->mc, 0) != CUDA_SUCCESS) good = 0;
        if (good && d.MemSetAccess(va, a->bytes, &acc, 1) != CUDA_SUCCESS) good = 0;
        if (sock_barrier(c, good, &all_good, timeout_ms) != 0) return bail(0);
        if (all_good) a->mc_ptr = va;
      }
    } else if (mc_fd >= 0) {
      close(mc_fd);
    }
  }
  // device-side pointer table
  if (cudaMalloc((void**)&a->dev_table, sizeof(void*) * 16) != cudaSuccess) { b200_fail(-1, "symm_alloc: cudaMalloc table"); return bail(0); }
  void* host_tab[16] = {nullptr};
  for (int p = 0; p < c->world && p < 16; ++p) host_tab[p] = (void*)a->ptrs[p];
  cudaMemcpy(a->dev_table, host_tab, sizeof(host_tab), cudaMemcpyHostToDevice);
  for (int p = 0; p < c->world; ++p) ptrs_out[p] = (void*)a->ptrs[p];
  *mc_out = (void*)a->mc_ptr;
  *padded_out = a->bytes;
  if (dev_table_out) *dev_table_out = a->dev_table;
  if (c->world > 1 && sock_barrier(c, 1, nullptr, timeout_ms) != 0) return bail(0);  // memset visible everywhere
  return a;
}

B200_API void b200_symm_free(void* comm, void* alloc) {
  if (!alloc) return;
  Comm* c = static_cast<Comm*>(comm);
  Alloc* a = static_cast<Alloc*>(alloc);
  Drv& d = drv();
  cudaDeviceSynchronize();
  if (a->mc_ptr) { d.MemUnmap(a->mc_ptr, a->bytes); d.MemAddressFree(a->mc_ptr, a->bytes); }
  if (a->mc) d.MemRelease(a->mc);
  for (int p = 0; p < (int)a->ptrs.size(); ++p) {
    if (a->ptrs[p]) { d.MemUnmap(a->ptrs[p], a->bytes); d.MemAddressFree(a->ptrs[p], a->bytes); }
    if (p != c->rank && a->imported[p]) d.MemRelease(a->imported[p]);
  }
  if (a->local) d.MemRelease(a->local);
  if (a->dev_table) cudaFree(a->dev_table);
  delete a;
}

// NVLink / P2P topology summary for rank-local logging (north star: "topology discovery").
// out[p] = 1 when `device` can access device p natively (P2P); returns the device count.
B200_API int b200_symm_topology(int device, int* p2p_out, int* native_atomics_out, int cap) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  for (int p = 0; p < n && p < cap; ++p) {
    int can = (p == device), at = (p == device);
    if (p != device) {
      cudaDeviceCanAccessPeer(&can, device, p);
      cudaDeviceGetP2PAttribute(&at, cudaDevP2PAttrNativeAtomicSupported, device, p);
    }
    p2p_out[p] = can;
    if (native_atomics_out) native_atomics_out[p] = at;
  }
  return n;
}

// ---------------------------------------------------------------------------------------------
// NVLink / NVSwitch discovery through NVML (loaded with dlopen: no link-time dependency, and a box
// without the library just reports "unknown").  The reference's bring-up (test_init.py:84-91) knows nothing
// about the fabric; here the process group reports it once and uses it to choose collective algorithms:
//   links_active       NVLink links of `device` in the active state (18 on a B200 HGX baseboard)
//   links_to_switch    how many of them end in an NVSwitch (all of them on HGX; 0 on bridge-connected boards)
//   peer_links[p]      active links whose remote end is GPU p directly (bridges); 0 behind a switch
//   link_gbs           per-link, per-direction bandwidth NVML states for the link generation (GB/s), 0 if unknown
//   multicast          CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED (NVLS)
// Returns 0 on success, 1 when NVML is unavailable (the cudaDeviceCanAccessPeer view still applies).
#include <dlfcn.h>
namespace {
struct NvmlPci { char busIdLegacy[16]; unsigned int domain, bus, device, pciDeviceId, pciSubSystemId; char busId[32]; };
typedef int (*nvmlInit_t)();
typedef int (*nvmlByPci_t)(const char*, void**);
typedef int (*nvmlLinkState_t)(void*, unsigned int, unsigned int*);
typedef int (*nvmlLinkRemotePci_t)(void*, unsigned int, NvmlPci*);
typedef int (*nvmlLinkRemoteType_t)(void*, unsigned int, unsigned int*);
typedef int (*nvmlFieldValues_t)(void*, int, void*);
}  // namespace

B200_API int b200_nvlink_report(int device, int* links_active, int* links_to_switch, int* peer_links, int cap,
                                int* link_gbs, int* multicast) {
  *links_active = *links_to_switch = *link_gbs = 0;
  for (int p = 0; p < cap; ++p) peer_links[p] = 0;
  int mc = 0;
  {
    CUdevice cd;
    if (drv().DeviceGet && drv().DeviceGet(&cd, device) == CUDA_SUCCESS && drv().DeviceGetAttribute)
      drv().DeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cd);
  }
  *multicast = mc;
  void* h = dlopen("libnvidia-ml.so.1", RTLD_NOW | RTLD_LOCAL);
  if (!h) return 1;
  auto init = (nvmlInit_t)dlsym(h, "nvmlInit_v2");
  auto byPci = (nvmlByPci_t)dlsym(h, "nvmlDeviceGetHandleByPciBusId_v2");
  auto linkState = (nvmlLinkState_t)dlsym(h, "nvmlDeviceGetNvLinkState");
  auto remotePci = (nvmlLinkRemotePci_t)dlsym(h, "nvmlDeviceGetNvLinkRemotePciInfo_v2");
  auto remoteType = (nvmlLinkRemoteType_t)dlsym(h, "nvmlDeviceGetNvLinkRemoteDeviceType");
  if (!init || !byPci || !linkState || init() != 0) return 1;
  char bus[64] = {0};
  if (cudaDeviceGetPCIBusId(bus, sizeof(bus), device) != cudaSuccess) return 1;
  void* dev = nullptr;
  if (byPci(bus, &dev) != 0) return 1;
  // PCI bus ids of the other visible CUDA devices, to recognise direct GPU-GPU links
  int n = 0;
  cudaGetDeviceCount(&n);
  char peers[64][64];
  for (int p = 0; p < n && p < 64; ++p) cudaDeviceGetPCIBusId(peers[p], 64, p);
  for (unsigned int link = 0; link < 18; ++link) {
    unsigned int active = 0;
    if (linkState(dev, link, &active) != 0 || !active) continue;
    ++*links_active;
    unsigned int type = 99;
    if (remoteType && remoteType(dev, link, &type) == 0 && type == 2 /* NVML_NVLINK_DEVICE_TYPE_SWITCH */) {
      ++*links_to_switch;
      continue;
    }
    NvmlPci pci;
    if (remotePci && remotePci(dev, link, &pci) == 0)
      for (int p = 0; p < n && p < cap && p < 64; ++p)
        if (p != device && strcasecmp(pci.busId, peers[p]) == 0) ++peer_links[p];
  }
  // NVML field NVML_FI_DEV_NVLINK_SPEED_MBPS_COMMON (90): per-link speed in MB/s
  struct FieldValue { unsigned int fieldId, scopeId; long long timestamp, latencyUsec; unsigned int valueType, nvmlReturn;
                      union { double d; unsigned int ui; unsigned long ul; unsigned long long ull; long long sll; } value; };
  auto fields = (nvmlFieldValues_t)dlsym(h, "nvmlDeviceGetFieldValues");
  if (fields) {
    FieldValue fv;
    memset(&fv, 0, sizeof(fv));
    fv.fieldId = 90;                                   // NVML_FI_DEV_NVLINK_SPEED_MBPS_COMMON
    if (fields(dev, 1, &fv) == 0 && fv.nvmlReturn == 0) *link_gbs = (int)(fv.value.ui / 1000);
    if (*link_gbs == 0) {
      memset(&fv, 0, sizeof(fv));
      fv.fieldId = 164;                                // NVML_FI_DEV_NVLINK_GET_SPEED, scope = link 0
      if (fields(dev, 1, &fv) == 0 && fv.nvmlReturn == 0) *link_gbs = (int)(fv.value.ui / 1000);
    }
  }
  return 0;
}
