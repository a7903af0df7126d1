// probe_exchange.cu -- symmetric segments between GPUs: the plumbing under the fused
// checksum + lanes exchange (exchange_lanes in probe_common.cuh) and under K4, the probe's
// own allreduce (probe_allreduce.cu).
//
// each GPU checksums its shard, then the 16-byte lanes are summed
// over GPUs.  The reference-style way is a separate ncclAllReduce(count=2, ncclUint64)
// launch after K2.  Here the last CTA of K2 (or of the fused K1+K2) writes its lanes into a
// mailbox slot on every peer over NVLink and adds up the slots that land in its own
// mailbox: compute and collective are ONE kernel.
//
// A segment is one physical allocation per rank (header + staging + payload, layout in
// probe_common.cuh) made with the driver's virtual-memory API, because that is the one kind
// of device memory that can BOTH be handed to another process (as a file descriptor) AND be
// bound into an NVSwitch multicast object (NVLS: multimem.ld_reduce / multimem.st, which let
// the switch do the reduction and the broadcast).  How ranks learn each other's segments:
//   * ranks in ONE process (the pod's C host): gpu_probe_exchange_connect_local -- peer
//     access is granted on the existing mappings, the multicast object is built in place;
//   * one process per GPU (bench.py under torchrun): gpu_probe_exchange_export gives 64
//     opaque bytes naming a unix datagram socket; the host carries all ranks' bytes to every
//     rank by any channel (torch.distributed all_gather in bench.py) and
//     gpu_probe_exchange_connect_ipc trades the file descriptors over those sockets
//     (SCM_RIGHTS), including the multicast object rank 0 creates.
// Multicast is an optimisation, never a requirement: if the device, the fabric or any rank
// cannot do it, the segments still connect and the P2P algorithms run.
// The driver entry points are looked up at run time (cudaGetDriverEntryPoint), so the
// library still loads on a machine without libcuda (the CPU-side ABI tests).
// No reference counterpart (slot: /root/reference/pods/nvidia-gpu-test-pod.yaml:9).
#include "gpu_probe.h"

#include <cerrno>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda.h>
#include <cuda_runtime.h>
#include <atomic>
#include <mutex>
#include <new>
#include <sys/random.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <sys/un.h>
#include <unistd.h>

#include "probe_common.cuh"
#include "probe_exchange.h"
#include "probe_launch.h"

using namespace probe;

static_assert(GPU_PROBE_EXCHANGE_MAX_RANKS == kMaxRanks, "header and device code disagree on the rank limit");

// ---------------------------------------------------------------------------
// driver API, resolved lazily
// ---------------------------------------------------------------------------
namespace {

struct DriverApi {
  int rc = -1;  // 0 once every symbol resolved
  CUresult (*DeviceGet)(CUdevice *, int);
  CUresult (*DeviceGetAttribute)(int *, CUdevice_attribute, CUdevice);
  CUresult (*MemCreate)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t);
  CUresult (*MemGetAllocationGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags);
  CUresult (*MemExportToShareableHandle)(void *, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
                                         unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle *, void *, CUmemAllocationHandleType);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle *, const CUmulticastObjectProp *);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                               unsigned long long);
  CUresult (*MulticastGetGranularity)(size_t *, const CUmulticastObjectProp *, CUmulticastGranularity_flags);
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
};

DriverApi g_drv;
std::mutex g_drv_mutex;

template <typename F>
bool resolve(F *slot, const char *name) {
  void *fn = nullptr;
  cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
  if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
    cudaGetLastError();
    return false;
  }
  *slot = reinterpret_cast<F>(fn);
  return true;
}

// 0, or a CUDA error when there is no usable driver (e.g. a machine without a GPU).
int driver(const DriverApi **out) {
  std::lock_guard<std::mutex> lock(g_drv_mutex);  // set-up paths only; kernels are launched elsewhere
  if (g_drv.rc != 0) {
    cudaError_t e = cudaFree(nullptr);  // brings the runtime (and the driver under it) up on the current device
    if (e != cudaSuccess) return int(e);
    DriverApi d;
    bool ok = resolve(&d.DeviceGet, "cuDeviceGet") && resolve(&d.DeviceGetAttribute, "cuDeviceGetAttribute") &&
              resolve(&d.MemCreate, "cuMemCreate") && resolve(&d.MemRelease, "cuMemRelease") &&
              resolve(&d.MemAddressReserve, "cuMemAddressReserve") && resolve(&d.MemAddressFree, "cuMemAddressFree") &&
              resolve(&d.MemMap, "cuMemMap") && resolve(&d.MemUnmap, "cuMemUnmap") &&
              resolve(&d.MemSetAccess, "cuMemSetAccess") &&
              resolve(&d.MemGetAllocationGranularity, "cuMemGetAllocationGranularity") &&
              resolve(&d.MemExportToShareableHandle, "cuMemExportToShareableHandle") &&
              resolve(&d.MemImportFromShareableHandle, "cuMemImportFromShareableHandle") &&
              resolve(&d.MulticastCreate, "cuMulticastCreate") && resolve(&d.MulticastAddDevice, "cuMulticastAddDevice") &&
              resolve(&d.MulticastBindMem, "cuMulticastBindMem") &&
              resolve(&d.MulticastGetGranularity, "cuMulticastGetGranularity") &&
              resolve(&d.MulticastUnbind, "cuMulticastUnbind");
    if (!ok) return int(cudaErrorNotSupported);
    d.rc = 0;
    g_drv = d;
  }
  *out = &g_drv;
  return 0;
}

inline int cu(CUresult r) { return int(r); }  // CUresult and cudaError_t share their numbering

CUmemAllocationProp segment_prop(int dev) {
  CUmemAllocationProp p;
  std::memset(&p, 0, sizeof(p));
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = dev;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

// Map `mem` (seg_bytes) into this process and let `devs[0..ndev)` read and write it.
int map_segment(const DriverApi *d, CUmemGenericAllocationHandle mem, size_t bytes, const int *devs, int ndev,
                CUdeviceptr *va_out) {
  CUdeviceptr va = 0;
  CUresult r = d->MemAddressReserve(&va, bytes, 2u << 20, 0, 0);
  if (r != CUDA_SUCCESS) return cu(r);
  r = d->MemMap(va, bytes, 0, mem, 0);
  if (r != CUDA_SUCCESS) { d->MemAddressFree(va, bytes); return cu(r); }
  CUmemAccessDesc ad[kMaxRanks];
  for (int i = 0; i < ndev; ++i) {
    std::memset(&ad[i], 0, sizeof(ad[i]));
    ad[i].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ad[i].location.id = devs[i];
    ad[i].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  }
  r = d->MemSetAccess(va, bytes, ad, size_t(ndev));
  if (r != CUDA_SUCCESS) { d->MemUnmap(va, bytes); d->MemAddressFree(va, bytes); return cu(r); }
  *va_out = va;
  return 0;
}

void unmap_segment(const DriverApi *d, void *va, size_t bytes) {
  if (!va) return;
  d->MemUnmap(reinterpret_cast<CUdeviceptr>(va), bytes);
  d->MemAddressFree(reinterpret_cast<CUdeviceptr>(va), bytes);
}

// ---------------------------------------------------------------------------
// unix datagram sockets (abstract namespace): file descriptors between processes
// ---------------------------------------------------------------------------
enum MsgType : int32_t { MSG_MEM = 1, MSG_MC = 2, MSG_MC_NONE = 3, MSG_ADDED = 4, MSG_ADD_FAILED = 5, MSG_BOUND = 6,
                         MSG_BIND_FAILED = 7 };
// `token` is the RECEIVER's secret from its exported handle: only a process that was given the
// handles over the host's own channel can name it.  The kernel also tells the receiver who sent
// each datagram (SO_PASSCRED), so a message is taken only from the user and -- when the ranks
// share a PID namespace -- the very process that exported handle `rank`.
struct Msg { uint32_t magic; int32_t type; int32_t rank; int32_t pad; uint64_t token; };
constexpr uint32_t kMsgMagic = 0x47505845u;

socklen_t abstract_addr(sockaddr_un *a, const char *name) {
  std::memset(a, 0, sizeof(*a));
  a->sun_family = AF_UNIX;
  const size_t n = std::strlen(name);
  std::memcpy(a->sun_path + 1, name, n);  // sun_path[0] == 0: abstract namespace, vanishes with the socket
  return socklen_t(offsetof(sockaddr_un, sun_path) + 1 + n);
}

int send_msg(int sock, const char *to, uint64_t to_token, int type, int rank, int fd) {
  Msg m{kMsgMagic, type, rank, 0, to_token};
  sockaddr_un addr;
  const socklen_t alen = abstract_addr(&addr, to);
  iovec iov{&m, sizeof(m)};
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
  msghdr h;
  std::memset(&h, 0, sizeof(h));
  h.msg_name = &addr;
  h.msg_namelen = alen;
  h.msg_iov = &iov;
  h.msg_iovlen = 1;
  if (fd >= 0) {
    std::memset(ctrl, 0, sizeof(ctrl));
    h.msg_control = ctrl;
    h.msg_controllen = sizeof(ctrl);
    cmsghdr *c = CMSG_FIRSTHDR(&h);
    c->cmsg_level = SOL_SOCKET;
    c->cmsg_type = SCM_RIGHTS;
    c->cmsg_len = CMSG_LEN(sizeof(int));
    std::memcpy(CMSG_DATA(c), &fd, sizeof(int));
  }
  for (int tries = 0; tries < 200; ++tries) {  // the peer's socket exists (it exported), but may be briefly full
    if (sendmsg(sock, &h, 0) == ssize_t(sizeof(m))) return 0;
    if (errno != EAGAIN && errno != ENOBUFS && errno != ECONNREFUSED && errno != EINTR) break;
    usleep(10000);
  }
  return int(cudaErrorSystemNotReady);
}

int recv_msg(int sock, Msg *m, int *fd, pid_t *from_pid, uid_t *from_uid) {
  *fd = -1;
  *from_pid = 0;
  *from_uid = uid_t(-1);
  iovec iov{m, sizeof(*m)};
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int)) + CMSG_SPACE(sizeof(ucred))];
  msghdr h;
  std::memset(&h, 0, sizeof(h));
  h.msg_iov = &iov;
  h.msg_iovlen = 1;
  h.msg_control = ctrl;
  h.msg_controllen = sizeof(ctrl);
  ssize_t n;
  do { n = recvmsg(sock, &h, 0); } while (n < 0 && errno == EINTR);
  if (n < 0) return GPU_PROBE_ETIMEDOUT;  // SO_RCVTIMEO ran out
  for (cmsghdr *c = CMSG_FIRSTHDR(&h); c; c = CMSG_NXTHDR(&h, c)) {
    if (c->cmsg_level != SOL_SOCKET) continue;
    if (c->cmsg_type == SCM_RIGHTS) std::memcpy(fd, CMSG_DATA(c), sizeof(int));
    if (c->cmsg_type == SCM_CREDENTIALS) {
      ucred u;
      std::memcpy(&u, CMSG_DATA(c), sizeof(u));
      *from_pid = u.pid;
      *from_uid = u.uid;
    }
  }
  if (n != ssize_t(sizeof(*m)) || m->magic != kMsgMagic) m->type = 0;  // not one of ours: the caller drops it (and its fd)
  return 0;
}

struct WireHandle {  // the 64 opaque bytes of gpu_probe_exchange_export
  uint32_t magic;
  int32_t rank, mc_supported, pid;
  uint64_t seg_bytes;
  uint64_t token;  // random per exchange; every message TO this rank must carry it
  char name[32];
};
static_assert(sizeof(WireHandle) == GPU_PROBE_EXCHANGE_HANDLE_BYTES, "wire handle must fill the 64 bytes exactly");

}  // namespace

// ---------------------------------------------------------------------------
// create / destroy
// ---------------------------------------------------------------------------
extern "C" int gpu_probe_exchange_create(gpu_probe_exchange **out, int rank, int world, uint64_t buffer_bytes) {
  if (!out) return GPU_PROBE_EINVAL;
  *out = nullptr;
  if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world) return GPU_PROBE_EINVAL;
  if (buffer_bytes > (1ull << 36)) return GPU_PROBE_ETOOBIG;
  buffer_bytes = (buffer_bytes + 255) & ~uint64_t(255);
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return int(e);
  const DriverApi *d = nullptr;
  int rc = driver(&d);
  if (rc) return rc;
  gpu_probe_exchange *x = new (std::nothrow) gpu_probe_exchange();
  if (!x) return int(cudaErrorMemoryAllocation);
  x->device = dev; x->rank = rank; x->world = world; x->sock = -1;
  x->timeout_ns = 10ull * 1000 * 1000 * 1000;  // a peer that is 10 s late is not coming
  x->buffer_bytes = buffer_bytes;
  x->stage_bytes = buffer_bytes < kMaxStageBytes ? buffer_bytes : kMaxStageBytes;
  x->payload_offset = kHeaderBytes + 2 * uint64_t(world) * x->stage_bytes;

  CUdevice cudev = 0;
  int mc = 0;
  CUresult r = d->DeviceGet(&cudev, dev);
  if (r == CUDA_SUCCESS) d->DeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cudev);
  const CUmemAllocationProp prop = segment_prop(dev);
  size_t gran = 0;
  if (r == CUDA_SUCCESS) r = d->MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM);
  if (r == CUDA_SUCCESS && mc) {  // the whole segment gets bound into the multicast object: honour its granularity too
    CUmulticastObjectProp mp;
    std::memset(&mp, 0, sizeof(mp));
    mp.numDevices = unsigned(world > 1 ? world : 2);
    mp.size = 2u << 20;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (d->MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_MINIMUM) == CUDA_SUCCESS && mg > 0) {
      if (mg > gran) gran = mg;
    } else {
      mc = 0;
    }
  }
  x->mc_supported = mc != 0 && getenv("GPU_PROBE_NO_MULTICAST") == nullptr;  // dev knob: plain peer mappings only
  if (r == CUDA_SUCCESS) {
    if (gran == 0) gran = 2u << 20;
    x->seg_bytes = ((x->payload_offset + buffer_bytes + gran - 1) / gran) * gran;
    r = d->MemCreate(&x->mem, x->seg_bytes, &prop, 0);
  }
  if (r != CUDA_SUCCESS) { delete x; return cu(r); }
  CUdeviceptr va = 0;
  rc = map_segment(d, x->mem, x->seg_bytes, &dev, 1, &va);
  if (rc) { d->MemRelease(x->mem); delete x; return rc; }
  x->mailbox = reinterpret_cast<uint64_t *>(va);
  e = cudaMemset(x->mailbox, 0, kHeaderBytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();  // zeroed before any peer can learn the address
  if (e != cudaSuccess) {
    unmap_segment(d, x->mailbox, x->seg_bytes);
    d->MemRelease(x->mem);
    delete x;
    return int(e);
  }
  x->peer[rank] = x->mailbox;
  x->connected = (world == 1);
  x->magic = kXMagic;
  *out = x;
  return 0;
}

extern "C" int gpu_probe_exchange_destroy(gpu_probe_exchange *x) {
  if (!live(x)) return GPU_PROBE_ENOTREADY;
  const DriverApi *d = nullptr;
  if (driver(&d)) return GPU_PROBE_ENOTREADY;
  DeviceGuard g(x->device);
  cudaDeviceSynchronize();
  if (x->mc_enabled) {
    CUdevice cudev = 0;
    if (d->DeviceGet(&cudev, x->device) == CUDA_SUCCESS) d->MulticastUnbind(x->mc_mem, cudev, 0, x->seg_bytes);
    if (x->mc_refs && --*x->mc_refs == 0) {  // one mapping per process: the last object sharing it takes it down
      unmap_segment(d, x->mc_base, x->seg_bytes);
      d->MemRelease(x->mc_mem);
      delete x->mc_refs;
    }
  }
  for (int p = 0; p < x->world; ++p)
    if (x->peer_imported[p]) {
      unmap_segment(d, x->peer[p], x->seg_bytes);
      d->MemRelease(x->peer_mem[p]);
    }
  unmap_segment(d, x->mailbox, x->seg_bytes);
  d->MemRelease(x->mem);
  if (x->sock >= 0) close(x->sock);
  x->magic = 0;
  delete x;
  return 0;
}

extern "C" int gpu_probe_exchange_set_timeout_ms(gpu_probe_exchange *x, uint64_t ms) {
  if (!live(x)) return GPU_PROBE_ENOTREADY;
  if (ms == 0 || ms > 600000) return GPU_PROBE_EINVAL;
  x->timeout_ns = ms * 1000000ull;
  return 0;
}

// ---------------------------------------------------------------------------
// ranks in one process
// ---------------------------------------------------------------------------
extern "C" int gpu_probe_exchange_connect_local(gpu_probe_exchange *const *xs, int world) {
  if (!xs || world < 1 || world > kMaxRanks) return GPU_PROBE_EINVAL;
  for (int r = 0; r < world; ++r)
    if (!live(xs[r])) return GPU_PROBE_ENOTREADY;
  const DriverApi *d = nullptr;
  int rc = driver(&d);
  if (rc) return rc;
  int devs[kMaxRanks], ndev = 0;
  bool distinct = true, all_mc = true;
  for (int r = 0; r < world; ++r) {
    if (xs[r]->world != world || xs[r]->rank != r || xs[r]->seg_bytes != xs[0]->seg_bytes) return GPU_PROBE_EINVAL;
    bool seen = false;
    for (int i = 0; i < ndev; ++i) seen = seen || devs[i] == xs[r]->device;
    if (seen) distinct = false; else devs[ndev++] = xs[r]->device;
    all_mc = all_mc && xs[r]->mc_supported;
  }
  for (int i = 0; i < ndev; ++i)
    for (int j = 0; j < ndev; ++j) {
      if (i == j) continue;
      int can = 0;
      cudaError_t e = cudaDeviceCanAccessPeer(&can, devs[i], devs[j]);
      if (e != cudaSuccess) return int(e);
      if (!can) return int(cudaErrorPeerAccessUnsupported);
    }
  // every device of the group may read and write every segment, at the address it already has
  CUmemAccessDesc ad[kMaxRanks];
  for (int i = 0; i < ndev; ++i) {
    std::memset(&ad[i], 0, sizeof(ad[i]));
    ad[i].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ad[i].location.id = devs[i];
    ad[i].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  }
  for (int r = 0; r < world; ++r) {
    CUresult cr = d->MemSetAccess(reinterpret_cast<CUdeviceptr>(xs[r]->mailbox), xs[r]->seg_bytes, ad, size_t(ndev));
    if (cr != CUDA_SUCCESS) return cu(cr);
  }
  for (int r = 0; r < world; ++r) {
    for (int p = 0; p < world; ++p) xs[r]->peer[p] = xs[p]->mailbox;
    xs[r]->connected = true;
  }
  // NVSwitch multicast over the same physical segments (needs one GPU per rank)
  if (world >= 2 && distinct && all_mc) {
    CUmulticastObjectProp mp;
    std::memset(&mp, 0, sizeof(mp));
    mp.numDevices = unsigned(world);
    mp.size = xs[0]->seg_bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUmemGenericAllocationHandle mch;
    if (d->MulticastCreate(&mch, &mp) == CUDA_SUCCESS) {
      bool ok = true;
      for (int r = 0; r < world && ok; ++r) {
        CUdevice cudev = 0;
        ok = d->DeviceGet(&cudev, xs[r]->device) == CUDA_SUCCESS && d->MulticastAddDevice(mch, cudev) == CUDA_SUCCESS;
      }
      int bound = 0;
      for (int r = 0; r < world && ok; ++r) {
        ok = d->MulticastBindMem(mch, 0, xs[r]->mem, 0, xs[r]->seg_bytes, 0) == CUDA_SUCCESS;
        if (ok) ++bound;
      }
      CUdeviceptr va = 0;
      if (ok) ok = map_segment(d, mch, xs[0]->seg_bytes, devs, ndev, &va) == 0;
      int *refs = ok ? new (std::nothrow) int(world) : nullptr;
      if (ok && !refs) { unmap_segment(d, reinterpret_cast<void *>(va), xs[0]->seg_bytes); ok = false; }
      if (ok) {
        for (int r = 0; r < world; ++r) {
          xs[r]->mc_mem = mch;
          xs[r]->mc_base = reinterpret_cast<unsigned char *>(va);
          xs[r]->mc_enabled = true;
          xs[r]->mc_refs = refs;
        }
      } else {
        for (int r = 0; r < bound; ++r) {
          CUdevice cudev = 0;
          if (d->DeviceGet(&cudev, xs[r]->device) == CUDA_SUCCESS) d->MulticastUnbind(mch, cudev, 0, xs[r]->seg_bytes);
        }
        d->MemRelease(mch);
      }
    }
  }
  return 0;
}

// ---------------------------------------------------------------------------
// one process per GPU
// ---------------------------------------------------------------------------
extern "C" int gpu_probe_exchange_export(gpu_probe_exchange *x, unsigned char handle[GPU_PROBE_EXCHANGE_HANDLE_BYTES]) {
  if (!live(x)) return GPU_PROBE_ENOTREADY;
  if (!handle) return GPU_PROBE_EINVAL;
  if (x->sock < 0) {
    static std::atomic<int> counter{0};
    int s = socket(AF_UNIX, SOCK_DGRAM | SOCK_CLOEXEC, 0);
    if (s < 0) return int(cudaErrorSystemNotReady);
    std::snprintf(x->sock_name, sizeof(x->sock_name), "gpuprobe-%d-%d-r%d", int(getpid()), ++counter, x->rank);
    sockaddr_un addr;
    const socklen_t alen = abstract_addr(&addr, x->sock_name);
    timeval tv{30, 0};  // nobody waits for a lost rank for ever
    const int on = 1;
    uint64_t token = 0;
    if (getrandom(&token, sizeof(token), 0) != ssize_t(sizeof(token)) ||
        bind(s, reinterpret_cast<sockaddr *>(&addr), alen) != 0 ||
        setsockopt(s, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv)) != 0 ||
        setsockopt(s, SOL_SOCKET, SO_PASSCRED, &on, sizeof(on)) != 0) {
      close(s);
      return int(cudaErrorSystemNotReady);
    }
    x->sock = s;
    x->token = token;
  }
  WireHandle w;
  std::memset(&w, 0, sizeof(w));
  w.magic = kXMagic;
  w.rank = x->rank;
  w.mc_supported = x->mc_supported ? 1 : 0;
  w.pid = int(getpid());
  w.seg_bytes = x->seg_bytes;
  w.token = x->token;
  std::memcpy(w.name, x->sock_name, sizeof(w.name));
  std::memcpy(handle, &w, sizeof(w));
  return 0;
}

extern "C" int gpu_probe_exchange_connect_ipc(gpu_probe_exchange *x, const unsigned char *handles) {
  if (!live(x)) return GPU_PROBE_ENOTREADY;
  if (!handles) return GPU_PROBE_EINVAL;
  if (x->connected) return 0;
  if (x->sock < 0) return GPU_PROBE_ENOTREADY;  // export first: that is what opens this rank's socket
  const DriverApi *d = nullptr;
  int rc = driver(&d);
  if (rc) return rc;
  DeviceGuard g(x->device);
  const int world = x->world, me = x->rank;
  WireHandle w[kMaxRanks];
  bool want_mc = world >= 2;
  for (int p = 0; p < world; ++p) {
    std::memcpy(&w[p], handles + size_t(p) * sizeof(WireHandle), sizeof(WireHandle));
    if (w[p].magic != kXMagic || w[p].rank != p || w[p].seg_bytes != x->seg_bytes) return GPU_PROBE_EINVAL;
    w[p].name[sizeof(w[p].name) - 1] = 0;
    want_mc = want_mc && w[p].mc_supported;
  }
  CUdevice cudev = 0;
  if (d->DeviceGet(&cudev, x->device) != CUDA_SUCCESS) return int(cudaErrorInvalidDevice);

  // 1. my segment's file descriptor to every peer; rank 0 also makes and hands out the multicast object
  int myfd = -1;
  CUresult r = d->MemExportToShareableHandle(&myfd, x->mem, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) return cu(r);
  for (int p = 0; p < world && rc == 0; ++p)
    if (p != me) rc = send_msg(x->sock, w[p].name, w[p].token, MSG_MEM, me, myfd);
  close(myfd);
  if (rc) return rc;
  bool have_mc = false;
  CUmemGenericAllocationHandle mch = 0;
  if (want_mc && me == 0) {
    CUmulticastObjectProp mp;
    std::memset(&mp, 0, sizeof(mp));
    mp.numDevices = unsigned(world);
    mp.size = x->seg_bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    int mcfd = -1;
    have_mc = d->MulticastCreate(&mch, &mp) == CUDA_SUCCESS;
    if (have_mc && d->MemExportToShareableHandle(&mcfd, mch, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) != CUDA_SUCCESS) {
      d->MemRelease(mch);
      have_mc = false;
    }
    for (int p = 1; p < world && rc == 0; ++p)
      rc = send_msg(x->sock, w[p].name, w[p].token, have_mc ? MSG_MC : MSG_MC_NONE, me, have_mc ? mcfd : -1);
    if (mcfd >= 0) close(mcfd);
    if (rc) return rc;
  }

  // 2. whatever arrives, in whatever order: peers may already be a phase ahead of this rank
  int mem_fd[kMaxRanks], mc_fd = -1;
  for (int p = 0; p < world; ++p) mem_fd[p] = -1;
  int n_mem = 0, n_added = 0, n_add_failed = 0, n_bound = 0, n_bind_failed = 0;
  bool mc_decided = !want_mc || me == 0;
  auto pump = [&]() -> int {  // one message
    Msg m;
    int fd = -1;
    pid_t from_pid = 0;
    uid_t from_uid = uid_t(-1);
    int prc = recv_msg(x->sock, &m, &fd, &from_pid, &from_uid);
    if (prc) return prc;
    // Only the peers this rank was told about may speak: right secret, a rank of this exchange, this
    // user, and the exporting process itself (pid 0: the sender lives in another PID namespace, which
    // the kernel cannot name here -- the secret and the uid still have to match).
    const bool known = m.type != 0 && m.token == x->token && m.rank >= 0 && m.rank < world && m.rank != me &&
                       from_uid == geteuid() && (from_pid == 0 || from_pid == pid_t(w[m.rank].pid));
    if (!known) {
      if (fd >= 0) close(fd);
      return 0;  // a stranger's datagram: dropped, never acted on; the wait for the real peers goes on
    }
    switch (m.type) {
      case MSG_MEM:
        if (m.rank >= 0 && m.rank < world && m.rank != me && mem_fd[m.rank] < 0 && fd >= 0) { mem_fd[m.rank] = fd; ++n_mem; fd = -1; }
        break;
      case MSG_MC: if (m.rank != 0 || mc_decided) break; if (fd >= 0) { mc_fd = fd; fd = -1; have_mc = true; } mc_decided = true; break;  // rank 0 owns the multicast object
      case MSG_MC_NONE: if (m.rank == 0) mc_decided = true; break;
      case MSG_ADDED: ++n_added; break;
      case MSG_ADD_FAILED: ++n_add_failed; break;
      case MSG_BOUND: ++n_bound; break;
      case MSG_BIND_FAILED: ++n_bind_failed; break;
    }
    if (fd >= 0) close(fd);
    return 0;
  };
  while (rc == 0 && (n_mem < world - 1 || !mc_decided)) rc = pump();
  for (int p = 0; p < world && rc == 0; ++p) {
    if (p == me) continue;
    CUmemGenericAllocationHandle h;
    r = d->MemImportFromShareableHandle(&h, reinterpret_cast<void *>(uintptr_t(mem_fd[p])), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (r != CUDA_SUCCESS) { rc = cu(r); break; }
    CUdeviceptr va = 0;
    rc = map_segment(d, h, x->seg_bytes, &x->device, 1, &va);
    if (rc) { d->MemRelease(h); break; }
    x->peer_mem[p] = h;
    x->peer[p] = reinterpret_cast<uint64_t *>(va);
    x->peer_imported[p] = true;
  }
  for (int p = 0; p < world; ++p)
    if (mem_fd[p] >= 0) close(mem_fd[p]);
  if (rc) { if (mc_fd >= 0) close(mc_fd); return rc; }

  // 3. multicast: every device joins, THEN every rank binds its segment, THEN anyone may use it
  if (want_mc) {
    bool ok = have_mc;
    if (ok && me != 0)
      ok = d->MemImportFromShareableHandle(&mch, reinterpret_cast<void *>(uintptr_t(mc_fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) == CUDA_SUCCESS;
    if (mc_fd >= 0) close(mc_fd);
    const bool holding = ok;
    if (ok) ok = d->MulticastAddDevice(mch, cudev) == CUDA_SUCCESS;
    for (int p = 0; p < world && rc == 0; ++p)
      if (p != me) rc = send_msg(x->sock, w[p].name, w[p].token, ok ? MSG_ADDED : MSG_ADD_FAILED, me, -1);
    while (rc == 0 && n_added + n_add_failed < world - 1) rc = pump();
    bool all = ok && n_add_failed == 0 && rc == 0;
    bool bound = false;
    CUdeviceptr va = 0;
    if (all) {
      bound = d->MulticastBindMem(mch, 0, x->mem, 0, x->seg_bytes, 0) == CUDA_SUCCESS;
      all = bound && map_segment(d, mch, x->seg_bytes, &x->device, 1, &va) == 0;
    }
    for (int p = 0; p < world && rc == 0; ++p)
      if (p != me) rc = send_msg(x->sock, w[p].name, w[p].token, all ? MSG_BOUND : MSG_BIND_FAILED, me, -1);
    while (rc == 0 && n_bound + n_bind_failed < world - 1) rc = pump();
    int *refs = (all && n_bind_failed == 0 && rc == 0) ? new (std::nothrow) int(1) : nullptr;
    if (refs) {
      x->mc_mem = mch;
      x->mc_base = reinterpret_cast<unsigned char *>(va);
      x->mc_enabled = true;
      x->mc_refs = refs;  // this process's own mapping and handle
    } else {
      if (va) unmap_segment(d, reinterpret_cast<void *>(va), x->seg_bytes);
      if (bound) d->MulticastUnbind(mch, cudev, 0, x->seg_bytes);
      if (holding) d->MemRelease(mch);
    }
    if (rc) return rc;
  }
  x->connected = true;
  return 0;
}

// ---------------------------------------------------------------------------
// queries
// ---------------------------------------------------------------------------
extern "C" int gpu_probe_exchange_status(gpu_probe_exchange *x) {
  if (!live(x)) return GPU_PROBE_ENOTREADY;
  DeviceGuard g(x->device);
  uint64_t head[2] = {0, 0};
  cudaError_t e = cudaMemcpy(head, x->mailbox, sizeof(head), cudaMemcpyDeviceToHost);  // blocking
  if (e != cudaSuccess) return int(e);
  return head[1] ? GPU_PROBE_ETIMEDOUT : 0;
}

extern "C" int gpu_probe_exchange_has_multicast(gpu_probe_exchange *x) { return live(x) && x->mc_enabled ? 1 : 0; }

extern "C" void *gpu_probe_exchange_buffer(gpu_probe_exchange *x) {
  if (!live(x) || x->buffer_bytes == 0) return nullptr;
  return reinterpret_cast<unsigned char *>(x->mailbox) + x->payload_offset;
}

extern "C" uint64_t gpu_probe_exchange_buffer_bytes(gpu_probe_exchange *x) { return live(x) ? x->buffer_bytes : 0; }

// ---------------------------------------------------------------------------
// the fused checksum + lanes exchange entry points
// ---------------------------------------------------------------------------
static int exchange_args(const gpu_probe_exchange *x, ExchangeArgs *a) {
  if (!live(x) || !x->connected) return GPU_PROBE_ENOTREADY;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return int(e);
  if (dev != x->device) return GPU_PROBE_EINVAL;  // the kernel must run on the mailbox's GPU
  std::memset(a, 0, sizeof(*a));
  for (int p = 0; p < x->world; ++p) a->mailbox[p] = x->peer[p];
  a->rank = x->rank;
  a->world = x->world;
  a->timeout_ns = x->timeout_ns;
  return 0;
}

extern "C" int gpu_probe_checksum_u64_exchange(const uint64_t *w, uint64_t n, uint64_t first_index, uint64_t *out_lanes,
                                               void *workspace, int variant, gpu_probe_exchange *x, void *stream) {
  ExchangeArgs a;
  int rc = exchange_args(x, &a);
  if (rc) return rc;
  return checked_checksum(w, n, first_index, out_lanes, workspace, variant, stream, a);
}

extern "C" int gpu_probe_triad_checksum_exchange(double *a_out, const double *b, const double *c, double s, uint64_t n,
                                                 uint64_t first_index, uint64_t *out_lanes, void *workspace, int variant,
                                                 gpu_probe_exchange *x, void *stream) {
  ExchangeArgs a;
  int rc = exchange_args(x, &a);
  if (rc) return rc;
  return checked_triad_checksum(a_out, b, c, s, n, first_index, out_lanes, workspace, variant, stream, a);
}

extern "C" int gpu_probe_checksum_u64_exchange_group(const uint64_t *const *w, const uint64_t *n,
                                                     const uint64_t *first_index, uint64_t *const *out_lanes,
                                                     void *const *workspace, gpu_probe_exchange *const *xs, int world,
                                                     void *stream) {
  if (!w || !n || !first_index || !out_lanes || !workspace || !xs || world < 1 || world > 8) return GPU_PROBE_EINVAL;
  ChecksumGroupArgs g;
  std::memset(&g, 0, sizeof(g));
  for (int r = 0; r < world; ++r) {
    if (!live(xs[r])) return GPU_PROBE_ENOTREADY;
    if (xs[r]->world != world || xs[r]->rank != r) return GPU_PROBE_EINVAL;
    int rc = exchange_args(xs[r], &g.x[r]);  // also checks: connected, and on the current device
    if (rc) return rc;
    g.w[r] = w[r]; g.n[r] = n[r]; g.first[r] = first_index[r];
    g.out[r] = out_lanes[r]; g.ws[r] = static_cast<uint64_t *>(workspace[r]);
  }
  return checked_checksum_group(g, world, stream);
}
