// probe_exchange.h -- the exchange object shared by probe_exchange.cu and probe_allreduce.cu.
// Internal: the C ABI only ever sees an opaque pointer (include/gpu_probe.h).
#pragma once

#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>

#include "probe_common.cuh"

using probe::kMaxRanks;

struct gpu_probe_exchange {
  uint32_t magic;
  int device, rank, world;
  bool connected;
  // This rank's symmetric segment: header + staging + payload, ONE physical allocation made
  // with the driver's virtual-memory API (cuMemCreate) so that it can be (a) shared with other
  // processes as a file descriptor and (b) bound into an NVSwitch multicast object.
  CUmemGenericAllocationHandle mem;
  size_t seg_bytes;             // physical size (multiple of the allocation and multicast granularity)
  uint64_t *mailbox;            // = the segment's base address on `device`
  uint64_t buffer_bytes;        // payload size (0: lanes exchange only)
  uint64_t stage_bytes;         // one staging slot of the one-shot allreduce (0 when there is no payload)
  uint64_t payload_offset;      // payload = segment + payload_offset (same on every rank)
  uint64_t *peer[kMaxRanks];    // every rank's segment as addressed from `device`
  CUmemGenericAllocationHandle peer_mem[kMaxRanks];  // imported handles (other processes' segments)
  bool peer_imported[kMaxRanks];
  // NVSwitch multicast (NVLS): one address that reads/writes ALL ranks' segments at once.
  bool mc_supported;            // this device can join a multicast object
  bool mc_enabled;              // multicast address is mapped and every rank has bound its segment
  int *mc_refs;                 // objects of this process that share mc_mem / mc_base; the last one out frees them
  CUmemGenericAllocationHandle mc_mem;
  unsigned char *mc_base;       // multicast alias of the segments (same layout)
  // one process per GPU: handles travel as file descriptors over a unix datagram socket
  int sock;
  char sock_name[32];
  uint64_t token;               // this rank's secret: travels inside the exported handle, demanded of every message
  unsigned long long timeout_ns;
  uint64_t push_calls;          // one-shot allreduce calls issued on this exchange (staging parity, flag epoch)
};

static constexpr uint32_t kXMagic = 0x58434847u;  // "XCHG"
static inline bool live(const gpu_probe_exchange *x) { return x && x->magic == kXMagic; }

namespace {
struct DeviceGuard {  // run a few calls with another device current, then put the old one back
  int prev = -1;
  explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
}  // namespace
