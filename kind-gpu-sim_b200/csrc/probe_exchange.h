// probe_exchange.h -- the exchange object shared by probe_exchange.cu and probe_allreduce.cu.
// Internal: the C ABI only ever sees an opaque pointer (include/gpu_probe.h).
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

#include "probe_common.cuh"

using probe::kMaxRanks;

struct gpu_probe_exchange {
  uint32_t magic;
  int device, rank, world;
  bool connected;
  uint64_t *mailbox;            // this rank's segment (cudaMalloc on `device`): header + payload
  uint64_t buffer_bytes;        // payload size (0: lanes exchange only)
  uint64_t stage_bytes;         // one staging slot of the one-shot allreduce (0 when there is no payload)
  uint64_t payload_offset;      // payload = segment + payload_offset (same on every rank)
  uint64_t *peer[kMaxRanks];    // every rank's mailbox as addressed from `device`
  bool ipc_opened[kMaxRanks];   // peer[p] came from cudaIpcOpenMemHandle
  unsigned long long timeout_ns;
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
