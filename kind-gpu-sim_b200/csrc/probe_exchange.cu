// probe_exchange.cu -- the checksum's one collective, fused into the checksum kernel.
//
// each GPU checksums its shard, then the 16-byte lanes are
// summed over GPUs.  The reference-style way is a separate ncclAllReduce(count=2,
// ncclUint64) launch after K2.  Here the last CTA of K2 (or of the fused K1+K2) writes
// its lanes into a mailbox slot on every peer over NVLink (plain P2P stores + a
// release flag) and adds up the slots that land in its own mailbox: compute and
// collective are ONE kernel, nothing else is launched (device code: exchange_lanes in
// probe_common.cuh).
//
// This file is the host side: the mailbox object, and how ranks learn each other's
// mailbox addresses --
//   * ranks in ONE process (the pod's C host, kind-gpu-sim_b200/host/): peer access +
//     raw pointers, gpu_probe_exchange_connect_local;
//   * one process per GPU (bench.py under torchrun): CUDA IPC handles, carried between
//     ranks by whatever the host has (torch.distributed all_gather in bench.py),
//     gpu_probe_exchange_export / _connect_ipc.
// No reference counterpart (slot: /root/reference/pods/nvidia-gpu-test-pod.yaml:9).
#include "gpu_probe.h"

#include <cstring>
#include <cuda_runtime.h>
#include <new>

#include "probe_common.cuh"
#include "probe_exchange.h"
#include "probe_launch.h"

using namespace probe;

static_assert(GPU_PROBE_EXCHANGE_MAX_RANKS == kMaxRanks, "header and device code disagree on the rank limit");
static_assert(GPU_PROBE_EXCHANGE_HANDLE_BYTES == sizeof(cudaIpcMemHandle_t), "IPC handle size");


extern "C" int gpu_probe_exchange_create(gpu_probe_exchange **out, int rank, int world, uint64_t buffer_bytes) {
  if (!out) return GPU_PROBE_EINVAL;
  *out = nullptr;
  if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world) return GPU_PROBE_EINVAL;
  if (buffer_bytes > (1ull << 36)) return GPU_PROBE_ETOOBIG;
  buffer_bytes = (buffer_bytes + 255) & ~uint64_t(255);
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return int(e);
  gpu_probe_exchange *x = new (std::nothrow) gpu_probe_exchange();
  if (!x) return int(cudaErrorMemoryAllocation);
  x->device = dev; x->rank = rank; x->world = world;
  x->timeout_ns = 10ull * 1000 * 1000 * 1000;  // a peer that is 10 s late is not coming
  x->buffer_bytes = buffer_bytes;
  x->stage_bytes = buffer_bytes < kMaxStageBytes ? buffer_bytes : kMaxStageBytes;
  x->payload_offset = kHeaderBytes + 2 * uint64_t(world) * x->stage_bytes;
  // header (mailbox, barrier epochs and flags), staging and payload are ONE allocation: one IPC handle maps all
  e = cudaMalloc(reinterpret_cast<void **>(&x->mailbox), x->payload_offset + buffer_bytes);
  if (e == cudaSuccess) e = cudaMemset(x->mailbox, 0, kHeaderBytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();  // zeroed before any peer can learn the address
  if (e != cudaSuccess) {
    if (x->mailbox) cudaFree(x->mailbox);
    delete x;
    return int(e);
  }
  x->peer[rank] = x->mailbox;
  x->connected = (world == 1);
  x->magic = kXMagic;
  *out = x;
  return 0;
}

extern "C" int gpu_probe_exchange_set_timeout_ms(gpu_probe_exchange *x, uint64_t ms) {
  if (!live(x)) return GPU_PROBE_ENOTREADY;
  if (ms == 0 || ms > 600000) return GPU_PROBE_EINVAL;
  x->timeout_ns = ms * 1000000ull;
  return 0;
}

extern "C" int gpu_probe_exchange_export(gpu_probe_exchange *x, unsigned char handle[GPU_PROBE_EXCHANGE_HANDLE_BYTES]) {
  if (!live(x)) return GPU_PROBE_ENOTREADY;
  if (!handle) return GPU_PROBE_EINVAL;
  DeviceGuard g(x->device);
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, x->mailbox);
  if (e != cudaSuccess) return int(e);
  std::memcpy(handle, &h, sizeof(h));
  return 0;
}

extern "C" int gpu_probe_exchange_connect_ipc(gpu_probe_exchange *x, const unsigned char *handles) {
  if (!live(x)) return GPU_PROBE_ENOTREADY;
  if (!handles) return GPU_PROBE_EINVAL;
  if (x->connected) return 0;
  DeviceGuard g(x->device);
  for (int p = 0; p < x->world; ++p) {
    if (p == x->rank) continue;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handles + size_t(p) * sizeof(h), sizeof(h));
    void *ptr = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return int(e);
    x->peer[p] = static_cast<uint64_t *>(ptr);
    x->ipc_opened[p] = true;
  }
  x->connected = true;
  return 0;
}

extern "C" int gpu_probe_exchange_connect_local(gpu_probe_exchange *const *xs, int world) {
  if (!xs || world < 1 || world > kMaxRanks) return GPU_PROBE_EINVAL;
  for (int r = 0; r < world; ++r) {
    if (!live(xs[r])) return GPU_PROBE_ENOTREADY;
    if (xs[r]->world != world || xs[r]->rank != r) return GPU_PROBE_EINVAL;
  }
  for (int r = 0; r < world; ++r) {
    DeviceGuard g(xs[r]->device);
    for (int p = 0; p < world; ++p) {
      if (p == r) continue;
      if (xs[p]->device != xs[r]->device) {
        int can = 0;
        cudaError_t e = cudaDeviceCanAccessPeer(&can, xs[r]->device, xs[p]->device);
        if (e != cudaSuccess) return int(e);
        if (!can) return int(cudaErrorPeerAccessUnsupported);
        e = cudaDeviceEnablePeerAccess(xs[p]->device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
        if (e != cudaSuccess) return int(e);
      }
      xs[r]->peer[p] = xs[p]->mailbox;  // unified addressing: same pointer, reached over NVLink
    }
    xs[r]->connected = true;
  }
  return 0;
}

extern "C" int gpu_probe_exchange_status(gpu_probe_exchange *x) {
  if (!live(x)) return GPU_PROBE_ENOTREADY;
  DeviceGuard g(x->device);
  uint64_t head[2] = {0, 0};
  cudaError_t e = cudaMemcpy(head, x->mailbox, sizeof(head), cudaMemcpyDeviceToHost);  // blocking
  if (e != cudaSuccess) return int(e);
  return head[1] ? GPU_PROBE_ETIMEDOUT : 0;
}

extern "C" int gpu_probe_exchange_destroy(gpu_probe_exchange *x) {
  if (!live(x)) return GPU_PROBE_ENOTREADY;
  DeviceGuard g(x->device);
  cudaDeviceSynchronize();
  for (int p = 0; p < x->world; ++p)
    if (x->ipc_opened[p] && x->peer[p]) cudaIpcCloseMemHandle(x->peer[p]);
  cudaFree(x->mailbox);
  x->magic = 0;
  delete x;
  return 0;
}

extern "C" void *gpu_probe_exchange_buffer(gpu_probe_exchange *x) {
  if (!live(x) || x->buffer_bytes == 0) return nullptr;
  return reinterpret_cast<unsigned char *>(x->mailbox) + x->payload_offset;
}

extern "C" uint64_t gpu_probe_exchange_buffer_bytes(gpu_probe_exchange *x) { return live(x) ? x->buffer_bytes : 0; }

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
