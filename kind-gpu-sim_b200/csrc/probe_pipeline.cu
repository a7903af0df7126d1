// probe_pipeline.cu -- host-buffer end-to-end path (include/gpu_probe.h, "pipeline").
//
// A host that already holds b[] and c[] (pinned) gets a[] and its lanes back.
// PCIe is the bound here (two 8-byte reads in, optionally one out, per element),
// so the job is overlap: chunk k+1 crosses the bus on the copy stream while the
// fused triad+checksum kernel runs chunk k and chunk k-1's results go home on a
// third stream.  Every chunk's lanes are computed with its GLOBAL first index, so
// the host just adds them (wrapping u64).
//
// All device memory is allocated in create() and reused; run_host() allocates
// nothing.  No reference counterpart (slot: /root/reference/pods/nvidia-gpu-test-pod.yaml:9).
#include "gpu_probe.h"

#include <cuda_runtime.h>
#include <new>

struct gpu_probe_pipeline {
  uint32_t magic;
  int device;
  int depth;
  uint64_t chunk;
  cudaStream_t s_in, s_run, s_out;
  double *d_b, *d_c, *d_a;  // depth * chunk each
  uint64_t *d_lanes;        // depth * 2
  unsigned char *d_ws;      // depth * workspace
  uint64_t *h_lanes;        // pinned, depth * 2
  cudaEvent_t *ev_in, *ev_run, *ev_out;
  bool *busy;
};

static constexpr uint32_t kMagic = 0x50524F42u;  // "PROB"

#define PIPE_TRY(x)                      \
  do {                                   \
    cudaError_t e__ = (x);               \
    if (e__ != cudaSuccess) { rc = int(e__); goto fail; } \
  } while (0)

extern "C" int gpu_probe_pipeline_destroy(gpu_probe_pipeline *p) {
  if (!p || p->magic != kMagic) return GPU_PROBE_ENOTREADY;
  int prev = -1;
  cudaGetDevice(&prev);
  cudaSetDevice(p->device);
  if (p->s_in) cudaStreamSynchronize(p->s_in);
  if (p->s_run) cudaStreamSynchronize(p->s_run);
  if (p->s_out) cudaStreamSynchronize(p->s_out);
  for (int i = 0; i < p->depth; ++i) {
    if (p->ev_in && p->ev_in[i]) cudaEventDestroy(p->ev_in[i]);
    if (p->ev_run && p->ev_run[i]) cudaEventDestroy(p->ev_run[i]);
    if (p->ev_out && p->ev_out[i]) cudaEventDestroy(p->ev_out[i]);
  }
  delete[] p->ev_in; delete[] p->ev_run; delete[] p->ev_out; delete[] p->busy;
  cudaFree(p->d_b); cudaFree(p->d_c); cudaFree(p->d_a); cudaFree(p->d_lanes); cudaFree(p->d_ws);
  if (p->h_lanes) cudaFreeHost(p->h_lanes);
  if (p->s_in) cudaStreamDestroy(p->s_in);
  if (p->s_run) cudaStreamDestroy(p->s_run);
  if (p->s_out) cudaStreamDestroy(p->s_out);
  p->magic = 0;
  delete p;
  if (prev >= 0) cudaSetDevice(prev);
  return 0;
}

extern "C" int gpu_probe_pipeline_create(gpu_probe_pipeline **out, uint64_t chunk_elems, int depth) {
  if (!out) return GPU_PROBE_EINVAL;
  *out = nullptr;
  if (chunk_elems == 0) chunk_elems = 1ull << 23;  // 64 MiB per array per chunk
  if (depth == 0) depth = 4;
  if (depth < 2 || depth > 16 || (chunk_elems & 1) || chunk_elems > (1ull << 30)) return GPU_PROBE_EINVAL;
  gpu_probe_pipeline *p = new (std::nothrow) gpu_probe_pipeline();
  if (!p) return int(cudaErrorMemoryAllocation);
  int rc = 0;
  const size_t ws = gpu_probe_checksum_workspace_bytes();
  p->magic = kMagic;
  p->depth = depth;
  p->chunk = chunk_elems;
  p->ev_in = new (std::nothrow) cudaEvent_t[depth]();
  p->ev_run = new (std::nothrow) cudaEvent_t[depth]();
  p->ev_out = new (std::nothrow) cudaEvent_t[depth]();
  p->busy = new (std::nothrow) bool[depth]();
  if (!p->ev_in || !p->ev_run || !p->ev_out || !p->busy) { rc = int(cudaErrorMemoryAllocation); goto fail; }
  PIPE_TRY(cudaGetDevice(&p->device));
  PIPE_TRY(cudaStreamCreateWithFlags(&p->s_in, cudaStreamNonBlocking));
  PIPE_TRY(cudaStreamCreateWithFlags(&p->s_run, cudaStreamNonBlocking));
  PIPE_TRY(cudaStreamCreateWithFlags(&p->s_out, cudaStreamNonBlocking));
  PIPE_TRY(cudaMalloc(&p->d_b, sizeof(double) * chunk_elems * depth));
  PIPE_TRY(cudaMalloc(&p->d_c, sizeof(double) * chunk_elems * depth));
  PIPE_TRY(cudaMalloc(&p->d_a, sizeof(double) * chunk_elems * depth));
  PIPE_TRY(cudaMalloc(&p->d_lanes, sizeof(uint64_t) * 2 * depth));
  PIPE_TRY(cudaMalloc(&p->d_ws, ws * depth));
  PIPE_TRY(cudaMemset(p->d_ws, 0, ws * depth));
  PIPE_TRY(cudaHostAlloc(&p->h_lanes, sizeof(uint64_t) * 2 * depth, cudaHostAllocDefault));
  for (int i = 0; i < depth; ++i) {
    PIPE_TRY(cudaEventCreateWithFlags(&p->ev_in[i], cudaEventDisableTiming));
    PIPE_TRY(cudaEventCreateWithFlags(&p->ev_run[i], cudaEventDisableTiming));
    PIPE_TRY(cudaEventCreateWithFlags(&p->ev_out[i], cudaEventDisableTiming));
  }
  PIPE_TRY(cudaDeviceSynchronize());
  *out = p;
  return 0;
fail:
  gpu_probe_pipeline_destroy(p);
  return rc;
}

extern "C" int gpu_probe_pipeline_run_host(gpu_probe_pipeline *p, const double *b_host, const double *c_host,
                                           double *a_host, double s, uint64_t n, uint64_t first_index,
                                           uint64_t lanes_out[2]) {
  if (!p || p->magic != kMagic) return GPU_PROBE_ENOTREADY;
  if (!lanes_out || (n && (!b_host || !c_host))) return GPU_PROBE_EINVAL;
  int rc = 0;
  int prev = -1;
  uint64_t l0 = 0, l1 = 0;
  const size_t ws = gpu_probe_checksum_workspace_bytes();
  cudaGetDevice(&prev);
  PIPE_TRY(cudaSetDevice(p->device));
  for (int i = 0; i < p->depth; ++i) p->busy[i] = false;

  for (uint64_t off = 0, k = 0; off < n; off += p->chunk, ++k) {
    const int slot = int(k % uint64_t(p->depth));
    const uint64_t len = (n - off < p->chunk) ? n - off : p->chunk;
    if (p->busy[slot]) {  // slot's previous chunk must be home before its buffers are reused
      PIPE_TRY(cudaEventSynchronize(p->ev_out[slot]));
      l0 += p->h_lanes[2 * slot];
      l1 += p->h_lanes[2 * slot + 1];
      p->busy[slot] = false;
    }
    double *db = p->d_b + size_t(slot) * p->chunk, *dc = p->d_c + size_t(slot) * p->chunk;
    double *da = p->d_a + size_t(slot) * p->chunk;
    uint64_t *dl = p->d_lanes + 2 * slot;
    PIPE_TRY(cudaMemcpyAsync(db, b_host + off, sizeof(double) * len, cudaMemcpyHostToDevice, p->s_in));
    PIPE_TRY(cudaMemcpyAsync(dc, c_host + off, sizeof(double) * len, cudaMemcpyHostToDevice, p->s_in));
    PIPE_TRY(cudaEventRecord(p->ev_in[slot], p->s_in));
    PIPE_TRY(cudaStreamWaitEvent(p->s_run, p->ev_in[slot], 0));
    rc = gpu_probe_triad_checksum(da, db, dc, s, len, first_index + off, dl, p->d_ws + size_t(slot) * ws,
                                  GPU_PROBE_VARIANT_AUTO, p->s_run);
    if (rc) goto fail;
    PIPE_TRY(cudaEventRecord(p->ev_run[slot], p->s_run));
    PIPE_TRY(cudaStreamWaitEvent(p->s_out, p->ev_run[slot], 0));
    PIPE_TRY(cudaMemcpyAsync(p->h_lanes + 2 * slot, dl, sizeof(uint64_t) * 2, cudaMemcpyDeviceToHost, p->s_out));
    if (a_host)
      PIPE_TRY(cudaMemcpyAsync(a_host + off, da, sizeof(double) * len, cudaMemcpyDeviceToHost, p->s_out));
    PIPE_TRY(cudaEventRecord(p->ev_out[slot], p->s_out));
    p->busy[slot] = true;
  }
  for (int slot = 0; slot < p->depth; ++slot) {
    if (!p->busy[slot]) continue;
    PIPE_TRY(cudaEventSynchronize(p->ev_out[slot]));
    l0 += p->h_lanes[2 * slot];
    l1 += p->h_lanes[2 * slot + 1];
    p->busy[slot] = false;
  }
  lanes_out[0] = l0;
  lanes_out[1] = l1;
  if (prev >= 0) cudaSetDevice(prev);
  return 0;
fail:
  cudaStreamSynchronize(p->s_in);
  cudaStreamSynchronize(p->s_run);
  cudaStreamSynchronize(p->s_out);
  if (prev >= 0) cudaSetDevice(prev);
  return rc;
}
