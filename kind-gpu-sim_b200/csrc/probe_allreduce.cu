// probe_allreduce.cu -- K4: fp32 sum allreduce over NVLink peer memory, one kernel.
//
// The NVLink half of the headline metric is an allreduce sweep (SURVEY.md section 8d,
// "Allreduce sweep").  NCCL is the baseline the host binary and bench.py time; this is
// the probe's own collective, written against the same symmetric segments the lanes
// exchange uses (probe_exchange.cu), so it needs no NCCL and no extra launches:
//
//   rank r owns slice r of the buffer.  For every 16-byte vector of its slice it PULLS the
//   vector from all G buffers (its own + G-1 peers over NVLink), adds them in rank order,
//   and PUSHES the sum back into all G buffers at the same offset.  The thread that
//   overwrites an element is the only one that ever reads it, so the reduce-scatter and
//   the all-gather need no barrier between them: one barrier before (every input is
//   ready) and one after (every push has landed) is all the synchronisation there is.
//
// Per direction each GPU moves 2(G-1)/G x bytes, so bus bandwidth = algbw x 2(G-1)/G is
// bounded by the NVLink rate (900 GB/s per direction nominal).  Every element is summed by
// exactly one rank in a fixed order, so all ranks end with bit-identical results.
// Barriers are per CTA: CTA b of every rank works on the same vectors of every slice, so it
// only has to meet CTA b of the peers (flags[b][rank] in the segment header) -- no
// grid-wide synchronisation, no co-residency requirement beyond "peer CTA b gets to run".
// No reference counterpart (slot: /root/reference/pods/nvidia-gpu-test-pod.yaml:9).
#include "gpu_probe.h"

#include <atomic>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

#include "probe_common.cuh"
#include "probe_exchange.h"

namespace probe {

constexpr int kArThreads = 512;

struct AllreduceArgs {
  unsigned char *seg[kMaxRanks];  // segment base of rank p as THIS GPU addresses it
  int rank, world;
  unsigned long long timeout_ns;
  uint64_t operand_off;   // operand = segment + operand_off (payload offset + caller's offset)
  uint64_t count;         // fp32 elements
  uint64_t stage_bytes;   // one staging slot (one-shot algorithm); staging = segment + kHeaderBytes
  unsigned char *mc;      // NVSwitch multicast alias of ALL ranks' segments (same layout), or null
  uint64_t push_call;     // one-shot algorithm: number of this call on the exchange, 1, 2, ... (host-side counter)
};

__device__ __forceinline__ float4 ld_peer(const float4 *p) {  // system-coherent, never from a stale L1 line
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer(float4 *p, const float4 &v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// NVLS: one load returns the sum over every rank's copy (added inside the switch), one store
// writes every rank's copy.  SASS: LDGMC.E.ADD.F32x4 / STG on a multicast address.
__device__ __forceinline__ float4 ld_reduce_mc(const float4 *mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void st_mc(float4 *mc, const float4 &v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float ld_peer(const float *p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer(float *p, float v) {
  asm volatile("st.relaxed.sys.global.f32 [%0], %1;" :: "l"(p), "f"(v) : "memory");
}

// CTA b of this rank meets CTA b of every peer.  Returns false on timeout.
// Release/acquire at system scope on the flag; the __syncthreads on either side extend
// that ordering to every thread of the CTA (cumulativity).
template <size_t FLAGS_OFFSET_WORDS = kFlagsOffsetWords>
__device__ __forceinline__ bool cta_barrier(const AllreduceArgs &a, uint64_t epoch) {
  __shared__ int timed_out;
  if (threadIdx.x == 0) timed_out = 0;
  __syncthreads();
  if (threadIdx.x < a.world) {
    const int p = threadIdx.x;
    uint64_t *theirs = reinterpret_cast<uint64_t *>(a.seg[p]) + FLAGS_OFFSET_WORDS + size_t(blockIdx.x) * kMaxRanks + a.rank;
    // st.release.sys alone orders everything the CTA wrote before the __syncthreads above; an
    // explicit __threadfence_system() in front of it cost 1-4 us per barrier (7.7 -> 6.5 us at
    // 1 KiB, 15.3 -> 11.5 us at 1 MiB, G = 2) and bought nothing.  Polling with ld.acquire beats
    // ld.relaxed + one fence after the loop by another 1-5 us.
    st_release_sys(theirs, epoch);
    const uint64_t *mine = reinterpret_cast<const uint64_t *>(a.seg[a.rank]) + FLAGS_OFFSET_WORDS +
                           size_t(blockIdx.x) * kMaxRanks + p;
    const unsigned long long t0 = global_ns();
    while (ld_acquire_sys(mine) < epoch) {
      if (global_ns() - t0 > a.timeout_ns) { timed_out = 1; break; }
    }
  }
  __syncthreads();
  return timed_out == 0;
}

template <int G, int U>
__device__ __forceinline__ void reduce_broadcast(const AllreduceArgs &a, uint64_t v_lo, uint64_t v_hi) {
  // vectors [v_lo, v_hi) of the operand belong to this rank; CTAs stride over them
  float4 *buf[G];
#pragma unroll
  for (int p = 0; p < G; ++p) buf[p] = reinterpret_cast<float4 *>(a.seg[p] + a.operand_off);
  const uint64_t tile = uint64_t(kArThreads) * U;
  for (uint64_t base = v_lo + blockIdx.x * tile; base < v_hi; base += uint64_t(gridDim.x) * tile) {
    float4 v[U][G];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t i = base + uint64_t(u) * kArThreads + threadIdx.x;
      if (i < v_hi) {
#pragma unroll
        for (int p = 0; p < G; ++p) v[u][p] = ld_peer(buf[p] + i);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t i = base + uint64_t(u) * kArThreads + threadIdx.x;
      if (i < v_hi) {
        float4 s = v[u][0];
#pragma unroll
        for (int p = 1; p < G; ++p) {  // fixed order 0,1,..,G-1: the defined result
          s.x += v[u][p].x; s.y += v[u][p].y; s.z += v[u][p].z; s.w += v[u][p].w;
        }
#pragma unroll
        for (int p = 0; p < G; ++p) st_peer(buf[p] + i, s);
      }
    }
  }
}

template <int G, int U>
__device__ __forceinline__ void pull_push_body(const AllreduceArgs &a) {
  uint64_t *hdr = reinterpret_cast<uint64_t *>(a.seg[a.rank]);
  __shared__ uint64_t epoch0;
  if (threadIdx.x == 0) epoch0 = hdr[kCtaEpochOffsetWords + blockIdx.x];
  __syncthreads();
  const uint64_t e = epoch0;
  bool ok = cta_barrier(a, e + 1);  // every rank's operand is ready to be read
  if (ok) {
    const uint64_t nvec = a.count / 4;
    const uint64_t v_lo = nvec * uint64_t(a.rank) / G, v_hi = nvec * uint64_t(a.rank + 1) / G;
    reduce_broadcast<G, U>(a, v_lo, v_hi);
    // fewer than 4 trailing elements: the last rank owns them
    if (a.rank == G - 1 && blockIdx.x == 0 && threadIdx.x < (a.count & 3)) {
      const uint64_t i = nvec * 4 + threadIdx.x;
      float s = 0.f;
#pragma unroll
      for (int p = 0; p < G; ++p) {
        const float x = ld_peer(reinterpret_cast<const float *>(a.seg[p] + a.operand_off) + i);
        s = p == 0 ? x : s + x;
      }
#pragma unroll
      for (int p = 0; p < G; ++p) st_peer(reinterpret_cast<float *>(a.seg[p] + a.operand_off) + i, s);
    }
    ok = cta_barrier(a, e + 2);  // every push has landed everywhere before any rank's kernel ends
  }
  if (threadIdx.x == 0) {
    hdr[kCtaEpochOffsetWords + blockIdx.x] = e + 2;
    if (!ok) hdr[1] = 1;  // sticky error flag, read by gpu_probe_exchange_status
  }
}

template <int G, int U>
__global__ void __launch_bounds__(kArThreads)
allreduce_f32_kernel(const AllreduceArgs a) { pull_push_body<G, U>(a); }

// ---- NVLS variant ----------------------------------------------------------------------
// Same ownership as above (rank r reduces and re-publishes slice r), but the reduction and
// the broadcast happen INSIDE the NVSwitch: multimem.ld_reduce pulls one vector and receives
// the sum of all G copies; multimem.st pushes one vector and the switch writes all G copies.
// Per direction a GPU now moves (1 + 1/G) x bytes instead of 2(G-1)/G x -- 1.56x less at
// G = 8 -- so bus bandwidth can exceed the link rate.  The switch's summation order is not
// ours to choose: ranks still agree bit for bit (one reducer per element), but the value may
// differ from the rank-order sum in the last place.
template <int U>
__global__ void __launch_bounds__(kArThreads)
allreduce_f32_multimem_kernel(const AllreduceArgs a) {
  uint64_t *hdr = reinterpret_cast<uint64_t *>(a.seg[a.rank]);
  __shared__ uint64_t epoch0;
  if (threadIdx.x == 0) epoch0 = hdr[kCtaEpochOffsetWords + blockIdx.x];
  __syncthreads();
  const uint64_t e = epoch0;
  bool ok = cta_barrier(a, e + 1);
  if (ok) {
    const int G = a.world;
    float4 *mc = reinterpret_cast<float4 *>(a.mc + a.operand_off);
    const uint64_t nvec = a.count / 4;
    const uint64_t v_lo = nvec * uint64_t(a.rank) / G, v_hi = nvec * uint64_t(a.rank + 1) / G;
    const uint64_t tile = uint64_t(kArThreads) * U;
    for (uint64_t base = v_lo + blockIdx.x * tile; base < v_hi; base += uint64_t(gridDim.x) * tile) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint64_t i = base + uint64_t(u) * kArThreads + threadIdx.x;
        if (i < v_hi) v[u] = ld_reduce_mc(mc + i);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint64_t i = base + uint64_t(u) * kArThreads + threadIdx.x;
        if (i < v_hi) st_mc(mc + i, v[u]);
      }
    }
    if (a.rank == G - 1 && blockIdx.x == 0 && threadIdx.x < (a.count & 3)) {  // tail: plain peer loads and stores
      const uint64_t i = nvec * 4 + threadIdx.x;
      float s = 0.f;
      for (int p = 0; p < G; ++p) {
        const float x = ld_peer(reinterpret_cast<const float *>(a.seg[p] + a.operand_off) + i);
        s = p == 0 ? x : s + x;
      }
      for (int p = 0; p < G; ++p) st_peer(reinterpret_cast<float *>(a.seg[p] + a.operand_off) + i, s);
    }
    ok = cta_barrier(a, e + 2);
  }
  if (threadIdx.x == 0) {
    hdr[kCtaEpochOffsetWords + blockIdx.x] = e + 2;
    if (!ok) hdr[1] = 1;
  }
}

// ---- one-shot variant for small operands ---------------------------------------------
// Latency, not bandwidth, decides below a few MiB, and the pull/push kernel above costs
// four NVLink traversals (barrier, pull round trip, push, barrier).  Here every rank PUSHES
// its whole operand into slot `rank` of a staging area on every peer, raises one flag per
// CTA, and then adds up the G slots that landed in its own staging area -- one traversal
// plus the fence.  Every rank computes every element itself, in rank order, so the ranks
// still agree bit for bit.  Staging is double-buffered by call parity (same argument as the
// lanes mailbox), so no trailing barrier is needed; peers never read the operand itself.
// The parity comes from ONE counter per exchange, kept by the host and identical in every CTA of
// every rank (all ranks make the same sequence of calls).  A per-CTA count would let CTAs of
// consecutive calls with different grids disagree about which staging buffer a vector lives in.
template <int G, int U>
__device__ __forceinline__ void push_body(const AllreduceArgs &a) {
  uint64_t *hdr = reinterpret_cast<uint64_t *>(a.seg[a.rank]);
  const uint64_t e = a.push_call;  // flags only ever grow: a CTA that sat out smaller calls simply jumps ahead
  const size_t slot_set = kHeaderBytes + size_t(e & 1) * G * a.stage_bytes;  // staging[e & 1]
  float4 *mine_op = reinterpret_cast<float4 *>(a.seg[a.rank] + a.operand_off);
  const uint64_t nvec = (a.count + 3) / 4;  // the operand is padded to 16 bytes inside the buffer (see entry point)
  float4 own[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const uint64_t i = (uint64_t(u) * gridDim.x + blockIdx.x) * kArThreads + threadIdx.x;
    if (i < nvec) {
      own[u] = mine_op[i];
      if (a.mc != nullptr && G > 2) {  // one store, the switch writes slot `rank` on every GPU (mine too: harmless)
        st_mc(reinterpret_cast<float4 *>(a.mc + slot_set + size_t(a.rank) * a.stage_bytes) + i, own[u]);
      } else {
#pragma unroll
        for (int q = 1; q < G; ++q) {  // start with the next rank: spread the first stores over the links
          const int p = (a.rank + q) % G;
          st_peer(reinterpret_cast<float4 *>(a.seg[p] + slot_set + size_t(a.rank) * a.stage_bytes) + i, own[u]);
        }
      }
    }
  }
  const bool ok = cta_barrier<kPushFlagsOffsetWords>(a, e);  // my stores have landed; the peers' too
  if (ok) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t i = (uint64_t(u) * gridDim.x + blockIdx.x) * kArThreads + threadIdx.x;
      if (i < nvec) {
        float4 s;
#pragma unroll
        for (int p = 0; p < G; ++p) {
          const float4 x = p == a.rank ? own[u]
                                       : ld_peer(reinterpret_cast<const float4 *>(a.seg[a.rank] + slot_set + size_t(p) * a.stage_bytes) + i);
          if (p == 0) s = x;
          else { s.x += x.x; s.y += x.y; s.z += x.z; s.w += x.w; }
        }
        mine_op[i] = s;
      }
    }
  }
  if (threadIdx.x == 0 && !ok) hdr[1] = 1;
}

template <int G, int U>
__global__ void __launch_bounds__(kArThreads)
allreduce_f32_push_kernel(const AllreduceArgs a) { push_body<G, U>(a); }

// ---- single-GPU group self-test ---------------------------------------------------------
// Ranks that wait for each other must be resident at the same time, and nothing guarantees
// that for SEPARATE launches on one GPU.  To validate the protocols on a box with fewer GPUs
// than ranks, all G ranks run as ONE cooperative launch: grid = (ctas, G), row y is rank y and
// executes exactly the bodies above (blockIdx.x / gridDim.x mean the same as in a per-rank
// launch).  Never faster than anything; a diagnostic.
struct GroupArgs { AllreduceArgs r[8]; };

template <int G>
__global__ void __launch_bounds__(kArThreads)
allreduce_f32_group_kernel(const GroupArgs ga, int algo) {
  const AllreduceArgs &a = ga.r[blockIdx.y];
  if (algo == GPU_PROBE_ALLREDUCE_PUSH) push_body<G, 2>(a);
  else pull_push_body<G, 2>(a);
}

constexpr int kPushU = 2;  // vectors per thread in the one-shot kernel: grid * 512 * kPushU * 16 B covers the operand

template <int G>
static int launch_allreduce(const AllreduceArgs &a, int algo, unsigned grid, cudaStream_t st) {
  if (algo == GPU_PROBE_ALLREDUCE_PUSH) {
    allreduce_f32_push_kernel<G, kPushU><<<grid, kArThreads, 0, st>>>(a);
  } else if (algo == GPU_PROBE_ALLREDUCE_MULTIMEM) {
    static const int u = [] { const char *e = getenv("GPU_PROBE_TUNE_MM_U"); return e ? atoi(e) : 0; }();  // dev knob
    if (u == 2) allreduce_f32_multimem_kernel<2><<<grid, kArThreads, 0, st>>>(a);
    else if (u == 8) allreduce_f32_multimem_kernel<8><<<grid, kArThreads, 0, st>>>(a);
    else if (u == 1) allreduce_f32_multimem_kernel<1><<<grid, kArThreads, 0, st>>>(a);
    else allreduce_f32_multimem_kernel<4><<<grid, kArThreads, 0, st>>>(a);
  } else {
    constexpr int U = G <= 2 ? 8 : (G <= 4 ? 4 : 2);  // G*U 16-byte loads in flight per thread
    allreduce_f32_kernel<G, U><<<grid, kArThreads, 0, st>>>(a);
  }
  return int(cudaGetLastError());
}

}  // namespace probe

using namespace probe;

static int device_sms(int dev, int *sms) {
  static std::atomic<int> cache[64];
  if (dev >= 0 && dev < 64) {
    const int known = cache[dev].load(std::memory_order_relaxed);
    if (known > 0) { *sms = known; return 0; }
  }
  int v = 0;
  cudaError_t e = cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) return int(e);
  if (dev >= 0 && dev < 64) cache[dev].store(v, std::memory_order_relaxed);
  *sms = v;
  return 0;
}

// Operand size (bytes) up to which AUTO takes the one-shot kernel.  It moves (G-1) x the
// operand per GPU instead of 2(G-1)/G x, so the crossover falls as G grows
// (profiles/allreduce_sweep_r1.md).
static uint64_t push_threshold_bytes(int world) {
  return world <= 4 ? (2u << 20) : (512u << 10);
}

// Argument checks, algorithm choice and grid size for one rank's call.  *count == 0 on return
// means "nothing to do".  The grid is a function of (count, world, algo, ctas, SM count) only,
// so every rank picks the same one -- CTA b must exist on every rank.
static int plan_allreduce(gpu_probe_exchange *x, uint64_t offset_bytes, uint64_t count, int *algo_io, int ctas,
                          AllreduceArgs *a_out, unsigned *grid_out) {
  int algo = *algo_io;
  if (!live(x) || !x->connected) return GPU_PROBE_ENOTREADY;
  if (x->buffer_bytes == 0 || (offset_bytes & 15)) return GPU_PROBE_EINVAL;
  if (offset_bytes > x->buffer_bytes || count > (x->buffer_bytes - offset_bytes) / 4) return GPU_PROBE_ETOOBIG;
  if (ctas < 0 || ctas > kMaxCommCtas) return GPU_PROBE_EINVAL;
  if (algo < GPU_PROBE_ALLREDUCE_AUTO || algo > GPU_PROBE_ALLREDUCE_MULTIMEM) return GPU_PROBE_EINVAL;
  if (algo == GPU_PROBE_ALLREDUCE_MULTIMEM && !x->mc_enabled) return GPU_PROBE_ENOTREADY;
  if (x->world > 8) return GPU_PROBE_EINVAL;  // more than 8 ranks: not one NVSwitch domain of this probe
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return int(e);
  if (dev != x->device) return GPU_PROBE_EINVAL;
  int sms = 0;
  int rc = device_sms(dev, &sms);
  if (rc) return rc;
  AllreduceArgs a;
  std::memset(&a, 0, sizeof(a));
  for (int p = 0; p < x->world; ++p) a.seg[p] = reinterpret_cast<unsigned char *>(x->peer[p]);
  a.rank = x->rank; a.world = x->world; a.timeout_ns = x->timeout_ns;
  a.operand_off = x->payload_offset + offset_bytes; a.count = count; a.stage_bytes = x->stage_bytes;
  a.mc = x->mc_enabled ? x->mc_base : nullptr;
  const uint64_t bytes = 4 * count;
  // The one-shot kernel works on whole 16-byte vectors; the buffer is padded to 256 bytes, so
  // the up-to-3 floats past `count` exist -- but they belong to the caller, so only take the
  // one-shot path when the operand is a whole number of vectors.
  const bool push_ok = (count % 4 == 0) && bytes <= x->stage_bytes;
  if (algo == GPU_PROBE_ALLREDUCE_PUSH && !push_ok) return GPU_PROBE_ETOOBIG;
  if (algo == GPU_PROBE_ALLREDUCE_AUTO)
    algo = (push_ok && bytes <= push_threshold_bytes(x->world)) ? GPU_PROBE_ALLREDUCE_PUSH
           : (x->mc_enabled && x->world >= 3)                          ? GPU_PROBE_ALLREDUCE_MULTIMEM  // (1+1/G) < 2(G-1)/G from G = 3
                                                                       : GPU_PROBE_ALLREDUCE_PULL_PUSH;
  unsigned grid;
  if (algo == GPU_PROBE_ALLREDUCE_PUSH) {
    const uint64_t nvec = count / 4;
    const uint64_t need = (nvec + uint64_t(kArThreads) * kPushU - 1) / (uint64_t(kArThreads) * kPushU);  // must cover the operand
    if (need > uint64_t(kMaxCommCtas)) return GPU_PROBE_ETOOBIG;
    uint64_t g = (nvec + kArThreads - 1) / kArThreads;   // one vector per thread while the SMs last
    if (g > uint64_t(sms)) g = uint64_t(sms);
    if (ctas > 0) g = uint64_t(ctas);
    if (g < need) g = need;
    if (g < 1) g = 1;
    grid = unsigned(g);
  } else {
    const uint64_t slice_vecs = (count / 4 + x->world - 1) / x->world;
    uint64_t want = (slice_vecs + kArThreads - 1) / kArThreads;  // at least one vector per thread
    if (want < 1) want = 1;
    // NVLS is bounded by the switch, not by SM issue rate: 16-48 CTAs already reach its ceiling
    // (835 GB/s bus at 8 GPUs) and more CTAs only add barrier traffic (679 -> 586 GB/s at 64 MiB
    // going from 16 to 148, profiles/allreduce_tuning_r1.md).  The P2P kernel wants every SM.
    const uint64_t cap = algo == GPU_PROBE_ALLREDUCE_MULTIMEM ? 32 : uint64_t(sms);
    grid = ctas > 0 ? unsigned(ctas) : unsigned(want < cap ? want : cap);
    if (grid > unsigned(sms)) grid = unsigned(sms);
  }
  *algo_io = algo;
  *a_out = a;
  *grid_out = grid;
  return 0;
}

extern "C" int gpu_probe_allreduce_f32(gpu_probe_exchange *x, uint64_t offset_bytes, uint64_t count, int algo,
                                       int ctas, void *stream) {
  AllreduceArgs a;
  unsigned grid = 1;
  int rc = plan_allreduce(x, offset_bytes, count, &algo, ctas, &a, &grid);
  if (rc) return rc;
  if (count == 0) return 0;  // every rank passes the same count, so every rank skips
  if (algo == GPU_PROBE_ALLREDUCE_PUSH) a.push_call = ++x->push_calls;  // counted only once nothing can fail any more
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (x->world) {
    case 1: return launch_allreduce<1>(a, algo, grid, st);
    case 2: return launch_allreduce<2>(a, algo, grid, st);
    case 3: return launch_allreduce<3>(a, algo, grid, st);
    case 4: return launch_allreduce<4>(a, algo, grid, st);
    case 5: return launch_allreduce<5>(a, algo, grid, st);
    case 6: return launch_allreduce<6>(a, algo, grid, st);
    case 7: return launch_allreduce<7>(a, algo, grid, st);
    default: return launch_allreduce<8>(a, algo, grid, st);
  }
}

template <int G>
static int launch_group(const GroupArgs &ga, int algo, unsigned grid, cudaStream_t st) {
  auto k = allreduce_f32_group_kernel<G>;
  int per_sm = 0, dev = 0, sms = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, kArThreads, 0);
  if (e == cudaSuccess) e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return int(e);
  int rc = device_sms(dev, &sms);
  if (rc) return rc;
  if (uint64_t(grid) * G > uint64_t(per_sm) * sms) return GPU_PROBE_ETOOBIG;  // all ranks' CTAs must be resident together
  void *params[] = {const_cast<GroupArgs *>(&ga), &algo};
  e = cudaLaunchCooperativeKernel(reinterpret_cast<const void *>(k), dim3(grid, G), dim3(kArThreads), params, 0, st);
  return int(e);
}

extern "C" int gpu_probe_allreduce_f32_group(gpu_probe_exchange *const *xs, int world, uint64_t offset_bytes,
                                             uint64_t count, int algo, int ctas, void *stream) {
  if (!xs || world < 1 || world > 8) return GPU_PROBE_EINVAL;
  if (algo == GPU_PROBE_ALLREDUCE_MULTIMEM) return GPU_PROBE_EINVAL;  // one device cannot form a multicast group
  GroupArgs ga;
  unsigned grid = 1;
  int algo0 = algo;
  for (int r = 0; r < world; ++r) {
    if (!live(xs[r])) return GPU_PROBE_ENOTREADY;
    if (xs[r]->world != world || xs[r]->rank != r) return GPU_PROBE_EINVAL;
    int algo_r = algo;
    unsigned grid_r = 1;
    int rc = plan_allreduce(xs[r], offset_bytes, count, &algo_r, ctas, &ga.r[r], &grid_r);  // also checks: all on the current device
    if (rc) return rc;
    if (r == 0) { algo0 = algo_r; grid = grid_r; }
    else if (algo_r != algo0 || grid_r != grid) return GPU_PROBE_EINVAL;
  }
  if (count == 0) return 0;
  if (algo0 == GPU_PROBE_ALLREDUCE_PUSH)
    for (int r = 0; r < world; ++r) ga.r[r].push_call = ++xs[r]->push_calls;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (world) {
    case 1: return launch_group<1>(ga, algo0, grid, st);
    case 2: return launch_group<2>(ga, algo0, grid, st);
    case 3: return launch_group<3>(ga, algo0, grid, st);
    case 4: return launch_group<4>(ga, algo0, grid, st);
    case 5: return launch_group<5>(ga, algo0, grid, st);
    case 6: return launch_group<6>(ga, algo0, grid, st);
    case 7: return launch_group<7>(ga, algo0, grid, st);
    default: return launch_group<8>(ga, algo0, grid, st);
  }
}
