// probe_common.cuh -- device helpers shared by the probe kernels (sm_100a only).
//
// Spec: SURVEY.md section 8(d).  No reference counterpart exists: the slot these
// kernels fill is the container command at
// /root/reference/pods/nvidia-gpu-test-pod.yaml:9.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace probe {

// ---- spec arithmetic --------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t splitmix64_at(uint64_t i, uint64_t seed) {
  uint64_t z = seed + (i + 1) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__device__ __forceinline__ double unit_double(uint64_t x) {
  // 53-bit integer -> double is exact; the power-of-two scale is exact.
  return __ull2double_rn(x >> 11) * 0x1.0p-53;
}

// ---- streaming global accesses ----------------------------------------------
// Everything on this path is touched once, so nothing may displace useful lines.
// 128-bit forms use .cs (evict-first in L1 and L2: SASS LDG.E.EF.128); the 256-bit
// forms are sm_100+ (LDG.E.NA.EFL2.256) -- ptxas accepts .L2::evict_first on ld
// only for the 256-bit types.
struct alignas(16) f64x2 { double x, y; };
struct alignas(32) f64x4 { double x, y, z, w; };
struct alignas(16) u64x2 { uint64_t x, y; };
struct alignas(32) u64x4 { uint64_t x, y, z, w; };

__device__ __forceinline__ f64x2 ld_stream(const f64x2 *p) {
  f64x2 v;
  asm volatile("ld.global.cs.v2.f64 {%0,%1}, [%2];"
               : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}
__device__ __forceinline__ f64x4 ld_stream(const f64x4 *p) {
  f64x4 v;
  asm volatile("ld.global.L1::no_allocate.L2::evict_first.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ u64x2 ld_stream(const u64x2 *p) {
  u64x2 v;
  asm volatile("ld.global.cs.v2.u64 {%0,%1}, [%2];"
               : "=l"(v.x), "=l"(v.y) : "l"(p));
  return v;
}
__device__ __forceinline__ u64x4 ld_stream(const u64x4 *p) {
  u64x4 v;
  asm volatile("ld.global.L1::no_allocate.L2::evict_first.v4.u64 {%0,%1,%2,%3}, [%4];"
               : "=l"(v.x), "=l"(v.y), "=l"(v.z), "=l"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream(f64x2 *p, const f64x2 &v) {
  asm volatile("st.global.cs.v2.f64 [%0], {%1,%2};"
               :: "l"(p), "d"(v.x), "d"(v.y) : "memory");
}
__device__ __forceinline__ void st_stream(f64x4 *p, const f64x4 &v) {
  asm volatile("st.global.L1::no_allocate.v4.f64 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "d"(v.x), "d"(v.y), "d"(v.z), "d"(v.w) : "memory");
}
__device__ __forceinline__ void st_stream(u64x2 *p, const u64x2 &v) {
  asm volatile("st.global.cs.v2.u64 [%0], {%1,%2};"
               :: "l"(p), "l"(v.x), "l"(v.y) : "memory");
}
__device__ __forceinline__ void st_stream(u64x4 *p, const u64x4 &v) {
  asm volatile("st.global.L1::no_allocate.v4.u64 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "l"(v.x), "l"(v.y), "l"(v.z), "l"(v.w) : "memory");
}

// ---- lanes ------------------------------------------------------------------
struct Lanes {
  uint64_t s0, s1;
  __device__ __forceinline__ void add(uint64_t w, uint64_t m) { s0 += w; s1 += w * m; }
};

__device__ __forceinline__ Lanes warp_sum(Lanes v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    v.s0 += __shfl_xor_sync(0xffffffffu, v.s0, o);
    v.s1 += __shfl_xor_sync(0xffffffffu, v.s1, o);
  }
  return v;
}

// Block sum; result valid in thread 0.  `smem` holds 2*32 words.
__device__ __forceinline__ Lanes block_sum(Lanes v, uint64_t *smem) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarps = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  if (lane == 0) { smem[2 * warp] = v.s0; smem[2 * warp + 1] = v.s1; }
  __syncthreads();
  if (warp == 0) {
    Lanes t{0, 0};
    if (lane < nwarps) { t.s0 = smem[2 * lane]; t.s1 = smem[2 * lane + 1]; }
    v = warp_sum(t);
  }
  __syncthreads();  // smem reusable afterwards
  return v;
}

// Workspace layout (device scratch, see gpu_probe_checksum_workspace_bytes):
//   word 0            : ticket counter (zero between launches)
//   words 32 ..       : per-CTA partial lanes, 2 words per CTA
constexpr int kMaxPartialCtas = 4096;
constexpr int kPartialsOffsetWords = 32;  // keep the ticket on its own 256-B line
constexpr size_t kWorkspaceBytes =
    sizeof(uint64_t) * (kPartialsOffsetWords + 2 * size_t(kMaxPartialCtas));

// ---- lanes exchange over peer memory (fused checksum + collective) ------------
// The path's one collective is a 16-byte sum.  Instead of a separate NCCL launch the
// last CTA of the checksum kernel stores this GPU's lanes straight into a mailbox slot
// on every peer (NVLink P2P stores), then waits for the peers' slots in its OWN mailbox
// and adds them.  Wrapping u64 sums commute, so arrival order cannot change a bit.
//
// Segment layout (one per GPU, allocated by gpu_probe_exchange_create; byte offsets):
//   0      mailbox   word 0: call counter of this GPU (local only)
//                    word 1: sticky error flag (local only; 1 = a peer never arrived)
//                    words 32..: slots[2][kMaxRanks] of {lane0, lane1, epoch, pad}; the buffer
//                    is chosen by epoch parity.  Two buffers suffice: a rank can only start
//                    call k+2 after every peer has written call k+1, which each peer does
//                    after it finished reading call k.
//   8 KiB   cta_epoch[kMaxCommCtas]           barrier epoch of CTA b on this GPU (local only)
//   16 KiB  (unused since round 2: the one-shot allreduce's call count lives on the host)
//   32 KiB  flags[kMaxCommCtas][kMaxRanks]    flags[b][p] = last barrier epoch CTA b of rank p
//                                             signalled to this GPU (written by p, polled here)
//   160 KiB push_flags[kMaxCommCtas][kMaxRanks]  same, for the one-shot allreduce
//   512 KiB staging[2][world][stage_bytes]    one-shot allreduce: rank p's operand lands in slot p
//   512 KiB + 2*world*stage_bytes: payload    the symmetric data buffer (allreduce operand)
constexpr int kMaxRanks = 16;
constexpr int kSlotWords = 4;
constexpr int kMailboxSlotsOffsetWords = 32;
constexpr size_t kMailboxBytes = sizeof(uint64_t) * (kMailboxSlotsOffsetWords + 2 * kMaxRanks * kSlotWords);
constexpr int kMaxCommCtas = 1024;
constexpr size_t kCtaEpochOffsetWords = 8192 / 8;
constexpr size_t kPushEpochOffsetWords = 16384 / 8;
constexpr size_t kFlagsOffsetWords = 32768 / 8;
constexpr size_t kPushFlagsOffsetWords = 163840 / 8;
constexpr size_t kHeaderBytes = 512 * 1024;
constexpr size_t kMaxStageBytes = 2u << 20;  // largest operand the one-shot allreduce takes
static_assert(kMailboxBytes <= 8192, "mailbox overlaps the epoch table");
static_assert(kFlagsOffsetWords + size_t(kMaxCommCtas) * kMaxRanks <= kPushFlagsOffsetWords, "flags overlap");
static_assert((kPushFlagsOffsetWords + size_t(kMaxCommCtas) * kMaxRanks) * 8 <= kHeaderBytes, "flags overlap the staging");

struct ExchangeArgs {           // passed to kernels by value; world == 0: no exchange
  uint64_t *mailbox[kMaxRanks]; // mailbox[p]: rank p's segment base as THIS GPU addresses it
  int rank, world;
  unsigned long long timeout_ns;
};

__device__ __forceinline__ void st_release_sys(uint64_t *p, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(uint64_t *p, uint64_t v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t ld_relaxed_sys(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Called by every thread of the LAST CTA; (s0, s1) = this GPU's final lanes, valid in
// thread 0.  Writes the sum over all ranks to out[0..1].  `smem` holds >= 4 words.
__device__ __forceinline__ void exchange_lanes(uint64_t s0, uint64_t s1, const ExchangeArgs &x,
                                               uint64_t *__restrict__ out, uint64_t *smem) {
  uint64_t *mine = x.mailbox[x.rank];
  if (threadIdx.x == 0) {
    const uint64_t epoch = mine[0] + 1;  // calls on one GPU are stream-ordered: no race
    mine[0] = epoch;
    smem[0] = s0; smem[1] = s1; smem[2] = epoch;
  }
  __syncthreads();
  if (threadIdx.x >= 32) return;  // warp 0 talks to the peers, one lane per rank
  const uint64_t epoch = smem[2];
  const size_t buf = kMailboxSlotsOffsetWords + size_t(epoch & 1) * kMaxRanks * kSlotWords;
  uint64_t v0 = 0, v1 = 0;
  int failed = 0;
  const int p = threadIdx.x;
  if (p < x.world) {
    uint64_t *theirs = x.mailbox[p] + buf + size_t(x.rank) * kSlotWords;  // my slot in p's mailbox
    st_relaxed_sys(theirs, smem[0]);
    st_relaxed_sys(theirs + 1, smem[1]);
    st_release_sys(theirs + 2, epoch);  // release: the two lanes are visible before the flag
    const uint64_t *slot = mine + buf + size_t(p) * kSlotWords;  // p's slot in my mailbox
    const unsigned long long t0 = global_ns();
    while (ld_acquire_sys(slot + 2) != epoch) {
      if (global_ns() - t0 > x.timeout_ns) { failed = 1; break; }
    }
    v0 = ld_relaxed_sys(slot);
    v1 = ld_relaxed_sys(slot + 1);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    v0 += __shfl_xor_sync(0xffffffffu, v0, o);
    v1 += __shfl_xor_sync(0xffffffffu, v1, o);
    failed |= __shfl_xor_sync(0xffffffffu, failed, o);
  }
  if (threadIdx.x == 0) {
    if (failed) { mine[1] = 1; v0 = ~0ull; v1 = ~0ull; }  // loud: no plausible lanes on a timeout
    out[0] = v0;
    out[1] = v1;
  }
}

// Every CTA publishes its partial; the last one to take a ticket adds them up,
// applies the global-index fix-up and overwrites out[0..1] -- with this GPU's lanes,
// or, when an exchange is attached, with the sum over all ranks' lanes.
// Wrapping u64 sums are associative and commutative, so the order CTAs (or GPUs)
// finish in cannot change a bit.
//   lane1 was accumulated with LOCAL indices k; sum w*(2(f+k)+1) = lane1 + 2f*lane0.
// `cta` / `nctas`: this CTA's index and the CTA count of ITS RANK -- blockIdx.x / gridDim.x in
// the product kernels (one rank per launch); the single-GPU group self-test runs several ranks
// in one launch and passes blockIdx.x / gridDim.x of the rank's row of the grid.
__device__ __forceinline__ void finish_lanes_at(Lanes mine, uint64_t first_index,
                                                uint64_t *__restrict__ out,
                                                uint64_t *__restrict__ ws, uint64_t *smem,
                                                const ExchangeArgs &x, unsigned cta, unsigned nctas) {
  __shared__ bool is_last;
  Lanes blk = block_sum(mine, smem);
  if (threadIdx.x == 0) {
    uint64_t *part = ws + kPartialsOffsetWords + 2 * size_t(cta);
    __stcg(part, blk.s0);
    __stcg(part + 1, blk.s1);
    __threadfence();
    unsigned long long t = atomicAdd(reinterpret_cast<unsigned long long *>(ws), 1ull);
    is_last = (t == nctas - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  Lanes acc{0, 0};
  for (unsigned i = threadIdx.x; i < nctas; i += blockDim.x) {
    const uint64_t *part = ws + kPartialsOffsetWords + 2 * size_t(i);
    acc.s0 += __ldcg(part);
    acc.s1 += __ldcg(part + 1);
  }
  acc = block_sum(acc, smem);
  const uint64_t l0 = acc.s0, l1 = acc.s1 + 2 * first_index * acc.s0;
  if (threadIdx.x == 0) ws[0] = 0;  // ticket back to zero for the next launch on this workspace
  if (x.world > 0) {
    exchange_lanes(l0, l1, x, out, smem);
  } else if (threadIdx.x == 0) {
    out[0] = l0;
    out[1] = l1;
  }
}

__device__ __forceinline__ void finish_lanes(Lanes mine, uint64_t first_index,
                                             uint64_t *__restrict__ out,
                                             uint64_t *__restrict__ ws, uint64_t *smem,
                                             const ExchangeArgs &x) {
  finish_lanes_at(mine, first_index, out, ws, smem, x, blockIdx.x, gridDim.x);
}

}  // namespace probe
