// tools/tune/tune_triad.cu -- parameter sweep for the triad kernels (development tool, not
// shipped).  Runtime-parameterised copies of the TMA ring and LDG kernels so one binary can
// sweep tile size, stage count, CTAs/SM, block size, store path and array placement.
// Build: make -C tools/tune     Run (GPU box): tools/tune/tune_triad [log2n] > gpurun_out/tune.csv
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>

#include "gpu_probe.h"
#include "probe_common.cuh"
#include "probe_tma.cuh"

using namespace probe;

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e__ = (x);                                                         \
    if (e__ != cudaSuccess) {                                                      \
      fprintf(stderr, "%s:%d %s -> %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e__)); \
      exit(1);                                                                     \
    }                                                                              \
  } while (0)

enum StoreMode { BULK_NONE = 0, BULK_EF = 1, BULK_EL = 2, STG_CS = 3, STG_DEF = 4, STG_NA = 5 };

__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_store_hint(void *dst, const void *src, uint32_t bytes, uint64_t pol) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
               :: "l"(dst), "r"(tma::smem_addr(src)), "r"(bytes), "l"(pol) : "memory");
}

// load_hint: 0 none(default policy via evict_normal), 1 evict_first
template <int STORE>
__global__ void triad_tma_rt(double *__restrict__ a, const double *__restrict__ b, const double *__restrict__ c,
                             double s, uint64_t n, int tile, int S, int SO, int blocked, int load_ef) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double *sb = reinterpret_cast<double *>(smem_raw);
  double *sc = sb + size_t(S) * tile;
  double *so = sc + size_t(S) * tile;
  constexpr bool kBulk = STORE <= BULK_EL;
  uint64_t *full = reinterpret_cast<uint64_t *>(so + (kBulk ? size_t(SO) * tile : 0));
  const uint32_t tile_bytes = uint32_t(tile) * 8u;
  const uint64_t ntiles = n / tile;
  const int tid = threadIdx.x;
  // tile assignment: interleaved t = cta + j*grid, or blocked contiguous ranges
  uint64_t my_tiles, t0, tstride;
  if (blocked) {
    const uint64_t per = (ntiles + gridDim.x - 1) / gridDim.x;
    t0 = per * blockIdx.x;
    my_tiles = t0 >= ntiles ? 0 : (ntiles - t0 < per ? ntiles - t0 : per);
    tstride = 1;
  } else {
    t0 = blockIdx.x;
    my_tiles = ntiles > blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    tstride = gridDim.x;
  }
  uint64_t lpol = 0, spol = 0;
  if (tid == 0) {
    for (int i = 0; i < S; ++i) tma::mbar_init(&full[i], 1);
    tma::fence_barrier_init();
  }
  __syncthreads();
  auto issue = [&](uint64_t j) {
    const int st = int(j % S);
    const uint64_t t = t0 + j * tstride;
    tma::mbar_arrive_expect_tx(&full[st], 2 * tile_bytes);
    tma::bulk_load(sb + size_t(st) * tile, b + t * tile, tile_bytes, &full[st], lpol);
    tma::bulk_load(sc + size_t(st) * tile, c + t * tile, tile_bytes, &full[st], lpol);
  };
  if (tid == 0) {
    if (load_ef) lpol = tma::policy_evict_first();
    else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(lpol));
    spol = STORE == BULK_EF ? tma::policy_evict_first() : policy_evict_last();
    for (uint64_t j = 0; j < my_tiles && j < uint64_t(S); ++j) issue(j);
  }
  for (uint64_t j = 0; j < my_tiles; ++j) {
    const int st = int(j % S), ob = int(j % SO);
    const uint32_t parity = uint32_t(j / S) & 1u;
    const uint64_t t = t0 + j * tstride;
    tma::mbar_wait(&full[st], parity);
    const f64x2 *pb = reinterpret_cast<const f64x2 *>(sb + size_t(st) * tile);
    const f64x2 *pc = reinterpret_cast<const f64x2 *>(sc + size_t(st) * tile);
    if (kBulk) {
      f64x2 *po = reinterpret_cast<f64x2 *>(so + size_t(ob) * tile);
      for (int v = tid; v < tile / 2; v += blockDim.x) {
        const f64x2 x = pb[v], y = pc[v];
        po[v] = f64x2{__fma_rn(s, y.x, x.x), __fma_rn(s, y.y, x.y)};
      }
      tma::fence_proxy_async_smem();
      if (tid == 0) {
        if (SO == 2) tma::bulk_wait_read<0>(); else if (SO == 3) tma::bulk_wait_read<1>(); else tma::bulk_wait_read<2>();
      }
      __syncthreads();
      if (tid == 0) {
        if (STORE == BULK_NONE) tma::bulk_store(a + t * tile, po, tile_bytes);
        else bulk_store_hint(a + t * tile, po, tile_bytes, spol);
        tma::bulk_commit();
        if (j + S < my_tiles) issue(j + S);
      }
    } else {
      f64x2 *ga = reinterpret_cast<f64x2 *>(a + t * tile);
      for (int v = tid; v < tile / 2; v += blockDim.x) {
        const f64x2 x = pb[v], y = pc[v];
        const f64x2 r{__fma_rn(s, y.x, x.x), __fma_rn(s, y.y, x.y)};
        if (STORE == STG_CS) st_stream(ga + v, r);
        else if (STORE == STG_NA) asm volatile("st.global.L1::no_allocate.v2.f64 [%0], {%1,%2};" :: "l"(ga + v), "d"(r.x), "d"(r.y) : "memory");
        else ga[v] = r;
      }
      __syncthreads();
      if (tid == 0 && j + S < my_tiles) issue(j + S);
    }
  }
  if (kBulk && tid == 0) tma::bulk_wait_all<0>();
}


// compile-time geometry copy of the shipped kernel, with a switchable consumer-loop shape
// MODE 0: fully unrolled (all LDS, then DFMA, then STG)   MODE 1: #pragma unroll 1
// MODE 2: unrolled, but the elected thread refills BEFORE the CTA barrier is replaced by
//         per-stage "empty" mbarrier (consumers arrive, producer waits) -- no __syncthreads
template <int TILE, int S, int MODE>
__global__ void __launch_bounds__(256)
triad_tma_ct(double *__restrict__ a, const double *__restrict__ b, const double *__restrict__ c, double s, uint64_t n) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double *sb = reinterpret_cast<double *>(smem_raw);
  double *sc = sb + size_t(S) * TILE;
  uint64_t *full = reinterpret_cast<uint64_t *>(sc + size_t(S) * TILE);
  constexpr uint32_t kTileBytes = TILE * 8;
  const uint64_t ntiles = n / TILE;
  const uint64_t my_tiles = ntiles > blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const int tid = threadIdx.x;
  uint64_t policy = 0;
  if (tid == 0) {
    for (int i = 0; i < S; ++i) tma::mbar_init(&full[i], 1);
    tma::fence_barrier_init();
  }
  __syncthreads();
  auto issue = [&](uint64_t j) {
    const int st = int(j % S);
    const uint64_t t = blockIdx.x + j * gridDim.x;
    tma::mbar_arrive_expect_tx(&full[st], 2 * kTileBytes);
    tma::bulk_load(sb + size_t(st) * TILE, b + t * TILE, kTileBytes, &full[st], policy);
    tma::bulk_load(sc + size_t(st) * TILE, c + t * TILE, kTileBytes, &full[st], policy);
  };
  if (tid == 0) {
    policy = tma::policy_evict_first();
    for (uint64_t j = 0; j < my_tiles && j < uint64_t(S); ++j) issue(j);
  }
  for (uint64_t j = 0; j < my_tiles; ++j) {
    const int st = int(j % S);
    const uint32_t parity = uint32_t(j / S) & 1u;
    const uint64_t t = blockIdx.x + j * gridDim.x;
    tma::mbar_wait(&full[st], parity);
    const f64x2 *pb = reinterpret_cast<const f64x2 *>(sb + size_t(st) * TILE);
    const f64x2 *pc = reinterpret_cast<const f64x2 *>(sc + size_t(st) * TILE);
    f64x2 *ga = reinterpret_cast<f64x2 *>(a + t * TILE);
    if (MODE == 1) {
#pragma unroll 1
      for (int v = tid; v < TILE / 2; v += 256) {
        const f64x2 x = pb[v], y = pc[v];
        st_stream(ga + v, f64x2{__fma_rn(s, y.x, x.x), __fma_rn(s, y.y, x.y)});
      }
    } else {
#pragma unroll
      for (int v = tid; v < TILE / 2; v += 256) {
        const f64x2 x = pb[v], y = pc[v];
        st_stream(ga + v, f64x2{__fma_rn(s, y.x, x.x), __fma_rn(s, y.y, x.y)});
      }
    }
    __syncthreads();
    if (tid == 0 && j + S < my_tiles) issue(j + S);
  }
}


// Division-free runtime-geometry kernel (stage / parity / tile index carried incrementally)
// so the consumer is fast at every geometry, plus an optional pacing knob: the elected
// thread will not refill a slot earlier than `pace` SM cycles after its previous refill.
template <int UNROLL>
__global__ void __launch_bounds__(256)
triad_tma_fast(double *__restrict__ a, const double *__restrict__ b, const double *__restrict__ c, double s,
               uint64_t n, int tile, int S, int pace) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double *sb = reinterpret_cast<double *>(smem_raw);
  double *sc = sb + size_t(S) * tile;
  uint64_t *full = reinterpret_cast<uint64_t *>(sc + size_t(S) * tile);
  const uint32_t tile_bytes = uint32_t(tile) * 8u;
  const uint64_t ntiles = n / tile;
  const uint64_t my_tiles = ntiles > blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const int tid = threadIdx.x;
  uint64_t policy = 0;
  if (tid == 0) {
    for (int i = 0; i < S; ++i) tma::mbar_init(&full[i], 1);
    tma::fence_barrier_init();
  }
  __syncthreads();
  // producer state (thread 0): next tile to request and the slot it goes to
  uint64_t pj = 0;
  int pst = 0;
  long long next_ok = 0;
  auto issue = [&]() {
    const uint64_t t = blockIdx.x + pj * gridDim.x;
    tma::mbar_arrive_expect_tx(&full[pst], 2 * tile_bytes);
    tma::bulk_load(sb + size_t(pst) * tile, b + t * tile, tile_bytes, &full[pst], policy);
    tma::bulk_load(sc + size_t(pst) * tile, c + t * tile, tile_bytes, &full[pst], policy);
    ++pj;
    if (++pst == S) pst = 0;
  };
  if (tid == 0) {
    policy = tma::policy_evict_first();
    for (int i = 0; i < S && pj < my_tiles; ++i) issue();
    next_ok = clock64() + pace;
  }
  int st = 0;
  uint32_t parity = 0;
  uint64_t t = blockIdx.x;
  for (uint64_t j = 0; j < my_tiles; ++j, t += gridDim.x) {
    tma::mbar_wait(&full[st], parity);
    const f64x2 *pb = reinterpret_cast<const f64x2 *>(sb + size_t(st) * tile);
    const f64x2 *pc = reinterpret_cast<const f64x2 *>(sc + size_t(st) * tile);
    f64x2 *ga = reinterpret_cast<f64x2 *>(a + t * tile);
#pragma unroll UNROLL
    for (int v = tid; v < tile / 2; v += 256) {
      const f64x2 x = pb[v], y = pc[v];
      st_stream(ga + v, f64x2{__fma_rn(s, y.x, x.x), __fma_rn(s, y.y, x.y)});
    }
    __syncthreads();
    if (tid == 0 && pj < my_tiles) {
      if (pace > 0) {
        while (clock64() < next_ok) { }
        next_ok += pace;
        const long long now = clock64();
        if (next_ok < now) next_ok = now;  // never bank credit: a late refill does not allow a burst
      }
      issue();
    }
    if (++st == S) { st = 0; parity ^= 1u; }
  }
}

// LDG variants: V = 16 or 32 bytes, U loads per array per thread
template <typename V, int U, int LD>  // LD: 0 = .cs / NA.EFL2, 1 = default ld.global, 2 = ld.global.nc
__global__ void triad_ldg_rt(V *__restrict__ a, const V *__restrict__ b, const V *__restrict__ c, double s, uint64_t nvec) {
  const uint64_t tile = uint64_t(blockDim.x) * U;
  for (uint64_t base = blockIdx.x * tile; base + tile <= nvec; base += gridDim.x * tile) {
    const uint64_t i0 = base + threadIdx.x;
    V vb[U], vc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) vb[u] = LD == 0 ? ld_stream(b + i0 + uint64_t(u) * blockDim.x) : b[i0 + uint64_t(u) * blockDim.x];
#pragma unroll
    for (int u = 0; u < U; ++u) vc[u] = LD == 0 ? ld_stream(c + i0 + uint64_t(u) * blockDim.x) : c[i0 + uint64_t(u) * blockDim.x];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      V r;
      double *pr = reinterpret_cast<double *>(&r);
      const double *pb = reinterpret_cast<const double *>(&vb[u]), *pc = reinterpret_cast<const double *>(&vc[u]);
#pragma unroll
      for (int e = 0; e < int(sizeof(V) / 8); ++e) pr[e] = __fma_rn(s, pc[e], pb[e]);
      if (LD == 0) st_stream(a + i0 + uint64_t(u) * blockDim.x, r); else a[i0 + uint64_t(u) * blockDim.x] = r;
    }
  }
}

// pure read / pure write references
__global__ void read_only(const u64x4 *__restrict__ w, uint64_t nvec, uint64_t *out) {
  uint64_t acc = 0;
  const uint64_t tile = uint64_t(blockDim.x) * 4;
  for (uint64_t base = blockIdx.x * tile; base + tile <= nvec; base += gridDim.x * tile) {
    u64x4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = ld_stream(w + base + threadIdx.x + uint64_t(u) * blockDim.x);
#pragma unroll
    for (int u = 0; u < 4; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x1234567) *out = acc;
}
__global__ void write_only(f64x4 *__restrict__ w, uint64_t nvec) {
  const uint64_t tile = uint64_t(blockDim.x) * 4;
  for (uint64_t base = blockIdx.x * tile; base + tile <= nvec; base += gridDim.x * tile) {
#pragma unroll
    for (int u = 0; u < 4; ++u) st_stream(w + base + threadIdx.x + uint64_t(u) * blockDim.x, f64x4{1.0, 2.0, 3.0, 4.0});
  }
}


// ---- checksum sweeps --------------------------------------------------------
__device__ __forceinline__ void acc2(Lanes &acc, const u64x2 &v, uint64_t k0) {
  const uint64_t m = 2 * k0 + 1; acc.add(v.x, m); acc.add(v.y, m + 2);
}
__device__ __forceinline__ void acc4(Lanes &acc, const u64x4 &v, uint64_t k0) {
  const uint64_t m = 2 * k0 + 1; acc.add(v.x, m); acc.add(v.y, m + 2); acc.add(v.z, m + 4); acc.add(v.w, m + 6);
}
__global__ void checksum_tma_rt(const uint64_t *__restrict__ w, uint64_t n, uint64_t *__restrict__ out,
                                uint64_t *__restrict__ ws, int tile, int S) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t *sw = reinterpret_cast<uint64_t *>(smem_raw);
  uint64_t *full = sw + size_t(S) * tile;
  __shared__ uint64_t red[64];
  const uint32_t tile_bytes = uint32_t(tile) * 8u;
  const uint64_t ntiles = n / tile;
  const uint64_t my_tiles = ntiles > blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const int tid = threadIdx.x;
  uint64_t policy = 0;
  if (tid == 0) {
    for (int i = 0; i < S; ++i) tma::mbar_init(&full[i], 1);
    tma::fence_barrier_init();
  }
  __syncthreads();
  auto issue = [&](uint64_t j) {
    const int st = int(j % S);
    const uint64_t t = blockIdx.x + j * gridDim.x;
    tma::mbar_arrive_expect_tx(&full[st], tile_bytes);
    tma::bulk_load(sw + size_t(st) * tile, w + t * tile, tile_bytes, &full[st], policy);
  };
  if (tid == 0) {
    policy = tma::policy_evict_first();
    for (uint64_t j = 0; j < my_tiles && j < uint64_t(S); ++j) issue(j);
  }
  Lanes acc{0, 0};
  for (uint64_t j = 0; j < my_tiles; ++j) {
    const int st = int(j % S);
    const uint32_t parity = uint32_t(j / S) & 1u;
    const uint64_t t = blockIdx.x + j * gridDim.x;
    tma::mbar_wait(&full[st], parity);
    const u64x2 *pw = reinterpret_cast<const u64x2 *>(sw + size_t(st) * tile);
    for (int v = tid; v < tile / 2; v += blockDim.x) acc2(acc, pw[v], t * tile + 2 * uint64_t(v));
    __syncthreads();
    if (tid == 0 && j + S < my_tiles) issue(j + S);
  }
  finish_lanes(acc, 0, out, ws, red, ExchangeArgs{});
}
template <typename V, int U>
__global__ void checksum_ldg_rt(const V *__restrict__ w, uint64_t nvec, uint64_t *__restrict__ out, uint64_t *__restrict__ ws) {
  constexpr int E = int(sizeof(V) / 8);
  __shared__ uint64_t red[64];
  Lanes acc{0, 0};
  const uint64_t tile = uint64_t(blockDim.x) * U;
  for (uint64_t base = blockIdx.x * tile; base + tile <= nvec; base += gridDim.x * tile) {
    const uint64_t i0 = base + threadIdx.x;
    V v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = ld_stream(w + i0 + uint64_t(u) * blockDim.x);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if constexpr (E == 2) acc2(acc, *reinterpret_cast<u64x2 *>(&v[u]), (i0 + uint64_t(u) * blockDim.x) * E);
      else acc4(acc, *reinterpret_cast<u64x4 *>(&v[u]), (i0 + uint64_t(u) * blockDim.x) * E);
    }
  }
  finish_lanes(acc, 0, out, ws, red, ExchangeArgs{});
}

struct Timer {
  cudaEvent_t e0, e1;
  Timer() { CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); }
  template <typename F>
  void run(F f, int warm, int iters, float *best, float *med) {
    for (int i = 0; i < warm; ++i) f();
    CK(cudaDeviceSynchronize());
    std::vector<float> ms(iters);
    for (int i = 0; i < iters; ++i) {
      CK(cudaEventRecord(e0));
      f();
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      CK(cudaEventElapsedTime(&ms[i], e0, e1));
    }
    std::sort(ms.begin(), ms.end());
    *best = ms[0];
    *med = ms[iters / 2];
  }
};


// Adaptive pacing ("congestion control" for HBM): thread 0 spaces its refills `pace` cycles apart
// and adapts the pace from what the consumers see.  If the data of the next tile is not there
// when the CTA needs it (stall at the mbarrier), demand exceeds what the memory system delivers
// -> slow down (additive increase of the pace); after `calm_n` tiles without a stall -> speed up.
// Fixed pacing reached 7110 GB/s at exactly the right pace and collapsed 2.5 % beyond it, and
// the right pace differs per GPU; the controller is meant to find it on its own.
__global__ void __launch_bounds__(256)
triad_tma_adapt(double *__restrict__ a, const double *__restrict__ b, const double *__restrict__ c, double s,
                uint64_t n, int tile, int S, int pace0, int inc, int dec, int thresh, int calm_n, int pace_min,
                int *__restrict__ pace_out) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double *sb = reinterpret_cast<double *>(smem_raw);
  double *sc = sb + size_t(S) * tile;
  uint64_t *full = reinterpret_cast<uint64_t *>(sc + size_t(S) * tile);
  const uint32_t tile_bytes = uint32_t(tile) * 8u;
  const uint64_t ntiles = n / tile;
  const uint64_t my_tiles = ntiles > blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const int tid = threadIdx.x;
  uint64_t policy = 0;
  if (tid == 0) {
    for (int i = 0; i < S; ++i) tma::mbar_init(&full[i], 1);
    tma::fence_barrier_init();
  }
  __syncthreads();
  uint64_t pj = 0;
  int pst = 0;
  long long next_ok = 0;
  int pace = pace0, calm = 0;
  long long pace_sum = 0;
  auto issue = [&]() {
    const uint64_t t = blockIdx.x + pj * gridDim.x;
    tma::mbar_arrive_expect_tx(&full[pst], 2 * tile_bytes);
    tma::bulk_load(sb + size_t(pst) * tile, b + t * tile, tile_bytes, &full[pst], policy);
    tma::bulk_load(sc + size_t(pst) * tile, c + t * tile, tile_bytes, &full[pst], policy);
    ++pj;
    if (++pst == S) pst = 0;
  };
  if (tid == 0) {
    policy = tma::policy_evict_first();
    for (int i = 0; i < S && pj < my_tiles; ++i) issue();
    next_ok = clock64() + pace;
  }
  int st = 0;
  uint32_t parity = 0;
  uint64_t t = blockIdx.x;
  for (uint64_t j = 0; j < my_tiles; ++j, t += gridDim.x) {
    long long t0 = 0;
    if (tid == 0) t0 = clock64();
    tma::mbar_wait(&full[st], parity);
    if (tid == 0) {
      const long long stall = clock64() - t0;
      if (pace_out) {
        const int bkt = stall < 100 ? 0 : stall < 200 ? 1 : stall < 400 ? 2 : stall < 800 ? 3 : stall < 1600 ? 4 : stall < 3200 ? 5 : 6;
        atomicAdd(&pace_out[8192 + bkt], 1);
      }
      if (stall > thresh) { pace += inc; calm = 0; }
      else if (++calm >= calm_n) { pace -= dec; calm = 0; if (pace < pace_min) pace = pace_min; }
      pace_sum += pace;
    }
    const f64x2 *pb = reinterpret_cast<const f64x2 *>(sb + size_t(st) * tile);
    const f64x2 *pc = reinterpret_cast<const f64x2 *>(sc + size_t(st) * tile);
    f64x2 *ga = reinterpret_cast<f64x2 *>(a + t * tile);
#pragma unroll 4
    for (int v = tid; v < tile / 2; v += 256) {
      const f64x2 x = pb[v], y = pc[v];
      st_stream(ga + v, f64x2{__fma_rn(s, y.x, x.x), __fma_rn(s, y.y, x.y)});
    }
    __syncthreads();
    if (tid == 0 && pj < my_tiles) {
      while (clock64() < next_ok) { }
      next_ok += pace;
      const long long now = clock64();
      if (next_ok < now) next_ok = now;  // never bank credit
      issue();
    }
    if (++st == S) { st = 0; parity ^= 1u; }
  }
  if (tid == 0 && pace_out) { pace_out[2 * blockIdx.x] = pace; pace_out[2 * blockIdx.x + 1] = my_tiles ? int(pace_sum / (long long)my_tiles) : 0; }
}


// Self-clocked pacing: refills are spaced by the CTA's own measured tile service time (EWMA of
// the intervals between tile completions), nudged so that the number of outstanding tiles
// stays at `target` (< S).  Demand can never run ahead of delivery for long (no fixed rate to
// overshoot), the ring bounds the damage, and arrivals at the memory system are smooth instead
// of ack-clocked bursts.
__global__ void __launch_bounds__(256)
triad_tma_selfclock(double *__restrict__ a, const double *__restrict__ b, const double *__restrict__ c, double s,
                    uint64_t n, int tile, int S, int target, int up1024, int down1024, int ewma_shift,
                    int *__restrict__ dbg) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double *sb = reinterpret_cast<double *>(smem_raw);
  double *sc = sb + size_t(S) * tile;
  uint64_t *full = reinterpret_cast<uint64_t *>(sc + size_t(S) * tile);
  const uint32_t tile_bytes = uint32_t(tile) * 8u;
  const uint64_t ntiles = n / tile;
  const uint64_t my_tiles = ntiles > blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const int tid = threadIdx.x;
  uint64_t policy = 0;
  if (tid == 0) {
    for (int i = 0; i < S; ++i) tma::mbar_init(&full[i], 1);
    tma::fence_barrier_init();
  }
  __syncthreads();
  uint64_t pj = 0;
  int pst = 0;
  auto issue = [&]() {
    const uint64_t t = blockIdx.x + pj * gridDim.x;
    tma::mbar_arrive_expect_tx(&full[pst], 2 * tile_bytes);
    tma::bulk_load(sb + size_t(pst) * tile, b + t * tile, tile_bytes, &full[pst], policy);
    tma::bulk_load(sc + size_t(pst) * tile, c + t * tile, tile_bytes, &full[pst], policy);
    ++pj;
    if (++pst == S) pst = 0;
  };
  long long last_done = 0, last_issue = 0, T = 0, out_sum = 0;
  if (tid == 0) {
    policy = tma::policy_evict_first();
    for (int i = 0; i < target && pj < my_tiles; ++i) issue();
    last_issue = clock64();
  }
  int st = 0;
  uint32_t parity = 0;
  uint64_t t = blockIdx.x;
  for (uint64_t j = 0; j < my_tiles; ++j, t += gridDim.x) {
    tma::mbar_wait(&full[st], parity);
    if (tid == 0) {
      const long long now = clock64();
      if (last_done) {
        const long long dt = now - last_done;
        T = T ? T + ((dt - T) >> ewma_shift) : dt;
      }
      last_done = now;
    }
    const f64x2 *pb = reinterpret_cast<const f64x2 *>(sb + size_t(st) * tile);
    const f64x2 *pc = reinterpret_cast<const f64x2 *>(sc + size_t(st) * tile);
    f64x2 *ga = reinterpret_cast<f64x2 *>(a + t * tile);
#pragma unroll 4
    for (int v = tid; v < tile / 2; v += 256) {
      const f64x2 x = pb[v], y = pc[v];
      st_stream(ga + v, f64x2{__fma_rn(s, y.x, x.x), __fma_rn(s, y.y, x.y)});
    }
    __syncthreads();  // slot st is free again
    if (tid == 0) {
      // outstanding = issued - consumed (tile j is consumed now)
      int outstanding = int(pj - (j + 1));
      out_sum += outstanding;
      // issue as many as the schedule allows, at most up to S-1 outstanding (one slot is being refilled next round)
      while (pj < my_tiles && outstanding < S - 1) {
        const long long gap = T ? (T * (outstanding >= target ? up1024 : down1024)) >> 10 : 0;
        const long long due = last_issue + gap;
        const long long now = clock64();
        if (now < due) {
          if (outstanding >= 1) break;            // data is on its way: do not sit here, come back after the next tile
          while (clock64() < due) { }              // nothing outstanding: wait out the gap rather than starve
        }
        issue();
        last_issue = clock64();
        ++outstanding;
      }
    }
    if (++st == S) { st = 0; parity ^= 1u; }
  }
  if (tid == 0 && dbg) { dbg[2 * blockIdx.x] = int(T); dbg[2 * blockIdx.x + 1] = my_tiles ? int(out_sum * 100 / (long long)my_tiles) : 0; }
}


// Optimistic pacing with a one-way ladder.  Refills are spaced `pace` cycles apart (smooth,
// slightly below what the memory system can sustain: 7100+ GB/s).  If the memory system cannot
// sustain it the kernel falls into the over-subscribed regime, which shows up as long consumer
// stalls (>= 800 cycles on ~16 % of the tiles instead of 0.3 %).  Then the CTA stops refilling,
// lets its ring drain, and restarts 2 % slower.  One-way: the pace never speeds up again.
__global__ void __launch_bounds__(256)
triad_tma_ladder(double *__restrict__ a, const double *__restrict__ b, const double *__restrict__ c, double s,
                 uint64_t n, int tile, int S, int pace0, int step1024, int stall_t, int bad_n, int win_n,
                 int *__restrict__ dbg) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double *sb = reinterpret_cast<double *>(smem_raw);
  double *sc = sb + size_t(S) * tile;
  uint64_t *full = reinterpret_cast<uint64_t *>(sc + size_t(S) * tile);
  const uint32_t tile_bytes = uint32_t(tile) * 8u;
  const uint64_t ntiles = n / tile;
  const uint64_t my_tiles = ntiles > blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const int tid = threadIdx.x;
  uint64_t policy = 0;
  if (tid == 0) {
    for (int i = 0; i < S; ++i) tma::mbar_init(&full[i], 1);
    tma::fence_barrier_init();
    policy = tma::policy_evict_first();
  }
  __syncthreads();
  uint64_t pj = 0;
  int pst = 0;
  auto issue = [&]() {
    const uint64_t t = blockIdx.x + pj * gridDim.x;
    tma::mbar_arrive_expect_tx(&full[pst], 2 * tile_bytes);
    tma::bulk_load(sb + size_t(pst) * tile, b + t * tile, tile_bytes, &full[pst], policy);
    tma::bulk_load(sc + size_t(pst) * tile, c + t * tile, tile_bytes, &full[pst], policy);
    ++pj;
    if (++pst == S) pst = 0;
  };
  long long next_ok = 0;
  int pace = pace0, bad = 0, win = 0, level = 0;
  bool draining = false;
  uint64_t grace_until = 0;
  int st = 0;
  uint32_t parity = 0;
  uint64_t t = blockIdx.x;
  for (uint64_t j = 0; j < my_tiles; ++j, t += gridDim.x) {
    long long t0 = 0;
    if (tid == 0) {
      if (pj <= j) {  // ring empty: first tile of the kernel, or the drain after a fallback is complete
        draining = false;
        for (int i = 0; i < S && pj < my_tiles; ++i) issue();
        next_ok = clock64() + pace;
        grace_until = j + 2 * S + 8;  // the start-up burst is not evidence of anything
        bad = win = 0;
      }
      t0 = clock64();
    }
    tma::mbar_wait(&full[st], parity);
    if (tid == 0 && !draining && j >= grace_until) {
      const long long stall = clock64() - t0;
      if (stall >= stall_t) ++bad;
      if (++win == win_n) {
        if (bad >= bad_n) {  // over-subscribed: drain, then restart one step slower
          draining = true;
          pace += (pace0 * step1024) >> 10;
          ++level;
        }
        bad = win = 0;
      }
    }
    const f64x2 *pb = reinterpret_cast<const f64x2 *>(sb + size_t(st) * tile);
    const f64x2 *pc = reinterpret_cast<const f64x2 *>(sc + size_t(st) * tile);
    f64x2 *ga = reinterpret_cast<f64x2 *>(a + t * tile);
#pragma unroll 4
    for (int v = tid; v < tile / 2; v += 256) {
      const f64x2 x = pb[v], y = pc[v];
      st_stream(ga + v, f64x2{__fma_rn(s, y.x, x.x), __fma_rn(s, y.y, x.y)});
    }
    __syncthreads();
    if (tid == 0 && !draining && pj < my_tiles && pj > j + 1) {  // (pj == j+1: ring just emptied, restart at the loop top)
      while (clock64() < next_ok) { }
      next_ok += pace;
      const long long now = clock64();
      if (next_ok < now) next_ok = now;  // never bank credit
      issue();
    }
    if (++st == S) { st = 0; parity ^= 1u; }
  }
  if (tid == 0 && dbg) { dbg[2 * blockIdx.x] = pace; dbg[2 * blockIdx.x + 1] = level; }
}


__global__ void globaltimer_probe(unsigned long long *out) {
  unsigned long long prev = 0, t;
  int k = 0;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(prev));
  const long long c0 = clock64();
  while (k < 64) {
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    if (t != prev) { out[k++] = t - prev; prev = t; }
  }
  out[64] = (unsigned long long)(clock64() - c0);
}

// Ladder kernel with the pace given in NANOSECONDS: thread 0 keeps a running estimate of the SM
// clock (clock64 against %globaltimer since kernel start) and converts, so that a power-capped
// clock does not silently lower the demand rate.
__global__ void __launch_bounds__(256)
triad_tma_ladder_ns(double *__restrict__ a, const double *__restrict__ b, const double *__restrict__ c, double s,
                    uint64_t n, int tile, int S, int pace_ns, int step1024, int stall_t, int bad_n, int win_n,
                    int *__restrict__ dbg) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double *sb = reinterpret_cast<double *>(smem_raw);
  double *sc = sb + size_t(S) * tile;
  uint64_t *full = reinterpret_cast<uint64_t *>(sc + size_t(S) * tile);
  const uint32_t tile_bytes = uint32_t(tile) * 8u;
  const uint64_t ntiles = n / tile;
  const uint64_t my_tiles = ntiles > blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const int tid = threadIdx.x;
  uint64_t policy = 0;
  if (tid == 0) {
    for (int i = 0; i < S; ++i) tma::mbar_init(&full[i], 1);
    tma::fence_barrier_init();
    policy = tma::policy_evict_first();
  }
  __syncthreads();
  uint64_t pj = 0;
  int pst = 0;
  auto issue = [&]() {
    const uint64_t t = blockIdx.x + pj * gridDim.x;
    tma::mbar_arrive_expect_tx(&full[pst], 2 * tile_bytes);
    tma::bulk_load(sb + size_t(pst) * tile, b + t * tile, tile_bytes, &full[pst], policy);
    tma::bulk_load(sc + size_t(pst) * tile, c + t * tile, tile_bytes, &full[pst], policy);
    ++pj;
    if (++pst == S) pst = 0;
  };
  long long next_ok = 0, c_start = 0;
  unsigned long long g_start = 0;
  int pace_cur_ns = pace_ns, bad = 0, win = 0, level = 0;
  long long pace = (long long)pace_ns * 1965 / 1000;  // first guess: 1.965 GHz
  bool draining = false;
  uint64_t grace_until = 0;
  if (tid == 0) { c_start = clock64(); asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g_start)); }
  int st = 0;
  uint32_t parity = 0;
  uint64_t t = blockIdx.x;
  for (uint64_t j = 0; j < my_tiles; ++j, t += gridDim.x) {
    long long t0 = 0;
    if (tid == 0) {
      if ((j & 63) == 63) {  // refresh the cycles-per-ns estimate
        unsigned long long g;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g));
        const long long dc = clock64() - c_start;
        const long long dg = (long long)(g - g_start);
        if (dg > 20000) pace = (long long)pace_cur_ns * dc / dg;
      }
      if (pj <= j) {
        draining = false;
        for (int i = 0; i < S && pj < my_tiles; ++i) issue();
        next_ok = clock64() + pace;
        grace_until = j + 2 * S + 8;
        bad = win = 0;
      }
      t0 = clock64();
    }
    tma::mbar_wait(&full[st], parity);
    if (tid == 0 && !draining && j >= grace_until) {
      const long long stall = clock64() - t0;
      if (stall >= stall_t) ++bad;
      if (++win == win_n) {
        if (bad >= bad_n) {
          draining = true;
          pace_cur_ns += (pace_ns * step1024) >> 10;
          pace = pace * pace_cur_ns / (pace_cur_ns - ((pace_ns * step1024) >> 10));
          ++level;
        }
        bad = win = 0;
      }
    }
    const f64x2 *pb = reinterpret_cast<const f64x2 *>(sb + size_t(st) * tile);
    const f64x2 *pc = reinterpret_cast<const f64x2 *>(sc + size_t(st) * tile);
    f64x2 *ga = reinterpret_cast<f64x2 *>(a + t * tile);
#pragma unroll 4
    for (int v = tid; v < tile / 2; v += 256) {
      const f64x2 x = pb[v], y = pc[v];
      st_stream(ga + v, f64x2{__fma_rn(s, y.x, x.x), __fma_rn(s, y.y, x.y)});
    }
    __syncthreads();
    if (tid == 0 && !draining && pj < my_tiles && pj > j + 1) {
      while (clock64() < next_ok) { }
      next_ok += pace;
      const long long now = clock64();
      if (next_ok < now) next_ok = now;
      issue();
    }
    if (++st == S) { st = 0; parity ^= 1u; }
  }
  if (tid == 0 && dbg) { dbg[2 * blockIdx.x] = int(pace); dbg[2 * blockIdx.x + 1] = level; }
}


// Write-only stream with paced tiles: does smoothing help pure writes too?
__global__ void __launch_bounds__(256)
fill_paced(f64x2 *__restrict__ dst, uint64_t nvec, int tile_vecs, int pace_ns) {
  const uint64_t ntiles = nvec / tile_vecs;
  __shared__ long long go;
  long long next_ok = 0, pace = (long long)pace_ns * 2, c0 = 0;
  unsigned long long g0 = 0;
  if (threadIdx.x == 0) { c0 = clock64(); g0 = global_ns(); next_ok = c0; }
  uint64_t j = 0;
  for (uint64_t t = blockIdx.x; t < ntiles; t += gridDim.x, ++j) {
    if (pace_ns > 0) {
      if (threadIdx.x == 0) {
        if ((j & 63) == 63) {
          const long long dg = (long long)(global_ns() - g0), dc = clock64() - c0;
          if (dg > 20000) pace = (long long)pace_ns * dc / dg;
        }
        while (clock64() < next_ok) { }
        next_ok += pace;
        const long long now = clock64();
        if (next_ok < now) next_ok = now;
        go = now;
      }
      __syncthreads();
    }
    f64x2 *p = dst + t * tile_vecs;
    for (int v = threadIdx.x; v < tile_vecs; v += 256) {
      const uint64_t k = 2 * (t * tile_vecs + v);
      st_stream(p + v, f64x2{unit_double(splitmix64_at(k, 1)), unit_double(splitmix64_at(k + 1, 1))});
    }
  }
}


// Read-only stream (the checksum's traffic) with paced refills: does smoothing help pure reads?
__global__ void __launch_bounds__(256)
csum_tma_paced(const uint64_t *__restrict__ w, uint64_t n, int tile, int S, int pace_ns, uint64_t *__restrict__ out) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t *sw = reinterpret_cast<uint64_t *>(smem_raw);
  uint64_t *full = sw + size_t(S) * tile;
  const uint32_t tile_bytes = uint32_t(tile) * 8u;
  const uint64_t ntiles = n / tile;
  const uint64_t my_tiles = ntiles > blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const int tid = threadIdx.x;
  uint64_t policy = 0;
  if (tid == 0) {
    for (int i = 0; i < S; ++i) tma::mbar_init(&full[i], 1);
    tma::fence_barrier_init();
    policy = tma::policy_evict_first();
  }
  __syncthreads();
  uint64_t pj = 0;
  int pst = 0;
  auto issue = [&]() {
    const uint64_t t = blockIdx.x + pj * gridDim.x;
    tma::mbar_arrive_expect_tx(&full[pst], tile_bytes);
    tma::bulk_load(sw + size_t(pst) * tile, w + t * tile, tile_bytes, &full[pst], policy);
    ++pj;
    if (++pst == S) pst = 0;
  };
  long long next_ok = 0, pace = (long long)pace_ns * 2, c0 = 0;
  unsigned long long g0 = 0;
  if (tid == 0) {
    c0 = clock64(); g0 = global_ns();
    for (int i = 0; i < S && pj < my_tiles; ++i) issue();
    next_ok = clock64() + pace;
  }
  uint64_t s0 = 0, s1 = 0;
  int st = 0;
  uint32_t parity = 0;
  uint64_t t = blockIdx.x;
  for (uint64_t j = 0; j < my_tiles; ++j, t += gridDim.x) {
    if (tid == 0 && (j & 63) == 63) {
      const long long dg = (long long)(global_ns() - g0), dc = clock64() - c0;
      if (dg > 20000) pace = (long long)pace_ns * dc / dg;
    }
    tma::mbar_wait(&full[st], parity);
    const u64x2 *pw = reinterpret_cast<const u64x2 *>(sw + size_t(st) * tile);
#pragma unroll 4
    for (int v = tid; v < tile / 2; v += 256) {
      const u64x2 x = pw[v];
      const uint64_t k = t * tile + 2 * uint64_t(v);
      s0 += x.x + x.y; s1 += x.x * (2 * k + 1) + x.y * (2 * k + 3);
    }
    __syncthreads();
    if (tid == 0 && pj < my_tiles) {
      if (pace_ns > 0) {
        while (clock64() < next_ok) { }
        next_ok += pace;
        const long long now = clock64();
        if (next_ok < now) next_ok = now;
      }
      issue();
    }
    if (++st == S) { st = 0; parity ^= 1u; }
  }
  if (s0 == 0x1234567 && s1 == 0x7654321) out[0] = s0;  // keep the sums alive
}


// K3 as a paced DMA pipeline (see copy_tma in probe_kernels.cu): one thread, bulk load -> bulk store.
__global__ void __launch_bounds__(32)
copy_tma_paced(unsigned char *__restrict__ dst, const unsigned char *__restrict__ src, uint64_t bytes, int tile_bytes, int S,
               int pace_ns) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + size_t(S) * tile_bytes);
  const uint64_t ntiles = bytes / tile_bytes;
  const uint64_t my_tiles = ntiles > blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  if (threadIdx.x != 0) return;
  for (int i = 0; i < S; ++i) tma::mbar_init(&full[i], 1);
  tma::fence_barrier_init();
  const uint64_t policy = tma::policy_evict_first();
  auto load = [&](uint64_t j) {
    const int st = int(j % S);
    const uint64_t t = blockIdx.x + j * gridDim.x;
    tma::mbar_arrive_expect_tx(&full[st], tile_bytes);
    tma::bulk_load(smem_raw + size_t(st) * tile_bytes, src + t * tile_bytes, tile_bytes, &full[st], policy);
  };
  long long pace = (long long)pace_ns * 2, next_ok = 0;
  const long long c0 = clock64();
  const unsigned long long g0 = global_ns();
  for (uint64_t j = 0; j < uint64_t(S - 1) && j < my_tiles; ++j) load(j);
  next_ok = clock64() + pace;
  for (uint64_t j = 0; j < my_tiles; ++j) {
    const int st = int(j % S);
    const uint64_t t = blockIdx.x + j * gridDim.x;
    if ((j & 63) == 63) {
      const long long dg = (long long)(global_ns() - g0), dc = clock64() - c0;
      if (dg > 20000) pace = (long long)pace_ns * dc / dg;
    }
    tma::mbar_wait(&full[st], uint32_t(j / S) & 1u);
    tma::bulk_store(dst + t * tile_bytes, smem_raw + size_t(st) * tile_bytes, tile_bytes);
    tma::bulk_commit();
    tma::bulk_wait_read<1>();
    if (j + S - 1 < my_tiles) {
      if (pace_ns > 0) {
        while (clock64() < next_ok) { }
        next_ok += pace;
        const long long now = clock64();
        if (next_ok < now) next_ok = now;
      }
      load(j + S - 1);
    }
  }
  tma::bulk_wait_all<0>();
}

int main(int argc, char **argv) {
  const int log2n = argc > 1 ? atoi(argv[1]) : 29;
  const char *only = argc > 2 ? argv[2] : "all";
  const uint64_t n = 1ull << log2n;
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  // one slab; arrays placed with a configurable pad between them
  const size_t slab_bytes = 3 * n * 8 + (64u << 20);
  unsigned char *slab = nullptr;
  CK(cudaMalloc(&slab, slab_bytes));
  CK(cudaMemset(slab, 0, slab_bytes));
  uint64_t *dout;
  CK(cudaMalloc(&dout, 8));
  Timer tm;
  printf("kind,param,best_ms,med_ms,GBps_med,frac8000\n");
  auto report = [&](const char *kind, const char *param, double bytes, float best, float med) {
    printf("%s,%s,%.4f,%.4f,%.1f,%.4f\n", kind, param, best, med, bytes / med / 1e6, bytes / med / 1e6 / 8000.0);
    fflush(stdout);
  };
  const size_t pads[] = {0, 256, 4096 + 256, (1u << 20) + 4096 + 256, (8u << 20) + 33 * 1024};
  auto place = [&](size_t pad, double **a, double **b, double **c) {
    *b = reinterpret_cast<double *>(slab);
    *c = reinterpret_cast<double *>(slab + n * 8 + pad);
    *a = reinterpret_cast<double *>(slab + 2 * (n * 8 + pad));
  };
  char param[256];
  float best, med;
  double *a, *b, *c;

  {  // references
    place(0, &a, &b, &c);
    tm.run([&] { read_only<<<sms * 6, 256>>>(reinterpret_cast<const u64x4 *>(b), n / 4, dout); }, 2, 10, &best, &med);
    report("read_only", "ldg256 U4 6cta", 8.0 * n, best, med);
    tm.run([&] { write_only<<<sms * 6, 256>>>(reinterpret_cast<f64x4 *>(a), n / 4); }, 2, 10, &best, &med);
    report("write_only", "stg256 U4 6cta", 8.0 * n, best, med);
    tm.run([&] { CK(cudaMemcpyAsync(a, b, n * 8, cudaMemcpyDeviceToDevice)); }, 2, 10, &best, &med);
    report("memcpy_d2d", "copy engine/kernel", 16.0 * n, best, med);
  }

  auto run_tma = [&](int store, int tile, int S, int SO, int ctas, int threads, int blocked, int load_ef, size_t pad) {
    const bool bulk = store <= BULK_EL;
    const size_t smem = size_t(tile) * 8 * (2 * S + (bulk ? SO : 0)) + 8 * S;
    if (smem > 227 * 1024 / size_t(ctas) - 1024) return;
    place(pad, &a, &b, &c);
    auto launch = [&](auto kern) {
      CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
      int occ = 0;
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, smem));
      if (occ < ctas) return false;
      tm.run([&] { kern<<<sms * ctas, threads, smem>>>(a, b, c, 3.0, n, tile, S, SO, blocked, load_ef); }, 2, 8, &best, &med);
      CK(cudaGetLastError());
      return true;
    };
    bool ok = false;
    switch (store) {
      case BULK_NONE: ok = launch(triad_tma_rt<BULK_NONE>); break;
      case BULK_EF: ok = launch(triad_tma_rt<BULK_EF>); break;
      case BULK_EL: ok = launch(triad_tma_rt<BULK_EL>); break;
      case STG_CS: ok = launch(triad_tma_rt<STG_CS>); break;
      case STG_DEF: ok = launch(triad_tma_rt<STG_DEF>); break;
      case STG_NA: ok = launch(triad_tma_rt<STG_NA>); break;
    }
    if (!ok) return;
    snprintf(param, sizeof param, "store=%d tileKB=%d S=%d SO=%d ctas=%d thr=%d blocked=%d loadEF=%d pad=%zu", store,
             tile * 8 / 1024, S, SO, ctas, threads, blocked, load_ef, pad);
    report("triad_tma", param, 24.0 * n, best, med);
  };

  const bool all = !strcmp(only, "all");
  if (all || !strcmp(only, "pad")) {
    for (size_t pad : pads) run_tma(BULK_NONE, 1024, 4, 2, 2, 256, 0, 1, pad);
    for (size_t pad : pads) run_tma(STG_CS, 1024, 4, 2, 2, 256, 0, 1, pad);
  }
  if (all || !strcmp(only, "store")) {
    for (int store = 0; store <= 5; ++store)
      for (int lef = 0; lef < 2; ++lef) run_tma(store, 1024, 4, 2, 2, 256, 0, lef, 0);
  }
  if (all || !strcmp(only, "geom")) {
    for (int store : {int(BULK_NONE), int(STG_CS)})
      for (int tile : {512, 1024, 2048, 4096})
        for (int S : {2, 3, 4, 6, 8})
          for (int ctas : {1, 2, 3, 4})
            run_tma(store, tile, S, 2, ctas, 256, 0, 1, 0);
    for (int threads : {128, 512})
      for (int tile : {1024, 2048})
        for (int ctas : {1, 2, 4}) run_tma(BULK_NONE, tile, 4, 2, ctas, threads, 0, 1, 0);
    for (int SO : {3, 4}) run_tma(BULK_NONE, 1024, 4, SO, 2, 256, 0, 1, 0);
    for (int tile : {1024, 2048, 4096}) run_tma(BULK_NONE, tile, 4, 2, 2, 256, 1, 1, 0);  // blocked assignment
  }
  if (all || !strcmp(only, "ldg")) {
    place(0, &a, &b, &c);
#define RUN_LDG(V, VB, U, LD, THREADS, CTAS)                                                                      \
  do {                                                                                                            \
    auto kern = triad_ldg_rt<V, U, LD>;                                                                           \
    int occ = 0;                                                                                                  \
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, THREADS, 0));                                    \
    if (occ >= CTAS) {                                                                                            \
      const uint64_t nvec = n * 8 / VB;                                                                           \
      tm.run([&] { kern<<<sms * CTAS, THREADS>>>(reinterpret_cast<V *>(a), reinterpret_cast<const V *>(b),        \
                                                  reinterpret_cast<const V *>(c), 3.0, nvec); }, 2, 8, &best, &med); \
      CK(cudaGetLastError());                                                                                     \
      snprintf(param, sizeof param, "vec=%d U=%d ld=%d thr=%d ctas=%d", VB, U, LD, THREADS, CTAS);                \
      report("triad_ldg", param, 24.0 * n, best, med);                                                            \
    }                                                                                                             \
  } while (0)
    for (int ctas : {1, 2, 3, 4, 6, 8}) {
      switch (ctas) {
#define CASES(C)                                                                                   \
  case C:                                                                                          \
    RUN_LDG(f64x4, 32, 1, 0, 256, C); RUN_LDG(f64x4, 32, 2, 0, 256, C); RUN_LDG(f64x4, 32, 4, 0, 256, C); \
    RUN_LDG(f64x2, 16, 2, 0, 256, C); RUN_LDG(f64x2, 16, 4, 0, 256, C); RUN_LDG(f64x2, 16, 8, 0, 256, C); \
    RUN_LDG(f64x4, 32, 2, 1, 256, C); RUN_LDG(f64x2, 16, 4, 1, 256, C);                               \
    RUN_LDG(f64x4, 32, 2, 0, 512, C); RUN_LDG(f64x4, 32, 2, 0, 128, C); RUN_LDG(f64x4, 32, 4, 0, 128, C); \
    break;
        CASES(1) CASES(2) CASES(3) CASES(4) CASES(6) CASES(8)
      }
    }
  }

  if (all || !strcmp(only, "csum")) {
    place(0, &a, &b, &c);
    uint64_t *ws;
    CK(cudaMalloc(&ws, kWorkspaceBytes));
    CK(cudaMemset(ws, 0, kWorkspaceBytes));
    const uint64_t *w = reinterpret_cast<const uint64_t *>(b);
    for (int tile : {512, 1024, 2048, 4096})
      for (int S : {2, 3, 4, 6, 8, 12})
        for (int ctas : {1, 2, 3, 4})
          for (int threads : {128, 256, 512}) {
            const size_t smem = size_t(tile) * 8 * S + 8 * S;
            if (smem > 227 * 1024 / size_t(ctas) - 1024) continue;
            CK(cudaFuncSetAttribute(checksum_tma_rt, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
            int occ = 0;
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, checksum_tma_rt, threads, smem));
            if (occ < ctas) continue;
            tm.run([&] { checksum_tma_rt<<<sms * ctas, threads, smem>>>(w, n, dout, ws, tile, S); }, 2, 8, &best, &med);
            CK(cudaGetLastError());
            snprintf(param, sizeof param, "tileKB=%d S=%d ctas=%d thr=%d", tile * 8 / 1024, S, ctas, threads);
            report("checksum_tma", param, 8.0 * n, best, med);
          }
#define RUN_CL(V, VB, U, THREADS, CTAS)                                                                    \
  do {                                                                                                     \
    auto kern = checksum_ldg_rt<V, U>;                                                                     \
    int occ = 0;                                                                                           \
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, THREADS, 0));                             \
    if (occ >= CTAS) {                                                                                     \
      const uint64_t nvec = n * 8 / VB;                                                                    \
      tm.run([&] { kern<<<sms * CTAS, THREADS>>>(reinterpret_cast<const V *>(w), nvec, dout, ws); }, 2, 8, &best, &med); \
      CK(cudaGetLastError());                                                                              \
      snprintf(param, sizeof param, "vec=%d U=%d thr=%d ctas=%d", VB, U, THREADS, CTAS);                   \
      report("checksum_ldg", param, 8.0 * n, best, med);                                                   \
    }                                                                                                      \
  } while (0)
#define CL_CASES(C)                                                                               \
    RUN_CL(u64x4, 32, 1, 256, C); RUN_CL(u64x4, 32, 2, 256, C); RUN_CL(u64x4, 32, 4, 256, C); RUN_CL(u64x4, 32, 8, 256, C); \
    RUN_CL(u64x2, 16, 4, 256, C); RUN_CL(u64x2, 16, 8, 256, C); RUN_CL(u64x2, 16, 16, 256, C);        \
    RUN_CL(u64x4, 32, 2, 128, C); RUN_CL(u64x4, 32, 4, 128, C); RUN_CL(u64x4, 32, 8, 128, C);        \
    RUN_CL(u64x4, 32, 2, 512, C); RUN_CL(u64x4, 32, 4, 512, C); RUN_CL(u64x4, 32, 4, 1024, C);
    CL_CASES(1) CL_CASES(2) CL_CASES(3) CL_CASES(4) CL_CASES(5) CL_CASES(6) CL_CASES(8)
    CK(cudaFree(ws));
  }
  if (all || !strcmp(only, "fine")) {  // around the triad optimum: TMA loads + STG stores, few CTAs, deep ring
    for (int store : {int(STG_CS), int(STG_DEF), int(STG_NA)})
      for (int tile : {512, 1024, 2048})
        for (int S : {4, 5, 6, 7, 8, 10, 12, 16})
          for (int ctas : {1, 2})
            for (int threads : {128, 256, 384, 512}) run_tma(store, tile, S, 2, ctas, threads, 0, 1, 0);
  }

  if (!strcmp(only, "b2b")) {  // does back-to-back launching or separate allocations change the picture?
    double *sa, *sb2, *sc2;
    CK(cudaMalloc(&sb2, n * 8)); CK(cudaMalloc(&sc2, n * 8)); CK(cudaMalloc(&sa, n * 8));
    CK(cudaMemset(sb2, 0, n * 8)); CK(cudaMemset(sc2, 0, n * 8));
    const int tile = 1024, S = 8;
    const size_t smem = size_t(tile) * 8 * 2 * S + 8 * S;
    auto kern = triad_tma_rt<STG_CS>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    for (int sep = 0; sep < 2; ++sep) {
      if (sep) { a = sa; b = sb2; c = sc2; } else place(0, &a, &b, &c);
      for (int batch : {1, 10, 50}) {
        auto f = [&] { for (int i = 0; i < batch; ++i) kern<<<sms, 256, smem>>>(a, b, c, 3.0, n, tile, S, 2, 0, 1); };
        tm.run(f, 2, 8, &best, &med);
        snprintf(param, sizeof param, "tma8K S8 sep_alloc=%d batch=%d (ms per launch)", sep, batch);
        report("triad_tma_b2b", param, 24.0 * n * batch, best, med);
      }
      for (int batch : {1, 10, 50}) {
        auto k2 = triad_ldg_rt<f64x2, 8, 0>;
        auto f = [&] { for (int i = 0; i < batch; ++i) k2<<<sms * 2, 256>>>(reinterpret_cast<f64x2 *>(a), reinterpret_cast<const f64x2 *>(b), reinterpret_cast<const f64x2 *>(c), 3.0, n / 2); };
        tm.run(f, 2, 8, &best, &med);
        snprintf(param, sizeof param, "ldg128 U8 2cta sep_alloc=%d batch=%d", sep, batch);
        report("triad_ldg_b2b", param, 24.0 * n * batch, best, med);
      }
    }
  }

  if (!strcmp(only, "ship")) {  // shipped library vs the tuner's copy, same buffers, same process
    double *sa, *sb2, *sc2;
    CK(cudaMalloc(&sb2, n * 8)); CK(cudaMalloc(&sc2, n * 8)); CK(cudaMalloc(&sa, n * 8));
    CK(cudaMemset(sb2, 0, n * 8)); CK(cudaMemset(sc2, 0, n * 8));
    const int tile = 1024, S = 8;
    const size_t smem = size_t(tile) * 8 * 2 * S + 8 * S;
    auto kern = triad_tma_rt<STG_CS>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    for (int rep = 0; rep < 2; ++rep)
      for (int batch : {1, 10}) {
        a = sa; b = sb2; c = sc2;
        tm.run([&] { for (int i = 0; i < batch; ++i) kern<<<sms, 256, smem>>>(a, b, c, 3.0, n, tile, S, 2, 0, 1); }, 2, 10, &best, &med);
        snprintf(param, sizeof param, "tuner copy batch=%d", batch);
        report("triad_tma", param, 24.0 * n * batch, best, med);
        for (int v = 0; v < 3; ++v) {
          tm.run([&] { for (int i = 0; i < batch; ++i) { int rc = gpu_probe_triad(a, b, c, 3.0, n, v, nullptr); if (rc) { fprintf(stderr, "rc=%d\n", rc); exit(1); } } }, 2, 10, &best, &med);
          snprintf(param, sizeof param, "shipped variant=%d batch=%d", v, batch);
          report("gpu_probe_triad", param, 24.0 * n * batch, best, med);
        }
      }
  }

  if (!strcmp(only, "shape")) {
    double *sa, *sb2, *sc2;
    CK(cudaMalloc(&sb2, n * 8)); CK(cudaMalloc(&sc2, n * 8)); CK(cudaMalloc(&sa, n * 8));
    CK(cudaMemset(sb2, 0, n * 8)); CK(cudaMemset(sc2, 0, n * 8));
    a = sa; b = sb2; c = sc2;
    const size_t smem = size_t(1024) * 8 * 2 * 8 + 8 * 8;
    auto k0 = triad_tma_ct<1024, 8, 0>; auto k1 = triad_tma_ct<1024, 8, 1>; auto kr = triad_tma_rt<STG_CS>;
    CK(cudaFuncSetAttribute(k0, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    CK(cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    CK(cudaFuncSetAttribute(kr, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    for (int rep = 0; rep < 3; ++rep) {
      tm.run([&] { kr<<<sms, 256, smem>>>(a, b, c, 3.0, n, 1024, 8, 2, 0, 1); }, 2, 10, &best, &med);
      report("triad_tma", "runtime-geometry (tuner)", 24.0 * n, best, med);
      tm.run([&] { k0<<<sms, 256, smem>>>(a, b, c, 3.0, n); }, 2, 10, &best, &med);
      report("triad_tma", "compile-time unrolled", 24.0 * n, best, med);
      tm.run([&] { k1<<<sms, 256, smem>>>(a, b, c, 3.0, n); }, 2, 10, &best, &med);
      report("triad_tma", "compile-time unroll 1", 24.0 * n, best, med);
      tm.run([&] { int rc = gpu_probe_triad(a, b, c, 3.0, n, 1, nullptr); if (rc) exit(1); }, 2, 10, &best, &med);
      report("triad_tma", "shipped", 24.0 * n, best, med);
    }
  }

  if (!strcmp(only, "fast")) {
    double *sa, *sb2, *sc2;
    CK(cudaMalloc(&sb2, n * 8)); CK(cudaMalloc(&sc2, n * 8)); CK(cudaMalloc(&sa, n * 8));
    CK(cudaMemset(sb2, 0, n * 8)); CK(cudaMemset(sc2, 0, n * 8));
    a = sa; b = sb2; c = sc2;
    auto kf = triad_tma_fast<4>;
    auto run_fast = [&](int tile, int S, int ctas, int pace) {
      const size_t smem = size_t(tile) * 8 * 2 * S + 8 * S;
      if (smem > 227 * 1024 / size_t(ctas) - 1024) return;
      CK(cudaFuncSetAttribute(kf, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
      int occ = 0;
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kf, 256, smem));
      if (occ < ctas) return;
      tm.run([&] { kf<<<sms * ctas, 256, smem>>>(a, b, c, 3.0, n, tile, S, pace); }, 2, 10, &best, &med);
      CK(cudaGetLastError());
      snprintf(param, sizeof param, "tileKB=%.1f S=%d ctas=%d pace=%d inflightKB=%d", tile * 8 / 1024.0, S, ctas, pace,
               tile * 16 * S * ctas / 1024);
      report("triad_fast", param, 24.0 * n, best, med);
    };
    for (int tile : {256, 512, 1024, 2048, 4096})
      for (int S : {2, 3, 4, 5, 6, 8, 12, 16})
        for (int ctas : {1, 2, 3, 4}) run_fast(tile, S, ctas, 0);
    // pacing: 8 KiB tiles, 1 CTA/SM, deep ring; pace in SM cycles per tile (24 KiB of traffic per tile per SM)
    for (int S : {4, 8})
      for (int pace : {700, 800, 850, 900, 925, 950, 975, 1000, 1025, 1050, 1100, 1200}) run_fast(1024, S, 1, pace);
    for (int pace : {350, 400, 425, 450, 475, 500, 525, 550, 600}) run_fast(512, 8, 1, pace);
    for (int pace : {1400, 1600, 1700, 1800, 1900, 2000, 2100, 2200, 2400}) run_fast(1024, 4, 2, pace);
  }

  if (!strcmp(only, "adapt")) {
    double *sa, *sb2, *sc2;
    int *pace_d;
    CK(cudaMalloc(&sb2, n * 8)); CK(cudaMalloc(&sc2, n * 8)); CK(cudaMalloc(&sa, n * 8));
    CK(cudaMalloc(&pace_d, sizeof(int) * (2 * 4096 + 16)));
    CK(cudaMemset(sb2, 0, n * 8)); CK(cudaMemset(sc2, 0, n * 8));
    a = sa; b = sb2; c = sc2;
    auto ka = triad_tma_adapt;
    auto run_adapt = [&](int tile, int S, int ctas, int pace0, int inc, int dec, int thresh, int calm_n) {
      const size_t smem = size_t(tile) * 8 * 2 * S + 8 * S;
      CK(cudaFuncSetAttribute(ka, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
      int occ = 0;
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ka, 256, smem));
      if (occ < ctas) return;
      const int grid = sms * ctas;
      CK(cudaMemset(pace_d + 8192, 0, sizeof(int) * 16));
      tm.run([&] { ka<<<grid, 256, smem>>>(a, b, c, 3.0, n, tile, S, pace0, inc, dec, thresh, calm_n, pace0 / 2, pace_d); }, 2, 10, &best, &med);
      CK(cudaGetLastError());
      std::vector<int> ph(2 * grid);
      CK(cudaMemcpy(ph.data(), pace_d, sizeof(int) * 2 * grid, cudaMemcpyDeviceToHost));
      int hist[8];
      CK(cudaMemcpy(hist, pace_d + 8192, sizeof(hist), cudaMemcpyDeviceToHost));
      double ht = 0;
      for (int i = 0; i < 7; ++i) ht += hist[i];
      fprintf(stderr, "stall_hist pace0=%d thresh=%d: <100 %.3f <200 %.3f <400 %.3f <800 %.3f <1600 %.3f <3200 %.3f >=3200 %.3f\n", pace0, thresh,
              hist[0] / ht, hist[1] / ht, hist[2] / ht, hist[3] / ht, hist[4] / ht, hist[5] / ht, hist[6] / ht);
      double fin = 0, avg = 0;
      for (int i = 0; i < grid; ++i) { fin += ph[2 * i]; avg += ph[2 * i + 1]; }
      snprintf(param, sizeof param, "tileKB=%.1f S=%d ctas=%d pace0=%d inc=%d dec=%d thresh=%d calm=%d final_pace=%.0f mean_pace=%.0f",
               tile * 8 / 1024.0, S, ctas, pace0, inc, dec, thresh, calm_n, fin / grid, avg / grid);
      report("triad_adapt", param, 24.0 * n, best, med);
    };
    // reference points: shipped kernel, fixed pacing at a few rates
    for (int rep = 0; rep < 2; ++rep) {
      tm.run([&] { int rc = gpu_probe_triad(a, b, c, 3.0, n, 1, nullptr); if (rc) exit(1); }, 2, 10, &best, &med);
      report("triad_tma", "shipped", 24.0 * n, best, med);
    }
    for (int pace : {1900, 1950, 2000, 2050, 2100, 2200}) run_adapt(1024, 4, 2, pace, 0, 0, 1 << 30, 1 << 30);
    for (int thresh : {50, 200, 600})
      for (int inc : {4, 16})
        for (int dec : {1, 4})
          for (int calm : {4, 16})
            run_adapt(1024, 4, 2, 2300, inc, dec, thresh, calm);
    for (int thresh : {50, 200, 600})
      for (int inc : {4, 16}) run_adapt(1024, 8, 1, 1150, inc, 1, thresh, 8);
  }

  if (!strcmp(only, "selfclock")) {
    double *sa, *sb2, *sc2;
    int *dbg_d;
    CK(cudaMalloc(&sb2, n * 8)); CK(cudaMalloc(&sc2, n * 8)); CK(cudaMalloc(&sa, n * 8));
    CK(cudaMalloc(&dbg_d, sizeof(int) * 2 * 4096));
    CK(cudaMemset(sb2, 0, n * 8)); CK(cudaMemset(sc2, 0, n * 8));
    a = sa; b = sb2; c = sc2;
    auto ks = triad_tma_selfclock;
    for (int rep = 0; rep < 2; ++rep) {
      tm.run([&] { int rc = gpu_probe_triad(a, b, c, 3.0, n, 1, nullptr); if (rc) exit(1); }, 2, 10, &best, &med);
      report("triad_tma", "shipped", 24.0 * n, best, med);
    }
    auto run_sc = [&](int tile, int S, int ctas, int target, int up, int down, int sh) {
      const size_t smem = size_t(tile) * 8 * 2 * S + 8 * S;
      if (smem > size_t(227 * 1024) / ctas - 1024) return;
      CK(cudaFuncSetAttribute(ks, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
      int occ = 0;
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ks, 256, smem));
      if (occ < ctas) return;
      const int grid = sms * ctas;
      tm.run([&] { ks<<<grid, 256, smem>>>(a, b, c, 3.0, n, tile, S, target, up, down, sh, dbg_d); }, 2, 10, &best, &med);
      CK(cudaGetLastError());
      std::vector<int> ph(2 * grid);
      CK(cudaMemcpy(ph.data(), dbg_d, sizeof(int) * 2 * grid, cudaMemcpyDeviceToHost));
      double T = 0, o = 0;
      for (int i = 0; i < grid; ++i) { T += ph[2 * i]; o += ph[2 * i + 1]; }
      snprintf(param, sizeof param, "tileKB=%.1f S=%d ctas=%d target=%d up=%d down=%d ewma=%d T=%.0f outstanding=%.2f",
               tile * 8 / 1024.0, S, ctas, target, up, down, sh, T / grid, o / grid / 100);
      report("triad_selfclock", param, 24.0 * n, best, med);
    };
    for (int tile : {512, 1024})
      for (int ctas : {1, 2})
        for (int S : {6, 8})
          for (int target : {2, 3, 4, 5})
            for (int up : {1034, 1075})
              for (int down : {1014, 973}) {
                if (target >= S - 1) continue;
                run_sc(tile, S, ctas, target, up, down, 4);
              }
  }

  if (!strcmp(only, "hist")) {
    double *sa, *sb2, *sc2;
    int *pace_d;
    CK(cudaMalloc(&sb2, n * 8)); CK(cudaMalloc(&sc2, n * 8)); CK(cudaMalloc(&sa, n * 8));
    CK(cudaMalloc(&pace_d, sizeof(int) * (2 * 4096 + 16)));
    CK(cudaMemset(sb2, 0, n * 8)); CK(cudaMemset(sc2, 0, n * 8));
    a = sa; b = sb2; c = sc2;
    auto ka = triad_tma_adapt;
    for (int rep = 0; rep < 2; ++rep) {
      tm.run([&] { int rc = gpu_probe_triad(a, b, c, 3.0, n, 1, nullptr); if (rc) exit(1); }, 2, 10, &best, &med);
      report("triad_tma", "shipped", 24.0 * n, best, med);
    }
    for (int tile : {1024, 512}) {
      const int S = 4, ctas = 2;
      const size_t smem = size_t(tile) * 8 * 2 * S + 8 * S;
      CK(cudaFuncSetAttribute(ka, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
      const int grid = sms * ctas;
      for (int p : {1900, 1950, 1975, 1990, 2000, 2010, 2025, 2050, 2100}) {
        const int pace = p * tile / 1024;
        CK(cudaMemset(pace_d + 8192, 0, sizeof(int) * 16));
        tm.run([&] { ka<<<grid, 256, smem>>>(a, b, c, 3.0, n, tile, S, pace, 0, 0, 1 << 30, 1 << 30, 1, pace_d); }, 2, 10, &best, &med);
        CK(cudaGetLastError());
        int hist[8];
        CK(cudaMemcpy(hist, pace_d + 8192, sizeof(hist), cudaMemcpyDeviceToHost));
        double ht = 0;
        for (int i = 0; i < 7; ++i) ht += hist[i];
        snprintf(param, sizeof param, "tileKB=%.1f pace=%d hist<100:%.3f <200:%.3f <400:%.3f <800:%.3f <1600:%.3f <3200:%.3f >=3200:%.3f",
                 tile * 8 / 1024.0, pace, hist[0] / ht, hist[1] / ht, hist[2] / ht, hist[3] / ht, hist[4] / ht, hist[5] / ht, hist[6] / ht);
        report("triad_paced", param, 24.0 * n, best, med);
      }
    }
  }

  if (!strcmp(only, "ladder")) {
    double *sa, *sb2, *sc2;
    int *dbg_d;
    CK(cudaMalloc(&sb2, n * 8)); CK(cudaMalloc(&sc2, n * 8)); CK(cudaMalloc(&sa, n * 8));
    CK(cudaMalloc(&dbg_d, sizeof(int) * 2 * 4096));
    CK(cudaMemset(sb2, 0, n * 8)); CK(cudaMemset(sc2, 0, n * 8));
    a = sa; b = sb2; c = sc2;
    auto kl = triad_tma_ladder;
    for (int rep = 0; rep < 2; ++rep) {
      tm.run([&] { int rc = gpu_probe_triad(a, b, c, 3.0, n, 1, nullptr); if (rc) exit(1); }, 2, 10, &best, &med);
      report("triad_tma", "shipped", 24.0 * n, best, med);
    }
    auto run_l = [&](int tile, int S, int ctas, int pace0, int step, int stall_t, int bad_n, int win_n) {
      const size_t smem = size_t(tile) * 8 * 2 * S + 8 * S;
      CK(cudaFuncSetAttribute(kl, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
      const int grid = sms * ctas;
      tm.run([&] { kl<<<grid, 256, smem>>>(a, b, c, 3.0, n, tile, S, pace0, step, stall_t, bad_n, win_n, dbg_d); }, 2, 10, &best, &med);
      CK(cudaGetLastError());
      std::vector<int> ph(2 * grid);
      CK(cudaMemcpy(ph.data(), dbg_d, sizeof(int) * 2 * grid, cudaMemcpyDeviceToHost));
      double pace = 0; int lmax = 0, l0 = 0;
      for (int i = 0; i < grid; ++i) { pace += ph[2 * i]; lmax = std::max(lmax, ph[2 * i + 1]); l0 += ph[2 * i + 1] == 0; }
      snprintf(param, sizeof param, "tileKB=%.1f S=%d ctas=%d pace0=%d step=%d stall_t=%d bad=%d/%d mean_final_pace=%.0f max_level=%d ctas_at_level0=%d",
               tile * 8 / 1024.0, S, ctas, pace0, step, stall_t, bad_n, win_n, pace / grid, lmax, l0);
      report("triad_ladder", param, 24.0 * n, best, med);
    };
    for (int pace0 : {1800, 1900, 1950, 1975, 1990, 2000, 2025})
      for (int step : {20, 10}) run_l(1024, 4, 2, pace0, step, 800, 5, 64);
    for (int pace0 : {1900, 1990}) {
      run_l(1024, 4, 2, pace0, 20, 800, 3, 32);
      run_l(1024, 4, 2, pace0, 20, 400, 8, 64);
      run_l(1024, 4, 2, pace0, 30, 800, 5, 64);
    }
  }

  if (!strcmp(only, "ladderns")) {
    unsigned long long *gt;
    CK(cudaMalloc(&gt, 8 * 80));
    globaltimer_probe<<<1, 1>>>(gt);
    unsigned long long gh[65];
    CK(cudaMemcpy(gh, gt, sizeof(gh), cudaMemcpyDeviceToHost));
    fprintf(stderr, "globaltimer increments (ns):");
    for (int i = 0; i < 12; ++i) fprintf(stderr, " %llu", gh[i]);
    unsigned long long tot = 0; for (int i = 0; i < 64; ++i) tot += gh[i];
    fprintf(stderr, " ... 64 ticks = %llu ns = %llu cycles\n", tot, gh[64]);
    double *sa, *sb2, *sc2;
    int *dbg_d;
    CK(cudaMalloc(&sb2, n * 8)); CK(cudaMalloc(&sc2, n * 8)); CK(cudaMalloc(&sa, n * 8));
    CK(cudaMalloc(&dbg_d, sizeof(int) * 2 * 4096));
    CK(cudaMemset(sb2, 0, n * 8)); CK(cudaMemset(sc2, 0, n * 8));
    a = sa; b = sb2; c = sc2;
    auto kl = triad_tma_ladder_ns;
    for (int rep = 0; rep < 2; ++rep) {
      tm.run([&] { int rc = gpu_probe_triad(a, b, c, 3.0, n, 1, nullptr); if (rc) exit(1); }, 2, 10, &best, &med);
      report("triad_tma", "shipped", 24.0 * n, best, med);
    }
    const int tile = 1024, S = 4, ctas = 2;
    const size_t smem = size_t(tile) * 8 * 2 * S + 8 * S;
    CK(cudaFuncSetAttribute(kl, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    const int grid = sms * ctas;
    for (int rep = 0; rep < 2; ++rep)
      for (int pace_ns : {1000, 1008, 1013, 1018, 1023, 1030, 1040, 1055, 1070}) {
        tm.run([&] { kl<<<grid, 256, smem>>>(a, b, c, 3.0, n, tile, S, pace_ns, 20, 800, 5, 64, dbg_d); }, 2, 20, &best, &med);
        CK(cudaGetLastError());
        std::vector<int> ph(2 * grid);
        CK(cudaMemcpy(ph.data(), dbg_d, sizeof(int) * 2 * grid, cudaMemcpyDeviceToHost));
        double pace = 0; int lmax = 0, l0 = 0;
        for (int i = 0; i < grid; ++i) { pace += ph[2 * i]; lmax = std::max(lmax, ph[2 * i + 1]); l0 += ph[2 * i + 1] == 0; }
        snprintf(param, sizeof param, "pace_ns=%d mean_final_pace_cycles=%.0f max_level=%d ctas_at_level0=%d", pace_ns, pace / grid, lmax, l0);
        report("triad_ladder_ns", param, 24.0 * n, best, med);
      }
  }

  if (!strcmp(only, "fillpace")) {
    double *sa;
    CK(cudaMalloc(&sa, n * 8));
    const uint64_t nvec = n / 2;
    for (int rep = 0; rep < 2; ++rep) {
      tm.run([&] { int rc = gpu_probe_init_fill(sa, n, 0, 1, nullptr); if (rc) exit(1); }, 2, 10, &best, &med);
      report("init_fill", "shipped", 8.0 * n, best, med);
    }
    for (int ctas : {2, 4, 8})
      for (int tile_kb : {4, 8, 16}) {
        const int tile_vecs = tile_kb * 1024 / 16;
        for (int gbs : {0, 6600, 6800, 7000, 7100, 7200, 7300, 7400}) {
          // pace = time for one tile per CTA at the target aggregate rate
          const int pace_ns = gbs ? int(double(tile_kb) * 1024 * sms * ctas / (gbs * 1e9) * 1e9) : 0;
          tm.run([&] { fill_paced<<<sms * ctas, 256>>>(reinterpret_cast<f64x2 *>(sa), nvec, tile_vecs, pace_ns); }, 2, 10, &best, &med);
          CK(cudaGetLastError());
          snprintf(param, sizeof param, "ctas=%d tileKB=%d target_GBps=%d pace_ns=%d", ctas, tile_kb, gbs, pace_ns);
          report("fill_paced", param, 8.0 * n, best, med);
        }
      }
  }

  if (!strcmp(only, "csumpace")) {
    double *sa; uint64_t *o;
    CK(cudaMalloc(&sa, n * 8)); CK(cudaMalloc(&o, 64));
    CK(cudaMemset(sa, 1, n * 8));
    auto kc = csum_tma_paced;
    for (int tile_kb : {8, 16, 32})
      for (int S : {4, 6})
        for (int ctas : {1, 2}) {
          const int tile = tile_kb * 128;
          const size_t smem = size_t(tile) * 8 * S + 8 * S;
          if (smem * ctas > 220 * 1024) continue;
          CK(cudaFuncSetAttribute(kc, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
          for (int gbs : {0, 7200, 7350, 7450, 7550, 7650, 7800}) {
            const int pace_ns = gbs ? int(double(tile_kb) * 1024 * sms * ctas / (gbs * 1e9) * 1e9) : 0;
            tm.run([&] { kc<<<sms * ctas, 256, smem>>>(reinterpret_cast<const uint64_t *>(sa), n, tile, S, pace_ns, o); }, 2, 10, &best, &med);
            CK(cudaGetLastError());
            snprintf(param, sizeof param, "tileKB=%d S=%d ctas=%d target_GBps=%d pace_ns=%d", tile_kb, S, ctas, gbs, pace_ns);
            report("csum_paced", param, 8.0 * n, best, med);
          }
        }
  }

  if (!strcmp(only, "copypace")) {
    unsigned char *src, *dst;
    const uint64_t bytes = n * 8;
    CK(cudaMalloc(&src, bytes)); CK(cudaMalloc(&dst, bytes));
    CK(cudaMemset(src, 3, bytes));
    auto kc = copy_tma_paced;
    for (int rep = 0; rep < 2; ++rep) {
      tm.run([&] { int rc = gpu_probe_p2p_copy(dst, src, bytes, nullptr); if (rc) exit(1); }, 2, 10, &best, &med);
      report("copy", "shipped", 2.0 * bytes, best, med);
    }
    for (int tile_kb : {16, 32})
      for (int S : {3, 4, 6}) {
        const size_t smem = size_t(tile_kb) * 1024 * S + 8 * S;
        if (smem > 200 * 1024) continue;
        CK(cudaFuncSetAttribute(kc, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        for (int gbs : {0, 6400, 6600, 6800, 6900, 7000, 7100, 7200}) {
          const int pace_ns = gbs ? int(2.0 * tile_kb * 1024 * sms / (gbs * 1e9) * 1e9) : 0;
          tm.run([&] { kc<<<sms, 32, smem>>>(dst, src, bytes, tile_kb * 1024, S, pace_ns); }, 2, 10, &best, &med);
          CK(cudaGetLastError());
          snprintf(param, sizeof param, "tileKB=%d S=%d target_GBps=%d pace_ns=%d", tile_kb, S, gbs, pace_ns);
          report("copy_paced", param, 2.0 * bytes, best, med);
        }
      }
  }
  CK(cudaFree(slab));
  return 0;
}
