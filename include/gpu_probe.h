/*
 * include/gpu_probe.h -- C ABI of the B200 validation probe (libgpuprobe.so).
 *
 * What this boundary replaces.  The reference (maryamtahhan/kind-gpu-sim) has no
 * FFI and no compute: a GPU-requesting pod runs
 *     /root/reference/pods/nvidia-gpu-test-pod.yaml:9    echo ... && sleep 3600
 * after the scheduler matched `limits nvidia.com/gpu` (same file :10-12) against
 * the capacity that /root/reference/kind-gpu-sim.sh:114-116 advertises.  The probe
 * fills that container-command slot with real work.  The entry points below are
 * what a host in the reference's own toolchain (bash -> a small C binary, see
 * kind-gpu-sim_b200/host/gpu_probe_main.c) binds; SURVEY.md section 8(b) lists
 * them.  Each comment names the reference line whose slot the call fills.
 *
 * Conventions
 *  - plain C: pointers, sizes, no CUDA/torch types.  `stream` is a cudaStream_t
 *    passed as void* (NULL = the legacy default stream).
 *  - every pointer called "device" must be device memory of the CURRENT device,
 *    8-byte aligned; 16-byte alignment enables the vector/TMA paths (cudaMalloc
 *    gives 256).  The caller owns every buffer; nothing here allocates unless
 *    its comment says so.
 *  - all calls are asynchronous on `stream` unless marked [blocking].
 *  - return value: 0 = ok; > 0 = cudaError_t; -1..-99 = -(ncclResult_t);
 *    <= -100 = GPU_PROBE_E* below.  gpu_probe_error_string() explains any of them.
 *  - there is NO CPU fallback: on a machine without an sm_100 GPU every compute
 *    entry point returns a CUDA error.
 */
#ifndef GPU_PROBE_H
#define GPU_PROBE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GPU_PROBE_ABI_VERSION 1

/* Spec constants. */
#define GPU_PROBE_SEED_B 0x0123456789ABCDEFull
#define GPU_PROBE_SEED_C 0xFEDCBA9876543210ull
#define GPU_PROBE_SCALAR 3.0

/* Kernel variants (how a streaming kernel moves its bytes). */
#define GPU_PROBE_VARIANT_AUTO (-1) /* the variant measured fastest on B200      */
#define GPU_PROBE_VARIANT_LDG 0     /* 128-bit coalesced ld/st, streaming hints   */
#define GPU_PROBE_VARIANT_TMA 1     /* cp.async.bulk + mbarrier smem ring         */
#define GPU_PROBE_VARIANT_LDG256 2  /* 256-bit ld/st (sm_100 v4.b64)              */
#define GPU_PROBE_VARIANT_TMA_PACED 3 /* TMA ring with paced refills (triad only)   */

/* Probe-level errors (<= -100). */
#define GPU_PROBE_EINVAL (-100)    /* null pointer, bad variant, bad size         */
#define GPU_PROBE_EALIGN (-101)    /* pointer not 8-byte aligned                  */
#define GPU_PROBE_ETOOBIG (-102)   /* size beyond a stated maximum                */
#define GPU_PROBE_ENOTREADY (-103) /* object used before create / after destroy   */
#define GPU_PROBE_ETIMEDOUT (-104) /* a peer GPU never reached the exchange        */

int gpu_probe_abi_version(void);

/* Static description: arch, variants compiled in.  Never NULL. */
const char *gpu_probe_kernel_info(void);

/* Text for any return value of this library.  Never NULL. */
const char *gpu_probe_error_string(int rc);

/* K0  dst[k] = (double)(splitmix64(first_index + k, seed) >> 11) * 2^-53.
 * Synthetic input generated on the device: the pod has no input to copy
 * (slot: pods/nvidia-gpu-test-pod.yaml:9; spec: SURVEY 8d "Generator"). */
int gpu_probe_init_fill(double *dst, uint64_t n, uint64_t first_index,
                        uint64_t seed, void *stream);

/* K1  a[k] = fma(s, c[k], b[k]), one rounding.  24 bytes of HBM traffic per
 * element.  (slot: pods/nvidia-gpu-test-pod.yaml:9; spec: SURVEY 8d "Triad")
 * Asynchronous for every variant, AUTO included: it never synchronises, times or calibrates
 * anything.  VARIANT_AUTO runs the TMA ring; on a device that gpu_probe_triad_calibrate() has
 * calibrated it runs the paced TMA kernel (n >= 2^26, 16-byte-aligned pointers) at that
 * device's pace instead, +3-4 % on this fleet.
 * Environment: GPU_PROBE_TRIAD_PACE_NS=0 disables the paced kernel, =N fixes the pace. */
int gpu_probe_triad(double *a, const double *b, const double *c, double s,
                    uint64_t n, int variant, void *stream);

/* [blocking, ALLOCATES 64 bytes of pinned host memory on a device's first calibration]
 * Finds the pace of the paced triad for the CURRENT device: 27 timed launches of the paced
 * kernel over min(n, 2^29) elements of the caller's own buffers plus a stream synchronise,
 * about 50 ms.  a[] is overwritten (with the triad for s = GPU_PROBE_SCALAR); b[] and c[] are
 * only read, and a[] must not overlap either of them (GPU_PROBE_EINVAL) -- an in-place triad
 * cannot be replayed.  Not inside a stream capture (cudaErrorStreamCaptureUnsupported).
 * With n < 2^26 or pointers that are not 16-byte aligned there is nothing to calibrate (AUTO
 * never paces such calls): returns 0 and changes nothing.
 * Call it once per device before the timed or captured part of the program -- and again after
 * anything that changes what the memory system sustains (creating an NCCL communicator costs
 * every local HBM kernel ~3.4 % on this stack).  Nothing else in this library ever runs a
 * calibration: a device that was never calibrated simply keeps the unpaced ring.
 * Do not launch other work on the device while it runs; calibrations of one device from
 * several threads are serialised.
 * Drift guard: every paced AUTO launch measures itself on the device, and its last CTA leaves the
 * verdict in that pinned host block, which the next call reads with a plain load (no
 * synchronisation).  After 3 launches in a row that took more than 1.04 x (tiles per CTA x pace)
 * + 10 us -- the memory system sustains less than when the device was calibrated -- AUTO goes
 * back to the ring on that device until it is calibrated again.  Paced AUTO launches that
 * overlap on one device share that record and may trip the guard early; that costs speed only.
 * The verdict and the addresses behind the guard are remembered per device for the life of the
 * process; after cudaDeviceReset() call gpu_probe_triad_recalibrate() and calibrate again. */
int gpu_probe_triad_calibrate(double *a, const double *b, const double *c, uint64_t n,
                              void *stream);

/* State of VARIANT_AUTO's pacing on the current device: 0 = never calibrated (ring), -1 =
 * calibrated, the unpaced ring is at least as fast here, -2 = calibrated, then sent back to the
 * ring by the drift guard, > 0 = nanoseconds between a CTA's refills. */
int gpu_probe_triad_pace_ns(void);

/* What the drift guard last saw on the current device, read from the pinned host block without
 * synchronising (so it describes the most recent paced AUTO launch that has FINISHED):
 * out[0] = that launch's duration in ns, first CTA start to last CTA end on the device's clock;
 * out[1] = how many launches in a row overran their limit.  Zeros before the first calibration. */
int gpu_probe_triad_pace_report(uint64_t out[2]);

/* Forget the current device's calibration: VARIANT_AUTO runs the ring until the next
 * gpu_probe_triad_calibrate(). */
int gpu_probe_triad_recalibrate(void);

/* Bytes of device scratch the checksum kernels need (a constant). */
size_t gpu_probe_checksum_workspace_bytes(void);

/* K2  out_lanes[0] = sum w[k], out_lanes[1] = sum w[k]*(2*(first_index+k)+1),
 * both mod 2^64.  One launch; out_lanes (device, 2 words) is overwritten.
 * `workspace` is device scratch of gpu_probe_checksum_workspace_bytes() bytes,
 * zeroed ONCE by the caller after allocation (the kernel leaves it zeroed);
 * calls that overlap on the GPU must not share a workspace.
 * For n >= 2^26 the TMA variant spaces its refills 650 ns apart (read-only traffic has no
 * over-subscription cliff, so no calibration); GPU_PROBE_CHECKSUM_PACE_NS=0 disables, =N sets.
 * (slot: pods/nvidia-gpu-test-pod.yaml:9; spec: SURVEY 8d "Checksum") */
int gpu_probe_checksum_u64(const uint64_t *w, uint64_t n, uint64_t first_index,
                           uint64_t *out_lanes, void *workspace, int variant,
                           void *stream);

/* K2 without any allocation, for hosts that cannot obtain device scratch.
 * No counterpart in the reference (it has no compute at all); it exists for callers
 * of THIS library that are not able to cudaMalloc the checksum workspace: a health
 * check embedded in a process whose device memory is wholly owned by another
 * allocator (a serving engine that reserved all of HBM at start-up, a plugin that
 * runs under a no-allocation-after-init policy), or a monitoring loop that must not
 * touch the allocator while the workload it watches is running.
 * Same result as gpu_probe_checksum_u64 with VARIANT_AUTO, but the scratch is a
 * 65,792-byte array that lives inside the library's device image: every GPU that
 * loads libgpuprobe.so carries it until the process exits.
 *  - limit: n <= GPU_PROBE_NOALLOC_MAX_WORDS, else GPU_PROBE_ETOOBIG (nothing runs).
 *  - makes no cudaMalloc/cudaFree/allocating call; launches one 256-byte memset and
 *    one kernel on `stream`.
 *  - the first call on each device looks the array's device address up (a one-time
 *    driver query that may also load the library's device code): make that first
 *    call before starting any stream capture on that device.  The address is remembered
 *    for the life of the process: a host that calls cudaDeviceReset() destroys the array
 *    the remembered address points at, and must not call this entry point afterwards.
 *  - all calls on a device share the one array, which still holds what the previous
 *    call wrote; the entry point re-zeroes the only word that must start at zero.
 *    Two calls that overlap on the GPU (different streams, no ordering) race on it
 *    and return garbage -- serialise them, or use gpu_probe_checksum_u64 with one
 *    workspace per stream. */
#define GPU_PROBE_NOALLOC_MAX_WORDS (1ull << 34) /* 128 GiB of words: above any B200 buffer */
int gpu_probe_checksum_u64_noalloc(const uint64_t *w, uint64_t n, uint64_t first_index,
                                   uint64_t *out_lanes, void *stream);

/* K1+K2 fused: writes a[] and the lanes of a[] in one pass (24 B/element instead
 * of 32).  Same lanes as gpu_probe_triad followed by gpu_probe_checksum_u64.
 *  */
int gpu_probe_triad_checksum(double *a, const double *b, const double *c,
                             double s, uint64_t n, uint64_t first_index,
                             uint64_t *out_lanes, void *workspace, int variant,
                             void *stream);

/* K3  SM-driven copy dst[k] = src[k]; dst and/or src may be a peer GPU's memory
 * once peer access is enabled (NVLink 5 path).  bytes need not be a multiple of
 * 16.  (spec: SURVEY 8d "P2P copy") */
int gpu_probe_p2p_copy(void *dst, const void *src, uint64_t bytes, void *stream);

/* ---- multi-GPU: the checksum fused with its collective ----
 * Sharded checksum: GPU g of G checksums words [g*N/G, (g+1)*N/G) of one global index
 * space, then the 16-byte lanes are summed over GPUs.  The two-launch form is K2
 * followed by ncclAllReduce(count=2, ncclUint64, ncclSum) (that is what
 * kind-gpu-sim_b200/host/ and bench.py time as the baseline).  The *_exchange entry
 * points do both in ONE kernel: the last CTA stores its lanes into a mailbox on every
 * peer over NVLink (P2P stores + a release flag), waits for the peers' stores in its
 * own mailbox and writes the global sum to out_lanes -- every rank ends with the same
 * two words, bit-identical to a single-GPU checksum of the whole buffer.
 *
 * A gpu_probe_exchange is one rank's symmetric segment -- a 512 KiB header (lanes mailbox,
 * barrier flags) plus, when buffer_bytes > 0, a data buffer (the allreduce operand) and
 * 2 x world x min(buffer_bytes, 2 MiB) of staging -- and the addresses of the peers' segments.
 *   create          [blocking, ALLOCATES the segment] on the current device.
 *   connect_local   ranks living in one process: enables peer access both ways and
 *                   hands raw pointers around.  xs[r] must be rank r.
 *   export / connect_ipc   one process per GPU: export gives 64 opaque bytes (they name a
 *                   unix socket of this rank); carry all `world` of them (rank order) to
 *                   every rank by any host channel and pass them to connect_ipc, which
 *                   trades the segments' file descriptors over those sockets [blocking;
 *                   every rank must call it].
 *   status          [blocking] 0, or GPU_PROBE_ETIMEDOUT if some call's wait for a peer
 *                   ran out (default 10 s; the kernel then writes all-ones lanes).
 * Rules: every rank makes the same sequence of *_exchange calls (it is a collective);
 * calls on one rank are stream-ordered (never two in flight on one mailbox); the
 * kernels of all ranks must be able to run at the same time (do not serialise the
 * ranks' streams behind each other). */
typedef struct gpu_probe_exchange gpu_probe_exchange;
#define GPU_PROBE_EXCHANGE_MAX_RANKS 16
#define GPU_PROBE_EXCHANGE_HANDLE_BYTES 64

int gpu_probe_exchange_create(gpu_probe_exchange **out, int rank, int world,
                              uint64_t buffer_bytes);
int gpu_probe_exchange_connect_local(gpu_probe_exchange *const *xs, int world);
int gpu_probe_exchange_export(gpu_probe_exchange *x,
                              unsigned char handle[GPU_PROBE_EXCHANGE_HANDLE_BYTES]);
int gpu_probe_exchange_connect_ipc(gpu_probe_exchange *x, const unsigned char *handles);
int gpu_probe_exchange_set_timeout_ms(gpu_probe_exchange *x, uint64_t ms);
int gpu_probe_exchange_status(gpu_probe_exchange *x);
int gpu_probe_exchange_destroy(gpu_probe_exchange *x);

/* 1 when the connected exchange also holds an NVSwitch multicast mapping of all ranks'
 * segments (set up by connect_local / connect_ipc when every device and the fabric allow it). */
int gpu_probe_exchange_has_multicast(gpu_probe_exchange *x);

/* This rank's data buffer inside the segment (device pointer; NULL when buffer_bytes was 0). */
void *gpu_probe_exchange_buffer(gpu_probe_exchange *x);
uint64_t gpu_probe_exchange_buffer_bytes(gpu_probe_exchange *x);

/* K4  in-place fp32 sum allreduce of `count` elements at buffer + offset_bytes (16-byte
 * aligned offset) over all ranks, in ONE kernel over NVLink peer memory.  All ranks end with
 * bit-identical data: every element is the fp32 sum of the ranks' values in rank order.
 *   PULL_PUSH  rank r pulls its slice from every rank, adds, pushes the sums back to every
 *              rank; 2(G-1)/G x bytes per direction per GPU -- the bandwidth algorithm.
 *   PUSH       every rank pushes its whole operand into staging on every peer and adds up
 *              what lands in its own; one NVLink traversal -- the latency algorithm.
 *              Needs count % 4 == 0 and 4*count <= 2 MiB, else GPU_PROBE_ETOOBIG.
 *   MULTIMEM   as PULL_PUSH, but reduction and broadcast happen inside the NVSwitch
 *              (multimem.ld_reduce / multimem.st on the multicast alias of the segments):
 *              (1 + 1/G) x bytes per direction.  Needs gpu_probe_exchange_has_multicast();
 *              else GPU_PROBE_ENOTREADY.  The switch picks the summation order, so the
 *              result may differ from the rank-order sum in the last place (ranks still
 *              agree with each other bit for bit).
 *   AUTO       PUSH below a per-world-size threshold; above it MULTIMEM when the exchange
 *              has multicast and G >= 3, else PULL_PUSH.
 * `ctas` = 0 lets the library size the grid.  Every rank must pass the same count, algo and
 * ctas.  World sizes 1..8.  This is the probe's own collective, timed beside NCCL's
 * allreduce (SURVEY 8d "Allreduce sweep"); same rules as the *_exchange calls above. */
#define GPU_PROBE_ALLREDUCE_AUTO 0
#define GPU_PROBE_ALLREDUCE_PULL_PUSH 1
#define GPU_PROBE_ALLREDUCE_PUSH 2
#define GPU_PROBE_ALLREDUCE_MULTIMEM 3
int gpu_probe_allreduce_f32(gpu_probe_exchange *x, uint64_t offset_bytes, uint64_t count,
                            int algo, int ctas, void *stream);

/* K2 + lanes exchange in one launch.  Arguments as gpu_probe_checksum_u64; out_lanes
 * receives the sum over ALL ranks.  Must be called with x's device current. */
int gpu_probe_checksum_u64_exchange(const uint64_t *w, uint64_t n, uint64_t first_index,
                                    uint64_t *out_lanes, void *workspace, int variant,
                                    gpu_probe_exchange *x, void *stream);

/* Fused K1+K2 + lanes exchange in one launch (24 B/element, no second kernel, no
 * NCCL launch).  Arguments as gpu_probe_triad_checksum. */
int gpu_probe_triad_checksum_exchange(double *a, const double *b, const double *c,
                                      double s, uint64_t n, uint64_t first_index,
                                      uint64_t *out_lanes, void *workspace, int variant,
                                      gpu_probe_exchange *x, void *stream);

/* Single-GPU self-test of the multi-GPU protocols.  Ranks that wait for each other must be
 * resident on the GPU at the same time, and nothing guarantees that for separate launches
 * on ONE device.  On a box with fewer GPUs than ranks, create all `world` exchanges on the
 * current device, connect_local them, and use these calls: every rank runs in the rows of
 * ONE launch (cooperative for the allreduce), executing the same device code as the
 * per-rank entry points.  Diagnostics, never faster than anything; world <= 8; MULTIMEM is
 * not available (one device cannot form a multicast group); the allreduce returns
 * GPU_PROBE_ETOOBIG when world x ctas CTAs cannot be resident together (pass a small ctas).
 * Array arguments are host arrays of `world` entries, index = rank. */
int gpu_probe_checksum_u64_exchange_group(const uint64_t *const *w, const uint64_t *n,
                                          const uint64_t *first_index, uint64_t *const *out_lanes,
                                          void *const *workspace, gpu_probe_exchange *const *xs,
                                          int world, void *stream);
int gpu_probe_allreduce_f32_group(gpu_probe_exchange *const *xs, int world, uint64_t offset_bytes,
                                  uint64_t count, int algo, int ctas, void *stream);

/* ---- host-buffer pipeline (the end-to-end call for a host that holds b, c) ----
 * create [blocking, ALLOCATES]: device staging for `depth` chunks of
 * `chunk_elems` doubles x3 arrays, two streams, events -- on the current device.
 * run_host [blocking]: streams pinned host b,c through the device in chunks
 * (H2D overlapped with the fused triad+checksum kernel), optionally copies a[]
 * back, and returns the lanes in host memory.  Host buffers should be pinned
 * (cudaHostAlloc / cudaHostRegister) or the copies serialise. */
typedef struct gpu_probe_pipeline gpu_probe_pipeline;

int gpu_probe_pipeline_create(gpu_probe_pipeline **out, uint64_t chunk_elems,
                              int depth);
int gpu_probe_pipeline_run_host(gpu_probe_pipeline *p, const double *b_host,
                                const double *c_host, double *a_host_or_null,
                                double s, uint64_t n, uint64_t first_index,
                                uint64_t lanes_out_host[2]);
int gpu_probe_pipeline_destroy(gpu_probe_pipeline *p);

#ifdef __cplusplus
}
#endif
#endif /* GPU_PROBE_H */
