/*
 * gpu-probe -- the container command of a GPU-requesting probe pod.
 *
 * Slot it fills: /root/reference/pods/nvidia-gpu-test-pod.yaml:9, where the reference
 * runs `echo Hello from fake NVIDIA GPU node && sleep 3600` after the scheduler matched
 * `limits nvidia.com/gpu` (same file :10-12).  This binary does the real work instead:
 *
 *   1. contract with the device plugin: cudaGetDeviceCount() must equal the pod's
 *      nvidia.com/gpu limit (--expect-gpus=N), else exit 3;
 *   2. per GPU, all GPUs at once: K0 fill, K1 triad, K2 checksum, fused K1+K2 -- HBM GB/s
 *      against the roofline, lanes against the known answers (golden_lanes.h);
 *   3. N > 1: the sharded checksum combined over GPUs, two ways -- K2 followed by an NCCL
 *      u64 allreduce, and the one-kernel checksum+exchange over peer memory;
 *   4. N > 1: fp32 allreduce sweep 1 KiB..1 GiB at 2/4/8 GPUs (bus GB/s against 900 GB/s per
 *      direction), NCCL and the probe's own one-kernel allreduce side by side, and P2P
 *      copies GPU0 -> GPUk (SM copy kernel and copy engine).
 *
 * Everything numeric goes through the C ABI in include/gpu_probe.h (libgpuprobe.so,
 * hand-written sm_100a kernels).  There is no CPU path in this program: the expected
 * lanes are constants.  NCCL is loaded with dlopen only when more than one GPU is probed,
 * so the 1-GPU pod image needs no libnccl.
 *
 * Output (stdout -> `kubectl logs`), SURVEY.md section 8(b):
 *   RESULT kind=... gpus=... gpu=... bytes=... ms=... ms_med=... gbs=... frac_nominal=...
 *          frac_measured=... lanes=<hex16>,<hex16> match=<1|0|na>
 *   PROBE <PASS|FAIL> rc=<n>
 * Exit codes: 0 pass; 2 usage; 3 device-count mismatch; 4 CUDA error; 5 NCCL error;
 * 6 parity mismatch.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <inttypes.h>
#include <pthread.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include <cuda_runtime_api.h>
#include <nccl.h> /* types and enums only; the functions are resolved with dlsym */

#include "golden_lanes.h"
#include "gpu_probe.h"

#define MAX_GPUS 16
#define NOMINAL_HBM_GBS 8000.0
#define NOMINAL_NVLINK_GBS 900.0

enum { RC_OK = 0, RC_USAGE = 2, RC_COUNT = 3, RC_CUDA = 4, RC_NCCL = 5, RC_PARITY = 6 };

/* ------------------------------------------------------------------ options */
typedef struct {
  int expect_gpus;      /* -1: accept whatever is visible */
  uint64_t bytes;       /* per array per GPU */
  int iters, warmup;
  int variant;
  const char *variant_name;
  uint64_t sweep_max_bytes, p2p_bytes;
  int do_sweep, do_p2p, do_sharded;
  double measured_hbm_gbs; /* 0: unknown -> frac_measured=na */
  int max_gpus;            /* use at most this many of the visible GPUs (0: all) */
  int allreduce_algo;      /* GPU_PROBE_ALLREDUCE_* for the probe's own allreduce */
  const char *allreduce_algo_name;
  int inject_fault;        /* flip one bit of a[] between K1 and K2: the run must end in rc 6 */
} options;

static int g_status = RC_OK; /* first failure wins */
static void fail(int rc) { if (g_status == RC_OK) g_status = rc; }

static double now_s(void) {
  struct timespec ts;
  clock_gettime(CLOCK_REALTIME, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static void die_usage(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  fprintf(stderr, "gpu-probe: ");
  vfprintf(stderr, fmt, ap);
  va_end(ap);
  fprintf(stderr,
          "\nusage: gpu-probe [--expect-gpus=N] [--bytes=B] [--iters=K] [--warmup=W]\n"
          "                 [--variant=auto|ldg128|tma|ldg256] [--max-gpus=N]\n"
          "                 [--sweep-max-bytes=B] [--p2p-bytes=B] [--no-sweep] [--no-p2p] [--no-sharded]\n"
          "                 [--measured-hbm-gbs=X] [--allreduce-algo=auto|pull_push|push|multimem]\n"
          "                 [--inject-fault]\n"
          "environment: GPU_PROBE_MEASURED_HBM_GBS=X (as the flag; the flag wins)\n");
  printf("PROBE FAIL rc=%d\n", RC_USAGE);
  exit(RC_USAGE);
}

static uint64_t parse_u64(const char *flag, const char *s) {
  char *end = NULL;
  if (!*s || *s == '-') die_usage("%s: not a number: '%s'", flag, s);
  unsigned long long v = strtoull(s, &end, 0);
  if (end && (*end == 'k' || *end == 'K')) { v <<= 10; ++end; }
  else if (end && (*end == 'm' || *end == 'M')) { v <<= 20; ++end; }
  else if (end && (*end == 'g' || *end == 'G')) { v <<= 30; ++end; }
  if (!end || *end) die_usage("%s: not a number: '%s'", flag, s);
  return (uint64_t)v;
}

static const char *flag_value(const char *arg, const char *name) {
  size_t n = strlen(name);
  return (strncmp(arg, name, n) == 0 && arg[n] == '=') ? arg + n + 1 : NULL;
}

static void parse_options(int argc, char **argv, options *o) {
  o->expect_gpus = -1;
  o->bytes = 4ull << 30; /* BASELINE.json configs[1]: 4 GiB arrays */
  o->iters = 20;
  o->warmup = 3;
  o->variant = GPU_PROBE_VARIANT_AUTO;
  o->variant_name = "auto";
  o->sweep_max_bytes = 1ull << 30;
  o->p2p_bytes = 1ull << 30;
  o->do_sweep = o->do_p2p = o->do_sharded = 1;
  o->measured_hbm_gbs = 0.0;
  o->max_gpus = 0;
  o->allreduce_algo = GPU_PROBE_ALLREDUCE_AUTO;
  o->allreduce_algo_name = "auto";
  o->inject_fault = 0;
  { /* the pod runner / operator exports what this box's HBM was measured to sustain */
    const char *m = getenv("GPU_PROBE_MEASURED_HBM_GBS");
    if (m && *m) o->measured_hbm_gbs = atof(m);
  }
  for (int i = 1; i < argc; ++i) {
    const char *a = argv[i], *v;
    if ((v = flag_value(a, "--expect-gpus"))) o->expect_gpus = (int)parse_u64("--expect-gpus", v);
    else if ((v = flag_value(a, "--bytes"))) o->bytes = parse_u64("--bytes", v);
    else if ((v = flag_value(a, "--iters"))) o->iters = (int)parse_u64("--iters", v);
    else if ((v = flag_value(a, "--warmup"))) o->warmup = (int)parse_u64("--warmup", v);
    else if ((v = flag_value(a, "--max-gpus"))) o->max_gpus = (int)parse_u64("--max-gpus", v);
    else if ((v = flag_value(a, "--sweep-max-bytes"))) o->sweep_max_bytes = parse_u64("--sweep-max-bytes", v);
    else if ((v = flag_value(a, "--p2p-bytes"))) o->p2p_bytes = parse_u64("--p2p-bytes", v);
    else if ((v = flag_value(a, "--measured-hbm-gbs"))) o->measured_hbm_gbs = atof(v);
    else if ((v = flag_value(a, "--variant"))) {
      o->variant_name = v;
      if (!strcmp(v, "auto")) o->variant = GPU_PROBE_VARIANT_AUTO;
      else if (!strcmp(v, "ldg128")) o->variant = GPU_PROBE_VARIANT_LDG;
      else if (!strcmp(v, "tma")) o->variant = GPU_PROBE_VARIANT_TMA;
      else if (!strcmp(v, "ldg256")) o->variant = GPU_PROBE_VARIANT_LDG256;
      else die_usage("--variant: unknown variant '%s'", v);
    } else if ((v = flag_value(a, "--allreduce-algo"))) {
      o->allreduce_algo_name = v;
      if (!strcmp(v, "auto")) o->allreduce_algo = GPU_PROBE_ALLREDUCE_AUTO;
      else if (!strcmp(v, "pull_push")) o->allreduce_algo = GPU_PROBE_ALLREDUCE_PULL_PUSH;
      else if (!strcmp(v, "push")) o->allreduce_algo = GPU_PROBE_ALLREDUCE_PUSH;
      else if (!strcmp(v, "multimem")) o->allreduce_algo = GPU_PROBE_ALLREDUCE_MULTIMEM;
      else die_usage("--allreduce-algo: unknown algorithm '%s'", v);
    } else if (!strcmp(a, "--no-sweep")) o->do_sweep = 0;
    else if (!strcmp(a, "--no-p2p")) o->do_p2p = 0;
    else if (!strcmp(a, "--no-sharded")) o->do_sharded = 0;
    else if (!strcmp(a, "--inject-fault")) o->inject_fault = 1;
    else if (!strcmp(a, "--help") || !strcmp(a, "-h")) die_usage("help");
    else die_usage("unknown argument '%s'", a);
  }
  if (o->bytes < 8 || (o->bytes & 7)) die_usage("--bytes must be a positive multiple of 8");
  if (o->iters < 1 || o->iters > 10000) die_usage("--iters must be in 1..10000");
  if (o->warmup < 0 || o->warmup > 10000) die_usage("--warmup must be in 0..10000");
  if (o->sweep_max_bytes < 1024 || (o->sweep_max_bytes & (o->sweep_max_bytes - 1)))
    die_usage("--sweep-max-bytes must be a power of two >= 1024");
  if (o->p2p_bytes < 16) die_usage("--p2p-bytes must be >= 16");
}

/* ------------------------------------------------------------- error plumbing */
#define CU(call)                                                                              \
  do {                                                                                        \
    cudaError_t e__ = (call);                                                                 \
    if (e__ != cudaSuccess) {                                                                 \
      fprintf(stderr, "gpu-probe: CUDA error %d (%s) at %s:%d: %s\n", (int)e__,               \
              cudaGetErrorString(e__), __FILE__, __LINE__, #call);                            \
      return RC_CUDA;                                                                         \
    }                                                                                         \
  } while (0)

#define PR(call)                                                                              \
  do {                                                                                        \
    int r__ = (call);                                                                         \
    if (r__ != 0) {                                                                           \
      fprintf(stderr, "gpu-probe: %s at %s:%d: %s\n", gpu_probe_error_string(r__), __FILE__,  \
              __LINE__, #call);                                                               \
      return RC_CUDA;                                                                         \
    }                                                                                         \
  } while (0)

/* ------------------------------------------------------------------- NCCL */
typedef struct {
  void *so;
  ncclResult_t (*GetVersion)(int *);
  ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*GroupStart)(void);
  ncclResult_t (*GroupEnd)(void);
  const char *(*GetErrorString)(ncclResult_t);
  /* optional (NCCL >= 2.19 / 2.27): user-buffer registration, what SURVEY 8(d) specifies for the sweep */
  ncclResult_t (*MemAlloc)(void **, size_t);
  ncclResult_t (*MemFree)(void *);
  ncclResult_t (*CommRegister)(const ncclComm_t, void *, size_t, void **);
  ncclResult_t (*CommDeregister)(const ncclComm_t, void *);
  ncclResult_t (*CommWindowRegister)(ncclComm_t, void *, size_t, ncclWindow_t *, int);
  ncclResult_t (*CommWindowDeregister)(ncclComm_t, ncclWindow_t);
} nccl_api;

static int nccl_load(nccl_api *n) {
  const char *names[] = {getenv("GPU_PROBE_NCCL_LIB"), "libnccl.so.2", "libnccl.so", NULL};
  memset(n, 0, sizeof(*n));
  for (int i = 0; i < 4 && !n->so; ++i)
    if (names[i] && *names[i]) n->so = dlopen(names[i], RTLD_NOW | RTLD_LOCAL);
  if (!n->so) {
    fprintf(stderr, "gpu-probe: cannot load NCCL (libnccl.so.2): %s\n", dlerror());
    return RC_NCCL;
  }
#define SYM(field, name)                                                     \
  do {                                                                       \
    *(void **)(&n->field) = dlsym(n->so, name);                              \
    if (!n->field) {                                                         \
      fprintf(stderr, "gpu-probe: NCCL symbol %s missing\n", name);          \
      dlclose(n->so);                                                        \
      memset(n, 0, sizeof(*n)); /* so == NULL: every phase skips NCCL */     \
      return RC_NCCL;                                                        \
    }                                                                        \
  } while (0)
  SYM(GetVersion, "ncclGetVersion");
  SYM(CommInitAll, "ncclCommInitAll");
  SYM(CommDestroy, "ncclCommDestroy");
  SYM(AllReduce, "ncclAllReduce");
  SYM(GroupStart, "ncclGroupStart");
  SYM(GroupEnd, "ncclGroupEnd");
  SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
  *(void **)(&n->MemAlloc) = dlsym(n->so, "ncclMemAlloc");
  *(void **)(&n->MemFree) = dlsym(n->so, "ncclMemFree");
  *(void **)(&n->CommRegister) = dlsym(n->so, "ncclCommRegister");
  *(void **)(&n->CommDeregister) = dlsym(n->so, "ncclCommDeregister");
  *(void **)(&n->CommWindowRegister) = dlsym(n->so, "ncclCommWindowRegister");
  *(void **)(&n->CommWindowDeregister) = dlsym(n->so, "ncclCommWindowDeregister");
  return 0;
}

#define NC(api, call)                                                                         \
  do {                                                                                        \
    ncclResult_t r__ = (call);                                                                \
    if (r__ != ncclSuccess) {                                                                 \
      fprintf(stderr, "gpu-probe: NCCL error %d (%s) at %s:%d: %s\n", (int)r__,               \
              (api)->GetErrorString(r__), __FILE__, __LINE__, #call);                         \
      return RC_NCCL;                                                                         \
    }                                                                                         \
  } while (0)

/* ------------------------------------------------------------- small helpers */
static int cmp_float(const void *a, const void *b) {
  float x = *(const float *)a, y = *(const float *)b;
  return (x > y) - (x < y);
}

typedef struct { double best_ms, med_ms; } timing;

static timing summarize(float *ms, int n) {
  timing t;
  qsort(ms, (size_t)n, sizeof(float), cmp_float);
  t.best_ms = ms[0];
  t.med_ms = (n & 1) ? ms[n / 2] : 0.5 * (ms[n / 2 - 1] + ms[n / 2]);
  return t;
}

static void frac_measured(char *buf, size_t len, double gbs, double measured) {
  if (measured > 0) snprintf(buf, len, "%.4f", gbs / measured);
  else snprintf(buf, len, "na");
}

/* match field: "1"/"0" against a known answer, "na" when this size has none. */
static const char *match_known(uint64_t n_total, const uint64_t lanes[2], int *known) {
  const golden_entry *g = golden_lookup(n_total);
  *known = g != NULL;
  if (!g) return "na";
  if (g->lane0 == lanes[0] && g->lane1 == lanes[1]) return "1";
  fail(RC_PARITY);
  return "0";
}

/* ------------------------------------------------------------ per-GPU state */
typedef struct {
  int dev;
  cudaStream_t st;
  double *a, *b, *c;
  uint64_t *lanes; /* device: [0..1] checksum, [2..3] fused, [4..5] combined */
  void *ws;
  cudaEvent_t *ev; /* 2 per timed launch */
  uint64_t h_lanes[6];
  gpu_probe_exchange *xchg;
} gpu_state;

static int gpu_setup(gpu_state *g, int dev, uint64_t n, int nevents) {
  memset(g, 0, sizeof(*g));
  g->dev = dev;
  CU(cudaSetDevice(dev));
  CU(cudaStreamCreateWithFlags(&g->st, cudaStreamNonBlocking));
  CU(cudaMalloc((void **)&g->a, 8 * n));
  CU(cudaMalloc((void **)&g->b, 8 * n));
  CU(cudaMalloc((void **)&g->c, 8 * n));
  CU(cudaMalloc((void **)&g->lanes, 6 * sizeof(uint64_t)));
  CU(cudaMemset(g->lanes, 0xff, 6 * sizeof(uint64_t)));
  CU(cudaMalloc(&g->ws, gpu_probe_checksum_workspace_bytes()));
  CU(cudaMemset(g->ws, 0, gpu_probe_checksum_workspace_bytes()));
  g->ev = (cudaEvent_t *)calloc((size_t)nevents, sizeof(cudaEvent_t));
  if (!g->ev) return RC_CUDA;
  for (int i = 0; i < nevents; ++i) CU(cudaEventCreate(&g->ev[i]));
  CU(cudaDeviceSynchronize());
  return 0;
}

static int sync_all(gpu_state *gs, int ng) {
  for (int d = 0; d < ng; ++d) {
    CU(cudaSetDevice(gs[d].dev));
    CU(cudaStreamSynchronize(gs[d].st));
  }
  return 0;
}

/* ----------------------------------------------- phase A: per-GPU HBM kernels */
enum { K_FILL, K_TRIAD, K_CHECKSUM, K_FUSED, K_COUNT };
static const char *k_name[K_COUNT] = {"init_fill", "triad", "checksum", "triad_checksum"};
static const uint64_t k_bytes_per_elem[K_COUNT] = {8, 24, 8, 24};

static int launch_kernel(int k, gpu_state *g, uint64_t n, uint64_t first, int variant) {
  switch (k) {
    case K_FILL: return gpu_probe_init_fill(g->b, n, first, GPU_PROBE_SEED_B, g->st);
    case K_TRIAD: return gpu_probe_triad(g->a, g->b, g->c, GPU_PROBE_SCALAR, n, variant, g->st);
    case K_CHECKSUM:
      return gpu_probe_checksum_u64((const uint64_t *)g->a, n, first, g->lanes, g->ws, variant, g->st);
    default:
      return gpu_probe_triad_checksum(g->a, g->b, g->c, GPU_PROBE_SCALAR, n, first, g->lanes + 2, g->ws, variant,
                                      g->st);
  }
}

static int phase_hbm(const options *o, gpu_state *gs, int ng, uint64_t n) {
  float *ms = (float *)malloc(sizeof(float) * (size_t)o->iters);
  if (!ms) return RC_CUDA;
  /* inputs: b, c from the counter-based generator, global index space shared by all GPUs */
  for (int d = 0; d < ng; ++d) {
    CU(cudaSetDevice(gs[d].dev));
    PR(gpu_probe_init_fill(gs[d].b, n, (uint64_t)d * n, GPU_PROBE_SEED_B, gs[d].st));
    PR(gpu_probe_init_fill(gs[d].c, n, (uint64_t)d * n, GPU_PROBE_SEED_C, gs[d].st));
  }
  int rc = sync_all(gs, ng);
  if (rc) return rc;
  /* AUTO's paced triad: the one blocking step, made here, one GPU at a time (a device calibrated
   * while its neighbours stream lands on a slower pace), before anything is timed */
  if (o->variant == GPU_PROBE_VARIANT_AUTO)
    for (int d = 0; d < ng; ++d) {
      CU(cudaSetDevice(gs[d].dev));
      const double t0 = now_s();
      PR(gpu_probe_triad_calibrate(gs[d].a, gs[d].b, gs[d].c, n, gs[d].st));
      printf("PROBE INFO triad_calibrate gpu=%d pace_ns=%d ms=%.1f\n", gs[d].dev, gpu_probe_triad_pace_ns(),
             (now_s() - t0) * 1e3);
    }

  for (int k = 0; k < K_COUNT; ++k) {
    /* every GPU runs the same kernel at the same time (config 4: 8 concurrent probes) */
    for (int d = 0; d < ng; ++d) {
      CU(cudaSetDevice(gs[d].dev));
      for (int i = 0; i < o->warmup; ++i) PR(launch_kernel(k, &gs[d], n, (uint64_t)d * n, o->variant));
      for (int i = 0; i < o->iters; ++i) {
        CU(cudaEventRecord(gs[d].ev[2 * i], gs[d].st));
        PR(launch_kernel(k, &gs[d], n, (uint64_t)d * n, o->variant));
        CU(cudaEventRecord(gs[d].ev[2 * i + 1], gs[d].st));
      }
    }
    if ((rc = sync_all(gs, ng))) return rc;
    if (k == K_TRIAD && o->inject_fault) /* the checksum that follows must notice: match=0, exit 6 */
      for (int d = 0; d < ng; ++d) {
        uint64_t w = 0;
        const uint64_t at = n / 2;
        CU(cudaSetDevice(gs[d].dev));
        CU(cudaMemcpy(&w, gs[d].a + at, 8, cudaMemcpyDeviceToHost));
        w ^= 1ull << 17;
        CU(cudaMemcpy(gs[d].a + at, &w, 8, cudaMemcpyHostToDevice));
        printf("PROBE INFO inject_fault gpu=%d element=%" PRIu64 " bit=17\n", gs[d].dev, at);
      }
    double sum_gbs = 0, worst_ms = 0;
    const uint64_t bytes = k_bytes_per_elem[k] * n;
    for (int d = 0; d < ng; ++d) {
      CU(cudaSetDevice(gs[d].dev));
      for (int i = 0; i < o->iters; ++i) CU(cudaEventElapsedTime(&ms[i], gs[d].ev[2 * i], gs[d].ev[2 * i + 1]));
      timing t = summarize(ms, o->iters);
      const double gbs = (double)bytes / (t.best_ms * 1e-3) / 1e9;
      char fm[32], lanes[64] = "na";
      frac_measured(fm, sizeof fm, gbs, o->measured_hbm_gbs);
      if (k == K_CHECKSUM || k == K_FUSED) {
        uint64_t *h = gs[d].h_lanes + (k == K_FUSED ? 2 : 0);
        CU(cudaMemcpy(h, gs[d].lanes + (k == K_FUSED ? 2 : 0), 16, cudaMemcpyDeviceToHost));
        snprintf(lanes, sizeof lanes, "%016" PRIx64 ",%016" PRIx64, h[0], h[1]);
      }
      /* per-GPU shard lanes have a known answer only for the single-GPU whole buffer */
      const char *match = "na";
      int known = 0;
      if ((k == K_CHECKSUM || k == K_FUSED) && ng == 1) match = match_known(n, gs[d].h_lanes + (k == K_FUSED ? 2 : 0), &known);
      printf("RESULT kind=%s gpus=%d gpu=%d bytes=%" PRIu64 " ms=%.4f ms_med=%.4f gbs=%.1f frac_nominal=%.4f "
             "frac_measured=%s lanes=%s match=%s variant=%s pace_ns=%d\n",
             k_name[k], ng, gs[d].dev, bytes, t.best_ms, t.med_ms, gbs, gbs / NOMINAL_HBM_GBS, fm, lanes, match,
             o->variant_name, (k == K_TRIAD || k == K_FUSED) ? gpu_probe_triad_pace_ns() : 0);
      sum_gbs += (double)bytes / (t.med_ms * 1e-3) / 1e9;
      if (t.med_ms > worst_ms) worst_ms = t.med_ms;
    }
    if (ng > 1)
      printf("RESULT kind=%s gpus=%d gpu=all bytes=%" PRIu64 " ms=%.4f ms_med=%.4f gbs=%.1f frac_nominal=%.4f "
             "frac_measured=na lanes=na match=na variant=%s\n",
             k_name[k], ng, bytes * (uint64_t)ng, worst_ms, worst_ms, sum_gbs, sum_gbs / (NOMINAL_HBM_GBS * ng),
             o->variant_name);
    if (k == K_FILL) { /* the timed fill overwrote b with itself: same generator, same seed -- nothing to redo */ }
  }
  if (ng > 1) { /* shards of one index space: the host-side sum of the per-GPU lanes has a known answer */
    uint64_t sum[2] = {0, 0};
    int known = 0;
    for (int d = 0; d < ng; ++d) { sum[0] += gs[d].h_lanes[0]; sum[1] += gs[d].h_lanes[1]; }
    const char *match = match_known(n * (uint64_t)ng, sum, &known);
    printf("RESULT kind=checksum_sum gpus=%d gpu=all bytes=%" PRIu64 " ms=na ms_med=na gbs=na frac_nominal=na "
           "frac_measured=na lanes=%016" PRIx64 ",%016" PRIx64 " match=%s combine=host_sum\n",
           ng, 8 * n * (uint64_t)ng, sum[0], sum[1], match);
  }
  /* fused and separate must agree on every GPU, known answer or not */
  for (int d = 0; d < ng; ++d)
    if (gs[d].h_lanes[0] != gs[d].h_lanes[2] || gs[d].h_lanes[1] != gs[d].h_lanes[3]) {
      fprintf(stderr, "gpu-probe: gpu %d: fused lanes differ from triad+checksum lanes\n", gs[d].dev);
      fail(RC_PARITY);
    }
  free(ms);
  return 0;
}

/* ------------------------- phase B: sharded checksum, combined over the GPUs */
static int phase_sharded(const options *o, gpu_state *gs, int ng, uint64_t n, nccl_api *nc, ncclComm_t *comms) {
  float *ms = (float *)malloc(sizeof(float) * (size_t)o->iters);
  if (!ms) return RC_CUDA;
  int rc;
  /* a[] already holds the triad output of shard d (phase A).  Host sum of the per-GPU lanes: */
  uint64_t sum[2] = {0, 0};
  for (int d = 0; d < ng; ++d) { sum[0] += gs[d].h_lanes[0]; sum[1] += gs[d].h_lanes[1]; }

  for (int mode = 0; mode < 2; ++mode) { /* 0: K2 + ncclAllReduce(2 x u64); 1: one kernel, P2P exchange */
    if (mode == 0 && !nc->so) continue;
    for (int i = -o->warmup; i < o->iters; ++i) {
      if (mode == 0) {
        for (int d = 0; d < ng; ++d) {
          CU(cudaSetDevice(gs[d].dev));
          if (i >= 0) CU(cudaEventRecord(gs[d].ev[2 * i], gs[d].st));
          PR(gpu_probe_checksum_u64((const uint64_t *)gs[d].a, n, (uint64_t)d * n, gs[d].lanes + 4, gs[d].ws,
                                    o->variant, gs[d].st));
        }
        NC(nc, nc->GroupStart());
        for (int d = 0; d < ng; ++d)
          NC(nc, nc->AllReduce(gs[d].lanes + 4, gs[d].lanes + 4, 2, ncclUint64, ncclSum, comms[d], gs[d].st));
        NC(nc, nc->GroupEnd());
        for (int d = 0; d < ng; ++d) {
          CU(cudaSetDevice(gs[d].dev));
          if (i >= 0) CU(cudaEventRecord(gs[d].ev[2 * i + 1], gs[d].st));
        }
      } else {
        for (int d = 0; d < ng; ++d) {
          CU(cudaSetDevice(gs[d].dev));
          if (i >= 0) CU(cudaEventRecord(gs[d].ev[2 * i], gs[d].st));
          PR(gpu_probe_checksum_u64_exchange((const uint64_t *)gs[d].a, n, (uint64_t)d * n, gs[d].lanes + 4,
                                             gs[d].ws, o->variant, gs[d].xchg, gs[d].st));
          if (i >= 0) CU(cudaEventRecord(gs[d].ev[2 * i + 1], gs[d].st));
        }
      }
    }
    if ((rc = sync_all(gs, ng))) return rc;
    /* per iteration: the slowest GPU decides */
    for (int i = 0; i < o->iters; ++i) {
      ms[i] = 0;
      for (int d = 0; d < ng; ++d) {
        float t;
        CU(cudaSetDevice(gs[d].dev));
        CU(cudaEventElapsedTime(&t, gs[d].ev[2 * i], gs[d].ev[2 * i + 1]));
        if (t > ms[i]) ms[i] = t;
      }
    }
    timing t = summarize(ms, o->iters);
    int all_equal = 1;
    for (int d = 0; d < ng; ++d) {
      CU(cudaSetDevice(gs[d].dev));
      CU(cudaMemcpy(gs[d].h_lanes + 4, gs[d].lanes + 4, 16, cudaMemcpyDeviceToHost));
      if (gs[d].h_lanes[4] != sum[0] || gs[d].h_lanes[5] != sum[1]) all_equal = 0;
      if (mode == 1 && gpu_probe_exchange_status(gs[d].xchg) != 0) all_equal = 0;
    }
    if (!all_equal) {
      fprintf(stderr, "gpu-probe: sharded checksum (%s): a rank's combined lanes differ from the sum of the shards\n",
              mode ? "exchange" : "nccl");
      fail(RC_PARITY);
    }
    int known = 0;
    const char *match = match_known(n * (uint64_t)ng, gs[0].h_lanes + 4, &known);
    if (!known) match = all_equal ? "na" : "0";
    const uint64_t bytes = 8 * n * (uint64_t)ng;
    const double gbs = (double)bytes / (t.best_ms * 1e-3) / 1e9;
    printf("RESULT kind=sharded_checksum gpus=%d gpu=all bytes=%" PRIu64 " ms=%.4f ms_med=%.4f gbs=%.1f "
           "frac_nominal=%.4f frac_measured=na lanes=%016" PRIx64 ",%016" PRIx64 " match=%s combine=%s launches_per_gpu=%d\n",
           ng, bytes, t.best_ms, t.med_ms, gbs, gbs / (NOMINAL_HBM_GBS * ng), gs[0].h_lanes[4], gs[0].h_lanes[5],
           match, mode ? "p2p_exchange_fused" : "nccl_allreduce_u64", mode ? 1 : 2);
  }
  free(ms);
  return 0;
}

/* ------------- phase C: allreduce sweep at 2/4/8 GPUs -- NCCL and the probe's own K4 */
/* One timed series: `iters` back-to-back in-place allreduces of S bytes on every GPU of the
 * group, bracketed by events on each GPU's stream; per-iteration time = slowest GPU. */
/* One host thread per GPU issues that GPU's launches, as one process per GPU would: a single
 * thread feeding G GPUs makes every small-message figure G launch overheads long. */
typedef struct {
  const options *o;
  int use_nccl, d, G, iters, rc;
  nccl_api *nc;
  ncclComm_t comm;
  gpu_state *g;
  gpu_probe_exchange *x;
  float *buf;
  uint64_t S;
  pthread_barrier_t *bar;
  float ms;
} ar_worker;

static int ar_worker_run(ar_worker *w) {
  CU(cudaSetDevice(w->g->dev));
  for (int i = -5; i < w->iters; ++i) {
    if (i == 0) {
      CU(cudaStreamSynchronize(w->g->st));
      pthread_barrier_wait(w->bar); /* all GPUs start the timed series together */
      CU(cudaEventRecord(w->g->ev[0], w->g->st));
    }
    if (w->use_nccl) NC(w->nc, w->nc->AllReduce(w->buf, w->buf, w->S / 4, ncclFloat, ncclSum, w->comm, w->g->st));
    else PR(gpu_probe_allreduce_f32(w->x, 0, w->S / 4, w->o->allreduce_algo, 0, w->g->st));
  }
  CU(cudaEventRecord(w->g->ev[1], w->g->st));
  CU(cudaStreamSynchronize(w->g->st));
  CU(cudaEventElapsedTime(&w->ms, w->g->ev[0], w->g->ev[1]));
  return 0;
}

static void *ar_worker_main(void *p) {
  ar_worker *w = (ar_worker *)p;
  w->rc = ar_worker_run(w);
  if (w->rc) { /* do not leave the others waiting at the barrier for ever */
    fprintf(stderr, "gpu-probe: allreduce worker for gpu %d failed (rc %d); aborting\n", w->g->dev, w->rc);
    printf("PROBE FAIL rc=%d\n", w->rc);
    exit(w->rc);
  }
  return NULL;
}

/* One timed series: `iters` back-to-back in-place allreduces of S bytes on every GPU of the
 * group, bracketed by events on each GPU's stream; per-iteration time = slowest GPU. */
static int time_allreduce(const options *o, int use_nccl, nccl_api *nc, ncclComm_t *comms, gpu_state *gs,
                          gpu_probe_exchange **xs, float **bufs, int G, uint64_t S, int iters, double *ms_out) {
  pthread_t th[MAX_GPUS];
  ar_worker w[MAX_GPUS];
  pthread_barrier_t bar;
  if (pthread_barrier_init(&bar, NULL, (unsigned)G)) return RC_CUDA;
  for (int d = 0; d < G; ++d) {
    ar_worker v = {o, use_nccl, d, G, iters, 0, nc, use_nccl ? comms[d] : NULL, &gs[d], xs[d], bufs[d], S, &bar, 0.f};
    w[d] = v;
    if (pthread_create(&th[d], NULL, ar_worker_main, &w[d])) return RC_CUDA;
  }
  double ms = 0;
  for (int d = 0; d < G; ++d) {
    pthread_join(th[d], NULL);
    if (w[d].ms / iters > ms) ms = w[d].ms / iters;
  }
  pthread_barrier_destroy(&bar);
  *ms_out = ms;
  return 0;
}

/* Integer-valued floats in, closed-form sum out: exact in fp32 whatever the reduction order. */
static int check_allreduce(const options *o, int use_nccl, const char *impl, nccl_api *nc, ncclComm_t *comms, gpu_state *gs,
                           gpu_probe_exchange **xs, float **bufs, int G, uint64_t max_bytes, int *ok_out) {
  enum { CHECK_ELEMS = (1 << 18) + 3 };
  int rc;
  float *h = (float *)malloc(sizeof(float) * CHECK_ELEMS);
  if (!h) return RC_CUDA;
  const uint64_t celems = (max_bytes / 4 < CHECK_ELEMS) ? max_bytes / 4 : CHECK_ELEMS;
  for (int d = 0; d < G; ++d) {
    for (uint64_t i = 0; i < celems; ++i) h[i] = (float)((d + 1) * (int)(1 + (i % 7)));
    CU(cudaSetDevice(gs[d].dev));
    CU(cudaMemcpyAsync(bufs[d], h, 4 * celems, cudaMemcpyHostToDevice, gs[d].st));
    CU(cudaStreamSynchronize(gs[d].st));
  }
  if (use_nccl) {
    NC(nc, nc->GroupStart());
    for (int d = 0; d < G; ++d)
      NC(nc, nc->AllReduce(bufs[d], bufs[d], celems, ncclFloat, ncclSum, comms[d], gs[d].st));
    NC(nc, nc->GroupEnd());
  } else {
    for (int d = 0; d < G; ++d) {
      CU(cudaSetDevice(gs[d].dev));
      PR(gpu_probe_allreduce_f32(xs[d], 0, celems, o->allreduce_algo == GPU_PROBE_ALLREDUCE_PUSH ? GPU_PROBE_ALLREDUCE_AUTO : o->allreduce_algo, 0, gs[d].st));
    }
  }
  if ((rc = sync_all(gs, G))) return rc;
  int ok = 1;
  for (int d = 0; d < G && ok; ++d) {
    CU(cudaSetDevice(gs[d].dev));
    CU(cudaMemcpy(h, bufs[d], 4 * celems, cudaMemcpyDeviceToHost));
    for (uint64_t i = 0; i < celems; ++i)
      if (h[i] != (float)((G * (G + 1) / 2) * (int)(1 + (i % 7)))) { ok = 0; break; }
    if (!use_nccl && gpu_probe_exchange_status(xs[d]) != 0) ok = 0;
  }
  free(h);
  printf("RESULT kind=allreduce_check gpus=%d gpu=all bytes=%" PRIu64 " ms=na ms_med=na gbs=na frac_nominal=na "
         "frac_measured=na lanes=na match=%d impl=%s\n", G, 4 * celems, ok, impl);
  if (!ok) fail(RC_PARITY);
  *ok_out = ok;
  return 0;
}

static int phase_sweep(const options *o, gpu_state *gs, int ng, nccl_api *nc) {
  int rc;
  const int group_sizes[] = {2, 4, 8};
  for (int gi = 0; gi < 3; ++gi) {
    const int G = group_sizes[gi];
    if (G > ng) break;
    /* operands live in symmetric segments (one per GPU of the group): both NCCL and K4 reduce them in place */
    gpu_probe_exchange *xs[MAX_GPUS];
    float *bufs[MAX_GPUS];
    ncclComm_t comms[MAX_GPUS];
    int devs[MAX_GPUS];
    const uint64_t max_bytes = o->sweep_max_bytes;
    for (int d = 0; d < G; ++d) {
      devs[d] = gs[d].dev;
      CU(cudaSetDevice(gs[d].dev));
      PR(gpu_probe_exchange_create(&xs[d], d, G, max_bytes));
      bufs[d] = (float *)gpu_probe_exchange_buffer(xs[d]);
    }
    PR(gpu_probe_exchange_connect_local(xs, G));
    for (int d = 0; d < G; ++d) bufs[d] = (float *)gpu_probe_exchange_buffer(xs[d]);
    printf("PROBE INFO allreduce_group gpus=%d nvswitch_multicast=%d own_algo=%s\n", G,
           gpu_probe_exchange_has_multicast(xs[0]), o->allreduce_algo_name);
    if (o->allreduce_algo == GPU_PROBE_ALLREDUCE_MULTIMEM && !gpu_probe_exchange_has_multicast(xs[0])) {
      fprintf(stderr, "gpu-probe: --allreduce-algo=multimem but this group has no NVSwitch multicast\n");
      return RC_USAGE;
    }
    if (nc->so) NC(nc, nc->CommInitAll(comms, G, devs));

    /* Four columns on the same sizes.  NCCL three ways -- on the probe's own segment as it is
     * (plain: NCCL stages through its own buffers), on ncclMemAlloc'd memory registered with
     * ncclCommRegister (zero-copy / NVLS user buffers) and, where the library has it,
     * registered as a symmetric window (NCCL 2.27) -- and the probe's one-kernel allreduce (K4). */
    enum { I_NCCL, I_NCCL_REG, I_NCCL_SYM, I_OWN, I_COUNT };
    static const char *impl_name[I_COUNT] = {"nccl", "nccl_registered", "nccl_symmetric", "p2p_kernel"};
    for (int impl = nc->so ? I_NCCL : I_OWN; impl < I_COUNT; ++impl) {
      const int use_nccl = impl != I_OWN;
      float *nbufs[MAX_GPUS] = {0};
      void *regs[MAX_GPUS] = {0};
      ncclWindow_t wins[MAX_GPUS] = {0};
      float **use = bufs;
      if (impl == I_NCCL_REG || impl == I_NCCL_SYM) {
        int ok = nc->MemAlloc && nc->MemFree &&
                 (impl == I_NCCL_REG ? (nc->CommRegister && nc->CommDeregister) : (nc->CommWindowRegister && nc->CommWindowDeregister));
        for (int d = 0; d < G && ok; ++d) {
          CU(cudaSetDevice(gs[d].dev));
          if (nc->MemAlloc((void **)&nbufs[d], max_bytes) != ncclSuccess) ok = 0;
        }
        if (ok && impl == I_NCCL_REG)
          for (int d = 0; d < G && ok; ++d) {
            CU(cudaSetDevice(gs[d].dev));
            if (nc->CommRegister(comms[d], nbufs[d], max_bytes, &regs[d]) != ncclSuccess) ok = 0;
          }
        if (ok && impl == I_NCCL_SYM) { /* collective: every rank of the communicator registers, grouped in one process */
          ok = nc->GroupStart() == ncclSuccess;
          for (int d = 0; d < G && ok; ++d) {
            CU(cudaSetDevice(gs[d].dev));
            if (nc->CommWindowRegister(comms[d], nbufs[d], max_bytes, &wins[d], NCCL_WIN_COLL_SYMMETRIC) != ncclSuccess) ok = 0;
          }
          if (nc->GroupEnd() != ncclSuccess) ok = 0;
        }
        printf("PROBE INFO allreduce_buffers gpus=%d impl=%s available=%d\n", G, impl_name[impl], ok);
        if (!ok) { /* an older NCCL, or a fabric that refuses: the column is absent, not a failure */
          cudaGetLastError();
          for (int d = 0; d < G; ++d) {
            CU(cudaSetDevice(gs[d].dev));
            if (regs[d]) nc->CommDeregister(comms[d], regs[d]);
            if (wins[d]) nc->CommWindowDeregister(comms[d], wins[d]);
            if (nbufs[d]) nc->MemFree(nbufs[d]);
          }
          continue;
        }
        use = nbufs;
      }
      int ok = 0;
      if ((rc = check_allreduce(o, use_nccl, impl_name[impl], nc, comms, gs, xs, use, G, max_bytes, &ok))) return rc;
      for (int d = 0; d < G; ++d) { /* zeros: sums stay finite however often the prefix is reduced */
        CU(cudaSetDevice(gs[d].dev));
        CU(cudaMemsetAsync(use[d], 0, max_bytes, gs[d].st));
      }
      double best_bus = 0;
      uint64_t best_bytes = 0;
      for (uint64_t S = 1024; S <= max_bytes; S <<= 1) {
        const int iters = S >= (1u << 20) ? o->iters : 2 * o->iters;
        double ms = 0;
        if (!use_nccl && o->allreduce_algo == GPU_PROBE_ALLREDUCE_PUSH && S > (2u << 20)) break; /* one-shot limit */
        if ((rc = time_allreduce(o, use_nccl, nc, comms, gs, xs, use, G, S, iters, &ms))) return rc;
        const double algbw = (double)S / (ms * 1e-3) / 1e9;
        const double busbw = algbw * 2.0 * (G - 1) / G;
        if (busbw > best_bus) { best_bus = busbw; best_bytes = S; }
        printf("RESULT kind=allreduce gpus=%d gpu=all bytes=%" PRIu64 " ms=%.5f ms_med=%.5f gbs=%.2f "
               "frac_nominal=%.4f frac_measured=na lanes=na match=na us=%.2f algbw=%.2f busbw=%.2f impl=%s\n",
               G, S, ms, ms, busbw, busbw / NOMINAL_NVLINK_GBS, ms * 1e3, algbw, busbw, impl_name[impl]);
      }
      printf("RESULT kind=allreduce_peak gpus=%d gpu=all bytes=%" PRIu64 " ms=na ms_med=na gbs=%.2f frac_nominal=%.4f "
             "frac_measured=na lanes=na match=na impl=%s\n", G, best_bytes, best_bus, best_bus / NOMINAL_NVLINK_GBS,
             impl_name[impl]);
      if (use == nbufs) {
        if ((rc = sync_all(gs, G))) return rc;
        if (impl == I_NCCL_SYM) NC(nc, nc->GroupStart());
        for (int d = 0; d < G; ++d) {
          CU(cudaSetDevice(gs[d].dev));
          if (regs[d]) NC(nc, nc->CommDeregister(comms[d], regs[d]));
          if (wins[d]) NC(nc, nc->CommWindowDeregister(comms[d], wins[d]));
        }
        if (impl == I_NCCL_SYM) NC(nc, nc->GroupEnd());
        for (int d = 0; d < G; ++d) {
          CU(cudaSetDevice(gs[d].dev));
          NC(nc, nc->MemFree(nbufs[d]));
        }
      }
    }
    if (nc->so)
      for (int d = 0; d < G; ++d) NC(nc, nc->CommDestroy(comms[d]));
    for (int d = 0; d < G; ++d) PR(gpu_probe_exchange_destroy(xs[d]));
  }
  return 0;
}

/* ----------------------------------- phase D: P2P copies GPU0 -> GPUk, 0 <-> k */
static int phase_p2p(const options *o, gpu_state *gs, int ng) {
  int rc;
  uint64_t S = o->p2p_bytes;
  if (S > o->bytes) S = o->bytes;
  S &= ~(uint64_t)7;
  float *ms = (float *)malloc(sizeof(float) * (size_t)o->iters);
  if (!ms) return RC_CUDA;
  gpu_state *g0 = &gs[0];
  /* source pattern: GPU0's b[], GPUk's b[] (their own generator shards) */
  uint64_t src_lanes[2];
  CU(cudaSetDevice(g0->dev));
  PR(gpu_probe_checksum_u64((const uint64_t *)g0->b, S / 8, 0, g0->lanes, g0->ws, o->variant, g0->st));
  CU(cudaStreamSynchronize(g0->st));
  CU(cudaMemcpy(src_lanes, g0->lanes, 16, cudaMemcpyDeviceToHost));

  for (int k = 1; k < ng; ++k) {
    gpu_state *gk = &gs[k];
    int can01 = 0, can10 = 0;
    CU(cudaDeviceCanAccessPeer(&can01, g0->dev, gk->dev));
    CU(cudaDeviceCanAccessPeer(&can10, gk->dev, g0->dev));
    if (!can01 || !can10) {
      printf("RESULT kind=p2p gpus=%d gpu=%d->%d bytes=%" PRIu64 " ms=na ms_med=na gbs=na frac_nominal=na "
             "frac_measured=na lanes=na match=na mode=unsupported\n", ng, g0->dev, gk->dev, S);
      continue;
    }
    cudaError_t e;
    CU(cudaSetDevice(g0->dev));
    e = cudaDeviceEnablePeerAccess(gk->dev, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
    CU(e);
    CU(cudaSetDevice(gk->dev));
    e = cudaDeviceEnablePeerAccess(g0->dev, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
    CU(e);

    /* 0: K3 launched on the SOURCE GPU, storing into the peer (push); 1: K3 launched on the
     * DESTINATION GPU, loading from the peer (pull, one-directional only); 2: the copy engine.
     * The same bytes cross the same link either way, but not the same packets: SM stores travel as
     * 128-byte writes with 24 bytes of protocol each (1 GiB of payload = 1.19 GiB on the wire,
     * 714 GB/s of payload on a link that is then 94 % busy), read responses carry 16 bytes per 128
     * (786 GB/s) at the price of 32-byte requests in the opposite direction -- which is why two
     * GPUs copying to each other at once are better off both pushing (profiles/nvlink_packets_r2.md). */
    for (int mode = 0; mode < 3; ++mode) {
      for (int bidir = 0; bidir < 2; ++bidir) {
        if (mode == 1 && bidir) continue;
        /* uni: GPU0's b[] lands in GPUk's a[].  bi: at the same time GPUk's b[] lands in GPU0's a[]. */
        gpu_state *drv0 = mode == 1 ? gk : g0; /* the GPU whose stream carries the 0 -> k copy */
        for (int i = -o->warmup; i < o->iters; ++i) {
          CU(cudaSetDevice(drv0->dev));
          if (i >= 0) CU(cudaEventRecord(drv0->ev[2 * i], drv0->st));
          if (mode < 2) PR(gpu_probe_p2p_copy(gk->a, g0->b, S, drv0->st));
          else CU(cudaMemcpyPeerAsync(gk->a, gk->dev, g0->b, g0->dev, S, g0->st));
          if (i >= 0) CU(cudaEventRecord(drv0->ev[2 * i + 1], drv0->st));
          if (bidir) {
            CU(cudaSetDevice(gk->dev));
            if (i >= 0) CU(cudaEventRecord(gk->ev[2 * i], gk->st));
            if (mode == 0) PR(gpu_probe_p2p_copy(g0->a, gk->b, S, gk->st));
            else CU(cudaMemcpyPeerAsync(g0->a, g0->dev, gk->b, gk->dev, S, gk->st));
            if (i >= 0) CU(cudaEventRecord(gk->ev[2 * i + 1], gk->st));
          }
        }
        if ((rc = sync_all(gs, ng))) return rc;
        for (int i = 0; i < o->iters; ++i) {
          float t0, t1 = 0;
          CU(cudaSetDevice(drv0->dev));
          CU(cudaEventElapsedTime(&t0, drv0->ev[2 * i], drv0->ev[2 * i + 1]));
          if (bidir) {
            CU(cudaSetDevice(gk->dev));
            CU(cudaEventElapsedTime(&t1, gk->ev[2 * i], gk->ev[2 * i + 1]));
          }
          ms[i] = t0 > t1 ? t0 : t1;
        }
        timing t = summarize(ms, o->iters);
        /* verify the destination with K2 on the destination GPU */
        uint64_t got[2];
        CU(cudaSetDevice(gk->dev));
        PR(gpu_probe_checksum_u64((const uint64_t *)gk->a, S / 8, 0, gk->lanes, gk->ws, o->variant, gk->st));
        CU(cudaStreamSynchronize(gk->st));
        CU(cudaMemcpy(got, gk->lanes, 16, cudaMemcpyDeviceToHost));
        const int ok = got[0] == src_lanes[0] && got[1] == src_lanes[1];
        if (!ok) fail(RC_PARITY);
        const uint64_t bytes = bidir ? 2 * S : S;
        const double gbs = (double)bytes / (t.best_ms * 1e-3) / 1e9;
        printf("RESULT kind=p2p gpus=%d gpu=%d%s%d bytes=%" PRIu64 " ms=%.4f ms_med=%.4f gbs=%.1f frac_nominal=%.4f "
               "frac_measured=na lanes=%016" PRIx64 ",%016" PRIx64 " match=%d mode=%s dir=%s\n",
               ng, g0->dev, bidir ? "<->" : "->", gk->dev, bytes, t.best_ms, t.med_ms, gbs,
               gbs / (NOMINAL_NVLINK_GBS * (bidir ? 2 : 1)), got[0], got[1], ok,
               mode == 0 ? "sm_kernel" : mode == 1 ? "sm_kernel_pull" : "copy_engine", bidir ? "bi" : "uni");
        /* the next pass must not pass on stale data */
        CU(cudaMemsetAsync(gk->a, 0, S, gk->st));
        CU(cudaStreamSynchronize(gk->st));
      }
    }
  }
  free(ms);
  return 0;
}

/* -------------------------------------------------------------------- main */
static int run(const options *o) {
  int count = 0, rc;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess) {
    fprintf(stderr, "gpu-probe: cudaGetDeviceCount: %s -- no usable GPU; this program has no CPU path\n",
            cudaGetErrorString(e));
    return o->expect_gpus > 0 ? RC_COUNT : RC_CUDA;
  }
  printf("PROBE INFO visible_gpus=%d expect_gpus=%d abi=%d %s\n", count, o->expect_gpus, gpu_probe_abi_version(),
         gpu_probe_kernel_info());
  if (o->expect_gpus >= 0 && count != o->expect_gpus) {
    fprintf(stderr, "gpu-probe: the pod was promised %d GPU(s) (nvidia.com/gpu limit) but CUDA sees %d\n",
            o->expect_gpus, count);
    return RC_COUNT;
  }
  int ng = count;
  if (o->max_gpus > 0 && ng > o->max_gpus) ng = o->max_gpus;
  if (ng > MAX_GPUS) ng = MAX_GPUS;
  if (ng < 1) return RC_COUNT;

  const uint64_t n = o->bytes / 8;
  gpu_state gs[MAX_GPUS];
  for (int d = 0; d < ng; ++d) {
    struct cudaDeviceProp p;
    CU(cudaGetDeviceProperties(&p, d));
    printf("PROBE GPU index=%d name=\"%s\" sm=%d.%d sms=%d mem_gib=%.1f pci=%04x:%02x:%02x\n", d, p.name, p.major,
           p.minor, p.multiProcessorCount, (double)p.totalGlobalMem / (1ull << 30), p.pciDomainID, p.pciBusID,
           p.pciDeviceID);
    if (p.major != 10) {
      fprintf(stderr, "gpu-probe: GPU %d is sm_%d%d; the kernels are built for sm_100a only\n", d, p.major, p.minor);
      return RC_CUDA;
    }
    if ((rc = gpu_setup(&gs[d], d, n, 2 * (o->iters > 1 ? o->iters : 1)))) return rc;
  }

  if ((rc = phase_hbm(o, gs, ng, n))) return rc;

  if (ng > 1) {
    nccl_api nc;
    ncclComm_t comms[MAX_GPUS];
    const int have_nccl = nccl_load(&nc) == 0;
    if (!have_nccl) fail(RC_NCCL);
    if (o->do_sharded) {
      gpu_probe_exchange *xs[MAX_GPUS];
      for (int d = 0; d < ng; ++d) {
        CU(cudaSetDevice(gs[d].dev));
        PR(gpu_probe_exchange_create(&gs[d].xchg, d, ng, 0));
        xs[d] = gs[d].xchg;
      }
      PR(gpu_probe_exchange_connect_local(xs, ng));
      if (have_nccl) {
        int v = 0, devs[MAX_GPUS];
        for (int d = 0; d < ng; ++d) devs[d] = gs[d].dev;
        NC(&nc, nc.GetVersion(&v));
        printf("PROBE INFO nccl_version=%d\n", v);
        NC(&nc, nc.CommInitAll(comms, ng, devs));
      }
      if ((rc = phase_sharded(o, gs, ng, n, &nc, comms))) return rc;
      if (have_nccl)
        for (int d = 0; d < ng; ++d) NC(&nc, nc.CommDestroy(comms[d]));
    }
    if (o->do_sweep && (rc = phase_sweep(o, gs, ng, &nc))) return rc;
    if (o->do_p2p && (rc = phase_p2p(o, gs, ng))) return rc;
  }
  return 0;
}

int main(int argc, char **argv) {
  options o;
  setvbuf(stdout, NULL, _IOLBF, 0);
  const double t0 = now_s();
  parse_options(argc, argv, &o);
  printf("PROBE START t_unix=%.6f bytes_per_array=%" PRIu64 " iters=%d warmup=%d variant=%s\n", t0, o.bytes, o.iters,
         o.warmup, o.variant_name);
  int rc = run(&o);
  if (rc == RC_OK) rc = g_status;
  const double t1 = now_s();
  printf("PROBE %s rc=%d t_unix=%.6f elapsed_s=%.3f\n", rc == RC_OK ? "PASS" : "FAIL", rc, t1, t1 - t0);
  return rc;
}
