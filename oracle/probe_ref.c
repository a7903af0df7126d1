/*
 * oracle/probe_ref.c -- CPU restatement of the probe workload (see probe_ref.h).
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT; PARITY UNPINNED against the reference
 * (the reference has no arithmetic: /root/reference/pods/nvidia-gpu-test-pod.yaml:9
 * is `echo ... && sleep 3600`).  Follows SURVEY.md section 8(d).
 *
 * Build: gcc -O3 -march=x86-64-v3 -ffp-contract=off -fopenmp  (oracle/Makefile).
 *   -ffp-contract=off  : nothing but the explicit fma() below may fuse, so the
 *                        triad rounds exactly once and the unfused KAT rounds twice.
 *   -march=x86-64-v3   : AVX2+FMA, so fma() is one vfmadd and the built .so
 *                        still runs on a GPU box whose host CPU is not this one.
 */
#include "probe_ref.h"

#include <math.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define GOLDEN 0x9E3779B97F4A7C15ull
#define MIX1 0xBF58476D1CE4E5B9ull
#define MIX2 0x94D049BB133111EBull

static inline uint64_t mix(uint64_t z) {
  z = (z ^ (z >> 30)) * MIX1;
  z = (z ^ (z >> 27)) * MIX2;
  return z ^ (z >> 31);
}

uint64_t ref_gen(uint64_t i, uint64_t seed) { return mix(seed + (i + 1) * GOLDEN); }

static inline double unit_double(uint64_t x) {
  /* top 53 bits -> [0,1); both steps exact in binary64 */
  return (double)(x >> 11) * 0x1.0p-53;
}

void ref_init_fill(double *dst, uint64_t n, uint64_t first_index, uint64_t seed) {
#pragma omp parallel for schedule(static)
  for (uint64_t k = 0; k < n; ++k) dst[k] = unit_double(ref_gen(first_index + k, seed));
}

void ref_triad(double *a, const double *b, const double *c, double s, uint64_t n) {
#pragma omp parallel for schedule(static)
  for (uint64_t k = 0; k < n; ++k) a[k] = fma(s, c[k], b[k]);
}

void ref_triad_unfused(double *a, const double *b, const double *c, double s,
                       uint64_t n) {
  for (uint64_t k = 0; k < n; ++k) {
    volatile double p = s * c[k]; /* force the intermediate rounding */
    a[k] = b[k] + p;
  }
}

void ref_checksum_u64(const uint64_t *w, uint64_t n, uint64_t first_index,
                      uint64_t lanes[2]) {
  uint64_t l0 = 0, l1 = 0;
#pragma omp parallel for schedule(static) reduction(+ : l0, l1)
  for (uint64_t k = 0; k < n; ++k) {
    uint64_t v = w[k];
    l0 += v;
    l1 += v * (2 * (first_index + k) + 1);
  }
  lanes[0] = l0;
  lanes[1] = l1;
}

void ref_stream_lanes(uint64_t n, uint64_t first_index, uint64_t seed_b,
                      uint64_t seed_c, double s, uint64_t lanes[2]) {
  uint64_t l0 = 0, l1 = 0;
#pragma omp parallel for schedule(static) reduction(+ : l0, l1)
  for (uint64_t k = 0; k < n; ++k) {
    uint64_t i = first_index + k;
    double b = unit_double(ref_gen(i, seed_b));
    double c = unit_double(ref_gen(i, seed_c));
    double a = fma(s, c, b);
    uint64_t v;
    memcpy(&v, &a, sizeof v);
    l0 += v;
    l1 += v * (2 * i + 1);
  }
  lanes[0] = l0;
  lanes[1] = l1;
}

int ref_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void ref_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}
