/*
 * oracle/probe_ref.h -- CPU restatement of the probe workload's arithmetic.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may build, load or call
 * anything in oracle/.  The shipped probe (kind-gpu-sim_b200/) never links it
 * and has no CPU fallback.
 *
 * PARITY UNPINNED against the reference: maryamtahhan/kind-gpu-sim holds no
 * arithmetic at all -- the slot this workload fills is
 *   /root/reference/pods/nvidia-gpu-test-pod.yaml:9   (echo ... && sleep 3600)
 * and the reference has no golden vectors, KATs or fixtures for it
 * (SURVEY.md section 8c).  The restatement therefore follows the spec in
 * SURVEY.md section 8(d) / BASELINE.md section 3, and is pinned instead to
 *   - the published SplitMix64 known-answer vector (seed 1234567), and
 *   - an independent pure-Python big-integer restatement (tests/pyspec.py),
 * both checked by tests/test_oracle.py.
 *
 * Argument meaning mirrors include/gpu_probe.h one to one so the parity tests
 * can diff the two sides call by call.
 */
#ifndef PROBE_REF_H
#define PROBE_REF_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Spec constants. */
#define PROBE_REF_SEED_B 0x0123456789ABCDEFull
#define PROBE_REF_SEED_C 0xFEDCBA9876543210ull
#define PROBE_REF_SCALAR 3.0

/* Counter-based SplitMix64: element i of the stream seeded with `seed`
 * (i is the GLOBAL element index; gen(i) == the (i+1)-th output of the
 * sequential SplitMix64 generator started at state `seed`). */
uint64_t ref_gen(uint64_t i, uint64_t seed);

/* dst[k] = (double)(gen(first_index + k, seed) >> 11) * 2^-53, k in [0, n). */
void ref_init_fill(double *dst, uint64_t n, uint64_t first_index, uint64_t seed);

/* a[k] = fma(s, c[k], b[k]) -- exactly one rounding per element. */
void ref_triad(double *a, const double *b, const double *c, double s, uint64_t n);

/* lanes[0] = sum w[k]; lanes[1] = sum w[k] * (2*(first_index+k)+1); mod 2^64. */
void ref_checksum_u64(const uint64_t *w, uint64_t n, uint64_t first_index,
                      uint64_t lanes[2]);

/* Streaming mode: generate -> fma -> accumulate with no arrays (any n).
 * Equals ref_checksum_u64 over ref_triad over two ref_init_fill arrays. */
void ref_stream_lanes(uint64_t n, uint64_t first_index, uint64_t seed_b,
                      uint64_t seed_c, double s, uint64_t lanes[2]);

/* Unfused variant a[k] = b[k] + s*c[k] (two roundings).  Exists only so the
 * FMA-discriminating KAT can show the two differ. */
void ref_triad_unfused(double *a, const double *b, const double *c, double s,
                       uint64_t n);

/* OpenMP thread control for the timed CPU baseline. */
int ref_max_threads(void);
void ref_set_threads(int n);

#ifdef __cplusplus
}
#endif
#endif /* PROBE_REF_H */
