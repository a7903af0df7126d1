// probe_launch.h -- launchers shared between the translation units of libgpuprobe.so.
// Internal: not part of the C ABI (include/gpu_probe.h), hidden from the dynamic symbol table.
#pragma once

#include <cstdint>

#include "probe_common.cuh"

namespace probe {

#define PROBE_INTERNAL __attribute__((visibility("hidden")))

// K2 and fused K1+K2 with an optional lanes exchange attached (x.world == 0: none).
PROBE_INTERNAL int checked_checksum(const uint64_t *w, uint64_t n, uint64_t first_index, uint64_t *out_lanes,
                                    void *workspace, int variant, void *stream, const ExchangeArgs &x);
PROBE_INTERNAL int checked_triad_checksum(double *a, const double *b, const double *c, double s, uint64_t n,
                                          uint64_t first_index, uint64_t *out_lanes, void *workspace, int variant,
                                          void *stream, const ExchangeArgs &x);

// Single-GPU group self-test: every rank's checksum + lanes exchange in one launch.
struct ChecksumGroupArgs {
  const uint64_t *w[8];
  uint64_t n[8], first[8];
  uint64_t *out[8], *ws[8];
  ExchangeArgs x[8];
};
PROBE_INTERNAL int checked_checksum_group(const ChecksumGroupArgs &g, int world, void *stream);

}  // namespace probe
