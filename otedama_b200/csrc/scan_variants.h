// scan_variants.h — the list of compiled scan-kernel variants (X-macro) and the
// launch-function table type.  Each variant is one (V, ITER, knobs) instantiation
// of scan_kernel, compiled in its own translation unit (scan_variant.cu with
// -DOTD_VIDX=<index>) so the build parallelises.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "scan_core.cuh"

namespace otd {
struct RingHeader;
struct RingEntry;
}  // namespace otd

// index, name, V, ITER, knobs
#define OTD_KN_FREE (otd::KN_E | otd::KN_W1 | otd::KN_SSHR)
#define OTD_SCAN_VARIANTS(X)                                                                        \
    X(0, "alu", 1, 16, 0u)                                                                          \
    X(1, "free", 1, 16, OTD_KN_FREE)                                                                \
    X(2, "hkw_bs1", 1, 16, OTD_KN_FREE | otd::KN_HKW | otd::KN_BS1)                                 \
    X(3, "bal", 1, 16, OTD_KN_FREE | otd::KN_HKW | otd::KN_BS1 | otd::KN_BS0 | otd::KN_SROT)        \
    X(4, "bal_nosrot", 1, 16, OTD_KN_FREE | otd::KN_HKW | otd::KN_BS1 | otd::KN_BS0)                \
    X(5, "bal_nobs0", 1, 16, OTD_KN_FREE | otd::KN_HKW | otd::KN_BS1 | otd::KN_SROT)                \
    X(6, "bal_nohkw", 1, 16, OTD_KN_FREE | otd::KN_BS1 | otd::KN_BS0 | otd::KN_SROT)                \
    X(7, "bal_w2", 1, 16, OTD_KN_FREE | otd::KN_HKW | otd::KN_BS1 | otd::KN_BS0 | otd::KN_SROT | otd::KN_W2) \
    X(8, "allfma", 1, 16, 1023u)                                                                    \
    X(9, "bal_v2", 2, 8, OTD_KN_FREE | otd::KN_HKW | otd::KN_BS1 | otd::KN_BS0 | otd::KN_SROT)      \
    X(10, "alu_v2", 2, 8, 0u)

#define OTD_DEFAULT_VARIANT 0

typedef cudaError_t (*otd_scan_launch_fn)(const otd::JobConsts&, uint32_t nonce_begin,
                                          unsigned long long count, uint32_t instance,
                                          otd::RingHeader*, otd::RingEntry*, uint32_t cap,
                                          const unsigned int* stop, unsigned int gen, cudaStream_t);

#define OTD_DECLARE_VARIANT(idx, name, v, iter, kn)                                                   \
    cudaError_t otd_scan_launch_##idx(const otd::JobConsts&, uint32_t, unsigned long long, uint32_t, \
                                      otd::RingHeader*, otd::RingEntry*, uint32_t, const unsigned int*, \
                                      unsigned int, cudaStream_t);
OTD_SCAN_VARIANTS(OTD_DECLARE_VARIANT)
#undef OTD_DECLARE_VARIANT
