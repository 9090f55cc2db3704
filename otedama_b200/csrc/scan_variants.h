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
struct VsParams;
}  // namespace otd

// index, name, V, ITER, knobs, HALF, MINB, BLOCK.  Production variants first (indices below
// OTD_NUM_PRODUCTION_VARIANTS); the rest are performance experiments: "h_*" variants
// stop after hash 1 (HALF) and do NOT compute SHA-256d — the library refuses to run
// them unless the environment says OTD_ALLOW_EXPERIMENTAL=1.
//
// Measured on B200 (profiles/r1_variant_sweep.log), GH/s per GPU over a full 2^32 scan:
//   add_z6_b4 9.55 | add_sp_b5 9.32 | add_t1x_b5 9.25 | c_add_t1x_b4 9.0-9.1 | alu_b5 8.25 | c_alu_b4 8.16
// Block shape at the same warps/SM: 256 = 512 threads > 1024 > 128; ITER 8..64: no effect.
#define OTD_KN_ADD1 (otd::KN_E | otd::KN_W1)               /* 2-input adds: one IMAD replaces one ALU add */
#define OTD_KN_ADDS (OTD_KN_ADD1 | otd::KN_A | otd::KN_W2) /* every immediate-free add except t1's */
#define OTD_KN_T1X (OTD_KN_ADDS | otd::KN_T1X)             /* ... plus t1 split without an immediate IMAD */
#define OTD_KN_SPX (OTD_KN_T1X | otd::KN_SP)               /* ... plus the reg+reg adds of the specialised parts */
#define OTD_KN_BEST (OTD_KN_ADDS | otd::KN_SP | otd::KN_T1Z | OTD_KN_Z_EIGHTHS(6)) /* ... and no ALU add at all in 6/8 of the rounds */
#define OTD_SCAN_VARIANTS(X)                                                                        \
    X(0, "add_z6_b4", 1, 16, OTD_KN_BEST, false, 4, 256)                                            \
    X(1, "add_sp_b5", 1, 16, OTD_KN_SPX, false, 5, 256)                                             \
    X(2, "alu_b5", 1, 16, 0u, false, 5, 256)                                                        \
    X(3, "c_add_t1x_b4", 1, 16, otd::KN_COMPACT | OTD_KN_T1X, false, 4, 256)                        \
    X(4, "c_alu_b4", 1, 16, otd::KN_COMPACT, false, 4, 256)                                         \
    X(5, "add_z6_v2_b2", 2, 8, OTD_KN_BEST, false, 2, 256)                                          \
    X(6, "h_alu", 1, 16, 0u, true, 5, 256)                                                          \
    X(7, "h_add_sp", 1, 16, OTD_KN_SPX, true, 5, 256)                                               \
    X(8, "x_z6_kc_b4", 1, 16, OTD_KN_BEST | otd::KN_KCONST, false, 4, 256)                          \
    X(9, "x_z6_t7z_b4", 1, 16, OTD_KN_BEST | otd::KN_T7Z, false, 4, 256)                            \
    X(10, "x_z6_kc_t7z_b4", 1, 16, OTD_KN_BEST | otd::KN_KCONST | otd::KN_T7Z, false, 4, 256)                      \
    X(11, "x_z6_s3", 1, 16, OTD_KN_BEST | OTD_KN_ZSEL(3), false, 4, 256)                            \
    X(12, "x_z6_late_b4", 1, 16, OTD_KN_BEST | otd::KN_LATE, false, 4, 256)                          \
    X(13, "x_z6_late_s3_b4", 1, 16, OTD_KN_BEST | otd::KN_LATE | OTD_KN_ZSEL(3), false, 4, 256)      \
    X(14, "x_z6_late_t7z_b4", 1, 16, OTD_KN_BEST | otd::KN_LATE | otd::KN_T7Z, false, 4, 256)

#define OTD_NUM_PRODUCTION_VARIANTS 6
#define OTD_DEFAULT_VARIANT 0

// Version-shared scan variants (scan_vs_kernel.cuh): index, name, ITER, knobs, MINB, BLOCK.  All of them compute
// the full SHA-256d.  OTD_KN_VS(n, m): the shipped add placement with n/8 of the immediate-constant rounds and
// m/8 of the shared-schedule rounds built from IMADs only (m = 0 reads as 8/8).
#define OTD_KN_VS(n, m) (OTD_KN_ADDS | otd::KN_SP | otd::KN_T1Z | OTD_KN_Z_EIGHTHS(n) | OTD_KN_VSZ_EIGHTHS(m))
// Measured on B200 (profiles/r2_vs_sweep_*.log), GH/s per GPU, 32 version instances x 2^27 nonces per launch
// (the plain kernel launched once per version: 9.52):
//   vs_z7_r8_s5_b3 11.455 (shipped: z7_r8_b3 with another choice of WHICH eighth of the immediate-constant rounds keeps
//   its IADD3 — KN_ZSEL 5; the other eleven choices 11.27-11.37, profiles/r2_vs_zsel*.log: an order lottery, see
//   DESIGN.md §3.2 point 5) | vs_z7_r8_b3 11.365 | z6_r8_b3 11.364 | z7_r7_b3 11.35 | z8_r8_b3 11.33 | z7_r8_b4 11.33 | z7_r7_b4 11.21 |
//   z6_r8_b4 11.19 | f20_b5 (5 blocks/SM) 11.19 | z6_r6_b4 11.11 | z7_r5_b4 11.07 | z8_r7_b4 11.02 |
//   2 blocks/SM (b2) 9.98 | 128-thread blocks x 6 10.88, x 5 9.71 | 192 x 6 10.55 | ITER 2 / 8 11.16 / 11.19.
// b3 = __launch_bounds__(256, 3): a register cap of 85, of which ptxas uses 56 — four blocks per SM stay resident
// (ncu: 49.5 % of the warp slots), the looser cap only buys a better schedule than b4's 64 (48 used); b5 buys
// nothing, b2 (more registers, fewer resident blocks) loses 12 %.  16 instances per launch instead of 32: 11.10 (z7_r7_b4).
// Around the shipped variant (profiles/r2_vs_lottery3.log): ITER 1 / 2 / 8 10.15 / 11.18 / 11.44, 44 words (f20) under
// b3 / b5 11.00 / 11.22, constant adds as IMADs 11.44 — nothing better.
// 64 instances per launch instead of 32 (shipped since; profiles/r2_vs_v64.log): 11.490 (s5), 11.434 (z7_r8_b3).
// -maxrregcount 50..64 changes nothing (launch bounds decide; profiles/r2_vs_maxrreg.log).
// ptxas -regUsageLevel 0..4 = the default build (11.455), 6..10 11.24 (profiles/r2_vs_ptxas_ru.log).
// Per-instance constants added as IMADs (cf, 14 fewer ALU instructions per hash): 11.355 — no gain, like every other
// IMAD that needs a vector-register 1 (profiles/r2_vs_cfma.log).
// Hash-1 rounds 24..63 rolled into an 8-round loop body (l1, 39 KB of code instead of 47): 10.96-11.05 — the
// instruction cache is not what holds the kernel back (ncu: hit rate 93 %, no_instruction 0.3 of 10.6 cycles per issue).
#define OTD_VS_VARIANTS(Y)                                  \
    Y(0, "vs_z7_r8_s5_b3", 4, OTD_KN_VS(7, 0) | OTD_KN_ZSEL(5), 3, 256) \
    Y(1, "vs_z7_r8_b3", 4, OTD_KN_VS(7, 0), 3, 256)         \
    Y(2, "vs_z6_r8_b3", 4, OTD_KN_VS(6, 0), 3, 256)         \
    Y(3, "vs_z7_r7_f20_b5", 4, OTD_KN_VS(7, 7) | otd::KN_VS_F20, 5, 256) \
    Y(4, "vs_z7_r8_l1_b4", 4, OTD_KN_VS(7, 0) | otd::KN_VS_LOOP1, 4, 256) \
    Y(5, "vs_z7_r8_cf_b3", 4, OTD_KN_VS(7, 0) | otd::KN_VS_CFMA, 3, 256) \
    Y(6, "vs_z7_r7_b4", 4, OTD_KN_VS(7, 7), 4, 256)
#define OTD_DEFAULT_VS_VARIANT 0

typedef cudaError_t (*otd_scan_launch_fn)(const otd::JobConsts&, uint32_t nonce_begin,
                                          unsigned long long count, uint32_t instance,
                                          otd::RingHeader*, otd::RingEntry*, uint32_t cap,
                                          const unsigned int* stop, unsigned int gen, cudaStream_t);

#define OTD_DECLARE_VARIANT(idx, name, v, iter, kn, half, minb, block)                                                   \
    cudaError_t otd_scan_launch_##idx(const otd::JobConsts&, uint32_t, unsigned long long, uint32_t, \
                                      otd::RingHeader*, otd::RingEntry*, uint32_t, const unsigned int*, \
                                      unsigned int, cudaStream_t);
OTD_SCAN_VARIANTS(OTD_DECLARE_VARIANT)
#undef OTD_DECLARE_VARIANT

// Version-shared launch: n_ver instances (p.jc[0..n_ver)) x the nonce window, instance index of p.jc[0] = instance0.
typedef cudaError_t (*otd_vs_launch_fn)(const otd::VsParams&, uint32_t n_ver, uint32_t nonce_begin,
                                        unsigned long long count, uint32_t instance0, otd::RingHeader*,
                                        otd::RingEntry*, uint32_t cap, const unsigned int* stop, unsigned int gen,
                                        cudaStream_t);
#define OTD_DECLARE_VS_VARIANT(idx, name, iter, kn, minb, block)                                              \
    cudaError_t otd_vs_launch_##idx(const otd::VsParams&, uint32_t, uint32_t, unsigned long long, uint32_t, \
                                    otd::RingHeader*, otd::RingEntry*, uint32_t, const unsigned int*,       \
                                    unsigned int, cudaStream_t);
OTD_VS_VARIANTS(OTD_DECLARE_VS_VARIANT)
#undef OTD_DECLARE_VS_VARIANT
