// scan_variant.cu — one scan-kernel variant per translation unit: compile with
// -DOTD_VIDX=<index into OTD_SCAN_VARIANTS>.
#include <cstdlib>

#include "scan_kernel.cuh"
#include "scan_variants.h"

#ifndef OTD_VIDX
#error "compile with -DOTD_VIDX=<variant index>"
#endif

namespace {
template <int V, int ITER, unsigned KN, bool HALF, int MINB, int BLOCK>
cudaError_t launch(const otd::JobConsts& jc, uint32_t nonce_begin, unsigned long long count, uint32_t instance,
                   otd::RingHeader* hdr, otd::RingEntry* ent, uint32_t cap, const unsigned int* stop,
                   unsigned int gen, cudaStream_t st)
{
    constexpr unsigned long long tile = (unsigned long long)BLOCK * V * ITER;
    const unsigned grid = (unsigned)((count + tile - 1) / tile);
    if (grid == 0) return cudaSuccess;
    // Occupancy experiment (performance studies only): OTD_K1_DYNSMEM=<bytes> of unused dynamic shared memory per
    // block lowers the number of resident blocks per SM without touching the code (DESIGN.md §3.2 point 4).
    static const int dyn = [] {
        const char* e = getenv("OTD_K1_DYNSMEM");
        return e ? atoi(e) : 0;
    }();
    if (dyn > 0) {
        cudaError_t e = cudaFuncSetAttribute(otd::scan_kernel<V, ITER, KN, HALF, MINB, BLOCK>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, dyn);
        if (e != cudaSuccess) return e;
    }
    otd::scan_kernel<V, ITER, KN, HALF, MINB, BLOCK><<<grid, BLOCK, dyn, st>>>(jc, otd::make_pipe_mults(), nonce_begin, count,
                                                                      instance, hdr, ent, cap, stop, gen);
    return cudaGetLastError();
}
}  // namespace

// The selected variant's parameters, picked out of the list at compile time.
namespace {
struct Sel {
    int v, iter;
    unsigned kn;
    bool half;
    int minb;
    int block;
};
constexpr Sel select(int want)
{
#define OTD_PICK(idx, name, v, iter, kn, half, minb, block) \
    if (want == idx) return Sel{v, iter, kn, half, minb, block};
    OTD_SCAN_VARIANTS(OTD_PICK)
#undef OTD_PICK
    return Sel{1, 16, 0u, false, 1, 256};
}
constexpr Sel kSel = select(OTD_VIDX);
}  // namespace

#define OTD_FN_NAME_(i) otd_scan_launch_##i
#define OTD_FN_NAME(i) OTD_FN_NAME_(i)

cudaError_t OTD_FN_NAME(OTD_VIDX)(const otd::JobConsts& jc, uint32_t nonce_begin, unsigned long long count,
                                  uint32_t instance, otd::RingHeader* hdr, otd::RingEntry* ent, uint32_t cap,
                                  const unsigned int* stop, unsigned int gen, cudaStream_t st)
{
    return launch<kSel.v, kSel.iter, kSel.kn, kSel.half, kSel.minb, kSel.block>(jc, nonce_begin, count, instance, hdr, ent, cap, stop, gen, st);
}
