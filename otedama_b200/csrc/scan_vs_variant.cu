// scan_vs_variant.cu — one version-shared scan variant per translation unit: compile
// with -DOTD_VSIDX=<index into OTD_VS_VARIANTS>.
#include "scan_vs_kernel.cuh"
#include "scan_variants.h"

#ifndef OTD_VSIDX
#error "compile with -DOTD_VSIDX=<variant index>"
#endif

namespace {
struct Sel {
    int iter;
    unsigned kn;
    int minb;
    int block;
};
constexpr Sel select(int want)
{
#define OTD_PICK(idx, name, iter, kn, minb, block) \
    if (want == idx) return Sel{iter, kn, minb, block};
    OTD_VS_VARIANTS(OTD_PICK)
#undef OTD_PICK
    return Sel{4, 0u, 1, 256};
}
constexpr Sel kSel = select(OTD_VSIDX);
}  // namespace

#define OTD_FN_NAME_(i) otd_vs_launch_##i
#define OTD_FN_NAME(i) OTD_FN_NAME_(i)

cudaError_t OTD_FN_NAME(OTD_VSIDX)(const otd::VsParams& p, uint32_t n_ver, uint32_t nonce_begin, unsigned long long count,
                                   uint32_t instance0, otd::RingHeader* hdr, otd::RingEntry* ent, uint32_t cap,
                                   const unsigned int* stop, unsigned int gen, cudaStream_t st)
{
    constexpr unsigned long long tile = (unsigned long long)kSel.block * kSel.iter;
    const unsigned grid = (unsigned)((count + tile - 1) / tile);
    if (grid == 0 || n_ver == 0) return cudaSuccess;
    if (n_ver > (uint32_t)otd::kVsMaxVersions) return cudaErrorInvalidValue;
    auto kern = otd::scan_vs_kernel<kSel.iter, kSel.kn, kSel.minb, kSel.block>;
    // 46 words x BLOCK threads of private scratch per block: ask for the largest shared-memory carve-out so that
    // MINB blocks stay resident (the kernel makes no use of L1)
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    kern<<<grid, kSel.block, 0, st>>>(p, otd::make_pipe_mults(), n_ver, nonce_begin, count, instance0, hdr, ent, cap, stop, gen);
    return cudaGetLastError();
}
