// scan_vs_kernel.cuh — K1v, the version-shared SHA-256d scan kernel for sm_100a:
// the nonce x BIP-320-version search of north_star (BASELINE configs[3], [4]).
//
// A thread takes ITER nonces; for each it expands the block-2 message schedule of
// hash 1 once (scan_vs_core.cuh: 46 words of K+W, parked in this thread's column
// of shared memory) and then walks the launch's n_ver version instances in a loop:
// block-2 state recurrence from that instance's midstate, hash 2, early reject on
// the final state word — the same test, slow path and warp-ballot ring as
// scan_kernel.cuh.  The reference's loop has no counterpart (one header, nonce
// only: internal/miner/worker.go:216-282).
//
// Bound: integer ALU issue, like K1.  Shared memory is private scratch (each
// thread reads back only what it wrote: no barrier), 46 (or 44) x BLOCK words per block;
// the instances' constants arrive as a __grid_constant__ array and are read
// through the constant bank with the (warp-uniform) instance index.
#pragma once
#include "scan_kernel.cuh"
#include "scan_vs_core.cuh"

namespace otd {

constexpr int kVsMaxVersions = 64;  // instances per launch: 64 x 172 B of kernel parameters

struct VsParams {
    JobConsts jc[kVsMaxVersions];
};

#if defined(__CUDACC__)

template <int ITER, unsigned KN, int MINB, int BLOCK>
__global__ void __launch_bounds__(BLOCK, MINB)
scan_vs_kernel(const __grid_constant__ VsParams p, const __grid_constant__ PipeMults pm, uint32_t n_ver,
               uint32_t nonce_begin, unsigned long long count, uint32_t instance0, RingHeader* ring,
               RingEntry* entries, uint32_t cap, const unsigned int* stop_gen, unsigned int my_gen)
{
    if (*reinterpret_cast<const volatile unsigned int*>(stop_gen) != my_gen) return;

    constexpr int kSkip = vs_first_round<KN>() - 18;  // leading words carried in registers instead
    __shared__ uint32_t kw_smem[vs_words<KN>() * BLOCK];  // word i of this thread: kw_smem[i * BLOCK + threadIdx.x]
    uint32_t* const col = kw_smem + threadIdx.x;

    VsOps<KN> op{{pm}};
    if (KN & KN_T1Z) op.one_v = pm.one | (threadIdx.x & pm.zero);
    constexpr unsigned long long kTile = (unsigned long long)BLOCK * ITER;
    const unsigned long long tile0 = (unsigned long long)blockIdx.x * kTile + threadIdx.x;

#pragma unroll 1
    for (int it = 0; it < ITER; it++) {
        const unsigned long long idx = tile0 + (unsigned long long)it * BLOCK;
        const uint32_t nonce = nonce_begin + (uint32_t)idx;
        const uint32_t w3 = __byte_perm(nonce, 0, 0x0123);
        uint32_t head[kSkip ? kSkip : 1];
        vs_schedule<KN>(p.jc[0], w3, op, [&](int i, uint32_t v) {
            if (i < kSkip) head[i] = v;
            else col[(i - kSkip) * BLOCK] = v;
        });

#pragma unroll 1
        for (uint32_t v = 0; v < n_ver; v++) {
            const JobConsts& jv = p.jc[v];
            // The schedule must be READ here, once per version, not kept in 46 registers across the loop: an
            // address the compiler cannot prove loop-invariant (pm.zero is a run-time 0) stops it hoisting the
            // loads out of the version loop and forwarding the stored values to them.
            const uint32_t* const colv = col + (v & pm.zero);
            const uint32_t e60 = vs_hash<KN>(jv, p.jc[0], w3, op, [&](int i) {
                return i < kSkip ? head[i] : colv[(i - kSkip) * BLOCK];
            });
            const bool cnd = __byte_perm(e60 + IV7, 0, 0x0123) <= p.jc[0].target[7];
            if (__any_sync(0xffffffffu, cnd)) {  // rare, warp-uniform
                uint32_t dig[8];
                bool hit = false;
                if (cnd && idx < count) {
                    atomicAdd(&ring->candidates, 1u);
                    hit = confirm_candidate(jv, nonce, dig);
                }
                ring_push_warp(hit, nonce, instance0 + v, dig, ring, entries, cap);
            }
        }
    }
    if (threadIdx.x == 0) atomicAdd(&ring->blocks_done, 1u);
}

#endif  // __CUDACC__

}  // namespace otd
