// scan_kernel.cuh — K1, the SHA-256d nonce-scan kernel for sm_100a.
//
// Replaces Worker.grind (internal/miner/worker.go:216-282): every thread takes
// V*ITER candidates of the window [nonce_begin, nonce_begin+count), runs the
// register-resident folded double compression (scan_core.cuh) and tests the
// most significant word of the little-endian hash integer against the target's
// (early reject on the final state word).  Survivors (about 2^-32 of candidates
// at difficulty >= 1) are re-hashed in full, compared
// against the whole 256-bit target exactly like Hash.LessOrEqual
// (internal/miner/sha256d.go:93-103) and compacted into the share ring with one
// warp ballot and one atomicAdd per warp.
//
// Bound: integer ALU issue.  The kernel reads no HBM in the hot loop (job
// constants arrive as a __grid_constant__ parameter = constant bank) and writes
// 40 bytes per share found.
#pragma once
#include "scan_core.cuh"

namespace otd {

struct RingHeader {
    unsigned int count;      // shares pushed (may exceed capacity: the excess was dropped)
    unsigned int candidates; // top-word survivors examined by the slow path
    unsigned int blocks_done; // thread blocks that ran to completion (progress when a launch is aborted)
    unsigned int pad;
};
struct RingEntry {
    uint32_t nonce;
    uint32_t instance;  // which header instance (version / extranonce2 roll) of the job
    uint8_t hash[32];   // SHA256d output bytes, same order as miner.Hash
};

constexpr int kScanBlock = 256;

#if defined(__CUDACC__)

// Full check of one surviving candidate; fills digest words when it is a share.
static __device__ __noinline__ bool confirm_candidate(const JobConsts& jc, uint32_t nonce, uint32_t dig[8])
{
    sha256d_from_consts(jc, nonce, dig);
    return digest_le_target(dig, jc.target);
}

__device__ __forceinline__ void ring_push_warp(bool hit, uint32_t nonce, uint32_t instance,
                                               const uint32_t dig[8], RingHeader* hdr,
                                               RingEntry* entries, uint32_t cap)
{
    const unsigned m = __ballot_sync(0xffffffffu, hit);
    if (m == 0) return;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(m) - 1;
    unsigned base = 0;
    if (lane == leader) base = atomicAdd(&hdr->count, (unsigned)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (hit) {
        const unsigned slot = base + __popc(m & ((1u << lane) - 1u));
        if (slot < cap) {
            RingEntry* en = &entries[slot];
            en->nonce = nonce;
            en->instance = instance;
            uint32_t* hw = reinterpret_cast<uint32_t*>(en->hash);
#pragma unroll
            for (int i = 0; i < 8; i++) hw[i] = bswap32(dig[i]);  // BE words -> byte order
        }
    }
}

// V: candidates interleaved per thread; ITER: sequential batches per thread;
// KN: instruction-placement knobs (scan_core.cuh).
// MINB: minimum resident blocks per SM handed to __launch_bounds__ (register budget = 65536 / (256*MINB)).
template <int V, int ITER, unsigned KN, bool HALF = false, int MINB = 1, int BLOCK = kScanBlock>
__global__ void __launch_bounds__(BLOCK, MINB)
scan_kernel(const __grid_constant__ JobConsts jc, const __grid_constant__ PipeMults pm,
            uint32_t nonce_begin, unsigned long long count,
            uint32_t instance, RingHeader* ring, RingEntry* entries, uint32_t cap,
            const unsigned int* stop_gen, unsigned int my_gen)
{
    // A newer job (or set_work(NULL)) bumps *stop_gen: blocks not yet started exit at once.
    if (*reinterpret_cast<const volatile unsigned int*>(stop_gen) != my_gen) return;

    Ops<KN> op{pm};
    if (KN & KN_T1Z) op.one_v = pm.one | (threadIdx.x & pm.zero);  // a 1 ptxas must keep in a vector register
    constexpr unsigned long long kTile = (unsigned long long)BLOCK * V * ITER;
    const unsigned long long tile0 = (unsigned long long)blockIdx.x * kTile + threadIdx.x;

    // early-reject test + rare path for the V candidates whose e60 words are `e60`, first index `idx0`
    auto test = [&](const vw<V>& e60, unsigned long long idx0) {
        unsigned cmask = 0;
#pragma unroll
        for (int v = 0; v < V; v++) {
            // most significant word of the little-endian hash integer vs the target's
            bool cnd;
            if ((KN & KN_T7Z) && jc.target[7] == 0u) cnd = e60.x[v] == kE60ForZeroTop;   // warp-uniform choice
            else cnd = __byte_perm(e60.x[v] + IV7, 0, 0x0123) <= jc.target[7];
            cmask |= (cnd ? 1u : 0u) << v;
        }
        if (__any_sync(0xffffffffu, cmask != 0)) {  // rare, warp-uniform
#pragma unroll 1
            for (int v = 0; v < V; v++) {
                const unsigned long long idx = idx0 + (unsigned long long)v * BLOCK;
                const uint32_t nonce = nonce_begin + (uint32_t)idx;
                uint32_t dig[8];
                bool hit = false;
                if (((cmask >> v) & 1u) && idx < count) {
                    atomicAdd(&ring->candidates, 1u);
                    hit = confirm_candidate(jc, nonce, dig);
                }
                ring_push_warp(hit, nonce, instance, dig, ring, entries, cap);
            }
        }
    };

    vw<V> pend = bc<V>(0u);  // KN_LATE: the previous iteration's result, tested inside the next one
#pragma unroll 1
    for (int it = 0; it < ITER; it++) {
        const unsigned long long idx0 = tile0 + (unsigned long long)it * (BLOCK * V);
        vw<V> w3;
#pragma unroll
        for (int v = 0; v < V; v++)
            w3.x[v] = __byte_perm(nonce_begin + (uint32_t)idx0 + (uint32_t)(v * BLOCK), 0, 0x0123);
        if (KN & KN_LATE) {
            const vw<V> cur = scan_core<V, KN, HALF>(jc, w3, op, [&] {
                if (it > 0) test(pend, idx0 - (unsigned long long)(BLOCK * V));
            });
            pend = cur;
        } else {
            const vw<V> e60 = (KN & KN_COMPACT) ? scan_core_compact<V, KN>(jc, w3, op) : scan_core<V, KN, HALF>(jc, w3, op);
            test(e60, idx0);
        }
    }
    if (KN & KN_LATE) test(pend, tile0 + (unsigned long long)(ITER - 1) * (BLOCK * V));
    if (threadIdx.x == 0) atomicAdd(&ring->blocks_done, 1u);
}

#endif  // __CUDACC__

}  // namespace otd
