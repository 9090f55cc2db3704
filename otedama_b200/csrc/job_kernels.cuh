// job_kernels.cuh — per-job preparation kernels (run once per header instance,
// never per candidate) and the batch HashHeader kernel.
//
//   K3 roll_extranonce2_kernel : coinbase = coinb1||extranonce1||extranonce2||coinb2,
//        root = SHA256d(coinbase), folded through the merkle branch.  The reference
//        drops these inputs (internal/poolproto/stratumv1/parse.go:52-54,91), so this
//        follows the public Stratum V1 convention ("parity unpinned").
//   K2 instance_consts_kernel  : BIP-320 version roll + block-1 midstate + the
//        nonce-independent folding (JobConsts) for each (extranonce2, version) instance.
//   hash_headers_kernel        : SHA256d of n 80-byte headers — the batch form of
//        miner.HashHeader (internal/miner/sha256d.go:114-117).
#pragma once
#include "sha256_core.cuh"

namespace otd {

// BIP-320: scatter the low bits of k into the set bits of mask (LSB first).
OTD_HD uint32_t roll_version(uint32_t base, uint32_t mask, uint32_t k)
{
    uint32_t v = base & ~mask;
    for (int bit = 0; bit < 32; bit++) {
        if (mask & (1u << bit)) {
            if (k & 1u) v |= (1u << bit);
            k >>= 1;
        }
    }
    return v;
}

struct CoinbaseDesc {
    const uint8_t* blob;  // coinb1 || extranonce1 || coinb2 || branches (device memory)
    uint32_t l1, le1, le2, l2, n_branches;
};

struct HeaderTemplate {
    uint32_t version, version_mask;
    uint8_t prev_hash[32];
    uint8_t merkle_root[32];  // used when there is no coinbase
    uint32_t time, bits;
    uint32_t target[8];
};

#if defined(__CUDACC__)

__device__ __forceinline__ uint32_t cb_byte(const CoinbaseDesc& cd, unsigned long long e, uint32_t pos,
                                            uint32_t total, uint32_t padded)
{
    if (pos < cd.l1) return cd.blob[pos];
    uint32_t p = pos - cd.l1;
    if (p < cd.le1) return cd.blob[cd.l1 + p];
    p -= cd.le1;
    if (p < cd.le2) return p < 8 ? (uint32_t)((e >> (8 * p)) & 0xffu) : 0u;
    p -= cd.le2;
    if (p < cd.l2) return cd.blob[cd.l1 + cd.le1 + p];
    if (pos == total) return 0x80u;
    if (pos >= padded - 8) {
        const unsigned long long bits = (unsigned long long)total * 8ull;
        return (uint32_t)((bits >> (8 * (padded - 1 - pos))) & 0xffu);
    }
    return 0u;
}

// second hash of a 32-byte digest held as 8 BE words (in place)
__device__ __forceinline__ void sha256_of_digest(uint32_t dig[8])
{
    uint32_t w[16], st[8];
    for (int i = 0; i < 8; i++) w[i] = dig[i];
    w[8] = 0x80000000u;
    for (int i = 9; i < 15; i++) w[i] = 0;
    w[15] = 0x100u;
    iv_init(st);
    compress(st, w);
    for (int i = 0; i < 8; i++) dig[i] = st[i];
}

__global__ void roll_extranonce2_kernel(CoinbaseDesc cd, unsigned long long e_begin,
                                        unsigned long long e_stride, uint32_t n, uint8_t* roots)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long e = e_begin + (unsigned long long)i * e_stride;
    const uint32_t total = cd.l1 + cd.le1 + cd.le2 + cd.l2;
    const uint32_t padded = ((total + 9 + 63) / 64) * 64;
    uint32_t st[8], w[16];
    iv_init(st);
    for (uint32_t blk = 0; blk < padded / 64; blk++) {
        for (int wi = 0; wi < 16; wi++) {
            const uint32_t p = blk * 64 + wi * 4;
            w[wi] = (cb_byte(cd, e, p, total, padded) << 24) | (cb_byte(cd, e, p + 1, total, padded) << 16) |
                    (cb_byte(cd, e, p + 2, total, padded) << 8) | cb_byte(cd, e, p + 3, total, padded);
        }
        compress(st, w);
    }
    sha256_of_digest(st);  // root = SHA256d(coinbase)
    const uint8_t* br = cd.blob + cd.l1 + cd.le1 + cd.l2;
    for (uint32_t b = 0; b < cd.n_branches; b++) {
        for (int k = 0; k < 8; k++) {
            w[k] = st[k];
            const uint8_t* q = br + 32 * b + 4 * k;
            w[8 + k] = ((uint32_t)q[0] << 24) | ((uint32_t)q[1] << 16) | ((uint32_t)q[2] << 8) | q[3];
        }
        iv_init(st);
        compress(st, w);
        w[0] = 0x80000000u;
        for (int k = 1; k < 15; k++) w[k] = 0;
        w[15] = 0x200u;
        compress(st, w);
        sha256_of_digest(st);
    }
    uint32_t* out = reinterpret_cast<uint32_t*>(roots + 32ull * i);
    for (int k = 0; k < 8; k++) out[k] = bswap32(st[k]);
}

// Instance j (0-based within the job): e_idx = j / n_ver, version index
// k = ver_begin + (j % n_ver) * ver_stride.
__global__ void instance_consts_kernel(HeaderTemplate ht, const uint8_t* roots, uint32_t root_first_eidx,
                                       uint32_t n_ver, uint32_t ver_begin, uint32_t ver_stride,
                                       unsigned long long first, uint32_t n, JobConsts* out,
                                       uint32_t* versions)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long j = first + i;
    const uint32_t e_idx = (uint32_t)(j / n_ver);
    const uint32_t k = ver_begin + (uint32_t)(j % n_ver) * ver_stride;
    const uint32_t version = ht.version_mask ? roll_version(ht.version, ht.version_mask, k) : ht.version;
    const uint8_t* root = roots ? roots + 32ull * (e_idx - root_first_eidx) : ht.merkle_root;
    uint32_t w[16];
    w[0] = bswap32(version);
    for (int q = 0; q < 8; q++)
        w[1 + q] = ((uint32_t)ht.prev_hash[4 * q] << 24) | ((uint32_t)ht.prev_hash[4 * q + 1] << 16) |
                   ((uint32_t)ht.prev_hash[4 * q + 2] << 8) | ht.prev_hash[4 * q + 3];
    for (int q = 0; q < 7; q++)
        w[9 + q] = ((uint32_t)root[4 * q] << 24) | ((uint32_t)root[4 * q + 1] << 16) |
                   ((uint32_t)root[4 * q + 2] << 8) | root[4 * q + 3];
    uint32_t mid[8];
    iv_init(mid);
    compress(mid, w);
    const uint32_t w0 = ((uint32_t)root[28] << 24) | ((uint32_t)root[29] << 16) | ((uint32_t)root[30] << 8) | root[31];
    JobConsts jc;
    fold_job_consts(mid, w0, bswap32(ht.time), bswap32(ht.bits), ht.target, &jc);
    out[i] = jc;
    if (versions) versions[i] = version;
}

__global__ void hash_headers_kernel(const uint8_t* __restrict__ hdrs, unsigned long long n,
                                    uint8_t* __restrict__ out)
{
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // 80-byte records are 16-byte aligned (cudaMalloc base, 80 = 5 x 16): five 128-bit loads per
    // header through the read-only path, two 128-bit stores per digest.
    const uint4* h4 = reinterpret_cast<const uint4*>(hdrs + 80ull * i);
    uint32_t raw[20];
#pragma unroll
    for (int k = 0; k < 5; k++) {
        const uint4 v = __ldg(h4 + k);
        raw[4 * k] = v.x; raw[4 * k + 1] = v.y; raw[4 * k + 2] = v.z; raw[4 * k + 3] = v.w;
    }
    uint32_t w[16], st[8];
    for (int k = 0; k < 16; k++) w[k] = bswap32(raw[k]);
    iv_init(st);
    compress(st, w);
    for (int k = 0; k < 4; k++) w[k] = bswap32(raw[16 + k]);
    w[4] = 0x80000000u;
    for (int k = 5; k < 15; k++) w[k] = 0;
    w[15] = 0x280u;
    compress(st, w);
    sha256_of_digest(st);
    uint4* o4 = reinterpret_cast<uint4*>(out + 32ull * i);
    o4[0] = make_uint4(bswap32(st[0]), bswap32(st[1]), bswap32(st[2]), bswap32(st[3]));
    o4[1] = make_uint4(bswap32(st[4]), bswap32(st[5]), bswap32(st[6]), bswap32(st[7]));
}

#endif  // __CUDACC__
}  // namespace otd
