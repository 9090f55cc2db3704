// sha256_core.cuh — SHA-256 building blocks shared by host and device code of
// libotedama_b200 (product code; independent of oracle/).
//
// Path being replaced: internal/miner/sha256d.go:107-117 (SHA256d/HashHeader),
// which the reference computes with Go's crypto/sha256 — three compressions per
// candidate, no midstate.  Here block 1 of the header (bytes 0..63) is
// compressed once per header instance (the "midstate"), the nonce-independent
// part of block 2 is folded into per-job constants (JobConsts), and the scan
// kernel does the remaining ~2 compressions per candidate.
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define OTD_HD __host__ __device__ __forceinline__
#define OTD_HD_NOINL __host__ __device__
#else
#define OTD_HD inline
#define OTD_HD_NOINL
#endif

namespace otd {

#define OTD_K_LIST                                                                                  \
    0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u,      \
        0xab1c5ed5u, 0xd807aa98u, 0x12835b01u, 0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu,  \
        0x9bdc06a7u, 0xc19bf174u, 0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu,  \
        0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau, 0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u,  \
        0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u, 0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu,  \
        0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u, 0xa2bfe8a1u, 0xa81a664bu,  \
        0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u, 0x19a4c116u,  \
        0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u,  \
        0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u, 0x90befffau, 0xa4506cebu, 0xbef9a3f7u,  \
        0xc67178f2u

// Compile-time round constant (folds to an immediate in unrolled device code).
OTD_HD constexpr uint32_t Kc(int i)
{
    constexpr uint32_t k[64] = {OTD_K_LIST};
    return k[i];
}

constexpr uint32_t IV0 = 0x6a09e667u, IV1 = 0xbb67ae85u, IV2 = 0x3c6ef372u, IV3 = 0xa54ff53au,
                   IV4 = 0x510e527fu, IV5 = 0x9b05688cu, IV6 = 0x1f83d9abu, IV7 = 0x5be0cd19u;

OTD_HD constexpr uint32_t rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
OTD_HD constexpr uint32_t bsig0(uint32_t x) { return rotr(x, 2) ^ rotr(x, 13) ^ rotr(x, 22); }
OTD_HD constexpr uint32_t bsig1(uint32_t x) { return rotr(x, 6) ^ rotr(x, 11) ^ rotr(x, 25); }
OTD_HD constexpr uint32_t ssig0(uint32_t x) { return rotr(x, 7) ^ rotr(x, 18) ^ (x >> 3); }
OTD_HD constexpr uint32_t ssig1(uint32_t x) { return rotr(x, 17) ^ rotr(x, 19) ^ (x >> 10); }
OTD_HD constexpr uint32_t ch(uint32_t e, uint32_t f, uint32_t g) { return (e & f) ^ (~e & g); }
OTD_HD constexpr uint32_t maj(uint32_t a, uint32_t b, uint32_t c) { return (a & b) ^ (a & c) ^ (b & c); }
OTD_HD constexpr uint32_t bswap32(uint32_t x)
{
    return (x >> 24) | ((x >> 8) & 0xff00u) | ((x << 8) & 0xff0000u) | (x << 24);
}

// Runtime-indexed K table for the looped (non-unrolled) compression.
#if defined(__CUDACC__)
__device__ __constant__ __align__(16) const uint32_t kKdev[64] = {OTD_K_LIST};
#endif
static const uint32_t kKhost[64] = {OTD_K_LIST};

OTD_HD uint32_t Kr(int i)
{
#if defined(__CUDA_ARCH__)
    return kKdev[i];
#else
    return kKhost[i];
#endif
}

// Generic compression: st += compress(st, w[16]) with w as big-endian words.
// Used once per header instance (midstate), by the job-preparation kernels, and
// by the rare-hit confirmation path — never in the per-candidate hot loop.
OTD_HD_NOINL inline void compress(uint32_t st[8], const uint32_t win[16])
{
    uint32_t w[16];
    for (int i = 0; i < 16; i++) w[i] = win[i];
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
    for (int i = 0; i < 64; i++) {
        uint32_t wi;
        if (i < 16) {
            wi = w[i];
        } else {
            wi = w[i & 15] + ssig0(w[(i + 1) & 15]) + w[(i + 9) & 15] + ssig1(w[(i + 14) & 15]);
            w[i & 15] = wi;
        }
        uint32_t t1 = h + bsig1(e) + ch(e, f, g) + Kr(i) + wi;
        uint32_t t2 = bsig0(a) + maj(a, b, c);
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

OTD_HD void iv_init(uint32_t st[8])
{
    st[0] = IV0; st[1] = IV1; st[2] = IV2; st[3] = IV3; st[4] = IV4; st[5] = IV5; st[6] = IV6; st[7] = IV7;
}

// ---------------------------------------------------------------------------
// JobConsts: everything about one header instance that does not depend on the
// nonce.  Produced once per instance (host for a plain header, the K2/K3
// preparation kernels for rolled versions / extranonce2) and handed to the scan
// kernel as a by-value kernel parameter, i.e. it lives in the constant bank and
// feeds ALU instructions as c[][] operands without occupying registers.
// ---------------------------------------------------------------------------
struct JobConsts {
    uint32_t mid[8];   // SHA-256 state after header bytes 0..63 (the midstate)
    uint32_t w0, w1, w2;  // header bytes 64..75 as big-endian words (merkle tail, time, bits)
    // state after block-2 rounds 0..2 (nonce enters at round 3):
    uint32_t a3, b3, c3, e3, f3, g3;
    uint32_t ca3, ce3;    // round 3: a = ca3 + W3, e = ce3 + W3
    uint32_t k4;          // g3 + K4 + W4            (T1 constant part of round 4)
    uint32_t k5;          // f3 + K5
    uint32_t k6;          // e3 + K6
    uint32_t w16, w17;    // nonce-independent schedule words
    uint32_t kw16, kw17;  // K16 + W16, K17 + W17
    uint32_t c18;         // W18 = c18 + ssig0(W3)
    uint32_t c19;         // W19 = c19 + W3
    uint32_t c31;         // W31 = ssig1(W29) + W24 + c31      (c31 = ssig0(W16) + W15)
    uint32_t c32;         // W32 = ssig1(W30) + W25 + c32      (c32 = ssig0(W17) + W16)
    uint32_t target[8];   // share target as LE words: target[i] = LE32(bytes 4i..4i+3); [7] most significant
};

// Fold the nonce-independent work of block 2 (reference has no counterpart: it
// re-hashes all 80 bytes per nonce, internal/miner/worker.go:250-253).
OTD_HD void fold_job_consts(const uint32_t mid[8], uint32_t w0, uint32_t w1, uint32_t w2,
                            const uint32_t target_le[8], JobConsts* jc)
{
    for (int i = 0; i < 8; i++) { jc->mid[i] = mid[i]; jc->target[i] = target_le[i]; }
    jc->w0 = w0; jc->w1 = w1; jc->w2 = w2;
    uint32_t a = mid[0], b = mid[1], c = mid[2], d = mid[3], e = mid[4], f = mid[5], g = mid[6], h = mid[7];
    const uint32_t w[3] = {w0, w1, w2};
    for (int i = 0; i < 3; i++) {
        uint32_t t1 = h + bsig1(e) + ch(e, f, g) + Kc(i) + w[i];
        uint32_t t2 = bsig0(a) + maj(a, b, c);
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    // round 3 with W3 = nonce word left symbolic
    uint32_t t1c = h + bsig1(e) + ch(e, f, g) + Kc(3);
    jc->ce3 = d + t1c;
    jc->ca3 = t1c + bsig0(a) + maj(a, b, c);
    jc->a3 = a; jc->b3 = b; jc->c3 = c; jc->e3 = e; jc->f3 = f; jc->g3 = g;
    jc->k4 = g + Kc(4) + 0x80000000u;
    jc->k5 = f + Kc(5);
    jc->k6 = e + Kc(6);
    const uint32_t W15 = 0x280u;  // bit length of the 80-byte message
    uint32_t w16 = w0 + ssig0(w1);              // + W9 (0) + ssig1(W14) (0)
    uint32_t w17 = w1 + ssig0(w2) + ssig1(W15); // + W10 (0)
    jc->w16 = w16; jc->w17 = w17;
    jc->kw16 = Kc(16) + w16; jc->kw17 = Kc(17) + w17;
    jc->c18 = w2 + ssig1(w16);                  // + W11 (0) + ssig0(W3)
    jc->c19 = ssig0(0x80000000u) + ssig1(w17);  // + W12 (0) + W3
    jc->c31 = ssig0(w16) + W15;
    jc->c32 = ssig0(w17) + w16;
}

// Midstate of an 80-byte header (bytes 0..63) and the three tail words.
OTD_HD void header_midstate(const uint8_t hdr[80], uint32_t mid[8], uint32_t tail[3])
{
    uint32_t w[16];
    for (int i = 0; i < 16; i++)
        w[i] = ((uint32_t)hdr[4 * i] << 24) | ((uint32_t)hdr[4 * i + 1] << 16) |
               ((uint32_t)hdr[4 * i + 2] << 8) | (uint32_t)hdr[4 * i + 3];
    iv_init(mid);
    compress(mid, w);
    for (int i = 0; i < 3; i++)
        tail[i] = ((uint32_t)hdr[64 + 4 * i] << 24) | ((uint32_t)hdr[65 + 4 * i] << 16) |
                  ((uint32_t)hdr[66 + 4 * i] << 8) | (uint32_t)hdr[67 + 4 * i];
}

// Full SHA256d of one candidate from its JobConsts (midstate path, looped
// code).  Digest is returned as 8 big-endian state words; byte j of the hash is
// byte (3 - j%4) of word j/4.
OTD_HD_NOINL inline void sha256d_from_consts(const JobConsts& jc, uint32_t nonce, uint32_t out[8])
{
    uint32_t st[8], w[16];
    for (int i = 0; i < 8; i++) st[i] = jc.mid[i];
    w[0] = jc.w0; w[1] = jc.w1; w[2] = jc.w2; w[3] = bswap32(nonce); w[4] = 0x80000000u;
    for (int i = 5; i < 15; i++) w[i] = 0;
    w[15] = 0x280u;
    compress(st, w);
    for (int i = 0; i < 8; i++) w[i] = st[i];
    w[8] = 0x80000000u;
    for (int i = 9; i < 15; i++) w[i] = 0;
    w[15] = 0x100u;
    iv_init(st);
    compress(st, w);
    for (int i = 0; i < 8; i++) out[i] = st[i];
}

// Hash.LessOrEqual (internal/miner/sha256d.go:93-103) on digest state words:
// the hash is a little-endian 256-bit integer, so the most significant 32 bits
// are bswap32(H7), then bswap32(H6) ... ; equality counts as a share.
OTD_HD bool digest_le_target(const uint32_t dig[8], const uint32_t target_le[8])
{
    for (int i = 7; i >= 0; i--) {
        uint32_t hw = bswap32(dig[i]);
        if (hw < target_le[i]) return true;
        if (hw > target_le[i]) return false;
    }
    return true;
}

}  // namespace otd
