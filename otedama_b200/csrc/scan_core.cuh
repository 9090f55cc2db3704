// scan_core.cuh — the per-candidate arithmetic of the SHA-256d nonce scan,
// fully unrolled and register resident.  Host+device so the folding can be
// unit-checked on the CPU (tests/test_fold_host.py) without a GPU; the shipped
// path is the device instantiation in scan_kernel.cuh.
//
// Replaces the body of Worker.grind's inner loop (internal/miner/worker.go:
// 248-280): h.Nonce = nonce; hash := HashHeader(h); hash.LessOrEqual(target).
//
// Per candidate (V candidates are interleaved per thread for ILP):
//   hash 1, block 2: rounds 3..63 (rounds 0..2 and the nonce-independent
//                    schedule terms come pre-folded in JobConsts);
//   hash 2:          rounds 0..60, only the e-chain in rounds 57..60 — enough
//                    to know digest word 7 (= IV7 + e after round 60), the most
//                    significant word of the little-endian hash integer.
// A candidate whose top word passes the cheap test is re-hashed in full by the
// caller (rare path).
#pragma once
#include "sha256_core.cuh"

namespace otd {

// ---- tiny SIMD-in-registers vector: V independent candidates per thread ----
template <int V>
struct vw {
    uint32_t x[V];
};

#define OTD_VEC_BINOP(OP)                                                                   \
    template <int V> OTD_HD vw<V> operator OP(const vw<V>& a, const vw<V>& b)               \
    {                                                                                       \
        vw<V> r;                                                                            \
        _Pragma("unroll") for (int i = 0; i < V; i++) r.x[i] = a.x[i] OP b.x[i];            \
        return r;                                                                           \
    }                                                                                       \
    template <int V> OTD_HD vw<V> operator OP(const vw<V>& a, uint32_t b)                   \
    {                                                                                       \
        vw<V> r;                                                                            \
        _Pragma("unroll") for (int i = 0; i < V; i++) r.x[i] = a.x[i] OP b;                 \
        return r;                                                                           \
    }
OTD_VEC_BINOP(+)
OTD_VEC_BINOP(^)
OTD_VEC_BINOP(&)
OTD_VEC_BINOP(|)
#undef OTD_VEC_BINOP

template <int V> OTD_HD vw<V> bc(uint32_t s)
{
    vw<V> r;
#pragma unroll
    for (int i = 0; i < V; i++) r.x[i] = s;
    return r;
}

#define OTD_VEC_UNARY(NAME)                                                  \
    template <int V> OTD_HD vw<V> NAME(const vw<V>& a)                       \
    {                                                                        \
        vw<V> r;                                                             \
        _Pragma("unroll") for (int i = 0; i < V; i++) r.x[i] = NAME(a.x[i]); \
        return r;                                                            \
    }
OTD_VEC_UNARY(bsig0)
OTD_VEC_UNARY(bsig1)
OTD_VEC_UNARY(ssig0)
OTD_VEC_UNARY(ssig1)
#undef OTD_VEC_UNARY

template <int V> OTD_HD vw<V> ch(const vw<V>& e, const vw<V>& f, const vw<V>& g)
{
    vw<V> r;
#pragma unroll
    for (int i = 0; i < V; i++) r.x[i] = ch(e.x[i], f.x[i], g.x[i]);
    return r;
}
template <int V> OTD_HD vw<V> maj(const vw<V>& a, const vw<V>& b, const vw<V>& c)
{
    vw<V> r;
#pragma unroll
    for (int i = 0; i < V; i++) r.x[i] = maj(a.x[i], b.x[i], c.x[i]);
    return r;
}

// One full round.  On return d holds the new e and h holds the new a; callers
// rotate the argument order instead of moving registers.
template <int V, typename KW>
OTD_HD void rnd(const vw<V>& a, const vw<V>& b, const vw<V>& c, vw<V>& d, const vw<V>& e,
                const vw<V>& f, const vw<V>& g, vw<V>& h, const KW& kw)
{
    vw<V> t1 = h + bsig1(e) + ch(e, f, g) + kw;
    d = d + t1;
    h = t1 + bsig0(a) + maj(a, b, c);
}
// e-chain only (the a value produced here is never consumed).
template <int V, typename KW>
OTD_HD void rnd_e(vw<V>& d, const vw<V>& e, const vw<V>& f, const vw<V>& g, const vw<V>& h,
                  const KW& kw)
{
    d = d + h + bsig1(e) + ch(e, f, g) + kw;
}

// 8 rounds starting at round index R with schedule words w[R..R+7] already final.
#define OTD_R8(R, a, b, c, d, e, f, g, h)      \
    rnd(a, b, c, d, e, f, g, h, w[(R) + 0] + Kc((R) + 0)); \
    rnd(h, a, b, c, d, e, f, g, w[(R) + 1] + Kc((R) + 1)); \
    rnd(g, h, a, b, c, d, e, f, w[(R) + 2] + Kc((R) + 2)); \
    rnd(f, g, h, a, b, c, d, e, w[(R) + 3] + Kc((R) + 3)); \
    rnd(e, f, g, h, a, b, c, d, w[(R) + 4] + Kc((R) + 4)); \
    rnd(d, e, f, g, h, a, b, c, w[(R) + 5] + Kc((R) + 5)); \
    rnd(c, d, e, f, g, h, a, b, w[(R) + 6] + Kc((R) + 6)); \
    rnd(b, c, d, e, f, g, h, a, w[(R) + 7] + Kc((R) + 7));

// Returns e after round 60 of the second hash for V candidates whose block-2
// word 3 (the big-endian nonce word) is w3.  Digest word 7 = IV7 + result.
template <int V>
OTD_HD vw<V> scan_core(const JobConsts& jc, const vw<V>& w3)
{
    vw<V> w[64];
    vw<V> a, b, c, d, e, f, g, h;

    // ---------------- hash 1, block 2 ----------------
    // state after round 3 (rounds 0..2 folded on the host / in K2-K3)
    a = w3 + jc.ca3;      b = bc<V>(jc.a3); c = bc<V>(jc.b3); d = bc<V>(jc.c3);
    e = w3 + jc.ce3;      f = bc<V>(jc.e3); g = bc<V>(jc.f3); h = bc<V>(jc.g3);

    // rounds 4..6: h and d are still job constants (k4/k5/k6 carry h + K + W)
    {
        vw<V> t1 = bsig1(e) + ch(e, f, g) + jc.k4;                  // round 4
        vw<V> e5 = t1 + jc.c3;
        vw<V> a5 = t1 + bsig0(a) + maj(a, b, c);
        t1 = bsig1(e5) + ch(e5, e, f) + jc.k5;                      // round 5
        vw<V> e6 = t1 + jc.b3;
        vw<V> a6 = t1 + bsig0(a5) + maj(a5, a, b);
        t1 = bsig1(e6) + ch(e6, e5, e) + jc.k6;                     // round 6
        vw<V> e7 = t1 + jc.a3;
        vw<V> a7 = t1 + bsig0(a6) + maj(a6, a5, a);
        // state entering round 7: (a7, a6, a5, a4, e7, e6, e5, e4)
        vw<V> a4 = a, e4 = e;
        a = a7; b = a6; c = a5; d = a4; e = e7; f = e6; g = e5; h = e4;
    }
    // rounds 7..15: W5..W14 = 0, W15 = 0x280
    rnd(a, b, c, d, e, f, g, h, Kc(7));
    rnd(h, a, b, c, d, e, f, g, Kc(8));
    rnd(g, h, a, b, c, d, e, f, Kc(9));
    rnd(f, g, h, a, b, c, d, e, Kc(10));
    rnd(e, f, g, h, a, b, c, d, Kc(11));
    rnd(d, e, f, g, h, a, b, c, Kc(12));
    rnd(c, d, e, f, g, h, a, b, Kc(13));
    rnd(b, c, d, e, f, g, h, a, Kc(14));
    rnd(a, b, c, d, e, f, g, h, Kc(15) + 0x280u);
    // schedule words 18..33 (16, 17 are job constants)
    w[18] = ssig0(w3) + jc.c18;
    w[19] = w3 + jc.c19;
    w[20] = ssig1(w[18]) + 0x80000000u;
    w[21] = ssig1(w[19]);
    w[22] = ssig1(w[20]) + 0x280u;
    w[23] = ssig1(w[21]) + jc.w16;
    w[24] = ssig1(w[22]) + jc.w17;
    w[25] = ssig1(w[23]) + w[18];
    w[26] = ssig1(w[24]) + w[19];
    w[27] = ssig1(w[25]) + w[20];
    w[28] = ssig1(w[26]) + w[21];
    w[29] = ssig1(w[27]) + w[22];
    w[30] = ssig1(w[28]) + w[23] + ssig0(0x280u);
    w[31] = ssig1(w[29]) + w[24] + jc.c31;
    w[32] = ssig1(w[30]) + w[25] + jc.c32;
    w[33] = ssig1(w[31]) + w[26] + ssig0(w[18]) + jc.w17;
#pragma unroll
    for (int i = 34; i < 64; i++) w[i] = ssig1(w[i - 2]) + w[i - 7] + ssig0(w[i - 15]) + w[i - 16];

    rnd(h, a, b, c, d, e, f, g, jc.kw16);
    rnd(g, h, a, b, c, d, e, f, jc.kw17);
    rnd(f, g, h, a, b, c, d, e, w[18] + Kc(18));
    rnd(e, f, g, h, a, b, c, d, w[19] + Kc(19));
    rnd(d, e, f, g, h, a, b, c, w[20] + Kc(20));
    rnd(c, d, e, f, g, h, a, b, w[21] + Kc(21));
    rnd(b, c, d, e, f, g, h, a, w[22] + Kc(22));
    rnd(a, b, c, d, e, f, g, h, w[23] + Kc(23));
    rnd(h, a, b, c, d, e, f, g, w[24] + Kc(24));
    rnd(g, h, a, b, c, d, e, f, w[25] + Kc(25));
    rnd(f, g, h, a, b, c, d, e, w[26] + Kc(26));
    rnd(e, f, g, h, a, b, c, d, w[27] + Kc(27));
    rnd(d, e, f, g, h, a, b, c, w[28] + Kc(28));
    rnd(c, d, e, f, g, h, a, b, w[29] + Kc(29));
    rnd(b, c, d, e, f, g, h, a, w[30] + Kc(30));
    rnd(a, b, c, d, e, f, g, h, w[31] + Kc(31));
    // after round 31 the "a" role is back on register h
    OTD_R8(32, h, a, b, c, d, e, f, g)
    OTD_R8(40, h, a, b, c, d, e, f, g)
    OTD_R8(48, h, a, b, c, d, e, f, g)
    OTD_R8(56, h, a, b, c, d, e, f, g)
    // 64 - 3 = 61 rotations since the state was named at round 3... the role
    // order now is (h,a,b,c,d,e,f,g) = (A,B,C,D,E,F,G,H).

    // ---------------- hash 2 ----------------
    // message = digest of hash 1 (state + midstate), padding, length 256 bits
    w[0] = h + jc.mid[0]; w[1] = a + jc.mid[1]; w[2] = b + jc.mid[2]; w[3] = c + jc.mid[3];
    w[4] = d + jc.mid[4]; w[5] = e + jc.mid[5]; w[6] = f + jc.mid[6]; w[7] = g + jc.mid[7];
    w[8] = bc<V>(0x80000000u);
#pragma unroll
    for (int i = 9; i < 15; i++) w[i] = bc<V>(0u);
    w[15] = bc<V>(0x100u);
#pragma unroll
    for (int i = 16; i < 61; i++) w[i] = ssig1(w[i - 2]) + w[i - 7] + ssig0(w[i - 15]) + w[i - 16];

    a = bc<V>(IV0); b = bc<V>(IV1); c = bc<V>(IV2); d = bc<V>(IV3);
    e = bc<V>(IV4); f = bc<V>(IV5); g = bc<V>(IV6); h = bc<V>(IV7);
    OTD_R8(0, a, b, c, d, e, f, g, h)
    OTD_R8(8, a, b, c, d, e, f, g, h)
    OTD_R8(16, a, b, c, d, e, f, g, h)
    OTD_R8(24, a, b, c, d, e, f, g, h)
    OTD_R8(32, a, b, c, d, e, f, g, h)
    OTD_R8(40, a, b, c, d, e, f, g, h)
    OTD_R8(48, a, b, c, d, e, f, g, h)
    // round 56 full, rounds 57..60 e-chain only
    rnd(a, b, c, d, e, f, g, h, w[56] + Kc(56));
    rnd_e(c, d, e, f, g, w[57] + Kc(57));
    rnd_e(b, c, d, e, f, w[58] + Kc(58));
    rnd_e(a, b, c, d, e, w[59] + Kc(59));
    rnd_e(h, a, b, c, d, w[60] + Kc(60));
    return h;
}

// e after round 60 for which digest word 7 is zero.
constexpr uint32_t kE60ForZeroTop = 0u - IV7;

}  // namespace otd
