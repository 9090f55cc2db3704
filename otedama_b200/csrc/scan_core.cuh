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

// ---------------------------------------------------------------------------
// Instruction placement.  sm_100 issues integer work on two pipes (measured on
// B200): the ALU pipe (SHF/LOP3/IADD3/PRMT, 64 lanes/clk/SM) and
// the FMA pipe (IMAD 64 lanes/clk/SM), ~119 lanes/clk/SM together.  A straight
// compile puts ~92% of SHA-256 on the ALU pipe and is bound by it.  The knobs
// below move work to the FMA pipe; what survived measurement:
//   * adds as IMAD x*1+y — with the 1 in a UNIFORM register (two vector-register
//     reads per IMAD; with the 1 in a vector register the register file cannot feed
//     both pipes and the kernel loses 11 %).  An IMAD takes only one non-vector
//     operand, so an IMAD that also carries an immediate needs a second, vector-
//     register 1 (one_v) of its own — otherwise ptxas moves THE 1 into a vector
//     register for every IMAD of the kernel;
//   * shipped policy (OTD_KN_BEST): E | A | W1 | W2 | SP everywhere, and T1Z (no ALU
//     add at all in the round) in 6/8 of the rounds, T1X (one IADD3 carrying the
//     immediate round constant) in the rest.
// Measured dead ends, kept so profiles/r1_variant_sweep.log stays reproducible:
// rotations/shifts through IMAD.WIDE (SSHR/SROT/BS1/BS0: IMAD.WIDE does not
// co-issue with ALU work), HKW and T1 (dependent IMAD chains with the 1 forced
// into a vector register).
// ---------------------------------------------------------------------------
#define OTD_KN_Z_EIGHTHS(n) ((unsigned)(n) << 16)
#define OTD_KN_ZSEL(s) ((unsigned)(s) << 28)
enum : unsigned {
    KN_E = 1u,       // e' = d + t1                 as IMAD
    KN_HKW = 2u,     // h + K + W                   as 2 IMAD instead of IADD3
    KN_T1 = 4u,      // (h+K+W) + Sigma1 + Ch       as 2 IMAD instead of IADD3
    KN_A = 8u,       // t1 + Sigma0 + Maj           as 2 IMAD instead of IADD3
    KN_W1 = 16u,     // schedule: last add          as IMAD
    KN_W2 = 32u,     // schedule: first 3-input add as 2 IMAD
    KN_SSHR = 64u,   // sigma0/1: the plain shift   via IMAD.WIDE (hi half)
    KN_SROT = 128u,  // sigma0/1: both rotations    via IMAD.WIDE
    KN_BS1 = 256u,   // Sigma1: two of three rotations via IMAD.WIDE
    KN_BS0 = 512u,   // Sigma0: two of three rotations via IMAD.WIDE
    KN_T1X = 1024u,  // t1 = IMAD(IADD3(IMAD(h,1,W), Sigma1, K), 1, Ch): one ALU add, no IMAD with an immediate
    KN_SP = 4096u,   // register+register adds of the specialised parts (first schedule words, rounds 4..6) as IMAD
                     // (adds with a constant-bank operand stay on the ALU: moving them loses, measured)
    KN_T1Z = 8192u,  // zero ALU adds in a round: t1 = IMAD(IMAD(h,1,IMAD(W,1v,K_imm)), 1, IMAD(Sigma1,1,Ch)); the
                     // immediate-carrying IMAD uses a separate VECTOR-register 1 (one_v) so that all the other
                     // IMADs keep their uniform-register 1 (an IMAD takes only one non-vector operand).
                     // Applied to KN_Z_EIGHTHS(n)/8 of the rounds (the rest use the KN_T1X form): the best
                     // FMA:ALU balance is an interior point (measured: 5/8..6/8).
    KN_Z_SHIFT = 16u,  // bits 16..19: n of KN_Z_EIGHTHS(n), 1..8
    KN_LATE = 1u << 14,    // kernel level: the early-reject test of candidate i runs inside candidate i+1 (after its rounds
                           // 4..6) instead of at the end of candidate i's dependency chain; the last one is flushed after the loop
    KN_KCONST = 1u << 26,  // rounds 16/17 of hash 1 (K + W is a job constant in the constant bank): fixed one-IADD3 form.
                           // Without it the KN_T1Z placement of these two rounds is picked from the VALUE of the
                           // constant, i.e. by a run-time (uniform) branch over two copies of the round
    KN_T7Z = 1u << 27,     // kernel level: when the target's top word is zero (share difficulty >= 1) the early-reject
                           // test is "e60 == -IV7" (one ISETP) instead of add + byte swap + compare
    KN_COMPACT = 2048u,  // partially rolled body (scan_core_compact): the three fully generic 16-round blocks
                         // share one loop body so the code stays well inside the instruction cache
};

struct PipeMults {  // run-time constants the compiler cannot fold (kernel parameter)
    uint32_t one;                                  // 1
    uint32_t zero;                                 // 0: threadIdx & zero makes one_v thread-variant for ptxas
    uint32_t m2, m13, m22, m6, m11, m25;           // 2^(32-n) for the Sigma rotations
    uint32_t m7, m18, m3, m17, m19, m10;           // ... and for sigma0 / sigma1
};
OTD_HD PipeMults make_pipe_mults()
{
    PipeMults m;
    m.one = 1u;
    m.zero = 0u;
    m.m2 = 1u << 30; m.m13 = 1u << 19; m.m22 = 1u << 10; m.m6 = 1u << 26; m.m11 = 1u << 21; m.m25 = 1u << 7;
    m.m7 = 1u << 25; m.m18 = 1u << 14; m.m3 = 1u << 29; m.m17 = 1u << 15; m.m19 = 1u << 13; m.m10 = 1u << 22;
    return m;
}

// a*1 + b on the FMA pipe
OTD_HD uint32_t fma_add(uint32_t a, uint32_t b, uint32_t one)
{
#if defined(__CUDA_ARCH__)
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b));
    return d;
#else
    return a * one + b;
#endif
}
// x * 2^(32-n) as a 64-bit product on the FMA pipe: hi = x >> n, lo = x << (32-n)
OTD_HD void fma_wide(uint32_t x, uint32_t m, uint32_t& hi, uint32_t& lo)
{
#if defined(__CUDA_ARCH__)
    asm("{ .reg .u64 t;\n mul.wide.u32 t, %2, %3;\n mov.b64 {%0, %1}, t; }" : "=r"(lo), "=r"(hi) : "r"(x), "r"(m));
#else
    const unsigned long long t = (unsigned long long)x * m;
    hi = (uint32_t)(t >> 32);
    lo = (uint32_t)t;
#endif
}

template <unsigned KN>
struct Ops {
    PipeMults m;
    uint32_t one_v = 1u;  // 1 in a vector register (KN_T1Z); set by the kernel
    OTD_HD uint32_t add_if(bool on_fma, uint32_t a, uint32_t b) const { return on_fma ? fma_add(a, b, m.one) : a + b; }
    OTD_HD uint32_t add3_if(bool on_fma, uint32_t a, uint32_t b, uint32_t c) const
    {
        return on_fma ? fma_add(fma_add(a, b, m.one), c, m.one) : a + b + c;
    }
    OTD_HD uint32_t S1(uint32_t x) const
    {
        if (KN & KN_BS1) {
            uint32_t h1, l1, h2, l2;
            fma_wide(x, m.m11, h1, l1);
            fma_wide(x, m.m25, h2, l2);
            return rotr(x, 6) ^ h1 ^ l1 ^ h2 ^ l2;
        }
        return bsig1(x);
    }
    OTD_HD uint32_t S0(uint32_t x) const
    {
        if (KN & KN_BS0) {
            uint32_t h1, l1, h2, l2;
            fma_wide(x, m.m13, h1, l1);
            fma_wide(x, m.m22, h2, l2);
            return rotr(x, 2) ^ h1 ^ l1 ^ h2 ^ l2;
        }
        return bsig0(x);
    }
    OTD_HD uint32_t s0(uint32_t x) const
    {
        if (KN & KN_SROT) {
            uint32_t h1, l1, h2, l2, h3, l3;
            fma_wide(x, m.m7, h1, l1);
            fma_wide(x, m.m18, h2, l2);
            fma_wide(x, m.m3, h3, l3);
            return h1 ^ l1 ^ h2 ^ l2 ^ h3;
        }
        if (KN & KN_SSHR) {
            uint32_t h3, l3;
            fma_wide(x, m.m3, h3, l3);
            return rotr(x, 7) ^ rotr(x, 18) ^ h3;
        }
        return ssig0(x);
    }
    OTD_HD uint32_t s1(uint32_t x) const
    {
        if (KN & KN_SROT) {
            uint32_t h1, l1, h2, l2, h3, l3;
            fma_wide(x, m.m17, h1, l1);
            fma_wide(x, m.m19, h2, l2);
            fma_wide(x, m.m10, h3, l3);
            return h1 ^ l1 ^ h2 ^ l2 ^ h3;
        }
        if (KN & KN_SSHR) {
            uint32_t h3, l3;
            fma_wide(x, m.m10, h3, l3);
            return rotr(x, 17) ^ rotr(x, 19) ^ h3;
        }
        return ssig1(x);
    }
    // one message-schedule word: s1(w2) + w7 + s0(w15) + w16
    OTD_HD uint32_t sched(uint32_t w2, uint32_t w7, uint32_t w15, uint32_t w16) const
    {
        const uint32_t t = add3_if(KN & KN_W2, s1(w2), w7, s0(w15));
        return add_if(KN & KN_W1, t, w16);
    }
    // t1 = h + Sigma1(e) + Ch(e,f,g) + K + W  (W a per-candidate word)
    OTD_HD uint32_t t1_kw(uint32_t h, uint32_t e, uint32_t f, uint32_t g, uint32_t w, uint32_t k) const
    {
        if ((KN & KN_T1Z) && z_round(k)) {
            const uint32_t kw = fma_add(w, k, one_v);
            const uint32_t hk = fma_add(h, kw, m.one);
            const uint32_t sc = fma_add(S1(e), ch(e, f, g), m.one);
            return fma_add(hk, sc, m.one);
        }
        if (KN & (KN_T1X | KN_T1Z)) {
            const uint32_t x = fma_add(h, w, m.one);
            const uint32_t y = x + S1(e) + k;
            return fma_add(y, ch(e, f, g), m.one);
        }
        const uint32_t kw = add_if(KN & KN_HKW, w, k);
        const uint32_t hk = add_if(KN & KN_HKW, h, kw);
        return add3_if(KN & KN_T1, hk, S1(e), ch(e, f, g));
    }
    // KN_T1Z is applied to a fraction of the rounds, picked by three bits of the round constant (a fixed
    // pseudo-random subset of the 64 constants): n eighths, n from the KN_Z_EIGHTHS field (0 reads as 8).
    OTD_HD static constexpr unsigned z_eighths()
    {
        return ((KN >> KN_Z_SHIFT) & 15u) ? ((KN >> KN_Z_SHIFT) & 15u) : 8u;
    }
    // Which rounds: three bits of the round constant starting at bit 7 + 2 * KN_ZSEL (bits 28..31 of KN; 0 = the
    // shipped subset).  The subsets are equivalent in opcode mix; they differ in the order ptxas ends up with.
    OTD_HD static constexpr unsigned z_shift()
    {
        static_assert(((KN >> 28) & 15u) <= 12u, "KN_ZSEL: the 3-bit window must stay inside the 32-bit constant");
        return 7u + 2u * ((KN >> 28) & 15u);
    }
    OTD_HD static constexpr bool z_round(uint32_t k)
    {
        return (((k >> z_shift()) | (k << (32u - z_shift()))) & 7u) < z_eighths();
    }
    // same with K + W a job constant or an immediate
    OTD_HD uint32_t t1_k(uint32_t h, uint32_t e, uint32_t f, uint32_t g, uint32_t k) const
    {
        if ((KN & KN_T1Z) && z_round(k)) {
            const uint32_t hk = fma_add(h, k, one_v);
            const uint32_t sc = fma_add(S1(e), ch(e, f, g), m.one);
            return fma_add(hk, sc, m.one);
        }
        if (KN & (KN_T1X | KN_T1Z)) {
            const uint32_t y = h + S1(e) + k;
            return fma_add(y, ch(e, f, g), m.one);
        }
        const uint32_t hk = h + k;
        return add3_if(KN & KN_T1, hk, S1(e), ch(e, f, g));
    }
    // same for a job constant whose value is not known at compile time
    OTD_HD uint32_t t1_kc(uint32_t h, uint32_t e, uint32_t f, uint32_t g, uint32_t k) const
    {
        if (!(KN & KN_KCONST)) return t1_k(h, e, f, g, k);
        if (KN & (KN_T1X | KN_T1Z)) {
            const uint32_t y = h + S1(e) + k;
            return fma_add(y, ch(e, f, g), m.one);
        }
        return add3_if(KN & KN_T1, h + k, S1(e), ch(e, f, g));
    }
    OTD_HD void finish(uint32_t t1, uint32_t a, uint32_t b, uint32_t c, uint32_t& d, uint32_t& h) const
    {
        d = add_if(KN & KN_E, d, t1);
        h = add3_if(KN & KN_A, t1, S0(a), maj(a, b, c));
    }
};

// ---- vector wrappers over Ops -------------------------------------------------
// One full round with a per-candidate message word w and round constant k.  On
// return d holds the new e and h holds the new a; callers rotate the argument
// order instead of moving registers.
template <int V, unsigned KN>
OTD_HD void rnd(const Ops<KN>& op, const vw<V>& a, const vw<V>& b, const vw<V>& c, vw<V>& d, const vw<V>& e,
                const vw<V>& f, const vw<V>& g, vw<V>& h, const vw<V>& w, uint32_t k)
{
#pragma unroll
    for (int i = 0; i < V; i++) {
        const uint32_t t1 = op.t1_kw(h.x[i], e.x[i], f.x[i], g.x[i], w.x[i], k);
        op.finish(t1, a.x[i], b.x[i], c.x[i], d.x[i], h.x[i]);
    }
}
// ... with K + W a job constant or an immediate
template <int V, unsigned KN>
OTD_HD void rnd(const Ops<KN>& op, const vw<V>& a, const vw<V>& b, const vw<V>& c, vw<V>& d, const vw<V>& e,
                const vw<V>& f, const vw<V>& g, vw<V>& h, uint32_t kw)
{
#pragma unroll
    for (int i = 0; i < V; i++) {
        const uint32_t t1 = op.t1_k(h.x[i], e.x[i], f.x[i], g.x[i], kw);
        op.finish(t1, a.x[i], b.x[i], c.x[i], d.x[i], h.x[i]);
    }
}
// ... with K + W a job constant (constant bank, value unknown at compile time)
template <int V, unsigned KN>
OTD_HD void rnd_kc(const Ops<KN>& op, const vw<V>& a, const vw<V>& b, const vw<V>& c, vw<V>& d, const vw<V>& e,
                   const vw<V>& f, const vw<V>& g, vw<V>& h, uint32_t kw)
{
#pragma unroll
    for (int i = 0; i < V; i++) {
        const uint32_t t1 = op.t1_kc(h.x[i], e.x[i], f.x[i], g.x[i], kw);
        op.finish(t1, a.x[i], b.x[i], c.x[i], d.x[i], h.x[i]);
    }
}
// e-chain only (the a value produced here is never consumed).
template <int V, unsigned KN>
OTD_HD void rnd_e(const Ops<KN>& op, vw<V>& d, const vw<V>& e, const vw<V>& f, const vw<V>& g, const vw<V>& h,
                  const vw<V>& w, uint32_t k)
{
#pragma unroll
    for (int i = 0; i < V; i++) {
        const uint32_t t1 = op.t1_kw(h.x[i], e.x[i], f.x[i], g.x[i], w.x[i], k);
        d.x[i] = op.add_if(KN & KN_E, d.x[i], t1);
    }
}
template <int V, unsigned KN>
OTD_HD vw<V> sched(const Ops<KN>& op, const vw<V>& w2, const vw<V>& w7, const vw<V>& w15, const vw<V>& w16)
{
    vw<V> r;
#pragma unroll
    for (int i = 0; i < V; i++) r.x[i] = op.sched(w2.x[i], w7.x[i], w15.x[i], w16.x[i]);
    return r;
}
template <int V, unsigned KN> OTD_HD vw<V> vs0(const Ops<KN>& op, const vw<V>& x)
{
    vw<V> r;
#pragma unroll
    for (int i = 0; i < V; i++) r.x[i] = op.s0(x.x[i]);
    return r;
}
template <int V, unsigned KN> OTD_HD vw<V> vs1(const Ops<KN>& op, const vw<V>& x)
{
    vw<V> r;
#pragma unroll
    for (int i = 0; i < V; i++) r.x[i] = op.s1(x.x[i]);
    return r;
}
template <int V, unsigned KN> OTD_HD vw<V> vS0(const Ops<KN>& op, const vw<V>& x)
{
    vw<V> r;
#pragma unroll
    for (int i = 0; i < V; i++) r.x[i] = op.S0(x.x[i]);
    return r;
}
template <int V, unsigned KN> OTD_HD vw<V> vS1(const Ops<KN>& op, const vw<V>& x)
{
    vw<V> r;
#pragma unroll
    for (int i = 0; i < V; i++) r.x[i] = op.S1(x.x[i]);
    return r;
}

// adds of the specialised parts, placed by KN_SP / KN_CADD
template <int V, unsigned KN> OTD_HD vw<V> padd(const Ops<KN>& op, const vw<V>& a, const vw<V>& b)
{
    vw<V> r;
#pragma unroll
    for (int i = 0; i < V; i++) r.x[i] = op.add_if(KN & KN_SP, a.x[i], b.x[i]);
    return r;
}
template <int V, unsigned KN> OTD_HD vw<V> cadd(const Ops<KN>&, const vw<V>& a, uint32_t c)
{
    return a + c;  // job-constant / immediate addend: ALU
}

// 8 rounds starting at round index R with schedule words w[R..R+7] already final.
#define OTD_R8(R, a, b, c, d, e, f, g, h)                                \
    rnd(op, a, b, c, d, e, f, g, h, w[(R) + 0], Kc((R) + 0));   \
    rnd(op, h, a, b, c, d, e, f, g, w[(R) + 1], Kc((R) + 1));   \
    rnd(op, g, h, a, b, c, d, e, f, w[(R) + 2], Kc((R) + 2));   \
    rnd(op, f, g, h, a, b, c, d, e, w[(R) + 3], Kc((R) + 3));   \
    rnd(op, e, f, g, h, a, b, c, d, w[(R) + 4], Kc((R) + 4));   \
    rnd(op, d, e, f, g, h, a, b, c, w[(R) + 5], Kc((R) + 5));   \
    rnd(op, c, d, e, f, g, h, a, b, w[(R) + 6], Kc((R) + 6));   \
    rnd(op, b, c, d, e, f, g, h, a, w[(R) + 7], Kc((R) + 7));

// Hash 2 of SHA-256d up to e after round 60, from the eight digest words of hash 1
// (big-endian state words, midstate already added).  Digest word 7 of the double
// hash = IV7 + result.  Shared by the plain scan (scan_core) and the version-shared
// scan (scan_vs_core.cuh).
template <int V, unsigned KN>
OTD_HD vw<V> hash2_e60(const Ops<KN>& op, const vw<V>& x0, const vw<V>& x1, const vw<V>& x2, const vw<V>& x3,
                       const vw<V>& x4, const vw<V>& x5, const vw<V>& x6, const vw<V>& x7)
{
    // message = digest of hash 1, padding, length 256 bits
    vw<V> w[64];
    vw<V> a, b, c, d, e, f, g, h;
    w[0] = x0; w[1] = x1; w[2] = x2; w[3] = x3; w[4] = x4; w[5] = x5; w[6] = x6; w[7] = x7;
    // W8 = 0x80000000, W9..14 = 0, W15 = 0x100: written out so the zero terms vanish
    w[16] = padd(op, vs0(op, w[1]), w[0]);
    w[17] = padd(op, vs0(op, w[2]), w[1]) + ssig1(0x100u);
    w[18] = padd(op, padd(op, vs1(op, w[16]), vs0(op, w[3])), w[2]);
    w[19] = padd(op, padd(op, vs1(op, w[17]), vs0(op, w[4])), w[3]);
    w[20] = padd(op, padd(op, vs1(op, w[18]), vs0(op, w[5])), w[4]);
    w[21] = padd(op, padd(op, vs1(op, w[19]), vs0(op, w[6])), w[5]);
    w[22] = sched(op, w[20], bc<V>(0x100u), w[7], w[6]);
    w[23] = padd(op, padd(op, vs1(op, w[21]), w[16]), w[7]) + ssig0(0x80000000u);
    w[24] = padd(op, vs1(op, w[22]), w[17]) + 0x80000000u;
    w[25] = padd(op, vs1(op, w[23]), w[18]);
    w[26] = padd(op, vs1(op, w[24]), w[19]);
    w[27] = padd(op, vs1(op, w[25]), w[20]);
    w[28] = padd(op, vs1(op, w[26]), w[21]);
    w[29] = padd(op, vs1(op, w[27]), w[22]);
    w[30] = padd(op, vs1(op, w[28]), w[23]) + ssig0(0x100u);
    w[31] = padd(op, padd(op, vs1(op, w[29]), w[24]), vs0(op, w[16])) + 0x100u;
#pragma unroll
    for (int i = 32; i < 61; i++) w[i] = sched(op, w[i - 2], w[i - 7], w[i - 15], w[i - 16]);

    a = bc<V>(IV0); b = bc<V>(IV1); c = bc<V>(IV2); d = bc<V>(IV3);
    e = bc<V>(IV4); f = bc<V>(IV5); g = bc<V>(IV6); h = bc<V>(IV7);
    // rounds 0..3 still see compile-time constants (IV words) in the state: plain C so they fold
    {
        const Ops<0> fold{};
        rnd(fold, a, b, c, d, e, f, g, h, w[0], Kc(0));
        rnd(fold, h, a, b, c, d, e, f, g, w[1], Kc(1));
        rnd(fold, g, h, a, b, c, d, e, f, w[2], Kc(2));
        rnd(fold, f, g, h, a, b, c, d, e, w[3], Kc(3));
    }
    rnd(op, e, f, g, h, a, b, c, d, w[4], Kc(4));
    rnd(op, d, e, f, g, h, a, b, c, w[5], Kc(5));
    rnd(op, c, d, e, f, g, h, a, b, w[6], Kc(6));
    rnd(op, b, c, d, e, f, g, h, a, w[7], Kc(7));
    rnd(op, a, b, c, d, e, f, g, h, Kc(8) + 0x80000000u);
    rnd(op, h, a, b, c, d, e, f, g, Kc(9));
    rnd(op, g, h, a, b, c, d, e, f, Kc(10));
    rnd(op, f, g, h, a, b, c, d, e, Kc(11));
    rnd(op, e, f, g, h, a, b, c, d, Kc(12));
    rnd(op, d, e, f, g, h, a, b, c, Kc(13));
    rnd(op, c, d, e, f, g, h, a, b, Kc(14));
    rnd(op, b, c, d, e, f, g, h, a, Kc(15) + 0x100u);
    OTD_R8(16, a, b, c, d, e, f, g, h)
    OTD_R8(24, a, b, c, d, e, f, g, h)
    OTD_R8(32, a, b, c, d, e, f, g, h)
    OTD_R8(40, a, b, c, d, e, f, g, h)
    OTD_R8(48, a, b, c, d, e, f, g, h)
    // round 56 full, rounds 57..60 e-chain only
    rnd(op, a, b, c, d, e, f, g, h, w[56], Kc(56));
    rnd_e(op, c, d, e, f, g, w[57], Kc(57));
    rnd_e(op, b, c, d, e, f, w[58], Kc(58));
    rnd_e(op, a, b, c, d, e, w[59], Kc(59));
    rnd_e(op, h, a, b, c, d, w[60], Kc(60));
    return h;
}

// Returns e after round 60 of the second hash for V candidates whose block-2
// word 3 (the big-endian nonce word) is w3.  Digest word 7 = IV7 + result.
// HALF (performance experiments only, never shipped as a scan variant's result):
// stop after hash 1 and return its digest word 7 minus IV7, so that instruction
// placement can be studied on a body that fits the instruction cache.
struct NoHook {
    OTD_HD void operator()() const {}
};
// `hook` runs once, early in hash 1 (after rounds 4..6): the kernel's KN_LATE form parks the previous candidate's
// early-reject test there, away from the end of the dependency chain.
template <int V, unsigned KN, bool HALF = false, class Hook = NoHook>
OTD_HD vw<V> scan_core(const JobConsts& jc, const vw<V>& w3, const Ops<KN>& op, Hook&& hook = Hook())
{
    vw<V> w[64];
    vw<V> a, b, c, d, e, f, g, h;

    // ---------------- hash 1, block 2 ----------------
    // state after round 3 (rounds 0..2 folded on the host / in K2-K3)
    a = w3 + jc.ca3;      b = bc<V>(jc.a3); c = bc<V>(jc.b3); d = bc<V>(jc.c3);
    e = w3 + jc.ce3;      f = bc<V>(jc.e3); g = bc<V>(jc.f3); h = bc<V>(jc.g3);

    // rounds 4..6: h and d are still job constants (k4/k5/k6 carry h + K + W)
    {
        vw<V> t1 = cadd(op, padd(op, vS1(op, e), ch(e, f, g)), jc.k4);                  // round 4
        vw<V> e5 = cadd(op, t1, jc.c3);
        vw<V> a5 = padd(op, padd(op, t1, vS0(op, a)), maj(a, b, c));
        t1 = cadd(op, padd(op, vS1(op, e5), ch(e5, e, f)), jc.k5);                      // round 5
        vw<V> e6 = cadd(op, t1, jc.b3);
        vw<V> a6 = padd(op, padd(op, t1, vS0(op, a5)), maj(a5, a, b));
        t1 = cadd(op, padd(op, vS1(op, e6), ch(e6, e5, e)), jc.k6);                     // round 6
        vw<V> e7 = cadd(op, t1, jc.a3);
        vw<V> a7 = padd(op, padd(op, t1, vS0(op, a6)), maj(a6, a5, a));
        // state entering round 7: (a7, a6, a5, a4, e7, e6, e5, e4)
        vw<V> a4 = a, e4 = e;
        a = a7; b = a6; c = a5; d = a4; e = e7; f = e6; g = e5; h = e4;
    }
    hook();
    // rounds 7..15: W5..W14 = 0, W15 = 0x280
    rnd(op, a, b, c, d, e, f, g, h, Kc(7));
    rnd(op, h, a, b, c, d, e, f, g, Kc(8));
    rnd(op, g, h, a, b, c, d, e, f, Kc(9));
    rnd(op, f, g, h, a, b, c, d, e, Kc(10));
    rnd(op, e, f, g, h, a, b, c, d, Kc(11));
    rnd(op, d, e, f, g, h, a, b, c, Kc(12));
    rnd(op, c, d, e, f, g, h, a, b, Kc(13));
    rnd(op, b, c, d, e, f, g, h, a, Kc(14));
    rnd(op, a, b, c, d, e, f, g, h, Kc(15) + 0x280u);
    // schedule words 18..33 (16, 17 are job constants)
    w[18] = cadd(op, vs0(op, w3), jc.c18);
    w[19] = cadd(op, w3, jc.c19);
    w[20] = vs1(op, w[18]) + 0x80000000u;
    w[21] = vs1(op, w[19]);
    w[22] = vs1(op, w[20]) + 0x280u;
    w[23] = cadd(op, vs1(op, w[21]), jc.w16);
    w[24] = cadd(op, vs1(op, w[22]), jc.w17);
    w[25] = padd(op, vs1(op, w[23]), w[18]);
    w[26] = padd(op, vs1(op, w[24]), w[19]);
    w[27] = padd(op, vs1(op, w[25]), w[20]);
    w[28] = padd(op, vs1(op, w[26]), w[21]);
    w[29] = padd(op, vs1(op, w[27]), w[22]);
    w[30] = padd(op, vs1(op, w[28]), w[23]) + ssig0(0x280u);
    w[31] = cadd(op, padd(op, vs1(op, w[29]), w[24]), jc.c31);
    w[32] = cadd(op, padd(op, vs1(op, w[30]), w[25]), jc.c32);
    w[33] = sched(op, w[31], w[26], w[18], bc<V>(jc.w17));
#pragma unroll
    for (int i = 34; i < 64; i++) w[i] = sched(op, w[i - 2], w[i - 7], w[i - 15], w[i - 16]);

    rnd_kc(op, h, a, b, c, d, e, f, g, jc.kw16);
    rnd_kc(op, g, h, a, b, c, d, e, f, jc.kw17);
    rnd(op, f, g, h, a, b, c, d, e, w[18], Kc(18));
    rnd(op, e, f, g, h, a, b, c, d, w[19], Kc(19));
    rnd(op, d, e, f, g, h, a, b, c, w[20], Kc(20));
    rnd(op, c, d, e, f, g, h, a, b, w[21], Kc(21));
    rnd(op, b, c, d, e, f, g, h, a, w[22], Kc(22));
    rnd(op, a, b, c, d, e, f, g, h, w[23], Kc(23));
    rnd(op, h, a, b, c, d, e, f, g, w[24], Kc(24));
    rnd(op, g, h, a, b, c, d, e, f, w[25], Kc(25));
    rnd(op, f, g, h, a, b, c, d, e, w[26], Kc(26));
    rnd(op, e, f, g, h, a, b, c, d, w[27], Kc(27));
    rnd(op, d, e, f, g, h, a, b, c, w[28], Kc(28));
    rnd(op, c, d, e, f, g, h, a, b, w[29], Kc(29));
    rnd(op, b, c, d, e, f, g, h, a, w[30], Kc(30));
    rnd(op, a, b, c, d, e, f, g, h, w[31], Kc(31));
    // after round 31 the "a" role is back on register h
    OTD_R8(32, h, a, b, c, d, e, f, g)
    OTD_R8(40, h, a, b, c, d, e, f, g)
    OTD_R8(48, h, a, b, c, d, e, f, g)
    OTD_R8(56, h, a, b, c, d, e, f, g)
    // the role order now is (h,a,b,c,d,e,f,g) = (A,B,C,D,E,F,G,H).

    if (HALF) return g + (jc.mid[7] - IV7);

    // ---------------- hash 2 ----------------
    // message = digest of hash 1 (state + midstate)
    return hash2_e60<V, KN>(op, cadd(op, h, jc.mid[0]), cadd(op, a, jc.mid[1]), cadd(op, b, jc.mid[2]),
                            cadd(op, c, jc.mid[3]), cadd(op, d, jc.mid[4]), cadd(op, e, jc.mid[5]),
                            cadd(op, f, jc.mid[6]), cadd(op, g, jc.mid[7]));
}

// ---------------------------------------------------------------------------
// Compact form.  Same arithmetic as scan_core, but the three 16-round blocks
// whose message schedule is fully generic (hash 1 rounds 32..47 and 48..63, hash 2
// rounds 32..47) run through ONE loop body with the round constants fetched from
// the constant table, and the two hashes are the two trips of an outer loop so
// that body exists once in the code.  Everything with folding opportunities
// (first 32 rounds of each hash, the e-chain tail) stays specialised and
// unrolled.  The fully unrolled scan_core is ~50 KB of SASS, right at the edge of
// what the SM's instruction cache sustains with 40 warps at different PCs; this
// form is ~35 KB and leaves room for the extra IMADs of the add-offload policies.
// ---------------------------------------------------------------------------
template <int V, unsigned KN>
OTD_HD void generic16(const Ops<KN>& op, vw<V> (&w)[16], vw<V>& a, vw<V>& b, vw<V>& c, vw<V>& d, vw<V>& e,
                      vw<V>& f, vw<V>& g, vw<V>& h, int kbase)
{
    uint32_t k[16];
#pragma unroll
    for (int r = 0; r < 16; r++) k[r] = Kr(kbase + r);
#define OTD_GEN_W(r) w[r] = sched(op, w[((r) + 14) & 15], w[((r) + 9) & 15], w[((r) + 1) & 15], w[r])
#define OTD_GEN8(R)                                                   \
    OTD_GEN_W((R) + 0); rnd(op, a, b, c, d, e, f, g, h, w[(R) + 0], k[(R) + 0]); \
    OTD_GEN_W((R) + 1); rnd(op, h, a, b, c, d, e, f, g, w[(R) + 1], k[(R) + 1]); \
    OTD_GEN_W((R) + 2); rnd(op, g, h, a, b, c, d, e, f, w[(R) + 2], k[(R) + 2]); \
    OTD_GEN_W((R) + 3); rnd(op, f, g, h, a, b, c, d, e, w[(R) + 3], k[(R) + 3]); \
    OTD_GEN_W((R) + 4); rnd(op, e, f, g, h, a, b, c, d, w[(R) + 4], k[(R) + 4]); \
    OTD_GEN_W((R) + 5); rnd(op, d, e, f, g, h, a, b, c, w[(R) + 5], k[(R) + 5]); \
    OTD_GEN_W((R) + 6); rnd(op, c, d, e, f, g, h, a, b, w[(R) + 6], k[(R) + 6]); \
    OTD_GEN_W((R) + 7); rnd(op, b, c, d, e, f, g, h, a, w[(R) + 7], k[(R) + 7]);
    OTD_GEN8(0)
    OTD_GEN8(8)
#undef OTD_GEN8
}

template <int V, unsigned KN>
OTD_HD vw<V> scan_core_compact(const JobConsts& jc, const vw<V>& w3, const Ops<KN>& op)
{
    vw<V> w[16];                  // rolling window: W[i] lives in w[i & 15]
    vw<V> a, b, c, d, e, f, g, h; // roles (A..H) at every loop boundary
    vw<V> result = bc<V>(0u);
#pragma unroll 1
    for (int phase = 0; phase < 2; phase++) {
        if (phase == 0) {
            // ---------------- hash 1, block 2, rounds 3..31 ----------------
            vw<V> x[32];  // W16..W31 as named values
            a = w3 + jc.ca3;      b = bc<V>(jc.a3); c = bc<V>(jc.b3); d = bc<V>(jc.c3);
            e = w3 + jc.ce3;      f = bc<V>(jc.e3); g = bc<V>(jc.f3); h = bc<V>(jc.g3);
            {
                vw<V> t1 = cadd(op, padd(op, vS1(op, e), ch(e, f, g)), jc.k4);                  // round 4
                vw<V> e5 = cadd(op, t1, jc.c3);
                vw<V> a5 = padd(op, padd(op, t1, vS0(op, a)), maj(a, b, c));
                t1 = cadd(op, padd(op, vS1(op, e5), ch(e5, e, f)), jc.k5);                      // round 5
                vw<V> e6 = cadd(op, t1, jc.b3);
                vw<V> a6 = padd(op, padd(op, t1, vS0(op, a5)), maj(a5, a, b));
                t1 = cadd(op, padd(op, vS1(op, e6), ch(e6, e5, e)), jc.k6);                     // round 6
                vw<V> e7 = cadd(op, t1, jc.a3);
                vw<V> a7 = padd(op, padd(op, t1, vS0(op, a6)), maj(a6, a5, a));
                vw<V> a4 = a, e4 = e;
                a = a7; b = a6; c = a5; d = a4; e = e7; f = e6; g = e5; h = e4;
            }
            rnd(op, a, b, c, d, e, f, g, h, Kc(7));
            rnd(op, h, a, b, c, d, e, f, g, Kc(8));
            rnd(op, g, h, a, b, c, d, e, f, Kc(9));
            rnd(op, f, g, h, a, b, c, d, e, Kc(10));
            rnd(op, e, f, g, h, a, b, c, d, Kc(11));
            rnd(op, d, e, f, g, h, a, b, c, Kc(12));
            rnd(op, c, d, e, f, g, h, a, b, Kc(13));
            rnd(op, b, c, d, e, f, g, h, a, Kc(14));
            rnd(op, a, b, c, d, e, f, g, h, Kc(15) + 0x280u);
            x[18] = vs0(op, w3) + jc.c18;
            x[19] = w3 + jc.c19;
            x[20] = vs1(op, x[18]) + 0x80000000u;
            x[21] = vs1(op, x[19]);
            x[22] = vs1(op, x[20]) + 0x280u;
            x[23] = vs1(op, x[21]) + jc.w16;
            x[24] = vs1(op, x[22]) + jc.w17;
            x[25] = padd(op, vs1(op, x[23]), x[18]);
            x[26] = padd(op, vs1(op, x[24]), x[19]);
            x[27] = padd(op, vs1(op, x[25]), x[20]);
            x[28] = padd(op, vs1(op, x[26]), x[21]);
            x[29] = padd(op, vs1(op, x[27]), x[22]);
            x[30] = padd(op, vs1(op, x[28]), x[23]) + ssig0(0x280u);
            x[31] = padd(op, vs1(op, x[29]), x[24]) + jc.c31;
            rnd(op, h, a, b, c, d, e, f, g, jc.kw16);
            rnd(op, g, h, a, b, c, d, e, f, jc.kw17);
            rnd(op, f, g, h, a, b, c, d, e, x[18], Kc(18));
            rnd(op, e, f, g, h, a, b, c, d, x[19], Kc(19));
            rnd(op, d, e, f, g, h, a, b, c, x[20], Kc(20));
            rnd(op, c, d, e, f, g, h, a, b, x[21], Kc(21));
            rnd(op, b, c, d, e, f, g, h, a, x[22], Kc(22));
            rnd(op, a, b, c, d, e, f, g, h, x[23], Kc(23));
            rnd(op, h, a, b, c, d, e, f, g, x[24], Kc(24));
            rnd(op, g, h, a, b, c, d, e, f, x[25], Kc(25));
            rnd(op, f, g, h, a, b, c, d, e, x[26], Kc(26));
            rnd(op, e, f, g, h, a, b, c, d, x[27], Kc(27));
            rnd(op, d, e, f, g, h, a, b, c, x[28], Kc(28));
            rnd(op, c, d, e, f, g, h, a, b, x[29], Kc(29));
            rnd(op, b, c, d, e, f, g, h, a, x[30], Kc(30));
            rnd(op, a, b, c, d, e, f, g, h, x[31], Kc(31));
            // roles after round 31 are (h,a,b,c,d,e,f,g): rename so (a..h) = (A..H)
            {
                const vw<V> t = h;
                h = g; g = f; f = e; e = d; d = c; c = b; b = a; a = t;
            }
            w[0] = bc<V>(jc.w16); w[1] = bc<V>(jc.w17);
#pragma unroll
            for (int i = 18; i < 32; i++) w[i - 16] = x[i];
        } else {
            // ---------------- hash 2, rounds 0..31 ----------------
            vw<V> x[32];
            x[0] = a + jc.mid[0]; x[1] = b + jc.mid[1]; x[2] = c + jc.mid[2]; x[3] = d + jc.mid[3];
            x[4] = e + jc.mid[4]; x[5] = f + jc.mid[5]; x[6] = g + jc.mid[6]; x[7] = h + jc.mid[7];
            x[16] = padd(op, vs0(op, x[1]), x[0]);
            x[17] = padd(op, vs0(op, x[2]), x[1]) + ssig1(0x100u);
            x[18] = padd(op, padd(op, vs1(op, x[16]), vs0(op, x[3])), x[2]);
            x[19] = padd(op, padd(op, vs1(op, x[17]), vs0(op, x[4])), x[3]);
            x[20] = padd(op, padd(op, vs1(op, x[18]), vs0(op, x[5])), x[4]);
            x[21] = padd(op, padd(op, vs1(op, x[19]), vs0(op, x[6])), x[5]);
            x[22] = sched(op, x[20], bc<V>(0x100u), x[7], x[6]);
            x[23] = padd(op, padd(op, vs1(op, x[21]), x[16]), x[7]) + ssig0(0x80000000u);
            x[24] = padd(op, vs1(op, x[22]), x[17]) + 0x80000000u;
            x[25] = padd(op, vs1(op, x[23]), x[18]);
            x[26] = padd(op, vs1(op, x[24]), x[19]);
            x[27] = padd(op, vs1(op, x[25]), x[20]);
            x[28] = padd(op, vs1(op, x[26]), x[21]);
            x[29] = padd(op, vs1(op, x[27]), x[22]);
            x[30] = padd(op, vs1(op, x[28]), x[23]) + ssig0(0x100u);
            x[31] = padd(op, padd(op, vs1(op, x[29]), x[24]), vs0(op, x[16])) + 0x100u;
            a = bc<V>(IV0); b = bc<V>(IV1); c = bc<V>(IV2); d = bc<V>(IV3);
            e = bc<V>(IV4); f = bc<V>(IV5); g = bc<V>(IV6); h = bc<V>(IV7);
            {
                const Ops<0> fold{};
                rnd(fold, a, b, c, d, e, f, g, h, x[0], Kc(0));
                rnd(fold, h, a, b, c, d, e, f, g, x[1], Kc(1));
                rnd(fold, g, h, a, b, c, d, e, f, x[2], Kc(2));
                rnd(fold, f, g, h, a, b, c, d, e, x[3], Kc(3));
            }
            rnd(op, e, f, g, h, a, b, c, d, x[4], Kc(4));
            rnd(op, d, e, f, g, h, a, b, c, x[5], Kc(5));
            rnd(op, c, d, e, f, g, h, a, b, x[6], Kc(6));
            rnd(op, b, c, d, e, f, g, h, a, x[7], Kc(7));
            rnd(op, a, b, c, d, e, f, g, h, Kc(8) + 0x80000000u);
            rnd(op, h, a, b, c, d, e, f, g, Kc(9));
            rnd(op, g, h, a, b, c, d, e, f, Kc(10));
            rnd(op, f, g, h, a, b, c, d, e, Kc(11));
            rnd(op, e, f, g, h, a, b, c, d, Kc(12));
            rnd(op, d, e, f, g, h, a, b, c, Kc(13));
            rnd(op, c, d, e, f, g, h, a, b, Kc(14));
            rnd(op, b, c, d, e, f, g, h, a, Kc(15) + 0x100u);
            rnd(op, a, b, c, d, e, f, g, h, x[16], Kc(16));
            rnd(op, h, a, b, c, d, e, f, g, x[17], Kc(17));
            rnd(op, g, h, a, b, c, d, e, f, x[18], Kc(18));
            rnd(op, f, g, h, a, b, c, d, e, x[19], Kc(19));
            rnd(op, e, f, g, h, a, b, c, d, x[20], Kc(20));
            rnd(op, d, e, f, g, h, a, b, c, x[21], Kc(21));
            rnd(op, c, d, e, f, g, h, a, b, x[22], Kc(22));
            rnd(op, b, c, d, e, f, g, h, a, x[23], Kc(23));
            rnd(op, a, b, c, d, e, f, g, h, x[24], Kc(24));
            rnd(op, h, a, b, c, d, e, f, g, x[25], Kc(25));
            rnd(op, g, h, a, b, c, d, e, f, x[26], Kc(26));
            rnd(op, f, g, h, a, b, c, d, e, x[27], Kc(27));
            rnd(op, e, f, g, h, a, b, c, d, x[28], Kc(28));
            rnd(op, d, e, f, g, h, a, b, c, x[29], Kc(29));
            rnd(op, c, d, e, f, g, h, a, b, x[30], Kc(30));
            rnd(op, b, c, d, e, f, g, h, a, x[31], Kc(31));
            // 32 rounds from the identity: (a..h) = (A..H) again
#pragma unroll
            for (int i = 16; i < 32; i++) w[i - 16] = x[i];
        }
        // ---------------- the shared generic blocks (rounds 32..47 [, 48..63]) ----------------
        const int nblk = 2 - phase;
#pragma unroll 1
        for (int blk = 0; blk < nblk; blk++) generic16(op, w, a, b, c, d, e, f, g, h, 32 + 16 * blk);

        if (phase == 1) {
            // ---------------- hash 2 tail: rounds 48..60, e-chain only in 57..60 ----------------
            OTD_GEN_W(0);  rnd(op, a, b, c, d, e, f, g, h, w[0], Kc(48));
            OTD_GEN_W(1);  rnd(op, h, a, b, c, d, e, f, g, w[1], Kc(49));
            OTD_GEN_W(2);  rnd(op, g, h, a, b, c, d, e, f, w[2], Kc(50));
            OTD_GEN_W(3);  rnd(op, f, g, h, a, b, c, d, e, w[3], Kc(51));
            OTD_GEN_W(4);  rnd(op, e, f, g, h, a, b, c, d, w[4], Kc(52));
            OTD_GEN_W(5);  rnd(op, d, e, f, g, h, a, b, c, w[5], Kc(53));
            OTD_GEN_W(6);  rnd(op, c, d, e, f, g, h, a, b, w[6], Kc(54));
            OTD_GEN_W(7);  rnd(op, b, c, d, e, f, g, h, a, w[7], Kc(55));
            OTD_GEN_W(8);  rnd(op, a, b, c, d, e, f, g, h, w[8], Kc(56));
            OTD_GEN_W(9);  rnd_e(op, c, d, e, f, g, w[9], Kc(57));
            OTD_GEN_W(10); rnd_e(op, b, c, d, e, f, w[10], Kc(58));
            OTD_GEN_W(11); rnd_e(op, a, b, c, d, e, w[11], Kc(59));
            OTD_GEN_W(12); rnd_e(op, h, a, b, c, d, w[12], Kc(60));
            result = h;
        }
    }
    return result;
}
#undef OTD_GEN_W

// e after round 60 for which digest word 7 is zero.
constexpr uint32_t kE60ForZeroTop = 0u - IV7;

}  // namespace otd
