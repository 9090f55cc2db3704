// scan_vs_core.cuh — the per-candidate arithmetic of the VERSION-SHARED scan:
// one nonce, many BIP-320 version instances of the same job.
//
// The reference has no version rolling (it ignores the BIP-310 extension,
// internal/poolproto/stratumv1/stratumv1.go:280-281) and re-hashes all 80 bytes
// per candidate (internal/miner/worker.go:250-253); north_star asks for the
// nonce x version space.  What this file exploits: the version sits in header
// bytes 0..3, i.e. in block 1 of hash 1.  Block 2 (merkle tail, time, bits,
// nonce, padding) is the same for every version of a job, so its whole message
// schedule W16..W63 — about 300 of the ~1850 ALU-pipe instructions of one
// candidate — depends on the nonce only.  Per nonce the schedule is expanded
// ONCE (vs_schedule: K[i] + W[i] for rounds 18..63, 46 words, parked in shared
// memory) and every version instance then runs only the state recurrence of
// block 2 from its own midstate plus hash 2 (vs_hash).
//
// Host+device like scan_core.cuh, so the folding is checked on the CPU against
// the oracle (tests/test_fold_host.py) before a GPU ever sees it.
#pragma once
#include "scan_core.cuh"

namespace otd {

// First round whose K+W is parked in shared memory: 18 (46 words per nonce), or 20 with KN_VS_F20 — rounds 18
// and 19 then travel in two registers across the version loop, and 44 words per nonce let five 256-thread blocks
// (40 warps) share an SM's 228 KB instead of four.
template <unsigned KN> OTD_HD constexpr int vs_first_round() { return (KN & (1u << 24)) ? 20 : 18; }
template <unsigned KN> OTD_HD constexpr int vs_words() { return 64 - vs_first_round<KN>(); }

// Knob bits private to the version-shared core (above scan_core.cuh's).
#define OTD_KN_VSZ_EIGHTHS(n) ((unsigned)(n) << 20)
enum : unsigned {
    KN_VSZ_SHIFT = 20u,  // bits 20..23: n of OTD_KN_VSZ_EIGHTHS(n), 1..8 (0 reads as 8): the share of the shared-schedule
                         // rounds whose t1 is built from IMADs only (h + kw, Sigma1 + Ch, sum); the rest keep one
                         // IADD3 (h + Sigma1 + kw) on the ALU pipe — the FMA:ALU balance knob of scan_core.cuh's
                         // KN_T1Z, for rounds that need no immediate
    KN_VS_F20 = 1u << 24,  // see vs_first_round()
    KN_VS_CFMA = 1u << 15,  // per-instance job constants (midstate words, round 3..6 folding) added as IMADs on the FMA pipe
                            // instead of IADD with a uniform-register operand on the ALU pipe
    KN_VS_LOOP1 = 1u << 25,  // hash 1 rounds 24..63 as five trips of one 8-round body instead of 40 unrolled rounds:
                             // these rounds carry no schedule window (K+W comes from shared memory), so rolling
                             // them costs no registers and takes ~10 KB off the instruction-cache footprint
};

template <unsigned KN>
struct VsOps : Ops<KN> {
    OTD_HD static constexpr unsigned vsz_eighths()
    {
        return ((KN >> KN_VSZ_SHIFT) & 15u) ? ((KN >> KN_VSZ_SHIFT) & 15u) : 8u;
    }
    OTD_HD static constexpr bool vsz_round(uint32_t k) { return ((k >> 7) & 7u) < vsz_eighths(); }
    // t1 = h + Sigma1(e) + Ch(e,f,g) + kw, kw = K + W already summed (a register, from the shared schedule).
    // ksel is the round constant, used only to pick the placement of this round at compile time.
    OTD_HD uint32_t t1_r(uint32_t h, uint32_t e, uint32_t f, uint32_t g, uint32_t kw, uint32_t ksel) const
    {
        if ((KN & KN_T1Z) && vsz_round(ksel)) {
            const uint32_t hk = fma_add(h, kw, this->m.one);
            const uint32_t sc = fma_add(this->S1(e), ch(e, f, g), this->m.one);
            return fma_add(hk, sc, this->m.one);
        }
        if (KN & (KN_T1X | KN_T1Z)) {
            const uint32_t y = h + this->S1(e) + kw;
            return fma_add(y, ch(e, f, g), this->m.one);
        }
        const uint32_t hk = this->add_if(KN & KN_HKW, h, kw);
        return this->add3_if(KN & KN_T1, hk, this->S1(e), ch(e, f, g));
    }
    // x + c for a per-instance job constant c
    OTD_HD vw<1> vadd(const vw<1>& x, uint32_t c) const
    {
        vw<1> r;
        r.x[0] = (KN & KN_VS_CFMA) ? fma_add(x.x[0], c, this->one_v) : x.x[0] + c;  // one_v: the constant is the one non-vector operand
        return r;
    }
    // a round whose K + W is a job constant (constant bank): the placement must not be picked from the value
    // (that would be a run-time branch), so the one-IADD3 form is used whenever t1 is split at all
    OTD_HD void rnd_c(const vw<1>& a, const vw<1>& b, const vw<1>& c, vw<1>& d, const vw<1>& e, const vw<1>& f,
                      const vw<1>& g, vw<1>& h, uint32_t kc) const
    {
        uint32_t t1;
        if (KN & (KN_T1X | KN_T1Z)) {
            const uint32_t y = h.x[0] + this->S1(e.x[0]) + kc;
            t1 = fma_add(y, ch(e.x[0], f.x[0], g.x[0]), this->m.one);
        } else {
            t1 = this->add3_if(KN & KN_T1, h.x[0] + kc, this->S1(e.x[0]), ch(e.x[0], f.x[0], g.x[0]));
        }
        this->finish(t1, a.x[0], b.x[0], c.x[0], d.x[0], h.x[0]);
    }
    OTD_HD void rnd_r(const vw<1>& a, const vw<1>& b, const vw<1>& c, vw<1>& d, const vw<1>& e, const vw<1>& f,
                      const vw<1>& g, vw<1>& h, uint32_t kw, uint32_t ksel) const
    {
        const uint32_t t1 = t1_r(h.x[0], e.x[0], f.x[0], g.x[0], kw, ksel);
        this->finish(t1, a.x[0], b.x[0], c.x[0], d.x[0], h.x[0]);
    }
};

// ---- phase A: once per nonce -------------------------------------------------
// K[i] + W[i] of hash 1's block 2 for i = 18..63, handed to put(i - 18, value).
// Only the version-independent members of jc are read (w16, w17, c18, c19, c31,
// c32): any instance of the job will do.
// put(i, value) receives i = round - 18 (so i = 0, 1 are the two register-carried words under KN_VS_F20).
template <unsigned KN, class Put>
OTD_HD void vs_schedule(const JobConsts& jc, uint32_t w3s, const Ops<KN>& op, Put&& put)
{
    typedef vw<1> W;
    W w[64];
    const W w3 = bc<1>(w3s);
    w[16] = bc<1>(jc.w16);
    w[17] = bc<1>(jc.w17);
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
    w[33] = sched(op, w[31], w[26], w[18], w[17]);
#pragma unroll
    for (int i = 34; i < 64; i++) w[i] = sched(op, w[i - 2], w[i - 7], w[i - 15], w[i - 16]);
#pragma unroll
    for (int i = 18; i < 64; i++) put(i - 18, w[i].x[0] + Kc(i));
}

// ---- phase B: once per (nonce, version instance) -------------------------------
// e after round 60 of hash 2 for the instance whose midstate-dependent constants
// are jv (mid, a3..g3, ca3, ce3, k4..k6); js supplies the version-independent
// ones (kw16, kw17) and may be any instance of the job; get(i) returns what
// vs_schedule put at i.
template <unsigned KN, class Get>
OTD_HD uint32_t vs_hash(const JobConsts& jv, const JobConsts& js, uint32_t w3s, const VsOps<KN>& op, Get&& get)
{
    typedef vw<1> W;
    const W w3 = bc<1>(w3s);
    W a, b, c, d, e, f, g, h;
    // state after round 3 (rounds 0..2 folded per instance by K2 / the host)
    a = op.vadd(w3, jv.ca3); b = bc<1>(jv.a3); c = bc<1>(jv.b3); d = bc<1>(jv.c3);
    e = op.vadd(w3, jv.ce3); f = bc<1>(jv.e3); g = bc<1>(jv.f3); h = bc<1>(jv.g3);
    {
        W t1 = op.vadd(padd(op, vS1(op, e), ch(e, f, g)), jv.k4);                       // round 4
        W e5 = op.vadd(t1, jv.c3);
        W a5 = padd(op, padd(op, t1, vS0(op, a)), maj(a, b, c));
        t1 = op.vadd(padd(op, vS1(op, e5), ch(e5, e, f)), jv.k5);                       // round 5
        W e6 = op.vadd(t1, jv.b3);
        W a6 = padd(op, padd(op, t1, vS0(op, a5)), maj(a5, a, b));
        t1 = op.vadd(padd(op, vS1(op, e6), ch(e6, e5, e)), jv.k6);                      // round 6
        W e7 = op.vadd(t1, jv.a3);
        W a7 = padd(op, padd(op, t1, vS0(op, a6)), maj(a6, a5, a));
        W a4 = a, e4 = e;
        a = a7; b = a6; c = a5; d = a4; e = e7; f = e6; g = e5; h = e4;
    }
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
    op.rnd_c(h, a, b, c, d, e, f, g, js.kw16);
    op.rnd_c(g, h, a, b, c, d, e, f, js.kw17);
#define OTD_VS_R(i, a, b, c, d, e, f, g, h) op.rnd_r(a, b, c, d, e, f, g, h, get((i) - 18), Kc(i))
    OTD_VS_R(18, f, g, h, a, b, c, d, e);
    OTD_VS_R(19, e, f, g, h, a, b, c, d);
    OTD_VS_R(20, d, e, f, g, h, a, b, c);
    OTD_VS_R(21, c, d, e, f, g, h, a, b);
    OTD_VS_R(22, b, c, d, e, f, g, h, a);
    OTD_VS_R(23, a, b, c, d, e, f, g, h);
#define OTD_VS_R8(R)                        \
    OTD_VS_R((R) + 0, h, a, b, c, d, e, f, g); \
    OTD_VS_R((R) + 1, g, h, a, b, c, d, e, f); \
    OTD_VS_R((R) + 2, f, g, h, a, b, c, d, e); \
    OTD_VS_R((R) + 3, e, f, g, h, a, b, c, d); \
    OTD_VS_R((R) + 4, d, e, f, g, h, a, b, c); \
    OTD_VS_R((R) + 5, c, d, e, f, g, h, a, b); \
    OTD_VS_R((R) + 6, b, c, d, e, f, g, h, a); \
    OTD_VS_R((R) + 7, a, b, c, d, e, f, g, h);
    if (KN & KN_VS_LOOP1) {
        // the placement pattern of rounds 24..31 repeats in every trip (ksel must be a compile-time constant)
#define OTD_VS_RL(j, a, b, c, d, e, f, g, h) op.rnd_r(a, b, c, d, e, f, g, h, get(base - 18 + (j)), Kc(24 + (j)))
#pragma unroll 1
        for (int base = 24; base < 64; base += 8) {
            OTD_VS_RL(0, h, a, b, c, d, e, f, g);
            OTD_VS_RL(1, g, h, a, b, c, d, e, f);
            OTD_VS_RL(2, f, g, h, a, b, c, d, e);
            OTD_VS_RL(3, e, f, g, h, a, b, c, d);
            OTD_VS_RL(4, d, e, f, g, h, a, b, c);
            OTD_VS_RL(5, c, d, e, f, g, h, a, b);
            OTD_VS_RL(6, b, c, d, e, f, g, h, a);
            OTD_VS_RL(7, a, b, c, d, e, f, g, h);
        }
#undef OTD_VS_RL
    } else {
        OTD_VS_R8(24)
        OTD_VS_R8(32)
        OTD_VS_R8(40)
        OTD_VS_R8(48)
        OTD_VS_R8(56)
    }
#undef OTD_VS_R8
#undef OTD_VS_R
    // the role order now is (h,a,b,c,d,e,f,g) = (A,B,C,D,E,F,G,H); hash 2 on the digest (state + midstate)
    return hash2_e60<1, KN>(op, op.vadd(h, jv.mid[0]), op.vadd(a, jv.mid[1]), op.vadd(b, jv.mid[2]),
                            op.vadd(c, jv.mid[3]), op.vadd(d, jv.mid[4]), op.vadd(e, jv.mid[5]),
                            op.vadd(f, jv.mid[6]), op.vadd(g, jv.mid[7]))
        .x[0];
}

}  // namespace otd
