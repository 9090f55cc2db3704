// Host instantiation of the product's scan_core.cuh / sha256_core.cuh so the
// constant folding, round bookkeeping and every instruction-placement policy
// can be checked on a CPU-only box (tests/test_fold_host.py).
// Test helper only — not shipped, not a fallback.
#include "../../otedama_b200/csrc/scan_core.cuh"
#include "../../otedama_b200/csrc/scan_vs_core.cuh"
using namespace otd;

template <int V, unsigned KN>
static void run_core(const JobConsts& jc, const uint32_t* nonces, uint32_t* out)
{
    vw<V> w3;
    for (int i = 0; i < V; i++) w3.x[i] = bswap32(nonces[i]);
    const Ops<KN> op{make_pipe_mults()};
    vw<V> r = (KN & KN_COMPACT) ? scan_core_compact<V, KN>(jc, w3, op) : scan_core<V, KN>(jc, w3, op);
    for (int i = 0; i < V; i++) out[i] = r.x[i];
}
template <unsigned KN>
static int run_v(const JobConsts& jc, const uint32_t* nonces, int v, uint32_t* out)
{
    if (v == 1) run_core<1, KN>(jc, nonces, out);
    else if (v == 2) run_core<2, KN>(jc, nonces, out);
    else if (v == 4) run_core<4, KN>(jc, nonces, out);
    else return -1;
    return 0;
}

static void consts_for(const uint8_t hdr76[76], const uint8_t target[32], JobConsts* jc)
{
    uint8_t hdr[80];
    memcpy(hdr, hdr76, 76);
    memset(hdr + 76, 0, 4);
    uint32_t mid[8], tail[3], tgt[8] = {0};
    if (target)
        for (int i = 0; i < 8; i++)
            tgt[i] = (uint32_t)target[4 * i] | (uint32_t)target[4 * i + 1] << 8 |
                     (uint32_t)target[4 * i + 2] << 16 | (uint32_t)target[4 * i + 3] << 24;
    header_midstate(hdr, mid, tail);
    fold_job_consts(mid, tail[0], tail[1], tail[2], tgt, jc);
}

// Version-shared core: the schedule of block 2 expanded once (from instance 0's constants),
// then one state recurrence + hash 2 per version instance.
template <unsigned KN>
static void run_vs(const JobConsts* jcs, int n_ver, uint32_t nonce, uint32_t* out)
{
    VsOps<KN> op{{make_pipe_mults()}};
    uint32_t kw[46];
    const uint32_t w3 = bswap32(nonce);
    vs_schedule<KN>(jcs[0], w3, op, [&](int i, uint32_t v) { kw[i] = v; });
    for (int v = 0; v < n_ver; v++) out[v] = vs_hash<KN>(jcs[v], jcs[0], w3, op, [&](int i) { return kw[i]; });
}

extern "C" {
// e after round 60 of hash 2 for n_ver headers that differ only in their version field
// (bytes 0..3 of hdr76 replaced by versions[v], little-endian), through the version-shared core.
int fc_vs_e60_kn(const uint8_t hdr76[76], const uint32_t* versions, int n_ver, uint32_t nonce, unsigned knobs,
                 uint32_t* e60_out)
{
    if (n_ver < 1 || n_ver > 64) return -1;
    JobConsts jcs[64];
    for (int v = 0; v < n_ver; v++) {
        uint8_t h[76];
        memcpy(h, hdr76, 76);
        h[0] = versions[v]; h[1] = versions[v] >> 8; h[2] = versions[v] >> 16; h[3] = versions[v] >> 24;
        consts_for(h, nullptr, &jcs[v]);
    }
#define VSSET(n, m) (KN_E | KN_W1 | KN_A | KN_W2 | KN_SP | KN_T1Z | OTD_KN_Z_EIGHTHS(n) | OTD_KN_VSZ_EIGHTHS(m))
    switch (knobs) {
        case 0: run_vs<0>(jcs, n_ver, nonce, e60_out); return 0;
        case KN_E | KN_W1 | KN_A | KN_W2 | KN_T1X | KN_SP:
            run_vs<KN_E | KN_W1 | KN_A | KN_W2 | KN_T1X | KN_SP>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(6, 0): run_vs<VSSET(6, 0)>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(6, 6): run_vs<VSSET(6, 6)>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(6, 4): run_vs<VSSET(6, 4)>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(6, 7): run_vs<VSSET(6, 7)>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(5, 6): run_vs<VSSET(5, 6)>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(7, 7): run_vs<VSSET(7, 7)>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(8, 0): run_vs<VSSET(8, 0)>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(7, 0): run_vs<VSSET(7, 0)>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(7, 0) | OTD_KN_ZSEL(5): run_vs<VSSET(7, 0) | OTD_KN_ZSEL(5)>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(7, 0) | KN_VS_CFMA: run_vs<VSSET(7, 0) | KN_VS_CFMA>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(8, 7): run_vs<VSSET(8, 7)>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(7, 6): run_vs<VSSET(7, 6)>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(5, 7): run_vs<VSSET(5, 7)>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(7, 5): run_vs<VSSET(7, 5)>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(7, 7) | KN_VS_F20: run_vs<VSSET(7, 7) | KN_VS_F20>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(6, 6) | KN_VS_F20: run_vs<VSSET(6, 6) | KN_VS_F20>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(7, 7) | KN_VS_LOOP1: run_vs<VSSET(7, 7) | KN_VS_LOOP1>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(7, 0) | KN_VS_LOOP1: run_vs<VSSET(7, 0) | KN_VS_LOOP1>(jcs, n_ver, nonce, e60_out); return 0;
        case VSSET(7, 7) | KN_VS_LOOP1 | KN_VS_F20:
            run_vs<VSSET(7, 7) | KN_VS_LOOP1 | KN_VS_F20>(jcs, n_ver, nonce, e60_out); return 0;
        case 1023: run_vs<1023>(jcs, n_ver, nonce, e60_out); return 0;
        default: return -2;
    }
#undef VSSET
}
// e after round 60 of hash 2 via the unrolled folded path, V lanes (1, 2, 4),
// under one of the knob sets below (the device variants use a subset of these).
int fc_scan_e60_kn(const uint8_t hdr76[76], const uint32_t* nonces, int v, unsigned knobs, uint32_t* e60_out)
{
    JobConsts jc;
    consts_for(hdr76, nullptr, &jc);
    switch (knobs) {
        case 0: return run_v<0>(jc, nonces, v, e60_out);
        case 1023: return run_v<1023>(jc, nonces, v, e60_out);
        case KN_E | KN_W1 | KN_SSHR: return run_v<KN_E | KN_W1 | KN_SSHR>(jc, nonces, v, e60_out);
        case KN_E | KN_W1 | KN_SSHR | KN_HKW | KN_BS1:
            return run_v<KN_E | KN_W1 | KN_SSHR | KN_HKW | KN_BS1>(jc, nonces, v, e60_out);
        case KN_E | KN_W1 | KN_SSHR | KN_HKW | KN_BS1 | KN_BS0 | KN_SROT:
            return run_v<KN_E | KN_W1 | KN_SSHR | KN_HKW | KN_BS1 | KN_BS0 | KN_SROT>(jc, nonces, v, e60_out);
        case KN_T1 | KN_A | KN_W2: return run_v<KN_T1 | KN_A | KN_W2>(jc, nonces, v, e60_out);
        case 63: return run_v<63>(jc, nonces, v, e60_out);
        case KN_E | KN_W1 | KN_A: return run_v<KN_E | KN_W1 | KN_A>(jc, nonces, v, e60_out);
        case KN_E | KN_W1 | KN_A | KN_T1X: return run_v<KN_E | KN_W1 | KN_A | KN_T1X>(jc, nonces, v, e60_out);
        case KN_E | KN_W1 | KN_A | KN_T1X | KN_W2: return run_v<KN_E | KN_W1 | KN_A | KN_T1X | KN_W2>(jc, nonces, v, e60_out);
        case KN_E | KN_W1 | KN_T1X: return run_v<KN_E | KN_W1 | KN_T1X>(jc, nonces, v, e60_out);
        case KN_E | KN_W1 | KN_A | KN_W2: return run_v<KN_E | KN_W1 | KN_A | KN_W2>(jc, nonces, v, e60_out);
        case 2047: return run_v<2047>(jc, nonces, v, e60_out);
        case KN_COMPACT: return run_v<KN_COMPACT>(jc, nonces, v, e60_out);
#define ZSET(n) (KN_E | KN_W1 | KN_A | KN_W2 | KN_SP | KN_T1Z | OTD_KN_Z_EIGHTHS(n))
        case ZSET(1): return run_v<ZSET(1)>(jc, nonces, v, e60_out);
        case ZSET(5): return run_v<ZSET(5)>(jc, nonces, v, e60_out);
        case ZSET(6): return run_v<ZSET(6)>(jc, nonces, v, e60_out);
        case ZSET(8): return run_v<ZSET(8)>(jc, nonces, v, e60_out);
        case KN_COMPACT | ZSET(6): return run_v<KN_COMPACT | ZSET(6)>(jc, nonces, v, e60_out);
        case ZSET(6) & ~KN_E: return run_v<ZSET(6) & ~KN_E>(jc, nonces, v, e60_out);
        case ZSET(6) & ~KN_A: return run_v<ZSET(6) & ~KN_A>(jc, nonces, v, e60_out);
        case ZSET(6) & ~KN_W2: return run_v<ZSET(6) & ~KN_W2>(jc, nonces, v, e60_out);
        case KN_E | KN_W1 | KN_A | KN_W2 | KN_T1X | KN_SP: return run_v<KN_E | KN_W1 | KN_A | KN_W2 | KN_T1X | KN_SP>(jc, nonces, v, e60_out);
        case KN_COMPACT | KN_E | KN_W1 | KN_A | KN_W2 | KN_T1X | KN_SP: return run_v<KN_COMPACT | KN_E | KN_W1 | KN_A | KN_W2 | KN_T1X | KN_SP>(jc, nonces, v, e60_out);
        case KN_COMPACT | KN_E | KN_W1 | KN_A | KN_W2 | KN_T1X: return run_v<KN_COMPACT | KN_E | KN_W1 | KN_A | KN_W2 | KN_T1X>(jc, nonces, v, e60_out);
        case KN_COMPACT | KN_E | KN_W1 | KN_A | KN_W2 | KN_T1: return run_v<KN_COMPACT | KN_E | KN_W1 | KN_A | KN_W2 | KN_T1>(jc, nonces, v, e60_out);
        case KN_COMPACT | KN_E | KN_W1 | KN_A | KN_W2: return run_v<KN_COMPACT | KN_E | KN_W1 | KN_A | KN_W2>(jc, nonces, v, e60_out);
        case KN_COMPACT | KN_E | KN_W1 | KN_A: return run_v<KN_COMPACT | KN_E | KN_W1 | KN_A>(jc, nonces, v, e60_out);
        case KN_E | KN_W1 | KN_A | KN_W2 | KN_T1: return run_v<KN_E | KN_W1 | KN_A | KN_W2 | KN_T1>(jc, nonces, v, e60_out);
        case KN_COMPACT | 1023u | KN_T1X: return run_v<KN_COMPACT | 1023u | KN_T1X>(jc, nonces, v, e60_out);
        default: return -2;
    }
}
// Full digest through the looped midstate path (what the rare-hit path runs).
void fc_sha256d_mid(const uint8_t hdr76[76], uint32_t nonce, const uint8_t target[32],
                    uint8_t hash_out[32], int* le_target)
{
    JobConsts jc;
    consts_for(hdr76, target, &jc);
    uint32_t dig[8];
    sha256d_from_consts(jc, nonce, dig);
    for (int i = 0; i < 8; i++) {
        hash_out[4 * i] = dig[i] >> 24; hash_out[4 * i + 1] = dig[i] >> 16;
        hash_out[4 * i + 2] = dig[i] >> 8; hash_out[4 * i + 3] = dig[i];
    }
    *le_target = digest_le_target(dig, jc.target) ? 1 : 0;
}
uint32_t fc_iv7(void) { return IV7; }
}
