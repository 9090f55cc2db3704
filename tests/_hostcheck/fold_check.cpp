// Host instantiation of the product's scan_core.cuh / sha256_core.cuh so the
// constant folding and round bookkeeping can be checked on a CPU-only box
// (tests/test_fold_host.py).  Test helper only — not shipped, not a fallback.
#include "../../otedama_b200/csrc/scan_core.cuh"
using namespace otd;

extern "C" {
// e after round 60 of hash 2 via the unrolled folded path, V lanes (1,2,4).
int fc_scan_e60(const uint8_t hdr76[76], const uint32_t* nonces, int v, uint32_t* e60_out)
{
    uint8_t hdr[80];
    memcpy(hdr, hdr76, 76);
    memset(hdr + 76, 0, 4);
    uint32_t mid[8], tail[3], tgt[8] = {0};
    header_midstate(hdr, mid, tail);
    JobConsts jc;
    fold_job_consts(mid, tail[0], tail[1], tail[2], tgt, &jc);
    if (v == 1) {
        vw<1> w3; w3.x[0] = bswap32(nonces[0]);
        vw<1> r = scan_core<1>(jc, w3); e60_out[0] = r.x[0];
    } else if (v == 2) {
        vw<2> w3; for (int i = 0; i < 2; i++) w3.x[i] = bswap32(nonces[i]);
        vw<2> r = scan_core<2>(jc, w3); for (int i = 0; i < 2; i++) e60_out[i] = r.x[i];
    } else if (v == 4) {
        vw<4> w3; for (int i = 0; i < 4; i++) w3.x[i] = bswap32(nonces[i]);
        vw<4> r = scan_core<4>(jc, w3); for (int i = 0; i < 4; i++) e60_out[i] = r.x[i];
    } else return -1;
    return 0;
}
// Full digest through the looped midstate path (what the rare-hit path runs).
void fc_sha256d_mid(const uint8_t hdr76[76], uint32_t nonce, const uint8_t target[32],
                    uint8_t hash_out[32], int* le_target)
{
    uint8_t hdr[80];
    memcpy(hdr, hdr76, 76);
    memset(hdr + 76, 0, 4);
    uint32_t mid[8], tail[3], tgt[8];
    for (int i = 0; i < 8; i++)
        tgt[i] = (uint32_t)target[4*i] | (uint32_t)target[4*i+1] << 8 | (uint32_t)target[4*i+2] << 16 | (uint32_t)target[4*i+3] << 24;
    header_midstate(hdr, mid, tail);
    JobConsts jc;
    fold_job_consts(mid, tail[0], tail[1], tail[2], tgt, &jc);
    uint32_t dig[8];
    sha256d_from_consts(jc, nonce, dig);
    for (int i = 0; i < 8; i++) {
        hash_out[4*i] = dig[i] >> 24; hash_out[4*i+1] = dig[i] >> 16; hash_out[4*i+2] = dig[i] >> 8; hash_out[4*i+3] = dig[i];
    }
    *le_target = digest_le_target(dig, jc.target) ? 1 : 0;
}
uint32_t fc_iv7(void) { return IV7; }
}
