/* Plain-C (C99) consumer of include/otedama_b200.h — what a cgo build sees.
 * Host-only entry points are exercised for real; compute entry points are
 * expected to fail with OTD_ENODEV when no CUDA device is visible, and to work
 * when one is (argv[1] = "gpu").  Test helper, built by tests/test_c_abi.py. */
#include <stdio.h>
#include <string.h>

#include "otedama_b200.h"

#define CHECK(cond)                                                                  \
    do {                                                                             \
        if (!(cond)) {                                                               \
            printf("FAIL %s:%d %s (last error: %s)\n", __FILE__, __LINE__, #cond, otd_last_error()); \
            return 1;                                                                \
        }                                                                            \
    } while (0)

static const char* kGenesisHex =
    "0100000000000000000000000000000000000000000000000000000000000000000000003ba3edfd7a7b12b27ac72c3e"
    "67768f617fc81bc3888a51323a9fb8aa4b1e5e4a29ab5f49ffff001d1dac2b7c";

static int unhex(const char* s, uint8_t* out, size_t n)
{
    for (size_t i = 0; i < n; i++) {
        unsigned v;
        if (sscanf(s + 2 * i, "%2x", &v) != 1) return -1;
        out[i] = (uint8_t)v;
    }
    return 0;
}

int main(int argc, char** argv)
{
    const int want_gpu = argc > 1 && strcmp(argv[1], "gpu") == 0;
    CHECK(otd_abi_version() == OTD_ABI_VERSION);
    CHECK(otd_abi_struct_size(1) == (int)sizeof(otd_header_t) && sizeof(otd_header_t) == 80);
    CHECK(otd_abi_struct_size(2) == (int)sizeof(otd_work_t));
    CHECK(otd_abi_struct_size(3) == (int)sizeof(otd_share_t));
    CHECK(otd_abi_struct_size(5) == (int)sizeof(otd_stats_t));

    uint8_t raw[80], back[80], target[32], t2[32];
    CHECK(unhex(kGenesisHex, raw, 80) == 0);
    otd_header_t h;
    otd_parse_header(raw, &h);
    CHECK(h.version == 1 && h.bits == 0x1d00ffffu && h.nonce == 2083236893u);
    otd_header_bytes(&h, back);
    CHECK(memcmp(raw, back, 80) == 0);

    CHECK(otd_target_from_nbits(0x1d00ffffu, target) == OTD_OK);
    CHECK(target[27] == 0xff && target[26] == 0xff && target[31] == 0 && target[28] == 0 && target[25] == 0);
    CHECK(otd_target_from_difficulty(1.0, t2) == OTD_OK && memcmp(target, t2, 32) == 0);
    CHECK(otd_nbits_from_target(target) == 0x1d00ffffu);
    CHECK(otd_target_from_nbits(0x1d800000u, t2) == OTD_EINVAL && strstr(otd_last_error(), "negative mantissa"));
    CHECK(otd_target_from_difficulty(0.0, t2) == OTD_EINVAL);
    CHECK(otd_hash_less_or_equal(target, target) == 1);
    CHECK(otd_v1_job_target(0x1d00ffffu, 0.0, t2) == OTD_OK && memcmp(target, t2, 32) == 0);

    otd_work_t w;
    uint8_t zero[32] = {0};
    CHECK(otd_work_from_v2_job(42, 1, 0x20000004u, h.merkle_root, h.prev_hash, 0x1d00ffffu, 0x60000000u, zero, &w) == OTD_OK);
    CHECK(w.job_id == 42 && w.header.version == 0x20000004u && memcmp(w.target, target, 32) == 0 && w.nonce_count == 0);
    CHECK(otd_work_from_v1_job("not-a-number", 1, zero, 1, 0x1d00ffffu, 0.0, &w) == OTD_EINVAL);

    /* the Stratum documentation's example, from plain C: notify -> work unit, share -> submit fields */
    {
        otd_v1_notify_t nt;
        memset(&nt, 0, sizeof nt);
        nt.job_id = "bf";   /* the documentation's own (hex) job id: opaque to the library */
        nt.prevhash_hex = "4d16b6f85af6e2198f44ae2a6de67f78487ae5611b77c6c0440b921e00000000";
        nt.coinb1_hex = "01000000010000000000000000000000000000000000000000000000000000000000000000ffffffff20020862062f503253482f04b8864e5008";
        nt.coinb2_hex = "072f736c7573682f000000000100f2052a010000001976a914d23fcdf86f7e756a64a7a9688ef9903327048ed988ac00000000";
        nt.version_hex = "00000002";
        nt.nbits_hex = "1c2ac4af";
        nt.ntime_hex = "504e86b9";
        uint8_t scratch[256];
        size_t need = 0;
        otd_work_t vw;
        CHECK(otd_work_from_v1_notify(&nt, "08000002", 4, 1.0, 0, &vw, scratch, 8, &need) == OTD_ERANGE && need > 8 && need <= sizeof scratch);
        CHECK(otd_work_from_v1_notify(&nt, "08000002", 4, 1.0, 0, &vw, scratch, sizeof scratch, &need) == OTD_OK);
        CHECK(vw.job_id == 0 && strcmp(vw.job_id_str, "bf") == 0 && vw.header.version == 2 && vw.header.bits == 0x1c2ac4afu && vw.header.time == 0x504e86b9u);
        CHECK(vw.header.prev_hash[0] == 0xf8 && vw.header.prev_hash[3] == 0x4d && vw.header.prev_hash[31] == 0x00);
        CHECK(vw.coinb1_len == 58 && vw.extranonce1_len == 4 && vw.extranonce1[0] == 0x08 && vw.extranonce2_size == 4 && vw.n_branches == 0);
        otd_share_t sh;
        memset(&sh, 0, sizeof sh);
        memcpy(sh.job_id_str, vw.job_id_str, sizeof sh.job_id_str);
        sh.ntime = 0x504e86edu; sh.nonce = 0xb2957c02u; sh.version = 2; sh.extranonce2 = 0x01000000u;
        otd_v1_submit_t sub;
        CHECK(otd_share_to_v1_submit(&sh, vw.extranonce2_size, 0, &sub) == OTD_OK);
        CHECK(strcmp(sub.extranonce2_hex, "00000001") == 0 && strcmp(sub.ntime_hex, "504e86ed") == 0 &&
              strcmp(sub.nonce_hex, "b2957c02") == 0 && sub.version_bits_hex[0] == 0 && strcmp(sub.job_id, "bf") == 0);
        sh.job_id_str[0] = 0; sh.job_id = 191;
        CHECK(otd_share_to_v1_submit(&sh, vw.extranonce2_size, 0, &sub) == OTD_OK && strcmp(sub.job_id, "191") == 0);
        CHECK(otd_share_to_v1_submit(&sh, 17, 0, &sub) == OTD_EINVAL);
    }

    const int ndev = otd_device_count();
    CHECK(ndev >= 0);
    memset(&w, 0, sizeof w);
    w.header = h;
    w.nbits = h.bits;
    memcpy(w.target, target, 32);
    w.nonce_begin = 2083236893u - 1000u;
    w.nonce_count = 2000;
    otd_share_t shares[4];
    uint64_t found = 0, hashes = 0;
    const int n = otd_scan(0, &w, -1, shares, 4, &found, &hashes, NULL, NULL);
    if (ndev == 0) {
        CHECK(!want_gpu);
        CHECK(n == OTD_ENODEV && strstr(otd_last_error(), "no CPU fallback"));
        CHECK(otd_worker_create(0, NULL) == NULL);
        printf("OK (no device: compute entry points refuse)\n");
        return 0;
    }
    CHECK(n == 1 && found == 1 && hashes == 2000 && shares[0].nonce == 2083236893u);
    CHECK(shares[0].hash[31] == 0 && shares[0].hash[0] == 0x6f);
    otd_worker_t* wk = otd_worker_create(0, NULL);
    CHECK(wk != NULL);
    CHECK(otd_worker_has_work(wk) == 0);
    CHECK(otd_worker_set_work(wk, &w) == 1);
    CHECK(otd_worker_wait_exhausted(wk, 10000) == 1);
    CHECK(otd_worker_poll_shares(wk, shares, 4, 0) == 1 && shares[0].nonce == 2083236893u);
    otd_stats_t st;
    CHECK(otd_worker_stats(wk, &st) == OTD_OK && st.hashes_total == 2000 && st.shares_found == 1 && st.exhausted == 1);
    CHECK(otd_worker_destroy(wk) == OTD_OK);
    printf("OK (device present: scan + worker through plain C)\n");
    return 0;
}
