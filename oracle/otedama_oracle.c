/*
 * otedama_oracle.c — CPU restatement of the reference's SHA-256d search path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under otedama_b200/ may include, link or
 * load this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker (or as
 * the timed CPU baseline), never as the product path.
 *
 * Parity status: PINNED for the core predicate.  The reference
 * (shizukutanaka/Otedama, pure Go) delegates the arithmetic to the Go standard
 * library crypto/sha256 (toolchain go1.24.0, go.mod:3-5), which is not in the
 * reference tree and cannot be built here (no Go toolchain).  crypto/sha256 is
 * FIPS 180-4 SHA-256; this file restates FIPS 180-4 and is checked by
 * tests/test_oracle.py against every known-answer vector the reference's own
 * tests hold for this path:
 *   - genesis header -> 6fe28c0a...0000000000   internal/miner/sha256d_test.go:24-59
 *                                               internal/btccrypto/btccrypto_extras_test.go:56-76
 *   - SHA256d("")    -> 5df6e0e2...5d4c9456     internal/miner/sha256d_test.go:61-70
 *   - header layout / nonce at offset 76        internal/miner/sha256d_test.go:74-117
 *   - LessOrEqual ordering                      internal/miner/sha256d_test.go:277-301
 * (target math lives in oracle/oracle.py: it needs big integers.)
 *
 * UNPINNED sub-paths (the reference has neither code nor tests for them; they
 * follow the public Stratum V1 / BIP-320 conventions):
 *   orc_v1_merkle_root, orc_extranonce2_bytes, orc_bip320_version.
 * orc_v1_merkle_root is additionally anchored to public chain data that the
 * tests re-derive with hashlib (genesis coinbase -> genesis merkle root; block
 * 170: coinbase -> txid -> merkle fold -> header -> published block hash; the
 * Stratum V1 documentation's worked example: notify + submit -> a hash with 34
 * leading zero bits).
 *
 * Two compression back-ends are provided because Go's crypto/sha256
 * dispatches to SHA-NI on amd64 when the CPU has it: a portable scalar
 * routine and an x86 SHA-NI routine.  Both give identical results (tested).
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <stdlib.h>
#include <pthread.h>

#if defined(__x86_64__)
#include <immintrin.h>
#include <cpuid.h>
#define ORC_X86 1
#else
#define ORC_X86 0
#endif

/* ------------------------------------------------------------------ */
/* SHA-256 (FIPS 180-4) — what Go's crypto/sha256.Sum256 computes.     */
/* Call sites in the reference: internal/miner/sha256d.go:107-110      */
/* ------------------------------------------------------------------ */

static const uint32_t K256[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
    0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
    0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
    0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
    0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
    0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
    0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};

static const uint32_t IV256[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a,
                                  0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};

static inline uint32_t rotr32(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
static inline uint32_t load_be32(const uint8_t *p)
{
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3];
}
static inline void store_be32(uint8_t *p, uint32_t v)
{
    p[0] = (uint8_t)(v >> 24); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t)v;
}
static inline void store_le32(uint8_t *p, uint32_t v)
{
    p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24);
}

static void compress_scalar(uint32_t st[8], const uint8_t blk[64])
{
    uint32_t w[64];
    for (int i = 0; i < 16; i++) w[i] = load_be32(blk + 4 * i);
    for (int i = 16; i < 64; i++) {
        uint32_t s0 = rotr32(w[i - 15], 7) ^ rotr32(w[i - 15], 18) ^ (w[i - 15] >> 3);
        uint32_t s1 = rotr32(w[i - 2], 17) ^ rotr32(w[i - 2], 19) ^ (w[i - 2] >> 10);
        w[i] = w[i - 16] + s0 + w[i - 7] + s1;
    }
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
    for (int i = 0; i < 64; i++) {
        uint32_t S1 = rotr32(e, 6) ^ rotr32(e, 11) ^ rotr32(e, 25);
        uint32_t ch = (e & f) ^ (~e & g);
        uint32_t t1 = h + S1 + ch + K256[i] + w[i];
        uint32_t S0 = rotr32(a, 2) ^ rotr32(a, 13) ^ rotr32(a, 22);
        uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
        uint32_t t2 = S0 + mj;
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

#if ORC_X86
__attribute__((target("sha,sse4.1,ssse3")))
static void compress_shani(uint32_t st[8], const uint8_t blk[64])
{
    /* Standard SHA-NI block function (two-lane ABEF/CDGH state layout). */
    const __m128i MASK = _mm_set_epi64x(0x0c0d0e0f08090a0bULL, 0x0405060700010203ULL);
    __m128i TMP = _mm_loadu_si128((const __m128i *)&st[0]);
    __m128i STATE1 = _mm_loadu_si128((const __m128i *)&st[4]);
    TMP = _mm_shuffle_epi32(TMP, 0xB1);          /* CDAB */
    STATE1 = _mm_shuffle_epi32(STATE1, 0x1B);    /* EFGH */
    __m128i STATE0 = _mm_alignr_epi8(TMP, STATE1, 8); /* ABEF */
    STATE1 = _mm_blend_epi16(STATE1, TMP, 0xF0);      /* CDGH */
    const __m128i ABEF_SAVE = STATE0, CDGH_SAVE = STATE1;
    __m128i M[4];
    for (int i = 0; i < 4; i++)
        M[i] = _mm_shuffle_epi8(_mm_loadu_si128((const __m128i *)(blk + 16 * i)), MASK);
    for (int r = 0; r < 16; r++) {
        __m128i cur = M[r & 3];
        __m128i MSG = _mm_add_epi32(cur, _mm_loadu_si128((const __m128i *)&K256[4 * r]));
        STATE1 = _mm_sha256rnds2_epu32(STATE1, STATE0, MSG);
        if (r >= 3 && r < 15) {
            /* finish the schedule word group that will be consumed at r+1 */
            __m128i t = _mm_alignr_epi8(cur, M[(r + 3) & 3], 4);
            M[(r + 1) & 3] = _mm_add_epi32(M[(r + 1) & 3], t);
            M[(r + 1) & 3] = _mm_sha256msg2_epu32(M[(r + 1) & 3], cur);
        }
        MSG = _mm_shuffle_epi32(MSG, 0x0E);
        STATE0 = _mm_sha256rnds2_epu32(STATE0, STATE1, MSG);
        if (r >= 1 && r < 13)
            M[(r + 3) & 3] = _mm_sha256msg1_epu32(M[(r + 3) & 3], cur);
    }
    STATE0 = _mm_add_epi32(STATE0, ABEF_SAVE);
    STATE1 = _mm_add_epi32(STATE1, CDGH_SAVE);
    TMP = _mm_shuffle_epi32(STATE0, 0x1B);       /* FEBA */
    STATE1 = _mm_shuffle_epi32(STATE1, 0xB1);    /* DCHG */
    STATE0 = _mm_blend_epi16(TMP, STATE1, 0xF0); /* DCBA */
    STATE1 = _mm_alignr_epi8(STATE1, TMP, 8);    /* ABEF -> HGFE */
    _mm_storeu_si128((__m128i *)&st[0], STATE0);
    _mm_storeu_si128((__m128i *)&st[4], STATE1);
}
#endif

typedef void (*compress_fn)(uint32_t st[8], const uint8_t blk[64]);
static compress_fn g_compress = compress_scalar;
static int g_backend = 0; /* 0 scalar, 1 sha-ni */

/* orc_cpu_has_shani: 1 when the CPU advertises SHA extensions. */
int orc_cpu_has_shani(void)
{
#if ORC_X86
    unsigned a, b, c, d;
    if (__get_cpuid_count(7, 0, &a, &b, &c, &d)) return (b >> 29) & 1;
#endif
    return 0;
}

/* orc_set_backend: 0 = portable scalar, 1 = SHA-NI (if available), -1 = best.
 * Returns the back-end now in use. */
int orc_set_backend(int which)
{
    int want = which;
    if (want < 0) want = orc_cpu_has_shani();
#if ORC_X86
    if (want == 1 && orc_cpu_has_shani()) { g_compress = compress_shani; g_backend = 1; return 1; }
#endif
    g_compress = compress_scalar; g_backend = 0;
    return 0;
}
int orc_backend(void) { return g_backend; }

/* orc_sha256: one-shot SHA-256 of an arbitrary byte string. */
void orc_sha256(const uint8_t *data, size_t len, uint8_t out[32])
{
    uint32_t st[8];
    memcpy(st, IV256, sizeof st);
    size_t off = 0;
    for (; off + 64 <= len; off += 64) g_compress(st, data + off);
    uint8_t tail[128];
    size_t rem = len - off;
    memset(tail, 0, sizeof tail);
    memcpy(tail, data + off, rem);
    tail[rem] = 0x80;
    size_t tl = (rem + 9 <= 64) ? 64 : 128;
    uint64_t bits = (uint64_t)len * 8;
    for (int i = 0; i < 8; i++) tail[tl - 1 - i] = (uint8_t)(bits >> (8 * i));
    g_compress(st, tail);
    if (tl == 128) g_compress(st, tail + 64);
    for (int i = 0; i < 8; i++) store_be32(out + 4 * i, st[i]);
}

/* orc_sha256d — reference: SHA256d, internal/miner/sha256d.go:107-110
 * (same function as btccrypto.Hash256, internal/btccrypto/btccrypto.go:352-355). */
void orc_sha256d(const uint8_t *data, size_t len, uint8_t out[32])
{
    uint8_t first[32];
    orc_sha256(data, len, first);
    orc_sha256(first, 32, out);
}

/* ------------------------------------------------------------------ */
/* Header — reference: internal/miner/sha256d.go:49-80                 */
/* ------------------------------------------------------------------ */
typedef struct {
    uint32_t version;
    uint8_t prev_hash[32];
    uint8_t merkle_root[32];
    uint32_t time;
    uint32_t bits;
    uint32_t nonce;
} orc_header_t;

/* Header.Bytes(), sha256d.go:59-68 */
void orc_header_bytes(const orc_header_t *h, uint8_t out[80])
{
    store_le32(out + 0, h->version);
    memcpy(out + 4, h->prev_hash, 32);
    memcpy(out + 36, h->merkle_root, 32);
    store_le32(out + 68, h->time);
    store_le32(out + 72, h->bits);
    store_le32(out + 76, h->nonce);
}

/* ParseHeader, sha256d.go:71-80 */
void orc_parse_header(const uint8_t b[80], orc_header_t *h)
{
    h->version = (uint32_t)b[0] | (uint32_t)b[1] << 8 | (uint32_t)b[2] << 16 | (uint32_t)b[3] << 24;
    memcpy(h->prev_hash, b + 4, 32);
    memcpy(h->merkle_root, b + 36, 32);
    h->time = (uint32_t)b[68] | (uint32_t)b[69] << 8 | (uint32_t)b[70] << 16 | (uint32_t)b[71] << 24;
    h->bits = (uint32_t)b[72] | (uint32_t)b[73] << 8 | (uint32_t)b[74] << 16 | (uint32_t)b[75] << 24;
    h->nonce = (uint32_t)b[76] | (uint32_t)b[77] << 8 | (uint32_t)b[78] << 16 | (uint32_t)b[79] << 24;
}

/* HashHeader, sha256d.go:114-117: serialise 80 bytes, then SHA256d —
 * three compressions per call, no midstate reuse, exactly like the reference. */
void orc_hash_header(const orc_header_t *h, uint8_t out[32])
{
    uint8_t b[80];
    orc_header_bytes(h, b);
    orc_sha256d(b, 80, out);
}

/* Hash.LessOrEqual, sha256d.go:93-103: bytes compared from index 31 down. */
int orc_less_or_equal(const uint8_t h[32], const uint8_t other[32])
{
    for (int i = 31; i >= 0; i--) {
        if (h[i] < other[i]) return 1;
        if (h[i] > other[i]) return 0;
    }
    return 1;
}

/* ------------------------------------------------------------------ */
/* grind — reference: Worker.grind, internal/miner/worker.go:216-282   */
/* ------------------------------------------------------------------ */
typedef struct {
    uint32_t nonce;
    uint32_t ntime;
    uint32_t version;
    uint8_t hash[32];
} orc_share_t;

typedef struct {
    const orc_header_t *tmpl;
    const uint8_t *target;
    uint32_t thread_id, step;
    uint64_t first, count;    /* window [first, first+count) of the 2^32 nonce space */
    orc_share_t *out;         /* per-thread output */
    size_t cap, n, dropped;
    uint64_t hashes;
    int midstate;             /* 0: reference path (3 compressions); 1: "strong" baseline */
} grind_arg_t;

static void *grind_thread(void *p)
{
    grind_arg_t *a = (grind_arg_t *)p;
    orc_header_t h = *a->tmpl;                    /* h := localWork.Header  (worker.go:250) */
    uint8_t hash[32];
    /* Thread i visits first+i, first+i+step, ... — one pass only: the
     * reference wraps uint32 and rescans forever (worker.go:279); the oracle
     * is the SET of qualifying nonces, so one pass is the whole answer. */
    uint32_t mid[8];
    uint8_t blk2[64];
    if (a->midstate) {
        uint8_t b[80];
        orc_header_bytes(&h, b);
        memcpy(mid, IV256, sizeof mid);
        g_compress(mid, b);
        memset(blk2, 0, 64);
        memcpy(blk2, b + 64, 12);
        blk2[16] = 0x80; blk2[62] = 0x02; blk2[63] = 0x80;
    }
    for (uint64_t off = a->thread_id; off < a->count; off += a->step) {
        uint32_t nonce = (uint32_t)(a->first + off);
        h.nonce = nonce;                          /* worker.go:252 */
        if (!a->midstate) {
            orc_hash_header(&h, hash);            /* worker.go:253 */
        } else {
            uint32_t st[8]; uint8_t first[64];
            memcpy(st, mid, sizeof st);
            store_le32(blk2 + 12, nonce);
            g_compress(st, blk2);
            memset(first, 0, 64);
            for (int i = 0; i < 8; i++) store_be32(first + 4 * i, st[i]);
            first[32] = 0x80; first[62] = 0x01; first[63] = 0x00;
            memcpy(st, IV256, sizeof st);
            g_compress(st, first);
            for (int i = 0; i < 8; i++) store_be32(hash + 4 * i, st[i]);
        }
        a->hashes++;                              /* worker.go:254 */
        if (orc_less_or_equal(hash, a->target)) { /* worker.go:256 */
            if (a->n < a->cap) {
                orc_share_t *s = &a->out[a->n++];
                s->nonce = nonce; s->ntime = h.time; s->version = h.version;
                memcpy(s->hash, hash, 32);
            } else {
                a->dropped++;                     /* worker.go:271-275 */
            }
        }
    }
    return NULL;
}

/*
 * orc_grind: restates one pass of Worker.grind over the nonce window
 * [first, first+count) with `threads` goroutine-equivalents striding by
 * `threads` (NonceStep defaults to Threads, worker.go:111-119).
 * Shares are returned sorted by nonce.  Returns the number written to `out`
 * (<= cap); *dropped gets the overflow, *hashes the candidates hashed.
 * midstate=0 is the reference's arithmetic (serialise + 3 compressions per
 * candidate); midstate=1 is the separately-labelled "strong" CPU baseline.
 */
static int share_cmp(const void *x, const void *y)
{
    uint32_t a = ((const orc_share_t *)x)->nonce, b = ((const orc_share_t *)y)->nonce;
    return (a > b) - (a < b);
}

size_t orc_grind(const orc_header_t *tmpl, const uint8_t target[32], uint64_t first, uint64_t count,
                 int threads, int midstate, orc_share_t *out, size_t cap, uint64_t *dropped,
                 uint64_t *hashes)
{
    if (threads < 1) threads = 1;
    pthread_t *tid = (pthread_t *)calloc((size_t)threads, sizeof *tid);
    grind_arg_t *args = (grind_arg_t *)calloc((size_t)threads, sizeof *args);
    for (int i = 0; i < threads; i++) {
        grind_arg_t *a = &args[i];
        a->tmpl = tmpl; a->target = target; a->thread_id = (uint32_t)i; a->step = (uint32_t)threads;
        a->first = first; a->count = count; a->cap = cap; a->midstate = midstate;
        a->out = (orc_share_t *)malloc((cap ? cap : 1) * sizeof(orc_share_t));
        if (threads == 1) grind_thread(a);
        else pthread_create(&tid[i], NULL, grind_thread, a);
    }
    size_t n = 0; uint64_t drop = 0, hs = 0;
    for (int i = 0; i < threads; i++) {
        if (threads > 1) pthread_join(tid[i], NULL);
        grind_arg_t *a = &args[i];
        for (size_t k = 0; k < a->n; k++) {
            if (n < cap) out[n++] = a->out[k]; else drop++;
        }
        drop += a->dropped; hs += a->hashes;
        free(a->out);
    }
    free(args); free(tid);
    qsort(out, n, sizeof *out, share_cmp);
    if (dropped) *dropped = drop;
    if (hashes) *hashes = hs;
    return n;
}

/* ------------------------------------------------------------------ */
/* UNPINNED: Stratum V1 coinbase -> merkle root, BIP-320 versions.      */
/* The reference discards these inputs (internal/poolproto/stratumv1/   */
/* parse.go:52-54,91); this follows the de-facto Stratum V1 rules.      */
/* ------------------------------------------------------------------ */

/* coinbase = coinb1 || extranonce1 || extranonce2 || coinb2;
 * root = SHA256d(coinbase); for each branch: root = SHA256d(root || branch). */
void orc_v1_merkle_root(const uint8_t *coinb1, size_t l1, const uint8_t *en1, size_t le1,
                        const uint8_t *en2, size_t le2, const uint8_t *coinb2, size_t l2,
                        const uint8_t *branches, size_t n_branch, uint8_t root[32])
{
    size_t total = l1 + le1 + le2 + l2;
    uint8_t *cb = (uint8_t *)malloc(total ? total : 1);
    memcpy(cb, coinb1, l1);
    memcpy(cb + l1, en1, le1);
    memcpy(cb + l1 + le1, en2, le2);
    memcpy(cb + l1 + le1 + le2, coinb2, l2);
    orc_sha256d(cb, total, root);
    free(cb);
    uint8_t cat[64];
    for (size_t i = 0; i < n_branch; i++) {
        memcpy(cat, root, 32);
        memcpy(cat + 32, branches + 32 * i, 32);
        orc_sha256d(cat, 64, root);
    }
}

/* extranonce2 value e -> its wire bytes: little-endian counter of `size` bytes. */
void orc_extranonce2_bytes(uint64_t e, size_t size, uint8_t *out)
{
    for (size_t i = 0; i < size; i++) out[i] = (i < 8) ? (uint8_t)(e >> (8 * i)) : 0;
}

/* BIP-320: k-th rolled version = base with the 16 mask bits (13..28 for the
 * standard mask 0x1fffe000) replaced by the bits of k scattered into the mask
 * from its least-significant set bit upward. */
uint32_t orc_bip320_version(uint32_t base, uint32_t mask, uint32_t k)
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

/* Batch helper for parity tests: SHA256d of n 80-byte headers. */
void orc_hash_headers(const uint8_t *hdrs, size_t n, uint8_t *out)
{
    for (size_t i = 0; i < n; i++) orc_sha256d(hdrs + 80 * i, 80, out + 32 * i);
}
