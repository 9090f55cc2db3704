"""Parity tests proper: the CUDA path, called through the C ABI, against the CPU
oracle on the same seeded inputs.  Bit-exact (integer work): share sets must be
identical — nonce, hash, ntime, version, extranonce2."""
import hashlib
import os
import random
import struct
import time

import pytest

from otedama_b200 import workloads as W

pytestmark = pytest.mark.gpu

NCPU = os.cpu_count() or 1
# The genesis header has TWO difficulty-1 solutions in its 2^32 nonce space: the mined one and
# 3197545707 (hash 00000000ea532922...); the CPU oracle's full scan found both
# (tests/golden/genesis_full_scan.json).
WANT_GENESIS_NONCES = [2083236893, 3197545707]


@pytest.fixture(scope="module")
def miner(product):
    from otedama_b200 import capi, miner as m
    if capi.lib().otd_device_count() <= 0:
        pytest.fail("no CUDA device: the gpu-marked tests must run on the B200 box")
    return m


def mheader(m, f):
    return m.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0)


def oheader(o, f, version=None, merkle_root=None):
    return o.Header(f["version"] if version is None else version, f["prev_hash"],
                    f["merkle_root"] if merkle_root is None else merkle_root, f["time"], f["bits"], 0)


def key(shares):
    return sorted((s.extranonce2, s.version, s.nonce, s.hash, s.ntime) for s in shares)


def okey(shares, extranonce2=0):
    return sorted((extranonce2, s.version, s.nonce, s.hash, s.ntime) for s in shares)


# ---------------------------------------------------------------- HashHeader --
def test_device_enumeration(miner):
    from otedama_b200 import hal
    devs = hal.CudaDriver().enumerate()
    assert devs
    ident = devs[0].identity()
    ident.validate()
    assert ident.id == "gpu-cuda0" and ident.family == "gpu" and ident.vendor == "nvidia"
    caps = devs[0].capabilities()
    assert caps.sha256d and caps.general_compute  # unlike the sysfs driver (gpu_linux_test.go:152-174)
    assert devs[0].info.sm_count >= 100 and devs[0].info.cc_major == 10
    devs[0].shutdown(); devs[0].shutdown()


@pytest.mark.parametrize("n", [1, 2, 127, 128, 129, 5000])
def test_hash_headers_matches_oracle(miner, oracle, n):
    rnd = random.Random(n)
    hdrs = [rnd.randbytes(80) for _ in range(n)]
    hdrs[0] = bytes.fromhex(W.GENESIS_HEADER_HEX)
    got = miner.hash_headers(hdrs)
    want = oracle.hash_headers(b"".join(hdrs))
    assert b"".join(got) == want
    assert got[0].hex() == W.GENESIS_HASH_HEX
    assert miner.hash_headers([]) == []


def test_hash_headers_across_staging_chunks(miner, oracle):
    """The entry point pipelines chunks of 2^18 headers over two streams and two kept device buffer pairs: ragged
    multi-chunk batches (1, 2 and 3+ chunks, back to back so slots are reused dirty) against the oracle."""
    rnd = random.Random(77)
    for n in ((1 << 18) + 5, 3 * (1 << 18) + 1, (1 << 18) - 1, 2 << 18):
        blob = rnd.randbytes(80 * n)
        got = b"".join(miner.hash_headers([blob[80 * i:80 * i + 80] for i in range(n)]))
        assert got == oracle.hash_headers(blob), n


def test_hash_header_single(miner, oracle):
    h = miner.parse_header(bytes.fromhex(W.GENESIS_HEADER_HEX))
    assert miner.hash_header(h).hex() == W.GENESIS_HASH_HEX


# ----------------------------------------------------------------- K1: scan --
def variants(miner):
    return list(range(len(miner.scan_variants())))


def test_genesis_window_every_variant(miner, oracle):
    f = W.genesis_fields()
    t = miner.target_from_nbits(W.DIFF1_NBITS)
    for v in variants(miner):
        r = miner.scan(miner.Work(job_id=7, channel_id=3, header=mheader(miner, f), nbits=W.DIFF1_NBITS, target=t,
                                  nonce_begin=W.GENESIS_NONCE - (1 << 23), nonce_count=1 << 24), variant=v)
        assert r.hashes == 1 << 24 and r.found == 1
        s = r.shares[0]
        assert (s.nonce, s.hash.hex()) == (W.GENESIS_NONCE, W.GENESIS_HASH_HEX)
        assert (s.job_id, s.channel_id, s.ntime, s.version, s.device_id) == (7, 3, f["time"], f["version"], "gpu-cuda0")


def test_equality_counts_as_a_share(miner):
    """Hash.LessOrEqual returns true on equality (sha256d.go:102)."""
    f = W.genesis_fields()
    hv = int.from_bytes(bytes.fromhex(W.GENESIS_HASH_HEX), "little")
    for tv, want in ((hv, 1), (hv - 1, 0), (hv + 1, 1)):
        r = miner.scan(miner.Work(header=mheader(miner, f), target=tv.to_bytes(32, "little"),
                                  nonce_begin=W.GENESIS_NONCE - 1000, nonce_count=2000))
        assert r.found == want


def test_smallest_windows_and_the_zero_target(miner, oracle):
    """Windows of one nonce (also the very last nonce, 0xFFFFFFFF), tiny ragged windows around a known
    share, and the all-zero target, which nothing but an all-zero hash meets."""
    f = W.genesis_fields()
    t = miner.target_from_nbits(W.DIFF1_NBITS)
    h = mheader(miner, f)
    for v in variants(miner):
        one = miner.scan(miner.Work(header=h, target=t, nonce_begin=W.GENESIS_NONCE, nonce_count=1), variant=v)
        assert (one.hashes, [s.nonce for s in one.shares]) == (1, [W.GENESIS_NONCE])
        miss = miner.scan(miner.Work(header=h, target=t, nonce_begin=W.GENESIS_NONCE + 1, nonce_count=1), variant=v)
        assert (miss.hashes, miss.found) == (1, 0)
        for before, count in ((0, 2), (1, 2), (255, 256), (256, 257), (299, 300), (4095, 4097)):
            r = miner.scan(miner.Work(header=h, target=t, nonce_begin=W.GENESIS_NONCE - before, nonce_count=count),
                           variant=v)
            assert (r.hashes, [s.nonce for s in r.shares]) == (count, [W.GENESIS_NONCE])
            r = miner.scan(miner.Work(header=h, target=t, nonce_begin=W.GENESIS_NONCE - before - count,
                                      nonce_count=count), variant=v)       # window ends just before the share
            assert (r.hashes, r.found) == (count, 0)
    easy = b"\xff" * 32
    last = miner.scan(miner.Work(header=h, target=easy, nonce_begin=0xFFFFFFFF, nonce_count=1))
    hdr = oracle.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0xFFFFFFFF)
    assert [(s.nonce, s.hash) for s in last.shares] == [(0xFFFFFFFF, oracle.hash_header(hdr))]
    zero = miner.scan(miner.Work(header=h, target=b"\x00" * 32, nonce_begin=W.GENESIS_NONCE - (1 << 20), nonce_count=1 << 21))
    assert (zero.hashes, zero.found, zero.shares) == (1 << 21, 0, [])


def test_dense_windows_match_golden_and_oracle(miner, oracle, golden):
    for case in golden["synthetic_windows"]["cases"]:
        f = W.config2_fields(case["index"])
        t = bytes.fromhex(case["target_hex"])
        for v in variants(miner):
            r = miner.scan(miner.Work(header=mheader(miner, f), target=t, nonce_begin=case["first"],
                                      nonce_count=case["count"]), variant=v)
            assert [s.nonce for s in r.shares] == case["nonces"]
            assert hashlib.sha256(b"".join(s.hash for s in r.shares)).hexdigest() == case["hash_digest"]
            assert r.hashes == case["count"] and r.found == len(case["nonces"])


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_random_headers_random_targets(miner, oracle, seed):
    rnd = random.Random(seed)
    for _ in range(6):
        f = {"version": rnd.getrandbits(32), "prev_hash": rnd.randbytes(32), "merkle_root": rnd.randbytes(32),
             "time": rnd.getrandbits(32), "bits": rnd.getrandbits(32)}
        # targets from "every hash" down to ~2^-20, in both kernel specialisations' ranges
        t = (rnd.getrandbits(256) >> rnd.choice([0, 1, 8, 12, 16, 20])).to_bytes(32, "little")
        first = rnd.choice([0, rnd.getrandbits(32) & ~0xFFFFF, (1 << 32) - (1 << 18)])
        count = rnd.choice([1, 255, 4097, 100003, 1 << 18])
        count = min(count, (1 << 32) - first)
        want, _, _ = oracle.grind(oheader(oracle, f), t, first, count, threads=NCPU)
        r = miner.scan(miner.Work(header=mheader(miner, f), target=t, nonce_begin=first, nonce_count=count),
                       variant=rnd.randrange(len(miner.scan_variants())), max_shares=1 << 18)
        assert r.hashes == count
        assert key(r.shares) == okey(want)


def test_all_ff_target_overflows_the_ring_without_hanging(miner):
    """integration_test.go:129-132 uses an all-0xFF target: every nonce is a share.
    The device ring is bounded; the excess is counted, never blocks."""
    f = W.config2_fields(5)
    r = miner.scan(miner.Work(header=mheader(miner, f), target=b"\xff" * 32, nonce_count=1 << 20), max_shares=1 << 17)
    assert r.found == 1 << 20 and r.hashes == 1 << 20
    assert len(r.shares) == 1 << 16  # ring capacity of the synchronous engine
    assert len({s.nonce for s in r.shares}) == len(r.shares)


def test_all_ff_target_overflows_the_ring_in_a_version_shared_launch(miner, oracle):
    """The same through K1v: every (version, nonce) of a 4-version launch is a share; the bounded ring keeps
    65536 of them (distinct, correctly attributed to their version and re-hashing right), counts the rest."""
    c4 = W.config4_v2()
    r = miner.scan(miner.Work(header=mheader(miner, c4), target=b"\xff" * 32, version_mask=c4["mask"], version_count=4,
                              nonce_count=1 << 16), max_shares=1 << 17)
    assert r.launches == 1 and r.found == 4 << 16 and r.hashes == 4 << 16
    assert len(r.shares) == 1 << 16 and len({(s.version, s.nonce) for s in r.shares}) == len(r.shares)
    versions = {oracle.bip320_version(c4["version"], c4["mask"], k) for k in range(4)}
    assert {s.version for s in r.shares} <= versions
    for s in r.shares[::997]:
        h = oracle.Header(s.version, c4["prev_hash"], c4["merkle_root"], s.ntime, c4["bits"], s.nonce)
        assert oracle.hash_header(h) == s.hash


def test_full_nonce_range_genesis(miner, golden):
    """configs[0]: full 2^32 scan must find exactly nonce 2083236893 — the same
    set the CPU oracle's full scan produced (tests/golden/genesis_full_scan.json)."""
    f = W.genesis_fields()
    r = miner.scan(miner.Work(header=mheader(miner, f), nbits=W.DIFF1_NBITS,
                              target=miner.target_from_nbits(W.DIFF1_NBITS)))
    assert r.hashes == 1 << 32
    assert (W.GENESIS_NONCE, W.GENESIS_HASH_HEX) in [(s.nonce, s.hash.hex()) for s in r.shares]
    g = golden["genesis_full_scan"]
    assert [(s.nonce, s.hash.hex()) for s in r.shares] == [(x["nonce"], x["hash"]) for x in g["shares"]]


def test_config2_full_range_matches_the_cpu_full_scans(miner):
    """Completeness at FULL range beyond genesis: synthetic headers of configs[1], all 2^32
    nonces at difficulty 1, against the CPU oracle's own full scans (committed fixtures,
    tests/golden/config2_full_scan_header*.json) — one burst, one launch per header."""
    import glob
    import json
    paths = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "config2_full_scan_header*.json")))
    if not paths:
        pytest.skip("no config2_full_scan_header*.json fixture")
    docs = [json.load(open(p)) for p in paths]
    hdrs = [mheader(miner, W.config2_fields(g["index"])) for g in docs]
    for h, g in zip(hdrs, docs):
        assert h.bytes()[:76].hex() == g["header76_hex"] and g["hashes"] == 1 << 32
    r = miner.scan_headers(hdrs, bytes.fromhex(docs[0]["target_hex"]))
    assert r.hashes == len(docs) << 32
    for i, g in enumerate(docs):
        got = [(s.nonce, s.hash.hex()) for s in r.shares if s.job_id == i]
        assert got == [(x["nonce"], x["hash"]) for x in g["shares"]], g["index"]


def test_all_257_headers_stratified_completeness(miner, oracle):
    """configs[1], every header: a 2^20 window at difficulty 2^-16 (dense hits) in one burst,
    share sets identical to the CPU oracle's (SURVEY §8d stratified completeness)."""
    t = oracle.target_from_difficulty(2.0 ** -16)
    hdrs = [mheader(miner, W.config2_fields(i)) for i in range(W.CONFIG2_HEADERS)]
    first, count = 0x7C000000, 1 << 20
    r = miner.scan_headers(hdrs, t, first, count)
    assert r.launches == 257 and r.hashes == 257 << 20
    by_header = {}
    for s in r.shares:
        by_header.setdefault(s.job_id, []).append((s.nonce, s.hash))
    total = 0
    for i in range(W.CONFIG2_HEADERS):
        want, _, _ = oracle.grind(oheader(oracle, W.config2_fields(i)), t, first, count, threads=NCPU, midstate=True)
        assert by_header.get(i, []) == [(s.nonce, s.hash) for s in want], i
        total += len(want)
    assert total == r.found and total > 3000


def test_config2_headers_full_range_soundness(miner, oracle):
    """configs[1] (subset): full nonce range at difficulty 1 on synthetic headers.
    Soundness bit-exact: every share re-hashed by the oracle meets the target."""
    t = miner.target_from_difficulty(1.0)
    total = 0
    for i in (1, 2, 3, 4):
        f = W.config2_fields(i)
        r = miner.scan(miner.Work(header=mheader(miner, f), target=t))
        assert r.hashes == 1 << 32
        for s in r.shares:
            oh = oheader(oracle, f)
            oh.nonce = s.nonce
            assert oracle.hash_header(oh) == s.hash and oracle.less_or_equal(s.hash, t)
        total += r.found
    assert total <= 20  # Poisson mean 4


def test_scan_headers_burst_matches_oracle(miner, oracle, golden):
    """configs[1] shape: many headers in one burst; dense windows against the golden sets."""
    cases = golden["synthetic_windows"]["cases"]
    hdrs = [mheader(miner, W.config2_fields(c["index"])) for c in cases]
    t = bytes.fromhex(cases[0]["target_hex"])
    r = miner.scan_headers(hdrs, t, 0, 1 << 24)
    assert r.launches == len(cases) and r.hashes == len(cases) << 24
    for i, c in enumerate(cases):
        mine = [s for s in r.shares if s.job_id == i]
        assert [s.nonce for s in mine] == c["nonces"]
        assert hashlib.sha256(b"".join(s.hash for s in mine)).hexdigest() == c["hash_digest"]
        f = W.config2_fields(c["index"])
        assert all(s.ntime == f["time"] and s.version == f["version"] for s in mine)


# ------------------------------------------------- K2: BIP-320 version roll --
def test_version_rolling_matches_oracle(miner, oracle):
    from tests import pysha
    c4 = W.config4_v2()
    f = dict(c4)
    t = oracle.target_from_difficulty(2.0 ** -18)
    work = miner.Work(job_id=9, header=mheader(miner, f), target=t, version_mask=c4["mask"], version_begin=3,
                      version_count=40, version_stride=5, nonce_begin=0x10000000, nonce_count=1 << 16)
    inst = miner.prepare_instances(work, 0, 40)
    want = []
    for i in range(40):
        ver = oracle.bip320_version(c4["version"], c4["mask"], 3 + 5 * i)
        assert inst[i].version == ver and inst[i].merkle_root == c4["merkle_root"]
        oh = oheader(oracle, f, version=ver)
        assert inst[i].midstate == pysha.midstate(oh.bytes())
        sh, _, _ = oracle.grind(oh, t, 0x10000000, 1 << 16, threads=NCPU)
        want += okey(sh)
    r = miner.scan(work)
    assert r.hashes == 40 << 16 and r.launches == 1       # one version-shared launch of the 40 instances
    assert key(r.shares) == sorted(want) and len(want) > 100
    assert all(s.job_id == 9 for s in r.shares)
    work.per_version_launch = True                        # the plain kernel, one launch per version instance
    r1 = miner.scan(work)
    assert r1.hashes == 40 << 16 and r1.launches == 40
    assert key(r1.shares) == sorted(want)


def test_version_rolling_default_count_covers_the_mask(miner, oracle):
    # a 3-bit mask: 8 versions, stride 3 from 1 -> k = 1, 4, 7
    f = W.config2_fields(9)
    t = oracle.target_from_difficulty(2.0 ** -20)
    r = miner.scan(miner.Work(header=mheader(miner, f), target=t, version_mask=0x0000E000, version_begin=1,
                              version_stride=3, nonce_count=1 << 14))
    assert r.launches == 1 and r.hashes == 3 << 14       # one version-shared launch of the three instances
    want = []
    for k in (1, 4, 7):
        ver = oracle.bip320_version(f["version"], 0xE000, k)
        sh, _, _ = oracle.grind(oheader(oracle, f, version=ver), t, 0, 1 << 14)
        want += okey(sh)
    assert key(r.shares) == sorted(want)


def test_version_shared_variants_match_oracle(miner, oracle):
    """K1v (csrc/scan_vs_kernel.cuh): every version-shared variant, group sizes 2..64 incl. ragged tails
    (65 = 64 + a lone instance on the plain kernel), ragged nonce windows, against the CPU oracle."""
    from otedama_b200 import capi
    VMAX = capi.OTD_VS_MAX_INSTANCES
    c4 = W.config4_v2()
    t = oracle.target_from_difficulty(2.0 ** -20)
    names = miner.vs_variants()
    assert names and miner.default_vs_variant() == 0
    try:
        for vi, name in enumerate(names):
            miner.vs_select(vi)
            for n_ver, begin, stride, nb, nc in ((2, 0, 1, 0, 5000), (16, 7, 3, 0xFFFF0000, 1 << 16),
                                                 (33, 100, 1, 123, 3 * 1024 + 1), (VMAX + 1, 64000, 16, 1 << 31, 4096)):
                if vi and n_ver not in (33, VMAX + 1):
                    continue
                work = miner.Work(job_id=5, header=mheader(miner, c4), target=t, version_mask=c4["mask"],
                                  version_begin=begin, version_count=n_ver, version_stride=stride, nonce_begin=nb,
                                  nonce_count=nc)
                want = []
                for i in range(n_ver):
                    ver = oracle.bip320_version(c4["version"], c4["mask"], begin + stride * i)
                    sh, _, _ = oracle.grind(oheader(oracle, c4, version=ver), t, nb, nc, threads=NCPU)
                    want += okey(sh)
                r = miner.scan(work)
                assert r.hashes == n_ver * nc, (name, n_ver)
                assert r.launches == (n_ver + VMAX - 1) // VMAX
                assert key(r.shares) == sorted(want), (name, n_ver)
    finally:
        miner.vs_select(-1)
    # switched off process-wide: the plain kernel per version gives the same set
    miner.vs_select(-2)
    try:
        work = miner.Work(header=mheader(miner, c4), target=t, version_mask=c4["mask"], version_count=5, nonce_count=1 << 14)
        r = miner.scan(work)
        assert r.launches == 5
    finally:
        miner.vs_select(-1)
    r2 = miner.scan(work)
    assert r2.launches == 1 and key(r2.shares) == key(r.shares)


def test_version_shared_groups_do_not_straddle_extranonce2(miner, oracle):
    """extranonce2 x version: instances of different extranonce2 values have different merkle roots, so a
    version-shared launch may only hold instances of ONE extranonce2 value."""
    c3 = W.config3_v1()
    t = oracle.target_from_difficulty(2.0 ** -20)
    work, f = v1_work(miner, c3, target=t, extranonce2_begin=5, extranonce2_count=3, extranonce2_stride=2,
                      version_mask=0x00006000, nonce_count=1 << 14)            # 4 versions per extranonce2
    want = []
    for e in (5, 7, 9):
        root = oracle.v1_merkle_root(c3["coinb1"], c3["extranonce1"], oracle.extranonce2_bytes(e, c3["extranonce2_size"]),
                                     c3["coinb2"], c3["branches"])
        for k in range(4):
            ver = oracle.bip320_version(c3["version"], 0x6000, k)
            sh, _, _ = oracle.grind(oheader(oracle, f, version=ver, merkle_root=root), t, 0, 1 << 14, threads=NCPU)
            want += okey(sh, extranonce2=e)
    r = miner.scan(work)
    assert r.launches == 3 and r.hashes == 12 << 14
    assert key(r.shares) == sorted(want) and len(want) > 20


# --------------------------------------- K3: extranonce2 roll + merkle root --
def v1_work(miner, c3, **kw):
    f = {"version": c3["version"], "prev_hash": c3["prev_hash"], "merkle_root": b"\x00" * 32,
         "time": c3["time"], "bits": c3["bits"]}
    return miner.Work(header=mheader(miner, f), coinb1=c3["coinb1"], extranonce1=c3["extranonce1"],
                      coinb2=c3["coinb2"], merkle_branches=c3["branches"],
                      extranonce2_size=c3["extranonce2_size"], **kw), f


def test_extranonce2_roll_and_merkle_root(miner, oracle, golden):
    c3 = W.config3_v1()
    work, f = v1_work(miner, c3, target=b"\xff" * 32, extranonce2_begin=0, extranonce2_count=8)
    inst = miner.prepare_instances(work, 0, 8)
    assert [i.merkle_root.hex() for i in inst] == golden["synthetic_windows"]["config3_roots_e0_7"]
    assert [i.extranonce2 for i in inst] == list(range(8))
    # the genesis coinbase split as coinb1|en1|en2|coinb2 reproduces the genesis merkle root
    g = golden["reference_vectors"]["genesis_coinbase_tx"]
    tx = bytes.fromhex(g["tx_hex"])
    en2 = int.from_bytes(tx[54:58], "little")
    gw = miner.Work(target=b"\xff" * 32, coinb1=tx[:50], extranonce1=tx[50:54], coinb2=tx[58:], extranonce2_size=4,
                    extranonce2_begin=en2, extranonce2_count=1)
    assert miner.prepare_instances(gw, 0, 1)[0].merkle_root.hex() == g["merkle_root_hex"]


@pytest.mark.parametrize("l1,l2,size,nb", [(1, 0, 1, 0), (41, 9, 4, 1), (42, 9, 4, 2), (50, 9, 4, 3), (51, 9, 4, 0),
                                           (58, 107, 8, 12), (120, 200, 8, 5), (0, 300, 2, 1), (500, 700, 16, 4)])
def test_coinbase_lengths_and_padding_edges(miner, oracle, l1, l2, size, nb):
    """Coinbase lengths straddling the 55/56/64-byte padding boundaries, 0..12 branches."""
    rnd = random.Random(l1 * 1000 + l2)
    cb1, en1, cb2 = rnd.randbytes(l1), rnd.randbytes(4), rnd.randbytes(l2)
    br = [rnd.randbytes(32) for _ in range(nb)]
    begin = 0 if size < 2 else 250
    work = miner.Work(target=b"\xff" * 32, coinb1=cb1, extranonce1=en1, coinb2=cb2, merkle_branches=br,
                      extranonce2_size=size, extranonce2_begin=begin, extranonce2_count=12, extranonce2_stride=1)
    inst = miner.prepare_instances(work, 0, 12)
    for i, it in enumerate(inst):
        e = begin + i
        want = oracle.v1_merkle_root(cb1, en1, oracle.extranonce2_bytes(e, size), cb2, br)
        assert it.merkle_root == want and it.extranonce2 == e


def test_v1_scan_with_extranonce2_matches_oracle(miner, oracle):
    c3 = W.config3_v1()
    t = oracle.target_from_difficulty(2.0 ** -17)
    work, f = v1_work(miner, c3, job_id=60, target=t, extranonce2_begin=5, extranonce2_count=20,
                      extranonce2_stride=3, nonce_begin=0xFFFF0000, nonce_count=1 << 16)
    r = miner.scan(work)
    want = []
    for i in range(20):
        e = 5 + 3 * i
        root = oracle.v1_merkle_root(c3["coinb1"], c3["extranonce1"], oracle.extranonce2_bytes(e, 8), c3["coinb2"],
                                     c3["branches"])
        sh, _, _ = oracle.grind(oheader(oracle, f, merkle_root=root), t, 0xFFFF0000, 1 << 16, threads=NCPU)
        want += okey(sh, extranonce2=e)
    assert key(r.shares) == sorted(want) and len(want) > 20
    assert r.hashes == 20 << 16


def test_extranonce2_times_version_space(miner, oracle):
    """nonce x extranonce2 x BIP-320 version bits, more instances than one preparation batch."""
    c3 = W.config3_v1()
    t = oracle.target_from_difficulty(2.0 ** -22)
    work, f = v1_work(miner, c3, target=t, extranonce2_begin=1, extranonce2_count=30, version_mask=W.BIP320_MASK,
                      version_begin=0, version_count=10, nonce_count=1 << 12)
    r = miner.scan(work)
    # one version-shared launch per extranonce2 value (10 instances each), plus one where the 256-instance
    # preparation batch boundary cuts a group in two (instances 250..255 | 256..259)
    assert r.launches == 31 and r.hashes == 300 << 12
    want = []
    for e in range(1, 31):
        root = oracle.v1_merkle_root(c3["coinb1"], c3["extranonce1"], oracle.extranonce2_bytes(e, 8), c3["coinb2"],
                                     c3["branches"])
        for k in range(10):
            ver = oracle.bip320_version(c3["version"], W.BIP320_MASK, k)
            sh, _, _ = oracle.grind(oheader(oracle, f, version=ver, merkle_root=root), t, 0, 1 << 12)
            want += okey(sh, extranonce2=e)
    assert key(r.shares) == sorted(want) and len(want) > 50


# ------------------------------------------------------------------- Worker --
def easy_work(miner, job_id=1, i=11, **kw):
    f = W.config2_fields(i)
    return miner.Work(job_id=job_id, channel_id=2, header=mheader(miner, f), nbits=f["bits"],
                      target=kw.pop("target", b"\xff" * 32), **kw), f


def test_worker_start_twice_raises_and_stop_is_idempotent(miner):
    w = miner.Worker().start()          # TestWorker_StartTwicePanics (worker_test.go:56-69)
    with pytest.raises(RuntimeError):
        w.start()
    w.stop(); w.stop()                  # TestWorker_StartAndStop (worker_test.go:71-91)


def test_worker_idle_until_set_work_then_finds_shares(miner, oracle):
    with miner.Worker(miner.WorkerConfig(launch_log2=16)) as w:
        assert not w.has_work()
        time.sleep(0.05)
        assert w.stats().hashes_total == 0
        work, f = easy_work(miner, job_id=42, nonce_count=1 << 16)
        w.set_work(work)                # TestWorker_FindsSharesWithEasyTarget (worker_test.go:93-140)
        assert w.has_work()
        assert w.wait_exhausted(10000)
        shares = []
        while True:
            got = w.poll_shares(4096, 0)
            if not got:
                break
            shares += got
        st = w.stats()
        assert st.exhausted and st.hashes_total == 1 << 16 and st.shares_found == 1 << 16
        assert st.shares_dropped == st.shares_found - len(shares)
        assert st.hash_rate > 0 and st.uptime > 0 and st.launches == 1
        s = shares[0]
        assert (s.job_id, s.channel_id, s.ntime, s.version, s.device_id) == (42, 2, f["time"], f["version"], "gpu-cuda0")
        oh = oheader(oracle, f); oh.nonce = s.nonce
        assert oracle.hash_header(oh) == s.hash


def test_worker_job_switch_is_observed_via_job_id(miner):
    """TestWorker_SetWorkJobChange (worker_test.go:142-175)."""
    t = (1 << 236).to_bytes(32, "little")     # ~2^-20 of hashes: a steady trickle
    with miner.Worker(miner.WorkerConfig(launch_log2=24)) as w:
        w1, _ = easy_work(miner, job_id=1, target=t, repeat=True)
        w.set_work(w1)
        first = w.poll_shares(16, 5000)
        assert first and all(s.job_id == 1 for s in first)
        w2, _ = easy_work(miner, job_id=2, i=12, target=t, repeat=True)
        seq = w.set_work(w2)
        deadline = time.time() + 10
        seen2 = False
        while time.time() < deadline and not seen2:
            seen2 = any(s.job_id == 2 and s.work_seq == seq for s in w.poll_shares(256, 200))
        assert seen2
        h0 = w.stats().hashes_total
        time.sleep(0.2)
        assert w.stats().hashes_total > h0   # monotone while working (HashrateMonitor, engine/stats.go:412-457)
        w.set_work(None)                     # arbitration pause / curtailment (arbitrate.go:255-262)
        assert not w.has_work()
        time.sleep(0.3)
        h1 = w.stats().hashes_total
        time.sleep(0.2)
        assert w.stats().hashes_total == h1


def test_poll_reports_the_end_of_a_work_once_then_honours_its_timeout(miner):
    """otd_worker_poll_shares: the end of a work (space exhausted) wakes ONE empty poll; after that an idle
    worker's poll waits for its timeout instead of returning at once (a forwarding loop must not spin), and a new
    set_work arms it again."""
    with miner.Worker() as w:
        impossible = bytes(32)                                   # no hash is <= 0 here
        a, _ = easy_work(miner, job_id=1, target=impossible, nonce_count=1 << 16)
        w.set_work(a)
        assert w.wait_exhausted(5000)
        t0 = time.time()
        assert w.poll_shares(16, 2000) == []                     # the end of the work: returns at once
        assert time.time() - t0 < 0.5
        for _ in range(2):
            t0 = time.time()
            assert w.poll_shares(16, 300) == []                  # idle and already reported: waits
            assert time.time() - t0 >= 0.28
        w.set_work(a)
        t0 = time.time()
        assert w.poll_shares(16, 2000) == [] and time.time() - t0 < 0.5
        t0 = time.time()
        assert w.poll_shares(16, 200) == [] and time.time() - t0 >= 0.18


def test_worker_set_work_aborts_a_long_launch_quickly(miner):
    with miner.Worker(miner.WorkerConfig(launch_log2=32)) as w:
        hard = (1).to_bytes(32, "little")
        a, _ = easy_work(miner, job_id=1, target=hard, repeat=True)
        w.set_work(a)
        time.sleep(0.3)
        b, f = easy_work(miner, job_id=2, i=13, nonce_count=1 << 16)
        t0 = time.time()
        w.set_work(b)
        got = w.poll_shares(16, 5000)
        dt = time.time() - t0
        assert got and got[0].job_id == 2
        assert dt < 0.25, f"job switch took {dt:.3f}s"   # a full 2^32 launch would take ~0.4 s


def test_worker_survives_rapid_job_switching(miner, oracle):
    """Host-runtime stress: 150 set_work calls (new job / pause / resume) at random moments of
    the launches in flight.  No hang, no stale-job confusion: every share carries the job id
    and work_seq of a job that was actually set, and re-hashes correctly on the CPU."""
    rnd = random.Random(123)
    t = (1 << 238).to_bytes(32, "little")      # ~2^-18 of hashes
    jobs = {}
    with miner.Worker(miner.WorkerConfig(launch_log2=22)) as w:
        shares = []
        for k in range(150):
            r = rnd.random()
            if r < 0.2:
                seq = w.set_work(None)
            else:
                i = rnd.randrange(1, 200)
                f = W.config2_fields(i)
                seq = w.set_work(miner.Work(job_id=1000 + k, header=mheader(miner, f), target=t, repeat=rnd.random() < 0.5,
                                            nonce_begin=rnd.randrange(0, 1 << 31), nonce_count=1 << rnd.randrange(16, 26)))
                jobs[seq] = (1000 + k, i)
            time.sleep(rnd.choice([0, 0, 0.001, 0.004, 0.015]))
            shares += w.poll_shares(256, 0)
        w.set_work(None)
        time.sleep(0.05)
        shares += w.poll_shares(4096, 0)
        st = w.stats()
    assert st.hashes_total > 0 and st.shares_dropped == 0
    assert len(shares) > 50
    for s in shares:
        assert s.work_seq in jobs and jobs[s.work_seq][0] == s.job_id
    for s in rnd.sample(shares, 40):
        f = W.config2_fields(jobs[s.work_seq][1])
        oh = oheader(oracle, f)
        oh.nonce = s.nonce
        assert oracle.hash_header(oh) == s.hash and oracle.less_or_equal(s.hash, t)


def test_worker_queue_overflow_counts_drops(miner):
    """Non-blocking send with a drop counter (worker.go:271-275)."""
    with miner.Worker(miner.WorkerConfig(queue_capacity=100, launch_log2=16)) as w:
        work, _ = easy_work(miner, nonce_count=1 << 12)
        w.set_work(work)
        assert w.wait_exhausted(10000)
        st = w.stats()
        assert st.shares_found == 1 << 12 and st.shares_dropped == (1 << 12) - 100
        assert len(w.poll_shares(4096, 0)) == 100


def test_worker_full_range_at_difficulty_one(miner):
    """TestUpdateWork_PopulatesFullHeaderAndShareTarget (engine/run_test.go:356-396) in
    spirit: the share echoes JobID / Version / NTime of the work it came from."""
    f = W.genesis_fields()
    with miner.Worker() as w:
        w.set_work(miner.Work(job_id=5, channel_id=1, header=mheader(miner, f), nbits=f["bits"],
                              target=miner.target_from_nbits(f["bits"])))
        assert w.wait_exhausted(30000)
        sh = w.poll_shares(16, 0)
        assert (W.GENESIS_NONCE, W.GENESIS_HASH_HEX, 5, 1, f["time"]) in \
               [(s.nonce, s.hash.hex(), s.job_id, s.version, s.ntime) for s in sh]
        assert sorted(s.nonce for s in sh) == WANT_GENESIS_NONCES
        st = w.stats()
        assert st.hashes_total == 1 << 32 and st.hash_rate > 1e9


def test_worker_runs_rolled_work_like_the_synchronous_scan(miner, oracle):
    """extranonce2 x version x nonce through the asynchronous worker path (K3 -> K2 -> K1
    with two launches in flight) gives exactly the synchronous scan's share set."""
    c3 = W.config3_v1()
    t = oracle.target_from_difficulty(2.0 ** -20)
    work, f = v1_work(miner, c3, job_id=77, target=t, extranonce2_begin=2, extranonce2_count=9, extranonce2_stride=4,
                      version_mask=W.BIP320_MASK, version_begin=1, version_count=5, version_stride=3,
                      nonce_begin=0x80000000, nonce_count=1 << 14)
    want = miner.scan(work)
    with miner.Worker(miner.WorkerConfig(launch_log2=16)) as w:
        w.set_work(work)
        assert w.wait_exhausted(20000)
        got = []
        while True:
            part = w.poll_shares(4096, 0)
            if not part:
                break
            got += part
        st = w.stats()
    assert key(got) == key(want.shares) and len(got) > 30
    # 9 groups of 5 version instances; 2^16 candidates per launch -> 12 blocks of 1024 nonces x 5, then the rest
    assert st.hashes_total == want.hashes == 45 << 14 and st.launches == 18 and st.shared_launches == 18
    assert all(s.job_id == 77 for s in got)
    # and a few of them against the oracle, end to end
    for s in got[:5]:
        root = oracle.v1_merkle_root(c3["coinb1"], c3["extranonce1"], oracle.extranonce2_bytes(s.extranonce2, 8),
                                     c3["coinb2"], c3["branches"])
        oh = oracle.Header(s.version, c3["prev_hash"], root, s.ntime, c3["bits"], s.nonce)
        assert oracle.hash_header(oh) == s.hash and oracle.less_or_equal(s.hash, t)


def test_two_workers_on_one_device_do_not_disturb_each_other(miner, oracle):
    t = oracle.target_from_difficulty(2.0 ** -18)
    jobs = []
    for i in (21, 22):
        f = W.config2_fields(i)
        jobs.append(miner.Work(job_id=i, header=mheader(miner, f), target=t, nonce_count=1 << 22))
    want = [miner.scan(j) for j in jobs]
    ws = [miner.Worker(miner.WorkerConfig(launch_log2=18)).start() for _ in jobs]
    try:
        for w, j in zip(ws, jobs):
            w.set_work(j)
        for w, j, r in zip(ws, jobs, want):
            assert w.wait_exhausted(20000)
            got = w.poll_shares(4096, 0)
            assert key(got) == key(r.shares) and all(s.job_id == j.job_id for s in got) and len(got) > 5
    finally:
        for w in ws:
            w.stop()


def test_two_workers_partition_the_version_space(miner, oracle):
    """static partition, no communication — version index k -> k mod G.
    Runs both halves on device 0 when only one GPU is visible."""
    from otedama_b200 import capi
    ndev = capi.lib().otd_device_count()
    c4 = W.config4_v2()
    t = oracle.target_from_difficulty(2.0 ** -18)
    base = dict(header=mheader(miner, c4), target=t, version_mask=c4["mask"], nonce_count=1 << 14)
    whole = miner.scan(miner.Work(version_begin=0, version_count=32, **base))
    parts = []
    for g in range(2):
        parts += miner.scan(miner.Work(version_begin=g, version_count=16, version_stride=2, **base),
                            device=g % ndev).shares
    assert sorted((s.version, s.nonce, s.hash) for s in parts) == sorted((s.version, s.nonce, s.hash) for s in whole.shares)
    assert len(whole.shares) > 20


def test_device_telemetry_feeds_joules_per_terahash(miner):
    """Observability parity: power / clock via NVML while the device hashes."""
    from otedama_b200 import hal
    dev = hal.CudaDriver().enumerate()[0]
    f = W.config2_fields(30)
    with miner.Worker() as w:
        w.set_work(miner.Work(header=mheader(miner, f), target=miner.target_from_difficulty(1.0), repeat=True))
        time.sleep(1.0)
        power, mhz, temp = dev.telemetry()
        h0, t0 = w.stats().hashes_total, time.time()
        time.sleep(1.0)
        rate = (w.stats().hashes_total - h0) / (time.time() - t0)
    assert power > 50 and 500 < mhz < 3000 and temp > 0
    assert rate > 1e9
    jpth = power / (rate / 1e12)
    assert 1e3 < jpth < 1e6, jpth     # ~55 kJ/TH at ~500 W and ~9 GH/s


def test_int_peak_microbenchmark_runs(miner):
    names = miner.int_peak_modes()
    assert "mix_alu" in names
    ops, mhz, ms = miner.int_peak(names.index("lop3"), iters=256)
    assert ops > 1e12 and 500 < mhz < 3000 and ms > 0
