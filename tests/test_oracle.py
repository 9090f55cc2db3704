"""The CPU oracle against the reference's own known-answer vectors
(tests/golden/reference_vectors.json, transcribed from the reference tests) and
against an independent implementation (hashlib/OpenSSL).  CPU only."""
import hashlib
import math
import os
import random
import struct

import pytest


def test_sha256d_reference_kats(oracle, golden):
    for backend in (0, 1):
        used = oracle.set_backend(backend)
        for v in golden["reference_vectors"]["sha256d"]:
            got = oracle.sha256d(bytes.fromhex(v["input_hex"]))
            assert got.hex() == v["hash_hex"], (v["cite"], used)
    oracle.set_backend(-1)


def test_genesis_nonce_and_header_layout(oracle, golden):
    rv = golden["reference_vectors"]
    raw = bytes.fromhex(rv["sha256d"][0]["input_hex"])
    h = oracle.Header.parse(raw)
    assert h.nonce == rv["genesis_nonce"]["nonce"] == 2083236893
    assert h.bytes() == raw
    assert oracle.hash_header(h).hex() == rv["sha256d"][0]["hash_hex"]
    # TestHeader_Bytes_Roundtrip / TestHeader_NonceIsAtOffset76
    hr = rv["header_roundtrip"]
    x = oracle.Header(hr["version"], bytes(range(32)), bytes(255 - i for i in range(32)), hr["time"],
                      hr["bits"], hr["nonce"])
    b = x.bytes()
    assert oracle.Header.parse(b) == x
    assert struct.unpack_from("<I", b, hr["nonce_offset"])[0] == hr["nonce"]
    assert struct.unpack_from("<I", oracle.Header(nonce=0x01020304).bytes(), 76)[0] == 0x01020304


def test_scalar_and_shani_agree_with_hashlib(oracle):
    rnd = random.Random(7)
    for backend in (0, 1):
        oracle.set_backend(backend)
        for n in list(range(0, 130)) + [200, 255, 256, 1000, 4096]:
            d = rnd.randbytes(n)
            assert oracle.sha256d(d) == oracle.sha256d_hashlib(d)
            assert oracle.sha256(d) == hashlib.sha256(d).digest()
    oracle.set_backend(-1)


def test_target_from_nbits_vectors(oracle, golden):
    rv = golden["reference_vectors"]
    for v in rv["target_from_nbits"]:
        t = oracle.target_from_nbits(int(v["nbits"], 16))
        assert t == bytes.fromhex(v["target_be_hex"])[::-1], v["cite"]
    for v in rv["target_from_nbits_errors"]:
        with pytest.raises(oracle.OracleError):
            oracle.target_from_nbits(int(v["nbits"], 16))
    for v in rv["target_from_nbits_ok_nonzero"]:
        assert any(oracle.target_from_nbits(int(v["nbits"], 16)))


def test_target_from_difficulty_vectors(oracle, golden):
    d = golden["reference_vectors"]["difficulty"]
    one = oracle.target_from_difficulty(d["one_equals_nbits"]["difficulty"])
    assert one == oracle.target_from_nbits(int(d["one_equals_nbits"]["nbits"], 16))
    easy = oracle.target_from_difficulty(d["easier"]["difficulty"])
    hard = oracle.target_from_difficulty(d["harder"]["difficulty"])
    assert oracle.less_or_equal(one, easy) and one != easy
    assert oracle.less_or_equal(hard, one) and hard != one
    for s in d["errors"]["values"]:
        with pytest.raises(oracle.OracleError):
            oracle.target_from_difficulty(float(s))
    # power-of-two difficulties divide exactly
    for k in range(-30, 200):
        want = (0xFFFF << 208) >> k if k >= 0 else (0xFFFF << 208) << -k
        if want == 0 or want.bit_length() > 256:
            with pytest.raises(oracle.OracleError):
                oracle.target_from_difficulty(2.0 ** k)
        else:
            assert int.from_bytes(oracle.target_from_difficulty(2.0 ** k), "little") == want
    # floor of the exact rational whenever the 256-bit rounding cannot reach an integer
    from fractions import Fraction
    rnd = random.Random(3)
    for _ in range(300):
        x = rnd.uniform(0.5, 1e9)
        exact = Fraction(0xFFFF << 208) / Fraction(x)
        assert int.from_bytes(oracle.target_from_difficulty(x), "little") in (math.floor(exact), math.floor(exact) + 1)


def test_nbits_from_target_vectors(oracle, golden):
    for v in golden["reference_vectors"]["nbits_from_target"]:
        t = oracle.target_from_nbits(int(v["from_nbits"], 16)) if "from_nbits" in v else bytes.fromhex(v["target_le_hex"])
        got = oracle.nbits_from_target(t)
        if "nbits" in v:
            assert got == int(v["nbits"], 16), v["cite"]
        if "exp" in v:
            assert got >> 24 == v["exp"], v["cite"]
        if "exp_min" in v:
            assert got >> 24 >= v["exp_min"], v["cite"]


def test_less_or_equal_and_meets_target(oracle, golden):
    lv = golden["reference_vectors"]["less_or_equal"]
    for a, b, want in lv["expect"]:
        assert oracle.less_or_equal(bytes.fromhex(lv[a]), bytes.fromhex(lv[b])) is want
    for v in golden["reference_vectors"]["meets_target"]:
        assert oracle.meets_target(bytes([v["hash_fill"]]) * 32, int(v["nbits"], 16)) is v["meets"]
    rnd = random.Random(11)
    for _ in range(2000):
        a, b = rnd.randbytes(32), rnd.randbytes(32)
        if rnd.random() < 0.3:
            k = rnd.randrange(33)          # share the k most significant bytes
            b = b[:32 - k] + a[32 - k:]
        assert oracle.less_or_equal(a, b) == (int.from_bytes(a, "little") <= int.from_bytes(b, "little"))


def test_grind_matches_python_loop_and_finds_genesis(oracle):
    from tests import workloads as wl
    h = wl.genesis_header()
    t1 = oracle.target_from_nbits(0x1D00FFFF)
    shares, dropped, hashes = oracle.grind(h, t1, 2083236893 - 50000, 100000, threads=3)
    assert [s.nonce for s in shares] == [2083236893] and dropped == 0 and hashes == 100000
    assert shares[0].hash.hex() == "6fe28c0ab6f1b372c1a6a246ae63f74f931e8365e15a089c68d6190000000000"
    assert shares[0].ntime == h.time and shares[0].version == h.version
    easy = oracle.target_from_difficulty(2.0 ** -24)
    for threads, mid in ((1, False), (4, False), (5, True)):
        got, _, n = oracle.grind(wl.config2_header(7), easy, 1000, 20000, threads=threads, midstate=mid)
        want = oracle.grind_py(wl.config2_header(7), easy, 1000, 20000)
        assert n == 20000 and [(s.nonce, s.hash) for s in got] == [(s.nonce, s.hash) for s in want]
        assert len(got) > 20
    # window that wraps the top of the nonce space is covered once
    got, _, n = oracle.grind(h, easy, (1 << 32) - 5000, 5000, threads=2)
    assert n == 5000 and all(s.nonce >= (1 << 32) - 5000 for s in got)
    # share cap -> dropped counter (worker.go:271-275)
    got, dropped, _ = oracle.grind(h, b"\xff" * 32, 0, 1000, threads=2, cap=10)
    assert len(got) == 10 and dropped == 990


def test_golden_windows_reproduce(oracle, golden):
    from tests import workloads as wl
    doc = golden["synthetic_windows"]
    case = doc["cases"][1]
    h = wl.config2_header(case["index"])
    assert h.bytes()[:76].hex() == case["header76_hex"]
    shares, _, _ = oracle.grind(h, bytes.fromhex(case["target_hex"]), 0, 1 << 20, threads=os.cpu_count() or 1)
    assert [s.nonce for s in shares] == [n for n in case["nonces"] if n < (1 << 20)]


def test_full_genesis_scan_fixture(golden, oracle):
    g = golden.get("genesis_full_scan")
    if g is None:
        pytest.skip("genesis_full_scan.json not generated yet")
    assert g["hashes"] == 1 << 32
    # the mined nonce plus a second, never-published difficulty-1 solution
    assert [s["nonce"] for s in g["shares"]] == [2083236893, 3197545707]
    assert g["shares"][0]["hash"] == "6fe28c0ab6f1b372c1a6a246ae63f74f931e8365e15a089c68d6190000000000"
    for s in g["shares"]:
        hdr = bytes.fromhex(g["header76_hex"]) + s["nonce"].to_bytes(4, "little")
        assert oracle.sha256d_hashlib(hdr).hex() == s["hash"]


def test_v1_merkle_root_unpinned_but_anchored(oracle, golden):
    """No reference code exists for this; anchor: SHA256d(genesis coinbase tx) is
    the merkle root inside the reference's genesis header vector."""
    g = golden["reference_vectors"]["genesis_coinbase_tx"]
    tx = bytes.fromhex(g["tx_hex"])
    # split the tx as a pool would: coinb1 | extranonce1(4) | extranonce2(4) | coinb2
    root = oracle.v1_merkle_root(tx[:50], tx[50:54], tx[54:58], tx[58:], [])
    assert root.hex() == g["merkle_root_hex"]
    assert golden["reference_vectors"]["sha256d"][0]["input_hex"][72:136] == g["merkle_root_hex"]
    from tests import workloads as wl
    v1 = wl.config3_v1()
    for e, want in enumerate(golden["synthetic_windows"]["config3_roots_e0_7"]):
        en2 = oracle.extranonce2_bytes(e, 8)
        assert en2 == struct.pack("<Q", e)
        r = oracle.v1_merkle_root(v1["coinb1"], v1["extranonce1"], en2, v1["coinb2"], v1["branches"])
        assert r == oracle.v1_merkle_root_py(v1["coinb1"], v1["extranonce1"], en2, v1["coinb2"], v1["branches"])
        assert r.hex() == want


def test_bip320_versions(oracle, golden):
    from tests import workloads as wl
    v2 = wl.config4_v2()
    seen = set()
    for k in list(range(300)) + [4095, 65535]:
        v = oracle.bip320_version(v2["version"], v2["mask"], k)
        assert v == oracle.bip320_version_py(v2["version"], v2["mask"], k)
        assert v & ~v2["mask"] == v2["version"] & ~v2["mask"]
        assert (v & v2["mask"]) >> 13 == k
        seen.add(v)
    assert len(seen) == 302
    for k, want in golden["synthetic_windows"]["config4_versions_k"].items():
        assert oracle.bip320_version(v2["version"], v2["mask"], int(k)) == want


def _rw(golden):
    return golden["reference_vectors"]["real_world_anchors"]


def test_real_world_anchors_pin_the_unpinned_subpaths(oracle, golden):
    """Public chain data (verified with hashlib here, not trusted): block 1 (prev-hash = the
    reference's genesis hash vector) and block 170 (coinbase -> txid -> merkle fold -> header)."""
    rw = _rw(golden)
    h1 = bytes.fromhex(rw["block1_header_hex"])
    genesis_hash = golden["reference_vectors"]["sha256d"][0]["hash_hex"]
    assert h1[4:36].hex() == genesis_hash                         # ties to the reference-pinned vector
    assert oracle.sha256d(h1)[::-1].hex() == oracle.sha256d_hashlib(h1)[::-1].hex() == rw["block1_hash_display"]
    assert oracle.Header.parse(h1).nonce == rw["block1_nonce"]
    b = rw["block170"]
    cb = bytes.fromhex(b["coinbase_tx_hex"])
    assert oracle.sha256d_hashlib(cb)[::-1].hex() == b["coinbase_txid_display"]
    sp = b["stratum_split"]
    k = sp["coinb1_len"]
    en2 = oracle.extranonce2_bytes(sp["extranonce2_value"], sp["extranonce2_size"])
    assert cb[k:k + 1] == en2
    branch = [bytes.fromhex(b["second_txid_display"])[::-1]]
    root = oracle.v1_merkle_root(cb[:k], b"", en2, cb[k + 1:], branch)
    assert root[::-1].hex() == b["merkle_root_display"]
    assert oracle.v1_merkle_root(cb[:k], b"", en2, cb[k + 1:], []) [::-1].hex() == b["coinbase_txid_display"]
    hdr = oracle.Header(b["version"], bytes.fromhex(b["prev_hash_display"])[::-1], root, b["time"],
                        int(b["bits"], 16), b["nonce"])
    hh = oracle.hash_header(hdr)
    assert hh[::-1].hex() == b["hash_display"]
    assert oracle.meets_target(hh, int(b["bits"], 16))
    # the same through the notify form (prev-hash words swapped, big-endian hex scalars)
    prev_wire = b"".join(hdr.prev_hash[4 * i:4 * i + 4][::-1] for i in range(8)).hex()
    params = ["170", prev_wire, cb[:k].hex(), cb[k + 1:].hex(), [branch[0].hex()], "%08x" % b["version"], b["bits"],
              "%08x" % b["time"], True]
    oh = oracle.v1_header(params, b"", en2)
    oh.nonce = b["nonce"]
    assert oracle.hash_header(oh)[::-1].hex() == b["hash_display"]


def test_stratum_documentation_example_pins_the_v1_wire_conventions(oracle, golden):
    """The protocol documentation's own accepted share: notify + subscribe + submit -> a header whose hash has
    34 leading zero bits, but only under the conventions the oracle (and the product) follow."""
    import hashlib
    ex = _rw(golden)["stratum_docs_example"]
    params, sub = ex["notify_params"], ex["submit_params"]
    en1, en2 = bytes.fromhex(ex["subscribe_extranonce1"]), bytes.fromhex(sub[2])
    assert len(en2) == ex["subscribe_extranonce2_size"] and sub[1] == params[0]
    h = oracle.v1_header(params, en1, en2)
    assert h.merkle_root.hex() == ex["merkle_root_header_order"]
    h.time, h.nonce = int(sub[3], 16), int(sub[4], 16)      # the miner rolled ntime; both big-endian hex of the u32
    hh = oracle.hash_header(h)
    assert hh[::-1].hex() == ex["hash_display"] == hashlib.sha256(hashlib.sha256(h.bytes()).digest()).digest()[::-1].hex()
    assert oracle.less_or_equal(hh, oracle.target_from_difficulty(1.0))
    # the extranonce2 bytes are the little-endian counter the product rolls: value 0x01000000 here
    assert oracle.extranonce2_bytes(int.from_bytes(en2, "little"), 4) == en2
    # any other reading of the wire form gives an ordinary hash
    raw_prev = bytes.fromhex(params[1])
    for prev in (raw_prev, raw_prev[::-1]):
        bad = oracle.Header(h.version, prev, h.merkle_root, h.time, h.bits, h.nonce)
        assert not oracle.hash_header(bad)[::-1].hex().startswith("0000")
    for t, n in ((int.from_bytes(bytes.fromhex(sub[3]), "little"), h.nonce), (h.time, int.from_bytes(bytes.fromhex(sub[4]), "little"))):
        bad = oracle.Header(h.version, h.prev_hash, h.merkle_root, t, h.bits, n)
        assert not oracle.hash_header(bad)[::-1].hex().startswith("0000")


def _hdr_from_anchor(oracle, b, root=None):
    return oracle.Header(b["version"], bytes.fromhex(b["prev_hash_display"])[::-1],
                         root if root is not None else bytes.fromhex(b["merkle_root_display"])[::-1], b["time"],
                         int(b["bits"], 16), b["nonce"])


def test_more_real_world_anchors(oracle, golden):
    """Block 125552 (the Bitcoin wiki's block-hashing example) and block 100000, whose coinbase has a two-level
    merkle branch and a 2-byte extranonce: coinbase -> txid -> fold twice -> root -> header -> published hash."""
    rw = _rw(golden)
    b = rw["block125552"]
    hh = oracle.hash_header(_hdr_from_anchor(oracle, b))
    assert hh[::-1].hex() == b["hash_display"] and oracle.meets_target(hh, int(b["bits"], 16))
    b = rw["block100000"]
    cb = bytes.fromhex(b["coinbase_tx_hex"])
    tx = [bytes.fromhex(t)[::-1] for t in b["txids_display"]]
    assert oracle.sha256d_hashlib(cb) == tx[0]
    sp = b["stratum_split"]
    k, size = sp["coinb1_len"], sp["extranonce2_size"]
    en2 = oracle.extranonce2_bytes(sp["extranonce2_value"], size)
    assert cb[k:k + size] == en2
    branch = [tx[1], oracle.sha256d_hashlib(tx[2] + tx[3])]
    root = oracle.v1_merkle_root(cb[:k], b"", en2, cb[k + size:], branch)
    assert root[::-1].hex() == b["merkle_root_display"] == oracle.v1_merkle_root_py(cb[:k], b"", en2, cb[k + size:], branch)[::-1].hex()
    # the branch order matters: swapped levels give another root
    assert oracle.v1_merkle_root(cb[:k], b"", en2, cb[k + size:], branch[::-1]) != root
    hh = oracle.hash_header(_hdr_from_anchor(oracle, b, root))
    assert hh[::-1].hex() == b["hash_display"] and oracle.meets_target(hh, int(b["bits"], 16))
