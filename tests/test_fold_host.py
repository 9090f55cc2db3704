"""The product's folded scan arithmetic (otedama_b200/csrc/scan_core.cuh,
sha256_core.cuh) instantiated for the HOST and compared with the oracle, so the
constant folding and round bookkeeping are checked on a CPU-only box.  The host
build is a test helper (tests/_hostcheck/fold_check.cpp), not a product path."""
import ctypes
import os
import random
import struct
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
IV7 = 0x5BE0CD19


@pytest.fixture(scope="module")
def fold():
    src = os.path.join(HERE, "_hostcheck", "fold_check.cpp")
    so = os.path.join(HERE, "_hostcheck", "libfoldcheck.so")
    deps = [src] + [os.path.join(HERE, "..", "otedama_b200", "csrc", f)
                    for f in ("scan_core.cuh", "scan_vs_core.cuh", "sha256_core.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        subprocess.check_call(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-Wno-unknown-pragmas", "-o", so, src])
    L = ctypes.CDLL(so)
    L.fc_iv7.restype = ctypes.c_uint32
    return L


KN = dict(E=1, HKW=2, T1=4, A=8, W1=16, W2=32, SSHR=64, SROT=128, BS1=256, BS0=512, T1X=1024, COMPACT=2048, SP=4096, T1Z=8192)
ADD1 = KN["E"] | KN["W1"]
ADDS = ADD1 | KN["A"] | KN["W2"]
CPT = KN["COMPACT"]


def ZSET(n):
    return ADDS | KN["SP"] | KN["T1Z"] | (n << 16)


KNOB_SETS = [0, 1023, 63, ADDS, ADDS | KN["T1"], ADDS | KN["T1X"], ADD1 | KN["A"], ADD1 | KN["T1X"],
             ADD1 | KN["A"] | KN["T1X"], KN["T1"] | KN["A"] | KN["W2"], KN["E"] | KN["W1"] | KN["SSHR"],
             KN["E"] | KN["W1"] | KN["SSHR"] | KN["HKW"] | KN["BS1"],
             KN["E"] | KN["W1"] | KN["SSHR"] | KN["HKW"] | KN["BS1"] | KN["BS0"] | KN["SROT"],
             ADDS | KN["T1X"] | KN["SP"], CPT | ADDS | KN["T1X"] | KN["SP"], ZSET(1), ZSET(5), ZSET(6), ZSET(8), CPT | ZSET(6), ZSET(6) & ~KN["E"], ZSET(6) & ~KN["A"],
             ZSET(6) & ~KN["W2"],
             CPT, CPT | ADDS | KN["T1X"], CPT | ADDS | KN["T1"], CPT | ADDS, CPT | ADD1 | KN["A"],
             CPT | 1023 | KN["T1X"]]


def test_every_production_variant_policy_is_host_checked():
    """The knob set of each production scan variant (csrc/scan_variants.h) is one the
    CPU-side check below exercises."""
    import re
    txt = open(os.path.join(HERE, "..", "otedama_b200", "csrc", "scan_variants.h")).read()
    nprod = int(re.search(r"#define OTD_NUM_PRODUCTION_VARIANTS (\d+)", txt).group(1))
    rows = re.findall(r"X\((\d+), \"(\w+)\", (\d+), (\d+), ([^,]+?), (true|false), \d+, \d+\)", txt)
    assert len(rows) >= nprod
    env = {"OTD_KN_ADD1": ADD1, "OTD_KN_ADDS": ADDS, "OTD_KN_T1X": ADDS | KN["T1X"], "OTD_KN_SPX": ADDS | KN["T1X"] | KN["SP"], "OTD_KN_BEST": ZSET(6), **{f"otd::KN_{k}": v for k, v in KN.items()}}
    for idx, name, v, it, kn, half in rows[:nprod]:
        assert half == "false", f"production variant {name} must compute the full SHA-256d"
        val = 0
        for tok in kn.replace("u", "").split("|"):
            tok = tok.strip()
            val |= env[tok] if tok in env else int(tok)
        assert val in KNOB_SETS, (name, val)


def _e60(L, h76, nonces, knobs=0):
    v = len(nonces)
    arr = (ctypes.c_uint32 * v)(*nonces)
    out = (ctypes.c_uint32 * v)()
    assert L.fc_scan_e60_kn(h76, arr, v, knobs, out) == 0
    return list(out)


def test_top_word_matches_oracle(fold, oracle):
    assert fold.fc_iv7() == IV7
    rnd = random.Random(1)
    genesis = bytes.fromhex("0100000000000000000000000000000000000000000000000000000000000000000000003ba3edfd7a7b12b27ac72c3e67768f617fc81bc3888a51323a9fb8aa4b1e5e4a29ab5f49ffff001d")
    cases = [(genesis, [2083236893])]
    for _ in range(200):
        v = rnd.choice([1, 2, 4])
        cases.append((rnd.randbytes(76), [rnd.getrandbits(32) for _ in range(v)]))
    cases.append((b"\x00" * 76, [0, 0xFFFFFFFF]))
    cases.append((b"\xff" * 76, [0xFFFFFFFF, 1, 2, 3]))
    for ci, (h76, nonces) in enumerate(cases):
        # every instruction-placement policy computes the same function
        for knobs in (KNOB_SETS if ci % 8 == 0 else [KNOB_SETS[ci % len(KNOB_SETS)]]):
            for n, e in zip(nonces, _e60(fold, h76, nonces, knobs)):
                d = oracle.sha256d(h76 + struct.pack("<I", n))
                h7 = struct.unpack(">I", d[28:32])[0]
                assert (e + IV7) & 0xFFFFFFFF == h7, knobs
    # the genesis nonce is exactly the "top word zero" constant the kernel tests for
    assert _e60(fold, genesis, [2083236893])[0] == (-IV7) & 0xFFFFFFFF


def test_midstate_digest_and_compare(fold, oracle):
    rnd = random.Random(2)
    out = ctypes.create_string_buffer(32)
    le = ctypes.c_int()
    for _ in range(200):
        h76, n = rnd.randbytes(76), rnd.getrandbits(32)
        want = oracle.sha256d(h76 + struct.pack("<I", n))
        # targets around the hash itself exercise every branch of the 256-bit compare
        hv = int.from_bytes(want, "little")
        for tv in (hv, hv - 1, hv + 1, hv ^ (1 << rnd.randrange(256)), rnd.getrandbits(256)):
            tv = max(0, min(tv, (1 << 256) - 1))
            t = tv.to_bytes(32, "little")
            fold.fc_sha256d_mid(h76, n, t, out, ctypes.byref(le))
            assert out.raw == want
            assert bool(le.value) == oracle.less_or_equal(want, t) == (hv <= tv)


def VSSET(n, m):
    return ZSET(n) | (m << 20)


VS_KNOB_SETS = [0, 1023, ADDS | KN["T1X"] | KN["SP"], VSSET(6, 0), VSSET(6, 6), VSSET(6, 4), VSSET(6, 7), VSSET(5, 6),
                VSSET(7, 7), VSSET(8, 0), VSSET(7, 7) | (1 << 24), VSSET(6, 6) | (1 << 24),
                VSSET(7, 7) | (1 << 25), VSSET(7, 0) | (1 << 25), VSSET(7, 7) | (3 << 24),
                VSSET(7, 0), VSSET(6, 7), VSSET(7, 6), VSSET(5, 7), VSSET(7, 5), VSSET(8, 7), VSSET(6, 0),
                VSSET(7, 0) | (1 << 15), VSSET(7, 0) | (5 << 28)]


def test_every_version_shared_variant_policy_is_host_checked():
    """Each version-shared scan variant's knob set (csrc/scan_variants.h, OTD_VS_VARIANTS) is exercised below."""
    import re
    txt = open(os.path.join(HERE, "..", "otedama_b200", "csrc", "scan_variants.h")).read()
    rows = re.findall(r"Y\((\d+), \"(\w+)\", (\d+), OTD_KN_VS\((\d+), (\d+)\)((?: \| (?:otd::KN_VS_\w+|OTD_KN_ZSEL\(\d+\)))*), \d+, \d+\)", txt)
    assert len(rows) == len(re.findall(r"^\s*Y\(\d+,", txt, flags=re.M)) > 0, "unparsed version-shared variant rows"
    for idx, name, it, n, m, extra in rows:
        bits = ((1 << 24) if "KN_VS_F20" in extra else 0) | ((1 << 25) if "KN_VS_LOOP1" in extra else 0) | \
               ((1 << 15) if "KN_VS_CFMA" in extra else 0)
        zsel = re.search(r"OTD_KN_ZSEL\((\d+)\)", extra)
        bits |= (int(zsel.group(1)) << 28) if zsel else 0
        assert VSSET(int(n), int(m)) | bits in VS_KNOB_SETS, name


def test_version_shared_core_matches_oracle(fold, oracle):
    """scan_vs_core.cuh: block 2's message schedule expanded once per nonce and reused by every BIP-320 version
    instance gives, per instance, the same top digest word as hashing that instance's 80 bytes from scratch."""
    rnd = random.Random(7)
    mask = 0x1FFFE000
    for case in range(120):
        h76 = bytearray(rnd.randbytes(76))
        n_ver = rnd.choice([1, 2, 3, 16, 33])
        base = rnd.getrandbits(32) & ~mask
        versions = [base | (rnd.getrandbits(32) & mask) for _ in range(n_ver)]
        nonce = rnd.choice([0, 0xFFFFFFFF, rnd.getrandbits(32)])
        arr = (ctypes.c_uint32 * n_ver)(*versions)
        out = (ctypes.c_uint32 * n_ver)()
        for knobs in (VS_KNOB_SETS if case % 6 == 0 else [VS_KNOB_SETS[case % len(VS_KNOB_SETS)]]):
            assert fold.fc_vs_e60_kn(bytes(h76), arr, n_ver, nonce, knobs, out) == 0, knobs
            for v, e in zip(versions, out):
                hdr = struct.pack("<I", v) + bytes(h76[4:]) + struct.pack("<I", nonce)
                h7 = struct.unpack(">I", oracle.sha256d(hdr)[28:32])[0]
                assert (e + IV7) & 0xFFFFFFFF == h7, (knobs, hex(v))
