"""The drop-in boundary without a GPU: the C-ABI library loads, exports every
symbol include/otedama_b200.h declares, its host-side helpers (the mirrors of
internal/miner/sha256d.go's target/header functions) agree with the oracle, the
compute entry points fail loudly when no device exists, and the product never
touches oracle/."""
import ctypes
import math
import os
import random
import re
import struct

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(product):
    from otedama_b200 import capi
    L = capi.lib()
    declared = capi.declared_symbols()
    assert len(declared) >= 25
    for sym in declared:
        assert hasattr(L, sym), f"{sym} declared in include/otedama_b200.h but not exported"
        assert sym in capi._SIGS, f"{sym} has no ctypes signature"
    assert L.otd_abi_version() == 2
    assert L.otd_last_error() is not None


def test_every_entry_point_cites_the_reference():
    txt = open(os.path.join(ROOT, "include", "otedama_b200.h")).read()
    for cite in ("internal/miner/sha256d.go", "internal/miner/worker.go", "internal/hal/device.go",
                 "internal/engine/setup.go"):
        assert cite in txt


def test_product_never_uses_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU arms may touch oracle/: not the package, not the
    headers, not the Go shim, not the measurement scripts."""
    for top in ("otedama_b200", "include", "go", "scripts"):
      for dirpath, _, files in os.walk(os.path.join(ROOT, top)):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                code = re.sub(r'""".*?"""', "", src, flags=re.S)
                code = "\n".join(l for l in code.splitlines() if not l.lstrip().startswith(("#", "//", "*", "/*")))
                assert not re.search(r"^\s*(from|import)\s+oracle\b", code, flags=re.M), f
                assert "libotedama_oracle" not in code and "orc_" not in code, f


def test_struct_layouts_match_the_header(product):
    from otedama_b200 import capi
    L = capi.lib()
    for which, cls in enumerate((capi.DeviceInfo, capi.CHeader, capi.CWork, capi.CShare, capi.CWorkerConfig,
                                 capi.CStats, capi.CTelemetry)):
        assert L.otd_abi_struct_size(which) == ctypes.sizeof(cls), cls.__name__
    assert L.otd_abi_struct_size(99) == -1
    assert ctypes.sizeof(capi.CHeader) == 80
    assert ctypes.sizeof(capi.CShare) == 80 + capi.OTD_JOB_ID_MAX
    assert ctypes.sizeof(capi.CWorkerConfig) == 16
    assert capi.CWork.header.offset == 8 and capi.CWork.target.offset == 92


def test_header_bytes_and_parse(product, oracle, golden):
    from otedama_b200 import miner
    raw = bytes.fromhex(golden["reference_vectors"]["sha256d"][0]["input_hex"])
    h = miner.parse_header(raw)
    assert h.nonce == 2083236893 and h.version == 1 and h.bits == 0x1D00FFFF
    assert h.bytes() == raw
    rnd = random.Random(4)
    for _ in range(200):
        b = rnd.randbytes(80)
        assert miner.parse_header(b).bytes() == b
        oh = oracle.Header.parse(b)
        mh = miner.parse_header(b)
        assert (oh.version, oh.prev_hash, oh.merkle_root, oh.time, oh.bits, oh.nonce) == \
               (mh.version, mh.prev_hash, mh.merkle_root, mh.time, mh.bits, mh.nonce)
    with pytest.raises(miner.MinerError):
        miner.parse_header(b"\x00" * 79)


def test_target_helpers_match_reference_vectors_and_oracle(product, oracle, golden):
    from otedama_b200 import miner
    rv = golden["reference_vectors"]
    for v in rv["target_from_nbits"]:
        assert miner.target_from_nbits(int(v["nbits"], 16)) == bytes.fromhex(v["target_be_hex"])[::-1]
    for v in rv["target_from_nbits_errors"]:
        with pytest.raises(miner.MinerError):
            miner.target_from_nbits(int(v["nbits"], 16))
    with pytest.raises(miner.MinerError):
        miner.meets_target(b"\x00" * 32, 0x00000001)
    for v in rv["meets_target"]:
        assert miner.meets_target(bytes([v["hash_fill"]]) * 32, int(v["nbits"], 16)) is v["meets"]
    assert miner.target_from_difficulty(1.0) == miner.target_from_nbits(0x1D00FFFF)
    for s in rv["difficulty"]["errors"]["values"]:
        with pytest.raises(miner.MinerError):
            miner.target_from_difficulty(float(s))
    for v in rv["nbits_from_target"]:
        t = miner.target_from_nbits(int(v["from_nbits"], 16)) if "from_nbits" in v else bytes.fromhex(v["target_le_hex"])
        assert miner.nbits_from_target(t) == oracle.nbits_from_target(t)
    rnd = random.Random(9)
    # every nBits class, against the oracle (same value or same error)
    for _ in range(3000):
        nb = rnd.getrandbits(32) if rnd.random() < 0.5 else (rnd.randrange(0, 40) << 24) | rnd.getrandbits(24)
        try:
            want = oracle.target_from_nbits(nb)
        except oracle.OracleError:
            want = None
        try:
            got = miner.target_from_nbits(nb)
        except miner.MinerError:
            got = None
        assert got == want, hex(nb)
    for _ in range(5000):
        r = rnd.random()
        if r < 0.4:
            d = rnd.uniform(1e-9, 1e9)
        elif r < 0.7:
            d = 2.0 ** rnd.randint(-60, 230) * rnd.choice([1, 3, 5, 0.75, 1.1])
        else:
            d = struct.unpack("<d", struct.pack("<Q", rnd.getrandbits(64)))[0]
        try:
            want = oracle.target_from_difficulty(d)
        except oracle.OracleError:
            want = None
        try:
            got = miner.target_from_difficulty(d)
        except miner.MinerError:
            got = None
        assert got == want, d
    for _ in range(500):
        a, b = rnd.randbytes(32), rnd.randbytes(32)
        assert miner.less_or_equal(a, b) == oracle.less_or_equal(a, b)
        assert miner.less_or_equal(a, a)


def test_compute_fails_loudly_without_a_device(product):
    from otedama_b200 import capi, miner
    if capi.lib().otd_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(capi.OtdError) as e:
        miner.hash_headers([b"\x00" * 80])
    assert e.value.code == capi.OTD_ENODEV and "no CPU fallback" in str(e.value)
    with pytest.raises(capi.OtdError):
        miner.scan(miner.Work(target=b"\xff" * 32))
    with pytest.raises(capi.OtdError):
        miner.Worker().start()
    assert miner.Worker().stats() == miner.Stats()  # TestWorker_StatsBeforeStart (worker_test.go:39-54)
    miner.Worker().stop()                           # Stop before Start is a no-op (worker.go:156-165)


def test_work_validation_happens_before_any_device_call(product):
    from otedama_b200 import capi, miner
    bad = [miner.Work(target=b"\xff" * 32, nonce_begin=10, nonce_count=1 << 32),
           miner.Work(target=b"\xff" * 32, version_count=4),
           miner.Work(target=b"\xff" * 32, version_mask=0x1FFFE000, version_begin=1 << 16),
           miner.Work(target=b"\xff" * 32, coinb1=b"ab", extranonce2_size=4, extranonce2_count=0),
           miner.Work(target=b"\xff" * 32, coinb1=b"ab", extranonce2_size=2, extranonce2_begin=65535, extranonce2_count=2)]
    for w in bad:
        with pytest.raises(capi.OtdError) as e:
            miner.scan(w)
        assert e.value.code == capi.OTD_EINVAL, str(e.value)


def test_hal_identity_rules(product):
    from otedama_b200 import hal
    hal.Identity("gpu-cuda0", "gpu", "nvidia", "NVIDIA B200").validate()
    for bad in ("", "gpu cuda0", "gpu/cuda0"):
        with pytest.raises(ValueError):
            hal.Identity(bad, "gpu", "nvidia", "x").validate()
    assert hal.CudaDriver().name() == "cuda"
    assert isinstance(hal.CudaDriver().enumerate(), list)


def test_hash_rate_string(product):
    """TestHashRateString (internal/miner/worker_test.go:246-265), same vectors."""
    from otedama_b200 import miner
    for hps, want in ((500, "500 H/s"), (1500, "1.50 kH/s"), (2.5e6, "2.50 MH/s"), (3.7e9, "3.70 GH/s"),
                      (120e12, "120.00 TH/s")):
        assert miner.hash_rate_string(hps) == want


def test_missing_library_fails_loudly(product, monkeypatch):
    """No silent fallback: if libotedama_b200.so is not there, loading raises."""
    from otedama_b200 import capi
    monkeypatch.setattr(capi, "_lib", None)
    monkeypatch.setattr(capi, "LIB_PATH", "/nonexistent/libotedama_b200.so")
    with pytest.raises(capi.OtdError, match="no CPU fallback"):
        capi.lib()


def test_no_undeclared_c_symbols_are_exported(product):
    """The other direction: every unmangled otd_* symbol the library exports is declared in the header."""
    import subprocess
    from otedama_b200 import capi
    out = subprocess.run(["nm", "-D", "--defined-only", capi.LIB_PATH], capture_output=True, text=True).stdout
    exported = {line.split()[-1] for line in out.splitlines() if line.split() and line.split()[-1].startswith("otd_")}
    assert exported == set(capi.declared_symbols())


# -- work validation (the checks every entry point runs before touching a device) ---------------------------------
def test_work_validation_errors(product):
    """SetWork cannot fail in the reference because bad jobs are rejected before it (run.go:1274-1279,
    1341-1344); the search-space extension adds fields that can be inconsistent, and those are refused with a
    message.  otd_work_partition runs the same validation without needing a device."""
    from otedama_b200 import engine, miner
    ok = dict(coinb1=b"\x01\x02", extranonce2_size=4, extranonce2_count=8)
    bad = [
        (dict(nonce_begin=0xFFFFFFFF, nonce_count=2), "nonce window exceeds 2^32"),
        (dict(nonce_begin=1, nonce_count=1 << 32), "nonce window exceeds 2^32"),
        (dict(version_count=2), "without a version_mask"),
        (dict(version_mask=0x6000, version_begin=4), "version_begin outside"),
        (dict(version_mask=0x6000, version_count=5), "version_count exceeds"),
        (dict(version_mask=0x6000, version_begin=1, version_stride=2, version_count=3), "version_count exceeds"),
        (dict(version_mask=0xFFFFFFFF), "larger than 2^31 indices"),       # 2^32 indices would wrap the device's 32-bit index
        (dict(ok, extranonce2_size=17), "extranonce2_size > 16"),
        (dict(ok, extranonce2_count=0), "extranonce2_count must be >= 1"),
        (dict(ok, extranonce2_size=1, extranonce2_count=257), "does not fit extranonce2_size"),
        (dict(ok, extranonce2_size=2, extranonce2_begin=65536, extranonce2_count=1), "does not fit extranonce2_size"),
        (dict(ok, extranonce2_size=2, extranonce2_begin=1, extranonce2_stride=1 << 15, extranonce2_count=3), "does not fit"),
        (dict(ok, coinb2=b"\x00" * ((1 << 20) + 1)), "coinbase too large"),
        (dict(ok, merkle_branches=[b"\x00" * 32] * 65), "coinbase too large"),
        (dict(ok, extranonce2_size=8, extranonce2_count=1 << 50, version_mask=0x1FFFE000), "search space too large"),
    ]
    for fields, msg in bad:
        with pytest.raises(engine.EngineError, match=msg.replace("^", r"\^")):
            engine.partition_work(miner.Work(**fields), 0, 1)
    # and the same work with consistent fields passes
    for fields in (dict(nonce_begin=0xFFFFFFFF, nonce_count=1), dict(version_mask=0x6000, version_begin=3),
                   dict(version_mask=0xFFFFFFFF, version_count=1 << 31), dict(version_mask=0x7FFFFFFF),
                   dict(version_mask=0x6000, version_begin=1, version_stride=2, version_count=2),
                   dict(ok, extranonce2_size=1, extranonce2_count=256), dict(ok, extranonce2_size=0, extranonce2_count=1)):
        engine.partition_work(miner.Work(**fields), 0, 1)


def test_public_difficulty_example(product, oracle):
    """The Bitcoin wiki's worked example for compact targets: nBits 0x1b0404cb <-> difficulty 16307.420938523983
    (public knowledge, not a reference vector; both directions, oracle and C ABI)."""
    from otedama_b200 import miner
    d = 16307.420938523983
    for m in (oracle, miner):
        t = m.target_from_nbits(0x1B0404CB)
        assert t[::-1].hex() == "00000000000404cb" + "00" * 24
        assert int.from_bytes(m.target_from_nbits(0x1D00FFFF), "little") / int.from_bytes(t, "little") == d
        td = m.target_from_difficulty(d)
        assert m.nbits_from_target(td) == 0x1B0404CB and td[::-1].hex().startswith("00000000000404cb00000000")
    assert miner.target_from_difficulty(d) == oracle.target_from_difficulty(d)
