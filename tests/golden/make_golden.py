#!/usr/bin/env python
"""Generates the committed golden fixtures with the CPU oracle.

    python tests/golden/make_golden.py [--full]

* reference_vectors.json is NOT generated: it is transcribed by hand from the
  reference's own tests (citations inside) and pins the oracle.
* genesis_full_scan.json  (--full, ~6 min on 8 cores): the reference path
  (serialise + 3 compressions per nonce, like Worker.grind) over ALL 2^32 nonces
  of the genesis header at nBits 0x1d00ffff — BASELINE.json configs[0].
* synthetic_windows.json: BASELINE configs[1]'s synthetic headers
  scanned over a 2^24 window at difficulty 2^-16 (dense hits), plus the V1 /
  BIP-320 workload definitions of configs[2]/[3] with their expected roots,
  versions and dense-window share sets.
"""
import hashlib
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import oracle as o  # noqa: E402
from tests import workloads as wl  # noqa: E402


def full_genesis():
    h = wl.genesis_header()
    h.nonce = 0
    target = o.target_from_nbits(0x1D00FFFF)
    threads = os.cpu_count() or 1
    t0 = time.time()
    shares, dropped, hashes = o.grind(h, target, 0, 1 << 32, threads=threads, midstate=False)
    dt = time.time() - t0
    doc = {
        "what": "full 2^32 nonce scan of the Bitcoin genesis header on the CPU oracle (reference arithmetic: 3 compressions per nonce)",
        "header76_hex": h.bytes()[:76].hex(),
        "nbits": 0x1D00FFFF,
        "target_hex": target.hex(),
        "hashes": hashes,
        "threads": threads,
        "backend": o.backend_name(),
        "seconds": round(dt, 1),
        "shares": [{"nonce": s.nonce, "hash": s.hash.hex()} for s in shares],
    }
    assert dropped == 0 and hashes == 1 << 32
    assert 2083236893 in [s.nonce for s in shares], shares
    with open(os.path.join(HERE, "genesis_full_scan.json"), "w") as f:
        json.dump(doc, f, indent=1)
    print("genesis_full_scan.json:", doc["shares"], f"{hashes / dt / 1e6:.1f} MH/s")


def full_synthetic(index=1):
    """Full 2^32 CPU scan of one synthetic configs[1] header at difficulty 1 (midstate path —
    cross-checked against the 3-compression path by tests/test_oracle.py on windows)."""
    h = wl.config2_header(index)
    target = o.target_from_difficulty(1.0)
    threads = os.cpu_count() or 1
    t0 = time.time()
    shares, dropped, hashes = o.grind(h, target, 0, 1 << 32, threads=threads, midstate=True)
    dt = time.time() - t0
    assert dropped == 0 and hashes == 1 << 32
    doc = {"what": f"full 2^32 nonce scan of synthetic configs[1] header {index} at difficulty 1 on the CPU oracle",
           "index": index, "header76_hex": h.bytes()[:76].hex(), "target_hex": target.hex(), "hashes": hashes,
           "threads": threads, "backend": o.backend_name(), "seconds": round(dt, 1),
           "shares": [{"nonce": s.nonce, "hash": s.hash.hex()} for s in shares]}
    with open(os.path.join(HERE, f"config2_full_scan_header{index}.json"), "w") as f:
        json.dump(doc, f, indent=1)
    print(f"config2_full_scan_header{index}.json:", doc["shares"], f"{hashes / dt / 1e6:.1f} MH/s")


def windows():
    doc = {"what": "dense-hit windows of the BASELINE workloads, from the CPU oracle", "cases": []}
    easy = o.target_from_difficulty(2.0 ** -16)
    for i in (0, 1, 2, 3, 128, 256):
        h = wl.config2_header(i)
        shares, _, _ = o.grind(h, easy, 0, 1 << 24, threads=os.cpu_count() or 1)
        doc["cases"].append({"kind": "config2", "index": i, "header76_hex": h.bytes()[:76].hex(),
                             "target_hex": easy.hex(), "first": 0, "count": 1 << 24,
                             "nonces": [s.nonce for s in shares],
                             "hash_digest": hashlib.sha256(b"".join(s.hash for s in shares)).hexdigest()})
    # config 3: V1 notify replay — roots for the first extranonce2 values
    v1 = wl.config3_v1()
    roots = []
    for e in range(8):
        en2 = o.extranonce2_bytes(e, v1["extranonce2_size"])
        roots.append(o.v1_merkle_root(v1["coinb1"], v1["extranonce1"], en2, v1["coinb2"], v1["branches"]).hex())
    doc["config3_roots_e0_7"] = roots
    # config 4: BIP-320 versions
    v2 = wl.config4_v2()
    doc["config4_versions_k"] = {str(k): o.bip320_version(v2["version"], v2["mask"], k)
                                 for k in (0, 1, 2, 3, 255, 256, 4095, 65535)}
    with open(os.path.join(HERE, "synthetic_windows.json"), "w") as f:
        json.dump(doc, f, indent=1)
    print("synthetic_windows.json:", [len(c["nonces"]) for c in doc["cases"]])


if __name__ == "__main__":
    if "--synthetic" in sys.argv:
        full_synthetic(int(sys.argv[sys.argv.index("--synthetic") + 1]))
        sys.exit(0)
    windows()
    if "--full" in sys.argv:
        full_genesis()
