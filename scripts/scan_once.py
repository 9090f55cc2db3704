#!/usr/bin/env python
"""One scan of 2^LOG2 nonces of the genesis header with one kernel variant (for ncu)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from otedama_b200 import miner, workloads as W  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--variant", default="-1")
ap.add_argument("--log2", type=int, default=28)
ap.add_argument("--reps", type=int, default=1)
a = ap.parse_args()
names = miner.scan_variants()
v = names.index(a.variant) if a.variant in names else int(a.variant)
f = W.genesis_fields()
hdr = miner.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0)
for _ in range(a.reps):
    r = miner.scan(miner.Work(header=hdr, target=miner.target_from_nbits(W.DIFF1_NBITS),
                              nonce_begin=max(0, W.GENESIS_NONCE - (1 << (a.log2 - 1))) if a.log2 < 32 else 0,
                              nonce_count=1 << a.log2), variant=v)
    print(f"variant={v} hashes={r.hashes} found={r.found} ms={r.device_ms:.3f} GH/s={r.hashes / r.device_ms / 1e6:.3f}")
    assert r.found >= 1
