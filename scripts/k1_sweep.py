#!/usr/bin/env python
"""Throughput of plain scan-kernel variants over a full 2^32 scan of the genesis header (share set checked).
    OTD_ALLOW_EXPERIMENTAL=1 python scripts/k1_sweep.py [--match x_z] [--reps 2]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from otedama_b200 import miner, workloads as W  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--match", nargs="*", default=None, help="only variants whose name contains one of these")
ap.add_argument("--reps", type=int, default=2)
a = ap.parse_args()
f = W.genesis_fields()
hdr = miner.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0)
t = miner.target_from_nbits(W.DIFF1_NBITS)
for v, name in enumerate(miner.scan_variants()):
    if name.startswith("h_") or (a.match and v != 0 and not any(m in name for m in a.match)):
        continue
    best = 0.0
    for _ in range(a.reps):
        r = miner.scan(miner.Work(header=hdr, target=t), variant=v)
        assert [s.nonce for s in r.shares] == [2083236893, 3197545707], (name, r.shares)
        best = max(best, r.hashes / r.device_ms / 1e6)
    print(f"{name:18s} GH/s={best:7.3f}", flush=True)
