#!/usr/bin/env python
"""Throughput of every version-shared scan variant on a BIP-320 rolled job (configs[3] shape),
next to the plain kernel launched once per version.  Share sets of all choices must be identical.

    python scripts/vs_sweep.py [--versions 16] [--log2 28] [--reps 2] [--only NAME ...]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from otedama_b200 import miner, workloads as W  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--versions", type=int, default=16)
ap.add_argument("--log2", type=int, default=28, help="nonces per version = 2^log2")
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--difficulty", type=float, default=1.0 / 64)
ap.add_argument("--only", nargs="*", default=None)
a = ap.parse_args()

c4 = W.config4_v2()
hdr = miner.Header(c4["version"], c4["prev_hash"], c4["merkle_root"], c4["time"], c4["bits"], 0)
work = miner.Work(job_id=4, header=hdr, target=miner.target_from_difficulty(a.difficulty), version_mask=c4["mask"],
                  version_count=a.versions, nonce_count=1 << a.log2)
names = miner.vs_variants()
choices = [(-2, "plain_per_version")] + list(enumerate(names))
if a.only:
    choices = [c for c in choices if c[1] in a.only]
ref = None
for idx, name in choices:
    miner.vs_select(idx)
    best = 0.0
    for _ in range(a.reps):
        r = miner.scan(work)
        best = max(best, r.hashes / r.device_ms / 1e6)
    k = sorted((s.version, s.nonce, s.hash) for s in r.shares)
    if ref is None:
        ref = k
    same = "same-shares" if k == ref else "SHARES-DIFFER"
    print(f"{name:20s} GH/s={best:7.3f} hashes={r.hashes} launches={r.launches} found={r.found} {same}", flush=True)
    assert k == ref, name
miner.vs_select(-1)
