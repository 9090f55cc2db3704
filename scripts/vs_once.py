#!/usr/bin/env python
"""One version-rolled scan (configs[3] shape: V BIP-320 versions x 2^LOG2 nonces) with one
version-shared variant (for ncu)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from otedama_b200 import miner, workloads as W  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--variant", default="-1", help="version-shared variant name or index; -2 = plain kernel per version")
ap.add_argument("--versions", type=int, default=32)
ap.add_argument("--log2", type=int, default=25)
ap.add_argument("--reps", type=int, default=1)
a = ap.parse_args()
names = miner.vs_variants()
v = names.index(a.variant) if a.variant in names else int(a.variant)
miner.vs_select(v)
c4 = W.config4_v2()
hdr = miner.Header(c4["version"], c4["prev_hash"], c4["merkle_root"], c4["time"], c4["bits"], 0)
work = miner.Work(header=hdr, target=miner.target_from_difficulty(1.0), version_mask=c4["mask"],
                  version_count=a.versions, nonce_count=1 << a.log2)
for _ in range(a.reps):
    r = miner.scan(work)
    print(f"vs_variant={v} hashes={r.hashes} launches={r.launches} found={r.found} ms={r.device_ms:.3f} "
          f"GH/s={r.hashes / r.device_ms / 1e6:.3f}")
