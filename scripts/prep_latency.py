#!/usr/bin/env python
"""Latency of the job-preparation kernels (K3 extranonce2 roll + merkle fold, K2 midstate + folding)
for one batch of 256 instances, host call to host results (includes the D2H copies and the sync)."""
import os
import statistics
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from otedama_b200 import miner, workloads as W  # noqa: E402

c3, c4 = W.config3_v1(), W.config4_v2()
t = miner.target_from_difficulty(1.0)
v1 = miner.Work(header=miner.Header(c3["version"], c3["prev_hash"], b"\x00" * 32, c3["time"], c3["bits"], 0), target=t,
                coinb1=c3["coinb1"], extranonce1=c3["extranonce1"], coinb2=c3["coinb2"], merkle_branches=c3["branches"],
                extranonce2_size=8, extranonce2_count=1 << 16)
v2 = miner.Work(header=miner.Header(c4["version"], c4["prev_hash"], c4["merkle_root"], c4["time"], c4["bits"], 0), target=t,
                version_mask=c4["mask"])
for name, w in (("K3+K2: V1 coinbase (165 B + 12 branches), 256 extranonce2 values", v1),
                ("K2 only: 256 BIP-320 versions", v2)):
    miner.prepare_instances(w, 0, 256)
    ts = []
    for i in range(20):
        t0 = time.perf_counter()
        miner.prepare_instances(w, 256 * i, 256)
        ts.append((time.perf_counter() - t0) * 1e3)
    print(f"{name}: median {statistics.median(ts):.3f} ms, min {min(ts):.3f} ms per batch of 256 "
          f"(a batch is {256 * 0.45:.0f} s of scanning)")
