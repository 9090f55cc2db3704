#!/usr/bin/env python
"""Host CPU time the worker thread costs while the GPU scans (completion is polled every 100 us)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from otedama_b200 import miner, workloads as W  # noqa: E402

f = W.genesis_fields()
hdr = miner.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0)
w = miner.Worker().start()
w.set_work(miner.Work(header=hdr, target=miner.target_from_nbits(W.DIFF1_NBITS), repeat=True))
time.sleep(1.0)
c0, t0, h0 = time.process_time(), time.time(), w.stats().hashes_total
time.sleep(10.0)
c1, t1, h1 = time.process_time(), time.time(), w.stats().hashes_total
w.stop()
print(f"worker host CPU: {(c1 - c0) / (t1 - t0) * 100:.1f} % of one core over {t1 - t0:.1f} s "
      f"at {(h1 - h0) / (t1 - t0) / 1e9:.2f} GH/s")
