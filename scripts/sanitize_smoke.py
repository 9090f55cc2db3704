#!/usr/bin/env python
"""Small pass over every kernel for compute-sanitizer (memcheck / racecheck / initcheck):
    compute-sanitizer --tool memcheck python scripts/sanitize_smoke.py
Dense-hit scans so the ballot + atomic ring path, ring overflow and the slow path all run."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes  # noqa: E402

from otedama_b200 import capi, miner, workloads as W  # noqa: E402

f = W.config2_fields(1)
hdr = miner.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0)
easy = miner.target_from_difficulty(2.0 ** -24)
for v in range(len(miner.scan_variants())):
    r = miner.scan(miner.Work(header=hdr, target=easy, nonce_begin=0xFFFF0000, nonce_count=1 << 16), variant=v)
    print("scan variant", v, r.found)
r = miner.scan(miner.Work(header=hdr, target=b"\xff" * 32, nonce_count=1 << 17))       # ring overflow
print("overflow", r.found, len(r.shares))
print("hash_headers", len(miner.hash_headers([os.urandom(80) for _ in range(1000)])))
c3 = W.config3_v1()
w = miner.Work(header=miner.Header(c3["version"], c3["prev_hash"], b"\x00" * 32, c3["time"], c3["bits"], 0), target=easy,
               coinb1=c3["coinb1"], extranonce1=c3["extranonce1"], coinb2=c3["coinb2"], merkle_branches=c3["branches"],
               extranonce2_size=8, extranonce2_count=5, version_mask=W.BIP320_MASK, version_count=3, nonce_count=1 << 12)
print("rolled", miner.scan(w).found, len(miner.prepare_instances(w, 0, 15)))
r = miner.scan_headers([hdr, hdr], easy, 0, 1 << 14)
print("burst", r.found)
cw, keep = miner.Work(header=hdr, target=easy, nonce_count=1 << 14)._c()
out = (capi.CShare * 64)()
fnd, hs = ctypes.c_uint64(), ctypes.c_uint64()
print("noalloc", capi.lib().otd_scan_noalloc(0, ctypes.byref(cw), -1, out, 64, ctypes.byref(fnd), ctypes.byref(hs)))
with miner.Worker(miner.WorkerConfig(launch_log2=16)) as wk:
    wk.set_work(miner.Work(header=hdr, target=easy, nonce_count=1 << 17))
    wk.wait_exhausted(60000)
    print("worker", len(wk.poll_shares(4096, 0)), wk.stats().hashes_total)
print("int_peak", miner.int_peak(0, iters=16)[0] > 0)
print("sanitize smoke done")
