#!/usr/bin/env python
"""Soak test of the worker runtime: random set_work / pause / resume at a high rate for --seconds,
checking every share against the job it claims to belong to (sampled re-hash on the CPU oracle),
that nothing hangs, and that host memory does not grow."""
import argparse
import os
import random
import resource
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from otedama_b200 import miner, workloads as W  # noqa: E402
from oracle import oracle as o  # noqa: E402  (checker only)

ap = argparse.ArgumentParser()
ap.add_argument("--seconds", type=float, default=60)
a = ap.parse_args()
rnd = random.Random(7)
t = (1 << 238).to_bytes(32, "little")
jobs, shares, switches, rolled_shares = {}, 0, 0, 0
bad = 0
rss0 = None
t0 = time.time()
with miner.Worker(miner.WorkerConfig(launch_log2=24)) as w:
    while time.time() - t0 < a.seconds:
        if rnd.random() < 0.15:
            w.set_work(None)
        else:
            i = rnd.randrange(1, 257)
            f = W.config2_fields(i)
            roll = {}
            if rnd.random() < 0.4:   # BIP-320 rolled work: goes through the version-shared kernel (groups of <= 32)
                roll = dict(version_mask=W.BIP320_MASK, version_begin=rnd.randrange(0, 60000),
                            version_count=rnd.choice([1, 2, 5, 32, 33, 70]), version_stride=rnd.choice([1, 3]))
            seq = w.set_work(miner.Work(job_id=switches & 0xFFFFFFFF, header=miner.Header(**f), target=t,
                                        repeat=rnd.random() < 0.5, nonce_begin=rnd.randrange(0, 1 << 31),
                                        nonce_count=1 << rnd.randrange(16, 30), **roll))
            jobs[seq] = (switches & 0xFFFFFFFF, i, roll)
        switches += 1
        time.sleep(rnd.choice([0, 0.0005, 0.002, 0.01]))
        for s in w.poll_shares(256, 0):
            shares += 1
            ok = s.work_seq in jobs and jobs[s.work_seq][0] == s.job_id
            if ok:
                f = W.config2_fields(jobs[s.work_seq][1])
                roll = jobs[s.work_seq][2]
                if roll:   # the version must be one of the slice's
                    want = {o.bip320_version(f["version"], W.BIP320_MASK, roll["version_begin"] + k * roll["version_stride"])
                            for k in range(roll["version_count"])}
                    ok = s.version in want
                    rolled_shares += 1
                else:
                    ok = s.version == f["version"]
                if ok and shares % 50 == 0:
                    oh = o.Header(s.version, f["prev_hash"], f["merkle_root"], f["time"], f["bits"], s.nonce)
                    ok = o.hash_header(oh) == s.hash and o.less_or_equal(s.hash, t)
            bad += not ok
        if switches == 2000:
            rss0 = resource.getrusage(resource.RUSAGE_SELF).ru_maxrss
        if len(jobs) > 5000:
            for k in sorted(jobs)[:2500]:
                del jobs[k]
    st = w.stats()
rss1 = resource.getrusage(resource.RUSAGE_SELF).ru_maxrss
print(f"{switches} set_work calls in {time.time() - t0:.0f} s, {shares} shares ({rolled_shares} of rolled works), {bad} bad, "
      f"hashes {st.hashes_total:.3e}, dropped {st.shares_dropped}, maxrss {rss0} -> {rss1} KiB")
assert bad == 0 and shares > 0
assert rss0 is None or rss1 - rss0 < 50_000, "host memory grew"
