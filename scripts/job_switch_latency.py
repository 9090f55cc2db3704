#!/usr/bin/env python
"""Job-switch latency of the GPU worker (SURVEY §7 "hard parts"): while a full 2^32-candidate launch
of job A is running, set_work(job B) and time until the first share of job B reaches the host."""
import os
import statistics
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from otedama_b200 import miner, workloads as W  # noqa: E402

args = [x for x in sys.argv[1:] if x != "--rolled"]
rolled = "--rolled" in sys.argv   # job A rolls 64 BIP-320 versions: its launches are version-shared ones (32 x 2^27)
variant = int(args[0]) if args else -1                        # kernel variant index (-1 = default)
f1, f2 = W.config2_fields(40), W.config2_fields(41)
hard = (1).to_bytes(32, "little")
lat = []
with miner.Worker(miner.WorkerConfig(kernel_variant=variant)) as w:
    for rep in range(12):
        a = miner.Work(job_id=1, header=miner.Header(**f1), target=hard, repeat=True,
                       **(dict(version_mask=W.BIP320_MASK, version_count=64) if rolled else {}))
        b = miner.Work(job_id=2 + rep, header=miner.Header(**f2), target=b"\xff" * 32, nonce_count=1 << 16)
        w.set_work(a)
        time.sleep(0.15 + 0.03 * rep)          # somewhere inside a 0.46 s launch
        t0 = time.perf_counter()
        w.set_work(b)
        got = w.poll_shares(1, 5000)
        dt = time.perf_counter() - t0
        assert got and got[0].job_id == 2 + rep
        lat.append(dt * 1e3)
        w.wait_exhausted(5000)
        while w.poll_shares(4096, 0):
            pass
print(("job A = version-shared launches; " if rolled else "") + "set_work -> first share of the new job, ms:", [round(x, 2) for x in lat])
print(f"median {statistics.median(lat):.2f} ms, max {max(lat):.2f} ms (a full 2^32 launch is ~464 ms)")
