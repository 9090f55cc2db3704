#!/usr/bin/env python
"""Time-boxed sustained sweep (BASELINE configs[4] shape) with an interval log.

Runs the V1 work of configs[2] with extranonce2 sharded e -> GPU e mod G, one worker per GPU in this
process, for --seconds; every --interval seconds it logs, per GPU, the hash rate over the interval
(from otd_worker_stats) and power / SM clock / temperature (otd_device_telemetry = NVML), plus the
joules-per-terahash the reference's efficiency gauge would show (engine/metrics.go:488).  Every share
is re-hashed by the CPU oracle at the end.
    python tests/full_size/sustained.py --gpus 1 --seconds 600 --interval 10 --out gpurun_out/sustained_1gpu.csv
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from otedama_b200 import capi, hal, miner, workloads as W  # noqa: E402
from oracle import oracle as o  # noqa: E402  (checker only)

ap = argparse.ArgumentParser()
ap.add_argument("--gpus", type=int, default=1)
ap.add_argument("--seconds", type=float, default=600.0)
ap.add_argument("--interval", type=float, default=10.0)
ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sustained.csv"))
ap.add_argument("--versions", type=int, default=0,
                help="also roll this many BIP-320 versions per extranonce2 value (0 = extranonce2 x nonce only); "
                     "rolled work goes through the version-shared kernel")
a = ap.parse_args()

devices = hal.CudaDriver().enumerate()[:a.gpus]
G = len(devices)
c3 = W.config3_v1()
diff1 = miner.target_from_difficulty(1.0)
workers = [miner.Worker(miner.WorkerConfig(device=d.index)).start() for d in devices]
for g, w in enumerate(workers):
    w.set_work(miner.Work(job_id=5, header=miner.Header(c3["version"], c3["prev_hash"], b"\x00" * 32, c3["time"],
                                                        c3["bits"], 0),
                          nbits=c3["bits"], target=diff1, coinb1=c3["coinb1"], extranonce1=c3["extranonce1"],
                          coinb2=c3["coinb2"], merkle_branches=c3["branches"], extranonce2_size=8,
                          extranonce2_begin=g, extranonce2_count=1 << 16, extranonce2_stride=G,
                          version_mask=W.BIP320_MASK if a.versions else 0, version_count=a.versions))

rows = ["t_s,gpu,ghs_interval,power_w,sm_mhz,temp_c,joules_per_terahash"]
shares = []
t0 = time.time()
last_t = [t0] * G
last_h = [0] * G
next_tick = t0 + a.interval
while time.time() - t0 < a.seconds:
    for w in workers:
        shares += w.poll_shares(256, 50)
    now = time.time()
    if now >= next_tick:
        next_tick += a.interval
        for g, (d, w) in enumerate(zip(devices, workers)):
            st = w.stats()
            try:
                p, mhz, temp = d.telemetry()
            except Exception:
                p = mhz = temp = float("nan")
            ghs = (st.hashes_total - last_h[g]) / (now - last_t[g]) / 1e9
            last_h[g], last_t[g] = st.hashes_total, now
            jth = p / (ghs / 1e3) if ghs > 0 else float("nan")
            rows.append(f"{now - t0:.1f},{g},{ghs:.4f},{p:.1f},{mhz:.0f},{temp:.0f},{jth:.0f}")
for w in workers:
    w.set_work(None)
time.sleep(0.3)
for w in workers:
    shares += w.poll_shares(4096, 0)
stats = [w.stats() for w in workers]
dt = time.time() - t0
for w in workers:
    w.stop()


def ok(s):
    root = o.v1_merkle_root(c3["coinb1"], c3["extranonce1"], o.extranonce2_bytes(s.extranonce2, 8), c3["coinb2"],
                            c3["branches"])
    h = o.Header(s.version, c3["prev_hash"], root, s.ntime, c3["bits"], s.nonce)
    return o.hash_header(h) == s.hash and o.less_or_equal(s.hash, diff1)


bad = [s for s in shares if not ok(s)]
hashes = sum(s.hashes_total for s in stats)
ghs_cols = [float(r.split(",")[2]) for r in rows[1:]]
summary = {"gpus": G, "versions_per_extranonce2": a.versions, "one_process": True, "seconds": round(dt, 1), "hashes": hashes, "ghs_wall": hashes / dt / 1e9,
           "ghs_device": hashes / max(1e-9, max(s.device_ms for s in stats) * 1e-3) / 1e9,
           "interval_ghs_min": min(ghs_cols) if ghs_cols else None, "interval_ghs_max": max(ghs_cols) if ghs_cols else None,
           "shares": len(shares), "shares_bad": len(bad), "expected_shares": hashes * (0xFFFF / 2 ** 48),
           "dropped": sum(s.shares_dropped for s in stats),
           "extrapolated_2p48_sweep_minutes": (1 << 48) / max(1.0, hashes / dt) / 60.0}
os.makedirs(os.path.dirname(a.out), exist_ok=True)
with open(a.out, "w") as f:
    f.write("# " + json.dumps(summary) + "\n")
    f.write("\n".join(rows) + "\n")
print(json.dumps(summary))
assert not bad, bad[:3]
