#!/usr/bin/env python
"""BASELINE.json configs[2..4] through the worker C ABI, time-boxed.

  config 3: replayed Stratum V1 notify — extranonce2 rolled on the device (K3), merkle root + midstate
            (K2), full 2^32 nonce range per extranonce2 value, 1 GPU.
  config 4: SV2 standard-channel job with BIP-320 version rolling (bits 13..28) x nonce, version index
            k -> GPU k mod G, one worker per GPU in this process.
  config 5: sustained sweep of config 3's space with extranonce2 sharded e -> GPU e mod G.

Every share that comes back is re-hashed by the CPU oracle (soundness, bit-exact).  The full spaces
(2^48 candidates) are far beyond a test run, so each config runs for --seconds and reports the rate.
    python tests/full_size/run_configs.py --gpus 1 --seconds 20
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from otedama_b200 import capi, miner, workloads as W  # noqa: E402
from oracle import oracle as o  # noqa: E402  (checker only)

ap = argparse.ArgumentParser()
ap.add_argument("--gpus", type=int, default=1)
ap.add_argument("--seconds", type=float, default=20.0)
ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "configs.json"))
ap.add_argument("--only", default="", help="comma list of 2,3,4 (BASELINE config indices) to run; default all")
a = ap.parse_args()
G = min(a.gpus, capi.check(capi.lib().otd_device_count()))
diff1 = miner.target_from_difficulty(1.0)


def run(name, works, check_share):
    """works[g] is the Work for GPU g. Runs for a.seconds, returns a result dict."""
    workers = [miner.Worker(miner.WorkerConfig(device=g)).start() for g in range(len(works))]
    t0 = time.time()
    for w, wk in zip(workers, works):
        w.set_work(wk)
    shares = []
    while time.time() - t0 < a.seconds:
        for w in workers:
            shares += w.poll_shares(256, 50)
    for w in workers:
        w.set_work(None)          # pause: aborts the launch in flight
    time.sleep(0.3)
    for w in workers:
        shares += w.poll_shares(4096, 0)
    stats = [w.stats() for w in workers]
    dt = time.time() - t0
    for w in workers:
        w.stop()
    bad = [s for s in shares if not check_share(s)]
    hashes = sum(s.hashes_total for s in stats)
    res = {"config": name, "gpus": len(works), "seconds": round(dt, 2), "hashes": hashes,
           "ghs_wall": hashes / dt / 1e9, "ghs_device": hashes / max(1e-9, max(s.device_ms for s in stats) * 1e-3) / 1e9,
           "launches": sum(s.launches for s in stats), "shares": len(shares), "shares_bad": len(bad),
           "expected_shares": hashes * (0xFFFF / 2 ** 48), "dropped": sum(s.shares_dropped for s in stats)}
    print(json.dumps(res), flush=True)
    assert not bad, bad[:3]
    return res


only = set(a.only.split(",")) if a.only else {"2", "3", "4"}
results = []
# ---- config 3 / 5: V1 coinbase, extranonce2 e -> GPU e mod G ----
c3 = W.config3_v1()


def v1_work(g, G):
    return miner.Work(job_id=3, header=miner.Header(c3["version"], c3["prev_hash"], b"\x00" * 32, c3["time"], c3["bits"], 0),
                      nbits=c3["bits"], target=diff1, coinb1=c3["coinb1"], extranonce1=c3["extranonce1"],
                      coinb2=c3["coinb2"], merkle_branches=c3["branches"], extranonce2_size=8,
                      extranonce2_begin=g, extranonce2_count=1 << 16, extranonce2_stride=G)


def check_v1(s):
    root = o.v1_merkle_root(c3["coinb1"], c3["extranonce1"], o.extranonce2_bytes(s.extranonce2, 8), c3["coinb2"],
                            c3["branches"])
    h = o.Header(s.version, c3["prev_hash"], root, s.ntime, c3["bits"], s.nonce)
    return o.hash_header(h) == s.hash and o.less_or_equal(s.hash, diff1) and s.version == c3["version"]


if "2" in only:
    results.append(run("configs[2] V1 notify replay, extranonce2 rolled on device, 1 GPU", [v1_work(0, 1)], check_v1))

# ---- config 4: BIP-320 version rolling, version index k -> GPU k mod G ----
c4 = W.config4_v2()


def v2_work(g, G):
    return miner.Work(job_id=4, channel_id=1, header=miner.Header(c4["version"], c4["prev_hash"], c4["merkle_root"],
                                                                  c4["time"], c4["bits"], 0),
                      nbits=c4["bits"], target=diff1, version_mask=c4["mask"], version_begin=g, version_stride=G)


def check_v2(s):
    h = o.Header(s.version, c4["prev_hash"], c4["merkle_root"], s.ntime, c4["bits"], s.nonce)
    ok_bits = (s.version & ~c4["mask"]) == (c4["version"] & ~c4["mask"])
    return ok_bits and o.hash_header(h) == s.hash and o.less_or_equal(s.hash, diff1)


if "3" in only:
    results.append(run(f"configs[3] SV2 + BIP-320 version rolling, {G} GPU(s)", [v2_work(g, G) for g in range(G)], check_v2))
if "4" in only and (G > 1 or a.only):
    results.append(run(f"configs[4] sustained sweep, extranonce2 sharded over {G} GPUs", [v1_work(g, G) for g in range(G)],
                       check_v1))
os.makedirs(os.path.dirname(a.out), exist_ok=True)
json.dump(results, open(a.out, "w"), indent=1)
