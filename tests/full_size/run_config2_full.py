#!/usr/bin/env python
"""BASELINE.json configs[1] in full: genesis + 256 synthetic headers, the whole 2^32 nonce range each,
share difficulty 1, on one B200 (~2 min).  Every share is re-hashed by the CPU oracle (soundness, bit
exact); headers with a committed full CPU scan (tests/golden/*full_scan*.json) are compared for
completeness.  Writes gpurun_out/config2_full.json (the share set).
With --worker the same job goes through the worker C ABI instead (set_work per header with host structs,
wait, poll_shares): the end-to-end path at full size, timed by wall clock -> gpurun_out/config2_full_worker.json."""
import glob
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from otedama_b200 import miner, workloads as W  # noqa: E402
from oracle import oracle as o  # noqa: E402  (checker only)

target = miner.target_from_difficulty(1.0)
hdrs = [miner.Header(**{k: v for k, v in W.config2_fields(i).items()}) for i in range(W.CONFIG2_HEADERS)]
WORKER = "--worker" in sys.argv
t0 = time.time()
shares, hashes, dev_ms = [], 0, 0.0
if WORKER:
    with miner.Worker() as w:
        t0 = time.time()
        for i, h in enumerate(hdrs):
            w.set_work(miner.Work(job_id=i, header=h, nbits=h.bits, target=target))
            assert w.wait_exhausted(-1)
            for s in sorted(w.poll_shares(4096, 0), key=lambda s: s.nonce):
                assert s.job_id == i
                shares.append((i, s.nonce, s.hash.hex(), s.ntime, s.version))
        wall = time.time() - t0
        st = w.stats()
        hashes, dev_ms = st.hashes_total, st.device_ms
        assert st.shares_dropped == 0 and hashes == len(hdrs) << 32
else:
    for first in range(0, len(hdrs), 32):      # bursts of 32 headers: the share ring is read back per burst
        r = miner.scan_headers(hdrs[first:first + 32], target)
        assert r.found == len(r.shares)
        for s in r.shares:
            shares.append((first + s.job_id, s.nonce, s.hash.hex(), s.ntime, s.version))
        hashes += r.hashes
        dev_ms += r.device_ms
    wall = time.time() - t0
bad = 0
for idx, nonce, hh, ntime, version in shares:
    f = W.config2_fields(idx)
    oh = o.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], nonce)
    ok = o.hash_header(oh).hex() == hh and o.less_or_equal(bytes.fromhex(hh), target) and ntime == f["time"] and version == f["version"]
    bad += not ok
checked = {}
for p in glob.glob(os.path.join(ROOT, "tests", "golden", "*full_scan*.json")):
    g = json.load(open(p))
    idx = g.get("index", 0)
    want = [(x["nonce"], x["hash"]) for x in g["shares"]]
    got = [(n, h) for i, n, h, _, _ in shares if i == idx]
    checked[idx] = got == want
doc = {"what": "BASELINE configs[1] in full on one B200" + (", through the worker C ABI (end to end)" if WORKER else ""),
       "ghs_wall": hashes / wall / 1e9, "headers": len(hdrs), "hashes": hashes, "shares": len(shares),
       "expected_shares_poisson_mean": hashes * (0xFFFF / 2 ** 48), "shares_failing_oracle_recheck": bad,
       "device_seconds": dev_ms / 1e3, "ghs_device": hashes / dev_ms / 1e6, "wall_seconds": wall,
       "headers_with_full_cpu_cross_scan": {str(k): v for k, v in sorted(checked.items())},
       "share_set": [{"header": i, "nonce": n, "hash": h} for i, n, h, _, _ in shares]}
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(doc, open(os.path.join(ROOT, "gpurun_out", "config2_full_worker.json" if WORKER else "config2_full.json"), "w"), indent=1)
print({k: v for k, v in doc.items() if k != "share_set"})
assert bad == 0 and all(checked.values())
