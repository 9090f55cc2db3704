#!/usr/bin/env python
"""GPU probe: integer-issue microbenchmarks + scan-kernel variant throughput.
Writes gpurun_out/probe.json. Run on the B200 box: python scripts/gpu_probe.py"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from otedama_b200 import miner, hal, workloads as W  # noqa: E402

out = {"when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime())}
dev = hal.CudaDriver().enumerate()[0]
out["device"] = {"model": dev.info.model.decode(), "sms": dev.info.sm_count, "clock_khz": dev.info.sm_clock_khz}
print(out["device"], flush=True)

peaks = {}
for i, name in enumerate(miner.int_peak_modes() if "--no-peak" not in sys.argv else []):
    best = None
    for rep in range(3):
        ops, mhz, ms = miner.int_peak(i, iters=4096)
        if best is None or ops > best[0]:
            best = (ops, mhz, ms)
    sms = dev.info.sm_count
    peaks[name] = {"ops_per_s": best[0], "sm_mhz": best[1], "ms": best[2],
                   "ops_per_clk_per_sm": best[0] / (best[1] * 1e6) / sms}
    print(f"intpeak {name:10s} {best[0]/1e12:8.3f} Tops/s  {best[1]:7.1f} MHz  {peaks[name]['ops_per_clk_per_sm']:6.1f} ops/clk/SM", flush=True)
out["int_peak"] = peaks

f = W.genesis_fields()
hdr = miner.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0)
t = miner.target_from_nbits(W.DIFF1_NBITS)
scan = {}
for v, name in enumerate(miner.scan_variants()):
    rates = []
    for rep in range(3):
        r = miner.scan(miner.Work(header=hdr, target=t), variant=v)
        if not name.startswith("h_"):   # half-kernel experiments do not compute SHA-256d
            assert [s.nonce for s in r.shares] == [2083236893, 3197545707], r.shares
        rates.append(r.hashes / (r.device_ms * 1e-3))
    scan[name] = {"ghs": [x / 1e9 for x in rates]}
    print(f"scan {name:8s} GH/s {[round(x/1e9,3) for x in rates]}", flush=True)
out["scan"] = scan
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "probe.json"), "w"), indent=1)
