#!/usr/bin/env python
"""bench.py — SHA-256d proof-of-work search throughput (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload: BASELINE.json configs[1] — the genesis header plus 256 synthetic
headers, full 2^32 nonce range each, share difficulty 1.  One STEP is one header:
2^32 candidates through the hot path (midstate-folded double SHA-256, early
reject on the top hash word, ballot + atomic ring for the shares).  K steps walk
the header list; under torchrun rank r takes headers r, r+N, ... (static
partition, zero data-path communication — "weak" scaling).

Prints ONE JSON line (rank 0).  `value` = candidates of all ranks / max-over-ranks
device time of the K steps, CUDA events on the launching stream, job constants
resident before the timed region.  `e2e` = the same through the worker C ABI
(set_work -> wait -> poll_shares), host buffers, wall clock.  `roofline` is the
integer-issue roofline measured live on the same GPU; `cpu_baseline` is the CPU
restatement of the reference's grind loop on this box's host cores.

--impl reference times the reference's CPU arithmetic (the oracle port: serialise
+ 3 compressions per nonce, all host threads); the Go reference itself cannot be
built here (no Go toolchain), so there is no oracle/_ref.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ALG_OPS_PER_HASH = 2450          # SHF 1176 + LOP3 628 + IADD3 646
METRIC = "sha256d_hashrate"
UNIT = "GH/s"
WORKLOAD = ("configs[1]: genesis + 256 synthetic headers, full 2^32 nonce range each, share difficulty 1; "
            "one step = one header = 2^32 candidates")


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


# ---------------------------------------------------------------- clocks --
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return self
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        rows = [r for ts, r in self.rows if t0 <= ts <= t1 + 0.2] or [r for _, r in self.rows]
        sm, smax, power, reasons = [], [], [], set()
        for r in rows:
            try:
                sm.append(float(r[1])); smax.append(float(r[2])); power.append(float(r[3]))
            except (ValueError, IndexError):
                continue
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7),
                              ("sw_power_cap", 8)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------- CPU arms --
def cpu_extra_numbers():
    """Single-thread and midstate ("strong") CPU numbers, informational, ~2 s each."""
    from oracle import oracle as o
    from tests import workloads as wl
    h = wl.genesis_header()
    target = o.target_from_nbits(0x1D00FFFF)
    out = {}
    for name, threads, mid, count in (("single_thread_mhs", 1, False, 1 << 22),
                                      ("midstate_all_threads_mhs", os.cpu_count() or 1, True, 1 << 26)):
        t0 = time.time()
        _, _, n = o.grind(h, target, 0, count, threads=threads, midstate=mid)
        out[name] = n / (time.time() - t0) / 1e6
    return out


def cpu_grind_sample(seconds_target: float, threads: int):
    """The oracle's restatement of Worker.grind (reference arithmetic: serialise +
    3 compressions per nonce) on a bounded nonce window of the genesis header."""
    from oracle import oracle as o
    from tests import workloads as wl
    o.build()
    o.set_backend(-1)
    h = wl.genesis_header()
    target = o.target_from_nbits(0x1D00FFFF)
    t0 = time.time()
    _, _, n = o.grind(h, target, 0, 1 << 22, threads=threads)           # calibration
    rate = n / max(time.time() - t0, 1e-6)
    count = int(min(1 << 31, max(1 << 22, rate * seconds_target)))
    t0 = time.time()
    shares, _, n = o.grind(h, target, 2083236893 - count // 2, count, threads=threads)
    dt = time.time() - t0
    assert any(s.nonce == 2083236893 for s in shares), "CPU baseline lost the genesis nonce"
    return n / dt, n, dt, o.backend_name()


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (oracle port), all host threads."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    per_step = args.ref_seconds or max(0.5, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_grind_sample(per_step, threads)
    tot_n, tot_t, backend = 0, 0.0, "?"
    for _ in range(args.steps):
        rate, n, dt, backend = cpu_grind_sample(per_step, threads)
        tot_n += n
        tot_t += dt
    value = tot_n / tot_t / 1e9
    sample = (f"{tot_n} nonces of the genesis header around nonce 2083236893 ({args.steps} steps), "
              f"3 compressions per nonce like Worker.grind, {backend} compression, {cpu_model()}")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot_t / max(1, args.steps) * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "reference_arm": "CPU oracle port of internal/miner Worker.grind "
                       "(Go toolchain absent: the reference itself cannot be built)", "sample": sample},
            "cpu_baseline": {"value": value * 1e3, "unit": "MH/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------- GPU arm --
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--variant", type=int, default=-1, help="scan kernel variant (-1 = library default)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU baseline sample length")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ref-seconds", type=float, default=0.0, help="--impl reference: CPU seconds per step (0 = auto)")
    args = ap.parse_args()

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    dist = None
    if args.impl == "reference":
        run_reference(args, rank, world)   # rank 0 only; the other ranks exit 0 without work
        return
    if world > 1:
        import torch.distributed as dist  # control plane only: barrier + max/sum of scalars
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group(backend="gloo", rank=rank, world_size=world)

    from otedama_b200 import build
    if rank == 0:
        build.build()          # no-op when the in-tree .so matches the sources (content hash)
    if dist:
        dist.barrier()         # nobody loads the library while rank 0 might still be linking it
    from otedama_b200 import capi, miner, workloads as W
    ndev = capi.check(capi.lib().otd_device_count())
    if ndev <= 0:
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    dev = local % ndev
    variant = args.variant if args.variant >= 0 else miner.default_variant()
    vname = miner.scan_variants()[variant]
    target = miner.target_from_difficulty(1.0)

    def header(i):
        f = W.config2_fields(i % W.CONFIG2_HEADERS)
        return miner.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0)

    my_steps = [header(rank + s * world) for s in range(args.warmup + args.steps)]
    warm, timed = my_steps[:args.warmup], my_steps[args.warmup:]

    def barrier():
        if dist:
            dist.barrier()

    def reduce(x, op):
        if not dist:
            return x
        import torch
        t = torch.tensor([float(x)], dtype=torch.float64)
        dist.all_reduce(t, op=getattr(dist.ReduceOp, op))
        return float(t[0])

    # ---- kernel leg: inputs resident, one CUDA-event pair around the K launches ----
    if warm:
        miner.scan_headers(warm, target, device=dev, variant=variant)
    sampler = ClockSampler(dev).start() if rank == 0 else None
    barrier()
    t0 = time.time()
    r = miner.scan_headers(timed, target, device=dev, variant=variant)   # synchronises on return
    barrier()
    t1 = time.time()
    clocks = sampler.stop(t0, t1) if sampler else None
    dev_ms = reduce(r.device_ms, "MAX")
    hashes = reduce(r.hashes, "SUM")
    launches = int(reduce(r.launches, "SUM"))
    shares_found = int(reduce(r.found, "SUM"))
    value = hashes / (dev_ms * 1e-3) / 1e9

    # ---- end-to-end leg: worker C ABI, host buffers, wall clock ----
    import ctypes
    w = miner.Worker(miner.WorkerConfig(device=dev, kernel_variant=variant)).start()
    e2e_shares = 0
    for h in warm[:1]:
        w.set_work(miner.Work(header=h, nbits=h.bits, target=target))
        w.wait_exhausted(-1)
        w.poll_shares(4096, 0)
    barrier()
    e0 = time.time()
    for i, h in enumerate(timed):
        w.set_work(miner.Work(job_id=i, header=h, nbits=h.bits, target=target))   # host struct -> device params
        w.wait_exhausted(-1)
        e2e_shares += len(w.poll_shares(4096, 0))                                 # ring -> host shares
    barrier()
    e1 = time.time()
    st = w.stats()
    w.stop()
    e2e_s = reduce(e1 - e0, "MAX")
    e2e_value = hashes / e2e_s / 1e9
    h2d = ctypes.sizeof(capi.CWork)                      # the work struct handed across the C ABI
    d2h = 16 + 40 * (e2e_shares / max(1, len(timed)))    # ring header + ring entries per step

    if rank == 0:
        # ---- roofline: measured integer-issue peak on this GPU, same process ----
        modes = miner.int_peak_modes()
        peak_alu = max(miner.int_peak(modes.index("mix_alu"), device=dev, iters=4096)[0] for _ in range(3))
        peak_dual = max(miner.int_peak(modes.index("alu_fma"), device=dev, iters=4096)[0] for _ in range(3))
        per_gpu_hps = value * 1e9 / world
        achieved = per_gpu_hps * ALG_OPS_PER_HASH
        # Two ways to read the same measurement:
        #  - contract figure: 2450 algorithmic 3-input ops/hash against the ALU-pipe rate -> frac > 1,
        #    because ~650 of those ops are adds that this kernel issues as IMAD on the FMA pipe;
        #  - two-pipe figure: the 1804 shift/rotate/logic ops per hash that ONLY the ALU pipe can run
        #    (IMAD.WIDE does not co-issue) against the same ALU-pipe rate -> the real ceiling.
        alu_only_ops = 1804
        roofline = {"bound": "int-issue", "achieved": achieved / 1e12, "peak": peak_alu / 1e12,
                    "unit": "Tint-op/s", "frac": achieved / peak_alu, "traffic": 127488,
                    "note": ("achieved = per-GPU H/s x 2450 algorithmic 3-input integer ops/hash (SURVEY §8d); "
                             "peak = ALU-pipe rate (SHF/LOP3/IADD3 at the SHA mix) measured on this GPU in this run; "
                             "frac exceeds 1 because the kernel issues almost all of the ~650 adds/hash as IMAD on the "
                             "FMA pipe, which a one-pipe roofline does not model; traffic = dram__bytes_read+write "
                             "(bytes) of one 2^30-candidate launch from ncu --set full "
                             "(profiles/r1_ncu_scan_kernel.csv): effectively none"),
                    "frac_two_pipe": per_gpu_hps * alu_only_ops / peak_alu,
                    "two_pipe_note": ("H/s x 1804 ALU-only ops/hash (SHF 1176 + LOP3 628) / measured ALU-pipe rate: "
                                      "the ceiling once every add runs on the FMA pipe; ncu shows the ALU pipe 95.1% busy "
                                      "with 1852 ALU-pipe instructions per hash"),
                    "peak_dual_issue": peak_dual / 1e12, "alg_ops_per_hash": ALG_OPS_PER_HASH,
                    "roofline_ghs_per_gpu": peak_alu / ALG_OPS_PER_HASH / 1e9,
                    "two_pipe_roofline_ghs_per_gpu": peak_alu / alu_only_ops / 1e9}
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            rate, n, dt, backend = cpu_grind_sample(args.cpu_seconds, threads)
            extra = cpu_extra_numbers()
            cpu = {"value": rate / 1e6, "unit": "MH/s", "cores": threads, "kind": "port", **extra,
                   "sample": f"{n} nonces of the genesis header around nonce 2083236893 in {dt:.1f}s, 3 compressions "
                             f"per nonce like Worker.grind (no midstate), {backend} compression, {cpu_model()}"}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_ms / max(1, args.steps), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "kernel_variant": vname, "partition": f"header i -> rank i mod {world}",
                           "control_plane": "gloo barrier/max only; no data-path collective",
                           "l2": "n/a: the kernel reads no HBM in the hot loop (job constants are kernel parameters)",
                           "shares_found": shares_found},
                "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "seconds": e2e_s, "worker_hashes_total": st.hashes_total},
                "gpu_launches": launches, "roofline": roofline}
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if dist:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
