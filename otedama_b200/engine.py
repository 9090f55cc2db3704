"""Host-side mirror of the engine functions that feed the hot path
(internal/engine/run.go:1271-1364): job -> miner.Work -> Worker.SetWork.  The
work-building logic lives behind the C ABI (otd_work_from_v2_job /
otd_work_from_v1_job / otd_v1_job_target) so the Go shim binds the same code."""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Iterable

from . import capi, miner
from .capi import CWork, lib


class EngineError(ValueError):
    """The Go functions' non-nil ``error`` return."""


@dataclass
class NewMiningJob:
    """stratum.NewMiningJob (internal/stratum/messages.go:89-96)."""
    channel_id: int = 0
    job_id: int = 0
    has_min_ntime: bool = False
    min_ntime: int = 0
    version: int = 0
    merkle_root: bytes = b"\x00" * 32


@dataclass
class Job:
    """poolproto.Job as applyJob uses it (internal/poolproto/poolproto.go:140-169)."""
    job_id: str = ""
    ntime: int = 0
    nbits: int = 0
    version: int = 0
    prev_hash: bytes = b"\x00" * 32
    merkle_root: bytes = b"\x00" * 32
    clean_jobs: bool = False


def _work_from_c(cw: CWork) -> miner.Work:
    h = cw.header
    return miner.Work(job_id=cw.job_id, channel_id=cw.channel_id, nbits=cw.nbits, target=bytes(cw.target),
                      job_id_str=cw.job_id_str.decode("utf-8", "replace"),
                      header=miner.Header(h.version, bytes(h.prev_hash), bytes(h.merkle_root), h.time, h.bits, 0))


def v1_job_target(nbits: int, difficulty: float) -> bytes:
    """v1JobTarget (run.go:1313-1324)."""
    out = ctypes.create_string_buffer(32)
    if lib().otd_v1_job_target(nbits & 0xFFFFFFFF, float(difficulty), out) < 0:
        raise EngineError(capi.last_error())
    return out.raw


def work_from_v2_job(job: NewMiningJob, chan_id: int, prev_hash: bytes, prev_nbits: int, ntime: int,
                     share_target: bytes):
    """The Work updateWork builds, or None where the reference silently returns."""
    cw = CWork()
    rc = lib().otd_work_from_v2_job(job.job_id, chan_id, job.version & 0xFFFFFFFF, bytes(job.merkle_root),
                                    bytes(prev_hash), prev_nbits & 0xFFFFFFFF, ntime & 0xFFFFFFFF,
                                    bytes(share_target), ctypes.byref(cw))
    return _work_from_c(cw) if rc == 0 else None


def update_work(workers: Iterable, job: NewMiningJob, chan_id: int, prev_hash: bytes, prev_nbits: int, ntime: int,
                share_target: bytes) -> None:
    """updateWork (run.go:1271-1297): point every worker at the job."""
    w = work_from_v2_job(job, chan_id, prev_hash, prev_nbits, ntime, share_target)
    if w is None:
        return
    for wr in workers:
        wr.set_work(w)


def work_from_v1_job(job: Job, chan_id: int, difficulty: float) -> miner.Work:
    cw = CWork()
    rc = lib().otd_work_from_v1_job(job.job_id.encode(), chan_id, bytes(job.merkle_root), job.ntime & 0xFFFFFFFF,
                                    job.nbits & 0xFFFFFFFF, float(difficulty), ctypes.byref(cw))
    if rc < 0:
        raise EngineError(capi.last_error())
    return _work_from_c(cw)


def apply_job(workers: Iterable, job: Job, chan_id: int, difficulty: float) -> None:
    """applyJob (run.go:1340-1364); raises where the reference returns an error."""
    w = work_from_v1_job(job, chan_id, difficulty)
    for wr in workers:
        wr.set_work(w)


def work_from_v1_notify(params, extranonce1_hex: str, extranonce2_size: int, difficulty: float, chan_id: int = 0,
                        extranonce2_begin: int = 0, extranonce2_count: int = 0, extranonce2_stride: int = 1) -> miner.Work:
    """mining.notify params [job_id, prevhash, coinb1, coinb2, merkle_branch, version, nbits, ntime, clean]
    -> a Work whose extranonce2 is rolled on the device (otd_work_from_v1_notify)."""
    job_id, prevhash, coinb1, coinb2, branch, version, nbits, ntime, clean = params[:9]
    arr = (ctypes.c_char_p * max(1, len(branch)))(*[b.encode() for b in branch])
    n = capi.CV1Notify(str(job_id).encode(), prevhash.encode(), coinb1.encode(), coinb2.encode(), arr, len(branch),
                       version.encode(), nbits.encode(), ntime.encode(), int(bool(clean)))
    need = ctypes.c_size_t()
    cw = CWork()
    scratch = ctypes.create_string_buffer(len(coinb1) // 2 + len(coinb2) // 2 + len(extranonce1_hex) // 2 + 32 * len(branch) + 1)
    rc = lib().otd_work_from_v1_notify(ctypes.byref(n), extranonce1_hex.encode(), extranonce2_size, float(difficulty),
                                       chan_id, ctypes.byref(cw), scratch, len(scratch), ctypes.byref(need))
    if rc < 0:
        raise EngineError(capi.last_error())
    raw = scratch.raw
    l1, le1, l2 = cw.coinb1_len, cw.extranonce1_len, cw.coinb2_len
    w = _work_from_c(cw)
    w.coinb1 = raw[:l1]
    w.extranonce1 = raw[l1:l1 + le1]
    w.coinb2 = raw[l1 + le1:l1 + le1 + l2]
    base = l1 + le1 + l2
    w.merkle_branches = [raw[base + 32 * i:base + 32 * i + 32] for i in range(cw.n_branches)]
    w.extranonce2_size = cw.extranonce2_size
    w.extranonce2_begin = extranonce2_begin
    w.extranonce2_count = extranonce2_count or cw.extranonce2_count
    w.extranonce2_stride = extranonce2_stride
    return w


def v1_submit_params(share: miner.Share, work: miner.Work, worker_name: str = "otedama"):
    """The mining.submit parameter list for a share of a V1 work unit (stratumv1.go:336-346 with the
    rolled extranonce2 instead of zero padding; a sixth BIP-310 version_bits field when rolling)."""
    cs = capi.CShare(share.channel_id, share.job_id, share.nonce, share.ntime, share.version,
                     (ctypes.c_uint8 * 32)(*share.hash), share.extranonce2, share.work_seq, 0, 0,
                     share.job_id_str.encode())
    out = capi.CV1Submit()
    size = work.extranonce2_size if work.coinb1 is not None else 0
    if lib().otd_share_to_v1_submit(ctypes.byref(cs), size, work.version_mask, ctypes.byref(out)) < 0:
        raise EngineError(capi.last_error())
    params = [worker_name, out.job_id.decode(), out.extranonce2_hex.decode(), out.ntime_hex.decode(),
              out.nonce_hex.decode()]
    if work.version_mask:
        params.append(out.version_bits_hex.decode())
    return params


def partition_work(work: miner.Work, g: int, G: int) -> miner.Work:
    """The slice of `work` that device g of G searches (otd_work_partition): version indices,
    else extranonce2 values, else a contiguous nonce window."""
    import copy
    cw, keep = work._c()
    out = CWork()
    if lib().otd_work_partition(ctypes.byref(cw), g, G, ctypes.byref(out)) < 0:
        raise EngineError(capi.last_error())
    del keep
    w = copy.copy(work)
    w.nonce_begin, w.nonce_count = out.nonce_begin, out.nonce_count
    w.version_begin, w.version_count, w.version_stride = out.version_begin, out.version_count, out.version_stride
    w.extranonce2_begin, w.extranonce2_count, w.extranonce2_stride = (out.extranonce2_begin, out.extranonce2_count,
                                                                        out.extranonce2_stride)
    return w


# ----------------------------------------------------------------------------
# startMinerWorkers / mergeShares  (internal/engine/setup.go:56-78, fanin.go:20-68)
# ----------------------------------------------------------------------------
def start_miner_workers(devices, log=lambda level, msg: None, cfg: "miner.WorkerConfig | None" = None,
                        allow=None, per_device_cfg=None):
    """One started worker per SHA256d-capable device, in device order, plus one merged share source
    (startMinerWorkers, setup.go:56-78).  Devices without the capability are skipped; none left is an
    error, as in the reference.  The caller owns shutdown (``for w in workers: w.stop()``).

    What the reference has no knob for (its WorkerConfig is just a name, config.go:225-229) and a GPU
    box needs:
      allow           device-ID allow-list (``hal.Identity.ID`` strings, e.g. ``{"gpu-cuda0", "gpu-cuda3"}``);
                      None = every SHA256d device.  A listed ID that no detected device has is an error —
                      a typo must not silently leave a GPU idle.
      per_device_cfg  {device ID: miner.WorkerConfig} overriding ``cfg`` for that device (ring capacity,
                      host queue, launch size, kernel variant -> otd_worker_config_t)."""
    import dataclasses
    devices = list(devices)
    ids = [d.identity().id for d in devices]
    if allow is not None:
        allow = set(allow)
        unknown = sorted(allow - set(ids))
        if unknown:
            raise EngineError(f"engine: allow-list names devices that were not detected: {unknown}")
    per_device_cfg = dict(per_device_cfg or {})
    unknown = sorted(set(per_device_cfg) - set(ids))
    if unknown:
        raise EngineError(f"engine: per-device config for devices that were not detected: {unknown}")
    workers = []
    try:
        for dev, dev_id in zip(devices, ids):
            if not dev.capabilities().sha256d:
                continue
            if allow is not None and dev_id not in allow:
                log("info", f"engine: {dev_id} not in the device allow-list, no worker")
                continue
            base = per_device_cfg.get(dev_id, cfg)
            c = dataclasses.replace(base) if base is not None else miner.default_worker_config()
            c.device = getattr(dev, "index", c.device)
            workers.append(miner.Worker(c).start())
            log("info", f"engine: worker for {dev.identity()}")
    except Exception:
        for w in workers:
            w.stop()
        raise
    if not workers:
        raise EngineError("engine: no SHA256d-capable devices found")
    return workers, MergedShares(workers)


def pause_device(workers, device_id: str) -> bool:
    """applyAllocation's pauseDevice (arbitrate.go:255-262): SetWork(nil) on the ONE worker whose
    DeviceID matches; every other device keeps hashing.  Returns whether a worker matched."""
    for w in workers:
        if w.device_id() == device_id:
            w.set_work(None)
            return True
    return False


def update_work_partitioned(workers, work: miner.Work) -> None:
    """updateWork for G devices: device g searches slice g of G of the work's space
    (otd_work_partition: version indices, else extranonce2 values, else nonce windows)."""
    G = len(workers)
    for g, wr in enumerate(workers):
        wr.set_work(partition_work(work, g, G) if G > 1 else work)


class MergedShares:
    """mergeShares: the fan-in of every worker's share queue.  The Go version is a channel fed by one
    goroutine per worker; here ``poll`` drains the workers round-robin so that no device starves
    another, and returns when something arrived or the timeout passed."""

    def __init__(self, workers):
        self._workers = list(workers)
        self._next = 0

    def poll(self, max_shares: int = 256, timeout_ms: int = 0):
        import time
        deadline = time.monotonic() + timeout_ms / 1000.0
        out = []
        n = len(self._workers)
        while True:
            for k in range(n):
                w = self._workers[(self._next + k) % n]
                room = max_shares - len(out)
                if room <= 0:
                    break
                out.extend(w.poll_shares(room, 0))
            self._next = (self._next + 1) % max(n, 1)
            if out or time.monotonic() >= deadline:
                return out
            # block briefly on one worker's queue instead of spinning
            w = self._workers[self._next]
            out.extend(w.poll_shares(max_shares, min(5, max(1, int((deadline - time.monotonic()) * 1000)))))
            if out:
                return out
