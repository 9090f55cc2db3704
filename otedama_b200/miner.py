"""Host-side mirror of the reference's ``internal/miner`` package over the C ABI.

Names, argument meaning and error behaviour follow the Go package
(internal/miner/sha256d.go, internal/miner/worker.go) so that the parity tests
read like the reference's own; all compute happens in libotedama_b200.so on the
GPU.  The Go-side shim a maintainer would add is shown in INTEGRATION.md — this
module is the Python harness bound to the same entry points.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

from . import capi
from .capi import CHeader, CShare, CStats, CWork, CWorkerConfig, OtdError, check, lib

HEADER_SIZE = 80  # miner.HeaderSize, sha256d.go:33


class MinerError(ValueError):
    """The Go functions' non-nil ``error`` return."""


# ----------------------------------------------------------------------------
# Header / Hash  (sha256d.go:49-117)
# ----------------------------------------------------------------------------
@dataclass
class Header:
    version: int = 0
    prev_hash: bytes = b"\x00" * 32
    merkle_root: bytes = b"\x00" * 32
    time: int = 0
    bits: int = 0
    nonce: int = 0

    def _c(self) -> CHeader:
        if len(self.prev_hash) != 32 or len(self.merkle_root) != 32:
            raise MinerError("prev_hash and merkle_root must be 32 bytes")
        h = CHeader()
        h.version, h.time, h.bits, h.nonce = (self.version & 0xFFFFFFFF, self.time & 0xFFFFFFFF,
                                              self.bits & 0xFFFFFFFF, self.nonce & 0xFFFFFFFF)
        h.prev_hash[:] = self.prev_hash
        h.merkle_root[:] = self.merkle_root
        return h

    def bytes(self) -> bytes:
        """Header.Bytes(): the canonical 80-byte wire form."""
        out = ctypes.create_string_buffer(80)
        c = self._c()
        lib().otd_header_bytes(ctypes.byref(c), out)
        return out.raw


def parse_header(b: bytes) -> Header:
    """ParseHeader."""
    if len(b) != HEADER_SIZE:
        raise MinerError(f"header must be {HEADER_SIZE} bytes")
    c = CHeader()
    lib().otd_parse_header(bytes(b), ctypes.byref(c))
    return Header(c.version, bytes(c.prev_hash), bytes(c.merkle_root), c.time, c.bits, c.nonce)


def less_or_equal(h: bytes, other: bytes) -> bool:
    """Hash.LessOrEqual: little-endian 256-bit compare, MSB at index 31."""
    return bool(lib().otd_hash_less_or_equal(bytes(h), bytes(other)))


def hash_headers(headers: Sequence[bytes], device: int = 0) -> List[bytes]:
    """Batch HashHeader / SHA256d over 80-byte serialised headers, on the GPU."""
    n = len(headers)
    if n == 0:
        return []
    blob = b"".join(headers)
    if len(blob) != 80 * n:
        raise MinerError("every header must be 80 bytes")
    out = ctypes.create_string_buffer(32 * n)
    check(lib().otd_hash_headers(device, blob, n, out))
    raw = out.raw  # one copy; .raw re-copies the whole buffer on every access
    return [raw[32 * i:32 * i + 32] for i in range(n)]


def hash_header(h: Header, device: int = 0) -> bytes:
    """HashHeader(h): SHA256d of the serialised header."""
    return hash_headers([h.bytes()], device)[0]


# ----------------------------------------------------------------------------
# target math  (sha256d.go:134-257)
# ----------------------------------------------------------------------------
def _target_call(fn, arg) -> bytes:
    out = ctypes.create_string_buffer(32)
    rc = fn(arg, out)
    if rc < 0:
        raise MinerError(capi.last_error())
    return out.raw


def target_from_nbits(nbits: int) -> bytes:
    return _target_call(lib().otd_target_from_nbits, nbits & 0xFFFFFFFF)


def target_from_difficulty(difficulty: float) -> bytes:
    return _target_call(lib().otd_target_from_difficulty, float(difficulty))


def nbits_from_target(target: bytes) -> int:
    return int(lib().otd_nbits_from_target(bytes(target)))


def meets_target(hash_: bytes, nbits: int) -> bool:
    rc = lib().otd_meets_target(bytes(hash_), nbits & 0xFFFFFFFF)
    if rc < 0:
        raise MinerError(capi.last_error())
    return bool(rc)


# ----------------------------------------------------------------------------
# Work / Share / Stats  (worker.go:18-86)
# ----------------------------------------------------------------------------
@dataclass
class Work:
    job_id: int = 0
    channel_id: int = 0
    header: Header = field(default_factory=Header)
    nbits: int = 0
    target: bytes = b"\x00" * 32
    # search-space extension (defaults = the reference's nonce-only search)
    repeat: bool = False
    per_version_launch: bool = False   # rolled versions: plain kernel per version instead of the version-shared one
    nonce_begin: int = 0
    nonce_count: int = 0            # 0 = 2^32
    version_mask: int = 0
    version_begin: int = 0
    version_count: int = 0
    version_stride: int = 0
    coinb1: Optional[bytes] = None
    extranonce1: bytes = b""
    coinb2: bytes = b""
    merkle_branches: Sequence[bytes] = ()
    extranonce2_size: int = 0
    extranonce2_begin: int = 0
    extranonce2_count: int = 0
    extranonce2_stride: int = 0
    job_id_str: str = ""            # Stratum V1: the pool's opaque job id, echoed into the shares

    def _c(self):
        """Returns (CWork, keepalive) — keepalive holds the buffers CWork points into."""
        if len(self.target) != 32:
            raise MinerError("target must be 32 bytes")
        w = CWork()
        w.job_id, w.channel_id, w.nbits = self.job_id & 0xFFFFFFFF, self.channel_id & 0xFFFFFFFF, self.nbits & 0xFFFFFFFF
        w.header = self.header._c()
        w.target[:] = self.target
        jid = self.job_id_str.encode()
        if len(jid) >= capi.OTD_JOB_ID_MAX:
            raise MinerError("job_id_str too long")
        w.job_id_str = jid
        w.flags = (capi.OTD_WORK_REPEAT if self.repeat else 0) | \
                  (capi.OTD_WORK_PER_VERSION_LAUNCH if self.per_version_launch else 0)
        w.nonce_begin, w.nonce_count = self.nonce_begin, self.nonce_count
        w.version_mask, w.version_begin = self.version_mask, self.version_begin
        w.version_count, w.version_stride = self.version_count, self.version_stride
        keep = []
        if self.coinb1 is not None:
            def buf(b):
                a = ctypes.create_string_buffer(bytes(b), max(len(b), 1))
                keep.append(a)
                return ctypes.cast(a, ctypes.c_void_p)
            br = b"".join(self.merkle_branches)
            if len(br) != 32 * len(self.merkle_branches):
                raise MinerError("merkle branches must be 32 bytes each")
            w.coinb1, w.coinb1_len = buf(self.coinb1), len(self.coinb1)
            w.extranonce1, w.extranonce1_len = buf(self.extranonce1), len(self.extranonce1)
            w.coinb2, w.coinb2_len = buf(self.coinb2), len(self.coinb2)
            w.merkle_branches, w.n_branches = buf(br), len(self.merkle_branches)
            w.extranonce2_size = self.extranonce2_size
            w.extranonce2_begin, w.extranonce2_count = self.extranonce2_begin, self.extranonce2_count
            w.extranonce2_stride = self.extranonce2_stride
        return w, keep


@dataclass(frozen=True)
class Share:
    channel_id: int
    job_id: int
    nonce: int
    ntime: int
    version: int
    hash: bytes
    device_id: str = ""
    extranonce2: int = 0
    work_seq: int = 0
    job_id_str: str = ""


@dataclass
class Stats:
    hashes_total: int = 0
    shares_found: int = 0
    shares_dropped: int = 0
    uptime: float = 0.0        # seconds
    hash_rate: float = 0.0
    launches: int = 0
    candidates: int = 0
    device_ms: float = 0.0
    work_seq: int = 0
    exhausted: bool = False
    shared_launches: int = 0   # of `launches`: version-shared ones (rolled work)


def _device_id(index: int) -> str:
    return f"gpu-cuda{index}"


def _share_from_c(s: CShare, device_id: str) -> Share:
    return Share(s.channel_id, s.job_id, s.nonce, s.ntime, s.version, bytes(s.hash), device_id,
                 s.extranonce2, s.work_seq, s.job_id_str.decode("utf-8", "replace"))


# ----------------------------------------------------------------------------
# synchronous one-pass grind (tests / benchmarks)
# ----------------------------------------------------------------------------
@dataclass
class ScanResult:
    shares: List[Share]
    found: int
    hashes: int
    device_ms: float
    launches: int


def scan(work: Work, device: int = 0, variant: int = -1, max_shares: int = 1 << 16) -> ScanResult:
    """One pass of grind (worker.go:216-282) over the work's whole search space."""
    cw, keep = work._c()
    out = (CShare * max_shares)()
    found, hashes, launches = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64()
    ms = ctypes.c_double()
    n = check(lib().otd_scan(device, ctypes.byref(cw), variant, out, max_shares, ctypes.byref(found),
                             ctypes.byref(hashes), ctypes.byref(ms), ctypes.byref(launches)))
    did = _device_id(device)
    del keep
    return ScanResult([_share_from_c(out[i], did) for i in range(n)], found.value, hashes.value,
                      ms.value, launches.value)


def scan_headers(headers: Sequence[Header], target: bytes, nonce_begin: int = 0, nonce_count: int = 0,
                 device: int = 0, variant: int = -1, max_shares: int = 1 << 16) -> ScanResult:
    """Burst scan of many independent headers (one launch each, one event pair
    around all of them). Shares carry the header's index in ``job_id``."""
    n = len(headers)
    arr = (CHeader * n)(*[h._c() for h in headers])
    out = (CShare * max_shares)()
    found, hashes, launches = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64()
    ms = ctypes.c_double()
    k = check(lib().otd_scan_headers(device, arr, n, bytes(target), nonce_begin, nonce_count, variant, out,
                                     max_shares, ctypes.byref(found), ctypes.byref(hashes), ctypes.byref(ms),
                                     ctypes.byref(launches)))
    did = _device_id(device)
    return ScanResult([_share_from_c(out[i], did) for i in range(k)], found.value, hashes.value, ms.value,
                      launches.value)


def scan_variants() -> List[str]:
    return capi.names_list(lib().otd_scan_variants)


def default_variant() -> int:
    return int(lib().otd_scan_default_variant())


def plan_launches(work: Work, launch_log2: int = 0, max_launches: int = 1 << 16):
    """The kernel launches one pass over `work` is cut into (otd_plan_launches): a list of
    (first_instance, n_instances, version_shared, nonce_begin, nonce_count, extranonce2, version_index_first)
    and the total number of launches.  Host arithmetic only — works without a device."""
    cw, keep = work._c()
    out = (capi.CLaunchPlan * max(1, max_launches))()
    total = ctypes.c_uint64()
    n = check(lib().otd_plan_launches(ctypes.byref(cw), launch_log2, out, max_launches, ctypes.byref(total)))
    del keep
    return [(p.first_instance, p.n_instances, bool(p.version_shared), p.nonce_begin, p.nonce_count, p.extranonce2,
             p.version_index_first) for p in out[:n]], total.value


def vs_variants() -> List[str]:
    """Version-shared scan variants (used for work that rolls BIP-320 version bits)."""
    return capi.names_list(lib().otd_vs_variants)


def default_vs_variant() -> int:
    return int(lib().otd_vs_default_variant())


def vs_select(variant: int) -> None:
    """Variant for rolled work from now on: index, -1 = library default, -2 = never share."""
    check(lib().otd_vs_select(variant))


# ----------------------------------------------------------------------------
# Worker  (worker.go:91-212)
# ----------------------------------------------------------------------------
@dataclass
class WorkerConfig:
    """miner.WorkerConfig: Threads/NonceStep have no GPU meaning (the device covers
    every nonce exactly once per pass); DeviceID comes from the device index."""
    device: int = 0
    ring_capacity: int = 0
    queue_capacity: int = 0
    launch_log2: int = 0
    kernel_variant: int = -1


def default_worker_config() -> WorkerConfig:
    return WorkerConfig()


class Worker:
    """GPU counterpart of *miner.Worker. ``start`` may be called once (a second
    call raises, where the reference panics); ``stop`` joins and frees."""

    def __init__(self, cfg: Optional[WorkerConfig] = None):
        self.cfg = cfg or default_worker_config()
        self._h = None
        self._started = False
        self._keep = None

    # -- lifecycle -----------------------------------------------------------
    def start(self) -> "Worker":
        if self._started:
            raise RuntimeError("miner: Worker.Start called more than once")
        c = CWorkerConfig(self.cfg.ring_capacity, self.cfg.queue_capacity, self.cfg.launch_log2,
                          self.cfg.kernel_variant)
        h = lib().otd_worker_create(self.cfg.device, ctypes.byref(c))
        if not h:
            raise OtdError(capi.OTD_ENODEV, capi.last_error())
        self._h = ctypes.c_void_p(h)
        self._started = True
        return self

    def stop(self) -> None:
        """Stop(): safe even if start was never called."""
        if self._h is not None:
            lib().otd_worker_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self.start()

    def __exit__(self, *exc):
        self.stop()

    # -- work ----------------------------------------------------------------
    def set_work(self, work: Optional[Work]) -> int:
        self._need()
        if work is None:
            return check(lib().otd_worker_set_work(self._h, None))
        cw, keep = work._c()
        seq = check(lib().otd_worker_set_work(self._h, ctypes.byref(cw)))
        del keep  # the library copied everything
        return seq

    def has_work(self) -> bool:
        self._need()
        return bool(check(lib().otd_worker_has_work(self._h)))

    def device_id(self) -> str:
        return _device_id(self.cfg.device)

    # -- shares (the `<-chan Share` side) --------------------------------------
    def poll_shares(self, max_shares: int = 256, timeout_ms: int = 0) -> List[Share]:
        self._need()
        out = (CShare * max_shares)()
        n = check(lib().otd_worker_poll_shares(self._h, out, max_shares, timeout_ms))
        did = self.device_id()
        return [_share_from_c(out[i], did) for i in range(n)]

    def wait_exhausted(self, timeout_ms: int = -1) -> bool:
        self._need()
        return bool(check(lib().otd_worker_wait_exhausted(self._h, timeout_ms)))

    def stats(self) -> Stats:
        """Stats(): zero-value before start, like the reference."""
        if self._h is None:
            return Stats()
        s = CStats()
        check(lib().otd_worker_stats(self._h, ctypes.byref(s)))
        return Stats(s.hashes_total, s.shares_found, s.shares_dropped, s.uptime_ns * 1e-9, s.hash_rate,
                     s.launches, s.candidates, s.device_ms, s.work_seq, bool(s.exhausted), s.shared_launches)

    def _need(self):
        if self._h is None:
            raise RuntimeError("miner: Worker is not running")


def hash_rate_string(hps: float) -> str:
    """HashRateString, worker.go:285-299."""
    for lim, unit in ((1e12, "TH/s"), (1e9, "GH/s"), (1e6, "MH/s"), (1e3, "kH/s")):
        if hps >= lim:
            return f"{hps / lim:.2f} {unit}"
    return f"{hps:.0f} H/s"


# ----------------------------------------------------------------------------
# job-preparation kernels (exposed for parity tests)
# ----------------------------------------------------------------------------
@dataclass
class Instance:
    merkle_root: bytes
    version: int
    extranonce2: int
    midstate: bytes


def prepare_instances(work: Work, first: int, n: int, device: int = 0) -> List[Instance]:
    cw, keep = work._c()
    roots = ctypes.create_string_buffer(32 * n)
    mids = ctypes.create_string_buffer(32 * n)
    vers = (ctypes.c_uint32 * n)()
    en2 = (ctypes.c_uint64 * n)()
    check(lib().otd_prepare_instances(device, ctypes.byref(cw), first, n, roots, vers, en2, mids))
    del keep
    r, m = roots.raw, mids.raw
    return [Instance(r[32 * i:32 * i + 32], vers[i], en2[i], m[32 * i:32 * i + 32]) for i in range(n)]


def int_peak(mode: int, device: int = 0, iters: int = 2048):
    """(ops_per_s, sm_mhz, ms) of the integer-issue microbenchmark `mode`."""
    ops, mhz, ms = ctypes.c_double(), ctypes.c_double(), ctypes.c_double()
    check(lib().otd_int_peak(device, mode, iters, ctypes.byref(ops), ctypes.byref(mhz), ctypes.byref(ms)))
    return ops.value, mhz.value, ms.value


def int_peak_modes() -> List[str]:
    return capi.names_list(lib().otd_int_peak_modes)
