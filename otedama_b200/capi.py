"""ctypes binding of include/otedama_b200.h (the C ABI of libotedama_b200.so).

Everything in this package goes through these entry points — the same ones a cgo
shim in the reference would bind.  There is no fallback: if the
shared library is missing or a symbol is absent, loading raises.
"""
from __future__ import annotations

import ctypes
import os
import re
from ctypes import (POINTER, Structure, c_char, c_char_p, c_double, c_int, c_int32, c_int64,
                    c_size_t, c_uint8, c_uint32, c_uint64, c_void_p)

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("OTD_LIB") or os.path.join(_PKG, "libotedama_b200.so")   # OTD_LIB: performance studies
HEADER_PATH = os.path.join(os.path.dirname(_PKG), "include", "otedama_b200.h")

OTD_OK, OTD_EINVAL, OTD_ENODEV, OTD_ENOMEM, OTD_EBUSY, OTD_ECUDA, OTD_ERANGE = 0, -22, -19, -12, -16, -5, -34
OTD_WORK_REPEAT = 1
OTD_WORK_PER_VERSION_LAUNCH = 2
OTD_NOALLOC_MAX_SHARES = 4096
OTD_JOB_ID_MAX = 64
OTD_VS_MAX_INSTANCES = 64      # version instances per version-shared launch (include/otedama_b200.h)


class OtdError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"[{code}] {msg}")
        self.code = code


class DeviceInfo(Structure):
    _fields_ = [("id", c_char * 32), ("model", c_char * 128), ("vendor", c_char * 16),
                ("pci_bus_id", c_char * 32), ("sm_count", c_int32), ("cc_major", c_int32),
                ("cc_minor", c_int32), ("sm_clock_khz", c_int32), ("total_mem_bytes", c_uint64),
                ("cap_sha256d", c_int32), ("cap_general_compute", c_int32)]


class CTelemetry(Structure):
    _fields_ = [("power_w", c_double), ("sm_clock_mhz", c_double), ("temperature_c", c_double)]


class CHeader(Structure):
    _fields_ = [("version", c_uint32), ("prev_hash", c_uint8 * 32), ("merkle_root", c_uint8 * 32),
                ("time", c_uint32), ("bits", c_uint32), ("nonce", c_uint32)]


class CWork(Structure):
    _fields_ = [("job_id", c_uint32), ("channel_id", c_uint32), ("header", CHeader), ("nbits", c_uint32),
                ("target", c_uint8 * 32), ("flags", c_uint32), ("nonce_begin", c_uint32),
                ("nonce_count", c_uint64), ("version_mask", c_uint32), ("version_begin", c_uint32),
                ("version_count", c_uint32), ("version_stride", c_uint32),
                ("coinb1", c_void_p), ("coinb1_len", c_uint32),
                ("extranonce1", c_void_p), ("extranonce1_len", c_uint32),
                ("coinb2", c_void_p), ("coinb2_len", c_uint32),
                ("merkle_branches", c_void_p), ("n_branches", c_uint32),
                ("extranonce2_size", c_uint32), ("extranonce2_begin", c_uint64),
                ("extranonce2_count", c_uint64), ("extranonce2_stride", c_uint64),
                ("job_id_str", c_char * OTD_JOB_ID_MAX)]


class CV1Notify(Structure):
    _fields_ = [("job_id", c_char_p), ("prevhash_hex", c_char_p), ("coinb1_hex", c_char_p), ("coinb2_hex", c_char_p),
                ("merkle_branch_hex", POINTER(c_char_p)), ("n_branches", c_uint32), ("version_hex", c_char_p),
                ("nbits_hex", c_char_p), ("ntime_hex", c_char_p), ("clean_jobs", c_int32)]


class CShare(Structure):
    _fields_ = [("channel_id", c_uint32), ("job_id", c_uint32), ("nonce", c_uint32), ("ntime", c_uint32),
                ("version", c_uint32), ("hash", c_uint8 * 32), ("extranonce2", c_uint64),
                ("work_seq", c_uint64), ("device_index", c_int32), ("reserved", c_int32),
                ("job_id_str", c_char * OTD_JOB_ID_MAX)]


class CV1Submit(Structure):
    _fields_ = [("extranonce2_hex", c_char * 33), ("ntime_hex", c_char * 9), ("nonce_hex", c_char * 9),
                ("version_bits_hex", c_char * 9), ("job_id", c_char * OTD_JOB_ID_MAX)]


class CLaunchPlan(Structure):
    _fields_ = [("first_instance", c_uint64), ("n_instances", c_uint32), ("version_shared", c_uint32),
                ("nonce_begin", c_uint32), ("version_index_first", c_uint32), ("nonce_count", c_uint64),
                ("extranonce2", c_uint64)]


class CWorkerConfig(Structure):
    _fields_ = [("ring_capacity", c_uint32), ("queue_capacity", c_uint32), ("launch_log2", c_uint32),
                ("kernel_variant", c_int32)]


class CStats(Structure):
    _fields_ = [("hashes_total", c_uint64), ("shares_found", c_uint64), ("shares_dropped", c_uint64),
                ("uptime_ns", c_uint64), ("hash_rate", c_double), ("launches", c_uint64),
                ("candidates", c_uint64), ("device_ms", c_double), ("work_seq", c_uint64),
                ("has_work", c_int32), ("exhausted", c_int32), ("shared_launches", c_uint64)]


_NAMES32 = (c_char * 32)

_SIGS = {
    "otd_last_error": (c_char_p, []),
    "otd_abi_version": (c_int, []),
    "otd_abi_struct_size": (c_int, [c_int]),
    "otd_device_count": (c_int, []),
    "otd_device_info": (c_int, [c_int, POINTER(DeviceInfo)]),
    "otd_device_telemetry": (c_int, [c_int, POINTER(CTelemetry)]),
    "otd_header_bytes": (None, [POINTER(CHeader), c_char_p]),
    "otd_parse_header": (None, [c_char_p, POINTER(CHeader)]),
    "otd_hash_less_or_equal": (c_int, [c_char_p, c_char_p]),
    "otd_target_from_nbits": (c_int, [c_uint32, c_char_p]),
    "otd_nbits_from_target": (c_uint32, [c_char_p]),
    "otd_target_from_difficulty": (c_int, [c_double, c_char_p]),
    "otd_meets_target": (c_int, [c_char_p, c_uint32]),
    "otd_hash_headers": (c_int, [c_int, c_char_p, c_size_t, c_char_p]),
    "otd_v1_job_target": (c_int, [c_uint32, c_double, c_char_p]),
    "otd_work_from_v2_job": (c_int, [c_uint32, c_uint32, c_uint32, c_char_p, c_char_p, c_uint32, c_uint32, c_char_p,
                                     POINTER(CWork)]),
    "otd_work_from_v1_job": (c_int, [c_char_p, c_uint32, c_char_p, c_uint32, c_uint32, c_double, POINTER(CWork)]),
    "otd_work_from_v1_notify": (c_int, [POINTER(CV1Notify), c_char_p, c_uint32, c_double, c_uint32, POINTER(CWork),
                                        c_char_p, c_size_t, POINTER(c_size_t)]),
    "otd_work_partition": (c_int, [POINTER(CWork), c_uint32, c_uint32, POINTER(CWork)]),
    "otd_share_to_v1_submit": (c_int, [POINTER(CShare), c_uint32, c_uint32, POINTER(CV1Submit)]),
    "otd_worker_create": (c_void_p, [c_int, POINTER(CWorkerConfig)]),
    "otd_worker_set_work": (c_int64, [c_void_p, POINTER(CWork)]),
    "otd_worker_poll_shares": (c_int, [c_void_p, POINTER(CShare), c_int, c_int]),
    "otd_worker_wait_exhausted": (c_int, [c_void_p, c_int]),
    "otd_worker_stats": (c_int, [c_void_p, POINTER(CStats)]),
    "otd_worker_has_work": (c_int, [c_void_p]),
    "otd_worker_device_index": (c_int, [c_void_p]),
    "otd_worker_destroy": (c_int, [c_void_p]),
    "otd_scan": (c_int, [c_int, POINTER(CWork), c_int, POINTER(CShare), c_uint32, POINTER(c_uint64),
                         POINTER(c_uint64), POINTER(c_double), POINTER(c_uint64)]),
    "otd_scan_headers": (c_int, [c_int, POINTER(CHeader), c_uint32, c_char_p, c_uint32, c_uint64, c_int,
                                 POINTER(CShare), c_uint32, POINTER(c_uint64), POINTER(c_uint64),
                                 POINTER(c_double), POINTER(c_uint64)]),
    "otd_scan_noalloc": (c_int, [c_int, POINTER(CWork), c_int, POINTER(CShare), c_uint32, POINTER(c_uint64),
                                 POINTER(c_uint64)]),
    "otd_scan_variants": (c_int, [c_void_p, c_int]),
    "otd_scan_default_variant": (c_int, []),
    "otd_plan_launches": (c_int, [POINTER(CWork), c_uint32, POINTER(CLaunchPlan), c_uint32, POINTER(c_uint64)]),
    "otd_vs_variants": (c_int, [c_void_p, c_int]),
    "otd_vs_default_variant": (c_int, []),
    "otd_vs_select": (c_int, [c_int]),
    "otd_prepare_instances": (c_int, [c_int, POINTER(CWork), c_uint32, c_uint32, c_char_p,
                                      POINTER(c_uint32), POINTER(c_uint64), c_char_p]),
    "otd_int_peak": (c_int, [c_int, c_int, c_int, POINTER(c_double), POINTER(c_double), POINTER(c_double)]),
    "otd_int_peak_modes": (c_int, [c_void_p, c_int]),
}


def declared_symbols() -> list:
    """Every function the public header declares (used by the export test)."""
    txt = open(HEADER_PATH).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(otd_[a-z0-9_]+)\s*\(", txt)))


_lib = None


def lib() -> ctypes.CDLL:
    """Load libotedama_b200.so (raises if it has not been built — no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OtdError(OTD_ENODEV, f"{LIB_PATH} is missing: run `python -m otedama_b200.build` "
                                       "(or __graft_entry__.build()); there is no CPU fallback")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def last_error() -> str:
    return lib().otd_last_error().decode("utf-8", "replace")


def check(rc: int) -> int:
    if rc < 0:
        raise OtdError(rc, last_error())
    return rc


def names_list(fn) -> list:
    n = fn(None, 0)
    buf = (_NAMES32 * n)()
    fn(ctypes.cast(buf, c_void_p), n)
    return [bytes(b.value).decode() for b in buf]
