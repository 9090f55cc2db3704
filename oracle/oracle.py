"""CPU oracle for the SHA-256d search path — TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module; the product package
``otedama_b200`` never does (tests/test_boundary.py checks that).

Two layers:

* a ``ctypes`` binding of ``oracle/otedama_oracle.c`` (hashing, header layout,
  ``LessOrEqual``, the striped ``grind`` loop, the V1 coinbase/merkle and BIP-320
  restatements) — the part that must be fast;
* the reference's big-number target math restated with Python integers
  (``TargetFromNBits``, ``NBitsFromTarget``, ``TargetFromDifficulty``,
  ``MeetsTarget`` — internal/miner/sha256d.go:134-257), plus a ``hashlib``-based
  second implementation of SHA256d used to cross-check the C code.

Parity: pinned by the reference's own vectors (see tests/test_oracle.py and
tests/golden/reference_vectors.json); the V1-coinbase and BIP-320 helpers are
"parity unpinned" (no reference code or test exists for them) — the V1 ones are
anchored to public data instead, re-derived with hashlib in the tests: blocks 170
and 100000 end to end, and the Stratum V1 documentation's own worked example (subscribe +
notify + accepted submit), which fixes the wire conventions.
"""
from __future__ import annotations

import ctypes
import hashlib
import math
import os
import struct
import subprocess
from dataclasses import dataclass
from typing import List, Sequence, Tuple

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libotedama_oracle.so")


class OracleError(ValueError):
    """Raised where the reference returns a non-nil error."""


# ----------------------------------------------------------------------------
# build / load
# ----------------------------------------------------------------------------
def build(force: bool = False) -> str:
    """Compile the C restatement in place (gcc only; a few hundred ms)."""
    src = os.path.join(_HERE, "otedama_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libotedama_oracle.so"],
                              stdout=subprocess.DEVNULL)
    return _LIB_PATH


class _Header(ctypes.Structure):
    _fields_ = [("version", ctypes.c_uint32), ("prev_hash", ctypes.c_uint8 * 32),
                ("merkle_root", ctypes.c_uint8 * 32), ("time", ctypes.c_uint32),
                ("bits", ctypes.c_uint32), ("nonce", ctypes.c_uint32)]


class _Share(ctypes.Structure):
    _fields_ = [("nonce", ctypes.c_uint32), ("ntime", ctypes.c_uint32),
                ("version", ctypes.c_uint32), ("hash", ctypes.c_uint8 * 32)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        L.orc_sha256.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p]
        L.orc_sha256d.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p]
        L.orc_header_bytes.argtypes = [ctypes.POINTER(_Header), ctypes.c_char_p]
        L.orc_parse_header.argtypes = [ctypes.c_char_p, ctypes.POINTER(_Header)]
        L.orc_hash_header.argtypes = [ctypes.POINTER(_Header), ctypes.c_char_p]
        L.orc_less_or_equal.argtypes = [ctypes.c_char_p, ctypes.c_char_p]
        L.orc_less_or_equal.restype = ctypes.c_int
        L.orc_grind.argtypes = [ctypes.POINTER(_Header), ctypes.c_char_p, ctypes.c_uint64,
                                ctypes.c_uint64, ctypes.c_int, ctypes.c_int,
                                ctypes.POINTER(_Share), ctypes.c_size_t,
                                ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)]
        L.orc_grind.restype = ctypes.c_size_t
        L.orc_v1_merkle_root.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p,
                                         ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t,
                                         ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p,
                                         ctypes.c_size_t, ctypes.c_char_p]
        L.orc_bip320_version.argtypes = [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32]
        L.orc_bip320_version.restype = ctypes.c_uint32
        L.orc_hash_headers.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p]
        L.orc_set_backend.argtypes = [ctypes.c_int]
        L.orc_set_backend.restype = ctypes.c_int
        L.orc_cpu_has_shani.restype = ctypes.c_int
        L.orc_backend.restype = ctypes.c_int
        L.orc_set_backend(-1)
        _lib = L
    return _lib


def set_backend(which: int) -> int:
    """0 = portable scalar, 1 = SHA-NI, -1 = best available. Returns the one in use."""
    return lib().orc_set_backend(which)


def backend_name() -> str:
    return "sha-ni" if lib().orc_backend() == 1 else "scalar"


# ----------------------------------------------------------------------------
# hashing  (reference: internal/miner/sha256d.go:107-117)
# ----------------------------------------------------------------------------
def sha256d(data: bytes) -> bytes:
    out = ctypes.create_string_buffer(32)
    lib().orc_sha256d(bytes(data), len(data), out)
    return out.raw


def sha256(data: bytes) -> bytes:
    out = ctypes.create_string_buffer(32)
    lib().orc_sha256(bytes(data), len(data), out)
    return out.raw


def sha256d_hashlib(data: bytes) -> bytes:
    """Independent second implementation (OpenSSL via hashlib) for cross-checks."""
    return hashlib.sha256(hashlib.sha256(data).digest()).digest()


@dataclass
class Header:
    """80-byte block header, reference layout internal/miner/sha256d.go:49-68."""
    version: int = 0
    prev_hash: bytes = b"\x00" * 32
    merkle_root: bytes = b"\x00" * 32
    time: int = 0
    bits: int = 0
    nonce: int = 0

    def _c(self) -> _Header:
        h = _Header()
        h.version, h.time, h.bits, h.nonce = (self.version & 0xFFFFFFFF, self.time & 0xFFFFFFFF,
                                              self.bits & 0xFFFFFFFF, self.nonce & 0xFFFFFFFF)
        h.prev_hash[:] = self.prev_hash
        h.merkle_root[:] = self.merkle_root
        return h

    def bytes(self) -> bytes:
        """Header.Bytes(), sha256d.go:59-68."""
        out = ctypes.create_string_buffer(80)
        c = self._c()
        lib().orc_header_bytes(ctypes.byref(c), out)
        return out.raw

    @staticmethod
    def parse(b: bytes) -> "Header":
        """ParseHeader, sha256d.go:71-80."""
        assert len(b) == 80
        c = _Header()
        lib().orc_parse_header(bytes(b), ctypes.byref(c))
        return Header(c.version, bytes(c.prev_hash), bytes(c.merkle_root), c.time, c.bits, c.nonce)


def hash_header(h: Header) -> bytes:
    """HashHeader, sha256d.go:114-117."""
    out = ctypes.create_string_buffer(32)
    c = h._c()
    lib().orc_hash_header(ctypes.byref(c), out)
    return out.raw


def hash_headers(headers: bytes) -> bytes:
    n = len(headers) // 80
    out = ctypes.create_string_buffer(32 * n)
    lib().orc_hash_headers(headers, n, out)
    return out.raw


def less_or_equal(h: bytes, other: bytes) -> bool:
    """Hash.LessOrEqual, sha256d.go:93-103 (MSB at index 31)."""
    return bool(lib().orc_less_or_equal(bytes(h), bytes(other)))


# ----------------------------------------------------------------------------
# target math  (reference: internal/miner/sha256d.go:134-257)
# ----------------------------------------------------------------------------
def _int_to_hash_le(v: int, what: str) -> bytes:
    if v.bit_length() > 256:
        raise OracleError(f"{what} target overflows 256 bits")
    return v.to_bytes(32, "little")


def target_from_nbits(nbits: int) -> bytes:
    """TargetFromNBits, sha256d.go:134-173: mant * 2^(8*(exp-3)) as 32 bytes LE."""
    nbits &= 0xFFFFFFFF
    exp = nbits >> 24
    mant = nbits & 0x007FFFFF
    if nbits & 0x00800000:
        raise OracleError(f"nBits 0x{nbits:08X} has negative mantissa bit set")
    if exp < 3:
        raise OracleError(f"nBits 0x{nbits:08X} exponent {exp} is below minimum 3")
    if mant == 0:
        raise OracleError(f"nBits 0x{nbits:08X} has zero mantissa")
    return _int_to_hash_le(mant << (8 * (exp - 3)), f"nBits 0x{nbits:08X}")


def nbits_from_target(target: bytes) -> int:
    """NBitsFromTarget, sha256d.go:178-205."""
    v = int.from_bytes(target, "little")
    if v == 0:
        return 0
    b = v.to_bytes((v.bit_length() + 7) // 8, "big")
    if b[0] & 0x80:
        b = b"\x00" + b
    exp = len(b) & 0xFF
    if len(b) == 1:
        mant = b[0]
    elif len(b) == 2:
        mant = b[0] << 8 | b[1]
    else:
        mant = b[0] << 16 | b[1] << 8 | b[2]
    return (exp << 24 | mant) & 0xFFFFFFFF


DIFF1_TARGET = 0xFFFF << 208  # sha256d.go:211


def _quo_round_prec(p: int, q: int, prec: int) -> Tuple[int, int]:
    """p/q (both > 0) correctly rounded (nearest-even) to a `prec`-bit mantissa,
    as math/big.Float.Quo does.  Returns (mant, exp2) with value = mant * 2**exp2."""
    s = max(0, prec + 4 - (p.bit_length() - q.bit_length()))
    t, r = divmod(p << s, q)
    sticky = r != 0
    shift = t.bit_length() - prec
    assert shift >= 2
    mant = t >> shift
    rem = t & ((1 << shift) - 1)
    half = 1 << (shift - 1)
    if rem > half or (rem == half and (sticky or (mant & 1))):
        mant += 1
    return mant, shift - s


def target_from_difficulty(difficulty: float) -> bytes:
    """TargetFromDifficulty, sha256d.go:226-247: diff1/difficulty in a 256-bit
    big.Float (round-to-nearest-even), then truncated toward zero."""
    if not (difficulty > 0) or math.isinf(difficulty):
        raise OracleError(f"invalid difficulty {difficulty}")
    num, den = float(difficulty).as_integer_ratio()      # exact
    mant, exp2 = _quo_round_prec(DIFF1_TARGET * den, num, 256)
    target = mant << exp2 if exp2 >= 0 else mant >> (-exp2)
    if target <= 0:
        raise OracleError(f"difficulty {difficulty} produces a non-positive target")
    return _int_to_hash_le(target, f"difficulty {difficulty}")


def meets_target(hash_: bytes, nbits: int) -> bool:
    """MeetsTarget, sha256d.go:251-257."""
    return less_or_equal(hash_, target_from_nbits(nbits))


# ----------------------------------------------------------------------------
# the search loop  (reference: Worker.grind, internal/miner/worker.go:216-282)
# ----------------------------------------------------------------------------
@dataclass(frozen=True)
class Share:
    nonce: int
    ntime: int
    version: int
    hash: bytes
    extranonce2: int = 0


def grind(tmpl: Header, target: bytes, first: int = 0, count: int = 1 << 32, threads: int = 1,
          midstate: bool = False, cap: int = 1 << 16) -> Tuple[List[Share], int, int]:
    """One pass of grind over nonces [first, first+count). Returns (shares sorted
    by nonce, dropped, hashes)."""
    out = (_Share * cap)()
    dropped = ctypes.c_uint64(0)
    hashes = ctypes.c_uint64(0)
    c = tmpl._c()
    n = lib().orc_grind(ctypes.byref(c), bytes(target), first, count, threads, int(midstate),
                        out, cap, ctypes.byref(dropped), ctypes.byref(hashes))
    shares = [Share(out[i].nonce, out[i].ntime, out[i].version, bytes(out[i].hash)) for i in range(n)]
    return shares, dropped.value, hashes.value


def grind_py(tmpl: Header, target: bytes, first: int, count: int) -> List[Share]:
    """Pure-Python/hashlib restatement for small windows (cross-checks the C loop)."""
    res = []
    prefix = tmpl.bytes()[:76]
    for n in range(first, first + count):
        n &= 0xFFFFFFFF
        h = sha256d_hashlib(prefix + struct.pack("<I", n))
        if int.from_bytes(h, "little") <= int.from_bytes(target, "little"):
            res.append(Share(n, tmpl.time, tmpl.version, h))
    return res


# ----------------------------------------------------------------------------
# UNPINNED helpers: Stratum V1 coinbase / merkle root, BIP-320 version rolling
# ----------------------------------------------------------------------------
def extranonce2_bytes(e: int, size: int) -> bytes:
    return bytes(((e >> (8 * i)) & 0xFF) if i < 8 else 0 for i in range(size))


def v1_merkle_root(coinb1: bytes, extranonce1: bytes, extranonce2: bytes, coinb2: bytes,
                   branches: Sequence[bytes]) -> bytes:
    root = ctypes.create_string_buffer(32)
    br = b"".join(branches)
    lib().orc_v1_merkle_root(coinb1, len(coinb1), extranonce1, len(extranonce1), extranonce2,
                             len(extranonce2), coinb2, len(coinb2), br, len(branches), root)
    return root.raw


def v1_merkle_root_py(coinb1, extranonce1, extranonce2, coinb2, branches) -> bytes:
    root = sha256d_hashlib(coinb1 + extranonce1 + extranonce2 + coinb2)
    for b in branches:
        root = sha256d_hashlib(root + b)
    return root


def bip320_version(base: int, mask: int, k: int) -> int:
    return lib().orc_bip320_version(base & 0xFFFFFFFF, mask & 0xFFFFFFFF, k & 0xFFFFFFFF)


def bip320_version_py(base: int, mask: int, k: int) -> int:
    v = base & ~mask & 0xFFFFFFFF
    for bit in range(32):
        if mask >> bit & 1:
            if k & 1:
                v |= 1 << bit
            k >>= 1
    return v


def v1_header(params, extranonce1: bytes, extranonce2: bytes) -> Header:
    """UNPINNED: the header a Stratum V1 miner hashes for mining.notify `params`
    ([job_id, prevhash, coinb1, coinb2, merkle_branch, version, nbits, ntime, clean]) —
    public convention: prevhash words byte-swapped, big-endian hex scalars."""
    _, prevhash, coinb1, coinb2, branch, version, nbits, ntime = params[:8]
    pw = bytes.fromhex(prevhash)
    prev = b"".join(pw[4 * i:4 * i + 4][::-1] for i in range(8))
    root = v1_merkle_root_py(bytes.fromhex(coinb1), extranonce1, extranonce2, bytes.fromhex(coinb2),
                             [bytes.fromhex(b) for b in branch])
    return Header(int(version, 16), prev, root, int(ntime, 16), int(nbits, 16), 0)
