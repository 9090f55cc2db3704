"""Deterministic synthetic jobs for BASELINE.json's configs.

Pure data: no hashing of candidates happens here.  ``S(x)`` is one SHA-256 of an
ASCII label (hashlib), used only to derive reproducible 32-byte fields.
"""
from __future__ import annotations

import hashlib
from typing import Dict, List

GENESIS_HEADER_HEX = ("0100000000000000000000000000000000000000000000000000000000000000000000003ba3edfd"
                      "7a7b12b27ac72c3e67768f617fc81bc3888a51323a9fb8aa4b1e5e4a29ab5f49ffff001d1dac2b7c")
GENESIS_NONCE = 2083236893
GENESIS_HASH_HEX = "6fe28c0ab6f1b372c1a6a246ae63f74f931e8365e15a089c68d6190000000000"
DIFF1_NBITS = 0x1D00FFFF
BIP320_MASK = 0x1FFFE000


def S(label: str) -> bytes:
    return hashlib.sha256(label.encode("ascii")).digest()


def s_stream(label: str, n: int) -> bytes:
    out = b""
    i = 0
    while len(out) < n:
        out += S(f"{label}/{i}")
        i += 1
    return out[:n]


def header_fields(raw80: bytes) -> Dict:
    return {"version": int.from_bytes(raw80[0:4], "little"), "prev_hash": raw80[4:36],
            "merkle_root": raw80[36:68], "time": int.from_bytes(raw80[68:72], "little"),
            "bits": int.from_bytes(raw80[72:76], "little"), "nonce": 0}


def genesis_fields() -> Dict:
    """configs[0]: the genesis header with the nonce zeroed."""
    return header_fields(bytes.fromhex(GENESIS_HEADER_HEX))


def config2_fields(i: int) -> Dict:
    """configs[1]: i = 0 is genesis, i in [1, 256] the synthetic headers."""
    if i == 0:
        return genesis_fields()
    return {"version": 0x20000000, "prev_hash": S(f"otedama-b200/prev/{i}"),
            "merkle_root": S(f"otedama-b200/merkle/{i}"), "time": 0x60000000 + i,
            "bits": DIFF1_NBITS, "nonce": 0}


CONFIG2_HEADERS = 257


def config3_v1() -> Dict:
    """configs[2]: replayed Stratum V1 mining.notify, 12-level merkle branch."""
    coinb1 = s_stream("otedama-b200/coinb1", 57) + bytes([0x20])  # ends in a script-length byte
    return {"coinb1": coinb1, "extranonce1": bytes.fromhex("deadbeef"), "extranonce2_size": 8,
            "coinb2": s_stream("otedama-b200/coinb2", 107),
            "branches": [S(f"otedama-b200/branch/{j}") for j in range(12)],
            "version": 0x20000000, "bits": DIFF1_NBITS, "time": 0x60000000,
            "prev_hash": S("otedama-b200/v1prev")}


def config4_v2() -> Dict:
    """configs[3]: SV2 standard-channel job with BIP-320 version rolling."""
    return {"version": 0x20000000, "merkle_root": S("otedama-b200/v2root"),
            "prev_hash": S("otedama-b200/v2prev"), "time": 0x60000000, "bits": DIFF1_NBITS,
            "mask": BIP320_MASK}


def partition(n_units: int, rank: int, world: int) -> List[int]:
    """Static zero-communication sharding: unit u -> rank u mod world."""
    return list(range(rank, n_units, world))
