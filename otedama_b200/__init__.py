"""otedama_b200 — B200-native SHA-256d proof-of-work search behind Otedama's
internal/miner worker / internal/hal device interface.

The product is ``libotedama_b200.so`` (hand-written sm_100a CUDA + C++ host
runtime behind the C ABI in ``include/otedama_b200.h``); this package is the thin
Python harness over that ABI.  There is no CPU fallback anywhere in it.
"""
from . import capi, engine, hal, miner  # noqa: F401
from .capi import OtdError  # noqa: F401

__all__ = ["capi", "engine", "hal", "miner", "OtdError"]
