#!/usr/bin/env python
"""Throughput of the batch HashHeader entry point (host buffers in, host buffers out)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes  # noqa: E402

from otedama_b200 import capi  # noqa: E402

n = 1 << 22
blob = os.urandom(80 * n)
out = ctypes.create_string_buffer(32 * n)
for _ in range(4):
    t = time.time()
    capi.check(capi.lib().otd_hash_headers(0, blob, n, out))  # host buffer -> host buffer
    dt = time.time() - t
    print(f"otd_hash_headers n=2^22 host->host: {n / dt / 1e6:.1f} M headers/s ({80 * n / dt / 1e9:.2f} GB/s in)")
