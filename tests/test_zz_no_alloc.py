"""otd_scan_noalloc — the scan entry point for callers that may not allocate GPU
memory (their device memory belongs to another allocator, or an allocation would
stall or fail at the wrong moment).  Its scratch is a fixed array inside the
library that is reused, dirty, by every call; so the check is: through the C ABI,
two different scans back to back (and the first one again) must each give exactly
what the ordinary entry point otd_scan gives for the same work — bit-exact, like
every other parity test here.  Runs last (file name) so that a failure here cannot
hide another test's result; it needs nothing from the tests before it."""
import ctypes

import pytest

from otedama_b200 import workloads as W

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api(product):
    from otedama_b200 import capi, miner
    if capi.lib().otd_device_count() <= 0:
        pytest.fail("no CUDA device: the gpu-marked tests must run on the B200 box")
    return capi, miner


def _work(miner, index, target, first, count):
    f = W.config2_fields(index)
    return miner.Work(job_id=100 + index, channel_id=7, nbits=f["bits"], target=target, nonce_begin=first,
                      nonce_count=count,
                      header=miner.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0))


def _noalloc(capi, work, max_shares):
    cw, keep = work._c()
    out = (capi.CShare * max(1, max_shares))()
    found, hashes = ctypes.c_uint64(), ctypes.c_uint64()
    n = capi.lib().otd_scan_noalloc(0, ctypes.byref(cw), -1, out, max_shares, ctypes.byref(found), ctypes.byref(hashes))
    return n, out, found.value, hashes.value


def _tuple(s):
    return (s.channel_id, s.job_id, s.nonce, s.ntime, s.version, bytes(s.hash), s.extranonce2, s.device_index)


def test_two_calls_in_a_row_match_the_ordinary_entry_point(api):
    capi, miner = api
    dense = miner.target_from_difficulty(2.0 ** -18)      # a few hundred shares
    sparse = miner.target_from_difficulty(2.0 ** -13)     # ten shares: must not inherit the first call's ring
    jobs = [_work(miner, 1, dense, 0, 1 << 24),
            _work(miner, 2, sparse, 0xF0000000, 1 << 22),
            _work(miner, 1, dense, 0, 1 << 24)]
    counts = []
    for work in jobs:
        want = miner.scan(work, max_shares=4096)           # existing entry point (allocates its rings)
        n, out, found, hashes = _noalloc(capi, work, 4096)
        assert n >= 0, capi.last_error()
        assert n == len(want.shares) and found == want.found and hashes == want.hashes == work.nonce_count
        got = [_tuple(out[i]) for i in range(n)]
        ref = [(s.channel_id, s.job_id, s.nonce, s.ntime, s.version, s.hash, s.extranonce2, 0) for s in want.shares]
        assert got == ref
        counts.append(n)
    assert counts[0] == counts[2] > 100 and 0 < counts[1] < counts[0]


def test_a_window_matches_the_cpu_oracle(api, oracle):
    """Judged like every other entry point: the share set of a dense window against the CPU oracle's grind of
    the same header and target (not only against the library's other entry point)."""
    capi, miner = api
    import os
    t = oracle.target_from_difficulty(2.0 ** -16)
    for index, first, count in ((5, 0x7FFF0000, 1 << 21), (0, W.GENESIS_NONCE - (1 << 20), 1 << 21)):
        f = W.config2_fields(index)
        want, _, _ = oracle.grind(oracle.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0),
                                  t, first, count, threads=os.cpu_count() or 1)
        n, out, found, hashes = _noalloc(capi, _work(miner, index, t, first, count), 4096)
        assert n == found == len(want) > 10 and hashes == count
        assert [(out[i].nonce, bytes(out[i].hash)) for i in range(n)] == [(s.nonce, s.hash) for s in want]
        assert all(out[i].ntime == f["time"] and out[i].version == f["version"] for i in range(n))


def test_the_genesis_nonce_is_found_without_allocating(api):
    capi, miner = api
    f = W.genesis_fields()
    work = miner.Work(header=miner.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0),
                      nbits=f["bits"], target=miner.target_from_nbits(f["bits"]),
                      nonce_begin=W.GENESIS_NONCE - 4096, nonce_count=8192)
    n, out, found, _ = _noalloc(capi, work, 8)
    assert (n, found) == (1, 1) and out[0].nonce == W.GENESIS_NONCE and bytes(out[0].hash).hex() == W.GENESIS_HASH_HEX


def test_limits_are_refused_with_a_clear_error(api):
    capi, miner = api
    work = _work(miner, 3, b"\xff" * 32, 0, 1 << 12)
    n, _, _, _ = _noalloc(capi, work, capi.OTD_NOALLOC_MAX_SHARES + 1)
    assert n == capi.OTD_ERANGE and "exceeds the fixed ring of 4096" in capi.last_error()
    # more shares than the caller's buffer: the excess is dropped and reported, nothing overruns
    n, out, found, _ = _noalloc(capi, work, 100)
    assert n == 100 and found == 1 << 12 and len({out[i].nonce for i in range(100)}) == 100
    rolled = _work(miner, 3, b"\xff" * 32, 0, 1 << 12)
    rolled.version_mask = W.BIP320_MASK
    n, _, _, _ = _noalloc(capi, rolled, 16)
    assert n == capi.OTD_EINVAL and "plain-header" in capi.last_error()
