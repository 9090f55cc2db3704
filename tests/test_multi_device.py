"""Every entry point on a device other than 0 gives what device 0 gives (one worker and one
stream per device, nothing shared).  Needs a box with at least two GPUs (gpurun --gpus 2); skipped otherwise."""
import ctypes

import pytest

from otedama_b200 import workloads as W

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api(product):
    from otedama_b200 import capi, engine, hal, miner
    n = capi.lib().otd_device_count()
    if n <= 0:
        pytest.fail("no CUDA device: the gpu-marked tests must run on the B200 box")
    if n < 2:
        pytest.skip("single-GPU box")
    return capi, engine, hal, miner, n


def _key(s):
    return (s.channel_id, s.job_id, s.nonce, s.ntime, s.version, s.hash, s.extranonce2)


def test_scans_on_the_last_device_equal_device_0(api, oracle):
    capi, engine, hal, miner, n = api
    last = n - 1
    target = miner.target_from_difficulty(2.0 ** -16)
    # plain header
    f = W.config2_fields(3)
    plain = miner.Work(job_id=9, channel_id=2, nbits=f["bits"], target=target, nonce_begin=0xFFF00000,
                       nonce_count=1 << 20,
                       header=miner.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0))
    # V1 coinbase work with version rolling: K3 + K2 + K1 on that device
    c3 = W.config3_v1()
    rolled = miner.Work(job_id=3, header=miner.Header(c3["version"], c3["prev_hash"], b"\x00" * 32, c3["time"],
                                                      c3["bits"], 0),
                        nbits=c3["bits"], target=target, nonce_count=1 << 18, coinb1=c3["coinb1"],
                        extranonce1=c3["extranonce1"], coinb2=c3["coinb2"], merkle_branches=c3["branches"],
                        extranonce2_size=8, extranonce2_begin=5, extranonce2_count=3, version_mask=0x1FFFE000,
                        version_count=4)
    for work in (plain, rolled):
        a = miner.scan(work, device=0)
        b = miner.scan(work, device=last)
        assert len(a.shares) > 0 and a.found == b.found and a.hashes == b.hashes
        assert [_key(s) for s in a.shares] == [_key(s) for s in b.shares]
        assert {s.device_id for s in b.shares} == {hal.CudaDriver().enumerate()[last].identity().id}
    assert [(i.merkle_root, i.version, i.extranonce2, i.midstate) for i in miner.prepare_instances(rolled, 0, 12, device=0)] == \
           [(i.merkle_root, i.version, i.extranonce2, i.midstate) for i in miner.prepare_instances(rolled, 0, 12, device=last)]
    # batch hashing
    hdrs = [W.s_stream(f"otedama-b200/mdev/{i}", 80) for i in range(300)]
    got = miner.hash_headers(hdrs, device=last)
    assert b"".join(got) == oracle.hash_headers(b"".join(hdrs))
    # the no-allocation entry point keeps one scratch array per device
    cw, keep = plain._c()
    out = (capi.CShare * 4096)()
    found, hashes = ctypes.c_uint64(), ctypes.c_uint64()
    want = miner.scan(plain, device=last, max_shares=4096)
    for _ in range(2):
        k = capi.lib().otd_scan_noalloc(last, ctypes.byref(cw), -1, out, 4096, ctypes.byref(found), ctypes.byref(hashes))
        assert k == len(want.shares), capi.last_error()
        assert [(out[i].nonce, bytes(out[i].hash), out[i].device_index) for i in range(k)] == \
               [(s.nonce, s.hash, last) for s in want.shares]


def test_one_worker_per_device_covers_the_partitioned_space(api, oracle):
    capi, engine, hal, miner, n = api
    devices = hal.CudaDriver().enumerate()
    workers, shares = engine.start_miner_workers(devices)
    try:
        c4 = W.config4_v2()
        target = miner.target_from_difficulty(2.0 ** -14)
        whole = miner.Work(job_id=4, channel_id=1, nbits=c4["bits"], target=target, nonce_count=1 << 20,
                           header=miner.Header(c4["version"], c4["prev_hash"], c4["merkle_root"], c4["time"], c4["bits"], 0),
                           version_mask=c4["mask"], version_count=4 * n)
        want = sorted(_key(s) for s in miner.scan(whole, device=0).shares)
        for g, w in enumerate(workers):
            w.set_work(engine.partition_work(whole, g, n))
        for w in workers:
            assert w.wait_exhausted(60_000)
        got = []
        while True:
            batch = shares.poll(4096, 0)
            if not batch:
                break
            got += batch
        assert sorted(_key(s) for s in got) == want and len(want) > 0
        by_dev = {}
        for s in got:
            by_dev.setdefault(s.device_id, set()).add(s.version)
        assert len(by_dev) == n                                   # every device contributed
        all_versions = [v for vs in by_dev.values() for v in vs]
        assert len(all_versions) == len(set(all_versions))        # no version searched by two devices
    finally:
        for w in workers:
            w.stop()
