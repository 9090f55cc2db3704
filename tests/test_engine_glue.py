"""engine.updateWork / applyJob / v1JobTarget mirrors (internal/engine/run.go:1271-1364),
modelled on internal/engine/run_test.go:356-515.  Host logic only — no GPU."""
import pytest


class FakeWorker:
    """SetWork is safe without Start in the reference; record what was set."""

    def __init__(self):
        self.work = None

    def set_work(self, w):
        self.work = w


@pytest.fixture()
def eng(product):
    from otedama_b200 import engine, miner
    return engine, miner


def test_update_work_populates_full_header_and_share_target(eng):
    engine, miner = eng
    job = engine.NewMiningJob(channel_id=1, job_id=42, version=0x20000004,
                              merkle_root=bytes((i * 7) & 0xFF for i in range(32)))
    prev = bytes(i + 1 for i in range(32))
    easiest = b"\xff" * 32
    ws = [FakeWorker(), FakeWorker()]
    engine.update_work(ws, job, 1, prev, 0x1D00FFFF, 0x60000000, easiest)
    for w in ws:
        wk = w.work
        assert (wk.job_id, wk.channel_id, wk.nbits, wk.target) == (42, 1, 0x1D00FFFF, easiest)
        h = wk.header
        assert (h.version, h.prev_hash, h.merkle_root, h.time, h.bits) == (0x20000004, prev, job.merkle_root,
                                                                            0x60000000, 0x1D00FFFF)


def test_update_work_zero_share_target_falls_back_to_network_target(eng, oracle):
    engine, miner = eng
    w = FakeWorker()
    engine.update_work([w], engine.NewMiningJob(channel_id=1, job_id=1, version=0x20000000), 1, b"\x00" * 32,
                       0x1D00FFFF, 0x495FAB29, b"\x00" * 32)
    assert w.work.target == oracle.target_from_nbits(0x1D00FFFF)
    # malformed nBits with no share target: the reference returns without touching the workers
    w2 = FakeWorker()
    engine.update_work([w2], engine.NewMiningJob(job_id=1), 1, b"\x00" * 32, 0x00000000, 1, b"\x00" * 32)
    assert w2.work is None


def test_apply_job_valid_unparseable_and_bad_nbits(eng):
    engine, miner = eng
    ws = [FakeWorker(), FakeWorker()]
    engine.apply_job(ws, engine.Job(job_id="42", ntime=0x60000000, nbits=0x1D00FFFF), 1, 0)
    assert all(w.work.job_id == 42 and w.work.header.time == 0x60000000 for w in ws)
    # the reference's applyJob copies merkle root / time / bits only (run.go:1349-1359)
    assert ws[0].work.header.version == 0 and ws[0].work.header.prev_hash == b"\x00" * 32
    with pytest.raises(engine.EngineError, match="unparseable job ID"):
        engine.apply_job([FakeWorker()], engine.Job(job_id="not-a-number", nbits=0x1D00FFFF), 1, 0)
    with pytest.raises(engine.EngineError, match="bad target"):
        engine.apply_job([FakeWorker()], engine.Job(job_id="1", nbits=0x00000000), 1, 0)
    engine.apply_job([FakeWorker()], engine.Job(job_id="1", nbits=0x1D00FFFF), 1, 0.001)
    # Sscanf("%d") accepts a numeric prefix
    w = FakeWorker()
    engine.apply_job([w], engine.Job(job_id="60abc", nbits=0x1D00FFFF), 1, 0)
    assert w.work.job_id == 60


def test_v1_job_target_selection(eng, oracle):
    engine, miner = eng
    nbits = 0x1D00FFFF
    assert engine.v1_job_target(nbits, 0) == oracle.target_from_nbits(nbits)
    got = engine.v1_job_target(nbits, 0.001)
    assert got == oracle.target_from_difficulty(0.001) and got != oracle.target_from_nbits(nbits)
    with pytest.raises(engine.EngineError):
        engine.v1_job_target(0x00000000, 0.001)
    # a difficulty whose conversion fails keeps the nBits target (derr != nil branch)
    assert engine.v1_job_target(nbits, 1e-300) == oracle.target_from_nbits(nbits)
    assert engine.v1_job_target(nbits, float("nan")) == oracle.target_from_nbits(nbits)


@pytest.mark.gpu
def test_update_work_share_echoes_version_and_ntime_on_the_gpu(eng):
    """TestUpdateWork_PopulatesFullHeaderAndShareTarget with a real (GPU) worker."""
    engine, miner = eng
    job = engine.NewMiningJob(channel_id=1, job_id=42, version=0x20000004,
                              merkle_root=bytes((i * 7) & 0xFF for i in range(32)))
    with miner.Worker(miner.WorkerConfig(launch_log2=16)) as w:
        engine.update_work([w], job, 1, bytes(i + 1 for i in range(32)), 0x1D00FFFF, 0x60000000, b"\xff" * 32)
        s = w.poll_shares(1, 3000)
        assert s, "no share within 3s at the easiest share target"
        assert (s[0].job_id, s[0].version, s[0].ntime, s[0].channel_id) == (42, 0x20000004, 0x60000000, 1)


# the reference's own (syntactic) notify fixture, internal/poolproto/stratumv1/stratumv1_test.go:71-81
REF_NOTIFY = ["60", "4d16b6f85af6e2198f44ae2a6de67f78487ae5611b77c6c0440b921e00000000",
              "01000000010000000000000000000000000000000000000000000000000000000000000000ffffffff20",
              "ffffffff0100f2052a010000004341041b0e8c2567c12536aa13357b79a073dc4444acb83c4ec7a0e2f99dd7457516c5817242da796924ca4e99947d087fedf9ce467cb9f7c6287078f801df276fdf84ac00000000",
              [], "00000002", "1d00ffff", "68d36c5e", True]


def test_v1_notify_to_work_unit(eng, oracle, golden):
    engine, miner = eng
    rv = golden["reference_vectors"]["v1_notify_fixture"]
    assert REF_NOTIFY[2] == rv["coinb1"] and REF_NOTIFY[3] == rv["coinb2"]
    w = engine.work_from_v1_notify(REF_NOTIFY, "deadbeef", 4, 1024.0, chan_id=3)
    # the fields the reference's parseNotify also extracts (parse.go:77-90)
    assert (w.job_id, w.channel_id, w.header.version, w.nbits, w.header.bits, w.header.time) == \
           (60, 3, 2, 0x1D00FFFF, 0x1D00FFFF, 0x68D36C5E)
    assert w.target == oracle.target_from_difficulty(1024.0)        # v1JobTarget with set_difficulty 1024
    want = oracle.v1_header(REF_NOTIFY, bytes.fromhex("deadbeef"), b"\x00" * 4)
    assert w.header.prev_hash == want.prev_hash and w.header.prev_hash.endswith(b"\x00" * 4)
    assert (w.coinb1.hex(), w.coinb2.hex(), w.extranonce1.hex(), list(w.merkle_branches)) == \
           (REF_NOTIFY[2], REF_NOTIFY[3], "deadbeef", [])
    assert (w.extranonce2_size, w.extranonce2_count) == (4, 1 << 32)
    assert engine.work_from_v1_notify(REF_NOTIFY, "", 8, 0).extranonce2_count == 1 << 40
    # with branches, random content: C decoder vs the Python restatement
    import random
    rnd = random.Random(5)
    params = ["7", rnd.randbytes(32).hex(), rnd.randbytes(41).hex(), rnd.randbytes(77).hex(),
              [rnd.randbytes(32).hex() for _ in range(12)], "20000000", "1d00ffff", "60000000", False]
    w = engine.work_from_v1_notify(params, "0a0b", 8, 0, extranonce2_begin=5, extranonce2_count=3)
    oh = oracle.v1_header(params, bytes.fromhex("0a0b"), (5).to_bytes(8, "little"))
    assert (w.header.version, w.header.prev_hash, w.header.time, w.header.bits) == (oh.version, oh.prev_hash, oh.time, oh.bits)
    assert [b.hex() for b in w.merkle_branches] == params[4] and w.target == oracle.target_from_nbits(0x1D00FFFF)
    assert engine.work_from_v1_notify(["x"] + params[1:], "", 4, 0).job_id_str == "x"     # job ids are opaque strings
    for bad in (["x" * 64] + params[1:], params[:1] + ["00"] + params[2:], params[:5] + ["2000000"] + params[6:],
                params[:4] + [["zz" * 32]] + params[5:], params[:2] + ["abc"] + params[3:]):
        with pytest.raises(engine.EngineError):
            engine.work_from_v1_notify(bad, "", 4, 0)


@pytest.mark.gpu
def test_v1_notify_work_unit_scans_like_the_oracle(eng, oracle):
    """notify -> work unit -> device (K3 roll + K2 midstate + K1 scan) vs the oracle's header."""
    engine, miner = eng
    import os
    import random
    rnd = random.Random(9)
    params = ["12", rnd.randbytes(32).hex(), rnd.randbytes(58).hex(), rnd.randbytes(107).hex(),
              [rnd.randbytes(32).hex() for _ in range(12)], "20000000", "1d00ffff", "60000000", True]
    w = engine.work_from_v1_notify(params, "deadbeef", 8, 2.0 ** -18, extranonce2_begin=100, extranonce2_count=6)
    w.nonce_count = 1 << 16
    r = miner.scan(w)
    want = []
    for e in range(100, 106):
        oh = oracle.v1_header(params, bytes.fromhex("deadbeef"), e.to_bytes(8, "little"))
        sh, _, _ = oracle.grind(oh, w.target, 0, 1 << 16, threads=os.cpu_count() or 1)
        want += [(e, s.nonce, s.hash) for s in sh]
    assert sorted((s.extranonce2, s.nonce, s.hash) for s in r.shares) == sorted(want) and len(want) > 10
    assert all(s.job_id == 12 and s.ntime == 0x60000000 and s.version == 0x20000000 for s in r.shares)


@pytest.mark.gpu
def test_block_170_is_found_from_its_notify_form(eng, oracle, golden):
    """Real chain data end to end: block 170 expressed as a Stratum V1 notify (coinbase split around
    its extranonce byte, one merkle branch) -> work unit -> K3/K2/K1 on the GPU must find exactly the
    mined (extranonce2, nonce) with the published block hash."""
    engine, miner = eng
    b = golden["reference_vectors"]["real_world_anchors"]["block170"]
    cb = bytes.fromhex(b["coinbase_tx_hex"])
    k = b["stratum_split"]["coinb1_len"]
    prev = bytes.fromhex(b["prev_hash_display"])[::-1]
    prev_wire = b"".join(prev[4 * i:4 * i + 4][::-1] for i in range(8)).hex()
    params = ["170", prev_wire, cb[:k].hex(), cb[k + 1:].hex(), [bytes.fromhex(b["second_txid_display"])[::-1].hex()],
              "%08x" % b["version"], b["bits"], "%08x" % b["time"], True]
    w = engine.work_from_v1_notify(params, "", 1, 1.0, extranonce2_begin=0, extranonce2_count=4)
    w.nonce_begin, w.nonce_count = b["nonce"] - (1 << 20), 1 << 21
    r = miner.scan(w)
    assert [(s.extranonce2, s.nonce, s.hash[::-1].hex()) for s in r.shares] == [(2, b["nonce"], b["hash_display"])]
    assert r.hashes == 4 << 21 and r.launches == 4
    inst = miner.prepare_instances(w, 2, 1)[0]
    assert inst.merkle_root[::-1].hex() == b["merkle_root_display"] and inst.extranonce2 == 2
    # ... and what goes back to the pool rebuilds the mined coinbase
    sub = engine.v1_submit_params(r.shares[0], w)
    assert sub == ["otedama", "170", "02", "%08x" % b["time"], "%08x" % b["nonce"]]
    assert cb[:k] + bytes.fromhex(sub[2]) + cb[k + 1:] == cb


def _docs_example_work(engine, golden):
    ex = golden["reference_vectors"]["real_world_anchors"]["stratum_docs_example"]
    params = list(ex["notify_params"])
    assert params[0] == "bf"        # a hex job id, as most pools send: opaque to the library (applyJob's
    #                                 Sscanf("%d"), run.go:1345-1348, would refuse it; otd_work_from_v1_job still does)
    with pytest.raises(engine.EngineError, match="unparseable job ID"):
        engine.work_from_v1_job(engine.Job(job_id="bf", nbits=0x1D00FFFF), 0, 1.0)
    with pytest.raises(engine.EngineError, match="job id longer"):
        engine.work_from_v1_notify(["x" * 64] + params[1:], ex["subscribe_extranonce1"], 4, 1.0)
    en2 = int.from_bytes(bytes.fromhex(ex["submit_params"][2]), "little")
    w = engine.work_from_v1_notify(params, ex["subscribe_extranonce1"], ex["subscribe_extranonce2_size"], 1.0,
                                   extranonce2_begin=en2 - 1, extranonce2_count=3)
    return ex, w


def test_stratum_documentation_example_as_a_work_unit(eng, golden):
    """The documentation's notify through otd_work_from_v1_notify: the header fields a miner must hash."""
    engine, miner = eng
    ex, w = _docs_example_work(engine, golden)
    wire = bytes.fromhex(ex["notify_params"][1])
    assert w.header.prev_hash == b"".join(wire[4 * i:4 * i + 4][::-1] for i in range(8))
    assert w.header.prev_hash[-4:] == b"\x00" * 4                # header order: the hash's zero bytes come last
    assert (w.header.version, w.header.bits, w.header.time) == (2, 0x1C2AC4AF, 0x504E86B9)
    assert (w.job_id, w.job_id_str, w.extranonce1.hex(), w.extranonce2_size) == (0, "bf", "08000002", 4)
    dec = engine.work_from_v1_notify(["191"] + list(ex["notify_params"][1:]), ex["subscribe_extranonce1"], 4, 1.0)
    assert (dec.job_id, dec.job_id_str) == (191, "191")      # a decimal id also lands in the numeric field
    for odd in ("1a", "4294967296", "-1", ""):
        assert engine.work_from_v1_notify([odd] + list(ex["notify_params"][1:]), "", 4, 1.0).job_id == 0
    assert w.coinb1.hex() == ex["notify_params"][2] and w.coinb2.hex() == ex["notify_params"][3] and not w.merkle_branches
    assert w.target == miner.target_from_difficulty(1.0)


@pytest.mark.gpu
def test_stratum_documentation_example_share_is_found_and_submitted(eng, golden):
    """... and on the device: rolling extranonce2 and scanning around the documented nonce finds exactly the
    share the documentation submits, and the submit fields come back as documented."""
    engine, miner = eng
    ex, w = _docs_example_work(engine, golden)
    sub = ex["submit_params"]
    w.header.time = int(sub[3], 16)                              # the miner's rolled ntime
    nonce = int(sub[4], 16)
    w.nonce_begin, w.nonce_count = nonce - (1 << 20), 1 << 21
    r = miner.scan(w)
    assert r.hashes == 3 << 21
    assert [(s.nonce, s.hash[::-1].hex()) for s in r.shares] == [(nonce, ex["hash_display"])]
    inst = miner.prepare_instances(w, 1, 1)[0]
    assert inst.merkle_root.hex() == ex["merkle_root_header_order"]
    assert r.shares[0].job_id_str == "bf"
    assert engine.v1_submit_params(r.shares[0], w, worker_name=sub[0]) == list(sub[:5])      # job id "bf" verbatim


@pytest.mark.gpu
def test_block_100000_is_found_from_its_notify_form(eng, oracle, golden):
    """Real chain data with a TWO-level merkle branch and a 2-byte extranonce, at the network target of its
    time (difficulty 14484): notify form -> K3 (coinbase hash + two folds) / K2 / K1 -> the mined nonce."""
    engine, miner = eng
    b = golden["reference_vectors"]["real_world_anchors"]["block100000"]
    cb = bytes.fromhex(b["coinbase_tx_hex"])
    tx = [bytes.fromhex(t)[::-1] for t in b["txids_display"]]
    sp = b["stratum_split"]
    k, size, e = sp["coinb1_len"], sp["extranonce2_size"], sp["extranonce2_value"]
    prev = bytes.fromhex(b["prev_hash_display"])[::-1]
    prev_wire = b"".join(prev[4 * i:4 * i + 4][::-1] for i in range(8)).hex()
    branch = [tx[1].hex(), oracle.sha256d_hashlib(tx[2] + tx[3]).hex()]
    params = ["100000", prev_wire, cb[:k].hex(), cb[k + size:].hex(), branch, "%08x" % b["version"], b["bits"],
              "%08x" % b["time"], True]
    w = engine.work_from_v1_notify(params, "", size, 0, extranonce2_begin=e - 2, extranonce2_count=4)
    assert w.target == miner.target_from_nbits(int(b["bits"], 16))          # difficulty 0: the network target
    w.nonce_begin, w.nonce_count = b["nonce"] - (1 << 20), 1 << 21
    r = miner.scan(w)
    assert [(s.extranonce2, s.nonce, s.hash[::-1].hex()) for s in r.shares] == [(e, b["nonce"], b["hash_display"])]
    assert miner.prepare_instances(w, 2, 1)[0].merkle_root[::-1].hex() == b["merkle_root_display"]
    assert engine.v1_submit_params(r.shares[0], w)[2:] == ["0602", "%08x" % b["time"], "%08x" % b["nonce"]]


@pytest.mark.gpu
def test_block_125552_nonce_is_found(eng, golden):
    """The Bitcoin wiki's block-hashing example at its network target (difficulty 244112)."""
    engine, miner = eng
    b = golden["reference_vectors"]["real_world_anchors"]["block125552"]
    h = miner.Header(b["version"], bytes.fromhex(b["prev_hash_display"])[::-1], bytes.fromhex(b["merkle_root_display"])[::-1],
                     b["time"], int(b["bits"], 16), 0)
    r = miner.scan(miner.Work(header=h, nbits=h.bits, target=miner.target_from_nbits(h.bits),
                              nonce_begin=b["nonce"] - (1 << 22), nonce_count=1 << 23))
    assert [(s.nonce, s.hash[::-1].hex()) for s in r.shares] == [(b["nonce"], b["hash_display"])]


@pytest.mark.gpu
def test_block_1_nonce_is_found(eng, golden):
    engine, miner = eng
    rw = golden["reference_vectors"]["real_world_anchors"]
    h = miner.parse_header(bytes.fromhex(rw["block1_header_hex"]))
    n = h.nonce
    h.nonce = 0
    r = miner.scan(miner.Work(header=h, nbits=h.bits, target=miner.target_from_nbits(h.bits), nonce_begin=n - 70000,
                              nonce_count=140000))
    assert [(s.nonce, s.hash[::-1].hex()) for s in r.shares] == [(n, rw["block1_hash_display"])]


def _instances(w):
    """(extranonce2, version index, nonce window) triples a Work covers, as sets of axis values."""
    vers = set(range(w.version_begin, w.version_begin + max(1, w.version_count) * max(1, w.version_stride),
                     max(1, w.version_stride))) if w.version_mask else {None}
    en2 = set(range(w.extranonce2_begin, w.extranonce2_begin + w.extranonce2_count * max(1, w.extranonce2_stride),
                    max(1, w.extranonce2_stride))) if w.coinb1 is not None else {None}
    return vers, en2, (w.nonce_begin, w.nonce_begin + (w.nonce_count or 1 << 32))


def test_partition_work_covers_the_space_exactly_once(eng):
    engine, miner = eng
    from otedama_b200 import workloads as W
    c3, c4 = W.config3_v1(), W.config4_v2()
    hdr = miner.Header(c4["version"], c4["prev_hash"], c4["merkle_root"], c4["time"], c4["bits"], 0)
    t = b"\xff" * 32
    for G in (1, 2, 3, 4, 8):
        # version axis (configs[3]): every G-th index, whole mask
        w = miner.Work(header=hdr, target=t, version_mask=c4["mask"])
        parts = [engine.partition_work(w, g, G) for g in range(G)]
        seen = [v for p in parts for v in _instances(p)[0]]
        assert sorted(seen) == list(range(1 << 16)) and all(p.nonce_count == 0 for p in parts)
        # ... also from an offset, with a stride, and a count not divisible by G
        w = miner.Work(header=hdr, target=t, version_mask=c4["mask"], version_begin=5, version_stride=3, version_count=101)
        seen = [v for g in range(G) for v in _instances(engine.partition_work(w, g, G))[0]]
        assert sorted(seen) == list(range(5, 5 + 3 * 101, 3))
        # extranonce2 axis (configs[2]/[4])
        w = miner.Work(header=hdr, target=t, coinb1=c3["coinb1"], extranonce1=c3["extranonce1"], coinb2=c3["coinb2"],
                       merkle_branches=c3["branches"], extranonce2_size=8, extranonce2_begin=7, extranonce2_count=1000,
                       extranonce2_stride=2)
        seen = [e for g in range(G) for e in _instances(engine.partition_work(w, g, G))[1]]
        assert sorted(seen) == list(range(7, 7 + 2000, 2))
        # plain header (configs[1] within one header): contiguous nonce windows
        w = miner.Work(header=hdr, target=t, nonce_begin=1000, nonce_count=(1 << 32) - 1000)
        wins = [_instances(engine.partition_work(w, g, G))[2] for g in range(G)]
        assert wins[0][0] == 1000 and wins[-1][1] == 1 << 32
        assert all(a[1] == b[0] for a, b in zip(wins, wins[1:]))
    with pytest.raises(engine.EngineError):
        engine.partition_work(miner.Work(header=hdr, target=t, nonce_count=3), 0, 4)
    with pytest.raises(engine.EngineError):
        engine.partition_work(miner.Work(header=hdr, target=t), 2, 2)


@pytest.mark.gpu
def test_partitioned_scans_reproduce_the_whole(eng, oracle):
    engine, miner = eng
    from otedama_b200 import capi, workloads as W
    ndev = capi.lib().otd_device_count()
    c3 = W.config3_v1()
    t = oracle.target_from_difficulty(2.0 ** -20)
    w = miner.Work(header=miner.Header(c3["version"], c3["prev_hash"], b"\x00" * 32, c3["time"], c3["bits"], 0), target=t,
                   coinb1=c3["coinb1"], extranonce1=c3["extranonce1"], coinb2=c3["coinb2"], merkle_branches=c3["branches"],
                   extranonce2_size=8, extranonce2_begin=3, extranonce2_count=21, nonce_count=1 << 14)
    whole = miner.scan(w)
    for G in (2, 3):
        got = []
        for g in range(G):
            got += miner.scan(engine.partition_work(w, g, G), device=g % ndev).shares
        assert sorted((s.extranonce2, s.nonce, s.hash) for s in got) == sorted((s.extranonce2, s.nonce, s.hash) for s in whole.shares)
    assert len(whole.shares) > 50


# -- startMinerWorkers / mergeShares (internal/engine/setup.go:56-78; run_test.go:2443-2468) -----------------------
class _NoSha256dDevice:
    """run_test.go:2443-2453: a GPU that can do general compute but not SHA-256d."""

    def identity(self):
        from otedama_b200 import hal
        return hal.Identity("gpu-0", hal.FAMILY_GPU, "", "")

    def capabilities(self):
        from otedama_b200 import hal
        return hal.Capabilities(sha256d=False, general_compute=True)


def test_start_miner_workers_no_sha256d_devices(eng):
    engine, miner = eng
    with pytest.raises(engine.EngineError, match="SHA256d"):
        engine.start_miner_workers([_NoSha256dDevice()])
    with pytest.raises(engine.EngineError, match="SHA256d"):
        engine.start_miner_workers([])


@pytest.mark.gpu
def test_start_miner_workers_one_per_device_and_merged_shares(eng, oracle):
    from otedama_b200 import hal
    engine, miner = eng
    devices = [_NoSha256dDevice()] + hal.CudaDriver().enumerate()
    logged = []
    workers, shares = engine.start_miner_workers(devices, log=lambda lv, m: logged.append((lv, m)))
    try:
        assert len(workers) == len(devices) - 1 and len(logged) == len(workers)
        assert [w.device_id() for w in workers] == [d.identity().id for d in devices[1:]]
        job = engine.NewMiningJob(channel_id=1, job_id=7, version=0x20000000, merkle_root=bytes(range(32)))
        target = oracle.target_from_difficulty(1 / 65536.0)
        engine.update_work(workers, job, 1, b"\x11" * 32, 0x1D00FFFF, 0x60000000, target)
        got = []
        for _ in range(400):
            got += shares.poll(256, 50)
            if len(got) >= 64:
                break
        assert len(got) >= 64
        ids = {w.device_id() for w in workers}
        for s in got:
            assert s.device_id in ids and s.job_id == 7
            h = oracle.Header(0x20000000, b"\x11" * 32, bytes(range(32)), s.ntime, 0x1D00FFFF, s.nonce)
            assert oracle.hash_header(h) == s.hash and oracle.less_or_equal(s.hash, target)
    finally:
        for w in workers:
            w.stop()


# -- share -> mining.submit fields (stratumv1.go:336-346) ----------------------------------------------------------
def test_v1_submit_params_format(eng, oracle):
    engine, miner = eng
    sh = miner.Share(channel_id=1, job_id=77, nonce=0x0000BEEF, ntime=0x60000001, version=0x2ABCE000, hash=b"\x00" * 32,
                     extranonce2=0x0102030405)
    for size in (4, 5, 8, 12, 16):
        w = miner.Work(coinb1=b"\x01", extranonce2_size=size)
        p = engine.v1_submit_params(sh, w)
        assert p[:2] == ["otedama", "77"] and p[3:] == ["60000001", "0000beef"]
        # the same bytes the coinbase was hashed with (oracle's insertion rule), low `size` bytes of the counter
        assert p[2] == oracle.extranonce2_bytes(0x0102030405 & ((1 << (8 * size)) - 1), size).hex()
    assert engine.v1_submit_params(sh, miner.Work())[2] == ""              # no coinbase in the work: nothing rolled
    rolled = engine.v1_submit_params(sh, miner.Work(coinb1=b"\x01", extranonce2_size=4, version_mask=0x1FFFE000))
    assert rolled[5] == "%08x" % (0x2ABCE000 & 0x1FFFE000) and len(rolled) == 6
    with pytest.raises(engine.EngineError, match="exceeds 16"):
        engine.v1_submit_params(sh, miner.Work(coinb1=b"\x01", extranonce2_size=17))


@pytest.mark.gpu
def test_pool_side_check_of_submitted_fields(eng, oracle):
    """What a pool does with mining.submit: rebuild coinbase, merkle root and header from the job it
    sent plus the five (six) submitted fields, hash, compare with the share target."""
    from otedama_b200 import workloads as W
    engine, miner = eng
    c3 = W.config3_v1()
    mask = 0x1FFFE000
    target = miner.target_from_difficulty(2.0 ** -16)
    w = miner.Work(job_id=31, header=miner.Header(c3["version"], c3["prev_hash"], b"\x00" * 32, c3["time"], c3["bits"], 0),
                   nbits=c3["bits"], target=target, nonce_count=1 << 18, coinb1=c3["coinb1"],
                   extranonce1=c3["extranonce1"], coinb2=c3["coinb2"], merkle_branches=c3["branches"],
                   extranonce2_size=8, extranonce2_begin=0xFFFFFFFE, extranonce2_count=4, version_mask=mask,
                   version_begin=65533, version_count=3)
    r = miner.scan(w)
    assert len(r.shares) > 10
    for s in r.shares:
        _, job, en2_hex, ntime_hex, nonce_hex, vbits_hex = engine.v1_submit_params(s, w)
        assert job == "31"
        root = oracle.v1_merkle_root(c3["coinb1"], c3["extranonce1"], bytes.fromhex(en2_hex), c3["coinb2"], c3["branches"])
        version = (c3["version"] & ~mask) | (int(vbits_hex, 16) & mask)
        h = oracle.Header(version, c3["prev_hash"], root, int(ntime_hex, 16), c3["bits"], int(nonce_hex, 16))
        assert oracle.hash_header(h) == s.hash and oracle.less_or_equal(s.hash, target)


@pytest.mark.gpu
def test_notify_to_submit_round_trip_with_rolled_versions(eng, oracle):
    """The whole V1 path a pool sees, end to end through the worker: mining.notify params with
    a HEX job id -> otd_work_from_v1_notify -> extranonce2 x BIP-320 version x nonce on the device -> share ->
    mining.submit fields -> the pool's own re-hash from nothing but the job it sent and the submitted strings."""
    from otedama_b200 import workloads as W
    engine, miner = eng
    c3 = W.config3_v1()
    mask = 0x1FFFE000
    prev_wire = b"".join(c3["prev_hash"][4 * i:4 * i + 4][::-1] for i in range(8)).hex()
    notify = ["1f9a", prev_wire, c3["coinb1"].hex(), c3["coinb2"].hex(), [b.hex() for b in c3["branches"]],
              "%08x" % c3["version"], "%08x" % c3["bits"], "%08x" % c3["time"], True]
    en1 = c3["extranonce1"].hex()
    w = engine.work_from_v1_notify(notify, en1, 4, 2.0 ** -16, chan_id=5, extranonce2_begin=0xFFFE, extranonce2_count=3)
    w.version_mask, w.version_begin, w.version_count = mask, 40, 65           # 64 + a lone instance per extranonce2
    w.nonce_begin, w.nonce_count = 0xFFFF0000, 1 << 16
    assert (w.job_id, w.job_id_str) == (0, "1f9a")
    with miner.Worker() as wk:
        wk.set_work(w)
        assert wk.wait_exhausted(30000)
        got = []
        while True:
            part = wk.poll_shares(4096, 0)
            if not part:
                break
            got += part
    assert len(got) > 50 and {s.extranonce2 for s in got} == {0xFFFE, 0xFFFF, 0x10000}
    target = oracle.target_from_difficulty(2.0 ** -16)
    seen_versions = set()
    for s in got:
        worker_name, job, en2_hex, ntime_hex, nonce_hex, vbits_hex = engine.v1_submit_params(s, w)
        assert job == "1f9a" and len(en2_hex) == 8
        # pool side: everything below comes from `notify`, the subscribe result and the submitted strings
        coinbase_root = oracle.v1_merkle_root(bytes.fromhex(notify[2]), bytes.fromhex(en1), bytes.fromhex(en2_hex),
                                              bytes.fromhex(notify[3]), [bytes.fromhex(b) for b in notify[4]])
        wire = bytes.fromhex(notify[1])
        prev = b"".join(wire[4 * i:4 * i + 4][::-1] for i in range(8))
        version = (int(notify[5], 16) & ~mask) | (int(vbits_hex, 16) & mask)
        h = oracle.Header(version, prev, coinbase_root, int(ntime_hex, 16), int(notify[6], 16), int(nonce_hex, 16))
        assert oracle.hash_header(h) == s.hash and oracle.less_or_equal(s.hash, target)
        seen_versions.add(version)
    assert len(seen_versions) > 30


# -- device allow-list, per-device configuration, pause by device id ---------------------------------
class _FakeCudaDevice:
    def __init__(self, index, dev_id):
        self.index, self._id = index, dev_id

    def identity(self):
        from otedama_b200 import hal
        return hal.Identity(self._id, hal.FAMILY_GPU, "nvidia", "B200")

    def capabilities(self):
        from otedama_b200 import hal
        return hal.Capabilities(sha256d=True, general_compute=True)


def test_allow_list_and_per_device_config_reach_the_worker_constructor(eng, monkeypatch):
    """setup.go:59-77 gives every SHA256d device a worker with default settings; a GPU box needs to leave devices
    out and to size rings / launches per device.  No GPU needed: the worker constructor is observed."""
    engine, miner = eng
    made = []

    class FakeWorker:
        def __init__(self, cfg):
            self.cfg = cfg
            made.append(cfg)

        def start(self):
            return self

        def stop(self):
            pass

        def device_id(self):
            return f"gpu-cuda{self.cfg.device}"

    monkeypatch.setattr(miner, "Worker", FakeWorker)
    devs = [_FakeCudaDevice(i, f"gpu-cuda{i}") for i in range(4)] + [_NoSha256dDevice()]
    base = miner.WorkerConfig(ring_capacity=1024, launch_log2=30)
    special = miner.WorkerConfig(ring_capacity=8192, queue_capacity=16, launch_log2=28, kernel_variant=1)
    logged = []
    workers, _ = engine.start_miner_workers(devs, log=lambda lv, m: logged.append(m), cfg=base,
                                            allow={"gpu-cuda0", "gpu-cuda2", "gpu-cuda3"},
                                            per_device_cfg={"gpu-cuda2": special})
    assert [w.cfg.device for w in workers] == [0, 2, 3]                     # gpu-cuda1 filtered, no worker
    assert any("gpu-cuda1 not in the device allow-list" in m for m in logged)
    c0, c2, c3 = made
    assert (c0.ring_capacity, c0.launch_log2, c0.kernel_variant) == (1024, 30, -1)
    assert (c2.ring_capacity, c2.queue_capacity, c2.launch_log2, c2.kernel_variant, c2.device) == (8192, 16, 28, 1, 2)
    assert (c3.ring_capacity, c3.launch_log2) == (1024, 30)
    assert base.device == 0 and special.device == 0                        # the caller's configs are not edited
    with pytest.raises(engine.EngineError, match="not detected"):
        engine.start_miner_workers(devs, allow={"gpu-cuda9"})
    with pytest.raises(engine.EngineError, match="not detected"):
        engine.start_miner_workers(devs, per_device_cfg={"gpu-cuda7": special})
    with pytest.raises(engine.EngineError, match="SHA256d"):
        engine.start_miner_workers(devs, allow=set())                       # everything filtered = nothing to mine on


@pytest.mark.gpu
def test_pause_by_device_id_stops_only_that_worker(eng, oracle):
    """applyAllocation's pauseDevice (arbitrate.go:255-262): SetWork(nil) on the named device only.  Two workers
    (two GPUs when the box has them, else both on device 0): the named one stops hashing, the other keeps going."""
    import time
    from otedama_b200 import hal, workloads as W
    engine, miner = eng
    ndev = len(hal.CudaDriver().enumerate())
    a = miner.Worker(miner.WorkerConfig(device=0, launch_log2=24)).start()
    b = miner.Worker(miner.WorkerConfig(device=1 % ndev, launch_log2=24)).start()
    b.device_id = lambda: "gpu-cuda-second"                                # distinct ids also on a 1-GPU box
    try:
        f = W.config2_fields(40)
        work = miner.Work(header=miner.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0),
                          target=bytes(32), repeat=True)
        engine.update_work_partitioned([a, b], work)
        time.sleep(0.5)
        assert a.stats().hashes_total > 0 and b.stats().hashes_total > 0
        assert engine.pause_device([a, b], "gpu-cuda-second") and not engine.pause_device([a, b], "gpu-cuda9")
        time.sleep(0.3)
        hb, ha = b.stats().hashes_total, a.stats().hashes_total
        time.sleep(0.5)
        assert b.stats().hashes_total == hb and not b.has_work()          # paused
        assert a.stats().hashes_total > ha and a.has_work()               # untouched
    finally:
        a.stop(); b.stop()
