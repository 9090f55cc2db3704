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
