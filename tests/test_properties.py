"""Property tests (hypothesis) of the host-side helpers behind the C ABI against the oracle's
restatement of internal/miner/sha256d.go:134-257 — same value or same error, for every input class
hypothesis likes to try (subnormals, huge exponents, NaN/inf, boundary nBits).  CPU only."""
import struct

import pytest

hyp = pytest.importorskip("hypothesis")
from hypothesis import given, settings, strategies as st  # noqa: E402

SET = settings(max_examples=400, deadline=None)


def _both(fn_c, fn_o, exc_c, exc_o, x):
    try:
        a = fn_c(x)
    except exc_c:
        a = None
    try:
        b = fn_o(x)
    except exc_o:
        b = None
    return a, b


@SET
@given(st.integers(min_value=0, max_value=0xFFFFFFFF))
def test_target_from_nbits_matches_oracle(product, oracle, nbits):
    from otedama_b200 import miner
    a, b = _both(miner.target_from_nbits, oracle.target_from_nbits, miner.MinerError, oracle.OracleError, nbits)
    assert a == b
    if a is not None:
        # NBitsFromTarget is a left inverse up to the reference's documented normalisation
        assert miner.nbits_from_target(a) == oracle.nbits_from_target(a)
        assert miner.less_or_equal(a, a)


@SET
@given(st.floats(allow_nan=True, allow_infinity=True, allow_subnormal=True, width=64))
def test_target_from_difficulty_matches_oracle(product, oracle, d):
    from otedama_b200 import miner
    a, b = _both(miner.target_from_difficulty, oracle.target_from_difficulty, miner.MinerError, oracle.OracleError, d)
    assert a == b, d


@SET
@given(st.floats(min_value=1e-9, max_value=1e12), st.integers(min_value=0, max_value=0xFFFFFFFF))
def test_v1_job_target_is_share_target_when_it_exists(product, oracle, diff, nbits):
    from otedama_b200 import engine
    try:
        block = oracle.target_from_nbits(nbits)
    except oracle.OracleError:
        with pytest.raises(engine.EngineError):
            engine.v1_job_target(nbits, diff)
        return
    try:
        want = oracle.target_from_difficulty(diff)
    except oracle.OracleError:
        want = block
    assert engine.v1_job_target(nbits, diff) == want


@SET
@given(st.binary(min_size=80, max_size=80))
def test_header_roundtrip_and_layout(product, oracle, raw):
    from otedama_b200 import miner
    h = miner.parse_header(raw)
    assert h.bytes() == raw
    assert h.nonce == struct.unpack_from("<I", raw, 76)[0] and h.version == struct.unpack_from("<I", raw, 0)[0]
    oh = oracle.Header.parse(raw)
    assert (oh.version, oh.time, oh.bits, oh.nonce) == (h.version, h.time, h.bits, h.nonce)


@SET
@given(st.binary(min_size=32, max_size=32), st.binary(min_size=32, max_size=32))
def test_less_or_equal_is_the_little_endian_integer_order(product, a, b):
    from otedama_b200 import miner
    assert miner.less_or_equal(a, b) == (int.from_bytes(a, "little") <= int.from_bytes(b, "little"))


# -- static partition: whatever the axis, the G slices are disjoint and cover the work ----------------
@settings(max_examples=300, deadline=None)
@given(st.integers(1, 16), st.integers(0, 200), st.integers(1, 7), st.integers(1, 300), st.sampled_from(["version", "en2", "nonce"]))
def test_partition_is_a_partition(product, G, begin, stride, count, axis):
    from otedama_b200 import engine, miner
    mask = 0x1FFFE000
    if axis == "version":
        w = miner.Work(version_mask=mask, version_begin=begin, version_stride=stride, version_count=count)
        whole = list(range(begin, begin + stride * count, stride))
    elif axis == "en2":
        w = miner.Work(coinb1=b"\x01", extranonce2_size=8, extranonce2_begin=begin, extranonce2_stride=stride,
                       extranonce2_count=count)
        whole = list(range(begin, begin + stride * count, stride))
    else:
        w = miner.Work(nonce_begin=begin, nonce_count=count)
        whole = list(range(begin, begin + count))
    if count < G:
        with pytest.raises(engine.EngineError, match="fewer"):
            engine.partition_work(w, 0, G)
        return
    seen = []
    for g in range(G):
        p = engine.partition_work(w, g, G)
        if axis == "version":
            seen += range(p.version_begin, p.version_begin + p.version_stride * p.version_count, p.version_stride)
            assert (p.nonce_begin, p.nonce_count) == (0, 0)
        elif axis == "en2":
            seen += range(p.extranonce2_begin, p.extranonce2_begin + p.extranonce2_stride * p.extranonce2_count,
                          p.extranonce2_stride)
        else:
            seen += range(p.nonce_begin, p.nonce_begin + p.nonce_count)
    assert sorted(seen) == whole
    with pytest.raises(engine.EngineError):
        engine.partition_work(w, G, G)


# -- the launch plan: one pass over a work = kernel launches that tile its (instance x nonce) space exactly once ------
@settings(max_examples=250, deadline=None)
@given(st.integers(0, 150), st.integers(0, 12), st.integers(1, 5), st.integers(0, 5000), st.integers(1, 1 << 22),
       st.sampled_from([16, 18, 20, 32]), st.booleans(), st.integers(0, 9), st.integers(1, 3))
def test_launch_plan_tiles_the_search_space(product, n_ver, n_en2, ver_stride, nonce_begin, nonce_count, launch_log2,
                                            per_version, en2_begin, en2_stride):
    """otd_plan_launches (the arithmetic Engine::run launches by): every (instance, nonce) of the work in exactly one
    launch; a version-shared launch holds 2..OTD_VS_MAX_INSTANCES version instances of ONE extranonce2 value and of
    one 256-instance preparation batch, with consecutive version indices; lone instances take the plain kernel;
    OTD_WORK_PER_VERSION_LAUNCH switches sharing off."""
    from otedama_b200 import capi, miner
    kw = dict(target=b"\xff" * 32, nonce_begin=nonce_begin, nonce_count=nonce_count, per_version_launch=per_version)
    if n_ver:
        kw.update(version_mask=0x1FFFE000, version_begin=3, version_stride=ver_stride, version_count=n_ver)
    if n_en2:
        kw.update(coinb1=b"\x01\x02", extranonce2_size=4, extranonce2_begin=en2_begin, extranonce2_stride=en2_stride,
                  extranonce2_count=n_en2)
    vers, en2s = max(1, n_ver), max(1, n_en2)
    plan, total = miner.plan_launches(miner.Work(**kw), launch_log2)
    assert total == len(plan)
    covered = {}
    for first, n, shared, nb, nc, en2, vidx in plan:
        assert 1 <= n <= capi.OTD_VS_MAX_INSTANCES and shared == (n >= 2) and nc >= 1
        if per_version or not n_ver:
            assert n == 1
        e_idx, slot = divmod(first, vers)
        assert slot + n <= vers                                   # one extranonce2 value per launch
        assert first // 256 == (first + n - 1) // 256             # one preparation batch per launch
        assert en2 == (en2_begin + e_idx * en2_stride if n_en2 else 0)
        assert vidx == (3 + slot * ver_stride if n_ver else 0)
        assert nonce_begin <= nb and nb + nc <= nonce_begin + nonce_count
        for inst in range(first, first + n):
            covered.setdefault(inst, []).append((nb, nb + nc))
    assert sorted(covered) == list(range(vers * en2s))
    for inst, spans in covered.items():                           # the nonce windows of an instance tile its range
        spans.sort()
        assert spans[0][0] == nonce_begin and spans[-1][1] == nonce_begin + nonce_count
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
    # a launch never exceeds the configured candidate count by more than one block of the version-shared kernel
    assert all(n * nc <= max(1 << launch_log2, n * 1024) for _, n, _, _, nc, _, _ in plan)


def test_launch_plan_of_an_unbounded_work_is_cut_short_not_walked(product):
    """A V1 work with every extranonce2 value (2^40 instances) would take forever to count: the total is refused,
    the first entries are still available."""
    import ctypes
    from otedama_b200 import capi, miner
    w = miner.Work(target=b"\xff" * 32, coinb1=b"\x01", extranonce2_size=8, extranonce2_count=1 << 40)
    with pytest.raises(capi.OtdError, match="refused"):
        miner.plan_launches(w)
    cw, keep = w._c()
    out = (capi.CLaunchPlan * 4)()
    assert capi.lib().otd_plan_launches(ctypes.byref(cw), 0, out, 4, None) == 4
    assert [out[i].extranonce2 for i in range(4)] == [0, 1, 2, 3] and out[0].nonce_count == 1 << 32
