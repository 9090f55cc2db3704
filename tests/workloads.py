"""Oracle-typed views of otedama_b200.workloads (test infrastructure)."""
from oracle import oracle as o
from otedama_b200 import workloads as w


def _oh(f) -> "o.Header":
    return o.Header(f["version"], f["prev_hash"], f["merkle_root"], f["time"], f["bits"], 0)


def genesis_header() -> "o.Header":
    return _oh(w.genesis_fields())


def config2_header(i: int) -> "o.Header":
    return _oh(w.config2_fields(i))


config3_v1 = w.config3_v1
config4_v2 = w.config4_v2
