"""The roofline denominators bench.py uses re-derived from the dataflow graph of one
candidate (scripts/count_ops.py): 2450 algorithmic 3-input ops per hash, of which 1804 are shift/logic."""
import importlib.util
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_algorithmic_op_count_matches_the_contract_figures():
    c = _load(os.path.join(ROOT, "scripts", "count_ops.py"), "count_ops")
    assert c.total(c.scan_graph(fold=False, early=False)) == [1344, 704, 720]
    shf, lop3, add = c.total(c.scan_graph(fold=True, early=True))
    assert (shf, lop3) == (1176, 628)
    assert abs(shf + lop3 + add - 2450) <= 1      # the survey counts one more add than the graph needs
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert int(re.search(r"^ALG_OPS_PER_HASH = (\d+)", src, re.M).group(1)) == 2450
    assert int(re.search(r"alu_only_ops = (\d+)", src).group(1)) == shf + lop3
