"""The N>1 path on CPU: world_size-2 gloo processes exercising the static,
zero-communication partition and the control-plane reductions
bench.py uses (barrier, max of times, sum of candidates), plus the reference arm's
"rank 0 only" rule.  No GPU, no data-path collective."""
import json
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

HELPER = r'''
import json, os, sys
sys.path.insert(0, os.environ["OTD_ROOT"])
import torch, torch.distributed as dist
from otedama_b200 import workloads as W
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group(backend="gloo", rank=rank, world_size=world)
units = W.partition(W.CONFIG2_HEADERS, rank, world)          # header i -> rank i mod world
vers = W.partition(1 << 16, rank, world)                    # version index k -> rank k mod world
t = torch.tensor([float(len(units))], dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.SUM)
mx = torch.tensor([1.0 + rank], dtype=torch.float64)
dist.all_reduce(mx, op=dist.ReduceOp.MAX)
gathered = [None] * world
dist.all_gather_object(gathered, units)
dist.barrier()
if rank == 0:
    print(json.dumps({"sum_units": t.item(), "max": mx.item(), "units": gathered, "n_vers": len(vers),
                      "vers_head": vers[:3]}))
dist.destroy_process_group()
'''


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _spawn(world, argv, extra_env=None):
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OTD_ROOT=ROOT)
        env.update(extra_env or {})
        procs.append(subprocess.Popen(argv, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, cwd=ROOT))
    outs = [p.communicate(timeout=240) for p in procs]
    for p, (o, e) in zip(procs, outs):
        assert p.returncode == 0, e[-2000:]
    return outs


def test_static_partition_and_control_plane_two_ranks():
    outs = _spawn(2, [sys.executable, "-c", HELPER])
    doc = json.loads(outs[0][0].strip().splitlines()[-1])
    assert outs[1][0].strip() == ""
    assert doc["sum_units"] == 257 and doc["max"] == 2.0
    u0, u1 = doc["units"]
    assert sorted(u0 + u1) == list(range(257)) and not set(u0) & set(u1)
    assert u0[:3] == [0, 2, 4] and u1[:3] == [1, 3, 5]
    assert doc["n_vers"] == 1 << 15 and doc["vers_head"] == [0, 2, 4]


def test_reference_arm_runs_on_rank0_only():
    outs = _spawn(2, [sys.executable, "bench.py", "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                      "--ref-seconds", "1.0"])
    line = json.loads(outs[0][0].strip().splitlines()[-1])
    assert outs[1][0].strip() == ""                      # the other rank exits 0 without work
    assert line["impl"] == "reference" and line["unit"] == "GH/s" and line["n_gpus"] == 2
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": "GH/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert 0.0005 < line["value"] < 1.0                  # MH/s-class CPU numbers expressed in GH/s


def test_partition_covers_every_unit_once_for_1_2_4_8():
    from otedama_b200 import workloads as W
    for world in (1, 2, 4, 8):
        parts = [W.partition(257, r, world) for r in range(world)]
        assert sorted(sum(parts, [])) == list(range(257))
        assert max(map(len, parts)) - min(map(len, parts)) <= 1
