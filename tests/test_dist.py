"""N > 1 host logic on CPU (gloo, world_size 2): the path shards with no data-path collective — each
rank owns its own stream; the only exchange is the timing reduction (MAX over ranks) — and the reference arm runs on
rank 0 alone."""
import json
import os
import socket
import subprocess
import sys

from conftest import ROOT

WORKER = r'''
import json, os, sys
sys.path.insert(0, os.environ["B2_ROOT"])
import torch.distributed as dist
import bench
dist.init_process_group("gloo")
r, w = dist.get_rank(), dist.get_world_size()
vals = bench.max_over_ranks([10.0 + r, 5.0 - r, 7.0])          # slowest rank wins, element-wise
print(json.dumps({"rank": r, "world": w, "max": vals, "seed": bench.stream_seed(r)}))
dist.destroy_process_group()
'''


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_timing_reduction_and_stream_sharding_world2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    port = _free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), B2_ROOT=ROOT)
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE))
    outs = []
    for p in procs:
        o, e = p.communicate(timeout=240)
        assert p.returncode == 0, e.decode()[-800:]
        outs.append(json.loads(o.decode().strip().splitlines()[-1]))
    assert {o["rank"] for o in outs} == {0, 1} and all(o["world"] == 2 for o in outs)
    for o in outs:
        assert o["max"] == [11.0, 5.0, 7.0]                     # MAX over ranks, the same on every rank
    assert outs[0]["seed"] != outs[1]["seed"]                   # every rank its own stream: weak scaling, nothing shared


def test_reference_arm_runs_on_rank0_only():
    """bench.py --impl reference under torchrun: rank != 0 exits 0 without work or output."""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(_free_port()))
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0"], env=env, capture_output=True, timeout=120)
    assert p.returncode == 0 and p.stdout.strip() == b""


def test_reference_arm_line_shape():
    """rank 0: one JSON line with the contract's keys (bounded sample so that it finishes in seconds here)."""
    env = dict(os.environ, RANK="0", WORLD_SIZE="1")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--size-mib", "64"], env=env, capture_output=True, timeout=600)
    assert p.returncode == 0, p.stderr.decode()[-500:]
    line = json.loads(p.stdout.decode().strip().splitlines()[-1])
    assert line["impl"] == "reference"
    if "unavailable" in line:
        return
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "config", "cpu_baseline", "e2e"):
        assert k in line, k
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert line["cpu_baseline"]["kind"] == "reference" and line["cpu_baseline"]["cores"] >= 1 and line["value"] > 0
