"""N>1 bookkeeping of bench.py on CPU: two gloo ranks, each owning an independent region shard (no data-path collective,
SURVEY.md §8(e)); the whole-job rate is all ranks' columns over the slowest rank's time, and the reference arm runs on
rank 0 only."""
import json
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import os, sys, json
sys.path.insert(0, %r)
import bench
dist, rank, local = bench.dist_setup(2, "gloo")
from somatic_sniper_b200 import synth
cols = synth.make_columns(2000 + 500 * rank, seed=bench.shard_seed(rank))
bench.dist_barrier(dist)
ms = 10.0 * (rank + 1)                       # rank 1 is the slow one
value, ms_max, total = bench.whole_job_rate(dist, cols["ncol"], 4, ms)
first = int(cols["bs"][:64].astype("int64").sum())
other = bench.dist_sum(dist, float(first))   # both shards took part, and they differ
print(json.dumps(dict(rank=rank, value=value, ms_max=ms_max, total=total, first=first, both=other)), flush=True)
dist.destroy_process_group()
""" % ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_gloo_ranks_shard_and_reduce(built, tmp_path):
    port = _free_port()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, "-c", WORKER], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        o, e = p.communicate(timeout=300)
        assert p.returncode == 0, e
        outs.append(json.loads(o.strip().splitlines()[-1]))
    a, b = sorted(outs, key=lambda d: d["rank"])
    assert a["total"] == b["total"] == 2000 + 2500
    assert a["ms_max"] == b["ms_max"] == 20.0
    assert abs(a["value"] - 4500 * 4 / 0.020) < 1e-6 and a["value"] == b["value"]
    assert a["first"] != b["first"] and a["both"] == a["first"] + b["first"]


def test_reference_arm_runs_on_rank0_only(built):
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                        "--columns", "2000"], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    assert p.returncode == 0 and p.stdout.strip() == ""
    env["RANK"] = "0"
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                        "--columns", "2000"], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    line = json.loads(p.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["value"] > 0 and line["e2e"]["h2d_bytes_per_step"] == 0
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["gpu_launches"] == 0
