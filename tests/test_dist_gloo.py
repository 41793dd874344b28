"""world_size-2 (and 4) CPU tests of the N > 1 host logic over the gloo backend."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import _oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def torchrun(nproc, port, script, *args, timeout=300):
    env = dict(os.environ, OMP_NUM_THREADS="1")
    return subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
                           "--master-addr", "127.0.0.1", "--master-port", str(port), script, *map(str, args)],
                          capture_output=True, text=True, timeout=timeout, env=env)


@pytest.mark.parametrize("world,port", [(2, 29611), (4, 29612)])
def test_spmd_partition_and_collectives_match_oracle(tmp_path, world, port):
    n, seed = 256, 4321
    out = tmp_path / "x.json"
    r = torchrun(world, port, os.path.join(HERE, "_gloo_worker.py"), n, seed, out)
    assert r.returncode == 0, r.stderr[-3000:]
    got = json.load(open(out))
    x = np.array([float.fromhex(v) for v in got["x"]])
    xo, so = O.conjugrad(O.gen_A(n, seed), O.gen_b(n, seed), np.zeros(n), ranks=world)
    assert got["iters"] == so["iterations"]
    if world == 2:
        assert x.tobytes() == xo.tobytes()          # two-term all-reduce: order cannot matter
    else:
        assert np.linalg.norm(x - xo) / np.linalg.norm(xo) < 1e-13


def test_bench_reference_arm_under_torchrun():
    r = torchrun(2, 29613, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", 2, "--steps", 3,
                 "--warmup", 1, "--matrix-n", 1024)
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1                           # rank 0 alone prints; rank 1 exits 0 without work
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["value"] > 0 and d["cpu_baseline"]["kind"] == "port"
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["unit"] == "iterations/s"
