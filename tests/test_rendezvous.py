"""CPU tests of the host-side rules around collectives:
the NCCL-id rendezvous file of conjugrad_mpi() ignores a file left by another launch, and the run-time ROWS
(cgb_set_dimension) belongs to the calling thread."""
import ctypes as C
import os
import subprocess
import sys
import threading
import time

import pytest

import conjugate_gradient_b200 as cg
from conjugate_gradient_b200 import _binding as B

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rendezvous(rank, np_, out, idx):
    buf = C.create_string_buffer(128)
    rc = B.lib().cgb_debug_rendezvous_id(rank, np_, buf)
    out[idx] = (rc, buf.raw)


@pytest.fixture()
def clean_env(monkeypatch):
    for k in ("CGB_ID_FILE", "CGB_RUN_ID", "TORCHELASTIC_RUN_ID", "SLURM_JOB_ID", "OMPI_MCA_ess_base_jobid", "PMIX_NAMESPACE", "PMI_JOBID"):
        monkeypatch.delenv(k, raising=False)
    return monkeypatch


def test_stale_id_file_of_another_launch_is_ignored(tmp_path, clean_env):
    path = tmp_path / "nccl.id"
    stale = bytes([0xEE]) * 128 + b"CGB_RUN_ID=old-launch".ljust(64, b"\0")
    path.write_bytes(stale)
    clean_env.setenv("CGB_ID_FILE", str(path))
    clean_env.setenv("CGB_RUN_ID", "new-launch")
    out = [None, None, None]
    readers = [threading.Thread(target=_rendezvous, args=(r, 3, out, r)) for r in (1, 2)]
    for t in readers:
        t.start()
    time.sleep(0.3)                       # the readers have been looking at the stale file for a while
    assert out[1] is None and out[2] is None, "a reader took the id of another launch"
    _rendezvous(0, 3, out, 0)
    for t in readers:
        t.join(timeout=30)
    assert all(o is not None and o[0] == 0 for o in out), out
    assert out[0][1] == out[1][1] == out[2][1] and out[0][1] != stale[:128]
    data = path.read_bytes()
    assert len(data) == 192 and data[:128] == out[0][1] and data[128:].startswith(b"CGB_RUN_ID=new-launch")


def test_without_a_launch_tag_only_a_fresh_file_counts(tmp_path, clean_env):
    path = tmp_path / "nccl.id"
    path.write_bytes(bytes([0xAB]) * 128 + bytes(64))
    old = time.time() - 3600
    os.utime(path, (old, old))            # written an hour ago: some earlier run's
    clean_env.setenv("CGB_ID_FILE", str(path))
    out = [None, None]
    t = threading.Thread(target=_rendezvous, args=(1, 2, out, 1))
    t.start()
    time.sleep(0.3)
    assert out[1] is None
    _rendezvous(0, 2, out, 0)
    t.join(timeout=30)
    assert out[0][0] == 0 and out[1][0] == 0 and out[0][1] == out[1][1] != bytes([0xAB]) * 128


def test_thread_ranks_meet_without_a_file(clean_env):
    out = [None] * 4
    th = [threading.Thread(target=_rendezvous, args=(r, 4, out, r)) for r in range(4)]
    for t in reversed(th):
        t.start()
    for t in th:
        t.join(timeout=30)
    assert all(o[0] == 0 for o in out) and len({o[1] for o in out}) == 1


def test_dimension_is_per_thread():
    """conjugrad() in two threads with different orders: each thread's cgb_set_dimension is its own.  Without a GPU
    the call fails after the argument checks -- what is observable here is which order each thread's call was made
    with (the divisibility message of CGB_GPUS=3 names it: parallel_cg.c:86-90's "%d is not divisible by %d")."""
    code = r'''
import ctypes as C, os, sys, threading
sys.path.insert(0, %r)
os.environ["CGB_GPUS"] = "3"
from conjugate_gradient_b200 import _binding as B
L = B.lib()
seen = {}
bar = threading.Barrier(2)
def run(n):
    L.cgb_set_dimension(n)
    bar.wait()                                   # both orders are set before either call is made
    a = (C.c_double * 1)()
    L.conjugrad(a, a, a)                         # prints the failure on stderr
th = [threading.Thread(target=run, args=(n,)) for n in (1000, 2000)]
[t.start() for t in th]; [t.join() for t in th]
''' % ROOT
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    assert "1000 is not divisible by 3" in r.stderr and "2000 is not divisible by 3" in r.stderr, r.stderr
