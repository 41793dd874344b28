"""Worker for tests/test_dist_gloo.py (launched by torch.distributed.run, backend gloo, CPU only).
Runs the N > 1 host logic of the solver -- row partition, in-place all-gather of p, all-reduce of
the two dot scalars, identical stop decision on every rank, final all-gather of x -- with the
oracle's per-rank primitives standing in for the CUDA kernels."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import _oracle as O  # noqa: E402


def main():
    n, seed, out = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
    dist.init_process_group(backend="gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    # rendezvous plumbing used by bench.py: rank 0 makes a 128-byte id, everyone receives it
    box = [bytes(range(128)) if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    assert box[0] == bytes(range(128))
    assert n % world == 0                                    # parallel_cg.c:86-90
    m, o = n // world, rank * (n // world)
    A = O.gen_A(n, seed, 0, o, m); b = O.gen_b(n, seed, 0, o, m)
    x_full = np.zeros(n); p_full = np.zeros(n)

    def allreduce(v):
        t = torch.tensor([v], dtype=torch.float64)
        dist.all_reduce(t)
        return float(t.item())

    def allgather_inplace(full):
        parts = [torch.empty(m, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(parts, torch.from_numpy(full[o:o + m].copy()))
        for g, t in enumerate(parts):
            full[g * m:(g + 1) * m] = t.numpy()

    Ap = O.matvec(A, x_full)
    r = b - Ap
    p_full[o:o + m] = r
    rho = allreduce(O.vecvec(r, r))
    iters, eps = 0, 1.0e-6
    for k in range(n):
        allgather_inplace(p_full)
        Ap = O.matvec(A, p_full)
        alpha = rho / allreduce(O.vecvec(p_full[o:o + m].copy(), Ap))
        x_full[o:o + m] = x_full[o:o + m] + p_full[o:o + m] * alpha
        r = r - Ap * alpha
        beta = allreduce(O.vecvec(r, r))
        iters = k + 1
        if np.sqrt(beta) < eps:
            break
        p_full[o:o + m] = r + p_full[o:o + m] * (beta / rho)
        rho = beta
    allgather_inplace(x_full)
    # every rank must have taken the same decision and hold the same x
    t = torch.tensor([float(iters)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    assert int(t.item()) == iters
    if rank == 0:
        with open(out, "w") as f:
            json.dump({"iters": iters, "x": [float(v).hex() for v in x_full], "world": world}, f)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
