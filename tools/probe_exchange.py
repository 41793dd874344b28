#!/usr/bin/env python
"""tools/probe_exchange.py -- where does an iteration of the row-sharded solve spend its time outside streaming A?

ncu cannot look at the exchange-fused kernels from inside a multi-rank run: a replayed pass restores the profiled
rank's device memory, its peer window included, and the kernel then waits for flags the peers never send again.
So the kernels carry their own probes (CGB_PROBE=1, cg_kernels.h): thread 0 of every CTA leaves %globaltimer
stamps at entry / after its own column block / after the flag wait / before the epilogue and the SM clocks it
spent polling.  This tool runs a fixed-iteration solve on every rank and prints, per rank, for the LAST iteration:

  product: span of the kernel (first CTA in -> last CTA out), how many CTAs polled at all, the longest poll,
           when the last own-block phase of the first wave ended relative to kernel start
  xr_update / p_update: span, and the time their CTAs waited for the peers' dot partials in the mailbox
  gaps:    end of product -> start of xr_update -> start of p_update (launch + drain between the kernels)

usage: python -m torch.distributed.run --nproc-per-node G ... tools/probe_exchange.py --order 65536 [--iters 50]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--order", dest="n", type=int, default=65536)
    ap.add_argument("--iters", type=int, default=51)
    args = ap.parse_args()
    os.environ["CGB_PROBE"] = "1"
    import numpy as np
    import torch
    import torch.distributed as dist
    import conjugate_gradient_b200 as cg
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group(backend="gloo", rank=rank, world_size=world)
    box = [cg.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    ctx = cg.Context(device=local, rank=rank, ranks=world, uid=box[0])
    n = args.n
    ctx.generate(n, 20170209)
    ctx.solve(ignore_stop=1, max_iter=4)
    ctx.set_x0(np.zeros(n))
    st = ctx.solve(ignore_stop=1, max_iter=args.iters)
    pr = ctx.probe_read().astype(np.int64)
    NP, NV = 16384, 592
    prod, xr, pu = pr[:NP], pr[NP:NP + NV], pr[NP + NV:NP + 2 * NV]
    prod, xr, pu = prod[prod[:, 7] == 1], xr[xr[:, 7] == 1], pu[pu[:, 7] == 1]
    sm_mhz = 1965.0
    t0 = prod[:, 0].min()
    waited = prod[prod[:, 2] > 0]                     # CTAs that had remote tiles (all, when ranks > 1)
    poll_us = waited[:, 4] / sm_mhz
    out = {
        "rank": rank, "n": n, "ranks": world, "iterations": int(st["iterations"]), "loop_us_per_iteration": st["loop_ms"] * 1e3 / st["iterations"],
        "kernels_per_iteration": (st["kernel_launches"] - 4) / st["iterations"], "graph": st["graph_chunk"], "pdl": st["pdl"],
        "product": {
            "ctas": int(len(prod)), "sms_used": int(len(set(prod[:, 5].tolist()))),
            "span_us": float(prod[:, 3].max() - t0) / 1e3,
            "first_cta_in_to_last_cta_in_us": float(prod[:, 0].max() - t0) / 1e3,
            "ctas_that_polled_longer_than_1us": int((poll_us > 1.0).sum()),
            "poll_us": {"max": float(poll_us.max()), "p50": float(np.median(poll_us)), "sum_over_ctas": float(poll_us.sum())},
            "first_wave": {"ctas": int((prod[:, 0] - t0 < 20000).sum()),
                           "own_block_done_us_after_start": float(np.median(prod[prod[:, 0] - t0 < 20000][:, 1] - t0) / 1e3),
                           "flags_seen_us_after_start": float(np.median(prod[prod[:, 0] - t0 < 20000][:, 2] - t0) / 1e3)},
        },
    }
    for name, v in (("xr_update", xr), ("p_update", pu)):
        if len(v):
            out[name] = {"ctas": int(len(v)), "start_after_product_end_us": float(v[:, 0].min() - prod[:, 3].max()) / 1e3,
                         "span_us": float(v[:, 2].max() - v[:, 0].min()) / 1e3 if (v[:, 2] > 0).any() else None,
                         "mailbox_wait_us": {"max": float((v[:, 4] / sm_mhz).max()), "p50": float(np.median(v[:, 4] / sm_mhz))},
                         "gather_done_us_after_start": float(np.median(v[:, 1] - v[:, 0])) / 1e3}
    if len(xr) and len(pu):
        out["p_update"]["start_after_xr_start_us"] = float(pu[:, 0].min() - xr[:, 0].min()) / 1e3
    allr = [None] * world
    dist.all_gather_object(allr, out)
    if rank == 0:
        print(json.dumps({"ranks": allr}), flush=True)
    ctx.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
