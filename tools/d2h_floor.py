#!/usr/bin/env python3
"""Platform floor of the host-buffer pass: N ranks (one per GPU, torchrun) each copy `--mb` MB device -> pinned host
per step, all at the same time, as the e2e leg of bench.py does with its rows.  Prints one JSON line: per-rank and
aggregate GB/s and the ms per step that alone would cost (is 73 ms at N = 8 1.1x or 3x the floor?)."""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument("--mb", type=float, default=585.0)
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--pieces", type=int, default=6, help="copies per step (the pipeline downloads sub-batch by sub-batch)")
a = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(a.mb * 1e6) // a.pieces
dev = [torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(a.pieces)]
host = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(a.pieces)]


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


for _ in range(3):
    for d, h in zip(dev, host):
        h.copy_(d, non_blocking=True)
barrier()
t0 = time.perf_counter()
for _ in range(a.steps):
    for d, h in zip(dev, host):
        h.copy_(d, non_blocking=True)
    torch.cuda.synchronize()
barrier()
ms = 1e3 * (time.perf_counter() - t0) / a.steps
t = torch.tensor([ms], dtype=torch.float64, device="cuda")
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    ms = float(t[0])
    print(json.dumps({"n_gpus": world, "mb_per_rank_per_step": a.mb, "ms_per_step_max_over_ranks": ms,
                      "gbs_per_rank": a.mb / ms, "gbs_aggregate": world * a.mb / ms}))
if world > 1:
    dist.destroy_process_group()
