#!/usr/bin/env python
"""Concurrent pinned host->device copies at N ranks (one process per GPU, torchrun): what the end-to-end leg is bound by.
Every rank copies a pinned buffer of image-sized pieces (6.2 MB, as pcp_submit_frame does) and one large piece (the cloud) to its GPU;
all ranks start together; reports per-rank and aggregate GB/s.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/h2d_bench.py"""
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    piece, n_pieces, big = 1920 * 1080 * 3, 200, 320_000_000
    h_small = torch.empty((n_pieces, piece), dtype=torch.uint8).pin_memory()
    h_big = torch.empty((big,), dtype=torch.uint8).pin_memory()
    d_small = torch.empty((16, piece), dtype=torch.uint8, device="cuda")
    d_big = torch.empty((big,), dtype=torch.uint8, device="cuda")
    out = {}
    for name, fn, nbytes in (("images_6MB_pieces", lambda: [d_small[i % 16].copy_(h_small[i], non_blocking=True) for i in range(n_pieces)], piece * n_pieces),
                             ("cloud_320MB", lambda: d_big.copy_(h_big, non_blocking=True), big)):
        fn(); torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / 3
        gbs = nbytes / dt / 1e9
        if world > 1:
            t = torch.tensor([gbs, -gbs, gbs], device="cuda", dtype=torch.float64)
            mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
            out[name] = {"per_rank_GBs_min": -float(mx[1]), "per_rank_GBs_max": float(mx[0]), "aggregate_GBs": float(sm[2])}
        else:
            out[name] = {"per_rank_GBs_min": gbs, "per_rank_GBs_max": gbs, "aggregate_GBs": gbs}
    if rank == 0:
        try:
            aff = len(os.sched_getaffinity(0))
        except Exception:
            aff = None
        print(json.dumps({"ranks": world, "host_threads_visible": aff, **out}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
