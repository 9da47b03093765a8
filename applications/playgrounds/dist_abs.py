#!/usr/bin/env python
"""Hello-world of the bootstrap layer: two (or more) local processes all-reduce a tensor over gloo
(or NCCL with --device cuda). Counterpart of the reference's rsrcs/playgrounds/dist_abs.py:37-46.

    python applications/playgrounds/dist_abs.py --world 2
"""
import argparse
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def node(rank, world, port, device):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("nccl" if device == "cuda" else "gloo", rank=rank, world_size=world)
    if device == "cuda":
        torch.cuda.set_device(rank % torch.cuda.device_count())
    t = torch.ones(4, device=device) * (rank + 1)
    dist.all_reduce(t)
    print("rank %d: %s" % (rank, t.tolist()), flush=True)
    assert t[0].item() == world * (world + 1) / 2
    dist.destroy_process_group()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=2)
    ap.add_argument("--device", default="cpu")
    args = ap.parse_args()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(node, args=(args.world, port, args.device), nprocs=args.world, join=True)
