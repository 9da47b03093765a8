#!/usr/bin/env python
"""Model-pull micro-benchmark: n peers each pull a d-length "model" from every peer, `num_iter` times.

Counterpart of the reference's applications/benchmarks/rpc_bench.py:96-117 (torch RPC, CPU
tensors). Transports:
  --transport rpc    torch RPC (TensorPipe), one process per peer  — the reference's mechanism
  --transport coll   torch.distributed all_gather (gloo | nccl) under torchrun
  --transport peer   device-to-device pulls from peer-mapped arenas (CUDA IPC + one copy kernel
                     per peer over NVLink) under torchrun — this framework's data plane
Prints the mean pull time and the effective aggregate bandwidth.
"""
import argparse
import os
import pathlib
import sys
import time

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[2]))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--transport", choices=["rpc", "coll", "peer"], default="coll")
    ap.add_argument("--d", type=int, default=1000000)
    ap.add_argument("--num_iter", type=int, default=10)
    ap.add_argument("--n", type=int, default=0, help="number of peers (rpc transport; default WORLD_SIZE)")
    ap.add_argument("--rank", type=int, default=0)
    ap.add_argument("--master", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=27800)
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    args = ap.parse_args(argv)
    world = args.n or int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", str(args.rank)))
    device = torch.device(args.device)
    if device.type == "cuda":
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        device = torch.device("cuda", torch.cuda.current_device())

    def sync():
        if device.type == "cuda":
            torch.cuda.synchronize()

    if args.transport == "rpc":
        from garfield_b200.apps import common
        from garfieldpp.server import Server
        rpc = common.init_rpc("node:%d" % rank, rank, world, args.master, args.port)
        ps = Server(rank, world, 0, world, 0, 0, "none:", "node:", 1, "pimanet", "pima", "sgd", device="cpu", lr=0.1)
        ps.model = torch.rand(args.d)
        ps.get_fake_model = lambda: ps.model
        t0 = time.perf_counter()
        for _ in range(args.num_iter):
            pulled = ps.get_fake_models()
            assert len(pulled) == world
        seconds = time.perf_counter() - t0
        rpc.shutdown()
    else:
        if world > 1 and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", args.master)
            os.environ.setdefault("MASTER_PORT", str(args.port))
            dist.init_process_group("nccl" if device.type == "cuda" else "gloo", rank=rank, world_size=world)
        model = torch.rand(args.d, device=device)
        if args.transport == "coll":
            out = [torch.empty_like(model) for _ in range(world)]
            pull = (lambda: dist.all_gather(out, model)) if world > 1 else (lambda: out[0].copy_(model))
        else:
            from garfield_b200.parallel.peer import PeerArena
            arena = PeerArena(args.d, device)
            arena.local.copy_(model)
            out = [torch.empty_like(model) for _ in range(world)]

            def pull():
                for r in range(world):
                    out[r].copy_(arena.peer(r), non_blocking=True)
        for _ in range(2):
            pull()
        if world > 1:
            dist.barrier()
        sync()
        t0 = time.perf_counter()
        for _ in range(args.num_iter):
            pull()
        sync()
        seconds = time.perf_counter() - t0
        if world > 1:
            dist.barrier()
    per = seconds / args.num_iter
    gbps = 4.0 * args.d * world / per / 1e9
    if rank == 0:
        print("transport=%s peers=%d d=%d: %.3f ms per pull-from-all, %.2f GB/s into each peer" % (args.transport, world, args.d, per * 1e3, gbps))
    return per


if __name__ == "__main__":
    main()
