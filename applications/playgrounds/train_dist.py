#!/usr/bin/env python
"""Plain (non-robust) synchronous data-parallel SGD with a parameter server — the "vanilla" baseline the robust topologies are
compared against. Counterpart of the reference's rsrcs/playgrounds/legacy_code/train_dist.py:155-218: rank 0 is the server, ranks
1..W-1 are workers; every step the workers' gradients are summed at the server, divided by the number of WORKERS (the server holds
a zero gradient), and the model is stepped.

What is different from the reference, on purpose:
  * one collective per step instead of one `dist.reduce` per parameter: the model lives in a flat arena (`FlatModel`), so the
    gradient of the whole network is ONE contiguous buffer (161 reduces per step for ResNet-50 become 1);
  * the reference leaves the workers stepping on their own local gradient (`dist.reduce` only defines the root's buffer), so its
    replicas drift apart; here the server steps and broadcasts the flat parameter vector, so every rank holds the same model
    (`--no-broadcast` reproduces the reference's behaviour);
  * SGD with momentum runs on the flat parameter / velocity vectors: two element-wise launches for the whole model instead of
    two per parameter.

    python applications/playgrounds/train_dist.py --world 3 --dataset mnist --model convnet --num_iter 50          # spawns 3 local processes
    torchrun --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 applications/playgrounds/train_dist.py --model resnet18 --dataset cifar10
"""
import argparse
import os
import pathlib
import socket
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[2]))

from garfield_b200.garfieldpp import tools  # noqa: E402
from garfield_b200.garfieldpp.datasets import DatasetManager  # noqa: E402
from garfield_b200.parallel.arena import FlatModel  # noqa: E402


def parse(argv):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--world", type=int, default=0, help="spawn this many local processes (0: take RANK / WORLD_SIZE from torchrun)")
    ap.add_argument("--dataset", default="mnist")
    ap.add_argument("--model", default="convnet")
    ap.add_argument("--loss", default="cross-entropy")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--num_iter", type=int, default=100)
    ap.add_argument("--lr", type=float, default=0.01)
    ap.add_argument("--momentum", type=float, default=0.5)
    ap.add_argument("--acc_freq", type=int, default=0, help="evaluate on the server every this many steps (0: at the end only)")
    ap.add_argument("--device", default="auto", choices=["auto", "cpu", "cuda"])
    ap.add_argument("--no-broadcast", action="store_true", help="workers keep stepping on their local gradient (the reference's behaviour)")
    ap.add_argument("--seed", type=int, default=1234)
    return ap.parse_args(argv)


def accuracy(module, test_set, device):
    module.eval()
    correct = total = 0
    with torch.no_grad():
        for x, y in test_set:
            out = module(x.to(device))
            correct += (out.argmax(1).cpu() == y.cpu()).sum().item()
            total += y.numel()
    module.train()
    return correct / max(total, 1)


def run(rank, world, args):
    """One rank of the job; returns (final flat parameters on the CPU, list of the losses this rank saw)."""
    use_cuda = args.device == "cuda" or (args.device == "auto" and torch.cuda.is_available())
    device = torch.device("cuda", rank % torch.cuda.device_count()) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(device)
    torch.manual_seed(args.seed)                                   # same initial model on every rank
    workers = world - 1
    if workers < 1:
        raise SystemExit("train_dist needs a server and at least one worker (world >= 2)")
    module = tools.select_model(args.model, device, args.dataset)
    flat = FlatModel(module, device)
    grad = torch.zeros(flat.numel, device=device)                  # the whole network's gradient: one buffer, one collective
    flat.bind_grads(grad)
    loss_fn = tools.select_loss(args.loss)
    velocity = torch.zeros_like(grad)
    manager = DatasetManager(args.dataset, args.batch * workers, workers, world, rank)     # --batch samples per worker and step
    train = manager.get_train_set() if rank > 0 else None
    test = manager.get_test_set() if rank == 0 else None
    losses = []
    started = time.time()
    for step in range(args.num_iter):
        grad.zero_()
        if rank > 0:
            x, y = train[step % len(train)]
            loss = loss_fn(module(x.to(device)), y.to(device))
            loss.backward()                                        # accumulates into `grad` through the bound views
            losses.append(float(loss.detach()))
        dist.reduce(grad, 0, op=dist.ReduceOp.SUM)                 # the server contributes zeros
        if rank == 0 or args.no_broadcast:
            if rank == 0:
                grad.div_(workers)
            with torch.no_grad():                                  # SGD with momentum on the flat vectors
                velocity.mul_(args.momentum).add_(grad)
                flat.flat.add_(velocity, alpha=-args.lr)
        if not args.no_broadcast:
            dist.broadcast(flat.flat, 0)
        if rank == 0 and args.acc_freq and (step + 1) % args.acc_freq == 0:
            print("step %d  accuracy %.4f  (%.1f s)" % (step + 1, accuracy(module, test, device), time.time() - started), flush=True)
    if rank == 0:
        print("final accuracy %.4f after %d steps with %d workers (%.1f s)" % (accuracy(module, test, device), args.num_iter, workers,
                                                                                 time.time() - started), flush=True)
    dist.barrier()
    return flat.flat.detach().cpu().clone(), losses


def _spawned(rank, world, port, argv):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    main(argv, spawned=True)


def main(argv=None, spawned=False):
    argv = sys.argv[1:] if argv is None else argv
    args = parse(argv)
    if args.world and not spawned:
        import torch.multiprocessing as mp
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            port = s.getsockname()[1]
        mp.spawn(_spawned, args=(args.world, port, argv), nprocs=args.world, join=True)
        return None
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    use_cuda = args.device == "cuda" or (args.device == "auto" and torch.cuda.is_available())
    dist.init_process_group("nccl" if use_cuda else "gloo", rank=rank, world_size=world)
    try:
        return run(rank, world, args)
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
