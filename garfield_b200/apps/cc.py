"""Garfield_CC — the collective-communication implementation: every rank holds a model replica,
gradients/models move with `torch.distributed` collectives (gloo on CPU, NCCL on GPU).

Counterpart of the reference's applications/Garfield_CC/trainer.py:55-208,252-324 with the same
modes — `--aggregator` (gradient GAR at the servers) and `--mar` (`vanilla` | `crash` | a GAR name
applied to the servers' aggregated gradients by every rank) — re-designed around FLAT vectors:
one all_gather of the d-length gradient per step instead of 1–4 blocking collectives per
parameter tensor (C9–C14 in SURVEY.md §2.4), which is also what makes the NCCL backend usable
for every mode (the reference falls back to gloo whenever it needs `gather`).

Rank layout (as in the reference): ranks < num_ps are servers, the rest are workers. Every rank
applies the same optimizer step locally, so replicas stay in sync by construction.

  vanilla      all_reduce(SUM)/num_workers of the workers' gradients
  aggregathor  (num_ps = 1) server aggregates the workers' gradients with --aggregator, broadcast
  byzsgd       (num_ps > 1) every server aggregates, then every rank aggregates the servers'
               results with --mar (Byzantine servers lie about theirs when --real_byzantine)
"""

from __future__ import annotations

import os
import sys
import time

import torch
import torch.distributed as dist

from . import common
from ..ops import attacks
from ..parallel.arena import FlatModel
from ..utils.tracing import Tracer


def parser():
    ap = common.base_parser("Garfield_CC on garfield-b200\n" + __doc__)
    ap.add_argument("--backend", default="", help="gloo | nccl (default: nccl on CUDA, gloo otherwise)")
    ap.add_argument("--aggregator", default="", help="gradient GAR (alias of --gar)")
    ap.add_argument("--mar", default="vanilla", help="vanilla | crash | GAR name for model/aggregate agreement")
    ap.add_argument("--epochs", type=int, default=0)
    ap.add_argument("--real_byzantine", type=common.str2bool, default=False, help="Byzantine ranks really attack")
    ap.add_argument("--world_size", type=int, default=0, help="override (default: env WORLD_SIZE)")
    return ap


def run(args):
    world = args.world_size or int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", str(args.rank)))
    device = torch.device(args.device)
    if device.type == "cuda":
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        device = torch.device("cuda", local)
    backend = args.backend or ("nccl" if device.type == "cuda" else "gloo")
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", args.master)
        os.environ.setdefault("MASTER_PORT", str(args.port))
        dist.init_process_group(backend, rank=rank, world_size=world)
    num_ps, n = args.num_ps, args.num_workers
    assert world == num_ps + n, "world size must be num_ps + num_workers"
    is_ps = rank < num_ps
    gar_name = args.aggregator or args.gar
    grad_gar = common.pick_gar(gar_name) if gar_name not in ("vanilla", "") else None
    model_gar = None if args.mar in ("vanilla", "crash") else common.pick_gar(args.mar)

    from garfieldpp import tools
    from garfieldpp.datasets import DatasetManager
    common.seed_everything()
    flat_model = FlatModel(tools.select_model(args.model, device, args.dataset), device)
    dist.broadcast(flat_model.flat, src=0)
    optimizer = tools.select_optimizer(flat_model, args.optimizer, **args.opt_args)
    loss_fn = tools.select_loss(args.loss)
    d = flat_model.numel
    batches = DatasetManager(args.dataset, args.batch, n, world, rank if not is_ps else num_ps, args.train_size).get_train_set()
    grad = torch.zeros(d, device=device)
    flat_model.bind_grads(grad)
    gathered = [torch.zeros(d, device=device) for _ in range(world)]
    byz_worker = args.real_byzantine and not is_ps and (rank - num_ps) >= n - args.fw and args.attack
    byz_server = args.real_byzantine and is_ps and rank >= num_ps - args.fps and args.server_attack
    tracer = Tracer(enabled=args.bench, device=device)
    num_iter = args.epochs * len(batches) if args.epochs else args.num_iter
    start = time.time()
    for i in range(num_iter):
        with tracer.phase("compute", i):
            grad.zero_()
            if not is_ps:
                data, target = batches[i % len(batches)]
                flat_model.module.train()
                loss = loss_fn(flat_model.module(data.to(device)), target.to(device))
                loss.backward()
                if byz_worker:
                    grad.copy_({"random": attacks.random_like, "reverse": attacks.reverse, "drop": attacks.drop}[args.attack](grad))
        with tracer.phase("reduce", i):
            if grad_gar is None:
                dist.all_reduce(grad)
                update = grad / n
            else:
                dist.all_gather(gathered, grad)
                worker_grads = gathered[num_ps:]
                mine = grad_gar(gradients=[g.clone() for g in worker_grads], f=args.fw) if is_ps else torch.zeros(d, device=device)
                if byz_server:
                    mine = {"random": attacks.random_like, "reverse": attacks.reverse, "drop": attacks.drop}[args.server_attack](mine)
                if num_ps == 1:
                    dist.broadcast(mine, src=0)
                    update = mine
                else:
                    dist.all_gather(gathered, mine)
                    from_servers = gathered[:num_ps]
                    if model_gar is None:
                        update = from_servers[0] if args.mar == "crash" else torch.stack(from_servers).mean(0)
                    else:
                        update = model_gar(gradients=[g.clone() for g in from_servers], f=args.fps)
        with tracer.phase("update", i):
            optimizer.step(update)
        if args.log and rank == num_ps and i % 10 == 0:
            print("Iteration {} loss {}".format(i, loss.item()))
            sys.stdout.flush()
    dist.barrier()
    digest = flat_model.flat.double().sum().item()
    if rank == 0:
        elapsed = time.time() - start
        print("{} iterations in {:.3f} s ({:.1f} images/s); model digest {:.6f}".format(num_iter, elapsed, num_iter * n * args.batch / elapsed, digest))
        if args.bench:
            for name, s in tracer.summary().items():
                print("phase %-8s mean %.3f ms (%d calls)" % (name, s["mean_ms"], s["count"]))
    return flat_model, digest


def main(argv=None):
    args = common.finish_args(parser().parse_args(argv))
    if int(os.environ.get("RANK", str(args.rank))) == 0:
        common.banner(args, 0)
    return run(args)
