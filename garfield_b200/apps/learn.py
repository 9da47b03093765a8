"""LEARN — decentralized Byzantine-resilient learning: every node is a Server + a Worker.

Loop per node (reference: applications/LEARN/trainer.py:208-287):

    gradients = ps.get_gradients(i, n − f);  aggr = gar(gradients, f)
    if non_iid:  repeat ⌈log2(i+1)⌉ times:  publish aggr; pull peers' aggregated gradients; aggr = gar(...)
    ps.update_model(aggr)
    models = ps.get_models(n − f);  ps.write_model(gar(models, f))          # "contract" step

local transport: the n nodes live in this process (the reference's web demo does the same with n
local processes); rpc transport: one node per process, named `node:<rank>`.
"""

from __future__ import annotations

import math
import sys
import time

import torch

from . import common
from ..utils.checkpoint import save_checkpoint
from ..utils.tracing import Tracer


def build_local(args):
    from garfieldpp.byzWorker import ByzWorker
    from garfieldpp.server import Server
    from garfieldpp.worker import Worker
    n = args.num_nodes
    workers, servers = [], []
    for r in range(n):
        byz = bool(args.attack) and r >= n - args.f
        if byz:
            workers.append(ByzWorker(r, n, n, args.batch, args.model, args.dataset, args.loss, args.attack, max(args.f, 1),
                                     args.train_size, device=args.device))
        else:
            workers.append(Worker(r, n, n, args.batch, args.model, args.dataset, args.loss, args.train_size, device=args.device))
    for r in range(n):
        common.seed_everything()
        ps = Server(r, 1, n, n, args.f, args.f, "node:", "node:", args.batch, args.model, args.dataset, args.optimizer,
                    args.train_size, device=args.device, **args.opt_args)
        ps.attach_workers(workers)
        servers.append(ps)
    for ps in servers:
        ps.attach_servers(servers)
    return servers, workers


def agreement_rounds(i):
    return int(math.ceil(math.log2(i + 1))) if i > 0 else 0


def node_step(args, gar, ps, peers, i, tracer):
    n, f = args.num_nodes, args.f
    with tracer.phase("get_gradients", i):
        gradients = ps.get_gradients(i, n - f)
    with tracer.phase("gar", i):
        aggr = gar(gradients=gradients, f=f)
    return aggr


def run_local(args):
    gar = common.pick_gar(args.gar)
    servers, _ = build_local(args)
    n, f = args.num_nodes, args.f
    honest = servers  # every node runs the server loop; Byzantine nodes only misbehave as workers (as in the reference)
    per_epoch = common.iterations_per_epoch(n, args.batch)
    tracer = Tracer(enabled=bool(args.trace) or args.bench, device=args.device)
    timer = common.StepTimer(args.bench, args.device)
    start = time.time()
    for i in range(args.num_iter):
        def train_step():
            aggrs = [node_step(args, gar, ps, servers, i, tracer) for ps in honest]
            if args.non_iid:
                for _ in range(agreement_rounds(i)):
                    for ps, aggr in zip(honest, aggrs):
                        ps.latest_aggr_grad = aggr
                    with tracer.phase("avg_agree", i):
                        aggrs = [gar(gradients=ps.get_aggr_grads(n - f), f=f) for ps in honest]
            for ps, aggr in zip(honest, aggrs):
                with tracer.phase("update_model", i):
                    ps.update_model(aggr)
            with tracer.phase("contract", i):
                pulled = [ps.get_models(n - f) for ps in honest]
                for ps, models in zip(honest, pulled):
                    ps.write_model(gar(gradients=models, f=f))

        timer.run(i, train_step)
        if i % per_epoch == 0 or (args.acc_freq and i % args.acc_freq == 0 and i != 0):
            ps = honest[0]
            acc = ps.compute_binary_accuracy() if args.loss == "binary-cross-entropy" else ps.compute_accuracy()
            print("Epoch: {} Accuracy: {} Time: {}".format(i / per_epoch, acc, time.time() - start))
            sys.stdout.flush()
    if args.checkpoint:
        save_checkpoint(args.checkpoint, honest[0].flat_model, honest[0].optimizer, step=args.num_iter, config=vars(args))
    if args.trace:
        tracer.dump_chrome_trace(args.trace)
    return honest[0]


def run_rpc(args):
    from garfieldpp.byzWorker import ByzWorker
    from garfieldpp.server import Server
    from garfieldpp.worker import Worker
    n, f, rank = args.num_nodes, args.f, args.rank
    gar = common.pick_gar(args.gar)
    rpc = common.init_rpc("node:{}".format(rank), rank, n, args.master, args.port)
    if args.attack and rank >= n - f:
        ByzWorker(rank, n, n, args.batch, args.model, args.dataset, args.loss, args.attack, max(f, 1), args.train_size, device=args.device)
    else:
        Worker(rank, n, n, args.batch, args.model, args.dataset, args.loss, args.train_size, device=args.device)
    common.seed_everything()
    ps = Server(rank, n, n, n, f, f, "node:", "node:", args.batch, args.model, args.dataset, args.optimizer, args.train_size,
                device=args.device, **args.opt_args)
    tracer = Tracer(enabled=False)
    for i in range(args.num_iter):
        aggr = node_step(args, gar, ps, None, i, tracer)
        if args.non_iid:
            for _ in range(agreement_rounds(i)):
                ps.latest_aggr_grad = aggr
                aggr = gar(gradients=ps.get_aggr_grads(n - f), f=f)
        ps.update_model(aggr)
        ps.write_model(gar(gradients=ps.get_models(n - f), f=f))
        if args.acc_freq and i % args.acc_freq == 0:
            print("Iteration {} Accuracy: {}".format(i, ps.compute_accuracy()))
            sys.stdout.flush()
    rpc.shutdown()


def run_spmd(args):
    """One process per GPU (torchrun): the device-resident `DecentralizedEngine` (node k lives on rank k mod N; gradient
    aggregation, the non-iid agreement rounds and the contract step are fused gather + GAR kernels over peer memory)."""
    import os
    import torch
    import torch.distributed as dist
    from ..parallel.topologies import DecentralizedEngine
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    byz = {args.num_nodes - 1 - k: args.attack for k in range(args.f)} if args.attack else {}
    engine = DecentralizedEngine(args.model, args.dataset, args.loss, args.num_nodes, args.f, args.gar, args.batch, optimizer=args.optimizer,
                                 opt_args=args.opt_args, byzantine=byz, non_iid=args.non_iid)
    start = time.time()
    for i in range(args.num_iter):
        loss = engine.step(i)
        if args.log and i % 10 == 0 and engine.rank == 0:
            print("Iteration {} loss {}".format(i, loss.item()))
    torch.cuda.synchronize()
    if engine.rank == 0:
        elapsed = time.time() - start
        print("{} iterations in {:.3f} s: {:.1f} images/s".format(args.num_iter, elapsed, args.num_iter * args.num_nodes * args.batch / elapsed))
    engine.close()
    if dist.is_initialized():
        dist.destroy_process_group()


def main(argv=None):
    args = common.finish_args(common.base_parser("LEARN (decentralized) on garfield-b200\n" + common.__doc__, decentralized=True).parse_args(argv))
    assert args.num_nodes >= 1 and args.f * 2 < args.num_nodes
    common.banner(args, args.rank)
    common.seed_everything()
    if args.transport == "rpc":
        return run_rpc(args)
    if args.transport == "spmd":
        return run_spmd(args)
    return run_local(args)
