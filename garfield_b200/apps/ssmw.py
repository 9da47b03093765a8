"""AggregaThor (SSMW) and ByzSGD (MSMW) training loops over the role objects.

Loops follow the reference applications (Aggregathor/trainer.py:217-271, ByzSGD/trainer.py:216-278):

    SSMW:  gradients = ps.get_gradients(i, n − fw);  aggr = gar(gradients, f=fw);  ps.update_model(aggr)
    MSMW:  (SSMW step on every server replica) then
           models = ps.get_models(num_ps − fps);  ps.write_model(gar(models, f=fps))

with the LR decay ×0.2 every 30 epochs and an accuracy report every epoch. Unlike the
reference ByzSGD app (which passes num_ps=1) the MSMW gather really queries
all `num_ps` replicas.
"""

from __future__ import annotations

import sys
import time

import torch

from . import common
from ..utils.checkpoint import load_checkpoint, save_checkpoint
from ..utils.tracing import Metrics, Tracer


def build_local(args, msmw):
    """All roles in this process: `num_ps` servers (first `fps` Byzantine when an attack is
    named) and `num_workers` workers (last `fw` Byzantine when an attack is named)."""
    from garfieldpp.byzServer import ByzServer
    from garfieldpp.byzWorker import ByzWorker
    from garfieldpp.server import Server
    from garfieldpp.worker import Worker
    n, world = args.num_workers, args.num_workers + args.num_ps
    workers = []
    for i in range(n):
        rank = args.num_ps + i
        byz = bool(args.attack) and i >= n - args.fw
        if byz:
            workers.append(ByzWorker(rank, world, n, args.batch, args.model, args.dataset, args.loss, args.attack, max(args.fw, 1),
                                     args.train_size, device=args.device))
        else:
            workers.append(Worker(rank, world, n, args.batch, args.model, args.dataset, args.loss, args.train_size, device=args.device))
    servers = []
    for r in range(args.num_ps if msmw else 1):
        common.seed_everything()  # replicas start from the same model
        byz = bool(args.server_attack) and r >= args.num_ps - args.fps and msmw
        if byz:
            ps = ByzServer(r, 1, n, args.num_ps, args.fw, args.fps, "worker:", "ps:", args.batch, args.model, args.dataset,
                           args.optimizer, args.server_attack, device=args.device, **args.opt_args)
        else:
            ps = Server(r, 1, n, args.num_ps, args.fw, args.fps, "worker:", "ps:", args.batch, args.model, args.dataset,
                        args.optimizer, args.train_size, device=args.device, **args.opt_args)
        ps.attach_workers(workers)
        servers.append(ps)
    for ps in servers:
        ps.attach_servers(servers)
    return servers, workers


def server_loop(args, servers, msmw, rank=0):
    """Run the PS-side loop on the honest replicas hosted by this process."""
    from garfieldpp.tools import adjust_learning_rate
    gar = common.pick_gar(args.gar)
    n, fw, fps = args.num_workers, args.fw, args.fps
    lr = args.opt_args["lr"]
    per_epoch = common.iterations_per_epoch(n, args.batch)
    print("One EPOCH consists of {} iterations".format(per_epoch))
    timer = common.StepTimer(args.bench, args.device)
    tracer = Tracer(enabled=bool(args.trace) or args.bench, device=args.device)
    metrics = Metrics()
    first = 0
    if args.resume:
        payload = load_checkpoint(args.resume, servers[0].flat_model, servers[0].optimizer)
        first = payload["step"]
        for ps in servers[1:]:
            ps.write_model(servers[0].flat_model.flat)
        print("Resumed from {} at step {}".format(args.resume, first))
    start = time.time()
    honest = [ps for ps in servers if not hasattr(ps, "attack")]
    for i in range(first, args.num_iter):
        if i % (per_epoch * 30) == 0 and i != 0:
            lr *= 0.2
            for ps in servers:
                adjust_learning_rate(ps.optimizer, lr)

        def train_step():
            for ps in servers:
                with tracer.phase("get_gradients", i):
                    gradients = ps.get_gradients(i, n - fw)
                with tracer.phase("gar", i):
                    aggr = gar(gradients=gradients, f=fw)
                with tracer.phase("update_model", i):
                    ps.update_model(aggr)
            if msmw:
                for ps in honest:
                    with tracer.phase("get_models", i):
                        models = ps.get_models(args.num_ps - fps)
                    with tracer.phase("gar_models", i):
                        agreed = gar(gradients=models, f=fps)
                    with tracer.phase("write_model", i):
                        ps.write_model(agreed)

        timer.run(i, train_step)
        if args.log and honest[0].last_losses:
            print("Iteration {} loss {}".format(i, sum(honest[0].last_losses) / len(honest[0].last_losses)))
        if i % per_epoch == 0 or (args.acc_freq and i % args.acc_freq == 0 and i != 0):
            acc = honest[0].compute_accuracy() if args.loss != "binary-cross-entropy" else honest[0].compute_binary_accuracy()
            metrics.log(step=i, accuracy=acc, elapsed=time.time() - start)
            print("Epoch: {} Accuracy: {} Time: {}".format(i / per_epoch, acc, time.time() - start))
            sys.stdout.flush()
        if args.checkpoint and args.ckpt_freq and (i + 1) % args.ckpt_freq == 0:
            save_checkpoint(args.checkpoint, honest[0].flat_model, honest[0].optimizer, step=i + 1, config=vars(args))
    if args.checkpoint:
        save_checkpoint(args.checkpoint, honest[0].flat_model, honest[0].optimizer, step=args.num_iter, config=vars(args))
    if args.trace:
        tracer.dump_chrome_trace(args.trace)
    if args.bench or args.trace:
        for name, s in tracer.summary().items():
            print("phase %-14s mean %.3f ms  max %.3f ms  (%d calls)" % (name, s["mean_ms"], s["max_ms"], s["count"]))
    return honest[0]


def run_rpc(args, msmw):
    """One role per process over torch RPC (the reference's deployment model)."""
    from garfieldpp.byzWorker import ByzWorker
    from garfieldpp.server import Server
    from garfieldpp.worker import Worker
    world = args.num_workers + args.num_ps
    rank = args.rank
    if rank < args.num_ps:
        rpc = common.init_rpc("ps:{}".format(rank), rank, world, args.master, args.port)
        common.seed_everything()
        ps = Server(rank, world, args.num_workers, args.num_ps if msmw else 1, args.fw, args.fps, "worker:", "ps:", args.batch,
                    args.model, args.dataset, args.optimizer, args.train_size, device=args.device, **args.opt_args)
        server_loop(args, [ps], msmw and args.num_ps > 1, rank)
    else:
        rpc = common.init_rpc("worker:{}".format(rank - args.num_ps), rank, world, args.master, args.port)
        index = rank - args.num_ps
        if args.attack and index >= args.num_workers - args.fw:
            ByzWorker(rank, world, args.num_workers, args.batch, args.model, args.dataset, args.loss, args.attack, max(args.fw, 1),
                      args.train_size, device=args.device)
        else:
            Worker(rank, world, args.num_workers, args.batch, args.model, args.dataset, args.loss, args.train_size, device=args.device)
    rpc.shutdown()


def run_spmd(args, msmw=False):
    """One process per GPU (torchrun): the device-resident engines — `SSMWEngine` (AggregaThor), or `MSMWEngine` (ByzSGD:
    `num_ps` replicated servers, the same rule aggregates gradients and — between servers — models, as in the reference's
    ByzSGD/trainer.py:237-244)."""
    import os
    import torch.distributed as dist
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    byz = {args.num_workers - 1 - k: args.attack for k in range(args.fw)} if args.attack else {}
    if msmw and args.num_ps > 1:
        from ..parallel.topologies import MSMWEngine
        byz_ps = {args.num_ps - 1 - k: args.server_attack for k in range(args.fps)} if args.server_attack else {}
        engine = MSMWEngine(args.model, args.dataset, args.loss, args.num_workers, args.num_ps, args.fw, args.fps, args.gar, args.batch,
                            mar=args.gar, optimizer=args.optimizer, opt_args=args.opt_args, byzantine_workers=byz, byzantine_servers=byz_ps)
    else:
        from ..parallel.engine import SSMWEngine
        engine = SSMWEngine(args.model, args.dataset, args.loss, args.num_workers, args.fw, args.gar, args.batch, args.optimizer,
                            args.opt_args, byzantine=byz)
    start = time.time()
    for i in range(args.num_iter):
        loss = engine.step(i)
        if args.log and i % 10 == 0 and engine.rank == 0:
            print("Iteration {} loss {}".format(i, loss.item()))
    torch.cuda.synchronize()
    if engine.rank == 0:
        elapsed = time.time() - start
        print("{} iterations in {:.3f} s: {:.1f} images/s".format(args.num_iter, elapsed, args.num_iter * args.num_workers * args.batch / elapsed))
    engine.close()
    if dist.is_initialized():
        dist.destroy_process_group()


def main(argv=None, msmw=False):
    title = "ByzSGD (MSMW)" if msmw else "AggregaThor (SSMW)"
    args = common.finish_args(common.base_parser(title + " on garfield-b200\n" + common.__doc__).parse_args(argv))
    assert args.num_ps >= 1 and args.num_workers >= 1
    assert args.fw * 2 < args.num_workers and args.fps * 2 < args.num_ps
    common.banner(args, args.rank)
    common.seed_everything()
    if args.bench:
        torch.backends.cudnn.benchmark = True
    if args.transport == "rpc":
        return run_rpc(args, msmw)
    if args.transport == "spmd":
        return run_spmd(args, msmw)
    servers, _ = build_local(args, msmw)
    return server_loop(args, servers, msmw)
