#!/usr/bin/env python
"""Throughput of the device engines on the BASELINE.json configurations (synthetic data):

  cifarnet-ssmw   CifarNet, 7 workers + 1 server, Multi-Krum f=1                      (config 2)
  resnet50-msmw   ResNet-50, 3 replicated servers (1 Byzantine) + 5 workers, MDA
                  gradients, median model pull                                          (config 4)
  vgg16-learn     VGG-16, 8 decentralized peers (1 running the 'lie' attack), median   (config 5)

    python applications/benchmarks/topology_bench.py --config resnet50-msmw --steps 20
    torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 applications/benchmarks/topology_bench.py --config vgg16-learn

Timing: CUDA events around `steps` iterations after 3 warm-up iterations, max over ranks.
"""
import argparse
import json
import os
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[2]))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def build(config, batch):
    from garfield_b200.parallel.engine import SSMWEngine
    from garfield_b200.parallel.topologies import DecentralizedEngine, MSMWEngine
    sgd = dict(lr=0.05, momentum=0.9, weight_decay=5e-4)
    if config == "cifarnet-ssmw":
        eng = SSMWEngine("cifarnet", "synthetic-cifar10", "cross-entropy", 7, 1, "krum", batch, "sgd", sgd, byzantine={6: "reverse"})
        return eng, 7 * batch
    if config == "resnet50-msmw":
        eng = MSMWEngine("resnet50", "synthetic-cifar10", "cross-entropy", 5, 3, 1, 1, "brute", batch, mar="median", opt_args=sgd,
                         byzantine_workers={4: "reverse"}, byzantine_servers={2: "random"})
        return eng, 5 * batch
    if config == "vgg16-learn":
        eng = DecentralizedEngine("vgg16", "synthetic-cifar10", "cross-entropy", 8, 1, "median", batch, opt_args=dict(lr=0.01, momentum=0.9),
                                  byzantine={7: "lie"})
        return eng, 8 * batch
    raise SystemExit("unknown config %r" % config)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--config", default="cifarnet-ssmw")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--batch", type=int, default=25)
    args = ap.parse_args(argv)
    os.environ.setdefault("GARFIELD_SYNTHETIC_SIZE", "4000")
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.backends.cudnn.benchmark = True
    engine, images = build(args.config, args.batch)
    for i in range(3):
        engine.step(i)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for i in range(args.steps):
        loss = engine.step(3 + i)
    stop.record()
    torch.cuda.synchronize()
    ms = torch.tensor([start.elapsed_time(stop)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if int(os.environ.get("RANK", "0")) == 0:
        per = ms.item() / args.steps
        print(json.dumps({"config": args.config, "n_gpus": world, "ms_per_step": round(per, 3), "images_per_s": round(images / per * 1e3, 1),
                          "d": engine.d, "loss": round(float(loss), 4), "steps": args.steps}), flush=True)
    engine.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
