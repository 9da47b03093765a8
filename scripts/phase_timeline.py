#!/usr/bin/env python
"""Per-phase device timeline of one SSMW step on N GPUs (torchrun), CUDA events on the launching stream, max over ranks:
wait for the previous parameters, gradient pass (CUDA graph: forward + backward + slice pushes), wait for all gradients,
aggregation (distances + slot reduction + selection + second pass with fused SGD and publication).

    torchrun --nproc-per-node N scripts/phase_timeline.py [--gar krum] [--steps 20] > profiles/timeline_nN.json
"""
import argparse
import json
import os
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gar", default="krum")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--model", default="resnet50")
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from garfield_b200.parallel.engine import SSMWEngine
    torch.backends.cudnn.benchmark = True
    eng = SSMWEngine(args.model, "synthetic-cifar10", "cross-entropy", 8, 1, args.gar, 25, "sgd", dict(lr=0.2, momentum=0.9, weight_decay=5e-4))
    eng._whole_step = False          # phases must be separate launches to be timed
    for i in range(6):
        eng.step(i)
    torch.cuda.synchronize()
    names = ["wait_params", "gradient_pass", "signal+wait_grads", "aggregate+sgd+publish"]
    sums = {k: 0.0 for k in names}
    total = 0.0
    for i in range(args.steps):
        eng.load_batches(6 + i)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        ev[0].record()
        if world > 1:
            eng.epoch += 1
            eng._bump()
            eng._wait(eng._param_flags, lag=1)
        ev[1].record()
        eng.compute_gradients()
        ev[2].record()
        if world > 1:
            eng._signal(eng._my_grad_flag)
            eng._wait(eng._grad_flags)
        ev[3].record()
        eng.aggregate_and_update()
        if world > 1:
            eng._signal(eng._my_param_flag)
        ev[4].record()
        torch.cuda.synchronize()
        for k, name in enumerate(names):
            sums[name] += ev[k].elapsed_time(ev[k + 1])
        total += ev[0].elapsed_time(ev[4])
        eng.steps_done += 1
    out = {k: v / args.steps for k, v in sums.items()}
    out["step_ms"] = total / args.steps
    t = torch.tensor([out[k] for k in names] + [out["step_ms"]], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if int(os.environ.get("RANK", "0")) == 0:
        vals = t.tolist()
        print(json.dumps({"n_gpus": world, "gar": args.gar, "model": args.model, "phases_ms_max_over_ranks": dict(zip(names, [round(v, 4) for v in vals[:-1]])),
                          "step_ms": round(vals[-1], 4), "note": "phases launched separately (no whole-step graph): the sum exceeds the graphed step"}))
    eng.close()
    if world > 1:
        dist.destroy_process_group()


main()
