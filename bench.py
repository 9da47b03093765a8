#!/usr/bin/env python
"""Headline benchmark: images/sec of Byzantine-resilient SSMW training of ResNet-50
(torchvision, 10 classes, 32x32 inputs — the reference's `--model resnet50 --dataset cifar10`),
n = 8 workers, f = 1 declared Byzantine worker, batch 25 per worker, SGD(lr 0.2, momentum 0.9,
weight-decay 5e-4), per aggregation rule (Krum / Median / Bulyan).

    python bench.py --gpus N --steps K --warmup W            # this framework
    python bench.py --impl reference --gpus N ...            # the unmodified reference (baseline/_ref)

For N > 1 launch under torchrun (one rank per GPU). The 8 workers are spread over the N GPUs
(rank r hosts workers r, r+N, …), rank 0 also hosts the server: total work is fixed → strong
scaling. Prints ONE JSON line (rank 0).

Timing: W warm-up steps (CUDA-graph capture and cuDNN autotuning happen there), then exactly K
steps between a barrier + torch.cuda.synchronize() on both sides, CUDA events on the launching
stream, max over ranks. `value` is measured with the step's inputs resident on the device;
`e2e` runs the same steps through the public engine API including, every step, the
pinned-host→device copy of the step's batches and a device→host read of the loss.
Every step streams the 8 x 94 MB gradient rows (753 MB > the 126 MB L2), so no step can be
served from cache.
"""

import argparse
import json
import os
import pathlib
import subprocess
import sys
import threading
import time

ROOT = pathlib.Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

MODEL, DATASET, LOSS = "resnet50", "synthetic-cifar10", "cross-entropy"
N_WORKERS, F_BYZ, BATCH = 8, 1, 25
OPT_ARGS = dict(lr=0.2, momentum=0.9, weight_decay=5e-4)
METRIC = "images/sec ResNet-50 n=8 f=1 per GAR (Krum/Median/Bulyan); GAR latency vs dim"


def parse_args(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--gar", default="krum", help="primary rule for `value` (krum | median | bulyan | ...)")
    ap.add_argument("--also", default="median,bulyan", help="other rules reported under per_gar ('' to skip)")
    ap.add_argument("--no-graphs", action="store_true")
    ap.add_argument("--fuse-workers", default="auto", choices=["auto", "on", "off"],
                    help="one batched forward/backward for the workers hosted by a GPU (auto: whenever the model is covered)")
    ap.add_argument("--fusion-chains", type=int, default=1, help="split a rank's fused workers into this many concurrent passes")
    ap.add_argument("--sync", default="flags", choices=["nccl", "flags"],
                    help="inter-rank synchronisation (N > 1): peer-mapped epoch flags (default, no NCCL call inside a step) or NCCL barriers")
    return ap.parse_args(argv)


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs."""

    QUERY = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.samples, self._stop, self._thread = index, [], threading.Event(), None

    def _loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.QUERY, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.05)

    def __enter__(self):
        self._thread = threading.Thread(target=self._loop, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._thread.join(timeout=6)

    def summary(self):
        sm = sorted(int(s[0]) for s in self.samples if s and s[0].isdigit())
        mx = [int(s[1]) for s in self.samples if len(s) > 1 and s[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(s) > 2 + i and s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.samples)}


def emit(obj):
    print(json.dumps(obj), flush=True)


# ------------------------------------------------------------------------------------- ours ---

def init_dist(gpus):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if world != gpus and rank == 0:
        print("warning: --gpus %d but WORLD_SIZE=%d (launch with torchrun --nproc-per-node %d)" % (gpus, world, gpus), file=sys.stderr)
    return rank, world, local


def timed_run(engine, steps, first_iter):
    """Device time (ms, max over ranks) of `engine.run(first_iter, steps)` — the public end-to-end loop
    (per-step pinned-host→device input copies and per-step loss read-backs included)."""
    import torch
    import torch.distributed as dist
    if engine.world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    losses = engine.run(first_iter, steps)
    stop.record()
    torch.cuda.synchronize()
    assert len(losses) == steps and all(l == l for l in losses), "every step must return a finite loss"
    ms = torch.tensor([start.elapsed_time(stop)], device=engine.device)
    if engine.world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.barrier()
    return float(ms.item()), 4 * steps


def timed_steps(engine, steps, first_iter, load, read_loss):
    """Device time (ms, max over ranks) of `steps` engine steps."""
    import torch
    import torch.distributed as dist
    if engine.world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    d2h = 0
    for i in range(steps):
        loss = engine.step(first_iter + i, load=load)
        if read_loss:
            loss.item()
            d2h += 4
    stop.record()
    torch.cuda.synchronize()
    ms = torch.tensor([start.elapsed_time(stop)], device=engine.device)
    if engine.world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.barrier()
    return float(ms.item()), d2h


def run_gar(gar, args, rank, world, local, full):
    import torch
    from garfield_b200.ops import _lib
    from garfield_b200.parallel.engine import SSMWEngine
    torch.backends.cudnn.benchmark = True
    engine = SSMWEngine(MODEL, DATASET, LOSS, N_WORKERS, F_BYZ, gar, BATCH, "sgd", OPT_ARGS, use_graphs=not args.no_graphs, sync=args.sync,
                        fuse_workers={"auto": None, "on": True, "off": False}[args.fuse_workers], fusion_chains=args.fusion_chains)
    it = 0
    for _ in range(max(args.warmup, 3)):
        engine.step(it).item()
        it += 1
    steps = args.steps if full else max(5, args.steps // 3)
    images = N_WORKERS * BATCH * steps
    launches0 = _lib.launches[0]
    with ClockSampler(local) as clocks:
        engine.load_batches(it)
        ms_dev, _ = timed_steps(engine, steps, it, load=False, read_loss=False)
        it += steps
        launches = _lib.launches[0] - launches0
        ms_e2e, d2h = timed_run(engine, steps, it)
        it += steps
    result = {
        "gar": gar, "steps": steps, "ms_per_step": ms_dev / steps, "images_per_s": images / (ms_dev / 1e3),
        "e2e_ms_per_step": ms_e2e / steps, "e2e_images_per_s": images / (ms_e2e / 1e3),
        "h2d_bytes_per_step": None,
        "d2h_bytes_per_step": d2h // steps, "gpu_launches": launches, "clocks": clocks.summary(), "d": engine.d,
        "worker_fusion": engine.fused is not None,
    }
    h2d = torch.tensor([float(engine.h2d_bytes)], device=engine.device)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(h2d)
    result["h2d_bytes_per_step"] = int(h2d.item())
    result["d2h_bytes_per_step"] = (d2h // steps) * world
    engine.close()
    del engine
    torch.cuda.empty_cache()
    return result


def run_ours(args):
    import torch
    rank, world, local = init_dist(args.gpus)
    primary = run_gar(args.gar, args, rank, world, local, full=True)
    per_gar = {args.gar: round(primary["images_per_s"], 2)}
    per_gar_e2e = {args.gar: round(primary["e2e_images_per_s"], 2)}
    for name in [g for g in args.also.split(",") if g and g != args.gar]:
        r = run_gar(name, args, rank, world, local, full=False)
        per_gar[name] = round(r["images_per_s"], 2)
        per_gar_e2e[name] = round(r["e2e_images_per_s"], 2)
    if rank == 0:
        baseline = None  # BASELINE.md publishes no number for this metric
        emit({
            "metric": METRIC, "value": round(primary["images_per_s"], 2), "unit": "images/s", "n_gpus": world,
            "steps": primary["steps"], "warmup": max(args.warmup, 3), "ms_per_step": round(primary["ms_per_step"], 4),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": baseline,
            "dtype": "fp32 (TF32 convolutions = torch default, same as the reference)", "data": "synthetic",
            "impl": "ours",
            "config": {"model": "resnet50 (torchvision, 10 classes, 3x32x32 inputs)", "global_batch": N_WORKERS * BATCH,
                       "seq_len": None, "parallelism": ("ssmw n=8 f=1 over %d gpu(s); workers r, r+N, ... on rank r; the server's aggregation + optimizer "
                                       "are sharded by coordinate range over the ranks (peer-memory loads/stores, %s)" % (
                                           world, "epoch-flag synchronisation, no NCCL call inside a step" if args.sync == "flags" else "NCCL barriers")) if world > 1
                       else "ssmw n=8 f=1, all roles on 1 gpu",
                       "gar": args.gar, "optimizer": "sgd lr=0.2 momentum=0.9 wd=5e-4 (fused in the GAR epilogue)",
                       "l2": "per-step working set 8 x 94 MB gradient rows = 753 MB > 126 MB L2 (no flush needed)",
                       "worker_fusion": ("one batched forward/backward per GPU for its %d workers (per-worker BatchNorm statistics, weight "
                                         "gradients and losses)" % (N_WORKERS // world)) if primary["worker_fusion"] else "off (one chain per worker)",
                       "d": primary["d"]},
            "per_gar": per_gar, "per_gar_e2e": per_gar_e2e,
            "e2e": {"value": round(primary["e2e_images_per_s"], 2), "unit": "images/s",
                    "h2d_bytes_per_step": primary["h2d_bytes_per_step"], "d2h_bytes_per_step": primary["d2h_bytes_per_step"],
                    "ms_per_step": round(primary["e2e_ms_per_step"], 4)},
            "gpu_launches": primary["gpu_launches"], "clocks": primary["clocks"],
        })
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


# -------------------------------------------------------------------------------- reference ---

def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    try:
        from baseline import run_reference as ref
    except Exception as err:
        if rank == 0:
            emit({"impl": "reference", "unavailable": "baseline harness import failed: %s" % (err,)})
        return
    ref.main(args, emit, METRIC, dict(model="resnet50", dataset="cifar10", loss=LOSS, n=N_WORKERS, f=F_BYZ, batch=BATCH,
                                      opt_args=OPT_ARGS), ClockSampler)


if __name__ == "__main__":
    cli = parse_args()
    if cli.impl == "reference":
        run_reference(cli)
    else:
        run_ours(cli)
