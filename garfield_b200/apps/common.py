"""Shared CLI and loop helpers of the trainers.

The flags are the reference's (applications/Aggregathor/trainer.py:62-183: --master --rank
--dataset --batch --num_ps --num_workers --fw --fps --model --loss --optimizer --opt_args
--num_iter --gar --acc_freq --bench --log) plus:

  --transport {local,rpc,spmd}   local = every role in this process on one device (default);
                                 rpc   = one role per process over torch RPC, as the reference;
                                 spmd  = one process per GPU under torchrun, device-resident engine
  --attack NAME                  attack used by the Byzantine workers (default: none → honest)
  --device DEV                   cuda | cpu (default: cuda when available)
  --checkpoint PATH / --resume PATH / --ckpt_freq N     state_dict checkpoints
  --trace PATH                   write a chrome-trace JSON of the step phases
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import time

import torch

CIFAR_NUM_SAMPLES = 50000


def str2bool(text):
    if isinstance(text, bool):
        return text
    return str(text).lower() not in ("", "0", "false", "no", "n")


def base_parser(description, decentralized=False):
    ap = argparse.ArgumentParser(description=description, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--master", type=str, default="127.0.0.1", help="Master node (rank 0) address.")
    ap.add_argument("--rank", type=int, default=0, help="Rank of this process (rpc transport).")
    ap.add_argument("--dataset", type=str, default="mnist")
    ap.add_argument("--batch", type=int, default=32, help="Minibatch size per worker.")
    if decentralized:
        ap.add_argument("--num_nodes", type=int, default=1, help="Number of nodes (each is server + worker).")
        ap.add_argument("--f", type=int, default=0, help="Number of declared Byzantine nodes.")
        ap.add_argument("--non_iid", type=str2bool, default=False, help="Run the multi-round agreement on aggregated gradients.")
    else:
        ap.add_argument("--num_ps", type=int, default=1)
        ap.add_argument("--num_workers", type=int, default=1)
        ap.add_argument("--fw", type=int, default=0, help="Declared Byzantine workers.")
        ap.add_argument("--fps", type=int, default=0, help="Declared Byzantine servers.")
    ap.add_argument("--model", type=str, default="convnet")
    ap.add_argument("--loss", type=str, default="nll")
    ap.add_argument("--optimizer", type=str, default="sgd")
    ap.add_argument("--opt_args", type=json.loads, default={"lr": "0.1"}, help="JSON dict of optimizer arguments.")
    ap.add_argument("--num_iter", type=int, default=5000)
    ap.add_argument("--gar", type=str, default="average")
    ap.add_argument("--acc_freq", type=int, default=100)
    ap.add_argument("--bench", type=str2bool, default=False)
    ap.add_argument("--log", type=str2bool, default=False)
    ap.add_argument("--transport", choices=["local", "rpc", "spmd"], default="local")
    ap.add_argument("--attack", type=str, default="", help="Attack of the Byzantine workers/nodes ('' = honest).")
    ap.add_argument("--server_attack", type=str, default="", help="Attack of the Byzantine servers ('' = honest).")
    ap.add_argument("--device", type=str, default="")
    ap.add_argument("--port", type=int, default=29500)
    ap.add_argument("--checkpoint", type=str, default="")
    ap.add_argument("--resume", type=str, default="")
    ap.add_argument("--ckpt_freq", type=int, default=0)
    ap.add_argument("--trace", type=str, default="")
    ap.add_argument("--train_size", type=int, default=None)
    return ap


def finish_args(args):
    args.opt_args = {k: float(v) for k, v in args.opt_args.items()}
    assert "lr" in args.opt_args and args.opt_args["lr"], "opt_args needs a learning rate"
    assert args.num_iter > 0 and args.batch >= 1
    if not args.device:
        args.device = "cuda" if torch.cuda.is_available() else "cpu"
    return args


def banner(args, rank):
    print("**** SETUP AT NODE {} ***".format(rank))
    for key, value in sorted(vars(args).items()):
        print("%-16s %s" % (key + ":", value))
    print("------------------------------------")
    sys.stdout.flush()


def seed_everything(seed=1234):
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)


def pick_gar(name):
    import aggregators
    gar = aggregators.gars.get(name)
    assert gar is not None, "unknown GAR %r; available: %s" % (name, sorted(aggregators.gars))
    return gar


def iterations_per_epoch(num_workers, batch):
    return max(1, CIFAR_NUM_SAMPLES // (num_workers * batch))


class StepTimer:
    """Per-step wall time (with a device sync when CUDA) and optional NIC byte delta (--bench)."""

    def __init__(self, enabled, device):
        self.enabled, self.cuda = enabled, str(device).startswith("cuda")

    def run(self, index, fn):
        if not self.enabled:
            return fn()
        from garfieldpp.tools import convert_to_gbit, get_bytes_com
        before = get_bytes_com()
        if self.cuda:
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = fn()
        if self.cuda:
            torch.cuda.synchronize()
        print("Training step {} takes {} seconds".format(index, time.perf_counter() - t0))
        print("Consumed bandwidth in this iteration: {} Gbits".format(convert_to_gbit(get_bytes_com() - before)))
        sys.stdout.flush()
        return out


def init_rpc(name, rank, world_size, master, port):
    import torch.distributed.rpc as rpc
    os.environ["MASTER_ADDR"] = master
    os.environ["MASTER_PORT"] = str(port)
    options = rpc.TensorPipeRpcBackendOptions(rpc_timeout=600, init_method="tcp://%s:%d" % (master, port))
    rpc.init_rpc(name, rank=rank, world_size=world_size, rpc_backend_options=options)
    return rpc
