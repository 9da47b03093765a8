#!/usr/bin/env python
"""Centralized baseline: one Server and one Worker in ONE process, no RPC, `average` aggregation.

Counterpart of the reference's applications/Centralized/trainer.py:150-185 (which breaks on
current torch because its Server constructor calls rpc.remote without init_rpc):

    grad = wrk.compute_gradients(i, ps.model)[1];  ps.update_model(gar(gradients=[grad], f=0))

with MultiStepLR([25, 50] epochs, 0.1) and an accuracy report per epoch.
"""
import pathlib
import sys
import time

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[2]))

import torch  # noqa: E402

from garfield_b200.apps import common  # noqa: E402


def main(argv=None):
    ap = common.base_parser(__doc__)
    ap.add_argument("--epochs", type=int, default=0, help="Train this many epochs (overrides --num_iter).")
    args = common.finish_args(ap.parse_args(argv))
    common.banner(args, 0)
    common.seed_everything()
    from garfieldpp.server import Server
    from garfieldpp.worker import Worker
    gar = common.pick_gar("average")
    ps = Server(0, 1, 1, 1, 0, 0, " ", " ", args.batch, args.model, args.dataset, args.optimizer, args.train_size,
                device=args.device, **args.opt_args)
    wrk = Worker(1, 2, 1, args.batch, args.model, args.dataset, args.loss, args.train_size, device=args.device)
    per_epoch = max(1, len(wrk.train_set))
    scheduler = torch.optim.lr_scheduler.MultiStepLR(ps.optimizer, milestones=[25, 50], gamma=0.1)
    num_iter = args.epochs * per_epoch if args.epochs else args.num_iter
    start = time.time()
    for i in range(num_iter):
        _, grad, loss = wrk.compute_gradients(i, ps.model)
        ps.update_model(gar(gradients=[grad], f=0))
        if args.log:
            print("Iteration {} loss {}".format(i, loss))
        if (i + 1) % per_epoch == 0:
            acc = ps.compute_binary_accuracy() if args.loss == "binary-cross-entropy" else ps.compute_accuracy()
            print("Epoch: {} Accuracy: {} Time: {}".format((i + 1) // per_epoch, acc, time.time() - start))
            scheduler.step()
            sys.stdout.flush()
    return ps


if __name__ == "__main__":
    main()
