#!/usr/bin/env python
"""Single-process CIFAR-10 training of a zoo model through the library's own pieces (DatasetManager, select_model,
FlatOptimizer, the sm_100a loss / optimizer kernels when a GPU is present) — the counterpart of the reference's
rsrcs/playgrounds/centralized/cifar10.py (the PyTorch tutorial loop the authors used as a sanity baseline).

    python applications/playgrounds/centralized/cifar10.py --model resnet18 --epochs 1 --batch 128
"""
import argparse
import pathlib
import sys
import time

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[3]))

import torch  # noqa: E402

from garfield_b200.garfieldpp import tools  # noqa: E402
from garfield_b200.garfieldpp.datasets import DatasetManager  # noqa: E402
from garfield_b200.ops.optim import FlatOptimizer  # noqa: E402
from garfield_b200.parallel.arena import FlatModel  # noqa: E402


def main(argv=None, model="cifarnet"):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--model", default=model)
    ap.add_argument("--dataset", default="cifar10")
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--lr", type=float, default=0.05)
    ap.add_argument("--max_steps", type=int, default=0, help="stop after this many steps (0 = whole epochs)")
    args = ap.parse_args(argv)
    device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
    manager = DatasetManager(args.dataset, args.batch, 1, 2, 1)
    train, test = manager.get_train_set(), manager.get_test_set()
    flat_model = FlatModel(tools.select_model(args.model, device, args.dataset), device)
    grad = torch.zeros_like(flat_model.flat)
    flat_model.bind_grads(grad)
    optimizer = FlatOptimizer(flat_model.flat, "sgd", lr=args.lr, momentum=0.9, weight_decay=5e-4)
    loss_fn = tools.select_loss("cross-entropy")
    steps = 0
    for epoch in range(args.epochs):
        flat_model.module.train()
        start, running = time.time(), 0.0
        for x, y in train:
            grad.zero_()
            loss = loss_fn(flat_model.module(x.to(device)), y.to(device))
            loss.backward()
            optimizer.step(grad)
            running += float(loss.detach())
            steps += 1
            if args.max_steps and steps >= args.max_steps:
                break
        flat_model.module.eval()
        correct = total = 0
        with torch.no_grad():
            for x, y in test:
                correct += int(flat_model.module(x.to(device)).argmax(1).eq(y.to(device)).sum())
                total += y.numel()
        print("epoch %d  loss %.4f  test accuracy %.2f %%  (%.1f s)" % (epoch, running / max(1, len(train)), 100.0 * correct / max(1, total),
                                                                      time.time() - start), flush=True)
        if args.max_steps and steps >= args.max_steps:
            break
    return 100.0 * correct / max(1, total)


if __name__ == "__main__":
    main()
