#!/usr/bin/env python
"""Garfield_CC trainer (collective-communication variant). Launch one process per rank, e.g.

    torchrun --nnodes=1 --nproc-per-node 9 --master-addr 127.0.0.1 applications/Garfield_CC/trainer.py \\
        --num_ps 3 --num_workers 6 --fw 1 --fps 1 --aggregator brute --mar median --model resnet50 --dataset cifar10 \\
        --loss cross-entropy --batch 25 --opt_args '{"lr":"0.2","momentum":"0.9"}'
"""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[2]))

from garfield_b200.apps.cc import main  # noqa: E402

if __name__ == "__main__":
    main()
