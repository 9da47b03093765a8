#!/usr/bin/env python
"""The same single-process loop on a CIFAR-style ResNet (reference: rsrcs/playgrounds/centralized/resnet.py).

    python applications/playgrounds/centralized/resnet.py --model resnet18 --epochs 1
"""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent))

from cifar10 import main  # noqa: E402

if __name__ == "__main__":
    main(model="resnet18")
