#!/usr/bin/env python
"""Playground trainer: the Aggregathor (SSMW) loop whose declared-Byzantine workers really run the `random` attack —
the counterpart of the reference's rsrcs/playgrounds/trainer.py (an Aggregathor variant built on `ByzWorker('random')`).

    python applications/playgrounds/trainer.py --dataset mnist --model convnet --num_workers 7 --fw 1 --gar krum --num_iter 200   (Krum needs n - f >= 2f + 3 replies)
"""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[2]))

from garfield_b200.apps.ssmw import main  # noqa: E402

if __name__ == "__main__":
    argv = sys.argv[1:]
    if not any(a == "--attack" or a.startswith("--attack=") for a in argv):
        argv += ["--attack", "random"]
    main(argv, msmw=False)
