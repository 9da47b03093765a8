#!/usr/bin/env python
"""Plain (non-robust) data-parallel SGD with one flat all-reduce per step — the "vanilla" baseline the
robust topologies are compared against. Counterpart of the reference's
rsrcs/playgrounds/legacy_code/train_dist.py (per-parameter dist.reduce). Thin wrapper over Garfield_CC
in vanilla mode:

    torchrun --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 applications/playgrounds/train_dist.py --num_workers 3
"""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[2]))

from garfield_b200.apps.cc import main  # noqa: E402

if __name__ == "__main__":
    main(sys.argv[1:] + ["--gar", "vanilla", "--mar", "vanilla"])
