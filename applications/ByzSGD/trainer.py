#!/usr/bin/env python
"""ByzSGD trainer — same command line as the reference's applications/ByzSGD/trainer.py.

    python trainer.py --dataset cifar10 --model resnet50 --loss cross-entropy --batch 25 \
        --num_workers 8 --fw 1 --gar krum --opt_args '{"lr":"0.2","momentum":"0.9","weight_decay":"0.0005"}'

--transport local (default) runs every role in this process on one device; --transport rpc runs
one role per process (--master/--rank as in the reference's run_exp.sh); --transport spmd runs
the device-resident engine under torchrun (one process per GPU).
"""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[2]))

from garfield_b200.apps.ssmw import main  # noqa: E402

if __name__ == "__main__":
    main(msmw=True)
