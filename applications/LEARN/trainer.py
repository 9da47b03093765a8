#!/usr/bin/env python
"""LEARN trainer — same command line as the reference's applications/LEARN/trainer.py
(--num_nodes --f --non_iid ... ) plus --transport/--attack (see garfield_b200/apps/common.py)."""
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[2]))

from garfield_b200.apps.learn import main  # noqa: E402

if __name__ == "__main__":
    main()
