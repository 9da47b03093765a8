#!/usr/bin/env python
"""Write one TF_CONFIG-style JSON per node (the reference's interactive config_generator.py, as flags).

    python config_generator.py --ps 1 --workers 4 --host 127.0.0.1 --base_port 7000 --strategy_gradient Krum \\
        --strategy_model Median --byz_workers 1 --attack Reverse --out ./configs
"""
import argparse
import json
import pathlib


def generate(ps, workers, host="127.0.0.1", base_port=7000, strategy_gradient="Average", strategy_model="Average", byz_workers=0,
             byz_ps=0, attack="Random"):
    cluster = {"ps": ["%s:%d" % (host, base_port + i) for i in range(ps)],
               "worker": ["%s:%d" % (host, base_port + ps + i) for i in range(workers)]}
    configs = {}
    for kind, count, byz in (("ps", ps, byz_ps), ("worker", workers, byz_workers)):
        for i in range(count):
            task = {"type": kind, "index": i, "strategy_gradient": strategy_gradient, "strategy_model": strategy_model,
                    "attack": attack if i >= count - byz else "None"}
            configs["%s%d" % (kind, i)] = {"cluster": cluster, "task": task}
    return configs


if __name__ == "__main__":
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--ps", type=int, default=1)
    ap.add_argument("--workers", type=int, default=3)
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--base_port", type=int, default=7000)
    ap.add_argument("--strategy_gradient", default="Average")
    ap.add_argument("--strategy_model", default="Average")
    ap.add_argument("--byz_workers", type=int, default=0)
    ap.add_argument("--byz_ps", type=int, default=0)
    ap.add_argument("--attack", default="Random")
    ap.add_argument("--out", default="configs")
    args = ap.parse_args()
    out = pathlib.Path(args.out)
    out.mkdir(parents=True, exist_ok=True)
    for name, cfg in generate(args.ps, args.workers, args.host, args.base_port, args.strategy_gradient, args.strategy_model,
                              args.byz_workers, args.byz_ps, args.attack).items():
        (out / (name + ".json")).write_text(json.dumps(cfg, indent=1))
        print(out / (name + ".json"))
