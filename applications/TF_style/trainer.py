#!/usr/bin/env python
"""Trainer in the style of the reference's TensorFlow applications (AggregaThor / ByzSGD / LEARN):
one process per node, configured by a TF_CONFIG-like JSON file, pull-based gRPC exchange.

    python applications/TF_style/config_generator.py --ps 1 --workers 3 --out /tmp/cfg      # writes one JSON per node
    python applications/TF_style/trainer.py --config /tmp/cfg/worker0.json --max_iter 100 &   # … one per node

Flags follow tensorflow_impl/applications/AggregaThor/trainer.py:103-138:
--config --log --max_iter --dataset --model --batch_size --nbbyzwrks --native.
"""
import argparse
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[2]))

from garfield_b200.compat import Network  # noqa: E402
from garfield_b200.compat.tf_roles import PS, ByzPS, ByzWorker, Worker  # noqa: E402


def run_node(args):
    network = Network(args.config)
    common = dict(model=args.model, dataset=args.dataset, batch_size=args.batch_size, nbbyzwrks=args.nbbyzwrks, native=args.native,
                  device=args.device or None, loss=args.loss)
    attack = network.get_my_attack()
    byz = attack not in ("None", "", None)
    n_ps, n_wrk = len(network.get_all_ps()), len(network.get_all_workers())
    if network.get_task_type() == "ps":
        node = (ByzPS if byz else PS)(network, **common)
        grad_agg = node.aggregator(network.get_gradient_strategy(), n_wrk)
        model_agg = node.aggregator(network.get_model_strategy(), n_ps)
        node.commit_model()                                    # iteration 0 model
        for i in range(args.max_iter):
            if n_ps > 1:
                node.write_model(model_agg.aggregate(node.get_models(i)))
            node.upate_model(grad_agg.aggregate(node.get_gradients(i)))
            node.commit_model()
            if args.log and i % 10 == 0:
                print("ps %d iteration %d accuracy %.3f" % (network.get_task_index(), i, node.compute_accuracy()), flush=True)
    else:
        node = (ByzWorker if byz else Worker)(network, **common)
        model_agg = node.aggregator(network.get_model_strategy(), n_ps)
        for i in range(args.max_iter):
            models = node.get_models(i)
            node.write_model(model_agg.aggregate(models) if len(models) > 1 else models[0])
            loss, grad = node.compute_gradients(i)
            node.commit_gradients(grad)
            if args.log and i % 10 == 0:
                print("worker %d iteration %d loss %.4f" % (network.get_task_index(), i, loss), flush=True)
    return node


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--config", required=True)
    ap.add_argument("--log", action="store_true")
    ap.add_argument("--max_iter", type=int, default=100)
    ap.add_argument("--dataset", default="mnist")
    ap.add_argument("--model", default="convnet")
    ap.add_argument("--loss", default="nll")
    ap.add_argument("--batch_size", type=int, default=32)
    ap.add_argument("--nbbyzwrks", type=int, default=0)
    ap.add_argument("--native", action="store_true")
    ap.add_argument("--device", default="")
    node = run_node(ap.parse_args(argv))
    node.stop()


if __name__ == "__main__":
    main()
