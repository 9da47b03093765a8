"""One node of the LEARN web demo as its own process (applications/LEARN/demo.py --transport processes)."""

import sys


def node_process(rank, n, f, gar, port, epochs, batch, events):
    """One node of the demo as its own process (reference: demo.py:95-241): RPC rendezvous on localhost, a [Byz]Worker and a
    Server named node:<rank>, the LEARN loop, progress through the queue."""
    import os
    import torch
    import aggregators
    from garfield_b200.apps import common
    from garfieldpp.byzWorker import ByzWorker
    from garfieldpp.server import Server
    from garfieldpp.worker import Worker
    torch.set_num_threads(1)
    os.environ["TP_SOCKET_IFNAME"] = os.environ["GLOO_SOCKET_IFNAME"] = "lo"
    try:
        rule = aggregators.gars[gar]
        rpc = common.init_rpc("node:%d" % rank, rank, n, "127.0.0.1", port)
        if rank < f:
            worker = ByzWorker(rank, n, n, batch, "pimanet", "pima", "binary-cross-entropy", "random", device="cpu")
        else:
            worker = Worker(rank, n, n, batch, "pimanet", "pima", "binary-cross-entropy", device="cpu")
        torch.manual_seed(1234)
        ps = Server(rank, n, n, n, f, f, "node:", "node:", batch, "pimanet", "pima", "rmsprop", device="cpu", lr=0.001)
        total = epochs * max(1, worker.num_train_samples)
        for i in range(total):
            ps.update_model(rule(gradients=ps.get_gradients(i, n - f), f=f))
            ps.write_model(rule(gradients=ps.get_models(n - f), f=f))
            if i % 5 == 0 or i + 1 == total:
                events.put({"rank": rank, "progress": 100.0 * (i + 1) / total})
        events.put({"rank": rank, "progress": 100.0, "result": ps.compute_binary_accuracy() if rank >= f else None})
        rpc.shutdown()
    except Exception as err:
        events.put({"rank": rank, "error": "%s: %s" % (type(err).__name__, err)})
