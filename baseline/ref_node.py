#!/usr/bin/env python
"""One node (parameter server or worker) of the reference SSMW deployment. Runs ONLY reference
code from baseline/_ref plus the two environment shims described in baseline/README.md."""

import json
import os
import pathlib
import sys
import time

HERE = pathlib.Path(__file__).resolve().parent
LIBS = HERE / "_ref" / "pytorch_impl" / "libs"
# the reference's libs first; this repository's root must NOT be importable from here
sys.path = [str(LIBS)] + [p for p in sys.path if p and pathlib.Path(p).resolve() not in (HERE.parent, HERE)]
sys.modules["native"] = None  # -> ImportError inside the reference -> pure-torch rules only

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torchvision  # noqa: E402


class _SyntheticCIFAR10(torchvision.datasets.CIFAR10):
    """CIFAR10 without download: random uint8 images of the real shape; stock __getitem__/transforms."""

    def __init__(self, root, train=True, transform=None, target_transform=None, download=False):
        torchvision.datasets.VisionDataset.__init__(self, root, transform=transform, target_transform=target_transform)
        self.train = train
        count = int(os.environ.get("REF_TRAIN_SAMPLES", "4000")) if train else int(os.environ.get("REF_TEST_SAMPLES", "500"))
        rng = np.random.RandomState(0 if train else 1)
        self.data = rng.randint(0, 256, size=(count, 32, 32, 3), dtype=np.uint8)
        self.targets = rng.randint(0, 10, size=(count,)).tolist()
        self.classes = [str(i) for i in range(10)]
        self.class_to_idx = {c: i for i, c in enumerate(self.classes)}


torchvision.datasets.CIFAR10 = _SyntheticCIFAR10


def main():
    cfg = json.loads(os.environ["REF_CONFIG"])
    rank, n, fw = int(os.environ["REF_RANK"]), cfg["n"], cfg["f"]
    world = n + 1
    torch.manual_seed(1234)
    torch.cuda.manual_seed_all(1234)
    torch.backends.cudnn.benchmark = True  # what the reference's --bench flag sets
    import torch.distributed.rpc as rpc
    import aggregators  # the REFERENCE's registry
    from garfieldpp.server import Server
    from garfieldpp.worker import Worker
    assert str(LIBS) in aggregators.__file__, aggregators.__file__
    options = rpc.TensorPipeRpcBackendOptions(rpc_timeout=600, init_method="tcp://127.0.0.1:%s" % os.environ["REF_PORT"])
    if rank == 0:
        rpc.init_rpc("ps:0", rank=0, world_size=world, rpc_backend_options=options)
        ps = Server(0, world, n, 1, fw, 0, "worker:", "ps:", cfg["batch"], cfg["model"], cfg["dataset"], "sgd", **cfg["opt_args"])
        gar = aggregators.gars.get(cfg["gar"])
        assert gar is not None

        def train_step(i):
            gradients = ps.get_gradients(i, n - fw)
            aggr_grad = gar(gradients=gradients, f=fw)
            ps.update_model(aggr_grad)

        it = 0
        for _ in range(cfg["warmup"]):
            train_step(it)
            it += 1
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(cfg["steps"]):
            train_step(it)
            it += 1
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        seconds = time.perf_counter() - t0
        d = sum(p.numel() for p in ps.model.parameters())
        with open(os.environ["REF_RESULT"], "w") as fh:
            json.dump({"seconds": seconds, "steps": cfg["steps"], "d": d, "torch": torch.__version__}, fh)
    else:
        rpc.init_rpc("worker:%d" % (rank - 1), rank=rank, world_size=world, rpc_backend_options=options)
        Worker(rank, world, n, cfg["batch"], cfg["model"], cfg["dataset"], cfg["loss"])
    rpc.shutdown()


if __name__ == "__main__":
    main()
