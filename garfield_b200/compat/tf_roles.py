"""Pull-based roles of the reference's TensorFlow implementation, framework-neutral.

Parity with tensorflow_impl/libs/{server,ps,worker,byz_ps,byz_worker}.py (SURVEY.md §2.7 T1–T3): every
node runs a gRPC `MessageExchange` service that serves its per-iteration history of models (PS) or
gradients (worker); peers PULL what they need for iteration i (blocking until it is committed):

    worker loop:  models = get_models(i) → aggregate (model strategy) → write_model → compute_gradients → commit_gradients
    PS loop:      [models = get_models(i) → aggregate → write_model] → gradients = get_gradients(i) → aggregate → upate_model → commit_model

Models are torch modules on this framework's flat arenas (TensorFlow is not installed); payloads are
float32 numpy vectors exactly as in the reference (`ndarray.tobytes()`), aggregation goes through
`Aggregator_tf`, attacks through `Attacker`.
"""

from __future__ import annotations

import numpy as np
import torch

from . import grpc_exchange
from .aggregator_tf import Aggregator_tf
from .attacker import Attacker
from .network import Network
from ..garfieldpp import tools
from ..garfieldpp.datasets import DatasetManager
from ..parallel.arena import FlatModel

__all__ = ["Server", "PS", "Worker", "ByzPS", "ByzWorker"]


class Server:
    """Base node: model, dataset, gRPC service and clients to every PS and worker of the cluster."""

    def __init__(self, network, model="convnet", dataset="mnist", batch_size=32, nbbyzwrks=0, native=False, device=None, loss="nll"):
        self.network = network if isinstance(network, Network) else Network(network)
        self.device = torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.batch_size, self.nbbyzwrks, self.native = batch_size, nbbyzwrks, native
        torch.manual_seed(1234)
        self.flat_model = FlatModel(tools.select_model(model, self.device, dataset), self.device)
        self.loss_fn = tools.select_loss(loss)
        self.dataset_name = dataset
        self.servicer = grpc_exchange.MessageExchangeServicer(timeout=120)
        self.grpc_server = grpc_exchange.serve(self.servicer, self.network.get_my_port())
        self._ps_clients = [grpc_exchange.Client(addr) for addr in self.network.get_all_ps()]
        self._worker_clients = [grpc_exchange.Client(addr) for addr in self.network.get_all_workers()]
        n_workers = max(1, len(self.network.get_all_workers()))
        self._manager = DatasetManager(dataset, batch_size, n_workers, n_workers + len(self.network.get_all_ps()),
                                       len(self.network.get_all_ps()) + (self.network.get_task_index() if self.network.get_task_type() == "worker" else 0))
        self.test_set = None

    def get_models(self, iteration):
        """Flat models of every PS for `iteration` (blocking pulls)."""
        return [client.get_model(iteration) for client in self._ps_clients]

    def write_model(self, flat):
        self.flat_model.write(torch.from_numpy(np.array(flat, dtype=np.float32, copy=True)).to(self.device))   # gRPC payloads are read-only buffers

    def flat(self):
        return self.flat_model.flat.detach().cpu().numpy().copy()

    def aggregator(self, strategy, count):
        return Aggregator_tf(strategy, count, self.nbbyzwrks, native=self.native, device=self.device)

    def compute_accuracy(self):
        if self.test_set is None:
            self.test_set = self._manager.get_test_set()
        module = self.flat_model.module
        module.eval()
        correct = total = 0
        with torch.no_grad():
            for x, y in self.test_set:
                pred = module(x.to(self.device)).max(1)[1]
                correct += int(pred.eq(y.to(self.device)).sum())
                total += y.numel()
        module.train()
        return correct / max(total, 1)

    def stop(self):
        for client in self._ps_clients + self._worker_clients:
            client.close()
        self.grpc_server.stop(0)


class PS(Server):
    def __init__(self, *args, lr=1e-3, **kwargs):
        super().__init__(*args, **kwargs)
        self.optimizer = tools.select_optimizer(self.flat_model, "adam", lr=lr)   # the TF PS uses Adam(1e-3)

    def get_gradients(self, iteration):
        return [client.get_gradient(iteration) for client in self._worker_clients]

    def upate_model(self, flat_grad):  # (sic) the reference spells it this way: tensorflow_impl/libs/ps.py
        self.optimizer.step(torch.from_numpy(np.ascontiguousarray(flat_grad, dtype=np.float32)).to(self.device))

    update_model = upate_model

    def commit_model(self):
        self.servicer.commit_model(self.flat())


class Worker(Server):
    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.train_set = self._manager.get_train_set()

    def compute_gradients(self, iteration):
        x, y = self.train_set[iteration % len(self.train_set)]
        module = self.flat_model.module
        for p in module.parameters():
            p.grad = None
        module.train()
        loss = self.loss_fn(module(x.to(self.device)), y.to(self.device))
        loss.backward()
        grad = torch.cat([p.grad.reshape(-1) for p in module.parameters()])
        return loss.item(), grad.detach().cpu().numpy()

    def commit_gradients(self, grad):
        self.servicer.commit_gradient(grad)


class ByzWorker(Worker):
    """Commits an attacked gradient (`task.attack` of the config: Random, Reverse, PartialDrop, …)."""

    def commit_gradients(self, grad, byz_gradients=()):
        name = self.network.get_my_attack()
        attacker = Attacker(name)
        if name in ("LittleIsEnough", "FallEmpires"):
            grad = attacker.attack(gradient=grad, byz_gradients=list(byz_gradients))
        elif name == "PartialDrop":
            grad = attacker.attack(gradient=grad, probability=0.3)
        else:
            grad = attacker.attack(gradient=grad)
        self.servicer.commit_gradient(grad)


class ByzPS(PS):
    """Commits an attacked model."""

    def commit_model(self):
        name = self.network.get_my_attack()
        attacker = Attacker(name)
        flat = self.flat()
        flat = attacker.attack(gradient=flat, probability=0.3) if name == "PartialDrop" else attacker.attack(gradient=flat)
        self.servicer.commit_model(flat)
