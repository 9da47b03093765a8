"""`Worker` — computes one mini-batch gradient of the model it is handed.

API parity with /root/reference/pytorch_impl/libs/garfieldpp/worker.py:52-96:

    Worker(rank, world_size, num_workers, batch_size, model, dataset, loss, train_size=None)
    .compute_gradients(iter_num, model) -> (rank, flat_grad, loss_float)
    .train_set (list of (data, target))   .num_train_samples

B200-native differences: nothing is pickled or copied to the host. `model` is either a plain
module / `FlatModel` living on this worker's device (co-located roles share it), or — over the
RPC compatibility transport — the flat parameter vector plus buffers (`compute_gradients_flat`),
loaded into a worker-local replica. When `out=` is given the gradient is accumulated straight
into that vector (a row of the [n, d] gradient matrix) by binding `param.grad` views to it.
"""

from __future__ import annotations

import logging

import torch

from . import tools
from .datasets import DatasetManager
from ..parallel.arena import FlatModel

logger = logging.getLogger(__name__)

__all__ = ["Worker"]


def _pick_device(device=None):
    if device is not None:
        return torch.device(device)
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


class Worker:
    """Byzantine-resilient worker (honest behaviour)."""

    def __init__(self, rank, world_size, num_workers, batch_size, model, dataset, loss, train_size=None, device=None):
        self.device = _pick_device(device)
        self.rank = rank
        self.world_size = world_size
        self.num_workers = num_workers
        self.batch_size = batch_size
        self.model_name, self.dataset_name = model, dataset
        self.loss = tools.select_loss(loss)
        manager = DatasetManager(dataset, batch_size, num_workers, world_size, rank, train_size)
        self.train_set = manager.get_train_set()
        self.num_train_samples = len(self.train_set)
        self._replica = None  # lazily built local model for the flat-parameter protocol
        tools.set_worker(self)

    # ------------------------------------------------------------------------------------------
    def batch(self, iter_num):
        data, target = self.train_set[iter_num % self.num_train_samples]
        return data.to(self.device, non_blocking=True), target.to(self.device, non_blocking=True)

    def forward_backward(self, iter_num, module):
        """One forward + backward of `module` on this worker's batch `iter_num`; returns the loss tensor."""
        module.train()
        data, target = self.batch(iter_num)
        loss = self.loss(module(data), target)
        loss.backward()
        return loss

    def compute_gradients(self, iter_num, model, out=None):
        """Gradient of the loss on batch `iter_num` w.r.t. the parameters of `model`.

        Returns `(rank, flat_grad, loss_value)`; `flat_grad` is on the worker's device in
        `parameters()` order (it is `out` when given).
        """
        flat_model = model if isinstance(model, FlatModel) else None
        module = flat_model.module if flat_model is not None else model
        if next(module.parameters()).device != self.device:
            module = module.to(self.device)
        if out is not None and flat_model is not None:
            out.zero_()
            flat_model.bind_grads(out)
            loss = self.forward_backward(iter_num, module)
            flat_model.unbind_grads()
            return self.rank, out, loss.item()
        for p in module.parameters():
            p.grad = None
        loss = self.forward_backward(iter_num, module)
        grads = [(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in module.parameters()]
        flat = torch.cat(grads)
        for p in module.parameters():
            p.grad = None
        if out is not None:
            out.copy_(flat)
            flat = out
        return self.rank, flat, loss.item()

    # -- flat-parameter protocol used by the RPC compatibility transport ---------------------------
    def _local_replica(self):
        if self._replica is None:
            module = tools.select_model(self.model_name, self.device, self.dataset_name)
            self._replica = FlatModel(module, self.device)
        return self._replica

    def compute_gradients_flat(self, iter_num, flat_params, buffers=None):
        """Same as `compute_gradients` but the model arrives as its flat parameter vector (+ buffers).
        The reply is a CPU tensor when the request was one (torch RPC moves CPU tensors)."""
        replica = self._local_replica()
        replica.write(flat_params)
        if buffers:
            own = dict(replica.module.named_buffers())
            with torch.no_grad():
                for name, value in buffers.items():
                    own[name].copy_(value)
        rank, grad, loss = self.compute_gradients(iter_num, replica)
        return rank, grad.to(flat_params.device), loss
