"""`Server` — the (possibly replicated) parameter server.

API parity with /root/reference/pytorch_impl/libs/garfieldpp/server.py:58-305:

    Server(rank, world_size, num_workers, num_ps, byz_wrk, byz_ps, wrk_base_name, ps_base_name,
           batch, model, dataset, optimizer, train_size=None, *opt_args, **opt_kwargs)
    get_gradients(iter_num, num_wait_wrk=-1)   get_models(num_wait_ps=-1)   get_aggr_grads(num_wait_ps=-1)
    get_model()  get_latest_aggr_grad()  update_model(flat_grad)  write_model(flat_model)
    build_graph(iter_num)  compute_accuracy()  compute_binary_accuracy()
    .model .optimizer .train_set .test_set .device .latest_aggr_grad

B200-native design: the model lives on the device as a `FlatModel` (parameters are views of one
flat fp32 arena); `update_model` is ONE fused optimizer kernel on that arena, `write_model` one
copy; gradients come back as rows of a device-resident [n, d] matrix that the aggregation
kernels read in place (no `.to(cpu)`, no pickled module, no per-parameter narrow/reshape).

Transports:
  * local  — peers are Python objects in this process (`attach_workers`, `attach_servers`, or the
             RPC-free single-process mode world_size <= 1 that the reference's Centralized app
             needs);
  * rpc    — `torch.distributed.rpc` (TensorPipe) with the reference's naming scheme
             (`wrk_base_name + i`, `ps_base_name + i`); the request carries the flat parameter
             vector and BN buffers instead of a pickled module.
The device-to-device SPMD data plane (peer-memory gather fused with the GAR) lives in
`garfield_b200.parallel`; this class is the reference-compatible control surface on top.
"""

from __future__ import annotations

import logging
import threading
import time

import torch

from . import tools
from .datasets import DatasetManager
from ..parallel.arena import FlatModel, GradMatrix

logger = logging.getLogger(__name__)

__all__ = ["Server"]


def _rpc_ready():
    try:
        from torch.distributed.rpc import api
        return api._is_current_rpc_agent_set()
    except Exception:
        return False


class Server:
    """Byzantine-resilient parameter server."""

    def __init__(self, rank, world_size, num_workers, num_ps, byz_wrk, byz_ps, wrk_base_name, ps_base_name, batch, model,
                 dataset, optimizer, train_size=None, *args, device=None, **kwargs):
        self.device = torch.device(device) if device is not None else (
            torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu"))
        self.rank, self.world_size = rank, world_size
        self.num_workers, self.num_ps = num_workers, num_ps
        self.byz_wrk, self.byz_ps = byz_wrk, byz_ps
        self.wrk_base_name, self.ps_base_name = wrk_base_name, ps_base_name
        self.model_name, self.dataset_name = model, dataset

        module = tools.select_model(model, self.device, dataset)
        self.flat_model = FlatModel(module, self.device)
        self.model = self.flat_model.module
        self.model._gb_flat = self.flat_model
        manager = DatasetManager(dataset, batch * max(1, num_workers), 1, 2, 1, train_size)
        self.test_set = manager.get_test_set()
        self.train_set = manager.get_train_set()
        self.optimizer = tools.select_optimizer(self.flat_model, optimizer, *args, **kwargs)
        self.latest_aggr_grad = None
        self._aggr_ready = threading.Condition()
        self._grads = None
        self.last_losses = []
        self.last_contributors = []

        # peers
        self._local_workers, self._local_servers = [], []
        self.workers_rref, self.workers_types, self.ps_rref, self.ps_types = [], [], [], []
        self._use_rpc = world_size > 1 and _rpc_ready()
        tools.set_server(self)  # announce before looking up peers (they may be looking us up)
        if self._use_rpc:
            self.workers_types, self.workers_rref = self.get_rrefs(wrk_base_name, 0, num_workers, True)
            if num_ps > 0:
                self.ps_types, self.ps_rref = self.get_rrefs(ps_base_name, 0, num_ps, False)

    # ------------------------------------------------------------------------------ peers ---
    def get_rrefs(self, base_name, base_id, num_nodes, worker=True):
        """Remote references (and their types) of `num_nodes` peers named `base_name + id`."""
        from torch.distributed.rpc import remote
        rrefs = [remote(base_name + str(i), tools.get_worker if worker else tools.get_server)
                 for i in range(base_id, base_id + num_nodes)]
        from .worker import Worker
        types = [Worker if worker else Server for _ in rrefs]  # duck-typed: only method names matter
        return types, rrefs

    def attach_workers(self, workers):
        """Local transport: the worker objects living in this process (co-located roles)."""
        self._local_workers = list(workers)
        self.num_workers = len(self._local_workers)

    def attach_servers(self, servers):
        """Local transport: the replica servers living in this process (including this one)."""
        self._local_servers = list(servers)
        self.num_ps = len(self._local_servers)

    def _grad_matrix(self):
        if self._grads is None or self._grads.n != self.num_workers:
            self._grads = GradMatrix(self.num_workers, self.flat_model.numel, self.device)
        return self._grads

    # ---------------------------------------------------------------------- gradient pull ---
    def get_gradients(self, iter_num, num_wait_wrk=-1):
        """Ask every worker for its gradient on iteration `iter_num`; return the first
        `num_wait_wrk` (default n − byz_wrk) as device tensors (rows of the gradient matrix)."""
        if num_wait_wrk < 0:
            num_wait_wrk = self.num_workers - self.byz_wrk
        self.model.train()
        grads = self._grad_matrix()
        if self._local_workers:
            return self._get_gradients_local(iter_num, num_wait_wrk, grads)
        if self._use_rpc:
            return self._get_gradients_rpc(iter_num, num_wait_wrk, grads)
        raise RuntimeError("no workers attached and RPC is not initialised")

    def _get_gradients_local(self, iter_num, q, grads):
        out, losses, who = [], [], []
        for i, wrk in enumerate(self._local_workers):
            rank, g, loss = wrk.compute_gradients(iter_num, self.flat_model, out=grads.row(i))
            out.append(g); losses.append(loss); who.append(rank)
        # synchronous in-process execution: the "fastest q" are the first q in worker order
        self.last_losses, self.last_contributors = losses[:q], who[:q]
        return out[:q]

    def _get_gradients_rpc(self, iter_num, q, grads):
        from .worker import Worker
        flat_cpu = self.flat_model.flat.detach().to("cpu")
        buffers = {k: v.detach().to("cpu") for k, v in self.flat_model.buffers_state().items()}
        n = self.num_workers
        slots, losses = [None] * n, [None] * n
        arrived = threading.Condition()

        def on_reply(index):
            def cb(fut):
                try:
                    rank, grad, loss = fut.wait()
                except Exception as err:  # a crashed/Byzantine peer must not take the server down
                    logger.warning("worker %d failed: %s", index, err)
                    return
                grads.row(index).copy_(grad, non_blocking=True)
                with arrived:
                    slots[index], losses[index] = grads.row(index), loss
                    arrived.notify_all()
            return cb

        for i, rref in enumerate(self.workers_rref):
            fut = tools._remote_method_async(Worker.compute_gradients_flat, rref, iter_num, flat_cpu, buffers)
            fut.then(on_reply(i))
        builder = threading.Thread(target=self.build_graph, args=(iter_num,))
        builder.start()
        with arrived:
            arrived.wait_for(lambda: sum(s is not None for s in slots) >= q)
            ready = [i for i in range(n) if slots[i] is not None][:max(q, 0)] if q < n else [i for i in range(n) if slots[i] is not None]
        builder.join()
        self.last_losses = [losses[i] for i in ready]
        self.last_contributors = ready
        return [slots[i] for i in ready]

    def build_graph(self, iter_num):
        """Server-side train-mode forward on its own batch (keeps the server's BN running
        statistics moving, as the reference does — server.py:186-194)."""
        if not any(True for _ in self.model.buffers()) or not self.train_set:
            return
        data, _ = self.train_set[iter_num % len(self.train_set)]
        with torch.no_grad():
            self.model.train()
            self.model(data.to(self.device, non_blocking=True))

    # ------------------------------------------------------------------------- model pull ---
    def get_model(self):
        """`(rank, flat parameter vector)` — a copy, in parameters() order."""
        return self.rank, self.flat_model.flat.detach().clone()

    def _get_model_cpu(self):
        rank, flat = self.get_model()
        return rank, flat.to("cpu")

    def _gather_from_servers(self, local_call, remote_method, num_wait_ps):
        if num_wait_ps < 0:
            num_wait_ps = self.num_ps - self.byz_ps
        if self._local_servers:
            return [local_call(ps)[1].to(self.device) for ps in self._local_servers][:num_wait_ps]
        if not self._use_rpc:
            raise RuntimeError("no servers attached and RPC is not initialised")
        n = self.num_ps
        slots = [None] * n
        arrived = threading.Condition()

        def on_reply(index):
            def cb(fut):
                try:
                    _, vec = fut.wait()
                except Exception as err:
                    logger.warning("server %d failed: %s", index, err)
                    return
                vec = vec.to(self.device)
                with arrived:
                    slots[index] = vec
                    arrived.notify_all()
            return cb

        for i, rref in enumerate(self.ps_rref):
            tools._remote_method_async(remote_method, rref).then(on_reply(i))
        with arrived:
            arrived.wait_for(lambda: sum(s is not None for s in slots) >= num_wait_ps)
            return [s for s in slots if s is not None]

    def get_models(self, num_wait_ps=-1):
        """Flat models of the server replicas (first `num_wait_ps`, default num_ps − byz_ps)."""
        return self._gather_from_servers(lambda ps: ps.get_model(), type(self)._get_model_cpu, num_wait_ps)

    def get_latest_aggr_grad(self, timeout=None):
        """`(rank, latest aggregated gradient)`; blocks until one has been published."""
        with self._aggr_ready:
            self._aggr_ready.wait_for(lambda: self.latest_aggr_grad is not None, timeout)
            return self.rank, self.latest_aggr_grad

    def _get_latest_aggr_grad_cpu(self):
        rank, grad = self.get_latest_aggr_grad()
        return rank, grad.to("cpu")

    def __setattr__(self, name, value):
        if name == "latest_aggr_grad" and "_aggr_ready" in self.__dict__:
            with self._aggr_ready:
                object.__setattr__(self, name, value)
                self._aggr_ready.notify_all()
        else:
            object.__setattr__(self, name, value)

    def get_aggr_grads(self, num_wait_ps=-1):
        """Latest aggregated gradients of the server replicas (LEARN agreement rounds)."""
        return self._gather_from_servers(lambda ps: ps.get_latest_aggr_grad(), type(self)._get_latest_aggr_grad_cpu, num_wait_ps)

    # -------------------------------------------------------------------------- updates ---
    def update_model(self, grad):
        """Apply one optimizer step with the flat gradient `grad` (one fused kernel on the arena)."""
        self.optimizer.step(grad)

    def write_model(self, model):
        """Overwrite the parameters with the flat vector `model`."""
        self.flat_model.write(model)

    # ------------------------------------------------------------------------- evaluation ---
    def _evaluate(self, predict):
        correct = total = 0
        was_training = self.model.training
        self.model.eval()
        with torch.no_grad():
            for inputs, targets in self.test_set:
                inputs = inputs.to(self.device, non_blocking=True)
                targets = targets.to(self.device, non_blocking=True)
                predicted = predict(self.model(inputs))
                total += targets.size(0)
                correct += predicted.eq(targets).sum()
        self.model.train(was_training)
        return float(correct) * 100 / max(total, 1)

    def compute_accuracy(self):
        """Top-1 accuracy (percent) of the current model on the test set, evaluated in place on
        the device (no deepcopy of the model)."""
        return self._evaluate(lambda out: out.max(1)[1])

    def compute_binary_accuracy(self):
        return self._evaluate(lambda out: out.round())

    # ---------------------------------------------------------------- RPC benchmark helpers ---
    def get_fake_models(self):
        """Pull the fake model of every server replica (rpc_bench); returns the list of tensors."""
        return self._gather_from_servers(lambda ps: (ps.rank, ps.get_fake_model()),
                                         type(self)._get_fake_model_pair, self.num_ps)

    def get_fake_model(self):
        return self.model if isinstance(self.model, torch.Tensor) else self.flat_model.flat

    def _get_fake_model_pair(self):
        vec = self.get_fake_model()
        return self.rank, vec.to("cpu")
