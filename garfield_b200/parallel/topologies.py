"""Device-resident engines for the replicated-server (MSMW / ByzSGD) and decentralized (LEARN)
topologies, built from the same fused gather+GAR kernels as the SSMW engine.

Every model replica lives in a peer-mapped parameter arena, every gradient in a peer-mapped
gradient row; "pulling" a peer's model or gradient is just putting its peer pointer in the row
table of an aggregation kernel, so the reference's

    models = ps.get_models(num_ps - fps);  ps.write_model(gar(models, f=fps))        (ByzSGD/trainer.py:241-244)
    models = ps.get_models(n - f) → gar → ps.write_model      ("contract", LEARN/trainer.py:252-258)
    aggr_grads = ps.get_aggr_grads(n - f) → gar               (avg_agree, LEARN/trainer.py:208-222)

each become ONE kernel whose loads go over NVLink (north-star paths 3 and 4). Results are written
into the second half of a double-buffered arena, so peers still reading the previous model are
never raced; buffers swap after a barrier.

Replica placement: replica k (server or node) lives on rank k mod world, so the engines also run
on a single GPU (all replicas co-located), which is how the GPU tests exercise them.
"""

from __future__ import annotations

import copy
import math

import torch
import torch.distributed as dist

from .arena import FlatModel
from .grouped import GroupedWorkers
from .peer import PeerArena, make_arena
from ..garfieldpp import tools
from ..garfieldpp.datasets import DatasetManager
from ..ops import attacks, gar as gar_ops
from ..ops.optim import FlatOptimizer

__all__ = ["DecentralizedEngine", "MSMWEngine"]


class _Base:
    def _init_dist(self, group, device, sync="flags", flag_timeout=20.0):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self._flag = torch.zeros(1, device=self.device)
        if not gar_ops.available():
            raise RuntimeError("the device engines need a CUDA device and the sm_100a library")
        # Phase boundaries between ranks ("all gradients written", "all models published", ...) are device-side: every rank
        # bumps a device-resident counter, releases its peer-mapped flag with it and polls the others' flags (one-warp kernel
        # with a deadline). No NCCL call inside a step; sync="nccl" keeps the 4-byte all-reduce as a baseline.
        self.sync = sync if self.world > 1 else "none"
        self._flag_arena = None
        if self.sync == "flags":
            self._flag_arena = PeerArena(64, self.device, group)
            self._seq = torch.zeros(1, dtype=torch.int32, device=self.device)
            self._late = torch.zeros(1, dtype=torch.int32, device=self.device)
            self._flag_timeout_ns = int(flag_timeout * 1e9)
            self._flags = [self._flag_arena.peer(r).data_ptr() for r in range(self.world)]

    def _barrier(self):
        if self.world <= 1:
            return
        if self.sync != "flags":
            dist.all_reduce(self._flag, group=self.group)
            return
        import ctypes
        from ..ops import _lib
        lib, st = _lib.require(), _lib.stream_ptr(self.device)
        _lib.check(lib.gb_epoch_bump(self._seq.data_ptr(), st))
        _lib.check(lib.gb_flag_signal_ep(ctypes.c_void_p(self._flags[self.rank]), self._seq.data_ptr(), st))
        table = (ctypes.c_void_p * self.world)(*self._flags)
        _lib.check(lib.gb_flag_wait_ep(table, self.world, self._seq.data_ptr(), 0, self._flag_timeout_ns, self._late.data_ptr(), None, st))
        _lib.count(3)

    @property
    def late_peers(self):
        return int(self._late.item()) if self.sync == "flags" else 0

    def _close_sync(self):
        if self._flag_arena is not None:
            self._flag_arena.close()
            self._flag_arena = None

    @staticmethod
    def _attack_row(kind, row, honest_rows):
        if kind == "random":
            row.copy_(attacks.random_like(row))
        elif kind == "reverse":
            row.copy_(attacks.reverse(row))
        elif kind == "drop":
            row.copy_(attacks.drop(row))
        elif kind == "lie":
            row.copy_(attacks.lie(honest_rows or [row]))
        elif kind == "empire":
            row.copy_(attacks.empire(honest_rows or [row]))
        else:
            raise ValueError("unknown attack %r" % (kind,))


class DecentralizedEngine(_Base):
    """LEARN: n nodes, each a worker + a server with ITS OWN model replica.

    Per step and node: gradient on the node's own model → every node aggregates ALL nodes' gradients
    (fused peer gather + GAR) → optional ⌈log2(i+1)⌉ agreement rounds on the aggregated gradients
    (`non_iid`) → SGD step → contract: every node replaces its model by the GAR of all nodes' models.
    Each worker computes ONE gradient per step (on its own node's model) that all nodes pull; the
    reference instead lets every server request a fresh gradient from every worker (n² backward
    passes) — equivalent when replicas agree, which the contract step enforces.
    """

    def __init__(self, model, dataset, loss, num_nodes, f, gar, batch, optimizer="sgd", opt_args=None, byzantine=None,
                 non_iid=False, device=None, group=None, seed=1234, fuse_workers=None, use_graphs=True, sync="flags"):
        self._init_dist(group, device, sync)
        self.n, self.f, self.gar_name, self.non_iid = num_nodes, f, gar, non_iid
        self.byzantine = dict(byzantine or {})
        torch.manual_seed(seed)
        module = tools.select_model(model, self.device, dataset)
        loss_fn = tools.select_loss(loss)
        # Worker fusion: the contract step leaves every node with the SAME model (each node runs the same deterministic
        # kernel over the same row table), so a rank computes the gradients of all its nodes in one pass on a copy of
        # its first node's live model (per-node BatchNorm statistics, weight gradients and losses).
        self.fuse = (GroupedWorkers.unsupported_reason(module, loss_fn) is None) if fuse_workers is None else bool(fuse_workers)
        self.channels_last = self.fuse
        if self.fuse:
            from ..models.fused_resnet import fuse_resnet_blocks
            from .engine import _skip_bn_counters
            _skip_bn_counters(module)
            fuse_resnet_blocks(module)
            module = module.to(memory_format=torch.channels_last)
        self.d = sum(p.numel() for p in module.parameters())
        self.pitch = (self.d + 63) // 64 * 64
        self.local_ids = [k for k in range(self.n) if k % self.world == self.rank]
        slots = (self.n + self.world - 1) // self.world
        self.slots = slots
        # per node: 2 parameter buffers (double-buffered), 1 gradient row, 2 aggregated-gradient buffers
        # gar = average: the agreement rounds and the contract step are in-switch reductions (multimem.ld_reduce / multimem.st)
        # over a multicast address — the double-buffered arenas then live in symmetric memory
        want_mc = gar == "average" and self.world > 1 and self.n % self.world == 0 and self.sync == "flags"
        self.params = make_arena(slots * 2 * self.pitch, self.device, group, multicast=want_mc)
        self.grads = PeerArena(slots * self.pitch, self.device, group)
        self.aggrs = make_arena(slots * 2 * self.pitch, self.device, group, multicast=want_mc)
        self._nvls = bool(want_mc and self.params.multicast_ptr and self.aggrs.multicast_ptr)
        init = FlatModel(module, self.device, channels_last=self.channels_last)
        self.layout = init
        if self.world > 1:
            dist.broadcast(init.flat, src=0, group=group)
        self.cur = 0            # which parameter buffer is live
        self.nodes = []
        for j, k in enumerate(self.local_ids):
            live = self._param(self.rank, j, 0)
            live.copy_(init.flat)
            self._param(self.rank, j, 1).copy_(init.flat)
            replica = FlatModel(copy.deepcopy(module), self.device, flat=live, channels_last=self.channels_last)
            batches = DatasetManager(dataset, batch, self.n, self.n, k).get_train_set()
            if self.fuse:   # static buffers: skip a shorter tail batch
                batches = [(MSMWEngine._host(x), y.pin_memory()) for x, y in batches if x.shape[0] == batches[0][0].shape[0]]
            opt = FlatOptimizer(live, optimizer, **(opt_args or {}))
            self.nodes.append(dict(id=k, slot=j, replica=replica, batches=batches, opt=opt, loss_fn=loss_fn,
                                   attack=self.byzantine.get(k), loss=torch.zeros((), device=self.device)))
        # Agreement rounds and the contract step of per-coordinate rules run inside ONE persistent kernel per GPU
        # (csrc/cuda/gar_gossip.cu): rounds separated by a grid barrier + peer flags instead of kernel relaunches and barriers.
        self._gossip_ok = (gar in gar_ops.GOSSIP_RULES and self.n <= 16 and 1 <= len(self.nodes) <= 8
                           and (self.world == 1 or self.sync == "flags"))
        self._gseq = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._gcounter = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.fused, self._graph, self._eager_passes, self.use_graphs = None, None, 0, use_graphs
        if self.fuse and self.nodes:
            G = len(self.nodes)
            self._wmodel = torch.empty(self.d, device=self.device)          # static copy of the rank's live model for the graphed pass
            self._wmodel.copy_(init.flat)
            shared = FlatModel(copy.deepcopy(module), self.device, flat=self._wmodel, channels_last=True)
            self.fused = GroupedWorkers(shared, [self._grad(self.rank, node["slot"]) for node in self.nodes], loss_fn)
            x0, y0 = self.nodes[0]["batches"][0]
            self._gdata = torch.zeros((G * x0.shape[0],) + tuple(x0.shape[1:]), device=self.device)
            if self._gdata.dim() == 4:
                self._gdata = self._gdata.contiguous(memory_format=torch.channels_last)
            self._gtarget = torch.zeros((G * y0.shape[0],) + tuple(y0.shape[1:]), dtype=y0.dtype, device=self.device)
            self._glosses = torch.zeros(G, device=self.device)
            for g, node in enumerate(self.nodes):
                node["loss"] = self._glosses[g]
                modules = dict(node["replica"].module.named_modules())
                for name, (mean, var) in self.fused.running_tables.items():
                    modules[name]._buffers["running_mean"], modules[name]._buffers["running_var"] = mean[g], var[g]
        self.steps_done = 0

    def _fused_pass(self, iter_num):
        self._wmodel.copy_(self._param(self.rank, self.nodes[0]["slot"], self.cur))
        bx, by = self._gdata.shape[0] // len(self.nodes), self._gtarget.shape[0] // len(self.nodes)
        for g, node in enumerate(self.nodes):
            data, target = node["batches"][iter_num % len(node["batches"])]
            self._gdata[g * bx:(g + 1) * bx].copy_(data, non_blocking=True)
            self._gtarget[g * by:(g + 1) * by].copy_(target, non_blocking=True)
        if not self.use_graphs or self._eager_passes < 2:
            self.fused.forward_backward(self._gdata, self._gtarget, self._glosses)
            self._eager_passes += 1
            return
        if self._graph is None:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                self.fused.forward_backward(self._gdata, self._gtarget, self._glosses)
            self._graph = graph
        self._graph.replay()

    def _nvls_average(self, arena, src_parity):
        """All nodes' parity-`src_parity` vectors of `arena` → their average in every node's other-parity vector: this rank
        reduces its coordinate slice inside the switch and multicasts the result (call between two phase barriers)."""
        import ctypes
        from ..ops import _lib
        per = (self.pitch // 4 + self.world - 1) // self.world * 4
        lo, hi = min(self.pitch, self.rank * per), min(self.pitch, self.rank * per + per)
        mc = arena.multicast_ptr
        # slot s, parity p starts at (2 s + p)·pitch: consecutive slots of one parity are 2·pitch apart
        src, dst = mc + 4 * src_parity * self.pitch, mc + 4 * (1 - src_parity) * self.pitch
        _lib.check(_lib.require().gb_multimem_average(ctypes.c_void_p(src), ctypes.c_void_p(dst), self.slots, 2 * self.pitch, lo, hi,
                                                     1.0 / self.n, _lib.stream_ptr(self.device)))
        _lib.count()

    def _gossip(self, getter, rounds, first):
        """`rounds` rounds of the rule over every node's double-buffered vectors (`getter(rank, slot, parity)`) in one launch;
        the result of a round reading parity p lands in parity 1 − p of every local node."""
        rows0, rows1 = self._all(getter, 0), self._all(getter, 1)
        out0 = [getter(self.rank, node["slot"], 1) for node in self.nodes]
        out1 = [getter(self.rank, node["slot"], 0) for node in self.nodes]
        my_flag, peers = None, []
        if self.world > 1:
            my_flag = self._flags[self.rank] + 4                   # word 1 of the flag arena (word 0 = phase barrier)
            peers = [self._flags[r] + 4 for r in range(self.world) if r != self.rank]
        gar_ops.gossip_rounds(self.gar_name, rows0, rows1, out0, out1, self.f, rounds, first, self._gseq, self._gcounter, my_flag, peers,
                              timeout_ns=getattr(self, "_flag_timeout_ns", 20_000_000_000), late=getattr(self, "_late", None))

    # arena views -----------------------------------------------------------------------------
    def _param(self, rank, slot, which):
        base = (slot * 2 + which) * self.pitch
        return self.params.peer(rank)[base:base + self.d]

    def _grad(self, rank, slot):
        return self.grads.peer(rank)[slot * self.pitch:slot * self.pitch + self.d]

    def _aggr(self, rank, slot, which):
        base = (slot * 2 + which) * self.pitch
        return self.aggrs.peer(rank)[base:base + self.d]

    def _all(self, getter, *extra):
        return [getter(k % self.world, k // self.world, *extra) for k in range(self.n)]

    def model_of(self, node_index):
        """Live flat model of local node `node_index` (position in local_ids)."""
        return self._param(self.rank, self.nodes[node_index]["slot"], self.cur)

    # step ------------------------------------------------------------------------------------
    def step(self, iter_num):
        n, f = self.n, self.f
        # 1. local gradients (on each node's own live model)
        if self.fused is not None:
            self._fused_pass(iter_num)
        for node in ([] if self.fused is not None else self.nodes):
            row = self._grad(self.rank, node["slot"])
            replica = node["replica"]
            row.zero_()
            replica.bind_grads(row)
            data, target = node["batches"][iter_num % len(node["batches"])]
            replica.module.train()
            loss = node["loss_fn"](replica.module(data.to(self.device, non_blocking=True)), target.to(self.device, non_blocking=True))
            loss.backward()
            replica.unbind_grads()
            node["loss"].copy_(loss.detach())
        for node in self.nodes:
            if node["attack"]:
                honest = [self._grad(self.rank, o["slot"]) for o in self.nodes if not o["attack"]]
                self._attack_row(node["attack"], self._grad(self.rank, node["slot"]), honest)
        self._barrier()
        # 2. every node aggregates all nodes' gradients (peer pointers → fused gather + GAR)
        rounds = int(math.ceil(math.log2(iter_num + 1))) if (self.non_iid and iter_num > 0) else 0
        rows = self._all(self._grad)
        which = 0
        fused_sgd = rounds == 0 and all(node["opt"].name == "sgd" for node in self.nodes)
        for node in self.nodes:
            if fused_sgd:    # no agreement rounds: the SGD step rides in the epilogue of the aggregation pass
                gar_ops.aggregate(self.gar_name, rows, f, update=node["opt"].sgd_fusion(), want_out=False)
                node["opt"].note_fused_step()
            else:
                gar_ops.aggregate(self.gar_name, rows, f, out=self._aggr(self.rank, node["slot"], which))
        # 3. agreement rounds on the aggregated gradients (double-buffered): one persistent kernel for all rounds, or — rules
        #    that need a selection step per round — one launch per round and node
        if rounds and self._nvls:
            for _ in range(rounds):
                self._barrier()
                self._nvls_average(self.aggrs, which)
                which = 1 - which
            self._barrier()          # the other ranks' slices of the last round have landed in this rank's buffers
            rounds = 0
        if rounds and self._gossip_ok and self.nodes:
            self._barrier()
            self._gossip(self._aggr, rounds, which)
            which = (which + rounds) & 1
            rounds = 0
        for _ in range(rounds):
            self._barrier()
            peers = self._all(self._aggr, which)
            for node in self.nodes:
                gar_ops.aggregate(self.gar_name, peers, f, out=self._aggr(self.rank, node["slot"], 1 - which))
            which = 1 - which
        # 4. optimizer step on each node's live model
        for node in ([] if fused_sgd else self.nodes):
            node["opt"].step(self._aggr(self.rank, node["slot"], which))
        self._barrier()
        # 5. contract: model := GAR(all nodes' live models), written into the other buffer (one pass for all local nodes)
        if self._nvls:
            self._nvls_average(self.params, self.cur)
        elif self._gossip_ok and self.nodes:
            self._gossip(self._param, 1, self.cur)
        else:
            models = self._all(self._param, self.cur)
            for node in self.nodes:
                gar_ops.aggregate(self.gar_name, models, f, out=self._param(self.rank, node["slot"], 1 - self.cur))
        self._barrier()
        self.cur = 1 - self.cur
        for node in self.nodes:  # re-alias replicas and optimizers onto the now-live buffer
            live = self._param(self.rank, node["slot"], self.cur)
            if self.fused is None:
                node["replica"] = FlatModel(node["replica"].module, self.device, flat=live)
            node["opt"].rebind(live)
        self.steps_done += 1
        return torch.stack([node["loss"] for node in self.nodes]).mean() if self.nodes else torch.zeros((), device=self.device)

    def close(self):
        torch.cuda.synchronize(self.device)
        self._close_sync()
        for arena in (self.aggrs, self.grads, self.params):
            arena.close()


class MSMWEngine(_Base):
    """ByzSGD: `num_ps` server replicas (possibly Byzantine) + `num_workers` workers.

    Per step: every worker first pulls the live models of ALL servers and takes their
    coordinate-wise median (fused peer gather + median straight into the worker's model — path 3),
    computes its gradient on that model; every server aggregates all workers' gradients with the
    gradient GAR and steps; then every server replaces its model by the model GAR of all servers'
    models (double-buffered). Byzantine servers publish an attacked model instead of their own.

    Worker fusion (default when the model is covered, `parallel/grouped.py`): every worker of a rank sees the same
    published models, so the rank takes their median ONCE into a shared channels-last model and computes all its
    workers' gradients in one CUDA-graphed pass (per-worker BatchNorm statistics, weight gradients and losses).
    """

    def __init__(self, model, dataset, loss, num_workers, num_ps, fw, fps, gar, batch, mar="median", optimizer="sgd", opt_args=None,
                 byzantine_workers=None, byzantine_servers=None, device=None, group=None, seed=1234, fuse_workers=None,
                 use_graphs=True, sync="flags"):
        self._init_dist(group, device, sync)
        self.n, self.num_ps, self.fw, self.fps = num_workers, num_ps, fw, fps
        self.gar_name, self.mar_name = gar, mar
        self.byz_w, self.byz_s = dict(byzantine_workers or {}), dict(byzantine_servers or {})
        torch.manual_seed(seed)
        module = tools.select_model(model, self.device, dataset)
        loss_fn = tools.select_loss(loss)
        self.fuse = (GroupedWorkers.unsupported_reason(module, loss_fn) is None) if fuse_workers is None else bool(fuse_workers)
        self.channels_last = self.fuse          # every arena of this engine then stores 4-D parameters in OHWI order
        if self.fuse:
            from ..models.fused_resnet import fuse_resnet_blocks
            from .engine import _skip_bn_counters
            _skip_bn_counters(module)
            fuse_resnet_blocks(module)
            module = module.to(memory_format=torch.channels_last)
        self.d = sum(p.numel() for p in module.parameters())
        self.pitch = (self.d + 63) // 64 * 64
        self.local_workers = [w for w in range(self.n) if w % self.world == self.rank]
        self.local_servers = [s for s in range(num_ps) if s % self.world == self.rank]
        w_slots = (self.n + self.world - 1) // self.world
        s_slots = (num_ps + self.world - 1) // self.world
        self.grads = PeerArena(w_slots * self.pitch, self.device, group)
        self.wmodels = PeerArena(max(1, w_slots) * self.pitch, self.device, group)      # workers' private models (local use only)
        self.smodels = PeerArena(s_slots * 3 * self.pitch, self.device, group)          # per server: 2 model buffers + 1 published copy
        init = FlatModel(module, self.device, channels_last=self.channels_last)
        self.layout = init                       # reference-order <-> arena-order conversions (`to_reference_order`)
        if self.world > 1:
            dist.broadcast(init.flat, src=0, group=group)
        self.cur = 0
        self.workers, self.servers = [], []
        for j, w in enumerate(self.local_workers):
            mine = self.wmodels.local[(0 if self.fuse else j) * self.pitch:(0 if self.fuse else j) * self.pitch + self.d]
            mine.copy_(init.flat)
            replica = FlatModel(copy.deepcopy(module), self.device, flat=mine, channels_last=self.channels_last)
            batches = DatasetManager(dataset, batch, self.n, self.n + num_ps, w + num_ps).get_train_set()
            if self.fuse:   # static buffers: skip a shorter tail batch
                batches = [(self._host(x), y.pin_memory()) for x, y in batches if x.shape[0] == batches[0][0].shape[0]]
            self.workers.append(dict(id=w, slot=j, replica=replica, batches=batches, loss_fn=loss_fn, attack=self.byz_w.get(w),
                                     loss=torch.zeros((), device=self.device)))
        self.fused, self._graph, self._eager_passes, self.use_graphs = None, None, 0, use_graphs
        if self.fuse and self.workers:
            G = len(self.workers)
            shared = FlatModel(copy.deepcopy(module), self.device, flat=self.wmodels.local[:self.d], channels_last=True)
            self.fused = GroupedWorkers(shared, [self._grad(self.rank, wrk["slot"]) for wrk in self.workers], loss_fn)
            x0, y0 = self.workers[0]["batches"][0]
            self._gdata = torch.zeros((G * x0.shape[0],) + tuple(x0.shape[1:]), device=self.device)
            if self._gdata.dim() == 4:
                self._gdata = self._gdata.contiguous(memory_format=torch.channels_last)
            self._gtarget = torch.zeros((G * y0.shape[0],) + tuple(y0.shape[1:]), dtype=y0.dtype, device=self.device)
            self._glosses = torch.zeros(G, device=self.device)
            for g, wrk in enumerate(self.workers):
                wrk["loss"] = self._glosses[g]
                modules = dict(wrk["replica"].module.named_modules())
                for name, (mean, var) in self.fused.running_tables.items():
                    modules[name]._buffers["running_mean"], modules[name]._buffers["running_var"] = mean[g], var[g]
        self._setup_pipelined_pull()
        for j, s in enumerate(self.local_servers):
            for which in range(3):
                self._smodel(self.rank, j, which).copy_(init.flat)
            opt = FlatOptimizer(self._smodel(self.rank, j, 0), optimizer, **(opt_args or {}))
            self.servers.append(dict(id=s, slot=j, opt=opt, attack=self.byz_s.get(s)))
        self._barrier()

    @staticmethod
    def _host(x):
        x = x.contiguous(memory_format=torch.channels_last) if x.dim() == 4 else x.contiguous()
        return x.pin_memory()

    def _setup_pipelined_pull(self):
        """Path 3: the model pull is cut into coordinate chunks along the top-level children of the network (forward order =
        parameter order). Chunk k's median kernel runs on a pull stream and records an event that child k's forward waits for,
        so the pull of the deep layers crosses NVLink while the first layers already compute. When the first layer is a
        convolution that needs an explicit zero-padded weight operand (the RGB stem), the head kernel of the pull emits that
        GEMM operand directly (gb_pull_median_operand)."""
        self._pull_chunks, self._pull_events, self._pull_stream, self._head = [], [], None, None
        if self.fused is None or self.num_ps < 2:
            return
        module = self.fused.module
        offsets = {id(p): (off, n) for p, (off, n, _) in zip(module.parameters(), self.fused.replica.slices)}
        covered = 0
        children = []
        for child in module.children():
            own = [offsets[id(p)] for p in child.parameters()]
            if not own:
                continue
            lo, hi = min(o for o, _ in own), max(o + n for o, n in own)
            if lo != covered:
                return                      # parameters are not laid out child by child: keep the single-launch pull
            covered = hi
            children.append((child, lo, hi))
        if covered != self.d or len(children) < 2:
            return
        self._pull_stream = torch.cuda.Stream(device=self.device)
        first = children[0][0]
        if (isinstance(first, torch.nn.Conv2d) and first.bias is None and first.groups == 1
                and (first.in_channels * first.kernel_size[0] * first.kernel_size[1]) % 4 != 0):
            k = first.in_channels * first.kernel_size[0] * first.kernel_size[1]
            self._head = (first, first.out_channels, k, torch.zeros((first.out_channels, (k + 3) // 4 * 4), device=self.device))
        for index, (child, lo, hi) in enumerate(children):
            self._pull_chunks.append((lo, hi))
            event = torch.cuda.Event()
            self._pull_events.append(event)
            child.register_forward_pre_hook(lambda _m, _inp, ev=event: torch.cuda.current_stream(self.device).wait_event(ev))

    def _pull(self, published):
        """Worker model := coordinate-wise median of the published server models (fused gather + median over peer rows)."""
        shared = self.wmodels.local[:self.d]
        if self.num_ps < 2:
            gar_ops.aggregate("average", published[:1], 0, out=shared)
            return
        if not self._pull_chunks:
            gar_ops.coordwise(published, "median_lower", out=shared)
            return
        main = torch.cuda.current_stream(self.device)
        self._pull_stream.wait_stream(main)
        with torch.cuda.stream(self._pull_stream):
            for index, ((lo, hi), event) in enumerate(zip(self._pull_chunks, self._pull_events)):
                rows = [p[lo:hi] for p in published]
                if index == 0 and self._head is not None:
                    conv, cout, k, operand = self._head
                    gar_ops.pull_median_operand(rows, cout, k, shared[lo:hi], operand)
                    conv._gb_w2 = operand
                else:
                    gar_ops.coordwise(rows, "median_lower", out=shared[lo:hi])
                event.record(self._pull_stream)

    def _fused_pass(self, iter_num, published):
        """Median of the published models once per rank (pipelined with the forward pass), then all local workers' gradients
        in one pass."""
        bx, by = self._gdata.shape[0] // len(self.workers), self._gtarget.shape[0] // len(self.workers)
        for g, wrk in enumerate(self.workers):
            data, target = wrk["batches"][iter_num % len(wrk["batches"])]
            self._gdata[g * bx:(g + 1) * bx].copy_(data, non_blocking=True)
            self._gtarget[g * by:(g + 1) * by].copy_(target, non_blocking=True)
        def body():
            self._pull(published)
            self.fused.forward_backward(self._gdata, self._gtarget, self._glosses)
            if self._pull_stream is not None:
                torch.cuda.current_stream(self.device).wait_stream(self._pull_stream)
        if not self.use_graphs or self._eager_passes < 2:
            body()
            self._eager_passes += 1
            return
        if self._graph is None:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                body()
            self._graph = graph
        self._graph.replay()

    def _smodel(self, rank, slot, which):
        base = (slot * 3 + which) * self.pitch
        return self.smodels.peer(rank)[base:base + self.d]

    def _published(self):
        """What every server shows to its peers (slot 2): its live model, or an attack."""
        return [self._smodel(s % self.world, s // self.world, 2) for s in range(self.num_ps)]

    def _grad(self, rank, slot):
        return self.grads.peer(rank)[slot * self.pitch:slot * self.pitch + self.d]

    def server_model(self, index=0):
        return self._smodel(self.rank, self.servers[index]["slot"], self.cur)

    def step(self, iter_num):
        # 0. servers publish (honest copy or attack) — previous step's barrier ordered the writes
        for srv in self.servers:
            live, shown = self._smodel(self.rank, srv["slot"], self.cur), self._smodel(self.rank, srv["slot"], 2)
            shown.copy_(live)
            if srv["attack"]:
                self._attack_row(srv["attack"], shown, [live])
        self._barrier()
        published = self._published()
        # 1. workers: model := median of all servers' published models (fused pull), then gradient
        if self.fused is not None:
            self._fused_pass(iter_num, published)
        for wrk in ([] if self.fused is not None else self.workers):
            mine = self.wmodels.local[wrk["slot"] * self.pitch:wrk["slot"] * self.pitch + self.d]
            if self.num_ps > 1:
                gar_ops.coordwise(published, "median_lower", out=mine)
            else:
                gar_ops.aggregate("average", published[:1], 0, out=mine)
            row = self._grad(self.rank, wrk["slot"])
            row.zero_()
            wrk["replica"].bind_grads(row)
            data, target = wrk["batches"][iter_num % len(wrk["batches"])]
            wrk["replica"].module.train()
            loss = wrk["loss_fn"](wrk["replica"].module(data.to(self.device, non_blocking=True)), target.to(self.device, non_blocking=True))
            loss.backward()
            wrk["replica"].unbind_grads()
            wrk["loss"].copy_(loss.detach())
        for wrk in self.workers:
            if wrk["attack"]:
                honest = [self._grad(self.rank, o["slot"]) for o in self.workers if not o["attack"]]
                self._attack_row(wrk["attack"], self._grad(self.rank, wrk["slot"]), honest)
        self._barrier()
        # 2. servers: aggregate all workers' gradients (peer rows) + fused SGD on the live model
        rows = [self._grad(w % self.world, w // self.world) for w in range(self.n)]
        for srv in self.servers:
            opt = srv["opt"]
            if opt.name == "sgd":
                gar_ops.aggregate(self.gar_name, rows, self.fw, update=opt.sgd_fusion(), want_out=False)
                opt.note_fused_step()
            else:
                opt.step(gar_ops.aggregate(self.gar_name, rows, self.fw))
            self._smodel(self.rank, srv["slot"], 2).copy_(self._smodel(self.rank, srv["slot"], self.cur))
            if srv["attack"]:
                self._attack_row(srv["attack"], self._smodel(self.rank, srv["slot"], 2), [self._smodel(self.rank, srv["slot"], self.cur)])
        self._barrier()
        # 3. servers agree: model := model-GAR of all servers' published models → other buffer
        if self.num_ps > 1:
            published = self._published()
            for srv in self.servers:
                nxt = self._smodel(self.rank, srv["slot"], 1 - self.cur)
                if self.mar_name == "vanilla":
                    gar_ops.aggregate("average", published, 0, out=nxt)
                elif self.mar_name == "crash":
                    gar_ops.aggregate("average", published[:1], 0, out=nxt)
                else:
                    gar_ops.aggregate(self.mar_name, published, self.fps, out=nxt)
            self._barrier()
            self.cur = 1 - self.cur
            for srv in self.servers:
                live = self._smodel(self.rank, srv["slot"], self.cur)
                srv["opt"].rebind(live)
        losses = [w["loss"] for w in self.workers if not w["attack"]]
        return torch.stack(losses).mean() if losses else torch.zeros((), device=self.device)

    def close(self):
        torch.cuda.synchronize(self.device)
        self._close_sync()
        for arena in (self.smodels, self.wmodels, self.grads):
            arena.close()
