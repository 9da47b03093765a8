"""Device-resident SSMW (AggregaThor) step engine: one process per GPU, n workers + 1 server.

What one `step()` does (reference loop: applications/Aggregathor/trainer.py:231-237 +
garfieldpp/server.py:112-159,277-287 + worker.py:77-96):

    S1/S2  every local worker runs forward+backward on ITS replica of the model (same flat
           parameter arena, own BatchNorm buffers) and writes its gradient straight into its
           row of the rank's peer-mapped gradient arena — all local workers are captured in ONE
           CUDA graph with one stream per worker so small kernels of different workers overlap;
    S3     Byzantine workers overwrite their row with an on-device attack;
    S4/S5  rank 0 (the server) runs the aggregation kernels over a row table that mixes local
           rows and peer-mapped rows of the other GPUs (NVLink P2P loads / TMA bulk copies from
           inside the kernel — no NCCL gather, no host copy);
    S6/S1' the SGD step is fused in the aggregation epilogue, and the same kernel stores the
           updated parameters into every rank's parameter arena (model publication).

Ranks synchronise twice per step (gradients ready / parameters published) with a 4-byte NCCL
all-reduce used purely as a barrier; everything else is device-to-device.
"""

from __future__ import annotations

import copy

import torch
import torch.distributed as dist

from .arena import FlatModel
from .peer import DeviceBuffer, PeerArena, make_arena
from ..garfieldpp import tools
from ..garfieldpp.datasets import DatasetManager
from ..ops import _lib, attacks, gar as gar_ops
from ..ops.optim import FlatOptimizer
from .grouped import GroupedWorkers, ScatterSink

__all__ = ["SSMWEngine"]


def _skip_bn_counters(module):
    """BatchNorm's `num_batches_tracked += 1` is a separate int64 kernel per layer and step (53 launches per
    ResNet-50 pass) that only matters when momentum is None; keep the running statistics, drop the counter."""
    import types

    def forward(self, x):
        return torch.nn.functional.batch_norm(x, self.running_mean, self.running_var, self.weight, self.bias,
                                              self.training or self.running_mean is None, self.momentum, self.eps)

    for m in module.modules():
        if isinstance(m, torch.nn.modules.batchnorm._BatchNorm) and m.momentum is not None and m.track_running_stats:
            m.forward = types.MethodType(forward, m)


class _SplitConv(torch.autograd.Function):
    """conv2d whose weight-gradient kernel runs on a side stream, off the data-gradient chain that the rest
    of the backward pass waits on (≈10 % shorter ResNet-50 step; the results are identical)."""

    @staticmethod
    def forward(ctx, x, w, b, stride, padding, dilation, groups, side):
        ctx.save_for_backward(x, w)
        ctx.cfg = (stride, padding, dilation, groups, side, None if b is None else list(b.shape))
        return torch.nn.functional.conv2d(x, w, b, stride, padding, dilation, groups)

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        stride, padding, dilation, groups, side, bias_sizes = ctx.cfg
        aten = torch.ops.aten
        gx = gw = gb = None
        if ctx.needs_input_grad[1]:
            side.wait_stream(torch.cuda.current_stream())
            gy.record_stream(side)
            x.record_stream(side)
            with torch.cuda.stream(side):
                _, gw, gb = aten.convolution_backward(gy, x, w, bias_sizes, stride, padding, dilation, False, [0, 0], groups,
                                                      [False, True, bias_sizes is not None])
        if ctx.needs_input_grad[0]:
            gx = aten.convolution_backward(gy, x, w, None, stride, padding, dilation, False, [0, 0], groups, [True, False, False])[0]
        return gx, gw, gb, None, None, None, None, None


def _split_conv_backward(module, side):
    import types

    def forward(self, x):
        return _SplitConv.apply(x, self.weight, self.bias, self.stride, self.padding, self.dilation, self.groups, side)

    for m in module.modules():
        if isinstance(m, torch.nn.Conv2d) and m.padding_mode == "zeros":
            m.forward = types.MethodType(forward, m)


class _LocalWorker:
    """One worker hosted by this rank: model replica, static input buffers, gradient row."""

    def __init__(self, wid, replica, row, loss_fn, batch_shape, target_shape, target_dtype, device, attack=None, channels_last=False,
                 gather=True, split_backward=True):
        self.wid, self.replica, self.row, self.loss_fn, self.attack = wid, replica, row, loss_fn, attack
        self.data = torch.zeros(batch_shape, device=device)
        if channels_last and len(batch_shape) == 4:
            self.data = self.data.contiguous(memory_format=torch.channels_last)
        self.target = torch.zeros(target_shape, dtype=target_dtype, device=device)
        self.loss = torch.zeros((), device=device)
        self.stream = torch.cuda.Stream(device=device) if device.type == "cuda" else None
        self.side = None
        if split_backward and gather and device.type == "cuda":
            self.side = torch.cuda.Stream(device=device)
            _split_conv_backward(replica.module, self.side)
        self.gather = gather
        if not gather:
            replica.bind_grads(row)  # autograd accumulates in place into the arena row (one add kernel per parameter)

    def forward_backward(self):
        module = self.replica.module
        module.train()
        if not self.gather:
            self.row.zero_()
            loss = self.loss_fn(module(self.data), self.target)
            loss.backward()
            self.loss.copy_(loss.detach())
            return
        # let autograd hand over its freshly produced gradient tensors, then gather all of them into the
        # row with ONE multi-tensor copy launch (the table lives in the kernel parameters → graph-safe)
        params = list(module.parameters())
        for p in params:
            p.grad = None
        loss = self.loss_fn(module(self.data), self.target)
        loss.backward()
        self.loss.copy_(loss.detach())
        if self.side is not None:
            torch.cuda.current_stream(self.row.device).wait_stream(self.side)   # weight gradients are complete
        self._gather(params)
        for p in params:
            p.grad = None

    def _gather(self, params):
        import ctypes
        srcs, offs, sizes, slow = [], [], [], []
        for p, (off, n, shape) in zip(params, self.replica.slices):
            g = p.grad
            if g is None:
                self.row[off:off + n].zero_()
            elif g.dtype == torch.float32 and g.stride() == p.stride() and g.is_cuda:
                srcs.append(g.data_ptr()); offs.append(off); sizes.append(n)
            else:
                slow.append((off, n, shape, g))
        if srcs:
            k = len(srcs)
            lib = _lib.require()
            _lib.check(lib.gb_multi_copy((ctypes.c_void_p * k)(*srcs), (ctypes.c_uint64 * k)(*offs), (ctypes.c_uint32 * k)(*sizes), k,
                                         self.row.data_ptr(), _lib.stream_ptr(self.row.device)))
            _lib.count((k + 111) // 112)
        for off, n, shape, g in slow:
            self.replica._view(self.row, off, n, shape).copy_(g)


class SSMWEngine:
    def __init__(self, model, dataset, loss, num_workers, f, gar, batch, optimizer="sgd", opt_args=None, byzantine=None,
                 device=None, group=None, use_graphs=True, train_set=None, seed=1234, gar_kwargs=None, sharded=None,
                 channels_last=True, sync=None, flag_timeout=5.0, gather_grads=True, fast_bn=True,
                 split_backward=None, fuse_workers=None, fusion_chains=1, scatter_grads=None, overlap_distances=None, num_wait=None):
        """`sync`: "flags" (default with several GPUs) — ranks synchronise through peer-mapped epoch flags and exchange the
        partial distance matrices through peer-visible slots, so that no NCCL call remains in a step (NCCL only bootstraps:
        IPC handles, initial broadcast); "nccl" keeps the 4-byte all-reduce barriers and the n×n all-reduce (baseline).
        `num_wait`: quorum — aggregate as soon as this many gradients are there and the deadline `flag_timeout` has passed;
        the rule is then given only the rows that made it (reference: server.get_gradients(iter, n − f))."""
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.group = group
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.n, self.f, self.gar_name, self.batch = num_workers, f, gar, batch
        self.gar_kwargs = dict(gar_kwargs or {})
        self.byzantine = dict(byzantine or {})
        # sharded=True: the (trusted) server's aggregation, optimizer state and publication are split by
        # coordinate range over all ranks — same result, balanced NVLink traffic.
        self.sharded = (self.world > 1) if sharded is None else (bool(sharded) and self.world > 1)
        self.is_server = self.rank == 0 or self.sharded
        if self.device.type != "cuda" or not gar_ops.available():
            raise RuntimeError("SSMWEngine needs a CUDA device and the sm_100a library")

        torch.manual_seed(seed)
        module = tools.select_model(model, self.device, dataset)
        if fast_bn:
            _skip_bn_counters(module)
            if channels_last:
                from ..models.fused_resnet import fuse_resnet_blocks
                fuse_resnet_blocks(module)   # BN + ReLU (+ residual add) in one hand-written kernel per block stage
        self.channels_last = bool(channels_last)
        if self.channels_last:
            module = module.to(memory_format=torch.channels_last)
        self.d = sum(p.numel() for p in module.parameters())
        self.pitch = (self.d + 63) // 64 * 64
        self.local_ids = [w for w in range(self.n) if w % self.world == self.rank]
        if split_backward is None:
            # weight gradients on a side stream shorten ONE worker's chain by ≈10 %, but with many co-located
            # workers the GPU is already saturated by their streams and the extra ones only add contention
            split_backward = len(self.local_ids) <= 2
        self.max_local = (self.n + self.world - 1) // self.world
        loss_fn = tools.select_loss(loss)
        if fuse_workers is None:     # also pays for a single hosted worker (in-place weight-gradient GEMMs, no gather copy)
            fuse_workers = gather_grads
        fuse_workers = bool(fuse_workers) and GroupedWorkers.unsupported_reason(module, loss_fn) is None
        # scatter: with the coordinate-sharded server, every worker's gradient is written slice by slice straight into
        # the arena of the rank that aggregates that range (a reduce-scatter by construction, overlapped with the
        # backward pass); the aggregation kernels then read local HBM only.
        self.per = (self.pitch // 64 + self.world - 1) // self.world * 64      # coordinates per rank (multiple of 64 floats)
        self.scatter = fuse_workers and self.sharded and (scatter_grads is None or bool(scatter_grads))

        # peer-mapped arenas: parameters (one per rank) and gradient rows (max_local full rows per rank, or — scatter —
        # one slice-long row per worker of the whole job)
        import os
        # parameters are published to every GPU through an NVLS multicast address when the platform has one (multimem.st in the
        # aggregation epilogue: one store per 16 bytes instead of one per peer); GARFIELD_MULTICAST=0 keeps per-peer stores
        self.param_arena = make_arena(self.pitch, self.device, group, multicast=os.environ.get("GARFIELD_MULTICAST", "1") != "0")
        self.grad_arena = PeerArena(self.world * self.max_local * self.per if self.scatter else self.max_local * self.pitch, self.device, group)
        self.flat = self.param_arena.local[:self.d]
        self.master = FlatModel(module, self.device, flat=None, channels_last=self.channels_last)
        if self.world > 1:  # everyone starts from rank 0's initialisation
            dist.broadcast(self.master.flat, src=0, group=group)
            for buf in self.master.module.buffers():
                dist.broadcast(buf, src=0, group=group)
        self.flat.copy_(self.master.flat)
        self.master = FlatModel(module, self.device, flat=self.flat, channels_last=self.channels_last)  # re-alias onto the arena

        self.train_sets = {}
        self.workers = []
        for j, wid in enumerate(self.local_ids):
            replica = FlatModel(copy.deepcopy(module), self.device, flat=self.flat, channels_last=self.channels_last)
            row = None if self.scatter else self.grad_arena.local[j * self.pitch:j * self.pitch + self.d]
            if train_set is not None:
                batches = train_set(wid)
            else:
                batches = DatasetManager(dataset, batch, self.n, self.n + 1, wid + 1).get_train_set()
            # static device buffers / CUDA graphs need one batch shape: a shorter tail batch of the partition is skipped
            batches = [b for b in batches if b[0].shape[0] == batches[0][0].shape[0]]
            if self.channels_last and batches[0][0].dim() == 4:  # host batches in NHWC so the H2D copy is a plain memcpy
                batches = [(self._pin(x.contiguous(memory_format=torch.channels_last)), y) for x, y in batches]
            self.train_sets[wid] = batches
            x0, y0 = batches[0]
            self.workers.append(_LocalWorker(wid, replica, row, loss_fn, tuple(x0.shape), tuple(y0.shape), y0.dtype, self.device,
                                             attack=self.byzantine.get(wid), channels_last=self.channels_last, gather=gather_grads,
                                             split_backward=split_backward))

        # worker fusion: the co-located workers share ONE forward/backward pass (parallel/grouped.py) when every
        # layer of the model is covered; otherwise each worker keeps its own chain and stream
        self.fused = None
        if fuse_workers and self.workers:
            self._fuse_workers(module, loss_fn, max(1, min(int(fusion_chains), len(self.workers))))

        # distances during the backward pass (one GPU, distance-based rule, honest local workers): squared distances are sums
        # over coordinates and the LAST layers' gradients are complete early in the backward pass, so the pairwise kernel
        # runs on that coordinate range on a side stream while the earlier layers are still being differentiated
        self._tail_at = None
        # opt-in: measured +0.6 % on the flagship (75.3 vs 74.8 k images/s) — the persistent pairwise kernel takes every SM
        # while it runs, so the backward pass it overlaps with stalls; it needs an SM-limited launch to pay off
        want_overlap = bool(overlap_distances) or (overlap_distances is None and os.environ.get("GARFIELD_OVERLAP_DISTANCES", "0") != "0")
        if (want_overlap and self.fused is not None and self.world == 1 and len(self._chains) == 1 and not self.byzantine
                and gar in gar_ops.distance_rules()):
            self._setup_distance_overlap()

        # server state (whole vector on rank 0, or this rank's coordinate slice when sharded)
        if self.sharded:
            self.lo, self.hi = self._bounds(self.rank)
        else:
            self.lo, self.hi = 0, self.d
        self.optimizer, self.rows = None, None
        if self.is_server and self.hi > self.lo:
            self.optimizer = FlatOptimizer(self.flat[self.lo:self.hi], optimizer, **(opt_args or {}))
            self.rows = []
            for wid in range(self.n):
                owner, j = wid % self.world, wid // self.world
                if self.scatter:    # this rank's slice of every worker's gradient is in its own arena
                    at = (owner * self.max_local + j) * self.per
                    self.rows.append(self.grad_arena.local[at:at + self.hi - self.lo])
                else:
                    self.rows.append(self.grad_arena.peer(owner)[j * self.pitch + self.lo:j * self.pitch + self.hi])
            self.mirrors = [self.param_arena.peer(r)[self.lo:self.hi] for r in range(self.world) if r != self.rank]
            self.multicast = (self.param_arena.multicast_ptr + 4 * self.lo) if self.param_arena.multicast_ptr else None
        self._flag = torch.zeros(1, device=self.device)
        # sync="flags": peer-mapped epoch flags polled by a one-warp kernel with a deadline instead of NCCL barriers, and
        # peer-visible slots + a fixed-order sum instead of the NCCL all-reduce of the partial distances. The epoch lives in
        # device memory (bumped by a kernel), so the whole multi-GPU iteration is one replayable CUDA graph.
        if sync is None:
            sync = "flags"
        self.sync = sync if self.world > 1 else "none"
        self.flag_timeout_ns = int(flag_timeout * 1e9)
        self.epoch = 0
        self.num_wait = int(num_wait) if (num_wait and num_wait < self.n and self.sync == "flags") else None
        self.last_arrived = None
        if self.sync == "flags":
            self.flag_arena = PeerArena(64, self.device, group)
            self.dist_arena = PeerArena((gar_ops.MAX_ROWS + 1) ** 2 + 63, self.device, group)
            self._late = torch.zeros(1, dtype=torch.int32, device=self.device)
            self._epoch_dev = torch.zeros(1, dtype=torch.int32, device=self.device)
            self._arrived = torch.zeros(64, dtype=torch.int32, device=self.device)
            word = lambda r, k: self.flag_arena.peer(r).data_ptr() + 4 * k      # flag k of rank r (uint32)
            self._word = word
            self._my_grad_flag, self._my_param_flag, self._my_dist_flag = word(self.rank, 0), word(self.rank, 1), word(self.rank, 2)
            self._grad_flags = [word(r, 0) for r in range(self.world)]
            servers = list(range(self.world)) if self.sharded else [0]
            self._servers = servers
            self._param_flags = [word(r, 1) for r in servers]
            self._dist_flags = [word(r, 2) for r in servers]
            self._reductions_this_step = 0
        self._graph = None
        self._whole_step = None          # None: try the whole-iteration graph on one GPU; False: it failed to capture
        self._copy_stream = None
        self.use_graphs = use_graphs
        self.steps_done = 0
        self.h2d_bytes = 0

    def _setup_distance_overlap(self):
        """Pick the top-level child whose parameters start the coordinate tail — the first one starting at >= 5 % of d, so
        that almost all distances leave the critical path while enough of the backward pass (the earlier children) is
        still to run when its gradients are complete — and hook its backward completion."""
        module = self.fused.module
        params = list(module.parameters())
        offsets = {id(p): off for p, (off, _, _) in zip(params, self.fused.replica.slices)}
        chosen = None
        for child in module.children():
            own = list(child.parameters())
            if own and 0.05 * self.d <= min(offsets[id(p)] for p in own) <= 0.9 * self.d:
                chosen = child
                break
        if chosen is None:
            return
        self._early_stream = torch.cuda.Stream(device=self.device)
        import os
        self._early_ctas = int(os.environ.get("GARFIELD_EARLY_DISTANCE_CTAS", "37"))
        start = self.fused.add_milestone(chosen, self._early_distances)
        self._tail_at = (start + 63) // 64 * 64                     # 256-byte aligned split of every row
        self._early_ws = None

    def _early_distances(self):
        """Backward-pass hook: the gradients of coordinates [tail_at, d) are in the rows (or in flight on the side
        streams) — compute their share of the pairwise distances now, off the critical path."""
        main = torch.cuda.current_stream(self.device)
        self._early_stream.wait_stream(main)
        self.fused.join_side_streams(self._early_stream)
        with torch.cuda.stream(self._early_stream):
            # a quarter of the SMs: the backward pass this hides behind keeps the rest
            self._early_ws = gar_ops.select([w.row[self._tail_at:] for w in self.workers], "none", workspace_tag=1, max_ctas=self._early_ctas)

    def _bounds(self, rank):
        lo = min(self.d, rank * self.per)
        return lo, min(self.d, lo + self.per)

    def _fuse_workers(self, module, loss_fn, chains=1):
        """Point every local worker's inputs, loss and BatchNorm running statistics at slices of the group's
        buffers, so loading batches, evaluation and checkpoints work on the worker objects as before.
        `chains` > 1 splits the workers into that many fused passes running concurrently on their own streams."""
        G = len(self.workers)
        first = self.workers[0]
        data = torch.zeros((G * first.data.shape[0],) + tuple(first.data.shape[1:]), device=self.device)
        if self.channels_last and data.dim() == 4:
            data = data.contiguous(memory_format=torch.channels_last)
        target = torch.zeros((G * first.target.shape[0],) + tuple(first.target.shape[1:]), dtype=first.target.dtype, device=self.device)
        self._group_data, self._group_target = data, target
        self._group_losses = torch.zeros(G, device=self.device)
        b, t = first.data.shape[0], first.target.shape[0]
        bounds = [G * c // chains for c in range(chains + 1)]
        self._chains = []
        for lo, hi in zip(bounds[:-1], bounds[1:]):
            replica = FlatModel(copy.deepcopy(module), self.device, flat=self.flat, channels_last=self.channels_last)
            sink = None
            if self.scatter:     # each chain scatters its own workers' rows (arena rows rank·max_local + lo ...)
                arenas = [self.grad_arena.peer(r) for r in range(self.world)]
                owned = [self._bounds(r) for r in range(self.world)]
                first_row, per, rank, count = self.rank * self.max_local + lo, self.per, self.rank, hi - lo
                sink = (lambda params, slices, view_of, first_row=first_row, count=count:
                        ScatterSink(params, slices, view_of, count, arenas, first_row, per, owned, rank))
            group = GroupedWorkers(replica, [w.row for w in self.workers[lo:hi]], loss_fn, sink=sink)
            stream = torch.cuda.Stream(device=self.device) if chains > 1 else None
            self._chains.append((group, data[lo * b:hi * b], target[lo * t:hi * t], self._group_losses[lo:hi], stream))
            for g, w in enumerate(self.workers[lo:hi]):
                w.data, w.target, w.loss = data[(lo + g) * b:(lo + g + 1) * b], target[(lo + g) * t:(lo + g + 1) * t], self._group_losses[lo + g]
                modules = dict(w.replica.module.named_modules())
                for name, (mean, var) in group.running_tables.items():
                    modules[name]._buffers["running_mean"], modules[name]._buffers["running_var"] = mean[g], var[g]
        self.fused = self._chains[0][0]
        # one pinned host tensor per step for the whole group: 2 H2D copies per step instead of 2 per worker
        counts = {len(self.train_sets[w.wid]) for w in self.workers}
        self._group_batches = None
        if len(counts) == 1 and len(self.workers) > 1:
            self._group_batches = []
            for k in range(counts.pop()):
                xs, ys = zip(*[self.train_sets[w.wid][k] for w in self.workers])
                x = torch.cat(xs)
                if self.channels_last and x.dim() == 4:
                    x = x.contiguous(memory_format=torch.channels_last)
                self._group_batches.append((self._pin(x), self._pin(torch.cat(ys))))

    # ------------------------------------------------------------------------------------------
    @staticmethod
    def _pin(tensor):
        return tensor.pin_memory() if torch.cuda.is_available() else tensor

    def _barrier(self):
        if self.world > 1:
            dist.all_reduce(self._flag, group=self.group)

    def load_batches(self, iter_num):
        """Host→device copy of this step's batch of every local worker (pinned source, async)."""
        nbytes = 0
        if getattr(self, "_group_batches", None):
            x, y = self._group_batches[iter_num % len(self._group_batches)]
            self._group_data.copy_(x, non_blocking=True)
            self._group_target.copy_(y, non_blocking=True)
            self.h2d_bytes = x.numel() * x.element_size() + y.numel() * y.element_size()
            return self.h2d_bytes
        for w in self.workers:
            batches = self.train_sets[w.wid]
            x, y = batches[iter_num % len(batches)]
            w.data.copy_(x, non_blocking=True)
            w.target.copy_(y, non_blocking=True)
            nbytes += x.numel() * x.element_size() + y.numel() * y.element_size()
        self.h2d_bytes = nbytes
        return nbytes

    # -- pipelined public loop: prefetch of step i+1's inputs and deferred loss read-back ---------
    def _prefetch(self, iter_num):
        """Start the pinned-host → device copy of step `iter_num`'s batches on the copy stream."""
        grouped = getattr(self, "_group_batches", None)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
            if grouped:
                self._stage_data, self._stage_target = torch.empty_like(self._group_data), torch.empty_like(self._group_target)
            for w in ([] if grouped else self.workers):
                w.stage_data, w.stage_target = torch.empty_like(w.data), torch.empty_like(w.target)
        nbytes = 0
        self._copy_stream.wait_stream(torch.cuda.current_stream(self.device))  # staging buffers are free again
        with torch.cuda.stream(self._copy_stream):
            if grouped:
                x, y = grouped[iter_num % len(grouped)]
                self._stage_data.copy_(x, non_blocking=True)
                self._stage_target.copy_(y, non_blocking=True)
                nbytes = x.numel() * x.element_size() + y.numel() * y.element_size()
            for w in ([] if grouped else self.workers):
                batches = self.train_sets[w.wid]
                x, y = batches[iter_num % len(batches)]
                w.stage_data.copy_(x, non_blocking=True)
                w.stage_target.copy_(y, non_blocking=True)
                nbytes += x.numel() * x.element_size() + y.numel() * y.element_size()
        self.h2d_bytes = nbytes

    def _commit_prefetch(self):
        """Make the prefetched batches the inputs of the next step (device-to-device, on the main stream)."""
        torch.cuda.current_stream(self.device).wait_stream(self._copy_stream)
        if getattr(self, "_group_batches", None):
            self._group_data.copy_(self._stage_data, non_blocking=True)
            self._group_target.copy_(self._stage_target, non_blocking=True)
            return
        for w in self.workers:
            w.data.copy_(w.stage_data, non_blocking=True)
            w.target.copy_(w.stage_target, non_blocking=True)

    def run(self, first_iter, steps):
        """Train `steps` iterations end to end and return the per-step mean losses (Python floats).

        Every step's batches are copied from pinned host memory while the previous step computes,
        and every step's loss is read back to the host one step later, so the host never stalls the
        device; all copies and reads happen inside this call."""
        losses, pending = [], None
        host = [torch.zeros((), pin_memory=True) for _ in range(2)]
        self._prefetch(first_iter)
        for i in range(steps):
            self._commit_prefetch()
            if i + 1 < steps:
                self._prefetch(first_iter + i + 1)
            loss = self.step(first_iter + i, load=False)
            slot = host[i % 2]
            slot.copy_(loss, non_blocking=True)
            done = torch.cuda.Event()
            done.record()
            if pending is not None:
                pending[0].synchronize()
                losses.append(float(pending[1]))
            pending = (done, slot)
        if pending is not None:
            pending[0].synchronize()
            losses.append(float(pending[1]))
        return losses

    def _run_workers(self):
        if self.fused is not None:
            if len(self._chains) == 1:
                group, data, target, losses, _ = self._chains[0]
                group.forward_backward(data, target, losses)
                if self._tail_at is not None:      # the early-distance branch rejoins here (also closes it inside a graph capture)
                    torch.cuda.current_stream(self.device).wait_stream(self._early_stream)
                return
            main = torch.cuda.current_stream(self.device)
            for group, data, target, losses, stream in self._chains:
                stream.wait_stream(main)
                with torch.cuda.stream(stream):
                    group.forward_backward(data, target, losses)
            for chain in self._chains:
                main.wait_stream(chain[4])
            return
        main = torch.cuda.current_stream(self.device)
        for w in self.workers:
            w.stream.wait_stream(main)
            with torch.cuda.stream(w.stream):
                w.forward_backward()
        for w in self.workers:
            main.wait_stream(w.stream)

    def _capture(self):
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(3):
                self._run_workers()
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        graph = torch.cuda.CUDAGraph()
        before = _lib.launches[0]
        with torch.cuda.graph(graph):
            self._run_workers()
        self._graph_launches = _lib.launches[0] - before   # this library's kernels inside the graph (counted per replay)
        self._graph = graph

    def compute_gradients(self):
        if not self.workers:
            return
        if self.use_graphs:
            if self._graph is None:
                self._capture()
            self._graph.replay()
            _lib.count(self._graph_launches)
        else:
            self._run_workers()
        for w in ([] if self.scatter else self.workers):
            if w.attack is not None:
                self._attack(w)

    def _attack_slices(self):
        """Scatter mode: every attack is coordinate-wise, so the rank that owns a coordinate range applies it to its
        slice of the Byzantine workers' rows (colluding attacks see the honest workers' slices)."""
        honest = [row for wid, row in enumerate(self.rows) if wid not in self.byzantine]
        for wid, kind in self.byzantine.items():
            row = self.rows[wid]
            if kind == "random":
                row.copy_(attacks.random_like(row))
            elif kind == "reverse":
                row.copy_(attacks.reverse(row))
            elif kind == "drop":
                row.copy_(attacks.drop(row))
            elif kind in ("lie", "empire"):
                row.copy_(attacks.lie(honest or [row]) if kind == "lie" else attacks.empire(honest or [row]))
            else:
                raise ValueError("unknown attack %r" % (kind,))

    def _attack(self, w):
        if w.attack == "random":
            w.row.copy_(attacks.random_like(w.row))
        elif w.attack == "reverse":
            w.row.copy_(attacks.reverse(w.row))
        elif w.attack == "drop":
            w.row.copy_(attacks.drop(w.row))
        elif w.attack in ("lie", "empire"):
            honest = [o.row for o in self.workers if o.attack is None] or [w.row]
            w.row.copy_(attacks.lie(honest) if w.attack == "lie" else attacks.empire(honest))
        else:
            raise ValueError("unknown attack %r" % (w.attack,))

    def _allreduce(self, tensor):
        """Sum a small device tensor (the n×n partial distances of this rank's coordinate slice) over the server ranks."""
        if self.sync != "flags":
            dist.all_reduce(tensor, group=self.group)        # baseline
            return
        # peer slots: publish my partial, raise my flag, wait for everyone's, sum the slots in rank order (bitwise the same on
        # every rank) — device to device, no collective library
        import ctypes
        if self._reductions_this_step:
            raise RuntimeError("one partial-distance reduction per step (the flags carry the step's epoch)")
        self._reductions_this_step += 1
        k = tensor.numel()
        self.dist_arena.local[:k].copy_(tensor.reshape(-1))
        self._signal(self._my_dist_flag)
        self._wait(self._dist_flags)
        slots = (ctypes.c_void_p * len(self._servers))(*[self.dist_arena.peer(r).data_ptr() for r in self._servers])
        _lib.check(_lib.require().gb_reduce_slots(slots, len(self._servers), k, tensor.data_ptr(), _lib.stream_ptr(self.device)))
        _lib.count()

    def aggregate_and_update(self, rows=None):
        """Server side: fused gather + GAR + optimizer step + parameter publication, on the whole
        vector (rank 0) or on this rank's coordinate slice (sharded). `rows`: the subset of gradient rows to hand to the
        rule (quorum mode: only those that arrived); default all n."""
        opt = self.optimizer
        if self.scatter and self.byzantine and self.rows:
            self._attack_slices()
        all_rows, self.rows = self.rows, (rows if rows is not None else self.rows)
        try:
            self._aggregate(opt)
        finally:
            self.rows = all_rows

    def _rule_kwargs(self):
        if self.gar_name == "condense":
            return dict(self.gar_kwargs, seed=self._condense_seed)
        return self.gar_kwargs

    def _aggregate(self, opt):
        if self.gar_name == "condense":
            # a fresh Bernoulli mask per step and per coordinate shard (the reference draws one per call, aggregators/condense.py:35)
            base = int(self.gar_kwargs.get("seed", 0))
            self._condense_seed = (base * 0x9E3779B97F4A7C15 + (self.steps_done + 1) * 0xBF58476D1CE4E5B9 + self.lo * 0x94D049BB133111EB) & (2 ** 63 - 1)
        if self._tail_at is not None and self._early_ws is not None:
            early, self._early_ws = self._early_ws["dist"], None
            head = [row[:self._tail_at] for row in self.rows]
            run = lambda **kw: gar_ops.aggregate_sharded(self.gar_name, head, self.f, lambda dist: dist.add_(early), apply_rows=self.rows,
                                                         **kw, **self.gar_kwargs)
        elif self.sharded:
            if opt is None:  # empty slice: still take part in the distance all-reduce
                raise RuntimeError("empty coordinate slice: use fewer ranks than d / 64")
            run = lambda **kw: gar_ops.aggregate_sharded(self.gar_name, self.rows, self.f, self._allreduce, **kw, **self._rule_kwargs())
        else:
            run = lambda **kw: gar_ops.aggregate(self.gar_name, self.rows, self.f, **kw, **self._rule_kwargs())
        if opt.name == "sgd":
            run(update=opt.sgd_fusion(self.mirrors, multicast=getattr(self, "multicast", None)), want_out=False)
            opt.note_fused_step()
        else:
            opt.step(run(), mirrors=self.mirrors)

    # -- flag synchronisation -------------------------------------------------------------------------
    def _bump(self):
        _lib.check(_lib.require().gb_epoch_bump(self._epoch_dev.data_ptr(), _lib.stream_ptr(self.device)))
        _lib.count()
        self._reductions_this_step = 0

    def _signal(self, flag_ptr):
        import ctypes
        _lib.check(_lib.require().gb_flag_signal_ep(ctypes.c_void_p(flag_ptr), self._epoch_dev.data_ptr(), _lib.stream_ptr(self.device)))
        _lib.count()

    def _wait(self, flag_ptrs, lag=0, arrived=None, patience=1):
        import ctypes
        table = (ctypes.c_void_p * len(flag_ptrs))(*flag_ptrs)
        _lib.check(_lib.require().gb_flag_wait_ep(table, len(flag_ptrs), self._epoch_dev.data_ptr(), lag, patience * self.flag_timeout_ns, self._late.data_ptr(),
                                                 arrived.data_ptr() if arrived is not None else None, _lib.stream_ptr(self.device)))
        _lib.count()

    def _quorum_rows(self):
        """Quorum mode: the lead server (rank 0) decides which ranks' gradients made it — all of them, or at least the
        quorum once the deadline has passed — and publishes that mask to every server rank; each server then hands the
        rule only those workers' rows (reference: garfieldpp/server.py:134-155). One host read per step."""
        import ctypes
        lib = _lib.require()
        need_ranks = max(1, min(self.world, -(-self.num_wait // self.max_local)))
        if self.rank == 0:
            k = len(self._servers)
            flags = (ctypes.c_void_p * self.world)(*self._grad_flags)
            words = (ctypes.c_void_p * k)(*[self._word(r, 4) for r in self._servers])
            dflags = (ctypes.c_void_p * k)(*[self._word(r, 3) for r in self._servers])
            _lib.check(lib.gb_quorum_decide(flags, self.world, self._epoch_dev.data_ptr(), need_ranks, self.flag_timeout_ns, words, dflags, k,
                                            self._late.data_ptr(), _lib.stream_ptr(self.device)))
            _lib.count()
        self._wait([self._word(self.rank, 3)], patience=20)       # the lead server decides after ITS deadline (and gives up after 16)
        mask_words = self.flag_arena.local[4:6].view(torch.int32).cpu()          # synchronises: the host picks the row subset
        mask = (int(mask_words[0]) & 0xffffffff) | ((int(mask_words[1]) & 0xffffffff) << 32)
        arrived = [bool((mask >> r) & 1) for r in range(self.world)]
        self.last_arrived = arrived
        if all(arrived):
            return None
        keep = [row for wid, row in enumerate(self.rows) if arrived[wid % self.world]]
        from .. import aggregators
        rule = aggregators.gars.get(self.gar_name)
        if rule is not None and rule.check(gradients=[torch.zeros(1)] * len(keep), f=self.f, **self.gar_kwargs) is not None:
            return None      # too few timely rows for this rule: fall back to everything that is in the arenas
        return keep

    def wait_published(self):
        """Block the stream until every shard's parameters of the last step have landed in this rank's arena
        (flags mode defers that wait to the next step; NCCL mode already ends each step with a barrier)."""
        if self.sync == "flags" and self.epoch > 0:
            self._wait(self._param_flags)

    @property
    def late_peers(self):
        """Number of (peer, wait) pairs that missed a deadline so far (flags mode; synchronises)."""
        return int(self._late.item()) if self.sync == "flags" else 0

    def step(self, iter_num, load=True, stall=False):
        """One training iteration. Returns the mean loss of this rank's workers (device scalar).
        `stall=True` (fault injection, flags mode) makes this rank withhold its "gradients ready" signal."""
        if load:
            self.load_batches(iter_num)
        if self._whole_step_capturable() and not stall:
            return self._step_whole()
        if self.sync == "flags":
            self.epoch += 1
            self._bump()
            self._wait(self._param_flags, lag=1)                   # every shard published the previous update
            self.compute_gradients()
            if not stall:
                self._signal(self._my_grad_flag)
            if self.is_server:
                rows = None
                if self.num_wait:
                    rows = self._quorum_rows()
                else:
                    self._wait(self._grad_flags)                   # all ranks' gradient rows are complete (or late)
                self.aggregate_and_update(rows)
                self._signal(self._my_param_flag)
            self.steps_done += 1
            if self.workers:
                return torch.stack([w.loss for w in self.workers]).mean()
            return torch.zeros((), device=self.device)
        self.compute_gradients()
        self._barrier()                      # all gradient rows are complete
        if self.is_server:
            self.aggregate_and_update()
        self._barrier()                      # parameters published to every arena
        self.steps_done += 1
        if self.workers:
            return torch.stack([w.loss for w in self.workers]).mean()
        return torch.zeros((), device=self.device)

    # -- single-GPU fast path: the WHOLE iteration (gradient pass, attacks, GAR + fused SGD, loss) is one CUDA graph ----
    def _whole_step_capturable(self):
        """One GPU, or several with flag synchronisation (no NCCL call in the step) and a server shard on every rank."""
        multi = self.world > 1
        if multi and not (self.sync == "flags" and self.sharded and not self.num_wait):
            return False
        attacks_ok = all(a in (None, "reverse", "lie", "empire") for a in
                         (self.byzantine.values() if self.scatter else [w.attack for w in self.workers]))
        return bool(self.use_graphs and self._whole_step is not False and self.workers and self.optimizer is not None
                    and self.optimizer.name == "sgd" and attacks_ok and self.gar_name != "condense")   # condense: per-step seed argument

    def _step_body(self):
        flags = self.sync == "flags"
        if flags:
            self._bump()
            self._wait(self._param_flags, lag=1)
        self._run_workers()
        for w in ([] if self.scatter else self.workers):
            if w.attack is not None:
                self._attack(w)
        if flags:
            self._signal(self._my_grad_flag)
            self._wait(self._grad_flags)
        self.aggregate_and_update()
        if flags:
            self._signal(self._my_param_flag)
        self._step_loss.copy_(self._group_losses.mean() if self.fused is not None else torch.stack([w.loss for w in self.workers]).mean())

    def _step_whole(self):
        opt = self.optimizer
        if not hasattr(self, "_step_loss"):
            self._step_loss = torch.zeros((), device=self.device)
            self._eager_steps, self._step_graph, self._step_key = 0, None, None
        key = tuple(sorted((k, v) for k, v in opt.hyper.items() if isinstance(v, (int, float, bool))))
        if self._eager_steps < 2 or opt.steps == 0:      # warm-up: cuDNN autotuning, allocator, momentum-buffer initialisation
            self._step_body()
            self._eager_steps += 1
        else:
            if self._step_graph is None or key != self._step_key:   # hyper-parameters live in kernel arguments: re-capture on change
                try:
                    graph = torch.cuda.CUDAGraph()
                    before, steps_before = _lib.launches[0], opt.steps
                    with torch.cuda.graph(graph):
                        self._step_body()
                    opt.steps = steps_before
                    self._step_launches = _lib.launches[0] - before
                    self._step_graph, self._step_key = graph, key
                except Exception:
                    self._whole_step = False            # something in the rule is not capturable: keep the two-phase path
                    torch.cuda.synchronize(self.device)
                    return self.step(self.steps_done, load=False)
            self._step_graph.replay()
            opt.note_fused_step()
            _lib.count(self._step_launches)
        self.steps_done += 1
        self.epoch += 1
        return self._step_loss

    # -- evaluation and checkpoints -----------------------------------------------------------------------
    def accuracy(self, batches):
        """Top-1 accuracy (percent) of the current parameters on `batches` [(x, y), ...], evaluated with
        local worker 0's BatchNorm statistics (the master module never runs a training forward)."""
        module = self.workers[0].replica.module if self.workers else self.master.module
        was_training = module.training
        module.eval()
        correct = total = 0
        with torch.no_grad():
            for x, y in batches:
                x = x.to(self.device, non_blocking=True)
                if self.channels_last and x.dim() == 4:
                    x = x.contiguous(memory_format=torch.channels_last)
                correct += int(module(x).max(1)[1].eq(y.to(self.device)).sum())
                total += y.numel()
        module.train(was_training)
        return 100.0 * correct / max(total, 1)

    def save_checkpoint(self, path):
        """Write `<path>.rank<r>`: parameters as a torch `state_dict` (logical layout), this rank's optimizer
        shard, its workers' BatchNorm buffers, the step counter and the RNG state."""
        self.wait_published()
        torch.cuda.synchronize(self.device)
        # BatchNorm's per-layer `num_batches_tracked += 1` kernels are skipped on the hot path (_skip_bn_counters); the counter is
        # a pure function of the number of training passes, so it is written here — checkpoints equal what torch would hold
        for w in self.workers:
            for name, buf in w.replica.module.named_buffers():
                if name.endswith("num_batches_tracked"):
                    buf.fill_(self.steps_done)
        payload = {"format": 1, "step": self.steps_done, "lo": self.lo, "hi": self.hi, "world": self.world,
                   "model": {k: v.detach().cpu().contiguous() for k, v in self.master.module.state_dict().items()},
                   "optimizer": self.optimizer.state_dict() if self.optimizer is not None else None,
                   "buffers": {w.wid: {k: v.detach().cpu() for k, v in w.replica.module.named_buffers()} for w in self.workers},
                   "rng": {"cpu": torch.get_rng_state(), "cuda": torch.cuda.get_rng_state(self.device)}}
        torch.save(payload, "%s.rank%d" % (path, self.rank))

    def load_checkpoint(self, path):
        """Restore what `save_checkpoint` wrote (same world size and sharding)."""
        payload = torch.load("%s.rank%d" % (path, self.rank), map_location="cpu", weights_only=False)
        if payload["world"] != self.world or (payload["lo"], payload["hi"]) != (self.lo, self.hi):
            raise ValueError("checkpoint was written with a different world size / sharding")
        self.master.load_state_dict(payload["model"])          # copies into the arena views
        if self.optimizer is not None and payload["optimizer"] is not None:
            self.optimizer.load_state_dict(payload["optimizer"])
        for w in self.workers:
            own = dict(w.replica.module.named_buffers())
            with torch.no_grad():
                for k, v in payload["buffers"].get(w.wid, {}).items():
                    own[k].copy_(v)
        self.steps_done = payload["step"]
        torch.set_rng_state(payload["rng"]["cpu"])
        torch.cuda.set_rng_state(payload["rng"]["cuda"], self.device)
        self._barrier()
        return payload["step"]

    def close(self):
        self._graph = None
        self.wait_published()
        torch.cuda.synchronize(self.device)
        if self.sync == "flags":
            self.dist_arena.close()
            self.flag_arena.close()
        self.grad_arena.close()
        self.param_arena.close()
