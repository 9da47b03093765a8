"""Run under torchrun by tests/test_multi_gpu.py: sharded vs single-server aggregation over peer memory."""
import os
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
os.environ.setdefault("GARFIELD_SYNTHETIC_SIZE", "1500")

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from garfield_b200.parallel.engine import SSMWEngine  # noqa: E402


def main():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    for gar, n, f in (("krum", 8, 1), ("median", 8, 1), ("bulyan", 8, 1), ("brute", 7, 2), ("aksel", 7, 1)):
        finals = []
        for sharded in (False, True):
            eng = SSMWEngine("cifarnet", "synthetic-cifar10", "cross-entropy", n, f, gar, 16, "sgd", dict(lr=0.05, momentum=0.9),
                             byzantine={n - 1: "reverse"}, use_graphs=False, sharded=sharded, seed=99)
            for i in range(3):
                eng.step(i)
            eng.wait_published()          # flag synchronisation defers "parameters landed" to the next step
            torch.cuda.synchronize()
            flat = eng.flat.clone()
            gathered = [torch.empty_like(flat) for _ in range(world)]
            dist.all_gather(gathered, flat)
            for other in gathered[1:]:
                assert torch.equal(gathered[0], other), "%s: parameter replicas diverged (sharded=%s)" % (gar, sharded)
            assert torch.isfinite(flat).all()
            finals.append(flat)
            eng.close()
        if gar != "aksel":  # aksel keeps the (n+1)//2 closest to the median: near-ties may legitimately flip with the summation order
            torch.testing.assert_close(finals[0], finals[1], rtol=1e-4, atol=1e-6, msg=lambda m: "%s sharded != centralized: %s" % (gar, m))
        if rank == 0:
            print("ok", gar, flush=True)
    # device-side flag synchronisation (no NCCL barrier) gives the same parameters, and a rank that
    # withholds its signal only costs the others one deadline
    finals = []
    for sync in ("nccl", "flags"):
        eng = SSMWEngine("cifarnet", "synthetic-cifar10", "cross-entropy", 8, 1, "median", 16, "sgd", dict(lr=0.05, momentum=0.9),
                         use_graphs=False, sharded=True, seed=5, sync=sync, flag_timeout=20.0)
        for i in range(4):
            eng.step(i)
        eng.wait_published()
        torch.cuda.synchronize()
        finals.append(eng.flat.clone())
        if sync == "flags":
            assert eng.late_peers == 0
            eng.flag_timeout_ns = int(0.3e9)               # short deadline for the fault-injection step
            eng.step(4, stall=(rank == world - 1))      # the last rank never announces its gradients
            torch.cuda.synchronize()
            late = torch.tensor([eng.late_peers], device="cuda")
            dist.all_reduce(late)
            assert late.item() >= 1, "the stalled rank must be reported late"
            assert torch.isfinite(eng.flat).all()
        eng.close()
    torch.testing.assert_close(finals[0], finals[1], rtol=1e-4, atol=1e-6)   # cuDNN wgrad is not bitwise reproducible
    if rank == 0:
        print("ok flags", flush=True)
    # quorum: the last rank withholds its "gradients ready" signal; the lead server decides after the deadline, publishes the
    # arrival mask, and every shard hands the rule ONLY the timely workers' rows (reference: server.get_gradients(i, n - f))
    eng = SSMWEngine("cifarnet", "synthetic-cifar10", "cross-entropy", 8, 1, "median", 16, "sgd", dict(lr=0.05), use_graphs=False,
                     sharded=True, seed=11, scatter_grads=False, flag_timeout=0.5, num_wait=8 - 8 // world)
    for i in range(2):
        eng.step(i)
    assert eng.last_arrived == [True] * world
    eng.wait_published()
    torch.cuda.synchronize()
    dist.barrier()
    before = eng.flat.clone()
    eng.step(2, stall=(rank == world - 1))
    eng.wait_published()
    torch.cuda.synchronize()
    dist.barrier()
    assert eng.last_arrived == [True] * (world - 1) + [False], eng.last_arrived
    timely = [row for wid, row in enumerate(eng.rows) if wid % world != world - 1]
    expect = before[eng.lo:eng.hi] - 0.05 * torch.stack([r.clone() for r in timely]).median(0)[0]
    torch.testing.assert_close(eng.flat[eng.lo:eng.hi], expect, rtol=1e-5, atol=1e-7)
    eng.close()
    if rank == 0:
        print("ok quorum", flush=True)
    # scatter mode (every gradient slice written straight into its aggregating rank's arena, owner-side attacks, CUDA graph)
    # against the pull-over-NVLink sharded server, on a model with BatchNorm and k×k / 1×1 / Linear weight-gradient paths
    finals = []
    for scatter in (False, True):
        eng = SSMWEngine("cnn", "synthetic-cifar10", "cross-entropy", 8, 1, "krum", 8, "sgd", dict(lr=0.02, momentum=0.9),
                         byzantine={7: "reverse"}, use_graphs=True, sharded=True, seed=17, scatter_grads=scatter)
        assert eng.scatter == scatter and eng.fused is not None
        for module in [w.replica.module for w in eng.workers] + [eng.fused.module]:
            for m in module.modules():
                if isinstance(m, (torch.nn.Dropout, torch.nn.Dropout2d)):
                    m.p = 0.0
        for i in range(4):
            eng.step(i)
        eng.wait_published()
        torch.cuda.synchronize()
        finals.append(eng.master.to_reference_order(eng.flat).clone())
        eng.close()
    diff = (finals[0] - finals[1]).abs()
    assert diff.max().item() < 2e-3 and (diff > 2e-4).float().mean().item() < 1e-3, "scatter != pull: max %g" % diff.max().item()
    # colluding attacks are applied by the owner of each coordinate range (they see every honest worker's slice)
    eng = SSMWEngine("cnn", "synthetic-cifar10", "cross-entropy", 8, 1, "median", 8, "sgd", dict(lr=0.02, momentum=0.9),
                     byzantine={7: "lie"}, use_graphs=True, sharded=True, seed=17)
    assert eng.scatter
    losses = [eng.step(i).item() for i in range(3)]
    eng.wait_published()
    torch.cuda.synchronize()
    assert all(l == l for l in losses) and torch.isfinite(eng.flat).all()
    eng.close()
    if rank == 0:
        print("ok scatter", flush=True)
    topologies(rank, world)
    dist.destroy_process_group()


def topologies(rank, world):
    """MSMW (ByzSGD) and decentralized (LEARN) engines across ranks: replicas of honest servers / nodes stay identical on
    every rank, a Byzantine server publishing garbage is voted out by the model median, flag and NCCL synchronisation agree."""
    from garfield_b200.parallel.topologies import DecentralizedEngine, MSMWEngine
    finals = []
    for sync in ("flags", "nccl"):
        eng = MSMWEngine("cifarnet", "synthetic-cifar10", "cross-entropy", 5, 3, 1, 1, "krum", 16, mar="median", opt_args=dict(lr=0.05, momentum=0.9),
                         byzantine_workers={4: "reverse"}, byzantine_servers={2: "random"}, seed=21, sync=sync)
        for i in range(4):
            eng.step(i)
        torch.cuda.synchronize()
        dist.barrier()
        # every honest server holds the same model (server k lives on rank k mod world)
        mine = [eng.server_model(j).clone() for j, srv in enumerate(eng.servers) if not srv["attack"]]
        ref = mine[0].clone() if (rank == 0 and mine) else torch.empty(eng.d, device="cuda")
        dist.broadcast(ref, src=0)
        for m in mine:
            assert torch.isfinite(m).all() and torch.equal(m, ref), "honest server replicas diverged (%s)" % sync
        assert eng.late_peers == 0
        finals.append(ref.clone())
        eng.close()
    torch.testing.assert_close(finals[0], finals[1], rtol=1e-4, atol=1e-6)
    if rank == 0:
        print("ok msmw", flush=True)
    finals = []
    for sync in ("flags", "nccl"):
        eng = DecentralizedEngine("cifarnet", "synthetic-cifar10", "cross-entropy", 8, 1, "median", 16, opt_args=dict(lr=0.05, momentum=0.9),
                                  byzantine={7: "lie"}, non_iid=True, seed=23, sync=sync)
        for i in range(4):
            eng.step(i)
        torch.cuda.synchronize()
        dist.barrier()
        models = [eng.model_of(j).clone() for j in range(len(eng.nodes))]
        ref = models[0].clone() if rank == 0 else torch.empty(eng.d, device="cuda")
        dist.broadcast(ref, src=0)
        for m in models:
            assert torch.isfinite(m).all() and torch.equal(m, ref), "node models diverged after the contract step (%s)" % sync
        finals.append(ref.clone())
        eng.close()
    torch.testing.assert_close(finals[0], finals[1], rtol=1e-4, atol=1e-6)
    if rank == 0:
        print("ok learn", flush=True)
    # gar = average: agreement rounds and contract are in-switch reductions over an NVLS multicast address (multimem.ld_reduce /
    # multimem.st) when the platform has one; the NCCL-synchronised engine keeps the per-peer loads — same models
    finals, used = [], []
    for sync in ("flags", "nccl"):
        eng = DecentralizedEngine("cifarnet", "synthetic-cifar10", "cross-entropy", 8, 0, "average", 16, opt_args=dict(lr=0.05, momentum=0.9),
                                  non_iid=True, seed=29, sync=sync)
        used.append(eng._nvls)
        for i in range(4):
            eng.step(i)
        torch.cuda.synchronize()
        dist.barrier()
        models = [eng.model_of(j).clone() for j in range(len(eng.nodes))]
        for m in models[1:]:
            assert torch.equal(m, models[0])
        assert torch.isfinite(models[0]).all()
        finals.append(models[0])
        eng.close()
    # the switch adds the ranks' copies in its own order: float rounding differs from the sequential sum, and four training
    # steps carry that through the network
    diff = (finals[0] - finals[1]).abs()
    assert diff.max().item() < 2e-4 and (diff > 2e-6).float().mean().item() < 0.02, diff.max().item()
    if rank == 0:
        print("ok nvls-average (multicast used: %s)" % used[0], flush=True)


if __name__ == "__main__":
    main()
