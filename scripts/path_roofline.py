#!/usr/bin/env python
"""The four fused gather + aggregation paths over peer memory, each as a fraction of its roofline (torchrun, N GPUs).

Every rank aggregates ITS coordinate slice of n = 8 ResNet-50-sized rows (d = 23 528 522) that live on the ranks round-robin, so
(N − 1)/N of the bytes it reads cross NVLink. Roofline time = max(bytes over NVLink / measured peer bandwidth (770 GB/s per
direction per GPU, B200_PROFILING.md), bytes touched locally / measured HBM copy bandwidth). Device-timed (CUDA events), max over ranks."""
import json
import os
import pathlib
import sys

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist

NVLINK_GBS = 770.0


def main():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    from garfield_b200.ops import gar
    from garfield_b200.parallel.peer import PeerArena
    peaks = pathlib.Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json"
    hbm = json.loads(peaks.read_text())["hbm_gbs"] if peaks.exists() else 6650.0
    n, d = 8, 23528522
    pitch = (d + 63) // 64 * 64
    per = (pitch // 64 + world - 1) // world * 64
    lo, hi = min(d, rank * per), min(d, rank * per + per)
    slots = (n + world - 1) // world
    arena = PeerArena(slots * pitch, torch.device("cuda", local))
    arena.local.normal_()
    torch.cuda.synchronize(); dist.barrier()
    rows = [arena.peer(w % world)[(w // world) * pitch + lo:(w // world) * pitch + hi] for w in range(n)]
    remote = sum(1 for w in range(n) if w % world != rank)
    out = torch.empty(hi - lo, device="cuda")
    results = {}

    def timed(name, fn, link_bytes, local_bytes, reps=10, egress_bytes=0.0):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        # ms: this GPU's NVLink ingress, the busiest source GPU's NVLink egress, local HBM traffic — whichever is slowest
        floor = max(link_bytes / (NVLINK_GBS * 1e6), egress_bytes / (NVLINK_GBS * 1e6), local_bytes / (hbm * 1e6))
        results[name] = {"ms": round(t.item(), 4), "roofline_ms": round(floor, 4), "fraction_of_roofline": round(floor / t.item(), 3),
                         "nvlink_bytes_per_gpu": int(link_bytes), "achieved_link_gbs": round(link_bytes / t.item() / 1e6, 1)}

    slice_bytes = 4.0 * (hi - lo)
    # path 2: gather + coordinate-wise median (one pass over the n rows' slices)
    timed("path2_gather_median", lambda: gar.coordwise(rows, "median_lower", out=out), remote * slice_bytes, (n - remote + 1) * slice_bytes)
    timed("path2_gather_trimmed_mean", lambda: gar.coordwise(rows, "trimmed_mean", f=1, out=out), remote * slice_bytes, (n - remote + 1) * slice_bytes)
    # path 1: gather + Krum (distance pass + weighted-average pass: the slices cross NVLink twice)
    timed("path1_gather_krum", lambda: gar.krum(rows, 1, 5, 6, False, out=out), 2 * remote * slice_bytes, 2 * (n - remote) * slice_bytes + slice_bytes)
    # path 3: model pull — median of 3 server replicas (on ranks 0..2 mod N) of the WHOLE vector into this worker's model
    servers = [arena.peer(s % world)[:d] for s in range(3)]
    far = sum(1 for s in range(3) if s % world != rank)
    model = torch.empty(d, device="cuda")
    # every rank pulls every replica: a server GPU's NVLink EGRESS carries its replica to the N − 1 other ranks
    timed("path3_pull_median_3_replicas", lambda: gar.coordwise(servers, "median_lower", out=model), far * 4.0 * d, (3 - far + 1) * 4.0 * d,
          egress_bytes=(world - 1) * 4.0 * d)
    # path 4: one gossip round over the 8 nodes' whole vectors (every rank reads all peers' vectors, writes its nodes' copies)
    full = [arena.peer(w % world)[(w // world) * pitch:(w // world) * pitch + d] for w in range(n)]
    far4 = sum(1 for w in range(n) if w % world != rank)
    outs0 = [torch.empty(d, device="cuda") for _ in range(min(slots, 8))]
    seq, counter = torch.zeros(1, dtype=torch.int32, device="cuda"), torch.zeros(1, dtype=torch.int32, device="cuda")
    timed("path4_gossip_round_median", lambda: gar.gossip_rounds("median", full, full, outs0, outs0, 1, 1, 0, seq, counter),
          far4 * 4.0 * d, (n - far4 + len(outs0)) * 4.0 * d, reps=5, egress_bytes=(world - 1) * slots * 4.0 * d)
    if rank == 0:
        print(json.dumps({"n_gpus": world, "n_rows": n, "d": d, "nvlink_gbs_per_direction": NVLINK_GBS, "hbm_gbs": hbm, "paths": results}))
    arena.close()
    dist.destroy_process_group()


main()
