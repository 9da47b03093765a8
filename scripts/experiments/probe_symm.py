"""Probe (2+ GPUs, torchrun): does torch symmetric memory give a multicast (NVLS) address on this box, and what is the measured
peer read / write bandwidth of our own copy kernel over NVLink?"""
import os
import sys
import pathlib
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent.parent))
import torch
import torch.distributed as dist


def main():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    try:
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(1 << 20, dtype=torch.float32, device="cuda")
        h = symm.rendezvous(t, dist.group.WORLD)
        mc = getattr(h, "multicast_ptr", 0)
        if rank == 0:
            print("symmetric memory ok: multicast_ptr=%#x buffer_ptrs=%s" % (mc, [hex(p) for p in h.buffer_ptrs][:4]), flush=True)
    except Exception as err:
        if rank == 0:
            print("symmetric memory unavailable:", repr(err)[:300], flush=True)
    # P2P bandwidth with our own kernels: coordwise "average" of ONE peer row = a pure peer read of d floats
    from garfield_b200.parallel.peer import PeerArena
    from garfield_b200.ops import gar
    d = 64 * 1024 * 1024
    arena = PeerArena(d, torch.device("cuda", local))
    out = torch.empty(d, device="cuda")
    peer = arena.peer((rank + 1) % world)
    for name, rows in (("local", [arena.local]), ("peer", [peer])):
        for _ in range(3):
            gar.aggregate("average", rows, 0, out=out)
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            gar.aggregate("average", rows, 0, out=out)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        t = torch.tensor([ms], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print("read %s row (TMA bulk pull, all ranks at once): %.3f ms -> %.1f GB/s per GPU" % (name, t.item(), d * 4 / t.item() / 1e6), flush=True)
    arena.close()
    dist.destroy_process_group()


main()
