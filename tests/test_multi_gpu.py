"""Multi-GPU data plane (needs >= 2 GPUs): peer-memory arenas, fused gather + GAR over P2P,
coordinate-sharded aggregation == single-server aggregation, replicas stay bit-identical."""

import pathlib
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = pathlib.Path(__file__).resolve().parent.parent


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
def test_sharded_equals_centralized_over_peer_memory():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    world = min(4, torch.cuda.device_count())
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(ROOT / "tests" / "_mgpu_worker.py")]
    proc = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=str(ROOT))
    out_dir = ROOT / "gpurun_out"
    if out_dir.is_dir():     # keep the per-section log of the run (what ran on how many GPUs) next to the other GPU artefacts
        (out_dir / ("multi_gpu_test_%dgpu.log" % world)).write_text(proc.stdout[-20000:])
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-3000:]
    for gar in ("krum", "median", "bulyan", "brute", "aksel", "flags", "quorum", "scatter", "msmw", "learn", "nvls-average"):
        assert "ok " + gar in proc.stdout
