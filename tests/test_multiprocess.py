"""Multi-process plumbing on CPU: gloo collectives (Garfield_CC modes, world_size > 1) and the
torch-RPC compatibility transport of the role objects (config 1: MNIST CNN SSMW, Krum f=1)."""

import os
import socket
import subprocess
import sys
import pathlib

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch(script, common_args, per_rank_args, world, timeout=240, env_extra=None):
    port = free_port()
    procs = []
    for rank in range(world):
        env = dict(os.environ, GARFIELD_SYNTHETIC="1", GARFIELD_SYNTHETIC_SIZE="800", RANK=str(rank), WORLD_SIZE=str(world),
                   LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), OMP_NUM_THREADS="1",
                   GLOO_SOCKET_IFNAME="lo", TP_SOCKET_IFNAME="lo", CUDA_VISIBLE_DEVICES="")
        env.update(env_extra or {})
        cmd = [sys.executable, str(ROOT / script)] + common_args + ["--port", str(port)] + per_rank_args(rank)
        procs.append(subprocess.Popen(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=str(ROOT)))
    outs = []
    try:
        for p in procs:
            out, _ = p.communicate(timeout=timeout)
            outs.append(out)
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    for rank, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, "rank %d failed:\n%s" % (rank, out[-3000:])
    return outs


@pytest.mark.parametrize("mode", ["vanilla", "aggregathor", "byzsgd"])
def test_garfield_cc_gloo(mode):
    base = ["--dataset", "mnist", "--model", "convnet", "--loss", "nll", "--batch", "8", "--num_iter", "4", "--device", "cpu",
            "--backend", "gloo"]
    if mode == "vanilla":
        args, world = base + ["--num_ps", "1", "--num_workers", "3", "--gar", "vanilla"], 4
    elif mode == "aggregathor":
        args, world = base + ["--num_ps", "1", "--num_workers", "5", "--fw", "1", "--aggregator", "krum", "--attack", "reverse",
                              "--real_byzantine", "true"], 6
    else:
        args, world = base + ["--num_ps", "3", "--num_workers", "5", "--fw", "1", "--fps", "1", "--aggregator", "median", "--mar",
                              "median", "--server_attack", "reverse", "--real_byzantine", "true"], 8
    outs = launch("applications/Garfield_CC/trainer.py", args, lambda r: [], world)
    assert "model digest" in outs[0]


def test_rpc_ssmw_config1_plumbing():
    """1 server + 6 workers over torch RPC on CPU, Krum f=1 (server waits for 5 — SURVEY.md Q8)."""
    args = ["--transport", "rpc", "--dataset", "mnist", "--model", "convnet", "--loss", "nll", "--batch", "8", "--num_workers", "6",
            "--fw", "1", "--gar", "krum", "--num_iter", "3", "--device", "cpu", "--acc_freq", "1000", "--log", "true"]
    outs = launch("applications/Aggregathor/trainer.py", args, lambda r: ["--rank", str(r)], 7)
    assert "Iteration 2 loss" in outs[0]


def _train_dist_rank(rank, world, port, queue, extra):
    import importlib.util
    import os

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "applications", "playgrounds", "train_dist.py")
    spec = importlib.util.spec_from_file_location("train_dist_playground", path)
    module = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(module)
    flat, losses = module.main(["--dataset", "synthetic-mnist", "--model", "convnet", "--num_iter", "8", "--batch", "8", "--device", "cpu"] + extra,
                               spawned=True)
    queue.put((rank, flat.numpy().copy(), losses))


@pytest.mark.parametrize("broadcast", [True, False])
def test_playground_train_dist_parameter_server(broadcast):
    """The vanilla data-parallel playground over gloo (1 server + 2 workers): with the broadcast every rank ends on the server's
    model and the loss falls; without it (the reference's behaviour) the workers drift away from the server."""
    import socket

    import numpy as np
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = [ctx.Process(target=_train_dist_rank, args=(r, 3, port, queue, [] if broadcast else ["--no-broadcast"])) for r in range(3)]
    for p in procs:
        p.start()
    results = {}
    for _ in procs:
        rank, flat, losses = queue.get(timeout=240)
        results[rank] = (flat, losses)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    server = results[0][0]
    assert np.isfinite(server).all() and results[0][1] == []
    assert len(results[1][1]) == 8 and len(results[2][1]) == 8
    if broadcast:
        assert np.array_equal(results[1][0], server) and np.array_equal(results[2][0], server)
    else:
        assert not np.array_equal(results[1][0], server)
