"""Launcher of the reference arm (see baseline/README.md): spawns the 9 reference processes, waits
for the server's timing, prints the bench JSON line."""

import json
import os
import pathlib
import shutil
import socket
import subprocess
import sys
import tempfile
import time

HERE = pathlib.Path(__file__).resolve().parent
REF = HERE / "_ref"
SRC = pathlib.Path("/root/reference")


def ensure_reference():
    """baseline/_ref must hold the unmodified reference; re-create it from /root/reference if absent."""
    marker = REF / "pytorch_impl" / "libs" / "garfieldpp" / "server.py"
    if marker.exists():
        return None
    if not (SRC / "pytorch_impl").exists():
        return "baseline/_ref is missing and /root/reference is not available to copy from (pip install is impossible: the reference has no setup.py/pyproject.toml)"
    REF.mkdir(parents=True, exist_ok=True)
    shutil.copytree(SRC / "pytorch_impl", REF / "pytorch_impl", symlinks=True, dirs_exist_ok=True)
    return None


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def run_once(gar, spec, gpus, steps, warmup, timeout):
    cfg = dict(spec, gar=gar, steps=steps, warmup=warmup)
    port = free_port()
    with tempfile.TemporaryDirectory() as tmp:
        result = os.path.join(tmp, "result.json")
        procs = []
        for rank in range(spec["n"] + 1):
            env = dict(os.environ)
            for key in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_PORT", "MASTER_ADDR", "TORCHELASTIC_RUN_ID", "GROUP_RANK",
                        "LOCAL_WORLD_SIZE", "ROLE_RANK", "ROLE_WORLD_SIZE", "PYTHONPATH"):
                env.pop(key, None)
            gpu = 0 if rank == 0 else (rank - 1) % gpus
            env.update(REF_CONFIG=json.dumps(cfg), REF_RANK=str(rank), REF_PORT=str(port), REF_RESULT=result,
                       CUDA_VISIBLE_DEVICES=str(gpu), TP_SOCKET_IFNAME="lo", GLOO_SOCKET_IFNAME="lo", OMP_NUM_THREADS="4")
            log = open(os.path.join(tmp, "node%d.log" % rank), "w")
            procs.append((subprocess.Popen([sys.executable, str(HERE / "ref_node.py")], env=env, stdout=log, stderr=subprocess.STDOUT,
                                           cwd=str(HERE)), log))
        deadline = time.time() + timeout
        failed = None
        while time.time() < deadline:
            if os.path.exists(result) and os.path.getsize(result) > 0:
                break
            dead = [i for i, (p, _) in enumerate(procs) if p.poll() not in (None, 0)]
            if dead:
                failed = "reference node %d exited with code %s" % (dead[0], procs[dead[0]][0].returncode)
                break
            time.sleep(0.5)
        else:
            failed = "reference run timed out after %ds" % timeout
        tail = ""
        if failed:
            for i, (p, log) in enumerate(procs):
                log.flush()
                text = open(log.name).read()[-1500:]
                if text.strip():
                    tail += "\n--- node %d ---\n%s" % (i, text)
        out = json.load(open(result)) if not failed else None
        for p, log in procs:  # exact PIDs we started
            try:
                p.wait(timeout=20 if not failed else 1)
            except subprocess.TimeoutExpired:
                p.kill()
            log.close()
        if failed:
            raise RuntimeError(failed + tail)
        return out


def main(args, emit, metric, spec, ClockSampler):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # torchrun started one process per GPU; the reference deployment is spawned by rank 0 only
    problem = ensure_reference()
    if problem:
        emit({"impl": "reference", "unavailable": problem})
        return
    gpus = max(1, args.gpus)
    steps, warmup = args.steps, max(args.warmup, 3)
    per_gar, primary, error = {}, None, None
    names = [args.gar] + [g for g in args.also.split(",") if g and g != args.gar]
    with ClockSampler(0) as clocks:
        for idx, name in enumerate(names):
            k = steps if idx == 0 else max(3, steps // 3)
            try:
                r = run_once(name, spec, gpus, k, warmup if idx == 0 else 3, timeout=1500)
            except Exception as err:
                error = str(err)[:1800]
                if idx == 0:
                    break
                continue
            ips = spec["n"] * spec["batch"] * r["steps"] / r["seconds"]
            per_gar[name] = round(ips, 2)
            if idx == 0:
                primary = dict(r, images_per_s=ips)
    if primary is None:
        emit({"impl": "reference", "unavailable": "reference run failed: %s" % (error,)})
        return
    emit({
        "metric": metric, "value": round(primary["images_per_s"], 2), "unit": "images/s", "n_gpus": gpus, "steps": primary["steps"],
        "warmup": warmup, "ms_per_step": round(primary["seconds"] / primary["steps"] * 1e3, 3), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "fp32 (TF32 convolutions = torch default)", "data": "synthetic",
        "impl": "reference",
        "config": {"model": "resnet50 (torchvision, 10 classes, 3x32x32 inputs)", "global_batch": spec["n"] * spec["batch"],
                   "seq_len": None, "parallelism": "reference SSMW over torch RPC: 1 server + 8 worker processes on %d gpu(s)" % gpus,
                   "gar": args.gar, "d": primary["d"], "torch": primary["torch"],
                   "timing": "wall clock around K train_steps on the server, cuda synchronize on both sides"},
        "per_gar": per_gar,
        "e2e": {"value": round(primary["images_per_s"], 2), "unit": "images/s", "h2d_bytes_per_step": None, "d2h_bytes_per_step": None,
                "note": "the reference step is end-to-end by construction (host-resident batches, CPU optimizer)"},
        "gpu_launches": None, "clocks": clocks.summary(), "error_other_gars": error,
    })
