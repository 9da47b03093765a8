"""LEARN web demo: POST / starts a run, GET /status reports progress then the final accuracy."""

import importlib.util
import json
import pathlib
import threading
import time
import urllib.request
from http.server import ThreadingHTTPServer

ROOT = pathlib.Path(__file__).resolve().parent.parent


def test_demo_http_roundtrip():
    spec = importlib.util.spec_from_file_location("learn_demo", ROOT / "applications" / "LEARN" / "demo.py")
    demo = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(demo)
    demo.EPOCHS = 1
    server = ThreadingHTTPServer(("127.0.0.1", 0), demo.Handler)
    port = server.server_address[1]
    threading.Thread(target=server.serve_forever, daemon=True).start()
    try:
        base = "http://127.0.0.1:%d" % port
        assert b"LEARN" in urllib.request.urlopen(base + "/", timeout=10).read()
        req = urllib.request.Request(base + "/", data=json.dumps({"n": 4, "f": 1, "gar": "median"}).encode(), method="POST")
        trainer_id = json.loads(urllib.request.urlopen(req, timeout=10).read())["trainer_id"]
        deadline = time.time() + 120
        status = {}
        while time.time() < deadline:
            status = json.loads(urllib.request.urlopen(base + "/status?trainer_id=%d" % trainer_id, timeout=10).read())
            assert "error" not in status, status
            if "result" in status:
                break
            time.sleep(0.2)
        assert 0.0 <= status.get("result", -1) <= 100.0, status
        bad = urllib.request.Request(base + "/", data=b"n=3&f=2&gar=median", method="POST")
        try:
            urllib.request.urlopen(bad, timeout=10)
            assert False, "invalid (n, f) must be rejected"
        except urllib.error.HTTPError as err:
            assert err.code == 400
    finally:
        server.shutdown()


def test_demo_runs_its_nodes_as_processes_over_rpc():
    """--transport processes: n local processes named node:<rank> over torch RPC on a fresh localhost port, progress through
    a multiprocessing queue (the reference's demo.py:264-335); /status names the transport and where the table came from."""
    import os
    os.environ["GARFIELD_DEMO_EPOCHS"] = "1"
    spec = importlib.util.spec_from_file_location("learn_demo_mp", ROOT / "applications" / "LEARN" / "demo.py")
    demo = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(demo)
    demo.EPOCHS = 1
    demo.TRANSPORT[0] = "processes"
    try:
        trainer = demo.Trainer(1, 3, 1, "median").start()
        trainer.thread.join(timeout=400)
        status = trainer.status()
        assert "error" not in status, status
        assert status["transport"] == "processes" and "dataset" in status
        assert 0.0 <= status.get("result", -1) <= 100.0, status
    finally:
        demo.TRANSPORT[0] = "threads"
        os.environ.pop("GARFIELD_DEMO_EPOCHS", None)


def test_playground_scripts_run(tmp_path):
    """rsrcs/playgrounds counterparts: the random-attack Aggregathor variant and the centralized tutorial loop."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, GARFIELD_SYNTHETIC_SIZE="400", CUDA_VISIBLE_DEVICES="")
    out = subprocess.run([sys.executable, os.path.join(root, "applications/playgrounds/centralized/cifar10.py"), "--dataset", "synthetic-cifar10",
                          "--epochs", "1", "--batch", "50", "--max_steps", "3"], env=env, capture_output=True, text=True, timeout=240)
    assert out.returncode == 0 and "test accuracy" in out.stdout, out.stderr[-800:]
    out = subprocess.run([sys.executable, os.path.join(root, "applications/playgrounds/trainer.py"), "--dataset", "synthetic-mnist", "--model", "convnet",
                          "--num_workers", "7", "--fw", "1", "--gar", "krum", "--num_iter", "4", "--acc_freq", "2", "--batch", "16"],
                         env=env, capture_output=True, text=True, timeout=240)
    assert out.returncode == 0, out.stderr[-800:]


def test_benchmark_scripts_run_on_cpu(tmp_path):
    """applications/benchmarks: the GAR sweep (CPU device, small budget) writes its JSON, and the collective model-pull
    benchmark runs in a single process."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = tmp_path / "gar.json"
    run = subprocess.run([sys.executable, os.path.join(root, "applications/benchmarks/gar_bench.py"), "--device", "cpu", "--sweep", "f",
                          "--gars", "median,krum,native-krum", "--reps", "1", "--max-bytes", "20000000", "--json", str(out)],
                         capture_output=True, text=True, timeout=600, env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
    assert run.returncode == 0, run.stderr[-1500:]
    rows = json.load(open(out))
    rows = rows["results"] if isinstance(rows, dict) and "results" in rows else rows
    assert len(rows) >= 3
