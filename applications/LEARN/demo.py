#!/usr/bin/env python
"""LEARN web demo — headless JSON API (no template engine, standard library only).

Counterpart of the reference's Quart demo (applications/LEARN/demo.py:95-446, README-demo.md):

    GET  /                      a small HTML form
    POST /   n=<=10, f, gar     starts a decentralized training of PimaNet on the Pima diabetes table
                                (n nodes, the first f are Byzantine 'random' workers, RMSprop, 15
                                epochs) and returns {"trainer_id": ...}
    GET  /status?trainer_id=ID  {"progress": percent} while running, then {"result": mean binary accuracy}

Two ways of running the n nodes, as in the reference:

    --transport processes   n local processes named node:<rank> that talk over torch RPC on a fresh localhost port and report
                            progress through a multiprocessing queue (what the reference does, demo.py:264-335);
    --transport threads     the n nodes are role objects in one worker thread of this process (local transport; default —
                            it needs no ports and starts in milliseconds).

    python applications/LEARN/demo.py --port 8080            # serve (http://127.0.0.1:8080/ is the front-end)
    python applications/LEARN/demo.py --init_demo            # warm-up: build natives, load the dataset
"""
import argparse
import json
import pathlib
import queue
import sys
import threading
import urllib.parse
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[2]))

import os

EPOCHS, BATCH, MAX_NODES = int(os.environ.get("GARFIELD_DEMO_EPOCHS", "15")), 16, 10
PAGE = (pathlib.Path(__file__).with_name("templates") / "index.html")
TRANSPORT = ["threads"]


def pima_source():
    """Where the demo's table comes from: the real CSV, or the synthetic stand-in used when it is not installed."""
    from garfield_b200.garfieldpp.datasets import PimaDiabetesDataset
    return "pima_diabetes.csv" if getattr(PimaDiabetesDataset(train=False), "source", None) else "synthetic stand-in (pima_diabetes.csv not installed)"


class Trainer:
    """One demo run: n nodes (first f Byzantine), LEARN loop without LR schedule."""

    def __init__(self, trainer_id, n, f, gar):
        self.id, self.n, self.f, self.gar = trainer_id, n, f, gar
        self.progress, self.result, self.error = 0.0, None, None
        self.events = queue.Queue()
        self.thread = threading.Thread(target=self._train, daemon=True)

    def start(self):
        self.thread.start()
        return self

    def _train_processes(self):
        """n processes over torch RPC; this thread only drains their progress queue."""
        import socket
        import torch.multiprocessing as mp
        ctx = mp.get_context("spawn")
        with socket.socket() as sock:
            sock.bind(("127.0.0.1", 0))
            port = sock.getsockname()[1]
        events = ctx.Queue()
        from garfield_b200.apps.learn_demo_node import node_process
        procs = [ctx.Process(target=node_process, args=(r, self.n, self.f, self.gar, port, EPOCHS, BATCH, events), daemon=True) for r in range(self.n)]
        progress, results, done = [0.0] * self.n, {}, 0
        try:
            for p in procs:
                p.start()
            while done < self.n:
                try:
                    msg = events.get(timeout=600)
                except queue.Empty:
                    raise TimeoutError("a node stopped reporting")
                if "error" in msg:
                    raise RuntimeError("node %d: %s" % (msg["rank"], msg["error"]))
                progress[msg["rank"]] = msg["progress"]
                self.progress = sum(progress) / self.n
                if "result" in msg:
                    done += 1
                    if msg["result"] is not None:
                        results[msg["rank"]] = msg["result"]
            self.result = sum(results.values()) / max(1, len(results))
        except Exception as err:
            self.error = "%s: %s" % (type(err).__name__, err)
        finally:
            for p in procs:          # exactly the processes started above
                p.join(timeout=30)
                if p.is_alive():
                    p.terminate()

    def _train(self):
        if TRANSPORT[0] == "processes":
            return self._train_processes()
        try:
            import torch
            import aggregators
            from garfieldpp.byzWorker import ByzWorker
            from garfieldpp.server import Server
            from garfieldpp.worker import Worker
            rule = aggregators.gars[self.gar]
            n, f = self.n, self.f
            workers = [ByzWorker(r, n, n, BATCH, "pimanet", "pima", "binary-cross-entropy", "random", device="cpu") if r < f
                       else Worker(r, n, n, BATCH, "pimanet", "pima", "binary-cross-entropy", device="cpu") for r in range(n)]
            servers = []
            for r in range(n):
                torch.manual_seed(1234)
                ps = Server(r, 1, n, n, f, f, "node:", "node:", BATCH, "pimanet", "pima", "rmsprop", device="cpu", lr=0.001)
                ps.attach_workers(workers)
                servers.append(ps)
            for ps in servers:
                ps.attach_servers(servers)
            per_epoch = max(1, workers[-1].num_train_samples)
            total = EPOCHS * per_epoch
            for i in range(total):
                aggrs = [rule(gradients=ps.get_gradients(i, n - f), f=f) for ps in servers]
                for ps, aggr in zip(servers, aggrs):
                    ps.update_model(aggr)
                pulled = [ps.get_models(n - f) for ps in servers]
                for ps, models in zip(servers, pulled):
                    ps.write_model(rule(gradients=models, f=f))
                self.progress = 100.0 * (i + 1) / total
                self.events.put({"progress": self.progress})
            accs = [ps.compute_binary_accuracy() for ps in servers[f:]]
            self.result = sum(accs) / len(accs)
            self.events.put({"result": self.result})
        except Exception as err:  # reported through /status
            self.error = "%s: %s" % (type(err).__name__, err)

    def status(self):
        base = {"n": self.n, "f": self.f, "gar": self.gar, "transport": TRANSPORT[0], "dataset": pima_source()}
        if self.error:
            return dict(base, error=self.error)
        if self.result is not None:
            return dict(base, result=self.result, progress=100.0)
        return dict(base, progress=round(self.progress, 2))


class Trainers:
    def __init__(self):
        self._lock, self._items, self._next = threading.Lock(), {}, 0

    def submit(self, n, f, gar):
        import aggregators
        if not (1 <= n <= MAX_NODES) or f < 0 or 2 * f >= n:
            raise ValueError("need 1 <= n <= %d and 2f < n" % MAX_NODES)
        if gar not in aggregators.gars:
            raise ValueError("unknown GAR %r" % gar)
        probe = [__import__("torch").zeros(1) for _ in range(n - f)]
        problem = aggregators.gars[gar].check(gradients=probe, f=f) if f > 0 or gar in ("average", "median") else None
        if problem:
            raise ValueError(problem)
        with self._lock:
            self._next += 1
            trainer = Trainer(self._next, n, f, gar)
            self._items[trainer.id] = trainer
        return trainer.start()

    def get(self, trainer_id):
        with self._lock:
            return self._items.get(trainer_id)


TRAINERS = Trainers()


class Handler(BaseHTTPRequestHandler):
    def _send(self, code, body, ctype="application/json"):
        data = body.encode() if isinstance(body, str) else json.dumps(body).encode()
        self.send_response(code)
        self.send_header("Content-Type", ctype)
        self.send_header("Content-Length", str(len(data)))
        self.end_headers()
        self.wfile.write(data)

    def log_message(self, fmt, *args):
        pass

    def do_GET(self):
        url = urllib.parse.urlparse(self.path)
        if url.path in ("/", "/index.html"):
            return self._send(200, PAGE.read_text(), "text/html; charset=utf-8")
        if url.path == "/status":
            query = urllib.parse.parse_qs(url.query)
            try:
                trainer = TRAINERS.get(int(query.get("trainer_id", ["-1"])[0]))
            except ValueError:
                trainer = None
            if trainer is None:
                return self._send(404, {"error": "unknown trainer_id"})
            return self._send(200, trainer.status())
        self._send(404, {"error": "not found"})

    def do_POST(self):
        length = int(self.headers.get("Content-Length", "0"))
        raw = self.rfile.read(length).decode()
        try:
            fields = json.loads(raw) if raw.lstrip().startswith("{") else {k: v[0] for k, v in urllib.parse.parse_qs(raw).items()}
            trainer = TRAINERS.submit(int(fields.get("n", 5)), int(fields.get("f", 1)), str(fields.get("gar", "median")))
        except (ValueError, KeyError) as err:
            return self._send(400, {"error": str(err)})
        self._send(200, {"trainer_id": trainer.id})


def init_demo():
    """Pre-build the native libraries and touch the dataset (the reference's `quart init_demo`)."""
    import native
    from garfield_b200.garfieldpp.datasets import PimaDiabetesDataset
    print("natives available:", native.available(), "| pima train rows:", len(PimaDiabetesDataset(train=True)))


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=8080)
    ap.add_argument("--init_demo", action="store_true")
    ap.add_argument("--transport", default="threads", choices=["threads", "processes"],
                    help="processes: n local processes over torch RPC (the reference's way); threads: in-process role objects")
    args = ap.parse_args(argv)
    TRANSPORT[0] = args.transport
    if args.init_demo:
        return init_demo()
    server = ThreadingHTTPServer((args.host, args.port), Handler)
    print("serving on http://%s:%d" % (args.host, args.port))
    try:
        server.serve_forever()
    except KeyboardInterrupt:
        server.shutdown()


if __name__ == "__main__":
    main()
