"""The `tools` toolbox and the auxiliary subsystems of §5 that run without a GPU:
registries, key:value parsing, timed contexts, cluster specs, access checks, flat-tensor helpers, tracing and metrics."""
import json
import os

import pytest
import torch

import tools
from garfield_b200.utils import access, cluster, flat, misc, tracing


def test_class_register_and_method_call_replicator():
    reg = misc.ClassRegister("optimizer")
    reg.register("sgd", dict)
    assert list(reg.itemize()) == ["sgd"] and reg.instantiate("sgd", lr=1) == {"lr": 1}
    with pytest.raises(AssertionError):
        reg.register("sgd", list)
    with pytest.raises(tools.UserException) as err:
        reg.instantiate("adam")
    assert "available optimizer(s): 'sgd'" in str(err.value)
    with pytest.raises(tools.UserException) as err:
        misc.ClassRegister("rule").instantiate("x")
    assert "no registered rule" in str(err.value)
    both = misc.MethodCallReplicator([1, 2, 2], [2, 2])
    assert both.count(2) == [2, 2]
    with pytest.raises(AssertionError):
        misc.MethodCallReplicator()


def test_parse_keyval_types_defaults_and_errors():
    got = misc.parse_keyval(["lr:0.1", "nesterov:true", "name:adam", "steps:3"])
    assert got == {"lr": 0.1, "nesterov": True, "name": "adam", "steps": 3}
    got = misc.parse_keyval(["lr:1", "flag:0"], defaults={"lr": 0.5, "flag": True, "momentum": 0.9})
    assert got == {"lr": 1.0, "flag": False, "momentum": 0.9} and isinstance(got["lr"], float)
    for bad in (["novalue"], ["a:1", "a:2"]):
        with pytest.raises(tools.UserException):
            misc.parse_keyval(bad)
    with pytest.raises(tools.UserException):
        misc.parse_keyval(["lr:abc"], defaults={"lr": 0.1})


def test_misc_helpers():
    assert list(misc.pairwise([1, 2, 3])) == [(1, 2), (1, 3), (2, 3)]
    assert misc.fullqual(dict).endswith("dict") and misc.fullqual({}).startswith("instance of")
    best = misc.line_maximize(lambda x: -(x - 3.0) ** 2, evals=40, start=0.0, delta=1.0)
    assert abs(best - 3.0) < 0.2
    is_set, set_it = misc.onetime("probe")
    assert is_set() is False
    set_it()
    assert is_set() is True and misc.onetime("probe")[0]() is True      # named flags are shared
    assert misc.onetime()[0]() is False                                  # anonymous ones are private
    with misc.TimedContext("block", "info"):
        pass
    assert isinstance(misc.deltatime_format(misc.deltatime_point(), misc.deltatime_point()), str)
    deps = misc.get_loaded_dependencies()
    assert any("torch" in str(d) for d in deps)


def test_cluster_parsers(tmp_path, monkeypatch):
    spec = {"ps": ["a:1"], "workers": ["b:2", "c:3"]}
    assert cluster.cluster_parse(json.dumps(spec)) == spec
    monkeypatch.setenv("MASTER_ADDR", "10.0.0.1")
    monkeypatch.setenv("MASTER_PORT", "29000")
    monkeypatch.setenv("WORLD_SIZE", "3")
    got = cluster.cluster_parse("torchrun")
    assert got["ps"] == ["10.0.0.1:29000"] and got["workers"] == ["10.0.0.1:29001", "10.0.0.1:29002", "10.0.0.1:29003"]
    nodes = tmp_path / "oar"
    nodes.write_text("n1\nn1\nn2\nn3\nn3\n")
    monkeypatch.setenv("OAR_FILE_NODES", str(nodes))
    cluster._cache.clear()
    assert cluster.cluster_parse("G5k") == {"ps": ["n1:7000"], "workers": ["n2:7000", "n3:7000"]}


def test_can_access(tmp_path):
    f = tmp_path / "d" / "x.txt"
    f.parent.mkdir()
    f.write_text("x")
    assert access.can_access(f, read=True) and access.can_access(tmp_path, read=True, write=True, recurse=True)
    assert not access.can_access(tmp_path / "missing", read=True)


def test_flat_helpers():
    a, b = torch.arange(6.0).reshape(2, 3), torch.arange(4.0)
    common = flat.flatten([a, b])
    assert common.numel() == 10
    common[0] = 42.0
    assert a[0, 0].item() == 42.0                      # relinked: the tensors are views of the common buffer
    p = torch.nn.Parameter(torch.ones(3))
    assert torch.equal(flat.grad_of(p), torch.zeros(3)) and p.grad is not None
    samples = [torch.tensor([1.0, 0.0]), torch.tensor([3.0, 0.0]), torch.tensor([2.0, 6.0])]
    mean, norm, dev, top = flat.compute_avg_dev_max(samples)
    assert torch.equal(mean, torch.tensor([2.0, 2.0])) and abs(norm - 8 ** 0.5) < 1e-6 and top == 2.0
    assert abs(dev - ((5 + 5 + 16) / 2) ** 0.5) < 1e-6               # sqrt(sum ||g - mean||^2 / (k - 1))
    assert flat.compute_avg_dev_max([])[0] is None
    import io
    buf = io.BytesIO()
    flat.pnm(buf, torch.zeros(2, 3))
    assert buf.getvalue().startswith(b"P5\n2 3 255\n") and len(buf.getvalue()) == len(b"P5\n2 3 255\n") + 6


def test_tracer_and_metrics(tmp_path):
    tracer = tracing.Tracer(enabled=True, device="cpu")
    for step in range(3):
        with tracer.phase("gradients", step):
            sum(range(1000))
        with tracer.phase("gar", step):
            pass
    stats = tracer.summary()
    assert stats["gradients"]["count"] == 3 and stats["gar"]["mean_ms"] >= 0.0
    path = tracer.dump_chrome_trace(str(tmp_path / "trace.json"))
    events = json.load(open(path))["traceEvents"]
    assert len(events) == 6 and {e["name"] for e in events} == {"gradients", "gar"} and all(e["ph"] == "X" for e in events)
    off = tracing.Tracer(enabled=False)
    with off.phase("x"):
        pass
    assert off.summary() == {}
    metrics = tracing.Metrics(str(tmp_path / "m.jsonl"))
    metrics.log(step=1, loss=0.5)
    metrics.log(step=2, loss=0.4)
    rows = [json.loads(line) for line in open(tmp_path / "m.jsonl")]
    assert [r["step"] for r in rows] == [1, 2] and "time" in rows[0]


def test_trace_function_survives_importing_the_tracing_module():
    """`tools.trace` is the console function whatever the import order (the tracing module has its own name)."""
    import importlib
    importlib.import_module("garfield_b200.utils.tracing")
    import garfield_b200.utils as utils
    assert callable(utils.trace) and callable(tools.trace)


def test_console_contexts_prefix_output(capsys):
    with tools.Context("train", "info"):
        with tools.Context("step", None):
            header = tools.Context._get()[0]
            print("hello")
    assert "train" in header and "step" in header and header.index("train") < header.index("step")     # outermost first
    assert "hello" in capsys.readouterr().out
    assert tools.Context._get()[0] == "" or "train" not in tools.Context._get()[0]
    with pytest.raises(AssertionError):
        tools.Context("x", "no-such-colour")
    with pytest.raises(tools.UserException):
        raise tools.UserException("boom")


def test_application_flag_system(capsys):
    """§5.6: the trainers' shared parser — reference flag names, real booleans (the reference's `type=bool` is truthy for
    any non-empty string), `--opt_args` JSON with float casting, sanity asserts."""
    from garfield_b200.apps import common
    ap = common.base_parser("t")
    args = common.finish_args(ap.parse_args(["--dataset", "cifar10", "--batch", "25", "--num_workers", "8", "--fw", "1", "--gar", "krum",
                                            "--opt_args", '{"lr":"0.2","momentum":"0.9","weight_decay":"0.0005"}', "--bench", "False", "--log", "true"]))
    assert args.opt_args == {"lr": 0.2, "momentum": 0.9, "weight_decay": 0.0005}
    assert args.bench is False and args.log is True and args.device in ("cpu", "cuda")
    assert common.str2bool("0") is False and common.str2bool("no") is False and common.str2bool("yes") is True and common.str2bool(True) is True
    dec = common.base_parser("t", decentralized=True).parse_args(["--num_nodes", "6", "--f", "1", "--non_iid", "true"])
    assert dec.num_nodes == 6 and dec.non_iid is True and not hasattr(dec, "num_ps")
    with pytest.raises(AssertionError):
        common.finish_args(ap.parse_args(["--opt_args", '{"momentum":"0.9"}']))
    assert common.pick_gar("median") is __import__("aggregators").gars["median"]
    with pytest.raises(AssertionError):
        common.pick_gar("nope")
    assert common.iterations_per_epoch(8, 25) == 250
    common.banner(args, 3)
    out = capsys.readouterr().out
    assert "SETUP AT NODE 3" in out and "gar:" in out
    timer = common.StepTimer(True, "cpu")
    assert timer.run(0, lambda: 5) == 5
    assert "Training step 0 takes" in capsys.readouterr().out


def test_spmd_transport_wires_the_device_engines(monkeypatch):
    """`--transport spmd` hands the trainers' flags to the right device engine (checked with stand-in engines: the real
    ones need a GPU and are covered by the GPU tests)."""
    import torch
    from garfield_b200.apps import learn, ssmw
    from garfield_b200.parallel import engine as engine_mod, topologies
    made = []

    class Fake:
        rank = 0

        def __init__(self, *args, **kwargs):
            made.append((type(self).__name__, args, kwargs))
            self.steps = 0

        def step(self, i):
            self.steps += 1
            return torch.zeros(())

        def close(self):
            made.append(("closed", self.steps))

    monkeypatch.setattr(engine_mod, "SSMWEngine", type("SSMWEngine", (Fake,), {}))
    monkeypatch.setattr(topologies, "MSMWEngine", type("MSMWEngine", (Fake,), {}))
    monkeypatch.setattr(topologies, "DecentralizedEngine", type("DecentralizedEngine", (Fake,), {}))
    monkeypatch.setattr(torch.cuda, "set_device", lambda index: None)
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a: None)
    monkeypatch.delenv("WORLD_SIZE", raising=False)
    common_flags = ["--dataset", "cifar10", "--model", "resnet50", "--loss", "cross-entropy", "--batch", "25", "--gar", "krum", "--num_iter", "3",
                    "--transport", "spmd", "--opt_args", '{"lr":"0.2","momentum":"0.9"}']
    ssmw.main(common_flags + ["--num_workers", "8", "--fw", "1", "--attack", "reverse"], msmw=False)
    name, args, kwargs = made[0]
    assert name == "SSMWEngine" and args[:7] == ("resnet50", "cifar10", "cross-entropy", 8, 1, "krum", 25)
    assert args[8] == {"lr": 0.2, "momentum": 0.9} and kwargs["byzantine"] == {7: "reverse"} and made[1] == ("closed", 3)
    made.clear()
    ssmw.main(common_flags + ["--num_workers", "5", "--fw", "1", "--num_ps", "3", "--fps", "1", "--attack", "reverse", "--server_attack", "random"],
              msmw=True)
    name, args, kwargs = made[0]
    assert name == "MSMWEngine" and args == ("resnet50", "cifar10", "cross-entropy", 5, 3, 1, 1, "krum", 25)
    assert kwargs["mar"] == "krum" and kwargs["byzantine_workers"] == {4: "reverse"} and kwargs["byzantine_servers"] == {2: "random"}
    made.clear()
    ssmw.main(common_flags + ["--num_workers", "5", "--fw", "1"], msmw=True)          # one server: ByzSGD degenerates to SSMW
    assert made[0][0] == "SSMWEngine" and made[0][2]["byzantine"] == {}
    made.clear()
    learn.main(["--dataset", "cifar10", "--model", "vgg16", "--loss", "cross-entropy", "--batch", "25", "--gar", "median", "--num_iter", "2",
                "--transport", "spmd", "--num_nodes", "8", "--f", "1", "--non_iid", "true", "--attack", "lie"])
    name, args, kwargs = made[0]
    assert name == "DecentralizedEngine" and args == ("vgg16", "cifar10", "cross-entropy", 8, 1, "median", 25)
    assert kwargs["byzantine"] == {7: "lie"} and kwargs["non_iid"] is True and made[1] == ("closed", 2)


def test_plugin_loader_and_interactive_console(tmp_path, monkeypatch, capsys):
    """`tools.import_directory` (the mechanism behind the auto-importing registries) loads every module of a directory,
    exports `__all__`, and survives a broken plugin; `tools.interactive` is a REPL over the caller's scope."""
    (tmp_path / "alpha.py").write_text("__all__ = ['answer']\nanswer = 42\nhidden = 1\n")
    (tmp_path / "beta.py").write_text("def shout():\n    return 'b'\n_private = 0\n")
    (tmp_path / "broken.py").write_text("raise RuntimeError('boom')\n")
    (tmp_path / "__init__.py").write_text("")
    (tmp_path / "notes.txt").write_text("ignored")
    scope = {"__name__": "plugins_under_test"}
    tools.import_directory(tmp_path, scope)
    assert scope["answer"] == 42 and "hidden" not in scope and scope["shout"]() == "b" and "_private" not in scope
    capsys.readouterr()                                                              # the failure is reported on the console, not raised
    import sys
    assert "plugins_under_test.alpha" in sys.modules and "plugins_under_test.broken" not in sys.modules
    for name in ("plugins_under_test.alpha", "plugins_under_test.beta"):
        sys.modules.pop(name, None)

    lines = iter(["result.append(seed + 1)"])

    def fake_input(prompt=""):
        try:
            return next(lines)
        except StopIteration:
            raise EOFError
    monkeypatch.setattr("builtins.input", fake_input)
    result, seed = [], 6
    misc.interactive()
    assert result == [7]
