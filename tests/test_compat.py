"""TF-side compatibility layer: numpy aggregator façade, numpy attacks, cluster JSON, gRPC pulls."""

import socket

import numpy as np
import pytest
import torch

from garfield_b200.compat import Aggregator_tf, Attacker, Network
from garfield_b200.compat import grpc_exchange
from garfield_b200.compat.aggregator_tf import average_nan


def test_aggregator_facade_matches_registry():
    import aggregators
    rng = np.random.RandomState(0)
    grads = [rng.randn(500).astype(np.float32) for _ in range(11)]
    rows = [torch.from_numpy(g) for g in grads]
    for agg, name in (("Average", "average"), ("Median", "median"), ("Krum", "krum"), ("Bulyan", "bulyan"), ("Brute", "brute"),
                      ("Aksel", "aksel")):
        got = Aggregator_tf(agg, 11, 2, device="cpu").aggregate(grads)
        np.testing.assert_allclose(got, aggregators.gars[name](gradients=rows, f=2).numpy(), rtol=1e-5, atol=1e-6)
    native = Aggregator_tf("Krum", 11, 2, native=True, device="cpu").aggregate(grads)
    np.testing.assert_allclose(native, aggregators.gars["native-krum"](gradients=rows, f=2).numpy(), rtol=1e-5, atol=1e-6)
    bad = [g.copy() for g in grads[:3]]
    bad[0][0] = np.nan
    assert average_nan(bad)[0] == pytest.approx((bad[1][0] + bad[2][0]) / 2)


def test_numpy_attacks():
    g = np.arange(8, dtype=np.float32)
    assert Attacker("Reverse").attack(gradient=g).tolist() == (-100 * g).tolist()
    assert Attacker("Random").attack(gradient=g).shape == g.shape
    dropped = Attacker("PartialDrop").attack(gradient=np.ones(10000, np.float32), probability=0.3)
    assert 0.25 < (dropped == 0).mean() < 0.35
    others = [g + 1, g + 2]
    stack = np.stack(others + [g])
    np.testing.assert_allclose(Attacker("LittleIsEnough").attack(gradient=g, byz_gradients=others), stack.mean(0) + 1.035 * stack.std(0), rtol=1e-6)
    np.testing.assert_allclose(Attacker("FallEmpires").attack(gradient=g, byz_gradients=others), -10 * stack.mean(0), rtol=1e-6)
    with pytest.raises(AssertionError):
        Attacker("Nope")


def test_network_config():
    cfg = {"cluster": {"ps": ["a:1000"], "worker": ["b:2000", "c:2001"]},
           "task": {"type": "worker", "index": 1, "strategy_gradient": "Krum", "strategy_model": "Median", "attack": "Random"}}
    net = Network(cfg)
    assert net.get_task_type() == "worker" and net.get_task_index() == 1 and net.get_my_node() == "c:2001"
    assert net.get_my_port() == 2001 and net.get_all_other_worker() == ["b:2000"] and net.get_all_ps() == ["a:1000"]
    assert net.get_gradient_strategy() == "Krum" and net.get_model_strategy() == "Median" and net.get_my_attack() == "Random"


def test_grpc_pull_service():
    pytest.importorskip("grpc")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    servicer = grpc_exchange.MessageExchangeServicer(timeout=10)
    server = grpc_exchange.serve(servicer, port)
    try:
        client = grpc_exchange.Client("127.0.0.1:%d" % port)
        model = np.linspace(0, 1, 1000, dtype=np.float32)
        servicer.commit_model(model)
        servicer.commit_gradient(model * 2)
        np.testing.assert_array_equal(client.get_model(0), model)
        np.testing.assert_array_equal(client.get_gradient(0), model * 2)
        import threading
        threading.Timer(0.2, lambda: servicer.commit_model(model + 1)).start()
        np.testing.assert_array_equal(client.get_model(1), model + 1)   # blocks until iteration 1 is committed
        client.close()
    finally:
        server.stop(0)


def test_tf_style_pull_deployment_in_threads():
    """1 PS + 4 workers (one Byzantine 'Reverse'), Krum gradients, pull-based gRPC on localhost."""
    pytest.importorskip("grpc")
    import importlib.util
    import os
    import pathlib
    import threading
    import types
    os.environ.setdefault("GARFIELD_SYNTHETIC", "1")
    os.environ.setdefault("GARFIELD_SYNTHETIC_SIZE", "600")
    root = pathlib.Path(__file__).resolve().parent.parent
    load = lambda name: importlib.util.spec_from_file_location(name, root / "applications" / "TF_style" / (name + ".py"))
    gen_spec = load("config_generator"); gen = importlib.util.module_from_spec(gen_spec); gen_spec.loader.exec_module(gen)
    tr_spec = load("trainer"); trainer = importlib.util.module_from_spec(tr_spec); tr_spec.loader.exec_module(trainer)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        base = s.getsockname()[1]
    configs = gen.generate(1, 5, base_port=base, strategy_gradient="Krum", byz_workers=1, attack="Reverse")
    results, errors = {}, []

    def run(name, cfg):
        args = types.SimpleNamespace(config=cfg, log=False, max_iter=3, dataset="mnist", model="convnet", loss="nll", batch_size=8,
                                     nbbyzwrks=1, native=False, device="cpu")
        try:
            results[name] = trainer.run_node(args)
        except Exception as err:  # surfaced below
            errors.append((name, repr(err)))

    threads = [threading.Thread(target=run, args=item) for item in configs.items()]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=180)
    assert not errors, errors
    assert len(results) == 6
    ps = results["ps0"]
    assert len(ps.servicer.model_history) == 4 and np.isfinite(ps.flat()).all()
    for node in results.values():
        node.stop()


def test_legacy_train_message_exchange_in_process_and_grpc():
    """Garfield_legacy/all.proto: every declared RPC answers; hashes and signatures verify; pulls block until published."""
    import threading
    import numpy as np
    from garfield_b200.compat import legacy_exchange as lx
    node = lx.TrainMessageExchangeServicer(index=3, secret=b"node-3", timeout=5.0)
    model = np.arange(10, dtype=np.float32)
    grad = np.linspace(-1, 1, 10).astype(np.float32)
    threading.Timer(0.05, lambda: node.publish_model(2, model, inputs=np.ones((2, 3), np.float32), labels=np.array([1, 0]), correct=[0, 2])).start()
    threading.Timer(0.05, lambda: node.publish_gradient(2, grad, lipschitz=0.25)).start()
    server = lx.serve_legacy(node, 0)
    try:
        for client in (lx.LegacyClient(node), lx.LegacyClient("127.0.0.1:%d" % server.bound_port)):
            index, pub = client.get_public_key()
            assert index == 3 and len(pub) == 32
            it, got = client.get_model(2)
            assert it == 2 and np.array_equal(got, model)
            assert np.array_equal(client.get_unified_model()[1], model)
            full = client.get_complete_model(1)
            assert full["iter"] == 2 and full["correct"] == [0, 2] and full["msg_hash"] == lx.model_hash(model)
            assert lx.Signer.verify(pub, full["msg_hash"], full["signature"])
            assert not lx.Signer.verify(pub, lx.model_hash(model + 1), full["signature"])
            _, digest, sig = client.get_only_hash(2)
            assert digest == lx.model_hash(model) and lx.Signer.verify(pub, digest, sig)
            it, g, lip = client.get_gradients(2)
            assert it == 2 and np.array_equal(g, grad) and abs(lip - 0.25) < 1e-7
            assert client.get_grad_hash(2) == (3, lx.model_hash(grad))
            node.record_peer_hash(2, 5, b"\\x01" * 32)
            hashes = client.get_grad_hashes(2)
            assert hashes[3] == lx.model_hash(grad) and 5 in hashes
            client.close()
    finally:
        server.stop(0)


def test_legacy_lipschitz_filter_and_update_model():
    import numpy as np
    import torch
    from garfield_b200.compat import legacy_exchange as lx
    flt = lx.LipschitzFilter(num_ps=4, num_byzps=1)
    rng = np.random.default_rng(0)
    for _ in range(120):
        ok, bound = flt.observe(abs(rng.normal()) + 0.1)
        assert bound > 0
    assert len(flt.history) <= 100
    assert flt.observe(1e9)[0] is False and flt.observe(0.0)[0] is True
    x0, x1 = np.zeros(4), np.array([3.0, 4.0, 0.0, 0.0])
    g0, g1 = np.zeros(4), np.array([0.0, 0.0, 6.0, 8.0])
    assert abs(lx.LipschitzFilter.coefficient(g1, g0, x1, x0) - 2.0) < 1e-12
    assert lx.drift_bound(0.1, 2.0, 10, 8, 2, 5) == 0.1 * 2.0 * ((32 * 6) / 8.0 + 8)
    assert lx.drift_bound(0.1, 2.0, 10, 8, 0, 5) == float("inf")
    m, g = np.ones(5, np.float32), np.full(5, 2.0, np.float32)
    assert np.allclose(lx.update_model(m, g), 0.8)
    assert torch.allclose(lx.update_model(torch.ones(5), torch.full((5,), 2.0)), torch.full((5,), 0.8))


def test_tf_model_manager_names():
    """tensorflow_impl/libs/model.py names resolve to torch modules producing [batch, classes] logits."""
    import torch
    from garfield_b200.compat import ModelManager
    info = {"num_classes": 7, "image_shape": (3, 32, 32)}
    for name in ("Small", "CNN", "Cifarnet", "MobileNetV2"):
        net = ModelManager(name, info).model
        net.eval()
        with torch.no_grad():
            assert net(torch.randn(2, 3, 32, 32)).shape == (2, 7), name
    assert ModelManager("Small", {"num_classes": 10, "image_shape": (1, 28, 28)}).model(torch.randn(3, 1, 28, 28)).shape == (3, 10)
    assert set(ModelManager.NAMES) == {"Small", "MobileNetV2", "Resnet50", "Resnet200", "VGG", "DenseNet", "Cifarnet", "Inception", "CNN"}
    import pytest
    with pytest.raises(AssertionError):
        ModelManager("Transformer", info)
