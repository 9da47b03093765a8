"""Role objects (Server / Worker / ByzWorker / ByzServer) in the RPC-free local mode, flat arenas,
checkpoints and the trainers' loops — all on CPU with synthetic data."""

import os

import pytest
import torch

os.environ.setdefault("GARFIELD_SYNTHETIC", "1")
os.environ.setdefault("GARFIELD_SYNTHETIC_SIZE", "1200")

import aggregators  # noqa: E402
from garfield_b200.parallel.arena import FlatModel, GradMatrix  # noqa: E402
from garfield_b200.utils.checkpoint import Checkpoints, load_checkpoint, save_checkpoint  # noqa: E402
from garfieldpp import tools  # noqa: E402
from garfieldpp.byzServer import ByzServer  # noqa: E402
from garfieldpp.byzWorker import ByzWorker  # noqa: E402
from garfieldpp.server import Server  # noqa: E402
from garfieldpp.worker import Worker  # noqa: E402


def deployment(n=7, f=1, attack=None, model="convnet", dataset="mnist", loss="nll", num_ps=1, server_attack=None, **opt):
    opt = opt or dict(lr=0.05, momentum=0.9)
    workers = []
    for i in range(n):
        if attack and i >= n - f:
            workers.append(ByzWorker(num_ps + i, n + num_ps, n, 16, model, dataset, loss, attack, fw=f, device="cpu"))
        else:
            workers.append(Worker(num_ps + i, n + num_ps, n, 16, model, dataset, loss, device="cpu"))
    servers = []
    for r in range(num_ps):
        torch.manual_seed(1234)
        if server_attack and r == num_ps - 1:
            ps = ByzServer(r, 1, n, num_ps, f, 1, "worker:", "ps:", 16, model, dataset, "sgd", server_attack, device="cpu", **opt)
        else:
            ps = Server(r, 1, n, num_ps, f, 1 if server_attack else 0, "worker:", "ps:", 16, model, dataset, "sgd", device="cpu", **opt)
        ps.attach_workers(workers)
        servers.append(ps)
    for ps in servers:
        ps.attach_servers(servers)
    return servers, workers


def test_flat_model_aliases_parameters_and_state_dict_roundtrip():
    module = tools.select_model("cifarnet", torch.device("cpu"), "cifar10")
    reference = [p.detach().clone() for p in module.parameters()]
    fm = FlatModel(module)
    assert fm.numel == 62006 == fm.flat.numel()
    torch.testing.assert_close(torch.cat([r.reshape(-1) for r in reference]), fm.flat)
    fm.flat.mul_(2.0)  # parameters are views: the module sees the change
    torch.testing.assert_close(next(module.parameters()), reference[0] * 2)
    sd = {k: v.clone() for k, v in fm.state_dict().items()}
    fm.flat.zero_()
    fm.load_state_dict(sd)
    torch.testing.assert_close(next(module.parameters()), reference[0] * 2)
    assert next(module.parameters()).data_ptr() == fm.flat.data_ptr()  # aliasing survived the load
    gm = GradMatrix(3, fm.numel, "cpu")
    fm.bind_grads(gm.row(1))
    module(torch.randn(4, 3, 32, 32)).sum().backward()
    assert gm.row(1).abs().sum() > 0 and gm.row(0).abs().sum() == 0
    assert gm.row(2).data_ptr() % 256 == gm.row(0).data_ptr() % 256


def test_worker_matches_reference_semantics():
    """compute_gradients == flat cat of param.grad after one forward/backward on batch iter % N."""
    wrk = Worker(1, 2, 1, 8, "cifarnet", "cifar10", "cross-entropy", device="cpu")
    module = tools.select_model("cifarnet", torch.device("cpu"), "cifar10")
    rank, grad, loss = wrk.compute_gradients(3, module)
    assert rank == 1 and grad.shape == (62006,) and isinstance(loss, float)
    data, target = wrk.train_set[3 % wrk.num_train_samples]
    module.zero_grad()
    out = torch.nn.CrossEntropyLoss()(module(data), target)
    out.backward()
    want = torch.cat([p.grad.reshape(-1) for p in module.parameters()])
    torch.testing.assert_close(grad, want)
    assert loss == pytest.approx(out.item())


def test_update_and_write_model_match_torch_sgd():
    (ps,), workers = deployment(n=3, f=0, lr=0.1, momentum=0.9, weight_decay=1e-3)
    twin = tools.select_model("convnet", torch.device("cpu"), "mnist")
    twin.load_state_dict(ps.model.state_dict())
    ref_opt = torch.optim.SGD(twin.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-3)
    for step in range(3):
        g = torch.randn(ps.flat_model.numel)
        ps.update_model(g)
        for p, view in zip(twin.parameters(), ps.flat_model.views_of(g)):
            p.grad = view.clone()
        ref_opt.step()
    torch.testing.assert_close(ps.get_model()[1], torch.cat([p.detach().reshape(-1) for p in twin.parameters()]), rtol=1e-5, atol=1e-6)
    rank, flat = ps.get_model()
    assert rank == 0 and flat.data_ptr() != ps.flat_model.flat.data_ptr()
    ps.write_model(torch.zeros_like(flat))
    assert all(float(p.detach().abs().sum()) == 0 for p in ps.model.parameters())


@pytest.mark.parametrize("gar,attack", [("krum", "reverse"), ("median", "random"), ("bulyan", "lie"), ("brute", "empire"),
                                         ("trimmed_mean", "drop")])
def test_ssmw_trains_under_attack(gar, attack):
    (ps,), workers = deployment(n=7 if gar != "bulyan" else 8, f=1, attack=attack)
    rule = aggregators.gars[gar]
    n = len(workers)
    first = last = None
    for i in range(40):
        grads = ps.get_gradients(i, n)
        assert len(grads) == n and all(g.shape == (ps.flat_model.numel,) for g in grads)
        ps.update_model(rule(gradients=grads, f=1))
        honest = sum(ps.last_losses[:-1]) / (n - 1)
        first = honest if first is None else first
        last = honest
    assert torch.isfinite(ps.flat_model.flat).all()
    assert last < first, (first, last)


def test_quorum_returns_n_minus_f_gradients():
    (ps,), workers = deployment(n=7, f=2)
    grads = ps.get_gradients(0)  # default: n - byz_wrk
    assert len(grads) == 5 and ps.last_contributors == [w.rank for w in workers[:5]]
    assert len(ps.get_gradients(0, 7)) == 7


def test_byzantine_attacks_are_applied():
    (ps,), workers = deployment(n=5, f=1, attack="reverse", model="cifarnet", dataset="cifar10", loss="cross-entropy")
    honest_twin = Worker(workers[-1].rank, 6, 5, 16, "cifarnet", "cifar10", "cross-entropy", device="cpu")
    grads = ps.get_gradients(0, 5)
    _, honest, _ = honest_twin.compute_gradients(0, ps.flat_model)
    torch.testing.assert_close(grads[-1], honest * -100)
    with pytest.raises(AssertionError):
        ByzWorker(1, 2, 1, 8, "cifarnet", "cifar10", "cross-entropy", "nonsense", device="cpu")


def test_msmw_replicas_agree_despite_byzantine_server():
    servers, workers = deployment(n=5, f=1, num_ps=3, server_attack="reverse")
    rule = aggregators.gars["median"]
    honest = [ps for ps in servers if not isinstance(ps, ByzServer)]
    for i in range(3):
        for ps in servers:
            ps.update_model(rule(gradients=ps.get_gradients(i, 5), f=1))
        for ps in honest:
            models = ps.get_models(3)
            assert len(models) == 3
            ps.write_model(rule(gradients=models, f=1))
    torch.testing.assert_close(honest[0].get_model()[1], honest[1].get_model()[1])
    rank, lied = servers[-1].get_model()
    assert rank == 2 and not torch.allclose(lied, honest[0].get_model()[1])


def test_latest_aggr_grad_handshake():
    servers, _ = deployment(n=3, f=0, num_ps=2)
    import threading
    out = {}
    th = threading.Thread(target=lambda: out.setdefault("v", servers[0].get_latest_aggr_grad(timeout=5)))
    th.start()
    servers[0].latest_aggr_grad = torch.ones(4)
    th.join()
    assert out["v"][0] == 0 and torch.equal(out["v"][1], torch.ones(4))
    servers[1].latest_aggr_grad = torch.zeros(4)
    assert len(servers[0].get_aggr_grads(2)) == 2


def test_accuracy_and_binary_accuracy():
    (ps,), _ = deployment(n=2, f=0)
    acc = ps.compute_accuracy()
    assert 0.0 <= acc <= 100.0 and ps.model.training in (True, False)
    pima = Server(0, 1, 1, 1, 0, 0, " ", " ", 16, "pimanet", "pima", "rmsprop", device="cpu", lr=0.001)
    assert 0.0 <= pima.compute_binary_accuracy() <= 100.0
    assert len(pima.test_set) == 2 and pima.test_set[0][0].shape == (100, 8)


def test_checkpoint_resume_is_bit_exact(tmp_path):
    def run(steps, ps, workers, start=0):
        rule = aggregators.gars["krum"]
        for i in range(start, steps):
            ps.update_model(rule(gradients=ps.get_gradients(i, len(workers)), f=1))
    torch.manual_seed(7)
    (a,), wa = deployment(n=5, f=1, model="cifarnet", dataset="cifar10", loss="cross-entropy")
    run(3, a, wa)
    path = save_checkpoint(tmp_path / "ck.pt", a.flat_model, a.optimizer, step=3, config={"gar": "krum"})
    run(6, a, wa, start=3)
    (b,), wb = deployment(n=5, f=1, model="cifarnet", dataset="cifar10", loss="cross-entropy")
    payload = load_checkpoint(path, b.flat_model, b.optimizer)
    assert payload["step"] == 3 and payload["config"]["gar"] == "krum"
    run(6, b, wb, start=3)
    assert torch.equal(a.flat_model.flat, b.flat_model.flat)
    manager = Checkpoints(tmp_path / "dir", keep=2)
    for step in (1, 2, 3):
        manager.save(b.flat_model, b.optimizer, step)
    assert [s for s, _ in manager.list()] == [2, 3] and manager.latest()[0] == 3
    assert manager.restore(a.flat_model, a.optimizer)["step"] == 3


def test_select_helpers():
    assert isinstance(tools.select_loss("nll"), torch.nn.NLLLoss)
    with pytest.raises(KeyError):
        tools.select_loss("hinge")
    for name, count in (("convnet", 21840), ("cifarnet", 62006), ("pimanet", 4801)):
        ds = {"convnet": "mnist", "cifarnet": "cifar10", "pimanet": "pima"}[name]
        assert sum(p.numel() for p in tools.select_model(name, torch.device("cpu"), ds).parameters()) == count
    model = tools.select_model("cifarnet", torch.device("cpu"), "cifar10")
    assert isinstance(tools.select_optimizer(model, "adam", lr=1e-3), torch.optim.Adam)
    flat = tools.select_optimizer(FlatModel(model), "sgd", lr=0.3)
    tools.adjust_learning_rate(flat, 0.01)
    assert flat.param_groups[0]["lr"] == 0.01
    assert tools.convert_to_gbit(1024 ** 3) == 8.0 and tools.get_bytes_com() >= 0
    assert set(("convnet", "cifarnet", "cnn", "resnet18", "resnet50", "vgg16", "pimanet")) <= set(tools.model_names())


def test_trainers_run_end_to_end(tmp_path):
    from garfield_b200.apps import learn, ssmw
    ck = tmp_path / "agg.pt"
    ps = ssmw.main(["--dataset", "mnist", "--model", "convnet", "--num_workers", "7", "--fw", "1", "--gar", "krum", "--num_iter", "4",
                    "--attack", "reverse", "--device", "cpu", "--checkpoint", str(ck), "--trace", str(tmp_path / "t.json")])
    assert ck.exists() and (tmp_path / "t.json").exists()
    resumed = ssmw.main(["--dataset", "mnist", "--model", "convnet", "--num_workers", "7", "--fw", "1", "--gar", "krum", "--num_iter", "4",
                         "--device", "cpu", "--resume", str(ck)])
    assert torch.equal(ps.flat_model.flat, resumed.flat_model.flat)
    ssmw.main(["--dataset", "mnist", "--model", "convnet", "--num_workers", "5", "--num_ps", "3", "--fps", "1", "--fw", "1", "--gar",
               "median", "--num_iter", "3", "--server_attack", "random", "--device", "cpu"], msmw=True)
    learn.main(["--dataset", "mnist", "--model", "convnet", "--num_nodes", "6", "--f", "1", "--gar", "median", "--num_iter", "3",
                "--attack", "random", "--non_iid", "true", "--device", "cpu"])


def test_ssmw_trajectory_matches_plain_torch_simulator():
    """End to end: the Server/Worker loop follows the same parameter trajectory as a from-scratch simulation of
    the reference's step (§3.2) written with plain torch only: per-worker forward/backward on a copy of the current model,
    `torch.cat` of the gradients, `grad * -100` for the Byzantine worker, Multi-Krum restated inline, `optim.SGD` on the
    unflattened aggregate."""
    n, f, steps = 7, 1, 5
    (ps,), workers = deployment(n=n, f=f, attack="reverse", model="cifarnet", dataset="cifar10", loss="cross-entropy", lr=0.05, momentum=0.9,
                                  weight_decay=5e-4)     # a model without dropout: both sides must be deterministic
    torch.manual_seed(1234)
    model = tools.select_model("cifarnet", torch.device("cpu"), "cifar10")
    with torch.no_grad():
        for p, q in zip(model.parameters(), ps.flat_model.module.parameters()):
            p.copy_(q)
    sim_opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9, weight_decay=5e-4)
    loss_fn = torch.nn.CrossEntropyLoss()
    import copy
    for it in range(steps):
        grads = []
        for k, w in enumerate(workers):
            replica = copy.deepcopy(model)
            replica.train()
            data, target = w.train_set[it % len(w.train_set)]
            loss_fn(replica(data), target).backward()
            g = torch.cat([p.grad.reshape(-1) for p in replica.parameters()])
            grads.append(g * -100 if k >= n - f else g)
        # Multi-Krum, python semantics: distances (not squared), n-f-1 neighbours, m = n-f-2 best averaged
        dist = [[(grads[i] - grads[j]).norm().item() if i != j else float("inf") for j in range(n)] for i in range(n)]
        scores = [sum(sorted(d for j, d in enumerate(dist[i]) if j != i)[:n - f - 1]) for i in range(n)]
        best = sorted(range(n), key=lambda i: scores[i])[:n - f - 2]
        aggregate = sum(grads[i] for i in best) / len(best)
        at = 0
        for p in model.parameters():
            p.grad = aggregate[at:at + p.numel()].view_as(p).clone()
            at += p.numel()
        sim_opt.step()
        # the framework's step
        ps.update_model(aggregators.gars["krum"](gradients=ps.get_gradients(it, n), f=f))
    want = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    torch.testing.assert_close(ps.flat_model.flat, want, rtol=1e-4, atol=1e-6)
