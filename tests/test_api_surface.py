"""every public name of the reference is importable from the repository root with the same shape."""
import inspect


def test_public_api_surface_of_the_reference():
    import aggregators
    import native
    from garfieldpp import datasets, tools
    from garfieldpp.byzServer import ByzServer
    from garfieldpp.byzWorker import ByzWorker
    from garfieldpp.server import Server
    from garfieldpp.worker import Worker

    for name in ("average median krum bulyan brute aksel condense native-median native-krum native-bulyan native-brute "
                 "trimmed_mean averaged_median mda").split():
        assert name in aggregators.gars, name
    for name in ("average", "median", "krum", "bulyan", "brute", "aksel", "condense"):
        rule = aggregators.gars[name]
        assert getattr(aggregators, name) is rule
        for attr in ("check", "checked", "unchecked"):
            assert callable(getattr(rule, attr)), (name, attr)
    assert callable(aggregators.register) and callable(aggregators.make_gar)
    for module in ("krum", "bulyan", "median", "brute"):
        assert callable(getattr(native, module).aggregate)

    for method in ("get_gradients", "get_models", "get_aggr_grads", "get_latest_aggr_grad", "get_model", "update_model", "write_model",
                   "build_graph", "compute_accuracy", "compute_binary_accuracy"):
        assert callable(getattr(Server, method)), method
    server_args = list(inspect.signature(Server.__init__).parameters)[1:14]
    assert server_args == ["rank", "world_size", "num_workers", "num_ps", "byz_wrk", "byz_ps", "wrk_base_name", "ps_base_name", "batch", "model",
                           "dataset", "optimizer", "train_size"]
    assert list(inspect.signature(Worker.__init__).parameters)[1:9] == ["rank", "world_size", "num_workers", "batch_size", "model", "dataset",
                                                                        "loss", "train_size"]
    assert "attack" in inspect.signature(ByzWorker.__init__).parameters and "fw" in inspect.signature(ByzWorker.__init__).parameters
    assert "attack" in inspect.signature(ByzServer.__init__).parameters
    assert callable(Worker.compute_gradients)

    for fn in ("select_loss", "select_model", "select_optimizer", "adjust_learning_rate", "get_bytes_com", "convert_to_gbit", "get_server", "get_worker"):
        assert callable(getattr(tools, fn)), fn
    for model in ("convnet cifarnet cnn resnet18 resnet34 resnet50 resnet152 inception vgg16 vgg19 preactresnet18 googlenet densenet121 resnext29 "
                  "mobilenet mobilenetv2 dpn92 shufflenetg2 senet18 efficientnetb0 regnetx200 pimanet").split():
        assert model in tools.model_names(), model
    for dataset, classes in (("mnist", 10), ("cifar10", 10), ("cifar100", 100), ("imagenet", 1000), ("pima", 1)):
        assert tools.num_classes_of(dataset) == classes
    for loss in ("nll", "cross-entropy", "binary-cross-entropy"):
        assert tools.select_loss(loss) is not None
    assert list(inspect.signature(datasets.DatasetManager.__init__).parameters)[1:] == ["dataset", "minibatch", "num_workers", "size", "rank", "train_size"]
    for method in ("fetch_dataset", "get_train_set", "get_test_set"):
        assert callable(getattr(datasets.DatasetManager, method))
