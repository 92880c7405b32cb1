"""Optimizers vs closed form, schedulers, losses, env config, fused-vs-autograd loss gradients -- all untested in the
reference."""
import math
import os

import pytest
import torch

from dcnn_b200 import nn as dnn
from dcnn_b200.utils import env


def test_sgd_and_momentum_closed_form():
    p, g = torch.tensor([1.0, -2.0]), torch.tensor([0.5, 0.25])
    o = dnn.SGD(0.1)
    o.attach([p], [g])
    o.update()
    assert torch.allclose(p, torch.tensor([0.95, -2.025]))
    p2, g2 = torch.tensor([1.0]), torch.tensor([1.0])
    o = dnn.SGD(0.1, 0.9)
    o.attach([p2], [g2])
    o.update()          # v = -0.1 ; p = 0.9
    o.update()          # v = 0.9*-0.1 - 0.1 = -0.19 ; p = 0.71
    assert torch.allclose(p2, torch.tensor([0.71]))
    o.clear_gradients()
    assert float(g2) == 0.0


@pytest.mark.parametrize("decoupled", [False, True])
def test_adam_matches_reference_formula(decoupled):
    torch.manual_seed(0)
    p = torch.randn(5)
    ref = p.clone().double()
    g = torch.randn(5)
    o = dnn.Adam(1e-2, 0.9, 0.999, 1e-8, 5e-3, decoupled)
    o.attach([p], [g])
    m = v = torch.zeros(5, dtype=torch.float64)
    for t in range(1, 4):
        o.update()
        gd = g.double()
        m = 0.9 * m + 0.1 * gd
        v = 0.999 * v + 0.001 * gd * gd
        upd = 1e-2 * (m / (1 - 0.9 ** t)) / (torch.sqrt(v / (1 - 0.999 ** t)) + 1e-8)
        if decoupled:
            ref = ref - 5e-3 * 1e-2 * ref
        else:
            upd = upd + 5e-3 * 1e-2 * ref
        ref = ref - upd
    assert torch.allclose(p.double(), ref, atol=1e-6)
    assert o.name() == ("AdamW" if decoupled else "Adam")
    clone = dnn.OptimizerFactory.create_from_config(o.get_config())
    assert clone.get_config() == o.get_config()


def test_schedulers_follow_reference_formulas():
    o = dnn.SGD(1.0)
    s = dnn.StepLR(o, 2, 0.5)
    lrs = []
    for _ in range(4):
        s.step()
        lrs.append(o.get_learning_rate())
    assert lrs == [1.0, 0.5, 0.5, 0.25]
    o = dnn.SGD(1.0)
    s = dnn.MultiStepLR(o, [3, 1], 0.1)
    got = []
    for _ in range(3):
        s.step()
        got.append(round(o.get_learning_rate(), 6))
    assert got == [0.1, 0.1, 0.01]
    o = dnn.SGD(1.0)
    s = dnn.CosineAnnealingLR(o, 4, 0.0)
    s.step()
    assert abs(o.get_learning_rate() - (1 + math.cos(math.pi / 4)) / 2) < 1e-6
    o = dnn.SGD(2.0)
    s = dnn.LinearWarmup(o, 4, 0.0)
    assert o.get_learning_rate() == 0.0
    s.step()
    assert abs(o.get_learning_rate() - 0.5) < 1e-9 and not s.is_warmup_complete()
    o = dnn.SGD(1.0)
    s = dnn.WarmupCosineAnnealing(o, 2, 6, 0.0, 0.0)
    vals = []
    for _ in range(6):
        s.step()
        vals.append(o.get_learning_rate())
    assert abs(vals[1] - 1.0) < 1e-9 and abs(vals[-1]) < 1e-9 and vals[2] > vals[3] > vals[4]
    o = dnn.SGD(1.0)
    s = dnn.ReduceLROnPlateau(o, "min", 0.5, 2)
    for metric in (1.0, 1.0, 1.0):
        s.step(metric)
    assert o.get_learning_rate() == 0.5
    o = dnn.SGD(1.0)
    s = dnn.PolynomialLR(o, 4, 2.0, 0.0)
    s.step()
    assert abs(o.get_learning_rate() - 0.75 ** 2) < 1e-9
    o = dnn.SGD(0.0)
    s = dnn.OneCycleLR(o, 1.0, 10, 0.3, 25.0, 1e4)
    assert abs(o.get_learning_rate() - 0.04) < 1e-9
    peak = 0
    for _ in range(10):
        s.step()
        peak = max(peak, o.get_learning_rate())
    assert abs(peak - 1.0) < 1e-9 and o.get_learning_rate() < 0.04
    o = dnn.SGD(1.0)
    s = dnn.CosineAnnealingWarmRestarts(o, 2, 2, 0.0)
    s.step()
    s.step()
    assert abs(o.get_learning_rate() - 1.0) < 1e-9      # restart
    for t in ("step_lr", "multi_step_lr", "exponential_lr", "cosine_annealing_lr", "cosine_annealing_warm_restarts",
              "linear_warmup", "warmup_cosine_annealing", "reduce_lr_on_plateau", "polynomial_lr", "one_cycle_lr"):
        sch = dnn.SchedulerFactory.create(t, dnn.SGD(0.1))
        assert dnn.SchedulerFactory.create_from_config(sch.get_config(), dnn.SGD(0.1)).get_config() == sch.get_config()


@pytest.mark.parametrize("kind,ctor", [("softmax_ce", dnn.SoftmaxCrossEntropyLoss), ("logsoftmax_ce", dnn.LogSoftmaxCrossEntropyLoss),
                                       ("mse", dnn.MSELoss), ("mae", dnn.MAELoss), ("huber", lambda: dnn.HuberLoss(0.7))])
def test_loss_gradients_match_autograd(kind, ctor):
    torch.manual_seed(1)
    B, C = 6, 9
    p = torch.randn(B, C, 1, 1)
    t = torch.nn.functional.one_hot(torch.randint(0, C, (B,)), C).float().reshape(B, C, 1, 1)
    loss_fn = ctor()
    loss, grad, _ = loss_fn.forward_backward(p, t)
    pr = p.clone().requires_grad_(True)
    if "ce" in kind:
        ref = torch.nn.functional.cross_entropy(pr.reshape(B, C), t.reshape(B, C).argmax(1))
    elif kind == "mse":
        ref = torch.nn.functional.mse_loss(pr, t)
    elif kind == "mae":
        ref = torch.nn.functional.l1_loss(pr, t)
    else:
        ref = torch.nn.functional.huber_loss(pr, t, delta=0.7)
    ref.backward()
    assert abs(float(loss) - float(ref.detach())) < 1e-5
    assert torch.allclose(grad, pr.grad, atol=1e-6)
    assert dnn.LossFactory.create(loss_fn.get_config()["type"]).name() == loss_fn.name()


def test_crossentropy_on_probabilities():
    p = torch.tensor([[0.7, 0.2, 0.1], [0.0, 1.0, 0.0]]).reshape(2, 3, 1, 1)
    t = torch.tensor([[1.0, 0, 0], [1.0, 0, 0]]).reshape(2, 3, 1, 1)
    l = dnn.CrossEntropyLoss(1e-7)
    loss, grad, correct = l.forward_backward(p, t)
    assert abs(float(loss) - (-(math.log(0.7) + math.log(1e-7)) / 2)) < 1e-4
    assert torch.allclose(grad, (p - t) / 2) and int(correct) == 1


def test_env_file_and_training_config(tmp_path, monkeypatch):
    f = tmp_path / ".env"
    f.write_text("# comment\nEPOCHS=3\nBATCH_SIZE = 48\nLR_INITIAL=\"0.005\"\nDEVICE_TYPE='GPU'\nPROFILER_TYPE=NORMAL\nbroken line\n")
    env.clear_loaded()
    for k in ("EPOCHS", "BATCH_SIZE", "LR_INITIAL", "DEVICE_TYPE", "PROFILER_TYPE"):
        monkeypatch.delenv(k, raising=False)
    assert env.load_env_file(str(f))
    assert env.get_env("EPOCHS", 10) == 3 and env.get_env("LR_INITIAL", 0.01) == 0.005
    assert os.environ["DEVICE_TYPE"] == "GPU" and env.get_env("MISSING", "dflt") == "dflt"
    from dcnn_b200.train import TrainingConfig
    cfg = TrainingConfig().load_from_env()
    assert cfg.epochs == 3 and cfg.device_type == "GPU" and cfg.profiler_type == "NORMAL" and cfg.num_microbatches == 2
    assert env.get_env("BATCH_SIZE", 32) == 32 or env.get_env("BATCH_SIZE", 32) == 48   # key with spaces is trimmed
    env.clear_loaded()


def test_profiling_summary_and_cache_accounting(capsys):
    from dcnn_b200.models import create_mnist_trainer
    m = create_mnist_trainer()
    m.initialize()
    m.enable_profiling(True)
    y = m.forward(torch.randn(4, 1, 28, 28))
    m.backward(torch.randn_like(y))
    m.print_profiling_summary()
    out = capsys.readouterr().out
    assert "Performance Profile: mnist_cnn_model" in out and "TOTAL" in out
    assert set(m.get_forward_times()) == set(m.get_backward_times())
    assert sum(l.cached_memory_bytes() for l in m.layers) > 0
    m.clear_cache()
    assert sum(l.cached_memory_bytes() for l in m.layers) == 0
    m.print_summary([4, 1, 28, 28])
    assert "Model Summary" in capsys.readouterr().out


def test_measured_layer_costs_drive_the_balanced_partitioner():
    import torch
    from dcnn_b200.models import create_resnet9_cifar10
    from dcnn_b200.partitioner import BalancedPartitioner, measure_layer_costs
    m = create_resnet9_cifar10()
    m.set_seed(0)
    m.initialize()
    before = {k: v.clone() for k, v in m.buffers().items()}
    costs = measure_layer_costs(m, [4, 3, 32, 32], warmup=0, iters=1)
    assert len(costs) == m.layer_size() and all(c > 0 for c in costs)
    for k, v in m.buffers().items():
        assert torch.equal(v, before[k])                      # running statistics untouched by the probe passes
    assert all(float(g.abs().sum()) == 0 for g in m.gradients()) and not m.is_profiling_enabled()
    parts = BalancedPartitioner(costs=costs).get_partitions(m.get_layers(), 3)
    assert parts[0].start_layer == 0 and parts[-1].end_layer == m.layer_size()
    loads = [sum(costs[p.start_layer:p.end_layer]) for p in parts]
    # min-max optimality against brute force over all contiguous 3-way splits
    n = len(costs)
    best = min(max(sum(costs[:i]), sum(costs[i:j]), sum(costs[j:])) for i in range(1, n - 1) for j in range(i + 1, n))
    assert abs(max(loads) - best) < 1e-9


def test_reference_named_helpers_exist_and_work():
    """Small API-parity items: data_loader.hpp factory / split / normalise, AugmentationStrategy.set_augmentations,
    HuberLoss delta accessors, Communicator recipient registry spellings, PipelineStage identity."""
    import numpy as np
    import torch
    from dcnn_b200 import nn
    from dcnn_b200.data import (AugmentationBuilder, AugmentationStrategy, MNISTDataLoader, create_data_loader,
                                normalize_to_unit_range, train_val_split)
    from dcnn_b200.parallel.pipeline import CommandType, Endpoint, InProcessCommunicator, Message, PipelineStage
    assert isinstance(create_data_loader("MNIST"), MNISTDataLoader) and create_data_loader("tiny-imagenet").size() == 0
    try:
        create_data_loader("imagenet21k")
        raise AssertionError("unknown dataset accepted")
    except ValueError:
        pass
    tr, va = train_val_split(103, 0.2, seed=1)
    assert len(va) == 20 and len(tr) == 83 and sorted(np.concatenate([tr, va]).tolist()) == list(range(103))
    assert np.array_equal(train_val_split(103, 0.2, seed=1)[1], va)
    assert float(normalize_to_unit_range(torch.full((3,), 255.0)).max()) == 1.0
    src = AugmentationBuilder().horizontal_flip(1.0).cutout(1.0, 2).build()
    dst = AugmentationStrategy()
    dst.set_augmentations(src.get_augmentations())
    assert dst.size() == 2 and dst.get_augmentations()[0] is not src.get_augmentations()[0]
    h = nn.HuberLoss(1.0)
    h.set_delta(0.5)
    assert h.get_delta() == 0.5 and h.get_config()["parameters"]["delta"] == 0.5 and h.num_parameters() == 1
    assert nn.MSELoss().num_parameters() == 0
    a, b = InProcessCommunicator("a"), InProcessCommunicator("b")
    a.register_recipient("b", Endpoint.in_process(b))
    a.enqueue_output_message(Message("b", CommandType.HEALTH_CHECK, "x"))
    assert b.message_count(CommandType.HEALTH_CHECK) == 1 and not a.has_output_message()
    a.unregister_recipient("b")
    try:
        a.send_message(Message("b", CommandType.HEALTH_CHECK))
        raise AssertionError("send to an unregistered recipient succeeded")
    except KeyError:
        pass
    st = PipelineStage(b, "stage_x")
    assert st.get_stage_id() == "stage_x" and not st.is_configured()


def test_more_reference_named_helpers(tmp_path, capsys):
    import numpy as np
    import torch
    from dcnn_b200.data import CIFAR100DataLoader, NormalizationAugmentation, WiFiDataLoader
    from dcnn_b200.ops import tensor_ops as T
    from dcnn_b200.parallel.pipeline import CommandType, Endpoint, InProcessCoordinator, Job, Message
    a = torch.tensor([[1.0, -2.0], [3.0, 4.0]])
    assert torch.equal(T.mul_add_scalar(a, 2.0, 1.0), a * 2 + 1) and torch.equal(T.sub_mul_scalar(a, 1.0, 3.0), (a - 1) * 3)
    assert torch.equal(T.scalar_max(a, 0.0), a.clamp_min(0)) and torch.equal(T.clamp(a, -1.0, 3.0), a.clamp(-1, 3))
    assert T.equal(a, a.clone()).sum() == 4 and T.greater(a, torch.zeros(2, 2)).sum() == 3
    assert abs(T.norm_squared(a) - 30.0) < 1e-9 and abs(T.mean(a) - 1.5) < 1e-9 and abs(T.variance(a) - 5.25) < 1e-9
    assert T.same_shape(a, a) and T.shape_str(a) == "(2, 2)" and not T.check_nan_and_inf(a)
    b = a.clone()
    b[1, 0] = float("inf")
    assert T.check_nan_and_inf(b, "b") and "non-finite" in capsys.readouterr().out
    dst = torch.zeros(3, 2)
    T.copy_batch(a, 1, dst, 2)
    assert torch.equal(dst[2], a[1]) and float(T.set_scalar(dst, 7.0).sum()) == 42.0
    T.print_data(a)
    assert "Tensor(2, 2)" in capsys.readouterr().out
    with __import__("pytest").raises(ZeroDivisionError):
        T.div_scalar(a, 0)
    # CIFAR-100 coarse / fine targets from a synthetic binary file: <coarse><fine><3072 bytes>
    rec = np.zeros((6, 3074), np.uint8)
    rec[:, 0] = [0, 1, 2, 3, 19, 19]
    rec[:, 1] = [5, 15, 25, 35, 99, 98]
    f = tmp_path / "train.bin"
    rec.tofile(f)
    ld = CIFAR100DataLoader()
    assert ld.load_data(str(f)) and not ld.is_using_coarse_labels() and ld.get_num_classes() == 100
    assert ld.labels.tolist() == [5, 15, 25, 35, 99, 98] and ld.get_label_counts()[99] == 1
    ld.set_use_coarse_labels(True)
    assert ld.is_using_coarse_labels() and ld.get_num_classes() == 20 and ld.labels.tolist() == [0, 1, 2, 3, 19, 19]
    ld.prepare_batches(6)
    _, t = ld.get_next_batch()
    assert t.shape == (6, 20, 1, 1) and t[4, 19, 0, 0] == 1
    ld.set_use_coarse_labels(False)
    assert ld.labels.tolist() == [5, 15, 25, 35, 99, 98]
    n = NormalizationAugmentation((0.1, 0.2, 0.3), (1, 2, 3))
    assert n.get_mean() == [0.1, 0.2, 0.3] and n.get_std() == [1.0, 2.0, 3.0]
    w = WiFiDataLoader()
    assert w.get_feature_means() is None
    w.print_statistics()
    ep = Endpoint.tcp("h", 9)
    assert ep.communication_type() == "tcp" and ep.get_parameter("host") == "h" and ep.get_parameter("port") == "9"
    ep.set_parameter("port", "10")
    assert ep.port == 10
    m = Message("x", CommandType.FORWARD_JOB, Job(a, 1))
    assert m.has_type(Job) and not m.has_type(str) and Message("x", CommandType.TRAIN_MODE, True).has_type(bool)
    from dcnn_b200 import nn
    from tests.test_pipeline_inprocess import _model
    c = InProcessCoordinator(_model(), nn.SGD(0.1), ["cpu", "cpu"])
    hits = []
    c.add_message_callback(lambda: hits.append(1))
    assert len(c.get_stages()) == 2 and c.get_stages()[0].get_hardware_info().get_logical_cores() >= 1
