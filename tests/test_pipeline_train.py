import torch

from dcnn_b200 import nn as dnn
from dcnn_b200.data import SyntheticDataLoader
from dcnn_b200.parallel.pipeline import InProcessCoordinator
from dcnn_b200.parallel.pipeline.train import train_model, validate_semi_async_epoch
from dcnn_b200.train import TrainingConfig
from tests.test_pipeline_inprocess import _model


def test_train_model_semi_async_and_validation():
    coord = InProcessCoordinator(_model(), dnn.Adam(5e-3), ["cpu", "cpu"])
    coord.seed = 1
    coord.set_loss_function(dnn.SoftmaxCrossEntropyLoss())
    coord.initialize()
    assert coord.deploy_stages()
    coord.start()
    train = SyntheticDataLoader(64, (3, 8, 8), 5, seed=0)
    test = SyntheticDataLoader(16, (3, 8, 8), 5, seed=1)
    cfg = TrainingConfig(epochs=3, batch_size=16, num_microbatches=2, progress_print_interval=1000)
    hist = train_model(coord, train, test, cfg)
    assert len(hist) == 3 and hist[-1]["loss"] < hist[0]["loss"]
    before = coord.collect_current_parameters() and [p.clone() for p in coord.model.parameters()]
    loss, acc = validate_semi_async_epoch(coord, test)
    assert loss == loss and 0.0 <= acc <= 1.0
    coord.collect_current_parameters()
    for a, b in zip(before, coord.model.parameters()):      # validation must not move the weights
        assert torch.equal(a, b)
    coord.stop()
