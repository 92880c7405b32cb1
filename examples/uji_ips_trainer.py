#!/usr/bin/env python
"""WiFi indoor positioning regression (UJIIndoorLoc) -- counterpart of examples/uji_ips_trainer.cpp: Dense network on
520 RSSI features -> (longitude, latitude), Huber/MSE loss, z-scored features and targets."""
import os

import numpy as np

from _common import get_env, setup

from dcnn_b200 import nn
from dcnn_b200.data import WiFiDataLoader
from dcnn_b200.nn import SequentialBuilder
from dcnn_b200.train import train_regression_model


def _synthetic(n, seed):
    ld = WiFiDataLoader()
    g = np.random.default_rng(seed)
    w = g.normal(size=(520, 2)).astype(np.float32) / 20
    ld.features = g.normal(size=(n, 520)).astype(np.float32)
    ld.targets = (ld.features @ w + 0.05 * g.normal(size=(n, 2))).astype(np.float32)
    return ld


def main():
    cfg = setup()
    train_csv, val_csv = "./data/uji/trainingData.csv", "./data/uji/validationData.csv"
    if os.path.exists(train_csv) and not get_env("SYNTHETIC", False):
        train, val = WiFiDataLoader(), WiFiDataLoader()
        train.load_data(train_csv, 0, 520, 520, 522, True)      # examples/uji_ips_trainer.cpp:455
        val.load_data(val_csv, 0, 520, 520, 522, True)
    else:
        print("UJI dataset not found: using synthetic regression data of the same shape")
        train, val = _synthetic(cfg.batch_size * 16, 0), _synthetic(cfg.batch_size * 4, 0)
    train.normalize_data()
    val.normalize_with(train)
    model = (SequentialBuilder("uji_ips_regressor").input([train.get_num_features(), 1, 1])
             .dense(256, True, "fc1").activation("relu", "relu1").dropout(0.1, "drop1")
             .dense(128, True, "fc2").activation("relu", "relu2")
             .dense(train.get_num_outputs(), True, "out").build())
    model.set_device(cfg.torch_device())
    model.initialize()
    loss = nn.LossFactory.create_huber(1.0) if get_env("LOSS", "huber") == "huber" else nn.LossFactory.create_mse()
    train_regression_model(model, train, val, nn.Adam(get_env("LR_INITIAL", 0.001)), loss, cfg)


if __name__ == "__main__":
    main()
