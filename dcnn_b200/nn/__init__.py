from .activations import (ActivationFactory, ActivationFunction, ELU, LeakyReLU, Linear, ReLU,  # noqa: F401
                          Sigmoid, Softmax, Tanh, create_activation)
from .blocks import ResidualBlock, residual_block  # noqa: F401
from .factory import (LayerBuilder, LayerFactory, activation_layer, avgpool2d_layer,  # noqa: F401
                      batchnorm_layer, conv2d_layer, dense_layer, dropout_layer, flatten_layer,
                      groupnorm_layer, maxpool2d_layer)
from .layers import (ActivationLayer, AvgPool2DLayer, BatchNormLayer, Conv2DLayer, DenseLayer,  # noqa: F401
                     DropoutLayer, FlattenLayer, GroupNormLayer, Layer, MaxPool2DLayer)
from .loss import (CrossEntropyLoss, HuberLoss, LogSoftmaxCrossEntropyLoss, Loss, LossFactory,  # noqa: F401
                   MAELoss, MSELoss, SoftmaxCrossEntropyLoss)
from .optimizers import SGD, Adam, Optimizer, OptimizerFactory  # noqa: F401
from .schedulers import (CosineAnnealingLR, CosineAnnealingWarmRestarts, ExponentialLR,  # noqa: F401
                         LinearWarmup, MultiStepLR, OneCycleLR, PolynomialLR, ReduceLROnPlateau,
                         Scheduler, SchedulerFactory, StepLR, WarmupCosineAnnealing)
from .sequential import Partition, Sequential, SequentialBuilder  # noqa: F401
from .mem_pool import MemPool, TempBuffer  # noqa: F401
