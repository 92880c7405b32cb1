"""Model zoo with the same function names, layer names and hyper-parameters as
/root/reference/include/nn/example_models.hpp:13-436, so the emitted JSON configs and the `.bin`
parameter order match the reference's."""
from __future__ import annotations

from ..nn.sequential import Sequential, SequentialBuilder


def _cbr(b, ch, bias, idx_conv, idx_bn, idx_relu, bn=True, eps=1e-5):
    b.conv2d(ch, 3, 3, 1, 1, 1, 1, bias, idx_conv)
    if bn:
        b.batchnorm(eps, 0.1, True, idx_bn)
    b.activation("relu", idx_relu)
    return b


def create_mnist_trainer() -> Sequential:
    return (SequentialBuilder("mnist_cnn_model").input([1, 28, 28])
            .conv2d(8, 5, 5, 1, 1, 0, 0, True, "conv1").batchnorm(1e-5, 0.1, True, "bn1")
            .activation("relu", "relu1").maxpool2d(3, 3, 3, 3, 0, 0, "pool1")
            .conv2d(16, 1, 1, 1, 1, 0, 0, True, "conv2_1x1").batchnorm(1e-5, 0.1, True, "bn2")
            .activation("relu", "relu2")
            .conv2d(48, 5, 5, 1, 1, 0, 0, True, "conv3").batchnorm(1e-5, 0.1, True, "bn3")
            .activation("relu", "relu3").maxpool2d(2, 2, 2, 2, 0, 0, "pool2")
            .flatten("flatten").dense(10, True, "output").build())


def create_cifar10_trainer_v1() -> Sequential:
    return (SequentialBuilder("cifar10_cnn_classifier_v1").input([3, 32, 32])
            .conv2d(16, 3, 3, 1, 1, 0, 0, True, "conv1").batchnorm(1e-5, 0.1, True, "bn1")
            .activation("relu", "relu1").maxpool2d(3, 3, 3, 3, 0, 0, "maxpool1")
            .conv2d(64, 3, 3, 1, 1, 0, 0, True, "conv2").activation("relu", "relu2")
            .maxpool2d(4, 4, 4, 4, 0, 0, "maxpool2").flatten("flatten").dense(10, True, "fc1").build())


def _vgg_like(name, in_hw, fc0, classes):
    b = SequentialBuilder(name) if name else SequentialBuilder()
    b.input([3, in_hw, in_hw])
    _cbr(b, 64, False, "conv0", "bn0", "relu0")
    _cbr(b, 64, False, "conv1", "bn1", "relu1")
    b.maxpool2d(2, 2, 2, 2, 0, 0, "pool0")
    _cbr(b, 128, False, "conv2", "bn2", "relu2")
    _cbr(b, 128, False, "conv3", "bn3", "relu3")
    b.maxpool2d(2, 2, 2, 2, 0, 0, "pool1")
    _cbr(b, 256, False, "conv4", "bn5", "relu5")
    _cbr(b, 256, False, "conv5", None, "relu6", bn=False)
    _cbr(b, 256, False, "conv6", "bn6", "relu6")
    b.maxpool2d(2, 2, 2, 2, 0, 0, "pool2")
    _cbr(b, 512, False, "conv7", "bn8", "relu7")
    _cbr(b, 512, False, "conv8", "bn9", "relu8")
    _cbr(b, 512, False, "conv9", "bn10", "relu9")
    b.maxpool2d(2, 2, 2, 2, 0, 0, "pool3")
    return (b.flatten("flatten").dense(fc0, True, "fc0").activation("relu", "relu10")
            .dense(classes, True, "fc1").build())


def create_cifar10_trainer_v2() -> Sequential:
    return _vgg_like("cifar10_cnn_classifier", 32, 512, 10)


def create_cifar100_trainer() -> Sequential:
    """CIFAR-100 CNN of examples/cifar100_cnn_trainer.cpp: the v2 VGG-style stack with a 100-way head."""
    return _vgg_like("cifar100_cnn_classifier", 32, 512, 100)


def create_cnn_tiny_imagenet() -> Sequential:
    return _vgg_like("", 64, 1024, 200)


def create_resnet9_cifar10() -> Sequential:
    b = SequentialBuilder("ResNet-9-CIFAR10").input([3, 32, 32])
    _cbr(b, 64, True, "conv1", "bn1", "relu1")
    _cbr(b, 128, True, "conv2", "bn2", "relu2")
    b.maxpool2d(2, 2, 2, 2, 0, 0, "pool1")
    b.basic_residual_block(128, 128, 1, "res_block1").basic_residual_block(128, 128, 1, "res_block2")
    _cbr(b, 256, True, "conv3", "bn3", "relu3")
    b.maxpool2d(2, 2, 2, 2, 0, 0, "pool2")
    b.basic_residual_block(256, 256, 1, "res_block3").basic_residual_block(256, 256, 1, "res_block4")
    _cbr(b, 512, True, "conv4", "bn4", "relu4")
    b.maxpool2d(2, 2, 2, 2, 0, 0, "pool3")
    b.basic_residual_block(512, 512, 1, "res_block5")
    return b.avgpool2d(4, 4, 1, 1, 0, 0, "avgpool").flatten("flatten").dense(10, True, "output").build()


def _basic_stack(b, spec):
    for cin, cout, stride, name in spec:
        b.basic_residual_block(cin, cout, stride, name)
    return b


def create_resnet18_cifar10() -> Sequential:
    b = SequentialBuilder("ResNet-18-CIFAR10").input([3, 32, 32])
    _cbr(b, 64, True, "conv1", "bn1", "relu1")
    _basic_stack(b, [(64, 64, 1, "layer1_block1"), (64, 64, 1, "layer1_block2"),
                     (64, 128, 2, "layer2_block1"), (128, 128, 1, "layer2_block2"), (128, 128, 1, "layer2_block3"),
                     (128, 256, 2, "layer3_block1"), (256, 256, 1, "layer3_block2"), (256, 256, 1, "layer3_block3"),
                     (256, 512, 2, "layer4_block1"), (512, 512, 1, "layer4_block2")])
    return b.avgpool2d(4, 4, 4, 4, 0, 0, "avgpool").flatten("flatten").dense(10, True, "output").build()


def create_resnet20_cifar10() -> Sequential:
    b = SequentialBuilder("ResNet-20-CIFAR10").input([3, 32, 32])
    _cbr(b, 64, True, "conv1", "bn1", "relu1")
    _basic_stack(b, [(64, 64, 1, "layer1_block1"), (64, 64, 1, "layer1_block2"), (64, 64, 1, "layer1_block3"),
                     (64, 128, 2, "layer2_block1"), (128, 128, 1, "layer2_block2"), (128, 128, 1, "layer2_block3"),
                     (128, 256, 2, "layer3_block1"), (256, 256, 1, "layer3_block2"), (256, 256, 1, "layer3_block3")])
    return b.avgpool2d(8, 8, 1, 1, 0, 0, "avgpool").flatten("flatten").dense(10, True, "output").build()


_RN50 = [(64, 64, 256, 1, "layer1_block1"), (256, 64, 256, 1, "layer1_block2"), (256, 64, 256, 1, "layer1_block3"),
         (256, 128, 512, 2, "layer2_block1"), (512, 128, 512, 1, "layer2_block2"),
         (512, 128, 512, 1, "layer2_block3"), (512, 128, 512, 1, "layer2_block4"),
         (512, 256, 1024, 2, "layer3_block1"), (1024, 256, 1024, 1, "layer3_block2"),
         (1024, 256, 1024, 1, "layer3_block3"), (1024, 256, 1024, 1, "layer3_block4"),
         (1024, 256, 1024, 1, "layer3_block5"), (1024, 256, 1024, 1, "layer3_block6"),
         (1024, 512, 2048, 2, "layer4_block1"), (2048, 512, 2048, 1, "layer4_block2"),
         (2048, 512, 2048, 1, "layer4_block3")]


def _bottleneck_stack(b):
    for cin, mid, cout, stride, name in _RN50:
        b.bottleneck_residual_block(cin, mid, cout, stride, name)
    return b


def create_resnet50_cifar10() -> Sequential:
    b = SequentialBuilder("ResNet-50-CIFAR10").input([3, 32, 32])
    _cbr(b, 64, True, "conv1", "bn1", "relu1")
    _bottleneck_stack(b)
    return b.flatten("flatten").dense(10, True, "fc").build()


def create_resnet9_tiny_imagenet() -> Sequential:
    b = SequentialBuilder("ResNet-9-Tiny-ImageNet").input([3, 64, 64])
    _cbr(b, 64, False, "conv1", "bn1", "relu1")
    _cbr(b, 128, False, "conv2", "bn2", "relu2")
    b.maxpool2d(2, 2, 2, 2, 0, 0, "pool1").basic_residual_block(128, 128, 1, "res1")
    _cbr(b, 256, False, "conv3", "bn3", "relu3")
    b.maxpool2d(2, 2, 2, 2, 0, 0, "pool2").basic_residual_block(256, 256, 1, "res2")
    _cbr(b, 512, False, "conv4", "bn4", "relu4")
    b.maxpool2d(2, 2, 2, 2, 0, 0, "pool3").basic_residual_block(512, 512, 1, "res3")
    return b.avgpool2d(4, 4, 1, 1, 0, 0, "avgpool").flatten("flatten").dense(200, True, "fc").build()


def create_resnet18_tiny_imagenet() -> Sequential:
    b = SequentialBuilder("ResNet-18-Tiny-ImageNet").input([3, 64, 64])
    _cbr(b, 32, False, "conv1", "bn1", "relu1", eps=1e-3)
    b.maxpool2d(2, 2, 2, 2, 0, 0, "maxpool")
    _basic_stack(b, [(32, 64, 1, "layer1_block1"), (64, 64, 1, "layer1_block2"),
                     (64, 128, 2, "layer2_block1"), (128, 128, 1, "layer2_block2"),
                     (128, 256, 2, "layer3_block1"), (256, 256, 1, "layer3_block2"),
                     (256, 512, 2, "layer4_block1"), (512, 512, 1, "layer4_block2")])
    return b.avgpool2d(4, 4, 1, 1, 0, 0, "avgpool").flatten("flatten").dense(200, True, "fc").build()


def create_resnet34_tiny_imagenet() -> Sequential:
    b = SequentialBuilder("ResNet-34-Tiny-ImageNet").input([3, 64, 64])
    _cbr(b, 32, False, "conv1", "bn1", "relu1", eps=1e-3)
    b.maxpool2d(2, 2, 2, 2, 0, 0, "maxpool")
    spec = [(32, 64, 1, "layer1_block1"), (64, 64, 1, "layer1_block2"), (64, 64, 1, "layer1_block3"),
            (64, 128, 2, "layer2_block1")] + [(128, 128, 1, f"layer2_block{i}") for i in (2, 3, 4)] + \
           [(128, 256, 2, "layer3_block1")] + [(256, 256, 1, f"layer3_block{i}") for i in (2, 3, 4, 5, 6)] + \
           [(256, 512, 2, "layer4_block1"), (512, 512, 1, "layer4_block2"), (512, 512, 1, "layer4_block3")]
    _basic_stack(b, spec)
    return b.avgpool2d(4, 4, 1, 1, 0, 0, "avgpool").flatten("flatten").dense(200, True, "fc").build()


def create_resnet50_tiny_imagenet() -> Sequential:
    b = SequentialBuilder("ResNet-50-Tiny-ImageNet").input([3, 64, 64])
    _cbr(b, 64, True, "conv1", "bn1", "relu1")
    b.maxpool2d(3, 3, 2, 2, 1, 1, "maxpool")
    _bottleneck_stack(b)
    return b.avgpool2d(4, 4, 1, 1, 0, 0, "avgpool").flatten("flatten").dense(200, True, "fc").build()


def create_resnet50_imagenet() -> Sequential:
    b = SequentialBuilder("ResNet-50-ImageNet").input([3, 224, 224])
    b.conv2d(64, 7, 7, 2, 2, 3, 3, True, "conv1").batchnorm(1e-5, 0.1, True, "bn1").activation("relu", "relu1")
    b.maxpool2d(3, 3, 2, 2, 1, 1, "maxpool")
    _bottleneck_stack(b)
    return b.avgpool2d(7, 7, 1, 1, 0, 0, "avgpool").flatten("flatten").dense(1000, True, "fc").build()


MODEL_ZOO = {
    "mnist_cnn": create_mnist_trainer,
    "cifar10_cnn_v1": create_cifar10_trainer_v1,
    "cifar10_cnn_v2": create_cifar10_trainer_v2,
    "cifar100_cnn": create_cifar100_trainer,
    "resnet9_cifar10": create_resnet9_cifar10,
    "resnet18_cifar10": create_resnet18_cifar10,
    "resnet20_cifar10": create_resnet20_cifar10,
    "resnet50_cifar10": create_resnet50_cifar10,
    "resnet9_tiny_imagenet": create_resnet9_tiny_imagenet,
    "cnn_tiny_imagenet": create_cnn_tiny_imagenet,
    "resnet18_tiny_imagenet": create_resnet18_tiny_imagenet,
    "resnet34_tiny_imagenet": create_resnet34_tiny_imagenet,
    "resnet50_tiny_imagenet": create_resnet50_tiny_imagenet,
    "resnet50_imagenet": create_resnet50_imagenet,
}
