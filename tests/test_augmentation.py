"""Augmentation tests: native (libdcnn_host.so) vs torch implementations agree given the same seed; semantics
match the reference formulas (rotation bilinear + zero fill, crop = zero-pad + window, clamp to [0,1])."""

import pytest
import torch

from dcnn_b200.data.augmentation import (AugmentationBuilder, BrightnessAugmentation, ContrastAugmentation,
                                         CutoutAugmentation, GaussianNoiseAugmentation, HorizontalFlipAugmentation,
                                         NormalizationAugmentation, RandomCropAugmentation, RotationAugmentation,
                                         VerticalFlipAugmentation)

AUGS = [lambda: HorizontalFlipAugmentation(0.5), lambda: VerticalFlipAugmentation(0.5),
        lambda: RotationAugmentation(0.7, 20.0), lambda: BrightnessAugmentation(0.6, 0.3),
        lambda: ContrastAugmentation(0.6, 0.4), lambda: GaussianNoiseAugmentation(0.5, 0.1),
        lambda: RandomCropAugmentation(0.7, 4), lambda: CutoutAugmentation(0.7, 5),
        lambda: NormalizationAugmentation()]


@pytest.mark.parametrize("make", AUGS)
def test_native_matches_torch(make):
    torch.manual_seed(0)
    x = torch.rand(16, 3, 12, 10)
    a, b = make(), make()
    a.set_seed(42)
    b.set_seed(42)
    b.use_native = False
    xa, xb = x.clone(), x.clone()
    a.apply(xa)
    b.apply(xb)
    assert torch.allclose(xa, xb, atol=1e-5), (xa - xb).abs().max()
    if not isinstance(a, NormalizationAugmentation):
        assert not torch.equal(xa, x)


def test_reference_semantics():
    x = torch.rand(2, 3, 8, 8)
    f = HorizontalFlipAugmentation(1.0)
    y = x.clone()
    f.apply(y)
    assert torch.equal(y, x.flip(3))
    c = RandomCropAugmentation(1.0, 2)
    c.set_seed(1)
    y = x.clone()
    c.apply(y)
    c2 = RandomCropAugmentation(1.0, 2)
    c2.set_seed(1)
    c2._mask(2, 1.0)
    sx, sy = c2.rng.integers(0, 5, 2), c2.rng.integers(0, 5, 2)
    padded = torch.nn.functional.pad(x, (2, 2, 2, 2))
    for b in range(2):
        assert torch.equal(y[b], padded[b, :, sy[b]:sy[b] + 8, sx[b]:sx[b] + 8])
    r = RotationAugmentation(1.0, 0.0)   # zero angle: interior is identity, last row/col zero (x2 < w rule)
    y = x.clone()
    r.apply(y)
    assert torch.allclose(y[:, :, :7, :7], x[:, :, :7, :7], atol=1e-6)
    assert float(y[:, :, 7, :].abs().max()) == 0.0
    n = NormalizationAugmentation((0.5, 0.4, 0.3), (0.2, 0.25, 0.5))
    y = x.clone()
    n.apply(y)
    assert torch.allclose(y[:, 1], (x[:, 1] - 0.4) / 0.25, atol=1e-6)


def test_builder_chain():
    s = (AugmentationBuilder().random_crop(1.0, 4).rotation(0.25, 5.0).horizontal_flip(0.5).brightness(0.2)
         .contrast(0.5, 0.2).gaussian_noise(0.3, 0.05).cutout(0.5, 8).vertical_flip(0.1)
         .normalize((0.485, 0.456, 0.406), (0.229, 0.224, 0.225)).build())
    assert s.size() == 9
    s.set_seed(3)
    x = torch.rand(8, 3, 32, 32)
    s.apply(x)
    assert torch.isfinite(x).all()
    s.remove_augmentation("Cutout")
    assert s.size() == 8
