"""Data augmentation (ref include/data_augmentation/augmentation.hpp:17-177 and the nine augmentation headers):
``Augmentation`` (apply in place on an NCHW batch in [0,1], own RNG), ``AugmentationStrategy``,
``AugmentationBuilder`` with horizontal_flip, vertical_flip, rotation (bilinear, zero fill), brightness (+delta,
clamp 0..1), contrast (*factor, clamp), gaussian_noise, random_crop (zero-pad then crop), cutout (zero square),
normalize (per-channel mean/std).

Each augmentation draws its per-sample decisions with a seeded NumPy generator, then applies them with the native
batched kernels of libdcnn_host.so (OpenMP over samples) on CPU batches, or with batched torch indexing when the
batch already lives on a GPU.  ``use_native=False`` forces the torch implementation (used to cross-check the two)."""
from __future__ import annotations

import ctypes as C
import math
from typing import List, Optional, Sequence

import numpy as np
import torch

from .. import _hostlib


def _native_ok(data: torch.Tensor, use_native: bool) -> bool:
    return use_native and data.device.type == "cpu" and data.dtype == torch.float32 and data.is_contiguous() \
        and _hostlib.available()


def _u8(mask: np.ndarray):
    m = np.ascontiguousarray(mask.astype(np.uint8))
    return m, m.ctypes.data_as(C.c_void_p)


def _f32(a: np.ndarray):
    m = np.ascontiguousarray(a.astype(np.float32))
    return m, m.ctypes.data_as(C.c_void_p)


def _i32(a: np.ndarray):
    m = np.ascontiguousarray(a.astype(np.int32))
    return m, m.ctypes.data_as(C.c_void_p)


class Augmentation:
    name = "Augmentation"

    def __init__(self):
        self.rng = np.random.default_rng()
        self.use_native = True

    def set_seed(self, seed: int):
        self.rng = np.random.default_rng(int(seed))

    def apply(self, data: torch.Tensor, labels: Optional[torch.Tensor] = None):
        raise NotImplementedError

    def get_name(self) -> str:
        return self.name

    def clone(self) -> "Augmentation":
        import copy
        c = copy.copy(self)
        c.rng = np.random.default_rng()
        return c

    def _mask(self, n: int, p: float) -> np.ndarray:
        return self.rng.random(n) < p


class _Flip(Augmentation):
    vertical = 0

    def __init__(self, probability: float = 0.5):
        super().__init__()
        self.probability = float(probability)

    def apply(self, data, labels=None):
        if data.dim() != 4:
            return
        n, c, h, w = data.shape
        mask = self._mask(n, self.probability)
        if _native_ok(data, self.use_native):
            m, mp = _u8(mask)
            _hostlib.lib().dcnn_aug_flip(C.c_void_p(data.data_ptr()), n, c, h, w, mp, self.vertical)
        else:
            idx = torch.from_numpy(np.nonzero(mask)[0]).to(data.device)
            if idx.numel():
                data[idx] = data[idx].flip(2 if self.vertical else 3)


class HorizontalFlipAugmentation(_Flip):
    name, vertical = "HorizontalFlip", 0


class VerticalFlipAugmentation(_Flip):
    name, vertical = "VerticalFlip", 1


class RotationAugmentation(Augmentation):
    name = "Rotation"

    def __init__(self, probability: float = 0.5, max_angle_degrees: float = 15.0):
        super().__init__()
        self.probability, self.max_angle_degrees = float(probability), float(max_angle_degrees)

    def apply(self, data, labels=None):
        if data.dim() != 4:
            return
        n, c, h, w = data.shape
        mask = self._mask(n, self.probability)
        angles = self.rng.uniform(-self.max_angle_degrees, self.max_angle_degrees, n).astype(np.float32)
        if _native_ok(data, self.use_native):
            m, mp = _u8(mask)
            a, ap = _f32(angles)
            _hostlib.lib().dcnn_aug_rotate(C.c_void_p(data.data_ptr()), n, c, h, w, mp, ap)
            return
        idx = np.nonzero(mask)[0]
        if idx.size == 0:
            return
        dev = data.device
        ang = torch.from_numpy(angles[idx]).to(dev) * (math.pi / 180.0)
        ca, sa = torch.cos(ang).view(-1, 1, 1), torch.sin(ang).view(-1, 1, 1)
        ys, xs = torch.meshgrid(torch.arange(h, device=dev, dtype=torch.float32),
                                torch.arange(w, device=dev, dtype=torch.float32), indexing="ij")
        cx, cy = w / 2.0, h / 2.0
        sx = (xs - cx) * ca - (ys - cy) * sa + cx
        sy = (xs - cx) * sa + (ys - cy) * ca + cy
        x1, y1 = torch.floor(sx).long(), torch.floor(sy).long()
        x2, y2 = x1 + 1, y1 + 1
        ok = (x1 >= 0) & (x2 < w) & (y1 >= 0) & (y2 < h)
        wx, wy = (sx - x1).unsqueeze(1), (sy - y1).unsqueeze(1)
        x1c, x2c, y1c, y2c = x1.clamp(0, w - 1), x2.clamp(0, w - 1), y1.clamp(0, h - 1), y2.clamp(0, h - 1)
        src = data[torch.from_numpy(idx).to(dev)]
        flat = src.reshape(len(idx), c, h * w)

        def g(yy, xx):
            lin = (yy * w + xx).view(len(idx), 1, h * w).expand(-1, c, -1)
            return flat.gather(2, lin).view(len(idx), c, h, w)
        out = g(y1c, x1c) * (1 - wx) * (1 - wy) + g(y1c, x2c) * wx * (1 - wy) + g(y2c, x1c) * (1 - wx) * wy + \
            g(y2c, x2c) * wx * wy
        data[torch.from_numpy(idx).to(dev)] = out * ok.unsqueeze(1).to(out.dtype)


class _AffineClamp(Augmentation):
    def _draw(self, n):
        raise NotImplementedError

    def apply(self, data, labels=None):
        if data.dim() != 4:
            return
        n = data.shape[0]
        mask = self._mask(n, self.probability)
        mul, add = self._draw(n)
        if _native_ok(data, self.use_native):
            m, mp = _u8(mask)
            mu, mup = _f32(mul)
            ad, adp = _f32(add)
            _hostlib.lib().dcnn_aug_affine_clamp(C.c_void_p(data.data_ptr()), n, C.c_size_t(data[0].numel()), mp, mup, adp)
        else:
            idx = torch.from_numpy(np.nonzero(mask)[0]).to(data.device)
            if idx.numel():
                mu = torch.from_numpy(mul[mask].astype(np.float32)).to(data.device).view(-1, 1, 1, 1)
                ad = torch.from_numpy(add[mask].astype(np.float32)).to(data.device).view(-1, 1, 1, 1)
                data[idx] = (data[idx] * mu + ad).clamp_(0, 1)


class BrightnessAugmentation(_AffineClamp):
    name = "Brightness"

    def __init__(self, probability: float = 0.5, brightness_range: float = 0.2):
        super().__init__()
        self.probability, self.brightness_range = float(probability), float(brightness_range)

    def _draw(self, n):
        return np.ones(n, np.float32), self.rng.uniform(-self.brightness_range, self.brightness_range, n)


class ContrastAugmentation(_AffineClamp):
    name = "Contrast"

    def __init__(self, probability: float = 0.5, contrast_range: float = 0.2):
        super().__init__()
        self.probability, self.contrast_range = float(probability), float(contrast_range)

    def _draw(self, n):
        return self.rng.uniform(1 - self.contrast_range, 1 + self.contrast_range, n), np.zeros(n, np.float32)


class GaussianNoiseAugmentation(Augmentation):
    name = "GaussianNoise"

    def __init__(self, probability: float = 0.3, noise_std: float = 0.05):
        super().__init__()
        self.probability, self.noise_std = float(probability), float(noise_std)

    def apply(self, data, labels=None):
        if data.dim() != 4:
            return
        n = data.shape[0]
        mask = self._mask(n, self.probability)
        gen = torch.Generator().manual_seed(int(self.rng.integers(0, 2 ** 31 - 1)))
        noise = torch.randn(data.shape, generator=gen) * self.noise_std
        if _native_ok(data, self.use_native):
            m, mp = _u8(mask)
            _hostlib.lib().dcnn_aug_add_noise_clamp(C.c_void_p(data.data_ptr()), C.c_void_p(noise.data_ptr()), n,
                                                    C.c_size_t(data[0].numel()), mp)
        else:
            idx = torch.from_numpy(np.nonzero(mask)[0]).to(data.device)
            if idx.numel():
                data[idx] = (data[idx] + noise.to(data.device)[idx]).clamp_(0, 1)


class RandomCropAugmentation(Augmentation):
    name = "RandomCrop"

    def __init__(self, probability: float = 0.5, padding: int = 4):
        super().__init__()
        self.probability, self.padding = float(probability), int(padding)

    def apply(self, data, labels=None):
        if data.dim() != 4:
            return
        n, c, h, w = data.shape
        mask = self._mask(n, self.probability)
        sx = self.rng.integers(0, 2 * self.padding + 1, n)
        sy = self.rng.integers(0, 2 * self.padding + 1, n)
        if _native_ok(data, self.use_native):
            m, mp = _u8(mask)
            xs, xp = _i32(sx)
            ys, yp = _i32(sy)
            _hostlib.lib().dcnn_aug_random_crop(C.c_void_p(data.data_ptr()), n, c, h, w, self.padding, mp, xp, yp)
        else:
            p = self.padding
            padded = torch.nn.functional.pad(data, (p, p, p, p))
            for b in np.nonzero(mask)[0]:
                data[b] = padded[b, :, sy[b]:sy[b] + h, sx[b]:sx[b] + w]


class CutoutAugmentation(Augmentation):
    name = "Cutout"

    def __init__(self, probability: float = 0.5, cutout_size: int = 8):
        super().__init__()
        self.probability, self.cutout_size = float(probability), int(cutout_size)

    def apply(self, data, labels=None):
        if data.dim() != 4:
            return
        n, c, h, w = data.shape
        mask = self._mask(n, self.probability)
        x0 = self.rng.integers(0, max(1, w - self.cutout_size + 1), n)
        y0 = self.rng.integers(0, max(1, h - self.cutout_size + 1), n)
        if _native_ok(data, self.use_native):
            m, mp = _u8(mask)
            xs, xp = _i32(x0)
            ys, yp = _i32(y0)
            _hostlib.lib().dcnn_aug_cutout(C.c_void_p(data.data_ptr()), n, c, h, w, self.cutout_size, mp, xp, yp)
        else:
            for b in np.nonzero(mask)[0]:
                data[b, :, y0[b]:y0[b] + self.cutout_size, x0[b]:x0[b] + self.cutout_size] = 0


class NormalizationAugmentation(Augmentation):
    name = "Normalization"

    def __init__(self, mean: Sequence[float] = (0.485, 0.456, 0.406), std: Sequence[float] = (0.229, 0.224, 0.225)):
        super().__init__()
        self.mean, self.std = [float(v) for v in mean], [float(v) for v in std]

    def get_mean(self):
        return list(self.mean)

    def get_std(self):
        return list(self.std)

    def apply(self, data, labels=None):
        if data.dim() != 4:
            return
        n, c, h, w = data.shape
        if c not in (1, 3):
            raise ValueError("NormalizationAugmentation: unsupported number of channels")
        if _native_ok(data, self.use_native):
            me, mep = _f32(np.array(self.mean))
            sd, sdp = _f32(np.array(self.std))
            _hostlib.lib().dcnn_aug_normalize(C.c_void_p(data.data_ptr()), n, c, h, w, mep, sdp)
        else:
            mean = torch.tensor(self.mean[:c] if c == 3 else self.mean[:1], device=data.device).view(1, c, 1, 1)
            std = torch.tensor(self.std[:c] if c == 3 else self.std[:1], device=data.device).view(1, c, 1, 1)
            data.sub_(mean).div_(std)


class AugmentationStrategy:
    def __init__(self, other: Optional["AugmentationStrategy"] = None):
        self.augmentations: List[Augmentation] = [a.clone() for a in other.augmentations] if other else []

    def apply(self, data, labels=None):
        for a in self.augmentations:
            a.apply(data, labels)

    def add_augmentation(self, aug: Augmentation):
        self.augmentations.append(aug)

    def remove_augmentation(self, index_or_name):
        if isinstance(index_or_name, int):
            if 0 <= index_or_name < len(self.augmentations):
                del self.augmentations[index_or_name]
        else:
            self.augmentations = [a for a in self.augmentations if a.get_name() != index_or_name]

    def clear_augmentations(self):
        self.augmentations.clear()

    def set_augmentations(self, augmentations):
        """Replace the whole chain (clones are stored, like the reference's copy semantics)."""
        self.augmentations = [a.clone() for a in augmentations]

    def set_seed(self, seed: int):
        for i, a in enumerate(self.augmentations):
            a.set_seed(seed + i)

    def set_use_native(self, flag: bool):
        for a in self.augmentations:
            a.use_native = flag

    def size(self) -> int:
        return len(self.augmentations)

    def get_augmentations(self):
        return self.augmentations


class AugmentationBuilder:
    def __init__(self):
        self.strategy = AugmentationStrategy()

    def horizontal_flip(self, probability=0.5):
        self.strategy.add_augmentation(HorizontalFlipAugmentation(probability))
        return self

    def vertical_flip(self, probability=0.5):
        self.strategy.add_augmentation(VerticalFlipAugmentation(probability))
        return self

    def rotation(self, probability=0.5, max_angle_degrees=15.0):
        self.strategy.add_augmentation(RotationAugmentation(probability, max_angle_degrees))
        return self

    def brightness(self, probability=0.5, range=0.2):
        self.strategy.add_augmentation(BrightnessAugmentation(probability, range))
        return self

    def contrast(self, probability=0.5, range=0.2):
        self.strategy.add_augmentation(ContrastAugmentation(probability, range))
        return self

    def gaussian_noise(self, probability=0.3, std_dev=0.05):
        self.strategy.add_augmentation(GaussianNoiseAugmentation(probability, std_dev))
        return self

    def random_crop(self, probability=0.5, padding=4):
        self.strategy.add_augmentation(RandomCropAugmentation(probability, padding))
        return self

    def cutout(self, probability=0.5, cutout_size=8):
        self.strategy.add_augmentation(CutoutAugmentation(probability, cutout_size))
        return self

    def normalize(self, mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225)):
        self.strategy.add_augmentation(NormalizationAugmentation(mean, std))
        return self

    def custom_augmentation(self, augmentation: Augmentation):
        self.strategy.add_augmentation(augmentation)
        return self

    def build(self) -> AugmentationStrategy:
        return AugmentationStrategy(self.strategy)
