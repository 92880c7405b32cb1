"""Data loaders (ref include/data_loading/*.hpp): whole dataset in RAM, ``prepare_batches(bs)`` builds every batch and
applies augmentation, labels are one-hot ``[N, classes, 1, 1]``, pixels are scaled by 1/255.

Loaders: MNIST (CSV), CIFAR-10 / CIFAR-100 (binary records), Tiny-ImageNet (JPEG tree with wnids.txt / words.txt /
val_annotations.txt), WiFi / UJI indoor positioning (CSV regression) and ``SyntheticDataLoader`` (named shape, fixed
seed; used by every benchmark since the GPU box has no datasets).  File parsing and batch gathering run in the native
host library when it is available (csrc/host/loaders.cpp); batches are handed out in pinned memory so the trainer's
host-to-device copies are asynchronous."""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from .. import _hostlib
from .augmentation import AugmentationStrategy


def _pin(t: torch.Tensor) -> torch.Tensor:
    if torch.cuda.is_available():
        try:
            return t.pin_memory()
        except RuntimeError:
            return t
    return t


class BaseDataLoader:
    def __init__(self):
        self.current_index = 0
        self.current_batch_index = 0
        self.batch_size = 32
        self.batches_prepared = False
        self.rng = np.random.default_rng()
        self.augmentation: Optional[AugmentationStrategy] = None
        self.batched_data: List[torch.Tensor] = []
        self.batched_labels: List[torch.Tensor] = []
        self._order: Optional[np.ndarray] = None
        self._shard: Optional[Tuple[int, int]] = None

    def set_shard(self, rank: int, world: int):
        """Data-parallel sharding: this loader serves samples ``rank, rank + world, ...`` of the (shuffled) order,
        truncated so that every replica sees the same number of samples (collectives need equal step counts).  All
        replicas must use the same seed (``set_seed``) so that they shuffle identically."""
        if not 0 <= int(rank) < int(world):
            raise ValueError(f"shard rank {rank} out of range for world size {world}")
        self._shard = (int(rank), int(world)) if int(world) > 1 else None
        self.batches_prepared = False
        self.current_index = self.current_batch_index = 0

    def get_rng_state(self) -> dict:
        """Everything that determines the next ``shuffle()`` / ``prepare_batches()``: the loader's generator, the sample
        order and the generators of the attached augmentations (for resumable training)."""
        aug = self.augmentation
        return {"rng": self.rng.bit_generator.state,
                "order": None if self._order is None else self._order.copy(),
                "aug": [a.rng.bit_generator.state for a in aug.augmentations] if aug is not None else []}

    def set_rng_state(self, state: dict):
        self.rng.bit_generator.state = state["rng"]
        self._order = None if state["order"] is None else np.asarray(state["order"]).copy()
        if self.augmentation is not None:
            for a, st in zip(self.augmentation.augmentations, state["aug"]):
                a.rng.bit_generator.state = st
        self.batches_prepared = False

    def shard_size(self) -> int:
        return self.size() // self._shard[1] if self._shard else self.size()

    # -- interface (ref data_loader.hpp:25-187) ----------------------------------------------------------
    def load_data(self, source: str) -> bool:
        raise NotImplementedError

    def size(self) -> int:
        raise NotImplementedError

    def _sample_arrays(self) -> Tuple[np.ndarray, np.ndarray]:
        """(data [N, ...] float32, labels) backing arrays."""
        raise NotImplementedError

    def _make_batch(self, idx: np.ndarray) -> Tuple[torch.Tensor, torch.Tensor]:
        raise NotImplementedError

    def set_seed(self, seed: int):
        self.rng = np.random.default_rng(int(seed))
        if self.augmentation is not None:
            self.augmentation.set_seed(int(seed) + 17)

    def shuffle(self):
        n = self.size()
        self._order = self.rng.permutation(n)
        self.current_index = 0
        if self.batches_prepared:  # the reference shuffles the prepared batches (sample order inside stays)
            perm = self.rng.permutation(len(self.batched_data))
            self.batched_data = [self.batched_data[i] for i in perm]
            self.batched_labels = [self.batched_labels[i] for i in perm]
            self.current_batch_index = 0

    def reset(self):
        self.current_index = 0
        self.current_batch_index = 0

    def _indices(self) -> np.ndarray:
        if self._order is None or len(self._order) != self.size():
            self._order = np.arange(self.size())
        if self._shard is not None:
            r, w = self._shard
            return self._order[r::w][:self.size() // w]
        return self._order

    def prepare_batches(self, batch_size: int):
        if self.size() == 0:
            print("Warning: No data loaded, cannot prepare batches")
            return
        self.batch_size = int(batch_size)
        order = self._indices()
        self.batched_data, self.batched_labels = [], []
        for s in range(0, len(order), self.batch_size):
            d, l = self._make_batch(order[s:s + self.batch_size])
            if self.augmentation is not None:
                self.augmentation.apply(d, l)
            self.batched_data.append(_pin(d))
            self.batched_labels.append(_pin(l))
        self.current_batch_index = 0
        self.batches_prepared = True

    def num_batches(self) -> int:
        return len(self.batched_data) if self.batches_prepared else (self.size() + self.batch_size - 1) // self.batch_size

    def are_batches_prepared(self) -> bool:
        return self.batches_prepared

    def get_batch_size(self) -> int:
        return self.batch_size

    def get_next_batch(self):
        """-> (data, labels) or None at the end of the epoch."""
        if not self.batches_prepared:
            raise RuntimeError("Batches not prepared! Call prepare_batches() first.")
        if self.current_batch_index >= len(self.batched_data):
            return None
        i = self.current_batch_index
        self.current_batch_index += 1
        return self.batched_data[i], self.batched_labels[i]

    def get_batch(self, batch_size: int):
        order = self._indices()
        if self.current_index >= len(order):
            return None
        idx = order[self.current_index:self.current_index + batch_size]
        self.current_index += len(idx)
        d, l = self._make_batch(idx)
        if self.augmentation is not None:
            self.augmentation.apply(d, l)
        return d, l

    def set_augmentation(self, aug: AugmentationStrategy):
        self.augmentation = aug

    def set_augmentation_strategy(self, aug: AugmentationStrategy):      # the reference's second spelling
        self.set_augmentation(aug)

    def generate_shuffled_indices(self, data_size: int) -> np.ndarray:
        return self.rng.permutation(int(data_size))

    def clear_augmentation(self):
        self.augmentation = None

    def has_augmentation(self) -> bool:
        return self.augmentation is not None

    def __iter__(self):
        self.reset()
        while True:
            b = self.get_next_batch()
            if b is None:
                return
            yield b


class ImageDataLoader(BaseDataLoader):
    """Classification images: ``images`` [N, C, H, W] float32 in [0,1], ``labels`` int32."""

    num_classes = 10
    image_shape = (1, 1, 1)
    class_names: Sequence[str] = ()

    def __init__(self):
        super().__init__()
        self.images = np.zeros((0,) + tuple(self.image_shape), np.float32)
        self.labels = np.zeros((0,), np.int32)

    def size(self):
        return int(self.images.shape[0])

    def get_image_shape(self):
        return list(self.image_shape)

    def get_num_classes(self):
        return self.num_classes

    def get_class_names(self):
        return list(self.class_names)

    def get_label_counts(self):
        """Samples per class (the reference prints these in print_data_stats)."""
        return np.bincount(self.labels, minlength=self.num_classes).tolist() if self.size() else [0] * self.num_classes

    def get_data_shape(self):
        return list(self.image_shape)

    def _make_batch(self, idx):
        n = len(idx)
        per = int(np.prod(self.image_shape))
        data = torch.empty((n,) + tuple(self.image_shape), dtype=torch.float32)
        onehot = torch.empty((n, self.num_classes, 1, 1), dtype=torch.float32)
        idx64 = np.ascontiguousarray(idx.astype(np.int64))
        if _hostlib.available() and self.images.flags["C_CONTIGUOUS"]:
            _hostlib.lib().dcnn_gather_batch(
                self.images.ctypes.data_as(C.c_void_p), self.labels.ctypes.data_as(C.c_void_p),
                idx64.ctypes.data_as(C.c_void_p), n, C.c_size_t(per), self.num_classes,
                C.c_void_p(data.data_ptr()), C.c_void_p(onehot.data_ptr()))
        else:
            data.copy_(torch.from_numpy(self.images[idx64]))
            onehot.zero_()
            lab = torch.from_numpy(self.labels[idx64].astype(np.int64))
            ok = (lab >= 0) & (lab < self.num_classes)
            onehot.view(n, -1)[torch.arange(n)[ok], lab[ok]] = 1.0
        return data, onehot

    def print_data_stats(self):
        if self.size() == 0:
            print("No data loaded")
            return
        counts = np.bincount(self.labels, minlength=self.num_classes)
        print(f"Dataset Statistics:\nTotal samples: {self.size()}\nImage shape: {tuple(self.image_shape)}")
        for i, c in enumerate(counts):
            nm = self.class_names[i] if i < len(self.class_names) else str(i)
            print(f"  {nm}: {c}")


class MNISTDataLoader(ImageDataLoader):
    """MNIST from CSV: header line, then ``label,p0..p783`` (ref mnist_data_loader.hpp:58-106)."""
    num_classes, image_shape = 10, (1, 28, 28)
    class_names = [str(i) for i in range(10)]

    def load_data(self, source: str) -> bool:
        if not os.path.exists(source):
            print(f"Error: Could not open file {source}")
            return False
        if _hostlib.available():
            with open(source, "rb") as f:
                max_rows = max(1, sum(1 for _ in f) - 1)
            img = np.empty((max_rows, 784), np.float32)
            lab = np.empty((max_rows,), np.int32)
            n = _hostlib.lib().dcnn_load_mnist_csv(source.encode(), img.ctypes.data_as(C.c_void_p),
                                                   lab.ctypes.data_as(C.c_void_p), C.c_longlong(max_rows), 784)
            if n < 0:
                return False
            img, lab = img[:n], lab[:n]
        else:
            raw = np.loadtxt(source, delimiter=",", skiprows=1, dtype=np.float64, ndmin=2)
            lab = raw[:, 0].astype(np.int32)
            img = (raw[:, 1:785] / 255.0).astype(np.float32)
        self.images = np.ascontiguousarray(img.reshape(-1, 1, 28, 28))
        self.labels = np.ascontiguousarray(lab)
        self.current_index = 0
        print(f"Loaded {self.size()} samples from {source}")
        return self.size() > 0


class _CIFARBase(ImageDataLoader):
    image_shape = (3, 32, 32)
    label_bytes = 1

    def __init__(self):
        super().__init__()
        self.coarse_labels = np.zeros((0,), np.int32)

    def load_data(self, source: str) -> bool:
        return self.load_multiple_files([source])

    def load_multiple_files(self, filenames: Sequence[str]) -> bool:
        imgs, labs, coarse = [], [], []
        rec = self.label_bytes + 3072
        for fn in filenames:
            if not os.path.exists(fn):
                print(f"Error: Could not open file {fn}")
                return False
            n_max = os.path.getsize(fn) // rec
            img = np.empty((n_max, 3072), np.float32)
            lab = np.empty((n_max,), np.int32)
            co = np.empty((n_max,), np.int32)
            if _hostlib.available():
                n = _hostlib.lib().dcnn_load_cifar_bin(fn.encode(), self.label_bytes, 3072,
                                                       img.ctypes.data_as(C.c_void_p), lab.ctypes.data_as(C.c_void_p),
                                                       co.ctypes.data_as(C.c_void_p), C.c_longlong(n_max))
            else:
                raw = np.fromfile(fn, dtype=np.uint8)[:n_max * rec].reshape(n_max, rec)
                lab[:] = raw[:, self.label_bytes - 1]
                co[:] = raw[:, 0]
                img[:] = raw[:, self.label_bytes:].astype(np.float32) / 255.0
                n = n_max
            imgs.append(img[:n])
            labs.append(lab[:n])
            coarse.append(co[:n])
            print(f"Loaded {n} samples from {fn}")
        self.images = np.ascontiguousarray(np.concatenate(imgs).reshape(-1, 3, 32, 32))
        self.labels = np.ascontiguousarray(np.concatenate(labs))
        self.coarse_labels = np.concatenate(coarse)
        self.current_index = 0
        return self.size() > 0


class CIFAR10DataLoader(_CIFARBase):
    num_classes, label_bytes = 10, 1
    class_names = ["airplane", "automobile", "bird", "cat", "deer", "dog", "frog", "horse", "ship", "truck"]


class CIFAR100DataLoader(_CIFARBase):
    """Records are <coarse label><fine label><3072 pixels>; the fine label is the target unless
    ``set_use_coarse_labels(True)`` selects the 20 superclasses (cifar100_data_loader.hpp)."""
    num_classes, label_bytes = 100, 2
    class_names = [f"class_{i}" for i in range(100)]

    def __init__(self, use_coarse_labels: bool = False):
        super().__init__()
        self.fine_labels = np.zeros((0,), np.int32)
        self._coarse = False
        self._want_coarse = bool(use_coarse_labels)

    def load_multiple_files(self, filenames):
        self._coarse = False
        self.num_classes = 100
        ok = super().load_multiple_files(filenames)
        if ok:
            self.fine_labels = self.labels
            self.set_use_coarse_labels(self._want_coarse)
        return ok

    def set_use_coarse_labels(self, use_coarse: bool):
        """Targets become the 20 coarse superclasses (True) or the 100 fine classes (False)."""
        self._want_coarse = self._coarse = bool(use_coarse)
        if self.fine_labels.shape[0] == 0:
            self.fine_labels = self.labels
        self.labels = np.ascontiguousarray(self.coarse_labels if self._coarse else self.fine_labels)
        self.num_classes = 20 if self._coarse else 100
        self.class_names = [f"{'superclass' if self._coarse else 'class'}_{i}" for i in range(self.num_classes)]
        self.batches_prepared = False

    def is_using_coarse_labels(self) -> bool:
        return self._coarse


class TinyImageNetDataLoader(ImageDataLoader):
    """Tiny-ImageNet-200 directory: wnids.txt, words.txt, train/<wnid>/images/*.JPEG, val/images + val_annotations.txt
    (ref tiny_imagenet_data_loader.hpp:55-626).  JPEG decode uses Pillow (the reference uses stb_image)."""
    num_classes, image_shape = 200, (3, 64, 64)

    def __init__(self):
        super().__init__()
        self.class_ids: List[str] = []
        self.class_id_to_index = {}
        self.class_id_to_name = {}

    def _load_class_ids(self, d):
        p = os.path.join(d, "wnids.txt")
        if not os.path.exists(p):
            print(f"Error: Could not open {p}")
            return False
        with open(p) as f:
            self.class_ids = [l.strip() for l in f if l.strip()]
        self.class_id_to_index = {w: i for i, w in enumerate(self.class_ids)}
        print(f"Loaded {len(self.class_ids)} class IDs")
        return len(self.class_ids) == self.num_classes

    def _load_class_names(self, d):
        p = os.path.join(d, "words.txt")
        if not os.path.exists(p):
            print(f"Warning: Could not open {p}")
            return
        with open(p) as f:
            for line in f:
                parts = line.rstrip("\n").split(None, 1)
                if parts and parts[0] in self.class_id_to_index:
                    self.class_id_to_name[parts[0]] = parts[1].strip() if len(parts) > 1 else ""
        self.class_names = [self.class_id_to_name.get(w, w) for w in self.class_ids]

    @staticmethod
    def _load_jpeg(path):
        from PIL import Image
        with Image.open(path) as im:
            im = im.convert("RGB")
            if im.size != (64, 64):
                return None
            a = np.asarray(im, dtype=np.float32) / 255.0
        return np.ascontiguousarray(a.transpose(2, 0, 1))

    def load_data(self, source: str, is_train: bool = True) -> bool:
        if not self._load_class_ids(source):
            return False
        self._load_class_names(source)
        imgs, labs = [], []
        if is_train:
            for wnid in self.class_ids:
                d = os.path.join(source, "train", wnid, "images")
                if not os.path.isdir(d):
                    continue
                for fn in sorted(os.listdir(d)):
                    if fn.lower().endswith((".jpeg", ".jpg")):
                        a = self._load_jpeg(os.path.join(d, fn))
                        if a is not None:
                            imgs.append(a)
                            labs.append(self.class_id_to_index[wnid])
        else:
            ann = os.path.join(source, "val", "val_annotations.txt")
            if not os.path.exists(ann):
                print(f"Error: Validation annotations not found: {ann}")
                return False
            with open(ann) as f:
                for line in f:
                    parts = line.split()
                    if len(parts) >= 2 and parts[1] in self.class_id_to_index:
                        a = self._load_jpeg(os.path.join(source, "val", "images", parts[0]))
                        if a is not None:
                            imgs.append(a)
                            labs.append(self.class_id_to_index[parts[1]])
        if not imgs:
            return False
        self.images = np.ascontiguousarray(np.stack(imgs))
        self.labels = np.asarray(labs, np.int32)
        print(f"Loaded {self.size()} {'training' if is_train else 'validation'} samples")
        return True


class SyntheticDataLoader(ImageDataLoader):
    """Random images / labels of a named shape with a fixed seed -- what every benchmark uses."""

    def __init__(self, num_samples: int, image_shape: Sequence[int], num_classes: int, seed: int = 0):
        self.image_shape = tuple(int(s) for s in image_shape)
        self.num_classes = int(num_classes)
        self.class_names = [str(i) for i in range(num_classes)]
        super().__init__()
        g = np.random.default_rng(seed)
        self.images = g.random((num_samples,) + self.image_shape, dtype=np.float32)
        self.labels = g.integers(0, num_classes, num_samples).astype(np.int32)

    def load_data(self, source: str = "") -> bool:
        return True


class RegressionDataLoader(BaseDataLoader):
    """features [N, F] -> targets [N, O]; batches are ``[n, F, 1, 1]`` / ``[n, O, 1, 1]`` (regression_data_loader.hpp)."""

    def __init__(self):
        super().__init__()
        self.features = np.zeros((0, 0), np.float32)
        self.targets = np.zeros((0, 0), np.float32)

    def size(self):
        return int(self.features.shape[0])

    def get_num_features(self):
        return int(self.features.shape[1])

    def get_num_outputs(self):
        return int(self.targets.shape[1])

    def _make_batch(self, idx):
        n = len(idx)
        return (torch.from_numpy(self.features[idx]).reshape(n, -1, 1, 1).contiguous(),
                torch.from_numpy(self.targets[idx]).reshape(n, -1, 1, 1).contiguous())


class WiFiDataLoader(RegressionDataLoader):
    """UJIIndoorLoc-style CSV: RSSI feature columns then target columns (ref wifi_data_loader.hpp:24-420).
    ``normalize_data()`` z-scores features (and targets for regression) and remembers the statistics so that the
    validation loader can reuse them (``normalize_with``) and predictions can be mapped back (``denormalize_targets``)."""

    def __init__(self, is_regression: bool = True):
        super().__init__()
        self.is_regression = is_regression
        self.is_normalized = False
        self.feature_means = self.feature_stds = self.target_means = self.target_stds = None

    def load_data(self, filename: str, feature_start_col: int = 0, feature_end_col: int = 0,
                  target_start_col: int = 0, target_end_col: int = 0, has_header: bool = True) -> bool:
        """CSV of RSSI features followed by the targets, parsed like the reference (wifi_data_loader.hpp:56-125):
        ``feature_end_col == 0`` means "the last two columns are the targets" (the last one for classification);
        a feature that reads 100 (the UJI "access point not seen" marker), 0 or is not a number becomes -100; a target
        that is not a number becomes 0; rows are never dropped."""
        if not os.path.exists(filename):
            print(f"Error: Could not open file {filename}")
            return False
        feats, tgts = [], []

        def num(cell):
            try:
                return float(cell)
            except ValueError:
                return None
        with open(filename) as f:
            for k, line in enumerate(f):
                if k == 0 and has_header:
                    continue
                row = line.rstrip("\n").rstrip("\r").split(",")
                if not line.strip():
                    continue
                if feature_end_col == 0:
                    tail = 2 if self.is_regression else 1
                    feature_end_col, target_start_col, target_end_col = len(row) - tail, len(row) - tail, len(row)
                fr = []
                for cell in row[feature_start_col:min(feature_end_col, len(row))]:
                    v = num(cell)
                    fr.append(-100.0 if v is None or v == 100.0 or v == 0.0 or v != v else v)
                tr = []
                for cell in row[target_start_col:min(target_end_col, len(row))]:
                    v = num(cell)
                    tr.append(0.0 if v is None else v)
                if fr and tr:
                    feats.append(fr)
                    tgts.append(tr)
        if not feats:
            print(f"Error: No valid data loaded from {filename}")
            return False
        self.features = np.ascontiguousarray(np.asarray(feats, dtype=np.float32))
        self.targets = np.ascontiguousarray(np.asarray(tgts, dtype=np.float32))
        print(f"Loaded {self.size()} samples from {filename}")
        print(f"Features: {self.get_num_features()}, Outputs: {self.get_num_outputs()}")
        return self.size() > 0

    def normalize_data(self):
        if self.size() == 0:
            print("Warning: No data to normalize!")
            return
        self.feature_means = self.features.mean(0)
        self.feature_stds = self.features.std(0)
        self.feature_stds[self.feature_stds < 1e-8] = 1.0
        self.features = ((self.features - self.feature_means) / self.feature_stds).astype(np.float32)
        if self.is_regression:
            self.target_means = self.targets.mean(0)
            self.target_stds = self.targets.std(0)
            self.target_stds[self.target_stds < 1e-8] = 1.0
            self.targets = ((self.targets - self.target_means) / self.target_stds).astype(np.float32)
        self.is_normalized = True

    def normalize_with(self, other: "WiFiDataLoader"):
        self.feature_means, self.feature_stds = other.feature_means, other.feature_stds
        self.features = ((self.features - self.feature_means) / self.feature_stds).astype(np.float32)
        if self.is_regression and other.target_means is not None:
            self.target_means, self.target_stds = other.target_means, other.target_stds
            self.targets = ((self.targets - self.target_means) / self.target_stds).astype(np.float32)
        self.is_normalized = True

    def get_feature_means(self):
        return None if self.feature_means is None else self.feature_means.copy()

    def get_feature_stds(self):
        return None if self.feature_stds is None else self.feature_stds.copy()

    def get_target_means(self):
        return None if self.target_means is None else self.target_means.copy()

    def get_target_stds(self):
        return None if self.target_stds is None else self.target_stds.copy()

    def print_statistics(self):
        print(f"WiFi dataset: {self.size()} samples, {self.get_num_features()} features, "
              f"{self.get_num_outputs()} outputs, normalized: {self.is_normalized}")
        if self.size():
            print(f"  feature range [{float(self.features.min()):.4g}, {float(self.features.max()):.4g}], "
                  f"target range [{float(self.targets.min()):.4g}, {float(self.targets.max()):.4g}]")

    def denormalize_targets(self, t: torch.Tensor) -> torch.Tensor:
        if not self.is_normalized or self.target_means is None:
            return t
        shape = t.shape
        m = torch.from_numpy(self.target_means).to(t.device)
        s = torch.from_numpy(self.target_stds).to(t.device)
        return (t.reshape(shape[0], -1) * s + m).reshape(shape)


def create_mnist_data_loaders(train_csv: str, test_csv: str):
    tr, te = MNISTDataLoader(), MNISTDataLoader()
    if not tr.load_data(train_csv) or not te.load_data(test_csv):
        raise RuntimeError("Failed to load MNIST data")
    return tr, te


def create_cifar10_data_loaders(data_dir: str):
    tr, te = CIFAR10DataLoader(), CIFAR10DataLoader()
    ok = tr.load_multiple_files([os.path.join(data_dir, f"data_batch_{i}.bin") for i in range(1, 6)])
    ok = ok and te.load_data(os.path.join(data_dir, "test_batch.bin"))
    if not ok:
        raise RuntimeError("Failed to load CIFAR-10 data")
    return tr, te


def create_cifar100_data_loaders(data_dir: str):
    tr, te = CIFAR100DataLoader(), CIFAR100DataLoader()
    if not tr.load_data(os.path.join(data_dir, "train.bin")) or not te.load_data(os.path.join(data_dir, "test.bin")):
        raise RuntimeError("Failed to load CIFAR-100 data")
    return tr, te


def create_tiny_imagenet_data_loaders(data_dir: str):
    tr, te = TinyImageNetDataLoader(), TinyImageNetDataLoader()
    if not tr.load_data(data_dir, True) or not te.load_data(data_dir, False):
        raise RuntimeError("Failed to load Tiny ImageNet data")
    return tr, te


def create_mnist_data_loaders_dir(data_dir: str):
    """``<dir>/train.csv`` + ``<dir>/test.csv`` (the reference examples read ./data/mnist/{train,test}.csv)."""
    return create_mnist_data_loaders(os.path.join(data_dir, "train.csv"), os.path.join(data_dir, "test.csv"))


def create_data_loader(dataset_type: str):
    """An empty loader for a dataset name (the reference declares this factory and returns nullptr from it,
    data_loader.hpp:192-196; here it works): mnist | cifar10 | cifar100 | tiny_imagenet | wifi | regression."""
    kinds = {"mnist": MNISTDataLoader, "cifar10": CIFAR10DataLoader, "cifar100": CIFAR100DataLoader,
             "tiny_imagenet": TinyImageNetDataLoader, "tinyimagenet": TinyImageNetDataLoader,
             "wifi": WiFiDataLoader, "uji": WiFiDataLoader, "regression": RegressionDataLoader}
    key = dataset_type.strip().lower().replace("-", "_")
    if key not in kinds:
        raise ValueError(f"unknown dataset type '{dataset_type}' (known: {sorted(kinds)})")
    return kinds[key]()


def train_val_split(dataset_size: int, val_ratio: float = 0.2, seed: int = 42):
    """(train indices, validation indices) of a shuffled split (ref data_loader.hpp:205-222: the last
    ``int(size * val_ratio)`` shuffled indices form the validation set)."""
    idx = np.random.default_rng(int(seed)).permutation(int(dataset_size))
    val = int(dataset_size * val_ratio)
    return idx[:dataset_size - val].copy(), idx[dataset_size - val:].copy()


def normalize_to_unit_range(data, max_value: float = 255.0):
    """In place ``data /= max_value`` for tensors and arrays (ref data_loader.hpp:182-186)."""
    data /= max_value
    return data
