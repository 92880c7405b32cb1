"""Loader tests on generated fixture files in the reference's on-disk formats (the reference only tests its
Tiny-ImageNet loader and needs the real dataset for it)."""

import numpy as np
import pytest
import torch

from dcnn_b200 import _hostlib
from dcnn_b200.data import (AugmentationBuilder, CIFAR10DataLoader, CIFAR100DataLoader, MNISTDataLoader,
                            SyntheticDataLoader, TinyImageNetDataLoader, WiFiDataLoader)


def _mnist_csv(path, n, seed=0):
    g = np.random.default_rng(seed)
    lab = g.integers(0, 10, n)
    pix = g.integers(0, 256, (n, 784))
    with open(path, "w") as f:
        f.write("label," + ",".join(f"p{i}" for i in range(784)) + "\n")
        for l, p in zip(lab, pix):
            f.write(str(l) + "," + ",".join(map(str, p)) + "\n")
    return lab, pix


def test_mnist_csv(tmp_path):
    p = str(tmp_path / "train.csv")
    lab, pix = _mnist_csv(p, 37)
    ld = MNISTDataLoader()
    assert ld.load_data(p) and ld.size() == 37
    assert np.array_equal(ld.labels, lab)
    assert np.allclose(ld.images.reshape(37, 784), pix / 255.0, atol=1e-7)
    ld.set_seed(1)
    ld.shuffle()
    ld.prepare_batches(16)
    assert ld.num_batches() == 3
    sizes, total = [], 0
    for d, l in ld:
        assert d.shape[1:] == (1, 28, 28) and l.shape[1:] == (10, 1, 1)
        assert torch.all(l.sum(dim=1) == 1)
        sizes.append(d.shape[0])
    assert sorted(sizes) == [5, 16, 16]
    assert ld.get_next_batch() is None


def test_native_and_numpy_parsers_agree(tmp_path, monkeypatch):
    p = str(tmp_path / "t.csv")
    _mnist_csv(p, 11, seed=3)
    a = MNISTDataLoader()
    a.load_data(p)
    monkeypatch.setattr(_hostlib, "available", lambda: False)
    b = MNISTDataLoader()
    b.load_data(p)
    assert np.array_equal(a.labels, b.labels) and np.allclose(a.images, b.images, atol=1e-7)
    d1, l1 = a._make_batch(np.array([3, 1, 7]))
    monkeypatch.undo()
    d2, l2 = a._make_batch(np.array([3, 1, 7]))
    assert torch.equal(d1, d2) and torch.equal(l1, l2)


@pytest.mark.parametrize("cls,label_bytes,classes", [(CIFAR10DataLoader, 1, 10), (CIFAR100DataLoader, 2, 100)])
def test_cifar_bin(tmp_path, cls, label_bytes, classes):
    g = np.random.default_rng(0)
    n = 23
    rec = np.zeros((n, label_bytes + 3072), np.uint8)
    fine = g.integers(0, classes, n)
    rec[:, label_bytes - 1] = fine
    if label_bytes == 2:
        rec[:, 0] = g.integers(0, 20, n)
    rec[:, label_bytes:] = g.integers(0, 256, (n, 3072))
    p = str(tmp_path / "batch.bin")
    rec.tofile(p)
    ld = cls()
    assert ld.load_data(p) and ld.size() == n
    assert np.array_equal(ld.labels, fine)
    assert np.allclose(ld.images.reshape(n, 3072), rec[:, label_bytes:] / 255.0, atol=1e-7)
    ld.set_augmentation(AugmentationBuilder().random_crop(1.0, 4).horizontal_flip(0.5).cutout(0.5, 8)
                        .normalize((0.49, 0.48, 0.45), (0.25, 0.24, 0.26)).build())
    ld.set_seed(5)
    ld.prepare_batches(8)
    d, l = ld.get_next_batch()
    assert d.shape == (8, 3, 32, 32) and l.shape == (8, classes, 1, 1)


def test_tiny_imagenet_tree(tmp_path):
    from PIL import Image
    root = tmp_path / "tiny"
    wnids = [f"n{i:08d}" for i in range(200)]
    (root / "val" / "images").mkdir(parents=True)
    (root / "wnids.txt").write_text("\n".join(wnids) + "\n")
    (root / "words.txt").write_text("\n".join(f"{w}\tthing {i}, object" for i, w in enumerate(wnids)) + "\nn99999999\tother\n")
    g = np.random.default_rng(0)
    for w in wnids[:3]:
        d = root / "train" / w / "images"
        d.mkdir(parents=True)
        for k in range(2):
            Image.fromarray(g.integers(0, 256, (64, 64, 3), dtype=np.uint8)).save(d / f"{w}_{k}.JPEG", quality=95)
    ann = []
    for k in range(4):
        Image.fromarray(g.integers(0, 256, (64, 64, 3), dtype=np.uint8)).save(root / "val" / "images" / f"val_{k}.JPEG")
        ann.append(f"val_{k}.JPEG\t{wnids[k]}\t0\t0\t62\t62")
    (root / "val" / "val_annotations.txt").write_text("\n".join(ann) + "\n")
    tr, va = TinyImageNetDataLoader(), TinyImageNetDataLoader()
    assert tr.load_data(str(root), True) and tr.size() == 6
    assert va.load_data(str(root), False) and va.size() == 4
    assert list(va.labels) == [0, 1, 2, 3]
    assert tr.class_names[1] == "thing 1, object"
    tr.prepare_batches(4)
    d, l = tr.get_next_batch()
    assert d.shape == (4, 3, 64, 64) and l.shape == (4, 200, 1, 1) and 0 <= float(d.min()) and float(d.max()) <= 1


def test_wifi_regression_loader(tmp_path):
    g = np.random.default_rng(0)
    n = 40
    feats = g.integers(-100, 0, (n, 520)).astype(float)
    feats[g.random((n, 520)) < 0.5] = 100
    tg = g.random((n, 2)) * 100
    p = str(tmp_path / "uji.csv")
    np.savetxt(p, np.hstack([feats, tg]), delimiter=",", header=",".join(f"c{i}" for i in range(522)), comments="")
    tr, va = WiFiDataLoader(), WiFiDataLoader()
    assert tr.load_data(p) and va.load_data(p)
    assert tr.get_num_features() == 520 and tr.get_num_outputs() == 2
    tr.normalize_data()
    va.normalize_with(tr)
    assert np.allclose(tr.features.mean(0), 0, atol=1e-4) and np.allclose(va.targets, tr.targets, atol=1e-6)
    tr.prepare_batches(16)
    d, t = tr.get_next_batch()
    assert d.shape == (16, 520, 1, 1) and t.shape == (16, 2, 1, 1)
    back = tr.denormalize_targets(t)
    assert np.allclose(back.reshape(16, 2).numpy(), tg[:16], atol=1e-3)


def test_synthetic_loader_is_deterministic():
    a, b = SyntheticDataLoader(64, (3, 64, 64), 200, seed=7), SyntheticDataLoader(64, (3, 64, 64), 200, seed=7)
    a.prepare_batches(32)
    b.prepare_batches(32)
    (d1, l1), (d2, l2) = a.get_next_batch(), b.get_next_batch()
    assert torch.equal(d1, d2) and torch.equal(l1, l2) and l1.shape == (32, 200, 1, 1)


def test_shards_are_disjoint_equal_and_cover_the_data():
    import numpy as np
    from dcnn_b200.data import SyntheticDataLoader
    world, n = 3, 50
    seen, sizes = [], []
    for r in range(world):
        ld = SyntheticDataLoader(n, (1, 4, 4), 5, seed=0)
        ld.set_seed(99)                      # same seed on every replica -> same permutation
        ld.shuffle()
        ld.set_shard(r, world)
        assert ld.shard_size() == n // world
        idx = ld._indices()
        sizes.append(len(idx))
        seen.append(set(int(i) for i in idx))
        ld.prepare_batches(4)
        assert ld.num_batches() == -(-(n // world) // 4)
        total = sum(b[0].shape[0] for b in ld)
        assert total == n // world
    assert len(set(sizes)) == 1                                   # equal step counts on every replica
    assert all(not (a & b) for i, a in enumerate(seen) for b in seen[i + 1:])
    assert len(set().union(*seen)) == (n // world) * world        # all but the n % world remainder
    ld.set_shard(0, 1)                                            # world 1: sharding off
    assert len(ld._indices()) == n
    import pytest
    with pytest.raises(ValueError):
        ld.set_shard(3, 3)
    assert isinstance(np.asarray(ld._indices()), np.ndarray)


def test_prefetch_loader_streams_the_same_batches_with_fresh_augmentation_per_epoch():
    """PrefetchLoader == the sequence of loader.get_batch() calls (order, sharding, augmentation draws), ends epochs
    cleanly, restarts after an early shuffle(), and re-draws the augmentation every epoch."""
    from dcnn_b200.data import AugmentationBuilder, PrefetchLoader, SyntheticDataLoader

    def make():
        ld = SyntheticDataLoader(50, (3, 8, 8), 5, seed=3)
        ld.set_augmentation(AugmentationBuilder().horizontal_flip(0.5).brightness(0.5, 0.2).random_crop(0.5, 2).build())
        ld.set_seed(11)
        return ld
    ref, ld = make(), make()
    pf = PrefetchLoader(ld, 16, depth=2)
    assert pf.num_batches() == 4 and pf.size() == 50 and pf.get_image_shape() == [3, 8, 8]
    epochs = []
    for _ in range(2):
        ref.shuffle(); ref.reset()
        pf.shuffle(); pf.reset()
        got = []
        while True:
            want = ref.get_batch(16)
            have = pf.get_next_batch()
            if want is None:
                assert have is None
                break
            assert torch.equal(want[0], have[0]) and torch.equal(want[1], have[1])
            got.append(have[0].clone())
        assert [g.shape[0] for g in got] == [16, 16, 16, 2]
        epochs.append(torch.cat(got))
    assert not torch.equal(epochs[0], epochs[1])
    # abandoning an epoch half way: the next one starts from the top
    pf.shuffle(); pf.reset()
    first = pf.get_next_batch()[0].clone()
    pf.reset()
    assert pf.get_next_batch()[0].shape == first.shape
    n = 1
    while pf.get_next_batch() is not None:
        n += 1
    assert n == 4
    # sharding and drop_last
    pf2 = PrefetchLoader(make(), 8, depth=1, drop_last=True)
    pf2.set_shard(1, 2)
    assert pf2.num_batches() == 3
    assert sum(1 for _ in pf2) == 3
    pf.close(); pf2.close()


def test_prefetch_loader_surfaces_producer_errors():
    from dcnn_b200.data import PrefetchLoader, SyntheticDataLoader

    class Broken(SyntheticDataLoader):
        def get_batch(self, bs):
            if self.current_index >= 8:
                raise RuntimeError("disk on fire")
            return super().get_batch(bs)
    pf = PrefetchLoader(Broken(32, (1, 4, 4), 2), 8)
    pf.reset()
    assert pf.get_next_batch() is not None
    import pytest as _pytest
    with _pytest.raises(RuntimeError, match="disk on fire"):
        while pf.get_next_batch() is not None:
            pass


def test_prefetch_loader_is_reusable_over_epochs_of_different_length():
    """bench.py's end-to-end arm keeps ONE PrefetchLoader for its warm-up epoch and its timed epoch (so that the pinned
    staging ring and the side-stream buffers exist before the timed region): an epoch abandoned half-way, then epochs of
    different lengths, must each deliver exactly their own batches, in order."""
    from dcnn_b200.data import PrefetchLoader

    class Counted:
        def __init__(self):
            self.n, self.i = 0, 0

        def size(self):
            return self.n * 4

        def reset(self):
            self.i = 0

        def get_batch(self, batch_size):
            if self.i >= self.n:
                return None
            self.i += 1
            return torch.full((batch_size, 2), float(self.i)), torch.full((batch_size, 1), float(-self.i))

    src = Counted()
    pf = PrefetchLoader(src, 4, depth=2)
    src.n = 6
    pf.reset()
    assert float(pf.get_next_batch()[0][0, 0]) == 1.0          # ... and abandon the epoch with batches still queued
    for n in (3, 1, 5):
        src.n = n
        pf.reset()
        seen = []
        while True:
            b = pf.get_next_batch()
            if b is None:
                break
            assert float(b[1][0, 0]) == -float(b[0][0, 0])
            seen.append(int(b[0][0, 0]))
        assert seen == list(range(1, n + 1)), seen
    pf.close()
