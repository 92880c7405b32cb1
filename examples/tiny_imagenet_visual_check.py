#!/usr/bin/env python
"""Visual sanity check of the Tiny-ImageNet loader and augmentations -- counterpart of the reference's
examples/tiny_imagenet_visual_check.cpp (which writes PNGs with stb_image_write; no window, no GL).  Writes, for the first
NUM_IMAGES samples, ``<OUT_DIR>/sample_<i>_<class>.png`` and ``..._aug.png`` (after crop / rotation / flip / brightness).
With no dataset on disk (or SYNTHETIC=1) synthetic images of the same shape are used so the path can still be checked."""
import os

import numpy as np
from _common import get_env, loaders, setup
from PIL import Image

from dcnn_b200.data import AugmentationBuilder, create_tiny_imagenet_data_loaders


def to_png(chw, path):
    a = np.clip(chw.permute(1, 2, 0).numpy(), 0.0, 1.0)
    Image.fromarray((a * 255.0 + 0.5).astype(np.uint8)).save(path)


def main():
    cfg = setup()
    out_dir = get_env("OUT_DIR", "./visual_check")
    n = get_env("NUM_IMAGES", 8)
    os.makedirs(out_dir, exist_ok=True)
    train_loader, _ = loaders(create_tiny_imagenet_data_loaders, "./data/tiny-imagenet-200", (3, 64, 64), 200, cfg,
                              n_train=max(n, 8), n_test=8)
    names = getattr(train_loader, "class_names", None) or []
    train_loader.prepare_batches(n)
    train_loader.reset()
    data, labels = train_loader.get_next_batch()
    aug = AugmentationBuilder().random_crop(1.0, 4).rotation(1.0, 10.0).horizontal_flip(0.5).brightness(1.0, 0.2).build()
    augmented = data.clone()
    aug.apply(augmented, labels)
    written = []
    for i in range(min(n, data.shape[0])):
        cls = int(labels[i].reshape(-1).argmax())
        tag = (names[cls].split(",")[0].replace(" ", "_") if cls < len(names) else f"class{cls}")
        for suffix, img in (("", data[i]), ("_aug", augmented[i])):
            path = os.path.join(out_dir, f"sample_{i}_{tag}{suffix}.png")
            to_png(img, path)
            written.append(path)
    print(f"Wrote {len(written)} PNG files to {out_dir}")
    print(f"pixel range: [{float(data.min()):.3f}, {float(data.max()):.3f}], mean {float(data.mean()):.3f}")


if __name__ == "__main__":
    main()
