"""Tiny-ImageNet folder dataset with an on-disk tensor cache (reference: ``tiny_imagenet/datasets.py:20-145``, which decodes
every JPEG with PIL once and pickles the arrays, ``:46-104``)."""
from __future__ import annotations

import os
import pickle
from typing import Optional, Tuple

import numpy as np
import torch


class TinyImageNet:
    """``root/train/<wnid>/images/*.JPEG`` and ``root/val/images/*.JPEG`` + ``val_annotations.txt``.  ``tensors()`` returns
    ``(x float [N,3,64,64] normalised, y int64 [N])``; decoded arrays are cached as ``<root>/<split>_cache.pkl``."""
    MEAN, STD = (0.5, 0.5, 0.5), (0.5, 0.5, 0.5)        # the reference's normalisation (tiny_imagenet/data_loader.py:48-49)

    def __init__(self, root: str, train: bool = True, dataidxs: Optional[np.ndarray] = None):
        self.root, self.split, self.dataidxs = root, "train" if train else "val", dataidxs
        self.data, self.targets = self._load()
        if dataidxs is not None:
            self.data, self.targets = self.data[dataidxs], self.targets[dataidxs]

    def _load(self) -> Tuple[np.ndarray, np.ndarray]:
        cache = os.path.join(self.root, f"{self.split}_cache.pkl")
        if os.path.exists(cache):
            with open(cache, "rb") as f:
                return pickle.load(f)
        from PIL import Image  # local import: only needed for the first decode
        wnids = sorted(os.listdir(os.path.join(self.root, "train")))
        cls = {w: i for i, w in enumerate(wnids)}
        items = []
        if self.split == "train":
            for w in wnids:
                d = os.path.join(self.root, "train", w, "images")
                items += [(os.path.join(d, f), cls[w]) for f in sorted(os.listdir(d))]
        else:
            with open(os.path.join(self.root, "val", "val_annotations.txt")) as f:
                for line in f:
                    name, w = line.split("\t")[:2]
                    items.append((os.path.join(self.root, "val", "images", name), cls[w]))
        data = np.stack([np.asarray(Image.open(p).convert("RGB"), dtype=np.uint8) for p, _ in items])
        targets = np.asarray([t for _, t in items], dtype=np.int64)
        with open(cache, "wb") as f:
            pickle.dump((data, targets), f)
        return data, targets

    def __len__(self):
        return len(self.targets)

    def tensors(self) -> Tuple[torch.Tensor, torch.Tensor]:
        x = torch.from_numpy(self.data).permute(0, 3, 1, 2).float() / 255.0
        m, s = torch.tensor(self.MEAN).view(1, 3, 1, 1), torch.tensor(self.STD).view(1, 3, 1, 1)
        return (x - m) / s, torch.from_numpy(self.targets)
