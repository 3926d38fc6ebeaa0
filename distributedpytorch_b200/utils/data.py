"""Datasets.

`BasicDataset` / `CarvanaDataset` keep the reference's contract (utils/dataloading.py:12-78): ids from the image
directory, one image + one mask file per id (mask name = id + suffix), PIL resize to `newsize = [W, H]`
(BICUBIC for images, NEAREST for masks), image HWC->CHW scaled to [0,1]; an item is
``{'image': float32 [C,H,W], 'mask': int64 [H,W]}``.

`SyntheticSegmentation` generates image/mask pairs of the same contract procedurally (no network, no Carvana
files on the training boxes): an ellipse "car" on a textured background, so a model can actually learn it and
convergence can be compared between implementations.  `write_synthetic_files` materialises such pairs as
`*.jpg` + `*_mask.gif` so the file-based path (and the unmodified reference) can be exercised end to end.
"""
from __future__ import annotations

import logging
import os
from os import listdir
from os.path import splitext
from pathlib import Path
from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch
from torch.utils.data import Dataset, Subset

DIR_IMG = Path("./data/train_hq/")
DIR_MASK = Path("./data/train_masks/")
DEFAULT_SIZE = (960, 640)  # (W, H), as in utils/train_utils.py:26


class BasicDataset(Dataset):
    def __init__(self, images_dir, masks_dir, newsize: Sequence[int] = DEFAULT_SIZE, mask_suffix: str = "",
                 raw_uint8: bool = False):
        self.raw_uint8 = raw_uint8  # True: image uint8 [C,H,W] (not divided by 255), mask uint8: normalised on the GPU
        self.images_dir = Path(images_dir)
        self.masks_dir = Path(masks_dir)
        self.newsize = (int(newsize[0]), int(newsize[1]))
        self.mask_suffix = mask_suffix
        if not self.images_dir.is_dir():
            raise RuntimeError(f"Image directory {images_dir} does not exist")
        self.ids = sorted(splitext(f)[0] for f in listdir(self.images_dir) if not f.startswith("."))
        if not self.ids:
            raise RuntimeError(f"No input file found in {images_dir}, make sure you put your images there")
        logging.info(f"Creating dataset with {len(self.ids)} examples")

    def __len__(self) -> int:
        return len(self.ids)

    @staticmethod
    def preprocess(pil_img, newsize: Sequence[int], is_mask: bool, raw: bool = False) -> np.ndarray:
        from PIL import Image

        new_w, new_h = int(newsize[0]), int(newsize[1])
        if new_w <= 0 or new_h <= 0:
            raise ValueError("Scale is too small, resized images would have no pixel")
        pil_img = pil_img.resize((new_w, new_h), resample=Image.NEAREST if is_mask else Image.BICUBIC)
        arr = np.asarray(pil_img)
        if is_mask:
            return arr
        arr = arr[np.newaxis, ...] if arr.ndim == 2 else arr.transpose((2, 0, 1))
        return arr if raw else arr / 255

    @staticmethod
    def load(filename):
        from PIL import Image

        ext = splitext(str(filename))[1]
        if ext in (".npz", ".npy"):
            return Image.fromarray(np.load(filename))
        if ext in (".pt", ".pth"):
            return Image.fromarray(torch.load(filename).numpy())
        return Image.open(filename)

    def __getitem__(self, idx: int) -> Dict[str, torch.Tensor]:
        name = self.ids[idx]
        mask_file = list(self.masks_dir.glob(name + self.mask_suffix + ".*"))
        img_file = list(self.images_dir.glob(name + ".*"))
        if len(mask_file) != 1:
            raise RuntimeError(f"Either no mask or multiple masks found for the ID {name}: {mask_file}")
        if len(img_file) != 1:
            raise RuntimeError(f"Either no image or multiple images found for the ID {name}: {img_file}")
        mask = self.load(mask_file[0])
        img = self.load(img_file[0])
        if img.size != mask.size:
            raise RuntimeError(f"Image and mask {name} should be the same size, but are {img.size} and {mask.size}")
        img_a = self.preprocess(img, self.newsize, is_mask=False, raw=self.raw_uint8)
        mask_a = self.preprocess(mask, self.newsize, is_mask=True)
        if self.raw_uint8 and img_a.dtype == np.uint8:
            return {"image": torch.from_numpy(np.ascontiguousarray(img_a)),
                    "mask": torch.from_numpy(np.ascontiguousarray(mask_a).astype(np.uint8))}
        if self.raw_uint8:  # not an 8-bit image after all: fall back to the normalised float form
            img_a = img_a / 255
        return {"image": torch.as_tensor(img_a.copy()).float().contiguous(),
                "mask": torch.as_tensor(mask_a.copy()).long().contiguous()}


class CarvanaDataset(BasicDataset):
    def __init__(self, images_dir, masks_dir, newsize: Sequence[int] = DEFAULT_SIZE, raw_uint8: bool = False):
        super().__init__(images_dir, masks_dir, newsize, mask_suffix="_mask", raw_uint8=raw_uint8)


def _synthetic_pair(index: int, width: int, height: int, seed: int) -> Tuple[np.ndarray, np.ndarray]:
    """Deterministic (image uint8 [H,W,3], mask uint8 [H,W] in {0,1}) for a given index."""
    rng = np.random.default_rng(seed * 1_000_003 + index)
    yy, xx = np.mgrid[0:height, 0:width].astype(np.float32)
    cx, cy = rng.uniform(0.3, 0.7) * width, rng.uniform(0.35, 0.65) * height
    ax, ay = rng.uniform(0.15, 0.3) * width, rng.uniform(0.12, 0.25) * height
    mask = (((xx - cx) / ax) ** 2 + ((yy - cy) / ay) ** 2 <= 1.0).astype(np.uint8)
    base = rng.uniform(40, 120, size=3).astype(np.float32)
    car = rng.uniform(130, 230, size=3).astype(np.float32)
    shade = 1.0 + 0.15 * np.sin(xx / width * rng.uniform(4, 9)) * np.cos(yy / height * rng.uniform(3, 7))
    img = base[None, None, :] * shade[..., None]
    img = np.where(mask[..., None] == 1, car[None, None, :] * (0.85 + 0.15 * shade[..., None]), img)
    img = img + rng.normal(0, 6.0, size=img.shape).astype(np.float32)
    return np.clip(img, 0, 255).astype(np.uint8), mask


class SyntheticSegmentation(Dataset):
    """Procedural stand-in for Carvana with the reference's item contract."""

    def __init__(self, length: int = 64, newsize: Sequence[int] = DEFAULT_SIZE, seed: int = 0, raw_uint8: bool = False):
        self.raw_uint8 = raw_uint8
        self.length = int(length)
        self.newsize = (int(newsize[0]), int(newsize[1]))
        self.seed = seed
        logging.info(f"Creating synthetic dataset with {self.length} examples")

    def __len__(self) -> int:
        return self.length

    def __getitem__(self, idx: int) -> Dict[str, torch.Tensor]:
        img, mask = _synthetic_pair(idx, self.newsize[0], self.newsize[1], self.seed)
        if self.raw_uint8:
            return {"image": torch.from_numpy(np.ascontiguousarray(img.transpose(2, 0, 1))),
                    "mask": torch.from_numpy(np.ascontiguousarray(mask))}
        return {"image": torch.from_numpy(img.transpose(2, 0, 1).astype(np.float32) / 255.0).contiguous(),
                "mask": torch.from_numpy(mask.astype(np.int64)).contiguous()}


def write_synthetic_files(images_dir, masks_dir, count: int, size: Sequence[int] = (96, 64), seed: int = 0,
                          image_ext: str = ".jpg", mask_suffix: str = "_mask", mask_ext: str = ".gif") -> List[str]:
    """Write `count` synthetic pairs as Carvana-style files (`<id>.jpg`, `<id>_mask.gif`, mask values {0,1})."""
    from PIL import Image

    os.makedirs(images_dir, exist_ok=True)
    os.makedirs(masks_dir, exist_ok=True)
    ids = []
    for i in range(count):
        img, mask = _synthetic_pair(i, int(size[0]), int(size[1]), seed)
        name = f"synth_{i:05d}"
        Image.fromarray(img).save(os.path.join(images_dir, name + image_ext))
        Image.fromarray(mask, mode="L").save(os.path.join(masks_dir, name + mask_suffix + mask_ext))
        ids.append(name)
    return ids


def build_dataset(images_dir=DIR_IMG, masks_dir=DIR_MASK, newsize: Sequence[int] = DEFAULT_SIZE,
                  synthetic_length: int = 0, seed: int = 0, raw_uint8: bool = False) -> Dataset:
    """Carvana files if present, plain image/mask files otherwise.  Synthetic data ONLY on request
    (`synthetic_length > 0`, i.e. DPT_SYNTHETIC=N for the trainers): a missing or mistyped data directory raises, as it
    does in the reference, instead of quietly training on procedural ellipses and writing a checkpoint.

    The reference's fallback (utils/train_utils.py:27-32) catches (AssertionError, RuntimeError) around the
    Carvana constructor, which can never succeed when the directory is missing; this one can."""
    if synthetic_length <= 0:
        try:
            ds = CarvanaDataset(images_dir, masks_dir, newsize, raw_uint8=raw_uint8)
            ds[0]
            logging.info("Carvana Dataset")
            return ds
        except (RuntimeError, OSError):
            try:
                ds = BasicDataset(images_dir, masks_dir, newsize, raw_uint8=raw_uint8)
                ds[0]
                logging.info("Basic Dataset")
                return ds
            except (RuntimeError, OSError) as e:
                raise RuntimeError(f"no usable images in {images_dir!r} / masks in {masks_dir!r} ({e}); put the dataset "
                                   "there or set DPT_SYNTHETIC=<number of samples> to train on synthetic data") from e
    logging.info("Synthetic Dataset")
    return SyntheticSegmentation(synthetic_length, newsize, seed, raw_uint8=raw_uint8)


def split_dataset(dataset: Dataset, val_percent: float, seed: int = 0) -> Tuple[Subset, Subset]:
    """Train/validation split; `val_percent` is a PERCENTAGE (train.py:18, utils/train_utils.py:35-37)."""
    n_val = int(len(dataset) * val_percent / 100)
    n_train = len(dataset) - n_val
    gen = torch.Generator().manual_seed(seed)
    train, val = torch.utils.data.random_split(dataset, [n_train, n_val], generator=gen)
    return train, val


class ShardedSampler(torch.utils.data.Sampler):
    """DistributedSampler equivalent (utils/train_utils.py:189) without a process-group dependency: pads to a
    multiple of `world`, shuffles with seed + epoch (the reference never calls set_epoch, so it replays the same
    order every epoch), rank r takes indices r, r+world, ..."""

    def __init__(self, length: int, rank: int, world: int, shuffle: bool = True, seed: int = 0):
        self.length, self.rank, self.world, self.shuffle, self.seed = length, rank, world, shuffle, seed
        self.epoch = 0
        self.num_samples = (length + world - 1) // world

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def __len__(self) -> int:
        return self.num_samples

    def __iter__(self):
        if self.shuffle:
            g = torch.Generator().manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.length, generator=g).tolist()
        else:
            idx = list(range(self.length))
        total = self.num_samples * self.world
        idx += idx[: total - len(idx)]
        return iter(idx[self.rank:total:self.world])
