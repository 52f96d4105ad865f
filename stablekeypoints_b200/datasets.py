"""Data side.  The sample-dict contract of the reference's `datasets/*.py`
(`datasets/celeba.py:93-113`, `custom_images.py:20-25`):
    {"img": float (3,S,S) in [0,1], "kpts": (n,2) as (row, col) / image size, ["visibility": (n,)]}.
 * `SyntheticKeypoints`  -- what BASELINE.json measures on (512x512 synthetic images);
 * `CustomImages`        -- a folder of images (`datasets/custom_images.py:8-28`, `--dataset_name custom`);
 * `CelebA`              -- aligned / in-the-wild CelebA with the MAFL split lists (`datasets/celeba.py:8-150`,
                            `--dataset_name celeba_aligned | celeba_wild`).
CUB / Taichi / Human3.6M / DeepFashion read `.mat` / `.h5` annotation files (scipy.io / h5py, the latter absent from this image):
pass such a dataset object via `dataset=`.  Decoding stays on the host (PIL, the reference's own resampling filter: the
pixels are then identical to the reference's), the hot path starts at the H2D copy of the (B,3,512,512) batch.
"""
from __future__ import annotations

import os

import torch


class SyntheticKeypoints(torch.utils.data.Dataset):
    """Deterministic per-index images: smooth random blobs (so attention maps have structure) +
    `n_kpts` ground-truth points at the blob centres.  Image i depends only on (seed, i)."""

    def __init__(self, length: int = 64, image_size: int = 512, n_kpts: int = 5, seed: int = 0):
        self.length, self.image_size, self.n_kpts, self.seed = length, image_size, n_kpts, seed

    def __len__(self):
        return self.length

    def __getitem__(self, i):
        g = torch.Generator().manual_seed(self.seed * 1_000_003 + int(i))
        s = self.image_size
        kp = torch.rand(self.n_kpts, 2, generator=g) * 0.8 + 0.1
        low = torch.rand(3, s // 16, s // 16, generator=g)
        img = torch.nn.functional.interpolate(low[None], size=(s, s), mode="bilinear", align_corners=False)[0] * 0.5
        ys = (torch.arange(s, dtype=torch.float32) + 0.5) / s
        for j in range(self.n_kpts):
            blob = torch.exp(-((ys[:, None] - kp[j, 0]) ** 2 + (ys[None, :] - kp[j, 1]) ** 2) / (2 * 0.03 ** 2))
            img[j % 3] = img[j % 3] + 0.5 * blob
        return {"img": img.clamp_(0, 1), "kpts": kp, "visibility": torch.ones(self.n_kpts)}


def _load_rgb(path: str, size: int) -> torch.Tensor:
    """RGB image file -> float (3, size, size) in [0, 1], PIL bilinear resize (the filter both reference loaders use:
    `celeba.py:105`, torchvision `Resize` on a PIL image in `custom_images.py:16`)."""
    import numpy as np
    from PIL import Image
    with Image.open(path) as im:
        arr = np.asarray(im.convert("RGB").resize((size, size), Image.BILINEAR))
    return torch.from_numpy(arr.copy()).permute(2, 0, 1).float().div_(255.0)


class CustomImages(torch.utils.data.Dataset):
    """Every regular file of `data_root`, sorted by name (`datasets/custom_images.py:11-13`).  No annotations: `kpts` and
    `visibility` are the reference's zero placeholders for 15 points (`custom_images.py:24`)."""

    def __init__(self, data_root: str, image_size: int = 512):
        self.data_root = os.path.expanduser(data_root)
        self.image_size = image_size
        self.image_files = sorted(f for f in os.listdir(self.data_root) if os.path.isfile(os.path.join(self.data_root, f)))

    def __len__(self):
        return len(self.image_files)

    def __getitem__(self, i):
        img = _load_rgb(os.path.join(self.data_root, self.image_files[i]), self.image_size)
        return {"img": img, "kpts": torch.zeros(15, 2), "visibility": torch.zeros(15)}


class CelebA(torch.utils.data.Dataset):
    """CelebA faces, five landmarks, MAFL split (`datasets/celeba.py`).  Layout under `dataset_loc`:
    `Anno/list_landmarks[_align]_celeba.txt` (two header lines, then `name x1 y1 ... x5 y5` per image in file-number order),
    `MAFL/training.txt | testing.txt` (one `NNNNNN.jpg` per line), `Img/img_align_celeba_png/NNNNNN.png` (aligned) or
    `Img/img_celeba/NNNNNN.jpg` + `Anno/list_bbox_celeba.txt` (in the wild: images whose face box covers less than
    `iou_threshold` of the picture are dropped, `celeba.py:52-72`).  `max_len` only caps `__len__` (`celeba.py:76-79`)."""

    num_kps = 5

    def __init__(self, max_len: int = -1, split: str = "train", align: bool = True, dataset_loc: str = "~", iou_threshold: float = 0.3,
                 image_size: int = 512):
        if split not in ("train", "test"):
            raise ValueError(f"split {split!r}: 'train' or 'test'")
        self.dataset_loc = os.path.expanduser(dataset_loc)
        self.align, self.split, self.max_len, self.image_size = align, split, max_len, image_size
        anno = os.path.join(self.dataset_loc, "Anno")
        with open(os.path.join(anno, "list_landmarks_align_celeba.txt" if align else "list_landmarks_celeba.txt")) as f:
            self._landmark_lines = f.read().splitlines()[2:]
        with open(os.path.join(self.dataset_loc, "MAFL", "training.txt" if split == "train" else "testing.txt")) as f:
            # file number (1-based in the name) -> 0-based row of the annotation tables
            self.indices = [int(line.split(".")[0]) - 1 for line in f.read().split()]
        if not align:
            from PIL import Image
            with open(os.path.join(anno, "list_bbox_celeba.txt")) as f:
                boxes = f.read().splitlines()[2:]
            keep = []
            for idx in self.indices:
                _, _, w, h = (int(v) for v in boxes[idx].split()[1:5])
                with Image.open(self.image_path(idx)) as im:
                    width, height = im.size
                if w * h >= width * height * iou_threshold:
                    keep.append(idx)
            self.indices = keep

    def image_path(self, idx: int) -> str:
        if self.align:
            return os.path.join(self.dataset_loc, "Img", "img_align_celeba_png", f"{idx + 1:06d}.png")
        return os.path.join(self.dataset_loc, "Img", "img_celeba", f"{idx + 1:06d}.jpg")

    def __len__(self):
        return len(self.indices) if self.max_len == -1 else self.max_len

    def __getitem__(self, i):
        from PIL import Image
        idx = self.indices[i]
        path = self.image_path(idx)
        with Image.open(path) as im:
            width, height = im.size
        xy = torch.tensor([float(v) for v in self._landmark_lines[idx].split()[1:11]]).reshape(5, 2)
        kpts = (xy / torch.tensor([float(width), float(height)]))[:, [1, 0]]          # (x, y) pixels -> (row, col) / size
        return {"img": _load_rgb(path, self.image_size), "kpts": kpts}


def make_dataset(dataset_name: str, dataset_loc: str, split: str = "train", max_len: int = -1):
    """`dataset_name` -> dataset, as the reference's stages construct them (`optimize.py:301-326` train, `eval.py:401-424` test)."""
    if dataset_name == "celeba_aligned":
        return CelebA(split=split, dataset_loc=dataset_loc, max_len=max_len)
    if dataset_name == "celeba_wild":
        return CelebA(split=split, dataset_loc=dataset_loc, align=False, max_len=max_len)
    if dataset_name == "custom":
        return CustomImages(data_root=dataset_loc, image_size=512)
    raise NotImplementedError(f"dataset {dataset_name!r} needs .mat / .h5 readers this package does not carry: pass a torch Dataset "
                              "via dataset= (sample dict with 'img', 'kpts'[, 'visibility'])")
