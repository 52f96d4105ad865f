"""Data side: `CustomImages` / `CelebA` against fixtures written here in the reference's on-disk layout
(`datasets/custom_images.py:8-28`, `datasets/celeba.py:8-150`), and against the reference's own classes where `/root/reference`
exists (this container; the GPU box skips that half)."""
import importlib.util
import os

import numpy as np
import pytest
import torch
from PIL import Image

from stablekeypoints_b200 import datasets as D
from stablekeypoints_b200.optimize import _make_dataset

REF = "/root/reference/datasets"


def _ref_module(name):
    path = os.path.join(REF, name + ".py")
    if not os.path.exists(path):
        pytest.skip("reference tree not present")
    spec = importlib.util.spec_from_file_location("ref_ds_" + name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _image(rng, w, h):
    return Image.fromarray(rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8), "RGB")


@pytest.fixture()
def celeba_tree(tmp_path):
    """Six numbered images (aligned .png 178x218, wild .jpg of varying size), landmark / bbox tables, MAFL lists."""
    rng = np.random.default_rng(0)
    root = tmp_path / "celeba"
    (root / "Anno").mkdir(parents=True)
    (root / "MAFL").mkdir()
    (root / "Img" / "img_align_celeba_png").mkdir(parents=True)
    (root / "Img" / "img_celeba").mkdir(parents=True)
    n = 6
    sizes = [(300, 400), (640, 480), (200, 200), (500, 350), (410, 333), (256, 512)]
    lm_a, lm_w, bb = ["%d" % n, "lefteye_x lefteye_y ..."], ["%d" % n, "lefteye_x lefteye_y ..."], ["%d" % n, "image_id x_1 y_1 width height"]
    for i in range(n):
        name = f"{i + 1:06d}"
        _image(rng, 178, 218).save(root / "Img" / "img_align_celeba_png" / (name + ".png"))
        w, h = sizes[i]
        _image(rng, w, h).save(root / "Img" / "img_celeba" / (name + ".jpg"), quality=95)
        lm_a.append(name + ".jpg " + " ".join(str(int(v)) for v in rng.integers(10, 170, size=10)))
        lm_w.append(name + ".jpg " + " ".join(str(int(v)) for v in rng.integers(10, 190, size=10)))
        frac = [0.9, 0.1, 0.5, 0.25, 0.35, 0.05][i]                 # share of the picture the face box covers
        bw = int(w * frac ** 0.5)
        bb.append(f"{name}.jpg 5 7 {bw} {int(h * frac ** 0.5) + 1}")
    (root / "Anno" / "list_landmarks_align_celeba.txt").write_text("\n".join(lm_a) + "\n")
    (root / "Anno" / "list_landmarks_celeba.txt").write_text("\n".join(lm_w) + "\n")
    (root / "Anno" / "list_bbox_celeba.txt").write_text("\n".join(bb) + "\n")
    (root / "MAFL" / "training.txt").write_text("000002.jpg\n000004.jpg\n000005.jpg\n000001.jpg\n")
    (root / "MAFL" / "testing.txt").write_text("000006.jpg\n000003.jpg\n")
    return str(root)


def test_custom_images_contract_and_reference(tmp_path):
    rng = np.random.default_rng(1)
    for name, (w, h) in {"b.png": (300, 200), "a.jpg": (640, 480), "c.png": (512, 512)}.items():
        _image(rng, w, h).save(tmp_path / name)
    (tmp_path / "sub").mkdir()                                      # directories are not images
    ds = D.CustomImages(str(tmp_path), 512)
    assert ds.image_files == ["a.jpg", "b.png", "c.png"] and len(ds) == 3
    s = ds[1]
    assert s["img"].shape == (3, 512, 512) and s["img"].dtype == torch.float32 and 0.0 <= float(s["img"].min()) and float(s["img"].max()) <= 1.0
    assert s["kpts"].shape == (15, 2) and s["visibility"].shape == (15,) and not s["kpts"].any()
    # an image already at the target size passes through unchanged
    want = torch.from_numpy(np.array(Image.open(tmp_path / "c.png"))).permute(2, 0, 1).float() / 255
    assert torch.equal(ds[2]["img"], want)
    ref = _ref_module("custom_images").CustomDataset(str(tmp_path), 512)
    for i in range(3):
        a, b = ds[i], ref[i]
        assert torch.equal(a["img"], b["img"]) and torch.equal(a["kpts"], b["kpts"]) and torch.equal(a["visibility"], b["visibility"])


def test_celeba_aligned_splits_keypoints_and_reference(celeba_tree):
    tr, te = D.CelebA(split="train", dataset_loc=celeba_tree), D.CelebA(split="test", dataset_loc=celeba_tree)
    assert tr.indices == [1, 3, 4, 0] and te.indices == [5, 2] and len(tr) == 4 and len(te) == 2
    s = tr[0]
    assert s["img"].shape == (3, 512, 512) and s["kpts"].shape == (5, 2)
    line = open(os.path.join(celeba_tree, "Anno", "list_landmarks_align_celeba.txt")).read().splitlines()[2 + 1].split()[1:]
    xy = np.array(line, dtype=np.float32).reshape(5, 2)
    assert np.allclose(s["kpts"].numpy(), np.stack([xy[:, 1] / 218, xy[:, 0] / 178], 1))      # (row, col) / (height, width)
    assert len(D.CelebA(split="train", dataset_loc=celeba_tree, max_len=3)) == 3
    with pytest.raises(ValueError):
        D.CelebA(split="val", dataset_loc=celeba_tree)
    ref = _ref_module("celeba").CelebA(split="train", dataset_loc=celeba_tree)
    for i in range(len(tr)):
        a, b = tr[i], ref[i]
        assert torch.equal(a["img"], b["img"]) and torch.allclose(a["kpts"], b["kpts"], atol=0, rtol=0)


def test_celeba_wild_drops_small_faces_like_the_reference(celeba_tree):
    tr = D.CelebA(split="train", dataset_loc=celeba_tree, align=False)
    # training list = images 2, 4, 5, 1 with box shares 0.1, 0.25, 0.35, 0.9: threshold 0.3 keeps 5 and 1
    assert tr.indices == [4, 0]
    assert D.CelebA(split="train", dataset_loc=celeba_tree, align=False, iou_threshold=0.05).indices == [1, 3, 4, 0]
    s = tr[0]
    w, h = Image.open(tr.image_path(4)).size
    line = open(os.path.join(celeba_tree, "Anno", "list_landmarks_celeba.txt")).read().splitlines()[2 + 4].split()[1:]
    xy = np.array(line, dtype=np.float32).reshape(5, 2)
    assert np.allclose(s["kpts"].numpy(), np.stack([xy[:, 1] / h, xy[:, 0] / w], 1))
    ref = _ref_module("celeba").CelebA(split="train", dataset_loc=celeba_tree, align=False)
    assert len(ref) == len(tr)
    for i in range(len(tr)):
        assert torch.equal(tr[i]["img"], ref[i]["img"]) and torch.equal(tr[i]["kpts"], ref[i]["kpts"])


def test_dataset_names_resolve_like_the_reference_stages(celeba_tree, tmp_path):
    assert isinstance(_make_dataset("celeba_aligned", celeba_tree), D.CelebA)
    assert _make_dataset("celeba_aligned", celeba_tree, split="test").indices == [5, 2]
    assert _make_dataset("celeba_wild", celeba_tree).align is False
    assert _make_dataset("celeba_aligned", celeba_tree, max_len=2).max_len == 2
    (tmp_path / "imgs").mkdir()
    _image(np.random.default_rng(2), 64, 64).save(tmp_path / "imgs" / "x.png")
    assert len(_make_dataset("custom", str(tmp_path / "imgs"))) == 1
    own = D.SyntheticKeypoints(length=2, image_size=32)
    assert _make_dataset("anything", "~", dataset=own) is own                 # an explicit dataset always wins
    with pytest.raises(NotImplementedError):
        _make_dataset("taichi", "~")
