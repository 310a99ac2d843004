"""Drop-in `view_synthesis.datasets`: the reference's `SRNDataset`
(view_synthesis/datasets/dataset.py:10-94) with the same constructor, attributes, printed lines and `__getitem__` samples
(bit-identical arrays), but every view is decoded ONCE.

The reference opens the intrinsics file, decodes the PNG, divides it by 255.0 in float64 and parses the pose text file on
every `__getitem__` (:62-82), and train.py:70 asks for `train_batch_size` samples per iteration through a DataLoader with
num_workers = 0 (utils/util.py:87-88) -- a few milliseconds of host work per image against a 2.2 ms GPU chunk step.  Here
a view is kept as the uint8 crop the PNG holds (27 KB for a 128 x 128 SRN view; all of SRN cars is 3.4 GB) plus its float32
pose; `__getitem__` turns the crop into the float32 image through a 256-entry table of `np.float32(v / 255.0)`.

    CODENERF_DATASET_CACHE=0   decode on every access, exactly as the reference does

`codenerf_b200.data.DeviceSRN(dataset, device)` moves the decoded views into HBM (the device-resident stage).
PNG decoding uses PIL (the image the reference's `imageio.imread` returns for a PNG is the same array).
"""
import os
from pathlib import Path
from typing import Literal

import numpy as np
import torch

__all__ = ["SRNDataset"]

_UNIT = (np.arange(256, dtype=np.float64) / 255.0).astype(np.float32)      # dataset.py:72,90: (u8 / 255.0).astype(float32)


def _imread(path):
    from PIL import Image
    with Image.open(path) as im:
        return np.asarray(im)


class SRNDataset(torch.utils.data.Dataset):
    """
    Dataset from Scene Representation Networks (V. Sitzmann et al 2020)
    """

    def __init__(self, path: str, stage: Literal["train", "val", "test"] = "train"):
        """Same attributes and printed lines as the reference constructor (dataset.py:15-57): base_path, dataset_name, stage,
        intrinsic (sorted intrinsics files = one per object), num_objects, rgb_all_filenames / pose_all_filenames (lists of
        (object index, path), sorted per object), num_views."""
        super().__init__()
        root = Path(path)
        self.dataset_name = root.stem.split("_")[-1]                       # ".../srn_cars" -> "cars"
        self.stage = stage
        self.base_path = root / f"{self.dataset_name}_{stage}"
        print("Loading SRN dataset", self.base_path, "name:", self.dataset_name)
        assert self.base_path.exists(), f"{self.base_path} does not exist"
        if stage == "train" and "chair" in self.dataset_name:               # the chairs training split sits one level deeper
            nested = self.base_path / "chairs_2.0_train"
            self.base_path = nested if nested.exists() else self.base_path

        self.intrinsic = sorted(self.base_path.glob("*/intrinsics.txt"))
        self.num_objects = len(self.intrinsic)

        def views(object_index, sub):
            return sorted((object_index, f) for f in (self.intrinsic[object_index].parent / sub).iterdir())

        self.rgb_all_filenames = [v for k in range(self.num_objects) for v in views(k, "rgb")]
        self.pose_all_filenames = [v for k in range(self.num_objects) for v in views(k, "pose")]
        assert len(self.rgb_all_filenames) == len(self.pose_all_filenames)
        self.num_views = len(self.rgb_all_filenames) // self.num_objects
        print(f"Total number of objects in the set: {self.num_objects}")
        print(f"Total number of views per object: {self.num_views}")

        self._use_cache = os.environ.get("CODENERF_DATASET_CACHE", "1") != "0"
        self._views = {}            # index -> (color_u8 crop, mask_u8 crop (0 / 255), pose float32, intrinsic float32)
        self._intrinsics = {}       # object index -> (focal, cx, cy, height, width)

    def __len__(self):
        return len(self.rgb_all_filenames)

    # ---- decoding (once per view) -------------------------------------------------------------------------------------
    def _object_intrinsics(self, object_index):
        got = self._intrinsics.get(object_index)
        if got is None:
            with Path(self.intrinsic[object_index]).open() as fh:                      # dataset.py:66-69
                lines = fh.readlines()
                focal, cx, cy, _ = map(float, lines[0].split())
                height, width = map(int, lines[-1].split())
            got = (focal, cx, cy, height, width)
            if self._use_cache:
                self._intrinsics[object_index] = got
        return got

    def decoded(self, index):
        """-> (color uint8 (h, w, C), mask uint8 (h, w, 1) of 0 / 255, pose float32 (4, 4), intrinsic float32 (4, 4)):
        the integer images behind the sample of dataset.py:70-92 (the float images are these / 255.0)."""
        got = self._views.get(index)
        if got is not None:
            return got
        object_index, rgb_filename = self.rgb_all_filenames[index]
        _, pose_filename = self.pose_all_filenames[index]
        focal, cx, cy, height, width = self._object_intrinsics(object_index)

        rgb = _imread(rgb_filename)                                                     # :71
        mask = (rgb != 255).all(axis=-1)[..., None].astype(np.uint8) * 255              # :72
        crop_height, crop_width = height // 8, width // 8                               # :76 (names as in the reference)
        rgb = np.ascontiguousarray(rgb[crop_width:width - crop_width, crop_height:height - crop_height, ...])      # :77
        mask = np.ascontiguousarray(mask[crop_width:width - crop_width, crop_height:height - crop_height, ...])   # :78

        pose = np.loadtxt(pose_filename).reshape(4, 4)                                  # :80
        pose = (pose @ np.diag([1, -1, -1, 1])).astype(np.float32)                      # :81, :91

        intrinsic = np.eye(4)                                                           # :83-85
        intrinsic[0, 0], intrinsic[1, 1] = focal, focal
        intrinsic[0, 2], intrinsic[1, 2] = cx - crop_width, cy - crop_height
        got = (rgb, mask, pose, intrinsic.astype(np.float32))
        if self._use_cache:
            self._views[index] = got
        return got

    def __getitem__(self, index):
        rgb, mask, pose, intrinsic = self.decoded(index)
        return {
            "object_id": self.rgb_all_filenames[index][0],
            "intrinsic": intrinsic.copy(),
            "color": _UNIT[rgb],                 # == (rgb / 255.0).astype(np.float32)
            "mask": _UNIT[mask],
            "pose": pose.copy(),
        }
