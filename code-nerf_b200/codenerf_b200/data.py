"""Device-resident dataset stage: every decoded view of an `SRNDataset` lives in HBM as the
uint8 crop its PNG holds, next to its pose, object id and intrinsics; a training iteration then moves B view indices to
the device instead of B float images.

    data = DeviceSRN(dataset, device)                  # dataset: view_synthesis.datasets.SRNDataset (drop-in or reference)
    batch = data.batch(indices)                        # == the reference's collated DataLoader batch, already on the device
    tgt = data.targets(indices, select_inds)           # == train.py:77-80 without materialising a float image

`batch()["color"]` and `targets()` are bit-identical to what the reference computes (csrc/data.cu).
Reference: view_synthesis/datasets/dataset.py:60-94, train.py:70-80, utils/util.py:72-88.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib


def _p(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class DeviceSRN:
    def __init__(self, dataset, device, indices=None):
        """Decode (once) and upload the views `indices` (default: all) of `dataset`."""
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("DeviceSRN keeps the views in GPU memory: it needs a CUDA device (there is no CPU fallback)")
        idx = list(range(len(dataset))) if indices is None else list(indices)
        if hasattr(dataset, "decoded"):
            dec = [dataset.decoded(i) for i in idx]
            colors = np.stack([d[0] for d in dec])
            masks = np.stack([d[1] for d in dec])
            poses = np.stack([d[2] for d in dec])
            intr = np.stack([d[3] for d in dec])
        else:       # the reference's own class: recover the integer images from its float samples (exact: v = round(255 f))
            smp = [dataset[i] for i in idx]
            colors = np.stack([np.rint(s["color"].astype(np.float64) * 255.0).astype(np.uint8) for s in smp])
            masks = np.stack([np.rint(s["mask"].astype(np.float64) * 255.0).astype(np.uint8) for s in smp])
            poses = np.stack([s["pose"] for s in smp])
            intr = np.stack([s["intrinsic"] for s in smp])
        oid = np.asarray([dataset.rgb_all_filenames[i][0] for i in idx], dtype=np.int64)
        self.n_views, self.height, self.width, self.channels = colors.shape
        self.colors_u8 = torch.from_numpy(colors).to(self.device)          # [V, h, w, C] uint8
        self.masks_u8 = torch.from_numpy(masks).to(self.device)            # [V, h, w, 1] uint8 (0 / 255)
        self.poses = torch.from_numpy(poses).to(self.device)               # [V, 4, 4] float32
        self.intrinsics = torch.from_numpy(intr).to(self.device)           # [V, 4, 4] float32
        self.object_ids = torch.from_numpy(oid).to(self.device)            # [V] int64
        self._bad = torch.zeros(1, dtype=torch.int32, device=self.device)

    def __len__(self):
        return self.n_views

    def _index(self, indices):
        if not torch.is_tensor(indices):
            indices = torch.as_tensor(np.asarray(indices, dtype=np.int64))
        return indices.to(self.device, torch.int64, non_blocking=True).contiguous()

    def _unit(self, u8):
        out = torch.empty(u8.shape, dtype=torch.float32, device=self.device)
        _lib.check(_lib.load().cnrf_u8_to_unit(_p(u8), u8.numel(), _p(out), _stream()), "cnrf_u8_to_unit")
        return out

    def batch(self, indices):
        """The dict `next(iter(DataLoader(dataset, batch_size=B)))` yields for these views (default collate), on the device."""
        ix = self._index(indices)
        return {"object_id": self.object_ids[ix], "intrinsic": self.intrinsics[ix], "color": self._unit(self.colors_u8[ix].contiguous()),
                "mask": self._unit(self.masks_u8[ix].contiguous()), "pose": self.poses[ix]}

    def targets(self, indices, select_inds, channels=None):
        """train.py:77-80: cat_k color[k].flatten(0, 1)[select_inds[k]] -> [B * R, channels] float32 (default: all channels
        of the PNG; train.py:103 then reads [..., :3])."""
        ix = self._index(indices)
        sel = select_inds if torch.is_tensor(select_inds) else torch.as_tensor(np.asarray(select_inds, dtype=np.int64))
        sel = sel.to(self.device, torch.int64, non_blocking=True).contiguous()
        B, R = sel.shape
        if B != ix.numel():
            raise ValueError(f"select_inds has {B} rows for {ix.numel()} views")
        co = self.channels if channels is None else int(channels)
        out = torch.empty(B * R, co, dtype=torch.float32, device=self.device)
        _lib.check(_lib.load().cnrf_gather_pixels_u8(_p(self.colors_u8), self.n_views, self.height * self.width, self.channels,
                                                     _p(ix), _p(sel), B, R, co, _p(out), _p(self._bad), _stream()),
                   "cnrf_gather_pixels_u8")
        return out

    def check(self):
        """Raise if any `targets` call since the last check saw a pixel index outside the image (one device read)."""
        if int(self._bad.item()):
            self._bad.zero_()
            raise IndexError("select_inds out of range in DeviceSRN.targets")
