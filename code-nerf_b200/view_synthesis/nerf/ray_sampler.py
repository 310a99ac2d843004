"""RaySampler -- same constructor, attributes and methods as the reference's view_synthesis/nerf/ray_sampler.py:7-99."""
from typing import Tuple, Union

import numpy as np
import torch

from codenerf_b200 import ops


class RaySampler(object):
    """Samples rays for a given image size and intrinsics (reference: ray_sampler.py:11-51)."""

    def __init__(self, height: int, width: int, intrinsics: Union[torch.Tensor, np.ndarray], sample_size: int,
                 device, datatype):
        assert height > 0 and width > 0, "Height and width must be positive integers"
        assert sample_size > 0 and sample_size <= height * width, "Sample size must be a positive number less than height * width"
        self.height = height
        self.width = width
        self.sample_size = sample_size
        self.device = device
        if isinstance(intrinsics, np.ndarray):
            intrinsics = torch.from_numpy(intrinsics).to(datatype)
        assert intrinsics.shape == torch.Size([4, 4]), "Incorrect intrinsics shape"
        host = intrinsics.detach().to("cpu", torch.float32)
        # scalars the kernels take by value (one host read at construction, never on the hot path)
        self._f, self._cx, self._cy = float(host[0, 0]), float(host[0, 2]), float(host[1, 2])
        self.intrinsics = intrinsics.to(device)
        self.focal_length = self.intrinsics[..., 0, 0]
        self.cx = self.intrinsics[..., 0, 2]
        self.cy = self.intrinsics[..., 1, 2]
        self._directions = None
        self._next = None                # (select_inds, device copy) drawn ahead of time by prefetch()
        self._pins = [None, None]        # two pinned staging buffers + the events of their last H2D copies
        self._pin_events = [None, None]
        self._pin_turn = 0
        self.select_inds_device = None   # int64 (B, n) device copy of the last sample()'s select_inds

    @property
    def directions(self) -> torch.Tensor:
        """(H, W, 3) camera-frame directions (ray_sampler.py:44-51).  The kernels recompute them per ray; this table
        is only materialised if somebody asks for the attribute (identity pose through the same kernel)."""
        if self._directions is None:
            eye = torch.eye(4, dtype=torch.float32, device=self.device)[None]
            _, rd = ops.ray_bundle(eye, self.height, self.width, self._f, self._cx, self._cy, want_ro=False)
            self._directions = rd.reshape(self.height, self.width, 3)
        return self._directions

    def sample(self, tform_cam2world: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, np.ndarray]:
        """Random subset of rays per image (ray_sampler.py:53-82) -> ro, rd (B*n, 3), select_inds np.ndarray (B, n).

        The pixel permutation stays on the host numpy RNG, call for call, so the stream is the reference's (:71-75);
        gather and rotation are one kernel writing only the selected rays."""
        batch_size = tform_cam2world.shape[0]
        if self._next is not None and self._next[0].shape[0] == batch_size and self._next[1].device == tform_cam2world.device:
            select_inds, sel = self._next
        else:
            select_inds = self._draw(batch_size)
            sel = self._to_device(select_inds, tform_cam2world.device)
        self._next = None
        self.select_inds_device = sel
        ro, rd = ops.ray_bundle(tform_cam2world, self.height, self.width, self._f, self._cx, self._cy, select=sel,
                                n=self.sample_size, want_ro=True)
        return ro, rd, select_inds

    def _draw(self, batch_size: int) -> np.ndarray:
        """One np.random.permutation per image from the global numpy RNG, exactly as ray_sampler.py:71-75."""
        pixel_range = np.arange(0, self.height * self.width)
        select_inds = []
        for _ in range(batch_size):
            select_inds.append(np.random.permutation(pixel_range)[:self.sample_size])
        return np.ascontiguousarray(np.asarray(select_inds), dtype=np.int64)

    def _to_device(self, select_inds: np.ndarray, device) -> torch.Tensor:
        """Asynchronous H2D copy through one of two pinned staging buffers (a buffer is rewritten only after the copy that
        last read it has completed)."""
        k = self._pin_turn
        self._pin_turn ^= 1
        if self._pins[k] is None or self._pins[k].shape != select_inds.shape:
            self._pins[k] = torch.empty(select_inds.shape, dtype=torch.int64).pin_memory()
            self._pin_events[k] = None
        if self._pin_events[k] is not None:
            self._pin_events[k].synchronize()
        self._pins[k].copy_(torch.from_numpy(select_inds))
        sel = self._pins[k].to(device, non_blocking=True)
        self._pin_events[k] = torch.cuda.Event()
        self._pin_events[k].record()
        return sel

    def prefetch(self, batch_size: int) -> None:
        """Optional (not in the reference): draw the NEXT sample() call's pixel permutations now -- same global numpy RNG,
        same order, so select_inds are bit-identical -- and start their H2D copy.  Called after a step's kernels have been
        enqueued and before its loss is read, the ~0.25 ms per image of host permutation runs under the GPU's work instead
        of in front of the next step.  The next sample() with this batch size consumes the prefetched draw."""
        if self._next is not None and self._next[0].shape[0] == batch_size:
            return
        select_inds = self._draw(batch_size)
        self._next = (select_inds, self._to_device(select_inds, torch.device(self.device)))

    def get_bundle(self, tform_cam2world: torch.Tensor):
        """All H*W rays (ray_sampler.py:84-99) -> ro (B,H,W,3) broadcast view of the translation, rd (B,H,W,3)."""
        B = tform_cam2world.shape[0]
        _, rd = ops.ray_bundle(tform_cam2world, self.height, self.width, self._f, self._cx, self._cy, want_ro=False)
        ray_directions = rd.reshape(B, self.height, self.width, 3)
        ray_origins = tform_cam2world[..., :3, -1][:, None, None, :].expand(ray_directions.shape)
        return ray_origins, ray_directions
