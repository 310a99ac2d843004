"""RaySampler.prefetch (drop-in extension, not in the reference): drawing the next call's pixel permutations ahead of time
must leave select_inds, the global numpy RNG stream and the rays bit-identical to plain sample() calls (ray_sampler.py:53-82)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_prefetch_is_bit_identical_to_plain_sampling():
    import view_synthesis.nerf as nerf
    assert torch.cuda.is_available()
    dev = torch.device("cuda", 0)
    H = W = 64
    K = torch.tensor([[65.6, 0, W / 2, 0], [0, 65.6, H / 2, 0], [0, 0, 1, 0], [0, 0, 0, 1]], dtype=torch.float32)
    g = torch.Generator().manual_seed(3)
    poses = torch.eye(4).repeat(3, 1, 1)
    poses[:, :3, :] += 0.1 * torch.randn(3, 3, 4, generator=g)
    poses = poses.to(dev)
    out = {}
    for mode in ("plain", "prefetch"):
        np.random.seed(77)
        rs = nerf.RaySampler(H, W, K, 500, dev, torch.float32)
        rows = []
        for it in range(4):
            ro, rd, sel = rs.sample(tform_cam2world=poses)
            assert isinstance(sel, np.ndarray) and sel.dtype == np.int64 and sel.shape == (3, 500)
            assert torch.equal(rs.select_inds_device.cpu(), torch.from_numpy(sel))
            rows.append((ro.cpu(), rd.cpu(), sel.copy()))
            if mode == "prefetch" and it != 2:          # (one step without a prefetch in between: falls back to drawing)
                rs.prefetch(3)
                rs.prefetch(3)                          # idempotent until consumed
        out[mode] = (rows, np.random.randint(1 << 30) if mode == "plain" else None)
    # the prefetching run has drawn one permutation set more (the one waiting for the next call)
    for a, b in zip(out["plain"][0], out["prefetch"][0]):
        assert np.array_equal(a[2], b[2])
        assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
