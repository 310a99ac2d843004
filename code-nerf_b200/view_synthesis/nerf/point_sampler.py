"""PointSampler -- signatures of the reference's view_synthesis/nerf/point_sampler.py:7-120."""
from typing import Tuple

import torch

from codenerf_b200 import ops


class PointSampler(object):
    """Sample 3D points along the given rays."""

    def __init__(self, num_samples_coarse, num_samples_fine, near: float, far: float, spacing_mode, perturb: bool,
                 dtype, device):
        assert near >= 0 and far > near, "Near and far ranges should be positive values, and far > near"
        assert num_samples_coarse > 0 and num_samples_fine > 0, "Number of samples must be greater than 0"
        self.num_samples_coarse = num_samples_coarse
        self.num_samples_fine = num_samples_fine
        self.near = near
        self.far = far
        self.spacing_mode = spacing_mode
        self.perturb = perturb
        self.dtype = dtype
        self.device = device
        # one-off (S_c,) tables, same expressions as point_sampler.py:33-47 (mode names swapped there: SURVEY F7)
        self.t_vals = torch.linspace(0.0, 1.0, self.num_samples_coarse, dtype=self.dtype, device=self.device)
        if self.spacing_mode == "lindisp":
            self.z_vals = self.near * (1.0 - self.t_vals) + self.far * self.t_vals
        else:
            self.z_vals = 1.0 / (1.0 / self.near * (1.0 - self.t_vals) + 1.0 / self.far * self.t_vals)
        self.mids = 0.5 * (self.z_vals[..., 1:] + self.z_vals[..., :-1])
        self.upper = torch.cat((self.mids, self.z_vals[..., -1:]), dim=-1)
        self.lower = torch.cat((self.z_vals[..., :1], self.mids), dim=-1)

    def sample_uniform(self, ro: torch.Tensor, rd: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """Stratified samples (point_sampler.py:49-71) -> pts (N, S_c, 3), z_vals (N, S_c)."""
        num_random_rays = ro.shape[-2]
        t_rand = None
        if self.perturb:
            # the same device RNG draw as point_sampler.py:64, handed to the kernel (parity of the random stream)
            t_rand = torch.rand((num_random_rays, self.num_samples_coarse), dtype=ro.dtype, device=ro.device)
        pts, z_vals = ops.sample_uniform(ro, rd, self.z_vals, self.lower, self.upper, t_rand)
        assert z_vals.shape == torch.Size([num_random_rays, self.num_samples_coarse]), "Incorrect shape of depth samples z_vals"
        return pts, z_vals

    def draw_offsets(self, num_rays: int, dtype, device):
        """The two device RNG draws of one coarse + fine pass, in the reference's order (point_sampler.py:64, then :93);
        (None, None) when perturb is off."""
        if not self.perturb:
            return None, None
        t_rand = torch.rand((num_rays, self.num_samples_coarse), dtype=dtype, device=device)
        u = torch.rand([num_rays, self.num_samples_fine], dtype=dtype, device=device)
        return t_rand, u

    def sample_pdf(self, ro: torch.Tensor, rd: torch.Tensor, weights: torch.Tensor, z_vals: torch.Tensor
                   ) -> Tuple[torch.Tensor, torch.Tensor]:
        """Inverse-CDF samples merged with the coarse ones (point_sampler.py:73-120) -> pts (N, S_c+N_f, 3), z (N, S_c+N_f)."""
        assert self.num_samples_coarse - 2 == weights.shape[-1], f"Weights size {weights.shape} should match {self.num_samples_coarse - 1}"
        u = None
        if self.perturb:
            u = torch.rand([weights.shape[0], self.num_samples_fine], dtype=weights.dtype, device=weights.device)  # :93
        return ops.sample_pdf(ro, rd, weights, z_vals, self.num_samples_fine, u)
