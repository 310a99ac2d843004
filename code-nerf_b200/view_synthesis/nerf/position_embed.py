"""PositionalEmbedder -- signatures of the reference's view_synthesis/nerf/position_embed.py:5-53."""
import torch

from codenerf_b200 import ops


class PositionalEmbedder(object):
    """Positionally encode the input vector through a fourier basis with the given frequency bands."""

    def __init__(self, num_freq: int, log_sampling: bool, include_input: bool, dtype, device) -> None:
        assert num_freq > 0, "Number of frequency samples should be a positive integer"
        self.num_freq = num_freq
        self.log_sampling = log_sampling
        self.include_input = include_input
        self.dtype = dtype
        self.device = device
        if self.log_sampling:   # position_embed.py:18-25
            self.frequency_bands = 2.0 ** torch.linspace(0.0, self.num_freq - 1, self.num_freq, dtype=self.dtype, device=self.device)
        else:                   # position_embed.py:26-33
            self.frequency_bands = torch.linspace(2.0 ** 0.0, 2.0 ** (self.num_freq - 1), self.num_freq, dtype=self.dtype, device=self.device)

    @property
    def out_dim_per_channel(self) -> int:
        return int(self.include_input) + 2 * self.num_freq

    def embed(self, tensor: torch.Tensor) -> torch.Tensor:
        """[..., D] -> [..., D * (include_input + 2 * num_freq)]  (position_embed.py:35-53)."""
        lead = tensor.shape[:-1]
        out = ops.pos_embed(tensor.reshape(-1, tensor.shape[-1]), self.frequency_bands, self.include_input)
        return out.reshape(*lead, out.shape[-1])
