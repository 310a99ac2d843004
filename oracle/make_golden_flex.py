"""Generate tests/golden/flexnerf.npz from the UNMODIFIED reference FlexibleNeRFModel (models/model.py:5-76) on CPU.

Test infrastructure, run in the build container only (`python oracle/make_golden_flex.py`); see oracle/make_golden.py for
the import recipe.  The reference's own drivers cannot reach this class (forward_pass calls model(z_s, z_t, x)),
so the forward_pass fixture goes through the 3-line adapter SURVEY section 8(c) describes.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import codenerf_oracle as orc  # noqa: E402
from oracle.make_golden import import_reference, OUT  # noqa: E402


class Adapter(torch.nn.Module):
    def __init__(self, inner):
        super().__init__()
        self.inner = inner

    def forward(self, z_s, z_t, x):
        return self.inner(x)


def main():
    nerf, models = import_reference()
    rng = np.random.default_rng(4321)
    out = {}
    for tag, (L, H, skips, M) in {"small": (4, 64, (2,), 37), "lego": (8, 256, (4,), 19)}.items():
        p = orc.make_flexible_nerf_params(31, L, H, skips)
        model = models.FlexibleNeRFModel(L, H, list(skips), 10, 4, True, True, True)
        model.load_state_dict({k: v.clone() for k, v in p.items()})
        x = torch.from_numpy(rng.normal(size=(M, 90)).astype(np.float32)).requires_grad_(True)
        y = model(x)
        g = torch.from_numpy(rng.normal(size=(M, 4)).astype(np.float32))
        (y * g).sum().backward()
        out.update({f"{tag}_L": L, f"{tag}_H": H, f"{tag}_skips": np.array(skips), f"{tag}_x": x.detach().numpy(),
                    f"{tag}_y": y.detach().numpy(), f"{tag}_gy": g.numpy(), f"{tag}_gx": x.grad.numpy()})
        for name, prm in model.named_parameters():
            out[f"{tag}_g_{name}"] = prm.grad.numpy().reshape(-1)[::97].copy()
            out[f"{tag}_gsum_{name}"] = np.float64(prm.grad.numpy().astype(np.float64).sum())
        # forward_pass with the reference's view-direction tiling, through the adapter
        N, S = 5, 3
        rd = torch.from_numpy(rng.normal(size=(N, 3)).astype(np.float32))
        pts = torch.from_numpy(rng.normal(size=(N, S, 3)).astype(np.float32))
        pe_x = nerf.PositionalEmbedder(10, True, True, torch.float32, torch.device("cpu"))
        pe_d = nerf.PositionalEmbedder(4, True, True, torch.float32, torch.device("cpu"))
        with torch.no_grad():
            raw = nerf.forward_pass(Adapter(model), (pe_x, pe_d), rd, pts, (torch.zeros(N, 1), torch.zeros(N, 1)))
        out.update({f"{tag}_fp_rd": rd.numpy(), f"{tag}_fp_pts": pts.numpy(), f"{tag}_fp_raw": raw.numpy()})
    # use_viewdirs=False (fc_out head, models/model.py:47-48,75-76): own generator, so the fixtures above keep their values
    rng = np.random.default_rng(8765)
    for tag, (L, H, skips, M) in {"small_nv": (4, 64, (2,), 37), "lego_nv": (8, 256, (4,), 19)}.items():
        p = orc.make_flexible_nerf_params(32, L, H, skips, use_viewdirs=False)
        model = models.FlexibleNeRFModel(L, H, list(skips), 10, 4, True, True, False)
        model.load_state_dict({k: v.clone() for k, v in p.items()})
        x = torch.from_numpy(rng.normal(size=(M, 90)).astype(np.float32)).requires_grad_(True)   # direction columns ignored
        y = model(x)
        g = torch.from_numpy(rng.normal(size=(M, 4)).astype(np.float32))
        (y * g).sum().backward()
        out.update({f"{tag}_L": L, f"{tag}_H": H, f"{tag}_skips": np.array(skips), f"{tag}_x": x.detach().numpy(),
                    f"{tag}_y": y.detach().numpy(), f"{tag}_gy": g.numpy(), f"{tag}_gx": x.grad.numpy()})
        for name, prm in model.named_parameters():
            out[f"{tag}_g_{name}"] = prm.grad.numpy().reshape(-1)[::97].copy()
            out[f"{tag}_gsum_{name}"] = np.float64(prm.grad.numpy().astype(np.float64).sum())
    np.savez(os.path.join(OUT, "flexnerf.npz"), **out)
    print("wrote flexnerf.npz", {k: v.shape for k, v in out.items() if hasattr(v, "shape") and v.ndim > 1})


if __name__ == "__main__":
    main()
