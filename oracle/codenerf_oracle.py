"""CPU oracle for the code-nerf volumetric-rendering hot path.

THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it.  The product path (``code-nerf_b200/``) never imports anything from
``oracle/`` and fails loudly when its CUDA library is missing.

It is an independent restatement, in plain fp32 torch-on-CPU tensor arithmetic (the
reference's own substrate: its arithmetic lives in PyTorch ATen, pinned as
``pytorch=1.8.2`` in ``environment.yml:9`` but effectively the installed torch, see
SURVEY.md section 8c), of the algorithm in the reference files cited per function.  Integer /
index pieces additionally have a numpy restatement (``*_np``) so they can be checked
bit-exactly without torch.

Parity pinning: the reference ships no tests, golden vectors or fixtures for this path
 -> "parity unpinned" by reference-owned tests.  The oracle is instead
pinned against outputs of the *live reference code* imported from ``/root/reference`` in
the build container: ``oracle/make_golden.py`` generates ``tests/golden/*.npz`` and
``tests/test_oracle_golden.py`` checks every function below against them.

All ``file:line`` citations are relative to the reference tree.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence, Tuple

import numpy as np
import torch

Tensor = torch.Tensor


# ----------------------------------------------------------------------------------------
# a1/a2/a3  ray generation                      view_synthesis/nerf/ray_sampler.py:11-99
# ----------------------------------------------------------------------------------------
def camera_directions(height: int, width: int, intrinsics: Tensor) -> Tensor:
    """Camera-frame direction per pixel, ``directions[h, w] = ((w-cx)/f, -(h-cy)/f, -1)``.

    ray_sampler.py:31-51 -- only ``K[0,0]`` is used as focal length, ``meshgrid(indexing='xy')``.
    """
    f, cx, cy = intrinsics[0, 0], intrinsics[0, 2], intrinsics[1, 2]
    w = torch.arange(width, dtype=intrinsics.dtype)
    h = torch.arange(height, dtype=intrinsics.dtype)
    ii = w[None, :].expand(height, width)
    jj = h[:, None].expand(height, width)
    return torch.stack([(ii - cx) / f, -(jj - cy) / f, -torch.ones_like(ii)], dim=-1)


def get_bundle(directions: Tensor, pose: Tensor) -> Tuple[Tensor, Tensor]:
    """``rd[b,h,w,j] = sum_i directions[h,w,i] * R_b[j,i]``; ``ro = t_b`` broadcast.

    ray_sampler.py:95-98.  rd is NOT normalised.
    """
    rot = pose[..., :3, :3]
    rd = (directions[None, :, :, None, :] * rot[:, None, None, :, :]).sum(-1)
    ro = pose[..., :3, -1][:, None, None, :].expand(rd.shape)
    return ro, rd.contiguous()


def sample_rays(directions: Tensor, pose: Tensor, select_inds: np.ndarray) -> Tuple[Tensor, Tensor]:
    """Gather the rays at flat pixel index ``h*W+w`` per image and concatenate over images.

    ray_sampler.py:66-80.  ``select_inds`` is the (B, n) int array the reference draws with
    ``np.random.permutation(H*W)[:n]`` per image (:71-75); it is an input here.
    """
    ro, rd = get_bundle(directions, pose)
    ro, rd = ro.flatten(1, 2), rd.flatten(1, 2)
    idx = torch.as_tensor(np.asarray(select_inds), dtype=torch.long)
    ro = torch.cat([ro[i, idx[i], :] for i in range(pose.shape[0])], dim=0)
    rd = torch.cat([rd[i, idx[i], :] for i in range(pose.shape[0])], dim=0)
    return ro, rd


def draw_select_inds(num_pixels: int, batch: int, n: int) -> np.ndarray:
    """The host RNG stream of ray_sampler.py:71-75 (global numpy RNG, one permutation per image)."""
    pixel_range = np.arange(0, num_pixels)
    return np.asarray([np.random.permutation(pixel_range)[:n] for _ in range(batch)])


def pixel_of_flat_index_np(flat: np.ndarray, width: int) -> Tuple[np.ndarray, np.ndarray]:
    """flat = h*W + w  (``flatten(1, 2)`` of a (B,H,W,3) tensor, ray_sampler.py:69)."""
    flat = np.asarray(flat, dtype=np.int64)
    return flat // width, flat % width


# ----------------------------------------------------------------------------------------
# a4/a5/a6  point sampling                    view_synthesis/nerf/point_sampler.py:11-120
# ----------------------------------------------------------------------------------------
def depth_tables(num_coarse: int, near: float, far: float, spacing_mode: str,
                 dtype=torch.float32) -> Tuple[Tensor, Tensor, Tensor]:
    """z table, lower and upper bin edges.  point_sampler.py:33-47.

    NB the mode names are swapped in the reference: ``"lindisp"`` is linear in
    depth, anything else is linear in disparity.
    """
    t = torch.linspace(0.0, 1.0, num_coarse, dtype=dtype)
    if spacing_mode == "lindisp":
        z = near * (1.0 - t) + far * t
    else:
        z = 1.0 / (1.0 / near * (1.0 - t) + 1.0 / far * t)
    mids = 0.5 * (z[1:] + z[:-1])
    upper = torch.cat((mids, z[-1:]))
    lower = torch.cat((z[:1], mids))
    return z, lower, upper


def sample_uniform(ro: Tensor, rd: Tensor, z_table: Tensor, lower: Tensor, upper: Tensor,
                   t_rand: Optional[Tensor]) -> Tuple[Tensor, Tensor]:
    """Stratified depths and points.  point_sampler.py:60-71.

    ``t_rand`` (N, S_c) replaces the ``torch.rand_like`` draw of :64; ``None`` = perturb off.
    """
    n = ro.shape[-2]
    if t_rand is not None:
        z = lower.expand(n, -1) + (upper - lower).expand(n, -1) * t_rand
    else:
        z = z_table.expand(n, -1)
    pts = ro[..., None, :] + rd[..., None, :] * z[..., :, None]
    return pts, z


def searchsorted_right_np(cdf: np.ndarray, u: np.ndarray) -> np.ndarray:
    """``torch.searchsorted(cdf, u, right=True)`` == number of cdf entries <= u, row-wise.

    point_sampler.py:101.  Pure-numpy loop restatement for the bit-exact index test.
    """
    cdf = np.asarray(cdf, dtype=np.float32)
    u = np.asarray(u, dtype=np.float32)
    out = np.zeros(u.shape, dtype=np.int64)
    for r in range(cdf.shape[0]):
        for k in range(u.shape[1]):
            out[r, k] = int(np.count_nonzero(cdf[r] <= u[r, k]))
    return out


def pdf_to_cdf(weights: Tensor) -> Tensor:
    """point_sampler.py:86-89: ``w+1e-5`` -> pdf -> cumsum -> prepend 0."""
    w = weights + 1e-5
    pdf = w / torch.sum(w, dim=-1, keepdim=True)
    cdf = torch.cumsum(pdf, dim=-1)
    return torch.cat([torch.zeros_like(cdf[..., :1]), cdf], dim=-1)


def invert_cdf(cdf: Tensor, bins: Tensor, u: Tensor) -> Tuple[Tensor, Tensor]:
    """point_sampler.py:98-113: searchsorted(right) -> clamp -> gather -> lerp.  Returns (samples, inds)."""
    u = u.contiguous()
    cdf = cdf.contiguous()
    inds = torch.searchsorted(cdf, u, right=True)
    below = torch.clamp(inds - 1, min=0)
    above = torch.clamp(inds, max=cdf.shape[-1] - 1)
    cdf_b, cdf_a = torch.gather(cdf, 1, below), torch.gather(cdf, 1, above)
    bin_b, bin_a = torch.gather(bins, 1, below), torch.gather(bins, 1, above)
    denom = cdf_a - cdf_b
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    t = (u - cdf_b) / denom
    return bin_b + t * (bin_a - bin_b), inds


def sample_pdf(ro: Tensor, rd: Tensor, weights: Tensor, z_vals: Tensor, num_fine: int,
               u: Optional[Tensor]) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """Hierarchical sampling.  point_sampler.py:84-120.

    ``weights`` is (N, S_c-2) (callers pass ``weights[..., 1:-1]``, nerf/__init__.py:87).
    ``u`` (N, N_f) replaces the ``torch.rand`` draw of :93; ``None`` = ``linspace(0,1,N_f)`` (:95-96).
    Samples are detached (:115).  Returns (pts, z_sorted, cdf, inds).
    """
    assert z_vals.shape[-1] - 2 == weights.shape[-1]
    bins = 0.5 * (z_vals[..., 1:] + z_vals[..., :-1])
    cdf = pdf_to_cdf(weights)
    if u is None:
        u = torch.linspace(0.0, 1.0, steps=num_fine, dtype=weights.dtype, device=weights.device).expand(cdf.shape[0], num_fine)
    samples, inds = invert_cdf(cdf, bins, u)
    z_all, _ = torch.sort(torch.cat((z_vals, samples.detach()), dim=-1), dim=-1)
    pts = ro[..., None, :] + rd[..., None, :] * z_all[..., :, None]
    return pts, z_all, cdf, inds


# ----------------------------------------------------------------------------------------
# a7  positional encoding                    view_synthesis/nerf/position_embed.py:9-53
# ----------------------------------------------------------------------------------------
def frequency_bands(num_freq: int, log_sampling: bool, dtype=torch.float32) -> Tensor:
    """position_embed.py:18-33."""
    if log_sampling:
        return 2.0 ** torch.linspace(0.0, num_freq - 1, num_freq, dtype=dtype)
    return torch.linspace(2.0 ** 0.0, 2.0 ** (num_freq - 1), num_freq, dtype=dtype)


def pos_embed(x: Tensor, freqs: Tensor, include_input: bool) -> Tensor:
    """``cat([x, sin(f0 x), cos(f0 x), sin(f1 x), ...])``, each block as wide as x.  position_embed.py:44-53."""
    enc = [x] if include_input else []
    for f in freqs:
        enc.append(torch.sin(x * f))
        enc.append(torch.cos(x * f))
    return enc[0] if len(enc) == 1 else torch.cat(enc, dim=-1)


# ----------------------------------------------------------------------------------------
# a9/a10  networks                              view_synthesis/models/model.py:87-194
# ----------------------------------------------------------------------------------------
CODENERF_LAYERS = ("layer_xyz1", "layer_xyz2", "fc_out", "shape_code_layer1", "shape_code_layer2",
                   "texture_code_layer1", "layer_dir1", "layer_dir2", "fc_rgb")


def codenerf_shapes(hidden: int, code: int, dim_xyz: int, dim_dir: int) -> Dict[str, Tuple[int, int]]:
    """(out, in) of every Linear.  models/model.py:145-156 (texture layer is sized by the *shape* code size)."""
    return {
        "layer_xyz1": (hidden, dim_xyz),
        "layer_xyz2": (hidden, hidden + code),
        "fc_out": (code + 1, hidden + code),
        "shape_code_layer1": (code, code),
        "shape_code_layer2": (code, code),
        "texture_code_layer1": (code, code),
        "layer_dir1": (hidden, dim_dir + code),
        "layer_dir2": (hidden, hidden),
        "fc_rgb": (3, hidden + code),
    }


def make_codenerf_params(seed: int, hidden: int = 256, code: int = 256, dim_xyz: int = 63,
                         dim_dir: int = 27, scale: float = 1.0) -> Dict[str, Tensor]:
    """Deterministic parameters from numpy's PCG64 stream (stable across machines), U(-1/sqrt(in), 1/sqrt(in))
    like torch.nn.Linear's default init range.  Keys follow the reference state_dict: ``<layer>.weight/.bias``."""
    rng = np.random.default_rng(seed)
    out: Dict[str, Tensor] = {}
    for name, (o, i) in codenerf_shapes(hidden, code, dim_xyz, dim_dir).items():
        bound = scale / math.sqrt(i)
        out[name + ".weight"] = torch.from_numpy(rng.uniform(-bound, bound, size=(o, i)).astype(np.float32))
        out[name + ".bias"] = torch.from_numpy(rng.uniform(-bound, bound, size=(o,)).astype(np.float32))
    return out


def _lin(p: Dict[str, Tensor], name: str, x: Tensor) -> Tensor:
    return x @ p[name + ".weight"].t() + p[name + ".bias"]


def codenerf_mlp(p: Dict[str, Tensor], z_s: Tensor, z_t: Tensor, x: Tensor, dim_xyz: int) -> Tensor:
    """CodeNeRFModel.forward as written (per-sample codes).  models/model.py:171-194.  -> (M, 4) = [rgb_raw | sigma_raw]."""
    xyz, view = x[..., :dim_xyz], x[..., dim_xyz:]
    a_s1 = torch.relu(_lin(p, "shape_code_layer1", z_s))
    a_s2 = torch.relu(_lin(p, "shape_code_layer2", z_s))
    a_t = torch.relu(_lin(p, "texture_code_layer1", z_t))
    h = torch.relu(_lin(p, "layer_xyz1", xyz))
    h = torch.relu(_lin(p, "layer_xyz2", torch.cat((h, a_s1), dim=-1)))
    o = _lin(p, "fc_out", torch.cat((h, a_s2), dim=-1))
    sigma, feat = o[..., :1], o[..., 1:]
    v = torch.relu(_lin(p, "layer_dir1", torch.cat((feat, view), dim=-1)))
    v = torch.relu(_lin(p, "layer_dir2", v))
    rgb = _lin(p, "fc_rgb", torch.cat((v, a_t), dim=-1))
    return torch.cat((rgb, sigma), dim=-1)


def embedding_lookup(table_s: Tensor, table_t: Tensor, ids: Tensor) -> Tuple[Tensor, Tensor]:
    """ShapeTextureEmbedding.forward.  models/model.py:102-105."""
    return table_s[ids], table_t[ids]


# ----------------------------------------------------------------------------------------
# a8  forward_pass                                  view_synthesis/nerf/__init__.py:94-134
# ----------------------------------------------------------------------------------------
def paired_ray_index_np(num_rays: int, num_samples: int) -> np.ndarray:
    """Which ray's view direction flat sample row k = n*S+s receives: ``k mod N``.

    nerf/__init__.py:127 tiles an (N,3) tensor with ``repeat([1, S, 1])`` -> (1, N*S, 3), so row k gets the
    direction of ray ``k mod N`` while the point of row k belongs to ray ``k // S`` (:120).  SURVEY F6: the
    drop-in reproduces this pairing by default.
    """
    return (np.arange(num_rays * num_samples, dtype=np.int64) % num_rays).reshape(num_rays, num_samples)


def forward_pass(p: Dict[str, Tensor], rd: Tensor, pts: Tensor, z_s: Tensor, z_t: Tensor,
                 freqs_xyz: Tensor, freqs_dir: Tensor, include_input_xyz: bool = True,
                 include_input_dir: bool = True, pairing: str = "reference") -> Tensor:
    """Embed points and (bug-compatibly paired) view directions, run the MLP.  nerf/__init__.py:116-134."""
    n, s = pts.shape[0], pts.shape[1]
    zs = z_s[:, None, :].expand(-1, s, -1).reshape(-1, z_s.shape[-1])
    zt = z_t[:, None, :].expand(-1, s, -1).reshape(-1, z_t.shape[-1])
    emb = pos_embed(pts.reshape(-1, 3), freqs_xyz, include_input_xyz)
    dim_xyz = emb.shape[-1]
    viewdirs = rd / rd.norm(p=2, dim=-1).unsqueeze(-1)
    if pairing == "reference":
        dirs_flat = viewdirs.repeat([1, s, 1]).reshape(-1, 3)      # row k -> ray k mod N
    else:
        dirs_flat = viewdirs[:, None, :].expand(-1, s, -1).reshape(-1, 3)  # row k -> ray k // S
    emb = torch.cat((emb, pos_embed(dirs_flat, freqs_dir, include_input_dir)), dim=-1)
    return codenerf_mlp(p, zs, zt, emb, dim_xyz).reshape(n, s, 4)


# ----------------------------------------------------------------------------------------
# a11  compositing                          view_synthesis/nerf/volumetric_render.py:28-66
# ----------------------------------------------------------------------------------------
def volume_render(raw: Tensor, z: Tensor, rd: Tensor) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """-> (rgb, disp, acc, weights, depth), the reference's tuple order.  volumetric_render.py:41-66."""
    dists = torch.cat((z[..., 1:] - z[..., :-1], torch.full_like(z[..., :1], 1e10)), dim=-1)
    delta = dists * rd[..., None, :].norm(p=2, dim=-1)
    sigma = torch.nn.functional.softplus(raw[..., 3] - 1)                 # shifted_softplus :32-33
    sd = sigma * delta
    rgb = torch.sigmoid(raw[..., :3]) * (1 + 2 * 0.001) - 0.001           # widened_sigmoid :28-29
    trans = torch.exp(-torch.cat([torch.zeros_like(sd[..., :1]), torch.cumsum(sd[..., :-1], dim=-1)], dim=-1))
    alpha = 1.0 - torch.exp(-sd)
    w = alpha * trans
    rgb_map = (w[..., None] * rgb).sum(dim=-2)
    depth = (w * z).sum(dim=-1)
    acc = w.sum(dim=-1)
    disp = 1.0 / torch.max(1e-10 * torch.ones_like(depth), depth / acc)
    return rgb_map, disp, acc, w, depth


# ----------------------------------------------------------------------------------------
# a12  predict_radiance_and_render                  view_synthesis/nerf/__init__.py:74-91
# ----------------------------------------------------------------------------------------
class RenderConfig:
    """The numbers of ``config/*.yml`` that the path reads (nerf/__init__.py:28-37,57-69)."""

    def __init__(self, num_coarse=32, num_fine=128, near=0.8, far=1.8, spacing_mode="lindepth",
                 num_freq_xyz=10, num_freq_dir=4, log_sampling=True, include_input=True):
        self.num_coarse, self.num_fine, self.near, self.far = num_coarse, num_fine, near, far
        self.spacing_mode = spacing_mode
        self.freqs_xyz = frequency_bands(num_freq_xyz, log_sampling)
        self.freqs_dir = frequency_bands(num_freq_dir, log_sampling)
        self.include_input = include_input
        self.z, self.lower, self.upper = depth_tables(num_coarse, near, far, spacing_mode)


def predict_radiance_and_render(cfg: RenderConfig, ro: Tensor, rd: Tensor, p_coarse: Dict[str, Tensor],
                                p_fine: Dict[str, Tensor], z_s: Tensor, z_t: Tensor,
                                t_rand: Optional[Tensor], u: Optional[Tensor],
                                pairing: str = "reference") -> Tuple[Tensor, Tensor, Dict[str, Tensor]]:
    """coarse sample -> MLP -> composite -> importance sample -> MLP -> composite.  nerf/__init__.py:81-91."""
    pts_c, z_c = sample_uniform(ro, rd, cfg.z, cfg.lower, cfg.upper, t_rand)
    raw_c = forward_pass(p_coarse, rd, pts_c, z_s, z_t, cfg.freqs_xyz, cfg.freqs_dir,
                         cfg.include_input, cfg.include_input, pairing)
    rgb_c, _, _, w_c, _ = volume_render(raw_c, z_c, rd)
    pts_f, z_f, cdf, inds = sample_pdf(ro, rd, w_c[..., 1:-1], z_c, cfg.num_fine, u)
    raw_f = forward_pass(p_fine, rd, pts_f, z_s, z_t, cfg.freqs_xyz, cfg.freqs_dir,
                         cfg.include_input, cfg.include_input, pairing)
    rgb_f, disp_f, acc_f, w_f, depth_f = volume_render(raw_f, z_f, rd)
    aux = {"z_coarse": z_c, "weights_coarse": w_c, "z_fine": z_f, "raw_coarse": raw_c, "raw_fine": raw_f,
           "cdf": cdf, "inds": inds, "depth_fine": depth_f, "acc_fine": acc_f, "disp_fine": disp_f}
    return rgb_c, rgb_f, aux


# ----------------------------------------------------------------------------------------
# a13  ray split of parallel_image_render           view_synthesis/nerf/__init__.py:179-187
# ----------------------------------------------------------------------------------------
def render_split_np(num_rays: int, n_gpus: int) -> Tuple[np.ndarray, np.ndarray]:
    """Rays per rank and zero-padding per rank: floor split, remainder on the last rank, others padded."""
    per = np.full([n_gpus], int(num_rays / n_gpus), dtype=np.int64)
    padding = num_rays - int(per.sum())
    per[-1] = num_rays - int(per[:-1].sum())
    pad = np.zeros([n_gpus], dtype=np.int64)
    if padding > 0:
        pad[:-1] = padding
    return per, pad


# ----------------------------------------------------------------------------------------
# a15  FlexibleNeRFModel ("next" row f1)              view_synthesis/models/model.py:5-76
# ----------------------------------------------------------------------------------------
def flexible_nerf_mlp(p: Dict[str, Tensor], x: Tensor, dim_xyz: int, num_layers: int,
                      skip_connect_ids: Sequence[int]) -> Tensor:
    """models/model.py:58-76; the use_viewdirs=False network is recognised by its fc_out head (:47-48,75-76)."""
    xyz, view = x[..., :dim_xyz], x[..., dim_xyz:]
    out = torch.relu(_lin(p, "layer1", xyz))
    for i in range(num_layers - 1):
        if i in skip_connect_ids:
            out = torch.cat((out, xyz), dim=-1)
        out = torch.relu(_lin(p, f"layers_xyz.{i}", out))
    if "fc_out.weight" in p:
        return _lin(p, "fc_out", out)
    feat = torch.relu(_lin(p, "fc_feat", out))
    sigma = _lin(p, "fc_alpha", feat)
    out = torch.relu(_lin(p, "layers_dir.0", torch.cat((feat, view), dim=-1)))
    rgb = _lin(p, "fc_rgb", out)
    return torch.cat((rgb, sigma), dim=-1)


def flexible_nerf_shapes(num_layers: int, hidden: int, skip_connect_ids: Sequence[int], dim_xyz: int,
                         dim_dir: int, use_viewdirs: bool = True) -> Dict[str, Tuple[int, int]]:
    """(out, in) of every Linear of FlexibleNeRFModel, in state_dict order.  models/model.py:31-48."""
    shapes = {"layer1": (hidden, dim_xyz)}
    for i in range(num_layers - 1):
        shapes[f"layers_xyz.{i}"] = (hidden, hidden + dim_xyz) if i in skip_connect_ids else (hidden, hidden)
    if not use_viewdirs:
        shapes["fc_out"] = (4, hidden)
        return shapes
    shapes["fc_feat"] = (hidden, hidden)
    shapes["layers_dir.0"] = (hidden // 2, dim_dir + hidden)
    shapes["fc_alpha"] = (1, hidden)
    shapes["fc_rgb"] = (3, hidden // 2)
    return shapes


def make_flexible_nerf_params(seed: int, num_layers: int = 8, hidden: int = 256, skip_connect_ids: Sequence[int] = (4,),
                              dim_xyz: int = 63, dim_dir: int = 27, use_viewdirs: bool = True) -> Dict[str, Tensor]:
    """Deterministic FlexibleNeRFModel parameters (numpy PCG64, torch.nn.Linear's default init range)."""
    rng = np.random.default_rng(seed)
    out: Dict[str, Tensor] = {}
    for name, (o, i) in flexible_nerf_shapes(num_layers, hidden, skip_connect_ids, dim_xyz, dim_dir, use_viewdirs).items():
        bound = 1.0 / math.sqrt(i)
        out[name + ".weight"] = torch.from_numpy(rng.uniform(-bound, bound, size=(o, i)).astype(np.float32))
        out[name + ".bias"] = torch.from_numpy(rng.uniform(-bound, bound, size=(o,)).astype(np.float32))
    return out


def flexible_forward_pass(p: Dict[str, Tensor], rd: Tensor, pts: Tensor, freqs_xyz: Tensor, freqs_dir: Tensor,
                          num_layers: int, skip_connect_ids: Sequence[int], pairing: str = "reference") -> Tensor:
    """forward_pass (nerf/__init__.py:116-134) around a FlexibleNeRFModel: same embedding and the same view-direction
    tiling (F6); the latent codes are expanded and then ignored by the 3-line adapter SURVEY F11 describes."""
    n, s = pts.shape[0], pts.shape[1]
    emb = pos_embed(pts.reshape(-1, 3), freqs_xyz, True)
    dim_xyz = emb.shape[-1]
    viewdirs = rd / rd.norm(p=2, dim=-1).unsqueeze(-1)
    if pairing == "reference":
        dirs_flat = viewdirs.repeat([1, s, 1]).reshape(-1, 3)
    else:
        dirs_flat = viewdirs[:, None, :].expand(-1, s, -1).reshape(-1, 3)
    emb = torch.cat((emb, pos_embed(dirs_flat, freqs_dir, True)), dim=-1)
    return flexible_nerf_mlp(p, emb, dim_xyz, num_layers, skip_connect_ids).reshape(n, s, 4)


def mse2psnr(mse_val: float) -> float:
    """utils/util.py:216-227."""
    if mse_val == 0:
        mse_val = 1e-5
    return -10.0 * math.log10(mse_val)
