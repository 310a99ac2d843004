"""Render pipeline API -- same names and signatures as the reference's view_synthesis/nerf/__init__.py:15-226."""
import os
from typing import List, Tuple, Union

import numpy as np
import torch
import torch.distributed as dist

from .ray_sampler import RaySampler
from .point_sampler import PointSampler
from .position_embed import PositionalEmbedder
from .volumetric_render import volume_render
from codenerf_b200 import _lib

# ---- knobs the reference does not have -----------------------------------------------------------------------
_PRECISIONS = {"fp32": _lib.PRECISION_FP32, "bf16": _lib.PRECISION_BF16, "split": _lib.PRECISION_SPLIT}
_PAIRINGS = {"reference": _lib.PAIRING_REFERENCE, "correct": _lib.PAIRING_CORRECT}
_state = {"precision": os.environ.get("CODENERF_PRECISION", "bf16"),
          "pairing": os.environ.get("CODENERF_PAIRING", "reference")}


def set_precision(name: str) -> None:
    """"bf16": fused tcgen05 MLP (throughput mode, default).  "fp32": CUDA-core fp32 MLP (1e-3 parity mode).
    "split": the fp32 path with its contractions on the tensor cores as split-bf16 (~16 mantissa bits)."""
    assert name in _PRECISIONS, f"precision must be one of {sorted(_PRECISIONS)}"
    _state["precision"] = name


def set_pairing(name: str) -> None:
    """"reference" (default) reproduces the view-direction tiling of nerf/__init__.py:127;
    "correct" pairs every sample with its own ray's direction."""
    assert name in _PAIRINGS, f"pairing must be one of {sorted(_PAIRINGS)}"
    _state["pairing"] = name


def get_precision() -> str:
    return _state["precision"]


def get_pairing() -> str:
    return _state["pairing"]


class FieldQuery(object):
    """What forward_pass hands to CodeNeRFModel.forward in place of the (N*S, 90) embedded tensor: the un-embedded
    points and ray directions plus the embedders, so the model can run encoding + MLP as fused kernels.  The call
    still goes through `model(...)`, i.e. through DistributedDataParallel.forward when the model is wrapped (H7)."""
    __slots__ = ("pts", "rd", "embedder_xyz", "embedder_dir", "precision", "pairing")

    def __init__(self, pts, rd, embedder_xyz, embedder_dir, precision, pairing):
        self.pts, self.rd, self.embedder_xyz, self.embedder_dir = pts, rd, embedder_xyz, embedder_dir
        self.precision, self.pairing = precision, pairing


class RenderQuery(object):
    """The fused form of predict_radiance_and_render (cnrf_render_fwd / _bwd: one library call per direction).  Passed
    through `coarse_model(z_s, z_t, query)` first (query.coarse is None -> the model returns its parameter token), then
    through `fine_model(z_s, z_t, query)` with the token attached -> (rgb_coarse, rgb_fine).  Both calls go through the
    modules' __call__, hence through DistributedDataParallel.forward when the caller wrapped them."""
    __slots__ = ("ro", "rd", "point_sampler", "embedder_xyz", "embedder_dir", "precision", "pairing", "t_rand", "u", "coarse",
                 "loss_target")

    def __init__(self, ro, rd, point_sampler, embedder_xyz, embedder_dir, precision, pairing, t_rand, u, loss_target=None):
        self.ro, self.rd, self.point_sampler = ro, rd, point_sampler
        self.embedder_xyz, self.embedder_dir, self.precision, self.pairing = embedder_xyz, embedder_dir, precision, pairing
        self.t_rand, self.u, self.coarse = t_rand, u, None
        self.loss_target = loss_target      # codenerf_b200.ops.LossTargetSpec: fold train.py:103-104's loss into the call


def _fused_render_ok(rays, point_sampler, embedders, coarse_model, fine_model) -> bool:
    """The one-call pipeline serves the training / rendering case: two CodeNeRFModels, both embedders, CUDA rays that
    carry no gradient (camera-pose optimisation, eval.py:129-168, keeps the per-stage path whose graph reaches the rays).
    CODENERF_FUSED_RENDER=0 forces the per-stage path."""
    from ..models.model import CodeNeRFModel
    if os.environ.get("CODENERF_FUSED_RENDER", "1") == "0":
        return False
    ro, rd = rays
    if not (torch.is_tensor(ro) and torch.is_tensor(rd) and ro.is_cuda and rd.is_cuda and ro.dim() == 2 and rd.dim() == 2):
        return False
    if torch.is_grad_enabled() and (ro.requires_grad or rd.requires_grad):
        return False
    if not (isinstance(_unwrap(coarse_model), CodeNeRFModel) and isinstance(_unwrap(fine_model), CodeNeRFModel)):
        return False
    if not isinstance(point_sampler, PointSampler) or embedders[0] is None or embedders[1] is None:
        return False
    return point_sampler.num_samples_coarse >= 3 and point_sampler.num_samples_fine > 0


def _unwrap(model):
    return model.module if hasattr(model, "module") and isinstance(model.module, torch.nn.Module) else model


def _get_minibatches(inputs: torch.Tensor, chunksize: int):
    """utils/util.py:230-235."""
    return [inputs[i: i + chunksize] for i in range(0, inputs.shape[0], chunksize)]


def _is_main_process(is_distributed: bool) -> bool:
    """utils/util.py:38-41."""
    return (not is_distributed) or dist.get_rank() == 0


def prepare_samplers(cfg, height: int, width: int, intrinsics, datatype, device) -> Tuple[RaySampler, PointSampler]:
    """nerf/__init__.py:15-39."""
    ray_sampler = RaySampler(height, width, intrinsics, sample_size=cfg.nerf.ray_sampler.num_random_rays,
                             device=device, datatype=datatype)
    point_sampler = PointSampler(cfg.nerf.point_sampler.num_coarse, cfg.nerf.point_sampler.num_fine,
                                 cfg.nerf.point_sampler.near_limit, cfg.nerf.point_sampler.far_limit,
                                 spacing_mode=cfg.nerf.point_sampler.spacing_mode,
                                 perturb=cfg.nerf.point_sampler.perturb, dtype=datatype, device=device)
    return ray_sampler, point_sampler


def prepare_embedders(cfg, datatype, device) -> Tuple[PositionalEmbedder, Union[PositionalEmbedder, None]]:
    """nerf/__init__.py:42-71."""
    embedder_xyz = PositionalEmbedder(num_freq=cfg.nerf.embedder.num_encoding_fn_xyz,
                                      log_sampling=cfg.nerf.embedder.log_sampling_xyz,
                                      include_input=cfg.nerf.embedder.include_input_xyz, dtype=datatype, device=device)
    embedder_dir = None
    if cfg.nerf.embedder.use_viewdirs:
        embedder_dir = PositionalEmbedder(num_freq=cfg.nerf.embedder.num_encoding_fn_dir,
                                          log_sampling=cfg.nerf.embedder.log_sampling_dir,
                                          include_input=cfg.nerf.embedder.include_input_dir, dtype=datatype,
                                          device=device)
    return embedder_xyz, embedder_dir


def predict_radiance_and_render(rays, point_sampler: PointSampler, embedders, coarse_model: torch.nn.Module,
                                fine_model: torch.nn.Module, latent_embedding) -> Tuple[torch.Tensor, torch.Tensor]:
    """coarse sample -> field -> composite -> importance sample -> field -> composite (nerf/__init__.py:74-91)."""
    (ro, rd) = rays
    if _fused_render_ok(rays, point_sampler, embedders, coarse_model, fine_model):
        # the same two torch.rand draws, in the same order, as sample_uniform (:64) and sample_pdf (:93) make
        t_rand, u = point_sampler.draw_offsets(ro.shape[0], ro.dtype, ro.device)
        query = RenderQuery(ro, rd, point_sampler, embedders[0], embedders[1], _state["precision"], _state["pairing"], t_rand, u)
        z_s, z_t = latent_embedding
        query.coarse = coarse_model(z_s, z_t, query)
        rgb_coarse, rgb_fine = fine_model(z_s, z_t, query)
        return rgb_coarse, rgb_fine
    pts_coarse, z_vals_coarse = point_sampler.sample_uniform(ro, rd)
    radiance_field_coarse = forward_pass(coarse_model, embedders, rd, pts_coarse, latent_embedding)
    (rgb_coarse, _, _, weights, _) = volume_render(radiance_field_coarse, z_vals_coarse, rd)

    pts_fine, z_vals_fine = point_sampler.sample_pdf(ro, rd, weights[..., 1:-1], z_vals_coarse)
    radiance_field_fine = forward_pass(fine_model, embedders, rd, pts_fine, latent_embedding)
    (rgb_fine, _, _, _, _) = volume_render(radiance_field_fine, z_vals_fine, rd)
    return rgb_coarse, rgb_fine


def predict_radiance_render_and_loss(rays, point_sampler: PointSampler, embedders, coarse_model: torch.nn.Module,
                                     fine_model: torch.nn.Module, latent_embedding, target_pixels=None, views_u8=None,
                                     pixel_index=None):
    """Not in the reference: predict_radiance_and_render plus the photometric loss of
    train.py:103-104 with the loss -- and, given `views_u8` / `pixel_index`, the target-pixel gather of train.py:77-80 --
    folded into the compositing epilogues.  -> (coarse_loss, fine_loss, loss, rgb_coarse, rgb_fine); the rgb tensors are
    detached (their gradient is the loss's).  Targets: `target_pixels` (N, >= 3) float rows, or uint8 views (..., C) resident
    on the device (codenerf_b200.data.DeviceSRN.colors_u8) with `pixel_index` (N,) int64 = view * H W + selected pixel.
    Callers the fused entry does not serve (pose gradients, other model classes) get the staged composition."""
    from codenerf_b200 import ops
    (ro, rd) = rays
    if _fused_render_ok(rays, point_sampler, embedders, coarse_model, fine_model):
        t_rand, u = point_sampler.draw_offsets(ro.shape[0], ro.dtype, ro.device)
        target = ops.LossTargetSpec(rows=target_pixels, views=views_u8, pixel=pixel_index)
        query = RenderQuery(ro, rd, point_sampler, embedders[0], embedders[1], _state["precision"], _state["pairing"], t_rand, u,
                            loss_target=target)
        z_s, z_t = latent_embedding
        query.coarse = coarse_model(z_s, z_t, query)
        coarse_loss, fine_loss, loss, rgb = fine_model(z_s, z_t, query)
        return coarse_loss, fine_loss, loss, rgb[0], rgb[1]
    rgb_coarse, rgb_fine = predict_radiance_and_render(rays, point_sampler, embedders, coarse_model, fine_model, latent_embedding)
    if target_pixels is None:
        flat = views_u8.reshape(-1, views_u8.shape[-1])
        target_pixels = flat[pixel_index].to(torch.float64).div(255.0).to(torch.float32)
    coarse_loss, fine_loss, loss = ops.mse_pair(rgb_coarse, rgb_fine, target_pixels)
    return coarse_loss, fine_loss, loss, rgb_coarse.detach(), rgb_fine.detach()


def forward_pass(model: torch.nn.Module, embedders, rd, pts: torch.Tensor, object_embedding) -> torch.Tensor:
    """Embed the sampled points (and the view directions), run the field: (N, S, 3) -> (N, S, 4).  nerf/__init__.py:94-134.

    For a CodeNeRFModel with view directions the per-sample expansion of the codes (:118-119), the (N*S, 90) embedded
    tensor (:123-130) and the per-sample evaluation of the code layers never materialise: the model receives a
    FieldQuery and runs the fused kernels.  Any other module gets the reference's composition verbatim (the encodings
    still come from the CUDA operator)."""
    from ..models.model import CodeNeRFModel, FlexibleNeRFModel
    num_rays, num_samples = pts.shape[0], pts.shape[1]
    [shape_embedding, texture_embedding] = object_embedding
    embedder_xyz, embedder_dir = embedders
    if isinstance(_unwrap(model), CodeNeRFModel) and rd is not None and embedder_dir is not None:
        query = FieldQuery(pts, rd, embedder_xyz, embedder_dir, _state["precision"], _state["pairing"])
        return model(shape_embedding, texture_embedding, query)
    if isinstance(_unwrap(model), FlexibleNeRFModel) and rd is not None and embedder_dir is not None:
        # The reference cannot reach this class from forward_pass (it calls model(z_s, z_t, x) on a one-argument forward:
        # TypeError); the drop-in runs it -- the codes are ignored, as the 3-line adapter of SURVEY 8(c) would.
        query = FieldQuery(pts, rd, embedder_xyz, embedder_dir, _state["precision"], _state["pairing"])
        return model(query)

    shape_embedding = shape_embedding[:, None, :].expand(-1, num_samples, -1)
    texture_embedding = texture_embedding[:, None, :].expand(-1, num_samples, -1)
    shape_embedding = shape_embedding.reshape(-1, shape_embedding.shape[-1])
    texture_embedding = texture_embedding.reshape(-1, texture_embedding.shape[-1])
    embedded = embedder_xyz.embed(pts.reshape(-1, pts.shape[-1]))
    if rd is not None:
        viewdirs = rd / rd.norm(p=2, dim=-1).unsqueeze(-1)
        if _state["pairing"] == "reference":
            input_dirs_flat = viewdirs.repeat([1, num_samples, 1]).reshape(-1, 3)          # :127 (F6)
        else:
            input_dirs_flat = viewdirs[:, None, :].expand(-1, num_samples, -1).reshape(-1, 3)
        embedded = torch.cat((embedded, embedder_dir.embed(input_dirs_flat)), dim=-1)
    radiance_field = model(shape_embedding, texture_embedding, embedded)
    return radiance_field.reshape([num_rays, num_samples, radiance_field.shape[-1]])


def render_split(num_rays: int, n_gpus: int) -> Tuple[List[int], List[int]]:
    """Rays per rank and zero-padding per rank of parallel_image_render (nerf/__init__.py:179-187): floor split,
    remainder on the last rank, every other rank padded up to it."""
    per = [int(num_rays / n_gpus)] * n_gpus
    padding = num_rays - sum(per)
    per[-1] = num_rays - sum(per[:-1])
    assert sum(per) == num_rays, "Mismatch in batchsize per process and total number of rays"
    pad = [0] * n_gpus
    if padding > 0:
        pad[:-1] = [padding] * (n_gpus - 1)
    assert padding + per[0] == per[-1], "Incorrect calculation of padding"
    return per, pad


def parallel_image_render(cfg, pose: torch.Tensor, object_embedding, models, samplers, embedders, device):
    """Ray-sharded no-grad render of one full image (nerf/__init__.py:137-226): contiguous split of the H*W rays over
    ranks, no communication while rendering, one all_gather of the padded rgb rows; rank 0 returns (H*W, 3), the
    others None."""
    rank = 0
    is_distributed = hasattr(cfg, "is_distributed") and cfg.is_distributed
    n_gpus = 1 if not hasattr(cfg, "gpus") else cfg.gpus
    if is_distributed:
        rank = dist.get_rank()
    for _, model in models.items():
        model.eval()

    with torch.no_grad():
        ray_sampler, point_sampler = samplers
        ray_origins, ray_directions = ray_sampler.get_bundle(tform_cam2world=pose)
        ro, rd = ray_origins.reshape(-1, 3), ray_directions.reshape(-1, 3)
        num_rays = ro.shape[0]
        [shape_embedding, texture_embedding] = object_embedding
        shape_embedding = shape_embedding.expand(num_rays, -1)
        texture_embedding = texture_embedding.expand(num_rays, -1)

        per, pad = render_split(num_rays, n_gpus)
        lo = sum(per[:rank])
        ro_batch, rd_batch = ro[lo: lo + per[rank]].to(device), rd[lo: lo + per[rank]].to(device)
        zs_batch = shape_embedding[lo: lo + per[rank]].to(device)      # stays a stride-0 view: never materialised
        zt_batch = texture_embedding[lo: lo + per[rank]].to(device)

        chunk = cfg.nerf.validation.chunksize
        rgb_batches = []
        for ro_c, rd_c, z_s, z_t in zip(_get_minibatches(ro_batch, chunk), _get_minibatches(rd_batch, chunk),
                                        _get_minibatches(zs_batch, chunk), _get_minibatches(zt_batch, chunk)):
            _, rgb_fine = predict_radiance_and_render((ro_c, rd_c), point_sampler, embedders, models["nerf_coarse"],
                                                      models["nerf_fine"], (z_s, z_t))
            rgb_batches.append(rgb_fine)
        rgb_batches = torch.cat(rgb_batches, dim=0)
        if not is_distributed:
            return rgb_batches

        padded_rgb = torch.zeros((pad[rank], rgb_batches.shape[-1]), dtype=rgb_batches.dtype, device=rgb_batches.device)
        rgb_batches = torch.cat([rgb_batches, padded_rgb], dim=0)
        all_rgb_batches = [torch.zeros_like(rgb_batches) for _ in range(n_gpus)]
        dist.all_gather(all_rgb_batches, rgb_batches)
        if _is_main_process(is_distributed):
            return torch.cat([all_rgb_batches[i][: per[i], ...] for i in range(n_gpus)], dim=0)
        return None
