"""Networks of the hot path -- constructor signatures, parameter names and shapes of the reference's
view_synthesis/models/model.py:79-194, so reference checkpoints load both ways and DDP / optimizer param groups see
the same modules.  The arithmetic runs in libcnrf.so."""
from typing import Tuple

import torch

from codenerf_b200 import _lib, ops

_PRECISION_IDS = {"fp32": _lib.PRECISION_FP32, "bf16": _lib.PRECISION_BF16, "split": _lib.PRECISION_SPLIT}

__all__ = ["FlexibleNeRFModel", "CodeNeRFModel", "ShapeTextureEmbedding", "get_params_tensor"]


class FlexibleNeRFModel(torch.nn.Module):
    """Vanilla NeRF MLP (models/model.py:5-76; config/lego.yml: 8 x 256, skip_connect_ids [4]).  Same constructor, module
    names (layer1, layers_xyz.N, fc_feat, layers_dir.0, fc_alpha, fc_rgb) and parameter shapes as the reference, so its
    checkpoints load both ways; the Linear modules only HOLD the parameters, the arithmetic runs in libcnrf.so
    (cnrf_flexnerf_fwd / _bwd).  forward(x) accepts
      * x: a tensor (..., dim_xyz + dim_dir) of embedded inputs -- the reference's literal signature -> (..., 4);
      * x: a view_synthesis.nerf.FieldQuery (what our forward_pass passes): points and ray directions are embedded by the
        CUDA operator and the view-direction half of layers_dir[0] is evaluated once per ray -> (N, S, 4).
    use_viewdirs=False (:47-48,75-76) keeps layer1 / layers_xyz and ends in fc_out (hidden -> 4); the direction columns of x,
    if any, are ignored as in the reference.  It runs on the per-layer toolkit GEMMs (cnrf_flexnerf_* with dim_dir = 0)."""

    def __init__(self, num_layers=4, hidden_size=128, skip_connect_ids=[4], num_encoding_fn_xyz=6, num_encoding_fn_dir=4,
                 include_input_xyz=True, include_input_dir=True, use_viewdirs=True):
        super(FlexibleNeRFModel, self).__init__()
        self.num_layers = num_layers
        self.hidden_size = hidden_size
        self.skip_connect_ids = skip_connect_ids
        include_input_xyz = 3 if include_input_xyz else 0
        include_input_dir = 3 if include_input_dir else 0
        self.use_viewdirs = use_viewdirs
        self.dim_xyz = include_input_xyz + 2 * 3 * num_encoding_fn_xyz
        self.dim_dir = include_input_dir + 2 * 3 * num_encoding_fn_dir
        assert num_layers - 1 <= _lib.FLEX_MAX_XYZ_LAYERS and hidden_size % 2 == 0
        self.layer1 = torch.nn.Linear(self.dim_xyz, hidden_size)
        self.layers_xyz = torch.nn.ModuleList()
        for i in range(self.num_layers - 1):
            if i in self.skip_connect_ids:
                self.layers_xyz.append(torch.nn.Linear(self.dim_xyz + hidden_size, hidden_size))
            else:
                self.layers_xyz.append(torch.nn.Linear(hidden_size, hidden_size))
        if use_viewdirs:
            self.fc_feat = torch.nn.Linear(hidden_size, hidden_size)
            self.layers_dir = torch.nn.ModuleList()
            self.layers_dir.append(torch.nn.Linear(self.dim_dir + hidden_size, hidden_size // 2))
            self.fc_alpha = torch.nn.Linear(hidden_size, 1)
            self.fc_rgb = torch.nn.Linear(hidden_size // 2, 3)
        else:
            self.fc_out = torch.nn.Linear(hidden_size, 4)
        self.relu = torch.nn.functional.relu

    def _param_list(self):
        head = [self.fc_feat, self.layers_dir[0], self.fc_alpha, self.fc_rgb] if self.use_viewdirs else [self.fc_out]
        layers = [self.layer1] + list(self.layers_xyz) + head
        out = []
        for layer in layers:
            out += [layer.weight, layer.bias]
        return out

    def _skip_mask(self):
        return sum(1 << i for i in range(self.num_layers - 1) if i in self.skip_connect_ids)

    def forward(self, x):
        from ..nerf import FieldQuery
        params = self._param_list()
        if isinstance(x, FieldQuery):
            ex, ed = x.embedder_xyz, x.embedder_dir
            assert 3 * ex.out_dim_per_channel == self.dim_xyz, "embedder width does not match the model's dim_xyz"
            N, S = x.pts.shape[0], x.pts.shape[1]
            precision = _PRECISION_IDS[x.precision]
            pairing = {"reference": _lib.PAIRING_REFERENCE, "correct": _lib.PAIRING_CORRECT}[x.pairing]
            if self.use_viewdirs and x.pts.shape[-1] == 3 and ops.flexnerf_fused_inference(
                    self.num_layers, self.hidden_size, self.dim_xyz, self.dim_dir, self._skip_mask(), precision):
                # rendering: the fused chain encodes the sample positions itself, gamma(x) never exists in HBM
                assert 3 * ed.out_dim_per_channel == self.dim_dir, "embedder width does not match the model's dim_dir"
                viewdirs = x.rd / x.rd.norm(p=2, dim=-1).unsqueeze(-1)                   # nerf/__init__.py:126
                raw = ops.flexnerf_pts(params, x.pts.reshape(-1, 3), ex.frequency_bands, ex.include_input, ed.embed(viewdirs),
                                       self.num_layers, self.hidden_size, self.dim_xyz, self.dim_dir, self._skip_mask(), S,
                                       pairing=pairing, precision=precision)
                return raw.reshape(N, S, 4)
            x_xyz = ex.embed(x.pts.reshape(-1, x.pts.shape[-1]))
            if self.use_viewdirs:
                assert 3 * ed.out_dim_per_channel == self.dim_dir, "embedder width does not match the model's dim_dir"
                viewdirs = x.rd / x.rd.norm(p=2, dim=-1).unsqueeze(-1)                   # nerf/__init__.py:126
                x_dir = ed.embed(viewdirs)
            else:
                x_dir = x_xyz.new_empty((N, 0))
            raw = ops.flexnerf(params, x_xyz, x_dir, self.num_layers, self.hidden_size, self.dim_xyz,
                               self.dim_dir if self.use_viewdirs else 0, self._skip_mask(), S, pairing=pairing,
                               precision=precision)
            return raw.reshape(N, S, 4)
        lead = x.shape[:-1]
        x2 = x.reshape(-1, x.shape[-1])
        x_dir = x2[:, self.dim_xyz:] if self.use_viewdirs else x2.new_empty((x2.shape[0], 0))
        raw = ops.flexnerf(params, x2[:, : self.dim_xyz], x_dir, self.num_layers, self.hidden_size, self.dim_xyz,
                           self.dim_dir if self.use_viewdirs else 0, self._skip_mask(), 1, pairing=_lib.PAIRING_CORRECT,
                           precision=_lib.PRECISION_FP32)
        return raw.reshape(*lead, 4)


def get_params_tensor(model, is_distributed):
    """models/model.py:79-84."""
    if is_distributed:
        shape_params, texture_params = model.module.get_params_tensor()
    else:
        shape_params, texture_params = model.get_params_tensor()
    return shape_params, texture_params


def _share_codes() -> bool:
    """CODENERF_SHARE_CODES=0 keeps the per-ray code path even when the ids are known (tests, A/B timing)."""
    import os
    return os.environ.get("CODENERF_SHARE_CODES", "1") != "0"


class ShapeTextureEmbedding(torch.nn.Module):
    """Two nn.Embedding tables (models/model.py:87-120).  The row gather and its sparse scatter backward stay in
    torch: the render kernels consume per-ray codes and return per-ray code gradients."""

    def __init__(self, num_embeddings, shape_code_size=128, texture_code_size=128):
        super(ShapeTextureEmbedding, self).__init__()
        self.num_embeddings = num_embeddings
        self.shape_code_size = shape_code_size
        self.texture_code_size = texture_code_size
        self.shape_embedding = torch.nn.Embedding(self.num_embeddings, self.shape_code_size)
        self.texture_embedding = torch.nn.Embedding(self.num_embeddings, self.texture_code_size)

    def forward(self, object_ids: torch.Tensor):
        z_s, z_t = self.shape_embedding(object_ids), self.texture_embedding(object_ids)
        # Where the codes came from travels with them: CodeNeRFModel's fused field then looks the rows up itself and
        # shares the code layers between the rays of one object (train.py:78 gives a whole image one id).  Any other
        # consumer sees two ordinary tensors.
        # The tag records the versions of the four tensors: an in-place edit of z_s / z_t (noise, clamping) or of the tables
        # after the lookup makes CodeNeRFModel fall back to the per-ray path, which reads the VALUES of z_s / z_t as the
        # reference does.  In shared-code mode the gradients flow to the tables directly (hooks on z_s / z_t do not fire).
        ws, wt = self.shape_embedding.weight, self.texture_embedding.weight
        src = (object_ids, ws, wt, (z_s._version, z_t._version, ws._version, wt._version))
        z_s._cnrf_code_source = src
        z_t._cnrf_code_source = src
        return z_s, z_t

    def get_all_embeddings(self, device):
        all_idx = torch.arange(0, self.num_embeddings, dtype=int, device=device)
        return self.shape_embedding(all_idx), self.texture_embedding(all_idx)

    def get_params_tensor(self) -> Tuple[torch.Tensor, torch.Tensor]:
        """models/model.py:107-119: the two tables, detached and flattened (F10).  The reference builds each with
        `torch.cat([x.view(-1) for x in params.data])`, i.e. one view per table ROW and a 2458-input cat -- ~6 ms of host
        time per call at SRN-cars size, paid twice per chunk by train.py:106.  `reshape(-1).clone()` of the detached table
        returns the same values in the same order as a fresh tensor, in one launch."""
        shape_params, texture_params = None, None
        for name, params in self.named_parameters():
            if 'shape' in name:
                shape_params = params.data.reshape(-1).clone()
            if 'texture' in name:
                texture_params = params.data.reshape(-1).clone()
        return shape_params, texture_params


class _RenderToken(object):
    """What the coarse model hands to the fine one in the fused render protocol: itself and its 18 parameters."""
    __slots__ = ("model", "params")

    def __init__(self, model, params):
        self.model, self.params = model, params


def _code_source(z_s, z_t):
    """(ids, shape table, texture table) when both code tensors are the untouched output of ShapeTextureEmbedding.forward."""
    src = getattr(z_s, "_cnrf_code_source", None)
    if src is None or getattr(z_t, "_cnrf_code_source", None) is not src or not _share_codes():
        return None
    if (z_s._version, z_t._version, src[1]._version, src[2]._version) != src[3]:
        return None                      # somebody edited the codes or the tables in place since the lookup
    return src[:3]


class CodeNeRFModel(torch.nn.Module):
    """CodeNeRF MLP conditioned on shape / texture codes (models/model.py:123-194).

    The nine torch.nn.Linear modules only HOLD the parameters (names, shapes and default init as the reference);
    forward never calls them.  forward(z_s, z_t, x) accepts
      * x: a view_synthesis.nerf.FieldQuery (what forward_pass passes): z_s / z_t are per-RAY (N, C) -> (N, S, 4);
      * x: a tensor (M, dim_xyz + dim_dir) with per-sample codes (M, C), the reference's literal signature -> (M, 4).
    """

    def __init__(self, hidden_size=128, num_embeddings=1, shape_code_size=128, texture_code_size=128,
                 num_encoding_fn_xyz=6, num_encoding_fn_dir=4, include_input_xyz=True, include_input_dir=True):
        super(CodeNeRFModel, self).__init__()
        self.hidden_size = hidden_size
        self.shape_code_size = shape_code_size
        self.texture_code_size = texture_code_size
        include_input_xyz = 3 if include_input_xyz else 0
        include_input_dir = 3 if include_input_dir else 0
        self.dim_xyz = include_input_xyz + 2 * 3 * num_encoding_fn_xyz
        self.dim_dir = include_input_dir + 2 * 3 * num_encoding_fn_dir

        self.layer_xyz1 = torch.nn.Linear(self.dim_xyz, self.hidden_size)
        self.layer_xyz2 = torch.nn.Linear(self.hidden_size + self.shape_code_size, self.hidden_size)
        self.fc_out = torch.nn.Linear(self.hidden_size + self.shape_code_size, self.shape_code_size + 1)
        self.shape_code_layer1 = torch.nn.Linear(self.shape_code_size, self.shape_code_size)
        self.shape_code_layer2 = torch.nn.Linear(self.shape_code_size, self.shape_code_size)
        self.texture_code_layer1 = torch.nn.Linear(self.shape_code_size, self.shape_code_size)
        self.layer_dir1 = torch.nn.Linear(self.dim_dir + self.shape_code_size, self.hidden_size)
        self.layer_dir2 = torch.nn.Linear(self.hidden_size, self.hidden_size)
        self.fc_rgb = torch.nn.Linear(self.hidden_size + self.texture_code_size, 3)
        self.activation = torch.nn.functional.relu

    def _param_list(self):
        out = []
        for name in _lib.LAYER_NAMES:
            layer = getattr(self, name)
            out += [layer.weight, layer.bias]
        return out

    def forward(self, z_s: torch.Tensor, z_t: torch.Tensor, x):
        from ..nerf import FieldQuery, RenderQuery
        if self.shape_code_size != self.texture_code_size:
            raise ValueError("CodeNeRFModel needs shape_code_size == texture_code_size (models/model.py:150,153)")
        params = self._param_list()
        if isinstance(x, RenderQuery):
            # predict_radiance_and_render as one library call (cnrf_render_fwd / _bwd).  The call still passes through
            # BOTH models' __call__ (and so through DistributedDataParallel.forward when they are wrapped, utils/util.py:
            # 139-142): the coarse model answers with its parameter list, the fine model runs the pipeline.
            if x.coarse is None:
                return _RenderToken(self, params)
            return self._render(z_s, z_t, x, params)
        if isinstance(x, FieldQuery):
            ex, ed = x.embedder_xyz, x.embedder_dir
            assert 3 * ex.out_dim_per_channel == self.dim_xyz and 3 * ed.out_dim_per_channel == self.dim_dir, \
                "embedder widths do not match the model's dim_xyz / dim_dir"
            precision = _PRECISION_IDS[x.precision]
            pairing = {"reference": _lib.PAIRING_REFERENCE, "correct": _lib.PAIRING_CORRECT}[x.pairing]
            src = _code_source(z_s, z_t)
            return ops.codenerf_field(params, x.pts, x.rd, z_s, z_t, ex.frequency_bands, ed.frequency_bands,
                                      self.hidden_size, self.shape_code_size, self.dim_xyz, self.dim_dir,
                                      pairing=pairing, precision=precision, code_source=src)
        lead = x.shape[:-1]
        return self._as_written(params, z_s, z_t, x, lead)

    def _render(self, z_s, z_t, q, params):
        """Fine model's half of the fused render protocol: q.coarse is the coarse model's _RenderToken."""
        coarse = q.coarse.model
        assert (coarse.hidden_size, coarse.shape_code_size, coarse.dim_xyz, coarse.dim_dir) == \
            (self.hidden_size, self.shape_code_size, self.dim_xyz, self.dim_dir), "coarse and fine models differ in shape"
        ex, ed, ps = q.embedder_xyz, q.embedder_dir, q.point_sampler
        assert 3 * ex.out_dim_per_channel == self.dim_xyz and 3 * ed.out_dim_per_channel == self.dim_dir, \
            "embedder widths do not match the model's dim_xyz / dim_dir"
        precision = _PRECISION_IDS[q.precision]
        pairing = {"reference": _lib.PAIRING_REFERENCE, "correct": _lib.PAIRING_CORRECT}[q.pairing]
        src = _code_source(z_s, z_t) if precision == _lib.PRECISION_BF16 else None
        ids = None
        if src is not None and src[0].dim() == 1 and src[0].shape[0] == q.ro.shape[0] and src[0].dtype == torch.int64:
            ids, z_s, z_t = src[0].contiguous(), src[1], src[2]
        spec = ops.RenderSpec(self.hidden_size, self.shape_code_size, self.dim_xyz, self.dim_dir, ps.num_samples_coarse,
                              ps.num_samples_fine, pairing, precision, ps.z_vals, ps.lower, ps.upper, ex.frequency_bands,
                              ed.frequency_bands, ids=ids)
        if q.loss_target is not None:
            return ops.render_loss(spec, q.loss_target, q.ro, q.rd, q.t_rand, q.u, z_s, z_t, q.coarse.params, params)
        return ops.render(spec, q.ro, q.rd, q.t_rand, q.u, z_s, z_t, q.coarse.params, params)

    def _as_written(self, params, z_s, z_t, x, lead):
        out = ops.codenerf_model(params, z_s.reshape(-1, z_s.shape[-1]), z_t.reshape(-1, z_t.shape[-1]),
                                 x.reshape(-1, x.shape[-1]), self.hidden_size, self.shape_code_size, self.dim_xyz,
                                 self.dim_dir, precision=_lib.PRECISION_FP32)
        return out.reshape(*lead, 4)
