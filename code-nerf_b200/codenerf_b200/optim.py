"""Fused AdamW for the training step of train.py:111-114 ("next" row f3).

`FusedAdamW` is a drop-in for `torch.optim.AdamW` (same constructor arguments, same per-parameter state keys `step`,
`exp_avg`, `exp_avg_sq`, so `state_dict()` / `load_state_dict()` interchange with the reference's checkpoints,
`utils/util.py:147-172`).  `step()` updates every CUDA fp32 parameter that has a gradient with ONE libcnrf launch
(csrc/adamw.cu); it never synchronises.  Anything else (CPU tensors, other dtypes, sparse gradients, amsgrad / maximize)
is refused loudly -- there is no silent fallback.

`capturable=True` (torch.optim.AdamW's name for it) keeps ONE step count on the device, shared by all parameters and
incremented by the launch itself (cnrf_adamw_step_dev), so `step()` can be captured in a CUDA graph together with the
forward and backward and replayed: no host value changes from step to step.  It requires every parameter to receive a
gradient in every step (true for the CodeNeRF training step) and a constant learning rate between captures.
"""
import ctypes as C

import torch

from . import _lib


class _AdamWTensor(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("numel", C.c_int64), ("lr", C.c_float), ("weight_decay", C.c_float), ("step", C.c_float)]


class FusedAdamW(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, amsgrad=False, maximize=False,
                 capturable=False):
        if amsgrad or maximize:
            raise ValueError("FusedAdamW implements plain AdamW only (amsgrad=False, maximize=False)")
        if not 0.0 <= lr or not 0.0 <= eps or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0 or not 0.0 <= weight_decay:
            raise ValueError("invalid AdamW hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False,
                                      maximize=False))
        self._n = {}                # python-side step counters (state["step"] tensors are refreshed lazily, see _sync_steps)
        self.capturable = bool(capturable)
        self._step_dev = None       # capturable: int64 device scalar, the step count of every parameter
        self._tabs = {}             # (b1, b2, eps, count) -> reusable ctypes table

    def _init_state(self, p):
        st = self.state[p]
        if len(st) == 0:
            st["step"] = torch.tensor(0.0, dtype=torch.float32)
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    def _sync_steps(self):
        """torch.optim.AdamW keeps `step` as a tensor in the per-parameter state; here the counters live in Python between
        checkpoints (38 tensor increments per step cost more host time than the optimiser launch itself)."""
        if self._step_dev is not None:          # capturable: graph replays advanced the device counter, not Python's (one read)
            n = int(self._step_dev.item())
            self._n = {p: n for p in self._n}
        for p, n in self._n.items():
            if p in self.state and len(self.state[p]):
                self.state[p]["step"] = torch.tensor(float(n), dtype=torch.float32)

    def state_dict(self):
        self._sync_steps()
        return super().state_dict()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._n = {p: int(float(st["step"])) for p, st in self.state.items() if "step" in st}

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        by_betas = {}
        for group in self.param_groups:
            key = (float(group["betas"][0]), float(group["betas"][1]), float(group["eps"]))
            rows = by_betas.setdefault(key, [])
            for p in group["params"]:
                if p.grad is None:
                    continue
                g = p.grad
                if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous()) or g.is_sparse or g.dtype != torch.float32:
                    raise _lib.CnrfError("FusedAdamW needs contiguous CUDA fp32 parameters with dense fp32 gradients")
                rows.append((p, g if g.is_contiguous() else g.contiguous(), None, float(group["lr"]),
                             float(group["weight_decay"])))
        if not any(by_betas.values()):
            return loss
        # every parameter passed validation: only now touch optimiser state (a raise above leaves all step counters alone)
        lib = _lib.load()
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        for (b1, b2, eps), rows in by_betas.items():
            if not rows:
                continue
            key = (b1, b2, eps, len(rows))
            tab = self._tabs.get(key)
            if tab is None:
                tab = self._tabs[key] = (_AdamWTensor * len(rows))()
            if self.capturable and self._step_dev is None:
                counts = {self._n.get(p, int(float(self._init_state(p)["step"]))) for (p, _, _, _, _) in rows}
                if len(counts) != 1 or len(by_betas) != 1:
                    raise _lib.CnrfError("FusedAdamW(capturable=True) needs one (betas, eps) setting and parameters that share a step count")
                self._step_dev = torch.tensor(counts.pop(), dtype=torch.int64, device=rows[0][0].device)
            for i, (p, g, _, lr, wd) in enumerate(rows):
                st = self._init_state(p)
                n = self._n.get(p)
                if n is None:
                    n = int(float(st["step"]))
                n += 1
                self._n[p] = n
                t = tab[i]
                t.param, t.grad = p.data_ptr(), g.data_ptr()
                t.exp_avg, t.exp_avg_sq = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
                t.numel, t.lr, t.weight_decay, t.step = p.numel(), lr, wd, float(n)
            if self.capturable:
                _lib.check(lib.cnrf_adamw_step_dev(C.cast(tab, C.c_void_p), len(rows), b1, b2, eps,
                                                   C.c_void_p(self._step_dev.data_ptr()), stream), "cnrf_adamw_step_dev")
            else:
                _lib.check(lib.cnrf_adamw_step(C.cast(tab, C.c_void_p), len(rows), b1, b2, eps, stream), "cnrf_adamw_step")
        return loss
